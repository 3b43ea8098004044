// tests/cpp/parse_dump.cpp -- CPU-only check of the C++ scene parser (include/sceneBuilder.h, RD::gltf::parse): writes the arrays
// behind the nine descriptors in the format of sample_scene's <prefix>.arrays, without creating a platform (no GPU needed).
// usage: parse_dump <scene.glb|gltf> <out.arrays> [texture dim: also appends embedded image 0 decoded and resized to dim x dim]
#include <cstdio>
#include <cstdlib>
#include <string>
#include <vector>
#include "sceneBuilder.h"

template <class T> static void dump(FILE* f, const std::vector<T>& v)
{
    const unsigned long long n = v.size() * sizeof(T);
    fwrite(&n, 8, 1, f);
    if (n) fwrite(v.data(), 1, n, f);
}

int main(int argc, char** argv)
{
    if (argc < 3) { fprintf(stderr, "usage: parse_dump <scene> <out.arrays>\n"); return 2; }
    const RD::gltf::Parsed P = RD::gltf::parse(argv[1]);
    FILE* f = fopen(argv[2], "wb");
    if (!f) return 4;
    dump(f, P.meshInfo); dump(f, P.vertices); dump(f, P.triangles); dump(f, P.uvs); dump(f, P.normals); dump(f, P.materials);
    std::vector<float> inst;
    for (auto& i : P.instances) { inst.push_back((float)i.mesh); inst.push_back((float)i.customInstanceID); for (int r = 0; r < 4; ++r) for (int c = 0; c < 4; ++c) inst.push_back(i.transform[r][c]); }
    dump(f, inst);
    std::vector<unsigned> imageSizes;
    for (auto& im : P.images) imageSizes.push_back((unsigned)im.size());
    dump(f, imageSizes);
    if (argc > 3 && !P.images.empty()) {
        std::vector<unsigned char> rgba; int w = 0, h = 0;
        if (!RD::gltf::decodePng(P.images[0].data(), P.images[0].size(), rgba, w, h)) { fprintf(stderr, "parse_dump: image 0 is not a PNG\n"); return 5; }
        dump(f, RD::gltf::resizeRgba(rgba, w, h, atoi(argv[3])));
    }
    fclose(f);
    printf("parse_dump: %zu meshes, %zu materials, %zu instances, %zu images\n", P.meshes.size(), P.materials.size(), P.instances.size(), P.images.size());
    return 0;
}
