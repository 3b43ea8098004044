// oracle/ref_scene.cpp -- TEST INFRASTRUCTURE (never linked into the product).
//
// Runs the reference's own scene loader, tools/sceneBuilder.cpp (RD::Scene::Load, :27-315), compiled UNCHANGED where it lies
// under /root/reference, on top of the reference's own vendored assimp glTF2 importer (external/assimp, compiled where it lies
// by oracle/Makefile: the ~30 source files the importer and the four post-processing steps of sceneBuilder.cpp:32-36 need).
// What the loader hands to the Radiance API -- CreateBuffer / WriteBuffer of the six scene arrays, WriteImage of the resized
// textures, BuildAccelStruct of every mesh and of the instance list -- is RECORDED here instead of going to a device, and
// read back through the C getters at the end of this file. tests/test_scene_loader.py compares the product's loaders
// (radiance-ray-tracing_b200/scene_loader.py, include/sceneBuilder.h) with these recordings array by array: that pins
// node-transform order, instance order, v flip, "*N" texture naming, default materials, normal generation and vertex joining to
// what assimp + sceneBuilder.cpp really do.
#include <sceneBuilder.h>

#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

namespace {

struct Recorded {            // stands in for a cl_mem
    std::vector<unsigned char> bytes;
    unsigned width = 0, height = 0, layers = 0;   // image arrays
};

struct RecordedScene {
    std::vector<Recorded*> all;
    std::vector<RD::Mesh> meshes;                  // BuildAccelStruct(Mesh) calls, in order
    std::vector<RD::_BottomAccelStruct*> blas;     // handle i <-> meshes[i]
    std::vector<float> instanceTransforms;         // 16 per instance, row major (RD::Mat4x4 = aiMatrix4x4: a1..d4)
    std::vector<unsigned> instanceMeta;            // {SBTOffset, customInstanceID, mesh index} per instance
    unsigned samplerAddress = 0, samplerFilter = 0;
    RD::Scene* scene = nullptr;
};

RecordedScene* g_cur = nullptr;

}  // namespace

namespace RD {

Buffer CreateBuffer(Platform*, unsigned int size)
{
    Recorded* r = new Recorded();
    r->bytes.resize(size);
    g_cur->all.push_back(r);
    return (Buffer)r;
}

void WriteBuffer(Platform*, Buffer handle, size_t size, void* data, size_t offset)
{
    Recorded* r = (Recorded*)handle;
    if (offset + size > r->bytes.size()) r->bytes.resize(offset + size);
    memcpy(r->bytes.data() + offset, data, size);
}

ImageArray CreateImageArray(Platform*, unsigned int width, unsigned int height, unsigned int arraySize)
{
    Recorded* r = new Recorded();
    r->width = width; r->height = height; r->layers = arraySize;
    r->bytes.resize((size_t)width * height * 4 * arraySize);
    g_cur->all.push_back(r);
    return (ImageArray)r;
}

void WriteImage(Platform*, ImageArray handle, unsigned int width, unsigned int height, size_t arrayIndex, void* data)
{
    Recorded* r = (Recorded*)handle;
    const size_t layer = (size_t)width * height * 4;
    memcpy(r->bytes.data() + arrayIndex * layer, data, layer);
}

Sampler CreateSampler(Platform*, unsigned int addressingMode, unsigned int filterMode)
{
    g_cur->samplerAddress = addressingMode; g_cur->samplerFilter = filterMode;
    return (Sampler)(uintptr_t)1;
}

BottomAccelStruct BuildAccelStruct(Platform*, Mesh& mesh)
{
    g_cur->meshes.push_back(mesh);
    _BottomAccelStruct* b = new _BottomAccelStruct();
    b->root = nullptr;
    g_cur->blas.push_back(b);
    return b;
}

TopAccelStruct BuildAccelStruct(Platform*, std::vector<Instance>& instances)
{
    for (const Instance& in : instances) {
        const float* m = &in.transform.a1;   // aiMatrix4x4: a1 a2 a3 a4 b1 ... d4, row major
        g_cur->instanceTransforms.insert(g_cur->instanceTransforms.end(), m, m + 16);
        unsigned meshIndex = ~0u;
        for (size_t k = 0; k < g_cur->blas.size(); ++k) if (g_cur->blas[k] == in.bottomAccelStruct) meshIndex = (unsigned)k;
        g_cur->instanceMeta.push_back(in.SBTOffset);
        g_cur->instanceMeta.push_back(in.customInstanceID);
        g_cur->instanceMeta.push_back(meshIndex);
    }
    Recorded* r = new Recorded();
    g_cur->all.push_back(r);
    return (TopAccelStruct)r;
}

void TopAccelStructToFile(Platform*, TopAccelStruct, const char*) {}   // the .cache file is not part of this comparison
void FileToTopAccelStruct(Platform*, const char*, TopAccelStruct* out) { *out = nullptr; }

}  // namespace RD

extern "C" {

// Runs RD::Scene::Load(path, nullptr, false) of the reference; returns an opaque recording or NULL.
void* ref_scene_load(const char* path)
{
    RecordedScene* rec = new RecordedScene();
    g_cur = rec;
    rec->scene = RD::Scene::Load(path, nullptr, false);
    g_cur = nullptr;
    if (!rec->scene) { delete rec; return nullptr; }
    return rec;
}

// which: 0 meshInfoData, 1 vertexData, 2 indexData, 3 uvData, 4 normalData, 5 materialData, 6 textureData
size_t ref_scene_array(void* h, int which, const void** data, unsigned dims[3])
{
    RecordedScene* rec = (RecordedScene*)h;
    RD::Scene* s = rec->scene;
    void* handles[7] = {s->meshInfoData, s->vertexData, s->indexData, s->uvData, s->normalData, s->materialData, s->textureData};
    if (which < 0 || which > 6) return 0;
    Recorded* r = (Recorded*)handles[which];
    *data = r->bytes.data();
    if (dims) { dims[0] = r->width; dims[1] = r->height; dims[2] = r->layers; }
    return r->bytes.size();
}

unsigned ref_scene_mesh_count(void* h) { return (unsigned)((RecordedScene*)h)->meshes.size(); }
// the arrays handed to BuildAccelStruct(Mesh) number i: vertices (3 floats each), triangles (3 indices each)
void ref_scene_mesh(void* h, unsigned i, const float** verts, unsigned* nv, const unsigned** idx, unsigned* nt)
{
    RD::Mesh& m = ((RecordedScene*)h)->meshes[i];
    *verts = (const float*)m.vertexData.data(); *nv = (unsigned)m.vertexData.size();
    *idx = (const unsigned*)m.indexData.data(); *nt = (unsigned)m.indexData.size();
}
unsigned ref_scene_instances(void* h, const float** transforms, const unsigned** meta)
{
    RecordedScene* rec = (RecordedScene*)h;
    *transforms = rec->instanceTransforms.data(); *meta = rec->instanceMeta.data();
    return (unsigned)(rec->instanceMeta.size() / 3);
}
void ref_scene_sampler(void* h, unsigned out[2]) { out[0] = ((RecordedScene*)h)->samplerAddress; out[1] = ((RecordedScene*)h)->samplerFilter; }
void ref_scene_free(void* h)
{
    RecordedScene* rec = (RecordedScene*)h;
    for (Recorded* r : rec->all) delete r;
    for (RD::_BottomAccelStruct* b : rec->blas) delete b;
    delete rec->scene;
    delete rec;
}
unsigned ref_scene_sizes(unsigned out[4])   // sizeof of the structs the arrays hold, for the Python side
{
    out[0] = sizeof(RD::MeshInfo); out[1] = sizeof(RD::Material); out[2] = sizeof(RD::Vec3); out[3] = sizeof(RD::Triangle);
    return 4;
}

}
