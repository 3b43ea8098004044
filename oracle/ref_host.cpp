// oracle/ref_host.cpp -- TEST INFRASTRUCTURE (never linked into the product library).
//
// C entry points around the reference's own host BVH builder. radiance/src/bvh.cpp is
// compiled UNCHANGED from /root/reference (see oracle/Makefile); this file only restates the
// two memcpy-style blob packers of radiance/src/radiance.cpp, which cannot be compiled as they
// are because they upload through clEnqueueWriteBuffer:
//   _buildBottomAccelStruct  radiance/src/radiance.cpp:318-364
//   _buildTopAccelStruct     radiance/src/radiance.cpp:366-425
// Blob layouts: radiance/src/core.h:34-101, radiance/shader/data.cl:237-278.
#include "radiance.h"
#include "bvh.h"

#include <cstring>
#include <cstdlib>
#include <map>
#include <vector>

using namespace RD;

extern "C" {

struct RefInstance {
    float transform[16];      // row-major object->world (RD::Instance::transform, radiance.h:50)
    unsigned int SBTOffset;
    unsigned int customInstanceID;
    void* blas;               // handle returned by ref_blas_build
};

// RD::BuildAccelStruct(Platform*, Mesh&)  radiance/src/radiance.cpp:20-52
void* ref_blas_build(const float* verts, unsigned int nv, const unsigned int* idx, unsigned int nt)
{
    Mesh mesh;
    mesh.vertexData.resize(nv);
    for (unsigned int i = 0; i < nv; ++i)
        mesh.vertexData[i] = Vec3(verts[3 * i], verts[3 * i + 1], verts[3 * i + 2]);
    mesh.indexData.resize(nt);
    for (unsigned int i = 0; i < nt; ++i)
        mesh.indexData[i] = Triangle{idx[3 * i], idx[3 * i + 1], idx[3 * i + 2]};

    BVHNode* root = CreateBVH(mesh.vertexData, mesh.indexData);           // reference code
    _BottomAccelStruct* as = new _BottomAccelStruct();
    as->root = root;
    std::vector<DeviceTriangle> faces;
    std::vector<DeviceBVHNode> nodes;
    CreateDeviceBVH(root, mesh.indexData, faces, nodes);                  // reference code

    // pack: [AccelStructBottom 16][nodes 48*Nn][faces 16*Nt][verts 16*Nv], vertex.w left 0
    unsigned int nodeBytes = (unsigned int)(nodes.size() * sizeof(DeviceBVHNode));
    unsigned int faceBytes = (unsigned int)(faces.size() * sizeof(DeviceTriangle));
    unsigned int vertBytes = (unsigned int)(mesh.vertexData.size() * sizeof(DeviceVertex));
    AccelStructBottom hdr;
    hdr.type = TYPE_BOT_AS;
    hdr.nodeByteOffset = sizeof(AccelStructBottom);
    hdr.faceByteOffset = hdr.nodeByteOffset + nodeBytes;
    hdr.vertexOffset = hdr.faceByteOffset + faceBytes;
    as->data.resize(sizeof(AccelStructBottom) + nodeBytes + faceBytes + vertBytes);
    char* p = as->data.data();
    memcpy(p, &hdr, sizeof(hdr));
    memcpy(p + hdr.nodeByteOffset, nodes.data(), nodeBytes);
    memcpy(p + hdr.faceByteOffset, faces.data(), faceBytes);
    DeviceVertex* dv = (DeviceVertex*)(p + hdr.vertexOffset);
    for (size_t i = 0; i < mesh.vertexData.size(); ++i) {
        dv[i].x = mesh.vertexData[i].x;
        dv[i].y = mesh.vertexData[i].y;
        dv[i].z = mesh.vertexData[i].z;
    }
    return as;
}

size_t ref_blas_size(void* h) { return ((_BottomAccelStruct*)h)->data.size(); }
const void* ref_blas_data(void* h) { return ((_BottomAccelStruct*)h)->data.data(); }
void ref_blas_free(void* h) { delete (_BottomAccelStruct*)h; }   // tree nodes leak, as in the reference

// RD::BuildAccelStruct(Platform*, std::vector<Instance>&)  radiance/src/radiance.cpp:54-84
// Returns a malloc'd TLAS blob (the content of the reference's single device buffer).
void* ref_tlas_build(const RefInstance* inst, unsigned int n, size_t* outSize)
{
    std::vector<Instance> instances(n);
    for (unsigned int i = 0; i < n; ++i) {
        const float* m = inst[i].transform;
        instances[i].transform = Mat4x4(m[0], m[1], m[2], m[3], m[4], m[5], m[6], m[7],
                                        m[8], m[9], m[10], m[11], m[12], m[13], m[14], m[15]);
        instances[i].SBTOffset = inst[i].SBTOffset;
        instances[i].customInstanceID = inst[i].customInstanceID;
        instances[i].bottomAccelStruct = (_BottomAccelStruct*)inst[i].blas;
    }
    BVHNode* root = CreateBVH(instances);                                 // reference code
    std::vector<DeviceInstance> dinst;
    std::vector<DeviceBVHNode> nodes;
    std::map<BottomAccelStruct, unsigned int> offsets;
    CreateDeviceBVH(root, instances, dinst, nodes, offsets);              // reference code

    unsigned int nodeBytes = (unsigned int)(nodes.size() * sizeof(DeviceBVHNode));
    unsigned int instBytes = (unsigned int)(dinst.size() * sizeof(DeviceInstance));
    unsigned int blasBytes = 0;
    for (auto& e : offsets) blasBytes += (unsigned int)e.first->data.size();
    AccelStructTop hdr;
    hdr.type = TYPE_TOP_AS;
    hdr.nodeByteOffset = sizeof(AccelStructTop);
    hdr.instByteOffset = hdr.nodeByteOffset + nodeBytes;
    hdr.totalBufferSize = (unsigned int)sizeof(AccelStructTop) + nodeBytes + instBytes + blasBytes;
    char* blob = (char*)calloc(hdr.totalBufferSize, 1);
    memcpy(blob, &hdr, sizeof(hdr));
    memcpy(blob + hdr.nodeByteOffset, nodes.data(), nodeBytes);
    memcpy(blob + hdr.instByteOffset, dinst.data(), instBytes);
    for (auto& e : offsets) memcpy(blob + e.second, e.first->data.data(), e.first->data.size());
    if (outSize) *outSize = hdr.totalBufferSize;
    return blob;
}

void ref_free(void* p) { free(p); }

unsigned int ref_sizeof(int which)
{
    switch (which) {
        case 0: return sizeof(AccelStructTop);
        case 1: return sizeof(AccelStructBottom);
        case 2: return sizeof(DeviceBVHNode);
        case 3: return sizeof(DeviceTriangle);
        case 4: return sizeof(DeviceVertex);
        case 5: return sizeof(DeviceInstance);
        case 6: return sizeof(RayTraceProperties);
        case 7: return sizeof(Material);
        case 8: return sizeof(MeshInfo);
        case 9: return sizeof(DirLight);
        case 10: return sizeof(SceneProperties);
        case 11: return sizeof(PhysicalCamera);
    }
    return 0;
}

}  // extern "C"
