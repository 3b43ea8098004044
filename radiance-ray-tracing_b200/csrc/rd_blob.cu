// rd_blob.cu -- the reference's acceleration-structure blob as an exchange format.
// Replaces RD::TopAccelStructToFile / RD::FileToTopAccelStruct (radiance/src/radiance.cpp:428-479) and
// reads / writes exactly the byte layout its packers produce (radiance.cpp:318-425, radiance/src/core.h:
// 34-101, radiance/shader/data.cl:237-278):
//   [AccelStructTop 16][DeviceBVHNode 48 x N][DeviceInstance 80 x I][bottom blobs ...]
//   bottom blob = [AccelStructBottom 16][DeviceBVHNode 48 x n][DeviceTriangle 16 x f][DeviceVertex 16 x v]
// Loading keeps the geometry and the instance table and rebuilds the native 8-wide structure on the
// GPU (the binary topology of the file is not traversed here). Saving emits a binary BVH in the
// reference's node format (median split, leaves below 8 primitives like bvh.cpp:300) over the
// triangle records of the native structure, so the reference's traversal can walk the file.
#include <stdio.h>
#include <string.h>
#include <algorithm>
#include <map>
#include <vector>
#include <stdlib.h>
#include "rd_internal.h"
#include "rd_build.h"

namespace {
struct BlobTop { uint32_t type, nodeByteOffset, instByteOffset, totalBufferSize; };
struct BlobBottom { uint32_t type, nodeByteOffset, faceByteOffset, vertexOffset; };
struct BlobInstance { float transform[16]; uint32_t SBTOffset, instanceID, customInstanceID, bottomAccelStructOffset; };
struct BlobNode { float lo[4], hi[4]; uint32_t a, b, type, pad; };   // inner: a/b = children; leaf: a = count | 1<<31, b = first
struct BlobFace { uint32_t idx0, idx1, idx2, primID; };
const uint32_t kTypeInst = 1, kTypeTri = 2, kTopAS = 1, kBotAS = 2;

// median-split binary BVH over boxes in the reference's node layout (root = node 0)
struct Bvh2 {
    std::vector<BlobNode> nodes; std::vector<uint32_t> order;
    const float* lo; const float* hi; uint32_t leafType;
    uint32_t build(std::vector<uint32_t>& ids, size_t b, size_t e)
    {
        const uint32_t me = (uint32_t)nodes.size();
        nodes.push_back(BlobNode());
        BlobNode nd; memset(&nd, 0, sizeof(nd));
        for (int a = 0; a < 3; ++a) { nd.lo[a] = 3.0e38f; nd.hi[a] = -3.0e38f; }
        for (size_t i = b; i < e; ++i)
            for (int a = 0; a < 3; ++a) { nd.lo[a] = std::min(nd.lo[a], lo[3 * ids[i] + a]); nd.hi[a] = std::max(nd.hi[a], hi[3 * ids[i] + a]); }
        if (e - b < 8) {
            nd.a = (uint32_t)(e - b) | 0x80000000u; nd.b = (uint32_t)order.size(); nd.type = leafType;
            for (size_t i = b; i < e; ++i) order.push_back(ids[i]);
        } else {
            int axis = 0;
            for (int a = 1; a < 3; ++a) if (nd.hi[a] - nd.lo[a] > nd.hi[axis] - nd.lo[axis]) axis = a;
            const size_t mid = (b + e) / 2;
            std::nth_element(ids.begin() + b, ids.begin() + mid, ids.begin() + e, [&](uint32_t x, uint32_t y) {
                return lo[3 * x + axis] + hi[3 * x + axis] < lo[3 * y + axis] + hi[3 * y + axis]; });
            nd.a = build(ids, b, mid);
            nd.b = build(ids, mid, e);
        }
        nodes[me] = nd;
        return me;
    }
};
}  // namespace

extern "C" int rd_tlas_load_memory(rd_platform* p, const void* data, size_t size, rd_handle* out, rd_blas*** blasOut, unsigned* blasCount)
{
    if (!p || !data || !out) { rd_set_error("rd_tlas_load: null argument"); return RD_E_ARG; }
    const unsigned char* base = (const unsigned char*)data;
    BlobTop top;
    if (size < sizeof(top)) { rd_set_error("rd_tlas_load: file shorter than its header"); return RD_E_IO; }
    memcpy(&top, base, sizeof(top));
    if (top.type != kTopAS || top.totalBufferSize > size || top.instByteOffset > top.totalBufferSize || top.nodeByteOffset != sizeof(top)) {
        rd_set_error("rd_tlas_load: not a top-level acceleration structure blob (type %u, %u of %zu bytes)", top.type, top.totalBufferSize, size);
        return RD_E_IO;
    }
    // the instance table ends where the first bottom-level blob starts
    std::vector<BlobInstance> inst;
    size_t firstBottom = top.totalBufferSize;
    for (size_t off = top.instByteOffset; off + sizeof(BlobInstance) <= firstBottom; off += sizeof(BlobInstance)) {
        BlobInstance I; memcpy(&I, base + off, sizeof(I));
        if (I.bottomAccelStructOffset < top.instByteOffset || I.bottomAccelStructOffset + sizeof(BlobBottom) > top.totalBufferSize) {
            rd_set_error("rd_tlas_load: instance %zu points outside the blob", inst.size()); return RD_E_IO;
        }
        firstBottom = std::min(firstBottom, (size_t)I.bottomAccelStructOffset);
        inst.push_back(I);
    }
    std::vector<uint32_t> offsets;
    for (auto& I : inst) offsets.push_back(I.bottomAccelStructOffset);
    std::sort(offsets.begin(), offsets.end());
    offsets.erase(std::unique(offsets.begin(), offsets.end()), offsets.end());
    std::map<uint32_t, rd_blas*> built;
    std::vector<rd_blas*> list;
    for (size_t k = 0; k < offsets.size(); ++k) {
        const size_t b0 = offsets[k], b1 = k + 1 < offsets.size() ? offsets[k + 1] : top.totalBufferSize;
        BlobBottom hdr; memcpy(&hdr, base + b0, sizeof(hdr));
        if (hdr.type != kBotAS || hdr.faceByteOffset > hdr.vertexOffset || b0 + hdr.vertexOffset > b1) {
            rd_set_error("rd_tlas_load: bottom-level blob at %zu is malformed", b0); return RD_E_IO;
        }
        const size_t nf = (hdr.vertexOffset - hdr.faceByteOffset) / sizeof(BlobFace), nv = (b1 - b0 - hdr.vertexOffset) / 16;
        const BlobFace* faces = (const BlobFace*)(base + b0 + hdr.faceByteOffset);
        const float* v4 = (const float*)(base + b0 + hdr.vertexOffset);
        std::vector<float> verts(3 * nv); std::vector<uint32_t> idx(3 * nf);
        for (size_t i = 0; i < nv; ++i) { verts[3 * i] = v4[4 * i]; verts[3 * i + 1] = v4[4 * i + 1]; verts[3 * i + 2] = v4[4 * i + 2]; }
        for (size_t i = 0; i < nf; ++i) {
            // faces are stored in the file's leaf order; primID is the index in the caller's mesh
            const BlobFace f = faces[i];
            if (f.primID >= nf || f.idx0 >= nv || f.idx1 >= nv || f.idx2 >= nv) { rd_set_error("rd_tlas_load: face %zu out of range", i); return RD_E_IO; }
            idx[3 * f.primID] = f.idx0; idx[3 * f.primID + 1] = f.idx1; idx[3 * f.primID + 2] = f.idx2;
        }
        rd_blas* bl = 0;
        int rc = rd_blas_build(p, verts.data(), (unsigned)nv, idx.data(), (unsigned)nf, &bl);
        if (rc) return rc;
        built[offsets[k]] = bl; list.push_back(bl);
    }
    std::vector<rd_instance_desc> desc(inst.size());
    for (auto& I : inst) {
        if (I.instanceID >= inst.size()) { rd_set_error("rd_tlas_load: instanceID %u out of range", I.instanceID); return RD_E_IO; }
        rd_instance_desc& d = desc[I.instanceID];
        memcpy(d.transform, I.transform, 64);
        d.SBTOffset = I.SBTOffset; d.customInstanceID = I.customInstanceID; d.bottomAccelStruct = built[I.bottomAccelStructOffset];
    }
    int rc = rd_tlas_build(p, desc.data(), (unsigned)desc.size(), out);
    if (rc) return rc;
    if (blasOut && blasCount) {
        *blasCount = (unsigned)list.size();
        *blasOut = (rd_blas**)malloc(sizeof(rd_blas*) * (list.size() ? list.size() : 1));
        for (size_t i = 0; i < list.size(); ++i) (*blasOut)[i] = list[i];
    }
    return RD_OK;
}

extern "C" int rd_tlas_load(rd_platform* p, const char* path, rd_handle* out)
{
    if (!path) { rd_set_error("rd_tlas_load: null path"); return RD_E_ARG; }
    FILE* f = fopen(path, "rb");
    if (!f) { rd_set_error("rd_tlas_load: cannot open %s", path); return RD_E_IO; }
    fseek(f, 0, SEEK_END); const long n = ftell(f); fseek(f, 0, SEEK_SET);
    std::vector<unsigned char> buf(n > 0 ? (size_t)n : 0);
    const size_t got = buf.empty() ? 0 : fread(buf.data(), 1, buf.size(), f);
    fclose(f);
    if (got != buf.size()) { rd_set_error("rd_tlas_load: short read on %s", path); return RD_E_IO; }   // radiance.cpp:458-459
    return rd_tlas_load_memory(p, buf.data(), buf.size(), out, 0, 0);
}

// Serialises a top-level structure into the reference blob. *sizeOut bytes are written to a malloc'ed
// buffer (*dataOut, caller frees with free()).
extern "C" int rd_tlas_save_memory(rd_platform* p, rd_handle tlas, void** dataOut, size_t* sizeOut)
{
    RdHandleObj* T = rd_obj(tlas);
    if (!p || !T || T->kind != RD_H_TLAS || !dataOut || !sizeOut) { rd_set_error("rd_tlas_save: bad handle"); return RD_E_ARG; }
    RD_CUDA(cudaSetDevice(p->device));
    RD_CUDA(cudaStreamSynchronize(p->stream));
    const unsigned ni = T->instanceCount;
    std::vector<RdInstance> inst(ni);
    if (ni) RD_CUDA(cudaMemcpy(inst.data(), T->dinst, (size_t)ni * sizeof(RdInstance), cudaMemcpyDeviceToHost));
    // one bottom blob per distinct triangle array
    std::map<const RdTri*, std::vector<unsigned char>> bottoms;
    std::vector<float> ilo(3 * (size_t)ni), ihi(3 * (size_t)ni);
    for (unsigned i = 0; i < ni; ++i) {
        const RdInstance& I = inst[i];
        std::vector<unsigned char>& blob = bottoms[I.tris];
        const uint32_t nt = I.flags;
        std::vector<RdTri> tris(nt);
        if (blob.empty()) {
            if (nt) RD_CUDA(cudaMemcpy(tris.data(), I.tris, (size_t)nt * sizeof(RdTri), cudaMemcpyDeviceToHost));
            std::vector<float> lo(3 * (size_t)nt), hi(3 * (size_t)nt);
            for (uint32_t t = 0; t < nt; ++t) {
                const float* v[3] = {&tris[t].v0.x, &tris[t].v1.x, &tris[t].v2.x};
                for (int a = 0; a < 3; ++a) {
                    lo[3 * t + a] = std::min(v[0][a], std::min(v[1][a], v[2][a]));
                    hi[3 * t + a] = std::max(v[0][a], std::max(v[1][a], v[2][a]));
                }
            }
            Bvh2 B; B.lo = lo.data(); B.hi = hi.data(); B.leafType = kTypeTri;
            std::vector<uint32_t> ids(nt);
            for (uint32_t t = 0; t < nt; ++t) ids[t] = t;
            B.build(ids, 0, nt);
            BlobBottom hdr;
            hdr.type = kBotAS; hdr.nodeByteOffset = sizeof(hdr);
            hdr.faceByteOffset = hdr.nodeByteOffset + (uint32_t)(B.nodes.size() * sizeof(BlobNode));
            hdr.vertexOffset = hdr.faceByteOffset + nt * (uint32_t)sizeof(BlobFace);
            blob.assign((size_t)hdr.vertexOffset + (size_t)nt * 3 * 16, 0);
            memcpy(blob.data(), &hdr, sizeof(hdr));
            memcpy(blob.data() + hdr.nodeByteOffset, B.nodes.data(), B.nodes.size() * sizeof(BlobNode));
            BlobFace* faces = (BlobFace*)(blob.data() + hdr.faceByteOffset);
            float* verts = (float*)(blob.data() + hdr.vertexOffset);
            for (uint32_t k = 0; k < nt; ++k) {
                const RdTri& tr = tris[B.order[k]];
                uint32_t prim; memcpy(&prim, &tr.v0.w, 4);
                // vertices are written per record (the native structure keeps no shared vertex array)
                faces[k].idx0 = 3 * k; faces[k].idx1 = 3 * k + 1; faces[k].idx2 = 3 * k + 2; faces[k].primID = prim;
                const float* v[3] = {&tr.v0.x, &tr.v1.x, &tr.v2.x};
                for (int c = 0; c < 3; ++c) { float* o = verts + 4 * (3 * (size_t)k + c); o[0] = v[c][0]; o[1] = v[c][1]; o[2] = v[c][2]; o[3] = 0.0f; }
            }
        }
        // world box of the instance from the root box of its bottom blob
        const BlobNode* root = (const BlobNode*)(blob.data() + sizeof(BlobBottom));
        RdF4 bl, bh, wl, wh;
        bl.x = root->lo[0]; bl.y = root->lo[1]; bl.z = root->lo[2]; bl.w = 0; bh.x = root->hi[0]; bh.y = root->hi[1]; bh.z = root->hi[2]; bh.w = 0;
        if (!nt) { bl.x = bl.y = bl.z = 0; bh = bl; }
        rd_instance_world_box(I.fwd, bl, bh, &wl, &wh);
        ilo[3 * i] = wl.x; ilo[3 * i + 1] = wl.y; ilo[3 * i + 2] = wl.z; ihi[3 * i] = wh.x; ihi[3 * i + 1] = wh.y; ihi[3 * i + 2] = wh.z;
    }
    Bvh2 B; B.lo = ilo.data(); B.hi = ihi.data(); B.leafType = kTypeInst;
    std::vector<uint32_t> ids(ni);
    for (unsigned i = 0; i < ni; ++i) ids[i] = i;
    B.build(ids, 0, ni);
    BlobTop top;
    top.type = kTopAS; top.nodeByteOffset = sizeof(top);
    top.instByteOffset = top.nodeByteOffset + (uint32_t)(B.nodes.size() * sizeof(BlobNode));
    size_t total = (size_t)top.instByteOffset + (size_t)ni * sizeof(BlobInstance);
    std::map<const RdTri*, uint32_t> offsetOf;
    for (auto& kv : bottoms) { offsetOf[kv.first] = (uint32_t)total; total += kv.second.size(); }
    if (total > 0xffffffffull) { rd_set_error("rd_tlas_save: %zu bytes do not fit the reference's 32-bit offsets (data.cl:92-96)", total); return RD_E_ARG; }
    top.totalBufferSize = (uint32_t)total;
    unsigned char* out = (unsigned char*)malloc(total);
    if (!out) { rd_set_error("rd_tlas_save: out of memory"); return RD_E_ARG; }
    memcpy(out, &top, sizeof(top));
    memcpy(out + top.nodeByteOffset, B.nodes.data(), B.nodes.size() * sizeof(BlobNode));
    BlobInstance* bi = (BlobInstance*)(out + top.instByteOffset);
    for (unsigned k = 0; k < ni; ++k) {
        const RdInstance& I = inst[B.order[k]];   // instance table in leaf order, instanceID = caller order
        memcpy(bi[k].transform, I.fwd, 64);
        bi[k].SBTOffset = I.sbtOffset; bi[k].instanceID = I.instanceID; bi[k].customInstanceID = I.customID;
        bi[k].bottomAccelStructOffset = offsetOf[I.tris];
    }
    for (auto& kv : bottoms) memcpy(out + offsetOf[kv.first], kv.second.data(), kv.second.size());
    *dataOut = out; *sizeOut = total;
    return RD_OK;
}

extern "C" int rd_tlas_save(rd_platform* p, rd_handle tlas, const char* path)
{
    if (!path) { rd_set_error("rd_tlas_save: null path"); return RD_E_ARG; }
    void* data = 0; size_t size = 0;
    int rc = rd_tlas_save_memory(p, tlas, &data, &size);
    if (rc) return rc;
    FILE* f = fopen(path, "wb");
    if (!f) { free(data); rd_set_error("rd_tlas_save: cannot open %s", path); return RD_E_IO; }
    const size_t put = fwrite(data, 1, size, f);
    fclose(f); free(data);
    if (put != size) { rd_set_error("rd_tlas_save: short write on %s", path); return RD_E_IO; }
    return RD_OK;
}

extern "C" void rd_free_host(void* ptr) { free(ptr); }
