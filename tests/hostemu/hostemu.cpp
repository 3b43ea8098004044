// tests/hostemu/hostemu.cpp -- CPU UNIT-TEST HARNESS, not part of the product.
//
// Runs the per-element build steps (csrc/rd_build.h) and the per-ray traversal (csrc/rd_traverse.h)
// serially on the host so the logic of the CUDA kernels can be unit-tested where there is no GPU
// (`pytest -m "not gpu"`). The product library (libradiance_b200.so) never links or loads this file
// and has no CPU path; kernels are launched by csrc/rd_build.cu / rd_trace.cu.
#include <algorithm>
#include <cstdlib>
#include <cstring>
#include <vector>
#include <cfloat>
#include "rd_build.h"

struct EmuBlas {
    std::vector<RdNode8> nodes;
    std::vector<RdTri> tris;
    RdF4 lo, hi;
    unsigned maxDepthItems = 0;
};

struct EmuBuildOut { std::vector<RdNode8> nodes; std::vector<rd_u32> leafPrim; unsigned levels; };

static void emu_build(const std::vector<RdF4>& plo, const std::vector<RdF4>& phi, EmuBuildOut& out, RdF4* rootLo, RdF4* rootHi)
{
    const rd_u32 n = (rd_u32)plo.size();
    out.nodes.clear(); out.leafPrim.clear(); out.levels = 0;
    if (n == 0) {
        RdNode8 e; memset(&e, 0, sizeof(e));
#if !RD_NODE_F32
        e.w[0].w = 1u | (1u << 8) | (1u << 16);
#endif
        out.nodes.push_back(e);
        rootLo->x = rootLo->y = rootLo->z = 0; rootHi->x = rootHi->y = rootHi->z = 0;
        return;
    }
    float cmin[3] = {FLT_MAX, FLT_MAX, FLT_MAX}, cmax[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX};
    for (rd_u32 i = 0; i < n; ++i) {
        float c[3] = {0.5f * (plo[i].x + phi[i].x), 0.5f * (plo[i].y + phi[i].y), 0.5f * (plo[i].z + phi[i].z)};
        for (int a = 0; a < 3; ++a) { cmin[a] = std::min(cmin[a], c[a]); cmax[a] = std::max(cmax[a], c[a]); }
    }
    float cinv[3];
    for (int a = 0; a < 3; ++a) cinv[a] = cmax[a] > cmin[a] ? 1.0f / (cmax[a] - cmin[a]) : 0.0f;
    std::vector<std::pair<rd_u64, rd_u32>> kv(n);
    for (rd_u32 i = 0; i < n; ++i)
        kv[i] = {rd_morton63(0.5f * (plo[i].x + phi[i].x), 0.5f * (plo[i].y + phi[i].y), 0.5f * (plo[i].z + phi[i].z), cmin, cinv), i};
    std::stable_sort(kv.begin(), kv.end(), [](auto& a, auto& b) { return a.first < b.first; });
    std::vector<rd_u64> keys(n); std::vector<rd_u32> prim(n);
    for (rd_u32 i = 0; i < n; ++i) { keys[i] = kv[i].first; prim[i] = kv[i].second; }
    const rd_u32 ni = n > 1 ? n - 1 : 1;
    std::vector<rd_u32> left(ni), right(ni), first(ni), last(ni), parent(2 * n), arrive(ni, 0), count(ni, 0);
    std::vector<RdF4> lo(2 * n), hi(2 * n);
    std::vector<float> cost(8 * (size_t)ni); std::vector<rd_u8> dec(8 * (size_t)ni);
    RdBinTree T; T.cost = cost.data(); T.dec = dec.data(); T.n = n; T.keys = keys.data(); T.prim = prim.data(); T.left = left.data(); T.right = right.data();
    T.first = first.data(); T.last = last.data(); T.parent = parent.data(); T.arrive = arrive.data(); T.count = count.data(); T.lo = lo.data(); T.hi = hi.data();
    for (rd_u32 i = 0; i < n; ++i) { lo[n - 1 + i] = plo[prim[i]]; hi[n - 1 + i] = phi[prim[i]]; }
    RdCollapseItem rootItem; rootItem.wide = 0;
    if (n == 1) { rootItem.bin = RD_BIN_LEAF | 0u; }
    else {
        for (rd_u32 i = 0; i + 1 < n; ++i) rd_karras_node(T, (int)i);
        for (rd_u32 i = 0; i < n; ++i) rd_refit_from_leaf(T, i);
        rootItem.bin = 0;
    }
    { rd_u32 s = rd_bin_slot(T, rootItem.bin); *rootLo = lo[s]; *rootHi = hi[s]; }
    out.nodes.resize(n + 1); out.leafPrim.resize(n);
    std::vector<RdCollapseItem> cur(1, rootItem), next(n + 1);
    rd_u32 counters[3] = {1, 0, 0};
    RdCollapseOut O; O.nodes = out.nodes.data(); O.leafPrim = out.leafPrim.data(); O.counters = counters; O.next = next.data();
    while (!cur.empty()) {
        counters[2] = 0;
        for (auto& it : cur) rd_collapse_item(T, O, it);
        cur.assign(next.begin(), next.begin() + counters[2]);
        out.levels++;
    }
    out.nodes.resize(counters[0]);
}

extern "C" {

void* emu_blas_build(const float* verts, unsigned nv, const unsigned* idx, unsigned nt)
{
    EmuBlas* b = new EmuBlas();
    std::vector<RdF4> plo(nt), phi(nt);
    for (unsigned p = 0; p < nt; ++p) rd_prim_box_tri(verts, idx, p, plo.data(), phi.data());
    EmuBuildOut o;
    emu_build(plo, phi, o, &b->lo, &b->hi);
    b->nodes = o.nodes; b->maxDepthItems = o.levels;
    b->tris.resize(nt);
    for (unsigned r = 0; r < nt; ++r) rd_gather_tri(verts, idx, o.leafPrim.data(), b->tris.data(), r);
    return b;
}
unsigned emu_blas_nodes(void* h) { return (unsigned)((EmuBlas*)h)->nodes.size(); }
unsigned emu_blas_levels(void* h) { return ((EmuBlas*)h)->maxDepthItems; }
void emu_blas_free(void* h) { delete (EmuBlas*)h; }

struct EmuInstanceIn { float transform[16]; unsigned sbt, custom; void* blas; };

struct EmuTlas {
    std::vector<RdNode8> nodes; std::vector<rd_u32> leafInst; std::vector<RdInstance> inst; RdScene scene; rd_u32 status[4];
};

void* emu_tlas_build(const EmuInstanceIn* in, unsigned n)
{
    EmuTlas* t = new EmuTlas();
    t->inst.resize(n);
    std::vector<RdF4> plo(n), phi(n);
    for (unsigned i = 0; i < n; ++i) {
        EmuBlas* b = (EmuBlas*)in[i].blas;
        RdInstance& I = t->inst[i];
        memcpy(I.fwd, in[i].transform, 64);
        float inv[16];
        if (!rd_inverse4x4(I.fwd, inv)) memset(inv, 0, sizeof(inv));
        memcpy(I.inv, inv, 48);
        I.sbtOffset = in[i].sbt; I.instanceID = i; I.customID = in[i].custom; I.flags = 0;
        I.maxAbs = fmaxf(fmaxf(fmaxf(fabsf(b->lo.x), fabsf(b->hi.x)), fmaxf(fabsf(b->lo.y), fabsf(b->hi.y))), fmaxf(fabsf(b->lo.z), fabsf(b->hi.z)));
        if (!(I.maxAbs < 3.0e38f)) I.maxAbs = 0.0f;
        I.nodes = b->nodes.data(); I.tris = b->tris.data();
        rd_instance_world_box(I.fwd, b->lo, b->hi, &plo[i], &phi[i]);
    }
    EmuBuildOut o; RdF4 rl, rh;
    emu_build(plo, phi, o, &rl, &rh);
    t->nodes = o.nodes; t->leafInst = o.leafPrim;
    memset(&t->scene, 0, sizeof(RdScene)); memset(t->status, 0, sizeof(t->status));
    t->scene.nodes = t->nodes.data(); t->scene.leafInstances = t->leafInst.data(); t->scene.instances = t->inst.data();
    t->scene.instanceCount = n; t->scene.nodeCount = (rd_u32)t->nodes.size(); t->scene.status = t->status;
    t->scene.maxAbs = fmaxf(fmaxf(fmaxf(fabsf(rl.x), fabsf(rh.x)), fmaxf(fabsf(rl.y), fabsf(rh.y))), fmaxf(fabsf(rl.z), fabsf(rh.z)));
    if (!(t->scene.maxAbs < 3.0e38f) || !n) t->scene.maxAbs = 0.0f;
    return t;
}
void emu_tlas_free(void* h) { delete (EmuTlas*)h; }
unsigned emu_tlas_status(void* h) { return ((EmuTlas*)h)->status[0]; }

// rays: n x 8 floats {o,tmin,d,tmax}; out: n x {t,b1,b2,prim,inst} (5 x 4 bytes)
void emu_trace(void* h, const float* rays, size_t n, int firstHit, void* out)
{
    EmuTlas* t = (EmuTlas*)h;
    RdHitRec* hits = (RdHitRec*)out;
    for (size_t i = 0; i < n; ++i) {
        const float* r = rays + 8 * i;
        if (firstHit) { RdFirstHit cb; rd_traverse(&t->scene, r[0], r[1], r[2], r[4], r[5], r[6], r[3], r[7], hits[i], cb); }
        else { RdNoAnyHit cb; rd_traverse(&t->scene, r[0], r[1], r[2], r[4], r[5], r[6], r[3], r[7], hits[i], cb); }
    }
}
int emu_inverse4x4(const float* m, float* out) { return rd_inverse4x4(m, out) ? 1 : 0; }
}
