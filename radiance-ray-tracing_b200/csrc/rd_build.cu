// rd_build.cu -- GPU acceleration-structure build: Morton codes -> radix sort (rd_sort.h) -> binary hierarchy by
// parallel locally-ordered clustering (PLOC; Karras' LBVH + refit kept as option) -> SAH-optimal
// level-synchronous collapse into 8-wide quantised nodes. Replaces the host SAH
// builder of the reference (radiance/src/bvh.cpp:46-597) and its blob packers
// (radiance/src/radiance.cpp:318-425). The per-element logic lives in rd_build.h.
#include <float.h>
#include <string.h>
#include <stdlib.h>
#include <mutex>
#include "rd_internal.h"
#include "rd_build.h"
#include "rd_sort.h"

#define RD_BT 256
static inline unsigned rd_grid(size_t n) { return (unsigned)((n + RD_BT - 1) / RD_BT); }

// order-preserving float <-> uint encoding for atomicMin/atomicMax
__device__ __forceinline__ unsigned rd_f2ord(float f) { unsigned u = __float_as_uint(f); return (u & 0x80000000u) ? ~u : (u | 0x80000000u); }
__host__ __device__ __forceinline__ float rd_ord2f(unsigned u)
{
    unsigned v = (u & 0x80000000u) ? (u & 0x7fffffffu) : ~u;
#ifdef __CUDA_ARCH__
    return __uint_as_float(v);
#else
    float f; memcpy(&f, &v, 4); return f;
#endif
}

__global__ void rd_k_tri_boxes(const float* verts, const rd_u32* idx, rd_u32 n, RdF4* lo, RdF4* hi)
{
    rd_u32 p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p < n) rd_prim_box_tri(verts, idx, p, lo, hi);
}

// bounds[0..2] = min centroid (ordered uint), bounds[3..5] = max centroid
__device__ __forceinline__ void rd_centroid_bounds_part(const RdF4* lo, const RdF4* hi, rd_u32 n, unsigned* bounds, rd_u32 first, rd_u32 stride)
{
    float mn[3] = {FLT_MAX, FLT_MAX, FLT_MAX}, mx[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX};
    for (rd_u32 i = first; i < n; i += stride) {
        const RdF4 l = lo[i], h = hi[i];
        const float c[3] = {0.5f * (l.x + h.x), 0.5f * (l.y + h.y), 0.5f * (l.z + h.z)};
        // a NaN / Inf coordinate (or a box whose centre overflows) is refused by the host: bounds[6] = count of such primitives
        if (!(fabsf(c[0]) < 3.0e38f && fabsf(c[1]) < 3.0e38f && fabsf(c[2]) < 3.0e38f &&
              fabsf(h.x - l.x) < 3.0e38f && fabsf(h.y - l.y) < 3.0e38f && fabsf(h.z - l.z) < 3.0e38f)) atomicAdd(&bounds[6], 1u);
#pragma unroll
        for (int a = 0; a < 3; ++a) { mn[a] = fminf(mn[a], c[a]); mx[a] = fmaxf(mx[a], c[a]); }
    }
#pragma unroll
    for (int a = 0; a < 3; ++a) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            mn[a] = fminf(mn[a], __shfl_xor_sync(0xffffffffu, mn[a], o));
            mx[a] = fmaxf(mx[a], __shfl_xor_sync(0xffffffffu, mx[a], o));
        }
    }
    if ((threadIdx.x & 31) == 0) {
#pragma unroll
        for (int a = 0; a < 3; ++a) { atomicMin(&bounds[a], rd_f2ord(mn[a])); atomicMax(&bounds[3 + a], rd_f2ord(mx[a])); }
    }
}

__global__ void rd_k_centroid_bounds(const RdF4* lo, const RdF4* hi, rd_u32 n, unsigned* bounds)
{
    rd_centroid_bounds_part(lo, hi, n, bounds, blockIdx.x * blockDim.x + threadIdx.x, gridDim.x * blockDim.x);
}

__device__ __forceinline__ void rd_morton_item(const RdF4* lo, const RdF4* hi, rd_u32 i, const unsigned* bounds, rd_u64* keys, rd_u32* vals)
{
    float cmin[3], cinv[3];
#pragma unroll
    for (int a = 0; a < 3; ++a) {
        cmin[a] = rd_ord2f(bounds[a]);
        const float ext = rd_ord2f(bounds[3 + a]) - cmin[a];
        cinv[a] = ext > 0.0f ? 1.0f / ext : 0.0f;
    }
    const RdF4 l = lo[i], h = hi[i];
    keys[i] = rd_morton63(0.5f * (l.x + h.x), 0.5f * (l.y + h.y), 0.5f * (l.z + h.z), cmin, cinv);
    vals[i] = i;
}

__global__ void rd_k_morton(const RdF4* lo, const RdF4* hi, rd_u32 n, const unsigned* bounds, rd_u64* keys, rd_u32* vals)
{
    rd_u32 i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) rd_morton_item(lo, hi, i, bounds, keys, vals);
}

__global__ void rd_k_leaf_boxes(RdBinTree T, const RdF4* plo, const RdF4* phi)
{
    rd_u32 i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= T.n) return;
    const rd_u32 p = T.prim[i];
    T.lo[T.n - 1 + i] = plo[p];
    T.hi[T.n - 1 + i] = phi[p];
}

__global__ void rd_k_karras(RdBinTree T)
{
    rd_u32 i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i + 1 < T.n) rd_karras_node(T, (int)i);
}

__global__ void rd_k_refit(RdBinTree T)
{
    rd_u32 i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < T.n) rd_refit_from_leaf(T, i);
}


// ---- PLOC: parallel locally-ordered clustering (Meister & Bittner 2018) ---------------------------
// The clusters start as the primitives in Morton order. Every round each cluster looks at its
// RD_PLOC radius neighbours on either side in the array and picks the one with which its merged box
// has the smallest surface area; pairs that pick each other merge into a new inner node that takes the
// left partner's place; the array is compacted in order. Unlike the Karras tree, whose topology is
// fixed by the Morton bits, merges are chosen by surface area, which is what the SAH counts.
// Inner node ids are handed out downwards from n-2 so that the last merge is node 0, the root.
// Boxes, primitive counts and the collapse DP costs of a new node are final when it is created (its
// children are older), so no refit pass is needed.
struct RdPlocSide { rd_u32* ref; RdF4* lo; RdF4* hi; };

__global__ void rd_k_ploc_init(RdBinTree T, RdPlocSide A)
{
    rd_u32 i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= T.n) return;
    A.ref[i] = RD_BIN_LEAF | i;
    A.lo[i] = T.lo[T.n - 1 + i];
    A.hi[i] = T.hi[T.n - 1 + i];
}

__global__ void rd_k_ploc_nearest(RdPlocSide A, rd_u32 m, int radius, rd_u32* nearest)
{
    const rd_u32 i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= m) return;
    const RdF4 l = A.lo[i], h = A.hi[i];
    const int j0 = (int)i - radius < 0 ? 0 : (int)i - radius;
    const int j1 = (int)i + radius >= (int)m ? (int)m - 1 : (int)i + radius;
    float best = __int_as_float(0x7f800000); rd_u32 bj = i;
    for (int j = j0; j <= j1; ++j) {
        if (j == (int)i) continue;
        const RdF4 a = A.lo[j], b = A.hi[j];
        const float ex = fmaxf(h.x, b.x) - fminf(l.x, a.x), ey = fmaxf(h.y, b.y) - fminf(l.y, a.y),
                    ez = fmaxf(h.z, b.z) - fminf(l.z, a.z);
        float area = ex * ey + ey * ez + ez * ex;
        // NaN / Inf / overflowing areas (a non-finite vertex) stay comparable, so the pair with the smallest (area, i, j)
        // is always mutual and every round merges at least once -- the clustering terminates on any input
        area = area < 3.0e38f ? area : 3.0e38f;
        // the same pair must score the same from both ends and ties must resolve the same way for a merge to be
        // mutual: area is symmetric, ties go to the lower index
        if (area < best) { best = area; bj = (rd_u32)j; }
    }
    nearest[i] = bj;
}

__global__ void rd_k_ploc_merge(RdBinTree T, RdPlocSide A, rd_u32 m, const rd_u32* nearest, rd_u32* nextId, rd_u32* keep)
{
    const rd_u32 i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= m) return;
    const rd_u32 j = nearest[i];
    if (j == i || nearest[j] != i) { keep[i] = 1u; return; }
    if (j < i) { keep[i] = 0u; return; }          // the right partner disappears
    const rd_u32 id = atomicSub(nextId, 1u) - 1u;
    const rd_u32 L = A.ref[i], R = A.ref[j];
    const RdF4 la = A.lo[i], ha = A.hi[i], lb = A.lo[j], hb = A.hi[j];
    RdF4 l, h;
    l.x = fminf(la.x, lb.x); l.y = fminf(la.y, lb.y); l.z = fminf(la.z, lb.z); l.w = 0.0f;
    h.x = fmaxf(ha.x, hb.x); h.y = fmaxf(ha.y, hb.y); h.z = fmaxf(ha.z, hb.z); h.w = 0.0f;
    T.left[id] = L; T.right[id] = R;
    T.parent[rd_bin_slot(T, L)] = id; T.parent[rd_bin_slot(T, R)] = id;
    if (id == 0u) T.parent[0] = RD_INVALID;
    T.lo[id] = l; T.hi[id] = h;
    T.count[id] = ((L & RD_BIN_LEAF) ? 1u : T.count[L]) + ((R & RD_BIN_LEAF) ? 1u : T.count[R]);
    if (T.cost) rd_dp_node(T, id, l, h);
    A.ref[i] = id; A.lo[i] = l; A.hi[i] = h;
    keep[i] = 1u;
}

__global__ void rd_k_ploc_compact(RdPlocSide A, RdPlocSide B, rd_u32 m, const rd_u32* keep, const rd_u32* pos)
{
    const rd_u32 i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= m) return;
    if (keep[i]) { const rd_u32 k = pos[i]; B.ref[k] = A.ref[i]; B.lo[k] = A.lo[i]; B.hi[k] = A.hi[i]; }
}

// Small inputs (the 4096 patch meshes of C4 have 2450 triangles each): all rounds in ONE block, the cluster arrays in
// global memory, rounds separated by __syncthreads -- no launch / host round trip per round.
#define RD_PLOC_SMALL_MAX 8192u
__device__ __forceinline__ void rd_ploc_small_block(RdBinTree T, RdPlocSide A, RdPlocSide B, rd_u32* nearest, int radius)
{
    __shared__ rd_u32 warpSum[32];
    __shared__ rd_u32 sBase, sNext;
    const rd_u32 tid = threadIdx.x, lane = tid & 31u, warp = tid >> 5;
    for (rd_u32 i = tid; i < T.n; i += blockDim.x) { A.ref[i] = RD_BIN_LEAF | i; A.lo[i] = T.lo[T.n - 1 + i]; A.hi[i] = T.hi[T.n - 1 + i]; }
    if (tid == 0) sNext = T.n - 1u;
    __syncthreads();
    rd_u32 m = T.n;
    while (m > 1u) {
        for (rd_u32 i = tid; i < m; i += blockDim.x) {
            const RdF4 l = A.lo[i], h = A.hi[i];
            const int j0 = (int)i - radius < 0 ? 0 : (int)i - radius;
            const int j1 = (int)i + radius >= (int)m ? (int)m - 1 : (int)i + radius;
            float best = __int_as_float(0x7f800000); rd_u32 bj = i;
            for (int j = j0; j <= j1; ++j) {
                if (j == (int)i) continue;
                const RdF4 a = A.lo[j], b = A.hi[j];
                const float ex = fmaxf(h.x, b.x) - fminf(l.x, a.x), ey = fmaxf(h.y, b.y) - fminf(l.y, a.y),
                            ez = fmaxf(h.z, b.z) - fminf(l.z, a.z);
                float area = ex * ey + ey * ez + ez * ex;
                area = area < 3.0e38f ? area : 3.0e38f;   // NaN / Inf / overflow: still comparable, so a round always has a mutual pair
                if (area < best) { best = area; bj = (rd_u32)j; }
            }
            nearest[i] = bj;
        }
        if (tid == 0) sBase = 0u;
        __syncthreads();
        for (rd_u32 c = 0; c < m; c += blockDim.x) {
            const rd_u32 i = c + tid;
            rd_u32 keep = 0u, ref = 0u;
            RdF4 l, h;
            l.x = l.y = l.z = l.w = 0.0f; h = l;
            if (i < m) {
                const rd_u32 j = nearest[i];
                if (j == i || nearest[j] != i) { keep = 1u; ref = A.ref[i]; l = A.lo[i]; h = A.hi[i]; }
                else if (j > i) {
                    const rd_u32 id = atomicSub(&sNext, 1u) - 1u;
                    const rd_u32 L = A.ref[i], R = A.ref[j];
                    const RdF4 la = A.lo[i], ha = A.hi[i], lb = A.lo[j], hb = A.hi[j];
                    l.x = fminf(la.x, lb.x); l.y = fminf(la.y, lb.y); l.z = fminf(la.z, lb.z);
                    h.x = fmaxf(ha.x, hb.x); h.y = fmaxf(ha.y, hb.y); h.z = fmaxf(ha.z, hb.z);
                    T.left[id] = L; T.right[id] = R;
                    T.parent[rd_bin_slot(T, L)] = id; T.parent[rd_bin_slot(T, R)] = id;
                    if (id == 0u) T.parent[0] = RD_INVALID;
                    T.lo[id] = l; T.hi[id] = h;
                    const volatile rd_u32* vc = (const volatile rd_u32*)T.count;
                    T.count[id] = ((L & RD_BIN_LEAF) ? 1u : vc[L]) + ((R & RD_BIN_LEAF) ? 1u : vc[R]);
                    if (T.cost) rd_dp_node(T, id, l, h);
                    keep = 1u; ref = id;
                }
            }
            // ordered compaction: block-wide exclusive scan of keep
            rd_u32 x = keep;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) { const rd_u32 y = __shfl_up_sync(0xffffffffu, x, o); if (lane >= (rd_u32)o) x += y; }
            if (lane == 31u) warpSum[warp] = x;
            __syncthreads();
            if (warp == 0) {
                rd_u32 s = warpSum[lane];
#pragma unroll
                for (int o = 1; o < 32; o <<= 1) { const rd_u32 y = __shfl_up_sync(0xffffffffu, s, o); if (lane >= (rd_u32)o) s += y; }
                warpSum[lane] = s;
            }
            __syncthreads();
            const rd_u32 pos = sBase + (x - keep) + (warp ? warpSum[warp - 1] : 0u);
            if (keep) { B.ref[pos] = ref; B.lo[pos] = l; B.hi[pos] = h; }
            __syncthreads();
            if (tid == 0) sBase += warpSum[31];
            __syncthreads();
        }
        m = sBase;
        RdPlocSide t = A; A = B; B = t;
        __syncthreads();
    }
}

__global__ void __launch_bounds__(1024) rd_k_ploc_small(RdBinTree T, RdPlocSide A, RdPlocSide B, rd_u32* nearest, int radius)
{
    rd_ploc_small_block(T, A, B, nearest, radius);
}

__global__ void rd_k_collapse(RdBinTree T, RdCollapseOut O, const RdCollapseItem* items, rd_u32 count)
{
    rd_u32 i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < count) rd_collapse_item(T, O, items[i]);
}

__global__ void rd_k_gather_tris(const float* verts, const rd_u32* idx, const rd_u32* leafPrim, RdTri* out, rd_u32 n)
{
    rd_u32 r = blockIdx.x * blockDim.x + threadIdx.x;
    if (r < n) rd_gather_tri(verts, idx, leafPrim, out, r);
}

struct RdTmp {   // scoped device scratch, stream-ordered: freed blocks go back to the device's memory pool
                 // (release threshold raised in rd_platform_create), so a second build -- or the 4096 bottom-level
                 // builds of C4 -- reuses them without a cudaMalloc / synchronising cudaFree per buffer
    std::vector<void*> ptrs;
    cudaStream_t st;
    explicit RdTmp(cudaStream_t s) : st(s) {}
    ~RdTmp() { for (void* q : ptrs) cudaFreeAsync(q, st); }
    template <class T> cudaError_t alloc(T** out, size_t count)
    {
        void* q = 0;
        cudaError_t e = cudaMallocAsync(&q, count * sizeof(T) ? count * sizeof(T) : 16, st);
        if (e == cudaSuccess) ptrs.push_back(q);
        *out = (T*)q;
        return e;
    }
};

// ---- small inputs (<= RD_PLOC_SMALL_MAX primitives): the whole build without a host round trip ---------------------
// A scene of many small meshes (C4: 4096 patch meshes of 2450 triangles) pays the general path's fixed costs per mesh:
// ~35 scratch allocations, a host synchronisation per collapse level, per sort and per clustering step, synchronous
// cudaMalloc / cudaFree of the results -- 0.77 ms per mesh, 3.1 s for the scene. Here: ONE scratch allocation, eight
// launches (every multi-round step runs its rounds inside one block), ONE synchronisation to learn the node count.
__global__ void rd_k_small_init(unsigned* bounds, rd_u32* counters)
{
    if (threadIdx.x < 3) bounds[threadIdx.x] = 0xffffffffu;
    else if (threadIdx.x < 8) bounds[threadIdx.x] = 0u;
    if (threadIdx.x == 0) { counters[0] = 1u; counters[1] = 0u; counters[2] = 0u; counters[3] = 0u; }
}

// the one-block sort leaves its result on side *where; the build wants it in (keysB, valsB)
__global__ void rd_k_small_settle(const rd_u32* where, const rd_u64* keysA, const rd_u32* valsA, rd_u64* keysB, rd_u32* valsB, rd_u32 n)
{
    if (*where != 0u) return;
    for (rd_u32 i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) { keysB[i] = keysA[i]; valsB[i] = valsA[i]; }
}

// all levels of the wide-node collapse in one block; summary = { nodes, leaf records, levels, non-finite boxes, root lo.xyz, root hi.xyz }
__device__ __forceinline__ void rd_collapse_small_block(RdBinTree T, RdCollapseOut O, RdCollapseItem* itemsA, RdCollapseItem* itemsB,
                                                        RdCollapseItem rootItem, const unsigned* bounds, rd_u32* summary)
{
    __shared__ rd_u32 sCount;
    RdCollapseItem* cur = itemsA;
    // non-finite input (flagged by rd_k_centroid_bounds): the build is refused by the host, nothing is collapsed
    if (threadIdx.x == 0) { itemsA[0] = rootItem; sCount = bounds[6] ? 0u : 1u; }
    O.next = itemsB;
    // ... but the leaf-record gather that follows on the stream still reads leafPrim: give it valid primitive numbers
    if (bounds[6]) for (rd_u32 i = threadIdx.x; i < T.n; i += blockDim.x) O.leafPrim[i] = i;
    __syncthreads();
    rd_u32 levels = 0;
    for (;;) {
        const rd_u32 count = sCount;
        if (!count || levels > 200u) break;
        __syncthreads();
        if (threadIdx.x == 0) O.counters[2] = 0u;
        __syncthreads();
        for (rd_u32 i = threadIdx.x; i < count; i += blockDim.x) rd_collapse_item(T, O, cur[i]);
        __syncthreads();
        if (threadIdx.x == 0) sCount = O.counters[2];
        RdCollapseItem* t = cur; cur = O.next; O.next = t;
        ++levels;
        __syncthreads();
    }
    if (threadIdx.x == 0) {
        const rd_u32 rootSlot = 0u;   // inner root = slot 0; single leaf = slot n - 1 = 0
        summary[0] = O.counters[0]; summary[1] = O.counters[1]; summary[2] = levels; summary[3] = bounds[6];
        summary[4] = RD_F2U(T.lo[rootSlot].x); summary[5] = RD_F2U(T.lo[rootSlot].y); summary[6] = RD_F2U(T.lo[rootSlot].z);
        summary[7] = RD_F2U(T.hi[rootSlot].x); summary[8] = RD_F2U(T.hi[rootSlot].y); summary[9] = RD_F2U(T.hi[rootSlot].z);
    }
}

__global__ void __launch_bounds__(256) rd_k_collapse_small(RdBinTree T, RdCollapseOut O, RdCollapseItem* itemsA, RdCollapseItem* itemsB,
                                                          RdCollapseItem rootItem, const unsigned* bounds, rd_u32* summary)
{
    rd_collapse_small_block(T, O, itemsA, itemsB, rootItem, bounds, summary);
}

// deferredSummary != 0: nothing waits for the device -- the nodes go straight into an allocation of the worst-case size
// (n + 1) and the summary lands in host-mapped memory, to be read by rd_blas_resolve once the build's event has fired.
static int rd_build_wide_bvh_small(rd_platform* p, cudaStream_t st, const RdF4* dlo, const RdF4* dhi, uint32_t n, RdBuildResult* out,
                                   rd_u32* deferredSummary = 0)
{
    const rd_u32 ni = n > 1 ? n - 1 : 1;
    // one scratch block, carved
    size_t off = 0;
    auto carve = [&off](size_t bytes) { const size_t at = off; off += (bytes + 255u) & ~(size_t)255u; return at; };
    const size_t oBounds = carve(8 * 4), oCounters = carve(4 * 4), oSummary = carve(16 * 4), oWhere = carve(4);
    const size_t oKeysA = carve((size_t)n * 8), oKeysB = carve((size_t)n * 8), oValsA = carve((size_t)n * 4), oValsB = carve((size_t)n * 4);
    const size_t oLeft = carve((size_t)ni * 4), oRight = carve((size_t)ni * 4), oFirst = carve((size_t)ni * 4), oLast = carve((size_t)ni * 4);
    const size_t oParent = carve(2 * (size_t)n * 4), oCount = carve((size_t)ni * 4);
    const size_t oLo = carve(2 * (size_t)n * 16), oHi = carve(2 * (size_t)n * 16), oCost = carve(8 * (size_t)ni * 4), oDec = carve(8 * (size_t)ni);
    const size_t oARef = carve((size_t)n * 4), oALo = carve((size_t)n * 16), oAHi = carve((size_t)n * 16);
    const size_t oBRef = carve((size_t)n * 4), oBLo = carve((size_t)n * 16), oBHi = carve((size_t)n * 16), oNearest = carve((size_t)n * 4);
    const size_t oWide = carve(deferredSummary ? 16 : ((size_t)n + 1) * sizeof(RdNode8)), oItemsA = carve(((size_t)n + 1) * sizeof(RdCollapseItem)),
                 oItemsB = carve(((size_t)n + 1) * sizeof(RdCollapseItem));
    RdTmp tmp(st);
    char* base;
    RD_CUDA(tmp.alloc(&base, off));
    unsigned* dbounds = (unsigned*)(base + oBounds);
    rd_u32* counters = (rd_u32*)(base + oCounters); rd_u32* summary = (rd_u32*)(base + oSummary); rd_u32* where = (rd_u32*)(base + oWhere);
    rd_u64 *keysA = (rd_u64*)(base + oKeysA), *keysB = (rd_u64*)(base + oKeysB);
    rd_u32 *valsA = (rd_u32*)(base + oValsA), *valsB = (rd_u32*)(base + oValsB);
    RdBinTree T;
    T.n = n; T.keys = keysB; T.prim = valsB;
    T.left = (rd_u32*)(base + oLeft); T.right = (rd_u32*)(base + oRight); T.first = (rd_u32*)(base + oFirst); T.last = (rd_u32*)(base + oLast);
    T.parent = (rd_u32*)(base + oParent); T.arrive = 0; T.count = (rd_u32*)(base + oCount);
    T.lo = (RdF4*)(base + oLo); T.hi = (RdF4*)(base + oHi); T.cost = (float*)(base + oCost); T.dec = (rd_u8*)(base + oDec);
    RdPlocSide A, B;
    A.ref = (rd_u32*)(base + oARef); A.lo = (RdF4*)(base + oALo); A.hi = (RdF4*)(base + oAHi);
    B.ref = (rd_u32*)(base + oBRef); B.lo = (RdF4*)(base + oBLo); B.hi = (RdF4*)(base + oBHi);
    RD_CUDA(cudaMallocAsync(&out->leafPrim, (size_t)n * sizeof(rd_u32), st));
    int radius = 32;
    if (const char* e = getenv("RD_PLOC_RADIUS")) { radius = atoi(e); if (radius < 1) radius = 1; }
    rd_k_small_init<<<1, 32, 0, st>>>(dbounds, counters);
    rd_k_centroid_bounds<<<rd_grid(n), RD_BT, 0, st>>>(dlo, dhi, n, dbounds);
    rd_k_morton<<<rd_grid(n), RD_BT, 0, st>>>(dlo, dhi, n, dbounds, keysA, valsA);
    if (n > 1) {
        rd_k_sort_small<<<1, 1024, 0, st>>>(keysA, valsA, keysB, valsB, n, 8, where);
        rd_k_small_settle<<<rd_grid(n), RD_BT, 0, st>>>(where, keysA, valsA, keysB, valsB, n);
    } else {
        RD_CUDA(cudaMemcpyAsync(keysB, keysA, 8, cudaMemcpyDeviceToDevice, st));
        RD_CUDA(cudaMemcpyAsync(valsB, valsA, 4, cudaMemcpyDeviceToDevice, st));
    }
    rd_k_leaf_boxes<<<rd_grid(n), RD_BT, 0, st>>>(T, dlo, dhi);
    if (n > 1) rd_k_ploc_small<<<1, 1024, 0, st>>>(T, A, B, (rd_u32*)(base + oNearest), radius);
    RdCollapseItem rootItem; rootItem.wide = 0; rootItem.bin = (n == 1) ? (RD_BIN_LEAF | 0u) : 0u;
    RdCollapseOut O; O.nodes = (RdNode8*)(base + oWide); O.leafPrim = out->leafPrim; O.counters = counters; O.next = 0;
    if (deferredSummary) {
        RD_CUDA(cudaMallocAsync(&out->nodes, ((size_t)n + 1) * sizeof(RdNode8), st));
        O.nodes = out->nodes; summary = deferredSummary;
    }
    rd_k_collapse_small<<<1, 256, 0, st>>>(T, O, (RdCollapseItem*)(base + oItemsA), (RdCollapseItem*)(base + oItemsB), rootItem, dbounds, summary);
    p->launches += n > 1 ? 8 : 5;
    if (deferredSummary) { RD_CUDA(cudaGetLastError()); out->nodeCount = 0; out->levels = 0; return RD_OK; }
    rd_u32 hs[10];
    RD_CUDA(cudaMemcpyAsync(hs, summary, sizeof(hs), cudaMemcpyDeviceToHost, st));
    RD_CUDA(cudaStreamSynchronize(st));
    RD_CUDA(cudaGetLastError());
    if (hs[3]) { cudaFreeAsync(out->leafPrim, st); out->leafPrim = 0; rd_set_error("BuildAccelStruct: %u primitive(s) with NaN / Inf coordinates", hs[3]); return RD_E_ARG; }
    if (hs[1] != n || hs[2] > 200u) { cudaFreeAsync(out->leafPrim, st); out->leafPrim = 0; rd_set_error("rd_build_wide_bvh: %u leaf records for %u primitives", hs[1], n); return RD_E_CUDA; }
    out->nodeCount = hs[0]; out->levels = hs[2];
    RD_CUDA(cudaMallocAsync(&out->nodes, (size_t)out->nodeCount * sizeof(RdNode8), st));
    RD_CUDA(cudaMemcpyAsync(out->nodes, O.nodes, (size_t)out->nodeCount * sizeof(RdNode8), cudaMemcpyDeviceToDevice, st));
    for (int k = 0; k < 3; ++k) { memcpy(&out->lo[k], &hs[4 + k], 4); memcpy(&out->hi[k], &hs[7 + k], 4); }
    return RD_OK;
}

int rd_build_wide_bvh(rd_platform* p, const RdF4* dlo, const RdF4* dhi, uint32_t n, RdBuildResult* out)
{
    cudaStream_t st = p->stream;
    memset(out, 0, sizeof(*out));
    if (n == 0) {
        // empty structure: one node without children
        RdNode8 e; memset(&e, 0, sizeof(e));
#if !RD_NODE_F32
        e.w[0].w = 1u | (1u << 8) | (1u << 16);
#endif
        RD_CUDA(cudaMallocAsync(&out->nodes, sizeof(RdNode8), st));
        RD_CUDA(cudaMemcpyAsync(out->nodes, &e, sizeof(e), cudaMemcpyHostToDevice, st));
        RD_CUDA(cudaMallocAsync(&out->leafPrim, 16, st));
        RD_CUDA(cudaStreamSynchronize(st));
        out->nodeCount = 1;
        return RD_OK;
    }
    if (n <= RD_PLOC_SMALL_MAX && n <= RD_SORT_SMALL_MAX && !getenv("RD_BUILD_GREEDY") && !getenv("RD_BUILD_LBVH") && !getenv("RD_BUILD_GENERAL"))
        return rd_build_wide_bvh_small(p, p->stream, dlo, dhi, n, out);
    RdTmp tmp(st);
    unsigned* dbounds; rd_u64 *keysA, *keysB; rd_u32 *valsA, *valsB;
    RD_CUDA(tmp.alloc(&dbounds, 8));
    RD_CUDA(tmp.alloc(&keysA, n)); RD_CUDA(tmp.alloc(&keysB, n));
    RD_CUDA(tmp.alloc(&valsA, n)); RD_CUDA(tmp.alloc(&valsB, n));
    const unsigned init[8] = {0xffffffffu, 0xffffffffu, 0xffffffffu, 0u, 0u, 0u, 0u, 0u};   // [6] = non-finite primitive boxes
    RD_CUDA(cudaMemcpyAsync(dbounds, init, sizeof(init), cudaMemcpyHostToDevice, st));
    unsigned rb = rd_grid(n); if (rb > (unsigned)p->smCount * 8) rb = p->smCount * 8;
    rd_k_centroid_bounds<<<rb, RD_BT, 0, st>>>(dlo, dhi, n, dbounds);
    rd_k_morton<<<rd_grid(n), RD_BT, 0, st>>>(dlo, dhi, n, dbounds, keysA, valsA);
    p->launches += 2;
    unsigned nonFinite = 0;   // device -> pageable host: returns once the value has arrived
    RD_CUDA(cudaMemcpyAsync(&nonFinite, dbounds + 6, sizeof(unsigned), cudaMemcpyDeviceToHost, st));
    RD_CUDA(cudaStreamSynchronize(st));
    if (nonFinite) { rd_set_error("BuildAccelStruct: %u primitive(s) with NaN / Inf coordinates", nonFinite); return RD_E_ARG; }
    RdSortScratch sortScratch;
    const size_t tableCount = 256u * (size_t)rd_sort_blocks(n);
    RD_CUDA(tmp.alloc(&sortScratch.hist, 8 * 256)); RD_CUDA(tmp.alloc(&sortScratch.table, tableCount));
    RD_CUDA(tmp.alloc(&sortScratch.scan, rd_scan_scratch_count(tableCount))); RD_CUDA(tmp.alloc(&sortScratch.where, 4));
    unsigned long long sortLaunches = 0;
    const int sortedSide = rd_radix_sort_pairs(st, keysA, valsA, keysB, valsB, n, 63, sortScratch, p->smCount, &sortLaunches);
    p->launches += sortLaunches;
    if (sortedSide < 0) { rd_set_error("rd_build_wide_bvh: radix sort failed: %s", cudaGetErrorString(cudaGetLastError())); return RD_E_CUDA; }
    if (sortedSide == 0) { rd_u64* tk = keysA; keysA = keysB; keysB = tk; rd_u32* tv = valsA; valsA = valsB; valsB = tv; }   // sorted pairs are (keysB, valsB) below

    const rd_u32 ni = n > 1 ? n - 1 : 1;
    RdBinTree T;
    T.n = n; T.keys = keysB; T.prim = valsB;
    RD_CUDA(tmp.alloc(&T.left, ni)); RD_CUDA(tmp.alloc(&T.right, ni));
    RD_CUDA(tmp.alloc(&T.first, ni)); RD_CUDA(tmp.alloc(&T.last, ni));
    RD_CUDA(tmp.alloc(&T.parent, 2 * (size_t)n)); RD_CUDA(tmp.alloc(&T.arrive, ni)); RD_CUDA(tmp.alloc(&T.count, ni));
    RD_CUDA(tmp.alloc(&T.lo, 2 * (size_t)n)); RD_CUDA(tmp.alloc(&T.hi, 2 * (size_t)n));
    T.cost = 0; T.dec = 0;
    if (!getenv("RD_BUILD_GREEDY")) {   // developer knob: greedy collapse instead of the SAH dynamic program
        RD_CUDA(tmp.alloc(&T.cost, 8 * (size_t)ni)); RD_CUDA(tmp.alloc(&T.dec, 8 * (size_t)ni));
    }
    RD_CUDA(cudaMemsetAsync(T.arrive, 0, ni * sizeof(rd_u32), st));
    rd_k_leaf_boxes<<<rd_grid(n), RD_BT, 0, st>>>(T, dlo, dhi);
    p->launches += 1;
    RdCollapseItem rootItem; rootItem.wide = 0; rootItem.bin = (n == 1) ? (RD_BIN_LEAF | 0u) : 0u;
    if (n > 1 && getenv("RD_BUILD_LBVH")) {   // developer knob: Karras' LBVH topology instead of PLOC
        rd_k_karras<<<rd_grid(n - 1), RD_BT, 0, st>>>(T);
        rd_k_refit<<<rd_grid(n), RD_BT, 0, st>>>(T);
        p->launches += 2;
    } else if (n > 1) {
        int radius = 32;   // 8 / 16 / 32 / 64 on C2: 4346 / 4345 / 4365 / 4272 Mrays/s, slowest tile row 0.257 / 0.260 / 0.226 / 0.228 ms
        if (const char* e = getenv("RD_PLOC_RADIUS")) { radius = atoi(e); if (radius < 1) radius = 1; }
        RdPlocSide A, B;
        RD_CUDA(tmp.alloc(&A.ref, n)); RD_CUDA(tmp.alloc(&A.lo, n)); RD_CUDA(tmp.alloc(&A.hi, n));
        RD_CUDA(tmp.alloc(&B.ref, n)); RD_CUDA(tmp.alloc(&B.lo, n)); RD_CUDA(tmp.alloc(&B.hi, n));
        rd_u32 *nearest, *keep, *pos, *pc;
        RD_CUDA(tmp.alloc(&nearest, n));
        if (n <= RD_PLOC_SMALL_MAX) {
            rd_k_ploc_small<<<1, 1024, 0, st>>>(T, A, B, nearest, radius);
            p->launches += 1;
        } else {
        RD_CUDA(tmp.alloc(&keep, n)); RD_CUDA(tmp.alloc(&pos, n)); RD_CUDA(tmp.alloc(&pc, 4));
        rd_u32* scanScratch;
        RD_CUDA(tmp.alloc(&scanScratch, rd_scan_scratch_count(n)));
        const rd_u32 pc0[2] = {n - 1u, n};   // [0] next inner id + 1 (ids are handed out downwards), [1] clusters
        RD_CUDA(cudaMemcpyAsync(pc, pc0, sizeof(pc0), cudaMemcpyHostToDevice, st));
        rd_k_ploc_init<<<rd_grid(n), RD_BT, 0, st>>>(T, A);
        p->launches += 1;
        rd_u32 m = n, rounds = 0;
        while (m > 1) {
            rd_k_ploc_nearest<<<rd_grid(m), RD_BT, 0, st>>>(A, m, radius, nearest);
            rd_k_ploc_merge<<<rd_grid(m), RD_BT, 0, st>>>(T, A, m, nearest, pc, keep);
            unsigned long long scanLaunches = 0;
            rd_exclusive_scan_u32(st, keep, pos, m, scanScratch, pc + 1, &scanLaunches);   // pc[1] = clusters that remain
            rd_k_ploc_compact<<<rd_grid(m), RD_BT, 0, st>>>(A, B, m, keep, pos);
            p->launches += 3 + scanLaunches;
            rd_u32 next = 0;
            RD_CUDA(cudaMemcpyAsync(&next, pc + 1, sizeof(next), cudaMemcpyDeviceToHost, st));
            RD_CUDA(cudaStreamSynchronize(st));
            if (next >= m || next == 0 || ++rounds > 100000u) { rd_set_error("rd_build_wide_bvh: clustering made no progress (%u -> %u clusters)", m, next); return RD_E_CUDA; }
            m = next;
            RdPlocSide t = A; A = B; B = t;
        }
        }
    }
    // collapse, one launch per level of the wide tree
    RdNode8* wide; rd_u32* counters; RdCollapseItem *itemsA, *itemsB;
    const size_t maxNodes = (size_t)n / 2 + 2;   // every wide node below the root has >= 4 primitives... see DESIGN.md
    RD_CUDA(tmp.alloc(&wide, (size_t)n + 1));
    (void)maxNodes;
    RD_CUDA(tmp.alloc(&counters, 4));
    RD_CUDA(tmp.alloc(&itemsA, (size_t)n + 1)); RD_CUDA(tmp.alloc(&itemsB, (size_t)n + 1));
    RD_CUDA(cudaMallocAsync(&out->leafPrim, (size_t)n * sizeof(rd_u32), st));
    const rd_u32 c0[4] = {1u, 0u, 0u, 0u};
    RD_CUDA(cudaMemcpyAsync(counters, c0, sizeof(c0), cudaMemcpyHostToDevice, st));
    RD_CUDA(cudaMemcpyAsync(itemsA, &rootItem, sizeof(rootItem), cudaMemcpyHostToDevice, st));
    RdCollapseOut O; O.nodes = wide; O.leafPrim = out->leafPrim; O.counters = counters; O.next = itemsB;
    rd_u32 count = 1, levels = 0;
    RdCollapseItem* cur = itemsA;
    rd_u32 hc[4];
    while (count) {
        RD_CUDA(cudaMemsetAsync(counters + 2, 0, sizeof(rd_u32), st));
        rd_k_collapse<<<rd_grid(count), RD_BT, 0, st>>>(T, O, cur, count);
        p->launches += 1;
        RD_CUDA(cudaMemcpyAsync(hc, counters, sizeof(hc), cudaMemcpyDeviceToHost, st));
        RD_CUDA(cudaStreamSynchronize(st));
        count = hc[2];
        RdCollapseItem* t = cur; cur = O.next; O.next = t;
        if (++levels > 200) { rd_set_error("rd_build_wide_bvh: collapse did not terminate"); return RD_E_CUDA; }
    }
    out->nodeCount = hc[0]; out->levels = levels;
    if (hc[1] != n) { rd_set_error("rd_build_wide_bvh: %u leaf records for %u primitives", hc[1], n); return RD_E_CUDA; }
    RD_CUDA(cudaMallocAsync(&out->nodes, (size_t)out->nodeCount * sizeof(RdNode8), st));
    RD_CUDA(cudaMemcpyAsync(out->nodes, wide, (size_t)out->nodeCount * sizeof(RdNode8), cudaMemcpyDeviceToDevice, st));
    RdF4 rl, rh;
    const rd_u32 rootSlot = (n == 1) ? 0u : 0u;   // inner root = slot 0; single leaf = slot n-1 = 0
    RD_CUDA(cudaMemcpyAsync(&rl, T.lo + rootSlot, sizeof(RdF4), cudaMemcpyDeviceToHost, st));
    RD_CUDA(cudaMemcpyAsync(&rh, T.hi + rootSlot, sizeof(RdF4), cudaMemcpyDeviceToHost, st));
    RD_CUDA(cudaStreamSynchronize(st));
    RD_CUDA(cudaGetLastError());
    out->lo[0] = rl.x; out->lo[1] = rl.y; out->lo[2] = rl.z;
    out->hi[0] = rh.x; out->hi[1] = rh.y; out->hi[2] = rh.z;
    return RD_OK;
}

// ---- bottom-level builds ---------------------------------------------------------------------------------------------
// Small meshes (<= RD_PLOC_SMALL_MAX triangles) build DEFERRED: upload, boxes, the one-block build kernels and the triangle
// gather are queued on one of RD_BUILD_STREAMS side streams and the call returns without waiting, so the one-block kernels
// of consecutive meshes overlap on the GPU (a scene of 4096 patch meshes: 3.1 s -> see profiles/README_r2.md). The node
// count, the root box and the non-finite-input flag arrive in host-mapped memory; rd_blas_resolve reads them the first time
// they are needed (rd_blas_stats, rd_tlas_build, rd_blas_free) and moves the nodes into an exact-size allocation. An error
// of a deferred build (NaN / Inf vertices) is therefore reported by that later call. Option "async_build" = 0 resolves at once.
#define RD_BUILD_STREAMS 32
#define RD_SUMMARY_WORDS 16
#define RD_SUMMARY_CHUNK 4096u

static int rd_build_stream(rd_platform* p, cudaStream_t* out, size_t* index = 0)
{
    if (p->buildStreams.empty()) {
        int count = RD_BUILD_STREAMS;
        if (const char* e = getenv("RD_BUILD_STREAMS")) { count = atoi(e); if (count < 1) count = 1; if (count > 64) count = 64; }   // developer knob
        for (int k = 0; k < count; ++k) { cudaStream_t s; RD_CUDA(cudaStreamCreateWithFlags(&s, cudaStreamNonBlocking)); p->buildStreams.push_back(s); }
    }
    const size_t k = p->buildNext++ % p->buildStreams.size();
    *out = p->buildStreams[k];
    if (index) *index = k;
    return RD_OK;
}

// ---- chunked device memory for small bottom-level structures (RdArenaChunk, rd_internal.h) --------------------------------
// Host-side bump allocation out of 64 MB chunks: no allocator call per mesh. A chunk whose structures have all been released
// (or, for the staging kind, resolved) is kept and reused once the platform stream has passed the point of its release.
#define RD_ARENA_CHUNK_BYTES ((size_t)64 << 20)
static int rd_arena_take(rd_platform* p, int kind, size_t bytes, void** ptr, RdArenaChunk** chunk)
{
    bytes = (bytes + 255u) & ~(size_t)255u;
    if (!bytes) bytes = 256;
    RdArenaChunk* c = p->arenaCur[kind];
    if (c && c->live == 0u && c->idleRecorded) {   // the current chunk ran empty: it starts over once the platform stream has let go of it
        RD_CUDA(cudaEventSynchronize(c->idle)); c->idleRecorded = false; c->used = 0;
    }
    if (!c || c->used + bytes > c->size) {
        c = 0;
        for (RdArenaChunk* q : p->arenaChunks)   // an empty chunk of this kind
            if (q->kind == kind && q->live == 0u && q != p->arenaCur[kind] && q->size >= bytes) { c = q; break; }
        if (c) {
            if (c->idleRecorded) { RD_CUDA(cudaEventSynchronize(c->idle)); c->idleRecorded = false; }
            c->used = 0;
        } else {
            c = new RdArenaChunk();
            memset(c, 0, sizeof(*c));
            c->size = bytes > RD_ARENA_CHUNK_BYTES ? bytes : RD_ARENA_CHUNK_BYTES; c->kind = kind;
            if (cudaMalloc(&c->base, c->size) != cudaSuccess || cudaEventCreateWithFlags(&c->idle, cudaEventDisableTiming) != cudaSuccess) {
                rd_set_error("BuildAccelStruct: %s", cudaGetErrorString(cudaGetLastError()));
                if (c->base) cudaFree(c->base);
                delete c; return RD_E_CUDA;
            }
            p->arenaChunks.push_back(c);
        }
        p->arenaCur[kind] = c;
    }
    *ptr = c->base + c->used; *chunk = c;
    c->used += bytes; c->live += 1u;
    return RD_OK;
}

// one structure less in the chunk; the platform stream may still be reading it (the resolve copy): mark the point
static void rd_arena_drop(rd_platform* p, RdArenaChunk* c)
{
    if (!c || !c->live) return;
    c->live -= 1u;
    if (c->live == 0u) c->idleRecorded = cudaEventRecord(c->idle, p->stream) == cudaSuccess;
}

// option trim_pool / tests: release the chunks nothing lives in
void rd_arena_trim(rd_platform* p)
{
    std::vector<RdArenaChunk*> keep;
    for (RdArenaChunk* c : p->arenaChunks) {
        if (c->live == 0u) {
            if (p->arenaCur[c->kind] == c) p->arenaCur[c->kind] = 0;
            cudaFree(c->base); cudaEventDestroy(c->idle); delete c;
        } else keep.push_back(c);
    }
    p->arenaChunks.swap(keep);
    for (size_t k = 0; k < p->buildScratch.size(); ++k) { if (p->buildScratch[k]) cudaFree(p->buildScratch[k]); p->buildScratch[k] = 0; p->buildScratchBytes[k] = 0; }
}

static int rd_summary_slot(rd_platform* p, volatile rd_u32** host, rd_u32** dev)
{
    if (p->summaryUsed == p->summaryHost.size() * RD_SUMMARY_CHUNK) {
        void* h = 0; void* d = 0;
        RD_CUDA(cudaHostAlloc(&h, (size_t)RD_SUMMARY_CHUNK * RD_SUMMARY_WORDS * 4, cudaHostAllocMapped));
        RD_CUDA(cudaHostGetDevicePointer(&d, h, 0));
        p->summaryHost.push_back((rd_u32*)h); p->summaryDev.push_back((rd_u32*)d);
    }
    const size_t chunk = p->summaryUsed / RD_SUMMARY_CHUNK, at = (p->summaryUsed % RD_SUMMARY_CHUNK) * RD_SUMMARY_WORDS;
    p->summaryUsed += 1;
    *host = p->summaryHost[chunk] + at; *dev = p->summaryDev[chunk] + at;
    return RD_OK;
}

int rd_blas_resolve(rd_platform* p, rd_blas* b)
{
    if (!b || !b->pending) return b && b->failed ? RD_E_ARG : RD_OK;
    RD_CUDA(cudaSetDevice(p->device));
    RD_CUDA(cudaEventSynchronize(b->done));
    cudaEventDestroy(b->done);
    b->pending = 0;
    const volatile rd_u32* hs = b->summary;
    if (hs[3]) { b->failed = 1; rd_set_error("BuildAccelStruct: %u primitive(s) with NaN / Inf coordinates", hs[3]); return RD_E_ARG; }
    if (hs[1] != b->triCount || hs[2] > 200u || hs[0] == 0u || hs[0] > b->triCount + 1u) {
        b->failed = 1; rd_set_error("rd_build_wide_bvh: %u leaf records for %u primitives", hs[1], b->triCount); return RD_E_CUDA;
    }
    b->nodeCount = hs[0]; b->levels = hs[2];
    for (int k = 0; k < 3; ++k) { rd_u32 u = hs[4 + k], v = hs[7 + k]; memcpy(&b->lo[k], &u, 4); memcpy(&b->hi[k], &v, 4); }
    // the build wrote into a worst-case allocation: keep what is used (the build has completed, so the platform stream may copy)
    RdNode8* exact = 0;
    if (b->nodesChunk) {   // staging chunk -> exact-size place in a final chunk; the staging chunk is reusable once all of it has moved
        RdArenaChunk* fin = 0;
        int rc = rd_arena_take(p, 1, (size_t)b->nodeCount * sizeof(RdNode8), (void**)&exact, &fin);
        if (rc) return rc;
        RD_CUDA(cudaMemcpyAsync(exact, b->nodes, (size_t)b->nodeCount * sizeof(RdNode8), cudaMemcpyDeviceToDevice, p->stream));
        rd_arena_drop(p, b->nodesChunk);
        b->nodes = exact; b->nodesChunk = fin;
        return RD_OK;
    }
    RD_CUDA(cudaMallocAsync(&exact, (size_t)b->nodeCount * sizeof(RdNode8), p->stream));
    RD_CUDA(cudaMemcpyAsync(exact, b->nodes, (size_t)b->nodeCount * sizeof(RdNode8), cudaMemcpyDeviceToDevice, p->stream));
    RD_CUDA(cudaFreeAsync(b->nodes, p->stream));
    b->nodes = exact;
    return RD_OK;
}

// ---- small triangle meshes: the WHOLE bottom-level build in one launch ----------------------------------------------------
// Every step of the small path above runs in one block already; a scene of thousands of small meshes (C4: 4096 patch meshes)
// then pays for the ~10 launches and ~12 stream-ordered allocations per mesh on the host. Here the steps are the phases of ONE
// 1024-thread block (primitive boxes, centroid bounds, Morton codes, sort, leaf boxes, PLOC, collapse, leaf-record gather;
// the same device functions, so the structure is bit-identical), over ONE scratch allocation that also holds the uploaded mesh.
struct RdSmallBuild {
    const float* verts; const rd_u32* idx; rd_u32 n; int radius;
    RdF4 *plo, *phi;
    unsigned* bounds; rd_u32* counters; rd_u32* where;
    rd_u64 *keysA, *keysB; rd_u32 *valsA, *valsB;
    RdBinTree T; RdPlocSide A, B; rd_u32* nearest;
    RdCollapseOut O; RdCollapseItem *itemsA, *itemsB;
    rd_u32* summary; RdTri* tris;
};

__global__ void __launch_bounds__(1024) rd_k_small_build_fused(const RdSmallBuild S)
{
    const rd_u32 tid = threadIdx.x, nth = blockDim.x, n = S.n;
    if (tid < 3) S.bounds[tid] = 0xffffffffu;
    else if (tid < 8) S.bounds[tid] = 0u;
    if (tid == 0) { S.counters[0] = 1u; S.counters[1] = 0u; S.counters[2] = 0u; S.counters[3] = 0u; }
    for (rd_u32 q = tid; q < n; q += nth) rd_prim_box_tri(S.verts, S.idx, q, S.plo, S.phi);
    __syncthreads();
    rd_centroid_bounds_part(S.plo, S.phi, n, S.bounds, tid, nth);
    __syncthreads();
    for (rd_u32 i = tid; i < n; i += nth) rd_morton_item(S.plo, S.phi, i, S.bounds, S.keysA, S.valsA);
    __syncthreads();
    if (n > 1u) {
        rd_sort_small_block(S.keysA, S.valsA, S.keysB, S.valsB, n, 8, S.where);
        __syncthreads();
        if (*S.where == 0u) for (rd_u32 i = tid; i < n; i += nth) { S.keysB[i] = S.keysA[i]; S.valsB[i] = S.valsA[i]; }
    } else if (tid == 0) { S.keysB[0] = S.keysA[0]; S.valsB[0] = S.valsA[0]; }
    __syncthreads();
    for (rd_u32 i = tid; i < n; i += nth) { const rd_u32 q = S.T.prim[i]; S.T.lo[n - 1u + i] = S.plo[q]; S.T.hi[n - 1u + i] = S.phi[q]; }
    __syncthreads();
    if (n > 1u) rd_ploc_small_block(S.T, S.A, S.B, S.nearest, S.radius);
    __syncthreads();
    RdCollapseItem rootItem; rootItem.wide = 0; rootItem.bin = (n == 1u) ? (RD_BIN_LEAF | 0u) : 0u;
    rd_collapse_small_block(S.T, S.O, S.itemsA, S.itemsB, rootItem, S.bounds, S.summary);
    __syncthreads();
    for (rd_u32 r = tid; r < n; r += nth) rd_gather_tri(S.verts, S.idx, S.O.leafPrim, S.tris, r);
}

// Page-locked staging for the uploads of small meshes. cudaMemcpyAsync from pageable memory costs ~20 us of host time per call
// (two per mesh: 40 of the ~50 us a 2450-triangle mesh took, profiles/exp_smallbuilds_r2g.log); a memcpy into a pinned ring
// and ONE asynchronous copy of vertices + indices costs ~5. A segment is reused only after everything copied out of it has
// been consumed: leaving a segment records a fence event on every side stream.
#define RD_UPLOAD_SEGMENTS 4
#define RD_UPLOAD_SEGMENT_BYTES ((size_t)8 << 20)
static int rd_upload_stage(rd_platform* p, size_t bytes, unsigned char** host)
{
    *host = 0;
    if (bytes > RD_UPLOAD_SEGMENT_BYTES || getenv("RD_BUILD_PAGEABLE")) return RD_OK;   // caller copies from pageable memory (developer knob)
    if (!p->upHost) {
        void* h = 0;
        RD_CUDA(cudaHostAlloc(&h, RD_UPLOAD_SEGMENTS * RD_UPLOAD_SEGMENT_BYTES, cudaHostAllocDefault));
        p->upHost = (unsigned char*)h; p->upSegBytes = RD_UPLOAD_SEGMENT_BYTES; p->upSeg = 0; p->upOff = 0;
    }
    if (p->upOff + bytes > p->upSegBytes) {
        std::vector<cudaEvent_t>& fence = p->upFence[p->upSeg];
        while (fence.size() < p->buildStreams.size()) { cudaEvent_t e; RD_CUDA(cudaEventCreateWithFlags(&e, cudaEventDisableTiming)); fence.push_back(e); }
        for (size_t k = 0; k < p->buildStreams.size(); ++k) RD_CUDA(cudaEventRecord(fence[k], p->buildStreams[k]));
        p->upFenced[p->upSeg] = true;
        p->upSeg = (p->upSeg + 1u) % RD_UPLOAD_SEGMENTS; p->upOff = 0;
        if (p->upFenced[p->upSeg]) {
            for (cudaEvent_t e : p->upFence[p->upSeg]) RD_CUDA(cudaEventSynchronize(e));
            p->upFenced[p->upSeg] = false;
        }
    }
    *host = p->upHost + (size_t)p->upSeg * p->upSegBytes + p->upOff;
    p->upOff += (bytes + 255u) & ~(size_t)255u;
    return RD_OK;
}

// hverts / hidx != 0: the mesh is uploaded into the head of the same scratch block (dverts / didx are ignored)
static int rd_blas_build_small_fused(rd_platform* p, cudaStream_t st, size_t streamIndex, const float* hverts, unsigned nv, const uint32_t* hidx,
                                     const float* dverts, const uint32_t* didx, unsigned nt, rd_blas** out)
{
    const size_t n = nt, ni = n > 1 ? n - 1 : 1;
    size_t off = 0;
    auto carve = [&off](size_t bytes) { const size_t at = off; off += (bytes + 255u) & ~(size_t)255u; return at; };
    const size_t oVerts = carve(hverts ? (size_t)nv * 12 : 0), oIdx = carve(hidx ? n * 12 : 0);
    const size_t oPlo = carve(n * 16), oPhi = carve(n * 16), oBounds = carve(8 * 4), oCounters = carve(4 * 4), oWhere = carve(4);
    const size_t oKeysA = carve(n * 8), oKeysB = carve(n * 8), oValsA = carve(n * 4), oValsB = carve(n * 4);
    const size_t oLeft = carve(ni * 4), oRight = carve(ni * 4), oFirst = carve(ni * 4), oLast = carve(ni * 4);
    const size_t oParent = carve(2 * n * 4), oCount = carve(ni * 4);
    const size_t oLo = carve(2 * n * 16), oHi = carve(2 * n * 16), oCost = carve(8 * ni * 4), oDec = carve(8 * ni);
    const size_t oARef = carve(n * 4), oALo = carve(n * 16), oAHi = carve(n * 16);
    const size_t oBRef = carve(n * 4), oBLo = carve(n * 16), oBHi = carve(n * 16), oNearest = carve(n * 4);
    const size_t oItemsA = carve((n + 1) * sizeof(RdCollapseItem)), oItemsB = carve((n + 1) * sizeof(RdCollapseItem)), oLeafPrim = carve(n * 4);
    // the scratch block of this side stream (stream order makes the reuse by consecutive meshes safe)
    if (p->buildScratch.size() < p->buildStreams.size()) { p->buildScratch.resize(p->buildStreams.size(), 0); p->buildScratchBytes.resize(p->buildStreams.size(), 0); }
    if (p->buildScratchBytes[streamIndex] < off) {
        RD_CUDA(cudaStreamSynchronize(st));
        if (p->buildScratch[streamIndex]) cudaFree(p->buildScratch[streamIndex]);
        p->buildScratch[streamIndex] = 0; p->buildScratchBytes[streamIndex] = 0;
        const size_t want = off + off / 2 > ((size_t)1 << 20) ? off + off / 2 : ((size_t)1 << 20);
        RD_CUDA(cudaMalloc(&p->buildScratch[streamIndex], want));
        p->buildScratchBytes[streamIndex] = want;
    }
    char* base = (char*)p->buildScratch[streamIndex];
    if (hidx) {
        // vertices and indices sit at the head of the block in the layout of the staging copy: one transfer
        unsigned char* stage = 0;
        int src = rd_upload_stage(p, oIdx + n * 12, &stage);
        if (src) return src;
        if (stage) {
            if (nv) memcpy(stage + oVerts, hverts, (size_t)nv * 12);
            memcpy(stage + oIdx, hidx, n * 12);
            RD_CUDA(cudaMemcpyAsync(base, stage, oIdx + n * 12, cudaMemcpyHostToDevice, st));
        } else {
            if (nv) RD_CUDA(cudaMemcpyAsync(base + oVerts, hverts, (size_t)nv * 12, cudaMemcpyHostToDevice, st));
            RD_CUDA(cudaMemcpyAsync(base + oIdx, hidx, n * 12, cudaMemcpyHostToDevice, st));
        }
    }
    volatile rd_u32* hostSummary = 0; rd_u32* devSummary = 0;
    int rc = rd_summary_slot(p, &hostSummary, &devSummary);
    if (rc) return rc;
    rd_blas* b = new rd_blas();
    memset(b, 0, sizeof(*b));
    b->owner = p; b->triCount = nt;
    // triangle records in their final place, the worst-case node array in a staging chunk (rd_blas_resolve keeps what is used)
    rc = rd_arena_take(p, 1, n * sizeof(RdTri), (void**)&b->tris, &b->trisChunk);
    if (!rc) rc = rd_arena_take(p, 0, (n + 1) * sizeof(RdNode8), (void**)&b->nodes, &b->nodesChunk);
    if (rc) { rd_arena_drop(p, b->trisChunk); delete b; return rc; }
    RdSmallBuild S;
    S.verts = hverts ? (const float*)(base + oVerts) : dverts; S.idx = hidx ? (const rd_u32*)(base + oIdx) : didx; S.n = nt;
    S.radius = 32;
    if (const char* e = getenv("RD_PLOC_RADIUS")) { S.radius = atoi(e); if (S.radius < 1) S.radius = 1; }
    S.plo = (RdF4*)(base + oPlo); S.phi = (RdF4*)(base + oPhi);
    S.bounds = (unsigned*)(base + oBounds); S.counters = (rd_u32*)(base + oCounters); S.where = (rd_u32*)(base + oWhere);
    S.keysA = (rd_u64*)(base + oKeysA); S.keysB = (rd_u64*)(base + oKeysB); S.valsA = (rd_u32*)(base + oValsA); S.valsB = (rd_u32*)(base + oValsB);
    RdBinTree& T = S.T;
    T.n = nt; T.keys = S.keysB; T.prim = S.valsB;
    T.left = (rd_u32*)(base + oLeft); T.right = (rd_u32*)(base + oRight); T.first = (rd_u32*)(base + oFirst); T.last = (rd_u32*)(base + oLast);
    T.parent = (rd_u32*)(base + oParent); T.arrive = 0; T.count = (rd_u32*)(base + oCount);
    T.lo = (RdF4*)(base + oLo); T.hi = (RdF4*)(base + oHi); T.cost = (float*)(base + oCost); T.dec = (rd_u8*)(base + oDec);
    S.A.ref = (rd_u32*)(base + oARef); S.A.lo = (RdF4*)(base + oALo); S.A.hi = (RdF4*)(base + oAHi);
    S.B.ref = (rd_u32*)(base + oBRef); S.B.lo = (RdF4*)(base + oBLo); S.B.hi = (RdF4*)(base + oBHi);
    S.nearest = (rd_u32*)(base + oNearest);
    S.O.nodes = b->nodes; S.O.leafPrim = (rd_u32*)(base + oLeafPrim); S.O.counters = S.counters; S.O.next = 0;
    S.itemsA = (RdCollapseItem*)(base + oItemsA); S.itemsB = (RdCollapseItem*)(base + oItemsB);
    S.summary = devSummary; S.tris = b->tris;
    rd_k_small_build_fused<<<1, 1024, 0, st>>>(S);
    p->launches += 1;
    b->pending = 1; b->summary = hostSummary;
    if (cudaGetLastError() != cudaSuccess || cudaEventCreateWithFlags(&b->done, cudaEventDisableTiming) != cudaSuccess ||
        cudaEventRecord(b->done, st) != cudaSuccess) {
        rd_set_error("rd_blas_build: %s", cudaGetErrorString(cudaGetLastError()));
        rd_arena_drop(p, b->nodesChunk); rd_arena_drop(p, b->trisChunk); delete b; return RD_E_CUDA;
    }
    if (!p->optAsyncBuild) {
        rc = rd_blas_resolve(p, b);
        if (rc) { rd_arena_drop(p, b->nodesChunk); rd_arena_drop(p, b->trisChunk); delete b; return rc; }
    }
    *out = b;
    return RD_OK;
}

// Many deferred builds resolved together (the instance build of a scene of small meshes): one kernel moves all node arrays from
// their staging places to exact-size final places instead of one cudaMemcpyAsync per mesh (4096 meshes: ~20 ms of host time).
struct RdMoveDesc { const uint4* src; uint4* dst; unsigned count16, pad; };   // count16 = 16-byte words
__global__ void rd_k_move_many(const RdMoveDesc* d)
{
    const RdMoveDesc m = d[blockIdx.x];
    for (unsigned i = threadIdx.x; i < m.count16; i += blockDim.x) m.dst[i] = m.src[i];
}

static int rd_blas_resolve_many(rd_platform* p, std::vector<rd_blas*>& list)
{
    std::vector<RdMoveDesc> moves;
    std::vector<rd_blas*> moved;
    std::vector<RdArenaChunk*> fins;
    for (rd_blas* b : list) {
        if (!b->pending || !b->nodesChunk) continue;
        RD_CUDA(cudaEventSynchronize(b->done));
        const volatile rd_u32* hs = b->summary;
        if (hs[3] || hs[1] != b->triCount || hs[2] > 200u || hs[0] == 0u || hs[0] > b->triCount + 1u) continue;   // rd_blas_resolve reports it
        cudaEventDestroy(b->done);
        b->pending = 0;
        b->nodeCount = hs[0]; b->levels = hs[2];
        for (int k = 0; k < 3; ++k) { rd_u32 u = hs[4 + k], v = hs[7 + k]; memcpy(&b->lo[k], &u, 4); memcpy(&b->hi[k], &v, 4); }
        RdMoveDesc m; RdArenaChunk* fin = 0; void* dst = 0;
        int rc = rd_arena_take(p, 1, (size_t)b->nodeCount * sizeof(RdNode8), &dst, &fin);
        if (rc) return rc;
        m.src = (const uint4*)b->nodes; m.dst = (uint4*)dst; m.count16 = b->nodeCount * (unsigned)(sizeof(RdNode8) / 16u); m.pad = 0;
        moves.push_back(m); moved.push_back(b); fins.push_back(fin);
    }
    if (moves.empty()) return RD_OK;
    RdTmp tmp(p->stream);
    RdMoveDesc* dm;
    RD_CUDA(tmp.alloc(&dm, moves.size()));
    RD_CUDA(cudaMemcpyAsync(dm, moves.data(), moves.size() * sizeof(RdMoveDesc), cudaMemcpyHostToDevice, p->stream));
    RD_CUDA(cudaStreamSynchronize(p->stream));   // `moves` is pageable host memory: make sure it has been read before it goes away
    rd_k_move_many<<<(unsigned)moves.size(), 128, 0, p->stream>>>(dm);
    p->launches += 1;
    RD_CUDA(cudaGetLastError());
    for (size_t k = 0; k < moved.size(); ++k) {
        rd_blas* b = moved[k];
        RdArenaChunk* stage = b->nodesChunk;
        b->nodes = (RdNode8*)moves[k].dst; b->nodesChunk = fins[k];
        rd_arena_drop(p, stage);   // records the platform stream's position behind the move when the chunk runs empty
    }
    return RD_OK;
}

static int rd_blas_build_on(rd_platform* p, cudaStream_t st, bool deferred, const float* dverts, const uint32_t* didx, unsigned nt, rd_blas** out)
{
    RdTmp tmp(st);
    RdF4 *plo, *phi;
    RD_CUDA(tmp.alloc(&plo, nt)); RD_CUDA(tmp.alloc(&phi, nt));
    if (nt) { rd_k_tri_boxes<<<rd_grid(nt), RD_BT, 0, st>>>(dverts, didx, nt, plo, phi); p->launches += 1; }
    RdBuildResult r;
    memset(&r, 0, sizeof(r));
    volatile rd_u32* hostSummary = 0; rd_u32* devSummary = 0;
    int rc;
    if (deferred) {
        rc = rd_summary_slot(p, &hostSummary, &devSummary);
        if (rc) return rc;
        rc = rd_build_wide_bvh_small(p, st, plo, phi, nt, &r, devSummary);
    } else rc = rd_build_wide_bvh(p, plo, phi, nt, &r);
    if (rc) return rc;
    rd_blas* b = new rd_blas();
    memset(b, 0, sizeof(*b));
    b->owner = p;
    b->nodes = r.nodes; b->nodeCount = r.nodeCount; b->triCount = nt; b->levels = r.levels;
    memcpy(b->lo, r.lo, 12); memcpy(b->hi, r.hi, 12);
    RD_CUDA(cudaMallocAsync(&b->tris, nt ? (size_t)nt * sizeof(RdTri) : 48, st));
    if (nt) { rd_k_gather_tris<<<rd_grid(nt), RD_BT, 0, st>>>(dverts, didx, r.leafPrim, b->tris, nt); p->launches += 1; }
    RD_CUDA(cudaGetLastError());
    RD_CUDA(cudaFreeAsync(r.leafPrim, st));   // everything that follows on the same stream is ordered behind the build
    if (deferred) {
        b->pending = 1; b->summary = hostSummary;
        RD_CUDA(cudaEventCreateWithFlags(&b->done, cudaEventDisableTiming));
        RD_CUDA(cudaEventRecord(b->done, st));
        if (!p->optAsyncBuild) {
            rc = rd_blas_resolve(p, b);
            if (rc) { cudaFree(b->nodes); cudaFree(b->tris); delete b; return rc; }
        }
    }
    *out = b;
    return RD_OK;
}

static bool rd_build_deferrable(unsigned nt)
{
    return nt > 0 && nt <= RD_PLOC_SMALL_MAX && nt <= RD_SORT_SMALL_MAX && !getenv("RD_BUILD_GREEDY") && !getenv("RD_BUILD_LBVH") && !getenv("RD_BUILD_GENERAL");
}

extern "C" int rd_blas_build_device(rd_platform* p, const float* dverts, unsigned nv, const uint32_t* didx, unsigned nt, rd_blas** out)
{
    if (!p || !out || (nt && (!dverts || !didx))) { rd_set_error("rd_blas_build: null argument"); return RD_E_ARG; }
    RD_CUDA(cudaSetDevice(p->device));
    (void)nv;
    if (!rd_build_deferrable(nt)) return rd_blas_build_on(p, p->stream, false, dverts, didx, nt, out);
    // the caller's arrays were produced on the platform stream: the side stream starts behind what is queued there now
    cudaStream_t st; size_t streamIndex = 0; int rc = rd_build_stream(p, &st, &streamIndex);
    if (rc) return rc;
    cudaEvent_t ready;
    RD_CUDA(cudaEventCreateWithFlags(&ready, cudaEventDisableTiming));
    RD_CUDA(cudaEventRecord(ready, p->stream));
    RD_CUDA(cudaStreamWaitEvent(st, ready, 0));
    cudaEventDestroy(ready);
    if (!getenv("RD_BUILD_UNFUSED")) return rd_blas_build_small_fused(p, st, streamIndex, 0, nv, 0, dverts, didx, nt, out);
    return rd_blas_build_on(p, st, true, dverts, didx, nt, out);
}

extern "C" int rd_blas_build(rd_platform* p, const float* verts, unsigned nv, const uint32_t* idx, unsigned nt, rd_blas** out)
{
    if (!p || !out || (nt && (!verts || !idx))) { rd_set_error("rd_blas_build: null argument"); return RD_E_ARG; }
    RD_CUDA(cudaSetDevice(p->device));
    for (size_t k = 0; k < (size_t)nt * 3; ++k)
        if (idx[k] >= nv) { rd_set_error("rd_blas_build: index %u out of range (nv=%u)", idx[k], nv); return RD_E_ARG; }
    const bool deferred = rd_build_deferrable(nt);
    cudaStream_t st = p->stream;
    size_t streamIndex = 0;
    if (deferred) { int rc = rd_build_stream(p, &st, &streamIndex); if (rc) return rc; }
    if (deferred && !getenv("RD_BUILD_UNFUSED")) return rd_blas_build_small_fused(p, st, streamIndex, verts, nv, idx, 0, 0, nt, out);   // developer knob: the multi-launch small path
    RdTmp tmp(st);
    float* dv; uint32_t* di;
    RD_CUDA(tmp.alloc(&dv, (size_t)nv * 3)); RD_CUDA(tmp.alloc(&di, (size_t)nt * 3));
    if (nv) RD_CUDA(cudaMemcpyAsync(dv, verts, (size_t)nv * 12, cudaMemcpyHostToDevice, st));
    if (nt) RD_CUDA(cudaMemcpyAsync(di, idx, (size_t)nt * 12, cudaMemcpyHostToDevice, st));
    return rd_blas_build_on(p, st, deferred, dv, di, nt, out);
}

extern "C" int rd_blas_free(rd_platform* p, rd_blas* b)
{
    if (!b) return RD_OK;
    if (p) cudaSetDevice(p->device);
    if (b->pending) { cudaEventSynchronize(b->done); cudaEventDestroy(b->done); b->pending = 0; }
    if (b->nodesChunk) rd_arena_drop(b->owner ? b->owner : p, b->nodesChunk); else cudaFree(b->nodes);
    if (b->trisChunk) rd_arena_drop(b->owner ? b->owner : p, b->trisChunk); else cudaFree(b->tris);
    delete b;
    return RD_OK;
}

extern "C" int rd_blas_stats(rd_blas* b, uint64_t out[4])
{
    if (!b || !out) { rd_set_error("rd_blas_stats: null argument"); return RD_E_ARG; }
    if (b->pending || b->failed) { int rc = rd_blas_resolve(b->owner, b); if (rc) return rc; }
    out[0] = b->nodeCount; out[1] = b->triCount; out[2] = b->levels;
    out[3] = (uint64_t)b->nodeCount * sizeof(RdNode8) + (uint64_t)b->triCount * sizeof(RdTri);
    return RD_OK;
}

extern "C" int rd_tlas_build(rd_platform* p, const rd_instance_desc* instances, unsigned n, rd_handle* out)
{
    if (!p || !out || (n && !instances)) { rd_set_error("rd_tlas_build: null argument"); return RD_E_ARG; }
    RD_CUDA(cudaSetDevice(p->device));
    cudaStream_t st = p->stream;
    std::vector<RdInstance> inst(n);
    std::vector<RdF4> lo(n), hi(n);
    {   // deferred small builds: resolve them together
        std::vector<rd_blas*> pend;
        for (unsigned i = 0; i < n; ++i) {
            rd_blas* b = instances[i].bottomAccelStruct;
            if (b && b->pending && b->nodesChunk && b->levels != 0xffffffffu) { b->levels = 0xffffffffu; pend.push_back(b); }   // levels doubles as the "listed" mark
        }
        for (rd_blas* b : pend) b->levels = 0;
        if (pend.size() > 4) { int prc = rd_blas_resolve_many(p, pend); if (prc) return prc; }
    }
    for (unsigned i = 0; i < n; ++i) {
        rd_blas* b = instances[i].bottomAccelStruct;
        if (!b) { rd_set_error("rd_tlas_build: instance %u has no bottom-level structure", i); return RD_E_ARG; }
        if (b->pending || b->failed) { int brc = rd_blas_resolve(p, b); if (brc) return brc; }   // deferred small builds: node count, root box, exact storage
        RdInstance& I = inst[i];
        memcpy(I.fwd, instances[i].transform, 64);
        float inv[16];
        if (!rd_inverse4x4(I.fwd, inv)) { rd_set_error("rd_tlas_build: instance %u transform is singular", i); return RD_E_ARG; }
        memcpy(I.inv, inv, 48);
        I.sbtOffset = instances[i].SBTOffset; I.instanceID = i; I.customID = instances[i].customInstanceID; I.flags = b->triCount;   // flags = triangle records of the BLAS (rd_tlas_save)
        I.nodes = b->nodes; I.tris = b->tris;
        I.maxAbs = 0.0f; I.pad[0] = I.pad[1] = I.pad[2] = 0u;
        for (int k = 0; k < 3; ++k) I.maxAbs = fmaxf(I.maxAbs, fmaxf(fabsf(b->lo[k]), fabsf(b->hi[k])));
        if (!(I.maxAbs < 3.0e38f)) I.maxAbs = 0.0f;   // empty structure: inverted root box
        RdF4 bl, bh; bl.x = b->lo[0]; bl.y = b->lo[1]; bl.z = b->lo[2]; bl.w = 0; bh.x = b->hi[0]; bh.y = b->hi[1]; bh.z = b->hi[2]; bh.w = 0;
        rd_instance_world_box(I.fwd, bl, bh, &lo[i], &hi[i]);
    }
    RdTmp tmp(st);
    RdF4 *dlo, *dhi;
    RD_CUDA(tmp.alloc(&dlo, n)); RD_CUDA(tmp.alloc(&dhi, n));
    if (n) {
        RD_CUDA(cudaMemcpyAsync(dlo, lo.data(), (size_t)n * sizeof(RdF4), cudaMemcpyHostToDevice, st));
        RD_CUDA(cudaMemcpyAsync(dhi, hi.data(), (size_t)n * sizeof(RdF4), cudaMemcpyHostToDevice, st));
    }
    RdBuildResult r;
    int rc = rd_build_wide_bvh(p, dlo, dhi, n, &r);
    if (rc) return rc;
    // L2 residency of the hot node array (option "l2_persist", bytes; 0 = off): the node array of the largest bottom-level
    // structure becomes a persisting access-policy window of the platform stream, its triangle records stream through --
    // on a structure larger than L2 the 10 node fetches of a ray then hit L2 while the 3 triangle fetches go to HBM.
    if (p->optL2Persist > 0 && n) {
        const rd_blas* big = 0;
        for (unsigned i = 0; i < n; ++i) { const rd_blas* b = instances[i].bottomAccelStruct; if (!big || b->nodeCount > big->nodeCount) big = b; }
        int maxWin = 0, maxPersist = 0;
        cudaDeviceGetAttribute(&maxWin, cudaDevAttrMaxAccessPolicyWindowSize, p->device);
        cudaDeviceGetAttribute(&maxPersist, cudaDevAttrMaxPersistingL2CacheSize, p->device);
        size_t want = (size_t)big->nodeCount * sizeof(RdNode8);
        size_t carve = (size_t)p->optL2Persist < (size_t)maxPersist ? (size_t)p->optL2Persist : (size_t)maxPersist;
        if (maxWin > 0 && carve > 0) {
            cudaDeviceSetLimit(cudaLimitPersistingL2CacheSize, carve);
            cudaStreamAttrValue attr; memset(&attr, 0, sizeof(attr));
            attr.accessPolicyWindow.base_ptr = (void*)big->nodes;
            attr.accessPolicyWindow.num_bytes = want < (size_t)maxWin ? want : (size_t)maxWin;
            attr.accessPolicyWindow.hitRatio = want <= carve ? 1.0f : (float)carve / (float)want;
            attr.accessPolicyWindow.hitProp = cudaAccessPropertyPersisting;
            attr.accessPolicyWindow.missProp = cudaAccessPropertyStreaming;
            if (cudaStreamSetAttribute(p->stream, cudaStreamAttributeAccessPolicyWindow, &attr) != cudaSuccess) cudaGetLastError();
        }
    }
    RdHandleObj* o = new RdHandleObj();
    memset(o, 0, sizeof(*o));
    o->magic = RD_HANDLE_MAGIC; o->kind = RD_H_TLAS;
    o->dnodes = r.nodes; o->dleaf = r.leafPrim; o->instanceCount = n; o->nodeCount = r.nodeCount;
    RD_CUDA(cudaMalloc(&o->dinst, n ? (size_t)n * sizeof(RdInstance) : 16));
    if (n) RD_CUDA(cudaMemcpyAsync(o->dinst, inst.data(), (size_t)n * sizeof(RdInstance), cudaMemcpyHostToDevice, st));
    RdScene S; memset(&S, 0, sizeof(S));
    S.nodes = r.nodes; S.leafInstances = r.leafPrim; S.instances = (const RdInstance*)o->dinst;
    S.instanceCount = n; S.nodeCount = r.nodeCount; S.status = p->dstatus; S.rayCounter = (rd_u64*)p->dRayCounter;
    S.maxAbs = 0.0f;
    for (int k = 0; k < 3; ++k) S.maxAbs = fmaxf(S.maxAbs, fmaxf(fabsf(r.lo[k]), fabsf(r.hi[k])));
    if (!(S.maxAbs < 3.0e38f) || !n) S.maxAbs = 0.0f;
    RD_CUDA(cudaMalloc(&o->dptr, sizeof(RdScene)));
    RD_CUDA(cudaMemcpyAsync(o->dptr, &S, sizeof(S), cudaMemcpyHostToDevice, st));
    RD_CUDA(cudaStreamSynchronize(st));
    o->size = sizeof(RdScene);
    o->arg = (uint64_t)(uintptr_t)o->dptr;
    *out = o;
    return RD_OK;
}

// ---- diagnostics: the build's sort and scan on host arrays (tests/test_gpu_parity.py checks them against numpy) ----
extern "C" int rd_debug_sort_pairs(rd_platform* p, uint64_t* keys, uint32_t* vals, unsigned n, int bits)
{
    if (!p || (n && (!keys || !vals)) || bits < 1 || bits > 64) { rd_set_error("rd_debug_sort_pairs: bad argument"); return RD_E_ARG; }
    RD_CUDA(cudaSetDevice(p->device));
    cudaStream_t st = p->stream;
    RdTmp tmp(st);
    rd_u64 *ka, *kb; rd_u32 *va, *vb;
    RD_CUDA(tmp.alloc(&ka, n)); RD_CUDA(tmp.alloc(&kb, n)); RD_CUDA(tmp.alloc(&va, n)); RD_CUDA(tmp.alloc(&vb, n));
    RdSortScratch S;
    const size_t tableCount = 256u * (size_t)rd_sort_blocks(n);
    RD_CUDA(tmp.alloc(&S.hist, 8 * 256)); RD_CUDA(tmp.alloc(&S.table, tableCount));
    RD_CUDA(tmp.alloc(&S.scan, rd_scan_scratch_count(tableCount))); RD_CUDA(tmp.alloc(&S.where, 4));
    if (n) {
        RD_CUDA(cudaMemcpyAsync(ka, keys, (size_t)n * 8, cudaMemcpyHostToDevice, st));
        RD_CUDA(cudaMemcpyAsync(va, vals, (size_t)n * 4, cudaMemcpyHostToDevice, st));
    }
    unsigned long long launches = 0;
    const int side = rd_radix_sort_pairs(st, ka, va, kb, vb, n, bits, S, p->smCount, &launches);
    p->launches += launches;
    if (side < 0) { rd_set_error("rd_debug_sort_pairs: %s", cudaGetErrorString(cudaGetLastError())); return RD_E_CUDA; }
    if (n) {
        RD_CUDA(cudaMemcpyAsync(keys, side ? kb : ka, (size_t)n * 8, cudaMemcpyDeviceToHost, st));
        RD_CUDA(cudaMemcpyAsync(vals, side ? vb : va, (size_t)n * 4, cudaMemcpyDeviceToHost, st));
    }
    RD_CUDA(cudaStreamSynchronize(st));
    return RD_OK;
}

extern "C" int rd_debug_exclusive_scan(rd_platform* p, uint32_t* data, unsigned n, uint32_t* total)
{
    if (!p || (n && !data)) { rd_set_error("rd_debug_exclusive_scan: bad argument"); return RD_E_ARG; }
    RD_CUDA(cudaSetDevice(p->device));
    cudaStream_t st = p->stream;
    RdTmp tmp(st);
    rd_u32 *d, *scratch, *dt;
    RD_CUDA(tmp.alloc(&d, n)); RD_CUDA(tmp.alloc(&scratch, rd_scan_scratch_count(n))); RD_CUDA(tmp.alloc(&dt, 4));
    RD_CUDA(cudaMemsetAsync(dt, 0, 4, st));
    if (n) RD_CUDA(cudaMemcpyAsync(d, data, (size_t)n * 4, cudaMemcpyHostToDevice, st));
    unsigned long long launches = 0;
    rd_exclusive_scan_u32(st, d, d, n, scratch, dt, &launches);
    p->launches += launches;
    if (n) RD_CUDA(cudaMemcpyAsync(data, d, (size_t)n * 4, cudaMemcpyDeviceToHost, st));
    if (total) RD_CUDA(cudaMemcpyAsync(total, dt, 4, cudaMemcpyDeviceToHost, st));
    RD_CUDA(cudaStreamSynchronize(st));
    RD_CUDA(cudaGetLastError());
    return RD_OK;
}

// ---- rd_sort_pairs_noalloc ------------------------------------------------------------------------------------------
// The pair sort of rd_sort.h (count / scan / scatter per 8-bit digit) for callers that cannot allocate: its ping-pong
// buffers, the digit table and the scan counters live in ONE file-scope device array, so the call touches no allocator.
// No counterpart in the reference (include/radiance_b200.h).
// Layout of the array (RD_NOALLOC_MAX_PAIRS = 2^18 pairs): keys A | keys B | values A | values B | table | scan counters.
// Nothing in it has to start at zero: every pass's count kernel writes all 256 x blocks table entries and the scan writes
// every tile sum it later reads, so stale contents of the previous call are never read. Two calls that overlap on the GPU
// (different streams / platforms of one device) race on the array.
#define RD_NOALLOC_MAX_PAIRS RD_SORT_NOALLOC_MAX_PAIRS
#define RD_NOALLOC_TABLE_COUNT (256u * (RD_NOALLOC_MAX_PAIRS / RD_SORT_TILE))        // 16384 counters
#define RD_NOALLOC_SCAN_COUNT 64u                                                     // >= rd_scan_scratch_count(16384) = 9
#define RD_NOALLOC_OFF_KEYS_B ((size_t)RD_NOALLOC_MAX_PAIRS * 8u)
#define RD_NOALLOC_OFF_VALS_A ((size_t)RD_NOALLOC_MAX_PAIRS * 16u)
#define RD_NOALLOC_OFF_VALS_B ((size_t)RD_NOALLOC_MAX_PAIRS * 20u)
#define RD_NOALLOC_OFF_TABLE ((size_t)RD_NOALLOC_MAX_PAIRS * 24u)
#define RD_NOALLOC_OFF_SCAN (RD_NOALLOC_OFF_TABLE + (size_t)RD_NOALLOC_TABLE_COUNT * 4u)
#define RD_NOALLOC_BYTES (RD_NOALLOC_OFF_SCAN + (size_t)RD_NOALLOC_SCAN_COUNT * 4u)  // 6 357 248 B per GPU that loads the library
static __device__ __align__(256) unsigned char rd_noalloc_scratch[RD_NOALLOC_BYTES];

#define RD_NOALLOC_MAX_DEVICES 64
static void* g_noallocBase[RD_NOALLOC_MAX_DEVICES];   // device address of rd_noalloc_scratch per device (the symbol itself is only a host stand-in)
static std::mutex g_noallocMutex;

// the array's address on `device`, which must be current; resolved once per device
static int rd_noalloc_base(int device, unsigned char** base)
{
    if (device < 0 || device >= RD_NOALLOC_MAX_DEVICES) { rd_set_error("rd_sort_pairs_noalloc: device %d out of range", device); return RD_E_ARG; }
    std::lock_guard<std::mutex> lock(g_noallocMutex);
    if (!g_noallocBase[device]) {
        void* ptr = 0;
        RD_CUDA(cudaGetSymbolAddress((void**)&ptr, rd_noalloc_scratch));
        if (!ptr || ((uintptr_t)ptr & 255u)) { rd_set_error("rd_sort_pairs_noalloc: scratch symbol resolved to %p", ptr); return RD_E_CUDA; }
        g_noallocBase[device] = ptr;
    }
    *base = (unsigned char*)g_noallocBase[device];
    return RD_OK;
}

extern "C" int rd_sort_pairs_noalloc(rd_platform* p, const uint64_t* dkeysIn, const uint32_t* dvalsIn, uint64_t* dkeysOut,
                                     uint32_t* dvalsOut, unsigned n, int bits)
{
    if (!p || bits < 1 || bits > 64 || (n && (!dkeysIn || !dvalsIn || !dkeysOut || !dvalsOut))) { rd_set_error("rd_sort_pairs_noalloc: bad argument"); return RD_E_ARG; }
    if (n > RD_NOALLOC_MAX_PAIRS) {
        rd_set_error("rd_sort_pairs_noalloc: n = %u exceeds the static scratch (at most %u pairs)", n, RD_NOALLOC_MAX_PAIRS);
        return RD_E_ARG;
    }
    if (n && ((const void*)dkeysIn == (const void*)dkeysOut || (const void*)dvalsIn == (const void*)dvalsOut)) {
        rd_set_error("rd_sort_pairs_noalloc: outputs must not alias inputs"); return RD_E_ARG;
    }
    if (!n) return RD_OK;
    RD_CUDA(cudaSetDevice(p->device));
    unsigned char* base = 0;
    const int rc = rd_noalloc_base(p->device, &base);
    if (rc != RD_OK) return rc;
    cudaStream_t st = p->stream;
    rd_u64* keysT[2] = { (rd_u64*)base, (rd_u64*)(base + RD_NOALLOC_OFF_KEYS_B) };
    rd_u32* valsT[2] = { (rd_u32*)(base + RD_NOALLOC_OFF_VALS_A), (rd_u32*)(base + RD_NOALLOC_OFF_VALS_B) };
    rd_u32* table = (rd_u32*)(base + RD_NOALLOC_OFF_TABLE);
    rd_u32* scan = (rd_u32*)(base + RD_NOALLOC_OFF_SCAN);
    // every pass runs (no histogram read-back, so no host synchronisation): pass 0 reads the caller's input, the last pass
    // scatters into the caller's output, the passes between alternate over the two halves of the static array
    const int passes = (bits + 7) / 8;
    const rd_u32 blocks = rd_sort_blocks(n);
    unsigned long long launches = 0;
    const rd_u64* kin = (const rd_u64*)dkeysIn; const rd_u32* vin = dvalsIn;
    for (int q = 0; q < passes; ++q) {
        rd_u64* kout = q == passes - 1 ? (rd_u64*)dkeysOut : keysT[q & 1];
        rd_u32* vout = q == passes - 1 ? dvalsOut : valsT[q & 1];
        rd_k_sort_count<<<blocks, RD_SORT_THREADS, 0, st>>>(kin, n, 8 * q, table);
        launches += 1;
        rd_exclusive_scan_u32(st, table, table, 256u * blocks, scan, 0, &launches);
        rd_k_sort_scatter<<<blocks, RD_SORT_THREADS, 0, st>>>(kin, vin, kout, vout, n, 8 * q, table);
        launches += 1;
        kin = kout; vin = vout;
    }
    p->launches += launches;
    RD_CUDA(cudaGetLastError());
    return RD_OK;
}
