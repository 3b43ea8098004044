// rd_build.h -- per-element steps of the GPU LBVH build and the 8-wide collapse.
//
// Replaces the reference's single-threaded recursive SAH builder and its depth-first flattening
// (radiance/src/bvh.cpp:46-285, :288-420, :463-597; 30 s for 1 M triangles on one core):
//   1. primitive boxes + centroid bounds          rd_prim_box_*            (one thread / primitive)
//   2. 63-bit Morton codes                         rd_morton63              (one thread / primitive)
//   3. radix sort of (code, primitive)             rd_radix_sort.cuh (device) / std::sort (host tests)
//   4. Karras 2012 hierarchy emission              rd_karras_node           (one thread / inner node)
//   5. bottom-up box refit with arrival counters   rd_refit_from_leaf       (one thread / leaf)
//   6. level-synchronous collapse into RdNode8     rd_collapse_item         (one thread / wide node)
//   7. leaf record gather                          rd_gather_tri            (one thread / record)
// Each step is a __host__ __device__ function of one work-item index so that the very same code is
// launched as CUDA kernels by rd_build.cu and looped serially by tests/hostemu (CPU unit tests).
#ifndef RD_BUILD_H
#define RD_BUILD_H
#include "rd_types.h"
#include "rd_traverse.h"   // RD_F2U / RD_U2F / RD_CLZ helpers

#ifdef __CUDA_ARCH__
#define RD_ATOMIC_ADD_U32(p, v) atomicAdd((p), (v))
#define RD_THREADFENCE() __threadfence()
#define RD_CLZ64(x) __clzll((long long)(x))
#else
static inline rd_u32 rd_host_atomic_add_(rd_u32* p, rd_u32 v) { rd_u32 o = *p; *p = o + v; return o; }
#define RD_ATOMIC_ADD_U32(p, v) rd_host_atomic_add_((p), (v))
#define RD_THREADFENCE()
#define RD_CLZ64(x) ((x) ? __builtin_clzll(x) : 64)
#endif

#ifndef RD_LEAF_MAX
#define RD_LEAF_MAX 3            // primitives per leaf slot (one bit of the slot's leaf nibble each: at most 4)
#endif
#define RD_BIN_LEAF 0x80000000u  // binary child reference: top bit = leaf, low bits = sorted position

// Binary hierarchy over n sorted primitives (Karras 2012 or PLOC, rd_build.cu): n-1 inner nodes [0, n-2], node 0 = root.
struct RdBinTree {
    rd_u32 n;               // primitives
    const rd_u64* keys;     // sorted Morton codes
    const rd_u32* prim;     // sorted primitive indices
    rd_u32* left;           // [n-1] child refs (RD_BIN_LEAF | pos for leaves)
    rd_u32* right;          // [n-1]
    rd_u32* first;          // [n-1] covered range of sorted positions
    rd_u32* last;           // [n-1]
    rd_u32* parent;         // [2n-1] parent inner node of inner i (slot i) / leaf pos (slot n-1+pos)
    rd_u32* arrive;         // [n-1] refit arrival counters (zeroed)
    rd_u32* count;          // [n-1] primitives below inner node i (subtrees need not be contiguous in sorted order: PLOC)
    RdF4* lo;               // [2n-1] boxes: inner i -> slot i, leaf pos -> slot n-1+pos
    RdF4* hi;
    float* cost;            // [8(n-1)] cost[8i + k-1] = SAH cost of subtree i laid out in at most k slots of its parent
    rd_u8* dec;             // [8(n-1)] decisions that reach those costs (see rd_dp_node)
};

// SAH constants of the wide-node cost model: one 8-wide node test vs one triangle test
#ifndef RD_SAH_NODE
#define RD_SAH_NODE 1.0f
#endif
#ifndef RD_SAH_PRIM
#define RD_SAH_PRIM 0.35f
#endif

RD_HD rd_u64 rd_expand21(rd_u32 v)
{
    rd_u64 x = (rd_u64)(v & 0x1fffffu);
    x = (x | (x << 32)) & 0x1f00000000ffffull;
    x = (x | (x << 16)) & 0x1f0000ff0000ffull;
    x = (x | (x << 8)) & 0x100f00f00f00f00full;
    x = (x | (x << 4)) & 0x10c30c30c30c30c3ull;
    x = (x | (x << 2)) & 0x1249249249249249ull;
    return x;
}

// centroid (cx,cy,cz) normalised by the scene's centroid bounds -> 63-bit Morton code
RD_HD rd_u64 rd_morton63(float cx, float cy, float cz, const float* cmin, const float* cinv)
{
    const float s = 2097152.0f;   // 2^21
    float fx = (cx - cmin[0]) * cinv[0] * s, fy = (cy - cmin[1]) * cinv[1] * s, fz = (cz - cmin[2]) * cinv[2] * s;
    fx = fx < 0.0f ? 0.0f : (fx > 2097151.0f ? 2097151.0f : fx);
    fy = fy < 0.0f ? 0.0f : (fy > 2097151.0f ? 2097151.0f : fy);
    fz = fz < 0.0f ? 0.0f : (fz > 2097151.0f ? 2097151.0f : fz);
    return (rd_expand21((rd_u32)fx) << 2) | (rd_expand21((rd_u32)fy) << 1) | rd_expand21((rd_u32)fz);
}

// common-prefix length of sorted positions i and j (ties broken by position), -1 out of range
RD_HD int rd_delta(const rd_u64* keys, int n, int i, int j)
{
    if (j < 0 || j >= n) return -1;
    const rd_u64 a = keys[i], b = keys[j];
    if (a != b) return RD_CLZ64(a ^ b);
    return 64 + RD_CLZ((rd_u32)i ^ (rd_u32)j);
}

// Karras 2012, one inner node.
RD_HD void rd_karras_node(const RdBinTree& T, int i)
{
    const int n = (int)T.n;
    const rd_u64* keys = T.keys;
    const int dNext = rd_delta(keys, n, i, i + 1), dPrev = rd_delta(keys, n, i, i - 1);
    const int d = dNext > dPrev ? 1 : -1;
    const int dMin = dNext > dPrev ? dPrev : dNext;
    int lmax = 2;
    while (rd_delta(keys, n, i, i + lmax * d) > dMin) lmax <<= 1;
    int l = 0;
    for (int t = lmax >> 1; t >= 1; t >>= 1)
        if (rd_delta(keys, n, i, i + (l + t) * d) > dMin) l += t;
    const int j = i + l * d;
    const int dNode = rd_delta(keys, n, i, j);
    int s = 0;
    int t = l;
    do {
        t = (t + 1) >> 1;
        if (rd_delta(keys, n, i, i + (s + t) * d) > dNode) s += t;
    } while (t > 1);
    const int gamma = i + s * d + (d < 0 ? -1 : 0);
    const int lo = i < j ? i : j, hi = i < j ? j : i;
    const rd_u32 L = (lo == gamma) ? (RD_BIN_LEAF | (rd_u32)gamma) : (rd_u32)gamma;
    const rd_u32 R = (hi == gamma + 1) ? (RD_BIN_LEAF | (rd_u32)(gamma + 1)) : (rd_u32)(gamma + 1);
    T.left[i] = L; T.right[i] = R; T.first[i] = (rd_u32)lo; T.last[i] = (rd_u32)hi;
    T.parent[(L & RD_BIN_LEAF) ? (rd_u32)(n - 1) + (L & ~RD_BIN_LEAF) : L] = (rd_u32)i;
    T.parent[(R & RD_BIN_LEAF) ? (rd_u32)(n - 1) + (R & ~RD_BIN_LEAF) : R] = (rd_u32)i;
    if (i == 0) T.parent[0] = RD_INVALID;
}

RD_HD rd_u32 rd_bin_slot(const RdBinTree& T, rd_u32 ref)
{
    return (ref & RD_BIN_LEAF) ? (T.n - 1u) + (ref & ~RD_BIN_LEAF) : ref;
}

RD_HD float rd_half_area(const RdF4& l, const RdF4& h)
{
    const float ex = h.x - l.x, ey = h.y - l.y, ez = h.z - l.z;
    return ex * ey + ey * ez + ez * ex;
}

// Optimal 8-wide layout of the binary topology by dynamic programming (Ylitie, Karras & Laine 2017,
// section 4.2). cost(n, k) = cheapest way to present subtree n to its parent in at most k child slots:
//   cost(n, 1) = min( leaf slot: A(n) * prims * c_prim   (prims <= RD_LEAF_MAX),
//                     own wide node: A(n) * c_node + dist(n, 8) )
//   cost(n, k) = min( dist(n, k), cost(n, k - 1) ),   dist(n, j) = min_i cost(left, i) + cost(right, j - i)
// dec[0] = 0: leaf slot, else the split i of dist(n, 8) used when n becomes a wide node;
// dec[k-1] (k >= 2) = split i of dist(n, k), or 0 for "no better than k - 1 slots".
RD_HD void rd_dp_node(const RdBinTree& T, rd_u32 cur, const RdF4& l, const RdF4& h)
{
    const float inf = 3.0e38f;
    float c[2][8];
    const rd_u32 kid[2] = {T.left[cur], T.right[cur]};
    for (int s = 0; s < 2; ++s) {
        if (kid[s] & RD_BIN_LEAF) {
            const rd_u32 b = rd_bin_slot(T, kid[s]);
            const float a = rd_half_area(T.lo[b], T.hi[b]) * RD_SAH_PRIM;
            for (int k = 0; k < 8; ++k) c[s][k] = a;
        } else {
#ifdef __CUDA_ARCH__
            const volatile float* src = (const volatile float*)(T.cost + 8u * (size_t)kid[s]);
#else
            const float* src = T.cost + 8u * (size_t)kid[s];
#endif
            for (int k = 0; k < 8; ++k) c[s][k] = src[k];
        }
    }
    float dist[9]; rd_u8 split[9];
    for (int j = 2; j <= 8; ++j) {
        float best = inf; rd_u8 bi = 1;
        for (int i = 1; i < j; ++i) {
            const float v = c[0][i - 1] + c[1][j - i - 1];
            if (v < best) { best = v; bi = (rd_u8)i; }
        }
        dist[j] = best; split[j] = bi;
    }
    const float A = rd_half_area(l, h);
    const rd_u32 prims = T.count[cur];
    const float cLeaf = prims <= RD_LEAF_MAX ? A * (float)prims * RD_SAH_PRIM : inf;
    const float cNode = A * RD_SAH_NODE + dist[8];
    float* outc = T.cost + 8u * (size_t)cur;
    rd_u8* outd = T.dec + 8u * (size_t)cur;
    float prev = cLeaf <= cNode ? cLeaf : cNode;
    outc[0] = prev; outd[0] = cLeaf <= cNode ? (rd_u8)0 : split[8];
    for (int k = 2; k <= 8; ++k) {
        if (dist[k] < prev) { prev = dist[k]; outd[k - 1] = split[k]; }
        else outd[k - 1] = 0;
        outc[k - 1] = prev;
    }
}

// Refit: the thread of leaf `pos` climbs; the second arrival at a node merges its children.
RD_HD void rd_refit_from_leaf(const RdBinTree& T, rd_u32 pos)
{
    rd_u32 cur = T.parent[(T.n - 1u) + pos];
    while (cur != RD_INVALID) {
        RD_THREADFENCE();
        if (RD_ATOMIC_ADD_U32(&T.arrive[cur], 1u) == 0u) return;
        RD_THREADFENCE();
        const rd_u32 a = rd_bin_slot(T, T.left[cur]), b = rd_bin_slot(T, T.right[cur]);
#ifdef __CUDA_ARCH__
        const volatile RdF4* vlo = (const volatile RdF4*)T.lo;
        const volatile RdF4* vhi = (const volatile RdF4*)T.hi;
        RdF4 la, lb, ha, hb;
        la.x = vlo[a].x; la.y = vlo[a].y; la.z = vlo[a].z; lb.x = vlo[b].x; lb.y = vlo[b].y; lb.z = vlo[b].z;
        ha.x = vhi[a].x; ha.y = vhi[a].y; ha.z = vhi[a].z; hb.x = vhi[b].x; hb.y = vhi[b].y; hb.z = vhi[b].z;
#else
        const RdF4 la = T.lo[a], lb = T.lo[b], ha = T.hi[a], hb = T.hi[b];
#endif
        RdF4 l, h;
        l.x = fminf(la.x, lb.x); l.y = fminf(la.y, lb.y); l.z = fminf(la.z, lb.z); l.w = 0.0f;
        h.x = fmaxf(ha.x, hb.x); h.y = fmaxf(ha.y, hb.y); h.z = fmaxf(ha.z, hb.z); h.w = 0.0f;
        T.lo[cur] = l; T.hi[cur] = h;
        {
#ifdef __CUDA_ARCH__
            const volatile rd_u32* vc = (const volatile rd_u32*)T.count;
#else
            const rd_u32* vc = T.count;
#endif
            const rd_u32 L = T.left[cur], R = T.right[cur];
            T.count[cur] = ((L & RD_BIN_LEAF) ? 1u : vc[L]) + ((R & RD_BIN_LEAF) ? 1u : vc[R]);
        }
        if (T.cost) rd_dp_node(T, cur, l, h);
        cur = T.parent[cur];
    }
}

RD_HD rd_u32 rd_bin_count(const RdBinTree& T, rd_u32 ref)
{
    return (ref & RD_BIN_LEAF) ? 1u : T.count[ref];
}

struct RdCollapseItem { rd_u32 wide, bin; };   // wide node index <- binary subtree reference

struct RdCollapseOut {
    RdNode8* nodes;          // wide nodes
    rd_u32* leafPrim;        // primitive indices in leaf-record order
    rd_u32* counters;        // [0] nodes allocated, [1] leaf records allocated, [2] next-level items
    RdCollapseItem* next;    // next-level work items
};

// smallest biased exponent byte e with extent / 2^(e-127) <= 255
RD_HD rd_u32 rd_grid_exponent(float extent)
{
    if (!(extent > 0.0f)) return 1u;
    const float step = extent * (1.0f / 255.0f);
    rd_u32 bits = RD_F2U(step);
    rd_u32 e = (bits >> 23) & 0xffu;
    if (bits & 0x7fffffu) e += 1u;      // round the step up to a power of two
    if (e < 1u) e = 1u;
    if (e > 254u) e = 254u;
    // guard against the rounding of extent * (1/255)
    while (e < 254u && extent / RD_U2F(e << 23) > 255.0f) e += 1u;
    return e;
}

// One wide node: pull up to 8 binary subtrees under `item.bin` (greedy by surface area), place
// them in octant-keyed slots, quantise their boxes, allocate children / leaf records, emit the
// next level's work items.
RD_HD void rd_collapse_item(const RdBinTree& T, const RdCollapseOut& O, RdCollapseItem item)
{
    rd_u32 ch[8];
    int n = 0;
    const rd_u32 root = item.bin;
    RdF4 nlo, nhi;
    {
        const rd_u32 s = rd_bin_slot(T, root);
        nlo = T.lo[s]; nhi = T.hi[s];
    }
    if ((root & RD_BIN_LEAF) || rd_bin_count(T, root) <= RD_LEAF_MAX) {
        ch[n++] = root;              // tiny mesh: the root itself is the only (leaf) child
    } else {
        if (T.cost && T.dec[8u * (size_t)root]) {
            // follow the decisions of the dynamic program from "root as a wide node with 8 slots"
            rd_u32 sref[16]; rd_u8 sbud[16]; int sp = 0;
            const rd_u8 k0 = T.dec[8u * (size_t)root];
            sref[sp] = T.right[root]; sbud[sp++] = (rd_u8)(8 - k0);
            sref[sp] = T.left[root]; sbud[sp++] = k0;
            while (sp > 0) {
                const rd_u32 r = sref[--sp]; rd_u8 bud = sbud[sp];
                if (r & RD_BIN_LEAF) { ch[n++] = r; continue; }
                const rd_u8* d = T.dec + 8u * (size_t)r;
                while (bud > 1 && d[bud - 1] == 0) --bud;
                if (bud == 1) { ch[n++] = r; continue; }
                const rd_u8 i = d[bud - 1];
                sref[sp] = T.right[r]; sbud[sp++] = (rd_u8)(bud - i);
                sref[sp] = T.left[r]; sbud[sp++] = i;
            }
        } else {
            ch[n++] = T.left[root]; ch[n++] = T.right[root];
            while (n < 8) {   // greedy by surface area (no cost table)
                int best = -1; float bestArea = -1.0f;
                for (int k = 0; k < n; ++k) {
                    if ((ch[k] & RD_BIN_LEAF) || rd_bin_count(T, ch[k]) <= RD_LEAF_MAX) continue;
                    const float area = rd_half_area(T.lo[ch[k]], T.hi[ch[k]]);
                    if (area > bestArea) { bestArea = area; best = k; }
                }
                if (best < 0) break;
                const rd_u32 b = ch[best];
                ch[best] = T.left[b]; ch[n++] = T.right[b];
            }
        }
    }
    // octant-keyed slot assignment (greedy): slot s prefers children displaced towards
    // (bit2 ? +x : -x, bit1 ? +y : -y, bit0 ? +z : -z) from the node centre.
    const float ccx = 0.5f * (nlo.x + nhi.x), ccy = 0.5f * (nlo.y + nhi.y), ccz = 0.5f * (nlo.z + nhi.z);
    float ox[8], oy[8], oz[8];
    for (int k = 0; k < n; ++k) {
        const rd_u32 s = rd_bin_slot(T, ch[k]);
        const RdF4 l = T.lo[s], h = T.hi[s];
        ox[k] = 0.5f * (l.x + h.x) - ccx; oy[k] = 0.5f * (l.y + h.y) - ccy; oz[k] = 0.5f * (l.z + h.z) - ccz;
    }
    int slotOf[8]; rd_u32 childAt[8];
    for (int s = 0; s < 8; ++s) childAt[s] = RD_INVALID;
    for (int k = 0; k < 8; ++k) slotOf[k] = -1;
    for (int round = 0; round < n; ++round) {
        int bk = -1, bs = -1; float bscore = -3.0e38f;
        for (int k = 0; k < n; ++k) {
            if (slotOf[k] >= 0) continue;
            for (int s = 0; s < 8; ++s) {
                if (childAt[s] != RD_INVALID) continue;
                const float score = ((s & 4) ? ox[k] : -ox[k]) + ((s & 2) ? oy[k] : -oy[k]) + ((s & 1) ? oz[k] : -oz[k]);
                if (score > bscore) { bscore = score; bk = k; bs = s; }
            }
        }
        if (bk < 0) {   // every remaining score is NaN (non-finite boxes): any free slot will do, nothing may index with -1
            for (int k = 0; k < n && bk < 0; ++k) if (slotOf[k] < 0) bk = k;
            for (int q = 0; q < 8 && bs < 0; ++q) if (childAt[q] == RD_INVALID) bs = q;
        }
        slotOf[bk] = bs; childAt[bs] = ch[bk];
    }
#if RD_NODE_F32
    rd_u32 imask = 0, nInner = 0, nLeafPrims = 0;
    for (int s = 0; s < 8; ++s) {
        if (childAt[s] == RD_INVALID) continue;
        const rd_u32 c = childAt[s];
        if (!(c & RD_BIN_LEAF) && rd_bin_count(T, c) > RD_LEAF_MAX) { imask |= 1u << s; nInner++; }
        else nLeafPrims += rd_bin_count(T, c);
    }
    const rd_u32 childBase = nInner ? RD_ATOMIC_ADD_U32(&O.counters[0], nInner) : 0u;
    const rd_u32 primBase = nLeafPrims ? RD_ATOMIC_ADD_U32(&O.counters[1], nLeafPrims) : 0u;
    const rd_u32 nextBase = nInner ? RD_ATOMIC_ADD_U32(&O.counters[2], nInner) : 0u;
    RdNode8 node;
    for (int k = 0; k < 16; ++k) { node.w[k].x = 0u; node.w[k].y = 0u; node.w[k].z = 0u; node.w[k].w = 0u; }
    float* planes = (float*)&node;    // float index = byte offset / 4
    rd_u32 leafNib = 0, innerNib = 0, leafOff = 0, innerRank = 0;
    for (int s = 0; s < 8; ++s) {
        float bl[3] = {3.0e38f, 3.0e38f, 3.0e38f}, bh[3] = {-3.0e38f, -3.0e38f, -3.0e38f};   // empty slot: inverted box
        if (childAt[s] != RD_INVALID) {
            const rd_u32 c = childAt[s];
            const rd_u32 slotBox = rd_bin_slot(T, c);
            const RdF4 l = T.lo[slotBox], h = T.hi[slotBox];
            // The boxes are the exact bounds of the triangles; the triangle test has rounding of its own (a ray can hit a
            // triangle numerically a few ulps outside its exact extent), so every child box grows by 4 ulps of its
            // coordinates. A parent's extreme coordinate is >= its children's, so padded children stay inside padded parents.
            const float pad = 4.76837158e-7f;
            bl[0] = l.x - pad * fabsf(l.x); bl[1] = l.y - pad * fabsf(l.y); bl[2] = l.z - pad * fabsf(l.z);
            bh[0] = h.x + pad * fabsf(h.x); bh[1] = h.y + pad * fabsf(h.y); bh[2] = h.z + pad * fabsf(h.z);
            if (imask & (1u << s)) {
                innerNib |= 1u << (4 * s);
                RdCollapseItem nx; nx.wide = childBase + innerRank; nx.bin = c;
                O.next[nextBase + innerRank] = nx;
                innerRank++;
            } else {
                const rd_u32 cnt = rd_bin_count(T, c);
                leafNib |= ((1u << cnt) - 1u) << (4 * s);
                // the <= RD_LEAF_MAX primitives of the subtree, left to right
                rd_u32 walk[RD_LEAF_MAX + 1]; int wp = 0; rd_u32 j = 0;
                walk[wp++] = c;
                while (wp > 0) {
                    const rd_u32 r = walk[--wp];
                    if (r & RD_BIN_LEAF) { O.leafPrim[primBase + leafOff + j] = T.prim[r & ~RD_BIN_LEAF]; ++j; }
                    else { walk[wp++] = T.right[r]; walk[wp++] = T.left[r]; }
                }
                leafOff += cnt;
            }
        }
        planes[RD_NODE_OFF_LOX / 4 + s] = bl[0]; planes[RD_NODE_OFF_LOY / 4 + s] = bl[1]; planes[RD_NODE_OFF_LOZ / 4 + s] = bl[2];
        planes[RD_NODE_OFF_HIX / 4 + s] = bh[0]; planes[RD_NODE_OFF_HIY / 4 + s] = bh[1]; planes[RD_NODE_OFF_HIZ / 4 + s] = bh[2];
    }
    node.w[0].x = childBase; node.w[0].y = primBase; node.w[0].z = leafNib; node.w[0].w = imask;
    node.w[1].x = innerNib;
    node.w[8].x = RD_F2U(nlo.x); node.w[8].y = RD_F2U(nlo.y); node.w[8].z = RD_F2U(nlo.z);
    node.w[9].x = RD_F2U(nhi.x); node.w[9].y = RD_F2U(nhi.y); node.w[9].z = RD_F2U(nhi.z);
#else
    // quantisation grid
    const rd_u32 ex = rd_grid_exponent(nhi.x - nlo.x), ey = rd_grid_exponent(nhi.y - nlo.y),
                 ez = rd_grid_exponent(nhi.z - nlo.z);
    const float isx = 1.0f / RD_U2F(ex << 23), isy = 1.0f / RD_U2F(ey << 23), isz = 1.0f / RD_U2F(ez << 23);
    rd_u32 imask = 0, nInner = 0, nLeafPrims = 0;
    for (int s = 0; s < 8; ++s) {
        if (childAt[s] == RD_INVALID) continue;
        const rd_u32 c = childAt[s];
        if (!(c & RD_BIN_LEAF) && rd_bin_count(T, c) > RD_LEAF_MAX) { imask |= 1u << s; nInner++; }
        else nLeafPrims += rd_bin_count(T, c);
    }
    const rd_u32 childBase = nInner ? RD_ATOMIC_ADD_U32(&O.counters[0], nInner) : 0u;
    const rd_u32 primBase = nLeafPrims ? RD_ATOMIC_ADD_U32(&O.counters[1], nLeafPrims) : 0u;
    const rd_u32 nextBase = nInner ? RD_ATOMIC_ADD_U32(&O.counters[2], nInner) : 0u;
    rd_u8 q[6][8];
    rd_u32 leafOff = 0, innerRank = 0, leafNib = 0, innerNib = 0;
    for (int s = 0; s < 8; ++s) {
        for (int a = 0; a < 6; ++a) q[a][s] = 0;
        if (childAt[s] == RD_INVALID) { q[0][s] = q[1][s] = q[2][s] = 255; continue; }   // inverted box
        const rd_u32 c = childAt[s];
        const rd_u32 slotBox = rd_bin_slot(T, c);
        const RdF4 l = T.lo[slotBox], h = T.hi[slotBox];
        float v;
        v = floorf((l.x - nlo.x) * isx); q[0][s] = (rd_u8)(v < 0.0f ? 0.0f : (v > 255.0f ? 255.0f : v));
        v = floorf((l.y - nlo.y) * isy); q[1][s] = (rd_u8)(v < 0.0f ? 0.0f : (v > 255.0f ? 255.0f : v));
        v = floorf((l.z - nlo.z) * isz); q[2][s] = (rd_u8)(v < 0.0f ? 0.0f : (v > 255.0f ? 255.0f : v));
        v = ceilf((h.x - nlo.x) * isx); q[3][s] = (rd_u8)(v < 0.0f ? 0.0f : (v > 255.0f ? 255.0f : v));
        v = ceilf((h.y - nlo.y) * isy); q[4][s] = (rd_u8)(v < 0.0f ? 0.0f : (v > 255.0f ? 255.0f : v));
        v = ceilf((h.z - nlo.z) * isz); q[5][s] = (rd_u8)(v < 0.0f ? 0.0f : (v > 255.0f ? 255.0f : v));
        if (imask & (1u << s)) {
            innerNib |= 1u << (4 * s);
            RdCollapseItem nx; nx.wide = childBase + innerRank; nx.bin = c;
            O.next[nextBase + innerRank] = nx;
            innerRank++;
        } else {
            const rd_u32 cnt = rd_bin_count(T, c);
            leafNib |= ((1u << cnt) - 1u) << (4 * s);
            // the <= RD_LEAF_MAX primitives of the subtree, left to right
            rd_u32 walk[RD_LEAF_MAX + 1]; int wp = 0; rd_u32 j = 0;
            walk[wp++] = c;
            while (wp > 0) {
                const rd_u32 r = walk[--wp];
                if (r & RD_BIN_LEAF) { O.leafPrim[primBase + leafOff + j] = T.prim[r & ~RD_BIN_LEAF]; ++j; }
                else { walk[wp++] = T.right[r]; walk[wp++] = T.left[r]; }
            }
            leafOff += cnt;
        }
    }
    RdNode8 node;
    node.w[0].x = RD_F2U(nlo.x); node.w[0].y = RD_F2U(nlo.y); node.w[0].z = RD_F2U(nlo.z);
    node.w[0].w = ex | (ey << 8) | (ez << 16) | (imask << 24);
    node.w[1].x = childBase; node.w[1].y = primBase;
#define RD_PACK4(a, o) ((rd_u32)(a)[o] | ((rd_u32)(a)[o + 1] << 8) | ((rd_u32)(a)[o + 2] << 16) | ((rd_u32)(a)[o + 3] << 24))
    node.w[1].z = leafNib; node.w[1].w = innerNib;
    node.w[2].x = RD_PACK4(q[0], 0); node.w[2].y = RD_PACK4(q[0], 4); node.w[2].z = RD_PACK4(q[1], 0); node.w[2].w = RD_PACK4(q[1], 4);
    node.w[3].x = RD_PACK4(q[2], 0); node.w[3].y = RD_PACK4(q[2], 4); node.w[3].z = RD_PACK4(q[3], 0); node.w[3].w = RD_PACK4(q[3], 4);
    node.w[4].x = RD_PACK4(q[4], 0); node.w[4].y = RD_PACK4(q[4], 4); node.w[4].z = RD_PACK4(q[5], 0); node.w[4].w = RD_PACK4(q[5], 4);
#undef RD_PACK4
#endif
    O.nodes[item.wide] = node;
}

// Leaf record r <- triangle leafPrim[r] of an indexed mesh (reference vertex/index arrays:
// radiance.h:32-36, Vec3 12 B, Triangle 3 x u32)
RD_HD void rd_gather_tri(const float* verts, const rd_u32* idx, const rd_u32* leafPrim, RdTri* out, rd_u32 r)
{
    const rd_u32 p = leafPrim[r];
    const rd_u32 i0 = idx[3 * p], i1 = idx[3 * p + 1], i2 = idx[3 * p + 2];
    RdTri t;
    t.v0.x = verts[3 * i0]; t.v0.y = verts[3 * i0 + 1]; t.v0.z = verts[3 * i0 + 2]; t.v0.w = RD_U2F(p);
    t.v1.x = verts[3 * i1]; t.v1.y = verts[3 * i1 + 1]; t.v1.z = verts[3 * i1 + 2]; t.v1.w = 0.0f;
    t.v2.x = verts[3 * i2]; t.v2.y = verts[3 * i2 + 1]; t.v2.z = verts[3 * i2 + 2]; t.v2.w = 0.0f;
    out[r] = t;
}

// primitive box of triangle p
RD_HD void rd_prim_box_tri(const float* verts, const rd_u32* idx, rd_u32 p, RdF4* lo, RdF4* hi)
{
    const rd_u32 i0 = idx[3 * p], i1 = idx[3 * p + 1], i2 = idx[3 * p + 2];
    RdF4 l, h;
    l.x = fminf(fminf(verts[3 * i0], verts[3 * i1]), verts[3 * i2]);
    l.y = fminf(fminf(verts[3 * i0 + 1], verts[3 * i1 + 1]), verts[3 * i2 + 1]);
    l.z = fminf(fminf(verts[3 * i0 + 2], verts[3 * i1 + 2]), verts[3 * i2 + 2]);
    h.x = fmaxf(fmaxf(verts[3 * i0], verts[3 * i1]), verts[3 * i2]);
    h.y = fmaxf(fmaxf(verts[3 * i0 + 1], verts[3 * i1 + 1]), verts[3 * i2 + 1]);
    h.z = fmaxf(fmaxf(verts[3 * i0 + 2], verts[3 * i1 + 2]), verts[3 * i2 + 2]);
    l.w = 0.0f; h.w = 0.0f;
    lo[p] = l; hi[p] = h;
}

// Row-major 4x4 inverse through 2x2 sub-determinants, evaluated in double and rounded once.
// The reference inverts the instance matrix in float for every ray at every instance visit
// (radiance/shader/radiance.cl:166-167, math.cl:56-183); here it happens once per instance at
// top-level build time. Returns false if the matrix is singular.
RD_HD bool rd_inverse4x4(const float* m, float* out)
{
    const double a00 = m[0], a01 = m[1], a02 = m[2], a03 = m[3], a10 = m[4], a11 = m[5], a12 = m[6], a13 = m[7];
    const double a20 = m[8], a21 = m[9], a22 = m[10], a23 = m[11], a30 = m[12], a31 = m[13], a32 = m[14], a33 = m[15];
    const double s0 = a00 * a11 - a10 * a01, s1 = a00 * a12 - a10 * a02, s2 = a00 * a13 - a10 * a03;
    const double s3 = a01 * a12 - a11 * a02, s4 = a01 * a13 - a11 * a03, s5 = a02 * a13 - a12 * a03;
    const double c5 = a22 * a33 - a32 * a23, c4 = a21 * a33 - a31 * a23, c3 = a21 * a32 - a31 * a22;
    const double c2 = a20 * a33 - a30 * a23, c1 = a20 * a32 - a30 * a22, c0 = a20 * a31 - a30 * a21;
    const double det = s0 * c5 - s1 * c4 + s2 * c3 + s3 * c2 - s4 * c1 + s5 * c0;
    if (det == 0.0) return false;
    const double r = 1.0 / det;
    out[0] = (float)((a11 * c5 - a12 * c4 + a13 * c3) * r);
    out[1] = (float)((-a01 * c5 + a02 * c4 - a03 * c3) * r);
    out[2] = (float)((a31 * s5 - a32 * s4 + a33 * s3) * r);
    out[3] = (float)((-a21 * s5 + a22 * s4 - a23 * s3) * r);
    out[4] = (float)((-a10 * c5 + a12 * c2 - a13 * c1) * r);
    out[5] = (float)((a00 * c5 - a02 * c2 + a03 * c1) * r);
    out[6] = (float)((-a30 * s5 + a32 * s2 - a33 * s1) * r);
    out[7] = (float)((a20 * s5 - a22 * s2 + a23 * s1) * r);
    out[8] = (float)((a10 * c4 - a11 * c2 + a13 * c0) * r);
    out[9] = (float)((-a00 * c4 + a01 * c2 - a03 * c0) * r);
    out[10] = (float)((a30 * s4 - a31 * s2 + a33 * s0) * r);
    out[11] = (float)((-a20 * s4 + a21 * s2 - a23 * s0) * r);
    out[12] = (float)((-a10 * c3 + a11 * c1 - a12 * c0) * r);
    out[13] = (float)((a00 * c3 - a01 * c1 + a02 * c0) * r);
    out[14] = (float)((-a30 * s3 + a31 * s1 - a32 * s0) * r);
    out[15] = (float)((a20 * s3 - a21 * s1 + a22 * s0) * r);
    return true;
}

// World-space box of an instance: transform the 8 corners of the bottom-level root box
// (reference: radiance/src/bvh.cpp:370-401).
RD_HD void rd_instance_world_box(const float* fwd, RdF4 blo, RdF4 bhi, RdF4* lo, RdF4* hi)
{
    RdF4 l, h;
    l.x = l.y = l.z = 3.402823466e+38f; h.x = h.y = h.z = -3.402823466e+38f; l.w = h.w = 0.0f;
    for (int c = 0; c < 8; ++c) {
        const float x = (c & 1) ? bhi.x : blo.x, y = (c & 2) ? bhi.y : blo.y, z = (c & 4) ? bhi.z : blo.z;
        const float wx = fwd[0] * x + fwd[1] * y + fwd[2] * z + fwd[3];
        const float wy = fwd[4] * x + fwd[5] * y + fwd[6] * z + fwd[7];
        const float wz = fwd[8] * x + fwd[9] * y + fwd[10] * z + fwd[11];
        l.x = fminf(l.x, wx); l.y = fminf(l.y, wy); l.z = fminf(l.z, wz);
        h.x = fmaxf(h.x, wx); h.y = fmaxf(h.y, wy); h.z = fmaxf(h.z, wz);
    }
    // pad by a few ulps of the magnitude: the instance-space traversal re-derives the ray with
    // rounded arithmetic, the world box must stay conservative
    const float px = 4e-7f * fmaxf(fabsf(l.x), fabsf(h.x)), py = 4e-7f * fmaxf(fabsf(l.y), fabsf(h.y)),
                pz = 4e-7f * fmaxf(fabsf(l.z), fabsf(h.z));
    l.x -= px; l.y -= py; l.z -= pz; h.x += px; h.y += py; h.z += pz;
    *lo = l; *hi = h;
}

#endif
