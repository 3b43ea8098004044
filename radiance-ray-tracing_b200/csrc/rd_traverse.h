// rd_traverse.h -- ray / wide-BVH traversal and ray/triangle intersection (one ray per thread).
//
// Replaces radiance/shader/radiance.cl:41-251 (intersectBot, intersectTop, intersectAABB,
// intersectTriangle). Same observable semantics:
//   * a candidate is accepted iff the triangle test passes and  Tmin < t < Tmax  and  t < best
//     (all strict, radiance.cl:90-91); two-sided triangles, inclusive edges, det == 0 rejected,
//     t > 0 (radiance.cl:221-240)
//   * the triangle test evaluates the same Moeller-Trumbore expressions in the same order with
//     separately rounded IEEE multiplies and adds, so t / barycentrics are bit-identical to the
//     reference compiled without contraction
//   * instance rays are transformed with w = 1 / w = 0 and NOT renormalised, so t is shared
//     between world and object space (radiance.cl:161-169)
//   * the candidate callback runs after the hit is committed and may only stop the traversal
//     (any-hit, radiance.cl:93-101)
// Different by design: 8-wide nodes, front-to-back order by ray octant, boxes culled
// against the current best t, precomputed instance inverse, two-level walk on one stack.
//
// The file compiles in three places: nvcc (library kernels), NVRTC (inlined into user programs)
// and g++ (tests/hostemu only, never the product).
//
// Measured and dropped in round 1 (profiles/README_r1.md has the numbers; the code is in the history before r2):
// warp-cooperative triangle phase (RD_COOP_TRI, -6 %), pending group in registers (RD_PEND_REG, -13 %), software
// prefetch (RD_PREFETCH, -2 %), triangle postponing by warp vote (RD_FLUSH_DIV, -3 %), PRMT / hybrid dequantisation
// (RD_DEQUANT_I2F 0 / 2, -3 %).
// Measured and dropped in round 2 (profiles/README_r2.md, code in the history at commit "Deferred multi-stream small ... RD_TRI_WAIT"):
// lanes that hold pending triangles WAIT until k/8 of the warp hold some (RD_TRI_WAIT, warp-synchronous loop with votes
// between the phases; the triangle phase is 40 % of the issued instructions at 5 of 32 lanes on the 10 M-triangle field):
// k = 2: 10 M 2991 -> 3050 Mrays/s, C2 4520 -> 4290; k >= 6: -7 % .. -20 % -- what the fuller triangle phase saves, the
// waiting lanes lose in extra node iterations. 256-byte nodes with float boxes: see RD_NODE_F32 in rd_types.h.
// The first 4 / 8 / 12 stack entries of every lane in shared memory ([entry][thread], the rest thread-local; RD_SMEM_STACK,
// code in the history at commit "north_star ablations with numbers"): 10 M 3022 -> 2918 / 2935 / 2923, C2 4665 -> 4561 / 4581 /
// 4581 Mrays/s, same image -- the thread-local stack lives in the same L1 and needs no index arithmetic per access.
#ifndef RD_TRAVERSE_H
#define RD_TRAVERSE_H
#include "rd_types.h"

#ifdef __CUDA_ARCH__
#define RD_MUL(a, b) __fmul_rn((a), (b))
#define RD_ADD(a, b) __fadd_rn((a), (b))
#define RD_SUB(a, b) __fsub_rn((a), (b))
#define RD_RCP(a) __frcp_rn(a)
#define RD_FMA(a, b, c) __fmaf_rn((a), (b), (c))
#define RD_FMIN(a, b) fminf((a), (b))
#define RD_FMAX(a, b) fmaxf((a), (b))
#define RD_POPC(x) __popc(x)
#define RD_CLZ(x) __clz((int)(x))
#define RD_FFS(x) __ffs((int)(x))
#define RD_F2U(f) __float_as_uint(f)
#define RD_U2F(u) __uint_as_float(u)
#define RD_PRMT(a, b, s) __byte_perm((a), (b), (s))
#else
#include <math.h>
#include <string.h>
// host build (unit tests): compiled with -ffp-contract=off so these stay separately rounded
#define RD_MUL(a, b) ((a) * (b))
#define RD_ADD(a, b) ((a) + (b))
#define RD_SUB(a, b) ((a) - (b))
#define RD_RCP(a) (1.0f / (a))
#define RD_FMA(a, b, c) fmaf((a), (b), (c))
#define RD_FMIN(a, b) fminf((a), (b))
#define RD_FMAX(a, b) fmaxf((a), (b))
#define RD_POPC(x) __builtin_popcount(x)
#define RD_CLZ(x) ((x) ? __builtin_clz(x) : 32)
#define RD_FFS(x) __builtin_ffs((int)(x))
static inline rd_u32 rd_f2u_(float f) { rd_u32 u; memcpy(&u, &f, 4); return u; }
static inline float rd_u2f_(rd_u32 u) { float f; memcpy(&f, &u, 4); return f; }
#define RD_F2U(f) rd_f2u_(f)
#define RD_U2F(u) rd_u2f_(u)
static inline rd_u32 rd_prmt_(rd_u32 a, rd_u32 b, rd_u32 s)
{
    const rd_u64 v = ((rd_u64)b << 32) | a;
    rd_u32 r = 0;
    for (int k = 0; k < 4; ++k) r |= (rd_u32)((v >> (8 * ((s >> (4 * k)) & 7u))) & 0xffu) << (8 * k);
    return r;
}
#define RD_PRMT(a, b, s) rd_prmt_((a), (b), (s))
#endif

#ifndef RD_RECONVERGE
#define RD_RECONVERGE 1
#endif
#ifndef RD_STACK_ENTRIES
#define RD_STACK_ENTRIES 48
#endif

struct RdHitRec {
    float t, b1, b2;     // t = RD_MISS_T when nothing was accepted
    rd_u32 prim, inst;   // primitive index inside the mesh, instanceID (caller order)
};

// RD_FAST_MT = 1 (ablation only): the same test with fused multiply-adds and the hardware reciprocal
// (about 2/3 of the instructions, +3.7 % Mrays/s on C2). Measured against the reference it moves
// barycentrics of grazing hits by up to 9e-4 and flips hit/miss on 2e-5 of the C3 primary rays, so the
// default keeps every operation separately rounded in the reference's order: t / barycentrics are then
// bit-identical to the reference built without contraction.
#ifndef RD_FAST_MT
#define RD_FAST_MT 0
#endif

#if RD_FAST_MT
RD_HD bool rd_intersect_triangle(float ox, float oy, float oz, float dx, float dy, float dz,
                                 float ax, float ay, float az, float bx, float by, float bz,
                                 float cx, float cy, float cz, float* t, float* b1, float* b2)
{
    const float e1x = bx - ax, e1y = by - ay, e1z = bz - az;
    const float e2x = cx - ax, e2y = cy - ay, e2z = cz - az;
    const float px = RD_FMA(dy, e2z, -dz * e2y), py = RD_FMA(dz, e2x, -dx * e2z), pz = RD_FMA(dx, e2y, -dy * e2x);
    const float det = RD_FMA(e1x, px, RD_FMA(e1y, py, e1z * pz));
    if (det == 0.0f) return false;
#ifdef __CUDA_ARCH__
    float inv;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(inv) : "f"(det));
#else
    const float inv = 1.0f / det;
#endif
    const float sx = ox - ax, sy = oy - ay, sz = oz - az;
    const float u = inv * RD_FMA(sx, px, RD_FMA(sy, py, sz * pz));
    const float qx = RD_FMA(sy, e1z, -sz * e1y), qy = RD_FMA(sz, e1x, -sx * e1z), qz = RD_FMA(sx, e1y, -sy * e1x);
    const float v = inv * RD_FMA(dx, qx, RD_FMA(dy, qy, dz * qz));
    const float tt = inv * RD_FMA(e2x, qx, RD_FMA(e2y, qy, e2z * qz));
    if (u < 0.0f || u > 1.0f) return false;
    if (v < 0.0f || u + v > 1.0f) return false;
    if (!(tt > 0.0f)) return false;
    *t = tt; *b1 = u; *b2 = v;
    return true;
}
#else
// Moeller-Trumbore with the reference's expression order (radiance/shader/radiance.cl:211-251).
RD_HD bool rd_intersect_triangle(float ox, float oy, float oz, float dx, float dy, float dz,
                                 float ax, float ay, float az, float bx, float by, float bz,
                                 float cx, float cy, float cz, float* t, float* b1, float* b2)
{
    const float e1x = RD_SUB(bx, ax), e1y = RD_SUB(by, ay), e1z = RD_SUB(bz, az);
    const float e2x = RD_SUB(cx, ax), e2y = RD_SUB(cy, ay), e2z = RD_SUB(cz, az);
    // p = d x e2
    const float px = RD_SUB(RD_MUL(dy, e2z), RD_MUL(dz, e2y));
    const float py = RD_SUB(RD_MUL(dz, e2x), RD_MUL(dx, e2z));
    const float pz = RD_SUB(RD_MUL(dx, e2y), RD_MUL(dy, e2x));
    const float det = RD_ADD(RD_ADD(RD_MUL(e1x, px), RD_MUL(e1y, py)), RD_MUL(e1z, pz));
    if (det == 0.0f) return false;
    const float inv = RD_RCP(det);
    const float sx = RD_SUB(ox, ax), sy = RD_SUB(oy, ay), sz = RD_SUB(oz, az);
    const float u = RD_MUL(inv, RD_ADD(RD_ADD(RD_MUL(sx, px), RD_MUL(sy, py)), RD_MUL(sz, pz)));
    // q = s x e1
    const float qx = RD_SUB(RD_MUL(sy, e1z), RD_MUL(sz, e1y));
    const float qy = RD_SUB(RD_MUL(sz, e1x), RD_MUL(sx, e1z));
    const float qz = RD_SUB(RD_MUL(sx, e1y), RD_MUL(sy, e1x));
    const float v = RD_MUL(inv, RD_ADD(RD_ADD(RD_MUL(dx, qx), RD_MUL(dy, qy)), RD_MUL(dz, qz)));
    const float tt = RD_MUL(inv, RD_ADD(RD_ADD(RD_MUL(e2x, qx), RD_MUL(e2y, qy)), RD_MUL(e2z, qz)));
    if (u < 0.0f || u > 1.0f) return false;
    if (v < 0.0f || RD_ADD(u, v) > 1.0f) return false;
    if (!(tt > 0.0f)) return false;
    *t = tt; *b1 = u; *b2 = v;
    return true;
}
#endif

// row-major 3x4 times (x,y,z,w) with the reference's left-to-right sum (math.cl:25-31)
RD_HD float rd_row_dot(const float* r, float x, float y, float z, float w)
{
    return RD_ADD(RD_ADD(RD_ADD(RD_MUL(r[0], x), RD_MUL(r[1], y)), RD_MUL(r[2], z)), RD_MUL(r[3], w));
}

RD_HD float rd_safe_rcp(float d)
{
    const float tiny = 1e-20f;
    float a = fabsf(d) > tiny ? d : (RD_F2U(d) >> 31 ? -tiny : tiny);
    return RD_RCP(a);   // correctly rounded whatever division mode the user program is compiled with
}

#ifdef __CUDA_ARCH__
#define RD_LOAD_U4(dst, ptr) { uint4 _v = __ldg((const uint4*)(ptr)); dst.x = _v.x; dst.y = _v.y; dst.z = _v.z; dst.w = _v.w; }
#define RD_LOAD_F4(dst, ptr) { float4 _v = __ldg((const float4*)(ptr)); dst.x = _v.x; dst.y = _v.y; dst.z = _v.z; dst.w = _v.w; }
#else
#define RD_LOAD_U4(dst, ptr) { dst = *(const RdU4*)(ptr); }
#define RD_LOAD_F4(dst, ptr) { dst = *(const RdF4*)(ptr); }
#endif

#if RD_NODE_F32
// ---------------------------------------------------------------------------------------------------------------
// Float nodes. Per-space ray constants: t(plane) = plane * i + n with n = -(o * i), one FFMA per plane. The near
// constants lean early and the far constants late by m = 2^-21 * (|o * i| + |i| * R), R = the largest |coordinate| of
// the structure being walked: that bounds the rounding of i, of o * i and of the FFMA itself for every plane of the
// structure, so the test is conservative (it may enter a box the exact ray just misses, never the reverse).
struct RdRayFrame {
    float ox, oy, oz, dx, dy, dz;   // the ray in this space (triangle test)
    float ix, iy, iz;               // 1/d with |d| clamped away from zero
    float nnx, nny, nnz;            // near-plane constants
    float nfx, nfy, nfz;            // far-plane constants
    rd_u32 offs;                    // byte offsets of the NEAR plane arrays inside a node: x | y << 8 | z << 16, and the
                                    // visiting-order key 7 - octant in the top byte (octant bit2 = x<0, bit1 = y<0, bit0 = z<0)
    rd_u32 selE;                    // PRMT selector that sorts the even-priority hit bytes (see rd_inner_priorities)
};

RD_HD void rd_frame_set(RdRayFrame& f, float ox, float oy, float oz, float dx, float dy, float dz, float maxAbs)
{
    f.ox = ox; f.oy = oy; f.oz = oz; f.dx = dx; f.dy = dy; f.dz = dz;
    f.ix = rd_safe_rcp(dx); f.iy = rd_safe_rcp(dy); f.iz = rd_safe_rcp(dz);
    const float k = 4.76837158e-7f;   // 2^-21
    const float bx = -(ox * f.ix), by = -(oy * f.iy), bz = -(oz * f.iz);
    const float mx = k * (fabsf(bx) + fabsf(f.ix) * maxAbs), my = k * (fabsf(by) + fabsf(f.iy) * maxAbs),
                mz = k * (fabsf(bz) + fabsf(f.iz) * maxAbs);
    f.nnx = bx - mx; f.nny = by - my; f.nnz = bz - mz;
    f.nfx = bx + mx; f.nfy = by + my; f.nfz = bz + mz;
    const rd_u32 nx = f.ix < 0.0f ? 1u : 0u, ny = f.iy < 0.0f ? 1u : 0u, nz = f.iz < 0.0f ? 1u : 0u;
    const rd_u32 octinv = 7u - ((nx << 2) | (ny << 1) | nz);
    f.offs = (nx ? RD_NODE_OFF_HIX : RD_NODE_OFF_LOX) | ((ny ? RD_NODE_OFF_HIY : RD_NODE_OFF_LOY) << 8) |
             ((nz ? RD_NODE_OFF_HIZ : RD_NODE_OFF_LOZ) << 16) | (octinv << 24);
    // priority p = slot ^ octinv (highest first). Even-priority byte k comes from slot (2k) ^ octinv: an even slot
    // (byte k ^ (octinv >> 1) of the even-slot word) when octinv is even, else the odd slot of the same pair.
    f.selE = (0x3210u ^ ((octinv >> 1) * 0x1111u)) | ((octinv & 1u) ? 0x4444u : 0u);
}

#define RD_OCTINV(f) ((f).offs >> 24)

// Tests the 8 child boxes of node n against the ray interval [tmin, tbest]. Returns the hit slots as nibbles:
// 0xf << 4s for every slot s whose box the ray enters.
RD_HD rd_u32 rd_node_test(const RdNode8* n, const RdRayFrame& f, float tmin, float tbest)
{
    const char* base = (const char*)n;
    const rd_u32 ox = f.offs & 0xffu, oy = (f.offs >> 8) & 0xffu, oz = (f.offs >> 16) & 0xffu;
    RdF4 nx[2], ny[2], nz[2], fx[2], fy[2], fz[2];
    RD_LOAD_F4(nx[0], base + ox); RD_LOAD_F4(nx[1], base + ox + 16);
    RD_LOAD_F4(ny[0], base + oy); RD_LOAD_F4(ny[1], base + oy + 16);
    RD_LOAD_F4(nz[0], base + oz); RD_LOAD_F4(nz[1], base + oz + 16);
    RD_LOAD_F4(fx[0], base + ((RD_NODE_OFF_LOX + RD_NODE_OFF_HIX) - ox)); RD_LOAD_F4(fx[1], base + ((RD_NODE_OFF_LOX + RD_NODE_OFF_HIX) - ox) + 16);
    RD_LOAD_F4(fy[0], base + ((RD_NODE_OFF_LOY + RD_NODE_OFF_HIY) - oy)); RD_LOAD_F4(fy[1], base + ((RD_NODE_OFF_LOY + RD_NODE_OFF_HIY) - oy) + 16);
    RD_LOAD_F4(fz[0], base + ((RD_NODE_OFF_LOZ + RD_NODE_OFF_HIZ) - oz)); RD_LOAD_F4(fz[1], base + ((RD_NODE_OFF_LOZ + RD_NODE_OFF_HIZ) - oz) + 16);
    rd_u32 acc = 0;
#define RD_CHILD(s, h, c) { \
        const float tn = RD_FMAX(RD_FMAX(RD_FMAX(RD_FMA(nx[h].c, f.ix, f.nnx), RD_FMA(ny[h].c, f.iy, f.nny)), RD_FMA(nz[h].c, f.iz, f.nnz)), tmin); \
        const float tf = RD_FMIN(RD_FMIN(RD_FMIN(RD_FMA(fx[h].c, f.ix, f.nfx), RD_FMA(fy[h].c, f.iy, f.nfy)), RD_FMA(fz[h].c, f.iz, f.nfz)), tbest); \
        if (tn <= tf) acc |= 0xfu << (4 * (s)); }
    RD_CHILD(0, 0, x) RD_CHILD(1, 0, y) RD_CHILD(2, 0, z) RD_CHILD(3, 0, w)
    RD_CHILD(4, 1, x) RD_CHILD(5, 1, y) RD_CHILD(6, 1, z) RD_CHILD(7, 1, w)
#undef RD_CHILD
    return acc;
}

#else
// ---------------------------------------------------------------------------------------------------------------
// 8-bit quantised nodes (round-1 form, kept for the A/B measurement).
struct RdRayFrame {   // per-space ray constants for the quantised box test
    float ox, oy, oz, dx, dy, dz;
    float ix, iy, iz;   // 1/d with |d| clamped away from zero
    rd_u32 octinv;      // 7 - octant; octant bit2=x<0, bit1=y<0, bit0=z<0
    rd_u32 selE;        // PRMT selector that sorts the even-priority hit bytes (see rd_inner_priorities)
};

RD_HD void rd_frame_set(RdRayFrame& f, float ox, float oy, float oz, float dx, float dy, float dz, float)
{
    f.ox = ox; f.oy = oy; f.oz = oz; f.dx = dx; f.dy = dy; f.dz = dz;
    f.ix = rd_safe_rcp(dx); f.iy = rd_safe_rcp(dy); f.iz = rd_safe_rcp(dz);
    rd_u32 oct = (f.ix < 0.0f ? 4u : 0u) | (f.iy < 0.0f ? 2u : 0u) | (f.iz < 0.0f ? 1u : 0u);
    f.octinv = 7u - oct;
    f.selE = (0x3210u ^ ((f.octinv >> 1) * 0x1111u)) | ((f.octinv & 1u) ? 0x4444u : 0u);
}
#define RD_OCTINV(f) ((f).octinv)

// Tests the 8 quantised child boxes of one node against the ray interval [tmin, tbest]. Returns the hit slots as
// nibbles: 0xf << 4s for every slot s whose box the ray enters.
// Plane distance of quantised coordinate c on one axis: t = (p + c*s - o) / d = c*a + b with a = s/d, b = (p - o)/d;
// the conversion of the byte runs on the XU pipe (I2F.U8), floor / ceil quantisation keeps the test conservative.
RD_HD rd_u32 rd_node_test(const RdU4& n0, const RdU4& n2, const RdU4& n3, const RdU4& n4,
                          const RdRayFrame& f, float tmin, float tbest)
{
    const float sx = RD_U2F((n0.w & 0xffu) << 23), sy = RD_U2F(((n0.w >> 8) & 0xffu) << 23),
                sz = RD_U2F(((n0.w >> 16) & 0xffu) << 23);
    const float Ax = sx * f.ix, Ay = sy * f.iy, Az = sz * f.iz;
    const float bx = (RD_U2F(n0.x) - f.ox) * f.ix, by = (RD_U2F(n0.y) - f.oy) * f.iy, bz = (RD_U2F(n0.z) - f.oz) * f.iz;
#define RD_PLANE(q, j, P, pb) RD_FMA((float)(((q) >> (8 * (j))) & 0xffu), P, pb)
    rd_u32 acc = 0;
#if defined(__CUDA_ARCH__)
#pragma unroll
#endif
    for (int h = 0; h < 2; ++h) {
        const rd_u32 qlx = h ? n2.y : n2.x, qly = h ? n2.w : n2.z, qlz = h ? n3.y : n3.x;
        const rd_u32 qhx = h ? n3.w : n3.z, qhy = h ? n4.y : n4.x, qhz = h ? n4.w : n4.z;
        const rd_u32 nx = f.ix < 0.0f ? qhx : qlx, fx = f.ix < 0.0f ? qlx : qhx;
        const rd_u32 ny = f.iy < 0.0f ? qhy : qly, fy = f.iy < 0.0f ? qly : qhy;
        const rd_u32 nz = f.iz < 0.0f ? qhz : qlz, fz = f.iz < 0.0f ? qlz : qhz;
#if defined(__CUDA_ARCH__)
#pragma unroll
#endif
        for (int j = 0; j < 4; ++j) {
            const float tn = RD_FMAX(RD_FMAX(RD_PLANE(nx, j, Ax, bx), RD_PLANE(ny, j, Ay, by)), RD_FMAX(RD_PLANE(nz, j, Az, bz), tmin));
            const float tf = RD_FMIN(RD_FMIN(RD_PLANE(fx, j, Ax, bx), RD_PLANE(fy, j, Ay, by)), RD_FMIN(RD_PLANE(fz, j, Az, bz), tbest));
            if (tn <= tf) acc |= 0xfu << (4 * (4 * h + j));
        }
    }
#undef RD_PLANE
    return acc;
}
#endif  // RD_NODE_F32

// hit nibbles of the inner slots -> priority bits 24..31 (bit 24 + p set <=> the child in slot p ^ octinv was hit)
RD_HD rd_u32 rd_inner_priorities(rd_u32 hitInner, const RdRayFrame& f)
{
    const rd_u32 E = hitInner & 0x0f0f0f0fu, O = (hitInner >> 4) & 0x0f0f0f0fu;   // byte k = slot 2k / slot 2k + 1, value 0 or 1
    const rd_u32 Ep = RD_PRMT(E, O, f.selE), Op = RD_PRMT(E, O, f.selE ^ 0x4444u); // byte k = priority 2k / 2k + 1
    // byte k (bit 8k) -> bit 24 + 2k (+1): every partial product lands on its own bit, the ones below bit 24 are masked off
    return ((Ep * 0x01041040u) | (Op * 0x02082080u)) & 0xff000000u;
}

// Traversal state of one ray over the two-level structure. begin() arms it, step() runs one
// iteration and returns false when the ray is finished. Kept as a struct so that a persistent kernel
// can interleave steps with ray fetching (csrc/rd_trace.cu); rd_traverse() below is the plain loop.
//
// One iteration has three phases that the lanes of a warp go through together:
//   1. node phase     every lane with an inner child pending tests one 8-wide node; the triangle
//                     records the test exposed are only NOTED in the lane's pending slot (thread-local memory:
//                     the round trip is what keeps ptxas from threading lanes straight into the triangle loop)
//   2. triangle phase every lane tests its pending triangles
//   3. pop phase      lanes without inner work take the next stack entry
// Top-level leaf records are instances and are entered at once.
struct RdStackEntry { rd_u32 x, y; };

#define RD_PEND_WORDS 2   // stack entries the pending slot occupies behind the stack
#define RD_STACK_ALLOC (RD_STACK_ENTRIES + RD_PEND_WORDS)

struct RdTraversal {
    RdStackEntry* stack;              // RD_STACK_ALLOC entries owned by the caller (thread-local memory);
                                      // kept outside so the scalar state below stays in registers
    RdStackEntry* pend;               // pending triangle group (the tail of the same allocation)
    int sp, blasSp, np;               // blasSp >= 0 while inside a bottom-level structure
    const RdNode8* nodes;
    const RdTri* tris;
    const RdInstance* curInst;
    RdRayFrame f;                     // ray in the current space (world or instance-local)
    float wox, woy, woz, wdx, wdy, wdz, tmin, tmax, tbest;
    RdStackEntry ng;                  // node group being worked on: x = childBase, y = priority bits 24..31 | imask
    RdHitRec hit;
    bool any;

    RD_HD void push(const RdScene* scene, RdStackEntry e)
    {
        if (sp < RD_STACK_ENTRIES) stack[sp++] = e;
        else if (scene->status) scene->status[0] = 1u;
    }
    RD_HD RdStackEntry pop() { return stack[--sp]; }

    RD_HD void begin(RdStackEntry* stackMemory, const RdScene* scene, float ox, float oy, float oz, float dx, float dy, float dz, float tmin_, float tmax_)
    {
        stack = stackMemory; pend = stackMemory + RD_STACK_ENTRIES; sp = 0; blasSp = -1; np = 0;
        nodes = scene->nodes; tris = 0; curInst = 0;
        wox = ox; woy = oy; woz = oz; wdx = dx; wdy = dy; wdz = dz; tmin = tmin_; tmax = tmax_;
        rd_frame_set(f, ox, oy, oz, dx, dy, dz, scene->maxAbs);
        // the interval is open at tmax in the reference (t < Tmax); boxes are culled with <= so tmax
        // itself is the first bound and strictness stays in the triangle accept test
        tbest = tmax_;
        any = false;
        hit.t = RD_MISS_T; hit.b1 = 0.0f; hit.b2 = 0.0f; hit.prim = RD_INVALID; hit.inst = RD_INVALID;
        ng.x = 0; ng.y = 0x80000000u;
    }

    // `onCandidate(hit)` runs after every accepted closer candidate (hit already committed) and
    // returns false to stop the ray (any-hit semantics).
    template <class OnCandidate>
    RD_HD bool step(const RdScene* scene, OnCandidate& onCandidate)
    {
        RdStackEntry tg; tg.x = 0; tg.y = 0;   // top-level leaf records (instances) to enter now: x = first record, y = dense bits
        // ---- 1. node phase
        if (ng.y > 0x00ffffffu) {
            const rd_u32 bit = 31u - (rd_u32)RD_CLZ(ng.y);
            const rd_u32 imask = ng.y & 0xffu;   // low byte of an inner group = imask
            ng.y &= ~(1u << bit);
            if (ng.y > 0x00ffffffu) push(scene, ng);
            const rd_u32 slot = (bit - 24u) ^ RD_OCTINV(f);
            const rd_u32 rel = (rd_u32)RD_POPC(imask & ~(0xffffffffu << slot));
            const RdNode8* n = nodes + (ng.x + rel);
            onCandidate.countNode();
#if RD_NODE_F32
            RdU4 n0, n1;
            RD_LOAD_U4(n0, &n->w[0]); RD_LOAD_U4(n1, &n->w[1]);
            const rd_u32 acc = rd_node_test(n, f, tmin, tbest);
            ng.x = n0.x; ng.y = rd_inner_priorities(acc & n1.x, f) | n0.w;
            const rd_u32 leafBits = acc & n0.z;
            const rd_u32 nodePrimBase = n0.y, nodeLeafNib = n0.z;
#else
            RdU4 n0, n1, n2, n3, n4;   // n1 = { childBase, primBase, leafNib, innerNib }
            RD_LOAD_U4(n0, &n->w[0]); RD_LOAD_U4(n1, &n->w[1]); RD_LOAD_U4(n2, &n->w[2]);
            RD_LOAD_U4(n3, &n->w[3]); RD_LOAD_U4(n4, &n->w[4]);
            const rd_u32 acc = rd_node_test(n0, n2, n3, n4, f, tmin, tbest);
            ng.x = n1.x; ng.y = rd_inner_priorities(acc & n1.w, f) | (n0.w >> 24);
            const rd_u32 leafBits = acc & n1.z;
            const rd_u32 nodePrimBase = n1.y, nodeLeafNib = n1.z;
#endif
            if (leafBits) {
                if (blasSp < 0) {
                    // top level: few nodes, few records -- densify the record bits so a parked group fits a stack entry
                    rd_u32 lb = leafBits, dense = 0;
                    while (lb) {
                        const rd_u32 low = lb & (0u - lb);
                        lb ^= low;
                        dense |= 1u << RD_POPC(nodeLeafNib & (low - 1u));
                    }
                    tg.x = nodePrimBase; tg.y = dense;
                } else {
                    // np indexes the slot on purpose: with a constant address ptxas forwards the stored values in registers and
                    // threads lanes straight from the node test into the triangle loop -- the warp then stops walking the
                    // phases together (measured -12 % on C2, -19 % at 10 M triangles, profiles/README_r2.md)
                    pend[2 * np].x = nodePrimBase; pend[2 * np].y = leafBits; pend[2 * np + 1].x = nodeLeafNib; ++np;
                }
            }
        } else if (ng.y) {
            tg = ng; ng.x = 0; ng.y = 0;   // a parked top-level instance group came off the stack
        }
        if (blasSp < 0) {
            if (tg.y) {
                // top level: the first record is an instance -> descend into its bottom-level structure
                const rd_u32 k = (rd_u32)RD_FFS(tg.y) - 1u;
                tg.y &= tg.y - 1u;
#if defined(RD_BOUNDS_CHECK) && RD_BOUNDS_CHECK && defined(__CUDA_ARCH__)
                if (tg.x + k >= scene->instanceCount) { if (scene->status) scene->status[2] = 1u; return false; }
#endif
                const rd_u32 ii = scene->leafInstances[tg.x + k];
                if (tg.y) push(scene, tg);
                if (ng.y > 0x00ffffffu) push(scene, ng);
                blasSp = sp;
                curInst = scene->instances + ii;
                const float* m = curInst->inv;
                const float lox = rd_row_dot(m + 0, wox, woy, woz, 1.0f), loy = rd_row_dot(m + 4, wox, woy, woz, 1.0f),
                            loz = rd_row_dot(m + 8, wox, woy, woz, 1.0f);
                const float ldx = rd_row_dot(m + 0, wdx, wdy, wdz, 0.0f), ldy = rd_row_dot(m + 4, wdx, wdy, wdz, 0.0f),
                            ldz = rd_row_dot(m + 8, wdx, wdy, wdz, 0.0f);
                rd_frame_set(f, lox, loy, loz, ldx, ldy, ldz, curInst->maxAbs);
                nodes = curInst->nodes; tris = curInst->tris;
                ng.x = 0; ng.y = 0x80000000u;
                return true;
            }
        } else if (np > 0) {
            // ---- 2. triangle phase
            RdStackEntry g; g.x = 0; g.y = 0;
            rd_u32 gnib = 0;
            while (np > 0 || g.y) {
                if (!g.y) { --np; g = pend[2 * np]; gnib = pend[2 * np + 1].x; }
                const rd_u32 low = g.y & (0u - g.y);
                g.y ^= low;
                const rd_u32 rec = g.x + (rd_u32)RD_POPC(gnib & (low - 1u));
#if defined(RD_BOUNDS_CHECK) && RD_BOUNDS_CHECK && defined(__CUDA_ARCH__)
                if (rec >= curInst->flags) { if (scene->status) scene->status[2] = 1u; continue; }   // flags = triangle records
#endif
                const RdTri* tr = tris + rec;
                RdF4 a, b, c;
                RD_LOAD_F4(a, &tr->v0); RD_LOAD_F4(b, &tr->v1); RD_LOAD_F4(c, &tr->v2);
                onCandidate.countTri();
                float t, b1, b2;
                if (rd_intersect_triangle(f.ox, f.oy, f.oz, f.dx, f.dy, f.dz, a.x, a.y, a.z, b.x, b.y, b.z,
                                          c.x, c.y, c.z, &t, &b1, &b2) &&
                    t > tmin && t < tmax &&
                    (t < hit.t || (t == hit.t && (curInst->instanceID < hit.inst ||
                                                  (curInst->instanceID == hit.inst && RD_F2U(a.w) < hit.prim))))) {
                    hit.t = t; hit.b1 = b1; hit.b2 = b2; hit.prim = RD_F2U(a.w); hit.inst = curInst->instanceID;
                    tbest = t; any = true;
                    if (!onCandidate(hit)) return false;
                }
            }
        }
        // ---- 3. pop phase
        if (ng.y <= 0x00ffffffu) {
            if (blasSp >= 0 && sp == blasSp) {
                // bottom-level structure exhausted (its pending group was tested above): back to world space
                blasSp = -1; nodes = scene->nodes;
                rd_frame_set(f, wox, woy, woz, wdx, wdy, wdz, scene->maxAbs);
            }
            if (sp == 0) return false;
            ng = pop();
        }
        return true;
    }
};

// Walks the two-level structure for one ray. Returns true iff something was accepted; `hit` holds
// the closest accepted candidate.
template <class OnCandidate>
RD_HD bool rd_traverse(const RdScene* scene, float ox, float oy, float oz, float dx, float dy, float dz,
                       float tmin, float tmax, RdHitRec& hit, OnCandidate& onCandidate)
{
    RdStackEntry stackMemory[RD_STACK_ALLOC];
    RdTraversal T;
#if defined(__CUDA_ARCH__) && RD_RECONVERGE
    // The walk leaves its loop from several places and ptxas puts no reconvergence point behind it:
    // lanes that finish early would run everything that follows (shading, the caller's bounce loop,
    // the next traceRay) on their own -- measured 2.2 of 32 lanes active in programs/path.cl. The lanes
    // that enter together therefore wait for each other here before returning.
    const unsigned together = __activemask();
#endif
    T.begin(stackMemory, scene, ox, oy, oz, dx, dy, dz, tmin, tmax);
    while (T.step(scene, onCandidate)) {}
#if defined(__CUDA_ARCH__) && RD_RECONVERGE
    __syncwarp(together);
#endif
    hit = T.hit;
    return T.any;
}

// callback interface: operator() = any-hit; countNode / countTri = optional fetch statistics
struct RdNoAnyHit {
    RD_HD bool operator()(const RdHitRec&) const { return true; }
    RD_HD void countNode() const {}
    RD_HD void countTri() const {}
};
struct RdFirstHit {
    RD_HD bool operator()(const RdHitRec&) const { return false; }
    RD_HD void countNode() const {}
    RD_HD void countTri() const {}
};
struct RdCountingHit {   // counters-on build of the same traversal (roofline bytes per ray)
    bool terminate; rd_u32 nodes, tris;
    RD_HD bool operator()(const RdHitRec&) const { return !terminate; }
    RD_HD void countNode() { nodes++; }
    RD_HD void countTri() { tris++; }
};

#endif
