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
// Different by design: 8-wide quantised nodes, front-to-back order by ray octant, boxes culled
// against the current best t, precomputed instance inverse, two-level walk on one stack.
//
// The file compiles in three places: nvcc (library kernels), NVRTC (inlined into user programs)
// and g++ (tests/hostemu only, never the product).
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
#define RD_LDG4(p) __ldg((const uint4*)(p))
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

struct RdRayFrame {   // per-space ray constants for the quantised box test
    float ox, oy, oz, dx, dy, dz;
    float ix, iy, iz;   // 1/d with |d| clamped away from zero
    rd_u32 octinv4;     // (7 - octant) replicated to 4 bytes; octant bit2=x<0, bit1=y<0, bit0=z<0
};

RD_HD float rd_safe_rcp(float d)
{
    const float tiny = 1e-30f;
    float a = fabsf(d) > tiny ? d : (RD_F2U(d) >> 31 ? -tiny : tiny);
    return 1.0f / a;
}

RD_HD void rd_frame_set(RdRayFrame& f, float ox, float oy, float oz, float dx, float dy, float dz)
{
    f.ox = ox; f.oy = oy; f.oz = oz; f.dx = dx; f.dy = dy; f.dz = dz;
    f.ix = rd_safe_rcp(dx); f.iy = rd_safe_rcp(dy); f.iz = rd_safe_rcp(dz);
    rd_u32 oct = (f.ix < 0.0f ? 4u : 0u) | (f.iy < 0.0f ? 2u : 0u) | (f.iz < 0.0f ? 1u : 0u);
    f.octinv4 = (7u - oct) * 0x01010101u;
}

// byte j of q -> float 1 + byte * 2^-15 without an int->float conversion: the byte is dropped into
// mantissa bits 8..15 of 1.0f by one PRMT. The 1.0f pattern is kept in a register (opaque to the
// compiler) so the byte selector can be the instruction's immediate operand.
#ifndef RD_DEQUANT_I2F
#define RD_DEQUANT_I2F 1
#endif
#ifndef RD_NODE_UNROLL_H
#define RD_NODE_UNROLL_H 2   // ablation: 1 keeps the loop over the two node halves of rd_node_test rolled (4 % less code in kernels that
#define RD_NODE_UNROLL_J 4   // inline several traversals). Measured on C3 path.cl, 4 spp: wavefront 1726 -> 1652, inline 805 -> 697 Mrays/s
#endif
#ifndef RD_INNER_FROM_IMASK
#define RD_INNER_FROM_IMASK 1   // inner-child byte mask of rd_node_test spread from the imask nibble (2 IMAD on the FMA pipe) instead of
                                // being derived from the meta bytes (4 ALU instructions): C2 4450 -> 4490 Mrays/s, 10 M unchanged, same image
#endif
#ifndef RD_MASK_V2
#define RD_MASK_V2 1       // hit-mask assembly of rd_node_test with the 4-wide words pinned (see there): 70 -> 54 instructions,
                           // C2 4374 -> 4460, 10 M triangles 2980 -> 3009 Mrays/s, same image; 0 = the plain expressions
#endif
#ifdef __CUDA_ARCH__
#define RD_BYTE_AS_UNIT_FLOAT(q, j) __uint_as_float(__byte_perm((q), rdOneBits, 0x7604u | ((j) << 4)))
#else
#define RD_BYTE_AS_UNIT_FLOAT(q, j) RD_U2F(0x3f800000u | ((((q) >> (8 * (j))) & 0xffu) << 8))
#endif

// Tests the 8 quantised child boxes of one node against the ray interval [tmin, tbest].
// Returns the hit mask: bits 24..31 = inner children keyed by visiting priority (highest first),
// bits 0..23 = leaf records (one bit per primitive record relative to primBase).
// Plane distance of quantised coordinate c on one axis: t = (p + c*s - o) / d = c*a + b with
// a = s/d, b = (p - o)/d. With u = 1 + c*2^-15 this is t = u*A + (b - A), A = a*2^15; the
// cancellation costs at most |A|*2^-23 (0.4 % of a grid step), which is given back by widening the
// near/far constants by |A|*2^-21, so the test stays conservative.
RD_HD rd_u32 rd_node_test(const RdU4& n0, const RdU4& n1, const RdU4& n2, const RdU4& n3, const RdU4& n4,
                          const RdRayFrame& f, float tmin, float tbest)
{
    const float sx = RD_U2F((n0.w & 0xffu) << 23), sy = RD_U2F(((n0.w >> 8) & 0xffu) << 23),
                sz = RD_U2F(((n0.w >> 16) & 0xffu) << 23);
    // Two ways to turn a quantised byte into a plane distance, on different pipes of the SM:
    //   I2F   t = float(q) * A + b                    conversion on the XU pipe (a warp instruction occupies it 8 cycles)
    //   PRMT  t = u * (A * 2^15) + (b - A * 2^15)     u = 1 + q * 2^-15 built by dropping the byte into the mantissa of
    //         1.0f (ALU pipe); the cancellation costs at most |A| * 2^-8 of a grid step... bounded by widening the
    //         constant by |A * 2^15| * 2^-21 so the test stays conservative
    // RD_DEQUANT_I2F: 1 = all 48 conversions I2F (XU 63 % busy, the limiter), 0 = all PRMT (ALU becomes the limiter),
    // 2 = near planes I2F, far planes PRMT. Measured on C2: 1 = 4018, 2 = 3835-3971, 0 = 3806-3933 Mrays/s -> default 1.
    const float Ax = sx * f.ix, Ay = sy * f.iy, Az = sz * f.iz;
    const float bx = (RD_U2F(n0.x) - f.ox) * f.ix, by = (RD_U2F(n0.y) - f.oy) * f.iy, bz = (RD_U2F(n0.z) - f.oz) * f.iz;
#if RD_DEQUANT_I2F != 1
#ifdef __CUDA_ARCH__
    rd_u32 rdOneBits;
    asm("mov.b32 %0, 0x3f800000;" : "=r"(rdOneBits));
#endif
    const float Px = Ax * 32768.0f, Py = Ay * 32768.0f, Pz = Az * 32768.0f;
    const float wx = fabsf(Px) * 4.76837158e-7f, wy = fabsf(Py) * 4.76837158e-7f, wz = fabsf(Pz) * 4.76837158e-7f;
    const float pfx = bx - Px + wx, pfy = by - Py + wy, pfz = bz - Pz + wz;   // far planes lean late
#define RD_FAR(q, j, P, pb) RD_FMA(RD_BYTE_AS_UNIT_FLOAT(q, j), P, pb)
#else
#define RD_FAR(q, j, P, pb) RD_FMA((float)(((q) >> (8 * (j))) & 0xffu), P, pb)
    const float Px = Ax, Py = Ay, Pz = Az, pfx = bx, pfy = by, pfz = bz;
#endif
#if RD_DEQUANT_I2F == 0
    const float Nx = Px, Ny = Py, Nz = Pz;
    const float pnx = bx - Px - wx, pny = by - Py - wy, pnz = bz - Pz - wz;   // near planes lean early
#define RD_NEAR(q, j, P, pb) RD_FMA(RD_BYTE_AS_UNIT_FLOAT(q, j), P, pb)
#else
    const float Nx = Ax, Ny = Ay, Nz = Az, pnx = bx, pny = by, pnz = bz;
#define RD_NEAR(q, j, P, pb) RD_FMA((float)(((q) >> (8 * (j))) & 0xffu), P, pb)
#endif
    rd_u32 hitmask = 0;
#if defined(__CUDA_ARCH__)
#if RD_NODE_UNROLL_H > 1
#pragma unroll
#else
#pragma unroll 1
#endif
#endif
    for (int h = 0; h < 2; ++h) {
        const rd_u32 meta4 = h ? n1.w : n1.z;
#if RD_INNER_FROM_IMASK
        // 0xff in the bytes of the inner children, spread from the node's imask nibble (bit j = slot j is inner) by two
        // integer multiply-adds (FMA pipe) instead of being derived from the meta bytes on the ALU pipe
        const rd_u32 innerMask4 = ((((n0.w >> (24 + 4 * h)) & 0xfu) * 0x00204081u) & 0x01010101u) * 0xffu;
#else
        const rd_u32 inner4 = (meta4 & (meta4 << 1)) & 0x10101010u;
        const rd_u32 innerMask4 = (inner4 >> 4) * 0xffu;
#endif
        const rd_u32 bitIndex4 = (meta4 ^ (f.octinv4 & innerMask4)) & 0x1f1f1f1fu;
#if defined(__CUDA_ARCH__) && RD_MASK_V2
        // the shift below takes its amount modulo 32, so the index bytes need no masking here
        rd_u32 childBits4 = (meta4 >> 5) & 0x07070707u, shift4 = meta4 ^ (f.octinv4 & innerMask4);
        // keep the two 4-wide words as computed: left alone, ptxas re-derives them per child under the hit predicate
        // (8 ALU instructions per child instead of 4)
        asm volatile("" : "+r"(childBits4), "+r"(shift4));
#else
        const rd_u32 childBits4 = (meta4 >> 5) & 0x07070707u;
#endif
        const rd_u32 qlx = h ? n2.y : n2.x, qly = h ? n2.w : n2.z, qlz = h ? n3.y : n3.x;
        const rd_u32 qhx = h ? n3.w : n3.z, qhy = h ? n4.y : n4.x, qhz = h ? n4.w : n4.z;
        const rd_u32 nx = f.ix < 0.0f ? qhx : qlx, fx = f.ix < 0.0f ? qlx : qhx;
        const rd_u32 ny = f.iy < 0.0f ? qhy : qly, fy = f.iy < 0.0f ? qly : qhy;
        const rd_u32 nz = f.iz < 0.0f ? qhz : qlz, fz = f.iz < 0.0f ? qlz : qhz;
#if defined(__CUDA_ARCH__)
#if RD_NODE_UNROLL_J > 1
#pragma unroll
#else
#pragma unroll 1
#endif
#endif
        for (int j = 0; j < 4; ++j) {
            const int sh = 8 * j;
            const float t0x = RD_NEAR(nx, j, Nx, pnx);
            const float t0y = RD_NEAR(ny, j, Ny, pny);
            const float t0z = RD_NEAR(nz, j, Nz, pnz);
            const float t1x = RD_FAR(fx, j, Px, pfx);
            const float t1y = RD_FAR(fy, j, Py, pfy);
            const float t1z = RD_FAR(fz, j, Pz, pfz);
            const float tn = RD_FMAX(RD_FMAX(t0x, t0y), RD_FMAX(t0z, tmin));
            const float tf = RD_FMIN(RD_FMIN(t1x, t1y), RD_FMIN(t1z, tbest));
#if defined(__CUDA_ARCH__) && RD_MASK_V2
            if (tn <= tf) hitmask |= __byte_perm(childBits4, 0u, 0x4440u | (unsigned)j) << ((shift4 >> sh) & 31u);
#else
            if (tn <= tf) hitmask |= ((childBits4 >> sh) & 0xffu) << ((bitIndex4 >> sh) & 0xffu);
#endif
        }
    }
#undef RD_NEAR
#undef RD_FAR
    return hitmask;
}

#ifdef __CUDA_ARCH__
#define RD_LOAD_U4(dst, ptr) { uint4 _v = __ldg((const uint4*)(ptr)); dst.x = _v.x; dst.y = _v.y; dst.z = _v.z; dst.w = _v.w; }
#define RD_LOAD_F4(dst, ptr) { float4 _v = __ldg((const float4*)(ptr)); dst.x = _v.x; dst.y = _v.y; dst.z = _v.z; dst.w = _v.w; }
#else
#define RD_LOAD_U4(dst, ptr) { dst = *(const RdU4*)(ptr); }
#define RD_LOAD_F4(dst, ptr) { dst = *(const RdF4*)(ptr); }
#endif

// Traversal state of one ray over the two-level structure. begin() arms it, step() runs one
// iteration and returns false when the ray is finished. Kept as a struct so that a persistent kernel
// can interleave steps with ray fetching (csrc/rd_trace.cu); rd_traverse() below is the plain loop.
//
// One iteration has three phases that the lanes of a warp go through together:
//   1. node phase     every lane with an inner child pending tests one 8-wide node; the triangle
//                     records the test exposed are only NOTED in the lane's pending list
//   2. triangle phase the warp votes; when at least 1/RD_FLUSH_DIV of the lanes hold pending records
//                     (or one lane must flush: list full / nothing else left in this bottom-level
//                     structure) every lane tests all its pending triangles
//   3. pop phase      lanes without inner work take the next stack entry
// Testing triangles as they appear ran the 87-instruction Moeller-Trumbore block with 6 of 32 lanes
// active (ncu, profiles/README_r1.md); collecting them first is the triangle postponing of Ylitie,
// Karras & Laine 2017 driven by a warp vote. Top-level leaf records are instances and are entered at
// once.
struct RdStackEntry { rd_u32 x, y; };

#ifndef RD_PREFETCH
#define RD_PREFETCH 0      // ablation: prefetch the next node and the group's last triangle record when a node exposes
                           // triangles. Measured on C2: 4364 -> 4272 Mrays/s, slowest tile row alone 0.228 -> 0.234 ms: a warp on
                           // its own is bound by dependent ALU latency, not by these loads (profiles/README_r1.md)
#endif
#ifndef RD_PEND_MAX
#define RD_PEND_MAX 1      // pending triangle groups per lane before it must flush
#endif
#ifndef RD_FLUSH_DIV
#define RD_FLUSH_DIV 1     // flush when >= 1/RD_FLUSH_DIV of the active lanes hold pending triangles; 1 = at once
                           // (measured fastest on C2: 4026 vs 3917 Mrays/s for 2..4, profiles/exp_variants_r1d.log)
#endif
#ifndef RD_COOP_TRI
#define RD_COOP_TRI 0      // ablation: warp-cooperative triangle phase (rd_traverse) -- the (ray, triangle) pairs the lanes of a
                           // warp hold are dealt out over all lanes through shared memory, one exact test per lane per round,
                           // finished lanes keep working for the others. Same image (px_diff 0), but slower everywhere:
                           // C2 primary 4373 -> 4091 Mrays/s (registers capped at 72; 3669 uncapped at 96), 10 M-triangle
                           // primary 2978 -> 2885, C3 path.cl inline 773 -> 494, wavefront (shadow rays in the shade
                           // stage) 1701 -> 1607: the ballots, three warp barriers and 3 x 16 B of staging per pair cost more
                           // than the exact test run at 6 of 32 lanes (profiles/README_r1.md)
#endif
#ifndef RD_COOP_WARPS
#define RD_COOP_WARPS 8    // warps per block the staging array is sized for (block <= 256)
#endif
#ifndef RD_PEND_REG
#define RD_PEND_REG 0      // ablation: keep the pending triangle group of an iteration in two registers instead of the thread's
                           // local-memory list (saves STL + LDL + address arithmetic in front of the triangle fetch). Same image,
                           // much slower: C2 4460 -> 3872, 10 M triangles 3000 -> 2431 Mrays/s. Without the memory round trip ptxas
                           // threads lanes straight from the node test into the triangle loop and the warp no longer walks the
                           // three phases together -- the phase structure is worth more than the instructions saved
#endif
#define RD_STACK_ALLOC (RD_STACK_ENTRIES + RD_PEND_MAX)

struct RdTraversal {
    RdStackEntry* stack;              // RD_STACK_ALLOC entries owned by the caller (thread-local memory);
                                      // kept outside so the scalar state below stays in registers
    RdStackEntry* pend;               // pending triangle groups (the tail of the same allocation)
    int sp, blasSp, np;               // blasSp >= 0 while inside a bottom-level structure
    const RdNode8* nodes;
    const RdTri* tris;
    const RdInstance* curInst;
    RdRayFrame f;                     // ray in the current space (world or instance-local)
    float wox, woy, woz, wdx, wdy, wdz, tmin, tmax, tbest;
    RdStackEntry ng;                  // node group being worked on
    RdHitRec hit;
    bool any;

    RD_HD void push(const RdScene* scene, RdStackEntry e)
    {
        if (sp < RD_STACK_ENTRIES) stack[sp++] = e;
        else if (scene->status) scene->status[0] = 1u;
    }

    RD_HD void begin(RdStackEntry* stackMemory, const RdScene* scene, float ox, float oy, float oz, float dx, float dy, float dz, float tmin_, float tmax_)
    {
        stack = stackMemory; pend = stackMemory + RD_STACK_ENTRIES; sp = 0; blasSp = -1; np = 0;
        nodes = scene->nodes; tris = 0; curInst = 0;
        wox = ox; woy = oy; woz = oz; wdx = dx; wdy = dy; wdz = dz; tmin = tmin_; tmax = tmax_;
        rd_frame_set(f, ox, oy, oz, dx, dy, dz);
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
        RdStackEntry tg; tg.x = 0; tg.y = 0;   // top-level leaf records (instances) to enter now
#if RD_PEND_REG
        RdStackEntry pg; pg.x = 0; pg.y = 0;   // triangle records exposed by this iteration's node test, tested below
#endif
        // ---- 1. node phase
        if (ng.y > 0x00ffffffu) {
            const rd_u32 bit = 31u - (rd_u32)RD_CLZ(ng.y);
            const rd_u32 imask = ng.y & 0xffu;   // low byte of an inner group = imask
            ng.y &= ~(1u << bit);
            if (ng.y > 0x00ffffffu) push(scene, ng);
            const rd_u32 slot = (bit - 24u) ^ (f.octinv4 & 0xffu);
            const rd_u32 rel = (rd_u32)RD_POPC(imask & ~(0xffffffffu << slot));
            const RdNode8* n = nodes + (ng.x + rel);
            RdU4 n0, n1, n2, n3, n4;
            RD_LOAD_U4(n0, &n->w[0]); RD_LOAD_U4(n1, &n->w[1]); RD_LOAD_U4(n2, &n->w[2]);
            RD_LOAD_U4(n3, &n->w[3]); RD_LOAD_U4(n4, &n->w[4]);
            onCandidate.countNode();
            const rd_u32 hm = rd_node_test(n0, n1, n2, n3, n4, f, tmin, tbest);
            ng.x = n1.x; ng.y = (hm & 0xff000000u) | (n0.w >> 24);
            const rd_u32 leafBits = hm & 0x00ffffffu;
            if (leafBits) {
                if (blasSp < 0) { tg.x = n1.y; tg.y = leafBits; }
                else {
#if RD_PEND_REG
                    pg.x = n1.y; pg.y = leafBits;
#else
                    pend[np].x = n1.y; pend[np].y = leafBits; ++np;
#endif
#if defined(__CUDA_ARCH__) && RD_PREFETCH
                    // The triangle phase runs before the next node is needed: start that node's fetch now, and the fetch of
                    // the last triangle record of this group (the first one is loaded at once). A warp on its own -- the tail
                    // of a frame, a frame split over several GPUs -- is bound by exactly this chain of dependent loads.
                    asm volatile("prefetch.global.L1 [%0];" :: "l"(tris + (n1.y + (31u - (rd_u32)RD_CLZ(leafBits)))));
                    if (ng.y > 0x00ffffffu) {
                        const rd_u32 nbit = 31u - (rd_u32)RD_CLZ(ng.y);
                        const rd_u32 nslot = (nbit - 24u) ^ (f.octinv4 & 0xffu);
                        const RdNode8* nx = nodes + (ng.x + (rd_u32)RD_POPC((ng.y & 0xffu) & ~(0xffffffffu << nslot)));
                        asm volatile("prefetch.global.L1 [%0];" :: "l"(nx));
                        asm volatile("prefetch.global.L1 [%0];" :: "l"((const char*)nx + 64));
                    }
#endif
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
                rd_frame_set(f, lox, loy, loz, ldx, ldy, ldz);
                nodes = curInst->nodes; tris = curInst->tris;
                ng.x = 0; ng.y = 0x80000000u;
                return true;
            }
        } else {
            // ---- 2. triangle phase
#if RD_PEND_REG
            {
                RdStackEntry g = pg;
                while (g.y) {
#else
            const bool exhausted = ng.y <= 0x00ffffffu && sp == blasSp;   // about to leave this structure
            bool flush = np > 0 && (np == RD_PEND_MAX || exhausted);
#if defined(__CUDA_ARCH__) && RD_FLUSH_DIV > 1
            {
                const unsigned act = __activemask();
                const unsigned forced = __ballot_sync(act, flush);
                const unsigned holding = __ballot_sync(act, np > 0);
                flush = np > 0 && (forced != 0u || __popc(holding) * RD_FLUSH_DIV >= __popc(act));
            }
#else
            flush = np > 0;
#endif
            if (flush) {
                RdStackEntry g; g.x = 0; g.y = 0;
                while (np > 0 || g.y) {
                    if (!g.y) g = pend[--np];
#endif
                    const rd_u32 k = (rd_u32)RD_FFS(g.y) - 1u;
                    g.y &= g.y - 1u;
                    const RdTri* tr = tris + (g.x + k);
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
        }
        // ---- 3. pop phase
        if (ng.y <= 0x00ffffffu) {
            if (blasSp >= 0 && sp == blasSp) {
                // bottom-level structure exhausted (its pending list was flushed above): back to world space
                blasSp = -1; nodes = scene->nodes;
                rd_frame_set(f, wox, woy, woz, wdx, wdy, wdz);
            }
            if (sp == 0) return false;
            ng = stack[--sp];
        }
        return true;
    }

#if defined(__CUDA_ARCH__) && RD_COOP_TRI
    // step() cut in two for the warp-cooperative triangle phase of rd_traverse (RD_COOP_TRI): head() = node phase and
    // instance entry, leaves the triangle records the node exposed in (*gbase, *gbits) and returns true when the lane
    // entered an instance (it then skips the rest of the iteration); tail() = pop phase, false when the ray is done.
    template <class OnCandidate>
    RD_HD bool head(const RdScene* scene, OnCandidate& onCandidate, rd_u32* gbase, rd_u32* gbits)
    {
        RdStackEntry tg; tg.x = 0; tg.y = 0;
        *gbase = 0; *gbits = 0;
        if (ng.y > 0x00ffffffu) {
            const rd_u32 bit = 31u - (rd_u32)RD_CLZ(ng.y);
            const rd_u32 imask = ng.y & 0xffu;
            ng.y &= ~(1u << bit);
            if (ng.y > 0x00ffffffu) push(scene, ng);
            const rd_u32 slot = (bit - 24u) ^ (f.octinv4 & 0xffu);
            const rd_u32 rel = (rd_u32)RD_POPC(imask & ~(0xffffffffu << slot));
            const RdNode8* n = nodes + (ng.x + rel);
            RdU4 n0, n1, n2, n3, n4;
            RD_LOAD_U4(n0, &n->w[0]); RD_LOAD_U4(n1, &n->w[1]); RD_LOAD_U4(n2, &n->w[2]);
            RD_LOAD_U4(n3, &n->w[3]); RD_LOAD_U4(n4, &n->w[4]);
            onCandidate.countNode();
            const rd_u32 hm = rd_node_test(n0, n1, n2, n3, n4, f, tmin, tbest);
            ng.x = n1.x; ng.y = (hm & 0xff000000u) | (n0.w >> 24);
            const rd_u32 leafBits = hm & 0x00ffffffu;
            if (leafBits) {
                if (blasSp < 0) { tg.x = n1.y; tg.y = leafBits; }
                else { *gbase = n1.y; *gbits = leafBits; }
            }
        } else if (ng.y) {
            tg = ng; ng.x = 0; ng.y = 0;
        }
        if (blasSp < 0 && tg.y) {
            const rd_u32 k = (rd_u32)RD_FFS(tg.y) - 1u;
            tg.y &= tg.y - 1u;
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
            rd_frame_set(f, lox, loy, loz, ldx, ldy, ldz);
            nodes = curInst->nodes; tris = curInst->tris;
            ng.x = 0; ng.y = 0x80000000u;
            return true;
        }
        return false;
    }

    RD_HD bool tail(const RdScene* scene)
    {
        if (ng.y <= 0x00ffffffu) {
            if (blasSp >= 0 && sp == blasSp) {
                blasSp = -1; nodes = scene->nodes;
                rd_frame_set(f, wox, woy, woz, wdx, wdy, wdz);
            }
            if (sp == 0) return false;
            ng = stack[--sp];
        }
        return true;
    }
#endif
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
#if defined(__CUDA_ARCH__) && RD_COOP_TRI && RD_RECONVERGE
    {
        // slot s of a warp: [0] = {o.xyz, d.x}, [1] = {d.y, d.z, triangle address}, [2] = result {t, b1, b2, prim}
        __shared__ float4 rdCoop[RD_COOP_WARPS][32][3];
        float4 (*slots)[3] = rdCoop[(threadIdx.x >> 5) % RD_COOP_WARPS];
        const unsigned lane = threadIdx.x & 31u, below = together & ((1u << lane) - 1u);
        const unsigned worker = (unsigned)__popc(below), workers = (unsigned)__popc(together);
        bool alive = true;
        for (;;) {
            rd_u32 gbase = 0, gbits = 0;
            bool entered = false;
            if (alive) entered = T.head(scene, onCandidate, &gbase, &gbits);
#if RD_COOP_TRI == 2
            // variant 2: no dealing out -- every lane tests its own records (held in registers), but the warp is made to
            // walk the phases together by explicit barriers. Same image; C2 4490 -> 4160, 10 M triangles 3005 -> 2920
            // Mrays/s (79 registers: one block per SM fewer, and finished lanes keep looping)
            __syncwarp(together);
            while (gbits) {
                const rd_u32 k = (rd_u32)__ffs((int)gbits) - 1u;
                gbits &= gbits - 1u;
                const RdTri* tr = T.tris + (gbase + k);
                RdF4 a, b, c;
                RD_LOAD_F4(a, &tr->v0); RD_LOAD_F4(b, &tr->v1); RD_LOAD_F4(c, &tr->v2);
                onCandidate.countTri();
                float t, b1, b2;
                if (rd_intersect_triangle(T.f.ox, T.f.oy, T.f.oz, T.f.dx, T.f.dy, T.f.dz, a.x, a.y, a.z, b.x, b.y, b.z,
                                          c.x, c.y, c.z, &t, &b1, &b2) &&
                    t > T.tmin && t < T.tmax &&
                    (t < T.hit.t || (t == T.hit.t && (T.curInst->instanceID < T.hit.inst ||
                                                      (T.curInst->instanceID == T.hit.inst && RD_F2U(a.w) < T.hit.prim))))) {
                    T.hit.t = t; T.hit.b1 = b1; T.hit.b2 = b2; T.hit.prim = RD_F2U(a.w); T.hit.inst = T.curInst->instanceID;
                    T.tbest = t; T.any = true;
                    if (!onCandidate(T.hit)) { alive = false; gbits = 0u; }
                }
            }
            __syncwarp(together);
#endif
            // ---- triangle phase, all lanes of the warp (finished ones work for the others)
            for (;;) {
                if (!__any_sync(together, gbits != 0u)) break;
                unsigned nslots = 0, mine[3] = {0u, 0u, 0u}, count = 0;
#pragma unroll
                for (int k = 0; k < 3; ++k) {
                    const unsigned m = __ballot_sync(together, gbits != 0u);
                    if (m == 0u || nslots + (unsigned)__popc(m) > workers) break;
                    if (gbits) {
                        const rd_u32 tri = gbase + ((rd_u32)__ffs((int)gbits) - 1u);
                        gbits &= gbits - 1u;
                        const unsigned s = nslots + (unsigned)__popc(m & below);
                        const unsigned long long addr = (unsigned long long)(T.tris + tri);
                        slots[s][0] = make_float4(T.f.ox, T.f.oy, T.f.oz, T.f.dx);
                        slots[s][1] = make_float4(T.f.dy, T.f.dz, __uint_as_float((unsigned)addr), __uint_as_float((unsigned)(addr >> 32)));
                        mine[k] = s; count = (unsigned)k + 1u;
                        onCandidate.countTri();
                    }
                    nslots += (unsigned)__popc(m);
                }
                __syncwarp(together);
                if (worker < nslots) {
                    const float4 ra = slots[worker][0], rb = slots[worker][1];
                    const RdTri* tr = (const RdTri*)(((unsigned long long)__float_as_uint(rb.w) << 32) | __float_as_uint(rb.z));
                    RdF4 a, b, c;
                    RD_LOAD_F4(a, &tr->v0); RD_LOAD_F4(b, &tr->v1); RD_LOAD_F4(c, &tr->v2);
                    float t, b1, b2;
                    const bool ok = rd_intersect_triangle(ra.x, ra.y, ra.z, ra.w, rb.x, rb.y, a.x, a.y, a.z, b.x, b.y, b.z,
                                                          c.x, c.y, c.z, &t, &b1, &b2);
                    slots[worker][2] = make_float4(ok ? t : -1.0f, b1, b2, a.w);
                }
                __syncwarp(together);
#pragma unroll
                for (int k = 0; k < 3; ++k) {
                    if (k < (int)count && alive) {
                        const float4 r = slots[mine[k]][2];
                        const float t = r.x; const rd_u32 prim = RD_F2U(r.w);
                        if (t > 0.0f && t > T.tmin && t < T.tmax &&
                            (t < T.hit.t || (t == T.hit.t && (T.curInst->instanceID < T.hit.inst ||
                                                              (T.curInst->instanceID == T.hit.inst && prim < T.hit.prim))))) {
                            T.hit.t = t; T.hit.b1 = r.y; T.hit.b2 = r.z; T.hit.prim = prim; T.hit.inst = T.curInst->instanceID;
                            T.tbest = t; T.any = true;
                            if (!onCandidate(T.hit)) { alive = false; gbits = 0u; }
                        }
                    }
                }
                __syncwarp(together);
            }
            if (alive && !entered) alive = T.tail(scene);
            if (!__any_sync(together, alive)) break;
        }
    }
#else
    while (T.step(scene, onCandidate)) {}
#if defined(__CUDA_ARCH__) && RD_RECONVERGE
    __syncwarp(together);
#endif
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
