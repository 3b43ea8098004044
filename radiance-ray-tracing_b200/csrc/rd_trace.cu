// rd_trace.cu -- wavefront building blocks compiled into the library:
//   rd_k_trace        persistent-warp closest-hit / first-hit traversal over an SoA ray queue
//   rd_k_primary_rays camera ray generation into the SoA queue (samples/shader.cl:111-173 semantics)
// The traversal itself is csrc/rd_traverse.h (shared with the NVRTC-compiled user programs).
#include <string.h>
#include "rd_internal.h"
#include "rd_traverse.h"

// Each warp grabs 32 consecutive rays at a time from a global counter (lane 0 atomics + shuffle),
// so the grid is sized to the machine (SM count x resident blocks), not to the ray count.
template <int MODE>   // 0 closest hit, 1 first hit, 2 closest + counters, 3 first + counters
__global__ void __launch_bounds__(128, 7) rd_k_trace(const RdScene* __restrict__ scene, const float4* __restrict__ rayA,
                                                   const float4* __restrict__ rayB, float4* __restrict__ hitA,
                                                   rd_u32* __restrict__ hitB, unsigned long long n,
                                                   unsigned long long* __restrict__ counters)
{
    const unsigned lane = threadIdx.x & 31u;
    unsigned long long cn = 0, ct = 0;
    for (;;) {
        unsigned long long base = 0;
        if (lane == 0) base = atomicAdd(&counters[4], 32ull);
        base = __shfl_sync(0xffffffffu, base, 0);
        if (base >= n) break;
        const unsigned long long i = base + lane;
        if (i < n) {
            const float4 a = __ldg(&rayA[i]), b = __ldg(&rayB[i]);
            RdHitRec h;
            if (MODE == 0) { RdNoAnyHit cb; rd_traverse(scene, a.x, a.y, a.z, b.x, b.y, b.z, a.w, b.w, h, cb); }
            else if (MODE == 1) { RdFirstHit cb; rd_traverse(scene, a.x, a.y, a.z, b.x, b.y, b.z, a.w, b.w, h, cb); }
            else {
                RdCountingHit cb; cb.terminate = (MODE == 3); cb.nodes = 0; cb.tris = 0;
                rd_traverse(scene, a.x, a.y, a.z, b.x, b.y, b.z, a.w, b.w, h, cb);
                cn += cb.nodes; ct += cb.tris;
            }
            hitA[i] = make_float4(h.t, h.b1, h.b2, __uint_as_float(h.prim));
            hitB[i] = h.inst;
        }
    }
    if (MODE >= 2) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) { cn += __shfl_xor_sync(0xffffffffu, cn, o); ct += __shfl_xor_sync(0xffffffffu, ct, o); }
        if (lane == 0) { atomicAdd(&counters[0], cn); atomicAdd(&counters[1], ct); }
    }
}

// Same traversal with dynamic ray fetch (Aila & Laine 2009): finished lanes do not wait for the
// slowest ray of their batch; whenever at least RD_FETCH_MIN lanes of the warp are idle the warp grabs
// that many new rays with one atomic and the idle lanes re-arm, so node tests run with (nearly) full
// warps. Lanes keep their ray's traversal state in RdTraversal between steps.
#ifndef RD_FETCH_MIN
#define RD_FETCH_MIN 8
#endif
template <int MODE>
__global__ void __launch_bounds__(128) rd_k_trace_dyn(const RdScene* __restrict__ scene, const float4* __restrict__ rayA,
                                                       const float4* __restrict__ rayB, float4* __restrict__ hitA,
                                                       rd_u32* __restrict__ hitB, unsigned long long n,
                                                       unsigned long long* __restrict__ counters)
{
    const unsigned lane = threadIdx.x & 31u;
    const unsigned ltMask = (1u << lane) - 1u;
    RdStackEntry stackMemory[RD_STACK_ALLOC];
    RdTraversal T;
    RdCountingHit cb; cb.terminate = (MODE & 1) != 0; cb.nodes = 0; cb.tris = 0;
    bool active = false, exhausted = false;
    unsigned long long mine = 0;
    for (;;) {
        unsigned act = __ballot_sync(0xffffffffu, active);
        if (!exhausted && __popc(act) <= 32 - RD_FETCH_MIN) {
            const unsigned idle = ~act;
            const unsigned cnt = (unsigned)__popc(idle);
            unsigned long long base = 0;
            if (lane == 0) base = atomicAdd(&counters[4], (unsigned long long)cnt);
            base = __shfl_sync(0xffffffffu, base, 0);
            if (!active) {
                const unsigned long long i = base + (unsigned)__popc(idle & ltMask);
                if (i < n) {
                    const float4 a = __ldg(&rayA[i]), b = __ldg(&rayB[i]);
                    T.begin(stackMemory, scene, a.x, a.y, a.z, b.x, b.y, b.z, a.w, b.w);
                    active = true; mine = i;
                }
            }
            if (base + cnt >= n) exhausted = true;
            act = __ballot_sync(0xffffffffu, active);
        }
        if (act == 0u) break;
        if (active) {
            bool more;
            if (MODE >= 2) more = T.step(scene, cb);
            else if (MODE == 1) { RdFirstHit f; more = T.step(scene, f); }
            else { RdNoAnyHit f; more = T.step(scene, f); }
            if (!more) {
                hitA[mine] = make_float4(T.hit.t, T.hit.b1, T.hit.b2, __uint_as_float(T.hit.prim));
                hitB[mine] = T.hit.inst;
                active = false;
            }
        }
    }
    if (MODE >= 2) {
        unsigned long long cn = cb.nodes, ct = cb.tris;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) { cn += __shfl_xor_sync(0xffffffffu, cn, o); ct += __shfl_xor_sync(0xffffffffu, ct, o); }
        if (lane == 0) { atomicAdd(&counters[0], cn); atomicAdd(&counters[1], ct); }
    }
}

extern "C" int rd_trace_ray_list(rd_platform* p, rd_handle tlas, rd_handle rayA, rd_handle rayB, rd_handle hitA,
                                 rd_handle hitB, size_t n, unsigned flags)
{
    RdHandleObj *T = rd_obj(tlas), *A = rd_obj(rayA), *B = rd_obj(rayB), *HA = rd_obj(hitA), *HB = rd_obj(hitB);
    if (!p || !T || T->kind != RD_H_TLAS || !A || !B || !HA || !HB) { rd_set_error("rd_trace_ray_list: bad handle"); return RD_E_ARG; }
    if (A->size < n * 16 || B->size < n * 16 || HA->size < n * 16 || HB->size < n * 4) {
        rd_set_error("rd_trace_ray_list: buffers too small for %zu rays", n);
        return RD_E_ARG;
    }
    if (!n) return RD_OK;
    RD_CUDA(cudaSetDevice(p->device));
    RD_CUDA(cudaMemsetAsync(p->dcounters, 0, 64, p->stream));
    unsigned long long rays = n;
    RD_CUDA(cudaMemcpyAsync(p->dcounters + 2, &rays, 8, cudaMemcpyHostToDevice, p->stream));
    const int mode = ((flags & RD_TRACE_TERMINATE_ON_FIRST_HIT) ? 1 : 0) | ((flags & RD_TRACE_COUNTERS) ? 2 : 0);
    size_t warps = (n + 31) / 32;
    unsigned blocks = (unsigned)((warps + 3) / 4);
    const unsigned maxBlocks = (unsigned)p->smCount * 16;
    if (blocks > maxBlocks) blocks = maxBlocks;
    const RdScene* S = (const RdScene*)T->dptr;
    const float4 *a = (const float4*)A->dptr, *b = (const float4*)B->dptr;
    float4* ha = (float4*)HA->dptr; rd_u32* hb = (rd_u32*)HB->dptr;
    if (!(flags & RD_TRACE_DYNAMIC_FETCH)) {   // static 32-ray batches measured faster (4.0 vs 2.9 Grays/s on C2)
        switch (mode) {
            case 0: rd_k_trace<0><<<blocks, 128, 0, p->stream>>>(S, a, b, ha, hb, n, p->dcounters); break;
            case 1: rd_k_trace<1><<<blocks, 128, 0, p->stream>>>(S, a, b, ha, hb, n, p->dcounters); break;
            case 2: rd_k_trace<2><<<blocks, 128, 0, p->stream>>>(S, a, b, ha, hb, n, p->dcounters); break;
            default: rd_k_trace<3><<<blocks, 128, 0, p->stream>>>(S, a, b, ha, hb, n, p->dcounters); break;
        }
    } else {
        // persistent: one wave of blocks sized to the machine
        unsigned pblocks = (unsigned)p->smCount * 6;
        if (pblocks > blocks) pblocks = blocks;
        switch (mode) {
            case 0: rd_k_trace_dyn<0><<<pblocks, 128, 0, p->stream>>>(S, a, b, ha, hb, n, p->dcounters); break;
            case 1: rd_k_trace_dyn<1><<<pblocks, 128, 0, p->stream>>>(S, a, b, ha, hb, n, p->dcounters); break;
            case 2: rd_k_trace_dyn<2><<<pblocks, 128, 0, p->stream>>>(S, a, b, ha, hb, n, p->dcounters); break;
            default: rd_k_trace_dyn<3><<<pblocks, 128, 0, p->stream>>>(S, a, b, ha, hb, n, p->dcounters); break;
        }
    }
    p->launches += 1;
    RD_CUDA(cudaGetLastError());
    return RD_OK;
}

extern "C" int rd_trace_counters(rd_platform* p, uint64_t out[3])
{
    if (!p || !out) { rd_set_error("rd_trace_counters: null argument"); return RD_E_ARG; }
    unsigned long long h[3];
    RD_CUDA(cudaSetDevice(p->device));
    RD_CUDA(cudaMemcpyAsync(h, p->dcounters, sizeof(h), cudaMemcpyDeviceToHost, p->stream));
    RD_CUDA(cudaStreamSynchronize(p->stream));
    out[0] = h[0]; out[1] = h[1]; out[2] = h[2];
    return rd_check_status(p, "rd_trace_counters");
}

// ---- primary rays ---------------------------------------------------------------------------
// PhysicalCamera, radiance/src/core.h:151-158 (48 bytes)
struct RdCamera { float widthPixel, heightPixel, focalLength, sensorWidth, focalDistance, fStop, x, y, z, wx, wy, wz; };

// PCG3D hash (Jarzynski & Olano 2020), as used by radiance/shader/math.cl:10-23: integer stage
// bit-exact, then division by (float)0xffffffff
__device__ __forceinline__ void rd_pcg3d(unsigned x, unsigned y, unsigned z, float* r)
{
    x = x * 1664525u + 1013904223u; y = y * 1664525u + 1013904223u; z = z * 1664525u + 1013904223u;
    x += y * z; y += z * x; z += x * y;
    x ^= x >> 16; y ^= y >> 16; z ^= z >> 16;
    x += y * z; y += z * x; z += x * y;
    const float d = (float)0xffffffffu;
    r[0] = __fdiv_rn((float)x, d); r[1] = __fdiv_rn((float)y, d); r[2] = __fdiv_rn((float)z, d);
}

__device__ __forceinline__ void rd_euler_zyx(const RdCamera& c, float& x, float& y, float& z)
{
    // v' = Rx * (Ry * (Rz * v)), radiance/shader/math.cl:185-252 and samples/shader.cl:139-147
    const float sz = sinf(c.wz), cz = cosf(c.wz), sy = sinf(c.wy), cy = cosf(c.wy), sx = sinf(c.wx), cx = cosf(c.wx);
    const float x1 = cz * x - sz * y, y1 = sz * x + cz * y, z1 = z;
    const float x2 = cy * x1 + sy * z1, y2 = y1, z2 = -sy * x1 + cy * z1;
    x = x2; y = cx * y2 - sx * z2; z = sx * y2 + cx * z2;
}

// Ray slot `slot` holds the ray of pixel `index`: with tileW != 0 the slots enumerate 8x4 pixel tiles
// row by row (32 consecutive slots = one tile = one warp of the trace kernel), the same mapping
// get_global_id() applies to raygen launches (shader/compat/cl_compat.h).
__global__ void rd_k_primary_rays(const RdCamera* __restrict__ cam, unsigned frameID, unsigned totalSamples,
                                  unsigned count, unsigned tileW, float4* __restrict__ rayA, float4* __restrict__ rayB)
{
    const unsigned slot = blockIdx.x * blockDim.x + threadIdx.x;
    if (slot >= count) return;
    const RdCamera c = *cam;
    const int w = (int)c.widthPixel;
    unsigned index = slot;
    if (tileW) {
        const unsigned tilesX = tileW >> 3, tile = slot >> 5, lane = slot & 31u;
        const unsigned ty = tile / tilesX, tx = tile - ty * tilesX;
        index = ((ty << 2) + (lane >> 3)) * tileW + (tx << 3) + (lane & 7u);
    }
    const int px = (int)index % w, py = (int)index / w;
    float rnd[3];
    rd_pcg3d(frameID, totalSamples, index, rnd);
    const float fx = ((float)px + rnd[0]) / c.widthPixel - 0.5f;
    const float fy = 0.5f - ((float)py + rnd[1]) / c.heightPixel;
    const float aspect = c.heightPixel / c.widthPixel;
    float dx = fx * c.sensorWidth, dy = fy * c.sensorWidth * aspect, dz = -c.focalLength;
    float len = sqrtf(dx * dx + dy * dy + dz * dz);
    dx /= len; dy /= len; dz /= len;
    const float time = -c.focalDistance / dz;
    rd_euler_zyx(c, dx, dy, dz);
    len = sqrtf(dx * dx + dy * dy + dz * dz);
    dx /= len; dy /= len; dz /= len;
    float ox = c.x, oy = c.y, oz = c.z;
    if (c.fStop != 0.0f) {
        // thin lens: concentric disk sample (pbrt) on the aperture, refocus on the focal plane
        const float lensRadius = (c.focalLength / c.fStop) / 2.0f;
        float ux = 2.0f * rnd[1] - 1.0f, uy = 2.0f * rnd[2] - 1.0f, lx = 0.0f, ly = 0.0f;
        if (!(ux == 0.0f && uy == 0.0f)) {
            const float pi = 3.14159265359f;
            float r, theta;
            if (fabsf(ux) > fabsf(uy)) { r = ux; theta = (pi / 4.0f) * (uy / ux); }
            else { r = uy; theta = (pi / 2.0f) - (pi / 4.0f) * (ux / uy); }
            lx = r * cosf(theta); ly = r * sinf(theta);
        }
        lx *= lensRadius; ly *= lensRadius;
        const float hx = ox + dx * time, hy = oy + dy * time, hz = oz + dz * time;
        float lz = 0.0f;
        rd_euler_zyx(c, lx, ly, lz);
        ox += lx; oy += ly; oz += lz;
        dx = hx - ox; dy = hy - oy; dz = hz - oz;
        len = sqrtf(dx * dx + dy * dy + dz * dz);
        dx /= len; dy /= len; dz /= len;
    }
    rayA[slot] = make_float4(ox, oy, oz, 0.001f);
    rayB[slot] = make_float4(dx, dy, dz, 1000.0f);
}

extern "C" int rd_generate_primary_rays(rd_platform* p, rd_handle camera, unsigned frameID, unsigned totalSamples,
                                        unsigned width, unsigned height, rd_handle rayA, rd_handle rayB)
{
    RdHandleObj *C = rd_obj(camera), *A = rd_obj(rayA), *B = rd_obj(rayB);
    const size_t n = (size_t)width * height;
    if (!p || !C || !A || !B || C->size < sizeof(RdCamera) || A->size < n * 16 || B->size < n * 16) {
        rd_set_error("rd_generate_primary_rays: bad handle or buffer too small");
        return RD_E_ARG;
    }
    if (!n) return RD_OK;
    RD_CUDA(cudaSetDevice(p->device));
    const unsigned tileW = (p->optTile && (width % 8u) == 0 && (height % 4u) == 0) ? width : 0u;
    rd_k_primary_rays<<<(unsigned)((n + 255) / 256), 256, 0, p->stream>>>((const RdCamera*)C->dptr, frameID, totalSamples,
                                                                            (unsigned)n, tileW, (float4*)A->dptr, (float4*)B->dptr);
    p->launches += 1;
    RD_CUDA(cudaGetLastError());
    return RD_OK;
}
