// rd_wavefront.cu -- library stages of wavefront mode: queue tracing, the hit-group / material sort
// and the host loop that sequences the stages (RD::TraceRays in wavefront mode).
// Replaces the loop nest inside the reference's raygen kernel (samples/shader.cl:196-281) by stage
// launches over queues in HBM; see rd_wavefront.h for the state and csrc/rd_program.cu for the
// generate / shade / resolve kernels that are compiled together with the user program.
#include <string.h>
#include "rd_internal.h"
#include "rd_traverse.h"
#include "rd_wavefront.h"

// ---- trace stage --------------------------------------------------------------------------------
// Closest hit for every entry of the current queue. Persistent warps take batches of 32 consecutive
// entries from a counter (queue size lives in HBM: no host round trip between stages). Also emits
// the shade stage's sort key: (hit group row, capped at 255) * 4 + (instanceCustomIndex & 3), where
// the row is instanceSBTOffset + sbtRecordOffset for hits (radiance.cl:23-28 dispatch rule) and the
// miss index for misses.
__global__ void __launch_bounds__(128, 7) rd_k_wf_trace(const RdScene* __restrict__ scene, const RdWfState* __restrict__ wfp)
{
    const RdWfState wf = *wfp;
    const unsigned n = wf.counts[wf.cur];
    const unsigned lane = threadIdx.x & 31u;
    const float4* __restrict__ rayA = (const float4*)wf.rayA[wf.cur];
    const float4* __restrict__ rayB = (const float4*)wf.rayB[wf.cur];
    for (;;) {
        unsigned base = 0;
        if (lane == 0) base = atomicAdd(&wf.counts[2], 32u);
        base = __shfl_sync(0xffffffffu, base, 0);
        if (base >= n) break;
        const unsigned j = base + lane;
        if (j < n) {
            const float4 a = __ldg(&rayA[j]), b = __ldg(&rayB[j]);
            RdHitRec h; RdNoAnyHit cb;
            rd_traverse(scene, a.x, a.y, a.z, b.x, b.y, b.z, a.w, b.w, h, cb);
            ((float4*)wf.hitA)[j] = make_float4(h.t, h.b1, h.b2, __uint_as_float(h.prim));
            wf.hitB[j] = h.inst;
            unsigned row = (unsigned)wf.missIndex, mat = 0u;
            if (h.t < RD_MISS_T) {
                const RdInstance* I = scene->instances + h.inst;
                row = I->sbtOffset + (unsigned)wf.sbtRecordOffset; mat = I->customID & 3u;
            }
            wf.keys[j] = ((row & 255u) << 2) | mat;   // rows beyond 255 share bins modulo 256: still few groups per bin
        }
    }
}


// Same stage with dynamic ray fetch (Aila & Laine 2009): bounce rays are incoherent and their walk lengths are
// heavy-tailed, so a warp that waits for the slowest ray of a fixed batch runs with 7 of 32 lanes active (ncu,
// profiles/README_r1.md). Here finished lanes do not wait: whenever at least RD_WF_FETCH_MIN lanes are idle the warp
// takes that many new entries with one atomic and the idle lanes re-arm; every lane keeps its ray's state in an
// RdTraversal between steps.
#ifndef RD_WF_FETCH_MIN
#define RD_WF_FETCH_MIN 8
#endif
__device__ __forceinline__ void rd_wf_store_hit(const RdScene* scene, const RdWfState& wf, unsigned j, const RdHitRec& h)
{
    ((float4*)wf.hitA)[j] = make_float4(h.t, h.b1, h.b2, __uint_as_float(h.prim));
    wf.hitB[j] = h.inst;
    unsigned row = (unsigned)wf.missIndex, mat = 0u;
    if (h.t < RD_MISS_T) {
        const RdInstance* I = scene->instances + h.inst;
        row = I->sbtOffset + (unsigned)wf.sbtRecordOffset; mat = I->customID & 3u;
    }
    wf.keys[j] = ((row & 255u) << 2) | mat;   // rows beyond 255 share bins modulo 256: still few groups per bin
}

template <int MINB>
__global__ void __launch_bounds__(128, MINB) rd_k_wf_trace_dyn(const RdScene* __restrict__ scene, const RdWfState* __restrict__ wfp)
{
    const RdWfState wf = *wfp;
    const unsigned n = wf.counts[wf.cur];
    const unsigned lane = threadIdx.x & 31u, ltMask = (1u << lane) - 1u;
    const float4* __restrict__ rayA = (const float4*)wf.rayA[wf.cur];
    const float4* __restrict__ rayB = (const float4*)wf.rayB[wf.cur];
    RdStackEntry stackMemory[RD_STACK_ALLOC];
    RdTraversal T;
    RdNoAnyHit cb;
    bool active = false, exhausted = false;
    unsigned mine = 0;
    for (;;) {
        unsigned act = __ballot_sync(0xffffffffu, active);
        if (!exhausted && __popc(act) <= 32 - RD_WF_FETCH_MIN) {
            const unsigned idle = ~act, cnt = (unsigned)__popc(idle);
            unsigned base = 0;
            if (lane == 0) base = atomicAdd(&wf.counts[2], cnt);
            base = __shfl_sync(0xffffffffu, base, 0);
            if (!active) {
                const unsigned j = base + (unsigned)__popc(idle & ltMask);
                if (j < n) {
                    const float4 a = __ldg(&rayA[j]), b = __ldg(&rayB[j]);
                    T.begin(stackMemory, scene, a.x, a.y, a.z, b.x, b.y, b.z, a.w, b.w);
                    active = true; mine = j;
                }
            }
            if (base + cnt >= n) exhausted = true;
            act = __ballot_sync(0xffffffffu, active);
        }
        if (act == 0u) break;
        if (active) {
            if (!T.step(scene, cb)) { rd_wf_store_hit(scene, wf, mine, T.hit); active = false; }
        }
    }
}

// ---- sort stage ---------------------------------------------------------------------------------
// Counting sort of the queue entries by key (RD_WF_KEY_BINS bins): block-local histograms in shared
// memory, one scan block, then a scatter in which every warp ranks its entries with match_any and the
// block claims a run of output positions per distinct key with one atomic (rd_k_wf_scatter).
__global__ void rd_k_wf_hist(const RdWfState* __restrict__ wfp)
{
    __shared__ unsigned bins[RD_WF_KEY_BINS];
    const RdWfState wf = *wfp;
    const unsigned n = wf.counts[wf.cur];
    for (unsigned b = threadIdx.x; b < RD_WF_KEY_BINS; b += blockDim.x) bins[b] = 0u;
    __syncthreads();
    for (unsigned j = blockIdx.x * blockDim.x + threadIdx.x; j < n; j += gridDim.x * blockDim.x)
    {
        const unsigned key = wf.keys[j];
        RD_BOUND(key < RD_WF_KEY_BINS, wf.status, continue);
        atomicAdd(&bins[key], 1u);
    }
    __syncthreads();
    for (unsigned b = threadIdx.x; b < RD_WF_KEY_BINS; b += blockDim.x)
        if (bins[b]) atomicAdd(&wf.hist[b], bins[b]);
}

__global__ void rd_k_wf_scan(const RdWfState* __restrict__ wfp)   // one block of RD_WF_KEY_BINS threads
{
    __shared__ unsigned warpSums[32];
    const RdWfState wf = *wfp;
    const unsigned t = threadIdx.x, lane = t & 31u, w = t >> 5;
    const unsigned v = wf.hist[t];
    unsigned x = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { const unsigned y = __shfl_up_sync(0xffffffffu, x, o); if (lane >= (unsigned)o) x += y; }
    if (lane == 31u) warpSums[w] = x;
    __syncthreads();
    if (w == 0) {
        unsigned s = warpSums[lane];
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) { const unsigned y = __shfl_up_sync(0xffffffffu, s, o); if (lane >= (unsigned)o) s += y; }
        warpSums[lane] = s;
    }
    __syncthreads();
    wf.hist[t] = x - v + (w ? warpSums[w - 1] : 0u);   // exclusive prefix = first output position of the bin
}

// Output positions are claimed per BLOCK: the warps of a block add their runs to a shared-memory table first, then one global
// atomic per non-empty bin of the block reserves the block's run. With a handful of distinct keys (C3: 5 rows x materials) a claim
// per warp put ~3 M atomics on a few addresses (0.7 ms for the 33 M entries of the first bounce).
__global__ void __launch_bounds__(256) rd_k_wf_scatter(const RdWfState* __restrict__ wfp)
{
    __shared__ unsigned cnt[RD_WF_KEY_BINS];
    const RdWfState wf = *wfp;
    const unsigned n = wf.counts[wf.cur];
    if (blockIdx.x * blockDim.x >= n) return;   // the grid is sized for the capacity of the queue
    for (unsigned b = threadIdx.x; b < RD_WF_KEY_BINS; b += blockDim.x) cnt[b] = 0u;
    __syncthreads();
    const unsigned j = blockIdx.x * blockDim.x + threadIdx.x;
    const unsigned lane = threadIdx.x & 31u;
    const bool valid = j < n;
    const unsigned act = __ballot_sync(0xffffffffu, valid);
    unsigned key = 0u, within = 0u;
    if (valid) {
        key = wf.keys[j];
        if (key >= RD_WF_KEY_BINS) { RD_BOUND(false, wf.status, (void)0); key = RD_WF_KEY_BINS - 1u; }
        const unsigned same = __match_any_sync(act, key);
        const unsigned leader = (unsigned)__ffs(same) - 1u;
        unsigned base = 0;
        if (lane == leader) base = atomicAdd(&cnt[key], (unsigned)__popc(same));
        base = __shfl_sync(same, base, leader);
        within = base + (unsigned)__popc(same & ((1u << lane) - 1u));
    }
    __syncthreads();
    for (unsigned b = threadIdx.x; b < RD_WF_KEY_BINS; b += blockDim.x) { const unsigned c = cnt[b]; if (c) cnt[b] = atomicAdd(&wf.hist[b], c); }
    __syncthreads();
    if (valid) {
        const unsigned at = cnt[key] + within;
        RD_BOUND(at < n, wf.status, return);
        wf.order[at] = j;
    }
}


// ---- ray reordering before a bounce trace -------------------------------------------------------------
// Bounce rays leave the shade stage in hit-group order: neighbours in the queue start anywhere and point
// anywhere, so a warp of the trace stage touches 32 unrelated parts of the structure. This pass re-sorts the
// queue by (direction octant, coarse origin cell) with the same counting sort: key = octant << 7 | 7-bit cell
// code (3 + 3 bits on the two widest axes of the scene box, 1 bit on the third), rays that start in the same
// region and head the same way become neighbours. The rays are moved (not indexed) into the idle queue buffer.
__global__ void rd_k_wf_ray_keys(const RdScene* __restrict__ scene, const RdWfState* __restrict__ wfp)
{
    const RdWfState wf = *wfp;
    const unsigned n = wf.counts[wf.cur];
    const unsigned j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= n) return;
    const float4 a = __ldg(&((const float4*)wf.rayA[wf.cur])[j]), b = __ldg(&((const float4*)wf.rayB[wf.cur])[j]);
#if RD_NODE_F32
    // scene box = the root node's own box (w[8], w[9]); cell coordinate in [0,8) per axis
    const float4 blo = __ldg((const float4*)&scene->nodes[0].w[8]), bhi = __ldg((const float4*)&scene->nodes[0].w[9]);
    const float wx = fmaxf(bhi.x - blo.x, 1e-30f), wy = fmaxf(bhi.y - blo.y, 1e-30f), wz = fmaxf(bhi.z - blo.z, 1e-30f);
    const float fx = (a.x - blo.x) / wx * 8.0f, fy = (a.y - blo.y) / wy * 8.0f, fz = (a.z - blo.z) / wz * 8.0f;
    const float ex = wx, ey = wy, ez = wz;   // extents decide which axis keeps one bit
#else
    // scene box from the root node of the top-level structure: origin p, 256 grid steps of 2^(e-127) per axis
    const uint4 n0 = __ldg((const uint4*)&scene->nodes[0].w[0]);
    const float px = __uint_as_float(n0.x), py = __uint_as_float(n0.y), pz = __uint_as_float(n0.z);
    const unsigned ex = n0.w & 0xffu, ey = (n0.w >> 8) & 0xffu, ez = (n0.w >> 16) & 0xffu;
    // cell coordinate in [0,8) per axis
    const float fx = (a.x - px) * __uint_as_float((254u - ex) << 23) * (1.0f / 32.0f);
    const float fy = (a.y - py) * __uint_as_float((254u - ey) << 23) * (1.0f / 32.0f);
    const float fz = (a.z - pz) * __uint_as_float((254u - ez) << 23) * (1.0f / 32.0f);
#endif
    unsigned cx = (unsigned)fminf(fmaxf(fx, 0.0f), 7.0f), cy = (unsigned)fminf(fmaxf(fy, 0.0f), 7.0f), cz = (unsigned)fminf(fmaxf(fz, 0.0f), 7.0f);
    // the axis with the smallest extent keeps one bit
    unsigned c0, c1, c2;   // c0, c1: 3 bits each; c2: 1 bit
    if (ey <= ex && ey <= ez) { c0 = cx; c1 = cz; c2 = cy >> 2; }
    else if (ex <= ey && ex <= ez) { c0 = cy; c1 = cz; c2 = cx >> 2; }
    else { c0 = cx; c1 = cy; c2 = cz >> 2; }
    // interleave c0 / c1 (Morton order keeps neighbouring cells close in the queue)
    unsigned m = 0;
#pragma unroll
    for (int k = 0; k < 3; ++k) m |= (((c0 >> k) & 1u) << (2 * k)) | (((c1 >> k) & 1u) << (2 * k + 1));
    const unsigned oct = (b.x < 0.0f ? 4u : 0u) | (b.y < 0.0f ? 2u : 0u) | (b.z < 0.0f ? 1u : 0u);
    wf.keys[j] = (oct << 7) | (c2 << 6) | m;
}

__global__ void rd_k_wf_ray_scatter(const RdWfState* __restrict__ wfp)
{
    const RdWfState wf = *wfp;
    const unsigned n = wf.counts[wf.cur];
    const unsigned j = blockIdx.x * blockDim.x + threadIdx.x;
    const unsigned lane = threadIdx.x & 31u;
    const bool valid = j < n;
    const unsigned act = __ballot_sync(0xffffffffu, valid);
    if (!valid) return;
    const unsigned key = wf.keys[j];
    const unsigned same = __match_any_sync(act, key);
    const unsigned leader = (unsigned)__ffs(same) - 1u;
    unsigned base = 0;
    if (lane == leader) base = atomicAdd(&wf.hist[key], (unsigned)__popc(same));
    base = __shfl_sync(same, base, leader);
    const unsigned pos = base + (unsigned)__popc(same & ((1u << lane) - 1u));
    const unsigned q = wf.cur, o = wf.cur ^ 1u;
    ((float4*)wf.rayA[o])[pos] = ((const float4*)wf.rayA[q])[j];
    ((float4*)wf.rayB[o])[pos] = ((const float4*)wf.rayB[q])[j];
    wf.pathOf[o][pos] = wf.pathOf[q][j];
}

// per-stage state edits on the device (keeps the stage sequence free of host synchronisation)
__global__ void rd_k_wf_advance(RdWfState* wf, unsigned op, unsigned a, unsigned b)
{
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    if (op == 0) { wf->frameID = a; wf->samples = b; wf->depth = 0; wf->cur = 0; wf->counts[2] = 0; wf->counts[3] = 0; }   // new pass: first frame id, samples in flight
    else if (op == 1) { wf->counts[2] = 0; }                                                           // before trace
    else if (op == 3) { const unsigned n = wf->counts[wf->cur]; wf->counts[wf->cur] = 0; wf->cur ^= 1u; wf->counts[wf->cur] = n; }   // after a ray reorder
    else if (op == 2) { wf->cur ^= 1u; wf->counts[wf->cur ^ 1u] = 0; wf->depth += 1; wf->counts[2] = 0; wf->counts[3] = 0; } // after shade
}

// ---- launch wrappers used by the stage sequencer in rd_program.cu -------------------------------
int rd_wf_reserve(rd_platform* p, unsigned count, unsigned payloadStride)
{
    RdWfBuffers& B = rd_wf_buffers(p);
    if (B.capacity >= count && B.payloadStride >= payloadStride && B.dstate) return RD_OK;
    RD_CUDA(cudaStreamSynchronize(rd_wf_stream(p)));
    void** all[] = {(void**)&B.payload, (void**)&B.color, (void**)&B.contrib, (void**)&B.rayA[0], (void**)&B.rayA[1], (void**)&B.rayB[0],
                    (void**)&B.rayB[1], (void**)&B.pathOf[0], (void**)&B.pathOf[1], (void**)&B.hitA, (void**)&B.hitB, (void**)&B.keys,
                    (void**)&B.order, (void**)&B.counts, (void**)&B.hist, (void**)&B.dstate};
    for (void** q : all) { if (*q) cudaFree(*q); *q = 0; }
    const size_t n = count ? count : 1;
    RD_CUDA(cudaMalloc(&B.payload, n * payloadStride));
    RD_CUDA(cudaMalloc(&B.color, n * 16)); RD_CUDA(cudaMalloc(&B.contrib, n * 16));
    for (int q = 0; q < 2; ++q) {
        RD_CUDA(cudaMalloc(&B.rayA[q], n * 16)); RD_CUDA(cudaMalloc(&B.rayB[q], n * 16)); RD_CUDA(cudaMalloc(&B.pathOf[q], n * 4));
    }
    RD_CUDA(cudaMalloc(&B.hitA, n * 16)); RD_CUDA(cudaMalloc(&B.hitB, n * 4));
    RD_CUDA(cudaMalloc(&B.keys, n * 4)); RD_CUDA(cudaMalloc(&B.order, n * 4));
    RD_CUDA(cudaMalloc(&B.counts, 64)); RD_CUDA(cudaMalloc(&B.hist, RD_WF_KEY_BINS * 4));
    RD_CUDA(cudaMalloc(&B.dstate, sizeof(RdWfState)));
    RD_CUDA(cudaMemset(B.counts, 0, 64));
    B.capacity = count; B.payloadStride = payloadStride;
    return RD_OK;
}

void rd_wf_release(rd_platform* p)
{
    for (int lane = 0; lane < RD_WF_MAX_LANES; ++lane) {
        RdWfBuffers& B = lane ? p->wfX[lane - 1] : p->wf;
        void* all[] = {B.payload, B.color, B.contrib, B.rayA[0], B.rayA[1], B.rayB[0], B.rayB[1], B.pathOf[0], B.pathOf[1],
                       B.hitA, B.hitB, B.keys, B.order, B.counts, B.hist, B.dstate};
        for (void* q : all) if (q) cudaFree(q);
        memset(&B, 0, sizeof(B));
    }
}

int rd_wf_upload_state(rd_platform* p, const RdWfState* host)
{
    RD_CUDA(cudaMemcpyAsync(rd_wf_buffers(p).dstate, host, sizeof(RdWfState), cudaMemcpyHostToDevice, rd_wf_stream(p)));
    return RD_OK;
}

int rd_wf_advance(rd_platform* p, unsigned op, unsigned a, unsigned b)
{
    rd_k_wf_advance<<<1, 32, 0, rd_wf_stream(p)>>>((RdWfState*)rd_wf_buffers(p).dstate, op, a, b);
    p->launches += 1;
    RD_CUDA(cudaGetLastError());
    return RD_OK;
}

int rd_wf_trace(rd_platform* p, const void* scene, int dynamicFetch)
{
    if (dynamicFetch) {
        static int perSm = -1, minb = 7;
        if (perSm < 0) {
            const char* e = getenv("RD_WF_BLOCKS_PER_SM"); perSm = e ? atoi(e) : 7;
            const char* m = getenv("RD_WF_MINB"); minb = m ? atoi(m) : 7;
        }
        if (minb >= 8)
            rd_k_wf_trace_dyn<8><<<(unsigned)p->smCount * (unsigned)perSm, 128, 0, rd_wf_stream(p)>>>((const RdScene*)scene, (const RdWfState*)rd_wf_buffers(p).dstate);
        else
            rd_k_wf_trace_dyn<7><<<(unsigned)p->smCount * (unsigned)perSm, 128, 0, rd_wf_stream(p)>>>((const RdScene*)scene, (const RdWfState*)rd_wf_buffers(p).dstate);
    }
    else
        rd_k_wf_trace<<<(unsigned)p->smCount * 12u, 128, 0, rd_wf_stream(p)>>>((const RdScene*)scene, (const RdWfState*)rd_wf_buffers(p).dstate);
    p->launches += 1;
    RD_CUDA(cudaGetLastError());
    return RD_OK;
}

int rd_wf_reorder_rays(rd_platform* p, const void* scene, unsigned count)
{
    const RdWfState* S = (const RdWfState*)rd_wf_buffers(p).dstate;
    RD_CUDA(cudaMemsetAsync(rd_wf_buffers(p).hist, 0, RD_WF_KEY_BINS * 4, rd_wf_stream(p)));
    unsigned hb = (count + 255u) / 256u; if (hb > (unsigned)p->smCount * 8u) hb = (unsigned)p->smCount * 8u; if (!hb) hb = 1;
    rd_k_wf_ray_keys<<<(count + 255u) / 256u, 256, 0, rd_wf_stream(p)>>>((const RdScene*)scene, S);
    rd_k_wf_hist<<<hb, 256, 0, rd_wf_stream(p)>>>(S);
    rd_k_wf_scan<<<1, RD_WF_KEY_BINS, 0, rd_wf_stream(p)>>>(S);
    rd_k_wf_ray_scatter<<<(count + 255u) / 256u, 256, 0, rd_wf_stream(p)>>>(S);
    rd_k_wf_advance<<<1, 32, 0, rd_wf_stream(p)>>>((RdWfState*)rd_wf_buffers(p).dstate, 3u, 0u, 0u);
    p->launches += 5;
    RD_CUDA(cudaGetLastError());
    return RD_OK;
}

int rd_wf_sort(rd_platform* p, unsigned count)
{
    const RdWfState* S = (const RdWfState*)rd_wf_buffers(p).dstate;
    RD_CUDA(cudaMemsetAsync(rd_wf_buffers(p).hist, 0, RD_WF_KEY_BINS * 4, rd_wf_stream(p)));
    unsigned hb = (count + 255u) / 256u; if (hb > (unsigned)p->smCount * 8u) hb = (unsigned)p->smCount * 8u; if (!hb) hb = 1;
    rd_k_wf_hist<<<hb, 256, 0, rd_wf_stream(p)>>>(S);
    rd_k_wf_scan<<<1, RD_WF_KEY_BINS, 0, rd_wf_stream(p)>>>(S);
    rd_k_wf_scatter<<<(count + 255u) / 256u, 256, 0, rd_wf_stream(p)>>>(S);
    p->launches += 3;
    RD_CUDA(cudaGetLastError());
    return RD_OK;
}
