// rd_sort.h -- the two data-parallel primitives of the build, written for this library (no CUB / Thrust):
//   rd_exclusive_scan_u32   exclusive prefix sum of 32-bit counters (used by the cluster compaction of PLOC and by the sort)
//   rd_radix_sort_pairs     stable LSD radix sort of (64-bit key, 32-bit value) pairs, 8 bits per pass
// Included by rd_build.cu only (device code; the host emulation of tests/hostemu sorts with std::stable_sort).
//
// Sort, per 8-bit digit: (A) every block counts the digits of its tile of RD_SORT_TILE keys into a column of a
// digit-major table [256][blocks]; (B) one exclusive scan over that table gives, for every (digit, block), where the
// block's run of that digit starts in the output -- digit-major order is what makes the pass a stable counting sort;
// (C) every block ranks its keys again, in input order (warp rounds with __match_any_sync, then a prefix over the
// block's warps), and scatters keys and values. A pass whose digit is the same for all keys is skipped: the
// digit histograms of all 8 passes are taken in one sweep up front (they do not depend on the order).
// Inputs of up to RD_SORT_SMALL_MAX pairs run all passes in one block (the 4096 small meshes of C4).
#ifndef RD_SORT_H
#define RD_SORT_H
#include "rd_types.h"

#define RD_SORT_THREADS 256u
#define RD_SORT_ITEMS 16u
#define RD_SORT_TILE (RD_SORT_THREADS * RD_SORT_ITEMS)
#define RD_SORT_SMALL_MAX 8192u
#define RD_SCAN_THREADS 256u
#define RD_SCAN_ITEMS 8u
#define RD_SCAN_TILE (RD_SCAN_THREADS * RD_SCAN_ITEMS)

// ---- block-wide exclusive scan of one value per thread (blockDim.x <= 1024); returns the exclusive prefix, total in *sum
__device__ __forceinline__ rd_u32 rd_block_exclusive_scan(rd_u32 v, rd_u32* warpSums /* [32] shared */, rd_u32* sum)
{
    const rd_u32 lane = threadIdx.x & 31u, warp = threadIdx.x >> 5, nwarps = (blockDim.x + 31u) >> 5;
    rd_u32 x = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { const rd_u32 y = __shfl_up_sync(0xffffffffu, x, o); if (lane >= (rd_u32)o) x += y; }
    __syncthreads();                      // warpSums may still be read from a previous call
    if (lane == 31u) warpSums[warp] = x;
    __syncthreads();
    if (warp == 0) {
        rd_u32 s = lane < nwarps ? warpSums[lane] : 0u;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) { const rd_u32 y = __shfl_up_sync(0xffffffffu, s, o); if (lane >= (rd_u32)o) s += y; }
        warpSums[lane] = s;
    }
    __syncthreads();
    *sum = warpSums[nwarps - 1u];
    return x - v + (warp ? warpSums[warp - 1u] : 0u);
}

// ---- exclusive scan of n counters: tile sums -> scan of the tile sums (recursive) -> tile-local scan + offset
__global__ void rd_k_scan_tile_sums(const rd_u32* __restrict__ in, rd_u32 n, rd_u32* __restrict__ tileSums)
{
    __shared__ rd_u32 warpSums[32];
    const rd_u32 base = blockIdx.x * RD_SCAN_TILE + threadIdx.x * RD_SCAN_ITEMS;
    rd_u32 s = 0;
#pragma unroll
    for (rd_u32 k = 0; k < RD_SCAN_ITEMS; ++k) if (base + k < n) s += in[base + k];
    rd_u32 total;
    rd_block_exclusive_scan(s, warpSums, &total);
    if (threadIdx.x == 0) tileSums[blockIdx.x] = total;
}

__global__ void rd_k_scan_tiles(const rd_u32* __restrict__ in, rd_u32* __restrict__ out, rd_u32 n, const rd_u32* __restrict__ tileOffsets,
                                rd_u32* __restrict__ total /* may be 0 */)
{
    __shared__ rd_u32 warpSums[32];
    const rd_u32 base = blockIdx.x * RD_SCAN_TILE + threadIdx.x * RD_SCAN_ITEMS;
    rd_u32 v[RD_SCAN_ITEMS], s = 0;
#pragma unroll
    for (rd_u32 k = 0; k < RD_SCAN_ITEMS; ++k) { v[k] = base + k < n ? in[base + k] : 0u; s += v[k]; }
    rd_u32 blockTotal;
    rd_u32 run = rd_block_exclusive_scan(s, warpSums, &blockTotal) + (tileOffsets ? tileOffsets[blockIdx.x] : 0u);
#pragma unroll
    for (rd_u32 k = 0; k < RD_SCAN_ITEMS; ++k) { if (base + k < n) out[base + k] = run; run += v[k]; }
    if (total && blockIdx.x == gridDim.x - 1u && threadIdx.x == blockDim.x - 1u) *total = run;   // the last thread ends on the grand total
}

// scratch: at least rd_scan_scratch_count(n) counters
static inline size_t rd_scan_scratch_count(size_t n)
{
    size_t c = 0;
    while (n > RD_SCAN_TILE) { n = (n + RD_SCAN_TILE - 1) / RD_SCAN_TILE; c += n; }
    return c + 1;
}

// out may alias in. `total` (device pointer or 0) receives the sum of all n counters.
static inline int rd_exclusive_scan_u32(cudaStream_t st, const rd_u32* in, rd_u32* out, rd_u32 n, rd_u32* scratch, rd_u32* total, unsigned long long* launches)
{
    if (!n) return 0;
    const rd_u32 tiles = (n + RD_SCAN_TILE - 1u) / RD_SCAN_TILE;
    if (tiles == 1u) {
        rd_k_scan_tiles<<<1, RD_SCAN_THREADS, 0, st>>>(in, out, n, 0, total);
        *launches += 1;
        return 0;
    }
    rd_k_scan_tile_sums<<<tiles, RD_SCAN_THREADS, 0, st>>>(in, n, scratch);
    *launches += 1;
    rd_exclusive_scan_u32(st, scratch, scratch, tiles, scratch + tiles, 0, launches);
    rd_k_scan_tiles<<<tiles, RD_SCAN_THREADS, 0, st>>>(in, out, n, scratch, total);
    *launches += 1;
    return 0;
}

// ---- radix sort ---------------------------------------------------------------------------------------------------
// digit histograms of all 8 passes in one sweep: hist[pass * 256 + digit]
__global__ void rd_k_sort_histograms(const rd_u64* __restrict__ keys, rd_u32 n, rd_u32* __restrict__ hist)
{
    __shared__ rd_u32 h[8 * 256];
    for (rd_u32 i = threadIdx.x; i < 8u * 256u; i += blockDim.x) h[i] = 0u;
    __syncthreads();
    for (rd_u32 i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const rd_u64 k = keys[i];
#pragma unroll
        for (int p = 0; p < 8; ++p) atomicAdd(&h[p * 256 + (int)((k >> (8 * p)) & 0xffull)], 1u);
    }
    __syncthreads();
    for (rd_u32 i = threadIdx.x; i < 8u * 256u; i += blockDim.x) if (h[i]) atomicAdd(&hist[i], h[i]);
}

// (A) digit counts of one tile -> table[digit * blocks + block]
__global__ void __launch_bounds__(RD_SORT_THREADS) rd_k_sort_count(const rd_u64* __restrict__ keys, rd_u32 n, int shift, rd_u32* __restrict__ table)
{
    __shared__ rd_u32 h[256];
    h[threadIdx.x] = 0u;
    __syncthreads();
    const rd_u32 base = blockIdx.x * RD_SORT_TILE;
#pragma unroll
    for (rd_u32 r = 0; r < RD_SORT_ITEMS; ++r) {
        const rd_u32 i = base + r * RD_SORT_THREADS + threadIdx.x;
        if (i < n) atomicAdd(&h[(rd_u32)((keys[i] >> shift) & 0xffull)], 1u);
    }
    __syncthreads();
    table[threadIdx.x * gridDim.x + blockIdx.x] = h[threadIdx.x];
}

// Stable rank of this thread's key among the keys of the block seen so far with the same digit.
// Warp w owns the keys [w * ITEMS * 32, (w + 1) * ITEMS * 32) of the tile, in rounds of 32 consecutive keys, so
// (warp, round, lane) order is input order. counts[w][d] runs over the rounds of warp w.
__device__ __forceinline__ void rd_sort_rank_tile(const rd_u64* __restrict__ keys, rd_u32 n, rd_u32 base, int shift,
                                                  rd_u32 (*counts)[256], rd_u32 digit[RD_SORT_ITEMS], rd_u32 rank[RD_SORT_ITEMS],
                                                  rd_u64 key[RD_SORT_ITEMS])
{
    const rd_u32 lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
    for (rd_u32 d = lane; d < 256u; d += 32u) counts[warp][d] = 0u;
    __syncwarp();
#pragma unroll
    for (rd_u32 r = 0; r < RD_SORT_ITEMS; ++r) {
        const rd_u32 i = base + (warp * RD_SORT_ITEMS + r) * 32u + lane;
        const bool valid = i < n;
        key[r] = valid ? keys[i] : ~0ull;
        digit[r] = valid ? (rd_u32)((key[r] >> shift) & 0xffull) : 256u;   // 256 = padding, never equal to a real digit
        const unsigned peers = __match_any_sync(0xffffffffu, digit[r]);
        const rd_u32 before = (rd_u32)__popc(peers & ((1u << lane) - 1u));
        rd_u32 start = 0u;
        if (valid) start = counts[warp][digit[r]];
        __syncwarp();
        if (valid && before == 0u) counts[warp][digit[r]] = start + (rd_u32)__popc(peers);   // the first lane of each group advances the counter
        __syncwarp();
        rank[r] = start + before;
    }
}

// (C) scatter one tile. offsets = exclusive scan of the table of (A).
__global__ void __launch_bounds__(RD_SORT_THREADS) rd_k_sort_scatter(const rd_u64* __restrict__ keysIn, const rd_u32* __restrict__ valsIn,
                                                                     rd_u64* __restrict__ keysOut, rd_u32* __restrict__ valsOut, rd_u32 n, int shift,
                                                                     const rd_u32* __restrict__ offsets)
{
    __shared__ rd_u32 counts[RD_SORT_THREADS / 32u][256];
    rd_u32 digit[RD_SORT_ITEMS], rank[RD_SORT_ITEMS];
    rd_u64 key[RD_SORT_ITEMS];
    const rd_u32 base = blockIdx.x * RD_SORT_TILE;
    rd_sort_rank_tile(keysIn, n, base, shift, counts, digit, rank, key);
    __syncthreads();
    {   // per digit: exclusive prefix over the warps, plus the block's global start
        const rd_u32 d = threadIdx.x;
        rd_u32 run = offsets[d * gridDim.x + blockIdx.x];
#pragma unroll
        for (rd_u32 w = 0; w < RD_SORT_THREADS / 32u; ++w) { const rd_u32 c = counts[w][d]; counts[w][d] = run; run += c; }
    }
    __syncthreads();
    const rd_u32 lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
#pragma unroll
    for (rd_u32 r = 0; r < RD_SORT_ITEMS; ++r) {
        const rd_u32 i = base + (warp * RD_SORT_ITEMS + r) * 32u + lane;
        if (i < n) {
            const rd_u32 o = counts[warp][digit[r]] + rank[r];
            keysOut[o] = key[r];
            valsOut[o] = valsIn[i];
        }
    }
}

// all passes in one block, for small inputs; result ends in (keysA, valsA) or (keysB, valsB): returns via *where (0 = A, 1 = B)
// (the body is a device function so that the fused small-mesh build of rd_build.cu can run it as one of its phases;
// every thread of a 1024-thread block must call it)
__device__ __forceinline__ void rd_sort_small_block(rd_u64* keysA, rd_u32* valsA, rd_u64* keysB, rd_u32* valsB, rd_u32 n, int passes, rd_u32* where)
{
    __shared__ rd_u32 counts[32][256];     // [warp][digit] of the current chunk
    __shared__ rd_u32 digitStart[256];     // running output position per digit
    __shared__ rd_u32 warpSums[32];
    const rd_u32 tid = threadIdx.x, lane = tid & 31u, warp = tid >> 5;
    rd_u64* kin = keysA; rd_u32* vin = valsA; rd_u64* kout = keysB; rd_u32* vout = valsB;
    rd_u32 side = 0u;
    for (int p = 0; p < passes; ++p) {
        const int shift = 8 * p;
        // histogram of the digit, then its exclusive scan = where each digit's run starts
        if (tid < 256u) digitStart[tid] = 0u;
        __syncthreads();
        for (rd_u32 i = tid; i < n; i += blockDim.x) atomicAdd(&digitStart[(rd_u32)((kin[i] >> shift) & 0xffull)], 1u);
        __syncthreads();
        rd_u32 total;
        const rd_u32 mine = tid < 256u ? digitStart[tid] : 0u;
        const rd_u32 ex = rd_block_exclusive_scan(mine, warpSums, &total);
        const bool uniform = __syncthreads_or(mine == n);      // every key has the same digit: nothing to move
        if (uniform) continue;
        if (tid < 256u) digitStart[tid] = ex;
        __syncthreads();
        for (rd_u32 c = 0; c < n; c += blockDim.x) {            // chunks in input order
            const rd_u32 i = c + tid;
            const bool valid = i < n;
            const rd_u64 k = valid ? kin[i] : ~0ull;
            const rd_u32 d = valid ? (rd_u32)((k >> shift) & 0xffull) : 256u;
            for (rd_u32 q = lane; q < 256u; q += 32u) counts[warp][q] = 0u;
            __syncwarp();
            const unsigned peers = __match_any_sync(0xffffffffu, d);
            const rd_u32 before = (rd_u32)__popc(peers & ((1u << lane) - 1u));
            if (valid && before == 0u) counts[warp][d] = (rd_u32)__popc(peers);
            __syncthreads();
            if (tid < 256u) {                                    // prefix over the warps, advance the digit's cursor
                rd_u32 run = digitStart[tid];
                for (rd_u32 w = 0; w < 32u; ++w) { const rd_u32 x = counts[w][tid]; counts[w][tid] = run; run += x; }
                digitStart[tid] = run;
            }
            __syncthreads();
            if (valid) { const rd_u32 o = counts[warp][d] + before; kout[o] = k; vout[o] = vin[i]; }
            __syncthreads();
        }
        rd_u64* tk = kin; kin = kout; kout = tk;
        rd_u32* tv = vin; vin = vout; vout = tv;
        side ^= 1u;
        __syncthreads();
    }
    if (tid == 0) *where = side;
}

__global__ void __launch_bounds__(1024) rd_k_sort_small(rd_u64* keysA, rd_u32* valsA, rd_u64* keysB, rd_u32* valsB, rd_u32 n, int passes, rd_u32* where)
{
    rd_sort_small_block(keysA, valsA, keysB, valsB, n, passes, where);
}

struct RdSortScratch {
    rd_u32* hist;      // 8 * 256
    rd_u32* table;     // 256 * blocks
    rd_u32* scan;      // rd_scan_scratch_count(256 * blocks)
    rd_u32* where;     // 1
};
static inline rd_u32 rd_sort_blocks(rd_u32 n) { return (n + RD_SORT_TILE - 1u) / RD_SORT_TILE; }

// Sorts n pairs by the low `bits` bits of the key. Returns 0 when the result is in (keysA, valsA), 1 for (keysB, valsB),
// negative on a CUDA error. Synchronises the stream once (to learn which passes can be skipped).
static inline int rd_radix_sort_pairs(cudaStream_t st, rd_u64* keysA, rd_u32* valsA, rd_u64* keysB, rd_u32* valsB, rd_u32 n, int bits,
                                      const RdSortScratch& S, int smCount, unsigned long long* launches)
{
    const int passes = (bits + 7) / 8;
    if (n <= 1u) return 0;
    if (n <= RD_SORT_SMALL_MAX) {
        rd_k_sort_small<<<1, 1024, 0, st>>>(keysA, valsA, keysB, valsB, n, passes, S.where);
        *launches += 1;
        rd_u32 side = 0;
        if (cudaMemcpyAsync(&side, S.where, 4, cudaMemcpyDeviceToHost, st) != cudaSuccess || cudaStreamSynchronize(st) != cudaSuccess) return -1;
        return (int)side;
    }
    if (cudaMemsetAsync(S.hist, 0, 8 * 256 * sizeof(rd_u32), st) != cudaSuccess) return -1;
    rd_k_sort_histograms<<<smCount * 4, 256, 0, st>>>(keysA, n, S.hist);
    *launches += 1;
    rd_u32 hist[8 * 256];
    if (cudaMemcpyAsync(hist, S.hist, sizeof(hist), cudaMemcpyDeviceToHost, st) != cudaSuccess || cudaStreamSynchronize(st) != cudaSuccess) return -1;
    const rd_u32 blocks = rd_sort_blocks(n);
    int side = 0;
    for (int p = 0; p < passes; ++p) {
        bool uniform = false;
        for (int d = 0; d < 256; ++d) if (hist[p * 256 + d] == n) uniform = true;
        if (uniform) continue;
        rd_u64* kin = side ? keysB : keysA; rd_u32* vin = side ? valsB : valsA;
        rd_u64* kout = side ? keysA : keysB; rd_u32* vout = side ? valsA : valsB;
        rd_k_sort_count<<<blocks, RD_SORT_THREADS, 0, st>>>(kin, n, 8 * p, S.table);
        *launches += 1;
        rd_exclusive_scan_u32(st, S.table, S.table, 256u * blocks, S.scan, 0, launches);
        rd_k_sort_scatter<<<blocks, RD_SORT_THREADS, 0, st>>>(kin, vin, kout, vout, n, 8 * p, S.table);
        *launches += 1;
        side ^= 1;
    }
    return cudaGetLastError() == cudaSuccess ? side : -1;
}

#endif
