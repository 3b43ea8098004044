// rd_types.h -- device data layout of the B200 tracer (shared by nvcc, NVRTC and the host-side
// unit-test build). Everything here is plain data; see DESIGN.md "Data layout in HBM".
//
// Replaces the reference's single-blob binary BVH (radiance/src/core.h:34-101,
// radiance/shader/data.cl:4-99): 48-byte binary nodes + indexed 16-byte faces + 16-byte vertices
// addressed by 32-bit byte offsets become
//   * RdNode8   80 B  8-wide node, child boxes quantised to 8 bits on a per-node power-of-two grid (a 256-byte form with
//                     plain float boxes is kept behind RD_NODE_F32=1 for the A/B measurement)
//   * RdTri     48 B  pre-gathered triangle (3 x float4, primID in v0.w), stored in leaf order
//   * RdInstance 160 B world->object 3x4 (precomputed once, the reference inverts a 4x4 per ray per
//                     instance visit: radiance/shader/radiance.cl:166-167), object->world 4x4, ids,
//                     64-bit pointers to the bottom-level arrays (no 4 GiB ceiling)
#ifndef RD_TYPES_H
#define RD_TYPES_H

#if defined(__CUDACC_RTC__)
typedef unsigned int rd_u32;
typedef unsigned long long rd_u64;
typedef unsigned char rd_u8;
#else
#include <stdint.h>
typedef uint32_t rd_u32;
typedef uint64_t rd_u64;
typedef uint8_t rd_u8;
#endif

#if defined(__CUDACC__) || defined(__CUDACC_RTC__)
#define RD_HD __host__ __device__ __forceinline__
#define RD_ALIGN(n) __align__(n)
#else
#define RD_HD inline
#define RD_ALIGN(n) __attribute__((aligned(n)))
#endif

struct RD_ALIGN(16) RdU4 { rd_u32 x, y, z, w; };
struct RD_ALIGN(16) RdF4 { float x, y, z, w; };

#ifndef RD_NODE_F32
#define RD_NODE_F32 0   // 0: 80-byte nodes, child boxes quantised to 8 bits (default); 1: 256-byte nodes with plain float child boxes.
                        // A/B on the B200 (profiles/README_r2.md): float boxes save the 48 conversions of a node test and cull
                        // exactly, but move 3.2x the bytes -- C2 (1 M triangles, L2-resident) 4644 vs 4540 Mrays/s at 72 registers,
                        // 10 M triangles 2812 vs 3000: the walk over a structure larger than L2 is bound by load latency as much
                        // as by instruction issue, so the small node stays the default
#endif

#if RD_NODE_F32
// 8-wide BVH node, 256 B (16 x 16 B), child boxes as plain floats in structure-of-arrays order. The walk is bound by
// instruction issue, not by bytes (profiles/README_r1.md: DRAM at 10 % of peak with 80-byte nodes), so the node trades
// bytes for instructions: no dequantisation (48 conversions + the grid set-up per node test), the near / far planes of an
// axis are picked by ADDRESS from the ray's direction signs instead of by 48 selects.
//  w[0]  = { childBase, primBase, leafNib, imask }
//          inner child in slot s lives at node index childBase + popc(imask & ((1<<s)-1))
//          leafNib nibble s = unary primitive count (1..3 -> 0b0001 / 0b0011 / 0b0111) of a leaf slot, 0 otherwise;
//          bit b of leafNib set  <=>  primitive record primBase + popc(leafNib & ((1<<b)-1))
//  w[1]  = { innerNib (0x1 in nibble s of an inner slot), 0, 0, 0 }
//  w[2..3] lo.x[0..7]   w[4..5] lo.y   w[6..7] lo.z        byte offsets  32 /  64 /  96
//  w[8]  = { node box lo.xyz, 0 }   w[9] = { node box hi.xyz, 0 }
//  w[10..11] hi.x[0..7] w[12..13] hi.y w[14..15] hi.z      byte offsets 160 / 192 / 224
// Empty slots hold an inverted box and have no bit in innerNib / leafNib.
struct RD_ALIGN(16) RdNode8 { RdU4 w[16]; };
#define RD_NODE_OFF_LOX 32u
#define RD_NODE_OFF_LOY 64u
#define RD_NODE_OFF_LOZ 96u
#define RD_NODE_OFF_HIX 160u
#define RD_NODE_OFF_HIY 192u
#define RD_NODE_OFF_HIZ 224u
#else
// 8-wide BVH node, 5 x 16 B.
//  w[0] = { p.x, p.y, p.z (float bits),  ex | ey<<8 | ez<<16 | imask<<24 }
//         p = lower corner of the node box, e* = biased float exponents of the grid step per axis,
//         imask bit s set <=> slot s holds an inner child
//  w[1] = { childBase, primBase, leafNib, innerNib }
//         inner child in slot s lives at node index childBase + popc(imask & ((1<<s)-1))
//         leafNib nibble s = unary primitive count of a leaf slot (0b0001 / 0b0011 / 0b0111), bit b set <=> primitive
//         record primBase + popc(leafNib & ((1<<b)-1)); innerNib = 0x1 in the nibble of every inner slot
//  w[2] = { qlo.x[0..3], qlo.x[4..7], qlo.y[0..3], qlo.y[4..7] }
//  w[3] = { qlo.z[0..3], qlo.z[4..7], qhi.x[0..3], qhi.x[4..7] }
//  w[4] = { qhi.y[0..3], qhi.y[4..7], qhi.z[0..3], qhi.z[4..7] }
struct RD_ALIGN(16) RdNode8 { RdU4 w[5]; };
#endif

// Triangle record: v0.w carries the primitive index (bits), v1.w / v2.w are zero.
struct RD_ALIGN(16) RdTri { RdF4 v0, v1, v2; };

struct RD_ALIGN(16) RdInstance {
    float inv[12];            // world->object, rows 0..2 of the 4x4 inverse
    float fwd[16];            // object->world, row-major 4x4 exactly as the caller gave it
    rd_u32 sbtOffset, instanceID, customID, flags;
    const RdNode8* nodes;     // bottom-level node array (root = node 0)
    const RdTri* tris;        // bottom-level triangle records
    float maxAbs;             // largest |coordinate| of the bottom-level root box (rounding margin of the box test)
    rd_u32 pad[3];
};

// Top-level acceleration structure header; lives in HBM, its address is the descriptor that a
// user program passes to traceRay as `topLevel` (reference: `__global struct AccelStruct*`).
struct RD_ALIGN(16) RdScene {
    const RdNode8* nodes;         // top-level nodes over instance boxes (root = node 0)
    const rd_u32* leafInstances;  // instance indices in top-level leaf order
    const RdInstance* instances;  // caller order: instanceID == array index
    rd_u32 instanceCount, nodeCount, flags;
    float maxAbs;                 // largest |coordinate| of the top-level root box
    rd_u32* status;               // host-mapped flags: [0] traversal-stack overflow, [1] traceRay nested too deep (must stay 0)
    rd_u64* rayCounter;           // traceRay invocations (option count_rays), device memory
};

// Ray / hit records of the wavefront queues (structure of arrays, 16-byte vector loads):
//   rayA[i] = { o.x, o.y, o.z, tmin }   rayB[i] = { d.x, d.y, d.z, tmax }
//   hitA[i] = { t, b1, b2, primID(bits) } (t = FLT_MAX: miss)   hitB[i] = instanceID
#define RD_MISS_T 3.402823466e+38f
#define RD_INVALID 0xffffffffu

// rd_trace flags
#define RD_TRACE_TERMINATE_ON_FIRST_HIT 1u
#define RD_TRACE_COUNTERS 2u
#define RD_TRACE_STATIC_BATCHES 4u   // default behaviour, kept for source compatibility
#define RD_TRACE_DYNAMIC_FETCH 8u

#endif
