// rd_wavefront.h -- state shared by the wavefront stages (library kernels in rd_wavefront.cu and the
// generate / shade / resolve kernels that rd_program.cu appends to the user program for NVRTC).
//
// Wavefront mode runs the sample loop and the bounce loop that the reference spells out inside its
// raygen kernel (samples/shader.cl:196-281) as a sequence of stage launches over queues in HBM:
//   [ generate -> [ trace -> sort by hit group / material -> shade ] x depth -> accumulate ] x passes -> resolve
// A pass carries `samples` samples of every work-item at once (option wf_paths caps paths in flight;
// HBM holds tens of millions of paths, and the number of stage launches -- each of which ends with the
// tail of its longest rays -- shrinks by that factor). A path owns one slot = sample * count +
// launch-local work-item index; the user's `struct Payload` lives in the payload array of that slot
// between stages. Finished samples stay in `color` and are folded into imageScratch in frame order
// by the accumulate stage, so the running mean is the same float sequence as the reference's loop. Ray / hit queues are structure-of-arrays, 16-byte
// records, compacted every bounce (only live paths are traced and shaded).
#ifndef RD_WAVEFRONT_H
#define RD_WAVEFRONT_H
#include "rd_types.h"

#define RD_WF_KEY_BINS 1024u      // (hit group row mod 256) * 4 + (instanceCustomIndex & 3): a coherence key, not an identity

struct RD_ALIGN(16) RdWfState {
    // per path slot
    unsigned char* payload;       // slot * payloadStride: the user's struct Payload
    RdF4* color;                  // accumulated sample colour (shader.cl:228 `color`)
    RdF4* contrib;                // path throughput (shader.cl:229 `contribution`)
    // compact queues, double buffered: entry j of queue q
    RdF4* rayA[2];                // { o.xyz, tmin }
    RdF4* rayB[2];                // { d.xyz, tmax }
    rd_u32* pathOf[2];            // slot of the path the ray belongs to
    RdF4* hitA;                   // { t, b1, b2, primID bits } of queue `cur`
    rd_u32* hitB;                 // instanceID
    rd_u32* keys;                 // sort key of entry j
    rd_u32* order;                // entries of queue `cur` in key order (0 = identity)
    rd_u32* counts;               // [0], [1] queue sizes; [2] batch fetch counter of the trace kernel
    rd_u32* hist;                 // RD_WF_KEY_BINS bin counters / cursors
    rd_u32 payloadStride, cur;
    rd_u32 first, count;          // work-item range of this TraceRays call (slot s = work-item first + s)
    rd_u32 total, tileW;          // NDRange size and the width of the 8x4 tile remap (0 = identity)
    rd_u32 frameID, totalSamples; // sample being traced, RTProp->totalSamples
    rd_u32 depth, maxDepth;       // bounce index of this stage, RTProp->depth
    rd_u32 debug, sortOn;
    rd_u32 ilWorld, ilRank, ilBandRows;   // interleaved multi-GPU partition (0 = whole frame), see rd_trace_rays_interleaved
    rd_u32 samples;               // samples in flight in this pass: path slot = sample * count + work-item, frame id = frameID + sample
    int sbtRecordOffset, missIndex;
    float tmin, tmax;
    rd_u32* status;               // host-mapped platform flags; [2] is raised by a bounds violation in RD_BOUNDS_CHECK builds
    rd_u32 capacity, pad;         // path slots / queue entries the buffers above were allocated for
};

// RD_BOUNDS_CHECK=1 builds (libradiance_b200_bounds.so, tests/test_gpu_bounds.py): every queue / order / record index the
// wavefront stages and the walk compute is checked against the size of what it indexes before the access; a violation
// raises status[2] (a plain store: an atomicAdd on the host-mapped word made the instrumented kernels hang although it never
// executed -- profiles/README_r2.md) and the access is skipped. compute-sanitizer is closed on this pool, this is the memcheck stand-in.
#ifndef RD_BOUNDS_CHECK
#define RD_BOUNDS_CHECK 0
#endif
#if RD_BOUNDS_CHECK && defined(__CUDA_ARCH__)
#define RD_BOUND(ok, st, action) do { if (!(ok)) { if (st) (st)[2] = 1u; action; } } while (0)
#else
#define RD_BOUND(ok, st, action) do { } while (0)
#endif

// work-item t of the launch -> global id (pixel index), same mapping as get_global_id in cl_compat.h
RD_HD rd_u32 rd_wf_global_id(rd_u32 t, rd_u32 tileW, rd_u32 world = 0, rd_u32 rank = 0, rd_u32 bandRows = 1)
{
    if (!tileW) return t;
    const rd_u32 tilesX = tileW >> 3, tile = t >> 5, lane = t & 31u;
    rd_u32 ty = tile / tilesX;
    const rd_u32 tx = tile - ty * tilesX;
    if (world > 1u) { const rd_u32 band = ty / bandRows; ty = (band * world + rank) * bandRows + (ty - band * bandRows); }
    return ((ty << 2) + (lane >> 3)) * tileW + (tx << 3) + (lane & 7u);
}

#endif
