// rd_internal.h -- host-side objects behind the C ABI (include/radiance_b200.h).
#ifndef RD_INTERNAL_H
#define RD_INTERNAL_H
#include <cuda_runtime.h>
#include <cuda.h>
#include <stdint.h>
#include <string>
#include <vector>
#include "../../include/radiance_b200.h"
#include "rd_types.h"

enum RdHandleKind { RD_H_BUFFER = 1, RD_H_IMAGE_ARRAY = 2, RD_H_SAMPLER = 3, RD_H_TLAS = 4 };

// device-side image descriptor, must match rd_image_array_desc in shader/cl_compat.h
struct RdImageArrayDesc { const unsigned char* texels; unsigned width, height, layers, pad; };

struct RdHandleObj {
    uint32_t magic;          // 'RDHO'
    RdHandleKind kind;
    void* dptr;              // buffer storage / texel storage / RdScene*
    size_t size;             // bytes of dptr
    uint64_t arg;            // the 8-byte kernel parameter this handle binds as
    // image array
    unsigned width, height, layers;
    void* ddesc;             // device RdImageArrayDesc
    // top-level structure
    void* dnodes; void* dleaf; void* dinst;
    unsigned instanceCount, nodeCount;
    int external;            // dptr is owned by the caller (rd_buffer_wrap)
};
#define RD_HANDLE_MAGIC 0x4f484452u

// Device memory of small bottom-level structures comes out of big chunks (rd_build.cu): a scene of thousands of small meshes
// otherwise makes tens of thousands of stream-ordered allocations, and from a cold pool those cost more than the builds.
// kind 0 = staging (worst-case node arrays of unresolved builds), kind 1 = final (triangle records, exact node arrays).
struct RdArenaChunk { char* base; size_t size, used; unsigned live; int kind; cudaEvent_t idle; bool idleRecorded; };

struct rd_blas {
    RdNode8* nodes; RdTri* tris;
    RdArenaChunk *nodesChunk, *trisChunk;   // 0: the array is an allocation of its own (cudaFree)
    uint32_t nodeCount, triCount, levels;
    float lo[3], hi[3];
    // deferred small build (rd_build.cu): node count / root box not read back yet
    int pending, failed; cudaEvent_t done; volatile uint32_t* summary; struct rd_platform* owner;
};
int rd_blas_resolve(struct rd_platform* p, rd_blas* b);
void rd_arena_trim(struct rd_platform* p);   // rd_build.cu: release the chunks nothing lives in and the per-stream build scratch

struct rd_module {
    CUmodule mod; CUfunction raygen;
    CUfunction stealRaygen;                      // persistent raygen with chunk stealing (option "stealing")
    CUfunction bandedRaygen;                     // raygen whose blocks report band completion (option "stream_image")
    CUfunction wfGenerate, wfShade, wfAccumulate, wfResolve;   // 0 when the program has no wavefront hooks
    unsigned wfPayloadSize; float wfRay[4];      // sizeof(struct Payload); sbtRecordOffset, missIndex, Tmin, Tmax
    int numRegs; int localBytes;
    std::string ptxInfo;
};

// device buffers of wavefront mode (rd_wavefront.h), grown on demand
struct RdWfBuffers {
    unsigned char* payload; void *color, *contrib, *rayA[2], *rayB[2]; unsigned *pathOf[2];
    void* hitA; unsigned *hitB, *keys, *order, *counts, *hist; void* dstate;
    unsigned capacity, payloadStride;
};

int rd_check_status(struct rd_platform* p, const char* where);   // rd_core.cu

#define RD_WF_MAX_LANES 4
struct rd_platform {
    int device;
    cudaStream_t stream;
    volatile uint32_t* hstatus;   // host-mapped: [0] stack overflow flag, [1] traceRay nesting exceeded -- the device only ever writes them
                              // when something went wrong, the host reads them for free at every call that synchronises (rd_check_status)
    uint32_t* dstatus;        // the device address of hstatus (RdScene::status)
    unsigned long long* dRayCounter;   // 64-bit traceRay counter (option count_rays), device memory
    unsigned long long* dcounters;   // trace counters [nodes, tris, rays] + work counter at [4]
    std::vector<uint64_t> boundArgs;
    std::vector<int> boundKinds;
    rd_module* boundModule;
    std::string sbtJson;
    long optFmad, optTile, optBlock, optCountRays, optMaxDepth, optWavefront, optWfSort, optSteal, optWfDynamic, optWfRaySort, optWfPaths;
    long optStreamImage;                         // modules get the band-reporting raygen (rd_buffer_stream_to_host)
    // streamed read-back (rd_buffer_stream_to_host): bands of the frame are copied to the host mirror while later bands render
    RdHandleObj* mirrorObj; void* mirrorHost; unsigned mirrorItemBytes, mirrorBands; size_t mirrorValidBytes; int mirrorFresh, copyPending;
    cudaStream_t copyStream; unsigned* bandDone; unsigned bandDoneCap;
    unsigned bandLayoutW, bandLayoutH, bandLayoutN, bandLayoutBlock, bandEpoch;
    volatile unsigned* bandFlags; unsigned* bandFlagsDev;   // host-mapped: band b of launch e is complete when bandFlags[b] == e
    unsigned bandCount, bandRows, bandTileRows; uint64_t bandCopied; size_t bandRowBytes;   // bandCopied: bit b = copy of band b queued
    unsigned ilWorld, ilRank, ilBandRows;        // interleaved multi-GPU partition of the launch being issued
    unsigned* stealCounters; unsigned* stealOwned; int stealImported; unsigned stealSlots, stealOwnedCap;
    RdWfBuffers wf;
    // further wavefront lanes (option "wf_lanes" = 2..4): the sample passes of a frame go round robin over that many sets of
    // queues, each on its own stream, so the tail of one pass's stage (its longest rays) overlaps the other passes' stages;
    // wfLane selects the set the rd_wf_* launch wrappers work on (0 = wf on the platform stream)
    RdWfBuffers wfX[RD_WF_MAX_LANES - 1]; cudaStream_t wfStreamX[RD_WF_MAX_LANES - 1]; cudaEvent_t wfEvents[RD_WF_MAX_LANES + 1]; int wfLane; long optWfLanes;
    std::vector<cudaStream_t> buildStreams; size_t buildNext;   // side streams of deferred small bottom-level builds
    std::vector<uint32_t*> summaryHost, summaryDev; size_t summaryUsed;   // host-mapped build summaries, RD_SUMMARY_CHUNK per chunk
    long optAsyncBuild, optPrecDiv, optL2Persist;
    // page-locked staging ring of the small-mesh uploads (rd_build.cu): RD_UPLOAD_SEGMENTS segments, a segment is reused once the
    // fence events recorded on every side stream when it was left have fired
    std::vector<RdArenaChunk*> arenaChunks; RdArenaChunk* arenaCur[2];
    std::vector<void*> buildScratch; std::vector<size_t> buildScratchBytes;   // one persistent scratch block per side stream
    unsigned char* upHost; size_t upSegBytes, upOff; unsigned upSeg;
    std::vector<cudaEvent_t> upFence[4]; bool upFenced[4];
    void* comm;                                  // RdComm* (rd_comm.cu): the NCCL communicator of the multi-GPU frame gather
    volatile unsigned* unitSlots; unsigned* unitSlotsDev; cudaEvent_t unitEvent;   // rd_trace_rays_units: unit numbers land in host-mapped memory
    long optWfTimeline; std::vector<cudaEvent_t> tlEvents; std::vector<int> tlCodes; size_t tlUsed;   // rd_wf_timeline
    uint64_t launches;
    int smCount;
};

void rd_set_error(const char* fmt, ...);
// rd_core.cu, streamed read-back: start the copy of every band the running frame has completed (waits for all of them);
// finish = drain + wait for the copies, after which the image buffer is free again
int rd_stream_bands_drain(rd_platform* p);
int rd_stream_bands_finish(rd_platform* p);
void rd_stream_bands_mark_launch();   // developer knob RD_BAND_TRACE: host time of the frame's launch
#define RD_CUDA(call)                                                                       \
    do {                                                                                    \
        cudaError_t _e = (call);                                                            \
        if (_e != cudaSuccess) {                                                            \
            rd_set_error("%s failed: %s (%s:%d)", #call, cudaGetErrorString(_e), __FILE__, __LINE__); \
            return RD_E_CUDA;                                                               \
        }                                                                                   \
    } while (0)
#define RD_CU(call)                                                                         \
    do {                                                                                    \
        CUresult _e = (call);                                                               \
        if (_e != CUDA_SUCCESS) {                                                           \
            const char* _s = 0; cuGetErrorString(_e, &_s);                                  \
            rd_set_error("%s failed: %s (%s:%d)", #call, _s ? _s : "?", __FILE__, __LINE__); \
            return RD_E_CUDA;                                                               \
        }                                                                                   \
    } while (0)

static inline RdHandleObj* rd_obj(rd_handle h)
{
    RdHandleObj* o = (RdHandleObj*)h;
    return (o && o->magic == RD_HANDLE_MAGIC) ? o : 0;
}

// rd_build.cu: generic device build over primitive boxes (dlo/dhi: n float4 each, device)
struct RdBuildResult { RdNode8* nodes; uint32_t* leafPrim; uint32_t nodeCount, levels; float lo[3], hi[3]; };
int rd_build_wide_bvh(rd_platform* p, const RdF4* dlo, const RdF4* dhi, uint32_t n, RdBuildResult* out);

// rd_wavefront.cu
struct RdWfState;
static inline RdWfBuffers& rd_wf_buffers(rd_platform* p) { return p->wfLane ? p->wfX[p->wfLane - 1] : p->wf; }
static inline cudaStream_t rd_wf_stream(rd_platform* p) { return p->wfLane ? p->wfStreamX[p->wfLane - 1] : p->stream; }
int rd_wf_reserve(rd_platform* p, unsigned count, unsigned payloadStride);
void rd_wf_release(rd_platform* p);
int rd_wf_upload_state(rd_platform* p, const RdWfState* host);
int rd_wf_advance(rd_platform* p, unsigned op, unsigned a, unsigned b = 0u);
int rd_wf_trace(rd_platform* p, const void* scene, int dynamicFetch);
int rd_wf_sort(rd_platform* p, unsigned count);
int rd_wf_reorder_rays(rd_platform* p, const void* scene, unsigned count);

#endif
