/* radiance_b200.h -- C ABI of libradiance_b200.so, the B200 (sm_100a) backend that replaces the
 * Radiance host library + OpenCL device runtime for the TraceRays hot path.
 *
 * Every entry point names the reference interface it replaces (paths relative to the reference
 * repository, zekailin00/Radiance-Ray-Tracing). Plain pointers and sizes only; no C++/torch types.
 * All functions return RD_OK (0) or a negative RD_E_* code; rd_last_error() holds the message of the
 * last failure on the calling thread. The reference aborts on every error (radiance/src/clcontext.h:27-47);
 * the C++ shim include/radiance.h converts non-zero returns into that behaviour.
 * There is no CPU execution path: every call needs a CUDA device.
 */
#ifndef RADIANCE_B200_H
#define RADIANCE_B200_H
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define RD_OK 0
#define RD_E_CUDA (-1)      /* CUDA runtime / driver error */
#define RD_E_ARG (-2)       /* invalid argument */
#define RD_E_COMPILE (-3)   /* NVRTC compile / link error, log in rd_last_error() */
#define RD_E_STATE (-4)     /* call order (no pipeline / descriptors bound, ...) */
#define RD_E_IO (-5)        /* file error */
#define RD_E_OVERFLOW (-6)  /* traversal stack overflow was detected on the device */

typedef struct rd_platform rd_platform;   /* RD::Platform            radiance/include/radiance.h:146-174 */
typedef struct rd_blas rd_blas;           /* RD::BottomAccelStruct   radiance/include/radiance.h:38-45  */
typedef struct rd_module rd_module;       /* RD::ShaderModule        radiance/include/radiance.h:60     */
typedef void* rd_handle;                  /* RD::Handle (Buffer / Image / ImageArray / Sampler / TopAccelStruct),
                                             radiance/include/radiance.h:12-19; one 8-byte kernel parameter each */

/* RD::Instance, radiance/include/radiance.h:48-55 (row-major object->world) */
typedef struct rd_instance_desc {
    float transform[16];
    uint32_t SBTOffset;
    uint32_t customInstanceID;
    rd_blas* bottomAccelStruct;
} rd_instance_desc;

const char* rd_last_error(void);

/* RD::Platform::GetPlatform() + CLContext::GetCLContext(), radiance.h:148-158, clcontext.cpp:12-40.
 * device < 0 selects the current CUDA device. */
int rd_platform_create(int device, rd_platform** out);
int rd_platform_destroy(rd_platform* p);
/* stream used for all work of this platform (cudaStream_t as void*), for external event timing */
void* rd_platform_stream(rd_platform* p);
int rd_platform_sync(rd_platform* p);
/* Device status flags: out[0] = traversal-stack overflow, out[1] = traceRay nested deeper than option max_trace_depth.
 * Every call that synchronises the platform stream (rd_trace_rays, rd_platform_sync, rd_buffer_read, rd_steal_owned,
 * rd_trace_counters, rd_ray_count) reports a raised flag as RD_E_OVERFLOW / RD_E_STATE and clears it, so work issued
 * through the asynchronous entry points cannot return incomplete results silently; this call reads them without
 * synchronising (clear != 0 resets them). The reference prints and returns a miss (radiance/shader/radiance.cl:66-74). */
int rd_status(rd_platform* p, uint32_t out[2], int clear);

/* RD::CreateBuffer, radiance/src/radiance.cpp:133-140 */
int rd_buffer_create(rd_platform* p, size_t size, rd_handle* out);
/* no reference counterpart: a buffer handle over device memory the caller owns (zero-copy interop with
 * other CUDA libraries, e.g. the NCCL framebuffer gather); rd_handle_free does not release the memory */
int rd_buffer_wrap(rd_platform* p, void* device_ptr, size_t size, rd_handle* out);
/* RD::CreateImage (a plain W*H*4-byte buffer), radiance/src/radiance.cpp:86-93 */
int rd_image_create(rd_platform* p, unsigned width, unsigned height, rd_handle* out);
/* RD::CreateImageArray (layered RGBA8), radiance/src/radiance.cpp:96-118 */
int rd_image_array_create(rd_platform* p, unsigned width, unsigned height, unsigned layers, rd_handle* out);
/* RD::CreateSampler, radiance/src/radiance.cpp:120-131; modes are the CL_ADDRESS_* / CL_FILTER_* values */
int rd_sampler_create(rd_platform* p, uint32_t addressingMode, uint32_t filterMode, rd_handle* out);
/* RD::WriteBuffer / RD::ReadBuffer (blocking), radiance/src/radiance.cpp:186-199 */
int rd_buffer_write(rd_platform* p, rd_handle h, size_t size, const void* data, size_t offset);
int rd_buffer_read(rd_platform* p, rd_handle h, size_t size, void* data, size_t offset);
/* Streamed read-back for the TraceRays + ReadBuffer pair of the reference's render loop (samples/sample1.cpp:466-490,
 * radiance/src/radiance.cpp:186-199, 242-267). Registers `host` (page-locked, at least width*height*itemBytes bytes) as
 * the mirror of buffer `h`, which must hold itemBytes bytes per work-item in launch order (a row-major image). From then
 * on a full-frame inline rd_trace_rays* of a module created under option "stream_image" renders the frame in `bands`
 * (1..64) equal runs of tile rows and copies every band to `host` as soon as its blocks have finished (in whatever order
 * they finish), overlapping PCIe with the rest of the frame; the following rd_buffer_read(p, h, size, host, 0) finds
 * the data in place and only waits for
 * the last band. Any other use of the buffer behaves as before. host == NULL ends the registration. */
int rd_buffer_stream_to_host(rd_platform* p, rd_handle h, void* host, unsigned itemBytes, unsigned bands);
/* RD::WriteImage / RD::ReadImage (one layer), radiance/src/radiance.cpp:202-224 */
int rd_image_write(rd_platform* p, rd_handle h, unsigned width, unsigned height, size_t layer, const void* data);
int rd_image_read(rd_platform* p, rd_handle h, unsigned width, unsigned height, size_t layer, void* data);
/* no reference counterpart (the reference never frees): release a buffer/image/sampler/TLAS handle */
int rd_handle_free(rd_platform* p, rd_handle h);
/* raw device address behind a buffer handle (for tests and zero-copy interop) */
void* rd_handle_device_ptr(rd_handle h);

/* RD::BuildAccelStruct(Platform*, Mesh&), radiance/src/radiance.cpp:20-52 (+ bvh.cpp:288-341, 502-522).
 * verts: nv x 3 floats (RD::Vec3), idx: nt x 3 uint32 (RD::Triangle). Host pointers. */
int rd_blas_build(rd_platform* p, const float* verts, unsigned nv, const uint32_t* idx, unsigned nt, rd_blas** out);
/* same, inputs already resident in HBM (device pointers): the GPU build without the upload */
int rd_blas_build_device(rd_platform* p, const float* dverts, unsigned nv, const uint32_t* didx, unsigned nt, rd_blas** out);
int rd_blas_free(rd_platform* p, rd_blas* b);
/* build statistics: [0] wide nodes, [1] triangle records, [2] collapse levels, [3] bytes in HBM */
int rd_blas_stats(rd_blas* b, uint64_t out[4]);

/* RD::BuildAccelStruct(Platform*, std::vector<Instance>&), radiance/src/radiance.cpp:54-84 */
int rd_tlas_build(rd_platform* p, const rd_instance_desc* instances, unsigned n, rd_handle* out);

/* RD::FileToTopAccelStruct, radiance/src/radiance.cpp:450-479: reads a blob in the reference's byte layout
 * (radiance.cpp:318-425, core.h:34-101), keeps its geometry and instance table and rebuilds the native
 * structure on the GPU. A short or malformed file is RD_E_IO (the reference throws, radiance.cpp:458-459).
 * _memory: same from a host buffer; blasOut/blasCount (optional) receive the bottom-level structures that were
 * built (array allocated with malloc, release with rd_free_host; the structures with rd_blas_free). */
int rd_tlas_load(rd_platform* p, const char* path, rd_handle* out);
int rd_tlas_load_memory(rd_platform* p, const void* data, size_t size, rd_handle* out, rd_blas*** blasOut, unsigned* blasCount);
/* RD::TopAccelStructToFile, radiance/src/radiance.cpp:428-448: writes the scene in the reference's blob format
 * (binary BVH in DeviceBVHNode layout, leaves below 8 primitives), walkable by the reference's traceRay.
 * _memory: into a malloc'ed buffer (release with rd_free_host). Fails with RD_E_ARG above 4 GiB (32-bit offsets). */
int rd_tlas_save(rd_platform* p, rd_handle tlas, const char* path);
int rd_tlas_save_memory(rd_platform* p, rd_handle tlas, void** dataOut, size_t* sizeOut);
void rd_free_host(void* ptr);

/* RD::CreateShaderModule, radiance/src/radiance.cpp:152-179: compiles an OpenCL-C user program
 * (samples/shader.cl conventions) with NVRTC for sm_100a; kernel "raygen" is extracted. `name` is
 * accepted and ignored like in the reference. */
int rd_shader_module_create(rd_platform* p, const char* code, unsigned size, const char* name, rd_module** out);
int rd_module_free(rd_platform* p, rd_module* m);
/* compile-only variant of the above (no device needed): RD_OK + cubin size, the ptxas statistics are
 * left in rd_last_error(); on failure RD_E_COMPILE with the compiler log there */
int rd_shader_compile_check(const char* code, unsigned size, const char* sbt_json, int fmad, size_t* cubinSize);
/* registers per thread / local bytes of the compiled raygen kernel */
int rd_module_info(rd_module* m, int out[2]);
/* samples/sbt.json: rows {raygen, closestHit, anyHit, miss}; row index = dispatcher case label.
 * When set before rd_shader_module_create and the program does not define callHit/callMiss/callAnyHit
 * itself, the dispatchers are generated from the table (tools/genSBT.py semantics). */
int rd_sbt_load_json(rd_platform* p, const char* json_text);
/* RD::BindPipeline, radiance/src/radiance.cpp:226-229 (module 0 of the pipeline is the one launched,
 * radiance.cpp:254) */
int rd_pipeline_bind(rd_platform* p, rd_module* const* modules, unsigned n);
/* RD::BindDescriptorSet, radiance/src/radiance.cpp:231-240: handle i becomes kernel argument i */
int rd_descriptors_bind(rd_platform* p, const rd_handle* handles, unsigned n);
/* RD::TraceRays, radiance/src/radiance.cpp:242-267: launches `raygen` over width*height work-items
 * and blocks. The three group indices are accepted and ignored like in the reference. */
int rd_trace_rays(rd_platform* p, unsigned raygenGroupIndex, unsigned missGroupIndex, unsigned hitGroupIndex,
                  unsigned width, unsigned height);
/* same without the final host synchronisation (for device-side event timing) */
int rd_trace_rays_async(rd_platform* p, unsigned width, unsigned height);
/* work-items [first, first+count) of the width*height NDRange only (count 0 = to the end); with the
 * 8x4 tile mapping a contiguous range is a contiguous run of screen tiles: the multi-GPU partition unit */
int rd_trace_rays_range(rd_platform* p, unsigned width, unsigned height, unsigned long long first,
                        unsigned long long count);

/* Multi-GPU static partition with interleaved bands: the frame's rows of 8x4 tiles are dealt to the ranks in
 * bands of bandTileRows rows, round robin; one launch renders every band of `rank` (the expensive screen
 * regions -- e.g. the horizon of a height-field -- are spread over all GPUs). No reference counterpart. */
int rd_trace_rays_interleaved(rd_platform* p, unsigned width, unsigned height, unsigned rank, unsigned world,
                              unsigned bandTileRows);

/* Multi-GPU dynamic tile stealing (one process per GPU, scene replicated): persistent raygen blocks take chunks
 * of `chunkItems` consecutive work-items (runs of 8x4 screen tiles) from one counter shared by all GPUs of the
 * job -- it lives on the exporting rank's GPU, the other ranks map it through CUDA IPC and increment it with
 * system-scope atomics over NVLink. Modules must be created under option "stealing" = 1. No reference
 * counterpart (the reference is single-device, radiance/src/clcontext.cpp:12-40).
 *   rank 0: rd_steal_counter_export -> 64-byte handle, sent to the other ranks (any transport)
 *   others: rd_steal_counter_import
 *   every rank, every frame: rd_trace_rays_steal(width, height, frame, chunkItems) with the same arguments,
 *   frame = 0, 1, 2, ... (it selects the counter slot; a slot is reusable after 512 further frames);
 *   each work-item is rendered by exactly one rank, pixels of the other ranks are left untouched.
 * rd_steal_owned: the chunk ids this rank rendered in its last frame. */
int rd_steal_counter_export(rd_platform* p, unsigned char handle[64]);
int rd_steal_counter_import(rd_platform* p, const unsigned char handle[64]);
/* ranks that are host threads of ONE process: share the owner platform's counters through peer access instead of IPC */
int rd_steal_counter_adopt(rd_platform* p, rd_platform* owner);
int rd_trace_rays_steal(rd_platform* p, unsigned width, unsigned height, unsigned frame, unsigned chunkItems);
int rd_steal_owned(rd_platform* p, unsigned* ids, unsigned capacity, unsigned* count);

/* The multi-GPU frame collective (csrc/rd_comm.cu). The reference renders one blocking frame on one device and reads it
 * back (radiance/src/radiance.cpp:242-267, :186-199; one cl_context on one device, clcontext.cpp:12-40); here the frame is
 * partitioned over the GPUs of one box (one process or host thread per GPU, scene replicated) and the finished framebuffer
 * is gathered on rank 0 over NVLink with NCCL -- the only exchange of a frame. A host in the reference's idiom
 * (samples/sample1.cpp:466-490) calls, per frame and on every rank:
 *     rd_trace_rays_interleaved(p, w, h, rank, world, bandTileRows);   rd_frame_gather(p, image, w, h, 4, bandTileRows);
 * and rank 0 then reads the frame with rd_buffer_read. NCCL is loaded on first use (libnccl.so.2).
 *   rd_comm_unique_id   rank 0 creates the 128-byte NCCL id and hands it to the other ranks (any transport)
 *   rd_comm_init        every rank: joins the communicator (collective call); rd_comm_adopt uses an ncclComm_t the host owns
 *   rd_frame_gather     static interleaved partition: each peer packs its bands, ONE send per peer, rank 0 unpacks
 *   rd_frame_gather_ranges  contiguous work-item ranges per rank (rd_trace_rays_range), sent in place
 *   rd_trace_rays_units + rd_frame_gather_units   DYNAMIC partition: the frame is cut into `units` (<= 64) equal runs of
 *                       8x4-tile rows, every rank takes the next unit from the job-wide counter of rd_steal_counter_export /
 *                       _import (system-scope atomic over NVLink) and renders it in whatever mode is on -- wavefront
 *                       included; `frame` = 0, 1, 2, ... selects the counter slot as in rd_trace_rays_steal; ownedMask
 *                       returns the units this rank rendered and is what rd_frame_gather_units exchanges. */
int rd_comm_unique_id(unsigned char id[128]);
int rd_comm_init(rd_platform* p, int rank, int world, const unsigned char id[128]);
int rd_comm_adopt(rd_platform* p, void* ncclComm, int rank, int world);
int rd_comm_destroy(rd_platform* p);
int rd_frame_gather(rd_platform* p, rd_handle image, unsigned width, unsigned height, unsigned itemBytes, unsigned bandTileRows);
int rd_frame_gather_ranges(rd_platform* p, rd_handle image, unsigned itemBytes, const unsigned long long* first,
                           const unsigned long long* count);
int rd_trace_rays_units(rd_platform* p, unsigned width, unsigned height, unsigned units, unsigned frame,
                        unsigned long long* ownedMask);
int rd_frame_gather_units(rd_platform* p, rd_handle image, unsigned width, unsigned height, unsigned itemBytes,
                          unsigned units, unsigned long long ownedMask);
/* test hook: the pack (peer side) and unpack (rank 0 side) kernels of rd_frame_gather on one GPU */
int rd_debug_band_roundtrip(rd_platform* p, rd_handle src, rd_handle dst, unsigned width, unsigned height, unsigned itemBytes,
                            unsigned bandTileRows, unsigned rank, unsigned world);
/* memcheck stand-in (compute-sanitizer is closed on the GPU pool): a library built with -DRD_BOUNDS_CHECK=1 checks every queue /
 * order / record index against the size of what it indexes; *count is non-zero if any check failed since the last call. */
int rd_debug_bounds_violations(rd_platform* p, unsigned* count, int* checkedBuild);
/* Stage timeline of the last wavefront-mode call (option "wf_timeline" = 1): ms[i] = device time of stage i, codes[i] =
 * 1 generate, 2 trace, 3 sort, 4 shade, 5 accumulate, 6 resolve, 7 ray re-sort. Synchronises the platform stream. */
int rd_wf_timeline(rd_platform* p, float* ms, int* codes, unsigned capacity, unsigned* count);

/* Explicit ray list through the traversal kernel (wavefront building block; used for hit-ID parity
 * against the reference's traceRay, radiance/shader/radiance.cl:254-275).
 *   rayA[i] = {o.xyz, tmin}, rayB[i] = {d.xyz, tmax}; hitA[i] = {t, b1, b2, primID bits}, hitB[i] = instanceID.
 * All four are buffer handles. flags: RD_TRACE_* below. Asynchronous on the platform stream. */
#define RD_TRACE_TERMINATE_ON_FIRST_HIT 1u
#define RD_TRACE_COUNTERS 2u   /* also count node / triangle fetches (see rd_trace_counters) */
#define RD_TRACE_STATIC_BATCHES 4u   /* 32-ray batches per warp (the default; flag kept for compatibility) */
#define RD_TRACE_DYNAMIC_FETCH 8u    /* ablation: persistent warps that re-fetch rays for idle lanes (slower) */
int rd_trace_ray_list(rd_platform* p, rd_handle tlas, rd_handle rayA, rd_handle rayB, rd_handle hitA,
                      rd_handle hitB, size_t n, unsigned flags);
/* totals of the last RD_TRACE_COUNTERS launch: [0] wide nodes fetched, [1] triangle records fetched,
 * [2] rays */
int rd_trace_counters(rd_platform* p, uint64_t out[3]);
/* pinhole / thin-lens primary rays of samples/shader.cl:111-173 into rayA/rayB (width*height rays). Ray slot
 * i belongs to work-item i; with option "tile" on and width % 8 == 0, height % 4 == 0 the slots enumerate
 * 8x4 pixel tiles row by row (slot -> pixel: tile = i / 32, lane = i % 32, x = 8*(tile % (W/8)) + lane % 8,
 * y = 4*(tile / (W/8)) + lane / 8), otherwise slot i is pixel i in row-major order */
int rd_generate_primary_rays(rd_platform* p, rd_handle camera, unsigned frameID, unsigned totalSamples,
                             unsigned width, unsigned height, rd_handle rayA, rd_handle rayB);

/* runtime options: "fmad" (0/1, NVRTC contraction for user code), "prec_div" (0/1, default 0: user-program division at the
 * 2.5 ulp OpenCL C asks for; 1 = correctly rounded), "async_build" (0/1, default 1: small bottom-level builds are deferred), "tile" (0/1 8x4 pixel tiles per warp),
 * "block" (threads per block of raygen launches), "count_rays" (0/1: modules created afterwards count
 * every traceRay invocation, the numerator of the Mrays/s metric; read with rd_ray_count),
 * "max_trace_depth" (1..8, default 2: how deep traceRay may nest -- raygen -> hit shader -> shadow ray is 2;
 * the user program is compiled once per level so every level inlines, PipelineCreateInfo::maxRayRecursionDepth
 * of radiance/include/radiance.h:74 is the reference's unused counterpart),
 * "wavefront" (0/1: rd_trace_rays runs the sample / bounce loop of samples/shader.cl:196-281 as stage launches
 * over ray queues in HBM -- generate, trace, sort by hit group, shade, resolve -- instead of one raygen
 * launch; needs the wavefrontScene / wavefrontPrimary hooks in the program),
 * "stealing" (0/1: modules created afterwards also carry the persistent chunk-stealing raygen),
 * "wf_dynamic" (0 never / 1 bounce stages (default) / 2 always: dynamic ray fetch in the wavefront trace stage),
 * "wf_ray_sort" (0/1, default 0: re-sort bounce rays by direction octant and origin cell before tracing them;
 * measured slower on C3 -- the pass costs more than the trace stage gains),
 * "wf_sort" (0/1, default 1: reorder each bounce's rays by hit-group row and material key before shading),
 * "wf_paths" (default 2^26: paths in flight per wavefront frame, shared by the lanes -- RTProp.batchSize samples of every
 * work-item are generated, traced and shaded together, as many as fit under this cap at ~200 B of HBM per path (13 GB at the
 * default); 1 = one sample per pass. 4K 64 spp: 2^25 426 ms, 2^26 419 ms, 2^27 419 ms. Images do not depend on it: finished samples are folded into imageScratch in frame order),
 * "wf_lanes" (1..4, default 2: the sample passes of a wavefront frame go round robin over that many queue sets, each on its own
 * stream, so the tail of one pass's stage -- its longest rays -- overlaps the other passes' stages; same image bit for bit, same
 * HBM; 3 and 4 lanes measured no faster than 2),
 * "stream_image" (default 0; modules created while it is 1 also carry the band-reporting raygen that
 * rd_buffer_stream_to_host needs),
 * "trim_pool" (any value: releases what the library keeps between calls -- the stream-ordered build scratch held in the
 * device's memory pool, the per-stream scratch blocks and the empty 64 MB chunks of the small-mesh builds, and the wavefront
 * queues, up to ~200 B x wf_paths -- they are allocated again on demand) */
int rd_set_option(rd_platform* p, const char* key, long value);
/* traceRay invocations counted by modules compiled with option count_rays (primary + bounce + shadow
 * rays, radiance/shader/radiance.cl:254); synchronises the platform stream; reset != 0 zeroes the counter */
int rd_ray_count(rd_platform* p, uint64_t* out, int reset);
/* diagnostics: the build's own radix sort (stable, LSD, 8 bits per pass over the low `bits` key bits; csrc/rd_sort.h
 * replaces a library sort in the reference-free GPU build that stands in for radiance/src/bvh.cpp:46-341) and its
 * exclusive scan, run on host arrays in place. The parity tests compare them with numpy. */
int rd_debug_sort_pairs(rd_platform* p, uint64_t* keys, uint32_t* vals, unsigned n, int bits);
int rd_debug_exclusive_scan(rd_platform* p, uint32_t* data, unsigned n, uint32_t* total);
/* No counterpart in the reference. The same pair sort (stable, by the low
 * `bits` key bits, 8 per pass) for callers that cannot allocate (CUDA graph capture, serving loops, harnesses that forbid
 * cudaMalloc): inputs and outputs are DEVICE pointers the caller owns (n keys / n values each, outputs distinct from the
 * inputs, inputs are left unchanged); the ping-pong buffers and digit tables are a file-scope static device array of the
 * library (6.4 MB, held by every GPU that loads the library until exit), so the call performs no cudaMalloc / cudaFree and
 * no host synchronisation -- it is asynchronous on the platform stream (rd_platform_sync to wait). n is limited to
 * RD_SORT_NOALLOC_MAX_PAIRS; more is RD_E_ARG with the limit in rd_last_error(). All calls on one GPU share the array:
 * calls that overlap on the GPU (two platforms / streams of one device) race; calls on one platform are stream-ordered.
 * The first call on a device resolves the array's address (cudaGetSymbolAddress); make it outside a capture. */
#define RD_SORT_NOALLOC_MAX_PAIRS (1u << 18)
int rd_sort_pairs_noalloc(rd_platform* p, const uint64_t* dkeysIn, const uint32_t* dvalsIn, uint64_t* dkeysOut,
                          uint32_t* dvalsOut, unsigned n, int bits);
/* number of kernels this library launched since platform creation */
uint64_t rd_launch_count(rd_platform* p);

#ifdef __cplusplus
}
#endif
#endif
