// rd_core.cu -- platform, resource handles and blocking transfers of the C ABI.
// Replaces radiance/src/clcontext.cpp:12-40 (context) and radiance/src/radiance.cpp:86-140,186-224
// (CreateBuffer / CreateImage / CreateImageArray / CreateSampler / Read* / Write*).
#include <time.h>
#include <stdio.h>
#include <stdarg.h>
#include <stdio.h>
#include <string.h>
#include "rd_internal.h"

static thread_local char g_err[16384] = "";

void rd_set_error(const char* fmt, ...)
{
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

extern "C" const char* rd_last_error(void) { return g_err; }

extern "C" int rd_platform_create(int device, rd_platform** out)
{
    if (!out) { rd_set_error("rd_platform_create: out is null"); return RD_E_ARG; }
    int count = 0;
    cudaError_t e = cudaGetDeviceCount(&count);
    if (e != cudaSuccess || count == 0) {
        rd_set_error("rd_platform_create: no CUDA device (%s); this library has no CPU path",
                     e != cudaSuccess ? cudaGetErrorString(e) : "device count 0");
        return RD_E_CUDA;
    }
    if (device < 0) RD_CUDA(cudaGetDevice(&device));
    RD_CUDA(cudaSetDevice(device));
    RD_CUDA(cudaFree(0));
    rd_platform* p = new rd_platform();
    p->device = device;
    p->boundModule = 0;
    p->optFmad = 1; p->optTile = 1; p->optBlock = 128; p->optCountRays = 0; p->optMaxDepth = 2; p->optWavefront = 0; p->optWfSort = 1; p->optSteal = 0; p->optWfDynamic = 1; p->optWfRaySort = 0; p->optWfPaths = 1l << 26; p->optWfLanes = 2;
    p->ilWorld = p->ilRank = p->ilBandRows = 0;
    p->optStreamImage = 0;
    p->mirrorObj = 0; p->mirrorHost = 0; p->mirrorItemBytes = 0; p->mirrorBands = 0; p->mirrorValidBytes = 0; p->mirrorFresh = 0; p->copyPending = 0;
    p->copyStream = 0; p->bandDone = 0; p->bandDoneCap = 0;
    p->bandLayoutW = p->bandLayoutH = p->bandLayoutN = p->bandLayoutBlock = p->bandEpoch = 0;
    p->bandFlags = 0; p->bandFlagsDev = 0; p->bandCount = p->bandRows = p->bandTileRows = 0; p->bandCopied = 0; p->bandRowBytes = 0;
    p->stealCounters = 0; p->stealOwned = 0; p->stealImported = 0; p->stealSlots = 0; p->stealOwnedCap = 0;
    memset(&p->wf, 0, sizeof(p->wf));
    p->buildNext = 0; p->summaryUsed = 0; p->optAsyncBuild = 1; p->optPrecDiv = 0; p->optL2Persist = 0;   // off by default: +3 % on the 10 M-triangle primary frame at 96 MB, but the wavefront queues lose the L2 they
    // stream through (C3 42.6 -> 58.0 ms, C5 488 -> 709 ms; profiles/README_r2.md)
    if (const char* e = getenv("RD_L2_PERSIST")) p->optL2Persist = atol(e);
    if (const char* e = getenv("RD_WF_LANES")) { p->optWfLanes = atol(e); if (p->optWfLanes < 1) p->optWfLanes = 1; if (p->optWfLanes > RD_WF_MAX_LANES) p->optWfLanes = RD_WF_MAX_LANES; }   // developer knob for A/B runs of bench.py
    p->comm = 0; p->unitSlots = 0; p->unitSlotsDev = 0; p->unitEvent = 0; p->optWfTimeline = 0; p->tlUsed = 0;
    p->launches = 0;
    RD_CUDA(cudaStreamCreateWithFlags(&p->stream, cudaStreamNonBlocking));
    {   // build scratch is stream-ordered (RdTmp in rd_build.cu): keep freed blocks in the pool instead of returning them to the OS
        cudaMemPool_t pool; unsigned long long keep = ~0ull;
        if (cudaDeviceGetDefaultMemPool(&pool, device) == cudaSuccess) cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &keep);
    }
    {
        void* hs = 0; void* ds = 0;
        RD_CUDA(cudaHostAlloc(&hs, 64, cudaHostAllocMapped));
        memset(hs, 0, 64);
        RD_CUDA(cudaHostGetDevicePointer(&ds, hs, 0));
        p->hstatus = (volatile uint32_t*)hs; p->dstatus = (uint32_t*)ds;
    }
    RD_CUDA(cudaMalloc(&p->dRayCounter, 64));
    RD_CUDA(cudaMemset(p->dRayCounter, 0, 64));
    RD_CUDA(cudaMalloc(&p->dcounters, 64));
    RD_CUDA(cudaMemset(p->dcounters, 0, 64));
    RD_CUDA(cudaDeviceGetAttribute(&p->smCount, cudaDevAttrMultiProcessorCount, device));
    // recursive traceRay -> callHit -> traceRay chains of user programs need a real call stack
    cudaDeviceSetLimit(cudaLimitStackSize, 8192);
    *out = p;
    return RD_OK;
}

extern "C" int rd_platform_destroy(rd_platform* p)
{
    if (!p) return RD_OK;
    cudaSetDevice(p->device);
    cudaStreamSynchronize(p->stream);
    rd_comm_destroy(p);
    for (cudaStream_t bs : p->buildStreams) { cudaStreamSynchronize(bs); cudaStreamDestroy(bs); }
    for (uint32_t* hsum : p->summaryHost) cudaFreeHost(hsum);
    for (RdArenaChunk* c : p->arenaChunks) { cudaFree(c->base); cudaEventDestroy(c->idle); delete c; }
    for (void* q : p->buildScratch) if (q) cudaFree(q);
    for (int k = 0; k < 4; ++k) for (cudaEvent_t e : p->upFence[k]) cudaEventDestroy(e);
    if (p->upHost) cudaFreeHost(p->upHost);
    for (cudaEvent_t e : p->tlEvents) cudaEventDestroy(e);
    if (p->unitSlots) { cudaFreeHost((void*)p->unitSlots); cudaEventDestroy(p->unitEvent); }
    rd_wf_release(p);
    if (p->wfStreamX[0]) { for (int k = 0; k < RD_WF_MAX_LANES - 1; ++k) cudaStreamDestroy(p->wfStreamX[k]); for (int k = 0; k <= RD_WF_MAX_LANES; ++k) cudaEventDestroy(p->wfEvents[k]); }
    { cudaMemPool_t pool; if (cudaDeviceGetDefaultMemPool(&pool, p->device) == cudaSuccess) cudaMemPoolTrimTo(pool, 0); }
    cudaFreeHost((void*)p->hstatus);
    cudaFree(p->dRayCounter);
    cudaFree(p->dcounters);
    if (p->copyStream) { cudaStreamSynchronize(p->copyStream); cudaStreamDestroy(p->copyStream); }
    if (p->bandDone) cudaFree(p->bandDone);
    if (p->bandFlags) cudaFreeHost((void*)p->bandFlags);
    cudaStreamDestroy(p->stream);
    delete p;
    return RD_OK;
}

extern "C" void* rd_platform_stream(rd_platform* p) { return p ? (void*)p->stream : 0; }

// Device status flags (traversal-stack overflow, traceRay nested deeper than option max_trace_depth). They live in
// host-mapped memory, so looking at them costs nothing; every entry point that has just synchronised the platform
// stream calls this, reports the condition once and clears it -- work launched through the asynchronous entry points
// (rd_trace_rays_async / _range / _interleaved / _steal, rd_trace_ray_list) cannot come back incomplete with RD_OK.
int rd_check_status(rd_platform* p, const char* where)
{
    const uint32_t s0 = p->hstatus[0], s1 = p->hstatus[1];
    if (!(s0 | s1)) return RD_OK;
    p->hstatus[0] = 0u; p->hstatus[1] = 0u;
    if (s0) { rd_set_error("%s: traversal stack overflow on the device (results incomplete)", where); return RD_E_OVERFLOW; }
    rd_set_error("%s: traceRay nested deeper than option max_trace_depth (%ld)", where, p->optMaxDepth);
    return RD_E_STATE;
}

extern "C" int rd_status(rd_platform* p, uint32_t out[2], int clear)
{
    if (!p || !out) { rd_set_error("rd_status: null argument"); return RD_E_ARG; }
    out[0] = p->hstatus[0]; out[1] = p->hstatus[1];
    if (clear) { p->hstatus[0] = 0u; p->hstatus[1] = 0u; }
    return RD_OK;
}

extern "C" int rd_platform_sync(rd_platform* p)
{
    if (!p) { rd_set_error("null platform"); return RD_E_ARG; }
    RD_CUDA(cudaStreamSynchronize(p->stream));
    return rd_check_status(p, "rd_platform_sync");
}

static RdHandleObj* new_handle(RdHandleKind kind)
{
    RdHandleObj* o = new RdHandleObj();
    memset(o, 0, sizeof(*o));
    o->magic = RD_HANDLE_MAGIC;
    o->kind = kind;
    return o;
}

extern "C" int rd_buffer_create(rd_platform* p, size_t size, rd_handle* out)
{
    if (!p || !out) { rd_set_error("rd_buffer_create: null argument"); return RD_E_ARG; }
    RD_CUDA(cudaSetDevice(p->device));
    RdHandleObj* o = new_handle(RD_H_BUFFER);
    size_t alloc = size ? size : 16;
    RD_CUDA(cudaMalloc(&o->dptr, alloc));
    RD_CUDA(cudaMemsetAsync(o->dptr, 0, alloc, p->stream));
    o->size = size;
    o->arg = (uint64_t)(uintptr_t)o->dptr;
    *out = o;
    return RD_OK;
}

extern "C" int rd_buffer_wrap(rd_platform* p, void* device_ptr, size_t size, rd_handle* out)
{
    if (!p || !out || !device_ptr) { rd_set_error("rd_buffer_wrap: null argument"); return RD_E_ARG; }
    RdHandleObj* o = new_handle(RD_H_BUFFER);
    o->dptr = device_ptr; o->size = size; o->external = 1;
    o->arg = (uint64_t)(uintptr_t)device_ptr;
    *out = o;
    return RD_OK;
}

extern "C" int rd_image_create(rd_platform* p, unsigned width, unsigned height, rd_handle* out)
{
    return rd_buffer_create(p, (size_t)width * height * 4, out);
}

extern "C" int rd_image_array_create(rd_platform* p, unsigned width, unsigned height, unsigned layers, rd_handle* out)
{
    if (!p || !out) { rd_set_error("rd_image_array_create: null argument"); return RD_E_ARG; }
    RD_CUDA(cudaSetDevice(p->device));
    RdHandleObj* o = new_handle(RD_H_IMAGE_ARRAY);
    o->width = width; o->height = height; o->layers = layers;
    o->size = (size_t)width * height * layers * 4;
    RD_CUDA(cudaMalloc(&o->dptr, o->size ? o->size : 16));
    RD_CUDA(cudaMemsetAsync(o->dptr, 0, o->size ? o->size : 16, p->stream));
    RdImageArrayDesc d;
    d.texels = (const unsigned char*)o->dptr; d.width = width; d.height = height; d.layers = layers; d.pad = 0;
    RD_CUDA(cudaMalloc(&o->ddesc, sizeof(d)));
    RD_CUDA(cudaMemcpyAsync(o->ddesc, &d, sizeof(d), cudaMemcpyHostToDevice, p->stream));
    RD_CUDA(cudaStreamSynchronize(p->stream));
    o->arg = (uint64_t)(uintptr_t)o->ddesc;
    *out = o;
    return RD_OK;
}

extern "C" int rd_sampler_create(rd_platform* p, uint32_t addressingMode, uint32_t filterMode, rd_handle* out)
{
    if (!p || !out) { rd_set_error("rd_sampler_create: null argument"); return RD_E_ARG; }
    RdHandleObj* o = new_handle(RD_H_SAMPLER);
    // token layout of shader/cl_compat.h: [15:0] address mode, [31:16] filter, [32] normalised coords
    // (the reference always creates normalised samplers, radiance/src/radiance.cpp:124)
    o->arg = (uint64_t)(addressingMode & 0xffffu) | ((uint64_t)(filterMode & 0xffffu) << 16) | (1ull << 32);
    *out = o;
    return RD_OK;
}

static int check_range(RdHandleObj* o, size_t size, size_t offset, const char* who)
{
    if (!o || (o->kind != RD_H_BUFFER && o->kind != RD_H_IMAGE_ARRAY)) { rd_set_error("%s: not a buffer handle", who); return RD_E_ARG; }
    if (offset > o->size || size > o->size - offset) {
        rd_set_error("%s: range [%zu, %zu) exceeds buffer of %zu bytes", who, offset, offset + size, o->size);
        return RD_E_ARG;
    }
    return RD_OK;
}

// developer knob RD_BAND_TRACE=1: host-clock microseconds since the launch at which each band's flag was seen, its copy
// was queued, and the last copy had arrived (stderr, one line per frame)
static double rd_now_us()
{
    timespec ts; clock_gettime(CLOCK_MONOTONIC, &ts);
    return ts.tv_sec * 1e6 + ts.tv_nsec * 1e-3;
}
static double rd_band_t0 = 0.0;
static double rd_band_seen[64];
void rd_stream_bands_mark_launch() { static const bool trace = getenv("RD_BAND_TRACE") != 0; if (trace) rd_band_t0 = rd_now_us(); }

int rd_stream_bands_drain(rd_platform* p)
{
    if (!p->copyPending) return RD_OK;
    static const bool trace = getenv("RD_BAND_TRACE") != 0;
    const uint64_t all = p->bandCount >= 64u ? ~0ull : ((1ull << p->bandCount) - 1ull);
    for (unsigned spins = 1; p->bandCopied != all; ++spins) {
        for (unsigned b = 0; b < p->bandCount; ++b) {
            if ((p->bandCopied >> b) & 1ull) continue;
            if (p->bandFlags[b] != p->bandEpoch) continue;
            const size_t r0 = (size_t)b * p->bandRows, r1 = r0 + p->bandRows < p->bandTileRows ? r0 + p->bandRows : p->bandTileRows;
            if (trace) rd_band_seen[b] = rd_now_us() - rd_band_t0;
            RD_CUDA(cudaMemcpyAsync((char*)p->mirrorHost + r0 * p->bandRowBytes, (const char*)p->mirrorObj->dptr + r0 * p->bandRowBytes,
                                    (r1 - r0) * p->bandRowBytes, cudaMemcpyDeviceToHost, p->copyStream));
            p->bandCopied |= 1ull << b;
        }
        if ((spins & 0xfffu) == 0u && p->bandCopied != all) {
            // the frame's launch is the last thing on the stream: once it is done every flag must be up
            const cudaError_t q = cudaStreamQuery(p->stream);
            bool up = true;
            for (unsigned b = 0; b < p->bandCount; ++b) up = up && p->bandFlags[b] == p->bandEpoch;
            if (q == cudaSuccess && !up) {
                rd_set_error("streamed read-back: the frame ended without completing all of its bands");
                p->copyPending = 0; p->mirrorFresh = 0; p->bandLayoutW = 0;   // counters start over with the next frame
                return RD_E_STATE;
            }
            if (q != cudaSuccess && q != cudaErrorNotReady) {
                rd_set_error("streamed read-back: %s", cudaGetErrorString(q));
                p->copyPending = 0; p->mirrorFresh = 0; p->bandLayoutW = 0;
                return RD_E_CUDA;
            }
        }
    }
    return RD_OK;
}

int rd_stream_bands_finish(rd_platform* p)
{
    if (!p->copyPending) return RD_OK;
    int rc = rd_stream_bands_drain(p);
    if (rc) return rc;
    RD_CUDA(cudaStreamSynchronize(p->copyStream));
    p->copyPending = 0;
    if (getenv("RD_BAND_TRACE")) {
        fprintf(stderr, "bands:");
        for (unsigned b = 0; b < p->bandCount; ++b) fprintf(stderr, " %u:%.0f", b, rd_band_seen[b]);
        fprintf(stderr, " arrived %.0f us\n", rd_now_us() - rd_band_t0);
    }
    return RD_OK;
}

extern "C" int rd_buffer_write(rd_platform* p, rd_handle h, size_t size, const void* data, size_t offset)
{
    RdHandleObj* o = rd_obj(h);
    if (!p || (!data && size)) { rd_set_error("rd_buffer_write: null argument"); return RD_E_ARG; }
    int rc = check_range(o, size, offset, "rd_buffer_write");
    if (rc) return rc;
    if (!size) return RD_OK;
    RD_CUDA(cudaSetDevice(p->device));
    if (o == p->mirrorObj) {   // the host mirror no longer matches; a streamed copy still in flight must read the old contents first
        if (p->copyPending) { rc = rd_stream_bands_finish(p); if (rc) return rc; }
        p->mirrorFresh = 0;
    }
    if (size <= 256 && !getenv("RD_SMALL_WRITE_SYNC")) {
        // Per-frame parameter blocks (RTProp, camera: samples/sample1.cpp:466-470). Blocking semantics only ask that `data`
        // may be reused on return: a copy from pageable memory returns once the driver has taken the bytes, so a stack
        // copy makes the call safe without draining the stream; stream order puts it ahead of the next launch.
        unsigned char staged[256];
        memcpy(staged, data, size);
        RD_CUDA(cudaMemcpyAsync((char*)o->dptr + offset, staged, size, cudaMemcpyHostToDevice, p->stream));
        return RD_OK;
    }
    RD_CUDA(cudaMemcpyAsync((char*)o->dptr + offset, data, size, cudaMemcpyHostToDevice, p->stream));
    RD_CUDA(cudaStreamSynchronize(p->stream));
    return RD_OK;
}

extern "C" int rd_buffer_read(rd_platform* p, rd_handle h, size_t size, void* data, size_t offset)
{
    RdHandleObj* o = rd_obj(h);
    if (!p || (!data && size)) { rd_set_error("rd_buffer_read: null argument"); return RD_E_ARG; }
    int rc = check_range(o, size, offset, "rd_buffer_read");
    if (rc) return rc;
    if (!size) return RD_OK;
    RD_CUDA(cudaSetDevice(p->device));
    if (p->mirrorFresh && o == p->mirrorObj && data == p->mirrorHost && offset == 0 && size <= p->mirrorValidBytes) {
        // the last TraceRays is streaming its bands into this very host buffer (rd_buffer_stream_to_host): start the
        // copies of the bands as they complete; behind the last ray only the small last band is left to move
        rc = rd_stream_bands_finish(p);
        p->mirrorFresh = 0;
        return rc ? rc : rd_check_status(p, "rd_buffer_read");
    }
    if (p->copyPending) { rc = rd_stream_bands_finish(p); if (rc) return rc; p->mirrorFresh = 0; }
    RD_CUDA(cudaMemcpyAsync(data, (const char*)o->dptr + offset, size, cudaMemcpyDeviceToHost, p->stream));
    RD_CUDA(cudaStreamSynchronize(p->stream));
    return rd_check_status(p, "rd_buffer_read");
}

// Streamed read-back. The reference's render loop is TraceRays (blocking) followed by ReadBuffer of the whole image
// (samples/sample1.cpp:466-490, radiance/src/radiance.cpp:186-199): the PCIe transfer starts when the last ray is done.
// After this call, full-frame inline launches of a module created under option "stream_image" split the frame into
// up to `bands` runs of tile rows; the raygen blocks count themselves off per band and the block that completes a band
// raises a flag in host-mapped memory. The host thread -- blocked in TraceRays / ReadBuffer anyway -- watches the flags
// and starts each finished band's copy to `host` on a second stream while the later bands are still rendering. The
// rd_buffer_read(handle, size, host, 0) with the same pointer then only waits for the last, small band. The buffer must
// hold itemBytes bytes per work-item in launch order (a row-major image); host should be page-locked. host == NULL ends it.
extern "C" int rd_buffer_stream_to_host(rd_platform* p, rd_handle h, void* host, unsigned itemBytes, unsigned bands)
{
    if (!p) { rd_set_error("rd_buffer_stream_to_host: null platform"); return RD_E_ARG; }
    RD_CUDA(cudaSetDevice(p->device));
    if (p->copyPending) { int frc = rd_stream_bands_finish(p); if (frc) return frc; }
    p->mirrorFresh = 0;
    if (!host) { p->mirrorObj = 0; p->mirrorHost = 0; return RD_OK; }
    RdHandleObj* o = rd_obj(h);
    if (!o || (o->kind != RD_H_BUFFER && o->kind != RD_H_IMAGE_ARRAY)) { rd_set_error("rd_buffer_stream_to_host: not a buffer handle"); return RD_E_ARG; }
    if (!itemBytes || bands < 1 || bands > 64) { rd_set_error("rd_buffer_stream_to_host: itemBytes must be positive and bands in [1,64]"); return RD_E_ARG; }
    if (!p->copyStream) {
        RD_CUDA(cudaStreamCreateWithFlags(&p->copyStream, cudaStreamNonBlocking));
        RD_CUDA(cudaMalloc(&p->bandDone, 64 * sizeof(unsigned)));
        RD_CUDA(cudaMemset(p->bandDone, 0, 64 * sizeof(unsigned)));
        p->bandDoneCap = 64;
        void* flags = 0;
        RD_CUDA(cudaHostAlloc(&flags, 64 * sizeof(unsigned), cudaHostAllocMapped));
        memset(flags, 0, 64 * sizeof(unsigned));
        p->bandFlags = (volatile unsigned*)flags;
        RD_CUDA(cudaHostGetDevicePointer((void**)&p->bandFlagsDev, flags, 0));
    }
    p->mirrorObj = o; p->mirrorHost = host; p->mirrorItemBytes = itemBytes; p->mirrorBands = bands;
    return RD_OK;
}

extern "C" int rd_image_write(rd_platform* p, rd_handle h, unsigned width, unsigned height, size_t layer, const void* data)
{
    RdHandleObj* o = rd_obj(h);
    if (!o || o->kind != RD_H_IMAGE_ARRAY || width != o->width || height != o->height || layer >= o->layers) {
        rd_set_error("rd_image_write: handle/size/layer mismatch");
        return RD_E_ARG;
    }
    size_t bytes = (size_t)width * height * 4;
    return rd_buffer_write(p, h, bytes, data, bytes * layer);
}

extern "C" int rd_image_read(rd_platform* p, rd_handle h, unsigned width, unsigned height, size_t layer, void* data)
{
    RdHandleObj* o = rd_obj(h);
    if (!o || o->kind != RD_H_IMAGE_ARRAY || width != o->width || height != o->height || layer >= o->layers) {
        rd_set_error("rd_image_read: handle/size/layer mismatch");
        return RD_E_ARG;
    }
    size_t bytes = (size_t)width * height * 4;
    return rd_buffer_read(p, h, bytes, data, bytes * layer);
}

extern "C" int rd_handle_free(rd_platform* p, rd_handle h)
{
    RdHandleObj* o = rd_obj(h);
    if (!o) return RD_OK;
    if (p) cudaSetDevice(p->device);
    if (p && o == p->mirrorObj) {   // streamed read-back ends with its buffer
        if (p->copyPending) rd_stream_bands_finish(p);
        p->mirrorObj = 0; p->mirrorHost = 0; p->mirrorFresh = 0; p->copyPending = 0;
    }
    if (o->kind == RD_H_TLAS) { cudaFree(o->dnodes); cudaFree(o->dleaf); cudaFree(o->dinst); }
    if (o->dptr && !o->external) cudaFree(o->dptr);
    if (o->ddesc) cudaFree(o->ddesc);
    o->magic = 0;
    delete o;
    return RD_OK;
}

extern "C" void* rd_handle_device_ptr(rd_handle h)
{
    RdHandleObj* o = rd_obj(h);
    return o ? o->dptr : 0;
}

extern "C" int rd_set_option(rd_platform* p, const char* key, long value)
{
    if (!p || !key) { rd_set_error("rd_set_option: null argument"); return RD_E_ARG; }
    if (!strcmp(key, "fmad")) p->optFmad = value;
    else if (!strcmp(key, "tile")) p->optTile = value;
    else if (!strcmp(key, "count_rays")) p->optCountRays = value;
    else if (!strcmp(key, "wavefront")) p->optWavefront = value;
    else if (!strcmp(key, "wf_sort")) p->optWfSort = value;
    else if (!strcmp(key, "wf_dynamic")) p->optWfDynamic = value;
    else if (!strcmp(key, "wf_ray_sort")) p->optWfRaySort = value;
    else if (!strcmp(key, "wf_timeline")) p->optWfTimeline = value;
    else if (!strcmp(key, "async_build")) p->optAsyncBuild = value;
    else if (!strcmp(key, "prec_div")) p->optPrecDiv = value;
    else if (!strcmp(key, "l2_persist")) p->optL2Persist = value;
    else if (!strcmp(key, "trim_pool")) {   // give the build scratch and the wavefront queues back to the driver (value ignored)
        cudaSetDevice(p->device);
        cudaStreamSynchronize(p->stream);
        rd_wf_release(p);
        cudaMemPool_t pool;
        if (cudaDeviceGetDefaultMemPool(&pool, p->device) == cudaSuccess) cudaMemPoolTrimTo(pool, 0);
        for (cudaStream_t bs : p->buildStreams) cudaStreamSynchronize(bs);
        rd_arena_trim(p);
    }
    else if (!strcmp(key, "wf_lanes")) p->optWfLanes = value < 1 ? 1 : (value > RD_WF_MAX_LANES ? RD_WF_MAX_LANES : value);
    else if (!strcmp(key, "wf_paths")) p->optWfPaths = value < 1 ? 1 : value;   // paths in flight per wavefront pass (0/1 = one sample per pass)
    else if (!strcmp(key, "stealing")) p->optSteal = value;
    else if (!strcmp(key, "stream_image")) p->optStreamImage = value;   // modules created from now on carry the band-reporting raygen
    else if (!strcmp(key, "max_trace_depth")) { if (value < 1 || value > 8) { rd_set_error("max_trace_depth must be in [1,8]"); return RD_E_ARG; } p->optMaxDepth = value; }
    else if (!strcmp(key, "block")) { if (value < 32 || value > 1024 || (value & 31)) { rd_set_error("block must be a multiple of 32 in [32,1024]"); return RD_E_ARG; } p->optBlock = value; }
    else { rd_set_error("rd_set_option: unknown key '%s'", key); return RD_E_ARG; }
    return RD_OK;
}

extern "C" uint64_t rd_launch_count(rd_platform* p) { return p ? p->launches : 0; }
