// rd_core.cu -- platform, resource handles and blocking transfers of the C ABI.
// Replaces radiance/src/clcontext.cpp:12-40 (context) and radiance/src/radiance.cpp:86-140,186-224
// (CreateBuffer / CreateImage / CreateImageArray / CreateSampler / Read* / Write*).
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
    p->optFmad = 1; p->optTile = 1; p->optBlock = 128; p->optCountRays = 0; p->optMaxDepth = 2; p->optWavefront = 0; p->optWfSort = 1; p->optSteal = 0; p->optWfDynamic = 1; p->optWfRaySort = 0; p->optWfPaths = 1l << 25;
    p->ilWorld = p->ilRank = p->ilBandRows = 0;
    p->stealCounters = 0; p->stealOwned = 0; p->stealImported = 0; p->stealSlots = 0; p->stealOwnedCap = 0;
    memset(&p->wf, 0, sizeof(p->wf));
    p->launches = 0;
    RD_CUDA(cudaStreamCreateWithFlags(&p->stream, cudaStreamNonBlocking));
    {   // build scratch is stream-ordered (RdTmp in rd_build.cu): keep freed blocks in the pool instead of returning them to the OS
        cudaMemPool_t pool; unsigned long long keep = ~0ull;
        if (cudaDeviceGetDefaultMemPool(&pool, device) == cudaSuccess) cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &keep);
    }
    RD_CUDA(cudaMalloc(&p->dstatus, 64));
    RD_CUDA(cudaMemset(p->dstatus, 0, 64));
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
    rd_wf_release(p);
    { cudaMemPool_t pool; if (cudaDeviceGetDefaultMemPool(&pool, p->device) == cudaSuccess) cudaMemPoolTrimTo(pool, 0); }
    cudaFree(p->dstatus);
    cudaFree(p->dcounters);
    cudaStreamDestroy(p->stream);
    delete p;
    return RD_OK;
}

extern "C" void* rd_platform_stream(rd_platform* p) { return p ? (void*)p->stream : 0; }

extern "C" int rd_platform_sync(rd_platform* p)
{
    if (!p) { rd_set_error("null platform"); return RD_E_ARG; }
    RD_CUDA(cudaStreamSynchronize(p->stream));
    return RD_OK;
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

extern "C" int rd_buffer_write(rd_platform* p, rd_handle h, size_t size, const void* data, size_t offset)
{
    RdHandleObj* o = rd_obj(h);
    if (!p || (!data && size)) { rd_set_error("rd_buffer_write: null argument"); return RD_E_ARG; }
    int rc = check_range(o, size, offset, "rd_buffer_write");
    if (rc) return rc;
    if (!size) return RD_OK;
    RD_CUDA(cudaSetDevice(p->device));
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
    RD_CUDA(cudaMemcpyAsync(data, (const char*)o->dptr + offset, size, cudaMemcpyDeviceToHost, p->stream));
    RD_CUDA(cudaStreamSynchronize(p->stream));
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
    else if (!strcmp(key, "trim_pool")) {   // give the build scratch and the wavefront queues back to the driver (value ignored)
        cudaSetDevice(p->device);
        cudaStreamSynchronize(p->stream);
        rd_wf_release(p);
        cudaMemPool_t pool;
        if (cudaDeviceGetDefaultMemPool(&pool, p->device) == cudaSuccess) cudaMemPoolTrimTo(pool, 0);
    }
    else if (!strcmp(key, "wf_paths")) p->optWfPaths = value < 1 ? 1 : value;   // paths in flight per wavefront pass (0/1 = one sample per pass)
    else if (!strcmp(key, "stealing")) p->optSteal = value;
    else if (!strcmp(key, "max_trace_depth")) { if (value < 1 || value > 8) { rd_set_error("max_trace_depth must be in [1,8]"); return RD_E_ARG; } p->optMaxDepth = value; }
    else if (!strcmp(key, "block")) { if (value < 32 || value > 1024 || (value & 31)) { rd_set_error("block must be a multiple of 32 in [32,1024]"); return RD_E_ARG; } p->optBlock = value; }
    else { rd_set_error("rd_set_option: unknown key '%s'", key); return RD_E_ARG; }
    return RD_OK;
}

extern "C" uint64_t rd_launch_count(rd_platform* p) { return p ? p->launches : 0; }
