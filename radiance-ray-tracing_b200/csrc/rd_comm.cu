// rd_comm.cu -- the multi-GPU frame collective behind the C ABI.
//
// The path shards by pixel: one process (or host thread) per GPU, the scene replicated, the frame
// partitioned into bands of 8x4-tile rows; the ONLY exchange of a frame is the gather of the finished framebuffer bands
// on rank 0. The reference renders one blocking frame on one device (radiance/src/radiance.cpp:242-267) and reads it
// back (radiance.cpp:186-199); a C++ host in its idiom (samples/sample1.cpp:363-443) gets the multi-GPU frame from
//     rd_comm_init(...)  once,  then per frame  rd_trace_rays_interleaved(...) + rd_frame_gather(...) + rd_buffer_read on rank 0.
//
// rd_frame_gather: every rank but 0 packs the bands it owns into one contiguous staging buffer (one kernel), sends it
// with ONE ncclSend; rank 0 posts one ncclRecv per peer in the same group and scatters the received bands into its
// framebuffer (one kernel). 1/N of the bytes of the sum-reduce it replaces, and no per-frame memset of the framebuffer.
// NCCL is reached through dlopen (libnccl.so.2 -- inside a torchrun process that is the copy torch already loaded), so
// the library still loads on machines without NCCL; every entry point fails loudly when it is missing.
#include <dlfcn.h>
#include <string.h>
#include "rd_internal.h"

typedef struct ncclComm* ncclComm_t;
typedef struct { char internal[128]; } ncclUniqueId;
typedef int ncclResult_t;
enum { ncclSuccess = 0 };
enum { ncclInt8 = 0, ncclChar = 0, ncclUint8 = 1, ncclInt32 = 2, ncclUint32 = 3, ncclUint64 = 5 };

struct RdNccl {
    void* so;
    ncclResult_t (*GetUniqueId)(ncclUniqueId*);
    ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int);
    ncclResult_t (*CommDestroy)(ncclComm_t);
    ncclResult_t (*Send)(const void*, size_t, int, int, ncclComm_t, cudaStream_t);
    ncclResult_t (*Recv)(void*, size_t, int, int, ncclComm_t, cudaStream_t);
    ncclResult_t (*AllGather)(const void*, void*, size_t, int, ncclComm_t, cudaStream_t);
    ncclResult_t (*GroupStart)();
    ncclResult_t (*GroupEnd)();
    const char* (*GetErrorString)(ncclResult_t);
    bool ok;
};
static RdNccl g_nccl;

static int rd_nccl_init()
{
    if (g_nccl.ok) return RD_OK;
    const char* names[] = {"libnccl.so.2", "libnccl.so", 0};
    void* so = 0;
    if (const char* e = getenv("RADIANCE_B200_NCCL")) so = dlopen(e, RTLD_NOW | RTLD_GLOBAL);
    for (int i = 0; !so && names[i]; ++i) so = dlopen(names[i], RTLD_NOW | RTLD_GLOBAL);
    if (!so) { rd_set_error("multi-GPU: libnccl.so.2 not found (%s)", dlerror()); return RD_E_STATE; }
    struct { const char* name; void** slot; } tab[] = {
        {"ncclGetUniqueId", (void**)&g_nccl.GetUniqueId}, {"ncclCommInitRank", (void**)&g_nccl.CommInitRank},
        {"ncclCommDestroy", (void**)&g_nccl.CommDestroy}, {"ncclSend", (void**)&g_nccl.Send}, {"ncclRecv", (void**)&g_nccl.Recv},
        {"ncclAllGather", (void**)&g_nccl.AllGather}, {"ncclGroupStart", (void**)&g_nccl.GroupStart},
        {"ncclGroupEnd", (void**)&g_nccl.GroupEnd}, {"ncclGetErrorString", (void**)&g_nccl.GetErrorString}};
    for (auto& t : tab) {
        *t.slot = dlsym(so, t.name);
        if (!*t.slot) { rd_set_error("multi-GPU: %s missing from libnccl", t.name); return RD_E_STATE; }
    }
    g_nccl.so = so; g_nccl.ok = true;
    return RD_OK;
}
#define RD_NCCL(call)                                                                                     \
    do {                                                                                                  \
        ncclResult_t _r = (call);                                                                         \
        if (_r != ncclSuccess) {                                                                          \
            rd_set_error("%s failed: %s (%s:%d)", #call, g_nccl.GetErrorString(_r), __FILE__, __LINE__);  \
            return RD_E_CUDA;                                                                             \
        }                                                                                                 \
    } while (0)

struct RdComm {
    ncclComm_t comm; int rank, world; bool owned;
    unsigned char* stage; size_t stageBytes;   // rank 0: the packed bands of all peers, peer r at (r - 1) * share; others: own bands
};

extern "C" int rd_comm_unique_id(unsigned char id[128])
{
    if (!id) { rd_set_error("rd_comm_unique_id: null argument"); return RD_E_ARG; }
    int rc = rd_nccl_init();
    if (rc) return rc;
    ncclUniqueId u;
    RD_NCCL(g_nccl.GetUniqueId(&u));
    memcpy(id, &u, 128);
    return RD_OK;
}

static int rd_comm_attach(rd_platform* p, ncclComm_t comm, int rank, int world, bool owned)
{
    RdComm* c = new RdComm();
    c->comm = comm; c->rank = rank; c->world = world; c->owned = owned; c->stage = 0; c->stageBytes = 0;
    p->comm = c;
    return RD_OK;
}

extern "C" int rd_comm_init(rd_platform* p, int rank, int world, const unsigned char id[128])
{
    if (!p || !id || world < 1 || rank < 0 || rank >= world) { rd_set_error("rd_comm_init: bad argument"); return RD_E_ARG; }
    if (p->comm) { rd_set_error("rd_comm_init: the platform already has a communicator"); return RD_E_STATE; }
    int rc = rd_nccl_init();
    if (rc) return rc;
    RD_CUDA(cudaSetDevice(p->device));
    ncclUniqueId u; memcpy(&u, id, 128);
    ncclComm_t comm = 0;
    RD_NCCL(g_nccl.CommInitRank(&comm, world, u, rank));
    return rd_comm_attach(p, comm, rank, world, true);
}

// adopt a communicator the host program already owns (ncclComm_t as void*); it is not destroyed by rd_comm_destroy
extern "C" int rd_comm_adopt(rd_platform* p, void* ncclComm, int rank, int world)
{
    if (!p || !ncclComm || world < 1 || rank < 0 || rank >= world) { rd_set_error("rd_comm_adopt: bad argument"); return RD_E_ARG; }
    if (p->comm) { rd_set_error("rd_comm_adopt: the platform already has a communicator"); return RD_E_STATE; }
    int rc = rd_nccl_init();
    if (rc) return rc;
    return rd_comm_attach(p, (ncclComm_t)ncclComm, rank, world, false);
}

extern "C" int rd_comm_destroy(rd_platform* p)
{
    if (!p || !p->comm) return RD_OK;
    RdComm* c = (RdComm*)p->comm;
    cudaSetDevice(p->device);
    cudaStreamSynchronize(p->stream);
    if (c->stage) cudaFree(c->stage);
    if (c->owned && g_nccl.ok) g_nccl.CommDestroy(c->comm);
    delete c;
    p->comm = 0;
    return RD_OK;
}

// ---- band bookkeeping of the interleaved partition (the same arithmetic as get_global_id in cl_compat.h) ------------
// band b = tile rows [b * bandRows, min((b + 1) * bandRows, tileRows)), owned by rank b % world
struct RdBandPlan { unsigned tileRows, bandRows, world, rowBytes /* bytes of one TILE row = 4 image rows */; };

static size_t rd_owned_tile_rows(const RdBandPlan& P, unsigned rank)
{
    size_t rows = 0;
    for (unsigned b = rank; (size_t)b * P.bandRows < P.tileRows; b += P.world) {
        const unsigned r0 = b * P.bandRows, r1 = r0 + P.bandRows < P.tileRows ? r0 + P.bandRows : P.tileRows;
        rows += r1 - r0;
    }
    return rows;
}

// pack (toStage = 1): stage[k-th owned tile row] <- image[its place in the frame]; unpack (toStage = 0): the reverse
__global__ void rd_k_band_copy(unsigned char* __restrict__ image, unsigned char* __restrict__ stage, RdBandPlan P, unsigned rank,
                               unsigned ownedRows, int toStage)
{
    const unsigned vecPerRow = P.rowBytes >> 4;   // 16-byte vectors of one tile row
    const size_t total = (size_t)ownedRows * vecPerRow;
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
        const unsigned k = (unsigned)(i / vecPerRow), v = (unsigned)(i - (size_t)k * vecPerRow);
        const unsigned band = k / P.bandRows, within = k - band * P.bandRows;     // k-th owned row: owned band `band`, row `within`
        const unsigned tileRow = (band * P.world + rank) * P.bandRows + within;
        if (tileRow >= P.tileRows) continue;   // cannot happen for the rows a rank owns; keeps a wrong plan from writing outside the frame
        uint4* a = (uint4*)(image + (size_t)tileRow * P.rowBytes) + v;
        uint4* s = (uint4*)(stage + (size_t)k * P.rowBytes) + v;
        if (toStage) *s = *a; else *a = *s;
    }
}

// Gathers the frame on rank 0 after every rank rendered its bands with rd_trace_rays_interleaved(width, height, rank,
// world, bandTileRows) into `image` (a buffer of itemBytes per pixel, row-major: the RGBA8 image, itemBytes = 4, or
// imageScratch, itemBytes = 16). Asynchronous on the platform stream; rank 0 reads the frame with rd_buffer_read.
extern "C" int rd_frame_gather(rd_platform* p, rd_handle image, unsigned width, unsigned height, unsigned itemBytes, unsigned bandTileRows)
{
    if (!p || !p->comm) { rd_set_error("rd_frame_gather: no communicator (rd_comm_init)"); return RD_E_STATE; }
    RdHandleObj* o = rd_obj(image);
    if (!o || !itemBytes || !bandTileRows || (width % 8u) || (height % 4u) || o->size < (size_t)width * height * itemBytes ||
        (((size_t)4 * width * itemBytes) & 15u)) {
        rd_set_error("rd_frame_gather: needs a buffer of width * height * itemBytes bytes, width %% 8 == 0, height %% 4 == 0");
        return RD_E_ARG;
    }
    RdComm* c = (RdComm*)p->comm;
    if (c->world == 1) return RD_OK;
    RD_CUDA(cudaSetDevice(p->device));
    RdBandPlan P; P.tileRows = height / 4u; P.bandRows = bandTileRows; P.world = (unsigned)c->world; P.rowBytes = 4u * width * itemBytes;
    // staging: the largest share any rank can own, so that peer r's bands land at a fixed offset on rank 0
    const size_t share = rd_owned_tile_rows(P, 0) * P.rowBytes;   // rank 0 owns the most rows (bands are dealt from rank 0 on)
    const size_t need = c->rank == 0 ? share * (size_t)(c->world - 1) : share;
    if (c->stageBytes < need) {
        if (c->stage) cudaFree(c->stage);
        RD_CUDA(cudaMalloc(&c->stage, need ? need : 16));
        c->stageBytes = need;
    }
    const unsigned blocks = (unsigned)p->smCount * 4u;
    if (c->rank != 0) {
        const unsigned rows = (unsigned)rd_owned_tile_rows(P, (unsigned)c->rank);
        if (rows) {
            rd_k_band_copy<<<blocks, 256, 0, p->stream>>>((unsigned char*)o->dptr, c->stage, P, (unsigned)c->rank, rows, 1);
            p->launches += 1;
            RD_NCCL(g_nccl.Send(c->stage, (size_t)rows * P.rowBytes, ncclUint8, 0, c->comm, p->stream));
        }
    } else {
        RD_NCCL(g_nccl.GroupStart());
        for (int r = 1; r < c->world; ++r) {
            const size_t rows = rd_owned_tile_rows(P, (unsigned)r);
            if (rows) RD_NCCL(g_nccl.Recv(c->stage + (size_t)(r - 1) * share, rows * P.rowBytes, ncclUint8, r, c->comm, p->stream));
        }
        RD_NCCL(g_nccl.GroupEnd());
        for (int r = 1; r < c->world; ++r) {
            const unsigned rows = (unsigned)rd_owned_tile_rows(P, (unsigned)r);
            if (!rows) continue;
            rd_k_band_copy<<<blocks, 256, 0, p->stream>>>((unsigned char*)o->dptr, c->stage + (size_t)(r - 1) * share, P, (unsigned)r, rows, 0);
            p->launches += 1;
        }
    }
    RD_CUDA(cudaGetLastError());
    return RD_OK;
}

// Gathers contiguous work-item ranges (the static partition of rd_trace_rays_range, and the work units of the dynamic
// wavefront partition): rank r owns `count[r]` items starting at `first[r]` (both arrays of `world` entries, identical on
// every rank); the ranges are contiguous in the buffer, so they travel without packing.
extern "C" int rd_frame_gather_ranges(rd_platform* p, rd_handle image, unsigned itemBytes, const unsigned long long* first,
                                      const unsigned long long* count)
{
    if (!p || !p->comm) { rd_set_error("rd_frame_gather_ranges: no communicator (rd_comm_init)"); return RD_E_STATE; }
    RdHandleObj* o = rd_obj(image);
    if (!o || !itemBytes || !first || !count) { rd_set_error("rd_frame_gather_ranges: bad argument"); return RD_E_ARG; }
    RdComm* c = (RdComm*)p->comm;
    if (c->world == 1) return RD_OK;
    for (int r = 0; r < c->world; ++r)
        if ((first[r] + count[r]) * itemBytes > o->size) { rd_set_error("rd_frame_gather_ranges: range of rank %d outside the buffer", r); return RD_E_ARG; }
    RD_CUDA(cudaSetDevice(p->device));
    unsigned char* base = (unsigned char*)o->dptr;
    if (c->rank != 0) {
        if (count[c->rank]) RD_NCCL(g_nccl.Send(base + first[c->rank] * itemBytes, count[c->rank] * itemBytes, ncclUint8, 0, c->comm, p->stream));
    } else {
        RD_NCCL(g_nccl.GroupStart());
        for (int r = 1; r < c->world; ++r)
            if (count[r]) RD_NCCL(g_nccl.Recv(base + first[r] * itemBytes, count[r] * itemBytes, ncclUint8, r, c->comm, p->stream));
        RD_NCCL(g_nccl.GroupEnd());
    }
    return RD_OK;
}

// Dynamic partition: work units (equal runs of tile rows) are taken by whichever rank is free (rd_trace_rays_units);
// unitOwner[u] on rank 0 is unknown until the frame is done, so every rank contributes the bit mask of the units it
// rendered (one 8-byte all-gather), then the units travel like ranges. Synchronises the platform stream once.
extern "C" int rd_frame_gather_units(rd_platform* p, rd_handle image, unsigned width, unsigned height, unsigned itemBytes,
                                     unsigned units, unsigned long long ownedMask)
{
    if (!p || !p->comm) { rd_set_error("rd_frame_gather_units: no communicator (rd_comm_init)"); return RD_E_STATE; }
    RdHandleObj* o = rd_obj(image);
    if (!o || !itemBytes || !units || units > 64u || (height % 4u) || o->size < (size_t)width * height * itemBytes) {
        rd_set_error("rd_frame_gather_units: bad argument (at most 64 units, height %% 4 == 0)"); return RD_E_ARG;
    }
    RdComm* c = (RdComm*)p->comm;
    if (c->world == 1) return RD_OK;
    RD_CUDA(cudaSetDevice(p->device));
    if (c->stageBytes < 16u * (size_t)c->world + 16u) {
        if (c->stage) cudaFree(c->stage);
        RD_CUDA(cudaMalloc(&c->stage, 16u * (size_t)c->world + 16u));
        c->stageBytes = 16u * (size_t)c->world + 16u;
    }
    unsigned long long* dmine = (unsigned long long*)c->stage;
    unsigned long long* dall = dmine + 1;
    RD_CUDA(cudaMemcpyAsync(dmine, &ownedMask, 8, cudaMemcpyHostToDevice, p->stream));
    RD_NCCL(g_nccl.AllGather(dmine, dall, 1, ncclUint64, c->comm, p->stream));
    std::vector<unsigned long long> masks((size_t)c->world);
    RD_CUDA(cudaMemcpyAsync(masks.data(), dall, 8u * (size_t)c->world, cudaMemcpyDeviceToHost, p->stream));
    RD_CUDA(cudaStreamSynchronize(p->stream));
    const unsigned tileRows = height / 4u;
    const size_t rowBytes = (size_t)4u * width * itemBytes;
    unsigned char* base = (unsigned char*)o->dptr;
    unsigned long long seen = 0;
    RD_NCCL(g_nccl.GroupStart());
    for (int r = 1; r < c->world; ++r) {
        if (c->rank != 0 && c->rank != r) continue;
        for (unsigned u = 0; u < units; ++u) {
            if (!((masks[(size_t)r] >> u) & 1ull)) continue;
            const unsigned r0 = (unsigned)((unsigned long long)tileRows * u / units), r1 = (unsigned)((unsigned long long)tileRows * (u + 1u) / units);
            if (r1 == r0) continue;
            if (c->rank == 0) RD_NCCL(g_nccl.Recv(base + r0 * rowBytes, (r1 - r0) * rowBytes, ncclUint8, r, c->comm, p->stream));
            else RD_NCCL(g_nccl.Send(base + r0 * rowBytes, (r1 - r0) * rowBytes, ncclUint8, 0, c->comm, p->stream));
        }
    }
    RD_NCCL(g_nccl.GroupEnd());
    for (int r = 0; r < c->world; ++r) {
        if (seen & masks[(size_t)r]) { rd_set_error("rd_frame_gather_units: a unit was rendered by two ranks"); return RD_E_STATE; }
        seen |= masks[(size_t)r];
    }
    if (seen != (units == 64u ? ~0ull : ((1ull << units) - 1ull))) { rd_set_error("rd_frame_gather_units: units missing from the frame (mask %llx)", seen); return RD_E_STATE; }
    return RD_OK;
}

// Test hook for the single-GPU suite: packs the bands rank `rank` of `world` owns out of `src` and scatters them into
// `dst` -- what rank `rank` and rank 0 do on either side of the NCCL transfer of rd_frame_gather.
extern "C" int rd_debug_band_roundtrip(rd_platform* p, rd_handle src, rd_handle dst, unsigned width, unsigned height, unsigned itemBytes,
                                       unsigned bandTileRows, unsigned rank, unsigned world)
{
    RdHandleObj *s = rd_obj(src), *d = rd_obj(dst);
    const size_t bytes = (size_t)width * height * itemBytes;
    if (!p || !s || !d || !itemBytes || !bandTileRows || !world || rank >= world || (width % 8u) || (height % 4u) || s->size < bytes ||
        d->size < bytes || (((size_t)4 * width * itemBytes) & 15u)) { rd_set_error("rd_debug_band_roundtrip: bad argument"); return RD_E_ARG; }
    RD_CUDA(cudaSetDevice(p->device));
    RdBandPlan P; P.tileRows = height / 4u; P.bandRows = bandTileRows; P.world = world; P.rowBytes = 4u * width * itemBytes;
    const unsigned rows = (unsigned)rd_owned_tile_rows(P, rank);
    if (!rows) return RD_OK;
    unsigned char* stage = 0;
    RD_CUDA(cudaMalloc(&stage, (size_t)rows * P.rowBytes));
    const unsigned blocks = (unsigned)p->smCount * 4u;
    rd_k_band_copy<<<blocks, 256, 0, p->stream>>>((unsigned char*)s->dptr, stage, P, rank, rows, 1);
    rd_k_band_copy<<<blocks, 256, 0, p->stream>>>((unsigned char*)d->dptr, stage, P, rank, rows, 0);
    p->launches += 2;
    RD_CUDA(cudaStreamSynchronize(p->stream));
    RD_CUDA(cudaGetLastError());
    cudaFree(stage);
    return RD_OK;
}
