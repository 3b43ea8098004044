// rd_program.cu -- user programs: NVRTC compilation of OpenCL-C shader files, sbt.json dispatch
// generation, descriptor binding and the TraceRays launch.
// Replaces radiance/src/radiance.cpp:142-184 (descriptor set / pipeline objects, CreateShaderModule
// through clBuildProgram with -I SHADER_LIB_PATH), :226-267 (BindPipeline, BindDescriptorSet,
// TraceRays) and tools/genSBT.py (switch generation from sbt.json).
#include <time.h>
#include <nvrtc.h>
#include <ctype.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <string>
#include <vector>
#include "rd_internal.h"
#include "rd_embedded.h"
#include "rd_wavefront.h"

// The driver API is reached through cudaGetDriverEntryPoint instead of linking libcuda.so.1, so the
// library (and its symbol table) can be loaded on a machine without a driver; every entry point is
// resolved on first use and fails loudly when there is no driver.
struct RdDriver {
    CUresult (*ModuleLoadData)(CUmodule*, const void*);
    CUresult (*ModuleUnload)(CUmodule);
    CUresult (*ModuleGetFunction)(CUfunction*, CUmodule, const char*);
    CUresult (*ModuleGetGlobal)(CUdeviceptr*, size_t*, CUmodule, const char*);
    CUresult (*FuncGetAttribute)(int*, CUfunction_attribute, CUfunction);
    CUresult (*LaunchKernel)(CUfunction, unsigned, unsigned, unsigned, unsigned, unsigned, unsigned, unsigned, CUstream, void**, void**);
    CUresult (*MemcpyHtoDAsync)(CUdeviceptr, const void*, size_t, CUstream);
    CUresult (*GetErrorString)(CUresult, const char**);
    CUresult (*OccupancyMaxActiveBlocksPerMultiprocessor)(int*, CUfunction, int, size_t);
    CUresult (*FuncSetAttribute)(CUfunction, CUfunction_attribute, int);
    bool ok;
};
static RdDriver g_drv;

static int rd_driver_init()
{
    if (g_drv.ok) return RD_OK;
    struct { const char* name; void** slot; } tab[] = {
        {"cuModuleLoadData", (void**)&g_drv.ModuleLoadData}, {"cuModuleUnload", (void**)&g_drv.ModuleUnload},
        {"cuModuleGetFunction", (void**)&g_drv.ModuleGetFunction}, {"cuModuleGetGlobal", (void**)&g_drv.ModuleGetGlobal},
        {"cuFuncGetAttribute", (void**)&g_drv.FuncGetAttribute}, {"cuLaunchKernel", (void**)&g_drv.LaunchKernel},
        {"cuMemcpyHtoDAsync", (void**)&g_drv.MemcpyHtoDAsync}, {"cuGetErrorString", (void**)&g_drv.GetErrorString},
        {"cuOccupancyMaxActiveBlocksPerMultiprocessor", (void**)&g_drv.OccupancyMaxActiveBlocksPerMultiprocessor},
        {"cuFuncSetAttribute", (void**)&g_drv.FuncSetAttribute}};
    for (auto& e : tab) {
        cudaDriverEntryPointQueryResult q;
        cudaError_t err = cudaGetDriverEntryPoint(e.name, e.slot, cudaEnableDefault, &q);
        if (err != cudaSuccess || q != cudaDriverEntryPointSuccess || !*e.slot) {
            rd_set_error("CUDA driver entry point %s unavailable (%s)", e.name, cudaGetErrorString(err));
            return RD_E_CUDA;
        }
    }
    g_drv.ok = true;
    return RD_OK;
}
#define cuModuleLoadData g_drv.ModuleLoadData
#define cuModuleUnload g_drv.ModuleUnload
#define cuModuleGetFunction g_drv.ModuleGetFunction
#undef cuModuleGetGlobal
#define cuModuleGetGlobal g_drv.ModuleGetGlobal
#define cuFuncGetAttribute g_drv.FuncGetAttribute
#define cuLaunchKernel g_drv.LaunchKernel
#undef cuMemcpyHtoDAsync
#define cuMemcpyHtoDAsync g_drv.MemcpyHtoDAsync
#define cuGetErrorString g_drv.GetErrorString
#define cuOccupancyMaxActiveBlocksPerMultiprocessor g_drv.OccupancyMaxActiveBlocksPerMultiprocessor
#define cuFuncSetAttribute g_drv.FuncSetAttribute

// ---- sbt.json ---------------------------------------------------------------------------------
struct SbtRow { std::string raygen, closestHit, anyHit, miss; };

static void skip_ws(const char*& s) { while (*s == ' ' || *s == '\t' || *s == '\n' || *s == '\r') ++s; }

static bool parse_string(const char*& s, std::string* out)
{
    skip_ws(s);
    if (*s != '"') return false;
    ++s; out->clear();
    while (*s && *s != '"') { if (*s == '\\' && s[1]) ++s; out->push_back(*s++); }
    if (*s != '"') return false;
    ++s;
    return true;
}

// samples/sbt.json: [ {"raygen": "...", "closestHit": "...", "anyHit": "...", "miss": "..."}, ... ]
static bool parse_sbt(const char* text, std::vector<SbtRow>* rows, std::string* err)
{
    const char* s = text;
    skip_ws(s);
    if (*s != '[') { *err = "sbt.json: expected '['"; return false; }
    ++s; skip_ws(s);
    if (*s == ']') return true;
    for (;;) {
        skip_ws(s);
        if (*s != '{') { *err = "sbt.json: expected '{'"; return false; }
        ++s;
        SbtRow row;
        skip_ws(s);
        while (*s != '}') {
            std::string key, val;
            if (!parse_string(s, &key)) { *err = "sbt.json: expected key string"; return false; }
            skip_ws(s);
            if (*s != ':') { *err = "sbt.json: expected ':'"; return false; }
            ++s;
            if (!parse_string(s, &val)) { *err = "sbt.json: expected string value for '" + key + "'"; return false; }
            if (key == "raygen") row.raygen = val;
            else if (key == "closestHit") row.closestHit = val;
            else if (key == "anyHit") row.anyHit = val;
            else if (key == "miss") row.miss = val;
            else { *err = "sbt.json: unknown key '" + key + "'"; return false; }
            skip_ws(s);
            if (*s == ',') { ++s; skip_ws(s); }
            else if (*s != '}') { *err = "sbt.json: expected ',' or '}'"; return false; }
        }
        ++s;
        rows->push_back(row);
        skip_ws(s);
        if (*s == ',') { ++s; continue; }
        if (*s == ']') return true;
        *err = "sbt.json: expected ',' or ']'";
        return false;
    }
}

static bool g_precDiv = false;   // set from the platform option right before a module is compiled

static std::string strip_comments(const std::string& src)
{
    std::string o; o.reserve(src.size());
    size_t i = 0, n = src.size();
    while (i < n) {
        if (src[i] == '/' && i + 1 < n && src[i + 1] == '/') { while (i < n && src[i] != '\n') ++i; }
        else if (src[i] == '/' && i + 1 < n && src[i + 1] == '*') { i += 2; while (i + 1 < n && !(src[i] == '*' && src[i + 1] == '/')) ++i; i += 2; o.push_back(' '); }
        else o.push_back(src[i++]);
    }
    return o;
}

// OpenCL address-space / kernel keywords without underscores -> their double-underscore spellings
// (which cl_compat.h maps to nothing / extern "C" __global__). Whole identifiers only, outside
// comments, string and character literals.
static std::string rewrite_bare_qualifiers(const std::string& src)
{
    std::string o; o.reserve(src.size() + 64);
    size_t i = 0, n = src.size();
    while (i < n) {
        const char c = src[i];
        if (c == '/' && i + 1 < n && src[i + 1] == '/') { while (i < n && src[i] != '\n') o.push_back(src[i++]); }
        else if (c == '/' && i + 1 < n && src[i + 1] == '*') {
            o.push_back(src[i++]); o.push_back(src[i++]);
            while (i < n && !(src[i] == '*' && i + 1 < n && src[i + 1] == '/')) o.push_back(src[i++]);
            if (i < n) { o.push_back(src[i++]); if (i < n) o.push_back(src[i++]); }
        }
        else if (c == '"' || c == '\'') {
            o.push_back(src[i++]);
            while (i < n && src[i] != c) { if (src[i] == '\\' && i + 1 < n) o.push_back(src[i++]); o.push_back(src[i++]); }
            if (i < n) o.push_back(src[i++]);
        }
        else if ((c >= 'a' && c <= 'z') || (c >= 'A' && c <= 'Z') || c == '_') {
            size_t b = i;
            while (i < n && ((src[i] >= 'a' && src[i] <= 'z') || (src[i] >= 'A' && src[i] <= 'Z') || (src[i] >= '0' && src[i] <= '9') || src[i] == '_')) ++i;
            const std::string id = src.substr(b, i - b);
            if (id == "global") o += "__global";
            else if (id == "constant") o += "__constant";
            else if (id == "kernel") o += "__kernel";
            else if (id == "local") o += "__local";
            else if (id == "private") o += "__private";
            else o += id;
        }
        else if (c >= '0' && c <= '9') {   // numbers (keeps suffixes such as 1.0f attached)
            while (i < n && ((src[i] >= '0' && src[i] <= '9') || (src[i] >= 'a' && src[i] <= 'z') || (src[i] >= 'A' && src[i] <= 'Z') || src[i] == '.' || src[i] == '_')) o.push_back(src[i++]);
        }
        else o.push_back(src[i++]);
    }
    return o;
}

// true if `name(` appears preceded by the `void` return type, i.e. the program defines the function
static bool defines_function(const std::string& code, const char* name)
{
    const std::string pat = std::string(name) + "(";
    size_t pos = 0;
    while ((pos = code.find(pat, pos)) != std::string::npos) {
        size_t b = pos;
        while (b > 0 && (code[b - 1] == ' ' || code[b - 1] == '\t' || code[b - 1] == '\n' || code[b - 1] == '\r')) --b;
        if (b >= 4 && code.compare(b - 4, 4, "void") == 0) return true;
        pos += pat.size();
    }
    return false;
}

// Row i of the table becomes `case i` (tools/genSBT.py:53-82 semantics) with the call shapes the
// current runtime uses (radiance/shader/radiance.cl:23-28, samples/shader.cl:574-605).
static std::string generate_dispatchers(const std::vector<SbtRow>& rows, bool hit, bool miss, bool any)
{
    std::string s = "\n// ---- dispatchers generated from sbt.json ----\n";
    char buf[512];
    if (hit) {
        s += "void callHit(int sbtRecordOffset, struct Payload* payload, struct HitData* hitData,\n"
             "    struct SceneData* sceneData, image2d_array_t imageArray, sampler_t sampler)\n{\n"
             "    int index = hitData->instanceSBTOffset + sbtRecordOffset;\n    switch (index)\n    {\n";
        for (size_t i = 0; i < rows.size(); ++i)
            if (!rows[i].closestHit.empty()) {
                snprintf(buf, sizeof(buf), "        case %zu: %s(payload, hitData, sceneData, imageArray, sampler); break;\n", i, rows[i].closestHit.c_str());
                s += buf;
            }
        s += "    }\n}\n";
    }
    if (miss) {
        s += "void callMiss(int missIndex, struct Payload* payload,\n"
             "    struct SceneData* sceneData, image2d_array_t imageArray, sampler_t sampler)\n{\n    switch (missIndex)\n    {\n";
        for (size_t i = 0; i < rows.size(); ++i)
            if (!rows[i].miss.empty()) {
                snprintf(buf, sizeof(buf), "        case %zu: %s(payload, sceneData, imageArray, sampler); break;\n", i, rows[i].miss.c_str());
                s += buf;
            }
        s += "    }\n}\n";
    }
    if (any) {
        s += "void callAnyHit(bool* cont, int sbtRecordOffset, struct Payload* payload, struct HitData* hitData,\n"
             "    struct SceneData* sceneData, image2d_array_t imageArray, sampler_t sampler)\n{\n"
             "    int index = hitData->instanceSBTOffset + sbtRecordOffset;\n    switch (index)\n    {\n";
        for (size_t i = 0; i < rows.size(); ++i)
            if (!rows[i].anyHit.empty()) {
                snprintf(buf, sizeof(buf), "        case %zu: %s(cont, payload, hitData, sceneData, imageArray, sampler); break;\n", i, rows[i].anyHit.c_str());
                s += buf;
            }
        s += "    }\n}\n";
    }
    return s;
}


// ---- wavefront stage kernels --------------------------------------------------------------------
// Parameter list of the program's raygen kernel: (type text, name) per parameter, in order. The
// generated stage kernels repeat it so that the same descriptor set binds to them.
struct RaygenParam { std::string type, name; };
static bool parse_raygen_params(const std::string& bare, std::vector<RaygenParam>* out)
{
    size_t pos = 0;
    while ((pos = bare.find("raygen", pos)) != std::string::npos) {
        size_t q = pos + 6;
        while (q < bare.size() && (bare[q] == ' ' || bare[q] == '\t' || bare[q] == '\n' || bare[q] == '\r')) ++q;
        if (q < bare.size() && bare[q] == '(' && bare.rfind("__kernel", pos) != std::string::npos) { pos = q; break; }
        pos += 6;
    }
    if (pos == std::string::npos) return false;
    size_t end = bare.find(')', pos);
    if (end == std::string::npos) return false;
    std::string list = bare.substr(pos + 1, end - pos - 1);
    size_t a = 0;
    while (a <= list.size()) {
        size_t b = list.find(',', a);
        if (b == std::string::npos) b = list.size();
        std::string item = list.substr(a, b - a);
        size_t e = item.find_last_not_of(" \t\r\n");
        if (e != std::string::npos) {
            size_t st = e;
            while (st > 0 && (isalnum((unsigned char)item[st - 1]) || item[st - 1] == '_')) --st;
            RaygenParam rp; rp.name = item.substr(st, e - st + 1); rp.type = item.substr(0, st);
            if (rp.name.empty()) return false;
            out->push_back(rp);
        }
        a = b + 1;
    }
    return !out->empty();
}

// generate / shade / resolve kernels around the user's wavefront hooks:
//   void wavefrontScene(struct SceneData*, <raygen parameters>)      fill SceneData from the descriptors
//   void wavefrontPrimary(struct Payload*, struct SceneData*, uint frameID, uint totalSamples)
//   float3 wavefrontTonemap(float3)                                  optional, default shader.cl:289-298
// The loop they implement is samples/shader.cl:228-281.
static std::string generate_wavefront_kernels(const std::vector<RaygenParam>& P, bool hasTonemap)
{
    std::string plist, names, image = "0", sampler = "0", top;
    for (size_t i = 0; i < P.size(); ++i) {
        plist += P[i].type + " " + P[i].name + ", ";
        names += (i ? ", " : "") + P[i].name;
        if (P[i].type.find("image2d_array_t") != std::string::npos) image = P[i].name;
        else if (P[i].type.find("sampler_t") != std::string::npos) sampler = P[i].name;
        else if (P[i].type.find("AccelStruct") != std::string::npos) top = P[i].name;
    }
    if (top.empty() || P.size() < 3) return "";
    std::string s = "\n// ---- wavefront stages generated by the runtime ----\n#include \"rd_wavefront.h\"\nnamespace rd_l0 {\n";
    s += "extern \"C\" __global__ void rd_wf_generate(" + plist + "const RdWfState* wfp)\n{\n"
         "    const RdWfState wf = *wfp;\n"
         "    const unsigned slot = blockIdx.x * blockDim.x + threadIdx.x;\n"
         "    if (slot >= wf.count * wf.samples) return;\n"
         "    RD_BOUND(slot < wf.capacity, wf.status, return);\n"
         "    const unsigned sample = slot / wf.count, item = slot - sample * wf.count, frameID = wf.frameID + sample;\n"
         "    rd_wf_gid[threadIdx.x] = rd_wf_global_id(wf.first + item, wf.tileW, wf.ilWorld, wf.ilRank, wf.ilBandRows);\n"
         "    struct SceneData sd; struct Payload pl;\n"
         "    wavefrontScene(&sd, " + names + ");\n"
         "    sd.depth = 0; sd.frameID = frameID;\n"
         "    wavefrontPrimary(&pl, &sd, frameID, wf.totalSamples);\n"
         "    pl.hit = false;\n"
         "    *(struct Payload*)(wf.payload + (size_t)slot * wf.payloadStride) = pl;\n"
         "    RdF4 a, b;\n"
         "    a.x = pl.nextRayOrigin.x; a.y = pl.nextRayOrigin.y; a.z = pl.nextRayOrigin.z; a.w = wf.tmin;\n"
         "    b.x = pl.nextRayDirection.x; b.y = pl.nextRayDirection.y; b.z = pl.nextRayDirection.z; b.w = wf.tmax;\n"
         "    wf.rayA[wf.cur][slot] = a; wf.rayB[wf.cur][slot] = b; wf.pathOf[wf.cur][slot] = slot;\n"
         "    // color = 0 and contribution = 1 (shader.cl:231-232) are not stored: the first shade stage starts from them\n"
         "    if (slot == 0) { wf.counts[wf.cur] = wf.count * wf.samples; wf.counts[wf.cur ^ 1u] = 0u; }\n"
         "}\n\n";
    // The shade stage carries the inline shadow traversal: 6 resident 128-thread blocks per SM (80 registers) measured
    // 5 % faster on C3 than the 128 registers ptxas takes unbounded (50.2 -> 47.6 ms; 8 blocks the same). Developer knob:
    // RD_WF_SHADE_MINB=0 removes the bound.
    int minb = 6;
    if (const char* e = getenv("RD_WF_SHADE_MINB")) minb = atoi(e);
    const std::string shadeBounds = minb > 0 ? std::string("__launch_bounds__(128, ") + std::to_string(minb) + ") " : std::string();
    // One queue entry. The kernel below hands entries out in runs of one block: the queue shrinks from bounce to bounce (33 M, 22 M,
    // 3 M, 1 M, 0.3 M entries in a C3 frame) while its length is known on the device only -- a grid sized for the first bounce spent
    // ~0.9 ms per later stage on blocks that found nothing to do (profiles/exp_wf_timeline_r2g.log).
    s += "static __device__ __forceinline__ void rd_wf_shade_entry(" + plist + "const RdWfState& wf, const unsigned n, const unsigned jj)\n{\n"
         "    const unsigned j = wf.sortOn ? wf.order[jj] : jj;\n"
         "    RD_BOUND(j < n, wf.status, return);\n"
         "    const unsigned slot = wf.pathOf[wf.cur][j];\n"
         "    RD_BOUND(slot < wf.capacity && slot < wf.count * wf.samples, wf.status, return);\n"
         "    const unsigned sample = slot / wf.count, item = slot - sample * wf.count;\n"
         "    rd_wf_gid[threadIdx.x] = rd_wf_global_id(wf.first + item, wf.tileW, wf.ilWorld, wf.ilRank, wf.ilBandRows);\n"
         "    struct SceneData sd;\n"
         "    wavefrontScene(&sd, " + names + ");\n"
         "    sd.depth = wf.depth; sd.frameID = wf.frameID + sample;\n"
         "    struct Payload pl = *(struct Payload*)(wf.payload + (size_t)slot * wf.payloadStride);\n"
         "    const RdF4 ha = wf.hitA[j];\n"
         "    if (ha.x < RD_MISS_T) {\n"
         "        const RdF4 ra = wf.rayA[wf.cur][j], rb = wf.rayB[wf.cur][j];\n"
         "        RdHitRec h; h.t = ha.x; h.b1 = ha.y; h.b2 = ha.z; h.prim = RD_F2U(ha.w); h.inst = wf.hitB[j];\n"
         "        struct HitData hd;\n"
         "        rd_fill_hit_data((const RdScene*)" + top + ", float3(ra.x, ra.y, ra.z), float3(rb.x, rb.y, rb.z), h, &hd);\n"
         "        rd_l1::callHit(wf.sbtRecordOffset, (rd_l1::Payload*)&pl, &hd, (rd_l1::SceneData*)&sd, " + image + ", " + sampler + ");\n"
         "    } else {\n"
         "        rd_l1::callMiss(wf.missIndex, (rd_l1::Payload*)&pl, (rd_l1::SceneData*)&sd, " + image + ", " + sampler + ");\n"
         "    }\n"
         "    // bounce loop body of samples/shader.cl:238-260\n"
         "    RdF4 col, con;\n"
         "    if (wf.depth == 0u) { col.x = col.y = col.z = col.w = 0.0f; con.x = con.y = con.z = con.w = 1.0f; }\n"
         "    else { col = wf.color[slot]; con = wf.contrib[slot]; }\n"
         "    bool more = false;\n"
         "    if (pl.hit) {\n"
         "        col.x += con.x * pl.color.x; col.y += con.y * pl.color.y; col.z += con.z * pl.color.z;\n"
         "        con.x *= pl.nextFactor.x; con.y *= pl.nextFactor.y; con.z *= pl.nextFactor.z;\n"
         "        more = true;\n"
         "    } else if (wf.depth == 0) { col.x = pl.color.x; col.y = pl.color.y; col.z = pl.color.z; }\n"
         "    if (wf.depth + 1u >= wf.maxDepth || wf.debug) more = false;\n"
         "    // live paths are compacted into the next queue: one atomic per warp\n"
         "    const unsigned act = __activemask();\n"
         "    const unsigned live = __ballot_sync(act, more);\n"
         "    unsigned base = 0;\n"
         "    const unsigned lane = threadIdx.x & 31u;\n"
         "    if (live) {\n"
         "        const unsigned leader = (unsigned)__ffs(live) - 1u;\n"
         "        if (lane == leader) base = atomicAdd(&wf.counts[wf.cur ^ 1u], (unsigned)__popc(live));\n"
         "        base = __shfl_sync(act, base, leader);\n"
         "    }\n"
         "    if (more) {\n"
         "        const unsigned k = base + (unsigned)__popc(live & ((1u << lane) - 1u));\n"
         "        RD_BOUND(k < wf.capacity && k < n, wf.status, return);\n"
         "        RdF4 a, b;\n"
         "        a.x = pl.nextRayOrigin.x; a.y = pl.nextRayOrigin.y; a.z = pl.nextRayOrigin.z; a.w = wf.tmin;\n"
         "        b.x = pl.nextRayDirection.x; b.y = pl.nextRayDirection.y; b.z = pl.nextRayDirection.z; b.w = wf.tmax;\n"
         "        wf.rayA[wf.cur ^ 1u][k] = a; wf.rayB[wf.cur ^ 1u][k] = b; wf.pathOf[wf.cur ^ 1u][k] = slot;\n"
         "        pl.hit = false;\n"
         "        *(struct Payload*)(wf.payload + (size_t)slot * wf.payloadStride) = pl;\n"
         "        wf.color[slot] = col; wf.contrib[slot] = con;\n"
         "    } else {\n"
         "        wf.color[slot] = col;   // the sample is complete; the accumulate stage folds it into imageScratch in frame order\n"
         "    }\n"
         "}\n\n";
    s += "extern \"C\" __global__ void " + shadeBounds + "rd_wf_shade(" + plist + "const RdWfState* wfp)\n{\n"
         "    const RdWfState wf = *wfp;\n"
         "    const unsigned n = wf.counts[wf.cur];\n"
         "    // persistent warps: every warp takes the next 32 entries until the queue is empty (counts[3]: entries handed out, zeroed by\n"
         "    // rd_k_wf_advance). Fetching per block (one atomic + two barriers per 128 entries) measured 0.3 % slower on C3.\n"
         "    for (;;) {\n"
         "        unsigned run = 0;\n"
         "        if ((threadIdx.x & 31u) == 0u) run = atomicAdd(&wf.counts[3], 32u);\n"
         "        run = __shfl_sync(0xffffffffu, run, 0);\n"
         "        if (run >= n) break;\n"
         "        const unsigned jj = run + (threadIdx.x & 31u);\n"
         "        if (jj < n) rd_wf_shade_entry(" + names + ", wf, n, jj);\n"
         "        __syncwarp();\n"
         "    }\n"
         "}\n\n";
    // running mean of samples/shader.cl:262-279 over the samples of this pass, in frame order
    s += "extern \"C\" __global__ void rd_wf_accumulate(" + plist + "const RdWfState* wfp)\n{\n"
         "    const RdWfState wf = *wfp;\n"
         "    const unsigned item = blockIdx.x * blockDim.x + threadIdx.x;\n"
         "    if (item >= wf.count) return;\n"
         "    const unsigned gid = rd_wf_global_id(wf.first + item, wf.tileW, wf.ilWorld, wf.ilRank, wf.ilBandRows);\n"
         "    float* px = (float*)" + P[1].name + " + 4 * (size_t)gid;\n"
         "    float r = px[0], g = px[1], b = px[2];\n"
         "    for (unsigned k = 0; k < wf.samples; ++k) {\n"
         "        const RdF4 col = wf.color[(size_t)k * wf.count + item];\n"
         "        const unsigned frameID = wf.frameID + k;\n"
         "        if (frameID == 0) { r = col.x; g = col.y; b = col.z; }\n"
         "        else {\n"
         "            r = (frameID * r + col.x) / (frameID + 1);\n"
         "            g = (frameID * g + col.y) / (frameID + 1);\n"
         "            b = (frameID * b + col.z) / (frameID + 1);\n"
         "        }\n"
         "    }\n"
         "    px[0] = r; px[1] = g; px[2] = b;\n"
         "}\n\n";
    s += "extern \"C\" __global__ void rd_wf_resolve(" + plist + "const RdWfState* wfp)\n{\n"
         "    const RdWfState wf = *wfp;\n"
         "    const unsigned slot = blockIdx.x * blockDim.x + threadIdx.x;\n"
         "    if (slot >= wf.count) return;\n"
         "    const unsigned gid = rd_wf_global_id(wf.first + slot, wf.tileW, wf.ilWorld, wf.ilRank, wf.ilBandRows);\n"
         "    const float* px = (const float*)" + P[1].name + " + 4 * (size_t)gid;\n"
         "    float3 color = {px[0], px[1], px[2]};\n"
         "    if (!wf.debug) {\n";
    if (hasTonemap) s += "        color = wavefrontTonemap(color);\n";
    else s += "        // samples/shader.cl:47-56, 289-298: ACES approximation, then pow 0.7\n"
              "        color *= 0.6f;\n"
              "        color = clamp((color * (2.51f * color + 0.03f)) / (color * (2.43f * color + 0.59f) + 0.14f), 0.0f, 1.0f);\n"
              "        color = pow(color, 0.7f);\n";
    s += "    }\n"
         "    uchar* out = (uchar*)" + P[2].name + " + 4 * (size_t)gid;\n"
         "    out[0] = (int)(color.x * 255); out[1] = (int)(color.y * 255); out[2] = (int)(color.z * 255); out[3] = 255;\n"
         "}\n"
         "#ifndef WAVEFRONT_SBT_RECORD_OFFSET\n#define WAVEFRONT_SBT_RECORD_OFFSET 1\n#endif\n"
         "#ifndef WAVEFRONT_MISS_INDEX\n#define WAVEFRONT_MISS_INDEX 3\n#endif\n"
         "#ifndef WAVEFRONT_TMIN\n#define WAVEFRONT_TMIN 0.001f\n#endif\n"
         "#ifndef WAVEFRONT_TMAX\n#define WAVEFRONT_TMAX 1000.0f\n#endif\n"
         "extern \"C\" __global__ void rd_wf_query(float* out)\n{\n"
         "    out[0] = (float)sizeof(struct Payload); out[1] = WAVEFRONT_SBT_RECORD_OFFSET; out[2] = WAVEFRONT_MISS_INDEX;\n"
         "    out[3] = WAVEFRONT_TMIN; out[4] = WAVEFRONT_TMAX;\n}\n"
         "}  // namespace rd_l0\n";
    return s;
}


// Persistent raygen for multi-GPU tile stealing: blocks take chunks of consecutive work-items (runs of 8x4
// screen tiles) from one counter that every GPU of the job increments with system-scope atomics (the counter
// lives on one GPU, the others reach it through a CUDA IPC mapping over NVLink), and run the user's raygen
// body -- a device-function copy of it in namespace rd_ls -- for each work-item of the chunk.
static std::string generate_steal_kernel(const std::vector<RaygenParam>& P)
{
    std::string plist, names;
    for (size_t i = 0; i < P.size(); ++i) { plist += P[i].type + " " + P[i].name + ", "; names += (i ? ", " : "") + P[i].name; }
    std::string s = "\n// ---- persistent raygen with chunk stealing, generated by the runtime ----\n"
        "struct RdStealArgs { unsigned int* counter; unsigned int* owned; unsigned int numChunks, chunkItems, first, count, tileW, pad; };\n"
        "namespace rd_ls {\n"
        "extern \"C\" __global__ void rd_raygen_steal(" + plist + "RdStealArgs a)\n{\n"
        "    // Every warp works on its own: lane 0 takes a chunk (a run of 8x4 tiles) from the job-wide counter with one\n"
        "    // system-scope atomic -- remote over NVLink for every rank but the owner -- and the NEXT chunk is requested\n"
        "    // before the current one is rendered, so the round trip hides behind the tracing.\n"
        "    const unsigned int lane = threadIdx.x & 31u;\n"
        "    unsigned int chunk = 0;\n"
        "    if (lane == 0) chunk = atomicAdd_system(a.counter, 1u);\n"
        "    for (;;) {\n"
        "        unsigned int next = 0;\n"
        "        if (lane == 0 && chunk < a.numChunks) {\n"
        "            next = atomicAdd_system(a.counter, 1u);\n"
        "            if (a.owned) a.owned[1u + atomicAdd(a.owned, 1u)] = chunk;\n"
        "        }\n"
        "        const unsigned int cur = __shfl_sync(0xffffffffu, chunk, 0);\n"
        "        if (cur >= a.numChunks) break;\n"
        "        for (unsigned int w = lane; w < a.chunkItems; w += 32u) {\n"
        "            const unsigned int t = cur * a.chunkItems + w;\n"
        "            if (t < a.count) {\n"
        "                rd_wf_gid[threadIdx.x] = rd_wf_global_id(a.first + t, a.tileW);\n"
        "                raygen(" + names + ");\n"
        "            }\n"
        "            __syncwarp();\n"
        "        }\n"
        "        chunk = next;\n"
        "    }\n"
        "}\n}  // namespace rd_ls\n";
    return s;
}

// Raygen for streamed read-back (rd_buffer_stream_to_host): the user's raygen body -- a device-function copy in
// namespace rd_lb -- for one work-item per thread as in the plain launch, after which the block counts itself off in its
// band's counter; the block that completes a band raises the band's flag in host-mapped memory. The host, which is
// blocked in TraceRays / ReadBuffer anyway, watches the flags and starts each band's copy on the copy stream, so the
// band's image rows go over PCIe while the rest of the frame is still being traced.
// Cost of the reporting, measured on C2 (tools/exp_variants.py, "banded" variants): plain raygen 0.467 ms, banded 0.496 ms;
// without the per-block __threadfence 0.481, without fence and barrier 0.476 (RD_BAND_NOFENCE / RD_BAND_NOSYNC exist for
// that timing experiment only -- without the fence a band could be copied before its last stores have landed). The fence
// holds each of the 16 200 blocks' SM slot about 1 us longer.
static std::string generate_banded_kernel(const std::vector<RaygenParam>& P)
{
    std::string plist, names;
    for (size_t i = 0; i < P.size(); ++i) { plist += P[i].type + " " + P[i].name + ", "; names += (i ? ", " : "") + P[i].name; }
    return "\n// ---- band-reporting raygen, generated by the runtime ----\n"
        "struct RdBandArgs { unsigned int* done; unsigned int* flags; unsigned int blocksPerBand, nBands, epoch, target, targetLast, pad; };\n"
        "namespace rd_lb {\n"
        "extern \"C\" __global__ void rd_raygen_banded(" + plist + "RdBandArgs a)\n{\n"
        "    raygen(" + names + ");\n"
        "#ifndef RD_BAND_NOSYNC\n    __syncthreads();\n#endif\n"
        "    if (threadIdx.x == 0) {\n"
        "        const unsigned int b = blockIdx.x / a.blocksPerBand;\n"
        "#ifndef RD_BAND_NOFENCE\n        __threadfence();\n#endif\n"
        "        if (atomicAdd(a.done + b, 1u) + 1u == (b + 1u == a.nBands ? a.targetLast : a.target)) {\n"
        "            __threadfence_system();   // every block of the band has passed its fence\n"
        "            *(volatile unsigned int*)(a.flags + b) = a.epoch;   // host-mapped memory\n"
        "        }\n"
        "    }\n"
        "}\n}  // namespace rd_lb\n";
}

extern "C" int rd_sbt_load_json(rd_platform* p, const char* json_text)
{
    if (!p || !json_text) { rd_set_error("rd_sbt_load_json: null argument"); return RD_E_ARG; }
    std::vector<SbtRow> rows; std::string err;
    if (!parse_sbt(json_text, &rows, &err)) { rd_set_error("%s", err.c_str()); return RD_E_ARG; }
    p->sbtJson = json_text;
    return RD_OK;
}

// ---- NVRTC ------------------------------------------------------------------------------------
// OpenCL-C user program (+ optional sbt.json) -> sm_100a cubin. Needs no device.
static int rd_compile_program(const char* code, unsigned size, const std::string& sbtJson, bool fmad,
                              std::vector<char>* cubin, std::string* log, bool countRays = false, int maxDepth = 2, bool steal = false,
                              bool banded = false)
{
    std::string src(code, size ? size : strlen(code));
    while (!src.empty() && src.back() == '\0') src.pop_back();
    src = rewrite_bare_qualifiers(src);
    const std::string bare = strip_comments(src);
    const bool needHit = !defines_function(bare, "callHit"), needMiss = !defines_function(bare, "callMiss"),
               needAny = !defines_function(bare, "callAnyHit");
    if (needHit || needMiss || needAny) {
        std::vector<SbtRow> rows; std::string err;
        if (!sbtJson.empty() && !parse_sbt(sbtJson.c_str(), &rows, &err)) { rd_set_error("%s", err.c_str()); return RD_E_ARG; }
        if (sbtJson.empty() && (needHit || needMiss)) {
            rd_set_error("rd_shader_module_create: program defines no callHit/callMiss and no sbt.json is loaded");
            return RD_E_STATE;
        }
        src += generate_dispatchers(rows, needHit, needMiss, needAny);
    }
    // One copy of the user program per traceRay nesting level, each in its own namespace, so that the
    // traceRay -> callHit -> <hit shader> -> traceRay chain has no cycle and inlines completely (see
    // shader/radiance.cl "per-level part"). Level 0 owns the raygen kernel; the last level's traceRay
    // is a stub that raises the depth-exceeded flag.
    const bool hasAny = !needAny || bare.find("callAnyHit") != std::string::npos || src.find("callAnyHit(bool* cont") != std::string::npos;
    std::string mainSrc = "#include \"radiance.cl\"\n#include \"pbr.cl\"\n";
    if (hasAny) mainSrc += "#define RD_HAS_ANYHIT 1\n";
    char line[1024];
    for (int k = 1; k <= maxDepth; ++k) {
        snprintf(line, sizeof(line),
                 "namespace rd_l%d { struct Payload; struct SceneData;\n"
                 "RD_DEV void callHit(int, struct Payload*, struct HitData*, struct SceneData*, image2d_array_t, sampler_t);\n"
                 "RD_DEV void callMiss(int, struct Payload*, struct SceneData*, image2d_array_t, sampler_t);\n"
                 "RD_DEV void callAnyHit(bool*, int, struct Payload*, struct HitData*, struct SceneData*, image2d_array_t, sampler_t); }\n", k);
        mainSrc += line;
    }
    for (int k = 0; k <= maxDepth; ++k) {
        snprintf(line, sizeof(line), "#undef RD_LEVEL\n#define RD_LEVEL %d\n#undef RD_LEVEL_PART_DONE\n#undef RD_NEXT_NS\n#define RD_NEXT_NS rd_l%d\n", k, k + 1);
        mainSrc += line;
        if (k == maxDepth) mainSrc += "#define RD_LAST_LEVEL 1\n";
        if (k == 1) mainSrc += "#undef __kernel\n#define __kernel static __device__\n";
        snprintf(line, sizeof(line), "namespace rd_l%d {\n#include \"rd_user_program.cl\"\n}\n", k);
        mainSrc += line;
    }
    // wavefront stages when the program provides the hooks
    if (defines_function(bare, "wavefrontScene") && defines_function(bare, "wavefrontPrimary")) {
        std::vector<RaygenParam> params;
        if (parse_raygen_params(bare, &params))
            mainSrc += generate_wavefront_kernels(params, bare.find("wavefrontTonemap") != std::string::npos);
    }
    if (steal) {
        std::vector<RaygenParam> params;
        if (parse_raygen_params(bare, &params)) {
            mainSrc += "#undef RD_LEVEL\n#define RD_LEVEL 0\n#undef RD_LEVEL_PART_DONE\n#undef RD_NEXT_NS\n#define RD_NEXT_NS rd_l1\n#undef RD_LAST_LEVEL\n"
                       "#include \"rd_wavefront.h\"\nnamespace rd_ls {\n#include \"rd_user_program.cl\"\n}\n";
            mainSrc += generate_steal_kernel(params);
        }
    }
    if (banded) {
        std::vector<RaygenParam> params;
        if (parse_raygen_params(bare, &params)) {
            mainSrc += "#undef RD_LEVEL\n#define RD_LEVEL 0\n#undef RD_LEVEL_PART_DONE\n#undef RD_NEXT_NS\n#define RD_NEXT_NS rd_l1\n#undef RD_LAST_LEVEL\n"
                       "namespace rd_lb {\n#include \"rd_user_program.cl\"\n}\n";
            mainSrc += generate_banded_kernel(params);
        }
    }
    std::vector<const char*> hdrNames(rd_embedded_names, rd_embedded_names + rd_embedded_count);
    std::vector<const char*> hdrSources(rd_embedded_sources, rd_embedded_sources + rd_embedded_count);
    hdrNames.push_back("rd_user_program.cl");
    hdrSources.push_back(src.c_str());
    nvrtcProgram prog;
    nvrtcResult r = nvrtcCreateProgram(&prog, mainSrc.c_str(), "rd_program_levels.cu", (int)hdrNames.size(), hdrSources.data(), hdrNames.data());
    if (r != NVRTC_SUCCESS) { rd_set_error("nvrtcCreateProgram: %s", nvrtcGetErrorString(r)); return RD_E_COMPILE; }
    std::vector<const char*> opts = {"--gpu-architecture=sm_100a", "--std=c++17", "-default-device", "--pre-include=cl_compat.h",
                                     "-lineinfo", "-w", "--ptxas-options=-v"};
    opts.push_back(fmad ? "--fmad=true" : "--fmad=false");
    // Single-precision division in user programs: OpenCL C requires 2.5 ulp of `/` (the reference builds its programs with
    // clBuildProgram and no -cl-fp32-correctly-rounded-divide-sqrt, radiance/src/radiance.cpp:165-167), and the ~12-instruction
    // IEEE sequence per division is 27 % of the path-tracing raygen (C3 inline 802 -> 1022, C1 910 -> 1296 Mrays/s, images as close
    // to the reference as before: profiles/README_r2.md). The traversal and random_pcg3d round explicitly and are not affected.
    // Option "prec_div" = 1 restores correctly rounded division.
    opts.push_back(g_precDiv ? "--prec-div=true" : "--prec-div=false");
    if (countRays) opts.push_back("-DRD_COUNT_RAYS=1");
    opts.push_back(RD_NODE_F32 ? "-DRD_NODE_F32=1" : "-DRD_NODE_F32=0");
    if (RD_BOUNDS_CHECK) opts.push_back("-DRD_BOUNDS_CHECK=1");   // user programs walk the node form this library builds
    // developer knob for ablations: extra NVRTC flags (e.g. "-DRD_FLUSH_DIV=2 -DRD_DEQUANT_I2F=0")
    std::vector<std::string> extra;
    if (const char* env = getenv("RD_NVRTC_FLAGS")) {
        std::string e(env); size_t a = 0;
        while (a < e.size()) { size_t b = e.find(' ', a); if (b == std::string::npos) b = e.size(); if (b > a) extra.push_back(e.substr(a, b - a)); a = b + 1; }
    }
    for (auto& x : extra) opts.push_back(x.c_str());
    r = nvrtcCompileProgram(prog, (int)opts.size(), opts.data());
    size_t logSize = 0;
    nvrtcGetProgramLogSize(prog, &logSize);
    log->assign(logSize, '\0');
    if (logSize > 1) nvrtcGetProgramLog(prog, &(*log)[0]);
    if (r != NVRTC_SUCCESS) {
        rd_set_error("shader compilation failed (%s):\n%.15000s", nvrtcGetErrorString(r), log->c_str());
        nvrtcDestroyProgram(&prog);
        return RD_E_COMPILE;
    }
    size_t cubinSize = 0;
    nvrtcGetCUBINSize(prog, &cubinSize);
    cubin->resize(cubinSize);
    r = nvrtcGetCUBIN(prog, cubin->data());
    nvrtcDestroyProgram(&prog);
    if (r != NVRTC_SUCCESS || !cubinSize) { rd_set_error("nvrtcGetCUBIN: %s", nvrtcGetErrorString(r)); return RD_E_COMPILE; }
    // developer knob: keep the cubin for cuobjdump -sass / nvdisasm
    if (const char* dump = getenv("RD_DUMP_CUBIN")) {
        if (FILE* f = fopen(dump, "wb")) { fwrite(cubin->data(), 1, cubin->size(), f); fclose(f); }
    }
    return RD_OK;
}

// Compile only (no device needed): returns RD_OK and the cubin size, or RD_E_COMPILE with the log in
// rd_last_error(). Lets build machines without a GPU validate user programs.
extern "C" int rd_shader_compile_check(const char* code, unsigned size, const char* sbt_json, int fmad, size_t* cubinSize)
{
    if (!code) { rd_set_error("rd_shader_compile_check: null argument"); return RD_E_ARG; }
    std::vector<char> cubin; std::string log;
    g_precDiv = getenv("RD_PREC_DIV") != 0;
    int rc = rd_compile_program(code, size, sbt_json ? sbt_json : "", fmad != 0, &cubin, &log, false, 2, getenv("RD_CHECK_STEAL") != 0, getenv("RD_CHECK_BANDED") != 0);
    if (rc) return rc;
    if (cubinSize) *cubinSize = cubin.size();
    rd_set_error("%.15000s", log.c_str());   // ptxas -v statistics of a successful build
    return RD_OK;
}

extern "C" int rd_shader_module_create(rd_platform* p, const char* code, unsigned size, const char* name, rd_module** out)
{
    (void)name;   // ignored like the reference (radiance/src/radiance.cpp:152,177)
    if (!p || !code || !out) { rd_set_error("rd_shader_module_create: null argument"); return RD_E_ARG; }
    RD_CUDA(cudaSetDevice(p->device));
    { int drc = rd_driver_init(); if (drc) return drc; }
    std::vector<char> cubin; std::string log;
    g_precDiv = p->optPrecDiv != 0;
    int rc = rd_compile_program(code, size, p->sbtJson, p->optFmad != 0, &cubin, &log, p->optCountRays != 0, (int)p->optMaxDepth, p->optSteal != 0,
                                p->optStreamImage != 0);
    if (rc) return rc;
    rd_module* m = new rd_module();
    m->ptxInfo = log;
    CUresult cr = cuModuleLoadData(&m->mod, cubin.data());
    if (cr != CUDA_SUCCESS) { const char* s = 0; cuGetErrorString(cr, &s); rd_set_error("cuModuleLoadData: %s", s ? s : "?"); delete m; return RD_E_CUDA; }
    cr = cuModuleGetFunction(&m->raygen, m->mod, "raygen");   // the reference always extracts "raygen" (radiance.cpp:177)
    if (cr != CUDA_SUCCESS) { rd_set_error("program has no __kernel named raygen"); cuModuleUnload(m->mod); delete m; return RD_E_COMPILE; }
    m->stealRaygen = 0;
    if (p->optSteal && cuModuleGetFunction(&m->stealRaygen, m->mod, "rd_raygen_steal") != CUDA_SUCCESS) m->stealRaygen = 0;
    m->bandedRaygen = 0;
    if (p->optStreamImage && cuModuleGetFunction(&m->bandedRaygen, m->mod, "rd_raygen_banded") != CUDA_SUCCESS) m->bandedRaygen = 0;
    // wavefront stages exist when the program defines the hooks (generate_wavefront_kernels)
    m->wfGenerate = m->wfShade = m->wfAccumulate = m->wfResolve = 0; m->wfPayloadSize = 0;
    if (cuModuleGetFunction(&m->wfGenerate, m->mod, "rd_wf_generate") == CUDA_SUCCESS &&
        cuModuleGetFunction(&m->wfShade, m->mod, "rd_wf_shade") == CUDA_SUCCESS &&
        cuModuleGetFunction(&m->wfAccumulate, m->mod, "rd_wf_accumulate") == CUDA_SUCCESS &&
        cuModuleGetFunction(&m->wfResolve, m->mod, "rd_wf_resolve") == CUDA_SUCCESS) {
        // sizeof(struct Payload) and the WAVEFRONT_* ray constants are known to the compiled program only
        CUfunction query = 0; float* dq = 0; float hq[5] = {0, 0, 0, 0, 0};
        if (cuModuleGetFunction(&query, m->mod, "rd_wf_query") == CUDA_SUCCESS && cudaMalloc(&dq, sizeof(hq)) == cudaSuccess) {
            void* qargv[] = {&dq};
            if (cuLaunchKernel(query, 1, 1, 1, 1, 1, 1, 0, (CUstream)p->stream, qargv, 0) == CUDA_SUCCESS &&
                cudaMemcpyAsync(hq, dq, sizeof(hq), cudaMemcpyDeviceToHost, p->stream) == cudaSuccess)
                cudaStreamSynchronize(p->stream);
            cudaFree(dq);
            p->launches += 1;
        }
        m->wfPayloadSize = (unsigned)hq[0];
        for (int i = 0; i < 4; ++i) m->wfRay[i] = hq[1 + i];
        if (!m->wfPayloadSize) m->wfGenerate = 0;
    } else m->wfGenerate = 0;
    if (const char* e = getenv("RD_L1_CARVEOUT"))   // developer knob: preferred shared-memory carve-out of the raygen kernel in percent (0 = all L1)
        cuFuncSetAttribute(m->raygen, CU_FUNC_ATTRIBUTE_PREFERRED_SHARED_MEMORY_CARVEOUT, atoi(e));
    cuFuncGetAttribute(&m->numRegs, CU_FUNC_ATTRIBUTE_NUM_REGS, m->raygen);
    cuFuncGetAttribute(&m->localBytes, CU_FUNC_ATTRIBUTE_LOCAL_SIZE_BYTES, m->raygen);
    *out = m;
    return RD_OK;
}

extern "C" int rd_module_free(rd_platform* p, rd_module* m)
{
    if (!m) return RD_OK;
    if (p) { cudaSetDevice(p->device); if (p->boundModule == m) p->boundModule = 0; }
    if (g_drv.ok) cuModuleUnload(m->mod);
    delete m;
    return RD_OK;
}

extern "C" int rd_module_info(rd_module* m, int out[2])
{
    if (!m || !out) { rd_set_error("rd_module_info: null argument"); return RD_E_ARG; }
    out[0] = m->numRegs; out[1] = m->localBytes;
    return RD_OK;
}

extern "C" int rd_pipeline_bind(rd_platform* p, rd_module* const* modules, unsigned n)
{
    if (!p || !modules || !n || !modules[0]) { rd_set_error("rd_pipeline_bind: pipeline has no module"); return RD_E_ARG; }
    p->boundModule = modules[0];   // radiance/src/radiance.cpp:254: only module 0 is ever launched
    return RD_OK;
}

extern "C" int rd_descriptors_bind(rd_platform* p, const rd_handle* handles, unsigned n)
{
    if (!p || (n && !handles)) { rd_set_error("rd_descriptors_bind: null argument"); return RD_E_ARG; }
    p->boundArgs.clear(); p->boundKinds.clear();
    for (unsigned i = 0; i < n; ++i) {
        RdHandleObj* o = rd_obj(handles[i]);
        if (!o) { rd_set_error("rd_descriptors_bind: descriptor %u is not a handle of this library", i); return RD_E_ARG; }
        p->boundArgs.push_back(o->arg); p->boundKinds.push_back((int)o->kind);
    }
    return RD_OK;
}


// ---- wavefront sequencer ------------------------------------------------------------------------
// RD::TraceRays in wavefront mode: the sample loop and the bounce loop of samples/shader.cl:196-281 as
// stage launches (rd_wavefront.h). The only host read is RTProp (16 bytes) before the first stage.
// Stage timeline (option "wf_timeline"): one event in front of every stage launch of the last wavefront call; read with
// rd_wf_timeline after the frame. Codes: 1 generate, 2 trace, 3 sort, 4 shade, 5 accumulate, 6 resolve, 7 ray re-sort.
static int rd_tl_mark(rd_platform* p, int code)
{
    if (!p->optWfTimeline) return RD_OK;
    if (p->tlUsed == p->tlEvents.size()) {
        cudaEvent_t e; RD_CUDA(cudaEventCreate(&e));
        p->tlEvents.push_back(e); p->tlCodes.push_back(0);
    }
    p->tlCodes[p->tlUsed] = code;
    RD_CUDA(cudaEventRecord(p->tlEvents[p->tlUsed], p->stream));
    p->tlUsed += 1;
    return RD_OK;
}

extern "C" int rd_wf_timeline(rd_platform* p, float* ms, int* codes, unsigned capacity, unsigned* count)
{
    if (!p || !count) { rd_set_error("rd_wf_timeline: null argument"); return RD_E_ARG; }
    RD_CUDA(cudaSetDevice(p->device));
    RD_CUDA(cudaStreamSynchronize(p->stream));
    const unsigned n = p->tlUsed > 0 ? (unsigned)p->tlUsed - 1u : 0u;   // the last mark closes the last stage
    *count = n;
    for (unsigned i = 0; i < n && i < capacity; ++i) {
        float t = 0.0f;
        RD_CUDA(cudaEventElapsedTime(&t, p->tlEvents[i], p->tlEvents[i + 1]));
        if (ms) ms[i] = t;
        if (codes) codes[i] = p->tlCodes[i];
    }
    return RD_OK;
}

static int rd_trace_rays_wavefront(rd_platform* p, unsigned width, unsigned height, unsigned long long first,
                                   unsigned long long count, const unsigned* rtKnown = 0, bool resetTimeline = true)
{
    rd_module* m = p->boundModule;
    p->wfLane = 0;
    if (!m->wfGenerate) {
        rd_set_error("rd_trace_rays: option wavefront is on but the program defines no wavefrontScene / wavefrontPrimary hooks");
        return RD_E_STATE;
    }
    if (p->boundArgs.size() < 3) { rd_set_error("rd_trace_rays: wavefront mode needs RTProp, imageScratch, image as descriptors 0..2"); return RD_E_STATE; }
    const unsigned total = width * height;
    unsigned rt[4] = {0, 0, 0, 0};   // RayTraceProperties: totalSamples, batchSize, depth, debug (core.h:103-109)
    if (rtKnown) memcpy(rt, rtKnown, 16);   // the caller read it already (one read per frame when the frame is rendered in units)
    else {
        RD_CUDA(cudaMemcpyAsync(rt, (const void*)p->boundArgs[0], 16, cudaMemcpyDeviceToHost, p->stream));
        RD_CUDA(cudaStreamSynchronize(p->stream));
    }
    if (resetTimeline) p->tlUsed = 0;
    const unsigned stride = (m->wfPayloadSize + 15u) & ~15u;
    // two lanes (option wf_lanes): the passes of the frame alternate between two queue sets on two streams
    int lanes = (p->optWfLanes >= 2 && !rt[3] && !p->optWfTimeline) ? (int)p->optWfLanes : 1;
    if (lanes > (int)rt[1]) lanes = rt[1] ? (int)rt[1] : 1;   // no more lanes than samples
    // samples in flight per pass: as many as option wf_paths allows (HBM: ~200 B per path)
    unsigned long long inFlight = (unsigned long long)p->optWfPaths / ((unsigned long long)lanes * count);
    if (inFlight < 1) inFlight = 1;
    if (inFlight > rt[1]) inFlight = rt[1] ? rt[1] : 1;
    if (lanes > 1 && inFlight > (rt[1] + (unsigned)lanes - 1u) / (unsigned)lanes) inFlight = (rt[1] + (unsigned)lanes - 1u) / (unsigned)lanes;   // at least one pass per lane
    if (rt[3]) inFlight = 1;                       // debug views stop after one bounce of one sample at a time
    while (inFlight * count > 0x7fffffffull) --inFlight;
    const unsigned paths = (unsigned)(inFlight * count);
    // the scene header is whichever descriptor is a top-level structure
    const void* scene = 0;
    for (size_t i = 0; i < p->boundKinds.size(); ++i) if (p->boundKinds[i] == RD_H_TLAS) scene = (const void*)p->boundArgs[i];
    if (!scene) { rd_set_error("rd_trace_rays: no top-level acceleration structure in the descriptor set"); return RD_E_STATE; }
    CUdeviceptr dinfo; size_t infoBytes = 0;
    RD_CU(cuModuleGetGlobal(&dinfo, &infoBytes, m->mod, "rd_launch_info"));
    const unsigned info[8] = {(unsigned)first, total, 0u, 1u, 0u, 0u, 0u, 0u};
    unsigned infoTiled[8]; memcpy(infoTiled, info, sizeof(info));
    infoTiled[2] = (p->optTile && (width % 8u) == 0 && (height % 4u) == 0) ? width : 0u;
    RD_CU(cuMemcpyHtoDAsync(dinfo, infoTiled, sizeof(infoTiled), (CUstream)p->stream));
    if (lanes > 1) {
        if (!p->wfStreamX[0]) {
            for (int k = 0; k < RD_WF_MAX_LANES - 1; ++k) RD_CUDA(cudaStreamCreateWithFlags(&p->wfStreamX[k], cudaStreamNonBlocking));
            for (int k = 0; k <= RD_WF_MAX_LANES; ++k) RD_CUDA(cudaEventCreateWithFlags(&p->wfEvents[k], cudaEventDisableTiming));
        }
        // the other lanes start behind everything queued on the platform stream so far (descriptor writes, the launch info above)
        RD_CUDA(cudaEventRecord(p->wfEvents[RD_WF_MAX_LANES], p->stream));
        for (int k = 1; k < lanes; ++k) RD_CUDA(cudaStreamWaitEvent(p->wfStreamX[k - 1], p->wfEvents[RD_WF_MAX_LANES], 0));
    }
    int rc = RD_OK;
    RdWfState S; memset(&S, 0, sizeof(S));
    std::vector<uint64_t> args[RD_WF_MAX_LANES];
    std::vector<void*> argv[RD_WF_MAX_LANES];
    for (int lane = 0; lane < lanes; ++lane) {
        p->wfLane = lane;
        rc = rd_wf_reserve(p, paths, stride);
        if (rc) { p->wfLane = 0; return rc; }
        RdWfBuffers& B = rd_wf_buffers(p);
        memset(&S, 0, sizeof(S));
        S.payload = B.payload; S.color = (RdF4*)B.color; S.contrib = (RdF4*)B.contrib;
        for (int q = 0; q < 2; ++q) { S.rayA[q] = (RdF4*)B.rayA[q]; S.rayB[q] = (RdF4*)B.rayB[q]; S.pathOf[q] = B.pathOf[q]; }
        S.hitA = (RdF4*)B.hitA; S.hitB = B.hitB; S.keys = B.keys; S.order = B.order; S.counts = B.counts; S.hist = B.hist;
        S.payloadStride = stride; S.cur = 0; S.first = (unsigned)first; S.count = (unsigned)count; S.total = total;
        S.tileW = infoTiled[2];
        S.frameID = rt[0]; S.samples = 1; S.totalSamples = rt[0]; S.depth = 0; S.maxDepth = rt[2]; S.debug = rt[3]; S.sortOn = p->optWfSort ? 1u : 0u;
        S.ilWorld = p->ilWorld; S.ilRank = p->ilRank; S.ilBandRows = p->ilBandRows ? p->ilBandRows : 1u;
        S.sbtRecordOffset = (int)m->wfRay[0]; S.missIndex = (int)m->wfRay[1]; S.tmin = m->wfRay[2]; S.tmax = m->wfRay[3];
        S.status = p->dstatus; S.capacity = B.capacity;
        rc = rd_wf_upload_state(p, &S);
        if (rc) { p->wfLane = 0; return rc; }
        args[lane] = p->boundArgs;
        args[lane].push_back((uint64_t)(uintptr_t)B.dstate);
        argv[lane].resize(args[lane].size());
        for (size_t i = 0; i < argv[lane].size(); ++i) argv[lane][i] = &args[lane][i];
    }
    const unsigned blocks = (unsigned)((count + 127) / 128);
    // the shade stage is persistent: as many blocks as fit the GPU, each fetching runs of queue entries until the queue is empty
    int shadeResident = 0;
    if (cuOccupancyMaxActiveBlocksPerMultiprocessor(&shadeResident, m->wfShade, 128, 0) != CUDA_SUCCESS || shadeResident < 1) shadeResident = 4;
    const unsigned shadeGrid = (unsigned)p->smCount * (unsigned)shadeResident;
    unsigned pass = 0;
    for (unsigned k = 0; k < rt[1]; k += (unsigned)inFlight, ++pass) {
        const int lane = (int)(pass % (unsigned)lanes);
        p->wfLane = lane;
        CUstream st = (CUstream)rd_wf_stream(p);
        const unsigned samples = rt[1] - k < (unsigned)inFlight ? rt[1] - k : (unsigned)inFlight;
        const unsigned live = samples * (unsigned)count, pathBlocks = (live + 127u) / 128u;
        rc = rd_wf_advance(p, 0u, rt[0] + k, samples); if (rc) break;
        rd_tl_mark(p, 1);
        RD_CU(cuLaunchKernel(m->wfGenerate, pathBlocks, 1, 1, 128, 1, 1, 0, st, argv[lane].data(), 0));
        p->launches += 1;
        for (unsigned d = 0; d < rt[2] && !rc; ++d) {
            if (d > 0 && p->optWfRaySort) { rd_tl_mark(p, 7); rc = rd_wf_reorder_rays(p, scene, live); if (rc) break; }
            rd_tl_mark(p, 2);
            // primary rays are coherent (static batches); bounce rays are not (dynamic fetch), option wf_dynamic: 0 never, 1 bounces, 2 always
            rc = rd_wf_trace(p, scene, p->optWfDynamic == 2 || (p->optWfDynamic == 1 && d > 0)); if (rc) break;
            if (S.sortOn) { rd_tl_mark(p, 3); rc = rd_wf_sort(p, live); if (rc) break; }
            rd_tl_mark(p, 4);
            RD_CU(cuLaunchKernel(m->wfShade, pathBlocks < shadeGrid ? pathBlocks : shadeGrid, 1, 1, 128, 1, 1, 0, st, argv[lane].data(), 0));
            p->launches += 1;
            rc = rd_wf_advance(p, 2u, 0u); if (rc) break;
            if (rt[3]) break;
        }
        if (rc) break;
        rd_tl_mark(p, 5);
        // the running mean is taken in frame order: this pass's accumulate runs behind the previous pass's (the other lane's)
        if (lanes > 1 && pass > 0) RD_CUDA(cudaStreamWaitEvent((cudaStream_t)st, p->wfEvents[(pass - 1u) % (unsigned)lanes], 0));
        RD_CU(cuLaunchKernel(m->wfAccumulate, blocks, 1, 1, 128, 1, 1, 0, st, argv[lane].data(), 0));
        p->launches += 1;
        if (lanes > 1) RD_CUDA(cudaEventRecord(p->wfEvents[lane], (cudaStream_t)st));
    }
    const int lastLane = pass ? (int)((pass - 1u) % (unsigned)lanes) : 0;
    p->wfLane = 0;
    if (rc) return rc;
    if (lanes > 1 && lastLane != 0) RD_CUDA(cudaStreamWaitEvent(p->stream, p->wfEvents[lastLane], 0));   // resolve (and whatever follows) behind the last pass
    rd_tl_mark(p, 6);
    RD_CU(cuLaunchKernel(m->wfResolve, blocks, 1, 1, 128, 1, 1, 0, (CUstream)p->stream, argv[0].data(), 0));
    p->launches += 1;
    rd_tl_mark(p, 0);
    return RD_OK;
}

// Threads per block for kernel f: option "block", lowered to what the kernel's register count allows (a 128-register
// path-tracing raygen cannot run 1024 threads per block; the work-item -> thread mapping does not depend on the block size).
static unsigned rd_block_for(rd_platform* p, CUfunction f)
{
    unsigned block = (unsigned)p->optBlock;
    int maxThreads = 0;
    if (cuFuncGetAttribute(&maxThreads, CU_FUNC_ATTRIBUTE_MAX_THREADS_PER_BLOCK, f) == CUDA_SUCCESS && maxThreads >= 32 &&
        block > (unsigned)maxThreads) block = (unsigned)maxThreads & ~31u;
    return block;
}

// Launches work-items [first, first + count) of the width*height NDRange (count == 0: all of them).
// With the 8x4 tile mapping on, consecutive work-items of a warp cover one 8x4 pixel tile, so a
// contiguous range is a contiguous run of screen tiles (the unit of the multi-GPU partition).
extern "C" int rd_trace_rays_range(rd_platform* p, unsigned width, unsigned height, unsigned long long first,
                                   unsigned long long count)
{
    if (!p) { rd_set_error("null platform"); return RD_E_ARG; }
    if (!p->boundModule) { rd_set_error("rd_trace_rays: no pipeline bound"); return RD_E_STATE; }
    const unsigned long long total = (unsigned long long)width * height;
    if (!total) return RD_OK;
    if (total > 0xffffffffull) { rd_set_error("rd_trace_rays: more than 2^32 work-items"); return RD_E_ARG; }
    if (first > total || count > total - first) { rd_set_error("rd_trace_rays_range: range outside the NDRange"); return RD_E_ARG; }
    if (!count) count = total - first;
    if (!count) return RD_OK;
    RD_CUDA(cudaSetDevice(p->device));
    if (p->copyPending) {   // the previous frame's streamed read-back still owns the image: render after it
        int drc = rd_stream_bands_finish(p);
        if (drc) return drc;
    }
    p->mirrorFresh = 0;
    if (p->optWavefront) return rd_trace_rays_wavefront(p, width, height, first, count);
    rd_module* m = p->boundModule;
    CUdeviceptr dinfo; size_t infoBytes = 0;
    RD_CU(cuModuleGetGlobal(&dinfo, &infoBytes, m->mod, "rd_launch_info"));
    std::vector<void*> argv(p->boundArgs.size());
    for (size_t i = 0; i < argv.size(); ++i) argv[i] = &p->boundArgs[i];
    const unsigned block = rd_block_for(p, m->raygen);
    const unsigned tileW = (p->optTile && (width % 8u) == 0 && (height % 4u) == 0) ? width : 0u;
    const unsigned full = (unsigned)(count / block), tail = (unsigned)(count % block);
    // a frame that is going to be read back into a registered host mirror: bands of tile rows, copied out as they finish
    if (p->mirrorObj && m->bandedRaygen && first == 0 && count == total && !tail && p->ilWorld <= 1u && (height % 4u) == 0 &&
        ((4ull * width) % block) == 0 && (unsigned long long)total * p->mirrorItemBytes <= p->mirrorObj->size) {
        const unsigned tileRows = height / 4u, blocksPerRow = (unsigned)(4ull * width / block);
        // Equal bands of whole tile rows. Bands complete out of order -- one warp grazing a ridge keeps its band open for
        // 0.2 ms -- so the host copies whichever band is ready (rd_stream_bands_drain) and the copy engine is busy from the
        // first band on; what is exposed behind the last ray is one band.
        struct { unsigned* done; unsigned* flags; unsigned blocksPerBand, nBands, epoch, target, targetLast, pad; } a;
        const unsigned rowsPerBand = (tileRows + p->mirrorBands - 1u) / p->mirrorBands;
        const unsigned nBands = (tileRows + rowsPerBand - 1u) / rowsPerBand;
        // Counters only grow: a band of launch number e is complete at e * (its block count), so nothing has to be
        // reset between frames. A new band layout starts over.
        if (p->bandLayoutW != width || p->bandLayoutH != height || p->bandLayoutN != p->mirrorBands || p->bandLayoutBlock != block ||
            p->bandEpoch >= 0x7fffffffu / (tileRows * blocksPerRow + 1u)) {
            RD_CUDA(cudaStreamSynchronize(p->copyStream));
            RD_CUDA(cudaMemsetAsync(p->bandDone, 0, 64 * sizeof(unsigned), p->stream));
            RD_CUDA(cudaStreamSynchronize(p->stream));
            for (unsigned b = 0; b < 64u; ++b) p->bandFlags[b] = 0u;
            p->bandLayoutW = width; p->bandLayoutH = height; p->bandLayoutN = p->mirrorBands; p->bandLayoutBlock = block; p->bandEpoch = 0;
        }
        p->bandEpoch += 1;
        a.done = p->bandDone; a.flags = p->bandFlagsDev; a.blocksPerBand = rowsPerBand * blocksPerRow; a.nBands = nBands; a.epoch = p->bandEpoch;
        a.target = p->bandEpoch * rowsPerBand * blocksPerRow;
        a.targetLast = p->bandEpoch * (tileRows - (nBands - 1u) * rowsPerBand) * blocksPerRow; a.pad = 0u;
        const unsigned info[8] = {0u, (unsigned)total, tileW, 0u, 0u, 0u, 0u, 0u};
        RD_CU(cuMemcpyHtoDAsync(dinfo, info, sizeof(info), (CUstream)p->stream));
        argv.push_back(&a);
        rd_stream_bands_mark_launch();
        RD_CU(cuLaunchKernel(m->bandedRaygen, full, 1, 1, block, 1, 1, 0, (CUstream)p->stream, argv.data(), 0));
        p->launches += 1;
        const size_t rowBytes = (size_t)4u * width * p->mirrorItemBytes;   // one row of tiles = 4 image rows
        p->bandCount = nBands; p->bandCopied = 0; p->bandRows = rowsPerBand; p->bandTileRows = tileRows; p->bandRowBytes = rowBytes;
        p->mirrorValidBytes = (size_t)tileRows * rowBytes; p->mirrorFresh = 1; p->copyPending = 1;
        return RD_OK;
    }
    if (full) {
        const unsigned info[8] = {(unsigned)first, (unsigned)total, tileW, 0u, p->ilWorld, p->ilRank, p->ilBandRows, (unsigned)p->smCount};
        RD_CU(cuMemcpyHtoDAsync(dinfo, info, sizeof(info), (CUstream)p->stream));
        RD_CU(cuLaunchKernel(m->raygen, full, 1, 1, block, 1, 1, 0, (CUstream)p->stream, argv.data(), 0));
        p->launches += 1;
    }
    if (tail) {
        const unsigned info[8] = {(unsigned)first + full * block, (unsigned)total, tileW, 0u, p->ilWorld, p->ilRank, p->ilBandRows, 0u};
        RD_CU(cuMemcpyHtoDAsync(dinfo, info, sizeof(info), (CUstream)p->stream));
        RD_CU(cuLaunchKernel(m->raygen, 1, 1, 1, tail, 1, 1, 0, (CUstream)p->stream, argv.data(), 0));
        p->launches += 1;
    }
    return RD_OK;
}


// ---- multi-GPU chunk stealing -------------------------------------------------------------------
// One array of frame counters lives on the GPU of the exporting rank; the other ranks (one process per GPU)
// map it with CUDA IPC and increment it with system-scope atomics from their persistent raygen blocks.
// Frame f uses counter f % slots (callers pass f = 0, 1, 2, ... , the same on every rank), the owner clears
// the slot half a ring ahead, so frames need no reset
// round trip; the per-frame collective of the caller keeps the ranks within a frame of each other.
#define RD_STEAL_SLOTS 1024u
struct RdStealArgsHost { unsigned* counter; unsigned* owned; unsigned numChunks, chunkItems, first, count, tileW, pad; };

extern "C" int rd_steal_counter_export(rd_platform* p, unsigned char handle[64])
{
    if (!p || !handle) { rd_set_error("rd_steal_counter_export: null argument"); return RD_E_ARG; }
    RD_CUDA(cudaSetDevice(p->device));
    if (!p->stealCounters) {
        RD_CUDA(cudaMalloc(&p->stealCounters, RD_STEAL_SLOTS * sizeof(unsigned)));
        RD_CUDA(cudaMemset(p->stealCounters, 0, RD_STEAL_SLOTS * sizeof(unsigned)));
        p->stealSlots = RD_STEAL_SLOTS; p->stealImported = 0;
    }
    cudaIpcMemHandle_t h;
    static_assert(sizeof(h) == 64, "cudaIpcMemHandle_t is 64 bytes");
    RD_CUDA(cudaIpcGetMemHandle(&h, p->stealCounters));
    memcpy(handle, &h, 64);
    return RD_OK;
}

extern "C" int rd_steal_counter_import(rd_platform* p, const unsigned char handle[64])
{
    if (!p || !handle) { rd_set_error("rd_steal_counter_import: null argument"); return RD_E_ARG; }
    RD_CUDA(cudaSetDevice(p->device));
    cudaIpcMemHandle_t h; memcpy(&h, handle, 64);
    void* ptr = 0;
    RD_CUDA(cudaIpcOpenMemHandle(&ptr, h, cudaIpcMemLazyEnablePeerAccess));
    p->stealCounters = (unsigned*)ptr; p->stealSlots = RD_STEAL_SLOTS; p->stealImported = 1;
    return RD_OK;
}

// Ranks that are host THREADS of one process (one platform per GPU) cannot open their own process's IPC handle; they
// share the owner's counter array directly: peer access over NVLink, the same system-scope atomics.
extern "C" int rd_steal_counter_adopt(rd_platform* p, rd_platform* owner)
{
    if (!p || !owner || p == owner) { rd_set_error("rd_steal_counter_adopt: bad argument"); return RD_E_ARG; }
    if (!owner->stealCounters) { rd_set_error("rd_steal_counter_adopt: the owner has not exported its counters (rd_steal_counter_export)"); return RD_E_STATE; }
    RD_CUDA(cudaSetDevice(p->device));
    if (p->device != owner->device) {
        cudaError_t e = cudaDeviceEnablePeerAccess(owner->device, 0);
        if (e == cudaErrorPeerAccessAlreadyEnabled) cudaGetLastError();
        else if (e != cudaSuccess) { rd_set_error("rd_steal_counter_adopt: no peer access from device %d to %d (%s)", p->device, owner->device, cudaGetErrorString(e)); return RD_E_CUDA; }
    }
    p->stealCounters = owner->stealCounters; p->stealSlots = owner->stealSlots; p->stealImported = 1;
    return RD_OK;
}

// RD::TraceRays with chunk stealing: every rank calls it with the same width, height, frame and chunkItems;
// each work-item of the NDRange is rendered by exactly one rank. Asynchronous on the platform stream.
extern "C" int rd_trace_rays_steal(rd_platform* p, unsigned width, unsigned height, unsigned frame, unsigned chunkItems)
{
    if (!p || !p->boundModule) { rd_set_error("rd_trace_rays_steal: no pipeline bound"); return RD_E_STATE; }
    rd_module* m = p->boundModule;
    if (!m->stealRaygen) { rd_set_error("rd_trace_rays_steal: module was not created under option stealing"); return RD_E_STATE; }
    RD_CUDA(cudaSetDevice(p->device));
    if (p->copyPending) { int drc = rd_stream_bands_finish(p); if (drc) return drc; }
    p->mirrorFresh = 0;
    if (!p->stealCounters) {   // single process: a private counter
        unsigned char tmp[64];
        int rc = rd_steal_counter_export(p, tmp);
        if (rc) return rc;
    }
    const unsigned long long total = (unsigned long long)width * height;
    if (!total || !chunkItems || (chunkItems & 31u) || total > 0xffffffffull) {
        rd_set_error("rd_trace_rays_steal: chunkItems must be a positive multiple of 32"); return RD_E_ARG;
    }
    const unsigned numChunks = (unsigned)((total + chunkItems - 1) / chunkItems);
    if (p->stealOwnedCap < numChunks + 1) {
        if (p->stealOwned) cudaFree(p->stealOwned);
        RD_CUDA(cudaMalloc(&p->stealOwned, ((size_t)numChunks + 1) * sizeof(unsigned)));
        p->stealOwnedCap = numChunks + 1;
    }
    RD_CUDA(cudaMemsetAsync(p->stealOwned, 0, sizeof(unsigned), p->stream));
    if (!p->stealImported)   // the owner clears the slot that comes into use half a ring from now
        RD_CUDA(cudaMemsetAsync(p->stealCounters + ((frame + p->stealSlots / 2) % p->stealSlots), 0, sizeof(unsigned), p->stream));
    CUdeviceptr dinfo; size_t infoBytes = 0;
    RD_CU(cuModuleGetGlobal(&dinfo, &infoBytes, m->mod, "rd_launch_info"));
    const unsigned tileW = (p->optTile && (width % 8u) == 0 && (height % 4u) == 0) ? width : 0u;
    const unsigned info[8] = {0u, (unsigned)total, tileW, 1u, 0u, 0u, 0u, 0u};
    RD_CU(cuMemcpyHtoDAsync(dinfo, info, sizeof(info), (CUstream)p->stream));
    RdStealArgsHost a;
    a.counter = p->stealCounters + (frame % p->stealSlots); a.owned = p->stealOwned;
    a.numChunks = numChunks; a.chunkItems = chunkItems; a.first = 0; a.count = (unsigned)total; a.tileW = tileW; a.pad = 0;
    std::vector<void*> argv(p->boundArgs.size() + 1);
    for (size_t i = 0; i < p->boundArgs.size(); ++i) argv[i] = &p->boundArgs[i];
    argv[p->boundArgs.size()] = &a;
    const unsigned block = rd_block_for(p, m->stealRaygen);
    const unsigned warpsPerBlock = block / 32u;
    unsigned blocks = (unsigned)p->smCount * 8u;
    if (blocks > (numChunks + warpsPerBlock - 1) / warpsPerBlock) blocks = (numChunks + warpsPerBlock - 1) / warpsPerBlock;
    RD_CU(cuLaunchKernel(m->stealRaygen, blocks, 1, 1, block, 1, 1, 0, (CUstream)p->stream, argv.data(), 0));
    p->launches += 1;
    return RD_OK;
}

// chunks this rank rendered in the last rd_trace_rays_steal (synchronises the stream); returns the count and
// copies up to `capacity` chunk ids
extern "C" int rd_steal_owned(rd_platform* p, unsigned* ids, unsigned capacity, unsigned* count)
{
    if (!p || !count) { rd_set_error("rd_steal_owned: null argument"); return RD_E_ARG; }
    *count = 0;
    if (!p->stealOwned) return RD_OK;
    RD_CUDA(cudaSetDevice(p->device));
    RD_CUDA(cudaStreamSynchronize(p->stream));
    RD_CUDA(cudaMemcpy(count, p->stealOwned, sizeof(unsigned), cudaMemcpyDeviceToHost));
    const unsigned n = *count < capacity ? *count : capacity;
    if (ids && n) RD_CUDA(cudaMemcpy(ids, p->stealOwned + 1, (size_t)n * sizeof(unsigned), cudaMemcpyDeviceToHost));
    return rd_check_status(p, "rd_steal_owned");
}

extern "C" int rd_ray_count(rd_platform* p, uint64_t* out, int reset)
{
    if (!p || !out) { rd_set_error("rd_ray_count: null argument"); return RD_E_ARG; }
    RD_CUDA(cudaSetDevice(p->device));
    RD_CUDA(cudaStreamSynchronize(p->stream));
    unsigned long long v = 0;
    RD_CUDA(cudaMemcpy(&v, p->dRayCounter, 8, cudaMemcpyDeviceToHost));
    if (reset) RD_CUDA(cudaMemset(p->dRayCounter, 0, 8));
    *out = v;
    return rd_check_status(p, "rd_ray_count");
}

// Multi-GPU static partition without load imbalance: the frame's 8x4-tile rows are dealt to the ranks in bands
// of `bandTileRows` rows, round robin (band b belongs to rank b % world); one launch renders all bands of
// `rank`. Needs the tile mapping (width % 8 == 0, height % 4 == 0). Asynchronous on the platform stream.
extern "C" int rd_trace_rays_interleaved(rd_platform* p, unsigned width, unsigned height, unsigned rank, unsigned world,
                                         unsigned bandTileRows)
{
    if (!p) { rd_set_error("null platform"); return RD_E_ARG; }
    if (!world || rank >= world || !bandTileRows || (width % 8u) || (height % 4u) || !p->optTile) {
        rd_set_error("rd_trace_rays_interleaved: needs rank < world, bandTileRows > 0, width %% 8 == 0, height %% 4 == 0, option tile");
        return RD_E_ARG;
    }
    const unsigned tileRows = height / 4u;
    unsigned owned = 0;
    for (unsigned b = rank; b * bandTileRows < tileRows; b += world) {
        const unsigned r0 = b * bandTileRows, r1 = r0 + bandTileRows < tileRows ? r0 + bandTileRows : tileRows;
        owned += r1 - r0;
    }
    if (!owned) return RD_OK;
    p->ilWorld = world; p->ilRank = rank; p->ilBandRows = bandTileRows;
    const int rc = rd_trace_rays_range(p, width, height, 0, (unsigned long long)owned * 4u * width);
    p->ilWorld = 0; p->ilRank = 0; p->ilBandRows = 0;
    return rc;
}

// ---- multi-GPU dynamic partition by work units ----------------------------------------------------------------
// The frame is cut into `units` equal runs of 8x4-tile rows; every rank of the job takes the next unit from the job-wide
// counter (the same IPC-shared counter array as chunk stealing: system-scope atomic over NVLink, one per unit) and renders
// it with the bound pipeline in whatever mode is on -- including wavefront mode, whose stage launches the host sequences.
// The NEXT unit is requested before the current one is enqueued, so the round trip (one tiny kernel + an event wait)
// hides behind the rendering of the current unit. ownedMask: bit u = this rank rendered unit u (rd_frame_gather_units).
__global__ void rd_k_take_unit(unsigned* counter, volatile unsigned* slot) { *slot = atomicAdd_system(counter, 1u); }

extern "C" int rd_trace_rays_units(rd_platform* p, unsigned width, unsigned height, unsigned units, unsigned frame,
                                   unsigned long long* ownedMask)
{
    if (!p || !p->boundModule || !ownedMask) { rd_set_error("rd_trace_rays_units: no pipeline bound / null argument"); return RD_E_STATE; }
    if (!units || units > 64u || (width % 8u) || (height % 4u) || !p->optTile) {
        rd_set_error("rd_trace_rays_units: needs 1..64 units, width %% 8 == 0, height %% 4 == 0, option tile"); return RD_E_ARG;
    }
    RD_CUDA(cudaSetDevice(p->device));
    if (!p->stealCounters) { unsigned char tmp[64]; int rc = rd_steal_counter_export(p, tmp); if (rc) return rc; }
    if (!p->unitSlots) {
        void* h = 0; void* d = 0;
        RD_CUDA(cudaHostAlloc(&h, 2 * sizeof(unsigned), cudaHostAllocMapped));
        RD_CUDA(cudaHostGetDevicePointer(&d, h, 0));
        p->unitSlots = (volatile unsigned*)h; p->unitSlotsDev = (unsigned*)d;
        RD_CUDA(cudaEventCreateWithFlags(&p->unitEvent, cudaEventDisableTiming));
    }
    if (!p->stealImported)   // the owner clears the slot that comes into use half a ring from now
        RD_CUDA(cudaMemsetAsync(p->stealCounters + ((frame + p->stealSlots / 2) % p->stealSlots), 0, sizeof(unsigned), p->stream));
    unsigned* counter = p->stealCounters + (frame % p->stealSlots);
    unsigned rt[4] = {0, 0, 0, 0};
    if (p->optWavefront) {
        if (p->boundArgs.size() < 3) { rd_set_error("rd_trace_rays_units: wavefront mode needs RTProp, imageScratch, image as descriptors 0..2"); return RD_E_STATE; }
        RD_CUDA(cudaMemcpyAsync(rt, (const void*)p->boundArgs[0], 16, cudaMemcpyDeviceToHost, p->stream));
    }
    const unsigned tileRows = height / 4u;
    *ownedMask = 0;
    p->tlUsed = 0;
    int which = 0;
    rd_k_take_unit<<<1, 1, 0, p->stream>>>(counter, p->unitSlotsDev + which);
    p->launches += 1;
    RD_CUDA(cudaEventRecord(p->unitEvent, p->stream));
    RD_CUDA(cudaEventSynchronize(p->unitEvent));
    unsigned cur = p->unitSlots[which];
    while (cur < units) {
        which ^= 1;
        rd_k_take_unit<<<1, 1, 0, p->stream>>>(counter, p->unitSlotsDev + which);   // ahead of the current unit's work in stream order
        p->launches += 1;
        RD_CUDA(cudaEventRecord(p->unitEvent, p->stream));
        const unsigned r0 = (unsigned)((unsigned long long)tileRows * cur / units), r1 = (unsigned)((unsigned long long)tileRows * (cur + 1u) / units);
        if (r1 > r0) {
            const unsigned long long first = (unsigned long long)r0 * 4u * width, count = (unsigned long long)(r1 - r0) * 4u * width;
            int rc = p->optWavefront ? rd_trace_rays_wavefront(p, width, height, first, count, rt, false) : rd_trace_rays_range(p, width, height, first, count);
            if (rc) return rc;
        }
        *ownedMask |= 1ull << cur;
        RD_CUDA(cudaEventSynchronize(p->unitEvent));   // the next unit's number arrived while this one was being enqueued / rendered
        cur = p->unitSlots[which];
    }
    return RD_OK;
}

// RD_BOUNDS_CHECK builds: non-zero if a bounds check failed on the device since the last call (always 0 in a normal build)
extern "C" int rd_debug_bounds_violations(rd_platform* p, unsigned* count, int* checkedBuild)
{
    if (!p || !count) { rd_set_error("rd_debug_bounds_violations: null argument"); return RD_E_ARG; }
    RD_CUDA(cudaSetDevice(p->device));
    RD_CUDA(cudaStreamSynchronize(p->stream));
    *count = p->hstatus[2]; p->hstatus[2] = 0u;
    if (checkedBuild) *checkedBuild = RD_BOUNDS_CHECK;
    return RD_OK;
}

extern "C" int rd_trace_rays_async(rd_platform* p, unsigned width, unsigned height)
{
    return rd_trace_rays_range(p, width, height, 0, 0);
}

extern "C" int rd_trace_rays(rd_platform* p, unsigned raygenGroupIndex, unsigned missGroupIndex, unsigned hitGroupIndex,
                             unsigned width, unsigned height)
{
    (void)raygenGroupIndex; (void)missGroupIndex; (void)hitGroupIndex;   // ignored by the reference too (radiance.cpp:242-247)
    int rc = rd_trace_rays_async(p, width, height);
    if (rc) return rc;
    if (p->copyPending) { rc = rd_stream_bands_drain(p); if (rc) return rc; }   // start the band copies while waiting
    RD_CUDA(cudaStreamSynchronize(p->stream));   // blocking like clFinish (radiance.cpp:261)
    return rd_check_status(p, "rd_trace_rays");
}
