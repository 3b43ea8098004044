// rd_program.cu -- user programs: NVRTC compilation of OpenCL-C shader files, sbt.json dispatch
// generation, descriptor binding and the TraceRays launch.
// Replaces radiance/src/radiance.cpp:142-184 (descriptor set / pipeline objects, CreateShaderModule
// through clBuildProgram with -I SHADER_LIB_PATH), :226-267 (BindPipeline, BindDescriptorSet,
// TraceRays) and tools/genSBT.py (switch generation from sbt.json).
#include <nvrtc.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <string>
#include <vector>
#include "rd_internal.h"
#include "rd_embedded.h"

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
        {"cuMemcpyHtoDAsync", (void**)&g_drv.MemcpyHtoDAsync}, {"cuGetErrorString", (void**)&g_drv.GetErrorString}};
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
                              std::vector<char>* cubin, std::string* log, bool countRays = false, int maxDepth = 2)
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
    if (countRays) opts.push_back("-DRD_COUNT_RAYS=1");
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
    int rc = rd_compile_program(code, size, sbt_json ? sbt_json : "", fmad != 0, &cubin, &log);
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
    int rc = rd_compile_program(code, size, p->sbtJson, p->optFmad != 0, &cubin, &log, p->optCountRays != 0, (int)p->optMaxDepth);
    if (rc) return rc;
    rd_module* m = new rd_module();
    m->ptxInfo = log;
    CUresult cr = cuModuleLoadData(&m->mod, cubin.data());
    if (cr != CUDA_SUCCESS) { const char* s = 0; cuGetErrorString(cr, &s); rd_set_error("cuModuleLoadData: %s", s ? s : "?"); delete m; return RD_E_CUDA; }
    cr = cuModuleGetFunction(&m->raygen, m->mod, "raygen");   // the reference always extracts "raygen" (radiance.cpp:177)
    if (cr != CUDA_SUCCESS) { rd_set_error("program has no __kernel named raygen"); cuModuleUnload(m->mod); delete m; return RD_E_COMPILE; }
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
    p->boundArgs.clear();
    for (unsigned i = 0; i < n; ++i) {
        RdHandleObj* o = rd_obj(handles[i]);
        if (!o) { rd_set_error("rd_descriptors_bind: descriptor %u is not a handle of this library", i); return RD_E_ARG; }
        p->boundArgs.push_back(o->arg);
    }
    return RD_OK;
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
    rd_module* m = p->boundModule;
    CUdeviceptr dinfo; size_t infoBytes = 0;
    RD_CU(cuModuleGetGlobal(&dinfo, &infoBytes, m->mod, "rd_launch_info"));
    std::vector<void*> argv(p->boundArgs.size());
    for (size_t i = 0; i < argv.size(); ++i) argv[i] = &p->boundArgs[i];
    const unsigned block = (unsigned)p->optBlock;
    const unsigned tileW = (p->optTile && (width % 8u) == 0 && (height % 4u) == 0) ? width : 0u;
    const unsigned full = (unsigned)(count / block), tail = (unsigned)(count % block);
    if (full) {
        const unsigned info[4] = {(unsigned)first, (unsigned)total, tileW, 0u};
        RD_CU(cuMemcpyHtoDAsync(dinfo, info, sizeof(info), (CUstream)p->stream));
        RD_CU(cuLaunchKernel(m->raygen, full, 1, 1, block, 1, 1, 0, (CUstream)p->stream, argv.data(), 0));
        p->launches += 1;
    }
    if (tail) {
        const unsigned info[4] = {(unsigned)first + full * block, (unsigned)total, tileW, 0u};
        RD_CU(cuMemcpyHtoDAsync(dinfo, info, sizeof(info), (CUstream)p->stream));
        RD_CU(cuLaunchKernel(m->raygen, 1, 1, 1, tail, 1, 1, 0, (CUstream)p->stream, argv.data(), 0));
        p->launches += 1;
    }
    return RD_OK;
}

extern "C" int rd_ray_count(rd_platform* p, uint64_t* out, int reset)
{
    if (!p || !out) { rd_set_error("rd_ray_count: null argument"); return RD_E_ARG; }
    RD_CUDA(cudaSetDevice(p->device));
    RD_CUDA(cudaStreamSynchronize(p->stream));
    unsigned long long v = 0;
    RD_CUDA(cudaMemcpy(&v, p->dstatus + 2, 8, cudaMemcpyDeviceToHost));
    if (reset) RD_CUDA(cudaMemset(p->dstatus + 2, 0, 8));
    *out = v;
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
    RD_CUDA(cudaStreamSynchronize(p->stream));   // blocking like clFinish (radiance.cpp:261)
    uint32_t status[2] = {0, 0};
    RD_CUDA(cudaMemcpy(status, p->dstatus, 8, cudaMemcpyDeviceToHost));
    if (status[0] || status[1]) {
        cudaMemset(p->dstatus, 0, 8);
        if (status[0]) { rd_set_error("rd_trace_rays: traversal stack overflow on the device (results incomplete)"); return RD_E_OVERFLOW; }
        rd_set_error("rd_trace_rays: traceRay nested deeper than option max_trace_depth (%ld)", p->optMaxDepth);
        return RD_E_STATE;
    }
    return RD_OK;
}
