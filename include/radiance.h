// radiance.h -- header-only C++ shim that re-creates `namespace RD` of the reference
// (radiance/include/radiance.h:9-176, host structs of radiance/src/core.h:11-158) on top of the C ABI of
// libradiance_b200.so (include/radiance_b200.h). Code written against the reference's host API
// (samples/sample1.cpp:363-443) recompiles against this header unchanged, with these differences:
//   * handles are opaque pointers of this library instead of cl_mem / cl_sampler / cl_kernel
//   * RD::Vec3 / RD::Mat4x4 are plain structs with assimp's member names (x y z, a1..d4 row-major)
//     instead of aiVector3f / aiMatrix4x4 (no assimp dependency on the hot path)
//   * every failure prints the library's message and exits(-1): the behaviour of CL_CHECK
//     (radiance/src/clcontext.h:27-47); nothing ever falls back to a CPU path
// Additions are marked "extension".
#pragma once
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <cmath>
#include <ctime>
#include <string>
#include <vector>
#include "radiance_b200.h"

// the OpenCL constants the reference's RD_ADDRESS_* / RD_FILTER_* macros expand to (CL/cl.h)
#ifndef CL_ADDRESS_CLAMP_TO_EDGE
#define CL_ADDRESS_CLAMP_TO_EDGE 0x1131
#define CL_ADDRESS_CLAMP 0x1132
#define CL_ADDRESS_REPEAT 0x1133
#define CL_ADDRESS_MIRRORED_REPEAT 0x1134
#define CL_FILTER_NEAREST 0x1140
#define CL_FILTER_LINEAR 0x1141
#endif

// The reference's radiance.h hands assimp's math types to every host through core.h (Vec3 = aiVector3D, Mat4x4 = aiMatrix4x4), and
// samples/sample1.cpp:74-81 builds its light direction with two more of them. Stand-ins for exactly what that host uses, so that
// the reference's own sample1.cpp compiles against this header unchanged (tests/test_cpp_shim.py); a host that includes assimp
// itself keeps assimp's types.
#ifndef AI_VECTOR3D_H_INC
struct aiVector3f
{
    float x = 0, y = 0, z = 0;
    aiVector3f() = default;
    aiVector3f(float X, float Y, float Z) : x(X), y(Y), z(Z) {}
    float operator[](unsigned i) const { return (&x)[i]; }
    float& operator[](unsigned i) { return (&x)[i]; }
};
struct aiMatrix3x3   // row major, identity by default
{
    float a1 = 1, a2 = 0, a3 = 0, b1 = 0, b2 = 1, b3 = 0, c1 = 0, c2 = 0, c3 = 1;
    // rotation by `a` radians about the unit axis (x, y, z), written into `out`
    static aiMatrix3x3& Rotation(float a, const aiVector3f& axis, aiMatrix3x3& out)
    {
        const float c = std::cos(a), s = std::sin(a), t = 1.0f - c, x = axis.x, y = axis.y, z = axis.z;
        out.a1 = t * x * x + c;     out.a2 = t * x * y - s * z; out.a3 = t * x * z + s * y;
        out.b1 = t * x * y + s * z; out.b2 = t * y * y + c;     out.b3 = t * y * z - s * x;
        out.c1 = t * x * z - s * y; out.c2 = t * y * z + s * x; out.c3 = t * z * z + c;
        return out;
    }
    aiMatrix3x3 operator*(const aiMatrix3x3& m) const
    {
        aiMatrix3x3 r;
        r.a1 = a1 * m.a1 + a2 * m.b1 + a3 * m.c1; r.a2 = a1 * m.a2 + a2 * m.b2 + a3 * m.c2; r.a3 = a1 * m.a3 + a2 * m.b3 + a3 * m.c3;
        r.b1 = b1 * m.a1 + b2 * m.b1 + b3 * m.c1; r.b2 = b1 * m.a2 + b2 * m.b2 + b3 * m.c2; r.b3 = b1 * m.a3 + b2 * m.b3 + b3 * m.c3;
        r.c1 = c1 * m.a1 + c2 * m.b1 + c3 * m.c1; r.c2 = c1 * m.a2 + c2 * m.b2 + c3 * m.c2; r.c3 = c1 * m.a3 + c2 * m.b3 + c3 * m.c3;
        return r;
    }
};
inline aiVector3f operator*(const aiMatrix3x3& m, const aiVector3f& v)
{
    return aiVector3f(m.a1 * v.x + m.a2 * v.y + m.a3 * v.z, m.b1 * v.x + m.b2 * v.y + m.b3 * v.z, m.c1 * v.x + m.c2 * v.y + m.c3 * v.z);
}
#endif

namespace RD
{

// ---- radiance/src/core.h -------------------------------------------------------------------------
struct Vec2 { float x = 0, y = 0; };
struct Vec3 { float x = 0, y = 0, z = 0; Vec3() = default; Vec3(float X, float Y, float Z) : x(X), y(Y), z(Z) {} };
struct Mat4x4   // row-major, identity by default like aiMatrix4x4
{
    float a1 = 1, a2 = 0, a3 = 0, a4 = 0, b1 = 0, b2 = 1, b3 = 0, b4 = 0, c1 = 0, c2 = 0, c3 = 1, c4 = 0, d1 = 0, d2 = 0, d3 = 0, d4 = 1;
    float* operator[](unsigned row) { return &a1 + 4 * row; }
    const float* operator[](unsigned row) const { return &a1 + 4 * row; }
};
struct Triangle { unsigned int idx0, idx1, idx2; };
struct RayTraceProperties { unsigned int totalSamples, batchSize, depth, debug; };
struct Material
{
    float albedo[4];
    float metallic, roughness, transmission, ior;
    int albedoTexIdx, metallicTexIdx, roughnessTexIdx, normalTexIdx;   // -1 := not used
};
struct MeshInfo { int vertexOffset, indexOffset, uvOffset, normalOffset, materialIndex, _0, _1, _2; };
struct DirLight { float direction[4]; float color[4]; };
struct SceneProperties { unsigned int lightCount[4]; DirLight lights[5]; };
struct PhysicalCamera
{
    float widthPixel, heightPixel, focalLength, sensorWidth, focalDistance, fStop;
    float x, y, z, wx, wy, wz;
};
static_assert(sizeof(Material) == 48 && sizeof(MeshInfo) == 32 && sizeof(SceneProperties) == 176 && sizeof(PhysicalCamera) == 48 &&
              sizeof(Mat4x4) == 64 && sizeof(Vec3) == 12 && sizeof(Triangle) == 12, "host structs must keep the reference layout");

// ---- radiance/include/radiance.h -----------------------------------------------------------------
typedef void* Handle;
typedef rd_handle TopAccelStruct;
typedef rd_handle Image;
typedef rd_handle ImageArray;
typedef rd_handle Sampler;
typedef unsigned int Uniform;
typedef rd_handle Buffer;

enum DescriptorType { ACCEL_STRUCT_TYPE, IMAGE_TYPE, IMAGE_ARRAY_TYPE, IMAGE_SAMPLER_TYPE, BUFFER_TYPE, TEX_ARRAY_TYPE };

struct Mesh
{
    std::vector<Vec3> vertexData;
    std::vector<Triangle> indexData;
};

typedef rd_blas* BottomAccelStruct;

struct Instance
{
    Mat4x4 transform;   // row major
    unsigned int SBTOffset;
    unsigned int customInstanceID;
    BottomAccelStruct bottomAccelStruct;
};

typedef std::vector<Handle> DescriptorSet;
typedef std::vector<DescriptorType> PipelineLayout;
typedef rd_module* ShaderModule;

#define SHADER_UNUSED (~0U)
struct ShaderGroup { ShaderModule generalShader, closestHitShader, anyHitShader; };

struct PipelineCreateInfo
{
    unsigned int maxRayRecursionDepth;
    PipelineLayout layout;
    std::vector<ShaderModule> modules;
    std::vector<ShaderGroup> groups;
};
typedef PipelineCreateInfo Pipeline;

#define CHANNEL 4
#define RD_CHANNEL CHANNEL

typedef uint32_t AddressingMode;
#define RD_ADDRESS_CLAMP_TO_EDGE CL_ADDRESS_CLAMP_TO_EDGE
#define RD_ADDRESS_CLAMP CL_ADDRESS_CLAMP
#define RD_ADDRESS_REPEAT CL_ADDRESS_REPEAT
#define RD_ADDRESS_MIRRORED_REPEAT CL_ADDRESS_MIRRORED_REPEAT
typedef uint32_t FilterMode;
#define RD_FILTER_NEAREST CL_FILTER_NEAREST
#define RD_FILTER_LINEAR CL_FILTER_LINEAR

struct Platform
{
    static Platform* GetPlatform()
    {
        static Platform ctx;
        if (!ctx.initialized) {
            if (rd_platform_create(-1, &ctx.impl) != RD_OK) {
                fprintf(stderr, "RD::Platform: %s\n", rd_last_error());
                exit(-1);
            }
            ctx.initialized = true;
            printf("Platform initialized.\n");
        }
        return &ctx;
    }
    ~Platform()
    {
        if (initialized) { printf("Platform destroyed.\n"); rd_platform_destroy(impl); }
    }
    Pipeline activePipeline;
    rd_platform* impl = nullptr;   // replaces CLContext* clContext
    bool initialized = false;

private:
    Platform() = default;
    void operator=(Platform&) = delete;
    Platform(Platform&) = delete;
};

// CL_CHECK behaviour (clcontext.h:27-47): report and leave
inline void _check(int rc, const char* what)
{
    if (rc != RD_OK) {
        fprintf(stderr, "RD::%s failed (%d): %s\n", what, rc, rd_last_error());
        exit(-1);
    }
}

// ---- the 22 functions of radiance/include/radiance.h:88-144 -----------------------------------
inline BottomAccelStruct BuildAccelStruct(Platform* platform, Mesh& mesh)   // radiance.cpp:20-52
{
    rd_blas* out = nullptr;
    _check(rd_blas_build(platform->impl, mesh.vertexData.empty() ? nullptr : &mesh.vertexData[0].x, (unsigned)mesh.vertexData.size(),
                         mesh.indexData.empty() ? nullptr : &mesh.indexData[0].idx0, (unsigned)mesh.indexData.size(), &out),
           "BuildAccelStruct(Mesh)");
    return out;
}

inline TopAccelStruct BuildAccelStruct(Platform* platform, std::vector<Instance>& instances)   // radiance.cpp:54-84
{
    std::vector<rd_instance_desc> desc(instances.size());
    for (size_t i = 0; i < instances.size(); ++i) {
        memcpy(desc[i].transform, &instances[i].transform.a1, 64);
        desc[i].SBTOffset = instances[i].SBTOffset;
        desc[i].customInstanceID = instances[i].customInstanceID;
        desc[i].bottomAccelStruct = instances[i].bottomAccelStruct;
    }
    rd_handle out = nullptr;
    _check(rd_tlas_build(platform->impl, desc.data(), (unsigned)desc.size(), &out), "BuildAccelStruct(Instances)");
    return out;
}

inline void TopAccelStructToFile(Platform* platform, TopAccelStruct accelStruct, const char* path)   // radiance.cpp:428-448
{
    _check(rd_tlas_save(platform->impl, accelStruct, path), "TopAccelStructToFile");
}

inline void FileToTopAccelStruct(Platform* platform, const char* path, TopAccelStruct* accelStruct)   // radiance.cpp:450-479
{
    _check(rd_tlas_load(platform->impl, path, accelStruct), "FileToTopAccelStruct");
}

inline Buffer CreateBuffer(Platform* platform, unsigned int size)
{
    rd_handle h = nullptr;
    _check(rd_buffer_create(platform->impl, size, &h), "CreateBuffer");
    return h;
}

inline Image CreateImage(Platform* platform, unsigned int width, unsigned int height)
{
    rd_handle h = nullptr;
    _check(rd_image_create(platform->impl, width, height, &h), "CreateImage");
    return h;
}

inline ImageArray CreateImageArray(Platform* platform, unsigned int width, unsigned int height, unsigned int arraySize)
{
    rd_handle h = nullptr;
    _check(rd_image_array_create(platform->impl, width, height, arraySize, &h), "CreateImageArray");
    return h;
}

inline Sampler CreateSampler(Platform* platform, AddressingMode addressingMode, FilterMode filterMode)
{
    rd_handle h = nullptr;
    _check(rd_sampler_create(platform->impl, addressingMode, filterMode, &h), "CreateSampler");
    return h;
}

inline void ReadImage(Platform* platform, ImageArray handle, unsigned int width, unsigned int height, size_t arrayIndex, void* data)
{
    _check(rd_image_read(platform->impl, handle, width, height, arrayIndex, data), "ReadImage");
}

inline void WriteImage(Platform* platform, ImageArray handle, unsigned int width, unsigned int height, size_t arrayIndex, void* data)
{
    _check(rd_image_write(platform->impl, handle, width, height, arrayIndex, data), "WriteImage");
}

inline void ReadBuffer(Platform* platform, Buffer handle, size_t size, void* data, size_t offset = 0)
{
    _check(rd_buffer_read(platform->impl, handle, size, data, offset), "ReadBuffer");
}

// B200 extension (no reference counterpart): frames traced after this call are copied into `host` band by band while
// they render, so the ReadBuffer(handle, size, host) that follows TraceRays only waits for the last band. Needs
// rd_set_option(platform->impl, "stream_image", 1) before CreateShaderModule; host must be page-locked.
inline void StreamBufferToHost(Platform* platform, Buffer handle, void* host, unsigned int itemBytes = 4, unsigned int bands = 4)
{
    _check(rd_buffer_stream_to_host(platform->impl, handle, host, itemBytes, bands), "StreamBufferToHost");
}

inline void WriteBuffer(Platform* platform, Buffer handle, size_t size, void* data, size_t offset = 0)
{
    _check(rd_buffer_write(platform->impl, handle, size, data, offset), "WriteBuffer");
}

inline DescriptorSet CreateDescriptorSet(std::vector<Handle> handles) { return handles; }                    // radiance.cpp:142-145
inline PipelineLayout CreatePipelineLayout(std::vector<DescriptorType> descriptorTypes) { return descriptorTypes; }   // :147-150

inline ShaderModule CreateShaderModule(Platform* platform, char* code, unsigned int size, char* name)      // radiance.cpp:152-179
{
    rd_module* m = nullptr;
    _check(rd_shader_module_create(platform->impl, code, size, name, &m), "CreateShaderModule");
    return m;
}

inline Pipeline CreatePipeline(PipelineCreateInfo pipelineCreateInfo) { return pipelineCreateInfo; }        // radiance.cpp:181-184

inline void BindPipeline(Platform* platform, Pipeline pipeline)                                             // radiance.cpp:226-229
{
    platform->activePipeline = pipeline;
    _check(rd_pipeline_bind(platform->impl, pipeline.modules.data(), (unsigned)pipeline.modules.size()), "BindPipeline");
}

inline void BindDescriptorSet(Platform* platform, DescriptorSet descriptorSet)                              // radiance.cpp:231-240
{
    _check(rd_descriptors_bind(platform->impl, descriptorSet.data(), (unsigned)descriptorSet.size()), "BindDescriptorSet");
}

inline void TraceRays(Platform* platform, unsigned int raygenGroupIndex, unsigned int missGroupIndex,
                      unsigned int hitGroupIndex, unsigned int width, unsigned int height)                 // radiance.cpp:242-267
{
    _check(rd_trace_rays(platform->impl, raygenGroupIndex, missGroupIndex, hitGroupIndex, width, height), "TraceRays");
}

// clcontext.cpp: whole file into a malloc'ed, NUL-terminated buffer; returns 0 on success
inline int read_kernel_file_str(const char* filename, char** data, size_t* size)
{
    FILE* fp = fopen(filename, "rb");
    if (!fp) { fprintf(stderr, "Could not open kernel file: %s\n", filename); return -1; }
    fseek(fp, 0, SEEK_END);
    const long n = ftell(fp);
    fseek(fp, 0, SEEK_SET);
    *data = (char*)malloc((size_t)n + 1);
    const size_t got = fread(*data, 1, (size_t)n, fp);
    fclose(fp);
    (*data)[got] = '\0';
    *size = got;
    return got == (size_t)n ? 0 : -1;
}

// ---- extensions ------------------------------------------------------------------------------
// samples/sbt.json for programs that do not hand-write callHit / callMiss / callAnyHit (the reference
// runs tools/genSBT.py offline instead); call before CreateShaderModule
inline void LoadShaderBindingTable(Platform* platform, const char* jsonText)
{
    _check(rd_sbt_load_json(platform->impl, jsonText), "LoadShaderBindingTable");
}
// runtime options of include/radiance_b200.h ("wavefront", "wf_sort", "fmad", "tile", ...)
inline void SetOption(Platform* platform, const char* key, long value)
{
    _check(rd_set_option(platform->impl, key, value), "SetOption");
}

} // namespace RD
