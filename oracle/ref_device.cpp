// oracle/ref_device.cpp -- TEST INFRASTRUCTURE (never linked into the product library).
//
// The reference's device sources compiled UNMODIFIED as C++ on the host, straight from
// /root/reference (oracle/Makefile passes -I for radiance/shader and samples):
//   namespace ref_trace   : radiance/shader/radiance.cl (+ data.cl, math.cl) with recording
//                           dispatchers -> per-ray hit records through the reference's own
//                           traceRay / intersectTop / intersectBot (radiance.cl:41-275)
//   namespace ref_shader1 : samples/shader.cl (+ radiance.cl, pbr.cl) -> the `raygen` kernel
//                           (shader.cl:175-305) run by an NDRange emulator with local size 1,
//                           exactly as RD::TraceRays launches it (radiance/src/radiance.cpp:250-259)
// The OpenCL-C language gap is closed by the host flavour of cl_compat.h (our own prelude).
// system headers first: cl_compat.h defines OpenCL keywords (global, constant, kernel) as macros
#include <thread>
#include <vector>
#include <cstring>
#include <cstdint>

#include "cl_compat.h"

thread_local size_t rd_host_global_id = 0;
size_t rd_host_global_size = 0;

// ----------------------------------------------------------------------------------------
namespace ref_trace {
#include "radiance.cl"

struct Payload {
    struct HitData hit;
    int wasHit;
    int anyHitCalls;
};
struct SceneData {
    int terminateOnFirstHit;
};

void callHit(int sbtRecordOffset, struct Payload* payload, struct HitData* hitData,
             struct SceneData* sceneData, image2d_array_t imageArray, sampler_t sampler)
{
    payload->hit = *hitData;
    payload->wasHit = 1;
}
void callMiss(int missIndex, struct Payload* payload, struct SceneData* sceneData,
              image2d_array_t imageArray, sampler_t sampler)
{
    payload->wasHit = 0;
}
void callAnyHit(bool* cont, int sbtRecordOffset, struct Payload* payload, struct HitData* hitData,
                struct SceneData* sceneData, image2d_array_t imageArray, sampler_t sampler)
{
    payload->anyHitCalls++;
    if (sceneData->terminateOnFirstHit) *cont = false;
}
}  // namespace ref_trace

// ----------------------------------------------------------------------------------------
// the reference headers carry include guards; the second namespace needs its own copy
#undef DATA_CL
#undef MATH_CL
namespace ref_shader1 {
#include "shader.cl"
}  // namespace ref_shader1

// ----------------------------------------------------------------------------------------
// our own benchmark user programs (radiance-ray-tracing_b200/programs/*.cl) compiled against the
// REFERENCE runtime headers: same program text on both sides of every image comparison
#undef DATA_CL
#undef MATH_CL
#undef PI
namespace ref_primary {
#include "primary.cl"
#include "primary_dispatch.inc"
}  // namespace ref_primary
#undef DATA_CL
#undef MATH_CL
#undef PI
namespace ref_path {
#include "path.cl"
#include "path_dispatch.inc"
}  // namespace ref_path
#undef DATA_CL
#undef MATH_CL
#undef PI
#undef WAVEFRONT_SBT_RECORD_OFFSET
#undef WAVEFRONT_MISS_INDEX
namespace ref_groups {
#include "groups.cl"
#include "groups_dispatch.inc"
}  // namespace ref_groups
#undef DATA_CL
#undef MATH_CL
#undef PI
namespace ref_textured {
#include "textured.cl"
#include "textured_dispatch.inc"
}  // namespace ref_textured

// ----------------------------------------------------------------------------------------
template <class F>
static void parallel_range(size_t n, int nthreads, F f)
{
    if (nthreads <= 1 || n < 64) { f(0, n); return; }
    std::vector<std::thread> th;
    size_t chunk = (n + nthreads - 1) / nthreads;
    for (int t = 0; t < nthreads; ++t) {
        size_t b = (size_t)t * chunk, e = b + chunk < n ? b + chunk : n;
        if (b >= e) break;
        th.emplace_back([=] { f(b, e); });
    }
    for (auto& t : th) t.join();
}

extern "C" {

struct RefHit {
    float t;
    unsigned int primitiveIndex, instanceIndex, instanceCustomIndex, instanceSBTOffset;
    float bary[3];
    float hitPoint[3];      // object space (radiance.cl:94,243)
    unsigned int hit;       // 1 = callHit ran, 0 = callMiss ran
    unsigned int anyHitCalls;
    float transform[16];    // object->world of the hit instance
};

// rays: n x 8 floats {ox,oy,oz,Tmin, dx,dy,dz,Tmax}
void ref_trace_rays(const void* tlasBlob, const float* rays, size_t n, int terminateOnFirstHit,
                    RefHit* out, int nthreads)
{
    parallel_range(n, nthreads, [=](size_t b, size_t e) {
        for (size_t i = b; i < e; ++i) {
            const float* r = rays + 8 * i;
            ref_trace::Payload payload;
            memset(&payload, 0, sizeof(payload));
            ref_trace::SceneData sd;
            sd.terminateOnFirstHit = terminateOnFirstHit;
            rd_float3 o(r[0], r[1], r[2]), d(r[4], r[5], r[6]);
            ref_trace::traceRay((ref_trace::AccelStruct*)tlasBlob, 0, 0, o, d, r[3], r[7],
                                &payload, &sd, nullptr, 0);
            RefHit& h = out[i];
            memset(&h, 0, sizeof(h));
            h.hit = payload.wasHit;
            h.anyHitCalls = payload.anyHitCalls;
            if (payload.wasHit) {
                h.t = payload.hit.distance;
                h.primitiveIndex = payload.hit.primitiveIndex;
                h.instanceIndex = payload.hit.instanceIndex;
                h.instanceCustomIndex = payload.hit.instanceCustomIndex;
                h.instanceSBTOffset = payload.hit.instanceSBTOffset;
                for (int k = 0; k < 3; ++k) {
                    h.bary[k] = payload.hit.barycentric[k];
                    h.hitPoint[k] = payload.hit.hitPoint[k];
                }
                for (int k = 0; k < 16; ++k) h.transform[k] = payload.hit.transform[k];
            }
        }
    });
}

// Known-answer helpers: the reference's scalar building blocks, one call each.
void ref_pcg3d(const unsigned int* in3, size_t n, float* out3)
{
    for (size_t i = 0; i < n; ++i) {
        rd_float3 r = ref_trace::random_pcg3d(rd_uint3(in3[3 * i], in3[3 * i + 1], in3[3 * i + 2]));
        out3[3 * i] = r.x; out3[3 * i + 1] = r.y; out3[3 * i + 2] = r.z;
    }
}
int ref_inverse4x4(const float* m16, float* out16)
{
    rd_float16 m, inv;
    for (int i = 0; i < 16; ++i) { m[i] = m16[i]; inv[i] = 0.0f; }
    bool ok = ref_trace::InverseMat4x4(&m, &inv);
    for (int i = 0; i < 16; ++i) out16[i] = inv[i];
    return ok ? 1 : 0;
}
int ref_intersect_aabb(const float* o3, const float* d3, const float* lo3, const float* hi3)
{
    return ref_trace::intersectAABB(rd_float3(o3[0], o3[1], o3[2]), rd_float3(d3[0], d3[1], d3[2]),
                                    rd_float3(lo3[0], lo3[1], lo3[2]), rd_float3(hi3[0], hi3[1], hi3[2])) ? 1 : 0;
}
// verts9: three vertices; out7: {t, px,py,pz, b0,b1,b2}
int ref_intersect_triangle(const float* o3, const float* d3, const float* verts9, float* out7)
{
    rd_float4 v[3];
    for (int k = 0; k < 3; ++k) v[k] = rd_float4(verts9[3 * k], verts9[3 * k + 1], verts9[3 * k + 2], 0.0f);
    ref_trace::Triangle tri; tri.idx0 = 0; tri.idx1 = 1; tri.idx2 = 2; tri.primID = 0;
    rd_float3 p, b; float t = 0.0f;
    bool ok = ref_trace::intersectTriangle(rd_float3(o3[0], o3[1], o3[2]), rd_float3(d3[0], d3[1], d3[2]),
                                           &tri, v, &p, &t, &b);
    if (ok) { out7[0] = t; out7[1] = p.x; out7[2] = p.y; out7[3] = p.z; out7[4] = b.x; out7[5] = b.y; out7[6] = b.z; }
    return ok ? 1 : 0;
}
// generateRay of samples/shader.cl:111-173 for work-item `index`; cam48 = PhysicalCamera bytes.
void ref_generate_ray(const void* cam48, const unsigned int* rand3, size_t index, float* o3, float* d3)
{
    rd_host_global_id = index;
    rd_float3 o, d;
    ref_shader1::generateRay((const ref_shader1::PhysicalCamera*)cam48,
                             rd_uint3(rand3[0], rand3[1], rand3[2]), &o, &d);
    for (int k = 0; k < 3; ++k) { o3[k] = o[k]; d3[k] = d[k]; }
}

// Known answers of the reference's library functions (radiance/shader/pbr.cl, math.cl) as samples/shader.cl sees them.
// fn: 0 BRDF(L,V,N,metallic,roughness,albedo)            in 14 -> out 3      pbr.cl:147-170
//     1 refract(V,H,eta)                                 in 7  -> out 3      pbr.cl:176-186
//     2 G_pbrt(wo,wi,N,roughness)                        in 10 -> out 1      pbr.cl:75-96
//     3 sampleMicrofacetBRDF_transm(V,N,base,m,r,tr,ior,random) in 16 -> out 6 (direction, nextFactor)  pbr.cl:289-385
//     4 microfacetBRDF(L,V,N,albedo,m,r,tr,ior)          in 16 -> out 3      pbr.cl:268-287
//     5 GetNormalSpace(normal)                           in 3  -> out 16     math.cl:269-298
//     6 sampleMicrofacetBRDF(V,N,base,m,r,random)        in 14 -> out 6      pbr.cl:199-266
// Returns the output stride (floats per item).
int ref_library_kat(int fn, const float* in, size_t n, float* out)
{
    using namespace ref_shader1;
    static const int inS[7] = {14, 7, 10, 16, 16, 3, 14}, outS[7] = {3, 3, 1, 6, 3, 16, 6};
    if (fn < 0 || fn > 6) return 0;
    for (size_t i = 0; i < n; ++i) {
        const float* a = in + i * inS[fn];
        float* o = out + i * outS[fn];
        rd_float3 r, f(0.0f, 0.0f, 0.0f);
        switch (fn) {
        case 0: r = BRDF(rd_float3(a[0], a[1], a[2]), rd_float3(a[3], a[4], a[5]), rd_float3(a[6], a[7], a[8]), a[9], a[10],
                         rd_float3(a[11], a[12], a[13]));
                o[0] = r.x; o[1] = r.y; o[2] = r.z; break;
        case 1: r = refract(rd_float3(a[0], a[1], a[2]), rd_float3(a[3], a[4], a[5]), a[6]);
                o[0] = r.x; o[1] = r.y; o[2] = r.z; break;
        case 2: o[0] = G_pbrt(rd_float3(a[0], a[1], a[2]), rd_float3(a[3], a[4], a[5]), rd_float3(a[6], a[7], a[8]), a[9]); break;
        case 3: r = sampleMicrofacetBRDF_transm(rd_float3(a[0], a[1], a[2]), rd_float3(a[3], a[4], a[5]), rd_float3(a[6], a[7], a[8]),
                                                a[9], a[10], a[11], a[12], rd_float3(a[13], a[14], a[15]), &f);
                o[0] = r.x; o[1] = r.y; o[2] = r.z; o[3] = f.x; o[4] = f.y; o[5] = f.z; break;
        case 4: r = microfacetBRDF(rd_float3(a[0], a[1], a[2]), rd_float3(a[3], a[4], a[5]), rd_float3(a[6], a[7], a[8]),
                                   rd_float3(a[9], a[10], a[11]), a[12], a[13], a[14], a[15]);
                o[0] = r.x; o[1] = r.y; o[2] = r.z; break;
        case 5: { mat4x4 m; for (int k = 0; k < 16; ++k) m[k] = 0.0f;
                  GetNormalSpace(rd_float3(a[0], a[1], a[2]), &m);
                  for (int k = 0; k < 16; ++k) o[k] = m[k]; break; }
        case 6: r = sampleMicrofacetBRDF(rd_float3(a[0], a[1], a[2]), rd_float3(a[3], a[4], a[5]), rd_float3(a[6], a[7], a[8]),
                                         a[9], a[10], rd_float3(a[11], a[12], a[13]), &f);
                o[0] = r.x; o[1] = r.y; o[2] = r.z; o[3] = f.x; o[4] = f.y; o[5] = f.z; break;
        }
    }
    return outS[fn];
}

#define RD_RAYGEN_ARGS(NS) \
    (NS::RayTraceProperties*)args[0], (float*)args[1], (uchar*)args[2], (NS::PhysicalCamera*)args[3], \
    (NS::SceneProperties*)args[4], (NS::MeshInfo*)args[5], (float*)args[6], (uint*)args[7], (float*)args[8], \
    (float*)args[9], (NS::Material*)args[10], (image2d_array_t)args[11], (sampler_t)(uintptr_t)args[12], \
    (NS::AccelStruct*)args[13]

// program: 0 = samples/shader.cl, 1 = programs/primary.cl, 2 = programs/path.cl, 3 = programs/groups.cl, 4 = programs/textured.cl
void ref_raygen(int program, void* const* args, size_t globalSize, size_t first, size_t count, int nthreads)
{
    rd_host_global_size = globalSize;
    parallel_range(count, nthreads, [=](size_t b, size_t e) {
        for (size_t i = first + b; i < first + e; ++i) {
            rd_host_global_id = i;
            if (program == 0) ref_shader1::raygen(RD_RAYGEN_ARGS(ref_shader1));
            else if (program == 1) ref_primary::raygen(RD_RAYGEN_ARGS(ref_primary));
            else if (program == 2) ref_path::raygen(RD_RAYGEN_ARGS(ref_path));
            else if (program == 3) ref_groups::raygen(RD_RAYGEN_ARGS(ref_groups));
            else if (program == 4) ref_textured::raygen(RD_RAYGEN_ARGS(ref_textured));
        }
    });
}

// NDRange emulation of the shader.cl `raygen` kernel: 14 positional arguments in descriptor-set
// order (shader.cl:176-190), NDRange size = globalSize, local size 1; work-items [first, first+count).
void ref_raygen_shader1(void* const* args, size_t globalSize, size_t first, size_t count, int nthreads)
{
    rd_host_global_size = globalSize;
    parallel_range(count, nthreads, [=](size_t b, size_t e) {
        for (size_t i = first + b; i < first + e; ++i) {
            rd_host_global_id = i;
            ref_shader1::raygen(
                (ref_shader1::RayTraceProperties*)args[0], (float*)args[1], (uchar*)args[2],
                (ref_shader1::PhysicalCamera*)args[3], (ref_shader1::SceneProperties*)args[4],
                (ref_shader1::MeshInfo*)args[5], (float*)args[6], (uint*)args[7], (float*)args[8],
                (float*)args[9], (ref_shader1::Material*)args[10], (image2d_array_t)args[11],
                (sampler_t)(uintptr_t)args[12], (ref_shader1::AccelStruct*)args[13]);
        }
    });
}

}  // extern "C"
