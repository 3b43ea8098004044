// tests/kat/kat.cl -- device-side known-answer harness (TEST INFRASTRUCTURE). A user program in the reference's
// conventions whose raygen evaluates ONE library function per work-item on the inputs of tests/golden/*.npz, so the
// device flavour of shader/math.cl, shader/pbr.cl and csrc/rd_traverse.h (the code user programs really run on sm_100a)
// is compared with the reference's functions (radiance/shader/math.cl, pbr.cl, radiance.cl:211-251).
#include "radiance.cl"
#include "pbr.cl"

struct Payload { int unused; };
struct SceneData { int unused; };
void callHit(int sbtRecordOffset, struct Payload* payload, struct HitData* hitData,
    struct SceneData* sceneData, image2d_array_t imageArray, sampler_t sampler) {}
void callMiss(int missIndex, struct Payload* payload,
    struct SceneData* sceneData, image2d_array_t imageArray, sampler_t sampler) {}
void callAnyHit(bool* cont, int sbtRecordOffset, struct Payload* payload, struct HitData* hitData,
    struct SceneData* sceneData, image2d_array_t imageArray, sampler_t sampler) {}

// header = {function, item count, input stride, output stride}
__kernel void raygen(__global uint* header, __global float* in, __global float* out)
{
    const uint i = get_global_id(0);
    const uint fn = header[0], n = header[1];
    if (i >= n) return;
    __global float* a = in + (size_t)i * header[2];
    __global float* o = out + (size_t)i * header[3];
    float3 r, f = {0.0f, 0.0f, 0.0f};
    switch (fn) {
    case 0: {   // random_pcg3d: inputs are the raw uint bits
        __global uint* u = (__global uint*)a;
        uint3 v = {u[0], u[1], u[2]};
        r = random_pcg3d(v);
        o[0] = r.x; o[1] = r.y; o[2] = r.z;
        break; }
    case 1: {   // InverseMat4x4
        mat4x4 m, inv;
        for (int k = 0; k < 16; ++k) { m.s[k] = a[k]; inv.s[k] = 0.0f; }
        const bool ok = InverseMat4x4(&m, &inv);
        for (int k = 0; k < 16; ++k) o[k] = inv.s[k];
        o[16] = ok ? 1.0f : 0.0f;
        break; }
    case 2: {   // the traversal's triangle test: {o, d, v0, v1, v2} -> {hit, t, b0, b1, b2}
        float t = 0.0f, b1 = 0.0f, b2 = 0.0f;
        const bool ok = rd_intersect_triangle(a[0], a[1], a[2], a[3], a[4], a[5], a[6], a[7], a[8], a[9], a[10], a[11],
                                              a[12], a[13], a[14], &t, &b1, &b2);
        o[0] = ok ? 1.0f : 0.0f; o[1] = t; o[2] = RD_SUB(RD_SUB(1.0f, b1), b2); o[3] = b1; o[4] = b2;
        break; }
    case 10: r = BRDF(float3(a[0], a[1], a[2]), float3(a[3], a[4], a[5]), float3(a[6], a[7], a[8]), a[9], a[10], float3(a[11], a[12], a[13]));
             o[0] = r.x; o[1] = r.y; o[2] = r.z; break;
    case 11: r = refract(float3(a[0], a[1], a[2]), float3(a[3], a[4], a[5]), a[6]);
             o[0] = r.x; o[1] = r.y; o[2] = r.z; break;
    case 12: o[0] = G_pbrt(float3(a[0], a[1], a[2]), float3(a[3], a[4], a[5]), float3(a[6], a[7], a[8]), a[9]); break;
    case 13: r = sampleMicrofacetBRDF_transm(float3(a[0], a[1], a[2]), float3(a[3], a[4], a[5]), float3(a[6], a[7], a[8]),
                                             a[9], a[10], a[11], a[12], float3(a[13], a[14], a[15]), &f);
             o[0] = r.x; o[1] = r.y; o[2] = r.z; o[3] = f.x; o[4] = f.y; o[5] = f.z; break;
    case 14: r = microfacetBRDF(float3(a[0], a[1], a[2]), float3(a[3], a[4], a[5]), float3(a[6], a[7], a[8]),
                                float3(a[9], a[10], a[11]), a[12], a[13], a[14], a[15]);
             o[0] = r.x; o[1] = r.y; o[2] = r.z; break;
    case 15: { mat4x4 m; for (int k = 0; k < 16; ++k) m.s[k] = 0.0f;
               GetNormalSpace(float3(a[0], a[1], a[2]), &m);
               for (int k = 0; k < 16; ++k) o[k] = m.s[k]; break; }
    case 16: r = sampleMicrofacetBRDF(float3(a[0], a[1], a[2]), float3(a[3], a[4], a[5]), float3(a[6], a[7], a[8]),
                                      a[9], a[10], float3(a[11], a[12], a[13]), &f);
             o[0] = r.x; o[1] = r.y; o[2] = r.z; o[3] = f.x; o[4] = f.y; o[5] = f.z; break;
    }
}
