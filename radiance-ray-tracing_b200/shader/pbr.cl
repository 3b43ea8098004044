// pbr.cl -- microfacet shading library visible to Radiance user programs (B200 backend).
//
// Same function names, parameters and formulas as the reference header of the same name
// (radiance/shader/pbr.cl: D_GGX :6-13, G1_GGX_Schlick/G_Smith_Disney :16-28, F_Schlick :31-37,
// Lambda/G_pbrt :66-96, G_SchlicksmithGGX :99-106, D_GGX_not_remapped :108-115, Roughness_Remap
// :118-123, G_SmithGGXCorrelated :126-132, Diffuse_Disney :135-143, BRDF :147-169, reflect/refract
// :171-197, sampleMicrofacetBRDF :199-266, microfacetBRDF :268-287, sampleMicrofacetBRDF_transm
// :289-385, Material/MeshInfo/DirLight/SceneProperties :387-425), written from scratch: the three
// hemisphere-sampling branches share rd_sample_* helpers instead of being spelled out per case.
#ifndef RD_B200_PBR_CL
#define RD_B200_PBR_CL
#include "math.cl"

#define PI 3.14159265359f

// ---- host-shared scene records (byte layouts of radiance/src/core.h:111-149) -------------------
struct Material {
    float4 albedo;
    float metallic, roughness, transmission, ior;
    int albedoTexIdx, metallicTexIdx, roughnessTexIdx, normalTexIdx;   // -1 = no texture
};
struct MeshInfo {
    int vertexOffset, indexOffset, uvOffset, normalOffset;              // in scalar elements
    int materialIndex;
    int _0, _1, _2;
};
struct DirLight { float4 direction; float4 color; };
struct SceneProperties { uint4 lightCount; struct DirLight lights[5]; };

// ---- distribution / masking / Fresnel terms ---------------------------------------------------
// GGX (Trowbridge-Reitz) with alpha = roughness^2
RD_F float D_GGX(float dotNH, float roughness)
{
    const float alpha = roughness * roughness;
    const float a2 = alpha * alpha;
    const float d = dotNH * dotNH * (a2 - 1.0f) + 1.0f;
    return a2 / (PI * d * d);
}

RD_F float G1_GGX_Schlick(float NdotV, float roughness)
{
    const float r = 0.5f + 0.5f * roughness;   // Disney remap
    const float k = (r * r) / 2.0f;
    return NdotV / (NdotV * (1.0f - k) + k);
}

RD_F float G_Smith_Disney(float NoL, float NoV, float roughness)
{
    return G1_GGX_Schlick(NoL, roughness) * G1_GGX_Schlick(NoV, roughness);
}

RD_F float3 F_Schlick(float cosTheta, float metallic, float3 albedo)
{
    const float3 F0 = mix(float3(0.04f, 0.04f, 0.04f), albedo, metallic);
    return F0 + (1.0f - F0) * (float)pow(1.0f - cosTheta, 5.0f);
}

// local-frame trigonometry (z = normal)
RD_F float Cos2Theta(float3 w) { return w.z * w.z; }
RD_F float Sin2Theta(float3 w) { return fmax(0.0f, 1.0f - Cos2Theta(w)); }
RD_F float SinTheta(float3 w) { return sqrt(Sin2Theta(w)); }
RD_F float CosPhi(float3 w)
{
    const float s = SinTheta(w);
    return (s == 0.0f) ? 1.0f : clamp(w.x / s, -1.0f, 1.0f);
}
RD_F float SinPhi(float3 w)
{
    const float s = SinTheta(w);
    return (s == 0.0f) ? 0.0f : clamp(w.y / s, -1.0f, 1.0f);
}
RD_F float Tan2Theta(float3 w) { return Sin2Theta(w) / Cos2Theta(w); }

// Smith Lambda for (isotropic use of) the anisotropic GGX form, pbrt-v4 9.6
RD_F float Lambda(float3 w, float a)
{
    const float t2 = Tan2Theta(w);
    if (isinf(t2)) return 0.0f;
    const float ca = CosPhi(w) * a, sa = SinPhi(w) * a;
    const float alpha2 = ca * ca + sa * sa;
    return (sqrt(1.0f + alpha2 * t2) - 1.0f) / 2.0f;
}

// height-correlated masking-shadowing G(wo, wi) evaluated in the frame of N
RD_F float G_pbrt(float3 wo, float3 wi, float3 N, float roughness)
{
    mat4x4 frame, toLocal;
    GetNormalSpace(N, &frame);
    InverseMat4x4(&frame, &toLocal);
    float4 gi = float4(wi, 0.0f), go = float4(wo, 0.0f), li, lo;
    MultiplyMat4Vec4(&toLocal, &go, &lo);
    MultiplyMat4Vec4(&toLocal, &gi, &li);
    if (li.z < 0 || lo.z < 0) return 0.0f;
    return 1 / (1 + Lambda(li.xyz, roughness) + Lambda(lo.xyz, roughness));
}

RD_F float G_SchlicksmithGGX(float dotNL, float dotNV, float roughness)
{
    const float r = roughness + 1.0f;
    const float k = (r * r) / 8.0f;
    return (dotNL / (dotNL * (1.0f - k) + k)) * (dotNV / (dotNV * (1.0f - k) + k));
}

RD_F float D_GGX_not_remapped(float NoH, float roughness)
{
    const float a = NoH * roughness;
    const float k = roughness / (1.0f - NoH * NoH + a * a);
    return k * k * (1.0f / PI);
}

RD_F float Roughness_Remap(float roughness)
{
    roughness = max(0.05f, roughness);
    return roughness * roughness;
}

RD_F float G_SmithGGXCorrelated(float NoL, float NoV, float roughness)
{
    const float a2 = roughness * roughness;
    const float gl = NoV * sqrt((-NoL * a2 + NoL) * NoL + a2);
    const float gv = NoL * sqrt((-NoV * a2 + NoV) * NoV + a2);
    return 0.5f / (gv + gl);
}

// (the reference computes the Disney retro-reflection factors and then returns the Lambert term)
RD_F float3 Diffuse_Disney(float3 c_diff, float roughness, float dotNL, float dotNV, float dotVH)
{
    return (1 / PI) * c_diff;
}

// ---- evaluation -------------------------------------------------------------------------------
RD_F float3 BRDF(float3 L, float3 V, float3 N, float metallic, float roughness, float3 albedo)
{
    roughness = max(0.05f, roughness);
    const float3 H = normalize(V + L);
    const float dotNV = clamp(dot(N, V), 0.0f, 1.0f), dotNL = clamp(dot(N, L), 0.0f, 1.0f);
    const float dotNH = clamp(dot(N, H), 0.0f, 1.0f), dotVH = clamp(dot(V, H), 0.0f, 1.0f);
    const float D = D_GGX(dotNH, roughness);
    const float G = G_pbrt(V, L, N, roughness);
    const float3 F = F_Schlick(dotVH, metallic, albedo);
    const float3 c_diff = albedo * (1.0f - metallic);
    const float3 f_diffuse = (1 - F) * (1 / PI) * c_diff;
    const float3 f_specular = F * D * G / max(4.0f * dotNL * dotNV, 0.001f);
    return (f_diffuse + f_specular) * dotNL;
}

RD_F float3 reflect(float3 in, float3 N) { return -in + 2 * dot(in, N) * N; }

RD_F float3 refract(float3 V, float3 H, float eta)
{
    const float ci = dot(H, V);
    const float s2i = max(0.0f, 1.0f - (ci * ci));
    const float s2t = s2i / (eta * eta);
    if ((1.0f - s2t) < 0.0f) return (ci * H - V) / eta;   // total internal reflection branch
    const float ct = sqrt(1.0f - s2t);
    return -V / eta + (ci / eta - ct) * H;
}

RD_F float3 refract_glsl(float3 I, float3 N, float eta)
{
    const float k = 1.0f - eta * eta * (1.0f - dot(N, I) * dot(N, I));
    if (k < 0.0f) return float3(0.0f);
    return eta * I - (eta * dot(N, I) + sqrt(k)) * N;
}

RD_F float3 microfacetBRDF(float3 L, float3 V, float3 N, float3 albedo,
                           float metallicness, float roughness, float transmission, float ior)
{
    const float3 H = normalize(V + L);
    const float NoV = clamp(dot(N, V), 0.0f, 1.0f), NoL = clamp(dot(N, L), 0.0f, 1.0f);
    const float NoH = clamp(dot(N, H), 0.0f, 1.0f), VoH = clamp(dot(V, H), 0.0f, 1.0f);
    const float3 F = F_Schlick(VoH, metallicness, albedo);
    const float D = D_GGX(NoH, roughness);
    const float G = G_pbrt(V, L, N, roughness);
    const float3 f_specular = (D * G * F) / max(4.0f * NoV * NoL, 0.001f);
    const float3 notSpec = (1.0f - F) * (1.0f - metallicness) * (1.0f - transmission);
    const float3 f_diffuse = notSpec * (albedo / PI);
    return (f_diffuse + f_specular) * NoL;
}

// ---- importance sampling ----------------------------------------------------------------------
// direction with polar angle theta / azimuth 2*pi*u around `axis`
RD_F float3 rd_sample_about(float3 axis, float theta, float u)
{
    const float phi = 2.0f * PI * u;
    float4 local = float4(sin(theta) * cos(phi), sin(theta) * sin(phi), cos(theta), 0.0f), world;
    mat4x4 frame;
    GetNormalSpace(axis, &frame);
    MultiplyMat4Vec4(&frame, &local, &world);
    return world.xyz;
}
// cosine-weighted hemisphere: theta = acos(sqrt(v))
RD_F float3 rd_sample_cosine(float3 N, float3 random) { return rd_sample_about(N, acos(sqrt(random.y)), random.x); }
// GGX half vector: theta = acos(sqrt((1 - v) / (1 + (a^2 - 1) v))), a = roughness^2
RD_F float3 rd_sample_ggx_half(float3 N, float roughness, float3 random)
{
    const float a = roughness * roughness;
    return rd_sample_about(N, acos(sqrt((1.0f - random.y) / (1.0f + (a * a - 1.0f) * random.y))), random.x);
}

// specular lobe shared by both samplers: reflect V about a sampled GGX half vector; weight
// F G VoH / (NoH NoV), doubled for the 50/50 lobe split
RD_F float3 rd_sample_specular(float3 V, float3 N, float3 baseColor, float metallicness, float roughnessSample,
                               float roughnessEval, float3 random, float3* nextFactor)
{
    const float3 H = rd_sample_ggx_half(N, roughnessSample, random);
    const float3 L = reflect(V, H);
    const float NoV = clamp(dot(N, V), 0.0f, 1.0f), NoH = clamp(dot(N, H), 0.0f, 1.0f), VoH = clamp(dot(V, H), 0.0f, 1.0f);
    const float G = G_pbrt(V, L, N, roughnessEval);
    const float3 F = F_Schlick(VoH, metallicness, baseColor);
    *nextFactor = F * G * VoH / max((NoH * NoV), 0.001f);
    *nextFactor *= 2.0f;
    return L;
}

RD_F float3 sampleMicrofacetBRDF(float3 V, float3 N, float3 baseColor, float metallicness, float roughness,
                                 float3 random, float3* nextFactor)
{
    if (random.z > 0.5f) {
        const float3 L = rd_sample_cosine(N, random);
        const float3 H = normalize(V + L);
        const float3 F = F_Schlick(clamp(dot(V, H), 0.0f, 1.0f), metallicness, baseColor);
        const float3 reflectance = baseColor * (1.0f - metallicness);
        *nextFactor = reflectance * (1 - F);
        *nextFactor *= 2.0f;
        return L;
    }
    // the half vector is sampled with the raw roughness, the masking term with the clamped one
    return rd_sample_specular(V, N, baseColor, metallicness, roughness, max(0.05f, roughness), random, nextFactor);
}

RD_F float3 sampleMicrofacetBRDF_transm(float3 V, float3 N, float3 baseColor, float metallicness, float roughness,
                                        float transmission, float ior, float3 random, float3* nextFactor)
{
    if (!(random.z < 0.5f))
        return rd_sample_specular(V, N, baseColor, metallicness, roughness, roughness, random, nextFactor);
    if ((2.0f * random.z) < transmission) {
        // rough refraction through the side of the surface that faces V
        float3 fn = N;
        float eta = ior;
        if (dot(V, N) < 0.0f) { fn = -N; eta = 1.0f / ior; }
        const float3 H = rd_sample_ggx_half(fn, roughness, random);
        const float3 L = refract(V, H, eta);
        const float NoV = clamp(dot(fn, V), 0.0f, 1.0f), NoH = clamp(dot(fn, H), 0.0f, 1.0f), VoH = clamp(dot(V, H), 0.0f, 1.0f);
        const float3 F = F_Schlick(VoH, metallicness, baseColor);
        const float G = G_pbrt(V, -L, fn, roughness);
        *nextFactor = baseColor * (1.0f - F) * G * VoH / max((NoH * NoV), 0.001f);
        *nextFactor *= 2.0f;
        return L;
    }
    const float3 L = rd_sample_cosine(N, random);
    const float3 H = normalize(V + L);
    const float3 F = F_Schlick(clamp(dot(V, H), 0.0f, 1.0f), metallicness, baseColor);
    *nextFactor = (1.0f - F) * (1.0f - metallicness) * baseColor;
    *nextFactor *= 2.0f;
    return L;
}

#endif
