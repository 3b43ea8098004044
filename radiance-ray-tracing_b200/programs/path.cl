// path.cl -- user program for BASELINE configs C3 / C5: diffuse + microfacet path tracing with
// any-hit shadow rays, written against the Radiance plugin conventions (samples/shader.cl,
// samples/shader2.cl): the 14-descriptor raygen argument list of samples/shader.cl:176-190, the
// sample loop / bounce loop / running-mean accumulation shape of shader2.cl:57-159, a closest-hit
// shader that traces a nested shadow ray (shader.cl:499-501) and samples the next direction with
// pbr.cl, any-hit termination for shadow rays, dispatchers generated from path_sbt.json.
// sbt rows: 0 raygen, 1 material, 2 shadow + anyShadow, 3 environment (miss), 4 shadowMiss (miss).
// The same file is compiled by NVRTC for the B200 and by g++ against the reference's own
// radiance.cl / pbr.cl for the CPU baseline and the image oracle (oracle/ref_device.cpp).
#include "radiance.cl"
#include "pbr.cl"

struct Payload {
    float3 color;
    float3 nextFactor;
    float3 nextRayOrigin;
    float3 nextRayDirection;
    bool hit;
};

struct PhysicalCamera {
    float widthPixel, heightPixel, focalLength, sensorWidth, focalDistance, fStop;
    float x, y, z, wx, wy, wz;
};

struct SceneData {
    __global struct PhysicalCamera* camData;
    __global struct SceneProperties* scene;
    __global struct MeshInfo* meshInfoData;
    __global float* vertexData;
    __global uint* indexData;
    __global float* normalData;
    __global struct Material* materials;
    __global struct AccelStruct* topLevel;
    unsigned int depth;
    unsigned int frameID;
};


// pinhole camera of samples/shader.cl:111-153 (jittered pixel, sensor point, Rx Ry Rz rotation)
void cameraRay(__global const struct PhysicalCamera* cam, uint3 seed, float3* origin, float3* direction)
{
    const int index = get_global_id(0);
    const int w = (int)cam->widthPixel;
    const float3 rnd = random_pcg3d(seed);
    const float fx = (((float)(index % w) + rnd.x) / cam->widthPixel) - 0.5f;
    const float fy = 0.5f - (((float)(index / w) + rnd.y) / cam->heightPixel);
    float4 d = {fx * cam->sensorWidth, fy * cam->sensorWidth * (cam->heightPixel / cam->widthPixel), -cam->focalLength, 0.0f};
    d = normalize(d);
    mat4x4 rx, ry, rz;
    float4 t0, t1;
    EulerXToMat4x4(cam->wx, &rx);
    EulerYToMat4x4(cam->wy, &ry);
    EulerZToMat4x4(cam->wz, &rz);
    MultiplyMat4Vec4(&rz, &d, &t0);
    MultiplyMat4Vec4(&ry, &t0, &t1);
    MultiplyMat4Vec4(&rx, &t1, &t0);
    t0 = normalize(t0);
    float3 o = {cam->x, cam->y, cam->z};
    *origin = o;
    *direction = t0.xyz;
}

// ---- wavefront hooks (INTEGRATION.md "Wavefront mode"): the two pieces of raygen that the runtime
// cannot guess, factored out so that the inline kernel below and the wavefront stages share them.
void wavefrontScene(struct SceneData* sceneData,
    __global struct RayTraceProperties* RTProp,
    __global float* imageScratch,
    __global uchar* image,
    __global struct PhysicalCamera* camData,
    __global struct SceneProperties* scene,
    __global struct MeshInfo* meshInfoData,
    __global float* vertexData,
    __global uint* indexData,
    __global float* uvData,
    __global float* normalData,
    __global struct Material* materials,
    image2d_array_t imageArray,
    sampler_t sampler,
    __global struct AccelStruct* topLevel)
{
    sceneData->camData = camData;
    sceneData->scene = scene;
    sceneData->meshInfoData = meshInfoData;
    sceneData->vertexData = vertexData;
    sceneData->indexData = indexData;
    sceneData->normalData = normalData;
    sceneData->materials = materials;
    sceneData->topLevel = topLevel;
    sceneData->depth = 0;
    sceneData->frameID = 0;
}

void wavefrontPrimary(struct Payload* payload, struct SceneData* sceneData, unsigned int frameID, unsigned int totalSamples)
{
    uint3 seed = {frameID, totalSamples, (uint)get_global_id(0)};
    cameraRay(sceneData->camData, seed, &payload->nextRayOrigin, &payload->nextRayDirection);
    payload->color = 0.0f;
    payload->nextFactor = 1.0f;
    payload->hit = false;
}

// Reinhard + gamma (shader2.cl:167-173)
float3 wavefrontTonemap(float3 color)
{
    color = color / (color + 1.0f);
    return pow(color, 0.4545f);
}

__kernel void raygen(
    __global struct RayTraceProperties* RTProp,
    __global float* imageScratch,
    __global uchar* image,
    __global struct PhysicalCamera* camData,
    __global struct SceneProperties* scene,
    __global struct MeshInfo* meshInfoData,
    __global float* vertexData,
    __global uint* indexData,
    __global float* uvData,
    __global float* normalData,
    __global struct Material* materials,
    image2d_array_t imageArray,
    sampler_t sampler,
    __global struct AccelStruct* topLevel)
{
    const int index = get_global_id(0);
    struct SceneData sceneData;
    wavefrontScene(&sceneData, RTProp, imageScratch, image, camData, scene, meshInfoData, vertexData, indexData,
                   uvData, normalData, materials, imageArray, sampler, topLevel);

    unsigned int frameID = RTProp->totalSamples;
    for (unsigned int k = 0; k < RTProp->batchSize; ++k, ++frameID) {
        struct Payload payload;
        wavefrontPrimary(&payload, &sceneData, frameID, RTProp->totalSamples);
        sceneData.depth = 0;
        sceneData.frameID = frameID;

        // bounce loop of samples/shader.cl:228-260 (a primary miss ends the path instead of re-tracing it)
        float3 color = 0.0f;
        float3 contribution = 1.0f;
        while (sceneData.depth < RTProp->depth) {
            traceRay(topLevel, 1, 3, payload.nextRayOrigin, payload.nextRayDirection,
                     0.001f, 1000.0f, &payload, &sceneData, imageArray, sampler);
            if (payload.hit) {
                color += contribution * payload.color;
                contribution *= payload.nextFactor;
            } else {
                if (sceneData.depth == 0) color = payload.color;
                break;
            }
            sceneData.depth++;
            payload.hit = false;
            if (RTProp->debug) break;
        }
        for (int c = 0; c < 3; ++c) {
            const float old = frameID == 0 ? 0.0f : imageScratch[4 * index + c];
            imageScratch[4 * index + c] = (frameID * old + color[c]) / (frameID + 1);
        }
    }
    float3 shown = {imageScratch[4 * index], imageScratch[4 * index + 1], imageScratch[4 * index + 2]};
    if (!RTProp->debug) shown = wavefrontTonemap(shown);
    image[4 * index + 0] = (int)(shown.x * 255);
    image[4 * index + 1] = (int)(shown.y * 255);
    image[4 * index + 2] = (int)(shown.z * 255);
    image[4 * index + 3] = 255;
}

float3 transformDirection(struct HitData* hitData, float3 v)
{
    float4 a = {v.x, v.y, v.z, 0.0f}, b;
    MultiplyMat4Vec4(&hitData->transform, &a, &b);
    return b.xyz;
}

// closest hit (row 1): smooth normal, one directional light behind a shadow ray, next direction
// sampled from the microfacet lobe (samples/shader.cl:482-541 structure)
void material(struct Payload* payload, struct HitData* hitData,
    struct SceneData* sceneData, image2d_array_t imageArray, sampler_t sampler)
{
    payload->hit = true;
    __global struct MeshInfo* info = &sceneData->meshInfoData[hitData->instanceIndex];
    __global uint* idx = sceneData->indexData + info->indexOffset + hitData->primitiveIndex * 3;
    __global float* v = sceneData->vertexData + info->vertexOffset;
    __global float* nd = sceneData->normalData + info->normalOffset;
    float3 p0 = {v[idx[0] * 3], v[idx[0] * 3 + 1], v[idx[0] * 3 + 2]};
    float3 p1 = {v[idx[1] * 3], v[idx[1] * 3 + 1], v[idx[1] * 3 + 2]};
    float3 p2 = {v[idx[2] * 3], v[idx[2] * 3 + 1], v[idx[2] * 3 + 2]};
    float3 n0 = {nd[idx[0] * 3], nd[idx[0] * 3 + 1], nd[idx[0] * 3 + 2]};
    float3 n1 = {nd[idx[1] * 3], nd[idx[1] * 3 + 1], nd[idx[1] * 3 + 2]};
    float3 n2 = {nd[idx[2] * 3], nd[idx[2] * 3 + 1], nd[idx[2] * 3 + 2]};
    const float3 bary = hitData->barycentric;

    float3 V = normalize(-payload->nextRayDirection);
    float3 faceN = normalize(transformDirection(hitData, cross(p1 - p0, p2 - p0)));
    if (dot(faceN, V) < 0.0f) faceN = -faceN;
    float3 N = normalize(transformDirection(hitData, n0 * bary.x + n1 * bary.y + n2 * bary.z));
    if (dot(N, faceN) < 0.0f) N = -N;

    float4 hp = {hitData->hitPoint.x, hitData->hitPoint.y, hitData->hitPoint.z, 1.0f}, hw;
    MultiplyMat4Vec4(&hitData->transform, &hp, &hw);
    float3 hitPos = hw.xyz + faceN * 1e-4f;

    __global struct Material* m = &sceneData->materials[hitData->instanceCustomIndex];
    float3 albedo = m->albedo.rgb;
    float3 L = normalize(-sceneData->scene->lights[0].direction.xyz);

    struct Payload shadowPayload;
    shadowPayload.hit = false;
    traceRay(sceneData->topLevel, 2, 4, hitPos, L, 0.001f, 1000.0f,
             &shadowPayload, sceneData, imageArray, sampler);

    float3 color = albedo * 0.1f;
    if (!shadowPayload.hit) {
        float3 radiance = sceneData->scene->lights[0].color.rgb;
        color += microfacetBRDF(L, V, N, albedo, m->metallic, m->roughness, 0.0f, m->ior) * radiance;
    }
    payload->color = color;

    uint3 randInput = {sceneData->frameID, (uint)get_global_id(0), sceneData->depth};
    float3 random = random_pcg3d(randInput);
    float3 nextFactor = {0.0f, 0.0f, 0.0f};
    float3 nextDir = sampleMicrofacetBRDF(V, N, albedo, m->metallic, m->roughness, random, &nextFactor);
    if (dot(nextDir, faceN) < 0.0f) hitPos = hw.xyz - faceN * 1e-4f;
    payload->nextRayOrigin = hitPos;
    payload->nextRayDirection = nextDir;
    payload->nextFactor = nextFactor;
}

// shadow rays (row 2): the first accepted candidate ends the ray
void shadow(struct Payload* payload, struct HitData* hitData,
    struct SceneData* sceneData, image2d_array_t imageArray, sampler_t sampler)
{
    payload->hit = true;
}

void anyShadow(bool* cont, struct Payload* payload, struct HitData* hitData,
    struct SceneData* sceneData, image2d_array_t imageArray, sampler_t sampler)
{
    payload->hit = true;
    *cont = false;
}

void environment(struct Payload* payload,
    struct SceneData* sceneData, image2d_array_t imageArray, sampler_t sampler)
{
    payload->hit = false;
    const float up = 0.5f * (payload->nextRayDirection.y + 1.0f);
    payload->color.x = 0.25f + 0.25f * (1.0f - up);
    payload->color.y = 0.35f + 0.25f * (1.0f - up);
    payload->color.z = 0.55f + 0.35f * up;
}

void shadowMiss(struct Payload* payload,
    struct SceneData* sceneData, image2d_array_t imageArray, sampler_t sampler)
{
    payload->hit = false;
}
