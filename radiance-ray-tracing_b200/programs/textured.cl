// textured.cl -- primary.cl with a textured closest-hit shader: the albedo comes from the layered RGBA8 image
// array through read_imageui with the scene's sampler (REPEAT + LINEAR), as samples/shader2.cl:255-265 does.
// Written against the Radiance plugin conventions (samples/shader.cl): includes "radiance.cl" /
// "pbr.cl", one __kernel raygen with the 14-descriptor argument list of samples/shader.cl:176-190,
// closest-hit / miss functions with the runtime's call shapes, dispatchers generated from
// primary_sbt.json (row index = case label). The same file is compiled by NVRTC for the B200 and by
// g++ against the reference's own radiance.cl for the CPU baseline (oracle/ref_device.cpp).
#include "radiance.cl"
#include "pbr.cl"

struct Payload {
    float3 color;
    bool hit;
};

struct SceneData {
    __global struct SceneProperties* scene;
    __global struct MeshInfo* meshInfoData;
    __global float* vertexData;
    __global uint* indexData;
    __global float* uvData;
    __global struct Material* materials;
};

struct PhysicalCamera {
    float widthPixel, heightPixel, focalLength, sensorWidth, focalDistance, fStop;
    float x, y, z, wx, wy, wz;
};

// pinhole camera: jittered pixel -> sensor point -> Rx Ry Rz rotation (same camera model and random
// stream as samples/shader.cl:111-153 so images are comparable)
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
    sceneData.scene = scene;
    sceneData.meshInfoData = meshInfoData;
    sceneData.vertexData = vertexData;
    sceneData.indexData = indexData;
    sceneData.uvData = uvData;
    sceneData.materials = materials;

    unsigned int frameID = RTProp->totalSamples;
    for (unsigned int k = 0; k < RTProp->batchSize; ++k, ++frameID) {
        float3 o, d;
        uint3 seed = {frameID, RTProp->totalSamples, (uint)index};
        cameraRay(camData, seed, &o, &d);
        struct Payload payload;
        payload.color = 0.0f;
        payload.hit = false;
        traceRay(topLevel, 1, 2, o, d, 0.001f, 1000.0f, &payload, &sceneData, imageArray, sampler);
        for (int c = 0; c < 3; ++c) {
            const float old = frameID == 0 ? 0.0f : imageScratch[4 * index + c];
            imageScratch[4 * index + c] = (frameID * old + payload.color[c]) / (frameID + 1);
        }
    }
    for (int c = 0; c < 3; ++c) {
        const float v = clamp(imageScratch[4 * index + c], 0.0f, 1.0f);
        image[4 * index + c] = (uchar)(int)(sqrt(v) * 255.0f);
    }
    image[4 * index + 3] = 255;
}

// closest hit: flat geometric normal from the mesh's own vertex / index arrays, one directional light
void surface(struct Payload* payload, struct HitData* hitData,
    struct SceneData* sceneData, image2d_array_t imageArray, sampler_t sampler)
{
    __global struct MeshInfo* info = &sceneData->meshInfoData[hitData->instanceIndex];
    __global uint* idx = sceneData->indexData + info->indexOffset + hitData->primitiveIndex * 3;
    __global float* v = sceneData->vertexData + info->vertexOffset;
    float3 a = {v[idx[0] * 3], v[idx[0] * 3 + 1], v[idx[0] * 3 + 2]};
    float3 b = {v[idx[1] * 3], v[idx[1] * 3 + 1], v[idx[1] * 3 + 2]};
    float3 c = {v[idx[2] * 3], v[idx[2] * 3 + 1], v[idx[2] * 3 + 2]};
    float4 n = {0.0f, 0.0f, 0.0f, 0.0f}, nw;
    n.xyz = cross(b - a, c - a);
    MultiplyMat4Vec4(&hitData->transform, &n, &nw);
    float3 N = normalize(nw.xyz);
    float3 L = normalize(-sceneData->scene->lights[0].direction.xyz);
    __global struct Material* m = &sceneData->materials[info->materialIndex];
    float3 albedo = m->albedo.rgb;
    if (m->albedoTexIdx >= 0) {
        // uv has stride 3 in the scene bundle (tools/sceneBuilder.cpp:73-79)
        __global float* uv = sceneData->uvData + info->uvOffset;
        const float3 w = hitData->barycentric;
        const float tu = w.x * uv[idx[0] * 3] + w.y * uv[idx[1] * 3] + w.z * uv[idx[2] * 3];
        const float tv = w.x * uv[idx[0] * 3 + 1] + w.y * uv[idx[1] * 3 + 1] + w.z * uv[idx[2] * 3 + 1];
        float4 coord = {tu, 1.0f - tv, (float)m->albedoTexIdx, 0.0f};
        uint4 texel = read_imageui(imageArray, sampler, coord);
        albedo.x = texel.x / 255.0f;
        albedo.y = texel.y / 255.0f;
        albedo.z = texel.z / 255.0f;
    }
    payload->hit = true;
    payload->color = albedo * (0.15f + 0.85f * fabs(dot(N, L)));
}

void sky(struct Payload* payload,
    struct SceneData* sceneData, image2d_array_t imageArray, sampler_t sampler)
{
    payload->hit = false;
    payload->color.x = 0.2f;
    payload->color.y = 0.2f;
    payload->color.z = 0.5f;
}
