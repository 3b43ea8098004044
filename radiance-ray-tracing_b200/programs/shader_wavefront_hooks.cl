// shader_wavefront_hooks.cl -- the two wavefront hooks for the reference's own samples/shader.cl.
// A host appends this text to the unmodified file before RD::CreateShaderModule; with option "wavefront" the runtime then
// runs the sample / bounce loop of shader.cl:196-281 as stage launches (generate, trace, sort by hit group, shade, accumulate,
// resolve) instead of one raygen launch. It states the two things the runtime cannot derive from the kernel body: how
// struct SceneData is filled from the descriptors (shader.cl:215-227) and how a sample's payload is armed (shader.cl:203-213).
// The literal traceRay arguments (1, 3, 0.001f, 1000) and the tone map (aces_approx + pow 0.7) are the defaults.
void wavefrontScene(struct SceneData* sceneData,
    __global struct RayTraceProperties* RTProp, __global float* imageScratch, __global uchar* image,
    __global struct PhysicalCamera* camData, __global struct SceneProperties* scene,
    __global struct MeshInfo* meshInfoData, __global float* vertexData, __global uint* indexData,
    __global float* uvData, __global float* normalData, __global struct Material* materials,
    image2d_array_t imageArray, sampler_t sampler, __global struct AccelStruct* topLevel)
{
    sceneData->camData = camData;
    sceneData->scene = scene;
    sceneData->meshInfoData = meshInfoData;
    sceneData->vertexData = vertexData;
    sceneData->indexData = indexData;
    sceneData->uvData = uvData;
    sceneData->normalData = normalData;
    sceneData->materials = materials;
    sceneData->topLevel = topLevel;
    sceneData->depth = 0;
    sceneData->frameID = 0;
    sceneData->debug = RTProp->debug;
}

void wavefrontPrimary(struct Payload* payload, struct SceneData* sceneData, unsigned int frameID, unsigned int totalSamples)
{
    float3 rayOrigin, rayDirection;
    uint3 randInput = {frameID, totalSamples, (uint)get_global_id(0)};
    generateRay(sceneData->camData, randInput, &rayOrigin, &rayDirection);
    payload->color = 0.0f;
    payload->nextFactor = 1.0f;
    payload->nextRayOrigin = rayOrigin;
    payload->nextRayDirection = rayDirection;
}
