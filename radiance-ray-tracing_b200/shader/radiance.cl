// radiance.cl -- device tracing runtime included by Radiance user programs (B200 backend).
//
// Drop-in for the reference header of the same name (radiance/shader/radiance.cl): same HitData
// record (:8-18), same user-supplied dispatcher prototypes (:23-28), same traceRay signature and
// semantics (:254-275). The walk itself is csrc/rd_traverse.h (8-wide quantised BVH, ordered and
// culled, precomputed instance inverse) instead of intersectTop / intersectBot (:41-192).
#ifndef RD_B200_RADIANCE_CL
#define RD_B200_RADIANCE_CL
#include "data.cl"
#include "math.cl"

struct HitData {
    float3 hitPoint;                    // object space (instance-local), as in the reference
    float distance;                     // ray parameter t (shared by world and object space)
    unsigned int primitiveIndex;        // triangle index inside the mesh
    unsigned int instanceIndex;         // position in the caller's instance vector
    unsigned int instanceCustomIndex;
    unsigned int instanceSBTOffset;
    float3 barycentric;                 // (1-b1-b2, b1, b2) for vertices (idx0, idx1, idx2)
    mat4x4 transform;                   // object -> world
};

// HitData of a committed candidate: the instance-local ray is re-derived with the same rounded
// operations the traversal used, so hitPoint = o_local + d_local * t matches it bit for bit.
RD_F void rd_fill_hit_data(const RdScene* scene, float3 o, float3 d, const RdHitRec& h, struct HitData* out)
{
    const RdInstance* I = scene->instances + h.inst;
    const float lox = rd_row_dot(I->inv + 0, o.x, o.y, o.z, 1.0f), loy = rd_row_dot(I->inv + 4, o.x, o.y, o.z, 1.0f),
                loz = rd_row_dot(I->inv + 8, o.x, o.y, o.z, 1.0f);
    const float ldx = rd_row_dot(I->inv + 0, d.x, d.y, d.z, 0.0f), ldy = rd_row_dot(I->inv + 4, d.x, d.y, d.z, 0.0f),
                ldz = rd_row_dot(I->inv + 8, d.x, d.y, d.z, 0.0f);
    out->hitPoint = float3(RD_ADD(lox, RD_MUL(ldx, h.t)), RD_ADD(loy, RD_MUL(ldy, h.t)), RD_ADD(loz, RD_MUL(ldz, h.t)));
    out->distance = h.t;
    out->primitiveIndex = h.prim;
    out->instanceIndex = I->instanceID;
    out->instanceCustomIndex = I->customID;
    out->instanceSBTOffset = I->sbtOffset;
    out->barycentric = float3(RD_SUB(RD_SUB(1.0f, h.b1), h.b2), h.b1, h.b2);
    for (int i = 0; i < 16; ++i) out->transform.s[i] = I->fwd[i];
}

#endif  // shared part

// ---- per-level part ------------------------------------------------------------------------------
// The reference's traceRay is reentrant: hit shaders call it again for shadow rays
// (samples/shader.cl:499-501), which makes traceRay -> callHit -> material -> traceRay a cycle in the
// call graph. A cycle forces real calls, and lanes that reach a real call at different times never
// reconverge (measured: 2.2 of 32 lanes active, profiles/README_r1.md). csrc/rd_program.cu therefore
// compiles the user program once per recursion level, each copy in its own namespace rd_l<k>: the
// raygen kernel is taken from level 0, traceRay of level k dispatches into the user functions of
// level k+1 (RD_NEXT_NS), and the last level's traceRay only raises the depth-exceeded flag. Every
// copy inlines completely. Payload / SceneData are per-level types with identical layout.
#if defined(RD_LEVEL) && !defined(RD_LEVEL_PART_DONE)
#define RD_LEVEL_PART_DONE

struct Payload;
struct SceneData;

// supplied by the user program (hand-written or generated from sbt.json)
RD_DEV void callHit(int sbtRecordOffset, struct Payload* payload, struct HitData* hitData,
                    struct SceneData* sceneData, image2d_array_t imageArray, sampler_t sampler);
RD_DEV void callMiss(int missIndex, struct Payload* payload,
                     struct SceneData* sceneData, image2d_array_t imageArray, sampler_t sampler);
RD_DEV void callAnyHit(bool* cont, int sbtRecordOffset, struct Payload* payload, struct HitData* hitData,
                       struct SceneData* sceneData, image2d_array_t imageArray, sampler_t sampler);

#ifdef RD_LAST_LEVEL
RD_F void traceRay(
    __global struct AccelStruct* topLevel,
    int sbtRecordOffset, int missIndex,
    float3 origin, float3 direction,
    float Tmin, float Tmax,
    struct Payload* payload, struct SceneData* sceneData,
    image2d_array_t imageArray, sampler_t sampler)
{
    // deeper than option "max_trace_depth": nothing is traced, rd_trace_rays reports RD_E_STATE
    const RdScene* scene = (const RdScene*)topLevel;
    if (scene->status) scene->status[1] = 1u;
}
#else
struct RdUserAnyHit {
    const RdScene* scene; float3 o, d; int sbtRecordOffset;
    struct Payload* payload; struct SceneData* sceneData; image2d_array_t imageArray; sampler_t sampler;
    bool cont;
    RD_F bool operator()(const RdHitRec& h)
    {
#ifdef RD_HAS_ANYHIT
        struct HitData hd;
        rd_fill_hit_data(scene, o, d, h, &hd);
        RD_NEXT_NS::callAnyHit(&cont, sbtRecordOffset, (RD_NEXT_NS::Payload*)payload, &hd,
                               (RD_NEXT_NS::SceneData*)sceneData, imageArray, sampler);
#endif
        return cont;
    }
    RD_F void countNode() const {}
    RD_F void countTri() const {}
};

RD_F void traceRay(
    __global struct AccelStruct* topLevel,
    int sbtRecordOffset, int missIndex,
    float3 origin, float3 direction,
    float Tmin, float Tmax,
    struct Payload* payload, struct SceneData* sceneData,
    image2d_array_t imageArray, sampler_t sampler)
{
    const RdScene* scene = (const RdScene*)topLevel;
#ifdef RD_COUNT_RAYS
    // metric pass only (option "count_rays"): every traceRay invocation is one ray of the Mrays/s figure
    atomicAdd((unsigned long long*)scene->rayCounter, 1ull);
#endif
    RdUserAnyHit cb;
    cb.scene = scene; cb.o = origin; cb.d = direction; cb.sbtRecordOffset = sbtRecordOffset;
    cb.payload = payload; cb.sceneData = sceneData; cb.imageArray = imageArray; cb.sampler = sampler; cb.cont = true;
    RdHitRec h;
    if (rd_traverse(scene, origin.x, origin.y, origin.z, direction.x, direction.y, direction.z, Tmin, Tmax, h, cb)) {
        struct HitData hd;
        rd_fill_hit_data(scene, origin, direction, h, &hd);
        RD_NEXT_NS::callHit(sbtRecordOffset, (RD_NEXT_NS::Payload*)payload, &hd, (RD_NEXT_NS::SceneData*)sceneData,
                            imageArray, sampler);
    } else {
        RD_NEXT_NS::callMiss(missIndex, (RD_NEXT_NS::Payload*)payload, (RD_NEXT_NS::SceneData*)sceneData,
                             imageArray, sampler);
    }
}
#endif  // RD_LAST_LEVEL
#endif  // per-level part
