// data.cl -- device-visible records shared with the host (B200 backend).
//
// The reference header of the same name (radiance/shader/data.cl) exposes its single-blob binary
// BVH (AccelStructTop/Bottom, BVHNode, Triangle, Instance and the TO_* pointer-arithmetic macros,
// :4-99). None of that layout exists here: the acceleration structure is the wide quantised BVH of
// csrc/rd_types.h and user programs only ever hold an opaque `struct AccelStruct*` that they pass
// to traceRay (samples/shader.cl:189,235,499). What stays is the host-shared launch record.
#ifndef RD_B200_DATA_CL
#define RD_B200_DATA_CL

struct AccelStruct;   // opaque: points at an RdScene header in HBM

// radiance/src/core.h:103-109
struct RayTraceProperties {
    unsigned int totalSamples;
    unsigned int batchSize;
    unsigned int depth;
    unsigned int debug;
};

typedef float4 Vertex;

#define TYPE_INST 1
#define TYPE_TRIG 2
#define TYPE_TOP_AS 1
#define TYPE_BOT_AS 2

#endif
