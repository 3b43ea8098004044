/* rd_oracle.h -- plain-C restatement of the reference's traceRay path. TEST INFRASTRUCTURE ONLY:
 * only tests/, __graft_entry__.smoke() and bench.py's CPU legs may call it; the product never does.
 * Pinned against tests/golden/ (outputs of the reference's own code, see tests/golden/make_golden.py). */
#ifndef RD_ORACLE_H
#define RD_ORACLE_H
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct rdo_hit {
    float t;
    uint32_t prim, inst, custom, sbt;
    float bary[3];
    float point[3];        /* object space */
    uint32_t hit;          /* 1 = something accepted */
    uint32_t anyhit_calls; /* accepted closer candidates seen */
} rdo_hit;

/* radiance/shader/math.cl:10-23 */
void rdo_pcg3d(const uint32_t in[3], float out[3]);
/* radiance/shader/math.cl:56-183; returns 0 if singular */
int rdo_inverse4x4(const float m[16], float out[16]);
/* radiance/shader/radiance.cl:195-208 */
int rdo_intersect_aabb(const float o[3], const float d[3], const float lo[3], const float hi[3]);
/* radiance/shader/radiance.cl:211-251; v = 3 vertices x 3 floats; out = {t, px,py,pz, b0,b1,b2} */
int rdo_intersect_triangle(const float o[3], const float d[3], const float v[9], float out[7]);
/* radiance/shader/radiance.cl:41-192,254-275 over a top-level blob in the reference's byte layout
 * (radiance/src/core.h:34-101). rays: n x {o.xyz, Tmin, d.xyz, Tmax}. */
/* as rdo_trace with a top-level stack of top_stack entries (8 = the reference); overflow[i] = 1 where the limit cut ray i */
void rdo_trace_ex(const void* tlas_blob, const float* rays, size_t n, int terminate_on_first_hit, rdo_hit* out, int nthreads,
                  int top_stack, unsigned char* overflow);
void rdo_trace(const void* tlas_blob, const float* rays, size_t n, int terminate_on_first_hit, rdo_hit* out, int nthreads);

/* Blob builders in the reference's layout (radiance/src/radiance.cpp:318-425) over a median-split
 * binary BVH with leaves of < 8 primitives (the reference's leaf rule, bvh.h:9). Topology differs
 * from the reference's SAH build; hit results do not depend on it except for exact ties. */
void* rdo_blas_build(const float* verts, uint32_t nv, const uint32_t* idx, uint32_t nt, size_t* size);
typedef struct rdo_instance { float transform[16]; uint32_t sbt, custom; const void* blas; size_t blas_size; } rdo_instance;
void* rdo_tlas_build(const rdo_instance* inst, uint32_t n, size_t* size);
void rdo_free(void* p);

#ifdef __cplusplus
}
#endif
#endif
