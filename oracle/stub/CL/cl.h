/* Minimal stand-in for <CL/cl.h>: opaque handle typedefs and the few constants that
 * radiance/include/radiance.h and radiance/src/clcontext.h mention. It exists only so the
 * reference's host BVH builder (radiance/src/bvh.cpp) compiles unchanged for oracle/_ref;
 * no OpenCL entry point is declared or called. Test infrastructure, not product code. */
#ifndef RD_ORACLE_STUB_CL_H
#define RD_ORACLE_STUB_CL_H
#include <stdint.h>
#include <stddef.h>
typedef struct _cl_platform_id* cl_platform_id;
typedef struct _cl_device_id* cl_device_id;
typedef struct _cl_context* cl_context;
typedef struct _cl_command_queue* cl_command_queue;
typedef struct _cl_mem* cl_mem;
typedef struct _cl_program* cl_program;
typedef struct _cl_kernel* cl_kernel;
typedef struct _cl_sampler* cl_sampler;
typedef int32_t cl_int;
typedef uint32_t cl_uint;
#define CL_SUCCESS 0
#define CL_INVALID_VALUE (-30)
#define CL_ADDRESS_NONE 0x1130
#define CL_ADDRESS_CLAMP_TO_EDGE 0x1131
#define CL_ADDRESS_CLAMP 0x1132
#define CL_ADDRESS_REPEAT 0x1133
#define CL_ADDRESS_MIRRORED_REPEAT 0x1134
#define CL_FILTER_NEAREST 0x1140
#define CL_FILTER_LINEAR 0x1141
#endif
