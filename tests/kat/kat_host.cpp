// tests/kat/kat_host.cpp -- TEST INFRASTRUCTURE: tests/kat/kat.cl compiled for the host against OUR shader library
// (shader/math.cl, pbr.cl, csrc/rd_traverse.h through the host flavour of cl_compat.h), so the CPU suite checks the
// restated library functions against the golden vectors of the reference without a GPU.
#include <cstddef>
#include <cstdint>
#include <cstring>
#include "cl_compat.h"
#include "rd_traverse.h"
thread_local size_t rd_host_global_id = 0;
size_t rd_host_global_size = 0;
#include "kat.cl"
extern "C" void kat_run(unsigned* header, float* in, float* out) {
    for (unsigned i = 0; i < header[1]; ++i) { rd_host_global_id = i; raygen(header, in, out); }
}
