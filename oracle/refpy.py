"""ctypes binding for oracle/_ref/libradiance_ref.so -- TEST INFRASTRUCTURE.

The .so holds the reference's own code (radiance/src/bvh.cpp unchanged; radiance/shader/*.cl and
samples/shader.cl compiled unmodified as C++), see oracle/Makefile, ref_host.cpp, ref_device.cpp.
Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
import this module; the product never does.
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "_ref", "libradiance_ref.so")


class RefInstance(C.Structure):
    _fields_ = [("transform", C.c_float * 16), ("SBTOffset", C.c_uint),
                ("customInstanceID", C.c_uint), ("blas", C.c_void_p)]


REF_HIT_DTYPE = np.dtype([
    ("t", "<f4"), ("prim", "<u4"), ("inst", "<u4"), ("custom", "<u4"), ("sbt", "<u4"),
    ("bary", "<f4", (3,)), ("point", "<f4", (3,)), ("hit", "<u4"), ("anyhit_calls", "<u4"),
    ("transform", "<f4", (16,))])


def available():
    return os.path.exists(_SO)


def sample_sources():
    """(samples/shader.cl, samples/sbt.json) of the reference, unmodified, as oracle/Makefile copied them into
    the git-ignored _ref/samples/ (config C1: the reference's own user program runs on the GPU); None if absent."""
    cl, sbt = os.path.join(_HERE, "_ref", "samples", "shader.cl"), os.path.join(_HERE, "_ref", "samples", "sbt.json")
    if not (os.path.exists(cl) and os.path.exists(sbt)):
        return None
    return open(cl).read(), open(sbt).read()


_lib = None


def lib():
    global _lib
    if _lib is None:
        L = C.CDLL(_SO)
        L.ref_blas_build.restype = C.c_void_p
        L.ref_blas_build.argtypes = [C.c_void_p, C.c_uint, C.c_void_p, C.c_uint]
        L.ref_blas_size.restype = C.c_size_t
        L.ref_blas_size.argtypes = [C.c_void_p]
        L.ref_blas_data.restype = C.c_void_p
        L.ref_blas_data.argtypes = [C.c_void_p]
        L.ref_blas_free.argtypes = [C.c_void_p]
        L.ref_tlas_build.restype = C.c_void_p
        L.ref_tlas_build.argtypes = [C.c_void_p, C.c_uint, C.POINTER(C.c_size_t)]
        L.ref_free.argtypes = [C.c_void_p]
        L.ref_sizeof.restype = C.c_uint
        L.ref_trace_rays.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_int, C.c_void_p, C.c_int]
        L.ref_pcg3d.argtypes = [C.c_void_p, C.c_size_t, C.c_void_p]
        L.ref_inverse4x4.argtypes = [C.c_void_p, C.c_void_p]
        L.ref_intersect_aabb.argtypes = [C.c_void_p] * 4
        L.ref_intersect_triangle.argtypes = [C.c_void_p] * 4
        L.ref_generate_ray.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p, C.c_void_p]
        L.ref_raygen_shader1.argtypes = [C.c_void_p, C.c_size_t, C.c_size_t, C.c_size_t, C.c_int]
        L.ref_library_kat.argtypes = [C.c_int, C.c_void_p, C.c_size_t, C.c_void_p]
        L.ref_library_kat.restype = C.c_int
        L.ref_raygen.argtypes = [C.c_int, C.c_void_p, C.c_size_t, C.c_size_t, C.c_size_t, C.c_int]
        assert C.sizeof(RefInstance) == 80
        _lib = L
    return _lib


def _f32(a):
    return np.ascontiguousarray(a, dtype=np.float32)


def _u32(a):
    return np.ascontiguousarray(a, dtype=np.uint32)


class RefBlas:
    """RD::BuildAccelStruct(Platform*, Mesh&) -- radiance/src/radiance.cpp:20-52."""

    def __init__(self, verts, tris):
        v, t = _f32(verts).reshape(-1, 3), _u32(tris).reshape(-1, 3)
        self.handle = lib().ref_blas_build(v.ctypes.data, v.shape[0], t.ctypes.data, t.shape[0])

    def blob(self):
        n = lib().ref_blas_size(self.handle)
        return np.ctypeslib.as_array(C.cast(lib().ref_blas_data(self.handle), C.POINTER(C.c_ubyte)), (n,)).copy()


class RefTlas:
    """RD::BuildAccelStruct(Platform*, vector<Instance>&) -- radiance/src/radiance.cpp:54-84.

    instances: list of (transform 4x4 row-major, SBTOffset, customInstanceID, RefBlas)."""

    def __init__(self, instances):
        arr = (RefInstance * len(instances))()
        self._keep = [i[3] for i in instances]
        for k, (m, sbt, custom, blas) in enumerate(instances):
            arr[k].transform[:] = list(_f32(m).reshape(16))
            arr[k].SBTOffset, arr[k].customInstanceID, arr[k].blas = int(sbt), int(custom), blas.handle
        size = C.c_size_t(0)
        p = lib().ref_tlas_build(arr, len(instances), C.byref(size))
        self.blob = np.ctypeslib.as_array(C.cast(p, C.POINTER(C.c_ubyte)), (size.value,)).copy()
        lib().ref_free(p)

    def trace(self, rays, terminate_on_first_hit=False, nthreads=None):
        """rays: (n,8) float32 {o.xyz,Tmin,d.xyz,Tmax} -> structured array REF_HIT_DTYPE."""
        r = _f32(rays).reshape(-1, 8)
        out = np.zeros(r.shape[0], dtype=REF_HIT_DTYPE)
        lib().ref_trace_rays(self.blob.ctypes.data, r.ctypes.data, r.shape[0],
                             int(terminate_on_first_hit), out.ctypes.data,
                             nthreads or os.cpu_count() or 1)
        return out


def build_scene(bundle):
    """The scene of a SceneBundle (meshes + instance list) through the reference's own host builder:
    RD::BuildAccelStruct per mesh, then over the instances (samples/sample1.cpp:363-443)."""
    blas = [RefBlas(v, t) for (v, t) in bundle.meshes]
    return RefTlas([(m, sbt, custom, blas[mi]) for (mi, m, sbt, custom) in bundle.instances])


def pcg3d(v):
    v = _u32(v).reshape(-1, 3)
    out = np.zeros(v.shape, np.float32)
    lib().ref_pcg3d(v.ctypes.data, v.shape[0], out.ctypes.data)
    return out


def inverse4x4(m):
    m = _f32(m).reshape(16)
    out = np.zeros(16, np.float32)
    ok = lib().ref_inverse4x4(m.ctypes.data, out.ctypes.data)
    return bool(ok), out.reshape(4, 4)


def intersect_aabb(o, d, lo, hi):
    o, d, lo, hi = _f32(o), _f32(d), _f32(lo), _f32(hi)
    return bool(lib().ref_intersect_aabb(o.ctypes.data, d.ctypes.data, lo.ctypes.data, hi.ctypes.data))


def intersect_triangle(o, d, verts):
    o, d, v = _f32(o), _f32(d), _f32(verts).reshape(9)
    out = np.zeros(7, np.float32)
    ok = lib().ref_intersect_triangle(o.ctypes.data, d.ctypes.data, v.ctypes.data, out.ctypes.data)
    return bool(ok), out


def generate_ray(cam48, rand3, index):
    cam = np.ascontiguousarray(cam48)
    r = _u32(rand3)
    o, d = np.zeros(3, np.float32), np.zeros(3, np.float32)
    lib().ref_generate_ray(cam.ctypes.data, r.ctypes.data, int(index), o.ctypes.data, d.ctypes.data)
    return o, d


LIBRARY_KAT = {"BRDF": (0, 14, 3), "refract": (1, 7, 3), "G_pbrt": (2, 10, 1), "sampleMicrofacetBRDF_transm": (3, 16, 6),
               "microfacetBRDF": (4, 16, 3), "GetNormalSpace": (5, 3, 16), "sampleMicrofacetBRDF": (6, 14, 6)}


def library_kat(name, inputs):
    """The reference's pbr.cl / math.cl function `name` on rows of flat float inputs (argument order of the
    function, vectors expanded); returns rows of outputs (oracle/ref_device.cpp ref_library_kat)."""
    fn, ins, outs = LIBRARY_KAT[name]
    a = _f32(inputs).reshape(-1, ins)
    out = np.zeros((a.shape[0], outs), np.float32)
    assert lib().ref_library_kat(fn, a.ctypes.data, a.shape[0], out.ctypes.data) == outs
    return out


def raygen_shader1(args, global_size, first=0, count=None, nthreads=None):
    """Run samples/shader.cl `raygen` over work-items [first, first+count) on the host.

    args: 14 entries in descriptor-set order (shader.cl:176-190); numpy arrays are passed by
    address, ints as raw 8-byte values (sampler token), None as NULL."""
    ptrs = (C.c_void_p * 14)()
    for i, a in enumerate(args):
        if a is None:
            ptrs[i] = None
        elif isinstance(a, np.ndarray):
            ptrs[i] = a.ctypes.data
        else:
            ptrs[i] = int(a)
    lib().ref_raygen_shader1(ptrs, int(global_size), int(first),
                             int(global_size - first if count is None else count),
                             nthreads or os.cpu_count() or 1)


PROGRAMS = {"shader": 0, "primary": 1, "path": 2, "groups": 3, "textured": 4}


def raygen(program, args, global_size, first=0, count=None, nthreads=None):
    """NDRange emulation (local size 1, as RD::TraceRays launches it: radiance/src/radiance.cpp:250-259)
    of a user program's `raygen` built against the reference runtime. program: "shader" =
    samples/shader.cl, "primary" = radiance-ray-tracing_b200/programs/primary.cl. args: 14 entries in
    descriptor-set order; numpy arrays by address, ints as raw 8-byte values, None as NULL."""
    ptrs = (C.c_void_p * 14)()
    for i, a in enumerate(args):
        if a is None:
            ptrs[i] = None
        elif isinstance(a, np.ndarray):
            ptrs[i] = a.ctypes.data
        else:
            ptrs[i] = int(a)
    lib().ref_raygen(PROGRAMS[program], ptrs, int(global_size), int(first),
                     int(global_size - first if count is None else count), nthreads or os.cpu_count() or 1)
