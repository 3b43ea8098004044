"""ctypes binding for oracle/librd_oracle.so (plain-C restatement, rd_oracle.c) -- TEST INFRASTRUCTURE.
Only tests/, __graft_entry__.smoke() and bench.py's CPU legs may import this module."""
import ctypes as C
import os

import numpy as np

_SO = os.path.join(os.path.dirname(os.path.abspath(__file__)), "librd_oracle.so")
HIT_DTYPE = np.dtype([("t", "<f4"), ("prim", "<u4"), ("inst", "<u4"), ("custom", "<u4"), ("sbt", "<u4"),
                      ("bary", "<f4", (3,)), ("point", "<f4", (3,)), ("hit", "<u4"), ("anyhit_calls", "<u4")])


class Instance(C.Structure):
    _fields_ = [("transform", C.c_float * 16), ("sbt", C.c_uint32), ("custom", C.c_uint32), ("blas", C.c_void_p),
                ("blas_size", C.c_size_t)]


def available():
    return os.path.exists(_SO)


_lib = None


def lib():
    global _lib
    if _lib is None:
        L = C.CDLL(_SO)
        L.rdo_blas_build.restype = C.c_void_p
        L.rdo_blas_build.argtypes = [C.c_void_p, C.c_uint32, C.c_void_p, C.c_uint32, C.POINTER(C.c_size_t)]
        L.rdo_tlas_build.restype = C.c_void_p
        L.rdo_tlas_build.argtypes = [C.c_void_p, C.c_uint32, C.POINTER(C.c_size_t)]
        L.rdo_free.argtypes = [C.c_void_p]
        L.rdo_trace.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_int, C.c_void_p, C.c_int]
        L.rdo_trace_ex.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_int, C.c_void_p, C.c_int, C.c_int, C.c_void_p]
        L.rdo_pcg3d.argtypes = [C.c_void_p, C.c_void_p]
        L.rdo_inverse4x4.argtypes = [C.c_void_p, C.c_void_p]
        L.rdo_intersect_aabb.argtypes = [C.c_void_p] * 4
        L.rdo_intersect_triangle.argtypes = [C.c_void_p] * 4
        assert HIT_DTYPE.itemsize == 52
        _lib = L
    return _lib


def _blob(ptr, size):
    out = np.ctypeslib.as_array(C.cast(ptr, C.POINTER(C.c_ubyte)), (size,)).copy()
    lib().rdo_free(ptr)
    return out


def blas_blob(verts, tris):
    v = np.ascontiguousarray(verts, np.float32).reshape(-1, 3)
    t = np.ascontiguousarray(tris, np.uint32).reshape(-1, 3)
    size = C.c_size_t(0)
    return _blob(lib().rdo_blas_build(v.ctypes.data, v.shape[0], t.ctypes.data, t.shape[0], C.byref(size)), size.value)


def tlas_blob(instances):
    """instances: list of (transform 4x4, SBTOffset, customInstanceID, blas blob ndarray)."""
    arr = (Instance * max(1, len(instances)))()
    for k, (m, sbt, custom, blob) in enumerate(instances):
        arr[k].transform[:] = [float(x) for x in np.asarray(m, np.float32).reshape(16)]
        arr[k].sbt, arr[k].custom, arr[k].blas, arr[k].blas_size = int(sbt), int(custom), blob.ctypes.data, blob.nbytes
    size = C.c_size_t(0)
    return _blob(lib().rdo_tlas_build(arr, len(instances), C.byref(size)), size.value)


def scene_blob(bundle):
    blobs = [blas_blob(v, t) for (v, t) in bundle.meshes]
    return tlas_blob([(m, sbt, custom, blobs[mi]) for (mi, m, sbt, custom) in bundle.instances])


def trace(blob, rays, terminate_on_first_hit=False, nthreads=None):
    """Works on any top-level blob in the reference layout (ours or oracle/_ref's)."""
    r = np.ascontiguousarray(rays, np.float32).reshape(-1, 8)
    out = np.zeros(r.shape[0], HIT_DTYPE)
    lib().rdo_trace(blob.ctypes.data, r.ctypes.data, r.shape[0], int(terminate_on_first_hit), out.ctypes.data,
                    nthreads or os.cpu_count() or 1)
    return out


def trace_ex(blob, rays, terminate_on_first_hit=False, top_stack=8, nthreads=None):
    """trace() with a top-level stack of `top_stack` entries (8 = the reference, radiance.cl:38). Returns (hits, overflow):
    overflow[i] = 1 where the limit made the walk give up and report a miss (the reference does that at 4096 instances)."""
    r = np.ascontiguousarray(rays, np.float32).reshape(-1, 8)
    out, ovf = np.zeros(r.shape[0], HIT_DTYPE), np.zeros(r.shape[0], np.uint8)
    lib().rdo_trace_ex(blob.ctypes.data, r.ctypes.data, r.shape[0], int(terminate_on_first_hit), out.ctypes.data,
                       nthreads or os.cpu_count() or 1, int(top_stack), ovf.ctypes.data)
    return out, ovf


def pcg3d(v):
    v = np.ascontiguousarray(v, np.uint32).reshape(-1, 3)
    out = np.zeros(v.shape, np.float32)
    for i in range(v.shape[0]):
        lib().rdo_pcg3d(v[i].ctypes.data, out[i].ctypes.data)
    return out


def inverse4x4(m):
    m = np.ascontiguousarray(m, np.float32).reshape(16)
    out = np.zeros(16, np.float32)
    return bool(lib().rdo_inverse4x4(m.ctypes.data, out.ctypes.data)), out.reshape(4, 4)


def intersect_triangle(o, d, verts):
    o, d, v = (np.ascontiguousarray(x, np.float32) for x in (o, d, verts))
    out = np.zeros(7, np.float32)
    return bool(lib().rdo_intersect_triangle(o.ctypes.data, d.ctypes.data, v.reshape(9).ctypes.data, out.ctypes.data)), out


def intersect_aabb(o, d, lo, hi):
    o, d, lo, hi = (np.ascontiguousarray(x, np.float32) for x in (o, d, lo, hi))
    return bool(lib().rdo_intersect_aabb(o.ctypes.data, d.ctypes.data, lo.ctypes.data, hi.ctypes.data))
