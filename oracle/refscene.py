"""ctypes view of oracle/_ref/libscene_ref.so (TEST INFRASTRUCTURE): the reference's own tools/sceneBuilder.cpp run on the
reference's vendored assimp glTF2 importer, with its Radiance API calls recorded (oracle/ref_scene.cpp). Only tests call this."""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "_ref", "libscene_ref.so")
_lib = None

MESH_INFO = np.dtype([("vertexOffset", "<i4"), ("indexOffset", "<i4"), ("uvOffset", "<i4"), ("normalOffset", "<i4"),
                      ("materialIndex", "<i4"), ("pad", "<i4", (3,))])                      # radiance/src/core.h MeshInfo, 32 B
MATERIAL = np.dtype([("albedo", "<f4", (4,)), ("metallic", "<f4"), ("roughness", "<f4"), ("transmission", "<f4"), ("ior", "<f4"),
                     ("albedoTexIdx", "<i4"), ("metallicTexIdx", "<i4"), ("roughnessTexIdx", "<i4"), ("normalTexIdx", "<i4")])  # 48 B


def available():
    return os.path.exists(_SO)


def lib():
    global _lib
    if _lib is None:
        L = C.CDLL(_SO)
        L.ref_scene_load.restype = C.c_void_p
        L.ref_scene_load.argtypes = [C.c_char_p]
        L.ref_scene_array.restype = C.c_size_t
        L.ref_scene_array.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_void_p), C.POINTER(C.c_uint)]
        L.ref_scene_mesh_count.argtypes = [C.c_void_p]
        L.ref_scene_mesh.argtypes = [C.c_void_p, C.c_uint, C.POINTER(C.c_void_p), C.POINTER(C.c_uint), C.POINTER(C.c_void_p), C.POINTER(C.c_uint)]
        L.ref_scene_instances.argtypes = [C.c_void_p, C.POINTER(C.c_void_p), C.POINTER(C.c_void_p)]
        L.ref_scene_sampler.argtypes = [C.c_void_p, C.POINTER(C.c_uint)]
        L.ref_scene_free.argtypes = [C.c_void_p]
        L.ref_scene_sizes.argtypes = [C.POINTER(C.c_uint)]
        _lib = L
    return _lib


def _copy(ptr, nbytes, dtype):
    if not nbytes:
        return np.zeros(0, dtype)
    return np.frombuffer(C.string_at(ptr, nbytes), dtype=dtype).copy()


def load(path, texture_crop=None):
    """RD::Scene::Load(path) of the reference -> dict of numpy arrays (what it wrote through the Radiance API)."""
    L = lib()
    sizes = (C.c_uint * 4)()
    L.ref_scene_sizes(sizes)
    assert tuple(sizes) == (MESH_INFO.itemsize, MATERIAL.itemsize, 12, 12), tuple(sizes)
    h = L.ref_scene_load(os.fsencode(path))
    if not h:
        raise RuntimeError("reference RD::Scene::Load failed for %s" % path)
    out = {}
    try:
        names = ["meshInfoData", "vertexData", "indexData", "uvData", "normalData", "materialData", "textureData"]
        dts = [MESH_INFO, np.float32, np.uint32, np.float32, np.float32, MATERIAL, np.uint8]
        for k, (name, dt) in enumerate(zip(names, dts)):
            ptr, dims = C.c_void_p(), (C.c_uint * 3)()
            n = L.ref_scene_array(h, k, C.byref(ptr), dims)
            a = _copy(ptr, n, dt)
            if name in ("vertexData", "uvData", "normalData", "indexData"):
                a = a.reshape(-1, 3)
            if name == "textureData":
                a = a.reshape(dims[2], dims[1], dims[0], 4) if n else a.reshape(0, 0, 0, 4)
            out[name] = a
        meshes = []
        for i in range(L.ref_scene_mesh_count(h)):
            pv, pi, nv, nt = C.c_void_p(), C.c_void_p(), C.c_uint(), C.c_uint()
            L.ref_scene_mesh(h, i, C.byref(pv), C.byref(nv), C.byref(pi), C.byref(nt))
            meshes.append((_copy(pv, nv.value * 12, np.float32).reshape(-1, 3), _copy(pi, nt.value * 12, np.uint32).reshape(-1, 3)))
        out["meshes"] = meshes
        pt, pm = C.c_void_p(), C.c_void_p()
        n = L.ref_scene_instances(h, C.byref(pt), C.byref(pm))
        out["instanceTransforms"] = _copy(pt, n * 64, np.float32).reshape(-1, 4, 4)
        out["instanceMeta"] = _copy(pm, n * 12, np.uint32).reshape(-1, 3)      # SBTOffset, customInstanceID, mesh index
        s = (C.c_uint * 2)()
        L.ref_scene_sampler(h, s)
        out["sampler"] = (int(s[0]), int(s[1]))
    finally:
        L.ref_scene_free(h)
    return out
