"""Host-side mirror of the reference's `namespace RD` API on top of the C ABI (include/radiance_b200.h).

Function names, argument order and error behaviour follow radiance/include/radiance.h:88-144 so that
drivers and tests read like the reference's samples (samples/sample1.cpp:363-443). Everything goes
through libradiance_b200.so with plain pointers; numpy is only used to hold host arrays. There is no
CPU execution path: without the CUDA library / a device every call raises.
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("RADIANCE_B200_LIB") or os.path.join(_HERE, "libradiance_b200.so")   # developer knob: ablation builds

RD_ADDRESS_CLAMP_TO_EDGE, RD_ADDRESS_CLAMP, RD_ADDRESS_REPEAT, RD_ADDRESS_MIRRORED_REPEAT = 0x1131, 0x1132, 0x1133, 0x1134
RD_FILTER_NEAREST, RD_FILTER_LINEAR = 0x1140, 0x1141
TRACE_TERMINATE_ON_FIRST_HIT, TRACE_COUNTERS, TRACE_STATIC_BATCHES, TRACE_DYNAMIC_FETCH = 1, 2, 4, 8

# every symbol include/radiance_b200.h declares (checked by tests/test_abi.py)
ABI_SYMBOLS = [
    "rd_last_error", "rd_platform_create", "rd_platform_destroy", "rd_platform_stream", "rd_platform_sync",
    "rd_buffer_create", "rd_buffer_wrap", "rd_image_create", "rd_image_array_create", "rd_sampler_create", "rd_buffer_write",
    "rd_buffer_read", "rd_image_write", "rd_image_read", "rd_handle_free", "rd_handle_device_ptr",
    "rd_blas_build", "rd_blas_build_device", "rd_blas_free", "rd_blas_stats", "rd_tlas_build",
    "rd_shader_module_create", "rd_module_free", "rd_shader_compile_check", "rd_module_info", "rd_sbt_load_json",
    "rd_pipeline_bind", "rd_descriptors_bind", "rd_trace_rays", "rd_trace_rays_async", "rd_trace_rays_range", "rd_trace_ray_list",
    "rd_trace_counters", "rd_generate_primary_rays", "rd_set_option", "rd_launch_count", "rd_ray_count",
    "rd_tlas_load", "rd_tlas_load_memory", "rd_tlas_save", "rd_tlas_save_memory", "rd_free_host",
    "rd_steal_counter_export", "rd_steal_counter_import", "rd_steal_counter_adopt", "rd_trace_rays_steal", "rd_steal_owned", "rd_trace_rays_interleaved",
    "rd_debug_sort_pairs", "rd_debug_exclusive_scan", "rd_buffer_stream_to_host", "rd_status",
    "rd_comm_unique_id", "rd_comm_init", "rd_comm_adopt", "rd_comm_destroy", "rd_frame_gather", "rd_frame_gather_ranges",
    "rd_trace_rays_units", "rd_frame_gather_units", "rd_wf_timeline", "rd_debug_band_roundtrip", "rd_debug_bounds_violations",
    "rd_sort_pairs_noalloc",
]


class RadianceError(RuntimeError):
    """The reference aborts the process on any error (clcontext.h:27-47); the mirror raises."""


class InstanceDesc(C.Structure):
    _fields_ = [("transform", C.c_float * 16), ("SBTOffset", C.c_uint32), ("customInstanceID", C.c_uint32),
                ("bottomAccelStruct", C.c_void_p)]


_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RadianceError("libradiance_b200.so is not built (run __graft_entry__.build()); there is no fallback path")
        L = C.CDLL(LIB_PATH)
        L.rd_last_error.restype = C.c_char_p
        L.rd_platform_stream.restype = C.c_void_p
        L.rd_platform_stream.argtypes = [C.c_void_p]
        L.rd_handle_device_ptr.restype = C.c_void_p
        L.rd_handle_device_ptr.argtypes = [C.c_void_p]
        L.rd_launch_count.restype = C.c_uint64
        L.rd_launch_count.argtypes = [C.c_void_p]
        vp, sz, u = C.c_void_p, C.c_size_t, C.c_uint
        sigs = {
            "rd_platform_create": [C.c_int, C.POINTER(vp)], "rd_platform_destroy": [vp], "rd_platform_sync": [vp],
            "rd_buffer_create": [vp, sz, C.POINTER(vp)], "rd_buffer_wrap": [vp, vp, sz, C.POINTER(vp)], "rd_image_create": [vp, u, u, C.POINTER(vp)],
            "rd_image_array_create": [vp, u, u, u, C.POINTER(vp)], "rd_sampler_create": [vp, C.c_uint32, C.c_uint32, C.POINTER(vp)],
            "rd_buffer_write": [vp, vp, sz, vp, sz], "rd_buffer_read": [vp, vp, sz, vp, sz],
            "rd_buffer_stream_to_host": [vp, vp, vp, u, u],
            "rd_image_write": [vp, vp, u, u, sz, vp], "rd_image_read": [vp, vp, u, u, sz, vp],
            "rd_handle_free": [vp, vp], "rd_blas_build": [vp, vp, u, vp, u, C.POINTER(vp)],
            "rd_blas_build_device": [vp, vp, u, vp, u, C.POINTER(vp)], "rd_blas_free": [vp, vp],
            "rd_blas_stats": [vp, C.POINTER(C.c_uint64)], "rd_tlas_build": [vp, vp, u, C.POINTER(vp)],
            "rd_shader_module_create": [vp, C.c_char_p, u, C.c_char_p, C.POINTER(vp)], "rd_module_free": [vp, vp],
            "rd_shader_compile_check": [C.c_char_p, u, C.c_char_p, C.c_int, C.POINTER(sz)],
            "rd_module_info": [vp, C.POINTER(C.c_int)], "rd_sbt_load_json": [vp, C.c_char_p],
            "rd_pipeline_bind": [vp, C.POINTER(vp), u], "rd_descriptors_bind": [vp, C.POINTER(vp), u],
            "rd_trace_rays": [vp, u, u, u, u, u], "rd_trace_rays_async": [vp, u, u], "rd_trace_rays_range": [vp, u, u, C.c_ulonglong, C.c_ulonglong],
            "rd_trace_ray_list": [vp, vp, vp, vp, vp, vp, sz, u], "rd_trace_counters": [vp, C.POINTER(C.c_uint64)],
            "rd_generate_primary_rays": [vp, vp, u, u, u, u, vp, vp], "rd_set_option": [vp, C.c_char_p, C.c_long],
            "rd_ray_count": [vp, C.POINTER(C.c_uint64), C.c_int],
            "rd_status": [vp, C.POINTER(C.c_uint32), C.c_int],
            "rd_comm_unique_id": [vp], "rd_comm_init": [vp, C.c_int, C.c_int, vp], "rd_comm_adopt": [vp, vp, C.c_int, C.c_int],
            "rd_comm_destroy": [vp], "rd_frame_gather": [vp, vp, C.c_uint, C.c_uint, C.c_uint, C.c_uint],
            "rd_frame_gather_ranges": [vp, vp, C.c_uint, vp, vp],
            "rd_trace_rays_units": [vp, C.c_uint, C.c_uint, C.c_uint, C.c_uint, C.POINTER(C.c_uint64)],
            "rd_frame_gather_units": [vp, vp, C.c_uint, C.c_uint, C.c_uint, C.c_uint, C.c_uint64],
            "rd_wf_timeline": [vp, vp, vp, C.c_uint, C.POINTER(C.c_uint)],
            "rd_debug_band_roundtrip": [vp, vp, vp] + [C.c_uint] * 6,
            "rd_debug_bounds_violations": [vp, C.POINTER(C.c_uint), C.POINTER(C.c_int)],
            "rd_tlas_load": [vp, C.c_char_p, C.POINTER(vp)], "rd_tlas_load_memory": [vp, vp, sz, C.POINTER(vp), vp, vp],
            "rd_steal_counter_export": [vp, vp], "rd_steal_counter_import": [vp, vp], "rd_trace_rays_steal": [vp, u, u, u, u],
            "rd_steal_owned": [vp, vp, u, C.POINTER(u)], "rd_trace_rays_interleaved": [vp, u, u, u, u, u],
            "rd_debug_sort_pairs": [vp, vp, vp, u, C.c_int], "rd_sort_pairs_noalloc": [vp, vp, vp, vp, vp, u, C.c_int], "rd_debug_exclusive_scan": [vp, vp, u, C.POINTER(u)],
            "rd_tlas_save": [vp, vp, C.c_char_p], "rd_tlas_save_memory": [vp, vp, C.POINTER(vp), C.POINTER(sz)],
        }
        L.rd_free_host.argtypes, L.rd_free_host.restype = [vp], None
        for name, args in sigs.items():
            f = getattr(L, name)
            f.argtypes, f.restype = args, C.c_int
        _lib = L
    return _lib


def _check(rc):
    if rc != 0:
        raise RadianceError("radiance_b200 error %d: %s" % (rc, lib().rd_last_error().decode(errors="replace")))


def last_message():
    return lib().rd_last_error().decode(errors="replace")


class Platform:
    """RD::Platform (radiance.h:146-174). GetPlatform() returns a per-process singleton per device."""
    _instances = {}

    def __init__(self, device=-1):
        self.handle = C.c_void_p()
        _check(lib().rd_platform_create(device, C.byref(self.handle)))
        self.activePipeline = None

    @staticmethod
    def GetPlatform(device=-1):
        if device not in Platform._instances:
            Platform._instances[device] = Platform(device)
        return Platform._instances[device]

    def stream(self):
        return lib().rd_platform_stream(self.handle)

    def sync(self):
        _check(lib().rd_platform_sync(self.handle))

    def status(self, clear=False):
        """(stack overflow, traceRay nesting exceeded) device flags, without synchronising."""
        out = (C.c_uint32 * 2)()
        _check(lib().rd_status(self.handle, out, int(bool(clear))))
        return int(out[0]), int(out[1])

    def bounds_violations(self):
        """(violations since the last call, 1 if the library is an RD_BOUNDS_CHECK build)."""
        n, chk = C.c_uint(0), C.c_int(0)
        _check(lib().rd_debug_bounds_violations(self.handle, C.byref(n), C.byref(chk)))
        return int(n.value), int(chk.value)

    def launch_count(self):
        return int(lib().rd_launch_count(self.handle))

    def set_option(self, key, value):
        _check(lib().rd_set_option(self.handle, key.encode(), int(value)))

    def ray_count(self, reset=True):
        """traceRay invocations counted by modules compiled under option count_rays."""
        v = C.c_uint64(0)
        _check(lib().rd_ray_count(self.handle, C.byref(v), int(bool(reset))))
        return int(v.value)


class Mesh:
    """RD::Mesh (radiance.h:32-36): vertexData (nv,3) float32, indexData (nt,3) uint32."""

    def __init__(self, vertexData, indexData):
        self.vertexData = np.ascontiguousarray(vertexData, np.float32).reshape(-1, 3)
        self.indexData = np.ascontiguousarray(indexData, np.uint32).reshape(-1, 3)


class Instance:
    """RD::Instance (radiance.h:48-55)."""

    def __init__(self, transform, SBTOffset, customInstanceID, bottomAccelStruct):
        self.transform = np.ascontiguousarray(transform, np.float32).reshape(4, 4)
        self.SBTOffset, self.customInstanceID, self.bottomAccelStruct = int(SBTOffset), int(customInstanceID), bottomAccelStruct


class BottomAccelStruct:
    def __init__(self, plt, handle):
        self.plt, self.handle = plt, handle

    def stats(self):
        out = (C.c_uint64 * 4)()
        _check(lib().rd_blas_stats(self.handle, out))
        return {"nodes": out[0], "triangles": out[1], "levels": out[2], "bytes": out[3]}

    def free(self):
        if self.handle:
            lib().rd_blas_free(self.plt.handle, self.handle)
            self.handle = None


class Handle:
    """Buffer / Image / ImageArray / Sampler / TopAccelStruct descriptor handle."""

    def __init__(self, plt, handle, size=0, keep=None):
        self.plt, self.handle, self.size, self._keep = plt, handle, size, keep

    def device_ptr(self):
        return lib().rd_handle_device_ptr(self.handle)

    def free(self):
        if self.handle:
            lib().rd_handle_free(self.plt.handle, self.handle)
            self.handle = None


def BuildAccelStruct(plt, arg):
    """Both overloads of RD::BuildAccelStruct (radiance.cpp:20-52 for a Mesh, :54-84 for instances)."""
    if isinstance(arg, Mesh):
        out = C.c_void_p()
        v, t = arg.vertexData, arg.indexData
        _check(lib().rd_blas_build(plt.handle, v.ctypes.data, v.shape[0], t.ctypes.data, t.shape[0], C.byref(out)))
        return BottomAccelStruct(plt, out)
    instances = list(arg)
    arr = (InstanceDesc * max(1, len(instances)))()
    for k, inst in enumerate(instances):
        C.memmove(arr[k].transform, inst.transform.ctypes.data, 64)   # (4,4) float32, C-contiguous (Instance.__init__)
        arr[k].SBTOffset, arr[k].customInstanceID = inst.SBTOffset, inst.customInstanceID
        arr[k].bottomAccelStruct = inst.bottomAccelStruct.handle
    out = C.c_void_p()
    _check(lib().rd_tlas_build(plt.handle, arr, len(instances), C.byref(out)))
    return Handle(plt, out, keep=[i.bottomAccelStruct for i in instances])


def StealCounterExport(plt):
    """64-byte IPC handle of this rank's chunk counter (rank 0 of a multi-GPU job)."""
    h = np.zeros(64, np.uint8)
    _check(lib().rd_steal_counter_export(plt.handle, h.ctypes.data))
    return h


def StealCounterImport(plt, handle):
    h = np.ascontiguousarray(handle, np.uint8)
    _check(lib().rd_steal_counter_import(plt.handle, h.ctypes.data))


def TraceRaysSteal(plt, width, height, frame, chunk_items=32):
    _check(lib().rd_trace_rays_steal(plt.handle, width, height, frame, chunk_items))


def TraceRaysInterleaved(plt, width, height, rank, world, band_tile_rows=1):
    _check(lib().rd_trace_rays_interleaved(plt.handle, width, height, rank, world, band_tile_rows))


def StealOwned(plt, capacity=1 << 20):
    ids, n = np.zeros(capacity, np.uint32), C.c_uint(0)
    _check(lib().rd_steal_owned(plt.handle, ids.ctypes.data, capacity, C.byref(n)))
    return ids[:min(capacity, n.value)].copy()


def CommUniqueId():
    """128-byte NCCL id created on rank 0; hand it to the other ranks (any transport)."""
    buf = (C.c_ubyte * 128)()
    _check(lib().rd_comm_unique_id(buf))
    return np.frombuffer(bytes(buf), np.uint8).copy()


def CommInit(plt, rank, world, unique_id):
    uid = np.ascontiguousarray(unique_id, np.uint8)
    assert uid.size == 128
    _check(lib().rd_comm_init(plt.handle, int(rank), int(world), uid.ctypes.data))


def CommDestroy(plt):
    _check(lib().rd_comm_destroy(plt.handle))


def FrameGather(plt, image, width, height, item_bytes=4, band_tile_rows=1):
    """After TraceRaysInterleaved on every rank: the bands of all ranks land in rank 0's buffer (async)."""
    _check(lib().rd_frame_gather(plt.handle, image.handle, width, height, item_bytes, band_tile_rows))


def FrameGatherRanges(plt, image, item_bytes, first, count):
    f, c = np.ascontiguousarray(first, np.uint64), np.ascontiguousarray(count, np.uint64)
    _check(lib().rd_frame_gather_ranges(plt.handle, image.handle, item_bytes, f.ctypes.data, c.ctypes.data))


def TraceRaysUnits(plt, width, height, units, frame):
    """Dynamic partition: renders the work units this rank wins; returns their bit mask."""
    mask = C.c_uint64(0)
    _check(lib().rd_trace_rays_units(plt.handle, width, height, units, frame, C.byref(mask)))
    return int(mask.value)


def FrameGatherUnits(plt, image, width, height, item_bytes, units, owned_mask):
    _check(lib().rd_frame_gather_units(plt.handle, image.handle, width, height, item_bytes, units, C.c_uint64(owned_mask)))


WF_STAGE_NAMES = {1: "generate", 2: "trace", 3: "sort", 4: "shade", 5: "accumulate", 6: "resolve", 7: "ray_sort"}


def WavefrontTimeline(plt, capacity=4096):
    """[(stage name, ms)] of the last wavefront-mode call made under option wf_timeline."""
    ms, codes, n = np.zeros(capacity, np.float32), np.zeros(capacity, np.int32), C.c_uint(0)
    _check(lib().rd_wf_timeline(plt.handle, ms.ctypes.data, codes.ctypes.data, capacity, C.byref(n)))
    return [(WF_STAGE_NAMES.get(int(codes[i]), "?"), float(ms[i])) for i in range(min(n.value, capacity))]


def TopAccelStructToFile(plt, accelStruct, path):
    """RD::TopAccelStructToFile (radiance.cpp:428-448): reference blob format."""
    _check(lib().rd_tlas_save(plt.handle, accelStruct.handle, str(path).encode()))


def FileToTopAccelStruct(plt, path):
    """RD::FileToTopAccelStruct (radiance.cpp:450-479); returns the handle (the reference fills an out-parameter)."""
    out = C.c_void_p()
    _check(lib().rd_tlas_load(plt.handle, str(path).encode(), C.byref(out)))
    return Handle(plt, out)


def TopAccelStructToBytes(plt, accelStruct):
    data, size = C.c_void_p(), C.c_size_t(0)
    _check(lib().rd_tlas_save_memory(plt.handle, accelStruct.handle, C.byref(data), C.byref(size)))
    blob = np.ctypeslib.as_array(C.cast(data, C.POINTER(C.c_ubyte)), (size.value,)).copy()
    lib().rd_free_host(data)
    return blob


def TopAccelStructFromBytes(plt, blob):
    blob = np.ascontiguousarray(blob, np.uint8)
    out = C.c_void_p()
    _check(lib().rd_tlas_load_memory(plt.handle, blob.ctypes.data, blob.nbytes, C.byref(out), None, None))
    return Handle(plt, out)


def CreateBuffer(plt, size):
    out = C.c_void_p()
    _check(lib().rd_buffer_create(plt.handle, int(size), C.byref(out)))
    return Handle(plt, out, int(size))


def WrapBuffer(plt, device_ptr, size, owner=None):
    """Buffer handle over caller-owned device memory (e.g. a torch tensor's data_ptr())."""
    out = C.c_void_p()
    _check(lib().rd_buffer_wrap(plt.handle, C.c_void_p(int(device_ptr)), int(size), C.byref(out)))
    return Handle(plt, out, int(size), keep=owner)


def CreateImage(plt, width, height):
    out = C.c_void_p()
    _check(lib().rd_image_create(plt.handle, width, height, C.byref(out)))
    return Handle(plt, out, width * height * 4)


def CreateImageArray(plt, width, height, arraySize):
    out = C.c_void_p()
    _check(lib().rd_image_array_create(plt.handle, width, height, arraySize, C.byref(out)))
    return Handle(plt, out, width * height * 4 * arraySize)


def CreateSampler(plt, addressingMode, filterMode):
    out = C.c_void_p()
    _check(lib().rd_sampler_create(plt.handle, addressingMode, filterMode, C.byref(out)))
    return Handle(plt, out)


def _host_ptr(data, size):
    if isinstance(data, np.ndarray):
        if not data.flags["C_CONTIGUOUS"]:
            raise RadianceError("host array must be C-contiguous")
        if data.nbytes < size:
            raise RadianceError("host array smaller than the requested transfer")
        return data.ctypes.data
    return data


def WriteBuffer(plt, handle, size, data, offset=0):
    _check(lib().rd_buffer_write(plt.handle, handle.handle, int(size), _host_ptr(data, size), int(offset)))


def ReadBuffer(plt, handle, size, data, offset=0):
    _check(lib().rd_buffer_read(plt.handle, handle.handle, int(size), _host_ptr(data, size), int(offset)))


def StreamBufferToHost(plt, handle, host_ptr, item_bytes=4, bands=4):
    """No reference counterpart. Registers page-locked host memory as the mirror of an image buffer: full-frame
    TraceRays of modules created under option `stream_image` then copy the frame out band by band while it renders,
    and ReadBuffer(handle, size, host_ptr) only waits for the last band. host_ptr None / 0 ends it."""
    _check(lib().rd_buffer_stream_to_host(plt.handle, handle.handle if handle is not None else None, host_ptr or None,
                                          int(item_bytes), int(bands)))


def WriteImage(plt, handle, width, height, arrayIndex, data):
    _check(lib().rd_image_write(plt.handle, handle.handle, width, height, arrayIndex, _host_ptr(data, width * height * 4)))


def ReadImage(plt, handle, width, height, arrayIndex, data):
    _check(lib().rd_image_read(plt.handle, handle.handle, width, height, arrayIndex, _host_ptr(data, width * height * 4)))


def CreateDescriptorSet(handles):
    return list(handles)


def CreatePipelineLayout(descriptorTypes):
    return list(descriptorTypes)


class ShaderModule:
    def __init__(self, plt, handle):
        self.plt, self.handle = plt, handle

    def info(self):
        out = (C.c_int * 2)()
        _check(lib().rd_module_info(self.handle, out))
        return {"registers": out[0], "local_bytes": out[1]}


def LoadSBT(plt, json_text):
    """samples/sbt.json text; dispatchers are generated when the program does not define them."""
    _check(lib().rd_sbt_load_json(plt.handle, json_text.encode() if isinstance(json_text, str) else json_text))


def CreateShaderModule(plt, code, size=0, name="raygen"):
    src = code.encode() if isinstance(code, str) else bytes(code)
    out = C.c_void_p()
    _check(lib().rd_shader_module_create(plt.handle, src, len(src) if not size else size, name.encode(), C.byref(out)))
    return ShaderModule(plt, out)


def CompileCheck(code, sbt_json=None, fmad=True):
    """NVRTC compile without a device; returns (cubin_bytes, ptxas_log)."""
    src = code.encode() if isinstance(code, str) else bytes(code)
    n = C.c_size_t(0)
    sbt = None if sbt_json is None else (sbt_json.encode() if isinstance(sbt_json, str) else sbt_json)
    _check(lib().rd_shader_compile_check(src, len(src), sbt, int(bool(fmad)), C.byref(n)))
    return n.value, last_message()


class PipelineCreateInfo:
    def __init__(self, maxRayRecursionDepth=1, layout=None, modules=None, groups=None):
        self.maxRayRecursionDepth, self.layout = maxRayRecursionDepth, layout or []
        self.modules, self.groups = modules or [], groups or []


def CreatePipeline(info):
    return info


def BindPipeline(plt, pipeline):
    mods = (C.c_void_p * max(1, len(pipeline.modules)))(*[m.handle for m in pipeline.modules])
    _check(lib().rd_pipeline_bind(plt.handle, mods, len(pipeline.modules)))
    plt.activePipeline = pipeline


def BindDescriptorSet(plt, descriptorSet):
    arr = (C.c_void_p * max(1, len(descriptorSet)))(*[h.handle for h in descriptorSet])
    _check(lib().rd_descriptors_bind(plt.handle, arr, len(descriptorSet)))
    plt._descriptors = list(descriptorSet)


def TraceRays(plt, raygenGroupIndex, missGroupIndex, hitGroupIndex, width, height):
    _check(lib().rd_trace_rays(plt.handle, raygenGroupIndex, missGroupIndex, hitGroupIndex, width, height))


def TraceRaysAsync(plt, width, height):
    _check(lib().rd_trace_rays_async(plt.handle, width, height))


def TraceRaysRange(plt, width, height, first, count):
    """Work-items [first, first+count) of the NDRange (asynchronous): one GPU's share of the screen tiles."""
    _check(lib().rd_trace_rays_range(plt.handle, width, height, int(first), int(count)))


# ---- wavefront building blocks (no reference counterpart) ----------------------------------------
HIT_DTYPE = np.dtype([("t", "<f4"), ("b1", "<f4"), ("b2", "<f4"), ("prim", "<u4")])


def TraceRayList(plt, tlas, rayA, rayB, hitA, hitB, n, flags=0):
    _check(lib().rd_trace_ray_list(plt.handle, tlas.handle, rayA.handle, rayB.handle, hitA.handle, hitB.handle, int(n), flags))


def debug_sort_pairs(plt, keys, vals, bits=64):
    """The build's radix sort on host arrays (uint64 keys, uint32 values), returns sorted copies."""
    k, v = np.ascontiguousarray(keys, np.uint64).copy(), np.ascontiguousarray(vals, np.uint32).copy()
    _check(lib().rd_debug_sort_pairs(plt.handle, k.ctypes.data, v.ctypes.data, k.size, bits))
    return k, v


SORT_NOALLOC_MAX_PAIRS = 1 << 18


def sort_pairs_noalloc(plt, keysIn, valsIn, keysOut, valsOut, n, bits=64):
    """rd_sort_pairs_noalloc: the same sort on buffers in HBM (Buffer handles or raw device addresses) with its scratch in
    the library's static device array -- no allocation, asynchronous on the platform stream."""
    ptr = lambda b: b if isinstance(b, int) else lib().rd_handle_device_ptr(b.handle)
    _check(lib().rd_sort_pairs_noalloc(plt.handle, ptr(keysIn), ptr(valsIn), ptr(keysOut), ptr(valsOut), n, bits))


def debug_exclusive_scan(plt, data):
    d = np.ascontiguousarray(data, np.uint32).copy()
    total = C.c_uint(0)
    _check(lib().rd_debug_exclusive_scan(plt.handle, d.ctypes.data, d.size, C.byref(total)))
    return d, int(total.value)


def TraceCounters(plt):
    out = (C.c_uint64 * 3)()
    _check(lib().rd_trace_counters(plt.handle, out))
    return {"nodes": out[0], "tris": out[1], "rays": out[2]}


def GeneratePrimaryRays(plt, camera, frameID, totalSamples, width, height, rayA, rayB):
    _check(lib().rd_generate_primary_rays(plt.handle, camera.handle, frameID, totalSamples, width, height, rayA.handle, rayB.handle))


def trace_host_rays(plt, tlas, rays, flags=0):
    """Convenience for tests: rays (n,8) float32 {o,tmin,d,tmax} on the host -> (hitA records, instanceIDs)."""
    rays = np.ascontiguousarray(rays, np.float32).reshape(-1, 8)
    n = rays.shape[0]
    a, b = np.ascontiguousarray(rays[:, 0:4]), np.ascontiguousarray(rays[:, 4:8])
    bufs = [CreateBuffer(plt, max(16, n * 16)) for _ in range(3)] + [CreateBuffer(plt, max(16, n * 4))]
    try:
        if n:
            WriteBuffer(plt, bufs[0], n * 16, a)
            WriteBuffer(plt, bufs[1], n * 16, b)
        TraceRayList(plt, tlas, bufs[0], bufs[1], bufs[2], bufs[3], n, flags)
        plt.sync()
        hitA, hitB = np.zeros(n, HIT_DTYPE), np.zeros(n, np.uint32)
        if n:
            ReadBuffer(plt, bufs[2], n * 16, hitA)
            ReadBuffer(plt, bufs[3], n * 4, hitB)
        return hitA, hitB
    finally:
        for h in bufs:
            h.free()
