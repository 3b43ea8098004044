"""CPU-side tests (no GPU): the oracle against the committed golden vectors, and the device code's
per-element logic (csrc/rd_build.h, csrc/rd_traverse.h looped on the host by tests/hostemu) against
the same vectors. The golden vectors were produced by the reference's own code (tests/golden/make_golden.py)."""
import ctypes as C
import os

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLD = os.path.join(ROOT, "tests", "golden")
EMU_SO = os.path.join(ROOT, "tests", "hostemu", "librd_hostemu.so")
HIT = np.dtype([("t", "<f4"), ("b1", "<f4"), ("b2", "<f4"), ("prim", "<u4"), ("inst", "<u4")])


class EmuInstance(C.Structure):
    _fields_ = [("t", C.c_float * 16), ("sbt", C.c_uint), ("custom", C.c_uint), ("blas", C.c_void_p)]


@pytest.fixture(scope="module")
def emu():
    if not os.path.exists(EMU_SO):
        pytest.skip("tests/hostemu not built (run __graft_entry__.build())")
    L = C.CDLL(EMU_SO)
    L.emu_blas_build.restype = C.c_void_p
    L.emu_blas_build.argtypes = [C.c_void_p, C.c_uint, C.c_void_p, C.c_uint]
    L.emu_tlas_build.restype = C.c_void_p
    L.emu_tlas_build.argtypes = [C.c_void_p, C.c_uint]
    L.emu_trace.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_int, C.c_void_p]
    L.emu_blas_nodes.argtypes = [C.c_void_p]
    L.emu_tlas_status.argtypes = [C.c_void_p]
    L.emu_inverse4x4.argtypes = [C.c_void_p, C.c_void_p]
    return L


def emu_scene(L, bundle):
    blas = []
    for v, t in bundle.meshes:
        v, t = np.ascontiguousarray(v, np.float32), np.ascontiguousarray(t, np.uint32)
        blas.append(L.emu_blas_build(v.ctypes.data, len(v), t.ctypes.data, len(t)))
    arr = (EmuInstance * len(bundle.instances))()
    for k, (mi, m, sbt, custom) in enumerate(bundle.instances):
        arr[k].t[:] = [float(x) for x in np.asarray(m, np.float32).reshape(16)]
        arr[k].sbt, arr[k].custom, arr[k].blas = sbt, custom, blas[mi]
    return L.emu_tlas_build(arr, len(bundle.instances))


def emu_trace(L, tlas, rays, first=False):
    out = np.zeros(len(rays), HIT)
    rays = np.ascontiguousarray(rays, np.float32)
    L.emu_trace(tlas, rays.ctypes.data, len(rays), int(first), out.ctypes.data)
    assert L.emu_tlas_status(tlas) == 0, "traversal stack overflow"
    return out


def cornell(scenes):
    b = scenes.cornell_like()
    b.camera["x"], b.camera["y"] = 0.013, 1.037
    return b


# ---- the oracle is pinned to the golden vectors -------------------------------------------------
def test_oracle_pcg3d_matches_golden(ref):
    g = np.load(os.path.join(GOLD, "pcg3d.npz"))
    assert np.array_equal(ref.pcg3d(g["input"]), g["output"])


def test_oracle_triangle_matches_golden(ref):
    g = np.load(os.path.join(GOLD, "triangle.npz"))
    for i in range(0, len(g["hit"]), 7):
        ok, r = ref.intersect_triangle(g["origin"][i], g["direction"][i], g["triangle"][i])
        assert ok == bool(g["hit"][i])
        if ok:
            assert np.array_equal(r, g["result"][i])
    assert list(g["hit"][:6]) == [1, 1, 1, 0, 0, 1]   # inside, edge, vertex, parallel, behind, back face


def test_oracle_traceray_matches_golden(ref, pkg):
    scenes = pkg.scenes
    g = np.load(os.path.join(GOLD, "cornell_hits.npz"))
    out = ref.build_scene(cornell(scenes)).trace(g["rays"])
    assert np.array_equal(out["hit"], g["hit"]) and np.array_equal(out["prim"], g["prim"])
    assert np.array_equal(out["t"], g["t"])


# ---- device logic on the host against the golden vectors ---------------------------------------
@pytest.mark.parametrize("name", ["cornell_hits.npz", "heightfield64_hits.npz"])
def test_device_logic_matches_golden_hits(emu, pkg, name):
    scenes = pkg.scenes
    g = np.load(os.path.join(GOLD, name))
    b = cornell(scenes) if name.startswith("cornell") else scenes.heightfield_scene(64, 160, 90)
    tl = emu_scene(emu, b)
    out = emu_trace(emu, tl, g["rays"])
    hit = out["t"] < 3e38
    assert np.array_equal(hit, g["hit"] == 1)
    bad = (out["prim"][hit] != g["prim"][hit]) | (out["inst"][hit] != g["inst"][hit])
    assert bad.mean() <= 1e-4
    ok = hit.copy()
    ok[hit] = ~bad
    assert (np.abs(out["t"][ok] - g["t"][ok]) / g["t"][ok]).max() <= 1e-5
    assert np.abs(out["b1"][ok] - g["bary"][ok, 1]).max() <= 1e-4
    occ = emu_trace(emu, tl, g["rays"], first=True)
    assert np.array_equal(occ["t"] < 3e38, g["occluded"] == 1)


def test_device_inverse_matches_golden(emu):
    g = np.load(os.path.join(GOLD, "inverse4x4.npz"))
    for m, inv in zip(g["input"], g["output"]):
        out = np.zeros(16, np.float32)
        m = np.ascontiguousarray(m, np.float32)
        assert emu.emu_inverse4x4(m.ctypes.data, out.ctypes.data) == 1
        assert np.allclose(out.reshape(4, 4), inv, rtol=2e-5, atol=2e-6)
    sing = np.zeros(16, np.float32)
    assert emu.emu_inverse4x4(sing.ctypes.data, out.ctypes.data) == 0


def test_device_logic_ragged_inputs(emu):
    # 1, 2, 3, 4 and 25 triangles (root leaf, partial node, one full leaf slot, first inner split)
    rng = np.random.default_rng(7)
    for nt in (1, 2, 3, 4, 25):
        v = rng.normal(size=(nt * 3, 3)).astype(np.float32)
        t = np.arange(nt * 3, dtype=np.uint32).reshape(nt, 3)
        blas = emu.emu_blas_build(v.ctypes.data, len(v), t.ctypes.data, nt)
        inst = (EmuInstance * 1)()
        inst[0].t[:] = [float(x) for x in np.eye(4, dtype=np.float32).reshape(16)]
        inst[0].blas = blas
        tl = emu.emu_tlas_build(inst, 1)
        c = v.reshape(nt, 3, 3).mean(axis=1)
        rays = np.zeros((nt, 8), np.float32)
        rays[:, 0:3] = c + np.array([0.01, 0.02, 5.0], np.float32)
        rays[:, 3], rays[:, 7] = 1e-3, 1e3
        rays[:, 4:7] = c - rays[:, 0:3]
        out = emu_trace(emu, tl, rays)
        assert (out["t"] < 3e38).all()          # every ray aims at a triangle centroid
        # brute force closest hit for comparison
        for i in range(nt):
            best = None
            for k in range(nt):
                a, b_, c_ = v[3 * k], v[3 * k + 1], v[3 * k + 2]
                e1, e2 = b_ - a, c_ - a
                p = np.cross(rays[i, 4:7], e2)
                det = e1 @ p
                if det == 0:
                    continue
                s = rays[i, 0:3] - a
                u = (s @ p) / det
                q = np.cross(s, e1)
                w = (rays[i, 4:7] @ q) / det
                tt = (e2 @ q) / det
                if u >= 0 and w >= 0 and u + w <= 1 and tt > 1e-3 and (best is None or tt < best[0]):
                    best = (tt, k)
            assert best is not None and out["prim"][i] == best[1]
            assert abs(out["t"][i] - best[0]) <= 1e-4 * best[0]


# ---- user programs compile for sm_100a without a device ----------------------------------------
def test_shipped_program_compiles_with_generated_dispatch(pkg):
    src, sbt = pkg.scenes.program_source("primary")
    size, log = pkg.rd.CompileCheck(src, sbt)
    assert size > 10000 and "raygen" in log


def test_reference_shader_cl_compiles_unmodified(pkg):
    path = "/root/reference/samples/shader.cl"
    if not os.path.exists(path):
        pytest.skip("reference sources not present on this machine")
    size, log = pkg.rd.CompileCheck(open(path).read(), None)
    assert size > 100000 and "raygen" in log


def test_reference_shader_cl_with_wavefront_hooks_compiles(pkg):
    """samples/shader.cl + programs/shader_wavefront_hooks.cl (the two functions a host appends): the raygen AND the generated
    wavefront stage kernels compile for sm_100a (the GPU test test_c1_shader_cl_in_wavefront_mode runs them)."""
    path = "/root/reference/samples/shader.cl"
    if not os.path.exists(path):
        pytest.skip("reference sources not present on this machine")
    hooks = open(os.path.join(os.path.dirname(pkg.scenes.__file__), "programs", "shader_wavefront_hooks.cl")).read()
    size, log = pkg.rd.CompileCheck(open(path).read() + "\n" + hooks, None)
    for kernel in ("raygen", "rd_wf_generate", "rd_wf_shade", "rd_wf_accumulate", "rd_wf_resolve"):
        assert kernel in log, kernel
    plain, _ = pkg.rd.CompileCheck(open(path).read(), None)
    assert size > plain


def test_compile_errors_and_bad_sbt_are_reported(pkg):
    rd = pkg.rd
    with pytest.raises(rd.RadianceError, match="compilation failed"):
        rd.CompileCheck('#include "radiance.cl"\n__kernel void raygen() { this is not C; }\n', "[]")
    with pytest.raises(rd.RadianceError, match="sbt.json"):
        rd.CompileCheck('#include "radiance.cl"\n__kernel void raygen() {}\n', '[{"raygen": 3}]')
    with pytest.raises(rd.RadianceError, match="no sbt.json"):
        rd.CompileCheck('#include "radiance.cl"\n__kernel void raygen() {}\n', None)


def test_library_kat_golden_is_the_reference(pkg):
    """tests/golden/library_kat.npz and generate_ray.npz reproduce from the reference's own functions (oracle/_ref)."""
    import __graft_entry__ as entry
    ref = entry.load_oracle()
    if not ref.available():
        pytest.skip("oracle/_ref not built")
    g = np.load(os.path.join(os.path.dirname(__file__), "golden", "library_kat.npz"))
    for name in ref.LIBRARY_KAT:
        out = ref.library_kat(name, g[name + "_in"])
        assert np.array_equal(out.view(np.uint32), g[name + "_out"].view(np.uint32)), name
    gr = np.load(os.path.join(os.path.dirname(__file__), "golden", "generate_ray.npz"))
    for k in range(0, len(gr["input"]), 17):
        cam = gr["cameras"][gr["camera_index"][k]]
        o, d = ref.generate_ray(np.array(cam).reshape(1), gr["input"][k], int(gr["input"][k][2]))
        assert np.array_equal(o, gr["origin"][k]) and np.array_equal(d, gr["direction"][k])


def test_restated_library_matches_reference_golden_on_host():
    """shader/math.cl, shader/pbr.cl and the traversal's triangle test, compiled for the host (tests/kat), on the
    golden vectors of the reference: integer / triangle results bit-exact, transcendental chains within 1e-6."""
    so = os.path.join(os.path.dirname(__file__), "kat", "libkat_host.so")
    if not os.path.exists(so):
        pytest.skip("tests/kat/libkat_host.so not built")
    L = C.CDLL(so)

    def kat(fn, a, stride):
        a = np.ascontiguousarray(a)
        hdr = np.array([fn, a.shape[0], a.shape[1], stride], np.uint32)
        out = np.zeros((a.shape[0], stride), np.float32)
        L.kat_run(C.c_void_p(hdr.ctypes.data), C.c_void_p(a.ctypes.data), C.c_void_p(out.ctypes.data))
        return out
    gold = os.path.join(os.path.dirname(__file__), "golden")
    p = np.load(os.path.join(gold, "pcg3d.npz"))
    assert np.array_equal(kat(0, p["input"].astype(np.uint32).view(np.float32), 3).view(np.uint32), p["output"].view(np.uint32))
    t = np.load(os.path.join(gold, "triangle.npz"))
    o = kat(2, np.hstack([t["origin"], t["direction"], t["triangle"]]).astype(np.float32), 5)
    hit = t["hit"].astype(bool)
    assert np.array_equal(o[:, 0] == 1, hit)
    assert np.array_equal(o[hit][:, 1].view(np.uint32), t["result"][hit][:, 0].view(np.uint32))
    assert np.array_equal(o[hit][:, 2:5].view(np.uint32), t["result"][hit][:, 4:7].view(np.uint32))
    g = np.load(os.path.join(gold, "library_kat.npz"))
    for name, fn in {"BRDF": 10, "refract": 11, "G_pbrt": 12, "sampleMicrofacetBRDF_transm": 13, "microfacetBRDF": 14,
                     "GetNormalSpace": 15, "sampleMicrofacetBRDF": 16}.items():
        exp = g[name + "_out"]
        out = kat(fn, g[name + "_in"], exp.shape[1])
        assert (np.abs(out - exp) / (1.0 + np.abs(exp))).max() <= 1e-6, name
