"""include/radiance.h: the namespace-RD C++ shim. A host program in the reference's own idiom
(tests/cpp/sample_shim.cpp, the flow of samples/sample1.cpp:363-490) builds against it; on the GPU its
image must equal the one the Python mirror renders for the same scene."""
import os
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CPP = os.path.join(ROOT, "tests", "cpp")
PROG = os.path.join(ROOT, "radiance-ray-tracing_b200", "programs")


def build_sample(name="sample_shim"):
    r = subprocess.run(["make", "-C", CPP], stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    assert r.returncode == 0, r.stdout
    return os.path.join(CPP, name)


def test_shim_compiles_and_has_no_cpu_path(pkg):
    exe = build_sample()
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present: covered by the gpu test")
    r = subprocess.run([exe, os.path.join(PROG, "primary.cl"), os.path.join(PROG, "primary_sbt.json"), "/tmp/shim.rgba"],
                       stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    assert r.returncode != 0 and "no CPU path" in r.stdout   # aborts like CL_CHECK, never falls back


def test_reference_sample1_cpp_builds_unchanged_against_the_shim(pkg, tmp_path):
    """the reference's OWN host program, samples/sample1.cpp (scene load, pipeline set-up, descriptor binding,
    TraceRays / ReadBuffer render loop, :363-548), compiled UNCHANGED against include/radiance.h + include/sceneBuilder.h +
    include/inspector.h (they carry the names of the headers it includes) and linked to libradiance_b200.so -- with the
    reference's vendored Dear ImGui core for its UI panel, which needs no graphics. Without a GPU the program stops at the
    platform ("no CPU path"); its asset paths are the author's, so it is a build test, not a run test."""
    ref = "/root/reference"
    src = os.path.join(ref, "samples", "sample1.cpp")
    if not os.path.exists(src):
        pytest.skip("reference sources not present on this machine")
    pkg.rd.lib()   # the library exists
    inc = ["-I", os.path.join(ROOT, "include"), "-I", os.path.join(ref, "external"), "-I", os.path.join(ref, "external", "imgui")]
    objs = []
    jobs = []
    for f, flags in [(src, ["-O1"])] + [(os.path.join(ref, "external", "imgui", n + ".cpp"), ["-O0"])
                                        for n in ("imgui", "imgui_draw", "imgui_tables", "imgui_widgets")]:
        o = str(tmp_path / (os.path.basename(f) + ".o"))
        objs.append(o)
        jobs.append(subprocess.Popen(["g++", "-std=c++17", "-w"] + flags + inc + ["-c", f, "-o", o], stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True))
    for j in jobs:
        out, _ = j.communicate()
        assert j.returncode == 0, out
    exe = str(tmp_path / "sample1")
    lib_dir = os.path.join(ROOT, "radiance-ray-tracing_b200")
    r = subprocess.run(["g++"] + objs + ["-o", exe, "-L", lib_dir, "-lradiance_b200", "-Wl,-rpath," + lib_dir],
                       stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    assert r.returncode == 0, r.stdout
    import torch
    if not torch.cuda.is_available():
        r = subprocess.run([exe], stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, cwd=str(tmp_path))
        assert "no CPU path" in r.stdout


def test_shim_declares_the_reference_api():
    text = open(os.path.join(ROOT, "include", "radiance.h")).read()
    for name in ("BuildAccelStruct", "TopAccelStructToFile", "FileToTopAccelStruct", "CreateBuffer", "CreateImage", "CreateImageArray",
                 "CreateSampler", "ReadImage", "WriteImage", "ReadBuffer", "WriteBuffer", "CreateDescriptorSet", "CreatePipelineLayout",
                 "CreateShaderModule", "CreatePipeline", "BindPipeline", "BindDescriptorSet", "TraceRays", "GetPlatform",
                 "read_kernel_file_str"):
        assert name + "(" in text, name


@pytest.mark.gpu
def test_cpp_sample_matches_python_mirror(pkg, plt, tmp_path):
    rd, scenes = pkg.rd, pkg.scenes
    exe = build_sample()
    out = str(tmp_path / "shim.rgba")
    w, h = 160, 120
    r = subprocess.run([exe, os.path.join(PROG, "primary.cl"), os.path.join(PROG, "primary_sbt.json"), out, str(w), str(h)],
                       stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    assert r.returncode == 0, r.stdout
    img_cpp = np.fromfile(out, np.uint8).reshape(h, w, 4)
    # the same program with the streamed read-back (RD::StreamBufferToHost) delivers the same bytes
    out2 = str(tmp_path / "shim_streamed.rgba")
    r = subprocess.run([exe, os.path.join(PROG, "primary.cl"), os.path.join(PROG, "primary_sbt.json"), out2, str(w), str(h)],
                       stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, env=dict(os.environ, RD_SAMPLE_STREAM="1"))
    assert r.returncode == 0, r.stdout
    assert np.array_equal(np.fromfile(out2, np.uint8).reshape(h, w, 4), img_cpp)
    # the same scene through the Python mirror
    lo, hi, t = -0.3, 0.3, 0.6
    bv = np.array([[lo, 0, lo], [lo, t, lo], [hi, t, lo], [hi, 0, lo], [lo, 0, hi], [hi, 0, hi], [hi, t, hi], [lo, t, hi],
                   [lo, 0, lo], [lo, 0, hi], [lo, t, hi], [lo, t, lo], [hi, 0, lo], [hi, t, lo], [hi, t, hi], [hi, 0, hi],
                   [lo, t, lo], [lo, t, hi], [hi, t, hi], [hi, t, lo]], np.float32)
    bt = np.array([[4 * q, 4 * q + 1, 4 * q + 2] for q in range(5)] + [[4 * q, 4 * q + 2, 4 * q + 3] for q in range(5)], np.uint32)
    bt = bt.reshape(2, 5, 3).transpose(1, 0, 2).reshape(-1, 3)
    floor = scenes.quad((-1, 0, -1), (1, 0, -1), (1, 0, 1), (-1, 0, 1))
    xf = np.eye(4, dtype=np.float32)
    xf[0, 3], xf[2, 3] = 0.2, -0.1
    mats = [scenes.material((0.7, 0.7, 0.7), 0.0, 0.8), scenes.material((0.8, 0.3, 0.2), 0.0, 0.5)]
    b = scenes.SceneBundle([floor, (bv, bt)], [(0, np.eye(4, dtype=np.float32), 0, 0), (1, xf, 0, 1)], mats, [0, 1])
    b.camera = scenes.camera(w, h, (0.0, 0.8, 2.6), (-0.2, 0.0, 0.0), 0.030, 0.036, 2.0, 0.0)
    b.scene = scenes.scene_properties((-0.4, -1.0, -0.3), (10.0, 10.0, 10.0))
    tlas, _ = scenes.build_scene(plt, b)
    ds = scenes.DeviceScene(plt, b, tlas, w, h, scenes.rtprop(0, 2, 1, 0))
    pipe, _ = scenes.make_pipeline(plt, "primary")
    rd.BindPipeline(plt, pipe)
    rd.BindDescriptorSet(plt, ds.descriptors)
    rd.TraceRays(plt, 0, 0, 0, w, h)
    ds.set_rtprop(2, 2, 1)
    rd.TraceRays(plt, 0, 0, 0, w, h)
    img_py = ds.read_image()
    assert img_cpp[..., :3].std() > 10
    assert np.array_equal(img_cpp, img_py)


def _asset(pkg, tmp_path):
    import sys
    sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
    import gltf_fixture
    doc, binary, expected = gltf_fixture.build_asset(pkg.scenes)
    glb = str(tmp_path / "scene.glb")
    gltf_fixture.write_glb(glb, doc, binary)
    return glb, expected


def test_cpp_scene_builder_rejects_a_broken_asset(pkg, tmp_path):
    """include/sceneBuilder.h aborts (prints, exit code != 0) on a truncated container. Without a GPU the platform
    singleton aborts first ("no CPU path"): either way the program must not continue."""
    exe = build_sample("sample_scene")
    glb, _ = _asset(pkg, tmp_path)
    bad = str(tmp_path / "short.glb")
    open(bad, "wb").write(open(glb, "rb").read()[:200])
    r = subprocess.run([exe, bad, os.path.join(PROG, "textured.cl"), os.path.join(PROG, "textured_sbt.json"), str(tmp_path / "x")],
                       stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    import torch
    assert r.returncode != 0 and ("RD::Scene::Load" if torch.cuda.is_available() else "no CPU path") in r.stdout, r.stdout


@pytest.mark.gpu
def test_cpp_scene_builder_and_render_loop_match_python(pkg, plt, tmp_path):
    """RD::Scene::Load (include/sceneBuilder.h) + renderLoop (include/inspector.h) in a C++ host program: the descriptor
    arrays equal the Python loader's, the rendered frame equals the Python mirror's, the PNG written by the loop
    decodes to the frame, the <scene>.cache blob renders the same image (checked inside the program)."""
    import importlib
    rd, scenes = pkg.rd, pkg.scenes
    sl = importlib.import_module("radiance_b200.scene_loader")
    exe = build_sample("sample_scene")
    glb, _ = _asset(pkg, tmp_path)
    prefix = str(tmp_path / "cpp")
    w, h, dim = 320, 240, 16            # dim = the PNG's own size: no resampling on either side
    r = subprocess.run([exe, glb, os.path.join(PROG, "textured.cl"), os.path.join(PROG, "textured_sbt.json"), prefix, str(dim), str(w), str(h)],
                       stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    assert r.returncode == 0, r.stdout
    p = sl.parse_gltf(glb)
    raw = open(prefix + ".arrays", "rb").read()
    at, arrays = 0, []
    while at < len(raw):
        n = int(np.frombuffer(raw, np.uint64, 1, at)[0])
        arrays.append(raw[at + 8: at + 8 + n])
        at += 8 + n
    info, verts, tris, uvs, normals, mats, inst = arrays
    assert info == p.meshInfoData.tobytes() and verts == p.vertexData.tobytes() and tris == p.indexData.tobytes()
    assert uvs == p.uvData.tobytes() and mats == p.materials.tobytes()
    assert np.allclose(np.frombuffer(normals, np.float32), p.normalData.reshape(-1), atol=2e-6)
    inst = np.frombuffer(inst, np.float32).reshape(-1, 18)
    assert len(inst) == len(p.instances)
    for row, (mesh, tf, sbt, custom) in zip(inst, p.instances):
        assert int(row[0]) == mesh and int(row[1]) == custom and np.allclose(row[2:].reshape(4, 4), tf, atol=1e-6)
    img_cpp = np.fromfile(prefix + ".rgba", np.uint8).reshape(h, w, 4)
    from PIL import Image
    assert np.array_equal(np.asarray(Image.open(prefix + "_1.png")), img_cpp)
    assert os.path.exists(prefix + "_0.png")
    # the same frames through the Python loader and mirror (two callbacks of one sample each)
    scene = sl.Scene.Load(glb, plt, tex_dim=dim)
    n = w * h

    def buf(arr):
        hnd = rd.CreateBuffer(plt, max(16, arr.nbytes))
        rd.WriteBuffer(plt, hnd, arr.nbytes, arr)
        return hnd
    image = rd.CreateImage(plt, w, h)
    rt = buf(np.array(scenes.rtprop(0, 1, 1, 0)).reshape(1))
    head = [rt, rd.CreateBuffer(plt, n * 16), image, buf(np.array(scenes.camera(w, h, (0.02, 0.9, 2.2), (-0.3, 0.0, 0.0))).reshape(1)),
            buf(np.array(scenes.scene_properties()).reshape(1))]
    pipe, _ = scenes.make_pipeline(plt, "textured")
    rd.BindPipeline(plt, pipe)
    rd.BindDescriptorSet(plt, rd.CreateDescriptorSet(head + scene.descriptors()))
    rd.TraceRays(plt, 0, 0, 0, w, h)
    rd.WriteBuffer(plt, rt, 16, np.array(scenes.rtprop(1, 1, 1, 0)).reshape(1))
    rd.TraceRays(plt, 0, 0, 0, w, h)
    img_py = np.zeros((h, w, 4), np.uint8)
    rd.ReadBuffer(plt, image, img_py.nbytes, img_py)
    assert img_cpp[..., :3].std() > 20
    d = img_cpp[..., :3].astype(np.float64) - img_py[..., :3].astype(np.float64)
    assert np.sqrt((d * d).mean()) <= 0.25 and (np.abs(d).max(axis=-1) > 2).mean() < 1e-3   # normals differ by float rounding only


@pytest.mark.gpu
def test_cpp_multigpu_host_through_the_c_abi():
    """tests/cpp/sample_multigpu.cpp: one host thread per visible GPU, rd_comm_init + rd_trace_rays_interleaved +
    rd_frame_gather and the dynamic rd_trace_rays_units + rd_frame_gather_units, all through include/radiance_b200.h.
    The program itself checks that both gathered frames equal the frame rank 0 renders alone."""
    import torch
    exe = build_sample("sample_multigpu")
    gpus = max(1, torch.cuda.device_count())
    r = subprocess.run([exe, os.path.join(PROG, "path.cl"), os.path.join(PROG, "path_sbt.json"), str(gpus), "320", "240"],
                       stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=300)
    assert r.returncode == 0, r.stdout
    assert "static partition identical, dynamic partition identical" in r.stdout
