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


def build_sample():
    r = subprocess.run(["make", "-C", CPP], stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    assert r.returncode == 0, r.stdout
    return os.path.join(CPP, "sample_shim")


def test_shim_compiles_and_has_no_cpu_path(pkg):
    exe = build_sample()
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present: covered by the gpu test")
    r = subprocess.run([exe, os.path.join(PROG, "primary.cl"), os.path.join(PROG, "primary_sbt.json"), "/tmp/shim.rgba"],
                       stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    assert r.returncode != 0 and "no CPU path" in r.stdout   # aborts like CL_CHECK, never falls back


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
