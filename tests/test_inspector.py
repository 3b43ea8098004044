"""Headless progressive viewer (`inspector.renderLoop` + the `render` callback of samples/sample1.cpp:447-499):
accumulation over several TraceRays calls, the restart after an edit, PNG output, in inline and wavefront mode,
against the same call sequence on the reference runtime (oracle/_ref)."""
import importlib

import numpy as np
import pytest

PATH_RMSE_U8 = 2.0   # as tests/test_gpu_parity.py: libm vs libdevice transcendentals perturb sampled directions


def reference_sequence(ref, scenes, b, w, h, steps):
    """steps: list of (camera xyz or None) -- the sample1.cpp protocol replayed on the reference runtime."""
    cam = np.array(b.camera).reshape(()).copy()
    cam["widthPixel"], cam["heightPixel"] = w, h
    n = w * h
    scratch, image = np.zeros(n * 4, np.float32), np.zeros(n * 4, np.uint8)
    tl = ref.build_scene(b)
    rt = np.array(scenes.rtprop(0, 2, 3, 0)).reshape(1).copy()
    out = []
    for edit in steps:
        updated = edit is not None
        if updated:
            cam["x"], cam["y"], cam["z"] = edit
        args = [rt, scratch, image, cam.reshape(1), np.array(b.scene).reshape(1), b.meshInfoData, b.vertexData, b.indexData,
                b.uvData, b.normalData, b.materials, None, 0, tl.blob]
        ref.raygen("path", args, n)
        rt[0]["totalSamples"] = 0 if updated else rt[0]["totalSamples"] + rt[0]["batchSize"]
        out.append(image.reshape(h, w, 4).copy())
    return out


@pytest.mark.gpu
@pytest.mark.parametrize("wavefront", [0, 1])
def test_progressive_callback_sequence_matches_reference(pkg, plt, ref, tmp_path, wavefront):
    rd, scenes = pkg.rd, pkg.scenes
    insp = importlib.import_module("radiance_b200.inspector")
    b = scenes.cornell_like()
    b.camera["x"], b.camera["y"] = 0.013, 1.037
    tlas, _ = scenes.build_scene(plt, b)
    w, h = 160, 120
    ds = scenes.DeviceScene(plt, b, tlas, w, h, scenes.rtprop(0, 2, 3, 0))
    plt.set_option("wavefront", wavefront)
    try:
        pipe, _ = scenes.make_pipeline(plt, "path")
        rd.BindPipeline(plt, pipe)
        rd.BindDescriptorSet(plt, ds.descriptors)
        d = insp.CbData(plt, (w, h), ds.rdImage, ds.rdRTProp, ds.rdCam, ds.rdScene)
        moved = (0.2, 1.1, float(b.camera["z"]))
        steps = [None, None, moved, None, None]
        got = []

        def callback(data):
            k = data.frames
            if steps[k] is not None:
                data.set_camera(position=steps[k])
            image, ww, hh = insp.render(data)
            got.append(image.copy())
            return image, ww, hh
        pattern = str(tmp_path / "frame_%04d.png")
        assert insp.renderLoop(callback, d, frames=len(steps), out_pattern=pattern, every=2) == 0
    finally:
        plt.set_option("wavefront", 0)
    assert d.total_samples() == 4          # 0,2,4 -> edit -> 0,2,4
    exp = reference_sequence(ref, scenes, b, w, h, steps)
    for k, (g, e) in enumerate(zip(got, exp)):
        diff = g[..., :3].astype(np.float64) - e[..., :3].astype(np.float64)
        assert np.sqrt((diff * diff).mean()) <= PATH_RMSE_U8, (k, np.sqrt((diff * diff).mean()))
    assert not np.array_equal(got[1], got[3])     # the camera moved
    from PIL import Image
    for k in (0, 2, 4):
        assert np.array_equal(np.asarray(Image.open(pattern % k)), got[k])
    assert not (tmp_path / "frame_0001.png").exists()


def test_render_loop_rejects_wrong_extent(pkg):
    insp = importlib.import_module("radiance_b200.inspector")
    with pytest.raises(pkg.rd.RadianceError):
        insp.renderLoop(lambda d: (np.zeros((4, 4, 4), np.uint8), 8, 4), None)
