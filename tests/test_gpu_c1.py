"""Config C1 (BASELINE.json configs[0]): the reference's own user program -- samples/shader.cl with
samples/sbt.json, both unmodified -- compiled by NVRTC and launched on the B200 through RD::TraceRays, against the
same file run by the reference runtime on the host (oracle/_ref: NDRange emulation, local size 1).

The two source files are plugin INPUT, not product: oracle/Makefile copies them into the git-ignored
oracle/_ref/samples/ (which travels to the GPU box like the reference .so), nothing is committed.
Scene: SURVEY.md section 8(d) C1 -- Cornell-like box, one instance per mesh, 800x600, PhysicalCamera
{800, 600, 0.030, 0.036, 2.0, fStop}, one DirLight of colour 10, RTProp {0, batch, depth 8, 0}. One material is
transmissive (sampleMicrofacetBRDF_transm's refraction branch, pbr.cl:289-385), one metallic.

Bars: identical RNG streams on both sides (random_pcg3d is bit-exact), so images differ only where libm and
libdevice transcendentals (sin / cos / acos / pow) round differently and a path then takes another branch.
  * RGBA8 image RMSE <= C1_RMSE_U8 8-bit levels, mean level within 0.25
  * float imageScratch: relative L1 <= 1 %
The achieved values are recorded in gpurun_out/parity_r2.json (profiles/parity_r2.json)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

C1_RMSE_U8 = 2.5      # one sample per pixel: a path that flips at a refraction / shadow edge moves a whole pixel
C1_RMSE_U8_16 = 1.0   # 16 samples per pixel
W, H = 800, 600


@pytest.fixture(scope="module")
def c1_sources(ref):
    s = ref.sample_sources()
    if s is None:
        pytest.skip("oracle/_ref/samples/ not present (built by oracle/Makefile from /root/reference)")
    return s


def c1_bundle(scenes, fstop=0.0):
    b = scenes.cornell_like()
    b.camera["x"], b.camera["y"] = 0.013, 1.037   # keeps pixel rows off the exactly shared wall edges
    b.camera["fStop"] = fstop
    b.camera["focalDistance"] = 3.1               # the boxes are in focus, the back wall is not
    return b


def strip_dispatchers(src):
    """samples/shader.cl without its hand-written callAnyHit / callHit / callMiss (shader.cl:574-605): the runtime then
    generates them from sbt.json (tools/genSBT.py semantics)."""
    out = src
    for name in ("callAnyHit", "callHit", "callMiss"):
        a = out.index("void " + name + "(")
        depth, i = 0, out.index("{", a)
        while True:
            depth += {"{": 1, "}": -1}.get(out[i], 0)
            i += 1
            if depth == 0:
                break
        out = out[:a] + out[i:]
    assert "switch" not in out[out.index("void anyShadow"):]
    return out


def reference_frame(ref, scenes, b, rt, first=0, count=None):
    cam = np.array(b.camera).reshape(()).copy()
    cam["widthPixel"], cam["heightPixel"] = W, H
    n = W * H
    scratch, image = np.zeros(n * 4, np.float32), np.zeros(n * 4, np.uint8)
    tl = ref.build_scene(b)
    args = [np.array(rt).reshape(1), scratch, image, cam.reshape(1), np.array(b.scene).reshape(1), b.meshInfoData,
            b.vertexData, b.indexData, b.uvData, b.normalData, b.materials, None, 0, tl.blob]
    ref.raygen("shader", args, n, first, count)
    return image.reshape(H, W, 4), scratch.reshape(H, W, 4)


def gpu_frame(pkg, plt, b, rt, source, sbt):
    rd, scenes = pkg.rd, pkg.scenes
    tlas, _ = scenes.build_scene(plt, b)
    ds = scenes.DeviceScene(plt, b, tlas, W, H, rt)
    plt.set_option("max_trace_depth", 2)   # raygen -> material -> shadow ray
    pipe, module = scenes.make_pipeline(plt, source=source, sbt=sbt)
    rd.BindPipeline(plt, pipe)
    rd.BindDescriptorSet(plt, ds.descriptors)
    launches = plt.launch_count()
    rd.TraceRays(plt, 0, 0, 0, W, H)
    assert plt.launch_count() > launches
    return ds.read_image(), ds.read_scratch(), ds


def compare(img, scratch, exp, exp_scratch, bar, parity_log, name):
    d = img[..., :3].astype(np.float64) - exp[..., :3].astype(np.float64)
    rmse = float(np.sqrt((d * d).mean()))
    mean_diff = float(abs(img[..., :3].mean() - exp[..., :3].mean()))
    rel = float(np.abs(scratch[..., :3] - exp_scratch[..., :3]).mean() / np.abs(exp_scratch[..., :3]).mean())
    exact = float((d == 0).all(axis=-1).mean())
    parity_log.record(name, rmse_u8=rmse, mean_level_diff=mean_diff, scratch_rel_l1=rel, pixels_identical=exact,
                      bar_rmse_u8=bar, width=W, height=H)
    assert rmse <= bar, "%s: RMSE %g" % (name, rmse)
    assert mean_diff <= 0.25
    assert rel <= 0.01, "%s: scratch relative L1 %g" % (name, rel)
    assert (img[..., 3] == 255).all()
    return rmse


def test_c1_shader_cl_as_shipped_matches_reference(pkg, plt, ref, c1_sources, parity_log):
    """shader.cl + its own hand-written dispatchers, pinhole camera, 1 and 16 samples, depth 8."""
    scenes = pkg.scenes
    src, sbt = c1_sources
    b = c1_bundle(scenes)
    for batch, bar in ((1, C1_RMSE_U8), (16, C1_RMSE_U8_16)):
        rt = scenes.rtprop(0, batch, 8, 0)
        img, scratch, _ = gpu_frame(pkg, plt, b, rt, src, sbt)
        exp, exp_scratch = reference_frame(ref, scenes, b, rt)
        compare(img, scratch, exp, exp_scratch, bar, parity_log, "C1_shader_cl_%dspp" % batch)
        # the scene really exercises every branch: lit and shadowed surfaces, the sky, refraction
        assert len(np.unique(img[..., :3].reshape(-1, 3), axis=0)) > 2000


def test_c1_shader_cl_in_wavefront_mode(pkg, plt, ref, c1_sources, parity_log):
    """samples/shader.cl, unmodified, with the two wavefront hooks a host appends (programs/shader_wavefront_hooks.cl): the
    stage launches render the frame of the inline raygen (same float sequence per sample; the shade stage evaluates a path's
    hit shader in another order than a neighbour's, not another arithmetic) and the reference's frame within the C1 bar."""
    import os
    rd, scenes = pkg.rd, pkg.scenes
    src, sbt = c1_sources
    hooks = open(os.path.join(os.path.dirname(scenes.__file__), "programs", "shader_wavefront_hooks.cl")).read()
    b = c1_bundle(scenes)
    rt = scenes.rtprop(0, 16, 8, 0)
    inline_img, inline_scr, _ = gpu_frame(pkg, plt, b, rt, src, sbt)
    plt.set_option("wavefront", 1)
    try:
        wf_img, wf_scr, ds = gpu_frame(pkg, plt, b, rt, src + "\n" + hooks, sbt)
        assert plt.status() == (0, 0)
    finally:
        plt.set_option("wavefront", 0)
    d = wf_img[..., :3].astype(np.float64) - inline_img[..., :3].astype(np.float64)
    rmse_inline = float(np.sqrt((d * d).mean()))
    exp, exp_scratch = reference_frame(ref, scenes, b, rt)
    rmse = compare(wf_img, wf_scr, exp, exp_scratch, C1_RMSE_U8_16, parity_log, "C1_shader_cl_wavefront_16spp")
    parity_log.record("C1_shader_cl_wavefront_16spp", rmse_u8_vs_inline=rmse_inline)
    assert rmse_inline <= 0.5, "wavefront vs inline RMSE %g" % rmse_inline


def test_c1_generated_dispatchers_equal_handwritten(pkg, plt, ref, c1_sources):
    """The same file with callHit / callMiss / callAnyHit removed: generated from sbt.json, bit-identical frame."""
    scenes = pkg.scenes
    src, sbt = c1_sources
    b = c1_bundle(scenes)
    rt = scenes.rtprop(0, 2, 8, 0)
    a_img, a_scr, _ = gpu_frame(pkg, plt, b, rt, src, sbt)
    g_img, g_scr, _ = gpu_frame(pkg, plt, b, rt, strip_dispatchers(src), sbt)
    assert np.array_equal(a_img, g_img) and np.array_equal(a_scr, g_scr)


def test_c1_thin_lens_matches_reference(pkg, plt, ref, c1_sources, parity_log):
    """fStop != 0: the thin-lens branch of generateRay (shader.cl:155-171: sampleUniformDisk, Euler rotation of the
    lens point, focal plane)."""
    scenes = pkg.scenes
    src, sbt = c1_sources
    b = c1_bundle(scenes, fstop=0.05)   # lens radius 0.3: a strong blur away from the focal plane
    rt = scenes.rtprop(0, 4, 8, 0)
    img, scratch, _ = gpu_frame(pkg, plt, b, rt, src, sbt)
    exp, exp_scratch = reference_frame(ref, scenes, b, rt)
    compare(img, scratch, exp, exp_scratch, C1_RMSE_U8, parity_log, "C1_shader_cl_thin_lens_4spp")
    sharp, _ = reference_frame(ref, scenes, c1_bundle(scenes), rt)
    assert np.abs(exp[..., :3].astype(int) - sharp[..., :3].astype(int)).mean() > 2.0   # the lens really blurs


def test_c1_debug_view_and_progressive_accumulation(pkg, plt, ref, c1_sources, parity_log):
    """RTProp.debug = 1 (first segment only, no tone mapping: shader.cl:256-259, 289) and a second TraceRays call that
    continues the running mean from totalSamples = 4 (shader.cl:262-279)."""
    rd, scenes = pkg.rd, pkg.scenes
    src, sbt = c1_sources
    b = c1_bundle(scenes)
    rt = scenes.rtprop(0, 1, 8, 1)
    img, scratch, _ = gpu_frame(pkg, plt, b, rt, src, sbt)
    exp, exp_scratch = reference_frame(ref, scenes, b, rt)
    compare(img, scratch, exp, exp_scratch, C1_RMSE_U8, parity_log, "C1_shader_cl_debug")
    rt = scenes.rtprop(0, 4, 8, 0)
    img, scratch, ds = gpu_frame(pkg, plt, b, rt, src, sbt)
    ds.set_rtprop(4, 4, 8)
    rd.TraceRays(plt, 0, 0, 0, W, H)
    img2, scratch2 = ds.read_image(), ds.read_scratch()
    # the reference side: same two calls on the same scratch buffer
    cam = np.array(b.camera).reshape(()).copy()
    cam["widthPixel"], cam["heightPixel"] = W, H
    n = W * H
    rscratch, rimage = np.zeros(n * 4, np.float32), np.zeros(n * 4, np.uint8)
    tl = ref.build_scene(b)
    for total in (0, 4):
        args = [np.array(scenes.rtprop(total, 4, 8, 0)).reshape(1), rscratch, rimage, cam.reshape(1), np.array(b.scene).reshape(1),
                b.meshInfoData, b.vertexData, b.indexData, b.uvData, b.normalData, b.materials, None, 0, tl.blob]
        ref.raygen("shader", args, n)
    compare(img2, scratch2, rimage.reshape(H, W, 4), rscratch.reshape(H, W, 4), C1_RMSE_U8_16 * 1.5, parity_log,
            "C1_shader_cl_progressive_4+4spp")
