"""GPU parity tests: the CUDA path (through the C ABI) against the reference's own code (oracle/_ref)
on the same seeded inputs. Bars (BASELINE.json north_star): hit / instance IDs bit-exact except a
<= 1e-4 fraction of near-tie rays, t within 1e-5 relative, images within a stated RMSE."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

ID_TIE_BUDGET = 1e-4
T_REL_TOL = 1e-5
IMAGE_RMSE_U8 = 1.0   # in 8-bit levels, equal sample counts and identical RNG streams


def compare_hits(hitA, hitB, exp, id_budget=ID_TIE_BUDGET, log=None, name=None, bary_tol=1e-4):
    hit_g = hitA["t"] < 3e38
    hit_r = exp["hit"] == 1
    mism = hit_g != hit_r
    both = hit_g & hit_r
    bad = (hitA["prim"][both] != exp["prim"][both]) | (hitB[both] != exp["inst"][both])
    ok = both.copy()
    ok[both] = ~bad
    rel = np.abs(hitA["t"][ok] - exp["t"][ok]) / exp["t"][ok] if ok.any() else np.zeros(1)
    db1 = np.abs(hitA["b1"][ok] - exp["bary"][ok, 1]).max() if ok.any() else 0.0
    db2 = np.abs(hitA["b2"][ok] - exp["bary"][ok, 2]).max() if ok.any() else 0.0
    if log is not None:
        # where IDs differ the two sides found different (coincident or near-tie) surfaces: how far apart in t?
        tie = np.abs(hitA["t"][both][bad] - exp["t"][both][bad]) / exp["t"][both][bad] if bad.any() else np.zeros(1)
        log.record(name, rays=len(mism), hits=int(both.sum()), hit_miss_mismatch_fraction=mism.mean(),
                   id_mismatch_fraction=bad.mean() if both.any() else 0.0, id_mismatch_max_rel_t_gap=tie.max(),
                   max_rel_t_err=rel.max(), t_bit_exact_fraction=(hitA["t"][ok] == exp["t"][ok]).mean() if ok.any() else 1.0,
                   max_bary_err=max(db1, db2), id_budget=id_budget, t_bar=T_REL_TOL, bary_bar=bary_tol)
    assert mism.mean() <= id_budget, "hit/miss disagreement on %d of %d rays" % (mism.sum(), len(mism))
    assert bad.mean() <= id_budget if both.any() else True, "ID mismatch fraction %g" % bad.mean()
    if ok.any():
        assert rel.max() <= T_REL_TOL
        # barycentrics come from the same expressions
        assert db1 <= bary_tol and db2 <= bary_tol
    return both.sum()


def offcentre_cornell(scenes):
    b = scenes.cornell_like()
    b.camera["x"], b.camera["y"] = 0.013, 1.037   # keeps pixel rows off the exactly shared wall edges
    return b


def test_cornell_ray_list_matches_reference(pkg, plt, ref, parity_log):
    rd, scenes = pkg.rd, pkg.scenes
    b = offcentre_cornell(scenes)
    tlas, _ = scenes.build_scene(plt, b)
    rays = scenes.pinhole_rays(b.camera, stride=2, jitter=(0.37, 0.61))
    hitA, hitB = rd.trace_host_rays(plt, tlas, rays)
    exp = ref.build_scene(b).trace(rays)
    assert compare_hits(hitA, hitB, exp, log=parity_log, name="C1_scene_ray_list") > 50000


def test_heightfield_ray_list_matches_reference(pkg, plt, ref):
    rd, scenes = pkg.rd, pkg.scenes
    b = scenes.heightfield_scene(224, 640, 360)   # 100 352 triangles
    tlas, blas = scenes.build_scene(plt, b)
    st = blas[0].stats()
    assert st["triangles"] == 2 * 224 * 224 and st["nodes"] > 1000
    rays = scenes.pinhole_rays(b.camera, jitter=(0.29, 0.71))
    hitA, hitB = rd.trace_host_rays(plt, tlas, rays)
    exp = ref.build_scene(b).trace(rays)
    assert compare_hits(hitA, hitB, exp) > 100000


def test_first_hit_mode_agrees_on_occlusion(pkg, plt, ref):
    rd, scenes = pkg.rd, pkg.scenes
    b = offcentre_cornell(scenes)
    tlas, _ = scenes.build_scene(plt, b)
    rays = scenes.pinhole_rays(b.camera, stride=4, jitter=(0.37, 0.61))
    hitA, _ = rd.trace_host_rays(plt, tlas, rays, rd.TRACE_TERMINATE_ON_FIRST_HIT)
    exp = ref.build_scene(b).trace(rays, terminate_on_first_hit=True)
    assert np.array_equal(hitA["t"] < 3e38, exp["hit"] == 1)
    assert (exp["anyhit_calls"][exp["hit"] == 1] >= 1).all()


def test_edge_cases(pkg, plt):
    rd, scenes = pkg.rd, pkg.scenes
    # single triangle, many instances of it, rays that miss everything, an empty ray list
    v = np.array([[0, 0, 0], [1, 0, 0], [0, 1, 0]], np.float32)
    t = np.array([[0, 1, 2]], np.uint32)
    blas = rd.BuildAccelStruct(plt, rd.Mesh(v, t))
    inst = [rd.Instance(scenes.translate(2.0 * i, 0, 0), 0, i, blas) for i in range(20)]
    tlas = rd.BuildAccelStruct(plt, inst)
    rays = np.zeros((20, 8), np.float32)
    rays[:, 0] = 2.0 * np.arange(20) + 0.25
    rays[:, 1], rays[:, 2], rays[:, 3] = 0.25, 1.0, 0.001
    rays[:, 6], rays[:, 7] = -1.0, 1000.0
    hitA, hitB = rd.trace_host_rays(plt, tlas, rays)
    assert np.array_equal(hitB, np.arange(20, dtype=np.uint32))
    assert np.allclose(hitA["t"], 1.0) and (hitA["prim"] == 0).all()
    rays[:, 6] = 1.0   # pointing away
    hitA, hitB = rd.trace_host_rays(plt, tlas, rays)
    assert (hitA["t"] > 3e38).all() and (hitB == 0xffffffff).all()
    hitA, hitB = rd.trace_host_rays(plt, tlas, np.zeros((0, 8), np.float32))
    assert len(hitA) == 0
    # Tmin / Tmax are exclusive (radiance/shader/radiance.cl:90-91)
    r = np.array([[0.25, 0.25, 1.0, 1.0, 0, 0, -1.0, 1000.0], [0.25, 0.25, 1.0, 0.001, 0, 0, -1.0, 1.0]], np.float32)
    hitA, _ = rd.trace_host_rays(plt, tlas, r)
    assert (hitA["t"] > 3e38).all()
    # empty mesh and empty instance list build and trace as all-miss
    eb = rd.BuildAccelStruct(plt, rd.Mesh(np.zeros((0, 3), np.float32), np.zeros((0, 3), np.uint32)))
    et = rd.BuildAccelStruct(plt, [rd.Instance(np.eye(4), 0, 0, eb)])
    hitA, _ = rd.trace_host_rays(plt, et, rays)
    assert (hitA["t"] > 3e38).all()
    et = rd.BuildAccelStruct(plt, [])
    hitA, _ = rd.trace_host_rays(plt, et, rays)
    assert (hitA["t"] > 3e38).all()


def reference_image(ref, scenes, b, program, width, height, rt):
    cam = np.array(b.camera).reshape(()).copy()
    cam["widthPixel"], cam["heightPixel"] = width, height
    n = width * height
    scratch, image = np.zeros(n * 4, np.float32), np.zeros(n * 4, np.uint8)
    tl = ref.build_scene(b)
    args = [np.array(rt).reshape(1), scratch, image, cam.reshape(1), np.array(b.scene).reshape(1), b.meshInfoData,
            b.vertexData, b.indexData, b.uvData, b.normalData, b.materials, None, 0, tl.blob]
    ref.raygen(program, args, n)
    return image.reshape(height, width, 4), scratch.reshape(height, width, 4)


def test_primary_program_image_matches_reference(pkg, plt, ref):
    rd, scenes = pkg.rd, pkg.scenes
    b = offcentre_cornell(scenes)
    tlas, _ = scenes.build_scene(plt, b)
    w, h = 320, 240
    launches = plt.launch_count()
    img = scenes.render_primary(plt, b, tlas, w, h, "primary", batch=2)
    assert plt.launch_count() > launches
    exp, _ = reference_image(ref, scenes, b, "primary", w, h, scenes.rtprop(0, 2, 1, 0))
    d = img[..., :3].astype(np.float64) - exp[..., :3].astype(np.float64)
    rmse = np.sqrt((d * d).mean())
    assert rmse <= IMAGE_RMSE_U8, "RMSE %g" % rmse
    assert (np.abs(d).max(axis=-1) > 8).mean() <= 1e-3   # isolated edge pixels only
    assert (img[..., 3] == 255).all()


PATH_RMSE_U8 = 2.0   # 8-bit levels; libm vs libdevice transcendentals perturb sampled directions by ULPs


def test_path_program_image_matches_reference(pkg, plt, ref):
    """programs/path.cl (bounce loop, nested shadow traceRay with any-hit termination, pbr.cl sampling):
    same program text through NVRTC on the GPU and against the reference runtime on the host."""
    rd, scenes = pkg.rd, pkg.scenes
    b = offcentre_cornell(scenes)
    tlas, _ = scenes.build_scene(plt, b)
    w, h = 200, 152
    rt = scenes.rtprop(0, 4, 4, 0)
    ds = scenes.DeviceScene(plt, b, tlas, w, h, rt)
    plt.set_option("count_rays", 1)
    try:
        pipe, _ = scenes.make_pipeline(plt, "path")
    finally:
        plt.set_option("count_rays", 0)
    rd.BindPipeline(plt, pipe)
    rd.BindDescriptorSet(plt, ds.descriptors)
    plt.ray_count()
    rd.TraceRays(plt, 0, 0, 0, w, h)
    rays = plt.ray_count()
    img, scratch = ds.read_image(), ds.read_scratch()
    exp, exp_scratch = reference_image(ref, scenes, b, "path", w, h, rt)
    # every sample traces a primary ray, every surface hit a shadow ray and (below the depth limit) a bounce ray
    assert w * h * 4 < rays <= w * h * 4 * 8
    d = img[..., :3].astype(np.float64) - exp[..., :3].astype(np.float64)
    rmse = np.sqrt((d * d).mean())
    assert rmse <= PATH_RMSE_U8, "RMSE %g" % rmse
    assert abs(img[..., :3].mean() - exp[..., :3].mean()) <= 0.25
    assert (np.abs(d).max(axis=-1) > 16).mean() <= 5e-3
    # the accumulated radiance agrees before tone mapping as well
    rel = np.abs(scratch[..., :3] - exp_scratch[..., :3]).mean() / np.abs(exp_scratch[..., :3]).mean()
    assert rel <= 0.01, "scratch relative L1 %g" % rel
    # progressive accumulation: a second call continues the running mean (samples/shader.cl:262-279)
    ds.set_rtprop(4, 4, 4)
    rd.TraceRays(plt, 0, 0, 0, w, h)
    s2 = ds.read_scratch()
    assert np.isfinite(s2).all() and not np.array_equal(s2, scratch)


def test_wavefront_mode_matches_inline_and_reference(pkg, plt, ref):
    """Wavefront mode (generate / trace / sort / shade / resolve stages over HBM queues) renders the same
    image as the inline raygen kernel of the same program, with the hit-group sort on and off, and both
    match the reference runtime on the host."""
    rd, scenes = pkg.rd, pkg.scenes
    b = offcentre_cornell(scenes)
    tlas, _ = scenes.build_scene(plt, b)
    w, h = 200, 152
    rt = scenes.rtprop(0, 3, 4, 0)
    ds = scenes.DeviceScene(plt, b, tlas, w, h, rt)
    pipe, _ = scenes.make_pipeline(plt, "path")
    rd.BindPipeline(plt, pipe)
    rd.BindDescriptorSet(plt, ds.descriptors)
    images = {}
    try:
        for mode, wf, srt in (("inline", 0, 1), ("sorted", 1, 1), ("unsorted", 1, 0)):
            plt.set_option("wavefront", wf)
            plt.set_option("wf_sort", srt)
            ds.set_rtprop(0, 3, 4)
            launches = plt.launch_count()
            rd.TraceRays(plt, 0, 0, 0, w, h)
            images[mode] = (ds.read_image(), ds.read_scratch(), plt.launch_count() - launches)
    finally:
        plt.set_option("wavefront", 0)
    assert images["inline"][2] <= 2 and images["sorted"][2] > images["unsorted"][2] > 3 * 4
    # the order in which paths are shaded cannot change any path
    assert np.array_equal(images["sorted"][1], images["unsorted"][1])
    assert np.array_equal(images["sorted"][0], images["unsorted"][0])
    d = images["sorted"][0][..., :3].astype(np.float64) - images["inline"][0][..., :3].astype(np.float64)
    assert np.sqrt((d * d).mean()) <= 0.5, "wavefront vs inline RMSE %g" % np.sqrt((d * d).mean())
    exp, _ = reference_image(ref, scenes, b, "path", w, h, rt)
    d = images["sorted"][0][..., :3].astype(np.float64) - exp[..., :3].astype(np.float64)
    assert np.sqrt((d * d).mean()) <= PATH_RMSE_U8
    # a program without the hooks refuses wavefront mode loudly
    pipe2, _ = scenes.make_pipeline(plt, "primary")
    rd.BindPipeline(plt, pipe2)
    plt.set_option("wavefront", 1)
    try:
        with pytest.raises(rd.RadianceError, match="wavefront"):
            rd.TraceRays(plt, 0, 0, 0, w, h)
    finally:
        plt.set_option("wavefront", 0)


def test_hit_group_dispatch_matches_reference(pkg, plt, ref):
    """programs/groups.cl (config C4: 64 hit groups, instance SBTOffset selects rows 1+k / 65+k of the
    131-row sbt.json): inline and wavefront (sorted / unsorted) images against the reference runtime."""
    rd, scenes = pkg.rd, pkg.scenes
    b = scenes.groups_scene(64, 8, 64, 192, 144)   # 64 instances, 8192 triangles, every group used once
    tlas, _ = scenes.build_scene(plt, b)
    w, h = 192, 144
    rt = scenes.rtprop(0, 2, 3, 0)
    ds = scenes.DeviceScene(plt, b, tlas, w, h, rt)
    pipe, _ = scenes.make_pipeline(plt, "groups")
    rd.BindPipeline(plt, pipe)
    rd.BindDescriptorSet(plt, ds.descriptors)
    out = {}
    try:
        for mode, wf, srt in (("inline", 0, 1), ("sorted", 1, 1), ("unsorted", 1, 0)):
            plt.set_option("wavefront", wf)
            plt.set_option("wf_sort", srt)
            ds.set_rtprop(0, 2, 3)
            rd.TraceRays(plt, 0, 0, 0, w, h)
            out[mode] = ds.read_image()
    finally:
        plt.set_option("wavefront", 0)
    assert np.array_equal(out["sorted"], out["unsorted"])
    exp, _ = reference_image(ref, scenes, b, "groups", w, h, rt)
    assert len(np.unique(exp[..., :3].reshape(-1, 3), axis=0)) > 500   # the groups really shade differently
    for mode in ("inline", "sorted"):
        d = out[mode][..., :3].astype(np.float64) - exp[..., :3].astype(np.float64)
        assert np.sqrt((d * d).mean()) <= PATH_RMSE_U8, "%s RMSE %g" % (mode, np.sqrt((d * d).mean()))


def test_reference_blob_import(pkg, plt, ref):
    """RD::FileToTopAccelStruct: the blob the reference's own builder + packers produce (oracle/_ref) loads
    into the native structure and traces to the same hits as the reference walking that blob."""
    rd, scenes = pkg.rd, pkg.scenes
    b = offcentre_cornell(scenes)
    tl = ref.build_scene(b)
    rays = scenes.pinhole_rays(b.camera, stride=3, jitter=(0.41, 0.27))
    exp = tl.trace(rays)
    tlas = rd.TopAccelStructFromBytes(plt, tl.blob)
    hitA, hitB = rd.trace_host_rays(plt, tlas, rays)
    assert compare_hits(hitA, hitB, exp) > 20000
    # truncated and foreign files are refused (the reference throws on a short read, radiance.cpp:458-459)
    with pytest.raises(rd.RadianceError):
        rd.TopAccelStructFromBytes(plt, tl.blob[: len(tl.blob) // 2])
    with pytest.raises(rd.RadianceError):
        rd.TopAccelStructFromBytes(plt, np.zeros(64, np.uint8))


def test_reference_blob_export_round_trip(pkg, plt, ref, tmp_path):
    """RD::TopAccelStructToFile: the exported file is a valid reference blob -- the plain-C port of the
    reference's traceRay (oracle/rd_oracle.c, pinned to the reference bit for bit) walks it to the hits the
    GPU finds, and loading it back gives the same hits again."""
    import importlib.util
    import os
    rd, scenes = pkg.rd, pkg.scenes
    spec = importlib.util.spec_from_file_location("rd_oracle_portpy", os.path.join(os.path.dirname(__file__), "..", "oracle", "portpy.py"))
    port = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(port)
    if not port.available():
        pytest.skip("oracle port not built")
    b = offcentre_cornell(scenes)
    tlas, _ = scenes.build_scene(plt, b)
    rays = scenes.pinhole_rays(b.camera, stride=3, jitter=(0.41, 0.27))
    hitA, hitB = rd.trace_host_rays(plt, tlas, rays)
    path = str(tmp_path / "scene.cache")
    rd.TopAccelStructToFile(plt, tlas, path)
    blob = np.fromfile(path, np.uint8)
    assert np.array_equal(blob, rd.TopAccelStructToBytes(plt, tlas))
    out = port.trace(blob, rays)
    assert compare_hits(hitA, hitB, out) > 20000
    hit = out["hit"] == 1
    assert np.array_equal(out["custom"][hit], np.array([i[3] for i in b.instances], np.uint32)[out["inst"][hit]])
    back = rd.FileToTopAccelStruct(plt, path)
    hitA2, hitB2 = rd.trace_host_rays(plt, back, rays)
    assert np.array_equal(hitA2["prim"], hitA["prim"]) and np.array_equal(hitB2, hitB) and np.array_equal(hitA2["t"], hitA["t"])


def test_chunk_stealing_renders_every_pixel_once(pkg, plt):
    """Persistent raygen with chunk stealing (multi-GPU partition, here with a private counter): same image as
    the plain launch, and the chunks it claimed are exactly all chunks, each once."""
    rd, scenes = pkg.rd, pkg.scenes
    b = offcentre_cornell(scenes)
    tlas, _ = scenes.build_scene(plt, b)
    w, h = 320, 240
    ds = scenes.DeviceScene(plt, b, tlas, w, h, scenes.rtprop(0, 2, 3, 0))
    plt.set_option("stealing", 1)
    try:
        pipe, _ = scenes.make_pipeline(plt, "path")
    finally:
        plt.set_option("stealing", 0)
    rd.BindPipeline(plt, pipe)
    rd.BindDescriptorSet(plt, ds.descriptors)
    rd.TraceRays(plt, 0, 0, 0, w, h)
    plain = ds.read_image()
    rd.WriteBuffer(plt, ds.rdImage, plain.nbytes, np.zeros_like(plain))
    for frame in (0, 1, 700):
        rd.TraceRaysSteal(plt, w, h, frame, 512)
        owned = rd.StealOwned(plt)
        assert np.array_equal(np.sort(owned), np.arange(w * h // 512))
    assert np.array_equal(ds.read_image(), plain)


def test_interleaved_partition_renders_every_pixel_once(pkg, plt):
    """rd_trace_rays_interleaved: the bands of all ranks together are the plain frame (ranks run one after the
    other on this GPU into one framebuffer), and one rank alone leaves the other ranks' rows untouched."""
    rd, scenes, mg = pkg.rd, pkg.scenes, pkg.multigpu
    b = offcentre_cornell(scenes)
    tlas, _ = scenes.build_scene(plt, b)
    w, h = 320, 240
    ds = scenes.DeviceScene(plt, b, tlas, w, h, scenes.rtprop(0, 1, 2, 0))
    pipe, _ = scenes.make_pipeline(plt, "path")
    rd.BindPipeline(plt, pipe)
    rd.BindDescriptorSet(plt, ds.descriptors)
    rd.TraceRays(plt, 0, 0, 0, w, h)
    plain = ds.read_image()
    for world, band in ((2, 1), (3, 2), (8, 1), (4, 7)):
        rd.WriteBuffer(plt, ds.rdImage, plain.nbytes, np.zeros_like(plain))
        rd.TraceRaysInterleaved(plt, w, h, world - 1, world, band)
        plt.sync()
        part = ds.read_image()
        mine = np.zeros(h, bool)
        for y0, ny in mg.interleaved_rows(world - 1, world, h, band):
            mine[y0:y0 + ny] = True
        assert np.array_equal(part[mine], plain[mine]) and (part[~mine] == 0).all()
        for r in range(world - 1):
            rd.TraceRaysInterleaved(plt, w, h, r, world, band)
        plt.sync()
        assert np.array_equal(ds.read_image(), plain)
    # the same partition through the wavefront stages
    plt.set_option("wavefront", 1)
    try:
        ds.set_rtprop(0, 1, 2)
        rd.TraceRays(plt, 0, 0, 0, w, h)
        whole = ds.read_image()
        rd.WriteBuffer(plt, ds.rdImage, plain.nbytes, np.zeros_like(plain))
        for r in range(3):
            ds.set_rtprop(0, 1, 2)
            rd.TraceRaysInterleaved(plt, w, h, r, 3, 2)
        plt.sync()
        assert np.array_equal(ds.read_image(), whole)
    finally:
        plt.set_option("wavefront", 0)


def test_textured_material_matches_reference_runtime(pkg, plt, ref):
    """RD::CreateImageArray / WriteImage / CreateSampler + read_imageui (samples/shader2.cl:255-265): layered RGBA8,
    REPEAT + LINEAR, normalised coordinates, uv outside [0,1]. GPU image against the same program built against the
    reference runtime on the host; WriteImage / ReadImage round trip per layer."""
    import ctypes as C
    rd, scenes = pkg.rd, pkg.scenes
    b = offcentre_cornell(scenes)
    for k in range(len(b.materials)):
        b.materials[k]["albedoTex"] = k % 2          # every mesh textured, two layers
    b.uvData[:, :2] *= 1.7                            # coordinates leave [0,1]: REPEAT wraps them
    tw, th, layers = 32, 16, 2
    yy, xx = np.meshgrid(np.arange(th), np.arange(tw), indexing="ij")
    tex = np.zeros((layers, th, tw, 4), np.uint8)
    tex[0, ..., 0], tex[0, ..., 1], tex[0, ..., 2] = (xx * 8) % 256, (yy * 16) % 256, ((xx ^ yy) & 1) * 255
    tex[1, ..., 0], tex[1, ..., 1], tex[1, ..., 2] = 255 - (yy * 16) % 256, ((xx // 4 + yy // 4) & 1) * 200 + 30, (xx * 5 + yy * 3) % 256
    tex[..., 3] = 255
    tlas, _ = scenes.build_scene(plt, b)
    w, h = 256, 192
    ds = scenes.DeviceScene(plt, b, tlas, w, h, scenes.rtprop(0, 1, 1, 0))
    ds.rdTextures = rd.CreateImageArray(plt, tw, th, layers)
    for layer in range(layers):
        rd.WriteImage(plt, ds.rdTextures, tw, th, layer, tex[layer])
    back = np.zeros((th, tw, 4), np.uint8)
    rd.ReadImage(plt, ds.rdTextures, tw, th, 1, back)
    assert np.array_equal(back, tex[1])
    ds.descriptors[11] = ds.rdTextures
    pipe, _ = scenes.make_pipeline(plt, "textured")
    rd.BindPipeline(plt, pipe)
    rd.BindDescriptorSet(plt, ds.descriptors)
    rd.TraceRays(plt, 0, 0, 0, w, h)
    img = ds.read_image()

    class Desc(C.Structure):   # rd_image_array_desc of shader/compat/cl_compat.h (host flavour)
        _fields_ = [("texels", C.c_void_p), ("width", C.c_uint), ("height", C.c_uint), ("layers", C.c_uint), ("pad", C.c_uint)]
    flat = np.ascontiguousarray(tex)
    desc = Desc(flat.ctypes.data, tw, th, layers, 0)
    sampler = rd.RD_ADDRESS_REPEAT | (rd.RD_FILTER_LINEAR << 16) | (1 << 32)
    cam = np.array(b.camera).reshape(()).copy()
    cam["widthPixel"], cam["heightPixel"] = w, h
    n = w * h
    scratch, image = np.zeros(n * 4, np.float32), np.zeros(n * 4, np.uint8)
    tl = ref.build_scene(b)
    args = [np.array(scenes.rtprop(0, 1, 1, 0)).reshape(1), scratch, image, cam.reshape(1), np.array(b.scene).reshape(1),
            b.meshInfoData, b.vertexData, b.indexData, b.uvData, b.normalData, b.materials, C.addressof(desc), sampler, tl.blob]
    ref.raygen("textured", args, n)
    exp = image.reshape(h, w, 4)
    d = img[..., :3].astype(np.float64) - exp[..., :3].astype(np.float64)
    assert np.sqrt((d * d).mean()) <= IMAGE_RMSE_U8, "RMSE %g" % np.sqrt((d * d).mean())
    assert img[..., :3].std() > 30 and len(np.unique(img[..., :3].reshape(-1, 3), axis=0)) > 2000   # the texture is visible


def test_ragged_frame_sizes_in_every_launch_mode(pkg, plt):
    """Frame sizes that are not multiples of the 8x4 tile or of the block: inline (full + tail launch), wavefront
    and chunk-stealing launches must all render the same image; a zero-sized frame is a no-op."""
    rd, scenes = pkg.rd, pkg.scenes
    b = offcentre_cornell(scenes)
    tlas, _ = scenes.build_scene(plt, b)
    plt.set_option("stealing", 1)
    try:
        pipe, _ = scenes.make_pipeline(plt, "path")
    finally:
        plt.set_option("stealing", 0)
    for w, h in ((203, 151), (8, 4), (1, 1), (129, 3)):
        ds = scenes.DeviceScene(plt, b, tlas, w, h, scenes.rtprop(0, 2, 3, 0))
        rd.BindPipeline(plt, pipe)
        rd.BindDescriptorSet(plt, ds.descriptors)
        rd.TraceRays(plt, 0, 0, 0, w, h)
        inline = ds.read_image()
        assert (inline[..., 3] == 255).all()
        plt.set_option("wavefront", 1)
        try:
            ds.set_rtprop(0, 2, 3)
            rd.TraceRays(plt, 0, 0, 0, w, h)
        finally:
            plt.set_option("wavefront", 0)
        wave = ds.read_image()
        d = wave[..., :3].astype(np.float64) - inline[..., :3].astype(np.float64)
        assert np.sqrt((d * d).mean()) <= 0.5, (w, h)
        rd.WriteBuffer(plt, ds.rdImage, inline.nbytes, np.zeros_like(inline))
        ds.set_rtprop(0, 2, 3)
        rd.TraceRaysSteal(plt, w, h, 900 + w, 64)
        plt.sync()
        assert np.array_equal(ds.read_image(), inline), (w, h)
    rd.TraceRays(plt, 0, 0, 0, 0, 0)
    # an empty scene renders the miss shader everywhere in both modes
    et = rd.BuildAccelStruct(plt, [])
    ds = scenes.DeviceScene(plt, b, et, 64, 32, scenes.rtprop(0, 1, 2, 0))
    rd.BindDescriptorSet(plt, ds.descriptors)
    rd.TraceRays(plt, 0, 0, 0, 64, 32)
    sky = ds.read_image()
    plt.set_option("wavefront", 1)
    try:
        rd.TraceRays(plt, 0, 0, 0, 64, 32)
    finally:
        plt.set_option("wavefront", 0)
    assert np.array_equal(ds.read_image(), sky) and sky[..., 2].min() > 100


def _brute_force_hits(verts, tris, rays, tmin=0.001, tmax=1000.0):
    """Moeller-Trumbore over every triangle in float64 (numpy), reference accept rule (radiance.cl:221-240, :90-91)."""
    v0, v1, v2 = (verts[tris[:, k]].astype(np.float64) for k in range(3))
    e1, e2 = v1 - v0, v2 - v0
    out = []
    for r in rays:
        o, d = r[0:3].astype(np.float64), r[4:7].astype(np.float64)
        p = np.cross(d, e2)
        det = (e1 * p).sum(1)
        ok = det != 0
        inv = np.where(ok, 1.0 / np.where(ok, det, 1.0), 0.0)
        s = o - v0
        u = inv * (s * p).sum(1)
        q = np.cross(s, e1)
        v = inv * (q @ d)
        t = inv * (e2 * q).sum(1)
        hit = ok & (u >= 0) & (u <= 1) & (v >= 0) & (u + v <= 1) & (t > tmin) & (t < tmax)
        if hit.any():
            k = np.where(hit)[0][np.argmin(t[hit])]
            out.append((int(k), float(t[k])))
        else:
            out.append((-1, np.inf))
    return out


def test_c2_full_size_hits_match_reference(pkg, plt, ref, parity_log):
    """BASELINE configs[1] at full size (1 002 528 triangles, the bench scene): 2^16 of the frame's own primary rays
    against the reference's traceRay over the reference's own SAH BVH."""
    rd, scenes = pkg.rd, pkg.scenes
    b = scenes.heightfield_scene(708, 1920, 1080)
    tlas, blas = scenes.build_scene(plt, b)
    rays = scenes.pinhole_rays(b.camera, jitter=(0.37, 0.61))
    rng = np.random.default_rng(7)
    rays = rays[rng.choice(len(rays), 1 << 16, replace=False)]
    hitA, hitB = rd.trace_host_rays(plt, tlas, rays)
    exp = ref.build_scene(b).trace(rays)
    assert compare_hits(hitA, hitB, exp, log=parity_log, name="C2_full_size_primary_rays") > 30000


def test_c3_full_size_properties(pkg, plt):
    """BASELINE configs[2] scene at full size (10 069 778 triangles, 4 instances): properties that need no 5-minute
    reference build -- brute force over all triangles for a few rays, closest-hit vs first-hit consistency,
    static vs dynamic-fetch kernels, determinism."""
    rd, scenes = pkg.rd, pkg.scenes
    b = scenes.blob_field_scene(2237, 1920, 1080)
    tlas, blas = scenes.build_scene(plt, b)
    assert blas[0].stats()["triangles"] == 2 * 2237 * 2237
    rays = scenes.pinhole_rays(b.camera, jitter=(0.37, 0.61))
    rng = np.random.default_rng(11)
    rays = rays[rng.choice(len(rays), 1 << 15, replace=False)]
    hitA, hitB = rd.trace_host_rays(plt, tlas, rays)
    hitA2, hitB2 = rd.trace_host_rays(plt, tlas, rays, rd.TRACE_DYNAMIC_FETCH)
    assert np.array_equal(hitA["t"], hitA2["t"]) and np.array_equal(hitA["prim"], hitA2["prim"]) and np.array_equal(hitB, hitB2)
    first, _ = rd.trace_host_rays(plt, tlas, rays, rd.TRACE_TERMINATE_ON_FIRST_HIT)
    assert np.array_equal(first["t"] < 3e38, hitA["t"] < 3e38)          # occluded <=> a closest hit exists
    assert (first["t"][first["t"] < 3e38] >= hitA["t"][first["t"] < 3e38]).all()
    # brute force: rays that hit the height-field instance (identity transform), all 10 M triangles each
    hv, ht = b.meshes[0]
    pick = np.where((hitB == 0) & (hitA["t"] < 3e38))[0][:12]
    for i, (k, t) in zip(pick, _brute_force_hits(hv, ht, rays[pick])):
        assert k == int(hitA["prim"][i]) and abs(t - hitA["t"][i]) <= 1e-5 * t, (i, k, t, hitA[i])
    miss = np.where(hitA["t"] > 3e38)[0][:4]
    for k, t in _brute_force_hits(hv, ht, rays[miss]):
        assert k == -1


def test_wavefront_two_lanes_render_the_single_lane_frame(pkg, plt):
    """Option wf_lanes = 2..4: the passes of a frame go round robin over that many queue sets on as many streams (stage tails of one pass overlap
    the other pass's stages); the running mean is still taken in frame order, so scratch and image are bit-identical with the
    single-lane frame -- for an even and an odd number of passes, with the hit-group sort on and off, and when a second call
    continues the accumulation."""
    rd, scenes = pkg.rd, pkg.scenes
    b = offcentre_cornell(scenes)
    tlas, _ = scenes.build_scene(plt, b)
    w, h = 168, 100
    ds = scenes.DeviceScene(plt, b, tlas, w, h, scenes.rtprop(0, 5, 4, 0))
    pipe, _ = scenes.make_pipeline(plt, "path")
    rd.BindPipeline(plt, pipe)
    rd.BindDescriptorSet(plt, ds.descriptors)
    plt.set_option("wavefront", 1)
    try:
        for paths, srt in ((1 << 25, 1), (2 * w * h, 1), (w * h, 0)):
            got = {}
            for lanes in (1, 2, 3, 4):
                plt.set_option("wf_lanes", lanes)
                plt.set_option("wf_paths", paths)
                plt.set_option("wf_sort", srt)
                ds.set_rtprop(0, 5, 4)
                rd.TraceRays(plt, 0, 0, 0, w, h)
                first = ds.read_scratch()
                ds.set_rtprop(5, 5, 4)
                rd.TraceRays(plt, 0, 0, 0, w, h)
                got[lanes] = (first, ds.read_scratch(), ds.read_image())
                assert plt.status() == (0, 0)
            for lanes in (2, 3, 4):
                for a, c in zip(got[1], got[lanes]):
                    assert np.array_equal(a, c), "%d-lane frame differs (wf_paths %d, sort %d)" % (lanes, paths, srt)
    finally:
        plt.set_option("wavefront", 0)
        plt.set_option("wf_lanes", 2)   # the default
        plt.set_option("wf_paths", 1 << 26)
        plt.set_option("wf_sort", 1)


def test_wavefront_samples_in_flight_do_not_change_the_image(pkg, plt):
    """Option wf_paths: 1 sample per pass, 2 of 5 (ragged last pass), all 5 at once -- the float scratch and the image
    are bit-identical, also when the frame continues an accumulation (totalSamples > 0)."""
    rd, scenes = pkg.rd, pkg.scenes
    b = offcentre_cornell(scenes)
    tlas, _ = scenes.build_scene(plt, b)
    w, h = 168, 100
    ds = scenes.DeviceScene(plt, b, tlas, w, h, scenes.rtprop(0, 5, 4, 0))
    pipe, _ = scenes.make_pipeline(plt, "path")
    rd.BindPipeline(plt, pipe)
    rd.BindDescriptorSet(plt, ds.descriptors)
    out = []
    plt.set_option("wavefront", 1)
    plt.set_option("wf_lanes", 1)   # one lane: the pass counts asserted below are those of a single queue set
    try:
        for paths in (1, 2 * w * h, 1 << 25):
            plt.set_option("wf_paths", paths)
            ds.set_rtprop(0, 5, 4)
            l0 = plt.launch_count()
            rd.TraceRays(plt, 0, 0, 0, w, h)
            first = ds.read_scratch()
            ds.set_rtprop(5, 5, 4)                      # second call continues the running mean
            rd.TraceRays(plt, 0, 0, 0, w, h)
            out.append((first, ds.read_scratch(), ds.read_image(), plt.launch_count() - l0))
    finally:
        plt.set_option("wavefront", 0)
        plt.set_option("wf_lanes", 2)
        plt.set_option("wf_paths", 1 << 26)
    for first, second, image, _ in out[1:]:
        assert np.array_equal(first, out[0][0]) and np.array_equal(second, out[0][1]) and np.array_equal(image, out[0][2])
    assert out[0][3] > out[1][3] > out[2][3]            # 5, 3 and 1 passes per call
    assert not np.array_equal(out[0][0], out[0][1])
    # option trim_pool drops the queues and the build scratch; the next frame allocates them again and is unchanged
    plt.set_option("trim_pool", 1)
    plt.set_option("wavefront", 1)
    try:
        ds.set_rtprop(0, 5, 4)
        rd.TraceRays(plt, 0, 0, 0, w, h)
        assert np.array_equal(ds.read_scratch(), out[0][0])
    finally:
        plt.set_option("wavefront", 0)


def test_build_sort_and_scan_match_numpy(pkg, plt):
    """csrc/rd_sort.h: stable LSD radix sort of (u64 key, u32 value) pairs and the exclusive scan, against numpy, at
    sizes on both sides of the single-block / multi-block switch, with heavy key duplication (stability) and with key
    bytes that are constant (skipped passes)."""
    rd = pkg.rd
    rng = np.random.default_rng(5)
    for n in (0, 1, 2, 31, 1000, 8192, 8193, 100003, 1 << 20):
        for kind in ("random63", "dups", "lowbits"):
            if kind == "random63":
                keys = rng.integers(0, 1 << 63, n, dtype=np.uint64)
            elif kind == "dups":
                keys = rng.integers(0, 37, n, dtype=np.uint64) << np.uint64(20)
            else:
                keys = rng.integers(0, 1 << 16, n, dtype=np.uint64) | np.uint64(0x1200000000000000)
            vals = np.arange(n, dtype=np.uint32)
            k, v = rd.debug_sort_pairs(plt, keys, vals, 63)
            order = np.argsort(keys, kind="stable")
            assert np.array_equal(k, keys[order]), (n, kind)
            assert np.array_equal(v, vals[order]), (n, kind)          # stable: equal keys keep their input order
    for n in (0, 1, 255, 2048, 2049, 70001, 5000000):
        d = rng.integers(0, 5, n, dtype=np.uint32)
        out, total = rd.debug_exclusive_scan(plt, d)
        exp = np.concatenate([[0], np.cumsum(d, dtype=np.uint64)[:-1]]).astype(np.uint32) if n else d
        assert np.array_equal(out, exp) and total == int(d.sum()), n


def test_build_is_deterministic_and_keeps_its_shape(pkg, plt):
    """Same mesh, two builds: identical statistics (stable sort, deterministic clustering up to the order in which inner
    ids are handed out); the C2-size numbers are pinned so that a change of the builder shows up here."""
    scenes = pkg.scenes
    b = scenes.heightfield_scene(224, 640, 360)
    s1 = scenes.build_scene(plt, b)[1][0].stats()
    s2 = scenes.build_scene(plt, b)[1][0].stats()
    assert s1 == s2 and s1["triangles"] == 2 * 224 * 224
    assert 0.08 * s1["triangles"] < s1["nodes"] < 0.13 * s1["triangles"]      # ~0.10 wide nodes per triangle (PLOC + DP collapse)


def test_streamed_readback_delivers_the_same_frame(pkg, plt):
    """rd_buffer_stream_to_host: a frame copied out band by band while it renders must arrive byte-identical to
    TraceRays + ReadBuffer, for band counts that do and do not divide the tile rows, frame after frame, and must
    fall back to the plain read for frame sizes the banding does not cover or a different destination."""
    import torch
    rd, scenes = pkg.rd, pkg.scenes
    b = scenes.heightfield_scene(96, 640, 360)
    tlas, _ = scenes.build_scene(plt, b)
    plt.set_option("stream_image", 1)
    try:
        pipe, module = scenes.make_pipeline(plt, "primary")
    finally:
        plt.set_option("stream_image", 0)
    for (w, h), bands in (((640, 360), 4), ((640, 360), 7), ((640, 360), 64), ((256, 64), 1), ((203, 151), 4)):
        ds = scenes.DeviceScene(plt, b, tlas, w, h, scenes.rtprop(0, 1, 1, 0))
        rd.BindPipeline(plt, pipe)
        rd.BindDescriptorSet(plt, ds.descriptors)
        rd.TraceRays(plt, 0, 0, 0, w, h)
        plain = ds.read_image()
        assert (plain[..., 3] == 255).all()
        host = torch.zeros(h * w * 4, dtype=torch.uint8).pin_memory()
        rd.StreamBufferToHost(plt, ds.rdImage, host.data_ptr(), 4, bands)
        try:
            for frame in range(3):
                host.fill_(7)
                rd.WriteBuffer(plt, ds.rdImage, plain.nbytes, np.zeros_like(plain))
                launches = plt.launch_count()
                rd.TraceRays(plt, 0, 0, 0, w, h)
                assert plt.launch_count() - launches == (1 if (4 * w) % 128 == 0 and h % 4 == 0 else 2)
                rd.ReadBuffer(plt, ds.rdImage, plain.nbytes, host.data_ptr())
                assert np.array_equal(host.numpy().reshape(h, w, 4), plain), (w, h, bands, frame)
            # a read into other memory right after a streamed frame is the plain copy
            rd.TraceRaysAsync(plt, w, h)
            assert np.array_equal(ds.read_image(), plain)
            # two frames back to back without a read in between: the second waits for the first one's copies
            rd.TraceRaysAsync(plt, w, h)
            rd.TraceRaysAsync(plt, w, h)
            rd.ReadBuffer(plt, ds.rdImage, plain.nbytes, host.data_ptr())
            assert np.array_equal(host.numpy().reshape(h, w, 4), plain)
        finally:
            rd.StreamBufferToHost(plt, None, None)
    # a module created without the option keeps working with a registered mirror (plain launch + plain read)
    pipe2, _ = scenes.make_pipeline(plt, "primary")
    ds = scenes.DeviceScene(plt, b, tlas, 640, 360, scenes.rtprop(0, 1, 1, 0))
    rd.BindPipeline(plt, pipe2)
    rd.BindDescriptorSet(plt, ds.descriptors)
    host = torch.zeros(360 * 640 * 4, dtype=torch.uint8).pin_memory()
    rd.StreamBufferToHost(plt, ds.rdImage, host.data_ptr(), 4, 4)
    try:
        rd.TraceRays(plt, 0, 0, 0, 640, 360)
        rd.ReadBuffer(plt, ds.rdImage, host.numel(), host.data_ptr())
    finally:
        rd.StreamBufferToHost(plt, None, None)
    assert np.array_equal(host.numpy().reshape(360, 640, 4), ds.read_image())


def test_fused_small_build_is_the_multi_launch_small_build(pkg, plt):
    """Meshes of up to 8192 triangles are built by ONE launch (csrc/rd_build.cu, rd_k_small_build_fused): the same device
    functions as the multi-launch small path (developer knob RD_BUILD_UNFUSED), so the structure must be the same tree --
    compared through node / level counts, the size of the exported blob and the hits of a ray list; also built from device arrays."""
    import os
    rd, scenes = pkg.rd, pkg.scenes
    rng = np.random.default_rng(11)
    for nt in (1, 2, 7, 300, 2450, 8192):
        v = rng.normal(size=(nt * 3, 3)).astype(np.float32)
        t = rng.permutation(nt * 3).astype(np.uint32).reshape(nt, 3)
        rays = np.zeros((256, 8), np.float32)
        rays[:, 0:3] = rng.normal(size=(256, 3)) * 3.0
        rays[:, 4:7] = v[rng.integers(0, nt * 3, 256)] - rays[:, 0:3]
        rays[:, 3], rays[:, 7] = 1e-3, 1e3
        got = {}
        for mode in ("fused", "unfused", "fused_device"):
            os.environ.pop("RD_BUILD_UNFUSED", None)
            if mode == "unfused":
                os.environ["RD_BUILD_UNFUSED"] = "1"
            try:
                if mode == "fused_device":
                    dv, di = rd.CreateBuffer(plt, v.nbytes), rd.CreateBuffer(plt, t.nbytes)
                    rd.WriteBuffer(plt, dv, v.nbytes, v)
                    rd.WriteBuffer(plt, di, t.nbytes, t)
                    out = rd.C.c_void_p()
                    rd._check(rd.lib().rd_blas_build_device(plt.handle, dv.device_ptr(), nt * 3, di.device_ptr(), nt, rd.C.byref(out)))
                    blas = rd.BottomAccelStruct(plt, out)
                else:
                    blas = rd.BuildAccelStruct(plt, rd.Mesh(v, t))
            finally:
                os.environ.pop("RD_BUILD_UNFUSED", None)
            tlas = rd.BuildAccelStruct(plt, [rd.Instance(np.eye(4), 0, 0, blas)])
            hitA, hitB = rd.trace_host_rays(plt, tlas, rays)
            # node and leaf-record NUMBERS come from atomic counters (allocation order may differ between a 256- and a
            # 1024-thread block), so the comparison is on what does not depend on them: counts, levels, blob size, hits
            got[mode] = (blas.stats(), rd.TopAccelStructToBytes(plt, tlas).size, hitA.tobytes(), hitB.tobytes())
        assert got["fused"] == got["unfused"], "fused small build differs from the multi-launch one at nt=%d" % nt
        assert got["fused_device"] == got["unfused"], "fused small build from device arrays differs at nt=%d" % nt


def test_non_finite_vertices_do_not_hang_the_build(pkg, plt):
    """a NaN / Inf vertex made a cluster unmergeable and the one-block PLOC loop of small meshes never ended.
    Areas are now sanitised (a round always merges), and primitives with non-finite boxes are refused with RD_E_ARG."""
    rd, scenes = pkg.rd, pkg.scenes
    rng = np.random.default_rng(3)
    for nt in (40, 9000):   # one-block path (<= 8192) and multi-launch path
        v = rng.normal(size=(nt * 3, 3)).astype(np.float32)
        t = np.arange(nt * 3, dtype=np.uint32).reshape(nt, 3)
        bad = v.copy()
        bad[5] = np.inf
        # small meshes build deferred: the error surfaces at the first call that needs the result (stats / the instance build)
        with pytest.raises(rd.RadianceError, match="NaN / Inf"):
            rd.BuildAccelStruct(plt, rd.Mesh(bad, t)).stats()
        bad = v.copy()
        bad[3:6] = np.nan                      # a triangle that is NaN in every vertex
        with pytest.raises(rd.RadianceError, match="NaN / Inf"):
            blas = rd.BuildAccelStruct(plt, rd.Mesh(bad, t))
            rd.BuildAccelStruct(plt, [rd.Instance(np.eye(4), 0, 0, blas)])
        bad = v.copy()
        bad[7, 1] = np.nan                     # one NaN coordinate: the triangle can never be hit, the rest still is
        bad[30:33] = 1.0e30 * np.array([[1, 0, 0], [0, 1, 0], [0, 0, 1]], np.float32)   # finite, but its area overflows
        blas = rd.BuildAccelStruct(plt, rd.Mesh(bad, t))
        tlas = rd.BuildAccelStruct(plt, [rd.Instance(np.eye(4), 0, 0, blas)])
        c = v.reshape(nt, 3, 3).mean(axis=1)[20:30]
        rays = np.zeros((len(c), 8), np.float32)
        rays[:, 0:3] = c + np.array([0.01, 0.02, 5.0], np.float32)
        rays[:, 3], rays[:, 7] = 1e-3, 1e3
        rays[:, 4:7] = c - rays[:, 0:3]
        hitA, _ = rd.trace_host_rays(plt, tlas, rays)
        assert (hitA["t"] < 3e38).all()


def test_chunk_stealing_with_large_blocks(pkg, plt):
    """the per-thread work-item slot of stage / stealing kernels was sized for 256 threads; option block accepts
    up to 1024."""
    rd, scenes = pkg.rd, pkg.scenes
    b = offcentre_cornell(scenes)
    tlas, _ = scenes.build_scene(plt, b)
    w, h = 320, 240
    ds = scenes.DeviceScene(plt, b, tlas, w, h, scenes.rtprop(0, 1, 2, 0))
    plt.set_option("stealing", 1)
    try:
        pipe, _ = scenes.make_pipeline(plt, "path")
    finally:
        plt.set_option("stealing", 0)
    rd.BindPipeline(plt, pipe)
    rd.BindDescriptorSet(plt, ds.descriptors)
    rd.TraceRays(plt, 0, 0, 0, w, h)
    plain = ds.read_image()
    try:
        for block in (512, 1024):
            plt.set_option("block", block)
            rd.WriteBuffer(plt, ds.rdImage, plain.nbytes, np.zeros_like(plain))
            rd.TraceRaysSteal(plt, w, h, 40 + block, 256)
            plt.sync()
            assert np.array_equal(ds.read_image(), plain), block
    finally:
        plt.set_option("block", 128)


def test_async_entry_points_report_device_status(pkg, plt, monkeypatch):
    """stack overflow / nesting flags were read by the blocking rd_trace_rays only. A program compiled with a
    2-entry traversal stack overflows on the height field; the asynchronous launch must not come back clean."""
    rd, scenes = pkg.rd, pkg.scenes
    b = scenes.heightfield_scene(96, 160, 96)
    tlas, _ = scenes.build_scene(plt, b)
    ds = scenes.DeviceScene(plt, b, tlas, 160, 96, scenes.rtprop(0, 1, 1, 0))
    monkeypatch.setenv("RD_NVRTC_FLAGS", "-DRD_STACK_ENTRIES=2")
    pipe, _ = scenes.make_pipeline(plt, "primary")
    monkeypatch.delenv("RD_NVRTC_FLAGS")
    rd.BindPipeline(plt, pipe)
    rd.BindDescriptorSet(plt, ds.descriptors)
    assert plt.status() == (0, 0)
    rd.TraceRaysAsync(plt, 160, 96)
    with pytest.raises(rd.RadianceError, match="stack overflow"):
        plt.sync()
    assert plt.status() == (0, 0)            # reported once, then cleared
    rd.TraceRaysRange(plt, 160, 96, 0, 160 * 48)
    img = np.zeros((96, 160, 4), np.uint8)
    with pytest.raises(rd.RadianceError, match="stack overflow"):
        rd.ReadBuffer(plt, ds.rdImage, img.nbytes, img)
    # a clean program on the same platform is not blamed afterwards
    pipe, _ = scenes.make_pipeline(plt, "primary")
    rd.BindPipeline(plt, pipe)
    rd.BindDescriptorSet(plt, ds.descriptors)
    rd.TraceRays(plt, 0, 0, 0, 160, 96)
    plt.sync()


def test_frame_gather_pack_unpack_and_work_units(pkg, plt):
    """The C-ABI multi-GPU pieces that need no second GPU: (1) the pack / unpack kernels of rd_frame_gather move exactly the
    bands of each rank (all ranks together rebuild the frame, RGBA8 and float scratch); (2) rd_trace_rays_units with a
    private counter renders every unit once, inline and in wavefront mode, bit-identical to the plain frame; (3) a
    one-rank communicator gathers trivially; (4) the wavefront stage timeline covers the frame."""
    import ctypes as C
    rd, scenes = pkg.rd, pkg.scenes
    b = offcentre_cornell(scenes)
    tlas, _ = scenes.build_scene(plt, b)
    w, h = 320, 240
    ds = scenes.DeviceScene(plt, b, tlas, w, h, scenes.rtprop(0, 2, 3, 0))
    pipe, _ = scenes.make_pipeline(plt, "path")
    rd.BindPipeline(plt, pipe)
    rd.BindDescriptorSet(plt, ds.descriptors)
    rd.TraceRays(plt, 0, 0, 0, w, h)
    plain, plain_scr = ds.read_image(), ds.read_scratch()
    for item, src in ((4, ds.rdImage), (16, ds.rdImageScratch)):
        dst = rd.CreateBuffer(plt, w * h * item)
        for world, band in ((2, 1), (8, 1), (3, 2), (4, 7)):
            rd.WriteBuffer(plt, dst, w * h * item, np.zeros(w * h * item, np.uint8))
            for r in range(world):
                rd._check(rd.lib().rd_debug_band_roundtrip(plt.handle, src.handle, dst.handle, w, h, item, band, r, world))
                if r == 0 and world > 1:
                    part = np.zeros(w * h * item, np.uint8)
                    rd.ReadBuffer(plt, dst, part.nbytes, part)
                    rows = part.reshape(h, -1)
                    mine = np.zeros(h, bool)
                    for y0, ny in pkg.multigpu.interleaved_rows(0, world, h, band):
                        mine[y0:y0 + ny] = True
                    assert (rows[~mine] == 0).all()
            out = np.zeros(w * h * item, np.uint8)
            rd.ReadBuffer(plt, dst, out.nbytes, out)
            ref_bytes = (plain if item == 4 else plain_scr).reshape(-1).view(np.uint8)
            assert np.array_equal(out, ref_bytes), (item, world, band)
        dst.free()
    # dynamic work units (one rank takes them all)
    for wf in (0, 1):
        plt.set_option("wavefront", wf)
        try:
            for units, frame in ((1, 200 + 10 * wf), (7, 201 + 10 * wf), (60, 202 + 10 * wf), (64, 203 + 10 * wf)):   # counter slots no other test used
                rd.WriteBuffer(plt, ds.rdImage, plain.nbytes, np.zeros_like(plain))
                ds.set_rtprop(0, 2, 3)
                mask = rd.TraceRaysUnits(plt, w, h, units, frame)
                assert mask == (1 << units) - 1
                plt.sync()
                img = ds.read_image()
                if wf == 0:
                    assert np.array_equal(img, plain), units
                else:
                    d = img[..., :3].astype(np.float64) - plain[..., :3].astype(np.float64)
                    assert np.sqrt((d * d).mean()) <= 0.5, units
        finally:
            plt.set_option("wavefront", 0)
    # stage timeline
    plt.set_option("wavefront", 1)
    plt.set_option("wf_timeline", 1)
    try:
        ds.set_rtprop(0, 2, 3)
        rd.TraceRays(plt, 0, 0, 0, w, h)
        tl = rd.WavefrontTimeline(plt)
    finally:
        plt.set_option("wavefront", 0)
        plt.set_option("wf_timeline", 0)
    names = [n for n, _ in tl]
    assert names[0] == "generate" and names[-1] == "resolve" and names.count("shade") == 3 and names.count("trace") == 3
    assert all(ms >= 0.0 for _, ms in tl) and sum(ms for _, ms in tl) > 0.0
    # a communicator of one rank: gather is a no-op that succeeds (NCCL present on the GPU box)
    uid = rd.CommUniqueId()
    rd.CommInit(plt, 0, 1, uid)
    try:
        rd.FrameGather(plt, ds.rdImage, w, h, 4, 1)
        rd.FrameGatherUnits(plt, ds.rdImage, w, h, 4, 8, 0xff)
        with pytest.raises(rd.RadianceError):
            rd.CommInit(plt, 0, 1, uid)
    finally:
        rd.CommDestroy(plt)


def test_sort_key_beyond_255_rows_and_4_materials(pkg, plt, ref, tmp_path):
    """An sbt.json with 603 rows (300 hit groups) and 289 materials: the wavefront sort key folds the row modulo 256 and
    keeps two material bits -- it only groups paths, so the sorted, unsorted and inline frames stay identical, and the
    frame matches the reference runtime running the same generated program."""
    import os
    import subprocess
    import sys
    rd, scenes = pkg.rd, pkg.scenes
    root = os.path.join(os.path.dirname(__file__), "..")
    env = dict(os.environ, RD_GROUPS="300", RD_GROUPS_OUT=str(tmp_path))
    subprocess.run([sys.executable, os.path.join(root, "tools", "gen_groups.py")], check=True, env=env, stdout=subprocess.DEVNULL)
    src, sbt = open(tmp_path / "groups.cl").read(), open(tmp_path / "groups_sbt.json").read()
    assert sbt.count("material_") == 300
    b = scenes.groups_scene(68, 34, 300, 192, 144)   # 1156 instances, SBTOffset 0..288, customInstanceID = SBTOffset
    assert max(i[2] for i in b.instances) > 255 and len({i[3] for i in b.instances}) > 255
    tlas, _ = scenes.build_scene(plt, b)
    w, h = 192, 144
    ds = scenes.DeviceScene(plt, b, tlas, w, h, scenes.rtprop(0, 2, 3, 0))
    pipe, _ = scenes.make_pipeline(plt, source=src, sbt=sbt)
    rd.BindPipeline(plt, pipe)
    rd.BindDescriptorSet(plt, ds.descriptors)
    out = {}
    try:
        for mode, wf, srt in (("inline", 0, 1), ("sorted", 1, 1), ("unsorted", 1, 0)):
            plt.set_option("wavefront", wf)
            plt.set_option("wf_sort", srt)
            ds.set_rtprop(0, 2, 3)
            rd.TraceRays(plt, 0, 0, 0, w, h)
            out[mode] = ds.read_image()
    finally:
        plt.set_option("wavefront", 0)
        plt.set_option("wf_sort", 1)
    assert np.array_equal(out["sorted"], out["unsorted"])
    d = out["sorted"][..., :3].astype(np.float64) - out["inline"][..., :3].astype(np.float64)
    assert np.sqrt((d * d).mean()) <= 0.5
    assert len(np.unique(out["inline"][..., :3].reshape(-1, 3), axis=0)) > 500
