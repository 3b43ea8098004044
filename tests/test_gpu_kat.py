"""Device-side known-answer tests: the functions user programs really execute on sm_100a (shader/math.cl, shader/pbr.cl,
csrc/rd_traverse.h's triangle test, the library's primary-ray generator) on the committed golden vectors that the
reference's own functions produced (tests/golden/make_golden.py through oracle/_ref).

Bars:
  * random_pcg3d (radiance/shader/math.cl:10-23)           bit-exact (integer hash + one IEEE division)
  * intersectTriangle (radiance/shader/radiance.cl:211-251) hit flag, t and all three barycentrics bit-exact
  * InverseMat4x4 (math.cl:56-183)                          <= 1e-5 of the largest element (different but equivalent
                                                            cofactor grouping, FMA contraction on the device)
  * generateRay (samples/shader.cl:111-173)                 origin / direction within 2e-6 absolute (libm vs libdevice
                                                            sin / cos), pinhole and thin lens
  * pbr.cl / GetNormalSpace                                 within 2e-4 (absolute + relative) on >= 99.5 % of the rows; the
                                                            remaining rows may sit on a branch (acos / sqrt arguments at
                                                            the clamp, hemisphere tests) that flips with one ulp
The achieved maxima are written to gpurun_out/parity_r2.json."""
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

HERE = os.path.dirname(os.path.abspath(__file__))
GOLD = os.path.join(HERE, "golden")


@pytest.fixture(scope="module")
def kat(pkg, plt):
    rd, scenes = pkg.rd, pkg.scenes
    pipe, _ = scenes.make_pipeline(plt, source=open(os.path.join(HERE, "kat", "kat.cl")).read(), sbt="[]")

    def run(fn, inputs, out_stride):
        a = np.ascontiguousarray(inputs)
        n, in_stride = a.shape[0], a.shape[1]
        hdr = np.array([fn, n, in_stride, out_stride], np.uint32)
        h_hdr, h_in, h_out = rd.CreateBuffer(plt, 16), rd.CreateBuffer(plt, a.nbytes), rd.CreateBuffer(plt, n * out_stride * 4)
        rd.WriteBuffer(plt, h_hdr, 16, hdr)
        rd.WriteBuffer(plt, h_in, a.nbytes, a)
        rd.BindPipeline(plt, pipe)
        rd.BindDescriptorSet(plt, rd.CreateDescriptorSet([h_hdr, h_in, h_out]))
        rd.TraceRays(plt, 0, 0, 0, n, 1)
        out = np.zeros((n, out_stride), np.float32)
        rd.ReadBuffer(plt, h_out, out.nbytes, out)
        for h in (h_hdr, h_in, h_out):
            h.free()
        return out
    return run


def test_pcg3d_bit_exact_on_device(kat, parity_log):
    g = np.load(os.path.join(GOLD, "pcg3d.npz"))
    out = kat(0, g["input"].astype(np.uint32), 3)
    assert np.array_equal(out.view(np.uint32), g["output"].view(np.uint32))
    parity_log.record("KAT_random_pcg3d", vectors=len(out), bit_exact=True)


def test_triangle_test_bit_exact_on_device(kat, parity_log):
    g = np.load(os.path.join(GOLD, "triangle.npz"))
    a = np.hstack([g["origin"], g["direction"], g["triangle"]]).astype(np.float32)
    out = kat(2, a, 5)
    hit = g["hit"].astype(bool)
    assert np.array_equal(out[:, 0] == 1.0, hit)
    assert hit.sum() > 300 and (~hit).sum() > 300
    exp = g["result"][hit]   # {t, point.xyz, bary.xyz}
    got = out[hit]
    assert np.array_equal(got[:, 1].view(np.uint32), exp[:, 0].view(np.uint32)), "t differs"
    assert np.array_equal(got[:, 2:5].view(np.uint32), exp[:, 4:7].view(np.uint32)), "barycentrics differ"
    parity_log.record("KAT_intersectTriangle", vectors=len(out), hits=int(hit.sum()), bit_exact=True)


def test_inverse4x4_on_device(kat, parity_log):
    g = np.load(os.path.join(GOLD, "inverse4x4.npz"))
    out = kat(1, g["input"].reshape(-1, 16), 17)
    assert (out[:, 16] == 1.0).all()
    exp = g["output"].reshape(-1, 16)
    err = np.abs(out[:, :16] - exp).max(axis=1) / np.abs(exp).max(axis=1)
    parity_log.record("KAT_InverseMat4x4", vectors=len(out), max_rel_err=err.max(), bar=1e-5)
    assert err.max() <= 1e-5
    # singular input: reports failure
    sing = np.zeros((1, 16), np.float32)
    sing[0, [0, 5]] = 1.0
    assert kat(1, sing, 17)[0, 16] == 0.0


LIB = {"BRDF": 10, "refract": 11, "G_pbrt": 12, "sampleMicrofacetBRDF_transm": 13, "microfacetBRDF": 14,
       "GetNormalSpace": 15, "sampleMicrofacetBRDF": 16}


@pytest.mark.parametrize("name", sorted(LIB))
def test_library_function_on_device(kat, parity_log, name):
    g = np.load(os.path.join(GOLD, "library_kat.npz"))
    a, exp = g[name + "_in"], g[name + "_out"]
    out = kat(LIB[name], a, exp.shape[1])
    fin = np.isfinite(exp).all(axis=1)
    assert np.array_equal(np.isfinite(out).all(axis=1), fin)
    err = np.abs(out[fin] - exp[fin]) / (1.0 + np.abs(exp[fin]))
    row = err.max(axis=1)
    ok = row <= 2e-4
    parity_log.record("KAT_" + name, vectors=len(out), rows_within_bar=ok.mean(), bar=2e-4, median_err=np.median(row),
                      max_err_of_rows_within=row[ok].max() if ok.any() else 0.0)
    assert ok.mean() >= 0.995, "%s: %d of %d rows off (max %g)" % (name, (~ok).sum(), len(ok), row.max())
    assert np.abs(exp[fin]).max() > 0.05   # the vectors are not degenerate


def test_generate_ray_on_device(pkg, plt, parity_log):
    """rd_generate_primary_rays (the library's restatement of generateRay, also what programs/*.cl inline) against
    the reference's generateRay, pinhole and thin lens, tilted cameras."""
    rd = pkg.rd
    g = np.load(os.path.join(GOLD, "generate_ray.npz"))
    worst_o = worst_d = 0.0
    plt.set_option("tile", 0)   # slot == work-item index
    try:
        for ci, cam in enumerate(g["cameras"]):
            w, h = int(cam["widthPixel"]), int(cam["heightPixel"])
            sel = np.nonzero(g["camera_index"] == ci)[0]
            hc = rd.CreateBuffer(plt, 48)
            rd.WriteBuffer(plt, hc, 48, np.array(cam).reshape(1))
            ra, rb = rd.CreateBuffer(plt, w * h * 16), rd.CreateBuffer(plt, w * h * 16)
            groups = {}
            for k in sel:
                frame, total, index = (int(x) for x in g["input"][k])
                groups.setdefault((frame, total), []).append((k, index))
            A, B = np.zeros((w * h, 4), np.float32), np.zeros((w * h, 4), np.float32)
            for (frame, total), items in groups.items():
                rd.GeneratePrimaryRays(plt, hc, frame, total, w, h, ra, rb)
                plt.sync()
                rd.ReadBuffer(plt, ra, A.nbytes, A)
                rd.ReadBuffer(plt, rb, B.nbytes, B)
                for k, index in items:
                    worst_o = max(worst_o, float(np.abs(A[index, :3] - g["origin"][k]).max()))
                    worst_d = max(worst_d, float(np.abs(B[index, :3] - g["direction"][k]).max()))
            for hdl in (hc, ra, rb):
                hdl.free()
    finally:
        plt.set_option("tile", 1)
    parity_log.record("KAT_generateRay", vectors=len(g["input"]), max_abs_origin_err=worst_o, max_abs_direction_err=worst_d, bar=2e-6)
    assert worst_o <= 2e-6 and worst_d <= 2e-6
