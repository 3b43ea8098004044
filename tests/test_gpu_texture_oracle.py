"""read_imageui (shader/compat/cl_compat.h) against an INDEPENDENT numpy statement of the OpenCL 1.2 image addressing and
filtering equations (specification section 8.2, normalised coordinates), texel by texel -- the image-parity test of
tests/test_gpu_parity.py compiles the reference program against the host flavour of the same prelude, so it cannot pin the
sampler conventions. Reference use: radiance/src/radiance.cpp:96-131 (CL_ADDRESS_REPEAT + CL_FILTER_LINEAR,
normalised, layered RGBA8), samples/shader2.cl:260-264.

What the specification fixes and this test checks: the wrap of the texel indices per addressing mode, the half-texel offset
and the weights of the linear filter, the layer selection clamp(rint(w)), the border colour of CL_ADDRESS_CLAMP.
What it leaves open: read_imageui with CL_FILTER_LINEAR is undefined in OpenCL; the convention here (and in the reference
images, which come from PoCL's float path) is the bilinear sum of the four 0..255 values rounded to nearest -- the bar is
|device - exact bilinear sum| <= 0.5 + 2e-3 (float32 weights)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

ADDRESS = {"CLAMP_TO_EDGE": 0x1131, "CLAMP": 0x1132, "REPEAT": 0x1133, "MIRRORED_REPEAT": 0x1134}
NEAREST, LINEAR = 0x1140, 0x1141

PROGRAM = r'''
#include "radiance.cl"
struct Payload { int unused; };
struct SceneData { int unused; };
void callHit(int sbtRecordOffset, struct Payload* payload, struct HitData* hitData,
    struct SceneData* sceneData, image2d_array_t imageArray, sampler_t sampler) {}
void callMiss(int missIndex, struct Payload* payload,
    struct SceneData* sceneData, image2d_array_t imageArray, sampler_t sampler) {}
void callAnyHit(bool* cont, int sbtRecordOffset, struct Payload* payload, struct HitData* hitData,
    struct SceneData* sceneData, image2d_array_t imageArray, sampler_t sampler) {}
__kernel void raygen(__global uint* header, __global float* coords, __global uint* out, image2d_array_t img, sampler_t smp)
{
    const uint i = get_global_id(0);
    if (i >= header[0]) return;
    float4 c = {coords[4 * i], coords[4 * i + 1], coords[4 * i + 2], 0.0f};
    uint4 t = read_imageui(img, smp, c);
    out[4 * i] = t.x; out[4 * i + 1] = t.y; out[4 * i + 2] = t.z; out[4 * i + 3] = t.w;
}
'''


def spec_indices(s, w, mode, linear):
    """OpenCL 1.2 section 8.2 for one normalised coordinate: (i0, i1, a, border0, border1); nearest uses i0 only."""
    s = s.astype(np.float64)
    if mode == "REPEAT":
        u = (s - np.floor(s)) * w
        if not linear:
            i = np.floor(u).astype(np.int64)
            i = np.where(i > w - 1, i - w, i)
            return i, i, np.zeros_like(u), np.zeros(len(u), bool), np.zeros(len(u), bool), u
        i0 = np.floor(u - 0.5).astype(np.int64)
        i1 = i0 + 1
        i0 = np.where(i0 < 0, w + i0, i0)
        i1 = np.where(i1 > w - 1, i1 - w, i1)
        return i0, i1, (u - 0.5) - np.floor(u - 0.5), np.zeros(len(u), bool), np.zeros(len(u), bool), u
    if mode == "MIRRORED_REPEAT":
        sp = np.abs(s - 2.0 * np.rint(0.5 * s))
        u = sp * w
        if not linear:
            i = np.minimum(np.floor(u).astype(np.int64), w - 1)
            return i, i, np.zeros_like(u), np.zeros(len(u), bool), np.zeros(len(u), bool), u
        i0 = np.floor(u - 0.5).astype(np.int64)
        i1 = i0 + 1
        return np.maximum(i0, 0), np.minimum(i1, w - 1), (u - 0.5) - np.floor(u - 0.5), np.zeros(len(u), bool), np.zeros(len(u), bool), u
    u = s * w
    if not linear:
        i0 = np.floor(u).astype(np.int64)
        i1 = i0
        a = np.zeros_like(u)
    else:
        i0 = np.floor(u - 0.5).astype(np.int64)
        i1 = i0 + 1
        a = (u - 0.5) - np.floor(u - 0.5)
    if mode == "CLAMP_TO_EDGE":
        return np.clip(i0, 0, w - 1), np.clip(i1, 0, w - 1), a, np.zeros(len(u), bool), np.zeros(len(u), bool), u
    b0, b1 = (i0 < 0) | (i0 > w - 1), (i1 < 0) | (i1 > w - 1)   # CL_ADDRESS_CLAMP: outside -> border colour (0,0,0,0) for RGBA
    return np.clip(i0, 0, w - 1), np.clip(i1, 0, w - 1), a, b0, b1, u


def spec_read(tex, coords, mode, linear):
    layers, h, w, _ = tex.shape
    layer = np.clip(np.rint(coords[:, 2].astype(np.float64)), 0, layers - 1).astype(np.int64)   # rint: half to even
    i0, i1, a, bi0, bi1, u = spec_indices(coords[:, 0], w, mode, linear)
    j0, j1, b, bj0, bj1, v = spec_indices(coords[:, 1], h, mode, linear)

    def fetch(i, j, bi, bj):
        t = tex[layer, j, i].astype(np.float64)
        t[bi | bj] = 0.0
        return t
    if not linear:
        return fetch(i0, j0, bi0, bj0), u, v
    a, b = a[:, None], b[:, None]
    return ((1 - a) * (1 - b) * fetch(i0, j0, bi0, bj0) + a * (1 - b) * fetch(i1, j0, bi1, bj0) +
            (1 - a) * b * fetch(i0, j1, bi0, bj1) + a * b * fetch(i1, j1, bi1, bj1)), u, v


def test_read_imageui_matches_the_opencl_addressing_and_filter_equations(pkg, plt, parity_log):
    rd, scenes = pkg.rd, pkg.scenes
    rng = np.random.default_rng(7)
    W, H, LAYERS = 13, 6, 3   # not powers of two, not square: wrap arithmetic has nowhere to hide
    tex = rng.integers(0, 256, (LAYERS, H, W, 4), dtype=np.uint8)
    img = rd.CreateImageArray(plt, W, H, LAYERS)
    for l in range(LAYERS):
        rd.WriteImage(plt, img, W, H, l, tex[l])
    n_rand = 20000
    coords = np.zeros((n_rand, 4), np.float32)
    coords[:, 0:2] = rng.uniform(-2.5, 3.5, (n_rand, 2))
    coords[:, 2] = rng.uniform(-1.0, LAYERS + 0.5, n_rand)
    # texel centres and texel edges of every texel, wrap seams included, and the half-way layers (rint ties go to even)
    cx, cy = np.meshgrid((np.arange(-W, 2 * W) + 0.5) / W, (np.arange(-H, 2 * H) + 0.5) / H)
    centres = np.zeros((cx.size, 4), np.float32)
    centres[:, 0], centres[:, 1], centres[:, 2] = cx.ravel(), cy.ravel(), 1.0
    ties = np.zeros((4, 4), np.float32)
    ties[:, 0:2] = 0.3
    ties[:, 2] = [0.5, 1.5, 2.5, -0.5]
    coords = np.vstack([coords, centres, ties])
    n = len(coords)
    pipe, _ = scenes.make_pipeline(plt, source=PROGRAM, sbt="[]")
    h_hdr, h_in, h_out = rd.CreateBuffer(plt, 16), rd.CreateBuffer(plt, coords.nbytes), rd.CreateBuffer(plt, n * 16)
    rd.WriteBuffer(plt, h_hdr, 16, np.array([n, 0, 0, 0], np.uint32))
    rd.WriteBuffer(plt, h_in, coords.nbytes, coords)
    worst = 0.0
    for mode, code in ADDRESS.items():
        for linear in (False, True):
            smp = rd.CreateSampler(plt, code, LINEAR if linear else NEAREST)
            rd.BindPipeline(plt, pipe)
            rd.BindDescriptorSet(plt, rd.CreateDescriptorSet([h_hdr, h_in, h_out, img, smp]))
            rd.TraceRays(plt, 0, 0, 0, n, 1)
            got = np.zeros((n, 4), np.uint32)
            rd.ReadBuffer(plt, h_out, got.nbytes, got)
            exp, u, v = spec_read(tex, coords, mode, linear)
            if linear:
                err = np.abs(got.astype(np.float64) - exp).max(axis=1)
                worst = max(worst, float(err.max()))
                assert err.max() <= 0.5 + 2e-3, "%s LINEAR: %g off the bilinear sum at %s" % (mode, err.max(), coords[err.argmax()])
            else:
                # a coordinate within float32 rounding of a texel boundary may land on either side: not a convention question
                fu, fv = u - np.floor(u), v - np.floor(v)
                safe = (np.minimum(fu, 1 - fu) > 1e-4) & (np.minimum(fv, 1 - fv) > 1e-4)
                assert safe.sum() > 0.99 * n
                bad = (got[safe] != exp[safe].astype(np.uint32)).any(axis=1)
                assert not bad.any(), "%s NEAREST: %d texels differ, first at %s" % (mode, bad.sum(), coords[safe][bad][0])
            smp.free()
    parity_log.record("texture_read_imageui_vs_opencl_spec", samples=n, modes=len(ADDRESS) * 2, max_abs_err_linear=worst, bar=0.5 + 2e-3)
    for hnd in (h_hdr, h_in, h_out, img):
        hnd.free()
