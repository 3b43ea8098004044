"""Generates tests/golden/*.npz by running the REFERENCE's own code (oracle/_ref: radiance/src/bvh.cpp
unchanged + radiance/shader/*.cl compiled as C++). The reference ships no golden vectors or tests
(SURVEY.md section 4), so these are the pinned known answers. Run in the build container:
    python tests/golden/make_golden.py
Inputs are seeded / closed-form; the script is committed with its outputs."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import __graft_entry__ as entry  # noqa: E402

ref = entry.load_oracle()
scenes = entry.load_package().scenes
out = os.path.dirname(os.path.abspath(__file__))
rng = np.random.default_rng(20261104)

# 1. random_pcg3d (radiance/shader/math.cl:10-23): integer hash, float stage
v = rng.integers(0, 2**32, size=(256, 3), dtype=np.uint64).astype(np.uint32)
v[:4] = [[0, 0, 0], [1, 2, 3], [0xffffffff, 0xffffffff, 0xffffffff], [7, 0, 479999]]
np.savez(os.path.join(out, "pcg3d.npz"), input=v, output=ref.pcg3d(v))

# 2. InverseMat4x4 (math.cl:56-183) on affine instance-like matrices
mats, invs = [], []
for k in range(32):
    m = np.eye(4, dtype=np.float32)
    a = rng.normal(size=(3, 3)).astype(np.float32)
    m[:3, :3] = a + 2.0 * np.eye(3, dtype=np.float32)
    m[:3, 3] = rng.normal(size=3).astype(np.float32) * 3
    ok, inv = ref.inverse4x4(m)
    assert ok
    mats.append(m), invs.append(inv)
np.savez(os.path.join(out, "inverse4x4.npz"), input=np.array(mats), output=np.array(invs))

# 3. intersectTriangle (radiance.cl:211-251): random rays against random triangles + edge cases
n = 2048
o = rng.normal(size=(n, 3)).astype(np.float32) * 2
tri = rng.normal(size=(n, 9)).astype(np.float32)
tgt = (tri[:, 0:3] + tri[:, 3:6] + tri[:, 6:9]) / 3 + rng.normal(size=(n, 3)).astype(np.float32) * 0.4
d = (tgt - o).astype(np.float32)
# exact edge / vertex / parallel / behind cases
o[0], d[0], tri[0] = [0.25, 0.25, 1], [0, 0, -1], [0, 0, 0, 1, 0, 0, 0, 1, 0]
o[1], d[1], tri[1] = [0.5, 0.0, 1], [0, 0, -1], [0, 0, 0, 1, 0, 0, 0, 1, 0]       # on an edge (inclusive)
o[2], d[2], tri[2] = [0.0, 0.0, 1], [0, 0, -1], [0, 0, 0, 1, 0, 0, 0, 1, 0]       # on a vertex
o[3], d[3], tri[3] = [0.25, 0.25, 1], [1, 0, 0], [0, 0, 0, 1, 0, 0, 0, 1, 0]      # parallel: det == 0
o[4], d[4], tri[4] = [0.25, 0.25, 1], [0, 0, 1], [0, 0, 0, 1, 0, 0, 0, 1, 0]      # behind: t < 0
o[5], d[5], tri[5] = [0.25, 0.25, -1], [0, 0, 1], [0, 0, 0, 1, 0, 0, 0, 1, 0]     # back face: two-sided
hit = np.zeros(n, np.uint8)
res = np.zeros((n, 7), np.float32)
for i in range(n):
    ok, r = ref.intersect_triangle(o[i], d[i], tri[i])
    hit[i], res[i] = ok, r
np.savez(os.path.join(out, "triangle.npz"), origin=o, direction=d, triangle=tri, hit=hit, result=res)

# 4. traceRay hit records (radiance.cl:254-275) on the two small parity scenes
def record(name, bundle, rays):
    tl = ref.build_scene(bundle)
    closest = tl.trace(rays)
    first = tl.trace(rays, terminate_on_first_hit=True)
    np.savez_compressed(os.path.join(out, name), rays=rays, hit=closest["hit"].astype(np.uint8), t=closest["t"],
                        prim=closest["prim"], inst=closest["inst"], bary=closest["bary"], point=closest["point"],
                        occluded=first["hit"].astype(np.uint8))

b = scenes.cornell_like()
b.camera["x"], b.camera["y"] = 0.013, 1.037
record("cornell_hits.npz", b, scenes.pinhole_rays(b.camera, stride=8, jitter=(0.37, 0.61)))
b = scenes.heightfield_scene(64, 160, 90)
record("heightfield64_hits.npz", b, scenes.pinhole_rays(b.camera, jitter=(0.29, 0.71)))

# 5. raygen image of programs/primary.cl through the reference runtime (Cornell, 96x72, 2 spp)
b = scenes.cornell_like()
b.camera["x"], b.camera["y"] = 0.013, 1.037
w, h = 96, 72
cam = np.array(b.camera).reshape(()).copy()
cam["widthPixel"], cam["heightPixel"] = w, h
scratch, image = np.zeros(w * h * 4, np.float32), np.zeros(w * h * 4, np.uint8)
tl = ref.build_scene(b)
ref.raygen("primary", [np.array(scenes.rtprop(0, 2, 1, 0)).reshape(1), scratch, image, cam.reshape(1),
                       np.array(b.scene).reshape(1), b.meshInfoData, b.vertexData, b.indexData, b.uvData, b.normalData,
                       b.materials, None, 0, tl.blob], w * h)
np.savez_compressed(os.path.join(out, "primary_cornell_96x72.npz"), image=image.reshape(h, w, 4), scratch=scratch.reshape(h, w, 4))

# 6. generateRay (samples/shader.cl:111-173): pinhole and thin lens, a tilted camera, work-items all over the frame
cams = [scenes.camera(800, 600, (0.013, 1.037, 3.4), (0.0, 0.0, 0.0), 0.030, 0.036, 2.0, 0.0),
        scenes.camera(800, 600, (0.0, 0.6, 0.9), (-0.6, 0.3, 0.1), 0.030, 0.036, 3.1, 1.4),
        scenes.camera(1920, 1080, (1.0, 2.0, -3.0), (0.2, -2.4, 0.05), 0.050, 0.036, 5.0, 2.8)]
gr_cam, gr_in, gr_o, gr_d = [], [], [], []
for ci, c in enumerate(cams):
    wpx, hpx = int(c["widthPixel"]), int(c["heightPixel"])
    for k in range(256):
        index = int(rng.integers(0, wpx * hpx))
        frame, total = int(rng.integers(0, 64)), int(rng.integers(0, 64))
        o3, d3 = ref.generate_ray(np.array(c).reshape(1), np.array([frame, total, index], np.uint32), index)
        gr_cam.append(ci), gr_in.append((frame, total, index)), gr_o.append(o3), gr_d.append(d3)
np.savez(os.path.join(out, "generate_ray.npz"), cameras=np.array(cams), camera_index=np.array(gr_cam, np.uint32),
         input=np.array(gr_in, np.uint32), origin=np.array(gr_o), direction=np.array(gr_d))

# 7. pbr.cl / math.cl library functions as samples/shader.cl calls them (oracle/ref_device.cpp ref_library_kat)
def unit(n):
    v = rng.normal(size=(n, 3))
    return (v / np.linalg.norm(v, axis=1, keepdims=True)).astype(np.float32)

def same_side(nrm, n):
    v = unit(n)
    flip = (v * nrm).sum(axis=1) < 0.05
    v[flip] = (v[flip] + 2.0 * nrm[flip] * np.abs((v[flip] * nrm[flip]).sum(axis=1, keepdims=True)) + 0.3 * nrm[flip]).astype(np.float32)
    return (v / np.linalg.norm(v, axis=1, keepdims=True)).astype(np.float32)

nk = 512
N = unit(nk); V = same_side(N, nk); L = same_side(N, nk)
alb = rng.uniform(0.05, 0.95, size=(nk, 3)).astype(np.float32)
met = rng.choice([0.0, 0.25, 0.9, 1.0], size=(nk, 1)).astype(np.float32)
rough = rng.uniform(0.02, 1.0, size=(nk, 1)).astype(np.float32)
trans = rng.choice([0.0, 0.5, 0.9], size=(nk, 1)).astype(np.float32)
ior = rng.uniform(1.05, 1.9, size=(nk, 1)).astype(np.float32)
rnd = rng.uniform(0.0, 1.0, size=(nk, 3)).astype(np.float32)
eta = rng.choice([1.0 / 1.45, 1.45, 1.0 / 1.1, 1.9], size=(nk, 1)).astype(np.float32)
kat_inputs = {
    "BRDF": np.hstack([L, V, N, met, rough, alb]),
    "refract": np.hstack([V, N, eta]),
    "G_pbrt": np.hstack([V, L, N, rough]),
    "sampleMicrofacetBRDF_transm": np.hstack([V, N, alb, met, rough, trans, ior, rnd]),
    "microfacetBRDF": np.hstack([L, V, N, alb, met, rough, trans, ior]),
    "GetNormalSpace": np.vstack([N, [[1, 0, 0], [-1, 0, 0], [0, 1, 0], [0, 0, 1]]]).astype(np.float32),
    "sampleMicrofacetBRDF": np.hstack([V, N, alb, met, rough, rnd]),
}
kat = {}
for name, a in kat_inputs.items():
    kat[name + "_in"] = np.ascontiguousarray(a, np.float32)
    kat[name + "_out"] = ref.library_kat(name, a)
np.savez_compressed(os.path.join(out, "library_kat.npz"), **kat)
print("golden vectors written to", out)
