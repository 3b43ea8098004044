"""Procedural scenes and descriptor bundles for tests and bench.py (synthetic data only).

Shapes follow SURVEY.md section 8(d): C1 Cornell-like box (~2k triangles, one instance per mesh),
C2 height-field n=708 (1 002 528 triangles), C3/C4 n=2237 (10 M), C5 = 5 instances of the 10 M mesh.
All byte layouts are the reference's host structs (radiance/src/core.h:103-158) and the
9-descriptor scene bundle of tools/sceneBuilder.h:8-17 / tools/sceneBuilder.cpp:69-101.
"""
import numpy as np

MATERIAL_DTYPE = np.dtype([("albedo", "<f4", (4,)), ("metallic", "<f4"), ("roughness", "<f4"),
                           ("transmission", "<f4"), ("ior", "<f4"), ("albedoTex", "<i4"),
                           ("metallicTex", "<i4"), ("roughnessTex", "<i4"), ("normalTex", "<i4")])
MESHINFO_DTYPE = np.dtype([("vertexOffset", "<i4"), ("indexOffset", "<i4"), ("uvOffset", "<i4"),
                           ("normalOffset", "<i4"), ("materialIndex", "<i4"), ("pad", "<i4", (3,))])
CAMERA_DTYPE = np.dtype([("widthPixel", "<f4"), ("heightPixel", "<f4"), ("focalLength", "<f4"),
                         ("sensorWidth", "<f4"), ("focalDistance", "<f4"), ("fStop", "<f4"),
                         ("x", "<f4"), ("y", "<f4"), ("z", "<f4"),
                         ("wx", "<f4"), ("wy", "<f4"), ("wz", "<f4")])
RTPROP_DTYPE = np.dtype([("totalSamples", "<u4"), ("batchSize", "<u4"), ("depth", "<u4"), ("debug", "<u4")])
SCENEPROP_DTYPE = np.dtype([("lightCount", "<u4", (4,)), ("lights", [("direction", "<f4", (4,)), ("color", "<f4", (4,))], (5,))])
assert MATERIAL_DTYPE.itemsize == 48 and MESHINFO_DTYPE.itemsize == 32
assert CAMERA_DTYPE.itemsize == 48 and SCENEPROP_DTYPE.itemsize == 176


def _hash01(ix, iz, seed):
    """Integer hash -> [0,1) float32, vectorised (jitter for the height-field)."""
    h = (ix.astype(np.uint64) * np.uint64(73856093)) ^ (iz.astype(np.uint64) * np.uint64(19349663)) ^ np.uint64(seed * 83492791 + 12345)
    h = (h ^ (h >> np.uint64(13))) * np.uint64(0x5bd1e995) & np.uint64(0xffffffff)
    h = h ^ (h >> np.uint64(15))
    return (h & np.uint64(0xffffff)).astype(np.float32) / np.float32(16777216.0)


def heightfield(n, seed=1, amplitude=0.05, jitter=0.25):
    """n x n quads on [-0.5,0.5]^2, y = A sin(13x) cos(11z) + hash jitter; 2 n^2 triangles."""
    g = np.arange(n + 1, dtype=np.int64)
    ix, iz = np.meshgrid(g, g, indexing="xy")
    x = (ix / n - 0.5).astype(np.float32)
    z = (iz / n - 0.5).astype(np.float32)
    y = (amplitude * np.sin(13.0 * x) * np.cos(11.0 * z)).astype(np.float32)
    y = y + (_hash01(ix, iz, seed) - np.float32(0.5)) * np.float32(jitter / n)
    verts = np.stack([x, y, z], axis=-1).reshape(-1, 3).astype(np.float32)
    q = (iz[:-1, :-1] * (n + 1) + ix[:-1, :-1]).reshape(-1).astype(np.uint32)
    a, b, c, d = q, q + 1, q + np.uint32(n + 1), q + np.uint32(n + 2)
    tris = np.empty((2 * n * n, 3), np.uint32)
    tris[0::2] = np.stack([a, c, b], axis=-1)
    tris[1::2] = np.stack([b, c, d], axis=-1)
    return verts, tris


def vertex_normals(verts, tris):
    v = verts.astype(np.float64)
    fn = np.cross(v[tris[:, 1]] - v[tris[:, 0]], v[tris[:, 2]] - v[tris[:, 0]])
    nrm = np.zeros_like(v)
    for k in range(3):
        for c in range(3):   # bincount scatter-add: seconds at 10 M triangles where np.add.at takes minutes
            nrm[:, c] += np.bincount(tris[:, k], weights=fn[:, c], minlength=v.shape[0])
    ln = np.linalg.norm(nrm, axis=1, keepdims=True)
    ln[ln == 0] = 1.0
    return (nrm / ln).astype(np.float32)


def icosphere(subdiv=3, radius=1.0, bumps=0.0, seed=0):
    t = (1.0 + 5.0 ** 0.5) / 2.0
    v = [(-1, t, 0), (1, t, 0), (-1, -t, 0), (1, -t, 0), (0, -1, t), (0, 1, t), (0, -1, -t), (0, 1, -t),
         (t, 0, -1), (t, 0, 1), (-t, 0, -1), (-t, 0, 1)]
    f = [(0, 11, 5), (0, 5, 1), (0, 1, 7), (0, 7, 10), (0, 10, 11), (1, 5, 9), (5, 11, 4), (11, 10, 2), (10, 7, 6),
         (7, 1, 8), (3, 9, 4), (3, 4, 2), (3, 2, 6), (3, 6, 8), (3, 8, 9), (4, 9, 5), (2, 4, 11), (6, 2, 10),
         (8, 6, 7), (9, 8, 1)]
    v = [np.array(p, np.float64) / np.linalg.norm(p) for p in v]
    for _ in range(subdiv):
        cache, nf = {}, []

        def mid(i, j):
            key = (min(i, j), max(i, j))
            if key not in cache:
                m = v[i] + v[j]
                v.append(m / np.linalg.norm(m))
                cache[key] = len(v) - 1
            return cache[key]
        for (a, b, c) in f:
            ab, bc, ca = mid(a, b), mid(b, c), mid(c, a)
            nf += [(a, ab, ca), (b, bc, ab), (c, ca, bc), (ab, bc, ca)]
        f = nf
    V = np.array(v, np.float64)
    if bumps:
        V = V * (1.0 + bumps * np.sin(7.0 * V[:, :1] + seed) * np.cos(5.0 * V[:, 1:2]) * np.sin(6.0 * V[:, 2:3]))
    return (V * radius).astype(np.float32), np.array(f, np.uint32)


def box(lo, hi):
    lo, hi = np.asarray(lo, np.float32), np.asarray(hi, np.float32)
    c = np.array([[lo[0], lo[1], lo[2]], [hi[0], lo[1], lo[2]], [hi[0], hi[1], lo[2]], [lo[0], hi[1], lo[2]],
                  [lo[0], lo[1], hi[2]], [hi[0], lo[1], hi[2]], [hi[0], hi[1], hi[2]], [lo[0], hi[1], hi[2]]], np.float32)
    f = np.array([[0, 2, 1], [0, 3, 2], [4, 5, 6], [4, 6, 7], [0, 1, 5], [0, 5, 4], [2, 3, 7], [2, 7, 6],
                  [1, 2, 6], [1, 6, 5], [0, 4, 7], [0, 7, 3]], np.uint32)
    return c, f


def quad(p0, p1, p2, p3):
    return np.array([p0, p1, p2, p3], np.float32), np.array([[0, 1, 2], [0, 2, 3]], np.uint32)


def translate(x, y, z):
    m = np.eye(4, dtype=np.float32)
    m[:3, 3] = (x, y, z)
    return m


def rotate_y(a):
    m = np.eye(4, dtype=np.float32)
    c, s = np.float32(np.cos(a)), np.float32(np.sin(a))
    m[0, 0], m[0, 2], m[2, 0], m[2, 2] = c, s, -s, c
    return m


def scale(s):
    m = np.eye(4, dtype=np.float32)
    m[0, 0] = m[1, 1] = m[2, 2] = s
    return m


def material(albedo=(0.8, 0.8, 0.8), metallic=0.0, roughness=0.5, transmission=0.0, ior=1.5):
    m = np.zeros((), MATERIAL_DTYPE)
    m["albedo"] = (*albedo, 1.0)
    m["metallic"], m["roughness"], m["transmission"], m["ior"] = metallic, roughness, transmission, ior
    m["albedoTex"] = m["metallicTex"] = m["roughnessTex"] = m["normalTex"] = -1
    return m


def camera(width, height, pos, rot=(0.0, 0.0, 0.0), focal=0.030, sensor=0.036, focal_distance=2.0, fstop=0.0):
    c = np.zeros((), CAMERA_DTYPE)
    c["widthPixel"], c["heightPixel"] = width, height
    c["focalLength"], c["sensorWidth"], c["focalDistance"], c["fStop"] = focal, sensor, focal_distance, fstop
    c["x"], c["y"], c["z"] = pos
    c["wx"], c["wy"], c["wz"] = rot
    return c


def rtprop(total=0, batch=1, depth=1, debug=0):
    return np.array((total, batch, depth, debug), RTPROP_DTYPE)


def scene_properties(direction=(-0.4, -1.0, -0.3), color=(10.0, 10.0, 10.0)):
    s = np.zeros((), SCENEPROP_DTYPE)
    s["lightCount"][0] = 1
    s["lights"][0]["direction"] = (*direction, 0.0)
    s["lights"][0]["color"] = (*color, 1.0)
    return s


class SceneBundle:
    """Host arrays of the sceneBuilder descriptor bundle (tools/sceneBuilder.cpp:69-101).

    meshes: list of (verts, tris); instances: list of (mesh_index, transform4x4, SBTOffset, customInstanceID).
    MeshInfo offsets are in scalar elements (x3), uv has stride 3, as in the reference."""

    def __init__(self, meshes, instances, materials, mesh_materials=None):
        self.meshes, self.instances = meshes, instances
        self.materials = np.array(materials, MATERIAL_DTYPE)
        nm = len(meshes)
        mesh_materials = mesh_materials or [0] * nm
        info = np.zeros(nm, MESHINFO_DTYPE)
        vs, ids, uvs, ns = [], [], [], []
        vo = io = 0
        for k, (v, t) in enumerate(meshes):
            info[k]["vertexOffset"] = info[k]["uvOffset"] = info[k]["normalOffset"] = vo * 3
            info[k]["indexOffset"] = io * 3
            info[k]["materialIndex"] = mesh_materials[k]
            vs.append(v)
            ids.append(t)
            ns.append(vertex_normals(v, t))
            uv = np.zeros((v.shape[0], 3), np.float32)
            uv[:, 0], uv[:, 1] = v[:, 0], v[:, 2]
            uvs.append(uv)
            vo += v.shape[0]
            io += t.shape[0]
        self.meshInfoData = info
        self.vertexData = np.ascontiguousarray(np.concatenate(vs), np.float32)
        self.indexData = np.ascontiguousarray(np.concatenate(ids), np.uint32)
        self.uvData = np.ascontiguousarray(np.concatenate(uvs), np.float32)
        self.normalData = np.ascontiguousarray(np.concatenate(ns), np.float32)

    def triangle_count(self):
        return sum(self.meshes[i[0]][1].shape[0] for i in self.instances)


def cornell_like(sphere_subdiv=3):
    """C1: 5 quad walls + 2 boxes + 1 UV-ish sphere, one instance per mesh (meshInfo[instanceIndex])."""
    w = 1.0
    meshes = [
        quad((-w, 0, -w), (w, 0, -w), (w, 0, w), (-w, 0, w)),          # floor
        quad((-w, 2 * w, -w), (-w, 2 * w, w), (w, 2 * w, w), (w, 2 * w, -w)),  # ceiling
        quad((-w, 0, -w), (-w, 2 * w, -w), (w, 2 * w, -w), (w, 0, -w)),  # back
        quad((-w, 0, -w), (-w, 0, w), (-w, 2 * w, w), (-w, 2 * w, -w)),  # left
        quad((w, 0, -w), (w, 2 * w, -w), (w, 2 * w, w), (w, 0, w)),      # right
        box((-0.3, -0.6, -0.3), (0.3, 0.6, 0.3)),
        box((-0.3, -0.3, -0.3), (0.3, 0.3, 0.3)),
        icosphere(sphere_subdiv, 0.35),
    ]
    mats = [material((0.73, 0.73, 0.73), 0.0, 0.9), material((0.65, 0.05, 0.05), 0.0, 0.9),
            material((0.12, 0.45, 0.15), 0.0, 0.9), material((0.8, 0.8, 0.85), 0.9, 0.25),
            material((0.9, 0.9, 0.9), 0.0, 0.1, 0.9, 1.45)]
    mesh_mats = [0, 0, 0, 1, 2, 0, 3, 4]
    eye = np.eye(4, dtype=np.float32)
    xf = [eye, eye, eye, eye, eye,
          translate(-0.35, 0.6, -0.3) @ rotate_y(0.3), translate(0.4, 0.3, 0.2) @ rotate_y(-0.35),
          translate(0.35, 0.95, 0.2)]
    instances = [(k, xf[k], 0, mesh_mats[k]) for k in range(len(meshes))]
    b = SceneBundle(meshes, instances, mats, mesh_mats)
    b.camera = camera(800, 600, (0.0, 1.0, 3.4), (0.0, 0.0, 0.0), 0.030, 0.036, 2.0, 0.0)
    b.scene = scene_properties((-0.35, -1.0, -0.45), (10.0, 10.0, 10.0))
    return b


def heightfield_scene(n, width=1920, height=1080, seed=1):
    """C2 (n=708) / C3 base (n=2237): one identity instance, camera of SURVEY 8(d)."""
    v, t = heightfield(n, seed)
    b = SceneBundle([(v, t)], [(0, np.eye(4, dtype=np.float32), 0, 0)], [material((0.7, 0.6, 0.5), 0.0, 0.6)])
    b.camera = camera(width, height, (0.0, 0.6, 0.9), (-0.6, 0.0, 0.0), 0.030, 0.036, 2.0, 0.0)
    b.scene = scene_properties()
    return b


def blob_field_scene(n=2237, width=1920, height=1080, seed=1, sphere_subdiv=5):
    """C3: n x n height-field (n=2237: 10 008 338 triangles) plus 3 instances of one displaced
    icosphere; 3 materials, customInstanceID = material index (tools/sceneBuilder.cpp:300-305)."""
    hv, ht = heightfield(n, seed)
    sv, st = icosphere(sphere_subdiv, 1.0, bumps=0.08, seed=seed)
    mats = [material((0.7, 0.6, 0.5), 0.0, 0.7), material((0.85, 0.85, 0.9), 0.9, 0.2), material((0.2, 0.5, 0.8), 0.0, 0.4)]
    inst = [(0, np.eye(4, dtype=np.float32), 0, 0),
            (1, translate(-0.22, 0.13, 0.05) @ scale(0.07), 0, 1),
            (1, translate(0.18, 0.12, -0.12) @ rotate_y(0.7) @ scale(0.06), 0, 2),
            (1, translate(0.02, 0.14, 0.22) @ rotate_y(-0.4) @ scale(0.05), 0, 1)]
    # MeshInfo is indexed by instanceIndex in user programs (samples/shader.cl:341): one record per instance
    b = SceneBundle([(hv, ht), (sv, st)], inst, mats, [0, 1])
    info = np.zeros(len(inst), MESHINFO_DTYPE)
    for k, (mi, _, _, custom) in enumerate(inst):
        info[k] = b.meshInfoData[mi]
        info[k]["materialIndex"] = custom
    b.meshInfoData = info
    b.camera = camera(width, height, (0.0, 0.6, 0.9), (-0.6, 0.0, 0.0), 0.030, 0.036, 2.0, 0.0)
    b.scene = scene_properties((-0.4, -1.0, -0.3), (6.0, 6.0, 6.0))
    return b


def groups_scene(n=2240, side=64, groups=64, width=1920, height=1080, seed=1):
    """C4: the n x n height-field cut into side x side spatial patches, one instance per patch,
    SBTOffset = hit group of the patch in a checkerboard (neighbouring patches never share a group),
    customInstanceID = material = group. n=2240, side=64: 4096 instances, 10 035 200 triangles."""
    assert n % side == 0
    c = n // side
    v, _ = heightfield(n, seed)
    grid = v.reshape(n + 1, n + 1, 3)
    # local index pattern of one (c+1) x (c+1) vertex patch, same winding as heightfield()
    g = np.arange(c, dtype=np.uint32)
    qx, qz = np.meshgrid(g, g, indexing="xy")
    q = (qz * (c + 1) + qx).reshape(-1)
    a, b, cc, d = q, q + 1, q + np.uint32(c + 1), q + np.uint32(c + 2)
    tris = np.empty((2 * c * c, 3), np.uint32)
    tris[0::2] = np.stack([a, cc, b], axis=-1)
    tris[1::2] = np.stack([b, cc, d], axis=-1)
    meshes, inst, mesh_mats = [], [], []
    gs = int(np.sqrt(groups))
    eye = np.eye(4, dtype=np.float32)
    for pz in range(side):
        for px in range(side):
            pv = np.ascontiguousarray(grid[pz * c:(pz + 1) * c + 1, px * c:(px + 1) * c + 1].reshape(-1, 3))
            k = ((px % gs) + gs * (pz % gs)) % groups
            meshes.append((pv, tris))
            inst.append((len(meshes) - 1, eye, k, k))
            mesh_mats.append(k)
    mats = [material((0.45 + 0.5 * ((k * 37) % 64) / 64.0, 0.45 + 0.5 * ((k * 11) % 64) / 64.0, 0.45 + 0.5 * ((k * 23) % 64) / 64.0),
                     0.9 if k % 8 == 0 else 0.0, 0.3 + 0.5 * ((k * 5) % 16) / 16.0) for k in range(groups)]
    bndl = SceneBundle(meshes, inst, mats, mesh_mats)
    bndl.camera = camera(width, height, (0.0, 0.6, 0.9), (-0.6, 0.0, 0.0), 0.030, 0.036, 2.0, 0.0)
    bndl.scene = scene_properties((-0.4, -1.0, -0.3), (6.0, 6.0, 6.0))
    return bndl


def replicated_scene(n=2237, copies=5, width=3840, height=2160, seed=1):
    """C5: `copies` instances of ONE n x n height-field BLAS laid out in a row along z (50 M
    triangles for n=2237, copies=5), 3 materials by instance, 4K camera looking down the row."""
    hv, ht = heightfield(n, seed)
    mats = [material((0.7, 0.6, 0.5), 0.0, 0.7), material((0.85, 0.85, 0.9), 0.9, 0.2), material((0.2, 0.5, 0.8), 0.0, 0.4)]
    inst = [(0, translate(0.0, 0.0, -1.0 * k), 0, k % 3) for k in range(copies)]
    b = SceneBundle([(hv, ht)], inst, mats, [0])
    info = np.zeros(len(inst), MESHINFO_DTYPE)
    for k, (mi, _, _, custom) in enumerate(inst):
        info[k] = b.meshInfoData[mi]
        info[k]["materialIndex"] = custom
    b.meshInfoData = info
    b.camera = camera(width, height, (0.0, 0.6, 0.9), (-0.35, 0.0, 0.0), 0.030, 0.036, 2.0, 0.0)
    b.scene = scene_properties((-0.4, -1.0, -0.3), (6.0, 6.0, 6.0))
    return b


def pinhole_rays(cam, tmin=0.001, tmax=1000.0, stride=1, jitter=None):
    """Explicit primary-ray list for hit-ID parity tests (float32, host-generated once and shared
    by oracle and GPU): pinhole of samples/shader.cl:111-153 with a fixed sub-pixel offset."""
    W, H = int(cam["widthPixel"]), int(cam["heightPixel"])
    ys, xs = np.meshgrid(np.arange(0, H, stride), np.arange(0, W, stride), indexing="ij")
    jx, jy = (0.5, 0.5) if jitter is None else jitter
    fx = ((xs + jx) / W - 0.5).astype(np.float64)
    fy = (0.5 - (ys + jy) / H).astype(np.float64)
    aspect = float(cam["heightPixel"]) / float(cam["widthPixel"])
    d = np.stack([fx * float(cam["sensorWidth"]), fy * float(cam["sensorWidth"]) * aspect,
                  np.full_like(fx, -float(cam["focalLength"]))], axis=-1)
    d /= np.linalg.norm(d, axis=-1, keepdims=True)

    def rx(a): c, s = np.cos(a), np.sin(a); return np.array([[1, 0, 0], [0, c, -s], [0, s, c]])
    def ry(a): c, s = np.cos(a), np.sin(a); return np.array([[c, 0, s], [0, 1, 0], [-s, 0, c]])
    def rz(a): c, s = np.cos(a), np.sin(a); return np.array([[c, -s, 0], [s, c, 0], [0, 0, 1]])
    R = rx(float(cam["wx"])) @ ry(float(cam["wy"])) @ rz(float(cam["wz"]))
    d = d.reshape(-1, 3) @ R.T
    rays = np.zeros((d.shape[0], 8), np.float32)
    rays[:, 0], rays[:, 1], rays[:, 2], rays[:, 3] = cam["x"], cam["y"], cam["z"], tmin
    rays[:, 4:7] = d.astype(np.float32)
    rays[:, 7] = tmax
    return rays


# ---- device upload / rendering helpers (through the RD mirror; used by tests, smoke() and bench.py) ----
import os as _os

PROGRAM_DIR = _os.path.join(_os.path.dirname(_os.path.abspath(__file__)), "programs")


def program_source(name):
    """(source, sbt_json or None) of a user program shipped in programs/."""
    src = open(_os.path.join(PROGRAM_DIR, name + ".cl")).read()
    sbt_path = _os.path.join(PROGRAM_DIR, name + "_sbt.json")
    return src, (open(sbt_path).read() if _os.path.exists(sbt_path) else None)


def build_scene(plt, bundle):
    """BuildAccelStruct for every mesh and for the instance list (samples/sample1.cpp-style)."""
    from . import rd
    blas = [rd.BuildAccelStruct(plt, rd.Mesh(v, t)) for (v, t) in bundle.meshes]
    inst = [rd.Instance(m, sbt, custom, blas[mi]) for (mi, m, sbt, custom) in bundle.instances]
    return rd.BuildAccelStruct(plt, inst), blas


class DeviceScene:
    """The 14-descriptor set of samples/sample1.cpp:387-394 / tools/sceneBuilder.h:8-17 on the device."""

    def __init__(self, plt, bundle, tlas, width, height, rt=None, cam=None):
        from . import rd
        self.plt, self.width, self.height = plt, width, height
        self.rtprop = np.array(rt if rt is not None else rtprop(0, 1, 1, 0), RTPROP_DTYPE).reshape(())
        self.cam = np.array(cam if cam is not None else bundle.camera, CAMERA_DTYPE).reshape(()).copy()
        self.cam["widthPixel"], self.cam["heightPixel"] = width, height
        n = width * height

        def buf(arr):
            arr = np.ascontiguousarray(arr)
            h = rd.CreateBuffer(plt, max(16, arr.nbytes))
            if arr.nbytes:
                rd.WriteBuffer(plt, h, arr.nbytes, arr)
            return h
        self.rdRTProp = buf(self.rtprop.reshape(1))
        self.rdImageScratch = rd.CreateBuffer(plt, n * 16)
        self.rdImage = rd.CreateImage(plt, width, height)
        self.rdCam = buf(self.cam.reshape(1))
        self.rdScene = buf(np.array(bundle.scene).reshape(1))
        self.rdMeshInfo, self.rdVertex, self.rdIndex = buf(bundle.meshInfoData), buf(bundle.vertexData), buf(bundle.indexData)
        self.rdUV, self.rdNormal, self.rdMaterial = buf(bundle.uvData), buf(bundle.normalData), buf(bundle.materials)
        self.rdTextures = rd.CreateImageArray(plt, 4, 4, 1)
        self.rdSampler = rd.CreateSampler(plt, rd.RD_ADDRESS_REPEAT, rd.RD_FILTER_LINEAR)
        self.tlas = tlas
        self.descriptors = rd.CreateDescriptorSet([
            self.rdRTProp, self.rdImageScratch, self.rdImage, self.rdCam, self.rdScene, self.rdMeshInfo, self.rdVertex,
            self.rdIndex, self.rdUV, self.rdNormal, self.rdMaterial, self.rdTextures, self.rdSampler, tlas])

    def set_rtprop(self, total, batch, depth, debug=0):
        from . import rd
        self.rtprop = rtprop(total, batch, depth, debug)
        rd.WriteBuffer(self.plt, self.rdRTProp, 16, self.rtprop.reshape(1))

    def read_image(self):
        from . import rd
        img = np.zeros((self.height, self.width, 4), np.uint8)
        rd.ReadBuffer(self.plt, self.rdImage, img.nbytes, img)
        return img

    def read_scratch(self):
        from . import rd
        s = np.zeros((self.height, self.width, 4), np.float32)
        rd.ReadBuffer(self.plt, self.rdImageScratch, s.nbytes, s)
        return s

    def host_args(self):
        """The same 14 arguments as host arrays for the CPU oracle's raygen (None = unused descriptor)."""
        n = self.width * self.height
        return {"rtprop": self.rtprop.reshape(1).copy(), "scratch": np.zeros(n * 4, np.float32), "image": np.zeros(n * 4, np.uint8),
                "cam": self.cam.reshape(1).copy()}


def make_pipeline(plt, program="primary", source=None, sbt=None):
    from . import rd
    if source is None:
        source, sbt = program_source(program)
    if sbt is not None:
        rd.LoadSBT(plt, sbt)
    module = rd.CreateShaderModule(plt, source, 0, "raygen")
    pipe = rd.CreatePipeline(rd.PipelineCreateInfo(2, [], [module], []))
    return pipe, module


def render_primary(plt, bundle, tlas, width, height, program="primary", batch=1, depth=1):
    """One TraceRays call of a shipped user program; returns the RGBA8 image (H,W,4)."""
    from . import rd
    ds = DeviceScene(plt, bundle, tlas, width, height, rtprop(0, batch, depth, 0))
    pipe, _ = make_pipeline(plt, program)
    rd.BindPipeline(plt, pipe)
    rd.BindDescriptorSet(plt, ds.descriptors)
    rd.TraceRays(plt, 0, 0, 0, width, height)
    return ds.read_image()
