"""glTF 2.0 / GLB scene ingestion: the host mirror of `RD::Scene::Load` (tools/sceneBuilder.cpp:27-315).

The reference reads assets through assimp (`aiProcess_GenSmoothNormals | Triangulate | JoinIdenticalVertices |
SortByPType`, sceneBuilder.cpp:32-36) and turns them into the 9-descriptor scene bundle of
tools/sceneBuilder.h:8-17. assimp is a link-time dependency that this image does not have, so the container format is
parsed here directly (glTF 2.0 is a JSON document plus binary buffers) and the importer conventions the bundle
depends on are restated from assimp's glTF2 importer (external/assimp/code/AssetLib/glTF2/glTF2Importer.cpp):

* one aiMesh per glTF *primitive*, numbered mesh by mesh (`meshOffsets`, glTF2Importer.cpp:463-477);
* the material list is the glTF list plus ONE default material at the end, used by primitives without a material
  (glTF2Importer.cpp:384-392, :880-882);
* TEXCOORD_0 with the v axis flipped, `v = 1 - v` (glTF2Importer.cpp:632); uv stored with stride 3;
* embedded images numbered in image order, materials refer to them as "*N" (glTF2Importer.cpp:1598-1620);
* node transform = `matrix`, else T * R * S (glTF2Importer.cpp:990-1015); several scene roots hang under an
  identity "ROOT" node (glTF2Importer.cpp:1259-1283);
* instances in depth-first order, a node's own primitives before its children, `customInstanceID = material index`,
  `SBTOffset = 0` (sceneBuilder.cpp:285-315).

Differences, stated: `meshInfoData` has one row per INSTANCE (the row of the instance's mesh), not one per mesh: the
shaders index it with `hitData->instanceIndex` (samples/shader.cl:310-435), which in the reference is only right
when instance i happens to use mesh i (sceneBuilder.cpp:69-101 pushes one row per mesh) -- for such assets the two
tables are the same bytes, for a mesh instanced twice or out of order the reference reads the wrong row or past the
end. The per-mesh table is kept as `ParsedScene.meshInfoPerMesh`. Identical vertices are not joined (index values differ from assimp's, geometry does not);
textures are resized with Pillow's bilinear filter instead of stb_image_resize2 (not bit-identical).
`parse_gltf` is pure host code (numpy); `Scene.Load` uploads through the `rd` mirror and builds on the GPU.
"""
import base64
import io
import json
import os
import struct

import numpy as np

from .scenes import MATERIAL_DTYPE, MESHINFO_DTYPE

TEX_DIM = 4096            # sceneBuilder.cpp:22
_COMPONENT = {5120: np.int8, 5121: np.uint8, 5122: np.int16, 5123: np.uint16, 5125: np.uint32, 5126: np.float32}
_WIDTH = {"SCALAR": 1, "VEC2": 2, "VEC3": 3, "VEC4": 4, "MAT2": 4, "MAT3": 9, "MAT4": 16}


class SceneFormatError(ValueError):
    pass


def _read_container(path):
    """Returns (json document, list of buffer bytes) for a .glb or a .gltf file."""
    raw = open(path, "rb").read()
    if raw[:4] == b"glTF":
        magic, version, length = struct.unpack_from("<4sII", raw, 0)
        if version != 2:
            raise SceneFormatError(f"{path}: GLB container version {version}, only 2 is supported")
        if length > len(raw):
            raise SceneFormatError(f"{path}: truncated GLB ({len(raw)} of {length} bytes)")
        doc, binary, at = None, None, 12
        while at + 8 <= length:
            clen, ctype = struct.unpack_from("<II", raw, at)
            body = raw[at + 8: at + 8 + clen]
            if len(body) != clen:
                raise SceneFormatError(f"{path}: truncated GLB chunk at byte {at}")
            if ctype == 0x4E4F534A:
                doc = json.loads(body.decode("utf-8"))
            elif ctype == 0x004E4942 and binary is None:
                binary = body
            at += 8 + clen + (-clen % 4)
        if doc is None:
            raise SceneFormatError(f"{path}: GLB without a JSON chunk")
    else:
        doc, binary = json.loads(raw.decode("utf-8")), None
    buffers = []
    for k, b in enumerate(doc.get("buffers", [])):
        uri = b.get("uri")
        if uri is None:
            if binary is None:
                raise SceneFormatError(f"{path}: buffer {k} has no uri and the file has no BIN chunk")
            data = binary
        elif uri.startswith("data:"):
            data = base64.b64decode(uri.split(",", 1)[1])
        else:
            data = open(os.path.join(os.path.dirname(path), uri), "rb").read()
        if len(data) < b["byteLength"]:
            raise SceneFormatError(f"{path}: buffer {k} holds {len(data)} bytes, {b['byteLength']} declared")
        buffers.append(data)
    return doc, buffers


def _accessor(doc, buffers, index):
    """Accessor -> numpy array (count, width), honouring byteOffset / byteStride / normalized."""
    a = doc["accessors"][index]
    if "sparse" in a:
        raise SceneFormatError(f"accessor {index}: sparse accessors are not supported")
    dt, width, count = np.dtype(_COMPONENT[a["componentType"]]), _WIDTH[a["type"]], a["count"]
    if "bufferView" not in a:
        return np.zeros((count, width), dt)
    view = doc["bufferViews"][a["bufferView"]]
    start = view.get("byteOffset", 0) + a.get("byteOffset", 0)
    elem = dt.itemsize * width
    stride = view.get("byteStride", 0) or elem
    buf = buffers[view["buffer"]]
    if start + stride * (count - 1) + elem > len(buf):
        raise SceneFormatError(f"accessor {index}: reads past the end of buffer {view['buffer']}")
    flat = np.frombuffer(buf, np.uint8, offset=start, count=stride * (count - 1) + elem)
    out = np.lib.stride_tricks.as_strided(flat, (count, elem), (stride, 1)).copy().view(dt).reshape(count, width)
    if a.get("normalized") and dt.kind in "iu":
        info = np.iinfo(dt)
        out = np.maximum(out.astype(np.float32) / np.float32(info.max), np.float32(-1.0))
    return out


def _smooth_normals(verts, tris):
    """aiProcess_GenSmoothNormals for meshes without NORMAL: area-weighted face normals summed over all vertices that
    share a position, then normalised."""
    v = verts.astype(np.float64)
    fn = np.cross(v[tris[:, 1]] - v[tris[:, 0]], v[tris[:, 2]] - v[tris[:, 0]])
    _, group = np.unique(verts, axis=0, return_inverse=True)
    group = group.reshape(-1)
    acc = np.zeros((group.max() + 1, 3))
    for k in range(3):
        for c in range(3):
            acc[:, c] += np.bincount(group[tris[:, k]], weights=fn[:, c], minlength=acc.shape[0])
    n = acc[group]
    ln = np.linalg.norm(n, axis=1, keepdims=True)
    ln[ln == 0] = 1.0
    return (n / ln).astype(np.float32)


def _quat_matrix(q):
    x, y, z, w = (float(c) for c in q)
    return np.array([[1 - 2 * (y * y + z * z), 2 * (x * y - z * w), 2 * (x * z + y * w), 0],
                     [2 * (x * y + z * w), 1 - 2 * (x * x + z * z), 2 * (y * z - x * w), 0],
                     [2 * (x * z - y * w), 2 * (y * z + x * w), 1 - 2 * (x * x + y * y), 0],
                     [0, 0, 0, 1]], np.float32)


def _node_transform(node):
    """Row-major 4x4 acting on column vectors, like aiMatrix4x4 (glTF2Importer.cpp:990-1015)."""
    if "matrix" in node:
        return np.array(node["matrix"], np.float32).reshape(4, 4).T.copy()    # glTF stores column-major
    m = np.eye(4, dtype=np.float32)
    if "translation" in node:
        t = np.eye(4, dtype=np.float32)
        t[:3, 3] = node["translation"]
        m = m @ t
    if "rotation" in node:
        m = m @ _quat_matrix(node["rotation"])
    if "scale" in node:
        m = m @ np.diag(np.array(list(node["scale"]) + [1.0], np.float32))
    return m.astype(np.float32)


class ParsedScene:
    """Host arrays of one asset in the layout of the 9-descriptor bundle (sceneBuilder.cpp:69-101, :199-221).

    meshes [(verts float32 (n,3), tris uint32 (m,3))], instances [(mesh, transform 4x4, SBTOffset, customInstanceID)],
    meshInfoData / vertexData / indexData / uvData / normalData / materials as uploaded, images = encoded bytes of
    the embedded images in "*N" order."""

    def triangle_count(self):
        return sum(self.meshes[i[0]][1].shape[0] for i in self.instances)


def _texture_image(doc, embedded, tex_info):
    if tex_info is None:
        return -1
    image = doc["textures"][tex_info["index"]].get("source")
    if image is None or embedded[image] < 0:
        raise SceneFormatError(f"texture {tex_info['index']}: only embedded images are supported "
                               "(sceneBuilder.cpp:127 asserts the \"*N\" form as well)")
    return embedded[image]


def _materials(doc, embedded):
    """One row per glTF material plus the importer's default material (sceneBuilder.cpp:111-197)."""
    out = np.zeros(len(doc.get("materials", [])) + 1, MATERIAL_DTYPE)
    for k, m in enumerate(list(doc.get("materials", [])) + [{}]):
        pbr, ext, row = m.get("pbrMetallicRoughness", {}), m.get("extensions", {}), out[k]
        row["albedo"] = pbr.get("baseColorFactor", [1.0, 1.0, 1.0, 1.0])
        row["albedoTex"] = _texture_image(doc, embedded, pbr.get("baseColorTexture"))
        row["metallic"] = pbr.get("metallicFactor", 1.0)
        row["roughness"] = pbr.get("roughnessFactor", 1.0)
        row["metallicTex"] = row["roughnessTex"] = _texture_image(doc, embedded, pbr.get("metallicRoughnessTexture"))
        row["transmission"] = ext.get("KHR_materials_transmission", {}).get("transmissionFactor", 0.0)
        row["ior"] = ext.get("KHR_materials_ior", {}).get("ior", 1.45)
        row["normalTex"] = _texture_image(doc, embedded, m.get("normalTexture"))
    return out


def parse_gltf(path):
    doc, buffers = _read_container(path)
    if str(doc.get("asset", {}).get("version", "2.0")).split(".")[0] != "2":
        raise SceneFormatError(f"{path}: glTF asset version {doc['asset'].get('version')} (2.x expected)")
    s = ParsedScene()

    # embedded images in image order -> "*N"
    embedded, s.images = [], []
    for img in doc.get("images", []):
        if "bufferView" in img:
            view = doc["bufferViews"][img["bufferView"]]
            at = view.get("byteOffset", 0)
            s.images.append(bytes(buffers[view["buffer"]][at: at + view["byteLength"]]))
        elif str(img.get("uri", "")).startswith("data:"):
            s.images.append(base64.b64decode(img["uri"].split(",", 1)[1]))
        else:
            embedded.append(-1)
            continue
        embedded.append(len(s.images) - 1)
    s.materials = _materials(doc, embedded)
    default_material = len(s.materials) - 1

    # one mesh per primitive
    s.meshes, mesh_offsets, info, uvs, normals = [], [], [], [], []
    vo = io_ = 0
    for mi, mesh in enumerate(doc.get("meshes", [])):
        mesh_offsets.append(len(s.meshes))
        for pi, prim in enumerate(mesh["primitives"]):
            if prim.get("mode", 4) != 4:
                raise SceneFormatError(f"mesh {mi} primitive {pi}: mode {prim['mode']}; only triangle lists (4) reach "
                                       "the tracer (sceneBuilder.cpp:93 asserts 3 indices per face)")
            attr = prim["attributes"]
            if "POSITION" not in attr:
                raise SceneFormatError(f"mesh {mi} primitive {pi}: no POSITION attribute")
            v = np.ascontiguousarray(_accessor(doc, buffers, attr["POSITION"]), np.float32)
            if "indices" in prim:
                idx = _accessor(doc, buffers, prim["indices"]).reshape(-1).astype(np.uint32)
            else:
                idx = np.arange(v.shape[0], dtype=np.uint32)
            if idx.size % 3:
                raise SceneFormatError(f"mesh {mi} primitive {pi}: {idx.size} indices is not a multiple of 3")
            t = idx.reshape(-1, 3)
            if t.size and int(t.max()) >= v.shape[0]:
                raise SceneFormatError(f"mesh {mi} primitive {pi}: index {int(t.max())} >= {v.shape[0]} vertices")
            if "NORMAL" in attr:
                n = np.ascontiguousarray(_accessor(doc, buffers, attr["NORMAL"]), np.float32)
            else:
                n = _smooth_normals(v, t)
            uv = np.zeros((v.shape[0], 3), np.float32)
            if "TEXCOORD_0" in attr:
                tc = _accessor(doc, buffers, attr["TEXCOORD_0"]).astype(np.float32)
                uv[:, 0], uv[:, 1] = tc[:, 0], np.float32(1.0) - tc[:, 1]
            material = prim.get("material", default_material)
            info.append((vo * 3, io_ * 3, vo * 3, vo * 3, material))
            s.meshes.append((v, np.ascontiguousarray(t)))
            uvs.append(uv)
            normals.append(n)
            vo += v.shape[0]
            io_ += t.shape[0]
    mesh_offsets.append(len(s.meshes))
    if not s.meshes:
        raise SceneFormatError(f"{path}: no meshes")
    s.meshInfoPerMesh = np.zeros(len(info), MESHINFO_DTYPE)
    for k, (a, b, c, d, m) in enumerate(info):
        s.meshInfoPerMesh[k] = (a, b, c, d, m, (0, 0, 0))
    s.vertexData = np.ascontiguousarray(np.concatenate([m[0] for m in s.meshes]), np.float32)
    s.indexData = np.ascontiguousarray(np.concatenate([m[1] for m in s.meshes]), np.uint32)
    s.uvData = np.ascontiguousarray(np.concatenate(uvs), np.float32)
    s.normalData = np.ascontiguousarray(np.concatenate(normals), np.float32)

    # instances: depth-first, own primitives first (sceneBuilder.cpp:285-315)
    s.instances = []
    nodes = doc.get("nodes", [])

    def visit(index, parent, trail):
        if index in trail:
            raise SceneFormatError(f"node {index}: cycle in the node hierarchy")
        node = nodes[index]
        current = (parent @ _node_transform(node)).astype(np.float32)
        if "mesh" in node:
            for m in range(mesh_offsets[node["mesh"]], mesh_offsets[node["mesh"] + 1]):
                s.instances.append((m, current, 0, int(s.meshInfoPerMesh[m]["materialIndex"])))
        for child in node.get("children", []):
            visit(child, current, trail | {index})

    scenes = doc.get("scenes", [])
    roots = scenes[doc.get("scene", 0)].get("nodes", []) if scenes else []
    for r in roots:
        visit(r, np.eye(4, dtype=np.float32), frozenset())
    if not s.instances:
        raise SceneFormatError(f"{path}: the default scene instantiates no mesh")
    s.meshInfoData = np.ascontiguousarray(s.meshInfoPerMesh[[i[0] for i in s.instances]])   # one row per instance
    return s


def decode_texture(encoded, dim=TEX_DIM):
    """Embedded PNG / JPEG -> (dim, dim, 4) uint8, as sceneBuilder.cpp:49-67 (decode to RGBA, resize to the fixed
    layer size)."""
    from PIL import Image
    img = Image.open(io.BytesIO(encoded)).convert("RGBA")
    if img.size != (dim, dim):
        img = img.resize((dim, dim), Image.BILINEAR)
    return np.ascontiguousarray(np.asarray(img, np.uint8))


class Scene:
    """`RD::Scene` (tools/sceneBuilder.h:34-56): the nine trailing descriptors of a raygen argument list."""

    LAYOUT = ("BUFFER_TYPE",) * 6 + ("TEX_ARRAY_TYPE", "IMAGE_SAMPLER_TYPE", "ACCEL_STRUCT_TYPE")   # INCLUDE_SCENE_LAYOUT

    def descriptors(self):
        """INCLUDE_SCENE_DESC(scene) (sceneBuilder.h:8-17)."""
        return [self.meshInfoData, self.vertexData, self.indexData, self.uvData, self.normalData, self.materialData,
                self.textureData, self.sampler, self.topAccelStruct]

    @staticmethod
    def Load(path, plt, loadFromCache=False, tex_dim=TEX_DIM):
        """Parse, upload, build (or read `<path>.cache`, sceneBuilder.cpp:224-229); writes `<path>.cache` in the
        reference's blob format after a build (sceneBuilder.cpp:262-263)."""
        from . import rd
        parsed = parse_gltf(path)
        s = Scene()
        s.parsed = parsed
        s.textureData = rd.CreateImageArray(plt, tex_dim, tex_dim, max(1, len(parsed.images)))
        s.sampler = rd.CreateSampler(plt, rd.RD_ADDRESS_REPEAT, rd.RD_FILTER_LINEAR)
        for k, enc in enumerate(parsed.images):
            rd.WriteImage(plt, s.textureData, tex_dim, tex_dim, k, decode_texture(enc, tex_dim))

        def buf(arr):
            h = rd.CreateBuffer(plt, max(16, arr.nbytes))
            rd.WriteBuffer(plt, h, arr.nbytes, arr)
            return h
        s.meshInfoData, s.vertexData, s.indexData = buf(parsed.meshInfoData), buf(parsed.vertexData), buf(parsed.indexData)
        s.uvData, s.normalData, s.materialData = buf(parsed.uvData), buf(parsed.normalData), buf(parsed.materials)
        cache = path + ".cache"
        if loadFromCache:
            s.topAccelStruct = rd.FileToTopAccelStruct(plt, cache)
            s.bottomAccelStructs = []
        else:
            s.bottomAccelStructs = [rd.BuildAccelStruct(plt, rd.Mesh(v, t)) for (v, t) in parsed.meshes]
            inst = [rd.Instance(m, sbt, custom, s.bottomAccelStructs[mi]) for (mi, m, sbt, custom) in parsed.instances]
            s.topAccelStruct = rd.BuildAccelStruct(plt, inst)
            rd.TopAccelStructToFile(plt, s.topAccelStruct, cache)
        return s
