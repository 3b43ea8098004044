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


def _materials(doc, embedded, order):
    """One row per glTF material the scene references, in the order the importer first meets them, plus the importer's
    default material (sceneBuilder.cpp:111-197). The transmission / ior extensions count only when the asset declares them in
    extensionsUsed (glTF2Asset.inl:1360, :1384)."""
    used = set(doc.get("extensionsUsed", []))
    listed = [doc["materials"][k] for k in order]
    out = np.zeros(len(listed) + 1, MATERIAL_DTYPE)
    for k, m in enumerate(listed + [{}]):
        pbr, ext, row = m.get("pbrMetallicRoughness", {}), m.get("extensions", {}), out[k]
        row["albedo"] = pbr.get("baseColorFactor", [1.0, 1.0, 1.0, 1.0])
        row["albedoTex"] = _texture_image(doc, embedded, pbr.get("baseColorTexture"))
        row["metallic"] = pbr.get("metallicFactor", 1.0)
        row["roughness"] = pbr.get("roughnessFactor", 1.0)
        row["metallicTex"] = row["roughnessTex"] = _texture_image(doc, embedded, pbr.get("metallicRoughnessTexture"))
        row["transmission"] = (ext.get("KHR_materials_transmission", {}).get("transmissionFactor", 0.0)
                               if "KHR_materials_transmission" in used else 0.0)
        row["ior"] = ext.get("KHR_materials_ior", {}).get("ior", 1.45) if "KHR_materials_ior" in used else 1.45
        row["normalTex"] = _texture_image(doc, embedded, m.get("normalTexture"))
    return out


def _importer_normals(v, t):
    """aiProcess_GenSmoothNormals as assimp runs it on the importer's INDEXED mesh (GenVertexNormalsProcess.cpp:104-235, default
    smoothing angle 175 degrees): every face writes its unit normal ((v1 - v0) x (v2 - v0), NormalizeSafe) to its three vertices --
    for a vertex shared by several faces the LAST face in index order wins -- and vertices at the same position then share the
    normalised sum of those per-vertex normals. (assimp groups positions within 1e-4 of the mesh's extent; exact equality here.)"""
    p0, p1, p2 = v[t[:, 0]].astype(np.float32), v[t[:, 1]].astype(np.float32), v[t[:, 2]].astype(np.float32)
    fn = np.cross(p1 - p0, p2 - p0).astype(np.float32)
    length = np.sqrt((fn * fn).sum(1, dtype=np.float32)).astype(np.float32)
    fn = np.where(length[:, None] > 0, fn / np.where(length > 0, length, 1)[:, None], fn).astype(np.float32)
    per_vertex = np.zeros((v.shape[0], 3), np.float32)
    flat = t.reshape(-1)
    per_vertex[flat] = np.repeat(fn, 3, axis=0)            # numpy keeps the last assignment of a repeated index: the last face wins
    keys = np.ascontiguousarray(v, np.float32).view([("", np.float32)] * 3).reshape(-1)
    _, inverse = np.unique(keys, return_inverse=True)
    inverse = inverse.reshape(-1)
    sums = np.zeros((inverse.max() + 1, 3), np.float32)
    np.add.at(sums, inverse, per_vertex)
    length = np.sqrt((sums * sums).sum(1, dtype=np.float32)).astype(np.float32)
    sums = np.where(length[:, None] > 0, sums / np.where(length > 0, length, 1)[:, None], sums).astype(np.float32)
    return np.ascontiguousarray(sums[inverse], np.float32)


def _importer_vertex_order(v, t, n, uv):
    """The vertex numbering assimp hands to sceneBuilder.cpp: the glTF2 importer keeps only the vertices the index list uses and
    numbers them in the order the index list first mentions them (glTF2Importer.cpp ImportMeshes, vertexRemappingTable); missing
    normals are generated on that mesh (_importer_normals; pass n = None); then aiProcess_JoinIdenticalVertices merges vertices
    whose position, normal and uv are identical, keeping the first (JoinVerticesProcess.cpp:258-381; its comparison has a 1e-5
    tolerance, exact equality is used here). Triangles keep their order."""
    flat = t.reshape(-1).astype(np.int64)
    if flat.size == 0:
        return v[:0], t, (n if n is not None else v)[:0], uv[:0]
    order = flat[np.sort(np.unique(flat, return_index=True)[1])]                 # vertex ids by first appearance
    remap = np.full(v.shape[0], -1, np.int64)
    remap[order] = np.arange(len(order))
    v, uv, t = v[order], uv[order], remap[t]
    n = n[order] if n is not None else _importer_normals(v, t)
    attr = np.ascontiguousarray(np.concatenate([v, n, uv], 1), np.float32) + np.float32(0.0)   # -0 and +0 are one value
    rows = attr.view([("", attr.dtype)] * attr.shape[1]).reshape(-1)
    _, first, inverse = np.unique(rows, return_index=True, return_inverse=True)
    keep = np.sort(first)                                                        # first occurrence of every distinct vertex
    rank = np.empty(len(first), np.int64)
    rank[np.argsort(first)] = np.arange(len(first))
    joined = rank[inverse.reshape(-1)]
    return (np.ascontiguousarray(v[keep]), np.ascontiguousarray(joined[t].astype(np.uint32)),
            np.ascontiguousarray(n[keep]), np.ascontiguousarray(uv[keep]))


def _retrieval_order(doc):
    """assimp's glTF2 importer loads only what the default scene references, in the order it first meets it (LazyDict::Retrieve):
    Node::Read (glTF2Asset.inl:1692-1760) retrieves a node's children -- depth first -- BEFORE the node's own mesh, Mesh::Read
    retrieves each primitive's material, Material::Read (:1283-1300) the baseColor, metallicRoughness, normal, occlusion and
    emissive textures and through them the images. aiMesh / aiMaterial / "*N" numbering follows that order, not the file's
    (pinned against the reference's own loader by tests/test_scene_loader.py). Returns (meshes, materials, images)."""
    nodes, meshes_d, materials_d, textures_d = doc.get("nodes", []), doc.get("meshes", []), doc.get("materials", []), doc.get("textures", [])
    meshes, materials, images, seen = [], [], [], set()

    def texture(info):
        if info is None:
            return
        source = textures_d[info["index"]].get("source")
        if source is not None and source not in images:
            images.append(source)

    def material(k):
        if k in materials:
            return
        materials.append(k)
        m = materials_d[k]
        pbr = m.get("pbrMetallicRoughness", {})
        for info in (pbr.get("baseColorTexture"), pbr.get("metallicRoughnessTexture"), m.get("normalTexture"),
                     m.get("occlusionTexture"), m.get("emissiveTexture")):
            texture(info)

    def node(i):
        if i in seen:
            return
        seen.add(i)
        for c in nodes[i].get("children", []):
            node(c)
        mi = nodes[i].get("mesh")
        if mi is not None and mi not in meshes:
            meshes.append(mi)
            for prim in meshes_d[mi]["primitives"]:
                if "material" in prim:
                    material(prim["material"])

    scenes = doc.get("scenes", [])
    for r in (scenes[doc.get("scene", 0)].get("nodes", []) if scenes else []):
        node(r)
    return meshes, materials, images


def parse_gltf(path):
    doc, buffers = _read_container(path)
    if str(doc.get("asset", {}).get("version", "2.0")).split(".")[0] != "2":
        raise SceneFormatError(f"{path}: glTF asset version {doc['asset'].get('version')} (2.x expected)")
    s = ParsedScene()

    mesh_order, material_order, image_order = _retrieval_order(doc)
    # embedded images, in the order the importer meets them -> "*N"
    embedded, s.images = [-1] * len(doc.get("images", [])), []
    for k in image_order:
        img = doc["images"][k]
        if "bufferView" in img:
            view = doc["bufferViews"][img["bufferView"]]
            at = view.get("byteOffset", 0)
            s.images.append(bytes(buffers[view["buffer"]][at: at + view["byteLength"]]))
        elif str(img.get("uri", "")).startswith("data:"):
            s.images.append(base64.b64decode(img["uri"].split(",", 1)[1]))
        else:
            continue
        embedded[k] = len(s.images) - 1
    s.materials = _materials(doc, embedded, material_order)
    default_material = len(s.materials) - 1
    material_row = {k: row for row, k in enumerate(material_order)}

    # one mesh per primitive
    s.meshes, mesh_offsets, info, uvs, normals = [], {}, [], [], []
    vo = io_ = 0
    for mi in mesh_order:
        mesh = doc["meshes"][mi]
        first = len(s.meshes)
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
            n = np.ascontiguousarray(_accessor(doc, buffers, attr["NORMAL"]), np.float32) if "NORMAL" in attr else None
            uv = np.zeros((v.shape[0], 3), np.float32)
            if "TEXCOORD_0" in attr:
                tc = _accessor(doc, buffers, attr["TEXCOORD_0"]).astype(np.float32)
                uv[:, 0], uv[:, 1] = tc[:, 0], np.float32(1.0) - tc[:, 1]
            v, t, n, uv = _importer_vertex_order(v, t, n, uv)
            material = material_row[prim["material"]] if "material" in prim else default_material
            info.append((vo * 3, io_ * 3, vo * 3, vo * 3, material))
            s.meshes.append((v, np.ascontiguousarray(t)))
            uvs.append(uv)
            normals.append(n)
            vo += v.shape[0]
            io_ += t.shape[0]
        mesh_offsets[mi] = (first, len(s.meshes))
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
            for m in range(*mesh_offsets[node["mesh"]]):
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


def _resize_kernel(x, upsample):
    """stb_image_resize2's default filters: Catmull-Rom when enlarging, Mitchell-Netravali (B = C = 1/3) when shrinking."""
    x = np.abs(x)
    if upsample:
        return np.where(x < 1, 1 - x * x * (2.5 - 1.5 * x), np.where(x < 2, 2 - x * (4 + x * (0.5 * x - 2.5)), 0.0))
    return np.where(x < 1, (16 + x * x * (21 * x - 36)) / 18, np.where(x < 2, (32 + x * (-60 + x * (36 - 7 * x))) / 18, 0.0))


def _resize_axis(a, n_out, axis):
    """One separable pass of stbir_resize_uint8_linear(..., STBIR_RGBA) as tools/sceneBuilder.cpp:60-62 calls it (edge clamp,
    texel centres at i + 0.5, filter widened by 1 / scale when shrinking, weights normalised). Against the reference's own stb
    code: identical for the 16 -> 4096 enlargement, within one level on 1.4e-5 of the texels for an 8192 -> 4096 reduction."""
    n_in = a.shape[axis]
    if n_in == n_out:
        return a
    scale = n_out / n_in
    upsample = scale >= 1
    support, kscale = (2.0, 1.0) if upsample else (2.0 / scale, scale)
    centre = (np.arange(n_out) + 0.5) / scale            # input-space position of every output texel centre
    lo = np.floor(centre - support + 0.5).astype(np.int64)
    taps = int(np.ceil(2 * support)) + 1
    shape = [1] * a.ndim
    shape[axis] = n_out
    out = np.zeros(a.shape[:axis] + (n_out,) + a.shape[axis + 1:], np.float32)
    wsum = np.zeros(n_out, np.float64)
    for k in range(taps):
        idx = lo + k
        w = _resize_kernel((centre - (idx + 0.5)) * kscale, upsample)
        out += np.take(a, np.clip(idx, 0, n_in - 1), axis=axis) * w.astype(np.float32).reshape(shape)
        wsum += w
    return out / wsum.astype(np.float32).reshape(shape)


def decode_texture(encoded, dim=TEX_DIM):
    """Embedded PNG / JPEG -> (dim, dim, 4) uint8, as sceneBuilder.cpp:49-67: decode to RGBA (stbi_load_from_memory), resize to
    the fixed layer size with stb_image_resize2's defaults (_resize_axis). Alpha is taken as opaque for the filter (STBIR_RGBA
    premultiplies; the reference's materials never use the alpha of these layers)."""
    from PIL import Image
    img = np.asarray(Image.open(io.BytesIO(encoded)).convert("RGBA"), np.uint8)
    if img.shape[:2] == (dim, dim):
        return np.ascontiguousarray(img)
    wide = _resize_axis(img.astype(np.float32), dim, 1)              # (h, dim, 4)
    out = np.empty((dim, dim, 4), np.uint8)
    block = max(1, (1 << 24) // (dim * 4))                           # rows of the result per step: bounds the float temporaries
    n_in = wide.shape[0]
    if n_in == dim:
        full = wide
        np.clip(np.floor(full + np.float32(0.5)), 0, 255, out=full)
        return np.ascontiguousarray(full.astype(np.uint8))
    scale = dim / n_in
    support = 2.0 if scale >= 1 else 2.0 / scale
    for r0 in range(0, dim, block):
        r1 = min(dim, r0 + block)
        # input rows this block of output rows can touch
        c0, c1 = (r0 + 0.5) / scale, (r1 - 0.5) / scale
        i0, i1 = max(0, int(np.floor(c0 - support - 1))), min(n_in, int(np.ceil(c1 + support + 1)) + 1)
        part = _resize_rows(wide, i0, i1, r0, r1, dim)
        out[r0:r1] = np.clip(np.floor(part + np.float32(0.5)), 0, 255).astype(np.uint8)
    return out


def _resize_rows(wide, i0, i1, r0, r1, dim):
    """output rows [r0, r1) of the vertical pass; rows outside [i0, i1) cannot contribute except through the edge clamp"""
    n_in = wide.shape[0]
    scale = dim / n_in
    upsample = scale >= 1
    support, kscale = (2.0, 1.0) if upsample else (2.0 / scale, scale)
    centre = (np.arange(r0, r1) + 0.5) / scale
    lo = np.floor(centre - support + 0.5).astype(np.int64)
    taps = int(np.ceil(2 * support)) + 1
    out = np.zeros((r1 - r0,) + wide.shape[1:], np.float32)
    wsum = np.zeros(r1 - r0, np.float64)
    for k in range(taps):
        idx = lo + k
        w = _resize_kernel((centre - (idx + 0.5)) * kscale, upsample)
        out += wide[np.clip(idx, 0, n_in - 1)] * w.astype(np.float32)[:, None, None]
        wsum += w
    return out / wsum.astype(np.float32)[:, None, None]


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
