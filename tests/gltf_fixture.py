"""Writes small glTF 2.0 assets for the scene-loader tests (the reference ships no .glb).

The asset exercises what `RD::Scene::Load` (tools/sceneBuilder.cpp:27-315) depends on: several primitives per mesh,
u8 / u16 / u32 indices, an interleaved buffer view with byteStride, a primitive without NORMAL and without material,
TRS and matrix nodes, a two-level hierarchy, two scene roots, an embedded PNG, the transmission / ior extensions.
`expected` holds the arrays the loader must produce, computed here from the source arrays, not by the loader."""
import base64
import io
import json
import struct

import numpy as np


def _pad4(b, fill=b"\0"):
    return b + fill * (-len(b) % 4)


def _quat(axis, angle):
    axis = np.asarray(axis, np.float64) / np.linalg.norm(axis)
    return [float(c) for c in axis * np.sin(angle / 2)] + [float(np.cos(angle / 2))]


def _rot(axis, angle):
    x, y, z = np.asarray(axis, np.float64) / np.linalg.norm(axis)
    c, s = np.cos(angle), np.sin(angle)
    return np.array([[c + x * x * (1 - c), x * y * (1 - c) - z * s, x * z * (1 - c) + y * s, 0],
                     [y * x * (1 - c) + z * s, c + y * y * (1 - c), y * z * (1 - c) - x * s, 0],
                     [z * x * (1 - c) - y * s, z * y * (1 - c) + x * s, c + z * z * (1 - c), 0],
                     [0, 0, 0, 1]])


def checker_png(size=16):
    from PIL import Image
    yy, xx = np.meshgrid(np.arange(size), np.arange(size), indexing="ij")
    px = np.zeros((size, size, 4), np.uint8)
    px[..., 0] = (xx * 255) // (size - 1)
    px[..., 1] = (yy * 255) // (size - 1)
    px[..., 2] = ((xx // 4 + yy // 4) & 1) * 255
    px[..., 3] = 255
    out = io.BytesIO()
    Image.fromarray(px, "RGBA").save(out, format="PNG")
    return out.getvalue(), px


def build_asset(scenes):
    """Returns (json document, binary chunk, expected dict). `scenes` is the package's scenes module (geometry)."""
    blob = bytearray()
    views, accessors = [], []

    def view(data, stride=None):
        while len(blob) % 4:
            blob.append(0)
        v = {"buffer": 0, "byteOffset": len(blob), "byteLength": len(data)}
        if stride:
            v["byteStride"] = stride
        blob.extend(data)
        views.append(v)
        return len(views) - 1

    def accessor(view_index, ctype, count, kind, offset=0, **extra):
        a = {"bufferView": view_index, "componentType": ctype, "count": count, "type": kind, "byteOffset": offset}
        a.update(extra)
        accessors.append(a)
        return len(accessors) - 1

    # mesh 0 / primitive 0: floor grid, NORMAL + TEXCOORD_0, u16 indices, textured material 0
    g = 6
    gx, gz = np.meshgrid(np.linspace(-1, 1, g + 1), np.linspace(-1, 1, g + 1), indexing="xy")
    fv = np.stack([gx, 0.05 * np.sin(3 * gx) * np.cos(2 * gz), gz], -1).reshape(-1, 3).astype(np.float32)
    q = (np.arange(g)[:, None] * (g + 1) + np.arange(g)[None, :]).reshape(-1)
    ft = np.concatenate([np.stack([q, q + g + 1, q + 1], -1), np.stack([q + 1, q + g + 1, q + g + 2], -1)]).astype(np.uint32)
    fn = scenes.vertex_normals(fv, ft)
    fuv = np.stack([(fv[:, 0] + 1) * 0.9, (fv[:, 2] + 1) * 0.7], -1).astype(np.float32)
    p_floor = {"attributes": {"POSITION": accessor(view(fv.tobytes()), 5126, len(fv), "VEC3", min=fv.min(0).tolist(), max=fv.max(0).tolist()),
                              "NORMAL": accessor(view(fn.tobytes()), 5126, len(fn), "VEC3"),
                              "TEXCOORD_0": accessor(view(fuv.tobytes()), 5126, len(fuv), "VEC2")},
               "indices": accessor(view(ft.astype(np.uint16).tobytes()), 5123, ft.size, "SCALAR"), "material": 0, "mode": 4}

    # mesh 0 / primitive 1: box, no NORMAL, no material, u8 indices
    bv, bt = scenes.box((-0.3, 0.0, -0.3), (0.3, 0.6, 0.3))
    p_box = {"attributes": {"POSITION": accessor(view(bv.tobytes()), 5126, len(bv), "VEC3")},
             "indices": accessor(view(bt.astype(np.uint8).tobytes()), 5121, bt.size, "SCALAR")}

    # mesh 1: icosphere, POSITION + NORMAL interleaved (byteStride 24), u32 indices, material 1
    sv, st = scenes.icosphere(2, 0.35)
    sn = scenes.vertex_normals(sv, st)
    inter = np.concatenate([sv, sn], 1).astype(np.float32)
    iv = view(inter.tobytes(), stride=24)
    p_sphere = {"attributes": {"POSITION": accessor(iv, 5126, len(sv), "VEC3"), "NORMAL": accessor(iv, 5126, len(sv), "VEC3", offset=12)},
                "indices": accessor(view(st.tobytes()), 5125, st.size, "SCALAR"), "material": 1}

    png, pixels = checker_png()
    image_view = view(png)

    m_b = np.array([[0.8, 0, 0.2, 0.1], [0, 1.1, 0, 0.05], [-0.2, 0, 0.8, -0.1], [0, 0, 0, 1]], np.float32)
    t_a, axis_a, angle_a, s_a = [0.5, 0.45, -0.2], (0.2, 1.0, 0.1), 0.7, [1.2, 0.9, 1.1]
    nodes = [{"name": "A", "translation": t_a, "rotation": _quat(axis_a, angle_a), "scale": s_a, "children": [1]},
             {"name": "B", "matrix": m_b.T.reshape(-1).tolist(), "mesh": 1},
             {"name": "C", "mesh": 0, "children": [3]},
             {"name": "D", "translation": [-0.6, 0.5, 0.3], "mesh": 1}]
    doc = {"asset": {"version": "2.0"}, "scene": 0, "scenes": [{"nodes": [0, 2]}], "nodes": nodes,
           # the specification requires used extensions to be declared; assimp ignores undeclared ones (glTF2Asset.inl:1360, :1384)
           "extensionsUsed": ["KHR_materials_transmission", "KHR_materials_ior"],
           "meshes": [{"primitives": [p_floor, p_box]}, {"primitives": [p_sphere]}],
           "materials": [{"pbrMetallicRoughness": {"baseColorTexture": {"index": 0}, "metallicFactor": 0.0, "roughnessFactor": 0.7}},
                         {"pbrMetallicRoughness": {"baseColorFactor": [0.9, 0.3, 0.2, 1.0], "metallicFactor": 0.2, "roughnessFactor": 0.4},
                          "extensions": {"KHR_materials_transmission": {"transmissionFactor": 0.25}, "KHR_materials_ior": {"ior": 1.33}}}],
           "textures": [{"source": 0}], "images": [{"bufferView": image_view, "mimeType": "image/png"}],
           "accessors": accessors, "bufferViews": views, "buffers": [{"byteLength": len(blob)}]}

    ta = np.eye(4)
    ta[:3, 3] = t_a
    tf_a = ta @ _rot(axis_a, angle_a) @ np.diag(s_a + [1.0])
    td = np.eye(4)
    td[:3, 3] = [-0.6, 0.5, 0.3]
    bsm = bv.copy()  # box normals: per-position area-weighted average (every corner position is unique in this mesh)
    # Order of the importer's meshes and materials = the order in which assimp first MEETS them while it loads the default scene
    # (LazyDict::Retrieve; a node's children are read before the node's own mesh, glTF2Asset.inl:1692-1760), not the order of the
    # file: root A has no mesh, its child B retrieves glTF mesh 1 (sphere, material 1) first; root C's child D uses mesh 1 again,
    # then C's own mesh 0 (floor with material 0, box without material) follows. Pinned against the reference's own loader on
    # the reference's vendored assimp by tests/test_scene_loader.py::test_loader_equals_reference_scenebuilder_on_assimp.
    # Vertices: the importer numbers them in the order the index list first mentions them (vertexRemappingTable in ImportMeshes).
    def first_use(v, t, *attrs):
        flat = t.reshape(-1)
        order = flat[np.sort(np.unique(flat, return_index=True)[1])]
        remap = np.zeros(len(v), np.uint32)
        remap[order] = np.arange(len(order), dtype=np.uint32)
        return (v[order], remap[t]) + tuple(a[order] for a in attrs)

    # generated normals (the box has no NORMAL): assimp's GenSmoothNormals on the importer's indexed mesh gives every vertex the
    # unit normal of the LAST face that uses it, then averages over vertices at the same position (GenVertexNormalsProcess.cpp:
    # 125-185) -- every corner position of this box is one vertex, so no averaging happens
    bn = np.zeros((len(bv), 3), np.float32)
    for tri in bt:
        a, b, c = bv[tri[0]], bv[tri[1]], bv[tri[2]]
        fn_ = np.cross(b - a, c - a).astype(np.float32)
        bn[tri] = fn_ / np.float32(np.sqrt((fn_ * fn_).sum(dtype=np.float32)))
    fuv3 = np.stack([fuv[:, 0], 1.0 - fuv[:, 1], np.zeros(len(fuv))], -1).astype(np.float32)
    sv_o, st_o, sn_o = first_use(sv, st, sn)
    fv_o, ft_o, fn_o, fuv_o = first_use(fv, ft, fn, fuv3)
    bv_o, bt_o, bn_o = first_use(bv, bt, bn)
    expected = {
        "meshes": [(sv_o, st_o), (fv_o, ft_o), (bv_o, bt_o)],
        "normals": [sn_o, fn_o, bn_o],
        "uv": [np.zeros((len(sv_o), 3), np.float32), fuv_o, np.zeros((len(bv_o), 3), np.float32)],
        "mesh_material": [0, 1, 2],             # sphere -> glTF material 1 = row 0, floor -> glTF material 0 = row 1, box -> default (row 2)
        # instances: depth-first, a node's primitives before its children (sceneBuilder.cpp:285-315): A (no mesh) -> B; C -> floor, box; D
        "instances": [(0, (tf_a @ m_b.astype(np.float64)).astype(np.float32), 0), (1, np.eye(4, dtype=np.float32), 1),
                      (2, np.eye(4, dtype=np.float32), 2), (0, td.astype(np.float32), 0)],
        "pixels": pixels,
    }
    return doc, bytes(blob), expected


def write_glb(path, doc, binary):
    js = _pad4(json.dumps(doc).encode("utf-8"), b" ")
    bn = _pad4(binary)
    with open(path, "wb") as f:
        f.write(struct.pack("<4sII", b"glTF", 2, 12 + 8 + len(js) + 8 + len(bn)))
        f.write(struct.pack("<II", len(js), 0x4E4F534A) + js)
        f.write(struct.pack("<II", len(bn), 0x004E4942) + bn)


def write_gltf(path, doc, binary):
    """Text flavour: the same document with the buffer as a data URI."""
    d = json.loads(json.dumps(doc))
    d["buffers"][0]["uri"] = "data:application/octet-stream;base64," + base64.b64encode(binary).decode("ascii")
    with open(path, "w") as f:
        json.dump(d, f)
