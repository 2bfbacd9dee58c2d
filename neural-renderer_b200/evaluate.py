"""Headless evaluate driver: the caller on one side of the hot path (reference src/evaluate.cu:118-310).

config JSON (the reference's schema, src/config_loader.cpp:10-144) -> scene -> ground-truth render and
neural render through RendererNeural, total_samples accumulated in batches of 8 spp -> PNGs, PSNR, FLIP.

  python tools/evaluate.py CONFIG.json [OUT_DIR] [--width W --height H]

Asset formats: the reference loads meshes through assimp/tinygltf and environments through stb/tinyexr
(out of scope).  Here: Wavefront OBJ, ASCII / binary-little-endian PLY and glTF 2.0 (.glb / .gltf, a
restatement of the reference's LoadGltfWithMaterials: node transforms, per-primitive materials, embedded images) meshes,
Radiance .hdr (RLE decoder below), OpenEXR (through OpenCV) and 8-bit .png / .jpg environments, raw fp16 .bin checkpoints, plus `synthetic:` pseudo-paths so the
driver runs without any asset on disk:
    "path": "synthetic:ico:5"         displaced icosphere, 5 subdivisions       (original mesh)
    "path": "synthetic:shell:3:1.12"  icosphere(3) scaled by 1.12               (shells)
    "path": "synthetic:ground"        the ground quad                            (additional mesh)
    "hdri_path": "synthetic:sky"      the seeded sun + gradient environment
    "checkpoint_path": "synthetic:42" seeded random weights of the configured network
"""
import argparse
import json
import os
import struct
import sys
import time

import numpy as np

from . import _capi as K
from . import metrics, scenes
from .renderer import RendererNeural, Scene, compute_camera_basis

BATCH_SPP = 8  # kBatchSizeGT / kBatchSizeNeural, src/evaluate.cu:22-23


# ---- config ---------------------------------------------------------------------------------------
def load_config(path):
    """LoadConfigFromFile with the reference's defaults; relative paths resolve against the config."""
    j = json.load(open(path))
    base = os.path.dirname(os.path.abspath(path))

    def resolve(p):
        if not p or p.startswith("synthetic:") or os.path.isabs(p):
            return p
        return os.path.normpath(os.path.join(base, p))

    def mesh(key):
        m = j.get("scene", {}).get(key, {})
        return {"path": resolve(m.get("path", "")), "scale": float(m.get("scale", 1.0))}

    cam, ren, mat, nn, env = (j.get(k, {}) for k in ("camera", "rendering", "material", "neural_network", "environment"))
    matrix = cam.get("matrix")
    if not (isinstance(matrix, list) and len(matrix) == 16):
        matrix = [1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1]
    return {
        "original_mesh": mesh("original_mesh"), "inner_shell": mesh("inner_shell"), "outer_shell": mesh("outer_shell"),
        "additional_mesh": mesh("additional_mesh"),
        "checkpoint_path": resolve(j.get("checkpoint_path", "")),
        "environment": {"hdri_path": resolve(env.get("hdri_path", "")), "rotation": float(env.get("rotation", 0.0)),
                        "strength": float(env.get("strength", 1.0))},
        "camera": {"matrix": [float(x) for x in matrix], "yfov": float(cam.get("yfov", 1.047198))},
        "rendering": {"total_samples": int(ren.get("total_samples", 2048)), "bounce_count": int(ren.get("bounce_count", 3)),
                      "width": int(ren.get("width", 1920)), "height": int(ren.get("height", 1080))},
        "material": {"base_color": tuple(mat.get("base_color", (1.0, 1.0, 1.0))),
                     **{k: float(mat.get(k, d)) for k, d in dict(roughness=1.0, metallic=0.0, specular=0.0, specular_tint=0.0,
                                                                 anisotropy=0.0, sheen=0.0, sheen_tint=0.0, clearcoat=0.0,
                                                                 clearcoat_gloss=0.0).items()},
                     "use_constant_neural_color": bool(mat.get("use_constant_neural_color", False)),
                     "constant_neural_color": tuple(mat.get("constant_neural_color", (0.8, 0.8, 0.8)))},
        "neural_network": {"log2_hashmap_size": int(nn.get("log2_hashmap_size", 14)),
                           "base_resolution": int(nn.get("base_resolution", 16)),
                           "use_neural_query": bool(nn.get("use_neural_query", False))},
    }


def matrix_to_camera_state(m):
    """MatrixToCameraState, src/config_loader.cpp:146-185: column-major world-to-camera matrix ->
    position, yaw, pitch (degrees)."""
    m = np.asarray(m, np.float32)
    t = m[12:15]
    r = np.array([[m[0], m[4], m[8]], [m[1], m[5], m[9]], [m[2], m[6], m[10]]], np.float32)
    pos = -(r.T @ t)
    fwd = -r[2]
    yaw = np.degrees(np.arctan2(fwd[2], fwd[0]))
    pitch = np.degrees(np.arctan2(fwd[1], np.sqrt(fwd[0] * fwd[0] + fwd[2] * fwd[2])))
    return pos.astype(np.float32), float(yaw), float(pitch)


def camera_state_to_matrix(pos, target):
    """Inverse helper for writing configs: a look-at world-to-camera matrix, column-major."""
    pos = np.asarray(pos, np.float64)
    f = np.asarray(target, np.float64) - pos
    f /= np.linalg.norm(f)
    r = np.cross(f, [0.0, 1.0, 0.0])
    r /= np.linalg.norm(r)
    u = np.cross(r, f)
    rot = np.stack([r, u, -f])  # rows of the view rotation
    t = -rot @ pos
    m = np.eye(4)
    m[:3, :3] = rot
    m[:3, 3] = t
    return [float(x) for x in m.T.reshape(-1)]


# ---- assets ---------------------------------------------------------------------------------------
def load_obj(path):
    vs, faces = [], []
    for line in open(path, errors="ignore"):
        p = line.split()
        if not p:
            continue
        if p[0] == "v":
            vs.append([float(x) for x in p[1:4]])
        elif p[0] == "f":
            idx = [int(tok.split("/")[0]) for tok in p[1:]]
            idx = [i - 1 if i > 0 else len(vs) + i for i in idx]
            for k in range(1, len(idx) - 1):  # fan triangulation
                faces.append([idx[0], idx[k], idx[k + 1]])
    return np.asarray(vs, np.float32), np.asarray(faces, np.uint32)


def flat_shaded(vertices, faces):
    """What assimp's GenNormals + JoinIdenticalVertices leave of a mesh WITHOUT normals (LoadMeshFromFile,
    src/mesh_loader.cpp:604-690): every face corner gets the face's normal normalize(cross(v1 - v0, v2 - v0)), then
    corners with the same position and normal become one vertex.  -> (vertices, faces, normals)."""
    v = np.asarray(vertices, np.float32)
    f = np.asarray(faces, np.int64)
    p = v[f]                                                     # [T, 3, 3]
    n = np.cross(p[:, 1] - p[:, 0], p[:, 2] - p[:, 0]).astype(np.float32)
    ln = np.linalg.norm(n, axis=1, keepdims=True)
    n = np.where(ln > 0, n / np.where(ln > 0, ln, 1), np.array([0, 1, 0], np.float32)).astype(np.float32)
    corners = np.concatenate([p.reshape(-1, 3), np.repeat(n, 3, axis=0)], 1)  # [3T, 6]
    corners = np.ascontiguousarray(corners + np.float32(0.0))    # -0 -> +0: equal values, equal bits
    keys = corners.view(np.uint64)                               # [3T, 3]: rows compare as three 64-bit words
    order = np.lexsort((keys[:, 2], keys[:, 1], keys[:, 0]))
    sorted_keys = keys[order]
    first = np.ones(len(order), bool)
    first[1:] = np.any(sorted_keys[1:] != sorted_keys[:-1], axis=1)
    inverse = np.empty(len(order), np.int64)
    inverse[order] = np.cumsum(first) - 1
    uniq = corners[order[first]]
    return uniq[:, :3].copy(), inverse.reshape(-1, 3).astype(np.uint32), uniq[:, 3:].copy()


def load_obj_shaded(path):
    """Wavefront OBJ as the reference's assimp path delivers it: positions and, where the file has `vn` records, the
    corners' normals (normalised; one vertex per distinct position / normal couple); without normals, flat_shaded.
    Texture coordinates and materials are not loaded on this path (`No textures, no materials`, :597).
    -> (vertices, faces, normals)."""
    vs, vns, corners, faces = [], [], {}, []
    out_v, out_n = [], []
    has_n = True
    for line in open(path, errors="ignore"):
        t = line.split()
        if not t:
            continue
        if t[0] == "v":
            vs.append([float(x) for x in t[1:4]])
        elif t[0] == "vn":
            vns.append([float(x) for x in t[1:4]])
        elif t[0] == "f":
            ids = []
            for tok in t[1:]:
                parts = tok.split("/")
                vi = int(parts[0])
                vi = vi - 1 if vi > 0 else len(vs) + vi
                ni = None
                if len(parts) > 2 and parts[2]:
                    ni = int(parts[2])
                    ni = ni - 1 if ni > 0 else len(vns) + ni
                else:
                    has_n = False
                key = (vi, ni)
                if key not in corners:
                    corners[key] = len(out_v)
                    out_v.append(vi)
                    out_n.append(ni)
                ids.append(corners[key])
            for k in range(1, len(ids) - 1):
                faces.append([ids[0], ids[k], ids[k + 1]])
    V = np.asarray(vs, np.float32)
    F = np.asarray(faces, np.uint32)
    if not has_n or not vns:
        v_only, f_only = load_obj(path)
        return flat_shaded(v_only, f_only)
    N = np.asarray(vns, np.float32)[np.asarray(out_n)]
    ln = np.linalg.norm(N, axis=1, keepdims=True)
    N = np.where(ln > 0, N / np.where(ln > 0, ln, 1), np.array([0, 1, 0], np.float32)).astype(np.float32)
    return V[np.asarray(out_v)], F, N


_PLY_TYPES = {"char": "i1", "uchar": "u1", "short": "i2", "ushort": "u2", "int": "i4", "uint": "u4", "float": "f4", "double": "f8",
              "int8": "i1", "uint8": "u1", "int16": "i2", "uint16": "u2", "int32": "i4", "uint32": "u4", "float32": "f4",
              "float64": "f8"}


def _ply_fan(polys):
    """[n, k] polygons -> [(k - 2) n, 3] triangle fans, face by face (assimp's Triangulate on convex input)."""
    polys = np.asarray(polys)
    k = polys.shape[1]
    if k == 3:
        return polys
    return np.stack([np.stack([polys[:, 0], polys[:, m], polys[:, m + 1]], 1) for m in range(1, k - 1)], 1).reshape(-1, 3)


def _ply_element_binary(f, count, props, endian):
    """One element of a binary PLY -> {property: array} (list properties: [count, k] when every row has k entries, else a
    list of arrays).  Whole-element numpy reads where the rows have one size; a row loop otherwise."""
    if count == 0:
        return {pr[-1]: np.zeros((0,)) for pr in props}
    lists = [pr for pr in props if pr[0] == "list"]
    start = f.tell()
    if lists:
        # probe the first row for the list lengths, then try one structured read with those lengths
        lens = {}
        for pr in props:
            if pr[0] == "list":
                n = int(np.frombuffer(f.read(np.dtype(_PLY_TYPES[pr[1]]).itemsize), endian + _PLY_TYPES[pr[1]])[0])
                lens[pr[-1]] = n
                f.seek(n * np.dtype(_PLY_TYPES[pr[2]]).itemsize, 1)
            else:
                f.seek(np.dtype(_PLY_TYPES[pr[0]]).itemsize, 1)
        f.seek(start)
    fields = []
    for pr in props:
        if pr[0] == "list":
            fields.append(("#" + pr[-1], endian + _PLY_TYPES[pr[1]]))
            fields.append((pr[-1], endian + _PLY_TYPES[pr[2]], (lens[pr[-1]],)))
        else:
            fields.append((pr[-1], endian + _PLY_TYPES[pr[0]]))
    dt = np.dtype(fields)
    raw = f.read(dt.itemsize * count)
    if len(raw) == dt.itemsize * count:
        rows = np.frombuffer(raw, dt)
        if all(np.all(rows["#" + pr[-1]] == lens[pr[-1]]) for pr in lists):
            return {pr[-1]: rows[pr[-1]] for pr in props}
    # ragged lists: row by row
    f.seek(start)
    out = {pr[-1]: [] for pr in props}
    for _ in range(count):
        for pr in props:
            if pr[0] == "list":
                n = int(np.frombuffer(f.read(np.dtype(_PLY_TYPES[pr[1]]).itemsize), endian + _PLY_TYPES[pr[1]])[0])
                out[pr[-1]].append(np.frombuffer(f.read(n * np.dtype(_PLY_TYPES[pr[2]]).itemsize), endian + _PLY_TYPES[pr[2]]))
            else:
                out[pr[-1]].append(np.frombuffer(f.read(np.dtype(_PLY_TYPES[pr[0]]).itemsize), endian + _PLY_TYPES[pr[0]])[0])
    return {k: (v if any(pr[0] == "list" and pr[-1] == k for pr in props) else np.asarray(v)) for k, v in out.items()}


def _ply_element_ascii(f, count, props):
    out = {pr[-1]: [] for pr in props}
    for _ in range(count):
        tok = f.readline().split()
        while not tok:
            tok = f.readline().split()
        at = 0
        for pr in props:
            if pr[0] == "list":
                n = int(tok[at])
                out[pr[-1]].append(np.asarray(tok[at + 1:at + 1 + n], np.float64).astype(_PLY_TYPES[pr[2]]))
                at += 1 + n
            else:
                out[pr[-1]].append(float(tok[at]))
                at += 1
    res = {}
    for pr in props:
        v = out[pr[-1]]
        if pr[0] == "list":
            res[pr[-1]] = np.stack(v) if v and all(len(x) == len(v[0]) for x in v) else v
        else:
            res[pr[-1]] = np.asarray(v, np.float64).astype(_PLY_TYPES[pr[0]])
    return res


def load_ply_full(path):
    """Stanford PLY (ascii, binary_little_endian, binary_big_endian) -> (vertices [V,3] f32, faces [T,3] u32 - polygons as
    fans -, normals [V,3] f32 or None when the vertex element has no nx / ny / nz).  Other elements and properties are
    skipped by their declared sizes."""
    f = open(path, "rb")
    if f.readline().strip() != b"ply":
        raise ValueError("not a PLY file")
    fmt, elems = None, []
    while True:
        line = f.readline()
        if not line:
            raise ValueError("PLY header without end_header")
        p = line.split()
        if not p:
            continue
        if p[0] == b"format":
            fmt = p[1].decode()
        elif p[0] == b"element":
            elems.append([p[1].decode(), int(p[2]), []])
        elif p[0] == b"property":
            elems[-1][2].append([x.decode() for x in p[1:]])
        elif p[0] == b"end_header":
            break
    if fmt not in ("ascii", "binary_little_endian", "binary_big_endian"):
        raise ValueError("unsupported PLY format " + str(fmt))
    verts = np.zeros((0, 3), np.float32)
    normals, tris = None, []
    for name, count, props in elems:
        if fmt == "ascii":
            data = _ply_element_ascii(f, count, props)
        else:
            data = _ply_element_binary(f, count, props, "<" if fmt == "binary_little_endian" else ">")
        if name == "vertex":
            verts = np.stack([np.asarray(data[c], np.float32) for c in "xyz"], 1) if count else verts
            if count and all(c in data for c in ("nx", "ny", "nz")):
                normals = np.stack([np.asarray(data[c], np.float32) for c in ("nx", "ny", "nz")], 1)
        elif name == "face" and count:
            key = next((k for k in ("vertex_indices", "vertex_index") if k in data), None)
            if key is None:
                key = next(pr[-1] for pr in props if pr[0] == "list")
            polys = data[key]
            if isinstance(polys, np.ndarray) and polys.ndim == 2:
                if polys.shape[1] >= 3:
                    tris.append(_ply_fan(polys))
            else:
                for poly in polys:
                    if len(poly) >= 3:
                        tris.append(_ply_fan(np.asarray(poly)[None, :]))
    faces = np.concatenate(tris).astype(np.uint32) if tris else np.zeros((0, 3), np.uint32)
    return verts, faces, normals


def load_ply(path):
    v, f, _ = load_ply_full(path)
    return v, f


def load_hdr(path):
    """Radiance RGBE (.hdr) with new-style RLE scanlines -> [H, W, 3] float32."""
    f = open(path, "rb")
    if not f.readline().startswith(b"#?"):
        raise ValueError("not a Radiance file")
    while f.readline().strip():
        pass
    dims = f.readline().split()
    h, w = int(dims[1]), int(dims[3])
    out = np.zeros((h, w, 4), np.uint8)
    for y in range(h):
        head = f.read(4)
        if w < 8 or w > 32767 or head[0] != 2 or head[1] != 2 or (head[2] << 8 | head[3]) != w:
            rest = np.frombuffer(head + f.read(4 * w - 4), np.uint8).reshape(w, 4)  # flat scanline
            out[y] = rest
            continue
        for c in range(4):
            x = 0
            while x < w:
                n = f.read(1)[0]
                if n > 128:
                    out[y, x:x + n - 128, c] = f.read(1)[0]
                    x += n - 128
                else:
                    out[y, x:x + n, c] = np.frombuffer(f.read(n), np.uint8)
                    x += n
    e = out[..., 3].astype(np.int32)
    scale = np.where(e > 0, np.ldexp(1.0, e - 136), 0.0).astype(np.float32)
    return out[..., :3].astype(np.float32) * scale[..., None]


# ---- glTF 2.0 (.glb / .gltf): LoadGltfWithMaterials, reference src/mesh_loader.cpp:160-624 ------------------------
_GLTF_COMPONENT = {5120: np.int8, 5121: np.uint8, 5122: np.int16, 5123: np.uint16, 5125: np.uint32, 5126: np.float32}
_GLTF_WIDTH = {"SCALAR": 1, "VEC2": 2, "VEC3": 3, "VEC4": 4, "MAT4": 16}


def _gltf_read(path):
    """-> (json dict, [buffer bytes]).  .glb: 12-byte header, then JSON and BIN chunks; .gltf: JSON with buffer URIs
    (files beside it or base64 data URIs)."""
    import base64
    import struct
    raw = open(path, "rb").read()
    base = os.path.dirname(os.path.abspath(path))
    bin_chunk = None
    if raw[:4] == b"glTF":
        _, _, total = struct.unpack_from("<4sII", raw, 0)
        off, doc = 12, None
        while off + 8 <= min(total, len(raw)):
            n, kind = struct.unpack_from("<II", raw, off)
            body = raw[off + 8: off + 8 + n]
            if kind == 0x4E4F534A:
                doc = json.loads(body.decode("utf-8"))
            elif kind == 0x004E4942 and bin_chunk is None:
                bin_chunk = body
            off += 8 + n  # chunk lengths include their padding to 4 bytes
        if doc is None:
            raise ValueError(f"{path}: no JSON chunk")
    else:
        doc = json.loads(raw.decode("utf-8"))
    buffers = []
    for b in doc.get("buffers", []):
        uri = b.get("uri")
        if uri is None:
            buffers.append(bin_chunk or b"")
        elif uri.startswith("data:"):
            buffers.append(base64.b64decode(uri.split(",", 1)[1]))
        else:
            buffers.append(open(os.path.join(base, uri), "rb").read())
    return doc, buffers, base


def _gltf_accessor(doc, buffers, index):
    """-> [count, width] array of the accessor's component type (byteStride honoured; sparse accessors not read)."""
    acc = doc["accessors"][index]
    view = doc["bufferViews"][acc["bufferView"]]
    dt = np.dtype(_GLTF_COMPONENT[acc["componentType"]]).newbyteorder("<")
    width = _GLTF_WIDTH[acc["type"]]
    start = view.get("byteOffset", 0) + acc.get("byteOffset", 0)
    stride = view.get("byteStride", 0) or dt.itemsize * width
    buf = np.frombuffer(buffers[view["buffer"]], np.uint8)
    count = acc["count"]
    rows = np.lib.stride_tricks.as_strided(buf[start:], shape=(count, dt.itemsize * width), strides=(stride, 1))
    return np.ascontiguousarray(rows).view(dt).reshape(count, width)


def load_gltf(path):
    """-> dict(vertices, normals, texcoords, indices, materials, material_map, material_ids, textures): one mesh out of
    every TRIANGLES primitive of every scene's node tree, positions and normals in world space (node matrices or TRS,
    column-major, fp32 like the reference), one material-map entry per primitive, images as 8-bit RGBA (tinygltf's
    default), materials from pbrMetallicRoughness: base colour factor / texture, metallic and roughness FACTORS (the
    reference maps the packed texture and then switches it off again, :263-301), normal and emissive textures,
    KHR_materials_transmission / _ior / _emissive_strength.  Non-indexed primitives contribute no triangles (:526)."""
    import io
    doc, buffers, base = _gltf_read(path)
    f32 = np.float32
    textures = []
    for img in doc.get("images", []):
        pixels = None
        try:
            from PIL import Image
            if "bufferView" in img:
                v = doc["bufferViews"][img["bufferView"]]
                o = v.get("byteOffset", 0)
                data = buffers[v["buffer"]][o: o + v["byteLength"]]
                pixels = np.asarray(Image.open(io.BytesIO(data)).convert("RGBA"), np.uint8)
            elif img.get("uri", "").startswith("data:"):
                import base64
                pixels = np.asarray(Image.open(io.BytesIO(base64.b64decode(img["uri"].split(",", 1)[1]))).convert("RGBA"), np.uint8)
            elif img.get("uri"):
                pixels = np.asarray(Image.open(os.path.join(base, img["uri"])).convert("RGBA"), np.uint8)
        except Exception as e:  # noqa: BLE001 - an unreadable image stays an empty texture, as in the reference
            print(f"[GLTF] image not loaded: {e}", file=sys.stderr)
        textures.append([pixels if pixels is not None else np.zeros((0, 0, 4), np.uint8), False])

    def source(tex_ref):
        if not tex_ref or tex_ref.get("index", -1) < 0:
            return -1
        src = doc["textures"][tex_ref["index"]].get("source", -1)
        return src if 0 <= src < len(textures) else -1

    materials = []
    for gm in doc.get("materials", []):
        m = K.Material.default()
        pbr = gm.get("pbrMetallicRoughness", {})
        m.base_color.value[:] = [f32(x) for x in pbr.get("baseColorFactor", [1, 1, 1, 1])[:3]]
        t = source(pbr.get("baseColorTexture"))
        if t >= 0:
            m.base_color.textured, m.base_color.texture_id = 1, t
            textures[t][1] = True
        m.metallic.value = f32(pbr.get("metallicFactor", 1.0))
        m.roughness.value = f32(pbr.get("roughnessFactor", 1.0))
        t = source(pbr.get("metallicRoughnessTexture"))
        if t >= 0:
            textures[t][1] = False
        t = source(gm.get("normalTexture"))
        if t >= 0:
            m.normal.textured, m.normal.texture_id = 1, t
            textures[t][1] = False
        m.base_emission.value[:] = [f32(x) for x in gm.get("emissiveFactor", [0, 0, 0])[:3]]
        t = source(gm.get("emissiveTexture"))
        if t >= 0:
            m.base_emission.textured, m.base_emission.texture_id = 1, t
            textures[t][1] = True
        ext = gm.get("extensions", {})
        if "transmissionFactor" in ext.get("KHR_materials_transmission", {}):
            m.specular_transmission = f32(ext["KHR_materials_transmission"]["transmissionFactor"])
        if "ior" in ext.get("KHR_materials_ior", {}):
            m.ior = f32(ext["KHR_materials_ior"]["ior"])
        if "emissiveStrength" in ext.get("KHR_materials_emissive_strength", {}):
            m.emission_scale = f32(ext["KHR_materials_emissive_strength"]["emissiveStrength"])
        materials.append(m)
    if not materials:
        materials.append(K.Material.default())

    def mat_mul(a, b):  # both column-major 4x4 as [16] fp32, fp32 accumulation in the reference's order
        out = np.zeros(16, f32)
        for col in range(4):
            for row in range(4):
                acc = f32(0.0)
                for k in range(4):
                    acc = f32(acc + f32(a[k * 4 + row] * b[col * 4 + k]))
                out[col * 4 + row] = acc
        return out

    ident = np.array([1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1], f32)
    V, N, UV, IDX, mmap, mids = [], [], [], [], [], []
    nv = [0]
    ntri = [0]

    def visit(ni, parent):
        node = doc["nodes"][ni]
        if node.get("matrix"):
            local = np.asarray(node["matrix"], f32)
        else:
            T, R, S = ident.copy(), ident.copy(), ident.copy()
            if node.get("translation"):
                T[12:15] = np.asarray(node["translation"], f32)
            if node.get("rotation"):
                qx, qy, qz, qw = (f32(x) for x in node["rotation"])
                one, two = f32(1), f32(2)
                R[0] = one - two * (qy * qy + qz * qz); R[1] = two * (qx * qy + qz * qw); R[2] = two * (qx * qz - qy * qw)
                R[4] = two * (qx * qy - qz * qw); R[5] = one - two * (qx * qx + qz * qz); R[6] = two * (qy * qz + qx * qw)
                R[8] = two * (qx * qz + qy * qw); R[9] = two * (qy * qz - qx * qw); R[10] = one - two * (qx * qx + qy * qy)
            if node.get("scale"):
                S[0], S[5], S[10] = (f32(x) for x in node["scale"])
            local = mat_mul(T, mat_mul(R, S))
        world = mat_mul(parent, local)
        if node.get("mesh", -1) >= 0:
            for prim in doc["meshes"][node["mesh"]].get("primitives", []):
                if prim.get("mode", 4) != 4 or "POSITION" not in prim.get("attributes", {}):
                    continue
                att = prim["attributes"]
                pos = _gltf_accessor(doc, buffers, att["POSITION"]).astype(f32)[:, :3]
                M = world.reshape(4, 4).T  # row-major view of the column-major matrix
                # transformPoint: m0*x + m4*y + m8*z + m12, summed left to right in fp32
                p = ((pos[:, 0:1] * M[:3, 0][None] + pos[:, 1:2] * M[:3, 1][None]) + pos[:, 2:3] * M[:3, 2][None]) + M[:3, 3][None]
                base_vertex = nv[0]
                V.append(p.astype(f32))
                if "NORMAL" in att:
                    nr = _gltf_accessor(doc, buffers, att["NORMAL"]).astype(f32)[:, :3]
                    q = (nr[:, 0:1] * M[:3, 0][None] + nr[:, 1:2] * M[:3, 1][None]) + nr[:, 2:3] * M[:3, 2][None]
                    ln = np.sqrt((q[:, 0] * q[:, 0] + q[:, 1] * q[:, 1]) + q[:, 2] * q[:, 2]).astype(f32)
                    ok = ln > 0
                    q = np.where(ok[:, None], q / np.where(ok, ln, 1)[:, None], np.array([0, 1, 0], f32)[None])
                    N.append(q.astype(f32))
                else:
                    N.append(np.tile(np.array([0, 1, 0], f32), (len(pos), 1)))
                if "TEXCOORD_0" in att:
                    UV.append(_gltf_accessor(doc, buffers, att["TEXCOORD_0"]).astype(f32)[:, :2])
                else:
                    UV.append(np.zeros((len(pos), 2), f32))
                nv[0] += len(pos)
                mmap.append(ntri[0])
                if prim.get("indices", -1) >= 0:
                    ii = _gltf_accessor(doc, buffers, prim["indices"]).reshape(-1).astype(np.uint32)
                    ii = ii[: len(ii) // 3 * 3].reshape(-1, 3) + np.uint32(base_vertex)
                    IDX.append(ii)
                    ntri[0] += len(ii)
                mids.append(prim.get("material", 0) if prim.get("material", -1) >= 0 else 0)
        for c in node.get("children", []):
            visit(c, world)

    for sc in doc.get("scenes", []):
        for ni in sc.get("nodes", []):
            visit(ni, ident)
    if not IDX or ntri[0] == 0:
        raise ValueError("No triangles found in GLTF file.")
    return dict(vertices=np.concatenate(V), normals=np.concatenate(N), texcoords=np.concatenate(UV),
                indices=np.concatenate(IDX), materials=materials, material_map=np.asarray(mmap, np.uint32),
                material_ids=np.asarray(mids, np.int32), textures=[(px, srgb) for px, srgb in textures])


def load_env(path):
    """EnvironmentMap::loadFromFile (reference src/scene.cu:289-316): .exr -> RGB of the float image (tinyexr LoadEXR
    there, OpenCV's OpenEXR reader here), .png / .jpg / .jpeg -> 8-bit RGB / 255 (no gamma, :245-280), anything else ->
    Radiance .hdr.  -> [H, W, 3] fp32."""
    low = path.lower()
    if low.endswith(".exr"):
        os.environ.setdefault("OPENCV_IO_ENABLE_OPENEXR", "1")  # must be set before cv2 is first imported
        import cv2
        img = cv2.imread(path, cv2.IMREAD_UNCHANGED)
        if img is None:
            raise ValueError(f"{path}: OpenCV could not read the EXR file (OPENCV_IO_ENABLE_OPENEXR must be 1 before cv2 is imported)")
        img = np.asarray(img, np.float32)
        if img.ndim == 2:
            img = np.repeat(img[..., None], 3, -1)
        return np.ascontiguousarray(img[..., 2::-1] if img.shape[-1] >= 3 else np.repeat(img[..., :1], 3, -1))  # BGR(A) -> RGB
    if low.endswith((".png", ".jpg", ".jpeg")):
        from PIL import Image
        return np.asarray(Image.open(path).convert("RGB"), np.float32) / np.float32(255.0)
    return load_hdr(path)


def load_mesh(spec, what):
    """-> MeshData or None.  `spec` = {"path", "scale"}; vertices are multiplied by scale (LoadMeshLabeled)."""
    path, scale = spec["path"], np.float32(spec["scale"])
    if not path:
        return None
    if path.startswith("synthetic:"):
        tok = path.split(":")
        if tok[1] == "ico":
            v0, idx = scenes.icosphere(int(tok[2]) if len(tok) > 2 else 5)
            v = scenes.displace(v0, 0.1)
            return scenes.mesh_from((v * scale).astype(np.float32), idx, normals=scenes.vertex_normals(v, idx))
        if tok[1] == "shell":
            v, idx = scenes.icosphere(int(tok[2]) if len(tok) > 2 else 3)
            r = np.float32(float(tok[3]) if len(tok) > 3 else 1.0)
            return scenes.mesh_from((v * r * scale).astype(np.float32), idx)
        if tok[1] == "ground":
            v, idx = scenes.ground_quad()
            return scenes.mesh_from((v * scale).astype(np.float32), idx)
        raise ValueError(f"unknown synthetic {what}: {path}")
    ext = os.path.splitext(path)[1].lower()
    if ext == ".obj":
        v, idx, nrm = load_obj_shaded(path)
        return scenes.mesh_from((v * scale).astype(np.float32), idx, normals=nrm)
    elif ext == ".ply":
        v, idx, nrm = load_ply_full(path)
        if nrm is None:
            v, idx, nrm = flat_shaded(v, idx)  # no nx / ny / nz: assimp's GenNormals
        else:
            ln = np.linalg.norm(nrm, axis=1, keepdims=True)
            nrm = np.where(ln > 0, nrm / np.where(ln > 0, ln, 1), np.array([0, 1, 0], np.float32)).astype(np.float32)
        return scenes.mesh_from((v * scale).astype(np.float32), idx, normals=nrm)
    elif ext in (".glb", ".gltf"):
        g = load_gltf(path)
        v = (g["vertices"] * scale).astype(np.float32)  # scaleMesh: positions only
        # LoadMeshLabeled: nearest-neighbour texture filtering on the original and the additional mesh (evaluate.cu:153-176)
        nearest = what in ("original mesh", "additional mesh")
        return scenes.mesh_from(v, g["indices"], normals=g["normals"], texcoords=g["texcoords"], materials=g["materials"],
                                material_map=g["material_map"], material_ids=g["material_ids"], textures=g["textures"],
                                has_mesh_materials=True, texture_nearest=nearest)
    else:
        raise ValueError(f"{what}: .obj / .ply / .glb / .gltf meshes are read here; other formats need assimp ({path})")
    v = (v * scale).astype(np.float32)
    return scenes.mesh_from(v, idx, normals=scenes.vertex_normals(v, idx))


def build_scene(cfg):
    m = cfg["material"]
    scene = Scene(original_mesh=load_mesh(cfg["original_mesh"], "original mesh"))
    scene.inner_shell = load_mesh(cfg["inner_shell"], "inner shell")
    scene.outer_shell = load_mesh(cfg["outer_shell"], "outer shell")
    scene.additional_mesh = load_mesh(cfg["additional_mesh"], "additional mesh")
    scene.global_material = K.Material.from_config(**{k: m[k] for k in (
        "base_color", "roughness", "metallic", "specular", "specular_tint", "anisotropy", "sheen", "sheen_tint", "clearcoat",
        "clearcoat_gloss")})
    # applyMaterialParamsOnly (evaluate.cu:194-209): the config's scalar parameters replace those of every mesh material;
    # base colour, normal and emission (and their textures) stay the file's
    for mesh in (scene.original_mesh, scene.inner_shell, scene.outer_shell, scene.additional_mesh):
        for mat in (mesh.materials if mesh is not None else []):
            for k in ("roughness", "metallic", "specular", "specular_tint", "anisotropy", "sheen", "sheen_tint", "clearcoat",
                      "clearcoat_gloss"):
                p = getattr(mat, k)
                p.value, p.textured, p.texture_id, p.channel = m[k], 0, 0, 0
    env = cfg["environment"]
    if env["hdri_path"].startswith("synthetic:"):
        scene.env_pixels = scenes.sky_env(1024, 512)
    elif env["hdri_path"]:
        try:
            scene.env_pixels = load_env(env["hdri_path"])
        except Exception as e:  # noqa: BLE001 - evaluate prints and goes on without an environment (evaluate.cu:213-216)
            print(f"Failed to load HDRI '{env['hdri_path']}': {e}", file=sys.stderr)
    scene.env_strength = env["strength"]
    return scene


def load_checkpoint(cfg, renderer):
    p = cfg["checkpoint_path"]
    if p.startswith("synthetic:"):
        nn = cfg["neural_network"]
        seed = int(p.split(":")[1]) if len(p.split(":")) > 1 and p.split(":")[1] else 42
        renderer.load_weights(scenes.make_weights(seed=seed, log2_hashmap_size=nn["log2_hashmap_size"],
                                                  base_resolution=nn["base_resolution"]))
        return True
    return renderer.load_weights_from_file(p)


def save_png(path, rgba):
    from PIL import Image
    Image.fromarray(np.ascontiguousarray(rgba[..., :3])).save(path)


def magma(t):
    """Coarse Magma ramp for the FLIP visualisation (the reference maps the error through FLIP::MapMagma)."""
    stops = np.array([[0, 0, 4], [28, 16, 68], [79, 18, 123], [129, 37, 129], [181, 54, 122], [229, 80, 100], [251, 135, 97],
                      [254, 194, 135], [252, 253, 191]], np.float64)
    x = np.clip(t, 0.0, 1.0) * (len(stops) - 1)
    i = np.minimum(x.astype(int), len(stops) - 2)
    f = (x - i)[..., None]
    return (stops[i] * (1 - f) + stops[i + 1] * f).astype(np.uint8)


def run(cfg, out_dir=None, width=None, height=None, lib=None, quiet=False):
    """The body of the reference's main(): returns {"psnr", "flip", "gt", "neural", seconds...}."""
    say = (lambda *a: None) if quiet else print
    W = width or 1920   # kWidth / kHeight are compile-time constants in the reference (src/evaluate.cu:20-21)
    H = height or 1080
    total, bounces = cfg["rendering"]["total_samples"], cfg["rendering"]["bounce_count"]
    pos, yaw, pitch = matrix_to_camera_state(cfg["camera"]["matrix"])
    pos = pos * np.float32(cfg["original_mesh"]["scale"])
    say(f"Loaded camera: pos=({pos[0]:.2f}, {pos[1]:.2f}, {pos[2]:.2f}), yaw={yaw:.2f}, pitch={pitch:.2f}, fovY={cfg['camera']['yfov']:.2f}")
    scene = build_scene(cfg)
    nn = cfg["neural_network"]
    r = RendererNeural(scene, nn_config=K.NNConfig(log2_hashmap_size=nn["log2_hashmap_size"], base_resolution=nn["base_resolution"]),
                       lib=lib)
    m = cfg["material"]
    r.set_constant_neural_color(m["use_constant_neural_color"], m["constant_neural_color"])
    r.resize(W, H)
    r.set_bounce_count(bounces)
    r.set_lambert_view(False)
    r.set_envmap_rotation(cfg["environment"]["rotation"])
    if not load_checkpoint(cfg, r):
        raise RuntimeError(f"Failed to load checkpoint: {cfg['checkpoint_path']}")
    r.set_camera_basis(compute_camera_basis(yaw, pitch, cfg["camera"]["yfov"]))
    frames, secs = {}, {}
    for name, neural in (("ground_truth", False), ("neural", True)):
        say(f"\n=== Rendering {name} ({total} samples) ===")
        r.set_use_neural_query(neural)
        r.set_classic_mesh_index(0)
        r.reset_samples()
        t0, left = time.perf_counter(), total
        while left > 0:
            r.set_samples_per_pixel(min(left, BATCH_SPP))
            r.render(pos)
            left -= BATCH_SPP
        frames[name] = r.read_pixels()
        secs[name] = time.perf_counter() - t0
    r.close()
    psnr = metrics.psnr8(frames["ground_truth"], frames["neural"])
    fmap = metrics.flip_ldr_map(frames["ground_truth"], frames["neural"])
    say("\n=== Metrics ===")
    say(f"PSNR: {psnr:.2f} dB")
    say(f"FLIP max error: {fmap.max():.4f}")
    say(f"FLIP: {fmap.mean():.4f} (mean)")
    if out_dir:
        os.makedirs(out_dir, exist_ok=True)
        save_png(os.path.join(out_dir, "ground_truth.png"), frames["ground_truth"])
        save_png(os.path.join(out_dir, "neural.png"), frames["neural"])
        save_png(os.path.join(out_dir, "flip_error.png"), magma(fmap))
    return {"psnr": float(psnr), "flip": float(fmap.mean()), "gt": frames["ground_truth"], "neural": frames["neural"],
            "seconds": secs, "paths_per_pass": W * H * total}


def main(argv=None):
    ap = argparse.ArgumentParser(description="headless evaluate driver (reference src/evaluate.cu)")
    ap.add_argument("config")
    ap.add_argument("out", nargs="?", default="comparison_output")
    ap.add_argument("--width", type=int, default=None)
    ap.add_argument("--height", type=int, default=None)
    a = ap.parse_args(argv)
    print("=== Comparison Renderer ===")
    cfg = load_config(a.config)
    res = run(cfg, a.out, a.width, a.height)
    for k, s in res["seconds"].items():
        print(f"{k}: {s:.3f} s, {res['paths_per_pass'] / s / 1e6:.1f} Mrays/s")
    print("\nComparison complete.")
    return 0


if __name__ == "__main__":
    sys.exit(main())
