"""glTF loader of the evaluate driver (neural-renderer_b200/evaluate.py:load_gltf), a restatement of the reference's
LoadGltfWithMaterials (src/mesh_loader.cpp:160-624), and the material-map remap of Mesh::buildBvh
(src/mesh_bvh.cpp:70-123).  CPU only: a GLB is assembled here byte by byte."""
import io
import json
import struct

import numpy as np
import pytest

from tests.util import nrb


def _glb(tmp_path, embed_png=True):
    """Two nodes: a quad (2 triangles, u16 indices, normals + uvs, material 1 with a base colour texture) under a TRS
    transform with a child holding a triangle (u32 indices, no normals / uvs, material 0, matrix transform)."""
    quad_pos = np.array([[0, 0, 0], [1, 0, 0], [1, 1, 0], [0, 1, 0]], np.float32)
    quad_nrm = np.tile(np.array([0, 0, 1], np.float32), (4, 1))
    quad_uv = np.array([[0, 0], [1, 0], [1, 1], [0, 1]], np.float32)
    quad_idx = np.array([0, 1, 2, 0, 2, 3], np.uint16)
    tri_pos = np.array([[0, 0, 0], [0, 2, 0], [0, 0, 2]], np.float32)
    tri_idx = np.array([0, 1, 2], np.uint32)
    blobs, views = [], []

    def add(b, target=None):
        off = sum(len(x) for x in blobs)
        pad = (4 - len(b) % 4) % 4
        blobs.append(b + b"\0" * pad)
        views.append({"buffer": 0, "byteOffset": off, "byteLength": len(b)})
        return len(views) - 1

    v_pos, v_nrm, v_uv, v_idx = add(quad_pos.tobytes()), add(quad_nrm.tobytes()), add(quad_uv.tobytes()), add(quad_idx.tobytes())
    v_tpos, v_tidx = add(tri_pos.tobytes()), add(tri_idx.tobytes())
    images = []
    if embed_png:
        from PIL import Image
        px = np.zeros((2, 3, 3), np.uint8)
        px[..., 0], px[0, 1, 1] = 200, 77
        buf = io.BytesIO()
        Image.fromarray(px, "RGB").save(buf, "PNG")
        images.append({"bufferView": add(buf.getvalue()), "mimeType": "image/png"})
    doc = {
        "asset": {"version": "2.0"},
        "scenes": [{"nodes": [0]}],
        "nodes": [
            {"mesh": 0, "translation": [1.0, 2.0, 3.0], "rotation": [0.0, 0.0, 0.70710678, 0.70710678], "scale": [2.0, 2.0, 2.0],
             "children": [1]},
            {"mesh": 1, "matrix": [1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0, 0.5, 0, 0, 1]},
        ],
        "meshes": [
            {"primitives": [{"attributes": {"POSITION": 0, "NORMAL": 1, "TEXCOORD_0": 2}, "indices": 3, "material": 1}]},
            {"primitives": [{"attributes": {"POSITION": 4}, "indices": 5}]},
        ],
        "accessors": [
            {"bufferView": v_pos, "componentType": 5126, "count": 4, "type": "VEC3"},
            {"bufferView": v_nrm, "componentType": 5126, "count": 4, "type": "VEC3"},
            {"bufferView": v_uv, "componentType": 5126, "count": 4, "type": "VEC2"},
            {"bufferView": v_idx, "componentType": 5123, "count": 6, "type": "SCALAR"},
            {"bufferView": v_tpos, "componentType": 5126, "count": 3, "type": "VEC3"},
            {"bufferView": v_tidx, "componentType": 5125, "count": 3, "type": "SCALAR"},
        ],
        "materials": [
            {"pbrMetallicRoughness": {"baseColorFactor": [0.1, 0.2, 0.3, 1.0], "metallicFactor": 0.25, "roughnessFactor": 0.75},
             "emissiveFactor": [0.5, 0.0, 0.0],
             "extensions": {"KHR_materials_ior": {"ior": 1.33}, "KHR_materials_transmission": {"transmissionFactor": 0.4},
                            "KHR_materials_emissive_strength": {"emissiveStrength": 3.0}}},
            {"pbrMetallicRoughness": dict({"metallicFactor": 0.0}, **({"baseColorTexture": {"index": 0}} if embed_png else {}))},
        ],
        "bufferViews": views,
        "buffers": [{"byteLength": 0}],
    }
    if embed_png:
        doc["images"] = images
        doc["textures"] = [{"source": 0}]
    binary = b"".join(blobs)
    doc["buffers"][0]["byteLength"] = len(binary)
    js = json.dumps(doc).encode()
    js += b" " * ((4 - len(js) % 4) % 4)
    glb = struct.pack("<4sII", b"glTF", 2, 12 + 8 + len(js) + 8 + len(binary)) + struct.pack("<II", len(js), 0x4E4F534A) + js + \
        struct.pack("<II", len(binary), 0x004E4942) + binary
    path = tmp_path / "scene.glb"
    path.write_bytes(glb)
    return str(path)


def test_gltf_geometry_transforms_and_materials(tmp_path):
    g = nrb.evaluate.load_gltf(_glb(tmp_path))
    v, idx = g["vertices"], g["indices"]
    assert v.shape == (7, 3) and idx.shape == (3, 3) and idx.dtype == np.uint32
    # node 0: scale 2, rotate 90 degrees about z (x -> y, y -> -x), translate (1, 2, 3)
    want_quad = np.array([[1, 2, 3], [1, 4, 3], [-1, 4, 3], [-1, 2, 3]], np.float32)
    assert np.allclose(v[:4], want_quad, atol=1e-5)
    # node 1 is a child: its matrix (translate x by 0.5) applies first, then the parent's transform
    local = np.array([[0.5, 0, 0], [0.5, 2, 0], [0.5, 0, 2]], np.float32)
    want_tri = np.stack([-2 * local[:, 1] + 1, 2 * local[:, 0] + 2, 2 * local[:, 2] + 3], 1)
    assert np.allclose(v[4:], want_tri, atol=1e-5)
    assert np.array_equal(idx, np.array([[0, 1, 2], [0, 2, 3], [4, 5, 6]], np.uint32))
    # normals: rotated (not scaled), default (0, 1, 0) where the primitive has none; uvs default (0, 0)
    assert np.allclose(g["normals"][:4], [[0, 0, 1]] * 4, atol=1e-6) and np.allclose(g["normals"][4:], [[0, 1, 0]] * 3)
    assert np.allclose(g["texcoords"][:4], [[0, 0], [1, 0], [1, 1], [0, 1]]) and np.all(g["texcoords"][4:] == 0)
    # one material-map entry per primitive, in node order; the second primitive has no material -> 0
    assert g["material_map"].tolist() == [0, 2] and g["material_ids"].tolist() == [1, 0]
    m0, m1 = g["materials"]
    assert np.allclose(list(m0.base_color.value), [0.1, 0.2, 0.3]) and not m0.base_color.textured
    assert m0.metallic.value == pytest.approx(0.25) and m0.roughness.value == pytest.approx(0.75)
    assert m0.ior == pytest.approx(1.33) and m0.specular_transmission == pytest.approx(0.4)
    assert m0.emission_scale == pytest.approx(3.0) and list(m0.base_emission.value) == [0.5, 0.0, 0.0]
    assert m1.base_color.textured == 1 and m1.base_color.texture_id == 0 and m1.roughness.value == pytest.approx(1.0)
    # images arrive as 8-bit RGBA (tinygltf's default); a base colour texture is tagged sRGB
    (px, srgb), = g["textures"]
    assert px.shape == (2, 3, 4) and srgb and px[0, 1].tolist() == [200, 77, 0, 255]


def test_gltf_without_materials_or_images_gets_the_default_material(tmp_path):
    path = _glb(tmp_path, embed_png=False)
    doc_path = tmp_path / "nomats.glb"
    raw = open(path, "rb").read()
    n = struct.unpack_from("<I", raw, 12)[0]
    doc = json.loads(raw[20:20 + n])
    del doc["materials"]
    for mesh in doc["meshes"]:
        for p in mesh["primitives"]:
            p.pop("material", None)
    js = json.dumps(doc).encode()
    js += b" " * ((4 - len(js) % 4) % 4)
    rest = raw[20 + n:]
    doc_path.write_bytes(struct.pack("<4sII", b"glTF", 2, 12 + 8 + len(js) + len(rest)) + struct.pack("<II", len(js), 0x4E4F534A) + js + rest)
    g = nrb.evaluate.load_gltf(str(doc_path))
    assert len(g["materials"]) == 1 and list(g["materials"][0].base_color.value) == pytest.approx([0.8, 0.8, 0.8])
    assert g["material_ids"].tolist() == [0, 0] and g["textures"] == []


def test_material_map_follows_the_bvh_reordering():
    """Mesh::buildBvh reorders triangles; the material of every triangle must travel with it (src/mesh_bvh.cpp:70-123)."""
    rng = np.random.RandomState(4)
    prim_ids = rng.permutation(20).astype(np.uint32)            # new -> old triangle
    mmap, mids = np.array([0, 5, 12], np.uint32), np.array([7, 3, 9], np.int32)
    new_map, new_ids = nrb.scenes.remap_materials(mmap, mids, prim_ids)
    old_mat = np.where(np.arange(20) < 5, 7, np.where(np.arange(20) < 12, 3, 9))
    per_new = old_mat[prim_ids]
    assert new_map[0] == 0 and np.all(np.diff(new_map.astype(int)) > 0)
    for k, start in enumerate(new_map):
        end = new_map[k + 1] if k + 1 < len(new_map) else 20
        assert np.all(per_new[start:end] == new_ids[k])
        if k + 1 < len(new_map):
            assert new_ids[k] != new_ids[k + 1]                  # runs are maximal


def test_gltf_mesh_goes_through_the_bvh_build_with_its_materials(tmp_path):
    """load_mesh on a .glb: scale, BVH build by libnrb200's host builder, remapped material map, nearest filtering on
    the original mesh (LoadMeshLabeled, evaluate.cu:153-176)."""
    mesh = nrb.evaluate.load_mesh({"path": _glb(tmp_path), "scale": 0.5}, "original mesh")
    assert mesh.num_triangles == 3 and mesh.has_mesh_materials and mesh.texture_nearest
    assert np.allclose(np.sort(mesh.vertices[:, 2]), np.sort(np.array([3, 3, 3, 3, 3, 3, 7], np.float32) * 0.5), atol=1e-5)
    # every triangle still carries the material of the primitive it came from: the quad's two triangles (vertices 0..3)
    # have material 1, the child triangle (vertices 4..6) material 0
    per_tri = np.zeros(3, int)
    for k, start in enumerate(mesh.material_map):
        end = mesh.material_map[k + 1] if k + 1 < len(mesh.material_map) else 3
        per_tri[start:end] = mesh.material_ids[k]
    for t in range(3):
        assert per_tri[t] == (1 if mesh.indices[t].max() <= 3 else 0)


def test_environment_formats(tmp_path):
    """EnvironmentMap::loadFromFile dispatch (src/scene.cu:289-316): .exr float RGB, 8-bit .png / 255, else Radiance .hdr."""
    import os
    os.environ.setdefault("OPENCV_IO_ENABLE_OPENEXR", "1")
    cv2 = pytest.importorskip("cv2")
    from PIL import Image
    rgb = (np.random.RandomState(1).rand(6, 9, 3) * 5).astype(np.float32)
    exr = str(tmp_path / "env.exr")
    if not cv2.imwrite(exr, rgb[..., ::-1]):
        pytest.skip("this OpenCV build has no OpenEXR writer")
    got = nrb.evaluate.load_env(exr)
    assert got.shape == (6, 9, 3) and got.dtype == np.float32 and np.array_equal(got, rgb)
    px = (np.random.RandomState(2).rand(4, 5, 3) * 255).astype(np.uint8)
    png = str(tmp_path / "env.png")
    Image.fromarray(px, "RGB").save(png)
    got = nrb.evaluate.load_env(png)
    assert np.array_equal(got, px.astype(np.float32) / np.float32(255.0))


def test_obj_normals_follow_the_assimp_path(tmp_path):
    """LoadMeshFromFile (src/mesh_loader.cpp:604-690): file normals where the OBJ has them, flat face normals (GenNormals +
    JoinIdenticalVertices) where it has none."""
    ev = nrb.evaluate
    smooth = tmp_path / "s.obj"
    smooth.write_text("v 0 0 0\nv 1 0 0\nv 0 1 0\nv 0 0 1\nvn 0 0 2\nvn 1 0 0\n"
                      "f 1//1 2//1 3//1\nf 1//2 3//2 4//2\n")
    v, f, n = ev.load_obj_shaded(str(smooth))
    assert len(v) == 6 and f.tolist() == [[0, 1, 2], [3, 4, 5]]          # vertex 1 and 3 appear with two normals each
    assert np.allclose(n[:3], [[0, 0, 1]] * 3) and np.allclose(n[3:], [[1, 0, 0]] * 3)   # normalised
    flat = tmp_path / "f.obj"
    flat.write_text("v 0 0 0\nv 1 0 0\nv 1 1 0\nv 0 1 0\nv 0 0 1\nf 1 2 3 4\nf 1 5 2\n")   # a quad in z = 0 and a triangle in y = 0
    v, f, n = ev.load_obj_shaded(str(flat))
    assert f.shape == (3, 3)
    tri_n = n[f]                                                           # every corner carries its face's normal
    geo = np.cross(v[f][:, 1] - v[f][:, 0], v[f][:, 2] - v[f][:, 0])
    geo /= np.linalg.norm(geo, axis=1, keepdims=True)
    assert np.allclose(tri_n, geo[:, None, :].repeat(3, 1), atol=1e-6)
    assert len(v) == 4 + 3                                                 # the quad's two triangles share their 4 corners; the other face has its own
    mesh = ev.load_mesh({"path": str(flat), "scale": 2.0}, "original mesh")
    assert mesh.num_triangles == 3 and np.isclose(np.abs(mesh.vertices).max(), 2.0) and mesh.normals.shape == mesh.vertices.shape


def test_ply_reader_variants(tmp_path):
    """Binary PLY in both byte orders with extra properties and elements, normals, quads and ragged polygon lists; a
    larger file goes through the whole-element numpy path (no per-row Python)."""
    import struct, time
    ev = nrb.evaluate
    head = ("ply\nformat {fmt} 1.0\ncomment made by a test\nelement vertex 5\nproperty float x\nproperty float y\nproperty float z\n"
            "property float nx\nproperty float ny\nproperty float nz\nproperty uchar red\n"
            "element face 2\nproperty uchar flags\nproperty list uchar int vertex_indices\n"
            "element edge 1\nproperty int a\nproperty int b\nend_header\n")
    pos = [(0, 0, 0), (1, 0, 0), (1, 1, 0), (0, 1, 0), (0, 0, 1)]
    for fmt, e in (("binary_little_endian", "<"), ("binary_big_endian", ">")):
        body = b"".join(struct.pack(e + "6fB", *p, 0, 0, 2, 7) for p in pos)
        body += struct.pack(e + "BB4i", 1, 4, 0, 1, 2, 3) + struct.pack(e + "BB3i", 0, 3, 0, 4, 1)   # a quad, then a triangle (ragged)
        body += struct.pack(e + "2i", 0, 1)
        path = tmp_path / (fmt + ".ply")
        path.write_bytes(head.format(fmt=fmt).encode() + body)
        v, f, n = ev.load_ply_full(str(path))
        assert v.tolist() == [list(map(float, p)) for p in pos]
        assert f.tolist() == [[0, 1, 2], [0, 2, 3], [0, 4, 1]]
        assert n.shape == (5, 3) and np.all(n[:, 2] == 2)
        mesh = ev.load_mesh({"path": str(path), "scale": 1.0}, "original mesh")
        assert mesh.num_triangles == 3 and np.allclose(np.linalg.norm(mesh.normals, axis=1), 1)     # file normals, normalised
    asc = tmp_path / "asc.ply"
    asc.write_text("ply\nformat ascii 1.0\nelement vertex 4\nproperty double x\nproperty double y\nproperty double z\n"
                   "element face 1\nproperty list uchar uint vertex_index\nend_header\n0 0 0\n1 0 0\n1 1 0\n\n0 1 0\n4 0 1 2 3\n")
    v, f, n = ev.load_ply_full(str(asc))
    assert n is None and v.dtype == np.float32 and f.tolist() == [[0, 1, 2], [0, 2, 3]]
    # 200 K triangles, uniform lists: one structured read
    rng = np.random.default_rng(0)
    nv, nf = 100_000, 200_000
    verts = rng.standard_normal((nv, 3)).astype("<f4")
    faces = np.zeros(nf, np.dtype([("n", "u1"), ("i", "<i4", (3,))]))
    faces["n"] = 3
    faces["i"] = rng.integers(0, nv, (nf, 3))
    big = tmp_path / "big.ply"
    big.write_bytes((f"ply\nformat binary_little_endian 1.0\nelement vertex {nv}\nproperty float x\nproperty float y\nproperty float z\n"
                     f"element face {nf}\nproperty list uchar int vertex_indices\nend_header\n").encode() + verts.tobytes() + faces.tobytes())
    t0 = time.time()
    v, f, n = ev.load_ply_full(str(big))
    assert time.time() - t0 < 2.0
    assert np.array_equal(v, verts) and np.array_equal(f, faces["i"].astype(np.uint32))


def _write_glb(path, prims, materials, png_images):
    """prims: (positions, normals, uvs, indices u32, material index) each its own mesh + node; images as PNG bytes."""
    blobs, views, accessors, meshes, nodes = [], [], [], [], []

    def add(b):
        off = sum(len(x) for x in blobs)
        blobs.append(b + b"\0" * ((4 - len(b) % 4) % 4))
        views.append({"buffer": 0, "byteOffset": off, "byteLength": len(b)})
        return len(views) - 1

    def acc(arr, ctype, kind):
        accessors.append({"bufferView": add(arr.tobytes()), "componentType": ctype, "count": len(arr), "type": kind})
        return len(accessors) - 1

    for pos, nrm, uv, idx, mat in prims:
        attrs = {"POSITION": acc(pos.astype(np.float32), 5126, "VEC3"), "NORMAL": acc(nrm.astype(np.float32), 5126, "VEC3"),
                 "TEXCOORD_0": acc(uv.astype(np.float32), 5126, "VEC2")}
        meshes.append({"primitives": [{"attributes": attrs, "indices": acc(idx.reshape(-1).astype(np.uint32), 5125, "SCALAR"),
                                       "material": mat}]})
        nodes.append({"mesh": len(meshes) - 1})
    doc = {"asset": {"version": "2.0"}, "scenes": [{"nodes": list(range(len(nodes)))}], "nodes": nodes, "meshes": meshes,
           "accessors": accessors, "materials": materials,
           "images": [{"bufferView": add(b), "mimeType": "image/png"} for b in png_images],
           "textures": [{"source": i} for i in range(len(png_images))], "bufferViews": views, "buffers": [{"byteLength": 0}]}
    binary = b"".join(blobs)
    doc["buffers"][0]["byteLength"] = len(binary)
    js = json.dumps(doc).encode()
    js += b" " * ((4 - len(js) % 4) % 4)
    open(path, "wb").write(struct.pack("<4sII", b"glTF", 2, 12 + 8 + len(js) + 8 + len(binary)) +
                           struct.pack("<II", len(js), 0x4E4F534A) + js + struct.pack("<II", len(binary), 0x004E4942) + binary)
    return str(path)


def _sphere_glb(tmp_path):
    """A UV sphere as two primitives: the northern half with a base-colour checker texture, the southern half a rough metal."""
    from PIL import Image
    nu, nv = 48, 24
    u, v = np.meshgrid(np.arange(nu + 1) / nu, np.arange(nv + 1) / nv)
    th, ph = v * np.pi, u * 2 * np.pi
    nrm = np.stack([np.sin(th) * np.cos(ph), np.cos(th), np.sin(th) * np.sin(ph)], -1).reshape(-1, 3)
    uv = np.stack([u, v], -1).reshape(-1, 2)
    a = (np.arange(nv)[:, None] * (nu + 1) + np.arange(nu)[None, :])
    quads = np.stack([a, a + nu + 1, a + nu + 2, a + 1], -1)               # [nv, nu, 4]
    tris = np.concatenate([quads[..., [0, 2, 1]], quads[..., [0, 3, 2]]], -1).reshape(nv, -1, 3)   # outward winding
    prims = []
    for rows, mat in ((slice(0, nv // 2), 0), (slice(nv // 2, nv), 1)):
        idx = tris[rows].reshape(-1, 3)
        idx = idx[(np.linalg.norm(np.cross(nrm[idx[:, 1]] - nrm[idx[:, 0]], nrm[idx[:, 2]] - nrm[idx[:, 0]]), axis=1) > 1e-9)]  # no pole slivers
        used, inv = np.unique(idx, return_inverse=True)
        prims.append((nrm[used], nrm[used], uv[used], inv.reshape(-1, 3), mat))
    yy, xx = np.mgrid[0:32, 0:64]
    checker = np.where(((yy // 4 + xx // 4) & 1)[..., None] == 1, [230, 60, 40], [40, 90, 220]).astype(np.uint8)
    buf = io.BytesIO()
    Image.fromarray(checker, "RGB").save(buf, "PNG")
    materials = [{"pbrMetallicRoughness": {"baseColorTexture": {"index": 0}, "metallicFactor": 0.0, "roughnessFactor": 0.6}},
                 {"pbrMetallicRoughness": {"baseColorFactor": [0.9, 0.7, 0.3, 1.0], "metallicFactor": 1.0, "roughnessFactor": 0.35}}]
    return _write_glb(tmp_path / "sphere.glb", prims, materials, [buf.getvalue()])


def _gltf_eval_config(tmp_path):
    import os
    ev = nrb.evaluate
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    cfg = ev.load_config(os.path.join(root, "configs", "synthetic_ico.json"))
    cfg["original_mesh"] = {"path": _sphere_glb(tmp_path), "scale": 1.0}
    cfg["rendering"]["total_samples"] = 8
    return cfg


def test_gltf_scene_through_the_evaluate_driver_on_the_oracle(tmp_path):
    """The loaded glTF scene (two primitives, per-primitive materials, an embedded texture) reaches the renderer: the CPU
    oracle's ground-truth frame shows both the checker's colours and the metal half."""
    from oracle import loader
    ev = nrb.evaluate
    cfg = _gltf_eval_config(tmp_path)
    scene = ev.build_scene(cfg)
    m = scene.original_mesh
    assert m.num_triangles > 2000 and len(m.materials) == 2 and len(m.textures) == 1 and m.has_mesh_materials
    b = ev.run(cfg, None, 96, 54, lib=loader.oracle_lib("auto"), quiet=True)
    img = b["gt"][..., :3].astype(int)
    reddish = (img[..., 0] > img[..., 2] + 40).mean()
    bluish = (img[..., 2] > img[..., 0] + 40).mean()
    assert reddish > 0.01 and bluish > 0.01 and img.std() > 5


@pytest.mark.gpu
def test_gltf_scene_renders_like_the_oracle(tmp_path):
    """The same glTF scene through libnrb200 and through the CPU oracle: ground-truth frames within the frame tests' bars."""
    from oracle import loader
    ev, metrics = nrb.evaluate, nrb.metrics
    cfg = _gltf_eval_config(tmp_path)
    a = ev.run(cfg, None, 192, 108, quiet=True)
    b = ev.run(cfg, None, 192, 108, lib=loader.oracle_lib("auto"), quiet=True)
    psnr, flip = metrics.psnr8(b["gt"], a["gt"]), metrics.flip_ldr(b["gt"], a["gt"])
    print(f"glTF scene: GT PSNR {psnr:.2f} dB, FLIP {flip:.5f}")
    assert psnr > 45.0 and flip <= 0.005
