"""Deterministic synthetic scenes, shells, environments and weights (SURVEY.md §8d / BASELINE.md §3.3).

The reference tree holds no scene data (dbrt_data/ is absent), so every test and benchmark input is
generated here with fixed seeds, in the formats the reference consumes: indexed triangle meshes,
fp32 equirect environment, raw fp16 weight block in [hg | mlp] order
(src/cuda_renderer_neural.cu:1752-1808).
"""
from __future__ import annotations

import numpy as np


def uv_sphere(stacks: int = 48, slices: int = 96, radius: float = 1.0):
    """GenerateUvSphere (reference src/mesh_loader.cpp:719-764): vertices, normals, uvs, indices."""
    kPi = np.float32(3.14159265358979323846)
    i = np.arange(stacks + 1, dtype=np.float32)[:, None] / np.float32(stacks)
    j = np.arange(slices + 1, dtype=np.float32)[None, :] / np.float32(slices)
    phi = (i * kPi).astype(np.float32)
    theta = (j * np.float32(2.0) * kPi).astype(np.float32)
    x = (np.sin(phi) * np.cos(theta)).astype(np.float32)
    y = (np.cos(phi) * np.ones_like(theta)).astype(np.float32)
    z = (np.sin(phi) * np.sin(theta)).astype(np.float32)
    n = np.stack([x, y, z], -1).reshape(-1, 3)
    v = (n * np.float32(radius)).astype(np.float32)
    uv = np.stack([np.broadcast_to(j, x.shape), np.broadcast_to(i, x.shape)], -1).reshape(-1, 2).astype(np.float32)
    ii, jj = np.meshgrid(np.arange(stacks), np.arange(slices), indexing="ij")
    p0 = (ii * (slices + 1) + jj).reshape(-1)
    p1, p2 = p0 + 1, p0 + (slices + 1)
    p3 = p2 + 1
    idx = np.stack([np.stack([p0, p2, p1], -1), np.stack([p1, p2, p3], -1)], 1).reshape(-1, 3).astype(np.uint32)
    return v, n.astype(np.float32), uv, idx


def icosphere(subdivisions: int):
    """Unit icosphere: 20 * 4**subdivisions triangles, outward-facing (CCW seen from outside)."""
    t = (1.0 + 5.0 ** 0.5) / 2.0
    v = np.array([[-1, t, 0], [1, t, 0], [-1, -t, 0], [1, -t, 0], [0, -1, t], [0, 1, t], [0, -1, -t], [0, 1, -t],
                  [t, 0, -1], [t, 0, 1], [-t, 0, -1], [-t, 0, 1]], np.float64)
    v /= np.linalg.norm(v, axis=1, keepdims=True)
    f = np.array([[0, 11, 5], [0, 5, 1], [0, 1, 7], [0, 7, 10], [0, 10, 11], [1, 5, 9], [5, 11, 4], [11, 10, 2],
                  [10, 7, 6], [7, 1, 8], [3, 9, 4], [3, 4, 2], [3, 2, 6], [3, 6, 8], [3, 8, 9], [4, 9, 5],
                  [2, 4, 11], [6, 2, 10], [8, 6, 7], [9, 8, 1]], np.int64)
    for _ in range(subdivisions):
        e = np.concatenate([f[:, [0, 1]], f[:, [1, 2]], f[:, [2, 0]]], 0)
        es = np.sort(e, 1)
        key = es[:, 0] * (len(v) + 1) + es[:, 1]
        uniq, inv = np.unique(key, return_inverse=True)
        a, b = uniq // (len(v) + 1), uniq % (len(v) + 1)
        mid = v[a] + v[b]
        mid /= np.linalg.norm(mid, axis=1, keepdims=True)
        base = len(v)
        v = np.concatenate([v, mid], 0)
        n = len(f)
        m01, m12, m20 = base + inv[:n], base + inv[n:2 * n], base + inv[2 * n:]
        f = np.concatenate([np.stack([f[:, 0], m01, m20], 1), np.stack([f[:, 1], m12, m01], 1),
                            np.stack([f[:, 2], m20, m12], 1), np.stack([m01, m12, m20], 1)], 0)
    return v.astype(np.float32), f.astype(np.uint32)


def displace(vertices: np.ndarray, amplitude: float = 0.1, seed: int = 1337, octaves: int = 4):
    """Radial displacement amplitude * mean_k sin(f_k . p + phi_k) with mt19937(seed) frequencies/phases."""
    rng = np.random.RandomState(seed)
    p = vertices.astype(np.float64)
    d = np.zeros(len(p))
    for k in range(octaves):
        fvec = rng.uniform(-1, 1, 3) * (3.0 * 2 ** k)
        phi = rng.uniform(0, 2 * np.pi)
        d += np.sin(p @ fvec + phi) / (k + 1)
    d /= sum(1.0 / (k + 1) for k in range(octaves))
    r = 1.0 + amplitude * d
    return (p * r[:, None]).astype(np.float32)


def vertex_normals(vertices, indices):
    v = vertices.astype(np.float64)
    fn = np.cross(v[indices[:, 1]] - v[indices[:, 0]], v[indices[:, 2]] - v[indices[:, 0]])
    n = np.zeros_like(v)
    for c in range(3):
        np.add.at(n, indices[:, c], fn)
    n /= np.maximum(np.linalg.norm(n, axis=1, keepdims=True), 1e-30)
    return n.astype(np.float32)


def ground_quad(y: float = -1.3, half: float = 4.0):
    v = np.array([[-half, y, -half], [half, y, -half], [half, y, half], [-half, y, half]], np.float32)
    idx = np.array([[0, 2, 1], [0, 3, 2]], np.uint32)  # normal +y
    return v, idx


def sky_env(width: int = 2048, height: int = 1024, seed: int = 7):
    """Synthetic fp32 equirect: vertical gradient + a bright sun lobe + seeded low-amplitude noise."""
    rng = np.random.RandomState(seed)
    v = (np.arange(height, dtype=np.float32) + 0.5) / height
    u = (np.arange(width, dtype=np.float32) + 0.5) / width
    theta = v[:, None] * np.pi
    phi = (u[None, :] - 0.5) * 2 * np.pi
    d = np.stack([np.sin(theta) * np.sin(phi), np.cos(theta) * np.ones_like(phi), -np.sin(theta) * np.cos(phi)], -1)
    sun = np.array([0.4, 0.7, -0.59])
    sun /= np.linalg.norm(sun)
    lobe = np.clip(d @ sun, 0, 1) ** 256
    grad = 0.5 * (d[..., 1:2] + 1.0)
    img = (1 - grad) * np.array([0.9, 0.85, 0.8]) + grad * np.array([0.35, 0.55, 1.0])
    img = img + lobe[..., None] * np.array([60.0, 55.0, 45.0])
    img = img * (1.0 + 0.05 * rng.standard_normal((height, width, 1)))
    return np.ascontiguousarray(np.maximum(img, 0.0), np.float32)


def network_sizes(log2_hashmap_size=14, base_resolution=16, n_points=3, n_levels=8, n_features=4,
                  per_level_scale=2.0, sh_degree=4, mlp_width=128, n_hidden=4):
    """(grid halfs, mlp halfs, per-level offsets) — GridEncodingTemplated ctor, tiny-cuda-nn grid.h:699-730."""
    offsets = [0]
    log2s = np.float32(np.log2(np.float32(per_level_scale)))
    for l in range(n_levels):
        scale = np.float32(np.exp2(np.float32(l) * log2s)) * np.float32(base_resolution) - np.float32(1.0)
        res = int(np.ceil(scale)) + 1
        n = min(res ** 3, 0xFFFFFFFF // 2)
        n = (n + 7) // 8 * 8
        n = min(n, 1 << log2_hashmap_size)
        offsets.append(offsets[-1] + n)
    enc = n_points * n_levels * n_features + sh_degree * sh_degree
    mlp = mlp_width * enc + (n_hidden - 1) * mlp_width * mlp_width + 16 * mlp_width
    return offsets[-1] * n_features, mlp, offsets


def make_weights(seed: int = 42, grid_range: float = 0.5, out_bias_hit: float = 0.0, **net):
    """Seeded weight file in the reference's default [hg | mlp] order: grid U(-r, r), Xavier-uniform MLP."""
    ng, nm, _ = network_sizes(**net)
    rng = np.random.RandomState(seed)
    grid = rng.uniform(-grid_range, grid_range, ng).astype(np.float16)
    n_points, n_levels, n_features = net.get("n_points", 3), net.get("n_levels", 8), net.get("n_features", 4)
    shd, w, nh = net.get("sh_degree", 4), net.get("mlp_width", 128), net.get("n_hidden", 4)
    enc = n_points * n_levels * n_features + shd * shd
    mats = []
    fan_in = enc
    for _ in range(nh):
        lim = np.sqrt(6.0 / (fan_in + w))
        mats.append(rng.uniform(-lim, lim, (w, fan_in)))
        fan_in = w
    lim = np.sqrt(6.0 / (w + 16))
    mats.append(rng.uniform(-lim, lim, (16, w)))
    mlp = np.concatenate([m.reshape(-1) for m in mats]).astype(np.float16)
    assert len(mlp) == nm
    return np.concatenate([grid, mlp])


# ---- scene assembly ---------------------------------------------------------------------------------

def build_bvh(vertices, indices):
    """nrb_build_bvh (host SAH builder of libnrb200): -> (nodes, leaf-ordered indices, prim ids)."""
    import ctypes as C
    from . import _capi as K
    lib = K.product_lib()
    v = np.ascontiguousarray(vertices, np.float32)
    idx = np.ascontiguousarray(indices, np.uint32).copy()
    nodes = np.zeros(2 * len(idx), K.BVH_NODE_DTYPE)
    prim = np.zeros(len(idx), np.uint32)
    n = C.c_int(0)
    rc = lib.fn("build_bvh")(v.ctypes.data, len(v), idx.ctypes.data, len(idx), nodes.ctypes.data, C.byref(n),
                             prim.ctypes.data)
    if rc != 0:
        raise K.NrbError(f"nrb_build_bvh failed ({rc})")
    return nodes[: n.value].copy(), idx, prim


def build_bvh_gpu(vertices, indices, max_leaf: int = 4, device: int = -1):
    """nrb_build_bvh_gpu (LBVH on the device): -> (nodes, leaf-ordered indices, prim ids, build ms)."""
    import ctypes as C
    from . import _capi as K
    lib = K.product_lib()
    v = np.ascontiguousarray(vertices, np.float32)
    idx = np.ascontiguousarray(indices, np.uint32).copy()
    nodes = np.zeros(2 * len(idx), K.BVH_NODE_DTYPE)
    prim = np.zeros(len(idx), np.uint32)
    n, ms = C.c_int(0), C.c_float(0.0)
    rc = lib.fn("build_bvh_gpu")(v.ctypes.data, len(v), idx.ctypes.data, len(idx), nodes.ctypes.data, C.byref(n),
                                 prim.ctypes.data, device, max_leaf, C.byref(ms))
    if rc != 0:
        raise K.NrbError(f"nrb_build_bvh_gpu failed ({rc}): {lib.fn('build_bvh_gpu_error')().decode()}")
    return nodes[: n.value].copy(), idx, prim, ms.value


def remap_materials(material_map, material_ids, prim_ids):
    """The material map after the BVH build has reordered the triangles (Mesh::buildBvh, reference
    src/mesh_bvh.cpp:70-123): per-triangle material ids carried through prim_ids (new -> old triangle), then primitives
    rebuilt as runs of consecutive triangles with the same material.  -> (material_map u32, material_ids i32)."""
    material_map = np.asarray(material_map, np.int64)
    material_ids = np.asarray(material_ids, np.int64)
    n = len(prim_ids)
    per_old = np.full(n, -1, np.int64)
    for p, start in enumerate(material_map):
        end = material_map[p + 1] if p + 1 < len(material_map) else n
        per_old[start:end] = material_ids[p]
    per_new = per_old[np.asarray(prim_ids, np.int64)]
    if n == 0:
        return np.zeros(1, np.uint32), np.array([-1], np.int32)
    starts = np.concatenate([[0], np.nonzero(per_new[1:] != per_new[:-1])[0] + 1])
    return starts.astype(np.uint32), per_new[starts].astype(np.int32)


def mesh_from(vertices, indices, normals=None, texcoords=None, gpu_bvh: bool = False, builder=None, **kw):
    """builder: (vertices, indices) -> (nodes, leaf-ordered indices, prim ids, ...); default = the host SAH builder of
    libnrb200.  Callers that must not load libnrb200 (bench.py's reference arm) pass the reference's own Mesh::buildBvh.
    A material map with more than one primitive is carried through the reordering (remap_materials)."""
    from .renderer import MeshData
    if builder is not None:
        built = builder(vertices, indices)
    else:
        built = build_bvh_gpu(vertices, indices) if gpu_bvh else build_bvh(vertices, indices)
    nodes, idx = built[0], built[1]
    mm, mi = kw.get("material_map"), kw.get("material_ids")
    if mm is not None and mi is not None and len(mm) > 1:
        if len(built) < 3:
            raise ValueError("a multi-primitive material map needs the builder's primitive ids")
        kw = dict(kw)
        kw["material_map"], kw["material_ids"] = remap_materials(mm, mi, built[2])
    return MeshData(vertices=vertices, indices=idx, nodes=nodes, normals=normals, texcoords=texcoords, **kw)


def checker_texture(size: int = 64, seed: int = 3):
    rng = np.random.RandomState(seed)
    y, x = np.mgrid[0:size, 0:size]
    base = (((x // 8) + (y // 8)) % 2).astype(np.float32)
    img = np.stack([0.25 + 0.6 * base, 0.3 + 0.4 * (1 - base), 0.35 + 0.3 * base], -1)
    img = img + 0.05 * rng.standard_normal(img.shape)
    return np.clip(img * 255.0, 0, 255).astype(np.uint8)


def param_textures(size: int = 32, seed: int = 5):
    """Two seeded 3-channel 8-bit textures: [0] a channel-packed parameter map (G = roughness, B = metallic, the glTF
    packing) and [1] a tangent-space normal map (mostly +z with a seeded tilt)."""
    rng = np.random.RandomState(seed)
    packed = np.zeros((size, size, 3), np.float32)
    y, x = np.mgrid[0:size, 0:size]
    packed[..., 0] = 1.0
    packed[..., 1] = 0.15 + 0.7 * (((x // 4) + (y // 4)) % 2) + 0.05 * rng.standard_normal((size, size))
    packed[..., 2] = 0.9 * ((x // 8) % 2) + 0.05 * rng.standard_normal((size, size))
    tilt = 0.35 * rng.standard_normal((size, size, 2))
    n = np.concatenate([tilt, np.ones((size, size, 1))], -1)
    n /= np.linalg.norm(n, axis=-1, keepdims=True)
    to8 = lambda a: np.clip(a * 255.0, 0, 255).astype(np.uint8)
    return to8(packed), to8(0.5 * (n + 1.0))


def make_scene(kind: str = "ico", subdivisions: int = 4, shells: bool = True, env: bool = True,
               additional: bool = False, shell_delta: float = 0.12, textured=False,
               env_size=(512, 256), roughness: float = 0.5, metallic: float = 0.0, specular: float = 0.5,
               material: dict = None, open_outer: float = 0.0, builder=None):
    """Synthetic scene in the reference's Scene layout: a displaced icosphere (or the
    reference's UV sphere) as original mesh, +-shell_delta icosphere hulls as inner/outer shells, an
    optional ground quad as the additional mesh and a seeded equirect environment."""
    from . import _capi as K
    from .renderer import Scene
    if kind == "sphere":
        v, n, uv, idx = uv_sphere(48, 96, 1.0)
    elif kind == "sphere_small":
        v, n, uv, idx = uv_sphere(12, 24, 1.0)
    else:
        v0, idx = icosphere(subdivisions)
        v = displace(v0, 0.1)
        n = vertex_normals(v, idx)
        uv = np.stack([np.arctan2(v0[:, 0], v0[:, 2]) / (2 * np.pi) + 0.5, np.arccos(np.clip(v0[:, 1], -1, 1)) / np.pi],
                      -1).astype(np.float32)
    blobs = kind == "blobs"
    if blobs:
        # two separate objects in two separate shell components, lined up with the default camera (a small near one in front of a larger far one): rays
        # that leave the first component without a hit re-enter the shell at the second one, so the
        # segment loop needs more than one iteration
        d = np.array([2.0, 1.2, 2.0], np.float32)
        d /= np.linalg.norm(d)
        nv = len(v)
        v = np.concatenate([v * np.float32(0.3) + d * np.float32(0.9), v * np.float32(0.6) - d * np.float32(0.6)]).astype(np.float32)
        n = np.concatenate([n, n]).astype(np.float32)
        uv = np.concatenate([uv, uv]).astype(np.float32)
        idx = np.concatenate([idx, idx + np.uint32(nv)]).astype(np.uint32)
    extra = {}
    mat_kw = dict(base_color=(0.8, 0.8, 0.8), roughness=roughness, metallic=metallic, specular=specular)
    extras_direct = {}
    if material:
        mat_kw.update({k: v for k, v in material.items() if k not in ("ior", "specular_transmission")})
        extras_direct = {k: v for k, v in material.items() if k in ("ior", "specular_transmission")}
    if textured:
        m = K.Material.from_config(**mat_kw)
        m.base_color.textured = 1
        m.base_color.texture_id = 0
        textures = [(checker_texture(), False)]
        nearest = False
        if textured == "packed":
            # everything resolveMaterial / computeHitData can read from textures (src/mesh_intersection.cuh:109-152,
            # 168-238, 244-337): base colour, channel-packed roughness (G) / metallic (B), a normal map, nearest filtering
            packed, nmap = param_textures()
            textures += [(packed, False), (nmap, False)]
            m.roughness.textured, m.roughness.texture_id, m.roughness.channel = 1, 1, 1
            m.metallic.textured, m.metallic.texture_id, m.metallic.channel = 1, 1, 2
            m.normal.textured, m.normal.texture_id = 1, 2
            nearest = True
        extra = dict(materials=[m], material_map=np.array([0], np.uint32), material_ids=np.array([0], np.int32),
                     textures=textures, has_mesh_materials=True, texture_nearest=nearest)
    scene = Scene(original_mesh=mesh_from(v, idx, normals=n, texcoords=uv, builder=builder, **extra))
    if shells:
        hv, hidx = icosphere(3)
        if blobs:
            two = lambda r: np.concatenate([hv * np.float32(0.3 * r) + d * np.float32(0.9), hv * np.float32(0.6 * r) - d * np.float32(0.6)]).astype(np.float32)
            hidx2 = np.concatenate([hidx, hidx + np.uint32(len(hv))]).astype(np.uint32)
            scene.inner_shell = mesh_from(two(1.0 - shell_delta), hidx2, builder=builder)
            scene.outer_shell = mesh_from(two(1.0 + shell_delta), hidx2, builder=builder)
        else:
            scene.inner_shell = mesh_from((hv * np.float32(1.0 - shell_delta)).astype(np.float32), hidx, builder=builder)
            if open_outer > 0.0:
                # an OPEN outer shell: a cap on the side away from the default camera is removed, so camera rays through
                # the middle enter the shell, reach the inner shell and have NO outer back face anywhere ahead - the
                # exitT = 1e-8 fallback of traceSegmentExitsKernel (src/cuda_renderer_neural.cu:899-905)
                c = hv[hidx].mean(1)
                c /= np.linalg.norm(c, axis=1, keepdims=True)
                d0 = np.array([2.0, 1.2, 2.0]) / np.linalg.norm([2.0, 1.2, 2.0])
                hidx = hidx[(c @ d0) > -(1.0 - open_outer)]
            scene.outer_shell = mesh_from((hv * np.float32(1.0 + shell_delta)).astype(np.float32), hidx, builder=builder)
    if additional:
        gv, gidx = ground_quad()
        scene.additional_mesh = mesh_from(gv, gidx, builder=builder)
    if env:
        scene.env_pixels = sky_env(*env_size)
    scene.global_material = K.Material.from_config(**mat_kw)
    for k, v in extras_direct.items():
        setattr(scene.global_material, k, v)
    return scene


def default_camera(radius: float = 1.0, fov_y: float = 0.785):
    """Look-at camera of SURVEY.md §8d: eye (2.0, 1.2, 2.0) * r toward the origin."""
    from .renderer import look_at_basis
    eye = np.array([2.0, 1.2, 2.0], np.float32) * np.float32(radius)
    return eye, look_at_basis(eye, (0.0, 0.0, 0.0), fov_y)
