"""Driver of the reference's own CUDA build (baseline/_ref/, built by baseline/Makefile where /root/reference exists).

Test / bench infrastructure only - the GPU arbiter and timed GPU baseline of SURVEY.md 8c row 2:
  * ``write_pack``   serialises a scene (the host mirror's ``Scene``), camera, options and weights into the flat file
                     ``ref_render`` reads;
  * ``ref_render``   runs the reference's ``RendererNeural`` on it (its own ``Mesh::buildBvh`` unless nodes are sent) and
                     returns its frame, hit buffers, timing line and the BVHs its builder produced;
  * ``tcnn_query``   runs tiny-cuda-nn's ``inference_mixed_precision`` configured like the renderer's network.
Nothing in neural-renderer_b200/ imports this module.  The binaries travel to the GPU box with the snapshot; nothing
here reads /root/reference at run time.
"""
from __future__ import annotations

import ctypes as C
import json
import os
import struct
import subprocess
import tempfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
BIN = os.path.join(HERE, "_ref")

BVH_NODE_DTYPE = np.dtype([("bounds_min", "<f4", 3), ("bounds_max", "<f4", 3), ("left", "<i4"), ("right", "<i4"),
                           ("first", "<i4"), ("count", "<i4"), ("is_leaf", "<i4")])


def have(name: str = "ref_render") -> bool:
    return os.path.exists(os.path.join(BIN, name))


def _mesh_bytes(m, send_nodes: bool) -> bytes:
    if m is None or len(m.indices) == 0:
        return struct.pack("<10i", *([0] * 10))
    nv, nt = len(m.vertices), len(m.indices)
    nodes = m.nodes if send_nodes else np.zeros(0, BVH_NODE_DTYPE)
    mats = list(m.materials)
    nprim = 0 if m.material_map is None else len(m.material_map)
    out = [struct.pack("<10i", nv, nt, int(m.normals is not None), int(m.texcoords is not None), len(nodes), len(mats),
                       nprim, len(m.textures), int(m.texture_nearest), int(m.has_mesh_materials))]
    out.append(np.ascontiguousarray(m.vertices, np.float32).tobytes())
    if m.normals is not None:
        out.append(np.ascontiguousarray(m.normals, np.float32).tobytes())
    if m.texcoords is not None:
        out.append(np.ascontiguousarray(m.texcoords, np.float32).tobytes())
    out.append(np.ascontiguousarray(m.indices, np.uint32).tobytes())
    out.append(np.ascontiguousarray(nodes).tobytes())
    for mat in mats:
        out.append(bytes(mat))
    if nprim:
        out.append(np.ascontiguousarray(m.material_map, np.uint32).tobytes())
        out.append(np.ascontiguousarray(m.material_ids, np.int32).tobytes())
    for px, srgb in m.textures:
        px = np.ascontiguousarray(px, np.uint8)
        h, w, c = px.shape
        out.append(struct.pack("<4i", w, h, c, int(srgb)))
        out.append(px.tobytes())
    return b"".join(out)


def write_pack(path, scene, eye, basis, width, height, spp=1, bounces=0, neural=False, weights=None, lambert=False,
               classic_mesh_index=0, use_additional=True, const_color=None, direct_env=None, log2_hashmap_size=14,
               base_resolution=16, accum_frames=1, swap_order=True, env_rotation=0.0, send_nodes=False):
    ints = [width, height, spp, bounces, int(neural), int(lambert), classic_mesh_index, int(use_additional),
            int(const_color is not None), int(direct_env is not None), log2_hashmap_size, base_resolution, accum_frames,
            int(swap_order), 0, 0]
    cc = const_color if const_color is not None else (0.8, 0.8, 0.8)
    de = direct_env if direct_env is not None else (0.0, 0.0, 0.0)
    floats = [*map(float, eye), *map(float, basis.forward), *map(float, basis.right), *map(float, basis.up),
              float(basis.fov_y), float(env_rotation), float(scene.env_strength), *map(float, cc), *map(float, de), 0.0, 0.0, 0.0]
    with open(path, "wb") as f:
        f.write(b"NRBPACK1")
        f.write(struct.pack("<16i", *ints))
        f.write(struct.pack("<24f", *floats))
        for m in (scene.original_mesh, scene.inner_shell, scene.outer_shell, scene.additional_mesh):
            f.write(_mesh_bytes(m, send_nodes))
        if scene.env_pixels is not None:
            e = np.ascontiguousarray(scene.env_pixels, np.float32)
            f.write(struct.pack("<2i", e.shape[1], e.shape[0]))
            f.write(e.tobytes())
        else:
            f.write(struct.pack("<2i", 0, 0))
        f.write(bytes(scene.global_material))
        w = b"" if weights is None else np.ascontiguousarray(weights).view(np.uint16).tobytes()
        f.write(struct.pack("<q", len(w)))
        f.write(w)


def read_bvh(path):
    """-> (nodes, leaf-ordered indices, material_map, material_ids) as Mesh::buildBvh left them."""
    raw = open(path, "rb").read()
    nn, nt, nprim, _ = struct.unpack_from("<4i", raw, 0)
    off = 16
    nodes = np.frombuffer(raw, BVH_NODE_DTYPE, nn, off).copy()
    off += nn * 44
    idx = np.frombuffer(raw, np.uint32, nt * 3, off).reshape(nt, 3).copy()
    off += nt * 12
    mmap = np.frombuffer(raw, np.uint32, nprim, off).copy()
    off += nprim * 4
    mids = np.frombuffer(raw, np.int32, nprim, off).copy()
    return nodes, idx, mmap, mids


def ref_render(scene, eye, basis, width, height, warmup=0, timed=0, prof=False, out_dir=None, keep=False, **kw):
    """Run the reference renderer; -> dict(pixels, hit_flags, ..., timing, bvh={slot: (nodes, indices, map, ids)})."""
    exe = os.path.join(BIN, "ref_render_prof" if prof else "ref_render")
    if not os.path.exists(exe):
        raise FileNotFoundError(f"{exe} is missing: run `make -C baseline` where /root/reference exists")
    tmp = out_dir or tempfile.mkdtemp(prefix="nrb_ref_")
    os.makedirs(tmp, exist_ok=True)
    pack = os.path.join(tmp, "scene.pack")
    write_pack(pack, scene, eye, basis, width, height, **kw)
    res = subprocess.run([exe, pack, tmp, str(warmup), str(timed)], capture_output=True, text=True, timeout=1800)
    if res.returncode != 0:
        raise RuntimeError(f"ref_render failed ({res.returncode}): {res.stdout[-2000:]} {res.stderr[-2000:]}")
    spp = max(1, kw.get("spp", 1))
    n = width * height * spp
    out = {"stdout": res.stdout}

    def rd(name, dt, shape):
        p = os.path.join(tmp, name)
        return np.fromfile(p, dt).reshape(shape) if os.path.exists(p) else None

    out["pixels"] = rd("pixels.rgba", np.uint8, (height, width, 4))
    out["hit_flags"] = rd("hit_flags.i32", np.int32, (n,))
    for k in ("hit_positions", "hit_normals", "hit_colors", "hit_material_params", "path_radiance"):
        out[k] = rd(k + ".f32", np.float32, (n, 3))
    out["hit_distances"] = rd("hit_distances.f32", np.float32, (n,))
    out["accum"] = rd("accum.f32", np.float32, (width * height, 3))
    out["timing"] = json.loads(open(os.path.join(tmp, "timing.json")).read())
    out["bvh"] = {}
    for slot in range(4):
        p = os.path.join(tmp, f"bvh_{slot}.bin")
        if os.path.exists(p):
            out["bvh"][slot] = read_bvh(p)
    if not keep and out_dir is None:
        import shutil
        shutil.rmtree(tmp, ignore_errors=True)
    return out


def scene_with_reference_bvh(scene, bvh):
    """The same scene with every mesh's BVH / triangle order replaced by what the reference's builder produced
    (slots: 0 original, 1 inner, 2 outer, 3 additional), so both implementations traverse ONE tree."""
    import copy
    import dataclasses
    s = copy.copy(scene)
    names = ("original_mesh", "inner_shell", "outer_shell", "additional_mesh")
    for slot, name in enumerate(names):
        m = getattr(scene, name)
        if m is None or slot not in bvh:
            continue
        nodes, idx, mmap, mids = bvh[slot]
        kw = dict(nodes=nodes, indices=idx)
        if len(mmap):
            kw.update(material_map=mmap, material_ids=mids)
        setattr(s, name, dataclasses.replace(m, **kw))
    return s


def tcnn_time(net: dict, weights, n: int, reps: int = 10):
    """Device time of tiny-cuda-nn's inference_mixed_precision on n seeded uniform queries -> timing dict."""
    exe = os.path.join(BIN, "tcnn_query")
    d = dict(log2_hashmap_size=14, base_resolution=16, n_points=3, n_levels=8, n_features=4, per_level_scale=2.0,
             sh_degree=4, mlp_width=128, n_hidden=4)
    d.update(net)
    with tempfile.TemporaryDirectory(prefix="nrb_tcnn_") as tmp:
        wp = os.path.join(tmp, "w.bin")
        np.ascontiguousarray(weights).view(np.uint16).tofile(wp)
        args = [exe, d["log2_hashmap_size"], d["base_resolution"], d["n_levels"], d["n_features"], d["per_level_scale"],
                d["n_points"], d["sh_degree"], d["mlp_width"], d["n_hidden"], wp, 1, "@rand", n, "@none", reps]
        res = subprocess.run([str(a) for a in args], capture_output=True, text=True, timeout=1800)
    if res.returncode != 0:
        raise RuntimeError(f"tcnn_query failed ({res.returncode}): {res.stdout[-1000:]} {res.stderr[-1000:]}")
    for line in res.stdout.splitlines():
        if line.startswith("{"):
            return json.loads(line)
    return None


def tcnn_query(net: dict, weights, inputs, swap_order=True, time_reps=0):
    """tiny-cuda-nn inference_mixed_precision on `inputs` [n, input_dims] fp32 -> ([n,16] fp16, timing dict or None).
    net keys as scenes.network_sizes: log2_hashmap_size, base_resolution, n_points, n_levels, n_features,
    per_level_scale, sh_degree, mlp_width, n_hidden."""
    exe = os.path.join(BIN, "tcnn_query")
    if not os.path.exists(exe):
        raise FileNotFoundError(f"{exe} is missing: run `make -C baseline` where /root/reference exists")
    d = dict(log2_hashmap_size=14, base_resolution=16, n_points=3, n_levels=8, n_features=4, per_level_scale=2.0,
             sh_degree=4, mlp_width=128, n_hidden=4)
    d.update(net)
    x = np.ascontiguousarray(inputs, np.float32)
    with tempfile.TemporaryDirectory(prefix="nrb_tcnn_") as tmp:
        wp, ip, op = (os.path.join(tmp, n) for n in ("w.bin", "in.f32", "out.f16"))
        np.ascontiguousarray(weights).view(np.uint16).tofile(wp)
        x.tofile(ip)
        args = [exe, d["log2_hashmap_size"], d["base_resolution"], d["n_levels"], d["n_features"], d["per_level_scale"],
                d["n_points"], d["sh_degree"], d["mlp_width"], d["n_hidden"], wp, int(swap_order), ip, len(x), op, time_reps]
        res = subprocess.run([str(a) for a in args], capture_output=True, text=True, timeout=1800)
        if res.returncode != 0:
            raise RuntimeError(f"tcnn_query failed ({res.returncode}): {res.stdout[-2000:]} {res.stderr[-2000:]}")
        out = np.fromfile(op, np.float16).reshape(len(x), -1)
    timing = None
    for line in res.stdout.splitlines():
        if line.startswith("{"):
            timing = json.loads(line)
    return out, timing
