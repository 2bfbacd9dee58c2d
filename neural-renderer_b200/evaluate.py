"""Headless evaluate driver: the caller on one side of the hot path (reference src/evaluate.cu:118-310).

config JSON (the reference's schema, src/config_loader.cpp:10-144) -> scene -> ground-truth render and
neural render through RendererNeural, total_samples accumulated in batches of 8 spp -> PNGs, PSNR, FLIP.

  python tools/evaluate.py CONFIG.json [OUT_DIR] [--width W --height H]

Asset formats: the reference loads meshes through assimp/tinygltf and environments through stb/tinyexr
(out of scope).  Here: Wavefront OBJ and ASCII / binary-little-endian PLY meshes, Radiance
.hdr environments (RLE decoder below), raw fp16 .bin checkpoints, plus `synthetic:` pseudo-paths so the
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


def load_ply(path):
    f = open(path, "rb")
    if f.readline().strip() != b"ply":
        raise ValueError("not a PLY file")
    fmt, elems = None, []
    while True:
        p = f.readline().split()
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
    sizes = {"char": "b", "uchar": "B", "short": "h", "ushort": "H", "int": "i", "uint": "I", "float": "f", "double": "d",
             "int8": "b", "uint8": "B", "int16": "h", "uint16": "H", "int32": "i", "uint32": "I", "float32": "f", "float64": "d"}
    verts, faces = [], []
    for name, count, props in elems:
        for _ in range(count):
            if fmt == "ascii":
                tok = f.readline().split()
                if name == "vertex":
                    verts.append([float(tok[i]) for i in range(3)])
                elif name == "face":
                    n = int(tok[0])
                    idx = [int(x) for x in tok[1:1 + n]]
                    faces += [[idx[0], idx[k], idx[k + 1]] for k in range(1, n - 1)]
            elif fmt == "binary_little_endian":
                row = []
                for pr in props:
                    if pr[0] == "list":
                        n = struct.unpack("<" + sizes[pr[1]], f.read(struct.calcsize(sizes[pr[1]])))[0]
                        row.append(list(struct.unpack("<" + sizes[pr[2]] * n, f.read(struct.calcsize(sizes[pr[2]]) * n))))
                    else:
                        row.append(struct.unpack("<" + sizes[pr[0]], f.read(struct.calcsize(sizes[pr[0]])))[0])
                if name == "vertex":
                    names = [pr[-1] for pr in props]
                    verts.append([row[names.index(c)] for c in "xyz"])
                elif name == "face":
                    idx = next(x for x in row if isinstance(x, list))
                    faces += [[idx[0], idx[k], idx[k + 1]] for k in range(1, len(idx) - 1)]
            else:
                raise ValueError("unsupported PLY format " + str(fmt))
    return np.asarray(verts, np.float32), np.asarray(faces, np.uint32)


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
        v, idx = load_obj(path)
    elif ext == ".ply":
        v, idx = load_ply(path)
    else:
        raise ValueError(f"{what}: only .obj / .ply meshes are supported without assimp ({path})")
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
    env = cfg["environment"]
    if env["hdri_path"].startswith("synthetic:"):
        scene.env_pixels = scenes.sky_env(1024, 512)
    elif env["hdri_path"].lower().endswith(".hdr"):
        scene.env_pixels = load_hdr(env["hdri_path"])
    elif env["hdri_path"]:
        print(f"Failed to load HDRI '{env['hdri_path']}': only Radiance .hdr is supported here", file=sys.stderr)
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
