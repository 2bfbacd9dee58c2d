#!/usr/bin/env python
"""tools/large_bvh_probe.py - the traversal's HBM / L2 claims on a BVH that does NOT fit L1 or L2 (SURVEY.md 8d scene S3:
displaced icosphere, 9 subdivisions = 5 242 880 triangles; 0.42 GB of node pairs + triangle blocks).

Prints one JSON line: tree size, node pairs / triangles fetched per ray (device counters of the instrumented
k_trace_rays), the algorithmic bytes per ray they imply (ours: 64 B per pair + 48 B per triangle + 16 B hit record; the
reference layout: 44 B per node x 2 + 48 B per triangle), rays/s of the GT primary pass and of a ground-truth frame.
Run it under `ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,lts__t_bytes.sum,gpu__time_duration.sum` for the
measured traffic beside it.
  python tools/large_bvh_probe.py [--subdiv 9] [--width 1920 --height 1080]
"""
import argparse
import importlib
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--subdiv", type=int, default=9)
    ap.add_argument("--width", type=int, default=1920)
    ap.add_argument("--height", type=int, default=1080)
    ap.add_argument("--steps", type=int, default=10)
    a = ap.parse_args()
    nrb = importlib.import_module("neural-renderer_b200")
    from tests.util import primary_rays, random_rays
    t0 = time.time()
    v0, f = nrb.scenes.icosphere(a.subdiv)
    v = nrb.scenes.displace(v0, 0.1)
    n = nrb.scenes.vertex_normals(v, f)
    nodes, idx, _ = nrb.scenes.build_bvh(v, f)
    build_s = time.time() - t0
    scene = nrb.Scene(original_mesh=nrb.MeshData(vertices=v, indices=idx, nodes=nodes, normals=n))
    scene.env_pixels = nrb.scenes.sky_env(2048, 1024)
    scene.global_material = nrb.Material.from_config(base_color=(0.8, 0.8, 0.8), roughness=0.5, metallic=0.0, specular=0.5)
    eye, basis = nrb.scenes.default_camera()
    r = nrb.RendererNeural(scene)
    out = {"triangles": int(len(f)), "nodes": int(len(nodes)), "node_pairs_MB": (len(nodes) // 2) * 64 / 1e6,
           "triangle_blocks_MB": len(f) * 48 / 1e6, "scene_build_s": round(build_s, 1)}
    for name, (o, d) in (("primary", primary_rays(a.width, a.height)), ("incoherent", random_rays(1 << 20, seed=3))):
        r.trace_rays(nrb.SLOT_ORIGINAL, nrb.TRACE_ANY, o[:1024], d[:1024])
        t0 = time.time()
        h = r.trace_rays(nrb.SLOT_ORIGINAL, nrb.TRACE_FORWARD_ONLY if name == "primary" else nrb.TRACE_ANY, o, d)
        st = r.trace_stats()
        pr, tr = st["node_pairs"] / len(o), st["triangles"] / len(o)
        out[name] = {"rays": int(len(o)), "hit_fraction": float((h["t"] < 1e30).mean()), "node_pairs_per_ray": pr, "triangles_per_ray": tr,
                     "algorithmic_bytes_per_ray": pr * 64 + tr * 48 + 16,
                     "algorithmic_bytes_per_ray_reference_layout": pr * 2 * 44 + tr * 48 + 16}
    r.resize(a.width, a.height)
    r.set_camera_basis(basis)
    r.set_samples_per_pixel(1)
    r.set_bounce_count(3)
    for _ in range(3):
        r.render(eye)
    r.sync()
    t0 = time.perf_counter()
    for _ in range(a.steps):
        r.render(eye)
    r.sync()
    ms = (time.perf_counter() - t0) * 1e3 / a.steps
    r.set_profile(True)
    r.reset_stages()
    for _ in range(a.steps):
        r.render(eye)
    r.sync()
    st = r.stages()
    out["gt_frame"] = {"ms": ms, "mrays_per_s": a.width * a.height / ms * 1e-3, "counters": r.counters(),
                       "kernels_ms": {k: v_[0] / a.steps for k, v_ in st.items()}}
    p_ms = st["k_gt_primary"][0] / a.steps
    out["gt_primary_algorithmic_GBs"] = out["primary"]["algorithmic_bytes_per_ray"] * a.width * a.height / (p_ms * 1e-3) / 1e9
    print(json.dumps(out))


if __name__ == "__main__":
    main()
