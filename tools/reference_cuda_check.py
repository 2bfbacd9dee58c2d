#!/usr/bin/env python
"""tools/reference_cuda_check.py - libnrb200 against the REFERENCE's own CUDA + tiny-cuda-nn build on the same GPU.

Runs baseline/_ref/ref_render (the reference's RendererNeural, its Mesh::buildBvh, tiny-cuda-nn; built by
baseline/Makefile) and libnrb200 on the same seeded scene, camera, weights and - through the BVH the reference's
builder dumped - the same trees, then reports
  * hit-flag / hit-position / normal agreement of the primary pass,
  * PSNR and mean LDR-FLIP between the two 8-bit frames (north_star bars: FLIP <= 0.005),
  * frame time and Mrays/s of both (the reference: 3 warm-up + 20 timed render() calls, cudaEvents on the legacy
    stream; ours: the same protocol on our stream), and the reference's 13 PROFILE_KERNELS buckets.
One JSON document on stdout / --out.  GPU only; reads nothing under /root/reference.

  python tools/reference_cuda_check.py [--width 1920 --height 1080 --subdiv 7 --bounces 3 --spp 1 8] [--out f.json]
"""
import argparse
import importlib
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def time_product(r, eye, warmup, timed):
    for _ in range(warmup):
        r.render(eye)
    r.sync()
    t0 = time.perf_counter()
    for _ in range(timed):
        r.render(eye)
    r.sync()
    return (time.perf_counter() - t0) * 1e3 / timed


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--width", type=int, default=1920)
    ap.add_argument("--height", type=int, default=1080)
    ap.add_argument("--subdiv", type=int, default=7)
    ap.add_argument("--bounces", type=int, default=3)
    ap.add_argument("--spp", type=int, nargs="+", default=[1, 8])
    ap.add_argument("--modes", nargs="+", default=["neural", "gt"])
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--timed", type=int, default=20)
    ap.add_argument("--metallic", type=float, default=0.0)
    ap.add_argument("--roughness", type=float, default=0.5)
    ap.add_argument("--out", default=None)
    ap.add_argument("--no-flip", action="store_true")
    a = ap.parse_args()

    nrb = importlib.import_module("neural-renderer_b200")
    from baseline import refpack
    scene = nrb.scenes.make_scene("ico", a.subdiv, shells=True, env=True, env_size=(2048, 1024), roughness=a.roughness,
                                  metallic=a.metallic)
    eye, basis = nrb.scenes.default_camera()
    weights = nrb.scenes.make_weights()
    doc = {"scene": f"displaced icosphere 20*4^{a.subdiv} tris, +-12% icosphere(3) shells, 2048x1024 env, roughness {a.roughness} metallic {a.metallic}",
           "width": a.width, "height": a.height, "bounces": a.bounces, "runs": []}
    ref_bvh = None
    for mode in a.modes:
        for spp in a.spp:
            neural = mode == "neural"
            kw = dict(spp=spp, bounces=a.bounces, neural=neural, weights=weights)
            ref = refpack.ref_render(scene, eye, basis, a.width, a.height, warmup=a.warmup, timed=a.timed, **kw)
            refp = refpack.ref_render(scene, eye, basis, a.width, a.height, warmup=a.warmup, timed=a.timed, prof=True, **kw)
            if ref_bvh is None:
                ref_bvh = ref["bvh"]
            # the same trees on our side: the reference builder is multithreaded, so use THIS run's dump
            scene2 = refpack.scene_with_reference_bvh(scene, ref["bvh"])
            r = nrb.RendererNeural(scene2)
            r.resize(a.width, a.height)
            r.set_camera_basis(basis)
            r.set_samples_per_pixel(spp)
            r.set_bounce_count(a.bounces)
            r.load_weights(weights)
            r.set_use_neural_query(neural)
            r.set_capture(True)
            r.render(eye)
            px = r.read_pixels()
            flags, pos, nrm = r.read_buffer("hit_flags"), r.read_buffer("hit_positions"), r.read_buffer("hit_normals")
            r.set_capture(False)
            ms = time_product(r, eye, a.warmup, a.timed)
            counters = r.counters()
            r.close()
            both = (flags != 0) & (ref["hit_flags"] != 0)
            dpos = np.abs(pos[both] - ref["hit_positions"][both]).max(axis=1) if both.any() else np.zeros(0)
            dn = np.abs(nrm[both] - ref["hit_normals"][both]).max(axis=1) if both.any() else np.zeros(0)
            d8 = np.abs(px[..., :3].astype(int) - ref["pixels"][..., :3].astype(int))
            run = {
                "mode": mode, "spp": spp,
                "reference_cuda": {k: ref["timing"][k] for k in ("ms_per_frame", "mrays_per_s", "host_ms_per_frame", "bvh_build_ms")},
                "reference_cuda_buckets_ms": refp["timing"].get("buckets_ms"),
                "reference_cuda_profiled_ms_per_frame": refp["timing"]["ms_per_frame"],
                "nrb200": {"ms_per_frame": ms, "mrays_per_s": a.width * a.height * spp / ms * 1e-3, "counters": counters},
                "speedup": ref["timing"]["ms_per_frame"] / ms,
                "hit_flag_mismatch": int((flags != ref["hit_flags"]).sum()), "elements": int(flags.size),
                "hit_flag_mismatch_rate": float((flags != ref["hit_flags"]).mean()),
                "hit_position_bit_exact_rate": float((dpos == 0).mean()) if dpos.size else None,
                "hit_position_max_abs": float(dpos.max()) if dpos.size else None,
                "hit_position_p999_abs": float(np.percentile(dpos, 99.9)) if dpos.size else None,
                "hit_normal_max_abs": float(dn.max()) if dn.size else None,
                "psnr_db": float(nrb.metrics.psnr8(ref["pixels"], px)),
                "pixels_differing": float((d8.max(axis=2) > 0).mean()), "pixel_max_abs": int(d8.max()),
                "pixels_off_by_more_than_1": float((d8.max(axis=2) > 1).mean()),
            }
            if not a.no_flip:
                run["flip_mean"] = float(nrb.metrics.flip_ldr(ref["pixels"], px))
            doc["runs"].append(run)
            print(json.dumps(run), file=sys.stderr, flush=True)
    text = json.dumps(doc, indent=1)
    print(text)
    if a.out:
        os.makedirs(os.path.dirname(os.path.abspath(a.out)), exist_ok=True)
        open(a.out, "w").write(text)


if __name__ == "__main__":
    main()
