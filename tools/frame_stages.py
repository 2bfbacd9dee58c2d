#!/usr/bin/env python
"""tools/frame_stages.py - frame time and per-kernel device times of one configuration (development aid).

  python tools/frame_stages.py [--mode neural|gt] [--width 1920 --height 1080 --spp 1 --bounces 3 --subdiv 7] [--steps 20]
Prints one JSON line: frame ms (CUDA events over `steps` frames after 5 warm-up frames) and, from a second pass with
NRB_OPT_PROFILE, the per-kernel ms per frame.
"""
import argparse
import importlib
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--mode", default="neural")
    ap.add_argument("--width", type=int, default=1920)
    ap.add_argument("--height", type=int, default=1080)
    ap.add_argument("--spp", type=int, default=1)
    ap.add_argument("--bounces", type=int, default=3)
    ap.add_argument("--subdiv", type=int, default=7)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--tag", default="")
    ap.add_argument("--query-kernel", type=int, default=0)
    a = ap.parse_args()
    nrb = importlib.import_module("neural-renderer_b200")
    scene = nrb.scenes.make_scene("ico", a.subdiv, shells=True, env=True, env_size=(2048, 1024))
    eye, basis = nrb.scenes.default_camera()
    r = nrb.RendererNeural(scene)
    r.resize(a.width, a.height)
    r.set_camera_basis(basis)
    r.set_samples_per_pixel(a.spp)
    r.set_bounce_count(a.bounces)
    if a.mode == "neural":
        r.load_weights(nrb.scenes.make_weights())
        r.set_use_neural_query(True)
        if a.query_kernel:
            r.set_query_kernel(a.query_kernel)
    for _ in range(5):
        r.render(eye)
    r.sync()
    best = 1e9
    for _ in range(3):
        t0 = time.perf_counter()
        for _ in range(a.steps):
            r.render(eye)
        r.sync()
        best = min(best, (time.perf_counter() - t0) * 1e3 / a.steps)
    c = r.counters()
    r.set_profile(True)
    r.reset_stages()
    for _ in range(a.steps):
        r.render(eye)
    r.sync()
    st = r.stages()
    out = {"tag": a.tag, "mode": a.mode, "frame_ms": round(best, 4), "mrays_per_s": round(a.width * a.height * a.spp / best * 1e-3, 1),
           "launches": c["kernels"], "counters": c,
           "kernels_ms": {k: [round(v[0] / a.steps, 4), int(v[1] / a.steps)] for k, v in sorted(st.items(), key=lambda kv: -kv[1][0])}}
    print(json.dumps(out))


if __name__ == "__main__":
    main()
