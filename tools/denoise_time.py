#!/usr/bin/env python
"""k_denoise at 1080p: CUDA-event time per launch (stage profiler) and the implied throughput."""
import importlib, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
nrb = importlib.import_module("neural-renderer_b200")
scene = nrb.scenes.make_scene("ico", 7, shells=True, env=True, env_size=(2048, 1024))
eye, basis = nrb.scenes.default_camera()
r = nrb.RendererNeural(scene)
r.resize(1920, 1080); r.set_camera_basis(basis); r.set_bounce_count(3); r.set_denoise(True)
for _ in range(3): r.render(eye)
r.set_profile(True); r.reset_stages()
for _ in range(10): r.render(eye)
r.sync()
for k, v in r.stages().items():
    print(f"{k:20s} {v[0] / v[1] * 1e3:9.1f} us per launch x{v[1]}")
