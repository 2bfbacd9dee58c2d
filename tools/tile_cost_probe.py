#!/usr/bin/env python
"""Distribution of the per-tile clocks of the ground-truth primary pass (development aid): how long is the longest tile
compared with the kernel?"""
import importlib, os, sys, json
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
nrb = importlib.import_module("neural-renderer_b200")
scene = nrb.scenes.make_scene("ico", 7, shells=True, env=True, env_size=(2048, 1024))
eye, basis = nrb.scenes.default_camera()
r = nrb.RendererNeural(scene)
W, H = 1920, 1080
r.resize(W, H); r.set_camera_basis(basis); r.set_samples_per_pixel(1); r.set_bounce_count(3)
r.set_frame_slices(1)
r.render(eye); r.sync()
tiles = ((W + 7) // 8) * ((H + 3) // 4)
buf = np.zeros(tiles, np.uint32)
r.lib.check(r.lib.fn("read_buffer")(r.handle, b"tile_cost", buf.ctypes.data, buf.nbytes), "read_buffer")
c = buf.astype(np.float64)
srt = np.sort(c)[::-1]
print(json.dumps({"tiles": tiles, "sum_Mclk": c.sum() / 1e6, "max_clk": srt[0], "top10": srt[:10].tolist(), "p99": float(np.percentile(c, 99)),
                  "p90": float(np.percentile(c, 90)), "median": float(np.median(c)), "mean": float(c.mean()),
                  "warps_resident": 148 * 40, "ideal_clk_if_balanced": c.sum() / (148 * 40)}))
