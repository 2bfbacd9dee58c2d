#!/usr/bin/env python
"""Experiment: does running the frame as S concurrent row-band slices (S renderers with nrb_set_tile on ONE
GPU, one stream each) shorten it?  Prints ms per full frame for S = 1..4."""
import importlib
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

nrb = importlib.import_module("neural-renderer_b200")


def main():
    mode = sys.argv[1] if len(sys.argv) > 1 else "neural"
    W, H = 1920, 1080
    scene = nrb.scenes.make_scene("ico", 7, shells=True, env=True, env_size=(2048, 1024))
    weights = nrb.scenes.make_weights()
    eye, basis = nrb.scenes.default_camera()
    for S in (1, 2, 3, 4, 6):
        rs = []
        for v in range(S):
            r = nrb.RendererNeural(scene)
            r.resize(W, H)
            r.set_camera_basis(basis)
            r.set_bounce_count(3)
            if mode == "neural":
                r.load_weights(weights)
                r.set_use_neural_query(True)
            if S > 1:
                r.set_tile(16, v, S)
            rs.append(r)
        for _ in range(3):
            for r in rs:
                r.render(eye)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        steps = 20
        t0 = time.time()
        e0.record()
        for _ in range(steps):
            for r in rs:
                r.render(eye)
        e1.record()
        host = time.time() - t0
        torch.cuda.synchronize()
        print(f"{mode} S={S}: {e0.elapsed_time(e1) / steps:.3f} ms per frame (host enqueue {host / steps * 1e3:.3f} ms)", flush=True)
        for r in rs:
            r.close()


if __name__ == "__main__":
    main()
