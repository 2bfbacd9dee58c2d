#!/usr/bin/env python
"""Synthetic batched neural-query sweep (BASELINE.json config 5): 2^18..2^24 queries through
nrb_neural_query (device pointers, CUDA events on the renderer's stream), reporting queries/s,
tensor TFLOP/s against the measured peak, compulsory HBM GB/s and gathered L2 GB/s.
  python tools/query_sweep.py [--kernel 0|1|2] [--log2 18 20 22] [--dist uniform|coherent] [--shape ref|w64]
"""
import argparse
import importlib
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--kernel", type=int, default=0)
    ap.add_argument("--log2", type=int, nargs="+", default=[18, 20, 22])
    ap.add_argument("--dist", default="uniform", choices=["uniform", "coherent"])
    ap.add_argument("--shape", default="ref", choices=["ref", "w64", "w128"])
    ap.add_argument("--tcnn", action="store_true", help="also time tiny-cuda-nn (baseline/_ref/tcnn_query) on the same shape and batch")
    ap.add_argument("--reps", type=int, default=10)
    a = ap.parse_args()
    import torch
    nrb = importlib.import_module("neural-renderer_b200")
    if a.shape == "ref":
        cfg, net = None, {}
        flops = 2 * (112 * 128 + 3 * 128 * 128 + 128 * 16)
        gathered, hbm = 3 * 8 * 8 * 8, 36 + 16
    else:
        W = 64 if a.shape == "w64" else 128
        cfg = nrb.NNConfig(log2_hashmap_size=19, base_resolution=16, n_points=1, n_levels=16, n_features_per_level=2,
                           per_level_scale=1.5, sh_degree=-1, mlp_width=W, n_hidden_layers=2)
        net = dict(log2_hashmap_size=19, base_resolution=16, n_points=1, n_levels=16, n_features=2, per_level_scale=1.5,
                   sh_degree=0, mlp_width=W, n_hidden=2)
        flops = 2 * (32 * W + W * W + W * 16)
        gathered, hbm = 16 * 8 * 4, 12 + 16
    r = nrb.RendererNeural(nrb.Scene(), nn_config=cfg)
    weights = nrb.scenes.make_weights(**net)
    r.load_weights(weights)
    r.set_query_kernel(a.kernel)
    D = r.input_dims()
    peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))) if os.path.exists(os.path.join(ROOT, "MEASURED_PEAKS.json")) else {"bf16_tflops": 1590.0, "hbm_gbs": 6650.0}
    stream = torch.cuda.ExternalStream(r.stream_ptr())
    fn = r.lib.fn("neural_query")
    for lg in a.log2:
        n = 1 << lg
        g = torch.Generator(device="cuda").manual_seed(99)
        x = torch.rand((n, D), device="cuda", generator=g)
        if a.dist == "coherent":  # queries ordered along a space-filling sweep: neighbours share grid cells
            t = torch.linspace(0, 1, n, device="cuda")
            base = torch.stack([t, (t * 37.0) % 1.0, (t * 1117.0) % 1.0], 1)
            x[:, :3] = base
            if D >= 9:
                x[:, 3:6] = (base + 0.01).clamp(0, 1)
                x[:, 6:9] = (base + 0.005).clamp(0, 1)
        out = torch.empty((n, 16), dtype=torch.float16, device="cuda")
        torch.cuda.synchronize()
        for _ in range(3):
            r.lib.check(fn(r.handle, x.data_ptr(), n, out.data_ptr()), "neural_query")
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        for _ in range(a.reps):
            r.lib.check(fn(r.handle, x.data_ptr(), n, out.data_ptr()), "neural_query")
        e1.record(stream)
        r.sync()
        ms = e0.elapsed_time(e1) / a.reps
        qps = n / (ms * 1e-3)
        extra = {}
        if a.tcnn:
            from baseline import refpack
            if refpack.have("tcnn_query"):
                t = refpack.tcnn_time(net, weights, n, reps=a.reps)
                extra = {"tcnn_ms": round(t["ms_per_call"], 4), "tcnn_Gqueries_per_s": round(t["gqueries_per_s"], 3),
                         "speedup_vs_tcnn": round(t["ms_per_call"] / ms, 2)}
        print(json.dumps({**extra, "queries": n, "kernel": a.kernel, "dist": a.dist, "shape": a.shape, "ms": round(ms, 4),
                          "Gqueries_per_s": round(qps / 1e9, 3), "tensor_TFLOPs": round(qps * flops / 1e12, 1),
                          "tensor_frac_of_measured_burst": round(qps * flops / 1e12 / peaks["bf16_tflops"], 4),
                          "hbm_GBs": round(qps * hbm / 1e9, 1), "hbm_frac": round(qps * hbm / 1e9 / peaks["hbm_gbs"], 4),
                          "l2_gather_GBs": round(qps * gathered / 1e9, 1)}))


if __name__ == "__main__":
    main()
