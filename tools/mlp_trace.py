#!/usr/bin/env python
"""Protocol timeline of k_query_mlp (CTA 0): clock64 stamps of the MMA issuer and of one epilogue
warp per tile context, via the debug bit 4 of the split query (experiments only).
  python tools/mlp_trace.py [--log2 20]"""
import argparse, importlib, os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--log2", type=int, default=20)
    a = ap.parse_args()
    import torch
    nrb = importlib.import_module("neural-renderer_b200")
    r = nrb.RendererNeural(nrb.Scene())
    r.load_weights(nrb.scenes.make_weights())
    r.set_query_kernel(3 | (16 << 8))
    n = 1 << a.log2
    x = torch.rand((n, r.input_dims()), device="cuda")
    out = torch.zeros((n + 4096, 16), dtype=torch.float16, device="cuda")
    fn = r.lib.fn("neural_query")
    for _ in range(2):
        r.lib.check(fn(r.handle, x.data_ptr(), n, out.data_ptr()), "neural_query")
    r.sync()
    t = out[n:].cpu().numpy().view(np.int64).reshape(-1)[: 3 * 64 * 4].reshape(3, 64, 4)
    t0 = t[0, 0, 0]
    print("issuer: step  wait_start  ready  issued   (cycles since first step)")
    for s in range(24):
        if t[0, s, 0] == 0: continue
        print(f"  s={s:2d} ctx={s&1} {t[0,s,0]-t0:8d} {t[0,s,1]-t0:8d} {t[0,s,2]-t0:8d}   wait={t[0,s,1]-t[0,s,0]:6d} issue={t[0,s,2]-t[0,s,1]:5d}")
    for c in (0, 1):
        print(f"epilogue ctx {c}: ev  wait_start  mma_done  ld_done  arrived")
        for e in range(10):
            if t[1 + c, e, 1] == 0: continue
            v = t[1 + c, e] - t0
            print(f"  ev={e:2d} {v[0]:8d} {v[1]:8d} {v[2]:8d} {v[3]:8d}   wait={v[1]-v[0]:6d} ld={v[2]-v[1]:5d} pack+sts+fence={v[3]-v[2]:5d}")


if __name__ == "__main__":
    main()
