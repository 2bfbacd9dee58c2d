"""Training-side twin of the network libnrb200 evaluates.

The reference trains its shell network elsewhere (`models.py`, not in the tree, PROJECT.md:130-153) and ships
weights as one raw fp16 file, `[hash grid | MLP]` by default or `[MLP | hash grid]`
(RendererNeural::loadWeightsFromFile, src/cuda_renderer_neural.cu:1752-1808).  `ShellNetwork` is a plain,
differentiable PyTorch module of the same function - n_points x HashGrid(L levels, F features, one shared
table) + spherical harmonics of the direction -> bias-free MLP with LeakyReLU(0.01) and a 16-wide linear head -
with the parameter layout of that file, so a fine-tuned network is exported with `save()` and loaded by
`nrb_load_weights_file` unchanged, and a weight file is imported with `load()`.

It computes in fp32 (the renderer in fp16 with fp32 accumulation): outputs agree to fp16 tolerance, not bit for
bit.  This is tooling beside the render path: nothing in libnrb200 or in the frame loop uses PyTorch.

Layout facts restated here (checked against the CPU oracle and the CUDA kernels in tests/test_torch_net.py):
  * level l: scale = 2^(l*log2(per_level_scale)) * base - 1, resolution = ceil(scale) + 1, entries =
    min(roundup8(res^3), 2^log2_hashmap_size); levels are concatenated in one table (tiny-cuda-nn grid.h:699-730);
  * a corner (x, y, z) of level l addresses entry (x + y*res + z*res^2) when the level is dense, else
    (x ^ y*2654435761 ^ z*805459861) mod entries in uint32 arithmetic (common_device.h:785-884);
  * position within the cell: v = scale*x + 0.5, cell = floor(v), weight = v - cell (pos_fract, :1017-1029);
  * encoded row order: point 0 levels 0..L-1 (F features each), point 1, ..., then the SH block
    (composite.h:215-258); SH degree 4 on 2*d - 1 (common_device.h:476-502);
  * MLP matrices row-major [out][in], W0 [width][encoded], hidden [width][width], head [16][width]
    (fully_fused_mlp.cu:659-671).
"""
import numpy as np
import torch

_P1, _P2 = 2654435761, 805459861


def _levels(log2_hashmap_size, base_resolution, n_levels, per_level_scale):
    out, offset = [], 0
    log2s = np.float32(np.log2(np.float32(per_level_scale)))
    for l in range(n_levels):
        scale = np.float32(np.exp2(np.float32(l) * log2s)) * np.float32(base_resolution) - np.float32(1.0)
        res = int(np.ceil(scale)) + 1
        n = min(res ** 3, 0xFFFFFFFF // 2)
        n = (n + 7) // 8 * 8
        n = min(n, 1 << log2_hashmap_size)
        dense = res <= 0x659 and n >= res ** 3
        out.append(dict(scale=float(scale), res=res, entries=n, offset=offset, dense=dense))
        offset += n
    return out, offset


class ShellNetwork(torch.nn.Module):
    def __init__(self, log2_hashmap_size=14, base_resolution=16, n_points=3, n_levels=8, n_features=4,
                 per_level_scale=2.0, sh_degree=4, mlp_width=128, n_hidden=4, seed=None):
        super().__init__()
        self.cfg = dict(log2_hashmap_size=log2_hashmap_size, base_resolution=base_resolution, n_points=n_points,
                        n_levels=n_levels, n_features=n_features, per_level_scale=per_level_scale,
                        sh_degree=sh_degree, mlp_width=mlp_width, n_hidden=n_hidden)
        self.levels, entries = _levels(log2_hashmap_size, base_resolution, n_levels, per_level_scale)
        self.n_points, self.n_features, self.sh_degree = n_points, n_features, sh_degree
        self.input_dims = 3 * n_points + (3 if sh_degree > 0 else 0)
        self.encoded_dims = n_points * n_levels * n_features + sh_degree * sh_degree
        gen = torch.Generator().manual_seed(seed) if seed is not None else None
        # tiny-cuda-nn's initial values: grid U(-1e-4, 1e-4), Xavier-uniform matrices
        self.grid = torch.nn.Parameter((torch.rand(entries, n_features, generator=gen) * 2 - 1) * 1e-4)
        shapes = [(mlp_width, self.encoded_dims)] + [(mlp_width, mlp_width)] * (n_hidden - 1) + [(16, mlp_width)]
        self.weights = torch.nn.ParameterList()
        for o, i in shapes:
            lim = (6.0 / (o + i)) ** 0.5
            self.weights.append(torch.nn.Parameter((torch.rand(o, i, generator=gen) * 2 - 1) * lim))

    # ---- the function -------------------------------------------------------------------------------
    def _level(self, x, lv):
        """x [N, 3] in [0, 1] -> [N, F]: trilinear interpolation of the level's table."""
        v = x * lv["scale"] + 0.5
        cell = torch.floor(v)
        w = v - cell
        cell = cell.to(torch.int64)
        table = self.grid[lv["offset"]:lv["offset"] + lv["entries"]]
        out = 0
        for corner in range(8):
            c, wt = [], 1.0
            for d in range(3):
                hi = (corner >> d) & 1
                c.append(cell[:, d] + hi)
                wt = wt * (w[:, d] if hi else 1.0 - w[:, d])
            if lv["dense"]:
                idx = (c[0] + c[1] * lv["res"] + c[2] * lv["res"] * lv["res"]) & 0xFFFFFFFF
            else:
                m = 0xFFFFFFFF
                idx = (c[0] & m) ^ ((c[1] * _P1) & m) ^ ((c[2] * _P2) & m)
            idx = idx % lv["entries"]
            out = out + wt[:, None] * table[idx]
        return out

    def _sh(self, d):
        x, y, z = (d * 2.0 - 1.0).unbind(-1)
        xy, xz, yz, x2, y2, z2 = x * y, x * z, y * z, x * x, y * y, z * z
        o = [torch.full_like(x, 0.28209479177387814)]
        if self.sh_degree > 1:
            o += [-0.48860251190291987 * y, 0.48860251190291987 * z, -0.48860251190291987 * x]
        if self.sh_degree > 2:
            o += [1.0925484305920792 * xy, -1.0925484305920792 * yz, 0.94617469575755997 * z2 - 0.31539156525251999,
                  -1.0925484305920792 * xz, 0.54627421529603959 * x2 - 0.54627421529603959 * y2]
        if self.sh_degree > 3:
            o += [0.59004358992664352 * y * (-3.0 * x2 + y2), 2.8906114426405538 * xy * z,
                  0.45704579946446572 * y * (1.0 - 5.0 * z2), 0.3731763325901154 * z * (5.0 * z2 - 3.0),
                  0.45704579946446572 * x * (1.0 - 5.0 * z2), 1.4453057213202769 * z * (x2 - y2),
                  0.59004358992664352 * x * (-x2 + 3.0 * y2)]
        return torch.stack(o[:self.sh_degree * self.sh_degree], -1)

    def encode(self, inputs):
        """[N, input_dims] -> [N, encoded_dims] in the renderer's row order."""
        cols = []
        for p in range(self.n_points):
            x = inputs[:, 3 * p:3 * p + 3]
            cols += [self._level(x, lv) for lv in self.levels]
        if self.sh_degree > 0:
            cols.append(self._sh(inputs[:, 3 * self.n_points:3 * self.n_points + 3]))
        return torch.cat(cols, -1)

    def forward(self, inputs):
        """[N, input_dims] fp32 -> [N, 16]: hit logit, distance, normal.xyz, colour.rgb, padding
        (decoded by applySegmentNeuralOutputKernel, cuda_renderer_neural.cu:1007-1123)."""
        h = self.encode(inputs)
        for w in self.weights[:-1]:
            h = torch.nn.functional.leaky_relu(h @ w.t(), 0.01)
        return h @ self.weights[-1].t()

    # ---- the weight file ------------------------------------------------------------------------------
    def file_halfs(self):
        return self.grid.numel() + sum(w.numel() for w in self.weights)

    def export_fp16(self, swap_param_order=True):
        """The raw fp16 weight file as a numpy array: [hg | mlp] (the reference's default) or [mlp | hg]."""
        with torch.no_grad():
            hg = self.grid.detach().reshape(-1).cpu().numpy().astype(np.float16)
            mlp = np.concatenate([w.detach().reshape(-1).cpu().numpy() for w in self.weights]).astype(np.float16)
        return np.concatenate([hg, mlp] if swap_param_order else [mlp, hg])

    def import_fp16(self, halfs, swap_param_order=True):
        halfs = np.asarray(halfs)
        if halfs.dtype != np.float16:
            halfs = halfs.view(np.float16)
        if halfs.size != self.file_halfs():
            raise ValueError(f"weight file has {halfs.size} halfs, this network needs {self.file_halfs()}")
        ng = self.grid.numel()
        nm = halfs.size - ng
        hg, mlp = (halfs[:ng], halfs[ng:]) if swap_param_order else (halfs[nm:], halfs[:nm])
        with torch.no_grad():
            self.grid.copy_(torch.from_numpy(hg.astype(np.float32)).reshape(self.grid.shape))
            o = 0
            for w in self.weights:
                w.copy_(torch.from_numpy(mlp[o:o + w.numel()].astype(np.float32)).reshape(w.shape))
                o += w.numel()
        return self

    def save(self, path, swap_param_order=True):
        self.export_fp16(swap_param_order).tofile(path)

    def load(self, path, swap_param_order=True):
        return self.import_fp16(np.fromfile(path, np.float16), swap_param_order)

    def nn_config(self):
        """The matching nrb_nn_config for RendererNeural(scene, nn_config=...)."""
        from . import _capi as K
        c = self.cfg
        return K.NNConfig(c["log2_hashmap_size"], c["base_resolution"], c["n_points"], c["n_levels"], c["n_features"],
                          c["per_level_scale"], c["sh_degree"] if c["sh_degree"] > 0 else -1, c["mlp_width"], c["n_hidden"])
