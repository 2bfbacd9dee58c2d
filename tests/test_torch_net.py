"""The PyTorch twin of the shell network and its weight-file exporter (neural-renderer_b200/torch_net.py)
against the CPU oracle (not gpu) and the CUDA kernels (gpu)."""
import importlib

import numpy as np
import pytest
import torch

from tests.util import nrb, rae_ok

tn = importlib.import_module("neural-renderer_b200.torch_net")


def _inputs(n, dims, seed=3):
    rng = np.random.RandomState(seed)
    return rng.uniform(0.02, 0.98, (n, dims)).astype(np.float32)


def _renderer(lib, net, halfs):
    r = nrb.RendererNeural(nrb.scenes.make_scene("sphere_small", 0), nn_config=net.nn_config(), lib=lib)
    r.load_weights(halfs)
    return r


@pytest.mark.parametrize("shape", [dict(), dict(log2_hashmap_size=12, base_resolution=4, n_points=2, n_levels=5,
                                                n_features=2, mlp_width=64, n_hidden=2)], ids=["reference", "small"])
def test_torch_twin_matches_oracle(oracle_lib, shape):
    net = tn.ShellNetwork(**shape)
    halfs = nrb.scenes.make_weights(seed=11, **shape)
    net.import_fp16(halfs)
    assert np.array_equal(net.export_fp16().view(np.uint16), halfs.view(np.uint16))       # file round trip
    swapped = net.export_fp16(swap_param_order=False)
    assert np.array_equal(tn.ShellNetwork(**shape).import_fp16(swapped, False).export_fp16().view(np.uint16),
                          halfs.view(np.uint16))
    x = _inputs(257, net.input_dims)
    with torch.no_grad():
        enc = net.encode(torch.from_numpy(x)).numpy()
        y = net(torch.from_numpy(x)).numpy()
    assert enc.shape == (257, net.encoded_dims)
    orc = _renderer(oracle_lib, net, halfs)
    want = orc.neural_query(x).astype(np.float32)
    err = rae_ok(y, want, floor=5e-2)
    assert np.percentile(err, 99) < 2e-2 and err.max() < 0.1, (np.percentile(err, 99), err.max())


def test_fine_tune_and_export(oracle_lib, tmp_path):
    """A few optimiser steps on the twin move the exported file's behaviour in the renderer's arithmetic."""
    shape = dict(log2_hashmap_size=10, base_resolution=4, n_points=2, n_levels=4, n_features=2, mlp_width=64, n_hidden=2)
    net = tn.ShellNetwork(seed=5, **shape)
    x = torch.from_numpy(_inputs(512, net.input_dims, seed=8))
    target = torch.zeros(512, 16)
    target[:, 0] = torch.sin(6.0 * x[:, 0]) * torch.cos(5.0 * x[:, 1])
    target[:, 1] = x[:, 2]
    opt = torch.optim.Adam(net.parameters(), lr=1e-2)
    first = None
    for _ in range(150):
        opt.zero_grad()
        loss = torch.nn.functional.mse_loss(net(x)[:, :2], target[:, :2])
        first = first if first is not None else float(loss.detach())
        loss.backward()
        opt.step()
    assert float(loss.detach()) < 0.25 * first
    path = tmp_path / "tuned.bin"
    net.save(str(path))
    assert path.stat().st_size == 2 * net.file_halfs()
    orc = nrb.RendererNeural(nrb.scenes.make_scene("sphere_small", 0), nn_config=net.nn_config(), lib=oracle_lib)
    orc.load_weights(np.fromfile(str(path), np.float16))
    got = orc.neural_query(x.numpy()).astype(np.float32)[:, :2]
    assert np.mean((got - target[:, :2].numpy()) ** 2) < 0.3 * first


@pytest.mark.gpu
def test_torch_twin_matches_cuda():
    net = tn.ShellNetwork()
    halfs = nrb.scenes.make_weights(seed=12)
    net.import_fp16(halfs)
    x = _inputs(4096, net.input_dims, seed=9)
    with torch.no_grad():
        y = net(torch.from_numpy(x)).numpy()
    prod = _renderer(None, net, net.export_fp16())
    err = rae_ok(y, prod.neural_query(x).astype(np.float32), floor=5e-2)
    assert np.percentile(err, 99) < 2e-2 and err.max() < 0.1, (np.percentile(err, 99), err.max())
