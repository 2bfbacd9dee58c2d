"""The neural query pinned to tiny-cuda-nn ITSELF.

tests/golden/tcnn_query_v1.npz holds outputs of tiny-cuda-nn's own
NetworkWithInputEncoding<__half>::inference_mixed_precision (ext/tiny-cuda-nn object.h:166-209) on a B200, configured
and loaded exactly as the reference renderer does it (src/cuda_renderer_neural.cu:1647-1688, :1752-1808, :2024-2031);
generator: tests/golden/make_tcnn_golden.py over baseline/_ref/tcnn_query.  Asserted against them:
  * oracle/neural_oracle.cpp (CPU, `-m "not gpu"`),
  * every query kernel of libnrb200 (mma.sync, fused tcgen05, split, split chunked; `-m gpu`),
  * and, when the reference build is present on the GPU box, tcnn run LIVE on fresh inputs.

Tolerance.  tiny-cuda-nn accumulates the MLP in fp16 (wmma accumulator __half, src/fully_fused_mlp.cu:68, 333, 436);
libnrb200 accumulates in fp32 and the oracle sums exactly, so the three agree to the rounding noise of tcnn's fp16
sums, not bit for bit.  tcnn's own convention is 1e-2 relative to the signal scale (vector_match_rae,
ext/tiny-cuda-nn/tests/test_common.h:63-113: |a-b| / (mean magnitude + scale * 1e-2)); with seeded random weights the
outputs are zero-mean with RMS 0.06-0.10, so the bars are stated against the RMS of tcnn's outputs:
    99.99 % of |a - b| <= 1e-2 * rms,  max |a - b| <= 2e-2 * rms,  mean |a - b| <= 2e-3 * rms,
    |mean (a - b)| <= 2e-4 * rms (no bias),  out[0] >= 0 (the hit decision, :1040) flips on <= 1e-3 of the rows.
Observed (oracle and kernels alike): max 0.5-1.3 % (650 K - 1.6 M values), mean 0.05-0.09 % of the RMS, flips 1-4 in 10 000.
"""
import os

import numpy as np
import pytest

from tests.golden.make_tcnn_golden import SHAPES, golden_inputs, golden_weights, input_dims
from tests.util import nrb

GOLD = np.load(os.path.join(os.path.dirname(__file__), "golden", "tcnn_query_v1.npz"))


def assert_matches_tcnn(got, want, what):
    a, b = got.astype(np.float64), want.astype(np.float64)
    assert a.shape == b.shape, (what, a.shape, b.shape)
    assert np.isfinite(a).all(), what
    rms = np.sqrt(np.mean(b * b))
    d = np.abs(a - b)
    flips = np.mean((a[:, 0] >= 0) != (b[:, 0] >= 0))
    msg = (what, f"max {d.max() / rms:.2e} mean {d.mean() / rms:.2e} bias {abs((a - b).mean()) / rms:.2e} of rms {rms:.4f}, flips {flips:.2e}")
    assert np.percentile(d, 99.99) <= 1e-2 * rms, msg
    assert d.max() <= 2e-2 * rms, msg
    assert d.mean() <= 2e-3 * rms, msg
    assert abs((a - b).mean()) <= 2e-4 * rms, msg
    assert flips <= 1e-3, msg
    return msg


def _renderer(shape, lib=None):
    net, cfgkw = SHAPES[shape]
    scene = nrb.scenes.make_scene("ico", 1, env=False)
    r = nrb.RendererNeural(scene, nn_config=nrb.NNConfig(**cfgkw) if cfgkw else None, lib=lib)
    if lib is not None:
        lib.fn("set_threads")(r.handle, 0)
    r.load_weights(golden_weights(nrb, net))
    return r, golden_inputs(input_dims(net))


def test_golden_file_is_what_the_generator_describes():
    for shape, (net, _) in SHAPES.items():
        assert GOLD[shape].shape == (512 + 8 + 40000, 16) and GOLD[shape].dtype == np.float16
        assert golden_inputs(input_dims(net)).shape == (40520, input_dims(net))


@pytest.mark.parametrize("shape", list(SHAPES))
def test_oracle_matches_tcnn(oracle_lib, shape):
    """oracle/neural_oracle.cpp (the CPU restatement every other neural test leans on) against tcnn's own outputs."""
    r, x = _renderer(shape, oracle_lib)
    print(assert_matches_tcnn(r.neural_query(x), GOLD[shape], f"oracle/{shape}"))
    r.close()


def test_oracle_weight_file_orders_agree(oracle_lib):
    """[hg | mlp] (swapParamOrder, the default) and [mlp | hg] describe the same network (:1764-1770)."""
    net, _ = SHAPES["ref"]
    w = golden_weights(nrb, net)
    ng = nrb.scenes.network_sizes(**net)[0]
    scene = nrb.scenes.make_scene("ico", 1, env=False)
    x = golden_inputs(input_dims(net))[:600]
    outs = []
    for weights, swap in ((w, True), (np.concatenate([w[ng:], w[:ng]]), False)):
        r = nrb.RendererNeural(scene, lib=oracle_lib)
        r.load_weights(weights, swap_param_order=swap)
        outs.append(r.neural_query(x))
        r.close()
    assert np.array_equal(outs[0].view(np.uint16), outs[1].view(np.uint16))
    assert_matches_tcnn(outs[0], GOLD["ref"][:600], "oracle/ref/[mlp|hg]")


KERNELS = [(1, "mma_sync"), (2, "tcgen05_fused"), (3, "split"), (3 | (8 << 8), "split_chunked")]


@pytest.mark.gpu
@pytest.mark.parametrize("shape", list(SHAPES))
@pytest.mark.parametrize("kernel", [k for k, _ in KERNELS], ids=[n for _, n in KERNELS])
def test_query_kernels_match_tcnn(kernel, shape):
    r, x = _renderer(shape)
    r.set_query_kernel(kernel)
    try:
        print(assert_matches_tcnn(r.neural_query(x), GOLD[shape], f"kernel {kernel}/{shape}"))
        # ragged batch sizes through the same kernels: a prefix of the batch must give the prefix of the outputs
        rms = np.sqrt(np.mean(GOLD[shape].astype(np.float64) ** 2))
        for n in (1, 127, 129, 520):
            y = r.neural_query(x[:n])
            assert np.abs(y.astype(np.float64) - GOLD[shape][:n].astype(np.float64)).max() <= 1e-2 * rms
    finally:
        r.set_query_kernel(0)
        r.close()


@pytest.mark.gpu
def test_weight_file_orders_agree_on_device():
    net, _ = SHAPES["ref"]
    w = golden_weights(nrb, net)
    ng = nrb.scenes.network_sizes(**net)[0]
    scene = nrb.scenes.make_scene("ico", 1, env=False)
    x = golden_inputs(input_dims(net))[:2000]
    outs = []
    for weights, swap in ((w, True), (np.concatenate([w[ng:], w[:ng]]), False)):
        r = nrb.RendererNeural(scene)
        r.load_weights(weights, swap_param_order=swap)
        outs.append(r.neural_query(x))
        r.close()
    assert np.array_equal(outs[0].view(np.uint16), outs[1].view(np.uint16))


@pytest.mark.gpu
@pytest.mark.parametrize("shape", ["ref", "c5w64"])
def test_query_matches_live_tcnn(shape):
    """tiny-cuda-nn run NOW on this GPU (baseline/_ref/tcnn_query, the reference's own code compiled for sm_100) on
    inputs and weights no fixture has seen, against the default query path."""
    from baseline import refpack
    if not refpack.have("tcnn_query"):
        pytest.skip("baseline/_ref/tcnn_query not built (needs /root/reference at build time)")
    net, cfgkw = SHAPES[shape]
    w = nrb.scenes.make_weights(seed=2025, **net)
    x = np.random.RandomState(31337).uniform(0, 1, (100000, input_dims(net))).astype(np.float32)
    want, _ = refpack.tcnn_query(net, w, x)
    scene = nrb.scenes.make_scene("ico", 1, env=False)
    r = nrb.RendererNeural(scene, nn_config=nrb.NNConfig(**cfgkw) if cfgkw else None)
    r.load_weights(w)
    print(assert_matches_tcnn(r.neural_query(x), want, f"live/{shape}"))
    r.close()
