"""nrb_neural_query_static (include/nrb200.h): the neural query with its scratch in a static __device__ array, for
callers that cannot call cudaMalloc.  No counterpart in the reference.  Collected last.

Checked through the C ABI on device pointers: two calls in a row agree with nrb_neural_query (the same two kernels on
the renderer's cudaMalloc'ed scratch, so the bar is bit equality, inside the 1e-2 relative error of the other query
tests), the driver's free-memory figure does not move across the calls, and sizes beyond the maximum are refused.
"""
import numpy as np
import pytest

from tests.util import nrb, rae_ok

pytestmark = pytest.mark.gpu

STATIC_MAX = 16384  # NRB_STATIC_QUERY_MAX


def _query(r, name, x_dev, n, out_dev):
    lib = r.lib
    lib.check(lib.fn(name)(r.handle, x_dev.data_ptr(), n, out_dev.data_ptr()), name)
    r.sync()


def test_static_scratch_query_matches_allocating_entry_point():
    import torch  # device memory for the caller's inputs and outputs only
    scene = nrb.scenes.make_scene("ico", 1, env=False)
    r = nrb.RendererNeural(scene)
    r.load_weights(nrb.scenes.make_weights(seed=42))
    D = r.input_dims()
    rng = np.random.RandomState(123)
    dev = torch.device("cuda:0")
    # the second call is smaller than the first and has a ragged last tile: stale tiles of call 1 stay in the array
    sizes = (STATIC_MAX, 5000 + 37)
    xs = [torch.from_numpy(rng.uniform(0, 1, (n, D)).astype(np.float32)).to(dev) for n in sizes]
    want = [torch.zeros((n, 16), dtype=torch.float16, device=dev) for n in sizes]
    got = [torch.full((n, 16), -7.0, dtype=torch.float16, device=dev) for n in sizes]
    for x, w in zip(xs, want):
        _query(r, "neural_query", x, len(x), w)  # existing entry point (allocates its scratch on first use)
    torch.cuda.synchronize()
    free0 = torch.cuda.mem_get_info()[0]
    for x, g in zip(xs, got):
        _query(r, "neural_query_static", x, len(x), g)
    torch.cuda.synchronize()
    assert torch.cuda.mem_get_info()[0] == free0, "the static-scratch call changed the device's free memory"
    for n, w, g in zip(sizes, want, got):
        a, b = g.cpu().numpy(), w.cpu().numpy()
        assert np.isfinite(a.astype(np.float32)).all()
        err = rae_ok(a, b)
        assert np.percentile(err, 99) < 1e-2 and err.max() < 5e-2, (n, err.max())
        assert np.array_equal(a.view(np.uint16), b.view(np.uint16)), (n, err.max())  # same kernels: same bits
    assert np.abs(want[0].cpu().numpy().astype(np.float32)).max() > 0

    # beyond the maximum: refused with a message, nothing launched
    big = torch.zeros((STATIC_MAX + 1, D), dtype=torch.float32, device=dev)
    out = torch.zeros((STATIC_MAX + 1, 16), dtype=torch.float16, device=dev)
    rc = r.lib.fn("neural_query_static")(r.handle, big.data_ptr(), STATIC_MAX + 1, out.data_ptr())
    assert rc != 0
    msg = r.lib.fn("last_error")().decode()
    assert "16384" in msg, msg
    # an empty batch is a no-op
    assert r.lib.fn("neural_query_static")(r.handle, big.data_ptr(), 0, out.data_ptr()) == 0
    # and the renderer still answers afterwards
    _query(r, "neural_query_static", xs[1], len(xs[1]), got[1])
    assert np.array_equal(got[1].cpu().numpy().view(np.uint16), want[1].cpu().numpy().view(np.uint16))
