"""GPU parity tests: libnrb200 (CUDA, through the C ABI) against the CPU oracle on the same inputs.

Bars (BASELINE.json north_star): hit triangle ids bit-exact; t within 1e-5 relative (we observe
bit-exact t/u/v because the kernels evaluate the reference's expressions without FMA contraction);
neural outputs within fp16 tolerance (tiny-cuda-nn's own 1e-2 relative-error convention); images
within 0.1 dB-equivalent PSNR floors stated per test.
"""
import numpy as np
import pytest

from tests.util import make_pair, nrb, psnr8, rae_ok, random_rays

pytestmark = pytest.mark.gpu

K = nrb._capi if hasattr(nrb, "_capi") else None


def _trace_both(prod, orc, slot, mode, o, d):
    return prod.trace_rays(slot, mode, o, d), orc.trace_rays(slot, mode, o, d)


@pytest.mark.parametrize("kind,sub", [("sphere_small", 0), ("ico", 2), ("ico", 5)])
def test_trace_rays_bit_exact(oracle_lib, kind, sub):
    scene = nrb.scenes.make_scene(kind, sub, additional=True, env=False)
    prod, orc, eye = make_pair(scene, oracle_lib, 8, 8)
    o, d = random_rays(20000, seed=11 + sub)
    for slot in (nrb.SLOT_ORIGINAL, nrb.SLOT_OUTER, nrb.SLOT_INNER, nrb.SLOT_ADDITIONAL):
        for mode in (nrb.TRACE_ANY, nrb.TRACE_FORWARD_ONLY, nrb.TRACE_BACKWARD_ONLY):
            a, b = _trace_both(prod, orc, slot, mode, o, d)
            hit = b["t"] < 1e30
            assert np.array_equal(a["t"] < 1e30, hit), (slot, mode)
            assert np.array_equal(a["prim"][hit], b["prim"][hit]), (slot, mode)
            assert np.array_equal(a["t"], b["t"]), (slot, mode)      # stronger than the 1e-5 bar
            assert np.array_equal(a["u"][hit], b["u"][hit]) and np.array_equal(a["v"][hit], b["v"][hit])
    assert hit.sum() > 0


def test_trace_edge_cases(oracle_lib):
    scene = nrb.scenes.make_scene("ico", 1, additional=True, env=False)
    prod, orc, eye = make_pair(scene, oracle_lib, 8, 8)
    # empty batch, axis-parallel directions (1/0 = inf slabs), rays starting on a vertex / inside the mesh
    out = prod.trace_rays(nrb.SLOT_ORIGINAL, nrb.TRACE_ANY, np.zeros((0, 3), np.float32), np.zeros((0, 3), np.float32))
    assert len(out) == 0
    v = scene.original_mesh.vertices
    o = np.concatenate([np.array([[0, 0, 3], [0, 3, 0], [3, 0, 0], [0, 0, 0], [0, 0, 0]], np.float32), v[:16] * 2])
    d = np.concatenate([np.array([[0, 0, -1], [0, -1, 0], [-1, 0, 0], [1, 0, 0], [0, 1, 0]], np.float32),
                        -v[:16] / np.linalg.norm(v[:16], axis=1, keepdims=True)]).astype(np.float32)
    for mode in (0, 1, 2):
        a, b = _trace_both(prod, orc, nrb.SLOT_ORIGINAL, mode, o, d)
        assert np.array_equal(a["t"], b["t"]) and np.array_equal(a["prim"], b["prim"])
    # an empty slot misses everything
    prod2, orc2, _ = make_pair(nrb.scenes.make_scene("ico", 1, shells=False, env=False), oracle_lib, 8, 8)
    a = prod2.trace_rays(nrb.SLOT_OUTER, 0, o, d)
    assert np.all(a["t"] >= 1e30)


def _weights(**net):
    return nrb.scenes.make_weights(seed=42, **net)


@pytest.mark.parametrize("kernel", [1, 2, 3, 3 | (8 << 8)], ids=["mma_sync", "tcgen05_fused", "split", "split_chunked"])
def test_neural_query_matches_oracle(oracle_lib, kernel):
    scene = nrb.scenes.make_scene("ico", 1, env=False)
    w = _weights()
    prod, orc, _ = make_pair(scene, oracle_lib, 8, 8, weights=w)
    prod.set_query_kernel(kernel)
    rng = np.random.RandomState(99)
    for n in (1, 15, 16, 127, 129, 257, 4096, 40000):
        x = rng.uniform(0, 1, (n, prod.input_dims())).astype(np.float32)
        a, b = prod.neural_query(x), orc.neural_query(x)
        err = rae_ok(a[:, :8], b[:, :8])
        # fp16 tolerance: fp32-accumulated tensor-core sums vs the oracle's exact sums, each
        # rounded to fp16 per layer -> a few fp16 ulps after 5 layers
        assert np.percentile(err, 99) < 1e-2, (n, np.percentile(err, 99))
        assert err.max() < 5e-2, (n, err.max())
    prod.set_query_kernel(0)


@pytest.mark.parametrize("kernel", [1, 2, 3, 3 | (8 << 8)], ids=["mma_sync", "tcgen05_fused", "split", "split_chunked"])
def test_neural_query_other_shape(oracle_lib, kernel):
    """BASELINE config 5 shape: 1 x HashGrid(16 levels, 2 features, T=2^15, scale 1.5) -> 32 -> 64 x 2 -> 16."""
    cfg = nrb.NNConfig(log2_hashmap_size=15, base_resolution=16, n_points=1, n_levels=16, n_features_per_level=2,
                       per_level_scale=1.5, sh_degree=-1, mlp_width=64, n_hidden_layers=2)
    net = dict(log2_hashmap_size=15, base_resolution=16, n_points=1, n_levels=16, n_features=2, per_level_scale=1.5,
               sh_degree=0, mlp_width=64, n_hidden=2)
    scene = nrb.scenes.make_scene("ico", 1, env=False)
    prod, orc, _ = make_pair(scene, oracle_lib, 8, 8, weights=_weights(**net), nn_config=cfg)
    prod.set_query_kernel(kernel)
    x = np.random.RandomState(5).uniform(0, 1, (1000, prod.input_dims())).astype(np.float32)
    err = rae_ok(prod.neural_query(x), orc.neural_query(x))
    prod.set_query_kernel(0)
    assert np.percentile(err, 99) < 1e-2 and err.max() < 5e-2


@pytest.mark.parametrize("spp,bounces,textured", [(1, 0, False), (2, 2, True)])
def test_gt_frame_parity(oracle_lib, spp, bounces, textured):
    scene = nrb.scenes.make_scene("ico", 4, additional=False, textured=textured)
    prod, orc, eye = make_pair(scene, oracle_lib, 160, 96, spp=spp, bounces=bounces)
    prod.render(eye)
    orc.render(eye)
    # primary hit buffers: bit-exact (ids decide positions; positions are barycentric sums)
    for name in ("hit_flags", "hit_positions", "hit_normals", "hit_colors", "hit_material_params"):
        a, b = prod.read_buffer(name), orc.read_buffer(name)
        assert np.array_equal(a, b), name
    pa, pb = prod.read_pixels(), orc.read_pixels()
    assert psnr8(pa, pb) > 45.0, psnr8(pa, pb)
    assert np.abs(pa.astype(int) - pb.astype(int)).max() <= (1 if bounces == 0 else 255)
    ca, cb = prod.counters(), orc.counters()
    if bounces == 0:
        assert ca["ray_segments"] == cb["ray_segments"]
    else:
        assert abs(ca["ray_segments"] - cb["ray_segments"]) <= 0.002 * cb["ray_segments"]


def test_gt_accumulation_and_lambert(oracle_lib):
    scene = nrb.scenes.make_scene("ico", 3)
    prod, orc, eye = make_pair(scene, oracle_lib, 96, 64, spp=1, bounces=1)
    for _ in range(3):
        prod.render(eye)
        orc.render(eye)
    assert prod.lib.fn("accum_sample_count")(prod.handle) == 3
    assert psnr8(prod.read_pixels(), orc.read_pixels()) > 45.0
    for r in (prod, orc):
        r.set_lambert_view(True)
        r.render(eye)
    assert psnr8(prod.read_pixels(), orc.read_pixels()) > 50.0
    assert prod.lib.fn("accum_sample_count")(prod.handle) == 0


@pytest.mark.parametrize("spp,bounces,additional", [(1, 0, False), (1, 2, True)])
def test_neural_frame_parity(oracle_lib, spp, bounces, additional):
    scene = nrb.scenes.make_scene("ico", 3, additional=additional)
    prod, orc, eye = make_pair(scene, oracle_lib, 128, 80, spp=spp, bounces=bounces, neural=True, weights=_weights())
    prod.render(eye)
    orc.render(eye)
    fa, fb = prod.read_buffer("hit_flags"), orc.read_buffer("hit_flags")
    # the hit decision is out[0] >= 0 on an fp16 logit: near-zero logits may flip
    assert np.mean(fa != fb) < 2e-3, np.mean(fa != fb)
    same = (fa == fb) & (fa != 0)
    pa, pb = prod.read_buffer("hit_positions"), orc.read_buffer("hit_positions")
    assert np.percentile(np.abs(pa[same] - pb[same]), 99) < 2e-3
    ia, ib = prod.read_pixels(), orc.read_pixels()
    assert psnr8(ia, ib) > (35.0 if bounces == 0 else 30.0), psnr8(ia, ib)
    ca, cb = prod.counters(), orc.counters()
    assert abs(ca["neural_queries"] - cb["neural_queries"]) <= 0.01 * cb["neural_queries"] + 2


@pytest.mark.parametrize("neural", [False, True], ids=["gt", "neural"])
def test_benchmarked_configuration_parity(oracle_lib, neural):
    """The configuration bench.py times (BASELINE configs 2 and 3): 1920x1080, spp 1, 3 bounces, displaced icosphere of
    327 680 triangles in its +-12 % shells, 2048x1024 environment, reference network - against the oracle on all host
    cores (~10 s per frame).  Ground truth: primary hit buffers bit-exact, frame within one step of 255 on >= 99.99 % of
    the pixels.  Neural: north_star's bars - mean FLIP <= 0.005 - and PSNR >= 45 dB."""
    scene = nrb.scenes.make_scene("ico", 7, shells=True, env=True, env_size=(2048, 1024))
    prod, orc, eye = make_pair(scene, oracle_lib, 1920, 1080, spp=1, bounces=3, neural=neural,
                               weights=_weights() if neural else None)
    prod.render(eye)
    orc.render(eye)
    ia, ib = prod.read_pixels(), orc.read_pixels()
    d = np.abs(ia.astype(int) - ib.astype(int)).max(axis=2)
    p, flip = psnr8(ia, ib), float(nrb.metrics.flip_ldr(ib, ia))
    fa, fb = prod.read_buffer("hit_flags"), orc.read_buffer("hit_flags")
    print(f"1080p {'neural' if neural else 'gt'}: psnr {p:.2f} dB, flip {flip:.5f}, within one step {np.mean(d <= 1):.6f}, "
          f"hit-flag mismatch {np.mean(fa != fb):.2e}")
    if neural:
        assert np.mean(fa != fb) < 1e-3
        assert p >= 45.0 and flip <= 0.005, (p, flip)
        ca, cb = prod.counters(), orc.counters()
        assert abs(ca["neural_queries"] - cb["neural_queries"]) <= 0.002 * cb["neural_queries"]
    else:
        for name in ("hit_flags", "hit_positions", "hit_normals", "hit_colors", "hit_material_params"):
            assert np.array_equal(prod.read_buffer(name), orc.read_buffer(name)), name
        assert p >= 60.0 and np.mean(d <= 1) >= 0.9999 and flip <= 0.001, (p, np.mean(d <= 1), flip)


def test_segment_tail_graph_matches_fixed_loop(oracle_lib):
    """Two shell components in line with the camera: rays that pass the first one re-enter the shell, so
    iterations >= 1 of the segment loop carry rays.  The device-driven WHILE graph must give the
    same frame, bit for bit, as the nine fixed launches, and both must match the oracle's host loop."""
    scene = nrb.scenes.make_scene("blobs", 3, additional=True, shell_delta=0.3)
    prod, orc, eye = make_pair(scene, oracle_lib, 160, 96, spp=1, bounces=2, neural=True, weights=_weights())
    frames, queries = [], []
    for mode in (1, 0, 1, 2):  # WHILE graph, fixed loop, graph again, fused tail kernel
        prod.set_tail_graph(mode)
        prod.reset_samples()
        prod.render(eye)
        frames.append(prod.read_pixels().copy())
        queries.append(prod.counters()["neural_queries"])
    assert np.array_equal(frames[0], frames[1]) and np.array_equal(frames[0], frames[2])
    assert queries[0] == queries[1] == queries[2]
    # the fused tail evaluates the network with the mma.sync tile (fp32 sums in another order than the tcgen05
    # kernels before the fp16 rounding): the same rays, outputs equal to fp16 tolerance
    assert abs(queries[3] - queries[0]) <= 2
    d = np.abs(frames[3].astype(int) - frames[0].astype(int))
    assert psnr8(frames[3], frames[0]) > 45.0 and (d > 2).mean() < 0.002, (psnr8(frames[3], frames[0]), (d > 2).mean())
    orc.render(eye)
    oq = orc.counters()["neural_queries"]
    print("tail test: queries", queries[0], "oracle", oq)
    assert abs(queries[0] - oq) <= 0.02 * oq + 2
    assert psnr8(frames[0], orc.read_pixels()) > 28.0, psnr8(frames[0], orc.read_pixels())


@pytest.mark.parametrize("kind,delta", [("ico", 0.12), ("blobs", 0.3)])
def test_decided_bounce_traces_are_skipped_without_changing_the_frame(oracle_lib, kind, delta):
    """The bounce entry does not run traces whose outcome is decided (inner-shell trace beyond the outer-shell hit,
    re-entry trace on a ray without any front face): same frame, radiance and counters as the literal sequence
    (NRB_OPT_LITERAL_BOUNCE_TRACES), on a convex-ish shell and on two separate blobs whose rays do re-enter."""
    scene = nrb.scenes.make_scene(kind, 4, additional=True, shell_delta=delta)
    prod, _, eye = make_pair(scene, oracle_lib, 192, 128, spp=2, bounces=3, neural=True, weights=_weights())
    prod.set_tail_graph(0)  # fixed loop: bit-reproducible network arithmetic
    out = []
    # (literal, entry kernel): NRB_OPT_BOUNCE_ENTRY 0 = separate pruned walks (the default), 1 = the fused walk under
    # the voting scheduler; two independent implementations of the same stage
    runs = ((True, 0), (False, 0), (True, 0), (True, 1), (False, 1))
    for literal, entry in runs:
        prod.set_literal_bounce_traces(literal)
        prod.set_bounce_entry(entry)
        prod.reset_samples()
        prod.render(eye)
        out.append((prod.read_pixels().copy(), prod.read_buffer("path_radiance").copy(), prod.counters()))
    prod.set_literal_bounce_traces(False)
    prod.set_bounce_entry(0)
    assert np.array_equal(out[0][0], out[2][0])  # the literal run is reproducible
    for k in (1, 3, 4):
        assert np.array_equal(out[0][0], out[k][0]), runs[k]
        assert np.array_equal(out[0][1], out[k][1]), runs[k]
        for key in ("neural_queries", "path_vertices", "ray_segments"):
            assert out[0][2][key] == out[k][2][key], (runs[k], key, out[0][2][key], out[k][2][key])
    assert out[0][2]["path_vertices"] > 1000


def test_tiles_reassemble_full_frame(oracle_lib):
    """Multi-GPU partition (row bands, global RNG streams): the union of ranks' rows is the 1-GPU image bit for bit."""
    scene = nrb.scenes.make_scene("ico", 3)
    eye, basis = nrb.scenes.default_camera()

    def render(rank, world):
        r = nrb.RendererNeural(scene)
        r.resize(96, 70)
        r.set_camera_basis(basis)
        r.set_samples_per_pixel(2)
        r.set_bounce_count(2)
        r.set_tile(8, rank, world)
        r.render(eye)
        return r.read_pixels()

    full = render(0, 1)
    parts = [render(k, 3) for k in range(3)]
    merged = np.zeros_like(full)
    for k, p in enumerate(parts):
        for y in range(full.shape[0]):
            if (y // 8) % 3 == k:
                merged[y] = p[y]
    assert np.array_equal(merged, full)


def test_matches_committed_reference_fixtures():
    """libnrb200 against tests/golden/frames_v1.npz (outputs of the reference's own headers)."""
    import os
    from tests.golden.make_golden import render_all
    gold = np.load(os.path.join(os.path.dirname(__file__), "golden", "frames_v1.npz"))
    got = render_all(nrb, None)
    for k in gold.files:
        if k.startswith("trace_"):
            assert np.array_equal(got[k], gold[k]), k
    for k in ("gt_hit_flags", "gt_hit_positions"):
        assert np.array_equal(got[k], gold[k]), k
    assert psnr8(got["gt_pixels"], gold["gt_pixels"]) > 45.0
    rad = np.abs(got["gt_radiance"] - gold["gt_radiance"])
    assert np.mean(rad > 1e-3) < 0.01  # bounce directions go through sinf/cosf/powf (CUDA vs glibc ulps)
    err = rae_ok(got["query"][:, :8], gold["query"][:, :8])
    assert np.percentile(err, 99) < 1e-2 and err.max() < 5e-2
    assert np.mean(got["neural_hit_flags"] != gold["neural_hit_flags"]) < 5e-3
    assert psnr8(got["neural_pixels"], gold["neural_pixels"]) > 30.0


def test_fused_three_mode_traversal_is_exact(oracle_lib):
    """trace_multi (one DFS for ANY + FORWARD + BACKWARD, used by the bounce entry) against three separate
    reference traversals: distances must be bit-identical."""
    scene = nrb.scenes.make_scene("ico", 4, additional=True, env=False)
    prod, orc, eye = make_pair(scene, oracle_lib, 8, 8)
    o, d = random_rays(30000, seed=321)
    o[:10000] *= 0.3  # many origins inside the shells
    for slot in (nrb.SLOT_OUTER, nrb.SLOT_INNER, nrb.SLOT_ORIGINAL):
        fused = prod.trace_rays(slot, 3, o, d)
        union = prod.trace_rays(slot, 4, o, d)  # the resumable engine of k_bounce_begin (prunes by the union of modes)
        for mode, key in ((nrb.TRACE_ANY, "t"), (nrb.TRACE_FORWARD_ONLY, "u"), (nrb.TRACE_BACKWARD_ONLY, "v")):
            ref = orc.trace_rays(slot, mode, o, d)
            assert np.array_equal(fused[key], ref["t"]), (slot, mode)
            # equal except where t, the best t and a box distance agree to the last bits (see nrb_trace.cuh)
            same = union[key] == ref["t"]
            assert same.mean() >= 1.0 - 1e-4, (slot, mode, same.mean())
            assert np.allclose(union[key], ref["t"], rtol=1e-5, atol=0)


@pytest.mark.parametrize("kind,additional", [("ico", True), ("blobs", False)])
def test_path_tail_kernel_matches_per_stage_bounces(oracle_lib, kind, additional):
    """NRB_OPT_PATH_TAIL: bounces >= 2 in ONE launch (a warp carries the surviving paths through sample -> entry ->
    segments -> integrate).  Same paths, same counters; radiance equal to the per-stage kernels up to the fp32
    summation order inside the network (mma.sync tile vs tcgen05), and both against the oracle."""
    scene = nrb.scenes.make_scene(kind, 3, additional=additional, shell_delta=0.3 if kind == "blobs" else 0.12,
                                  material=dict(metallic=0.6, roughness=0.3))  # glossy: more paths survive to bounces 2..4
    prod, orc, eye = make_pair(scene, oracle_lib, 192, 128, spp=2, bounces=4, neural=True, weights=_weights())
    out = []
    for below in (0, 1 << 30, 0):
        prod.set_path_tail(below)
        prod.reset_samples()
        prod.render(eye)
        prod.render(eye)  # the second frame of a setting has the feedback of the first
        out.append((prod.read_pixels().copy(), prod.read_buffer("path_radiance").copy(), prod.counters()))
    prod.set_path_tail(49152)
    assert np.array_equal(out[0][0], out[2][0])
    assert out[1][2]["kernels"] < out[0][2]["kernels"] - 10, (out[0][2], out[1][2])
    for key in ("path_vertices", "neural_queries"):
        assert abs(out[0][2][key] - out[1][2][key]) <= 0.002 * out[0][2][key] + 2, (key, out[0][2], out[1][2])
    assert out[0][2]["path_vertices"] > 2000
    d = np.abs(out[0][0].astype(int) - out[1][0].astype(int)).max(axis=2)
    print(f"path tail vs per-stage: psnr {psnr8(out[0][0], out[1][0]):.2f} dB, pixels differing {np.mean(d > 0):.5f}, counters {out[1][2]}")
    assert psnr8(out[0][0], out[1][0]) > 45.0 and np.mean(d > 1) < 0.005
    orc.render(eye)
    orc.render(eye)
    assert psnr8(out[1][0], orc.read_pixels()) > 30.0


def test_gt_tail_kernel_is_bit_identical(oracle_lib):
    """Ground-truth mode: bounces >= 2 as one launch (k_gt_tail) - same arithmetic per path, so the frame, the radiance
    buffer and the counters equal the per-bounce launches exactly."""
    scene = nrb.scenes.make_scene("ico", 4, additional=False, material=dict(metallic=0.8, roughness=0.25))
    prod, orc, eye = make_pair(scene, oracle_lib, 192, 128, spp=2, bounces=5)
    out = []
    for below in (0, 1 << 30):
        prod.set_path_tail(below)
        prod.reset_samples()
        prod.render(eye)
        prod.render(eye)
        out.append((prod.read_pixels().copy(), prod.read_buffer("path_radiance").copy(), prod.counters()))
    prod.set_path_tail(49152)
    assert np.array_equal(out[0][0], out[1][0]) and np.array_equal(out[0][1], out[1][1])
    assert out[1][2]["kernels"] < out[0][2]["kernels"]
    for key in ("path_vertices", "ray_segments"):
        assert out[0][2][key] == out[1][2][key], (key, out[0][2], out[1][2])
    assert out[0][2]["path_vertices"] > 10000


@pytest.mark.parametrize("neural", [False, True], ids=["gt", "neural"])
def test_frame_slices_give_the_unsliced_frame(neural):
    """NRB_OPT_FRAME_SLICES: the frame as 2 or 3 concurrent row-band slices on as many streams - every pixel is computed
    by the same kernels on the same inputs (global pixel indices and RNG streams), so frames, accumulation and counters
    equal the unsliced renderer's exactly, frame after frame, and also combined with the multi-GPU tile partition."""
    scene = nrb.scenes.make_scene("ico", 4, additional=True)
    eye, basis = nrb.scenes.default_camera()

    def run(slices, tile=None):
        r = nrb.RendererNeural(scene)
        r.resize(256, 208)
        r.set_camera_basis(basis)
        r.set_samples_per_pixel(2)
        r.set_bounce_count(3)
        r.set_path_tail(0)  # the same kernels in every slice whatever its path count
        if neural:
            r.load_weights(_weights())
            r.set_use_neural_query(True)
        if tile:
            r.set_tile(16, *tile)
        r.set_frame_slices(slices)
        frames = []
        for _ in range(3):
            r.render(eye)
            frames.append(r.read_pixels().copy())
        c = r.counters()
        r.close()
        return frames, c

    base, cb = run(1)
    for s in (2, 3):
        got, cg = run(s)
        for a, b in zip(base, got):
            assert np.array_equal(a, b), s
        for key in ("ray_segments", "neural_queries", "path_vertices"):
            assert cb[key] == cg[key], (key, cb, cg)
    # slices inside a rank of a 2-GPU partition: still that rank's rows of the full frame
    part, _ = run(2, tile=(1, 2))
    rows = [y for y in range(208) if (y // 16) % 2 == 1]
    assert np.array_equal(part[2][rows], base[2][rows])


def test_tile_order_is_scheduling_only():
    """NRB_OPT_TILE_ORDER: the ground-truth primary pass takes its tiles most-expensive-first from the second frame of
    a camera on.  Pixels, hit buffers and accumulation are those of the row-major order, frame after frame, also
    across a camera change and with several samples per pixel."""
    scene = nrb.scenes.make_scene("ico", 5)
    eye, basis = nrb.scenes.default_camera()
    outs = []
    for on in (False, True):
        r = nrb.RendererNeural(scene)
        r.resize(200, 120)
        r.set_camera_basis(basis)
        r.set_samples_per_pixel(2)
        r.set_bounce_count(2)
        r.set_tile_order(on)
        r.set_capture(True)
        frames = []
        for k in range(5):
            cam = eye if k < 3 else eye * np.float32(1.1)  # the camera moves before the fourth frame
            r.render(cam)
            frames.append((r.read_pixels().copy(), r.read_buffer("hit_positions").copy(), r.read_buffer("accum").copy()))
        outs.append(frames)
        r.close()
    for a, b in zip(*outs):
        assert np.array_equal(a[0], b[0]) and np.array_equal(a[1], b[1]) and np.array_equal(a[2], b[2])


def test_normal_cone_culling_changes_nothing(oracle_lib):
    """The shells' node pairs carry normal cones (csrc/nrb_scene.cuh): facing-mode traversals skip subtrees in which
    every triangle would fail the facing test.  The run-time-mode loop of the segment stages (parity hook modes 5-7)
    must return exactly what the reference traversal returns - ids included - on convex and on non-convex shells,
    from inside and outside, along axis-parallel and unnormalised directions."""
    for kind, sub in (("ico", 3), ("blobs", 3)):
        scene = nrb.scenes.make_scene(kind, sub, additional=True, env=False, shell_delta=0.3 if kind == "blobs" else 0.12)
        # a bumpy (non-convex) outer shell: cones of coarse nodes are wide, of fine nodes narrow
        hv, hidx = nrb.scenes.icosphere(4)
        scene.outer_shell = nrb.scenes.mesh_from(nrb.scenes.displace(hv * np.float32(1.3), 0.25, seed=9), hidx) if kind == "ico" else scene.outer_shell
        prod, orc, eye = make_pair(scene, oracle_lib, 8, 8)
        o, d = random_rays(30000, seed=99)
        o[:12000] *= 0.35            # origins inside the shells
        d[:3000] *= np.float32(2.5)  # unnormalised directions
        d[3000:3006] = np.eye(3, dtype=np.float32).repeat(2, 0) * np.array([1, -1, 1, -1, 1, -1], np.float32)[:, None]
        for slot in (nrb.SLOT_OUTER, nrb.SLOT_INNER):
            for mode in (0, 1, 2):
                b = orc.trace_rays(slot, mode, o, d)
                hit = b["t"] < 1e30
                for hook in (5, 8, 0):  # the segment stages' loop, the first engine, the round-reconverged walk
                    a = prod.trace_rays(slot, hook + mode, o, d)
                    assert np.array_equal(a["t"], b["t"]), (kind, slot, mode, hook)
                    assert np.array_equal(a["prim"][hit], b["prim"][hit]) and hit.sum() > 500, (kind, slot, mode, hook)
        prod.close()
        orc.close()


def test_async_read_back_is_the_frame():
    """nrb_read_pixels_async (second stream, staged copy): frames streamed back while the next one renders are
    the frames a blocking read returns."""
    import torch
    scene = nrb.scenes.make_scene("ico", 3)
    eye, basis = nrb.scenes.default_camera()

    def renderer():
        r = nrb.RendererNeural(scene)
        r.resize(320, 200)
        r.set_camera_basis(basis)
        r.set_bounce_count(1)
        return r

    a, b = renderer(), renderer()
    n = 320 * 200 * 4
    host = [torch.empty(n, dtype=torch.uint8).pin_memory() for _ in range(4)]
    want = []
    for k in range(4):
        a.render(eye)
        a.read_pixels_async(host[k].data_ptr(), n)  # no wait: the next render is enqueued behind it
        b.render(eye)
        want.append(b.read_pixels().reshape(-1))
    a.sync()
    for k in range(4):
        assert np.array_equal(host[k].numpy(), want[k]), k
    assert not np.array_equal(want[0], want[3])  # accumulation moved on between the frames
    with pytest.raises(nrb.NrbError):
        a.read_pixels_async(host[0].data_ptr(), n - 4)


@pytest.mark.parametrize("path_tail", [0, 1 << 30], ids=["per_stage", "path_tail"])
def test_pipelined_frames_equal_unpipelined(path_tail):
    """NRB_OPT_PIPELINE: frame k+1 starts on a second stream and a second state buffer under the end of frame k.  Every
    frame - enqueued back to back with streamed read-backs, across a camera move, a weight reload, a batch query on the
    renderer's scratch and a frame of the ground-truth mode in between - equals the unpipelined renderer's bit for bit,
    and so do the counters.  (path_tail fixes which kernels run bounces >= 2, so both runs launch the same ones.)"""
    import torch
    scene = nrb.scenes.make_scene("ico", 4, additional=True)
    eye, basis = nrb.scenes.default_camera()
    eye2 = (np.asarray(eye, np.float32) * np.float32(1.07)).astype(np.float32)
    w1, w2 = _weights(), nrb.scenes.make_weights(seed=7)
    W, H = 288, 200
    n = W * H * 4
    x = np.random.RandomState(3).uniform(0, 1, (3000, 12)).astype(np.float32)

    def run(pipeline):
        r = nrb.RendererNeural(scene)
        r.resize(W, H)
        r.set_camera_basis(basis)
        r.set_samples_per_pixel(2)
        r.set_bounce_count(3)
        r.set_path_tail(path_tail)
        r.set_pipeline(pipeline)
        r.load_weights(w1)
        r.set_use_neural_query(True)
        host = [torch.empty(n, dtype=torch.uint8).pin_memory() for _ in range(12)]
        k = 0
        q = None
        for phase in range(4):
            for _ in range(3):
                r.render(eye if phase < 2 else eye2)
                r.read_pixels_async(host[k].data_ptr(), n)
                k += 1
            if phase == 0:
                q = r.neural_query(x)          # the split query's scratch between two pipelined frames
            elif phase == 1:
                r.load_weights(w2)             # new parameters: accumulation continues, the frames change
            elif phase == 2:
                r.set_use_neural_query(False)  # one ground-truth frame (unpipelined), then back
                r.render(eye2)
                r.set_use_neural_query(True)
        r.sync()
        c = r.counters()
        out = [h.numpy().copy() for h in host]
        r.close()
        return out, c, q

    base, cb, qb = run(0)
    for mode in (1,):
        got, cg, qg = run(mode)
        for k, (a, b) in enumerate(zip(base, got)):
            assert np.array_equal(a, b), (mode, k)
        assert np.array_equal(qb, qg)
        for key in ("ray_segments", "neural_queries", "path_vertices"):
            assert cb[key] == cg[key], (mode, key, cb, cg)
    assert not np.array_equal(base[0], base[2]) and not np.array_equal(base[5], base[6])


def test_pipelined_ground_truth_frames_equal_unpipelined():
    """The same for ground-truth frames (unsliced: k_gt_tail and finalize of frame k under k_gt_primary of frame k+1)."""
    import torch
    scene = nrb.scenes.make_scene("ico", 5, additional=True)
    eye, basis = nrb.scenes.default_camera()
    eye2 = (np.asarray(eye, np.float32) * np.float32(0.93)).astype(np.float32)
    W, H = 320, 208
    n = W * H * 4

    def run(pipeline):
        r = nrb.RendererNeural(scene)
        r.resize(W, H)
        r.set_camera_basis(basis)
        r.set_samples_per_pixel(2)
        r.set_bounce_count(3)
        r.set_frame_slices(1)
        r.set_pipeline(pipeline)
        host = [torch.empty(n, dtype=torch.uint8).pin_memory() for _ in range(8)]
        for k in range(8):
            r.render(eye if k < 4 else eye2)
            r.read_pixels_async(host[k].data_ptr(), n)
        r.sync()
        c = r.counters()
        out = [h.numpy().copy() for h in host]
        r.close()
        return out, c

    base, cb = run(False)
    got, cg = run(True)
    for k, (a, b) in enumerate(zip(base, got)):
        assert np.array_equal(a, b), k
    for key in ("ray_segments", "path_vertices"):
        assert cb[key] == cg[key], (key, cb, cg)
    assert not np.array_equal(base[0], base[3]) and not np.array_equal(base[3], base[4])


def test_denoiser_matches_oracle(oracle_lib):
    """NRB_OPT_DENOISE: bilateralDenoiseKernel of the reference (src/denoiser.cuh:47-113) behind finalize, GPU vs
    the oracle's restatement on the same noisy 2-spp frame (8-bit output; device expf/powf may move a value
    by one step), and the property it exists for: closer to a converged render than the unfiltered frame."""
    scene = nrb.scenes.make_scene("ico", 3)
    prod, orc, eye = make_pair(scene, oracle_lib, 160, 96, spp=2, bounces=2)
    for r in (prod, orc):
        r.set_denoise(True)
        r.render(eye)
    a, b = prod.read_pixels(), orc.read_pixels()
    d = np.abs(a.astype(int) - b.astype(int))
    assert d.max() <= 1 and (d > 0).mean() < 0.01, (d.max(), (d > 0).mean())
    prod.set_denoise(False)
    prod.reset_samples()
    prod.render(eye)
    noisy = prod.read_pixels()
    assert not np.array_equal(noisy, a)
    prod.set_samples_per_pixel(64)
    for _ in range(4):
        prod.render(eye)
    converged = prod.read_pixels()
    assert psnr8(a, converged) > psnr8(noisy, converged) + 1.0
    # accumulation is untouched by the filter: turning it off later shows the plain running mean
    prod.set_tile(16, 0, 2)
    prod.set_denoise(True)
    with pytest.raises(nrb.NrbError):
        prod.render(eye)


def _check_tree(nodes, idx, prim, v, src_idx, max_leaf):
    """Structural validity of a reference-format BVH: every triangle in exactly one reachable leaf, children inside
    their parent's box, triangles inside their leaf's box, leaf sizes bounded, index buffer = permuted source."""
    nt = len(idx)
    assert np.array_equal(np.sort(prim), np.arange(nt)) and np.array_equal(idx, src_idx[prim])
    covered = np.zeros(nt, np.int32)
    todo, depth_max = [(0, 1)], 1
    while todo:
        i, depth = todo.pop()
        n = nodes[i]
        depth_max = max(depth_max, depth)
        if n["is_leaf"]:
            first, count = int(n["first"]), int(n["count"])
            assert 1 <= count <= max_leaf
            covered[first:first + count] += 1
            tri = v[idx[first:first + count].reshape(-1)]
            assert np.all(tri >= n["bounds_min"] - 0) and np.all(tri <= n["bounds_max"] + 0)
        else:
            for c in (int(n["left"]), int(n["right"])):
                assert 0 < c < len(nodes)
                assert np.all(nodes[c]["bounds_min"] >= n["bounds_min"]) and np.all(nodes[c]["bounds_max"] <= n["bounds_max"])
                todo.append((c, depth + 1))
    assert np.all(covered == 1)
    return depth_max


@pytest.mark.parametrize("sub,max_leaf", [(0, 4), (3, 1), (5, 4)])
def test_gpu_bvh_build(sub, max_leaf):
    """nrb_build_bvh_gpu (LBVH on the device): a valid tree in the reference's node format, and rays traced through it
    find the same source triangle at the bit-identical t as through the host SAH build."""
    v0, f = nrb.scenes.icosphere(sub)
    v = nrb.scenes.displace(v0, 0.1)
    nodes, idx, prim, ms = nrb.scenes.build_bvh_gpu(v, f, max_leaf=max_leaf)
    depth = _check_tree(nodes, idx, prim, v, np.ascontiguousarray(f, np.uint32), max_leaf)
    assert depth <= 64
    h_nodes, h_idx, h_prim = nrb.scenes.build_bvh(v, f)
    o, d = random_rays(20000, seed=31 + sub)
    hits = []
    for nd, ix in ((nodes, idx), (h_nodes, h_idx)):
        scene = nrb.Scene(original_mesh=nrb.MeshData(vertices=v, indices=ix, nodes=nd))
        r = nrb.RendererNeural(scene)
        hits.append([r.trace_rays(nrb.SLOT_ORIGINAL, mode, o, d) for mode in (0, 1, 2)])
        r.close()
    for a, b in zip(*hits):
        assert np.array_equal(a["t"], b["t"])
        hit = b["t"] < 1e30
        assert hit.sum() > 100
        same = prim[a["prim"][hit]] == h_prim[b["prim"][hit]]
        assert same.mean() > 0.999  # equal-t ties (shared edges) may resolve to the neighbouring triangle
    print(f"LBVH build of {len(f)} triangles: {ms:.3f} ms on the device, depth {depth}, {len(nodes)} nodes")


def test_gpu_bvh_build_large_and_degenerate():
    v0, f = nrb.scenes.icosphere(7)
    v = nrb.scenes.displace(v0, 0.1)
    import time
    t0 = time.time()
    nodes, idx, prim, ms = nrb.scenes.build_bvh_gpu(v, f)
    wall = time.time() - t0
    t0 = time.time()
    nrb.scenes.build_bvh(v, f)
    host = time.time() - t0
    print(f"327680 triangles: device build {ms:.2f} ms ({wall * 1e3:.0f} ms with copies) vs host SAH build {host * 1e3:.0f} ms")
    assert ms < 50.0
    # all triangles identical (every Morton code equal): the index half of the key keeps the tree well formed
    tri = np.tile(np.array([[0, 1, 2]], np.uint32), (1000, 1))
    pts = np.array([[0, 0, 0], [1, 0, 0], [0, 1, 0]], np.float32)
    nodes, idx, prim, _ = nrb.scenes.build_bvh_gpu(pts, tri, max_leaf=2)
    assert _check_tree(nodes, idx, prim, pts, tri, 2) <= 64
    with pytest.raises(nrb.NrbError):
        nrb.scenes.build_bvh_gpu(pts, np.array([[0, 1, 7]], np.uint32))


@pytest.mark.parametrize("neural", [False, True], ids=["gt", "neural"])
def test_peer_memory_gather_two_processes(neural):
    """nrb_gather_export / nrb_gather_attach: a second PROCESS (its own renderer, here on the same GPU, as a second
    GPU's process would be) stores the row bands it owns into the root's frame buffer through CUDA IPC peer memory;
    after the root's render the root holds the frame a single full-frame renderer produces, bit for bit."""
    import multiprocessing as mp
    from tests.util import peer_rank_main
    W, H, frames = 320, 208, 3
    scene = nrb.scenes.make_scene("ico", 3)
    eye, basis = nrb.scenes.default_camera()

    def renderer():
        r = nrb.RendererNeural(scene)
        r.resize(W, H)
        r.set_camera_basis(basis)
        r.set_bounce_count(2)
        if neural:
            r.load_weights(nrb.scenes.make_weights())
            r.set_use_neural_query(True)
        return r

    full = renderer()
    for _ in range(frames):
        full.render(eye)
    want = full.read_pixels()
    root = renderer()
    root.set_tile(16, 0, 2)
    ctx = mp.get_context("spawn")
    here, there = ctx.Pipe()
    p = ctx.Process(target=peer_rank_main, args=(there, 1, 2, W, H, frames, neural))
    p.start()
    try:
        here.send(root.gather_export(1))
        assert here.poll(120) and here.recv() == "attached"
        for _ in range(frames):
            root.render(eye)
        root.sync()
        got = root.read_pixels()
        assert here.poll(120) and here.recv() == "done"
        here.send("bye")
    finally:
        p.join(60)
        if p.is_alive():
            p.kill()
    assert np.array_equal(got, want)
    # a root whose peer never renders gives up instead of hanging the GPU, and says so
    root.gather_export(1)
    root.render(eye)
    with pytest.raises(nrb.NrbError):
        root.sync()
    root.gather_detach()
    root.render(eye)
    root.sync()
