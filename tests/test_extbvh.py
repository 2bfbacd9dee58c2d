"""BASELINE config 1: the vendored ext/bvh CPU traversal (bvh::v2::Bvh::intersect + PrecomputedTri::intersect, configured
as ext/bvh/test/benchmark.cpp; ext/bvh/src/bvh/v2/bvh.h:124-156, tri.h:58-74) on primary rays of a scene mesh -> hit-id /
t / normal buffers, as the second oracle north_star names.  It is built from the sources under /root/reference into
oracle/_ref/libnrb_refbvh.so (oracle/Makefile ref) and travels as a binary.

Its triangle test is not the renderer's (tolerance -eps on the barycentrics, no determinant cut, no facing cull,
different edge arithmetic), so it is a cross-check, not the bit-exact arbiter: >= 99.9 % of the rays must report the
same source triangle as traceMeshWithMode (mode ANY); on those t agrees within 1e-5 relative for 99.9 % and 1e-4 for all
(two different fp32 formulas on grazing hits; observed max 1.3e-5), normals equal to 1e-5.
"""
import numpy as np
import pytest

from tests.util import nrb, primary_rays


def _mesh(sub):
    v0, f = nrb.scenes.icosphere(sub)
    return nrb.scenes.displace(v0, 0.1), np.ascontiguousarray(f, np.uint32)


def _loader():
    from oracle import loader
    if not loader.have_refbvh():
        pytest.skip("oracle/_ref/libnrb_refbvh.so not built (needs /root/reference at build time)")
    return loader


def _agreement(ext, hits, prim_ids, what):
    """ext: ref_extbvh_trace result (source ids); hits: nrb/orc trace result (leaf-order ids of a tree with prim_ids)."""
    h_ext, h_ref = ext["prim"] != 0xFFFFFFFF, hits["t"] < 1e30
    src = prim_ids[np.where(h_ref, hits["prim"], 0)]
    same = (h_ext == h_ref) & (~h_ref | (src == ext["prim"]))
    both = h_ext & h_ref & (src == ext["prim"])
    rel = np.abs(ext["t"][both] - hits["t"][both]) / hits["t"][both]
    print(f"{what}: id agreement {same.mean():.6f} over {len(same)} rays ({h_ref.mean():.3f} hit), t rel max {rel.max():.2e}")
    assert same.mean() >= 0.999, same.mean()
    assert np.percentile(rel, 99.9) <= 1e-5 and rel.max() <= 1e-4, (np.percentile(rel, 99.9), rel.max())
    return both


def test_simple_example_known_answer():
    """ext/bvh/test/simple_example.cpp: distance 1, barycentrics (-0, 0.5)."""
    loader = _loader()
    v = np.array([[1, -1, 1], [1, 1, 1], [-1, 1, 1], [-1, -1, 1]], np.float32)
    f = np.array([[0, 1, 2], [0, 3, 2]], np.uint32)
    r = loader.ref_extbvh_trace(v, f, [[0, 0, 0]], [[0, 0, 1]], tmin=0.0)
    assert r["t"][0] == 1.0 and r["uv"][0, 1] == 0.5 and abs(r["uv"][0, 0]) == 0.0 and r["prim"][0] in (0, 1)
    assert np.allclose(r["normal"][0], [0, 0, 1])
    miss = loader.ref_extbvh_trace(v, f, [[0, 0, 0]], [[0, 0, -1]])
    assert miss["prim"][0] == 0xFFFFFFFF and miss["t"][0] >= 1e30


def test_config1_extbvh_against_reference_traversal_cpu(oracle_lib):
    """512 x 512 primary rays of the displaced icosphere (81 920 triangles): ext/bvh vs the reference's traceMeshWithMode
    compiled for the host."""
    loader = _loader()
    v, f = _mesh(6)
    nodes, idx, prim = nrb.scenes.build_bvh(v, f)
    scene = nrb.Scene(original_mesh=nrb.MeshData(vertices=v, indices=idx, nodes=nodes))
    orc = nrb.RendererNeural(scene, lib=oracle_lib)
    oracle_lib.fn("set_threads")(orc.handle, 0)
    o, d = primary_rays(512, 512)
    ext = loader.ref_extbvh_trace(v, f, o, d)
    hits = orc.trace_rays(nrb.SLOT_ORIGINAL, nrb.TRACE_ANY, o, d)
    both = _agreement(ext, hits, prim, "config 1 (CPU)")
    # hit-id / t / normal buffers: the normal of the reported triangle is the geometric normal of that source triangle
    tri = v[f[ext["prim"][both]]]
    n = np.cross(tri[:, 1] - tri[:, 0], tri[:, 2] - tri[:, 0])
    n /= np.linalg.norm(n, axis=1, keepdims=True)
    assert np.abs(n - ext["normal"][both]).max() <= 1e-5
    orc.close()


@pytest.mark.gpu
@pytest.mark.parametrize("size,sub", [((512, 512), 6), ((1920, 1080), 7)], ids=["512x512-S6", "1080p-S2"])
def test_config1_extbvh_against_gpu_traversal(size, sub):
    """The same cross-check against nrb_trace_rays on the GPU, with the timed CPU figures of SURVEY.md 8d
    (all host cores and one core; median of 3 passes)."""
    loader = _loader()
    v, f = _mesh(sub)
    nodes, idx, prim = nrb.scenes.build_bvh(v, f)
    scene = nrb.Scene(original_mesh=nrb.MeshData(vertices=v, indices=idx, nodes=nodes))
    r = nrb.RendererNeural(scene)
    o, d = primary_rays(*size)
    ext = loader.ref_extbvh_trace(v, f, o, d, threads=0, reps=3)
    one = loader.ref_extbvh_trace(v, f, o[::16], d[::16], threads=1, reps=3)
    hits = r.trace_rays(nrb.SLOT_ORIGINAL, nrb.TRACE_ANY, o, d)
    _agreement(ext, hits, prim, f"config 1 (GPU) {size[0]}x{size[1]}")
    print(f"ext/bvh CPU traversal, {len(f)} triangles, {len(o)} primary rays: {len(o) / ext['trace_ms'] * 1e-3:.2f} Mrays/s on "
          f"{ext['threads']} threads, {len(o[::16]) / one['trace_ms'] * 1e-3:.2f} Mrays/s on 1 thread; build {ext['build_ms']:.0f} ms")
    r.close()
