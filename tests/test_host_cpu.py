"""CPU tests of the host side: the C-ABI library loads and exports every declared symbol, fails
loudly without a device, the host BVH builder is sound, tile partition + gather work over gloo."""
import ctypes as C
import os
import re
import subprocess
import sys

import numpy as np
import pytest

from tests.util import nrb

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _has_gpu():
    try:
        import torch
        return torch.cuda.is_available()
    except Exception:
        return False


def test_library_exports_every_declared_symbol():
    hdr = open(os.path.join(ROOT, "include", "nrb200.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    declared = set(re.findall(r"\b(nrb_[a-z0-9_]+)\s*\(", hdr))
    assert len(declared) >= 38
    lib = nrb.product_lib()
    missing = [name for name in sorted(declared) if not hasattr(lib.dll, name)]
    assert not missing, missing
    assert lib.fn("abi_version")() == 1


def test_no_cpu_fallback_create_fails_without_device():
    if _has_gpu():
        pytest.skip("a CUDA device is present")
    lib = nrb.product_lib()
    h = C.c_void_p()
    rc = lib.fn("create")(None, -1, C.byref(h))
    assert rc == -2 and not h.value
    assert b"no usable CUDA device" in lib.fn("last_error")()
    with pytest.raises(nrb.NrbError):
        nrb.RendererNeural(nrb.Scene())


def test_missing_library_is_an_error(tmp_path):
    with pytest.raises(nrb.NrbError):
        nrb.Lib(str(tmp_path / "libnrb200.so"), "nrb_")


@pytest.mark.parametrize("sub", [0, 2, 4])
def test_host_bvh_builder_is_sound(sub):
    v, idx = nrb.scenes.icosphere(sub)
    nodes, order, prim = nrb.scenes.build_bvh(v, idx)
    assert sorted(prim.tolist()) == list(range(len(idx)))
    assert np.array_equal(order, idx[prim])
    assert nodes[0]["is_leaf"] == (len(idx) <= 2) or len(nodes) > 1
    covered = np.zeros(len(idx), int)
    stack, depth = [(0, 1)], 0
    while stack:
        i, d = stack.pop()
        depth = max(depth, d)
        n = nodes[i]
        if n["is_leaf"]:
            f, c = int(n["first"]), int(n["count"])
            assert 1 <= c <= 32
            covered[f:f + c] += 1
            tri = v[order[f:f + c].reshape(-1)]
            assert np.all(tri >= n["bounds_min"] - 1e-6) and np.all(tri <= n["bounds_max"] + 1e-6)
        else:
            assert n["right"] == n["left"] + 1  # the reference's flattening (src/mesh_bvh.cpp:131-133)
            for c in (int(n["left"]), int(n["right"])):
                assert np.all(nodes[c]["bounds_min"] >= n["bounds_min"]) and np.all(nodes[c]["bounds_max"] <= n["bounds_max"])
                stack.append((c, d + 1))
    assert np.all(covered == 1) and depth < 64


def test_struct_layouts_match_reference_sizes():
    K = nrb._capi
    assert C.sizeof(K.BvhNode) == 44 and C.sizeof(K.Material) == 180 and C.sizeof(K.MaterialParam) == 12
    assert C.sizeof(K.MaterialParam3) == 20 and K.BVH_NODE_DTYPE.itemsize == 44 and K.HIT_DTYPE.itemsize == 16


def test_tile_partition_covers_frame():
    for h, band, world in ((1080, 16, 8), (70, 8, 3), (5, 16, 4), (2160, 16, 2)):
        rows = [nrb.tiles.owned_rows(h, band, r, world) for r in range(world)]
        assert sorted(np.concatenate(rows).tolist()) == list(range(h))
        for r in range(world):
            flat = [y for a, b in nrb.tiles.owned_bands(h, band, r, world) for y in range(a, b)]
            assert flat == rows[r].tolist()


def test_camera_basis_matches_evaluate_driver():
    b = nrb.compute_camera_basis(-90.0, 0.0, 0.785)
    assert np.allclose(b.forward, [0, 0, -1], atol=1e-6) and np.allclose(b.up, [0, 1, 0], atol=1e-6)
    assert np.allclose(np.cross(b.right, b.forward), b.up, atol=1e-6)


WORKER = r'''
import importlib, os, sys
import numpy as np, torch, torch.distributed as dist
sys.path.insert(0, sys.argv[1])
nrb = importlib.import_module("neural-renderer_b200")
from oracle import loader
rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
dist.init_process_group("gloo")
lib = loader.oracle_lib("auto")
scene = nrb.scenes.make_scene("ico", 2, env_size=(64, 32))
eye, basis = nrb.scenes.default_camera()
W, H, BAND = 48, 36, 8
def renderer():
    r = nrb.RendererNeural(scene, lib=lib)
    r.resize(W, H); r.set_camera_basis(basis); r.set_samples_per_pixel(2); r.set_bounce_count(2)
    return r
r = renderer()
frame = np.zeros((H, W, 4), np.uint8)
for (y0, y1) in nrb.tiles.owned_bands(H, BAND, rank, world):   # what one GPU rank renders
    r.reset_samples()
    lib.fn("set_row_range")(r.handle, y0, y1)
    r.render(eye)
    frame[y0:y1] = r.read_pixels()[y0:y1]
full = nrb.tiles.gather_frame(torch.from_numpy(frame), dst=0)
if rank == 0:
    ref = renderer(); ref.render(eye)
    assert np.array_equal(full.numpy(), ref.read_pixels()), "gathered tiles differ from the single-rank frame"
    print("GATHER_OK")
dist.barrier(); dist.destroy_process_group()
'''


def test_two_rank_tile_gather_over_gloo(tmp_path):
    from oracle import loader
    if not (loader.have("ref") or loader.have("port")):
        import __graft_entry__ as g
        g.build_oracle()
    script = tmp_path / "worker.py"
    script.write_text(WORKER)
    env = dict(os.environ, MASTER_ADDR="127.0.0.1", MASTER_PORT="29533", OMP_NUM_THREADS="1")
    p = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2",
                        "--master-addr", "127.0.0.1", "--master-port", "29533", str(script), ROOT],
                       env=env, capture_output=True, text=True, timeout=600)
    assert p.returncode == 0, p.stdout[-2000:] + p.stderr[-3000:]
    assert "GATHER_OK" in p.stdout


def test_cpp_shim_compiles_against_reference_headers(tmp_path):
    """include/nrb200_renderer.hpp (class RendererNeural over the C ABI) must compile against the
    reference's own, unmodified scene.h / mesh.h / renderer.h / config_loader.h and run its evaluate
    call sequence.  Only possible where /root/reference exists."""
    ref = "/root/reference/src"
    if not os.path.isdir(ref):
        pytest.skip("/root/reference is not present on this machine")
    out = tmp_path / "shim.o"
    p = subprocess.run(["nvcc", "-std=c++17", "-x", "cu", "-c", "-o", str(out), f"-I{ref}",
                        "-I/root/reference/ext/tiny-cuda-nn/dependencies", f"-I{ROOT}/include",
                        f"{ROOT}/tests/cpp/shim_compile_check.cpp"], capture_output=True, text=True, timeout=600)
    assert p.returncode == 0 and out.exists(), p.stderr[-3000:]
