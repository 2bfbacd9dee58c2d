"""Host-side checks of the evaluate driver and its metrics (reference src/evaluate.cu, src/image_utils.h)."""
import importlib
import json
import os
import struct

import numpy as np
import pytest

nrb = importlib.import_module("neural-renderer_b200")
metrics = importlib.import_module("neural-renderer_b200.metrics")
ev = importlib.import_module("neural-renderer_b200.evaluate")
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _image(seed=0, h=72, w=96):
    rng = np.random.RandomState(seed)
    yy, xx = np.mgrid[0:h, 0:w]
    base = np.stack([128 + 100 * np.sin(xx / 9.0), 128 + 90 * np.cos(yy / 7.0), 100 + 0.8 * xx], -1)
    img = np.clip(base + rng.normal(0, 4, base.shape), 0, 255).astype(np.uint8)
    return np.concatenate([img, np.full((h, w, 1), 255, np.uint8)], -1)


def test_psnr_definition():
    a = _image()
    assert metrics.psnr8(a, a) == 100.0
    b = a.copy()
    b[..., :3] = np.where(a[..., :3] < 255, a[..., :3] + 1, a[..., :3] - 1)
    assert abs(metrics.psnr8(a, b) - 10 * np.log10(255.0 ** 2)) < 1e-9  # MSE = 1


def test_flip_properties():
    a = _image()
    assert metrics.flip_ldr(a, a) == 0.0
    rng = np.random.RandomState(1)
    small = np.clip(a.astype(int) + rng.randint(-3, 4, a.shape), 0, 255).astype(np.uint8)
    large = np.clip(a.astype(int) + rng.randint(-40, 41, a.shape), 0, 255).astype(np.uint8)
    fs, fl = metrics.flip_ldr(a, small), metrics.flip_ldr(a, large)
    assert 0.0 < fs < fl < 1.0
    assert abs(metrics.flip_ldr(a, small) - metrics.flip_ldr(small, a)) < 1e-12  # symmetric
    black, white = np.zeros_like(a), np.full_like(a, 255)
    m = metrics.flip_ldr_map(black, white)
    assert m.min() > 0.9 and m.max() <= 1.0
    # an edit is seen where it is made, not elsewhere (filters have a finite radius)
    c = a.copy()
    c[10:20, 10:20, :3] = 255 - c[10:20, 10:20, :3]
    fm = metrics.flip_ldr_map(a, c)
    assert fm[12:18, 12:18].mean() > 0.3 and fm[40:, 50:].max() == 0.0
    assert abs(metrics.pixels_per_degree() - 0.7 * 3840 / 0.7 * np.pi / 180) < 1e-12


def test_config_defaults_and_camera(tmp_path):
    p = tmp_path / "c.json"
    p.write_text("{}")
    c = ev.load_config(str(p))
    assert c["rendering"] == {"total_samples": 2048, "bounce_count": 3, "width": 1920, "height": 1080}
    assert abs(c["camera"]["yfov"] - 1.047198) < 1e-9 and c["neural_network"]["log2_hashmap_size"] == 14
    assert c["neural_network"]["base_resolution"] == 16 and c["material"]["roughness"] == 1.0
    q = tmp_path / "d.json"
    q.write_text(json.dumps({"scene": {"original_mesh": {"path": "m/a.obj", "scale": 2.0}}, "checkpoint_path": "w.bin"}))
    d = ev.load_config(str(q))
    assert d["original_mesh"] == {"path": str(tmp_path / "m" / "a.obj"), "scale": 2.0} and d["checkpoint_path"] == str(tmp_path / "w.bin")
    eye, tgt = np.array([2.0, 1.2, 2.0]), np.zeros(3)
    pos, yaw, pitch = ev.matrix_to_camera_state(ev.camera_state_to_matrix(eye, tgt))
    assert np.allclose(pos, eye, atol=1e-5)
    b = nrb.renderer.compute_camera_basis(yaw, pitch, 0.785)
    assert np.allclose(b.forward, (tgt - eye) / np.linalg.norm(tgt - eye), atol=1e-5)


def test_asset_loaders(tmp_path):
    obj = tmp_path / "q.obj"
    obj.write_text("v 0 0 0\nv 1 0 0\nv 1 1 0\nv 0 1 0\nvn 0 0 1\nf 1//1 2//1 3//1 4//1\nf -4 -3 -2\n")
    v, f = ev.load_obj(str(obj))
    assert v.shape == (4, 3) and f.tolist() == [[0, 1, 2], [0, 2, 3], [0, 1, 2]]
    ply = tmp_path / "a.ply"
    ply.write_text("ply\nformat ascii 1.0\nelement vertex 3\nproperty float x\nproperty float y\nproperty float z\n"
                   "element face 1\nproperty list uchar int vertex_indices\nend_header\n0 0 0\n1 0 0\n0 1 0\n3 0 1 2\n")
    v, f = ev.load_ply(str(ply))
    assert v.shape == (3, 3) and f.tolist() == [[0, 1, 2]]
    bply = tmp_path / "b.ply"
    head = ("ply\nformat binary_little_endian 1.0\nelement vertex 3\nproperty float x\nproperty float y\nproperty float z\n"
            "element face 1\nproperty list uchar uint vertex_indices\nend_header\n").encode()
    bply.write_bytes(head + struct.pack("<9f", 0, 0, 0, 1, 0, 0, 0, 1, 0) + struct.pack("<B3I", 3, 0, 1, 2))
    v2, f2 = ev.load_ply(str(bply))
    assert np.array_equal(v, v2) and np.array_equal(f, f2)
    # Radiance .hdr: one RLE scanline (a run and a literal span per channel) and one flat image
    w = 10
    scan = bytes([2, 2, 0, w])
    for ch, val in enumerate((128, 64, 32, 129)):  # mantissas 128, 64, 32; exponent 129 -> 2^(129-136)
        scan += bytes([128 + 6, val]) + bytes([4, val, val, val, val])
    hdr = tmp_path / "e.hdr"
    hdr.write_bytes(b"#?RADIANCE\nFORMAT=32-bit_rle_rgbe\n\n-Y 1 +X 10\n" + scan)
    img = ev.load_hdr(str(hdr))
    assert img.shape == (1, 10, 3) and np.allclose(img[0, :, 0], 128 / 128.0) and np.allclose(img[0, :, 2], 32 / 128.0)
    flat = tmp_path / "f.hdr"
    flat.write_bytes(b"#?RADIANCE\n\n-Y 1 +X 2\n" + bytes([128, 0, 0, 129, 0, 0, 0, 0]))
    img = ev.load_hdr(str(flat))
    assert img.shape == (1, 2, 3) and img[0, 0, 0] == 1.0 and img[0, 1].sum() == 0.0


def test_evaluate_runs_on_the_oracle():
    """The whole driver (config -> scene -> GT + neural renders -> PSNR, FLIP) on the CPU oracle, tiny frame."""
    from oracle import loader
    cfg = ev.load_config(os.path.join(ROOT, "configs", "synthetic_ico.json"))
    cfg["original_mesh"]["path"] = "synthetic:ico:2"
    cfg["rendering"]["total_samples"] = 10  # a batch of 8 and a batch of 2
    cfg["rendering"]["bounce_count"] = 1
    res = ev.run(cfg, None, 48, 32, lib=loader.oracle_lib("auto"), quiet=True)
    assert res["gt"].shape == (32, 48, 4) and res["neural"].shape == (32, 48, 4)
    assert 5.0 < res["psnr"] < 100.0 and 0.0 < res["flip"] < 1.0
    assert res["gt"][..., :3].std() > 5  # an image, not a constant


@pytest.mark.gpu
def test_evaluate_gpu_matches_oracle():
    """Image parity at the caller's level: libnrb200 and the CPU oracle run the same evaluate config; the
    frames must agree within the north-star tolerances (0.1 dB-scale PSNR, mean FLIP <= 0.005) for the
    ground-truth pass; the neural pass is reported with the fp16 hit-logit caveat of the frame tests."""
    from oracle import loader
    cfg = ev.load_config(os.path.join(ROOT, "configs", "synthetic_ico.json"))
    cfg["original_mesh"]["path"] = "synthetic:ico:3"
    cfg["rendering"]["total_samples"] = 16
    a = ev.run(cfg, None, 192, 108, quiet=True)
    b = ev.run(cfg, None, 192, 108, lib=loader.oracle_lib("auto"), quiet=True)
    gt_flip, gt_psnr = metrics.flip_ldr(b["gt"], a["gt"]), metrics.psnr8(b["gt"], a["gt"])
    nn_flip, nn_psnr = metrics.flip_ldr(b["neural"], a["neural"]), metrics.psnr8(b["neural"], a["neural"])
    print(f"evaluate parity: GT PSNR {gt_psnr:.2f} dB FLIP {gt_flip:.5f}; neural PSNR {nn_psnr:.2f} dB FLIP {nn_flip:.5f}; "
          f"driver metrics GPU {a['psnr']:.2f} dB / {a['flip']:.4f}, oracle {b['psnr']:.2f} dB / {b['flip']:.4f}")
    assert gt_psnr > 45.0 and gt_flip <= 0.005
    assert nn_psnr > 30.0 and nn_flip <= 0.02
    assert abs(a["psnr"] - b["psnr"]) < 0.5 and abs(a["flip"] - b["flip"]) < 0.01
