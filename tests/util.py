"""Shared helpers of the parity tests: paired product/oracle renderers on the same scene."""
import importlib

import numpy as np

nrb = importlib.import_module("neural-renderer_b200")


def make_pair(scene, oracle_lib, width, height, spp=1, bounces=0, neural=False, weights=None, nn_config=None,
              threads=0):
    eye, basis = nrb.scenes.default_camera()
    out = []
    for lib in (None, oracle_lib):
        r = nrb.RendererNeural(scene, nn_config=nn_config, lib=lib)
        if lib is not None:
            lib.fn("set_threads")(r.handle, threads)
        r.resize(width, height)
        r.set_camera_basis(basis)
        r.set_samples_per_pixel(spp)
        r.set_bounce_count(bounces)
        if weights is not None:
            r.load_weights(weights)
        r.set_use_neural_query(neural)
        if lib is None:
            r.set_capture(True)  # the parity tests read per-element buffers (nrb_read_buffer): whole-frame element order, unsliced
        out.append(r)
    return out[0], out[1], eye


def psnr8(a, b):
    """computePsnr of the reference (src/image_utils.h:84-99): 8-bit sRGB MSE over RGB."""
    d = a[..., :3].astype(np.float64) - b[..., :3].astype(np.float64)
    mse = np.mean(d * d)
    return 99.0 if mse == 0 else 10.0 * np.log10(255.0 * 255.0 / mse)


def rae_ok(a, b, tol=1e-2, floor=1e-2):
    """Symmetric relative error of tiny-cuda-nn's tests (tests/test_common.h:63-113), with an absolute floor."""
    a = a.astype(np.float64)
    b = b.astype(np.float64)
    err = np.abs(a - b) / np.maximum(np.maximum(np.abs(a), np.abs(b)), floor)
    return err


def random_rays(n, seed, radius=3.0):
    rng = np.random.RandomState(seed)
    o = rng.standard_normal((n, 3))
    o = o / np.linalg.norm(o, axis=1, keepdims=True) * radius * rng.uniform(0.2, 1.0, (n, 1))
    tgt = rng.uniform(-0.8, 0.8, (n, 3))
    d = tgt - o
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    return o.astype(np.float32), d.astype(np.float32)


def peer_rank_main(conn, rank, world, width, height, frames, neural):
    """Body of a non-root rank of the peer-memory gather test: its own process, its own renderer."""
    try:
        scene = nrb.scenes.make_scene("ico", 3)
        eye, basis = nrb.scenes.default_camera()
        r = nrb.RendererNeural(scene)
        r.resize(width, height)
        r.set_camera_basis(basis)
        r.set_bounce_count(2)
        if neural:
            r.load_weights(nrb.scenes.make_weights())
            r.set_use_neural_query(True)
        r.set_tile(16, rank, world)
        r.gather_attach(conn.recv())
        conn.send("attached")
        for _ in range(frames):
            r.render(eye)
        r.sync()
        conn.send("done")
        conn.recv()  # keep the mapping alive until the root has read the frame
        r.close()
    except Exception as e:  # pragma: no cover
        conn.send(f"error: {e!r}")


def primary_rays(width, height, radius=1.0):
    """Pixel-centre pinhole rays of the default camera (generatePrimaryRay, src/mesh_traversal.cuh:40-55, without jitter)."""
    eye, basis = nrb.scenes.default_camera(radius)
    x, y = np.meshgrid(np.arange(width, dtype=np.float32), np.arange(height, dtype=np.float32))
    u = (x + 0.5) / width * 2.0 - 1.0
    v = (1.0 - (y + 0.5) / height) * 2.0 - 1.0
    th = np.float32(np.tan(basis.fov_y * 0.5))
    d = (basis.forward[None, None, :] + basis.right[None, None, :] * (u * (width / height) * th)[..., None]
         + basis.up[None, None, :] * (v * th)[..., None])
    d = (d / np.linalg.norm(d, axis=-1, keepdims=True)).astype(np.float32).reshape(-1, 3)
    o = np.broadcast_to(np.asarray(eye, np.float32), d.shape).copy()
    return o, d
