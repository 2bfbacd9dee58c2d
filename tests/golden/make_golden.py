"""Regenerates tests/golden/*.npz from the reference's own code (oracle/_ref: the reference headers
compiled for the host where they lie under /root/reference).  Run from the repo root in the build
container:  python tests/golden/make_golden.py
Inputs are the seeded synthetic scenes of neural-renderer_b200/scenes.py, so only outputs are stored.
"""
import importlib
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
HERE = os.path.dirname(os.path.abspath(__file__))


def golden_scene(nrb):
    return nrb.scenes.make_scene("ico", 2, additional=True, textured=True, env_size=(128, 64))


def golden_rays():
    from tests.util import random_rays
    return random_rays(1000, seed=2024)


def golden_query_inputs(dims=12, n=512):
    return np.random.RandomState(77).uniform(0, 1, (n, dims)).astype(np.float32)


FRAME = dict(width=64, height=40)


def render_all(nrb, lib):
    """Everything the fixtures hold, computed through `lib` (an orc_ or nrb_ library)."""
    out = {}
    scene = golden_scene(nrb)
    eye, basis = nrb.scenes.default_camera()
    r = nrb.RendererNeural(scene, lib=lib)
    if lib is not None and lib.has("set_threads"):
        lib.fn("set_threads")(r.handle, 0)
    o, d = golden_rays()
    for slot, sname in ((nrb.SLOT_ORIGINAL, "orig"), (nrb.SLOT_OUTER, "outer"), (nrb.SLOT_ADDITIONAL, "add")):
        for mode in (0, 1, 2):
            out[f"trace_{sname}_{mode}"] = r.trace_rays(slot, mode, o, d)
    r.resize(FRAME["width"], FRAME["height"])
    r.set_camera_basis(basis)
    r.set_samples_per_pixel(2)
    r.set_bounce_count(2)
    r.render(eye)
    out["gt_pixels"] = r.read_pixels()
    out["gt_hit_flags"] = r.read_buffer("hit_flags")
    out["gt_hit_positions"] = r.read_buffer("hit_positions")
    out["gt_radiance"] = r.read_buffer("path_radiance")
    r.load_weights(nrb.scenes.make_weights())
    out["query"] = r.neural_query(golden_query_inputs(r.input_dims()))
    r.set_use_neural_query(True)
    r.set_samples_per_pixel(1)
    r.set_bounce_count(1)
    r.render(eye)
    out["neural_pixels"] = r.read_pixels()
    out["neural_hit_flags"] = r.read_buffer("hit_flags")
    r.close()
    return out


if __name__ == "__main__":
    nrb = importlib.import_module("neural-renderer_b200")
    from oracle import loader
    assert loader.have("ref"), "oracle/_ref is missing: run `make -C oracle ref` where /root/reference exists"
    lib = loader.oracle_lib("ref")
    assert lib.fn("backend")().decode() == "reference"
    data = render_all(nrb, lib)
    np.savez_compressed(os.path.join(HERE, "frames_v1.npz"), **data)
    print({k: (v.shape, str(v.dtype)) for k, v in data.items()})
    print(os.path.getsize(os.path.join(HERE, "frames_v1.npz")), "bytes")
