#!/usr/bin/env python
"""bench.py - neural-render Mrays/s at 1080p (BASELINE.json metric) for libnrb200 on N B200s.

One "step" = one RendererNeural::render() call = one frame of the hot path (ray generation, shell BVH traversal, neural
query on the shell segments, Disney-BRDF bounces, accumulation + sRGB) over W*H*spp camera paths; Mrays/s = W*H*spp /
frame time (the reference's KernelTimings.rayCount convention, src/cuda_renderer_neural.cu:2741).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference] [--scaling weak|strong]
                  [--mode neural|gt] [--width 1920 --height 1080 --spp 1 --bounces 3 --subdiv 7]

JSON keys.  value = device-timed, scene and state resident in HBM.  e2e = the same metric through the public host API:
camera in from host memory every step, the RGBA frame out to pinned host memory every step, copies inside the timed
region.  roofline = the kernel with the largest share of the frame (per-launch CUDA events on the renderer's stream,
a second pass over the same K steps with NRB_OPT_PROFILE) against the roofline that bounds it, with the tensor-pipe
figure of the MLP kernel beside it.  gpu_baseline = the reference's OWN CUDA + tiny-cuda-nn build (baseline/_ref,
sm_100) rendering the same scene pack on the same GPU right after (3 warm-up + 20 timed render() calls, cudaEvents).
cpu_baseline = the CPU oracle (the reference's headers compiled for the host) on a bounded uniform sample of the frame's
rows; cpu_extbvh = ext/bvh's own CPU traversal (BASELINE config 1) on the primary rays.
--impl reference times the reference's CPU implementation of the path on all host cores, on a uniform sample of row
bands per step (what is rendered is what is reported: no extrapolation).
"""
import argparse
import importlib
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402,F401

FLOPS_PER_QUERY = 2 * (112 * 128 + 3 * 128 * 128 + 128 * 16)  # 131072 for the reference network
HBM_BYTES_PER_QUERY = 36 + 16

# Algorithmic HBM bytes per unit of work of each kernel: what the stage must read and write
# per element of its queue when everything re-used (BVH, tables, weights) stays on chip.  unit -> which device counter
# gives the units per step.  The traversal kernels gather their BVH from L1/L2 (two 1280-triangle shells = 0.2 MB; the
# 327 680-triangle mesh of gt mode = 26 MB, L2-resident), so these figures are far below the HBM roofline by
# construction: those kernels are bound by instruction issue and load latency (ncu: profiles/), which is what the
# fraction shows.
KERNEL_BYTES = {
    "k_bounce_begin": (32 + 88, "path_vertices"), "k_bounce_sample": (72 + 48, "path_vertices"),
    "k_bounce_finish": (48 + 52, "path_vertices"), "k_seg_exits": (32 + 72, "neural_queries"),
    "k_seg_apply": (88 + 56, "neural_queries"), "k_query_encode": (48 + 224, "neural_queries"),
    "k_query_fused": (HBM_BYTES_PER_QUERY, "neural_queries"),
    "k_entry_camera": (32, "rays"), "k_primary_finish": (48 + 52, "rays"), "k_finalize": (16 + 12 + 12 + 4, "rays"),
    "k_gt_primary": (88, "rays"), "k_gt_bounce": (72 + 104, "path_vertices"), "k_bounce_all": (120 + 104, "path_vertices"),
}


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="nrb200", choices=["nrb200", "reference"])
    ap.add_argument("--mode", default="neural", choices=["neural", "gt"])
    ap.add_argument("--width", type=int, default=1920)
    ap.add_argument("--height", type=int, default=1080)
    ap.add_argument("--spp", type=int, default=1)
    ap.add_argument("--bounces", type=int, default=3)
    ap.add_argument("--subdiv", type=int, default=7)
    ap.add_argument("--scaling", default="weak", choices=["weak", "strong"],
                    help="weak (default): spp = --spp x GPUs, every GPU keeps the 1-GPU element count; strong: the frame is fixed")
    ap.add_argument("--cpu-seconds", type=float, default=10.0, help="budget of the cpu_baseline sample")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-gpu-baseline", action="store_true")
    ap.add_argument("--gather", default="peer", choices=["peer", "nccl"],
                    help="multi-GPU frame gather: peer-memory stores fused into finalize (default) or an NCCL reduce")
    ap.add_argument("--no-tail-graph", action="store_true", help="fixed nine-iteration segment tail (profiling aid)")
    ap.add_argument("--slices", type=int, default=0, help="NRB_OPT_FRAME_SLICES (0 = the library's automatic choice)")
    ap.add_argument("--pipeline", type=int, default=1, choices=[0, 1],
                    help="neural mode: frame k+1 starts under the end of frame k (NRB_OPT_PIPELINE, the library default)")
    return ap.parse_args()


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return dict(hbm=d["hbm_gbs"], tf_burst=d["bf16_tflops"], tf_sustained=d["bf16_tflops_sustained"], src="measured (MEASURED_PEAKS.json)")
    return dict(hbm=6650.0, tf_burst=1590.0, tf_sustained=1400.0, src="fallback (B200_PROFILING.md)")


def workload_name(a):
    net = "3xHashGrid(L8,F4,T2^14)+SH4->112->128x4->16"
    if a.mode == "neural":
        return (f"neural render {a.width}x{a.height} spp{a.base_spp} bounces{a.bounces}: rays traverse the two 1280-triangle shell hulls "
                f"(+-12% icosphere(3)) around a displaced icosphere (20*4^{a.subdiv} tris, traversed in gt mode only), "
                f"neural query {net} on every shell segment, 2048x1024 fp32 environment")
    return (f"ground-truth BVH path trace {a.width}x{a.height} spp{a.base_spp} bounces{a.bounces}, displaced icosphere 20*4^{a.subdiv} tris, "
            f"2048x1024 fp32 environment")


def config_of(a, n_gpus):
    """The same dict in both arms (the driver compares them)."""
    return {"workload": workload_name(a), "mode": a.mode, "n_gpus": n_gpus, "scaling": a.scaling,
            "spp_per_step": a.spp,
            "partition": "16-row bands round-robin over the GPUs, scene and weights replicated",
            "l2": "per-frame wavefront state (~0.4 KB/path, 0.8 GB at 1080p) exceeds the 126 MB L2 and every step draws new "
                  "samples (progressive accumulation); BVH and hash tables are L2-resident by design"}


def build_scene(nrb, a, builder=None):
    return nrb.scenes.make_scene("ico", a.subdiv, shells=True, env=True, env_size=(2048, 1024), builder=builder)


def setup(nrb, a, scene, lib=None):
    r = nrb.RendererNeural(scene, lib=lib)
    eye, basis = nrb.scenes.default_camera()
    r.resize(a.width, a.height)
    r.set_camera_basis(basis)
    r.set_samples_per_pixel(a.spp)
    r.set_bounce_count(a.bounces)
    if lib is None and getattr(a, "slices", 0):
        r.set_frame_slices(a.slices)
    if lib is None:
        r.set_pipeline(int(getattr(a, "pipeline", 1)))
    if a.mode == "neural":
        r.load_weights(nrb.scenes.make_weights())
        r.set_use_neural_query(True)
        if getattr(a, "no_tail_graph", False) and lib is None:
            r.set_tail_graph(False)

    return r, eye, basis


class ClockSampler:
    """nvidia-smi clocks/throttle reasons sampled during the timed region (B200_PROFILING.md recipe)."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.rows, self.proc = [], None
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(index), f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append((time.time(), [c.strip() for c in line.split(",")]))

    def stop(self, t0, t1):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        rows = [r for (t, r) in self.rows if t0 - 0.05 <= t <= t1 + 0.15 and len(r) >= 8] or [r for _, r in self.rows if len(r) >= 8]
        if not rows:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["no samples"]}
        sm = sorted(float(r[0]) for r in rows)
        reasons = set()
        for r in rows:
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[4:8]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": sm[len(sm) // 2], "sm_max_mhz": float(rows[0][1]), "reasons": sorted(reasons),
                "samples": len(rows), "power_w_max": max(float(r[2]) for r in rows)}


def oracle_renderer(a):
    """The CPU oracle on the bench scene.  Scene BVHs come from the reference's own Mesh::buildBvh (oracle/_ref) when
    that build is present, so this leg never maps libnrb200.so."""
    nrb = importlib.import_module("neural-renderer_b200")
    from oracle import loader
    lib = loader.oracle_lib("auto")
    kind = "reference" if lib.fn("backend")().decode() == "reference" else "port"
    builder = None
    if loader.have_refbvh():
        builder = lambda v, i: loader.ref_build_bvh(v, i)[:2]  # noqa: E731
    scene = build_scene(nrb, a, builder=builder)
    r, eye, _ = setup(nrb, a, scene, lib=lib)
    cores = lib.fn("set_threads")(r.handle, 0)
    return lib, r, eye, kind, int(cores), ("the reference's Mesh::buildBvh" if builder else "nrb_build_bvh")


def sample_rows(lib, r, eye, a, per_frame_seconds, band=8):
    """Choose a uniform sample of 8-row bands (every k-th band of the frame) whose frame costs about
    `per_frame_seconds`; returns (rows covered, stride)."""
    bands = (a.height + band - 1) // band
    stride = max(1, bands // 8)                      # probe: 8 bands spread over the frame
    rows = lib.fn("set_row_sample")(r.handle, band, stride, a.height)
    t0 = time.perf_counter()
    r.render(eye)
    per_row = (time.perf_counter() - t0) / max(1, rows)
    want_rows = max(band, min(a.height, int(per_frame_seconds / max(per_row, 1e-9))))
    stride = max(1, int(round(a.height / want_rows)))
    rows = lib.fn("set_row_sample")(r.handle, band, stride, a.height)
    return rows, stride


def cpu_sample(a, seconds):
    """cpu_baseline: the oracle on a uniform row-band sample of the same workload, one frame of about `seconds`."""
    lib, r, eye, kind, cores, bvh = oracle_renderer(a)
    rows, stride = sample_rows(lib, r, eye, a, seconds)
    r.reset_samples()
    t0 = time.perf_counter()
    r.render(eye)
    dt = time.perf_counter() - t0
    r.close()
    rays = rows * a.width * a.spp
    return {"value": rays / dt / 1e6, "unit": "Mrays/s", "cores": cores, "kind": kind,
            "sample": f"every {stride}-th 8-row band of the frame ({rows} of {a.height} rows, {rays} paths) of the same workload, "
                      f"one frame in {dt:.2f} s; scene BVHs by {bvh}"}


def extbvh_sample(a):
    """cpu_extbvh (BASELINE config 1): ext/bvh's own traversal on the primary rays of the scene mesh, 512x512 and the
    bench resolution (every 4th ray), all cores and one core."""
    from oracle import loader
    if not loader.have_refbvh():
        return None
    nrb = importlib.import_module("neural-renderer_b200")
    from tests.util import primary_rays
    v0, f = nrb.scenes.icosphere(a.subdiv)
    v = nrb.scenes.displace(v0, 0.1)
    out = {"unit": "Mrays/s", "kind": "reference (ext/bvh bvh::v2::Bvh::intersect + PrecomputedTri, as ext/bvh/test/benchmark.cpp)",
           "triangles": int(len(f))}
    for name, (w, h), step in (("512x512", (512, 512), 1), (f"{a.width}x{a.height}", (a.width, a.height), 4)):
        o, d = primary_rays(w, h)
        o, d = o[::step], d[::step]
        allc = loader.ref_extbvh_trace(v, f, o, d, threads=0, reps=3)
        one = loader.ref_extbvh_trace(v, f, o[::8], d[::8], threads=1, reps=3)
        out[name] = {"all_cores": len(o) / allc["trace_ms"] * 1e-3, "cores": allc["threads"],
                     "one_core": len(o[::8]) / one["trace_ms"] * 1e-3, "rays": int(len(o)), "bvh_build_ms": allc["build_ms"]}
    return out


def gpu_baseline(nrb, a, scene, eye, basis):
    """The reference's own CUDA + tiny-cuda-nn build on this GPU, same scene pack."""
    from baseline import refpack
    if not refpack.have("ref_render"):
        return {"unavailable": "baseline/_ref/ref_render not built (needs /root/reference at build time)"}
    w = nrb.scenes.make_weights() if a.mode == "neural" else None
    ref = refpack.ref_render(scene, eye, basis, a.width, a.height, warmup=3, timed=20, spp=a.spp, bounces=a.bounces,
                             neural=a.mode == "neural", weights=w)
    t = ref["timing"]
    return {"value": t["mrays_per_s"], "unit": "Mrays/s", "ms_per_step": t["ms_per_frame"], "host_ms_per_step": t["host_ms_per_frame"],
            "kind": "reference-cuda: the reference's RendererNeural + tiny-cuda-nn compiled for sm_100 with its own flags "
                    "(--use_fast_math), its own Mesh::buildBvh, same scene/camera/weights",
            "protocol": "3 warm-up + 20 timed render() calls, cudaEvents on the legacy default stream"}


def run_reference(a, rank):
    """--impl reference: the reference's CPU implementation of the path (its headers compiled for the host) on all
    host cores; each step renders a uniform sample of row bands of the frame."""
    if rank != 0:
        return
    lib, r, eye, kind, cores, bvh = oracle_renderer(a)
    budget = 100.0 / max(1, a.steps + a.warmup)      # the whole run ends within a few minutes
    rows, stride = sample_rows(lib, r, eye, a, budget)
    for _ in range(a.warmup):
        r.render(eye)
    t0 = time.perf_counter()
    for _ in range(a.steps):
        r.render(eye)
    dt = (time.perf_counter() - t0) / a.steps
    rays = rows * a.width * a.spp
    v = rays / dt / 1e6
    sample = (f"every {stride}-th 8-row band of the frame per step ({rows} of {a.height} rows, {rays} paths), {cores} host threads; "
              f"scene BVHs by {bvh}")
    emit({
        "impl": "reference", "metric": "neural-render Mrays/s at 1080p", "value": v, "unit": "Mrays/s", "n_gpus": a.gpus,
        "steps": a.steps, "warmup": a.warmup, "ms_per_step": dt * 1e3, "higher_is_better": True,
        "scaling": a.scaling, "vs_baseline": None, "dtype": "f32 geometry + f16 network", "data": "synthetic",
        "config": config_of(a, a.gpus),
        "cpu_baseline": {"value": v, "unit": "Mrays/s", "cores": cores, "kind": kind, "sample": sample},
        "e2e": {"value": v, "unit": "Mrays/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    })


_REAL_STDOUT = None


def emit(line):
    """The ONE JSON line of the contract, on the process's real stdout."""
    os.write(_REAL_STDOUT if _REAL_STDOUT is not None else 1, (json.dumps(line) + "\n").encode())


def kernel_roofline(name, ms, counters, rays, pk):
    """HBM roofline of a non-tensor kernel from its algorithmic bytes per unit (KERNEL_BYTES)."""
    if name not in KERNEL_BYTES or not ms:
        return None
    per, unit = KERNEL_BYTES[name]
    units = rays if unit == "rays" else counters[unit]
    gbs = units * per / (ms * 1e-3) / 1e9
    return {"kernel": name, "bound": "hbm", "achieved": gbs, "peak": pk["hbm"], "unit": "GB/s", "frac": gbs / pk["hbm"],
            "algorithmic_bytes_per_unit": per, "unit_of_work": unit, "units_per_step": units, "ms_per_step": ms}


def main():
    # Libraries write to fd 1 behind Python's back (NCCL prints its version banner there): everything but
    # the result line goes to stderr, so stdout carries exactly one JSON line.
    global _REAL_STDOUT
    sys.stdout.flush()
    _REAL_STDOUT = os.dup(1)
    os.dup2(2, 1)
    a = parse()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    # Weak scaling (default): the frame is sharded by row bands and the samples per pixel grow with the GPU count, so
    # every GPU keeps the element count of the 1-GPU frame.  Strong scaling: the frame (W x H x spp) is fixed.
    n_scale = max(world, a.gpus if a.impl == "reference" else 1)
    a.base_spp = a.spp
    if n_scale > 1 and a.scaling == "weak":
        a.spp = a.spp * n_scale
    if a.impl == "reference":
        run_reference(a, rank)
        return

    import torch
    import torch.distributed as dist
    assert torch.cuda.is_available(), "bench.py needs a CUDA device (there is no CPU path)"
    torch.cuda.set_device(local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    nrb = importlib.import_module("neural-renderer_b200")
    scene = build_scene(nrb, a)
    r, eye, basis = setup(nrb, a, scene)
    band = 16
    if world > 1:
        r.set_tile(band, rank, world)
    stream = torch.cuda.ExternalStream(r.stream_ptr(), device=torch.device("cuda", local_rank))
    rays_per_step = a.width * a.height * a.spp

    class Pixels:  # zero-copy torch view of devicePixels()
        def __init__(self, ptr, n):
            self.__cuda_array_interface__ = {"shape": (n,), "typestr": "|u1", "data": (ptr, False), "version": 2}

    def frame_tensor():
        return torch.as_tensor(Pixels(r.device_pixels(), a.width * a.height * 4), device=torch.device("cuda", local_rank))

    gathered = None
    # The one exchange step of the path: every rank owns row bands and rank 0 ends up with the whole frame.
    # Default: fused into the frame - each rank's finalize kernel stores its bands into rank 0's frame buffer
    # over NVLink peer memory (nrb_gather_export / nrb_gather_attach) and rank 0's render ends with the wait
    # for their signals.  --gather nccl keeps the collective (a sum-reduce of the zero-padded frames).
    peer_gather = world > 1 and a.gather == "peer"
    if peer_gather:
        ok = 1
        box = [None]
        try:
            if rank == 0:
                box[0] = r.gather_export(world - 1)
        except Exception as e:  # noqa: BLE001
            print(f"[bench] peer gather unavailable on rank 0: {e}", file=sys.stderr)
            ok = 0
        dist.broadcast_object_list(box, src=0)
        if rank != 0:
            try:
                if box[0] is None:
                    raise RuntimeError("no handle from rank 0")
                r.gather_attach(box[0])
            except Exception as e:  # noqa: BLE001
                print(f"[bench] peer gather unavailable on rank {rank}: {e}", file=sys.stderr)
                ok = 0
        flag = torch.tensor([ok], device="cuda")
        dist.all_reduce(flag, op=dist.ReduceOp.MIN)
        if int(flag.item()) == 0:  # every rank agrees to fall back to the collective
            r.gather_detach()
            peer_gather = False

    def gather():
        nonlocal gathered
        if world == 1 or peer_gather:
            return
        t = frame_tensor()
        if gathered is None:
            gathered = torch.empty_like(t)
        gathered.copy_(t)  # non-owned rows are zero on every rank, so a sum over ranks is the gather
        stream.wait_stream(torch.cuda.current_stream())  # the next frame's finalize must not overtake this read
        dist.reduce(gathered, dst=0, op=dist.ReduceOp.SUM)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(steps, body, finish=None):
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0 = time.time()
        e0.record(torch.cuda.current_stream())
        for _ in range(steps):
            body()
        if finish:
            finish()  # e2e: the frame read-backs run on the renderer's copy stream; wait for the last of them
        e1.record(torch.cuda.current_stream())  # render() joins its stream into the legacy default stream
        barrier()
        ms = e0.elapsed_time(e1)
        if world > 1:
            t = torch.tensor([ms], device="cuda")
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t.item())
        return ms, t0, time.time()

    def step_device():
        r.render(eye)
        gather()

    # e2e: the step's host inputs (camera position + basis + fov, 52 bytes: what viewer / evaluate hand over per frame)
    # live in pinned host memory and are read from there every step; the frame goes back to pinned host memory every step.
    cam_host = torch.empty(16, dtype=torch.float32).pin_memory()
    cam_host[:3] = torch.tensor([float(x) for x in eye])
    cam_host[3:6] = torch.tensor([float(x) for x in basis.forward])
    cam_host[6:9] = torch.tensor([float(x) for x in basis.right])
    cam_host[9:12] = torch.tensor([float(x) for x in basis.up])
    cam_host[12] = float(basis.fov_y)
    host_frames = [torch.empty(a.width * a.height * 4, dtype=torch.uint8).pin_memory() for _ in range(2)]
    e2e_step = [0]

    def step_e2e():
        """One frame through the public API with host buffers on both sides: camera in, RGBA frame out.  The
        device->host read of frame k is enqueued right behind it (nrb_read_pixels_async: second stream, double-buffered
        pinned destination) and so travels under the kernels of frame k+1; the closing barrier of the timed region
        waits for every transfer."""
        c = cam_host.numpy()
        r.set_camera_basis(nrb.RenderBasis(c[3:6], c[6:9], c[9:12], float(c[12])))
        r.render(c[:3])
        gather()
        dst = host_frames[e2e_step[0] & 1]
        e2e_step[0] += 1
        if world == 1 or (peer_gather and rank == 0):
            r.read_pixels_async(dst.data_ptr(), dst.numel())
        elif rank == 0:
            dst.copy_(gathered, non_blocking=True)

    for _ in range(max(3, a.warmup)):
        step_device()
    sampler = ClockSampler(local_rank) if rank == 0 else None
    ms, t0, t1 = timed(a.steps, step_device)
    clocks = sampler.stop(t0, t1) if sampler else None
    counters = r.counters()
    for _ in range(2):
        step_e2e()
    ms_e2e, _, _ = timed(a.steps, step_e2e, finish=r.sync)

    # second pass with per-launch events -> per-kernel times
    r.set_profile(True)
    r.reset_stages()
    for _ in range(a.steps):
        r.render(eye)
    r.sync()
    stages = r.stages()
    r.set_profile(False)
    total_stage_ms = sum(v[0] for v in stages.values()) or 1.0
    neural_q = counters["neural_queries"]
    launches_per_step = counters["kernels"]

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    pk = peaks()
    value = rays_per_step * a.steps / (ms * 1e-3) / 1e6
    e2e = rays_per_step * a.steps / (ms_e2e * 1e-3) / 1e6
    local_rays = rays_per_step // world
    per_kernel = {k: {"ms_per_step": v[0] / a.steps, "launches_per_step": v[1] / a.steps, "share": v[0] / total_stage_ms}
                  for k, v in sorted(stages.items(), key=lambda kv: -kv[1][0])}
    stage_ms = lambda name: stages[name][0] / a.steps if name in stages else None  # noqa: E731
    traffic_db = {}
    tpath = os.path.join(ROOT, "profiles", "roofline_traffic.json")
    if os.path.exists(tpath):
        traffic_db = json.load(open(tpath))

    # ---- roofline: the kernel with the largest share of the frame, against the roofline that bounds it -------------
    top = max(stages.items(), key=lambda kv: kv[1][0])[0]
    tensor_kernels = [k for k in stages if k in ("k_query_mlp", "k_query_fused", "k_neural_query_tc", "k_neural_query")]
    mlp = None
    if tensor_kernels and a.mode == "neural" and neural_q:
        mk = max(tensor_kernels, key=lambda k: stages[k][0])
        m_ms = stage_ms(mk)
        tf = neural_q * FLOPS_PER_QUERY / (m_ms * 1e-3) / 1e12
        tr = traffic_db.get(mk)
        # the timed region is tens of milliseconds at full clocks, not a power-limited long step: the BURST peak applies
        mlp = {"kernel": mk, "bound": "tensor", "achieved": tf, "peak": pk["tf_burst"], "unit": "TFLOP/s", "frac": tf / pk["tf_burst"],
               "frac_of_sustained_peak": tf / pk["tf_sustained"],
               "traffic": (tr["dram_bytes_per_query"] * neural_q / max(1.0, stages[mk][1] / a.steps)) if tr and "dram_bytes_per_query" in tr else None,
               "peak_source": pk["src"] + ": burst cuBLAS bf16 (fp16 runs at the same pipe rate)",
               "queries_per_step": neural_q, "flops_per_query": FLOPS_PER_QUERY, "ms_per_step": m_ms,
               "launches_per_step": stages[mk][1] / a.steps, "share_of_frame": stages[mk][0] / total_stage_ms}
        qkeys = [k for k in stages if k.startswith("k_query_") or k.startswith("k_neural_query")]
        q_ms = sum(stages[k][0] for k in qkeys) / a.steps
        mlp["whole_query"] = {"kernels": sorted(qkeys), "ms_per_step": q_ms, "Gqueries_per_s": neural_q / (q_ms * 1e-3) / 1e9,
                              "tensor_TFLOPs": neural_q * FLOPS_PER_QUERY / (q_ms * 1e-3) / 1e12,
                              "hbm_gbs": neural_q * HBM_BYTES_PER_QUERY / (q_ms * 1e-3) / 1e9}
    if mlp and top == mlp["kernel"]:
        roof = dict(mlp)
    else:
        roof = kernel_roofline(top, stage_ms(top), counters, local_rays, pk) or \
            {"kernel": top, "bound": "hbm", "achieved": None, "peak": pk["hbm"], "unit": "GB/s", "frac": None, "ms_per_step": stage_ms(top)}
        tr = traffic_db.get(top)
        roof["traffic"] = tr.get("dram_bytes_per_launch") if tr else None
        roof["peak_source"] = pk["src"] + ": HBM copy bandwidth"
        roof["note"] = ("a traversal / shading stage: its BVH is L1/L2-resident, so the HBM roofline it is measured against is far "
                        "away by construction; ncu (profiles/) shows what bounds it - instruction issue and load latency "
                        "(active lanes per instruction and issue-slot utilisation are listed there)")
    roof["share_of_frame"] = stages[top][0] / total_stage_ms
    roof["top_kernel_by_time"] = top
    if mlp and top != mlp["kernel"]:
        roof["mlp"] = mlp
    others = {}
    for k in list(per_kernel)[:6]:
        kr = kernel_roofline(k, stage_ms(k), counters, local_rays, pk)
        if kr and k != top:
            others[k] = {"achieved_GBs": kr["achieved"], "frac_of_hbm_peak": kr["frac"], "algorithmic_bytes_per_unit": kr["algorithmic_bytes_per_unit"]}
    roof["other_kernels_hbm"] = others

    line = {
        "metric": "neural-render Mrays/s at 1080p", "value": value, "unit": "Mrays/s", "n_gpus": world, "steps": a.steps,
        "warmup": max(3, a.warmup), "ms_per_step": ms / a.steps, "higher_is_better": True, "scaling": a.scaling,
        "vs_baseline": None, "dtype": "f32 geometry + f16 network (f32 accumulate)", "data": "synthetic",
        "config": config_of(a, world),
        "frames": {"pipelined": bool(a.pipeline), "slices": a.slices or "auto",
                   "note": "NRB_OPT_PIPELINE: frame k+1 starts on a second stream and state buffer under the end of frame k; "
                           "all K frames complete inside the timed region"},
        "gather": ("none (1 GPU)" if world == 1 else ("finalize stores into rank 0's frame over NVLink peer memory" if peer_gather else "NCCL sum-reduce of zero-padded frames")),
        "clocks": clocks,
        "e2e": {"value": e2e, "unit": "Mrays/s", "h2d_bytes_per_step": 52, "d2h_bytes_per_step": a.width * a.height * 4,
                "ms_per_step": ms_e2e / a.steps,
                "note": "host inputs of a step = camera position + basis + fov (13 floats, read from pinned host memory each step; the "
                        "scene is resident, as in the reference); output = the RGBA frame into pinned host memory each step"},
        "gpu_launches": int(launches_per_step) * a.steps,
        "launches_per_step": int(launches_per_step),
        "roofline": roof,
        "kernels": per_kernel,
        "per_step": {"ray_segments": counters["ray_segments"], "neural_queries": neural_q, "path_vertices": counters["path_vertices"],
                     "segments_per_s_M": counters["ray_segments"] * a.steps / (ms * 1e-3) / 1e6},
    }
    if world == 1:
        r.close()
        del r
        torch.cuda.synchronize()
        if not a.no_gpu_baseline:
            try:
                line["gpu_baseline"] = gpu_baseline(nrb, a, scene, eye, basis)
                if "value" in line["gpu_baseline"]:
                    line["gpu_baseline"]["speedup_device"] = value / line["gpu_baseline"]["value"]
                    line["gpu_baseline"]["speedup_e2e"] = e2e / line["gpu_baseline"]["value"]
            except Exception as e:  # noqa: BLE001
                line["gpu_baseline"] = {"unavailable": f"{type(e).__name__}: {e}"[:300]}
        if not a.no_cpu_baseline:
            line["cpu_baseline"] = cpu_sample(a, a.cpu_seconds)
            try:
                ext = extbvh_sample(a)
                if ext:
                    line["cpu_extbvh"] = ext
            except Exception as e:  # noqa: BLE001
                line["cpu_extbvh"] = {"unavailable": f"{type(e).__name__}: {e}"[:300]}
    emit(line)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
