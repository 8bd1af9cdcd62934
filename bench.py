#!/usr/bin/env python
"""Benchmark of the image->voxel hot path (BASELINE.json's metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload NAME]
    python bench.py --impl reference ...        # the reference's CPU path (port)

One "step" = one pass of the hot path over one batch of synthetic scenes of
the named configuration.  Default workload: ``imvoxelnet_kitti`` -- 1280x384
monocular, 216x248x12 voxels, 64 channels, batch 8 (BASELINE.json configs[1]).
For N > 1 every rank processes its own batch of scenes (plain data-parallel,
no collective on the data path: "scaling": "weak"); ``--workload
scannet_fast`` instead shards the 50 views of each scene over the ranks and
all-reduces the partial sums/counts over NCCL.

Prints ONE JSON line (rank 0).  See DESIGN.md "Measurement" for every key.
"""
import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

from imvoxelnet_b200 import scenes  # noqa: E402
from imvoxelnet_b200 import projection as pj  # noqa: E402

WORKLOADS = {
    # name -> (scene builder key, scenes per rank)
    "kitti": ("kitti", 8),
    "scannet_fast": ("scannet_fast", 8),
    "sunrgbd_fast": ("sunrgbd_fast", 8),
    "nuscenes": ("nuscenes", 8),
    "total_sunrgbd": ("total_sunrgbd", 32),
}
METRIC = "backproject_voxel_views_per_s"
UNIT = "voxel-views/s"


def load_peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.isfile(path):
        with open(path) as f:
            return json.load(f), "measured"
    return {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0}, "fallback"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons while the timed region runs."""
    QUERY = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
             "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
             "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.rows, self.proc, self.index = [], None, index

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.QUERY}",
                 "--format=csv,noheader,nounits", "-lms", "100"],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        self.thread.join(timeout=2)
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            if len(r) < 7:
                continue
            try:
                sm.append(float(r[0]))
                mx.append(float(r[1]))
            except ValueError:
                continue
            for name, flag in zip(names, r[3:7]):
                if flag.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": statistics.median(sm) if sm else None,
                "sm_max_mhz": max(mx) if mx else None, "reasons": sorted(reasons),
                "samples": len(sm)}


def scene_setup(name):
    key, batch = WORKLOADS[name]
    sc = scenes.build(key)
    P = pj.compute_projection(sc.img_meta, scenes.STRIDE)
    origin = sc.img_meta['lidar2img']['origin']
    return sc, batch, P, origin


def config_of(sc, batch, world, extra=None):
    hf, wf = sc.map_hw
    cfg = {"workload": sc.name, "scenes_per_gpu": batch, "views": sc.n_views,
           "n_voxels": list(sc.n_voxels), "channels": sc.channels, "feature_map": [hf, wf],
           "global_scenes": batch * world, "stages": ["backproject+view-mean"]}
    cfg.update(extra or {})
    return cfg


# --------------------------------------------------------------------------
# the reference arm: the reference's CPU path (torch port), all host threads
# --------------------------------------------------------------------------
def cpu_reference_time(sc, P, origin, repeats, warmup, seed=0):
    """Seconds per scene of the reference's torch-CPU path (best of repeats)."""
    from oracle import torch_port
    torch.set_num_threads(os.cpu_count() or 1)
    full = scenes.random_features(sc, seed=seed, channels_last=False)
    h, w = sc.feat_hw
    feats = full[:, :, :h, :w]
    pts = pj.get_points(sc.n_voxels, sc.voxel_size, origin)
    times = []
    with torch.no_grad():
        for i in range(warmup + repeats):
            t0 = time.perf_counter()
            torch_port.scene_volume(feats, pts, P)
            dt = time.perf_counter() - t0
            if i >= warmup:
                times.append(dt)
    return times


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    sc, batch, P, origin = scene_setup(args.workload)
    # a step = ONE scene of the batch (bounded sample of the workload)
    times = cpu_reference_time(sc, P, origin, repeats=args.steps, warmup=max(1, min(args.warmup, 2)))
    t = sum(times) / len(times)
    value = sc.voxel_views / t
    cores = os.cpu_count() or 1
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": t * 1e3,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
        "data": "synthetic",
        "config": config_of(sc, batch, 1, {"note": "one scene per step (bounded sample of the batch)"}),
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": "port",
                         "sample": f"{args.steps} x 1 scene of {sc.name}, torch {torch.__version__} "
                                   f"CPU ops of the reference's backproject + view mean "
                                   f"(oracle/torch_port.py), {cores} threads"},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    print(json.dumps(line), flush=True)


# --------------------------------------------------------------------------
# our arm
# --------------------------------------------------------------------------
def algorithmic_bytes(sc, P, origin, device):
    """B_algo per scene: U*C*4 + N*C*4 + N + V*48, with U the
    number of distinct feature pixels hit by valid voxel-views (exact)."""
    from imvoxelnet_b200 import ops
    h, w = sc.feat_hw
    x, y, valid = ops.project_indices(None, P, h, w, n_voxels=sc.n_voxels, voxel_size=sc.voxel_size,
                                      origin=origin, device=device)
    v = torch.arange(sc.n_views, device=device).view(-1, 1).expand_as(x)
    key = ((v * h + y) * w + x)[valid]
    U = int(torch.unique(key).numel())
    n_valid = int(valid.sum())
    N, C, V = sc.n_total_voxels, sc.channels, sc.n_views
    return U * C * 4 + N * C * 4 + N + V * 48, U, n_valid


def run_ours(args):
    import torch.distributed as dist
    from imvoxelnet_b200 import _lib, ops, sharding

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device (no CPU fallback)")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    _lib.lib()  # fail loudly if the CUDA library is missing

    sc, batch, P, origin = scene_setup(args.workload)
    view_sharded = args.workload == "scannet_fast" and world > 1
    h, w = sc.feat_hw
    N, C, V = sc.n_total_voxels, sc.channels, sc.n_views
    b_algo, U, n_valid = algorithmic_bytes(sc, P, origin, dev)

    if view_sharded:
        my_views = list(range(rank, V, world))
    else:
        my_views = list(range(V))
    Vr = len(my_views)

    # rotating device input sets (> L2 in total, so no step finds its map in L2
    # from the previous one; the 1.3 GB of output per step evicts it anyway)
    n_sets = 3
    host = scenes.random_features(sc, seed=rank, batch=batch, channels_last=True)
    if view_sharded:
        host = _channels_last_5d(host[:, my_views])
    host_pinned = _pinned_like(host)  # keeps the channels-last physical layout
    host_pinned.copy_(host)
    dev_sets = [host_pinned.to(dev) for _ in range(n_sets)]
    for i, d in enumerate(dev_sets):
        d.add_(float(i))  # different contents, same layout
    Pb = P[my_views].unsqueeze(0).expand(batch, Vr, 3, 4).contiguous()
    P_dev = Pb.to(dev)
    hws = [(h, w)] * batch
    origins = [origin] * batch

    def step(i):
        if view_sharded:
            return sharding.backproject_scenes_view_sharded(
                dev_sets[i % n_sets], hws, P_dev, sc.n_voxels, sc.voxel_size, origins)
        return ops.backproject_scenes(dev_sets[i % n_sets], hws, P_dev, sc.n_voxels, sc.voxel_size,
                                      origins)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for i in range(args.warmup):
        step(i)
    barrier()

    # ---- device-resident timing ----
    # The K steps are captured once into a CUDA graph and replayed, so that the
    # timed region holds K back-to-back passes with no host launch gaps (a step
    # of the small configurations is shorter than one Python call).  With a
    # collective in the step (view-sharded mode) the steps are issued eagerly.
    sampler = ClockSampler(local)
    use_graph = not view_sharded
    launches0 = ops.launch_count()
    if use_graph:
        graph = torch.cuda.CUDAGraph()
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            with torch.cuda.graph(graph, stream=side):
                for i in range(args.steps):
                    step(i)
        torch.cuda.current_stream().wait_stream(side)
        launches = ops.launch_count() - launches0   # kernels captured = launched per replay
        graph.replay()                              # untimed: first replay uploads the graph
        barrier()
    if rank == 0:
        sampler.start()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    repeats = 5 if use_graph else 1
    totals = []
    for _ in range(repeats):
        barrier()
        ev0.record()
        if use_graph:
            graph.replay()
        else:
            for i in range(args.steps):
                step(i)
        ev1.record()
        barrier()
        totals.append(ev0.elapsed_time(ev1))
    if not use_graph:
        launches = ops.launch_count() - launches0
    total_ms = statistics.median(totals)            # of 5 timed runs of exactly K steps each
    t = torch.tensor([total_ms], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    total_ms = float(t.item())
    ms_per_step = total_ms / args.steps
    clocks = sampler.stop() if rank == 0 else None
    # eager issue of the same K steps, for the record (host-launch bound when short)
    barrier()
    ev0.record()
    for i in range(args.steps):
        step(i)
    ev1.record()
    barrier()
    eager_ms_per_step = ev0.elapsed_time(ev1) / args.steps

    scenes_global = batch * (1 if view_sharded else world)
    units_per_step = scenes_global * V * N
    value = units_per_step / (ms_per_step * 1e-3)

    # dominant kernel: the fused back-projection; its launch duration is the
    # step (one launch per step) unless the step also holds the collective
    kern_ms = ms_per_step if not view_sharded else None
    peaks, peak_kind = load_peaks()
    if kern_ms is not None:
        achieved = b_algo * batch / (kern_ms * 1e-3) / 1e9
        roofline = {"bound": "hbm", "achieved": achieved, "peak": peaks["hbm_gbs"], "unit": "GB/s",
                    "frac": achieved / peaks["hbm_gbs"], "traffic": _ncu_traffic(sc.name),
                    "kernel": "bp_vec_kernel", "peak_kind": peak_kind,
                    "algorithmic_bytes_per_launch": b_algo * batch,
                    "kernel_ms": kern_ms, "distinct_pixels_per_scene": U,
                    "valid_voxel_views_per_scene": n_valid}
    else:
        roofline = None

    # ---- end to end: host buffers in, host buffers out, every step ----
    out0 = step(0)
    vol_host = _pinned_like(out0[0])
    val_host = _pinned_like(out0[1])
    h2d = host_pinned.numel() * 4 + Pb.numel() * 4
    d2h = vol_host.numel() * vol_host.element_size() + val_host.numel()
    P_pinned = Pb.pin_memory()
    e2e_steps = max(3, min(args.steps, 10))

    def e2e_step():
        feats = host_pinned.to(dev, non_blocking=True)
        pd = P_pinned.to(dev, non_blocking=True)
        if view_sharded:
            out = sharding.backproject_scenes_view_sharded(feats, hws, pd, sc.n_voxels,
                                                           sc.voxel_size, origins)
        else:
            out = ops.backproject_scenes(feats, hws, pd, sc.n_voxels, sc.voxel_size, origins)
        vol_host.copy_(out[0], non_blocking=True)
        val_host.copy_(out[1], non_blocking=True)

    e2e_step()
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(e2e_steps):
        e2e_step()
    e1.record()
    barrier()
    t = torch.tensor([e0.elapsed_time(e1)], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    e2e_ms = float(t.item()) / e2e_steps
    e2e = {"value": units_per_step / (e2e_ms * 1e-3), "unit": UNIT, "h2d_bytes_per_step": h2d,
           "d2h_bytes_per_step": d2h, "ms_per_step": e2e_ms, "steps": e2e_steps,
           "api": "imvoxelnet_b200.ops.backproject_scenes (ivx_backproject_scenes), pinned host "
                  "features + projections in, fp32 volume + valid mask out"}

    if rank == 0:
        cpu = None
        if world == 1:
            times = cpu_reference_time(sc, P, origin, repeats=3, warmup=1)
            tcpu = min(times)
            cores = os.cpu_count() or 1
            cpu = {"value": sc.voxel_views / tcpu, "unit": UNIT, "cores": cores, "kind": "port",
                   "sample": f"best of 3 x 1 scene of {sc.name} (torch-CPU port of the reference's "
                             f"backproject + view mean, oracle/torch_port.py)",
                   "ms_per_scene": tcpu * 1e3}
            try:
                from oracle import backproject as ob
                full = scenes.random_features(sc, seed=0, channels_last=False).numpy()
                pts = ob.get_points(sc.n_voxels, sc.voxel_size, origin)
                t0 = time.perf_counter()
                ob.backproject_mean(full[:, :, :h, :w], pts, P.numpy())
                cpu["c_oracle_openmp_value"] = sc.voxel_views / (time.perf_counter() - t0)
            except Exception as exc:  # the C oracle is optional for the bench
                cpu["c_oracle_openmp_value"] = f"unavailable: {exc}"
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms_per_step, "higher_is_better": True,
            "scaling": "strong" if view_sharded else "weak", "vs_baseline": None, "dtype": "f32",
            "data": "synthetic",
            "config": config_of(sc, batch, world, {
                "parallelism": (f"views sharded over {world} ranks + NCCL all-reduce" if view_sharded
                                else f"scenes data-parallel x{world}, no collective"),
                "l2": f"{n_sets} rotating input sets + {d2h / 1e6:.0f} MB written per step (> 126 MB L2)",
                "layout": "features NHWC fp32, volume NDHWC fp32"}),
            "scenes_per_s": scenes_global / (ms_per_step * 1e-3),
            "timing": {"mode": "cuda-graph replay of K steps, median of 5" if use_graph else "eager",
                       "eager_ms_per_step": eager_ms_per_step},
            "roofline": roofline, "cpu_baseline": cpu, "e2e": e2e, "clocks": clocks,
            "gpu_launches": launches, "lib": _lib.version(),
        }
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def _channels_last_5d(x):
    """[B,V,C,H,W] tensor re-laid so that every [C,H,W] image is channels-last."""
    b, v, c, h, w = x.shape
    y = x.reshape(b * v, c, h, w).contiguous(memory_format=torch.channels_last)
    return y.as_strided((b, v, c, h, w), (v * y.stride(0),) + tuple(y.stride()))


def _pinned_like(t):
    """Pinned host tensor with the same shape and (dense) strides as ``t``."""
    flat = torch.empty(t.numel(), dtype=t.dtype).pin_memory()
    return flat.as_strided(t.shape, t.stride())


def _ncu_traffic(workload):
    """dram bytes per launch from the committed ncu capture of this workload
    (profiles/*.json written by tools/ncu_summary.py), or None."""
    path = os.path.join(ROOT, "profiles", f"traffic_{workload}.json")
    if os.path.isfile(path):
        with open(path) as f:
            return json.load(f).get("dram_bytes_per_launch")
    return None


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="kitti", choices=sorted(WORKLOADS))
    args = ap.parse_args()
    if args.warmup < 3 and args.impl == "ours":
        args.warmup = 3
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
