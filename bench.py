#!/usr/bin/env python
"""Benchmark of the image->voxel hot path (BASELINE.json's metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload NAME]
    python bench.py --impl reference ...        # the reference's CPU path (port)

One "step" = one pass of the hot path -- fused back-projection + view mean,
then the configuration's 3D neck -- over one batch of synthetic scenes of the
named configuration.  Default workload: ``imvoxelnet_kitti`` -- 1280x384
monocular, 216x248x12 voxels, 64 channels, batch 8 (BASELINE.json configs[1]).
For N > 1 every rank processes its own batch of scenes (plain data-parallel,
no collective on the data path: "scaling": "weak").  The same line carries, under
``view_sharded``, the strong-scaling figures of the multi-view configurations
(50-view ScanNet, 6-camera nuScenes) whose VIEWS are sharded over the N ranks
with the reduction fused into the finalize kernel; ``--workload scannet_fast``
with N > 1 makes that mode the headline of the line instead.

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
    "nuscenes_refgrid": ("nuscenes_refgrid", 4),   # the reference's own 312x312x12 grid (neck-compatible)
    "total_sunrgbd": ("total_sunrgbd", 32),
}
METRIC = "voxel_views_per_s"   # through the whole hot path: back-projection + 3D neck
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


NECKS = {
    # workload -> (neck class, constructor kwargs, dense GFLOP per scene: SURVEY.md appendix A)
    "kitti": ("KittiImVoxelNeck", dict(in_channels=64, out_channels=256), 2603.37),
    "scannet_fast": ("FastIndoorImVoxelNeck", dict(in_channels=256, n_blocks=[1, 1, 1], out_channels=128), 489.79),
    "sunrgbd_fast": ("FastIndoorImVoxelNeck", dict(in_channels=256, n_blocks=[1, 1, 1], out_channels=128), 489.79),
    "total_sunrgbd": ("ImVoxelNeck", dict(channels=[64, 128, 256, 512], out_channels=64,
                                          down_layers=[1, 2, 3, 4], up_layers=[3, 2, 1], conditional=False), 631.66),
    "nuscenes": None,   # BASELINE's 200x200x4 grid: every reference neck rejects Z=4
    "nuscenes_refgrid": ("NuScenesImVoxelNeck", dict(in_channels=64, out_channels=256), 1571.76),
}


def randomize_norms(module, seed=0):
    """Random BN affine / running statistics and non-zero bn2 weights, so the
    random-init neck is not degenerate (hard parts)."""
    g = torch.Generator().manual_seed(seed)
    for m in module.modules():
        if isinstance(m, torch.nn.BatchNorm3d):
            with torch.no_grad():
                m.running_mean.copy_(torch.randn(m.num_features, generator=g) * 0.1)
                m.running_var.copy_(torch.rand(m.num_features, generator=g) + 0.5)
                m.weight.copy_(torch.rand(m.num_features, generator=g) + 0.5)
                m.bias.copy_(torch.randn(m.num_features, generator=g) * 0.1)
    return module


# --------------------------------------------------------------------------
# the reference arm: the reference's CPU path (torch port), all host threads
# --------------------------------------------------------------------------
def reference_modules():
    """(scene_volume, neck classes, kind, description) of the CPU arm: the UNMODIFIED reference files
    (mmdet3d/models/detectors/imvoxelnet.py, necks/imvoxelnet.py, imported by oracle/ref_loader.py) when a
    copy of the reference tree is on this machine, else the op-for-op port pinned to them."""
    from oracle import ref_loader, torch_necks, torch_port
    for root in (os.environ.get("IVX_REFERENCE_ROOT"), "/root/reference", os.path.join(ROOT, "baseline", "_ref")):
        if root and os.path.isfile(os.path.join(root, "mmdet3d/models/detectors/imvoxelnet.py")):
            ref_loader.REFERENCE_ROOT = root
            det, nk = ref_loader.load()

            def scene_volume(features, points, projection):   # detectors/imvoxelnet.py:69-74, verbatim statements
                volume, valid = det.backproject(features, points, projection)
                volume = volume.sum(dim=0)
                valid = valid.sum(dim=0)
                volume = volume / valid
                valid = valid > 0
                volume[:, ~valid[0]] = .0
                return volume, valid

            return scene_volume, (lambda name: getattr(nk, name)), "reference", \
                f"the unmodified reference files under {root} (oracle/ref_loader.py)"
    return torch_port.scene_volume, (lambda name: torch_necks.NECKS[name]), "port", \
        "torch-CPU port of the reference (oracle/torch_port.py, oracle/torch_necks.py; no reference tree here)"


def cpu_reference_time(sc, P, origin, workload, repeats, warmup, with_neck, seed=0):
    """Seconds per scene of the reference's CPU path: backproject + view mean and, if asked,
    the neck, on all host threads.  Returns (times, kind, description)."""
    scene_volume, neck_class, kind, what = reference_modules()
    torch.set_num_threads(os.cpu_count() or 1)
    full = scenes.random_features(sc, seed=seed, channels_last=False)
    h, w = sc.feat_hw
    feats = full[:, :, :h, :w]
    pts = pj.get_points(sc.n_voxels, sc.voxel_size, origin)
    neck = None
    if with_neck and NECKS[workload] is not None:
        name, kw, _ = NECKS[workload]
        torch.manual_seed(0)
        neck = randomize_norms(neck_class(name)(**kw).eval())
    times = []
    with torch.no_grad():
        for i in range(warmup + repeats):
            t0 = time.perf_counter()
            vol, _ = scene_volume(feats, pts, P)
            if neck is not None:
                neck(vol[None])
            dt = time.perf_counter() - t0
            if i >= warmup:
                times.append(dt)
    return times, kind, what


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    sc, batch, P, origin = scene_setup(args.workload)
    with_neck = not args.no_neck and NECKS[args.workload] is not None
    # a step = ONE scene of the batch (bounded sample of the workload)
    times, kind, what = cpu_reference_time(sc, P, origin, args.workload, repeats=args.steps,
                                           warmup=1 if with_neck else max(1, min(args.warmup, 2)),
                                           with_neck=with_neck)
    t = sum(times) / len(times)
    value = sc.voxel_views / t
    cores = os.cpu_count() or 1
    stages = ["backproject+view-mean"] + (["3d-neck"] if with_neck else [])
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": t * 1e3,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
        "data": "synthetic",
        "config": config_of(sc, batch, 1, {"stages": stages,
                                           "note": "one scene per step (bounded sample of the batch)"}),
        "scenes_per_s": 1.0 / t,
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": kind,
                         "sample": f"{args.steps} x 1 scene of {sc.name}: {' + '.join(stages)} by {what}, "
                                   f"torch {torch.__version__} CPU, {cores} threads"},
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


def timed_graph(fn, steps, barrier, repeats=5):
    """Capture ``steps`` calls of fn(i) in a CUDA graph; median ms of ``repeats``
    timed replays (each exactly ``steps`` steps, synchronised on both sides)."""
    graph = torch.cuda.CUDAGraph()
    side = torch.cuda.Stream()
    side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side):
        with torch.cuda.graph(graph, stream=side):
            for i in range(steps):
                fn(i)
    torch.cuda.current_stream().wait_stream(side)
    graph.replay()      # untimed: the first replay uploads the graph
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    totals = []
    for _ in range(repeats):
        barrier()
        ev0.record()
        graph.replay()
        ev1.record()
        barrier()
        totals.append(ev0.elapsed_time(ev1))
    return statistics.median(totals), graph


def pin_to_gpu_cpus(index):
    """Bind this process (and the pinned buffers it touches first) to the CPUs NVML reports as
    nearest to GPU `index`.  Returns the number of CPUs in the set, or None when unavailable."""
    try:
        import pynvml
        pynvml.nvmlInit()
        handle = pynvml.nvmlDeviceGetHandleByIndex(index)
        words = pynvml.nvmlDeviceGetCpuAffinity(handle, (os.cpu_count() + 63) // 64)
        cpus = {64 * i + b for i, w in enumerate(words) for b in range(64) if (int(w) >> b) & 1}
        cpus &= os.sched_getaffinity(0)
        if cpus:
            os.sched_setaffinity(0, cpus)
            return len(cpus)
    except Exception:
        pass
    return None


def tensor_peak(peaks, timed_region_ms):
    """Measured bf16 tensor peak to compare with: the burst figure for a timed region shorter than
    about a second (it never reaches the power-capped steady state the sustained figure was taken
    in), the sustained one from 2 s on."""
    burst = peaks["bf16_tflops"]
    sustained = peaks.get("bf16_tflops_sustained", burst)
    if timed_region_ms >= 2000.0:
        return sustained, f"sustained (timed region {timed_region_ms / 1e3:.2f} s)"
    return burst, f"burst (timed region {timed_region_ms / 1e3:.2f} s < 2 s)"


def _event_time(fn, reps, warm=1):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


def gpu_reference(sc, batch, P, origin, workload, dev):
    """Same-box bar, a reported baseline only: the reference's own op sequence run by stock PyTorch on
    this GPU.  Back-projection: the reference's ATen ops (bmm, round, per-view boolean-mask gather /
    scatter, sum / count; detectors/imvoxelnet.py:145-160, :70-74) per scene on NCHW maps, eager, fp32.
    Neck: the reference module structure through cuDNN with TF32 (torch's default for fp32
    convolutions on this hardware) and in bf16 channels_last_3d.  ms per batch."""
    scene_volume, neck_class, kind, what = reference_modules()
    h, w = sc.feat_hw
    nchw = scenes.random_features(sc, seed=0, batch=batch, channels_last=False).to(dev)
    pts = pj.get_points(sc.n_voxels, sc.voxel_size, origin).to(dev)
    Pd = P.to(dev)
    out = {"what": f"stock PyTorch {torch.__version__} / cuDNN {torch.backends.cudnn.version()} on the same GPU, "
                   f"eager; modules: {what}", "scenes": batch}
    old = torch.backends.cudnn.allow_tf32
    with torch.no_grad():
        def ref_bp():
            for b in range(batch):
                scene_volume(nchw[b][:, :, :h, :w], pts, Pd)
        out["backproject_ms"] = _event_time(ref_bp, reps=2)
        del nchw
        if NECKS[workload] is not None:
            name, kw, _ = NECKS[workload]
            torch.manual_seed(0)
            ref = randomize_norms(neck_class(name)(**kw).eval()).to(dev)
            x = torch.randn((batch, sc.channels) + tuple(sc.n_voxels), device=dev)
            torch.backends.cudnn.allow_tf32 = True
            out["neck_cudnn_tf32_ms"] = _event_time(lambda: ref(x), reps=2)
            refb = ref.to(torch.bfloat16).to(memory_format=torch.channels_last_3d)
            xb = x.to(torch.bfloat16).contiguous(memory_format=torch.channels_last_3d)
            del x
            out["neck_cudnn_bf16_channels_last_3d_ms"] = _event_time(lambda: refb(xb), reps=2)
            del ref, refb, xb
    torch.backends.cudnn.allow_tf32 = old
    torch.cuda.empty_cache()
    return out


def view_sharded_record(name, steps, warmup, collective, dist, world, rank, dev, peaks, scenes_per_rank=None,
                        view_group=None):
    """Strong scaling of one multi-view configuration: the V views of every scene of ONE fixed batch are
    sharded over the `world` ranks (rank r takes views r, r + world, ...), every rank back-projects its
    views into partial sums / counts in symmetric memory, `ivx_finalize_mean_peers` reduces them over
    NVLink and divides (reduce-scatter semantics: each rank finalizes, and runs the bf16 neck on, its own
    slice of the scenes).  world == 1 gives the single-GPU point of the same curve.  Inputs are generated
    on the device (seeded per view, identical on every rank); K steps are replayed from one CUDA graph.
    `scenes_per_rank` (throughput mode): the step holds scenes_per_rank * world scenes, so
    that after the scatter every rank runs its neck on a full batch; compare with the one-GPU record
    through scenes_per_s.
    `view_group` = (process group, ways): hybrid of the two axes -- the ranks form world / ways groups, the views
    of a scene are sharded over the `ways` ranks of one group only (its own symmetric-memory buffers and
    barriers) and every group works on its own scenes, so the NVLink exchange per GPU shrinks from
    (world - 1) / world to (ways - 1) / ways of the partial volumes while all `world` GPUs stay busy."""
    from imvoxelnet_b200 import necks, ops, sharding
    sc, batch, P, origin = scene_setup(name)
    n_groups, group_index, pg, all_ranks = 1, 0, None, world
    if view_group is not None:
        pg, ways = view_group
        n_groups, group_index = world // ways, rank // ways
        world, rank = ways, rank % ways            # from here on: position inside the view group
    if scenes_per_rank is not None:
        batch = scenes_per_rank * world
    V, C, N = sc.n_views, sc.channels, sc.n_total_voxels
    h, w = sc.feat_hw
    hf, wf = sc.map_hw
    my_views = sharding.shard_views(V, rank, world)
    mine = sharding.shard_scenes(batch, rank, world)
    my_scenes = mine.stop - mine.start

    def view_features(v, scene_slice=slice(None)):      # [b, hf, wf, C]: channels-last maps of view v
        g = torch.Generator(device=dev)
        g.manual_seed(1234 + v + 1000 * group_index)
        return torch.randn((batch, hf, wf, C), generator=g, device=dev)[scene_slice]

    def stack(views, scene_slice=slice(None)):          # logical [b, V', C, hf, wf], NHWC memory
        if not views:
            return torch.empty((batch, 0, hf, wf, C), device=dev).permute(0, 1, 4, 2, 3)
        return torch.stack([view_features(v, scene_slice) for v in views], 1).permute(0, 1, 4, 2, 3)

    feats = stack(my_views)
    proj = P[my_views].unsqueeze(0).expand(batch, len(my_views), 3, 4).contiguous().to(dev)
    hws, origins = [(h, w)] * batch, [origin] * batch
    cls, kw, neck_gflop = NECKS[name]
    torch.manual_seed(0)
    neck = randomize_norms(getattr(necks, cls)(**kw).eval()).to(dev)
    neck.precision = "bf16"
    fused, lanes = None, []
    if world > 1:
        use_mc = collective == "multimem" or (collective == "auto" and world > 2)
        # two sets of partial-sum buffers ("lanes"): consecutive steps alternate between them, so the
        # exchange of one step can overlap the neck of the step before it (pipelined timing below)
        lanes = [sharding.FusedViewSharded(batch, C, sc.n_voxels, group=pg, use_multicast=use_mc) for _ in range(2)]
        fused = lanes[0]

    def backproject(out_dtype, lane=0):
        if fused is not None:
            return lanes[lane](feats, hws, proj, sc.voxel_size, origins, out_dtype=out_dtype, scatter_scenes=True)
        return ops.backproject_scenes(feats, hws, proj, sc.n_voxels, sc.voxel_size, origins, out_dtype=out_dtype)

    def bp_only(i):
        return backproject(torch.bfloat16)

    def step(i, lane=0):
        vol, valid, _ = backproject(torch.bfloat16, lane)
        return neck(vol) if vol.shape[0] > 0 else None

    def timed_pipelined(repeats=5):
        """K steps in one CUDA graph, alternating between two streams (and the two buffer sets): step i + 1's
        back-projection, barriers and NVLink reduction may run while step i's neck still does."""
        graph = torch.cuda.CUDAGraph()
        main = torch.cuda.Stream()
        side = [torch.cuda.Stream(), torch.cuda.Stream()]
        main.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(main):
            with torch.cuda.graph(graph, stream=main):
                fork = torch.cuda.Event()
                fork.record(main)
                for st in side:
                    st.wait_event(fork)
                for i in range(steps):
                    with torch.cuda.stream(side[i % 2]):
                        step(i, i % 2)
                for st in side:
                    join = torch.cuda.Event()
                    join.record(st)
                    main.wait_event(join)
        torch.cuda.current_stream().wait_stream(main)
        graph.replay()
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        totals = []
        for _ in range(repeats):
            barrier()
            ev0.record()
            graph.replay()
            ev1.record()
            barrier()
            totals.append(ev0.elapsed_time(ev1))
        return statistics.median(totals), graph

    def barrier():
        if all_ranks > 1:
            dist.barrier()
        torch.cuda.synchronize()

    rec = {"workload": sc.name, "scenes_per_step": batch * n_groups, "views": V, "n_gpus": all_ranks,
           "view_sharding_ways": world, "scene_groups": n_groups,
           "mode": ("strong scaling: one fixed batch" if scenes_per_rank is None else
                    f"throughput: {scenes_per_rank} scenes per rank in flight, views sharded {world}-way"
                    + (f" inside each of {n_groups} groups of GPUs" if n_groups > 1 else "")),
           "views_on_rank0": len(my_views), "scenes_finalized_on_rank0": my_scenes, "neck": cls + " (bf16)"}
    with torch.no_grad():
        # self-check on rank 0 (all ranks take part in the collective): the sharded fp32 volume of its
        # scenes against the unsharded kernel on all V views of the same scenes
        vol32, valid, count = backproject(torch.float32)
        if world > 1 and rank == 0 and group_index == 0 and my_scenes > 0:
            full = stack(list(range(V)), mine)
            pf = P.unsqueeze(0).expand(my_scenes, V, 3, 4).contiguous().to(dev)
            ref, ref_valid, ref_count = ops.backproject_scenes(full, hws[:my_scenes], pf, sc.n_voxels, sc.voxel_size,
                                                               origins[:my_scenes])
            err = (vol32 - ref).abs().max().item()
            scale = ref.abs().max().item()
            rec["self_check"] = {"max_err": err, "max_ref": scale, "within_1e-5": bool(err <= 1e-5 * scale + 1e-6),
                                 "counts_equal": bool(torch.equal(count, ref_count)),
                                 "valid_equal": bool(torch.equal(valid, ref_valid))}
            del full, ref, ref_valid, ref_count
        del vol32, valid, count
        for i in range(max(warmup, 3)):
            step(i, i % 2 if lanes else 0)
        barrier()
        mode = "cuda-graph replay of K steps, median of 5"
        pipelined_total = None
        try:
            bp_total, _g1 = timed_graph(bp_only, steps, barrier)
            total, _g2 = timed_graph(step, steps, barrier)
            if lanes and os.environ.get("IVX_BENCH_PIPELINED", "1") != "0":
                pipelined_total, _g3 = timed_pipelined()
        except Exception as exc:            # capture refused (e.g. the symmetric-memory barrier): eager issue
            torch.cuda.synchronize()
            mode = f"eager ({type(exc).__name__} during graph capture)"
            ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            res = []
            for fn in (bp_only, step):
                barrier()
                ev0.record()
                for i in range(steps):
                    fn(i)
                ev1.record()
                barrier()
                res.append(ev0.elapsed_time(ev1))
            bp_total, total = res
    t = torch.tensor([total, bp_total, pipelined_total if pipelined_total is not None else -1.0], device=dev,
                     dtype=torch.float64)
    if all_ranks > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    total, bp_total, pipelined_total = (float(v) for v in t.tolist())
    ms, bp_ms = total / steps, bp_total / steps
    batch_all = batch * n_groups
    if pipelined_total > 0:
        rec["pipelined"] = {"ms_per_step": pipelined_total / steps,
                            "scenes_per_s": batch_all / (pipelined_total / steps * 1e-3),
                            "what": "the same K steps, consecutive steps on two streams with two sets of "
                                    "partial-sum buffers: a step's NVLink reduction overlaps the previous step's neck"}
    neck_ms = max(ms - bp_ms, 1e-6)
    per_rank = -(-batch // world)                                   # scenes the busiest rank ran the neck on
    tflops = neck_gflop * per_rank / neck_ms
    peak, peak_kind = tensor_peak(peaks, total)
    partial_bytes = batch * N * (C + 1) * 4                          # what every rank writes
    nvlink_in = per_rank * N * (C + 1) * 4 * (world - 1)              # peers' partials of this rank's scenes
    rec.update({
        "ms_per_step": ms, "scenes_per_s": batch_all / (ms * 1e-3), "voxel_views_per_s": batch_all * V * N / (ms * 1e-3),
        "timing": mode,
        "stages": {"backproject_reduce_finalize_ms": bp_ms, "neck_ms": neck_ms},
        "collective": ("none (one GPU)" if fused is None else
                       "fused into ivx_finalize_mean_peers: " +
                       ("multimem.ld_reduce (NVLS)" if fused.multicast else "peer loads over NVLink")),
        "partial_bytes_written_per_rank": partial_bytes, "nvlink_bytes_read_per_rank": nvlink_in,
        "nvlink_read_gbs_if_all_of_stage": nvlink_in / (bp_ms * 1e-3) / 1e9 if world > 1 else 0.0,
        "roofline_neck": {"bound": "tensor", "achieved": tflops, "peak": peak, "unit": "TFLOP/s",
                          "frac": tflops / peak, "peak_kind": "measured " + peak_kind,
                          "scenes_on_busiest_rank": per_rank}})
    del fused, lanes, neck, feats
    torch.cuda.empty_cache()
    return rec


def detector_record(steps, warmup, collective, dist, world, rank, dev):
    """Context for the view-sharded numbers, NOT part of the hot path's claims: the whole `extract_feat` of a
    50-view ScanNet batch -- 2D backbone + FPN on the images, back-projection, 3D neck -- with the views of
    every scene sharded over the ranks, 2D network included (each rank runs it on its own views only).  The 2D
    network is out of scope of this library (the reference builds mmdet's ResNet-50 + FPN); the stand-in of
    `imvoxelnet_b200/backbone2d.py` (torchvision ResNet-50 + a textbook FPN, stock cuDNN, bf16 channels_last,
    random weights) is timed as it is.  Eager issue, CUDA events, max over ranks."""
    from imvoxelnet_b200 import backbone2d, necks, ops, sharding
    sc, batch, P, origin = scene_setup("scannet_fast")
    V, C = sc.n_views, sc.channels
    h, w = sc.feat_hw
    hp, wp = sc.map_hw[0] * 4, sc.map_hw[1] * 4
    my_views = sharding.shard_views(V, rank, world)
    torch.manual_seed(0)
    backbone = backbone2d.ResNet(depth=50).eval().to(dev).to(torch.bfloat16).to(memory_format=torch.channels_last)
    fpn = backbone2d.FPN([256, 512, 1024, 2048], C, 4).eval().to(dev).to(torch.bfloat16).to(memory_format=torch.channels_last)
    cls, kw, _ = NECKS["scannet_fast"]
    neck = randomize_norms(getattr(necks, cls)(**kw).eval()).to(dev)
    neck.precision = "bf16"
    g = torch.Generator(device=dev)
    g.manual_seed(77 + rank)
    imgs = torch.randn((batch * len(my_views), 3, hp, wp), generator=g, device=dev).to(torch.bfloat16)
    imgs = imgs.contiguous(memory_format=torch.channels_last)
    proj = P[my_views].unsqueeze(0).expand(batch, len(my_views), 3, 4).contiguous().to(dev)
    hws, origins = [(h, w)] * batch, [origin] * batch
    fused = None
    if world > 1:
        use_mc = collective == "multimem" or (collective == "auto" and world > 2)
        fused = sharding.FusedViewSharded(batch, C, sc.n_voxels, use_multicast=use_mc)

    def features():                    # [B, V_local, C, Hf, Wf] bf16, channels-last per image
        maps = [fpn(backbone(chunk))[0] for chunk in imgs.split(50)]
        return torch.cat(maps).unflatten(0, (batch, len(my_views)))

    def step():
        feats = features()
        if fused is not None:
            vol, _, _ = fused(feats, hws, proj, sc.voxel_size, origins, out_dtype=torch.bfloat16, scatter_scenes=True)
        else:
            vol, _, _ = ops.backproject_scenes(feats, hws, proj, sc.n_voxels, sc.voxel_size, origins,
                                               out_dtype=torch.bfloat16)
        return neck(vol) if vol.shape[0] > 0 else None

    def timed(fn, reps):
        for _ in range(2):
            fn()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(reps):
            fn()
        e1.record()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) / reps

    reps = max(3, min(steps, 5))
    with torch.no_grad():
        t = torch.tensor([timed(step, reps), timed(features, reps)], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms, ms_2d = (float(v) for v in t.tolist())
    del fused, backbone, fpn, neck, imgs
    torch.cuda.empty_cache()
    return {"workload": sc.name, "what": "whole extract_feat with the 2D network sharded by views too; the 2D network "
            "is an out-of-scope stand-in (torchvision ResNet-50 + textbook FPN, cuDNN, bf16 channels_last)",
            "scenes_per_step": batch, "views": V, "image": [hp, wp], "n_gpus": world,
            "images_on_rank0": batch * len(my_views), "ms_per_step": ms, "scenes_per_s": batch / (ms * 1e-3),
            "stages": {"backbone_fpn_ms": ms_2d, "hot_path_ms": max(ms - ms_2d, 0.0)}, "timing": "eager, CUDA events"}


def run_ours(args):
    import torch.distributed as dist
    from imvoxelnet_b200 import _lib, conv, necks, ops, sharding

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device (no CPU fallback)")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    n_cpus = pin_to_gpu_cpus(local)          # before any pinned allocation: first touch on the near node
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    _lib.lib()  # fail loudly if the CUDA library is missing

    sc, batch, P, origin = scene_setup(args.workload)
    view_sharded = args.workload == "scannet_fast" and world > 1
    h, w = sc.feat_hw
    N, C, V = sc.n_total_voxels, sc.channels, sc.n_views
    b_algo, U, n_valid = algorithmic_bytes(sc, P, origin, dev)
    my_views = list(range(rank, V, world)) if view_sharded else list(range(V))
    Vr = len(my_views)

    neck = None
    if not args.no_neck and NECKS[args.workload] is not None:
        name, kw, neck_gflop = NECKS[args.workload]
        torch.manual_seed(0)
        neck = randomize_norms(getattr(necks, name)(**kw).eval()).to(dev)
        neck.precision = args.neck_precision
    vol_dtype = torch.bfloat16 if (neck is not None and args.neck_precision == "bf16") else torch.float32

    # rotating device input sets (> L2 in total, so no step finds its map in L2
    # from the previous one; the output written per step evicts it anyway)
    n_sets = 3
    host = scenes.random_features(sc, seed=rank, batch=batch, channels_last=True)
    if view_sharded:
        host = _channels_last_5d(host[:, my_views])
    host_pinned = _pinned_like(host)  # keeps the channels-last physical layout
    host_pinned.copy_(host)
    dev_sets = [host_pinned.to(dev) for _ in range(n_sets)]
    for i, d in enumerate(dev_sets):
        d.add_(float(i))  # different contents, same layout
    feat_bytes = 4
    if args.feature_dtype == "bf16":
        dev_sets = [d.to(torch.bfloat16) for d in dev_sets]
        host_pinned = _pinned_like(dev_sets[0])
        host_pinned.copy_(dev_sets[0])
        feat_bytes = 2
        b_algo -= U * C * 2            # the gathered pixels are half as wide
    Pb = P[my_views].unsqueeze(0).expand(batch, Vr, 3, 4).contiguous()
    P_dev = Pb.to(dev)
    hws = [(h, w)] * batch
    origins = [origin] * batch

    fused = None
    if view_sharded and args.collective != "nccl":
        try:
            # two ranks: one peer copy to read, plain NVLink loads win; more ranks: let the switch add
            use_mc = args.collective == "multimem" or (args.collective == "auto" and world > 2)
            fused = sharding.FusedViewSharded(batch, C, sc.n_voxels, use_multicast=use_mc)
        except Exception as exc:
            if rank == 0:
                print(f"[bench] symmetric memory unavailable ({exc}); using the NCCL all-reduce", file=sys.stderr)

    def backproject(feats, proj, out_dtype):
        if fused is not None:
            # reduce-scatter semantics: this rank gets (and runs the neck on) its slice of the scenes
            return fused(feats, hws, proj, sc.voxel_size, origins, out_dtype=out_dtype, scatter_scenes=True)
        if view_sharded:
            return sharding.backproject_scenes_view_sharded(feats, hws, proj, sc.n_voxels, sc.voxel_size,
                                                            origins, out_dtype=out_dtype)
        return ops.backproject_scenes(feats, hws, proj, sc.n_voxels, sc.voxel_size, origins,
                                      out_dtype=out_dtype)

    def bp_step(i):                      # stage 1 alone, fp32 volume (the parity dtype)
        return backproject(dev_sets[i % n_sets], P_dev, torch.float32)

    def bp_step_as_in_step(i):           # stage 1 alone, volume dtype of the full step
        return backproject(dev_sets[i % n_sets], P_dev, vol_dtype)

    def step(i):                         # the whole hot path
        vol, valid, _ = backproject(dev_sets[i % n_sets], P_dev, vol_dtype)
        return (neck(vol) if neck is not None else [vol]), valid

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def count_launches():
        return ops.launch_count() + conv.launch_count()

    with torch.no_grad():
        for i in range(args.warmup):
            step(i)
        barrier()

        # ---- device-resident timing ----
        # K steps are captured once into a CUDA graph and replayed: the timed region
        # holds K back-to-back passes with no host launch gaps (a back-projection of
        # the small configurations is shorter than one Python call).  With a
        # collective in the step (view-sharded mode) the steps are issued eagerly.
        sampler = ClockSampler(local)
        use_graph = not view_sharded
        launches0 = count_launches()
        if use_graph:
            bp_ms_total, _g1 = timed_graph(bp_step, args.steps, barrier)
            bp_in_step_total = bp_ms_total
            if vol_dtype != torch.float32:
                bp_in_step_total, _g3 = timed_graph(bp_step_as_in_step, args.steps, barrier)
            launches0 = count_launches()
            if rank == 0:
                sampler.start()
            total_ms, _g2 = timed_graph(step, args.steps, barrier)
            launches = count_launches() - launches0      # kernels captured = launched per replay
        else:
            if rank == 0:
                sampler.start()
            ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            barrier()
            ev0.record()
            for i in range(args.steps):
                bp_step(i)
            ev1.record()
            barrier()
            bp_ms_total = bp_in_step_total = ev0.elapsed_time(ev1)
            launches0 = count_launches()
            barrier()
            ev0.record()
            for i in range(args.steps):
                step(i)
            ev1.record()
            barrier()
            total_ms = ev0.elapsed_time(ev1)
            launches = count_launches() - launches0
        t = torch.tensor([total_ms, bp_ms_total, bp_in_step_total], device=dev, dtype=torch.float64)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        total_ms, bp_ms_total, bp_in_step_total = (float(v) for v in t.tolist())
        ms_per_step = total_ms / args.steps
        bp_ms = bp_ms_total / args.steps
        bp_in_step_ms = bp_in_step_total / args.steps
        clocks = sampler.stop() if rank == 0 else None
        # eager issue of the same K steps, for the record (host-launch bound when short)
        barrier()
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        ev0.record()
        for i in range(args.steps):
            step(i)
        ev1.record()
        barrier()
        eager_ms_per_step = ev0.elapsed_time(ev1) / args.steps

    scenes_global = batch * (1 if view_sharded else world)
    # scenes whose neck the busiest rank ran: its scatter slice when the views are sharded through the fused
    # kernel (reduce-scatter semantics), the whole batch otherwise (own scenes, or replicated after all-reduce)
    neck_scenes = -(-batch // world) if (view_sharded and fused is not None) else batch
    units_per_step = scenes_global * V * N
    value = units_per_step / (ms_per_step * 1e-3)
    peaks, peak_kind = load_peaks()

    roofline_bp = None
    if not view_sharded:
        achieved = b_algo * batch / (bp_ms * 1e-3) / 1e9
        roofline_bp = {"bound": "hbm", "achieved": achieved, "peak": peaks["hbm_gbs"], "unit": "GB/s",
                       "frac": achieved / peaks["hbm_gbs"], "traffic": _ncu_traffic(sc.name),
                       "kernel": "bp_vec_kernel", "peak_kind": peak_kind,
                       "algorithmic_bytes_per_launch": b_algo * batch, "kernel_ms": bp_ms,
                       "distinct_pixels_per_scene": U, "valid_voxel_views_per_scene": n_valid}
    stages = {"backproject": {"ms": bp_ms, "voxel_views_per_s": units_per_step / (bp_ms * 1e-3),
                              "volume_dtype": "f32", "ms_inside_step": bp_in_step_ms,
                              "volume_dtype_inside_step": str(vol_dtype).replace("torch.", "")}}
    roofline = roofline_bp
    if neck is not None:
        neck_ms = max(ms_per_step - bp_in_step_ms, 1e-6)   # full step minus its back-projection
        tflops = neck_gflop * neck_scenes / neck_ms                # GFLOP / ms == TFLOP/s, of one rank
        peak, which_peak = tensor_peak(peaks, total_ms)
        sustained = peaks.get("bf16_tflops_sustained", peaks["bf16_tflops"])
        stages["neck"] = {"ms": neck_ms, "tflops": tflops, "precision": args.neck_precision,
                          "gflop_per_scene": neck_gflop, "module": type(neck).__name__,
                          "scenes_per_rank": neck_scenes}
        if args.neck_precision == "bf16":
            # dominant kernel of the step: conv3d_tc_kernel (tcgen05 implicit GEMM), timed inside a long step
            roofline = {"bound": "tensor", "achieved": tflops, "peak": peak, "unit": "TFLOP/s",
                        "frac": tflops / peak, "frac_of_sustained_peak": tflops / sustained,
                        "traffic": _ncu_traffic("neck_" + sc.name),
                        "kernel": "conv3d_tc_halo_kernel / conv3d_tc_kernel", "peak_kind": f"{peak_kind} {which_peak}",
                        "algorithmic_flops_per_step": neck_gflop * neck_scenes * 1e9,
                        "launches_per_step": launches // max(args.steps, 1) - 1}
        elif args.neck_precision == "tf32":
            # same kernel with kind::tf32 MMAs; no measured tf32 peak in MEASURED_PEAKS.json: half the
            # bf16 figure is the nominal ratio of the two tensor rates (1.1 vs 2.25 PFLOP/s dense)
            roofline = {"bound": "tensor", "achieved": tflops, "peak": peak / 2, "unit": "TFLOP/s",
                        "frac": tflops / (peak / 2), "traffic": None, "kernel": "conv3d_tc_halo_kernel<tf32> / conv3d_tc_kernel<tf32>",
                        "peak_kind": f"{peak_kind} bf16 {which_peak} / 2: nominal tf32 : bf16 ratio",
                        "algorithmic_flops_per_step": neck_gflop * neck_scenes * 1e9,
                        "launches_per_step": launches // max(args.steps, 1) - 1}
        elif args.neck_precision == "fp32":
            # fp32-grade results from three tf32 products per multiply-add (3xTF32): the tensor pipe
            # does 3x the algorithmic flops at the tf32 rate, so the bound is bf16 peak / 2 / 3
            roofline = {"bound": "tensor", "achieved": tflops, "peak": peak / 6, "unit": "TFLOP/s",
                        "frac": tflops / (peak / 6), "traffic": None,
                        "kernel": "conv3d_tc_halo_kernel<tf32, x3> / conv3d_tc_kernel<tf32, x3>",
                        "peak_kind": f"{peak_kind} bf16 {which_peak} / 2 (tf32 : bf16) / 3 (tf32 products per fp32-grade product)",
                        "algorithmic_flops_per_step": neck_gflop * neck_scenes * 1e9,
                        "tensor_pipe_tflops": 3 * tflops,
                        "launches_per_step": launches // max(args.steps, 1) - 1}

    # ---- end to end: host buffers in, host buffers out, every step ----
    with torch.no_grad():
        out0 = step(0)
        h2d = host_pinned.numel() * feat_bytes + Pb.numel() * 4
        d2h = sum(o.numel() * o.element_size() for o in out0[0]) + out0[1].numel()
        del out0
        P_pinned = Pb.pin_memory()
        e2e_steps = max(10, args.steps)      # long enough to amortise the fill / drain of the 2-deep pipeline

        # the call a user with host buffers makes: StreamedHotPath uploads batch i+1, computes
        # batch i and downloads batch i-1 concurrently (copy-in / compute / copy-out streams)
        from imvoxelnet_b200.pipeline import StreamedHotPath

        def device_step(feats, pd):
            vol, valid, _ = backproject(feats, pd, vol_dtype)
            return (neck(vol) if neck is not None else [vol]), valid

        runner = StreamedHotPath(device_step)

        def batches(n):
            for _ in range(n):
                yield host_pinned, P_pinned

        for _ in runner.run(batches(2)):       # warm-up: allocates the device / pinned buffers
            pass
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in runner.run(batches(e2e_steps)):
            pass
        e1.record()
        barrier()
        e2e_total = e0.elapsed_time(e1)
        # the same copies with no kernel between them: what the host link alone allows at this N
        # (all ranks copy at once, as in the run above)
        static_out = device_step(runner._dev_in[0][0], runner._dev_in[0][1])
        copier = StreamedHotPath(lambda feats, pd: static_out)
        for _ in copier.run(batches(2)):
            pass
        barrier()
        e0.record()
        for _ in copier.run(batches(e2e_steps)):
            pass
        e1.record()
        barrier()
        copy_total = e0.elapsed_time(e1)
        del static_out, copier
    t = torch.tensor([e2e_total, copy_total], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    e2e_ms, copy_ms = (float(v) / e2e_steps for v in t.tolist())
    e2e = {"value": units_per_step / (e2e_ms * 1e-3), "unit": UNIT, "h2d_bytes_per_step": h2d,
           "d2h_bytes_per_step": d2h, "ms_per_step": e2e_ms, "steps": e2e_steps,
           "scenes_per_s": scenes_global / (e2e_ms * 1e-3),
           "copy_ceiling": {"ms_per_step": copy_ms, "h2d_gbs_per_gpu": h2d / (copy_ms * 1e-3) / 1e9,
                            "d2h_gbs_per_gpu": d2h / (copy_ms * 1e-3) / 1e9,
                            "d2h_gbs_all_gpus": world * d2h / (copy_ms * 1e-3) / 1e9,
                            "what": "the same pinned-host copies in and out on the same streams with no kernel "
                                    "in between, all ranks at once: the step cannot be shorter than this"},
           "api": "pipeline.StreamedHotPath over ops.backproject_scenes (ivx_backproject_scenes) + neck "
                  "forward (ivx_conv3d_ndhwc*): every step copies pinned host features + projections in "
                  "and the neck outputs + valid mask out to pinned host memory; copies overlap compute "
                  "on side streams"}

    extras = {}

    def guarded(fn, *a, **k):
        """A side record must never take the headline line down with it."""
        try:
            return fn(*a, **k)
        except Exception as exc:          # e.g. no symmetric memory on this machine
            torch.cuda.synchronize()
            return {"error": f"{type(exc).__name__}: {exc}"[:300]}

    if not args.no_extras:
        with torch.no_grad():
            if neck is not None and args.neck_precision == "bf16" and not view_sharded:
                # the same neck at the reference's own precision (fp32, 1e-5 of the reference): 3xTF32 on tcgen05
                neck.precision = "fp32"
                vol32 = backproject(dev_sets[0], P_dev, torch.float32)[0]
                reps = 3
                fp32_ms = _event_time(lambda: neck(vol32), reps=reps, warm=2)
                neck.precision = args.neck_precision
                del vol32
                burst = peaks["bf16_tflops"]
                tf = neck_gflop * batch / fp32_ms
                extras["neck_fp32"] = {
                    "ms": fp32_ms, "tflops": tf, "precision": "fp32 results (1e-5) from 3 tf32 products on tcgen05",
                    "roofline": {"bound": "tensor", "achieved": tf, "peak": burst / 6, "unit": "TFLOP/s",
                                 "frac": tf / (burst / 6),
                                 "peak_kind": f"{peak_kind} bf16 burst (timed region {reps * fp32_ms / 1e3:.2f} s) / 2 "
                                              "(tf32 : bf16) / 3 (tf32 products per fp32-grade product)",
                                 "kernel": "conv3d_tc_halo_kernel<tf32, x3> / conv3d_tc_kernel<tf32, x3> + split_tf32_kernel"}}
            if world == 1:
                extras["gpu_reference"] = guarded(gpu_reference, sc, batch, P, origin, args.workload, dev)
            if not view_sharded:
                del dev_sets
                torch.cuda.empty_cache()
                common = (args.steps, args.warmup, args.collective, dist, world, rank, dev, peaks)
                extras["view_sharded"] = [guarded(view_sharded_record, name, *common)
                                          for name in ("scannet_fast", "nuscenes_refgrid")]
                if world > 1:
                    extras["view_sharded"].append(guarded(view_sharded_record, "scannet_fast", *common,
                                                          scenes_per_rank=WORKLOADS["scannet_fast"][1]))
                extras["detector_view_sharded"] = guarded(detector_record, args.steps, args.warmup, args.collective,
                                                          dist, world, rank, dev)
                if world >= 4 and world % 2 == 0:
                    # hybrid: pairs of GPUs share the views of their scenes, the pairs are data-parallel
                    pair, _, _ = sharding.view_groups(2, world, rank)
                    extras["view_sharded"].append(guarded(view_sharded_record, "scannet_fast", *common,
                                                          scenes_per_rank=WORKLOADS["scannet_fast"][1],
                                                          view_group=(pair, 2)))

    if rank == 0:
        cpu = None
        if world == 1:
            times, kind, what = cpu_reference_time(sc, P, origin, args.workload,
                                                   repeats=2 if neck is not None else 3, warmup=1,
                                                   with_neck=neck is not None)
            tcpu = min(times)
            cores = os.cpu_count() or 1
            cpu = {"value": sc.voxel_views / tcpu, "unit": UNIT, "cores": cores, "kind": kind,
                   "sample": f"best of {len(times)} x 1 scene of {sc.name}: backproject + view mean"
                             f"{' + 3D neck' if neck is not None else ''} by {what}",
                   "ms_per_scene": tcpu * 1e3, "scenes_per_s": 1.0 / tcpu}
        stage_names = ["backproject+view-mean"] + (["3d-neck"] if neck is not None else [])
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms_per_step, "higher_is_better": True,
            "scaling": "strong" if view_sharded else "weak", "vs_baseline": None,
            "dtype": "f32 back-projection" + (f", {args.neck_precision} neck" if neck is not None else ""),
            "data": "synthetic",
            "config": config_of(sc, batch, 1 if view_sharded else world, {
                "stages": stage_names,
                "cpus_bound_to_this_gpu": n_cpus,
                "parallelism": (f"views sharded over {world} ranks + " +
                                ("NCCL all-reduce" if fused is None else
                                 "reduction fused into the finalize kernel (" +
                                 ("multimem.ld_reduce over the NVSwitch multicast address"
                                  if fused.multicast else "peer loads over NVLink") +
                                 "), scenes scattered: each rank runs the neck on its own scenes")
                                if view_sharded
                                else f"scenes data-parallel x{world}, no collective"),
                "l2": f"{n_sets} rotating input sets; activations written per step >> 126 MB L2",
                "layout": f"features NHWC {args.feature_dtype}, volume and neck activations NDHWC"}),
            "scenes_per_s": scenes_global / (ms_per_step * 1e-3),
            "stages": stages,
            "timing": {"mode": "cuda-graph replay of K steps, median of 5" if use_graph else "eager",
                       "eager_ms_per_step": eager_ms_per_step},
            "roofline": roofline, "roofline_backproject": roofline_bp, "cpu_baseline": cpu, "e2e": e2e,
            "clocks": clocks, "gpu_launches": launches, "lib": _lib.version(),
        }
        line.update(extras)
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
    ap.add_argument("--neck-precision", default="bf16", choices=["bf16", "tf32", "fp32"],
                    help="bf16: tcgen05 kernels (1e-2 of the reference); tf32: fp32 tensors, tf32 tensor-core math "
                         "(5e-3); fp32: fp32-grade results (1e-5) on the tensor cores, three tf32 products (3xTF32)")
    ap.add_argument("--no-neck", action="store_true", help="time the back-projection stage only")
    ap.add_argument("--no-extras", action="store_true",
                    help="skip the side records (neck_fp32, gpu_reference, view_sharded)")
    ap.add_argument("--feature-dtype", default="f32", choices=["f32", "bf16"],
                    help="dtype of the 2D feature maps (the reference's are fp32; bf16 halves the gather)")
    ap.add_argument("--collective", default="auto", choices=["auto", "multimem", "peer", "nccl"],
                    help="view-sharded mode (scannet_fast, N > 1): how the partial sums are reduced")
    args = ap.parse_args()
    if args.warmup < 3 and args.impl == "ours":
        args.warmup = 3
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
