#!/usr/bin/env python
"""Same-box GPU bar: the reference's own op sequence run by stock PyTorch on the
B200 next to this library, per configuration.

    python tools/ref_gpu_bench.py            # prints one JSON object per workload

back-projection: oracle/torch_port.scene_volume (the reference's ATen ops: bmm,
round, boolean-mask gather/scatter per view, sum / count) on CUDA tensors, per
scene, fp32 -- versus ops.backproject_scenes on the whole batch.
neck: oracle/torch_necks (cuDNN) in fp32 (TF32 off), TF32 and bf16
channels_last_3d -- versus the fused kernels in bf16.
"""
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import bench  # noqa: E402
from imvoxelnet_b200 import necks, ops, scenes  # noqa: E402
from imvoxelnet_b200 import projection as pj  # noqa: E402
from oracle import torch_necks, torch_port  # noqa: E402


def timed(fn, reps=3, warm=1):
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


def main():
    dev = "cuda:0"
    for name in ("kitti", "scannet_fast", "sunrgbd_fast", "total_sunrgbd", "nuscenes"):
        sc, batch, P, origin = bench.scene_setup(name)
        h, w = sc.feat_hw
        full = scenes.random_features(sc, seed=0, batch=batch, channels_last=True).to(dev)
        nchw = full.contiguous()                         # the reference holds NCHW maps
        pts = pj.get_points(sc.n_voxels, sc.voxel_size, origin).to(dev)
        Pd = P.to(dev)
        out = {"workload": sc.name, "scenes": batch}
        with torch.no_grad():
            def ref_bp():
                for b in range(batch):
                    torch_port.scene_volume(nchw[b][:, :, :h, :w], pts, Pd)
            out["backproject_ref_torch_gpu_ms"] = timed(ref_bp)
            Pb = Pd.unsqueeze(0).expand(batch, -1, 3, 4).contiguous()
            out["backproject_ours_ms"] = timed(lambda: ops.backproject_scenes(
                full, [(h, w)] * batch, Pb, sc.n_voxels, sc.voxel_size, [origin] * batch), reps=10, warm=3)
            out["backproject_ours_from_nchw_ms"] = timed(lambda: ops.backproject_scenes(
                nchw, [(h, w)] * batch, Pb, sc.n_voxels, sc.voxel_size, [origin] * batch), reps=10, warm=3)
            if bench.NECKS[name] is not None:
                cls, kw, gflop = bench.NECKS[name]
                torch.manual_seed(0)
                ref = bench.randomize_norms(torch_necks.NECKS[cls](**kw).eval()).to(dev)
                mine = getattr(necks, cls)(**kw).eval()
                mine.load_state_dict(ref.state_dict())
                mine.to(dev)
                mine.precision = "bf16"
                x = torch.randn((batch, sc.channels) + tuple(sc.n_voxels), device=dev)
                xcl = x.contiguous(memory_format=torch.channels_last_3d)
                torch.backends.cudnn.allow_tf32 = False
                torch.backends.cuda.matmul.allow_tf32 = False
                out["neck_ref_cudnn_fp32_ms"] = timed(lambda: ref(x), reps=2)
                torch.backends.cudnn.allow_tf32 = True
                out["neck_ref_cudnn_tf32_ms"] = timed(lambda: ref(x), reps=2)
                refb = ref.to(torch.bfloat16).to(memory_format=torch.channels_last_3d)
                xb = xcl.to(torch.bfloat16)
                out["neck_ref_cudnn_bf16_cl3d_ms"] = timed(lambda: refb(xb), reps=2)
                out["neck_ours_bf16_ms"] = timed(lambda: mine(xb), reps=5, warm=2)
                mine.precision = "fp32"
                out["neck_ours_fp32_ms"] = timed(lambda: mine(xcl), reps=2)
                del ref, refb, mine
        print(json.dumps({k: (round(v, 4) if isinstance(v, float) else v) for k, v in out.items()}), flush=True)
        torch.cuda.empty_cache()


if __name__ == "__main__":
    main()
