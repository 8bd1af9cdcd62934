#!/usr/bin/env python
"""End-to-end ``ImVoxelNet.extract_feat`` on synthetic images (GPU box): 2D backbone + FPN
(stock PyTorch / cuDNN stand-ins, out of scope of this library) -> fused back-projection ->
3D neck.  Prints scenes/s and the share of the step spent in the hot path.

    python tools/detector_bench.py [--config kitti|scannet_fast|sunrgbd_fast] [--batch B]
"""
import argparse
import json
import os
import sys
import warnings

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import bench  # noqa: E402
from imvoxelnet_b200 import detector, scenes  # noqa: E402

MODELS = {
    "kitti": dict(fpn_out=64, neck_3d=dict(type='KittiImVoxelNeck', in_channels=64, out_channels=256)),
    "scannet_fast": dict(fpn_out=256, neck_3d=dict(type='FastIndoorImVoxelNeck', in_channels=256,
                                                   n_blocks=[1, 1, 1], out_channels=128)),
    "sunrgbd_fast": dict(fpn_out=256, neck_3d=dict(type='FastIndoorImVoxelNeck', in_channels=256,
                                                   n_blocks=[1, 1, 1], out_channels=128)),
}


def timed(fn, reps=5, warm=2):
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
    ap = argparse.ArgumentParser()
    ap.add_argument("--config", default="kitti", choices=sorted(MODELS))
    ap.add_argument("--batch", type=int, default=None)
    args = ap.parse_args()
    sc, batch, P, origin = bench.scene_setup(args.config)
    batch = args.batch or batch
    spec = MODELS[args.config]
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        model = detector.ImVoxelNet(
            backbone=dict(type='ResNet', depth=50, num_stages=4, out_indices=(0, 1, 2, 3)),
            neck=dict(type='FPN', in_channels=[256, 512, 1024, 2048], out_channels=spec["fpn_out"], num_outs=4),
            neck_3d=spec["neck_3d"], bbox_head=None, n_voxels=sc.n_voxels, voxel_size=sc.voxel_size)
    bench.randomize_norms(model.neck_3d)
    model = model.eval().cuda()
    model.backbone.to(torch.bfloat16).to(memory_format=torch.channels_last)
    model.neck.to(torch.bfloat16).to(memory_format=torch.channels_last)
    model.set_precision("bf16")
    H, W = sc.img_meta['img_shape'][:2]
    Hp, Wp = sc.map_hw[0] * 4, sc.map_hw[1] * 4
    img = torch.randn(batch, sc.n_views, 3, Hp, Wp, device="cuda", dtype=torch.bfloat16)
    metas = [sc.img_meta] * batch

    def full():
        with torch.no_grad():
            return model.extract_feat(img, metas, 'test')

    def backbone_only():
        with torch.no_grad():
            x = img.reshape(-1, 3, Hp, Wp).contiguous(memory_format=torch.channels_last)
            return model.neck(model.backbone(x))[0]

    t_full = timed(full)
    t_2d = timed(backbone_only)
    print(json.dumps({"config": sc.name, "scenes": batch, "views": sc.n_views, "image": [Hp, Wp],
                      "extract_feat_ms": round(t_full, 3), "backbone_fpn_ms": round(t_2d, 3),
                      "hot_path_ms": round(t_full - t_2d, 3), "scenes_per_s": round(batch / t_full * 1e3, 1),
                      "note": "2D part: torchvision ResNet-50 + textbook FPN in bf16 channels_last (stock cuDNN)"}))


if __name__ == "__main__":
    main()
