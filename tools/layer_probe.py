#!/usr/bin/env python
"""Run ONE fused bf16 neck layer a few times (for ncu captures and quick A/B timings on the GPU box).

    python tools/layer_probe.py --cin 64 --cout 64 --dims 80 80 32 --batch 32 --kernel 1 --residual
"""
import argparse
import json
import os
import sys

import torch
from torch import nn

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from imvoxelnet_b200 import conv  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--cin", type=int, default=64)
    ap.add_argument("--cout", type=int, default=64)
    ap.add_argument("--dims", type=int, nargs=3, default=[80, 80, 32])
    ap.add_argument("--batch", type=int, default=32)
    ap.add_argument("--kernel", type=int, default=1)
    ap.add_argument("--stride", type=int, default=1)
    ap.add_argument("--transposed", action="store_true")
    ap.add_argument("--residual", action="store_true")
    ap.add_argument("--reps", type=int, default=5)
    ap.add_argument("--precision", default="bf16", choices=["bf16", "tf32"])
    args = ap.parse_args()
    dev = "cuda:0"
    torch.manual_seed(0)
    if args.transposed:
        c = nn.ConvTranspose3d(args.cin, args.cout, args.kernel, args.stride, bias=False)
    else:
        c = nn.Conv3d(args.cin, args.cout, args.kernel, args.stride, args.kernel // 2, bias=False)
    bn = nn.BatchNorm3d(args.cout)
    layer = conv.FusedConv(c.to(dev).eval(), bn.to(dev).eval(), relu=True)
    x = torch.randn((args.batch, *args.dims, args.cin), device=dev)
    if args.precision == "bf16":
        x = x.to(torch.bfloat16)
    with torch.no_grad(), conv.tf32(args.precision == "tf32"):
        y = layer(x)
        res = torch.randn_like(y) if args.residual else None
        for _ in range(2):
            layer(x, residual=res)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(args.reps):
            layer(x, residual=res)
        e1.record()
        torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / args.reps
    taps = args.kernel ** 3
    rows = (x.numel() // args.cin) if args.transposed else (y.numel() // args.cout)
    nbytes = (x.numel() + y.numel() * (2 if args.residual else 1)) * x.element_size()
    print(json.dumps({"ms": ms, "tflops": 2.0 * rows * args.cin * args.cout * taps / ms / 1e9,
                      "gbytes_per_s": nbytes / ms / 1e6, "args": vars(args)}))


if __name__ == "__main__":
    main()
