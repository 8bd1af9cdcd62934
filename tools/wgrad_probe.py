#!/usr/bin/env python
"""Weight gradient of one Conv3d layer on the tensor cores (ivx_conv3d_wgrad_tf32_tc) against autograd, with
timings of both wgrad kernels (GPU box).

    python tools/wgrad_probe.py --cin 64 --cout 64 --dims 216 248 12 --batch 2 [--stride 1 1 2]
"""
import argparse
import json
import os
import sys

import torch
from torch import nn

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from imvoxelnet_b200 import _lib, conv, train  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--cin", type=int, default=64)
    ap.add_argument("--cout", type=int, default=64)
    ap.add_argument("--dims", type=int, nargs=3, default=[16, 16, 12])
    ap.add_argument("--batch", type=int, default=1)
    ap.add_argument("--kernel", type=int, default=3)
    ap.add_argument("--stride", type=int, nargs=3, default=[1, 1, 1])
    ap.add_argument("--reps", type=int, default=3)
    args = ap.parse_args()
    dev = "cuda:0"
    torch.backends.cudnn.allow_tf32 = False
    torch.manual_seed(0)
    layer = nn.Conv3d(args.cin, args.cout, args.kernel, tuple(args.stride), args.kernel // 2, bias=False).to(dev)
    fc = conv.FusedConv(layer, None)
    x = torch.randn((args.batch, args.cin, *args.dims), device=dev)
    y = layer(x)
    dz = torch.randn_like(y)
    want = torch.autograd.grad(y, layer.weight, dz)[0]
    xn, dzn = conv.to_ndhwc(x), conv.to_ndhwc(dz)
    p = fc.make_params(args.batch, tuple(args.dims), tuple(y.shape[2:]), args.cin, args.cout, relu=0, after=0,
                       post=1.0, in_dtype=_lib.IVX_F32, out_dtype=_lib.IVX_F32)

    def timed(fn):
        fn()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(args.reps):
            out = fn()
        e1.record()
        torch.cuda.synchronize()
        return out, e0.elapsed_time(e1) / args.reps

    with conv.tf32():
        got, ms_tc = timed(lambda: train._wgrad_tf32(p, xn, dzn))
    out = {"args": vars(args)}
    if got is None:
        out["tensor_cores"] = "layer not supported"
    else:
        err = ((got - want).norm() / want.norm()).item()
        flops = 2.0 * dz.numel() * args.cin * args.kernel ** 3
        out.update(tc_ms=ms_tc, tc_tflops=flops / ms_tc / 1e9, rel_l2_err=err,
                   max_err_over_max=((got - want).abs().max() / want.abs().max()).item())

    def simt():
        dwp = torch.zeros((args.kernel,) * 3 + (args.cin, args.cout), device=dev)
        _lib.check(_lib.lib().ivx_conv3d_wgrad_ndhwc(p, xn.data_ptr(), dzn.data_ptr(), dwp.data_ptr(),
                                                     torch.cuda.current_stream().cuda_stream), "wgrad")
        return dwp.permute(4, 3, 0, 1, 2)

    ref, ms_simt = timed(simt)
    out.update(simt_ms=ms_simt, simt_rel_l2_err=((ref - want).norm() / want.norm()).item())
    print(json.dumps(out))


if __name__ == "__main__":
    main()
