#!/usr/bin/env python
"""Time the 3D necks layer by layer and as a whole (GPU box).

    python tools/neck_bench.py [--neck KittiImVoxelNeck] [--batch 8] [--precision bf16]

Prints per-layer ms / TFLOP/s of the fused conv kernels, the whole-neck time
and the same neck in stock PyTorch (cuDNN) fp32 / TF32 / bf16 channels_last_3d
as the same-box reference bar."""
import argparse
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from imvoxelnet_b200 import conv, necks  # noqa: E402

CONFIGS = {
    "KittiImVoxelNeck": (dict(in_channels=64, out_channels=256), (64, 216, 248, 12), 2603.37),
    "NuScenesImVoxelNeck": (dict(in_channels=64, out_channels=256), (64, 312, 312, 12), 1571.76),
    "FastIndoorImVoxelNeck": (dict(in_channels=256, n_blocks=[1, 1, 1], out_channels=128), (256, 40, 40, 16), 489.79),
    "ImVoxelNeck": (dict(channels=[64, 128, 256, 512], out_channels=64, down_layers=[1, 2, 3, 4],
                         up_layers=[3, 2, 1], conditional=False), (64, 80, 80, 32), 631.66),
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
    ap.add_argument("--neck", default="KittiImVoxelNeck")
    ap.add_argument("--batch", type=int, default=8)
    ap.add_argument("--precision", default="bf16")
    ap.add_argument("--layers", action="store_true")
    ap.add_argument("--graph", action="store_true", help="also time the forward replayed from a CUDA graph")
    ap.add_argument("--torch-ref", action="store_true")
    ap.add_argument("--train", action="store_true", help="time forward + backward in training mode (fp32)")
    args = ap.parse_args()
    kw, shape, gflop = CONFIGS[args.neck]
    dev = "cuda:0"
    if args.train:
        from oracle import torch_necks
        torch.backends.cudnn.allow_tf32 = False
        torch.manual_seed(0)
        ref = torch_necks.NECKS[args.neck](**kw).to(dev).train()
        mine = getattr(necks, args.neck)(**kw).to(dev).train()
        mine.load_state_dict(ref.state_dict())
        mine.precision = args.precision if args.precision in ("fp32", "tf32") else "fp32"
        x = torch.randn((args.batch,) + shape, device=dev).contiguous(memory_format=torch.channels_last_3d)

        def step(m, inp):
            m.zero_grad(set_to_none=True)
            xin = inp.clone().requires_grad_(True)
            sum(o.float().square().mean() for o in m(xin)).backward()

        out = {"neck": args.neck, "batch": args.batch, "mode": "train fwd+bwd, " + mine.precision}
        out["ours_ms"] = timed(lambda: step(mine, x), reps=3, warm=1)
        from imvoxelnet_b200 import train
        train.PROFILE = {}
        t0 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0[0].record()
        step(mine, x)
        t0[1].record()
        torch.cuda.synchronize()
        out["profiled_step_ms"] = t0[0].elapsed_time(t0[1])
        out["kernel_ms"] = {k: [v[0], round(v[1], 2)] for k, v in sorted(train.PROFILE.items(), key=lambda kv: -kv[1][1])}
        train.PROFILE = None
        out["torch_cudnn_fp32_ms"] = timed(lambda: step(ref, x.contiguous()), reps=3, warm=1)
        torch.backends.cudnn.allow_tf32 = True
        out["torch_cudnn_tf32_ms"] = timed(lambda: step(ref, x.contiguous()), reps=3, warm=1)
        print(json.dumps(out))
        return
    torch.manual_seed(0)
    m = getattr(necks, args.neck)(**kw).eval().to(dev)
    m.precision = args.precision
    x = torch.randn((args.batch,) + shape, device=dev).contiguous(memory_format=torch.channels_last_3d)
    if args.precision == "bf16":
        x = x.to(torch.bfloat16)
    # (fp32 and tf32 keep fp32 tensors)
    out = {"neck": args.neck, "batch": args.batch, "precision": args.precision}
    if args.layers:
        records = {}
        order = []
        orig, orig_up = conv.FusedConv.__call__, conv.upsample2x_trilinear

        def note(key, desc, ms, fl):
            if key not in records:
                records[key] = [desc, ms, fl]
                order.append(key)
            records[key][1] = min(records[key][1], ms)

        counter = [0]

        def wrapped(self, xx, **k):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            y = orig(self, xx, **k)
            e1.record()
            torch.cuda.synchronize()
            ks = self.conv.kernel_size
            taps = ks[0] * ks[1] * ks[2]
            rows = y.numel() // y.shape[-1] if not self.transposed else xx.numel() // xx.shape[-1]
            fl = 2.0 * rows * self.conv.in_channels * self.conv.out_channels * taps
            desc = "%-28s -> %-28s k%s s%s %s%s" % (tuple(xx.shape), tuple(y.shape), ks[0], self.conv.stride,
                                                  "tc" if self.tc_eligible(xx, y.dtype) else ("tf32" if self.tf32_eligible(xx, y.dtype) else
                                                                                                 ("3xtf32" if self.x3_eligible(xx, y.dtype) else "simt")),
                                                  " +res" if k.get("residual") is not None else "")
            note(counter[0], desc, e0.elapsed_time(e1), fl)
            counter[0] += 1
            return y

        def wrapped_up(xx):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            y = orig_up(xx)
            e1.record()
            torch.cuda.synchronize()
            note(counter[0], "%-28s -> %-28s trilinear x2" % (tuple(xx.shape), tuple(y.shape)), e0.elapsed_time(e1), 0.0)
            counter[0] += 1
            return y

        with torch.no_grad():
            m(x)
            conv.FusedConv.__call__ = wrapped
            necks.upsample2x_trilinear = wrapped_up
            for _ in range(3):          # best of three passes: the first one pays for allocator misses
                counter[0] = 0
                m(x)
            conv.FusedConv.__call__ = orig
            necks.upsample2x_trilinear = orig_up
        total = 0.0
        for key in order:
            desc, ms, fl = records[key]
            total += ms
            print("%-84s %8.3f ms %8.1f TFLOP/s" % (desc, ms, fl / ms / 1e9))
        print("sum of layers %.3f ms" % total)
    with torch.no_grad():
        ms = timed(lambda: m(x))
        if args.graph:       # the same forward replayed from a CUDA graph (no host launch gaps)
            g = torch.cuda.CUDAGraph()
            side = torch.cuda.Stream()
            side.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(side):
                m(x)
                with torch.cuda.graph(g, stream=side):
                    for _ in range(10):
                        m(x)
            torch.cuda.current_stream().wait_stream(side)
            out["graph_ms"] = timed(g.replay, reps=5, warm=2) / 10
    out["ms"] = ms
    out["tflops"] = gflop * args.batch / ms
    out["frac_of_bf16_burst_peak"] = out["tflops"] / 1616.4
    if args.torch_ref:
        sys.path.insert(0, ROOT)
        from oracle import torch_necks
        ref = torch_necks.NECKS[args.neck](**kw).eval().to(dev)
        xr = x.float()
        with torch.no_grad():
            torch.backends.cudnn.allow_tf32 = False
            out["torch_fp32_ms"] = timed(lambda: ref(xr), reps=2, warm=1)
            torch.backends.cudnn.allow_tf32 = True
            out["torch_tf32_ms"] = timed(lambda: ref(xr), reps=2, warm=1)
            refb = ref.to(torch.bfloat16).to(memory_format=torch.channels_last_3d)
            xb = xr.to(torch.bfloat16).contiguous(memory_format=torch.channels_last_3d)
            out["torch_bf16_cl3d_ms"] = timed(lambda: refb(xb), reps=2, warm=1)
    print(json.dumps(out))


if __name__ == "__main__":
    main()
