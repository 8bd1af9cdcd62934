#!/usr/bin/env python
"""Does ivx_finalize_mean_peers co-run with the 3D-neck kernels on one GPU?  (one-GPU stand-in for the
pipelined view-sharded step: the peer path with world = 1 reads local partial sums.)

    IVX_FINALIZE_SMALL_GRID=1 python tools/overlap_probe.py

Prints the time of the neck alone, of the finalize alone, and of both issued on two streams."""
import ctypes
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from imvoxelnet_b200 import _lib, necks  # noqa: E402

dev = "cuda:0"
B, N, C = 8, 25600, 256
neck = necks.FastIndoorImVoxelNeck(256, [1, 1, 1], 128).eval().to(dev)
neck.precision = "bf16"
x = torch.randn(B, 256, 40, 40, 16, device=dev).to(torch.bfloat16).contiguous(memory_format=torch.channels_last_3d)
reps = int(os.environ.get("FINALIZE_REPS", "8"))
sums = torch.randn(B * reps, N, C, device=dev)
counts = torch.randint(0, 5, (B * reps, N), dtype=torch.int32, device=dev)
vol = torch.empty((B * reps, N, C), dtype=torch.bfloat16, device=dev).permute(0, 2, 1)
valid = torch.empty((B * reps, N), dtype=torch.bool, device=dev)
cnt = torch.empty((B * reps, N), dtype=torch.int32, device=dev)


def finalize():
    ps = (ctypes.c_void_p * 1)(sums.data_ptr())
    pc = (ctypes.c_void_p * 1)(counts.data_ptr())
    _lib.check(_lib.lib().ivx_finalize_mean_peers(ps, pc, 1, None, None, B * reps, C, N, vol.data_ptr(),
                                                  _lib.i64(vol.stride()), _lib.IVX_BF16, valid.data_ptr(),
                                                  cnt.data_ptr(), torch.cuda.current_stream().cuda_stream),
               "finalize")


def timed(fn, n=5):
    fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n


s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()


def both():
    cur = torch.cuda.current_stream()
    s1.wait_stream(cur), s2.wait_stream(cur)
    with torch.cuda.stream(s1):
        neck(x)
    with torch.cuda.stream(s2):
        finalize()
    cur.wait_stream(s1), cur.wait_stream(s2)


big = torch.randn(1 << 28, device=dev)


def pair(f, g):
    def run():
        cur = torch.cuda.current_stream()
        s1.wait_stream(cur), s2.wait_stream(cur)
        with torch.cuda.stream(s1):
            f()
        with torch.cuda.stream(s2):
            g()
        cur.wait_stream(s1), cur.wait_stream(s2)
    return run


def ew():
    big.mul_(1.0001)


with torch.no_grad():
    print("elementwise alone %.3f ms; finalize || finalize %.3f; finalize || elementwise %.3f; neck || elementwise %.3f" % (
        timed(ew), timed(pair(finalize, finalize)), timed(pair(finalize, ew)), timed(pair(lambda: neck(x), ew))))
    one = neck.down_layer_0[0]._f1
    xin = x.permute(0, 2, 3, 4, 1).contiguous()
    print("one layer alone %.3f ms; layer || finalize %.3f" % (timed(lambda: one(xin)), timed(pair(lambda: one(xin), finalize))))
    print("neck alone %.3f ms, finalize alone %.3f ms, both on two streams %.3f ms" % (
        timed(lambda: neck(x)), timed(finalize), timed(both)))
