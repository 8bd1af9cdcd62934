"""Per-tap check of the halo conv kernel (GPU box)."""
import os, sys, itertools
import torch, torch.nn.functional as F
from torch import nn
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from imvoxelnet_b200 import conv
dev = "cuda:0"
torch.manual_seed(0)
cin, cout, dims = 64, 64, (16, 32, 6)
x = torch.randn((1, cin) + dims, device=dev).to(torch.bfloat16)
xn = conv.to_ndhwc(x)
for tap in [(1, 1, 1), (0, 1, 1), (2, 1, 1), (1, 0, 1), (1, 2, 1), (1, 1, 0), (1, 1, 2), (0, 0, 0), (2, 2, 2)]:
    layer = nn.Conv3d(cin, cout, 3, 1, 1, bias=False).to(dev)
    with torch.no_grad():
        w = torch.zeros_like(layer.weight)
        w[:, :, tap[0], tap[1], tap[2]] = torch.randn(cout, cin, device=dev) * 0.1
        layer.weight.copy_(w)
    fused = conv.FusedConv(layer, None, relu=False)
    with torch.no_grad():
        got = conv.from_ndhwc(fused(xn)).float()
        want = F.conv3d(x.float(), layer.weight.to(torch.bfloat16).float(), None, 1, 1)
    err = (got - want).abs()
    bad = (err > 0.05).float()
    # where are the bad outputs?  fraction per index along each axis
    print("tap", tap, "max err %.3f" % err.max().item(), "bad frac %.3f" % bad.mean().item(),
          "| bad by x%8:", [round(bad[0, :, i::8].mean().item(), 2) for i in range(8)],
          "| by y%16:", [round(bad[0, :, :, i::16].mean().item(), 2) for i in (0, 1, 8, 15)],
          "| by z:", [round(bad[0, :, :, :, i].mean().item(), 2) for i in range(6)])
