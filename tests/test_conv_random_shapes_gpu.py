"""Seeded sweep of layer shapes through every tensor-core precision (bf16, tf32, fp32-grade 3xTF32).

The kernel a layer takes -- halo-reuse or per-tap, tile orientation, tiles per CTA, split-K -- is decided by a
cost model (csrc/ivx_conv3d_tc.cu use_halo / plan_tc), so odd extents, batch sizes and
channel counts must land on a correct plan whichever one it is.  Reference: torch fp32 convolution (TF32 off);
bars as in the other layer tests (1e-5 fp32, 2e-3 tf32, 1e-2 bf16, of max|ref|)."""
import random

import pytest
import torch
import torch.nn.functional as F
from torch import nn

from imvoxelnet_b200 import conv
from tests.util import assert_volume_close

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _cases(n, seed):
    rng = random.Random(seed)
    out = []
    while len(out) < n:
        cin = rng.choice([32, 64, 64, 128, 256])
        cout = rng.choice([64, 64, 128, 256])
        kind = rng.choice(["k3", "k3", "k3", "k3sz", "k3s2", "k3p0", "k3pz", "k1", "k1s2", "up"])
        dims = (rng.randint(2, 21), rng.randint(2, 41), rng.choice([1, 2, 3, 4, 6, 8, 12, 16, 33]))
        batch = rng.choice([1, 1, 2, 3, 5])
        k, s, p = {"k3": (3, 1, 1), "k3sz": (3, (1, 1, 2), 1), "k3s2": (3, 2, 1), "k3p0": (3, 1, 0),
                   "k3pz": (3, 1, (1, 1, 0)), "k1": (1, 1, 0), "k1s2": (1, 2, 0), "up": (2, 2, 0)}[kind]
        if kind in ("k3p0", "k3pz") and min(dims) < 3:
            continue
        if kind == "up" and dims[0] * dims[1] * dims[2] > 2000:
            continue
        out.append((kind, cin, cout, dims, k, s, p, batch, rng.choice(["plain", "res", "res_after"])))
    return out


CASES = _cases(48, seed=20261107)


@pytest.fixture(autouse=True)
def _no_tf32_in_torch():
    old = (torch.backends.cudnn.allow_tf32, torch.backends.cuda.matmul.allow_tf32)
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    yield
    torch.backends.cudnn.allow_tf32, torch.backends.cuda.matmul.allow_tf32 = old


@pytest.mark.parametrize("case", CASES, ids=[f"{c[0]}-{c[1]}-{c[2]}-{'x'.join(map(str, c[3]))}-b{c[7]}-{c[8]}" for c in CASES])
def test_random_layer_in_every_precision(case):
    kind, cin, cout, dims, k, s, p, batch, variant = case
    torch.manual_seed(1000 + CASES.index(case))
    if kind == "up":
        layer = nn.ConvTranspose3d(cin, cout, k, s, bias=False).to(DEV)
        variant = "plain"                       # the transposed kernel takes no residual
    else:
        layer = nn.Conv3d(cin, cout, k, s, p, bias=True).to(DEV)
    bn = nn.BatchNorm3d(cout).to(DEV).eval()
    with torch.no_grad():
        bn.running_mean.normal_(0, .1), bn.running_var.uniform_(.5, 1.5)
        bn.weight.uniform_(.5, 1.5), bn.bias.normal_(0, .1)
    fused = conv.FusedConv(layer, bn, relu=True)
    x = torch.randn((batch, cin) + dims, device=DEV)
    od = fused.out_dims(dims)
    res = torch.randn((batch, cout) + od, device=DEV) if variant != "plain" else None
    after = variant == "res_after"

    def reference(xr, rr, wq=lambda t: t):
        w = wq(layer.weight)
        y = F.conv_transpose3d(xr, w, None, s) if kind == "up" else F.conv3d(xr, w, layer.bias, s, p)
        y = F.batch_norm(y, bn.running_mean, bn.running_var, bn.weight, bn.bias, False, 0.0, bn.eps)
        if rr is not None and not after:
            y = y + rr
        y = F.relu(y)
        return y + rr if (rr is not None and after) else y

    with torch.no_grad():
        want = reference(x, res)
        xn = conv.to_ndhwc(x)
        rn = conv.to_ndhwc(res) if res is not None else None
        got32 = conv.from_ndhwc(fused(xn, residual=rn, residual_after_act=after))
        with conv.tf32():
            got_tf = conv.from_ndhwc(fused(xn, residual=rn, residual_after_act=after))
        xb, rb = x.to(torch.bfloat16), (res.to(torch.bfloat16) if res is not None else None)
        want16 = reference(xb.float(), rb.float() if rb is not None else None, lambda t: t.to(torch.bfloat16).float())
        got16 = conv.from_ndhwc(fused(conv.to_ndhwc(xb), residual=conv.to_ndhwc(rb) if rb is not None else None,
                                      residual_after_act=after)).float()
    assert_volume_close(got32.cpu().numpy(), want.cpu().numpy())
    assert_volume_close(got_tf.cpu().numpy(), want.cpu().numpy(), rel=2e-3, atol=1e-4)
    assert_volume_close(got16.cpu().numpy(), want16.cpu().numpy(), rel=1e-2, atol=1e-3)
