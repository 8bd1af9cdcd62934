"""The tcgen05 kernel in tf32 mode (fp32 tensors, tf32 math: `ivx_conv3d_ndhwc_tf32_tc`) against torch fp32
convolution with TF32 off.  Tolerance: both operands are rounded to 10 mantissa bits (2^-11 relative each),
the sums are fp32: |err| <= 2e-3 * max|ref|, the "1e-3 class" between the fp32 (1e-5) and bf16 (1e-2) modes."""
import copy

import pytest
import torch
import torch.nn.functional as F
from torch import nn

from imvoxelnet_b200 import conv, necks
from oracle import torch_necks
from tests.util import assert_volume_close

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
TF32_RTOL = 2e-3

CASES = [
    # cin, cout, dims(X,Y,Z), kernel, stride, pad, batch
    (64, 64, (6, 10, 12), 3, 1, 1, 1),
    (32, 64, (7, 5, 6), 3, 1, 1, 2),
    (64, 128, (9, 11, 12), 3, (1, 1, 2), 1, 2),
    (128, 256, (7, 9, 6), 3, (1, 1, 2), 1, 1),
    (256, 256, (6, 7, 3), 3, 1, 0, 1),
    (64, 128, (12, 10, 12), 3, 2, 1, 1),
    (256, 512, (8, 8, 4), 3, 2, 1, 2),
    (512, 512, (5, 5, 4), 3, 1, 1, 1),
    (256, 128, (10, 10, 16), 3, 1, 1, 1),
    (256, 512, (8, 8, 4), 1, 2, 0, 1),
    (128, 64, (8, 8, 8), 1, 1, 0, 1),
    (64, 64, (20, 50, 4), 3, 1, 0, 1),
    (1024, 128, (10, 10, 4), 3, 1, 1, 2),      # split-K
    # shapes the halo-reuse kernel takes (3x3x3, stride 1 or strided along one axis, well-filled tiles)
    (64, 64, (16, 32, 6), 3, 1, 1, 1),
    (64, 128, (5, 24, 32), 3, 1, 1, 2),
    (128, 128, (40, 7, 16), 3, 1, 1, 1),
    (128, 64, (18, 34, 3), 3, 1, (1, 1, 0), 1),
    (32, 64, (20, 50, 4), 3, 1, 0, 1),
    (96, 64, (12, 248, 3), 3, 1, 1, 1),
    (64, 128, (16, 32, 12), 3, (1, 1, 2), 1, 1),
    (128, 256, (8, 32, 6), 3, (1, 1, 2), 1, 2),
    (64, 64, (13, 16, 32), 3, (2, 1, 1), 1, 1),
    (256, 256, (16, 16, 3), 3, 1, 0, 1),
    (128, 512, (8, 32, 4), 3, 1, 1, 1),
]


@pytest.fixture(autouse=True)
def _no_tf32_in_torch():
    old = (torch.backends.cudnn.allow_tf32, torch.backends.cuda.matmul.allow_tf32)
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    yield
    torch.backends.cudnn.allow_tf32, torch.backends.cuda.matmul.allow_tf32 = old


def _bn(cout):
    bn = nn.BatchNorm3d(cout).to(DEV).eval()
    with torch.no_grad():
        bn.running_mean.normal_(0, .1), bn.running_var.uniform_(.5, 1.5)
        bn.weight.uniform_(.5, 1.5), bn.bias.normal_(0, .1)
    return bn


@pytest.mark.parametrize("case", CASES, ids=[str(c) for c in CASES])
@pytest.mark.parametrize("variant", ["plain", "bn_relu_res", "res_after_half"])
def test_tf32_conv(case, variant):
    cin, cout, dims, k, stride, pad, batch = case
    torch.manual_seed(CASES.index(case) * 7 + ["plain", "bn_relu_res", "res_after_half"].index(variant))
    layer = nn.Conv3d(cin, cout, k, stride, pad, bias=(variant != "plain")).to(DEV)
    bn = _bn(cout) if variant != "plain" else None
    relu = variant != "plain"
    fused = conv.FusedConv(layer, bn, relu=relu)
    x = torch.randn((batch, cin) + dims, device=DEV)
    xn = conv.to_ndhwc(x)
    od = fused.out_dims(dims)
    residual = torch.randn((batch, cout) + od, device=DEV) if variant != "plain" else None
    after, post = (variant == "res_after_half"), (0.5 if variant == "res_after_half" else 1.0)
    with torch.no_grad():
        with conv.tf32():
            assert fused.tf32_eligible(xn, torch.float32)
            got = fused(xn, residual=conv.to_ndhwc(residual) if residual is not None else None,
                        residual_after_act=after, post_scale=post)
        assert not fused.tf32_eligible(xn, torch.float32)          # off outside the context
        y = F.conv3d(x, layer.weight, layer.bias, layer.stride, layer.padding)
        if bn is not None:
            y = F.batch_norm(y, bn.running_mean, bn.running_var, bn.weight, bn.bias, False, 0.0, bn.eps)
        if residual is not None and not after:
            y = y + residual
        if relu:
            y = F.relu(y)
        if residual is not None and after:
            y = y + residual
        want = y * post
    got = conv.from_ndhwc(got)
    assert got.dtype == torch.float32 and got.shape == want.shape
    assert_volume_close(got.cpu().numpy(), want.cpu().numpy(), rel=TF32_RTOL, atol=1e-4)
    # and it is NOT the fp32 path in disguise: tf32 rounding must be visible
    assert (got - want).abs().max() > 1e-6 * want.abs().max()


@pytest.mark.parametrize("cin,cout,dims", [(1024, 512, (5, 5, 2)), (512, 256, (10, 10, 4)), (64, 64, (6, 5, 4))])
def test_tf32_transposed_conv(cin, cout, dims):
    torch.manual_seed(cin + dims[0])
    layer = nn.ConvTranspose3d(cin, cout, 2, 2, bias=False).to(DEV)
    bn = _bn(cout)
    fused = conv.FusedConv(layer, bn, relu=True)
    x = torch.randn((2, cin) + dims, device=DEV)
    with torch.no_grad():
        with conv.tf32():
            got = conv.from_ndhwc(fused(conv.to_ndhwc(x)))
        want = F.relu(F.batch_norm(F.conv_transpose3d(x, layer.weight, None, 2), bn.running_mean,
                                   bn.running_var, bn.weight, bn.bias, False, 0.0, bn.eps))
    assert_volume_close(got.cpu().numpy(), want.cpu().numpy(), rel=TF32_RTOL, atol=1e-4)


NECKS = {
    "FastIndoorImVoxelNeck": (dict(in_channels=64, n_blocks=[1, 1, 1], out_channels=64), (2, 64, 12, 8, 8)),
    "ImVoxelNeck": (dict(channels=[32, 64, 128, 256], out_channels=64, down_layers=[1, 2, 3, 4],
                         up_layers=[3, 2, 1], conditional=False), (1, 32, 16, 24, 8)),
    "KittiImVoxelNeck": (dict(in_channels=64, out_channels=256), (2, 64, 13, 18, 12)),
    "NuScenesImVoxelNeck": (dict(in_channels=64, out_channels=256), (1, 64, 16, 14, 12)),
}


@pytest.mark.parametrize("name", sorted(NECKS))
def test_neck_in_tf32_precision(name):
    kw, shape = NECKS[name]
    torch.manual_seed(17)
    ref = torch_necks.randomize_norms(torch_necks.NECKS[name](**kw).eval(), 5).to(DEV)
    m = getattr(necks, name)(**kw).eval()
    m.load_state_dict(ref.state_dict())
    m.to(DEV)
    m.precision = "tf32"
    x = torch.randn(shape, device=DEV)
    with torch.no_grad():
        want, got = ref(x), m(x)
    from tests.util import rel_l2
    for a, b in zip(got, want):
        assert a.dtype == torch.float32
        l2 = rel_l2(a, b)
        print(name, tuple(b.shape), "tf32 relative L2 = %.2e" % l2)
        assert l2 <= 2e-3
        assert_volume_close(a.cpu().numpy(), b.cpu().numpy(), rel=5e-3, atol=1e-4)


@pytest.mark.parametrize("kind,cin,cout,dims", [("k3", 64, 64, (6, 7, 5)), ("k3", 128, 64, (5, 4, 6)),
                                                ("k1", 64, 128, (4, 4, 4)), ("k3s2", 64, 128, (8, 6, 4)),
                                                ("up", 128, 64, (3, 4, 2)), ("k3p0", 64, 64, (6, 6, 3)),
                                                ("k3sz", 64, 128, (6, 9, 6)), ("k3sz", 128, 256, (5, 8, 3)),
                                                ("k3", 32, 96, (9, 4, 10)), ("k3", 256, 256, (4, 4, 4))])
def test_single_layer_gradients_in_tf32(kind, cin, cout, dims):
    """One fused layer (conv + batch-statistics BN, no activation: a ReLU mask flipped by the tf32 forward would
    dominate the comparison) differentiated with the forward and the data gradient on the tensor cores
    (stride-1 and transposed layers) against PyTorch autograd in fp32."""
    from imvoxelnet_b200 import train
    torch.manual_seed(cin + cout + dims[0])
    if kind == "up":
        layer = nn.ConvTranspose3d(cin, cout, 2, 2, bias=False).to(DEV)
    else:
        k, s, p = {"k3": (3, 1, 1), "k1": (1, 1, 0), "k3s2": (3, 2, 1), "k3p0": (3, 1, 0),
                   "k3sz": (3, (1, 1, 2), 1)}[kind]
        layer = nn.Conv3d(cin, cout, k, s, p, bias=False).to(DEV)
    bn = nn.BatchNorm3d(cout).to(DEV).train()
    with torch.no_grad():
        bn.weight.uniform_(.5, 1.5), bn.bias.normal_(0, .1)
    ref_layer, ref_bn = copy.deepcopy(layer), copy.deepcopy(bn)
    fused = conv.FusedConv(layer, bn, relu=False)
    xa = torch.randn((2, cin) + dims, device=DEV, requires_grad=True)
    xb = conv.to_ndhwc(xa.detach()).clone().requires_grad_(True)
    ya = ref_bn(ref_layer(xa))
    seed = torch.randn_like(ya)
    (ya * seed).sum().backward()
    with conv.tf32():
        yb = train.fused_layer_train(fused, xb)
        (conv.from_ndhwc(yb) * seed).sum().backward()
    assert_volume_close(conv.from_ndhwc(yb).detach().cpu().numpy(), ya.detach().cpu().numpy(), rel=TF32_RTOL, atol=1e-4)
    def close(got, want, what):
        err = (got - want).norm() / want.norm()
        assert err <= 5e-3, f"{what}: relative L2 error {err:.3e}"
    close(conv.from_ndhwc(xb.grad), xa.grad, "grad_x")
    close(layer.weight.grad, ref_layer.weight.grad, "grad_w")
    close(bn.weight.grad, ref_bn.weight.grad, "grad_gamma")
    close(bn.bias.grad, ref_bn.bias.grad, "grad_beta")


@pytest.mark.parametrize("name", ["KittiImVoxelNeck", "FastIndoorImVoxelNeck"])
def test_neck_training_step_in_tf32_precision(name):
    """forward and data gradients on the tensor cores, weight gradients in fp32: every gradient within
    1e-2 of PyTorch autograd (fp32, TF32 off) relative to its own max."""
    kw, shape = NECKS[name]
    torch.manual_seed(23)
    ref = torch_necks.randomize_norms(torch_necks.NECKS[name](**kw), 6).to(DEV).train()
    m = getattr(necks, name)(**kw)
    m.load_state_dict(ref.state_dict())
    m.to(DEV).train()
    m.precision = "tf32"
    xa = torch.randn(shape, device=DEV, requires_grad=True)
    xb = xa.detach().clone().requires_grad_(True)
    seeds = None

    def loss(outs):
        nonlocal seeds
        if seeds is None:
            seeds = [torch.randn_like(o) for o in outs]
        return sum((o * s).sum() for o, s in zip(outs, seeds))

    loss(ref(xa)).backward()
    loss(m(xb)).backward()
    # the yardstick: the same net in stock PyTorch with cuDNN's tf32 convolutions switched on
    torch.backends.cudnn.allow_tf32 = True
    twin = copy.deepcopy(ref)
    twin.zero_grad(set_to_none=True)
    xc = xa.detach().clone().requires_grad_(True)
    loss(twin(xc)).backward()
    torch.backends.cudnn.allow_tf32 = False

    # tf32 moves pre-activations by ~1e-3, which flips some ReLU masks of a deep net: isolated
    # gradient entries differ a lot, so the bar is the relative L2 error of every gradient tensor,
    # and it is anchored on what cuDNN's tf32 does to the same gradients (ours <= 3x theirs; floor 0.1 in case
    # cuDNN picks a non-tf32 algorithm on some box: measured 0.03-0.07 for both, a wrong kernel gives O(1))
    def rel_l2(got, want):
        return ((got - want).norm() / want.norm()).item()

    assert rel_l2(xb.grad, xa.grad) <= max(3 * rel_l2(xc.grad, xa.grad), 1e-1)
    for (k, p), q, t in zip(ref.named_parameters(), m.parameters(), twin.parameters()):
        if p.grad is None or p.grad.abs().max().item() < 1e-3:
            continue    # conv bias in front of a batch-statistics norm: analytically zero, pure noise
        assert rel_l2(q.grad, p.grad) <= max(3 * rel_l2(t.grad, p.grad), 1e-1), k
