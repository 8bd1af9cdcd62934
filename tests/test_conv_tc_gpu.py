"""The tcgen05 Conv3d kernel against torch fp32 convolution of the same
bf16-rounded operands (TF32 off).  Tolerance: the output is rounded to bf16
(2^-9 relative), so |err| <= 1e-2 * max|ref| as the stated bf16 bar."""
import numpy as np
import pytest
import torch
import torch.nn.functional as F
from torch import nn

from imvoxelnet_b200 import conv
from tests.util import assert_volume_close

pytestmark = pytest.mark.gpu
DEV = "cuda:0"

CASES = [
    # cin, cout, dims(X,Y,Z), kernel, stride, pad, batch
    (64, 64, (6, 10, 12), 3, 1, 1, 1),
    (64, 128, (9, 11, 12), 3, (1, 1, 2), 1, 2),
    (128, 256, (7, 9, 6), 3, (1, 1, 2), 1, 1),
    (256, 256, (6, 7, 3), 3, 1, 0, 1),
    (64, 128, (12, 10, 12), 3, 2, 1, 1),
    (256, 256, (8, 8, 3), 3, 1, (1, 1, 0), 1),
    (256, 512, (8, 8, 4), 3, 2, 1, 2),
    (512, 512, (5, 5, 4), 3, 1, 1, 1),
    (256, 128, (10, 10, 16), 3, 1, 1, 1),
    (256, 512, (8, 8, 4), 1, 2, 0, 1),
    (128, 64, (8, 8, 8), 1, 1, 0, 1),
    # shapes the halo-reuse kernel takes (3x3x3, stride 1, c_out <= 128, well-filled tiles)
    (64, 64, (16, 32, 6), 3, 1, 1, 1),
    (64, 128, (5, 24, 32), 3, 1, 1, 2),
    (128, 128, (40, 7, 16), 3, 1, 1, 1),
    (128, 64, (18, 34, 3), 3, 1, (1, 1, 0), 1),
    (64, 64, (20, 50, 4), 3, 1, 0, 1),
    (192, 64, (12, 248, 3), 3, 1, 1, 1),
    (64, 128, (16, 32, 12), 3, (1, 1, 2), 1, 1),
    (128, 256, (8, 32, 6), 3, (1, 1, 2), 1, 2),
    (64, 64, (13, 16, 32), 3, (2, 1, 1), 1, 1),
    (256, 256, (16, 16, 3), 3, 1, 0, 1),
    (128, 512, (8, 32, 4), 3, 1, 1, 1),
]


def _reference(x, layer, bn, relu, residual, after, post):
    w = layer.weight.to(torch.bfloat16).float()
    y = F.conv3d(x.float(), w, layer.bias, layer.stride, layer.padding)
    if bn is not None:
        y = F.batch_norm(y, bn.running_mean, bn.running_var, bn.weight, bn.bias, False, 0.0, bn.eps)
    if residual is not None and not after:
        y = y + residual.float()
    if relu:
        y = F.relu(y)
    if residual is not None and after:
        y = y + residual.float()
    return y * post


@pytest.fixture(autouse=True)
def _no_tf32():
    old = (torch.backends.cudnn.allow_tf32, torch.backends.cuda.matmul.allow_tf32)
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    yield
    torch.backends.cudnn.allow_tf32, torch.backends.cuda.matmul.allow_tf32 = old


@pytest.mark.parametrize("case", CASES, ids=[str(c) for c in CASES])
@pytest.mark.parametrize("variant", ["plain", "bn_relu_res", "res_after_half"])
def test_tensor_core_conv(case, variant):
    cin, cout, dims, k, stride, pad, batch = case
    torch.manual_seed(CASES.index(case) * 7 + ["plain", "bn_relu_res", "res_after_half"].index(variant))
    layer = nn.Conv3d(cin, cout, k, stride, pad, bias=(variant != "plain")).to(DEV)
    bn = None
    if variant != "plain":
        bn = nn.BatchNorm3d(cout).to(DEV).eval()
        with torch.no_grad():
            bn.running_mean.normal_(0, .1), bn.running_var.uniform_(.5, 1.5)
            bn.weight.uniform_(.5, 1.5), bn.bias.normal_(0, .1)
    relu = variant != "plain"
    fused = conv.FusedConv(layer, bn, relu=relu)
    x = torch.randn((batch, cin) + dims, device=DEV).to(torch.bfloat16)
    xn = conv.to_ndhwc(x)
    od = fused.out_dims(dims)
    residual = None
    if variant != "plain":
        residual = torch.randn((batch, cout) + od, device=DEV).to(torch.bfloat16)
    after, post = (variant == "res_after_half"), (0.5 if variant == "res_after_half" else 1.0)
    assert fused.tc_eligible(xn, torch.bfloat16)
    with torch.no_grad():
        got = fused(xn, residual=conv.to_ndhwc(residual) if residual is not None else None,
                    residual_after_act=after, post_scale=post)
        want = _reference(x, layer, bn, relu, residual, after, post)
    got = conv.from_ndhwc(got).float()
    assert got.shape == want.shape
    assert_volume_close(got.cpu().numpy(), want.cpu().numpy(), rel=1e-2, atol=1e-3)


def test_cuda_core_bf16_path_agrees_with_tensor_cores():
    torch.manual_seed(0)
    layer = nn.Conv3d(64, 64, 3, 1, 1, bias=False).to(DEV)
    fused = conv.FusedConv(layer, None, relu=True)
    x = conv.to_ndhwc(torch.randn(1, 64, 8, 8, 8, device=DEV).to(torch.bfloat16))
    with torch.no_grad():
        a = fused(x)
        conv.USE_TENSOR_CORES = False
        try:
            b = fused(x)
        finally:
            conv.USE_TENSOR_CORES = True
    assert_volume_close(a.float().cpu().numpy(), b.float().cpu().numpy(), rel=1e-2, atol=1e-3)


@pytest.mark.parametrize("cin,cout,dims,batch", [(128, 64, (5, 6, 4), 2), (1024, 512, (10, 10, 4), 1),
                                                  (512, 256, (20, 20, 8), 1)])
def test_tensor_core_transposed_conv(cin, cout, dims, batch):
    """ConvTranspose3d(k=2, s=2) + BN + ReLU of the FastIndoor up blocks (:57-59)."""
    torch.manual_seed(cin)
    layer = nn.ConvTranspose3d(cin, cout, 2, 2, bias=False).to(DEV)
    bn = nn.BatchNorm3d(cout).to(DEV).eval()
    with torch.no_grad():
        bn.running_mean.normal_(0, .1), bn.running_var.uniform_(.5, 1.5)
        bn.weight.uniform_(.5, 1.5), bn.bias.normal_(0, .1)
    fused = conv.FusedConv(layer, bn, relu=True)
    x = torch.randn((batch, cin) + dims, device=DEV).to(torch.bfloat16)
    xn = conv.to_ndhwc(x)
    assert fused.tc_eligible(xn, torch.bfloat16)
    with torch.no_grad():
        got = conv.from_ndhwc(fused(xn)).float()
        w = layer.weight.to(torch.bfloat16).float()
        want = F.relu(F.batch_norm(F.conv_transpose3d(x.float(), w, None, 2), bn.running_mean,
                                   bn.running_var, bn.weight, bn.bias, False, 0.0, bn.eps))
    assert got.shape == want.shape
    assert_volume_close(got.cpu().numpy(), want.cpu().numpy(), rel=1e-2, atol=1e-3)


@pytest.mark.parametrize("cin,cout,dims,stride,batch", [(1024, 1024, (10, 10, 4), 1, 1), (1024, 128, (10, 10, 4), 1, 2),
                                                         (512, 1024, (20, 20, 8), 2, 1), (512, 128, (20, 20, 8), 1, 1)])
def test_split_k_tail_layers_and_workspace_reuse(cin, cout, dims, stride, batch):
    """Few-tile layers split their k-steps over several CTAs through an fp32 workspace;
    the workspace must come back zeroed: three calls in a row give the same result."""
    torch.manual_seed(cout + dims[0])
    layer = nn.Conv3d(cin, cout, 3, stride, 1, bias=False).to(DEV)
    bn = nn.BatchNorm3d(cout).to(DEV).eval()
    with torch.no_grad():
        bn.running_mean.normal_(0, .1), bn.running_var.uniform_(.5, 1.5)
        bn.weight.uniform_(.5, 1.5), bn.bias.normal_(0, .1)
    fused = conv.FusedConv(layer, bn, relu=True)
    x = torch.randn((batch, cin) + dims, device=DEV).to(torch.bfloat16)
    xn = conv.to_ndhwc(x)
    od = fused.out_dims(dims)
    res = torch.randn((batch, cout) + od, device=DEV).to(torch.bfloat16)
    with torch.no_grad():
        want = _reference(x, layer, bn, True, res, False, 1.0)
        outs = [conv.from_ndhwc(fused(xn, residual=conv.to_ndhwc(res))).float() for _ in range(3)]
    from imvoxelnet_b200 import _lib
    assert any(w is not False for w in fused._ws.values()), "this shape was expected to use split-K"
    for got in outs:
        assert_volume_close(got.cpu().numpy(), want.cpu().numpy(), rel=1e-2, atol=1e-3)
    # the order of the fp32 atomics is not fixed: two calls may differ by one bf16 rounding (2^-8 relative)
    assert torch.equal(outs[0], outs[2]) or (outs[0] - outs[2]).abs().max() <= 2.0 ** -7 * outs[0].abs().max()
