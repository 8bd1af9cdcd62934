"""fp32-grade layers on the tensor cores (3xTF32: `ivx_split_tf32_ndhwc` + `ivx_conv3d_ndhwc_f32x3_tc`)
against a float64 convolution of the same fp32 tensors.

Bar: the fp32 bar of the whole suite, |err| <= 1e-5 * max|ref| + 1e-6 (tests/util.py) -- the reference neck is
an fp32 nn.Conv3d stack (mmdet3d/models/necks/imvoxelnet.py:116-120, no fp16 key in any config).  The error
of cuDNN's own fp32 convolution against the same float64 result is printed beside ours."""
import pytest
import torch
import torch.nn.functional as F
from torch import nn

from imvoxelnet_b200 import conv
from tests.test_conv_tf32_gpu import CASES, _bn
from tests.util import assert_volume_close

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


@pytest.fixture(autouse=True)
def _no_tf32_in_torch():
    old = (torch.backends.cudnn.allow_tf32, torch.backends.cuda.matmul.allow_tf32)
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    yield
    torch.backends.cudnn.allow_tf32, torch.backends.cuda.matmul.allow_tf32 = old


def test_split_tf32_halves():
    """hi has 10 explicit mantissa bits, lo too, and hi + lo reproduces x to 2^-22."""
    torch.manual_seed(0)
    x = (torch.randn(3, 5, 7, 2, 64, device=DEV) * torch.logspace(-6, 6, 64, device=DEV)).contiguous()
    y = conv.split_tf32(x)
    assert y.shape == (3, 5, 7, 2, 128)
    hi, lo = y[..., :64], y[..., 64:]
    assert int((hi.contiguous().view(torch.int32) & 0x1FFF).abs().max()) == 0
    assert int((lo.contiguous().view(torch.int32) & 0x1FFF).abs().max()) == 0
    assert torch.equal(hi, conv._rna_tf32(x))
    assert torch.equal(lo, conv._rna_tf32(x - hi))
    assert ((hi.double() + lo.double() - x.double()).abs() <= x.double().abs() * 2.0 ** -21).all()


def _reference64(x, layer, bn, residual, relu, after, post, transposed=False):
    bias = layer.bias.double() if layer.bias is not None else None
    if transposed:
        y = F.conv_transpose3d(x.double(), layer.weight.double(), bias, layer.stride)
    else:
        y = F.conv3d(x.double(), layer.weight.double(), bias, layer.stride, layer.padding)
    if bn is not None:
        y = F.batch_norm(y, bn.running_mean.double(), bn.running_var.double(), bn.weight.double(),
                         bn.bias.double(), False, 0.0, bn.eps)
    if residual is not None and not after:
        y = y + residual.double()
    if relu:
        y = F.relu(y)
    if residual is not None and after:
        y = y + residual.double()
    return y * post


@pytest.mark.parametrize("case", CASES, ids=[str(c) for c in CASES])
@pytest.mark.parametrize("variant", ["plain", "bn_relu_res", "res_after_half"])
def test_x3_conv(case, variant):
    cin, cout, dims, k, stride, pad, batch = case
    torch.manual_seed(CASES.index(case) * 3 + ["plain", "bn_relu_res", "res_after_half"].index(variant))
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
        assert fused.x3_eligible(xn, torch.float32)
        n0 = conv.launch_count()
        got = fused(xn, residual=conv.to_ndhwc(residual) if residual is not None else None,
                    residual_after_act=after, post_scale=post)
        assert conv.launch_count() == n0 + 2           # split + the tensor-core layer
        with conv.fp32x3(False):
            assert not fused.x3_eligible(xn, torch.float32)
        want = _reference64(x, layer, bn, residual, relu, after, post)
        y32 = F.conv3d(x, layer.weight, layer.bias, layer.stride, layer.padding)
    got = conv.from_ndhwc(got)
    assert got.dtype == torch.float32 and got.shape == want.shape
    scale = want.abs().max().item()
    ours = (got.double() - want).abs().max().item() / scale
    if variant == "plain":
        cudnn = (y32.double() - want).abs().max().item() / scale
        print(f"{case}: 3xTF32 {ours:.2e}, cuDNN fp32 {cudnn:.2e} of max|ref| against float64")
    assert_volume_close(got.cpu().numpy(), want.cpu().numpy())


@pytest.mark.parametrize("cin,cout,dims", [(1024, 512, (5, 5, 2)), (512, 256, (10, 10, 4)), (64, 64, (6, 5, 4))])
def test_x3_transposed_conv(cin, cout, dims):
    torch.manual_seed(cin + dims[0])
    layer = nn.ConvTranspose3d(cin, cout, 2, 2, bias=False).to(DEV)
    bn = _bn(cout)
    fused = conv.FusedConv(layer, bn, relu=True)
    x = torch.randn((2, cin) + dims, device=DEV)
    with torch.no_grad():
        assert fused.x3_eligible(conv.to_ndhwc(x), torch.float32)
        got = conv.from_ndhwc(fused(conv.to_ndhwc(x)))
        want = _reference64(x, layer, bn, None, True, False, 1.0, transposed=True)
    assert_volume_close(got.cpu().numpy(), want.cpu().numpy())


def test_x3_large_k_accumulation():
    """K = 27 * 512 = 13 824 products per output with inputs of one sign (no cancellation: the
    accumulator grows monotonically, the worst case for a truncating fp32 accumulate)."""
    torch.manual_seed(7)
    layer = nn.Conv3d(512, 128, 3, 1, 1, bias=False).to(DEV)
    with torch.no_grad():
        layer.weight.abs_()
    fused = conv.FusedConv(layer, None, relu=False)
    x = torch.rand(1, 512, 16, 16, 8, device=DEV)
    with torch.no_grad():
        got = conv.from_ndhwc(fused(conv.to_ndhwc(x)))
        want = F.conv3d(x.double(), layer.weight.double(), None, 1, 1)
        y32 = F.conv3d(x, layer.weight, None, 1, 1)
    scale = want.abs().max().item()
    print("one-signed K=13824: 3xTF32 %.2e, cuDNN fp32 %.2e of max|ref|" % (
        (got.double() - want).abs().max().item() / scale, (y32.double() - want).abs().max().item() / scale))
    assert_volume_close(got.cpu().numpy(), want.cpu().numpy())
