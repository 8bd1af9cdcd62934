"""Training mode of the drop-in necks: forward with batch-statistics
BatchNorm, running-statistics update and every gradient against PyTorch autograd through
the oracle modules (cuDNN fp32, TF32 off) with the same parameters.

Bars: gradients element-wise within 1e-4 of max|ref|, or -- where ReLU-mask flips make fp32 runs
of the same network differ -- as close to a float64 run as PyTorch's own fp32 run is (x3).
Forward and running statistics 5e-5 of max|ref| (batch-statistics normalisation divides by
the batch standard deviation of as few as 32 samples at the coarsest level, which amplifies the
1e-7 rounding differences of the convolutions; eval mode keeps 1e-5); gradients 1e-4 of max|ref|
per tensor (they pass through up to 44 layers and atomics-ordered fp32 reductions)."""
import copy

import pytest
import torch

from imvoxelnet_b200 import necks
from oracle import torch_necks
from tests.golden.cases import NECK_CASES
from tests.util import assert_volume_close

import numpy as np

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
GRAD_RTOL = 1e-4
FWD_RTOL = 5e-5


def l2_rel(a, b):
    a, b = np.asarray(a, np.float64), np.asarray(b, np.float64)
    return float(np.linalg.norm(a - b) / max(np.linalg.norm(b), 1e-30))


def assert_grad_close(got, ref32, ref64, what=""):
    """Gradients through batch-statistics BN and ReLU are ill-conditioned in fp32: a pre-activation
    within rounding of zero flips its ReLU mask, and through the batch sums of the BN backward that
    one flip moves every element a little.  So the bar is set by a float64 run of the same network:
    our fp32 gradient must be as close to it (relative L2) as PyTorch/cuDNN's fp32 gradient is, up to
    a factor 3 -- or simply within GRAD_RTOL of the fp32 reference element-wise."""
    got, ref32 = np.asarray(got, np.float64), np.asarray(ref32, np.float64)
    scale = float(np.abs(ref32).max()) or 1.0
    if float(np.abs(got - ref32).max()) <= GRAD_RTOL * scale + 1e-6:
        return
    ours, theirs = l2_rel(got, ref64), l2_rel(ref32, ref64)
    assert ours <= max(3 * theirs, 1e-5), \
        f"{what}: relative L2 error vs float64 {ours:.2e}, PyTorch fp32 has {theirs:.2e}"


CASES = dict(NECK_CASES)
CASES["ImVoxelNeck"] = (NECK_CASES["ImVoxelNeck"][0], (2, 4, 32, 16, 16))   # 32 voxels at the coarsest level
CASES["KittiImVoxelNeck_64"] = (dict(in_channels=64, out_channels=64), (2, 64, 16, 18, 12))


@pytest.fixture(autouse=True)
def _no_tf32():
    old = (torch.backends.cudnn.allow_tf32, torch.backends.cuda.matmul.allow_tf32)
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    yield
    torch.backends.cudnn.allow_tf32, torch.backends.cuda.matmul.allow_tf32 = old


@pytest.mark.parametrize("case", sorted(CASES))
def test_training_forward_backward_match_autograd(case):
    name = case.split("_")[0]
    kw, shape = CASES[case]
    torch.manual_seed(21)
    ref = torch_necks.NECKS[name](**kw)
    torch_necks.randomize_norms(ref, 7)
    mine = getattr(necks, name)(**kw)
    mine.load_state_dict(ref.state_dict())
    ref.to(DEV).train(), mine.to(DEV).train()
    x = torch.randn(shape, device=DEV)
    ref64 = copy.deepcopy(ref).double()
    xr = x.clone().requires_grad_(True)
    xm = x.clone().requires_grad_(True)
    xd = x.double().requires_grad_(True)
    want = ref(xr)
    got = mine(xm)
    want64 = ref64(xd)
    assert len(got) == len(want)
    loss_w = loss_g = loss_d = 0.
    for i, (a, b) in enumerate(zip(got, want)):
        assert a.shape == b.shape and a.requires_grad
        assert_volume_close(a.detach().cpu().numpy(), b.detach().cpu().numpy(), rel=FWD_RTOL)
        w = torch.randn(b.shape, device=DEV, generator=torch.Generator(DEV).manual_seed(i))
        loss_w = loss_w + (b * w).sum()
        loss_g = loss_g + (a * w).sum()
        loss_d = loss_d + (want64[i] * w.double()).sum()
    loss_w.backward()
    loss_g.backward()
    loss_d.backward()
    assert_grad_close(xm.grad.cpu().numpy(), xr.grad.cpu().numpy(), xd.grad.cpu().numpy(), "input")
    dp = dict(ref64.named_parameters())
    rp, mp = dict(ref.named_parameters()), dict(mine.named_parameters())
    assert rp.keys() == mp.keys()
    checked = 0
    # a conv bias in front of a batch-statistics BN has an analytically zero gradient: both sides
    # return rounding noise, so errors are measured against at least 1 % of the largest gradient
    floor = 1e-2 * max(float(p.grad.abs().max()) for p in rp.values() if p.grad is not None)
    for k, p in rp.items():
        if p.grad is None:                       # e.g. ImVoxelNeck.conv_blocks.3 is never executed
            assert mp[k].grad is None or float(mp[k].grad.abs().max()) == 0, k
            continue
        assert mp[k].grad is not None, k
        scale = float(p.grad.abs().max())
        if scale < floor:                        # rounding noise around an (analytically) zero gradient
            assert float((mp[k].grad - p.grad).abs().max()) <= GRAD_RTOL * floor + 1e-6, k
        else:
            assert_grad_close(mp[k].grad.cpu().numpy(), p.grad.cpu().numpy(), dp[k].grad.cpu().numpy(), k)
        checked += 1
    assert checked > 10
    # running statistics were updated like nn.BatchNorm3d does
    rb, mb = dict(ref.named_buffers()), dict(mine.named_buffers())
    for k, v in rb.items():
        if k.endswith("num_batches_tracked"):
            assert int(mb[k]) == int(v), k
        else:
            assert_volume_close(mb[k].cpu().numpy(), v.cpu().numpy(), rel=FWD_RTOL)


def test_eval_mode_with_input_gradient_uses_running_statistics():
    """Frozen BatchNorm (eval) with a gradient wanted for the input: fine-tuning style."""
    kw, shape = NECK_CASES["KittiImVoxelNeck"]
    torch.manual_seed(3)
    ref = torch_necks.KittiImVoxelNeck(**kw)
    torch_necks.randomize_norms(ref, 2)
    mine = necks.KittiImVoxelNeck(**kw)
    mine.load_state_dict(ref.state_dict())
    ref.to(DEV).eval(), mine.to(DEV).eval()
    before = copy.deepcopy(mine.state_dict())
    x = torch.randn(shape, device=DEV)
    xr, xm = x.clone().requires_grad_(True), x.clone().requires_grad_(True)
    xd = x.double().requires_grad_(True)
    want, got = ref(xr)[0], mine(xm)[0]
    assert_volume_close(got.detach().cpu().numpy(), want.detach().cpu().numpy())
    want.square().sum().backward()
    got.square().sum().backward()
    copy.deepcopy(ref).double()(xd)[0].square().sum().backward()
    assert_grad_close(xm.grad.cpu().numpy(), xr.grad.cpu().numpy(), xd.grad.cpu().numpy(), "input")
    for k, v in mine.state_dict().items():
        assert torch.equal(v, before[k]), k      # nothing updated in eval mode
