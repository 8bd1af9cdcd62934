"""Parity of the dense-head tail kernels (imvoxelnet_b200/heads.py, csrc/ivx_head.cu) with the reference.

Anchors: golden fixtures produced by the reference's own head classes
(tests/golden/head_*.npz, mask_*.npz), and the torch oracle (pinned to the
reference in tests/test_head_oracle.py) run on the GPU with TF32 disabled.
Tolerances: fp32 1e-5 relative (2e-5 on the exp()-ed distances: the exponent
carries the convolution's 1e-5), masks bit-exact, bf16 mode 2e-2."""
import copy

import numpy as np
import pytest
import torch

from imvoxelnet_b200 import heads
from oracle import torch_heads
from tests.golden.cases import HEAD_CASES, MASK_CASES
from tests.util import assert_volume_close, load_golden

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


@pytest.fixture(autouse=True)
def _no_tf32():
    old = (torch.backends.cudnn.allow_tf32, torch.backends.cuda.matmul.allow_tf32)
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    yield
    torch.backends.cudnn.allow_tf32, torch.backends.cuda.matmul.allow_tf32 = old


def _state_dict(g):
    return {k[3:]: torch.from_numpy(g[k]) for k in g.files if k.startswith("sd/")}


def _holder(name):
    _, (cls_name, kw), shapes = HEAD_CASES[name]
    kw = {k: v for k, v in kw.items() if k != "box_angle"}
    return (heads.HeadConvsV2 if cls_name == "HeadV2" else heads.HeadConvsV1)(**kw).eval(), shapes


@pytest.mark.parametrize("name", sorted(HEAD_CASES))
def test_head_forward_matches_reference_golden(name):
    g = load_golden(f"head_{name}")
    m, shapes = _holder(name)
    m.load_state_dict(_state_dict(g))
    m.to(DEV)
    before = heads.launch_count()
    with torch.no_grad():
        ctr, box, cls = m([torch.from_numpy(g[f"x{j}"]).to(DEV) for j in range(len(shapes))])
    assert heads.launch_count() - before == 3 * len(shapes)
    for j in range(len(shapes)):
        for got, key, rel in ((ctr[j], f"centerness{j}", 1e-5), (box[j], f"bbox{j}", 2e-5), (cls[j], f"cls{j}", 1e-5)):
            assert got.dtype == torch.float32 and got.is_contiguous() and tuple(got.shape) == g[key].shape
            assert_volume_close(got.cpu().numpy(), g[key], rel=rel)


@pytest.mark.parametrize("name", sorted(HEAD_CASES))
def test_head_forward_on_a_reference_shaped_module(name):
    """head_forward() only needs the reference's attribute names: drive it with the oracle module itself."""
    _, (cls_name, kw), shapes = HEAD_CASES[name]
    torch.manual_seed(3)
    ref = torch_heads.randomise(getattr(torch_heads, cls_name)(**kw), 9).eval().to(DEV)
    xs = [torch.randn(s, device=DEV) for s in shapes]
    with torch.no_grad():
        want = ref(xs)
        got = heads.head_forward(ref, xs)
    for u, v in zip(want, got):
        for p, q in zip(u, v):
            assert_volume_close(q.cpu().numpy(), p.cpu().numpy(), rel=2e-5)
    assert "_ivx_head_plan" not in ref.state_dict() and not any("ivx" in k for k in ref.state_dict())


def test_head_forward_follows_parameter_updates():
    m = torch_heads.randomise(heads.HeadConvsV2(5, 8, 6, 1), 1).eval().to(DEV)
    x = [torch.randn(1, 8, 6, 6, 4, device=DEV)]
    with torch.no_grad():
        a = m(x)[2][0].clone()
        m.cls_conv.bias.add_(1.0)
        m.scales[0].scale.mul_(0.5)
        b = m(x)
    assert_volume_close((b[2][0] - a).cpu().numpy(), np.ones(a.shape, np.float32), rel=1e-5)
    ref = torch_heads.HeadV2(5, 8, 6, 1).to(DEV).eval()
    ref.load_state_dict(m.state_dict())
    with torch.no_grad():
        assert_volume_close(b[1][0].cpu().numpy(), ref(x)[1][0].cpu().numpy(), rel=2e-5)


@pytest.mark.parametrize("version", ["v2", "v1"])
def test_head_forward_bf16_on_tensor_cores(version):
    """bf16 channels-last neck outputs at the configs' channel counts: the fused final layer runs on tcgen05
    (c_out padded to 64); compared with the fp32 oracle on the same bf16-rounded inputs."""
    torch.manual_seed(5)
    if version == "v2":     # imvoxelnet_scannet_fast.py: 18 classes, 128 channels
        ref = torch_heads.HeadV2(18, 128, 6, 3).eval()
        shapes = [(2, 128, 16, 16, 8), (2, 128, 8, 8, 4), (2, 128, 4, 4, 2)]
    else:                   # imvoxelnet_total_sunrgbd.py-like: 64 channels, 7 box outputs
        ref = torch_heads.HeadV1(33, 64, 1, 7, box_angle=True).eval()
        shapes = [(1, 64, 16, 16, 8), (1, 64, 8, 8, 4), (1, 64, 4, 4, 2)]
    torch_heads.randomise(ref, 2).to(DEV)
    xs = [torch.randn(s, device=DEV).to(torch.bfloat16).to(memory_format=torch.channels_last_3d) for s in shapes]
    with torch.no_grad():
        want = ref([x.float() for x in xs])
        got = heads.head_forward(ref, xs)
    for kind, (u, v) in enumerate(zip(want, got)):
        for p, q in zip(u, v):
            assert q.dtype == torch.float32
            if kind == 1:       # distances: exp() turns the bf16 rounding of the logit into a relative error
                p, q = p.clone(), q.clone()
                p[:, :6], q[:, :6] = p[:, :6].log(), q[:, :6].log()
            assert_volume_close(q.cpu().numpy(), p.cpu().numpy(), rel=2e-2)


@pytest.mark.parametrize("name", sorted(MASK_CASES))
@pytest.mark.parametrize("as_bool", [False, True])
def test_mask_resize_matches_reference_golden(name, as_bool):
    g = load_golden(f"mask_{name}")
    _, sizes = MASK_CASES[name]
    valid = torch.from_numpy(g["valid"]).to(DEV)
    if not as_bool:
        valid = valid.float()
    for j, size in enumerate(sizes):
        got = heads.resize_valid(valid, size)
        assert got.dtype == torch.bool and tuple(got.shape) == g[f"out{j}"].shape
        np.testing.assert_array_equal(got.cpu().numpy(), g[f"out{j}"])


def test_mask_resize_at_config_sizes_matches_torch_on_the_gpu():
    g = torch.Generator().manual_seed(4)
    for shape, sizes in (((4, 1, 80, 80, 32), [(80, 80, 32), (40, 40, 16), (20, 20, 8)]),
                         ((8, 1, 40, 40, 16), [(40, 40, 16), (20, 20, 8), (10, 10, 4)])):
        valid = (torch.rand(shape, generator=g) > 0.5).to(DEV)
        for size in sizes:
            want = torch.nn.Upsample(size=size, mode='trilinear')(valid.float()).round().bool()
            assert torch.equal(heads.resize_valid(valid, size), want)


@pytest.mark.parametrize("name", sorted(HEAD_CASES))
def test_head_scores_match_reference_golden(name):
    g = load_golden(f"head_{name}")
    ctr = torch.from_numpy(g["centerness0"][:1]).to(DEV)
    cls = torch.from_numpy(g["cls0"][:1]).to(DEV)
    valid = torch.from_numpy(g["score_valid"])[None].to(DEV)
    scores, best = heads.head_scores(ctr, cls, valid)
    assert_volume_close(scores[0].cpu().numpy(), g["scores"], rel=1e-6, atol=1e-7)
    assert_volume_close(best[0].cpu().numpy(), g["max_scores"], rel=1e-6, atol=1e-7)
    assert (scores[0].cpu().numpy()[~g["score_valid"].reshape(-1)] == 0).all()


def test_head_scores_batched_and_topk_agrees_with_the_reference_statements():
    torch.manual_seed(8)
    b, n, dims = 3, 18, (20, 20, 8)
    ctr, cls = torch.randn((b, 1) + dims, device=DEV), torch.randn((b, n) + dims, device=DEV)
    valid = torch.rand((b, 1) + dims, device=DEV) > 0.4
    scores, best = heads.head_scores(ctr, cls, valid)
    for i in range(b):
        s, m = torch_heads.scores(ctr[i], cls[i], valid[i])
        assert_volume_close(scores[i].cpu().numpy(), s.cpu().numpy(), rel=1e-6, atol=1e-7)
        # nms_pre selection (imvoxel_head_v2.py:272-276): same voxels up to ties within the tolerance
        ours, theirs = best[i].topk(1000)[0], m.topk(1000)[0]
        assert_volume_close(ours.cpu().numpy(), theirs.cpu().numpy(), rel=1e-6, atol=1e-7)


@pytest.mark.parametrize("version", ["v2", "v1"])
def test_head_training_gradients_match_autograd(version):
    torch.manual_seed(21)
    if version == "v2":
        ref = torch_heads.HeadV2(5, 8, 7, 2, box_angle=True)
        shapes = [(2, 8, 8, 6, 4), (2, 8, 4, 3, 2)]
    else:
        ref = torch_heads.HeadV1(4, 8, 1, 6, n_scales=2)
        shapes = [(2, 8, 8, 8, 4), (2, 8, 4, 4, 2)]
    torch_heads.randomise(ref, 6).to(DEV).train()
    ours = copy.deepcopy(ref)
    xs_a = [torch.randn(s, device=DEV, requires_grad=True) for s in shapes]
    xs_b = [x.detach().clone().requires_grad_(True) for x in xs_a]
    seeds = None

    def loss(outs):
        nonlocal seeds
        flat = [t for level in zip(*outs) for t in level]
        if seeds is None:
            seeds = [torch.randn_like(t) for t in flat]
        return sum((t * s).sum() for t, s in zip(flat, seeds))

    loss(ref(xs_a)).backward()
    loss(heads.head_forward(ours, xs_b)).backward()
    for a, b in zip(xs_a, xs_b):
        assert_volume_close(b.grad.cpu().numpy(), a.grad.cpu().numpy(), rel=1e-4)
    for (k, p), q in zip(ref.named_parameters(), ours.parameters()):
        assert q.grad is not None, k
        assert_volume_close(q.grad.cpu().numpy(), p.grad.cpu().numpy(), rel=1e-4)



def test_accelerate_routes_forward_and_survives_deepcopy():
    ref = torch_heads.randomise(torch_heads.HeadV2(4, 8, 6, 2), 12).eval().to(DEV)
    plain = copy.deepcopy(ref)
    assert heads.supports(ref) and not heads.supports(torch.nn.Conv3d(1, 1, 3))
    heads.accelerate(ref)
    xs = [torch.randn(1, 8, 6, 6, 4, device=DEV), torch.randn(1, 8, 3, 3, 2, device=DEV)]
    before = heads.launch_count()
    with torch.no_grad():
        got, want = ref(xs), plain(xs)
    assert heads.launch_count() - before == 6          # the module call went through the fused tail
    clone = copy.deepcopy(ref)
    assert clone.forward.__self__ is clone             # the copy runs on its own parameters
    with torch.no_grad():
        clone.cls_conv.bias.add_(2.0)
        moved = clone(xs)
    for u, v in zip(got, want):
        for p, q in zip(u, v):
            assert_volume_close(p.cpu().numpy(), q.cpu().numpy(), rel=2e-5)
    assert_volume_close((moved[2][0] - got[2][0]).cpu().numpy(), np.full(got[2][0].shape, 2.0, np.float32), rel=1e-5)
    assert list(ref.state_dict().keys()) == list(plain.state_dict().keys())
