"""Parity of the CUDA 3D necks with the reference (fp32 mode: 1e-5 relative).

Anchors: golden fixtures produced by the reference's own modules, and the
torch oracle (pinned to the reference in tests/test_neck_oracle.py) run on
the GPU in fp32 with TF32 disabled at larger sizes."""
import numpy as np
import pytest
import torch

from imvoxelnet_b200 import necks
from oracle import torch_necks
from tests.golden.cases import NECK_CASES
from tests.util import assert_volume_close, load_golden, rel_l2

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


@pytest.mark.parametrize("name", sorted(NECK_CASES))
def test_neck_matches_reference_golden(name):
    g = load_golden(f"neck_{name}")
    kw, shape = NECK_CASES[name]
    m = getattr(necks, name)(**kw).eval()
    m.load_state_dict(_state_dict(g))
    m.to(DEV)
    with torch.no_grad():
        outs = m(torch.from_numpy(g["x"]).to(DEV))
    n_out = sum(1 for k in g.files if k.startswith("out"))
    assert len(outs) == n_out
    for j, o in enumerate(outs):
        assert o.shape == g[f"out{j}"].shape and o.dtype == torch.float32
        assert_volume_close(o.cpu().numpy(), g[f"out{j}"])


BIGGER = {
    "FastIndoorImVoxelNeck": (dict(in_channels=64, n_blocks=[1, 1, 1], out_channels=32), (2, 64, 12, 8, 8)),
    "ImVoxelNeck": (dict(channels=[16, 32, 64, 128], out_channels=16, down_layers=[1, 2, 3, 4],
                         up_layers=[3, 2, 1], conditional=False), (1, 16, 16, 24, 8)),
    "KittiImVoxelNeck": (dict(in_channels=64, out_channels=256), (2, 64, 13, 18, 12)),
    "NuScenesImVoxelNeck": (dict(in_channels=64, out_channels=256), (1, 64, 16, 14, 12)),
}


@pytest.mark.parametrize("name", sorted(BIGGER))
def test_neck_matches_oracle_at_real_channel_counts(name):
    kw, shape = BIGGER[name]
    torch.manual_seed(11)
    ref = torch_necks.NECKS[name](**kw).eval()
    torch_necks.randomize_norms(ref, 5)
    m = getattr(necks, name)(**kw).eval()
    m.load_state_dict(ref.state_dict())
    ref.to(DEV), m.to(DEV)
    x = torch.randn(shape, device=DEV)
    with torch.no_grad():
        want = ref(x)
        got = m(x.contiguous(memory_format=torch.channels_last_3d))   # the layout backproject emits
    for a, b in zip(got, want):
        assert a.shape == b.shape
        assert_volume_close(a.cpu().numpy(), b.cpu().numpy())


def test_weights_are_repacked_after_a_state_dict_load():
    kw, shape = NECK_CASES["KittiImVoxelNeck"]
    m = necks.KittiImVoxelNeck(**kw).eval().to(DEV)
    x = torch.randn(shape, device=DEV)
    with torch.no_grad():
        a = m(x)[0].clone()
        g = load_golden("neck_KittiImVoxelNeck")
        m.load_state_dict(_state_dict(g))
        b = m(torch.from_numpy(g["x"]).to(DEV))[0]
    assert_volume_close(b.cpu().numpy(), g["out0"])
    assert not torch.equal(a, b)


# bf16 mode: bf16 activations between layers, tcgen05 kernels for every layer with
# channel counts that are multiples of 64.  Stated bar: 1e-2 of max|ref| per output
# (north_star), with a small atol for outputs that are mostly zero after the ReLU.
BF16_RTOL = 1e-2


@pytest.mark.parametrize("name", sorted(BIGGER))
def test_neck_bf16_mode(name):
    kw, shape = BIGGER[name]
    torch.manual_seed(13)
    ref = torch_necks.NECKS[name](**kw).eval()
    torch_necks.randomize_norms(ref, 6)
    m = getattr(necks, name)(**kw).eval()
    m.load_state_dict(ref.state_dict())
    m.precision = "bf16"
    ref.to(DEV), m.to(DEV)
    x = torch.randn(shape, device=DEV)
    with torch.no_grad():
        want = ref(x)
        got = m(x)
    for a, b in zip(got, want):
        assert a.dtype == torch.bfloat16 and a.shape == b.shape
        err = (a.float() - b).abs().max().item() / b.abs().max().item()
        l2 = rel_l2(a.float(), b)
        print(name, tuple(b.shape), "bf16 max err / max|ref| = %.4f, relative L2 = %.4f" % (err, l2))
        assert l2 <= 1e-2
        assert_volume_close(a.float().cpu().numpy(), b.cpu().numpy(), rel=BF16_RTOL, atol=1e-3)


# Full-size grids of the reference configs, one scene each, against the oracle run
# by cuDNN in fp32 with TF32 off.  Bars per precision, as (max-norm relative to max|ref|, relative L2):
#   fp32 / fp32_simt  1e-5 (+1e-6 absolute) -- north_star's first tolerance; "fp32" is the 3xTF32
#                     tensor-core path for every eligible layer, "fp32_simt" the CUDA-core cross-check
#   tf32              5e-3 max-norm, 2e-3 L2
#   bf16              1e-2 max-norm (+1e-3 absolute), 1e-2 L2
FULL = {
    "KittiImVoxelNeck": (dict(in_channels=64, out_channels=256), (1, 64, 216, 248, 12)),             # cfg2
    "FastIndoorImVoxelNeck": (dict(in_channels=256, n_blocks=[1, 1, 1], out_channels=128),
                              (2, 256, 40, 40, 16)),                                               # cfg1, cfg3
    "ImVoxelNeck": (dict(channels=[64, 128, 256, 512], out_channels=64, down_layers=[1, 2, 3, 4],
                         up_layers=[3, 2, 1], conditional=False), (1, 64, 80, 80, 32)),             # cfg5
    "NuScenesImVoxelNeck": (dict(in_channels=64, out_channels=256), (1, 64, 312, 312, 12)),         # cfg4 (reference grid)
}
L2_BAR = {"fp32": 1e-5, "fp32_simt": 1e-5, "tf32": 2e-3, "bf16": 1e-2}
_full_cache = {}


def _full_case(name):
    """(reference module's state dict, input, oracle outputs) of a full-size case, computed once."""
    if name not in _full_cache:
        _full_cache.clear()                       # one case resident at a time
        kw, shape = FULL[name]
        torch.manual_seed(3)
        ref = torch_necks.NECKS[name](**kw).eval()
        torch_necks.randomize_norms(ref, 9)
        ref.to(DEV)
        x = torch.randn(shape, device=DEV)
        with torch.no_grad():
            want = [w.clone() for w in ref(x)]
        _full_cache[name] = (ref.state_dict(), x, want)
        del ref
        torch.cuda.empty_cache()
    return _full_cache[name]


@pytest.mark.parametrize("precision", ["fp32", "fp32_simt", "tf32", "bf16"])
@pytest.mark.parametrize("name", list(FULL))
def test_neck_at_full_size(name, precision):
    sd, x, want = _full_case(name)
    m = getattr(necks, name)(**FULL[name][0]).eval()
    m.load_state_dict(sd)
    m.to(DEV)
    m.precision = precision
    with torch.no_grad():
        got = m(x.contiguous(memory_format=torch.channels_last_3d))   # the layout backproject emits
    torch.cuda.synchronize()
    assert len(got) == len(want)
    for a, b in zip(got, want):
        assert a.shape == b.shape
        assert a.dtype == (torch.bfloat16 if precision == "bf16" else torch.float32)
        a = a.float()
        worst = (a - b).abs().max().item() / b.abs().max().item()
        l2 = rel_l2(a, b)
        print(f"{name} {tuple(b.shape)} {precision}: max err / max|ref| = {worst:.2e}, relative L2 = {l2:.2e}")
        assert l2 <= L2_BAR[precision], f"relative L2 {l2:.3e} above the {precision} bar {L2_BAR[precision]:.0e}"
        if precision in ("fp32", "fp32_simt"):
            assert_volume_close(a.cpu().numpy(), b.cpu().numpy())
        elif precision == "tf32":
            assert_volume_close(a.cpu().numpy(), b.cpu().numpy(), rel=5e-3, atol=1e-4)
        else:
            assert_volume_close(a.cpu().numpy(), b.cpu().numpy(), rel=BF16_RTOL, atol=1e-3)
    if name == "KittiImVoxelNeck":
        assert got[0].shape == (1, 256, 246, 214)


def test_nuscenes_neck_on_a_small_odd_grid_bf16():
    """All-axis stride-2 layer and the pad (1,1,0) tail on a grid that does not fill the tiles."""
    name, kw, shape = "NuScenesImVoxelNeck", dict(in_channels=64, out_channels=256), (1, 64, 96, 104, 12)
    torch.manual_seed(5)
    ref = torch_necks.NECKS[name](**kw).eval()
    torch_necks.randomize_norms(ref, 11)
    m = getattr(necks, name)(**kw).eval()
    m.load_state_dict(ref.state_dict())
    m.precision = "bf16"
    ref.to(DEV), m.to(DEV)
    x = torch.randn(shape, device=DEV)
    with torch.no_grad():
        want = ref(x)
        got = m(x)
    assert len(got) == len(want)
    for a, b in zip(got, want):
        assert a.shape == b.shape
        assert rel_l2(a.float(), b) <= L2_BAR["bf16"]
        assert_volume_close(a.float().cpu().numpy(), b.cpu().numpy(), rel=BF16_RTOL, atol=1e-3)


def test_graphed_hot_path_replays_the_same_results():
    """CUDA-graph replay of back-projection + neck equals the eager calls."""
    from imvoxelnet_b200 import ops, scenes
    from imvoxelnet_b200 import projection as pj
    from imvoxelnet_b200.pipeline import GraphedHotPath
    sc = scenes.scannet(n_views=5, seed=3)
    sc.n_voxels, sc.channels = (16, 16, 8), 64
    origin = sc.img_meta['lidar2img']['origin']
    P = pj.compute_projection(sc.img_meta, scenes.STRIDE).to(DEV)[None]
    h, w = sc.feat_hw
    neck = necks.FastIndoorImVoxelNeck(64, [1, 1, 1], 64).eval().to(DEV)
    torch_necks.randomize_norms(neck, 2)
    neck.precision = "bf16"

    def step(feats, proj):
        vol, valid, _ = ops.backproject_scenes(feats, [(h, w)], proj, sc.n_voxels, sc.voxel_size, [origin],
                                               out_dtype=torch.bfloat16)
        return neck(vol), valid

    graphed = GraphedHotPath(step)
    for seed in (1, 2, 3):
        feats = scenes.random_features(sc, seed=seed, batch=1, channels_last=True).to(DEV)
        with torch.no_grad():
            want, want_valid = step(feats, P)
        got, got_valid = graphed(feats, P)
        assert torch.equal(got_valid, want_valid)
        for a, b in zip(got, want):
            # not bit-equal: the few-tile layers of this small neck use split-K, whose fp32
            # partial sums are added with atomics in arrival order
            assert_volume_close(a.float().cpu().numpy(), b.float().cpu().numpy(), rel=BF16_RTOL, atol=1e-3)
    assert len(graphed._graphs) == 1


@pytest.mark.parametrize("shape", [(2, 8, 5, 7, 3), (1, 64, 10, 10, 4), (1, 16, 1, 1, 1), (1, 24, 1, 6, 2),
                                   (3, 4, 4, 3, 5), (1, 12, 3, 4, 2)])
@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_upsample2x_trilinear_matches_interpolate(shape, dtype):
    """Both up-sampling kernels (16-byte channel groups per input voxel; the 4-channel per-output one for
    other channel counts) against F.interpolate(scale_factor=2, 'trilinear') (necks/imvoxelnet.py:359),
    including one-voxel axes and the clamped borders."""
    import torch.nn.functional as F
    from imvoxelnet_b200 import conv
    torch.manual_seed(13)
    x = torch.randn(shape, device=DEV)
    if dtype == torch.bfloat16:
        x = x.to(dtype).float()
    want = F.interpolate(x, scale_factor=2, mode='trilinear', align_corners=False)
    got = conv.from_ndhwc(conv.upsample2x_trilinear(conv.to_ndhwc(x.to(dtype))))
    assert got.shape == want.shape and got.dtype == dtype
    if dtype == torch.float32:
        assert_volume_close(got.cpu().numpy(), want.cpu().numpy(), rel=1e-6)
    else:   # one rounding to bf16 of the fp32 result
        assert_volume_close(got.float().cpu().numpy(), want.cpu().numpy(), rel=4e-3)


@pytest.mark.parametrize("precision,rel", [("fp32", 1e-5), ("bf16", BF16_RTOL)])
def test_conditional_projection(precision, rel):
    """ImVoxelNeck(conditional=True) against the oracle (pinned to the live reference in
    tests/test_neck_oracle.py), at toy and at tensor-core channel counts."""
    for channels, shape in (([4, 8, 16, 32], (1, 4, 16, 8, 8)), ([64, 128, 256, 512], (1, 64, 16, 16, 8))):
        if precision == "bf16" and channels[0] < 64:
            continue
        kw = dict(channels=channels, out_channels=channels[0], down_layers=[1, 2, 3, 4], up_layers=[3, 2, 1],
                  conditional=True)
        torch.manual_seed(21)
        ref = torch_necks.ImVoxelNeck(**kw).eval()
        torch_necks.randomize_norms(ref, 4)
        m = necks.ImVoxelNeck(**kw).eval()
        m.load_state_dict(ref.state_dict())
        m.precision = precision
        ref.to(DEV), m.to(DEV)
        x = torch.randn(shape, device=DEV)
        x[:, :, torch.rand(shape[2:], device=DEV) < 0.4] = 0
        with torch.no_grad():
            want, got = ref(x), m(x)
        for a, b in zip(got, want):
            assert_volume_close(a.float().cpu().numpy(), b.cpu().numpy(), rel=rel, atol=1e-6 if precision == "fp32" else 1e-3)


def test_train_mode_without_grad_uses_batch_statistics():
    """A neck left in .train() under torch.no_grad() behaves as nn.BatchNorm3d does: batch statistics
    and updated running buffers (a validation loop that forgot .eval())."""
    kw, shape = BIGGER["KittiImVoxelNeck"]
    torch.manual_seed(2)
    ref = torch_necks.KittiImVoxelNeck(**kw)
    torch_necks.randomize_norms(ref, 5)
    m = necks.KittiImVoxelNeck(**kw)
    m.load_state_dict(ref.state_dict())
    ref.to(DEV).train(), m.to(DEV).train()
    x = torch.randn(shape, device=DEV)
    with torch.no_grad():
        want, got = ref(x), m(x)
    assert_volume_close(got[0].cpu().numpy(), want[0].cpu().numpy(), rel=5e-5)
    a, b = m.state_dict(), ref.state_dict()
    for k in a:
        if "running_mean" in k or "running_var" in k:
            assert_volume_close(a[k].cpu().numpy(), b[k].cpu().numpy(), rel=5e-5)


def test_x3_neck_on_a_batch_matches_the_cuda_core_path():
    """The 3xTF32 tensor-core path against the CUDA-core kernel on a batch of four KITTI-grid scenes:
    persistent CTAs loop over ~80 tiles each and every tile sums many accumulation periods."""
    kw = FULL["KittiImVoxelNeck"][0]
    torch.manual_seed(12)
    m = necks.KittiImVoxelNeck(**kw).eval()
    torch_necks.randomize_norms(m, 7)
    m.to(DEV)
    x = torch.randn(4, 64, 216, 248, 12, device=DEV).contiguous(memory_format=torch.channels_last_3d)
    with torch.no_grad():
        m.precision = "fp32"
        got = m(x)[0]
        m.precision = "fp32_simt"
        want = m(x)[0]
    assert rel_l2(got, want) <= 1e-5
    assert_volume_close(got.cpu().numpy(), want.cpu().numpy())
