"""BatchNorm3d folding as a checkpoint transform (imvoxelnet_b200/checkpoint.py)."""
import copy

import pytest
import torch
from torch import nn

from imvoxelnet_b200 import checkpoint, necks
from oracle import ref_loader, torch_heads, torch_necks
from tests.golden.cases import NECK_CASES
from tests.util import assert_volume_close, load_golden


def _state_dict(g):
    return {k[3:]: torch.from_numpy(g[k]) for k in g.files if k.startswith("sd/")}


def _n_norms(m):
    return sum(isinstance(c, nn.BatchNorm3d) for c in m.modules())


@pytest.mark.parametrize("name", sorted(NECK_CASES))
def test_folded_oracle_neck_reproduces_the_reference_golden(name):
    """CPU: fold the oracle module (reference-shaped tree) and evaluate it with plain torch."""
    g = load_golden(f"neck_{name}")
    kw, _ = NECK_CASES[name]
    m = torch_necks.NECKS[name](**kw).eval()
    m.load_state_dict(_state_dict(g))
    n = _n_norms(m)
    assert checkpoint.fold_conv_bn3d(m) == n and _n_norms(m) == 0
    assert not any("running_mean" in k for k in m.state_dict())
    with torch.no_grad():
        outs = m(torch.from_numpy(g["x"]))
    for j, o in enumerate(outs):
        assert_volume_close(o.numpy(), g[f"out{j}"], rel=2e-5)


@pytest.mark.skipif(not ref_loader.available(), reason="reference tree not present on this machine")
def test_fold_walks_the_live_reference_modules():
    _, ref = ref_loader.load()
    v1, _ = ref_loader.load_heads()
    torch.manual_seed(2)
    for module, x in ((ref.KittiImVoxelNeck(4, 8), torch.randn(1, 4, 6, 6, 12)),
                      (ref.FastIndoorImVoxelNeck(8, [1, 1, 1], 8), torch.randn(1, 8, 8, 8, 4))):
        torch_necks.randomize_norms(module.eval(), 4)
        with torch.no_grad():
            want = module(x)
            folded = copy.deepcopy(module)
            assert checkpoint.fold_conv_bn3d(folded) == _n_norms(module) and _n_norms(folded) == 0
            got = folded(x)
        for a, b in zip(want, got):
            assert_volume_close(b.numpy(), a.numpy(), rel=2e-5)
    head = torch_heads.randomise(v1.ScanNetImVoxelHead(n_classes=4, n_channels=8, n_convs=2, n_reg_outs=6), 3).eval()
    xs = [torch.randn(1, 8, 4, 4, 4)] * 3
    with torch.no_grad():
        want = head(xs)
        assert checkpoint.fold_conv_bn3d(head) == 4
        got = head(xs)
    for u, v in zip(want, got):
        assert_volume_close(v[0].numpy(), u[0].numpy(), rel=2e-5)


def test_fold_state_dict_is_a_pure_file_level_transform():
    g = load_golden("neck_KittiImVoxelNeck")
    kw, _ = NECK_CASES["KittiImVoxelNeck"]
    sd = {"neck_3d." + k: v for k, v in _state_dict(g).items()}
    sd["backbone.conv1.weight"] = torch.ones(3)
    out = checkpoint.fold_state_dict(lambda: necks.KittiImVoxelNeck(**kw), sd, prefix="neck_3d.")
    assert out["backbone.conv1.weight"] is sd["backbone.conv1.weight"]
    assert not any("running_var" in k or "num_batches" in k for k in out)
    convs = [k for k in out if k.startswith("neck_3d.") and k.endswith(".weight")]
    assert all(k[:-len("weight")] + "bias" in out for k in convs)      # every folded conv carries a bias
    # a folded product neck loads it; an unfolded one does not
    m = necks.KittiImVoxelNeck(**kw)
    with pytest.raises(RuntimeError):
        m.load_state_dict({k[len("neck_3d."):]: v for k, v in out.items() if k.startswith("neck_3d.")})
    checkpoint.fold_conv_bn3d(m)
    m.load_state_dict({k[len("neck_3d."):]: v for k, v in out.items() if k.startswith("neck_3d.")})


@pytest.mark.gpu
@pytest.mark.parametrize("name", sorted(NECK_CASES))
@pytest.mark.parametrize("precision", ["fp32", "bf16"])
def test_folded_product_neck_matches_the_unfolded_one(name, precision):
    g = load_golden(f"neck_{name}")
    kw, _ = NECK_CASES[name]
    m = getattr(necks, name)(**kw).eval()
    m.load_state_dict(_state_dict(g))
    m.to("cuda:0")
    m.precision = precision
    x = torch.from_numpy(g["x"]).to("cuda:0")
    with torch.no_grad():
        want = [o.float().clone() for o in m(x)]
        assert checkpoint.fold_conv_bn3d(m) > 0
        got = m(x)
    for a, b in zip(want, got):
        assert_volume_close(b.float().cpu().numpy(), a.cpu().numpy(), rel=2e-5 if precision == "fp32" else 2e-2)
    if precision == "fp32":
        for j, o in enumerate(got):
            assert_volume_close(o.cpu().numpy(), g[f"out{j}"], rel=2e-5)


def test_fold_refuses_a_conditional_projection():
    """ConditionalProjection(condition=True) applies its norm to where(mask, conv(x), y)
    (necks/imvoxelnet.py:289-291): folding the norm into the convolution would be wrong."""
    import pytest

    from imvoxelnet_b200 import checkpoint, necks
    m = necks.ImVoxelNeck([4, 8, 16, 32], 4, [1, 1, 1, 1], [1, 1, 1], True)
    with pytest.raises(ValueError, match="condition=True"):
        checkpoint.fold_conv_bn3d(m)
    assert checkpoint.fold_conv_bn3d(necks.ImVoxelNeck([4, 8, 16, 32], 4, [1, 1, 1, 1], [1, 1, 1], False)) > 0
