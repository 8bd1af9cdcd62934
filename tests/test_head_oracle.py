"""Pin the torch restatement of the dense-head tail (oracle/torch_heads.py) to the
reference: golden fixtures made from the reference head classes, and the live
classes when /root/reference is present."""
import numpy as np
import pytest
import torch

from oracle import ref_loader, torch_heads
from tests.golden.cases import HEAD_CASES, MASK_CASES
from tests.util import assert_volume_close, load_golden


def _state_dict(g):
    return {k[3:]: torch.from_numpy(g[k]) for k in g.files if k.startswith("sd/")}


def _oracle_head(name):
    _, (cls_name, kw), shapes = HEAD_CASES[name]
    return getattr(torch_heads, cls_name)(**kw).eval(), shapes


@pytest.mark.parametrize("name", sorted(HEAD_CASES))
def test_oracle_head_matches_golden(name):
    g = load_golden(f"head_{name}")
    head, shapes = _oracle_head(name)
    sd = _state_dict(g)
    assert list(head.state_dict().keys()) == list(sd.keys())
    head.load_state_dict(sd)
    with torch.no_grad():
        ctr, box, cls = head([torch.from_numpy(g[f"x{j}"]) for j in range(len(shapes))])
    for j in range(len(shapes)):
        assert_volume_close(ctr[j].numpy(), g[f"centerness{j}"])
        assert_volume_close(box[j].numpy(), g[f"bbox{j}"])
        assert_volume_close(cls[j].numpy(), g[f"cls{j}"])
    s, m = torch_heads.scores(torch.from_numpy(g["centerness0"][0]), torch.from_numpy(g["cls0"][0]),
                              torch.from_numpy(g["score_valid"]))
    np.testing.assert_array_equal(s.numpy(), g["scores"])
    np.testing.assert_array_equal(m.numpy(), g["max_scores"])


@pytest.mark.parametrize("name", sorted(MASK_CASES))
def test_oracle_mask_resize_matches_golden(name):
    g = load_golden(f"mask_{name}")
    _, sizes = MASK_CASES[name]
    valid = torch.from_numpy(g["valid"]).float()
    for j, size in enumerate(sizes):
        np.testing.assert_array_equal(torch_heads.resize_valid(valid, size).numpy(), g[f"out{j}"])


def test_mask_resize_by_two_is_a_majority_vote_of_eight():
    """With the configs' integer ratios the trilinear value is the mean of a 2x2x2 block around the
    sample point, so the rounded mask is 'more than half of the eight' (round-half-even sends 4/8 to 0)."""
    g = load_golden("mask_indoor_fast")
    v = g["valid"].astype(np.float32)[:, 0]
    b, X, Y, Z = v.shape
    blocks = v.reshape(b, X // 2, 2, Y // 2, 2, Z // 2, 2).sum(axis=(2, 4, 6))
    np.testing.assert_array_equal(blocks > 4, g["out1"][:, 0])
    inner = v[:, 1::4][:, :, 1::4][:, :, :, 1::4] + 0    # voxels 4d+1 and 4d+2 on every axis
    acc = np.zeros_like(inner)
    for dx in (1, 2):
        for dy in (1, 2):
            for dz in (1, 2):
                acc += v[:, dx::4][:, :, dy::4][:, :, :, dz::4]
    np.testing.assert_array_equal(acc > 4, g["out2"][:, 0])


@pytest.mark.skipif(not ref_loader.available(), reason="reference tree not present on this machine")
@pytest.mark.parametrize("name", sorted(HEAD_CASES))
def test_oracle_head_matches_live_reference(name):
    v1, v2 = ref_loader.load_heads()
    kw = HEAD_CASES[name][0]
    torch.manual_seed(11)
    ref = getattr(v2 if name.endswith("V2") else v1, name)(**kw).eval()
    torch_heads.randomise(ref, 5)
    head, shapes = _oracle_head(name)
    head.load_state_dict(ref.state_dict())
    xs = [torch.randn(s) for s in shapes]
    with torch.no_grad():
        a, b = ref(xs), head(xs)
    for u, v in zip(a, b):
        for p, q in zip(u, v):
            assert torch.equal(p, q)


def test_product_head_holders_share_the_reference_state_dict_layout():
    from imvoxelnet_b200 import heads
    for name, (_, (cls_name, kw), _) in HEAD_CASES.items():
        sd = _state_dict(load_golden(f"head_{name}"))
        kw = {k: v for k, v in kw.items() if k != "box_angle"}
        m = (heads.HeadConvsV2 if cls_name == "HeadV2" else heads.HeadConvsV1)(**kw)
        assert list(m.state_dict().keys()) == list(sd.keys()), name
        m.load_state_dict(sd)


def test_head_ops_refuse_cpu_tensors():
    from imvoxelnet_b200 import _lib, heads
    m = heads.HeadConvsV2(3, 8, 6, 1).eval()
    with pytest.raises(_lib.IvxError):
        m([torch.zeros(1, 8, 4, 4, 4)])
    with pytest.raises(_lib.IvxError):
        heads.resize_valid(torch.zeros(1, 1, 4, 4, 4), (2, 2, 2))
    with pytest.raises(_lib.IvxError):
        heads.head_scores(torch.zeros(1, 1, 2, 2, 2), torch.zeros(1, 3, 2, 2, 2),
                          torch.zeros(1, 1, 2, 2, 2, dtype=torch.bool))
