"""Pin the torch restatement of the necks (oracle/torch_necks.py) to the
reference: golden fixtures made from the reference modules, and the live
reference modules when /root/reference is present."""
import numpy as np
import pytest
import torch

from oracle import ref_loader, torch_necks
from tests.golden.cases import NECK_CASES  # the construction arguments of the fixtures
from tests.util import assert_volume_close, load_golden


def _state_dict(g):
    return {k[3:]: torch.from_numpy(g[k]) for k in g.files if k.startswith("sd/")}


@pytest.mark.parametrize("name", sorted(NECK_CASES))
def test_oracle_neck_matches_golden(name):
    g = load_golden(f"neck_{name}")
    kw, shape = NECK_CASES[name]
    m = torch_necks.NECKS[name](**kw).eval()
    sd = _state_dict(g)
    assert list(m.state_dict().keys()) == list(sd.keys())
    m.load_state_dict(sd)
    with torch.no_grad():
        outs = m(torch.from_numpy(g["x"]))
    for j, o in enumerate(outs):
        assert_volume_close(o.numpy(), g[f"out{j}"])


@pytest.mark.skipif(not ref_loader.available(), reason="reference tree not present on this machine")
@pytest.mark.parametrize("name", sorted(NECK_CASES))
def test_oracle_neck_matches_live_reference(name):
    _, ref = ref_loader.load()
    kw, shape = NECK_CASES[name]
    torch.manual_seed(7)
    r = getattr(ref, name)(**kw).eval()
    torch_necks.randomize_norms(r, 3)
    m = torch_necks.NECKS[name](**kw).eval()
    m.load_state_dict(r.state_dict())
    x = torch.randn(shape)
    with torch.no_grad():
        a, b = r(x), m(x)
    assert len(a) == len(b)
    for u, v in zip(a, b):
        assert torch.equal(u, v)


def test_product_necks_share_the_reference_state_dict_layout():
    """The drop-in modules must load the reference's checkpoints: same keys and shapes."""
    from imvoxelnet_b200 import necks
    for name, (kw, _) in NECK_CASES.items():
        g = load_golden(f"neck_{name}")
        sd = _state_dict(g)
        m = getattr(necks, name)(**kw)
        assert list(m.state_dict().keys()) == list(sd.keys()), name
        m.load_state_dict(sd)
    full = {"FastIndoorImVoxelNeck": (dict(in_channels=256, n_blocks=[1, 1, 1], out_channels=128), 90),
            "KittiImVoxelNeck": (dict(in_channels=64, out_channels=256), 57),
            "NuScenesImVoxelNeck": (dict(in_channels=64, out_channels=256), 57)}
    for name, (kw, n_keys) in full.items():
        with torch.device("meta"):
            m = getattr(necks, name)(**kw)
        assert len(m.state_dict()) == n_keys


def test_necks_refuse_cpu_and_training_mode():
    from imvoxelnet_b200 import _lib, necks
    m = necks.KittiImVoxelNeck(4, 8)
    m.precision = "bf16"                      # training is fp32 only
    with pytest.raises(NotImplementedError):
        m.train()(torch.zeros(1, 4, 4, 4, 12))
    m.precision = "fp32"
    with pytest.raises(_lib.IvxError):
        m.eval()(torch.zeros(1, 4, 4, 4, 12))


def _conditional_case():
    kw = dict(channels=[4, 8, 16, 32], out_channels=4, down_layers=[1, 2, 3, 4], up_layers=[3, 2, 1],
              conditional=True)
    torch.manual_seed(21)
    x = torch.randn(1, 4, 16, 8, 8)
    x[:, :, torch.rand(16, 8, 8) < 0.4] = 0          # unobserved voxels: every channel zero (:347-348)
    return kw, x


@pytest.mark.skipif(not ref_loader.available(), reason="reference tree not present on this machine")
def test_oracle_conditional_projection_matches_live_reference():
    """ImVoxelNeck(conditional=True) (necks/imvoxelnet.py:281-294, :347-348, :361-363): no reference
    config sets it, but it is a constructor argument of the drop-in class."""
    _, ref = ref_loader.load()
    kw, x = _conditional_case()
    r = ref.ImVoxelNeck(**kw).eval()
    torch_necks.randomize_norms(r, 4)
    m = torch_necks.ImVoxelNeck(**kw).eval()
    m.load_state_dict(r.state_dict())
    with torch.no_grad():
        a, b = r(x), m(x)
        plain = torch_necks.ImVoxelNeck(**dict(kw, conditional=False)).eval()
        plain.load_state_dict(r.state_dict())
        c = plain(x)
    for u, v in zip(a, b):
        assert torch.equal(u, v)
    assert not torch.equal(a[0], c[0])               # the condition changes the result


def test_detector_module_mirrors_the_reference_surface():
    """detectors/imvoxelnet.py exposes get_points (:133), backproject (:145), get_extrinsics (:164)
    at module level and ImVoxelNet._compute_projection as a static method (:114)."""
    import inspect

    from imvoxelnet_b200 import detector, ops, projection, scenes
    assert detector.backproject is ops.backproject
    assert detector.get_extrinsics is projection.get_extrinsics
    assert detector.get_points is projection.get_points
    assert isinstance(inspect.getattr_static(detector.ImVoxelNet, "_compute_projection"), staticmethod)
    sc = scenes.build("scannet_fast")
    a = detector.ImVoxelNet._compute_projection(sc.img_meta, 4, None)
    assert torch.equal(a, projection.compute_projection(sc.img_meta, 4, None))
