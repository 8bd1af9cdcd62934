"""Host-side pieces of bench.py that need no GPU: the peak chooser and the CPU arm's module source."""
import os
import sys

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import bench  # noqa: E402
from imvoxelnet_b200 import scenes  # noqa: E402
from imvoxelnet_b200 import projection as pj  # noqa: E402
from oracle import ref_loader, torch_port  # noqa: E402


def test_tensor_peak_follows_the_length_of_the_timed_region():
    peaks = {"bf16_tflops": 1600.0, "bf16_tflops_sustained": 1300.0}
    assert bench.tensor_peak(peaks, 280.0)[0] == 1600.0          # a 0.28 s replay never reaches the steady state
    assert "burst" in bench.tensor_peak(peaks, 280.0)[1]
    assert bench.tensor_peak(peaks, 1999.0)[0] == 1600.0
    assert bench.tensor_peak(peaks, 2000.0)[0] == 1300.0
    assert "sustained" in bench.tensor_peak(peaks, 5000.0)[1]
    assert bench.tensor_peak({"bf16_tflops": 1500.0}, 9000.0)[0] == 1500.0   # no sustained figure on file


def test_every_workload_has_a_scene_and_a_neck_entry():
    for name in bench.WORKLOADS:
        assert name in bench.NECKS
        sc, batch, P, origin = bench.scene_setup(name)
        assert P.shape == (sc.n_views, 3, 4) and batch >= 1 and len(origin) == 3
        cfg = bench.config_of(sc, batch, 2)
        assert cfg["global_scenes"] == 2 * batch and cfg["workload"] == sc.name


@pytest.mark.skipif(not ref_loader.available(), reason="reference tree not present on this machine")
def test_cpu_arm_uses_the_unmodified_reference_when_it_is_here():
    """bench.py --impl reference / cpu_baseline: with /root/reference on the machine the timed functions are the
    reference's own (kind "reference"), and they agree bit for bit with the port used elsewhere (kind "port")."""
    scene_volume, neck_class, kind, what = bench.reference_modules()
    assert kind == "reference" and "unmodified reference" in what
    sc = scenes.scannet(n_views=3, seed=5)
    sc.n_voxels, sc.channels, sc.map_hw = (6, 5, 4), 8, (30, 40)
    full = scenes.random_features(sc, seed=1, channels_last=False)
    h, w = sc.feat_hw
    pts = pj.get_points(sc.n_voxels, sc.voxel_size, sc.img_meta['lidar2img']['origin'])
    P = pj.compute_projection(sc.img_meta, scenes.STRIDE)
    a, av = scene_volume(full[:, :, :h, :w], pts, P)
    b, bv = torch_port.scene_volume(full[:, :, :h, :w], pts, P)
    assert torch.equal(a, b) and torch.equal(av, bv)
    assert neck_class("KittiImVoxelNeck").__module__ != "oracle.torch_necks"
