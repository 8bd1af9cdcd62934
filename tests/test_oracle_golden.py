"""Pin the CPU oracle to the reference.

The fixtures under tests/golden/ are outputs of the reference's own
functions (made by tests/golden/make_golden.py).  Pixel indices and valid
masks must match bit for bit, per-view volumes bit for bit (pure copies), the
view mean within 1e-6 (only the summation order over views may differ).
When /root/reference is present (build container) the oracle is additionally
checked against the live reference on fresh random scenes.
"""
import hashlib
import os

import numpy as np
import pytest
import torch

from imvoxelnet_b200 import projection as pj
from imvoxelnet_b200 import scenes
from oracle import backproject as ob
from oracle import ref_loader

SMALL = ["bp_ring5", "bp_kitti_small", "bp_nus6", "bp_ties"]


def _load(golden_dir, name):
    return np.load(os.path.join(golden_dir, name + ".npz"))


def _sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def assert_volume_close(got, ref, rel=1e-5, atol=1e-6):
    """north_star tolerance: 1e-5 relative, measured against max|ref| per
    tensor, plus an elementwise atol that absorbs cancellation."""
    scale = float(np.abs(ref).max()) or 1.0
    err = np.abs(got.astype(np.float64) - ref.astype(np.float64))
    assert err.max() <= rel * scale + atol, f"max err {err.max():.3e} (scale {scale:.3e})"


@pytest.mark.parametrize("name", SMALL)
def test_indices_bit_exact(golden_dir, name):
    g = _load(golden_dir, name)
    x, y, valid = ob.project(g["points"], g["projection"], int(g["h"]), int(g["w"]))
    ref_valid = g["valid_views"].reshape(valid.shape)
    assert np.array_equal(valid, ref_valid)
    # all lanes, not only valid ones: the oracle also restates the x86 cast
    assert np.array_equal(x, g["x"])
    assert np.array_equal(y, g["y"])


@pytest.mark.parametrize("name", SMALL)
def test_per_view_volume_bit_exact(golden_dir, name):
    g = _load(golden_dir, name)
    h, w = int(g["h"]), int(g["w"])
    feats = g["features"][:, :, :h, :w]          # strided slice, as :69
    vol, valid = ob.backproject(feats, g["points"], g["projection"])
    assert np.array_equal(valid, g["valid_views"])
    assert np.array_equal(vol, g["volume_views"])
    # channels-last physical layout must give the same values
    cl = np.ascontiguousarray(g["features"].transpose(0, 2, 3, 1)).transpose(0, 3, 1, 2)
    vol2, _ = ob.backproject(cl[:, :, :h, :w], g["points"], g["projection"])
    assert np.array_equal(vol2, vol)


@pytest.mark.parametrize("name", SMALL)
def test_view_mean(golden_dir, name):
    g = _load(golden_dir, name)
    h, w = int(g["h"]), int(g["w"])
    feats = g["features"][:, :, :h, :w]
    mean, valid, cnt = ob.view_mean(g["volume_views"], g["valid_views"])
    assert np.array_equal(valid, g["valid"])
    assert_volume_close(mean, g["volume"], rel=1e-6)
    assert not np.isnan(mean).any()
    assert np.all(mean[:, ~valid[0]] == 0)
    mean2, valid2, cnt2 = ob.backproject_mean(feats, g["points"], g["projection"])
    assert np.array_equal(valid2, valid) and np.array_equal(cnt2, cnt)
    assert np.array_equal(mean2, mean)


def test_ties_are_round_half_even(golden_dir):
    g = _load(golden_dir, "bp_ties")
    pts = g["points"].reshape(3, -1)
    x, y, valid = ob.project(g["points"], g["projection"], int(g["h"]), int(g["w"]))
    v0 = 0  # view 0 maps the point coordinate straight to the pixel coordinate
    sel = lambda cx, cy: np.where((pts[0] == cx) & (pts[1] == cy) & (pts[2] == 1.0))[0][0]
    n = sel(-0.5, 0.5)           # -0.5 -> -0 -> pixel 0, valid; 0.5 -> 0
    assert x[v0, n] == 0 and y[v0, n] == 0 and valid[v0, n]
    n = sel(2.5, 7.5)            # 2.5 -> 2, 7.5 -> 8
    assert x[v0, n] == 2 and y[v0, n] == 8 and valid[v0, n]
    n = sel(11.5, 3.5)           # 11.5 -> 12 == w: invalid
    assert x[v0, n] == 12 and not valid[v0, n]
    n = sel(10.5, 8.5)           # 8.5 -> 8 (< h=9) valid; 10.5 -> 10
    assert x[v0, n] == 10 and y[v0, n] == 8 and valid[v0, n]
    # view 1 uses z as depth: z <= 0 lanes are invalid, never NaN in the mean
    assert not valid[1][pts[2] <= 0].any()


def test_host_setup_golden(golden_dir):
    g = _load(golden_dir, "host_setup")
    for name in scenes.BUILDERS:
        sc = scenes.build(name)
        P = pj.compute_projection(sc.img_meta, scenes.STRIDE).numpy()
        assert P.dtype == np.float32 and np.array_equal(P, g[f"P_{name}"]), name
        origin = sc.img_meta['lidar2img']['origin']
        for pts in (pj.get_points(sc.n_voxels, sc.voxel_size, origin).numpy(),
                    ob.get_points(sc.n_voxels, sc.voxel_size, origin)):
            assert _sha(pts) == str(g[f"points_sha_{name}"]), name
            assert np.array_equal(pts[:, 0, 0, 0], g[f"points_first_{name}"])
            assert np.array_equal(pts[:, -1, -1, -1], g[f"points_last_{name}"])
        no = pj.grid_new_origin(sc.n_voxels, sc.voxel_size, origin).numpy()
        assert np.array_equal(no, ob.new_origin(sc.n_voxels, sc.voxel_size, origin))
        assert np.array_equal(no, g[f"points_first_{name}"])
    angles = torch.from_numpy(g["angles"])
    ext = torch.stack([pj.get_extrinsics(a) for a in angles]).numpy()
    assert np.array_equal(ext, g["extrinsics"])
    sc = scenes.build("total_sunrgbd")
    P = pj.compute_projection(sc.img_meta, scenes.STRIDE, angles[:1]).numpy()
    assert np.array_equal(P, g["P_total_angles"])


@pytest.mark.parametrize("name", ["sunrgbd_fast", "kitti", "scannet_fast", "nuscenes", "total_sunrgbd", "nuscenes_refgrid"])
def test_fullsize_digests(golden_dir, name):
    """BASELINE.json's full-size configurations against digests of the
    reference's outputs on the same seeded inputs."""
    g = _load(golden_dir, "fullsize_digests")
    sc = scenes.build(name)
    full = scenes.random_features(sc, seed=100, channels_last=False)
    assert _sha(full.numpy()) == str(g[f"{name}_features_sha"]), "seeded inputs drifted"
    h, w = sc.feat_hw
    feats = full.numpy()[:, :, :h, :w]
    origin = sc.img_meta['lidar2img']['origin']
    pts = ob.get_points(sc.n_voxels, sc.voxel_size, origin)
    P = pj.compute_projection(sc.img_meta, scenes.STRIDE).numpy()
    x, y, valid = ob.project(pts, P, h, w)
    assert int(valid.sum()) == int(g[f"{name}_n_valid"])
    assert _sha(valid) == str(g[f"{name}_valid_sha"])
    assert _sha(np.where(valid, x, 0)) == str(g[f"{name}_x_sha"])
    assert _sha(np.where(valid, y, 0)) == str(g[f"{name}_y_sha"])
    mean, val, _ = ob.backproject_mean(feats, pts, P)
    assert int(val.sum()) == int(g[f"{name}_covered"])
    v = mean.reshape(mean.shape[0], -1)
    ref_abs = g[f"{name}_chan_abs_sum"]
    assert np.allclose(v.astype(np.float64).sum(1), g[f"{name}_chan_sum"], rtol=0, atol=1e-6 * ref_abs.max())
    assert np.allclose(np.abs(v.astype(np.float64)).sum(1), ref_abs, rtol=1e-6)
    sample = v.reshape(-1)[g[f"{name}_sample_idx"]]
    assert_volume_close(sample, g[f"{name}_sample_val"], rel=1e-6)


@pytest.mark.skipif(not ref_loader.available(), reason="reference tree not present on this machine")
@pytest.mark.parametrize("seed", [1, 2, 3])
def test_against_live_reference(seed):
    """Fresh random rigs through the live reference (build container only)."""
    det, _ = ref_loader.load()
    rng = np.random.default_rng(seed)
    sc = scenes.scannet(n_views=int(rng.integers(1, 9)), seed=seed)
    sc.n_voxels = tuple(int(v) for v in rng.integers(3, 14, 3))
    sc.voxel_size = tuple(float(v) for v in rng.uniform(.2, .6, 3))
    sc.channels = int(rng.integers(1, 9))
    sc.img_meta['lidar2img']['origin'] = rng.normal(0, .3, 3).astype(np.float32)
    full = scenes.random_features(sc, seed=seed, channels_last=bool(seed % 2))
    h, w = sc.feat_hw
    h, w = h - int(rng.integers(0, 3)), w - int(rng.integers(0, 3))
    feats = full[:, :, :h, :w]
    origin = sc.img_meta['lidar2img']['origin']
    pts = pj.get_points(sc.n_voxels, sc.voxel_size, origin)
    P = pj.compute_projection(sc.img_meta, scenes.STRIDE)
    vol_ref, valid_ref = det.backproject(feats, pts, P)
    vol, valid = ob.backproject(feats.numpy(), pts.numpy(), P.numpy())
    assert np.array_equal(valid, valid_ref.numpy())
    assert np.array_equal(vol, vol_ref.numpy())


@pytest.mark.parametrize("name", SMALL)
def test_torch_port_matches_golden(golden_dir, name):
    """The torch-CPU port timed as the CPU baseline computes the reference's
    numbers exactly (same ATen ops)."""
    from oracle import torch_port
    g = _load(golden_dir, name)
    h, w = int(g["h"]), int(g["w"])
    feats = torch.from_numpy(g["features"])[:, :, :h, :w]
    pts, P = torch.from_numpy(g["points"]), torch.from_numpy(g["projection"])
    vol, valid = torch_port.backproject(feats, pts, P)
    assert np.array_equal(vol.numpy(), g["volume_views"])
    assert np.array_equal(valid.numpy(), g["valid_views"])
    mean, seen = torch_port.scene_volume(feats, pts, P)
    assert np.array_equal(seen.numpy(), g["valid"])
    assert np.array_equal(mean.numpy(), g["volume"])
