"""Parity of the CUDA back-projection with the reference (through the C ABI).

Three anchors:
  * the committed golden fixtures (outputs of the reference's own functions);
  * the CPU oracle on the same seeded inputs, at BASELINE.json's full sizes;
  * size-independent properties (ones -> valid mask, view permutation,
    partial + finalize == mean, count == sum of per-view valids).
Bars: pixel indices and valid masks bit-exact (indices on valid lanes);
per-view volumes bit-exact (they are copies); fp32 means within 1e-5
relative (tests/util.py).
"""
import numpy as np
import pytest
import torch

from imvoxelnet_b200 import ops, scenes
from imvoxelnet_b200 import projection as pj
from oracle import backproject as ob
from tests.util import assert_volume_close, load_golden, sha

pytestmark = pytest.mark.gpu

SMALL = ["bp_ring5", "bp_kitti_small", "bp_nus6", "bp_ties"]
FULL = ["sunrgbd_fast", "kitti", "scannet_fast", "nuscenes", "total_sunrgbd", "nuscenes_refgrid"]
DEV = "cuda:0"


def cuda(a):
    return torch.from_numpy(np.ascontiguousarray(a)).to(DEV)


def channels_last(t):
    return t.contiguous(memory_format=torch.channels_last)


@pytest.mark.parametrize("name", SMALL)
def test_golden_indices(name):
    g = load_golden(name)
    x, y, valid = ops.project_indices(cuda(g["points"]), g["projection"], int(g["h"]), int(g["w"]))
    valid = valid.cpu().numpy()
    ref_valid = g["valid_views"].reshape(valid.shape)
    assert np.array_equal(valid, ref_valid)
    assert np.array_equal(x.cpu().numpy()[valid], g["x"][valid])
    assert np.array_equal(y.cpu().numpy()[valid], g["y"][valid])


@pytest.mark.parametrize("layout", ["nchw_slice", "channels_last"])
@pytest.mark.parametrize("name", SMALL)
def test_golden_per_view_backproject(name, layout):
    """Drop-in ``backproject(features, points, projection)``: bit-exact."""
    g = load_golden(name)
    h, w = int(g["h"]), int(g["w"])
    full = cuda(g["features"])
    if layout == "channels_last":
        full = channels_last(full)
    feats = full[:, :, :h, :w]                      # the reference's slice (:69)
    vol, valid = ops.backproject(feats, cuda(g["points"]), cuda(g["projection"]))
    assert vol.shape == g["volume_views"].shape and valid.shape == g["valid_views"].shape
    assert valid.dtype == torch.bool and vol.dtype == torch.float32
    assert np.array_equal(valid.cpu().numpy(), g["valid_views"])
    assert np.array_equal(vol.cpu().numpy(), g["volume_views"])


@pytest.mark.parametrize("layout", ["nchw_slice", "channels_last"])
@pytest.mark.parametrize("name", SMALL)
def test_golden_scene_mean(name, layout):
    g = load_golden(name)
    h, w = int(g["h"]), int(g["w"])
    full = cuda(g["features"])
    if layout == "channels_last":
        full = channels_last(full)
    shape = g["points"].shape[1:]
    vol, valid, count = ops.backproject_scenes(
        full[None], [(h, w)], g["projection"][None], shape, None, None, points=cuda(g["points"]),
        channels_last=(layout == "channels_last"))
    assert np.array_equal(valid[0].cpu().numpy(), g["valid"])
    assert np.array_equal(count[0].cpu().numpy(), g["valid_views"].sum(0)[0])
    out = vol[0].cpu().numpy()
    assert not np.isnan(out).any()
    assert_volume_close(out, g["volume"])
    assert np.all(out[:, ~g["valid"][0]] == 0)


def _scene_inputs(name, seed=100):
    sc = scenes.build(name)
    full = scenes.random_features(sc, seed=seed, channels_last=True)
    origin = sc.img_meta['lidar2img']['origin']
    P = pj.compute_projection(sc.img_meta, scenes.STRIDE)
    return sc, full, origin, P


@pytest.mark.parametrize("name", FULL)
def test_fullsize_against_oracle_and_reference_digests(name):
    g = load_golden("fullsize_digests")
    sc, full, origin, P = _scene_inputs(name)
    assert sha(full.contiguous().numpy()) == str(g[f"{name}_features_sha"])
    h, w = sc.feat_hw
    x, y, valid = ops.project_indices(None, P, h, w, n_voxels=sc.n_voxels, voxel_size=sc.voxel_size,
                                      origin=origin, device=DEV)
    valid = valid.cpu().numpy()
    x = np.where(valid, x.cpu().numpy(), 0).astype(np.int64)
    y = np.where(valid, y.cpu().numpy(), 0).astype(np.int64)
    # digests of the reference's own outputs
    assert sha(valid) == str(g[f"{name}_valid_sha"])
    assert sha(x) == str(g[f"{name}_x_sha"])
    assert sha(y) == str(g[f"{name}_y_sha"])

    feats = full.to(DEV)
    vol, val, cnt = ops.backproject_scenes(feats[None], [(h, w)], P[None], sc.n_voxels,
                                           sc.voxel_size, [origin])
    pts = ob.get_points(sc.n_voxels, sc.voxel_size, origin)
    ref, ref_valid, ref_cnt = ob.backproject_mean(full.numpy()[:, :, :h, :w], pts, P.numpy())
    assert np.array_equal(val[0].cpu().numpy(), ref_valid)
    assert np.array_equal(cnt[0].cpu().numpy(), ref_cnt)
    out = vol[0].cpu().numpy()
    assert_volume_close(out, ref)
    sample = out.reshape(-1)[g[f"{name}_sample_idx"]]
    assert_volume_close(sample, g[f"{name}_sample_val"])
    # the reference layout ([B,C,X,Y,Z] contiguous) gives the same numbers
    vol2, val2, _ = ops.backproject_scenes(feats[None], [(h, w)], P[None], sc.n_voxels,
                                           sc.voxel_size, [origin], channels_last=False)
    assert vol2.is_contiguous()
    assert torch.equal(vol2, vol) and torch.equal(val2, val)


def test_grid_and_points_tensor_agree():
    sc, full, origin, P = _scene_inputs("sunrgbd_fast", seed=3)
    h, w = sc.feat_hw
    feats = full.to(DEV)[None]
    a = ops.backproject_scenes(feats, [(h, w)], P[None], sc.n_voxels, sc.voxel_size, [origin])
    pts = pj.get_points(sc.n_voxels, sc.voxel_size, origin).to(DEV)
    b = ops.backproject_scenes(feats, [(h, w)], P[None], sc.n_voxels, None, None, points=pts)
    for u, v in zip(a, b):
        assert torch.equal(u, v)


def test_batch_with_per_scene_origin_size_and_projection():
    """Scenes of one batch differ in origin, (h, w) and cameras (multi-scale
    training and RandomShiftOrigin in the reference's pipelines)."""
    rng = np.random.default_rng(7)
    base = scenes.scannet(n_views=7, seed=1)
    base.n_voxels, base.channels = (16, 12, 8), 32
    B = 5
    full = scenes.random_features(base, seed=9, batch=B, channels_last=True)
    metas, hws, origins, Ps = [], [], [], []
    for b in range(B):
        sc = scenes.scannet(n_views=7, seed=10 + b)
        origin = (rng.normal(0, .4, 3) + [0, 0, .5]).astype(np.float32)
        hh, ww = 120 - int(rng.integers(0, 9)), 160 - int(rng.integers(0, 9))
        sc.img_meta['img_shape'] = (hh * 4, ww * 4, 3)
        sc.img_meta['ori_shape'] = (hh * 4, ww * 4, 3)
        Ps.append(pj.compute_projection(sc.img_meta, 4))
        hws.append((hh, ww))
        origins.append(origin)
    P = torch.stack(Ps)
    vol, val, cnt = ops.backproject_scenes(full.to(DEV), hws, P, base.n_voxels, base.voxel_size, origins)
    for b in range(B):
        pts = ob.get_points(base.n_voxels, base.voxel_size, origins[b])
        hh, ww = hws[b]
        ref, rv, rc = ob.backproject_mean(full[b].numpy()[:, :, :hh, :ww], pts, P[b].numpy())
        assert np.array_equal(val[b].cpu().numpy(), rv)
        assert np.array_equal(cnt[b].cpu().numpy(), rc)
        assert_volume_close(vol[b].cpu().numpy(), ref)


@pytest.mark.parametrize("channels", [1, 3, 4, 6, 8, 20, 32, 64, 96, 128, 256, 384, 512, 1024])
def test_channel_counts(channels):
    """Every dispatch of the vectorised kernel, and the strided one for odd C."""
    sc = scenes.scannet(n_views=9, seed=4)
    sc.n_voxels, sc.channels, sc.map_hw = (9, 7, 5), channels, (24, 32)
    sc.img_meta['img_shape'] = (92, 124, 3)
    sc.img_meta['ori_shape'] = (480, 640, 3)
    full = scenes.random_features(sc, seed=channels, channels_last=True)
    origin = sc.img_meta['lidar2img']['origin']
    P = pj.compute_projection(sc.img_meta, 4)
    h, w = sc.feat_hw
    vol, val, cnt = ops.backproject_scenes(full.to(DEV)[None], [(h, w)], P[None], sc.n_voxels,
                                           sc.voxel_size, [origin])
    pts = ob.get_points(sc.n_voxels, sc.voxel_size, origin)
    ref, rv, rc = ob.backproject_mean(full.numpy()[:, :, :h, :w], pts, P.numpy())
    assert np.array_equal(val[0].cpu().numpy(), rv) and np.array_equal(cnt[0].cpu().numpy(), rc)
    assert rv.any()
    assert_volume_close(vol[0].cpu().numpy(), ref)


@pytest.mark.parametrize("n_views", [1, 2, 31, 32, 33, 64, 65, 100, 200])
def test_view_counts(n_views):
    sc = scenes.scannet(n_views=n_views, seed=n_views)
    sc.n_voxels, sc.channels, sc.map_hw = (10, 10, 4), 16, (30, 40)
    sc.img_meta['img_shape'] = (120, 160, 3)
    sc.img_meta['ori_shape'] = (480, 640, 3)
    full = scenes.random_features(sc, seed=1, channels_last=True)
    origin = sc.img_meta['lidar2img']['origin']
    P = pj.compute_projection(sc.img_meta, 4)
    h, w = sc.feat_hw
    feats = full.to(DEV)
    vol, val, cnt = ops.backproject_scenes(feats[None], [(h, w)], P[None], sc.n_voxels,
                                           sc.voxel_size, [origin])
    pts = ob.get_points(sc.n_voxels, sc.voxel_size, origin)
    ref, rv, rc = ob.backproject_mean(full.numpy(), pts, P.numpy())
    assert np.array_equal(cnt[0].cpu().numpy(), rc)
    assert_volume_close(vol[0].cpu().numpy(), ref)
    # per-view drop-in with more views than one launch holds (64)
    pv, pvalid = ops.backproject(feats, torch.from_numpy(pts).to(DEV), P.to(DEV))
    rvol, rvalid = ob.backproject(full.numpy(), pts, P.numpy())
    assert np.array_equal(pvalid.cpu().numpy(), rvalid)
    assert np.array_equal(pv.cpu().numpy(), rvol)
    assert np.array_equal(pvalid.sum(0)[0].cpu().numpy(), rc)


def test_ones_give_the_valid_mask_and_views_permute():
    sc, full, origin, P = _scene_inputs("scannet_fast", seed=5)
    h, w = sc.feat_hw
    ones = torch.ones_like(full).to(DEV)
    vol, val, cnt = ops.backproject_scenes(ones[None], [(h, w)], P[None], sc.n_voxels,
                                           sc.voxel_size, [origin])
    assert torch.equal(vol[0], val[0].float().expand_as(vol[0]))
    perm = torch.randperm(sc.n_views, generator=torch.Generator().manual_seed(0))
    feats = full.to(DEV)
    a = ops.backproject_scenes(feats[None], [(h, w)], P[None], sc.n_voxels, sc.voxel_size, [origin])
    b = ops.backproject_scenes(channels_last(feats[perm])[None], [(h, w)], P[perm][None],
                               sc.n_voxels, sc.voxel_size, [origin])
    assert torch.equal(a[1], b[1]) and torch.equal(a[2], b[2])
    assert_volume_close(b[0].cpu().numpy(), a[0].cpu().numpy())


def test_partial_then_finalize_equals_mean_and_shards_add_up():
    sc, full, origin, P = _scene_inputs("scannet_fast", seed=6)
    h, w = sc.feat_hw
    feats = full.to(DEV)
    mean, val, cnt = ops.backproject_scenes(feats[None], [(h, w)], P[None], sc.n_voxels,
                                            sc.voxel_size, [origin])
    s, _, c = ops.backproject_scenes(feats[None], [(h, w)], P[None], sc.n_voxels, sc.voxel_size,
                                     [origin], reduce="partial")
    m2, v2 = ops.finalize_mean(s, c)
    assert torch.equal(m2, mean) and torch.equal(v2, val) and torch.equal(c, cnt)
    # two "ranks" with interleaved views, summed like an all-reduce would
    tot_s, tot_c = None, None
    for r in range(2):
        idx = torch.arange(r, sc.n_views, 2)
        ps, _, pc = ops.backproject_scenes(channels_last(feats[idx])[None], [(h, w)], P[idx][None],
                                           sc.n_voxels, sc.voxel_size, [origin], reduce="partial")
        tot_s = ps if tot_s is None else tot_s + ps
        tot_c = pc if tot_c is None else tot_c + pc
    m3, v3 = ops.finalize_mean(tot_s, tot_c)
    assert torch.equal(tot_c, cnt) and torch.equal(v3, val)
    assert_volume_close(m3.cpu().numpy(), mean.cpu().numpy())
    # reference layout out of finalize
    m4, _ = ops.finalize_mean(s, c, channels_last=False)
    assert m4.is_contiguous() and torch.equal(m4, mean)


def test_bf16_volume_is_the_rounded_fp32_volume():
    sc, full, origin, P = _scene_inputs("kitti", seed=8)
    h, w = sc.feat_hw
    feats = full.to(DEV)
    f32, val, _ = ops.backproject_scenes(feats[None], [(h, w)], P[None], sc.n_voxels, sc.voxel_size,
                                         [origin])
    b16, val2, _ = ops.backproject_scenes(feats[None], [(h, w)], P[None], sc.n_voxels, sc.voxel_size,
                                          [origin], out_dtype=torch.bfloat16)
    assert b16.dtype == torch.bfloat16 and torch.equal(val, val2)
    assert torch.equal(b16, f32.to(torch.bfloat16))


def test_repack_is_exact_and_feeds_the_fast_kernel():
    g = torch.Generator().manual_seed(3)
    x = torch.randn(3, 37, 19, 45, generator=g).to(DEV)
    y = ops.repack_nchw_to_nhwc(x[:, :, :17, :41])
    assert y.shape == (3, 37, 17, 41) and y.stride(1) == 1
    assert torch.equal(y, x[:, :, :17, :41])


def test_too_many_views_is_a_clear_error():
    from imvoxelnet_b200 import _lib
    f = torch.zeros(1, 1025, 4, 4, 4, device=DEV)
    with pytest.raises(_lib.IvxError, match="n_views"):
        ops.backproject_scenes(f, [(4, 4)], torch.zeros(1, 1025, 3, 4), (2, 2, 2), (1., 1., 1.),
                               [(0., 0., 0.)])


@pytest.mark.parametrize("name", ["kitti", "scannet_fast"])
def test_nchw_maps_through_scratch_match_channels_last(name):
    """The reference hands NCHW maps: ``ivx_backproject_scenes_nchw`` re-lays
    them in a scratch workspace and must give the channels-last result exactly."""
    sc, full, origin, P = _scene_inputs(name, seed=21)
    h, w = sc.feat_hw
    hws = [(h, w), (h - 3, w - 5)]
    nchw = torch.stack([full.contiguous(), full.contiguous() * 0.5 + 1.0]).to(DEV)   # [2,V,C,Hf,Wf]
    assert nchw.stride(2) != 1
    Pb = torch.stack([P, P])
    origins = [origin, origin + np.float32(0.07)]
    a = ops.backproject_scenes(nchw, hws, Pb, sc.n_voxels, sc.voxel_size, origins)
    ref = []
    for b in range(2):
        ref.append(ops.backproject_scenes(channels_last(nchw[b])[None], [hws[b]], Pb[b][None],
                                          sc.n_voxels, sc.voxel_size, [origins[b]]))
    for b in range(2):
        for got, want in zip(a, ref[b]):
            assert torch.equal(got[b], want[0])
    # gathering from the strided maps directly (no scratch) agrees as well
    c = ops.backproject_scenes(nchw, hws, Pb, sc.n_voxels, sc.voxel_size, origins, repack=False)
    assert torch.equal(c[1], a[1]) and torch.equal(c[2], a[2])
    assert_volume_close(c[0].cpu().numpy(), a[0].cpu().numpy())


def test_nchw_workspace_is_checked():
    from imvoxelnet_b200 import _lib
    lib = _lib.lib()
    hw = _lib.i32([30, 40])
    need = lib.ivx_backproject_nchw_workspace_bytes(1, 2, 8, hw)
    assert need == 2 * 30 * 40 * 8 * 4
    f = torch.zeros(1, 2, 8, 32, 40, device=DEV)
    ws = torch.empty(need, dtype=torch.uint8, device=DEV)
    vol = torch.empty(1, 8, 8, device=DEV)
    g = _lib.grid((2, 2, 2), (1., 1., 1.))
    P = torch.zeros(1, 2, 12, device=DEV)
    args = (f.data_ptr(), _lib.i64(f.stride()), 1, 2, 8, hw, P.data_ptr(), None, g,
            _lib.f32([0, 0, 0]), 8, vol.data_ptr(), _lib.i64([64, 1, 8]), 0, 0, None, None)
    assert lib.ivx_backproject_scenes_nchw(*args, ws.data_ptr(), need - 1, None) == -1
    assert b"too small" in lib.ivx_last_error()
    assert lib.ivx_backproject_scenes_nchw(*args, ws.data_ptr(), need, None) == 0
    torch.cuda.synchronize()
    assert torch.count_nonzero(vol) == 0


def test_streamed_hot_path_equals_serial_calls():
    """Copies on side streams, double-buffered: same results as plain calls."""
    from imvoxelnet_b200.pipeline import StreamedHotPath
    sc, full, origin, P = _scene_inputs("sunrgbd_fast", seed=31)
    h, w = sc.feat_hw
    hws, origins = [(h, w)] * 2, [origin] * 2

    def compute(feats, proj):
        vol, valid, _ = ops.backproject_scenes(feats, hws, proj, sc.n_voxels, sc.voxel_size, origins)
        return [vol], valid

    batches = []
    for i in range(5):
        f = scenes.random_features(sc, seed=40 + i, batch=2, channels_last=True)
        pinned = StreamedHotPath._pinned_like(f)
        pinned.copy_(f)
        batches.append((pinned, torch.stack([P, P]).pin_memory()))
    runner = StreamedHotPath(compute)
    n = 0
    for i, (outs, valid) in enumerate(runner.run(iter(batches))):
        ref_outs, ref_valid = compute(batches[i][0].to(DEV), batches[i][1].to(DEV))
        assert torch.equal(outs[0], ref_outs[0].cpu()) and torch.equal(valid, ref_valid.cpu())
        n += 1
    assert n == 5


@pytest.mark.parametrize("name,channels", [("scannet_fast", 32), ("kitti", 64), ("nuscenes", 16)])
def test_gradient_wrt_features_matches_autograd_of_the_reference_ops(name, channels):
    """Training support (back-projection part): d(loss)/d(features)
    through the fused kernels == torch autograd through the reference's op sequence."""
    from oracle import torch_port
    sc = scenes.build(name)
    if name != "nuscenes":
        sc.n_voxels = tuple(max(4, v // 4) for v in sc.n_voxels)
        sc.voxel_size = tuple(v * 4 for v in sc.voxel_size)
    sc.channels = channels
    origin = sc.img_meta['lidar2img']['origin']
    P = pj.compute_projection(sc.img_meta, scenes.STRIDE)
    h, w = sc.feat_hw
    full = scenes.random_features(sc, seed=77, channels_last=True)
    g = torch.Generator().manual_seed(5)
    weight = torch.randn((channels,) + tuple(sc.n_voxels), generator=g)

    feats = full.to(DEV)[None].requires_grad_(True)
    vol, valid, cnt = ops.backproject_scenes(feats, [(h, w)], P[None], sc.n_voxels, sc.voxel_size, [origin])
    assert vol.requires_grad and not valid.requires_grad
    (vol[0] * weight.to(DEV)).sum().backward()
    got = feats.grad[0].cpu()

    ref_in = full.clone().contiguous().requires_grad_(True)
    pts = pj.get_points(sc.n_voxels, sc.voxel_size, origin)
    ref_vol, _ = torch_port.scene_volume(ref_in[:, :, :h, :w], pts, P)
    (ref_vol * weight).sum().backward()
    want = ref_in.grad
    assert got.shape == want.shape
    assert float(want.abs().max()) > 0
    assert_volume_close(got.numpy(), want.numpy())
    assert_volume_close(vol[0].detach().cpu().numpy(), ref_vol.detach().numpy())


@pytest.mark.parametrize("name", ["scannet_fast", "kitti", "nuscenes"])
def test_bf16_feature_maps_equal_the_widened_maps(name):
    """bf16 maps are widened exactly and summed in fp32: same result as fp32 maps holding
    the same (bf16-representable) values -- deep, wide and single-view schedules."""
    sc, full, origin, P = _scene_inputs(name, seed=41)
    h, w = sc.feat_hw
    maps16 = full.to(DEV).to(torch.bfloat16)[None]
    assert maps16.stride(2) == 1
    a = ops.backproject_scenes(maps16, [(h, w)], P[None], sc.n_voxels, sc.voxel_size, [origin])
    b = ops.backproject_scenes(maps16.float(), [(h, w)], P[None], sc.n_voxels, sc.voxel_size, [origin])
    for u, v in zip(a, b):
        assert torch.equal(u, v)
    c = ops.backproject_scenes(maps16, [(h, w)], P[None], sc.n_voxels, sc.voxel_size, [origin],
                               out_dtype=torch.bfloat16)
    assert torch.equal(c[0], b[0].to(torch.bfloat16))


def test_hw_beyond_the_feature_map_is_refused():
    """The reference's slice feature[:, :, :h, :w] (:69) clamps to the tensor; the kernels only see
    strides, so the Python mirror refuses an hw outside the map instead of reading out of bounds."""
    sc, full, origin, P = _scene_inputs("sunrgbd_fast", seed=3)
    feats = full.to(DEV)[None]
    hf, wf = feats.shape[3:]
    for bad in ((hf + 1, wf), (hf, wf + 1), (0, wf)):
        with pytest.raises(ValueError, match="outside the feature map"):
            ops.backproject_scenes(feats, [bad], P[None], sc.n_voxels, sc.voxel_size, [origin])
