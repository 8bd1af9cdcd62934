"""View-sharded back-projection over NCCL on two GPUs (skipped with fewer)."""
import os

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from imvoxelnet_b200 import scenes, sharding
from imvoxelnet_b200 import projection as pj
from tests.util import assert_volume_close

pytestmark = pytest.mark.gpu


def _inputs():
    sc = scenes.scannet(n_views=9, seed=2)
    sc.n_voxels, sc.channels = (20, 16, 8), 64
    full = scenes.random_features(sc, seed=3, batch=2, channels_last=True)
    P = pj.compute_projection(sc.img_meta, scenes.STRIDE)
    origin = sc.img_meta['lidar2img']['origin']
    return sc, full, torch.stack([P, P]), origin


def _cl5(x):
    b, v, c, h, w = x.shape
    y = x.reshape(b * v, c, h, w).contiguous(memory_format=torch.channels_last)
    return y.as_strided((b, v, c, h, w), (v * y.stride(0),) + tuple(y.stride()))


def _worker(rank, world, port, out):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    try:
        sc, full, P, origin = _inputs()
        mine = sharding.shard_views(sc.n_views, rank, world)
        feats = _cl5(full[:, mine]).cuda()
        h, w = sc.feat_hw
        vol, valid, cnt = sharding.backproject_scenes_view_sharded(
            feats, [(h, w)] * 2, P[:, mine].contiguous(), sc.n_voxels, sc.voxel_size, [origin] * 2)
        torch.cuda.synchronize()
        blob = dict(vol=vol.cpu().numpy(), valid=valid.cpu().numpy(), cnt=cnt.cpu().numpy())
        # the same step with the reduction fused into the finalize kernel (peer loads, then multicast)
        for name, use_mc in (("peer", False), ("mc", True)):
            try:
                fused = sharding.FusedViewSharded(2, sc.channels, sc.n_voxels, use_multicast=use_mc)
            except Exception as exc:   # symmetric memory may be unavailable in a sandbox
                blob[f"{name}_error"] = np.array(str(exc)[:200])
                continue
            for _ in range(2):         # twice: the buffers are reused
                v2, ok2, c2 = fused(feats, [(h, w)] * 2, P[:, mine].contiguous(), sc.voxel_size, [origin] * 2)
            torch.cuda.synchronize()
            blob[f"{name}_vol"], blob[f"{name}_valid"], blob[f"{name}_cnt"] = \
                v2.cpu().numpy(), ok2.cpu().numpy(), c2.cpu().numpy()
            blob[f"{name}_multicast"] = np.array(fused.multicast)
            v3, ok3, c3 = fused(feats, [(h, w)] * 2, P[:, mine].contiguous(), sc.voxel_size, [origin] * 2,
                                scatter_scenes=True)
            torch.cuda.synchronize()
            blob[f"{name}_slice_vol"], blob[f"{name}_slice_cnt"] = v3.cpu().numpy(), c3.cpu().numpy()
        np.savez(out % rank, **blob)
    finally:
        dist.destroy_process_group()


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs")
def test_view_sharded_equals_single_gpu(tmp_path):
    from imvoxelnet_b200 import ops
    out = str(tmp_path / "rank%d.npz")
    mp.spawn(_worker, args=(2, 29650 + os.getpid() % 1000, out), nprocs=2, join=True)
    sc, full, P, origin = _inputs()
    h, w = sc.feat_hw
    ref = ops.backproject_scenes(full.cuda(), [(h, w)] * 2, P, sc.n_voxels, sc.voxel_size, [origin] * 2)
    for rank in range(2):
        got = np.load(out % rank)
        assert np.array_equal(got["valid"], ref[1].cpu().numpy())
        assert np.array_equal(got["cnt"], ref[2].cpu().numpy())
        assert_volume_close(got["vol"], ref[0].cpu().numpy())
        for name in ("peer", "mc"):
            if f"{name}_error" in got.files:
                print(f"fused path ({name}) unavailable here:", got[f"{name}_error"])
                continue
            print(f"fused path ({name}): multicast used =", bool(got[f"{name}_multicast"]))
            assert np.array_equal(got[f"{name}_valid"], ref[1].cpu().numpy())
            assert np.array_equal(got[f"{name}_cnt"], ref[2].cpu().numpy())
            assert_volume_close(got[f"{name}_vol"], ref[0].cpu().numpy())
            # reduce-scatter flavour: rank r holds scene r of the two
            assert got[f"{name}_slice_vol"].shape[0] == 1
            assert np.array_equal(got[f"{name}_slice_cnt"][0], ref[2][rank].cpu().numpy())
            assert_volume_close(got[f"{name}_slice_vol"][0], ref[0][rank].cpu().numpy())
    a, b = np.load(out % 0), np.load(out % 1)
    for name in ("peer", "mc"):                       # replicated result: bit-identical on both ranks
        if f"{name}_vol" in a.files and f"{name}_vol" in b.files:
            assert np.array_equal(a[f"{name}_vol"], b[f"{name}_vol"])
