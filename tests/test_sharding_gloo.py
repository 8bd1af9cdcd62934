"""Host-side logic of the multi-GPU path on CPU: world_size-2 gloo processes.

Each rank computes the partial sums / counts of its view shard (with the CPU
oracle standing in for the kernel, which needs a GPU), the product's
``allreduce_partials`` adds them over gloo, and the result must equal the
unsharded mean.  Also covers the two partitioners."""
import os

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from imvoxelnet_b200 import scenes, sharding
from imvoxelnet_b200 import projection as pj
from oracle import backproject as ob
from tests.util import assert_volume_close


def test_partitioners():
    for n, world in [(50, 8), (6, 8), (1, 2), (7, 3), (0, 4)]:
        owned = [sharding.shard_views(n, r, world) for r in range(world)]
        assert sorted(v for o in owned for v in o) == list(range(n))
        assert max(len(o) for o in owned) - min(len(o) for o in owned) <= 1
        slices = [sharding.shard_scenes(n, r, world) for r in range(world)]
        covered = [i for s in slices for i in range(n)[s]]
        assert covered == list(range(n))
    assert sharding.shard_views(50, 1, 8) == [1, 9, 17, 25, 33, 41, 49]
    with pytest.raises(ValueError):
        sharding.shard_views(4, 2, 2)


def _scene():
    sc = scenes.scannet(n_views=5, seed=3)
    sc.n_voxels, sc.channels, sc.map_hw = (8, 7, 5), 8, (30, 40)
    sc.img_meta['img_shape'] = (120, 160, 3)
    full = scenes.random_features(sc, seed=4, channels_last=False)
    origin = sc.img_meta['lidar2img']['origin']
    P = pj.compute_projection(sc.img_meta, scenes.STRIDE)
    pts = ob.get_points(sc.n_voxels, sc.voxel_size, origin)
    return sc, full.numpy(), P.numpy(), pts


def _worker(rank, world, port, out):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        sc, full, P, pts = _scene()
        mine = sharding.shard_views(sc.n_views, rank, world)
        vol, valid = ob.backproject(full[mine], pts, P[mine])          # [v, C, X, Y, Z]
        s = torch.from_numpy(vol.sum(0))
        c = torch.from_numpy(valid.sum(0)[0].astype(np.int32))
        sharding.allreduce_partials(s, c)
        if rank == 0:
            np.savez(out, s=s.numpy(), c=c.numpy())
    finally:
        dist.destroy_process_group()


def test_view_sharded_sum_over_gloo(tmp_path):
    out = str(tmp_path / "reduced.npz")
    port = 29500 + os.getpid() % 2000
    mp.spawn(_worker, args=(2, port, out), nprocs=2, join=True)
    got = np.load(out)
    sc, full, P, pts = _scene()
    ref, ref_valid, ref_cnt = ob.backproject_mean(full, pts, P)
    assert np.array_equal(got["c"], ref_cnt)
    cnt = got["c"].astype(np.float32)
    mean = np.where(cnt > 0, got["s"] / np.maximum(cnt, 1), 0).astype(np.float32)
    assert_volume_close(mean, ref)
    assert np.array_equal(cnt > 0, ref_valid[0])


def _group_worker(rank, world, port, out):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        group, gi, lr = sharding.view_groups(2)
        assert (gi, lr) == (rank // 2, rank % 2) and dist.get_world_size(group) == 2
        sc, full, P, pts = _scene()
        full = full + gi                                   # every group has its own scene
        mine = sharding.shard_views(sc.n_views, lr, 2)
        vol, valid = ob.backproject(full[mine], pts, P[mine])
        s = torch.from_numpy(vol.sum(0))
        c = torch.from_numpy(valid.sum(0)[0].astype(np.int32))
        sharding.allreduce_partials(s, c, group)           # only inside the pair
        if lr == 0:
            np.savez(out % gi, s=s.numpy(), c=c.numpy())
    finally:
        dist.destroy_process_group()


def test_view_groups_reduce_inside_each_group_only(tmp_path):
    """4 gloo ranks = 2 pairs: the views of a scene are sharded over one pair, the pairs hold different scenes."""
    out = str(tmp_path / "group%d.npz")
    port = 31500 + os.getpid() % 2000
    mp.spawn(_group_worker, args=(4, port, out), nprocs=4, join=True)
    sc, full, P, pts = _scene()
    for gi in range(2):
        got = np.load(out % gi)
        ref, ref_valid, ref_cnt = ob.backproject_mean(full + gi, pts, P)
        assert np.array_equal(got["c"], ref_cnt)
        cnt = got["c"].astype(np.float32)
        mean = np.where(cnt > 0, got["s"] / np.maximum(cnt, 1), 0).astype(np.float32)
        assert_volume_close(mean, ref)
    with pytest.raises(ValueError):
        sharding.view_groups(3, world_size=8, rank=0)
