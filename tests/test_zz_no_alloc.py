"""ivx_backproject_scenes_nchw_noalloc: the entry point for callers that cannot
allocate device memory (host applications that own all GPU memory through a
pool of their own, threads that must stay clear of the driver's allocation
lock).  Its scratch is a fixed 32 MiB array inside the library.

Checked here, through the C ABI: its volume, mask and counts equal those of
ivx_backproject_scenes_nchw (which takes the scratch from the caller) on two
calls in a row -- the second with other maps and a smaller region, so stale
contents of the shared array would show -- and a request beyond the 32 MiB is
refused without touching the outputs.  This file is named to be collected
last; it passes on its own in a fresh process as well as after the others.
"""
import numpy as np
import pytest
import torch

from imvoxelnet_b200 import _lib, scenes
from imvoxelnet_b200 import projection as pj
from tests.util import assert_volume_close

pytestmark = pytest.mark.gpu

DEV = "cuda:0"


def _outputs(b, c, n):
    vol = torch.full((b, n, c), float("nan"), device=DEV)
    valid = torch.full((b, n), 7, dtype=torch.uint8, device=DEV)
    count = torch.full((b, n), -1, dtype=torch.int32, device=DEV)
    return vol, valid, count


def _call(entry, feats, hws, proj, grid, new_origins, n, with_workspace):
    lib = _lib.lib()
    b, v, c = feats.shape[:3]
    vol, valid, count = _outputs(b, c, n)
    hw = _lib.i32([t for p in hws for t in p])
    args = [feats.data_ptr(), _lib.i64(feats.stride()), b, v, c, hw, proj.data_ptr(), None, grid,
            _lib.f32(new_origins), n, vol.data_ptr(), _lib.i64([n * c, 1, c]), _lib.IVX_F32,
            _lib.IVX_REDUCE_MEAN, valid.data_ptr(), count.data_ptr()]
    keep = None
    if with_workspace:
        nbytes = lib.ivx_backproject_nchw_workspace_bytes(b, v, c, hw)
        keep = torch.empty(nbytes, dtype=torch.uint8, device=DEV)
        args += [keep.data_ptr(), nbytes]
    rc = getattr(lib, entry)(*args, torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    return rc, vol, valid, count


def test_noalloc_matches_workspace_entry_on_two_calls_in_a_row():
    sc = scenes.scannet(n_views=3, seed=5)
    n_voxels, voxel_size = (10, 9, 6), (.5, .5, .4)
    n = int(np.prod(n_voxels))
    sc.img_meta['img_shape'] = (96, 128, 3)
    sc.img_meta['ori_shape'] = (480, 640, 3)
    P = pj.compute_projection(sc.img_meta, scenes.STRIDE)
    proj = torch.stack([P, P]).reshape(2, 3, 12).to(DEV).contiguous()
    grid = _lib.grid(n_voxels, voxel_size)
    origin = sc.img_meta['lidar2img']['origin']
    no = pj.grid_new_origin(n_voxels, voxel_size, origin).tolist()
    g = torch.Generator().manual_seed(0)
    rounds = [
        (torch.randn(2, 3, 16, 24, 32, generator=g).to(DEV), [(24, 32), (22, 30)]),
        (torch.randn(2, 3, 16, 24, 32, generator=g).to(DEV) * 3 - 1, [(20, 25), (17, 32)]),
    ]
    for feats, hws in rounds:                      # NCHW, as the reference holds its maps
        assert feats.stride(2) != 1
        rc, vol, valid, count = _call("ivx_backproject_scenes_nchw", feats, hws, proj, grid,
                                      no + no, n, True)
        assert rc == 0, _lib.lib().ivx_last_error()
        rc, vol2, valid2, count2 = _call("ivx_backproject_scenes_nchw_noalloc", feats, hws, proj,
                                         grid, no + no, n, False)
        assert rc == 0, _lib.lib().ivx_last_error()
        assert valid.max() <= 1 and valid.any()
        assert torch.equal(valid2, valid) and torch.equal(count2, count)
        assert not torch.isnan(vol2).any()
        assert_volume_close(vol2.cpu().numpy(), vol.cpu().numpy())


def test_noalloc_refuses_what_does_not_fit():
    feats = torch.zeros(1, 1, 64, 400, 400, device=DEV)          # 41 MB of maps > 32 MiB
    proj = torch.zeros(1, 1, 12, device=DEV)
    grid = _lib.grid((2, 2, 2), (1., 1., 1.))
    rc, vol, valid, count = _call("ivx_backproject_scenes_nchw_noalloc", feats, [(400, 400)], proj,
                                  grid, [0., 0., 0.], 8, False)
    assert rc == -2
    msg = _lib.lib().ivx_last_error().decode()
    assert "built-in scratch" in msg and str(_lib.IVX_NOALLOC_SCRATCH_BYTES) in msg
    assert torch.isnan(vol).all() and (valid == 7).all() and (count == -1).all()
