"""Device-side projection set-up (ivx_compute_projection) is bit-identical to the reference's host loop.

Anchors: tests/golden/host_setup.npz (projection matrices produced by the reference's own
``ImVoxelNet._compute_projection``) and the host mirror ``projection.compute_projection`` (pinned to the
same fixture in tests/test_oracle_golden.py) on every synthetic scene family and on random metadata."""
import numpy as np
import pytest
import torch

from imvoxelnet_b200 import projection, scenes
from tests.util import load_golden

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _bits(t):
    return t.detach().cpu().contiguous().view(torch.int32)


@pytest.mark.parametrize("name", sorted(scenes.BUILDERS))
def test_batched_projection_matches_reference_golden(name):
    """P_<scene> of host_setup.npz was produced by the reference's own ImVoxelNet._compute_projection."""
    g = load_golden("host_setup")
    meta = scenes.BUILDERS[name]().img_meta
    got = projection.compute_projection_batch([meta, meta, meta], 4, DEV)
    assert got.dtype == torch.float32 and tuple(got.shape[1:]) == g["P_" + name].shape
    for i in range(3):
        assert np.array_equal(got[i].cpu().numpy().view(np.uint32), g["P_" + name].view(np.uint32))
    want = projection.compute_projection(meta, 4)
    assert torch.equal(_bits(got[0]), _bits(want))


def test_batched_projection_on_random_metadata_is_bit_exact():
    rng = np.random.default_rng(3)
    for n_views, kshape in ((1, (3, 3)), (6, (4, 4)), (50, (3, 4))):
        metas = []
        for _ in range(7):
            k = np.zeros(kshape, np.float32)
            k[:3, :3] = rng.standard_normal((3, 3)) * 300
            ext = [rng.standard_normal((4, 4)).astype(np.float32) for _ in range(n_views)]
            h = int(rng.integers(200, 1000))
            metas.append(dict(img_shape=(h, 640, 3), ori_shape=(int(rng.integers(200, 1400)), 1242, 3),
                              lidar2img=dict(intrinsic=k, extrinsic=ext, origin=[0., 0., 0.])))
        want = torch.stack([projection.compute_projection(m, 4) for m in metas])
        got = projection.compute_projection_batch(metas, 4, DEV)
        assert torch.equal(_bits(got), _bits(want))


def test_batched_projection_needs_cuda_and_equal_view_counts():
    from imvoxelnet_b200 import _lib
    scene = scenes.BUILDERS["kitti"]()
    with pytest.raises(_lib.IvxError):
        projection.compute_projection_batch([scene.img_meta], 4, "cpu")
    other = scenes.BUILDERS["nuscenes"]()
    with pytest.raises(ValueError):
        projection.compute_projection_batch([scene.img_meta, other.img_meta], 4, DEV)
