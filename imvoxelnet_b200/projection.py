"""Host-side set-up of the per-view projection matrices and the voxel grid.

Mirrors, by name and argument meaning, the helpers the reference keeps in
``mmdet3d/models/detectors/imvoxelnet.py``:

* ``compute_projection``  <- ``ImVoxelNet._compute_projection`` (:114-129)
* ``get_extrinsics``      <- ``get_extrinsics``                 (:164-187)
* ``get_points``          <- ``get_points``                     (:132-141)
* ``grid_new_origin``     <- the ``new_origin`` line            (:139)

These are a few dozen flops per scene; the host versions use the same fp32
torch operations, in the same order, as the reference, because they *define
the bits* of ``P`` and of the voxel coordinates that the CUDA kernel then has
to reproduce.  ``compute_projection_batch`` is the device version for a whole
batch (one packed host->device copy and one launch of
``ivx_compute_projection``, bit-identical to the host loop).
"""
import numpy as np
import torch

__all__ = ["compute_projection", "compute_projection_batch", "get_extrinsics", "get_points", "grid_new_origin"]


def get_extrinsics(angles):
    """Camera extrinsic [4, 4] from predicted ``(pitch, roll)`` (yaw = 0).

    Used only by the SUN RGB-D "Total" models at test time
    (reference :164-187).  The rotation is the Total3DUnderstanding one,
    re-expressed in the depth-camera axis order of ``DepthInstance3DBoxes``.
    """
    pitch, roll = angles
    yaw = angles.new_zeros(())
    cy, sy = torch.cos(yaw), torch.sin(yaw)
    cp, sp = torch.cos(pitch), torch.sin(pitch)
    cr, sr = torch.cos(roll), torch.sin(roll)
    rot = torch.stack([
        torch.stack([cy * cp, sy * sr - cy * cr * sp, cr * sy + cy * sp * sr]),
        torch.stack([sp, cp * cr, -cp * sr]),
        torch.stack([-cp * sy, cy * sr + cr * sy * sp, cy * cr - sy * sp * sr]),
    ])
    # world axes of Total3D -> (x right, y forward, z up) of the depth frame
    axes = angles.new_tensor([[0., 0., 1.], [0., -1., 0.], [-1., 0., 0.]])
    rot = (axes @ rot.T)[:, [2, 0, 1]]
    rot[2] *= -1
    out = angles.new_zeros((4, 4))
    out[:3, :3] = rot
    out[3, 3] = 1.
    return out


def compute_projection(img_meta, stride, angles=None):
    """``[V, 3, 4]`` projection matrices of one scene (reference :114-129).

    ``img_meta['lidar2img']`` holds ``intrinsic`` ([3,3]/[3,4]/[4,4] array),
    ``extrinsic`` (list of [4,4] world->camera arrays, one per view) and
    ``origin``.  The intrinsic's first two rows are scaled from the original
    image to the stride-``stride`` feature map; ``angles`` (SUN RGB-D Total,
    test mode) replaces the metadata extrinsics by predicted pitch/roll.
    """
    lidar2img = img_meta['lidar2img']
    intrinsic = torch.tensor(lidar2img['intrinsic'][:3, :3])
    ratio = img_meta['ori_shape'][0] / (img_meta['img_shape'][0] / stride)
    intrinsic[:2] /= ratio
    if angles is not None:
        extrinsics = [get_extrinsics(a).to(intrinsic.device) for a in angles]
    else:
        extrinsics = [torch.tensor(e) for e in lidar2img['extrinsic']]
    return torch.stack([intrinsic @ e[:3] for e in extrinsics])


def compute_projection_batch(img_metas, stride, device):
    """``[B, V, 3, 4]`` CUDA tensor of every scene's matrices, metadata extrinsics only.

    Equals ``torch.stack([compute_projection(m, stride) for m in img_metas]).to(device)`` bit for
    bit (reference :57-61 with :114-129), without the per-scene CPU tensor ops: the raw
    intrinsics, ratios and extrinsics travel in one pinned buffer and the products are formed on
    the device.  All scenes must have the same number of views."""
    from . import _lib
    device = torch.device(device)
    if device.type != "cuda":
        raise _lib.IvxError("compute_projection_batch needs a CUDA device: use compute_projection on the host")
    b = len(img_metas)
    n_views = len(img_metas[0]['lidar2img']['extrinsic'])
    packed = np.empty(b * 10 + b * n_views * 16, dtype=np.float32)
    k, r, e = packed[:b * 9].reshape(b, 9), packed[b * 9:b * 10], packed[b * 10:].reshape(b, n_views, 16)
    for i, meta in enumerate(img_metas):
        lidar2img = meta['lidar2img']
        if len(lidar2img['extrinsic']) != n_views:
            raise ValueError("scenes of one batch must have the same number of views")
        k[i] = np.asarray(lidar2img['intrinsic'], dtype=np.float32)[:3, :3].reshape(9)
        r[i] = meta['ori_shape'][0] / (meta['img_shape'][0] / stride)     # double, then one rounding to fp32
        for v, ext in enumerate(lidar2img['extrinsic']):
            e[i, v] = np.asarray(ext, dtype=np.float32).reshape(16)
    host = torch.from_numpy(packed).pin_memory()     # torch's pinned and device caches are stream-ordered
    with torch.cuda.device(device):
        dev = host.to(device, non_blocking=True)
        proj = torch.empty((b, n_views, 3, 4), dtype=torch.float32, device=device)
        _lib.check(_lib.lib().ivx_compute_projection(
            dev.data_ptr(), dev.data_ptr() + 4 * b * 9, dev.data_ptr() + 4 * b * 10, b, n_views, proj.data_ptr(),
            torch.cuda.current_stream().cuda_stream), "ivx_compute_projection")
    return proj


def grid_new_origin(n_voxels, voxel_size, origin):
    """Lower corner of the grid, ``origin - n_voxels / 2. * voxel_size``, fp32
    (reference :139).  Voxel ``(i, j, k)`` sits at ``idx * voxel_size +
    new_origin`` -- the lower corner, not the centre, exactly as upstream."""
    n_voxels = torch.as_tensor(n_voxels)
    voxel_size = torch.as_tensor(voxel_size, dtype=torch.float32)
    origin = torch.as_tensor(origin, dtype=torch.float32)
    return origin - n_voxels / 2. * voxel_size


@torch.no_grad()
def get_points(n_voxels, voxel_size, origin):
    """Voxel coordinates ``[3, X, Y, Z]`` fp32, Z fastest (reference :132-141)."""
    n_voxels = torch.as_tensor(n_voxels)
    voxel_size = torch.as_tensor(voxel_size, dtype=torch.float32)
    new_origin = grid_new_origin(n_voxels, voxel_size, origin)
    idx = torch.stack(torch.meshgrid(
        [torch.arange(int(n)) for n in n_voxels], indexing='ij'))
    return idx * voxel_size.view(3, 1, 1, 1) + new_origin.view(3, 1, 1, 1)
