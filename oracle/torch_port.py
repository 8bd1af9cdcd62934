"""PyTorch-CPU port of the reference's image->voxel path, op for op.

TEST / BASELINE INFRASTRUCTURE ONLY (see oracle/__init__.py).  The reference
cannot travel to the GPU box (it needs /root/reference plus mmcv/mmdet), so
``bench.py --impl reference`` and the ``cpu_baseline`` leg time this port on
the box's host cores instead: it issues the same ATen operations, with the
same temporaries (the homogeneous [V,4,N] points, the zero-filled [V,C,N]
volume, the per-view boolean-mask gather/scatter, the re-read for the sum),
as mmdet3d/models/detectors/imvoxelnet.py:145-160 and :70-74, so its cost is
the reference's cost.  tests/test_oracle_golden.py pins it to the goldens.
"""
import torch


def pixel_indices(points, projection, height, width):
    """:148-154 -- homogeneous projection, rounded pixel, validity."""
    n_views = projection.shape[0]
    flat = points.reshape(1, 3, -1).expand(n_views, 3, -1)
    homogeneous = torch.cat([flat, torch.ones_like(flat[:, :1])], dim=1)
    cam = torch.bmm(projection, homogeneous)
    depth = cam[:, 2]
    col = (cam[:, 0] / depth).round().long()
    row = (cam[:, 1] / depth).round().long()
    inside = (col >= 0) & (row >= 0) & (col < width) & (row < height) & (depth > 0)
    return col, row, inside


def backproject(features, points, projection):
    """:145-160 -- per-view volumes [V,C,X,Y,Z] and valid [V,1,X,Y,Z]."""
    n_views, n_channels, height, width = features.shape
    grid_shape = tuple(points.shape[-3:])
    col, row, inside = pixel_indices(points, projection, height, width)
    volume = torch.zeros((n_views, n_channels, col.shape[-1]), device=features.device)
    for view in range(n_views):
        hit = inside[view]
        volume[view, :, hit] = features[view, :, row[view, hit], col[view, hit]]
    return (volume.view((n_views, n_channels) + grid_shape),
            inside.view((n_views, 1) + grid_shape))


def scene_volume(features, points, projection):
    """:69-74 -- mean over valid views, zero where no view sees the voxel."""
    volume, seen = backproject(features, points, projection)
    total = volume.sum(dim=0)
    hits = seen.sum(dim=0)
    mean = total / hits
    seen_any = hits > 0
    mean[:, ~seen_any[0]] = .0
    return mean, seen_any
