"""Multi-GPU decomposition of the image->voxel path.

The reference has no such thing: it loops over the views of a scene on one
device (mmdet3d/models/detectors/imvoxelnet.py:156) and is data-parallel over
scenes only (tools/dist_train.sh).  Two independent axes are offered here,
one process per GPU, ``torch.distributed`` for the plumbing:

* scenes   -- each rank takes a slice of the batch; no collective at all.
* views    -- the per-view contributions to ``sum[C,N]`` and ``count[N]`` are
  independent (:156-157, :70-71), so rank r back-projects views r, r+G, ...
  with ``reduce="partial"``, one all-reduce adds the fp32 sums and the int32
  counts (NCCL over NVLink on the GPUs, gloo in the CPU tests), and
  ``finalize_mean`` divides.  Only the summation order changes (<= 1e-5).
"""
import torch
import torch.distributed as dist

__all__ = ["shard_views", "shard_scenes", "allreduce_partials", "backproject_scenes_view_sharded"]


def shard_views(n_views, rank, world_size):
    """Indices of the views rank ``rank`` owns: ``rank, rank + G, ...``.
    Interleaving keeps neighbouring cameras (similar hit counts) on different
    ranks; with fewer views than ranks the tail ranks own nothing."""
    if not 0 <= rank < world_size:
        raise ValueError(f"rank {rank} outside world of {world_size}")
    return list(range(rank, n_views, world_size))


def shard_scenes(n_scenes, rank, world_size):
    """Contiguous, balanced slice of the batch for this rank (sizes differ by
    at most one)."""
    if not 0 <= rank < world_size:
        raise ValueError(f"rank {rank} outside world of {world_size}")
    base, extra = divmod(n_scenes, world_size)
    start = rank * base + min(rank, extra)
    return slice(start, start + base + (1 if rank < extra else 0))


def allreduce_partials(partial_sum, count, group=None):
    """In-place sum over ranks of the fp32 partial sums and the int32 view
    counts.  Two collectives on the caller's stream; works for any backend
    (nccl on GPUs; gloo on CPU tensors in the tests)."""
    if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.all_reduce(partial_sum, op=dist.ReduceOp.SUM, group=group)
        dist.all_reduce(count, op=dist.ReduceOp.SUM, group=group)
    return partial_sum, count


def backproject_scenes_view_sharded(features, hw, projection, n_voxels, voxel_size, origins, *,
                                    group=None, out_dtype=torch.float32, channels_last=True):
    """View-sharded ``ops.backproject_scenes``.

    ``features`` [B, V_local, C, Hf, Wf] and ``projection`` [B, V_local, 3, 4]
    hold only this rank's views (``shard_views``).  A rank that owns no view
    passes tensors with ``V_local == 0`` and contributes zeros.  Returns the
    same ``(volume, valid, count)`` on every rank."""
    from . import ops
    b, v_local, c = features.shape[:3]
    xs, ys, zs = (int(t) for t in n_voxels)
    if v_local > 0:
        s, _, cnt = ops.backproject_scenes(features, hw, projection, n_voxels, voxel_size, origins,
                                           reduce="partial")
    else:
        n = xs * ys * zs
        s = torch.zeros((b, n, c), device=features.device).permute(0, 2, 1).unflatten(2, (xs, ys, zs))
        cnt = torch.zeros((b, xs, ys, zs), dtype=torch.int32, device=features.device)
    allreduce_partials(s, cnt, group)
    volume, valid = ops.finalize_mean(s, cnt, out_dtype=out_dtype, channels_last=channels_last)
    return volume, valid, cnt
