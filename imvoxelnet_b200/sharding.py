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
* both     -- ``view_groups(ways)``: the views are sharded inside groups of
  ``ways`` ranks, the groups hold different scenes; the exchange shrinks with
  ``ways`` (8 GPUs as 4 pairs: 7.4x of one GPU on 50-view scenes).
"""
import torch
import torch.distributed as dist

__all__ = ["shard_views", "shard_scenes", "view_groups", "allreduce_partials", "backproject_scenes_view_sharded",
           "FusedViewSharded"]


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


def view_groups(ways, world_size=None, rank=None):
    """Hybrid of the two axes: the ranks form ``world_size // ways`` groups of ``ways`` consecutive ranks;
    the views of a scene are sharded over the ranks of ONE group (pass the returned process group to
    ``backproject_scenes_view_sharded`` / ``FusedViewSharded``), the groups work on different scenes.  The
    exchange per GPU is then ``(ways - 1) / ways`` of the partial volumes instead of
    ``(world_size - 1) / world_size``, which is what lets 50-view scenes scale.
    Collective: every rank must call it.  Returns ``(group, group_index, rank_in_group)``."""
    world_size = dist.get_world_size() if world_size is None else world_size
    rank = dist.get_rank() if rank is None else rank
    if ways < 1 or world_size % ways:
        raise ValueError(f"{ways}-way view sharding does not divide a world of {world_size}")
    groups = [dist.new_group(list(range(g * ways, (g + 1) * ways))) for g in range(world_size // ways)]
    return groups[rank // ways], rank // ways, rank % ways


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


class FusedViewSharded:
    """View-sharded back-projection whose reduction is fused into the finalize kernel.

    The partial sums and counts of every rank live in torch symmetric memory
    (same offset on every GPU of the NVLink domain).  A step is: the fused
    back-projection kernel writes this rank's partials there; a barrier over
    the ranks; ``ivx_finalize_mean_peers`` reads all ranks' partials in place --
    through the NVSwitch multicast address with ``multimem.ld_reduce`` when the
    allocation has one (the switch does the sum), else peer by peer over NVLink
    -- and divides; a second barrier frees the buffers for the next step.  No
    NCCL call, no reduced copy of the partials in HBM.
    """

    def __init__(self, n_scenes, n_channels, n_voxels, group=None, device=None, use_multicast=True):
        import torch.distributed._symmetric_memory as symm_mem
        self.group = group if group is not None else dist.group.WORLD
        self.device = torch.device(device if device is not None else torch.cuda.current_device())
        self.shape = (int(n_scenes), int(n_channels), tuple(int(v) for v in n_voxels))
        n = self.shape[2][0] * self.shape[2][1] * self.shape[2][2]
        self.n_points = n
        self.sum = symm_mem.empty((n_scenes, n, n_channels), dtype=torch.float32, device=self.device)
        self.count = symm_mem.empty((n_scenes, n), dtype=torch.int32, device=self.device)
        self._hs = symm_mem.rendezvous(self.sum, self.group)
        self._hc = symm_mem.rendezvous(self.count, self.group)
        self.world = self._hs.world_size
        self.rank = self._hs.rank
        self.multicast = bool(use_multicast and self._hs.multicast_ptr and self._hc.multicast_ptr)
        self._valid = torch.empty((n_scenes, n), dtype=torch.bool, device=self.device)

    def __call__(self, features, hw, projection, voxel_size, origins, *, out_dtype=torch.float32,
                 scatter_scenes=False):
        """``features`` / ``projection`` hold this rank's views (``shard_views``); a rank
        without views passes ``V_local == 0`` tensors.  Returns ``(volume, valid, count)``:
        of all scenes, identical on every rank (all-reduce semantics), or with
        ``scatter_scenes`` only of this rank's ``shard_scenes`` slice (reduce-scatter
        semantics: every rank then runs the neck on its own scenes, none is replicated)."""
        import ctypes

        from . import _lib, ops
        b, c, nv = self.shape
        n = self.n_points
        if features.shape[1] > 0:
            ops.backproject_scenes(features, hw, projection, nv, voxel_size, origins, reduce="partial",
                                   out=(self.sum, self._valid, self.count))
        else:
            self.sum.zero_()
            self.count.zero_()
        self._hs.barrier(channel=0)          # every rank's partials are written and visible
        mine = shard_scenes(b, self.rank, self.world) if scatter_scenes else slice(0, b)
        b0, b = mine.start, mine.stop - mine.start
        volume = torch.empty((b, n, c), dtype=out_dtype, device=self.device).permute(0, 2, 1)
        valid = torch.empty((b, n), dtype=torch.bool, device=self.device)
        count = torch.empty((b, n), dtype=torch.int32, device=self.device)
        off_s, off_c = b0 * n * c * 4, b0 * n * 4      # byte offsets of the first scene of the slice
        if b == 0:
            self._hs.barrier(channel=1)
            return (volume.unflatten(2, nv), valid.view((0, 1) + nv), count.view((0,) + nv))
        if self.multicast:
            sums = counts = None
            mc_sum, mc_count = self._hs.multicast_ptr + off_s, self._hc.multicast_ptr + off_c
        else:
            sums = (ctypes.c_void_p * self.world)(*[int(p) + off_s for p in self._hs.buffer_ptrs])
            counts = (ctypes.c_void_p * self.world)(*[int(p) + off_c for p in self._hc.buffer_ptrs])
            mc_sum = mc_count = None
        dt = {torch.float32: _lib.IVX_F32, torch.bfloat16: _lib.IVX_BF16}[out_dtype]
        with torch.cuda.device(self.device):
            _lib.check(_lib.lib().ivx_finalize_mean_peers(
                sums, counts, self.world, mc_sum, mc_count, b, c, n, volume.data_ptr(),
                _lib.i64(volume.stride()), dt, valid.data_ptr(), count.data_ptr(),
                torch.cuda.current_stream().cuda_stream), "ivx_finalize_mean_peers")
        self._hs.barrier(channel=1)          # nobody overwrites partials that a peer still reads
        return (volume.unflatten(2, nv), valid.view((b, 1) + nv), count.view((b,) + nv))
