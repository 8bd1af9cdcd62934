"""Python face of the fused back-projection kernels.

Same names, argument meaning and returns as the reference's functions in
``mmdet3d/models/detectors/imvoxelnet.py`` (``backproject`` :145-160, the view
reduction of ``extract_feat`` :70-74), implemented by the C-ABI library
through ``_lib``.  PyTorch is used for device memory and streams only.
"""
import functools

import numpy as np
import torch

from . import _lib
from .projection import grid_new_origin

__all__ = ["backproject", "backproject_scenes", "project_indices", "finalize_mean",
           "repack_nchw_to_nhwc", "launch_count"]

# number of kernels this module has enqueued (bench.py reports it as gpu_launches)
_launches = 0


def launch_count():
    return _launches


@functools.lru_cache(maxsize=4096)
def _new_origin_cached(n_voxels, voxel_size, origin):
    return tuple(grid_new_origin(n_voxels, voxel_size, origin).tolist())


def _new_origin(n_voxels, voxel_size, origin):
    """fp32 ``origin - n_voxels / 2 * voxel_size`` (projection.grid_new_origin),
    memoised: the same few origins come back every batch."""
    return _new_origin_cached(tuple(int(v) for v in n_voxels), tuple(float(v) for v in voxel_size),
                              tuple(float(v) for v in origin))


def _stream():
    return torch.cuda.current_stream().cuda_stream


def _need_cuda(t, name):
    if not (isinstance(t, torch.Tensor) and t.is_cuda):
        raise _lib.IvxError(f"{name} must be a CUDA tensor: this path has no CPU implementation")
    return t


def _no_grad(features):
    if torch.is_grad_enabled() and features.requires_grad:
        raise NotImplementedError(
            "this entry point has no autograd support; backproject_scenes(reduce='mean' or 'partial', "
            "fp32, channels-last) is differentiable with respect to the features")


class _BackprojectScenes(torch.autograd.Function):
    """Differentiable (w.r.t. the features) fused back-projection: forward
    ``ivx_backproject_scenes``, backward ``ivx_backproject_scenes_backward``."""

    @staticmethod
    def forward(ctx, features, call):
        volume, valid, count = call(features.detach())
        ctx.call_backward = call.backward
        ctx.feat_meta = (features.shape, features.stride())
        ctx.mark_non_differentiable(valid, count)
        return volume, valid, count

    @staticmethod
    def backward(ctx, grad_volume, _gv, _gc):
        shape, stride = ctx.feat_meta
        grad_feat = torch.zeros(shape, dtype=torch.float32, device=grad_volume.device)
        if stride[2] == 1:      # same channels-last layout as the features
            grad_feat = grad_feat.permute(0, 1, 3, 4, 2).contiguous().permute(0, 1, 4, 2, 3)
        ctx.call_backward(grad_volume, grad_feat)
        return grad_feat, None


def _proj_dev(projection, device):
    p = torch.as_tensor(projection, dtype=torch.float32)
    if not p.is_cuda:
        p = p.to(device, non_blocking=True)
    return p.contiguous()


def _points_arg(points, device):
    pts = _need_cuda(points, "points") if points.is_cuda else points.to(device)
    pts = pts.to(torch.float32).reshape(3, -1).contiguous()
    return pts


def project_indices(points, projection, height, width, *, n_voxels=None, voxel_size=None,
                    origin=None, device=None):
    """``x, y, valid`` of :151-154 -> (int32 [V,N], int32 [V,N], bool [V,N]).

    Either ``points`` ([3,X,Y,Z] CUDA tensor) or the grid description
    (``n_voxels``, ``voxel_size``, ``origin``) selects the voxels."""
    global _launches
    if points is not None:
        device = points.device if points.is_cuda else device
        pts = _points_arg(points, device)
        n, g, no = pts.shape[1], None, None
        pp = pts.data_ptr()
    else:
        device = torch.device(device or "cuda")
        g = _lib.grid(n_voxels, voxel_size)
        no = _lib.f32(_new_origin(n_voxels, voxel_size, origin))
        n, pp = int(np.prod(n_voxels)), None
    proj = _proj_dev(projection, device).reshape(-1, 12)
    v = proj.shape[0]
    x = torch.empty((v, n), dtype=torch.int32, device=device)
    y = torch.empty((v, n), dtype=torch.int32, device=device)
    valid = torch.empty((v, n), dtype=torch.bool, device=device)
    with torch.cuda.device(device):
        _lib.check(_lib.lib().ivx_project_indices(
            pp, g, no, n, proj.data_ptr(), v, int(height), int(width), x.data_ptr(), y.data_ptr(),
            valid.data_ptr(), _stream()), "ivx_project_indices")
    _launches += 1
    return x, y, valid


def backproject(features, points, projection):
    """Drop-in for the reference ``backproject(features, points, projection)``.

    features [V,C,h,w] fp32 CUDA (any strides: the reference passes an NCHW
    slice, a channels-last map takes the vectorised kernel), points
    [3,X,Y,Z], projection [V,3,4].  Returns ``(volume [V,C,X,Y,Z] fp32, valid
    [V,1,X,Y,Z] bool)``.  The volume's memory format follows the features':
    channels-last features give a channels-last (NDHWC-strided) volume."""
    global _launches
    _need_cuda(features, "features")
    _no_grad(features)
    if features.dtype != torch.float32:
        raise TypeError("features must be float32")
    v, c, h, w = features.shape
    xs, ys, zs = points.shape[-3:]
    pts = _points_arg(points, features.device)
    n = pts.shape[1]
    proj = _proj_dev(projection, features.device).reshape(-1, 12)
    if proj.shape[0] != v:
        raise ValueError("projection and features disagree on the number of views")
    if features.stride(1) == 1 and c % 4 == 0:
        volume = torch.empty((v, n, c), dtype=torch.float32, device=features.device).permute(0, 2, 1)
    else:
        volume = torch.empty((v, c, n), dtype=torch.float32, device=features.device)
    valid = torch.empty((v, n), dtype=torch.bool, device=features.device)
    with torch.cuda.device(features.device):
        _lib.check(_lib.lib().ivx_backproject_views(
            features.data_ptr(), _lib.i64(features.stride()), v, c, h, w, proj.data_ptr(),
            pts.data_ptr(), None, None, n, volume.data_ptr(), _lib.i64(volume.stride()),
            valid.data_ptr(), _stream()), "ivx_backproject_views")
    _launches += (v + 63) // 64
    return volume.view(v, c, xs, ys, zs) if volume.is_contiguous() else \
        volume.unflatten(2, (xs, ys, zs)), valid.view(v, 1, xs, ys, zs)


def backproject_scenes(features, hw, projection, n_voxels, voxel_size, origins, *, points=None,
                       out_dtype=torch.float32, channels_last=True, reduce="mean", repack=True, out=None):
    """Fused back-projection + view reduction of a batch of scenes.

    Replaces the per-scene loop of ``extract_feat`` (:57-78): returns the
    stacked ``(volume [B,C,X,Y,Z], valid [B,1,X,Y,Z] bool, count [B,X,Y,Z]
    int32)``.

    features    [B,V,C,Hf,Wf] fp32 CUDA, any strides (channels-last is fast), or
                channels-last bf16 (half the gather traffic; summed in fp32)
    hw          B pairs ``img_shape // stride`` (:67-68)
    projection  [B,V,3,4] (CPU or CUDA)
    origins     B grid origins (``img_meta['lidar2img']['origin']``); ignored
                if ``points`` ([3,X,Y,Z]) is given
    reduce      "mean" (the reference's result) or "partial" (fp32 sums and
                counts for the view-sharded multi-GPU path)
    out         optional pre-allocated ``(volume [B,N,C] dense, valid [B,N] bool, count [B,N]
                int32)`` to write into (e.g. peer-addressable symmetric memory)
    repack      NCHW features with C % 4 == 0 are first re-laid to channels-last
                in a scratch buffer (``ivx_backproject_scenes_nchw``); False
                gathers from the strided maps directly
    """
    global _launches
    _need_cuda(features, "features")
    if torch.is_grad_enabled() and features.requires_grad:
        # training: the same kernels under autograd (gradient w.r.t. the features only, as upstream:
        # get_points is @no_grad and the projection comes from the metadata)
        if out is not None or out_dtype != torch.float32 or not channels_last or features.shape[2] % 4:
            raise NotImplementedError("autograd needs fp32, channels-last volumes and C % 4 == 0")
        kw = dict(points=points, out_dtype=out_dtype, channels_last=True, reduce=reduce, repack=repack)
        feats_cl = features if features.stride(2) == 1 else \
            features.permute(0, 1, 3, 4, 2).contiguous().permute(0, 1, 4, 2, 3)

        def call(f):
            with torch.no_grad():
                return backproject_scenes(f, hw, projection, n_voxels, voxel_size, origins, **kw)

        def call_backward(grad_volume, grad_feat):
            _backproject_scenes_backward(grad_volume, grad_feat, hw, projection, n_voxels, voxel_size,
                                         origins, points, reduce)

        call.backward = call_backward
        return _BackprojectScenes.apply(feats_cl, call)
    if features.dtype not in (torch.float32, torch.bfloat16) or features.dim() != 5:
        raise TypeError("features must be a float32 (or channels-last bfloat16) [B,V,C,Hf,Wf] tensor")
    bf16_maps = features.dtype == torch.bfloat16
    if bf16_maps and (features.stride(2) != 1 or features.shape[2] % 8):
        raise TypeError("bfloat16 features must be channels-last with C % 8 == 0")
    b, v, c = features.shape[:3]
    dev = features.device
    xs, ys, zs = (int(t) for t in n_voxels)
    n = xs * ys * zs
    proj = _proj_dev(projection, dev).reshape(b, v, 12)
    hw_arr = _hw_arg(hw, b, features.shape[3], features.shape[4])
    if points is not None:
        pts = _points_arg(points, dev)
        pp, g, no = pts.data_ptr(), None, None
    else:
        pp, g = None, _lib.grid(n_voxels, voxel_size)
        flat = []
        for o in origins:
            flat += _new_origin(n_voxels, voxel_size, o)
        if len(flat) != 3 * b:
            raise ValueError("origins must hold one origin per scene")
        no = _lib.f32(flat)
    mode = {"mean": _lib.IVX_REDUCE_MEAN, "partial": _lib.IVX_REDUCE_PARTIAL}[reduce]
    dt = {torch.float32: _lib.IVX_F32, torch.bfloat16: _lib.IVX_BF16}[out_dtype]
    if out is not None:
        vol_nc, valid, count = out
        if tuple(vol_nc.shape) != (b, n, c) or not vol_nc.is_contiguous() or vol_nc.dtype != out_dtype:
            raise ValueError("out[0] must be a dense [B, N, C] tensor of the output dtype")
        volume = vol_nc.permute(0, 2, 1)
    elif channels_last:
        volume = torch.empty((b, n, c), dtype=out_dtype, device=dev).permute(0, 2, 1)
    else:
        volume = torch.empty((b, c, n), dtype=out_dtype, device=dev)
    if out is None:
        valid = torch.empty((b, n), dtype=torch.bool, device=dev)
        count = torch.empty((b, n), dtype=torch.int32, device=dev)
    args = (b, v, c, hw_arr, proj.data_ptr(), pp, g, no, n, volume.data_ptr(),
            _lib.i64(volume.stride()), dt, mode, valid.data_ptr(), count.data_ptr())
    with torch.cuda.device(dev):
        if bf16_maps:
            _lib.check(_lib.lib().ivx_backproject_scenes_bf16(
                features.data_ptr(), _lib.i64(features.stride()), *args, _stream()),
                "ivx_backproject_scenes_bf16")
        elif features.stride(2) != 1 and c % 4 == 0 and repack:
            # the reference's NCHW maps: re-lay them in scratch, then the vectorised kernel
            nbytes = _lib.lib().ivx_backproject_nchw_workspace_bytes(b, v, c, hw_arr)
            workspace = torch.empty(nbytes, dtype=torch.uint8, device=dev)
            _lib.check(_lib.lib().ivx_backproject_scenes_nchw(
                features.data_ptr(), _lib.i64(features.stride()), *args, workspace.data_ptr(),
                nbytes, _stream()), "ivx_backproject_scenes_nchw")
            _launches += 1
        else:
            _lib.check(_lib.lib().ivx_backproject_scenes(
                features.data_ptr(), _lib.i64(features.stride()), *args, _stream()),
                "ivx_backproject_scenes")
    _launches += (b + 63) // 64
    return (volume.unflatten(2, (xs, ys, zs)), valid.view(b, 1, xs, ys, zs),
            count.view(b, xs, ys, zs))


def _hw_arg(hw, n_scenes, map_h, map_w):
    """Host int32 [B][2] array of the per-scene (height, width).  The kernels only get strides, so the
    extents are checked here against the real feature map: the reference's slice
    ``feature[:, :, :height, :width]`` (detectors/imvoxelnet.py:69) clamps to the tensor, an
    unchecked hw beyond it would read (forward) or atomically add (backward) out of bounds."""
    flat = [int(t) for pair in hw for t in pair]
    if len(flat) != 2 * n_scenes:
        raise ValueError("hw must hold one (height, width) pair per scene")
    for i in range(n_scenes):
        h, w = flat[2 * i], flat[2 * i + 1]
        if not (0 < h <= map_h and 0 < w <= map_w):
            raise ValueError(f"scene {i}: hw = ({h}, {w}) outside the feature map ({map_h}, {map_w})")
    return _lib.i32(flat)


def _backproject_scenes_backward(grad_volume, grad_feat, hw, projection, n_voxels, voxel_size, origins,
                                 points, reduce):
    """grad_feat (zero-filled, channels-last [B,V,C,Hf,Wf]) += scatter of grad_volume [B,C,X,Y,Z]."""
    global _launches
    b, v, c = grad_feat.shape[:3]
    dev = grad_feat.device
    xs, ys, zs = (int(t) for t in n_voxels)
    n = xs * ys * zs
    gv = grad_volume.reshape(b, c, n)
    if gv.stride(1) != 1 or gv.stride(2) != c or gv.dtype != torch.float32:
        gv = gv.float().permute(0, 2, 1).contiguous().permute(0, 2, 1)      # NDHWC, dense
    proj = _proj_dev(projection, dev).reshape(b, v, 12)
    hw_arr = _hw_arg(hw, b, grad_feat.shape[3], grad_feat.shape[4])
    if points is not None:
        pts = _points_arg(points, dev)
        pp, g, no = pts.data_ptr(), None, None
    else:
        pp, g = None, _lib.grid(n_voxels, voxel_size)
        flat = []
        for o in origins:
            flat += _new_origin(n_voxels, voxel_size, o)
        no = _lib.f32(flat)
    mode = {"mean": _lib.IVX_REDUCE_MEAN, "partial": _lib.IVX_REDUCE_PARTIAL}[reduce]
    with torch.cuda.device(dev):
        _lib.check(_lib.lib().ivx_backproject_scenes_backward(
            gv.data_ptr(), _lib.i64(gv.stride()), grad_feat.data_ptr(), _lib.i64(grad_feat.stride()), b, v, c,
            hw_arr, proj.data_ptr(), pp, g, no, n, mode, _stream()), "ivx_backproject_scenes_backward")
    _launches += (b + 63) // 64


def finalize_mean(partial_sum, count, *, out_dtype=torch.float32, channels_last=True):
    """``sum / count`` with unseen voxels zeroed (:72-74) for all-reduced
    partial results.  partial_sum: [B,C,X,Y,Z] fp32 with channels-last strides
    (as ``backproject_scenes(reduce="partial")`` returns it), count [B,X,Y,Z]."""
    global _launches
    _need_cuda(partial_sum, "partial_sum")
    b, c = partial_sum.shape[:2]
    shp = tuple(partial_sum.shape[2:])
    n = int(np.prod(shp))
    s = partial_sum.reshape(b, c, n)
    if s.stride(1) != 1 or s.stride(2) != c:
        raise ValueError("partial_sum must be channels-last (NDHWC-strided)")
    cnt = count.reshape(b, n).contiguous()
    dev = partial_sum.device
    if channels_last:
        volume = torch.empty((b, n, c), dtype=out_dtype, device=dev).permute(0, 2, 1)
    else:
        volume = torch.empty((b, c, n), dtype=out_dtype, device=dev)
    valid = torch.empty((b, n), dtype=torch.bool, device=dev)
    dt = {torch.float32: _lib.IVX_F32, torch.bfloat16: _lib.IVX_BF16}[out_dtype]
    with torch.cuda.device(dev):
        _lib.check(_lib.lib().ivx_finalize_mean(
            s.data_ptr(), cnt.data_ptr(), b, c, n, volume.data_ptr(), _lib.i64(volume.stride()), dt,
            valid.data_ptr(), _stream()), "ivx_finalize_mean")
    _launches += 1
    return volume.unflatten(2, shp), valid.view((b, 1) + shp)


def repack_nchw_to_nhwc(features):
    """[..., C, H, W] fp32 CUDA tensor of any strides -> same logical tensor in
    channels-last memory (one tiled transpose kernel)."""
    global _launches
    _need_cuda(features, "features")
    lead = features.shape[:-3]
    c, h, w = features.shape[-3:]
    f4 = features.reshape((-1, c, h, w)) if features.dim() != 4 else features
    out = torch.empty((f4.shape[0], h, w, c), dtype=torch.float32, device=features.device)
    with torch.cuda.device(features.device):
        _lib.check(_lib.lib().ivx_repack_nchw_to_nhwc(
            f4.data_ptr(), _lib.i64(f4.stride()), f4.shape[0], c, h, w, out.data_ptr(), _stream()),
            "ivx_repack_nchw_to_nhwc")
    _launches += 1
    out = out.permute(0, 3, 1, 2)
    return out.unflatten(0, lead) if len(lead) != 1 else out
