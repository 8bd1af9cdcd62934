"""ctypes binding of libivx.so (the C ABI declared in include/ivx.h).

There is no fallback: if the CUDA library has not been built, importing a
symbol raises, and every op refuses CPU tensors.
"""
import ctypes
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libivx.so")

IVX_F32, IVX_BF16, IVX_U8, IVX_F32X2 = 0, 1, 2, 3
IVX_REDUCE_MEAN, IVX_REDUCE_PARTIAL = 0, 1
IVX_MAX_VIEWS = 1024

c_void_p, c_int, c_int64 = ctypes.c_void_p, ctypes.c_int, ctypes.c_int64
_i64p = ctypes.POINTER(ctypes.c_int64)
_i32p = ctypes.POINTER(ctypes.c_int32)
_f32p = ctypes.POINTER(ctypes.c_float)


class Grid(ctypes.Structure):
    """``ivx_grid`` of include/ivx.h."""
    _fields_ = [("n_voxels", ctypes.c_int32 * 3), ("voxel_size", ctypes.c_float * 3)]


_GridP = ctypes.POINTER(Grid)


class ConvParams(ctypes.Structure):
    """``ivx_conv3d_params`` of include/ivx.h."""
    _fields_ = [("batch", ctypes.c_int32), ("c_in", ctypes.c_int32), ("c_out", ctypes.c_int32),
                ("in_dims", ctypes.c_int32 * 3), ("out_dims", ctypes.c_int32 * 3),
                ("kernel", ctypes.c_int32 * 3), ("stride", ctypes.c_int32 * 3),
                ("pad", ctypes.c_int32 * 3), ("transposed", ctypes.c_int32),
                ("relu", ctypes.c_int32), ("residual_after_act", ctypes.c_int32),
                ("post_scale", ctypes.c_float), ("in_dtype", ctypes.c_int32),
                ("out_dtype", ctypes.c_int32)]


_ConvP = ctypes.POINTER(ConvParams)

# name -> argtypes, exactly the prototypes of include/ivx.h
PROTOTYPES = {
    "ivx_project_indices": [c_void_p, _GridP, _f32p, c_int64, c_void_p, c_int, c_int, c_int,
                            c_void_p, c_void_p, c_void_p, c_void_p],
    "ivx_backproject_scenes": [c_void_p, _i64p, c_int, c_int, c_int, _i32p, c_void_p, c_void_p,
                               _GridP, _f32p, c_int64, c_void_p, _i64p, c_int, c_int, c_void_p,
                               c_void_p, c_void_p],
    "ivx_backproject_views": [c_void_p, _i64p, c_int, c_int, c_int, c_int, c_void_p, c_void_p,
                              _GridP, _f32p, c_int64, c_void_p, _i64p, c_void_p, c_void_p],
    "ivx_finalize_mean": [c_void_p, c_void_p, c_int, c_int, c_int64, c_void_p, _i64p, c_int,
                          c_void_p, c_void_p],
    "ivx_repack_nchw_to_nhwc": [c_void_p, _i64p, c_int, c_int, c_int, c_int, c_void_p, c_void_p],
    "ivx_backproject_nchw_workspace_bytes": [c_int, c_int, c_int, _i32p],
    "ivx_backproject_scenes_nchw": [c_void_p, _i64p, c_int, c_int, c_int, _i32p, c_void_p, c_void_p,
                                    _GridP, _f32p, c_int64, c_void_p, _i64p, c_int, c_int, c_void_p,
                                    c_void_p, c_void_p, c_int64, c_void_p],
}
PROTOTYPES["ivx_backproject_scenes_nchw_noalloc"] = PROTOTYPES["ivx_backproject_scenes_nchw"][:17] + [c_void_p]
IVX_NOALLOC_SCRATCH_BYTES = 32 * 1024 * 1024
PROTOTYPES["ivx_backproject_scenes_bf16"] = list(PROTOTYPES["ivx_backproject_scenes"])
PROTOTYPES["ivx_backproject_scenes_backward"] = [c_void_p, _i64p, c_void_p, _i64p, c_int, c_int, c_int, _i32p,
                                                 c_void_p, c_void_p, _GridP, _f32p, c_int64, c_int, c_void_p]
PROTOTYPES["ivx_finalize_mean_peers"] = [c_void_p, c_void_p, c_int, c_void_p, c_void_p, c_int, c_int, c_int64,
                                         c_void_p, _i64p, c_int, c_void_p, c_void_p, c_void_p]
PROTOTYPES["ivx_conv3d_ndhwc"] = [_ConvP, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p,
                                  c_void_p]
PROTOTYPES["ivx_conv3d_ndhwc_bf16_tc"] = PROTOTYPES["ivx_conv3d_ndhwc"][:7] + [c_void_p, c_int64, c_void_p]
PROTOTYPES["ivx_conv3d_ndhwc_tf32_tc"] = PROTOTYPES["ivx_conv3d_ndhwc_bf16_tc"]
PROTOTYPES["ivx_conv3d_ndhwc_f32x3_tc"] = PROTOTYPES["ivx_conv3d_ndhwc_bf16_tc"]
PROTOTYPES["ivx_split_tf32_ndhwc"] = [c_void_p, c_void_p, c_int64, c_int, c_void_p]
PROTOTYPES["ivx_conv3d_tc_workspace_bytes"] = [_ConvP]
c_float = ctypes.c_float
PROTOTYPES["ivx_channel_moments_ndhwc"] = [c_void_p, c_void_p, c_int64, c_int, c_void_p, c_void_p]
PROTOTYPES["ivx_affine_act_ndhwc"] = [c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, c_float, c_int64,
                                      c_int, c_void_p, c_void_p]
PROTOTYPES["ivx_affine_act_backward_ndhwc"] = [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int,
                                               c_float, c_int64, c_int, c_void_p, c_void_p, c_void_p]
PROTOTYPES["ivx_channel_lincomb_ndhwc"] = [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int64, c_int,
                                           c_void_p, c_void_p]
PROTOTYPES["ivx_conv3d_dgrad_ndhwc"] = [_ConvP, c_void_p, c_void_p, c_void_p, c_void_p]
PROTOTYPES["ivx_conv3d_wgrad_ndhwc"] = [_ConvP, c_void_p, c_void_p, c_void_p, c_void_p]
PROTOTYPES["ivx_upsample2x_trilinear_backward_ndhwc"] = [c_void_p, c_void_p, c_int, c_int, _i32p, c_void_p]
PROTOTYPES["ivx_upsample2x_trilinear_ndhwc"] = [c_void_p, c_void_p, c_int, c_int, _i32p, c_int, c_void_p]
PROTOTYPES["ivx_head_split_ndhwc"] = [c_void_p, c_int, c_int, c_int64, c_int, c_int, c_int, c_int, c_void_p,
                                      c_void_p, c_void_p]
PROTOTYPES["ivx_resize_mask_trilinear"] = [c_void_p, c_int, c_int, _i32p, _i32p, c_void_p, c_void_p]
PROTOTYPES["ivx_head_scores"] = [c_void_p, c_void_p, c_void_p, c_int, c_int64, c_int, c_void_p, c_void_p,
                                 c_void_p]
PROTOTYPES["ivx_compute_projection"] = [c_void_p, c_void_p, c_void_p, c_int, c_int, c_void_p, c_void_p]
PROTOTYPES["ivx_conv3d_wgrad_tf32_axes"] = [_ConvP, _i32p]
PROTOTYPES["ivx_conv3d_wgrad_tf32_tc"] = [_ConvP, c_void_p, c_void_p, c_void_p, c_void_p]
PROTOTYPES["ivx_conv3d_wgrad_f32x3_tc"] = PROTOTYPES["ivx_conv3d_wgrad_tf32_tc"]
PROTOTYPES["ivx_conv3d_ndhwc_tf32_tc_view"] = [_ConvP, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, _i64p,
                                               c_void_p, c_int64, c_void_p]
PROTOTYPES["ivx_conv3d_ndhwc_f32x3_tc_view"] = PROTOTYPES["ivx_conv3d_ndhwc_tf32_tc_view"]
PROTOTYPES["ivx_channel_major_ndhwc"] = [c_void_p, c_int, _i32p, c_int, _i32p, c_int, c_int, c_int, c_int, c_int,
                                         c_void_p, c_void_p]
RESTYPES = {"ivx_backproject_nchw_workspace_bytes": c_int64, "ivx_conv3d_tc_workspace_bytes": c_int64}

_lib = None


class IvxError(RuntimeError):
    pass


def lib():
    """The loaded library.  Raises if libivx.so is missing -- build it with
    ``python -m imvoxelnet_b200.build`` (needs nvcc; targets sm_100a)."""
    global _lib
    if _lib is None:
        if not os.path.isfile(LIB_PATH):
            raise IvxError(
                f"{LIB_PATH} not found: the CUDA library is required (there is no CPU or "
                "PyTorch fallback). Build it with `python -m imvoxelnet_b200.build`.")
        handle = ctypes.CDLL(LIB_PATH)
        handle.ivx_version.restype = ctypes.c_char_p
        handle.ivx_last_error.restype = ctypes.c_char_p
        for name, argtypes in PROTOTYPES.items():
            fn = getattr(handle, name)
            fn.argtypes = argtypes
            fn.restype = RESTYPES.get(name, c_int)
        _lib = handle
    return _lib


def check(rc, what):
    if rc != 0:
        msg = lib().ivx_last_error().decode()
        raise IvxError(f"{what} failed ({rc}): {msg}")


def version():
    return lib().ivx_version().decode()


def i64(values):
    return (ctypes.c_int64 * len(values))(*[int(v) for v in values])


def i32(values):
    return (ctypes.c_int32 * len(values))(*[int(v) for v in values])


def f32(values):
    return (ctypes.c_float * len(values))(*[float(v) for v in values])


def grid(n_voxels, voxel_size):
    g = Grid()
    for k in range(3):
        g.n_voxels[k] = int(n_voxels[k])
        g.voxel_size[k] = float(voxel_size[k])
    return g
