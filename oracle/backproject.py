"""numpy front-end of the C oracle (oracle/backproject_oracle.c).

TEST INFRASTRUCTURE ONLY (see oracle/__init__.py).  Each function names the
reference lines it restates; all paths are relative to
/root/reference/mmdet3d/models/detectors/imvoxelnet.py.
"""
import ctypes
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "libivx_oracle.so")
_lib = None

_f32p = ctypes.POINTER(ctypes.c_float)
_i64p = ctypes.POINTER(ctypes.c_int64)
_i32p = ctypes.POINTER(ctypes.c_int32)
_u8p = ctypes.POINTER(ctypes.c_uint8)


def build(force=False):
    """Compile the C oracle with the committed Makefile (gcc only)."""
    if force or not os.path.isfile(_SO) or (
            os.path.getmtime(_SO) < os.path.getmtime(os.path.join(_HERE, "backproject_oracle.c"))):
        subprocess.run(["make", "-C", _HERE, "-B"], check=True, capture_output=True)
    return _SO


def lib():
    global _lib
    if _lib is None:
        if not os.path.isfile(_SO):
            build()
        _lib = ctypes.CDLL(_SO)
        _lib.ivx_oracle_max_threads.restype = ctypes.c_int
    return _lib


def _p(a, t):
    return a.ctypes.data_as(t)


def _c(a, dtype):
    return np.ascontiguousarray(a, dtype=dtype)


def set_threads(n):
    lib().ivx_oracle_set_threads(ctypes.c_int(int(n)))


def max_threads():
    return int(lib().ivx_oracle_max_threads())


def new_origin(n_voxels, voxel_size, origin):
    """``origin - n_voxels / 2. * voxel_size`` in fp32 (:139)."""
    nv, vs, og = _c(n_voxels, np.int32), _c(voxel_size, np.float32), _c(origin, np.float32)
    out = np.empty(3, np.float32)
    lib().ivx_oracle_new_origin(_p(nv, _i32p), _p(vs, _f32p), _p(og, _f32p), _p(out, _f32p))
    return out


def get_points(n_voxels, voxel_size, origin):
    """get_points (:132-141) -> [3, X, Y, Z] fp32."""
    nv, vs, og = _c(n_voxels, np.int32), _c(voxel_size, np.float32), _c(origin, np.float32)
    pts = np.empty((3, int(nv[0]), int(nv[1]), int(nv[2])), np.float32)
    lib().ivx_oracle_points(_p(nv, _i32p), _p(vs, _f32p), _p(og, _f32p), _p(pts, _f32p))
    return pts


def project(points, projection, height, width):
    """:148-154 -> (x [V,N] int64, y [V,N] int64, valid [V,N] bool)."""
    pts = _c(points, np.float32).reshape(3, -1)
    proj = _c(projection, np.float32).reshape(-1, 12)
    V, N = proj.shape[0], pts.shape[1]
    x = np.empty((V, N), np.int64)
    y = np.empty((V, N), np.int64)
    valid = np.empty((V, N), np.uint8)
    lib().ivx_oracle_project(_p(pts, _f32p), ctypes.c_int64(N), _p(proj, _f32p),
                             ctypes.c_int(V), ctypes.c_int(int(height)), ctypes.c_int(int(width)),
                             _p(x, _i64p), _p(y, _i64p), _p(valid, _u8p))
    return x, y, valid.astype(bool)


def _feat_args(features):
    f = np.asarray(features)
    assert f.dtype == np.float32 and f.ndim == 4
    st = [s // 4 for s in f.strides]
    return f, [ctypes.c_int64(s) for s in st]


def backproject(features, points, projection):
    """backproject (:145-160).  ``features`` may be any strided fp32 view
    [V, C, h, w]; returns (volume [V,C,X,Y,Z] fp32, valid [V,1,X,Y,Z] bool)."""
    f, st = _feat_args(features)
    V, C, h, w = f.shape
    pts = _c(points, np.float32)
    X, Y, Z = pts.shape[-3:]
    pts = pts.reshape(3, -1)
    N = pts.shape[1]
    proj = _c(projection, np.float32).reshape(-1, 12)
    assert proj.shape[0] == V
    vol = np.empty((V, C, N), np.float32)
    valid = np.empty((V, N), np.uint8)
    lib().ivx_oracle_backproject(_p(f, _f32p), *st, ctypes.c_int(V), ctypes.c_int(C),
                                 ctypes.c_int(h), ctypes.c_int(w), _p(pts, _f32p),
                                 ctypes.c_int64(N), _p(proj, _f32p), _p(vol, _f32p), _p(valid, _u8p))
    return vol.reshape(V, C, X, Y, Z), valid.astype(bool).reshape(V, 1, X, Y, Z)


def view_mean(volume, valid):
    """:70-74 -> (mean [C,X,Y,Z] fp32, valid [1,X,Y,Z] bool, count [X,Y,Z] int32)."""
    vol = _c(volume, np.float32)
    V, C = vol.shape[:2]
    shp = vol.shape[2:]
    N = int(np.prod(shp))
    va = _c(np.asarray(valid).reshape(V, N), np.uint8)
    out = np.empty((C, N), np.float32)
    vo = np.empty(N, np.uint8)
    cnt = np.empty(N, np.int32)
    lib().ivx_oracle_view_mean(_p(vol, _f32p), _p(va, _u8p), ctypes.c_int(V), ctypes.c_int(C),
                               ctypes.c_int64(N), _p(out, _f32p), _p(vo, _u8p), _p(cnt, _i32p))
    return out.reshape((C,) + shp), vo.astype(bool).reshape((1,) + shp), cnt.reshape(shp)


def backproject_mean(features, points, projection):
    """:69-74 fused (no [V,C,N] temporary).  Same returns as ``view_mean``."""
    f, st = _feat_args(features)
    V, C, h, w = f.shape
    pts = _c(points, np.float32)
    shp = pts.shape[-3:]
    pts = pts.reshape(3, -1)
    N = pts.shape[1]
    proj = _c(projection, np.float32).reshape(-1, 12)
    assert proj.shape[0] == V
    out = np.empty((C, N), np.float32)
    vo = np.empty(N, np.uint8)
    cnt = np.empty(N, np.int32)
    lib().ivx_oracle_backproject_mean(_p(f, _f32p), *st, ctypes.c_int(V), ctypes.c_int(C),
                                      ctypes.c_int(h), ctypes.c_int(w), _p(pts, _f32p),
                                      ctypes.c_int64(N), _p(proj, _f32p), _p(out, _f32p),
                                      _p(vo, _u8p), _p(cnt, _i32p))
    return out.reshape((C,) + shp), vo.astype(bool).reshape((1,) + shp), cnt.reshape(shp)


def distinct_pixels(x, y, valid, height, width):
    """U of SURVEY.md section 8(d): number of distinct (view, y, x) feature
    pixels referenced by valid voxel-views (exact, from the oracle's indices)."""
    V = x.shape[0]
    v = np.broadcast_to(np.arange(V, dtype=np.int64)[:, None], x.shape)
    key = (v[valid] * height + y[valid]) * width + x[valid]
    return int(np.unique(key).size)
