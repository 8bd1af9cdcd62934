"""The C-ABI library loads and exports every symbol include/ivx.h declares
(no GPU needed: nothing is launched), and the Python mirror refuses to run
without CUDA tensors instead of falling back."""
import ctypes
import os
import re

import pytest
import torch

from imvoxelnet_b200 import _lib, build, ops

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    text = open(os.path.join(ROOT, "include", "ivx.h")).read()
    return sorted(set(re.findall(r"IVX_API[^;(]*?\b(ivx_\w+)\s*\(", text)))


def test_library_is_built_for_sm100a():
    path = build.build()
    assert os.path.isfile(path)
    assert "sm_100a" in _lib.version()


def test_every_declared_symbol_is_exported():
    names = declared_symbols()
    assert len(names) >= 7
    handle = ctypes.CDLL(_lib.LIB_PATH)
    for name in names:
        assert hasattr(handle, name), f"{name} declared in ivx.h but not exported"
    # and the ctypes prototypes cover every compute entry point
    for name in names:
        if name not in ("ivx_version", "ivx_last_error"):
            assert name in _lib.PROTOTYPES, f"no ctypes prototype for {name}"


def test_no_unexpected_exports():
    import subprocess
    out = subprocess.run(["nm", "-D", "--defined-only", _lib.LIB_PATH], capture_output=True, text=True)
    exported = {l.split()[-1] for l in out.stdout.splitlines() if " T " in l}
    assert exported == set(declared_symbols())


def test_argument_errors_come_back_as_status_codes():
    # invalid arguments are rejected before anything touches the GPU
    rc = _lib.lib().ivx_backproject_scenes(None, None, 1, 1, 1, None, None, None, None, None, 10,
                                           None, None, 0, 0, None, None, None)
    assert rc == -1
    assert b"points or grid" in _lib.lib().ivx_last_error()
    g = _lib.grid((2, 2, 2), (1., 1., 1.))
    rc = _lib.lib().ivx_project_indices(None, g, _lib.f32([0, 0, 0]), 9, None, 1, 4, 4, None, None,
                                        None, None)
    assert rc == -1 and b"does not hold" in _lib.lib().ivx_last_error()


def test_cpu_tensors_are_refused():
    f = torch.zeros(1, 4, 4, 4)
    pts = torch.zeros(3, 2, 2, 2)
    P = torch.zeros(1, 3, 4)
    with pytest.raises(_lib.IvxError):
        ops.backproject(f, pts, P)
    with pytest.raises(_lib.IvxError):
        ops.backproject_scenes(f[None], [(4, 4)], P[None], (2, 2, 2), (1., 1., 1.), [(0., 0., 0.)])


def test_3xtf32_entry_points_validate_their_arguments():
    """ivx_conv3d_ndhwc_f32x3_tc / ivx_split_tf32_ndhwc / the fp32-training variants reject bad calls
    before anything is launched (host-side checks only: dummy non-NULL pointers are never dereferenced)."""
    L = _lib.lib()
    p = _lib.ConvParams()
    p.batch, p.c_in, p.c_out = 1, 64, 64
    for d in range(3):
        p.in_dims[d] = p.out_dims[d] = 8
        p.kernel[d], p.stride[d], p.pad[d] = 3, 1, 1
    p.post_scale = 1.0
    p.in_dtype, p.out_dtype = _lib.IVX_F32, _lib.IVX_F32           # plain fp32 in: not the split format
    dummy = 1 << 20
    assert L.ivx_conv3d_ndhwc_f32x3_tc(p, dummy, dummy, None, None, None, dummy, None, 0, None) == -1
    assert b"IVX_F32X2" in L.ivx_last_error()
    p.in_dtype = _lib.IVX_F32X2
    p.c_in = 48                                                    # not a multiple of 32
    assert L.ivx_conv3d_ndhwc_f32x3_tc(p, dummy, dummy, None, None, None, dummy, None, 0, None) == -2
    p.c_in = 64
    assert L.ivx_conv3d_ndhwc_f32x3_tc(p, dummy + 4, dummy, None, None, None, dummy, None, 0, None) == -1
    assert b"16-byte aligned" in L.ivx_last_error()
    assert L.ivx_conv3d_ndhwc_f32x3_tc(p, None, dummy, None, None, None, dummy, None, 0, None) == -1
    assert L.ivx_conv3d_ndhwc_f32x3_tc_view(p, dummy, dummy, None, None, dummy, None, None, 0, None) == -1
    assert L.ivx_conv3d_wgrad_f32x3_tc(p, None, dummy, dummy, None) == -1
    assert L.ivx_split_tf32_ndhwc(dummy, dummy, 10, 6, None) == -1  # channels not a multiple of 4
    assert b"multiple of 4" in L.ivx_last_error()
    assert L.ivx_split_tf32_ndhwc(None, dummy, 10, 8, None) == -1
    # the split-K workspace query knows the split format (and needs no GPU)
    assert L.ivx_conv3d_tc_workspace_bytes(p) >= 0
