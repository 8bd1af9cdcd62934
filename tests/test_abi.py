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
