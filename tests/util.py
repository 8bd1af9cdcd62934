"""Shared helpers of the test-suite."""
import hashlib
import os

import numpy as np

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")

# north_star tolerance for fp32 volumes: 1e-5 relative.  "Relative" is taken
# against max|ref| of the tensor, plus an elementwise atol of 1e-6 so that
# cancellation in a sum over views cannot blow the ratio up.
VOLUME_RTOL = 1e-5
VOLUME_ATOL = 1e-6


def load_golden(name):
    return np.load(os.path.join(GOLDEN, name + ".npz"))


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def assert_volume_close(got, ref, rel=VOLUME_RTOL, atol=VOLUME_ATOL):
    got = np.asarray(got, dtype=np.float64)
    ref = np.asarray(ref, dtype=np.float64)
    assert got.shape == ref.shape, (got.shape, ref.shape)
    scale = float(np.abs(ref).max()) if ref.size else 1.0
    err = np.abs(got - ref)
    worst = float(err.max()) if err.size else 0.0
    assert worst <= rel * (scale or 1.0) + atol, f"max err {worst:.3e} (scale {scale:.3e})"


def rel_l2(got, ref):
    """||got - ref||_2 / ||ref||_2 (float64) of two arrays or torch tensors."""
    if hasattr(got, "detach"):
        d = (got.detach().double() - ref.detach().double()).norm().item()
        return d / max(ref.detach().double().norm().item(), 1e-300)
    got, ref = np.asarray(got, dtype=np.float64), np.asarray(ref, dtype=np.float64)
    return float(np.linalg.norm(got - ref) / max(np.linalg.norm(ref), 1e-300))
