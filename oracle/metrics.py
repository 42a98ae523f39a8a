"""The two error metrics of SURVEY.md F3 (TEST INFRASTRUCTURE)."""
import numpy as np


def raw_relative_error(a, b):
    """max |a-b| / |b| over elements with b != 0 (reported; gated only on positive data)."""
    a = np.asarray(a, dtype=np.float64).ravel()
    b = np.asarray(b, dtype=np.float64).ravel()
    nz = b != 0
    if not nz.any():
        return 0.0
    return float(np.max(np.abs(a[nz] - b[nz]) / np.abs(b[nz])))


def backward_error(a, b, per_rank_inputs):
    """max |a-b| / mean_r |g_r|: error normalised by the size of the summands (gated <= 1e-6)."""
    a = np.asarray(a, dtype=np.float64).ravel()
    b = np.asarray(b, dtype=np.float64).ravel()
    scale = np.mean([np.abs(np.asarray(g, dtype=np.float64).ravel()) for g in per_rank_inputs], axis=0)
    nz = scale != 0
    if not nz.any():
        return 0.0
    return float(np.max(np.abs(a[nz] - b[nz]) / scale[nz]))
