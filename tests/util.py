import numpy as np


def ulp_diff(a, b):
    """Distance in units in the last place between two float64 arrays."""
    a = np.ascontiguousarray(a, dtype=np.float64).view(np.int64).astype(np.int64)
    b = np.ascontiguousarray(b, dtype=np.float64).view(np.int64).astype(np.int64)
    a = np.where(a < 0, np.int64(-2**63) - a, a)
    b = np.where(b < 0, np.int64(-2**63) - b, b)
    return np.abs(a - b)


def rel_fro(got, want):
    got = np.asarray(got, dtype=np.float64)
    want = np.asarray(want, dtype=np.float64)
    den = np.linalg.norm(want)
    return np.linalg.norm(got - want) / (den if den > 0 else 1.0)
