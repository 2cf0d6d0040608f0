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


def golden_input(n, d):
    """The fixed input of tests/golden/golden.json["applies"] (exact binary fractions)."""
    import numpy as np
    A = (((7 * np.arange(n)[:, None] + 3 * np.arange(d)[None, :]) % 11) - 5) / 4.0
    Asp = A * (((np.arange(n)[:, None] + 2 * np.arange(d)[None, :]) % 3) == 0)
    return A, Asp


def unhex(lst, shape):
    import numpy as np
    return np.array([float.fromhex(x) for x in lst]).reshape(shape, order="F")
