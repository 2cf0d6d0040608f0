"""CWT / MMT / WZT apply on the GPU (through the C ABI) against the oracle."""
import ctypes

import numpy as np
import pytest
import scipy.sparse

from util import rel_fro

pytestmark = pytest.mark.gpu
TOL = 1e-12   # BASELINE.json: SA within 1e-12 relative Frobenius error in fp64


def _pair(sk, oracle, T, n, s, seed=11):
    from libskylark_b200 import lib
    c, oc = lib.Context(seed), oracle.Context(seed)
    if T == "CWT":
        return sk.CWT(n, s, c), oracle.CWT(n, s, oc)
    if T == "MMT":
        return sk.MMT(n, s, c), oracle.MMT(n, s, oc)
    return sk.WZT(n, s, 1.5, c), oracle.WZT(n, s, 1.5, oc)


@pytest.mark.parametrize("T", ["CWT", "MMT", "WZT"])
@pytest.mark.parametrize("shape", [(1000, 37, 64), (257, 8, 1), (5000, 7, 130), (300, 1000, 17),
                                   (4000, 7000, 5)])
def test_numpy_both_orders_both_dims(sk, oracle, T, shape):
    n, s, d = shape
    S, O = _pair(sk, oracle, T, n, s)
    rng = np.random.default_rng(1)
    A = rng.standard_normal((n, d))
    want = O.apply(A, "columnwise")
    assert rel_fro(S.apply(A.copy(), None, "columnwise"), want) < TOL          # C order
    assert rel_fro(S.apply(np.asfortranarray(A), None, 0), want) < TOL         # F order
    assert rel_fro(S * A.copy(), want) < TOL
    B = rng.standard_normal((d, n))
    want = O.apply(B, "rowwise")
    assert rel_fro(S.apply(B.copy(), None, "rowwise"), want) < TOL
    assert rel_fro(S.apply(np.asfortranarray(B), None, 1), want) < TOL
    assert rel_fro(S / B.copy(), want) < TOL


def test_cwt_rows_layout_is_bit_exact_when_buckets_are_short(sk, oracle):
    # every bucket <= 32 rows -> no atomics, rows added in ascending order with separate
    # multiply and add: the reference loop's exact result (hash_transform_Elemental.hpp:106-123)
    n, s, d = 20000, 4000, 96
    S, O = _pair(sk, oracle, "CWT", n, s, seed=3)
    assert np.bincount(O.row_idx.astype(int), minlength=s).max() <= 32
    A = np.random.default_rng(2).standard_normal((n, d))
    got = S.apply(A.copy(), None, "columnwise")
    assert np.array_equal(got, O.apply(A, "columnwise"))
    S2, O2 = _pair(sk, oracle, "MMT", n, s, seed=3)
    got, want = S2.apply(A.copy(), None, 0), O2.apply(A, "columnwise")
    assert rel_fro(got, want) < 1e-15     # values differ by <= 2 ulp, sums follow


@pytest.mark.parametrize("shape", [(3000, 50, 40), (4000, 3, 333), (70, 64, 9), (3500, 2000, 20),
                                   (50000, 2000, 6)])
def test_cwt_strided_layout(sk, oracle, shape):
    # column-major operand (sketched index contiguous) and its twin, a row-major operand sketched
    # rowwise: shared-memory accumulation, rows of a bucket arrive in no fixed order
    n, s, d = shape
    S, O = _pair(sk, oracle, "CWT", n, s, seed=5)
    A = np.random.default_rng(4).standard_normal((n, d))
    want = O.apply(A, "columnwise")
    assert rel_fro(S.apply(np.asfortranarray(A), None, "columnwise"), want) < 1e-14
    B = np.ascontiguousarray(A.T)
    assert rel_fro(S.apply(B, None, "rowwise"), want.T) < 1e-14


@pytest.mark.parametrize("fmt", ["csc", "csr"])
@pytest.mark.parametrize("T", ["CWT", "MMT"])
def test_scipy_sparse_input(sk, oracle, T, fmt):
    n, s, d = 3000, 64, 41
    S, O = _pair(sk, oracle, T, n, s)
    A = scipy.sparse.random(n, d, 0.05, format=fmt, random_state=3, dtype=np.float64)
    order = "F" if fmt == "csc" else "C"
    want = O.apply(A.toarray(), "columnwise")
    got = S.apply(A, np.empty((s, d), order=order), "columnwise")     # sparse in, dense out
    assert isinstance(got, np.ndarray) and got.shape == (s, d)
    assert rel_fro(got, want) < TOL
    B = scipy.sparse.random(d, n, 0.05, format=fmt, random_state=4, dtype=np.float64)
    assert rel_fro(S.apply(B, np.empty((d, s), order=order), "rowwise"),
                   O.apply(B.toarray(), "rowwise")) < TOL
    # empty matrix and empty columns
    E = scipy.sparse.csc_matrix((n, 5)) if fmt == "csc" else scipy.sparse.csr_matrix((n, 5))
    assert np.count_nonzero(S.apply(E, np.empty((s, 5), order=order), 0)) == 0


@pytest.mark.parametrize("T", ["CWT", "MMT"])
@pytest.mark.parametrize("case", ["few long columns", "all columns long"])
def test_csc_to_dense_long_columns(sk, oracle, T, case):
    """CSC -> dense in both kernel forms: the register dump (default) and the bulk-store slices, the latter
    with columns beyond the 2048 nonzeros its CTA holds in registers (overflow loops) and, when columns are
    long throughout, falling back to the dump by the average column length; s % 4 == 0 and a column-major
    result, as the bulk form needs."""
    n, s = 12000, 128
    rng = np.random.default_rng(5)
    lens = [60] * 30 + [3000, 5000, 2049] if case == "few long columns" else [2600] * 7
    cols = [np.sort(rng.choice(n, size=L, replace=False)) for L in lens]
    indptr = np.concatenate([[0], np.cumsum(lens)]).astype(np.int32)
    A = scipy.sparse.csc_matrix((rng.standard_normal(indptr[-1]), np.concatenate(cols).astype(np.int32), indptr),
                                shape=(n, len(lens)))
    S, O = _pair(sk, oracle, T, n, s)
    want = O.apply(A.toarray(), "columnwise")
    got = S.apply(A, np.empty((s, len(lens)), order="F"), "columnwise")
    assert rel_fro(got, want) < TOL
    from libskylark_b200 import lib
    lib.set_option("csc_form", "bulk")                   # the slices-through-the-copy-engine form
    try:
        for _ in range(2):                               # twice: the slice buffers are left clean
            assert rel_fro(S.apply(A, np.empty((s, len(lens)), order="F"), "columnwise"), want) < TOL
    finally:
        lib.set_option("csc_form", "dump")
    # a second apply starts from clean shared accumulators (sums of a slot are taken with shared-memory
    # atomics, so the two results agree to rounding, not bitwise)
    got2 = S.apply(A, np.empty((s, len(lens)), order="F"), "columnwise")
    assert rel_fro(got2, want) < TOL


def _canon(M):
    M = scipy.sparse.csc_matrix(M).copy()
    M.sort_indices()
    return M


@pytest.mark.parametrize("fmt", ["csc", "csr"])
@pytest.mark.parametrize("T", ["CWT", "MMT", "WZT"])
@pytest.mark.parametrize("shape", [(3000, 64, 41, 0.05), (500, 3, 200, 0.3), (40, 1, 7, 0.9),
                                   (2000, 5000, 30, 0.02),
                                   # >= 256 entries per column: the per-column shared-memory sort
                                   (6000, 64, 12, 0.2), (20000, 5000, 8, 0.09), (2100, 7, 5, 0.95),
                                   (30000, 900, 3, 0.15), (12000, 20000, 4, 0.8)])     # 4500 / 9600 per column
def test_scipy_sparse_to_sparse_is_the_reference_result(sk, oracle, T, fmt, shape):
    # sparse -> sparse is the default for scipy inputs, as in the reference (lib.py:257-266).
    # Entries of one output slot are added in input order on both sides, so values are
    # compared exactly (CWT) / to the 2-ulp value tolerance (MMT, WZT); the stored pattern,
    # including sums that cancel to zero, is identical after sorting rows inside columns.
    n, s, d, dens = shape
    S, O = _pair(sk, oracle, T, n, s)
    A = scipy.sparse.random(n, d, dens, format=fmt, random_state=3, dtype=np.float64)
    if dens > 0.2:                       # force cancellations: +-1 entries only
        A.data[:] = 1.0
    got = S.apply(A, None, "columnwise")
    assert scipy.sparse.issparse(got) and got.format == fmt and got.shape == (s, d)
    want = _canon(O.apply_csc_sparse(A, "columnwise"))
    g = _canon(got)
    assert np.array_equal(g.indptr, want.indptr) and np.array_equal(g.indices, want.indices)
    if T == "CWT":
        assert np.array_equal(g.data, want.data)
    else:
        assert np.allclose(g.data, want.data, rtol=1e-13, atol=1e-13 * np.abs(want.data).max())
    B = scipy.sparse.random(d, n, dens, format=fmt, random_state=4, dtype=np.float64)
    got = S.apply(B, None, "rowwise")
    assert got.format == fmt and got.shape == (d, s)
    want = _canon(O.apply_csc_sparse(B, "rowwise"))
    g = _canon(got)
    assert np.array_equal(g.indptr, want.indptr) and np.array_equal(g.indices, want.indices)
    if T == "CWT":
        assert np.array_equal(g.data, want.data)
    else:
        assert np.allclose(g.data, want.data, rtol=1e-13, atol=1e-13 * np.abs(want.data).max())
    # operators and an all-zero input
    assert rel_fro((S * A).toarray(), O.apply(A.toarray(), "columnwise")) < TOL
    E = scipy.sparse.csc_matrix((n, 5)) if fmt == "csc" else scipy.sparse.csr_matrix((n, 5))
    Z = S.apply(E, None, 0)
    assert Z.shape == (s, 5) and Z.nnz == 0
    # a preallocated sparse SA is overwritten in place
    out = scipy.sparse.csc_matrix((s, d)) if fmt == "csc" else scipy.sparse.csr_matrix((s, d))
    ret = S.apply(A, out, 0)
    assert ret is out and rel_fro(out.toarray(), O.apply(A.toarray(), "columnwise")) < TOL


def test_device_coalesce(sk):
    # merge step of the sparse exchange: equal (row, column) entries added in storage order
    import torch
    from libskylark_b200 import lib
    rng = np.random.default_rng(9)
    s, d, per = 50, 37, 400                       # 400 entries per column over 50 rows: long runs
    rows = rng.integers(0, s, size=(d, per)).astype(np.int32)
    vals = rng.standard_normal((d, per))
    indptr = np.arange(0, d * per + 1, per, dtype=np.int32)
    A = lib.DeviceCSC(torch.tensor(indptr, device="cuda"), torch.tensor(rows.reshape(-1), device="cuda"),
                      torch.tensor(vals.reshape(-1), device="cuda"), (s, d))
    got = sk.coalesce(A).to_scipy()
    want = np.zeros((s, d))
    for j in range(d):                            # left-to-right sums, as the kernel promises
        for r, v in zip(rows[j], vals[j]):
            want[r, j] += v
    assert got.has_sorted_indices or np.all(np.diff(got.indices[got.indptr[0]:got.indptr[1]]) > 0)
    assert np.array_equal(got.toarray(), want)
    short = lib.DeviceCSC(torch.tensor([0, 2, 2, 5], dtype=torch.int32, device="cuda"),
                          torch.tensor([3, 3, 0, 4, 0], dtype=torch.int32, device="cuda"),
                          torch.tensor([1.0, 2.0, 5.0, 7.0, -5.0], device="cuda", dtype=torch.float64), (6, 3))
    g = sk.coalesce(short).to_scipy()
    assert g.nnz == 3 and g[3, 0] == 3.0 and g[4, 2] == 7.0 and g[0, 2] == 0.0


def test_sparse_to_sparse_rejects_other_transforms(sk):
    from libskylark_b200 import errors
    A = scipy.sparse.random(100, 5, 0.1, format="csc", dtype=np.float64)
    with pytest.raises(errors.UnsupportedError):
        sk.JLT(100, 10).apply(A, None, 0)
    with pytest.raises(errors.UnsupportedError):      # dense in, sparse out
        sk.CWT(100, 10).apply(np.zeros((100, 5), order="F"), scipy.sparse.csc_matrix((10, 5)), 0)


def test_sparse_large_sketch_uses_global_path(sk, oracle):
    n, s, d = 2000, 40000, 9           # S*8 > shared memory
    S, O = _pair(sk, oracle, "CWT", n, s)
    A = scipy.sparse.random(n, d, 0.1, format="csc", random_state=5, dtype=np.float64)
    assert rel_fro(S.apply(A, np.empty((s, d), order="F"), 0), O.apply(A.toarray(), "columnwise")) < TOL


def test_torch_tensors_and_views(sk, oracle):
    import torch
    n, s, d = 6000, 100, 250
    S, O = _pair(sk, oracle, "CWT", n, s)
    A = torch.randn(n, d, dtype=torch.float64, device="cuda", generator=torch.Generator("cuda").manual_seed(0))
    want = O.apply(A.cpu().numpy(), "columnwise")
    got = S.apply(A, None, "columnwise")
    assert got.is_cuda and got.shape == (s, d) and got.is_contiguous()
    assert rel_fro(got.cpu().numpy(), want) < TOL
    At = A.t().contiguous().t()                     # column-major tensor
    got = S.apply(At, None, 0)
    assert got.stride(0) == 1
    assert rel_fro(got.cpu().numpy(), want) < TOL
    # strided view (leading dimension > width) and preallocated output
    big = torch.zeros(n, d + 6, dtype=torch.float64, device="cuda")
    big[:, :d] = A
    out = torch.empty(s, d, dtype=torch.float64, device="cuda")
    S.apply(big[:, :d], out, 0)
    assert rel_fro(out.cpu().numpy(), want) < TOL
    B = A.t().contiguous()                          # d x n row-major, rowwise
    assert rel_fro(S.apply(B, None, "rowwise").cpu().numpy(), want.T) < TOL


def test_device_csc(sk, oracle):
    import torch
    from libskylark_b200 import lib
    n, s, d = 5000, 300, 77
    S, O = _pair(sk, oracle, "MMT", n, s)
    A = scipy.sparse.random(n, d, 0.02, format="csc", random_state=6, dtype=np.float64)
    dev = lib.DeviceCSC(torch.tensor(A.indptr, dtype=torch.int32, device="cuda"),
                        torch.tensor(A.indices, dtype=torch.int32, device="cuda"),
                        torch.tensor(A.data, dtype=torch.float64, device="cuda"), A.shape)
    got = S.apply(dev, None, "columnwise")
    assert got.shape == (s, d) and got.stride(0) == 1
    assert rel_fro(got.cpu().numpy(), O.apply(A.toarray(), "columnwise")) < TOL
    # device sparse -> device sparse
    out = S.apply(dev, lib.DeviceCSC.empty((s, d)), "columnwise")
    assert out.values.is_cuda and out.shape == (s, d)
    want = _canon(O.apply_csc_sparse(A, "columnwise"))
    g = out.to_scipy()
    assert np.array_equal(g.indptr, want.indptr) and np.array_equal(g.indices, want.indices)
    assert np.allclose(g.data, want.data, rtol=1e-13, atol=1e-13 * np.abs(want.data).max())
    S2 = sk.MMT(n, s, context=lib.Context(11), defouttype="DeviceSpMatrix")
    assert np.array_equal(S2.apply(dev, None, 0).to_scipy().indices, want.indices)


def test_config1_full_size_parity_and_properties(sk, oracle):
    # BASELINE.json configs[0]: CWT columnwise, dense fp64 100k x 500 -> s = 2000
    import torch
    n, d, s = 100000, 500, 2000
    S, O = _pair(sk, oracle, "CWT", n, s, seed=123834)
    A = torch.randn(n, d, dtype=torch.float64, device="cuda", generator=torch.Generator("cuda").manual_seed(0))
    SA = S.apply(A, None, "columnwise")
    assert rel_fro(SA.cpu().numpy(), O.apply(A.cpu().numpy(), "columnwise")) < TOL
    # size-independent properties: column sums are preserved up to sign pattern ...
    signs = torch.tensor(O.row_value, device="cuda")
    assert torch.allclose(SA.sum(0), (signs[:, None] * A).sum(0), rtol=1e-9, atol=1e-9)
    # ... linearity, and shard additivity
    B = torch.randn(n, d, dtype=torch.float64, device="cuda")
    lhs = S.apply(2.0 * A - 3.0 * B, None, 0)
    rhs = 2.0 * SA - 3.0 * S.apply(B, None, 0)
    assert rel_fro(lhs.cpu().numpy(), rhs.cpu().numpy()) < 1e-13
    parts = sum(S.apply_shard(A[o:o + c], None, 0, o) for o, c in ((0, 33333), (33333, 40000), (73333, 26667)))
    assert rel_fro(parts.cpu().numpy(), SA.cpu().numpy()) < 1e-13
    # apply twice (tables and plan are cached): same result up to the order of the atomics
    assert rel_fro(S.apply(A, None, 0).cpu().numpy(), SA.cpu().numpy()) < 1e-15


def test_raw_c_abi_call(sk, oracle):
    """The same entry points the reference's lib.py binds, called with bare ctypes."""
    from libskylark_b200 import _build
    L = ctypes.CDLL(_build.SO)
    ctx = ctypes.c_void_p(); S = ctypes.c_void_p(); Aw = ctypes.c_void_p(); SAw = ctypes.c_void_p()
    assert L.sl_create_default_context(42, ctypes.byref(ctx)) == 0
    assert L.sl_create_sketch_transform(ctx, b"CWT", 500, 20, ctypes.byref(S)) == 0
    A = np.asfortranarray(np.random.default_rng(0).standard_normal((500, 30)))
    SA = np.empty((20, 30), order="F")
    dp = ctypes.POINTER(ctypes.c_double)
    assert L.sl_wrap_raw_matrix(A.ctypes.data_as(dp), 500, 30, ctypes.byref(Aw)) == 0
    assert L.sl_wrap_raw_matrix(SA.ctypes.data_as(dp), 20, 30, ctypes.byref(SAw)) == 0
    assert L.sl_apply_sketch_transform(S, b"Matrix", Aw, b"Matrix", SAw, 1) == 0
    want = oracle.CWT(500, 20, oracle.Context(42)).apply(A, "columnwise")
    assert rel_fro(SA, want) < TOL
    assert L.sl_apply_sketch_transform(S, b"DistMatrix", Aw, b"Matrix", SAw, 1) == 103
    assert L.sl_apply_sketch_transform(S, b"Matrix", Aw, b"Matrix", SAw, 7) == 111
    for w in (Aw, SAw):
        L.sl_free_raw_matrix_wrap(w)
    L.sl_free_sketch_transform(S)
    L.sl_free_context(ctx)


def test_edge_shapes(sk, oracle):
    """Empty and degenerate operands: no columns, one row, one bucket, more buckets than rows."""
    import torch
    S, O = _pair(sk, oracle, "CWT", 50, 7)
    out = S.apply(np.zeros((50, 0)), None, "columnwise")
    assert out.shape == (7, 0)
    out = S.apply(np.zeros((0, 50)), None, "rowwise")
    assert out.shape == (0, 7)
    S1, O1 = _pair(sk, oracle, "MMT", 1, 5)
    a = np.array([[2.5, -1.0, 0.0]])
    assert rel_fro(S1.apply(a.copy(), None, 0), O1.apply(a, "columnwise")) < TOL
    S2, O2 = _pair(sk, oracle, "CWT", 500, 1)          # everything lands in one bucket
    A = np.random.default_rng(0).standard_normal((500, 33))
    assert rel_fro(S2.apply(A.copy(), None, 0), O2.apply(A, "columnwise")) < TOL
    assert rel_fro(S2.apply(np.asfortranarray(A), None, 0), O2.apply(A, "columnwise")) < TOL
    S3, O3 = _pair(sk, oracle, "WZT", 40, 3000)        # mostly empty buckets
    B = np.random.default_rng(1).standard_normal((40, 5))
    got = S3.apply(B.copy(), None, 0)
    assert rel_fro(got, O3.apply(B, "columnwise")) < TOL
    assert np.count_nonzero(got.any(axis=1)) <= 40
    # a single column vector, both as (n,1) numpy and as a torch tensor
    S4, O4 = _pair(sk, oracle, "CWT", 3000, 64)
    v = np.random.default_rng(2).standard_normal((3000, 1))
    assert rel_fro(S4.apply(v.copy(), None, 0), O4.apply(v, "columnwise")) < TOL
    tv = torch.tensor(v, device="cuda")
    assert rel_fro(S4.apply(tv, None, 0).cpu().numpy(), O4.apply(v, "columnwise")) < TOL


def test_serialized_sketch_replays_identically(sk, oracle):
    """serialize -> deserialize (and pickle) rebuilds the same random state: identical sketch
    bits (the reference's SerializationTest.cpp:13-87 round-trip check)."""
    import pickle
    import torch
    from libskylark_b200 import lib
    c = lib.Context(2024, 77)
    A = torch.randn(3000, 40, dtype=torch.float64, device="cuda")
    for make in (lambda: sk.CWT(3000, 4000, c), lambda: sk.MMT(3000, 4000, c),
                 lambda: sk.WZT(3000, 4000, 1.5, c)):
        S = make()
        want = S.apply(A, None, 0)
        S2 = sk.deserialize_sketch(S.serialize())
        S3 = pickle.loads(pickle.dumps(S))
        assert torch.equal(S2.apply(A, None, 0), want)      # buckets <= 32 rows: no atomics
        assert torch.equal(S3.apply(A, None, 0), want)
        assert np.array_equal(S2.row_idx(), S.row_idx()) and np.array_equal(S2.row_value(), S.row_value())
    J = sk.JLT(3000, 64, c)
    J2 = sk.deserialize_sketch(J.serialize())
    assert np.array_equal(J2.export_dense(0, 0, 64, 50), J.export_dense(0, 0, 64, 50))
    F = sk.FJLT(4096, 100, c)
    F2 = pickle.loads(pickle.dumps(F))
    assert np.array_equal(F2.samples(), F.samples()) and np.array_equal(F2.diagonal(), F.diagonal())


@pytest.mark.parametrize("T,shape", [("CWT", (100000, 500, 2000)), ("CWT", (30001, 70, 513)), ("MMT", (40000, 37, 5000)),
                                     ("WZT", (50000, 26, 20000)), ("CWT", (4097, 300, 256)), ("CWT", (4097, 300, 64))])
def test_strided_layout_planned_kernel(sk, oracle, T, shape):
    """Column-major operands (the reference-native Elemental layout, hash_transform_Elemental.hpp:83-100)
    large enough for the atomics-free kernel: buckets owned by warps, rounds matched to bank pairs, tiles
    staged in shared memory.  C1's shape, 4 / 2 / 1 columns per CTA (by S), ragged column groups, odd n,
    valued transforms, the smallest S the kernel takes (256) and one below it (the shared-atomics kernel);
    CWT sums follow the reference's order, so the result is bit-identical to the oracle's."""
    import torch
    n, d, s = shape
    S, O = _pair(sk, oracle, T, n, s, seed=4711)
    g = torch.Generator("cuda").manual_seed(3)
    A = torch.randn(d, n, dtype=torch.float64, device="cuda", generator=g).t()        # column-major n x d
    want = O.apply(A.cpu().numpy(), "columnwise")
    SA = torch.full((d, s), float("nan"), dtype=torch.float64, device="cuda").t()
    for _ in range(2):
        S.apply(A, SA, "columnwise")
        got = SA.cpu().numpy()
        assert rel_fro(got, want) < TOL
    if T == "CWT" and d >= 480:                             # one index range per CTA: the reference's summation order
        assert np.array_equal(got, want)
    # rowwise on a row-major operand is the same kernel (the transposed view)
    B = A.t().contiguous()                                  # d x n row-major
    assert rel_fro(S.apply(B, None, "rowwise").cpu().numpy(), want.T) < TOL
    # row shards with global indices (own plans per shard), one of them at an odd offset
    cut = (n // 3) | 1
    parts = S.apply_shard(A[:cut], None, 0, 0) + S.apply_shard(A[cut:], None, 0, cut)
    assert rel_fro(parts.cpu().numpy(), want) < TOL


@pytest.mark.parametrize("nseg,shape", [(3, (500, 40)), (8, (20000, 64)), (2, (7, 3))])
def test_device_merge_segments(sk, nseg, shape):
    """sl_sp_merge_segments: rank-major CSC pieces (the receive buffers of the sparse exchange) added
    without re-ordering them; pattern and values against scipy's sum of the pieces."""
    import scipy.sparse
    import torch
    m, n = shape
    rng = np.random.default_rng(5)
    dens = min(0.5, 600.0 / m)
    pieces = [scipy.sparse.random(m, n, dens, format="csc", random_state=int(rng.integers(1 << 30)), dtype=np.float64)
              for _ in range(nseg)]
    for p in pieces:
        p.sort_indices()
    want = sum(pieces[1:], pieces[0]).tocsc()
    want.sort_indices()
    idx = np.concatenate([p.indices for p in pieces]).astype(np.int32)
    val = np.concatenate([p.data for p in pieces])
    vptr = np.zeros(nseg * n + 1, dtype=np.int32)
    vptr[1:] = np.cumsum(np.concatenate([np.diff(p.indptr) for p in pieces]))
    got = sk.merge_segments(torch.from_numpy(vptr).cuda(), torch.from_numpy(idx).cuda(), torch.from_numpy(val).cuda(),
                            nseg, (m, n)).to_scipy()
    got.sort_indices()
    # entries that cancel to zero stay stored, as in the sketch itself: compare as dense plus pattern superset
    assert np.allclose(got.toarray(), want.toarray(), rtol=1e-13, atol=1e-13)
    assert got.indptr[-1] >= want.indptr[-1] and np.all(np.diff(got.indptr) >= 0)
    for j in range(n):
        rows = got.indices[got.indptr[j]:got.indptr[j + 1]]
        assert np.all(np.diff(rows) > 0)                    # sorted, merged (no duplicates)
