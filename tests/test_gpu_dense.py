"""JLT / CT / random Fourier features on the GPU (through the C ABI) against the oracle."""
import numpy as np
import pytest

from util import rel_fro

pytestmark = pytest.mark.gpu
TOL64 = 1e-12   # BASELINE.json: 1e-12 relative Frobenius in fp64
TOL32 = 1e-5    #                1e-5 in fp32


def _ctx(sk, oracle, seed, counter=0):
    from libskylark_b200 import lib
    return lib.Context(seed, counter), oracle.Context(seed, counter)


@pytest.mark.parametrize("shape", [(300, 70, 45), (128, 128, 128), (1000, 257, 3), (17, 5, 200),
                                   (2048, 130, 131)])
@pytest.mark.parametrize("T", ["JLT", "CT"])
def test_numpy_both_orders_both_dims(sk, oracle, T, shape):
    n, s, d = shape
    c, oc = _ctx(sk, oracle, 21)
    S, O = (sk.JLT(n, s, c), oracle.JLT(n, s, oc)) if T == "JLT" else \
           (sk.CT(n, s, 1.5, c), oracle.CT(n, s, 1.5, oc))
    rng = np.random.default_rng(7)
    A = rng.standard_normal((n, d))
    want = O.apply(A, "columnwise")
    assert rel_fro(S.apply(A.copy(), None, "columnwise"), want) < TOL64
    assert rel_fro(S.apply(np.asfortranarray(A), None, 0), want) < TOL64
    B = rng.standard_normal((d, n))
    want = O.apply(B, "rowwise")
    assert rel_fro(S.apply(B.copy(), None, "rowwise"), want) < TOL64
    assert rel_fro(S.apply(np.asfortranarray(B), None, 1), want) < TOL64


def test_many_panels_and_shards(sk, oracle):
    import torch
    n, s, d = 20000, 48, 72      # one panel (cap 32768 rows); several panels: next test
    c, oc = _ctx(sk, oracle, 5, 12345)
    S, O = sk.JLT(n, s, c), oracle.JLT(n, s, oc)
    A = torch.randn(n, d, dtype=torch.float64, device="cuda", generator=torch.Generator("cuda").manual_seed(1))
    want = O.apply(A.cpu().numpy(), "columnwise")
    SA = S.apply(A, None, "columnwise")
    assert rel_fro(SA.cpu().numpy(), want) < TOL64
    # column-major tensor takes the other kernel layout
    SAf = S.apply(A.t().contiguous().t(), None, 0)
    assert rel_fro(SAf.cpu().numpy(), want) < TOL64
    # row shards with global input indices sum to the full sketch
    parts = sum(S.apply_shard(A[o:o + k], None, 0, o) for o, k in ((0, 7001), (7001, 9999), (17000, 3000)))
    assert rel_fro(parts.cpu().numpy(), want) < TOL64
    # linearity
    B = torch.randn(n, d, dtype=torch.float64, device="cuda")
    lhs = S.apply(A + 0.5 * B, None, 0)
    assert rel_fro(lhs.cpu().numpy(), (SA + 0.5 * S.apply(B, None, 0)).cpu().numpy()) < 1e-13


@pytest.mark.parametrize("shape", [(70000, 48, 72), (20000, 4096, 64)])
def test_benchmarked_multi_panel_device_path(sk, oracle, shape):
    """The path bench.py times: device tensors, several generated panels (cap 32768 rows, or
    256 MiB: 8192 rows at s = 4096 as in config C2), double-buffered on the forked side
    streams, not accumulating.  Both operand layouts, twice in a row (buffers are reused)."""
    import torch
    n, s, d = shape
    c, oc = _ctx(sk, oracle, 5, 999)
    S, O = sk.JLT(n, s, c), oracle.JLT(n, s, oc)
    A = torch.randn(n, d, dtype=torch.float64, device="cuda", generator=torch.Generator("cuda").manual_seed(2))
    want = O.apply(A.cpu().numpy(), "columnwise")
    SA = torch.empty(s, d, dtype=torch.float64, device="cuda")
    for _ in range(2):
        SA.fill_(float("nan"))
        S.apply(A, SA, "columnwise")
        assert rel_fro(SA.cpu().numpy(), want) < TOL64
    Af = A.t().contiguous().t()                       # column-major operand
    SAf = torch.empty(d, s, dtype=torch.float64, device="cuda").t()
    S.apply(Af, SAf, "columnwise")
    assert rel_fro(SAf.cpu().numpy(), want) < TOL64
    half = (n // 2) // 16 * 16 + 8                    # shard boundary inside a panel, odd offset/8
    parts = S.apply_shard(A[:half], None, 0, 0) + S.apply_shard(A[half:], None, 0, half)
    assert rel_fro(parts.cpu().numpy(), want) < TOL64


def test_tile_mode_direct_store(sk, oracle):
    import torch
    n, s, m = 40, 256, 38400      # 2 x 300 output tiles >= 4*148 -> whole-tile partition
    c, oc = _ctx(sk, oracle, 8)
    S, O = sk.JLT(n, s, c), oracle.JLT(n, s, oc)
    A = torch.randn(m, n, dtype=torch.float64, device="cuda")
    got = S.apply(A, None, "rowwise")
    assert rel_fro(got.cpu().numpy(), O.apply(A.cpu().numpy(), "rowwise")) < TOL64


@pytest.mark.parametrize("T", ["GaussianRFT", "LaplacianRFT"])
def test_rft_fp64(sk, oracle, T):
    import torch
    n, s = 60, 200
    c, oc = _ctx(sk, oracle, 13)
    S = getattr(sk, T)(n, s, 2.5, context=c)
    O = getattr(oracle, T)(n, s, 2.5, oc)
    rng = np.random.default_rng(3)
    X = rng.uniform(0, 1, (500, n))
    want = O.apply(X, "rowwise")
    # Laplacian features take cos() of Cauchy-weighted sums whose magnitude reaches 1e4..1e5:
    # the summation order alone (numpy's blocked dgemm vs any other order) moves the argument
    # by eps*|arg|, so the features are only defined to ~1e-11 there.  Gaussian: 1e-12.
    tol = TOL64 if T == "GaussianRFT" else 1e-10
    assert rel_fro(S.apply(X.copy(), None, "rowwise"), want) < tol
    assert rel_fro(S.apply(np.asfortranarray(X), None, 1), want) < tol
    Xc = rng.uniform(0, 1, (n, 33))
    assert rel_fro(S.apply(Xc.copy(), None, "columnwise"), O.apply(Xc, "columnwise")) < tol
    # feature-map shape that takes the fused-epilogue (whole tile) path
    Xb = torch.rand(80000, n, dtype=torch.float64, device="cuda")
    got = S.apply(Xb, None, "rowwise")
    assert rel_fro(got.cpu().numpy(), O.apply(Xb.cpu().numpy(), "rowwise")) < tol
    assert float(got.abs().max()) <= np.sqrt(2.0 / s) * (1 + 1e-15)
    with pytest.raises(Exception):
        S.apply_shard(Xc[:30].copy(), None, 0, 0)       # cosine is not additive over shards


def test_fp32_tensors(sk, oracle):
    import torch
    n, s = 784, 512
    c, oc = _ctx(sk, oracle, 99)
    S, O = sk.GaussianRFT(n, s, 10.0, context=c), oracle.GaussianRFT(n, s, 10.0, oc)
    X = torch.rand(3000, n, dtype=torch.float32, device="cuda")
    got = S.apply(X, None, "rowwise")
    assert got.dtype == torch.float32
    want = O.apply(X.cpu().numpy().astype(np.float64), "rowwise")
    assert rel_fro(got.cpu().numpy(), want) < TOL32
    # the reference's own float path (S cast to float, sgemm, float epilogue) is no closer
    ref32 = O.apply(X.cpu().numpy(), "rowwise", dtype=np.float32)
    assert rel_fro(got.cpu().numpy(), ref32) < TOL32
    Xl = torch.rand(80000, 64, dtype=torch.float32, device="cuda")    # whole-tile path, float
    S2, O2 = sk.GaussianRFT(64, 256, 1.0, context=c), oracle.GaussianRFT(64, 256, 1.0, oc)
    got = S2.apply(Xl, None, "rowwise")
    assert rel_fro(got.cpu().numpy(), O2.apply(Xl.cpu().numpy().astype(np.float64), "rowwise")) < TOL32
    J, OJ = sk.JLT(5000, 40, context=c), oracle.JLT(5000, 40, oc)
    A = torch.randn(5000, 100, dtype=torch.float32, device="cuda")
    got = J.apply(A, None, "columnwise")
    assert rel_fro(got.cpu().numpy(), OJ.apply(A.cpu().numpy().astype(np.float64), "columnwise")) < TOL32


def test_large_host_arrays_take_the_pipelined_path(sk, oracle):
    """numpy operands above 64 MB are staged chunk by chunk (copy overlapped with compute)."""
    n, s, d = 70000, 48, 160        # 89.6 MB
    c, oc = _ctx(sk, oracle, 77)
    S, O = sk.JLT(n, s, c), oracle.JLT(n, s, oc)
    A = np.random.default_rng(9).standard_normal((n, d))
    want = O.apply(A, "columnwise")
    assert rel_fro(S.apply(A.copy(), None, "columnwise"), want) < TOL64          # C order
    assert rel_fro(S.apply(np.asfortranarray(A), None, "columnwise"), want) < TOL64   # F order
    B = np.ascontiguousarray(A.T)
    assert rel_fro(S.apply(B, None, "rowwise"), want.T) < TOL64
    half = n // 2
    part = S.apply_shard(A[:half].copy(), None, 0, 0) + S.apply_shard(A[half:].copy(), None, 0, half)
    assert rel_fro(part, want) < TOL64


def test_large_pageable_operands_and_results(sk, oracle):
    """Host arrays above 64 MiB on the generic path (hash transforms, feature maps): the operand goes to
    the device, and the result comes back, in chunks through pinned memory filled by worker threads."""
    c, oc = _ctx(sk, oracle, 78)
    n, s, d = 60000, 64, 200                                     # 96 MB in
    A = np.random.default_rng(10).standard_normal((n, d))
    S, O = sk.CWT(n, s, c), oracle.CWT(n, s, oc)
    want = O.apply(A, "columnwise")
    assert rel_fro(S.apply(np.asfortranarray(A), None, "columnwise"), want) < TOL64
    assert rel_fro(S.apply(A.copy(), None, "columnwise"), want) < TOL64
    m, k, f = 45000, 40, 256                                     # 92 MB out
    X = np.random.default_rng(11).random((m, k))
    R, Ro = sk.GaussianRFT(k, f, 3.0, context=c), oracle.GaussianRFT(k, f, 3.0, oc)
    assert rel_fro(R.apply(X, None, "rowwise"), Ro.apply(X, "rowwise")) < TOL64


def test_edge_shapes(sk, oracle):
    c, oc = _ctx(sk, oracle, 3)
    J, OJ = sk.JLT(37, 1, c), oracle.JLT(37, 1, oc)              # one sketch row
    A = np.random.default_rng(0).standard_normal((37, 9))
    assert rel_fro(J.apply(A.copy(), None, 0), OJ.apply(A, "columnwise")) < TOL64
    assert J.apply(np.zeros((37, 0)), None, 0).shape == (1, 0)   # no columns
    J1, OJ1 = sk.JLT(1, 5, c), oracle.JLT(1, 5, oc)              # one input row
    a = np.array([[1.5, -2.0]])
    assert rel_fro(J1.apply(a.copy(), None, 0), OJ1.apply(a, "columnwise")) < TOL64
    v = np.random.default_rng(1).standard_normal((37, 1))        # a single vector
    assert rel_fro(J.apply(v.copy(), None, 0), OJ.apply(v, "columnwise")) < TOL64


@pytest.mark.parametrize("fmt", ["csc", "csr"])
@pytest.mark.parametrize("T", ["JLT", "CT", "GaussianRFT", "LaplacianRFT"])
def test_dense_transform_of_sparse_operand(sk, oracle, T, fmt):
    # (T, SparseMatrix, Matrix) of capi/csketch.cpp:80,100,177,197: same result as for the dense
    # matrix with the same entries
    import scipy.sparse
    from libskylark_b200 import lib
    n, s, d = 700, 48, 23
    mk = {"JLT": lambda m, c: m.JLT(n, s, c), "CT": lambda m, c: m.CT(n, s, 2.5, c),
          "GaussianRFT": lambda m, c: m.GaussianRFT(n, s, 3.0, c),
          "LaplacianRFT": lambda m, c: m.LaplacianRFT(n, s, 3.0, c)}[T]
    S, O = mk(sk, lib.Context(31)), mk(oracle, oracle.Context(31))
    tol = 1e-10 if T == "LaplacianRFT" else 1e-12
    order = "F" if fmt == "csc" else "C"
    A = scipy.sparse.random(n, d, 0.1, format=fmt, random_state=2, dtype=np.float64)
    got = S.apply(A, np.empty((s, d), order=order), "columnwise")
    assert rel_fro(got, O.apply(A.toarray(), "columnwise")) < tol
    B = scipy.sparse.random(d, n, 0.1, format=fmt, random_state=3, dtype=np.float64)
    got = S.apply(B, np.empty((d, s), order=order), "rowwise")
    assert rel_fro(got, O.apply(B.toarray(), "rowwise")) < tol
    params = {"JLT": (), "CT": (2.5,), "GaussianRFT": (3.0,), "LaplacianRFT": (3.0,)}[T]
    S2 = getattr(sk, T)(n, s, *params, defouttype="LocalMatrix", context=lib.Context(31))
    got = S2.apply(A, None, 0)                          # the reference's way to ask for a dense result
    assert isinstance(got, np.ndarray) and rel_fro(got, O.apply(A.toarray(), "columnwise")) < tol


@pytest.mark.parametrize("shape", [(784, 16384, 700), (100, 520, 133), (36, 12, 5), (16, 4, 1), (1028, 260, 300),
                                   (2000, 128, 129)])      # the last two: k > 1024 -> the draining form
def test_fp32_feature_map_keeps_fp32_accuracy(sk, oracle, shape):
    # the strided fp32 case runs on the TF32 tensor cores with operand splitting.  Tensor-core
    # accumulation truncates, so one chain over all three products would sit at ~9e-6; the
    # tcgen05 kernel keeps the large term in its own accumulator (3.4e-6 at k = 784) and, for
    # k > 1024, drains it every 32 MMAs (1.4e-6); the mma.sync kernel restarts its accumulator
    # every 8 (6.5e-7) -- asserted inside the 1e-5 bar, edge tiles included
    import torch
    n, s, m = shape
    c, oc = _ctx(sk, oracle, 7)
    S, O = sk.GaussianRFT(n, s, 10.0, context=c), oracle.GaussianRFT(n, s, 10.0, oc)
    X = torch.rand(m, n, dtype=torch.float32, device="cuda", generator=torch.Generator("cuda").manual_seed(1))
    got = S.apply(X, None, "rowwise").cpu().numpy()
    want = O.apply(X.cpu().numpy().astype(np.float64), "rowwise")
    assert rel_fro(got, want) < 5e-6
    J, OJ = sk.JLT(n, s, context=c), oracle.JLT(n, s, oc)        # no epilogue, same kernel
    got = J.apply(X, None, "rowwise").cpu().numpy()
    assert rel_fro(got, OJ.apply(X.cpu().numpy().astype(np.float64), "rowwise")) < 5e-6


@pytest.mark.parametrize("env", [{"SKB_TC_FORM": "persist"}, {"SKB_TC_FORM": "nonpersistent"},
                                 {"SKB_DEBUG_NO_TMA": "1"}, {"SKB_DEBUG_NO_TCGEN05": "1"}, {"SKB_DEBUG_NO_TF32": "1"}])
def test_fp32_feature_map_fallback_forms(env):
    # the other forms of the fp32 feature-map kernel (128 x 128 persistent / per-tile tcgen05 kernels, the
    # tcgen05 kernel without TMA, the mma.sync kernel, FP64 DMMA) are selected by environment variables that
    # the library reads once, so each runs in its own process
    import os, subprocess, sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    e = dict(os.environ); e.update(env)
    r = subprocess.run([sys.executable, "-m", "pytest", os.path.join(root, "tests", "test_gpu_dense.py"), "-x", "-q",
                        "-k", "fp32_feature_map_keeps or fp32_tensors"], env=e, cwd=root, capture_output=True, text=True,
                       timeout=300)
    assert r.returncode == 0, r.stdout[-1500:] + r.stderr[-500:]
