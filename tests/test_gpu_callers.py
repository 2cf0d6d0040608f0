"""The callers one level above the sketch path, on the device: randomized SVD
and sketched least squares (nla/svd.hpp, nla/least_squares.hpp) and the kernel objects whose
create_rft builds the feature transforms (ml/kernels.hpp:53-66).  The reference tests the SVD by
reconstruction (tests/unit/SVDElementalTest.cpp); so do these, plus parity of the random part."""
import numpy as np
import pytest

from util import rel_fro

pytestmark = pytest.mark.gpu


def test_approximate_svd_reconstructs_low_rank(sk, oracle):
    import torch
    from libskylark_b200 import lib, nla
    g = torch.Generator("cuda").manual_seed(3)
    m, n, r = 3000, 400, 12
    A = torch.randn(m, r, dtype=torch.float64, device="cuda", generator=g) @ \
        torch.randn(r, n, dtype=torch.float64, device="cuda", generator=g)
    for M in (A, A.t().contiguous()):                     # tall and wide
        U, S, V = nla.approximate_svd(M, r, context=lib.Context(7))
        assert rel_fro(((U * S) @ V.t()).cpu().numpy(), M.cpu().numpy()) < 1e-10
        assert rel_fro((U.t() @ U).cpu().numpy(), np.eye(r)) < 1e-10
        exact = torch.linalg.svdvals(M)[:r]
        assert float((S - exact).abs().max() / exact[0]) < 1e-10
    # the range finder is this package's JLT from the given context: same seed, same Q as the oracle's S
    c = lib.Context(7)
    nla.approximate_svd(A, r, nla.SVDParams(num_iterations=0), context=c)
    assert c.counter == n * min(n, 2 * r)                 # JLT(n, k) consumed n*k counters (App. B)
    Asym = A[:n].t() @ A[:n]
    Vs, w = nla.approximate_symmetric_svd(Asym, r, context=lib.Context(8))
    assert rel_fro(((Vs * w) @ Vs.t()).cpu().numpy(), Asym.cpu().numpy()) < 1e-9


def test_sketched_least_squares(sk):
    import torch
    from libskylark_b200 import lib, nla
    g = torch.Generator("cuda").manual_seed(5)
    m, n = 16384, 60
    A = torch.randn(m, n, dtype=torch.float64, device="cuda", generator=g)
    xtrue = torch.randn(n, 3, dtype=torch.float64, device="cuda", generator=g)
    B = A @ xtrue + 1e-3 * torch.randn(m, 3, dtype=torch.float64, device="cuda", generator=g)
    exact = torch.linalg.lstsq(A, B).solution
    X = nla.faster_least_squares(A, B, context=lib.Context(11))
    assert rel_fro(X.cpu().numpy(), exact.cpu().numpy()) < 1e-10          # preconditioned LSQR: full accuracy
    Xa = nla.approximate_least_squares(A, B, context=lib.Context(12))
    res = lambda Z: float((A @ Z - B).norm())
    assert res(Xa) <= 1.5 * res(exact)                                     # sketch-and-solve: (1+eps) residual


def test_kernels_build_the_feature_transforms(sk, oracle, tmp_path):
    import torch
    from libskylark_b200 import lib
    from libskylark_b200.ml import io, kernels
    d, s, npts = 20, 4096, 300
    g = torch.Generator("cuda").manual_seed(9)
    X = torch.rand(npts, d, dtype=torch.float64, device="cuda", generator=g)
    for K, O in ((kernels.Gaussian(d, 2.0), oracle.GaussianRFT), (kernels.Laplacian(d, 3.0), oracle.LaplacianRFT)):
        Z, T = K.features(X, s, context=lib.Context(21))
        want = O(d, s, K._sigma, oracle.Context(21)).apply(X.cpu().numpy(), "rowwise")
        assert rel_fro(Z.cpu().numpy(), want) < 1e-10                      # same features as the reference path
        G = K.gram(X)
        approx = Z @ Z.t()                                                 # E[z(x) z(y)] = k(x, y)
        assert float((approx - G).abs().max()) < 0.12
    assert isinstance(kernels.kernel("linear", d).rft(8, context=lib.Context(1)), sk.JLT)
    assert isinstance(kernels.kernel("linear", d).rft(8, "hash", context=lib.Context(1)), sk.CWT)
    # LIBSVM text into device buffers, dense and sparse, then through a sketch
    path = tmp_path / "toy.libsvm"
    path.write_text("1 1:0.5 3:2.0\n-1 2:1.5\n1 1:-1.0 4:4.0\n")
    Xd, y = io.read_libsvm(str(path))
    assert Xd.shape == (3, 4) and y.tolist() == [1.0, -1.0, 1.0] and float(Xd[2, 3]) == 4.0
    Xs, _ = io.read_libsvm(str(path), sparse=True)
    C = sk.CWT(4, 3, context=lib.Context(2))
    dense = C.apply(Xd.t().contiguous(), None, "columnwise")
    sp = C.apply(Xs, torch.empty(3, 3, dtype=torch.float64, device="cuda").t(), "columnwise")
    assert rel_fro(sp.cpu().numpy(), dense.cpu().numpy()) < 1e-14
