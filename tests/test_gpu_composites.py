"""Fastfood (FastGaussianRFT, sketch/FRFT_data.hpp + FRFT_Elemental.hpp) and TensorSketch (PPT,
sketch/PPT_data.hpp + PPT_Elemental.hpp): composites whose random tables come from the counter
stream through the C ABI and whose CountSketches run on the hash kernels, against the oracle."""
import numpy as np
import pytest

from util import rel_fro

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("shape", [(64, 200, 9), (100, 100, 5), (37, 30, 4)])     # N, S, columns; several blocks, ragged last block
def test_fastfood_matches_the_oracle(sk, oracle, shape):
    import torch
    from libskylark_b200 import lib
    n, s, m = shape
    c, oc = lib.Context(77, 9), oracle.Context(77, 9)
    F, O = sk.FastGaussianRFT(n, s, 2.5, context=c), oracle.FastGaussianRFT(n, s, 2.5, oc)
    assert c.counter == oc.counter                       # S + 2 nb N + nb (N - 1) counters (FRFT_data.hpp:91-114)
    assert sk.Fastfood is sk.FastGaussianRFT
    X = np.random.default_rng(1).uniform(0, 1, (n, m))
    want = O.apply(X, "columnwise")
    assert rel_fro(F.apply(X, None, "columnwise"), want) < 1e-12
    Xd = torch.from_numpy(np.ascontiguousarray(X.T)).cuda()                       # points x features, rowwise
    got = F.apply(Xd, None, "rowwise")
    assert got.is_cuda and rel_fro(got.cpu().numpy(), want.T) < 1e-12
    # the features approximate the Gaussian kernel (sigma = 2.5)
    if s >= 200:
        Z = got
        K = torch.exp(-torch.cdist(Xd, Xd) ** 2 / (2 * 2.5 ** 2))
        assert float((Z @ Z.t() - K).abs().max()) < 0.35


@pytest.mark.parametrize("params", [(3, 1.0, 1.0), (2, 0.0, 0.5), (4, 2.0, 1.5)])
def test_tensorsketch_matches_the_oracle(sk, oracle, params):
    import torch
    from libskylark_b200 import lib
    q, cc, gamma = params
    n, s, m = 300, 128, 11
    c, oc = lib.Context(5, 3), oracle.Context(5, 3)
    T, O = sk.PPT(n, s, q, cc, gamma, context=c), oracle.PPT(n, s, q, cc, gamma, oc)
    assert c.counter == oc.counter == 3 + q * 2 * n + 2 * q
    assert sk.TensorSketch is sk.PPT
    X = np.random.default_rng(2).standard_normal((n, m)) / np.sqrt(n)
    want = O.apply(X, "columnwise")
    assert rel_fro(T.apply(X, None, "columnwise"), want) < 1e-12
    Xd = torch.from_numpy(np.ascontiguousarray(X.T)).cuda()
    assert rel_fro(T.apply(Xd, None, "rowwise").cpu().numpy(), want.T) < 1e-12
    SA = torch.empty(s, m, dtype=torch.float64, device="cuda")
    T.apply(torch.from_numpy(X).cuda(), SA, 0)
    assert rel_fro(SA.cpu().numpy(), want) < 1e-12
