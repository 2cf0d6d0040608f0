"""Randomized NLA on the device, driven by the sketch path -- the callers SURVEY section 8(f)#2 names:
nla::ApproximateSVD (nla/svd.hpp:222-320, python-skylark/skylark/nla/nla.py:27-73) and the two
sketched least-squares solvers (nla/least_squares.hpp:42-180 ApproximateLeastSquares = sketch and
solve with an FJLT; :237-262 FasterLeastSquares = Blendenpik: FJLT sketch -> QR -> LSQR
preconditioned by R, algorithms/regression/accelerated_linearl2_regression_solver_Elemental.hpp).

Operands are torch CUDA tensors (float64).  The random range finder / the sketches run through this
package's kernels (JLT, FJLT) from a `lib.Context`, so the random part of every result is the
reference's for the same seed and counter; the dense factorizations (QR, SVD, triangular solves)
are torch.linalg, i.e. cuSOLVER / cuBLAS -- library calls outside the sketch hot path, as
Elemental's LAPACK calls are in the reference."""
from . import errors, lib, sketch

__all__ = ["SVDParams", "FasterLeastSquaresParams", "approximate_svd", "approximate_symmetric_svd",
           "approximate_least_squares", "faster_least_squares"]


class SVDParams(object):
    """nla.py:8-18 (Python defaults: two power iterations)."""

    def __init__(self, oversampling_ratio=2, oversampling_additive=0, num_iterations=2, skip_qr=False):
        self.oversampling_ratio = oversampling_ratio
        self.oversampling_additive = oversampling_additive
        self.num_iterations = num_iterations
        self.skip_qr = skip_qr


class FasterLeastSquaresParams(object):
    def __init__(self, tolerance=1e-14, iter_lim=100):
        self.tolerance = tolerance
        self.iter_lim = iter_lim


def _torch():
    import torch
    return torch


def _check(A):
    torch = _torch()
    if not (isinstance(A, torch.Tensor) and A.is_cuda and A.dim() == 2 and A.dtype == torch.float64):
        raise errors.UnsupportedError("nla: operands must be 2-D float64 CUDA tensors")


def _orth(M):
    return _torch().linalg.qr(M, mode="reduced")[0]


def _power_iteration(A, Q, iters, ortho):
    """PowerIteration(ADJOINT, NORMAL, NORMAL, A, V=Q, U), nla/svd.hpp:104-114: returns (Q, A^T Q)."""
    if ortho:
        Q = _orth(Q)
    for _ in range(iters):
        U = A.t() @ Q
        if ortho:
            U = _orth(U)
        Q = A @ U
        if ortho:
            Q = _orth(Q)
    return Q, A.t() @ Q


def approximate_svd(A, k=10, params=None, context=None):
    """A ~= U diag(S) V^T, rank k (nla/svd.hpp:222-320).  Returns (U, S, V) on A's device."""
    torch = _torch()
    _check(A)
    p = params or SVDParams()
    m, n = A.shape
    if k > min(m, n):
        raise errors.InvalidParamterError("Incompatible matrix dimensions (%d) and target rank (%d)" % (min(m, n), k))
    if m < n:                      # svd.hpp:296-318 is the m >= n branch applied to A^T
        V, S, U = approximate_svd(A.t(), k, p, context)
        return U, S, V
    kk = max(k, min(n, p.oversampling_ratio * k + p.oversampling_additive))
    Omega = sketch.JLT(n, kk, context=context)
    Q = Omega.apply(A, None, "rowwise")                       # m x kk = A * Omega^T  (svd.hpp:256-257)
    Q, Vt = _power_iteration(A, Q, p.num_iterations, not p.skip_qr)
    if p.skip_qr:
        Q = _orth(Q)
        Vt = A.t() @ Q
    Vs, S, Bh = torch.linalg.svd(Vt, full_matrices=False)     # V <- left factors, B = right factors
    U = Q @ Bh.t()[:, :k]
    return U, S[:k], Vs[:, :k]


def approximate_symmetric_svd(A, k=10, params=None, context=None):
    """A = A^T ~= V diag(S) V^T (nla/svd.hpp:322-400: range finder + Rayleigh-Ritz)."""
    torch = _torch()
    _check(A)
    p = params or SVDParams()
    n = A.shape[0]
    kk = max(k, min(n, p.oversampling_ratio * k + p.oversampling_additive))
    Q = sketch.JLT(n, kk, context=context).apply(A, None, "rowwise")
    Q = _orth(Q)
    for _ in range(p.num_iterations):
        Q = _orth(A @ Q)
    T = Q.t() @ (A @ Q)
    w, Z = torch.linalg.eigh(0.5 * (T + T.t()))
    order = torch.argsort(w.abs(), descending=True)[:k]
    return Q @ Z[:, order], w[order]


def approximate_least_squares(A, B, sketch_size=-1, context=None):
    """argmin ||A X - B|| by sketch-and-solve with an FJLT of 4 * width(A) rows
    (nla/least_squares.hpp:42-66)."""
    torch = _torch()
    _check(A); _check(B)
    m, n = A.shape
    s = 4 * n if sketch_size == -1 else int(sketch_size)
    F = sketch.FJLT(m, s, context=context)
    SA = F.apply(A, None, "columnwise")
    SB = F.apply(B, None, "columnwise")
    return torch.linalg.lstsq(SA, SB).solution


def faster_least_squares(A, B, params=None, context=None):
    """argmin ||A X - B|| to full accuracy: R from the QR of an FJLT sketch of A preconditions LSQR
    (Blendenpik; nla/least_squares.hpp:237-262).  Returns X."""
    torch = _torch()
    _check(A); _check(B)
    p = params or FasterLeastSquaresParams()
    m, n = A.shape
    if m < n:
        raise errors.UnsupportedError("faster_least_squares needs a tall matrix")
    s = min(m, 4 * n)
    SA = sketch.FJLT(m, s, context=context).apply(A, None, "columnwise")
    R = torch.linalg.qr(SA, mode="r")[1]

    def op(X):          # A R^-1 X
        return A @ torch.linalg.solve_triangular(R, X, upper=True)

    def opt(Y):         # R^-T A^T Y
        return torch.linalg.solve_triangular(R.t(), A.t() @ Y, upper=False)

    # LSQR (Paige & Saunders) on the preconditioned operator, all right-hand sides at once
    X = torch.zeros(n, B.shape[1], dtype=A.dtype, device=A.device)
    u = B.clone()
    beta = u.norm(dim=0)
    u = u / beta.clamp_min(1e-300)
    v = opt(u)
    alpha = v.norm(dim=0)
    v = v / alpha.clamp_min(1e-300)
    w = v.clone()
    phibar, rhobar = beta.clone(), alpha.clone()
    bnorm = beta.clone()
    for _ in range(p.iter_lim):
        u = op(v) - alpha * u
        beta = u.norm(dim=0)
        u = u / beta.clamp_min(1e-300)
        v = opt(u) - beta * v
        alpha = v.norm(dim=0)
        v = v / alpha.clamp_min(1e-300)
        rho = torch.sqrt(rhobar * rhobar + beta * beta)
        c, sn = rhobar / rho, beta / rho
        theta = sn * alpha
        rhobar = -c * alpha
        phi = c * phibar
        phibar = sn * phibar
        X = X + (phi / rho) * w
        w = v - (theta / rho) * w
        if bool(((phibar * alpha * c.abs()) <= p.tolerance * bnorm.clamp_min(1e-300)).all()):
            break
    return torch.linalg.solve_triangular(R, X, upper=True)
