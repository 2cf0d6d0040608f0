"""Kernels whose `rft` builds the random-feature transform of this package -- the reference's
ml::kernels create_rft (ml/kernels.hpp:53-66, gaussian_t :190-300, laplacian_t, linear_t) and
python-skylark/skylark/ml/kernels.py -- plus `gram` and `features` on the device.

`k.rft(s)` returns a sketch transform (GaussianRFT / LaplacianRFT / JLT ...) created from the given
`lib.Context`, so features are the reference's for the same seed and counter; `k.features(X, s)`
applies it rowwise to a device tensor (one data point per row) and leaves Z on the device, which is
the input format of the solvers one level up (ml/krr.hpp:199-451 forms Z^T Z from it)."""
import math

from .. import errors, sketch


def _torch():
    import torch
    return torch


class Kernel(object):
    def __init__(self, d):
        self._d = int(d)

    def get_dim(self):
        return self._d

    def rft(self, s, subtype=None, defouttype=None, **kwargs):
        raise NotImplementedError

    def features(self, X, s, subtype=None, context=None, out=None):
        """Z = rft(s) / X : one row of `s` random features per data point (rows of X), on X's device."""
        T = self.rft(s, subtype, context=context)
        return T.apply(X, out, "rowwise"), T

    def gram(self, X, K=None, dirX="rows", dirY="rows", Y=None):
        """Dense Gram matrix K[i, j] = k(x_i, y_j) on the device (kernels.py:60-88).  dir = "rows": one
        data point per row; "columns": per column."""
        torch = _torch()
        if Y is None:
            Y, dirY = X, dirX
        Xr = X if dirX in (1, "rows") else X.t()
        Yr = Y if dirY in (1, "rows") else Y.t()
        if Xr.shape[1] != self._d or Yr.shape[1] != self._d:
            raise errors.DimensionMistmatchError("data dimension differs from the kernel's")
        G = self._gram(torch, Xr, Yr)
        if K is not None:
            K.copy_(G)
            return K
        return G


class Linear(Kernel):
    """k(x, y) = x^T y; random features = JLT / FJLT / CWT (kernels.py:94-122)."""

    def _gram(self, torch, X, Y):
        return X @ Y.t()

    def rft(self, s, subtype=None, defouttype=None, **kwargs):
        if subtype is None:
            return sketch.JLT(self._d, s, defouttype, **kwargs)
        if subtype == "fast":
            return sketch.FJLT(self._d, s, defouttype, **kwargs)
        if subtype == "hash":
            return sketch.CWT(self._d, s, defouttype, **kwargs)
        raise ValueError("invalid subtype supplied")


class Gaussian(Kernel):
    """k(x, y) = exp(-||x - y||^2 / (2 sigma^2)) (ml/kernels.hpp gaussian_t); rft = GaussianRFT."""

    def __init__(self, d, sigma):
        super().__init__(d)
        self._sigma = float(sigma)

    def _gram(self, torch, X, Y):
        d2 = (X * X).sum(1)[:, None] + (Y * Y).sum(1)[None, :] - 2.0 * (X @ Y.t())
        return torch.exp(-d2.clamp_min(0.0) / (2.0 * self._sigma ** 2))

    def rft(self, s, subtype=None, defouttype=None, **kwargs):
        if subtype == "fast":
            return sketch.FastGaussianRFT(self._d, s, self._sigma, defouttype, **kwargs)
        return sketch.GaussianRFT(self._d, s, self._sigma, defouttype, **kwargs)


class Laplacian(Kernel):
    """k(x, y) = exp(-||x - y||_1 / sigma) (ml/kernels.hpp laplacian_t); rft = LaplacianRFT."""

    def __init__(self, d, sigma):
        super().__init__(d)
        self._sigma = float(sigma)

    def _gram(self, torch, X, Y):
        return torch.exp(-torch.cdist(X, Y, p=1) / self._sigma)

    def rft(self, s, subtype=None, defouttype=None, **kwargs):
        if subtype == "quasirandom":
            return sketch.LaplacianQRFT(self._d, s, self._sigma, defouttype, **kwargs)
        return sketch.LaplacianRFT(self._d, s, self._sigma, defouttype, **kwargs)


class Polynomial(Kernel):
    """k(x, y) = (gamma x^T y + c)^q (ml/kernels.hpp polynomial_t); rft = TensorSketch (kernels.py:163-199)."""

    def __init__(self, d, q=3, c=0, gamma=1):
        if c < 0:
            raise ValueError("kernel paramter must be >= 0")
        if type(q) is not int:
            raise errors.InvalidParamterError("exponent must be integer")
        super().__init__(d)
        self._q, self._c, self._gamma = q, c, gamma

    def _gram(self, torch, X, Y):
        return (self._gamma * (X @ Y.t()) + self._c) ** self._q

    def rft(self, s, subtype=None, defouttype=None, **kwargs):
        return sketch.PPT(self._d, s, self._q, self._c, self._gamma, defouttype, **kwargs)


def kernel(kerneltype, *args):
    """Factory (kernels.py:13-38)."""
    table = {"linear": Linear, "gaussian": Gaussian, "laplacian": Laplacian, "polynomial": Polynomial}
    if not isinstance(kerneltype, str):
        raise ValueError("kerneltype must be a string")
    try:
        return table[kerneltype.lower()](*args)
    except KeyError:
        raise ValueError("kerneltype not recognized")
