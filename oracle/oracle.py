"""CPU oracle for the sketch-apply hot path (Python face of skylark_oracle.c).

TEST INFRASTRUCTURE ONLY: imported by tests/, __graft_entry__.smoke() and the
cpu_baseline / --impl reference legs of bench.py.  The product package
(libskylark_b200) never imports it.

PARITY UNPINNED -- see the header of skylark_oracle.c: the reference's tests hold
no golden vector for this path and the reference cannot be built here, so this
restatement is pinned only by Random123's published Threefry2x64-20 KATs and by
structural properties (apply == explicit S*A, serialize/replay, counter rules).

The classes mirror the reference data classes one to one:
  Context          base/context.hpp:19-183
  CWT / MMT / WZT  sketch/{CWT,MMT,WZT}_data.hpp + hash_transform_data.hpp:64-76
  JLT / CT         sketch/{JLT,CT}_data.hpp + random_dense_transform_data.hpp:66-73
  GaussianRFT / LaplacianRFT   sketch/RFT_data.hpp:71-89,100-175
  FJLT             sketch/FJLT_data.hpp:79-86 + RFUT_data.hpp:27-50
`apply(A, dim)` takes a 2-D numpy array in mathematical terms: columnwise gives
S*A (s x w), rowwise gives A*S^T (h x s); layouts are handled internally the way
Elemental would (column-major).
"""
import ctypes
import math
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "libskylark_oracle.so")
_SRC = os.path.join(_HERE, "skylark_oracle.c")

BOOST_PI = 3.14159265358979323846
BOOST_160 = 160
BOOST_155 = 155


def build(force=False):
    """Compile the C restatement (gcc, no reference sources involved)."""
    if (not force and os.path.exists(_SO)
            and os.path.getmtime(_SO) >= os.path.getmtime(_SRC)):
        return _SO
    cmd = ["gcc", "-O2", "-fopenmp", "-ffp-contract=off", "-fPIC", "-shared",
           "-o", _SO, _SRC, "-lquadmath", "-lm"]
    subprocess.check_call(cmd)
    return _SO


_lib = None


def lib():
    global _lib
    if _lib is None:
        build()
        _lib = ctypes.CDLL(_SO)
        _lib.sko_random_int.restype = ctypes.c_int
        _lib.sko_num_threads.restype = ctypes.c_int
        _lib.sko_hash_apply_csc_sparse.restype = ctypes.c_size_t
    return _lib


def _p(a):
    return a.ctypes.data_as(ctypes.c_void_p)


_u64 = ctypes.c_uint64
_sz = ctypes.c_size_t
_dbl = ctypes.c_double
_int = ctypes.c_int


def num_threads():
    return lib().sko_num_threads()


def set_libm(correctly_rounded):
    """True (default): log / cos / tan / pow inside the variates are correctly rounded, like the
    libm of the reference's era (see skylark_oracle.c); False: this box's glibc, as the timed CPU
    baseline uses.  Returns the previous setting."""
    prev = bool(lib().sko_get_libm())
    lib().sko_set_libm(1 if correctly_rounded else 0)
    return prev


# ---------------------------------------------------------------------------
# RNG primitives
# ---------------------------------------------------------------------------
def threefry2x64(rounds, ctr, key):
    c = (ctypes.c_uint64 * 2)(*ctr)
    k = (ctypes.c_uint64 * 2)(*key)
    o = (ctypes.c_uint64 * 2)()
    lib().sko_threefry2x64(_int(rounds), c, k, o)
    return int(o[0]), int(o[1])


def raw_draws(seed, base, count):
    """First and second 64-bit draw of each sample's MicroURNG (X1, then X0)."""
    a = np.empty(count, dtype=np.uint64)
    b = np.empty(count, dtype=np.uint64)
    lib().sko_raw_draws(_int(seed), _u64(base), _sz(count), _p(a), _p(b))
    return a, b


def uniform_int(seed, base, count, hi):
    out = np.empty(count, dtype=np.uint64)
    lib().sko_uniform_int(_int(seed), _u64(base), _sz(count), _u64(hi), _p(out))
    return out


def rademacher(seed, base, count, order=BOOST_160):
    out = np.empty(count, dtype=np.float64)
    lib().sko_rademacher(_int(seed), _u64(base), _sz(count), _int(order), _p(out))
    return out


def normal(seed, base, count):
    out = np.empty(count, dtype=np.float64)
    lib().sko_normal(_int(seed), _u64(base), _sz(count), _p(out))
    return out


def cauchy(seed, base, count):
    out = np.empty(count, dtype=np.float64)
    lib().sko_cauchy(_int(seed), _u64(base), _sz(count), _p(out))
    return out


def exponential(seed, base, count):
    out = np.empty(count, dtype=np.float64)
    lib().sko_exponential(_int(seed), _u64(base), _sz(count), _p(out))
    return out


def uniform_real(seed, base, count, a, b):
    out = np.empty(count, dtype=np.float64)
    lib().sko_uniform_real(_int(seed), _u64(base), _sz(count), _dbl(a), _dbl(b),
                           _p(out))
    return out


def realize_dense(seed, base, dist, scale, S, i, j, height, width,
                  col_stride=1, row_stride=1, dtype=np.float64):
    """dense_transform_data_t::realize_matrix_view; returns height x width (F order)."""
    if dtype == np.float32:
        assert col_stride == 1 and row_stride == 1
        R = np.empty((height, width), dtype=np.float32, order="F")
        lib().sko_realize_dense_f32(_int(seed), _u64(base), _int(dist), _dbl(scale),
                                    _sz(S), _sz(i), _sz(j), _sz(height), _sz(width),
                                    _p(R))
        return R
    R = np.empty((height, width), dtype=np.float64, order="F")
    lib().sko_realize_dense(_int(seed), _u64(base), _int(dist), _dbl(scale), _sz(S),
                            _sz(i), _sz(j), _sz(height), _sz(width),
                            _sz(col_stride), _sz(row_stride), _p(R))
    return R


def dct2(x, axis=0):
    """FFTW REDFT10 (utility/fft/fftw_futs.h:117-126) == scipy dct type 2, norm=None."""
    import scipy.fft
    return scipy.fft.dct(x, type=2, norm=None, axis=axis)


def dct2_naive(x):
    x = np.ascontiguousarray(x, dtype=np.float64)
    y = np.empty_like(x)
    lib().sko_dct2_naive(_p(x), _sz(x.shape[0]), _p(y))
    return y


def wht(x):
    """Unnormalised natural-order Walsh-Hadamard transform along axis 0."""
    x = np.array(x, dtype=np.float64, order="F", copy=True)
    if x.ndim == 1:
        lib().sko_wht_inplace(_p(x), _sz(x.shape[0]))
        return x
    for j in range(x.shape[1]):
        col = np.ascontiguousarray(x[:, j])
        lib().sko_wht_inplace(_p(col), _sz(col.shape[0]))
        x[:, j] = col
    return x


# ---------------------------------------------------------------------------
# Context and transforms
# ---------------------------------------------------------------------------
class Context:
    """base/context.hpp:19-183 -- {seed, counter}; arrays reserve counter ranges."""

    def __init__(self, seed, counter=0):
        self.seed = int(seed)
        self.counter = int(counter)

    def allocate(self, size):
        base = self.counter
        self.counter += int(size)
        return base

    def random_int(self):
        base = self.allocate(1)
        return lib().sko_random_int(_int(self.seed), _u64(base))

    def copy(self):
        return Context(self.seed, self.counter)

    def to_dict(self):
        return {"skylark_object_type": "context", "skylark_version": "0.20",
                "seed": self.seed, "counter": self.counter}


def _as_f(A, dtype=np.float64):
    return np.asfortranarray(A, dtype=dtype)


class _Transform:
    sketch_type = None

    def __init__(self, N, S, context):
        self.N, self.S = int(N), int(S)
        self.creation_context = context.copy()

    def to_dict(self):
        d = {"skylark_object_type": "sketch", "sketch_type": self.sketch_type,
             "skylark_version": "0.20", "N": self.N, "S": self.S,
             "creation_context": self.creation_context.to_dict()}
        d.update(self._extra())
        return d

    def _extra(self):
        return {}


class _Hash(_Transform):
    def _build_idx(self, ctx):
        # hash_transform_data.hpp:64-76
        self.idx_base = ctx.allocate(self.N)
        self.row_idx = uniform_int(ctx.seed, self.idx_base, self.N, self.S - 1)
        self.val_base = ctx.allocate(self.N)

    def apply(self, A, dim="columnwise"):
        A = _as_f(A)
        L = lib()
        if dim == "columnwise":
            N, w = A.shape
            assert N == self.N
            SA = np.empty((self.S, w), order="F")
            L.sko_hash_apply_columnwise(_p(A), _sz(N), _sz(w), _sz(N),
                                        _p(self.row_idx), _p(self.row_value),
                                        _p(SA), _sz(self.S), _sz(self.S))
        else:
            h, N = A.shape
            assert N == self.N
            SA = np.empty((h, self.S), order="F")
            L.sko_hash_apply_rowwise(_p(A), _sz(h), _sz(N), _sz(h),
                                     _p(self.row_idx), _p(self.row_value),
                                     _p(SA), _sz(self.S), _sz(h))
        return SA

    def apply_csc(self, indptr, indices, values, shape, dim="columnwise"):
        indptr = np.ascontiguousarray(indptr, dtype=np.int32)
        indices = np.ascontiguousarray(indices, dtype=np.int32)
        values = np.ascontiguousarray(values, dtype=np.float64)
        L = lib()
        h, w = shape
        if dim == "columnwise":
            assert h == self.N
            SA = np.empty((self.S, w), order="F")
            L.sko_hash_apply_csc_columnwise(_p(indptr), _p(indices), _p(values),
                                            _sz(w), _p(self.row_idx),
                                            _p(self.row_value), _p(SA),
                                            _sz(self.S), _sz(self.S))
        else:
            assert w == self.N
            SA = np.empty((h, self.S), order="F")
            L.sko_hash_apply_csc_rowwise(_p(indptr), _p(indices), _p(values),
                                         _sz(h), _sz(w), _p(self.row_idx),
                                         _p(self.row_value), _p(SA), _sz(self.S),
                                         _sz(h))
        return SA

    def apply_csc_sparse(self, A_csc, dim="columnwise"):
        """sparse -> sparse (hash_transform_local_sparse.hpp:88-223); returns scipy CSC."""
        import scipy.sparse
        A = scipy.sparse.csc_matrix(A_csc)
        indptr = np.ascontiguousarray(A.indptr, dtype=np.int32)
        indices = np.ascontiguousarray(A.indices, dtype=np.int32)
        values = np.ascontiguousarray(A.data, dtype=np.float64)
        h, w = A.shape
        rowwise = 0 if dim == "columnwise" else 1
        ncols_out = w if not rowwise else self.S
        oi = np.zeros(ncols_out + 1, dtype=np.int32)
        ox = np.zeros(max(1, len(indices)), dtype=np.int32)
        ov = np.zeros(max(1, len(indices)), dtype=np.float64)
        nnz = lib().sko_hash_apply_csc_sparse(_p(indptr), _p(indices), _p(values), _sz(h), _sz(w),
                                              _int(rowwise), _p(self.row_idx), _p(self.row_value),
                                              _sz(self.S), _p(oi), _p(ox), _p(ov))
        shape = (self.S, w) if not rowwise else (h, self.S)
        return scipy.sparse.csc_matrix((ov[:nnz], ox[:nnz], oi), shape=shape)

    def explicit(self):
        """The S x N matrix Pi (one nonzero per column)."""
        P = np.zeros((self.S, self.N))
        P[self.row_idx.astype(np.int64), np.arange(self.N)] = self.row_value
        return P


class CWT(_Hash):
    sketch_type = "CWT"

    def __init__(self, N, S, context, order=BOOST_160):
        super().__init__(N, S, context)
        self._build_idx(context)
        self.row_value = rademacher(context.seed, self.val_base, self.N, order)


class MMT(_Hash):
    sketch_type = "MMT"

    def __init__(self, N, S, context):
        super().__init__(N, S, context)
        self._build_idx(context)
        self.row_value = cauchy(context.seed, self.val_base, self.N)


class WZT(_Hash):
    sketch_type = "WZT"

    def __init__(self, N, S, p, context, order=BOOST_160):
        super().__init__(N, S, context)
        if p < 1 or 2 < p:
            raise ValueError("WZT parameter p has to be in (1, 2)")
        self.p = float(p)
        self._build_idx(context)
        self.pm_base = context.allocate(self.N)
        out = np.empty(self.N)
        lib().sko_wzt_values(_int(context.seed), _u64(self.val_base),
                             _u64(self.pm_base), _sz(self.N), _dbl(self.p),
                             _int(order), _p(out))
        self.row_value = out

    def _extra(self):
        return {"P": self.p}


class _Dense(_Transform):
    dist = 0

    def _build(self, ctx, scale):
        self.scale = scale
        self.seed = ctx.seed
        self.base = ctx.allocate(self.N * self.S)  # lazy, N*S counters

    def realize(self, dtype=np.float64):
        return realize_dense(self.seed, self.base, self.dist, self.scale, self.S,
                             0, 0, self.S, self.N, dtype=dtype)

    def apply(self, A, dim="columnwise", dtype=np.float64):
        # dense_transform_Elemental_local.hpp:79-111
        R = self.realize(dtype)
        A = np.asarray(A, dtype=dtype)
        if dim == "columnwise":
            return R @ A
        return A @ R.T


class JLT(_Dense):
    sketch_type = "JLT"

    def __init__(self, N, S, context):
        super().__init__(N, S, context)
        self._build(context, math.sqrt(1.0 / float(S)))


class CT(_Dense):
    sketch_type = "CT"
    dist = 1

    def __init__(self, N, S, C, context):
        super().__init__(N, S, context)
        self.C = float(C)
        self._build(context, self.C / float(S))

    def _extra(self):
        return {"C": self.C}


class _RFT(_Transform):
    dist = 0

    def __init__(self, N, S, sigma, context):
        super().__init__(N, S, context)
        self.sigma = float(sigma)
        self.inscale = 1.0 / self.sigma
        self.outscale = math.sqrt(2.0 / S)
        self.seed = context.seed
        self.base = context.allocate(self.N * self.S)
        self.shift_base = context.allocate(self.S)
        pi = math.pi  # boost::math::constants::pi<double>()
        self.shifts = uniform_real(self.seed, self.shift_base, self.S, 0.0, 2 * pi)
        self.scales = np.ones(self.S)

    def _extra(self):
        return {"sigma": self.sigma}

    def realize(self, dtype=np.float64):
        return realize_dense(self.seed, self.base, self.dist, self.inscale, self.S,
                             0, 0, self.S, self.N, dtype=dtype)

    def apply(self, A, dim="columnwise", dtype=np.float64):
        R = self.realize(dtype)
        A = np.asarray(A, dtype=dtype)
        SA = np.asfortranarray(R @ A if dim == "columnwise" else A @ R.T)
        fn = lib().sko_rft_epilogue if dtype == np.float64 else lib().sko_rft_epilogue_f32
        fn(_p(SA), _sz(SA.shape[0]), _sz(SA.shape[1]), _sz(SA.shape[0]),
           _int(0 if dim == "columnwise" else 1), _p(self.scales), _p(self.shifts),
           _dbl(self.outscale))
        return SA


class GaussianRFT(_RFT):
    sketch_type = "GaussianRFT"
    dist = 0


class LaplacianRFT(_RFT):
    sketch_type = "LaplacianRFT"
    dist = 1


class FJLT(_Transform):
    """FJLT_data_t + RFUT_data_t: D uses [c, c+N) WITHOUT advancing the context
    (RFUT_data.hpp:27-31 discards build()'s return), samples use [c, c+S), the
    context advances by S only (FJLT_data.hpp:79-86)."""
    sketch_type = "FJLT"

    def __init__(self, N, S, context, order=BOOST_160):
        super().__init__(N, S, context)
        c = context.counter
        self.D = rademacher(context.seed, c, self.N, order)
        base = context.allocate(self.S)
        assert base == c
        self.samples = uniform_int(context.seed, base, self.S, self.N - 1)

    def apply(self, A, dim="columnwise", fut="dct"):
        # FJLT_Elemental.hpp:144-186; rowwise = transpose, columnwise, transpose.
        A = np.asarray(A, dtype=np.float64)
        if dim != "columnwise":
            return np.ascontiguousarray(self.apply(A.T, "columnwise", fut).T)
        N, w = A.shape
        assert N == self.N
        if fut == "dct":
            scale = 1.0 / math.sqrt(2.0 * N)      # fftw_futs.h:44-46, ScaleVal 2
            TA = dct2((scale * self.D)[:, None] * A, axis=0)
        else:
            scale = 1.0 / math.sqrt(float(N))     # FUT.hpp WHT_t::scale
            TA = wht((scale * self.D)[:, None] * A)
        out = math.sqrt(float(N) / float(self.S))
        return out * TA[self.samples.astype(np.int64), :]


class UST(_Transform):
    """sketch/UST_data.hpp:81-102 (samples) and sketch/UST_Elemental.hpp (apply): uniform row /
    column sampling, with replacement (S draws of uniform_int(0, N-1)) or without (inside-out
    shuffle over N single draws uniform_int<int>(0, i), first S entries)."""
    sketch_type = "UST"

    def __init__(self, N, S, replace, context):
        super().__init__(N, S, context)
        self.replace = bool(replace)
        seed = context.seed
        if self.replace:
            self.base = context.allocate(S)
            self.samples = uniform_int(seed, self.base, S, N - 1).astype(np.int64)
        else:
            self.base = context.allocate(N)
            self.samples = np.empty(S, dtype=np.int64)
            lib().sko_ust_noreplace(_int(seed), _u64(self.base), _sz(N), _sz(S), _p(self.samples))

    def _extra(self):
        return {"replace": "true" if self.replace else "false"}

    def apply(self, A, dim="columnwise"):
        A = np.asarray(A)
        return A[self.samples, :] if dim in (0, "columnwise") else A[:, self.samples]


class FastGaussianRFT(_Transform):
    """Fastfood: sketch/FRFT_data.hpp:91-120 (tables: shifts, B, G, P in that order; P reduced modulo
    j+1 for the Fisher-Yates swaps), FastGaussianRFT_data_t::build (:190-196, Sm = 1/(sigma sqrt(N))),
    and the columnwise apply of sketch/FRFT_Elemental.hpp:70-135 with the FFTW DCT (block size N)."""
    sketch_type = "FastGaussianRFT"

    def __init__(self, N, S, sigma, context, order=BOOST_160):
        super().__init__(N, S, context)
        self.sigma = float(sigma)
        seed = context.seed
        nb = 1 + (S - 1) // N
        pi = 3.141592653589793238462643383279502884
        self.shifts = uniform_real(seed, context.allocate(S), S, 0.0, 2 * pi)
        self.B = rademacher(seed, context.allocate(nb * N), nb * N, order)
        self.G = normal(seed, context.allocate(nb * N), nb * N)
        praw, _ = raw_draws(seed, context.allocate(nb * (N - 1)), nb * (N - 1))
        self.P = np.array([int(praw[i * (N - 1) + l]) % (N - l) for i in range(nb) for l in range(N - 1)],
                          dtype=np.int64)
        self.nb, self.scale = nb, math.sqrt(2.0 / S)
        self.Sm = np.full(nb * N, 1.0 / (self.sigma * math.sqrt(N)))

    def apply(self, A, dim="columnwise"):
        import scipy.fft
        A = np.asarray(A, dtype=np.float64)
        X = A if dim in (0, "columnwise") else A.T
        N, S = self.N, self.S
        scal = math.sqrt(N) * (1.0 / math.sqrt(2.0 * N))
        out = np.empty((S, X.shape[1]))
        for c in range(X.shape[1]):
            ac = X[:, c]
            for i in range(self.nb):
                s0, e0 = i * N, min((i + 1) * N, S)
                w = self.B[i * N:(i + 1) * N] * ac
                w = scipy.fft.dct(w, type=2)
                for l in range(N - 1):
                    i1, i2 = N - 1 - l, self.P[i * (N - 1) + l]
                    w[i1], w[i2] = w[i2], w[i1]
                w = (scal * self.G[i * N:(i + 1) * N]) * w
                w = scipy.fft.dct(w, type=2)
                w = (scal * self.Sm[i * N:(i + 1) * N]) * w
                out[s0:e0, c] = self.scale * np.cos(w[:e0 - s0] + self.shifts[s0:e0])
        return out if dim in (0, "columnwise") else out.T


class PPT(_Transform):
    """TensorSketch: sketch/PPT_data.hpp:83-99 (q CWTs, then q hash positions, then q signs) and the
    apply of sketch/PPT_Elemental.hpp:140-185 (r2c FFT of every CountSketch, product, c2r / S)."""
    sketch_type = "PPT"

    def __init__(self, N, S, q, c, gamma, context, order=BOOST_160):
        super().__init__(N, S, context)
        self.q, self.c, self.gamma = int(q), float(c), float(gamma)
        self.cwts = [CWT(N, S, context, order) for _ in range(self.q)]
        seed = context.seed
        self.hash_idx = uniform_int(seed, context.allocate(self.q), self.q, S - 1).astype(np.int64)
        self.hash_val = rademacher(seed, context.allocate(self.q), self.q, order)

    def apply(self, A, dim="columnwise"):
        A = np.asarray(A, dtype=np.float64)
        X = A if dim in (0, "columnwise") else A.T
        P = np.ones((self.S // 2 + 1, X.shape[1]), dtype=np.complex128)
        for i, cw in enumerate(self.cwts):
            W = cw.apply(X, "columnwise") * math.sqrt(self.gamma)
            W[self.hash_idx[i], :] += math.sqrt(self.c) * self.hash_val[i]
            P *= np.fft.rfft(W, axis=0)
        out = np.fft.irfft(P, n=self.S, axis=0)
        return out if dim in (0, "columnwise") else out.T
