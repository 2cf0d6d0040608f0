"""Sketch transforms -- Python-3 counterpart of python-skylark/skylark/sketch.py.

Same classes, constructor arguments, `apply(A, SA, dim)` protocol, `*` and `/` operators,
serialisation and aliases as the reference (sketch.py:47-232, 238-620, 962-1003), for the
transforms on the hot path: JLT, CT, FJLT, CWT, MMT, WZT, GaussianRFT, LaplacianRFT.
Everything executes in libcskylark_b200.so on the GPU: `forceppy=True` (the reference's
pure-Python fallback) is refused, because this package has no CPU path.

Additions over the reference (all optional, existing calls keep their meaning):
  * `context=` keyword -- or a `Context` as third positional argument -- selects an
    explicit `lib.Context(seed, counter)` instead of the global one;
  * torch CUDA tensors (float64; float32 for JLT/CT/RFT) and `lib.DeviceCSC` as A / SA;
  * `apply_shard(A_local, SA, dim, index_offset)` for a row shard of the sketched
    dimension and `apply_distributed(...)` which also sums the partial sketches over a
    torch.distributed process group;
  * `RFT` alias for `GaussianRFT`; table accessors used by the parity tests.
"""
import ctypes
import json
import math
from ctypes import byref, c_char_p, c_double, c_void_p

import numpy

from . import errors, lib

_libc = lib._libc


def initialize(seed=-1):
    """examples/example_sketch.py:53 calls sketch.initialize(seed); forward to lib."""
    lib.initialize(seed)


def finalize():
    lib.finalize()


def _supported():
    raw = lib.lib.sl_supported_sketch_transforms().decode()
    return [eval(tok) for tok in raw.split()]     # sketch.py:21


SUPPORTED_SKETCH_TRANSFORMS = _supported()


def coalesce(A):
    """Adds entries of a lib.DeviceCSC that share (row, column), in storage order, and sorts the
    rows of every column (sl_sp_matrix_coalesce).  Returns a new lib.DeviceCSC."""
    out = lib.DeviceCSC.empty(A.shape, A.format, A.values.device)
    a, o = lib.DeviceCSCAdapter(A), lib.DeviceCSCAdapter(out)
    lib.set_stream_from_torch()
    ap = a.ptr()
    try:
        op = o.ptr()
        try:
            lib.callsl("sl_sp_matrix_coalesce", ap, op)
        finally:
            o.ptrcleaner()
    finally:
        a.ptrcleaner()
    return out


def merge_segments(vptr, indices, values, nseg, shape):
    """Adds `nseg` rank-major CSC pieces as they arrive from an all-to-all (sl_sp_merge_segments;
    see dist.exchange_sparse_partials(reorder=False)).  Returns a lib.DeviceCSC of `shape`."""
    out = lib.DeviceCSC.empty(shape, "csc", values.device)
    o = lib.DeviceCSCAdapter(out)
    lib.set_stream_from_torch()
    op = o.ptr()
    try:
        lib.callsl("sl_sp_merge_segments", c_void_p(vptr.data_ptr()), c_void_p(indices.data_ptr()),
                   c_void_p(values.data_ptr()), ctypes.c_int(int(nseg)), ctypes.c_int64(int(shape[0])),
                   ctypes.c_int64(int(shape[1])), ctypes.c_int64(int(indices.numel())), op)
    finally:
        o.ptrcleaner()
    return out


def deserialize_sketch(sketch_dict):
    """Rebuild a transform from `.serialize()` output (sketch.py:33-42)."""
    sketch_transform = c_void_p()
    lib.callsl("sl_deserialize_sketch_transform", json.dumps(sketch_dict), byref(sketch_transform))
    sketch_name = str(sketch_dict["sketch_type"])
    return _map_csketch_type_to_cfun[sketch_name](sketch_dict, sketch_transform)


def _split_context(defouttype, context):
    """`T(n, s, context)` (third positional) is accepted next to `T(n, s, defouttype)`."""
    if isinstance(defouttype, lib.Context):
        if context is not None:
            raise errors.ParameterMistmatchError("context given twice")
        return None, defouttype
    return defouttype, context


class _SketchTransform(object):
    """Base class of all sketch transforms (sketch.py:47-232)."""

    def __init__(self, ttype, n, s, defouttype=None, forceppy=False, sketch_transform=None,
                 context=None, params=()):
        defouttype, context = _split_context(defouttype, context)
        self._baseinit(ttype, n, s, defouttype, forceppy)
        if sketch_transform is None:
            sketch_transform = c_void_p()
            ctx = lib.ctxt_obj if context is None else context._obj
            lib.callsl("sl_create_sketch_transform", ctx, ttype, int(n), int(s),
                       byref(sketch_transform),
                       *[ctypes.c_int(int(p)) if isinstance(p, bool) else c_double(float(p)) for p in params])
        self._obj = sketch_transform

    def _baseinit(self, ttype, n, s, defouttype, forceppy):
        if defouttype is not None and defouttype not in lib.map_to_ctor:
            raise errors.UnsupportedError("Unsupported default output type (%s)" % defouttype)
        if forceppy:
            raise errors.UnsupportedError(
                "forceppy: this package has no pure-Python / CPU implementation")
        self._ttype = ttype
        self._n = n
        self._s = s
        self._defouttype = defouttype
        self._ppy = False

    def __del__(self):
        try:    # module globals may already be gone at interpreter shutdown
            obj = getattr(self, "_obj", None)
            if obj is not None and lib.lib is not None and getattr(obj, "value", obj):
                lib.lib.sl_free_sketch_transform(obj)
                self._obj = None
        except Exception:
            pass

    def serialize(self):
        """Dictionary from which `deserialize_sketch` rebuilds the sketch (sketch.py:94-113)."""
        json_data = c_char_p()
        lib.callsl("sl_serialize_sketch_transform", self._obj, byref(json_data))
        try:
            return json.loads(json_data.value.decode())
        finally:
            _libc.free(json_data)

    def __getstate__(self):
        d = self.__dict__.copy()
        if "_obj" in d:
            d["_obj"] = self.serialize()
        return d

    def __setstate__(self, d):
        self.__dict__ = d
        sketch_transform = c_void_p()
        lib.callsl("sl_deserialize_sketch_transform", json.dumps(d["_obj"]), byref(sketch_transform))
        self._obj = sketch_transform

    # -- apply ---------------------------------------------------------------
    @staticmethod
    def _normdim(dim):
        if dim == 0 or dim == "columnwise" or dim == "left":
            return 0
        if dim == 1 or dim == "rowwise" or dim == "right":
            return 1
        raise ValueError("Dimension must be either columnwise/rowwise or left/right or 0/1")

    def _prepare(self, A, SA, dim, local_n=None):
        A = lib.adapt(A)
        if SA is None:
            ctor = A.getctor() if self._defouttype is None else lib.map_to_ctor[self._defouttype]
            if dim == 0:
                SA = ctor(self._s, A.getdim(1), A)
            else:
                SA = ctor(A.getdim(0), self._s, A)
        SA = lib.adapt(SA)
        reqcomb = (self._ttype, A.ctype(), SA.ctype())
        if reqcomb not in SUPPORTED_SKETCH_TRANSFORMS:
            raise errors.UnsupportedError("Unsupported transform-input-output combination: "
                                          + str(reqcomb))
        incomp, cinvert = A.iscompatible(SA)
        if incomp is not None:
            raise errors.UnsupportedError("Input and output are incompatible: " + incomp)
        want_n = self._n if local_n is None else local_n
        if A.getdim(dim) != want_n:
            raise errors.DimensionMistmatchError("Sketched dimension is incorrect (input)")
        if SA.getdim(dim) != self._s:
            raise errors.DimensionMistmatchError("Sketched dimension is incorrect (output)")
        if A.getdim(1 - dim) != SA.getdim(1 - dim):
            raise errors.DimensionMistmatchError("Sketched dimension is incorrect (input != output)")
        return A, SA, cinvert

    def apply(self, A, SA=None, dim=0):
        """Apply the transform to A along `dim` and write SA (allocated when None).
        dim: 0 / "columnwise" / "left" gives S*A; 1 / "rowwise" / "right" gives A*S^T
        (sketch.py:130-200)."""
        dim = self._normdim(dim)
        A, SA, cinvert = self._prepare(A, SA, dim)
        self._call(A, SA, (1 - dim) if cinvert else dim, None)
        return SA.getobj()

    def apply_shard(self, A_local, SA=None, dim=0, index_offset=0):
        """A_local holds sketched indices [index_offset, index_offset + local) of the
        transform's input dimension; SA receives this shard's partial sketch.  Summing the
        partial sketches of a partition of [0, n) gives `apply` of the whole matrix -- the
        local step of sketch/hash_transform_Elemental.hpp:543-561 and
        sketch/dense_transform_Elemental_coldist_star.hpp:200-229."""
        dim = self._normdim(dim)
        A = lib.adapt(A_local)
        A, SA, cinvert = self._prepare(A, SA, dim, local_n=A.getdim(dim))
        self._call(A, SA, (1 - dim) if cinvert else dim, int(index_offset))
        return SA.getobj()

    def _call(self, A, SA, cdim, offset):
        if isinstance(A, (lib.TorchAdapter, lib.DeviceCSCAdapter)):
            lib.set_stream_from_torch()
        Aobj = A.ptr()
        try:
            SAobj = SA.ptr()
            try:
                if offset is None:
                    lib.callsl("sl_apply_sketch_transform", self._obj, A.ctype(), Aobj,
                               SA.ctype(), SAobj, cdim + 1)
                else:
                    lib.callsl("sl_apply_sketch_transform_shard", self._obj, A.ctype(), Aobj,
                               SA.ctype(), SAobj, cdim + 1, offset)
            finally:
                SA.ptrcleaner()
        finally:
            A.ptrcleaner()

    def apply_distributed(self, A_local, SA=None, dim=0, index_offset=0, group=None,
                          reduce="allreduce", dst=0):
        """Row-sharded sketch over a torch.distributed group: every rank sketches its shard
        of the sketched dimension, then the partial sketches are summed.
        reduce = "allreduce"      replicated result  (reference: -> SharedMatrix [*,*],
                                  sketch/hash_transform_Elemental.hpp:557-561)
                 "reduce"         result on rank `dst` (-> RootMatrix [o,o], :360-366)
                 "reduce_scatter" each rank keeps a contiguous slice of the rows of SA's
                                  storage (-> [VC,*]-like, El::AxpyContract at
                                  sketch/dense_transform_Elemental_coldist_star.hpp:228)
                 None             no collective (caller reduces)"""
        from . import dist
        return dist.apply_distributed(self, A_local, SA, dim, index_offset, group, reduce, dst)

    def __mul__(self, A):
        return self.apply(A, None, dim=0)

    def __truediv__(self, A):
        return self.apply(A, None, dim=1)

    __div__ = __truediv__

    def getindim(self):
        return self._n

    def getsketchdim(self):
        return self._s

    # -- generated state, for parity tests -------------------------------------
    def _export(self, what, count, dtype):
        out = numpy.empty(count, dtype=dtype)
        lib.callsl("sl_sketch_export", self._obj, what, 0, int(count), out.ctypes.data_as(c_void_p))
        return out

    def export_dense(self, i0=0, j0=0, h=None, w=None):
        """S[i0:i0+h, j0:j0+w] of a dense transform, generated on the device."""
        h = self._s - i0 if h is None else h
        w = self._n - j0 if w is None else w
        out = numpy.empty((h, w), dtype=numpy.float64, order="F")
        lib.callsl("sl_sketch_export_dense", self._obj, int(i0), int(j0), int(h), int(w),
                   out.ctypes.data_as(c_void_p))
        return out


class JLT(_SketchTransform):
    """Johnson-Lindenstrauss: dense i.i.d. N(0,1)/sqrt(s) map (sketch.py:238-301)."""

    def __init__(self, n, s, defouttype=None, forceppy=False, sketch_transform=None, context=None):
        super().__init__("JLT", n, s, defouttype, forceppy, sketch_transform, context)


class CT(_SketchTransform):
    """Cauchy transform: dense i.i.d. Cauchy * C/s (sketch.py:340-375)."""

    def __init__(self, n, s, C, defouttype=None, forceppy=False, sketch_transform=None, context=None):
        self._C = C
        super().__init__("CT", n, s, defouttype, forceppy, sketch_transform, context, params=(C,))


class FJLT(_SketchTransform):
    """Fast Johnson-Lindenstrauss: sqrt(n/s) * sample(DCT-II(D x)) (sketch.py:377-407)."""

    def __init__(self, n, s, defouttype=None, forceppy=False, sketch_transform=None, context=None):
        super().__init__("FJLT", n, s, defouttype, forceppy, sketch_transform, context)

    def diagonal(self):
        return self._export("D", self._n, numpy.float64)

    def samples(self):
        return self._export("samples", self._s, numpy.int64)


class UST(_SketchTransform):
    """Uniform sampling transform: SA = A[samples, :] (columnwise) with `s` indices drawn uniformly
    from range(n), with or without replacement (sketch/UST_data.hpp:81-102, UST_Elemental.hpp).
    The reference's Python layer only has the numpy-random `URST`; this is its C++ UST_t, whose
    samples come from the context's counter stream."""

    def __init__(self, n, s, replace=True, defouttype=None, forceppy=False, sketch_transform=None, context=None):
        self._replace = bool(replace)
        super().__init__("UST", n, s, defouttype, forceppy, sketch_transform, context, params=(bool(replace),))

    def samples(self):
        return self._export("samples", self._s, numpy.int64)


class _HashTransform(_SketchTransform):
    def row_idx(self):
        return self._export("row_idx", self._n, numpy.int64)

    def row_value(self):
        return self._export("row_value", self._n, numpy.float64)


class CWT(_HashTransform):
    """Clarkson-Woodruff (CountSketch) transform (sketch.py:409-440)."""

    def __init__(self, n, s, defouttype=None, forceppy=False, sketch_transform=None, context=None):
        super().__init__("CWT", n, s, defouttype, forceppy, sketch_transform, context)


class MMT(_HashTransform):
    """Meng-Mahoney transform: CountSketch with Cauchy values (sketch.py:442-471)."""

    def __init__(self, n, s, defouttype=None, forceppy=False, sketch_transform=None, context=None):
        super().__init__("MMT", n, s, defouttype, forceppy, sketch_transform, context)


class WZT(_HashTransform):
    """Woodruff-Zhang transform for l_p, 1 <= p <= 2 (sketch.py:473-526)."""

    def __init__(self, n, s, p, defouttype=None, forceppy=False, sketch_transform=None, context=None):
        self._p = p
        super().__init__("WZT", n, s, defouttype, forceppy, sketch_transform, context, params=(p,))


class _RFT(_SketchTransform):
    def shifts(self):
        return self._export("shifts", self._s, numpy.float64)

    def scales(self):
        return self._export("scales", self._s, numpy.float64)


class GaussianRFT(_RFT):
    """Random Fourier features of the Gaussian kernel (sketch.py:528-563)."""

    def __init__(self, n, s, sigma=1.0, defouttype=None, forceppy=False, sketch_transform=None,
                 context=None):
        self._sigma = sigma
        super().__init__("GaussianRFT", n, s, defouttype, forceppy, sketch_transform, context,
                         params=(sigma,))


class LaplacianRFT(_RFT):
    """Random Fourier features of the Laplacian kernel (sketch.py:565-590)."""

    def __init__(self, n, s, sigma=1.0, defouttype=None, forceppy=False, sketch_transform=None,
                 context=None):
        self._sigma = sigma
        super().__init__("LaplacianRFT", n, s, defouttype, forceppy, sketch_transform, context,
                         params=(sigma,))


# ---------------------------------------------------------------------------
# Composite transforms: their random tables come from the context's counter
# stream through the C ABI (bit for bit the reference's), their CountSketches run on the hash
# kernels, and the FFT / DCT steps go through torch.fft (cuFFT) -- library code, as FFTW is in the
# reference.  Device tensors in, device tensors out; numpy operands are copied over.
# ---------------------------------------------------------------------------
def _torch():
    import torch
    return torch


def _to_device_f64(A):
    torch = _torch()
    if isinstance(A, torch.Tensor):
        if not A.is_cuda or A.dtype != torch.float64:
            raise errors.UnsupportedError("composite transforms take float64 CUDA tensors or numpy arrays")
        return A, True
    return torch.from_numpy(numpy.ascontiguousarray(A, dtype=numpy.float64)).cuda(), False


def _dct2_unnormalized(X):
    """FFTW REDFT10 along dim 0: Y_k = 2 sum_n x_n cos(pi (n + 1/2) k / N) (utility/fft/fftw_futs.h:117-126),
    by Makhoul's even/odd reordering and one complex FFT of length N."""
    torch = _torch()
    N = X.shape[0]
    v = torch.cat([X[0::2], X[1::2].flip(0)], dim=0)
    V = torch.fft.fft(v, dim=0)
    k = torch.arange(N, device=X.device, dtype=torch.float64)
    w = torch.polar(torch.ones_like(k), -math.pi * k / (2.0 * N))
    return 2.0 * (V * w.reshape((N,) + (1,) * (X.dim() - 1))).real


class _Composite(object):
    def _ctx(self, context):
        return context if context is not None else lib._GlobalContext()

    def getindim(self):
        return self._n

    def getsketchdim(self):
        return self._s

    def __mul__(self, A):
        return self.apply(A, None, dim=0)

    def __truediv__(self, A):
        return self.apply(A, None, dim=1)

    def _finish(self, out, SA, was_device, dim):
        """out: s x m (one column per data vector).  Hands back what the caller's operands call for."""
        res = out if dim == 0 else out.t()
        if SA is not None:
            if isinstance(SA, _torch().Tensor):
                SA.copy_(res)
            else:
                SA[...] = res.cpu().numpy()
            return SA
        return res.contiguous() if was_device else res.cpu().numpy()


class FastGaussianRFT(_Composite):
    """Fastfood features of the Gaussian kernel (sketch/FRFT_data.hpp:19-160, FRFT_Elemental.hpp:70-135;
    sketch.py:679-741).  Per block of N features: scale * cos(Sm .* DCT(G .* Perm(DCT(B .* x))) + shifts),
    DCT = the unnormalised DCT-II of the FFTW build, B Rademacher, G Gaussian, Perm a Fisher-Yates
    shuffle, Sm = 1 / (sigma sqrt(N)); tables drawn in the reference's order (shifts, B, G, P)."""

    def __init__(self, n, s, sigma=1.0, defouttype=None, forceppy=False, sketch_transform=None, context=None):
        defouttype, context = _split_context(defouttype, context)
        if forceppy:
            raise errors.UnsupportedError("forceppy: this package has no pure-Python / CPU implementation")
        torch = _torch()
        self._n, self._s, self._sigma = int(n), int(s), float(sigma)
        ctx = self._ctx(context)
        N, S = self._n, self._s
        nb = 1 + (S - 1) // N                                    # numblks, block size NB = N (FFTW build)
        seed = ctx.seed
        self._creation = (seed, ctx.counter)
        two_pi = 2.0 * 3.141592653589793238462643383279502884      # boost::math::constants::pi
        shifts = lib.rng_export("uniform_real", seed, ctx.allocate(S), S, 0.0, two_pi)
        B = lib.rng_export("rademacher", seed, ctx.allocate(nb * N), nb * N)
        G = lib.rng_export("normal", seed, ctx.allocate(nb * N), nb * N)
        praw = lib.rng_export_raw(seed, ctx.allocate(nb * (N - 1)), nb * (N - 1))[0]
        scal = math.sqrt(N) * (1.0 / math.sqrt(2.0 * N))         # sqrt(NB) * fut.scale()
        perms = numpy.empty((nb, N), dtype=numpy.int64)
        for i in range(nb):                                      # FRFT_data.hpp:111-114 + the swap loop of apply
            arr = list(range(N))
            for l in range(N - 1):
                idx1 = N - 1 - l
                idx2 = int(praw[i * (N - 1) + l] % numpy.uint64(idx1 + 1))
                arr[idx1], arr[idx2] = arr[idx2], arr[idx1]
            perms[i] = arr
        dev = "cuda"
        self._B = torch.from_numpy(B.reshape(nb, N)).to(dev)
        self._G = torch.from_numpy((scal * G).reshape(nb, N)).to(dev)
        self._Sm = scal * (1.0 / (self._sigma * math.sqrt(N)))
        self._perm = torch.from_numpy(perms).to(dev)
        self._shifts = torch.from_numpy(shifts).to(dev)
        self._scale = math.sqrt(2.0 / S)
        self._nb = nb

    def apply(self, A, SA=None, dim=0):
        torch = _torch()
        dim = _SketchTransform._normdim(dim)
        X, was_device = _to_device_f64(A)
        X = X if dim == 0 else X.t()                             # N x m, one data vector per column
        if X.shape[0] != self._n:
            raise errors.DimensionMistmatchError("Sketched dimension is incorrect (input)")
        N, S = self._n, self._s
        out = torch.empty(S, X.shape[1], dtype=torch.float64, device=X.device)
        for i in range(self._nb):
            s0, e0 = i * N, min((i + 1) * N, S)
            W = _dct2_unnormalized(X * self._B[i][:, None])
            W = W[self._perm[i]]
            W = _dct2_unnormalized(W * self._G[i][:, None])
            W = W * self._Sm
            out[s0:e0] = self._scale * torch.cos(W[:e0 - s0] + self._shifts[s0:e0, None])
        return self._finish(out, SA, was_device, dim)

    def serialize(self):
        return {"skylark_object_type": "sketch", "sketch_type": "FastGaussianRFT", "skylark_version": "0.20",
                "N": str(self._n), "S": str(self._s), "sigma": repr(self._sigma),
                "creation_context": {"skylark_object_type": "context", "skylark_version": "0.20",
                                     "seed": str(self._creation[0]), "counter": str(self._creation[1])}}


class PPT(_Composite):
    """Pham-Pagh TensorSketch of the polynomial kernel (gamma x^T y + c)^q (sketch/PPT_data.hpp:19-105,
    PPT_Elemental.hpp:140-185; sketch.py:829-881): q CountSketches of sqrt(gamma) x, each with
    sqrt(c) * (+-1) added at one random position, multiplied in the Fourier domain, transformed back."""

    def __init__(self, n, s, q=3, c=0.0, gamma=1.0, defouttype=None, forceppy=False, sketch_transform=None,
                 context=None):
        defouttype, context = _split_context(defouttype, context)
        if forceppy:
            raise errors.UnsupportedError("forceppy: this package has no pure-Python / CPU implementation")
        if c < 0:
            raise ValueError("kernel parameter must be >= 0")
        self._n, self._s, self._q, self._c, self._gamma = int(n), int(s), int(q), float(c), float(gamma)
        ctx = self._ctx(context)
        seed = ctx.seed
        self._creation = (seed, ctx.counter)
        passed = context                                         # None = the global context, as CWT expects
        self._cwts = [CWT(self._n, self._s, context=passed) for _ in range(self._q)]     # PPT_data.hpp:85-88
        self._hash_idx = lib.rng_export("uniform_int", seed, ctx.allocate(self._q), self._q, self._s - 1).astype(numpy.int64)
        self._hash_val = lib.rng_export("rademacher", seed, ctx.allocate(self._q), self._q)

    def apply(self, A, SA=None, dim=0):
        torch = _torch()
        dim = _SketchTransform._normdim(dim)
        X, was_device = _to_device_f64(A)
        X = X if dim == 0 else X.t()
        if X.shape[0] != self._n:
            raise errors.DimensionMistmatchError("Sketched dimension is incorrect (input)")
        Xc = X.t().contiguous().t() if X.stride(0) != 1 else X   # column-major operand for the hash kernels
        P = None
        for i, cw in enumerate(self._cwts):
            W = cw.apply(Xc, None, "columnwise") * math.sqrt(self._gamma)
            W[self._hash_idx[i], :] += math.sqrt(self._c) * self._hash_val[i]
            FW = torch.fft.rfft(W, dim=0)
            P = FW if P is None else P * FW
        out = torch.fft.irfft(P, n=self._s, dim=0)               # FFTW's c2r is unscaled; the reference divides by S
        return self._finish(out, SA, was_device, dim)

    def serialize(self):
        return {"skylark_object_type": "sketch", "sketch_type": "PPT", "skylark_version": "0.20",
                "N": str(self._n), "S": str(self._s), "q": str(self._q), "c": repr(self._c), "gamma": repr(self._gamma),
                "creation_context": {"skylark_object_type": "context", "skylark_version": "0.20",
                                     "seed": str(self._creation[0]), "counter": str(self._creation[1])}}


def _outside_hot_path(name, where):
    """Reference transforms that are not part of the accelerated path keep their names, so that
    code importing them fails with a clear message instead of an AttributeError."""
    class _Unsupported(object):
        def __init__(self, *args, **kwargs):
            raise errors.UnsupportedError(
                "%s (%s) is outside the GPU sketch-apply path of this package; see DESIGN.md "
                "section 7" % (name, where))
    _Unsupported.__name__ = name
    return _Unsupported


SJLT = _outside_hot_path("SJLT", "sketch.py:303-338, pure Python in the reference")
MaternRFT = _outside_hot_path("MaternRFT", "sketch.py:592-619")
GaussianQRFT = _outside_hot_path("GaussianQRFT", "sketch.py:621-648")
LaplacianQRFT = _outside_hot_path("LaplacianQRFT", "sketch.py:650-677")
FastMaternRFT = _outside_hot_path("FastMaternRFT", "sketch.py:743-772")
ExpSemigroupRLT = _outside_hot_path("ExpSemigroupRLT", "sketch.py:774-800")
ExpSemigroupQRLT = _outside_hot_path("ExpSemigroupQRLT", "sketch.py:802-827")
URST = UST      # sketch.py:883-904 is a numpy.random sampler; here it is the C++ UST_t
NURST = _outside_hot_path("NURST", "sketch.py:906-931, pure Python in the reference")

# Additional names (sketch.py:962-970); FastJLT = JLT is the reference's own aliasing.
SparseJLT = SJLT
Fastfood = FastGaussianRFT
MaternFastfood = FastMaternRFT
TensorSketch = PPT
UniformSampler = URST
NonUniformSampler = NURST
FastJLT = JLT
CountSketch = CWT
RRT = GaussianRFT
RFT = GaussianRFT

_map_csketch_type_to_cfun = {
    "JLT": lambda sd, obj: JLT(int(sd["N"]), int(sd["S"]), None, False, obj),
    "CT": lambda sd, obj: CT(int(sd["N"]), int(sd["S"]), float(sd["C"]), None, False, obj),
    "FJLT": lambda sd, obj: FJLT(int(sd["N"]), int(sd["S"]), None, False, obj),
    "UST": lambda sd, obj: UST(int(sd["N"]), int(sd["S"]), str(sd["replace"]) in ("true", "1", "True"), None, False, obj),
    "CWT": lambda sd, obj: CWT(int(sd["N"]), int(sd["S"]), None, False, obj),
    "MMT": lambda sd, obj: MMT(int(sd["N"]), int(sd["S"]), None, False, obj),
    "WZT": lambda sd, obj: WZT(int(sd["N"]), int(sd["S"]), float(sd["P"]), None, False, obj),
    "GaussianRFT": lambda sd, obj: GaussianRFT(int(sd["N"]), int(sd["S"]), float(sd["sigma"]),
                                               None, False, obj),
    "LaplacianRFT": lambda sd, obj: LaplacianRFT(int(sd["N"]), int(sd["S"]), float(sd["sigma"]),
                                                 None, False, obj),
}
