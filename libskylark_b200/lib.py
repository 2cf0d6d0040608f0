"""ctypes loader, global context and matrix adapters.

Python-3 counterpart of python-skylark/skylark/lib.py: same names and call protocol
(`initialize`, `finalize`, `callsl`, `adapt`, `map_to_ctor`, `NumpyAdapter`,
`ScipyAdapter`), binding libcskylark_b200.so instead of libcskylark.so.  Additions:
`Context(seed, counter)` objects, a `TorchAdapter` for CUDA tensors (zero copy) and a
`DeviceCSC` container for device-resident CSC matrices.
"""
import atexit
import ctypes
import ctypes.util
import os
import time
from ctypes import (POINTER, byref, c_bool, c_char_p, c_double, c_int, c_int64,  # noqa: F401
                    c_size_t, c_uint64, c_void_p)

import numpy
import scipy.sparse

from . import _build, errors

_libc = ctypes.CDLL(ctypes.util.find_library("c"))
_libc.free.argtypes = (ctypes.c_void_p,)

lib = None
ctxt_obj = None
ELEM_INSTALLED = False
KDT_INSTALLED = False

SL_DTYPE_F64 = 0
SL_DTYPE_F32 = 1
KERNEL_CLASSES = ("tables", "hash", "dense_gen", "dense_gemm", "fjlt", "misc")


def _load():
    """Load the CUDA library; fail loudly when it is missing (there is no fallback)."""
    so = _build.SO
    if not os.path.exists(so):
        try:
            _build.build()
        except Exception as e:  # pragma: no cover
            raise ImportError(
                "libskylark_b200: %s is missing and could not be built (%s). Run "
                "`python -m libskylark_b200._build`; this package has no CPU path." % (so, e))
    L = ctypes.CDLL(so)
    # restypes exactly as lib.py:27-48 sets them
    for name in ("sl_create_context", "sl_create_default_context", "sl_free_context",
                 "sl_create_sketch_transform", "sl_serialize_sketch_transform",
                 "sl_deserialize_sketch_transform", "sl_wrap_raw_matrix",
                 "sl_free_raw_matrix_wrap", "sl_wrap_raw_sp_matrix",
                 "sl_free_raw_sp_matrix_wrap", "sl_raw_sp_matrix_nnz", "sl_raw_sp_matrix_height",
                 "sl_raw_sp_matrix_width", "sl_raw_sp_matrix_struct_updated",
                 "sl_raw_sp_matrix_reset_update_flag", "sl_raw_sp_matrix_data",
                 "sl_apply_sketch_transform", "sl_free_sketch_transform",
                 "sl_create_context_at", "sl_context_get_seed", "sl_context_get_counter",
                 "sl_wrap_device_matrix", "sl_wrap_device_sp_matrix", "sl_set_stream",
                 "sl_apply_sketch_transform_shard", "sl_set_option", "sl_sketch_export",
                 "sl_sketch_export_dense", "sl_rng_export_raw", "sl_profile_enable",
                 "sl_profile_reset", "sl_profile_read", "sl_host_alloc_pinned",
                 "sl_host_free_pinned", "sl_device_count", "sl_rng_export", "sl_context_allocate",
                 "sl_fjlt_apply_shard_peer", "sl_sp_merge_segments", "sl_sp_matrix_coalesce"):
        getattr(L, name).restype = c_int
    L.sl_strerror.restype = c_char_p
    L.sl_supported_sketch_transforms.restype = c_char_p
    L.sl_has_elemental.restype = c_bool
    L.sl_has_combblas.restype = c_bool
    L.sl_get_exception_info.restype = None
    L.sl_print_exception_trace.restype = None
    # argtypes where ctypes' defaults would truncate 64-bit values
    L.sl_create_context_at.argtypes = (c_int, c_uint64, c_void_p)
    L.sl_wrap_device_matrix.argtypes = (c_void_p, c_int64, c_int64, c_int64, c_int, c_int, c_void_p)
    L.sl_wrap_device_sp_matrix.argtypes = (c_void_p, c_void_p, c_void_p, c_int64, c_int64,
                                           c_int64, c_int, c_void_p)
    L.sl_apply_sketch_transform_shard.argtypes = (c_void_p, c_char_p, c_void_p, c_char_p,
                                                  c_void_p, c_int, c_int64)
    L.sl_apply_sketch_transform.argtypes = (c_void_p, c_char_p, c_void_p, c_char_p, c_void_p, c_int)
    L.sl_sketch_export.argtypes = (c_void_p, c_char_p, c_int64, c_int64, c_void_p)
    L.sl_sketch_export_dense.argtypes = (c_void_p, c_int64, c_int64, c_int64, c_int64, c_void_p)
    L.sl_rng_export_raw.argtypes = (c_int, c_uint64, c_int64, c_void_p, c_void_p)
    L.sl_rng_export.argtypes = (c_int, c_int, c_uint64, c_int64, c_double, c_double, c_void_p)
    L.sl_context_allocate.argtypes = (c_void_p, c_uint64, c_void_p)
    L.sl_sp_merge_segments.argtypes = (c_void_p, c_void_p, c_void_p, c_int, c_int64, c_int64, c_int64, c_void_p)
    L.sl_fjlt_apply_shard_peer.argtypes = (c_void_p, c_void_p, c_int64, c_void_p, c_int, c_int, c_void_p)
    L.sl_set_stream.argtypes = (c_void_p,)
    L.sl_host_alloc_pinned.argtypes = (c_size_t, c_void_p)
    L.sl_host_free_pinned.argtypes = (c_void_p,)
    L.sl_free_sketch_transform.argtypes = (c_void_p,)
    L.sl_free_context.argtypes = (c_void_p,)
    L.sl_free_raw_matrix_wrap.argtypes = (c_void_p,)
    L.sl_free_raw_sp_matrix_wrap.argtypes = (c_void_p,)
    return L


def initialize(seed=-1):
    """(Re)initialise the library with a seed; -1 uses the system time (lib.py:15-60)."""
    global lib, ctxt_obj, ELEM_INSTALLED, KDT_INSTALLED
    if lib is None:
        lib = _load()
    ELEM_INSTALLED = lib.sl_has_elemental()
    KDT_INSTALLED = lib.sl_has_combblas()
    if seed == -1:
        seed = int(time.time())
    if ctxt_obj is not None and ctxt_obj.value:
        lib.sl_free_context(ctxt_obj)
    ctxt_obj = c_void_p()
    lib.sl_create_default_context(c_int(int(seed)), byref(ctxt_obj))


def finalize():
    """Free the global context (lib.py:63-74).  Sketch objects are freed by the GC."""
    global ctxt_obj
    if lib is not None and ctxt_obj is not None and ctxt_obj.value:
        lib.sl_free_context(ctxt_obj)
    ctxt_obj = None


atexit.register(finalize)


def callsl(fname, *args):
    """Call the C function `fname`; raise LowerLayerError on a non-zero code (lib.py:82-93)."""
    f = getattr(lib, fname)
    args = [a.encode() if isinstance(a, str) else a for a in args]
    errno = f(*args)
    if errno != 0:
        c_errinfo = c_char_p()
        lib.sl_get_exception_info(byref(c_errinfo))
        errinfo = c_errinfo.value.decode() if c_errinfo.value else ""
        _libc.free(c_errinfo)
        err = errors.LowerLayerError("%s (code %d: %s)" % (
            errinfo, errno, lib.sl_strerror(errno).decode()))
        err.code = errno
        raise err


class Context:
    """Explicit `context_t(seed, counter)` (base/context.hpp:25-27); an extension over the
    reference's single global context.  Creating a sketch from it advances its counter."""

    def __init__(self, seed, counter=0):
        self._obj = c_void_p()
        callsl("sl_create_context_at", int(seed), int(counter), byref(self._obj))

    def __del__(self):
        try:
            if lib is not None and getattr(self, "_obj", None) is not None and self._obj.value:
                lib.sl_free_context(self._obj)
                self._obj = None
        except Exception:
            pass

    @property
    def seed(self):
        v = c_int()
        callsl("sl_context_get_seed", self._obj, byref(v))
        return v.value

    @property
    def counter(self):
        v = c_uint64()
        callsl("sl_context_get_counter", self._obj, byref(v))
        return v.value

    def allocate(self, size):
        """Reserve `size` stream indices (base/context.hpp:94-101); returns the first."""
        v = c_uint64()
        callsl("sl_context_allocate", self._obj, c_uint64(int(size)), byref(v))
        return v.value


class _GlobalContext(object):
    """The reference's single global context (lib.py:80) behind the Context interface."""

    @property
    def _obj(self):
        return ctxt_obj

    seed = Context.seed
    counter = Context.counter
    allocate = Context.allocate


def rng_export(kind, seed, base, count, a=0.0, b=0.0):
    """Variates of stream indices [base, base + count) generated on the device (sl_rng_export):
    kind "rademacher" | "normal" | "cauchy" | "uniform_real" (a, b) | "uniform_int" (0..a)."""
    code = {"rademacher": 0, "normal": 1, "cauchy": 2, "uniform_real": 3, "uniform_int": 4}[kind]
    out = numpy.empty(int(count), dtype=numpy.float64)
    callsl("sl_rng_export", c_int(code), c_int(int(seed)), c_uint64(int(base)), ctypes.c_int64(int(count)),
           c_double(float(a)), c_double(float(b)), out.ctypes.data_as(c_void_p))
    return out


def rng_export_raw(seed, base, count):
    """First and second 64-bit draw of every stream index in [base, base + count)."""
    first = numpy.empty(int(count), dtype=numpy.uint64)
    second = numpy.empty(int(count), dtype=numpy.uint64)
    callsl("sl_rng_export_raw", c_int(int(seed)), c_uint64(int(base)), ctypes.c_int64(int(count)),
           first.ctypes.data_as(c_void_p), second.ctypes.data_as(c_void_p))
    return first, second


def global_counter():
    v = c_uint64()
    callsl("sl_context_get_counter", ctxt_obj, byref(v))
    return v.value


def set_option(key, value):
    callsl("sl_set_option", key, str(value))


# ---------------------------------------------------------------------------
# Matrix adapters (lib.py:99-266).  `iscompatible` returns (error, cinvert): C-ordered
# data is handed to the C layer as its transpose, with the sketch direction flipped.
# ---------------------------------------------------------------------------
class NumpyAdapter:
    def __init__(self, A):
        if A.dtype.type is not numpy.float64:
            raise errors.UnsupportedError("Only float64 matrices are supported.")
        if A.base is not None:
            raise errors.UnsupportedError("Passing a numpy matrix view is not supported.")
        self._A = A

    def ctype(self):
        return "Matrix"

    def ptr(self):
        data = c_void_p()
        A = self._A
        dptr = A.ctypes.data_as(ctypes.POINTER(ctypes.c_double))
        if self.getorder() == "F":
            callsl("sl_wrap_raw_matrix", dptr, A.shape[0], A.shape[1] if A.ndim > 1 else 1,
                   byref(data))
        else:
            callsl("sl_wrap_raw_matrix", dptr, A.shape[1] if A.ndim > 1 else A.shape[0],
                   A.shape[0] if A.ndim > 1 else 1, byref(data))
        self._ptr = data
        return data

    def ptrcleaner(self):
        callsl("sl_free_raw_matrix_wrap", self._ptr)

    def getdim(self, dim):
        return self._A.shape[dim]

    def getobj(self):
        return self._A

    def orders(self):
        """Element orders this array satisfies; degenerate shapes (one row / one column /
        empty) are both C- and F-contiguous, which lib.py:137-141 resolves to 'F' blindly."""
        o = set()
        if self._A.flags.f_contiguous:
            o.add("F")
        if self._A.flags.c_contiguous:
            o.add("C")
        return o

    def getorder(self):
        if getattr(self, "_order", None):
            return self._order
        return "F" if self._A.flags.f_contiguous else "C"

    def iscompatible(self, B):
        if not isinstance(B, (NumpyAdapter, ScipyAdapter)):
            return "numpy input needs numpy output", None
        common = self.orders() & B.orders()
        if not common:
            return "sketching numpy array to numpy/scipy array requires same element ordering", None
        # agree on one order for both operands (matters only for degenerate shapes)
        order = "F" if "F" in common else "C"
        self._order = order
        B._order = order
        return None, order == "C"

    def getctor(self):
        return NumpyAdapter.ctor

    @staticmethod
    def ctor(m, n, B):
        if isinstance(B, (NumpyAdapter, ScipyAdapter)):
            return numpy.empty((m, n), order=B.getorder())
        return numpy.empty((m, n), order="F")


class ScipyAdapter:
    def __init__(self, A):
        if isinstance(A, (scipy.sparse.csr_matrix, scipy.sparse.csc_matrix)):
            self._A = A
        else:
            self._A = scipy.sparse.csr_matrix(A)
        if self._A.dtype != numpy.float64:
            raise errors.UnsupportedError("Only float64 matrices are supported.")
        self._keep = None

    def ctype(self):
        return "SparseMatrix"

    def ptr(self):
        data = c_void_p()
        A = self._A
        indptr = numpy.ascontiguousarray(A.indptr, dtype=numpy.int32)
        indices = numpy.ascontiguousarray(A.indices, dtype=numpy.int32)
        values = numpy.ascontiguousarray(A.data, dtype=numpy.float64)
        self._keep = (indptr, indices, values)
        iptr = indptr.ctypes.data_as(ctypes.POINTER(ctypes.c_int))
        cols = indices.ctypes.data_as(ctypes.POINTER(ctypes.c_int))
        dptr = values.ctypes.data_as(ctypes.POINTER(ctypes.c_double))
        if self.getorder() == "F":
            callsl("sl_wrap_raw_sp_matrix", iptr, cols, dptr, len(indices), A.shape[0], A.shape[1],
                   byref(data))
        else:
            callsl("sl_wrap_raw_sp_matrix", iptr, cols, dptr, len(indices), A.shape[1], A.shape[0],
                   byref(data))
        callsl("sl_raw_sp_matrix_reset_update_flag", data)
        self._ptr = data
        return data

    def ptrcleaner(self):
        # A sparse result replaces the wrapped structure inside the library; fetch it before
        # the wrapper goes away (same protocol as the reference's ScipyAdapter, lib.py:200-231).
        try:
            updated = ctypes.c_bool()
            callsl("sl_raw_sp_matrix_struct_updated", self._ptr, byref(updated))
            if updated.value:
                nnz, h, w = ctypes.c_int(), ctypes.c_int(), ctypes.c_int()
                callsl("sl_raw_sp_matrix_nnz", self._ptr, byref(nnz))
                callsl("sl_raw_sp_matrix_height", self._ptr, byref(h))
                callsl("sl_raw_sp_matrix_width", self._ptr, byref(w))
                indptr = numpy.zeros(w.value + 1, dtype=numpy.int32)
                indices = numpy.zeros(nnz.value, dtype=numpy.int32)
                values = numpy.zeros(nnz.value, dtype=numpy.float64)
                callsl("sl_raw_sp_matrix_data", self._ptr, indptr.ctypes.data_as(c_void_p),
                       indices.ctypes.data_as(c_void_p), values.ctypes.data_as(c_void_p))
                A = self._A
                A._shape = (h.value, w.value) if self.getorder() == "F" else (w.value, h.value)
                A.indptr, A.indices, A.data = indptr, indices, values
        finally:
            callsl("sl_free_raw_sp_matrix_wrap", self._ptr)
            self._keep = None

    def getdim(self, dim):
        return self._A.shape[dim]

    def getobj(self):
        return self._A

    def orders(self):
        return {self.getorder()}

    def getorder(self):
        return "C" if isinstance(self._A, scipy.sparse.csr_matrix) else "F"

    def iscompatible(self, B):
        if not isinstance(B, (NumpyAdapter, ScipyAdapter)):
            return "scipy input needs numpy output", None
        if self.getorder() not in B.orders():
            return "sketching scipy matrix to numpy/scipy matrix requires same format", None
        B._order = self.getorder()
        return None, self.getorder() == "C"

    def getctor(self):
        # sparse in -> sparse out by default, as in the reference (lib.py:257-266); pass a numpy
        # SA or defouttype="LocalMatrix" for a dense sketch of a sparse input.
        return ScipyAdapter.ctor

    @staticmethod
    def ctor(m, n, B):
        if isinstance(B, (ScipyAdapter, NumpyAdapter)) and B.getorder() == "C":
            return scipy.sparse.csr_matrix((m, n))
        return scipy.sparse.csc_matrix((m, n))


def _torch():
    import torch
    return torch


class TorchAdapter:
    """2-D CUDA tensor, float64 (float32 for JLT/CT/RFT), row- or column-major, zero copy."""

    def __init__(self, A):
        torch = _torch()
        if not A.is_cuda:
            raise errors.UnsupportedError("torch tensors must live on a CUDA device")
        if A.dim() != 2:
            raise errors.UnsupportedError("only 2-D tensors are supported")
        if A.dtype not in (torch.float64, torch.float32):
            raise errors.UnsupportedError("Only float64/float32 tensors are supported.")
        if A.stride(1) == 1 and A.stride(0) >= max(1, A.shape[1]):
            self._order = "C"
        elif A.stride(0) == 1 and A.stride(1) >= max(1, A.shape[0]):
            self._order = "F"
        else:
            raise errors.UnsupportedError("tensor must be row-major or column-major with unit inner stride")
        if A.shape[0] == 1 and A.shape[1] == 1:
            self._order = "C"
        self._A = A

    def ctype(self):
        return "Matrix"

    def dtype_code(self):
        return SL_DTYPE_F64 if self._A.dtype == _torch().float64 else SL_DTYPE_F32

    def ptr(self):
        A = self._A
        data = c_void_p()
        dev = A.device.index if A.device.index is not None else _torch().cuda.current_device()
        if self._order == "F":
            m, n, ld = A.shape[0], A.shape[1], A.stride(1) if A.shape[1] > 1 else max(1, A.shape[0])
        else:
            m, n, ld = A.shape[1], A.shape[0], A.stride(0) if A.shape[0] > 1 else max(1, A.shape[1])
        callsl("sl_wrap_device_matrix", c_void_p(A.data_ptr()), m, n, max(ld, m, 1),
               self.dtype_code(), dev, byref(data))
        self._ptr = data
        return data

    def ptrcleaner(self):
        callsl("sl_free_raw_matrix_wrap", self._ptr)

    def getdim(self, dim):
        return self._A.shape[dim]

    def getobj(self):
        return self._A

    def getorder(self):
        return self._order

    def iscompatible(self, B):
        if not isinstance(B, TorchAdapter):
            return "device input needs a device (torch CUDA) output", None
        if self.getorder() != B.getorder():
            return "sketching a tensor to a tensor requires same element ordering", None
        if self._A.device != B._A.device or self._A.dtype != B._A.dtype:
            return "input and output tensors must share device and dtype", None
        return None, self.getorder() == "C"

    def getctor(self):
        return TorchAdapter.ctor

    @staticmethod
    def ctor(m, n, B):
        torch = _torch()
        A = B.getobj() if isinstance(B, (TorchAdapter, DeviceCSCAdapter)) else None
        dev = A.device if A is not None else "cuda"
        dt = A.dtype if isinstance(B, TorchAdapter) else torch.float64
        if isinstance(B, (TorchAdapter, DeviceCSCAdapter)) and B.getorder() == "F":
            return torch.empty((n, m), device=dev, dtype=dt).t()
        return torch.empty((m, n), device=dev, dtype=dt)


class DeviceCSC:
    """Device-resident sparse matrix in compressed form: `indptr`, `indices` (int32 CUDA
    tensors) and `values` (float64 CUDA tensor).  format="csc" stores columns (what the
    reference's sparse_matrix_t holds, base/sparse_matrix.hpp:24-140); format="csr" stores
    rows and is passed down as the CSC of the transpose, like scipy CSR in lib.py:189-197."""

    def __init__(self, indptr, indices, values, shape, format="csc"):
        self.indptr, self.indices, self.values = indptr, indices, values
        self.shape = tuple(shape)
        self.format = format
        self.device = values.device
        self.dtype = values.dtype

    @classmethod
    def empty(cls, shape, format="csc", device="cuda"):
        """All-zero matrix; an output operand for sparse -> sparse CWT/MMT/WZT."""
        torch = _torch()
        nptr = (shape[1] if format == "csc" else shape[0]) + 1
        return cls(torch.zeros(nptr, dtype=torch.int32, device=device),
                   torch.zeros(0, dtype=torch.int32, device=device),
                   torch.zeros(0, dtype=torch.float64, device=device), shape, format)

    def to_scipy(self):
        ctor = scipy.sparse.csc_matrix if self.format == "csc" else scipy.sparse.csr_matrix
        return ctor((self.values.cpu().numpy(), self.indices.cpu().numpy(),
                     self.indptr.cpu().numpy()), shape=self.shape)


class DeviceCSCAdapter:
    def __init__(self, A):
        torch = _torch()
        if A.indptr.dtype != torch.int32 or A.indices.dtype != torch.int32 \
                or A.values.dtype != torch.float64:
            raise errors.UnsupportedError("DeviceCSC needs int32 indptr/indices and float64 values")
        self._A = A

    def ctype(self):
        return "SparseMatrix"

    def ptr(self):
        A = self._A
        data = c_void_p()
        dev = A.values.device.index if A.values.device.index is not None else 0
        rows, cols = A.shape if A.format == "csc" else (A.shape[1], A.shape[0])
        callsl("sl_wrap_device_sp_matrix", c_void_p(A.indptr.data_ptr()),
               c_void_p(A.indices.data_ptr()), c_void_p(A.values.data_ptr()),
               int(A.values.numel()), rows, cols, dev, byref(data))
        self._ptr = data
        return data

    def ptrcleaner(self):
        try:
            updated = ctypes.c_bool()
            callsl("sl_raw_sp_matrix_struct_updated", self._ptr, byref(updated))
            if updated.value:     # sparse result: copy it into tensors torch owns
                torch = _torch()
                nnz, h, w = ctypes.c_int(), ctypes.c_int(), ctypes.c_int()
                callsl("sl_raw_sp_matrix_nnz", self._ptr, byref(nnz))
                callsl("sl_raw_sp_matrix_height", self._ptr, byref(h))
                callsl("sl_raw_sp_matrix_width", self._ptr, byref(w))
                A = self._A
                dev = A.values.device
                A.indptr = torch.empty(w.value + 1, dtype=torch.int32, device=dev)
                A.indices = torch.empty(nnz.value, dtype=torch.int32, device=dev)
                A.values = torch.empty(nnz.value, dtype=torch.float64, device=dev)
                callsl("sl_raw_sp_matrix_data", self._ptr, c_void_p(A.indptr.data_ptr()),
                       c_void_p(A.indices.data_ptr()), c_void_p(A.values.data_ptr()))
                A.shape = (h.value, w.value) if A.format == "csc" else (w.value, h.value)
        finally:
            callsl("sl_free_raw_sp_matrix_wrap", self._ptr)

    def getdim(self, dim):
        return self._A.shape[dim]

    def getobj(self):
        return self._A

    def getorder(self):
        return "F" if self._A.format == "csc" else "C"

    def iscompatible(self, B):
        if isinstance(B, DeviceCSCAdapter):
            if self.getorder() != B.getorder():
                return "sketching device CSC to device CSC requires the same format", None
            return None, self.getorder() == "C"
        if not isinstance(B, TorchAdapter):
            return "device sparse input needs a device (torch CUDA) output", None
        if self.getorder() != B.getorder():
            return "sketching CSC needs a column-major output, CSR a row-major one", None
        return None, self.getorder() == "C"

    def getctor(self):
        return TorchAdapter.ctor

    @staticmethod
    def ctor(m, n, B):
        A = B.getobj() if isinstance(B, (TorchAdapter, DeviceCSCAdapter)) else None
        fmt = "csr" if (A is not None and B.getorder() == "C") else "csc"
        return DeviceCSC.empty((m, n), fmt, A.device if A is not None else "cuda")


def adapt(obj):
    """Wrap a supported matrix object in its adapter (lib.py:405-441)."""
    if isinstance(obj, (NumpyAdapter, ScipyAdapter, TorchAdapter, DeviceCSCAdapter)):
        return obj
    if isinstance(obj, numpy.ndarray):
        return NumpyAdapter(obj)
    if isinstance(obj, (scipy.sparse.csr_matrix, scipy.sparse.csc_matrix)):
        return ScipyAdapter(obj)
    if isinstance(obj, DeviceCSC):
        return DeviceCSCAdapter(obj)
    try:
        torch = _torch()
        if isinstance(obj, torch.Tensor):
            return TorchAdapter(obj)
    except ImportError:  # pragma: no cover
        pass
    raise errors.InvalidObjectError("Invalid/unsupported object passed as parameter")


map_to_ctor = {}
map_to_ctor["LocalMatrix"] = NumpyAdapter.ctor
map_to_ctor["LocalSpMatrix"] = ScipyAdapter.ctor
map_to_ctor["DeviceMatrix"] = TorchAdapter.ctor
map_to_ctor["DeviceSpMatrix"] = DeviceCSCAdapter.ctor


def set_stream_from_torch():
    """Make the library launch on torch's current CUDA stream."""
    torch = _torch()
    callsl("sl_set_stream", c_void_p(torch.cuda.current_stream().cuda_stream))


def profile_enable(timing=True):
    callsl("sl_profile_enable", 1 if timing else 0)


def profile_reset():
    callsl("sl_profile_reset")


def profile_read():
    """Returns {class: (launches, ms)}; synchronises pending event pairs."""
    n = (ctypes.c_ulonglong * 6)()
    ms = (ctypes.c_double * 6)()
    callsl("sl_profile_read", n, ms)
    return {k: (int(n[i]), float(ms[i])) for i, k in enumerate(KERNEL_CLASSES)}


def device_count():
    n = c_int()
    callsl("sl_device_count", byref(n))
    return n.value


# Load the C API on import, like the reference (lib.py:80).
initialize(int(time.time()) & 0x7FFFFFFF)
