"""sl_dense_sketch_apply_noalloc: the dense sketch for callers that may not allocate GPU
memory while they run (real-time loops, processes whose memory pool belongs to someone
else, hosts that forbid cudaMalloc after start-up).  It must give what sl_apply_sketch_transform gives, call after call,
although every call reuses the same built-in scratch array.  Collected last on purpose."""
import ctypes
import math

import pytest

from util import rel_fro

pytestmark = pytest.mark.gpu
TOL = 1e-12    # same bound as the other fp64 apply tests


def _noalloc(lib, seed, base, dist, scale, n, s, A, SA, dim):
    import torch
    fn = lib.lib.sl_dense_sketch_apply_noalloc
    fn.restype = ctypes.c_int
    fn.argtypes = (ctypes.c_int, ctypes.c_uint64, ctypes.c_int, ctypes.c_double, ctypes.c_int64,
                   ctypes.c_int64, ctypes.c_void_p, ctypes.c_int64, ctypes.c_int64, ctypes.c_void_p,
                   ctypes.c_int64, ctypes.c_int, ctypes.c_void_p)
    # column-major views: a row-major (r x c) tensor is the column-major (c x r) matrix
    m = A.shape[1] if dim == 2 else A.shape[0]
    return fn(seed, base, dist, scale, n, s, A.data_ptr(), A.stride(0), m, SA.data_ptr(),
              SA.stride(0), dim, ctypes.c_void_p(torch.cuda.current_stream().cuda_stream))


def test_matches_regular_entry_point_twice_in_a_row(sk):
    import torch
    from libskylark_b200 import lib
    n, s, d = 3000, 200, 96
    ctx = lib.Context(4242, 1000)
    J = sk.JLT(n, s, ctx)                       # uses counters [1000, 1000 + n*s)
    gen = torch.Generator("cuda").manual_seed(3)
    for call in range(2):                       # second call finds the first call's panel in place
        A = torch.randn(n, d, dtype=torch.float64, device="cuda", generator=gen)
        want = J.apply(A, None, "columnwise")                       # existing entry point
        got = torch.full((s, d), float("nan"), dtype=torch.float64, device="cuda")
        # row-major n x d == column-major d x n  ->  rowwise on the transposed view (dim = 2)
        rc = _noalloc(lib, 4242, 1000, 0, math.sqrt(1.0 / s), n, s, A, got, 2)
        assert rc == 0
        torch.cuda.synchronize()
        assert rel_fro(got.cpu().numpy(), want.cpu().numpy()) < TOL, call
    # the other layout (column-major operands, dim = 1), and a Cauchy sketch, after the above
    C = sk.CT(n, s, 2.0, lib.Context(7, 5))
    At = torch.randn(d, n, dtype=torch.float64, device="cuda", generator=gen)   # col-major n x d
    want = C.apply(At.t(), None, "columnwise")
    got = torch.empty(d, s, dtype=torch.float64, device="cuda")                 # col-major s x d
    assert _noalloc(lib, 7, 5, 1, 2.0 / s, n, s, At, got, 1) == 0
    torch.cuda.synchronize()
    assert rel_fro(got.t().cpu().numpy(), want.cpu().numpy()) < 1e-11   # Cauchy tails: cf. test_gpu_dense


def test_panel_reuse_inside_one_call_against_the_oracle(sk, oracle):
    """s = 4096 leaves room for 1024 input rows in the built-in array, so n = 3000 makes the call
    regenerate the panel three times in place; odd s and padded leading dimensions take the
    8-byte copy variants and the strided clear.  Checked against the CPU oracle as well as the
    regular entry point, and device memory in use must not move across a call."""
    import torch
    from libskylark_b200 import lib
    gen = torch.Generator("cuda").manual_seed(11)
    for (n, s, d, pad) in ((3000, 4096, 40, 0), (2500, 1023, 33, 5)):
        J = sk.JLT(n, s, lib.Context(99, 77))
        O = oracle.JLT(n, s, oracle.Context(99, 77))
        A = torch.randn(n, d, dtype=torch.float64, device="cuda", generator=gen)
        want = J.apply(A, None, "columnwise")
        want_cpu = O.apply(A.cpu().numpy(), "columnwise")
        buf = torch.full((s, d + pad), float("nan"), dtype=torch.float64, device="cuda")
        got = buf[:, :d]                      # row-major s x d with row stride d + pad
        torch.cuda.synchronize()
        for call in range(2):
            free0 = torch.cuda.mem_get_info()[0]
            assert _noalloc(lib, 99, 77, 0, math.sqrt(1.0 / s), n, s, A, got, 2) == 0
            torch.cuda.synchronize()
            if call == 1:                     # the first call may load kernels; the second may not move anything
                assert torch.cuda.mem_get_info()[0] == free0
            assert rel_fro(got.cpu().numpy(), want.cpu().numpy()) < TOL, (n, s, call)
            assert rel_fro(got.cpu().numpy(), want_cpu) < TOL, (n, s, call)
        if pad:
            assert bool(torch.isnan(buf[:, d:]).all())      # the padding is left alone


def test_size_limit_is_refused_with_a_message(sk):
    import torch
    from libskylark_b200 import lib
    A = torch.zeros(4, 8, dtype=torch.float64, device="cuda")
    SA = torch.zeros(300000, 8, dtype=torch.float64, device="cuda")
    rc = _noalloc(lib, 1, 0, 0, 1.0, 4, 300000, A, SA, 2)
    assert rc == 111
    info = ctypes.c_char_p()
    lib.lib.sl_get_exception_info(ctypes.byref(info))
    assert b"exceeds" in info.value
    assert _noalloc(lib, 1, 0, 0, 1.0, 4, 16, A, SA, 7) == 111          # bad direction
