"""FJLT on the GPU against the oracle: DCT-II mode (what the reference computes) and the
Walsh-Hadamard mode named by BASELINE.json."""
import numpy as np
import pytest

from util import rel_fro

pytestmark = pytest.mark.gpu
TOL = 1e-12


@pytest.mark.parametrize("fut", ["dct", "wht"])
@pytest.mark.parametrize("shape", [(64, 16, 5), (1024, 100, 33), (4096, 512, 8), (16384, 300, 9),
                                   (65536, 9000, 5), (2048, 64, 150)])
def test_fjlt_power_of_two(sk, oracle, fut, shape):
    from libskylark_b200 import lib
    n, s, d = shape
    lib.set_option("fjlt_fut", fut)
    try:
        S = sk.FJLT(n, s, lib.Context(17))
    finally:
        lib.set_option("fjlt_fut", "dct")
    O = oracle.FJLT(n, s, oracle.Context(17))
    A = np.random.default_rng(5).standard_normal((n, d))
    want = O.apply(A, "columnwise", fut=fut)
    assert rel_fro(S.apply(np.asfortranarray(A), None, "columnwise"), want) < TOL
    assert rel_fro(S.apply(A.copy(), None, 0), want) < TOL
    B = np.ascontiguousarray(A.T)
    assert rel_fro(S.apply(B, None, "rowwise"), want.T) < TOL


@pytest.mark.parametrize("shape", [(8192, 700, 1000), (32768, 9000, 450), (16384, 100, 149),
                                   (8192, 50, 1), (16384, 50, 2), (65536, 8192, 3)])
def test_wht_many_columns_per_cta(sk, oracle, shape):
    # more columns than SMs: every CTA of the streaming kernel walks several columns, so its
    # ring of shared buffers and both butterfly groups wrap across column boundaries
    # (n = 8192: one segment per column, the groups alternate columns)
    from libskylark_b200 import lib
    n, s, d = shape
    lib.set_option("fjlt_fut", "wht")
    try:
        S = sk.FJLT(n, s, lib.Context(23))
    finally:
        lib.set_option("fjlt_fut", "dct")
    O = oracle.FJLT(n, s, oracle.Context(23))
    A = np.asfortranarray(np.random.default_rng(8).standard_normal((n, d)))
    want = O.apply(A, "columnwise", fut="wht")
    for _ in range(3):            # repeated: the hand-off between warps is timing dependent
        assert rel_fro(S.apply(A, None, "columnwise"), want) < TOL
    import torch
    odd = torch.zeros(d, n + 1, dtype=torch.float64, device="cuda").t()[1:, :]   # columns only 8-byte aligned
    odd.copy_(torch.from_numpy(A))
    assert rel_fro(S.apply(odd, None, "columnwise").cpu().numpy(), want) < TOL


def test_fjlt_general_n_dct(sk, oracle):
    from libskylark_b200 import lib
    n, s, d = 300, 40, 7
    S, O = sk.FJLT(n, s, lib.Context(3)), oracle.FJLT(n, s, oracle.Context(3))
    A = np.random.default_rng(6).standard_normal((n, d))
    assert rel_fro(S.apply(np.asfortranarray(A), None, 0), O.apply(A, "columnwise")) < TOL


def test_fjlt_wht_needs_power_of_two(sk):
    from libskylark_b200 import errors, lib
    lib.set_option("fjlt_fut", "wht")
    try:
        S = sk.FJLT(300, 40, lib.Context(3))
    finally:
        lib.set_option("fjlt_fut", "dct")
    with pytest.raises(errors.LowerLayerError):
        S.apply(np.zeros((300, 4), order="F"), None, 0)


def test_wht_row_shards_add_up(sk, oracle):
    """H_N = H_P (x) H_{N/P}: partial sketches of 8192-aligned row blocks sum to the full one
    (BASELINE.json config 3: FJLT row-sharded over GPUs with a reduce)."""
    import torch
    from libskylark_b200 import errors, lib
    n, s, d = 65536, 700, 24
    lib.set_option("fjlt_fut", "wht")
    try:
        S = sk.FJLT(n, s, lib.Context(5))
    finally:
        lib.set_option("fjlt_fut", "dct")
    O = oracle.FJLT(n, s, oracle.Context(5))
    A = torch.randn(d, n, dtype=torch.float64, device="cuda").t()        # column-major n x d
    want = O.apply(A.cpu().numpy(), "columnwise", fut="wht")
    full = S.apply(A, None, "columnwise")
    assert rel_fro(full.cpu().numpy(), want) < TOL
    for cuts in ((0, 32768, 65536), (0, 8192, 24576, 57344, 65536)):
        parts = sum(S.apply_shard(A[a:b], None, 0, a) for a, b in zip(cuts, cuts[1:]))
        assert rel_fro(parts.cpu().numpy(), want) < TOL
    Ar = A.contiguous()                                                   # row-major: transposed path
    parts = sum(S.apply_shard(Ar[a:b], None, 0, a) for a, b in ((0, 16384), (16384, 65536)))
    assert rel_fro(parts.cpu().numpy(), want) < TOL
    with pytest.raises(errors.LowerLayerError):
        S.apply_shard(A[:1000], None, 0, 0)                               # not 8192-aligned
    Sd = sk.FJLT(n, s, lib.Context(5))                                    # DCT mode cannot
    with pytest.raises(errors.LowerLayerError):
        Sd.apply_shard(A[:32768], None, 0, 0)


@pytest.mark.parametrize("fut", ["dct", "wht"])
def test_fjlt_at_the_baseline_length(sk, oracle, fut):
    """Config C3's N = 2^20 and s = 8192 on a few columns: the DCT runs its 2^19-point split
    and the Walsh-Hadamard kernel walks 128 segments per column, as in the benchmark."""
    import torch
    from libskylark_b200 import lib
    n, s, d = 2 ** 20, 8192, 4
    lib.set_option("fjlt_fut", fut)
    try:
        S = sk.FJLT(n, s, lib.Context(123834))
    finally:
        lib.set_option("fjlt_fut", "dct")
    O = oracle.FJLT(n, s, oracle.Context(123834))
    A = torch.randn(d, n, dtype=torch.float64, device="cuda", generator=torch.Generator("cuda").manual_seed(4)).t()
    want = O.apply(A.cpu().numpy(), "columnwise", fut=fut)
    got = S.apply(A, None, "columnwise")
    assert rel_fro(got.cpu().numpy(), want) < TOL
    got_r = S.apply(A.contiguous(), None, "columnwise")          # row-major operand
    assert rel_fro(got_r.cpu().numpy(), want) < TOL


def test_fjlt_long_and_wide_shapes(sk, oracle):
    """Shapes whose tile counts exceed the 65535 limit of a launch grid's y extent: a rows-layout
    operand with N = 2^21 (65536 row tiles in the transpose) and N = 1024 with 70000 columns
    (one batch of more than 65535 columns in the small-N DCT)."""
    import torch
    from libskylark_b200 import lib
    for fut in ("dct", "wht"):
        n, s, d = 2 ** 21, 64, 2
        lib.set_option("fjlt_fut", fut)
        try:
            S = sk.FJLT(n, s, lib.Context(8))
        finally:
            lib.set_option("fjlt_fut", "dct")
        O = oracle.FJLT(n, s, oracle.Context(8))
        A = torch.randn(n, d, dtype=torch.float64, device="cuda", generator=torch.Generator("cuda").manual_seed(6))
        got = S.apply(A, None, "columnwise")                     # row-major n x d: the transposing path
        assert rel_fro(got.cpu().numpy(), O.apply(A.cpu().numpy(), "columnwise", fut=fut)) < TOL
    n, s, d = 1024, 16, 70000
    S, O = sk.FJLT(n, s, lib.Context(9)), oracle.FJLT(n, s, oracle.Context(9))
    A = torch.randn(d, n, dtype=torch.float64, device="cuda", generator=torch.Generator("cuda").manual_seed(7)).t()
    got = S.apply(A, None, "columnwise")
    assert rel_fro(got.cpu().numpy(), O.apply(A.cpu().numpy(), "columnwise")) < TOL
