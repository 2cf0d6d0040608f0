"""Writes tests/golden/golden.json from the CPU oracle (oracle/skylark_oracle.c).

The reference holds no golden vectors for the sketch path and cannot be built here
(SURVEY.md section 8c), so these fixtures are derived, not reference-pinned ("parity
unpinned").  Only the Threefry2x64-20 vectors are published by Random123 itself.
Run from the repo root:  python tests/golden/make_golden.py
"""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import oracle as o  # noqa: E402


def hexs(a):
    return ["%016x" % int(x) for x in a]


def main():
    g = {"note": "derived from oracle/skylark_oracle.c; parity unpinned (no reference vectors exist)",
         "libm": "log / cos / tan / pow inside the variates correctly rounded (oracle libm mode 1: the "
                 "behaviour of the reference era's glibc <= 2.27), not this box's glibc"}
    o.set_libm(True)
    g["threefry2x64_20_published"] = [
        {"ctr": ["0" * 16, "0" * 16], "key": ["0" * 16, "0" * 16],
         "out": ["c2b6e3a8c2c69865", "6f81ed42f350084d"]},
        {"ctr": ["f" * 16, "f" * 16], "key": ["f" * 16, "f" * 16],
         "out": ["e02cb7c4d95d277a", "d06633d0893b8b68"]},
        {"ctr": ["243f6a8885a308d3", "13198a2e03707344"],
         "key": ["a4093822299f31d0", "082efa98ec4e6c89"],
         "out": ["263c7d30bb0f0af1", "56be8361d3311526"]}]
    g["threefry2x64_13_derived"] = []
    for e in g["threefry2x64_20_published"]:
        c = [int(x, 16) for x in e["ctr"]]
        k = [int(x, 16) for x in e["key"]]
        g["threefry2x64_13_derived"].append(
            {"ctr": e["ctr"], "key": e["key"], "out": hexs(o.threefry2x64(13, c, k))})
    cases = []
    for seed in (0, 123834, -7):
        for base in (0, 1000003, 2**40 + 17):
            a, b = o.raw_draws(seed, base, 8)
            cases.append({"seed": seed, "base": base, "first": hexs(a), "second": hexs(b),
                          "uniform_int_S1999": [int(x) for x in o.uniform_int(seed, base, 8, 1999)],
                          "rademacher_160": o.rademacher(seed, base, 8, 160).tolist(),
                          "rademacher_155": o.rademacher(seed, base, 8, 155).tolist(),
                          "normal": [x.hex() for x in o.normal(seed, base, 8).tolist()],
                          "cauchy": [x.hex() for x in o.cauchy(seed, base, 8).tolist()],
                          "exponential": [x.hex() for x in o.exponential(seed, base, 8).tolist()],
                          "uniform_real_0_2pi": [x.hex() for x in o.uniform_real(
                              seed, base, 8, 0.0, 2 * np.pi).tolist()]})
    g["streams"] = cases
    tr = []
    for seed in (0, 123834):
        c = o.Context(seed)
        t = o.CWT(10, 6, c)
        m = o.MMT(7, 3, c)
        w = o.WZT(5, 4, 1.5, c)
        j = o.JLT(4, 3, c)
        r = o.GaussianRFT(3, 5, 2.0, c)
        f = o.FJLT(16, 4, c)
        tr.append({"seed": seed,
                   "cwt_row_idx": [int(x) for x in t.row_idx], "cwt_row_value": t.row_value.tolist(),
                   "mmt_row_idx": [int(x) for x in m.row_idx],
                   "mmt_row_value": [x.hex() for x in m.row_value.tolist()],
                   "wzt_row_value": [x.hex() for x in w.row_value.tolist()],
                   "jlt_S": [x.hex() for x in j.realize().ravel(order="F").tolist()],
                   "rft_shifts": [x.hex() for x in r.shifts.tolist()],
                   "fjlt_D": f.D.tolist(), "fjlt_samples": [int(x) for x in f.samples],
                   "counter_after": c.counter})
    g["transforms_in_sequence"] = tr
    # end results on a fixed small input: A[i, j] = ((7 i + 3 j) mod 11 - 5) / 4 (exact in binary)
    n, d = 64, 3
    A = (((7 * np.arange(n)[:, None] + 3 * np.arange(d)[None, :]) % 11) - 5) / 4.0
    Asp = A * (((np.arange(n)[:, None] + 2 * np.arange(d)[None, :]) % 3) == 0)       # about a third kept
    ap = {"n": n, "d": d, "seed": 123834, "input": "A[i,j] = ((7i+3j) mod 11 - 5)/4; sparse: kept where (i+2j) mod 3 == 0"}

    def hx(M):
        return [x.hex() for x in np.asarray(M, dtype=np.float64).ravel(order="F").tolist()]
    import scipy.sparse
    c = o.Context(123834)
    cw = o.CWT(n, 5, c)
    ap["cwt_columnwise"] = hx(cw.apply(A, "columnwise"))
    ap["cwt_rowwise_of_transpose"] = hx(cw.apply(np.ascontiguousarray(A.T), "rowwise"))
    sp = cw.apply_csc_sparse(scipy.sparse.csc_matrix(Asp), "columnwise")
    sp.sort_indices()
    ap["cwt_sparse_columnwise"] = {"indptr": sp.indptr.tolist(), "indices": sp.indices.tolist(),
                                   "data": [x.hex() for x in sp.data.tolist()]}
    ap["mmt_columnwise"] = hx(o.MMT(n, 4, c).apply(A, "columnwise"))
    ap["wzt_columnwise"] = hx(o.WZT(n, 4, 1.5, c).apply(A, "columnwise"))
    ap["jlt_columnwise"] = hx(o.JLT(n, 6, c).apply(A, "columnwise"))
    ap["ct_columnwise"] = hx(o.CT(n, 6, 2.0, c).apply(A, "columnwise"))
    ap["gaussian_rft_columnwise"] = hx(o.GaussianRFT(n, 7, 3.0, c).apply(A, "columnwise"))
    ap["laplacian_rft_columnwise"] = hx(o.LaplacianRFT(n, 7, 3.0, c).apply(A, "columnwise"))
    fj = o.FJLT(n, 8, c)
    ap["fjlt_dct_columnwise"] = hx(fj.apply(A, "columnwise", fut="dct"))
    ap["fjlt_wht_columnwise"] = hx(fj.apply(A, "columnwise", fut="wht"))
    ap["counter_after"] = c.counter
    g["applies"] = ap
    with open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden.json"), "w") as fh:
        json.dump(g, fh, indent=1)
    print("wrote golden.json")


if __name__ == "__main__":
    main()
