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
    g = {"note": "derived from oracle/skylark_oracle.c; parity unpinned (no reference vectors exist)"}
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
    with open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden.json"), "w") as fh:
        json.dump(g, fh, indent=1)
    print("wrote golden.json")


if __name__ == "__main__":
    main()
