"""Distribution of the ulp distance between device-generated variates and the CPU oracle
(correctly rounded libm mode, and this box's glibc).  Needs a GPU.  python tools/ulp_stats.py"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
from util import ulp_diff
from libskylark_b200 import lib, sketch
from oracle import oracle as o

n, s = 2000, 1000
for name, mk, omk in (("normal (JLT entries)", lambda c: sketch.JLT(n, s, c), lambda c: o.JLT(n, s, c)),
                      ("cauchy (CT entries)", lambda c: sketch.CT(n, s, 1.0, c), lambda c: o.CT(n, s, 1.0, c))):
    got = mk(lib.Context(123834, 5)).export_dense()
    for mode, label in ((True, "oracle libm correctly rounded"), (False, "oracle libm = this box's glibc")):
        o.set_libm(mode)
        d = ulp_diff(got, omk(o.Context(123834, 5)).realize()).ravel()
        h = np.bincount(np.minimum(d, 8).astype(int), minlength=9)
        print("%-22s vs %-32s n=%d  identical %.5f%%  <=2ulp %.6f%%  max %d  hist(0..8+) %s" %
              (name, label, d.size, 100.0 * (d == 0).mean(), 100.0 * (d <= 2).mean(), int(d.max()), h.tolist()))
o.set_libm(True)
