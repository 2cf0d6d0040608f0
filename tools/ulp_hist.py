"""Distribution of the ulp distance between device-generated variates and the CPU oracle."""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
from util import ulp_diff
from libskylark_b200 import lib, sketch
from oracle import oracle as o
n, s = 4000, 500
for name, mk, omk in (("normal", lambda c: sketch.JLT(n, s, c), lambda c: o.JLT(n, s, c)),
                      ("cauchy", lambda c: sketch.CT(n, s, 1.0, c), lambda c: o.CT(n, s, 1.0, c))):
    d = ulp_diff(mk(lib.Context(1)).export_dense(), omk(o.Context(1)).realize()).ravel()
    print(name, "n=%d" % d.size, "hist(0..8):", np.bincount(np.minimum(d, 8), minlength=9).tolist(), "max", int(d.max()))
