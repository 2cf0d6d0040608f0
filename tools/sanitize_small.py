"""Small run of every kernel family (odd shapes included) for compute-sanitizer."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, scipy.sparse, torch
from libskylark_b200 import lib, sketch
rng = np.random.default_rng(0)
c = lib.Context(7)
# hash: rows layout (C order), cols layout (F order), odd widths, CSC both directions
for n, s, d in ((300, 17, 33), (1000, 200, 64), (257, 3000, 5)):
    S = sketch.CWT(n, s, c); A = rng.standard_normal((n, d))
    S.apply(A.copy(), None, 0); S.apply(np.asfortranarray(A), None, 0); S.apply(np.ascontiguousarray(A.T), None, 1)
    M = sketch.MMT(n, s, c)
    Asp = scipy.sparse.random(n, d, 0.2, format="csc", random_state=1); M.apply(Asp, None, 0)
    Bsp = scipy.sparse.random(d, n, 0.2, format="csc", random_state=2); M.apply(Bsp, None, 1)
# dense: JLT / CT / RFT, fp64 and fp32, both layouts, ragged tiles, several panels
for n, s, d in ((130, 70, 45), (2050, 129, 131)):
    J = sketch.JLT(n, s, c); A = rng.standard_normal((n, d))
    J.apply(A.copy(), None, 0); J.apply(np.asfortranarray(A), None, 0)
    R = sketch.GaussianRFT(n, s, 2.0, context=c)
    R.apply(rng.random((d, n)), None, 1)
    X32 = torch.rand(700, n, dtype=torch.float32, device="cuda"); R.apply(X32, None, 1)
    Xt = torch.randn(n, 300, dtype=torch.float64, device="cuda"); J.apply(Xt, None, 0)
os.environ["SKB_PANEL_MIB"] = "1"
# FJLT: direct, WHT fast path (N >= 8192) incl. shards, DCT radix-4 (1024) and radix-16 (16384) engines
for fut in ("dct", "wht"):
    lib.set_option("fjlt_fut", fut)
    for n, s, d in ((96, 10, 3), (1024, 50, 7), (16384, 300, 5)):
        if fut == "wht" and n & (n - 1):
            continue
        F = sketch.FJLT(n, s, lib.Context(3))
        A = rng.standard_normal((n, d))
        F.apply(np.asfortranarray(A), None, 0); F.apply(A.copy(), None, 0)
        if fut == "wht" and n >= 16384:
            At = torch.tensor(np.asfortranarray(A).T.copy(), device="cuda").t()
            F.apply_shard(At[:8192], None, 0, 0); F.apply_shard(At[8192:], None, 0, 8192)
lib.set_option("fjlt_fut", "dct")
# allocation-free entry point
import ctypes, math
fn = lib.lib.sl_dense_sketch_apply_noalloc
fn.argtypes = (ctypes.c_int, ctypes.c_uint64, ctypes.c_int, ctypes.c_double, ctypes.c_int64, ctypes.c_int64,
               ctypes.c_void_p, ctypes.c_int64, ctypes.c_int64, ctypes.c_void_p, ctypes.c_int64, ctypes.c_int, ctypes.c_void_p)
A = torch.randn(500, 33, dtype=torch.float64, device="cuda"); SA = torch.empty(40, 33, dtype=torch.float64, device="cuda")
assert fn(1, 0, 0, math.sqrt(1 / 40), 500, 40, A.data_ptr(), A.stride(0), 33, SA.data_ptr(), SA.stride(0), 2, None) == 0
torch.cuda.synchronize()
print("sanitize_small: done")
