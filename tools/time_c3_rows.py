"""C3 timing with a ROW-major operand (the numpy / torch default), both modes.  python tools/time_c3_rows.py [d] [fut...]"""
import os, sys, statistics
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from libskylark_b200 import lib, sketch
d = int(sys.argv[1]) if len(sys.argv) > 1 else 2048
futs = sys.argv[2:] or ["dct", "wht"]
n, s = 2 ** 20, 8192
A = torch.randn(n, d, dtype=torch.float64, device="cuda")
for fut in futs:
    lib.set_option("fjlt_fut", fut)
    S = sketch.FJLT(n, s, lib.Context(123834))
    lib.set_option("fjlt_fut", "dct")
    SA = S.apply(A, None, "columnwise")
    torch.cuda.synchronize()
    ts = []
    for _ in range(5):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); S.apply(A, SA, "columnwise"); e1.record(); torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    ms = statistics.median(ts)
    print("C3 rows-layout %s d=%d: %.3f ms (min %.3f)  %.0f GB/s" % (fut, d, ms, min(ts), (n * d * 8 + s * d * 8) / ms / 1e6), flush=True)
