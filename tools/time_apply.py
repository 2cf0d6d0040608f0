"""Quick device-resident timing of one transform apply (tuning helper, not the bench).
usage: python tools/time_apply.py JLT|CWT|CWTF|FJLT|RFT [steps]"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from libskylark_b200 import lib, sketch

kind = sys.argv[1] if len(sys.argv) > 1 else "JLT"
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 3
dev = "cuda"
if kind == "JLT":
    n, d, s = 1_000_000, 1024, 4096
    A = torch.randn(n, d, dtype=torch.float64, device=dev)
    S = sketch.JLT(n, s, lib.Context(123834)); dim = "columnwise"
    work = 2.0 * s * n * d / 1e12; unit = "TFLOP/s"
elif kind in ("CWT", "CWTF"):
    n, d, s = 100_000, 500, 2000
    A = torch.randn(n, d, dtype=torch.float64, device=dev)
    if kind == "CWTF":
        A = A.t().contiguous().t()
    S = sketch.CWT(n, s, lib.Context(123834)); dim = "columnwise"
    work = (n * d * 8 + s * d * 8) / 1e9; unit = "GB/s"
elif kind == "FJLT":
    n, d, s = 2**20, int(os.environ.get("FJLT_D", "2048")), 8192
    lib.set_option("fjlt_fut", os.environ.get("FJLT_FUT", "dct"))
    A = torch.randn(d, n, dtype=torch.float64, device=dev).t()      # column-major n x d
    S = sketch.FJLT(n, s, lib.Context(123834)); dim = "columnwise"
    work = (n * d * 8 + s * d * 8) / 1e9; unit = "GB/s"
elif kind == "RFT":
    m, n, s = int(os.environ.get("RFT_M", "500000")), 784, 16384
    A = torch.rand(m, n, dtype=torch.float32, device=dev)
    S = sketch.GaussianRFT(n, s, 10.0, context=lib.Context(123834)); dim = "rowwise"
    work = 2.0 * m * n * s / 1e12; unit = "TFLOP/s"
SA = S.apply(A, None, dim)
torch.cuda.synchronize()
# L2 flush buffer for the small-input cases
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
ts = []
for _ in range(steps):
    flush.zero_()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); S.apply(A, SA, dim); e1.record()
    torch.cuda.synchronize()
    ts.append(e0.elapsed_time(e1))
best = min(ts)
print("%s env=%s: best %.3f ms  median %.3f ms -> %.2f %s" % (
    kind, {k: v for k, v in os.environ.items() if k.startswith(("SKB_", "FJLT_", "RFT_"))}, best,
    sorted(ts)[len(ts) // 2], work / (best * 1e-3), unit))
