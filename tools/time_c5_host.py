"""Gaussian RFT end to end on PAGEABLE numpy operands: 100k x 784 fp64 points -> 4096 features (3.3 GB back; numpy operands are fp64, as in the reference).
python tools/time_c5_host.py   (SKB_COPY_THREADS=0: plain copies)"""
import os, sys, time, statistics
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from libskylark_b200 import lib, sketch
m, n, s = 100_000, 784, 4096
A = np.random.default_rng(0).random((m, n))
SA = np.empty((m, s))
S = sketch.GaussianRFT(n, s, 10.0, context=lib.Context(123834))
S.apply(A, SA, "rowwise")
ts = []
for _ in range(3):
    t0 = time.perf_counter(); S.apply(A, SA, "rowwise"); ts.append(time.perf_counter() - t0)
ref = sketch.GaussianRFT(n, s, 10.0, context=lib.Context(123834)).apply(torch.from_numpy(A).cuda(), None, "rowwise").cpu().numpy()
ms = statistics.median(ts) * 1e3
print("RFT %dx%d -> %d from / to pageable numpy: %.1f ms (%.2f GB moved, %.1f GB/s); max |diff| vs device path %.1e; threads=%s" % (
    m, n, s, ms, (A.nbytes + SA.nbytes) / 1e9, (A.nbytes + SA.nbytes) / ms / 1e6, float(np.abs(SA - ref).max()),
    os.environ.get("SKB_COPY_THREADS", "default")))
