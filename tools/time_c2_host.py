"""JLT end to end on PAGEABLE numpy operands (the reference's calling convention), a quarter of C2's rows.
python tools/time_c2_host.py   (SKB_COPY_THREADS=0: plain copies)"""
import os, sys, time, statistics
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from libskylark_b200 import lib, sketch
n, d, s = 250_000, 1024, 4096
A = np.asfortranarray(np.random.default_rng(0).standard_normal((n, d)))     # column-major n x d (the reference's layout)
SA = np.empty((s, d), order="F")
S = sketch.JLT(n, s, lib.Context(123834))
S.apply(A, SA, "columnwise")
ts = []
for _ in range(3):
    t0 = time.perf_counter(); S.apply(A, SA, "columnwise"); ts.append(time.perf_counter() - t0)
ref = sketch.JLT(n, s, lib.Context(123834)).apply(torch.from_numpy(A.T.copy()).cuda().t(), None, "columnwise").cpu().numpy()
err = float(np.linalg.norm(SA - ref) / np.linalg.norm(ref))
ms = statistics.median(ts) * 1e3
print("JLT %dx%d -> %d from pageable numpy: %.1f ms = %.2f TFLOP/s; rel. diff vs device path %.1e; threads=%s" % (
    n, d, s, ms, 2.0 * s * n * d / ms / 1e9, err, os.environ.get("SKB_COPY_THREADS", "default")))
