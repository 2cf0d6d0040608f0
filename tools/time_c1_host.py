import os, sys, time, statistics
sys.path.insert(0, "/root/repo")
import numpy as np, torch
from libskylark_b200 import lib, sketch
n, d, s = 100000, 500, 2000
A = np.random.default_rng(0).standard_normal((n, d))
SA = np.empty((s, d))
S = sketch.CWT(n, s, lib.Context(123834))
S.apply(A, SA, "columnwise")
ts = []
for _ in range(7):
    t0 = time.perf_counter(); S.apply(A, SA, "columnwise"); ts.append(time.perf_counter() - t0)
At = torch.from_numpy(A); dA = torch.empty_like(At, device="cuda")
dA.copy_(At); torch.cuda.synchronize()
cs = []
for _ in range(5):
    t0 = time.perf_counter(); dA.copy_(At); torch.cuda.synchronize(); cs.append(time.perf_counter() - t0)
ref = sketch.CWT(n, s, lib.Context(123834)).apply(torch.from_numpy(A).cuda(), None, "columnwise").cpu().numpy()
print("C1 host operands: %.2f ms (plain pageable copy of A alone: %.2f ms); max |diff| vs device path %.1e; threads=%s" % (
    statistics.median(ts) * 1e3, statistics.median(cs) * 1e3, float(np.abs(SA - ref).max()), os.environ.get("SKB_COPY_THREADS", "default")))
