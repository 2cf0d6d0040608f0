"""Times the C4 shard (CWT on a device CSC, 1.25e8 nnz -> dense 20000 x 1e5) on one GPU."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from libskylark_b200 import lib, sketch
dev = "cuda"
n_total, n_loc, d, s, nnz_col = 10_000_000, 1_250_000, 100_000, 20_000, 1250
g = torch.Generator(device=dev).manual_seed(0)
indptr = (torch.arange(d + 1, device=dev, dtype=torch.int64) * nnz_col).to(torch.int32)
indices = torch.randint(0, n_loc, (d * nnz_col,), device=dev, dtype=torch.int32, generator=g)
values = torch.randn(d * nnz_col, dtype=torch.float64, device=dev, generator=g)
A = lib.DeviceCSC(indptr, indices, values, (n_loc, d))
S = sketch.CWT(n_total, s, lib.Context(123834))
SA = torch.empty(d, s, dtype=torch.float64, device=dev).t()
S.apply_shard(A, SA, "columnwise", 0); torch.cuda.synchronize()
ts = []
for _ in range(3):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); S.apply_shard(A, SA, "columnwise", 0); e1.record(); torch.cuda.synchronize()
    ts.append(e0.elapsed_time(e1))
nbytes = d * nnz_col * 12 + (d + 1) * 4 + s * d * 8
print("CSC shard: best %.3f ms -> %.1f GB/s" % (min(ts), nbytes / min(ts) / 1e6))
