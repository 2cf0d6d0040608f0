"""One or two applies of each BASELINE config, for `ncu` captures (kernel filters on the command line).
python tools/profile_kernels.py [c1 c1f c2 c3wht c3dct c4 c5]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from libskylark_b200 import lib, sketch

which = sys.argv[1:] or ["c1", "c1f", "c2", "c3wht", "c3dct", "c4"]
dev = "cuda"
SEED = 123834
for w in which:
    if w in ("c1", "c1f"):
        n, d, s = 100_000, 500, 2000
        A = torch.randn(n, d, dtype=torch.float64, device=dev)
        if w == "c1f":
            A = A.t().contiguous().t()
        S = sketch.CWT(n, s, lib.Context(SEED))
        SA = S.apply(A, None, "columnwise"); S.apply(A, SA, "columnwise")
    elif w == "c2":
        n, d, s = 65536, 1024, 4096            # 8 panels of the C2 shape (same tiles, same panel size)
        A = torch.randn(n, d, dtype=torch.float64, device=dev)
        S = sketch.JLT(1_000_000, s, lib.Context(SEED))
        SA = S.apply_shard(A, None, "columnwise", 0); S.apply_shard(A, SA, "columnwise", 0)
    elif w in ("c3wht", "c3dct"):
        n, s, d = 2 ** 20, 8192, 592
        lib.set_option("fjlt_fut", w[2:])
        S = sketch.FJLT(n, s, lib.Context(SEED))
        lib.set_option("fjlt_fut", "dct")
        A = torch.randn(d, n, dtype=torch.float64, device=dev).t()
        SA = S.apply(A, None, "columnwise"); S.apply(A, SA, "columnwise")
    elif w == "c4":
        n_total, n_loc, d, s, nnz_col = 10_000_000, 1_250_000, 25_000, 20_000, 1250     # a quarter of the shard's columns
        indptr = (torch.arange(d + 1, device=dev, dtype=torch.int64) * nnz_col).to(torch.int32)
        indices = torch.randint(0, n_loc, (d * nnz_col,), device=dev, dtype=torch.int32)
        values = torch.randn(d * nnz_col, dtype=torch.float64, device=dev)
        A = lib.DeviceCSC(indptr, indices, values, (n_loc, d))
        S = sketch.CWT(n_total, s, lib.Context(SEED))
        SA = torch.empty(d, s, dtype=torch.float64, device=dev).t()
        S.apply_shard(A, SA, "columnwise", 0); S.apply_shard(A, SA, "columnwise", 0)
    elif w == "c5":
        m, n, s = 65536, 784, 16384
        S = sketch.GaussianRFT(n, s, 10.0, context=lib.Context(SEED))
        A = torch.rand(m, n, dtype=torch.float32, device=dev)
        SA = S.apply(A, None, "rowwise"); S.apply(A, SA, "rowwise")
    torch.cuda.synchronize()
    del A, SA, S
    torch.cuda.empty_cache()
print("profile_kernels: done", which)
