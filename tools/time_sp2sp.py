"""Time sparse -> sparse CWT on device-resident CSC (C4-like shard)."""
import sys, time
import numpy as np, torch
sys.path.insert(0, ".")
from libskylark_b200 import lib, sketch as sk

n, d, s, nnz_col = 1 << 20, 65536, 20000, 100
if len(sys.argv) > 1:
    n, d, s, nnz_col = (int(x) for x in sys.argv[1:5])
g = torch.Generator("cuda").manual_seed(0)
indices = torch.randint(0, n, (d, nnz_col), device="cuda", generator=g, dtype=torch.int32)
indices, _ = torch.sort(indices, dim=1)
indptr = torch.arange(0, d * nnz_col + 1, nnz_col, device="cuda", dtype=torch.int32)
vals = torch.randn(d * nnz_col, device="cuda", dtype=torch.float64, generator=g)
A = lib.DeviceCSC(indptr, indices.reshape(-1).contiguous(), vals, (n, d))
S = sk.CWT(n, s, lib.Context(1))
for name, mk in (("sparse", lambda: lib.DeviceCSC.empty((s, d))),
                 ("dense", lambda: torch.empty(d, s, device="cuda", dtype=torch.float64).t())):
    out = mk()
    S.apply(A, out, 0)
    torch.cuda.synchronize()
    ts = []
    for _ in range(5):
        out = mk()
        torch.cuda.synchronize(); t0 = time.perf_counter()
        S.apply(A, out, 0)
        torch.cuda.synchronize(); ts.append(time.perf_counter() - t0)
    nnz_out = out.values.numel() if name == "sparse" else s * d
    print(f"{name}: best {min(ts)*1e3:.2f} ms  (nnz in {d*nnz_col}, out entries {nnz_out})")
