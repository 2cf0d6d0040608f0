"""C4 (CWT of a row-sharded CSC 10M x 100k, 1.25e8 nnz per GPU -> 20000) with the sparse exchange:
one-shot exchange + merge against the pipelined rounds.
python -m torch.distributed.run --nproc-per-node N tools/time_c4_multi.py"""
import os, sys, statistics
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, torch.distributed as dist
rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", rank)))
dist.init_process_group("nccl", device_id=torch.device("cuda", torch.cuda.current_device()))
from libskylark_b200 import lib, sketch, dist as sd
n_loc, d, s, nnz_col = 1_250_000, 100_000, 20_000, 1250
dev = "cuda"
g = torch.Generator(dev).manual_seed(1000 + rank)
indptr = (torch.arange(d + 1, device=dev, dtype=torch.int64) * nnz_col).to(torch.int32)
indices = torch.randint(0, n_loc, (d * nnz_col,), device=dev, dtype=torch.int32, generator=g)
values = torch.randn(d * nnz_col, dtype=torch.float64, device=dev, generator=g)
A = lib.DeviceCSC(indptr, indices, values, (n_loc, d))
S = sketch.CWT(n_loc * world, s, lib.Context(123834))

def timed(fn, reps=5):
    fn(); fn(); torch.cuda.synchronize(); dist.barrier()
    ts = []
    for _ in range(reps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize()
        t = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX); ts.append(float(t.item())); dist.barrier()
    return statistics.median(ts)

res, outs = {}, {}
def run(tag, nblk):
    os.environ["SKB_SPARSE_PIPE"] = str(nblk)
    res[tag] = timed(lambda: outs.__setitem__(tag, sd.apply_distributed_sparse(S, A, rank * n_loc, None)[0]))
res["local sketch only"] = timed(lambda: S.apply_shard(A, lib.DeviceCSC.empty((s, d), "csc", dev), 0, rank * n_loc))
part = S.apply_shard(A, lib.DeviceCSC.empty((s, d), "csc", dev), 0, rank * n_loc)
res["exchange only (one shot)"] = timed(lambda: outs.__setitem__("x", sd.exchange_sparse_partials(part.indptr, part.indices, part.values, None, reorder=False)))
vptr, ix, vv, (c0, nc) = outs["x"]
res["merge only (one shot)"] = timed(lambda: sketch.merge_segments(vptr, ix, vv, world, (s, nc)))
run("one shot", 1)
for nb in (2, 4, 8):
    run("pipelined, %d rounds" % nb, nb)
os.environ["SKB_DIST_DEBUG"] = "1"; os.environ["SKB_SPARSE_PIPE"] = "4"
torch.cuda.synchronize(); dist.barrier()
sd.apply_distributed_sparse(S, A, rank * n_loc, None); torch.cuda.synchronize(); dist.barrier()
sd.apply_distributed_sparse(S, A, rank * n_loc, None); torch.cuda.synchronize(); dist.barrier()
os.environ.pop("SKB_DIST_DEBUG")
a, b = outs["one shot"], outs["pipelined, 4 rounds"]
same = bool(torch.equal(a.indptr, b.indptr) and torch.equal(a.indices, b.indices))
err = float((a.values - b.values).abs().max()) if same else float("nan")
flag = torch.tensor([1.0 if same and err == 0.0 else 0.0], device=dev); dist.all_reduce(flag, op=dist.ReduceOp.MIN)
if rank == 0:
    for k, v in res.items():
        print("world %d  %-26s %.3f ms" % (world, k, v), flush=True)
    print("pipelined result identical to the one-shot result on every rank:", bool(flag.item()), flush=True)
dist.barrier(); dist.destroy_process_group()
