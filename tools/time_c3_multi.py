"""C3 (FJLT-WHT 2^20 x 2048 -> 8192) on row shards: exchange variants timed under torchrun.
python -m torch.distributed.run --nproc-per-node N tools/time_c3_multi.py"""
import os, sys, statistics
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, torch.distributed as dist
rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", rank)))
dist.init_process_group("nccl", device_id=torch.device("cuda", torch.cuda.current_device()))
from libskylark_b200 import lib, sketch, dist as sd
n, d, s = 2 ** 20, 2048, 8192
nloc = n // world
lib.set_option("fjlt_fut", "wht"); S = sketch.FJLT(n, s, lib.Context(123834)); lib.set_option("fjlt_fut", "dct")
A = torch.randn(d, nloc, dtype=torch.float64, device="cuda").t()
SA = torch.empty(d, s, dtype=torch.float64, device="cuda").t()

def timed(fn, reps=7):
    fn(); fn(); torch.cuda.synchronize(); dist.barrier()
    ts = []
    for _ in range(reps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize()
        t = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX); ts.append(float(t.item())); dist.barrier()
    return statistics.median(ts)

res = {}
res["kernel only"] = timed(lambda: S.apply_shard(A, SA, 0, rank * nloc))
res["kernel + all-reduce"] = timed(lambda: sd.apply_distributed(S, A, SA, 0, rank * nloc, reduce="allreduce"))
res["kernel + reduce-scatter"] = timed(lambda: sd.apply_distributed(S, A, SA, 0, rank * nloc, reduce="reduce_scatter"))
res["fused peer-store reduce-scatter"] = timed(lambda: sd.fjlt_wht_fused_reduce_scatter(S, A, rank * nloc))
if rank == 0:
    for k, v in res.items():
        print("world %d  %-34s %.3f ms" % (world, k, v), flush=True)
dist.barrier(); dist.destroy_process_group()
