"""Worker of tests/test_gpu_multi.py: one process per GPU under torchrun (NCCL).  Every check is
"distributed == local for the same seed" (tests/unit/DenseSketchApplyElementalTest.cpp:28-106):
the sharded result must match the single-GPU result of the same transform on the whole matrix,
and that one the CPU oracle."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import numpy as np
import torch
import torch.distributed as dist


def rel(a, b):
    return float((a - b).norm() / b.norm())


def main():
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", rank)))
    dist.init_process_group("nccl", device_id=torch.device("cuda", torch.cuda.current_device()))
    from libskylark_b200 import lib, sketch
    from libskylark_b200 import dist as sd
    from oracle import oracle as o
    dev = "cuda"
    gen = torch.Generator(dev).manual_seed(5)            # same data on every rank
    checks = {}

    # --- hash and dense transforms: row shards + all-reduce / reduce-scatter / reduce
    n, d, s = 8192 * world * 2, 64 * world, 96
    A = torch.randn(n, d, dtype=torch.float64, device=dev, generator=gen)
    off, cnt = sd.shard_range(n, rank, world)
    for name, mk, omk in (("CWT", lambda c: sketch.CWT(n, s, c), lambda c: o.CWT(n, s, c)),
                          ("JLT", lambda c: sketch.JLT(n, s, c), lambda c: o.JLT(n, s, c))):
        S = mk(lib.Context(31))
        full = S.apply(A, None, 0)
        if rank == 0:
            want = torch.from_numpy(omk(o.Context(31)).apply(A.cpu().numpy(), "columnwise")).to(dev)
            checks[name + " local vs oracle"] = rel(full, want)
        got = S.apply_distributed(A[off:off + cnt], None, 0, off, reduce="allreduce")
        checks[name + " allreduce"] = rel(got, full)
        got = S.apply_distributed(A[off:off + cnt].contiguous(), None, 0, off, reduce="reduce_scatter")
        rows = s // world
        checks[name + " reduce_scatter"] = rel(got, full[rank * rows:(rank + 1) * rows])
        got = S.apply_distributed(A[off:off + cnt], None, 0, off, reduce="reduce", dst=0)
        if rank == 0:
            checks[name + " reduce"] = rel(got, full)

    # --- FJLT, Walsh-Hadamard mode: row shards, overlapped exchange (both collectives)
    Ac = A.t().contiguous().t()                          # column-major
    lib.set_option("fjlt_fut", "wht")
    F = sketch.FJLT(n, s, lib.Context(77))
    lib.set_option("fjlt_fut", "dct")
    full = F.apply(Ac, None, 0)
    if rank == 0:
        want = torch.from_numpy(o.FJLT(n, s, o.Context(77)).apply(A.cpu().numpy(), "columnwise", fut="wht")).to(dev)
        checks["FJLT-WHT local vs oracle"] = rel(full, want)
    got = sd.apply_distributed_overlapped(F, Ac[off:off + cnt], None, 0, off, reduce="allreduce", blocks=4)
    checks["FJLT-WHT overlapped allreduce"] = rel(got, full)
    got = sd.apply_distributed_overlapped(F, Ac[off:off + cnt], None, 0, off, reduce="reduce_scatter", blocks=4)
    edges = sd._block_edges(d, 4, world)
    mine = torch.cat([full[:, b0 + rank * (b1 - b0) // world: b0 + (rank + 1) * (b1 - b0) // world]
                      for b0, b1 in zip(edges, edges[1:])], dim=1)
    checks["FJLT-WHT overlapped reduce_scatter"] = rel(got, mine)

    try:        # reduce-scatter fused into the kernel (stores into peer memory)
        got = sd.fjlt_wht_fused_reduce_scatter(F, Ac[off:off + cnt], off)
        cpo = d // world
        for _ in range(2):      # twice: the staging slabs are reused
            got = sd.fjlt_wht_fused_reduce_scatter(F, Ac[off:off + cnt], off)
        checks["FJLT-WHT fused peer-store reduce-scatter"] = rel(got, full[:, rank * cpo:(rank + 1) * cpo])
    except Exception as e:
        print("rank %d fused peer path failed: %r" % (rank, e), flush=True)
        checks["FJLT-WHT fused peer-store reduce-scatter"] = 1.0

    # --- FJLT, DCT mode (what the reference computes): rows -> columns all-to-all, gather
    Fd = sketch.FJLT(n, s, lib.Context(78))
    full = Fd.apply(Ac, None, 0)
    if rank == 0:
        want = torch.from_numpy(o.FJLT(n, s, o.Context(78)).apply(A.cpu().numpy(), "columnwise")).to(dev)
        checks["FJLT-DCT local vs oracle"] = rel(full, want)
    got = sd.fjlt_apply_row_sharded(Fd, Ac[off:off + cnt], gather=True)
    checks["FJLT-DCT row-sharded (all-to-all + all-gather)"] = rel(got, full)
    got = sd.fjlt_apply_row_sharded(Fd, A[off:off + cnt], gather=True)        # row-major shard
    checks["FJLT-DCT row-sharded, row-major shard"] = rel(got, full)

    # --- sparse exchange (CSC -> CSC partials, column-block all-to-all, merge)
    import scipy.sparse
    ns, ds, ss = 4000 * world, 40 * world, 64
    Asp = scipy.sparse.random(ns, ds, 0.05, format="csc", random_state=9, dtype=np.float64)
    C = sketch.CWT(ns, ss, lib.Context(41))
    want = o.CWT(ns, ss, o.Context(41)).apply_csc_sparse(Asp, "columnwise")
    off2, cnt2 = sd.shard_range(ns, rank, world)
    sh = Asp[off2:off2 + cnt2].tocsc()
    Ad = lib.DeviceCSC(torch.from_numpy(sh.indptr.astype(np.int32)).to(dev), torch.from_numpy(sh.indices.astype(np.int32)).to(dev),
                       torch.from_numpy(sh.data).to(dev), (cnt2, ds))
    res, (c0, nc) = sd.apply_distributed_sparse(C, Ad, off2)
    got = scipy.sparse.csc_matrix((res.values.cpu().numpy(), res.indices.cpu().numpy(), res.indptr.cpu().numpy()), shape=(ss, nc))
    w = want[:, c0:c0 + nc].tocsc(); w.sort_indices(); got.sort_indices()
    # same stored pattern; values to rounding (every rank adds its own rows first, so a slot's sum
    # is associated differently from the single left-to-right chain)
    same = got.indptr.tolist() == w.indptr.tolist() and got.indices.tolist() == w.indices.tolist()
    checks["CWT sparse exchange: pattern identical"] = 0.0 if same else 1.0
    checks["CWT sparse exchange: values"] = float(np.linalg.norm(got.data - w.data) / np.linalg.norm(w.data)) if same else 1.0
    # --- the same through the pipelined rounds (enough columns for dist.exchange_merge_pipelined):
    # entry for entry what the one-shot exchange gives
    ds2 = 320 * world
    Asp2 = scipy.sparse.random(ns, ds2, 0.02, format="csc", random_state=10, dtype=np.float64)
    sh2 = Asp2[off2:off2 + cnt2].tocsc()
    Ad2 = lib.DeviceCSC(torch.from_numpy(sh2.indptr.astype(np.int32)).to(dev), torch.from_numpy(sh2.indices.astype(np.int32)).to(dev),
                        torch.from_numpy(sh2.data).to(dev), (cnt2, ds2))
    os.environ["SKB_SPARSE_PIPE"] = "1"
    one, (c0, nc) = sd.apply_distributed_sparse(C, Ad2, off2)
    os.environ["SKB_SPARSE_PIPE"] = "4"
    pip, (c0p, ncp) = sd.apply_distributed_sparse(C, Ad2, off2)
    os.environ.pop("SKB_SPARSE_PIPE")
    ident = (c0, nc) == (c0p, ncp) and torch.equal(one.indptr, pip.indptr) and torch.equal(one.indices, pip.indices) \
        and torch.equal(one.values, pip.values)
    checks["CWT sparse exchange in rounds == one shot (bitwise)"] = 0.0 if ident else 1.0
    want2 = o.CWT(ns, ss, o.Context(41)).apply_csc_sparse(Asp2, "columnwise")[:, c0:c0 + nc].tocsc()
    got2 = scipy.sparse.csc_matrix((pip.values.cpu().numpy(), pip.indices.cpu().numpy(), pip.indptr.cpu().numpy()), shape=(ss, nc))
    checks["CWT sparse exchange in rounds vs oracle"] = float(abs(got2 - want2).max()) / float(abs(want2).max())
    bad = {k: v for k, v in checks.items() if not v < 1e-12}
    for k, v in checks.items():
        print("rank %d  %-50s %.3e" % (rank, k, v), flush=True)
    flag = torch.tensor([len(bad)], device=dev)
    dist.all_reduce(flag)
    dist.barrier()
    dist.destroy_process_group()
    if int(flag.item()):
        print("rank %d FAILED: %s" % (rank, bad), flush=True)
        sys.exit(1)
    if rank == 0:
        print("NCCL_WORKER_OK world=%d" % world, flush=True)


if __name__ == "__main__":
    main()
