"""N>1 host path on CPU: world_size-2 gloo.  Each rank computes the partial sketch of its
row shard with the ORACLE (the product has no CPU path) using global input indices, then
the product's collective code sums them; the result must equal the unsharded sketch --
the reference's "distributed == local for the same seed" test
(tests/unit/DenseSketchApplyElementalTest.cpp:28-106)."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, mode, out_dir):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from libskylark_b200.dist import reduce_partial, shard_range
    from oracle import oracle as o
    n, d, s = 101, 6, 8
    A = np.random.default_rng(0).standard_normal((n, d))
    off, cnt = shard_range(n, rank, world)
    ctx = o.Context(77)
    if mode == "cwt":
        t = o.CWT(n, s, ctx)
        P = t.explicit()
        full = P @ A
        part = P[:, off:off + cnt] @ A[off:off + cnt]
    else:
        t = o.JLT(n, s, ctx)
        R = t.realize()
        full = R @ A
        part = R[:, off:off + cnt] @ A[off:off + cnt]
    part_c = np.ascontiguousarray(part)
    res = reduce_partial(part_c.copy(), None, "allreduce")
    np.testing.assert_allclose(res, full, rtol=1e-13, atol=1e-13)
    res = reduce_partial(np.asfortranarray(part).copy(order="F"), None, "allreduce")
    np.testing.assert_allclose(res, full, rtol=1e-13, atol=1e-13)
    res = reduce_partial(part_c.copy(), None, "reduce", dst=1)
    if rank == 1:
        np.testing.assert_allclose(res, full, rtol=1e-13, atol=1e-13)
    res = reduce_partial(torch.from_numpy(part_c.copy()), None, "reduce_scatter")
    rows = s // world
    np.testing.assert_allclose(res.numpy(), full[rank * rows:(rank + 1) * rows], rtol=1e-13, atol=1e-13)
    assert reduce_partial(part_c, None, None) is part_c
    dist.barrier()
    dist.destroy_process_group()
    open(os.path.join(out_dir, "ok%d" % rank), "w").write("ok")


@pytest.mark.parametrize("mode", ["cwt", "jlt"])
def test_row_sharded_sum_world2(tmp_path, mode):
    port = _free_port()
    mp.spawn(_worker, args=(2, port, mode, str(tmp_path)), nprocs=2, join=True)
    assert os.path.exists(tmp_path / "ok0") and os.path.exists(tmp_path / "ok1")
