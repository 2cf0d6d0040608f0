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


def _sparse_worker(rank, world, port, out_dir):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    import scipy.sparse
    from libskylark_b200.dist import exchange_sparse_partials, shard_range
    from oracle import oracle as o
    n, d, s = 300, 11, 7                      # s small: many equal (row, column) entries to merge
    A = scipy.sparse.random(n, d, 0.3, format="csc", random_state=5, dtype=np.float64)
    t = o.CWT(n, s, o.Context(41))
    full = t.apply_csc_sparse(A, "columnwise")
    off, cnt = shard_range(n, rank, world)
    # this rank's rows only (global row ids kept) -> its sparse partial sketch
    mask = (A.indices >= off) & (A.indices < off + cnt)
    colid = np.repeat(np.arange(d), np.diff(A.indptr))
    Ash = scipy.sparse.csc_matrix((A.data[mask], (A.indices[mask], colid[mask])), shape=(n, d))
    part = t.apply_csc_sparse(Ash, "columnwise")
    part.sort_indices()
    ip, ix, vv, (c0, nc) = exchange_sparse_partials(
        torch.from_numpy(part.indptr.astype(np.int32)), torch.from_numpy(part.indices.astype(np.int32)),
        torch.from_numpy(part.data.copy()), None)
    assert (c0, nc) == shard_range(d, rank, world)
    assert int(ip[-1]) == ix.numel() == vv.numel()
    # entries of one column arrive in rank order: rank 0's rows first
    merged = scipy.sparse.csc_matrix((vv.numpy().copy(), ix.numpy().copy(), ip.numpy().copy()), shape=(s, nc))
    merged.sum_duplicates()
    want = full[:, c0:c0 + nc]
    np.testing.assert_allclose(merged.toarray(), want.toarray(), rtol=1e-13, atol=1e-13)
    dist.barrier()
    dist.destroy_process_group()
    open(os.path.join(out_dir, "ok%d" % rank), "w").write("ok")


def test_sparse_partials_exchange_world2(tmp_path):
    # the sparse form of the reduce-scatter: column-block all-to-all of CSC partial sketches
    port = _free_port()
    mp.spawn(_sparse_worker, args=(2, port, str(tmp_path)), nprocs=2, join=True)
    assert os.path.exists(tmp_path / "ok0") and os.path.exists(tmp_path / "ok1")


def _reshard_worker(rank, world, port, out_dir):
    """FJLT of a row-sharded matrix the reference's way (FJLT_Elemental.hpp:144-169): rows ->
    columns all-to-all, whole columns transformed locally (here by the oracle: the product has no
    CPU path), all-gather.  Uneven row and column blocks on purpose."""
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from libskylark_b200.dist import gather_columns, reshard_rows_to_columns, shard_range, _block_edges
    from oracle import oracle as o
    n, d, s = 37, 7, 5
    A = np.random.default_rng(3).standard_normal((n, d))
    f = o.FJLT(n, s, o.Context(19))
    full = f.apply(A, "columnwise")
    off, cnt = shard_range(n, rank, world)
    for local in (torch.from_numpy(A[off:off + cnt].copy()),                      # row-major shard
                  torch.from_numpy(np.ascontiguousarray(A[off:off + cnt].T)).t()):  # column-major shard
        cols, (c0, dc) = reshard_rows_to_columns(local, n, None)
        assert (c0, dc) == shard_range(d, rank, world) and tuple(cols.shape) == (n, dc)
        np.testing.assert_array_equal(cols.numpy(), A[:, c0:c0 + dc])
        part = f.apply(cols.numpy(), "columnwise")
        got = gather_columns(torch.from_numpy(np.ascontiguousarray(part.T)).t(), d, None)
        np.testing.assert_allclose(got.numpy(), full, rtol=1e-13, atol=1e-13)
    assert _block_edges(24, 4, 2) == [0, 6, 12, 18, 24] and _block_edges(10, 4, 1) == [0, 3, 6, 8, 10]
    assert _block_edges(6, 8, 2) == [0, 2, 4, 6]
    with pytest.raises(ValueError):
        _block_edges(7, 2, 2)
    dist.barrier()
    dist.destroy_process_group()
    open(os.path.join(out_dir, "ok%d" % rank), "w").write("ok")


def test_fjlt_rows_to_columns_reshard_world2(tmp_path):
    port = _free_port()
    mp.spawn(_reshard_worker, args=(2, port, str(tmp_path)), nprocs=2, join=True)
    assert os.path.exists(tmp_path / "ok0") and os.path.exists(tmp_path / "ok1")
