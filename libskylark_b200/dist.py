"""Row-sharded sketches over a torch.distributed process group (one process per GPU).

What the reference does with MPI: a columnwise sketch of a matrix
whose sketched dimension is distributed is a local partial sketch that uses GLOBAL input
indices (sketch/hash_transform_Elemental.hpp:543-555,
sketch/dense_transform_data.hpp:113-152), followed by a sum over ranks
(boost::mpi::all_reduce at hash_transform_Elemental.hpp:557-561, boost::mpi::reduce at
:360-366, El::AxpyContract = reduce-scatter at
dense_transform_Elemental_coldist_star.hpp:228).  Here the partition is contiguous blocks
(any partition of the index range is valid because S depends on the global index only) and
the sum is an NCCL collective over NVLink.  Rowwise sketches of column-sharded data (feature
maps over sharded data points) need no collective: call `apply` on each rank.
"""
import os

import numpy


def shard_range(n, rank, world):
    """Contiguous block partition of range(n): returns (offset, count) of `rank`."""
    base, rem = divmod(int(n), int(world))
    count = base + (1 if rank < rem else 0)
    offset = rank * base + min(rank, rem)
    return offset, count


def _as_tensor(SA):
    import torch
    if isinstance(SA, torch.Tensor):
        return SA, None
    t = torch.from_numpy(SA if SA.flags.c_contiguous else SA.T)   # shares memory
    return t, SA


def reduce_partial(SA, group=None, reduce="allreduce", dst=0):
    """Sum partial sketches over the group.  SA: torch tensor (CUDA with NCCL, CPU with gloo)
    or numpy array (CPU, gloo).  Returns the reduced sketch: SA itself for allreduce /
    reduce (valid on `dst` only), this rank's slice for reduce_scatter."""
    import torch
    import torch.distributed as dist
    if reduce is None or not dist.is_available() or not dist.is_initialized():
        return SA
    world = dist.get_world_size(group)
    if world == 1:
        return SA
    t, np_owner = _as_tensor(SA)
    if reduce == "allreduce":
        base = t if t.is_contiguous() else t.t()
        dist.all_reduce(base, op=dist.ReduceOp.SUM, group=group)
        return SA
    if reduce == "reduce":
        base = t if t.is_contiguous() else t.t()
        dist.reduce(base, dst=dst, op=dist.ReduceOp.SUM, group=group)
        return SA
    if reduce == "reduce_scatter":
        transposed = not t.is_contiguous()
        base = t.t() if transposed else t
        if not base.is_contiguous():
            raise ValueError("reduce_scatter needs a row- or column-major sketch")
        if base.shape[0] % world != 0:
            raise ValueError("reduce_scatter needs the leading storage dimension (%d) to be a "
                             "multiple of the group size (%d)" % (base.shape[0], world))
        rows = base.shape[0] // world
        rank = dist.get_rank(group)
        out = torch.empty((rows,) + tuple(base.shape[1:]), dtype=base.dtype, device=base.device)
        if base.is_cuda:
            dist.reduce_scatter_tensor(out, base, op=dist.ReduceOp.SUM, group=group)
        else:   # gloo has no reduce_scatter: all-reduce, then keep the local slice
            dist.all_reduce(base, op=dist.ReduceOp.SUM, group=group)
            out.copy_(base[rank * rows:(rank + 1) * rows])
        out = out.t() if transposed else out
        if np_owner is not None:
            res = out.numpy()
            return res if np_owner.flags.c_contiguous else res
        return out
    raise ValueError("reduce must be allreduce, reduce, reduce_scatter or None")


def apply_distributed(S, A_local, SA=None, dim=0, index_offset=0, group=None,
                      reduce="allreduce", dst=0):
    """Partial sketch of this rank's shard, then the collective (see reduce_partial)."""
    SA = S.apply_shard(A_local, SA, dim, index_offset)
    return reduce_partial(SA, group, reduce, dst)


def exchange_sparse_partials(indptr, indices, values, group=None, reorder=True):
    """Column-block all-to-all of sparse partial sketches (the sparse form of the reduce-scatter
    above).  Every rank passes its partial s x d sketch as CSC tensors (int32 indptr/indices,
    float64 values; CUDA with NCCL, CPU with gloo).  Rank r receives, for its block of columns
    `shard_range(d, r, world)`, the entries of all ranks, stored per column in rank order --
    the order of the global input index when rank r holds the r-th block of rows, so adding
    equal (row, column) entries left to right reproduces the single-process sum.  Returns
    (indptr, indices, values, (first_column, columns)); equal (row, column) entries are NOT yet
    added (see sketch.coalesce / sl_sp_matrix_coalesce)."""
    import torch
    import torch.distributed as dist
    d = indptr.numel() - 1
    if not dist.is_available() or not dist.is_initialized() or dist.get_world_size(group) == 1:
        return indptr, indices, values, (0, d)
    world, rank = dist.get_world_size(group), dist.get_rank(group)
    dev = indptr.device
    blocks = [shard_range(d, r, world) for r in range(world)]
    col0, ncols = blocks[rank]
    ip = indptr.to(torch.int64)
    edges = torch.tensor([b[0] for b in blocks] + [d], device=dev)
    send_entries = (ip[edges[1:]] - ip[edges[:-1]]).tolist()            # entries per destination
    counts = (ip[1:] - ip[:-1]).to(torch.int32).contiguous()            # entries per column
    recv_counts = torch.empty(world * ncols, dtype=torch.int32, device=dev)
    dist.all_to_all_single(recv_counts, counts, output_split_sizes=[ncols] * world,
                           input_split_sizes=[b[1] for b in blocks], group=group)
    C = recv_counts.view(world, ncols).to(torch.int64)                  # C[s, j]: entries of column j from rank s
    recv_entries = C.sum(dim=1).tolist()
    total = int(sum(recv_entries))
    ridx = torch.empty(total, dtype=indices.dtype, device=dev)
    rval = torch.empty(total, dtype=values.dtype, device=dev)
    dist.all_to_all_single(ridx, indices.contiguous(), output_split_sizes=recv_entries,
                           input_split_sizes=send_entries, group=group)
    dist.all_to_all_single(rval, values.contiguous(), output_split_sizes=recv_entries,
                           input_split_sizes=send_entries, group=group)
    if not reorder:
        # leave the pieces as they arrived (rank-major) and describe them: entries of column j from
        # rank r are [vptr[r * ncols + j], vptr[r * ncols + j + 1]) -- what sl_sp_merge_segments reads
        vptr = torch.zeros(world * ncols + 1, dtype=torch.int32, device=dev)
        vptr[1:] = torch.cumsum(recv_counts, 0, dtype=torch.int32) if total < 2 ** 31 else 0
        return vptr, ridx, rval, (col0, ncols)
    # received order is (rank, column); wanted order is (column, rank)
    per_col = C.sum(dim=0)
    out_indptr = torch.zeros(ncols + 1, dtype=torch.int64, device=dev)
    out_indptr[1:] = torch.cumsum(per_col, 0)
    flat = C.reshape(-1)
    src_start = torch.cumsum(flat, 0) - flat
    dst_start = (out_indptr[:-1].unsqueeze(0) + torch.cumsum(C, 0) - C).reshape(-1)
    seg = torch.repeat_interleave(torch.arange(flat.numel(), device=dev), flat)
    dest = dst_start[seg] + (torch.arange(total, device=dev) - src_start[seg])
    oidx = torch.empty_like(ridx)
    oval = torch.empty_like(rval)
    oidx[dest] = ridx
    oval[dest] = rval
    return out_indptr.to(torch.int32), oidx, oval, (col0, ncols)


def exchange_merge_pipelined(indptr, indices, values, nrows, group=None, nblocks=4):
    """exchange_sparse_partials(reorder=False) + sketch.merge_segments, cut in `nblocks` rounds: round
    q moves, for every destination, the q-th slice of that destination's columns.  All rounds are
    queued on NCCL at once (one count exchange and one host read of the split sizes for all of
    them; grouped send/recv straight from slices of the partial, no packing copy), and the merge
    kernel of round q runs while round q + 1 is on the wire.  Returns (lib.DeviceCSC of shape
    nrows x columns_of_this_rank, (first_column, columns)).  CUDA + NCCL only."""
    import torch
    import torch.distributed as dist
    from . import lib, sketch
    world, rank = dist.get_world_size(group), dist.get_rank(group)
    dev = indptr.device
    dbg = os.environ.get("SKB_DIST_DEBUG") and rank == 0
    if dbg:
        import time
        marks = []

        def mark(tag):
            ev = torch.cuda.Event(enable_timing=True); ev.record()
            marks.append((tag, time.perf_counter(), ev))
        mark("start")
    d = indptr.numel() - 1
    blocks = [shard_range(d, r, world) for r in range(world)]
    col0, ncols = blocks[rank]
    # column edges of (destination r, round q): edge[r][q] .. edge[r][q + 1]
    edge = [[b0 + (bn * q) // nblocks for q in range(nblocks + 1)] for (b0, bn) in blocks]
    my = [e - col0 for e in edge[rank]]                                # this rank's slices, local columns
    ip = indptr.to(torch.int64)
    counts = (ip[1:] - ip[:-1]).to(torch.int32).contiguous()            # entries per column
    recv_counts = torch.empty(world * ncols, dtype=torch.int32, device=dev)
    dist.all_to_all_single(recv_counts, counts, output_split_sizes=[ncols] * world,
                           input_split_sizes=[b[1] for b in blocks], group=group)
    C = recv_counts.view(world, ncols)                                  # C[s, j]: entries of column j from rank s
    flat_edges = torch.tensor([e for r in range(world) for e in edge[r]], device=dev)
    Ccum = torch.zeros(world, ncols + 1, dtype=torch.int64, device=dev)
    Ccum[:, 1:] = torch.cumsum(C, 1)
    my_t = torch.tensor(my, device=dev)
    # ONE host read: send offsets at every (r, q) edge, receive sizes of every (source, round)
    host = torch.cat([ip[flat_edges], (Ccum[:, my_t[1:]] - Ccum[:, my_t[:-1]]).reshape(-1)]).tolist()
    send_off = [host[r * (nblocks + 1):(r + 1) * (nblocks + 1)] for r in range(world)]
    recv_n = host[world * (nblocks + 1):]                               # [source][round]
    if dbg:
        mark("sizes on host")
    works, bufs = [], []
    for q in range(nblocks):
        n_from = [int(recv_n[sr * nblocks + q]) for sr in range(world)]
        tot = sum(n_from)
        ridx = torch.empty(tot, dtype=indices.dtype, device=dev)
        rval = torch.empty(tot, dtype=values.dtype, device=dev)
        o_idx = list(torch.split(ridx, n_from)); o_val = list(torch.split(rval, n_from))
        i_idx = [indices[int(send_off[r][q]):int(send_off[r][q + 1])] for r in range(world)]
        i_val = [values[int(send_off[r][q]):int(send_off[r][q + 1])] for r in range(world)]
        w1 = dist.all_to_all(o_idx, i_idx, group=group, async_op=True)
        w2 = dist.all_to_all(o_val, i_val, group=group, async_op=True)
        works.append((w1, w2)); bufs.append((ridx, rval))
    if dbg:
        mark("rounds queued")
    parts = []
    for q in range(nblocks):
        works[q][0].wait(); works[q][1].wait()
        if dbg:
            mark("round %d arrived" % q)
        ncq = my[q + 1] - my[q]
        vptr = torch.zeros(world * ncq + 1, dtype=torch.int32, device=dev)
        vptr[1:] = torch.cumsum(C[:, my[q]:my[q + 1]].reshape(-1), 0, dtype=torch.int32)
        parts.append(sketch.merge_segments(vptr, bufs[q][0], bufs[q][1], world, (nrows, ncq)))
        if dbg:
            mark("round %d merged" % q)
    sizes = torch.tensor([0] + [int(p_.indices.numel()) for p_ in parts]).cumsum(0).tolist()
    out_ip = torch.cat([parts[0].indptr.to(torch.int64)] +
                       [parts[q].indptr[1:].to(torch.int64) + sizes[q] for q in range(1, nblocks)]).to(torch.int32)
    merged = lib.DeviceCSC(out_ip, torch.cat([p_.indices for p_ in parts]), torch.cat([p_.values for p_ in parts]),
                           (nrows, ncols), "csc")
    if dbg:
        mark("joined")
        torch.cuda.synchronize()
        print("  ".join("%s host %.2f gpu %.2f |" % (t, (h - marks[0][1]) * 1e3, marks[0][2].elapsed_time(e))
                        for (t, h, e) in marks[1:]), flush=True)
    return merged, (col0, ncols)


def apply_distributed_sparse(S, A_local, index_offset=0, group=None):
    """Columnwise CWT/MMT/WZT of a row-sharded device CSC matrix with a SPARSE exchange: every
    rank sketches its rows into a sparse partial (lib.DeviceCSC, format "csc"), the partials
    are exchanged by column block (exchange_sparse_partials) and merged on the device.
    Returns (lib.DeviceCSC of shape s x columns_of_this_rank, (first_column, columns)).  For
    sketches much sparser than s x d this moves nnz*12 bytes instead of the s*d*8 bytes of the
    dense reduce-scatter (config C4: 1.5 GB instead of 16 GB per GPU)."""
    import torch
    from . import lib, sketch
    part = S.apply_shard(A_local, lib.DeviceCSC.empty((S.getsketchdim(), A_local.shape[1]), "csc",
                                                       A_local.values.device), 0, index_offset)
    import torch.distributed as dist
    world = dist.get_world_size(group) if dist.is_available() and dist.is_initialized() else 1
    nblk = int(os.environ.get("SKB_SPARSE_PIPE", "4"))
    if world > 1 and part.values.is_cuda and nblk > 1 and A_local.shape[1] >= world * nblk * 64 \
            and part.values.numel() < 2 ** 31:
        try:
            return exchange_merge_pipelined(part.indptr, part.indices, part.values, S.getsketchdim(), group, nblk)
        except lib.errors.LowerLayerError:
            pass                          # a merged column too long for the shared-memory sort: below
    if world > 1 and part.values.is_cuda:
        # merge straight from the receive buffers (no re-ordering pass); falls back when a merged
        # column is too long for the shared-memory sort
        vptr, ix, vv, (c0, nc) = exchange_sparse_partials(part.indptr, part.indices, part.values, group,
                                                          reorder=False)
        try:
            return sketch.merge_segments(vptr, ix, vv, world, (S.getsketchdim(), nc)), (c0, nc)
        except Exception:
            flat = (vptr[1:] - vptr[:-1]).to(torch.int64).view(world, nc)
            ip, ix, vv = _reorder_segments(flat, ix, vv)
            return sketch.coalesce(lib.DeviceCSC(ip, ix, vv, (S.getsketchdim(), nc), "csc")), (c0, nc)
    ip, ix, vv, (c0, nc) = exchange_sparse_partials(part.indptr, part.indices, part.values, group)
    merged = lib.DeviceCSC(ip, ix, vv, (S.getsketchdim(), nc), "csc")
    return sketch.coalesce(merged), (c0, nc)


def _reorder_segments(C, ridx, rval):
    """(rank, column)-ordered entries -> CSC with every column's entries in rank order."""
    import torch
    dev = ridx.device
    world, ncols = C.shape
    total = int(ridx.numel())
    per_col = C.sum(dim=0)
    out_indptr = torch.zeros(ncols + 1, dtype=torch.int64, device=dev)
    out_indptr[1:] = torch.cumsum(per_col, 0)
    flat = C.reshape(-1)
    src_start = torch.cumsum(flat, 0) - flat
    dst_start = (out_indptr[:-1].unsqueeze(0) + torch.cumsum(C, 0) - C).reshape(-1)
    seg = torch.repeat_interleave(torch.arange(flat.numel(), device=dev), flat)
    dest = dst_start[seg] + (torch.arange(total, device=dev) - src_start[seg])
    oidx = torch.empty_like(ridx)
    oval = torch.empty_like(rval)
    oidx[dest] = ridx
    oval[dest] = rval
    return out_indptr.to(torch.int32), oidx, oval


# ---------------------------------------------------------------------------
# Overlapped exchange, FJLT in DCT mode on row shards
# ---------------------------------------------------------------------------
def _block_edges(extent, blocks, multiple):
    """Cuts range(extent) in at most `blocks` pieces whose sizes are multiples of `multiple`
    (the last one takes the remainder, which must itself be a multiple)."""
    unit = max(1, int(multiple))
    if extent % unit:
        raise ValueError("extent %d is not a multiple of %d" % (extent, unit))
    nunits = extent // unit
    blocks = max(1, min(int(blocks), nunits))
    per, rem = divmod(nunits, blocks)
    edges = [0]
    for b in range(blocks):
        edges.append(edges[-1] + (per + (1 if b < rem else 0)) * unit)
    return edges


def apply_distributed_overlapped(S, A_local, SA=None, dim=0, index_offset=0, group=None,
                                 reduce="allreduce", blocks=4, sm_reserve=8):
    """apply_distributed with the exchange hidden behind the kernels: the NON-sketched dimension
    is cut in `blocks` blocks; the sketch of block b+1 runs on the current stream while NCCL
    reduces block b on its own stream (what SURVEY 2.3 / hash_transform_Elemental.hpp:557-561
    do in one step at the end).  `sm_reserve` SMs are left free by the persistent FJLT kernels
    so that the collective's CTAs can be resident next to them.

    Needs device tensors whose blocks are contiguous: column-major operands for dim 0,
    row-major for dim 1 (anything else falls back to apply_distributed).  With
    reduce="reduce_scatter" every block is scattered on its own, so a rank ends up owning the
    r-th 1/world of EVERY block (block-cyclic), returned concatenated in block order."""
    import torch
    import torch.distributed as dist
    dim = S._normdim(dim)
    world = dist.get_world_size(group) if dist.is_available() and dist.is_initialized() else 1
    ok = isinstance(A_local, torch.Tensor) and A_local.is_cuda and A_local.dim() == 2 and reduce in (
        "allreduce", "reduce_scatter")
    if ok:
        col_major = A_local.stride(0) == 1 and A_local.shape[0] > 1
        ok = (dim == 0 and col_major) or (dim == 1 and not col_major and A_local.stride(1) == 1)
    if not ok or world == 1 or blocks <= 1:
        return apply_distributed(S, A_local, SA, dim, index_offset, group, reduce)
    from . import lib
    s, other = S.getsketchdim(), A_local.shape[1 - dim]
    if SA is None:
        SA = (torch.empty(other, s, dtype=A_local.dtype, device=A_local.device).t() if dim == 0
              else torch.empty(other, s, dtype=A_local.dtype, device=A_local.device))
    # the persistent FJLT kernels take one column per CTA and round: blocks that are whole rounds of
    # the CTAs left after the reservation lose nothing to a ragged last round
    unit = world if reduce == "reduce_scatter" else 1
    ctas = 148 - int(sm_reserve)
    if other % unit == 0 and other >= 2 * ctas:
        rounds = -(-other // ctas)                       # ceil
        per = max(1, rounds // blocks)
        step = per * ctas // unit * unit
        if step > 0 and other - (blocks - 1) * step >= step:
            first = other - (blocks - 1) * step          # the ragged (largest) block goes first:
            edges = [0] + [first + b * step for b in range(blocks)]   # the last reduction is a short one
        else:
            edges = _block_edges(other, blocks, unit)
    else:
        edges = _block_edges(other, blocks, unit)
    works, outs = [], []
    lib.set_option("sm_reserve", int(sm_reserve))
    try:
        for b0, b1 in zip(edges, edges[1:]):
            Ab = A_local[:, b0:b1] if dim == 0 else A_local[b0:b1, :]
            SAb = SA[:, b0:b1] if dim == 0 else SA[b0:b1, :]
            S.apply_shard(Ab, SAb, dim, index_offset)
            slab = SAb.t() if dim == 0 else SAb            # contiguous (b1 - b0) x s
            if reduce == "allreduce":
                works.append(dist.all_reduce(slab, op=dist.ReduceOp.SUM, group=group, async_op=True))
            else:
                out = torch.empty((slab.shape[0] // world, s), dtype=slab.dtype, device=slab.device)
                works.append(dist.reduce_scatter_tensor(out, slab, op=dist.ReduceOp.SUM, group=group,
                                                        async_op=True))
                outs.append(out)
    finally:
        lib.set_option("sm_reserve", 0)
    for w in works:
        w.wait()
    if reduce == "allreduce":
        return SA
    res = torch.cat(outs, dim=0)
    return res.t() if dim == 0 else res


def reshard_rows_to_columns(A_local, n_total, group=None):
    """[VC,*] -> [*,VC] of sketch/FJLT_Elemental.hpp:149-150 as one all-to-all: every rank passes
    its contiguous block of ROWS (shard_range(n_total, rank, world)) of an n_total x d matrix and
    receives all rows of its block of COLUMNS (shard_range(d, rank, world)).  Returns the
    column-major n_total x d_local tensor (a transposed view of contiguous storage) and
    (first_column, columns)."""
    import torch
    import torch.distributed as dist
    d = A_local.shape[1]
    if not dist.is_available() or not dist.is_initialized() or dist.get_world_size(group) == 1:
        return A_local, (0, d)
    world, rank = dist.get_world_size(group), dist.get_rank(group)
    r0, nloc = shard_range(n_total, rank, world)
    if A_local.shape[0] != nloc:
        raise ValueError("rank %d must hold rows [%d, %d) of the matrix" % (rank, r0, r0 + nloc))
    send = A_local.t().contiguous()                      # (d, nloc): column blocks are row slabs
    c0, dloc = shard_range(d, rank, world)
    rows = [shard_range(n_total, q, world)[1] for q in range(world)]
    cols = [shard_range(d, q, world)[1] for q in range(world)]
    recv = torch.empty(dloc * n_total, dtype=A_local.dtype, device=A_local.device)
    dist.all_to_all_single(recv, send.view(-1), output_split_sizes=[dloc * rq for rq in rows],
                           input_split_sizes=[cq * nloc for cq in cols], group=group)
    out = torch.empty(dloc, n_total, dtype=A_local.dtype, device=A_local.device)
    pos = 0
    for q in range(world):                               # source q sent (dloc, rows[q])
        off = shard_range(n_total, q, world)[0]
        out[:, off:off + rows[q]] = recv[pos:pos + dloc * rows[q]].view(dloc, rows[q])
        pos += dloc * rows[q]
    return out.t(), (c0, dloc)


def gather_columns(SA_local, d_total, group=None):
    """[*,VC] -> [*,*] (sketch/FJLT_Elemental.hpp:168): every rank holds the s x d_local sketch of
    its column block (column-major); returns the replicated s x d_total sketch (column-major)."""
    import torch
    import torch.distributed as dist
    if not dist.is_available() or not dist.is_initialized() or dist.get_world_size(group) == 1:
        return SA_local
    world = dist.get_world_size(group)
    s = SA_local.shape[0]
    slab = SA_local.t().contiguous()                     # (d_local, s)
    cols = [shard_range(d_total, q, world)[1] for q in range(world)]
    out = torch.empty(d_total, s, dtype=slab.dtype, device=slab.device)
    if len(set(cols)) == 1:
        dist.all_gather_into_tensor(out, slab, group=group)
    else:           # uneven blocks: pad to the widest, gather, drop the padding
        cmax = max(cols)
        padded = torch.zeros(cmax, s, dtype=slab.dtype, device=slab.device)
        padded[:slab.shape[0]] = slab
        allp = torch.empty(world * cmax, s, dtype=slab.dtype, device=slab.device)
        dist.all_gather_into_tensor(allp, padded, group=group)
        pos = 0
        for q in range(world):
            out[pos:pos + cols[q]] = allp[q * cmax:q * cmax + cols[q]]
            pos += cols[q]
    return out.t()


def fjlt_apply_row_sharded(S, A_local, group=None, gather=True):
    """Columnwise FJLT of a row-sharded matrix the way the reference does it
    (sketch/FJLT_Elemental.hpp:144-169): re-shard rows -> columns, transform and sample whole
    columns locally, gather.  Works in both FUT modes (the DCT has no other way to use row
    shards); in the Walsh-Hadamard mode apply_distributed_overlapped moves s*d instead of n*d.
    Returns the replicated s x d sketch, or (s x d_local sketch, (first_column, columns)) with
    gather=False."""
    A_cols, (c0, dloc) = reshard_rows_to_columns(A_local, S.getindim(), group)
    SA_loc = S.apply(A_cols, None, 0)
    if not gather:
        return SA_loc, (c0, dloc)
    return gather_columns(SA_loc, A_local.shape[1], group)


# ---------------------------------------------------------------------------
# FJLT (Walsh-Hadamard mode) on row shards with the reduce-scatter fused into the kernel
# ---------------------------------------------------------------------------
_fused_state = {}


def fjlt_wht_fused_reduce_scatter(S, A_local, index_offset, group=None):
    """Columnwise FJLT (fjlt_fut = wht) of a row-sharded matrix in ONE kernel per rank: every rank
    sketches its rows (A_local: n_loc x d CUDA tensor, column-major) and the kernel itself sends each
    finished column of its partial sketch to the GPU that owns that column (stores into peer memory
    over NVLink, torch symmetric memory for the mapping) -- no separate collective.  After a group
    barrier each rank adds the `world` slabs it received.  Returns this rank's s x (d / world) slice
    of the sketch (column-major), columns [rank * d / world, ...).  Needs d % world == 0 and an even
    s <= 8192; the reference reaches the same result through two redistributions
    (sketch/FJLT_Elemental.hpp:144-169)."""
    import ctypes
    import torch
    import torch.distributed as dist
    import torch.distributed._symmetric_memory as symm
    from . import lib
    world, rank = dist.get_world_size(group), dist.get_rank(group)
    s, d = S.getsketchdim(), A_local.shape[1]
    if d % world:
        raise ValueError("the number of columns (%d) must be a multiple of the group size (%d)" % (d, world))
    if not (A_local.is_cuda and A_local.dtype == torch.float64 and A_local.stride(0) == 1):
        raise ValueError("A_local must be a column-major float64 CUDA tensor")
    cpo = d // world
    pg = group if group is not None else dist.group.WORLD
    key = (id(pg), s, d, A_local.device.index)
    st = _fused_state.get(key)
    if st is None:
        # two staging sets used alternately: a rank may start writing set k % 2 of step k as soon as it
        # has passed the barrier of step k - 1, by which time every owner has finished adding its slabs
        # of step k - 2 (stream order) -- one barrier per step instead of two
        sets = []
        for _ in range(2):
            stage = symm.empty((world, cpo, s), dtype=torch.float64, device=A_local.device)
            sets.append((stage, symm.rendezvous(stage, pg.group_name)))
        scratch = torch.empty(148 * s, dtype=torch.float64, device=A_local.device)
        sets[0][1].barrier(channel=0)
        st = _fused_state[key] = {"sets": sets, "scratch": scratch, "step": 0}
    stage, hdl = st["sets"][st["step"] & 1]
    st["step"] += 1
    A = lib.adapt(A_local)
    lib.set_stream_from_torch()
    Aobj = A.ptr()
    try:
        lib.callsl("sl_fjlt_apply_shard_peer", S._obj, Aobj, ctypes.c_int64(int(index_offset)),
                   ctypes.c_void_p(int(hdl.buffer_ptrs_dev)), ctypes.c_int(world), ctypes.c_int(rank),
                   ctypes.c_void_p(st["scratch"].data_ptr()))
    finally:
        A.ptrcleaner()
    hdl.barrier(channel=0)                       # every rank's columns have arrived
    return stage.sum(dim=0).t()                  # (d/world, s) storage = column-major s x d/world
