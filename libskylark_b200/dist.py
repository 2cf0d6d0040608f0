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
