"""Sharding of a batch of independent OCPs over the ranks of one box.

The B problems of a batch share nothing (models are read-only), so the multi-GPU path is: every rank owns a
contiguous block of problems, solves it with its own solver handle on its own GPU, and ONE collective at the end
gathers the per-problem record {cost, iter, stop, converged} to every rank (4 doubles per problem: 256 KB for
B = 8192). No collective sits inside the solve. The functions below are backend-agnostic (`nccl` on the GPUs,
`gloo` in the CPU tests) and never touch the data path.
"""
import numpy as np

RECORD_FIELDS = ("cost", "iter", "stop", "converged")
_SCALAR_COLUMNS = (0, 9, 1, 11)  # columns of cgx_solver_get_scalars (include/crocoddyl_b200.h) in RECORD_FIELDS order


def shard_range(n_total, rank, world):
    """Contiguous block [lo, hi) of rank `rank`: sizes differ by at most one, the first n_total % world ranks
    get the extra problem."""
    if world < 1 or not (0 <= rank < world):
        raise ValueError("invalid rank/world")
    base, rem = divmod(int(n_total), world)
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def shard_sizes(n_total, world):
    return [shard_range(n_total, r, world)[1] - shard_range(n_total, r, world)[0] for r in range(world)]


def records_from_scalars(scalars):
    """[B, CGX_NSCALARS] scalar block of a solver -> [B, 4] records {cost, iter, stop, converged}."""
    return np.ascontiguousarray(np.asarray(scalars)[:, list(_SCALAR_COLUMNS)], dtype=np.float64)


def gather_records(local_records, n_total=None, dist=None, device=None):
    """All-gather of the per-problem records. local_records: [B_local, 4] (numpy). Returns [n_total, 4] on every
    rank, in global problem order. Uneven shards are padded to the largest one for the collective.
    dist: an initialised torch.distributed module (None = single process: returns the input)."""
    local_records = np.ascontiguousarray(local_records, dtype=np.float64)
    if dist is None or not dist.is_initialized() or dist.get_world_size() == 1:
        return local_records
    import torch

    world, rank = dist.get_world_size(), dist.get_rank()
    if n_total is None:
        cnt = torch.tensor([local_records.shape[0]], dtype=torch.int64, device=device)
        dist.all_reduce(cnt)
        n_total = int(cnt.item())
    sizes = shard_sizes(n_total, world)
    if sizes[rank] != local_records.shape[0]:
        raise ValueError("rank %d holds %d problems, its contiguous shard has %d" % (rank, local_records.shape[0], sizes[rank]))
    width = local_records.shape[1]
    pad = max(sizes)
    send = torch.zeros((pad, width), dtype=torch.float64, device=device)
    send[:sizes[rank]] = torch.from_numpy(local_records).to(send.device)
    recv = torch.empty((world * pad, width), dtype=torch.float64, device=device)  # concatenated form: nccl and gloo
    dist.all_gather_into_tensor(recv, send)
    recv = recv.cpu().numpy().reshape(world, pad, width)
    return np.concatenate([recv[r, :sizes[r]] for r in range(world)], 0)
