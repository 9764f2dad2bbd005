"""Multi-GPU plumbing: one process per GPU, independent ciphertext shards, keys broadcast ONCE.

The reference has no distributed runtime; gates are independent, so the only exchange step is the
one-time broadcast of the packed keys (NCCL over NVLink on GPUs; gloo on CPU for the host-logic tests).  There is no
collective on the per-gate path."""
import numpy as np
import torch
import torch.distributed as dist


def shard_range(count, rank, world):
    """Contiguous [lo, hi) slice of a batch of `count` ciphertexts owned by `rank` (sizes differ by at most 1)."""
    base, rem = divmod(count, world)
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def shard_circuit_inputs(in_a, in_b, rank, world):
    """A circuit batch (luxb200_circuit_batch) shards over its independent input sets: wire-major buffers in_a
    [wires][count][n], in_b [wires][count] -> this rank's contiguous slice of the count axis (every rank keeps all wires
    and the whole gate list; no exchange between ranks at any level)."""
    lo, hi = shard_range(in_b.shape[1], rank, world)
    return np.ascontiguousarray(in_a[:, lo:hi]), np.ascontiguousarray(in_b[:, lo:hi]), (lo, hi)


def broadcast_keys(arrays, lengths, src=0, device="cpu"):
    """Broadcast the canonical flat key arrays from rank `src`.

    arrays : dict name -> numpy array (uint64 or int64) on rank src, None elsewhere
    lengths: dict name -> element count (known on every rank from the parameters)
    returns dict name -> torch.int64 tensor on `device` (bit pattern of the uint64 data)"""
    out = {}
    rank = dist.get_rank() if dist.is_initialized() else 0
    for name in sorted(lengths):
        t = torch.empty(lengths[name], dtype=torch.int64, device=device)
        if rank == src:
            a = np.ascontiguousarray(arrays[name])
            t.copy_(torch.from_numpy(a.view(np.int64)))
        if dist.is_initialized() and dist.get_world_size() > 1:
            dist.broadcast(t, src)
        out[name] = t
    return out


def gather_counts(local_count, device="cpu"):
    """Total units processed over all ranks (for whole-job throughput)."""
    t = torch.tensor([local_count], dtype=torch.int64, device=device)
    if dist.is_initialized() and dist.get_world_size() > 1:
        dist.all_reduce(t)
    return int(t.item())


def max_over_ranks(x, device="cpu"):
    t = torch.tensor([x], dtype=torch.float64, device=device)
    if dist.is_initialized() and dist.get_world_size() > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())
