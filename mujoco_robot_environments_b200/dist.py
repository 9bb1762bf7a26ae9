"""Multi-GPU plumbing: scenes are independent, so the batch is cut into contiguous shards, one process
per GPU, one ``b2p_handle`` per rank and NO per-step communication.  The only collectives on this path are

  * ``reduce_stats``  -- all_reduce(SUM) of a small episode-statistics vector (episodes, successes, steps, failures), and
  * ``gather_demos``  -- gather of demonstration records ``[B_local, ...]`` to rank 0 for the transporter data generator
                         (transporter_network_data_generation.py:97-143 writes the dataset from one process),

over NCCL (NVLink) on GPUs and gloo in the CPU tests.  Launch with torchrun; rendezvous on 127.0.0.1.
"""
from __future__ import annotations

import os
from typing import List, Optional, Tuple

import torch
import torch.distributed as dist

MAX_MSG_BYTES = 256 << 20       # chunk size for large demo gathers (obs frames once the ray-cast row exists)


def init(backend: Optional[str] = None) -> Tuple[int, int, int]:
    """-> (rank, world_size, local_rank) from the torchrun environment; no-op when WORLD_SIZE is 1 / unset."""
    rank, world, local = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1")), int(os.environ.get("LOCAL_RANK", "0"))
    if world > 1 and not dist.is_initialized():
        if backend is None:
            backend = "nccl" if torch.cuda.is_available() else "gloo"
        if backend == "nccl":
            torch.cuda.set_device(local)
            dist.init_process_group(backend, device_id=torch.device("cuda", local))
        else:
            dist.init_process_group(backend)
    return rank, world, local


def shard_range(n_total: int, rank: int, world: int) -> Tuple[int, int]:
    """Contiguous shard [lo, hi) of ``n_total`` scenes owned by ``rank``; sizes differ by at most one."""
    base, rem = divmod(int(n_total), int(world))
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def _world() -> int:
    return dist.get_world_size() if dist.is_available() and dist.is_initialized() else 1


def reduce_stats(stats: torch.Tensor) -> torch.Tensor:
    """Sum a small statistics vector over the ranks (in place; returns it)."""
    if _world() > 1:
        dist.all_reduce(stats, op=dist.ReduceOp.SUM)
    return stats


def max_over_ranks(x: torch.Tensor) -> torch.Tensor:
    """Device-timed durations are reported as the max over ranks (bench contract)."""
    if _world() > 1:
        dist.all_reduce(x, op=dist.ReduceOp.MAX)
    return x


HOST_ABOVE_BYTES = 4 << 30      # a gathered field larger than this is assembled in rank 0's HOST memory, message by message


def gather_demos(local: torch.Tensor, dst: int = 0) -> Optional[torch.Tensor]:
    """Concatenate per-rank demo records ``[B_rank, ...]`` (same trailing shape, any B_rank) on ``dst`` in rank order.
    Returns the concatenation on ``dst`` and None elsewhere.  Messages are cut to <= MAX_MSG_BYTES.  The frames of a whole box do
    not fit one GPU next to its own (8 x 1024 scenes x 3 rounds of 480 x 640 rgb + depth = 56 GiB on rank 0: the first 8-GPU run
    died there), so a field above HOST_ABOVE_BYTES is copied to the host as its messages arrive and returned as a CPU tensor."""
    world = _world()
    if world == 1:
        return local
    rank = dist.get_rank()
    local = local.contiguous()
    n = torch.tensor([local.shape[0]], device=local.device, dtype=torch.int64)
    sizes = [torch.zeros_like(n) for _ in range(world)]
    dist.all_gather(sizes, n)
    sizes = [int(s.item()) for s in sizes]
    row = 1
    for d in local.shape[1:]:
        row *= int(d)
    rows_per_msg = max(1, MAX_MSG_BYTES // max(1, row * local.element_size()))
    nmax = max(sizes)
    to_host = sum(sizes) * row * local.element_size() > HOST_ABOVE_BYTES
    parts: List[List[torch.Tensor]] = [[] for _ in range(world)]
    host, starts = None, [0] * world
    if to_host and rank == dst:      # one pinned buffer for the whole field: message parts land in it with asynchronous copies
        for r in range(1, world):
            starts[r] = starts[r - 1] + sizes[r - 1]
        try:
            host = torch.empty((sum(sizes),) + tuple(local.shape[1:]), dtype=local.dtype, pin_memory=local.is_cuda)
        except RuntimeError:
            host = torch.empty((sum(sizes),) + tuple(local.shape[1:]), dtype=local.dtype)
    for lo in range(0, nmax, rows_per_msg):
        hi = min(nmax, lo + rows_per_msg)
        buf = torch.zeros((hi - lo,) + tuple(local.shape[1:]), device=local.device, dtype=local.dtype)
        k = max(0, min(local.shape[0], hi) - lo)
        if k > 0:
            buf[:k] = local[lo:lo + k]
        out = [torch.empty_like(buf) for _ in range(world)] if rank == dst else None
        dist.gather(buf, out, dst=dst)
        if rank == dst:
            for r in range(world):
                kr = max(0, min(sizes[r], hi) - lo)
                if kr > 0 and to_host:
                    host[starts[r] + lo:starts[r] + lo + kr].copy_(out[r][:kr], non_blocking=True)
                elif kr > 0:
                    parts[r].append(out[r][:kr])
            if to_host and local.is_cuda:
                torch.cuda.current_stream(local.device).synchronize()     # the message buffers are reused by the next round
    if rank != dst:
        return None
    if to_host:
        return host
    flat = [t for r in range(world) for t in parts[r]]
    return torch.cat(flat, 0) if flat else local[:0]
