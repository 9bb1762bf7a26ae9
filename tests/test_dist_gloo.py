"""world_size-2 gloo tests of the N>1 host logic: contiguous scene shards, the episode-statistics
all_reduce and the demo gather to rank 0.  CPU only; the data path itself has no collective."""
import os
import socket
import sys

import pytest
import torch
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close(); return p


def _worker(rank, world, port, q):
    sys.path.insert(0, ROOT)
    os.environ.update(RANK=str(rank), WORLD_SIZE=str(world), LOCAL_RANK=str(rank), MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    from mujoco_robot_environments_b200 import dist as D
    import torch.distributed as dist
    r, w, _ = D.init("gloo")
    assert (r, w) == (rank, world)
    n_total = 11
    lo, hi = D.shard_range(n_total, r, w)
    # every rank "rolls out" its shard: demo record = [scene id, step, 7 action reals]; stats = [episodes, successes, steps]
    demos = torch.stack([torch.full((9,), float(e)) for e in range(lo, hi)]).reshape(hi - lo, 9)
    stats = torch.tensor([hi - lo, (hi - lo) // 2, 100.0 * (hi - lo)], dtype=torch.float64)
    D.reduce_stats(stats)
    t = torch.tensor([1.0 + rank], dtype=torch.float64); D.max_over_ranks(t)
    old = D.MAX_MSG_BYTES
    D.MAX_MSG_BYTES = 2 * 9 * 4          # force chunking: 2 rows per message
    got = D.gather_demos(demos, dst=0)
    D.HOST_ABOVE_BYTES, keep = 1, D.HOST_ABOVE_BYTES          # force the host-assembly path (fields too large for rank 0's GPU)
    got_host = D.gather_demos(demos, dst=0)
    D.HOST_ABOVE_BYTES = keep
    assert (got is None) == (got_host is None) and (got is None or (got_host.device.type == "cpu" and bool((got_host == got.cpu()).all())))
    D.MAX_MSG_BYTES = old
    empty = D.gather_demos(demos[:0] if rank == 1 else demos[:1], dst=0)
    q.put((rank, lo, hi, stats.tolist(), t.item(), None if got is None else got[:, 0].tolist(), None if empty is None else empty[:, 0].tolist()))
    dist.barrier(); dist.destroy_process_group()


def test_shard_range_covers_everything():
    sys.path.insert(0, ROOT)
    from mujoco_robot_environments_b200.dist import shard_range
    for n in (1, 7, 16384, 16385):
        for w in (1, 2, 3, 8):
            r = [shard_range(n, k, w) for k in range(w)]
            assert r[0][0] == 0 and r[-1][1] == n
            assert all(r[k][1] == r[k + 1][0] for k in range(w - 1))
            assert max(b - a for a, b in r) - min(b - a for a, b in r) <= 1


@pytest.mark.timeout(120)
def test_two_rank_stats_and_gather():
    world, port = 2, _free_port()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    ps = [ctx.Process(target=_worker, args=(r, world, port, q)) for r in range(world)]
    [p.start() for p in ps]
    res = sorted(q.get(timeout=100) for _ in range(world))
    [p.join(timeout=30) for p in ps]
    assert all(p.exitcode == 0 for p in ps)
    (r0, lo0, hi0, s0, t0, g0, e0), (r1, lo1, hi1, s1, t1, g1, e1) = res
    assert (lo0, hi0, lo1, hi1) == (0, 6, 6, 11)
    assert s0 == s1 == [11.0, 5.0, 1100.0]
    assert t0 == t1 == 2.0
    assert g0 == [float(e) for e in range(11)] and g1 is None          # rank order preserved across chunked messages
    assert e0 == [0.0] and e1 is None                                    # a rank may contribute zero records
