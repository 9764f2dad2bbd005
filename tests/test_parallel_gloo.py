"""CPU, world_size 2 over gloo: the N>1 host logic (shard ownership, one-time key broadcast, max-over-ranks timing)."""
import os

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp


def _worker(rank, world, port, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from fhe_b200 import parallel as par
    lengths = {"bsk": 4096, "kskA": 1000, "kskB": 10}
    arrays = None
    if rank == 0:
        rng = np.random.default_rng(0)
        arrays = {k: rng.integers(0, 2**63, size=v, dtype=np.uint64) for k, v in lengths.items()}
    keys = par.broadcast_keys(arrays, lengths, src=0)
    chk = {k: int(v.sum().item()) for k, v in keys.items()}
    lo, hi = par.shard_range(1001, rank, world)
    total = par.gather_counts(hi - lo)
    tmax = par.max_over_ranks(1.0 + rank)
    # a circuit batch shards over its input sets (count axis), all wires on every rank
    ia = np.arange(3 * 11 * 4, dtype=np.uint64).reshape(3, 11, 4); ib = np.arange(33, dtype=np.uint64).reshape(3, 11)
    sa, sb, (clo, chi) = par.shard_circuit_inputs(ia, ib, rank, world)
    assert sa.shape == (3, chi - clo, 4) and (sa == ia[:, clo:chi]).all() and (sb == ib[:, clo:chi]).all()
    assert par.gather_counts(chi - clo) == 11
    q.put((rank, chk, lo, hi, total, tmax))
    dist.destroy_process_group()


def test_shard_and_broadcast_world2():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29500 + os.getpid() % 2000
    ps = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in ps:
        p.start()
    res = sorted(q.get(timeout=120) for _ in range(2))
    for p in ps:
        p.join(timeout=60)
        assert p.exitcode == 0
    (r0, chk0, lo0, hi0, tot0, t0), (r1, chk1, lo1, hi1, tot1, t1) = res
    assert chk0 == chk1                       # every rank holds rank 0's keys
    assert (lo0, hi0, lo1, hi1) == (0, 501, 501, 1001) and tot0 == tot1 == 1001
    assert t0 == t1 == 2.0                    # max over ranks


def test_shard_range_covers_batch():
    from fhe_b200.parallel import shard_range
    for count in (0, 1, 7, 65536, 65537):
        for world in (1, 2, 4, 8):
            spans = [shard_range(count, r, world) for r in range(world)]
            assert spans[0][0] == 0 and spans[-1][1] == count
            assert all(spans[i][1] == spans[i + 1][0] for i in range(world - 1))
            assert max(h - l for l, h in spans) - min(h - l for l, h in spans) <= 1
