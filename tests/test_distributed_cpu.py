"""CPU, world_size 2 over gloo: the multi-rank plumbing of the baseline-statistics exchange (all-gather of (n, mean, M2)
accumulators, rank-major order) and session sharding.  The merge arithmetic itself is a CUDA kernel (GPU tests); here the
gathered blocks are checked with the oracle's Chan merge."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from oracle import hga_oracle as O


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from delayed_speech_synthesis_b200.corpus import gather_accumulators, shard_sessions
    C, days = 5, 3
    rng = np.random.default_rng(100)
    sessions = [rng.normal(3, 0.5, (40 + 7 * s, C)) for s in range(6)]      # same on every rank
    mine = shard_sessions(len(sessions), rank, world)
    acc = torch.zeros((days, 3, C), dtype=torch.float64)
    for s in mine:                                                            # day = s % days
        f = sessions[s]
        cur = tuple(acc[s % days, i].numpy() for i in range(3))
        n, m, m2 = O.chan_merge(cur, (np.full(C, float(len(f))), f.mean(0), ((f - f.mean(0)) ** 2).sum(0)))
        acc[s % days] = torch.from_numpy(np.stack([n, m, m2]))
    g = gather_accumulators(acc)                                              # (world * days, 3, C), rank-major
    assert g.shape == (world * days, 3, C)
    g = g.reshape(world, days, 3, C)
    assert torch.equal(g[rank], acc)
    out = []
    for d in range(days):
        tot = (np.zeros(C), np.zeros(C), np.zeros(C))
        for r in range(world):
            tot = O.chan_merge(tot, tuple(g[r, d, i].numpy() for i in range(3)))
        out.append(np.stack([tot[1], np.sqrt(tot[2] / tot[0])]))
    q.put((rank, np.stack(out)))
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.timeout(120)
def test_stats_exchange_world2():
    world, port = 2, _free_port()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = dict(q.get(timeout=100) for _ in range(world))
    for p in procs:
        p.join(timeout=30)
        assert p.exitcode == 0
    assert np.array_equal(res[0], res[1])                                     # fixed order -> identical on every rank
    rng = np.random.default_rng(100)
    sessions = [rng.normal(3, 0.5, (40 + 7 * s, 5)) for s in range(6)]
    for d in range(3):
        mu, sd = O.baseline_statistics([sessions[s] for s in range(6) if s % 3 == d])
        np.testing.assert_allclose(res[0][d, 0], mu, rtol=1e-13)
        np.testing.assert_allclose(res[0][d, 1], sd, rtol=1e-12)
