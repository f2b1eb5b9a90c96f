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


# ----------------------------------------------------------------------------------------------------------------------
# extract_corpus over two ranks: sharding, per-day accumulation, ONE all-gather for all days, fixed-order merge, per-session
# normalisation (prepare_corpus.py:179-236).  The extraction / accumulation callables are numpy-level stand-ins here (the
# CUDA ones are covered by the GPU tests and by bench.py's parity check at N > 1).
# ----------------------------------------------------------------------------------------------------------------------
def _corpus_sessions(C=6, n=10):
    rng = np.random.default_rng(7)
    return [torch.from_numpy(rng.normal(3 + 0.1 * s, 0.5, (50 + 11 * s, C))) for s in range(n)]


def _corpus_worker(rank, world, port, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from delayed_speech_synthesis_b200 import corpus
    C = 6
    sess = _corpus_sessions(C)
    day_of = [s // 4 for s in range(len(sess))]                 # days 0, 0, 0, 0, 1, 1, 1, 1, 2, 2
    baseline = [0, 1, 4, 9]                                     # day 0 has two baseline sessions, on different ranks

    def extract(x, mu, sigma):
        return x if mu is None else (x - mu) / sigma

    def accumulate(f):
        return torch.stack([torch.full((C,), float(len(f)), dtype=torch.float64), f.mean(0), ((f - f.mean(0)) ** 2).sum(0)])

    got = {}
    mu, sigma, slots = corpus.extract_corpus(len(sess), day_of, baseline, lambda s: sess[s], extract, accumulate, n_channels=C,
                                             on_features=lambda s, f: got.__setitem__(s, f.numpy().copy()))
    # the one-pass baseline callback (HgaEngine.extract_offline_stats on a GPU) replaces accumulate(extract(x)): same statistics
    calls = []
    mu2, sigma2, _ = corpus.extract_corpus(len(sess), day_of, baseline, lambda s: sess[s], extract, None, n_channels=C,
                                           extract_stats=lambda x: (calls.append(len(x)), accumulate(x))[1])
    assert torch.equal(mu, mu2) and torch.equal(sigma, sigma2)
    assert len(calls) == sum(1 for s in corpus.shard_weighted([2 if s in baseline else 1 for s in range(len(sess))], rank, world) if s in baseline)
    assert sorted(got) == corpus.shard_weighted([2 if s in baseline else 1 for s in range(len(sess))], rank, world)
    q.put((rank, mu.numpy().copy(), sigma.numpy().copy(), got))
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.timeout(120)
def test_extract_corpus_world2():
    world, port = 2, _free_port()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_corpus_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = {r: (mu, sg, got) for r, mu, sg, got in (q.get(timeout=100) for _ in range(world))}
    for p in procs:
        p.join(timeout=30)
        assert p.exitcode == 0
    assert np.array_equal(res[0][0], res[1][0]) and np.array_equal(res[0][1], res[1][1])      # identical on every rank
    sess = [s.numpy() for s in _corpus_sessions()]
    mu, sg = res[0][0], res[0][1]
    for d, members in enumerate(([0, 1], [4], [9])):
        cat = np.concatenate([sess[s] for s in members])
        np.testing.assert_allclose(mu[d], cat.mean(0), rtol=1e-13)
        np.testing.assert_allclose(sg[d], cat.std(0), rtol=1e-12)
    feats = {**res[0][2], **res[1][2]}
    assert sorted(feats) == list(range(len(sess)))
    for s, f in feats.items():
        np.testing.assert_allclose(f, (sess[s] - mu[s // 4]) / sg[s // 4], rtol=1e-13)
