"""World-size-2 (and 3) gloo runs of the row-sharded protocol on the CPU: the numpy model of the kernels
(tests/_sharded_model.py) with torch.distributed all_reduce as the exchange, every rank checked against
oracle B on the rows it owns.  Covers: equal-area split, side-state sync, speculative column exchange,
rounds longer than the panel (m > max_round), failed speculation (fault injection), new landmarks."""
import os
import sys

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)


def _scenario(N=30, count=26, seed=3):
    rng = np.random.default_rng(seed)
    n = 3 + 2 * N
    side = int(np.ceil(np.sqrt(count)))
    state = np.zeros(n)
    k = np.arange(count)
    state[3:3 + 2 * count:2] = k % side + rng.uniform(-0.2, 0.2, count)
    state[4:4 + 2 * count:2] = k // side + rng.uniform(-0.2, 0.2, count)
    U = 0.01 * rng.uniform(-1, 1, (n, 6))
    S = U @ U.T + np.diag(1e-3 * rng.uniform(1, 2, n))
    act = 3 + 2 * count
    S[act:, :] = 0.0; S[:, act:] = 0.0
    S[np.arange(act, n), np.arange(act, n)] = 1e7
    return state, S, count, side


def _worker(rank, world, port, drop_guess, max_round, m, out):
    sys.path[:0] = [ROOT, HERE, os.path.join(ROOT, "oracle")]
    import torch
    import torch.distributed as dist
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        import struct_oracle as so
        from _sharded_model import ShardedModel

        def allreduce(buf):
            t = torch.from_numpy(buf)
            dist.all_reduce(t, op=dist.ReduceOp.SUM)

        state, S, count, side = _scenario()
        N = (len(state) - 3) // 2
        g = ShardedModel(rank, world, allreduce, n_landmarks=N, max_round=max_round, drop_guess=drop_guess)
        g.mu[:] = state; g.cnt = count; g.set_sigma(S); g.sync()
        o = so.StructEkf(n_landmarks=N)
        o.state[:] = state; o.sigma[:, :] = S; o.count = count
        rng = np.random.default_rng(11)
        pose = np.array([0.3, side / 2.0, side / 2.0])
        g.prev = pose[1:3].copy(); o.set_prev(pose[1], pose[2])
        for s in range(3):
            pose = pose + np.array([0.05, 0.1, 0.07])
            lm = o.state[3:3 + 2 * count].reshape(-1, 2)
            near = rng.permutation(count)[:m]
            c, sn = np.cos(pose[0]), np.sin(pose[0])
            dx, dy = lm[near, 0] - pose[1], lm[near, 1] - pose[2]
            xy = np.stack([c * dx + sn * dy, -sn * dx + c * dy], axis=1) + rng.normal(0, 1e-3, (m, 2))
            if s == 2:
                xy[:2] += 0.45                      # two sightings of nothing mapped -> new landmarks
            aG = g.scan(pose, xy)
            o.predict(*pose)
            aO = o.update_unknown(xy)
            assert list(aG) == list(aO), (rank, s, aG, list(aO))
            assert g.cnt == o.count
        rb, rows = g.owned_rows()
        ref = np.tril(o.sigma)[rb:rb + rows.shape[0]]
        err = np.linalg.norm(rows - ref) / np.linalg.norm(ref)
        assert err <= 1e-9, (rank, err)
        assert np.linalg.norm(g.mu - o.state) / np.linalg.norm(o.state) <= 1e-9
        # the rows of all ranks tile the matrix exactly
        lens = [None] * world
        dist.all_gather_object(lens, (g.rb, g.re))
        assert lens[0][0] == 0 and lens[-1][1] == g.n and all(lens[i][1] == lens[i + 1][0] for i in range(world - 1))
        from ekf_slam_b200 import shard
        rng_f = [None] * world
        dist.all_gather_object(rng_f, shard.filter_range(65536, rank, world))
        assert rng_f[0][0] == 0 and rng_f[-1][1] == 65536 and all(rng_f[i][1] == rng_f[i + 1][0] for i in range(world - 1))
        out.put((rank, g.rounds))
    finally:
        dist.destroy_process_group()


def _run(world, drop_guess=-1, max_round=24, m=12, port=29631):
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.SimpleQueue()
    procs = [ctx.Process(target=_worker, args=(r, world, port, drop_guess, max_round, m, q)) for r in range(world)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(180)
        assert p.exitcode == 0, p.exitcode
    res = dict(q.get() for _ in range(world))
    assert len(set(res.values())) == 1          # every rank ran the same number of rounds
    return res[0]


def test_world2_one_round_per_scan():
    assert _run(2, port=29631) == 3


def test_world3_rounds_split_by_panel_capacity():
    assert _run(3, max_round=5, m=12, port=29641) == 9          # 12 sightings = rounds of 5 + 5 + 2


def test_world2_failed_speculation_resumes():
    assert _run(2, drop_guess=4, m=12, port=29651) > 3
