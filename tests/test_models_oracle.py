"""Pins oracle/models_oracle.py: its analytic SGD step must agree with central finite differences of its own loss
(the reference has no golden vectors for whole models; its autodiff rules are pinned op by op in test_oracle_golden.py),
and a 2-rank data-parallel step (batch shards, SUM all-reduce over gloo) must equal the full-batch step."""
import os
import socket
import sys

import numpy as np
import pytest

from oracle import models_oracle as MO

LR = 0.5


def fd_check(loss_of, params, new_params, picks, eps=2e-3, tol=3e-2):
    rng = np.random.RandomState(0)
    for p, q in zip(params, new_params):
        if p is None:
            continue
        grad = (p.astype(np.float64) - q.astype(np.float64)) / LR
        for _ in range(picks):
            idx = tuple(rng.randint(0, s) for s in p.shape)
            old = p[idx]
            p[idx] = old + eps
            up = loss_of()
            p[idx] = old - eps
            down = loss_of()
            p[idx] = old
            fd = (up - down) / (2 * eps)
            assert abs(fd - grad[idx]) <= tol * max(1.0, abs(fd)), (idx, fd, grad[idx])


def test_mlp_oracle_matches_finite_differences():
    rng = np.random.RandomState(1)
    sizes, batch = (6, 5, 4), 3
    weights = [(rng.randn(o, i).astype(np.float32) * 0.5, rng.randn(o).astype(np.float32) * 0.1)
               for i, o in zip(sizes[:-1], sizes[1:])]
    x = rng.rand(sizes[0], batch).astype(np.float32)
    t = rng.rand(sizes[-1], batch).astype(np.float32)
    _, new, _ = MO.mlp_step(weights, x, t, LR)
    flat = [a for pair in weights for a in pair]
    flat_new = [a for pair in new for a in pair]
    fd_check(lambda: MO.mlp_step(weights, x, t, LR)[0], flat, flat_new, picks=6)


def test_cnn_oracle_matches_finite_differences():
    rng = np.random.RandomState(2)
    chans = (2, 3, 4)
    convs = [(rng.randn(3, 3, ci, co).astype(np.float32) * 0.3, rng.rand(co).astype(np.float32) * 0.2 + 0.1)
             for ci, co in zip(chans[:-1], chans[1:])]
    flat_n = 2 * 2 * chans[-1]
    fc1 = (rng.randn(flat_n, 5).astype(np.float32) * 0.3, rng.rand(5).astype(np.float32) * 0.2 + 0.1)
    fc2 = (rng.randn(5, 3).astype(np.float32) * 0.3, rng.randn(3).astype(np.float32) * 0.1)
    x = rng.rand(2, 4, 4, chans[0]).astype(np.float32)
    t = rng.rand(2, 3).astype(np.float32)
    _, nconvs, nfc1, nfc2 = MO.cnn_step(convs, fc1, fc2, x, t, LR)
    flat = [a for pair in convs + [fc1, fc2] for a in pair]
    flat_new = [a for pair in nconvs + [nfc1, nfc2] for a in pair]
    fd_check(lambda: MO.cnn_step(convs, fc1, fc2, x, t, LR)[0], flat, flat_new, picks=5)


def _dp_worker(rank, world, port, queue):
    import torch
    import torch.distributed as dist
    sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    from deeplearning.scala_b200.harness import shard
    dist.init_process_group("gloo", init_method="tcp://127.0.0.1:%d" % port, rank=rank, world_size=world)
    rng = np.random.RandomState(3)
    sizes, batch = (8, 7, 5), 8
    weights = [(rng.randn(o, i).astype(np.float32) * 0.5, rng.randn(o).astype(np.float32) * 0.1)
               for i, o in zip(sizes[:-1], sizes[1:])]
    x = rng.rand(sizes[0], batch).astype(np.float32)
    t = rng.rand(sizes[-1], batch).astype(np.float32)
    lr = 0.1
    _, full, _ = MO.mlp_step(weights, x, t, lr)
    _, local, _ = MO.mlp_step(weights, shard(x, 1, rank, world), shard(t, 1, rank, world), lr)
    worst = 0.0
    for (w, b), (lw, lb), (fw, fb) in zip(weights, local, full):
        for p, l, f in ((w, lw, fw), (b, lb, fb)):
            delta = torch.from_numpy((p.astype(np.float64) - l.astype(np.float64)))
            dist.all_reduce(delta, op=dist.ReduceOp.SUM)      # SUM, not mean: the implicit loss is a sum over the batch
            dp = p.astype(np.float64) - delta.numpy()
            worst = max(worst, float(np.abs(dp - f).max() / (np.abs(f).max() + 1e-30)))
    dist.destroy_process_group()
    queue.put((rank, worst))


def test_data_parallel_shards_sum_to_the_full_batch_step():
    import torch.multiprocessing as mp
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    ctx = mp.get_context("spawn")
    queue = ctx.Queue()
    procs = [ctx.Process(target=_dp_worker, args=(r, 2, port, queue)) for r in range(2)]
    for p in procs:
        p.start()
    results = [queue.get(timeout=120) for _ in procs]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert all(w <= 1e-5 for _, w in results), results


def test_elementwise_rule_oracle_matches_finite_differences():
    """divide / min / max / abs / sqrt / sum / permute / broadcast -- the composite of MO.ops_step, pinned by
    central finite differences (inputs kept away from the kinks of abs / min / max and from b = 0)."""
    rng = np.random.RandomState(3)
    a = (rng.rand(4, 6).astype(np.float32) + 0.5) * np.where(rng.rand(4, 6) < 0.5, -1, 1).astype(np.float32)
    b = rng.rand(4, 6).astype(np.float32) + 0.7
    c = (rng.rand(4).astype(np.float32) - 0.5) * 2
    # keep every comparison decided by a margin larger than the finite-difference step
    assert np.abs(a - c[:, None]).min() > 0.02 and np.abs(b - np.maximum(a, c[:, None])).min() > 0.02
    _, na, nb, nc, _ = MO.ops_step(a, b, c, LR)
    fd_check(lambda: MO.ops_step(a, b, c, LR)[0], [a, b, c], [na, nb, nc], picks=8, eps=2e-3, tol=3e-2)


def test_left_broadcast_and_sum_as_are_adjoint():
    rng = np.random.RandomState(4)
    v = rng.randn(3).astype(np.float32)
    d = rng.randn(3, 5, 2).astype(np.float32)
    lhs = float((MO.left_broadcast(v, d.shape).astype(np.float64) * d).sum())
    rhs = float((v.astype(np.float64) * MO.sum_as(d, (3,)).astype(np.float64)).sum())
    assert abs(lhs - rhs) <= 1e-5 * (abs(lhs) + 1)
