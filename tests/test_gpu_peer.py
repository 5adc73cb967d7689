"""The one-launch end-of-step exchange (csrc/peer.cu, dls_peer_sgd_step): multi-tensor SGD on one GPU, and on two GPUs the
sum over NVLink peer memory -- bit-exact against numpy (rank-ordered fp32 adds, `w - g * lr` with one rounding per
operation), across segment splits, ragged lengths, staging overflow and graph replays."""
import ctypes as C
import os
import sys

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
F = np.float32
SIZES = [1, 3, 4, 1023, 1024, 1025, 4099, 70001, 5, 17, 300000, 2, 8192, 33]      # 14 segments: more than one launch holds


def bits(a):
    return np.ascontiguousarray(a, dtype=F).view(np.uint32)


def arrays(seed, sizes):
    rng = np.random.RandomState(seed)
    return [rng.standard_normal(n).astype(F) for n in sizes]


def sgd(w, g, lr):
    return (w - (g * F(lr)).astype(F)).astype(F)


def call_step(be, N, weights, grads, sizes, exchange, lr):
    n = len(grads)
    wh = (C.c_void_p * n)(*[b.handle for b in weights]) if weights is not None else None
    gh = (C.c_void_p * n)(*[b.handle for b in grads])
    counts = (C.c_size_t * n)(*sizes)
    flags = (C.c_int * n)(*exchange) if exchange is not None else None
    N.check(N.lib.dls_peer_sgd_step(be.ctx, wh, gh, counts, flags, n, lr))


def test_multi_tensor_sgd_on_one_gpu_is_bit_exact():
    import deeplearning.scala_b200.tensors as B
    from deeplearning.scala_b200 import _native as N
    be = B.Backend()
    try:
        ws, gs = arrays(1, SIZES), arrays(2, SIZES)
        wb, gb = [be.upload(w) for w in ws], [be.upload(g) for g in gs]
        launches0 = be.stats()["kernel_launches"]
        call_step(be, N, wb, gb, SIZES, None, 0.125)
        assert be.stats()["kernel_launches"] - launches0 == 2          # 14 segments, 12 per launch
        for w, g, b, n in zip(ws, gs, wb, SIZES):
            assert (bits(be.download(b, n)) == bits(sgd(w, g, 0.125))).all()
        for g, b, n in zip(gs, gb, SIZES):
            assert (bits(be.download(b, n)) == bits(g)).all()          # gradients untouched
        with pytest.raises(ValueError):
            call_step(be, N, None, gb, SIZES, None, 0.1)                # nothing to do
        with pytest.raises(N.NativeError):
            call_step(be, N, wb, gb, SIZES, [1] * len(SIZES), 0.1)      # exchange without dls_peer_init
    finally:
        be.close()


def _setup(rank, world, uid_queue, staging):
    import faulthandler
    faulthandler.dump_traceback_later(120, exit=True)                   # a rank that hangs says where (stderr), then exits 1
    sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    import deeplearning.scala_b200.tensors as B
    import deeplearning.scala_b200.harness as H
    from deeplearning.scala_b200 import _native as N
    be = B.Backend(device=rank, gemm_mode="fp32")
    be.ctx
    if rank == 0:
        uid = H.Exchange.unique_id()
        for _ in range(world - 1):
            uid_queue.put(uid)
    else:
        uid = uid_queue.get(timeout=120)
    if staging is not None:
        H.Exchange.peer_staging_floats = staging
    ex = H.Exchange(be, rank, world, uid)
    ex.verify()
    return B, H, N, be, ex


def _abi_rank(rank, world, uid_queue, out_queue):
    try:
        B, H, N, be, ex = _setup(rank, world, uid_queue, 8192)          # 8 chunks per staging half: the 70001 / 300000 segments split
        assert ex.peer, N.last_error()
        ws = arrays(1, SIZES)                                           # replicas: the same weights on every rank
        gs = arrays(10 + rank, SIZES)                                   # a different gradient per rank
        wb, gb = [be.upload(w) for w in ws], [be.upload(g) for g in gs]
        flags = [0 if n == 1024 else 1 for n in SIZES]                  # one segment is "already reduced"
        call_step(be, N, wb, gb, SIZES, flags, 0.25)
        first = [be.download(b, n) for b, n in zip(wb, SIZES)]
        call_step(be, N, None, gb, SIZES, [1] * len(SIZES), 0.0)        # all-reduce only: sums land in the gradients
        sums = [be.download(b, n) for b, n in zip(gb, SIZES)]
        # replays: the epoch counters advance on the device
        g2 = [be.upload(g) for g in arrays(20 + rank, SIZES)]
        with be.capture() as graph:
            call_step(be, N, wb, g2, SIZES, flags, 0.25)
        for _ in range(3):
            graph.launch()
        be.synchronize()
        replayed = [be.download(b, n) for b, n in zip(wb, SIZES)]
        out_queue.put((rank, first, sums, replayed))
        be.close()
    except BaseException as exc:                                        # the parent must not wait for a dead rank
        out_queue.put((rank, repr(exc), None, None))
        raise


def _model_rank(rank, world, uid_queue, out_queue, peer_bytes, compiled, optimizer=None):
    try:
        os.environ["DLS_PEER_BYTES"] = str(peer_bytes)
        os.environ["DLS_PEER_REDUCE"] = "1"
        B, H, N, be, ex = _setup(rank, world, uid_queue, None)
        assert ex.peer == (peer_bytes > 0)
        opt = None
        if optimizer == "adam":
            import deeplearning.scala_b200.optimizers as Opt
            opt = Opt.Adam(1e-2)
        model = H.Mlp((24, 16, 6), bias=True, backend=be, learning_rate=0.05, exchange=ex, optimizer=opt)
        rng = np.random.RandomState(5)
        x = rng.rand(24, 16).astype(F)
        t = H.one_hot(rng.randint(0, 6, size=16), 6, feature_major=True)
        xd, td = be.tensor(H.shard(x, 1, rank, world)).do_cache(), be.tensor(H.shard(t, 1, rank, world)).do_cache()
        losses = []
        if compiled:
            step = model.compile(xd, td)                                # one eager step, then the recording
            losses = [step() for _ in range(3)]
        else:
            losses = [model.step(xd, td) for _ in range(4)][1:]
        launches0 = be.stats()["kernel_launches"]
        if not compiled:
            model.step(xd, td)
        per_step = be.stats()["kernel_launches"] - launches0
        out_queue.put((rank, losses, [p.copy() for p in model.parameters()], per_step))
        be.close()
    except BaseException as exc:
        out_queue.put((rank, repr(exc), None, None))
        raise


def spawn(target, *args):
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs (ranks must not share a device)")
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    uid_q, out_q = ctx.Queue(), ctx.Queue()
    procs = [ctx.Process(target=target, args=(r, 2, uid_q, out_q) + args) for r in range(2)]
    for p in procs:
        p.start()
    results = sorted([out_q.get(timeout=300) for _ in procs], key=lambda r: r[0])
    stuck = []
    for p in procs:
        p.join(timeout=150)
        if p.is_alive():                                                # never leave a child behind: pytest would wait for it at exit
            stuck.append(p.pid)
            p.kill()
            p.join(timeout=10)
    assert not stuck, "rank process(es) %r did not exit after delivering their results" % stuck
    assert [p.exitcode for p in procs] == [0, 0], "a rank did not shut down cleanly (see its traceback on stderr)"
    for r in results:
        assert not isinstance(r[1], str), "rank %d failed: %s" % (r[0], r[1])
    return results


def test_peer_exchange_abi_on_two_gpus():
    (_, first0, sums0, rep0), (_, first1, sums1, rep1) = spawn(_abi_rank)
    ws, g0, g1 = arrays(1, SIZES), arrays(10, SIZES), arrays(11, SIZES)
    h0, h1 = arrays(20, SIZES), arrays(21, SIZES)
    for i, n in enumerate(SIZES):
        total = (g0[i] + g1[i]).astype(F)
        # the "already reduced" segment is updated from each rank's own gradient
        want0 = sgd(ws[i], g0[i] if n == 1024 else total, 0.25)
        want1 = sgd(ws[i], g1[i] if n == 1024 else total, 0.25)
        assert (bits(first0[i]) == bits(want0)).all(), "segment %d (%d floats), rank 0" % (i, n)
        assert (bits(first1[i]) == bits(want1)).all(), "segment %d (%d floats), rank 1" % (i, n)
        assert (bits(sums0[i]) == bits(total)).all() and (bits(sums1[i]) == bits(total)).all()
        w0, w1 = want0, want1
        for _ in range(3):
            t2 = (h0[i] + h1[i]).astype(F)
            w0, w1 = sgd(w0, h0[i] if n == 1024 else t2, 0.25), sgd(w1, h1[i] if n == 1024 else t2, 0.25)
        assert (bits(rep0[i]) == bits(w0)).all() and (bits(rep1[i]) == bits(w1)).all(), "replays, segment %d" % i


@pytest.mark.parametrize("compiled", [False, True])
def test_peer_and_nccl_exchange_train_the_same_replicas(compiled):
    """Two ranks: a + b is commutative and the update arithmetic is the same, so the one-launch peer path and the NCCL
    path produce the same bits; the peer path needs one launch for the whole end of the step."""
    peer = spawn(_model_rank, 1 << 20, compiled)
    nccl = spawn(_model_rank, 0, compiled)
    (_, pl0, pp0, launches_peer), (_, pl1, pp1, _) = peer
    (_, nl0, np0, launches_nccl), (_, nl1, np1, _) = nccl
    assert pl0 == nl0 and pl1 == nl1 and len(set(pl0)) > 1
    for a, b, c, d in zip(pp0, pp1, np0, np1):
        assert (bits(a) == bits(b)).all()                               # replicas
        assert (bits(a) == bits(c)).all() and (bits(b) == bits(d)).all()  # paths
    if not compiled:
        assert launches_peer < launches_nccl


@pytest.mark.parametrize("compiled", [False, True])
def test_stateful_optimizer_takes_its_small_gradients_from_the_peer_sum(compiled):
    """Adam under data parallelism: `Exchange.reduce` holds the small gradients for one peer-memory all-reduce launch
    (dls_peer_sgd_step without weights) instead of NCCL; at two ranks both routes give the same bits."""
    peer = spawn(_model_rank, 1 << 20, compiled, "adam")
    nccl = spawn(_model_rank, 0, compiled, "adam")
    (_, pl0, pp0, _), (_, pl1, pp1, _) = peer
    (_, nl0, np0, _), (_, nl1, np1, _) = nccl
    assert pl0 == nl0 and pl1 == nl1 and len(set(pl0)) > 1
    for a, b, c, d in zip(pp0, pp1, np0, np1):
        assert (bits(a) == bits(b)).all()
        assert (bits(a) == bits(c)).all() and (bits(b) == bits(d)).all()
