"""Training-step legs of bench.py: BASELINE.json configs C1 (MLP 784-500-10, batch 128), C3 (CIFAR-shaped CNN, batch 256),
C4 (wide MLP 4 x 4096, batch 8192) and C5 (C3 data parallel).  One step = forward + backward + SGD update of every
weight (`loss.train`, TL:70-76; `data - lr * delta`, TW:50-74) on synthetic data of the named shapes.

`run(...)` returns {"samples_per_s": device-resident inputs, "e2e_samples_per_s": host inputs in pinned memory copied in
every step + the scalar loss read back, ...}.  Timing: CUDA events on the launch stream for the resident leg, host clock
around blocking steps for the e2e leg (each step ends with a D2H read).
"""
import ctypes
import time

import numpy as np


def pinned_like(N, array):
    p = ctypes.POINTER(ctypes.c_float)()
    N.check(N.lib.dls_host_alloc(array.size, ctypes.byref(p)))
    view = np.ctypeslib.as_array(p, shape=(array.size,)).reshape(array.shape)
    view[...] = array
    return p, view


# `bench.py --workload c5 --strong`: SURVEY 8(d)'s strong-scaling variant -- a fixed global batch sharded over the ranks
# (None = weak scaling, 256 samples per GPU)
C5_GLOBAL_BATCH = None


def build(workload, be, exchange=None, rank=0, world=1, gemm_mode=None):
    """-> (model, x_host, target_host or None, per-rank batch, flop per step per rank, description)"""
    import deeplearning.scala_b200.harness as H
    rng = np.random.RandomState(7 + rank)
    if workload == "c1":
        batch = 128
        model = H.Mlp((784, 500, 10), bias=True, last_activation="none", backend=be, learning_rate=0.01 / batch, exchange=exchange)
        x = rng.rand(784, batch).astype(np.float32)
        t = H.one_hot(rng.randint(0, 10, size=batch), 10, feature_major=True)
        flop = 2.0 * batch * (784 * 500 + 500 * 10) * 3 - 2.0 * batch * 784 * 500          # no dX for layer 1
        desc = "C1: MLP 784-500-10 (tanh hidden, squared error), batch 128, feature-major, SGD lr 0.01 on the mean loss (= 0.01/128 on the reference's implicit sum loss)"
    elif workload == "c4":
        batch, width, depth = 8192, 4096, 4
        model = H.Mlp((width,) * (depth + 1), bias=False, last_activation="tanh", backend=be, learning_rate=1e-4, exchange=exchange)
        x = rng.standard_normal((width, batch)).astype(np.float32)
        t = None
        flop = 2.0 * batch * width * width * (3 * depth - 1)
        desc = "C4: wide MLP 4 x (4096x4096) tanh, batch 8192, loss = sum(out), 11 GEMMs of 8192x4096x4096 per step"
    elif workload in ("c3", "c5"):
        batch = 256
        if workload == "c5" and C5_GLOBAL_BATCH is not None:
            if C5_GLOBAL_BATCH % world:
                raise ValueError("global batch %d is not divisible by %d ranks" % (C5_GLOBAL_BATCH, world))
            batch = C5_GLOBAL_BATCH // world
        model = H.Cnn(image=32, channels=(3, 32, 64, 128), hidden=256, classes=10, backend=be, learning_rate=1e-3 / world, exchange=exchange)
        x = rng.rand(batch, 32, 32, 3).astype(np.float32)
        t = H.one_hot(rng.randint(0, 10, size=batch), 10, feature_major=False)
        conv = [(32 * 32, 3, 32), (16 * 16, 32, 64), (8 * 8, 64, 128)]
        fwd = sum(2.0 * batch * hw * 9 * ci * co for hw, ci, co in conv) + 2.0 * batch * (8192 * 256 + 256 * 10)
        flop = 3.0 * fwd - 2.0 * batch * 32 * 32 * 27 * 32
        desc = ("C3: CNN 3 conv3x3 (3-32-64-128, ReLU, 2x nearest down-sampling) via im2col + 2 FC (8192-256-10), "
                "batch %d NHWC 32x32x3 per GPU, squared error, SGD" % batch)
        if workload == "c5":
            desc = "C5: " + desc[4:] + "; data parallel, " + ("weak" if C5_GLOBAL_BATCH is None else "strong") + " scaling, SUM all-reduce of every weight gradient (NCCL + NVLink peer memory)"
    else:
        raise ValueError(workload)
    return model, x, t, batch, flop, desc


def run(workload, be, steps, warmup, e2e_steps=5, exchange=None, rank=0, world=1, barrier=None, compiled=True):
    """Times the step twice: eagerly (one host dispatch per launch, `model.step`) and, with `compiled`, as a captured step
    (`model.compile`: the same user code recorded once into a CUDA graph, one driver call per step).  The headline
    numbers (`ms_total`, `e2e_ms_per_step`) are the captured step's when it is available, else the eager ones."""
    import deeplearning.scala_b200.tensors as B
    from deeplearning.scala_b200 import _native as N
    model, x, t, batch, flop, desc = build(workload, be, exchange, rank, world)
    barrier = barrier or be.synchronize
    xd = be.tensor(x).do_cache()
    td = be.tensor(t).do_cache() if t is not None else None
    losses = []
    for _ in range(warmup):
        losses.append(model.step(xd, td))
    barrier()
    st0 = be.stats()
    e0 = be.record_event()
    for _ in range(steps):
        losses.append(model.step(xd, td))
    e1 = be.record_event()
    barrier()
    eager_ms = B.elapsed_ms(e0, e1)
    st1 = be.stats()
    # e2e: the step a user makes -- host batch (pinned) -> Tensor.apply -> train -> loss on the host
    px, hx = pinned_like(N, x)
    pt, ht = pinned_like(N, t) if t is not None else (None, None)
    model.step(be.tensor(hx), be.tensor(ht) if t is not None else None)
    barrier()
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        losses.append(model.step(be.tensor(hx), be.tensor(ht) if t is not None else None))
    be.synchronize()
    eager_e2e_s = time.perf_counter() - t0
    out = {"workload": desc, "batch_per_gpu": batch, "flop_per_step_per_gpu": flop,
           "eager": {"ms_per_step": eager_ms / steps, "e2e_ms_per_step": eager_e2e_s * 1e3 / e2e_steps,
                     "launches_per_step": (st1["kernel_launches"] - st0["kernel_launches"]) / steps},
           "driver_allocs_in_timed_region": st1["driver_allocs"] - st0["driver_allocs"],
           "h2d_bytes_per_step": int(x.nbytes + (t.nbytes if t is not None else 0)), "d2h_bytes_per_step": 4,
           "steps": steps, "e2e_steps": e2e_steps, "mode": "eager"}
    ms, e2e_s, launches = eager_ms, eager_e2e_s, out["eager"]["launches_per_step"]
    if compiled:
        step = model.compile(xd, td)                           # one eager step, then the recording
        for _ in range(warmup):
            step.launch()
        losses.append(step.loss())
        barrier()
        st0 = be.stats()
        e0 = be.record_event()
        for _ in range(steps):
            step.launch()
        e1 = be.record_event()
        barrier()
        ms = B.elapsed_ms(e0, e1)
        st1 = be.stats()
        losses.append(step.loss())
        # e2e: pinned host batch written into the step's input buffers, one replay, the scalar loss read back
        step(hx, ht)
        barrier()
        t0 = time.perf_counter()
        for _ in range(e2e_steps):
            losses.append(step(hx, ht))
        be.synchronize()
        e2e_s = time.perf_counter() - t0
        info = step.graph.info()
        launches = (st1["kernel_launches"] - st0["kernel_launches"]) / steps
        out["captured"] = {"ms_per_step": ms / steps, "e2e_ms_per_step": e2e_s * 1e3 / e2e_steps, "launches_per_step": launches,
                           "graph_nodes": info["nodes"], "arena_bytes": info["arena_bytes"]}
        out["driver_allocs_in_timed_region"] += st1["driver_allocs"] - st0["driver_allocs"]
        if ms <= eager_ms:
            out["mode"] = "captured step (CUDA graph replay of the recorded loss.train)"
        else:                                                  # device-bound steps gain nothing from the recording and pay
            ms, e2e_s, launches = eager_ms, eager_e2e_s, out["eager"]["launches_per_step"]    # the weight copy-back
        del step
    N.lib.dls_host_free(px)
    if pt is not None:
        N.lib.dls_host_free(pt)
    n_params = int(sum(np.prod(w.data.shape) for w in model.weights))
    if exchange is not None:
        out["exchange"] = ("gradients >= %d bytes: grouped NCCL all-reduce on the exchange stream; smaller ones summed over NVLink peer "
                           "memory and every weight updated by one launch (dls_peer_sgd_step)" % exchange.peer_bytes
                           if exchange.peer else "grouped NCCL all-reduce on the exchange stream, one update launch per weight")
    out.update({"ms_per_step": ms / steps, "ms_total": ms, "e2e_ms_per_step": e2e_s * 1e3 / e2e_steps, "parameters": n_params,
                "launches_per_step": launches, "first_loss": losses[0], "last_loss": losses[-1]})
    return out
