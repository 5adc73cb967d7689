"""Training-step legs of bench.py: BASELINE.json configs C1 (MLP 784-500-10, batch 128), C3 (CIFAR-shaped CNN, batch 256),
C4 (wide MLP 4 x 4096, batch 8192) and C5 (C3 data parallel).  One step = forward + backward + SGD update of every
weight (`loss.train`, TL:70-76; `data - lr * delta`, TW:50-74) on synthetic data of the named shapes.

`run(...)` returns {"samples_per_s": device-resident inputs, "e2e_samples_per_s": host batches in pinned memory uploaded
every step + the scalar loss read back, ...}.  Timing: CUDA events on the launch stream for the resident leg, host
clock around the pipelined `CompiledStep.submit` loop for the e2e leg (every step uploads its batch and returns a loss).
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


# `bench.py --strong`: SURVEY 8(d)'s strong-scaling variant of C5 -- a fixed global batch sharded over the ranks
# (None = weak scaling, 256 samples per GPU)
C5_GLOBAL_BATCH = None

C3_CHANNELS, C3_IMAGE, C3_HIDDEN, C3_CLASSES = (3, 32, 64, 128), 32, 256, 10


def synthetic_batch(workload, rank=0, batch=None):
    """The synthetic host batch of a config (same generator for the GPU arm and the CPU arm): -> (x, target or None)"""
    import deeplearning.scala_b200.harness as H
    rng = np.random.RandomState(7 + rank)
    if workload == "c1":
        batch = batch or 128
        return rng.rand(784, batch).astype(np.float32), H.one_hot(rng.randint(0, 10, size=batch), 10, feature_major=True)
    if workload == "c4":
        return rng.standard_normal((4096, batch or 8192)).astype(np.float32), None
    batch = batch or 256
    x = rng.rand(batch, C3_IMAGE, C3_IMAGE, C3_CHANNELS[0]).astype(np.float32)
    return x, H.one_hot(rng.randint(0, C3_CLASSES, size=batch), C3_CLASSES, feature_major=False)


def build(workload, be, exchange=None, rank=0, world=1, gemm_mode=None, learning_rate=None):
    """-> (model, x_host, target_host or None, per-rank batch, flop per step per rank, description)"""
    import deeplearning.scala_b200.harness as H
    lr = (lambda default: default if learning_rate is None else learning_rate)     # parity tests use a larger step
    if workload == "c1":
        batch = 128
        model = H.Mlp((784, 500, 10), bias=True, last_activation="none", backend=be, learning_rate=lr(0.01 / batch), exchange=exchange)
        flop = 2.0 * batch * (784 * 500 + 500 * 10) * 3 - 2.0 * batch * 784 * 500          # no dX for layer 1
        desc = "C1: MLP 784-500-10 (tanh hidden, squared error), batch 128, feature-major, SGD lr 0.01 on the mean loss (= 0.01/128 on the reference's implicit sum loss)"
    elif workload == "c4":
        batch, width, depth = 8192, 4096, 4
        model = H.Mlp((width,) * (depth + 1), bias=False, last_activation="tanh", backend=be, learning_rate=lr(1e-4), exchange=exchange)
        flop = 2.0 * batch * width * width * (3 * depth - 1)
        desc = "C4: wide MLP 4 x (4096x4096) tanh, batch 8192, loss = sum(out), 11 GEMMs of 8192x4096x4096 per step"
    elif workload in ("c3", "c5"):
        batch = 256
        if workload == "c5" and C5_GLOBAL_BATCH is not None:
            if C5_GLOBAL_BATCH % world:
                raise ValueError("global batch %d is not divisible by %d ranks" % (C5_GLOBAL_BATCH, world))
            batch = C5_GLOBAL_BATCH // world
        model = H.Cnn(image=C3_IMAGE, channels=C3_CHANNELS, hidden=C3_HIDDEN, classes=C3_CLASSES, backend=be,
                      learning_rate=lr(1e-3 / world), exchange=exchange)
        flop = c3_flop(batch)
        desc = ("C3: CNN 3 conv3x3 (3-32-64-128, ReLU, 2x nearest down-sampling between them) + 2 FC (8192-256-10), "
                "batch %d NHWC 32x32x3 per GPU, squared error, SGD" % batch)
        if workload == "c5":
            desc = "C5: " + desc[4:] + "; data parallel, " + ("weak" if C5_GLOBAL_BATCH is None else "strong") + " scaling, SUM all-reduce of every weight gradient (NCCL + NVLink peer memory)"
    else:
        raise ValueError(workload)
    x, t = synthetic_batch(workload, rank, batch)
    return model, x, t, batch, flop, desc


def c3_flop(batch):
    conv = [(32 * 32, 3, 32), (16 * 16, 32, 64), (8 * 8, 64, 128)]
    fwd = sum(2.0 * batch * hw * 9 * ci * co for hw, ci, co in conv) + 2.0 * batch * (8192 * 256 + 256 * 10)
    return 3.0 * fwd - 2.0 * batch * 32 * 32 * 27 * 32          # forward + dX + dW, no dX for the first layer


def c3_algorithmic_bytes(batch):
    """HBM bytes one C3 step has to move with fp32 tensors when every kernel pass reads its inputs and writes its output
    exactly once (layer-by-layer schedule, nothing fused ACROSS layers; BASELINE.md section 4 spells the arithmetic out).
    It is the denominator-side figure of `roofline.achieved` for the training headline: bytes / step time."""
    f = 4
    x = batch * 32 * 32 * 3 * f
    y1, y2, y3 = batch * 32 * 32 * 32 * f, batch * 16 * 16 * 64 * f, batch * 8 * 8 * 128 * f
    d1, d2 = y1 // 4, y2 // 4                                   # the down-sampled inputs of conv2 / conv3
    h1, out = batch * 256 * f, batch * 10 * f
    w = (27 * 32 + 32 + 288 * 64 + 64 + 576 * 128 + 128 + 8192 * 256 + 256 + 2560 + 10) * f
    forward = (x + y1) + (d1 + y2) + (d2 + y3) + (y3 + h1) + (h1 + out) + w       # per layer: read input, write output
    backward = (3 * out                        # d_out = 2 (out - target): read out, target, write d_out
                + 2 * h1 + out                 # fc2: dW reads h1 and d_out; dX reads d_out, writes d_h1
                + 3 * h1                       # ReLU mask of h1: read d_h1, h1, write dz
                + 2 * y3 + 2 * h1              # fc1: dW reads y3, dz; dX reads dz, writes d_flat
                + 3 * y3                       # mask of y3
                + 2 * d2 + 2 * y3              # conv3: dW reads its input and dz3; dX reads dz3, writes the delta of its input
                + d2 + 2 * y2                  # up-sampling + mask of y2: read the delta, y2, write dz2
                + 2 * d1 + 2 * y2              # conv2
                + d1 + 2 * y1                  # up-sampling + mask of y1
                + x + y1                       # conv1: dW only
                + 2 * w)                       # weights read again, weight gradients written
    update = 3 * w                             # read weight + gradient, write weight
    return forward + backward + update


def cpu_model(workload):
    """Oracle-side parameters of a config, drawn from the same counter-based generators and seeds as harness.Cnn / Mlp
    (`TensorWeight.allocate(Tensor.randomNormal(shape, seed) * scale)`), for the CPU arm."""
    from oracle import tensors_oracle as O

    def normal(shape, seed, scale):
        n = int(np.prod(shape))
        return (O.random_normal_flat(n, seed | 0x40000000).reshape(shape) * np.float32(scale)).astype(np.float32)
    if workload in ("c3", "c5"):
        convs, seed = [], 51
        for ci, co in zip(C3_CHANNELS[:-1], C3_CHANNELS[1:]):
            convs.append((normal([3, 3, ci, co], seed, 0.05), normal([co], seed + 100, 0.01)))
            seed += 1
        fc1 = (normal([8192, C3_HIDDEN], seed, 0.01), normal([C3_HIDDEN], seed + 100, 0.01))
        fc2 = (normal([C3_HIDDEN, C3_CLASSES], seed + 1, 0.01), normal([C3_CLASSES], seed + 101, 0.01))
        return convs, fc1, fc2
    sizes, bias = ((784, 500, 10), True) if workload == "c1" else ((4096,) * 5, False)
    layers, seed = [], 11
    for fan_in, fan_out in zip(sizes[:-1], sizes[1:]):
        layers.append((normal([fan_out, fan_in], seed, 0.01), normal([fan_out], seed + 100, 0.01) if bias else None))
        seed += 1
    return layers


def run(workload, be, steps, warmup, e2e_steps=None, exchange=None, rank=0, world=1, barrier=None, compiled=True, eager_steps=None):
    """Times the step as a captured step (`model.compile`: the user code recorded once into a CUDA graph, one driver call
    per step) for exactly `steps` replays after `warmup` untimed ones, and a short eager leg (one host dispatch per launch,
    `model.step`) for comparison.  `compiled=False` makes the eager leg the headline."""
    import deeplearning.scala_b200.tensors as B
    from deeplearning.scala_b200 import _native as N
    model, x, t, batch, flop, desc = build(workload, be, exchange, rank, world)
    barrier = barrier or be.synchronize
    e2e_steps = steps if e2e_steps is None else e2e_steps
    eager_steps = (steps if not compiled else min(steps, 10)) if eager_steps is None else eager_steps
    xd = be.tensor(x).do_cache()
    td = be.tensor(t).do_cache() if t is not None else None
    losses = []
    for _ in range(min(warmup, 3) if compiled else warmup):
        losses.append(model.step(xd, td))
    barrier()
    st0 = be.stats()
    e0 = be.record_event()
    for _ in range(eager_steps):
        losses.append(model.step(xd, td))
    e1 = be.record_event()
    barrier()
    eager_ms = B.elapsed_ms(e0, e1)
    st1 = be.stats()
    # two host batches in page-locked memory, alternating like a data loader's double buffer
    x2, t2 = synthetic_batch(workload, rank + 1000, batch)
    pins = [pinned_like(N, x), pinned_like(N, x2)]
    tpins = [pinned_like(N, t), pinned_like(N, t2)] if t is not None else None
    out = {"workload": desc, "batch_per_gpu": batch, "flop_per_step_per_gpu": flop,
           "eager": {"ms_per_step": eager_ms / max(eager_steps, 1), "steps": eager_steps,
                     "launches_per_step": (st1["kernel_launches"] - st0["kernel_launches"]) / max(eager_steps, 1)},
           "driver_allocs_in_timed_region": st1["driver_allocs"] - st0["driver_allocs"],
           "h2d_bytes_per_step": int(x.nbytes + (t.nbytes if t is not None else 0)), "d2h_bytes_per_step": 4,
           "steps": steps, "e2e_steps": e2e_steps, "mode": "eager"}
    ms, launches = eager_ms * steps / max(eager_steps, 1), out["eager"]["launches_per_step"]
    if not compiled:
        # e2e of the eager path: the step a user makes -- host batch (pinned) -> Tensor.apply -> train -> loss on the host
        model.step(be.tensor(pins[0][1]), be.tensor(tpins[0][1]) if t is not None else None)
        barrier()
        t0 = time.perf_counter()
        for i in range(e2e_steps):
            losses.append(model.step(be.tensor(pins[i & 1][1]), be.tensor(tpins[i & 1][1]) if t is not None else None))
        be.synchronize()
        e2e_s = time.perf_counter() - t0
        out["e2e_mode"] = "eager: Tensor.apply upload, train, blocking loss read-back, serial"
    else:
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
        # e2e: every step uploads its pinned host batch (copy stream), replays, and hands a loss back to the host
        step.submit(pins[0][1], tpins[0][1] if t is not None else None)
        step.submit(pins[1][1], tpins[1][1] if t is not None else None)
        step.drain()
        barrier()
        t0 = time.perf_counter()
        for i in range(e2e_steps):
            loss = step.submit(pins[i & 1][1], tpins[i & 1][1] if t is not None else None)
            if loss is not None:
                losses.append(loss)
        losses.append(step.drain())
        be.synchronize()
        e2e_s = time.perf_counter() - t0
        info = step.graph.info()
        launches = (st1["kernel_launches"] - st0["kernel_launches"]) / steps
        out["captured"] = {"ms_per_step": ms / steps, "e2e_ms_per_step": e2e_s * 1e3 / e2e_steps, "launches_per_step": launches,
                           "graph_nodes": info["nodes"], "arena_bytes": info["arena_bytes"]}
        out["driver_allocs_in_timed_region"] = st1["driver_allocs"] - st0["driver_allocs"]
        out["mode"] = "captured step (CUDA graph replay of the recorded loss.train)"
        out["e2e_mode"] = ("pipelined CompiledStep.submit: batch k+1 uploaded on the copy stream while step k replays, device-to-"
                           "device into the step's inputs, loss of step k-1 read back asynchronously")
        del step
    for p, _ in pins + (tpins or []):
        N.lib.dls_host_free(p)
    n_params = int(sum(np.prod(w.data.shape) for w in model.weights))
    if exchange is not None:
        out["exchange"] = ("gradients >= %d bytes: grouped NCCL all-reduce on the exchange stream; smaller ones summed over NVLink peer "
                           "memory and every weight updated by one launch (dls_peer_sgd_step)" % exchange.peer_bytes
                           if exchange.peer else "grouped NCCL all-reduce on the exchange stream, one update launch per weight")
    out.update({"ms_per_step": ms / steps, "ms_total": ms, "e2e_ms_per_step": e2e_s * 1e3 / e2e_steps, "parameters": n_params,
                "launches_per_step": launches, "first_loss": losses[0], "last_loss": losses[-1]})
    return out
