"""Latency of the exchange primitives on C5's gradient set, each timed as a graph replay (device time, max over ranks):

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29541 \
        tools/exchange_probe.py

  peer_tail      dls_peer_sgd_step on C5's ten weights: nine small gradients summed over peer memory, fc1 updated locally
  peer_sum_only  the same nine gradients, all-reduce only (weights = NULL)
  nccl_small     one grouped NCCL all-reduce of the nine small gradients (the tail it replaces)
  nccl_fc1       NCCL all-reduce of the 8.4 MB fc1 gradient alone (what still overlaps the backward pass)
  sgd_local      the ten updates as one multi-tensor launch without any exchange (lower bound of the tail)

Every rank must run the same sequence; rank 0 prints one JSON line per primitive.
"""
import ctypes as C
import json
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import deeplearning.scala_b200.tensors as B  # noqa: E402
from deeplearning.scala_b200 import _native as N  # noqa: E402

# C5 / C3: conv weights [3,3,Cin,Cout] + biases, fc1 8192x256, fc2 256x10, in the order their gradients arrive
SIZES = [2560, 10, 8192 * 256, 256, 3 * 3 * 64 * 128, 128, 3 * 3 * 32 * 64, 64, 3 * 3 * 3 * 32, 32]
FC1 = 2


def main():
    import torch
    import torch.distributed as dist
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    reps = int(sys.argv[1]) if len(sys.argv) > 1 else 50
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    from deeplearning.scala_b200.harness import Exchange
    be = B.Backend(device=local)
    uid = torch.zeros(128, dtype=torch.uint8, device="cuda")
    if rank == 0:
        uid = torch.tensor(list(Exchange.unique_id()), dtype=torch.uint8, device="cuda")
    dist.broadcast(uid, 0)
    ex = Exchange(be, rank, world, bytes(uid.cpu().tolist()))
    ex.verify()
    rng = np.random.RandomState(rank)
    weights = [be.upload(np.zeros(n, np.float32)) for n in SIZES]
    grads = [be.upload((rng.standard_normal(n) * 1e-3).astype(np.float32)) for n in SIZES]
    small = [i for i in range(len(SIZES)) if i != FC1]

    def arr(bufs):
        return (C.c_void_p * len(bufs))(*[b.handle for b in bufs])

    def counts(idx):
        return (C.c_size_t * len(idx))(*[SIZES[i] for i in idx])

    def peer_tail():
        flags = (C.c_int * len(SIZES))(*[0 if i == FC1 else 1 for i in range(len(SIZES))])
        N.check(N.lib.dls_peer_sgd_step(be.ctx, arr(weights), arr(grads), counts(range(len(SIZES))), flags, len(SIZES), 1e-3))

    def peer_sum_only():
        flags = (C.c_int * len(small))(*([1] * len(small)))
        N.check(N.lib.dls_peer_sgd_step(be.ctx, None, arr([grads[i] for i in small]), counts(small), flags, len(small), 0.0))

    def nccl_small():
        N.check(N.lib.dls_allreduce_sum_many(be.ctx, arr([grads[i] for i in small]), counts(small), len(small)))

    def nccl_fc1():
        N.check(N.lib.dls_allreduce_sum_many(be.ctx, arr([grads[FC1]]), counts([FC1]), 1))

    def sgd_local():
        N.check(N.lib.dls_peer_sgd_step(be.ctx, arr(weights), arr(grads), counts(range(len(SIZES))), None, len(SIZES), 1e-3))

    for name, fn in (("peer_tail", peer_tail), ("peer_sum_only", peer_sum_only), ("nccl_small", nccl_small),
                     ("nccl_fc1", nccl_fc1), ("sgd_local", sgd_local)):
        if not ex.peer and name.startswith("peer"):
            continue
        fn()                                                   # eager once (NCCL connects lazily)
        be.synchronize()
        with be.capture() as graph:
            fn()
        for _ in range(5):
            graph.launch()
        be.synchronize()
        dist.barrier()
        e0 = be.record_event()
        for _ in range(reps):
            graph.launch()
        e1 = be.record_event()
        be.synchronize()
        t = torch.tensor([B.elapsed_ms(e0, e1) * 1e3 / reps], device="cuda", dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        if rank == 0:
            print(json.dumps({"primitive": name, "world": world, "us_per_call": t.item(), "reps": reps,
                              "bytes_small": 4 * sum(SIZES[i] for i in small), "bytes_fc1": 4 * SIZES[FC1]}))
        del graph
    del ex
    be.close()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
