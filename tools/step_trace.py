"""Device timeline of one replay of a captured training step (diagnosis only -- never a bench number: it runs under
CUPTI).  nsys is not in the image; torch.profiler (kineto over CUPTI) sees every kernel of the process, including the ones
libdlsb200 launches through ctypes and NCCL's, with start time, duration and stream.

    python tools/step_trace.py c3                       # one GPU
    python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29540 \
        tools/step_trace.py c5                          # every rank writes gpurun_out/trace_c5_rank<r>.txt

Output per rank: the kernels of the LAST traced replay in start order -- offset from the first kernel, duration, the
gap to the previous kernel on the same stream, stream id, name -- followed by totals per kernel name and the busy / idle
split of the compute stream.  The gaps are where an exchange tail, a join or launch latency shows up.
"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench_train  # noqa: E402
import deeplearning.scala_b200.tensors as B  # noqa: E402


def main():
    import torch
    from torch.profiler import ProfilerActivity, profile
    workload = sys.argv[1] if len(sys.argv) > 1 else "c3"
    replays = int(sys.argv[2]) if len(sys.argv) > 2 else 5
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    dist = None
    exchange = None
    torch.cuda.set_device(local)
    be = B.Backend(device=local, gemm_mode=os.environ.get("DLS_GEMM_MODE", "bf16"))
    if world > 1:
        import torch.distributed as dist
        from deeplearning.scala_b200.harness import Exchange
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
        uid = torch.zeros(128, dtype=torch.uint8, device="cuda")
        if rank == 0:
            uid = torch.tensor(list(Exchange.unique_id()), dtype=torch.uint8, device="cuda")
        dist.broadcast(uid, 0)
        exchange = Exchange(be, rank, world, bytes(uid.cpu().tolist()))
        exchange.verify()
        workload = "c5" if workload == "c3" else workload
    model, x, t, batch, flop, desc = bench_train.build(workload, be, exchange, rank, world)
    xd = be.tensor(x).do_cache()
    td = be.tensor(t).do_cache() if t is not None else None
    step = model.compile(xd, td)
    for _ in range(20):
        step.launch()
    be.synchronize()
    if dist is not None:
        dist.barrier()
    with profile(activities=[ProfilerActivity.CUDA]) as prof:
        for _ in range(replays):
            step.launch()
        be.synchronize()
        torch.cuda.synchronize()
    events = [e for e in prof.events() if e.device_type == torch.autograd.DeviceType.CUDA]
    events.sort(key=lambda e: e.time_range.start)
    if not events:
        print("no device activity recorded (CUPTI unavailable?)")
        return
    # the last replay = the last len(events) / replays kernels (every replay launches the same sequence)
    per = len(events) // replays
    last = events[-per:]
    t0 = last[0].time_range.start
    out_dir = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "gpurun_out")
    os.makedirs(out_dir, exist_ok=True)
    path = os.path.join(out_dir, "trace_%s_rank%d.txt" % (workload, rank))
    with open(path, "w") as f:
        f.write("# %s, world %d, rank %d, %d kernels per replay (diagnosis under CUPTI, not a bench number)\n" % (desc, world, rank, per))
        f.write("# start_us  dur_us  gap_us  stream  name\n")
        prev_end = {}
        busy = {}
        totals = {}
        for e in last:
            s, d = e.time_range.start - t0, e.time_range.end - e.time_range.start
            stream = getattr(e, "device_index", 0), getattr(e, "device_resource_id", 0)
            gap = s - prev_end[stream] if stream in prev_end else 0.0
            prev_end[stream] = s + d
            busy[stream] = busy.get(stream, 0.0) + d
            name = e.name[:110]
            totals.setdefault(name, [0, 0.0])
            totals[name][0] += 1
            totals[name][1] += d
            f.write("%9.1f %7.1f %7.1f  %s  %s\n" % (s, d, gap, stream[1], name))
        span = max(prev_end.values())
        f.write("\n# replay span %.1f us\n" % span)
        for stream, b in sorted(busy.items()):
            f.write("# stream %s: busy %.1f us (%.0f %% of the span)\n" % (stream[1], b, 100.0 * b / span))
        f.write("\n# per kernel name: launches, total us\n")
        for name, (count, total) in sorted(totals.items(), key=lambda kv: -kv[1][1]):
            f.write("%4d %9.1f  %s\n" % (count, total, name))
    if rank == 0:
        print(open(path).read()[-3000:])
    del step
    be.close()
    if dist is not None:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
