#!/usr/bin/env python
"""bench.py -- BASELINE.json config C2: `tanh(a*b+c)` with a broadcast operand, materialised, then `.sum`, over 2^28 fp32.

  python bench.py --gpus N --steps K --warmup W            (N > 1: launched by torch.distributed.run, one rank per GPU)
  python bench.py --impl reference ...                      CPU arm: the oracle's C restatement of the reference path

One step = one pass of the hot path over one batch: y = tanh(a*b+c).doCache (ew_vec_kernel, 12 B/element) followed by
y.sum (sum_kernel, 4 B/element) -- the 16 B/element the reference's path moves (it materialises, then reduces,
Tensors.scala:673-766).  `value` is algorithmic GB/s with inputs resident in HBM; `e2e` is the same metric through the
public Tensor API with HOST inputs (pinned), H2D of a, b, c and D2H of the sum inside the timed region.
Each rank owns an independent 2^28-element problem (weak scaling, no data-path collective).
"""
import argparse
import ctypes
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

ROWS, COLS = 16384, 16384                    # 2^28 fp32 per full operand (1 GiB)
BYTES_MAP, BYTES_SUM = 12, 4                 # algorithmic bytes per element
METRIC = "fused elementwise + reduce HBM throughput (tanh(a*b+c) materialised, then sum)"


def config_dict(rows):
    n = rows * COLS
    return {"workload": "C2: y = tanh(a*b+c) materialised (c broadcast per row), then y.sum; a,b = %dx%d fp32 per GPU" % (rows, COLS),
            "elements_per_gpu": n, "algorithmic_bytes_per_element": BYTES_MAP + BYTES_SUM,
            "l2": "inputs (2 x %.2f GiB) and output exceed the 126 MB L2, no flush needed" % (n * 4 / 2 ** 30),
            "rng": "reference Wang-hash uniform mapped to [-2,2), seeds 1,2,3", "parallelism": "independent shard per GPU"}


def load_traffic():
    """dram__bytes_read.sum + dram__bytes_write.sum per launch of the dominant kernel, from the committed
    `ncu --set full` capture (profiles/summarise.py); None when no capture exists for this build."""
    path = os.path.join(ROOT, "profiles", "r1_c2_traffic.json")
    try:
        with open(path) as f:
            return float(json.load(f)["ew_vec_kernel<1, 4>"])       # rows x inner launch, 4 tiles per lane: the C2a map
    except (OSError, KeyError, ValueError):
        return None


def load_peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        with open(path) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


class ClockSampler:
    """SM clock and throttle reasons sampled DURING the timed region (B200_PROFILING.md recipe), through NVML every
    ~5 ms so that even a 30 ms region gets samples; falls back to `nvidia-smi -lms` when NVML is unusable."""
    NAMES = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]

    def __init__(self, gpu_index):
        self.rows, self.stop_flag, self.smi = [], False, None
        self.max_mhz = None
        try:
            import pynvml
            pynvml.nvmlInit()
            phys = gpu_index
            vis = os.environ.get("CUDA_VISIBLE_DEVICES")
            if vis:
                try:
                    phys = int(vis.split(",")[gpu_index])
                except ValueError:
                    phys = gpu_index
            self.nv, self.h = pynvml, pynvml.nvmlDeviceGetHandleByIndex(phys)
            self.max_mhz = float(pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM))
            self.thread = threading.Thread(target=self._pump_nvml, daemon=True)
        except Exception:
            self.nv = None
            q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
                 "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
            try:
                self.smi = subprocess.Popen(["nvidia-smi", "-i", str(gpu_index), "--query-gpu=" + q,
                                             "--format=csv,noheader,nounits", "-lms", "20"],
                                            stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            except OSError:
                self.smi = None
            self.thread = threading.Thread(target=self._pump_smi, daemon=True)
        self.thread.start()

    def _pump_nvml(self):
        nv = self.nv
        bits = [(nv.nvmlClocksEventReasonHwSlowdown, "hw_slowdown"), (nv.nvmlClocksEventReasonHwThermalSlowdown, "hw_thermal_slowdown"),
                (nv.nvmlClocksEventReasonSwThermalSlowdown, "sw_thermal_slowdown"), (nv.nvmlClocksEventReasonSwPowerCap, "sw_power_cap")]
        while not self.stop_flag:
            try:
                mhz = float(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                mask = int(nv.nvmlDeviceGetCurrentClocksEventReasons(self.h))
                self.rows.append((time.time(), mhz, self.max_mhz, [n for b, n in bits if mask & b]))
            except Exception:
                pass
            time.sleep(0.004)

    def _pump_smi(self):
        if not self.smi:
            return
        for line in self.smi.stdout:
            parts = [p.strip() for p in line.split(",")]
            try:
                reasons = [n for n, v in zip(self.NAMES, parts[2:6]) if v.lower().startswith("active")]
                self.rows.append((time.time(), float(parts[0]), float(parts[1]), reasons))
            except (ValueError, IndexError):
                continue

    def window(self, t0, t1):
        sm, mx, reasons = [], 0.0, set()
        for t, mhz, mxz, rs in list(self.rows):
            if t < t0 or t > t1:
                continue
            sm.append(mhz)
            mx = max(mx, mxz or 0.0)
            reasons.update(rs)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": mx or self.max_mhz, "reasons": sorted(reasons),
                "samples": len(sm), "source": "nvml" if self.nv else "nvidia-smi"}

    def stop(self):
        self.stop_flag = True
        if self.smi:
            self.smi.terminate()


def c_oracle():
    lib = ctypes.CDLL(os.path.join(ROOT, "oracle", "_build", "libc2oracle.so"))
    fp = ctypes.POINTER(ctypes.c_float)
    lib.c2_map_rows.argtypes = [fp, fp, fp, fp, ctypes.c_size_t, ctypes.c_size_t, ctypes.c_int]
    lib.c2_sum16.argtypes = [fp, ctypes.c_size_t]
    lib.c2_sum16.restype = ctypes.c_float
    lib.c2_threads.restype = ctypes.c_int
    return lib


def cpu_step_runner(rows):
    """The reference's CPU path for C2 on a bounded sample of `rows` rows: generated map kernel on all cores, then the
    single-work-item 16-lane sum (what a CPU OpenCL device executes, Tensors.scala:683-696)."""
    lib = c_oracle()
    fp = ctypes.POINTER(ctypes.c_float)
    rng = np.random.RandomState(1)
    a = rng.standard_normal((rows, COLS)).astype(np.float32)
    b = rng.standard_normal((rows, COLS)).astype(np.float32)
    c = rng.standard_normal(rows).astype(np.float32)
    y = np.empty_like(a)
    threads = lib.c2_threads()

    def step():
        lib.c2_map_rows(a.ctypes.data_as(fp), b.ctypes.data_as(fp), c.ctypes.data_as(fp), y.ctypes.data_as(fp), rows, COLS, threads)
        return lib.c2_sum16(y.ctypes.data_as(fp), rows * COLS)
    return step, threads


def run_reference(args, rank):
    if rank != 0:
        return
    rows = 1024                                              # bounded sample: 2^24 of the 2^28 elements per step
    step, threads = cpu_step_runner(rows)
    for _ in range(args.warmup):
        step()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        step()
    dt = time.perf_counter() - t0
    elems = rows * COLS * args.steps
    gbs = elems * (BYTES_MAP + BYTES_SUM) / dt / 1e9
    sample = "%d of %d rows (2^24 elements) per step; map on %d threads + single-thread 16-lane sum" % (rows, ROWS, threads)
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": gbs, "unit": "GB/s", "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": dt / args.steps * 1e3, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": dict(config_dict(ROWS), cpu_sample=sample),
        "cpu_baseline": {"value": gbs, "unit": "GB/s", "cores": threads, "kind": "port", "sample": sample},
        "e2e": {"value": gbs, "unit": "GB/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "note": "reference's own CPU path (Compute.scala on a CPU OpenCL device) is unavailable on this box (no JVM, "
                "no OpenCL ICD); this arm times oracle/c2_oracle.c, the C restatement of that path"}))


def main():
    # NCCL prints its version banner to stdout at NCCL_DEBUG=VERSION and above; stdout carries exactly one JSON line
    if os.environ.get("NCCL_DEBUG", "").upper() in ("VERSION", "INFO", "TRACE"):
        os.environ["NCCL_DEBUG"] = "WARN"
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=100)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--e2e-steps", type=int, default=5)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--rows", type=int, default=ROWS, help="debug: smaller problem (INVALID as a bench number)")
    ap.add_argument("--workload", default="c2", choices=["c2", "c1", "c3", "c4", "c5"],
                    help="c2 (default) = BASELINE configs[1]; c1/c3/c4/c5 = the training configs (samples/s)")
    ap.add_argument("--gemm-mode", default="bf16", choices=["fp32", "tf32", "bf16"], help="contraction precision of the training legs")
    ap.add_argument("--no-train", action="store_true", help="skip the short training legs appended to the C2 line")
    ap.add_argument("--strong", action="store_true", help="c5 only: strong scaling, global batch 2048 sharded over the ranks")
    ap.add_argument("--eager", action="store_true", help="training legs: time only the eager path (no captured step)")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if args.workload != "c2":
        main_train(args, rank, world, local)
        return
    if args.impl == "reference":
        run_reference(args, rank)
        return
    if args.warmup < 3:
        args.warmup = 3

    import deeplearning.scala_b200.tensors as B
    from deeplearning.scala_b200 import _native as N

    dist = None
    if world > 1:
        import torch
        import torch.distributed as dist
        torch.cuda.set_device(local)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    be = B.Backend(device=local)
    rows = args.rows
    n = rows * COLS

    # ---- synthetic inputs generated on the device by the reference's own Wang-hash Box-Muller RNG -------------------
    # uniform, mapped to [-2, 2): the reference's Box-Muller normal does not guard u1 == 0 (Tensors.scala:409-425), so
    # 2^28 draws contain inf/NaN and would turn the checksum into NaN
    a = (be.random([rows, COLS], seed=1 + 16 * rank) * 4.0 - 2.0).do_cache()
    b = (be.random([rows, COLS], seed=2 + 16 * rank) * 4.0 - 2.0).do_cache()
    c = (be.random([rows], seed=3 + 16 * rank) * 4.0 - 2.0).do_cache()

    def step():
        y = be.tanh(a * b + c).do_cache()                    # ew_vec_kernel: read a, b (+c), write y
        s = y.sum()._buffer()                                # sum_kernel: read y
        return y, s

    def barrier():
        be.synchronize()
        if dist is not None:
            dist.barrier()
        be.synchronize()

    for _ in range(args.warmup):
        step()
    barrier()
    sampler = ClockSampler(local)
    time.sleep(0.25)
    st0 = be.stats()
    wall0 = time.time()
    e0 = be.record_event()
    for _ in range(args.steps):
        keep = step()
    e1 = be.record_event()
    barrier()
    wall1 = time.time()
    ms_total = B.elapsed_ms(e0, e1)
    launches = be.stats()["kernel_launches"] - st0["kernel_launches"]
    clocks = sampler.window(wall0, wall1)
    check_sum = float(be.download(keep[1], 1)[0])
    del keep

    # ---- dominant kernel alone (roofline): K launches of the elementwise kernel between two events -----------------
    barrier()
    k0 = be.record_event()
    for _ in range(args.steps):
        y = be.tanh(a * b + c).do_cache()
    k1 = be.record_event()
    be.synchronize()
    ms_map = B.elapsed_ms(k0, k1) / args.steps
    k0 = be.record_event()
    for _ in range(args.steps):
        s = y.sum()._buffer()
    k1 = be.record_event()
    be.synchronize()
    ms_sum = B.elapsed_ms(k0, k1) / args.steps
    k0 = be.record_event()
    for _ in range(args.steps):
        s = be.tanh(a * b + c).sum()._buffer()               # fused variant (C2c): nothing materialised, 8 B/element
    k1 = be.record_event()
    be.synchronize()
    ms_fused = B.elapsed_ms(k0, k1) / args.steps
    del y, s
    # SURVEY 8(d) also asks for the per-COLUMN broadcast of c (`c.reshape([1, COLS])`: one value per column, repeated down
    # the rows).  A side measurement: it must never cost the headline line, so any failure is recorded instead of raised.
    variants = {}
    try:
        c_col = (be.random([COLS], seed=4 + 16 * rank) * 4.0 - 2.0).do_cache().reshape([1, COLS])
        y = be.tanh(a * b + c_col).do_cache()
        be.synchronize()
        k0 = be.record_event()
        for _ in range(args.steps):
            y = be.tanh(a * b + c_col).do_cache()
        k1 = be.record_event()
        be.synchronize()
        ms_col = B.elapsed_ms(k0, k1) / args.steps
        variants["c2a_column_broadcast"] = {"ms": ms_col, "gbs": n * BYTES_MAP / (ms_col / 1e3) / 1e9}
        del y, c_col
    except Exception as exc:                                  # noqa: BLE001
        variants["c2a_column_broadcast"] = {"error": repr(exc)}

    # ---- e2e: host inputs in pinned memory, H2D + compute + D2H of the sum inside the timed region -------------------
    def pinned(count):
        p = ctypes.POINTER(ctypes.c_float)()
        N.check(N.lib.dls_host_alloc(count, ctypes.byref(p)))
        return p, np.ctypeslib.as_array(p, shape=(count,))
    pa, ha = pinned(n)
    pb, hb = pinned(n)
    pc, hc = pinned(rows)
    pr, hr = pinned(1)
    ha[:] = be.download(a._buffer(), n)
    hb[:] = be.download(b._buffer(), n)
    hc[:] = be.download(c._buffer(), rows)
    del a, b, c

    def e2e_step():
        ta = be.tensor(ha.reshape(rows, COLS))               # Tensor.apply: host data -> device (dls_buf_from_host)
        tb = be.tensor(hb.reshape(rows, COLS))
        tc = be.tensor(hc)
        y = be.tanh(ta * tb + tc).do_cache()
        sbuf = y.sum()._buffer()
        N.check(N.lib.dls_buf_read(be.ctx, sbuf.handle, 0, pr, 1, None, 0, None))     # blocking D2H of the result
        return float(hr[0])
    e2e_step()
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.e2e_steps):
        e2e_sum = e2e_step()
    be.synchronize()
    e2e_s = time.perf_counter() - t0
    for p in (pa, pb, pc, pr):
        N.lib.dls_host_free(p)

    # ---- max over ranks -------------------------------------------------------------------------------------------------
    times = [ms_total, e2e_s * 1e3]
    if dist is not None:
        import torch
        t = torch.tensor(times, device="cuda", dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        times = t.tolist()
    ms_total, e2e_ms = times
    bytes_step = n * (BYTES_MAP + BYTES_SUM)
    value = world * bytes_step * args.steps / (ms_total / 1e3) / 1e9
    e2e_value = world * bytes_step * args.e2e_steps / (e2e_ms / 1e3) / 1e9

    peak, peak_src = load_peaks()
    achieved = n * BYTES_MAP / (ms_map / 1e3) / 1e9
    out = {
        "metric": METRIC, "value": value, "unit": "GB/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": ms_total / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f32", "data": "synthetic",
        "config": config_dict(rows),
        "clocks": clocks,
        "e2e": {"value": e2e_value, "unit": "GB/s", "h2d_bytes_per_step": (2 * n + rows) * 4, "d2h_bytes_per_step": 4,
                "steps": args.e2e_steps, "ms_per_step": e2e_ms / args.e2e_steps},
        "gpu_launches": launches,
        "roofline": {"kernel": "ew_vec_kernel", "bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s",
                     "frac": achieved / peak, "traffic": load_traffic() if rows == ROWS else None, "peak_source": peak_src, "ms_per_launch": ms_map,
                     "algorithmic_bytes_per_launch": n * BYTES_MAP},
        "kernels": {"sum_kernel": {"ms": ms_sum, "gbs": n * BYTES_SUM / (ms_sum / 1e3) / 1e9, "frac": n * BYTES_SUM / (ms_sum / 1e3) / 1e9 / peak},
                    "ew_vec_kernel+reduce (fused tanh(a*b+c).sum)": {"ms": ms_fused, "gbs": n * 8 / (ms_fused / 1e3) / 1e9,
                                                                      "frac": n * 8 / (ms_fused / 1e3) / 1e9 / peak}},
        "checksum": {"sum": check_sum, "e2e_sum": e2e_sum},
    }
    for name, v in variants.items():
        if "gbs" in v:
            v["frac"] = v["gbs"] / peak
        out["kernels"][name] = v
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        sample_rows = 1024
        cpu_step, threads = cpu_step_runner(sample_rows)
        cpu_step()
        t0, reps = time.perf_counter(), 0
        while time.perf_counter() - t0 < 10.0 and reps < 200:
            cpu_step()
            reps += 1
        dt = time.perf_counter() - t0
        out["cpu_baseline"] = {"value": sample_rows * COLS * reps * (BYTES_MAP + BYTES_SUM) / dt / 1e9, "unit": "GB/s",
                               "cores": threads, "kind": "port",
                               "sample": "%d reps over %d of %d rows (2^24 elements): C restatement of the generated map "
                                         "kernel on %d threads + the single-work-item 16-lane sum" % (reps, sample_rows, ROWS, threads)}
    sampler.stop()
    be.close()
    if not args.no_train and rows == ROWS:
        out["train"] = train_legs(args, rank, world, local, dist)
    if rank == 0:
        print(json.dumps(out))
    if dist is not None:
        dist.destroy_process_group()


# ---- training legs (BASELINE metric "train samples/sec at 1/2/4/8 B200") ----------------------------------------------------

def make_exchange(be, rank, world, dist):
    """NCCL communicator of libdlsb200 itself (dls_comm_init); the unique id travels over torch.distributed."""
    if world == 1:
        return None
    import torch
    from deeplearning.scala_b200.harness import Exchange
    uid = torch.zeros(128, dtype=torch.uint8, device="cuda")
    if rank == 0:
        uid = torch.tensor(list(Exchange.unique_id()), dtype=torch.uint8, device="cuda")
    dist.broadcast(uid, 0)
    exchange = Exchange(be, rank, world, bytes(uid.cpu().tolist()))
    exchange.verify()                                          # known-answer sum over every rank's peer segment (untimed)
    return exchange


def train_once(workload, args, rank, world, local, dist, steps, warmup, e2e_steps):
    import deeplearning.scala_b200.tensors as B
    import bench_train
    be = B.Backend(device=local, gemm_mode=args.gemm_mode)
    exchange = make_exchange(be, rank, world, dist) if workload == "c5" else None

    def barrier():
        be.synchronize()
        if dist is not None:
            dist.barrier()
        be.synchronize()
    r = bench_train.run(workload, be, steps, warmup, e2e_steps, exchange, rank, world, barrier, compiled=not args.eager)
    times = [r["ms_total"], r["e2e_ms_per_step"]]
    if dist is not None:
        import torch
        t = torch.tensor(times, device="cuda", dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        times = t.tolist()
    r["ms_total"], r["e2e_ms_per_step"] = times
    r["ms_per_step"] = times[0] / steps
    r["samples_per_s"] = world * r["batch_per_gpu"] * steps / (times[0] / 1e3)
    r["e2e_samples_per_s"] = world * r["batch_per_gpu"] / (times[1] / 1e3)
    r["tflops"] = world * r["flop_per_step_per_gpu"] / (r["ms_per_step"] / 1e3) / 1e12
    r["gemm_mode"] = args.gemm_mode
    del exchange
    be.close()
    return r


def train_legs(args, rank, world, local, dist):
    """Short runs of the training configs appended to the C2 line: C1, C3 (as C5 when world > 1) and C4."""
    legs = {}
    for workload, steps in (("c1", 50), ("c5" if world > 1 else "c3", 10), ("c4", 5)):
        try:
            r = train_once(workload, args, rank, world, local, dist, steps, 3, 3)
            legs[workload] = {k: r[k] for k in ("workload", "samples_per_s", "e2e_samples_per_s", "ms_per_step", "tflops",
                                                "launches_per_step", "gemm_mode", "last_loss", "mode", "eager")}
        except Exception as exc:                      # a training leg must never take the headline line down
            legs[workload] = {"error": "%s: %s" % (type(exc).__name__, exc)}
    return legs


def train_reference(args, rank):
    """CPU arm of a training workload: the numpy restatement of the reference's autodiff rules (oracle/models_oracle.py)
    on the host cores (BLAS threads), one step per timed step on the full batch of the config."""
    if rank != 0:
        return
    from oracle import models_oracle as MO
    import deeplearning.scala_b200.harness as H
    rng = np.random.RandomState(7)
    w = args.workload
    if w == "c1":
        batch, sizes = 128, (784, 500, 10)
        weights = [(rng.randn(o, i).astype(np.float32) * 0.01, rng.randn(o).astype(np.float32) * 0.01) for i, o in zip(sizes[:-1], sizes[1:])]
        x = rng.rand(784, batch).astype(np.float32)
        t = H.one_hot(rng.randint(0, 10, size=batch), 10, True)
        step = lambda: MO.mlp_step(weights, x, t, 0.01)
        sample = "full C1 step (batch 128)"
    elif w == "c4":
        batch, width = 512, 4096
        weights = [(rng.randn(width, width).astype(np.float32) * 0.01, None) for _ in range(4)]
        x = rng.randn(width, batch).astype(np.float32)
        step = lambda: MO.mlp_step(weights, x, None, 1e-4, "tanh")
        sample = "C4 model on a 512-sample slice of the 8192 batch (fp64 BLAS accumulate)"
    else:
        batch = 32
        chans = (3, 32, 64, 128)
        convs = [(rng.randn(3, 3, ci, co).astype(np.float32) * 0.05, rng.randn(co).astype(np.float32) * 0.01) for ci, co in zip(chans[:-1], chans[1:])]
        fc1 = (rng.randn(8192, 256).astype(np.float32) * 0.01, np.zeros(256, np.float32))
        fc2 = (rng.randn(256, 10).astype(np.float32) * 0.01, np.zeros(10, np.float32))
        x = rng.rand(batch, 32, 32, 3).astype(np.float32)
        t = H.one_hot(rng.randint(0, 10, size=batch), 10, False)
        step = lambda: MO.cnn_step(convs, fc1, fc2, x, t, 1e-3)
        sample = "C3 model on a 32-sample slice of the 256 batch"
    for _ in range(max(1, min(args.warmup, 2))):
        step()
    steps = max(1, min(args.steps, 20))
    t0 = time.perf_counter()
    for _ in range(steps):
        step()
    dt = time.perf_counter() - t0
    value = batch * steps / dt
    cores = os.cpu_count() or 1
    print(json.dumps({"impl": "reference", "metric": "train samples/sec", "value": value, "unit": "samples/s", "n_gpus": args.gpus,
                      "steps": steps, "warmup": args.warmup, "ms_per_step": dt * 1e3 / steps, "higher_is_better": True,
                      "scaling": "weak", "vs_baseline": None, "dtype": "f32 (fp64 accumulate)", "data": "synthetic",
                      "config": {"workload": args.workload},
                      "cpu_baseline": {"value": value, "unit": "samples/s", "cores": cores, "kind": "port", "sample": sample},
                      "e2e": {"value": value, "unit": "samples/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}))


def main_train(args, rank, world, local):
    if args.impl == "reference":
        train_reference(args, rank)
        return
    if args.warmup < 3:
        args.warmup = 3
    dist = None
    if world > 1:
        import torch
        import torch.distributed as dist
        torch.cuda.set_device(local)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    workload = args.workload
    if workload == "c3" and world > 1:
        workload = "c5"
    if args.strong and workload == "c5":
        import bench_train
        bench_train.C5_GLOBAL_BATCH = 2048
    sampler = ClockSampler(local)
    time.sleep(0.25)
    wall0 = time.time()
    r = train_once(workload, args, rank, world, local, dist, args.steps, args.warmup, args.e2e_steps)
    wall1 = time.time()
    clocks = sampler.window(wall0, wall1)
    sampler.stop()
    peaks = {}
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            peaks = json.load(f)
    except OSError:
        pass
    tensor_peak = float(peaks.get("bf16_tflops_sustained", 1400.0))
    out = {"metric": "train samples/sec", "value": r["samples_per_s"], "unit": "samples/s", "n_gpus": world, "steps": args.steps,
           "warmup": args.warmup, "ms_per_step": r["ms_per_step"], "higher_is_better": True,
           "scaling": "strong" if args.strong and workload == "c5" else "weak", "vs_baseline": None,
           "dtype": {"fp32": "f32", "tf32": "tf32 (fp32 accumulate)", "bf16": "bf16 (fp32 accumulate)"}[args.gemm_mode],
           "data": "synthetic",
           "config": {"workload": r["workload"], "batch_per_gpu": r["batch_per_gpu"], "parameters": r["parameters"],
                      "parallelism": "dp%d" % world if workload == "c5" else "single GPU replica",
                      "l2": "working set per step exceeds L2 for c3/c4; c1 (3.6 MB) is L2 resident by nature"},
           "clocks": clocks,
           "e2e": {"value": r["e2e_samples_per_s"], "unit": "samples/s", "h2d_bytes_per_step": r["h2d_bytes_per_step"],
                   "d2h_bytes_per_step": r["d2h_bytes_per_step"], "ms_per_step": r["e2e_ms_per_step"], "steps": r["e2e_steps"]},
           "gpu_launches": int(r["launches_per_step"] * args.steps),
           "roofline": {"kernel": "gemm_tc_kernel (all contractions of the step)", "bound": "tensor", "achieved": r["tflops"] / world,
                        "peak": tensor_peak, "unit": "TFLOP/s", "frac": r["tflops"] / world / tensor_peak, "traffic": None,
                        "peak_source": "MEASURED_PEAKS.json bf16_tflops_sustained" if peaks else "fallback 1.4 PFLOP/s sustained",
                        "note": "whole-step FLOP / whole-step time (includes elementwise, staging and host launch gaps)"},
           "launches_per_step": r["launches_per_step"], "driver_allocs_in_timed_region": r["driver_allocs_in_timed_region"],
           "mode": r["mode"], "eager": r["eager"], "captured": r.get("captured"),
           "loss": {"first": r["first_loss"], "last": r["last_loss"]}}
    if "exchange" in r:
        out["config"]["exchange"] = r["exchange"]
    if rank == 0:
        print(json.dumps(out))
    if dist is not None:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
