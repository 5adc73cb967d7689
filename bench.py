#!/usr/bin/env python
"""bench.py -- BASELINE.json metric "train samples/sec at 1/2/4/8 B200; fused-elementwise/reduce HBM GB/s vs peak".

  python bench.py --gpus N --steps K --warmup W            (N > 1: launched by torch.distributed.run, one rank per GPU)
  python bench.py --impl reference ...                      CPU arm: the oracle's restatement of the same step on the host cores

Headline (`metric`, `value`): one training step = forward + backward + SGD update of every weight (`loss.train`,
TL:70-76; TW:50-74) of the CIFAR-shaped CNN -- config C3 at N = 1, config C5 (C3 data parallel, 256 samples per GPU,
SUM all-reduce of every weight gradient over NCCL + NVLink peer memory) at N > 1 -- replayed as a captured step,
exactly K timed replays after W untimed ones, CUDA events on the launch stream, max over ranks.
`e2e` = the same steps fed from HOST batches (page-locked), upload + loss read-back inside the timed region.
`roofline` = the step's algorithmic HBM bytes (bench_train.c3_algorithmic_bytes) / step time
against the measured HBM peak.  `kernels` (N = 1) carries the metric's second half: config C2 -- tanh(a*b+c) with a
broadcast operand over 2^28 fp32, materialised (12 B/element), summed (4 B/element), fused (8 B/element) -- and the
tcgen05 GEMM rates.  `--workload c1|c2|c3|c4|c5` selects another config as the headline.
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

ROWS, COLS = 16384, 16384                    # C2: 2^28 fp32 per full operand (1 GiB)
BYTES_MAP, BYTES_SUM = 12, 4                 # algorithmic bytes per element
METRIC = "train samples/sec"
C2_METRIC = "fused elementwise + reduce HBM throughput (tanh(a*b+c) materialised, then sum)"
TOLERANCE = {"fp32": 1e-5, "tf32": 1e-3, "bf16": 4e-3}       # stated relative bound of the contractions per mode
DTYPE = {"fp32": "f32", "tf32": "tf32 (fp32 accumulate)", "bf16": "bf16 (fp32 accumulate)"}


def load_peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        with open(path) as f:
            p = json.load(f)
        return {"hbm": float(p["hbm_gbs"]), "bf16": float(p.get("bf16_tflops", 1653.0)),
                "bf16_sustained": float(p.get("bf16_tflops_sustained", 1384.0)), "source": "measured (MEASURED_PEAKS.json)"}
    return {"hbm": 6650.0, "bf16": 1650.0, "bf16_sustained": 1400.0, "source": "fallback (B200_PROFILING.md)"}


def load_traffic(name, key):
    """dram__bytes_read.sum + dram__bytes_write.sum from a committed `ncu --set full` capture (profiles/summarise.py);
    None when no capture exists for this build."""
    try:
        with open(os.path.join(ROOT, "profiles", name)) as f:
            return float(json.load(f)[key])
    except (OSError, KeyError, ValueError, TypeError):
        return None


def train_config(workload, world, gemm_mode, strong=False):
    """`config` of a training line -- built by the same function for the GPU arm and the CPU arm, so both describe the
    same workload."""
    import bench_train
    per_gpu = 256 if not strong else 2048 // world
    names = {"c1": "C1: MLP 784-500-10 (tanh hidden, squared error), batch 128, feature-major",
             "c4": "C4: wide MLP 4 x (4096x4096) tanh, batch 8192, loss = sum(out)"}
    cfg = {"workload": names.get(workload) or
           ("%s: CNN 3 conv3x3 (3-32-64-128, ReLU, 2x nearest down-sampling between them) + 2 FC (8192-256-10), batch %d NHWC "
            "32x32x3 per GPU, squared error, SGD step of every weight" % (workload.upper(), per_gpu)),
           "global_batch": {"c1": 128, "c4": 8192}.get(workload, per_gpu * world),
           "parallelism": "dp%d" % world if workload == "c5" else "single GPU",
           "contractions": gemm_mode, "tolerance": TOLERANCE[gemm_mode],
           "rng": "inputs numpy RandomState(7 + rank); weights from the reference's Wang-hash Box-Muller generator, seeds 51-55 / 151-155",
           "l2": "activations + gradients of one step (~0.4 GB for C3) exceed the 126 MB L2; no flush between steps"}
    if workload == "c5":
        cfg["exchange"] = ("SUM all-reduce of every weight gradient: >= 1 MiB over a grouped NCCL launch on the exchange stream, "
                           "the rest + every update in one NVLink peer-memory launch")
        cfg["scaling"] = "strong (global batch 2048)" if strong else "weak (256 per GPU)"
    if workload in ("c3", "c5"):
        cfg["algorithmic_bytes_per_step_per_gpu"] = bench_train.c3_algorithmic_bytes(per_gpu)
    return cfg


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


# ---- config C2: the fused elementwise + reduce kernels (the metric's second half) ------------------------------------------

def c2_config(rows):
    n = rows * COLS
    return {"workload": "C2: y = tanh(a*b+c) materialised (c broadcast per row), then y.sum; a,b = %dx%d fp32 per GPU" % (rows, COLS),
            "elements_per_gpu": n, "algorithmic_bytes_per_element": BYTES_MAP + BYTES_SUM,
            "l2": "inputs (2 x %.2f GiB) and output exceed the 126 MB L2, no flush needed" % (n * 4 / 2 ** 30),
            "rng": "reference Wang-hash uniform mapped to [-2,2), seeds 1,2,3", "parallelism": "independent shard per GPU"}


def c2_run(be, rows, steps, warmup, rank=0, e2e_steps=0, barrier=None):
    """One C2 'step' = tanh(a*b+c).doCache (ew_vec_kernel) then .sum (sum_kernel).  -> dict with the step time, the
    per-kernel times (map, sum, fused map+sum, column-broadcast variant) and, with e2e_steps, the host-input leg."""
    import deeplearning.scala_b200.tensors as B
    from deeplearning.scala_b200 import _native as N
    barrier = barrier or be.synchronize
    n = rows * COLS
    # uniform, mapped to [-2, 2): the reference's Box-Muller normal does not guard u1 == 0 (Tensors.scala:409-425), so
    # 2^28 draws contain inf/NaN and would turn the checksum into NaN
    a = (be.random([rows, COLS], seed=1 + 16 * rank) * 4.0 - 2.0).do_cache()
    b = (be.random([rows, COLS], seed=2 + 16 * rank) * 4.0 - 2.0).do_cache()
    c = (be.random([rows], seed=3 + 16 * rank) * 4.0 - 2.0).do_cache()

    def step():
        y = be.tanh(a * b + c).do_cache()                    # ew_vec_kernel: read a, b (+c), write y
        return y, y.sum()._buffer()                          # sum_kernel: read y

    def timed(fn, reps):
        fn()
        be.synchronize()
        k0 = be.record_event()
        for _ in range(reps):
            keep = fn()
        k1 = be.record_event()
        be.synchronize()
        return B.elapsed_ms(k0, k1) / reps, keep

    for _ in range(warmup):
        step()
    barrier()
    st0 = be.stats()
    e0 = be.record_event()
    for _ in range(steps):
        keep = step()
    e1 = be.record_event()
    barrier()
    out = {"ms_total": B.elapsed_ms(e0, e1), "launches": be.stats()["kernel_launches"] - st0["kernel_launches"],
           "checksum": float(be.download(keep[1], 1)[0]), "n": n}
    del keep
    reps = max(5, min(steps, 50))
    out["ms_map"], y = timed(lambda: be.tanh(a * b + c).do_cache(), reps)
    out["ms_sum"], _ = timed(lambda: y.sum()._buffer(), reps)
    out["ms_fused"], _ = timed(lambda: be.tanh(a * b + c).sum()._buffer(), reps)      # C2c: nothing materialised, 8 B/element
    del y
    try:                                                     # the per-COLUMN broadcast of c (`c.reshape([1, COLS])`)
        c_col = (be.random([COLS], seed=4 + 16 * rank) * 4.0 - 2.0).do_cache().reshape([1, COLS])
        out["ms_col"], _ = timed(lambda: be.tanh(a * b + c_col).do_cache(), reps)
        del c_col
    except Exception as exc:                                 # noqa: BLE001 -- a side measurement never costs the line
        out["col_error"] = repr(exc)
    if e2e_steps:
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
            ta = be.tensor(ha.reshape(rows, COLS))           # Tensor.apply: host data -> device (dls_buf_from_host)
            tb = be.tensor(hb.reshape(rows, COLS))
            tc = be.tensor(hc)
            sbuf = be.tanh(ta * tb + tc).do_cache().sum()._buffer()
            N.check(N.lib.dls_buf_read(be.ctx, sbuf.handle, 0, pr, 1, None, 0, None))     # blocking D2H of the result
            return float(hr[0])
        e2e_step()
        barrier()
        t0 = time.perf_counter()
        for _ in range(e2e_steps):
            out["e2e_sum"] = e2e_step()
        be.synchronize()
        out["e2e_s"] = time.perf_counter() - t0
        for p in (pa, pb, pc, pr):
            N.lib.dls_host_free(p)
    return out


def c2_kernel_table(r, peak):
    n = r["n"]
    table = {}
    for key, name, nbytes in (("ms_map", "C2a ew_vec_kernel tanh(a*b+c) materialised", BYTES_MAP), ("ms_sum", "C2b sum_kernel", BYTES_SUM),
                              ("ms_fused", "C2c ew_vec_kernel+reduce (fused tanh(a*b+c).sum)", 8),
                              ("ms_col", "C2a column broadcast of c", BYTES_MAP)):
        if key in r:
            gbs = n * nbytes / (r[key] / 1e3) / 1e9
            table[name] = {"ms": r[key], "gbs": gbs, "frac": gbs / peak, "bound": "hbm", "algorithmic_bytes": n * nbytes}
    if "col_error" in r:
        table["C2a column broadcast of c"] = {"error": r["col_error"]}
    return table


def gemm_probe(be, peaks, reps=10):
    """tcgen05 GEMM rate on C4's shape (8192x4096x4096) and 8192^3, bf16 and tf32, operands already staged (the second
    and later launches reuse the bf16 / tf32 shadows, as a training step does for its weights)."""
    import deeplearning.scala_b200.tensors as B
    out = {}
    for m, n, k in ((8192, 4096, 4096), (8192, 8192, 8192)):
        a = (be.random([m, k], seed=91) - 0.5).do_cache()
        b = (be.random([k, n], seed=92) - 0.5).do_cache()
        for mode in ("bf16", "tf32"):
            try:
                for _ in range(2):
                    keep = be.matmul(a, b, mode=mode)._buffer()
                be.synchronize()
                e0 = be.record_event()
                for _ in range(reps):
                    keep = be.matmul(a, b, mode=mode)._buffer()
                e1 = be.record_event()
                be.synchronize()
                ms = B.elapsed_ms(e0, e1) / reps
                tf = 2.0 * m * n * k / (ms / 1e3) / 1e12
                rec = {"ms": ms, "tflops": tf, "bound": "tensor"}
                if mode == "bf16":
                    rec["frac"] = tf / peaks["bf16"]
                    rec["peak"] = peaks["bf16"]
                else:                                        # cuBLAS TF32 8192^3 on this pool, builder-measured
                    rec["frac"] = tf / 742.2                 # (tools/tf32_peak_probe.py, profiles/r2_cublas_peaks.jsonl)
                    rec["peak"] = 742.2
                    rec["peak_source"] = "builder-measured cuBLAS tf32 burst (MEASURED_PEAKS.json has no tf32 entry)"
                out["gemm2_tc_kernel (CTA pair) %s %dx%dx%d" % (mode, m, n, k)] = rec
                del keep
            except Exception as exc:                         # noqa: BLE001
                out["gemm2_tc_kernel (CTA pair) %s %dx%dx%d" % (mode, m, n, k)] = {"error": repr(exc)}
        del a, b
    return out


def c_oracle():
    lib = ctypes.CDLL(os.path.join(ROOT, "oracle", "_build", "libc2oracle.so"))
    fp = ctypes.POINTER(ctypes.c_float)
    lib.c2_map_rows.argtypes = [fp, fp, fp, fp, ctypes.c_size_t, ctypes.c_size_t, ctypes.c_int]
    lib.c2_sum16.argtypes = [fp, ctypes.c_size_t]
    lib.c2_sum16.restype = ctypes.c_float
    lib.c2_threads.restype = ctypes.c_int
    lib.c2_wang_uniform.argtypes = [fp, ctypes.c_size_t, ctypes.c_uint32, ctypes.c_float, ctypes.c_float, ctypes.c_int]
    lib.c2_fused_sum.argtypes = [fp, fp, fp, ctypes.c_size_t, ctypes.c_size_t, ctypes.c_int]
    lib.c2_fused_sum.restype = ctypes.c_double
    return lib


def c2_cpu_runner(rows):
    """The reference's CPU path for C2 on the first `rows` rows of the GPU arm's own inputs (same Wang-hash generator, same
    seeds, same `* 4 - 2` map): (literal step, fused step, threads).  Literal = generated map kernel on all cores, then the
    single-work-item 16-lane sum a CPU OpenCL device executes (Tensors.scala:683-696); fused = BASELINE.md section 5's
    best-effort mode (map and sum in one pass on every core)."""
    lib = c_oracle()
    fp = ctypes.POINTER(ctypes.c_float)
    threads = lib.c2_threads()
    a, b = np.empty((rows, COLS), np.float32), np.empty((rows, COLS), np.float32)
    c, y = np.empty(rows, np.float32), np.empty((rows, COLS), np.float32)
    P = lambda v: v.ctypes.data_as(fp)
    lib.c2_wang_uniform(P(a), a.size, 1, 4.0, -2.0, threads)
    lib.c2_wang_uniform(P(b), b.size, 2, 4.0, -2.0, threads)
    lib.c2_wang_uniform(P(c), c.size, 3, 4.0, -2.0, threads)

    def literal():
        lib.c2_map_rows(P(a), P(b), P(c), P(y), rows, COLS, threads)
        return lib.c2_sum16(P(y), rows * COLS)

    def fused():
        return lib.c2_fused_sum(P(a), P(b), P(c), rows, COLS, threads)
    return literal, fused, threads


# ---- CPU arms (`--impl reference`, and `cpu_baseline` of the GPU line) ----------------------------------------------------

REF_NOTE = ("reference's own CPU path (Compute.scala on a CPU OpenCL device) is unavailable on this box (no JVM, no OpenCL ICD); "
            "this arm times the oracle's restatement of that path")


def cpu_train_runner(workload, batch):
    """One training step of the config on the host cores: oracle/models_oracle.py (numpy + BLAS, fp64 accumulation) on the
    GPU arm's own synthetic batch and initial weights.  -> (step, description)"""
    from oracle import models_oracle as MO
    import bench_train
    w = "c3" if workload == "c5" else workload
    x, t = bench_train.synthetic_batch(w, 0, batch)
    params = bench_train.cpu_model(w)
    state = {"p": params}
    if w == "c3":
        def step():
            convs, fc1, fc2 = state["p"]
            loss, nc, n1, n2 = MO.cnn_step(convs, fc1, fc2, x, t, 1e-3)
            state["p"] = (nc, n1, n2)
            return loss
        return step, "full C3 step (forward, backward, SGD) on %d samples" % batch
    lr, act = (0.01 / 128, "none") if w == "c1" else (1e-4, "tanh")

    def step():
        loss, new, _ = MO.mlp_step(state["p"], x, t, lr, act)
        state["p"] = new
        return loss
    return step, "full %s step on %d samples" % (w.upper(), batch)


def cpu_sample_batch(workload, total_steps):
    """Bounded CPU sample: the full per-GPU batch when the requested steps finish within ~2 minutes, else a slice."""
    full, per_sample_s = {"c1": (128, 2e-5), "c4": (8192, 8e-3)}.get(workload, (256, 0.014))
    fit = int(120.0 / max(total_steps, 1) / per_sample_s)
    batch = full
    while batch > 8 and batch > fit:
        batch //= 2
    return batch, full


def run_reference(args, rank, world):
    """`--impl reference`: rank 0 alone times the CPU restatement of the same workload, K timed steps after W warm-ups."""
    if rank != 0:
        return
    workload = resolve_workload(args, world)
    cores = os.cpu_count() or 1
    if workload == "c2":
        rows = 1024                                          # bounded sample: 2^24 of the 2^28 elements per step
        literal, fused, threads = c2_cpu_runner(rows)
        for _ in range(args.warmup):
            literal()
        t0 = time.perf_counter()
        for _ in range(args.steps):
            literal()
        dt = time.perf_counter() - t0
        value = rows * COLS * args.steps * (BYTES_MAP + BYTES_SUM) / dt / 1e9
        sample = "first %d of %d rows (2^24 elements) of the GPU arm's own inputs per step; map on %d threads + single-thread 16-lane sum" % (rows, ROWS, threads)
        metric, unit, cfg, cores, dtype = C2_METRIC, "GB/s", c2_config(ROWS), threads, "f32"
    else:
        batch, full = cpu_sample_batch(workload, args.steps + args.warmup)
        step, what = cpu_train_runner(workload, batch)
        for _ in range(args.warmup):
            step()
        t0 = time.perf_counter()
        for _ in range(args.steps):
            loss = step()
        dt = time.perf_counter() - t0
        value = batch * args.steps / dt
        sample = what + (" (the whole per-GPU batch)" if batch == full else " (a %d-sample slice of the %d batch)" % (batch, full)) + \
            "; numpy + BLAS on %d host threads, fp64 accumulation; last loss %.6g" % (cores, loss)
        metric, unit, cfg, dtype = METRIC, "samples/s", train_config(workload, world, args.gemm_mode, args.strong), "f32 (fp64 accumulate)"
    emit(json.dumps({
        "impl": "reference", "metric": metric, "value": value, "unit": unit, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": dt / args.steps * 1e3, "higher_is_better": True,
        "scaling": "strong" if (args.strong and workload == "c5") else "weak", "vs_baseline": None, "dtype": dtype, "data": "synthetic",
        "config": cfg, "cpu_baseline": {"value": value, "unit": unit, "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": value, "unit": unit, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}, "note": REF_NOTE}))


def cpu_baseline_train(workload, budget_s=12.0):
    """`cpu_baseline` of the GPU line: the same CPU restatement, full per-GPU batch, ~10-30 s of host work."""
    batch = {"c1": 128, "c4": 512}.get(workload, 256)
    step, what = cpu_train_runner(workload, batch)
    step()
    t0, reps = time.perf_counter(), 0
    while reps < 3 or (time.perf_counter() - t0 < budget_s and reps < 200):
        step()
        reps += 1
    dt = time.perf_counter() - t0
    cores = os.cpu_count() or 1
    return {"value": batch * reps / dt, "unit": "samples/s", "cores": cores, "kind": "port",
            "sample": "%d reps of the %s; oracle/models_oracle.py (numpy + BLAS, fp64 accumulation) on %d host threads" % (reps, what, cores)}


def cpu_baseline_c2(budget_s=8.0):
    rows = 1024
    literal, fused, threads = c2_cpu_runner(rows)
    out = {}
    for name, fn, nbytes in (("literal", literal, BYTES_MAP + BYTES_SUM), ("fused", fused, 8)):
        fn()
        t0, reps = time.perf_counter(), 0
        while reps < 3 or (time.perf_counter() - t0 < budget_s / 2 and reps < 200):
            fn()
            reps += 1
        dt = time.perf_counter() - t0
        out[name] = {"gbs": rows * COLS * reps * nbytes / dt / 1e9, "elements_per_s": rows * COLS * reps / dt, "reps": reps}
    return {"value": out["literal"]["gbs"], "unit": "GB/s", "cores": threads, "kind": "port",
            "sample": "%d reps over the first %d of %d rows (2^24 elements) of the GPU arm's own inputs: C restatement of the generated "
                      "map kernel on %d threads + the single-work-item 16-lane sum (what the reference executes on a CPU device)"
                      % (out["literal"]["reps"], rows, ROWS, threads),
            "best_effort": {"value": out["fused"]["gbs"], "unit": "GB/s at 8 B/element", "elements_per_s": out["fused"]["elements_per_s"],
                            "literal_elements_per_s": out["literal"]["elements_per_s"],
                            "what": "map and sum fused, every core, fp64 partials (BASELINE.md section 5); not what the reference runs"}}


# ---- GPU training legs ---------------------------------------------------------------------------------------------------

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


def train_once(workload, gemm_mode, rank, world, local, dist, steps, warmup, e2e_steps, eager=False, strong=False):
    import deeplearning.scala_b200.tensors as B
    import bench_train
    bench_train.C5_GLOBAL_BATCH = 2048 if (strong and workload == "c5") else None
    be = B.Backend(device=local, gemm_mode=gemm_mode)
    exchange = make_exchange(be, rank, world, dist) if workload == "c5" else None

    def barrier():
        be.synchronize()
        if dist is not None:
            dist.barrier()
        be.synchronize()
    try:
        r = bench_train.run(workload, be, steps, warmup, e2e_steps, exchange, rank, world, barrier, compiled=not eager)
    finally:
        bench_train.C5_GLOBAL_BATCH = None
    times = [r["ms_total"], r["e2e_ms_per_step"]]
    if dist is not None:
        import torch
        t = torch.tensor(times, device="cuda", dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        times = t.tolist()
    r["ms_total"], r["e2e_ms_per_step"] = times
    r["ms_per_step"] = times[0] / steps
    ranks = world if workload == "c5" else 1
    r["samples_per_s"] = ranks * r["batch_per_gpu"] * steps / (times[0] / 1e3)
    r["e2e_samples_per_s"] = ranks * r["batch_per_gpu"] / (times[1] / 1e3)
    r["tflops"] = r["flop_per_step_per_gpu"] / (r["ms_per_step"] / 1e3) / 1e12       # per GPU
    r["gemm_mode"] = gemm_mode
    if exchange is not None:
        r["nccl_ranks"] = world
    del exchange
    be.close()
    return r


def brief(r):
    keys = ("samples_per_s", "e2e_samples_per_s", "ms_per_step", "e2e_ms_per_step", "tflops", "launches_per_step", "gemm_mode",
            "batch_per_gpu", "steps", "last_loss", "mode")
    return {k: r[k] for k in keys if k in r}


def resolve_workload(args, world):
    w = args.workload
    if w == "train":
        return "c3" if world == 1 else "c5"
    if w == "c3" and world > 1:
        return "c5"
    return w


_REAL_STDOUT = None


def setup_nccl_logging():
    """stdout carries exactly one JSON line, but NCCL prints to the process's stdout at every level from WARN up (the
    version banner) and at INFO the `nranks` lines the driver looks for.  Nothing is muted or re-levelled: file descriptor 1
    is pointed at stderr for the rest of the run (so NCCL's and any other native library's stdout text lands on stderr, or in
    the caller's NCCL_DEBUG_FILE if one is set), and the JSON line is written to the original stdout kept in `_REAL_STDOUT`."""
    global _REAL_STDOUT
    if _REAL_STDOUT is None:
        sys.stdout.flush()
        _REAL_STDOUT = os.fdopen(os.dup(1), "w")
        os.dup2(2, 1)


def emit(line: str):
    out = _REAL_STDOUT or sys.stdout
    out.write(line + "\n")
    out.flush()


def main():
    setup_nccl_logging()
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=20)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="train", choices=["train", "c1", "c2", "c3", "c4", "c5"],
                    help="train (default) = C3 on one GPU, C5 (C3 data parallel) on several; c2 = the fused elementwise + reduce config")
    ap.add_argument("--gemm-mode", default="bf16", choices=["fp32", "tf32", "bf16"], help="contraction precision of the training step")
    ap.add_argument("--strong", action="store_true", help="c5: strong scaling, global batch 2048 sharded over the ranks")
    ap.add_argument("--eager", action="store_true", help="time the eager path (one host dispatch per launch) instead of the captured step")
    ap.add_argument("--e2e-steps", type=int, default=None, help="steps of the host-input leg (default: --steps)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-extras", action="store_true", help="headline only: skip the kernel table and the side legs")
    ap.add_argument("--rows", type=int, default=ROWS, help="c2 debug: smaller problem (INVALID as a bench number)")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        run_reference(args, rank, world)
        return
    if args.warmup < 3:
        args.warmup = 3
    dist = None
    if world > 1:
        import torch
        import torch.distributed as dist
        torch.cuda.set_device(local)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    workload = resolve_workload(args, world)
    out = main_c2(args, rank, world, local, dist) if workload == "c2" else main_train(args, workload, rank, world, local, dist)
    if rank == 0:
        emit(json.dumps(out))
    if dist is not None:
        dist.destroy_process_group()


def main_train(args, workload, rank, world, local, dist):
    import bench_train
    peaks = load_peaks()
    strong = args.strong and workload == "c5"
    e2e_steps = args.steps if args.e2e_steps is None else args.e2e_steps
    sampler = ClockSampler(local)
    time.sleep(0.25)
    wall0 = time.time()
    r = train_once(workload, args.gemm_mode, rank, world, local, dist, args.steps, args.warmup, e2e_steps, args.eager, strong)
    wall1 = time.time()
    clocks = sampler.window(wall0, wall1)
    ranks = world if workload == "c5" else 1
    out = {"metric": METRIC, "value": r["samples_per_s"], "unit": "samples/s", "n_gpus": world, "steps": args.steps,
           "warmup": args.warmup, "ms_per_step": r["ms_per_step"], "higher_is_better": True,
           "scaling": "strong" if strong else "weak", "vs_baseline": None, "dtype": DTYPE[args.gemm_mode], "data": "synthetic",
           "tolerance": {"contractions_relative": TOLERANCE[args.gemm_mode],
                         "note": "north_star states 1e-3 for tensor-core GEMMs: tf32 (operands rounded to nearest) meets it, bf16's 8-bit "
                                 "significand cannot (declared deviation, BASELINE.md section 3); the tf32 run of the same step is under `modes`"},
           "config": train_config(workload, world, args.gemm_mode, strong),
           "clocks": clocks,
           "e2e": {"value": r["e2e_samples_per_s"], "unit": "samples/s", "h2d_bytes_per_step": r["h2d_bytes_per_step"] * ranks,
                   "d2h_bytes_per_step": r["d2h_bytes_per_step"] * ranks, "ms_per_step": r["e2e_ms_per_step"], "steps": r["e2e_steps"],
                   "how": r.get("e2e_mode")},
           "gpu_launches": int(round(r["launches_per_step"] * args.steps)) * ranks,
           "launches_per_step": r["launches_per_step"], "driver_allocs_in_timed_region": r["driver_allocs_in_timed_region"],
           "mode": r["mode"], "eager": r["eager"], "captured": r.get("captured"),
           "tflops_per_gpu": r["tflops"], "loss": {"first": r["first_loss"], "last": r["last_loss"]}}
    if "exchange" in r:
        out["config"]["exchange_detail"] = r["exchange"]
        out["nccl"] = {"communicator_ranks": r.get("nccl_ranks"), "log": os.environ.get("NCCL_DEBUG_FILE") or "stderr (fd 1 is pointed at stderr; stdout carries only this line)"}
    if workload in ("c3", "c5"):
        nbytes = bench_train.c3_algorithmic_bytes(r["batch_per_gpu"])
        achieved = nbytes / (r["ms_per_step"] / 1e3) / 1e9
        out["roofline"] = {"kernel": "whole training step (%d launches: conv / GEMM / gather / reduction / update kernels)" % round(r["launches_per_step"]),
                           "bound": "hbm", "achieved": achieved, "peak": peaks["hbm"], "unit": "GB/s", "frac": achieved / peaks["hbm"],
                           "traffic": load_traffic("r2_c3_traffic.json", "step_dram_bytes"), "peak_source": peaks["source"],
                           "algorithmic_bytes_per_step": nbytes, "ms_per_step": r["ms_per_step"],
                           "note": "per GPU; algorithmic bytes = every tensor of the step read / written once per kernel pass "
                                   "(bench_train.c3_algorithmic_bytes, BASELINE.md section 4); per-kernel rooflines under `kernels`"}
    else:
        out["roofline"] = {"kernel": "gemm_tc_kernel (all contractions of the step)", "bound": "tensor", "achieved": r["tflops"],
                           "peak": peaks["bf16_sustained"], "unit": "TFLOP/s", "frac": r["tflops"] / peaks["bf16_sustained"], "traffic": None,
                           "peak_source": peaks["source"] + " bf16_tflops_sustained",
                           "note": "whole-step FLOP / whole-step time (includes elementwise, staging and launch gaps)"}
    extras = not args.no_extras
    if extras and args.steps < 200 and not args.eager:
        # the driver's K can be small: the same captured step over a longer window (sustained clocks), reported beside it
        try:
            s = train_once(workload, args.gemm_mode, rank, world, local, dist, 500, 20, 0 if world > 1 else 50, False, strong)
            out["sustained"] = dict(brief(s), warmup=20)
        except Exception as exc:                             # noqa: BLE001
            out["sustained"] = {"error": "%s: %s" % (type(exc).__name__, exc)}
    if extras and workload == "c5" and not strong:
        try:
            s = train_once("c5", args.gemm_mode, rank, world, local, dist, max(args.steps, 100), 20, 20, False, True)
            out["strong"] = dict(brief(s), global_batch=2048, warmup=20,
                                 note="same model, fixed global batch 2048 sharded over the ranks (SURVEY 8(d)); the driver's "
                                      "efficiency is computed on the weak headline")
        except Exception as exc:                             # noqa: BLE001
            out["strong"] = {"error": "%s: %s" % (type(exc).__name__, exc)}
    if extras and world == 1:
        modes = {}
        for mode in ("tf32", "bf16", "fp32"):
            if mode == args.gemm_mode:
                continue
            try:
                n = 10 if mode == "fp32" else max(args.steps, 100)          # the CUDA-core parity mode is ~100 ms per C3 step
                modes[mode] = dict(brief(train_once(workload, mode, rank, world, local, dist, n, 5 if mode == "fp32" else 10, min(n, 20))),
                                   tolerance=TOLERANCE[mode])
            except Exception as exc:                         # noqa: BLE001
                modes[mode] = {"error": "%s: %s" % (type(exc).__name__, exc)}
        out["modes"] = modes
        legs = {}
        for w, steps in (("c1", 200), ("c4", 20)):
            if w == workload:
                continue
            try:
                s = train_once(w, args.gemm_mode, rank, world, local, dist, steps, 5, min(steps, 20))
                legs[w] = dict(brief(s), workload=s["workload"])
                if w == "c4":
                    legs[w]["frac_of_sustained_bf16_peak"] = s["tflops"] / peaks["bf16_sustained"]
            except Exception as exc:                         # noqa: BLE001 -- a side leg must never take the headline line down
                legs[w] = {"error": "%s: %s" % (type(exc).__name__, exc)}
        out["train"] = legs
        try:
            import deeplearning.scala_b200.tensors as B
            be = B.Backend(device=local)
            c2 = c2_run(be, ROWS, 20, 3)
            kernels = c2_kernel_table(c2, peaks["hbm"])
            kernels["C2 step (map + sum, 16 B/element)"] = {"ms": c2["ms_total"] / 20, "gbs": c2["n"] * 16 / (c2["ms_total"] / 20 / 1e3) / 1e9,
                                                            "checksum": c2["checksum"]}
            kernels["C2a ew_vec_kernel tanh(a*b+c) materialised"]["traffic"] = load_traffic("r1_c2_traffic.json", "ew_vec_kernel<1, 4>")
            be.close()
            be = B.Backend(device=local)
            kernels.update(gemm_probe(be, peaks))
            be.close()
            out["kernels"] = kernels
        except Exception as exc:                             # noqa: BLE001
            out["kernels"] = {"error": "%s: %s" % (type(exc).__name__, exc)}
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        out["cpu_baseline"] = cpu_baseline_train(workload)
    sampler.stop()
    return out


def main_c2(args, rank, world, local, dist):
    import deeplearning.scala_b200.tensors as B
    peaks = load_peaks()
    be = B.Backend(device=local)

    def barrier():
        be.synchronize()
        if dist is not None:
            dist.barrier()
        be.synchronize()
    sampler = ClockSampler(local)
    time.sleep(0.25)
    wall0 = time.time()
    e2e_steps = 5 if args.e2e_steps is None else args.e2e_steps
    r = c2_run(be, args.rows, args.steps, args.warmup, rank, e2e_steps, barrier)
    wall1 = time.time()
    clocks = sampler.window(wall0, wall1)
    sampler.stop()
    be.close()
    times = [r["ms_total"], r.get("e2e_s", 0.0) * 1e3]
    if dist is not None:
        import torch
        t = torch.tensor(times, device="cuda", dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        times = t.tolist()
    n = r["n"]
    bytes_step = n * (BYTES_MAP + BYTES_SUM)
    kernels = c2_kernel_table(r, peaks["hbm"])
    top = kernels["C2a ew_vec_kernel tanh(a*b+c) materialised"]
    out = {"metric": C2_METRIC, "value": world * bytes_step * args.steps / (times[0] / 1e3) / 1e9, "unit": "GB/s", "n_gpus": world,
           "steps": args.steps, "warmup": args.warmup, "ms_per_step": times[0] / args.steps, "higher_is_better": True, "scaling": "weak",
           "vs_baseline": None, "dtype": "f32", "data": "synthetic", "config": c2_config(args.rows), "clocks": clocks,
           "e2e": {"value": world * bytes_step * e2e_steps / (times[1] / 1e3) / 1e9 if e2e_steps else None, "unit": "GB/s",
                   "h2d_bytes_per_step": (2 * n + args.rows) * 4, "d2h_bytes_per_step": 4, "steps": e2e_steps},
           "gpu_launches": r["launches"],
           "roofline": {"kernel": "ew_vec_kernel", "bound": "hbm", "achieved": top["gbs"], "peak": peaks["hbm"], "unit": "GB/s",
                        "frac": top["frac"], "traffic": load_traffic("r1_c2_traffic.json", "ew_vec_kernel<1, 4>") if args.rows == ROWS else None,
                        "peak_source": peaks["source"], "ms_per_launch": top["ms"], "algorithmic_bytes_per_launch": n * BYTES_MAP},
           "kernels": kernels, "checksum": {"sum": r["checksum"], "e2e_sum": r.get("e2e_sum")}}
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        out["cpu_baseline"] = cpu_baseline_c2()
    return out


if __name__ == "__main__":
    main()
