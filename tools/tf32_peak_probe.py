"""Builder-measured library peaks that MEASURED_PEAKS.json (driver-written: HBM copy + cuBLAS bf16) does not hold: cuBLAS
TF32 and, for cross-check, cuBLAS bf16 at 8192^3 through torch.matmul -- burst (best of 10) and sustained (back to back for
2 s), CUDA events.  Measurement aid only: torch / cuBLAS never appear on the product path.  One JSON line per dtype."""
import json
import time

import torch


def run(dtype, tf32, uniform=False):
    torch.backends.cuda.matmul.allow_tf32 = tf32
    n = 8192
    # operand statistics matter on a power-limited part: zero-mean (randn) operands toggle more bits than uniform [0, 1)
    make = (lambda: torch.rand(n, n, device="cuda", dtype=torch.float32).to(dtype)) if uniform else \
        (lambda: torch.randn(n, n, device="cuda", dtype=dtype))
    a, b = make(), make()
    for _ in range(5):
        torch.matmul(a, b)
    torch.cuda.synchronize()
    best = 1e9
    for _ in range(10):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        torch.matmul(a, b)
        e1.record()
        e1.synchronize()
        best = min(best, e0.elapsed_time(e1))
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    reps, t0 = 0, time.time()
    e0.record()
    while time.time() - t0 < 2.0:
        for _ in range(20):
            torch.matmul(a, b)
        reps += 20
        torch.cuda.synchronize()
    e1.record()
    e1.synchronize()
    sustained = e0.elapsed_time(e1) / reps
    flop = 2.0 * n ** 3
    return {"dtype": ("tf32" if tf32 else str(dtype).split(".")[-1]), "operands": "uniform [0,1)" if uniform else "randn", "n": n, "burst_ms": round(best, 4),
            "burst_tflops": round(flop / best / 1e9, 1), "sustained_ms": round(sustained, 4),
            "sustained_tflops": round(flop / sustained / 1e9, 1), "how": "torch.matmul (cuBLAS), builder-measured"}


if __name__ == "__main__":
    print(json.dumps(run(torch.float32, True)), flush=True)
    print(json.dumps(run(torch.bfloat16, False)), flush=True)
    print(json.dumps(run(torch.float32, True, uniform=True)), flush=True)
    print(json.dumps(run(torch.bfloat16, False, uniform=True)), flush=True)
