"""Times dls_matmul_ex (tcgen05 GEMM incl. operand staging) on the current device; prints one JSON line per case."""
import json
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import deeplearning.scala_b200.tensors as B  # noqa: E402


def main():
    be = B.default_backend()
    cases = [(4096, 8192, 4096, False, True), (4096, 8192, 4096, False, False), (4096, 4096, 8192, False, True),
             (4096, 8192, 4096, True, False), (8192, 8192, 8192, False, True)]
    args = [a for a in sys.argv[1:] if not a.startswith("--")]
    modes = args or ["bf16", "tf32"]
    for a in sys.argv[1:]:
        if a.startswith("--case="):
            cases = [cases[int(a.split("=")[1])]]
    for mode in modes:
        for (m, n, k, ta, tb) in cases:
            a = B.random([k, m] if ta else [m, k], seed=1).do_cache()
            b = B.random([n, k] if tb else [k, n], seed=2).do_cache()
            for _ in range(3):
                B.matmul(a, b, transpose_a=ta, transpose_b=tb, mode=mode)._buffer()
            be.synchronize()
            reps = 10
            e0 = be.record_event()
            for _ in range(reps):
                B.matmul(a, b, transpose_a=ta, transpose_b=tb, mode=mode)._buffer()
            e1 = be.record_event()
            e1.wait()
            ms = B.elapsed_ms(e0, e1) / reps
            print(json.dumps({"mode": mode, "m": m, "n": n, "k": k, "ta": ta, "tb": tb, "ms": round(ms, 4),
                              "tflops": round(2.0 * m * n * k / ms / 1e9, 1)}), flush=True)
            del a, b
            be.synchronize()


if __name__ == "__main__":
    main()
