"""Device-bound timings (graph replay) of the C3 / C1 contraction shapes: GEMM only -- the operands' staged copies are
cached on the buffers after the warm-up launches.  One JSON line per case with the HBM-roofline time beside it."""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import deeplearning.scala_b200.tensors as B  # noqa: E402

CASES = [  # name, m, n, k, transpose_a, transpose_b
    ("conv1 fwd", 262144, 32, 27, False, False), ("conv2 fwd", 65536, 64, 288, False, False),
    ("conv3 fwd", 16384, 128, 576, False, False), ("conv2 dX", 65536, 288, 64, False, True),
    ("conv3 dX", 16384, 576, 128, False, True), ("conv1 dW", 27, 32, 262144, True, False),
    ("conv2 dW", 288, 64, 65536, True, False), ("conv3 dW", 576, 128, 16384, True, False),
    ("fc1 fwd", 256, 256, 8192, False, False), ("fc1 dW", 8192, 256, 256, True, False),
    ("c1 fwd", 500, 128, 784, False, False), ("c1 dW", 500, 784, 128, False, True),
]


def main():
    be = B.Backend(gemm_mode="bf16")
    only = sys.argv[1:]
    for name, m, n, k, ta, tb in CASES:
        if only and name not in only:
            continue
        a = be.random([k, m] if ta else [m, k], seed=1).do_cache()
        b = be.random([n, k] if tb else [k, n], seed=2).do_cache()
        for _ in range(2):
            be.matmul(a, b, transpose_a=ta, transpose_b=tb)._buffer()
        be.synchronize()
        reps = 20
        with be.capture() as step:
            keep = [be.matmul(a, b, transpose_a=ta, transpose_b=tb)._buffer() for _ in range(reps)]
        step.launch()
        be.synchronize()
        e0 = be.record_event()
        step.launch()
        e1 = be.record_event()
        e1.wait()
        us = B.elapsed_ms(e0, e1) / reps * 1e3
        nodes = step.info()["kernel_nodes"] / reps
        bytes_min = 2 * (m * k + k * n) + 4 * m * n
        print(json.dumps({"case": name, "m": m, "n": n, "k": k, "us": round(us, 2), "kernels": nodes,
                          "tflops": round(2.0 * m * n * k / us / 1e6, 1), "hbm_floor_us": round(bytes_min / 6.5e6, 2)}), flush=True)
        del keep, step, a, b


if __name__ == "__main__":
    main()
