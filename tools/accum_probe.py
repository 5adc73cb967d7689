"""How exact is the fp32 accumulation of the tcgen05 contraction over a very long K?  (GPU probe, not a test.)

The conv1 weight gradient of C3 contracts over K = 262 144 pixels.  This probe runs that shape (M = 27, N = 32) and a few
others on operands that look like the real ones (x uniform in [0, 1), dz 75 % zeros with mixed signs), in every mode, and
prints the error against the float64 product of the SAME rounded operands three ways:
  rel_fro  = |got - ref|_F / |ref|_F                  (what the model tests use; conditioning dependent)
  rel_abs  = max |got - ref| / (|a|^T |b|)            (conditioning independent: the bound tests/test_gpu_gemm.py uses)
  numpy32  = the same two numbers for a plain fp32 BLAS product on the host, as a yardstick
One JSON line per case."""
import json
import sys
import os

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    import deeplearning.scala_b200.tensors as B
    from oracle import tensors_oracle as O
    rng = np.random.RandomState(0)
    for (m, n, k) in ((27, 32, 262144), (27, 32, 20000), (288, 64, 65536), (576, 128, 16384), (500, 784, 128)):
        a = rng.rand(k, m).astype(np.float32)
        b = (rng.standard_normal((k, n)) * (rng.rand(k, n) < 0.25)).astype(np.float32)
        for mode in ("fp32", "tf32", "bf16"):
            ar = O.round_operand(a, None if mode == "fp32" else mode).astype(np.float64)
            br = O.round_operand(b, None if mode == "fp32" else mode).astype(np.float64)
            ref = ar.T @ br
            scale = np.abs(ar).T @ np.abs(br) + 1e-30
            got = B.matmul(B.tensor(a), B.tensor(b), transpose_a=True, mode=mode).flat_array().reshape(m, n).astype(np.float64)
            host = (ar.astype(np.float32).T @ br.astype(np.float32)).astype(np.float64)
            rec = {"m": m, "n": n, "k": k, "mode": mode,
                   "rel_fro": float(np.linalg.norm(got - ref) / np.linalg.norm(ref)),
                   "rel_abs": float((np.abs(got - ref) / scale).max()),
                   "mean_signed_over_scale": float(((got - ref) / scale).mean()),
                   "numpy32_rel_fro": float(np.linalg.norm(host - ref) / np.linalg.norm(ref)),
                   "numpy32_rel_abs": float((np.abs(host - ref) / scale).max())}
            print(json.dumps(rec), flush=True)


if __name__ == "__main__":
    main()
