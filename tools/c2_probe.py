"""Times the C2 kernels alone (map materialised, fused map+sum, a few other N1 programs) -- one JSON line per case.
DLS_VEC_UNROLL=1|2|4 forces the unroll of ew_vec_kernel (debug knob of libdlsb200)."""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import deeplearning.scala_b200.tensors as B  # noqa: E402

ROWS = int(os.environ.get("ROWS", "16384"))
COLS = 16384
STEPS = int(os.environ.get("STEPS", "30"))
NCASES = int(os.environ.get("NCASES", "99"))


def main():
    be = B.Backend(device=0)
    a = (be.random([ROWS, COLS], seed=1) * 4.0 - 2.0).do_cache()
    b = (be.random([ROWS, COLS], seed=2) * 4.0 - 2.0).do_cache()
    c = (be.random([ROWS], seed=3) * 4.0 - 2.0).do_cache()
    n = ROWS * COLS
    cases = [
        ("map tanh(a*b+c)", 12, lambda: be.tanh(a * b + c).do_cache()),
        ("fused tanh(a*b+c).sum", 8, lambda: be.tanh(a * b + c).sum()._buffer()),
        ("map a*b+c", 12, lambda: (a * b + c).do_cache()),
        ("fused (a*b+c).sum", 8, lambda: (a * b + c).sum()._buffer()),
        ("map a+b", 12, lambda: (a + b).do_cache()),
        ("map exp(a)*b", 12, lambda: (be.exp(a) * b).do_cache()),
        ("map shared (a+b)*(a+b)", 12, lambda: ((a + b) * (a + b)).do_cache()),
    ]
    for name, bpe, fn in cases[:NCASES]:
        for _ in range(3):
            keep = fn()
        be.synchronize()
        e0 = be.record_event()
        for _ in range(STEPS):
            keep = fn()
        e1 = be.record_event()
        be.synchronize()
        ms = B.elapsed_ms(e0, e1) / STEPS
        print(json.dumps({"case": name, "unroll": os.environ.get("DLS_VEC_UNROLL", "auto"), "ms": round(ms, 4),
                          "gbs": round(n * bpe / ms / 1e6, 1)}), flush=True)
        del keep


if __name__ == "__main__":
    main()
