"""Times the N2 gather kernel on representative views; prints one JSON line per case (algorithmic bytes = 4 B written per
output element + 4 B per distinct input element read)."""
import json
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import deeplearning.scala_b200.tensors as B  # noqa: E402


def main():
    be = B.default_backend()
    only = sys.argv[1:] 
    cases = []
    a = B.random([8192, 8192], 1).do_cache()
    cases.append(("transpose 8192x8192", lambda: a.transpose() + 0.0, 2 * a.size * 4))
    x = B.random([256, 32, 32, 64], 2).do_cache()
    cases.append(("permute NHWC->NCHW 256x32x32x64", lambda: x.permute([0, 3, 1, 2]) + 0.0, 2 * x.size * 4))
    cases.append(("translate (1,1) with padding 256x32x32x64", lambda: x.translate([0, 1, 1, 0]) + 0.0, 2 * x.size * 4))
    cases.append(("scale 2x down 256x32x32x64", lambda: x.scale([256, 16, 16, 64]) + 0.0, 2 * (x.size // 4) * 4))
    y = B.random([256, 16, 16, 32], 3).do_cache()
    offs = [(ky - 1, kx - 1) for ky in range(3) for kx in range(3)]
    cases.append(("im2col join of 9 translates 256x16x16x32", lambda: B.join([y.translate([0, oy, ox, 0]) for oy, ox in offs]), (9 + 1) * y.size * 4))
    cases.append(("column broadcast a[i,j]*c[1,j] 8192x8192", None, 0))
    c = B.random([1, 8192], 4).do_cache()
    cases[-1] = ("row-vector broadcast a*c[1,j] 8192x8192", lambda: a * c, 2 * a.size * 4)
    for name, build, nbytes in cases:
        if only and not any(o in name for o in only):
            continue
        for _ in range(3):
            build()._buffer()
        be.synchronize()
        reps = 10
        # the launches are recorded and replayed as one graph, so the host's dispatch cost (tens of microseconds per
        # launch through the Python mirror) stays out of the measurement
        with be.capture() as step:
            keep = [build()._buffer() for _ in range(reps)]
        step.launch()
        be.synchronize()
        e0 = be.record_event()
        step.launch()
        e1 = be.record_event()
        e1.wait()
        ms = B.elapsed_ms(e0, e1) / reps
        del keep, step
        t = build()
        plan = be.describe_plan([t.closure()], t.shape) if not hasattr(t, "_produce") else "join"
        print(json.dumps({"case": name, "ms": round(ms, 4), "gbs": round(nbytes / ms / 1e6, 1), "plan": plan[:60]}), flush=True)


if __name__ == "__main__":
    main()
