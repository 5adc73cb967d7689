"""cProfile of the host side of N training steps (where the Python mirror spends its time between launches)."""
import cProfile
import os
import pstats
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench_train  # noqa: E402
import deeplearning.scala_b200.tensors as B  # noqa: E402

workload = sys.argv[1] if len(sys.argv) > 1 else "c1"
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 200
be = B.Backend(device=0)
model, x, t, batch, flop, desc = bench_train.build(workload, be)
xd = be.tensor(x).do_cache()
td = be.tensor(t).do_cache() if t is not None else None
for _ in range(10):
    model.step(xd, td)
be.synchronize()
pr = cProfile.Profile()
pr.enable()
for _ in range(steps):
    model.step(xd, td)
be.synchronize()
pr.disable()
st = pstats.Stats(pr)
st.sort_stats("tottime").print_stats(35)
