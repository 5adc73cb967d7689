"""Lists every distinct plan one training step launches (kernel family, program size, leaf modes) with its launch count."""
import collections
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench_train  # noqa: E402
import deeplearning.scala_b200.tensors as B  # noqa: E402

workload = sys.argv[1] if len(sys.argv) > 1 else "c3"
be = B.Backend(device=0)
model, x, t, batch, flop, desc = bench_train.build(workload, be)
xd = be.tensor(x).do_cache()
td = be.tensor(t).do_cache() if t is not None else None
for _ in range(3):
    model.step(xd, td)
seen = collections.OrderedDict()
orig = B.Backend.launch


def traced(self, roots, out_shape, reduce_sum=False):
    text = self.describe_plan(roots, out_shape)
    key = (text, tuple(out_shape), reduce_sum)
    seen[key] = seen.get(key, 0) + 1
    return orig(self, roots, out_shape, reduce_sum)


B.Backend.launch = traced
model.step(xd, td)
for (text, shape, red), n in seen.items():
    print("x%d shape=%s reduce=%s\n    %s" % (n, list(shape), red, text[:600]))
