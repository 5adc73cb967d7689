"""Debug: pipeline timeline of CTA 0 of one tcgen05 GEMM launch.  Needs a trace build:
   make -C deeplearning/scala_b200/csrc clean all NVFLAGS_EXTRA=-DDLS_GEMM_TRACE"""
import ctypes
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import deeplearning.scala_b200.tensors as B  # noqa: E402
from deeplearning.scala_b200 import _native as N  # noqa: E402

m, n, k = (int(v) for v in sys.argv[1:4])
ta, tb = sys.argv[4] == "T", sys.argv[5] == "T"
be = B.Backend(gemm_mode="bf16")
a = be.random([k, m] if ta else [m, k], seed=1).do_cache()
b = be.random([n, k] if tb else [k, n], seed=2).do_cache()
fn = N.lib.dls_debug_gemm_trace
fn.restype = ctypes.c_int
buf = (ctypes.c_ulonglong * 1024)()
for _ in range(3):
    be.matmul(a, b, transpose_a=ta, transpose_b=tb)._buffer()
be.synchronize()
fn(buf, 1024)
be.matmul(a, b, transpose_a=ta, transpose_b=tb)._buffer()
be.synchronize()
cnt = fn(buf, 1024)
ev = sorted(((buf[i] & 0xFFFFFFFFFFFF, buf[i] >> 48) for i in range(cnt)))
names = {1: "start", 2: "end", 10: "producer: slot free", 20: "mma: operands landed", 21: "mma: tile committed",
         30: "epilogue: accumulator ready", 32: "  epilogue: chunk in registers", 33: "  epilogue: chunk finished math", 31: "epilogue: tile stored"}
t0 = ev[0][0]
for t, i in ev[:60]:
    print("%8.2f us  %s" % ((t - t0) / 1e3, names.get(i, i)))
print("events", cnt)
