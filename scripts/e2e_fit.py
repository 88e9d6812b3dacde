#!/usr/bin/env python
"""End-to-end validity call on page-locked float64 states at several batch sizes: time = fixed + n * per_state.
The slope says what rate the chunked upload really runs at, the intercept what a call costs besides the upload."""
import ctypes as C
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402
from puzzletunneldiscovery_b200 import pyosr, workloads as W  # noqa: E402
from puzzletunneldiscovery_b200._lib import check  # noqa: E402

uw = bench.make_world(pyosr.UnitWorld)
L, h = uw._L, uw._h
N = 1 << 22
q = pyosr.pinned_empty((N, 7))
q[:] = W.random_states(N, 5)
out = pyosr.pinned_empty(N, np.uint8)
qp = q.ctypes.data_as(C.POINTER(C.c_double))
op = out.ctypes.data_as(C.POINTER(C.c_uint8))
xs, ys = [], []
for n in (1 << 17, 1 << 18, 1 << 19, 1 << 20, 3 << 19, 1 << 21, 1 << 22):
    for _ in range(5):
        check(L.osr_is_valid_states(h, qp, n, op))
    reps = 30
    t = time.perf_counter()
    for _ in range(reps):
        check(L.osr_is_valid_states(h, qp, n, op))
    dt = (time.perf_counter() - t) / reps
    xs.append(n); ys.append(dt)
    print("n=%8d: %8.1f us per call, %.3e configs/s, %.1f GB/s of states" % (n, dt * 1e6, n / dt, n * 56 / dt / 1e9))
a, b = np.polyfit(np.array(xs[2:], float), np.array(ys[2:]), 1)
print("fit over n >= 2^19: fixed %.1f us + %.4f ns per state (= %.1f GB/s)" % (b * 1e6, a * 1e9, 56 / a / 1e9))
