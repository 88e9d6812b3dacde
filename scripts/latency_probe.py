#!/usr/bin/env python
"""Latency of small host calls through the C ABI (what a planner that asks one state at a time pays)."""
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
q = W.random_states(1 << 16, 0)
out = np.empty(1 << 16, np.uint8)
for n in (1, 32, 1024, 2048, 4096, 8192, 16384, 65536):
    qq = np.ascontiguousarray(q[:n])
    fn = lambda: check(L.osr_is_valid_states(h, qq.ctypes.data_as(C.POINTER(C.c_double)), n, out.ctypes.data_as(C.POINTER(C.c_uint8))))
    for _ in range(20):
        fn()
    reps = 300
    t = time.perf_counter()
    for _ in range(reps):
        fn()
    dt = (time.perf_counter() - t) / reps
    print("osr_is_valid_states n=%6d: %8.1f us per call  (%.3e configs/s)" % (n, dt * 1e6, n / dt))
v = uw.is_valid_states(q[:4096])
for name, qs in (("random", q[:256]), ("free", q[:4096][v][:256])):
    qs = np.ascontiguousarray(qs)
    o1 = np.empty(1, np.uint8)
    ptrs = [(qs[i:i + 1].ctypes.data_as(C.POINTER(C.c_double))) for i in range(len(qs))]
    op = o1.ctypes.data_as(C.POINTER(C.c_uint8))
    for rep in range(3):
        t = time.perf_counter()
        for p_ in ptrs:
            L.osr_is_valid_states(h, p_, 1, op)
        dt = (time.perf_counter() - t) / len(ptrs)
    print("osr_is_valid_states, one %s state per call over %d different states: %.1f us per call" % (name, len(ptrs), dt * 1e6))
fr, to = W.touch_rays(q[:4096][v][:64], 64, seed=1)
for n in (1, 64):
    a, b = np.ascontiguousarray(fr[:n]), np.ascontiguousarray(to[:n])
    fn = lambda: uw.transit_states_to_with_contact(a, b, 0.0125 / 8)
    for _ in range(10):
        fn()
    t = time.perf_counter()
    for _ in range(100):
        fn()
    print("transit_states_to_with_contact n=%3d: %8.1f us per call" % (n, (time.perf_counter() - t) / 100 * 1e6))
t = time.perf_counter()
for i in range(300):
    uw.is_disentangled(q[i])
print("pyosr.is_disentangled (one state, Python): %.1f us per call" % ((time.perf_counter() - t) / 300 * 1e6))
t = time.perf_counter()
for i in range(300):
    uw.is_valid_state(q[i])
print("pyosr.is_valid_state (one state, Python): %.1f us per call" % ((time.perf_counter() - t) / 300 * 1e6))
