#!/usr/bin/env python
"""Development probe: what is the ~100 us that a resident K1 launch costs whatever its size?  Far-away states (free at
the root test: no traversal) against random ones, and single states one launch each."""
import ctypes as C, os, sys
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
from puzzletunneldiscovery_b200 import pyosr, workloads as W
from puzzletunneldiscovery_b200._lib import check
uw = bench.make_world(pyosr.UnitWorld); L, h = uw._L, uw._h
N = 1 << 16
q = W.random_states(N, 5)
far = q.copy(); far[:, 0] += 10.0
do = torch.empty(N, dtype=torch.uint8, device="cuda")


def kernel_us(d, n, reps=5):
    ms = []
    for _ in range(reps):
        check(L.osr_is_valid_states_device(h, C.c_void_p(d.data_ptr()), n, C.c_void_p(do.data_ptr())))
        check(L.osr_sync(h))
        f = C.c_float(0); check(L.osr_last_kernel_ms(h, C.byref(f))); ms.append(f.value * 1e3)
    return float(np.median(ms[1:]))


dq = torch.from_numpy(q).cuda(); dfar = torch.from_numpy(far).cuda()
for n in (32, 1024, 8192, 65536):
    print("n=%6d: random %.0f us, far (root-separated) %.0f us" % (n, kernel_us(dq, n), kernel_us(dfar, n)))
one = []
for i in range(256):
    d1 = dq[i:i + 1].contiguous()
    uw.counters(reset=True)
    us = kernel_us(d1, 1, reps=3)
    c = uw.counters(reset=True)
    one.append((us, c["bv_tests"] // 2, c["leaf_tests"] // 2, i))
one.sort()
print("one state per launch over 256 states: min %.0f, median %.0f, max %.0f us" % (one[0][0], one[128][0], one[-1][0]))
for us, bv, lf, i in one[-5:]:
    print("   state %d: %.0f us, %d box tests, %d triangle tests per launch" % (i, us, bv, lf))
for us, bv, lf, i in one[:3]:
    print("   state %d: %.0f us, %d box tests, %d triangle tests per launch" % (i, us, bv, lf))
