#!/usr/bin/env python
"""Development probe: OSR_FEED_TRACE=1 prints where a stream-fed host call's time goes; also the resident kernel's
time on small batches (what processing the LAST chunk of a fed call costs at least)."""
import ctypes as C, os, sys, time
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
from puzzletunneldiscovery_b200 import pyosr, workloads as W
from puzzletunneldiscovery_b200._lib import check
uw = bench.make_world(pyosr.UnitWorld); L, h = uw._L, uw._h
N = 1 << 20
q = pyosr.pinned_empty((N, 7)); q[:] = W.random_states(N, 5)
out = pyosr.pinned_empty(N, np.uint8)
qp = q.ctypes.data_as(C.POINTER(C.c_double)); op = out.ctypes.data_as(C.POINTER(C.c_uint8))
for n in (1 << 20,):
    for _ in range(5):
        t = time.perf_counter(); check(L.osr_is_valid_states(h, qp, n, op)); dt = time.perf_counter() - t
        print("call n=%d: %.0f us" % (n, dt * 1e6), file=sys.stderr)
if "--resident" in sys.argv:
    dq = torch.from_numpy(q).cuda(); do = torch.empty(N, dtype=torch.uint8, device="cuda")
    for n in (32, 1024, 8192, 16384, 65536, 262144, 1 << 20):
        ms = []
        for rep in range(6):
            off = (rep * 65536) % (N - n + 1)
            check(L.osr_is_valid_states_device(h, C.c_void_p(dq.data_ptr() + off * 56), n, C.c_void_p(do.data_ptr())))
            check(L.osr_sync(h))
            f = C.c_float(0); check(L.osr_last_kernel_ms(h, C.byref(f))); ms.append(f.value * 1e3)
        print("resident kernel n=%7d: %s us" % (n, " ".join("%.0f" % x for x in ms[1:])), file=sys.stderr)
