"""Where does the time of a stream-fed host call go?  (development probe)"""
import ctypes as C, os, sys, time
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from puzzletunneldiscovery_b200 import pyosr, workloads as W
from puzzletunneldiscovery_b200._lib import check
import bench
uw = bench.make_world(pyosr.UnitWorld, devices=[0]); L, h = uw._L, uw._h
n = 1 << 20
hq = [torch.from_numpy(W.random_states(n, b)).pin_memory() for b in range(4)]
ho = torch.empty(n, dtype=torch.uint8).pin_memory()
dq = hq[0].cuda(); do = torch.empty(n, dtype=torch.uint8, device="cuda")
def call(b):
    check(L.osr_is_valid_states(h, C.cast(hq[b].data_ptr(), C.POINTER(C.c_double)), n, C.cast(ho.data_ptr(), C.POINTER(C.c_uint8))))
for i in range(3): call(i)
for rep in range(3):
    t = time.perf_counter(); call(rep); dt = time.perf_counter() - t
    f = C.c_float(0); check(L.osr_last_kernel_ms(h, C.byref(f)))
    print("host call %.3f ms, kernel span %.3f ms" % (dt * 1e3, f.value))
# plain copies
s = torch.cuda.Stream()
d = torch.empty_like(dq)
torch.cuda.synchronize()
for ck in (n, n // 16):
    t = time.perf_counter()
    with torch.cuda.stream(s):
        for c0 in range(0, n, ck):
            d[c0:c0 + ck].copy_(hq[1][c0:c0 + ck], non_blocking=True)
    s.synchronize(); print("H2D in chunks of %d: %.3f ms" % (ck, (time.perf_counter() - t) * 1e3))
t = time.perf_counter(); ho.copy_(do, non_blocking=True); torch.cuda.synchronize(); print("D2H %.3f ms" % ((time.perf_counter() - t) * 1e3))
for rep in range(2):
    check(L.osr_is_valid_states_device(h, C.c_void_p(dq.data_ptr()), n, C.c_void_p(do.data_ptr()))); check(L.osr_sync(h))
    f = C.c_float(0); check(L.osr_last_kernel_ms(h, C.byref(f))); print("resident kernel %.3f ms" % f.value)
# resident kernel while an unrelated H2D copy runs on another stream
for rep in range(3):
    with torch.cuda.stream(s):
        for c0 in range(0, n, n // 16):
            d[c0:c0 + n // 16].copy_(hq[2][c0:c0 + n // 16], non_blocking=True)
    check(L.osr_is_valid_states_device(h, C.c_void_p(dq.data_ptr()), n, C.c_void_p(do.data_ptr()))); check(L.osr_sync(h)); s.synchronize()
    f = C.c_float(0); check(L.osr_last_kernel_ms(h, C.byref(f))); print("resident kernel under concurrent H2D %.3f ms" % f.value)
