"""Raw pinned H2D / D2H bandwidth of the box (the roofline of the end-to-end leg) and the effect of the upload chunk
size on osr_is_valid_states with host buffers.  Development aid."""
import ctypes as C, json, os, subprocess, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

n = 1 << 20
if len(sys.argv) > 1 and sys.argv[1] == "child":
    import bench
    from puzzletunneldiscovery_b200 import pyosr, workloads as W
    from puzzletunneldiscovery_b200._lib import check
    uw = bench.make_world(pyosr.UnitWorld)
    L, h = uw._L, uw._h
    hq = [torch.from_numpy(W.random_states(n, s)).pin_memory() for s in range(4)]
    hq32 = [q.to(torch.float32).pin_memory() for q in hq]
    out = torch.empty(n, dtype=torch.uint8).pin_memory()
    for name, fn, arr, ty in (("f64", L.osr_is_valid_states, hq, C.c_double), ("f32", L.osr_is_valid_states_f32, hq32, C.c_float)):
        call = lambda b: check(fn(h, C.cast(arr[b].data_ptr(), C.POINTER(ty)), n, C.cast(out.data_ptr(), C.POINTER(C.c_uint8))))
        for i in range(5): call(i % 4)
        best = 1e9
        for rep in range(3):
            t = time.perf_counter()
            for i in range(20): call(i % 4)
            best = min(best, (time.perf_counter() - t) / 20)
        print("  chunk %s %s: %.3f ms/call  %.3e configs/s" % (os.environ.get("OSR_HOST_CHUNK", "default"), name, best * 1e3, n / best))
    sys.exit(0)

a = torch.empty(n * 7, dtype=torch.float64).pin_memory()
d = torch.empty(n * 7, dtype=torch.float64, device="cuda")
for nb, label in ((n * 56, "58.7 MB"), (n * 28, "29.4 MB")):
    k = nb // 8
    for _ in range(3): d[:k].copy_(a[:k], non_blocking=True)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(10): d[:k].copy_(a[:k], non_blocking=True)
    e1.record(); torch.cuda.synchronize()
    print("raw pinned H2D %s: %.1f GB/s" % (label, 10 * nb / (e0.elapsed_time(e1) * 1e-3) / 1e9))
for ck in ("32768", "65536", "131072", "262144", "524288"):
    env = dict(os.environ, OSR_HOST_CHUNK=ck)
    subprocess.run([sys.executable, __file__, "child"], env=env)
