#!/usr/bin/env python
"""A/B of the shared-memory node budget of K1 on geometry that does not fit (G-wt2, a subdivided alpha puzzle):
OSR_DFS_SMEM_NODES=<n> python scripts/k1_cache_probe.py  ->  kernel-only configs/s from the engine's own events."""
import ctypes as C
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from puzzletunneldiscovery_b200 import pyosr, workloads as W  # noqa: E402

A = W.assets_dir()


def run(name, uw, n):
    q = W.random_states(n, seed=3)
    uw.is_valid_states(q)
    best = 1e9
    for _ in range(3):
        t = time.perf_counter()
        v = uw.is_valid_states(q)
        best = min(best, time.perf_counter() - t)
    ms = C.c_float(0)
    uw._L.osr_last_kernel_ms(uw._h, C.byref(ms))
    qp = pyosr.pinned_empty((n, 7)); qp[:] = q
    op = pyosr.pinned_empty(n, np.uint8)
    uw.is_valid_states(qp, out=op)
    bestp = 1e9
    for _ in range(3):
        t = time.perf_counter()
        vp = uw.is_valid_states(qp, out=op)
        bestp = min(bestp, time.perf_counter() - t)
    assert np.array_equal(vp, v)
    print("%s: %d states, e2e pageable %.3e cfg/s, e2e pinned %.3e cfg/s, kernel %.3f ms = %.3e cfg/s, valid %.3f" %
          (name, n, n / best, n / bestp, ms.value, n / (ms.value * 1e-3), v.mean()), flush=True)


uw = pyosr.UnitWorld()
envf, robf = W.REF_ALPHA_FILES["wt2"]
uw.loadModelFromFile(os.path.join(A, envf)); uw.loadRobotFromFile(os.path.join(A, robf))
uw.enforceRobotCenter(W.ROB_OMPL_CENTER); uw.scaleToUnit(); uw.angleModel(0.0, 0.0)
run("G-wt2", uw, 1 << 21)
eV, eF = W.load_obj_vf(os.path.join(A, "alpha_env-1.2.org.obj"))
rV, rF = W.load_obj_vf(os.path.join(A, "alpha-1.2.org.obj"))
EV, EF = W.subdivide(eV, eF, 22); RV, RF = W.subdivide(rV, rF, 10)
hi = pyosr.UnitWorld(); hi.set_meshes(EV, EF, RV, RF); hi.scaleToUnit(); hi.angleModel(0.0, 0.0)
run("G-hi", hi, 1 << 21)
