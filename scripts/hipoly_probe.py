#!/usr/bin/env python
"""K1 on G-hi (robot 100 800 x env 487 872 triangles) with DEVICE-RESIDENT states: the launch to profile with ncu
(the host call runs the stream-fed kernel, which must not be captured: under ncu's serialisation it waits for its
input until the feed times out).   python scripts/hipoly_probe.py [n_states]"""
import ctypes as C
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from puzzletunneldiscovery_b200 import pyosr, workloads as W  # noqa: E402
from puzzletunneldiscovery_b200._lib import check  # noqa: E402

A = W.assets_dir()
n = int(sys.argv[1]) if len(sys.argv) > 1 else 1 << 20
eV, eF = W.load_obj_vf(os.path.join(A, "alpha_env-1.2.org.obj"))
rV, rF = W.load_obj_vf(os.path.join(A, "alpha-1.2.org.obj"))
EV, EF = W.subdivide(eV, eF, 22)
RV, RF = W.subdivide(rV, rF, 10)
u0 = pyosr.UnitWorld()
u0.loadModelFromFile(os.path.join(A, "alpha_env-1.2.org.obj")); u0.loadRobotFromFile(os.path.join(A, "alpha-1.2.org.obj"))
u0.scaleToUnit(); u0.angleModel(0.0, 0.0)
uw = pyosr.UnitWorld()
uw.set_meshes(EV, EF, RV, RF)
uw.enforceRobotCenter(u0.robot_center)
uw.scaleToUnit(); uw.angleModel(0.0, 0.0)
L, h = uw._L, uw._h
dq = torch.from_numpy(W.random_states(n, 3)).cuda()
do = torch.empty(n, dtype=torch.uint8, device="cuda")
for rep in range(4):
    uw.counters(reset=True)
    check(L.osr_is_valid_states_device(h, C.c_void_p(dq.data_ptr()), n, C.c_void_p(do.data_ptr())))
    check(L.osr_sync(h))
    f = C.c_float(0)
    check(L.osr_last_kernel_ms(h, C.byref(f)))
    c = uw.counters(reset=True)
    print("G-hi resident: %.3f ms for %d states = %.3e configs/s; box tests/state %.1f, triangle tests/state %.2f" %
          (f.value, n, n / (f.value * 1e-3), c["bv_tests"] / n, c["leaf_tests"] / n), flush=True)
