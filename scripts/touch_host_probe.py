#!/usr/bin/env python
"""Config-2 touch segments (G-wt2 tunnel rays, vd 0.0125/8): the host call (numpy arrays in and out) against the same
kernel on device-resident segments -- how much of the call is copies."""
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
uw = pyosr.UnitWorld()
envf, robf = W.REF_ALPHA_FILES["wt2"]
uw.loadModelFromFile(os.path.join(A, envf)); uw.loadRobotFromFile(os.path.join(A, robf))
uw.enforceRobotCenter(W.ROB_OMPL_CENTER); uw.scaleToUnit(); uw.angleModel(0.0, 0.0)
tv = np.concatenate([np.load(os.path.join(A, "alpha-1.2.org.tunnel.npz"))["TUNNEL_V"],
                     np.load(os.path.join(A, "alpha-1.2.org.tunnel-more.npz"))["TUNNEL_V"]])
n = int(sys.argv[1]) if len(sys.argv) > 1 else 1 << 20
fr, to = W.touch_rays(uw.translate_to_unit_states(tv), n, seed=1)
vd = 0.0125 / 8
uw.transit_states_to_with_contact(fr[:4096], to[:4096], vd)
for rep in range(2):
    t = time.perf_counter(); r = uw.transit_states_to_with_contact(fr, to, vd, with_steps=True); dt = time.perf_counter() - t
    print("host call: %.1f ms  %.3e segments/s" % (dt * 1e3, n / dt))
L, h = uw._L, uw._h
dfr, dto = torch.from_numpy(fr).cuda(), torch.from_numpy(to).cuda()
lf = torch.empty((n, 7), dtype=torch.float64, device="cuda"); fh = torch.empty_like(lf)
isf = torch.empty(n, dtype=torch.uint8, device="cuda"); ft = torch.empty(n, dtype=torch.float64, device="cuda"); tt = torch.empty_like(ft)
st = torch.empty(n, dtype=torch.int64, device="cuda")
p = lambda x: C.c_void_p(x.data_ptr())
for rep in range(2):
    check(L.osr_transit_states_device(h, p(dfr), p(dto), n, vd, p(lf), p(fh), p(isf), p(ft), p(tt), p(st))); check(L.osr_sync(h))
    f = C.c_float(0); check(L.osr_last_kernel_ms(h, C.byref(f)))
    print("device resident: %.1f ms  %.3e segments/s" % (f.value, n / (f.value * 1e-3)))
print("same outputs:", bool(np.array_equal(lf.cpu().numpy(), r[0]) and np.array_equal(st.cpu().numpy(), r[5])))
