#!/usr/bin/env python
"""Development aid: run a few config-2 touch segments (G-wt2 tunnel rays) on a trace build of the library
(python puzzletunneldiscovery_b200/build.py --variant trace OSR_K2_TRACE; OSR_LIB=.../libosrgpu_trace.so)."""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from puzzletunneldiscovery_b200 import pyosr, workloads as W  # noqa: E402

A = W.assets_dir()
which = os.environ.get("GEO", "wt2")
uw = pyosr.UnitWorld()
envf, robf = W.REF_ALPHA_FILES[which]
uw.loadModelFromFile(os.path.join(A, envf)); uw.loadRobotFromFile(os.path.join(A, robf))
if which == "wt2":
    uw.enforceRobotCenter(W.ROB_OMPL_CENTER)
uw.scaleToUnit(); uw.angleModel(0.0, 0.0)
if which == "wt2":
    tv = np.concatenate([np.load(os.path.join(A, "alpha-1.2.org.tunnel.npz"))["TUNNEL_V"],
                         np.load(os.path.join(A, "alpha-1.2.org.tunnel-more.npz"))["TUNNEL_V"]])
    fr0 = uw.translate_to_unit_states(tv)
else:
    q0 = W.random_states(1 << 14, seed=0)
    fr0 = q0[uw.is_valid_states(q0) & ~uw.is_disentangled_states(q0)][:1024]
fr, to = W.touch_rays(fr0, int(os.environ.get("TOUCH_N", "48")), seed=1)
r = uw.transit_states_to_with_contact(fr, to, float(os.environ.get("VD", 0.0125 / 8)), with_steps=True)
print("steps", r[5].tolist())
