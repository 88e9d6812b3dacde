#!/usr/bin/env python
"""bench.py's touch leg (G-org, from-states valid and not disentangled, verify_delta 0.0125/8) through the host call,
with the engine's counters: segments/s, box tests per reference step.  Knobs come from the environment
(OSR_SWEEP_TILES / OSR_SWEEP_RMAX / OSR_SWEEP_RSTART / OSR_SWEEP_SHORT_TILE)."""
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from puzzletunneldiscovery_b200 import pyosr, workloads as W  # noqa: E402

A = W.assets_dir()
uw = pyosr.UnitWorld()
envf, robf = W.REF_ALPHA_FILES["org"]
uw.loadModelFromFile(os.path.join(A, envf)); uw.loadRobotFromFile(os.path.join(A, robf))
uw.scaleToUnit(); uw.angleModel(0.0, 0.0)
q0 = W.random_states(1 << 16, seed=0)
ok = uw.is_valid_states(q0) & ~uw.is_disentangled_states(q0)
fr, to = W.touch_rays(q0[ok][:1024], int(os.environ.get("TOUCH_N", "65536")), seed=1)
uw.transit_states_to_with_contact(fr, to, 0.0125 / 8)
uw.counters(reset=True)
t = time.perf_counter()
r = uw.transit_states_to_with_contact(fr, to, 0.0125 / 8, with_steps=True)
dt = time.perf_counter() - t
c = uw.counters(reset=True)
st = r[5]
print("touch G-org: %.3e seg/s, %.1f steps/seg, free %.3f, box tests/step %.2f, leaf tests/step %.3f" %
      (len(fr) / dt, st.mean(), r[2].mean(), c["bv_tests"] / st.sum(), c["leaf_tests"] / st.sum()))
