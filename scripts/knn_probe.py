#!/usr/bin/env python
"""K6 timing and checksums (uniform, scaled and clustered states): the neighbour lists must not change when the kernel does."""
import os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
from puzzletunneldiscovery_b200 import pyosr, workloads as W
uw = bench.make_world(pyosr.UnitWorld)
q = W.random_states(1 << 18, 2)
for k, win in ((8, 4096), (4, 256), (16, 4096)):
    uw.knn_windowed(q[:8192], k, win)
    t = time.perf_counter(); idx, d = uw.knn_windowed(q, k, win, with_distance=True); dt = time.perf_counter() - t
    print("k=%d window=%d: %.1f ms  %.3e edges/s  checksum %d %.12f" % (k, win, dt * 1e3, len(q) * k / dt, int(idx.astype(np.int64).sum()), float(d.sum())))
# non-unit frame (OMPL-like coordinates) and clustered states
q2 = q.copy(); q2[:, :3] *= 65.9
idx2 = uw.knn_windowed(q2[:65536], 8, 4096)
print("scaled checksum", int(idx2.astype(np.int64).sum()))
q3 = q[:65536].copy(); q3[:, :3] = q3[:, :3] * 1e-6; q3[:, 3:] = q3[0, 3:] + 1e-7 * q3[:, 3:]
idx3 = uw.knn_windowed(q3, 8, 4096)
print("clustered checksum", int(idx3.astype(np.int64).sum()))
