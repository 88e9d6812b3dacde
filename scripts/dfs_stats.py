"""Lane-occupancy counters of the depth-first validity kernel on seeded batches (development aid)."""
import sys, os, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
from puzzletunneldiscovery_b200 import pyosr, workloads as W
uw = bench.make_world(pyosr.UnitWorld)
for n in [int(a) for a in sys.argv[1:]] or [1 << 20]:
    q = W.random_states(n, 0)
    uw.is_valid_states(q[:4096]); uw.counters(reset=True)
    t = time.time(); v = uw.is_valid_states(q); dt = time.time() - t
    c = uw.counters(reset=True)
    it = max(1, c["warp_iters"])
    print("n=%d  %.2f ms e2e  bv/pose %.1f leaf/pose %.2f fp64/pose %.3f | warp iters %d  active %.1f waiting %.1f idle %.1f (of 32)" %
          (n, dt * 1e3, c["bv_tests"] / n, c["leaf_tests"] / n, c["fp64_rechecks"] / n, it, c["lanes_active"] / it,
           c["lanes_waiting"] / it, c["lanes_idle"] / it))
