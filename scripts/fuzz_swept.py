#!/usr/bin/env python
"""Fuzz of K2's swept windows: the same segments walked with windows on (default) and with every step traversed
(OSR_SWEEP_TILES=0) must give bit-identical outputs (last free state, first hit state, free flag, both taus, step
count).  Regimes: touch rays, random pairs, k-NN edges, segments shorter than a step, walks that START at a
first-contact distance (from = last free state of an earlier walk), per-segment verify_delta, a perturbed
environment; verify_delta from 1e-5 to 1e-2; G-org and G-wt2.   python scripts/fuzz_swept.py [--rounds 3]"""
import argparse
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from puzzletunneldiscovery_b200 import pyosr, workloads as W  # noqa: E402

A = W.assets_dir()


def world(which, sweep, perturb=None):
    if sweep:
        os.environ.pop("OSR_SWEEP_TILES", None)
    else:
        os.environ["OSR_SWEEP_TILES"] = "0"
    w = pyosr.UnitWorld()
    envf, robf = W.REF_ALPHA_FILES[which]
    w.loadModelFromFile(os.path.join(A, envf))
    w.loadRobotFromFile(os.path.join(A, robf))
    if which == "wt2":
        w.enforceRobotCenter(W.ROB_OMPL_CENTER)
    if perturb is not None:
        w.set_perturbation(perturb)
    w.scaleToUnit()
    w.angleModel(0.0, 0.0)
    os.environ.pop("OSR_SWEEP_TILES", None)
    return w


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--rounds", type=int, default=3)
    args = ap.parse_args()
    total = bad = 0
    t0 = time.perf_counter()
    for rnd in range(args.rounds):
        for which in ("org", "wt2"):
            pert = None if rnd % 2 == 0 else np.array([0.01, -0.02, 0.015, 0.9987, 0.03, -0.02, 0.03])
            sw, pl = world(which, True, pert), world(which, False, pert)
            rs = np.random.RandomState(100 + rnd)
            q = W.random_states(60000, seed=200 + rnd)
            free = q[sw.is_valid_states(q)]
            if which == "wt2":
                tv = np.load(os.path.join(A, "alpha-1.2.org.tunnel.npz"))["TUNNEL_V"]
                tun = sw.translate_to_unit_states(tv)
                free = np.concatenate([tun[sw.is_valid_states(tun)], free[:2000]])
            cases = {}
            cases["touch"] = W.touch_rays(free, 4096, seed=300 + rnd)
            a = free[rs.randint(0, len(free), 2048)]
            cases["pairs"] = (a, free[rs.randint(0, len(free), 2048)])
            cases["knn"] = sw.knn_edges(free[:1024], k=4, window=256)
            tiny = a.copy(); tiny[:, 0] += rs.uniform(0, 2e-5, len(a))
            cases["tiny"] = (a, tiny)
            lf = sw.transit_states_to_with_contact(*cases["touch"], 1e-4)[0]
            cases["from_contact"] = W.touch_rays(lf[:2048], 2048, seed=400 + rnd)
            for name, (fr, to) in cases.items():
                for vd in (1e-5, 1e-4, 1e-3, 1e-2, "per-segment"):
                    if vd == 1e-5 and name in ("pairs", "touch"):
                        fr_, to_ = fr[:512], to[:512]
                    else:
                        fr_, to_ = fr, to
                    v = 10.0 ** rs.uniform(-4.5, -2, len(fr_)) if vd == "per-segment" else vd
                    x = sw.transit_states_to_with_contact(fr_, to_, v, with_steps=True)
                    y = pl.transit_states_to_with_contact(fr_, to_, v, with_steps=True)
                    same = all(np.array_equal(p, r) for p, r in zip(x, y))
                    if not isinstance(v, np.ndarray):
                        same = same and np.array_equal(sw.is_valid_transitions(fr_, to_, v),
                                                       pl.is_valid_transitions(fr_, to_, v))
                    total += len(fr_)
                    if not same:
                        bad += 1
                        print("MISMATCH", which, name, vd, "perturbed" if pert is not None else "", flush=True)
            if pert is None:   # K2b: OMPL-frame motion validation, windows on / off
                fr, to = cases["knn"]
                oa, ob = sw.translate_unit_to_ompl(fr), sw.translate_unit_to_ompl(to)
                oc, od = sw.translate_unit_to_ompl(cases["from_contact"][0]), sw.translate_unit_to_ompl(cases["from_contact"][1])
                allp = np.concatenate([oa[:, :3], ob[:, :3], oc[:, :3], od[:, :3]])
                bounds = (allp.min(axis=0) - 1.0, allp.max(axis=0) + 1.0)   # OmplDriver.set_bounds: every state inside
                for res in (1e-4, 1e-3, 1e-2):
                    for (u, v2) in ((oa, ob), (oc[:1024], od[:1024])):
                        x = sw.ompl_validate_motion_pairs(u, v2, bounds, res, with_counts=True)
                        y = pl.ompl_validate_motion_pairs(u, v2, bounds, res, with_counts=True)
                        total += len(u)
                        if not (np.array_equal(x[0], y[0]) and np.array_equal(x[1], y[1])):
                            bad += 1
                            print("MISMATCH ompl", which, res, flush=True)
            print("round %d %s done: %d segments so far, %d mismatching batches, %.0f s" %
                  (rnd, which, total, bad, time.perf_counter() - t0), flush=True)
    print("fuzz_swept: %d segments, %d mismatching batches" % (total, bad))
    return 1 if bad else 0


if __name__ == "__main__":
    sys.exit(main())
