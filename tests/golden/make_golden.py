"""Regenerates the golden fixtures of tests/golden/ from the CPU oracle.

The reference binary (pyosr + FCL) cannot be built in this image, so the vectors come from the
oracle, whose label mode used here is the BRUTE-FORCE all-pairs SAT-17 (no BVH involved): the
ground truth both the oracle's own BVH mode and the CUDA engine must reproduce bit for bit.
Run from the repo root:  python tests/golden/make_golden.py
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import oracle as O  # noqa: E402
from puzzletunneldiscovery_b200 import workloads as W  # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))
A = W.assets_dir()


def world(which, centre=None):
    envf, robf = W.REF_ALPHA_FILES[which]
    w = O.OracleWorld()
    w.loadModelFromFile(os.path.join(A, envf)); w.loadRobotFromFile(os.path.join(A, robf))
    if centre is not None:
        w.enforceRobotCenter(centre)
    w.angleModel(0.0, 0.0)
    return w


def main():
    ow = world("org")
    q = W.random_states(4096, seed=0)
    brute = ow.is_valid_states(q, mode="brute")
    bvh, cnt = ow.is_valid_states(q, counters=True)
    assert (brute == bvh).all()
    dis = ow.is_disentangled_states(q)
    np.savez_compressed(os.path.join(HERE, "validity_org_seed0.npz"), q=q, valid=brute, disentangled=dis,
                        n_bv=cnt[:, 0], n_leaf=cnt[:, 1], scene_scale=ow.scene_scale)
    free = q[(brute == 1) & (dis == 0)][:64]
    fr, to = W.touch_rays(free, 512, seed=1)
    vd = 0.0125 / 8
    lf, fh, isf, ft, tt, c = ow.transit_states_to_with_contact(fr, to, vd, counters=True)
    vt = ow.is_valid_transitions(fr, to, 0.01)
    np.savez_compressed(os.path.join(HERE, "segments_org_seed1.npz"), fr=fr, to=to, vd=vd, last_free=lf, first_hit=fh,
                        is_free=isf, free_tau=ft, touch_tau=tt, steps=c[:, 0], valid_transition_0p01=vt)
    hit = fh[isf == 0][:48]
    off, segs, faces, cop, total = ow.intersecting_segments_batch(hit)
    np.savez_compressed(os.path.join(HERE, "contacts_org.npz"), q=hit, offsets=off, segs=segs, faces=faces,
                        coplanar=cop, total=total)
    ow2 = world("wt2", W.ROB_OMPL_CENTER)
    q2 = W.random_states(1024, seed=5)
    b2 = ow2.is_valid_states(q2, mode="brute")
    np.savez_compressed(os.path.join(HERE, "validity_wt2_seed5.npz"), q=q2, valid=b2, scene_scale=ow2.scene_scale)
    print("golden fixtures written:", sorted(f for f in os.listdir(HERE) if f.endswith(".npz")))


if __name__ == "__main__":
    main()
