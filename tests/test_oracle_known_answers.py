"""Pins the CPU oracle on every known answer the reference holds for this path
and on the committed golden vectors.  CPU only."""
import os

import numpy as np
import pytest

from conftest import ASSETS, golden, make_world
from puzzletunneldiscovery_b200 import workloads as W


def test_sancheck_fclcontacts_triangle_pair(oracle_mod):
    # /root/reference/sancheck/fclcontacts/main.cc:21-22,46-54: collide == true, 2 contacts
    P = np.array([[0, 0, 0], [1, 0, 0], [0, 0, 1]], float)
    Q = np.array([[0, 0, 0], [1, 0, 0], [0, 1, 0]], float) + np.array([-0.49, -0.5, 0.5])
    assert oracle_mod.sat17(P, Q)
    assert oracle_mod.contact_multiplicity(P, Q) == 2
    if oracle_mod.has_reference_moller():  # same pair through the reference's own Moller97d (main.cc:98-127)
        hit, cop, p0, p1 = oracle_mod.moller_isectline(P, Q)
        assert hit == 1 and cop == 0
        assert np.allclose(sorted([p0[0], p1[0]]), [0.0, 0.01]) and np.allclose([p0[2], p1[2]], 0.5)


def test_sat17_touching_counts_as_intersecting_and_separated_does_not(oracle_mod):
    P = np.array([[0, 0, 0], [1, 0, 0], [0, 1, 0]], float)
    assert oracle_mod.sat17(P, P + np.array([1.0, 0, 0]))        # shares exactly one vertex
    assert not oracle_mod.sat17(P, P + np.array([1.0 + 1e-12, 0, 0]))
    assert not oracle_mod.sat17(P, P + np.array([0, 0, 1e-9]))   # parallel planes
    assert oracle_mod.sat17(P, P)                                # coplanar, identical


def test_rl_test_collide_expecting_true(oracle_mod):
    # /root/reference/src/RL/test_collide.py:40-43 on res/alpha/env-1.2.obj + robot.obj
    w = oracle_mod.OracleWorld()
    w.loadModelFromFile(os.path.join(ASSETS, "env-1.2.obj"))
    w.loadRobotFromFile(os.path.join(ASSETS, "robot.obj"))
    w.angleModel(0.0, 0.0)
    q = np.array([[0.04, 0, 0, .5, .5, .5, .5], [0.7, 0, 0, .5, .5, .5, .5]])
    assert w.is_valid_states(q).tolist() == [1, 1]


def _ompl_path_unit(scale):
    P = np.loadtxt(os.path.join(ASSETS, "alpha-1.2.org.path"))  # x y z qx qy qz qw (OMPL, w last)
    return np.concatenate([P[:, :3] * scale, P[:, 6:7], P[:, 3:6]], axis=1)


def test_solution_path_is_valid_and_connected(ow_org_centred, ow_wt2):
    # preprocess_key.py:129-133 / add_puzzle.py:86-89: every solution vertex must be valid.  The path is in
    # OMPL coordinates, whose robot origin is rob_ompl_center (src/GP/aniconf12_2.py:19).
    for w in (ow_org_centred, ow_wt2):
        up = _ompl_path_unit(w.scene_scale)
        assert len(up) == 81
        assert w.is_valid_states(up).all()
    up = _ompl_path_unit(ow_org_centred.scene_scale)
    brute = ow_org_centred.is_valid_states(up, mode="brute")
    assert brute.all()
    # consecutive vertices connect at the pipeline's resolution (recommended_cres = scale * 0.001)
    vd = ow_org_centred.scene_scale * 0.001
    ok = ow_org_centred.is_valid_transitions(up[:-1], up[1:], vd)
    assert ok.mean() > 0.9  # OMPL validated at its own (coarser) resolution; report, do not require 100%
    wf = np.load(os.path.join(ASSETS, "alpha-1.2.org.w-first.npz"))["VS"]
    uw = np.concatenate([wf[:, :3] * ow_org_centred.scene_scale, wf[:, 3:]], axis=1)
    assert ow_org_centred.is_valid_states(uw).all()


def test_tunnel_vertices_valid_on_wt2(ow_wt2):
    # alpha-1.2.org.tunnel.npz TUNNEL_V: from-states of condor_touch_configuration.py:71-92
    tv = np.load(os.path.join(ASSETS, "alpha-1.2.org.tunnel.npz"))["TUNNEL_V"]
    u = np.concatenate([tv[:, :3] * ow_wt2.scene_scale, tv[:, 3:]], axis=1)
    assert ow_wt2.is_valid_states(u).all()
    assert ow_wt2.is_valid_states(u, mode="brute").all()


def test_unit_scale_matches_survey(ow_org):
    assert abs(1.0 / ow_org.scene_scale - 65.945251) < 1e-4


def test_golden_validity_bvh_equals_bruteforce_truth(ow_org, ow_wt2):
    g = golden("validity_org_seed0.npz")
    assert np.array_equal(W.random_states(4096, seed=0), g["q"])  # generator is reproducible
    v, cnt = ow_org.is_valid_states(g["q"], counters=True)
    assert np.array_equal(v, g["valid"])
    assert np.array_equal(cnt[:, 0], g["n_bv"]) and np.array_equal(cnt[:, 1], g["n_leaf"])
    assert np.array_equal(ow_org.is_disentangled_states(g["q"]), g["disentangled"])
    g2 = golden("validity_wt2_seed5.npz")
    assert np.array_equal(ow_wt2.is_valid_states(g2["q"]), g2["valid"])


def test_golden_segments_and_invariants(ow_org):
    g = golden("segments_org_seed1.npz")
    lf, fh, isf, ft, tt, c = ow_org.transit_states_to_with_contact(g["fr"], g["to"], float(g["vd"]), counters=True)
    assert np.array_equal(isf, g["is_free"]) and np.array_equal(c[:, 0], g["steps"])
    assert np.array_equal(lf, g["last_free"]) and np.array_equal(tt, g["touch_tau"])
    # runtime invariant of unit_world.cc:325-326: last_free is valid; first_hit of a blocked ray is not
    assert ow_org.is_valid_states(lf).all()
    assert not ow_org.is_valid_states(fh[isf == 0]).any()
    assert (tt[isf == 1] == 0.0).all()  # shadowed tau on the free path (unit_world.cc:452-455,473)


def test_transit_degenerate_returns(ow_org):
    q = golden("validity_org_seed0.npz")["q"]
    a, b = q[0:1], q[1:2]
    lf, fh, isf, ft, tt = ow_org.transit_states_to_with_contact(a, b, 100.0)  # vd >= dist
    assert np.array_equal(lf, a) and np.array_equal(fh, b) and isf[0] == 0 and ft[0] == 0.0 and tt[0] == 1.0
    assert ow_org.is_valid_transitions(a, a, 0.01)[0] == 0  # from == to: dist 0 -> (from,to,false,...)


def test_state_math_against_closed_forms(oracle_mod):
    ident = np.array([0, 0, 0, 1, 0, 0, 0.0])
    qz90 = np.array([1, 2, 3, np.cos(np.pi / 4), 0, 0, np.sin(np.pi / 4)])
    assert abs(oracle_mod.distance(ident, qz90) - (np.sqrt(14) + np.pi / 2)) < 1e-12
    mid = oracle_mod.interpolate(ident, qz90, 0.5)
    assert np.allclose(mid, [0.5, 1, 1.5, np.cos(np.pi / 8), 0, 0, np.sin(np.pi / 8)], atol=1e-12)
    ap = oracle_mod.apply(ident, [1, 0, 0], [0, 0, np.pi / 2])
    assert np.allclose(ap, [1, 0, 0, np.cos(np.pi / 4), 0, 0, np.sin(np.pi / 4)], atol=1e-12)
    assert np.allclose(oracle_mod.apply(qz90, [0, 0, 0], [0, 0, 0]), qz90, atol=1e-12)  # zero angle -> axis (1,0,0)


def test_contacts_golden_and_moller(ow_org, oracle_mod):
    g = golden("contacts_org.npz")
    off, segs, faces, cop, total = ow_org.intersecting_segments_batch(g["q"])
    assert total == int(g["total"]) and np.array_equal(off, g["offsets"]) and np.array_equal(faces, g["faces"])
    assert np.array_equal(segs, g["segs"])
    assert total > 0 and (np.diff(off) > 0).all()  # every first-hit state is in collision
    # property (FCL gtest test_fcl_collision.cpp:560-640): the pair set does not depend on the BVH --
    # compare with a brute-force enumeration for one state
    V0, F0 = ow_org.unit_geometry(0)
    V1, F1 = ow_org.unit_geometry(1)
    q = g["q"][0]
    w, x, y, z = q[3:7]
    R = np.array([[1 - 2 * (y * y + z * z), 2 * (x * y - z * w), 2 * (x * z + y * w)],
                  [2 * (x * y + z * w), 1 - 2 * (x * x + z * z), 2 * (y * z - x * w)],
                  [2 * (x * z - y * w), 2 * (y * z + x * w), 1 - 2 * (x * x + y * y)]])
    Q = V1 @ R.T + q[:3]
    got = set(map(tuple, faces[off[0]:off[1]]))
    lo = Q[F1].min(axis=1); hi = Q[F1].max(axis=1)
    brute = set()
    for i, f in enumerate(F0):
        P = V0[f]
        cand = np.nonzero((lo <= P.max(axis=0) + 1e-9).all(axis=1) & (hi >= P.min(axis=0) - 1e-9).all(axis=1))[0]
        for j in cand:
            if oracle_mod.sat17(P, Q[F1[j]]) and oracle_mod.contact_multiplicity(P, Q[F1[j]]) > 0:
                brute.add((i, int(j)))
    assert got == brute


def test_surface_map_closed_form(oracle_mod):
    """A16 (unit_world.cc:1092-1176): coplanar overlapping pair -> one row with the analytic barycentrics;
    a crossing (non-coplanar) pair and a coplanar but disjoint pair -> no row.  Rows are ordered mesh face
    outer, surface face inner (Eigen column-major SparseMatrix iteration)."""
    O = oracle_mod
    if not O.has_reference_moller():
        pytest.skip("oracle/_ref/libmoller97d.so not built")
    TV = np.array([[0, 0, 0], [1, 0, 0], [0, 1, 0], [5, 5, 0], [6, 5, 0], [5, 6, 0]], float)
    TF = np.array([[0, 1, 2], [3, 4, 5]], np.int32)
    V = np.array([[0.25, 0.25, 0], [0.5, 0.25, 0], [0.25, 0.5, 0],          # inside surface face 0, coplanar
                  [0.2, 0.2, -1], [0.2, 0.2, 1], [0.4, 0.4, 1],             # crosses face 0, not coplanar
                  [5.1, 5.1, 0], [5.5, 5.1, 0], [5.1, 5.5, 0]], float)      # inside surface face 1
    F = np.array([[6, 7, 8], [3, 4, 5], [0, 1, 2]], np.int32)
    f, L = O.surface_map(TV, TF, V, F)
    assert f.tolist() == [[3, 4, 5], [0, 1, 2]]
    assert np.allclose(L[0:3], [[0.8, 0.1, 0.1], [0.4, 0.5, 0.1], [0.4, 0.1, 0.5]], atol=1e-15)
    assert np.allclose(L[3:6], [[0.5, 0.25, 0.25], [0.25, 0.5, 0.25], [0.25, 0.25, 0.5]], atol=1e-15)


def test_chart_splat_rules_closed_form(oracle_mod):
    """N4 restatement (bary.geom / bary.frag, osr_render.cc:745-748): pixel-centre coverage, shared edges counted
    once (additive blending would show a double hit), discard of the part outside the surface face."""
    O = oracle_mod
    eye = np.eye(3, dtype=np.float32)
    one = np.ones(3, np.float32)
    tri = np.array([[0, 0], [1, 0], [0, 1]], np.float32)
    img = O.rasterize_barycentric(tri, eye, one, 8, 8)
    assert img.sum() == 28 and img[0, 0] == 1 and img[7, 0] == 0 and img[0, 7] == 0   # centres with x + y < 8
    quad = np.array([[[0, 0], [1, 0], [1, 1]], [[0, 0], [1, 1], [0, 1]]], np.float32)
    img = O.rasterize_barycentric(quad, np.stack([eye, eye]), np.stack([one, one]) * 2.5, 16, 16)
    assert (img == 2.5).all()                                                         # every pixel exactly once
    # barycentrics (2,-1,0) .. : the primitive sticks out of its face; only the part with all b in [0,1] is drawn
    b = np.array([[1, 0, 0], [-1, 2, 0], [-1, 0, 2]], np.float32)
    img = O.rasterize_barycentric(tri, b, one, 8, 8)
    full = O.rasterize_barycentric(tri, eye, one, 8, 8)
    # ... plus the 8 centres lying exactly ON the face's hypotenuse: interior to the big primitive, b0 == 0 is not < 0
    assert img.sum() == 36 and (img >= full).all() and img[7, 7] == 0
    assert O.rasterize_barycentric(tri, eye * 0, one, 8, 8).sum() == 0                # degenerate: no fragments
