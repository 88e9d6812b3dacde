"""GPU parity tests: libosrgpu.so (through its C ABI, via the pyosr-compatible front end) against the CPU
oracle on identical seeded inputs, against the committed golden fixtures, and -- at BASELINE.json's full
sizes -- through size-independent properties.  Integer / boolean / index results must be bit-exact; fp64
outputs (states, taus, Moeller segments) are produced with the same operation order as the oracle and are
compared bit-exactly as well unless a tolerance is written in the test."""
import numpy as np
import pytest

from conftest import golden, make_world
from puzzletunneldiscovery_b200 import workloads as W

pytestmark = pytest.mark.gpu

VD_TOUCH = 0.0125 / 8  # src/GP/condor_touch_configuration.py:35,92
TOL = 1e-14            # absolute tolerance on fp64 outputs that pass through libm (see _close)


def _close(a, b):
    return np.abs(np.asarray(a, float) - np.asarray(b, float)).max() <= TOL


# ---- config 1: batched is_valid_state --------------------------------------------------------------
def test_validity_golden_fixtures(uw_org, uw_wt2):
    g = golden("validity_org_seed0.npz")
    assert np.array_equal(uw_org.is_valid_states(g["q"]), g["valid"].astype(bool))
    assert np.array_equal(uw_org.is_disentangled_states(g["q"]), g["disentangled"].astype(bool))
    g2 = golden("validity_wt2_seed5.npz")
    assert np.array_equal(uw_wt2.is_valid_states(g2["q"]), g2["valid"].astype(bool))
    assert abs(uw_org.scene_scale - float(g["scene_scale"])) == 0.0
    assert abs(uw_wt2.scene_scale - float(g2["scene_scale"])) == 0.0


def test_reference_known_answers_through_the_engine(uw_wt2, ow_org_centred):
    """SURVEY 8(c) C7, asked of the ENGINE through the drop-in module (tests/test_oracle_known_answers.py asks the
    same of the oracle): the reference's own fixtures and the answers its scripts expect."""
    import os
    from conftest import ASSETS
    from puzzletunneldiscovery_b200 import pyosr
    # (2) src/RL/test_collide.py:40-43 on res/alpha/env-1.2.obj + robot.obj (no scaleToUnit there): "Expecting True"
    w = pyosr.UnitWorld()
    w.loadModelFromFile(os.path.join(ASSETS, "env-1.2.obj"))
    w.loadRobotFromFile(os.path.join(ASSETS, "robot.obj"))
    w.angleModel(0.0, 0.0)
    q = np.array([[0.04, 0, 0, .5, .5, .5, .5], [0.7, 0, 0, .5, .5, .5, .5]])
    assert w.is_valid_states(q).tolist() == [True, True]
    assert w.is_valid_state(q[0]) and w.is_valid_state(q[1])
    # (3) res/alpha/alpha-1.2.org.path (81 OMPL states, w last) and alpha-1.2.org.w-first.npz['VS']: every solution
    # vertex is valid on G-org and on G-wt2 with the robot centre of the OMPL problem (preprocess_key.py:129-133);
    # consecutive vertices connect at the pipeline's resolution exactly where the oracle says they do
    uc = make_world(pyosr.UnitWorld, "org", W.ROB_OMPL_CENTER)
    P = np.loadtxt(os.path.join(ASSETS, "alpha-1.2.org.path"))
    for u in (uc, uw_wt2):
        up = np.concatenate([P[:, :3] * u.scene_scale, P[:, 6:7], P[:, 3:6]], axis=1)
        assert len(up) == 81 and u.is_valid_states(up).all()
    up = np.concatenate([P[:, :3] * uc.scene_scale, P[:, 6:7], P[:, 3:6]], axis=1)
    vd = uc.scene_scale * 0.001
    ok = uc.is_valid_transitions(up[:-1], up[1:], vd)
    assert np.array_equal(ok, ow_org_centred.is_valid_transitions(up[:-1], up[1:], vd).astype(bool)) and ok.mean() > 0.9
    wf = np.load(os.path.join(ASSETS, "alpha-1.2.org.w-first.npz"))["VS"]
    assert uc.is_valid_states(np.concatenate([wf[:, :3] * uc.scene_scale, wf[:, 3:]], axis=1)).all()
    # (4) TUNNEL_V of alpha-1.2.org.tunnel.npz / tunnel-more.npz: valid narrow-tunnel states on G-wt2, the from-states
    # of the reference's touch sampling (condor_touch_configuration.py:71-92), through translate_to_unit_state
    for f in ("alpha-1.2.org.tunnel.npz", "alpha-1.2.org.tunnel-more.npz"):
        tv = np.load(os.path.join(ASSETS, f))["TUNNEL_V"]
        assert uw_wt2.is_valid_states(uw_wt2.translate_to_unit_states(tv)).all()
    # (5) the run-time invariant of transitStateTo (unit_world.cc:325-326): the returned state is valid
    tv = np.load(os.path.join(ASSETS, "alpha-1.2.org.tunnel.npz"))["TUNNEL_V"]
    fr, to = W.touch_rays(uw_wt2.translate_to_unit_states(tv), 2048, seed=99)
    lf, fh, isf, ft, tt = uw_wt2.transit_states_to_with_contact(fr, to, VD_TOUCH)
    assert uw_wt2.is_valid_states(lf).all() and not uw_wt2.is_valid_states(fh[~isf & (tt > 0)]).any()
    nq, free, tau = uw_wt2.transit_state_to(fr[0], to[0], VD_TOUCH)
    # (the single-state call returns tau as a C float, like the reference's std::tuple<StateVector, bool, float>)
    assert np.array_equal(nq, lf[0]) and bool(free) == bool(isf[0]) and tau == float(np.float32(ft[0]))


@pytest.mark.parametrize("which,n,seed", [("org", 1 << 16, 11), ("wt2", 1 << 15, 12)])
def test_validity_matches_oracle_bit_exact(which, n, seed, uw_org, uw_wt2, ow_org, ow_wt2):
    uw, ow = (uw_org, ow_org) if which == "org" else (uw_wt2, ow_wt2)
    q = W.random_states(n, seed)
    uw.counters(reset=True)
    v = uw.is_valid_states(q)
    c = uw.counters()
    vo = ow.is_valid_states(q).astype(bool)
    assert np.array_equal(v, vo)  # ambiguity set is empty: the fp64 re-test decides every uncertain pair
    assert c["poses"] == n and c["bv_tests"] > 0 and c["spills"] == 0


def test_validity_near_contact_states_match_oracle(uw_org, ow_org):
    """States inside the contact band: the two end states of blocked touch rays (last free / first hit)
    and their midpoints -- where the fp32 filter is most often undecided."""
    g = golden("segments_org_seed1.npz")
    blocked = g["is_free"] == 0
    lf, fh = g["last_free"][blocked], g["first_hit"][blocked]
    mid = 0.5 * (lf + fh)
    q = np.concatenate([lf, fh, mid])
    assert np.array_equal(uw_org.is_valid_states(q), ow_org.is_valid_states(q).astype(bool))
    assert uw_org.is_valid_states(lf).all() and not uw_org.is_valid_states(fh).any()


def test_validity_edge_cases(uw_org, ow_org):
    assert uw_org.is_valid_states(np.zeros((0, 7))).shape == (0,)           # empty batch
    q = W.random_states(97, 3)                                               # ragged: 3 tiles + 1 state
    vo = ow_org.is_valid_states(q).astype(bool)
    assert np.array_equal(uw_org.is_valid_states(q), vo)
    for n in (1, 31, 32, 33):
        assert np.array_equal(uw_org.is_valid_states(q[:n]), vo[:n])
    assert uw_org.is_valid_state(q[0]) == bool(vo[0])
    # float32 input widens like pybind11 does (uw_random.py:15-17)
    assert np.array_equal(uw_org.is_valid_states(q.astype(np.float32)), vo)
    # the quaternion is NOT renormalised (osr_state.cc:16-25): a non-unit quaternion gives a non-rigid map
    # 1 + |q|^2 (Rot - 1).  FCL's OBB pruning is not conservative for such maps (its result then depends on
    # its tree), so outside the unit-quaternion domain the engine is held to the tree-independent truth:
    # the all-pairs SAT-17 of the oracle's brute-force mode.
    q2 = q.copy()
    q2[:, 3:] *= 1.25
    assert np.array_equal(uw_org.is_valid_states(q2), ow_org.is_valid_states(q2, mode="brute").astype(bool))
    q3 = q.copy()
    q3[:, 3:] *= 1.0 + 1e-4
    assert np.array_equal(uw_org.is_valid_states(q3), ow_org.is_valid_states(q3, mode="brute").astype(bool))
    # far away: everything separated at the root
    far = q.copy()
    far[:, 0] += 10.0
    assert uw_org.is_valid_states(far).all()
    with pytest.raises(RuntimeError):
        uw_org.is_valid_states(np.zeros((4, 6)))


def test_validity_with_perturbation(oracle_mod):
    from puzzletunneldiscovery_b200 import pyosr
    pert = np.array([0.01, -0.02, 0.015, 0.9980, 0.0366, -0.0366, 0.0329])
    uw = make_world(pyosr.UnitWorld, "org", perturb=pert)
    ow = make_world(oracle_mod.OracleWorld, "org", perturb=pert)
    q = W.random_states(8192, 21)
    assert np.array_equal(uw.is_valid_states(q), ow.is_valid_states(q).astype(bool))


def test_validity_full_size_properties(uw_org, ow_org):
    """BASELINE config 1 at full size (2^20 states): determinism, batch-split invariance, quaternion sign
    invariance (R(q) == R(-q) bit for bit), and the oracle on a 32 Ki random subset."""
    n = 1 << 20
    q = W.random_states(n, 0)
    v = uw_org.is_valid_states(q)
    assert np.array_equal(v, uw_org.is_valid_states(q))
    parts = np.concatenate([uw_org.is_valid_states(q[a:b]) for a, b in ((0, 333333), (333333, 700001), (700001, n))])
    assert np.array_equal(v, parts)
    qn = q.copy()
    qn[:, 3:] = -qn[:, 3:]
    assert np.array_equal(v, uw_org.is_valid_states(qn))
    idx = np.random.RandomState(5).choice(n, 1 << 15, replace=False)
    assert np.array_equal(v[idx], ow_org.is_valid_states(q[idx]).astype(bool))
    g = golden("validity_org_seed0.npz")
    assert np.array_equal(v[:4096], g["valid"].astype(bool))
    assert 0.76 < v.mean() < 0.775


def test_bvh_boxes_contain_their_triangles(uw_org):
    """Conservativeness of the engine's own fp32 OBB tree (labels only equal FCL's if no box prunes a
    triangle it should contain): every leaf box contains its triangle, every parent box its children's
    triangles, evaluated in fp64 on the fp32 node data."""
    for which in (0, 1):
        V, F = uw_org._geometry(which)
        nodes = uw_org.bvh_nodes(which).astype(np.float64)
        link = uw_org.bvh_nodes(which)[:, 15].copy().view(np.int32)
        tris = [None] * len(nodes)
        for i in range(len(nodes) - 1, -1, -1):  # breadth-first numbering: children after parents
            if link[i] < 0:
                tris[i] = [-(int(link[i]) + 1)]
            else:
                tris[i] = tris[link[i]] + tris[link[i] + 1]
        assert sorted(tris[0]) == list(range(len(F)))
        for i in range(len(nodes)):
            A = nodes[i, :12].reshape(3, 4)
            P = V[F[tris[i]].reshape(-1)] - A[:, 3]
            proj = np.abs(P @ A[:, :3].T)
            assert (proj <= nodes[i, 12:15] + 1e-12).all()


# ---- config 2: segment walk / first contact ------------------------------------------------------------
def test_segments_golden_fixture(uw_org):
    g = golden("segments_org_seed1.npz")
    lf, fh, isf, ft, tt, st = uw_org.transit_states_to_with_contact(g["fr"], g["to"], float(g["vd"]), with_steps=True)
    assert np.array_equal(isf, g["is_free"].astype(bool))
    assert np.array_equal(st, g["steps"])                      # first contact at the very same step
    # fp64 outputs: same formulas and operation order; the device libm (atan2/acos/sin) may differ from
    # glibc in the last bit, hence 1e-14 absolute instead of bit equality
    assert _close(tt, g["touch_tau"]) and _close(ft, g["free_tau"])
    assert _close(lf, g["last_free"]) and _close(fh, g["first_hit"])
    vt = uw_org.is_valid_transitions(g["fr"], g["to"], 0.01)
    assert np.array_equal(vt, g["valid_transition_0p01"].astype(bool))


def test_segments_match_oracle_wt2_tunnel(uw_wt2, ow_wt2):
    """D3: from-states = TUNNEL_V on G-wt2 (the reference's alpha-1.2 touch sampling set-up)."""
    import os
    tv = np.load(os.path.join(W.assets_dir(), "alpha-1.2.org.tunnel.npz"))["TUNNEL_V"]
    free = np.concatenate([tv[:, :3] * uw_wt2.scene_scale, tv[:, 3:]], axis=1)
    assert uw_wt2.is_valid_states(free).all()
    fr, to = W.touch_rays(free, 768, seed=31)
    lf, fh, isf, ft, tt, st = uw_wt2.transit_states_to_with_contact(fr, to, VD_TOUCH, with_steps=True)
    olf, ofh, oisf, oft, ott, oc = ow_wt2.transit_states_to_with_contact(fr, to, VD_TOUCH, counters=True)
    assert np.array_equal(isf, oisf.astype(bool)) and np.array_equal(st, oc[:, 0])
    assert _close(tt, ott) and _close(ft, oft) and _close(lf, olf) and _close(fh, ofh)
    # runtime invariant unit_world.cc:325-326
    assert uw_wt2.is_valid_states(lf).all()
    assert not uw_wt2.is_valid_states(fh[~isf]).any()


def test_segments_per_segment_verify_delta_and_degenerate(uw_org, ow_org):
    g = golden("segments_org_seed1.npz")
    fr, to = g["fr"][:128], g["to"][:128]
    vd = np.linspace(0.001, 0.05, 128)
    vd[5] = 100.0  # vd >= dist -> (from, to, false, 0, 1)   unit_world.cc:445-447
    r = uw_org.transit_states_to_with_contact(fr, to, vd, with_steps=True)
    o = ow_org.transit_states_to_with_contact(fr, to, vd, counters=True)
    assert np.array_equal(r[2], o[2].astype(bool)) and np.array_equal(r[5], o[5][:, 0])
    assert _close(r[3], o[3]) and _close(r[4], o[4])
    assert np.array_equal(r[0][5], fr[5]) and np.array_equal(r[1][5], to[5]) and not r[2][5]
    assert r[3][5] == 0.0 and r[4][5] == 1.0 and r[5][5] == 0
    # from == to: dist = 0 <= vd
    assert not uw_org.is_valid_transition(fr[0], fr[0], 0.01)
    # single-state wrappers return the reference's tuple shapes (taus narrowed to C float)
    t = uw_org.transit_state_to_with_contact(fr[0], to[0], VD_TOUCH)
    assert len(t) == 5 and t[0].shape == (7,) and isinstance(t[2], bool)
    t3 = uw_org.transit_state_to(fr[0], to[0], VD_TOUCH)
    assert np.array_equal(t3[0], t[0]) and t3[1] == t[2] and t3[2] == t[3]
    assert uw_org.transit_states_to_with_contact(np.zeros((0, 7)), np.zeros((0, 7)), 0.1)[2].shape == (0,)


def test_segments_fine_verify_delta_long_walks(uw_org, ow_org):
    """recommended_cres-sized steps (D3 ii): tens of thousands of steps per segment."""
    g = golden("segments_org_seed1.npz")
    fr, to = g["fr"][:24], g["to"][:24]
    vd = uw_org.scene_scale * 0.001
    r = uw_org.transit_states_to_with_contact(fr, to, vd, with_steps=True)
    o = ow_org.transit_states_to_with_contact(fr, to, vd, counters=True)
    assert np.array_equal(r[5], o[5][:, 0]) and np.array_equal(r[2], o[2].astype(bool))
    assert _close(r[4], o[4]) and r[5].max() > 10000


# ---- config 3: edge validation -----------------------------------------------------------------------
def test_edge_validation_knn_edges_match_oracle(uw_org, ow_org):
    ms = W.random_states(4096, 2)
    fr, to = W.knn_edges(ms, k=4, window=1024)
    vd = uw_org.scene_scale * 0.001 * 8
    v = uw_org.is_valid_transitions(fr, to, vd)
    assert np.array_equal(v, ow_org.is_valid_transitions(fr, to, vd).astype(bool))
    assert 0 < v.mean() < 1


def test_visibility_matrices(uw_org, ow_org):
    q0 = W.random_states(24, 41)
    q1 = W.random_states(17, 42)
    vd = 0.02
    m2 = uw_org.calculate_visibility_matrix2(q0, True, q1, True, vd)
    assert m2.dtype == np.int8 and m2.shape == (24, 17)
    fr = np.repeat(q0, 17, axis=0)
    to = np.tile(q1, (24, 1))
    assert np.array_equal(m2.reshape(-1), ow_org.is_valid_transitions(fr, to, vd).astype(np.int8))
    vp = uw_org.calculate_visibility_pair(q0, True, q1, True, vd)
    assert vp.shape == (24, 1)
    idx = np.arange(24)
    assert np.array_equal(vp[:, 0], ow_org.is_valid_transitions(q0[idx % 24], q1[idx % 17], vd).astype(np.int8))
    ms = uw_org.calculate_visibility_matrix(q0, True, vd)
    assert np.array_equal(ms, ms.T) and (np.diag(ms) == 0).all()
    # non-unit rows go through translate_to_unit_state first (unit_world.cc:810-817)
    big = q0.copy()
    big[:, :3] /= uw_org.scene_scale
    assert np.array_equal(uw_org.calculate_visibility_matrix2(big, False, q1, True, vd), m2)


# ---- config 5: contacts -> surface ----------------------------------------------------------------------
def _canon(off, faces, seg, cop):
    rows = []
    for i in range(len(off) - 1):
        for k in range(off[i], off[i + 1]):
            rows.append((i, int(faces[k][0]), int(faces[k][1])) + tuple(seg[k]) + (int(cop[k]),))
    return np.array(sorted(rows))


def test_contacts_golden_fixture(uw_org):
    g = golden("contacts_org.npz")
    r = uw_org.intersecting_segments_batch(g["q"], cap=1 << 16, with_bary=True)
    assert r["total"] == int(g["total"]) and np.array_equal(r["offsets"], g["offsets"])
    a = _canon(r["offsets"], r["faces"], r["seg"], r["coplanar"])
    b = _canon(g["offsets"], g["faces"], g["segs"], g["coplanar"])
    assert a.shape == b.shape
    assert np.array_equal(a[:, :3], b[:, :3])          # (state, env triangle, robot triangle) rows
    assert np.array_equal(a[:, 9], b[:, 9])            # coplanar flags
    assert np.array_equal(a[:, 3:9], b[:, 3:9])        # Moeller segment end points, same operation order
    # surface points: barycentric coordinates of each segment's mid point sum to 1 on both triangles
    seg = r["seg"]
    nz = np.abs(seg).sum(axis=1) > 0
    bary = r["bary"][nz]
    assert np.allclose(bary[:, :3].sum(axis=1), 1.0, atol=1e-9) and np.allclose(bary[:, 3:].sum(axis=1), 1.0, atol=1e-9)
    # the mid point reconstructed from the env-triangle barycentrics lies on the segment
    V0, F0 = uw_org._geometry(0)
    mid = 0.5 * (seg[nz, :3] + seg[nz, 3:])
    tri = V0[F0[r["faces"][nz, 0]]]
    rec = (bary[:, :3, None] * tri).sum(axis=1)
    # Moeller snaps plane distances below 1e-6 (unnormalised) to zero (Moller97d.cc:41,603-606), so an end point
    # may sit up to ~1e-3 off the other triangle's plane; typical rows reconstruct to rounding error
    err = np.abs(rec - mid).max(axis=1)
    assert np.median(err) < 1e-12 and err.max() < 1e-2
    p0, p1, mag, faces = uw_org.intersecting_segments(g["q"][0])
    assert p0.shape[1] == 3 and mag.shape == (len(p0), 1) and faces.shape == (len(p0), 2)


def test_contacts_capacity_overflow_flag(uw_org):
    g = golden("contacts_org.npz")
    full = uw_org.intersecting_segments_batch(g["q"], cap=1 << 16)
    cap = int(full["offsets"][5]) + 1
    r = uw_org.intersecting_segments_batch(g["q"], cap=cap)
    assert r["total"] == full["total"] and r["offsets"][-1] == cap
    assert r["overflow"][5] and r["overflow"][6:].all() and not r["overflow"][:5].any()
    free = golden("validity_org_seed0.npz")
    fq = free["q"][free["valid"] == 1][:64]
    e = uw_org.intersecting_segments_batch(fq, cap=16)
    assert e["total"] == 0 and (e["offsets"] == 0).all()


def test_contacts_of_deep_penetrations_large_buckets(oracle_mod, monkeypatch):
    """States with many touching triangle pairs: buckets beyond a threshold (2 048 records; lowered to 64 here so
    that states of the subdivided puzzle cross it, with buckets spanning several 1 024-key tiles) are ranked by a
    whole block instead of a warp (kernels_contacts.cuh).  Rows, flags and segments equal the oracle's, and the
    rows of a state are the same alone and inside a batch, ordered by (env triangle, robot triangle)."""
    import os
    from puzzletunneldiscovery_b200 import pyosr
    A = W.assets_dir()
    eV, eF = W.load_obj_vf(os.path.join(A, "alpha_env-1.2.org.obj"))
    rV, rF = W.load_obj_vf(os.path.join(A, "alpha-1.2.org.obj"))
    EV, EF = W.subdivide(eV, eF, 8)
    RV, RF = W.subdivide(rV, rF, 8)

    def mk(cls):
        w = cls()
        w.set_meshes(EV, EF, RV, RF)
        w.scaleToUnit()
        w.angleModel(0.0, 0.0)
        return w
    monkeypatch.setenv("OSR_REC_WARP_MAX", "64")
    uw = mk(pyosr.UnitWorld)
    monkeypatch.delenv("OSR_REC_WARP_MAX")
    uw_warp = mk(pyosr.UnitWorld)
    q = W.random_states(1024, seed=301)
    r = uw.intersecting_segments_batch(q, cap=1 << 22)
    rw = uw_warp.intersecting_segments_batch(q, cap=1 << 22)
    for k in ("offsets", "faces", "seg", "coplanar"):
        assert np.array_equal(r[k], rw[k])           # block-ranked = warp-ranked, row for row
    rows = np.diff(r["offsets"])
    top = np.argsort(-rows)[:3]
    assert rows[top].min() > 2 * 1024 + 8            # > 1 024 records each: more than one key tile
    ow = mk(oracle_mod.OracleWorld)
    ooff, osegs, ofaces, ocop, ototal = ow.intersecting_segments_batch(q[top], cap=1 << 20)
    r2 = uw.intersecting_segments_batch(q[top], cap=1 << 20)
    assert r2["total"] == ototal and np.array_equal(r2["offsets"], ooff)
    assert np.array_equal(_canon(r2["offsets"], r2["faces"], r2["seg"], r2["coplanar"]), _canon(ooff, ofaces, osegs, ocop))
    for k, i in enumerate(top):                      # same rows, same order, alone and inside the batch
        sl = slice(r["offsets"][i], r["offsets"][i + 1]); s2 = slice(r2["offsets"][k], r2["offsets"][k + 1])
        assert np.array_equal(r["faces"][sl], r2["faces"][s2]) and np.array_equal(r["seg"][sl], r2["seg"][s2])
        f = r["faces"][sl].astype(np.int64)
        assert (np.diff(f[:, 0] * (1 << 32) + f[:, 1]) >= 0).all()


# ---- other hot-path entry points -------------------------------------------------------------------------
def test_enum_free_configurations_labels_are_validity(uw_org):
    V0, F0 = uw_org._geometry(0)
    V1, F1 = uw_org._geometry(1)
    rs = np.random.RandomState(9)

    def sample(V, F, n):
        f = F[rs.randint(0, len(F), n)]
        a, b, c = V[f[:, 0]], V[f[:, 1]], V[f[:, 2]]
        nrm = np.cross(b - a, c - a)
        return (a + b + c) / 3.0, nrm / np.linalg.norm(nrm, axis=1, keepdims=True)

    rp, rn = sample(V1, F1, 16)
    ep, en = sample(V0, F0, 16)
    pts = np.stack([rp, rn, ep, en], axis=1)
    q, valid, counts = uw_org.enum_free_configurations(pts, 1e-4, 64)
    assert q.shape == (16, 64, 7) and np.array_equal(counts, valid.sum(axis=1))
    assert np.array_equal(valid.reshape(-1), uw_org.is_valid_states(q.reshape(-1, 7)))
    # contact construction: R * (rob point + margin * n) lands on env point + margin * n
    from puzzletunneldiscovery_b200.pyosr import _rotate_points
    for b in range(4):
        for i in (0, 17):
            p = _rotate_points(q[b, i], (rp[b] + 1e-4 * rn[b])[None])[0]
            assert np.abs(p - (ep[b] + 1e-4 * en[b])).max() < 1e-9
    _, vm, cm = uw_org.enum_free_configurations(pts, 1e-4, 64, only_median=True)
    assert (cm <= counts).all() and (vm & ~valid).sum() == 0
    # N2: candidates are built on the device -- bit for bit the host construction (unit_world.cc:1340-1373)
    assert np.array_equal(q, uw_org.enum_configuration_states(pts, 1e-4, 64))
    # the run filter (unit_world.cc:1375-1386): middle of every valid run that an invalid sample closes
    want = np.zeros_like(valid)
    for r in range(len(valid)):
        i = 0
        while i < 64:
            if not valid[r, i]:
                i += 1
                continue
            j = i
            while j + 1 < 64 and valid[r, j + 1]:
                j += 1
            if j + 1 < 64:
                want[r, i + (j - i + 1) // 2] = True
            i = j + 1
    assert np.array_equal(vm, want) and np.array_equal(cm, want.sum(axis=1))
    # a larger ragged batch (den not a multiple of anything): labels = validity of the returned states
    rp2, rn2 = sample(V1, F1, 300)
    ep2, en2 = sample(V0, F0, 300)
    pts2 = np.stack([rp2, rn2, ep2, en2], axis=1)
    q2, v2, c2 = uw_org.enum_free_configurations(pts2, 1e-4, 37)
    assert np.array_equal(v2.reshape(-1), uw_org.is_valid_states(q2.reshape(-1, 7))) and 0 < v2.mean() < 1
    assert np.array_equal(q2, uw_org.enum_configuration_states(pts2, 1e-4, 37))


def test_state_frame_maps_round_trip(uw_wt2):
    q = W.random_states(64, 77)
    back = uw_wt2._map(uw_wt2._L.osr_translate_from_unit, uw_wt2.translate_to_unit_states(q))
    # both maps renormalise the quaternion (unit_world.cc:713-738); the seeded states are float32-rounded
    qn = q.copy()
    qn[:, 3:] /= np.linalg.norm(qn[:, 3:], axis=1, keepdims=True)
    assert np.abs(back - qn).max() < 1e-12
    o = uw_wt2.translate_unit_to_ompl(q)
    assert np.abs(uw_wt2.translate_ompl_to_unit(o) - qn).max() < 1e-12
    assert np.allclose(o[:, 6], qn[:, 3]) and np.allclose(o[:, 3:6], qn[:, 4:7])  # w-last <-> w-first


def test_no_cpu_fallback_counters_show_gpu_work(uw_org):
    uw_org.counters(reset=True)
    uw_org.is_valid_states(W.random_states(1024, 1))
    c = uw_org.counters()
    assert c["kernel_launches"] == 1 and c["poses"] == 1024 and c["bv_tests"] >= 1024


# ---- A16: intersection mesh -> surface faces + barycentrics -------------------------------------------
def _world_vertices(q, V):
    """translate_state_to_transform(q) * v with the engine's / Eigen's operation order (no FMA in numpy)."""
    w, x, y, z = q[3:7]
    tx, ty, tz = 2 * x, 2 * y, 2 * z
    twx, twy, twz = tx * w, ty * w, tz * w
    txx, txy, txz = tx * x, ty * x, tz * x
    tyy, tyz, tzz = ty * y, tz * y, tz * z
    R = [[1 - (tyy + tzz), txy - twz, txz + twy], [txy + twz, 1 - (txx + tzz), tyz - twx],
         [txz - twy, tyz + twx, 1 - (txx + tyy)]]
    return np.stack([(R[k][0] * V[:, 0] + R[k][1] * V[:, 1] + R[k][2] * V[:, 2]) + q[k] for k in range(3)], axis=1)


@pytest.mark.gpu
def test_intersecting_to_surface_matches_oracle(uw_org, oracle_mod):
    g = golden("contacts_org.npz")
    q = g["q"][0]
    r = uw_org.intersecting_segments_batch(q, cap=1 << 14)
    assert r["total"] > 0
    Ve, Fe = uw_org._geometry(0)
    Vr, Fr = uw_org._geometry(1)
    Wr = _world_vertices(q, Vr)
    # a stand-in for the CGAL boolean result: shrunken copies of the touching robot and environment triangles
    # (coplanar with, and inside, their source faces) in world space
    rt = np.unique(r["faces"][:, 1])[:24]
    et = np.unique(r["faces"][:, 0])[:24]
    tri = np.concatenate([Wr[Fr[rt]], Ve[Fe[et]]])
    cen = tri.mean(axis=1, keepdims=True)
    V = (cen + 0.5 * (tri - cen)).reshape(-1, 3)
    F = np.arange(len(V), dtype=np.int32).reshape(-1, 3)
    for which, TV, TF, src in ((1, Wr, Fr, rt), (0, Ve, Fe, et)):
        fn = uw_org.intersecting_to_robot_surface if which else uw_org.intersecting_to_model_surface
        f, L = fn(q, True, V, F)
        of, oL = oracle_mod.surface_map(TV, TF, V, F)
        assert np.array_equal(f, of) and len(f) >= len(src)
        assert np.abs(L - oL).max() < 1e-12
        # every shrunken triangle lands on its source face with barycentrics inside [0, 1]
        rows = {tuple(x) for x in f.tolist()}
        assert all(tuple(TF[t].tolist()) in rows for t in src)
        assert np.allclose(L.sum(axis=1), 1.0, atol=1e-9)
    # empty mesh, capacity handling and bad indices
    f, L = uw_org.intersecting_to_robot_surface(q, True, np.zeros((0, 3)), np.zeros((0, 3), np.int32))
    assert f.shape == (0, 3) and L.shape == (0, 3)
    with pytest.raises(RuntimeError):
        uw_org.intersecting_to_robot_surface(q, True, V, F + 10 ** 6)


@pytest.mark.gpu
def test_transit_state_actions_and_sampled_configurations(uw_org, ow_org):
    g = golden("validity_org_seed0.npz")
    free = g["q"][g["valid"] == 1]
    q0 = free[3]
    # 12 discrete actions (unit_world.cc:267-336): the returned state is valid and moved along one axis only
    for action in range(12):
        st, done, prog = uw_org.transit_state(q0, action, 0.05, 0.0125)
        assert uw_org.is_valid_state(st) and 0.0 <= prog <= 1.0
        if action < 6:
            d = st[:3] - q0[:3]
            assert np.abs(np.delete(d, action // 2)).max() < 1e-12 and d[action // 2] * (1 if action % 2 == 0 else -1) >= 0
            assert np.abs(st[3:] - q0[3:] / np.linalg.norm(q0[3:])).max() < 1e-12
        else:
            assert np.abs(st[:3] - q0[:3]).max() < 1e-12
        # same walk as transit_state_to on the explicitly built target, checked against the oracle
        if action == 0:
            to = q0.copy(); to[3:] /= np.linalg.norm(to[3:]); to[0] += float(np.float32(0.0125))
            olf, ofh, oisf, oft, ott = ow_org.transit_state_to_with_contact(q0, to, float(np.float32(0.0125)))
            if oft > 0:
                assert np.abs(st - olf).max() < 1e-12 and done == oisf
    st, done, prog = uw_org.transit_state(q0, 12, 0.05, 0.0125)       # invalid action id
    assert np.array_equal(st, q0) and not done and prog == 0.0
    bad = g["q"][g["valid"] == 0][0]
    st, done, prog = uw_org.transit_state(bad, 0, 0.05, 0.0125)       # invalid initial state
    assert np.array_equal(st, bad) and not done
    # sample_free_configuration (unit_world.cc:1295-1337): contact construction + validity of the result
    V0, F0 = uw_org._geometry(0)
    V1, F1 = uw_org._geometry(1)

    def face(V, F, i):
        a, b, c = V[F[i]]
        n = np.cross(b - a, c - a)
        return (a + b + c) / 3, n / np.linalg.norm(n)
    ep, en = face(V0, F0, 100)
    rp, rn = face(V1, F1, 200)
    q = uw_org.sample_free_configuration(rp, rn, ep, en, 1e-3, max_trials=255, rng=np.random.default_rng(3))
    from puzzletunneldiscovery_b200.pyosr import _rotate_points
    p = _rotate_points(q, (rp + 1e-3 * rn)[None])[0]
    assert np.abs(p - (ep + 1e-3 * en)).max() < 1e-9
    n_world = _rotate_points(np.concatenate([[0, 0, 0], q[3:]]), rn[None])[0]
    assert np.abs(n_world + en).max() < 1e-9                            # robot normal opposes the env normal


# ---- multi-GPU: one handle, several devices (index-range shards, host gather) ---------------------------
@pytest.mark.gpu
def test_two_devices_one_handle_equal_one_device(uw_org):
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    from conftest import make_world as mk
    from puzzletunneldiscovery_b200 import pyosr
    import functools
    uw2 = mk(functools.partial(pyosr.UnitWorld, devices=[0, 1]))
    q = W.random_states(100001, seed=41)
    v1 = uw_org.is_valid_states(q)
    assert np.array_equal(uw2.is_valid_states(q), v1)
    assert np.array_equal(uw2.is_disentangled_states(q), uw_org.is_disentangled_states(q))
    fr, to = W.touch_rays(q[v1][:64], 1001, seed=42)
    a = uw_org.transit_states_to_with_contact(fr, to, VD_TOUCH, with_steps=True)
    b = uw2.transit_states_to_with_contact(fr, to, VD_TOUCH, with_steps=True)
    assert all(np.array_equal(x, y) for x, y in zip(a, b))
    hit = a[1][~a[2]][:257]
    ra = uw_org.intersecting_segments_batch(hit, cap=1 << 16)
    rb = uw2.intersecting_segments_batch(hit, cap=1 << 16)
    assert ra["total"] == rb["total"] and np.array_equal(ra["offsets"], rb["offsets"])
    assert np.array_equal(ra["faces"], rb["faces"]) and np.array_equal(ra["seg"], rb["seg"])
    from puzzletunneldiscovery_b200 import sharding as S
    assert np.array_equal(S.sharded_map(uw2.is_valid_states, [q], np.bool_), v1)   # no process group: plain call
    # candidate-sharded entry points: clearance statistics (K8) and key-configuration candidates (K9)
    cq = np.ascontiguousarray(q[v1][:7])
    cto = W.touch_rays(np.repeat(cq, 33, axis=0), 7 * 33, seed=43)[1].reshape(7, 33, 7)
    sa, da = uw_org.clearance_stats(cq, cto, VD_TOUCH, with_distances=True)
    sb, db = uw2.clearance_stats(cq, cto, VD_TOUCH, with_distances=True)
    assert np.array_equal(sa, sb) and np.array_equal(da, db)
    # window-sharded: k-NN edges validated where they are (three windows over two devices, the last one short)
    ms = q[v1][:2 * 256 + 100]
    ia, oa = uw_org.knn_validate_edges(ms, k=4, window=256, verify_delta=2e-3)
    ib, ob = uw2.knn_validate_edges(ms, k=4, window=256, verify_delta=2e-3)
    assert np.array_equal(ia, ib) and np.array_equal(oa, ob) and ib[256:].min() >= 256
    pts = np.random.RandomState(44).randn(9, 4, 3)
    ka = uw_org.enum_free_configurations(pts, 1e-4, 19, only_median=True)
    kb = uw2.enum_free_configurations(pts, 1e-4, 19, only_median=True)
    assert all(np.array_equal(x, y) for x, y in zip(ka, kb))


@pytest.mark.gpu
def test_float32_states_give_the_same_labels(uw_org, uw_wt2):
    """osr_is_valid_states_f32: float32 samples (uw_random.py:15-17) widened on the device -- exact, so the
    labels equal those of the float64 call on the widened array; ragged sizes cross the upload chunks."""
    for uw, n in ((uw_org, 200003), (uw_wt2, 70001), (uw_org, 5), (uw_org, 65536)):
        q32 = W.random_states(n, seed=n % 97).astype(np.float32)
        assert np.array_equal(uw.is_valid_states(q32), uw.is_valid_states(q32.astype(np.float64)))
    assert uw_org.is_valid_states(np.zeros((0, 7), np.float32)).shape == (0,)


@pytest.mark.gpu
def test_knn_windowed_matches_host_restatement(uw_org):
    """K6: neighbour sets equal the numpy restatement (workloads.state_distance, osr_state.cc:130-154) per
    milestone; distances ascending; ragged last window; a window with fewer than k other states pads with -1."""
    q = W.random_states(2 * 512 + 77, seed=51)
    idx, dist = uw_org.knn_windowed(q, k=8, window=512, with_distance=True)
    assert idx.shape == (len(q), 8) and (np.diff(dist, axis=1) >= 0).all()
    for w0 in range(0, len(q), 512):
        blk = q[w0:w0 + 512]
        m = len(blk)
        for i in range(0, m, 37):
            d = W.state_distance(np.repeat(blk[i:i + 1], m, axis=0), blk)
            d[i] = np.inf
            want = set((np.argsort(d, kind="stable")[:8] + w0).tolist())
            assert set(idx[w0 + i].tolist()) == want
            assert np.abs(np.sort(d)[:8] - dist[w0 + i]).max() < 1e-12
    assert (idx >= 0).all() and (idx != np.arange(len(q))[:, None]).all()
    small = uw_org.knn_windowed(q[:5], k=8, window=256)
    assert (small[:, :4] >= 0).all() and (small[:, 4:] == -1).all()
    fr, to = uw_org.knn_edges(q[:5], k=8, window=256)
    assert fr.shape == (20, 7) and to.shape == (20, 7)
    with pytest.raises(RuntimeError):
        uw_org.knn_windowed(q, k=17, window=512)


@pytest.mark.gpu
def test_chart_splat_matches_restatement_and_chains_from_surface_mapping(oracle_mod):
    """K7: the CUDA rasteriser covers exactly the pixels of the numpy restatement of the
    reference's shaders and adds the same weights; then the whole chain contact state -> intersecting_to_*_surface
    -> add_barycentric -> render_barycentric on a mesh with synthetic per-vertex UVs."""
    from puzzletunneldiscovery_b200 import pyosr
    r = make_world(pyosr.Renderer)
    rs = np.random.RandomState(7)
    P = 300
    uv = rs.rand(P, 3, 2).astype(np.float32)
    bary = rs.dirichlet([1, 1, 1], size=(P, 3)).astype(np.float32)
    bary[::5] *= 1.3                                        # some primitives leave their face: discard path
    w = np.repeat(rs.rand(P, 1).astype(np.float32) + 0.5, 3, axis=1)
    out = np.zeros((64, 64), np.float32)
    from puzzletunneldiscovery_b200._lib import check, c_fp
    check(r._L.osr_rasterize_barycentric(r._h, uv.ctypes.data_as(c_fp), bary.ctypes.data_as(c_fp),
                                         w.ctypes.data_as(c_fp), P, 64, 64, out.ctypes.data_as(c_fp)))
    ref = oracle_mod.rasterize_barycentric(uv, bary, w, 64, 64)
    assert np.array_equal(out > 0, ref > 0) and (ref > 0).sum() > 500
    assert np.abs(out - ref).max() <= 1e-4 * max(1.0, ref.max())   # fp32 additive blending, order of adds differs
    # chain on the alpha robot with a synthetic planar UV unwrap
    g = golden("contacts_org.npz")
    q = g["q"][0]
    V1, F1 = r._geometry(1)
    lo, hi = V1.min(axis=0), V1.max(axis=0)
    r.set_uv(r.BARY_RENDERING_ROBOT, ((V1 - lo) / (hi - lo))[:, :2])
    Wr = _world_vertices(q, V1)
    c = r.intersecting_segments_batch(q, cap=1 << 14)
    rt = np.unique(c["faces"][:, 1])[:16]
    tri = Wr[F1[rt]]
    cen = tri.mean(axis=1, keepdims=True)
    IV = (cen + 0.5 * (tri - cen)).reshape(-1, 3)
    IFc = np.arange(len(IV), dtype=np.int32).reshape(-1, 3)
    IF, IBV = r.intersecting_to_robot_surface(q, True, IV, IFc)
    r.clear_barycentric(r.BARY_RENDERING_ROBOT)
    r.add_barycentric(IF, IBV, r.BARY_RENDERING_ROBOT, 127.0)
    fb = r.render_barycentric(r.BARY_RENDERING_ROBOT, np.array([512, 512], np.int32))
    assert fb.shape == (512, 512) and fb.dtype == np.float32 and fb.max() >= 127.0 and (fb >= 0).all()
    ref = oracle_mod.rasterize_barycentric(r._uv[0][IF.reshape(-1)], IBV, np.full(len(IBV), 127.0), 512, 512)
    assert np.array_equal(fb > 0, ref > 0) and np.abs(fb - ref).max() < 1e-2
    r.clear_barycentric(r.BARY_RENDERING_ROBOT)
    assert r.render_barycentric(r.BARY_RENDERING_ROBOT, [32, 32]).sum() == 0
    with pytest.raises(RuntimeError):
        r.add_barycentric(IF, IBV, r.BARY_RENDERING_SCENE, 1.0)      # no UVs set for the environment


@pytest.mark.gpu
@pytest.mark.parametrize("shift", [None, (1.0, -2.0, 3.0)])
def test_ompl_frame_validators_match_restatement(shift, oracle_mod):
    """N1: OmplDriver.validate_states / validate_motion_pairs semantics (DiscreteMotionValidator::checkMotion with
    OMPL's own SE(3) interpolation and segment count) on OMPL-frame rows, against the oracle's literal
    restatement; `shift` moves the robot centre away from the OMPL centre so the frame map is not a pure scale."""
    from puzzletunneldiscovery_b200 import pyosr
    import functools
    probe = make_world(pyosr.UnitWorld)
    centre = None if shift is None else probe.get_ompl_center(1) + np.asarray(shift)
    uw = make_world(pyosr.UnitWorld, "org", centre)
    ow = make_world(oracle_mod.OracleWorld, "org", centre)
    q = W.random_states(6000, seed=61)
    valid = uw.is_valid_states(q)
    qo = uw.translate_unit_to_ompl(q)
    assert np.array_equal(uw.ompl_validate_states(qo).astype(bool), uw.is_valid_states(uw.translate_ompl_to_unit(qo)))
    assert (uw.ompl_validate_states(qo).astype(bool) == valid).mean() > 0.999     # round trip of the frame maps
    ms = q[valid][:1024]
    fr, to = uw.knn_edges(ms, k=4, window=256)
    a, b = uw.translate_unit_to_ompl(fr), uw.translate_unit_to_ompl(to)
    scale = uw.scene_scale
    dc = (uw.get_ompl_center(1).astype(np.float32) - uw.robot_center.astype(np.float32)).astype(np.float64) * scale
    # translation bounds of the planner (OmplDriver.set_bounds); every milestone lies inside them here
    lo = np.floor(np.minimum(a[:, :3].min(axis=0), qo[:, :3].min(axis=0))) - 1.0
    extent = (lo, lo + np.array([90.0, 80.0, 90.0]))
    assert (qo[:, :3].max(axis=0) < extent[1] - 1.0).all()
    res = 0.01
    got, nd = uw.ompl_validate_motion_pairs(a, b, extent, res, with_counts=True)
    want, ond = ow.ompl_check_motions(a, b, scale, dc, extent, res)
    assert np.array_equal(nd, ond) and nd.max() > 32 and nd.min() >= 1
    assert np.array_equal(got.astype(bool), want.astype(bool)) and 0.05 < got.mean() < 0.999
    # the planner's own resolution (collision_resolution fallback 1e-4, solve.py:265): thousands of interior points
    # per motion, cleared in swept windows -- same booleans as the oracle's point-by-point checkMotion
    fine = 2e-4
    gf, nf = uw.ompl_validate_motion_pairs(a[:192], b[:192], extent, fine, with_counts=True)
    wf, onf = ow.ompl_check_motions(a[:192], b[:192], scale, dc, extent, fine)
    assert np.array_equal(nf, onf) and nf.max() > 1000
    assert np.array_equal(gf.astype(bool), wf.astype(bool)) and 0.05 < gf.mean() < 0.999
    # an invalid goal fails whatever lies in between; coincident states need one check only
    bad = uw.translate_unit_to_ompl(q[~valid][:8])
    g2, n2 = uw.ompl_validate_motion_pairs(a[:8], bad, extent, res, with_counts=True)
    assert not g2.any()
    g3, n3 = uw.ompl_validate_motion_pairs(a[:8], a[:8], extent, res, with_counts=True)
    assert g3.all() and (n3 == 0).all()
    # FCLStateValidityChecker::isValid = satisfiesBounds && collision free: bounds that cut through the milestones
    # (translation), goals whose quaternion is off the unit sphere by more than 1e-9, start states outside the bounds
    # (checkMotion never tests s1, but the interior points next to it are outside too), NaN rows
    mid = np.median(a[:, :3], axis=0)
    tight = (extent[0], np.array([mid[0], extent[1][1], extent[1][2]]))
    gs = uw.ompl_validate_states(qo, tight).astype(bool)
    ws = ow.ompl_validate_states(qo, scale, dc, tight).astype(bool)
    inside = qo[:, 0] <= mid[0]
    assert np.array_equal(gs, ws) and np.array_equal(gs, valid & inside) and 0 < gs.sum() < valid.sum()
    assert np.array_equal(uw.ompl_validate_states(qo, extent).astype(bool), valid)
    gt, nt = uw.ompl_validate_motion_pairs(a, b, tight, res, with_counts=True)
    wt, ont = ow.ompl_check_motions(a, b, scale, dc, tight, res)
    assert np.array_equal(nt, ont) and np.array_equal(gt.astype(bool), wt.astype(bool))
    assert not gt[b[:, 0] > mid[0] + 1e-6].any()                       # a goal outside the bounds is invalid
    straddle = (b[:, 0] < mid[0] - 1e-6) & (nt >= 2) & (a[:, 0] + (b[:, 0] - a[:, 0]) / np.maximum(nt, 1) > mid[0] + 1e-6)
    assert straddle.any() and not gt[straddle].any()                   # interior points next to s1 are outside
    assert gt.sum() < got.sum() and gt.any()
    b_off = b[:64].copy(); b_off[::2, 3:7] *= 1.0 + 1e-8             # every other goal fails the SO(3) norm test
    go = uw.ompl_validate_motion_pairs(a[:64], b_off, extent, res)
    wo, _ = ow.ompl_check_motions(a[:64], b_off, scale, dc, extent, res)
    assert np.array_equal(go.astype(bool), wo.astype(bool)) and not go[::2].any()
    assert np.array_equal(go[1::2], got[:64][1::2])
    a_off = a[:64].copy(); a_off[::2, 3:7] *= 1.0 + 1e-8             # start states are not tested, interior points are
    ga = uw.ompl_validate_motion_pairs(a_off, b[:64], extent, res)
    wa, _ = ow.ompl_check_motions(a_off, b[:64], scale, dc, extent, res)
    assert np.array_equal(ga.astype(bool), wa.astype(bool))
    b_nan = b[:8].copy(); b_nan[0, 0] = np.nan; b_nan[1, 4] = np.nan; b_nan[2, :3] = np.inf
    gn = uw.ompl_validate_motion_pairs(a[:8], b_nan, extent, res)
    assert not gn[:3].any() and np.array_equal(gn[3:], got[3:8])
    with pytest.raises(RuntimeError):
        uw.ompl_validate_motion_pairs(a[:8], b[:8], (extent[1], extent[0]), res)   # hi < lo


@pytest.mark.gpu
@pytest.mark.parametrize("npoint", [96, 257, 4096, 8192])
def test_clearance_stats_match_oracle_and_numpy(npoint, uw_wt2, ow_wt2, oracle_mod):
    """N3: calc_touch + multi_distance + [median, max, min, mean, std] per candidate (touchq_util.py:8-27,
    preprocess_key.py:275-281) reduced on the device, against the oracle's segment walk + numpy statistics.
    Order statistics are exact; mean / std to 1e-13 relative (summation order); even, odd and power-of-two counts."""
    import os
    tv = np.load(os.path.join(W.assets_dir(), "alpha-1.2.org.tunnel.npz"))["TUNNEL_V"]
    free = np.concatenate([tv[:, :3] * uw_wt2.scene_scale, tv[:, 3:]], axis=1)
    m = 5 if npoint < 4096 else 2
    mo = m if npoint < 4096 else 1          # the oracle walks a bounded prefix of the big cases
    q0 = np.ascontiguousarray(free[3:3 + m])
    fr, to = W.touch_rays(np.repeat(q0, npoint, axis=0), m * npoint, seed=71 + npoint)
    assert np.array_equal(fr.reshape(m, npoint, 7)[:, 0], q0)
    stats, dist = uw_wt2.clearance_stats(q0, to.reshape(m, npoint, 7), VD_TOUCH, with_distances=True)
    # the same reduction from the product's own segment outputs: exact
    lf = uw_wt2.transit_states_to_with_contact(fr, to, VD_TOUCH)[0]
    olf = ow_wt2.transit_states_to_with_contact(fr[:mo * npoint], to[:mo * npoint], VD_TOUCH)[0]
    assert _close(lf[:mo * npoint], olf)
    for c in range(m):
        d = np.array([oracle_mod.distance(q0[c], s) for s in lf[c * npoint:(c + 1) * npoint]])
        assert np.abs(dist[c] - d).max() < 1e-13
        want = np.array([np.median(dist[c]), dist[c].max(), dist[c].min(), dist[c].mean(), dist[c].std()])
        assert np.array_equal(stats[c, :3], want[:3])
        assert np.abs(stats[c, 3:] - want[3:]).max() < 1e-13 * max(1.0, want[3])
        if c < mo:
            od = np.array([oracle_mod.distance(q0[c], s) for s in olf[c * npoint:(c + 1) * npoint]])
            ow = np.array([np.median(od), od.max(), od.min(), od.mean(), od.std()])
            assert np.abs(stats[c] - ow).max() < 1e-12
    assert stats[:, 2].min() >= 0.0 and (stats[:, 1] > stats[:, 2]).all()
    with pytest.raises(Exception):
        uw_wt2.clearance_stats(q0, np.zeros((m, 8193, 7)), VD_TOUCH)


@pytest.mark.gpu
@pytest.mark.parametrize("which,vd,n", [("wt2", 0.0125 / 8, 4096), ("wt2", 1.5164e-5, 512), ("org", 1.5164e-5, 512),
                                         ("org", 3e-4, 2048)])
def test_swept_windows_leave_every_output_unchanged(which, vd, n, monkeypatch):
    """K2 advances over windows of steps that one grown-box / triangle-gap traversal proves free
    (swept windows).  That is a pruning of work, not of semantics: against the same walk with every
    step traversed (OSR_SWEEP_TILES=0) all six outputs -- first-contact step included -- are bit-identical, on
    tunnel rays (contact after a long approach) and on k-NN edges between free milestones (mostly free)."""
    import os
    from puzzletunneldiscovery_b200 import pyosr
    centre = W.ROB_OMPL_CENTER if which == "wt2" else None
    monkeypatch.setenv("OSR_SWEEP_TILES", "0")
    plain = make_world(pyosr.UnitWorld, which, centre)
    monkeypatch.delenv("OSR_SWEEP_TILES")
    swept = make_world(pyosr.UnitWorld, which, centre)
    if which == "wt2":
        tv = np.load(os.path.join(W.assets_dir(), "alpha-1.2.org.tunnel.npz"))["TUNNEL_V"]
        fr, to = W.touch_rays(swept.translate_to_unit_states(tv), n, seed=91)
    else:
        q = W.random_states(8 * n, seed=92)
        ms = q[swept.is_valid_states(q)][: n // 4]
        fr, to = swept.knn_edges(ms, k=4, window=256)
    a = swept.transit_states_to_with_contact(fr, to, vd, with_steps=True)
    swept_boxes = swept.counters(reset=True)["bv_tests"]
    b = plain.transit_states_to_with_contact(fr, to, vd, with_steps=True)
    plain_boxes = plain.counters(reset=True)["bv_tests"]
    for x, y in zip(a, b):
        assert np.array_equal(x, y)
    assert 0 < a[2].sum() < len(fr) or which == "wt2"
    assert np.array_equal(swept.is_valid_transitions(fr, to, vd), plain.is_valid_transitions(fr, to, vd))
    if vd < 1e-4:
        assert swept_boxes * 8 < plain_boxes        # and it is where the work goes away


@pytest.mark.gpu
@pytest.mark.parametrize("which,vd,resolve,rmax", [("wt2", 0.0125 / 8, "4096", "1"), ("wt2", 1.5164e-5, "64", "1"),
                                                   ("wt2", 2e-4, "100000", "1"), ("org", 1.5164e-5, "2048", "1e-3"),
                                                   ("org", 0.0125 / 8, "512", "1")])
def test_windows_that_keep_pairs_leave_every_output_unchanged(which, vd, resolve, rmax, monkeypatch):
    """A swept window may KEEP the triangle pairs it cannot clear and have its steps decided by the exact predicate
    on exactly those pairs (windows that keep pairs).  With the gates forced wide open (every window
    keeps pairs, also wide ones resolved in several 32-step chunks, also at the coarse verify_delta where a window
    keeps a hundred pairs) the six outputs equal those of the walk that splits such windows (OSR_SWEEP_RESOLVE=0)
    and of the walk that traverses every step, bit for bit."""
    import os
    from puzzletunneldiscovery_b200 import pyosr
    centre = W.ROB_OMPL_CENTER if which == "wt2" else None
    monkeypatch.setenv("OSR_SWEEP_TILES", "0")
    plain = make_world(pyosr.UnitWorld, which, centre)
    monkeypatch.delenv("OSR_SWEEP_TILES")
    monkeypatch.setenv("OSR_SWEEP_RESOLVE", "0")
    split = make_world(pyosr.UnitWorld, which, centre)
    monkeypatch.setenv("OSR_SWEEP_RESOLVE", resolve)
    monkeypatch.setenv("OSR_SWEEP_KEEP_RMAX", rmax)
    keep = make_world(pyosr.UnitWorld, which, centre)
    if which == "wt2":
        tv = np.load(os.path.join(W.assets_dir(), "alpha-1.2.org.tunnel-more.npz"))["TUNNEL_V"]
        fr, to = W.touch_rays(keep.translate_to_unit_states(tv), 1024 if vd < 1e-4 else 4096, seed=95)
    else:
        q = W.random_states(1 << 14, seed=96)
        ok = q[keep.is_valid_states(q) & ~keep.is_disentangled_states(q)][:256]
        fr, to = W.touch_rays(ok, 1024 if vd < 1e-4 else 4096, seed=97)
    a = keep.transit_states_to_with_contact(fr, to, vd, with_steps=True)
    kept_leaf = keep.counters(reset=True)["leaf_tests"]
    b = split.transit_states_to_with_contact(fr, to, vd, with_steps=True)
    split_leaf = split.counters(reset=True)["leaf_tests"]
    c = plain.transit_states_to_with_contact(fr, to, vd, with_steps=True)
    for x, y, z in zip(a, b, c):
        assert np.array_equal(x, y) and np.array_equal(x, z)
    assert (~a[2]).any() and kept_leaf != split_leaf     # walks end in contacts, and the listed path did run


@pytest.mark.gpu
def test_swept_windows_on_deep_trees(monkeypatch):
    """Same equality on a subdivided alpha puzzle (64 512 x 36 288 triangles: trees too large for shared memory,
    deeper work stacks): swept windows on = every step traversed, outputs bit-identical."""
    import os
    from puzzletunneldiscovery_b200 import pyosr
    A = W.assets_dir()
    eV, eF = W.load_obj_vf(os.path.join(A, "alpha_env-1.2.org.obj"))
    rV, rF = W.load_obj_vf(os.path.join(A, "alpha-1.2.org.obj"))
    EV, EF = W.subdivide(eV, eF, 8)
    RV, RF = W.subdivide(rV, rF, 6)

    def mk():
        w = pyosr.UnitWorld()
        w.set_meshes(EV, EF, RV, RF)
        w.scaleToUnit()
        w.angleModel(0.0, 0.0)
        return w
    monkeypatch.setenv("OSR_SWEEP_TILES", "0")
    plain = mk()
    monkeypatch.delenv("OSR_SWEEP_TILES")
    swept = mk()
    q = W.random_states(4096, seed=93)
    ms = q[swept.is_valid_states(q)][:256]
    fr, to = swept.knn_edges(ms, k=4, window=256)
    for vd in (1.5164e-5, 2e-3):
        a = swept.transit_states_to_with_contact(fr, to, vd, with_steps=True)
        b = plain.transit_states_to_with_contact(fr, to, vd, with_steps=True)
        assert all(np.array_equal(x, y) for x, y in zip(a, b))
    fr2, to2 = W.touch_rays(ms[:32], 512, seed=94)
    a = swept.transit_states_to_with_contact(fr2, to2, 1e-4, with_steps=True)
    b = plain.transit_states_to_with_contact(fr2, to2, 1e-4, with_steps=True)
    assert all(np.array_equal(x, y) for x, y in zip(a, b)) and 0 < a[2].sum() < len(fr2)


@pytest.mark.gpu
def test_pinned_host_arrays_give_the_same_labels(uw_org):
    """pyosr.pinned_empty (osr_host_alloc): page-locked sample / label arrays; same labels, in-place result."""
    from puzzletunneldiscovery_b200 import pyosr
    q = W.random_states(300001, seed=95)
    qp = pyosr.pinned_empty((len(q), 7))
    qp[:] = q
    out = pyosr.pinned_empty(len(q), np.uint8)
    out[:] = 7
    v = uw_org.is_valid_states(qp, out=out)
    assert v.dtype == np.bool_ and np.array_equal(v, uw_org.is_valid_states(q)) and out.max() <= 1
    assert np.shares_memory(v, out)
    with pytest.raises(Exception):
        uw_org.is_valid_states(q, out=np.empty(5, np.uint8))
    del qp, out, v      # the allocation is released with the last view


@pytest.mark.gpu
def test_numa_local_host_blocks_give_the_same_labels(uw_org, monkeypatch):
    """osr_host_alloc puts a block on the NUMA node of the current GPU (anonymous memory first touched there, then
    cudaHostRegister) when sysfs names the node; this pool's guests hide it (numa_node = -1, one node), so the path
    is forced with OSR_NUMA_NODE=0: same labels as pageable arrays, blocks can be freed and made again."""
    from puzzletunneldiscovery_b200 import pyosr
    monkeypatch.setenv("OSR_NUMA_NODE", "0")
    q = W.random_states(200003, seed=96)
    want = uw_org.is_valid_states(q)
    for _ in range(3):
        qp = pyosr.pinned_empty((len(q), 7))
        line = [l for l in open("/proc/self/numa_maps") if int(l.split()[0], 16) == qp.ctypes.data]
        assert line and "file=" not in line[0]      # the registered anonymous mapping, not cudaHostAlloc's
        qp[:] = q
        out = pyosr.pinned_empty(len(q), np.uint8)
        assert np.array_equal(uw_org.is_valid_states(qp, out=out), want)
        del qp, out


@pytest.mark.gpu
@pytest.mark.parametrize("n", [65536, 65537, 69631, 73729, 131071, 200003, 1048576 + 4097])
def test_host_call_chunk_schedule_ragged_sizes(uw_org, n):
    """The host validity call uploads in chunks (64 Ki states, then half of what is left, down to 8 Ki) under ONE
    stream-fed kernel: whatever the batch size does to the schedule -- a single chunk, a one-state second chunk, a
    remainder folded into the last chunk -- page-locked and pageable float64 samples and float32 samples give the
    labels of one batch of small single-chunk calls."""
    from puzzletunneldiscovery_b200 import pyosr
    q = W.random_states(n, seed=98)
    want = np.concatenate([uw_org.is_valid_states(q[i:i + 32768]) for i in range(0, n, 32768)])
    assert np.array_equal(uw_org.is_valid_states(q), want)                        # pageable float64
    qp = pyosr.pinned_empty((n, 7)); qp[:] = q
    assert np.array_equal(uw_org.is_valid_states(qp), want)                       # page-locked float64
    assert np.array_equal(uw_org.is_valid_states(q.astype(np.float32)), want)     # float32 (the states ARE float32 values)
    del qp


def test_enum_2drot_free_configuration_closed_form_properties(uw_org):
    """enum2DRotationFreeConfiguration (unit_world.cc:1391-1523; validity is stubbed to 1 there, so the output is pure
    geometry): every state puts the robot surface point on the pushed-out environment point, the first altitude row
    aligns the normals, no vertex of the robot triangle ends up below the environment triangle's plane (:1443-1459), the
    azimuth step is 2 pi / altitude_divider (the reference's own arithmetic), and the row-indexed component pass of
    return_all=False returns altitude rows of the same table."""
    from puzzletunneldiscovery_b200.pyosr import _quat_rotate, _angular_distance
    Vr, Fr = uw_org._geometry(uw_org.GEO_ROB)
    Ve, Fe = uw_org._geometry(uw_org.GEO_ENV)
    rs = np.random.RandomState(5)
    for trial in range(4):
        rprim, eprim = int(rs.randint(len(Fr))), int(rs.randint(len(Fe)))
        rt, et = Vr[Fr[rprim]], Ve[Fe[eprim]]
        rp, ep = rt.mean(0), et.mean(0)
        rn = np.cross(rt[1] - rt[0], rt[2] - rt[0]); rn /= np.linalg.norm(rn)
        en = np.cross(et[1] - et[0], et[2] - et[0]); en /= np.linalg.norm(en)
        margin, az, al = 1e-3, 5, 6
        qs = uw_org.enum_2drot_free_configuration(rp, rn, rprim, ep, en, eprim, margin, az, al, return_all=True)
        assert qs.shape == (az * al, 7) and np.abs(np.linalg.norm(qs[:, 3:], axis=1) - 1).max() < 1e-12
        for i in range(al):
            for j in range(az):
                q = qs[i * az + j]
                p = _quat_rotate(q[3:], rp) + q[:3]
                depth = float(en @ (p - ep))
                assert depth >= margin - 1e-12 and np.abs((p - ep) - en * depth).max() < 1e-12   # on the normal ray
                tri = np.array([_quat_rotate(q[3:], v) + q[:3] for v in rt])
                assert ((tri - ep) @ en).min() >= -1e-12       # the robot triangle stays on the outer side of the plane
                if i == 0:
                    assert np.abs(_quat_rotate(q[3:], rn) + en).max() < 1e-9                       # normals opposed
                if j > 0:   # consecutive azimuth samples differ by a turn of 2 pi / altitude_divider about en
                    turn = _angular_distance(q[3:], qs[i * az + j - 1][3:])
                    assert abs(turn - 2 * np.pi / al) < 1e-9
        some = uw_org.enum_2drot_free_configuration(rp, rn, rprim, ep, en, eprim, margin, az, al, return_all=False)
        assert len(some) >= 1 and all(any(np.array_equal(r, qs[k]) for k in range(al)) for r in some)


def test_knn_validate_edges_equals_the_two_separate_calls(uw_org):
    """osr_knn_validate_edges (k-NN edges validated where they are, only the milestones uploaded) = osr_knn_windowed
    followed by osr_is_valid_transitions on the materialised (from, to) arrays: same indices, same labels; windows
    with fewer than k other states give index -1 and label 0."""
    q = W.random_states(6000, seed=203)
    ms = q[uw_org.is_valid_states(q)][:4096 + 256 + 3]      # a full window, a short one, and one of three states
    for (k, window, vd) in ((8, 4096, None), (4, 256, 2e-3)):
        if k == 8:
            sel = np.concatenate([ms[:4096], ms[4096:4096 + 3]])   # second window holds 3 states: 2 neighbours each
        else:
            sel = ms
        idx, ok = uw_org.knn_validate_edges(sel, k=k, window=window, verify_delta=vd)
        want_idx = uw_org.knn_windowed(sel, k, window)
        assert np.array_equal(idx, want_idx)
        src = np.repeat(np.arange(len(sel)), k); dst = idx.reshape(-1)
        have = dst >= 0
        lab = uw_org.is_valid_transitions(sel[src[have]], sel[dst[have]], uw_org.recommended_cres if vd is None else vd)
        assert np.array_equal(ok.reshape(-1)[have], lab) and not ok.reshape(-1)[~have].any()
        assert 0 < lab.sum() < len(lab)
        if k == 8:
            assert (~have).sum() == 3 * (8 - 2)
    i0, ok0 = uw_org.knn_validate_edges(np.zeros((0, 7)), k=4, window=256)
    assert i0.shape == (0, 4) and ok0.shape == (0, 4)


def test_small_host_calls_equal_the_batch(uw_org, uw_wt2, ow_org):
    """Host calls of up to 16 384 states take the small-call path (mapped buffers, one warp per state or per few
    states on the warp-shared frontier, a grid sized to the states); their
    labels are those of the same states inside one large batch and of the oracle; a single state through
    is_valid_state too.  Repeated calls reuse the mapped buffers: a later, shorter call must not see an earlier one."""
    q = W.random_states(40000, seed=204)
    for uw in (uw_org, uw_wt2):
        big = uw.is_valid_states(q)
        for n in (1024, 1, 31, 33, 384, 385, 1023, 1025, 7, 1777, 5000, 16384, 16385, 3):
            o = (n * 37) % 20000
            assert np.array_equal(uw.is_valid_states(q[o:o + n]), big[o:o + n])
        assert [bool(uw.is_valid_state(q[i])) for i in range(20)] == big[:20].tolist()
        dis = uw.is_disentangled_states(q)
        assert np.array_equal(uw.is_disentangled_states(q[100:1100]), dis[100:1100])
        assert [bool(uw.is_disentangled(q[i])) for i in range(20)] == dis[:20].tolist() and 0 < dis.sum() < len(dis)
    assert np.array_equal(uw_org.is_valid_states(q[:1024]), ow_org.is_valid_states(q[:1024]).astype(bool))


def test_batched_touch_stages_equal_the_reference_loop(uw_wt2, tmp_path):
    """stages.calc_touch / stages.sample_touch (one device call per batch) against the reference's loops restated
    literally -- touchq_util.calc_touch (:15-27) and the worker loop of preprocess_surface.sample_touch (:76-96),
    one scalar transit_state_to_with_contact per sample -- under the same seed: same rays, same rows."""
    import math
    import os
    from puzzletunneldiscovery_b200 import pyosr, sharding, stages
    tv = np.load(os.path.join(W.assets_dir(), "alpha-1.2.org.tunnel.npz"))["TUNNEL_V"]
    keys = uw_wt2.translate_to_unit_states(tv)

    def sample_one_touch(uw, q0, stepping):              # touchq_util.py:8-13 on numpy's global generator
        tr = stages.random_on_sphere(1.0)
        aa = stages.random_within_sphere(2 * math.pi)
        return uw.transit_state_to_with_contact(q0, pyosr.apply(q0, tr, aa), stepping)
    np.random.seed(21)
    rows = [sample_one_touch(uw_wt2, keys[3], VD_TOUCH) for _ in range(24)]
    np.random.seed(21)
    got = stages.calc_touch(uw_wt2, keys[3], 24, VD_TOUCH)
    for k in range(5):   # the scalar call returns its taus as C floats (std::tuple<..., float, float>)
        want = np.array([r[k] for r in rows])
        have = got[k] if k < 3 else got[k].astype(np.float32).astype(np.float64)
        assert np.array_equal(have, want)
    tind = sharding.get_task_chunk((len(keys), 3), 4, 1)
    np.random.seed(22)
    ref = [(ki,) + tuple(sample_one_touch(uw_wt2, keys[ki], 2e-4)[:3]) for ki, _ in tind]
    np.random.seed(22)
    d = stages.sample_touch(uw_wt2, keys, tind, stepping=2e-4, out_fn=str(tmp_path / "touchq_batch-1.npz"))
    assert np.array_equal(d["FROM_VI"], [r[0] for r in ref]) and np.array_equal(d["FREE_V"], np.array([r[1] for r in ref]))
    assert np.array_equal(d["TOUCH_V"], np.array([r[2] for r in ref])) and np.array_equal(d["IS_INF"], [r[3] for r in ref])
    assert np.array_equal(np.load(str(tmp_path / "touchq_batch-1.npz"))["TOUCH_V"], d["TOUCH_V"])
    # keyconf.estimate_keyconf_clearance's worker loop (:300-315): per candidate calc_touch + multi_distance
    oskey = uw_wt2.translate_unit_to_ompl(keys)
    tk = [(2,), (5,), (7,)]
    np.random.seed(23)
    want = []
    for (k,) in tk:
        uq = uw_wt2.translate_ompl_to_unit(oskey[k:k + 1])[0]
        fv = stages.calc_touch(uw_wt2, uq, 16, uw_wt2.recommended_cres)[0]
        want.append(pyosr.multi_distance(uq, fv))
    np.random.seed(23)
    got = stages.estimate_keyconf_clearance(uw_wt2, oskey, tk, 16, out_fn=str(tmp_path / "clearance_batch-0.npz"))
    assert got.shape == (3, 16) and np.abs(got - np.array(want)).max() < 1e-13
    assert np.array_equal(np.load(str(tmp_path / "clearance_batch-0.npz"))["DISTANCE_BATCH"], got)


def test_small_segment_calls_equal_the_batch(uw_wt2, uw_org):
    """Segment calls of up to 1 024 segments take the small-call path (one mapped arena for inputs and outputs, a
    grid sized to the segments); all six outputs are those of the same segments inside one large batch -- scalar or
    per-segment verify_delta, transit and is_valid_transition semantics, repeated calls on the shared arena."""
    import os
    tv = np.load(os.path.join(W.assets_dir(), "alpha-1.2.org.tunnel.npz"))["TUNNEL_V"]
    fr, to = W.touch_rays(uw_wt2.translate_to_unit_states(tv), 3000, seed=205)
    big = uw_wt2.transit_states_to_with_contact(fr, to, VD_TOUCH, with_steps=True)
    vds = np.where(np.arange(3000) % 2 == 0, VD_TOUCH, 2e-4)
    bigv = uw_wt2.transit_states_to_with_contact(fr, to, vds, with_steps=True)
    lab = uw_wt2.is_valid_transitions(fr, to, 1e-3)
    for n in (1024, 1, 5, 64, 1023, 1025, 2):
        o = (n * 13) % 1500
        sl = slice(o, o + n)
        a = uw_wt2.transit_states_to_with_contact(fr[sl], to[sl], VD_TOUCH, with_steps=True)
        assert all(np.array_equal(x, y[sl]) for x, y in zip(a, big))
        b = uw_wt2.transit_states_to_with_contact(fr[sl], to[sl], vds[sl], with_steps=True)
        assert all(np.array_equal(x, y[sl]) for x, y in zip(b, bigv))
        assert np.array_equal(uw_wt2.is_valid_transitions(fr[sl], to[sl], 1e-3), lab[sl])
    one = uw_wt2.transit_state_to_with_contact(fr[7], to[7], VD_TOUCH)
    assert np.array_equal(one[0], big[0][7]) and np.array_equal(one[1], big[1][7]) and bool(one[2]) == bool(big[2][7])


@pytest.mark.gpu
def test_clearance_stats_edge_cases(uw_org, ow_org, oracle_mod):
    """One and two points per candidate (median = the value / the mean of both), targets closer than one step
    (the walk does not move: distance 0), an empty batch, and bad arguments raising like every other call."""
    q = W.random_states(4096, seed=96)
    cq = np.ascontiguousarray(q[uw_org.is_valid_states(q)][:3])
    for npoint in (1, 2):
        to = W.touch_rays(np.repeat(cq, npoint, axis=0), 3 * npoint, seed=97)[1].reshape(3, npoint, 7)
        st, d = uw_org.clearance_stats(cq, to, VD_TOUCH, with_distances=True)
        lf = ow_org.transit_states_to_with_contact(np.repeat(cq, npoint, axis=0), to.reshape(-1, 7), VD_TOUCH)[0]
        od = np.array([oracle_mod.distance(cq[i // npoint], s) for i, s in enumerate(lf)]).reshape(3, npoint)
        assert np.abs(d - od).max() < 1e-13
        assert np.array_equal(st[:, 0], np.median(d, axis=1)) and np.array_equal(st[:, 1], d.max(1))
        assert np.array_equal(st[:, 2], d.min(1)) and np.abs(st[:, 3] - d.mean(1)).max() < 1e-15
        assert np.abs(st[:, 4] - d.std(1)).max() < 1e-15
    near = np.repeat(cq[:, None, :], 4, axis=1).copy()
    near[:, :, 0] += 1e-6                       # closer than verify_delta: transitStateTo returns `from`
    st = uw_org.clearance_stats(cq, near, VD_TOUCH)
    assert np.array_equal(st, np.zeros((3, 5)))
    assert uw_org.clearance_stats(np.zeros((0, 7)), np.zeros((0, 4, 7)), VD_TOUCH).shape == (0, 5)
    with pytest.raises(Exception):
        uw_org.clearance_stats(cq, near, 0.0)
    with pytest.raises(Exception):
        uw_org.clearance_stats(cq, near[:2], VD_TOUCH)


@pytest.mark.timeout(120)
def test_transit_state_by_pure_translation(uw_org, ow_org):
    """transit_state_by with a zero angle-axis (what aapath-walker / rlsampler pass for translation actions):
    Eigen's normalized() leaves the zero vector alone and AngleAxis(0, 0) is the identity rotation
    (unit_world.cc:489-506), so the target is the translated state -- not a NaN state."""
    q = W.random_states(4096, seed=71)
    free = q[uw_org.is_valid_states(q)][:16]
    for s in free[:6]:
        tr = np.array([0.03, -0.02, 0.01])
        got = uw_org.transit_state_by(s, tr, np.zeros(3), VD_TOUCH)
        to = np.concatenate([s[:3] + tr, s[3:7]])
        want = uw_org.transit_state_to(s, to, VD_TOUCH)
        assert np.isfinite(got[0]).all() and np.array_equal(got[0], want[0]) and got[1:] == want[1:]
        olf, ofh, oisf, oft, ott = ow_org.transit_states_to_with_contact(s[None], to[None], VD_TOUCH)
        assert np.array_equal(got[0], olf[0]) and got[1] == bool(oisf[0])
        # a rotation action still rotates
        rot = uw_org.transit_state_by(s, np.zeros(3), np.array([0.0, 0.0, 0.05]), VD_TOUCH)
        assert np.isfinite(rot[0]).all()


@pytest.mark.timeout(120)
def test_degenerate_walks_return_instead_of_hanging(uw_org, ow_org):
    """NaN states / NaN verify_delta fall through both comparisons of the reference's walk (unit_world.cc:443-447):
    its loop body never runs and it returns (from, from, false, 0, 0) -- the kernel must do the same instead of
    spinning in its window loop.  verify_delta <= 0 on a segment of non-zero length never ends in
    the reference; here it is an argument error.  Neighbouring segments of the same batch are unaffected."""
    q = W.random_states(4096, seed=72)
    free = q[uw_org.is_valid_states(q)][:64]
    fr, to = W.touch_rays(free, 64, seed=73)
    base = uw_org.transit_states_to_with_contact(fr, to, VD_TOUCH, with_steps=True)
    fr2, to2 = fr.copy(), to.copy()
    to2[0, 3:7] = np.nan          # what a 0/0 axis used to produce
    fr2[1, 0] = np.nan
    to2[2, :] = np.nan
    to2[3, 1] = np.inf            # infinite distance: the first step is a NaN pose, reported as the first hit
    got = uw_org.transit_states_to_with_contact(fr2, to2, VD_TOUCH, with_steps=True)
    want = ow_org.transit_states_to_with_contact(fr2[:3], to2[:3], VD_TOUCH)
    for i in range(3):
        assert np.array_equal(got[0][i], fr2[i], equal_nan=True) and np.array_equal(got[1][i], fr2[i], equal_nan=True)
        assert not got[2][i] and got[3][i] == 0.0 and got[4][i] == 0.0 and got[5][i] == 0
        assert np.array_equal(want[0][i], fr2[i], equal_nan=True) and np.array_equal(want[1][i], fr2[i], equal_nan=True)
        assert not want[2][i] and want[3][i] == 0.0 and want[4][i] == 0.0
    assert not got[2][3] and got[5][3] == 1
    for k in range(6):
        assert np.array_equal(np.asarray(got[k])[4:], np.asarray(base[k])[4:])
    # per-segment verify_delta with a NaN entry
    vd = np.full(64, VD_TOUCH); vd[5] = np.nan
    gv = uw_org.transit_states_to_with_contact(fr, to, vd, with_steps=True)
    assert np.array_equal(gv[0][5], fr[5]) and np.array_equal(gv[1][5], fr[5]) and not gv[2][5] and gv[5][5] == 0
    assert np.array_equal(gv[5][6:], base[5][6:])
    assert not uw_org.is_valid_transitions(fr2[:3], to2[:3], VD_TOUCH).any()
    # verify_delta <= 0
    for bad in (0.0, -1e-3):
        with pytest.raises(RuntimeError):
            uw_org.transit_states_to_with_contact(fr, to, bad)
        with pytest.raises(RuntimeError):
            uw_org.is_valid_transitions(fr, to, bad)
    # ... but 0 on zero-length segments is the reference's `verify_delta >= dist` early return
    z = uw_org.transit_states_to_with_contact(fr[:4], fr[:4], 0.0)
    assert not z[2].any() and (z[4] == 1.0).all() and np.array_equal(z[1], fr[:4])
    assert not uw_org.is_valid_transitions(fr[:4], fr[:4], VD_TOUCH).any()
    # the engine still works afterwards (error flag cleared)
    again = uw_org.transit_states_to_with_contact(fr, to, VD_TOUCH, with_steps=True)
    for k in range(6):
        assert np.array_equal(np.asarray(again[k]), np.asarray(base[k]))
    # discrete actions whose verify_delta underflows to 0 return instead of walking forever
    s, done, prog = uw_org.transit_state(free[0], 0, 0.01, 0.0)
    assert np.array_equal(s, free[0]) and not done and prog == 0.0
