#!/usr/bin/env python
"""Bounded runs of BASELINE.json configs 2-5 on one B200 (bench.py measures config 1, the headline).

Every leg goes through the C ABI (`libosrgpu.so`) with HOST buffers, is timed by wall clock around the call
(copies included), and a prefix of its output is compared with the CPU oracle (test infrastructure, used here
only as the checker and as the timed CPU baseline).  One JSON line per leg; keep the output under profiles/.

    python scripts/bench_configs.py [--legs validity_wt2,touch,touch_fine,edges,hipoly,contacts,clearance,keyconf,ompl_motions] [--scale 1.0]
"""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from puzzletunneldiscovery_b200 import pyosr, workloads as W  # noqa: E402

A = W.assets_dir()
VD_TOUCH = 0.0125 / 8   # /root/reference/src/GP/condor_touch_configuration.py:35,92
PREFIX_SCALE = 1        # --prefix-scale: multiplies the size of every oracle-checked prefix


def world(cls, which="org", centre=None, meshes=None, devices=None):
    w = cls(devices=devices) if devices is not None else cls()
    if meshes is not None:
        w.set_meshes(*meshes)
    else:
        envf, robf = W.REF_ALPHA_FILES[which]
        w.loadModelFromFile(os.path.join(A, envf))
        w.loadRobotFromFile(os.path.join(A, robf))
    if centre is not None:
        w.enforceRobotCenter(centre)
    w.scaleToUnit()
    w.angleModel(0.0, 0.0)
    return w


def timed(fn, reps=1):
    fn()  # warm-up (buffers, clocks)
    t = time.perf_counter()
    for _ in range(reps):
        r = fn()
    return r, (time.perf_counter() - t) / reps


def leg_touch(O, scale, fine=False):
    """Config 2: touch segments from the tunnel vertices on G-wt2, verify_delta 0.0125/8 (or the
    pipeline's recommended_cres on a small subset when fine=True)."""
    uw = world(pyosr.UnitWorld, "wt2", W.ROB_OMPL_CENTER)
    ow = world(O.OracleWorld, "wt2", W.ROB_OMPL_CENTER)
    tv = np.concatenate([np.load(os.path.join(A, "alpha-1.2.org.tunnel.npz"))["TUNNEL_V"],
                         np.load(os.path.join(A, "alpha-1.2.org.tunnel-more.npz"))["TUNNEL_V"]])
    fr0 = uw.translate_to_unit_states(tv)
    vd = 1.5164e-5 if fine else VD_TOUCH
    n = int((1 << 12 if fine else 1 << 18) * scale)
    fr, to = W.touch_rays(fr0, n, seed=1)
    (lf, fh, isf, ft, tt, st), dt = timed(lambda: uw.transit_states_to_with_contact(fr, to, vd, with_steps=True))
    m = min(n, (64 if fine else 2048) * PREFIX_SCALE)
    tc = time.perf_counter()
    olf, ofh, oisf, oft, ott, ocnt = ow.transit_states_to_with_contact(fr[:m], to[:m], vd, counters=True)
    dtc = time.perf_counter() - tc
    # first contact at the very same step; taus to 1e-12 (device libm vs glibc may differ in the last bit of dist)
    same = bool((isf[:m] == oisf.astype(bool)).all() and (st[:m] == ocnt[:, 0]).all() and
                np.abs(tt[:m] - ott).max() < 1e-12 and np.abs(ft[:m] - oft).max() < 1e-12 and
                np.abs(fh[:m] - ofh).max() < 1e-12)
    # size-independent properties on the WHOLE batch (the reference's own invariant, unit_world.cc:325-326, and the
    # structure of its return tuple): last_free is a valid state, first_hit of a blocked walk is not, a blocked walk
    # stops strictly after its last free tau, a free one ends at tau = 1 with the shadowed touch_tau = 0
    lf_ok = uw.is_valid_states(lf)
    blocked = ~isf
    walked = st > 0
    props = {"last_free_states_all_valid": bool(lf_ok.all()),
             "first_hit_states_of_blocked_walks_all_invalid": bool((~uw.is_valid_states(fh[blocked & walked])).all()),
             "blocked_walks_have_free_tau_lt_touch_tau": bool((ft[blocked & walked] < tt[blocked & walked]).all()),
             "free_walks_end_at_tau_1": bool((np.abs(ft[isf] - 1.0) <= 2.3e-16).all() and (tt[isf] == 0.0).all()),
             "label_checksum": int(np.dot(isf.astype(np.int64), np.arange(1, n + 1, dtype=np.int64) % 1000003)),
             "steps_checksum": int(st.sum())}
    half = n // 2     # the same batch evaluated as two halves gives the same rows (no dependence on batch layout)
    h1 = uw.transit_states_to_with_contact(fr[:half], to[:half], vd, with_steps=True)
    h2 = uw.transit_states_to_with_contact(fr[half:], to[half:], vd, with_steps=True)
    props["two_half_batches_equal_the_whole"] = bool(np.array_equal(np.concatenate([h1[5], h2[5]]), st) and
                                                      np.array_equal(np.concatenate([h1[0], h2[0]]), lf))
    return {"leg": "touch_fine" if fine else "touch", "geometry": "G-wt2", "segments": n, "verify_delta": vd,
            "properties_full_size": props,
            "segments_per_s": n / dt, "validity_checks_per_s": float(st.sum()) / dt,
            "mean_steps_per_segment": float(st.mean()), "free_fraction": float(isf.mean()), "seconds": dt,
            "oracle_prefix": m, "same_first_contact_step_as_oracle_on_prefix": same,
            "cpu_port": {"segments_per_s": m / dtc, "validity_checks_per_s": float(ocnt[:, 0].sum()) / dtc,
                         "threads": O.lib().orc_max_threads()}}


def leg_validity_wt2(O, scale):
    """Config 1 repeated on G-wt2 (the mesh set of the reference's alpha-1.2 touch scripts, 5 690 x 5 948
    triangles -- the trees no longer fit into shared memory): 2^20 seeded uniform states, host call with page-locked
    and pageable arrays, the kernel alone on device-resident states, oracle-checked prefix."""
    import ctypes as C
    import torch
    from puzzletunneldiscovery_b200._lib import check
    uw = world(pyosr.UnitWorld, "wt2", W.ROB_OMPL_CENTER)
    ow = world(O.OracleWorld, "wt2", W.ROB_OMPL_CENTER)
    n = int((1 << 20) * scale)
    q = W.random_states(n, seed=0)
    qp = pyosr.pinned_empty((n, 7)); qp[:] = q
    op = pyosr.pinned_empty(n, np.uint8)
    v, dt = timed(lambda: uw.is_valid_states(qp, out=op), reps=3)
    v = np.array(v, copy=True)
    _, dt_pageable = timed(lambda: uw.is_valid_states(q), reps=2)
    dq = torch.from_numpy(q).cuda(); do = torch.empty(n, dtype=torch.uint8, device="cuda")
    ms = []
    uw.counters(reset=True)
    for _ in range(4):
        check(uw._L.osr_is_valid_states_device(uw._h, C.c_void_p(dq.data_ptr()), n, C.c_void_p(do.data_ptr())))
        check(uw._L.osr_sync(uw._h))
        f = C.c_float(0); check(uw._L.osr_last_kernel_ms(uw._h, C.byref(f))); ms.append(f.value)
    c = uw.counters(reset=True)
    m = min(n, 65536 * PREFIX_SCALE)
    tc = time.perf_counter()
    ov = ow.is_valid_states(q[:m], threads=0)
    dtc = time.perf_counter() - tc
    return {"leg": "validity_wt2", "geometry": "G-wt2", "states": n, "configs_per_s": n / dt, "seconds": dt,
            "configs_per_s_pageable_host_arrays": n / dt_pageable,
            "configs_per_s_device_resident": n / (min(ms[1:]) * 1e-3), "kernel_ms": min(ms[1:]),
            "valid_fraction": float(v.mean()), "box_tests_per_state": c["bv_tests"] / (4.0 * n),
            "triangle_tests_per_state": c["leaf_tests"] / (4.0 * n), "fp64_rechecks_per_state": c["fp64_rechecks"] / (4.0 * n),
            "device_labels_equal_host_call": bool(np.array_equal(do.cpu().numpy().astype(bool), v)),
            "oracle_prefix": m, "labels_equal_oracle_on_prefix": bool(np.array_equal(v[:m], ov.astype(bool))),
            "cpu_port": {"configs_per_s": m / dtc, "threads": O.lib().orc_max_threads()}}


def leg_edges(O, scale):
    """Config 3: is_valid_transition on windowed k-NN edges among valid milestones, G-org."""
    uw = world(pyosr.UnitWorld)
    ow = world(O.OracleWorld)
    nm = int((1 << 16) * scale)
    q = W.random_states(2 * nm, seed=2)
    q = q[uw.is_valid_states(q)][:nm]
    uw.knn_edges(q[:4096], k=8, window=4096)
    t = time.perf_counter()
    fr, to = uw.knn_edges(q, k=8, window=4096)      # K6: windowed k-NN under osr::distance on the GPU
    t_knn = time.perf_counter() - t
    hfr, hto = W.knn_edges(q[:4096], k=8, window=4096)   # host restatement on the first window
    same_edges = {(a.tobytes(), b.tobytes()) for a, b in zip(hfr, hto)} == \
        {(a.tobytes(), b.tobytes()) for a, b in zip(fr[:8 * 4096], to[:8 * 4096])}
    vd = 1.51641e-5
    uw.counters(reset=True)     # the timed call runs twice (warm-up + measurement): counters below are halved
    res, dt = timed(lambda: uw.is_valid_transitions(fr, to, vd))
    m = min(len(fr), 256 * PREFIX_SCALE)
    tc = time.perf_counter()
    ores, ocnt = ow.is_valid_transitions(fr[:m], to[:m], vd, counters=True)
    dtc = time.perf_counter() - tc
    c = uw.counters(reset=True)
    # the same workload without building the edges: k-NN and validation in one call, only the milestones uploaded
    (fidx, fok), dt_fused = timed(lambda: uw.knn_validate_edges(q, k=8, window=4096, verify_delta=vd))
    fdst = fidx.reshape(-1)
    fused_same = bool(np.array_equal(fok.reshape(-1)[fdst >= 0], res) and np.array_equal(q[fdst[fdst >= 0]], to))
    # full-size properties: a random subset re-evaluated on its own gives the same labels; an edge reversed end for
    # end is walked at different taus but between the same two valid milestones, so the two directions may only
    # disagree when a contact lies inside one verify_delta of the path (reported, not asserted)
    rs = np.random.RandomState(11)
    sub = rs.choice(len(fr), min(len(fr), 1 << 16), replace=False)
    again = uw.is_valid_transitions(fr[sub], to[sub], vd)
    rev = uw.is_valid_transitions(to[sub], fr[sub], vd)
    props = {"subset_relabelled_alone_equal": bool(np.array_equal(again, res[sub])),
             "reversed_edges_disagree_fraction": float((rev != res[sub]).mean()),
             "label_checksum": int(np.dot(res.astype(np.int64), np.arange(1, len(res) + 1, dtype=np.int64) % 1000003))}
    return {"leg": "edges", "geometry": "G-org", "milestones": len(q), "edges": len(fr), "verify_delta": vd,
            "properties_full_size": props,
            "edges_per_s": len(fr) / dt, "seconds": dt, "valid_edge_fraction": float(res.mean()),
            "gpu_knn_seconds": t_knn, "knn_edges_per_s": len(fr) / t_knn,
            "fused_knn_and_validation": {"seconds": dt_fused, "edges_per_s": len(fr) / dt_fused,
                                         "same_edges_and_labels": fused_same,
                                         "separate_calls_seconds (k-NN + edge arrays + validation)": t_knn + dt,
                                         "h2d_bytes": int(q.nbytes), "h2d_bytes_separate_calls": int(q.nbytes + fr.nbytes + to.nbytes)},
            "knn_first_window_equals_host_restatement": bool(same_edges), "oracle_prefix": m,
            "labels_equal_oracle_on_prefix": bool(np.array_equal(res[:m], ores.astype(bool))),
            "steps_reference_semantics": int(c["poses"]) // 2, "box_tests_per_step": c["bv_tests"] / max(1, c["poses"]),
            "validity_checks_per_s": c["poses"] / dt / 2,
            "cpu_port": {"edges_per_s": m / dtc, "validity_checks_per_s": float(ocnt[:, 0].sum()) / dtc,
                         "threads": O.lib().orc_max_threads()}}


def leg_hipoly(O, scale, device_counts=(1,)):
    """Config 4: validity on the subdivided alpha puzzle (robot 100 800 x env 487 872 triangles), one
    handle driving 1 / 2 / 4 / 8 GPUs of the box (strong scaling: the same batch, geometry replicated, index ranges
    per device, host gather)."""
    import torch
    eV, eF = W.load_obj_vf(os.path.join(A, "alpha_env-1.2.org.obj"))
    rV, rF = W.load_obj_vf(os.path.join(A, "alpha-1.2.org.obj"))
    EV, EF = W.subdivide(eV, eF, 22)
    RV, RF = W.subdivide(rV, rF, 10)
    n = int((1 << 22) * scale)
    q = W.random_states(n, seed=3)
    qp = pyosr.pinned_empty((n, 7)); qp[:] = q          # page-locked sample / label arrays (osr_host_alloc)
    op = pyosr.pinned_empty(n, np.uint8)
    navail = torch.cuda.device_count()
    # the robot frame's origin is the mean of the mesh vertices (Scene::mean_of_vertices_), which subdivision moves:
    # G-hi is pinned to the centre of the original mesh so that the same states mean the same poses on both
    u0 = world(pyosr.UnitWorld)
    centre = u0.robot_center
    scaling = []
    v = None
    for nd in device_counts:
        if nd > navail:
            continue
        t = time.perf_counter()
        uw = world(pyosr.UnitWorld, centre=centre, meshes=(EV, EF, RV, RF), devices=list(range(nd)))
        t_build = time.perf_counter() - t
        vv, dt = timed(lambda: uw.is_valid_states(qp, out=op), reps=2)
        row = {"gpus": nd, "configs_per_s": n / dt, "seconds": dt, "finalize_seconds (host BVH build + upload)": t_build}
        if nd == 1:
            _, dt_pageable = timed(lambda: uw.is_valid_states(q))
            row["configs_per_s_pageable_host_arrays"] = n / dt_pageable
            c = uw.counters(reset=True)
            v = np.array(vv, copy=True)
        else:
            row["labels_equal_one_gpu"] = bool(np.array_equal(np.asarray(vv), v))
        scaling.append(row)
        del uw
    # the subdivided meshes describe the same surfaces: labels on G-hi equal the labels of the same states on the
    # original 1 008-triangle meshes except where the rounding of the new vertices decides (counted)
    v_org = u0.is_valid_states(qp)
    ow = world(O.OracleWorld, centre=centre, meshes=(EV, EF, RV, RF))
    m = 1 << 16
    tc = time.perf_counter()
    vo = ow.is_valid_states(q[:m])
    dtc = time.perf_counter() - tc
    return {"leg": "hipoly", "geometry": "G-hi (env %d, robot %d triangles)" % (len(EF), len(RF)), "states": n,
            "configs_per_s": scaling[0]["configs_per_s"], "seconds": scaling[0]["seconds"],
            "strong_scaling_one_handle": scaling, "valid_fraction": float(v.mean()),
            "labels_differing_from_G-org_same_states": int((v != v_org).sum()), "oracle_prefix": m,
            "labels_equal_oracle_on_prefix": bool(np.array_equal(v[:m], vo.astype(bool))),
            "cpu_port": {"configs_per_s": m / dtc, "threads": O.lib().orc_max_threads()}}


def leg_contacts(O, scale):
    """Config 5: first-hit states of the touch search -> every touching triangle pair, its Moeller
    segment, coplanar flag and the barycentric coordinates of the segment midpoint on both triangles."""
    uw = world(pyosr.UnitWorld, "wt2", W.ROB_OMPL_CENTER)
    ow = world(O.OracleWorld, "wt2", W.ROB_OMPL_CENTER)
    tv = np.load(os.path.join(A, "alpha-1.2.org.tunnel-more.npz"))["TUNNEL_V"]
    fr0 = uw.translate_to_unit_states(tv)
    ns = int((1 << 16) * scale)
    fr, to = W.touch_rays(fr0, ns, seed=4)
    lf, fh, isf, ft, tt = uw.transit_states_to_with_contact(fr, to, VD_TOUCH)
    hit = fh[~isf]
    cap = max(1 << 21, 24 * len(hit))
    r, dt = timed(lambda: uw.intersecting_segments_batch(hit, cap=cap, with_bary=True))
    m = min(len(hit), 1024)
    tc = time.perf_counter()
    off, segs, faces, cop, total = ow.intersecting_segments_batch(hit[:m])
    dtc = time.perf_counter() - tc
    k = int(off[-1])

    def canon(off_, faces_, seg_, cop_):  # the oracle lists a state's rows in FCL's DFS order, the engine sorted
        rows = [(i, int(faces_[j][0]), int(faces_[j][1])) + tuple(seg_[j]) + (int(cop_[j]),)
                for i in range(m) for j in range(off_[i], off_[i + 1])]
        return np.array(sorted(rows)).reshape(-1, 10)
    a = canon(r["offsets"], r["faces"], r["seg"], r["coplanar"])
    b = canon(off, faces, segs, cop)
    same_faces = bool(a.shape == b.shape and np.array_equal(a[:, :3], b[:, :3]) and np.array_equal(a[:, 9], b[:, 9]))
    err = float(np.abs(a[:, 3:9] - b[:, 3:9]).max()) if same_faces and k else None
    # full-size properties: every first-hit state has at least one row; a random subset of states re-evaluated alone
    # yields the same rows; both surface points of a row (barycentrics on each triangle) are the segment midpoint
    cnt = np.diff(r["offsets"])
    rs = np.random.RandomState(12)
    sub = np.sort(rs.choice(len(hit), min(len(hit), 4096), replace=False))
    r2 = uw.intersecting_segments_batch(hit[sub], cap=cap, with_bary=True)
    rows_sub = np.concatenate([np.arange(r["offsets"][i], r["offsets"][i + 1]) for i in sub]) if len(sub) else np.zeros(0, int)
    props = {"every_first_hit_state_has_a_contact_row": bool((cnt > 0).all()), "no_overflow": bool(r["total"] <= cap),
             "subset_states_alone_give_the_same_rows": bool(np.array_equal(np.diff(r2["offsets"]), cnt[sub]) and
                                                             np.array_equal(r2["faces"][:r2["total"]], r["faces"][rows_sub]) and
                                                             np.array_equal(r2["seg"][:r2["total"]], r["seg"][rows_sub]))}
    return {"leg": "contacts", "geometry": "G-wt2", "states": int(len(hit)), "contact_rows": int(r["total"]),
            "properties_full_size": props,
            "contacts_per_s": r["total"] / dt, "states_per_s": len(hit) / dt, "seconds": dt,
            "rows_per_state": r["total"] / max(1, len(hit)), "oracle_prefix_states": m,
            "rows_equal_oracle_on_prefix": same_faces, "max_abs_segment_error_on_prefix": err,
            "cpu_port": {"states_per_s": m / dtc, "contacts_per_s": k / dtc, "threads": 1}}


def leg_clearance(O, scale):
    """clearance statistics of key-configuration candidates on G-wt2 -- npoint touch segments per
    candidate at the pipeline's recommended_cres, reduced on the device to [median, max, min, mean, std]."""
    uw = world(pyosr.UnitWorld, "wt2", W.ROB_OMPL_CENTER)
    ow = world(O.OracleWorld, "wt2", W.ROB_OMPL_CENTER)
    tv = np.load(os.path.join(A, "alpha-1.2.org.tunnel-more.npz"))["TUNNEL_V"]
    m, npoint, vd = max(1, int(64 * scale)), 1024, 1.5164e-5
    q0 = np.ascontiguousarray(uw.translate_to_unit_states(tv)[:m])
    m = len(q0)
    fr, to = W.touch_rays(np.repeat(q0, npoint, axis=0), m * npoint, seed=5)
    to = to.reshape(m, npoint, 7)
    (stats, dist), dt = timed(lambda: uw.clearance_stats(q0, to, vd, with_distances=True))
    _, dt_stats_only = timed(lambda: uw.clearance_stats(q0, to, vd))
    k = 64
    tc = time.perf_counter()
    olf = ow.transit_states_to_with_contact(fr[:k], to[0, :k], vd)[0]
    od = np.array([O.distance(q0[0], s) for s in olf])
    dtc = time.perf_counter() - tc
    want = np.stack([np.median(dist, 1), dist.max(1), dist.min(1), dist.mean(1), dist.std(1)], axis=1)
    return {"leg": "clearance", "geometry": "G-wt2", "candidates": m, "points_per_candidate": npoint,
            "verify_delta": vd, "candidates_per_s": m / dt_stats_only, "segments_per_s": m * npoint / dt_stats_only,
            "seconds": dt_stats_only, "seconds_with_distances": dt,
            "d2h_bytes": 40 * m, "d2h_bytes_reference_layout": 15 * 8 * m * npoint,
            "max_abs_stat_error_vs_numpy": float(np.abs(stats - want).max()),
            "oracle_prefix_segments": k, "max_abs_distance_error_vs_oracle_on_prefix": float(np.abs(dist[0, :k] - od).max()),
            "cpu_port": {"segments_per_s": k / dtc, "threads": O.lib().orc_max_threads()}}


def leg_keyconf(O, scale):
    """key-configuration candidates -- SurfacePairsToSample=1024 surface-point pairs x
    NumberOfRotations=256 spins (src/GP/pipeline/keyconf.py:117-131), built + labelled + run-filtered on the device."""
    uw = world(pyosr.UnitWorld)
    ow = world(O.OracleWorld)
    rs = np.random.RandomState(6)

    def sample(which, n):
        V, F = uw._geometry(which)
        f = F[rs.randint(0, len(F), n)]
        a, b, c = V[f[:, 0]], V[f[:, 1]], V[f[:, 2]]
        nrm = np.cross(b - a, c - a)
        return (a + b + c) / 3.0, nrm / np.linalg.norm(nrm, axis=1, keepdims=True)

    b, den = max(1, int(1024 * scale)), 256
    rp, rn = sample(1, b)
    ep, en = sample(0, b)
    pts = np.stack([rp, rn, ep, en], axis=1)
    (q, valid, counts), dt = timed(lambda: uw.enum_free_configurations(pts, 1e-4, den, only_median=True))
    q_all, v_all, _ = uw.enum_free_configurations(pts, 1e-4, den)
    m = min(b, 64)
    tc = time.perf_counter()
    qh = uw.enum_configuration_states(pts[:m], 1e-4, den)
    vo = ow.is_valid_states(qh.reshape(-1, 7)).astype(bool).reshape(m, den)
    dtc = time.perf_counter() - tc
    return {"leg": "keyconf", "geometry": "G-org", "surface_pairs": b, "rotations": den,
            "candidates_per_s": b * den / dt, "pairs_per_s": b / dt, "seconds": dt,
            "valid_fraction": float(v_all.mean()), "kept_after_run_filter": int(counts.sum()),
            "h2d_bytes": 96 * b + 16 * den, "oracle_prefix_pairs": m,
            "states_equal_host_construction_on_prefix": bool(np.array_equal(q_all[:m], qh)),
            "labels_equal_oracle_on_prefix": bool(np.array_equal(v_all[:m], vo)),
            "cpu_port": {"candidates_per_s": m * den / dtc, "threads": O.lib().orc_max_threads()}}


def leg_ompl(O, scale):
    """OmplDriver.validate_motion_pairs semantics (DiscreteMotionValidator::checkMotion, OMPL's own
    interpolation and segment count) on the windowed k-NN edges of valid milestones, OMPL-frame rows, at the
    pipeline's collision_resolution fallback 1e-4 (src/GP/pipeline/solve.py:265)."""
    uw = world(pyosr.UnitWorld)
    ow = world(O.OracleWorld)
    nm = int((1 << 14) * scale)
    q = W.random_states(2 * nm, seed=7)
    q = q[uw.is_valid_states(q)][:nm]
    fr, to = uw.knn_edges(q, k=8, window=4096)
    a, b = uw.translate_unit_to_ompl(fr), uw.translate_unit_to_ompl(to)
    # planner bounds (OmplDriver.set_bounds): a cube around the milestones whose diagonal is 150 OMPL units
    mid = 0.5 * (np.minimum(a[:, :3].min(axis=0), b[:, :3].min(axis=0)) + np.maximum(a[:, :3].max(axis=0), b[:, :3].max(axis=0)))
    half = 0.5 * 150.0 / np.sqrt(3.0)
    extent, res = (mid - half, mid + half), 1e-4
    assert (a[:, :3] > extent[0]).all() and (b[:, :3] < extent[1]).all()
    uw.counters(reset=True)
    (ok, nd), dt = timed(lambda: uw.ompl_validate_motion_pairs(a, b, extent, res, with_counts=True))
    c = uw.counters(reset=True)
    m = min(len(a), 64 * PREFIX_SCALE)
    dc = (uw.get_ompl_center(1).astype(np.float32) - uw.robot_center.astype(np.float32)).astype(np.float64) * uw.scene_scale
    tc = time.perf_counter()
    want, ond = ow.ompl_check_motions(a[:m], b[:m], uw.scene_scale, dc, extent, res)
    dtc = time.perf_counter() - tc
    return {"leg": "ompl_motions", "geometry": "G-org", "motions": len(a), "resolution": res,
            "motions_per_s": len(a) / dt, "seconds": dt, "mean_points_per_motion": float(nd.mean()),
            "points_per_s_reference_semantics": float(nd.sum()) / dt, "valid_fraction": float(ok.mean()),
            "box_tests_per_point": c["bv_tests"] / 2 / max(1.0, float(nd.sum())), "oracle_prefix": m,
            "labels_and_counts_equal_oracle_on_prefix": bool(np.array_equal(ok[:m].astype(bool), want.astype(bool)) and
                                                             np.array_equal(nd[:m], ond)),
            "cpu_port": {"motions_per_s": m / dtc, "points_per_s": float(ond.sum()) / dtc, "threads": O.lib().orc_max_threads()}}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--legs", default="validity_wt2,touch,touch_fine,edges,hipoly,contacts,clearance,keyconf,ompl_motions")
    ap.add_argument("--scale", type=float, default=1.0)
    ap.add_argument("--prefix-scale", type=int, default=1)
    ap.add_argument("--full", action="store_true",
                    help="BASELINE.json's sizes: 2^20 touch segments (2^16 at recommended_cres), 4 Mi milestones / 3.3e7 "
                         "k-NN edges, 2^24 states on G-hi, >= 2^20 contact rows")
    ap.add_argument("--devices", default="1", help="device counts of the hipoly leg, e.g. 1,2,4,8")
    args = ap.parse_args()
    global PREFIX_SCALE
    PREFIX_SCALE = max(1, args.prefix_scale)
    from oracle import oracle as O
    mult = {"touch": 4, "touch_fine": 16, "edges": 64, "hipoly": 4, "contacts": 2} if args.full else {}
    sc = lambda leg: args.scale * mult.get(leg, 1)
    devs = tuple(int(x) for x in args.devices.split(","))
    legs = {"validity_wt2": lambda: leg_validity_wt2(O, args.scale), "touch": lambda: leg_touch(O, sc("touch")), "touch_fine": lambda: leg_touch(O, sc("touch_fine"), fine=True),
            "edges": lambda: leg_edges(O, sc("edges")), "hipoly": lambda: leg_hipoly(O, sc("hipoly"), devs),
            "contacts": lambda: leg_contacts(O, sc("contacts")), "clearance": lambda: leg_clearance(O, args.scale),
            "keyconf": lambda: leg_keyconf(O, args.scale), "ompl_motions": lambda: leg_ompl(O, args.scale)}
    for name in args.legs.split(","):
        t = time.perf_counter()
        try:
            out = legs[name]()
        except Exception as e:  # keep the other legs running; the failure is part of the record
            out = {"leg": name, "error": repr(e)}
        out["leg_wall_seconds"] = time.perf_counter() - t
        print(json.dumps(out), flush=True)


if __name__ == "__main__":
    main()
