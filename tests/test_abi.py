"""CPU-only checks of the drop-in boundary: libosrgpu.so loads, exports every symbol include/osrgpu.h
declares (and the ctypes table binds exactly those), the host-side geometry front end matches the
oracle's unit-world frame, and -- without a CUDA device -- every query fails loudly (no CPU fallback)."""
import ctypes as C
import os
import re

import numpy as np
import pytest

from conftest import ASSETS, ROOT, _has_cuda, make_world
from puzzletunneldiscovery_b200 import _lib, pyosr, workloads as W


def _header_symbols():
    src = open(os.path.join(ROOT, "include", "osrgpu.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(osr_[a-z0-9_]+)\s*\(", src)))


def test_library_exports_every_declared_symbol():
    names = _header_symbols()
    assert len(names) >= 35
    L = C.CDLL(_lib.library_path())
    for n in names:
        assert hasattr(L, n), "libosrgpu.so does not export " + n
    assert sorted(_lib.SIGNATURES) == names  # the Python binding covers the header, nothing more


def test_no_torch_types_in_the_abi():
    src = open(os.path.join(ROOT, "include", "osrgpu.h")).read()
    assert "torch" not in src and "at::" not in src and 'extern "C"' in src


def test_module_level_state_functions_match_oracle(oracle_mod):
    rs = np.random.RandomState(0)
    qs = W.random_states(64, 5)
    for i in range(0, 64, 2):
        a, b = qs[i], qs[i + 1]
        assert pyosr.distance(a, b) == oracle_mod.distance(a, b)
        tau = rs.rand()
        assert np.array_equal(pyosr.interpolate(a, b, tau), oracle_mod.interpolate(a, b, tau))
        tr, aa = rs.randn(3), rs.randn(3)
        assert np.array_equal(pyosr.apply(a, tr, aa), oracle_mod.apply(a, tr, aa))
    assert np.array_equal(pyosr.multi_distance(qs[0], qs[1:5]),
                          np.array([oracle_mod.distance(qs[0], q) for q in qs[1:5]]))
    assert pyosr.STATE_DIMENSION == 7
    assert np.array_equal(pyosr.get_identity_state_vector(), [0, 0, 0, 1, 0, 0, 0])


def _host_world(which, centre=None):
    """Geometry front end only: osr_finalize builds the unit meshes + BVH on the host and, on a box
    without a GPU, returns OSR_ENODEV after doing so."""
    try:
        return make_world(pyosr.UnitWorld, which, centre), True
    except _lib.OsrError as e:
        assert "no CUDA device" in str(e)
    envf, robf = W.REF_ALPHA_FILES[which]
    w = pyosr.UnitWorld()
    w.loadModelFromFile(os.path.join(ASSETS, envf))
    w.loadRobotFromFile(os.path.join(ASSETS, robf))
    if centre is not None:
        w.enforceRobotCenter(centre)
    with pytest.raises(_lib.OsrError):
        w.angleModel(0.0, 0.0)
    return w, False


@pytest.mark.parametrize("which,centre", [("org", None), ("wt2", W.ROB_OMPL_CENTER)])
def test_unit_world_frame_equals_oracle(which, centre, oracle_mod):
    w, _ = _host_world(which, centre)
    ow = make_world(oracle_mod.OracleWorld, which, centre)
    assert w.scene_scale == ow.scene_scale
    assert np.array_equal(w.robot_center, ow.robot_center)
    for g in (0, 1):
        V, F = w._geometry(g)
        Vo, Fo = ow.unit_geometry(g)
        # same triangle soup, bit for bit (vertex numbering may differ: the engine drops unreferenced v rows)
        assert np.array_equal(V[F], Vo[Fo])
        nv, nf, nn, depth = w.mesh_sizes(g)
        assert nn == 2 * nf - 1 and depth < 40


def test_bvh_is_conservative_on_host():
    w, _ = _host_world("org")
    for g in (0, 1):
        V, F = w._geometry(g)
        nodes = w.bvh_nodes(g)
        link = nodes[:, 15].copy().view(np.int32)
        nd = nodes.astype(np.float64)
        leaves = np.nonzero(link < 0)[0]
        assert len(leaves) == len(F)
        assert sorted(-(link[leaves] + 1)) == list(range(len(F)))
        for i in leaves:
            A = nd[i, :12].reshape(3, 4)
            P = V[F[-(link[i] + 1)]] - A[:, 3]
            assert (np.abs(P @ A[:, :3].T) <= nd[i, 12:15] + 1e-12).all()


def test_errors_are_runtime_errors_like_the_reference():
    w = pyosr.UnitWorld()
    with pytest.raises(RuntimeError):  # unit_world.cc:140-142
        w.angleModel(0.1, 0.0)
    with pytest.raises(RuntimeError):
        w.loadModelFromFile("/nonexistent/mesh.obj")
    with pytest.raises(RuntimeError):  # query before scaleToUnit/angleModel
        w.is_valid_states(np.zeros((1, 7)))
    with pytest.raises(RuntimeError):  # robot missing
        w.loadModelFromFile(os.path.join(ASSETS, "alpha_env-1.2.org.obj"))
        w.angleModel(0.0, 0.0)


@pytest.mark.skipif(_has_cuda(), reason="checks the behaviour on a box without a GPU")
def test_no_cpu_fallback_without_a_device():
    w, on_gpu = _host_world("org")
    assert not on_gpu
    q = W.random_states(8, 0)
    for call in (lambda: w.is_valid_states(q), lambda: w.is_disentangled_states(q),
                 lambda: w.transit_states_to_with_contact(q[:4], q[4:], 0.01),
                 lambda: w.is_valid_transitions(q[:4], q[4:], 0.01),
                 lambda: w.intersecting_segments_batch(q, cap=16)):
        with pytest.raises(_lib.OsrError):
            call()


def test_product_package_never_imports_the_oracle():
    pkg = os.path.join(ROOT, "puzzletunneldiscovery_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".cc", ".h")):
                src = open(os.path.join(dirpath, f)).read()
                assert "oracle" not in src.replace("test oracle", "").replace("the oracle", "") or f == "workloads.py" \
                    or "import oracle" not in src and "from oracle" not in src
                assert "from oracle" not in src and "import oracle" not in src and "libosr_oracle" not in src


def test_state_helpers_round_trips(oracle_mod):
    """osr_state.cc helpers mirrored on the host: differential inverts apply, decompose_3 inverts
    compose_from_angleaxis, path_interpolate hits the way points of a path at its own metrics."""
    rs = np.random.RandomState(5)
    a = W.random_states(1, 9)[0]
    for _ in range(16):
        tr = rs.rand(3) - 0.5
        aa = (rs.rand(3) - 0.5) * 2.0
        b = pyosr.apply(a, tr, aa)
        dtr, daa, se3 = pyosr.differential(a, b)
        assert np.abs(dtr - tr).max() < 1e-12 and np.abs(daa - aa).max() < 1e-9
        assert abs(np.linalg.norm(se3[3:]) - 1.0) < 1e-12
    t, ang, ax = pyosr.decompose_3(pyosr.compose_from_angleaxis([1, 2, 3], 0.7, [0, 0.6, 0.8]))
    assert np.allclose(t, [1, 2, 3]) and abs(ang - 0.7) < 1e-12 and np.allclose(ax, [0, 0.6, 0.8])
    t, R = pyosr.decompose_2(a)
    assert np.allclose(R @ R.T, np.eye(3), atol=1e-12) and np.allclose(t, a[:3])
    assert pyosr.multi_differential(a, W.random_states(5, 3), with_se3=True).shape == (5, 13)
    path = np.load(os.path.join(ASSETS, "alpha-1.2.org.w-first.npz"))["VS"][:12]
    m = pyosr.path_metrics(path)
    assert m[0] == 0.0 and (np.diff(m) >= 0).all()
    assert abs(m[3] - sum(oracle_mod.distance(path[i], path[i + 1]) for i in range(3))) < 1e-12
    for i in (0, 4, 10):
        got = pyosr.path_interpolate(path, m, m[i])
        want = path[i].copy(); want[3:] /= np.linalg.norm(want[3:])
        assert np.abs(got - want).max() < 1e-12 or np.abs(got[3:] + want[3:]).max() < 1e-12
    assert np.array_equal(pyosr.path_interpolate(path, m, m[-1] + 1.0), path[0])   # outside: first row (:73-75)


def test_vanilla_frames_and_matrices():
    w, _ = _host_world("wt2", W.ROB_OMPL_CENTER)
    q = W.random_states(16, 4)
    q = np.concatenate([q[:, :3] * 30, q[:, 4:7], q[:, 3:4]], axis=1)     # OMPL-style: w-last
    back = w.translate_ompl_to_vanilla(w.translate_vanilla_to_ompl(q))
    assert np.abs(back - q).max() < 1e-12
    oc = w.get_ompl_center(1)
    V, _ = W.load_obj_vf(os.path.join(ASSETS, W.REF_ALPHA_FILES["wt2"][1]))
    assert np.abs(oc - V.mean(axis=0)).max() < 1e-3
    s = w.scene_scale
    assert np.allclose(np.diag(w.scene_matrix), [s, s, s, 1.0])
    assert np.allclose(w.robot_matrix[:3, 3], -s * np.asarray(W.ROB_OMPL_CENTER), rtol=1e-6)
    p = w.translate_vanilla_pts_to_unit(w.GEO_ROB, oc[None])
    assert np.abs(p).max() < 1e-12 and w.translate_vanilla_pts_to_unit(w.GEO_ENV, np.ones((2, 3))).shape == (2, 3)


def _serial_walk(d, vd, dist, ns):
    """unit_world.cc:445-470's running sum, one fp64 add per step (Python floats are IEEE doubles)."""
    valid, last = 0, d
    for _ in range(ns):
        t = d + vd
        d = min(dist, t) if d < dist else t
        if not d <= dist:
            break
        valid += 1
        last = d
    return d, valid, last


def test_walk_deltas_reproduce_every_rounding_of_the_running_sum():
    """K2 advances over swept windows with osr::advance_chain (k steps inside one binade = one integer
    multiply-add); here its host build is compared bit for bit with the one-add-per-step loop: random walks, walks
    that cross binades and reach dist, and the tie case (vd / ulp(d) ending in exactly one half) where the
    per-step increment is NOT constant."""
    L = _lib.load()
    rs = np.random.RandomState(5)
    cases = []
    for _ in range(400):
        vd = float(10.0 ** rs.uniform(-6, -1) * rs.uniform(1, 2))
        dist = float(vd * 10.0 ** rs.uniform(0.01, 4.5))
        start = float(rs.choice([0.0, vd, vd * rs.randint(1, 2000)]))
        cases.append((start, vd, dist, int(rs.choice([1, 4, 32, 128, 2048, 65536]))))
    for e in range(-12, 1):       # ties: the lowest set bit of vd is half an ulp of d in the binade above vd's
        vd = float(np.ldexp(1.0 + 2.0 ** -52 * (2 * rs.randint(1, 1 << 20) + 1), e))
        cases += [(0.0, vd, vd * 37.3, 64), (vd, vd, vd * 1000.0, 4096), (2 * vd, vd, 2.5 * vd, 8)]
    cases += [(0.0, 0.25, 1.0, 16), (0.0, 0.1, 0.3, 16), (0.7, 0.1, 0.5, 4), (1.0, 1e-17, 2.0, 100),
              (0.0, 1.5164e-5, 0.0831, 8192), (0.05, 1.5164e-5, 0.05 + 1.5164e-5 * 0.5, 64)]
    fast_steps = 0
    for start, vd, dist, ns in cases:
        d = C.c_double(start)
        last = C.c_double(-1.0)
        valid = L.osr_walk_deltas(C.byref(d), vd, dist, ns, C.byref(last))
        wd, wvalid, wlast = _serial_walk(start, vd, dist, ns)
        assert (valid, d.value.hex(), last.value.hex()) == (wvalid, wd.hex(), wlast.hex()), (start, vd, dist, ns)
        fast_steps += valid
    assert fast_steps > 100000
