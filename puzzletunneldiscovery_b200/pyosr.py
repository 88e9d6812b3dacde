"""`pyosr`-compatible front end over libosrgpu.so.

Mirrors the hot subset of the reference's pybind11 module (/root/reference/lib/pyosr/module.cc:77-99,
127-250): same class name, method names, argument meaning and return shapes, so pipeline code such
as src/GP/pipeline/touchq_util.py or solve.py can `import puzzletunneldiscovery_b200.pyosr as pyosr`.
Single-state methods forward to the batched C ABI with N = 1; the batched methods
(`is_valid_states`, `transit_states_to_with_contact`, `is_valid_transitions`,
`intersecting_segments_batch`) are the additions SURVEY.md section 8(b) B7 asks for.
Every query runs on the GPU through the C ABI; nothing here computes collisions.
"""
import ctypes as C

import numpy as np

from . import _lib
from ._lib import OsrError, c_dp, c_fp, c_i8p, c_i32p, c_i64p, c_u8p, check

STATE_DIMENSION = 7


def _p(a, t):
    return a.ctypes.data_as(t) if a is not None else None


def _states(q):
    q = np.ascontiguousarray(np.asarray(q, dtype=np.float64))
    if q.ndim == 1:
        q = q.reshape(1, -1)
    if q.ndim != 2 or q.shape[1] != 7:
        raise OsrError("states must have shape (N, 7)")
    return q


# ---- module-level functions (module.cc:77-99) ----------------------------------------------------
def get_identity_state_vector():
    return np.array([0, 0, 0, 1, 0, 0, 0], dtype=np.float64)


def distance(a, b):
    a = _states(a)[0]; b = _states(b)[0]
    return _lib.load().osr_distance(_p(a, c_dp), _p(b, c_dp))


def multi_distance(origin, targets):
    """osr_state.cc:157-167: one-vs-many distance."""
    origin = _states(origin)[0]
    targets = _states(targets)
    L = _lib.load()
    return np.array([L.osr_distance(_p(origin, c_dp), _p(t, c_dp)) for t in targets])


def interpolate(a, b, tau):
    a = _states(a)[0]; b = _states(b)[0]
    out = np.zeros(7)
    _lib.load().osr_interpolate(_p(a, c_dp), _p(b, c_dp), float(tau), _p(out, c_dp))
    return out


def apply(q, tr, aa):
    q = _states(q)[0]
    tr = np.ascontiguousarray(tr, dtype=np.float64).reshape(3)
    aa = np.ascontiguousarray(aa, dtype=np.float64).reshape(3)
    out = np.zeros(7)
    _lib.load().osr_apply(_p(q, c_dp), _p(tr, c_dp), _p(aa, c_dp), _p(out, c_dp))
    return out


def compose_from_angleaxis(tr, angle, axis):
    """osr_state.cc:121-128 (the axis is used as given, like Eigen::AngleAxis)."""
    axis = np.asarray(axis, dtype=np.float64)
    q = np.concatenate([[np.cos(0.5 * angle)], np.sin(0.5 * angle) * axis])
    q = q / np.linalg.norm(q)
    return np.concatenate([np.asarray(tr, dtype=np.float64), q])


class UnitWorld:
    """Drop-in for pyosr.UnitWorld (hot-path subset).  See include/osrgpu.h for the ABI below it."""

    GEO_ENV = 0
    GEO_ROB = 1

    def __init__(self, devices=None):
        self._L = _lib.load()
        h = C.c_void_p()
        check(self._L.osr_create(C.byref(h)))
        self._h = h
        self._devices = devices
        self._scaled = False
        self.state = get_identity_state_vector()

    def __del__(self):
        try:
            if self._h:
                self._L.osr_destroy(self._h)
                self._h = None
        except Exception:
            pass

    # ---- setup (module.cc:128-137) ---------------------------------------------------------------
    def loadModelFromFile(self, fn):
        check(self._L.osr_load_env_obj(self._h, str(fn).encode()))

    def loadRobotFromFile(self, fn):
        check(self._L.osr_load_rob_obj(self._h, str(fn).encode()))

    def set_meshes(self, env_V, env_F, rob_V, rob_F):
        ev = np.ascontiguousarray(env_V, np.float32); ef = np.ascontiguousarray(env_F, np.int32)
        rv = np.ascontiguousarray(rob_V, np.float32); rf = np.ascontiguousarray(rob_F, np.int32)
        check(self._L.osr_set_env_mesh(self._h, _p(ev, c_fp), len(ev), _p(ef, c_i32p), len(ef)))
        check(self._L.osr_set_rob_mesh(self._h, _p(rv, c_fp), len(rv), _p(rf, c_i32p), len(rf)))

    def enforceRobotCenter(self, center):
        c = np.ascontiguousarray(center, dtype=np.float64).reshape(3)
        check(self._L.osr_enforce_robot_center(self._h, _p(c, c_dp)))

    def scaleToUnit(self):
        self._scaled = True  # the scale is computed together with the unit meshes in angleModel()

    def angleModel(self, latitude, longitude):
        if latitude != 0.0 or longitude != 0.0:
            raise OsrError("The arguments of UnitWorld::angleModel are resevered for compatibility reason. "
                           "Both values must be 0.")
        if self._devices is None:
            check(self._L.osr_finalize(self._h, None, 0))
        else:
            d = (C.c_int * len(self._devices))(*self._devices)
            check(self._L.osr_finalize(self._h, d, len(self._devices)))

    def set_perturbation(self, q):
        q = _states(q)[0]
        check(self._L.osr_set_perturbation(self._h, _p(q, c_dp)))

    @property
    def scene_scale(self):
        return self._L.osr_scene_scale(self._h)

    @property
    def recommended_cres(self):
        return self._L.osr_get_recommended_cres(self._h)

    @recommended_cres.setter
    def recommended_cres(self, v):
        check(self._L.osr_set_recommended_cres(self._h, float(v)))

    @property
    def robot_center(self):
        c = np.zeros(3)
        check(self._L.osr_get_robot_center(self._h, _p(c, c_dp)))
        return c

    def mesh_sizes(self, which):
        v = [C.c_int() for _ in range(4)]
        check(self._L.osr_mesh_sizes(self._h, which, *[C.byref(x) for x in v]))
        return tuple(x.value for x in v)  # nV, nF, nNodes, depth

    def _geometry(self, which):
        nv, nf, _, _ = self.mesh_sizes(which)
        V = np.zeros((nv, 3)); F = np.zeros((nf, 3), np.int32)
        check(self._L.osr_get_unit_mesh(self._h, which, _p(V, c_dp), _p(F, c_i32p)))
        return V, F

    def bvh_nodes(self, which):
        nn = self.mesh_sizes(which)[2]
        out = np.zeros((nn, 16), np.float32)
        check(self._L.osr_get_bvh_nodes(self._h, which, _p(out, c_fp)))
        return out

    def counters(self, reset=False):
        c = _lib.Counters()
        check(self._L.osr_get_counters(self._h, C.byref(c), 1 if reset else 0))
        return {k: int(getattr(c, k)) for k, _ in _lib.Counters._fields_}

    # ---- validity (module.cc:138-139) ------------------------------------------------------------
    def is_valid_states(self, qs):
        q = _states(qs)
        out = np.empty(len(q), np.uint8)
        check(self._L.osr_is_valid_states(self._h, _p(q, c_dp), len(q), _p(out, c_u8p)))
        return out.astype(bool)

    def is_valid_state(self, q):
        return bool(self.is_valid_states(q)[0])

    def is_disentangled_states(self, qs):
        q = _states(qs)
        out = np.empty(len(q), np.uint8)
        check(self._L.osr_is_disentangled_states(self._h, _p(q, c_dp), len(q), _p(out, c_u8p)))
        return out.astype(bool)

    def is_disentangled(self, q):
        return bool(self.is_disentangled_states(q)[0])

    # ---- segments (module.cc:141-156) ------------------------------------------------------------
    def transit_states_to_with_contact(self, froms, tos, verify_delta, with_steps=False):
        fr = _states(froms); to = _states(tos)
        if fr.shape != to.shape:
            raise OsrError("from/to shape mismatch")
        n = len(fr)
        vd = np.ascontiguousarray(np.atleast_1d(np.asarray(verify_delta, dtype=np.float64)))
        if len(vd) not in (1, n):
            raise OsrError("verify_delta must be a scalar or have one value per segment")
        lf = np.empty((n, 7)); fh = np.empty((n, 7)); isf = np.empty(n, np.uint8)
        ft = np.empty(n); tt = np.empty(n)
        st = np.empty(n, np.int64) if with_steps else None
        check(self._L.osr_transit_states_to_with_contact(self._h, _p(fr, c_dp), _p(to, c_dp), n, _p(vd, c_dp), len(vd),
                                                         _p(lf, c_dp), _p(fh, c_dp), _p(isf, c_u8p), _p(ft, c_dp),
                                                         _p(tt, c_dp), _p(st, c_i64p)))
        r = (lf, fh, isf.astype(bool), ft, tt)
        return r + (st,) if with_steps else r

    def transit_state_to_with_contact(self, from_state, to_state, verify_delta):
        lf, fh, isf, ft, tt = self.transit_states_to_with_contact(from_state, to_state, verify_delta)
        # the reference's tuple carries the taus as C float (unit_world.h)
        return lf[0], fh[0], bool(isf[0]), float(np.float32(ft[0])), float(np.float32(tt[0]))

    def transit_state_to(self, from_state, to_state, verify_delta):
        t = self.transit_state_to_with_contact(from_state, to_state, verify_delta)
        return t[0], t[2], t[3]

    def transit_state_by(self, from_state, tr, aa, verify_delta):
        """unit_world.cc:489-506 (target quaternion is NOT renormalised there)."""
        q = _states(from_state)[0]
        aa = np.asarray(aa, dtype=np.float64)
        ang = np.linalg.norm(aa)
        axis = aa / ang
        qa = np.concatenate([[np.cos(0.5 * ang)], np.sin(0.5 * ang) * axis])
        w0, x0, y0, z0 = qa
        w1, x1, y1, z1 = q[3:7]
        rot = np.array([w0 * w1 - x0 * x1 - y0 * y1 - z0 * z1, w0 * x1 + x0 * w1 + y0 * z1 - z0 * y1,
                        w0 * y1 + y0 * w1 + z0 * x1 - x0 * z1, w0 * z1 + z0 * w1 + x0 * y1 - y0 * x1])
        to = np.concatenate([q[0:3] + np.asarray(tr, dtype=np.float64), rot])
        return self.transit_state_to(q, to, verify_delta)

    def is_valid_transitions(self, froms, tos, initial_verify_delta):
        fr = _states(froms); to = _states(tos)
        out = np.empty(len(fr), np.uint8)
        check(self._L.osr_is_valid_transitions(self._h, _p(fr, c_dp), _p(to, c_dp), len(fr),
                                               float(initial_verify_delta), _p(out, c_u8p)))
        return out.astype(bool)

    def is_valid_transition(self, from_state, to_state, initial_verify_delta):
        return bool(self.is_valid_transitions(from_state, to_state, initial_verify_delta)[0])

    # ---- visibility (module.cc:170-184) ----------------------------------------------------------
    def calculate_visibility_matrix2(self, qs0, qs0_is_unit, qs1, qs1_is_unit, verify_magnitude, enable_mt=True):
        a = _states(qs0); b = _states(qs1)
        out = np.empty((len(a), len(b)), np.int8)
        check(self._L.osr_visibility_matrix2(self._h, _p(a, c_dp), len(a), int(bool(qs0_is_unit)), _p(b, c_dp), len(b),
                                             int(bool(qs1_is_unit)), float(verify_magnitude), _p(out, c_i8p)))
        return out

    def calculate_visibility_matrix(self, qs, is_unit, verify_magnitude):
        """unit_world.cc:740-762: symmetric N x N (upper triangle computed, mirrored, zero diagonal)."""
        a = _states(qs)
        n = len(a)
        iu, ju = np.triu_indices(n, 1)
        if not is_unit:
            a = self.translate_to_unit_states(a)
        out = np.zeros((n, n), np.int8)
        if len(iu):
            v = self.is_valid_transitions(a[iu], a[ju], verify_magnitude).astype(np.int8)
            out[iu, ju] = v
            out[ju, iu] = v
        return out

    def calculate_visibility_pair(self, qs0, qs0_is_unit, qs1, qs1_is_unit, verify_magnitude, enable_mt=True):
        a = _states(qs0); b = _states(qs1)
        m = max(len(a), len(b))
        out = np.empty((m, 1), np.int8)
        check(self._L.osr_visibility_pair(self._h, _p(a, c_dp), len(a), int(bool(qs0_is_unit)), _p(b, c_dp), len(b),
                                          int(bool(qs1_is_unit)), float(verify_magnitude), _p(out, c_i8p)))
        return out

    # ---- contacts (module.cc:193) ----------------------------------------------------------------
    def intersecting_segments_batch(self, qs, cap=1 << 20, with_bary=False):
        q = _states(qs)
        n = len(q)
        off = np.zeros(n + 1, np.int64)
        seg = np.zeros((cap, 6)); faces = np.zeros((cap, 2), np.int32); cop = np.zeros(cap, np.uint8)
        bary = np.zeros((cap, 6)) if with_bary else None
        ovf = np.zeros(n, np.uint8)
        total = C.c_int64(0)
        check(self._L.osr_intersecting_segments(self._h, _p(q, c_dp), n, cap, _p(off, c_i64p), _p(seg, c_dp),
                                                _p(faces, c_i32p), _p(cop, c_u8p), _p(bary, c_dp), _p(ovf, c_u8p),
                                                C.byref(total)))
        m = int(off[-1])
        r = {"offsets": off, "seg": seg[:m], "faces": faces[:m], "coplanar": cop[:m], "overflow": ovf.astype(bool),
             "total": int(total.value)}
        if with_bary:
            r["bary"] = bary[:m]
        return r

    def intersecting_segments(self, q):
        r = self.intersecting_segments_batch(q, cap=1 << 16)
        seg = r["seg"]
        mag = np.sqrt(((seg[:, 3:6] - seg[:, 0:3]) ** 2).sum(axis=1)).reshape(-1, 1)
        return seg[:, 0:3].copy(), seg[:, 3:6].copy(), mag, r["faces"]

    # ---- key configurations (module.cc:219-227) --------------------------------------------------
    def enum_free_configurations(self, pts, margin, denominator, only_median=False):
        pts = np.ascontiguousarray(np.asarray(pts, dtype=np.float64).reshape(-1, 12))
        b = len(pts)
        q = np.empty((b, denominator, 7)); valid = np.empty((b, denominator), np.uint8)
        counts = np.empty(b, np.int32)
        check(self._L.osr_enum_free_configurations(self._h, _p(pts, c_dp), b, float(margin), int(denominator),
                                                   int(bool(only_median)), _p(q, c_dp), _p(valid, c_u8p),
                                                   _p(counts, c_i32p)))
        return q, valid.astype(bool), counts

    def enum_free_configuration(self, rob_surface_point, rob_surface_normal, env_surface_point, env_surface_normal,
                                margin, denominator, only_median=False):
        pts = np.concatenate([np.asarray(x, dtype=np.float64).reshape(3) for x in
                              (rob_surface_point, rob_surface_normal, env_surface_point, env_surface_normal)])
        q, valid, _ = self.enum_free_configurations(pts, margin, denominator, only_median)
        return q[0][valid[0]]

    # ---- state-frame maps (module.cc:157-169) ----------------------------------------------------
    def _map(self, fn, qs):
        q = _states(qs)
        out = np.empty_like(q)
        check(fn(self._h, _p(q, c_dp), len(q), _p(out, c_dp)))
        return out

    def translate_to_unit_states(self, qs):
        return self._map(self._L.osr_translate_to_unit, qs)

    def translate_to_unit_state(self, q):
        return self.translate_to_unit_states(q)[0]

    def translate_from_unit_state(self, q):
        return self._map(self._L.osr_translate_from_unit, q)[0]

    def translate_ompl_to_unit(self, qs):
        return self._map(self._L.osr_translate_ompl_to_unit, qs)

    def translate_unit_to_ompl(self, qs, to_angle_axis=False):
        if to_angle_axis:
            raise OsrError("to_angle_axis=True is outside the accelerated path")
        return self._map(self._L.osr_translate_unit_to_ompl, qs)

    def get_robot_geometry(self, q, is_unit=True):
        """unit_world.cc:1525-1555: world-space robot vertices for a state, and faces."""
        V, F = self._geometry(self.GEO_ROB)
        q = _states(q)[0] if is_unit else self.translate_to_unit_state(q)
        return _rotate_points(q, V), F

    def get_scene_geometry(self, q=None, is_unit=True):
        V, F = self._geometry(self.GEO_ENV)
        return V, F


def _rotate_points(q, V):
    w, x, y, z = q[3:7]
    R = np.array([[1 - 2 * (y * y + z * z), 2 * (x * y - z * w), 2 * (x * z + y * w)],
                  [2 * (x * y + z * w), 1 - 2 * (x * x + z * z), 2 * (y * z - x * w)],
                  [2 * (x * z - y * w), 2 * (y * z + x * w), 1 - 2 * (x * x + y * y)]])
    return V @ R.T + q[0:3]
