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


class _PinnedBlock:
    """Owner of one page-locked allocation; freed when the last numpy view of it goes away."""

    def __init__(self, nbytes):
        self._L = _lib.load()
        self.ptr = self._L.osr_host_alloc(C.c_size_t(max(1, nbytes)))
        if not self.ptr:
            raise OsrError("osr_host_alloc(%d) failed (no CUDA device?)" % nbytes)

    def __del__(self):
        if getattr(self, "ptr", None):
            self._L.osr_host_free(C.c_void_p(self.ptr))
            self.ptr = None


def pinned_empty(shape, dtype=np.float64):
    """numpy array in page-locked host memory (osr_host_alloc): sample / result arrays built in it make the batched
    calls run at PCIe speed instead of the driver's pageable staging speed."""
    dtype = np.dtype(dtype)
    shape = (shape,) if np.isscalar(shape) else tuple(shape)
    nbytes = int(np.prod(shape, dtype=np.int64)) * dtype.itemsize
    block = _PinnedBlock(nbytes)
    buf = (C.c_char * max(1, nbytes)).from_address(block.ptr)
    buf._osr_block = block          # the ctypes buffer keeps the allocation alive; numpy keeps the buffer
    return np.frombuffer(buf, dtype=dtype, count=int(np.prod(shape, dtype=np.int64))).reshape(shape)


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


def _quat_normalized(q):
    q = np.asarray(q, dtype=np.float64)
    n2 = float((q * q).sum())
    return q / np.sqrt(n2) if n2 > 0 else q.copy()


def _quat_mul(a, b):
    aw, ax, ay, az = a
    bw, bx, by, bz = b
    return np.array([aw * bw - ax * bx - ay * by - az * bz, aw * bx + ax * bw + ay * bz - az * by,
                     aw * by + ay * bw + az * bx - ax * bz, aw * bz + az * bw + ax * by - ay * bx])


def _quat_to_rot(q):
    w, x, y, z = q
    tx, ty, tz = 2 * x, 2 * y, 2 * z
    twx, twy, twz = tx * w, ty * w, tz * w
    txx, txy, txz = tx * x, ty * x, tz * x
    tyy, tyz, tzz = ty * y, tz * y, tz * z
    return np.array([[1 - (tyy + tzz), txy - twz, txz + twy], [txy + twz, 1 - (txx + tzz), tyz - twx],
                     [txz - twy, tyz + twx, 1 - (txx + tyy)]])


def _quat_rotate(q, v):
    """Eigen Quaternion::_transformVector: v + w (2 u x v) + u x (2 u x v)."""
    u = np.asarray(q[1:4], dtype=np.float64)
    uv = 2.0 * np.cross(u, v)
    return v + q[0] * uv + np.cross(u, uv)


def _angle_axis_from_quat(q):
    """Eigen 3.3 AngleAxis(Quaternion): angle = 2 atan2(|vec|, |w|), axis = vec / (+-|vec|); (0, x) if vec = 0."""
    n = float(np.linalg.norm(q[1:4]))
    if n != 0.0:
        angle = 2.0 * np.arctan2(n, abs(q[0]))
        if q[0] < 0:
            n = -n
        return angle, np.asarray(q[1:4], dtype=np.float64) / n
    return 0.0, np.array([1.0, 0.0, 0.0])


def _angular_distance(a, b):
    """Eigen 3.3 Quaternion::angularDistance: 2 atan2(|vec(d)|, |w(d)|), d = a * conj(b)."""
    d = _quat_mul(a, np.array([b[0], -b[1], -b[2], -b[3]]))
    return 2.0 * np.arctan2(np.linalg.norm(d[1:4]), abs(d[0]))


def _quat_from_two_vectors(a, b):
    """Eigen 3.3 Quaternion::setFromTwoVectors (generic branch; antiparallel: any perpendicular axis)."""
    v0 = np.asarray(a, dtype=np.float64) / np.linalg.norm(a)
    v1 = np.asarray(b, dtype=np.float64) / np.linalg.norm(b)
    c = float(v1 @ v0)
    if c < -1.0 + 1e-12:
        h = np.array([0.0, 1.0, 0.0]) if abs(v0[0]) > 0.9 else np.array([1.0, 0.0, 0.0])
        ax = np.cross(v0, h)
        ax /= np.linalg.norm(ax)
        return np.concatenate([[0.0], ax])
    ax = np.cross(v0, v1)
    s = np.sqrt((1.0 + c) * 2.0)
    return np.concatenate([[s * 0.5], ax / s])


def decompose_2(state):
    """osr_state.cc:94-99 -> (translation, rotation matrix of the normalised quaternion)."""
    s = _states(state)[0]
    return s[0:3].copy(), _quat_to_rot(_quat_normalized(s[3:7]))


def decompose_3(state):
    """osr_state.cc:101-108 -> (translation, angle, axis)."""
    s = _states(state)[0]
    angle, axis = _angle_axis_from_quat(_quat_normalized(s[3:7]))
    return s[0:3].copy(), angle, axis


def extract_rotation_matrix(state):
    """osr_state.cc:169-174 (quaternion used as given)."""
    return _quat_to_rot(_states(state)[0][3:7])


def differential(from_state, to_state):
    """osr_state.cc:176-222 -> (translation, angle-axis vector, SE(3) delta state)."""
    a = _states(from_state)[0]; b = _states(to_state)[0]
    tr = b[0:3] - a[0:3]
    qa, qb = a[3:7], b[3:7]
    inv = np.array([qa[0], -qa[1], -qa[2], -qa[3]]) / float((qa * qa).sum())   # Eigen inverse(): conj / |q|^2
    delta = _quat_mul(qb, inv)
    angle, axis = _angle_axis_from_quat(delta)
    aav = axis * angle
    half = 0.5 * angle
    rec = _quat_mul(np.concatenate([[np.cos(half)], np.sin(half) * axis]), qa)
    if _angular_distance(qb, rec) > 1e-3:
        raise OsrError("differential is buggy")
    nrm = np.linalg.norm(aav)
    if nrm > 0:
        rec2 = _quat_mul(np.concatenate([[np.cos(0.5 * nrm)], np.sin(0.5 * nrm) * (aav / nrm)]), qa)
        if _angular_distance(qb, rec2) > 1e-3:
            raise OsrError("differential is buggy in aa->quat")
    return tr, aav, np.concatenate([tr, _quat_normalized(delta)])


def multi_differential(from_state, to_states, with_se3=False):
    """osr_state.cc:224-241."""
    tos = _states(to_states)
    rows = []
    for t in tos:
        tr, aav, se3 = differential(from_state, t)
        rows.append(np.concatenate([tr, aav, se3]) if with_se3 else np.concatenate([tr, aav]))
    return np.array(rows).reshape(len(tos), 13 if with_se3 else 6)


def path_metrics(qs):
    """osr_state.cc:46-59: cumulative osr distance along a path."""
    qs = _states(qs)
    out = np.zeros(len(qs))
    d = 0.0
    for i in range(len(qs)):
        if i > 0:
            d += distance(qs[i - 1], qs[i])
        out[i] = d
    return out


def path_interpolate(qs, metrics, tau):
    """osr_state.cc:61-84."""
    qs = _states(qs)
    metrics = np.asarray(metrics, dtype=np.float64)
    key = -1
    for i in range(len(qs) - 1):
        if metrics[i] <= tau < metrics[i + 1]:
            key = i
            break
    if key < 0:
        return qs[0].copy()
    d = metrics[key + 1] - metrics[key]
    t = 0.0 if d < 1e-6 else (tau - metrics[key]) / d
    return interpolate(qs[key], qs[key + 1], t)


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
        self._perturbation = q.copy()

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
    def is_valid_states(self, qs, out=None):
        """Batched is_valid_state.  `out` (optional): a contiguous uint8 / bool array of len(qs) that receives the
        labels in place (e.g. from pyosr.pinned_empty, so that the read-back is a pinned copy too)."""
        n = len(qs)
        if out is None:
            res = np.empty(n, np.uint8)
        else:
            if not (isinstance(out, np.ndarray) and out.dtype.itemsize == 1 and out.shape == (n,) and
                    out.flags.c_contiguous):
                raise OsrError("out must be a contiguous 1-byte array of len(qs)")
            res = out.view(np.uint8)
        if isinstance(qs, np.ndarray) and qs.dtype == np.float32 and qs.ndim == 2 and qs.shape[1] == 7:
            q = np.ascontiguousarray(qs)      # float32 samples go up as they are and are widened on the device
            check(self._L.osr_is_valid_states_f32(self._h, _p(q, c_fp), n, _p(res, c_u8p)))
        else:
            q = _states(qs)
            check(self._L.osr_is_valid_states(self._h, _p(q, c_dp), n, _p(res, c_u8p)))
        return res.view(np.bool_) if out is not None else res.astype(bool)

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
        # Eigen 3.3 normalized() leaves a zero vector unchanged: a pure translation (aa = 0) is the identity rotation
        axis = aa / ang if ang > 0 else aa
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

    # ---- discrete actions (module.cc:140; unit_world.cc:267-336) ---------------------------------
    def transit_state(self, state, action, transit_magnitude, verify_delta):
        """12 discrete actions: 0-5 = +-x, +-y, +-z translation, 6-11 = rotation about x, y, z.  The reference
        keeps magnitudes and deltas in C float and halves verify_delta until the walk makes progress."""
        state = _states(state)[0]
        if action >= 12:
            return state.copy(), False, 0.0
        if not self.is_valid_state(state):
            return state.copy(), False, 0.0
        action_type = 0 if action < 6 else 1
        deltacap = float(np.float32(verify_delta) if action_type == 0 else np.float32(verify_delta * 2))
        sym = 1.0 if action % 2 == 0 else -1.0
        axis_id = (action - 6 * action_type) // 2
        tfvec = np.zeros(3); tfvec[axis_id] = 1.0
        trans = state[0:3].copy()
        rot = _quat_normalized(state[3:7])
        if action_type == 0:
            trans = trans + deltacap * tfvec * sym
        else:
            half = 0.5 * deltacap * sym
            rot = _quat_mul(np.concatenate([[np.cos(half)], np.sin(half) * tfvec]), rot)
        to_state = np.concatenate([trans, _quat_normalized(rot)])
        cur = np.float32(verify_delta)
        while True:
            if cur == 0.0:   # verify_delta vanished: the reference would now walk forever (unit_world.cc:445-470)
                return state.copy(), False, 0.0
            free_state, done, prog = self.transit_state_to(state, to_state, float(cur))
            if not self.is_valid_state(free_state):
                raise OsrError("SAN check failed, invalid state from transitStateTo")
            if prog > 0.0:
                return free_state, done, prog
            cur = np.float32(cur / np.float32(2.0))

    # ---- frames (module.cc:136-137, 164-167, 241, 249-250) ----------------------------------------
    @property
    def perturbation(self):
        return getattr(self, "_perturbation", get_identity_state_vector())

    def get_ompl_center(self, geo=1):
        c = np.zeros(3)
        check(self._L.osr_get_ompl_center(self._h, int(geo), _p(c, c_dp)))
        return c

    @property
    def scene_matrix(self):
        """Scene::getCalibrationTransform after angleModel(0,0): glm::scale(scene_scale) in C float."""
        s = np.float32(self.scene_scale)
        return np.diag([s, s, s, np.float32(1)]).astype(np.float64)

    @property
    def robot_matrix(self):
        """scale(scene_scale) * translate(-center), float arithmetic (unit_world.cc:152-158, scene.cc:211-230)."""
        s = np.float32(self.scene_scale)
        m = np.diag([s, s, s, np.float32(1)]).astype(np.float32)
        m[0:3, 3] = s * (-self.robot_center.astype(np.float32))
        return m.astype(np.float64)

    def translate_vanilla_to_ompl(self, qs):
        """unit_world.cc:606-631: w-last quaternions, t_ompl = t_vanilla + R * ompl_center."""
        q = _states(qs)
        oc = self.get_ompl_center(self.GEO_ROB)
        out = q.copy()
        for i in range(len(q)):
            rot = _quat_normalized(np.array([q[i, 6], q[i, 3], q[i, 4], q[i, 5]]))
            out[i, 0:3] = q[i, 0:3] + _quat_rotate(rot, oc)
        return out

    def translate_ompl_to_vanilla(self, qs):
        """unit_world.cc:641-661."""
        q = _states(qs)
        oc = self.get_ompl_center(self.GEO_ROB)
        out = q.copy()
        for i in range(len(q)):
            rot = _quat_normalized(np.array([q[i, 6], q[i, 3], q[i, 4], q[i, 5]]))
            out[i, 0:3] = q[i, 0:3] - _quat_rotate(rot, oc)
        return out

    def translate_vanilla_to_unit(self, qs):
        return self.translate_ompl_to_unit(self.translate_vanilla_to_ompl(qs))

    def translate_vanilla_pts_to_unit(self, geo, pts):
        """unit_world.cc:669-687: (pts - ompl_center [robot only]) scaled by the calibration matrix."""
        pts = np.asarray(pts, dtype=np.float64).reshape(-1, 3)
        if len(pts) == 0:
            return np.zeros((0, 3))
        if geo != self.GEO_ENV:
            pts = pts - self.get_ompl_center(self.GEO_ROB)
        return pts * float(np.float32(self.scene_scale))

    # ---- sampled key configurations (module.cc:211-218; unit_world.cc:1295-1337) ------------------
    def sample_free_configuration(self, rob_surface_point, rob_surface_normal, env_surface_point, env_surface_normal,
                                  margin, max_trials=-1, rng=None, batch=64):
        """Random rotation about the environment normal until the placed robot is free.  The reference draws
        one angle per trial from std::random_device; here trials are evaluated `batch` at a time on the GPU
        (angles from `rng`, a numpy Generator) and the first free one in draw order is returned, or the last
        trial when max_trials is exhausted -- the reference's return value in that case."""
        rng = rng or np.random.default_rng()
        rp, rn, ep, en = (np.asarray(x, dtype=np.float64).reshape(3) for x in
                          (rob_surface_point, rob_surface_normal, env_surface_point, env_surface_normal))
        rob_o = rp + rn * margin
        env_o = ep + en * margin
        rot1 = _quat_from_two_vectors(rn, -en)
        done = 0
        while True:
            n = batch if max_trials < 0 else min(batch, max_trials + 1 - done)
            ang = rng.uniform(0.0, 2.0 * np.pi, n)
            qs = np.empty((n, 7))
            for i, a in enumerate(ang):
                rot2 = np.concatenate([[np.cos(0.5 * a)], np.sin(0.5 * a) * en])   # AngleAxis: axis used as given
                acc = _quat_mul(rot2, rot1)
                qs[i, 0:3] = env_o - _quat_rotate(acc, rob_o)
                qs[i, 3:7] = _quat_normalized(acc)
            v = self.is_valid_states(qs)
            if v.any():
                return qs[int(np.argmax(v))]
            done += n
            if max_trials >= 0 and done > max_trials:
                return qs[-1]

    # ---- OMPL-frame batched validators (lib/pyse3ompl/pyse3ompl.cc:77-81) --------------------------
    @staticmethod
    def _ompl_bounds(bounds):
        """(lo[3], hi[3]) of the planner's RealVectorBounds (OmplDriver.set_bounds) -> float64[6]."""
        if bounds is None:
            return None
        b = np.ascontiguousarray(np.concatenate([np.asarray(bounds[0], np.float64).reshape(3),
                                                 np.asarray(bounds[1], np.float64).reshape(3)]))
        return b

    def ompl_validate_states(self, qs, bounds=None):
        """OmplDriver.validate_states: int32[N], rows x y z qx qy qz qw in the OMPL frame.  `bounds` = (mins, maxs)
        as given to OmplDriver.set_bounds: a state outside them is invalid (FCLStateValidityChecker::isValid);
        None = collision only."""
        q = _states(qs)
        out = np.empty(len(q), np.uint8)
        b = self._ompl_bounds(bounds)
        check(self._L.osr_ompl_validate_states(self._h, _p(q, c_dp), len(q), _p(b, c_dp), _p(out, c_u8p)))
        return out.astype(np.int32)

    def ompl_validate_motion_pairs(self, qs0, qs1, bounds, resolution, with_counts=False):
        """OmplDriver.validate_motion_pairs under OMPL's DiscreteMotionValidator: int32[min(N0, N1)].
        `bounds` = (mins, maxs) of OmplDriver.set_bounds, `resolution` = OmplDriver.set_cdres."""
        a = _states(qs0); b = _states(qs1)
        n = min(len(a), len(b))
        a = np.ascontiguousarray(a[:n]); b = np.ascontiguousarray(b[:n])
        out = np.empty(n, np.uint8)
        nd = np.empty(n, np.int64) if with_counts else None
        check(self._L.osr_ompl_check_motions(self._h, _p(a, c_dp), _p(b, c_dp), n, _p(self._ompl_bounds(bounds), c_dp),
                                             float(resolution), _p(out, c_u8p), _p(nd, c_i64p)))
        return (out.astype(np.int32), nd) if with_counts else out.astype(np.int32)

    # ---- k-NN edge generation -----------------------------------------
    def knn_windowed(self, qs, k=8, window=4096, with_distance=False):
        """k nearest states under pyosr.distance inside windows of consecutive indices -> int32[N,k]."""
        q = _states(qs)
        idx = np.empty((len(q), k), np.int32)
        dist = np.empty((len(q), k)) if with_distance else None
        check(self._L.osr_knn_windowed(self._h, _p(q, c_dp), len(q), int(k), int(window), _p(idx, c_i32p),
                                       _p(dist, c_dp)))
        return (idx, dist) if with_distance else idx

    def knn_validate_edges(self, qs, k=8, window=4096, verify_delta=None):
        """Windowed k-NN and is_valid_transition of every edge found, in one call that uploads only the milestones
        (osr_knn_validate_edges): (idx int32[N, k], valid bool[N, k]); valid is False where idx is -1."""
        q = _states(qs)
        n = len(q)
        vd = self.recommended_cres if verify_delta is None else float(verify_delta)
        idx = np.empty((n, k), np.int32); ok = np.empty((n, k), np.uint8)
        check(self._L.osr_knn_validate_edges(self._h, _p(q, c_dp), n, int(k), int(window), vd, _p(idx, c_i32p),
                                             _p(ok, c_u8p)))
        return idx, ok.astype(bool)

    def knn_edges(self, qs, k=8, window=4096):
        """(from, to) state arrays of the k-NN edges, milestone-major like workloads.knn_edges."""
        q = _states(qs)
        idx = self.knn_windowed(q, k, window)
        src = np.repeat(np.arange(len(q)), k)
        dst = idx.reshape(-1)
        keep = dst >= 0
        return np.ascontiguousarray(q[src[keep]]), np.ascontiguousarray(q[dst[keep]])

    # ---- clearance statistics (touchq_util.calc_touch + preprocess_key.py:275-281 on the device) ----
    def clearance_stats(self, q0, to, verify_delta=None, with_distances=False):
        """For every candidate unit state q0[c] and its touch targets to[c, i] (shape [M, npoint, 7]; build them
        with pyosr.apply as touchq_util.sample_one_touch does): walk q0[c] -> to[c, i] until first contact and
        return float64[M, 5] = [median, max, min, mean, std] of pyosr.multi_distance(q0[c], FREE_V) -- the row
        preprocess_key.py keeps per candidate -- and, when asked, the distances themselves
        (keyconf.estimate_keyconf_clearance's DISTANCE_BATCH)."""
        q = _states(q0)
        t = np.ascontiguousarray(np.asarray(to, dtype=np.float64))
        if t.ndim != 3 or t.shape[0] != len(q) or t.shape[2] != 7:
            raise ValueError("`to` must have shape [len(q0), npoint, 7]")
        npoint = t.shape[1]
        vd = self.recommended_cres if verify_delta is None else float(verify_delta)
        stats = np.empty((len(q), 5))
        dist = np.empty((len(q), npoint)) if with_distances else None
        check(self._L.osr_clearance_stats(self._h, _p(q, c_dp), len(q), _p(t, c_dp), int(npoint), vd,
                                          _p(stats, c_dp), _p(dist, c_dp)))
        return (stats, dist) if with_distances else stats

    # ---- surface mapping (module.cc: intersecting_to_robot_surface / intersecting_to_model_surface) ----
    def _to_surface(self, which, q, q_is_unit, V, F):
        """unit_world.cc:928-957 + 1092-1176 -> (faces int32[K,3], barycentrics float64[3K,3])."""
        qu = _states(q)[0] if q_is_unit else self.translate_to_unit_state(q)
        V = np.ascontiguousarray(np.asarray(V, dtype=np.float64).reshape(-1, 3))
        F = np.ascontiguousarray(np.asarray(F, dtype=np.int32).reshape(-1, 3))
        total = C.c_int64(0)
        cap = 1 << 14
        while True:
            of = np.zeros((cap, 3), np.int32); ob = np.zeros((3 * cap, 3))
            check(self._L.osr_intersecting_to_surface(self._h, which, _p(np.ascontiguousarray(qu), c_dp), _p(V, c_dp),
                                                      len(V), _p(F, c_i32p), len(F), cap, _p(of, c_i32p),
                                                      _p(ob, c_dp), C.byref(total)))
            k = int(total.value)
            if k <= cap:
                return of[:k].copy(), ob[:3 * k].copy()
            cap = k

    def intersecting_to_robot_surface(self, q, q_is_unit, V, F):
        return self._to_surface(self.GEO_ROB, q, q_is_unit, V, F)

    def intersecting_to_model_surface(self, q, q_is_unit, V, F):
        return self._to_surface(self.GEO_ENV, q, q_is_unit, V, F)

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

    @staticmethod
    def enum_configuration_states(pts, margin, denominator):
        """The candidate states of enum_free_configurations built on the host (no labels): float64[B, den, 7]."""
        pts = np.ascontiguousarray(np.asarray(pts, dtype=np.float64).reshape(-1, 12))
        q = np.empty((len(pts), denominator, 7))
        check(_lib.load().osr_enum_configuration_states(_p(pts, c_dp), len(pts), float(margin), int(denominator),
                                                        _p(q, c_dp)))
        return q

    def enum_free_configuration(self, rob_surface_point, rob_surface_normal, env_surface_point, env_surface_normal,
                                margin, denominator, only_median=False):
        pts = np.concatenate([np.asarray(x, dtype=np.float64).reshape(3) for x in
                              (rob_surface_point, rob_surface_normal, env_surface_point, env_surface_normal)])
        q, valid, _ = self.enum_free_configurations(pts, margin, denominator, only_median)
        return q[0][valid[0]]

    # ---- state-frame maps (module.cc:157-169) ----------------------------------------------------
    def enum_2drot_free_configuration(self, rob_surface_point, rob_surface_normal, rob_prim_id, env_surface_point,
                                      env_surface_normal, env_prim_id, margin, azimuth_divider, altitude_divider,
                                      return_all=False, enable_mt=True):
        """UnitWorld::enum2DRotationFreeConfiguration (unit_world.cc:1391-1523), restated with its quirks because
        src/GP/pipeline/keyconf.py:244 consumes exactly this output: the azimuth step is 2 pi / ALTITUDE_divider
        (:1435), validity is stubbed to 1 in the reference (:1471-1476: no collision query is issued, so there is
        nothing for the GPU to do), and the connected-component pass indexes `valids` and `qs` by the altitude ROW
        (:1493, :1517).  Host only (numpy, Eigen 3.3 semantics).  enable_mt only gated the stubbed loop."""
        rp = np.asarray(rob_surface_point, np.float64); rn = np.asarray(rob_surface_normal, np.float64)
        ep = np.asarray(env_surface_point, np.float64); en = np.asarray(env_surface_normal, np.float64)
        Vr, Fr = self._geometry(self.GEO_ROB)
        Ve, Fe = self._geometry(self.GEO_ENV)
        rob_v = Vr[Fr[int(rob_prim_id)]]
        env_v = Ve[Fe[int(env_prim_id)]]
        tan_env = env_v[1] - env_v[0]
        nt = np.linalg.norm(tan_env)
        if nt > 0:
            tan_env = tan_env / nt
        az, al = int(azimuth_divider), int(altitude_divider)
        env_o = ep + en * margin
        rot_1 = _quat_from_two_vectors(rn, -en)

        def aa_quat(angle, axis):     # Eigen Quaternion(AngleAxis): the axis is used as given
            return np.concatenate([[np.cos(0.5 * angle)], np.sin(0.5 * angle) * axis])
        qs = np.zeros((az * al, 7))
        d_alt = 0.5 * np.pi / float(al)
        d_azi = 2.0 * np.pi / float(al)
        for i in range(al):
            rot_2 = aa_quat(i * d_alt, tan_env)
            acc = _quat_mul(rot_2, rot_1)
            align = env_o - _quat_rotate(acc, rp)
            pen = min(float(en @ (_quat_rotate(acc, rob_v[k]) + align - ep)) for k in range(3))
            alt_margin = margin if pen > 0 else -pen + margin
            alt_env_o = ep + en * alt_margin
            for j in range(az):
                acc3 = _quat_mul(_quat_mul(aa_quat(j * d_azi, en), rot_2), rot_1)
                qs[i * az + j] = np.concatenate([alt_env_o - _quat_rotate(acc3, rp), _quat_normalized(acc3)])
        valids = np.ones(az * al, np.int32)
        if return_all:
            return qs[valids != 0]
        out = []
        for i in range(al):
            for j in range(az):
                if not valids[i]:
                    continue
                queue = [i * az + j]
                while queue:
                    cur = queue.pop(0)
                    ci, cj = cur // az, cur % az
                    for dx in (-1, 1):
                        for dy in (-1, 1):
                            pi_, pj = (ci + dx + al) % al, (cj + dy + az) % az
                            k = pi_ * az + pj
                            if valids[k]:
                                queue.append(k)
                                valids[k] = 0
                out.append(qs[i].copy())
        return np.array(out).reshape(-1, 7)

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


class Renderer(UnitWorld):
    """Headless subset of pyosr.Renderer (module.cc:253-300): the barycentric UV-chart splat that turns
    `intersecting_to_*_surface` output into a surface weight chart (src/GP/pipeline/preprocess_surface.py:248-311).
    The GL context, depth / RGB rendering and the other Renderer members are outside the accelerated path; the
    splat itself runs as a CUDA rasteriser (csrc/kernels_raster.cuh)."""

    BARY_RENDERING_ROBOT = 0
    BARY_RENDERING_SCENE = 1

    def __init__(self, devices=None):
        super().__init__(devices)
        self._uv = {0: None, 1: None}
        self._brd = {0: ([], [], []), 1: ([], [], [])}

    def set_uv(self, target, uv):
        """Per-vertex texture coordinates of the target mesh (Mesh::getUV), rows in the vertex numbering of
        get_robot_geometry / get_scene_geometry."""
        self._uv[self._target(target)] = np.ascontiguousarray(np.asarray(uv, dtype=np.float32).reshape(-1, 2))

    def loadRobotTextureImage(self, fn):   # kept for call compatibility; the chart splat needs no texture image
        pass

    @staticmethod
    def _target(target):
        if target not in (0, 1):
            raise OsrError("invalid barycentric rendering target")
        return int(target)

    def add_barycentric(self, F, V, target, weight=1.0):
        """Renderer::addBarycentric (osr_render.cc:538-569): F int[P,3] = vertex ids of the surface face of each
        primitive, V float[3P,3] = barycentrics of its three vertices, one weight for the whole batch."""
        t = self._target(target)
        if self._uv[t] is None:
            raise OsrError("the target mesh has no UV coordinates (set_uv)")
        F = np.asarray(F, dtype=np.int64).reshape(-1, 3)
        V = np.asarray(V, dtype=np.float32).reshape(-1, 3)
        if len(V) != 3 * len(F):
            raise OsrError("V must hold three barycentric rows per face of F")
        uvs, barys, weights = self._brd[t]
        uvs.append(self._uv[t][F.reshape(-1)])
        barys.append(V)
        weights.append(np.full(len(V), weight, dtype=np.float32))

    def clear_barycentric(self, target):
        self._brd[self._target(target)] = ([], [], [])

    def render_barycentric(self, target, res, svg_fn=""):
        """Renderer::renderBarycentric (osr_render.cc:590-759) -> float32[res[0], res[1]], rows bottom-up."""
        t = self._target(target)
        res = np.asarray(res, dtype=np.int64).reshape(2)
        uvs, barys, weights = self._brd[t]
        out = np.zeros((int(res[1]), int(res[0])), np.float32)
        if uvs:
            uv = np.ascontiguousarray(np.concatenate(uvs)); ba = np.ascontiguousarray(np.concatenate(barys))
            we = np.ascontiguousarray(np.concatenate(weights))
        else:
            uv = ba = we = None
        n = 0 if uv is None else len(uv) // 3
        check(self._L.osr_rasterize_barycentric(self._h, _p(uv, c_fp), _p(ba, c_fp), _p(we, c_fp), n, int(res[0]),
                                                int(res[1]), _p(out, c_fp)))
        return out
