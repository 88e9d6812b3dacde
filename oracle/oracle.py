"""ctypes front-end of the CPU oracle (oracle/osr_oracle.cc).

TEST INFRASTRUCTURE ONLY: imported by tests/, by __graft_entry__.smoke() and by bench.py's
cpu_baseline / --impl reference legs.  The product package never imports this module.
Every method cites the reference function it restates (paths relative to /root/reference).
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None
_MOLLER = None
_MOLLER_SYM = "_Z32tri_tri_intersect_with_isectlinePKdS0_S0_S0_S0_S0_PiPdS2_"

_dp = C.POINTER(C.c_double)
_u8p = C.POINTER(C.c_uint8)
_ip = C.POINTER(C.c_int)
_llp = C.POINTER(C.c_longlong)


def build(force=False):
    """Compile the oracle (and oracle/_ref when /root/reference exists)."""
    so = os.path.join(_HERE, "libosr_oracle.so")
    src = os.path.join(_HERE, "osr_oracle.cc")
    stale = (not os.path.exists(so)) or os.path.getmtime(so) < os.path.getmtime(src)
    if force or stale or not os.path.exists(os.path.join(_HERE, "_ref", "libmoller97d.so")):
        subprocess.check_call(["make", "-C", _HERE, "-s"], stdout=subprocess.DEVNULL)
    return so


def _ptr(a, t):
    return a.ctypes.data_as(t) if a is not None else None


def lib():
    global _LIB, _MOLLER
    if _LIB is not None:
        return _LIB
    so = build()
    L = C.CDLL(so)
    L.orc_create.restype = C.c_void_p
    L.orc_destroy.argtypes = [C.c_void_p]
    L.orc_load_env_obj.argtypes = [C.c_void_p, C.c_char_p]
    L.orc_load_rob_obj.argtypes = [C.c_void_p, C.c_char_p]
    L.orc_set_meshes.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_void_p, C.c_int,
                                 C.c_void_p, C.c_int]
    L.orc_enforce_robot_center.argtypes = [C.c_void_p, _dp]
    L.orc_set_perturbation.argtypes = [C.c_void_p, _dp]
    L.orc_finalize.argtypes = [C.c_void_p]
    L.orc_scene_scale.argtypes = [C.c_void_p]
    L.orc_scene_scale.restype = C.c_double
    L.orc_get_center.argtypes = [C.c_void_p, _dp]
    L.orc_num_tris.argtypes = [C.c_void_p, C.c_int]
    L.orc_num_verts.argtypes = [C.c_void_p, C.c_int]
    L.orc_get_unit_vertices.argtypes = [C.c_void_p, C.c_int, _dp]
    L.orc_get_faces.argtypes = [C.c_void_p, C.c_int, _ip]
    L.orc_set_moller.argtypes = [C.c_void_p]
    L.orc_is_valid_states.argtypes = [C.c_void_p, _dp, C.c_longlong, _u8p, C.c_int, C.c_int, _ip]
    L.orc_transit_states.argtypes = [C.c_void_p, _dp, _dp, C.c_longlong, _dp, C.c_int, _dp, _dp, _u8p, _dp, _dp,
                                     _llp, C.c_int]
    L.orc_is_valid_transitions.argtypes = [C.c_void_p, _dp, _dp, C.c_longlong, C.c_double, _u8p, _llp, C.c_int]
    L.orc_intersecting_segments.argtypes = [C.c_void_p, _dp, C.c_longlong, C.c_longlong, _llp, _dp, _ip, _u8p, _llp]
    L.orc_intersecting_segments.restype = C.c_longlong
    L.orc_is_disentangled_states.argtypes = [C.c_void_p, _dp, C.c_longlong, _u8p]
    L.orc_interpolate.argtypes = [_dp, _dp, C.c_double, _dp]
    L.orc_distance.argtypes = [_dp, _dp]
    L.orc_distance.restype = C.c_double
    L.orc_apply.argtypes = [_dp, _dp, _dp, _dp]
    L.orc_sat17.argtypes = [_dp]
    L.orc_contact_multiplicity.argtypes = [_dp]
    L.orc_surface_map.argtypes = [_dp, _ip, C.c_longlong, _dp, _ip, C.c_longlong, C.c_longlong, _ip, _dp]
    L.orc_surface_map.restype = C.c_longlong
    L.orc_ompl_check_motions.argtypes = [C.c_void_p, _dp, _dp, C.c_longlong, C.c_double, _dp, _dp, C.c_double,
                                         C.c_double, _u8p, _llp, C.c_int]
    L.orc_ompl_validate_states.argtypes = [C.c_void_p, _dp, C.c_longlong, C.c_double, _dp, _dp, _u8p]
    ref = os.path.join(_HERE, "_ref", "libmoller97d.so")
    if os.path.exists(ref):
        _MOLLER = C.CDLL(ref)
        fn = getattr(_MOLLER, _MOLLER_SYM)
        L.orc_set_moller(C.cast(fn, C.c_void_p))
    _LIB = L
    return L


def has_reference_moller():
    lib()
    return _MOLLER is not None


def _states(q):
    q = np.ascontiguousarray(np.asarray(q, dtype=np.float64))
    if q.ndim == 1:
        q = q.reshape(1, 7)
    assert q.shape[1] == 7
    return q


def moller_isectline(V, U):
    """The reference's own lib/tritri/Moller97d.cc:579-717, called from oracle/_ref."""
    lib()
    V = np.ascontiguousarray(V, dtype=np.float64)
    U = np.ascontiguousarray(U, dtype=np.float64)
    cop = C.c_int(0)
    p0 = np.zeros(3)
    p1 = np.zeros(3)
    fn = getattr(_MOLLER, _MOLLER_SYM)
    fn.argtypes = [_dp] * 6 + [C.POINTER(C.c_int), _dp, _dp]
    r = fn(_ptr(V[0], _dp), _ptr(V[1], _dp), _ptr(V[2], _dp), _ptr(U[0], _dp), _ptr(U[1], _dp), _ptr(U[2], _dp),
           C.byref(cop), _ptr(p0, _dp), _ptr(p1, _dp))
    return r, cop.value, p0, p1


def sat17(P, Q):
    """intersect-inl.h:728-799 on raw triangles (Q already in P's frame)."""
    t = np.ascontiguousarray(np.concatenate([np.asarray(P, float).reshape(9), np.asarray(Q, float).reshape(9)]))
    return bool(lib().orc_sat17(_ptr(t, _dp)))


def contact_multiplicity(P, Q):
    t = np.ascontiguousarray(np.concatenate([np.asarray(P, float).reshape(9), np.asarray(Q, float).reshape(9)]))
    return int(lib().orc_contact_multiplicity(_ptr(t, _dp)))


def surface_map(TV, TF, V, F, cap=1 << 16):
    """unit_world.cc:1092-1176 on world-space target vertices TV -> (faces[K,3], barycentrics[3K,3])."""
    TV = np.ascontiguousarray(TV, dtype=np.float64); TF = np.ascontiguousarray(TF, dtype=np.int32)
    V = np.ascontiguousarray(V, dtype=np.float64); F = np.ascontiguousarray(F, dtype=np.int32)
    of = np.zeros((cap, 3), np.int32); ol = np.zeros((3 * cap, 3))
    k = lib().orc_surface_map(_ptr(TV, _dp), _ptr(TF, _ip), len(TF), _ptr(V, _dp), _ptr(F, _ip), len(F), cap,
                              _ptr(of, _ip), _ptr(ol, _dp))
    if k < 0:
        raise RuntimeError("oracle/_ref/libmoller97d.so (the reference's Moeller routine) is not built")
    assert k <= cap
    return of[:k].copy(), ol[:3 * k].copy()


def interpolate(a, b, tau):
    a = _states(a)[0]; b = _states(b)[0]
    out = np.zeros(7)
    lib().orc_interpolate(_ptr(a, _dp), _ptr(b, _dp), float(tau), _ptr(out, _dp))
    return out


def distance(a, b):
    a = _states(a)[0]; b = _states(b)[0]
    return lib().orc_distance(_ptr(a, _dp), _ptr(b, _dp))


def apply(q, tr, aa):
    q = _states(q)[0]
    tr = np.ascontiguousarray(tr, dtype=np.float64); aa = np.ascontiguousarray(aa, dtype=np.float64)
    out = np.zeros(7)
    lib().orc_apply(_ptr(q, _dp), _ptr(tr, _dp), _ptr(aa, _dp), _ptr(out, _dp))
    return out


class OracleWorld:
    """Restatement of osr::UnitWorld (lib/osr/unit_world.cc) for the hot-path queries."""

    GEO_ENV = 0
    GEO_ROB = 1

    def __init__(self):
        self.L = lib()
        self.h = C.c_void_p(self.L.orc_create())

    def __del__(self):
        try:
            self.L.orc_destroy(self.h)
        except Exception:
            pass

    # setup ------------------------------------------------------------------------------------
    def loadModelFromFile(self, fn):
        if self.L.orc_load_env_obj(self.h, fn.encode()) != 0:
            raise RuntimeError("cannot load " + fn)

    def loadRobotFromFile(self, fn):
        if self.L.orc_load_rob_obj(self.h, fn.encode()) != 0:
            raise RuntimeError("cannot load " + fn)

    def set_meshes(self, Ve, Fe, Vr, Fr):
        Ve = np.ascontiguousarray(Ve, np.float32); Vr = np.ascontiguousarray(Vr, np.float32)
        Fe = np.ascontiguousarray(Fe, np.int32); Fr = np.ascontiguousarray(Fr, np.int32)
        self.L.orc_set_meshes(self.h, Ve.ctypes.data, len(Ve), Fe.ctypes.data, len(Fe), Vr.ctypes.data, len(Vr),
                              Fr.ctypes.data, len(Fr))

    def enforceRobotCenter(self, c):
        c = np.ascontiguousarray(c, dtype=np.float64)
        self.L.orc_enforce_robot_center(self.h, _ptr(c, _dp))

    def set_perturbation(self, q):
        q = _states(q)[0]
        self.L.orc_set_perturbation(self.h, _ptr(q, _dp))

    def scaleToUnit(self):
        pass

    def angleModel(self, lat, lon):
        if lat != 0.0 or lon != 0.0:
            raise RuntimeError("angleModel arguments must be 0")
        self.L.orc_finalize(self.h)

    @property
    def scene_scale(self):
        return self.L.orc_scene_scale(self.h)

    @property
    def robot_center(self):
        c = np.zeros(3)
        self.L.orc_get_center(self.h, _ptr(c, _dp))
        return c

    def unit_geometry(self, which):
        nv = self.L.orc_num_verts(self.h, which); nf = self.L.orc_num_tris(self.h, which)
        V = np.zeros((nv, 3)); F = np.zeros((nf, 3), np.int32)
        self.L.orc_get_unit_vertices(self.h, which, _ptr(V, _dp))
        self.L.orc_get_faces(self.h, which, _ptr(F, _ip))
        return V, F

    # queries ----------------------------------------------------------------------------------
    def is_valid_states(self, q, mode="bvh", threads=0, counters=False):
        """unit_world.cc:215-247 per row.  mode 'brute' = all-pairs SAT-17 ground truth."""
        q = _states(q)
        out = np.zeros(len(q), np.uint8)
        cnt = np.zeros((len(q), 2), np.int32) if counters else None
        self.L.orc_is_valid_states(self.h, _ptr(q, _dp), len(q), _ptr(out, _u8p), 1 if mode == "brute" else 0,
                                   threads, _ptr(cnt, _ip))
        return (out, cnt) if counters else out

    def is_valid_state(self, q):
        return bool(self.is_valid_states(q)[0])

    def is_disentangled_states(self, q):
        """unit_world.cc:249-264 per row."""
        q = _states(q)
        out = np.zeros(len(q), np.uint8)
        self.L.orc_is_disentangled_states(self.h, _ptr(q, _dp), len(q), _ptr(out, _u8p))
        return out

    def transit_states_to_with_contact(self, fr, to, vd, threads=0, counters=False):
        """unit_world.cc:405-474 per row -> (last_free, first_hit, is_free, free_tau, touch_tau)."""
        fr = _states(fr); to = _states(to)
        n = len(fr)
        vd = np.ascontiguousarray(np.atleast_1d(np.asarray(vd, dtype=np.float64)))
        lf = np.zeros((n, 7)); fh = np.zeros((n, 7)); isf = np.zeros(n, np.uint8)
        ft = np.zeros(n); tt = np.zeros(n)
        cnt = np.zeros((n, 3), np.int64) if counters else None
        self.L.orc_transit_states(self.h, _ptr(fr, _dp), _ptr(to, _dp), n, _ptr(vd, _dp), 1 if len(vd) > 1 else 0,
                                  _ptr(lf, _dp), _ptr(fh, _dp), _ptr(isf, _u8p), _ptr(ft, _dp), _ptr(tt, _dp),
                                  _ptr(cnt, _llp), threads)
        r = (lf, fh, isf, ft, tt)
        return r + (cnt,) if counters else r

    def transit_state_to_with_contact(self, fr, to, vd):
        lf, fh, isf, ft, tt = self.transit_states_to_with_contact(fr, to, vd)
        return lf[0], fh[0], bool(isf[0]), float(np.float32(ft[0])), float(np.float32(tt[0]))

    def is_valid_transitions(self, fr, to, ivd, threads=0, counters=False):
        """unit_world.cc:477-486 per row."""
        fr = _states(fr); to = _states(to)
        out = np.zeros(len(fr), np.uint8)
        cnt = np.zeros((len(fr), 3), np.int64) if counters else None
        self.L.orc_is_valid_transitions(self.h, _ptr(fr, _dp), _ptr(to, _dp), len(fr), float(ivd), _ptr(out, _u8p),
                                        _ptr(cnt, _llp), threads)
        return (out, cnt) if counters else out

    @staticmethod
    def _bounds(bounds):
        if bounds is None:
            return None
        return np.ascontiguousarray(np.concatenate([np.asarray(bounds[0], np.float64).reshape(3),
                                                    np.asarray(bounds[1], np.float64).reshape(3)]))

    def ompl_validate_states(self, q, scale, dc, bounds=None):
        """OmplDriver::validateStates: satisfiesBounds && collision free per OMPL-frame row -> uint8[N]."""
        q = _states(q)
        dc = np.ascontiguousarray(dc, dtype=np.float64)
        b = self._bounds(bounds)
        out = np.zeros(len(q), np.uint8)
        self.L.orc_ompl_validate_states(self.h, _ptr(q, _dp), len(q), float(scale), _ptr(dc, _dp), _ptr(b, _dp),
                                        _ptr(out, _u8p))
        return out

    def ompl_check_motions(self, q0, q1, scale, dc, bounds, resolution, threads=0):
        """DiscreteMotionValidator::checkMotion per OMPL-frame row pair -> (valid uint8[N], nd int64[N]).
        bounds = (mins, maxs) of the translation; the segment length is resolution x their diagonal."""
        q0 = _states(q0); q1 = _states(q1)
        dc = np.ascontiguousarray(dc, dtype=np.float64)
        b = self._bounds(bounds)
        extent = float(np.sqrt(((b[3:] - b[:3]) ** 2).sum()))
        out = np.zeros(len(q0), np.uint8); nd = np.zeros(len(q0), np.int64)
        self.L.orc_ompl_check_motions(self.h, _ptr(q0, _dp), _ptr(q1, _dp), len(q0), float(scale), _ptr(dc, _dp),
                                      _ptr(b, _dp), float(resolution * extent), float(resolution * 0.5 * np.pi),
                                      _ptr(out, _u8p), _ptr(nd, _llp), threads)
        return out, nd

    def intersecting_segments_batch(self, q, cap=1 << 22, counters=False):
        """unit_world.cc:959-1012 per row: CSR offsets, segments[.,6], faces[.,2], coplanar."""
        q = _states(q)
        n = len(q)
        off = np.zeros(n + 1, np.int64); segs = np.zeros((cap, 6)); faces = np.zeros((cap, 2), np.int32)
        cop = np.zeros(cap, np.uint8)
        cnt = np.zeros((n, 2), np.int64) if counters else None
        total = self.L.orc_intersecting_segments(self.h, _ptr(q, _dp), n, cap, _ptr(off, _llp), _ptr(segs, _dp),
                                                 _ptr(faces, _ip), _ptr(cop, _u8p), _ptr(cnt, _llp))
        m = int(off[-1])
        r = (off, segs[:m].copy(), faces[:m].copy(), cop[:m].copy(), int(total))
        return r + (cnt,) if counters else r


def rasterize_barycentric(uv, bary, weight, res_x, res_y):
    """CPU restatement of the reference's barycentric chart splat (lib/osr/shader/bary.{vert,geom,frag} +
    lib/osr/osr_render.cc:590-759): vertex positions blended in float32 like the geometry shader, coverage at
    pixel centres with the top-left rule, linear attribute interpolation, discard outside [0,1], additive
    blending.  PARITY UNPINNED: the GL path cannot run on this box (no graphics stack); this states the
    rasterisation rules the engine's kernel (csrc/kernels_raster.cuh) follows, pixel loops in numpy."""
    uv = np.asarray(uv, np.float32).reshape(-1, 3, 2)
    bary = np.asarray(bary, np.float32).reshape(-1, 3, 3)
    weight = np.asarray(weight, np.float32).reshape(-1, 3)
    out = np.zeros((res_y, res_x), np.float64)

    def top_left(ex, ey):
        return ey < 0.0 or (ey == 0.0 and ex < 0.0)
    for p in range(len(uv)):
        u, b, wt = uv[p], bary[p], weight[p]
        X = np.zeros(3); Y = np.zeros(3)
        for v in range(3):
            bx = np.float32(u[0, 0] * b[v, 0]); bx = np.float32(bx + np.float32(u[1, 0] * b[v, 1]))
            bx = np.float32(bx + np.float32(u[2, 0] * b[v, 2]))
            by = np.float32(u[0, 1] * b[v, 0]); by = np.float32(by + np.float32(u[1, 1] * b[v, 1]))
            by = np.float32(by + np.float32(u[2, 1] * b[v, 2]))
            X[v] = float(bx) * res_x; Y[v] = float(by) * res_y
        area = (X[1] - X[0]) * (Y[2] - Y[0]) - (X[2] - X[0]) * (Y[1] - Y[0])
        if not np.isfinite(area) or area == 0.0:
            continue
        i1, i2 = (1, 2) if area > 0 else (2, 1)
        area = abs(area)
        x0, y0, x1, y1, x2, y2 = X[0], Y[0], X[i1], Y[i1], X[i2], Y[i2]
        tl = (top_left(x2 - x1, y2 - y1), top_left(x0 - x2, y0 - y2), top_left(x1 - x0, y1 - y0))
        px0 = int(max(np.floor(min(x0, x1, x2)), 0)); px1 = int(min(np.ceil(max(x0, x1, x2)), res_x))
        py0 = int(max(np.floor(min(y0, y1, y2)), 0)); py1 = int(min(np.ceil(max(y0, y1, y2)), res_y))
        if px1 <= px0 or py1 <= py0:
            continue
        cx, cy = np.meshgrid(np.arange(px0, px1) + 0.5, np.arange(py0, py1) + 0.5)
        e0 = (x2 - x1) * (cy - y1) - (y2 - y1) * (cx - x1)
        e1 = (x0 - x2) * (cy - y2) - (y0 - y2) * (cx - x2)
        e2 = (x1 - x0) * (cy - y0) - (y1 - y0) * (cx - x0)
        inside = ((e0 > 0) | ((e0 == 0) & tl[0])) & ((e1 > 0) | ((e1 == 0) & tl[1])) & ((e2 > 0) | ((e2 == 0) & tl[2]))
        l0, l1, l2 = e0 / area, e1 / area, e2 / area
        fb = [l0 * float(b[0, k]) + l1 * float(b[i1, k]) + l2 * float(b[i2, k]) for k in range(3)]
        mn = np.minimum(np.minimum(fb[0], fb[1]), fb[2]); mx = np.maximum(np.maximum(fb[0], fb[1]), fb[2])
        keep = inside & ~(mn < 0.0) & ~(mx > 1.0)
        fw = l0 * float(wt[0]) + l1 * float(wt[i1]) + l2 * float(wt[i2])
        out[py0:py1, px0:px1] += np.where(keep, fw.astype(np.float32), 0.0)
    return out
