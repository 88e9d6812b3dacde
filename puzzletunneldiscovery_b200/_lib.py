"""ctypes binding of libosrgpu.so (include/osrgpu.h).  Fails loudly if the library is missing:
there is no Python / CPU fallback for any query."""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
# OSR_LIB selects a development build of the same library (build.py --variant) for A/B runs
_SO = os.environ.get("OSR_LIB") or os.path.join(_HERE, "libosrgpu.so")
_L = None

c_dp = C.POINTER(C.c_double)
c_u8p = C.POINTER(C.c_uint8)
c_i8p = C.POINTER(C.c_int8)
c_i32p = C.POINTER(C.c_int32)
c_i64p = C.POINTER(C.c_int64)
c_fp = C.POINTER(C.c_float)


class Counters(C.Structure):
    _fields_ = [("bv_tests", C.c_uint64), ("leaf_tests", C.c_uint64), ("fp64_rechecks", C.c_uint64),
                ("poses", C.c_uint64), ("spills", C.c_uint64), ("kernel_launches", C.c_uint64),
                ("warp_iters", C.c_uint64), ("lanes_active", C.c_uint64), ("lanes_waiting", C.c_uint64),
                ("lanes_idle", C.c_uint64)]


# name -> (restype, argtypes): exactly the symbols include/osrgpu.h declares
SIGNATURES = {
    "osr_last_error": (C.c_char_p, []),
    "osr_create": (C.c_int, [C.POINTER(C.c_void_p)]),
    "osr_destroy": (None, [C.c_void_p]),
    "osr_load_env_obj": (C.c_int, [C.c_void_p, C.c_char_p]),
    "osr_load_rob_obj": (C.c_int, [C.c_void_p, C.c_char_p]),
    "osr_set_env_mesh": (C.c_int, [C.c_void_p, c_fp, C.c_int, c_i32p, C.c_int]),
    "osr_set_rob_mesh": (C.c_int, [C.c_void_p, c_fp, C.c_int, c_i32p, C.c_int]),
    "osr_enforce_robot_center": (C.c_int, [C.c_void_p, c_dp]),
    "osr_finalize": (C.c_int, [C.c_void_p, C.POINTER(C.c_int), C.c_int]),
    "osr_set_perturbation": (C.c_int, [C.c_void_p, c_dp]),
    "osr_scene_scale": (C.c_double, [C.c_void_p]),
    "osr_get_recommended_cres": (C.c_double, [C.c_void_p]),
    "osr_set_recommended_cres": (C.c_int, [C.c_void_p, C.c_double]),
    "osr_get_robot_center": (C.c_int, [C.c_void_p, c_dp]),
    "osr_get_ompl_center": (C.c_int, [C.c_void_p, C.c_int, c_dp]),
    "osr_num_devices": (C.c_int, [C.c_void_p]),
    "osr_mesh_sizes": (C.c_int, [C.c_void_p, C.c_int] + [C.POINTER(C.c_int)] * 4),
    "osr_get_unit_mesh": (C.c_int, [C.c_void_p, C.c_int, c_dp, c_i32p]),
    "osr_get_bvh_nodes": (C.c_int, [C.c_void_p, C.c_int, c_fp]),
    "osr_get_counters": (C.c_int, [C.c_void_p, C.POINTER(Counters), C.c_int]),
    "osr_is_valid_states": (C.c_int, [C.c_void_p, c_dp, C.c_int64, c_u8p]),
    "osr_is_valid_states_f32": (C.c_int, [C.c_void_p, c_fp, C.c_int64, c_u8p]),
    "osr_is_disentangled_states": (C.c_int, [C.c_void_p, c_dp, C.c_int64, c_u8p]),
    "osr_transit_states_to_with_contact": (C.c_int, [C.c_void_p, c_dp, c_dp, C.c_int64, c_dp, C.c_int64, c_dp, c_dp,
                                                     c_u8p, c_dp, c_dp, c_i64p]),
    "osr_is_valid_transitions": (C.c_int, [C.c_void_p, c_dp, c_dp, C.c_int64, C.c_double, c_u8p]),
    "osr_visibility_matrix2": (C.c_int, [C.c_void_p, c_dp, C.c_int64, C.c_int, c_dp, C.c_int64, C.c_int, C.c_double,
                                         c_i8p]),
    "osr_visibility_pair": (C.c_int, [C.c_void_p, c_dp, C.c_int64, C.c_int, c_dp, C.c_int64, C.c_int, C.c_double,
                                      c_i8p]),
    "osr_intersecting_segments": (C.c_int, [C.c_void_p, c_dp, C.c_int64, C.c_int64, c_i64p, c_dp, c_i32p, c_u8p, c_dp,
                                            c_u8p, c_i64p]),
    "osr_intersecting_to_surface": (C.c_int, [C.c_void_p, C.c_int, c_dp, c_dp, C.c_int64, c_i32p, C.c_int64, C.c_int64,
                                              c_i32p, c_dp, c_i64p]),
    "osr_rasterize_barycentric": (C.c_int, [C.c_void_p, c_fp, c_fp, c_fp, C.c_int64, C.c_int, C.c_int, c_fp]),
    "osr_ompl_validate_states": (C.c_int, [C.c_void_p, c_dp, C.c_int64, c_dp, c_u8p]),
    "osr_ompl_check_motions": (C.c_int, [C.c_void_p, c_dp, c_dp, C.c_int64, c_dp, C.c_double, c_u8p, c_i64p]),
    "osr_knn_windowed": (C.c_int, [C.c_void_p, c_dp, C.c_int64, C.c_int, C.c_int64, c_i32p, c_dp]),
    "osr_knn_validate_edges": (C.c_int, [C.c_void_p, c_dp, C.c_int64, C.c_int, C.c_int64, C.c_double, c_i32p, c_u8p]),
    "osr_clearance_stats": (C.c_int, [C.c_void_p, c_dp, C.c_int64, c_dp, C.c_int, C.c_double, c_dp, c_dp]),
    "osr_enum_free_configurations": (C.c_int, [C.c_void_p, c_dp, C.c_int64, C.c_double, C.c_int, C.c_int, c_dp, c_u8p,
                                               c_i32p]),
    "osr_enum_configuration_states": (C.c_int, [c_dp, C.c_int64, C.c_double, C.c_int, c_dp]),
    "osr_translate_to_unit": (C.c_int, [C.c_void_p, c_dp, C.c_int64, c_dp]),
    "osr_translate_from_unit": (C.c_int, [C.c_void_p, c_dp, C.c_int64, c_dp]),
    "osr_translate_ompl_to_unit": (C.c_int, [C.c_void_p, c_dp, C.c_int64, c_dp]),
    "osr_translate_unit_to_ompl": (C.c_int, [C.c_void_p, c_dp, C.c_int64, c_dp]),
    "osr_host_alloc": (C.c_void_p, [C.c_size_t]),
    "osr_host_free": (None, [C.c_void_p]),
    "osr_distance": (C.c_double, [c_dp, c_dp]),
    "osr_walk_deltas": (C.c_int, [c_dp, C.c_double, C.c_double, C.c_int, c_dp]),
    "osr_triangle_pair_predicates": (C.c_int, [c_dp, C.c_int64, c_u8p, c_u8p]),
    "osr_interpolate": (None, [c_dp, c_dp, C.c_double, c_dp]),
    "osr_apply": (None, [c_dp, c_dp, c_dp, c_dp]),
    "osr_is_valid_states_device": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p]),
    "osr_transit_states_device": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_double] +
                                  [C.c_void_p] * 6),
    "osr_is_valid_transitions_device": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_double,
                                                  C.c_void_p]),
    "osr_is_valid_transitions_device_noalloc": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_double,
                                                          C.c_void_p]),
    "osr_stream": (C.c_void_p, [C.c_void_p]),
    "osr_sync": (C.c_int, [C.c_void_p]),
    "osr_last_kernel_ms": (C.c_int, [C.c_void_p, C.POINTER(C.c_float)]),
}


def library_path():
    return _SO


def load():
    """Load libosrgpu.so.  Raises if it has not been built (python -m puzzletunneldiscovery_b200.build)."""
    global _L
    if _L is not None:
        return _L
    if not os.path.exists(_SO):
        raise RuntimeError("libosrgpu.so is missing: run `python __graft_entry__.py` or "
                           "`python puzzletunneldiscovery_b200/build.py` (the engine has no CPU fallback)")
    L = C.CDLL(_SO)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(L, name)  # AttributeError here = header/library mismatch
        fn.restype = res
        fn.argtypes = args
    _L = L
    return L


class OsrError(RuntimeError):
    """Mirrors the std::runtime_error -> RuntimeError translation of the reference's pybind11 module."""


def check(rc):
    if rc != 0:
        msg = load().osr_last_error()
        raise OsrError("osrgpu error %d: %s" % (rc, msg.decode() if msg else "?"))
