"""The C++ host side (include/osrgpu.hpp: osr::UnitWorld's member names over the C ABI) used the way a C++ caller of
the reference would: tests/cxx/host_check.cc is compiled with g++ against the header alone, linked to libosrgpu.so and
run as its own process; what it writes is compared with the Python layer's answers on the same states (both sit on
the same C ABI; the Python layer is the one the oracle parity tests cover)."""
import os
import subprocess

import numpy as np
import pytest

from conftest import ASSETS, ROOT
from puzzletunneldiscovery_b200 import workloads as W

PKG = os.path.join(ROOT, "puzzletunneldiscovery_b200")


def _build(tmp_path):
    exe = str(tmp_path / "host_check")
    cmd = ["g++", "-std=c++14", "-O1", "-Wall", "-Wextra", "-Werror", "-I" + os.path.join(ROOT, "include"),
           os.path.join(ROOT, "tests", "cxx", "host_check.cc"), "-o", exe, "-L" + PKG, "-losrgpu",
           "-Wl,-rpath," + PKG]
    r = subprocess.run(cmd, capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    return exe


def _meshes(which="org"):
    envf, robf = W.REF_ALPHA_FILES[which]
    return os.path.join(ASSETS, envf), os.path.join(ASSETS, robf)


def test_cxx_header_compiles_and_fails_loudly_without_a_device(tmp_path):
    """Header + library are enough to build a caller (no CUDA toolkit, no Eigen); set-up errors and the reference's
    angleModel argument check throw std::runtime_error; without a device every query throws (no CPU path)."""
    exe = _build(tmp_path)
    env, rob = _meshes()
    r = subprocess.run([exe, "--no-device", env, rob], capture_output=True, text=True, timeout=120)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "no-device checks passed" in r.stdout


@pytest.mark.gpu
def test_cxx_caller_gets_the_python_layers_answers(tmp_path, uw_org):
    exe = _build(tmp_path)
    env, rob = _meshes()
    n = 4096
    q = np.ascontiguousarray(W.random_states(n, seed=77))
    q[:8, 3:7] *= 1.5            # non-unit quaternions go through the same normalisation either side
    (tmp_path / "states.f64").write_bytes(q.tobytes())
    out = tmp_path / "out.bin"
    r = subprocess.run([exe, env, rob, str(tmp_path / "states.f64"), str(n), str(out)], capture_output=True,
                       text=True, timeout=300)
    assert r.returncode == 0, r.stdout + r.stderr
    buf = out.read_bytes()
    pos = 0

    def take(dtype, count):
        nonlocal pos
        a = np.frombuffer(buf, dtype=dtype, count=count, offset=pos)
        pos += a.nbytes
        return a

    uw = uw_org
    scale, cres = take(np.float64, 2)
    assert scale == uw.scene_scale and cres == uw.recommended_cres
    valid = take(np.uint8, n).astype(bool)
    assert np.array_equal(valid, uw.is_valid_states(q)) and 0 < valid.sum() < n
    assert np.array_equal(take(np.uint8, n).astype(bool), uw.is_disentangled_states(q))
    m = int(take(np.uint64, 1)[0])
    assert m == min(256, valid.sum() - 1)
    fr = take(np.float64, 7 * m).reshape(m, 7)
    to = take(np.float64, 7 * m).reshape(m, 7)
    free = q[valid]
    assert np.array_equal(fr, free[:m]) and np.array_equal(to, free[1:m + 1])
    vd = 0.0125 / 8
    lf, fh, isf, ft, tt = uw.transit_states_to_with_contact(fr, to, vd)
    assert np.array_equal(take(np.float64, 7 * m).reshape(m, 7), lf)         # bit for bit: the same library call
    assert np.array_equal(take(np.float64, 7 * m).reshape(m, 7), fh)
    assert np.array_equal(take(np.uint8, m).astype(bool), isf)
    assert np.array_equal(take(np.float64, m), ft)
    assert np.array_equal(take(np.float64, m), tt)
    assert np.array_equal(take(np.uint8, m).astype(bool), uw.is_valid_transitions(fr, to, vd))
    nh, nr = (int(x) for x in take(np.uint64, 2))
    hit = q[:64][~valid[:64]]
    assert nh == len(hit)
    offsets = take(np.int64, nh + 1)
    faces = take(np.int32, 2 * nr).reshape(nr, 2)
    segs = take(np.float64, 6 * nr).reshape(nr, 6)
    want = uw.intersecting_segments_batch(hit)
    assert np.array_equal(offsets, want["offsets"]) and np.array_equal(faces, want["faces"])
    assert np.array_equal(segs, want["seg"])
    assert nr == offsets[-1] and nr > 0
    u = take(np.float64, 7)
    back = take(np.float64, 7)
    assert np.array_equal(u, uw.translate_to_unit_state(q[0]))
    assert np.array_equal(back, uw.translate_from_unit_state(u))
    om = take(np.float64, 28).reshape(4, 7)
    un = take(np.float64, 28).reshape(4, 7)
    assert np.array_equal(om, uw.translate_unit_to_ompl(q[:4]))
    assert np.array_equal(un, uw.translate_ompl_to_unit(om))
    k = int(take(np.uint64, 1)[0])
    assert k == 512
    oq = uw.translate_unit_to_ompl(q[:k])
    bounds = ((-100.0, -100.0, -100.0), (100.0, 100.0, 100.0))
    v0 = take(np.int32, k)
    v1 = take(np.int32, k)
    mv = take(np.int32, m)
    assert np.array_equal(v0, uw.ompl_validate_states(oq)) and 0 < v0.sum() < k
    assert np.array_equal(v1, uw.ompl_validate_states(oq, bounds))
    assert np.array_equal(mv, uw.ompl_validate_motion_pairs(uw.translate_unit_to_ompl(fr),
                                                            uw.translate_unit_to_ompl(to), bounds, 0.002))
    assert pos == len(buf)
