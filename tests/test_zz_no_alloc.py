"""The allocation-free entry point osr_is_valid_transitions_device_noalloc (include/osrgpu.h; no counterpart in the
reference) against the existing osr_is_valid_transitions_device, through
the C ABI.  Collected last on purpose.  Passes alone in a fresh process and after
the other tests: it builds its own world and touches nothing the other tests hold."""
import ctypes as C

import numpy as np
import pytest

from conftest import make_world
from puzzletunneldiscovery_b200 import workloads as W

pytestmark = pytest.mark.gpu

MAX_SEGMENTS = 1 << 22   # OSR_NOALLOC_MAX_SEGMENTS of include/osrgpu.h


def _labels(L, h, fn, d_from, d_to, n, vd, d_out):
    from puzzletunneldiscovery_b200._lib import check
    check(fn(h, C.c_void_p(d_from.data_ptr()), C.c_void_p(d_to.data_ptr()), n, vd, C.c_void_p(d_out.data_ptr())))
    check(L.osr_sync(h))
    return d_out.cpu().numpy().copy()


def test_noalloc_entry_point_equals_the_allocating_one_on_two_calls_in_a_row():
    import torch
    from puzzletunneldiscovery_b200 import pyosr
    from puzzletunneldiscovery_b200._lib import OsrError, check
    uw = make_world(pyosr.UnitWorld, "org")
    L, h = uw._L, uw._h
    q = W.random_states(1 << 14, seed=201)
    ms = q[uw.is_valid_states(q)]
    # call 1: k-NN edges between free milestones (mostly free walks of ~40 000 steps at this resolution);
    # call 2: fewer, different segments -- touch rays into the environment (mostly blocked) at a coarser resolution.
    # The static scratch keeps the first call's contents; the second call must not see them.
    fr1, to1 = uw.knn_edges(ms[:2048], k=4, window=256)
    fr2, to2 = W.touch_rays(ms[:64], 3001, seed=202)
    for (fr, to, vd) in ((fr1, to1, uw.recommended_cres), (fr2, to2, 0.0125 / 8)):
        n = len(fr)
        d_from = torch.from_numpy(np.ascontiguousarray(fr)).cuda()
        d_to = torch.from_numpy(np.ascontiguousarray(to)).cuda()
        d_a = torch.full((n,), 7, dtype=torch.uint8, device="cuda")
        d_b = torch.full((n,), 9, dtype=torch.uint8, device="cuda")
        torch.cuda.synchronize()
        want = _labels(L, h, L.osr_is_valid_transitions_device, d_from, d_to, n, vd, d_a)
        got = _labels(L, h, L.osr_is_valid_transitions_device_noalloc, d_from, d_to, n, vd, d_b)
        assert got.max() <= 1 and np.array_equal(got, want)            # labels are integers: bit for bit
        assert np.array_equal(want.astype(bool), uw.is_valid_transitions(fr, to, vd))   # ... and the host call's
        assert 0 < want.sum() < n                                       # both outcomes occur
        # the same call again gives the same labels (nothing left over in the scratch matters)
        again = _labels(L, h, L.osr_is_valid_transitions_device_noalloc, d_from, d_to, n, vd, d_b)
        assert np.array_equal(again, want)
    # sizes beyond the stated maximum are rejected with a clear error, before anything is launched
    rc = L.osr_is_valid_transitions_device_noalloc(h, C.c_void_p(d_from.data_ptr()), C.c_void_p(d_to.data_ptr()),
                                                   MAX_SEGMENTS + 1, 0.01, C.c_void_p(d_b.data_ptr()))
    assert rc != 0
    with pytest.raises(OsrError, match="OSR_NOALLOC_MAX_SEGMENTS"):
        check(rc)
    # an empty call is a no-op
    check(L.osr_is_valid_transitions_device_noalloc(h, None, None, 0, 0.01, None))
    check(L.osr_sync(h))
