"""Host logic of the batched touch-sampling stages (puzzletunneldiscovery_b200/stages.py): the rays a seeded run
draws are those of the reference's per-sample loop (touchq_util.py:8-27 over uw_random.py:19-44), and the file layout
is preprocess_surface.sample_touch's (`touchq_batch-*.npz`, :95).  CPU only: the walk is a recorded stub."""
import math

import numpy as np

from puzzletunneldiscovery_b200 import pyosr, sharding, stages


class _RecordingWorld:
    recommended_cres = 1.5e-5

    def __init__(self):
        self.calls = []

    def transit_states_to_with_contact(self, fr, to, vd):
        self.calls.append((np.array(fr), np.array(to), vd))
        n = len(fr)
        return (np.array(fr) + 0.25, np.array(to) - 0.25, np.arange(n) % 2 == 0, np.full(n, 0.5), np.full(n, 0.75))


def _reference_loop_targets(q0s):
    """touchq_util.sample_one_touch's draws, sample by sample, on numpy's global generator (literal restatement)."""
    out = []
    for q0 in q0s:
        while True:
            x1, x2 = 2.0 * (np.random.rand(2) - 0.5)
            l2 = x1 * x1 + x2 * x2
            if l2 < 1:
                break
        tr = np.array([2 * x1 * math.sqrt(1 - l2), 2 * x2 * math.sqrt(1 - l2), 1 - 2 * l2]) * 1.0
        while True:
            y1, y2, y3 = 2.0 * (np.random.rand(3) - 0.5)
            m2 = y1 * y1 + y2 * y2 + y3 * y3
            if m2 <= 1:
                break
        aa = np.array([y1, y2, y3]) * (2 * math.pi)
        out.append(pyosr.apply(q0, tr, aa))
    return np.array(out)


def test_calc_touch_draws_the_reference_rays_and_makes_one_call():
    q0 = np.array([0.1, -0.2, 0.3, 0.5, 0.5, -0.5, 0.5])
    np.random.seed(11)
    want = _reference_loop_targets(np.repeat(q0[None], 37, axis=0))
    after_ref = np.random.rand()
    np.random.seed(11)
    uw = _RecordingWorld()
    free_v, touch_v, to_inf, free_tau, touch_tau = stages.calc_touch(uw, q0, 37, 0.002)
    assert np.random.rand() == after_ref                 # the generator is left where the reference's loop leaves it
    assert len(uw.calls) == 1                            # one device call for the batch
    fr, to, vd = uw.calls[0]
    assert vd == 0.002 and np.array_equal(fr, np.repeat(q0[None], 37, axis=0)) and np.array_equal(to, want)
    assert free_v.shape == (37, 7) and touch_v.shape == (37, 7) and to_inf.dtype == np.bool_ and to_inf.shape == (37,)
    assert free_tau.shape == (37,) and touch_tau.shape == (37,)
    # a private generator gives the same stream as the global one seeded alike
    uw2 = _RecordingWorld()
    stages.calc_touch(uw2, q0, 37, 0.002, rng=np.random.RandomState(11))
    assert np.array_equal(uw2.calls[0][1], want)


def test_sample_touch_layout_and_task_chunks(tmp_path):
    rs = np.random.RandomState(3)
    keys = np.concatenate([rs.randn(5, 3), rs.randn(5, 4)], axis=1)
    keys[:, 3:] /= np.linalg.norm(keys[:, 3:], axis=1, keepdims=True)
    shape = (5, 7)                                       # (keys, touch samples per key), as in sample_touch
    seen = []
    for task in range(3):
        tind = sharding.get_task_chunk(shape, 3, task)
        np.random.seed(100 + task)
        want = _reference_loop_targets(keys[[t[0] for t in tind]])
        np.random.seed(100 + task)
        uw = _RecordingWorld()
        fn = tmp_path / ("touchq_batch-%d.npz" % task)
        d = stages.sample_touch(uw, keys, tind, out_fn=str(fn))
        assert uw.calls[0][2] == uw.recommended_cres and np.array_equal(uw.calls[0][1], want)
        f = np.load(str(fn))
        assert sorted(f.files) == ["FREE_V", "FROM_V", "FROM_VI", "IS_INF", "TOUCH_V"]
        n = len(tind)
        assert f["FROM_VI"].shape == (n,) and f["FROM_V"].shape == (n, 7) and f["FREE_V"].shape == (n, 7)
        assert f["TOUCH_V"].shape == (n, 7) and f["IS_INF"].shape == (n,) and f["IS_INF"].dtype == np.bool_
        assert np.array_equal(f["FROM_V"], keys[f["FROM_VI"]]) and np.array_equal(f["FREE_V"], d["FREE_V"])
        seen += [tuple(t) for t in tind]
    assert sorted(seen) == [(i, j) for i in range(5) for j in range(7)]   # the chunks cover the task grid once
    empty = stages.sample_touch(_RecordingWorld(), keys, [])
    assert empty["FREE_V"].shape == (0, 7) and empty["IS_INF"].shape == (0,)
