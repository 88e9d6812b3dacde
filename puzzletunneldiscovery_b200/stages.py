"""The pipeline's touch-sampling loops with ONE device call per batch instead of one per sample.

Mirrors, with the same names, argument meaning, random stream and output layout:
  * touchq_util.sample_one_touch / calc_touch      (/root/reference/src/GP/pipeline/touchq_util.py:8-27)
  * the worker loop of preprocess_surface.sample_touch and its `touchq_batch-*.npz` files
                                                   (/root/reference/src/GP/pipeline/preprocess_surface.py:76-96)
  * uw_random.random_on_sphere / random_within_sphere (/root/reference/src/GP/pipeline/uw_random.py:19-44)

The reference draws a ray and walks it, sample by sample.  Here the rays of a whole batch are drawn first -- with the
same calls on the same generator in the same order, so a seeded run draws the same rays -- and walked by one
`transit_states_to_with_contact` call (K2).  The walk consumes no random numbers, so the results are those of the
reference's loop, row for row.
"""
import math

import numpy as np

from . import pyosr


def random_on_sphere(scale=1.0, rng=np.random):
    """uw_random.py:19-34 (Marsaglia; rejection in the unit disc), drawing from `rng` like the reference does from
    numpy's global generator."""
    while True:
        x1, x2 = 2.0 * (rng.rand(2) - 0.5)
        l2 = x1 * x1 + x2 * x2
        if l2 < 1:
            break
    x = 2 * x1 * math.sqrt(1 - l2)
    y = 2 * x2 * math.sqrt(1 - l2)
    z = 1 - 2 * l2
    return np.array([x, y, z]) * scale


def random_within_sphere(scale=1.0, rng=np.random):
    """uw_random.py:36-44 (rejection in the unit ball)."""
    while True:
        x1, x2, x3 = 2.0 * (rng.rand(3) - 0.5)
        l2 = x1 * x1 + x2 * x2 + x3 * x3
        if l2 <= 1:
            break
    return np.array([x1, x2, x3]) * scale


def touch_targets(q0s, rng=np.random):
    """The `to` states of sample_one_touch (touchq_util.py:8-13) for every row of q0s, drawn in row order:
    tr = random_on_sphere(1.0), aa = random_within_sphere(2 pi), to = pyosr.apply(q0, tr, aa)."""
    q0s = np.asarray(q0s, dtype=np.float64).reshape(-1, 7)
    to = np.empty_like(q0s)
    for i, q0 in enumerate(q0s):
        tr = random_on_sphere(1.0, rng)
        aa = random_within_sphere(2 * math.pi, rng)
        to[i] = pyosr.apply(q0, tr, aa)
    return to


def calc_touch(uw, q0, batch_size, stepping, rng=np.random):
    """touchq_util.calc_touch: batch_size touch rays from q0.  Returns the reference's five arrays
    [free_vertices (B,7), touch_vertices (B,7), to_inf (B,), free_tau (B,), touch_tau (B,)]."""
    q0 = np.asarray(q0, dtype=np.float64).reshape(7)
    fr = np.repeat(q0[None, :], batch_size, axis=0)
    to = touch_targets(fr, rng)
    lf, fh, isf, ft, tt = uw.transit_states_to_with_contact(fr, to, stepping)[:5]
    return [lf, fh, np.asarray(isf, dtype=bool), ft, tt]


def sample_touch(uw, keys, tindices, stepping=None, rng=np.random, out_fn=None):
    """The worker loop of preprocess_surface.sample_touch (:76-96): one touch ray per (key index, sample index) pair
    of `tindices` (partt.get_task_chunk order) at `stepping` (default uw.recommended_cres, as there).  Returns the
    dictionary the reference saves -- FROM_VI, FROM_V, FREE_V, TOUCH_V, IS_INF -- and writes it with numpy.savez
    when out_fn is given (`touchq_batch-<task>.npz`)."""
    keys = np.asarray(keys, dtype=np.float64).reshape(-1, 7)
    ki = np.array([int(t[0]) for t in tindices], dtype=np.int64)
    fr = keys[ki] if len(ki) else np.zeros((0, 7))
    to = touch_targets(fr, rng)
    vd = uw.recommended_cres if stepping is None else stepping
    if len(ki):
        lf, fh, isf = uw.transit_states_to_with_contact(fr, to, vd)[:3]
    else:
        lf, fh, isf = np.zeros((0, 7)), np.zeros((0, 7)), np.zeros(0, bool)
    d = {"FROM_VI": ki, "FROM_V": fr, "FREE_V": lf, "TOUCH_V": fh, "IS_INF": np.asarray(isf, dtype=bool)}
    if out_fn is not None:
        np.savez(out_fn, **d)
    return d


def estimate_keyconf_clearance(uw, oskey_ompl, tindices, nsample, stepping=None, rng=np.random, out_fn=None):
    """The worker loop of keyconf.estimate_keyconf_clearance (/root/reference/src/GP/pipeline/keyconf.py:300-315):
    for every (key_index,) of `tindices` the OMPL-frame candidate oskey_ompl[key_index] is mapped to the unit frame,
    `nsample` touch rays are shot from it at `stepping` (default uw.recommended_cres) and the distances from the
    candidate to the rays' last free states are kept.  Returns DISTANCE_BATCH (len(tindices), nsample) and writes
    `clearance_batch-<task>.npz` (key DISTANCE_BATCH) when out_fn is given.  The rays are drawn candidate by
    candidate, sample by sample, like the reference's nested loops; the walks and the distances of ALL candidates
    are one device call (K2 + K8, `UnitWorld.clearance_stats`), and only the distances come back."""
    oskey_ompl = np.asarray(oskey_ompl, dtype=np.float64)
    ki = [int(t[0]) for t in tindices]
    if not ki:
        d = np.zeros((0, nsample))
    else:
        q0 = np.asarray(uw.translate_ompl_to_unit(oskey_ompl[ki]), dtype=np.float64).reshape(-1, 7)
        to = touch_targets(np.repeat(q0, nsample, axis=0), rng).reshape(len(ki), nsample, 7)
        vd = uw.recommended_cres if stepping is None else stepping
        _, d = uw.clearance_stats(q0, to, vd, with_distances=True)
    if out_fn is not None:
        np.savez(out_fn, DISTANCE_BATCH=d)
    return d
