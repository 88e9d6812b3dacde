"""Synthetic, seeded input batches for the collision hot path.

Vectorised restatements of the reference's samplers (paths relative to /root/reference):
  * random_states          src/GP/pipeline/uw_random.py:8-17  (float32 values widened to float64)
  * random_on_sphere       src/GP/pipeline/uw_random.py:19-34
  * random_within_sphere   src/GP/pipeline/uw_random.py:36-44
  * touch_rays             src/GP/pipeline/touchq_util.py:8-13 (to = pyosr.apply(q0, tr, aa))
The per-sample formulas are the reference's; the random stream is consumed in bulk
(rows of RandomState.rand) instead of one Python call per sample, which only changes the
interleaving of the rejection samplers' draws.  Host-side numpy only: no collision work here.
"""
import math
import os

import numpy as np

REF_ALPHA_FILES = {
    "org": ("alpha_env-1.2.org.obj", "alpha-1.2.org.obj"),
    "wt2": ("alpha_env-1.2.wt2.obj", "alpha-1.2.wt2.obj"),
}
# src/GP/aniconf12_2.py:19
ROB_OMPL_CENTER = np.array([16.973146438598633, 1.2278236150741577, 10.204807281494141])


def assets_dir():
    """Directory with the alpha puzzle meshes: tests/assets in the repo (committed copies of
    the reference's res/alpha data files, needed on the GPU box where /root/reference is absent)."""
    here = os.path.dirname(os.path.abspath(__file__))
    return os.path.join(os.path.dirname(here), "tests", "assets")


def random_states(n, seed=0, scale=1.0, rs=None):
    rs = rs or np.random.RandomState(seed)
    u = rs.rand(n, 6)
    tr = scale * (u[:, 0:3] - 0.5)
    u1, u2, u3 = u[:, 3], u[:, 4], u[:, 5]
    quat = np.stack([np.sqrt(1 - u1) * np.sin(2 * math.pi * u2), np.sqrt(1 - u1) * np.cos(2 * math.pi * u2),
                     np.sqrt(u1) * np.sin(2 * math.pi * u3), np.sqrt(u1) * np.cos(2 * math.pi * u3)], axis=1)
    q = np.concatenate([tr, quat], axis=1).astype(np.float32)
    return np.ascontiguousarray(q.astype(np.float64))


def _rejection(rs, n, dim, strict):
    out = np.zeros((0, dim))
    while len(out) < n:
        x = 2.0 * (rs.rand(max(1024, int(2.2 * (n - len(out)))), dim) - 0.5)
        l2 = (x * x).sum(axis=1)
        keep = (l2 < 1) if strict else (l2 <= 1)
        out = np.concatenate([out, x[keep]], axis=0)
    return out[:n]


def random_on_sphere(n, rs, scale=1.0):
    x = _rejection(rs, n, 2, True)
    l2 = (x * x).sum(axis=1)
    s = np.sqrt(1 - l2)
    return np.stack([2 * x[:, 0] * s, 2 * x[:, 1] * s, 1 - 2 * l2], axis=1) * scale


def random_within_sphere(n, rs, scale=1.0):
    return _rejection(rs, n, 3, False) * scale


def quat_mul(a, b):
    aw, ax, ay, az = a[:, 0], a[:, 1], a[:, 2], a[:, 3]
    bw, bx, by, bz = b[:, 0], b[:, 1], b[:, 2], b[:, 3]
    return np.stack([aw * bw - ax * bx - ay * by - az * bz, aw * bx + ax * bw + ay * bz - az * by,
                     aw * by + ay * bw + az * bx - ax * bz, aw * bz + az * bw + ax * by - ay * bx], axis=1)


def apply_batch(q0, tr, aa):
    """Batched lib/osr/osr_state.cc:244-259."""
    q0 = np.asarray(q0, np.float64)
    rot0 = q0[:, 3:7] / np.linalg.norm(q0[:, 3:7], axis=1, keepdims=True)
    ang = np.linalg.norm(aa, axis=1)
    axis = np.where(ang[:, None] == 0.0, np.array([[1.0, 0.0, 0.0]]), aa / np.where(ang == 0, 1.0, ang)[:, None])
    qa = np.concatenate([np.cos(0.5 * ang)[:, None], np.sin(0.5 * ang)[:, None] * axis], axis=1)
    rot = quat_mul(qa, rot0)
    rot = rot / np.linalg.norm(rot, axis=1, keepdims=True)
    return np.ascontiguousarray(np.concatenate([q0[:, 0:3] + tr, rot], axis=1))


def touch_rays(from_states, n, seed=1):
    """n segments (from, to): from-states cycled, to = apply(from, on_sphere(1), within_sphere(2pi))."""
    rs = np.random.RandomState(seed)
    from_states = np.asarray(from_states, np.float64).reshape(-1, 7)
    fr = np.ascontiguousarray(from_states[np.arange(n) % len(from_states)])
    tr = random_on_sphere(n, rs, 1.0)
    aa = random_within_sphere(n, rs, 2 * math.pi)
    return fr, apply_batch(fr, tr, aa)


def state_distance(a, b):
    """Batched lib/osr/osr_state.cc:130-154 (host numpy; used only to choose k-NN edges)."""
    qa = a[:, 3:7] / np.linalg.norm(a[:, 3:7], axis=1, keepdims=True)
    qb = b[:, 3:7] / np.linalg.norm(b[:, 3:7], axis=1, keepdims=True)
    d = quat_mul(qa, qb * np.array([[1.0, -1.0, -1.0, -1.0]]))
    return np.linalg.norm(a[:, 0:3] - b[:, 0:3], axis=1) + 2.0 * np.arctan2(np.linalg.norm(d[:, 1:4], axis=1),
                                                                           np.abs(d[:, 0]))


def knn_edges(milestones, k=8, window=4096):
    """D4: for milestone i the k nearest (osr distance) inside its window of consecutive indices."""
    n = len(milestones)
    src = []
    dst = []
    for w0 in range(0, n, window):
        blk = milestones[w0:w0 + window]
        m = len(blk)
        D = np.empty((m, m))
        for i in range(m):
            D[i] = state_distance(np.repeat(blk[i:i + 1], m, axis=0), blk)
            D[i, i] = np.inf
        nn = np.argpartition(D, min(k, m - 1) - 1, axis=1)[:, :k]
        src.append(np.repeat(np.arange(m), nn.shape[1]) + w0)
        dst.append(nn.reshape(-1) + w0)
    src = np.concatenate(src)
    dst = np.concatenate(dst)
    return np.ascontiguousarray(milestones[src]), np.ascontiguousarray(milestones[dst])


def subdivide(V, F, n):
    """G-hi: split every triangle into n^2 congruent sub-triangles on a barycentric grid."""
    V = np.asarray(V, np.float64)
    F = np.asarray(F, np.int64)
    idx = [(i, j) for i in range(n + 1) for j in range(n + 1 - i)]
    pos = {ij: t for t, ij in enumerate(idx)}
    bary = np.array([[(n - i - j) / n, i / n, j / n] for i, j in idx])
    tris = []
    for i in range(n):
        for j in range(n - i):
            tris.append((pos[(i, j)], pos[(i + 1, j)], pos[(i, j + 1)]))
            if i + j < n - 1:
                tris.append((pos[(i + 1, j)], pos[(i + 1, j + 1)], pos[(i, j + 1)]))
    tris = np.array(tris, np.int64)
    P = np.einsum("gk,fkd->fgd", bary, V[F])  # (nf, G, 3)
    G = len(idx)
    newV = P.reshape(-1, 3).astype(np.float32)
    newF = (tris[None, :, :] + (np.arange(len(F)) * G)[:, None, None]).reshape(-1, 3).astype(np.int32)
    return newV, newF


def load_obj_vf(path):
    """Minimal OBJ reader (v / f only, fan triangulation), same rules as csrc/geometry.cc."""
    V = []
    F = []
    with open(path) as f:
        for line in f:
            if line.startswith("v ") or line.startswith("v\t"):
                V.append([float(x) for x in line.split()[1:4]])
            elif line.startswith("f ") or line.startswith("f\t"):
                idx = [int(t.split("/")[0]) for t in line.split()[1:]]
                idx = [i - 1 if i > 0 else len(V) + i for i in idx]
                for k in range(2, len(idx)):
                    F.append([idx[0], idx[k - 1], idx[k]])
    return np.array(V, np.float32), np.array(F, np.int32)
