"""Host build of the kernels' triangle-pair predicates (csrc/sat17.h through osr_triangle_pair_predicates) against
the oracle's SAT-17 (intersect-inl.h:728-799): the exact predicate IS the label and the fp32 filter may only answer
when it agrees with it -- on generic pairs and on the degenerate families where SAT axes vanish (parallel edges,
shared vertices, touching, needles, coordinates on a coarse grid).  Runs without a GPU."""
import numpy as np
import pytest

from puzzletunneldiscovery_b200 import _lib


def _run(T):
    L = _lib.load()
    n = len(T)
    T = np.ascontiguousarray(T.reshape(n, 18))
    f = np.zeros(n, np.uint8); e = np.zeros(n, np.uint8)
    assert L.osr_triangle_pair_predicates(T.ctypes.data_as(_lib.c_dp), n, f.ctypes.data_as(_lib.c_u8p),
                                          e.ctypes.data_as(_lib.c_u8p)) == 0
    return f, e.astype(bool)


def _pairs(kind, n, rs):
    P = rs.randn(n, 3, 3) * 0.02 + rs.randn(n, 1, 3) * 0.3
    Q = P.mean(1, keepdims=True) + rs.randn(n, 3, 3) * 0.02
    if kind == "near":
        Q = Q + rs.randn(n, 1, 3) * 0.01
    elif kind == "parallel_edges":
        Q[:, 1] = Q[:, 0] + (P[:, 1] - P[:, 0]) * rs.rand(n, 1) * 2
    elif kind == "shared_vertex":
        Q[:, 0] = P[:, 0]
    elif kind == "touching":       # a vertex of Q on P's plane, inside P
        b = rs.rand(n, 3); b /= b.sum(1, keepdims=True)
        Q[:, 0] = np.einsum("nj,njk->nk", b, P)
    elif kind == "needle":
        P[:, 2] = P[:, 0] + (P[:, 1] - P[:, 0]) * rs.rand(n, 1) + rs.randn(n, 3) * 1e-5
    elif kind == "grid":           # coordinates on a coarse grid: exact ties in the projections
        P = np.round(P * 64) / 64
        Q = np.round((P.mean(1, keepdims=True) + rs.randn(n, 3, 3) * 0.05) * 64) / 64
    elif kind == "tiny":           # the scale of G-hi leaf triangles, far from the origin of the pair
        P = rs.randn(n, 3, 3) * 2e-4 + rs.randn(n, 1, 3) * 0.3
        Q = P.mean(1, keepdims=True) + rs.randn(n, 3, 3) * 2e-4
    elif kind.startswith("thin"):  # long thin triangles (aspect 1 .. 1000) like the faces of the alpha puzzle: the
        # per-axis error bounds of the filter scale with the edges involved, and these are the pairs that probe them
        def thin(c):
            a = 10 ** rs.uniform(-2, 0, size=(n, 1, 1)); r = 10 ** rs.uniform(-3, 0, size=(n, 1, 1))
            d1 = rs.randn(n, 1, 3); d1 /= np.linalg.norm(d1, axis=2, keepdims=True)
            d2 = rs.randn(n, 1, 3); d2 -= (d2 * d1).sum(2, keepdims=True) * d1
            d2 /= np.linalg.norm(d2, axis=2, keepdims=True)
            o = c - 0.5 * a * d1
            return np.concatenate([o, o + a * d1, o + a * rs.rand(n, 1, 1) * d1 + a * r * d2], axis=1), a * r
        P, w = thin(rs.randn(n, 1, 3) * 0.3)
        off = rs.randn(n, 1, 3) * w * rs.choice([0, 1e-7, 1e-5, 1e-3, 1e-1, 1], size=(n, 1, 1))
        Q, _ = thin(P.mean(1, keepdims=True) + off)
        if kind == "thin_touching":    # a vertex of Q in P's plane
            b = rs.rand(n, 3); b /= b.sum(1, keepdims=True)
            Q[:, 0] = np.einsum("nj,njk->nk", b, P)
        elif kind == "thin_shared":    # a shared vertex
            Q[:, 0] = P[:, 1]
    return np.concatenate([P, Q], axis=1)


@pytest.mark.parametrize("kind", ["random", "near", "parallel_edges", "shared_vertex", "touching", "needle", "grid", "tiny",
                                  "thin", "thin_touching", "thin_shared"])
def test_filter_never_contradicts_the_exact_predicate(kind, oracle_mod):
    rs = np.random.RandomState(hash(kind) % 1000)
    T = _pairs(kind, 60000, rs)
    filt, exact = _run(T)
    assert not ((filt == 0) & exact).any()            # "certainly separated" is never an intersecting pair
    assert not ((filt == 1) & ~exact).any()           # "certainly intersecting" is never a separated pair
    sub = rs.choice(len(T), 1500, replace=False)
    want = np.array([oracle_mod.sat17(t[:3], t[3:]) for t in T[sub]])
    assert np.array_equal(exact[sub], want)           # the exact predicate is the oracle's SAT-17, bit for bit
    if kind in ("random", "near"):
        assert (filt != 2).mean() > 0.99              # ... and generic pairs are decided in fp32
    if kind == "thin":
        assert (filt != 2).mean() > 0.9               # ... and so are long thin triangles (one L^4 bound left half open)


def test_predicate_known_answers():
    # the reference's own triangle pair (sancheck/fclcontacts/main.cc:21-22,46-54): intersecting
    P = np.array([[0, 0, 0], [1, 0, 0], [0, 0, 1.0]])
    Q = np.array([[0, 0, 0], [1, 0, 0], [0, 1.0, 0]]) + np.array([-0.49, -0.5, 0.5])
    filt, exact = _run(np.concatenate([P, Q])[None])
    assert exact[0] and filt[0] in (1, 2)
    # parallel, separated by a hair: nobody may claim intersection
    P = np.array([[0, 0, 0], [1, 0, 0], [0, 1.0, 0]])
    Q = P + np.array([0, 0, 1e-9])
    filt, exact = _run(np.concatenate([P, Q])[None])
    assert not exact[0] and filt[0] != 1
    # empty batch
    assert _lib.load().osr_triangle_pair_predicates(None, 0, None, None) == 0
