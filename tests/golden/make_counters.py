"""Writes tests/golden/counters_org.json: the ORACLE's FCL-order DFS work counters (OBB-pair tests, triangle-pair
tests; FCL's num_bv_tests / num_leaf_tests) summed over each seeded 2^20-state batch bench.py uses on rank 0,
plus the per-state means.  These define the algorithmic bytes of SURVEY.md 8(d) D8 that bench.py's roofline
uses.  Run from the repo root:  python tests/golden/make_counters.py"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import oracle as O  # noqa: E402
from puzzletunneldiscovery_b200 import workloads as W  # noqa: E402
import bench  # noqa: E402

N = 1 << 20


def main():
    ow = bench.make_world(O.OracleWorld)
    per = {}
    tb = tl = tn = tv = 0
    for b in range(bench.NBATCH):
        seed = bench.batch_seed(0, b)
        q = W.random_states(N, seed)
        v, c = ow.is_valid_states(q, counters=True)
        per[str(seed)] = {"n": N, "sum_bv": int(c[:, 0].sum()), "sum_leaf": int(c[:, 1].sum()), "valid": int(v.sum())}
        tb += int(c[:, 0].sum()); tl += int(c[:, 1].sum()); tn += N; tv += int(v.sum())
    out = {"geometry": "G-org (alpha_env-1.2.org.obj x alpha-1.2.org.obj)", "generator": "workloads.random_states(2^20, seed)",
           "mean_bv": tb / tn, "mean_leaf": tl / tn, "valid_fraction": tv / tn, "per_seed": per}
    with open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "counters_org.json"), "w") as f:
        json.dump(out, f, indent=1)
    print(json.dumps({k: out[k] for k in ("mean_bv", "mean_leaf", "valid_fraction")}))


if __name__ == "__main__":
    main()
