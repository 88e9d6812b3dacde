"""Multi-GPU host logic on CPU: index-range shards, cost-balanced segment ranges and the gather over
torch.distributed (gloo, world_size 2).  The per-rank evaluator is the CPU oracle here (test infrastructure);
on the GPU box the same `sharded_map` wraps the engine's batched queries (test_gpu_parity.py)."""
import os
import socket

import numpy as np
import pytest

from conftest import make_world
from puzzletunneldiscovery_b200 import sharding as S, workloads as W


def _ref_chunk_bounds(n, num):
    """partt.chunk_it (/root/reference/src/GP/pipeline/partt.py:10-20) on range(n), as (lo, hi) pairs."""
    seq = list(range(n))
    avg = len(seq) / float(num)
    out = []
    last = 0.0
    while last < len(seq) and len(out) < num:
        out.append(seq[int(last):int(last + avg)])
        last += avg
    out[-1] += seq[int(last):]
    return out


@pytest.mark.parametrize("n,parts", [(10, 3), (1 << 20, 8), (1000003, 7), (17, 4), (64, 8)])
def test_task_chunk_is_the_reference_partition(n, parts):
    ref = _ref_chunk_bounds(n, parts)
    assert len(ref) == parts
    covered = []
    for i in range(parts):
        lo, hi = S.task_chunk(n, parts, i)
        assert list(range(lo, hi)) == ref[i]
        covered += list(range(lo, hi)) if n < 100 else []
        elo, ehi = S.engine_shard(n, parts, i)
        assert 0 <= elo <= ehi <= n
    if n < 100:
        assert covered == list(range(n))
    assert S.engine_shard(n, parts, 0)[0] == 0 and S.engine_shard(n, parts, parts - 1)[1] == n


def test_balanced_ranges_cover_and_balance():
    rs = np.random.RandomState(0)
    cost = S.expected_steps(rs.rand(10000) * 3.0, 0.0125 / 8)
    r = S.balanced_ranges(cost, 8)
    assert r[0][0] == 0 and r[-1][1] == len(cost) and all(r[k][1] == r[k + 1][0] for k in range(7))
    sums = np.array([cost[a:b].sum() for a, b in r])
    assert sums.max() / sums.mean() < 1.01
    assert S.balanced_ranges([], 4) == [(0, 0)] * 4
    assert S.expected_steps([0.001], 0.01)[0] == 0      # verify_delta >= dist: no walk (unit_world.cc:445-447)


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, q, fr, to, ret):
    import torch.distributed as dist
    from oracle import oracle as O
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    ow = make_world(O.OracleWorld)
    labels = S.sharded_map(lambda x: ow.is_valid_states(x, threads=1), [q], np.uint8)
    dists = np.array([O.distance(a, b) for a, b in zip(fr, to)])
    cost = S.expected_steps(dists, 0.05)
    taus = S.sharded_map(lambda a, b: np.stack(ow.transit_states_to_with_contact(a, b, 0.05, threads=1)[3:5], axis=1),
                         [fr, to], np.float64, cost=cost)
    t = S.max_over_ranks(1.0 + rank)
    # a rank with an EMPTY range and 2-D rows: one segment over two ranks; all the cost on the last
    # segment, so the first rank's cost-balanced range is everything but it ... and the degenerate n < world case
    lf1 = S.sharded_map(lambda a, b: ow.transit_states_to_with_contact(a, b, 0.05, threads=1)[0], [fr[:1], to[:1]],
                        np.float64)
    skew = np.zeros(len(fr)); skew[-1] = 1.0
    lf2 = S.sharded_map(lambda a, b: ow.transit_states_to_with_contact(a, b, 0.05, threads=1)[0], [fr, to],
                        np.float64, cost=skew)
    empty = S.sharded_map(lambda a: a, [np.zeros((0, 7))], np.float64)
    if rank == 0:
        ret["labels"] = labels
        ret["taus"] = taus
        ret["t"] = t
        ret["lf1"] = lf1
        ret["lf2"] = lf2
        ret["empty_shape"] = tuple(empty.shape)
        ret["ranges_skew"] = S.balanced_ranges(skew, world)
    dist.barrier()
    dist.destroy_process_group()


def test_world_size_2_gloo_shards_equal_single_process(oracle_mod):
    import torch.multiprocessing as mp
    q = W.random_states(1001, seed=21)
    free = q[make_world(oracle_mod.OracleWorld).is_valid_states(q).astype(bool)][:16]
    fr, to = W.touch_rays(free, 97, seed=22)
    mgr = mp.Manager()
    ret = mgr.dict()
    port = _free_port()
    ctx = mp.get_context("spawn")
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q, fr, to, ret)) for r in range(2)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(300)
        assert p.exitcode == 0
    ow = make_world(oracle_mod.OracleWorld)
    assert np.array_equal(ret["labels"], ow.is_valid_states(q))
    ref = ow.transit_states_to_with_contact(fr, to, 0.05)
    assert np.array_equal(ret["taus"][:, 0], ref[3]) and np.array_equal(ret["taus"][:, 1], ref[4])
    assert ret["t"] == 2.0                      # max over ranks
    assert ret["lf1"].shape == (1, 7) and np.array_equal(ret["lf1"], ref[0][:1])
    assert any(h == l for l, h in ret["ranges_skew"])      # one rank really had nothing to do
    assert ret["lf2"].shape == (len(fr), 7) and np.array_equal(ret["lf2"], ref[0])
    assert ret["empty_shape"][0] == 0
    # no process group: plain call
    assert np.array_equal(S.sharded_map(lambda x: ow.is_valid_states(x), [q], np.uint8), ret["labels"])
