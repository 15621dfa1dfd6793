"""The N > 1 host path on CPU: two ranks over gloo shard the analysed individuals (cnf2freq_b200.shard),
each computes the statistics of its shard (with the CPU oracle standing in for the GPU engine -- the oracle
is the checker here, the thing under test is the sharding and the reduction), the shared rows are
sum-all-reduced, and the result equals the unsharded E-step."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import oracle_lib as ol
import randomped
from cases import STATS, assert_close
from cnf2freq_b200 import shard


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, maker, kw, outdir):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    pb = getattr(randomped, maker)(**kw)
    shards = shard.shard_dous(pb.par, pb.dous, world)
    shared = shard.shared_rows(pb.par, shards)
    mine = ol.Problem(pb.par, pb.empty, pb.founder, pb.gen, pb.descendants, shards[rank], pb.markerpos, pb.chromstarts,
                      pb.genrec, pb.alleles, pb.sure, pb.haploweight)
    out = ol.oracle_estep(mine)
    # what cnf2_estep's all-reduce hook does: sum the shared rows' statistics over ranks, in place
    for k in STATS:
        t = torch.from_numpy(np.ascontiguousarray(out[k][shared]))
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
        out[k][shared] = t.numpy()
    np.savez(os.path.join(outdir, f"rank{rank}.npz"), shard=shards[rank], shared=shared, **{k: out[k] for k in STATS},
             loglik=out["loglik"])
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("maker,kw", [
    ("f2_cross", dict(n_f2=7, n_chrom=2, mpc=11, seed=31)),
    ("fullsib", dict(n_fam=2, n_off=3, n_chrom=1, mpc=9, seed=32)),
])
def test_two_ranks_reduce_to_the_unsharded_estep(tmp_path, maker, kw):
    world = 2
    mp.spawn(_worker, args=(world, _free_port(), maker, kw, str(tmp_path)), nprocs=world, join=True)
    pb = getattr(randomped, maker)(**kw)
    whole = ol.oracle_estep(pb)
    parts = [dict(np.load(tmp_path / f"rank{r}.npz")) for r in range(world)]
    assert sorted(np.concatenate([p["shard"] for p in parts]).tolist()) == sorted(pb.dous.tolist())
    shared = parts[0]["shared"]
    assert np.array_equal(shared, parts[1]["shared"])
    touched = [set(i for d in p["shard"] for i in shard.focus_tree(pb.par, d)) for p in parts]
    assert set(shared.tolist()) == touched[0] & touched[1]
    for k in STATS:
        # shared rows: every rank holds the reduced sum; other rows: exactly one rank contributed
        for p in parts:
            assert_close(p[k][shared], whole[k][shared], f"shared:{k}", rtol=1e-12, atol=1e-13)
        only = np.zeros_like(whole[k])
        for p, tch in zip(parts, touched):
            rows = np.array(sorted(tch - set(shared.tolist())), np.int64)
            only[rows] = p[k][rows]
        only[shared] = parts[0][k][shared]
        assert_close(only, whole[k], f"merged:{k}", rtol=1e-12, atol=1e-13)
    pos = {int(d): i for i, d in enumerate(pb.dous)}
    for p in parts:
        assert_close(p["loglik"], whole["loglik"][[pos[int(d)] for d in p["shard"]]], "loglik", rtol=1e-12)


def test_sibships_stay_together():
    pb = randomped.fullsib(n_fam=4, n_off=3, n_chrom=1, mpc=2, seed=5)
    shards = shard.shard_dous(pb.par, pb.dous, 2)
    assert sorted(len(s) for s in shards) == [6, 6]
    for s in shards:
        assert len({tuple(pb.par[d]) for d in s}) == 2
    assert len(shard.shared_rows(pb.par, shards)) == 0
