"""The N > 1 path on real GPUs: two ranks (one engine per GPU, NCCL), analysed individuals sharded with
cnf2freq_b200.shard, the shared statistics rows summed by the engine's all-reduce hook on the device;
the merged result equals the CPU oracle's unsharded E-step (1e-9 relative); then whole EM passes (E-step + device
M-step per chromosome) on both ranks against the oracle's M-step on the unsharded statistics.  Skipped with fewer than
2 GPUs."""
import os
import socket

import numpy as np
import pytest
import torch

import oracle_lib as ol
import randomped
from cases import RTOL, STATS, assert_close, engine_problem
from cnf2freq_b200 import shard

pytestmark = pytest.mark.gpu


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, maker, kw, outdir):
    import torch.distributed as dist
    from cnf2freq_b200.engine import Engine
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    pb = getattr(randomped, maker)(**kw)
    shards = shard.shard_dous(pb.par, pb.dous, world)
    shared = shard.shared_rows(pb.par, shards)
    eng = Engine(device=rank)

    class DevBuf:
        def __init__(self, ptr, n):
            self.__cuda_array_interface__ = {"shape": (n,), "typestr": "<f8", "data": (ptr, False), "version": 2}

    def allreduce(ptr, count, stream):
        with torch.cuda.stream(torch.cuda.ExternalStream(stream)):
            t = torch.as_tensor(DevBuf(ptr, count), device=torch.device("cuda", rank))
            dist.all_reduce(t, op=dist.ReduceOp.SUM)

    eng.set_allreduce(allreduce)
    engine_problem(eng, pb, dous=shards[rank], all_dous=pb.dous, shared=shared)
    out = eng.estep_all(pb.alleles, pb.sure, pb.haploweight)
    eng.sync()
    np.savez(os.path.join(outdir, f"rank{rank}.npz"), shard=shards[rank], shared=shared, **{k: out[k] for k in STATS + ("negshift",)},
             loglik=out["loglik"])
    eng.close()
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("maker,kw", [
    ("f2_cross", dict(n_f2=9, n_chrom=2, mpc=23, seed=41)),
    ("fullsib", dict(n_fam=2, n_off=3, n_chrom=1, mpc=17, seed=42)),
])
def test_two_gpus_reduce_to_the_unsharded_estep(tmp_path, maker, kw):
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    import torch.multiprocessing as mp
    world = 2
    mp.spawn(_worker, args=(world, _free_port(), maker, kw, str(tmp_path)), nprocs=world, join=True)
    pb = getattr(randomped, maker)(**kw)
    whole = ol.oracle_estep(pb)
    parts = [dict(np.load(tmp_path / f"rank{r}.npz")) for r in range(world)]
    shared = parts[0]["shared"]
    touched = [set(i for d in p["shard"] for i in shard.focus_tree(pb.par, d)) for p in parts]
    for k in STATS + ("negshift",):
        for p in parts:  # every rank holds the reduced sum of the shared rows
            assert_close(p[k][shared], whole[k][shared], f"shared:{k}", rtol=RTOL, atol=1e-11 if k == "negshift" else 1e-13)
        merged = np.zeros_like(whole[k])
        for p, tch in zip(parts, touched):
            rows = np.array(sorted(tch - set(shared.tolist())), np.int64)
            merged[rows] = p[k][rows]
        merged[shared] = parts[0][k][shared]
        assert_close(merged, whole[k], f"merged:{k}", rtol=RTOL, atol=1e-11 if k == "negshift" else 1e-13)
    pos = {int(d): i for i, d in enumerate(pb.dous)}
    for p in parts:
        assert_close(p["loglik"], whole["loglik"][[pos[int(d)] for d in p["shard"]]], "loglik", rtol=RTOL, atol=1e-9)


def _mstep_inputs(pb, seed):
    rng = np.random.default_rng(seed)
    K, M = pb.K, pb.M
    relhaplo = rng.uniform(0.6, 1.0, (K, M))
    hasprior = (rng.random(K) < 0.7).astype(np.uint8)
    prioral = pb.alleles.copy()
    prioral[rng.random((K, M)) < 0.2] = 0
    priorsure = rng.uniform(0.0, 0.3, (K, M, 2))
    children = rng.integers(0, 4, K).astype(np.int32)
    return relhaplo, prioral, priorsure, hasprior, children


SF, ENT = 0.013, 1.0


def _em_worker(rank, world, port, maker, kw, outdir):
    import torch.distributed as dist
    from cnf2freq_b200.engine import Engine
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    pb = getattr(randomped, maker)(**kw)
    shards = shard.shard_dous(pb.par, pb.dous, world)
    shared = shard.shared_rows(pb.par, shards)
    eng = Engine(device=rank)

    class DevBuf:
        def __init__(self, ptr, n):
            self.__cuda_array_interface__ = {"shape": (n,), "typestr": "<f8", "data": (ptr, False), "version": 2}

    def allreduce(ptr, count, stream):
        with torch.cuda.stream(torch.cuda.ExternalStream(stream)):
            t = torch.as_tensor(DevBuf(ptr, count), device=torch.device("cuda", rank))
            dist.all_reduce(t, op=dist.ReduceOp.SUM)

    eng.set_allreduce(allreduce)
    engine_problem(eng, pb, dous=shards[rank], all_dous=pb.dous, shared=shared)
    al, su, hw = pb.alleles.copy(), pb.sure.copy(), pb.haploweight.copy()
    for c in range(pb.C):
        eng.set_params(c, al, su, hw)
    relhaplo, prioral, priorsure, hasprior, children = _mstep_inputs(pb, kw["seed"])
    eng.set_mstep_inputs(relhaplo, prioral, priorsure, hasprior, children, pb.empty, pb.descendants)
    eng.begin_iteration()
    hits = []
    for c in range(pb.C):
        eng.estep(c)
        hits.append(eng.mstep(c, SF, ENT, split=True))
    for c in range(pb.C):
        eng.get_params(c, al, su, hw)
    np.savez(os.path.join(outdir, f"em{rank}.npz"), shard=shards[rank], shared=shared, alleles=al, sure=su, haploweight=hw,
             hits=np.array(hits, np.int64))
    eng.close()
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("maker,kw", [
    ("f2_cross", dict(n_f2=9, n_chrom=2, mpc=19, seed=71)),
    ("tangled", dict(n_chrom=2, mpc=15, seed=72)),
])
def test_two_gpus_em_pass_equals_the_unsharded_mstep(tmp_path, maker, kw):
    """Every rank updates its own rows and its copy of the shared rows from the all-reduced statistics: the shared
    rows end up identical on both ranks, the merged parameters equal the oracle's M-step on the whole pedigree, and
    the split hit counters add up to the reference's hitnnn."""
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    import torch.multiprocessing as mp
    from test_mstep_gpu import ATOL, MAX_MASKED, stationary
    from test_mstep_gpu import RTOL as MRTOL
    world = 2
    mp.spawn(_em_worker, args=(world, _free_port(), maker, kw, str(tmp_path)), nprocs=world, join=True)
    pb = getattr(randomped, maker)(**kw)
    K, M, C = pb.K, pb.M, pb.C
    relhaplo, prioral, priorsure, hasprior, children = _mstep_inputs(pb, kw["seed"])
    whole = ol.oracle_estep(pb)  # an M-step pass only rewrites chromosomes already analysed: one E-step serves all passes
    if any(np.isnan(whole[k]).any() for k in STATS):
        pytest.skip("impossible unit in this random draw (NaN statistics)")
    active = np.ones(K, np.uint8)
    active[0] = 0
    hb, hc = np.zeros((K, M)), np.zeros((K, M))
    ral, rsu, rhw = pb.alleles.copy(), pb.sure.copy(), pb.haploweight.copy()
    tie = np.zeros((K, M), bool)
    ref_hits = []
    for c in range(C):
        cols = slice(int(pb.chromstarts[c]), int(pb.chromstarts[c + 1]))
        hb[:, cols], hc[:, cols] = whole["haplobase"][:, cols], whole["haplocount"][:, cols]
        ip, hz = np.zeros((K, M, 2, 2)), np.zeros((K, M, 2))
        ip[:, cols], hz[:, cols] = whole["infprobs"][:, cols], whole["homozyg"][:, cols]
        tie |= stationary(ral, rsu, ip, prioral, hasprior)
        r = ol.oracle_mstep(pb.chromstarts, active, pb.empty, children, pb.descendants, ral, rsu, rhw, relhaplo, hasprior, prioral,
                            priorsure, hb, hc, ip, hz, SF, ENT)
        ral, rsu, rhw, hb, hc = r["alleles"], r["sure"], r["haploweight"], r["haplobase"], r["haplocount"]
        ref_hits.append(r["hitnnn"])
    parts = [dict(np.load(tmp_path / f"em{r}.npz")) for r in range(world)]
    shared = parts[0]["shared"]
    for k in ("alleles", "sure", "haploweight"):
        assert np.array_equal(parts[0][k][shared], parts[1][k][shared]), f"shared rows differ between ranks: {k}"
    touched = [set(i for d in p["shard"] for i in shard.focus_tree(pb.par, d)) for p in parts]
    merged = dict(alleles=pb.alleles.copy(), sure=pb.sure.copy(), haploweight=pb.haploweight.copy())
    for p, tch in zip(parts, touched):
        rows = np.array(sorted(tch), np.int64)
        for k in merged:
            merged[k][rows] = p[k][rows]
    ok = ~tie
    print(f"masked as exact ties: {int(tie.sum())} of {tie.size} cells ({tie.mean():.4%})")
    assert tie.mean() < MAX_MASKED, f"{tie.mean():.3%} of the cells masked as ties"
    assert np.array_equal(merged["alleles"][ok], ral[ok])
    assert_close(merged["sure"][ok], rsu[ok], "markersure", rtol=MRTOL, atol=ATOL)
    assert_close(merged["haploweight"][ok], rhw[ok], "haploweight", rtol=MRTOL, atol=ATOL)
    for c in range(C):
        total = sum(int(p["hits"][c][0] - p["hits"][c][1]) for p in parts) + int(parts[0]["hits"][c][1])
        assert int(parts[0]["hits"][c][1]) == int(parts[1]["hits"][c][1])
        assert abs(total - ref_hits[c]) <= 2 * int(tie.sum()), (c, total, ref_hits[c])
