"""The N > 1 path on real GPUs: two ranks (one engine per GPU, NCCL), analysed individuals sharded with
cnf2freq_b200.shard, the shared statistics rows summed by the engine's all-reduce hook on the device;
the merged result equals the CPU oracle's unsharded E-step (1e-9 relative).  Skipped with fewer than 2 GPUs."""
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
    np.savez(os.path.join(outdir, f"rank{rank}.npz"), shard=shards[rank], shared=shared, **{k: out[k] for k in STATS},
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
    for k in STATS:
        for p in parts:  # every rank holds the reduced sum of the shared rows
            assert_close(p[k][shared], whole[k][shared], f"shared:{k}", rtol=RTOL, atol=1e-13)
        merged = np.zeros_like(whole[k])
        for p, tch in zip(parts, touched):
            rows = np.array(sorted(tch - set(shared.tolist())), np.int64)
            merged[rows] = p[k][rows]
        merged[shared] = parts[0][k][shared]
        assert_close(merged, whole[k], f"merged:{k}", rtol=RTOL, atol=1e-13)
    pos = {int(d): i for i, d in enumerate(pb.dous)}
    for p in parts:
        assert_close(p["loglik"], whole["loglik"][[pos[int(d)] for d in p["shard"]]], "loglik", rtol=RTOL, atol=1e-9)
