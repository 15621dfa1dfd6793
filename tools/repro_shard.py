import sys, os
sys.path.insert(0, "tests"); sys.path.insert(0, ".")
import numpy as np
import oracle_lib as ol, randomped
from cases import STATS, assert_close, engine_problem
from cnf2freq_b200 import shard
from cnf2freq_b200.engine import Engine
pb = randomped.fullsib(n_fam=2, n_off=3, n_chrom=1, mpc=17, seed=42)
shards = shard.shard_dous(pb.par, pb.dous, 2)
shared = shard.shared_rows(pb.par, shards)
print("par", pb.par.tolist()); print("dous", pb.dous, "shards", shards, "shared", shared)
eng = Engine(device=0)
for r in range(2):
    for sh in (None, shared):
        engine_problem(eng, pb, dous=shards[r], all_dous=pb.dous, shared=sh)
        out = eng.estep_all(pb.alleles, pb.sure, pb.haploweight)
        mine = ol.Problem(pb.par, pb.empty, pb.founder, pb.gen, pb.descendants, shards[r], pb.markerpos, pb.chromstarts, pb.genrec, pb.alleles, pb.sure, pb.haploweight)
        mine.founder = ol.steady_founder(pb) if hasattr(ol, "steady_founder") else mine.founder
        ref = ol.oracle_estep(mine)
        for k in STATS:
            try:
                assert_close(out[k], ref[k], f"rank{r} shared={'set' if sh is not None else 'unset'} {k}", rtol=1e-9, atol=1e-13)
            except AssertionError as ex:
                print("FAIL", ex)
print("done")
