import sys
sys.path.insert(0, "tests"); sys.path.insert(0, ".")
import numpy as np
import oracle_lib as ol, randomped
from cases import engine_problem
from cnf2freq_b200.engine import Engine
kw = dict(n_f2=9, n_chrom=2, mpc=21, seed=61)
pb = randomped.f2_cross(**kw)
rng = np.random.default_rng(kw["seed"])
K, M, C = pb.K, pb.M, pb.C
relhaplo = rng.uniform(0.6, 1.0, (K, M))
hasprior = (rng.random(K) < 0.7).astype(np.uint8)
prioral = pb.alleles.copy()
prioral[rng.random((K, M)) < 0.2] = 0
priorsure = np.where(rng.random((K, M, 2)) < 0.1, 0.0, rng.uniform(0.0, 0.3, (K, M, 2)))
children = rng.integers(0, 4, K).astype(np.int32)
eng = Engine(device=0)
engine_problem(eng, pb, shared=np.zeros(0, np.int32))
al, su, hw = pb.alleles.copy(), pb.sure.copy(), pb.haploweight.copy()
for c in range(C):
    eng.set_params(c, al, su, hw)
eng.set_mstep_inputs(relhaplo, prioral, priorsure, hasprior, children, pb.empty, pb.descendants)
eng.begin_iteration()
active = np.ones(K, np.uint8); active[0] = 0
c = 0
eng.estep(c)
st = dict(haplobase=np.zeros((K, M)), haplocount=np.zeros((K, M)), infprobs=np.zeros((K, M, 2, 2)), homozyg=np.zeros((K, M, 2)))
eng.get_stats(c, **st)
r = ol.oracle_mstep(pb.chromstarts, active, pb.empty, children, pb.descendants, al, su, hw, relhaplo, hasprior, prioral, priorsure,
                    st["haplobase"], st["haplocount"], st["infprobs"], st["homozyg"], 0.013, 1.0)
hit = eng.mstep(c, 0.013, 1.0)
eng.get_params(c, al, su, hw)
print("hit", hit, r["hitnnn"])
bad = np.argwhere(np.abs(su - r["sure"]) > 1e-9 * np.abs(r["sure"]) + 1e-12)
print(len(bad), "sure mismatches;", np.sum(np.abs(hw - r["haploweight"]) > 1e-9), "hw mismatches")
for i, m, s in bad[:8]:
    print(i, m, s, "before al", pb.alleles[i, m], "sure", pb.sure[i, m], "ip", st["infprobs"][i, m, s], "prior", hasprior[i], prioral[i, m], priorsure[i, m],
          "children", children[i], "empty", pb.empty[i], "| dev", al[i, m], su[i, m, s], "oracle", r["alleles"][i, m], r["sure"][i, m, s])
