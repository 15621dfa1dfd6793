"""Engine vs the reference's own statistics and M-step result for one golden iteration (debugging aid)."""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
from cnf2freq_b200.engine import Engine
import oracle_lib as ol
np.set_printoptions(precision=12, linewidth=220)
name, it = sys.argv[1], int(sys.argv[2])
g = dict(np.load(os.path.join(ROOT, "tests/golden", name + ".npz")))
b = lambda k: g[f"it{it}_before_{k}"]
eng = Engine(device=0)
cs = g["chromstarts"]; C = len(cs) - 1
eng.set_map(g["markerpos"], cs, g["genrec"]); eng.set_shared_rows(np.zeros(0, np.int32))
eng.set_pedigree(g["par"], g["empty"], b("founder"), g["gen"], g["descendants"], g["dous"])
al, su, hw = b("alleles").copy(), b("sure").copy(), b("haploweight").copy()
K, M = hw.shape
for c in range(C): eng.set_params(c, al, su, hw)
eng.set_mstep_inputs(b("relhaplo"), b("prioralleles"), b("priorsure"), g["hasprior"], g[f"it{it}_after_children"], g["empty"], g["descendants"])
eng.begin_iteration()
sf, ent, o1, o2 = b("scalars")
for c in range(C):
    eng.estep(c)
    st = dict(haplobase=np.zeros((K, M)), haplocount=np.zeros((K, M)), infprobs=np.zeros((K, M, 2, 2)), homozyg=np.zeros((K, M, 2)))
    eng.get_stats(c, **st)
    for k in st:
        ref = g[f"it{it}_c{c}_{k}"].reshape(st[k].shape)
        d = np.abs(st[k] - ref); bad = d > 1e-9 * np.abs(ref) + 1e-13
        print("chrom", c, k, int(bad.sum()), "of", bad.size, np.argwhere(bad)[:4].tolist())
    hit = eng.mstep(c, sf, ent)
    sf, o1, o2 = ol.oracle_scalefactor(sf, hit, len(g["dous"]), o1, o2)
for c in range(C): eng.get_params(c, al, su, hw)
for k, v in (("alleles", al), ("sure", su), ("haploweight", hw)):
    ref = g[f"it{it}_after_{k}"]
    bad = np.argwhere(np.abs(v.astype(float) - ref) > 1e-9)
    print(k, len(bad), bad[:8].tolist())
    for i in bad[:4]:
        i = tuple(i); print("   ", i, "got", v[i], "ref", ref[i], "before", b(k)[i], "alleles", b("alleles")[i[0], i[1]], "sure", b("sure")[i[0], i[1]], "prior", b("prioralleles")[i[0], i[1]],
                            "hasprior", g["hasprior"][i[0]], "infprobs", g[f"it{it}_c0_infprobs"].reshape(K, M, 2, 2)[i[0], i[1]])
