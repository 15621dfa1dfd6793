import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import oracle_lib as ol, randomped
from test_preprocess_gpu import raw_inputs
from cnf2freq_b200.engine import Engine
kw = dict(n_fam=2, n_off=4, n_chrom=2, mpc=8, seed=84)
pb = randomped.fullsib(**kw)
exists, al, su, hw = raw_inputs(pb, 84, 0.08)
ref = ol.oracle_preprocess(pb.chromstarts, pb.par, exists, pb.empty, al, su, hw)
eng = Engine(device=0)
eng.set_map(pb.markerpos, pb.chromstarts, pb.genrec)
a, s, h = al.copy(), su.copy(), hw.copy()
lines = []
eng.preprocess(pb.par, exists, pb.empty, a, s, h, log=lambda st, l: lines.append(l))
bad = np.argwhere(~np.isclose(s, ref["sure"], rtol=1e-12, atol=0, equal_nan=True))
print(len(bad), bad[:10].tolist())
for x in bad[:6]:
    x = tuple(x); print(x, "engine", repr(s[x]), "oracle", repr(ref["sure"][x]), "alleles", a[x[0], x[1]], ref["alleles"][x[0], x[1]], "input", al[x[0], x[1]], "par", pb.par[x[0]])
print([l for l in lines if l.startswith(("Clearing", "Number"))][:20])
