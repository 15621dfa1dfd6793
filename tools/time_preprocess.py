"""Times cnf2_preprocess (row f3) on an F2 cross as the readers leave it: ranked (factorised ranking + literal sums for the
markers that can hold a chromosome's maximum) against CNF2_PRE_LITERAL=1 (the literal sums for every cell).
usage: python tools/time_preprocess.py N_F2 N_CHROM MARKERS_PER_CHROM [literal]"""
import os
import sys
import time
import zlib

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from cnf2freq_b200 import synth  # noqa: E402
from cnf2freq_b200.engine import Engine  # noqa: E402

n, C, mpc = (int(x) for x in sys.argv[1:4])
if len(sys.argv) > 4 and sys.argv[4] == "literal":
    os.environ["CNF2_PRE_LITERAL"] = "1"
sim = synth.simulate_f2(n_f2=n, n_chrom=C, markers_per_chrom=mpc, spacing_cM=0.1, missing=0.2, seed=3)
M = C * mpc
K = 3 + 3 * n  # 0, A, B, then per F2: itself and its two implicit (empty) F1 parents (readalphaped, cnF2freq.cpp:6515-6527)
par = np.zeros((K, 2), np.int32)
empty = np.zeros(K, np.uint8)
exists = np.ones(K, np.uint8)
exists[0] = 0
empty[0] = 1
al = np.zeros((K, M, 2), np.uint8)
tok = sim["tokens"]


def geno(t):
    a = np.zeros(t.shape + (2,), np.uint8)
    a[t == 0] = (1, 1)
    a[t == 1] = (1, 2)
    a[t == 2] = (2, 2)
    return a


al[1], al[2] = geno(tok[0]), geno(tok[1])
for k in range(n):
    me = 3 + 3 * k
    par[me] = (me + 1, me + 2)
    par[me + 1] = par[me + 2] = (1, 2)
    empty[me + 1] = empty[me + 2] = 1
    al[me] = geno(tok[2 + k])
su = np.where(al != 0, 0.02, 0.0)
hw = np.full((K, M), 0.5)
eng = Engine(device=0)
eng.set_map(sim["markerpos"], sim["chromstarts"], np.full(3, -1.0 / 50))
t0 = time.perf_counter()
nfix = [0]


def log(st, line):
    nfix[0] += line.startswith("Fixing point")


eng.preprocess(par, exists, empty, al, su, hw, log=log)
dt = time.perf_counter() - t0
print(f"{'literal' if os.environ.get('CNF2_PRE_LITERAL') == '1' else 'ranked'}: {K} individuals x {M} markers ({C} x {mpc}): cnf2_preprocess {dt:.2f} s, "
      f"literal sums for {eng.preprocess_literal_cells()} of {K * M} cells, {nfix[0]} phases locked, "
      f"checksum of the locked phases {zlib.crc32(hw.tobytes()):08x}, of the genotypes {zlib.crc32(al.tobytes()):08x}")
eng.close()
