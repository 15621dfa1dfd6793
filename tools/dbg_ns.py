"""Prints the engine's phase-inversion scores next to the reference's for one golden iteration (near-tie analysis)."""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
from cnf2freq_b200.engine import Engine
from test_em_loop_gpu import setup_iteration
np.set_printoptions(precision=17, linewidth=200)
name, it, ind, a, b = sys.argv[1], int(sys.argv[2]), int(sys.argv[3]), int(sys.argv[4]), int(sys.argv[5])
g = dict(np.load(os.path.join(ROOT, "tests/golden", name + ".npz")))
eng = Engine(device=0)
bufs = setup_iteration(eng, g, it, first=True)
K, M = bufs[2].shape
ns = np.zeros((K, M))
for rep in range(3):
    eng.estep(0)
    eng.get_negshift(0, ns)
    print("ours", ns[ind, a:b])
print("ref ", g[f"it{it}_negshift"][ind, a:b])
print("diff", ns[ind, a:b] - g[f"it{it}_negshift"][ind, a:b])
