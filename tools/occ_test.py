"""Throughput of the E-step kernel against resident warps per SM (whole waves only, so the tail does not blur it)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from cnf2freq_b200 import synth
from cnf2freq_b200.engine import Engine
waves = 2
for per_sm in [int(x) for x in (sys.argv[1:] or "4 7 9 10 11 12".split())]:
    n = 148 * per_sm * waves
    d = synth.f2_flat(n, 1, 1000, spacing_cM=0.1, missing=0.2, seed=3)
    eng = Engine(device=0, ctas_per_sm=per_sm)
    eng.set_map(d["markerpos"], d["chromstarts"], d["genrec"]); eng.set_shared_rows(d["shared"])
    eng.set_pedigree(d["par"], d["empty"], d["founder"], d["gen"], d["descendants"], d["dous"])
    eng.set_params(0, d["alleles"], d["sure"], d["haploweight"])
    for _ in range(2): eng.estep(0)
    eng.sync(); ms = []
    for _ in range(3):
        eng.estep(0); ms.append(eng.last_estep_ms())
    print(per_sm, "warps/SM:", n, "units in", min(ms), "ms ->", n * 1000 * 64 / min(ms) / 1e6, "G states/s; ms per wave", min(ms) / waves, flush=True)
    eng.close()
