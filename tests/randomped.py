"""Seeded random pedigrees + parameters as flat arrays (test inputs; sizes the CPU oracle finishes in seconds)."""
import numpy as np

import oracle_lib as ol


def _finish(par, empty, gen, dous, n_chrom, mpc, rng, spacing, missing, samepos=False):
    K = len(par)
    par = np.asarray(par, np.int32)
    M = n_chrom * mpc
    pos = np.tile(np.cumsum(rng.uniform(0.2 * spacing, 2 * spacing, mpc)), n_chrom)
    if samepos and mpc > 3:
        pos[2] = pos[1]  # two markers at the same position: the transition is skipped (cnF2freq.cpp:2273)
    cs = np.arange(n_chrom + 1, dtype=np.uint32) * mpc
    alleles = rng.integers(1, 3, size=(K, M, 2)).astype(np.uint8)
    alleles[rng.random((K, M)) < missing] = 0
    alleles[rng.random((K, M, 2)) < missing / 4] = 0  # half-missing genotypes
    sure = np.where(rng.random((K, M, 2)) < 0.5, 0.02, rng.uniform(0, 0.3, (K, M, 2)))
    sure[rng.random((K, M)) < 0.3] = 0.02           # equal sureness: the homozygous rule applies
    sure[rng.random((K, M, 2)) < 0.05] = 0.0
    sure[alleles == 0] = 0.0
    hw = rng.uniform(0.05, 0.95, (K, M))
    hw[rng.random((K, M)) < 0.3] = 0.5
    hw[rng.random((K, M)) < 0.05] = 0.0               # locked phase
    hw[rng.random((K, M)) < 0.05] = 1.0
    empty = np.asarray(empty, np.uint8)
    alleles[empty == 1] = 0
    sure[empty == 1] = 0
    alleles[0] = 0
    descendants = np.zeros(K, np.int32)
    for d in dous:  # any positive weights do; the reference counts descendants in postmarkerdata
        descendants[d] = 1 + (d % 3)
    founder = np.zeros(K, np.uint8)
    pb = ol.Problem(par, empty, founder, np.asarray(gen, np.int32), descendants, np.asarray(dous, np.int32), pos, cs,
                    np.full(3, -1.0 / 50), alleles, sure, hw)
    pb.founder = ol.steady_founder(pb)
    return pb


def f2_cross(n_f2, n_chrom, mpc, seed, spacing=1.0, missing=0.2):
    """Founders A, B; every F2 has two implicit, empty F1 parents (readalphaped, cnF2freq.cpp:6515-6527)."""
    rng = np.random.default_rng(seed)
    par, empty, gen, dous = [(0, 0), (0, 0), (0, 0)], [1, 0, 0], [0, 0, 0], []
    for k in range(n_f2):
        base = len(par)
        par += [(base + 1, base + 2), (1, 2), (1, 2)]
        empty += [0, 1, 1]
        gen += [2, 1, 1]
        dous.append(base)
    return _finish(par, empty, gen, dous, n_chrom, mpc, rng, spacing, missing, samepos=True)


def fullsib(n_fam, n_off, n_chrom, mpc, seed, spacing=1.0, missing=0.1):
    """Three generations, no inbreeding: all 7 slots distinct, 128 orientations, 8 shift modes."""
    rng = np.random.default_rng(seed)
    par, empty, gen, dous = [(0, 0)], [1], [0], []
    for f in range(n_fam):
        b = len(par)
        par += [(0, 0)] * 4 + [(b, b + 1), (b + 2, b + 3)]
        empty += [0] * 6
        gen += [0] * 4 + [1, 1]
        for k in range(n_off):
            dous.append(len(par))
            par.append((b + 4, b + 5))
            empty.append(0)
            gen.append(2)
    return _finish(par, empty, gen, dous, n_chrom, mpc, rng, spacing, missing)


def tangled(n_chrom, mpc, seed, spacing=1.0, missing=0.15):
    """Missing parents, a shared grandparent on both sides (inbreeding), a half-sib sire, gen < 2 analysed."""
    rng = np.random.default_rng(seed)
    #        0       1 G1    2 G2    3 G3     4 P1     5 P2     6 X       7 Y       8 Z       9 H1     10 W      11 E(empty) 12 V
    par = [(0, 0), (0, 0), (0, 0), (0, 0), (1, 2), (1, 3), (4, 5), (4, 0), (0, 5), (4, 3), (6, 9), (1, 2), (11, 5)]
    empty = [1, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 1, 0]
    gen = [0, 0, 0, 0, 1, 1, 2, 2, 2, 1, 3, 1, 2]
    dous = [6, 7, 8, 9, 10, 12, 4, 1]
    return _finish(par, empty, gen, dous, n_chrom, mpc, rng, spacing, missing, samepos=True)
