"""cnf2_preprocess (row f3: postmarkerdata with its emission scans on the GPU) against the CPU oracle -- which is pinned on
the reference binary's own post-postmarkerdata state (tests/test_preprocess_oracle.py) -- on random pedigrees as the readers
would leave them: Mendelian genotypes with missing and half-missing entries, injected impossible genotypes (cleared by
fixparents, :1432-1437), untyped `empty` individuals, markersure 0.02 / 0, haploweight 0.5."""
import numpy as np
import pytest

import oracle_lib as ol
import randomped

pytestmark = pytest.mark.gpu


def raw_inputs(pb, seed, impossible=0.0):
    """Replaces the random E-step parameters of a randomped problem by reader-like raw data: genotypes dropped down the
    pedigree from founder alleles, `missing` of them unknown."""
    rng = np.random.default_rng(seed)
    K, M = pb.K, pb.M
    hap = np.zeros((K, M, 2), np.uint8)
    order = np.argsort(pb.gen, kind="stable")
    for i in order:
        if i == 0:
            continue
        for s in range(2):
            p = pb.par[i, s]
            if p and hap[p].any():
                pick = rng.integers(0, 2, M)
                hap[i, :, s] = np.where(pick == 0, hap[p, :, 0], hap[p, :, 1])
            else:
                hap[i, :, s] = rng.integers(1, 3, M)
    al = hap.copy()
    al[rng.random((K, M)) < 0.25] = 0
    al[rng.random((K, M, 2)) < 0.05] = 0
    if impossible:
        bad = rng.random((K, M)) < impossible
        al[bad] = 3 - np.where(al[bad] == 0, 1, al[bad])  # flip alleles: some of these contradict the relatives
    al[pb.empty == 1] = 0
    al[0] = 0
    su = np.where(al != 0, 0.02, 0.0)
    if impossible:
        su[...] = 0.0  # Gigi-style data (markersure 0, :7497-7498): a Mendelian inconsistency has no admissible interpretation
    hw = np.full((K, M), 0.5)
    exists = np.ones(K, np.uint8)
    exists[0] = 0
    return exists, al, su, hw


@pytest.mark.parametrize("maker,kw,impossible,selfing", [
    ("f2_cross", dict(n_f2=7, n_chrom=2, mpc=11, seed=81), 0.0, False),
    ("fullsib", dict(n_fam=2, n_off=3, n_chrom=1, mpc=13, seed=82), 0.0, False),
    ("tangled", dict(n_chrom=2, mpc=9, seed=83), 0.0, False),
    ("fullsib", dict(n_fam=2, n_off=4, n_chrom=2, mpc=8, seed=84), 0.08, False),
    ("tangled", dict(n_chrom=1, mpc=21, seed=85), 0.1, False),
    # row g1: the scans over the 256 states of SELFING = true (engine and oracle both built / compiled for it).  No injected
    # impossible genotypes here: with markersure 0 everywhere the restated postmarkerdata loop then clears and re-infers the same
    # genotypes for ever under SELFING (any stays > 0, :3364) -- an input the reference's loop is not known to leave either
    ("f2_cross", dict(n_f2=5, n_chrom=2, mpc=9, seed=86), 0.0, True),
    ("tangled", dict(n_chrom=1, mpc=13, seed=87), 0.0, True),
])
def test_preprocess_matches_the_oracle(maker, kw, impossible, selfing):
    from cnf2freq_b200.engine import Engine
    pb = getattr(randomped, maker)(**kw)
    exists, al, su, hw = raw_inputs(pb, kw["seed"], impossible)
    ref = ol.oracle_preprocess(pb.chromstarts, pb.par, exists, pb.empty, al, su, hw, selfing=selfing)
    eng = Engine(device=0, selfing=selfing)
    try:
        eng.set_map(pb.markerpos, pb.chromstarts, pb.genrec)
        a, s, h = al.copy(), su.copy(), hw.copy()
        lines = []
        desc, children, founder = eng.preprocess(pb.par, exists, pb.empty, a, s, h, log=lambda st, l: lines.append((st, l)))
    finally:
        eng.close()
    assert np.array_equal(a, ref["alleles"]), np.argwhere(a != ref["alleles"])[:5]
    assert np.allclose(s, ref["sure"], rtol=1e-12, atol=0, equal_nan=True)  # dosureval(0, .) is NaN in the reference too
    assert np.array_equal(h, ref["haploweight"]), np.argwhere(h != ref["haploweight"])[:5]
    assert np.array_equal(desc, ref["descendants"]) and np.array_equal(children, ref["children"]) and np.array_equal(founder, ref["founder"])
    assert sum(1 for st, l in lines if l.startswith("Correction at")) == ref["corrections"]
    assert sum(1 for st, l in lines if l.startswith("Fixing point")) == int((h != hw).sum())
    assert sum(1 for st, l in lines if l.startswith("Clearing")) == ref["cleared"]
    if impossible:
        assert ref["cleared"] > 0, "no impossible genotype in this draw"


@pytest.mark.parametrize("maker,kw,impossible", [
    ("fullsib", dict(n_fam=3, n_off=4, n_chrom=2, mpc=40, seed=91), 0.0),
    ("f2_cross", dict(n_f2=12, n_chrom=3, mpc=50, seed=92), 0.0),   # few distinct genotype patterns: many exact ties per chromosome
    ("tangled", dict(n_chrom=2, mpc=35, seed=93), 0.05),
])
def test_factorised_ranking_gives_the_literal_decisions(monkeypatch, maker, kw, impossible):
    """addvariance's sums only decide where lockhaplos locks the phase: cnf2_preprocess ranks the markers with the factorised
    form and walks the literal sums for the few markers that can hold the maximum (include/cnf2_engine.h,
    cnf2_preprocess_literal_cells).  Same locks, same lines as with the literal kernel on every cell (CNF2_PRE_LITERAL=1)."""
    from cnf2freq_b200.engine import Engine
    pb = getattr(randomped, maker)(**kw)
    exists, al, su, hw = raw_inputs(pb, kw["seed"], impossible)
    runs = {}
    for mode in ("ranked", "literal"):
        if mode == "literal":
            monkeypatch.setenv("CNF2_PRE_LITERAL", "1")
        else:
            monkeypatch.delenv("CNF2_PRE_LITERAL", raising=False)
        eng = Engine(device=0)
        try:
            eng.set_map(pb.markerpos, pb.chromstarts, pb.genrec)
            a, s, h = al.copy(), su.copy(), hw.copy()
            lines = []
            out = eng.preprocess(pb.par, exists, pb.empty, a, s, h, log=lambda st, l: lines.append((st, l)))
            runs[mode] = (a, s, h, out, lines, eng.preprocess_literal_cells())
        finally:
            eng.close()
    r, l = runs["ranked"], runs["literal"]
    assert np.array_equal(r[0], l[0]) and np.array_equal(r[1], l[1], equal_nan=True) and np.array_equal(r[2], l[2])
    assert all(np.array_equal(x, y) for x, y in zip(r[3], l[3]))
    assert r[4] == l[4]
    assert sum(1 for st, line in r[4] if line.startswith("Fixing point")) > 0
    cells = pb.K * pb.M
    assert l[5] == cells
    print(f"{maker}: literal sums walked for {r[5]} of {cells} cells")
    assert 0 < r[5] <= 0.2 * cells
