"""Synthetic pedigrees of the shapes BASELINE.json names, in the reference's PlantImpute text formats
(.map/.ped/.gen, readers at /root/reference/cnF2freq.cpp:6495-6685) and as flat arrays.

Generators follow SURVEY.md section 8(d): Haldane crossovers r = (1 - exp(-0.02 d)) / 2 with d in cM
(the model the reference's transition uses, cnF2freq.cpp:2286 with genrec = -1/50, :7927-7943).
"""
import numpy as np


def make_map(n_chrom, markers_per_chrom, spacing_cM):
    """Marker positions restart at 0 per chromosome: that is what splits chromosomes (cnF2freq.cpp:6676-6679)."""
    pos = np.tile(np.arange(markers_per_chrom, dtype=np.float64) * spacing_cM, n_chrom)
    chromstarts = np.arange(n_chrom + 1, dtype=np.uint32) * markers_per_chrom
    return pos, chromstarts


def _gametes(rng, hap_pair, pos, chromstarts, n):
    """n gametes from a parent with haplotypes hap_pair [2][M] (alleles 1/2)."""
    M = pos.shape[0]
    d = np.diff(pos, prepend=pos[0])
    r = 0.5 * (1.0 - np.exp(-0.02 * np.abs(d)))
    r[np.asarray(chromstarts[:-1], dtype=np.int64)] = 0.5  # free recombination between chromosomes
    cross = rng.random((n, M)) < r[None, :]
    strand = np.cumsum(cross, axis=1) & 1
    strand ^= rng.integers(0, 2, size=(n, 1))
    return np.where(strand == 0, hap_pair[0][None, :], hap_pair[1][None, :]).astype(np.uint8)


def _tokens(geno_pair, rng, missing):
    """PlantImpute .gen tokens: 0 -> (1,1), 1 -> (1,2), 2 -> (2,2), 9 -> unknown (cnF2freq.cpp:6568-6582)."""
    tok = (geno_pair[0] == 2).astype(np.int8) + (geno_pair[1] == 2).astype(np.int8)
    if missing > 0:
        tok = np.where(rng.random(tok.shape) < missing, 9, tok)
    return tok.astype(np.int8)


def simulate_f2(n_f2, n_chrom, markers_per_chrom, spacing_cM=0.1, missing=0.2, seed=3):
    """cfg3: founders A (all 2) and B (all 0), n_f2 lines ``I<k> A B 2``."""
    rng = np.random.default_rng(seed)
    pos, cs = make_map(n_chrom, markers_per_chrom, spacing_cM)
    M = pos.shape[0]
    f1 = np.stack([np.full(M, 2, np.uint8), np.full(M, 1, np.uint8)])
    g0 = _gametes(rng, f1, pos, cs, n_f2)
    g1 = _gametes(rng, f1, pos, cs, n_f2)
    names = ["A", "B"] + [f"I{k}" for k in range(n_f2)]
    ped = [("A", "0", "0", 0), ("B", "0", "0", 0)] + [(f"I{k}", "A", "B", 2) for k in range(n_f2)]
    tokens = np.concatenate([np.full((1, M), 2, np.int8), np.full((1, M), 0, np.int8), _tokens((g0, g1), rng, missing)])
    return dict(names=names, ped=ped, tokens=tokens, markerpos=pos, chromstarts=cs)


def simulate_fullsib(n_fam, n_off, n_chrom, markers_per_chrom, spacing_cM=0.02, missing=0.05, seed=4):
    """cfg4: per family 4 founders (gen 0, i.i.d. alleles), 2 parents (gen 1), n_off offspring (gen 2)."""
    rng = np.random.default_rng(seed)
    pos, cs = make_map(n_chrom, markers_per_chrom, spacing_cM)
    M = pos.shape[0]
    names, ped, toks = [], [], []
    for f in range(n_fam):
        founders = rng.integers(1, 3, size=(4, 2, M)).astype(np.uint8)
        pars = []
        for p in range(2):
            a = _gametes(rng, founders[2 * p], pos, cs, 1)[0]
            b = _gametes(rng, founders[2 * p + 1], pos, cs, 1)[0]
            pars.append(np.stack([a, b]))
        o0 = _gametes(rng, pars[0], pos, cs, n_off)
        o1 = _gametes(rng, pars[1], pos, cs, n_off)
        for q in range(4):
            names.append(f"F{f}_{q}")
            ped.append((names[-1], "0", "0", 0))
            toks.append(_tokens((founders[q][0][None], founders[q][1][None]), rng, missing)[0])
        for p in range(2):
            names.append(f"P{f}_{p}")
            ped.append((names[-1], f"F{f}_{2 * p}", f"F{f}_{2 * p + 1}", 1))
            toks.append(_tokens((pars[p][0][None], pars[p][1][None]), rng, missing)[0])
        for k in range(n_off):
            names.append(f"O{f}_{k}")
            ped.append((names[-1], f"P{f}_0", f"P{f}_1", 2))
        toks.extend(_tokens((o0, o1), rng, missing))
    return dict(names=names, ped=ped, tokens=np.stack(toks), markerpos=pos, chromstarts=cs)


def write_alpha_files(prefix, sim):
    """Writes <prefix>.map/.ped/.gen as readalphamap/readalphaped/readalphadata expect them."""
    with open(prefix + ".map", "w") as f:
        f.write("\n".join(repr(float(x)) for x in sim["markerpos"]) + "\n")
    with open(prefix + ".ped", "w") as f:
        for me, fa, mo, gen in sim["ped"]:
            f.write(f"{me} {fa} {mo} {gen}\n")
    with open(prefix + ".gen", "w") as f:
        for name, row in zip(sim["names"], sim["tokens"]):
            f.write(name + " " + " ".join(str(int(t)) for t in row) + "\n")


def f2_flat(n_f2, n_chrom, markers_per_chrom, spacing_cM=0.1, missing=0.2, seed=3, pinned=None):
    """cfg3 as the flat arrays the engine's C ABI takes, in the state the reference holds them during its EM
    iterations (see tests/golden/f2small_iter2.npz): founders A (2,2) and B (1,1) with markersure 0.02; two
    implicit, `empty` F1 parents per F2 (readalphaped, cnF2freq.cpp:6515-6527) inferred as (2,1) with
    markersure 0.5 and their phase locked (haploweight 0) at the first marker of each chromosome
    (lockhaplos, :3045-3079); F2 genotypes with markersure 0.02, missing ones unknown.
    Individual ids: 0 none, 1 A, 2 B, then per F2: itself, F1 father, F1 mother.
    pinned: optional callable (shape, dtype) -> ndarray allocating page-locked host memory."""
    rng = np.random.default_rng(seed)
    alloc = pinned or (lambda shape, dtype: np.zeros(shape, dtype))
    pos, cs = make_map(n_chrom, markers_per_chrom, spacing_cM)
    M = pos.shape[0]
    K = 3 + 3 * n_f2
    f1 = np.stack([np.full(M, 2, np.uint8), np.full(M, 1, np.uint8)])
    g0 = _gametes(rng, f1, pos, cs, n_f2)
    g1 = _gametes(rng, f1, pos, cs, n_f2)
    miss = rng.random((n_f2, M)) < missing
    alleles = alloc((K, M, 2), np.uint8)
    sure = alloc((K, M, 2), np.float64)
    hw = alloc((K, M), np.float64)
    alleles[...] = 0
    sure[...] = 0
    hw[...] = 0.5
    alleles[1] = 2
    alleles[2] = 1
    sure[1:3] = 0.02
    f2 = 3 + 3 * np.arange(n_f2)
    a = np.stack([g0, g1], axis=2)
    a[miss] = 0
    alleles[f2] = a
    sure[f2] = np.where(miss[:, :, None], 0.0, 0.02)
    for off in (1, 2):
        alleles[f2 + off, :, 0] = 2
        alleles[f2 + off, :, 1] = 1
        sure[f2 + off] = 0.5
        hw[(f2 + off)[:, None], np.asarray(cs[:-1], np.int64)[None, :]] = 0.0
    par = np.zeros((K, 2), np.int32)
    par[f2, 0] = f2 + 1
    par[f2, 1] = f2 + 2
    par[f2 + 1] = (1, 2)
    par[f2 + 2] = (1, 2)
    empty = np.zeros(K, np.uint8)
    empty[0] = 1
    empty[f2 + 1] = 1
    empty[f2 + 2] = 1
    gen = np.zeros(K, np.int32)
    gen[f2] = 2
    gen[f2 + 1] = 1
    gen[f2 + 2] = 1
    desc = np.zeros(K, np.int32)
    desc[f2] = 1
    founder = np.zeros(K, np.uint8)
    founder[1:3] = 1
    return dict(par=par, empty=empty, founder=founder, gen=gen, descendants=desc, dous=f2.astype(np.int32), markerpos=pos,
                chromstarts=cs, genrec=np.full(3, -1.0 / 50), alleles=alleles, sure=sure, haploweight=hw,
                shared=np.array([1, 2], np.int32))


def fullsib_flat(n_fam, n_off, n_chrom, markers_per_chrom, spacing_cM=0.02, missing=0.05, seed=4, pinned=None):
    """cfg4 (outbred full-sib families, three generations, no inbreeding: 7 distinct tree members, all 128
    orientations and 8 shift modes in play) as the flat arrays of the C ABI.  Per family: 4 founders (gen 0,
    i.i.d. alleles at frequency 0.5), 2 parents (gen 1), n_off analysed offspring (gen 2); genotypes Mendelian,
    `missing` of them unknown in every generation, known ones with markersure 0.02 (readalphadata,
    cnF2freq.cpp:6584-6587); haploweight 0.5.  Ids: 0 none, then per family 4 founders, 2 parents, the offspring.
    Families are independent, so sharding whole families over ranks needs no shared rows."""
    rng = np.random.default_rng(seed)
    alloc = pinned or (lambda shape, dtype: np.zeros(shape, dtype))
    pos, cs = make_map(n_chrom, markers_per_chrom, spacing_cM)
    M = pos.shape[0]
    per = 6 + n_off
    K = 1 + n_fam * per
    alleles = alloc((K, M, 2), np.uint8)
    sure = alloc((K, M, 2), np.float64)
    hw = alloc((K, M), np.float64)
    alleles[...] = 0
    sure[...] = 0
    hw[...] = 0.5
    par = np.zeros((K, 2), np.int32)
    gen = np.zeros(K, np.int32)
    desc = np.zeros(K, np.int32)
    dous = []
    for f in range(n_fam):
        b = 1 + f * per
        founders = rng.integers(1, 3, size=(4, 2, M)).astype(np.uint8)
        pars = []
        for p in range(2):
            x = _gametes(rng, founders[2 * p], pos, cs, 1)[0]
            y = _gametes(rng, founders[2 * p + 1], pos, cs, 1)[0]
            pars.append(np.stack([x, y]))
        o0 = _gametes(rng, pars[0], pos, cs, n_off)
        o1 = _gametes(rng, pars[1], pos, cs, n_off)
        for q in range(4):
            alleles[b + q] = founders[q].T
        for p in range(2):
            alleles[b + 4 + p] = pars[p].T
            par[b + 4 + p] = (b + 2 * p, b + 2 * p + 1)
            gen[b + 4 + p] = 1
        ids = b + 6 + np.arange(n_off)
        alleles[ids] = np.stack([o0, o1], axis=2)
        par[ids] = (b + 4, b + 5)
        gen[ids] = 2
        desc[ids] = 1
        dous.extend(ids.tolist())
    miss = rng.random((K, M)) < missing
    alleles[miss] = 0
    known = alleles[:, :, 0] != 0
    sure[known] = 0.02
    empty = np.zeros(K, np.uint8)
    empty[0] = 1
    founder = np.zeros(K, np.uint8)
    founder[(gen == 0) & (np.arange(K) > 0)] = 1
    return dict(par=par, empty=empty, founder=founder, gen=gen, descendants=desc, dous=np.asarray(dous, np.int32),
                markerpos=pos, chromstarts=cs, genrec=np.full(3, -1.0 / 50), alleles=alleles, sure=sure, haploweight=hw,
                shared=np.zeros(0, np.int32))
