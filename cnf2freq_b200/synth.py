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


def simulate_selfed_lines(n, n_chrom, markers_per_chrom, spacing_cM, seed, gens, missing=0.2):
    """PlantImpute lines for the SELFING = true instantiation (demo.sh:18-20): an F2 cross whose line k is listed with pedigree
    generation gens[k] >= 2, i.e. selfed gens[k] - 2 times behind the listed pedigree; a heterozygous call survives each selfing
    with probability 1/2."""
    sim = simulate_f2(n_f2=n, n_chrom=n_chrom, markers_per_chrom=markers_per_chrom, spacing_cM=spacing_cM, missing=missing, seed=seed)
    rng = np.random.default_rng(seed)
    sim["ped"] = sim["ped"][:2] + [(me, fa, mo, g) for (me, fa, mo, _), g in zip(sim["ped"][2:], gens)]
    for k, g in enumerate(gens):
        row = sim["tokens"][2 + k]
        fix = (row == 1) & (rng.random(row.shape) > 0.5 ** (g - 2))
        row[fix] = rng.choice([0, 2], size=int(fix.sum()))
    return sim


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


class F2Shard:
    """One rank's part of an F2 design of `n_total` analysed individuals (cfg3 / cfg5 shapes): `lines` inbred founder lines
    (2: line A all `2`, line B all `1` -- the F2 intercross of cfg3; more: homozygous with allele frequency 0.5, biparental
    families of `fam_size` F2s each -- cfg5).  The pedigree arrays cover the whole design (ids: 0 none, 1..lines the
    founders, then per F2 itself, its F1 father, its F1 mother: the implicit `empty` F1s of readalphaped,
    cnF2freq.cpp:6515-6527); genotypes exist only for the `count` F2s from `first` on, and parameters are produced per
    chromosome in a caller-given row order (the engine's cnf2_row_ids) -- the packed form of the C ABI, so no
    [n_ind][n_markers] array is ever built.  State as the reference holds it during its EM iterations (see f2_flat)."""

    def __init__(self, n_total, first, count, n_chrom, markers_per_chrom, spacing_cM=0.1, missing=0.2, seed=3, lines=2,
                 fam_size=None):
        self.pos, self.cs = make_map(n_chrom, markers_per_chrom, spacing_cM)
        M = self.pos.shape[0]
        L = self.L = lines
        self.first, self.count, self.M = first, count, M
        fam_size = fam_size or n_total
        if L == 2:
            self.line = np.stack([np.full(M, 2, np.uint8), np.full(M, 1, np.uint8)])
        else:
            self.line = np.random.default_rng([seed, 1 << 20]).integers(1, 3, size=(L, M)).astype(np.uint8)
        fam = np.arange(n_total) // fam_size
        self.pa = (fam % L).astype(np.int64)
        self.pb = ((fam + 1 + fam // L) % L).astype(np.int64)
        self.pb = np.where(self.pb == self.pa, (self.pb + 1) % L, self.pb)
        rng = np.random.default_rng([seed, first])
        self.g = np.zeros((count, M, 2), np.uint8)
        loc = np.arange(first, first + count)
        for f in np.unique(fam[loc]):
            sel = np.nonzero(fam[loc] == f)[0]
            f1 = np.stack([self.line[self.pa[loc[sel[0]]]], self.line[self.pb[loc[sel[0]]]]])
            self.g[sel, :, 0] = _gametes(rng, f1, self.pos, self.cs, len(sel))
            self.g[sel, :, 1] = _gametes(rng, f1, self.pos, self.cs, len(sel))
        self.g[rng.random((count, M)) < missing] = 0
        K = 1 + L + 3 * n_total
        self.base = 1 + L
        f2 = self.base + 3 * np.arange(n_total)
        self.par = np.zeros((K, 2), np.int32)
        self.par[f2, 0], self.par[f2, 1] = f2 + 1, f2 + 2
        for off in (1, 2):
            self.par[f2 + off, 0], self.par[f2 + off, 1] = 1 + self.pa, 1 + self.pb
        self.empty = np.zeros(K, np.uint8)
        self.empty[0] = self.empty[f2 + 1] = self.empty[f2 + 2] = 1
        self.gen = np.zeros(K, np.int32)
        self.gen[f2], self.gen[f2 + 1], self.gen[f2 + 2] = 2, 1, 1
        self.descendants = np.zeros(K, np.int32)
        self.descendants[f2] = 1
        self.founder = np.zeros(K, np.uint8)
        self.founder[1:1 + L] = 1
        self.all_dous = f2.astype(np.int32)
        self.dous = self.all_dous[first:first + count].copy()
        self.shared = np.arange(1, 1 + L, dtype=np.int32)
        self.genrec = np.full(3, -1.0 / 50)
        self.hasprior = (1 - self.empty).astype(np.uint8)   # readalphadata copies the genotypes it reads (:6664-6665)
        # individ::children as every doit recounts it: analysed children only (:5226, :5246-5259)
        self.children = np.bincount(self.par[self.all_dous].ravel(), minlength=K).astype(np.int32)
        self.children[0] = 0

    def params(self, ids, c, out=None):
        """(alleles u8 [R][Mc][2], sure f64 [R][Mc][2], haploweight f64 [R][Mc]) of chromosome c for the rows `ids`."""
        cols = slice(int(self.cs[c]), int(self.cs[c + 1]))
        Mc, R = cols.stop - cols.start, len(ids)
        al, su, hw = out if out is not None else (np.zeros((R, Mc, 2), np.uint8), np.zeros((R, Mc, 2)), np.zeros((R, Mc)))
        ids = np.asarray(ids, np.int64)
        al[...], su[...], hw[...] = 0, 0, 0.5
        rf = np.nonzero((ids >= 1) & (ids < self.base))[0]
        al[rf] = self.line[ids[rf] - 1][:, cols, None]
        su[rf] = 0.02
        gk, off = (ids - self.base) // 3, (ids - self.base) % 3
        k = gk - self.first
        own = (ids >= self.base) & (k >= 0) & (k < self.count)
        r2 = np.nonzero(own & (off == 0))[0]
        a = self.g[k[r2], cols]
        al[r2] = a
        su[r2] = np.where((a == 0), 0.0, 0.02)
        r1 = np.nonzero(own & (off != 0))[0]
        al[r1, :, 0] = self.line[self.pa[gk[r1]]][:, cols]   # fixkid (:1469-1487): a homozygous parent's allele, sureness 0.5
        al[r1, :, 1] = self.line[self.pb[gk[r1]]][:, cols]
        su[r1] = 0.5
        hw[r1, 0] = 0.0   # phase locked at the first marker of the chromosome (lockhaplos, :3045-3079)
        return al, su, hw

    def priors(self, ids, c, out=None):
        """(relhaplo, prior alleles, prior markersure) of chromosome c: relhaplo 0.5 (:2496), priors = the genotypes as read."""
        cols = int(self.cs[c + 1]) - int(self.cs[c])
        rel, al, su = out if out is not None else (np.zeros((len(ids), cols)), np.zeros((len(ids), cols, 2), np.uint8), np.zeros((len(ids), cols, 2)))
        self.params(ids, c, out=(al, su, rel))   # rel doubles as the haploweight scratch
        noprior = self.hasprior[np.asarray(ids, np.int64)] == 0
        al[noprior], su[noprior] = 0, 0
        rel[...] = 0.5
        return rel, al, su
