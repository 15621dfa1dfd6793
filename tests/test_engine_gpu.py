"""Parity of the CUDA engine (through the C ABI) with the reference: against the golden fixtures the
reference binary wrote, and against the CPU oracle on seeded inputs.  Tolerance: 1e-9 relative (fp64),
as BASELINE.json's north_star states."""
import os

import numpy as np
import pytest

import oracle_lib as ol
import randomped
from cases import GOLDEN_CASES, NS_ATOL, RTOL, STATS, assert_close, engine_problem, load_golden

pytestmark = pytest.mark.gpu
ATOL = 1e-13  # statistics are sums of O(1) terms; entries that are exact zeros in one order of summation


@pytest.fixture(scope="module")
def eng():
    from cnf2freq_b200.engine import Engine
    e = Engine(device=0)
    yield e
    e.close()


def run_engine(eng, pb, **kw):
    engine_problem(eng, pb, **kw)
    return eng.estep_all(pb.alleles, pb.sure, pb.haploweight)


def compare(out, ref, what):
    for k in STATS:
        assert_close(out[k], ref[k], f"{what}:{k}", rtol=RTOL, atol=ATOL)
    assert_close(out["negshift"], ref["negshift"], f"{what}:negshift", rtol=RTOL, atol=NS_ATOL)
    assert_close(out["loglik"], ref["loglik"], f"{what}:loglik", rtol=RTOL, atol=1e-9)


@pytest.mark.parametrize("name", GOLDEN_CASES)
def test_engine_matches_reference_golden(eng, name):
    """The reference's sticky founder flags (fixtrees) are steady from its pre-processing on (postmarkerdata calls fixtrees
    for every individual, :3377), which is the state the engine works in (include/cnf2_engine.h, cnf2_set_pedigree): the
    first EM iteration is reproduced like the later ones."""
    g, pb = load_golden(name)
    out = run_engine(eng, pb)
    for k in STATS:
        assert_close(out[k], g[k], f"{name}:{k}", rtol=RTOL, atol=ATOL)
    assert_close(out["negshift"], g["negshift"], f"{name}:negshift", rtol=RTOL, atol=NS_ATOL)
    assert_close(out["loglik"], g["loglik_printed"], f"{name}:loglik", rtol=0, atol=6e-7)


@pytest.mark.parametrize("name", GOLDEN_CASES)
def test_engine_matches_oracle_on_golden_inputs(eng, name):
    g, pb = load_golden(name)
    pb.founder = ol.steady_founder(pb)
    compare(run_engine(eng, pb), ol.oracle_estep(pb), name)


@pytest.mark.parametrize("maker,kw", [
    (randomped.f2_cross, dict(n_f2=12, n_chrom=2, mpc=37, seed=11)),
    (randomped.f2_cross, dict(n_f2=3, n_chrom=1, mpc=1, seed=12)),       # single-marker chromosome
    (randomped.f2_cross, dict(n_f2=5, n_chrom=3, mpc=33, seed=13, missing=0.6)),
    (randomped.fullsib, dict(n_fam=1, n_off=2, n_chrom=1, mpc=9, seed=14)),
    (randomped.tangled, dict(n_chrom=2, mpc=21, seed=15)),
    (randomped.tangled, dict(n_chrom=1, mpc=70, seed=16, missing=0.5)),
    # two of the six units are impossible under every shift mode (zero likelihood): the reference carries on with
    # factor = MINFACTOR + log 8 and scatters NaNs into the relatives' statistics; so do we
    (randomped.fullsib, dict(n_fam=2, n_off=3, n_chrom=1, mpc=17, seed=42)),
])
def test_engine_matches_oracle_on_random_pedigrees(eng, maker, kw):
    pb = maker(**kw)
    compare(run_engine(eng, pb), ol.oracle_estep(pb), f"{maker.__name__}{kw}")


def test_sharded_engines_sum_to_the_whole(eng):
    """Two shards of the analysed individuals (what two ranks hold) add up to the unsharded statistics."""
    pb = randomped.f2_cross(n_f2=10, n_chrom=2, mpc=19, seed=21)
    whole = run_engine(eng, pb)
    acc = {k: np.zeros_like(whole[k]) for k in STATS + ("negshift",)}
    for r in range(2):
        part = run_engine(eng, pb, dous=pb.dous[r::2], all_dous=pb.dous, shared=[1, 2])
        for k in acc:
            acc[k] += part[k]
        assert_close(part["loglik"], whole["loglik"][r::2], "shard loglik", rtol=1e-12)
    for k in acc:
        assert_close(acc[k], whole[k], f"sharded:{k}", rtol=1e-12, atol=ATOL)


def test_call_order_is_enforced(eng):
    from cnf2freq_b200.engine import EngineError
    pb = randomped.f2_cross(n_f2=2, n_chrom=2, mpc=5, seed=1)
    engine_problem(eng, pb)
    with pytest.raises(EngineError, match="cnf2_set_params"):
        eng.estep(0)
    eng.set_params(0, pb.alleles, pb.sure, pb.haploweight)
    with pytest.raises(EngineError, match="cnf2_set_params"):
        eng.estep(1)
    eng.estep(0)
    with pytest.raises(EngineError, match="cnf2_estep"):
        eng.get_stats(1)
    eng.sync()


def test_one_kernel_launch_per_estep(eng):
    pb = randomped.f2_cross(n_f2=4, n_chrom=1, mpc=8, seed=2)
    engine_problem(eng, pb)
    eng.set_params(0, pb.alleles, pb.sure, pb.haploweight)
    eng.estep(0)
    eng.sync()
    assert eng.last_estep_launches() >= 1
    assert eng.last_estep_ms() > 0


@pytest.mark.parametrize("maker,kw", [
    ("f2_flat", dict(n_f2=6, n_chrom=2, markers_per_chrom=30, seed=3)),            # bench workload (configs[2]) in small
    ("fullsib_flat", dict(n_fam=2, n_off=3, n_chrom=1, markers_per_chrom=25, seed=4)),  # configs[3]'s pedigree shape
])
def test_engine_matches_oracle_on_bench_workloads(eng, maker, kw):
    from cnf2freq_b200 import synth
    d = getattr(synth, maker)(**kw)
    pb = ol.Problem(d["par"], d["empty"], d["founder"], d["gen"], d["descendants"], d["dous"], d["markerpos"],
                    d["chromstarts"], d["genrec"], d["alleles"], d["sure"], d["haploweight"])
    compare(run_engine(eng, pb), ol.oracle_estep(pb), maker)


def test_full_size_chromosome_properties(eng):
    """BASELINE.json configs[2] at full size for one chromosome (2,000 F2 x 1,000 markers; the oracle would need
    hours): size-independent properties of the statistics.
      * moveinfprobs normalises by the root's side-0 sum (cnF2freq.cpp:5879-5885, :3577-3597): an analysed individual
        that occurs once in its own tree gets infprobs[m][0][1] + infprobs[m][0][2] == descendants at every marker;
      * movehaplos (:3599-3616) adds `descendants` to haplocount exactly where the individual's phase matters
        (not homozygous-with-equal-sureness, :1235-1239, which includes unknown genotypes) and is not locked;
      * linearity: the statistics of two disjoint shards of the analysed individuals add up to the whole
        (what the all-reduce relies on);
      * the two likelihood columns differ by at most log 8 (8 shift modes)."""
    from cnf2freq_b200 import synth
    d = synth.f2_flat(2000, 1, 1000, spacing_cM=0.1, missing=0.2, seed=3)
    pb = ol.Problem(d["par"], d["empty"], d["founder"], d["gen"], d["descendants"], d["dous"], d["markerpos"],
                    d["chromstarts"], d["genrec"], d["alleles"], d["sure"], d["haploweight"])
    whole = run_engine(eng, pb, shared=[1, 2])
    f2 = d["dous"]
    side0 = whole["infprobs"][f2][:, :, 0, :].sum(axis=2)
    assert_close(side0, np.ones_like(side0), "root side-0 infprobs sum", rtol=1e-9)
    side1 = whole["infprobs"][f2][:, :, 1, :].sum(axis=2)
    assert_close(side1, np.ones_like(side1), "root side-1 infprobs sum", rtol=1e-9)
    al, su = d["alleles"][f2], d["sure"][f2]
    phased = (al[:, :, 0] != al[:, :, 1]) | (su[:, :, 0] != su[:, :, 1])
    assert_close(whole["haplocount"][f2], phased.astype(np.float64), "root haplocount", rtol=0, atol=0)
    assert np.all(whole["haplobase"][f2] >= 0) and np.all(whole["haplobase"][f2] <= 1)
    # the founders are homozygous everywhere: no phase statistics; every F2 adds to their genotype statistics
    assert_close(whole["haplocount"][1:3], np.zeros((2, 1000)), "founder haplocount", rtol=0, atol=0)
    assert np.all(whole["infprobs"][1:3].sum(axis=(2, 3)) > 0)
    ll = whole["loglik"][:, 0, :]
    assert np.all(np.isfinite(ll)) and np.all(ll[:, 1] >= ll[:, 0]) and np.all(ll[:, 1] - ll[:, 0] <= np.log(8) + 1e-12)
    acc = {k: np.zeros_like(whole[k]) for k in STATS}
    for r in range(2):
        part = run_engine(eng, pb, dous=f2[r::2], all_dous=f2, shared=[1, 2])
        for k in STATS:
            acc[k] += part[k]
    for k in STATS:
        assert_close(acc[k], whole[k], f"linearity:{k}", rtol=1e-9, atol=1e-12)


@pytest.mark.parametrize("maker,kw,n_sample,own", [
    # BASELINE.json configs[2] at full size for one chromosome; an F2's own row and its two implicit F1 parents' rows
    # receive contributions from its own unit only (cnF2freq.cpp:5283-5295, :5893-5902)
    ("f2_flat", dict(n_f2=2000, n_chrom=1, markers_per_chrom=1000, spacing_cM=0.1, missing=0.2, seed=3), 16, 3),
    # configs[3]'s pedigree shape at the size bench.py --workload fullsib runs per chromosome
    ("fullsib_flat", dict(n_fam=66, n_off=20, n_chrom=1, markers_per_chrom=400, spacing_cM=0.02, missing=0.05, seed=4), 16, 1),
])
def test_full_size_sampled_parity(eng, monkeypatch, maker, kw, n_sample, own):
    """Full-size E-step on the GPU; the oracle on the pedigree reduced to a seeded sample of the analysed individuals.
    The log-likelihood pair and the statistics rows only the sampled unit adds to must agree to 1e-9."""
    from cnf2freq_b200 import synth
    d = getattr(synth, maker)(**kw)
    pb = ol.Problem(d["par"], d["empty"], d["founder"], d["gen"], d["descendants"], d["dous"], d["markerpos"],
                    d["chromstarts"], d["genrec"], d["alleles"], d["sure"], d["haploweight"])
    whole = run_engine(eng, pb, shared=d["shared"])
    rng = np.random.default_rng(1234)
    pick = np.sort(rng.choice(len(pb.dous), n_sample, replace=False))
    sub = ol.Problem(d["par"], d["empty"], d["founder"], d["gen"], d["descendants"], pb.dous[pick], d["markerpos"],
                     d["chromstarts"], d["genrec"], d["alleles"], d["sure"], d["haploweight"])
    monkeypatch.setenv("CNF2_ORACLE_THREADS", str(os.cpu_count() or 1))  # units in parallel; own rows are unaffected by the order
    ref = ol.oracle_estep(sub, all_dous=pb.dous)
    assert_close(whole["loglik"][pick], ref["loglik"], f"{maker}: sampled loglik", rtol=RTOL, atol=1e-9)
    rows = np.concatenate([pb.dous[pick] + k for k in range(own)])  # f2_flat ids: the F2, its F1 father, its F1 mother
    for k in STATS:
        assert_close(whole[k][rows], ref[k][rows], f"{maker}: sampled own rows {k}", rtol=RTOL, atol=ATOL)
    # the reference forms a score from log-factors of the size of the log-likelihood (F+ + B + log(sum) - factor, :1994-2016),
    # so its own rounding noise grows with the chromosome: ~ 1e-13 |loglik| where a turn changes nothing
    ns_atol = max(NS_ATOL, 1e-13 * float(np.abs(ref["loglik"]).max()))
    assert_close(whole["negshift"][rows], ref["negshift"][rows], f"{maker}: sampled own rows negshift", rtol=RTOL, atol=ns_atol)


def test_no_analysed_individuals(eng):
    """An empty shard (a rank with nothing to do) is legal: the E-step launches nothing and touches nothing."""
    pb = randomped.f2_cross(n_f2=2, n_chrom=1, mpc=5, seed=3)
    engine_problem(eng, pb, dous=np.zeros(0, np.int32), all_dous=pb.dous, shared=np.zeros(0, np.int32))
    eng.set_params(0, pb.alleles, pb.sure, pb.haploweight)
    eng.estep(0)
    out = dict(haplobase=np.full((len(pb.par), 5), 7.0), haplocount=np.full((len(pb.par), 5), 7.0))
    eng.get_stats(0, out["haplobase"], out["haplocount"])
    eng.sync()
    assert eng.last_estep_launches() == 0
    assert np.all(out["haplobase"] == 7.0) and np.all(out["haplocount"] == 7.0)


def test_packed_statistics_equal_the_scattered_ones(eng):
    pb = randomped.tangled(n_chrom=2, mpc=13, seed=51)
    engine_problem(eng, pb, shared=np.array([1, 4], np.int32))
    ids = eng.row_ids()
    assert len(set(ids.tolist())) == len(ids) and ids[0] == 1 and ids[1] == 4
    K, M = len(pb.par), len(pb.markerpos)
    for c in range(2):
        eng.set_params(c, pb.alleles, pb.sure, pb.haploweight)
        eng.estep(c)
        full = dict(haplobase=np.zeros((K, M)), haplocount=np.zeros((K, M)), infprobs=np.zeros((K, M, 2, 2)),
                    homozyg=np.zeros((K, M, 2)), loglik=np.zeros((len(pb.dous), 2, 2)))
        eng.get_stats(c, **full)
        Mc = 13
        pk = dict(haplobase=np.zeros((len(ids), Mc)), haplocount=np.zeros((len(ids), Mc)), infprobs=np.zeros((len(ids), Mc, 2, 2)),
                  homozyg=np.zeros((len(ids), Mc, 2)), loglik=np.zeros((len(pb.dous), 2)))
        eng.get_stats_packed(c, **pk)
        for k in STATS:
            assert np.array_equal(pk[k], full[k][ids, c * Mc:(c + 1) * Mc], equal_nan=True), k
        assert np.array_equal(pk["loglik"], full["loglik"][:, c, :])
        # the same E-step from row-ordered parameter blocks of this chromosome
        cols = slice(c * Mc, (c + 1) * Mc)
        eng.set_params_packed(c, np.ascontiguousarray(pb.alleles[ids][:, cols]), np.ascontiguousarray(pb.sure[ids][:, cols]),
                              np.ascontiguousarray(pb.haploweight[ids][:, cols]))
        eng.estep(c)
        pk2 = {k: np.zeros_like(v) for k, v in pk.items()}
        eng.get_stats_packed(c, **pk2)
        for k in pk:
            assert_close(pk2[k], pk[k], "packed parameters:" + k, rtol=1e-12, atol=1e-15)  # atomics: order of summation


@pytest.mark.parametrize("segment", [4, 8, 20])
@pytest.mark.parametrize("maker,kw", [
    (randomped.f2_cross, dict(n_f2=12, n_chrom=2, mpc=37, seed=11)),
    (randomped.tangled, dict(n_chrom=1, mpc=70, seed=16, missing=0.5)),
    (randomped.fullsib, dict(n_fam=1, n_off=3, n_chrom=2, mpc=26, seed=17)),
])
def test_segmented_sweeps_equal_the_oracle(maker, kw, segment):
    """Row a5: alpha checkpoints + recompute.  With a segment shorter than the chromosome the forward pass keeps alpha at
    segment starts only and the backward pass re-runs the forward sweep over each segment before it walks it; the
    statistics, scores and log-likelihoods must not depend on the segment length (ragged last segments included)."""
    from cnf2freq_b200.engine import Engine
    pb = maker(**kw)
    e = Engine(device=0, segment=segment)
    try:
        compare(run_engine(e, pb), ol.oracle_estep(pb), f"{maker.__name__}{kw} segment {segment}")
    finally:
        e.close()


def test_multi_line_f2_design_matches_oracle(eng):
    """The cfg5-shaped bench workload in small (synth.F2Shard: biparental F2 families from several inbred lines, founders
    shared between families, F1s homozygous where their lines agree) against the oracle, through the packed-parameter
    calls the bench uses."""
    from cnf2freq_b200 import synth
    s = synth.F2Shard(12, 0, 12, 2, 21, spacing_cM=0.7, missing=0.2, seed=5, lines=5, fam_size=3)
    K, M, C = s.par.shape[0], s.M, 2
    ids = np.arange(K)
    al, su, hw = np.zeros((K, M, 2), np.uint8), np.zeros((K, M, 2)), np.zeros((K, M))
    for c in range(C):
        cols = slice(int(s.cs[c]), int(s.cs[c + 1]))
        al[:, cols], su[:, cols], hw[:, cols] = s.params(ids, c)
    al[0], su[0] = 0, 0
    pb = ol.Problem(s.par, s.empty, s.founder, s.gen, s.descendants, s.dous, s.pos, s.cs, s.genrec, al, su, hw)
    ref = ol.oracle_estep(pb)
    engine_problem(eng, pb, shared=s.shared)
    rows = eng.row_ids()
    out = dict(haplobase=np.zeros((K, M)), haplocount=np.zeros((K, M)), infprobs=np.zeros((K, M, 2, 2)), homozyg=np.zeros((K, M, 2)),
               loglik=np.zeros((len(s.dous), C, 2)), negshift=np.zeros((K, M)))
    for c in range(C):
        eng.set_params_packed(c, *[np.ascontiguousarray(x) for x in s.params(rows, c)])
        eng.estep(c)
        eng.get_stats(c, out["haplobase"], out["haplocount"], out["infprobs"], out["homozyg"], out["loglik"])
        eng.get_negshift(c, out["negshift"])
    compare(out, ref, "F2Shard lines=5")
    eng.set_shared_rows(np.zeros(0, np.int32))
