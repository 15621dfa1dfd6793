"""Parity of the CUDA engine (through the C ABI) with the reference: against the golden fixtures the
reference binary wrote, and against the CPU oracle on seeded inputs.  Tolerance: 1e-9 relative (fp64),
as BASELINE.json's north_star states."""
import numpy as np
import pytest

import oracle_lib as ol
import randomped
from cases import GOLDEN_CASES, RTOL, STATS, assert_close, engine_problem, load_golden

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
    assert_close(out["loglik"], ref["loglik"], f"{what}:loglik", rtol=RTOL, atol=1e-9)


@pytest.mark.parametrize("name", [n for n in GOLDEN_CASES if not n.endswith("iter1")])
def test_engine_matches_reference_golden(eng, name):
    """From the second EM iteration on the reference's sticky founder flags are steady, which is the state
    the engine works in (include/cnf2_engine.h, cnf2_set_pedigree)."""
    g, pb = load_golden(name)
    out = run_engine(eng, pb)
    for k in STATS:
        assert_close(out[k], g[k], f"{name}:{k}", rtol=RTOL, atol=ATOL)
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
    acc = {k: np.zeros_like(whole[k]) for k in STATS}
    for r in range(2):
        part = run_engine(eng, pb, dous=pb.dous[r::2], all_dous=pb.dous, shared=[1, 2])
        for k in STATS:
            acc[k] += part[k]
        assert_close(part["loglik"], whole["loglik"][r::2], "shard loglik", rtol=1e-12)
    for k in STATS:
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
