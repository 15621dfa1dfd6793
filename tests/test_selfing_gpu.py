"""Row g1: the SELFING = true instantiation (settings.h:14; 256 states) of the CUDA engine through the C ABI
(cnf2_settings.selfing = 1) against the reference compiled with that knob (oracle/_ref/cnf2ref_selfing -> tests/golden/selfing_*.npz,
mstep_run_selfing.npz) and against the CPU restatement compiled the same way, on seeded pedigrees whose analysed individuals
carry generations 2..7.  Tolerance: 1e-9 relative (fp64), as everywhere."""
import numpy as np
import pytest

import oracle_lib as ol
import randomped
import test_em_loop_gpu as eml
from cases import NS_ATOL, RTOL, SELFING_CASES, STATS, assert_close, engine_problem, load_golden

pytestmark = pytest.mark.gpu
ATOL = 1e-13


@pytest.fixture(scope="module")
def eng():
    from cnf2freq_b200.engine import Engine
    e = Engine(device=0, selfing=True)
    yield e
    e.close()


def run_engine(eng, pb):
    engine_problem(eng, pb)
    return eng.estep_all(pb.alleles, pb.sure, pb.haploweight)


@pytest.mark.parametrize("name", SELFING_CASES)
def test_selfing_engine_matches_reference_golden(eng, name):
    g, pb = load_golden(name)
    out = run_engine(eng, pb)
    for k in STATS:
        assert_close(out[k], g[k], f"{name}:{k}", rtol=RTOL, atol=ATOL)
    assert_close(out["negshift"], g["negshift"], f"{name}:negshift", rtol=RTOL, atol=NS_ATOL)
    assert_close(out["loglik"], g["loglik_printed"], f"{name}:loglik", rtol=0, atol=6e-7)


def selfed(pb, seed):
    """The analysed individuals of a seeded pedigree put at generations 2..7 (gen - 2 selfings behind the listed pedigree)."""
    rng = np.random.default_rng(1000 + seed)
    keep = np.array([d for d in pb.dous if pb.gen[d] >= 2], np.int32)
    pb.dous = keep
    pb.gen = pb.gen.copy()
    pb.gen[keep] = rng.integers(2, 8, len(keep)).astype(np.int32)
    return pb


@pytest.mark.parametrize("maker,kw", [
    (randomped.f2_cross, dict(n_f2=8, n_chrom=2, mpc=19, seed=31)),
    (randomped.f2_cross, dict(n_f2=3, n_chrom=1, mpc=1, seed=32)),  # single-marker chromosome
    (randomped.fullsib, dict(n_fam=1, n_off=3, n_chrom=1, mpc=9, seed=34)),
    (randomped.tangled, dict(n_chrom=1, mpc=15, seed=35)),
])
def test_selfing_engine_matches_oracle_on_random_pedigrees(eng, maker, kw):
    pb = selfed(maker(**kw), kw["seed"])
    ref = ol.oracle_estep(pb, selfing=True)
    out = run_engine(eng, pb)
    what = f"selfing {maker.__name__}{kw}"
    for k in STATS:
        assert_close(out[k], ref[k], f"{what}:{k}", rtol=RTOL, atol=ATOL)
    assert_close(out["negshift"], ref["negshift"], f"{what}:negshift", rtol=RTOL, atol=NS_ATOL)
    assert_close(out["loglik"], ref["loglik"], f"{what}:loglik", rtol=RTOL, atol=1e-9)


def test_selfing_generation_two_shifts_the_likelihood_by_the_prior(eng):
    """With every analysed individual at generation 2 the selfing states carry no mass (:2095-2099, selfprec[0][0] = 1): the
    sweeps are those of the SELFING = false engine and the log-likelihood is lower by log 4 (EVENGEN = 1 / 256 instead of
    1 / 64).  The statistics differ: rule 3 of ignoreflag2 is off under SELFING (:3489)."""
    from cnf2freq_b200.engine import Engine
    pb = randomped.f2_cross(n_f2=5, n_chrom=1, mpc=21, seed=36)
    out = run_engine(eng, pb)
    plain = Engine(device=0)
    try:
        ref = run_engine(plain, pb)
    finally:
        plain.close()
    assert_close(out["loglik"], ref["loglik"] - np.log(4.0), "gen 2:loglik", rtol=RTOL, atol=1e-9)


def test_selfing_rejects_generations_it_cannot_hold(eng):
    from cnf2freq_b200.engine import EngineError
    pb = randomped.f2_cross(n_f2=2, n_chrom=1, mpc=4, seed=37)
    pb.gen = pb.gen.copy()
    pb.gen[pb.dous[0]] = 40
    with pytest.raises(EngineError, match="SELFING"):
        engine_problem(eng, pb)


def test_selfing_em_iterations_teacher_forced(eng):
    """Whole EM iterations of the SELFING build (E-steps, phase inversions, M-step passes, step-size control) per iteration."""
    eml.test_reference_inversions_teacher_forced(eng, "mstep_run_selfing")


def test_selfing_em_pass_makes_the_reference_decisions(eng):
    eml.test_em_pass_makes_the_reference_decisions(eng, "mstep_run_selfing")


def test_selfing_shards_sum_to_the_whole(eng):
    """Two shards of the analysed individuals (what two ranks hold) add up to the unsharded statistics under SELFING too."""
    pb = selfed(randomped.f2_cross(n_f2=6, n_chrom=1, mpc=17, seed=38), 38)
    whole = run_engine(eng, pb)
    keys = STATS + ("negshift",)
    acc = {k: np.zeros_like(whole[k]) for k in keys}
    for r in range(2):
        engine_problem(eng, pb, dous=pb.dous[r::2], all_dous=pb.dous, shared=[1, 2])
        part = eng.estep_all(pb.alleles, pb.sure, pb.haploweight)
        for k in keys:
            acc[k] += part[k]
        assert_close(part["loglik"], whole["loglik"][r::2], "selfing shard loglik", rtol=1e-12)
    for k in keys:
        assert_close(acc[k], whole[k], f"selfing sharded:{k}", rtol=1e-12, atol=ATOL)
