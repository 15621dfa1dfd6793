"""The CPU oracle (oracle/cnf2_oracle.cpp) against what the reference's own code computed (tests/golden/*.npz,
written by tests/golden/make_golden.py from oracle/_ref/cnf2ref).  This is what pins the oracle."""
import numpy as np
import pytest

import oracle_lib as ol
from cases import GOLDEN_CASES, SELFING_CASES, STATS, assert_close, load_golden


def test_fixtures_present():
    assert {"demo_iter1", "demo_iter3", "halfsib_iter2", "f2small_iter2", "fullsib_iter2"} <= set(GOLDEN_CASES)


def test_selfing_fixtures_present():
    assert {"selfing_demo_iter2", "selfing_lines_iter2"} <= set(SELFING_CASES)


@pytest.mark.parametrize("name", GOLDEN_CASES + SELFING_CASES)
def test_oracle_statistics_match_reference(name):
    g, pb = load_golden(name)
    out = ol.oracle_estep(pb, selfing=name in SELFING_CASES)
    for k in STATS:
        # same algorithm, same operation order: expected to be bit-identical, asserted at 1e-12
        assert_close(out[k], g[k], f"{name}:{k}", rtol=1e-12)
    # the reference prints the log-likelihood pair with 6 decimals (cnF2freq.cpp:5399-5401)
    assert_close(out["loglik"], g["loglik_printed"], f"{name}:loglik", rtol=0, atol=6e-7)
    # turn scoring + updatenegshifts (row f2; :5668-5757, :3640-3715).  Scores are sums of log(val) terms that cancel to
    # rounding where a turn changes nothing, hence the absolute part
    assert_close(out["negshift"], g["negshift"], f"{name}:negshift", rtol=1e-12, atol=1e-13)


@pytest.mark.parametrize("name", GOLDEN_CASES + SELFING_CASES)
def test_oracle_sweeps_match_reference(name):
    g, pb = load_golden(name)
    keys = [k for k in g if k.startswith("fwbw_c")]
    assert keys
    for k in keys:
        c, n = (int(x[1:]) for x in k.split("_")[1:])
        j = int(np.nonzero(pb.dous == n)[0][0])
        out = ol.oracle_estep(pb, fwbw_for=(j, c), selfing=name in SELFING_CASES)
        mask = g[f"fwbwmask_c{c}_n{n}"]
        for s in range(8):
            if mask[s] != 3:
                continue
            assert out["fwbwmask"][s] == 3
            # pad 0 = alpha before emission, 1 = beta, 2 = alpha after emission (cnF2freq.cpp:2172-2178)
            assert_close(out["fwbw"][s], g[k][s], f"{name}:{k}[shift {s}]", rtol=1e-12, atol=1e-300)
            assert_close(out["fwbwfactors"][s], g[f"fwbwfactors_c{c}_n{n}"][s], f"{name}:factors[shift {s}]", rtol=1e-12)


def test_posterior_mass_is_not_normalised():
    """the surviving (g, shift, flag2) terms do not sum to one; the statistics do after the
    transfer's own normalisation.  infprobs[root][side 0] sums to descendants over the two alleles."""
    g, pb = load_golden("f2small_iter2")
    out = ol.oracle_estep(pb, counters=True)
    tried, kept = out["counters"]
    assert 0 < kept <= tried
    for n in pb.dous:
        tot = out["infprobs"][n, :, 0, :].sum(axis=1)
        np.testing.assert_allclose(tot, pb.descendants[n], rtol=1e-9)
