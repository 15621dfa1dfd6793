"""The EM loop on the device (rows a6 / f1 / f2) through the C ABI against whole iterations of the reference binary in
which it inverts phases (tests/golden/mstep_inv_*.npz, made by tests/golden/make_mstep_golden.py): the phase-inversion
scores and their minima, negshifter's inversions, the pass that follows them with scalefactor 0, the step-size control
on the device, and the candidate selection with libc rand()."""
import ctypes
import os

import numpy as np
import pytest

from cases import GOLDEN, NS_ATOL, assert_close

pytestmark = pytest.mark.gpu
RTOL, ATOL = 1e-9, 1e-12
# markersure and haploweight are probabilities p (or 1 - p): the bar is 1e-9 relative on the larger of the two, i.e. the
# bisection's end point itself; for a markersure of 1e-3 that is 1e-6 relative
PTOL = 1e-9


def assert_prob_close(a, b, what):
    err = np.abs(np.asarray(a) - np.asarray(b))
    tol = PTOL * np.maximum(np.abs(b), np.abs(1 - np.asarray(b)))
    if (err > tol).any():
        i = np.unravel_index(np.argmax(err - tol), err.shape)
        raise AssertionError(f"{what}: {(err > tol).sum()} of {err.size} differ; worst at {i}: got {a[i]!r} expected {b[i]!r}")


@pytest.fixture(scope="module")
def eng():
    from cnf2freq_b200.engine import Engine
    e = Engine(device=0)
    yield e
    e.close()


def setup_iteration(eng, g, it, first=False):
    """Parameters and control state as the reference held them right before doit() of iteration `it`."""
    b = lambda k: g[f"it{it}_before_{k}"]
    cs = g["chromstarts"]
    C = len(cs) - 1
    if first:
        eng.set_map(g["markerpos"], cs, g["genrec"])
        eng.set_shared_rows(np.zeros(0, np.int32))
    # steady founder flags (iteration 1 is analysed with them too; the reference sets them while it goes, :3170-3177)
    eng.set_pedigree(g["par"], g["empty"], g["it2_before_founder"], g["gen"], g["descendants"], g["dous"])
    al, su, hw = b("alleles").copy(), b("sure").copy(), b("haploweight").copy()
    for c in range(C):
        eng.set_params(c, al, su, hw)
    eng.set_mstep_inputs(b("relhaplo"), b("prioralleles"), b("priorsure"), g["hasprior"], g[f"it{it}_after_children"], g["empty"],
                         g["descendants"])
    sf, ent, o1, o2 = b("scalars")
    eng.set_em_state(sf, ent, int(o1), int(o2))
    eng.begin_iteration()
    return al, su, hw


def check_after(eng, g, it, bufs, name):
    al, su, hw = bufs
    C = len(g["chromstarts"]) - 1
    for c in range(C):
        eng.get_params(c, al, su, hw)
    sf, ent, o1, o2, hit = eng.get_em_state()
    assert np.array_equal(al, g[f"it{it}_after_alleles"]), f"{name} it{it}: imputed alleles"
    assert_prob_close(su, g[f"it{it}_after_sure"], f"{name} it{it} markersure")
    assert_prob_close(hw, g[f"it{it}_after_haploweight"], f"{name} it{it} haploweight")
    assert_close(np.array([sf, ent, o1, o2]), g[f"it{it}_after_scalars"], f"{name} it{it} scalefactor / hitnnn history", rtol=1e-12)


@pytest.mark.parametrize("name", ["mstep_inv_demo", "mstep_inv_fullsib"])
def test_reference_inversions_teacher_forced(eng, name):
    """Per iteration: the reference's parameters before doit(), the engine's own E-steps, the inversions the reference
    applied in each pass handed to cnf2_apply_inversions, cnf2_mstep_pass with the control on the device; the state after
    the iteration equals the reference's.  Also: negshift and its minima against the reference's scores."""
    g = dict(np.load(os.path.join(GOLDEN, name + ".npz")))
    cs = g["chromstarts"].astype(np.int64)
    C = len(cs) - 1
    n_inv = 0
    for k, it in enumerate(int(i) for i in g["iterations"] if i >= 2):
        bufs = setup_iteration(eng, g, it, first=k == 0)
        ids = eng.row_ids()
        K, M = bufs[2].shape
        ns = np.zeros((K, M))
        for c in range(C):
            eng.estep(c)
            eng.get_negshift(c, ns)
            ref_ns = g[f"it{it}_negshift"][:, cs[c]:cs[c + 1]]
            assert_close(ns[:, cs[c]:cs[c + 1]], ref_ns, f"{name} it{it} c{c} negshift", rtol=RTOL, atol=NS_ATOL)
            mv, ms = eng.negshift_minima(c)
            for r, i in enumerate(ids):  # minval / minstart of cnF2freq.cpp:6272-6284 from the reference's own scores
                row = ref_ns[i]
                cand = np.nonzero(row < -1e-10)[0]
                if len(cand) == 0 or row.min() > -1e-9:  # scores within rounding of the threshold are not compared
                    continue
                assert ms[r] == int(np.argmin(row)), (it, c, i, ms[r], int(np.argmin(row)))
                assert abs(mv[r] - row.min()) <= 1e-9 * abs(row.min()) + NS_ATOL
            inv = g[f"it{it}_c{c}_inv"]
            for cc in range(C):  # negshifter runs over the candidates of every chromosome in every pass (:6336-6340)
                sel = [(int(n), int(p) - 1 - int(cs[cc])) for n, p in inv if cs[cc] <= p - 1 < cs[cc + 1]]
                if sel:
                    eng.apply_inversions(cc, [s[0] for s in sel], [s[1] for s in sel])
            n_inv += len(inv)
            eng.mstep_pass(c, any_inversion=len(inv) > 0)
        check_after(eng, g, it, bufs, name)
    assert n_inv > 0


@pytest.mark.parametrize("name", ["mstep_inv_demo", "mstep_inv_fullsib"])
def test_em_pass_makes_the_reference_decisions(eng, name):
    """cnf2_em_pass (candidate selection with inferiorrelated and libc rand() as at :6270-6327, negshifter, the pass, the
    control) over all iterations of the run in order, the random stream started where the reference binary's starts
    (seed 1); parameters are re-set to the reference's at every iteration, decisions are the engine's own."""
    g = dict(np.load(os.path.join(GOLDEN, name + ".npz")))
    cs = g["chromstarts"].astype(np.int64)
    C = len(cs) - 1
    ctypes.CDLL(None).srand(1)
    for k, it in enumerate(int(i) for i in g["iterations"]):
        bufs = setup_iteration(eng, g, it, first=k == 0)
        for c in range(C):
            eng.estep(c)
            inv = eng.em_pass(c)
            assert sorted(inv) == sorted((int(n), int(p)) for n, p in g[f"it{it}_c{c}_inv"]), (it, c, inv)
        if it >= 2:  # the first iteration is analysed with steady founder flags here (include/cnf2_engine.h)
            check_after(eng, g, it, bufs, name)
