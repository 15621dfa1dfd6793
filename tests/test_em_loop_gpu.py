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


def tied(row, a, b):
    """Markers a and b (chromosome-relative) hold the same phase-inversion score up to rounding in the REFERENCE's own row: the
    minimum is then taken where the reference's accumulated log-factors happen to round (they carry ulp(|F|) ~ 3.6e-15 of
    noise, the engine's power-of-two scaling does not), a documented exact tie (negshift_minima_kernel)."""
    return abs(row[a] - row[b]) <= 1e-13 + 1e-9 * abs(row[a])


def same_decisions(mine, ref, ns_ref, cs):
    """Inversion lists [(id, first inverted marker)] agree up to exact ties of the minimum (see tied)."""
    mine, ref = sorted(mine), sorted(ref)
    if len(mine) != len(ref) or [m[0] for m in mine] != [r[0] for r in ref]:
        return False, False
    exact = True
    for (n, p), (_, q) in zip(mine, ref):
        if p == q:
            continue
        exact = False
        c = int(np.searchsorted(cs, p - 1, side="right") - 1)
        if c != int(np.searchsorted(cs, q - 1, side="right") - 1):
            return False, False
        row = ns_ref[n, cs[c]:cs[c + 1]]
        if not tied(row, p - 1 - cs[c], q - 1 - cs[c]):
            return False, False
    return True, exact


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


RUNS = ["mstep_inv_demo", "mstep_inv_fullsib", "mstep_run_f2", "mstep_run_fullsib"]


@pytest.mark.parametrize("name", RUNS)
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
                assert ms[r] == int(np.argmin(row)) or tied(row, ms[r], int(np.argmin(row))), (it, c, i, ms[r], int(np.argmin(row)))
                assert abs(mv[r] - row.min()) <= 1e-9 * abs(row.min()) + NS_ATOL
            inv = g[f"it{it}_c{c}_inv"]
            for cc in range(C):  # negshifter runs over the candidates of every chromosome in every pass (:6336-6340)
                sel = [(int(n), int(p) - 1 - int(cs[cc])) for n, p in inv if cs[cc] <= p - 1 < cs[cc + 1]]
                if sel:
                    eng.apply_inversions(cc, [s[0] for s in sel], [s[1] for s in sel])
            n_inv += len(inv)
            eng.mstep_pass(c, any_inversion=len(inv) > 0)
        check_after(eng, g, it, bufs, name)
    assert n_inv > 0 or not name.startswith("mstep_inv")


@pytest.mark.parametrize("name", RUNS)
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
        exact = True
        for c in range(C):
            eng.estep(c)
            inv = eng.em_pass(c)
            ok, ex = same_decisions(inv, [(int(n), int(p)) for n, p in g[f"it{it}_c{c}_inv"]], g[f"it{it}_negshift"], cs)
            assert ok, (it, c, inv, g[f"it{it}_c{c}_inv"])
            exact &= ex
        # an inversion started one marker away at an exact tie leaves that marker's phase different; the first iteration is
        # analysed with steady founder flags here (include/cnf2_engine.h) -- it reproduces the reference's in three of the
        # four runs, in mstep_run_fullsib a founder's unknown genotype ends elsewhere
        if exact and (it >= 2 or name != "mstep_run_fullsib"):
            check_after(eng, g, it, bufs, name)


@pytest.mark.parametrize("name", RUNS)
def test_free_running_em_loop(eng, name):
    """The device EM loop left alone: the reference's state before its second iteration is uploaded ONCE; from there every
    iteration is cnf2_begin_iteration + per chromosome cnf2_estep + cnf2_em_pass (own inversion decisions, libc rand()
    seeded as the reference binary's stream stood at that point), parameters and the step-size control never leaving the
    device.  Only individ::children is refreshed per iteration (doit increments it on every call, :5246-5259).  After
    every iteration the state is read back and compared with the reference's: imputed calls must be identical and
    markersure / haploweight within 1e-9 for the whole run (6 to 10 iterations, with inversions); the first iteration
    and cell where anything differs is reported."""
    g = dict(np.load(os.path.join(GOLDEN, name + ".npz")))
    cs = g["chromstarts"].astype(np.int64)
    C = len(cs) - 1
    its = [int(i) for i in g["iterations"]]
    # the reference's rand() stream: replay the decisions of iteration 1 (every candidate that reaches :6319 draws once)
    ctypes.CDLL(None).srand(1)
    bufs = setup_iteration(eng, g, its[0], first=True)
    for c in range(C):
        eng.estep(c)
        eng.em_pass(c)
    bufs = setup_iteration(eng, g, its[1])
    K, M = bufs[2].shape
    report = []
    for it in its[1:]:
        if it != its[1]:
            eng.set_mstep_rows(g["hasprior"], g[f"it{it}_after_children"], g["empty"], g["descendants"])
            eng.begin_iteration()
        invs = []
        for c in range(C):
            eng.estep(c)
            invs.append(sorted(eng.em_pass(c)))
        al, su, hw = bufs
        for c in range(C):
            eng.get_params(c, al, su, hw)
        sf, ent, o1, o2, hit = eng.get_em_state()
        same = [same_decisions(invs[c], [(int(n), int(p)) for n, p in g[f"it{it}_c{c}_inv"]], g[f"it{it}_negshift"], cs) for c in range(C)]
        assert all(ok for ok, _ in same), f"{name}: inversion decisions differ in iteration {it}: {invs}"
        if not all(ex for _, ex in same):
            print(f"iteration {it}: an inversion starts at the other marker of an exact tie of the reference's scores; the runs part here")
            break
        d_al = int((al != g[f"it{it}_after_alleles"]).sum())
        d_su = float(np.abs(su - g[f"it{it}_after_sure"]).max())
        d_hw = float(np.abs(hw - g[f"it{it}_after_haploweight"]).max())
        report.append((it, d_al, d_su, d_hw, sf, float(g[f"it{it}_after_scalars"][0])))
    for r in report:
        print("iteration %d: inversions equal, calls differing %d, max |d markersure| %.2e, max |d haploweight| %.2e, scalefactor %.9g (ref %.9g)" % r)
    assert len(report) >= 5 or name == "mstep_run_fullsib", f"{name}: only {len(report)} iterations compared"
    for it, d_al, d_su, d_hw, sf, rsf in report:
        assert d_al == 0, f"{name}: {d_al} imputed calls differ after iteration {it}"
        assert d_su <= 1e-9 and d_hw <= 1e-9, f"{name}: drift after iteration {it}: {d_su:.2e} / {d_hw:.2e}"
        assert abs(sf - rsf) <= 1e-12 * abs(rsf), f"{name}: scalefactor after iteration {it}"


def test_async_readback_equals_the_final_state(eng):
    """cnf2_get_params_packed_async as bench.py's e2e loop and a dumping host use it: alleles and markersure of a chromosome
    right behind its own pass (later passes and inversions of the iteration leave them alone), the haploweights at the
    iteration's end; after cnf2_sync the buffers hold what cnf2_get_params_packed returns."""
    g = dict(np.load(os.path.join(GOLDEN, "mstep_run_f2.npz")))
    cs = g["chromstarts"].astype(np.int64)
    C = len(cs) - 1
    ctypes.CDLL(None).srand(1)
    setup_iteration(eng, g, 7, first=True)  # the reference inverts ten phases per pass in this iteration
    R = len(eng.row_ids())
    mk = lambda: [(np.zeros((R, int(cs[c + 1] - cs[c]), 2), np.uint8), np.zeros((R, int(cs[c + 1] - cs[c]), 2)), np.zeros((R, int(cs[c + 1] - cs[c]))))
                  for c in range(C)]
    back, ref = mk(), mk()
    n_inv = 0
    for c in range(C):
        eng.estep(c)
        n_inv += len(eng.em_pass(c))
        eng.get_params_packed_async(c, back[c][0], back[c][1], None)
    for c in range(C):
        eng.get_params_packed_async(c, None, None, back[c][2])
    eng.sync()
    for c in range(C):
        eng.get_params_packed(c, *ref[c])
        for a, b in zip(back[c], ref[c]):
            assert np.array_equal(a, b)
    print(f"async read-back: {n_inv} inversions applied in the iteration")
