"""Shared helpers of the parity tests: golden fixtures -> oracle Problem, comparison of E-step outputs."""
import glob
import os

import numpy as np

import oracle_lib as ol

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
GOLDEN_CASES = sorted(os.path.splitext(os.path.basename(p))[0] for p in glob.glob(os.path.join(GOLDEN, "*.npz"))
                      if not os.path.basename(p).startswith(("mstep_", "selfing_")))
# written by oracle/_ref/cnf2ref_selfing, the SELFING = true instantiation of the reference (row g1)
SELFING_CASES = sorted(os.path.splitext(os.path.basename(p))[0] for p in glob.glob(os.path.join(GOLDEN, "selfing_*.npz")))
STATS = ("haplobase", "haplocount", "infprobs", "homozyg")
RTOL = 1e-9  # BASELINE.json north_star: posteriors within 1e-9 relative (fp64)
# phase-inversion scores (row f2): differences of logarithms that cancel to rounding where a turn changes nothing, summed
# over the units an individual is a relative of -- hence an absolute part (the reference acts on scores < -1e-10, :6277)
NS_ATOL = 1e-11


def load_golden(name):
    g = dict(np.load(os.path.join(GOLDEN, name + ".npz")))
    pb = ol.Problem(g["par"], g["empty"], g["founder"], g["gen"], g["descendants"], g["dous"], g["markerpos"],
                    g["chromstarts"], g["genrec"], g["alleles"], g["sure"], g["haploweight"])
    return g, pb


def assert_close(a, b, what, rtol=RTOL, atol=0.0):
    """a (tested) against b (expected): NaNs must coincide, the rest within rtol relative / atol absolute."""
    a = np.asarray(a, np.float64)
    b = np.asarray(b, np.float64)
    assert a.shape == b.shape, (what, a.shape, b.shape)
    na, nb = np.isnan(a), np.isnan(b)
    assert np.array_equal(na, nb), f"{what}: NaN pattern differs ({na.sum()} vs {nb.sum()})"
    a = np.where(na, 0.0, a)
    b = np.where(nb, 0.0, b)
    err = np.abs(a - b)
    tol = atol + rtol * np.abs(b)
    bad = err > tol
    if bad.any():
        i = np.unravel_index(np.argmax(err - tol), a.shape)
        raise AssertionError(f"{what}: {bad.sum()} of {a.size} differ; worst at {i}: got {a[i]!r} expected {b[i]!r}")


def engine_problem(eng, pb, dous=None, all_dous=None, shared=None):
    """Feeds an oracle_lib.Problem to a cnf2freq_b200.engine.Engine."""
    eng.set_map(pb.markerpos, pb.chromstarts, pb.genrec)
    if shared is not None:
        eng.set_shared_rows(shared)
    eng.set_pedigree(pb.par, pb.empty, pb.founder, pb.gen, pb.descendants, pb.dous if dous is None else dous,
                     pb.dous if all_dous is None else all_dous)
