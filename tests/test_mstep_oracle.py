"""The M-step restatement (oracle/cnf2_mstep_oracle.cpp) against what the reference binary itself did in whole EM
iterations (tests/golden/mstep_*.npz, made by tests/golden/make_mstep_golden.py): new alleles, markersure,
haploweight, hitnnn history and scalefactor, bit for bit -- both builds run the same libm on the same host."""
import os

import numpy as np
import pytest

import oracle_lib as ol
from cases import GOLDEN


def replay(g, it, mstep):
    """Replays iteration `it`: one M-step pass per chromosome, in order, as doit() does; returns the final state."""
    cs = g["chromstarts"]
    C = len(cs) - 1
    b = lambda k: g[f"it{it}_before_{k}"]
    al, su, hw = b("alleles"), b("sure"), b("haploweight")
    sf, ent, o1, o2 = b("scalars")
    for c in range(C):
        st = {k: g[f"it{it}_c{c}_{k}"] for k in ("haplobase", "haplocount", "infprobs", "homozyg")}
        r = mstep(cs, g["exists"], g["empty"], g[f"it{it}_after_children"], g["descendants"], al, su, hw, b("relhaplo"),
                  g["hasprior"], b("prioralleles"), b("priorsure"), st["haplobase"], st["haplocount"], st["infprobs"],
                  st["homozyg"], sf, ent)
        al, su, hw = r["alleles"], r["sure"], r["haploweight"]
        sf, o1, o2 = ol.oracle_scalefactor(sf, r["hitnnn"], len(g["dous"]), o1, o2)
    return al, su, hw, np.array([sf, ent, o1, o2])


@pytest.mark.parametrize("name", ["mstep_demo", "mstep_f2"])
def test_oracle_mstep_reproduces_the_reference(name):
    g = dict(np.load(os.path.join(GOLDEN, name + ".npz")))
    for it in g["iterations"]:
        al, su, hw, sc = replay(g, it, ol.oracle_mstep)
        assert np.array_equal(al, g[f"it{it}_after_alleles"])
        assert np.array_equal(su, g[f"it{it}_after_sure"])
        assert np.array_equal(hw, g[f"it{it}_after_haploweight"])
        assert np.array_equal(sc, g[f"it{it}_after_scalars"])
        assert not np.array_equal(hw, g[f"it{it}_before_haploweight"])  # the pass did something
