"""The CPU restatement of postmarkerdata (oracle/cnf2_preprocess_oracle.cpp, row f3) against the reference binary: the
inputs as the readers leave them (host/cnf2freq_gpu --parse-only, itself checked against the reference's readers in
test_host_readers.py) -> the state the reference holds before its first EM iteration (tests/golden/mstep_*.npz:
it1_before_*, descendants, sticky founder flags) resp. its iteration-0 state dump (tests/golden/host/*.statedump)."""
import os

import numpy as np
import pytest

import oracle_lib as ol
from test_host_readers import GOLD, HOST, parse


def arrays(args):
    M, cs, inds = parse(args)
    K = max(inds) + 1
    par, exists, empty = np.zeros((K, 2), np.int32), np.zeros(K, np.uint8), np.zeros(K, np.uint8)
    al, su, hw = np.zeros((K, M, 2), np.uint8), np.zeros((K, M, 2)), np.full((K, M), 0.5)
    empty[0] = 1
    for n, d in inds.items():
        par[n], exists[n], empty[n] = d["pars"], 1, d["empty"]
        al[n], su[n], hw[n] = d["alleles"], d["sure"], d["hw"]
    return cs, par, exists, empty, al, su, hw, inds


@pytest.mark.parametrize("name,args", [
    ("mstep_demo", ["--mapfile", HOST + "/refdata/demoplantimpute.map", "--pedfile", HOST + "/refdata/demoplantimpute.ped", "--genfile",
                    HOST + "/refdata/demoplantimpute.gen"]),
    ("mstep_host_fullsib", ["--mapfile", HOST + "/fullsib.map", "--pedfile", HOST + "/fullsib.ped", "--genfile", HOST + "/fullsib.gen"]),
])
def test_oracle_preprocess_reproduces_the_reference_state(name, args):
    g = dict(np.load(os.path.join(GOLD, name + ".npz")))
    cs, par, exists, empty, al, su, hw, _ = arrays(args)
    r = ol.oracle_preprocess(cs, par, exists, empty, al, su, hw)
    ex = exists.astype(bool)
    assert np.array_equal(r["alleles"][ex], g["it1_before_alleles"][ex])
    # (the read-count genotypes of the demo pass through --parse-only's six printed decimals)
    assert np.allclose(r["sure"][ex], g["it1_before_sure"][ex], rtol=1e-12, atol=1e-6)
    assert np.array_equal(r["haploweight"][ex], g["it1_before_haploweight"][ex])
    assert np.array_equal(r["descendants"][ex], g["descendants"][ex])
    assert np.array_equal(r["founder"][ex], g["it1_before_founder"][ex])
    assert not np.array_equal(r["haploweight"], hw)  # phases were locked


@pytest.mark.parametrize("name,args,selfing", [
    ("halfsib", ["--gigimapfile", HOST + "/refdata/halfsibdemo.map", "--gigipedfile", HOST + "/refdata/halfsibdemo.ped"], False),
    # row g1: selfed lines through the reference built with SELFING = true (fixparents / addvariance then run over 256 states)
    ("selfing", ["--mapfile", HOST + "/selfing.map", "--pedfile", HOST + "/selfing.ped", "--genfile", HOST + "/selfing.gen"], True),
])
def test_oracle_preprocess_against_the_iteration_0_dump(name, args, selfing):
    cs, par, exists, empty, al, su, hw, inds = arrays(args)
    r = ol.oracle_preprocess(cs, par, exists, empty, al, su, hw, selfing=selfing)
    lines = [l for l in open(os.path.join(HOST, name + ".statedump")).read().splitlines() if l.strip()]
    M, k = hw.shape[1], 0
    for n in sorted(inds):
        assert lines[k].split() == [str(n), inds[n]["name"]]
        for j in range(M):
            t = lines[k + 1 + j].split()
            assert abs(float(t[0]) - r["haploweight"][n, j]) < 1e-6 and [int(t[1]), int(t[2])] == r["alleles"][n, j].tolist()
            assert abs(float(t[4]) - r["sure"][n, j, 0]) < 2e-6 and abs(float(t[5]) - r["sure"][n, j, 1]) < 2e-6
        k += 1 + M
