"""The readers of the C++ host program (host/cnf2freq_gpu --parse-only: no GPU involved) against what the REFERENCE's readers
built from the same files -- pedigree links, implicit F1s, generations, empty flags, the analysed set, chromosome starts
and every individual's genotypes / error probabilities as read (incl. the read-count genotypes `a/b` of the demo, which go
through the binomial mixture of cnF2freq.cpp:6599-6657) -- taken from the reference binary's dumps in tests/golden/mstep_*.npz
(priors = the genotypes as read, :6664-6665)."""
import os
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLD = os.path.join(ROOT, "tests", "golden")
HOST = os.path.join(GOLD, "host")


def parse(args):
    subprocess.check_call(["make", "-C", os.path.join(ROOT, "host"), "cnf2freq_gpu"], stdout=subprocess.DEVNULL)
    r = subprocess.run([os.path.join(ROOT, "host", "cnf2freq_gpu")] + args + ["--parse-only"], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    lines = [l for l in r.stdout.splitlines()]
    at = next(i for i, l in enumerate(lines) if l.startswith("markers "))
    head = lines[at].split()
    M, C, K1 = int(head[1]), int(head[3]), int(head[5])
    cs = [int(l.split()[1]) for l in lines[at + 1:at + 2 + C]]
    inds, k = {}, at + 2 + C
    while k < len(lines):
        t = lines[k].split()
        n = int(t[0])
        rows = [lines[k + 1 + j].split() for j in range(M)]
        inds[n] = dict(name=t[1], pars=(int(t[3]), int(t[4])), gen=int(t[6]), empty=int(t[8]), prior=int(t[10]), analysed=int(t[12]),
                       alleles=np.array([[int(r[1]), int(r[2])] for r in rows]), sure=np.array([[float(r[4]), float(r[5])] for r in rows]),
                       hw=np.array([float(r[0]) for r in rows]))
        k += 1 + M
    assert len(inds) == K1
    return M, cs, inds


@pytest.mark.parametrize("name,args", [
    ("mstep_demo", ["--mapfile", HOST + "/refdata/demoplantimpute.map", "--pedfile", HOST + "/refdata/demoplantimpute.ped", "--genfile",
                    HOST + "/refdata/demoplantimpute.gen"]),
    ("mstep_host_fullsib", ["--mapfile", HOST + "/fullsib.map", "--pedfile", HOST + "/fullsib.ped", "--genfile", HOST + "/fullsib.gen"]),
])
def test_plantimpute_readers_build_what_the_reference_builds(name, args):
    g = dict(np.load(os.path.join(GOLD, name + ".npz")))
    M, cs, inds = parse(args)
    assert cs == g["chromstarts"].tolist() and M == len(g["markerpos"])
    K = g["par"].shape[0]
    assert sorted(inds) == [i for i in range(1, K) if g["exists"][i]]
    assert [n for n in sorted(inds) if inds[n]["analysed"]] == sorted(g["dous"].tolist())
    for n, d in inds.items():
        assert d["pars"] == tuple(int(x) for x in g["par"][n]), (n, d["name"])
        assert d["empty"] == int(g["empty"][n]) and d["prior"] == int(g["hasprior"][n]), (n, d["name"])
        if 0 <= int(g["gen"][n]) < 10:  # individuals the pedigree file never lists keep an uninitialised gen in the reference
            assert d["gen"] == int(g["gen"][n]), (n, d["name"])
        if d["prior"]:
            assert np.array_equal(d["alleles"], g["it1_before_prioralleles"][n]), (n, d["name"])
            assert np.allclose(d["sure"], g["it1_before_priorsure"][n], rtol=0, atol=1e-6), (n, d["name"])
        assert np.all(d["hw"] == 0.5)


def test_gigi_reader():
    """halfsibdemo (readgigidata, :7450-7515): everybody is analysed, genotypes 0/1/2 with markersure 0, one chromosome."""
    M, cs, inds = parse(["--gigimapfile", HOST + "/refdata/halfsibdemo.map", "--gigipedfile", HOST + "/refdata/halfsibdemo.ped"])
    assert M == 5 and cs == [0, 5] and len(inds) == 5 and all(d["analysed"] and d["prior"] and not d["empty"] for d in inds.values())
    names = {d["name"]: n for n, d in inds.items()}
    assert inds[names["o1"]]["pars"] == (names["sire"], names["d1"]) and inds[names["sire"]]["pars"] == (0, 0)
    assert inds[names["sire"]]["alleles"].tolist()[:3] == [[1, 1], [1, 1], [1, 2]] and np.all(inds[names["sire"]]["sure"] == 0)
