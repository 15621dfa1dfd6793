"""Generates the golden fixtures in this directory from the REFERENCE's own code.

Runs oracle/_ref/cnf2ref (the unmodified reference translation unit + our dump driver, built by
oracle/build_ref.sh from /root/reference) single-threaded on
  * the reference's two demo data sets (demoplantimpute.*, halfsibdemo.*),
  * small synthetic pedigrees written in the reference's .map/.ped/.gen format by cnf2freq_b200.synth,
and stores, per case, the E-step inputs of one EM iteration (pedigree, map, parameters as the reference
held them right before doit()) and what the reference computed from them: the per-shift forward/backward
arrays of some individuals, the EM sufficient statistics of every chromosome as they stood before the
M-step, and the log-likelihood lines it printed.  Only runs where /root/reference exists; the fixtures
are committed.

usage: python tests/golden/make_golden.py [case ...]
"""
import os
import re
import subprocess
import sys
import tempfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from cnf2dump import read_dump  # noqa: E402
from cnf2freq_b200 import synth  # noqa: E402

REF = os.environ.get("CNF2_REFERENCE_DIR", "/root/reference")
BIN = os.path.join(ROOT, "oracle", "_ref", "cnf2ref")
BIN_SELFING = os.path.join(ROOT, "oracle", "_ref", "cnf2ref_selfing")  # settings.h with SELFING = true (row g1)
LINE = re.compile(r"^(\d+),(\d+),(\d+): (\S+)\t(\S+) (\d+)$")


def run_ref(args, prefix, count, binary=None):
    env = dict(os.environ, OMP_NUM_THREADS="1")
    cmd = [binary or BIN] + args + ["--count", str(count), "--dump", prefix, "--dumpfwbw"]
    res = subprocess.run(cmd, env=env, capture_output=True, text=True, check=True)
    return res.stdout


def collect(prefix, stdout, it, fwbw_individuals):
    prob = read_dump(prefix + ".problem")
    par = read_dump(f"{prefix}.iter{it}.params")
    cs = prob["chromstarts"].astype(np.int64)
    C = len(cs) - 1
    K, M = par["haploweight"].shape
    out = dict(
        par=np.stack([prob["par0"], prob["par1"]], axis=1).astype(np.int32), empty=prob["empty"], founder=par["founder"],
        gen=prob["gen"], descendants=prob["descendants"], dous=prob["dous"], markerpos=prob["markerposes"],
        chromstarts=prob["chromstarts"].astype(np.uint32), genrec=prob["genrec"], alleles=par["alleles"],
        sure=par["sure"], haploweight=par["haploweight"], iteration=np.array(it))
    stats = dict(haplobase=np.zeros((K, M)), haplocount=np.zeros((K, M)), infprobs=np.zeros((K, M, 2, 2)),
                 homozyg=np.zeros((K, M, 2)))
    for c in range(C):
        st = read_dump(f"{prefix}.iter{it}.chrom{c}.stats")
        assert int(st["oddkeys"][0]) == 0
        for k in stats:
            stats[k][:, cs[c]:cs[c + 1]] = st[k][:, cs[c]:cs[c + 1]]
    out.update(stats)
    # phase-inversion scores as the whole doit() left them (accumulated by updatenegshifts, cnF2freq.cpp:3640-3715;
    # nothing after the E-step touches them until main() zeroes them at :8187)
    out["negshift"] = read_dump(f"{prefix}.iter{it}.state")["negshift"]
    # printed log-likelihood pairs of this iteration: lines come chromosome by chromosome, in dous order
    lines = [LINE.match(l) for l in stdout.splitlines()]
    lines = [m for m in lines if m]
    nd = len(prob["dous"])
    per_iter = nd * C
    sel = lines[(it - 1) * per_iter: it * per_iter]
    assert len(sel) == per_iter, (len(lines), per_iter, it)
    ll = np.zeros((nd, C, 2))
    for i, m in enumerate(sel):
        c, j = divmod(i, nd)
        assert int(m.group(1)) == int(prob["dous"][j])
        ll[j, c] = float(m.group(4)), float(m.group(5))
    out["loglik_printed"] = ll
    fw = read_dump(f"{prefix}.iter{it}.fwbw")
    for (c, n) in fwbw_individuals:
        out[f"fwbw_c{c}_n{n}"] = fw[f"fwbw_c{c}_n{n}"]
        out[f"fwbwfactors_c{c}_n{n}"] = fw[f"fwbwfactors_c{c}_n{n}"]
        out[f"fwbwmask_c{c}_n{n}"] = fw[f"fwbwmask_c{c}_n{n}"]
    return out


def case_demo(tmp):
    args = ["--mapfile", f"{REF}/demoplantimpute.map", "--pedfile", f"{REF}/demoplantimpute.ped", "--genfile",
            f"{REF}/demoplantimpute.gen"]
    out = run_ref(args, tmp + "/demo", 4)
    return {"demo_iter1": collect(tmp + "/demo", out, 1, [(0, 3)]), "demo_iter3": collect(tmp + "/demo", out, 3, [(0, 10)])}


def case_halfsib(tmp):
    args = ["--gigimapfile", f"{REF}/halfsibdemo.map", "--gigipedfile", f"{REF}/halfsibdemo.ped"]
    out = run_ref(args, tmp + "/hs", 3)
    return {"halfsib_iter2": collect(tmp + "/hs", out, 2, [(0, 3), (0, 4)])}


def case_f2(tmp):
    sim = synth.simulate_f2(n_f2=6, n_chrom=2, markers_per_chrom=24, spacing_cM=1.5, missing=0.2, seed=3)
    synth.write_alpha_files(tmp + "/f2", sim)
    args = ["--mapfile", tmp + "/f2.map", "--pedfile", tmp + "/f2.ped", "--genfile", tmp + "/f2.gen"]
    out = run_ref(args, tmp + "/f2", 3)
    prob = read_dump(tmp + "/f2.problem")
    n0 = int(prob["dous"][0])
    return {"f2small_iter2": collect(tmp + "/f2", out, 2, [(0, n0), (1, n0)])}


def case_fullsib(tmp):
    sim = synth.simulate_fullsib(n_fam=1, n_off=2, n_chrom=1, markers_per_chrom=10, spacing_cM=2.0, missing=0.1, seed=4)
    synth.write_alpha_files(tmp + "/fs", sim)
    args = ["--mapfile", tmp + "/fs.map", "--pedfile", tmp + "/fs.ped", "--genfile", tmp + "/fs.gen"]
    out = run_ref(args, tmp + "/fs", 3)
    prob = read_dump(tmp + "/fs.problem")
    n0 = int(prob["dous"][0])
    return {"fullsib_iter2": collect(tmp + "/fs", out, 2, [(0, n0)])}


def case_selfing_demo(tmp):
    """The PlantImpute demo under SELFING = true: individual F is listed with generation 20 (18 generations of selfing)."""
    args = ["--mapfile", f"{REF}/demoplantimpute.map", "--pedfile", f"{REF}/demoplantimpute.ped", "--genfile",
            f"{REF}/demoplantimpute.gen"]
    out = run_ref(args, tmp + "/sdemo", 3, BIN_SELFING)
    return {"selfing_demo_iter2": collect(tmp + "/sdemo", out, 2, [(0, 10)])}


def case_selfing_lines(tmp):
    """An F2 cross whose lines were selfed for 0-4 further generations (ped generation 2-6), SELFING = true."""
    sim = synth.simulate_f2(n_f2=6, n_chrom=2, markers_per_chrom=14, spacing_cM=2.5, missing=0.2, seed=8)
    rng = np.random.default_rng(8)
    gens = [2, 3, 4, 4, 6, 3]
    sim["ped"] = sim["ped"][:2] + [(me, fa, mo, g) for (me, fa, mo, _), g in zip(sim["ped"][2:], gens)]
    tok = sim["tokens"]
    for k, g in enumerate(gens):
        row = tok[2 + k]
        fix = (row == 1) & (rng.random(row.shape) > 0.5 ** (g - 2))  # a heterozygote survives a selfing with probability 1/2
        row[fix] = rng.choice([0, 2], size=int(fix.sum()))
    synth.write_alpha_files(tmp + "/sl", sim)
    args = ["--mapfile", tmp + "/sl.map", "--pedfile", tmp + "/sl.ped", "--genfile", tmp + "/sl.gen"]
    out = run_ref(args, tmp + "/sl", 3, BIN_SELFING)
    prob = read_dump(tmp + "/sl.problem")
    d = prob["dous"]
    return {"selfing_lines_iter2": collect(tmp + "/sl", out, 2, [(0, int(d[2])), (1, int(d[4]))])}


CASES = dict(demo=case_demo, halfsib=case_halfsib, f2=case_f2, fullsib=case_fullsib, selfing_demo=case_selfing_demo,
             selfing_lines=case_selfing_lines)

if __name__ == "__main__":
    if not os.path.exists(BIN):
        subprocess.check_call(["sh", os.path.join(ROOT, "oracle", "build_ref.sh")])
    which = sys.argv[1:] or list(CASES)
    for name in which:
        with tempfile.TemporaryDirectory() as tmp:
            for key, arrays in CASES[name](tmp).items():
                path = os.path.join(HERE, key + ".npz")
                np.savez_compressed(path, **arrays)
                print(key, os.path.getsize(path), "bytes")
