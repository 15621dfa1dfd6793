"""Golden fixtures of the M-step (rows a6 / f1 / f2) from the REFERENCE's own code: for whole EM iterations of
oracle/_ref/cnf2ref (single-threaded) the parameters before doit(), the EM statistics as they stood before each
chromosome's M-step pass ("FIRST PASS" hook of oracle/ref_driver.cpp), the phase-inversion scores (negshift) and the
inversions the reference applied in each pass (its "Inv: <id> <marker>" lines, cnF2freq.cpp:3450), and the
parameters + step-size scalars after doit().  mstep_demo / mstep_f2 hold iterations without inversions,
mstep_inv_* iterations with them.  Only runs where /root/reference exists; the fixtures are committed.

usage: python tests/golden/make_mstep_golden.py [selfing]     (selfing: only mstep_run_selfing, from oracle/_ref/cnf2ref_selfing)
"""
import os
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


def passes(stdout, C):
    """Per pass (one per chromosome per iteration, closed by the "Scale factor now" line, :6392): the inversions applied."""
    out, cur = [], []
    for line in stdout.splitlines():
        if line.startswith("Inv: "):
            n, p = line.split()[1:3]
            cur.append((int(n), int(p)))
        elif line.startswith("Scale factor now"):
            out.append(cur)
            cur = []
    return out


def collect(prefix, iterations, stdout=None, allow_inversions=False, stats=True):
    prob = read_dump(prefix + ".problem")
    C = len(prob["chromstarts"]) - 1
    inv = passes(stdout, C) if stdout is not None else None
    out = dict(par=np.stack([prob["par0"], prob["par1"]], axis=1).astype(np.int32), exists=prob["exists"], empty=prob["empty"],
               hasprior=prob["hasprior"], descendants=prob["descendants"], dous=prob["dous"],
               chromstarts=prob["chromstarts"].astype(np.uint32), iterations=np.asarray(iterations), gen=prob["gen"],
               markerpos=prob["markerposes"], genrec=prob["genrec"])
    for it in iterations:
        a, b = read_dump(f"{prefix}.iter{it}.params"), read_dump(f"{prefix}.iter{it}.state")
        assert allow_inversions or b["negshift"].min() > -1e-10, "phase inversion in this iteration"
        out[f"it{it}_negshift"] = b["negshift"]
        if inv is not None:
            for c in range(C):
                lst = inv[(it - 1) * C + c]
                out[f"it{it}_c{c}_inv"] = np.asarray(lst, np.int32).reshape(-1, 2)
        for k in ("alleles", "sure", "haploweight", "relhaplo", "prioralleles", "priorsure", "scalars", "founder"):
            out[f"it{it}_before_{k}"] = a[k]
        for k in ("alleles", "sure", "haploweight", "scalars", "children"):  # children is rebuilt inside doit (:5246-5259)
            out[f"it{it}_after_{k}"] = b[k]
        for c in range(C if stats else 0):
            st = read_dump(f"{prefix}.iter{it}.chrom{c}.stats")
            for k in ("haplobase", "haplocount", "infprobs", "homozyg"):
                out[f"it{it}_c{c}_{k}"] = st[k]
    return out


def main_selfing():
    """mstep_run_selfing: 8 EM iterations of the SELFING = true build on selfed lines of generation 2-6."""
    env = dict(os.environ, OMP_NUM_THREADS="1")
    with tempfile.TemporaryDirectory() as tmp:
        sim = synth.simulate_selfed_lines(10, 2, 30, 1.2, 23, [2, 3, 4, 5, 6, 3, 4, 2, 5, 4])
        synth.write_alpha_files(tmp + "/sl", sim)
        args = ["--mapfile", tmp + "/sl.map", "--pedfile", tmp + "/sl.ped", "--genfile", tmp + "/sl.gen"]
        r = subprocess.run([BIN_SELFING] + args + ["--count", "9", "--dump", tmp + "/sl"], env=env, capture_output=True, text=True, check=True)
        g = collect(tmp + "/sl", list(range(1, 9)), r.stdout, True, stats=False)
        print("inversions per iteration:", [sum(len(g[f"it{it}_c{c}_inv"]) for c in range(2)) for it in range(1, 9)])
        np.savez_compressed(os.path.join(HERE, "mstep_run_selfing.npz"), **g)
    print("mstep_run_selfing.npz", os.path.getsize(os.path.join(HERE, "mstep_run_selfing.npz")), "bytes")


def main():
    env = dict(os.environ, OMP_NUM_THREADS="1")
    with tempfile.TemporaryDirectory() as tmp:
        args = ["--mapfile", f"{REF}/demoplantimpute.map", "--pedfile", f"{REF}/demoplantimpute.ped", "--genfile",
                f"{REF}/demoplantimpute.gen"]
        subprocess.run([BIN] + args + ["--count", "4", "--dump", tmp + "/demo"], env=env, capture_output=True, check=True)
        np.savez_compressed(os.path.join(HERE, "mstep_demo.npz"), **collect(tmp + "/demo", [1, 2, 3]))
        sim = synth.simulate_f2(n_f2=6, n_chrom=2, markers_per_chrom=24, spacing_cM=1.5, missing=0.2, seed=3)
        synth.write_alpha_files(tmp + "/f2", sim)
        args = ["--mapfile", tmp + "/f2.map", "--pedfile", tmp + "/f2.ped", "--genfile", tmp + "/f2.gen"]
        subprocess.run([BIN] + args + ["--count", "3", "--dump", tmp + "/f2"], env=env, capture_output=True, check=True)
        np.savez_compressed(os.path.join(HERE, "mstep_f2.npz"), **collect(tmp + "/f2", [1, 2]))
        # iterations in which the reference inverts phases (demo: 6, 8, 9, 10; two-chromosome full-sib families: from 2 on)
        args = ["--mapfile", f"{REF}/demoplantimpute.map", "--pedfile", f"{REF}/demoplantimpute.ped", "--genfile",
                f"{REF}/demoplantimpute.gen"]
        r = subprocess.run([BIN] + args + ["--count", "12", "--dump", tmp + "/inv"], env=env, capture_output=True, text=True, check=True)
        g = collect(tmp + "/inv", list(range(1, 12)), r.stdout, True)
        assert any(len(g[f"it{it}_c0_inv"]) for it in range(1, 12))
        np.savez_compressed(os.path.join(HERE, "mstep_inv_demo.npz"), **g)
        sim = synth.simulate_fullsib(n_fam=2, n_off=3, n_chrom=2, markers_per_chrom=12, spacing_cM=2.0, missing=0.1, seed=4)
        synth.write_alpha_files(tmp + "/fs", sim)
        args = ["--mapfile", tmp + "/fs.map", "--pedfile", tmp + "/fs.ped", "--genfile", tmp + "/fs.gen"]
        r = subprocess.run([BIN] + args + ["--count", "7", "--dump", tmp + "/fs"], env=env, capture_output=True, text=True, check=True)
        g = collect(tmp + "/fs", list(range(1, 7)), r.stdout, True)
        assert any(len(g[f"it{it}_c{c}_inv"]) for it in range(1, 7) for c in range(2))
        np.savez_compressed(os.path.join(HERE, "mstep_inv_fullsib.npz"), **g)
        # longer runs for the free-running EM loop test (no per-chromosome statistics: before / after / scores / inversions only)
        sim = synth.simulate_f2(n_f2=16, n_chrom=2, markers_per_chrom=60, spacing_cM=0.8, missing=0.2, seed=13)
        synth.write_alpha_files(tmp + "/rf2", sim)
        args = ["--mapfile", tmp + "/rf2.map", "--pedfile", tmp + "/rf2.ped", "--genfile", tmp + "/rf2.gen"]
        r = subprocess.run([BIN] + args + ["--count", "9", "--dump", tmp + "/rf2"], env=env, capture_output=True, text=True, check=True)
        np.savez_compressed(os.path.join(HERE, "mstep_run_f2.npz"), **collect(tmp + "/rf2", list(range(1, 9)), r.stdout, True, stats=False))
        sim = synth.simulate_fullsib(n_fam=6, n_off=4, n_chrom=1, markers_per_chrom=40, spacing_cM=1.0, missing=0.1, seed=14)
        synth.write_alpha_files(tmp + "/rfs", sim)
        args = ["--mapfile", tmp + "/rfs.map", "--pedfile", tmp + "/rfs.ped", "--genfile", tmp + "/rfs.gen"]
        r = subprocess.run([BIN] + args + ["--count", "8", "--dump", tmp + "/rfs"], env=env, capture_output=True, text=True, check=True)
        np.savez_compressed(os.path.join(HERE, "mstep_run_fullsib.npz"), **collect(tmp + "/rfs", list(range(1, 8)), r.stdout, True, stats=False))
    for n in ("mstep_demo.npz", "mstep_f2.npz", "mstep_inv_demo.npz", "mstep_inv_fullsib.npz", "mstep_run_f2.npz", "mstep_run_fullsib.npz"):
        print(n, os.path.getsize(os.path.join(HERE, n)), "bytes")


if __name__ == "__main__":
    if sys.argv[1:] == ["selfing"]:
        main_selfing()
    else:
        main()
        main_selfing()
