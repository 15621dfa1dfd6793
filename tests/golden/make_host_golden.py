"""Golden text outputs for the C++ host program (host/cnf2freq_gpu, row f4): what the REFERENCE's main writes per
iteration (the state dump of cnF2freq.cpp:8157-8192, via oracle/ref_driver.cpp --statedump, which restates those printf
lines over the reference's own data structures) and its stdout progress lines, for the two data sets the reference ships
(demo.sh:37, halfsibdemo) and two small synthetic pedigrees.  Single-threaded (deterministic).  Only runs where
/root/reference exists; the outputs are committed.

usage: python tests/golden/make_host_golden.py [selfing]      (selfing: only the SELFING = true run)
"""
import os
import subprocess
import sys
import tempfile

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
from cnf2freq_b200 import synth  # noqa: E402

REF = os.environ.get("CNF2_REFERENCE_DIR", "/root/reference")
BIN = os.path.join(ROOT, "oracle", "_ref", "cnf2ref")
BIN_SELFING = os.path.join(ROOT, "oracle", "_ref", "cnf2ref_selfing")  # settings.h with SELFING = true (row g1)
OUT = os.path.join(HERE, "host")


def run(name, args, count, binary=BIN):
    env = dict(os.environ, OMP_NUM_THREADS="1")
    sd = os.path.join(OUT, name + ".statedump")
    r = subprocess.run([binary] + args + ["--count", str(count), "--statedump", sd, "--quiet"], env=env, capture_output=True, text=True, check=True)
    keep = [l for l in r.stdout.splitlines() if l.startswith(("Correction at", "Fixing point", "Inv:", "Scale factor now")) or
            (l[:1].isdigit() and "," in l.split(":")[0])]
    open(os.path.join(OUT, name + ".stdout"), "w").write("\n".join(keep) + "\n")
    print(name, os.path.getsize(sd), "bytes,", len(keep), "progress lines")


def main_selfing():
    """Selfed PlantImpute lines (generation 2-6) through the reference built with SELFING = true."""
    os.makedirs(OUT, exist_ok=True)
    prefix = os.path.join(OUT, "selfing")
    synth.write_alpha_files(prefix, synth.simulate_selfed_lines(6, 2, 14, 2.0, 25, [3, 4, 2, 6, 5, 4]))
    run("selfing", ["--mapfile", prefix + ".map", "--pedfile", prefix + ".ped", "--genfile", prefix + ".gen"], 4, BIN_SELFING)


def main():
    os.makedirs(OUT, exist_ok=True)
    run("demo", ["--mapfile", f"{REF}/demoplantimpute.map", "--pedfile", f"{REF}/demoplantimpute.ped", "--genfile", f"{REF}/demoplantimpute.gen"], 6)
    run("halfsib", ["--gigimapfile", f"{REF}/halfsibdemo.map", "--gigipedfile", f"{REF}/halfsibdemo.ped"], 4)
    for name, sim in (("f2", synth.simulate_f2(n_f2=5, n_chrom=2, markers_per_chrom=14, spacing_cM=1.5, missing=0.25, seed=23)),
                      ("fullsib", synth.simulate_fullsib(n_fam=2, n_off=3, n_chrom=1, markers_per_chrom=16, spacing_cM=2.0, missing=0.15, seed=24))):
        prefix = os.path.join(OUT, name)
        synth.write_alpha_files(prefix, sim)
        run(name, ["--mapfile", prefix + ".map", "--pedfile", prefix + ".ped", "--genfile", prefix + ".gen"], 4)


if __name__ == "__main__":
    if sys.argv[1:] == ["selfing"]:
        main_selfing()
    else:
        main()
        main_selfing()
