#!/usr/bin/env python
"""Benchmark of the cnF2freq hot path on B200: individual-marker-states per second for forward-backward + EM.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
    torchrun --nproc-per-node N bench.py --gpus N ...          (one rank per GPU, NCCL)

A "step" is one whole EM iteration on the device = what the reference's doit() does (cnF2freq.cpp:5189-6410): per
chromosome the E-step (sweeps for 8 shift modes, posterior scatter, phase-inversion scores) and the block behind it
(:6239-6392: phase-inversion candidates, negshifter, processinfprobs + updatehaploweights, step-size control), the
parameters staying on the device and EVOLVING from iteration to iteration.  W warm-up iterations are not timed.

Workload (BASELINE.json configs[2]): synthetic F2 intercross, 2,000 analysed individuals x 10,000 markers (10 chromosomes
x 1,000), 20 % missing genotypes, PER GPU (weak scaling: the cross has N x 2,000 F2s with distinct ids, every rank
analyses its 2,000; the two founders' statistics rows are all-reduced over NCCL once per chromosome).

value        K EM iterations, parameters resident in HBM; CUDA events on the engine's streams; max over ranks.
e2e          the same through the C ABI with pinned HOST buffers inside the timed region: the upload of all parameters and
             M-step inputs (once, amortised over the K iterations) and, every iteration, the read-back of every
             chromosome's parameters (what the per-iteration dump of :8157-8192 prints).
estep        sub-key: the E-step alone (no M-step), as last round's headline.
roofline     per dominant kernel.  warp_estep_kernel: algorithmic flops of SURVEY.md 8(d) / the launch
             duration (E-steps back to back, launches in flight; the isolated launch beside it); peaks: fp64 measured in
             this run (MEASURED_PEAKS.json has no fp64 figure), HBM from MEASURED_PEAKS.json.  mstep_infprobs_kernel /
             mstep_haploweight_kernel: gradient evaluations counted on the device x flops per evaluation.
configs      (N = 1 only) two more timed workloads in the same line: the full-sib shape of configs[3] and one
             cfg5-shaped chromosome.
parity_check before any timing: small pedigrees through the SAME N-rank path (sharding, all-reduce hook, device M-step)
             against the CPU oracle (test infrastructure; used here only as the checker).  Non-zero exit on mismatch.
cpu_baseline / --impl reference: the reference's own code (oracle/_ref/cnf2ref_fast, built from /root/reference by
             oracle/build_ref.sh with demo.sh's flags) on all host cores, on a bounded sample of the same workload, same
             warm-up / iteration window; it times the reference's whole doit().
"""
import argparse
import json
import os
import re
import statistics
import subprocess
import sys
import tempfile
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

NUMTYPES = 64
# per (individual-marker, shift mode) E + 2 sweep steps + P + 2 N T = 640 + 1,920 + 5,888 + 128 T flop;
# T = 8 turn masks are scored now that row f2 is built (the DOTOULBAR 0 path: no turn + the 7 single-individual turns)
N_TURNS = 8
FLOP_PER_INDMARKER = (8448 + 2 * NUMTYPES * N_TURNS) * 8
BYTES_PER_INDMARKER = 238 + 8 + 448 + 7 * 8  # 34 B x 7 read, 8 B interval, 64 B x 7 accumulated (+ 8 B negshift x 7)
# one gradient evaluation of the step solver (ms_actual, cnf2_mstep_kernel.cuh): 14 (MODE 0) / 20 (MODE 1) fp64 operations
# besides one logarithm (own series: 34 fp64 operations) and 2 / 3 reciprocals (8 each: seed + two Newton steps)
FLOP_PER_GRADIENT = (14 + 34 + 2 * 8, 20 + 34 + 3 * 8)
METRIC = "individual-marker-states/sec fwd-bwd+EM"
UNIT = "individual-marker-states/s"


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="f2", choices=["f2", "fullsib", "cfg5chrom", "selfing"],
                    help="f2: BASELINE.json configs[2] (the bench workload); fullsib: configs[3]'s pedigree shape (families of 20 "
                         "offspring); cfg5chrom: one chromosome of configs[4]'s shape (biparental F2 families of 200 from 100 lines)")
    ap.add_argument("--individuals", type=int, default=None, help="analysed individuals per GPU")
    ap.add_argument("--chroms", type=int, default=None)
    ap.add_argument("--markers-per-chrom", type=int, default=None)
    ap.add_argument("--cpu-sample-individuals", type=int, default=64)
    ap.add_argument("--cpu-sample-markers", type=int, default=500)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-extra-configs", action="store_true", help="skip the fullsib and cfg5chrom lines of `configs`")
    ap.add_argument("--no-parity-check", action="store_true")
    ap.add_argument("--strong", action="store_true",
                    help="strong scaling: --individuals is the size of the whole pedigree, dealt to the ranks (default: per GPU, weak scaling)")
    a = ap.parse_args()
    a.explicit_shape = a.individuals is not None or a.chroms is not None or a.markers_per_chrom is not None
    return a


SHAPES = {"f2": (2000, 10, 1000), "fullsib": (1320, 1, 2000), "cfg5chrom": (20000, 1, 5000), "selfing": (296, 1, 500)}


def shape_of(a, workload):
    n, c, m = SHAPES[workload]
    if workload == a.workload:
        n, c, m = a.individuals or n, a.chroms or c, a.markers_per_chrom or m
    return n, c, m


def workload_name(workload, n, c, m):
    if workload == "fullsib":
        return (f"synthetic outbred full-sib families (3 generations, 20 offspring each), {n} individuals x {c * m} markers "
                f"({c} chromosomes x {m}) per GPU, 5% missing")
    if workload == "cfg5chrom":
        return (f"synthetic PlantImpute-shaped chromosome: biparental F2 families of 200 from 100 inbred lines, {n} individuals x "
                f"{c * m} markers ({c} chromosomes x {m}, 0.02 cM apart) per GPU, 20% missing")
    if workload == "selfing":
        return (f"SELFING = true (settings.h:14, 256 states, literal kernel): synthetic F2 lines listed at generation 4 (selfed twice), "
                f"{n} individuals x {c * m} markers ({c} chromosomes x {m}) per GPU, 20% missing")
    return f"synthetic F2 intercross, {n} individuals x {c * m} markers ({c} chromosomes x {m}) per GPU, 20% missing"


# ---------------------------------------------------------------------------------------------------------
# the reference's CPU build on a bounded sample
def run_reference_cpu(a, iterations, warm):
    """Times oracle/_ref/cnf2ref_fast (reference code, -O2 -ffast-math -fopenmp as demo.sh:6) on the first
    cpu-sample-individuals x cpu-sample-markers of the F2 workload.  Returns (rate, seconds per iteration list,
    cores, sample description) or (None, reason)."""
    from cnf2freq_b200 import synth
    binary = os.path.join(ROOT, "oracle", "_ref", "cnf2ref_fast")
    if not os.path.exists(binary):
        return None, f"{binary} not built (oracle/build_ref.sh needs /root/reference)"
    cores = os.cpu_count() or 1
    n, m = a.cpu_sample_individuals, a.cpu_sample_markers
    sim = synth.simulate_f2(n_f2=n, n_chrom=1, markers_per_chrom=m, spacing_cM=0.1, missing=0.2, seed=3)
    with tempfile.TemporaryDirectory() as tmp:
        synth.write_alpha_files(tmp + "/s", sim)
        env = dict(os.environ, OMP_NUM_THREADS=str(cores), OMP_STACKSIZE="128M")
        cmd = [binary, "--mapfile", tmp + "/s.map", "--pedfile", tmp + "/s.ped", "--genfile", tmp + "/s.gen", "--count",
               str(1 + warm + iterations), "--mode", "time"]
        res = subprocess.run(cmd, env=env, capture_output=True, text=True)
    times = [float(x) for x in re.findall(r"TIME iter \d+ seconds (\S+)", res.stdout)]
    if res.returncode != 0 or len(times) < warm + iterations:
        return None, f"reference binary failed (rc {res.returncode}): {res.stderr[-200:]}"
    all_times = times
    times = times[warm:]
    states = n * m * NUMTYPES
    sample = (f"first {n} individuals x {m} markers of chromosome 1 of the F2 workload; reference doit() = E-step + turn scoring + "
              f"M-step per iteration; iterations {warm + 1}..{warm + len(times)} timed after {warm} warm-up iterations (parameters "
              f"evolving); seconds per iteration incl. warm-up: " + " ".join("%.2f" % t for t in all_times))
    return (states / statistics.median(times), times, cores, sample), None


def reference_arm(a, rank):
    if rank != 0:
        return
    r, why = run_reference_cpu(a, a.steps, a.warmup)
    if r is None:
        print(json.dumps({"impl": "reference", "unavailable": why}))
        return
    rate, times, cores, sample = r
    n, c, m = shape_of(a, "f2")
    line = {
        "impl": "reference", "metric": METRIC, "value": rate, "unit": UNIT, "n_gpus": a.gpus, "steps": a.steps,
        "warmup": a.warmup, "ms_per_step": 1e3 * statistics.median(times), "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": workload_name("f2", n, c, m), "same_config": False,
                   "timing": "host wall clock inside the reference binary (omp_get_wtime) around doit(); ms_per_step is the sample's"},
        "cpu_baseline": {"value": rate, "unit": UNIT, "cores": cores, "kind": "reference", "sample": sample},
        "e2e": {"value": rate, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line))


# ---------------------------------------------------------------------------------------------------------
class ClockSampler:
    """nvidia-smi clocks and throttle reasons during the timed region (B200_PROFILING.md)."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.rows, self.proc, self.thread = [], None, None
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(index), f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "200"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append((time.perf_counter(), [x.strip() for x in line.split(",")]))

    def wait_ready(self, timeout=20.0):
        """nvidia-smi's start-up can stall the driver for seconds: wait for its first sample before timing."""
        t0 = time.perf_counter()
        while self.proc and not self.rows and time.perf_counter() - t0 < timeout:
            time.sleep(0.05)

    def stop(self, t_from=None, t_to=None):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        self.thread.join(timeout=2)
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        inside = [r for ts, r in self.rows if (t_from is None or ts >= t_from) and (t_to is None or ts <= t_to + 0.25)]
        for r in (inside or [r for _, r in self.rows]):
            if len(r) < 7:
                continue
            try:
                sm.append(float(r[0]))
                mx.append(float(r[1]))
            except ValueError:
                continue
            for nm, v in zip(names, r[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def traffic_from_profile():
    """dram__bytes_read.sum + dram__bytes_write.sum of warp_estep_kernel per individual-marker from the newest committed
    ncu --set full capture (profiles/r*_warp_kernel_traffic.json names the capture it came from)."""
    import glob
    for p in sorted(glob.glob(os.path.join(ROOT, "profiles", "r*_warp_kernel_traffic.json")), reverse=True):
        try:
            t = json.load(open(p))
            return t["dram_bytes_per_individual_marker"], os.path.basename(p) + ": " + str(t.get("capture", ""))
        except (OSError, ValueError, KeyError):
            continue
    return None, None


def measured_peaks():
    try:
        return json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except (OSError, ValueError):
        return None


# ---------------------------------------------------------------------------------------------------------
class Dist:
    """torch.distributed plumbing of one rank (NCCL): barrier, max over ranks, the engine's all-reduce hook."""

    def __init__(self, rank, world, local_rank):
        import torch
        self.torch, self.rank, self.world, self.local_rank = torch, rank, world, local_rank
        self.dist = None
        if world > 1:
            import torch.distributed as dist
            dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
            self.dist = dist
            # a communicator of its own for the engine's control stream (include/cnf2_engine.h, cnf2_set_allreduce): the
            # passes' one-double sums do not queue behind the statistics of the E-steps issued ahead
            self.control_group = dist.new_group(backend="nccl")

    def hook(self, eng):
        torch, dist, dev = self.torch, self.dist, self.local_rank
        control_stream, control_group = eng.control_stream(), self.control_group

        class DevBuf:  # zero-copy view of an engine buffer for torch.distributed
            def __init__(self, ptr, n):
                self.__cuda_array_interface__ = {"shape": (n,), "typestr": "<f8", "data": (ptr, False), "version": 2}

        def allreduce(ptr, count, stream):
            with torch.cuda.stream(torch.cuda.ExternalStream(stream)):
                t = torch.as_tensor(DevBuf(ptr, count), device=torch.device("cuda", dev))
                dist.all_reduce(t, op=dist.ReduceOp.SUM, group=control_group if stream == control_stream else None)
        return allreduce

    def barrier(self, eng=None):
        if eng is not None:
            eng.sync()
        self.torch.cuda.synchronize()
        if self.dist:
            self.dist.barrier()

    def max(self, x):
        if not self.dist:
            return x
        t = self.torch.tensor([x], dtype=self.torch.float64, device="cuda")
        self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX)
        return float(t.item())

    def close(self):
        if self.dist:
            self.dist.barrier()
            self.dist.destroy_process_group()


def pinned_alloc(torch):
    def pinned(shape, dtype):
        t = torch.zeros(tuple(int(x) for x in shape), dtype={np.uint8: torch.uint8, np.float64: torch.float64}[dtype], pin_memory=True)
        return t.numpy()
    return pinned


# ---------------------------------------------------------------------------------------------------------
def parity_check(D):
    """Small pedigrees through the same N-rank engine path as the timed runs -- analysed individuals sharded with
    cnf2freq_b200.shard, shared rows summed by the all-reduce hook, E-step + device M-step pass per chromosome -- against
    the CPU oracle on the whole pedigree.  Every rank checks the rows it holds; the verdict is the worst over ranks."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import oracle_lib as ol
    import randomped
    from cnf2freq_b200 import shard
    from cnf2freq_b200.engine import Engine

    def rel_err(a, b, atol):
        a, b = np.asarray(a, np.float64), np.asarray(b, np.float64)
        if not np.array_equal(np.isnan(a), np.isnan(b)):
            return np.inf
        a, b = np.nan_to_num(a), np.nan_to_num(b)
        return float((np.abs(a - b) / (atol / 1e-9 + np.abs(b))).max()) if a.size else 0.0

    worst_e, worst_m, calls_differ, cells, ties = 0.0, 0.0, 0, 0, 0
    SF, ENT = 0.013, 1.0
    for maker, kw in (("f2_cross", dict(n_f2=9, n_chrom=2, mpc=23, seed=41)), ("tangled", dict(n_chrom=2, mpc=15, seed=72))):
        pb = getattr(randomped, maker)(**kw)
        K, M, C = pb.K, pb.M, pb.C
        shards = shard.shard_dous(pb.par, pb.dous, D.world)
        shared = shard.shared_rows(pb.par, shards)
        mine = sorted(set(i for d in shards[D.rank] for i in shard.focus_tree(pb.par, d)))
        rows = np.array(mine, np.int64)
        eng = Engine(device=D.local_rank)
        if D.world > 1:
            eng.set_allreduce(D.hook(eng))
        eng.set_map(pb.markerpos, pb.chromstarts, pb.genrec)
        eng.set_shared_rows(shared)
        eng.set_pedigree(pb.par, pb.empty, pb.founder, pb.gen, pb.descendants, shards[D.rank], pb.dous)
        whole = ol.oracle_estep(pb)
        # E-step: statistics, phase-inversion scores and log-likelihoods of this rank's rows
        out = eng.estep_all(pb.alleles, pb.sure, pb.haploweight)
        sh = set(shared.tolist())
        own = np.array([i for i in mine if i not in sh], np.int64)   # rows only this rank adds to; shared ones hold the global sum
        for k in ("haplobase", "haplocount", "infprobs", "homozyg", "negshift"):
            atol = 1e-11 if k == "negshift" else 1e-13
            for sel in (own, shared.astype(np.int64)):
                if len(sel) and (len(shards[D.rank]) or sel is not own):
                    worst_e = max(worst_e, rel_err(out[k][sel], whole[k][sel], atol))
        pos = {int(d): i for i, d in enumerate(pb.dous)}
        if len(shards[D.rank]):
            worst_e = max(worst_e, rel_err(out["loglik"], whole["loglik"][[pos[int(d)] for d in shards[D.rank]]], 1e-9))
        # M-step: one pass per chromosome on the device against the oracle's M-step on the unsharded statistics
        rng = np.random.default_rng(kw["seed"])
        relhaplo = rng.uniform(0.6, 1.0, (K, M))
        hasprior = (rng.random(K) < 0.7).astype(np.uint8)
        prioral = pb.alleles.copy()
        prioral[rng.random((K, M)) < 0.2] = 0
        priorsure = rng.uniform(0.0, 0.3, (K, M, 2))
        children = rng.integers(0, 4, K).astype(np.int32)
        al, su, hw = pb.alleles.copy(), pb.sure.copy(), pb.haploweight.copy()
        for c in range(C):
            eng.set_params(c, al, su, hw)
        eng.set_mstep_inputs(relhaplo, prioral, priorsure, hasprior, children, pb.empty, pb.descendants)
        eng.begin_iteration()
        for c in range(C):
            eng.estep(c)
            eng.mstep(c, SF, ENT)
        for c in range(C):
            eng.get_params(c, al, su, hw)
        eng.close()
        if any(np.isnan(whole[k]).any() for k in ("haplobase", "infprobs")):
            continue  # an impossible unit in this draw: the E-step comparison above covers its NaN pattern
        active = np.ones(K, np.uint8)
        active[0] = 0
        hb, hc = np.zeros((K, M)), np.zeros((K, M))
        ral, rsu, rhw = pb.alleles.copy(), pb.sure.copy(), pb.haploweight.copy()
        tie = np.zeros((K, M), bool)
        for c in range(C):
            cols = slice(int(pb.chromstarts[c]), int(pb.chromstarts[c + 1]))
            hb[:, cols], hc[:, cols] = whole["haplobase"][:, cols], whole["haplocount"][:, cols]
            ip, hz = np.zeros((K, M, 2, 2)), np.zeros((K, M, 2))
            ip[:, cols], hz[:, cols] = whole["infprobs"][:, cols], whole["homozyg"][:, cols]
            # exact ties of processinfprobs (statistics equal to the current parameter, no prior): see tests/test_mstep_gpu.py
            for side in range(2):
                tot = ip[:, :, side, :].sum(axis=2)
                for k in range(2):
                    cur = np.where(ral[:, :, side] != 0, np.abs((ral[:, :, side] == k + 1).astype(float) - rsu[:, :, side]), 0.5)
                    noprior = (prioral[:, :, side] == 0) | (hasprior[:, None] == 0)
                    tie |= (ip[:, :, side, k] != 0) & noprior & (np.abs(ip[:, :, side, k] - tot * cur) <= 1e-12 * tot)
            r = ol.oracle_mstep(pb.chromstarts, active, pb.empty, children, pb.descendants, ral, rsu, rhw, relhaplo, hasprior, prioral,
                                priorsure, hb, hc, ip, hz, SF, ENT)
            ral, rsu, rhw, hb, hc = r["alleles"], r["sure"], r["haploweight"], r["haplobase"], r["haplocount"]
        if len(rows):
            ok = ~tie[rows]
            calls_differ += int((al[rows][ok] != ral[rows][ok]).sum())
            worst_m = max(worst_m, rel_err(su[rows][ok], rsu[rows][ok], 1e-12), rel_err(hw[rows][ok], rhw[rows][ok], 1e-12))
            cells += int(ok.size)
            ties += int(tie[rows].sum())
    worst_e, worst_m, calls_differ = D.max(worst_e), D.max(worst_m), int(D.max(calls_differ))
    ok_e, ok_m = worst_e <= 1e-9, worst_m <= 1e-9 and calls_differ == 0
    return {"n_ranks": D.world, "estep": "ok" if ok_e else "MISMATCH", "mstep": "ok" if ok_m else "MISMATCH",
            "max_rel": max(worst_e, worst_m), "max_rel_estep": worst_e, "max_rel_mstep": worst_m, "calls_differing": calls_differ,
            "cells_rank0": cells, "exact_ties_left_out_rank0": ties,
            "cases": "randomped.f2_cross(9 F2, 2 x 23 markers) and randomped.tangled(2 x 15 markers): statistics, negshift, loglik "
                     "(1e-9 relative) and one EM pass per chromosome (alleles exact, markersure / haploweight 1e-9) vs the CPU oracle"}


# ---------------------------------------------------------------------------------------------------------
def make_workload(a, workload, D, eng, pinned):
    """Feeds the engine map, pedigree, parameters and M-step inputs of `workload`; returns what the timed loops need."""
    from cnf2freq_b200 import synth
    n, C, mc = shape_of(a, workload)
    rank, world = D.rank, D.world
    if workload == "fullsib":
        if world > 1:
            raise SystemExit("bench.py: --workload fullsib is a single-GPU line")
        d = synth.fullsib_flat(max(1, n // 20), 20, C, mc, spacing_cM=0.02, missing=0.05, seed=4)
        K = d["par"].shape[0]
        children = np.bincount(d["par"][d["dous"]].ravel(), minlength=K).astype(np.int32)
        children[0] = 0
        src = dict(par=d["par"], empty=d["empty"], founder=d["founder"], gen=d["gen"], descendants=d["descendants"], dous=d["dous"],
                   all_dous=d["dous"], shared=d["shared"], pos=d["markerpos"], cs=d["chromstarts"], genrec=d["genrec"],
                   hasprior=np.ones(K, np.uint8), children=children)

        def params(ids, c, out):
            cols = slice(int(d["chromstarts"][c]), int(d["chromstarts"][c + 1]))
            out[0][...], out[1][...], out[2][...] = d["alleles"][ids][:, cols], d["sure"][ids][:, cols], d["haploweight"][ids][:, cols]

        def priors(ids, c, out):
            cols = slice(int(d["chromstarts"][c]), int(d["chromstarts"][c + 1]))
            out[0][...], out[1][...], out[2][...] = 0.5, d["alleles"][ids][:, cols], d["sure"][ids][:, cols]
    else:
        total, mine = (n, n // world) if a.strong else (n * world, n)
        first = mine * rank
        if a.strong and rank == world - 1:
            mine = total - first
        if workload == "cfg5chrom":
            s = synth.F2Shard(total, first, mine, C, mc, spacing_cM=0.02, missing=0.2, seed=5, lines=100, fam_size=200)
        else:
            s = synth.F2Shard(total, first, mine, C, mc, spacing_cM=0.1, missing=0.2, seed=3)
            if workload == "selfing":  # the same cross, its lines listed two selfings further down (demo.sh:18-20)
                s.gen[s.all_dous] = 4
        src = dict(par=s.par, empty=s.empty, founder=s.founder, gen=s.gen, descendants=s.descendants, dous=s.dous, all_dous=s.all_dous,
                   shared=s.shared, pos=s.pos, cs=s.cs, genrec=s.genrec, hasprior=s.hasprior, children=s.children)
        params, priors = s.params, s.priors
    eng.set_map(src["pos"], src["cs"], src["genrec"])
    eng.set_shared_rows(src["shared"])
    if world > 1:
        eng.set_allreduce(D.hook(eng))
    eng.set_pedigree(src["par"], src["empty"], src["founder"], src["gen"], src["descendants"], src["dous"], src["all_dous"])
    ids = eng.row_ids()
    R = len(ids)
    P, Q = [], []
    for c in range(C):
        blk = (pinned((R, mc, 2), np.uint8), pinned((R, mc, 2), np.float64), pinned((R, mc), np.float64))
        params(ids, c, blk)
        P.append(blk)
        q = (pinned((R, mc), np.float64), pinned((R, mc, 2), np.uint8), pinned((R, mc, 2), np.float64))
        priors(ids, c, q)
        Q.append(q)
    return dict(src=src, ids=ids, R=R, C=C, mc=mc, n=len(src["dous"]), P=P, Q=Q, name=workload_name(workload, n, C, mc))


def run_workload(a, workload, D, steps, warmup, with_clocks):
    """Times `steps` EM iterations (after `warmup`) of `workload` on this rank's engine; returns the numbers of one
    bench line (rank 0; other ranks return None)."""
    import torch
    from cnf2freq_b200.engine import Engine
    pinned = pinned_alloc(torch)
    rank, world = D.rank, D.world
    n_chrom = shape_of(a, workload)[1]
    # three E-steps in flight (each with its own scratch and statistics block) once there are chromosomes to overlap
    eng = Engine(device=D.local_rank, async_params=True, in_flight=3 if n_chrom >= 3 else 1, selfing=workload == "selfing")
    w = make_workload(a, workload, D, eng, pinned)
    C, mc, R, P, Q, src = w["C"], w["mc"], w["R"], w["P"], w["Q"], w["src"]
    M = C * mc
    AHEAD = 2 if C >= 3 else 0  # E-steps issued ahead of the block behind a chromosome (3 slots in flight)

    def upload(Pset):
        for c in range(C):
            eng.set_params_packed(c, *Pset[c])
        eng.set_mstep_rows(src["hasprior"], src["children"], src["empty"], src["descendants"])
        for c in range(C):
            eng.set_mstep_chrom_packed(c, *Q[c])

    def em_iteration(readback=None):
        """doit(): per chromosome the E-step and the block behind it; E-steps of the next chromosomes run beside the block
        (they read their own chromosome only; the block writes chromosomes <= c).  readback: pinned (alleles, sure, haploweight)
        per chromosome for the state dump -- a chromosome's alleles and markersure are final once its own pass is through and
        are taken then, beside the remaining chromosomes' kernels; the haploweights (later passes and inversions still move
        them) at the end."""
        eng.begin_iteration()
        n_inv = 0
        for c in range(min(AHEAD, C)):
            eng.estep(c)
        for c in range(C):
            if c + AHEAD < C or AHEAD == 0:
                eng.estep(c + AHEAD)
            n_inv += len(eng.em_pass(c))
            if readback is not None:
                eng.get_params_packed_async(c, readback[c][0], readback[c][1], None)
        if readback is not None:
            for c in range(C):
                eng.get_params_packed_async(c, None, None, readback[c][2])
            eng.sync()
        return n_inv

    upload(P)
    eng.set_em_state(0.013, 1.0, 0, 0)
    inv_log, it_ms = [], []
    for _ in range(warmup):
        inv_log.append(em_iteration())
    D.barrier(eng)
    # the evolved state, read back once: the e2e region below starts from it
    Pev = [(pinned((R, mc, 2), np.uint8), pinned((R, mc, 2), np.float64), pinned((R, mc), np.float64)) for _ in range(C)]
    back = [(pinned((R, mc, 2), np.uint8), pinned((R, mc, 2), np.float64), pinned((R, mc), np.float64)) for _ in range(C)]
    for c in range(C):
        eng.get_params_packed(c, *Pev[c])
    state = eng.get_em_state()
    ind_markers = w["n"] * M
    # the whole job: every rank's analysed individuals; 256 states under SELFING (settings.h:27)
    states_per_step = len(src["all_dous"]) * M * (4 * NUMTYPES if workload == "selfing" else NUMTYPES)

    # ---- e2e: host buffers; upload of everything once + per iteration the read-back of every chromosome's parameters ----
    D.barrier(eng)
    t0 = time.perf_counter()
    upload(Pev)
    eng.set_em_state(state[0], state[1], state[2], state[3])
    for _ in range(steps):
        em_iteration(readback=back)
    D.barrier(eng)
    e2e_s = D.max(time.perf_counter() - t0) / steps
    h2d = (R * M * (26 + 26) + 2 * M * 8) / steps          # parameters + (relhaplo, priors), once, amortised
    d2h = R * M * 26 + 2 * R * C * 12                       # parameters back; the phase-inversion minima of every pass

    # ---- value: the same state again, parameters resident, device time ----
    upload(Pev)
    eng.set_em_state(state[0], state[1], state[2], state[3])
    eng.mstep_counters(reset=True)
    D.barrier(eng)
    eng.estep_times(reset=True)
    sampler = ClockSampler(D.local_rank) if (rank == 0 and with_clocks) else None
    if sampler:
        sampler.wait_ready()
    D.barrier(eng)
    t_from = time.perf_counter()
    eng.mark(0)
    for _ in range(steps):
        eng.mark(2)
        inv_log.append(em_iteration())
        eng.mark(3)
        it_ms.append(eng.elapsed_ms(2, 3))
    eng.mark(1)
    D.barrier(eng)
    step_ms = D.max(eng.elapsed_ms(0, 1)) / steps
    clocks = sampler.stop(t_from, time.perf_counter()) if sampler else None
    live_ms, live_n, n_kernels = eng.estep_times(reset=True)
    evals = eng.mstep_counters(reset=True)
    launches_per_estep = eng.last_estep_launches()
    sf_end = eng.get_em_state()

    # ---- E-step alone (sub-key `estep`): launches back to back, then one launch at a time ----
    for _ in range(2):
        for c in range(C):
            eng.estep(c)
    D.barrier(eng)
    eng.mark(0)
    for _ in range(steps):
        for c in range(C):
            eng.estep(c)
    eng.mark(1)
    D.barrier(eng)
    estep_ms = D.max(eng.elapsed_ms(0, 1)) / steps
    iso = []
    for c in range(C):
        eng.estep(c)
        iso.append(eng.last_estep_ms())
    D.barrier(eng)

    res = None
    if rank == 0:
        per_launch_im = w["n"] * mc
        launch_ms = estep_ms / C
        res = dict(name=w["name"], states_per_step=states_per_step, step_ms=step_ms, it_ms=it_ms, e2e_s=e2e_s, h2d=h2d, d2h=d2h,
                   clocks=clocks, estep_ms=estep_ms, launch_ms=launch_ms, launch_ms_isolated=statistics.mean(iso),
                   launch_ms_live=live_ms / max(1, live_n), per_launch_im=per_launch_im, evals=evals, inversions=inv_log,
                   scalefactor_end=sf_end[0], hitnnn_last=sf_end[4], launches_per_estep=launches_per_estep, rows=R, C=C, mc=mc,
                   gpu_launches=n_kernels)
        res["peaks"] = eng.measure_peaks()
    eng.close()
    return res


def roofline_of(res, hbm_peak, peaks_present):
    dfma_tflops, copy_gbs = res["peaks"]
    im = res["per_launch_im"]

    def frac(ms):
        return im * FLOP_PER_INDMARKER / (ms * 1e-3) * 1e-12 / dfma_tflops

    fl = im * FLOP_PER_INDMARKER / (res["launch_ms"] * 1e-3) * 1e-12
    by = im * BYTES_PER_INDMARKER / (res["launch_ms"] * 1e-3) * 1e-9
    per_im, cap = traffic_from_profile()
    # M-step kernels: gradient evaluations counted on the device over the timed iterations
    ev0, ev1, n_conf, n_rej = res["evals"]
    steps = len(res["it_ms"])
    ms_fl = (ev0 * FLOP_PER_GRADIENT[0] + ev1 * FLOP_PER_GRADIENT[1]) / max(1, steps)
    t_left = max(1e-9, res["step_ms"] - res["estep_ms"])  # what an iteration costs beyond its E-steps run back to back
    return {
        "bound": "fp64", "kernel": "cnf2::warp_estep_kernel (+ the literal cnf2::estep_kernel for the units it hands over; none in this workload)",
        "achieved": fl, "peak": dfma_tflops, "unit": "TFLOP/s", "frac": fl / dfma_tflops,
        "traffic": per_im * im if per_im else None, "traffic_source": cap,
        "algorithmic": {"flop_per_individual_marker": FLOP_PER_INDMARKER, "bytes_per_individual_marker": BYTES_PER_INDMARKER,
                        "individual_markers_per_launch": im},
        "launch_ms": res["launch_ms"], "launch_ms_isolated": res["launch_ms_isolated"], "launch_ms_in_em_iteration": res["launch_ms_live"],
        "frac_isolated": frac(res["launch_ms_isolated"]),
        "launch_ms_note": "launch_ms = E-steps back to back / launches (3 launches in flight, the tail of one overlaps the next); isolated = "
                          "one launch alone, CUDA events around it; in_em_iteration = mean event duration of the launches inside the timed EM "
                          "iterations, where up to 3 launches and the M-step kernels share the GPU",
        "peak_source": "fp64: register-resident DFMA kernel measured in this run (MEASURED_PEAKS.json has no fp64 figure); hbm: "
                       "MEASURED_PEAKS.json" + ("" if peaks_present else " absent -> fallback 6650 GB/s"),
        "hbm": {"achieved": by, "peak": hbm_peak, "unit": "GB/s", "frac": by / hbm_peak, "copy_gbs_this_run": copy_gbs},
        "fp64": {"achieved": fl, "peak": dfma_tflops, "unit": "TFLOP/s", "frac": fl / dfma_tflops},
        "mstep": {"kernel": "cnf2::mstep_infprobs_kernel + cnf2::mstep_haploweight_kernel (cappedgd solves)", "bound": "fp64",
                  "gradient_evaluations_per_iteration": [ev0 / max(1, steps), ev1 / max(1, steps)],
                  "flop_per_gradient": list(FLOP_PER_GRADIENT), "flop_per_iteration": ms_fl,
                  "solves_on_integral_free_steps_per_iteration": {"confirmed": n_conf / max(1, steps), "sent_back": n_rej / max(1, steps)},
                  "ms_per_iteration_beyond_esteps": t_left, "achieved": ms_fl / (t_left * 1e-3) * 1e-12, "peak": dfma_tflops,
                  "unit": "TFLOP/s", "frac": ms_fl / (t_left * 1e-3) * 1e-12 / dfma_tflops,
                  "note": "time = EM iteration - its E-steps run back to back (the passes overlap the next chromosomes' E-steps, so this is "
                          "the passes' exposed time; their ncu launch list is under profiles/)"},
    }


def ours(a, D):
    import torch
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the engine has no CPU path (use --impl reference for the CPU arm)")
    torch.cuda.set_device(D.local_rank)
    rank, world = D.rank, D.world
    parity = None
    if not a.no_parity_check:
        parity = parity_check(D)
        if parity["estep"] != "ok" or parity["mstep"] != "ok":
            if rank == 0:
                print(json.dumps({"parity_check": parity}))
            raise SystemExit("bench.py: parity check against the oracle failed: " + json.dumps(parity))
    res = run_workload(a, a.workload, D, a.steps, a.warmup, True)
    extra = {}
    if world == 1 and not a.no_extra_configs and not a.explicit_shape and a.workload == "f2":
        for wl in ("fullsib", "cfg5chrom", "selfing"):
            extra[wl] = run_workload(a, wl, D, max(2, min(3, a.steps)), 2, False)
    if rank == 0:
        peaks = measured_peaks()
        hbm_peak = peaks["hbm_gbs"] if peaks else 6650.0
        roof = roofline_of(res, hbm_peak, bool(peaks))
        cpu = None
        if not a.no_cpu_baseline:
            r, why = run_reference_cpu(a, a.steps, a.warmup)
            cpu = ({"value": r[0], "unit": UNIT, "cores": r[2], "kind": "reference", "sample": r[3]} if r else
                   {"value": None, "unit": UNIT, "cores": os.cpu_count(), "kind": "reference", "sample": "unavailable: " + why})
        S = res["states_per_step"]
        configs = {}
        for wl, x in extra.items():
            rf = roofline_of(x, hbm_peak, bool(peaks))
            configs[wl] = {"workload": x["name"], "value": x["states_per_step"] / (x["step_ms"] * 1e-3), "unit": UNIT,
                           "ms_per_step": x["step_ms"], "ms_per_iteration": x["it_ms"], "estep": {"value": x["states_per_step"] / (x["estep_ms"] * 1e-3),
                                                                                              "ms_per_step": x["estep_ms"]},
                           "e2e": {"value": x["states_per_step"] / x["e2e_s"], "ms_per_step": x["e2e_s"] * 1e3},
                           "frac": rf["frac"], "frac_isolated": rf["frac_isolated"], "launch_ms": rf["launch_ms"],
                           "launch_ms_isolated": rf["launch_ms_isolated"], "mstep_frac": rf["mstep"]["frac"], "inversions": x["inversions"]}
            if wl == "selfing":  # every unit through the literal kernel estep_kernel<true>: no roofline model is claimed for it
                configs[wl].update(frac=None, frac_isolated=None, kernel="cnf2::estep_kernel<true> (literal, one CTA per unit, 256 states)")
        line = {
            "metric": METRIC, "value": S / (res["step_ms"] * 1e-3), "unit": UNIT, "n_gpus": world, "steps": a.steps,
            "warmup": a.warmup, "ms_per_step": res["step_ms"], "higher_is_better": True, "scaling": "strong" if a.strong else "weak", "vs_baseline": None,
            "dtype": "f64", "data": "synthetic",
            "config": {"workload": res["name"], "parallelism": f"analysed individuals sharded over {world} GPU(s); founders' rows all-reduced (NCCL)",
                       "l2": "parameters + statistics of one chromosome (%.0f MB) exceed the 126 MB L2" % (res["rows"] * res["mc"] * (26 + 72) / 1e6),
                       "step": "one EM iteration = doit(): per chromosome E-step (sweeps for 8 shift modes, posterior scatter, phase-inversion "
                               "scores) + the block behind it (inversion candidates, negshifter, processinfprobs, updatehaploweights, step-size "
                               "control); parameters stay on the device and evolve; iterations %d..%d timed" % (a.warmup + 1, a.warmup + a.steps)},
            "ms_per_iteration": res["it_ms"], "inversions_per_iteration": res["inversions"], "scalefactor_end": res["scalefactor_end"],
            "e2e": {"value": S / res["e2e_s"], "unit": UNIT, "h2d_bytes_per_step": res["h2d"], "d2h_bytes_per_step": res["d2h"],
                    "ms_per_step": res["e2e_s"] * 1e3,
                    "note": "timed region: upload of every chromosome's parameters and M-step inputs from pinned host memory (once), K EM "
                            "iterations, each with the read-back of every chromosome's parameters into pinned host memory (alleles and markersure "
                            "of a chromosome behind its own pass, beside the later chromosomes' kernels; haploweights at the iteration's end; "
                            "cnf2_get_params_packed_async + cnf2_sync)"},
            "estep": {"value": S / (res["estep_ms"] * 1e-3), "unit": UNIT, "ms_per_step": res["estep_ms"],
                      "note": "the E-step of all chromosomes alone (no M-step), launches back to back"},
            "gpu_launches": res["gpu_launches"], "parity_check": parity, "roofline": roof, "configs": configs, "cpu_baseline": cpu,
            "clocks": res["clocks"],
        }
        print(json.dumps(line))


def main():
    a = parse()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if a.impl == "reference":
        reference_arm(a, rank)
        return
    if world != a.gpus:
        raise SystemExit(f"bench.py: --gpus {a.gpus} needs {a.gpus} ranks (launch with torch.distributed.run); WORLD_SIZE={world}")
    import torch
    if torch.cuda.is_available():
        torch.cuda.set_device(local_rank)
    D = Dist(rank, world, local_rank)
    try:
        ours(a, D)
    finally:
        D.close()


if __name__ == "__main__":
    main()
