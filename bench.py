#!/usr/bin/env python
"""Benchmark of the cnF2freq hot path (forward-backward + posterior statistics, the E-step of one EM
iteration) on B200: individual-marker-states per second.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
    torchrun --nproc-per-node N bench.py --gpus N ...          (one rank per GPU, NCCL)

A "step" is one E-step pass over every chromosome of the workload.  Workload (BASELINE.json configs[2], the
largest configuration that one GPU finishes in seconds at the current kernel speed): synthetic F2 intercross,
2,000 analysed individuals x 10,000 markers (10 chromosomes x 1,000), 20 % missing genotypes, PER GPU (weak
scaling: every rank holds its own 2,000 F2s of the same cross; the two founders' statistics rows are
all-reduced over NCCL once per chromosome).

value      inputs resident in HBM; CUDA events on the engine's stream around the K steps; max over ranks.
e2e        through the C ABI with pinned HOST buffers: every step copies the parameters host->device
           (cnf2_set_params_packed) and the statistics device->host (cnf2_get_stats_packed), row-ordered
           contiguous blocks per chromosome; three E-steps in flight.
roofline   fp64-pipe and HBM bounds from the algorithmic work of SURVEY.md 8(d): 67,584 flop and
           694 B per individual-marker; peaks: fp64 measured in this run (register-resident DFMA kernel,
           MEASURED_PEAKS.json has no fp64 figure), HBM from MEASURED_PEAKS.json.
cpu_baseline / --impl reference: the reference's own code (oracle/_ref/cnf2ref_fast, built from
           /root/reference by oracle/build_ref.sh with demo.sh's flags) on all host cores, on a bounded sample
           of the same workload; it times the reference's whole doit() (E-step + M-step).
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
FLOP_PER_INDMARKER = 8448 * 8      # 8,448 flop per (individual-marker, shift mode), 8 shift modes
BYTES_PER_INDMARKER = 238 + 8 + 448  # 34 B x 7 relatives read, 8 B interval, 64 B x 7 accumulated
METRIC = "individual-marker-states/sec fwd-bwd+EM"
UNIT = "individual-marker-states/s"


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="f2", choices=["f2", "fullsib"],
                    help="f2: BASELINE.json configs[2] (the bench workload); fullsib: configs[3]'s pedigree shape (families of 20 offspring)")
    ap.add_argument("--individuals", type=int, default=2000, help="analysed individuals per GPU")
    ap.add_argument("--chroms", type=int, default=10)
    ap.add_argument("--markers-per-chrom", type=int, default=1000)
    ap.add_argument("--cpu-sample-individuals", type=int, default=64)
    ap.add_argument("--cpu-sample-markers", type=int, default=500)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-mstep", action="store_true", help="skip the device M-step timing (em_iteration)")
    return ap.parse_args()


def workload_name(a):
    if a.workload == "fullsib":
        return (f"synthetic outbred full-sib families (3 generations, 20 offspring each), {a.individuals} individuals x "
                f"{a.chroms * a.markers_per_chrom} markers ({a.chroms} chromosomes x {a.markers_per_chrom}) per GPU, 5% missing")
    return (f"synthetic F2 intercross, {a.individuals} individuals x {a.chroms * a.markers_per_chrom} markers "
            f"({a.chroms} chromosomes x {a.markers_per_chrom}) per GPU, 20% missing")


# ---------------------------------------------------------------------------------------------------------
# the reference's CPU build on a bounded sample
def run_reference_cpu(a, iterations, warm):
    """Times oracle/_ref/cnf2ref_fast (reference code, -O2 -ffast-math -fopenmp as demo.sh:6) on the first
    cpu-sample-individuals x cpu-sample-markers of the workload.  Returns (rate, seconds per iteration list,
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
    times = times[warm:]
    states = n * m * NUMTYPES
    sample = (f"first {n} individuals x {m} markers of chromosome 1 of the workload; reference doit() = E-step + "
              f"M-step per iteration; {len(times)} iterations after {warm} warm-up")
    return (states / statistics.median(times), times, cores, sample), None


def reference_arm(a, rank):
    if rank != 0:
        return
    r, why = run_reference_cpu(a, a.steps, a.warmup)
    if r is None:
        print(json.dumps({"impl": "reference", "unavailable": why}))
        return
    rate, times, cores, sample = r
    line = {
        "impl": "reference", "metric": METRIC, "value": rate, "unit": UNIT, "n_gpus": a.gpus, "steps": a.steps,
        "warmup": a.warmup, "ms_per_step": 1e3 * statistics.median(times), "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": workload_name(a), "timing": "host wall clock inside the reference binary (omp_get_wtime)"},
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


def traffic_from_profile(per_launch_im):
    """dram__bytes_read.sum + dram__bytes_write.sum of warp_estep_kernel from the committed ncu --set full capture
    (profiles/r1_warp_kernel_traffic.json: bytes per individual-marker of that capture), scaled to this launch."""
    try:
        t = json.load(open(os.path.join(ROOT, "profiles", "r1_warp_kernel_traffic.json")))
        return t["dram_bytes_per_individual_marker"] * per_launch_im
    except (OSError, ValueError, KeyError):
        return None


def measured_peaks():
    try:
        return json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except (OSError, ValueError):
        return None


def ours(a, rank, world, local_rank):
    import torch
    import torch.distributed as dist
    from cnf2freq_b200 import synth
    from cnf2freq_b200.engine import Engine

    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the engine has no CPU path (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))

    def pinned(shape, dtype):
        t = torch.zeros(tuple(shape), dtype={np.uint8: torch.uint8, np.float64: torch.float64}[dtype], pin_memory=True)
        return t.numpy()

    if a.workload == "fullsib":
        d = synth.fullsib_flat(max(1, a.individuals // 20), 20, a.chroms, a.markers_per_chrom, spacing_cM=0.02, missing=0.05,
                               seed=4 + rank, pinned=pinned)
    else:
        d = synth.f2_flat(a.individuals, a.chroms, a.markers_per_chrom, spacing_cM=0.1, missing=0.2, seed=3 + rank, pinned=pinned)
    K, M, C = d["par"].shape[0], d["markerpos"].shape[0], a.chroms
    # three E-steps in flight; the page-locked parameter arrays are not touched while a step runs
    eng = Engine(device=local_rank, async_params=True, in_flight=3)
    eng.set_map(d["markerpos"], d["chromstarts"], d["genrec"])
    eng.set_shared_rows(d["shared"])
    if world > 1:
        class DevBuf:  # zero-copy view of the engine's shared statistics block for torch.distributed
            def __init__(self, ptr, n):
                self.__cuda_array_interface__ = {"shape": (n,), "typestr": "<f8", "data": (ptr, False), "version": 2}

        def allreduce(ptr, count, stream):
            with torch.cuda.stream(torch.cuda.ExternalStream(stream)):
                t = torch.as_tensor(DevBuf(ptr, count), device=torch.device("cuda", local_rank))
                dist.all_reduce(t, op=dist.ReduceOp.SUM)
        eng.set_allreduce(allreduce)
    eng.set_pedigree(d["par"], d["empty"], d["founder"], d["gen"], d["descendants"], d["dous"])
    # statistics come back in the engine's row order, one contiguous block per quantity and chromosome
    # (cnf2_get_stats_packed): unstrided device -> host copies into page-locked memory; two buffers, alternating
    R, mc, LAG = len(eng.row_ids()), a.markers_per_chrom, 2
    out = [dict(haplobase=pinned((R, mc), np.float64), haplocount=pinned((R, mc), np.float64),
                infprobs=pinned((R, mc, 2, 2), np.float64), homozyg=pinned((R, mc, 2), np.float64),
                loglik=pinned((len(d["dous"]), 2), np.float64)) for _ in range(3)]

    def barrier():
        eng.sync()
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()

    def maxranks(x):
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    # parameters per chromosome in the engine's row order (cnf2_set_params_packed): unstrided host -> device copies
    ids = eng.row_ids()
    P = []
    for c in range(C):
        cols = slice(int(d["chromstarts"][c]), int(d["chromstarts"][c + 1]))
        blk = (pinned((R, mc, 2), np.uint8), pinned((R, mc, 2), np.float64), pinned((R, mc), np.float64))
        blk[0][...] = d["alleles"][ids][:, cols]
        blk[1][...] = d["sure"][ids][:, cols]
        blk[2][...] = d["haploweight"][ids][:, cols]
        P.append(blk)

    def e2e_step():
        # E-steps in flight: chromosome c is handed over and started before chromosome c - LAG's statistics are read
        for c in range(C + LAG):
            if c < C:
                eng.set_params_packed(c, *P[c])
                eng.estep(c)
            if c >= LAG:
                eng.get_stats_packed(c - LAG, **out[(c - LAG) % 3])

    def e2e_trace():
        # CNF2_DEBUG: host time spent inside each call of one e2e step (ms per chromosome)
        tt = {"set_params": [], "estep": [], "get_stats": []}
        for c in range(C + LAG):
            t = time.perf_counter()
            if c < C:
                eng.set_params_packed(c, *P[c])
            t1 = time.perf_counter()
            if c < C:
                eng.estep(c)
            t2 = time.perf_counter()
            if c >= LAG:
                eng.get_stats_packed(c - LAG, **out[(c - LAG) % 3])
            t3 = time.perf_counter()
            tt["set_params"].append(1e3 * (t1 - t)), tt["estep"].append(1e3 * (t2 - t1)), tt["get_stats"].append(1e3 * (t3 - t2))
        print("cnf2: e2e host ms per call: " + "; ".join(k + " " + " ".join("%.1f" % x for x in v) for k, v in tt.items()), file=sys.stderr)

    def resident_step():
        for c in range(C):
            eng.estep(c)

    # warm-up: the e2e steps also leave every chromosome's parameters resident for the device-timed region
    for _ in range(max(1, a.warmup)):
        e2e_step()
    barrier()
    if os.environ.get("CNF2_DEBUG") and rank == 0:
        e2e_trace()
        barrier()
    ind_markers = len(d["dous"]) * M
    states_per_step = ind_markers * NUMTYPES * world

    # ---- e2e: host buffers, copies inside the timed region ----
    t0 = time.perf_counter()
    for _ in range(a.steps):
        e2e_step()
    barrier()
    e2e_s = maxranks(time.perf_counter() - t0) / a.steps
    pars = d["par"][d["dous"]].ravel()
    n_rows = int((np.unique(np.concatenate([d["dous"], pars, d["par"][pars].ravel(), d["shared"]])) > 0).sum())
    h2d = n_rows * M * 26 + 2 * M * 8
    d2h = n_rows * M * 64 + len(d["dous"]) * C * 16

    # ---- value: inputs resident in HBM, CUDA events on the launching stream ----
    sampler = ClockSampler(local_rank) if rank == 0 else None
    for _ in range(a.warmup):
        resident_step()
    if sampler:
        sampler.wait_ready()
    barrier()
    t_from = time.perf_counter()
    eng.mark(0)
    kernel_ms = []
    for _ in range(a.steps):
        resident_step()
    eng.mark(1)
    barrier()
    step_ms = maxranks(eng.elapsed_ms(0, 1)) / a.steps
    clocks = sampler.stop(t_from, time.perf_counter()) if sampler else None
    # per-launch duration of the dominant kernel (CUDA events around each launch, engine stream)
    for c in range(C):
        eng.estep(c)
        kernel_ms.append(eng.last_estep_ms())
    barrier()
    launch_ms_isolated = statistics.mean(kernel_ms)
    # launches of consecutive chromosomes overlap (two in flight): the average duration over the timed region
    launch_ms = step_ms / C
    launches_per_estep = eng.last_estep_launches()

    # ---- the M-step on the device (row a6), timed after everything else because it updates the parameters in place:
    # one EM iteration = per chromosome E-step + the reference's M-step pass (cnF2freq.cpp:6356-6367)
    em = None
    if not a.no_mstep:
        K_ind = d["par"].shape[0]
        rel = np.full((K_ind, M), 0.9)                         # relhaplo: the reference's start value is a constant too
        children = np.bincount(d["par"].ravel(), minlength=K_ind).astype(np.int32)
        children[0] = 0
        eng.set_mstep_inputs(rel, d["alleles"], d["sure"], np.ones(K_ind, np.uint8), children, d["empty"], d["descendants"])
        barrier()
        hits = 0
        t0 = time.perf_counter()
        eng.begin_iteration()
        AHEAD = 2                                              # E-steps issued ahead of the pass (engine in_flight = 3 slots)
        for c in range(min(AHEAD, C)):
            eng.estep(c)
        for c in range(C):
            if c + AHEAD < C:
                eng.estep(c + AHEAD)                           # reads its own chromosome only; the pass below writes chromosomes <= c
            hits += eng.mstep(c, 0.013, 1.0)                   # blocks until the pass is done
        barrier()
        em_s = maxranks(time.perf_counter() - t0)
        em = {"value": states_per_step / em_s, "unit": UNIT, "ms_per_iteration": em_s * 1e3,
              "hitnnn": hits, "note": "one EM iteration on the device: per chromosome the E-step and the reference's M-step pass "
              "(processinfprobs + updatehaploweights, bisection with 15-point Gauss-Legendre per cell), the next two chromosomes' E-steps "
              "running beside it; the M-step of chromosome c "
              "also re-runs updatehaploweights on chromosomes < c as the reference does"}

    if rank == 0:
        dfma_tflops, copy_gbs = eng.measure_peaks()
        peaks = measured_peaks()
        hbm_peak = peaks["hbm_gbs"] if peaks else 6650.0
        per_launch_im = len(d["dous"]) * a.markers_per_chrom
        fl = per_launch_im * FLOP_PER_INDMARKER / (launch_ms * 1e-3) * 1e-12
        by = per_launch_im * BYTES_PER_INDMARKER / (launch_ms * 1e-3) * 1e-9
        t_fp64 = per_launch_im * FLOP_PER_INDMARKER / (dfma_tflops * 1e12)
        t_hbm = per_launch_im * BYTES_PER_INDMARKER / (hbm_peak * 1e9)
        roof = {"bound": "fp64" if t_fp64 >= t_hbm else "hbm", "kernel": "cnf2::warp_estep_kernel (+ the literal cnf2::estep_kernel for the units it hands over; none in this workload)",
                "achieved": fl if t_fp64 >= t_hbm else by, "peak": dfma_tflops if t_fp64 >= t_hbm else hbm_peak,
                "unit": "TFLOP/s" if t_fp64 >= t_hbm else "GB/s",
                "frac": (fl / dfma_tflops) if t_fp64 >= t_hbm else by / hbm_peak, "traffic": traffic_from_profile(per_launch_im),
                "launch_ms": launch_ms, "launch_ms_isolated": launch_ms_isolated,
                "launch_ms_note": "launch_ms = timed region / launches (launches in flight, the tail of one overlaps the next); isolated = one launch alone, CUDA events around it",
                "peak_source": "fp64: register-resident DFMA kernel measured in this run (MEASURED_PEAKS.json has no fp64 "
                               "figure); hbm: MEASURED_PEAKS.json" + ("" if peaks else " absent -> fallback 6650 GB/s"),
                "hbm": {"achieved": by, "peak": hbm_peak, "unit": "GB/s", "frac": by / hbm_peak, "copy_gbs_this_run": copy_gbs},
                "fp64": {"achieved": fl, "peak": dfma_tflops, "unit": "TFLOP/s", "frac": fl / dfma_tflops}}
        cpu = None
        if not a.no_cpu_baseline:
            r, why = run_reference_cpu(a, 2, 0)
            cpu = ({"value": r[0], "unit": UNIT, "cores": r[2], "kind": "reference", "sample": r[3]} if r else
                   {"value": None, "unit": UNIT, "cores": os.cpu_count(), "kind": "reference", "sample": "unavailable: " + why})
        line = {
            "metric": METRIC, "value": states_per_step / (step_ms * 1e-3), "unit": UNIT, "n_gpus": world, "steps": a.steps,
            "warmup": a.warmup, "ms_per_step": step_ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f64", "data": "synthetic",
            "config": {"workload": workload_name(a), "parallelism": f"individuals sharded over {world} GPU(s)",
                       "l2": "inputs + statistics per launch (%.0f MB) exceed the 126 MB L2" % ((h2d + d2h) / C / 1e6),
                       "step": "E-step of all chromosomes: fwd/bwd sweeps for 8 shift modes + posterior scatter into EM statistics"},
            "e2e": {"value": states_per_step / e2e_s, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                    "ms_per_step": e2e_s * 1e3},
            "gpu_launches": a.steps * C * launches_per_estep, "em_iteration": em, "roofline": roof, "cpu_baseline": cpu, "clocks": clocks,
        }
        print(json.dumps(line))
    eng.close()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


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
    ours(a, rank, world, local_rank)


if __name__ == "__main__":
    main()
