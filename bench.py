#!/usr/bin/env python
"""bench.py -- BSW hot-path benchmark (BASELINE.json config[1]).

Workload ("config.workload"): the BSW-only micro-bench of BASELINE.json -- ~10M synthetic
SeqPairs from 150 bp reads (one seed per read, left + right extension), band w=100, routed by
the caller's tier rule to getScores8 / getScores16 ("8-bit with 16-bit overflow re-run"),
right-extension h0 chained from the left score.  Generator: oracle/workload.c.

One "step" = one pass of the hot path over the whole batch: four boundary calls
(left tier-8, left tier-16, right tier-8, right tier-16).

  value  : GCUPS with the inputs already resident in HBM (bsw_dev_getScores), CUDA events.
  e2e    : GCUPS through the reference-facing C-ABI calls bsw_getScores8/16 on pinned HOST
           buffers -- H2D of SeqPairs + sequences and D2H of results inside the timed region.
  cells  : DP cells as the oracle counts them (sum over rows of end-beg); the library's
           counter is cross-checked against the oracle on the CPU-baseline sample.
  roofline: packed-integer/DPX issue rate: achieved = algorithmic packed ops
           (8 per cell) / DP-kernel time;  peak = measured VIMNMX3/VIADDMNMX.S16x2 issue rate x 2.

Multi-GPU (torchrun, one rank per GPU): every rank owns an independent workload of the same
size (weak scaling); no collective on the data path, only the barrier/max-reduce for timing.

`--impl reference` times the CPU implementation of the path (the oracle port: the reference's
own sources are not in /root/reference) with all host threads on a bounded sample.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
for p in (os.path.join(ROOT, "bwa-mem2-arm_b200"), os.path.join(ROOT, "oracle")):
    if p not in sys.path:
        sys.path.insert(0, p)

import numpy as np

OPS_PER_CELL = 8           # algorithmic packed-integer ops per BSW cell
W_BAND = 100


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--streams", type=int, default=int(os.environ.get("BENCH_STREAMS", "4")),
                    help="caller streams the calls of a resident step are issued on (1 = strictly serial calls)")
    ap.add_argument("--reads", type=int, default=int(os.environ.get("BENCH_READS", "5060000")),
                    help="150 bp reads per GPU (5.06M reads -> ~10M SeqPairs)")
    ap.add_argument("--cpu-sample", type=int, default=int(os.environ.get("BENCH_CPU_SAMPLE", "400000")))
    return ap.parse_args()


def rank_info():
    return int(os.environ.get("RANK", "0")), int(os.environ.get("LOCAL_RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))


def host_threads():
    try:
        return len(os.sched_getaffinity(0))
    except AttributeError:
        return os.cpu_count() or 1


class ClockSampler(threading.Thread):
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md): one
    streaming `nvidia-smi -lms 50` process, read line by line."""

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.stop_flag, self.rows, self.proc = index, False, [], None

    def run(self):
        q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
             "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
             "clocks_event_reasons.sw_power_cap")
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={q}", "--format=csv,noheader,nounits",
                                          "-i", str(self.index), "-lms", "50"], stdout=subprocess.PIPE, text=True)
            for line in self.proc.stdout:
                if self.stop_flag:
                    break
                self.rows.append([x.strip() for x in line.strip().split(",")])
        except Exception:
            pass

    def stop(self):
        self.stop_flag = True
        if self.proc is not None:
            try:
                self.proc.terminate()
            except Exception:
                pass
        self.join(timeout=2)

    def summary(self):
        sm, mx, reasons = [], 0, set()
        for r in self.rows:
            try:
                sm.append(float(r[0])); mx = max(mx, float(r[1]))
            except (ValueError, IndexError):
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": mx or None,
                "reasons": sorted(reasons), "samples": len(sm)}


def build_workload(n_reads, seed, nthreads):
    """Returns per side the tier-8 and tier-16 call inputs (numpy, pageable) + the pair count."""
    import ora
    d = ora.gen_bsw(n_reads, read_len=150, seed=seed, div=0.01, w=W_BAND, nthreads=nthreads)
    return d


def split_tiers(pairs):
    import ora
    t = ora.classify(pairs)
    return np.ascontiguousarray(pairs[t == 0]), np.ascontiguousarray(pairs[t != 0])


def pinned_copy(a):
    import torch
    t = torch.empty(a.nbytes, dtype=torch.uint8, pin_memory=True)
    out = t.numpy().view(a.dtype).reshape(a.shape)
    out[...] = a
    return out, t            # keep t alive


def run_reference(args):
    """CPU implementation of the path (oracle port) on all host threads, bounded sample per step."""
    import ora
    rank, _, world = rank_info()
    if rank != 0:
        return
    cores = host_threads()
    n_reads = max(1000, args.cpu_sample // 2)
    d = build_workload(n_reads, 0xC2, cores)
    prm = ora.make_params()
    L, lr, lq = d["left"]; R, rr, rq = d["right"]
    Lo = L.copy(); ora.bsw_batch(Lo, lr, lq, prm, W_BAND, cores); ora.chain_h0(Lo, R)
    npairs = len(L) + len(R)
    times, cells = [], 0
    for it in range(args.warmup + args.steps):
        a, b = L.copy(), R.copy()
        t0 = time.perf_counter()
        c = ora.bsw_batch(a, lr, lq, prm, W_BAND, cores) + ora.bsw_batch(b, rr, rq, prm, W_BAND, cores)
        dt = time.perf_counter() - t0
        if it >= args.warmup:
            times.append(dt); cells = c
    ms = 1e3 * sum(times) / len(times)
    val = cells / (ms * 1e-3) / 1e9
    line = {
        "impl": "reference", "metric": "bsw_gcups", "value": val, "unit": "GCUPS", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "int32 (scalar CPU)", "data": "synthetic",
        "pairs_per_s": npairs / (ms * 1e-3),
        "config": {"workload": "BASELINE config[1] shape (150bp reads, w=100, 8/16-bit tiers), bounded sample",
                   "pairs_per_step": npairs, "cells_per_step": int(cells), "band_w": W_BAND},
        "cpu_baseline": {"value": val, "unit": "GCUPS", "cores": cores, "kind": "port",
                         "sample": f"{npairs} pairs ({n_reads} reads) per step, oracle restatement (ksw_extend2 scalar), "
                                   f"pthreads x{cores}; the reference's SSE/AVX2 sources are not in /root/reference"},
        "e2e": {"value": val, "unit": "GCUPS", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    print(json.dumps(line), flush=True)


def main():
    args = parse()
    if args.impl == "reference":
        run_reference(args)
        return
    import torch
    import torch.distributed as dist
    import bswb200
    import ora

    rank, local_rank, world = rank_info()
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    torch.cuda.set_device(local_rank)
    os.environ["BSW_CUDA_DEVICES"] = str(local_rank)
    cores = host_threads()
    gen_threads = max(1, cores // max(1, world))

    # ---------------- workload (per rank: weak scaling) ----------------
    t_setup = time.perf_counter()
    d = build_workload(args.reads, 0xC2 + 7919 * rank, gen_threads)
    L, lr, lq = d["left"]; R, rr, rq = d["right"]
    sw = bswb200.BandedPairWiseSW(6, 1, 6, 1, 100, 5, None, 1, 4, 1)
    stream = torch.cuda.current_stream().cuda_stream

    def to_dev(a):
        return torch.from_numpy(a.view(np.uint8).reshape(-1)).cuda()

    d_lr, d_lq, d_rr, d_rq = to_dev(lr), to_dev(lq), to_dev(rr), to_dev(rq)
    # right-extension h0 = left score (caller protocol): computed once, outside the timed region
    dL = to_dev(L)
    sw.dev_getScores(dL, d_lr, d_lq, len(L), W_BAND, stream)
    torch.cuda.synchronize()
    Lres = dL.cpu().numpy().view(bswb200.SEQPAIR)
    ora.chain_h0(Lres, R)
    del dL
    L8, L16 = split_tiers(L)
    R8, R16 = split_tiers(R)
    calls = [(L8, d_lr, d_lq, lr, lq, 8), (L16, d_lr, d_lq, lr, lq, 16), (R8, d_rr, d_rq, rr, rq, 8), (R16, d_rr, d_rq, rr, rq, 16)]
    calls = [c for c in calls if len(c[0])]
    npairs = sum(len(c[0]) for c in calls)
    dev_pairs = [to_dev(c[0]) for c in calls]

    # ---------------- CPU baseline + cell-counter cross-check (rank 0, bounded sample) ----------------
    cpu_baseline = None
    if rank == 0:
        prm = ora.make_params()
        ns = min(args.cpu_sample, len(L))
        samp = np.ascontiguousarray(Lres[:ns])          # first ns left pairs (generator order is random)
        samp_in = np.ascontiguousarray(L[:ns])
        t0 = time.perf_counter()
        o = samp_in.copy()
        cpu_cells = ora.bsw_batch(o, lr, lq, prm, W_BAND, cores)
        cpu_dt = time.perf_counter() - t0
        for f in ora.BSW_OUT:                            # parity of the GPU results on the sample
            if not (o[f] == samp[f]).all():
                raise SystemExit(f"PARITY FAILURE on the CPU-baseline sample, field {f}")
        chk = bswb200.BandedPairWiseSW(6, 1, 6, 1, 100, 5, None, 1, 4, 1)
        chk.set_cell_counting(True)
        g = samp_in.copy()
        chk.getScores16(g, lr, lq, ns, 1, W_BAND)
        if chk.stats()["cells"] != cpu_cells:
            raise SystemExit("cell counter of the library disagrees with the oracle on the sample")
        chk.close()
        cpu_baseline = {"value": cpu_cells / cpu_dt / 1e9, "unit": "GCUPS", "cores": cores, "kind": "port",
                        "pairs_per_s": ns / cpu_dt,
                        "sample": f"first {ns} left-extension pairs of the workload ({cpu_cells} cells, {cpu_dt:.1f} s), "
                                  f"oracle restatement of ksw_extend2 (scalar C, pthreads x{cores}); the reference's "
                                  f"SSE4.1/AVX2 build is unavailable (no source in /root/reference)"}

    # The calls of a step are independent (tier 8 / tier 16, left / right with h0 already chained), like the batches of
    # the reference's worker threads.  They are issued round-robin on `args.streams` caller streams: calls on different
    # streams overlap inside the library (slot ring of bsw_dev_getScores), so the next call's prep / sort / DP blocks
    # fill the SMs while the previous call's last warps drain.  --streams 1 = strictly serial calls.
    side = [torch.cuda.Stream(priority=int(os.environ.get("BENCH_STREAM_PRIO", "-1"))) for _ in range(max(1, args.streams))]

    def one_step_resident(serial=False):
        for i, (c, dp) in enumerate(zip(calls, dev_pairs)):
            s = stream if serial or len(side) == 1 else side[i % len(side)].cuda_stream
            sw.dev_getScores(dp, c[1], c[2], len(c[0]), W_BAND, s)

    def join_side():
        cur = torch.cuda.current_stream()
        for s in side:
            cur.wait_stream(s)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---------------- DP cells of one step (counting variant of the kernel, outside the timed region) ----
    sw.set_cell_counting(True)
    sw.reset_stats()
    one_step_resident(serial=True)
    torch.cuda.synchronize()
    cells_step = sw.stats()["cells"]
    sw.set_cell_counting(False)

    # ---------------- resident timing ("value") ----------------
    for _ in range(args.warmup):
        one_step_resident()
    barrier()
    sw.reset_stats()
    sampler = ClockSampler(local_rank)
    sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    e0.record()
    for _ in range(args.steps):
        one_step_resident()
    join_side()
    e1.record()
    barrier()
    ms_total = e0.elapsed_time(e1)
    st = sw.stats()
    launches = st["kernel_launches"]

    # DP-kernel-only time for the roofline (CUDA events around the DP kernels on the library's stream)
    sw.set_timing(True)
    sw.reset_stats()
    one_step_resident(serial=True)
    torch.cuda.synchronize()
    kst = sw.stats()
    sw.set_timing(False)
    kernel_ms_step = kst["kernel_ms"]

    # ---------------- end-to-end through the C ABI on pinned host buffers ("e2e") ----------------
    keep = []
    host_calls = []
    pin_seq = {}
    for c in calls:
        hp, t1 = pinned_copy(c[0]); keep.append(t1)
        for key, arr in (("r" + str(id(c[3])), c[3]), ("q" + str(id(c[4])), c[4])):
            if key not in pin_seq:
                pa, t2 = pinned_copy(arr); keep.append(t2); pin_seq[key] = pa
        host_calls.append((hp, pin_seq["r" + str(id(c[3]))], pin_seq["q" + str(id(c[4]))], c[5]))

    def one_step_e2e():
        for hp, hr, hq, tier in host_calls:
            (sw.getScores8 if tier == 8 else sw.getScores16)(hp, hr, hq, len(hp), 1, W_BAND)

    for _ in range(max(1, min(2, args.warmup))):
        one_step_e2e()
    barrier()
    sw.reset_stats()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        one_step_e2e()
    torch.cuda.synchronize()
    e2e_ms = 1e3 * (time.perf_counter() - t0)
    est = sw.stats()
    barrier()

    sampler.stop()
    # ---------------- reduce over ranks: max of the times, sum of the work ----------------
    import sharding
    (ms_total, e2e_ms, kernel_ms_step), (cells_all, pairs_all) = sharding.reduce_step_stats(
        dist if world > 1 else None, torch, "cuda", [ms_total, e2e_ms, kernel_ms_step], [float(cells_step), float(npairs)])

    if rank == 0:
        ms_step = ms_total / args.steps
        value = cells_all / (ms_step * 1e-3) / 1e9
        e2e_step = e2e_ms / args.steps
        e2e_val = cells_all / (e2e_step * 1e-3) / 1e9
        peak_instr = bswb200.measure_dpx_peak(local_rank, 2)          # thread-level DPX instr/s, live
        peak_gcups = peak_instr * 2 / OPS_PER_CELL / 1e9
        # DP-kernel duration of a step = sum of the calls' fork->join CUDA-event windows around the DP launches (timing
        # mode: the calls of the step strictly one after another, a sync per call).  The timed resident run above issues
        # the calls on several caller streams, where these windows overlap each other and the next call's prep / sort;
        # its step time (prep + sort included) is reported beside it as `overlapped_calls`.
        ach = cells_step / (kernel_ms_step * 1e-3) / 1e9 if kernel_ms_step > 0 else 0.0
        ach_step = cells_step / (ms_step * 1e-3) / 1e9
        traffic = None                      # DRAM bytes per launch of the dominant kernel, from the committed ncu pass
        try:
            with open(os.path.join(ROOT, "profiles", "r01_launches_summary.json")) as f:
                traffic = json.load(f)["bsw_packed_kernel<false>"]["dram_bytes_per_launch"]
        except Exception:
            pass
        line = {
            "metric": "bsw_gcups", "value": value, "unit": "GCUPS", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "s16x2 (packed int16 DPX)", "data": "synthetic",
            "pairs_per_s": pairs_all / (ms_step * 1e-3),
            "config": {"workload": "BASELINE config[1]: BSW micro-bench, 150bp reads, w=100, getScores8 + getScores16 tiers, "
                                   "right h0 chained from left score",
                       "pairs_per_step_per_gpu": npairs, "cells_per_step_per_gpu": int(cells_step), "band_w": W_BAND,
                       "calls_per_step": len(calls), "caller_streams": len(side), "l2": "inputs (>1.5 GB per GPU) larger than L2; no flush needed",
                       "sharding": "independent pair slices per GPU, no collective"},
            "e2e": {"value": e2e_val, "unit": "GCUPS", "ms_per_step": e2e_step,
                    "pairs_per_s": pairs_all / (e2e_step * 1e-3),
                    "h2d_bytes_per_step": est["h2d_bytes"] // args.steps, "d2h_bytes_per_step": est["d2h_bytes"] // args.steps,
                    "api": "bsw_getScores8/bsw_getScores16 on pinned host SeqPair/sequence buffers"},
            "gpu_launches": int(launches),
            "roofline": {"bound": "dpx-int", "achieved": ach, "peak": peak_gcups, "unit": "GCUPS",
                         "frac": ach / peak_gcups if peak_gcups else None, "traffic": traffic,
                         "traffic_note": "dram__bytes_read+write per bsw_packed_kernel launch (one launch per query-length bin), "
                                         "ncu pass profiles/r01_launches_bench_steps1.csv; algorithmic ~0.055 B/cell",
                         "kernel": "bsw_packed_kernel", "kernel_ms_per_step": kernel_ms_step,
                         "kernel_ms_note": "sum of the calls' fork->join CUDA-event windows around the DP launches, calls "
                                           "strictly one after another (every call pays its own ramp-up and drain)",
                         "overlapped_calls": {"caller_streams": len(side), "step_ms": ms_step, "achieved": ach_step,
                                              "frac": ach_step / peak_gcups if peak_gcups else None,
                                              "note": "whole step of the timed resident run, prep + sort included"},
                         "ops_per_cell": OPS_PER_CELL,
                         "peak_source": "measured live: issue rate of independent VIMNMX3/VIADDMNMX.S16x2 streams "
                                        f"({peak_instr / 1e12:.2f} T lane-instr/s) x 2 cells / 8 ops; MEASURED_PEAKS.json has "
                                        "no integer-pipe figure; HBM is not the bound (~0.03 B/cell)"},
            "cpu_baseline": cpu_baseline,
            "clocks": sampler.summary(),
            "setup_s": time.perf_counter() - t_setup,
        }
        print(json.dumps(line), flush=True)
    sw.close()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
