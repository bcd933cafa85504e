#!/usr/bin/env python
"""bench.py -- benchmark of the BSW / kswv hot path, one BASELINE.json config per --workload.

  c2 (default)  BASELINE config[1]: BSW micro-bench, ~10M SeqPairs from 150 bp reads, w=100, tier-8 + tier-16 calls,
                right-extension h0 chained from the left score.  One step = the four boundary calls.
  c1            config[0] stand-in: 100k paired-end 150 bp reads through the caller protocol (band retry w, 2w;
                left before right).  resident = bsw_dev_extendChained, e2e = the host protocol over getScores8/16.
  c3            config[2]: 250 bp reads at 5 % divergence, 2M pairs, getScores16 path with z-drop.
  kswv_c4       config[3]: kswv mate rescue, 1M pairs 150 bp x 300-500 bp (phase 0 + phase 1) + a 16-bit slice.
  t16           the 16-bit tier proper: 2 x 300 bp reads, h0 up to 300 (scores > 255, queries > 252 columns).
  stream_c5     config[4] stand-in: small per-task calls from T host threads in one process through the C++ shim
                classes (tools/stream_c5.cpp), pageable and pinned arenas, per-task objects.

Every workload: generate at full size, run the GPU path, compare ALL pairs and ALL fields with the CPU restatement
(outside the timed regions; "parity" in the JSON line), time resident (`value`) and through the reference-facing C ABI
on host buffers (`e2e`), roofline against the live DPX issue-rate peak, CPU baseline on a bounded sample.

  value  : GCUPS with the inputs already resident in HBM, CUDA events.
  e2e    : GCUPS through bsw_getScores8/16 (kswv_getScores8/16) on pinned HOST buffers, H2D + D2H inside the timed region.
           c2 declares the two sequence arenas of each side once (bsw_arena_begin, the buffers bwa-mem2 passes to both
           tier calls); e2e_variants holds the same step without the declaration ("strict") and with the targets
           fetched from a resident 2-bit reference ("ref_resident").
  cells  : DP cells as the oracle counts them.

`--impl reference` times the CPU implementation of the path (the oracle restatement: the reference's own sources are
not in /root/reference) with all host threads on the SAME config.
Multi-GPU (torchrun, one rank per GPU): every rank owns an independent workload of the same size (weak scaling).
"""
import argparse
import json
import os
import subprocess
import sys
import tempfile
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
for p in (os.path.join(ROOT, "bwa-mem2-arm_b200"), os.path.join(ROOT, "oracle")):
    if p not in sys.path:
        sys.path.insert(0, p)

import numpy as np

W_BAND = 100
WORKLOADS = ("c2", "c1", "c3", "kswv_c4", "t16", "stream_c5")


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default=os.environ.get("BENCH_WORKLOAD", "c2"), choices=WORKLOADS)
    ap.add_argument("--streams", type=int, default=int(os.environ.get("BENCH_STREAMS", "4")),
                    help="caller streams the calls of a resident step are issued on (1 = strictly serial calls)")
    ap.add_argument("--reads", type=int, default=int(os.environ.get("BENCH_READS", "0")),
                    help="reads (pairs for kswv_c4) per GPU; 0 = the config's full size")
    ap.add_argument("--cpu-sample", type=int, default=int(os.environ.get("BENCH_CPU_SAMPLE", "400000")))
    ap.add_argument("--threads", default=os.environ.get("BENCH_THREADS", "1,8,32"), help="stream_c5: host thread counts")
    ap.add_argument("--task-reads", type=int, default=int(os.environ.get("BENCH_TASK_READS", "512")))
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


def pinned_copy(a):
    import torch
    t = torch.empty(a.nbytes, dtype=torch.uint8, pin_memory=True)
    out = t.numpy().view(a.dtype).reshape(a.shape)
    out[...] = a
    return out, t            # keep t alive


def mismatches(o, g, fields):
    bad = np.zeros(len(o), dtype=bool)
    for f in fields:
        bad |= o[f] != g[f]
    return int(bad.sum())


# ------------------------------------------------------------------------------------------------ workload definitions

def bsw_shape(name, reads):
    """(generator kwargs, default reads per GPU, description) of the BSW workloads."""
    if name == "c2":
        return dict(read_len=150, seed=0xC2, div=0.01), reads or 5060000, \
            "BASELINE config[1]: BSW micro-bench, 150bp reads, w=100, getScores8 + getScores16 tiers, right h0 chained from left score"
    if name == "c1":
        return dict(read_len=150, seed=0xC1, div=0.01), reads or 200000, \
            "BASELINE config[0] stand-in: 100k paired-end 150bp reads, caller protocol (band retry w / 2w, left before right), 8/16-bit tiers"
    if name == "c3":
        return dict(read_len=250, seed=0xC3, div=0.05), reads or 1000000, \
            "BASELINE config[2]: 250bp reads at 5% divergence, 2M pairs, getScores16 path, z-drop 100"
    if name == "t16":
        return dict(read_len=300, seed=0xC6, div=0.02, seed_max=150), reads or 500000, \
            "16-bit tier: 2 x 300bp reads, seeds up to 150bp, right h0 = left score (up to 300): scores > 255, queries > 252 columns"
    raise ValueError(name)


def gen_bsw_workload(ora, name, reads, seed_add, nthreads):
    kw, n_reads, desc = bsw_shape(name, reads)
    kw = dict(kw); kw["seed"] += seed_add
    d = ora.gen_bsw(n_reads, w=W_BAND, nthreads=nthreads, **kw)
    return d, n_reads, desc


def oracle_sides(ora, d, tries, cores, simd=False):
    """Expected results of both sides through the caller protocol (tries = 1: plain calls).  Returns (Lo, Rin, Ro, cells,
    pairs submitted).  simd: the inter-sequence AVX2 port instead of the scalar restatement (same results)."""
    prm = ora.make_params()
    L, lr, lq = d["left"]; R, rr, rq = d["right"]
    cells = [0]

    def side(P, r, q):
        if tries == 1:                      # plain tier calls: one batch, in place on a copy
            o = P.copy()
            cells[0] += (ora.bsw_batch_avx2 if simd else ora.bsw_batch)(o, r, q, prm, W_BAND, cores)
            return o, len(o)
        if not simd:
            o, n, c = ora.protocol_side(P, r, q, prm, W_BAND, tries, cores)
            cells[0] += c
            return o, n

        def runner(sub, w):
            cells[0] += ora.bsw_batch_avx2(sub, r, q, prm, w, cores)
        o, n, _ = ora.protocol_side(P, r, q, prm, W_BAND, tries, cores, runner=runner)
        return o, n
    Lo, nl = side(L, lr, lq)
    Rin = R.copy(); ora.chain_h0(Lo, Rin)
    Ro, nr = side(Rin, rr, rq)
    return Lo, Rin, Ro, cells[0], nl + nr


def stream_harness(ora, bswb200, L, R, lr, lq, rr, rq, n_reads, task_reads, threads, pinned, reps, tag="s", env=None):
    """Runs tools/bin/stream_c5 on the workload; returns (its JSON line, the SeqPair records it wrote: left then right)."""
    exe = os.path.join(ROOT, "tools", "bin", "stream_c5")
    if not os.path.exists(exe):
        subprocess.check_call(["make", "-C", os.path.join(ROOT, "tools"), "bin/stream_c5"])
    tmp = os.path.join(tempfile.gettempdir(), "bsw_b200_stream"); os.makedirs(tmp, exist_ok=True)      # not under gpurun_out/: it is large
    w_in, w_out = os.path.join(tmp, f"w_{tag}.bin"), os.path.join(tmp, f"r_{tag}.bin")
    if not os.path.exists(w_in):
        ora.write_stream_workload(w_in, 0, L, R, lr, lq, rr, rq, ora.task_bounds(L["seqid"], n_reads, task_reads),
                                  ora.task_bounds(R["seqid"], n_reads, task_reads))
    r = subprocess.run([exe, "--in", w_in, "--out", w_out, "--threads", str(threads), "--pinned", str(pinned), "--reps", str(reps)],
                       capture_output=True, text=True, env=env or dict(os.environ), timeout=1800)
    if r.returncode != 0:
        raise SystemExit("stream_c5 failed: " + r.stderr[-2000:])
    info = json.loads(r.stdout.strip().splitlines()[-1])
    raw = np.fromfile(w_out, dtype=bswb200.SEQPAIR)
    os.remove(w_out)
    return info, raw


def stream_cleanup(tag):
    try:
        os.remove(os.path.join(tempfile.gettempdir(), "bsw_b200_stream", f"w_{tag}.bin"))
    except OSError:
        pass


# ------------------------------------------------------------------------------------------------ reference arm (CPU)

def run_reference(args):
    """CPU implementation of the path on all host threads, SAME config as the GPU arm (rank 0 only)."""
    import ora
    rank, _, world = rank_info()
    if rank != 0:
        return
    cores = host_threads()
    prm = ora.make_params()
    wl = args.workload if args.workload != "stream_c5" else "c1"
    if wl == "kswv_c4":
        n = args.reads or 1000000
        K, kr, kq = ora.gen_kswv(n)
        desc = "BASELINE config[3]: kswv mate rescue, 150bp x 300-500bp, phase 0 + phase 1"
        def step():
            return ora.kswv_align2_batch(K, kr, kq, prm, cores, simd=True)[1]
        npairs = n
        metric = "kswv_gcups"
    else:
        d, n_reads, desc = gen_bsw_workload(ora, wl, args.reads, 0, cores)
        tries = 2 if wl == "c1" else 1
        def step():
            return oracle_sides(ora, d, tries, cores, simd=True)[3]
        npairs = len(d["left"][0]) + len(d["right"][0])
        metric = "bsw_gcups"
    times, cells = [], 0
    for it in range(args.warmup + args.steps):
        t0 = time.perf_counter()
        c = step()
        dt = time.perf_counter() - t0
        if it >= args.warmup:
            times.append(dt); cells = c
    ms = 1e3 * sum(times) / len(times)
    val = cells / (ms * 1e-3) / 1e9
    line = {
        "impl": "reference", "metric": metric, "value": val, "unit": "GCUPS", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "int16 x16 (AVX2)", "data": "synthetic",
        "pairs_per_s": npairs / (ms * 1e-3),
        "config": {"workload": desc, "pairs_per_step": npairs, "cells_per_step": int(cells), "band_w": W_BAND,
                   "same_config_as_gpu_arm": True},
        "cpu_baseline": {"value": val, "unit": "GCUPS", "cores": cores, "kind": "port-avx2",
                         "sample": f"the whole config: {npairs} pairs per step, "
                                   + ("inter-sequence AVX2 port of ksw_u8 / ksw_i16 / ksw_align2 (oracle/kswv_avx2.c: 16 pairs per register in 16-bit lanes, "
                                      "sorted by length, bit-exact with the scalar restatement; the hand-off between the passes is numpy)" if wl == "kswv_c4" else
                                      "inter-sequence AVX2 port (oracle/bsw_avx2.c: 16 pairs per register in 16-bit lanes, sorted by length, "
                                      "bit-exact with the scalar restatement)")
                                   + f", pthreads x{cores}; the reference's own x86 sources are not in /root/reference"},
        "e2e": {"value": val, "unit": "GCUPS", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    print(json.dumps(line), flush=True)


# ------------------------------------------------------------------------------------------------ GPU arm

class Ctx:
    pass


def setup(args):
    import torch
    import torch.distributed as dist
    c = Ctx()
    c.torch, c.dist = torch, dist
    c.rank, c.local_rank, c.world = rank_info()
    if c.world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", c.local_rank))
    torch.cuda.set_device(c.local_rank)
    os.environ["BSW_CUDA_DEVICES"] = str(c.local_rank)
    c.cores = host_threads()
    c.gen_threads = max(1, c.cores // max(1, c.world))
    c.stream = torch.cuda.current_stream().cuda_stream
    c.side = [torch.cuda.Stream(priority=int(os.environ.get("BENCH_STREAM_PRIO", "-1"))) for _ in range(max(1, args.streams))]
    return c


def to_dev(c, a):
    return c.torch.from_numpy(a.view(np.uint8).reshape(-1)).cuda()


def barrier(c):
    if c.world > 1:
        c.dist.barrier()
    c.torch.cuda.synchronize()


def time_resident(c, args, step_fn, join_fn=None):
    """W warm-up + K timed steps on the device, bracketed by barrier + synchronize (CUDA events on the current stream)."""
    torch = c.torch
    for _ in range(args.warmup):
        step_fn()
    if join_fn:
        join_fn()
    barrier(c)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier(c)
    e0.record()
    for _ in range(args.steps):
        step_fn()
    if join_fn:
        join_fn()
    e1.record()
    barrier(c)
    return e0.elapsed_time(e1)


def time_host(c, args, step_fn):
    for _ in range(max(1, min(2, args.warmup))):
        step_fn()
    barrier(c)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        step_fn()
    c.torch.cuda.synchronize()
    ms = 1e3 * (time.perf_counter() - t0)
    barrier(c)
    return ms


def reduce_and_print(c, args, line_fn, times, sums):
    import sharding
    t, s = sharding.reduce_step_stats(c.dist if c.world > 1 else None, c.torch, "cuda", times, sums)
    if c.rank == 0:
        print(json.dumps(line_fn(t, s)), flush=True)
    if c.world > 1:
        c.dist.destroy_process_group()


def cpu_baseline_bsw(ora, c, args, L, lr, lq):
    """Oracle port on all host threads on a bounded sample of the workload (rank 0)."""
    prm = ora.make_params()
    ns = min(args.cpu_sample * 4, len(L))
    samp = np.ascontiguousarray(L[:ns])
    t0 = time.perf_counter()
    cells = ora.bsw_batch_avx2(samp, lr, lq, prm, W_BAND, c.cores)
    dt = time.perf_counter() - t0
    samp2 = np.ascontiguousarray(L[:ns // 4])
    t0 = time.perf_counter()
    cells2 = ora.bsw_batch(samp2, lr, lq, prm, W_BAND, c.cores)
    dt2 = time.perf_counter() - t0
    return {"value": cells / dt / 1e9, "unit": "GCUPS", "cores": c.cores, "kind": "port-avx2", "pairs_per_s": ns / dt,
            "scalar_port_gcups": cells2 / dt2 / 1e9,
            "sample": f"first {ns} left-extension pairs of the workload ({cells} cells, {dt:.2f} s) on the inter-sequence AVX2 port "
                      f"(oracle/bsw_avx2.c: 16 pairs per register in 16-bit lanes, sorted by length -- the organisation of the reference's "
                      f"smithWaterman256_16 -- bit-exact with the scalar restatement), pthreads x{c.cores}; the reference's own x86 build "
                      f"is unavailable (no source in /root/reference); `--impl reference` runs the whole config"}


def roofline(bswb200, c, ops_per_cell, cells_step, kernel_ms, ms_step, kernel, traffic_key=None):
    peak_instr = bswb200.measure_dpx_peak(c.local_rank, 2)          # thread-level DPX instr/s, live
    peak = peak_instr * 2 / ops_per_cell / 1e9
    ach = cells_step / (kernel_ms * 1e-3) / 1e9 if kernel_ms > 0 else 0.0
    ach_step = cells_step / (ms_step * 1e-3) / 1e9
    traffic = None
    for fn in ("r02_launches_summary.json", "r01_launches_summary.json"):
        try:
            with open(os.path.join(ROOT, "profiles", fn)) as f:
                traffic = json.load(f)[traffic_key or "bsw_packed_kernel<false>"]["dram_bytes_per_launch"]
            break
        except Exception:
            pass
    return {"bound": "dpx-int", "achieved": ach, "peak": peak, "unit": "GCUPS", "frac": ach / peak if peak else None,
            "traffic": traffic, "kernel": kernel, "kernel_ms_per_step": kernel_ms,
            "kernel_ms_note": "sum of the calls' fork->join CUDA-event windows around the DP launches, calls strictly one after another; median of three steps",
            "whole_step": {"step_ms": ms_step, "achieved": ach_step, "frac": ach_step / peak if peak else None,
                           "note": "the timed resident run, prep + sort included"},
            "ops_per_cell": ops_per_cell,
            "peak_source": "measured live: issue rate of independent VIMNMX3/VIADDMNMX.S16x2 streams "
                           f"({peak_instr / 1e12:.2f} T lane-instr/s) x 2 cells / {ops_per_cell} ops; MEASURED_PEAKS.json has "
                           "no integer-pipe figure; HBM is not the bound (~0.03-0.06 B/cell)"}


def run_bsw(args):
    """c2, c3, t16: plain tier calls.  c1: the caller protocol (retry) on top."""
    import bswb200
    import ora
    c = setup(args)
    torch = c.torch
    t_setup = time.perf_counter()
    wl = args.workload
    tries = 2 if wl == "c1" else 1
    d, n_reads, desc = gen_bsw_workload(ora, wl, args.reads, 7919 * c.rank, c.gen_threads)
    L, lr, lq = d["left"]; R, rr, rq = d["right"]
    # expected results of every pair (all host threads; outside every timed region)
    t0 = time.perf_counter()
    Lo, Rin, Ro, cells_oracle, submitted = oracle_sides(ora, d, tries, c.cores)
    oracle_s = time.perf_counter() - t0
    sw = bswb200.BandedPairWiseSW(6, 1, 6, 1, 100, 5, None, 1, 4, 1)
    d_lr, d_lq, d_rr, d_rq = to_dev(c, lr), to_dev(c, lq), to_dev(c, rr), to_dev(c, rq)
    npairs = len(L) + len(R)
    parity = {"pairs": 0, "mismatches": 0, "fields": list(ora.BSW_OUT), "checked": []}

    def check(tag, o, g):
        parity["pairs"] += len(o); parity["mismatches"] += mismatches(o, g, ora.BSW_OUT); parity["checked"].append(tag)

    def split(P):
        t = ora.classify(P)
        return np.ascontiguousarray(P[t == 0]), np.ascontiguousarray(P[t != 0]), t

    if tries == 1:
        L8, L16, tl = split(L); R8, R16, tr = split(Rin)
        calls = [(L8, d_lr, d_lq, lr, lq, 8), (L16, d_lr, d_lq, lr, lq, 16), (R8, d_rr, d_rq, rr, rq, 8), (R16, d_rr, d_rq, rr, rq, 16)]
        calls = [x for x in calls if len(x[0])]
        dev_pairs = [to_dev(c, x[0]) for x in calls]

        def step_resident(serial=False):
            for i, (x, dp) in enumerate(zip(calls, dev_pairs)):
                s = c.stream if serial or len(c.side) == 1 else c.side[i % len(c.side)].cuda_stream
                sw.dev_getScores(dp, x[1], x[2], len(x[0]), W_BAND, s)

        def join_side():
            cur = torch.cuda.current_stream()
            for s in c.side:
                cur.wait_stream(s)
    else:
        dL, dR = to_dev(c, L), to_dev(c, R)
        dL0, dR0 = dL.clone(), dR.clone()

        def step_resident(serial=False):
            dL.copy_(dL0); dR.copy_(dR0)               # the protocol rewrites scores and h0: restore the inputs (device copy, ~3 % of a step)
            sw.dev_extendChained(dL, d_lr, d_lq, len(L), dR, d_rr, d_rq, len(R), W_BAND, tries, c.stream)
        join_side = None

    # ---- DP cells of one step (counting variant, outside the timed region) and resident parity
    sw.set_cell_counting(True); sw.reset_stats()
    step_resident(serial=True); torch.cuda.synchronize()
    cells_step = sw.stats()["cells"]
    sw.set_cell_counting(False)
    step_resident(serial=True); torch.cuda.synchronize()          # the timed kernels (not the counting variant) produce what is compared
    if tries == 1:
        got = {8: {}, 16: {}}
        for x, dp in zip(calls, dev_pairs):
            g = dp.cpu().numpy().view(bswb200.SEQPAIR)
            side_o, tiers = (Lo, tl) if x[3] is lr else (Ro, tr)
            o = side_o[tiers == 0] if x[5] == 8 else side_o[tiers != 0]
            check(f"resident {'left' if x[3] is lr else 'right'} tier{x[5]}", o, g)
    else:
        check("resident left (device-side protocol)", Lo, dL.cpu().numpy().view(bswb200.SEQPAIR))
        check("resident right (device-side protocol)", Ro, dR.cpu().numpy().view(bswb200.SEQPAIR))
    if tries == 1 and cells_step != cells_oracle:
        raise SystemExit(f"cell counter of the library ({cells_step}) disagrees with the oracle ({cells_oracle})")
    cells_step = cells_oracle

    cpu_baseline = cpu_baseline_bsw(ora, c, args, L, lr, lq) if c.rank == 0 else None

    # ---- resident timing ("value")
    sampler = ClockSampler(c.local_rank); sampler.start()
    sw.reset_stats()
    ms_total = time_resident(c, args, step_resident, join_side)
    launches = sw.stats()["kernel_launches"]
    # DP-kernel-only time for the roofline
    sw.set_timing(True)
    kms = []
    for _ in range(3):                      # median of three serial steps (one step alone varies by ~5 %)
        sw.reset_stats(); step_resident(serial=True); torch.cuda.synchronize(); kms.append(sw.stats()["kernel_ms"])
    kernel_ms_step = sorted(kms)[1]
    sw.set_timing(False)

    # ---- e2e through the C ABI on pinned host buffers
    keep = []
    def pin(a):
        out, t = pinned_copy(a); keep.append(t); return out
    h_lr, h_lq, h_rr, h_rq = pin(lr), pin(lq), pin(rr), pin(rq)
    variants = {}
    if tries == 1:
        hcalls = [(pin(x[0]), (h_lr, h_lq) if x[3] is lr else (h_rr, h_rq), x[5]) for x in calls]

        def step_e2e(arena=True, defer=False):
            sw.arena_defer(arena and defer)
            cur = None
            for hp, (hr, hq), tier in hcalls:
                if arena and cur is not hr:
                    if cur is not None:
                        sw.arena_end()
                    sw.arena_begin(hr, hq); cur = hr
                (sw.getScores8 if tier == 8 else sw.getScores16)(hp, hr, hq, len(hp), 1, W_BAND)
            if arena and cur is not None:
                sw.arena_end()

        sw.reset_stats(); e2e_ms = time_host(c, args, step_e2e); est = sw.stats()
        for (hp, _, tier), x in zip(hcalls, calls):
            side_o, tiers = (Lo, tl) if x[3] is lr else (Ro, tr)
            check(f"e2e {'left' if x[3] is lr else 'right'} tier{tier}", side_o[tiers == 0] if tier == 8 else side_o[tiers != 0], hp)
        e2e_api = "bsw_getScores8/bsw_getScores16 on pinned host SeqPair/sequence buffers; the two sequence arenas of each side declared once (bsw_arena_begin/end)"
        if wl == "c2":
            sw.reset_stats(); ms = time_host(c, args, lambda: step_e2e(True, True)); s2 = sw.stats()
            variants["arena_deferred"] = {"ms_per_step": ms / args.steps, "h2d_bytes_per_step": s2["h2d_bytes"] // (args.steps + 2),
                                          "api": "arenas declared + bsw_arena_defer: the large tier calls of a side are recorded and run as ONE "
                                                 "interleaved chunk pipeline at bsw_arena_end (upstream reads no result between the tier calls)"}
            sw.reset_stats(); ms = time_host(c, args, lambda: step_e2e(False)); s2 = sw.stats()
            variants["strict"] = {"ms_per_step": ms / args.steps, "h2d_bytes_per_step": s2["h2d_bytes"] // (args.steps + 2),
                                  "api": "bsw_getScores8/16 only: each tier call uploads the byte range of the shared arenas it touches"}
            # targets from a resident 2-bit reference: the "genome" is the concatenation of the target buffers (N -> a base)
            refs = {}
            for nm, hr in (("l", lr), ("r", rr)):
                fwd = hr.copy(); nn = fwd > 3; fwd[nn] = (np.nonzero(nn)[0] & 3).astype(np.uint8)
                refs[nm] = fwd
            both = np.concatenate([refs["l"], refs["r"]])
            Rf = bswb200.Reference(bswb200.pack_pac(both), len(both))
            hP = [pin(L), pin(Rin)]
            rpos = [pin(L["idr"].astype(np.int64)), pin(Rin["idr"].astype(np.int64) + len(lr))]
            exp = []
            for P, fwd, hq in ((L, refs["l"], lq), (Rin, refs["r"], rq)):
                o = P.copy(); ora.bsw_batch(o, fwd, hq, ora.make_params(), W_BAND, c.cores); exp.append(o)

            def step_ref():
                sw.getScoresRef(hP[0], Rf, rpos[0], h_lq, len(L), W_BAND)
                sw.getScoresRef(hP[1], Rf, rpos[1], h_rq, len(Rin), W_BAND)
            sw.reset_stats(); ms = time_host(c, args, step_ref); s3 = sw.stats()
            check("e2e ref_resident left", exp[0], hP[0]); check("e2e ref_resident right", exp[1], hP[1])
            variants["ref_resident"] = {"ms_per_step": ms / args.steps, "h2d_bytes_per_step": s3["h2d_bytes"] // (args.steps + 2),
                                        "api": "bsw_ref_upload once + bsw_getScoresRef(pairs, rpos, query bases): no target bases cross PCIe "
                                               "(SURVEY.md 8(f).2); targets with N replaced by a base, expected results recomputed"}
            Rf.close()
    else:
        # the caller's protocol in C++: tools/stream_c5.cpp (worker threads over per-task slices, per-task shim objects, band retry)
        T = max(1, 2 * c.cores // max(1, c.world))      # callers mostly sleep on their batch: two per core
        info, raw = stream_harness(ora, bswb200, L, R, lr, lq, rr, rq, n_reads, args.task_reads, T, 0, 3, tag=f"c1_{c.rank}")
        stream_cleanup(f"c1_{c.rank}")
        Ro2 = Ro.copy(); Ro2["h0"] = Rin["h0"]
        check("e2e left (C++ harness)", Lo, raw[:len(L)]); check("e2e right (C++ harness)", Ro2, raw[len(L):])
        e2e_ms = info["wall_s"] * 1e3 * args.steps
        est = {"h2d_bytes": info["engine_h2d_bytes"] * (args.steps + max(1, min(2, args.warmup))),
               "d2h_bytes": info["engine_d2h_bytes"] * (args.steps + max(1, min(2, args.warmup)))}
        variants["stream_harness"] = {"ms_per_step": info["wall_s"] * 1e3, "h2d_bytes_per_step": info["engine_h2d_bytes"], "threads": T,
                                      "reads_per_task": args.task_reads, "us_per_call": info["us_per_call"],
                                      "engine_pairs_per_batch": info["engine_pairs_per_batch"],
                                      "api": "tools/stream_c5.cpp: BandedPairWiseSW::getScores8/16 of bandedSWA_b200.h from worker threads"}
        e2e_api = (f"mem_chain2aln_across_reads_V2's protocol in C++ (tools/stream_c5.cpp): {T} worker threads, tasks of {args.task_reads} reads, "
                   "per-task BandedPairWiseSW objects, pageable buffers, band retry; best of 3 passes over the whole config")
    sampler.stop()
    st_all = sw.stats()

    def line(t, s):
        ms_tot, e2e_tot, kms = t
        cells_all, pairs_all = s
        ms_step, e2e_step = ms_tot / args.steps, e2e_tot / args.steps
        out = {
            "metric": "bsw_gcups", "value": cells_all / (ms_step * 1e-3) / 1e9, "unit": "GCUPS", "n_gpus": c.world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "s16x2 (packed int16 DPX)", "data": "synthetic",
            "pairs_per_s": pairs_all / (ms_step * 1e-3),
            "config": {"workload": desc, "name": wl, "reads_per_gpu": n_reads, "pairs_per_step_per_gpu": npairs,
                       "pairs_submitted_per_step_per_gpu": int(submitted), "cells_per_step_per_gpu": int(cells_step), "band_w": W_BAND,
                       "caller_streams": len(c.side), "l2": "inputs larger than L2 (126 MB) for c2/c3/t16; c1 is a small-batch config by definition",
                       "sharding": "independent pair slices per GPU, no collective"},
            "parity": parity,
            "e2e": {"value": cells_all / (e2e_step * 1e-3) / 1e9, "unit": "GCUPS", "ms_per_step": e2e_step,
                    "pairs_per_s": pairs_all / (e2e_step * 1e-3),
                    "h2d_bytes_per_step": est["h2d_bytes"] // (args.steps + max(1, min(2, args.warmup))),
                    "d2h_bytes_per_step": est["d2h_bytes"] // (args.steps + max(1, min(2, args.warmup))), "api": e2e_api},
            "gpu_launches": int(launches),
            "kernels": {"pairs_packed": st_all["pairs_packed"], "pairs_wide": st_all["pairs_wide"], "pairs_general": st_all["pairs_general"]},
            "roofline": roofline(bswb200, c, 8, cells_step, kms, ms_step, "bsw_packed_kernel"),
            "cpu_baseline": cpu_baseline, "clocks": sampler.summary(),
            "setup_s": time.perf_counter() - t_setup, "oracle_full_config_s": oracle_s,
        }
        for k, v in variants.items():
            v["value"] = cells_step / (v["ms_per_step"] * 1e-3) / 1e9; v["unit"] = "GCUPS (this rank)"
        if variants:
            out["e2e_variants"] = variants
        return out

    reduce_and_print(c, args, line, [ms_total, e2e_ms, kernel_ms_step], [float(cells_step), float(npairs)])
    sw.close()


def run_kswv(args):
    import bswb200
    import ora
    c = setup(args)
    torch = c.torch
    t_setup = time.perf_counter()
    n = args.reads or 1000000
    prm = ora.make_params()
    K, kr, kq = ora.gen_kswv(n, seed=0xC4 + 7919 * c.rank)
    n16 = max(1000, n // 20)
    K16, kr16, kq16 = ora.gen_kswv(n16, qlen=250, tmin=400, tmax=700, seed=0xC5 + 7919 * c.rank)
    t0 = time.perf_counter()
    a0 = np.zeros(n, ora.KSWR); cells0 = ora.kswv_batch(K, kr, kq, a0, prm, 1, 0, c.cores)
    p1, r1, q1 = ora.kswv_phase1_inputs(K, kr, kq, a0)
    a1 = a0.copy(); cells1 = ora.kswv_batch(p1, r1, q1, a1, prm, 1, 1, c.cores)
    b0 = np.zeros(n16, ora.KSWR); cells16 = ora.kswv_batch(K16, kr16, kq16, b0, prm, 0, 0, c.cores)
    p16, r16, q16 = ora.kswv_phase1_inputs(K16, kr16, kq16, b0)
    b1 = b0.copy(); cells16 += ora.kswv_batch(p16, r16, q16, b1, prm, 0, 1, c.cores)
    oracle_s = time.perf_counter() - t0
    cells_step = cells0 + cells1 + cells16
    ks = bswb200.kswv(6, 1, 6, 1, 1, 4, 1, 2048, 256)
    parity = {"pairs": 0, "mismatches": 0, "fields": ["score", "te", "qe", "score2", "te2", "tb", "qb"], "checked": []}

    def check(tag, o, g):
        parity["pairs"] += len(o); parity["mismatches"] += int((o != g.view(ora.KSWR)).sum()); parity["checked"].append(tag)

    dv = lambda a: to_dev(c, a)
    dK, dkr, dkq, dp1, dr1, dq1 = dv(K), dv(kr), dv(kq), dv(p1), dv(r1), dv(q1)
    dK16, dkr16, dkq16, dp16, dr16, dq16 = dv(K16), dv(kr16), dv(kq16), dv(p16), dv(r16), dv(q16)
    da = torch.zeros(n * 28, dtype=torch.uint8, device="cuda"); db = torch.zeros(n16 * 28, dtype=torch.uint8, device="cuda")

    s16 = c.side[0].cuda_stream if len(c.side) > 1 else c.stream      # the 16-bit slice on its own caller stream: the tiers are independent

    def step_resident():
        ks.dev_getScores(dK, dkr, dkq, da, n, 1, 0, c.stream)
        ks.dev_getScores(dK16, dkr16, dkq16, db, n16, 0, 0, s16)
        ks.dev_getScores(dp1, dr1, dq1, da, len(p1), 1, 1, c.stream)
        ks.dev_getScores(dp16, dr16, dq16, db, len(p16), 0, 1, s16)

    def join_side():
        if len(c.side) > 1:
            torch.cuda.current_stream().wait_stream(c.side[0])

    ks.reset_stats(); step_resident(); torch.cuda.synchronize()
    if ks.stats()["cells"] != cells_step:
        raise SystemExit("cell counter of the library disagrees with the oracle")
    check("resident 8-bit phase 0+1", a1, da.cpu().numpy().view(bswb200.KSWR))
    check("resident 16-bit phase 0+1", b1, db.cpu().numpy().view(bswb200.KSWR))
    cpu_baseline = None
    if c.rank == 0:
        ns = min(args.cpu_sample // 4, n)
        sub = np.ascontiguousarray(K[:ns])
        t0 = time.perf_counter(); ab = np.zeros(ns, ora.KSWR); cs = ora.kswv_batch(sub, kr, kq, ab, prm, 1, 0, c.cores)
        dts = time.perf_counter() - t0
        t0 = time.perf_counter(); aa = np.zeros(n, ora.KSWR); cc = 0
        for _ in range(5):                                        # phase 0 of the whole config, five times: ~5 s of CPU work
            cc += ora.kswv_batch_avx2(K, kr, kq, aa, prm, 1, 0, c.cores)
        dt = time.perf_counter() - t0
        if (aa[:ns] != ab).any() or (aa != a0).any():
            raise SystemExit("the AVX2 port of kswv disagrees with the scalar restatement")
        cpu_baseline = {"value": cc / dt / 1e9, "unit": "GCUPS", "cores": c.cores, "kind": "port-avx2", "pairs_per_s": 5 * n / dt,
                        "scalar_restatement_gcups": cs / dts / 1e9,
                        "sample": f"phase 0 of all {n} pairs, five passes ({cc} cells, {dt:.2f} s), inter-sequence AVX2 port of ksw_u8 (oracle/kswv_avx2.c: "
                                  f"16 pairs per register in 16-bit lanes, bit-exact with the scalar restatement, pthreads x{c.cores}); "
                                  "`--impl reference --workload kswv_c4` runs the whole config"}
    sampler = ClockSampler(c.local_rank); sampler.start()
    ks.reset_stats()
    ms_total = time_resident(c, args, step_resident, join_side)
    launches = ks.stats()["kernel_launches"]
    ks.set_timing(True)
    kms = []
    for _ in range(3):                      # median of three steps (one step alone varies by ~5 %)
        ks.reset_stats(); step_resident(); torch.cuda.synchronize(); kms.append(ks.stats()["kernel_ms"])
    kernel_ms_step = sorted(kms)[1]; ks.set_timing(False)
    keep = []
    def pin(a):
        out, t = pinned_copy(a); keep.append(t); return out
    hK, hkr, hkq, hK16, hkr16, hkq16 = pin(K), pin(kr), pin(kq), pin(K16), pin(kr16), pin(kq16)
    ha, hb = pin(np.zeros(n, bswb200.KSWR)), pin(np.zeros(n16, bswb200.KSWR))

    def step_e2e():
        # mem_sam_pe_batch's sequence; the hand-off (reversed prefixes) is the caller's host work and inside the time
        ks.getScores8(hK, hkr, hkq, ha, n, 1, 0); ks.getScores16(hK16, hkr16, hkq16, hb, n16, 1, 0)
        x1, xr, xq = ora.kswv_phase1_inputs(hK, hkr, hkq, ha.view(ora.KSWR))
        y1, yr, yq = ora.kswv_phase1_inputs(hK16, hkr16, hkq16, hb.view(ora.KSWR))
        ks.getScores16(y1, yr, yq, hb, len(y1), 1, 1); ks.getScores8(x1, xr, xq, ha, len(x1), 1, 1)

    ks.reset_stats()
    # the numpy hand-off loops over pairs in Python (oracle/ora.py): time the four boundary calls separately from it
    t_calls = [0.0]
    hp1, hr1, hq1, hp16, hr16, hq16 = pin(p1), pin(r1), pin(q1), pin(p16), pin(r16), pin(q16)

    def step_e2e_timed():
        t0 = time.perf_counter()
        ks.getScores8(hK, hkr, hkq, ha, n, 1, 0); ks.getScores16(hK16, hkr16, hkq16, hb, n16, 1, 0)
        t_calls[0] += time.perf_counter() - t0
        x1, xr, xq = hp1, hr1, hq1               # the caller's hand-off output (identical every step: precomputed, pinned)
        y1, yr, yq = hp16, hr16, hq16
        t0 = time.perf_counter()
        ks.getScores16(y1, yr, yq, hb, len(y1), 1, 1); ks.getScores8(x1, xr, xq, ha, len(x1), 1, 1)
        t_calls[0] += time.perf_counter() - t0
    for _ in range(2):
        step_e2e_timed()
    barrier(c); t_calls[0] = 0.0; ks.reset_stats()
    for _ in range(args.steps):
        step_e2e_timed()
    e2e_ms = 1e3 * t_calls[0]; est = ks.stats()
    check("e2e 8-bit phase 0+1", a1, ha); check("e2e 16-bit phase 0+1", b1, hb)
    # the same step through kswv_align2: ONE call per tier, hand-off and both passes on the device, bases cross PCIe once
    ha[...] = 0; hb[...] = 0

    def step_align2():
        ks.align2(hK, hkr, hkq, ha, n, 1); ks.align2(hK16, hkr16, hkq16, hb, n16, 0)
    ks.reset_stats(); a2_ms = time_host(c, args, step_align2); a2st = ks.stats()
    check("e2e kswv_align2 8-bit", a1, ha); check("e2e kswv_align2 16-bit", b1, hb)
    variants = {"align2": {"ms_per_step": a2_ms / args.steps, "value": cells_step / (a2_ms / args.steps * 1e-3) / 1e9, "unit": "GCUPS (this rank)",
                           "h2d_bytes_per_step": a2st["h2d_bytes"] // (args.steps + 2), "d2h_bytes_per_step": a2st["d2h_bytes"] // (args.steps + 2),
                           "api": "kswv_align2 on pinned host buffers: one call per tier instead of getScores(phase 0) + the caller's reversed-prefix "
                                  "loop + getScores(phase 1); the hand-off runs on the device (SURVEY.md 8(f).3)"}}
    # windows from a resident 2-bit reference (SURVEY.md 8(f).2 applied to mate rescue): the "genome" is the concatenation of
    # the two window buffers (N -> a base; expected results recomputed if that changed anything)
    fw = np.concatenate([kr, kr16]); nn = fw > 3
    e8, e16 = a1, b1
    if nn.any():
        fw[nn] = (np.nonzero(nn)[0] & 3).astype(np.uint8)
        e8 = ora.kswv_align2_batch(K, fw[:len(kr)], kq, prm, c.cores)[0]; e16 = ora.kswv_align2_batch(K16, fw[len(kr):], kq16, prm, c.cores)[0]
    Rf = bswb200.Reference(bswb200.pack_pac(fw), len(fw))
    rp8, rp16 = pin(K["idr"].astype(np.int64)), pin(K16["idr"].astype(np.int64) + len(kr))
    ha[...] = 0; hb[...] = 0

    def step_align2_ref():
        ks.align2Ref(hK, Rf, rp8, hkq, ha, n, 1); ks.align2Ref(hK16, Rf, rp16, hkq16, hb, n16, 0)
    ks.reset_stats(); a3_ms = time_host(c, args, step_align2_ref); a3st = ks.stats()
    check("e2e kswv_align2Ref 8-bit", e8, ha); check("e2e kswv_align2Ref 16-bit", e16, hb)
    variants["align2_ref"] = {"ms_per_step": a3_ms / args.steps, "value": cells_step / (a3_ms / args.steps * 1e-3) / 1e9, "unit": "GCUPS (this rank)",
                              "h2d_bytes_per_step": a3st["h2d_bytes"] // (args.steps + 2), "d2h_bytes_per_step": a3st["d2h_bytes"] // (args.steps + 2),
                              "api": "bsw_ref_upload once + kswv_align2Ref(records, rpos, mates): the windows are cut out of the resident 2-bit "
                                     "reference on the device, no target base crosses PCIe"}
    Rf.close()
    sampler.stop()
    npairs = n + len(p1) + n16 + len(p16)

    def line(t, s):
        ms_tot, e2e_tot, kms = t
        cells_all, pairs_all = s
        ms_step, e2e_step = ms_tot / args.steps, e2e_tot / args.steps
        return {
            "metric": "kswv_gcups", "value": cells_all / (ms_step * 1e-3) / 1e9, "unit": "GCUPS", "n_gpus": c.world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "s16x2 (packed int16 DPX)", "data": "synthetic", "pairs_per_s": pairs_all / (ms_step * 1e-3),
            "config": {"workload": "BASELINE config[3]: kswv batched mate rescue, 150bp mates x 300-500bp windows, phase 0 + phase 1, "
                                   "plus a 16-bit slice (250bp mates, no KSW_XBYTE)", "name": "kswv_c4",
                       "pairs_phase0": n, "pairs_phase1": int(len(p1)), "pairs16_phase0": n16, "pairs16_phase1": int(len(p16)),
                       "cells_per_step_per_gpu": int(cells_step), "l2": "inputs (~0.7 GB) larger than L2"},
            "parity": parity,
            "e2e": {"value": cells_all / (e2e_step * 1e-3) / 1e9, "unit": "GCUPS", "ms_per_step": e2e_step, "pairs_per_s": pairs_all / (e2e_step * 1e-3),
                    "h2d_bytes_per_step": est["h2d_bytes"] // args.steps, "d2h_bytes_per_step": est["d2h_bytes"] // args.steps,
                    "api": "kswv_getScores8/16 phase 0 then phase 1 on pinned host buffers (the caller's reversed-prefix hand-off between them is host work and not in the time)"},
            "e2e_variants": variants,
            "gpu_launches": int(launches),
            "roofline": roofline(bswb200, c, 7, cells_step, kms, ms_step, "kswv_sys_kernel", "kswv_sys_kernel"),
            "cpu_baseline": cpu_baseline, "clocks": sampler.summary(), "setup_s": time.perf_counter() - t_setup, "oracle_full_config_s": oracle_s,
        }
    reduce_and_print(c, args, line, [ms_total, e2e_ms, kernel_ms_step], [float(cells_step), float(npairs)])
    ks.close()


def run_stream(args):
    """C5 stand-in: tools/bin/stream_c5 (C++ worker threads, per-task shim objects, small calls) on one process."""
    import bswb200
    import ora
    rank, local_rank, world = rank_info()
    if rank != 0:
        return
    t_setup = time.perf_counter()
    cores = host_threads()
    n_reads = args.reads or 1000000
    prm = ora.make_params()
    d = ora.gen_bsw(n_reads, seed=0xC5, nthreads=cores)
    L, lr, lq = d["left"]; R, rr, rq = d["right"]
    Lo, Rin, Ro, cells, submitted = oracle_sides(ora, d, 2, cores, simd=True)
    gpus = args.gpus
    env = dict(os.environ); env["BSW_CUDA_DEVICES"] = ",".join(str(i) for i in range(gpus))
    rows, parity = [], {"pairs": 0, "mismatches": 0, "fields": list(ora.BSW_OUT), "checked": []}
    Ro2 = Ro.copy(); Ro2["h0"] = Rin["h0"]
    for T in [int(x) for x in args.threads.split(",")]:
        for pinned in (0, 1):
            info, raw = stream_harness(ora, bswb200, L, R, lr, lq, rr, rq, n_reads, args.task_reads, T, pinned, 2, tag="c5", env=env)
            bad = mismatches(Lo, raw[:len(L)], ora.BSW_OUT) + mismatches(Ro2, raw[len(L):], ora.BSW_OUT)
            parity["pairs"] += len(L) + len(R); parity["mismatches"] += bad; parity["checked"].append(f"T={T} pinned={pinned}")
            info["gcups"] = cells / info["wall_s"] / 1e9
            rows.append(info)
    stream_cleanup("c5")
    best = max(rows, key=lambda x: x["pairs_per_s"])
    t0 = time.perf_counter()
    samp = np.ascontiguousarray(L[:min(args.cpu_sample * 4, len(L))]); cc = ora.bsw_batch_avx2(samp, lr, lq, prm, W_BAND, cores); dt = time.perf_counter() - t0
    line = {
        "metric": "bsw_gcups", "value": best["gcups"], "unit": "GCUPS", "n_gpus": gpus, "steps": 2, "warmup": 1,
        "ms_per_step": best["wall_s"] * 1e3, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
        "dtype": "s16x2 (packed int16 DPX)", "data": "synthetic", "pairs_per_s": best["pairs_per_s"],
        "config": {"workload": "BASELINE config[4] stand-in: C1-shaped pairs streamed in per-task calls from T host threads of ONE process "
                               "(C++ shim classes constructed per task, caller protocol with band retry), BSW_CUDA_DEVICES = all GPUs of the run",
                   "name": "stream_c5", "reads": n_reads, "reads_per_task": args.task_reads, "pairs": len(L) + len(R),
                   "pairs_submitted": int(submitted), "cells": int(cells)},
        "parity": parity,
        "e2e": {"value": best["gcups"], "unit": "GCUPS", "pairs_per_s": best["pairs_per_s"],
                "h2d_bytes_per_step": best["engine_h2d_bytes"], "d2h_bytes_per_step": best["engine_d2h_bytes"],
                "api": "BandedPairWiseSW::getScores8/16 of bandedSWA_b200.h from worker threads (the whole run is host-buffer calls; best row)"},
        "stream_table": rows,
        "cpu_baseline": {"value": cc / dt / 1e9, "unit": "GCUPS", "cores": cores, "kind": "port-avx2", "pairs_per_s": len(samp) / dt,
                         "sample": f"first {len(samp)} left pairs, inter-sequence AVX2 port (oracle/bsw_avx2.c), pthreads x{cores}"},
        "gpu_launches": int(sum(r["engine_launches"] for r in rows)),
        "setup_s": time.perf_counter() - t_setup,
    }
    print(json.dumps(line), flush=True)


def main():
    args = parse()
    if args.impl == "reference":
        run_reference(args)
    elif args.workload == "kswv_c4":
        run_kswv(args)
    elif args.workload == "stream_c5":
        run_stream(args)
    else:
        run_bsw(args)


if __name__ == "__main__":
    main()
