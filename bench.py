#!/usr/bin/env python3
"""bench.py -- throughput of the cellularity x ploidy grid fit (baf.model.fit) on B200.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
                    [--workload config3|config2|config4|config5|mufreq_grid]

One "step" is one complete grid fit (R/bayes.R:106-116) over one synthetic segment
table.  Metric (BASELINE.json): grid-point x segment evaluations per second, fp64; one
evaluation = the full max-over-types posterior of one (cellularity, ploidy, segment).

Workloads
  config3 (default, every N; BASELINE configs[2], the north-star target): 1e5 segments,
      CNt.max = 7, grid 901 x 601, cut into units of 32 columns x 32 ploidies that are
      dealt to the N GPUs (strong scaling: total work fixed).  At N = 1 the line also
      carries `extra.config2` (BASELINE configs[1]: 3e4 segments, CNt.max = 20, 91 x 61,
      the single-GPU configuration) and `extra.worst_case` (config 3 with every sd x 0.1,
      where candidate pruning helps least).
  config2: BASELINE configs[1] as the headline (one GPU).
  config4: mufreq.bayes on 1e7 SNV positions (BASELINE configs[3]); metric positions/s.
  config5: cohort of 512 samples in one call (BASELINE configs[4]), sqz_baf_grid_batch
      with n_gpus = N from one process.
  mufreq_grid: mufreq.model.fit (R/bayes.R:175-194) on 1e4 positions x 91 x 61.
  segstats: segment.breaks' per-segment statistics (R/segments.R:113-198) over a synthetic
      genome of 2.4e7 seqz rows, 24 chromosomes, 2,400 segments; metric rows/s.

Arms
  ours       value: inputs resident in HBM, CUDA events on the launch stream, max over
             ranks; e2e: the same fit through the public API with HOST buffers (H2D of the
             segment table and D2H of result$LPP inside the timing).  For N > 1 also
             `e2e_single_process`: rank 0 alone calls sqz_baf_grid(n_gpus = N) on pageable
             host vectors -- the call the R shim makes.
  reference  the CPU restatement of the reference (oracle/, kind "port": R is not
             installed, so oracle/_ref cannot exist; with Rscript + sequenza on PATH the
             real package is timed instead, kind "reference-R") on all host cores, on a
             bounded sample of the same workload's grid points.  Rank 0 only.
"""
from __future__ import annotations

import argparse
import ctypes as C
import json
import os
import shutil
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = "grid-point x segment evaluations/sec (fp64), baf.model.fit"
PRI = {"CN": [2], "value": [2]}  # sequenza.fit default priors.table (R/fit.R:6)
KERNEL_NAMES = {0: "baf_grid_fast_kernel (fused)", 1: "baf_grid_split_kernel",
                2: "baf_grid_pruned_kernel", 3: "baf_grid_colwarp_kernel"}


def make_workload(name: str):
    from sequenza_b200.bayes import r_seq
    from sequenza_b200.synth import DEFAULT_SEED, baf_args, synth_segments
    if name == "config2":
        n, cn_max = 30000, 20
        pl, ce = r_seq(1, 7, 0.1), r_seq(0.1, 1, 0.01)
        desc = ("BASELINE configs[1]: 3e4 synthetic segments, CNt.max=20 (T=121), default "
                "grid 91 cellularities x 61 ploidies")
    elif name in ("config3", "worst_case"):
        n, cn_max = 100000, 7
        pl, ce = r_seq(1, 7, 0.01), r_seq(0.1, 1, 0.001)
        desc = ("BASELINE configs[2]: 1e5 synthetic segments, CNt.max=7 (T=20), grid 901 x "
                "601, split over the GPUs by units of 32 columns x 32 ploidies")
    else:
        raise SystemExit("unknown grid workload %s" % name)
    seg = synth_segments(n, cn_max=7, seed=DEFAULT_SEED)
    if name == "worst_case":
        seg["sd_ratio"] = seg["sd_ratio"] * 0.1
        seg["sd_Bf"] = seg["sd_Bf"] * 0.1
        desc = ("config 3 with sd.ratio and sd.BAF x 0.1 (sharp likelihood: the candidate "
                "pruning keeps the most copy numbers; x 10 is faster than x 1)")
    return {"name": name, "args": baf_args(seg), "ce": ce, "pl": pl, "cn_max": cn_max,
            "n": n, "scaling": "strong", "desc": desc,
            "n_c": cn_max + 1, "T": sum(c // 2 + 1 for c in range(cn_max + 1))}


def config_block(w, world):
    """The `config` object: identical in both arms (what is specific to an arm goes to the
    line's `config_detail`)."""
    return {"workload": w["desc"], "segments": w["n"], "CNt.max": w["cn_max"],
            "grid": [int(w["ce"].size), int(w["pl"].size)],
            "evals_per_step": float(w["ce"].size) * w["pl"].size * w["n"],
            "parallelism": "grid units x%d" % world,
            "l2": "256 MiB buffer written between timed iterations of the GPU arm (L2 flush)"}


# ---- clocks during the timed region (NVML, polled from a thread) ----------------------
class ClockSampler:
    REASONS = {0x4: "sw_power_cap", 0x8: "hw_slowdown", 0x20: "sw_thermal_slowdown",
               0x40: "hw_thermal_slowdown", 0x80: "hw_power_brake_slowdown",
               0x2: "applications_clocks_setting", 0x10: "sync_boost"}

    def __init__(self, cuda_index: int):
        self.samples, self.stop, self.thread, self.h, self.nv = [], False, None, None, None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            try:
                import torch
                uuid = str(torch.cuda.get_device_properties(cuda_index).uuid)
                if not uuid.startswith("GPU-"):
                    uuid = "GPU-" + uuid
                self.h = pynvml.nvmlDeviceGetHandleByUUID(uuid.encode())
            except Exception:
                self.h = pynvml.nvmlDeviceGetHandleByIndex(cuda_index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception as e:  # no NVML: report that instead of inventing clocks
            self.err = repr(e)

    def _loop(self):
        nv = self.nv
        while not self.stop:
            try:
                mhz = nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM)
                try:
                    rs = nv.nvmlDeviceGetCurrentClocksEventReasons(self.h)
                except Exception:
                    rs = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                self.samples.append((time.perf_counter(), mhz, rs))
            except Exception:
                pass
            time.sleep(0.002)

    def start(self):
        if self.h is not None:
            self.thread = threading.Thread(target=self._loop, daemon=True)
            self.thread.start()

    def finish(self, t0, t1):
        self.stop = True
        if self.thread:
            self.thread.join(timeout=2)
        if self.h is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [],
                    "note": "NVML unavailable: %s" % getattr(self, "err", "?")}
        inside = [s for s in self.samples if t0 <= s[0] <= t1]
        note = None
        if not inside:  # timed region shorter than one poll: take the closest samples
            inside = sorted(self.samples, key=lambda s: abs(s[0] - 0.5 * (t0 + t1)))[:3]
            note = "timed region shorter than the NVML poll period; nearest samples used"
        mask = 0
        for s in inside:
            mask |= s[2]
        out = {"sm_mhz": float(np.median([s[1] for s in inside])) if inside else None,
               "sm_max_mhz": float(self.max_mhz), "samples": len(inside),
               "reasons": [v for k, v in self.REASONS.items() if mask & k]}
        if note:
            out["note"] = note
        return out


def ncu_traffic(workload, kernel_path):
    """dram__bytes_read.sum + dram__bytes_write.sum of the dominant kernel, per launch, from
    the committed `ncu --set full` capture of this workload (profiles/r02_traffic.json: it
    names the kernel and the git commit it was captured at; a capture of another kernel
    path is not reported)."""
    try:
        with open(os.path.join(ROOT, "profiles", "r02_traffic.json")) as f:
            d = json.load(f)
        e = d[workload]
        if int(e["kernel_path"]) != int(kernel_path):
            return None, "capture in profiles/r02_traffic.json is of another kernel path"
        return e["dram_bytes_per_launch"], "ncu capture at commit %s (%s)" % (
            e.get("commit", "?"), e.get("kernel", "?"))
    except Exception:
        return None, "no capture committed for this workload"


def ncu_instructions(workload, kernel_path):
    """Warp instructions one launch of the dominant kernel executes (same capture as
    ncu_traffic), or None."""
    try:
        with open(os.path.join(ROOT, "profiles", "r02_traffic.json")) as f:
            e = json.load(f)[workload]
        if int(e["kernel_path"]) != int(kernel_path):
            return None
        return float(e["warp_instructions_per_launch"])
    except Exception:
        return None


def cpu_baseline(w, steps=1, warmup=0):
    """The oracle (CPU restatement, all host threads) on a bounded contiguous block of
    this workload's grid points.  Returns (evals/s, description dict, seconds per step)."""
    from oracle import oracle as O
    cores = os.cpu_count() or 1
    G = w["ce"].size * w["pl"].size
    per_point = w["n"] * (w["n_c"] + w["T"]) * 5.5e-8  # ~55 ns per dt() call, one thread
    pts = int(max(cores, min(G, round(12.0 * cores / max(per_point, 1e-9) / 8)))) \
        if per_point > 0 else cores
    pts = max(cores, min(pts, 64 * cores, G))
    g0 = max(0, G // 2 - pts // 2)
    kw = dict(cellularity=w["ce"], ploidy=w["pl"], CNt_max=w["cn_max"], priors_table=PRI,
              g0=g0, g1=g0 + pts, nthreads=cores, **w["args"])
    for _ in range(warmup):
        O.baf_grid_sums(**kw)
    t0 = time.perf_counter()
    for _ in range(steps):
        O.baf_grid_sums(**kw)
    dt = (time.perf_counter() - t0) / steps
    evals = pts * w["n"]
    desc = {"value": evals / dt, "unit": "evals/s", "cores": cores, "kind": "port",
            "sample": "%d contiguous grid points (from index %d of %d) x %d segments of the "
                      "same workload, %.1f s per pass; oracle/sqz_oracle.c, gcc -O2, "
                      "pthreads" % (pts, g0, G, w["n"], dt)}
    return evals / dt, desc, dt


def r_baseline(w, steps, warmup):
    """The UNMODIFIED R package, when Rscript and sequenza are installed (they are not in
    this image; tools/r_reference_dump.R is the recipe): baf.model.fit on a block of
    grid points with mc.cores = all cores.  Returns None when R is unavailable."""
    rscript = shutil.which("Rscript")
    if not rscript:
        return None
    probe = subprocess.run([rscript, "-e", "suppressMessages(library(sequenza))"],
                           capture_output=True, text=True)
    if probe.returncode != 0:
        return None
    import tempfile
    cores = os.cpu_count() or 1
    d = tempfile.mkdtemp(prefix="sqz_r_")
    tsv = os.path.join(d, "segments.tsv")
    keys = ("Bf", "depth_ratio", "sd_Bf", "sd_ratio", "weight_Bf", "weight_ratio")
    cols = [np.broadcast_to(np.asarray(w["args"][k], dtype=float), (w["n"],)) for k in keys]
    np.savetxt(tsv, np.column_stack(cols), delimiter="\t", header="\t".join(keys),
               comments="", fmt="%.17g")
    # a block of grid points: `cores` ploidies x 1 cellularity around the truth; R needs
    # seconds per grid point at this N
    n_pl = max(2, min(cores, 16))
    pl = w["pl"][w["pl"].size // 2:w["pl"].size // 2 + n_pl]
    ce = w["ce"][w["ce"].size // 2:w["ce"].size // 2 + 1]
    script = os.path.join(ROOT, "tools", "r_reference_dump.R")
    cmd = [rscript, script, "--time-grid", tsv, ",".join("%.17g" % v for v in ce),
           ",".join("%.17g" % v for v in pl), str(w["cn_max"]), str(cores),
           str(max(1, steps)), str(warmup)]
    out = subprocess.run(cmd, capture_output=True, text=True)
    if out.returncode != 0:
        return None
    try:
        secs = float(out.stdout.strip().splitlines()[-1])
    except Exception:
        return None
    evals = float(ce.size * pl.size * w["n"])
    desc = {"value": evals / secs, "unit": "evals/s", "cores": cores, "kind": "reference-R",
            "sample": "%d grid points x %d segments through sequenza::baf.model.fit("
                      "mc.cores = %d), %.1f s per pass" % (ce.size * pl.size, w["n"], cores,
                                                           secs)}
    return evals / secs, desc, secs


def run_reference(args, rank, world):
    if rank != 0:
        return
    if args.workload not in ("config2", "config3"):
        print(json.dumps({"impl": "reference", "unavailable":
                          "the reference arm times the grid fit (config2 / config3) only"}))
        return
    w = make_workload(args.workload)
    got = r_baseline(w, max(1, args.steps), args.warmup)
    if got is None:
        got = cpu_baseline(w, steps=max(1, args.steps), warmup=args.warmup)
    val, desc, dt = got
    line = {"impl": "reference", "metric": METRIC, "value": val, "unit": "evals/s",
            "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": dt * 1e3, "higher_is_better": True, "scaling": w["scaling"],
            "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": config_block(w, args.gpus),
            "config_detail": {
                "note": "reference R path cannot run here (R not installed); this is the "
                        "compiled CPU restatement on the host cores" if desc["kind"] == "port"
                else "the unmodified R package on the host cores"},
            "cpu_baseline": desc,
            "e2e": {"value": val, "unit": "evals/s", "h2d_bytes_per_step": 0,
                    "d2h_bytes_per_step": 0}}
    print(json.dumps(line), flush=True)


def roofline_model(w, plan, kpath, n_cand, ncall, cols, nP, n):
    """Executed work per evaluation of the kernel path that ran.
    Returns (candidates per (segment, tile), fp64-pipe instructions, math-dispatch slots,
    fp64 flops, fp32 flops) per evaluation.  Per evaluated (copy number, ploidy): 2 DFMA (+ DSETP and 2 FSEL
    from the second candidate on); per ploidy 1 DMUL into the running product; the
    per-column pre-pass is shared by all nP ploidies of a column."""
    PT, nPT = plan["PT"], plan["nPT"]
    n_c, T = w["n_c"], w["T"]
    pre = (2.0 * T + 14.0 * n_c) / nP
    f32 = 0.0
    if kpath == 3 and ncall > 0:
        # column-per-lane kernel, per (segment, tile): the bounds (per copy number 7 fp64 + 4
        # integer instructions for CNt 0..20; 6 fp32 + 6 integer ones, off the fp64 pipe, for
        # CNt 0..7) and ~10 fp64 + ~10 others to set up the first candidate; the first
        # candidate costs 2 DFMA + DMUL per ploidy (no minimum), every further one that the
        # warp evaluates 2 DFMA + DSETP + 2 FSEL
        cand = n_cand / (float(ncall) * n * cols * nPT)
        if n_c <= 8:
            inst = 3.0 + (cand - 1.0) * 3.0 + 10.0 / PT + pre
            math = inst + (cand - 1.0) * 2.0 + (12.0 * n_c + 10.0) / PT
            flop = 5.0 + (cand - 1.0) * 5.0 + 14.0 / PT + 1.5 * pre
            f32 = 8.0 * n_c / PT
        else:
            inst = 3.0 + (cand - 1.0) * 3.0 + (7.0 * n_c + 10.0) / PT + pre
            math = inst + (cand - 1.0) * 2.0 + (4.0 * n_c + 10.0) / PT
            flop = 5.0 + (cand - 1.0) * 5.0 + (9.0 * n_c + 14.0) / PT + 1.5 * pre
    elif kpath == 2 and ncall > 0:
        cand = n_cand / (float(ncall) * n * cols * nPT)
        inst = cand * (3.0 + 2.0 / PT) + 1 + 6.0 * n_c / PT + pre
        math = inst + cand * 2.0 + 4.0 * n_c / PT
        flop = cand * (5.0 + 2.0 / PT) + 1 + 8.0 * n_c / PT + 1.5 * pre
    else:
        cand = float(n_c)
        inst = 3.0 * n_c + 1 + 3.0 * n_c / PT + pre         # fp64-pipe issue slots
        math = inst + 2.0 * n_c                             # + the FSEL pair of each minimum
        flop = 5.0 * n_c + 1 + 3.0 * n_c / PT + 1.5 * pre   # DFMA = 2 flops, others 1
    return cand, inst, math, flop, f32


def measure_grid(w, env, steps, warmup, with_e2e=True):
    """Times `steps` fits of workload w on this rank's share of the grid.  env: torch, dist,
    L, device, world, rank, flush, host_group."""
    torch, dist, L = env["torch"], env["dist"], env["L"]
    import sequenza_b200 as S
    from sequenza_b200 import _lib
    from sequenza_b200.dist import BafGridDevice, DistributedGridFit
    device, world, rank = env["device"], env["world"], env["rank"]
    ce, pl = w["ce"], w["pl"]
    nC, nP, n = ce.size, pl.size, w["n"]
    evals_total = float(nC) * nP * n

    dev = BafGridDevice(device=device, **w["args"])
    fit = DistributedGridFit(dev, pl, ce, peer=os.environ.get("SQZ_NO_PEER", "") != "1",
                             CNt_max=w["cn_max"], priors_table=PRI)
    plan = fit.plan
    flush = env["flush"]

    def barrier():
        if world > 1:
            dist.barrier()

    for _ in range(max(warmup, 3)):
        res = fit.step()
    torch.cuda.synchronize()
    L.sqz_profile_enable(1)
    ev0 = [torch.cuda.Event(enable_timing=True) for _ in range(steps)]
    ev1 = [torch.cuda.Event(enable_timing=True) for _ in range(steps)]
    barrier()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for k in range(steps):
        flush.zero_()  # L2 flush between timed iterations (outside the event pair)
        ev0[k].record()
        res = fit.step()
        ev1[k].record()
    torch.cuda.synchronize()
    barrier()
    t1 = time.perf_counter()
    ms3 = (C.c_double * 3)()
    ncall, kpath = C.c_int(0), C.c_int(0)
    n_cand, n_launch = C.c_ulonglong(0), C.c_longlong(0)
    L.sqz_profile_counters(C.byref(kpath), C.byref(n_cand), C.byref(n_launch))
    L.sqz_profile_collect(ms3, C.byref(ncall))
    L.sqz_profile_enable(0)
    step_ms = sum(a.elapsed_time(b) for a, b in zip(ev0, ev1))
    tm = torch.tensor([step_ms], dtype=torch.float64, device=device)
    if world > 1:
        dist.all_reduce(tm, op=dist.ReduceOp.MAX)
    ms_per_step = float(tm.item()) / steps
    lpp_sum = res[0].cpu().numpy().copy()
    # this rank's own phases, and everybody's (the step ends when the slowest rank is done)
    mine = torch.tensor([ms3[0], ms3[1], ms3[2], step_ms], dtype=torch.float64,
                        device=device) / max(ncall.value, 1)
    if world > 1:
        allr = [torch.zeros_like(mine) for _ in range(world)]
        dist.all_gather(allr, mine)
    else:
        allr = [mine]
    per_rank = {k: [round(float(t[i]), 4) for t in allr]
                for i, k in enumerate(("prep_ms", "grid_kernels_ms", "finalize_ms", "step_ms"))}
    out = {"plan": plan, "t0": t0, "t1": t1, "ms_per_step": ms_per_step,
           "value": evals_total / (ms_per_step * 1e-3), "lpp_sum": lpp_sum,
           "kpath": kpath.value, "n_cand": n_cand.value, "ncall": ncall.value,
           "ms3": [ms3[0], ms3[1], ms3[2]], "collective": fit.collective,
           "units": (fit.u0, fit.u1), "per_rank": per_rank,
           # kernels launched inside the timed region: the library's own count (preparation,
           # pre-pass, grid kernel, finalize, repair per fit) + the L2-flush fill + the two
           # kernels of the assembly (flag store + wait, or fill + NCCL all-reduce)
           "launches": int(n_launch.value) + steps * (1 + (2 if world > 1 else 0))}

    if with_e2e:
        # ---- end to end through the public API, host buffers, copies inside the timing ----
        if world == 1:
            pinned = [torch.from_numpy(_lib.vec(w["args"][k], n)).pin_memory()
                      for k in ("Bf", "depth_ratio", "sd_Bf", "sd_ratio", "weight_Bf",
                                "weight_ratio")]
            hout = torch.empty(nC * nP, dtype=torch.float64).pin_memory()
            pc, pv = np.array([2.0]), np.array([2.0])

            def e2e_step():
                _lib.check(L.sqz_baf_grid(*[t.data_ptr() for t in pinned], n, _lib.ptr(pl), nP,
                                          _lib.ptr(ce), nC, 1.0, 0, w["cn_max"], 2,
                                          _lib.ptr(pc), _lib.ptr(pv), 1, 0, 1, S.SQZ_MODE_FAST,
                                          hout.data_ptr()))
            h2d = 6 * n * 8 + 16 + 8 + (nC + nP) * 8
            d2h = nC * nP * 8
            api = "sqz_baf_grid (host pointers)"
        else:
            hout = torch.empty((1, nC * nP), dtype=torch.float64).pin_memory()

            def e2e_step():
                dev.upload()
                full = fit.step()
                hout.copy_(full, non_blocking=True)
                torch.cuda.synchronize()
            h2d = world * (6 * n * 8 + 16 + 8 + (nP + nC) * 8)
            d2h = world * nC * nP * 8
            api = ("BafGridDevice.upload + sqz_baf_grid_dev_units + " + fit.collective +
                   " + D2H on every rank")
        for _ in range(3):
            e2e_step()
        barrier()
        torch.cuda.synchronize()
        te = time.perf_counter()
        for _ in range(steps):
            e2e_step()
        torch.cuda.synchronize()
        barrier()
        e2e_s = time.perf_counter() - te
        tm = torch.tensor([e2e_s], dtype=torch.float64, device=device)
        if world > 1:
            dist.all_reduce(tm, op=dist.ReduceOp.MAX)
        e2e_ms = float(tm.item()) * 1e3 / steps
        assert np.array_equal(hout.numpy().reshape(-1), lpp_sum), \
            "e2e and resident results differ"
        out["e2e"] = {"value": evals_total / (e2e_ms * 1e-3), "unit": "evals/s",
                      "ms_per_step": e2e_ms, "h2d_bytes_per_step": h2d,
                      "d2h_bytes_per_step": d2h, "api": api}
    fit.close()
    return out


def single_process_e2e(w, env, steps):
    """Rank 0 alone drives all N GPUs through sqz_baf_grid(n_gpus = N) with pageable host
    vectors (what the R shim passes); the other ranks wait on a host-side barrier."""
    torch, dist, L = env["torch"], env["dist"], env["L"]
    import sequenza_b200 as S
    from sequenza_b200 import _lib
    world, rank = env["world"], env["rank"]
    out = None
    dist.barrier(group=env["host_group"])
    if rank == 0:
        ce, pl, n = w["ce"], w["pl"], w["n"]
        v = [_lib.vec(w["args"][k], n) for k in ("Bf", "depth_ratio", "sd_Bf", "sd_ratio",
                                                 "weight_Bf", "weight_ratio")]
        res = np.empty(ce.size * pl.size)
        pc, pv = np.array([2.0]), np.array([2.0])

        def call():
            _lib.check(L.sqz_baf_grid(*[_lib.ptr(a) for a in v], n, _lib.ptr(pl), pl.size,
                                      _lib.ptr(ce), ce.size, 1.0, 0, w["cn_max"], 2,
                                      _lib.ptr(pc), _lib.ptr(pv), 1, 0, world,
                                      S.SQZ_MODE_FAST, _lib.ptr(res)))
        for _ in range(3):
            call()
        t = time.perf_counter()
        for _ in range(steps):
            call()
        ms = (time.perf_counter() - t) * 1e3 / steps
        out = {"value": float(ce.size) * pl.size * n / (ms * 1e-3), "unit": "evals/s",
               "ms_per_step": ms, "api": "sqz_baf_grid(n_gpus=%d), one process, pageable "
               "host vectors in and out" % world,
               "h2d_bytes_per_step": 6 * n * 8 + 16 + 8 + world * (ce.size + pl.size) * 8,
               "d2h_bytes_per_step": ce.size * pl.size * 8, "_res": res}
        L.sqz_release()
    dist.barrier(group=env["host_group"])
    return out


def roofline_block(w, m, peak, world, local_rank):
    """The `roofline` object of one measured workload (m = measure_grid's result)."""
    plan, kpath, ncall = m["plan"], m["kpath"], m["ncall"]
    nC, nP, n = w["ce"].size, w["pl"].size, w["n"]
    n_c, T = w["n_c"], w["T"]
    PT, nPT = plan["PT"], plan["nPT"]
    u0, u1 = m["units"]
    # grid points of this rank: units x W x PT, minus the ragged edges (reported per rank 0)
    from sequenza_b200.dist import unit_points
    local_points = float(unit_points(plan, nC, nP, u0, u1).sum())
    cols = local_points / nP  # column equivalents
    cand, inst, math, flop, f32 = roofline_model(w, plan, kpath, m["n_cand"], ncall, cols, nP,
                                                 n)
    fast_ms = m["ms3"][1] / max(ncall, 1)
    local_evals = local_points * n
    padded = float(nPT * PT) / nP  # ragged last ploidy tile is computed and discarded
    ach = local_evals * flop / (fast_ms * 1e-3) / 1e12 if fast_ms > 0 else None
    ref_flops = 11.0 * n_c + 13.0 * T  # reference formulation
    ref_logs = n_c + T
    traffic, traffic_src = ncu_traffic(w["name"], kpath) if world == 1 else (None, "N > 1")
    winst = ncu_instructions(w["name"], kpath) if world == 1 else None
    # what bounds the column-per-lane kernel since its bounds moved to fp32: instruction issue
    # (one warp instruction per scheduler and cycle; the pre-pass is included in kernel_ms)
    sm_hz = 1e6 * float(m.get("sm_mhz") or 1965.0)
    issue = {"warp_instructions_per_launch": winst,
             "achieved_per_scheduler_cycle": winst / (fast_ms * 1e-3 * 148 * 4 * sm_hz),
             "peak": 1.0, "source": traffic_src,
             "note": "grid kernel alone, ncu at the capture: smsp__issue_active in "
                     "profiles/r02_ncu_summary.md"} if winst and fast_ms > 0 else None
    return {
        "bound": "fp64",
        "kernel": "%s<%d> (+ baf_colterm%s pre-pass)" % (
            KERNEL_NAMES.get(kpath, "?"), PT, "_cg" if kpath == 3 else ""),
        "copy_numbers_evaluated_per_segment_and_tile": cand, "copy_numbers": n_c,
        "achieved": ach, "peak": peak, "unit": "TFLOP/s",
        "frac": (ach / peak) if ach and peak else None,
        "peak_source": "measured live: sqz_fp64_peak, dependent-free DFMA chains on all "
                       "SMs (MEASURED_PEAKS.json carries no fp64 figure)",
        "kernel_ms": fast_ms,
        "kernel_share_of_step": fast_ms / m["ms_per_step"] if m["ms_per_step"] else None,
        "fp64_issue_slots_per_eval": inst, "flops_per_eval_executed": flop,
        "fp32_flops_per_eval_executed": f32,
        "frac_note": "fp64 flops only: the candidate bounds of the CNt 0..7 kernel run in fp32 "
                     "since round 2 (fp64 flops per evaluation 9.88 -> 7.63), so the fraction "
                     "fell while the fit got faster; `issue` is the resource that binds",
        "issue": issue,
        "pipe_util": (local_evals * inst / (fast_ms * 1e-3)) / (peak * 1e12 / 2)
        if fast_ms > 0 and peak else None,
        "math_issue_slots_per_eval": math,
        "tile_padding": padded,
        "reference_formulation": {
            "flops_per_eval": ref_flops, "logs_per_eval": ref_logs,
            "tflops": local_evals * ref_flops / (fast_ms * 1e-3) / 1e12 if fast_ms > 0 else None,
            "tlogs_per_s": local_evals * ref_logs / (fast_ms * 1e-3) / 1e12 if fast_ms > 0 else None,
            "note": "the kernel reaches the same result with %.0fx fewer fp64 flops and "
                    "no per-type log (exact factorisation, DESIGN.md)" % (ref_flops / flop)},
        "traffic": traffic, "traffic_source": traffic_src,
        "algorithmic_hbm_bytes": 48 * n + 8 * int(local_points),
        "traffic_note": "the grid kernel reads the pre-pass scratch (8 B x copy numbers x "
                        "segments per column) once per ploidy tile, mostly from L2; that "
                        "scratch replaces recomputing the BAF term in every ploidy tile",
    }


def run_ours(args, rank, world, local_rank):
    import torch
    import torch.distributed as dist
    import sequenza_b200 as S

    if not torch.cuda.is_available() or S.lib().sqz_device_count() < 1:
        raise SystemExit("bench.py --impl ours needs a CUDA device (no CPU fallback)")
    torch.cuda.set_device(local_rank)
    device = torch.device("cuda", local_rank)
    L = S.lib()
    env = {"torch": torch, "dist": dist, "L": L, "device": device, "world": world,
           "rank": rank, "host_group": None,
           "flush": torch.empty(256 << 20, dtype=torch.uint8, device=device)}  # > 126 MB L2
    if world > 1:
        env["host_group"] = dist.new_group(backend="gloo")
    if args.workload in ("config4", "config5", "mufreq_grid", "segstats"):
        return run_other(args, env, local_rank)
    w = make_workload(args.workload)
    sampler = ClockSampler(local_rank)
    sampler.start()
    m = measure_grid(w, env, args.steps, args.warmup)
    clocks = sampler.finish(m["t0"], m["t1"])
    sp = single_process_e2e(w, env, min(args.steps, 10)) if world > 1 else None
    if rank != 0:
        return
    peak = C.c_double(0)
    L.sqz_fp64_peak(local_rank, 3, C.byref(peak))
    ce, pl = w["ce"], w["pl"]
    nP = pl.size
    lpp_sum = m["lpp_sum"]
    k = int(np.argmax(lpp_sum))
    plan = m["plan"]
    line = {
        "metric": METRIC, "value": m["value"], "unit": "evals/s", "n_gpus": world,
        "steps": args.steps, "warmup": max(args.warmup, 3), "ms_per_step": m["ms_per_step"],
        "higher_is_better": True, "scaling": w["scaling"], "vs_baseline": None,
        "dtype": "f64", "data": "synthetic",
        "config": config_block(w, world),
        "config_detail": {
            "mode": "fast + faithful repair", "ploidy_tile": plan["PT"],
            "segment_chunks": plan["chunks"], "threads_per_cta": plan["threads"],
            "units": plan["units"], "unit_columns": plan["W"], "assembly": m["collective"]},
        "clocks": clocks,
        "e2e": m["e2e"],
        "gpu_launches": m["launches"],
        "gpu_launches_note": "counted at the launch sites inside the library for the timed "
                             "fits (sqz_profile_counters), plus per step the L2-flush fill "
                             "and, for N > 1, the two kernels of the assembly",
        "roofline": roofline_block(w, m, peak.value, world, local_rank),
        "per_rank": m["per_rank"],
        "fit_wall_ms": m["e2e"]["ms_per_step"],
        "solution": {"cellularity": float(ce[k // nP]), "ploidy": float(pl[k % nP])},
    }
    if sp is not None:
        same = bool(np.array_equal(sp.pop("_res"), lpp_sum))
        sp["identical_to_per_rank_result"] = same
        line["e2e_single_process"] = sp
    if world == 1:
        _, cpu_desc, _ = cpu_baseline(w)
        line["cpu_baseline"] = cpu_desc
        extra = {}
        for name in (("config2", "worst_case") if w["name"] == "config3" else ()):
            w2 = make_workload(name)
            m2 = measure_grid(w2, env, max(5, min(args.steps, 50)), 3,
                              with_e2e=(name == "config2"))
            k2 = int(np.argmax(m2["lpp_sum"]))
            e = {"workload": w2["desc"], "value": m2["value"], "unit": "evals/s",
                 "ms_per_step": m2["ms_per_step"],
                 "roofline": roofline_block(w2, m2, peak.value, 1, local_rank),
                 "solution": {"cellularity": float(w2["ce"][k2 // w2["pl"].size]),
                              "ploidy": float(w2["pl"][k2 % w2["pl"].size])}}
            if "e2e" in m2:
                e["e2e"] = m2["e2e"]
            if name == "config2":
                _, d2, _ = cpu_baseline(w2)
                e["cpu_baseline"] = d2
            extra[name] = e
        if extra:
            line["extra"] = extra
    print(json.dumps(line), flush=True)


def hbm_peak():
    """Measured HBM copy bandwidth of this pool's B200s (GB/s) and where it came from."""
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "MEASURED_PEAKS.json"
    except Exception:
        return 6650.0, "fallback of /opt/skills/guides/B200_PROFILING.md (6.65 TB/s)"


def timed_ms(torch, fn, steps, warm, flush=None):
    for _ in range(max(warm, 3)):
        fn()
    torch.cuda.synchronize()
    tot = 0.0
    for _ in range(steps):
        if flush is not None:
            flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        tot += e0.elapsed_time(e1)
    return tot / steps


def run_other(args, env, local_rank):
    """config4 (mufreq.bayes, 1e7 positions), config5 (cohort) and mufreq_grid
    (mufreq.model.fit): single-process workloads; under torchrun rank 0 runs them (config5
    with n_gpus = N) while the other ranks wait."""
    torch, dist, L = env["torch"], env["dist"], env["L"]
    import sequenza_b200 as S
    from sequenza_b200 import _lib
    from sequenza_b200.synth import baf_args, synth_mutations, synth_segments
    world, rank, device = env["world"], env["rank"], env["device"]
    if world > 1:
        dist.barrier(group=env["host_group"])
    n_dev = max(world, args.gpus)  # plain `python bench.py --gpus N`: N devices, one process
    if rank == 0:
        sampler = ClockSampler(local_rank)
        sampler.start()
        t0 = time.perf_counter()
        stream = torch.cuda.current_stream().cuda_stream
        steps = max(3, args.steps)
        base = {"n_gpus": n_dev, "steps": steps, "warmup": max(args.warmup, 3),
                "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
                "dtype": "f64", "data": "synthetic"}
        if args.workload == "config4":
            from oracle import oracle as O
            n = 10_000_000
            mut = synth_mutations(n)
            f = torch.from_numpy(mut["mufreq"]).to(device)
            r = torch.from_numpy(mut["depth_ratio"]).to(device)
            out = torch.empty(4 * n, dtype=torch.float64, device=device)
            peak, peak_src = hbm_peak()
            res = {}
            for cn_max in (20, 7):
                tb = L.sqz_mufreq_scalar_table_bytes(100.0, 100.0)
                table = torch.empty(tb + 64, dtype=torch.uint8, device=device)
                tptr = (table.data_ptr() + 31) & ~31
                ms = timed_ms(torch, lambda: _lib.check(L.sqz_mufreq_point_scalar_dev(
                    f.data_ptr(), r.data_ptr(), n, 100.0, 100.0, 0.7, 2.4, 1.0, 1, cn_max, 2,
                    None, None, 0, tptr, tb, stream, out.data_ptr())), steps, args.warmup,
                    env["flush"])
                wv = torch.full((n,), 100.0, dtype=torch.float64, device=device)
                ms_pt = timed_ms(torch, lambda: _lib.check(L.sqz_mufreq_point_dev(
                    f.data_ptr(), r.data_ptr(), wv.data_ptr(), wv.data_ptr(), n, 0.7, 2.4, 1.0,
                    1, cn_max, 2, None, None, 0, stream, out.data_ptr())), 3, 1)
                res[cn_max] = (ms, ms_pt)
            te = time.perf_counter()
            host = S.mufreq_bayes(mut["mufreq"], mut["depth_ratio"], 0.7, 2.4, 1.0, CNt_max=20)
            te = time.perf_counter() - te
            # the same call on PINNED host vectors handed to the _dev entry point as they are
            # (unified addressing: the kernel reads the inputs and writes the N x 4 result
            # over PCIe, no staging copy) -- what SURVEY 8(e) suggests for the 320 MB result
            fp = torch.from_numpy(mut["mufreq"]).pin_memory()
            rp = torch.from_numpy(mut["depth_ratio"]).pin_memory()
            op = torch.empty(4 * n, dtype=torch.float64).pin_memory()
            ms_pin = timed_ms(torch, lambda: _lib.check(L.sqz_mufreq_point_scalar_dev(
                fp.data_ptr(), rp.data_ptr(), n, 100.0, 100.0, 0.7, 2.4, 1.0, 1, 20, 2,
                None, None, 0, tptr, tb, stream, op.data_ptr())), 5, 2)
            torch.cuda.synchronize()
            pin_ok = bool(np.array_equal(op.numpy().reshape(4, n).T[:, :3], host[:, :3],
                                         equal_nan=True))
            ns = 200_000  # CPU sample: ~10-20 s single thread at T = 230
            tc = time.perf_counter()
            ref = O.mufreq_bayes(mut["mufreq"][:ns], mut["depth_ratio"][:ns], 0.7, 2.4, 1.0,
                                 CNt_max=20)
            tc = time.perf_counter() - tc
            assert np.array_equal(host[:ns, :3], ref[:, :3])
            assert np.max(np.abs(host[:ns, 3] - ref[:, 3]) / np.abs(ref[:, 3])) < 1e-9
            ms = res[20][0]
            line = dict(base, metric="mufreq.bayes positions/sec (fp64), scalar weights 100/100 "
                        "as sequenza.results calls it", value=n / (ms * 1e-3), unit="positions/s",
                        ms_per_step=ms,
                        config={"workload": "BASELINE configs[3]: mufreq.bayes on 1e7 synthetic "
                                "SNV positions at the fitted solution (0.7, 2.4), CNt.max=20 "
                                "(T=230)", "positions": n,
                                "l2": "256 MiB buffer written between timed iterations"},
                        e2e={"value": n / te, "unit": "positions/s", "ms_per_step": te * 1e3,
                             "h2d_bytes_per_step": 16 * n, "d2h_bytes_per_step": 32 * n,
                             "api": "sequenza_b200.mufreq_bayes -> sqz_mufreq_point_scalar "
                                    "(pageable numpy arrays in and out)"},
                        gpu_launches=3 * steps,
                        e2e_pinned_zero_copy={
                            "value": n / (ms_pin * 1e-3), "unit": "positions/s",
                            "ms_per_step": ms_pin, "equals_host_api_result": pin_ok,
                            "api": "sqz_mufreq_point_scalar_dev on pinned host vectors (inputs "
                                   "read and result written over PCIe by the kernel)"},
                        roofline={"bound": "hbm",
                                  "kernel": "mufreq_lookup_smem_kernel (+ scores and table "
                                            "kernels: two small launches)",
                                  "achieved": 48.0 * n / (ms * 1e-3) / 1e9, "peak": peak,
                                  "unit": "GB/s", "frac": 48.0 * n / (ms * 1e-3) / 1e9 / peak,
                                  "peak_source": peak_src, "traffic": None,
                                  "bytes_per_position": 48},
                        cpu_baseline={"value": ns / tc, "unit": "positions/s", "cores": 1,
                                      "kind": "port", "sample": "first %d positions, oracle "
                                      "mufreq_bayes (single thread), %.1f s" % (ns, tc)},
                        extra={"CNt.max=7 (T=35) table path ms": res[7][0],
                               "per-position weights (mufreq_point_kernel) ms, T=230": res[20][1],
                               "per-position weights ms, T=35": res[7][1]})
        elif args.workload == "config5":
            from oracle import oracle as O
            S_n, n_seg = 512, 3000
            ce, pl = S.r_seq(0.1, 1, 0.01), S.r_seq(1, 7, 0.1)
            samples = [baf_args(synth_segments(n_seg, cellularity=0.3 + 0.6 * (k % 50) / 50,
                                               ploidy=1.5 + 3.0 * (k % 37) / 37, seed=1000 + k))
                       for k in range(S_n)]
            keys = ("Bf", "depth_ratio", "sd_Bf", "sd_ratio", "weight_Bf", "weight_ratio")
            cat = {k: np.concatenate([s[k] for s in samples]) for k in keys}
            offs = np.arange(S_n + 1, dtype=np.int64) * n_seg
            adr = np.array([s["avg_depth_ratio"] for s in samples])
            evals = float(S_n) * n_seg * ce.size * pl.size
            out = np.empty(S_n * ce.size * pl.size)
            pc, pv = np.array([2.0]), np.array([2.0])

            def call():
                _lib.check(L.sqz_baf_grid_batch(
                    S_n, offs.ctypes.data, *[_lib.ptr(cat[k]) for k in keys], _lib.ptr(pl),
                    pl.size, _lib.ptr(ce), ce.size, _lib.ptr(adr), 0, 7, 2, _lib.ptr(pc),
                    _lib.ptr(pv), 1, 0, n_dev, 0, _lib.ptr(out)))
            for _ in range(3):
                call()
            t = time.perf_counter()
            for _ in range(steps):
                call()
            ms = (time.perf_counter() - t) * 1e3 / steps
            tc = time.perf_counter()
            ref = O.baf_grid_sums(cellularity=ce, ploidy=pl, priors_table=PRI, g0=2000,
                                  g1=2000 + 64 * (os.cpu_count() or 1),
                                  nthreads=os.cpu_count() or 1, **samples[5])
            tc = time.perf_counter() - tc
            G = ce.size * pl.size
            g = out[5 * G + 2000:5 * G + 2000 + ref.size]
            assert np.max(np.abs(g - ref) / np.abs(ref)) < 1e-9
            line = dict(base, metric=METRIC + " (cohort)", value=evals / (ms * 1e-3),
                        unit="evals/s", ms_per_step=ms,
                        config={"workload": "BASELINE configs[4]: cohort of 512 synthetic "
                                "samples x 3000 segments, default grid 91 x 61, CNt.max=7, one "
                                "sqz_baf_grid_batch call, samples dealt to the GPUs",
                                "samples": S_n, "segments": S_n * n_seg, "grid": [91, 61]},
                        e2e={"value": evals / (ms * 1e-3), "unit": "evals/s", "ms_per_step": ms,
                             "h2d_bytes_per_step": 48 * S_n * n_seg, "d2h_bytes_per_step": 8 * S_n * G,
                             "api": "sqz_baf_grid_batch(n_gpus=%d), pageable host vectors" % n_dev},
                        value_note="timed through the host API (there is no resident-input "
                                   "form of the cohort call in this bench): value == e2e",
                        gpu_launches=5 * steps * n_dev,
                        cpu_baseline={"value": ref.size * n_seg / tc, "unit": "evals/s",
                                      "cores": os.cpu_count() or 1, "kind": "port",
                                      "sample": "%d grid points of one sample, %.1f s" % (ref.size, tc)})
        elif args.workload == "segstats":
            from oracle import oracle as O
            from sequenza_b200 import segments as G
            g = G.synth_seqz(n_chr=24, rows_per_chr=1_000_000, segs_per_chr=100, seed=11)
            n_rows, n_seg = int(g["row_off"][-1]), int(g["seg_off"][-1])
            keys = ("position", "adjusted_ratio", "depth_normal", "good_reads", "Af", "Bf")
            dv = [torch.from_numpy(g[k]).to(device) for k in keys]
            dhet = torch.from_numpy(g["is_het"]).to(device)
            dro = torch.from_numpy(g["row_off"]).to(device)
            dso = torch.from_numpy(g["seg_off"]).to(device)
            dbrk = torch.from_numpy(g["breaks_vect"]).to(device)
            wsb = L.sqz_segment_stats_workspace_bytes(n_rows, n_seg)
            ws = torch.empty(wsb + 256, dtype=torch.uint8, device=device)
            wsp = (ws.data_ptr() + 255) & ~255
            dout = torch.empty(6 * n_seg, dtype=torch.float64, device=device)

            def call():
                _lib.check(L.sqz_segment_stats_dev(
                    24, dro.data_ptr(), dso.data_ptr(), n_rows, n_seg,
                    *[t.data_ptr() for t in dv],
                    dhet.data_ptr(), dbrk.data_ptr(), 1.0, 1, wsp, wsb, stream, dout.data_ptr()))
            ms = timed_ms(torch, call, steps, args.warmup, env["flush"])
            got = dout.cpu().numpy().reshape(6, n_seg).T
            te = time.perf_counter()
            host = G.segment_stats(g["row_off"], g["seg_off"], *[g[k] for k in keys], g["is_het"],
                                   g["breaks_vect"])
            te = time.perf_counter() - te
            assert np.array_equal(host, got, equal_nan=True)
            c = 5  # CPU sample: one chromosome (1e6 rows, 100 segments)
            r = slice(g["row_off"][c], g["row_off"][c + 1])
            tc = time.perf_counter()
            ref = O.segment_stats([0, 1_000_000], [0, 100], *[g[k][r] for k in keys],
                                  g["is_het"][r], g["breaks_vect"][200 * c:200 * (c + 1)])
            tc = time.perf_counter() - tc
            mine = got[100 * c:100 * (c + 1)]
            assert np.array_equal(mine[:, [1, 4]], ref[:, [1, 4]])
            ok = ~np.isnan(ref)
            assert np.max(np.abs(mine[ok] - ref[ok]) / np.maximum(np.abs(ref[ok]), 1e-300)) < 1e-12
            peak, peak_src = hbm_peak()
            # algorithmic bytes: every seqz column of every row once (6 x 8 + 1) + the table out
            bytes_alg = 49.0 * n_rows + 48.0 * n_seg
            line = dict(base, metric="seqz rows/sec, segment.breaks per-segment statistics (fp64)",
                        value=n_rows / (ms * 1e-3), unit="rows/s", ms_per_step=ms,
                        config={"workload": "segment.breaks statistics (R/segments.R:113-198, "
                                "weighted.mean = TRUE) + b_allele_freq mode (R/model.R:69-98): "
                                "2.4e7 synthetic seqz rows, 24 chromosomes, 2,400 segments, 30 % "
                                "het", "rows": n_rows, "segments": n_seg,
                                "l2": "256 MiB buffer written between timed iterations"},
                        e2e={"value": n_rows / te, "unit": "rows/s", "ms_per_step": te * 1e3,
                             "h2d_bytes_per_step": 49 * n_rows, "d2h_bytes_per_step": 48 * n_seg,
                             "api": "sqz_segment_stats (pageable host vectors)"},
                        gpu_launches=2 * steps,
                        roofline={"bound": "hbm", "kernel": "segstat_reduce_kernel + "
                                  "segstat_kde_kernel", "achieved": bytes_alg / (ms * 1e-3) / 1e9,
                                  "peak": peak, "unit": "GB/s",
                                  "frac": bytes_alg / (ms * 1e-3) / 1e9 / peak,
                                  "peak_source": peak_src, "traffic": None,
                                  "bytes_per_row": 49},
                        cpu_baseline={"value": 1_000_000 / tc, "unit": "rows/s", "cores": 1,
                                      "kind": "port", "sample": "one chromosome (1e6 rows, 100 "
                                      "segments), oracle segment_stats, %.1f s" % tc})
        else:  # mufreq_grid
            from oracle import oracle as O
            n = 10000
            mut = synth_mutations(n, seed=7)
            wv = np.round(np.random.default_rng(1).uniform(20, 200, n))  # round(good.reads)
            ce, pl = S.r_seq(0.1, 1, 0.01), S.r_seq(1, 7, 0.1)
            kw = dict(mufreq=mut["mufreq"], depth_ratio=mut["depth_ratio"], cellularity=ce,
                      ploidy=pl, avg_depth_ratio=1.0, weight_mufreq=wv, weight_ratio=2 * wv)
            S.mufreq_grid_sums(n_gpus=n_dev, **kw)
            t = time.perf_counter()
            for _ in range(steps):
                got = S.mufreq_grid_sums(n_gpus=n_dev, **kw)
            ms = (time.perf_counter() - t) * 1e3 / steps
            cores = os.cpu_count() or 1
            tc = time.perf_counter()
            ref = O.mufreq_grid_sums(g0=2700, g1=2700 + 2 * cores, nthreads=cores, **kw)
            tc = time.perf_counter() - tc
            assert np.max(np.abs(got[2700:2700 + ref.size] - ref) / np.abs(ref)) < 1e-9
            evals = float(n) * ce.size * pl.size
            line = dict(base, metric="grid-point x position evaluations/sec (fp64), "
                        "mufreq.model.fit", value=evals / (ms * 1e-3), unit="evals/s",
                        ms_per_step=ms,
                        config={"workload": "mufreq.model.fit (R/bayes.R:175-194) as "
                                "sequenza.fit(method='mufreq') calls it: 1e4 positions, "
                                "per-position weights, grid 91 x 61, CNt 1..7 (T=35)",
                                "positions": n, "grid": [91, 61]},
                        e2e={"value": evals / (ms * 1e-3), "unit": "evals/s", "ms_per_step": ms,
                             "h2d_bytes_per_step": 32 * n, "d2h_bytes_per_step": 8 * 5551,
                             "api": "sqz_mufreq_grid (host pointers)"},
                        value_note="timed through the host API: value == e2e",
                        gpu_launches=2 * steps * n_dev,
                        cpu_baseline={"value": ref.size * n / tc, "unit": "evals/s",
                                      "cores": cores, "kind": "port",
                                      "sample": "%d grid points x 1e4 positions, %.1f s" % (ref.size, tc)})
        line["clocks"] = sampler.finish(t0, time.perf_counter())
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier(group=env["host_group"])


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="config3",
                    choices=["config3", "config2", "config4", "config5", "mufreq_grid",
                             "segstats"])
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        run_reference(args, rank, world)
        return
    if world > 1:
        import torch
        import torch.distributed as dist
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        torch.cuda.set_device(local_rank)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    try:
        run_ours(args, rank, world, local_rank)
    finally:
        if world > 1:
            import torch.distributed as dist
            dist.destroy_process_group()


if __name__ == "__main__":
    main()
