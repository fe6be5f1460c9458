#!/usr/bin/env python3
"""bench.py -- throughput of the cellularity x ploidy grid fit (baf.model.fit) on B200.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
                    [--workload config2|config3]

One "step" is one complete grid fit (R/bayes.R:106-116) over one synthetic segment
table.  Metric (BASELINE.json): grid-point x segment evaluations per second, fp64; one
evaluation = the full max-over-types posterior of one (cellularity, ploidy, segment).

Workloads
  config2 (default; BASELINE configs[1], the single-GPU configuration): 3e4 synthetic
      segments, CNt.max = 20 (121 types), grid 91 cellularities x 61 ploidies per GPU.
      With N GPUs the cellularity axis has 91*N values (weak scaling: per-GPU work fixed).
  config3 (BASELINE configs[2]): 1e5 segments, CNt.max = 7, grid 901 x 601, split over
      the N GPUs by cellularity column (strong scaling).

Arms
  ours       value: inputs resident in HBM, CUDA events on the launch stream, max over
             ranks; e2e: the same fit through the public host API with pinned HOST
             buffers (H2D of the segment table and D2H of result$LPP inside the timing).
  reference  the CPU restatement of the reference (oracle/, kind "port": R is not
             installed, so oracle/_ref cannot exist) on all host cores, on a bounded
             sample of the same workload's grid points.  Rank 0 only.
"""
from __future__ import annotations

import argparse
import ctypes as C
import json
import os
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = "grid-point x segment evaluations/sec (fp64), baf.model.fit"
PRI = {"CN": [2], "value": [2]}  # sequenza.fit default priors.table (R/fit.R:6)


def make_workload(name: str, world: int):
    from sequenza_b200.bayes import r_seq
    from sequenza_b200.synth import DEFAULT_SEED, baf_args, synth_segments
    if name == "config2":
        n, cn_max = 30000, 20
        pl = r_seq(1, 7, 0.1)
        ce = r_seq(0.1, 1, 0.01) if world == 1 else np.linspace(0.1, 1.0, 91 * world)
        scaling = "weak"
        desc = ("BASELINE configs[1]: 3e4 synthetic segments, CNt.max=20 (T=121), grid %d "
                "cellularities x 61 ploidies (91 columns per GPU)" % ce.size)
    elif name == "config3":
        n, cn_max = 100000, 7
        pl, ce = r_seq(1, 7, 0.01), r_seq(0.1, 1, 0.001)
        scaling = "strong"
        desc = ("BASELINE configs[2]: 1e5 synthetic segments, CNt.max=7 (T=20), grid 901 x "
                "601 split by cellularity column")
    else:
        raise SystemExit("unknown workload %s" % name)
    seg = synth_segments(n, cn_max=7, seed=DEFAULT_SEED)
    return {"name": name, "args": baf_args(seg), "ce": ce, "pl": pl, "cn_max": cn_max,
            "n": n, "scaling": scaling, "desc": desc,
            "n_c": cn_max + 1, "T": sum(c // 2 + 1 for c in range(cn_max + 1))}


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


def ncu_traffic(workload):
    """dram__bytes_read.sum + dram__bytes_write.sum of the dominant kernel, per launch, from
    the committed ncu --set full capture of this workload (profiles/r01_traffic.json)."""
    try:
        with open(os.path.join(ROOT, "profiles", "r01_traffic.json")) as f:
            return json.load(f)[workload]["dram_bytes_per_launch"]
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


def run_reference(args, rank, world):
    if rank != 0:
        return
    w = make_workload(args.workload, world)
    val, desc, dt = cpu_baseline(w, steps=max(1, args.steps), warmup=args.warmup)
    line = {"impl": "reference", "metric": METRIC, "value": val, "unit": "evals/s",
            "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": dt * 1e3, "higher_is_better": True, "scaling": w["scaling"],
            "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": w["desc"], "segments": w["n"], "CNt.max": w["cn_max"],
                       "grid": [int(w["ce"].size), int(w["pl"].size)],
                       "note": "reference R path cannot run here (R not installed); this "
                               "is the compiled CPU restatement on the host cores"},
            "cpu_baseline": desc,
            "e2e": {"value": val, "unit": "evals/s", "h2d_bytes_per_step": 0,
                    "d2h_bytes_per_step": 0}}
    print(json.dumps(line), flush=True)


def run_ours(args, rank, world, local_rank):
    import torch
    import torch.distributed as dist
    import sequenza_b200 as S
    from sequenza_b200 import _lib
    from sequenza_b200.dist import BafGridDevice, gather_columns, shard_columns

    if not torch.cuda.is_available() or S.lib().sqz_device_count() < 1:
        raise SystemExit("bench.py --impl ours needs a CUDA device (no CPU fallback)")
    torch.cuda.set_device(local_rank)
    device = torch.device("cuda", local_rank)
    L = S.lib()
    w = make_workload(args.workload, world)
    ce, pl = w["ce"], w["pl"]
    nC, nP, n = ce.size, pl.size, w["n"]
    c0, c1 = shard_columns(nC, world, rank)
    evals_total = float(nC) * nP * n

    dev = BafGridDevice(device=device, **w["args"])
    out = torch.empty((1, (c1 - c0) * nP), dtype=torch.float64, device=device)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=device)  # > 126 MB L2
    kw = dict(CNt_max=w["cn_max"], priors_table=PRI)

    def step():
        local = dev.run(pl, ce, c0, c1, out=out, **kw)
        return gather_columns(local, nC, nP, world, rank) if world > 1 else local

    def barrier():
        if world > 1:
            dist.barrier()

    sampler = ClockSampler(local_rank)
    sampler.start()
    for _ in range(max(args.warmup, 3)):
        step()
    torch.cuda.synchronize()
    L.sqz_profile_enable(1)
    ev0 = [torch.cuda.Event(enable_timing=True) for _ in range(args.steps)]
    ev1 = [torch.cuda.Event(enable_timing=True) for _ in range(args.steps)]
    barrier()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for k in range(args.steps):
        flush.zero_()  # L2 flush between timed iterations (outside the event pair)
        ev0[k].record()
        res = step()
        ev1[k].record()
    torch.cuda.synchronize()
    barrier()
    t1 = time.perf_counter()
    ms3 = (C.c_double * 3)()
    ncall = C.c_int(0)
    kpath = C.c_int(0)
    n_cand = C.c_ulonglong(0)
    L.sqz_profile_counters(C.byref(kpath), C.byref(n_cand))
    L.sqz_profile_collect(ms3, C.byref(ncall))
    L.sqz_profile_enable(0)
    clocks = sampler.finish(t0, t1)
    step_ms = sum(a.elapsed_time(b) for a, b in zip(ev0, ev1))
    tm = torch.tensor([step_ms], dtype=torch.float64, device=device)
    if world > 1:
        dist.all_reduce(tm, op=dist.ReduceOp.MAX)
    total_ms = float(tm.item())
    ms_per_step = total_ms / args.steps
    value = evals_total / (ms_per_step * 1e-3)
    lpp_sum = res[0].cpu().numpy()

    # ---- end to end through the public API, host buffers, copies inside the timing ----
    e2e_steps = args.steps
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
    else:
        hout = torch.empty((1, nC * nP), dtype=torch.float64).pin_memory()

        def e2e_step():
            dev.upload()
            full = step()
            hout.copy_(full, non_blocking=True)
            torch.cuda.synchronize()
        h2d = world * (6 * n * 8 + 16 + 8 + (nP + (c1 - c0)) * 8)
        d2h = world * nC * nP * 8
    for _ in range(3):
        e2e_step()
    barrier()
    torch.cuda.synchronize()
    te = time.perf_counter()
    for _ in range(e2e_steps):
        e2e_step()
    torch.cuda.synchronize()
    barrier()
    e2e_s = time.perf_counter() - te
    tm = torch.tensor([e2e_s], dtype=torch.float64, device=device)
    if world > 1:
        dist.all_reduce(tm, op=dist.ReduceOp.MAX)
    e2e_ms = float(tm.item()) * 1e3 / e2e_steps
    if world == 1:
        assert np.array_equal(hout.numpy(), lpp_sum), "e2e and resident results differ"

    if rank != 0:
        return
    # ---- roofline of the dominant kernel (baf_grid_fast_kernel) ------------------------
    plan = (C.c_int * 5)()
    L.sqz_baf_grid_plan(n, 1, nP, c1 - c0, 0, w["cn_max"], 2, 0, plan)
    PT, nPT = plan[0], plan[1]
    n_c, T = w["n_c"], w["T"]
    peak = C.c_double(0)
    L.sqz_fp64_peak(local_rank, 3, C.byref(peak))
    fast_ms = ms3[1] / max(ncall.value, 1)
    local_evals = float(c1 - c0) * nP * n
    padded = float(nPT * PT) / nP  # ragged last ploidy tile is computed and discarded
    # Executed work per evaluation.  Per evaluated (copy number, ploidy):
    # 2 DFMA + DSETP on the fp64 pipe and 2 FSEL; per evaluated copy number 2 DMUL; per
    # ploidy 1 DMUL; the per-column pre-pass (2 T compares + ~14 n_c for sqrt and scaling)
    # is shared by all nP ploidies of a column.  The split kernel evaluates every copy
    # number; the pruned kernel only those its bounds cannot exclude (counted by the
    # kernel itself while profiling) plus ~6 fp64 / ~4 integer instructions per copy number
    # and tile for the bounds.
    pre = (2.0 * T + 14.0 * n_c) / nP
    kernel_names = {0: "baf_grid_fast_kernel (fused)", 1: "baf_grid_split_kernel",
                    2: "baf_grid_pruned_kernel", 3: "baf_grid_colwarp_kernel"}
    if kpath.value == 3 and ncall.value > 0:
        # column-per-lane kernel: per (segment, tile) 6 fp64 + 4 integer instructions per
        # copy number for the bounds and ~10 + ~10 to set up the first candidate; the first
        # candidate costs 2 DFMA + DMUL per ploidy (no minimum), every further one that the
        # warp evaluates 2 DFMA + DSETP + 2 FSEL
        cand = n_cand.value / (float(ncall.value) * n * (c1 - c0) * nPT)  # per segment, tile
        inst = 3.0 + (cand - 1.0) * 3.0 + (6.0 * n_c + 10.0) / PT + pre
        math = inst + (cand - 1.0) * 2.0 + (4.0 * n_c + 10.0) / PT
        flop = 5.0 + (cand - 1.0) * 5.0 + (8.0 * n_c + 14.0) / PT + 1.5 * pre
    elif kpath.value == 2 and ncall.value > 0:
        cand = n_cand.value / (float(ncall.value) * n * (c1 - c0) * nPT)  # per segment, tile
        inst = cand * (3.0 + 2.0 / PT) + 1 + 6.0 * n_c / PT + pre
        math = inst + cand * 2.0 + 4.0 * n_c / PT
        flop = cand * (5.0 + 2.0 / PT) + 1 + 8.0 * n_c / PT + 1.5 * pre
    else:
        cand = float(n_c)
        inst = 3.0 * n_c + 1 + 2.0 * n_c / PT + pre         # fp64-pipe issue slots
        math = inst + 2.0 * n_c                             # + the FSEL pair of each minimum
        flop = 5.0 * n_c + 1 + 2.0 * n_c / PT + 1.5 * pre   # DFMA = 2 flops, others 1
    ach = local_evals * flop / (fast_ms * 1e-3) / 1e12 if fast_ms > 0 else None
    # reference formulation: 11 n_c + 13 T flops and n_c + T logs per eval
    ref_flops = 11.0 * n_c + 13.0 * T
    ref_logs = n_c + T
    roofline = {
        "bound": "fp64",
        "kernel": "%s<%d> (+ baf_colterm%s pre-pass)" % (
            kernel_names.get(kpath.value, "?"), PT, "_cg" if kpath.value == 3 else ""),
        "copy_numbers_evaluated_per_segment_and_tile": cand, "copy_numbers": n_c,
        "achieved": ach, "peak": peak.value, "unit": "TFLOP/s",
        "frac": (ach / peak.value) if ach and peak.value else None,
        "peak_source": "measured live: sqz_fp64_peak, dependent-free DFMA chains on all "
                       "SMs (MEASURED_PEAKS.json carries no fp64 figure)",
        "kernel_ms": fast_ms, "kernel_share_of_step": fast_ms / ms_per_step if ms_per_step else None,
        "fp64_issue_slots_per_eval": inst, "flops_per_eval_executed": flop,
        "pipe_util": (local_evals * inst / (fast_ms * 1e-3)) / (peak.value * 1e12 / 2)
        if fast_ms > 0 and peak.value else None,
        "math_issue_slots_per_eval": math,
        "math_dispatch_util": (local_evals * math / (fast_ms * 1e-3)) / (peak.value * 1e12 / 2)
        if fast_ms > 0 and peak.value else None,
        "math_dispatch_note": "on B200 every warp-wide math instruction (fp64 or the FSEL "
                              "pair of a minimum) holds the SMSP math dispatch ~2.2 cycles "
                              "(tools/microbench*.cu); this is the binding resource",
        "tile_padding": padded,
        "reference_formulation": {
            "flops_per_eval": ref_flops, "logs_per_eval": ref_logs,
            "tflops": local_evals * ref_flops / (fast_ms * 1e-3) / 1e12 if fast_ms > 0 else None,
            "tlogs_per_s": local_evals * ref_logs / (fast_ms * 1e-3) / 1e12 if fast_ms > 0 else None,
            "note": "the kernel reaches the same result with %.0fx fewer fp64 flops and "
                    "no per-type log (exact factorisation, DESIGN.md)" % (ref_flops / flop)},
        "traffic": ncu_traffic(w["name"]) if world == 1 else None,
        "algorithmic_hbm_bytes": 48 * n + 8 * (c1 - c0) * nP,
        "traffic_note": "the grid kernel reads the pre-pass scratch (16 B, or 8 B for the "
                        "column-per-lane kernel, x copy numbers x segments per column) once "
                        "per ploidy tile, mostly from L2; that scratch replaces recomputing "
                        "the BAF term in every ploidy tile; its DRAM traffic is far below the "
                        "HBM peak: the kernel is bound by math dispatch, not by this traffic",
    }
    cpu_val, cpu_desc, _ = cpu_baseline(w) if world == 1 else (None, None, None)
    line = {
        "metric": METRIC, "value": value, "unit": "evals/s", "n_gpus": world,
        "steps": args.steps, "warmup": max(args.warmup, 3), "ms_per_step": ms_per_step,
        "higher_is_better": True, "scaling": w["scaling"], "vs_baseline": None,
        "dtype": "f64", "data": "synthetic",
        "config": {"workload": w["desc"], "segments": n, "CNt.max": w["cn_max"],
                   "grid": [int(nC), int(nP)], "evals_per_step": evals_total,
                   "parallelism": "grid columns x%d" % world,
                   "l2": "256 MiB buffer written between timed iterations (L2 flush)",
                   "mode": "fast + faithful repair", "ploidy_tile": PT,
                   "segment_chunks": plan[2], "threads_per_cta": plan[3]},
        "clocks": clocks,
        "e2e": {"value": evals_total / (e2e_ms * 1e-3), "unit": "evals/s",
                "ms_per_step": e2e_ms, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                "api": "sqz_baf_grid (host pointers)" if world == 1 else
                       "BafGridDevice.upload + sqz_baf_grid_dev + NCCL all-gather + D2H"},
        "gpu_launches": 5 * args.steps,
        "gpu_launches_note": "per step: baf_prep, the pre-pass (baf_colterm or baf_colterm_cg), "
                             "the grid kernel named in roofline.kernel, baf_grid_finalize, "
                             "baf_grid_faithful (repair); all columns fit one batch of the "
                             "pre-pass scratch here; + torch copy kernels for N>1",
        "roofline": roofline,
        "fit_wall_ms": e2e_ms,
        "solution": {"cellularity": float(ce[int(np.argmax(lpp_sum)) // nP]) if world == 1 else None,
                     "ploidy": float(pl[int(np.argmax(lpp_sum)) % nP]) if world == 1 else None},
    }
    if cpu_desc:
        line["cpu_baseline"] = cpu_desc
    print(json.dumps(line), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="config2", choices=["config2", "config3"])
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
