#!/usr/bin/env python3
"""Measurements for the BASELINE configs that are not the headline bench line:
config 4 (mufreq.bayes on 1e7 positions) and config 5 (cohort of 512 samples, one launch).
Prints one JSON line per measurement.  python tools/bench_extra.py [config4] [config5]"""
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
import sequenza_b200 as S  # noqa: E402
from sequenza_b200 import _lib  # noqa: E402
from sequenza_b200.synth import baf_args, synth_mutations, synth_segments  # noqa: E402

PRI = {"CN": [2], "value": [2]}
which = sys.argv[1:] or ["config4", "config5"]
L = S.lib()
dev = torch.device("cuda", 0)


def timed(fn, steps=10, warm=3):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / steps


if "config4" in which:
    n = 10_000_000
    mut = synth_mutations(n)
    f = torch.from_numpy(mut["mufreq"]).to(dev)
    r = torch.from_numpy(mut["depth_ratio"]).to(dev)
    out = torch.empty(4 * n, dtype=torch.float64, device=dev)
    stream = torch.cuda.current_stream().cuda_stream
    for cn_max in (7, 20):
        T = sum(c + 1 for c in range(1, cn_max + 1))
        tb = L.sqz_mufreq_scalar_table_bytes(100.0, 100.0)
        table = torch.empty(tb + 64, dtype=torch.uint8, device=dev)
        tptr = (table.data_ptr() + 31) & ~31
        ms = timed(lambda: _lib.check(L.sqz_mufreq_point_scalar_dev(
            f.data_ptr(), r.data_ptr(), n, 100.0, 100.0, 0.7, 2.4, 1.0, 1, cn_max, 2, None, None,
            0, tptr, tb, stream, out.data_ptr())))
        hbm = 48.0 * n / (ms * 1e-3) / 1e9
        print(json.dumps({"config": "config4 mufreq.bayes, 1e7 positions, scalar weights 100/100, "
                          "CNt.max=%d (T=%d)" % (cn_max, T), "kernel": "mufreq_table + mufreq_lookup",
                          "ms": ms, "positions_per_s": n / (ms * 1e-3), "hbm_GBps_algorithmic": hbm,
                          "hbm_frac_of_measured_6549.8": hbm / 6549.8}))
        wv = torch.full((n,), 100.0, dtype=torch.float64, device=dev)
        ms = timed(lambda: _lib.check(L.sqz_mufreq_point_dev(
            f.data_ptr(), r.data_ptr(), wv.data_ptr(), wv.data_ptr(), n, 0.7, 2.4, 1.0, 1, cn_max,
            2, None, None, 0, stream, out.data_ptr())), steps=3, warm=1)
        print(json.dumps({"config": "config4 mufreq.bayes, 1e7 positions, per-position weights, "
                          "CNt.max=%d (T=%d)" % (cn_max, T), "kernel": "mufreq_point (direct nmath)",
                          "ms": ms, "positions_per_s": n / (ms * 1e-3),
                          "type_evals_per_s": n * T / (ms * 1e-3)}))
    # end to end through the host API (pageable numpy arrays, as R would pass them)
    t0 = time.perf_counter()
    res = S.mufreq_bayes(mut["mufreq"], mut["depth_ratio"], 0.7, 2.4, 1.0, CNt_max=20)
    dt = time.perf_counter() - t0
    print(json.dumps({"config": "config4 host API mufreq_bayes(1e7, CNt.max=20), H2D 160 MB + "
                      "D2H 320 MB included", "s": dt, "positions_per_s": n / dt}))

if "config5" in which:
    S_n, n_seg = 512, 3000
    ce, pl = S.r_seq(0.1, 1, 0.01), S.r_seq(1, 7, 0.1)
    samples = [baf_args(synth_segments(n_seg, cellularity=0.3 + 0.6 * (k % 50) / 50,
                                       ploidy=1.5 + 3.0 * (k % 37) / 37, seed=1000 + k))
               for k in range(S_n)]
    ndev = L.sqz_device_count()
    keys = ("Bf", "depth_ratio", "sd_Bf", "sd_ratio", "weight_Bf", "weight_ratio")
    cohort = {k: np.concatenate([s[k] for s in samples]) for k in keys}
    cohort["seg_offsets"] = np.arange(S_n + 1, dtype=np.int64) * n_seg
    cohort["avg_depth_ratio"] = np.array([s["avg_depth_ratio"] for s in samples])
    evals = float(S_n) * n_seg * ce.size * pl.size
    for g in sorted({1, ndev}):
        S.baf_model_fit_batch(samples[:8], cellularity=ce, ploidy=pl, priors_table=PRI, n_gpus=g)
        S.baf_model_fit_batch(cohort, cellularity=ce, ploidy=pl, priors_table=PRI, n_gpus=g)
        t0 = time.perf_counter()
        fits = S.baf_model_fit_batch(cohort, cellularity=ce, ploidy=pl, priors_table=PRI,
                                     n_gpus=g)
        dt = time.perf_counter() - t0
        # the C call alone (host vectors in, host result out)
        out = np.empty(S_n * ce.size * pl.size)
        pc, pv = np.array([2.0]), np.array([2.0])
        def call():
            _lib.check(L.sqz_baf_grid_batch(
                S_n, cohort["seg_offsets"].ctypes.data, *[_lib.ptr(cohort[k]) for k in keys],
                _lib.ptr(pl), pl.size, _lib.ptr(ce), ce.size, _lib.ptr(cohort["avg_depth_ratio"]),
                0, 7, 2, _lib.ptr(pc), _lib.ptr(pv), 1, 0, g, 0, _lib.ptr(out)))
        call()
        t0 = time.perf_counter()
        call()
        dc = time.perf_counter() - t0
        ok = sum(abs(S.get_ci(f)["max.cellularity"] - (0.3 + 0.6 * (k % 50) / 50)) <= 0.02
                 for k, f in enumerate(fits))
        print(json.dumps({"config": "config5 cohort: 512 samples x %d segments, grid 91 x 61, "
                          "CNt.max=7, one launch per device" % n_seg, "n_gpus": g,
                          "s_python_fit_and_finish": dt, "evals_per_s_python": evals / dt,
                          "s_sqz_baf_grid_batch": dc, "evals_per_s_c_call": evals / dc,
                          "samples_with_recovered_cellularity": int(ok)}))
