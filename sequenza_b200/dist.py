"""One-process-per-GPU plumbing for the grid fit (torch for device memory, streams and
torch.distributed; the arithmetic is all in libsequenza_b200.so).

The cellularity x ploidy grid is laid out ploidy-fastest (R/bayes.R:106-107), so a
contiguous block of cellularity columns is a contiguous slice of result$LPP.  Rank r
evaluates columns shard_columns(nC, world, r) with the segment table replicated on
every GPU, and one all-gather (NCCL over NVLink) assembles the LPP vector on every rank.
"""
from __future__ import annotations

import numpy as np
import torch

from . import _lib
from .bayes import _priors, finish_model_fit


def shard_columns(nC: int, world: int, rank: int):
    """Contiguous, balanced column blocks: same rule as the C library's n_gpus split."""
    return nC * rank // world, nC * (rank + 1) // world


def max_shard_columns(nC: int, world: int) -> int:
    return max(shard_columns(nC, world, r)[1] - shard_columns(nC, world, r)[0]
               for r in range(world))


def gather_columns(local: torch.Tensor, nC: int, nP: int, world: int, rank: int,
                   group=None) -> torch.Tensor:
    """All-gather per-rank slices [S, nCs_r * nP] into [S, nC * nP] on every rank.
    Shards are padded to the widest one so a single all_gather_into_tensor suffices."""
    import torch.distributed as dist
    S = local.shape[0]
    width = max_shard_columns(nC, world) * nP
    send = local
    if local.shape[1] != width:
        send = torch.zeros((S, width), dtype=local.dtype, device=local.device)
        send[:, :local.shape[1]] = local
    send = send.contiguous()
    if world == 1:
        recv = send.unsqueeze(0)
    else:
        flat = torch.empty((world * S, width), dtype=local.dtype, device=local.device)
        dist.all_gather_into_tensor(flat, send, group=group)  # rank-major concatenation
        recv = flat.view(world, S, width)
    full = torch.empty((S, nC * nP), dtype=local.dtype, device=local.device)
    for r in range(world):
        c0, c1 = shard_columns(nC, world, r)
        full[:, c0 * nP:c1 * nP] = recv[r, :, :(c1 - c0) * nP]
    return full


class BafGridDevice:
    """Segment tables resident in HBM plus a reusable workspace; `run` launches the grid
    kernels for a block of cellularity columns on the current CUDA stream."""

    def __init__(self, Bf, depth_ratio, sd_Bf, sd_ratio, weight_Bf, weight_ratio,
                 avg_depth_ratio, seg_offsets=None, device=None):
        self.device = torch.device("cuda", torch.cuda.current_device()) if device is None \
            else torch.device(device)
        r = _lib.vec(depth_ratio)
        n = r.size
        host = [_lib.vec(a, n) for a in (Bf, r, sd_Bf, sd_ratio, weight_Bf, weight_ratio)]
        if seg_offsets is None:
            seg_offsets = np.array([0, n], dtype=np.int64)
        self.h_offsets = np.ascontiguousarray(seg_offsets, dtype=np.int64)
        self.S = self.h_offsets.size - 1
        self.N = n
        self.h_inputs = [torch.from_numpy(a).pin_memory() for a in host]
        self.h_adr = torch.from_numpy(_lib.vec(avg_depth_ratio, self.S)).pin_memory()
        self.h_off_t = torch.from_numpy(self.h_offsets).pin_memory()
        self.upload()
        self._ws = None
        self.ws_cap = 16 << 30  # bytes of pre-pass scratch the grid fit may use per call

    def upload(self):
        """Host (pinned) -> device copies of every per-segment input; returns bytes moved."""
        self.d_inputs = [t.to(self.device, non_blocking=True) for t in self.h_inputs]
        self.d_adr = self.h_adr.to(self.device, non_blocking=True)
        self.d_off = self.h_off_t.to(self.device, non_blocking=True)
        return sum(t.numel() * t.element_size()
                   for t in self.h_inputs + [self.h_adr, self.h_off_t])

    def run(self, ploidy, cellularity, ic0=0, ic1=None, CNt_min=0, CNt_max=7, CNn=2,
            priors_table=None, ratio_priority=False, mode=_lib.SQZ_MODE_FAST, out=None):
        pl, ce = _lib.vec(ploidy), _lib.vec(cellularity)
        ic1 = ce.size if ic1 is None else ic1
        pc, pv = _priors(priors_table, CNt_min, CNt_max)
        L = _lib.lib()
        need = L.sqz_baf_grid_workspace_preferred_bytes(
            self.N, self.S, pl.size, ic1 - ic0, int(CNt_max) - int(CNt_min) + 1, self.ws_cap)
        if self._ws is None or self._ws.numel() < need + 256:
            self._ws = torch.empty(need + 256, dtype=torch.uint8, device=self.device)
        base = self._ws.data_ptr()
        ws_ptr = (base + 255) & ~255
        if out is None:
            out = torch.empty((self.S, (ic1 - ic0) * pl.size), dtype=torch.float64,
                              device=self.device)
        stream = torch.cuda.current_stream(self.device).cuda_stream
        with torch.cuda.device(self.device):
            _lib.check(L.sqz_baf_grid_dev(
                self.S, self.d_off.data_ptr(), self.N,
                *[t.data_ptr() for t in self.d_inputs], _lib.ptr(pl), pl.size, _lib.ptr(ce),
                ce.size, ic0, ic1, self.d_adr.data_ptr(), int(CNt_min), int(CNt_max),
                int(CNn), _lib.ptr(pc), _lib.ptr(pv), pc.size, int(bool(ratio_priority)),
                int(mode), ws_ptr, need, stream, out.data_ptr()))
        return out


def run_noalloc(dev: "BafGridDevice", ploidy, cellularity, ic0=0, ic1=None, CNt_min=0,
                CNt_max=7, CNn=2, priors_table=None, ratio_priority=False,
                mode=_lib.SQZ_MODE_FAST, out=None):
    """BafGridDevice.run through sqz_baf_grid_dev_noalloc: the library's built-in 16 MiB
    scratch replaces the torch-allocated workspace, for embedders that must not trigger
    device allocations once they are running.  `out` should be preallocated by such a
    caller (it is allocated here only when omitted).  Calls on one device share the
    scratch: keep them on one stream."""
    pl, ce = _lib.vec(ploidy), _lib.vec(cellularity)
    ic1 = ce.size if ic1 is None else ic1
    pc, pv = _priors(priors_table, CNt_min, CNt_max)
    if out is None:
        out = torch.empty((dev.S, (ic1 - ic0) * pl.size), dtype=torch.float64,
                          device=dev.device)
    stream = torch.cuda.current_stream(dev.device).cuda_stream
    with torch.cuda.device(dev.device):
        _lib.check(_lib.lib().sqz_baf_grid_dev_noalloc(
            dev.S, dev.d_off.data_ptr(), dev.N, *[t.data_ptr() for t in dev.d_inputs],
            _lib.ptr(pl), pl.size, _lib.ptr(ce), ce.size, ic0, ic1, dev.d_adr.data_ptr(),
            int(CNt_min), int(CNt_max), int(CNn), _lib.ptr(pc), _lib.ptr(pv), pc.size,
            int(bool(ratio_priority)), int(mode), stream, out.data_ptr()))
    return out


def baf_model_fit_distributed(cellularity, ploidy, group=None, **kwargs):
    """baf.model.fit (R/bayes.R:103-124) across the ranks of an initialised
    torch.distributed job: every rank passes the same arguments, evaluates its block of
    cellularity columns on its own GPU, and receives the full result after one
    all-gather.  Returns list(ploidy, cellularity, lpp) as a dict on every rank."""
    import torch.distributed as dist
    world = dist.get_world_size(group) if dist.is_initialized() else 1
    rank = dist.get_rank(group) if dist.is_initialized() else 0
    fit_keys = ("CNt_min", "CNt_max", "CNn", "priors_table", "ratio_priority", "mode")
    run_kw = {k: kwargs.pop(k) for k in fit_keys if k in kwargs}
    kwargs.setdefault("sd_Bf", 0.1)
    kwargs.setdefault("sd_ratio", 0.5)
    kwargs.setdefault("weight_Bf", 1)
    kwargs.setdefault("weight_ratio", 1)
    dev = BafGridDevice(**kwargs)
    pl, ce = _lib.vec(ploidy), _lib.vec(cellularity)
    c0, c1 = shard_columns(ce.size, world, rank)
    local = dev.run(pl, ce, c0, c1, **run_kw)
    full = gather_columns(local, ce.size, pl.size, world, rank, group)
    return finish_model_fit(full[0].cpu().numpy(), pl, ce)
