"""One-process-per-GPU plumbing for the grid fit (torch for device memory, streams and
torch.distributed; the arithmetic is all in libsequenza_b200.so).

The cellularity x ploidy grid (R/bayes.R:106-107, ploidy fastest) is cut into UNITS of
W columns x PT ploidies (sqz_baf_grid_plan); rank r evaluates the units
shard_units(n_units, world, r) with the segment table replicated on every GPU.  Kernel,
tile width and segment chunks are planned from the whole grid, so every grid point has the
same bits for any world size.  Each rank's last kernel writes its results at their final
place of a whole-grid buffer:

  * PeerGather (default when CUDA IPC works between the ranks' GPUs): that buffer is every
    rank's, mapped over NVLink, so the write IS the all-gather; one flag per rank and a
    wait kernel order the readers (sqz_peer_*).
  * otherwise: the rank's own zero-filled buffer, then ONE NCCL all-reduce(SUM) -- every
    grid point has exactly one non-zero contributor, so the sum is exact.
"""
from __future__ import annotations

import ctypes as C

import numpy as np
import torch

from . import _lib
from .bayes import _priors, finish_model_fit


def shard_columns(nC: int, world: int, rank: int):
    """Contiguous, balanced column blocks."""
    return nC * rank // world, nC * (rank + 1) // world


def shard_units(n_units: int, world: int, rank: int):
    """Contiguous, balanced unit ranges: same rule as the C library's n_gpus split."""
    return n_units * rank // world, n_units * (rank + 1) // world


def max_shard_columns(nC: int, world: int) -> int:
    return max(shard_columns(nC, world, r)[1] - shard_columns(nC, world, r)[0]
               for r in range(world))


def grid_plan(n_segments, n_samples, ploidy, nC, CNt_min=0, CNt_max=7, CNn=2,
              ratio_priority=False):
    """sqz_baf_grid_plan as a dict (host code only: needs no GPU)."""
    pl = _lib.vec(ploidy)
    out = (C.c_int64 * 8)()
    _lib.check(_lib.lib().sqz_baf_grid_plan(int(n_segments), int(n_samples), _lib.ptr(pl),
                                            pl.size, int(nC), int(CNt_min), int(CNt_max),
                                            int(CNn), int(bool(ratio_priority)), out))
    keys = ("PT", "nPT", "chunks", "threads", "T", "family", "W", "units")
    return dict(zip(keys, (int(v) for v in out)))


def unit_points(plan, nC, nP, u0, u1):
    """Boolean [nC, nP] mask of the grid points in units [u0, u1) (host logic mirror of
    in_units() in csrc/sqz_baf_grid.cuh; used by the CPU tests and the fallback gather)."""
    ic = np.arange(nC)[:, None] // plan["W"]
    it = np.arange(nP)[None, :] // plan["PT"]
    u = ic * plan["nPT"] + it
    return (u >= u0) & (u < u1)


def gather_columns(local: torch.Tensor, nC: int, nP: int, world: int, rank: int,
                   group=None) -> torch.Tensor:
    """All-gather per-rank COLUMN slices [S, nCs_r * nP] into [S, nC * nP] on every rank.
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


def assemble_units(full: torch.Tensor, group=None) -> torch.Tensor:
    """`full` [S, nC * nP] holds this rank's units and zeros elsewhere: one all-reduce
    (every point has a single non-zero contributor, so SUM copies it exactly)."""
    import torch.distributed as dist
    if dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.all_reduce(full, op=dist.ReduceOp.SUM, group=group)
    return full


class _RawCuda:
    """A raw device pointer as a __cuda_array_interface__ object (float64 vector)."""

    def __init__(self, ptr, n):
        self.__cuda_array_interface__ = {"shape": (int(n),), "typestr": "<f8",
                                         "data": (int(ptr), False), "version": 3}


class PeerGather:
    """Whole-grid result buffers of all ranks mapped into every rank (CUDA IPC over
    NVLink): the grid fit's last kernel stores each result into all of them, then
    `finish` raises this rank's flag in every peer and waits for the peers' flags.
    Two buffers alternate between consecutive fits, so a fast rank never overwrites what
    a slow rank is still reading."""

    def __init__(self, n_doubles: int, device, group=None):
        import torch.distributed as dist
        self.L = _lib.lib()
        self.group = group
        self.world = dist.get_world_size(group)
        self.rank = dist.get_rank(group)
        self.device = torch.device(device)
        self.n = int(n_doubles)
        self.epoch = 0
        self.ok = False
        self.local = C.c_void_p()
        self.peers = []
        handle = (C.c_ubyte * 64)()
        nbytes = C.c_size_t(0)
        with torch.cuda.device(self.device):
            rc = self.L.sqz_peer_alloc(2 * self.n * 8, C.byref(self.local), handle,
                                       C.byref(nbytes))
        mine = (rc, bytes(handle), int(self.device.index))
        everyone = [None] * self.world
        dist.all_gather_object(everyone, mine, group=group)
        good = all(e[0] == 0 for e in everyone)
        if good:
            with torch.cuda.device(self.device):
                for r, (_, h, _dev) in enumerate(everyone):
                    if r == self.rank:
                        self.peers.append(self.local.value)
                        continue
                    p = C.c_void_p()
                    hb = (C.c_ubyte * 64).from_buffer_copy(h)
                    if self.L.sqz_peer_open(hb, C.byref(p)) != 0:
                        good = False
                        break
                    self.peers.append(p.value)
        flags = [None] * self.world
        dist.all_gather_object(flags, bool(good), group=group)
        self.ok = all(flags)
        if self.ok:
            self.data_off = int(self.L.sqz_peer_data_offset())
            self._ptr_arr = (C.c_void_p * self.world)(*self.peers)

    def out_pointers(self, parity: int):
        """Host array of the ranks' data buffers for this fit (own buffer first)."""
        order = [self.rank] + [r for r in range(self.world) if r != self.rank]
        ptrs = [self.peers[r] + self.data_off + parity * self.n * 8 for r in order]
        return (C.c_void_p * self.world)(*ptrs), self.world

    def view(self, parity: int, shape):
        t = torch.as_tensor(_RawCuda(self.local.value + self.data_off + parity * self.n * 8,
                                     self.n), device=self.device)
        return t.view(shape)

    def finish(self, stream):
        """Signal every peer, then wait for all of them (both on `stream`)."""
        self.epoch += 1
        with torch.cuda.device(self.device):
            _lib.check(self.L.sqz_peer_signal(self._ptr_arr, self.world, self.rank,
                                              self.epoch, stream))
            _lib.check(self.L.sqz_peer_wait(self.local, self.world, self.epoch, stream))

    def close(self):
        with torch.cuda.device(self.device):
            for r, p in enumerate(self.peers):
                if r != self.rank and p:
                    self.L.sqz_peer_close(C.c_void_p(p))
            if self.local:
                self.L.sqz_peer_free(self.local)
        self.peers, self.local, self.ok = [], C.c_void_p(), False


class BafGridDevice:
    """Segment tables resident in HBM plus a reusable workspace; `run` launches the grid
    kernels for a block of cellularity columns, `run_units` for a range of units, on the
    current CUDA stream."""

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
        self.d_inputs = None
        self.upload()
        self._ws = None
        self.ws_cap = 16 << 30  # bytes of pre-pass scratch the grid fit may use per call

    def upload(self):
        """Host (pinned) -> device copies of every per-segment input; returns bytes moved.
        The device buffers are reused from call to call."""
        src = self.h_inputs + [self.h_adr, self.h_off_t]
        if self.d_inputs is None:
            dst = [torch.empty_like(t, device=self.device) for t in src]
            self.d_inputs, self.d_adr, self.d_off = dst[:6], dst[6], dst[7]
        for d, s in zip(self.d_inputs + [self.d_adr, self.d_off], src):
            d.copy_(s, non_blocking=True)
        return sum(t.numel() * t.element_size() for t in src)

    def _workspace(self, need):
        if self._ws is None or self._ws.numel() < need + 256:
            self._ws = torch.empty(need + 256, dtype=torch.uint8, device=self.device)
        return (self._ws.data_ptr() + 255) & ~255

    def run(self, ploidy, cellularity, ic0=0, ic1=None, CNt_min=0, CNt_max=7, CNn=2,
            priors_table=None, ratio_priority=False, mode=_lib.SQZ_MODE_FAST, out=None):
        pl, ce = _lib.vec(ploidy), _lib.vec(cellularity)
        ic1 = ce.size if ic1 is None else ic1
        pc, pv = _priors(priors_table, CNt_min, CNt_max)
        L = _lib.lib()
        need = L.sqz_baf_grid_workspace_preferred_bytes(
            self.N, self.S, pl.size, ce.size, ic1 - ic0, int(CNt_max) - int(CNt_min) + 1,
            self.ws_cap)
        ws_ptr = self._workspace(need)
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

    def run_units(self, ploidy, cellularity, u0, u1, out=None, out_ptrs=None, CNt_min=0,
                  CNt_max=7, CNn=2, priors_table=None, ratio_priority=False,
                  mode=_lib.SQZ_MODE_FAST):
        """Units [u0, u1) of the whole grid; results at their final place of `out`
        ([S, nC * nP], untouched elsewhere) or of every buffer in `out_ptrs` (a ctypes
        array of device pointers and its length: PeerGather.out_pointers)."""
        pl, ce = _lib.vec(ploidy), _lib.vec(cellularity)
        pc, pv = _priors(priors_table, CNt_min, CNt_max)
        L = _lib.lib()
        rp = int(bool(ratio_priority))
        need = L.sqz_baf_grid_units_workspace_bytes(
            self.N, self.S, _lib.ptr(pl), pl.size, ce.size, int(CNt_min), int(CNt_max),
            int(CNn), rp, int(u0), int(u1), self.ws_cap)
        if need == 0:
            raise ValueError("bad unit range [%d, %d)" % (u0, u1))
        ws_ptr = self._workspace(need)
        if out_ptrs is None:
            if out is None:
                out = torch.zeros((self.S, ce.size * pl.size), dtype=torch.float64,
                                  device=self.device)
            out_ptrs = ((C.c_void_p * 1)(out.data_ptr()), 1)
        stream = torch.cuda.current_stream(self.device).cuda_stream
        with torch.cuda.device(self.device):
            _lib.check(L.sqz_baf_grid_dev_units(
                self.S, self.d_off.data_ptr(), self.N,
                *[t.data_ptr() for t in self.d_inputs], _lib.ptr(pl), pl.size, _lib.ptr(ce),
                ce.size, int(u0), int(u1), self.d_adr.data_ptr(), int(CNt_min), int(CNt_max),
                int(CNn), _lib.ptr(pc), _lib.ptr(pv), pc.size, rp, int(mode), ws_ptr, need,
                stream, out_ptrs[0], out_ptrs[1]))
        return out


class DistributedGridFit:
    """The grid fit of one sample table (or cohort) across the ranks of a torch.distributed
    job: plan once, then `step()` = this rank's units + the assembly of the whole LPP
    vector on every rank.  `peer=True` tries the NVLink peer-store path first."""

    def __init__(self, dev: BafGridDevice, ploidy, cellularity, group=None, peer=True,
                 **fit_kw):
        import torch.distributed as dist
        self.dev, self.group, self.kw = dev, group, fit_kw
        self.pl, self.ce = _lib.vec(ploidy), _lib.vec(cellularity)
        on = dist.is_initialized()
        self.world = dist.get_world_size(group) if on else 1
        self.rank = dist.get_rank(group) if on else 0
        self.plan = grid_plan(dev.N, dev.S, self.pl, self.ce.size,
                              fit_kw.get("CNt_min", 0), fit_kw.get("CNt_max", 7),
                              fit_kw.get("CNn", 2), fit_kw.get("ratio_priority", False))
        self.u0, self.u1 = shard_units(self.plan["units"], self.world, self.rank)
        self.shape = (dev.S, self.ce.size * self.pl.size)
        self.peer = None
        self.steps = 0
        if peer and self.world > 1 and dev.device.type == "cuda":
            pg = PeerGather(self.shape[0] * self.shape[1], dev.device, group)
            if pg.ok:
                self.peer = pg
            else:
                pg.close()
        self.full = None if self.peer else torch.zeros(self.shape, dtype=torch.float64,
                                                       device=dev.device)

    @property
    def collective(self):
        if self.world == 1:
            return "none"
        return "peer stores over NVLink (sqz_peer_*)" if self.peer else "NCCL all-reduce"

    def step(self):
        if self.peer:
            parity = self.steps & 1
            self.steps += 1
            self.dev.run_units(self.pl, self.ce, self.u0, self.u1,
                               out_ptrs=self.peer.out_pointers(parity), **self.kw)
            self.peer.finish(torch.cuda.current_stream(self.dev.device).cuda_stream)
            return self.peer.view(parity, self.shape)
        if self.world > 1:
            self.full.zero_()
        self.dev.run_units(self.pl, self.ce, self.u0, self.u1, out=self.full, **self.kw)
        return assemble_units(self.full, self.group)

    def close(self):
        if self.peer:
            self.peer.close()
            self.peer = None


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


def baf_model_fit_distributed(cellularity, ploidy, group=None, peer=True, **kwargs):
    """baf.model.fit (R/bayes.R:103-124) across the ranks of an initialised
    torch.distributed job: every rank passes the same arguments, evaluates its units of
    the grid on its own GPU, and holds the full result after the assembly.  Returns
    list(ploidy, cellularity, lpp) as a dict on every rank."""
    fit_keys = ("CNt_min", "CNt_max", "CNn", "priors_table", "ratio_priority", "mode")
    run_kw = {k: kwargs.pop(k) for k in fit_keys if k in kwargs}
    kwargs.setdefault("sd_Bf", 0.1)
    kwargs.setdefault("sd_ratio", 0.5)
    kwargs.setdefault("weight_Bf", 1)
    kwargs.setdefault("weight_ratio", 1)
    dev = BafGridDevice(**kwargs)
    fit = DistributedGridFit(dev, ploidy, cellularity, group=group, peer=peer, **run_kw)
    try:
        full = fit.step()
        host = full[0].cpu().numpy()
    finally:
        fit.close()
    return finish_model_fit(host, fit.pl, fit.ce)
