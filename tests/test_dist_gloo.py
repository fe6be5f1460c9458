"""world_size-2 checks of the multi-GPU plumbing (sequenza_b200/dist.py) on CPU with the
gloo backend: the column and unit sharding rules, the padded all-gather of column shards
and the one-collective assembly of unit shards.  The per-rank compute is replaced by a
closed-form stand-in; the kernels themselves are covered by the -m gpu tests."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from sequenza_b200.dist import (assemble_units, gather_columns, grid_plan, max_shard_columns,
                                shard_columns, shard_units, unit_points)


def test_shard_columns_cover_and_balance():
    for nC in (1, 2, 7, 91, 901):
        for world in (1, 2, 3, 4, 8):
            blocks = [shard_columns(nC, world, r) for r in range(world)]
            assert blocks[0][0] == 0 and blocks[-1][1] == nC
            assert all(blocks[r][1] == blocks[r + 1][0] for r in range(world - 1))
            sizes = [b - a for a, b in blocks]
            assert max(sizes) - min(sizes) <= 1
            assert max_shard_columns(nC, world) == max(sizes)


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, nC, nP, S, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        c0, c1 = shard_columns(nC, world, rank)
        # stand-in for the kernel: value encodes (sample, column, ploidy index)
        s = torch.arange(S, dtype=torch.float64)[:, None, None] * 1e6
        c = torch.arange(c0, c1, dtype=torch.float64)[None, :, None] * 1e3
        p = torch.arange(nP, dtype=torch.float64)[None, None, :]
        local = (s + c + p).reshape(S, (c1 - c0) * nP)
        full = gather_columns(local, nC, nP, world, rank)
        q.put((rank, full.numpy()))
    except Exception as e:  # surface the failure instead of letting the parent time out
        q.put((rank, repr(e)))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("nC,nP,S", [(7, 5, 1), (4, 3, 2), (1, 6, 1)])
def test_gather_columns_world2(nC, nP, S):
    world = 2
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, nC, nP, S, q))
             for r in range(world)]
    for p in procs:
        p.start()
    got = dict(q.get(timeout=120) for _ in range(world))
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    s = np.arange(S)[:, None, None] * 1e6
    c = np.arange(nC)[None, :, None] * 1e3
    p_ = np.arange(nP)[None, None, :]
    want = (s + c + p_).reshape(S, nC * nP)
    for r in range(world):
        assert not isinstance(got[r], str), got[r]
        np.testing.assert_array_equal(got[r], want)


def _unit_worker(rank, world, port, nC, nP, S, n_seg, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        plan = grid_plan(n_seg, S, np.linspace(1, 7, nP), nC)
        u0, u1 = shard_units(plan["units"], world, rank)
        mine = unit_points(plan, nC, nP, u0, u1).reshape(-1)
        # stand-in for sqz_baf_grid_dev_units: this rank's points, zeros elsewhere
        value = -(np.arange(S)[:, None] * 1e7 + np.arange(nC * nP)[None, :] + 0.25)
        full = torch.from_numpy(np.where(mine[None, :], value, 0.0))
        q.put((rank, assemble_units(full).numpy(), plan["W"]))
    except Exception as e:
        q.put((rank, repr(e), 0))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("nC,nP,S,n_seg", [(901, 601, 1, 100000), (91, 61, 2, 3000),
                                            (5, 3, 1, 10)])
def test_assemble_units_world2(nC, nP, S, n_seg):
    """Unit shards (column-per-lane units of 32 x 32 for the fine grid, (column, tile) units
    otherwise) assembled by one all-reduce: every grid point arrives exactly once."""
    world = 2
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_unit_worker, args=(r, world, port, nC, nP, S, n_seg, q))
             for r in range(world)]
    for p in procs:
        p.start()
    got = {}
    for _ in range(world):
        r, arr, w = q.get(timeout=180)
        got[r] = (arr, w)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    want = -(np.arange(S)[:, None] * 1e7 + np.arange(nC * nP)[None, :] + 0.25)
    for r in range(world):
        assert not isinstance(got[r][0], str), got[r][0]
        np.testing.assert_array_equal(got[r][0], want)
    assert got[0][1] == (32 if nC == 901 else 1)
