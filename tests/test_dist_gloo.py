"""world_size-2 checks of the multi-GPU plumbing (sequenza_b200/dist.py) on CPU with the
gloo backend: the column sharding rule and the padded all-gather that assembles
result$LPP.  The per-rank compute is replaced by a closed-form stand-in; the kernels
themselves are covered by the -m gpu tests."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from sequenza_b200.dist import gather_columns, max_shard_columns, shard_columns


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
