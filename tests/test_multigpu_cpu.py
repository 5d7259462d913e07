"""Host-side logic of the multi-GPU slab decomposition, on CPU: slab/neighbour tables and the point-to-point
exchange protocol under torch.distributed with the gloo backend, world sizes 2 and 3 (the same function that
runs over NCCL on the GPUs).  The CUDA pack/unpack kernels are covered by tests/test_gpu_multirank.py."""
import math
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from sph_b200 import multigpu


def test_slab_tables():
    bounds = [0.0, 1.0, 2.5, 4.0]
    s0, s1, s2 = (multigpu.make_slab(r, 3, bounds, 0.4) for r in range(3))
    assert (s0.lo, s0.hi, s0.has_left, s0.has_right) == (-math.inf, 1.0, 0, 1)
    assert (s1.lo, s1.hi, s1.has_left, s1.has_right) == (1.0, 2.5, 1, 1)
    assert (s2.lo, s2.hi, s2.has_left, s2.has_right) == (2.5, math.inf, 1, 0)
    assert s0.shift_left == s0.shift_right == s2.shift_right == 0.0
    p0, p2 = multigpu.make_slab(0, 3, bounds, 0.4, period=4.0), multigpu.make_slab(2, 3, bounds, 0.4, period=4.0)
    assert (p0.lo, p0.has_left, p0.shift_left) == (0.0, 1, 4.0)          # records leaving rank 0 leftwards reappear at x + L
    assert (p2.hi, p2.has_right, p2.shift_right) == (4.0, 1, -4.0)
    assert multigpu.neighbours(0, 3) == (None, 1) and multigpu.neighbours(2, 3) == (1, None)
    assert multigpu.neighbours(0, 3, periodic=True) == (2, 1) and multigpu.neighbours(2, 3, periodic=True) == (1, 0)
    assert multigpu.neighbours(0, 1, periodic=True) == (0, 0)            # a lone rank in a periodic box is its own neighbour
    one = multigpu.make_slab(0, 1, [0.0, 4.0], 0.4)
    assert (one.lo, one.hi, one.has_left, one.has_right) == (-math.inf, math.inf, 0, 0)
    solo = multigpu.make_slab(0, 1, [0.0, 4.0], 0.4, period=4.0, yz_box=((0.0, 2.0), None))
    assert (solo.lo, solo.hi, solo.has_left, solo.has_right, solo.shift_left, solo.shift_right) == (0.0, 4.0, 1, 1, 4.0, -4.0)
    assert (solo.periodic_y, solo.ylo, solo.ylen, solo.periodic_z) == (1, 0.0, 2.0, 0)


def test_quantile_bounds_balance():
    rng = np.random.default_rng(0)
    x = np.concatenate([rng.uniform(0, 1, 30000), rng.uniform(0, 4, 10000)])     # dam-break-like: dense on the left
    b = multigpu.quantile_bounds(x, 4, 0.2025, 0.0, 4.0)
    assert b[0] == 0.0 and b[-1] == 4.0 and all(b[i] < b[i + 1] for i in range(4))
    counts = np.histogram(x, b)[0]
    assert counts.max() / counts.min() < 2.2      # snapping to a cell plane costs up to half a cell of particles
    assert all(abs((v / 0.2025) - round(v / 0.2025)) < 1e-9 for v in b[1:-1])       # snapped to cell planes


def test_balanced_bounds_even_out_owned_plus_ghosts():
    rng = np.random.default_rng(1)
    x = np.concatenate([rng.uniform(0, 24, 80000), rng.uniform(9, 15, 60000)])    # a dense column in the middle of a pool
    halo, cell = 0.41, 0.2025
    def imbalance(b):
        w = [((x >= b[r] - (halo if r else 0)) & (x < b[r + 1] + (halo if r < 7 else 0))).sum() for r in range(8)]
        return max(w) / np.mean(w)
    q = multigpu.quantile_bounds(x, 8, cell, 0.0, 24.0)
    b = multigpu.balanced_bounds(x, 8, cell, 0.0, 24.0, halo)
    assert b[0] == 0.0 and b[-1] == 24.0 and all(b[i] < b[i + 1] for i in range(8))
    assert all(abs((v / cell) - round(v / cell)) < 1e-9 for v in b[1:-1])
    assert imbalance(b) < imbalance(q) and imbalance(b) < 1.12
    assert multigpu.balanced_bounds(x, 1, cell, 0.0, 24.0, halo) == [0.0, 24.0]


def _worker(rank, world, port, periodic, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        left, right = multigpu.neighbours(rank, world, periodic)
        n = 64
        send_left = torch.full((n,), float(10 * rank + 1))
        send_right = torch.full((n,), float(10 * rank + 2))
        recv_left, recv_right = torch.full((n,), -1.0), torch.full((n,), -1.0)
        for _ in range(3):      # several steps in a row: matching must stay in lock-step
            multigpu.exchange(dist, send_left, send_right, recv_left, recv_right, left, right)
        q.put((rank, left, right, float(recv_left[0]), float(recv_right[-1])))
    finally:
        dist.destroy_process_group()


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


@pytest.mark.parametrize("world,periodic", [(2, False), (3, False), (3, True), (2, True)])
def test_exchange_protocol_gloo(world, periodic):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, periodic, q)) for r in range(world)]
    for p in procs:
        p.start()
    got = {}
    for _ in range(world):
        rank, left, right, from_left, from_right = q.get(timeout=120)
        got[rank] = (left, right, from_left, from_right)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    for rank, (left, right, from_left, from_right) in got.items():
        # what arrives from the left neighbour is what it sent to ITS right, and vice versa
        assert from_left == (10 * left + 2 if left is not None else -1.0)
        assert from_right == (10 * right + 1 if right is not None else -1.0)


def _limit(capacity, last):
    """SlabRank._limit without an engine: the size both ends of a link derive from the last message's record count."""
    class _S:
        adaptive = True
    s = _S()
    s.capacity = capacity
    return multigpu.SlabRank._limit(s, last)


def _worker_prefix(rank, world, port, q):
    """Messages travel as prefixes whose length each end derives on its own: the sender from what it sent last step,
    the receiver from what it received last step (the same number).  Bytes beyond the prefix must stay untouched."""
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        left, right = multigpu.neighbours(rank, world, periodic=True)
        capacity = 65536
        nbytes_full = 32 * (capacity + 1)
        mk = lambda v: torch.full((nbytes_full,), v, dtype=torch.uint8)
        out_l = out_r = in_l = in_r = capacity
        ok = True
        for step in range(4):
            sent_l, sent_r = 100 + 7 * rank + step, 200 + 5 * rank + step         # records "packed" this step
            send_left, send_right = mk(1 + rank), mk(101 + rank)
            send_left[:8] = torch.frombuffer(np.array([sent_l, out_l], np.int32).tobytes(), dtype=torch.uint8)
            send_right[:8] = torch.frombuffer(np.array([sent_r, out_r], np.int32).tobytes(), dtype=torch.uint8)
            recv_left, recv_right = mk(255), mk(255)
            nb = tuple(32 * (k + 1) for k in (out_l, out_r, in_l, in_r))
            multigpu.exchange(dist, send_left, send_right, recv_left, recv_right, left, right, nb)
            got_l = int(np.frombuffer(recv_left[:4].numpy().tobytes(), np.int32)[0])   # header: count from the left neighbour
            got_r = int(np.frombuffer(recv_right[:4].numpy().tobytes(), np.int32)[0])
            ok &= got_l == 200 + 5 * left + step and got_r == 100 + 7 * right + step
            ok &= bool((recv_left[nb[2]:] == 255).all()) and bool((recv_right[nb[3]:] == 255).all())
            ok &= int(recv_left[nb[2] - 1]) == 101 + left and int(recv_right[nb[3] - 1]) == 1 + right
            out_l, out_r = _limit(capacity, sent_l), _limit(capacity, sent_r)
            in_l, in_r = _limit(capacity, got_l), _limit(capacity, got_r)
        q.put((rank, ok, out_l, out_r, in_l, in_r))
    finally:
        dist.destroy_process_group()


def test_prefix_exchange_sizes_agree_gloo():
    world = 3
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker_prefix, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    got = {}
    for _ in range(world):
        rank, ok, out_l, out_r, in_l, in_r = q.get(timeout=120)
        got[rank] = (ok, out_l, out_r, in_l, in_r)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    for rank, (ok, out_l, out_r, in_l, in_r) in got.items():
        left, right = multigpu.neighbours(rank, world, periodic=True)
        assert ok, f"rank {rank}: payload or untouched tail wrong"
        assert out_r == got[right][3] and out_l == got[left][4]      # my next send size == what that neighbour will post
        assert out_l < 65536 and out_r < 65536                       # and it really is a prefix
