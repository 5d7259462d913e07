"""Multi-GPU driver: x-slab domain decomposition, one process (one engine) per GPU.

Nothing like this exists in the reference (one thread).  Each rank owns the particles with
x in [lo, hi) and mirrors a ghost layer of 2 support radii from its two x-neighbours, so that ghost
densities are recomputed locally and ONE exchange per step suffices:

    engine.mg_pack(slab, msg_left, msg_right)      # CUDA: classify + deterministic compaction (libsphb200)
    exchange(...)                                  # torch.distributed point-to-point (NCCL over NVLink)
    engine.mg_unpack(slab, from_left, from_right)  # CUDA: append, refresh ownership flags, adopt
    engine.step(1)                                 # sort + density + fused force/integrate/collide

There is no collective on the data path: only the two neighbour messages.  torch is plumbing here (device
buffers for the messages, the process group); every particle operation is a kernel behind the C ABI.
"""
from __future__ import annotations

import math
import os

import numpy as np

from . import capi

RECORD = 32  # bytes per particle in a message: float4 posm + float4 vel


def support_radius(h=0.45) -> float:
    return float(np.float32(h) * np.float32(h))


def make_slab(rank: int, world: int, bounds, halo: float, period: float | None = None, yz_box=None) -> capi.Slab:
    """bounds: world+1 increasing x values; the outer two are replaced by -inf/+inf unless x is periodic
    (period = the x length of the whole box; with one rank the rank is then its own neighbour on both sides).
    yz_box = ((ylo, ylen) or None, (zlo, zlen) or None) makes y and/or z periodic (ghost images, see sphb_slab)."""
    s = capi.Slab()
    periodic = period is not None
    s.lo = float(bounds[rank]) if (rank > 0 or periodic) else -math.inf
    s.hi = float(bounds[rank + 1]) if (rank < world - 1 or periodic) else math.inf
    s.halo = float(halo)
    s.has_left = int(rank > 0 or periodic)
    s.has_right = int(rank < world - 1 or periodic)
    s.shift_left = float(period) if (periodic and rank == 0) else 0.0
    s.shift_right = -float(period) if (periodic and rank == world - 1) else 0.0
    if yz_box is not None:
        yb, zb = yz_box
        if yb is not None:
            s.periodic_y, s.ylo, s.ylen = 1, float(yb[0]), float(yb[1])
        if zb is not None:
            s.periodic_z, s.zlo, s.zlen = 1, float(zb[0]), float(zb[1])
    return s


def neighbours(rank: int, world: int, periodic: bool = False):
    """(left, right) ranks; with a periodic x a lone rank is its own neighbour."""
    left = rank - 1 if rank > 0 else (world - 1 if periodic else None)
    right = rank + 1 if rank < world - 1 else (0 if periodic else None)
    return left, right


def quantile_bounds(x: np.ndarray, world: int, cell: float, lo: float, hi: float):
    """Slab boundaries at particle-count quantiles of x, snapped to multiples of `cell`."""
    qs = np.quantile(x, np.linspace(0, 1, world + 1)[1:-1]) if world > 1 else []
    inner = [lo + round((q - lo) / cell) * cell for q in qs]
    return [lo] + inner + [hi]


def balanced_bounds(x: np.ndarray, world: int, cell: float, lo: float, hi: float, halo: float, iters: int = 40):
    """Slab boundaries that equalise what a rank WORKS on -- its own particles plus the ghost layers of width `halo`
    on both sides -- rather than what it owns: where particles are dense in x the quantile slabs get narrow and
    their ghosts can be half a slab again (8-way strong scaling of the block drop).  Starts from the count
    quantiles and moves every inner boundary towards the heavier neighbour; snapped to cell planes at the end."""
    b = np.array(quantile_bounds(x, world, cell, lo, hi), np.float64)
    if world < 2:
        return list(b)
    xs = np.sort(x.astype(np.float64))
    count = lambda a, c: np.searchsorted(xs, c) - np.searchsorted(xs, a)
    load = lambda bb: np.array([count(bb[r] - (halo if r else 0.0), bb[r + 1] + (halo if r < world - 1 else 0.0)) for r in range(world)], np.float64)
    best, best_imb = b.copy(), np.inf
    for _ in range(iters):
        w = load(b)
        imb = w.max() / w.mean()
        if imb < best_imb:
            best, best_imb = b.copy(), imb
        for k in range(1, world):                       # boundary k separates ranks k-1 and k
            width = min(b[k] - b[k - 1], b[k + 1] - b[k])
            b[k] += 0.25 * width * (w[k] - w[k - 1]) / (w[k] + w[k - 1])   # rank k heavier: give rank k-1 more room
        b[1:-1] = np.clip(b[1:-1], b[0] + 2 * halo, b[-1] - 2 * halo)
        b[1:-1] = np.maximum.accumulate(b[1:-1])
    inner = [lo + round((q - lo) / cell) * cell for q in best[1:-1]]
    return [lo] + inner + [hi]


def exchange(dist, send_left, send_right, recv_left, recv_right, left, right, nbytes=None):
    """Post the (at most) four point-to-point operations of one halo exchange and wait for them.
    Works on any backend (NCCL on device tensors, gloo on CPU tensors in the tests).  A rank that is its own
    neighbour (one rank, periodic x) copies locally.  nbytes = (out_left, out_right, in_left, in_right): transfer
    only these prefixes of the buffers (both ends of a link must pass the same number, see SlabRank)."""
    if nbytes is not None:
        send_left, send_right = send_left[:nbytes[0]], send_right[:nbytes[1]]
        recv_left, recv_right = recv_left[:nbytes[2]], recv_right[:nbytes[3]]
    me = dist.get_rank() if dist is not None and dist.is_initialized() else 0
    if left == me and right == me:       # world size 1, periodic x
        recv_left.copy_(send_right)
        recv_right.copy_(send_left)
        return
    ops = []
    if right is not None:
        ops.append(dist.P2POp(dist.isend, send_right, right))
    if left is not None:
        ops.append(dist.P2POp(dist.irecv, recv_left, left))
    if left is not None:
        ops.append(dist.P2POp(dist.isend, send_left, left))
    if right is not None:
        ops.append(dist.P2POp(dist.irecv, recv_right, right))
    if not ops:
        return
    for w in dist.batch_isend_irecv(ops):
        w.wait()


class SlabRank:
    """One rank's engine plus its halo transport.

    transport "nccl": four message buffers here, exchanged with torch.distributed point-to-point (NCCL over NVLink).
    transport "p2p":  a sphb_mg_link -- the pack kernel writes straight into the neighbours' receive buffers (peer memory over
                      NVLink), flags instead of a communication call; connect() wires the links up (CUDA IPC handles gathered with
                      torch.distributed between processes, or connect_local() for several ranks living in one process)."""

    def __init__(self, engine: capi.Engine, slab: capi.Slab, capacity: int, device_index: int, transport: str = "nccl"):
        import torch
        self.torch = torch
        self.engine = engine
        self.slab = slab
        self.capacity = int(capacity)
        self.transport = transport
        self.link = None
        dev = torch.device("cuda", device_index)
        self.stream = torch.cuda.ExternalStream(engine.stream(), device=dev)
        if transport == "p2p":
            self.link = engine.mg_link_create(self.capacity)
        else:
            nbytes = int(engine.lib.sphb_mg_message_bytes(self.capacity))
            mk = lambda: torch.zeros(nbytes, dtype=torch.uint8, device=dev)
            self.send_left, self.send_right, self.recv_left, self.recv_right = mk(), mk(), mk(), mk()
        self.counts = (0, 0, 0, 0)
        # Records each of the four messages may carry in the NEXT exchange.  The buffers are sized for the worst case
        # (capacity), but sending them whole every step costs more than the step's compute once slabs get thin
        # (8-way strong scaling: 2 x 33 MB per rank and step).  Both ends of a link know how many records the last
        # message held -- the sender from sphb_mg_sent, the receiver from the unpack counts -- so both apply the same
        # rule to that number and agree on the size without talking: 1.5 x last + 4096, at most the capacity.
        # Counts drift by well under a per cent per step; a message that would not fit is reported by unpack.
        self.adaptive = True
        self.out_left = self.out_right = self.in_left = self.in_right = self.capacity
        # After the first exchange the unpack no longer waits for the GPU (sphb_mg_unpack_async): its counts reach the host one
        # call late -- read at the start of the next step, when they have long arrived -- which is still before the next pack
        # needs them for the message limits, so both ends of a link keep deriving the same limits from the same exchange.
        self.use_async = True
        self.exchanges = 0
        self.status_pending = False

    # --- peer-to-peer link wiring
    def connect(self, dist, rank, world, left, right):
        """Between processes: everybody publishes its link's IPC handle, then opens its two neighbours'."""
        mine = self.engine.mg_link_export(self.link)
        if world > 1:
            handles = [None] * world
            dist.all_gather_object(handles, mine)
        else:
            handles = [mine]
        if world == 1:                 # a lone rank in a periodic box is its own neighbour: plain pointers, no IPC
            self.engine.mg_link_connect_local(self.link, self.link if left is not None else None, self.link if right is not None else None)
        else:
            self.engine.mg_link_connect_ipc(self.link, handles[left] if left is not None else None, handles[right] if right is not None else None)

    def connect_local(self, left_rank, right_rank):
        """Several ranks in one process."""
        self.engine.mg_link_connect_local(self.link, left_rank.link if left_rank is not None else None, right_rank.link if right_rank is not None else None)

    def close(self):
        if self.link is not None:
            self.engine.mg_link_destroy(self.link)
            self.link = None

    def _limit(self, last: int) -> int:
        return min(self.capacity, int(last * 1.5) + 4096) if self.adaptive else self.capacity

    def pack(self):
        self.poll()     # the message limits come from the last exchange's counts
        s = self.slab
        self.engine.mg_pack_limits(s, self.send_left.data_ptr() if s.has_left else None, self.out_left,
                                   self.send_right.data_ptr() if s.has_right else None, self.out_right)

    def _adopt(self, counts, sent):
        self.counts = tuple(counts)
        self.out_left, self.out_right = self._limit(sent[0]), self._limit(sent[1])
        self.in_left, self.in_right = self._limit(self.counts[1]), self._limit(self.counts[2])

    def poll(self):
        """Adopt the counts of the last asynchronous unpack (no-op if there is none outstanding)."""
        if self.status_pending:
            st = self.engine.mg_poll(True)
            self._adopt(st[:4], st[5:7])
            self.status_pending = False

    def unpack(self, sync=None):
        """Append the received messages.  The first exchange reads its counts back (they size everything that follows); later ones
        do not wait for the GPU unless sync=True (or use_async is off): call poll() before the next pack, as step() does."""
        s = self.slab
        ptrs = (self.recv_left.data_ptr() if s.has_left else None, self.recv_right.data_ptr() if s.has_right else None)
        if sync is None:
            sync = not (self.use_async and self.exchanges >= 1)
        self.exchanges += 1
        if sync:
            self.poll()
            counts = self.engine.mg_unpack(s, *ptrs)
            self._adopt(counts, self.engine.mg_sent())
        else:
            self.engine.mg_unpack_async(s, *ptrs)
            self.status_pending = True
        return self.counts

    def exchange_distributed(self, dist, left, right):
        # NCCL enqueues after the current stream: make that the engine's stream so pack -> send and recv -> unpack order
        nbytes = tuple(32 * (k + 1) for k in (self.out_left, self.out_right, self.in_left, self.in_right))
        with self.torch.cuda.stream(self.stream):
            exchange(dist, self.send_left, self.send_right, self.recv_left, self.recv_right, left, right, nbytes)

    def exchange_p2p(self):
        self.poll()
        self.engine.mg_exchange(self.link, self.slab)
        self.exchanges += 1
        self.status_pending = True

    def step(self, dist=None, left=None, right=None):
        if self.transport == "p2p":
            self.exchange_p2p()
            self.engine.step(1)
            return
        self.poll()
        self.pack()
        if dist is not None:
            self.exchange_distributed(dist, left, right)
        self.unpack()
        self.engine.step(1)

    def owned(self):
        """(posm, vel) of the owned particles in sorted order, as numpy arrays (diagnostics / tests)."""
        posm, vel = self.engine.download_soa()[:2]
        keep = vel[:, 3] != 0
        return posm[keep], vel[keep]


def exchange_in_process(ranks, periodic=False):
    """Emulates the neighbour exchange for several SlabRanks living in ONE process on one GPU (tests on a
    single-GPU box): the send buffers are copied into the neighbours' receive buffers."""
    world = len(ranks)
    for r in ranks:
        r.engine.synchronize()
    for i, r in enumerate(ranks):
        left, right = neighbours(i, world, periodic)
        if right is not None:   # only what a real exchange would transfer, and both ends must have agreed on it
            assert r.out_right == ranks[right].in_left, "the two ends of a link disagree on the message size"
            nb = 32 * (r.out_right + 1)
            ranks[right].recv_left[:nb].copy_(r.send_right[:nb])
        if left is not None:
            assert r.out_left == ranks[left].in_right, "the two ends of a link disagree on the message size"
            nb = 32 * (r.out_left + 1)
            ranks[left].recv_right[:nb].copy_(r.send_left[:nb])
    ranks[0].torch.cuda.synchronize()


# ---------------------------------------------------------------------------------------------------------
# bench.py --gpus N (launched under torchrun)

class _Solo:
    """world size 1: no process group; the only "collectives" are identities."""
    P2POp = isend = irecv = batch_isend_irecv = None

    class ReduceOp:
        MAX = SUM = None

    @staticmethod
    def is_initialized(): return False
    @staticmethod
    def get_rank(): return 0
    @staticmethod
    def barrier(): pass
    @staticmethod
    def all_reduce(t, op=None): return t
    @staticmethod
    def broadcast(t, src=0): return t
    @staticmethod
    def gather_object(obj, object_gather_list=None, dst=0):
        if object_gather_list is not None:
            object_gather_list[0] = obj


SCALE_OF = {"drop16m": 1, "drop2m": 2, "drop256k": 4, "periodic16m": 1, "periodic2m": 2, "periodic256k": 4}


def setup_slab_run(mode, scale, world, rank, local, accel=True, transport="nccl", dist=None):
    """One rank's share of a slab-decomposed workload: (engine, SlabRank, own host records, neighbours, description).
    mode: "sections" = config-4 pattern, one tank section per GPU (weak); "strong" = the ONE config-4 tank cut into `world`
    slabs of equal owned-plus-ghost counts; "periodic" = config 5, a fully periodic box per GPU, `world` side by side in x."""
    from . import scenes
    R = support_radius()
    halo = 2.0 * R * 1.001
    if mode == "periodic":
        side = 256 // scale
        L = side * 0.1
        pos, vel = scenes.lattice_block(side, side, side, 0.1, (0.05 + L * rank, 0.05, 0.05), seed=1234 + rank, jitter=0.2, vel_amp=0.5)
        ps = scenes.make_particles(pos, vel)
        del pos, vel
        bounds = [L * r for r in range(world + 1)]
        slab = make_slab(rank, world, bounds, halo, period=L * world, yz_box=((0.0, L), (0.0, L)))
        left, right = neighbours(rank, world, periodic=True)
        m = halo + 0.35
        dlo, dhi = (bounds[rank] - m, -m, -m), (bounds[rank + 1] + m, L + m, L + m)
        eng = capi.Engine(device=local, domain_mode=capi.DOMAIN_FIXED, domain_lo=dlo, domain_hi=dhi, gravity=(0.0, 0.0, 0.0),
                          write_accel=1 if accel else 0)
        n_own = len(ps)
        capacity = int((L + 2 * halo) ** 2 * 1100.0 * halo * 1.6) + 4096
        eng.reserve(int(n_own * ((L + 2 * halo) / L) ** 3 * 1.08) * 2 - n_own + 4 * capacity)   # old ghosts stay (dead) until the next sort
        what = f"weak-scaling periodic box {L:.1f}^3 of {n_own} particles per GPU ({world} side by side in x), gravity off, no walls (config 5)"
    else:
        section = 256 // scale * 0.1                     # x-length of one tank section
        strong = mode == "strong"
        if strong:
            ps, walls = scenes.block_drop(scale=scale)
            bounds = balanced_bounds(ps["pos"][:, 0], world, R, 0.0, section + 0.2, halo)
            ps = ps[(ps["pos"][:, 0] >= (bounds[rank] if rank else -1e30)) & (ps["pos"][:, 0] < (bounds[rank + 1] if rank < world - 1 else 1e30))].copy()
        else:
            ps = scenes.block_drop_section(rank, world, scale=scale)
            walls = scenes.tank_walls(section * world + 0.2, section + 0.2)
            bounds = [0.1 + section * r for r in range(world + 1)]
        n_own = len(ps)
        slab = make_slab(rank, world, bounds, halo)
        left, right = neighbours(rank, world)
        top = 0.1 + (192 + 256) // scale * 0.1 + 0.3 + 1.0
        dlo = (bounds[rank] - halo - 0.35, -0.25, -0.25)
        dhi = (bounds[rank + 1] + halo + 0.35, top, section + 0.45)
        eng = capi.Engine(device=local, domain_mode=capi.DOMAIN_FIXED, domain_lo=dlo, domain_hi=dhi, write_accel=1 if accel else 0)
        eng.set_walls(walls)
        column = (192 + (256 if strong else 0)) // scale * 0.1 + 0.5        # a strong-scaling cut may pass through the falling block
        face = (section + 0.2) * column * 1100.0                             # particles per unit x near a face (upper bound)
        capacity = int(face * halo * 2.0) + 4096
        eng.reserve(n_own + 4 * capacity)   # last step's ghosts stay (marked dead) until the next sort drops them
        what = (f"block drop into tank (config 4), ONE tank of {2 ** 24 // scale ** 3} particles cut into {world} x-slabs of equal owned-plus-ghost counts" if strong
                else f"block drop into tank (config 4 pattern), one tank section of {n_own} particles per GPU ({world} sections side by side)")
    sr = SlabRank(eng, slab, capacity, local, transport=transport)
    if transport == "p2p":
        # every rank must end up on the same transport: if mapping a neighbour's block fails anywhere (no peer access, IPC
        # forbidden in this container, ...), all ranks fall back to NCCL message buffers together
        ok = 1
        try:
            sr.connect(dist, rank, world, left, right)
        except capi.SphbError as err:
            ok = 0
            print(f"[rank {rank}] peer-to-peer halo link unavailable ({err}); falling back to NCCL send/recv", flush=True)
        if world > 1:
            import torch
            flag = torch.tensor([ok], device="cuda", dtype=torch.int32)
            dist.all_reduce(flag, op=dist.ReduceOp.MIN)
            ok = int(flag.item())
        if not ok:
            sr.close()
            sr = SlabRank(eng, slab, capacity, local, transport="nccl")
    return eng, sr, ps, (left, right), what, capacity


def run_slab_workload(args, mode, scale, world, rank, local, dist, steps, warmup, e2e_steps=0, sample_clocks=False, reset_every=16):
    """Times `steps` slab-decomposed steps (pack -> exchange -> unpack -> engine step) after `warmup`: CUDA events on the engine's
    stream between two barriers + synchronisations, max over ranks.  Returns a dict; every rank gets the same timing values."""
    import time

    import torch

    import bench as B

    accel = not args.no_accel
    transport = getattr(args, "transport", "nccl")
    eng, sr, ps, (left, right), what, capacity = setup_slab_run(mode, scale, world, rank, local, accel, transport, dist)
    n_own = len(ps)
    host = torch.empty(n_own * 52, dtype=torch.uint8).pin_memory()
    host_np = host.numpy().view(capi.PARTICLE)
    host_np[:] = ps
    del ps
    eng.upload(host_np)
    eng.update_cells()
    init = [torch.from_numpy(a).cuda() for a in eng.download_soa()[:2]]   # this rank's initial owned state, on the device
    init[1][:, 3] = 1.0
    since_reset = [0]
    xdist = dist      # world size 1: the _Solo shim (a lone rank in a periodic box is its own neighbour and copies locally)

    def one_step():
        # the reference's scheme blows up on deep pools after ~35 steps (bench.py --reset-every): start over in time
        if reset_every > 0 and since_reset[0] >= reset_every:
            eng.upload_soa_device(init[0].data_ptr(), init[1].data_ptr(), n_own)
            since_reset[0] = 0
        sr.step(xdist, left, right)
        since_reset[0] += 1

    for _ in range(max(3, warmup)):
        one_step()
    eng.synchronize()
    launches0 = eng.launch_count()
    eng.enable_timing(True)
    sampler = B.ClockSampler(local) if (sample_clocks and rank == 0) else None
    start, stop = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    dist.barrier()
    torch.cuda.synchronize()
    start.record(sr.stream)
    for _ in range(steps):
        one_step()
    stop.record(sr.stream)
    eng.synchronize()
    torch.cuda.synchronize()
    dist.barrier()
    ms = torch.tensor([start.elapsed_time(stop)], device="cuda")
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    ms = float(ms.item())
    clocks = sampler.stop() if sampler else None
    launches = eng.launch_count() - launches0
    phase, timed = eng.get_timing()
    eng.enable_timing(False)
    bad = eng.count_nonfinite()
    assert bad == 0, f"rank {rank}: {bad} non-finite particles after the timed steps"
    _, vel = sr.owned()
    cnt = torch.tensor([len(vel), n_own, int(launches)], device="cuda", dtype=torch.int64)
    step_max = torch.tensor([float(phase[3])], device="cuda")
    if world > 1:
        dist.all_reduce(cnt)
        dist.all_reduce(step_max, op=dist.ReduceOp.MAX)
    total, expect, launches_all = (int(x) for x in cnt.tolist())
    assert total == expect, f"particles lost or duplicated by the exchange: {total} owned, {expect} generated"
    value = total * steps / (ms * 1e-3)

    # where the time between two engine steps goes (a few extra steps, events on the engine's stream; not part of `value`)
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(4)]
    parts = np.zeros(3)
    nb = 4
    for _ in range(nb):
        if sr.transport == "p2p":           # one call: pack into the neighbours' buffers, flags, append; only the whole is timed
            sr.poll()
            ev[0].record(sr.stream)
            sr.exchange_p2p()
            ev[3].record(sr.stream)
            eng.step(1)
            since_reset[0] += 1
            eng.synchronize()
            parts += [ev[0].elapsed_time(ev[3]), 0.0, 0.0]
            continue
        sr.poll()
        ev[0].record(sr.stream)
        sr.pack()
        ev[1].record(sr.stream)
        sr.exchange_distributed(xdist, left, right)
        ev[2].record(sr.stream)
        sr.unpack(sync=True)
        ev[3].record(sr.stream)
        eng.step(1)
        since_reset[0] += 1
        eng.synchronize()
        parts += [ev[0].elapsed_time(ev[1]), ev[1].elapsed_time(ev[2]), ev[2].elapsed_time(ev[3])]
    parts /= nb

    out = {"workload": what, "particles": total, "particles_per_gpu_rank0": n_own, "n_gpus": world, "steps": steps, "warmup": max(3, warmup),
           "ms_per_step": ms / steps, "value": value, "gpu_launches": launches_all,
           "phase_ms": {"sort": float(phase[0]), "density": float(phase[1]), "force": float(phase[2]), "engine_step_rank0": float(phase[3]),
                        "engine_step_max_over_ranks": float(step_max.item()),
                        "breakdown_rank0": ({"exchange_p2p (pack + flags + append)": float(parts[0])} if sr.transport == "p2p" else
                                            {"pack": float(parts[0]), "send_recv": float(parts[1]), "unpack": float(parts[2])})},
           "transport": sr.transport,
           "exchange_exposed_ms": ms / steps - float(step_max.item()),
           "halo": {"counts_rank0": {"kept": sr.counts[0], "from_left": sr.counts[1], "from_right": sr.counts[2]}, "message_capacity": capacity},
           "local_particles_incl_ghosts_rank0": eng.count(), "clocks": clocks, "_phase": phase}

    if e2e_steps:
        # end to end on host records: upload own records, exchange ghosts, step, download own records
        res = torch.empty(n_own * 52, dtype=torch.uint8).pin_memory()
        res_np = res.numpy().view(capi.PARTICLE)

        def e2e_step():
            eng.upload(host_np)             # the rank's own 52-byte records (pinned)
            sr.step(xdist, left, right)     # ghosts come from the neighbours
            eng.download_first(res_np)      # own records back (ghosts are not returned)

        e2e_step()
        dist.barrier()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for _ in range(e2e_steps):
            e2e_step()
        torch.cuda.synchronize()
        t = torch.tensor([(time.perf_counter() - t0) / e2e_steps], device="cuda")
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        t_e2e = float(t.item())
        out["e2e"] = {"value": total / t_e2e, "unit": B.UNIT, "h2d_bytes_per_step": expect * 52, "d2h_bytes_per_step": expect * 52,
                      "ms_per_step": 1e3 * t_e2e, "call": "per rank: sphb_upload_aos + halo exchange + sphb_step + sphb_download_aos"}
    dist.barrier()          # nobody frees a block a neighbour may still be writing flags into
    sr.close()
    eng.close()
    del init, host_np, host
    torch.cuda.empty_cache()
    return out


def parity_vs_single(args, world, rank, local, dist, steps=12, scale=4):
    """Hardware check of the NCCL path (SURVEY.md 8(e) "Check"): the config-4 scene at 262,144 particles stepped `steps` times through
    SlabRank.step on all ranks AND through one engine on rank 0; owned particles gathered to rank 0 and paired with the single
    engine's by position.  Same bounds as tests/test_gpu_multirank.py: positions <= 2e-5, density <= 1e-5 relative."""
    import torch

    from . import scenes
    eng, sr, ps, (left, right), _, _ = setup_slab_run("strong", scale, world, rank, local, True, getattr(args, "transport", "nccl"), dist)
    eng.upload(ps)
    xdist = dist
    for _ in range(steps):
        sr.step(xdist, left, right)
    posm, vel, aux = eng.download_soa()[:3]
    keep = vel[:, 3] != 0
    mine = (posm[keep], aux[keep, 0].copy())
    dist.barrier()
    sr.close()
    eng.close()
    gathered = [None] * world if rank == 0 else None
    if world > 1:
        dist.gather_object(mine, gathered, dst=0)
    else:
        gathered = [mine]
    out = None
    if rank == 0:
        from scipy.spatial import cKDTree
        full, walls = scenes.block_drop(scale=scale)
        section = 256 // scale * 0.1
        top = 0.1 + (192 + 256) // scale * 0.1 + 0.3 + 1.0
        with capi.Engine(device=local, domain_mode=capi.DOMAIN_FIXED, domain_lo=(-0.25, -0.25, -0.25), domain_hi=(section + 0.45, top, section + 0.45)) as e1:
            e1.set_walls(walls)
            e1.upload(full)
            e1.step(steps)
            p1, _, a1 = e1.download_soa()[:3]
        pm = np.concatenate([g[0] for g in gathered])
        rm = np.concatenate([g[1] for g in gathered])
        ok_count = len(pm) == len(p1)
        d, idx = cKDTree(pm[:, :3].astype(np.float64)).query(p1[:, :3].astype(np.float64))
        one_to_one = len(np.unique(idx)) == len(p1)
        drho = np.abs(rm[idx].astype(np.float64) - a1[:, 0]) / a1[:, 0]
        out = {"particles": int(len(p1)), "steps": steps, "count_conserved": bool(ok_count), "pairing_one_to_one": bool(one_to_one),
               "compressed_fraction": float((a1[:, 0] > 10.0).mean()),      # > 0: the falling block has hit the pool, pressure forces are active
               "max_dpos": float(d.max()), "max_drho_rel": float(drho.max()), "bounds": {"max_dpos": 2e-5, "max_drho_rel": 1e-5}}
        assert ok_count and one_to_one and out["max_dpos"] <= 2e-5 and out["max_drho_rel"] <= 1e-5, f"multi-GPU result differs from the single engine: {out}"
    torch.cuda.empty_cache()
    return out


def n1_baselines(args, scale, local, rank, dist, steps=8):
    """Rank 0 alone runs both named configs on ONE GPU (the others wait): the denominators of efficiency_vs_n1."""
    import bench as B
    out = None
    if rank == 0:
        name = {1: "drop16m", 2: "drop2m", 4: "drop256k"}[scale]
        r = B.run_resident(name, steps, 3, local, accel=not args.no_accel)
        p = run_slab_workload(args, "periodic", scale, 1, 0, local, _Solo, steps, 3)
        out = {"config4_ms_per_step": r["ms_per_step"], "config4_value": r["value"], "config5_ms_per_step": p["ms_per_step"], "config5_value": p["value"]}
    dist.barrier()
    return out


def strip(d):
    return {k: v for k, v in d.items() if not k.startswith("_")}


def bench_main(args, world, rank, local):
    import json

    import torch
    import torch.distributed as tdist

    import bench as B

    torch.cuda.set_device(local)
    dist = _Solo
    if world > 1:
        tdist.init_process_group("nccl", device_id=torch.device("cuda", local))
        dist = tdist
    periodic = args.workload.startswith("periodic")
    scale = SCALE_OF.get(args.workload)
    if scale is None:
        raise SystemExit("multi-GPU bench runs the block-drop (drop16m / drop2m / drop256k) and periodic workloads")
    mode = "periodic" if periodic else ("strong" if getattr(args, "scaling", "weak") == "strong" else "sections")
    reset = args.reset_every if args.reset_every is not None else 16
    head = run_slab_workload(args, mode, scale, world, rank, local, dist, args.steps, args.warmup, e2e_steps=max(1, args.e2e_steps),
                             sample_clocks=True, reset_every=reset)
    sub = {}
    if not getattr(args, "quick", False):
        k = max(4, min(args.steps, 10))
        n1 = n1_baselines(args, scale, local, rank, dist) if world > 1 else None
        for key, m in (("config4_strong", "strong"), ("config5_periodic", "periodic")):
            if m == mode:
                r = head
            else:
                r = run_slab_workload(args, m, scale, world, rank, local, dist, k, 3, reset_every=reset)
            rec = strip(r)
            rec["scaling"] = "strong" if m == "strong" else "weak"
            if n1 is not None:
                base = n1["config4_value"] if m == "strong" else n1["config5_value"]
                rec["n1_value_same_run_rank0"] = base
                rec["efficiency_vs_n1"] = r["value"] / (world * base)   # aggregate throughput over N x the one-GPU throughput (weak and strong alike)
            sub[key] = rec
        sub["parity_vs_single"] = parity_vs_single(args, world, rank, local, dist)
    if rank == 0:
        peak, peak_src = B.measured_peak()
        phase = head["_phase"]
        n_local = head["local_particles_incl_ghosts_rank0"]
        rl = B.rooflines(n_local, phase, not args.no_accel, head["clocks"].get("sm_mhz") if head["clocks"] else None)
        rl["roofline"]["kernel"] = "k_force (rank 0, local particles incl. ghosts)"
        line = {
            "metric": B.METRIC, "value": head["value"], "unit": B.UNIT, "n_gpus": world, "steps": args.steps, "warmup": max(3, args.warmup),
            "ms_per_step": head["ms_per_step"], "higher_is_better": True, "scaling": "strong" if mode == "strong" else "weak", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic",
            "config": {"workload": head["workload"], "particles": head["particles"], "particles_per_gpu": head["particles_per_gpu_rank0"],
                       "decomposition": "x slabs, halo = 2 support radii, one exchange with each x-neighbour per step",
                       "transport": ("pack kernel writes into the neighbours' receive buffers over NVLink peer memory (CUDA IPC), flag kernels"
                                     if head.get("transport") == "p2p" else "NCCL send/recv of message buffers (torch.distributed)"),
                       "l2": "state exceeds the 126 MB L2", "accel_output": not args.no_accel, "reset_every": reset},
            "clocks": head["clocks"], "e2e": head.get("e2e"), "gpu_launches": head["gpu_launches"],
            "phase_ms": head["phase_ms"], "exchange_exposed_ms": head["exchange_exposed_ms"], "halo": head["halo"],
            "cpu_baseline": None,
        }
        line.update(rl)
        line.update(sub)
        print(json.dumps(line, default=float), flush=True)
    dist.barrier()
    if world > 1:
        tdist.destroy_process_group()
