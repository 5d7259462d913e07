#!/usr/bin/env python
"""Headline benchmark: particle-updates/s of the SPH step (grid sort + density + force/integrate/collide).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload drop16m|dam1m|dam64k|...]
    python bench.py --impl reference ...       # the reference's own CPU loop, same JSON shape

One JSON line on stdout (rank 0).  A "step" is one pass of the hot path -- what step_scene does per frame
(/root/reference SPH/main.cpp:293-300) -- over the whole synthetic particle set.

* value    : N_particles * K / t, state resident in HBM (uploaded once, before the timed region), timed
             with CUDA events between two synchronisations, max over ranks.
* e2e      : same metric through the reference-facing call on HOST records (sphb_step_aos = upload of the
             52-byte Particle records from pinned memory + one step + download), copies inside the timing.
* roofline : dominant kernel (k_force: pair force + integrate + wall collision, fused); achieved =
             88 B * N / its average CUDA-event duration over the timed steps (algorithmic bytes).
* cpu_baseline : the oracle port (cell-list restatement, OpenMP, all host cores) on a bounded sample.
Inputs are far larger than the 126 MB L2 at the default workload (16.7 M particles ~ 2 GB of state), so
no explicit L2 flush is needed; config.l2 says which applies.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import tempfile
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "particle-updates/sec (density+force+integrate)"
UNIT = "particle-updates/s"

# algorithmic bytes per particle-update, SURVEY.md 8(d) / DESIGN.md
B_DENSITY = 24      # read posm 16, write (rho,P) 8
B_FORCE = 88        # read posm 16 + vel 16 + (rho,P) 8; write posm 16 + vel 16 + accel 16
B_FORCE_NOACC = 72
B_FUSED = B_DENSITY + B_FORCE   # 112
B_STEP = 196        # incl. key / counting sort / reorder

WORKLOADS = {
    # name: (description, generator kwargs)
    "default972": "reference default main.cpp scene, 972 particles (config 1)",
    "dam64k": "3D dam break, 65,536 particles, stressed (config 2)",
    "dam1m": "3D dam break, 1,048,576 particles (config 3)",
    "drop16m": "block drop into tank, 16,777,216 particles (config 4)",
    "drop2m": "block drop into tank, 2,097,152 particles (config 4 at scale 2)",
    "drop256k": "block drop into tank, 262,144 particles (config 4 at scale 4)",
    "periodic16m": "weak-scaling periodic box, 16,777,216 particles per GPU (config 5)",
    "periodic2m": "weak-scaling periodic box, 2,097,152 particles per GPU (config 5 at scale 2)",
    "periodic256k": "weak-scaling periodic box, 262,144 particles per GPU (config 5 at scale 4)",
}


def make_workload(name):
    from sph_b200 import scenes
    if name == "default972":
        ps, walls = scenes.default_scene()
        return ps, walls, ((-1.3, -1.3, -1.3), (1.3, 3.0, 1.3))
    if name == "dam64k":
        ps, walls = scenes.dam_break(32, 64, 32, stressed=True)
        return ps, walls, scenes.dam_break_domain(32, 64, 32, stressed=True)
    if name == "dam1m":
        ps, walls = scenes.dam_break(128, 64, 128)
        return ps, walls, scenes.dam_break_domain(128, 64, 128)
    if name.startswith("drop"):
        scale = {"drop16m": 1, "drop2m": 2, "drop256k": 4}[name]
        ps, walls = scenes.block_drop(scale=scale)
        t = 256 // scale * 0.1 + 0.2
        top = 0.1 + (192 + 256) // scale * 0.1 + 0.3 + 1.0
        return ps, walls, ((-0.25, -0.25, -0.25), (t + 0.25, top, t + 0.25))
    raise SystemExit(f"unknown workload {name}")


class ClockSampler:
    """nvidia-smi clocks/throttle reasons during the timed region (B200_PROFILING.md recipe)."""

    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.f = tempfile.NamedTemporaryFile("w+", suffix=".csv", delete=False)
        self.p = None
        try:
            self.p = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "100",
                                       "-i", str(gpu_index)], stdout=self.f, stderr=subprocess.DEVNULL)
        except OSError:
            pass

    def stop(self):
        out = {"sm_mhz": None, "sm_max_mhz": None, "reasons": []}
        if self.p is None:
            return out
        time.sleep(0.15)
        self.p.terminate()
        try:
            self.p.wait(timeout=5)
        except Exception:
            self.p.kill()
        self.f.flush()
        self.f.seek(0)
        sm, mx, reasons = [], [], set()
        for line in self.f.read().splitlines():
            c = [x.strip() for x in line.split(",")]
            if len(c) < 9:
                continue
            try:
                sm.append(float(c[1])); mx.append(float(c[2]))
            except ValueError:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), c[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        self.f.close()
        os.unlink(self.f.name)
        if sm:
            out = {"sm_mhz": float(np.median(sm)), "sm_max_mhz": float(max(mx)), "reasons": sorted(reasons), "samples": len(sm)}
        return out


def measured_peak():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        try:
            return float(json.load(open(path))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md)"


def cpu_baseline_port(sample_side=144, steps=8):
    """The oracle's cell-list restatement (bit-identical to the reference's loops) on all host cores, on a bounded
    sample of the same kind of fluid: a jittered sample_side^3 block at the bench's spacing, `steps` steps
    (sized for roughly 10 s of CPU work)."""
    from oracle.oracle import Oracle
    from sph_b200 import scenes
    orc = Oracle()
    ps, walls = scenes.dam_break(sample_side, sample_side, sample_side)
    threads = orc.max_threads()
    t0 = time.perf_counter()
    orc.step(ps, walls, steps, cells=True, threads=threads)
    dt = time.perf_counter() - t0
    return {"value": len(ps) * steps / dt, "unit": UNIT, "cores": threads, "kind": "port",
            "sample": f"{steps} steps of a jittered {sample_side}^3 = {len(ps)} particle block (same spacing/jitter as the workload), "
                      f"oracle cell-list restatement with OpenMP on {threads} threads, {dt:.1f} s"}


def run_reference_arm(args):
    """--impl reference: the reference's own CPU implementation (unmodified SphUtils.cpp, all-pairs, 1 thread -- it has
    no threading) when oracle/_ref was built, else the oracle port; bounded sample per step."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from oracle.oracle import Oracle, Reference
    from sph_b200 import scenes
    side = 24       # 13,824 particles: ~6 s per all-pairs step
    ps, walls = scenes.dam_break(side, side, side)
    if Reference.available():
        eng, kind, cores = Reference(), "reference", 1
        stepper = lambda: eng.step(ps, walls, 1)
        what = "unmodified reference SphUtils.cpp (all-pairs O(N^2), single thread: it has none)"
    else:
        eng, kind = Oracle(), "port"
        cores = eng.max_threads()
        stepper = lambda: eng.step(ps, walls, 1, cells=True, threads=cores)
        what = "oracle port (cell list, OpenMP)"
    steps = max(1, min(args.steps, 3))
    warm = max(0, min(args.warmup, 1))
    for _ in range(warm):
        stepper()
    t0 = time.perf_counter()
    for _ in range(steps):
        stepper()
    dt = time.perf_counter() - t0
    val = len(ps) * steps / dt
    sample = f"{steps} step(s) of a jittered {side}^3 = {len(ps)} particle block, {what}; all-pairs cost grows as N^2, so the " \
             f"full workload would be far slower per particle"
    # the reference's OWN config (config 1: its built-in 972-particle scene) on the same code path: the one workload both arms can run
    # whole, so that bench.py's `configs.default972` / `config1_vs_reference` have a same-config counterpart in this line
    ps1, walls1 = scenes.default_scene()
    k1 = 100
    t0 = time.perf_counter()
    if kind == "reference":
        eng.step(ps1, walls1, k1)
    else:
        eng.step(ps1, walls1, k1, cells=False)
    dt1 = time.perf_counter() - t0
    config1 = {"workload": WORKLOADS["default972"], "particles": len(ps1), "steps": k1, "ms_per_step": 1e3 * dt1 / k1, "value": len(ps1) * k1 / dt1}
    line = {"impl": "reference", "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": args.gpus, "steps": steps, "warmup": warm,
            "configs": {"default972": config1},
            "ms_per_step": 1e3 * dt / steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
            "data": "synthetic", "config": {"workload": WORKLOADS[args.workload], "sample_particles": len(ps)},
            "cpu_baseline": {"value": val, "unit": UNIT, "cores": cores, "kind": kind, "sample": sample},
            "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}, "gpu_launches": 0}
    print(json.dumps(line, default=float), flush=True)


def issue_profile():
    try:
        return json.load(open(os.path.join(ROOT, "profiles", "issue.json")))
    except Exception:
        return None


def rooflines(n, phase, accel, sm_mhz=None):
    """The three HBM rooflines (contract: algorithmic bytes / kernel time / measured copy peak) plus the roofline of what
    actually binds these kernels: instruction issue and L1 data-pipe (LSU) wavefronts (SURVEY.md 8(d) "Binding roofline")."""
    peak, peak_src = measured_peak()
    bforce = B_FORCE if accel else B_FORCE_NOACC
    ach = bforce * n / (phase[2] * 1e-3) / 1e9
    traffic, traffic_src = None, None
    try:
        tj = json.load(open(os.path.join(ROOT, "profiles", "traffic.json")))["k_force"]
        traffic = tj["dram_bytes_per_particle"] * n / (phase[2] * 1e-3) / 1e9      # GB/s, per launch like `achieved`
        traffic_src = tj["source"] + "; per-particle figure scaled by this run's particle count"
    except Exception:
        pass
    out = {}
    out["roofline"] = {"bound": "hbm", "kernel": "k_force (pair force + integrate + wall collision)", "achieved": ach, "peak": peak,
                       "unit": "GB/s", "frac": ach / peak, "traffic": traffic, "traffic_source": traffic_src, "peak_source": peak_src,
                       "bytes_per_particle": bforce, "kernel_ms": float(phase[2])}
    fused_ms = float(phase[1] + phase[2])
    ach_f = (B_DENSITY + bforce) * n / (fused_ms * 1e-3) / 1e9
    out["roofline_fused"] = {"what": "density + force/integrate/collide kernels together (north_star's fused step)",
                             "bytes_per_particle": B_DENSITY + bforce, "achieved": ach_f, "frac": ach_f / peak, "ms": fused_ms}
    ach_s = B_STEP * n / (float(phase[3]) * 1e-3) / 1e9
    out["roofline_step"] = {"what": "whole step incl. grid/sort/reorder", "bytes_per_particle": B_STEP, "achieved": ach_s,
                            "frac": ach_s / peak, "ms": float(phase[3])}
    ip = issue_profile()
    if ip and sm_mhz:
        lanes_per_s = 148 * 4 * 32 * sm_mhz * 1e6          # SMs x schedulers x lanes x measured SM clock: thread-instructions/s at 1 warp-instruction per scheduler and cycle
        ri = {"what": "thread-instructions per particle (ncu) x particles / kernel time, against 148 SMs x 4 schedulers x 32 lanes x the SM clock "
                      "sampled during this run; lsu_wavefront_frac and the stall note are the profile's", "unit": "thread-instructions/s",
              "peak": lanes_per_s, "sm_mhz": sm_mhz, "source": ip.get("source")}
        for kname, ms, bound in (("k_density", float(phase[1]), "issue + lsu + fp32 pipe (co-limited)"), ("k_force", float(phase[2]), "lsu (L1 data-pipe wavefronts)")):
            k = ip.get(kname)
            if not k or ms <= 0:
                continue
            a = k["thread_inst_per_particle"] * n / (ms * 1e-3)
            ri[kname] = {"bound": bound, "achieved": a, "frac": a / lanes_per_s, "thread_inst_per_particle": k["thread_inst_per_particle"],
                         "lsu_wavefront_frac": k["lsu_wavefront_frac_of_peak"], "kernel_ms": ms, "top_stall": k.get("top_stall"),
                         "floor_ms_at_full_issue": k["thread_inst_per_particle"] * n / lanes_per_s * 1e3}
        out["roofline_issue"] = ri
    return out


def run_resident(name, steps, warmup, local, accel=True, domain="fixed", cell_x_div=0, cell_frac=1.0, reset_every=None, sample_clocks=False,
                 keep=False):
    """State-resident throughput of one workload on one GPU: W warm-up steps, K timed steps (CUDA events on the engine's
    stream, synchronised on both sides), the reference scheme's blow-up avoided by restoring the initial state every
    `reset_every` steps inside the timed region.  Returns a dict (and the engine + host records when keep=True)."""
    import torch
    import sph_b200
    ps, walls, (dlo, dhi) = make_workload(name)
    n = len(ps)
    if reset_every is None:
        reset_every = 4 if name == "dam64k" else 16   # the pre-compressed 64 K block explodes within a dozen steps, in the reference too
    kw = dict(write_accel=1 if accel else 0)
    xdiv = cell_x_div if cell_x_div > 0 else 2
    cell = float(np.float32(0.45) * np.float32(0.45)) * cell_frac
    if domain == "fixed":
        ncell = ((dhi[0] - dlo[0]) / (cell / xdiv) + 2) * ((dhi[1] - dlo[1]) / cell + 2) * ((dhi[2] - dlo[2]) / cell + 2)
        kw.update(max_cells=int(ncell) + 4096, domain_mode=sph_b200.DOMAIN_FIXED, domain_lo=dlo, domain_hi=dhi)
    kw.update(cell_size=cell if cell_frac != 1.0 else 0.0, cell_x_div=cell_x_div)
    eng = sph_b200.Engine(device=local, **kw)
    eng.set_walls(walls)
    host = torch.empty(n * 52, dtype=torch.uint8).pin_memory()   # host records in pinned memory (torch is plumbing: it owns the page-locked buffer)
    host_np = host.numpy().view(sph_b200.PARTICLE)
    host_np[:] = ps
    del ps
    eng.upload(host_np)
    eng.update_cells()
    init = [torch.from_numpy(a).cuda() for a in eng.download_soa()[:2]]   # grid-sorted initial state, kept on the device
    init[1][:, 3] = 1.0
    since = [0]

    def step_once():
        if reset_every > 0 and since[0] >= reset_every:
            eng.upload_soa_device(init[0].data_ptr(), init[1].data_ptr(), n)
            since[0] = 0
        eng.step(1)
        since[0] += 1

    for _ in range(warmup):
        step_once()
    eng.synchronize()
    launches0 = eng.launch_count()
    eng.enable_timing(True)
    sampler = ClockSampler(local) if sample_clocks else None
    start, stop = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    stream = torch.cuda.ExternalStream(eng.stream(), device=local)
    torch.cuda.synchronize()
    start.record(stream)
    for _ in range(steps):
        step_once()
    stop.record(stream)
    eng.synchronize()
    torch.cuda.synchronize()
    ms = start.elapsed_time(stop)
    clocks = sampler.stop() if sampler else None
    launches = eng.launch_count() - launches0
    phase, timed = eng.get_timing()
    eng.enable_timing(False)
    bad = eng.count_nonfinite()      # a blown-up state would make the timing meaningless
    assert bad == 0, f"{name}: {bad} non-finite particles after the timed steps: lower --reset-every"
    out = {"workload": WORKLOADS[name], "particles": n, "steps": steps, "warmup": warmup, "ms_per_step": ms / steps, "value": n * steps / (ms * 1e-3),
           "domain": domain, "reset_every": reset_every, "cell_x_div": xdiv, "accel_output": accel, "gpu_launches": int(launches),
           "phase_ms": {"sort": float(phase[0]), "density": float(phase[1]), "force": float(phase[2]), "step": float(phase[3])},
           "steps_averaged": int(timed), "clocks": clocks, "arrival_order_cells": eng.stats()["arrival_order_cells"]}
    out["_phase"] = phase
    if keep:
        out["_keep"] = (eng, host_np, host, init)
    else:
        eng.close()
    return out


def config1_vs_reference(local, ours_steps=100):
    """Config 1 on BOTH sides: the reference's own 972-particle scene (main.cpp:490-540), the unmodified reference
    (oracle/_ref, its all-pairs loops, one thread: it has no threading) for 100 steps on this box's host, and this
    engine as the drop-in runs it (bounding-box grid, acceleration output on) for 100 resident steps after 10 warm-up."""
    from oracle.oracle import Oracle, Reference
    from sph_b200 import scenes
    ps, walls = scenes.default_scene()
    ref_steps = 100
    if Reference.available():
        eng, kind = Reference(), "reference (unmodified SphUtils.cpp, all-pairs, 1 thread)"
        t0 = time.perf_counter()
        eng.step(ps.copy(), walls, ref_steps)
    else:
        eng, kind = Oracle(), "port (oracle all-pairs restatement, 1 thread)"
        t0 = time.perf_counter()
        eng.step(ps.copy(), walls, ref_steps, cells=False)
    ref_ms = 1e3 * (time.perf_counter() - t0) / ref_steps
    ours = run_resident("default972", ours_steps, 10, local, domain="bbox", reset_every=0)
    us = 1e3 * ours["ms_per_step"]
    return {"particles": len(ps), "ours_us_per_step": us, "ours_steps": ours_steps, "reference_ms_per_step": ref_ms, "reference_steps": ref_steps,
            "reference_kind": kind, "ratio": ref_ms * 1e3 / us, "ours_phase_ms": ours["phase_ms"]}


def strip(d):
    return {k: v for k, v in d.items() if not k.startswith("_")}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="drop16m", choices=sorted(WORKLOADS))
    ap.add_argument("--e2e-steps", type=int, default=3)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-accel", action="store_true", help="skip the per-particle acceleration output (72 B force kernel)")
    ap.add_argument("--domain", default="fixed", choices=["fixed", "bbox"])
    ap.add_argument("--scaling", default="weak", choices=["weak", "strong"],
                    help="multi-GPU headline: weak = one 16.7 M-particle tank section per GPU (default); strong = the ONE config-4 tank "
                         "cut into N slabs.  Both named configs (config 4 strong, config 5 periodic) are measured as sub-records either way")
    ap.add_argument("--reset-every", type=int, default=None,
                    help="restore the initial state (device-to-device, inside the timed region) every this many steps: the "
                         "reference's scheme (dt=0.033, k=100) blows up on deep pools after ~35 steps, in its own arithmetic too "
                         "(default 16; 4 for the pre-compressed 64 K dam break, which explodes within a dozen steps)")
    ap.add_argument("--cell-frac", type=float, default=1.0, help="grid cell edge as a fraction of the support radius h*h (>= 0.5)")
    ap.add_argument("--cell-x-div", type=int, default=0, help="x pitch of the grid = cell / this (0 = engine default, 2)")
    ap.add_argument("--transport", default="p2p", choices=["p2p", "nccl"],
                    help="multi-GPU halo transport: p2p = the pack kernel writes into the neighbours' receive buffers over NVLink peer memory "
                         "(CUDA IPC between the ranks, flags instead of a communication call); nccl = message buffers exchanged with NCCL send/recv")
    ap.add_argument("--quick", action="store_true", help="headline only: skip the sub-records (other configs, reference comparison, bbox mode)")
    args = ap.parse_args()
    if args.impl == "reference":
        return run_reference_arm(args)
    if args.warmup < 3:
        args.warmup = 3

    import torch
    import sph_b200

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world > 1 or args.workload.startswith("periodic"):
        from sph_b200 import multigpu
        return multigpu.bench_main(args, world, rank, local)
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: sph_b200 has no CPU path")
    torch.cuda.set_device(local)

    # ---- headline: resident-state throughput
    head = run_resident(args.workload, args.steps, args.warmup, local, accel=not args.no_accel, domain=args.domain, cell_x_div=args.cell_x_div,
                        cell_frac=args.cell_frac, reset_every=args.reset_every, sample_clocks=True, keep=True)
    eng, host_np, host, init = head["_keep"]
    n = head["particles"]
    phase = head["_phase"]
    clocks = head["clocks"]

    # ---- end to end through the reference-facing call on host records
    e2e_steps = max(1, args.e2e_steps)
    eng.step_aos(host_np, 1)            # warm the staging buffers
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        eng.step_aos(host_np, 1)        # upload + step + download, synchronised
    t_e2e = (time.perf_counter() - t0) / e2e_steps
    e2e = {"value": n / t_e2e, "unit": UNIT, "h2d_bytes_per_step": n * 52, "d2h_bytes_per_step": n * 52,
           "ms_per_step": 1e3 * t_e2e, "call": "sphb_step_aos (52-byte Particle records, pinned host memory, SPHB_F_ALL)"}
    # positions only on the way back (what main.cpp's viewer reads, main.cpp:97-110)
    eng.step_aos(host_np, 1, mask=sph_b200.F_POS)
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        eng.step_aos(host_np, 1, mask=sph_b200.F_POS)
    t_pos = (time.perf_counter() - t0) / e2e_steps
    e2e["pos_only"] = {"value": n / t_pos, "unit": UNIT, "ms_per_step": 1e3 * t_pos, "call": "sphb_step_aos with SPHB_F_POS"}

    # ---- the viewer's loop (main.cpp idle -> step_scene -> update_positions): state resident, positions back every step
    xyz = torch.empty(n * 3, dtype=torch.float32).pin_memory()
    eng.upload_soa_device(init[0].data_ptr(), init[1].data_ptr(), n)
    eng.step(1); eng.download_positions(ptr=xyz.data_ptr(), n=n)
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        eng.step(1)
        eng.download_positions(ptr=xyz.data_ptr(), n=n)
    t_view = (time.perf_counter() - t0) / e2e_steps
    e2e["viewer_loop"] = {"value": n / t_view, "unit": UNIT, "ms_per_step": 1e3 * t_view, "h2d_bytes_per_step": 0,
                          "d2h_bytes_per_step": n * 12, "call": "sphb_step + sphb_download_positions (state stays on the GPU, "
                          "12 B/particle come back: what main.cpp's update_positions needs each frame)"}
    eng.close()
    del eng, host_np, host, init, xyz

    rl = rooflines(n, phase, not args.no_accel, clocks.get("sm_mhz") if clocks else None)
    rl["roofline"]["steps_averaged"] = head["steps_averaged"]

    line = {
        "metric": METRIC, "value": head["value"], "unit": UNIT, "n_gpus": 1, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": head["ms_per_step"], "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
        "data": "synthetic",
        "scaling_note": "the --gpus N series replicates this workload per GPU (weak); BASELINE.json's own multi-GPU configs -- config 4 as ONE "
                        "tank (strong) and config 5 (periodic box, weak) -- are the config4_strong / config5_periodic sub-records of every --gpus N line",
        "config": {"workload": WORKLOADS[args.workload], "particles": n, "domain": args.domain, "reset_every": head["reset_every"], "cell_frac": args.cell_frac,
                   "cell_x_div": head["cell_x_div"],
                   "l2": "state (>= 32 B/particle x 2 buffers) exceeds the 126 MB L2" if n * 64 > 126e6 else
                         "state fits in L2: no flush between steps (steps are dependent, not repeats)",
                   "accel_output": not args.no_accel},
        "clocks": clocks, "e2e": e2e, "gpu_launches": head["gpu_launches"],
        "phase_ms": head["phase_ms"],
    }
    line.update(rl)

    if not args.quick:
        # ---- the other single-GPU configs BASELINE.json names, the drop-in's default grid mode, and config 1 on both sides
        configs = {}
        for name, k, w in (("default972", 100, 10), ("dam64k", 40, 8), ("dam1m", 100, 10)):     # config 3: 100 timed steps after 10 warm-up
            r = run_resident(name, k, w, local, domain="bbox" if name == "default972" else "fixed")
            r.update(rooflines(r["particles"], r["_phase"], True, clocks.get("sm_mhz") if clocks else None))
            configs[name] = strip(r)
        line["configs"] = configs
        bb = run_resident(args.workload, min(args.steps, 10), 3, local, accel=not args.no_accel, domain="bbox")
        line["bbox_mode"] = {"what": "the headline workload with SPHB_DOMAIN_BBOX (the drop-in default: bounding box folded into the force kernel, grid "
                                     "re-derived on the device every step, like update_cells)", "ms_per_step": bb["ms_per_step"], "value": bb["value"],
                             "phase_ms": bb["phase_ms"]}
        line["config1_vs_reference"] = config1_vs_reference(local)
        # ---- the two multi-GPU configs at N = 1: config 4 "strong" IS the headline here; config 5 runs its slab path with this GPU as
        # its own periodic neighbour (ghost images + self-exchange), which is what one rank of the 8-GPU run does
        from sph_b200 import multigpu
        scale = multigpu.SCALE_OF.get(args.workload)
        if scale is not None:
            line["config4_strong"] = {"workload": WORKLOADS[args.workload] + ", one GPU = the whole tank", "n_gpus": 1, "scaling": "strong",
                                      "ms_per_step": head["ms_per_step"], "value": head["value"], "phase_ms": head["phase_ms"], "efficiency_vs_n1": 1.0}
            p5 = multigpu.run_slab_workload(args, "periodic", scale, 1, 0, local, multigpu._Solo, max(4, min(args.steps, 10)), 3)
            p5 = multigpu.strip(p5)
            p5.update(scaling="weak", efficiency_vs_n1=1.0)
            line["config5_periodic"] = p5
    line["cpu_baseline"] = None if args.no_cpu_baseline else cpu_baseline_port()
    print(json.dumps(line, default=float), flush=True)


if __name__ == "__main__":
    main()
