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
    line = {"impl": "reference", "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": args.gpus, "steps": steps, "warmup": warm,
            "ms_per_step": 1e3 * dt / steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
            "data": "synthetic", "config": {"workload": WORKLOADS[args.workload], "sample_particles": len(ps)},
            "cpu_baseline": {"value": val, "unit": UNIT, "cores": cores, "kind": kind, "sample": sample},
            "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}, "gpu_launches": 0}
    print(json.dumps(line, default=float), flush=True)


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
                    help="multi-GPU: weak = one 16.7 M-particle tank section per GPU (default); strong = the ONE config-4 tank "
                         "cut into N slabs at particle-count quantiles")
    ap.add_argument("--reset-every", type=int, default=None,
                    help="restore the initial state (device-to-device, inside the timed region) every this many steps: the "
                         "reference's scheme (dt=0.033, k=100) blows up on deep pools after ~35 steps, in its own arithmetic too "
                         "(default 16; 8 for the pre-compressed 64 K dam break, which explodes within a dozen steps)")
    ap.add_argument("--cell-frac", type=float, default=1.0, help="grid cell edge as a fraction of the support radius h*h (>= 0.5)")
    ap.add_argument("--cell-x-div", type=int, default=0, help="x pitch of the grid = cell / this (0 = engine default, 2)")
    args = ap.parse_args()
    if args.reset_every is None:
        args.reset_every = 8 if args.workload == "dam64k" else 16
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

    ps, walls, (dlo, dhi) = make_workload(args.workload)
    n = len(ps)
    kw = dict(write_accel=0 if args.no_accel else 1)
    xdiv = args.cell_x_div if args.cell_x_div > 0 else 2
    cell = float(np.float32(0.45) * np.float32(0.45)) * args.cell_frac
    if args.domain == "fixed":
        ncell = ((dhi[0] - dlo[0]) / (cell / xdiv) + 2) * ((dhi[1] - dlo[1]) / cell + 2) * ((dhi[2] - dlo[2]) / cell + 2)
        kw.update(max_cells=int(ncell) + 4096)
    kw.update(cell_size=cell if args.cell_frac != 1.0 else 0.0, cell_x_div=args.cell_x_div)
    if args.domain == "fixed":
        kw.update(domain_mode=sph_b200.DOMAIN_FIXED, domain_lo=dlo, domain_hi=dhi)
    eng = sph_b200.Engine(device=local, **kw)
    eng.set_walls(walls)

    # host records in pinned memory (torch is plumbing: it owns the page-locked buffer)
    host = torch.empty(n * 52, dtype=torch.uint8).pin_memory()
    host_np = host.numpy().view(sph_b200.PARTICLE)
    host_np[:] = ps
    del ps

    # ---- resident-state throughput
    eng.upload(host_np)
    eng.update_cells()
    init = [torch.from_numpy(a).cuda() for a in eng.download_soa()[:2]]   # grid-sorted initial state, kept on the device
    init[1][:, 3] = 1.0
    since_reset = 0

    def step_once():
        nonlocal since_reset
        if args.reset_every > 0 and since_reset >= args.reset_every:
            eng.upload_soa_device(init[0].data_ptr(), init[1].data_ptr(), n)
            since_reset = 0
        eng.step(1)
        since_reset += 1

    for _ in range(args.warmup):
        step_once()
    eng.synchronize()
    launches0 = eng.launch_count()
    eng.enable_timing(True)
    sampler = ClockSampler(local)
    start, stop = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    stream = torch.cuda.ExternalStream(eng.stream(), device=local)
    torch.cuda.synchronize()
    start.record(stream)
    for _ in range(args.steps):
        step_once()
    stop.record(stream)
    eng.synchronize()
    torch.cuda.synchronize()
    ms = start.elapsed_time(stop)
    clocks = sampler.stop()
    launches = eng.launch_count() - launches0
    phase, timed = eng.get_timing()
    eng.enable_timing(False)
    value = n * args.steps / (ms * 1e-3)

    # sanity: the state after the timed steps is finite (a blown-up state would make the timing meaningless)
    bad = eng.count_nonfinite()
    assert bad == 0, f"{bad} non-finite particles after the timed steps: lower --reset-every"

    # ---- end to end through the reference-facing call on host records
    e2e_steps = max(1, args.e2e_steps)
    eng.step_aos(host_np, 1)            # warm the staging buffers
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        eng.step_aos(host_np, 1)        # upload + step + download, synchronised
    t_e2e = (time.perf_counter() - t0) / e2e_steps
    e2e = {"value": n / t_e2e, "unit": UNIT, "h2d_bytes_per_step": n * 52, "d2h_bytes_per_step": n * 52,
           "ms_per_step": 1e3 * t_e2e, "call": "sphb_step_aos (52-byte Particle records, pinned host memory)"}

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

    # ---- roofline of the dominant kernel
    peak, peak_src = measured_peak()
    bforce = B_FORCE_NOACC if args.no_accel else B_FORCE
    ach = bforce * n / (phase[2] * 1e-3) / 1e9
    fused_ms = float(phase[1] + phase[2])
    traffic, traffic_src = None, None
    try:
        tj = json.load(open(os.path.join(ROOT, "profiles", "traffic.json")))["k_force"]
        traffic = tj["dram_bytes_per_particle"] * n / (phase[2] * 1e-3) / 1e9      # GB/s, per launch like `achieved`
        traffic_src = tj["source"]
    except Exception:
        pass
    roofline = {"bound": "hbm", "kernel": "k_force (pair force + integrate + wall collision)", "achieved": ach, "peak": peak,
                "unit": "GB/s", "frac": ach / peak, "traffic": traffic, "traffic_source": traffic_src, "peak_source": peak_src,
                "bytes_per_particle": bforce, "kernel_ms": float(phase[2]), "steps_averaged": int(timed)}
    ach_f = (B_DENSITY + bforce) * n / (fused_ms * 1e-3) / 1e9
    roofline_fused = {"what": "density + force/integrate/collide kernels together (north_star's fused step)",
                      "bytes_per_particle": B_DENSITY + bforce, "achieved": ach_f, "frac": ach_f / peak, "ms": fused_ms}
    ach_s = B_STEP * n / (float(phase[3]) * 1e-3) / 1e9
    roofline_step = {"what": "whole step incl. grid/sort/reorder", "bytes_per_particle": B_STEP, "achieved": ach_s,
                     "frac": ach_s / peak, "ms": float(phase[3])}

    cpu = None if args.no_cpu_baseline else cpu_baseline_port()

    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": 1, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
        "data": "synthetic",
        "config": {"workload": WORKLOADS[args.workload], "particles": n, "domain": args.domain, "reset_every": args.reset_every, "cell_frac": args.cell_frac, "cell_x_div": xdiv,
                   "l2": "state (>= 32 B/particle x 2 buffers) exceeds the 126 MB L2" if n * 64 > 126e6 else
                         "state fits in L2: no flush between steps (steps are dependent, not repeats)",
                   "accel_output": not args.no_accel},
        "clocks": clocks, "e2e": e2e, "gpu_launches": int(launches),
        "roofline": roofline, "roofline_fused": roofline_fused, "roofline_step": roofline_step,
        "phase_ms": {"sort": float(phase[0]), "density": float(phase[1]), "force": float(phase[2]), "step": float(phase[3])},
        "cpu_baseline": cpu,
    }
    print(json.dumps(line, default=float), flush=True)
    eng.close()


if __name__ == "__main__":
    main()
