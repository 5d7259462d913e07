#!/usr/bin/env python
"""A/B timing of experimental builds of libsphb200.so (sph_b200/variants/, `make -C sph_b200/csrc variant ...`).

    python tools/kbench.py --libs default,pm0,pm2 [--workload drop16m] [--steps 20] [--cell-x-div 0]

One child process per library (SPHB_LIB), the bench's own workload and reset rule, state resident; prints one JSON
line per library with the per-phase CUDA-event times of sphb_step (sort / density / force / step, ms).
Development tool: the numbers that count are bench.py's.
"""
import argparse
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def child(args):
    import numpy as np
    import torch
    import bench
    import sph_b200
    ps, walls, (dlo, dhi) = bench.make_workload(args.workload)
    n = len(ps)
    xdiv = args.cell_x_div if args.cell_x_div > 0 else 2
    cell = float(np.float32(0.45) * np.float32(0.45))
    ncell = ((dhi[0] - dlo[0]) / (cell / xdiv) + 2) * ((dhi[1] - dlo[1]) / cell + 2) * ((dhi[2] - dlo[2]) / cell + 2)
    eng = sph_b200.Engine(device=0, write_accel=0 if args.no_accel else 1, max_cells=int(ncell) + 4096, cell_x_div=args.cell_x_div,
                          domain_mode=sph_b200.DOMAIN_FIXED, domain_lo=dlo, domain_hi=dhi)
    eng.set_walls(walls)
    eng.upload(ps)
    eng.update_cells()
    init = [torch.from_numpy(a).cuda() for a in eng.download_soa()[:2]]
    init[1][:, 3] = 1.0
    reset = 4 if args.workload == "dam64k" else 16
    since = 0

    def step():
        nonlocal since
        if since >= reset:
            eng.upload_soa_device(init[0].data_ptr(), init[1].data_ptr(), n)
            since = 0
        eng.step(1)
        since += 1

    for _ in range(5):
        step()
    eng.synchronize()
    eng.enable_timing(True)
    for _ in range(args.steps):
        step()
    phase, timed = eng.get_timing()
    bad = eng.count_nonfinite()
    print(json.dumps({"lib": os.environ.get("SPHB_LIB", "default"), "workload": args.workload, "n": n, "cell_x_div": xdiv, "nonfinite": bad,
                      "sort": round(float(phase[0]), 4), "density": round(float(phase[1]), 4), "force": round(float(phase[2]), 4),
                      "step": round(float(phase[3]), 4), "steps": int(timed)}), flush=True)
    eng.close()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--libs", default="default")
    ap.add_argument("--workload", default="drop16m")
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--cell-x-div", type=int, default=0)
    ap.add_argument("--no-accel", action="store_true")
    ap.add_argument("--child", action="store_true")
    args = ap.parse_args()
    if args.child:
        return child(args)
    for name in args.libs.split(","):
        env = dict(os.environ)
        if name != "default":
            path = name if os.path.exists(name) else os.path.join(ROOT, "sph_b200", "variants", f"libsphb200_{name}.so")
            env["SPHB_LIB"] = path
        else:
            env.pop("SPHB_LIB", None)
        cmd = [sys.executable, os.path.abspath(__file__), "--child", "--workload", args.workload, "--steps", str(args.steps),
               "--cell-x-div", str(args.cell_x_div)] + (["--no-accel"] if args.no_accel else [])
        subprocess.run(cmd, env=env, check=False)


if __name__ == "__main__":
    main()
