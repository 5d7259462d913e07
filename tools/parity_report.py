#!/usr/bin/env python
"""Per-particle parity table: density and acceleration of the CUDA path against the reference arithmetic (the
oracle's cell-list restatement, bit-identical to the reference's loops), with an FP64 evaluation of the same
formulas as referee, for the sizes of BASELINE.json's configs.

    python tools/parity_report.py [--sizes 972,4096,64k,1m,16m] [--out gpurun_out/parity_report.json]

Runs on a GPU box (the CUDA path through the C ABI; oracle/ is the checker).  SPHB_LIB selects an experimental
build of the library.  For every size it prints
  * density: max |rho - rho_ref| / rho_ref, and the same for reference-vs-FP64 and ours-vs-FP64
  * acceleration with the force kernel fed the REFERENCE's rho and P (isolates the force arithmetic):
      plain relative error |a - a_ref| / |a_ref|: max, 99.9th percentile, number of particles above 1e-5
      normalised error |a - a_ref| / (1 + sum_j |temp_ij| |gradW_ij|): max
    each for ours-vs-reference, reference-vs-FP64 and ours-vs-FP64.
"""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import sph_b200  # noqa: E402
from oracle.oracle import Oracle  # noqa: E402
from sph_b200 import F_ALL, scenes  # noqa: E402


def make_state(name, orc):
    """(label, particles in a state with rho > rho0 and approaching pairs, engine kwargs)"""
    if name == "972":
        ps, walls = scenes.default_scene()
        orc.step(ps, walls, 20, cells=True)                 # step 20: first compressed state of main.cpp's scene
        return "972 (main.cpp scene after 20 reference steps)", ps, {}
    if name == "4096":
        g = np.load(os.path.join(ROOT, "tests", "golden", "block4096.npz"))
        return "4,096 (stressed block, golden)", g["initial"].copy(), {}
    if name == "64k":
        ps, _ = scenes.dam_break(32, 64, 32, stressed=True)
        return "65,536 (config 2, stressed dam break)", ps, {}
    import bench
    wl = {"1m": "dam1m", "16m": "drop16m", "2m": "drop2m"}[name]
    ps, walls, (dlo, dhi) = bench.make_workload(wl)
    kw = dict(domain_mode=sph_b200.DOMAIN_FIXED, domain_lo=dlo, domain_hi=dhi)
    with sph_b200.Engine(**kw) as e:                        # advance on the GPU to just before the scheme blows up
        e.set_walls(walls)
        e.upload(ps)
        e.step(28)
        e.download(ps, F_ALL)
        assert e.count_nonfinite() == 0
    label = {"1m": "1,048,576 (config 3 after 28 steps)", "16m": "16,777,216 (config 4 after 28 steps)",
             "2m": "2,097,152 (config 4 at scale 2 after 28 steps)"}[name]
    return label, ps, kw


def stats(err):
    return {"max": float(err.max()), "p99.9": float(np.quantile(err, 0.999)), "above_1e-5": int((err > 1e-5).sum())}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--sizes", default="972,4096,64k,1m")
    ap.add_argument("--out", default=os.path.join(ROOT, "gpurun_out", "parity_report.json"))
    args = ap.parse_args()
    orc = Oracle()
    rows = []
    for name in args.sizes.split(","):
        t0 = time.time()
        label, ps, kw = make_state(name, orc)
        n = len(ps)
        want = ps.copy()
        orc.update_density(want, cells=True)
        rho64 = np.maximum(orc.density_f64(ps), float(orc.params.rho0))
        with sph_b200.Engine(**kw) as e:
            got = ps.copy()
            e.upload(got)
            e.update_density()
            e.download(got, sph_b200.F_DENSITY)
            orc.update_forces(want, cells=True)             # the reference's accel from the reference's rho, P
            fed = want.copy()
            e.upload(fed)
            e.update_forces()
            e.download(fed, sph_b200.F_ACCEL)
        a64, scale = orc.forces_f64(want)
        wd = want["density"].astype(np.float64)
        row = {"config": label, "n": n, "compressed_fraction": float((want["density"] > orc.params.rho0).mean()),
               "density": {"ours_vs_ref": float(np.max(np.abs(got["density"] - wd) / wd)),
                           "ref_vs_f64": float(np.max(np.abs(wd - rho64) / rho64)),
                           "ours_vs_f64": float(np.max(np.abs(got["density"] - rho64) / rho64))}}
        aw = want["accel"].astype(np.float64)
        ao = fed["accel"].astype(np.float64)
        for key, x, y in (("ours_vs_ref", ao, aw), ("ref_vs_f64", aw, a64), ("ours_vs_f64", ao, a64)):
            d = np.linalg.norm(x - y, axis=1)
            row["accel_" + key] = {"plain": stats(d / np.maximum(np.linalg.norm(y, axis=1), 1e-30)), "normalised_max": float(np.max(d / scale))}
        row["seconds"] = time.time() - t0
        rows.append(row)
        print(json.dumps(row), flush=True)
    os.makedirs(os.path.dirname(args.out), exist_ok=True)
    with open(args.out, "w") as f:
        json.dump({"library": os.environ.get("SPHB_LIB", "sph_b200/libsphb200.so"), "rows": rows}, f, indent=1)
    print("\n| config | rho: ours-ref | rho: ref-f64 | accel plain max / p99.9 / #>1e-5 (ours-ref) | (ref-f64) | (ours-f64) | accel normalised max ours-ref / ref-f64 / ours-f64 |")
    print("|---|---|---|---|---|---|---|")
    for r in rows:
        f = lambda k: "%.1e / %.1e / %d" % (r[k]["plain"]["max"], r[k]["plain"]["p99.9"], r[k]["plain"]["above_1e-5"])
        print(f"| {r['config']} | {r['density']['ours_vs_ref']:.1e} | {r['density']['ref_vs_f64']:.1e} | {f('accel_ours_vs_ref')} | {f('accel_ref_vs_f64')} | "
              f"{f('accel_ours_vs_f64')} | {r['accel_ours_vs_ref']['normalised_max']:.1e} / {r['accel_ref_vs_f64']['normalised_max']:.1e} / {r['accel_ours_vs_f64']['normalised_max']:.1e} |")


if __name__ == "__main__":
    main()
