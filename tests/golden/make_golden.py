"""Generates tests/golden/*.npz from the UNMODIFIED reference (oracle/_ref/libsphref.so).

Run in the build container only (needs /root/reference to have been present for `make -C oracle ref`):
    python tests/golden/make_golden.py
The fixtures travel with the repository; the GPU box has no /root/reference.  Contents:
  kernels.npz        W / gradW / lapW known answers (SphUtils.cpp:55-95) at fixed point pairs
  default_scene.npz  main.cpp's built-in scene: full records after steps 1 and 20, FNV-1a-64 hashes of the
                     positions and statistics after steps 1, 2, 10, 20, 100
  block4096.npz      stressed 16^3 block: input records and the records after each phase of one step
  dam64k.npz         config 2 (65,536 particles, stressed): SHA-256 of the reference's all-pairs density /
                     pressure / accel arrays after one density+force pass, plus every 16th particle's values
"""
import hashlib
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle.oracle import Oracle, Reference  # noqa: E402
from sph_b200 import scenes  # noqa: E402

OUT = os.path.dirname(os.path.abspath(__file__))


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def main():
    ref = Reference()
    orc = Oracle()

    # --- the viewer's camera matrix (main.cpp:48-60) from the reference's vendored GLM
    np.savez_compressed(os.path.join(OUT, "camera.npz"), mvp=ref.default_mvp())
    if "--only-camera" in sys.argv:
        return

    # --- kernels
    rng = np.random.default_rng(7)
    xi = rng.uniform(-1, 1, (256, 3)).astype(np.float32)
    r = np.concatenate([np.linspace(0.0, 0.25, 200), [0.1, 0.2, 0.2025, 0.21, 0.20249999]]).astype(np.float32)
    d = rng.normal(size=(len(r), 3))
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    xi = np.resize(xi, (len(r), 3)).astype(np.float32)
    xj = (xi + (d * r[:, None]).astype(np.float32)).astype(np.float32)
    xi[0] = (0, 0, 0); xj[0] = (0.1, 0, 0)      # SURVEY.md section 4 known answers
    w = np.array([ref.kernel_function(a, b) for a, b in zip(xi, xj)], np.float32)
    g = np.array([ref.grad_kernel(a, b) for a, b in zip(xi, xj)], np.float32)
    lap = np.array([ref.lap_kernel(a, b) for a, b in zip(xi, xj)], np.float32)
    np.savez_compressed(os.path.join(OUT, "kernels.npz"), xi=xi, xj=xj, w=w, grad=g, lap=lap)

    # --- default scene
    ps, walls = scenes.default_scene()
    cur = ps.copy()
    done = 0
    steps, hashes, stats, full = [1, 2, 10, 20, 100], [], [], {}
    for s in steps:
        ref.step(cur, walls, s - done)
        done = s
        hashes.append(orc.fnv1a_pos(cur))
        m = cur["mass"].astype(np.float64)
        v = cur["vel"].astype(np.float64)
        stats.append([cur["density"].mean(dtype=np.float64), float(cur["density"].max()),
                      float((0.5 * m * (v ** 2).sum(1)).sum()), float((m * (cur["pos"][:, 1].astype(np.float64) + 1)).sum())])
        if s in (1, 20):
            full[f"after{s}"] = cur.copy()
    np.savez_compressed(os.path.join(OUT, "default_scene.npz"), initial=ps, walls=walls, steps=np.array(steps),
                        fnv=np.array(hashes, np.uint64), stats=np.array(stats), **full)

    # --- stressed 16^3 block, one step phase by phase
    ps, walls = scenes.dam_break(16, 16, 16, stressed=True)
    a = ps.copy(); ref.update_density(a)
    b = a.copy(); ref.update_forces(b)
    c = b.copy(); ref.update_posvel(c)
    d_ = c.copy(); ref.collision(d_, walls)
    np.savez_compressed(os.path.join(OUT, "block4096.npz"), initial=ps, walls=walls, density=a, forces=b, posvel=c, collision=d_)

    # --- config 2: 65,536 particles, stressed, all-pairs density + forces on the reference (minutes)
    ps, walls = scenes.dam_break(32, 64, 32, stressed=True)
    t = time.time()
    a = ps.copy(); ref.update_density(a); ref.update_forces(a)
    print("reference all-pairs 64K density+forces: %.1f s" % (time.time() - t))
    np.savez_compressed(os.path.join(OUT, "dam64k.npz"), n=len(ps), sha_density=sha(a["density"]), sha_pressure=sha(a["pressure"]),
                        sha_accel=sha(a["accel"]), sub_density=a["density"][::16].copy(), sub_pressure=a["pressure"][::16].copy(),
                        sub_accel=a["accel"][::16].copy(), seconds=time.time() - t)
    for f in sorted(os.listdir(OUT)):
        if f.endswith(".npz"):
            print(f, os.path.getsize(os.path.join(OUT, f)))


if __name__ == "__main__":
    main()
