"""The oracle against the committed golden vectors (generated from the compiled reference by
tests/golden/make_golden.py).  Runs anywhere: needs neither /root/reference nor a GPU."""
import hashlib

import numpy as np

from sph_b200 import scenes


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def test_kernels(oracle, golden):
    g = golden("kernels.npz")
    w = np.array([oracle.kernel_function(a, b) for a, b in zip(g["xi"], g["xj"])], np.float32)
    gr = np.array([oracle.grad_kernel(a, b) for a, b in zip(g["xi"], g["xj"])], np.float32)
    lap = np.array([oracle.lap_kernel(a, b) for a, b in zip(g["xi"], g["xj"])], np.float32)
    assert w.tobytes() == g["w"].tobytes()
    assert gr.tobytes() == g["grad"].tobytes()
    assert lap.tobytes() == g["lap"].tobytes()


def test_default_scene_trajectory(oracle, golden):
    g = golden("default_scene.npz")
    ps, walls = scenes.default_scene()
    assert ps.tobytes() == g["initial"].tobytes() and walls.tobytes() == g["walls"].tobytes()
    cur, done = ps.copy(), 0
    for s, h in zip(g["steps"], g["fnv"]):
        oracle.step(cur, walls, int(s - done), cells=True)
        done = int(s)
        assert oracle.fnv1a_pos(cur) == int(h), f"step {s}"
        if f"after{s}" in g.files:
            assert cur.tobytes() == g[f"after{s}"].tobytes()


def test_block4096_phases(oracle, golden):
    g = golden("block4096.npz")
    a = g["initial"].copy()
    oracle.update_density(a, cells=True)
    assert a.tobytes() == g["density"].tobytes()
    oracle.update_forces(a, cells=True)
    assert a.tobytes() == g["forces"].tobytes()
    oracle.update_posvel(a)
    assert a.tobytes() == g["posvel"].tobytes()
    oracle.collision(a, g["walls"])
    assert a.tobytes() == g["collision"].tobytes()


def test_dam64k_cell_list_equals_reference_all_pairs(oracle, golden):
    """Config 2 at full size: the oracle's cell-list pass reproduces the reference's O(N^2) loops bit for bit."""
    g = golden("dam64k.npz")
    ps, _ = scenes.dam_break(32, 64, 32, stressed=True)
    assert len(ps) == int(g["n"]) == 65536
    oracle.update_density(ps, cells=True)
    oracle.update_forces(ps, cells=True)
    assert sha(ps["density"]) == str(g["sha_density"])
    assert sha(ps["pressure"]) == str(g["sha_pressure"])
    assert sha(ps["accel"]) == str(g["sha_accel"])
    assert ps["density"][::16].tobytes() == g["sub_density"].tobytes()


def test_fp64_referee_brackets_the_reference_rounding(oracle, golden):
    """The FP64 evaluation of the reference's formulas (oracle.density_f64 / forces_f64, the referee of SURVEY.md H2)
    agrees with the reference's own float results to the level float rounding explains: density <= 1e-6 relative,
    acceleration <= 1e-6 of 1 + sum|terms| (plain relative error is larger where terms cancel, <= 5e-5 here)."""
    g = golden("block4096.npz")
    ps = g["density"].copy()                       # state after the reference's update_density
    rho64 = np.maximum(oracle.density_f64(g["initial"]), 10.0)
    assert np.max(np.abs(ps["density"] - rho64) / rho64) <= 1e-6
    a64, scale = oracle.forces_f64(ps)
    d = np.linalg.norm(g["forces"]["accel"].astype(np.float64) - a64, axis=1)
    assert np.all(scale >= 1.0)
    assert np.max(d / scale) <= 1e-6
    assert np.max(d / np.linalg.norm(a64, axis=1)) <= 5e-5


def test_camera_matrix_matches_the_references_glm(oracle, golden):
    """Row f2, the part of it the reference pins: MVP = perspective(45, 4/3, 0.1, 100) * lookAt(...) * I of main.cpp:48-60,
    as computed by the reference's vendored GLM 0.9.6.1 (tests/golden/camera.npz, from oracle/_ref).  Both the oracle's
    restatement and the library's sphb_default_camera (host arithmetic, no GPU involved) reproduce it bit for bit."""
    import sph_b200
    want = golden("camera.npz")["mvp"]
    assert np.array_equal(oracle.default_mvp(), want)
    cam = sph_b200.capi.default_camera()
    assert np.array_equal(np.array(cam.mvp[:], np.float32), want)
    assert (cam.width, cam.height) == (1080, 920)       # main.cpp:13-14
