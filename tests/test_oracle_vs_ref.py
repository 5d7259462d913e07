"""Pins the C restatement (oracle/sph_oracle.c) bit-for-bit against the UNMODIFIED reference
(oracle/_ref/libsphref.so = /root/reference/SPH/SphUtils.cpp compiled in place).  CPU only."""
import numpy as np
import pytest

from sph_b200 import scenes


def same_bits(a, b):
    return np.ascontiguousarray(a).tobytes() == np.ascontiguousarray(b).tobytes()


def test_record_layout(reference):
    # SPH/Particle.hpp:42-50, Wall.hpp:28-31, Sphere.hpp:29-33
    off = reference.offsets()
    assert [off[k] for k in (0, 1, 2, 3, 4, 5, 6)] == [0, 12, 24, 36, 40, 44, 48]
    assert (off[10], off[11]) == (0, 12)
    assert [off[k] for k in (20, 21, 22, 23, 24)] == [0, 12, 24, 36, 40]


def test_default_scene_is_972(oracle):
    ps, walls = oracle.scene_default()
    assert len(ps) == 972 and len(walls) == 5
    assert ps["mass"][0] == np.float32(0.0911249891)
    mine, mywalls = scenes.default_scene()
    assert same_bits(ps, mine) and same_bits(walls, mywalls)


def test_kernels_bit_exact(oracle, reference):
    rng = np.random.default_rng(1)
    xi = rng.uniform(-1, 1, (2000, 3)).astype(np.float32)
    xj = (xi + rng.normal(scale=0.12, size=(2000, 3))).astype(np.float32)
    xj[:10] = xi[:10]                                              # coincident: W(0), NaN gradient
    for a, b in zip(xi, xj):
        assert np.float32(oracle.kernel_function(a, b)).tobytes() == np.float32(reference.kernel_function(a, b)).tobytes()
        assert same_bits(oracle.grad_kernel(a, b), reference.grad_kernel(a, b))
        assert np.float32(oracle.lap_kernel(a, b)).tobytes() == np.float32(reference.lap_kernel(a, b)).tobytes()


def test_kernel_known_answers(reference):
    # SURVEY.md section 4, h = 0.45
    assert reference.kernel_function([0, 0, 0], [0, 0, 0]) == pytest.approx(17.1926636, rel=1e-7)
    assert reference.kernel_function([0, 0, 0], [0.1, 0, 0]) == pytest.approx(7.43259907, rel=1e-7)
    assert reference.kernel_function([0, 0, 0], [0.2, 0, 0]) == pytest.approx(2.54038081e-4, rel=1e-6)
    assert reference.kernel_function([0, 0, 0], [0.21, 0, 0]) == 0.0
    g = reference.grad_kernel([0, 0, 0], [0.1, 0, 0])
    assert g[0] == pytest.approx(89.4973602, rel=1e-7)
    assert np.isnan(reference.grad_kernel([0, 0, 0], [0, 0, 0])).all()


@pytest.mark.parametrize("steps", [1, 20, 60])
def test_default_scene_steps_bit_exact(oracle, reference, steps):
    ps, walls = scenes.default_scene()
    a, b, c = ps.copy(), ps.copy(), ps.copy()
    oracle.step(a, walls, steps)
    reference.step(b, walls, steps)
    oracle.step(c, walls, steps, cells=True, threads=0)
    assert same_bits(a, b)
    assert same_bits(c, b)      # cell list in ascending original index == all pairs, bit for bit


@pytest.mark.parametrize("stressed", [False, True])
def test_phases_bit_exact_on_jittered_block(oracle, reference, stressed):
    ps, walls = scenes.dam_break(12, 14, 10, stressed=stressed, seed=99)
    a, b, c = ps.copy(), ps.copy(), ps.copy()
    for phase in ("update_density", "update_forces", "update_posvel"):
        getattr(oracle, phase)(a)
        getattr(reference, phase)(b)
        assert same_bits(a, b), phase
    oracle.collision(a, walls)
    reference.collision(b, walls)
    assert same_bits(a, b)
    oracle.update_density(c, cells=True, threads=3)
    oracle.update_forces(c, cells=True, threads=3)
    d = ps.copy()
    reference.update_density(d)
    reference.update_forces(d)
    assert same_bits(c, d)
    if stressed:
        assert (d["density"] > 10.0).mean() > 0.5   # the P > 0 and v.r < 0 branches are exercised


def test_collision_thresholds_bit_exact(oracle, reference):
    # particles straddling the hit test dot-0.001 <= 0.01 of every wall (SphUtils.cpp:349)
    _, walls = scenes.default_scene()
    rng = np.random.default_rng(3)
    pos = rng.uniform(-1.02, 1.02, (4000, 3)).astype(np.float32)
    pos[:500, 0] = np.float32(-1.0) + rng.uniform(0.0109, 0.0111, 500).astype(np.float32)
    pos[500:1000, 1] = np.float32(-1.0) + rng.uniform(0.0109, 0.0111, 500).astype(np.float32)
    vel = rng.normal(size=(4000, 3)).astype(np.float32)
    ps = scenes.make_particles(pos, vel)
    a, b = ps.copy(), ps.copy()
    oracle.collision(a, walls)
    reference.collision(b, walls)
    assert same_bits(a, b)


def test_sphere_phases_bit_exact(oracle, reference):
    from oracle.oracle import SPHERE
    ps, walls = scenes.dam_break(6, 6, 6, stressed=True)
    sp = np.zeros(1, SPHERE)
    sp[0]["pos"] = (0.4, 0.4, 0.4); sp[0]["vel"] = (0, 2, 0); sp[0]["radius"] = 0.1; sp[0]["mass"] = 1.2
    a, b, sa, sb = ps.copy(), ps.copy(), sp.copy(), sp.copy()
    oracle.update_posvel(a, sa); reference.update_posvel(b, sb)
    oracle.collision(a, walls, sa); reference.collision(b, walls, sb)
    assert same_bits(a, b) and same_bits(sa, sb)


def test_update_cells_restatement(oracle, reference):
    # the reference's dead update_cells (SphUtils.cpp:8-47) incl. its buggy flat index; default scene keeps it in bounds
    ps, _ = scenes.default_scene()
    n1, bbox1, dims1, cell1 = oracle.ref_update_cells(ps)
    n2, bbox2, dims2, cell2 = reference.update_cells(ps)
    assert n1 == n2 == 18 and same_bits(bbox1, bbox2) and same_bits(dims1, dims2) and same_bits(cell1, cell2)
    assert sorted(set(cell1.tolist())) == [0, 1, 8, 9]


def test_dead_viscosity_forces_pair_body(oracle, reference):
    """viscosity_forces (SphUtils.cpp:244-290) never runs in the reference (:175) and scans only half a stencil of
    2h-wide cells; on a block that fits inside ONE such cell it degenerates to all pairs in ascending index, which
    pins the oracle's restatement of its pair body (:277-278) and of lap_kernel (:84-95)."""
    ps, _ = scenes.dam_break(8, 8, 8, stressed=True)          # 0.68 wide < 2h = 0.9
    oracle.update_density(ps); oracle.update_forces(ps)
    a, b, c = ps.copy(), ps.copy(), ps.copy()
    oracle.viscosity_forces(a)
    reference.viscosity_forces(b)
    oracle.viscosity_forces(c, cells=True, threads=2)
    assert same_bits(a, b) and same_bits(a, c)
    assert np.abs(a["accel"] - ps["accel"]).max() > 0.05


def test_camera_golden_is_what_the_reference_glm_computes(reference, golden):
    assert np.array_equal(reference.default_mvp(), golden("camera.npz")["mvp"])
