"""sphb_params.parity_mode: the CUDA path evaluating every pair in the reference's own order and rounding (SURVEY.md H2's "parity
mode").  Bar: BIT-EXACT against the unmodified reference's results held in tests/golden/ (generated from oracle/_ref) -- density,
pressure and acceleration at 4,096 particles phase by phase, the reference's own 972-particle scene over 100 steps (its position
hashes at steps 1 / 2 / 10 / 20 / 100), and config 2 at full size: the SHA-256 of the reference's all-pairs density, pressure and
acceleration of 65,536 particles."""
import hashlib

import numpy as np
import pytest

import sph_b200
from sph_b200 import F_ALL, scenes

pytestmark = pytest.mark.gpu


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def fnv1a_pos(ps):
    h = 0xcbf29ce484222325
    for w in np.ascontiguousarray(ps["pos"]).view(np.uint32).ravel().tolist():
        h = ((h ^ w) * 0x100000001b3) & 0xFFFFFFFFFFFFFFFF
    return h


def test_parity_mode_phases_bit_exact_block4096(golden):
    g = golden("block4096.npz")
    with sph_b200.Engine(parity_mode=1) as e:
        e.set_walls(g["walls"])
        ps = g["initial"].copy()
        e.upload(ps)
        e.update_density()
        e.download(ps, F_ALL)
        assert ps["density"].tobytes() == g["density"]["density"].tobytes()
        assert ps["pressure"].tobytes() == g["density"]["pressure"].tobytes()
        e.update_forces()
        e.download(ps, F_ALL)
        assert ps["accel"].tobytes() == g["forces"]["accel"].tobytes()
        # and a whole step: positions and velocities after posvel + collision
        ps = g["initial"].copy()
        e.step_aos(ps, 1)
        want = g["collision"]
        for f in ("pos", "vel", "accel", "density", "pressure"):
            assert ps[f].tobytes() == want[f].tobytes(), f


def test_parity_mode_default_scene_100_steps_bit_exact(golden):
    g = golden("default_scene.npz")
    ps = g["initial"].copy()
    with sph_b200.Engine(parity_mode=1) as e:
        e.set_walls(g["walls"])
        done = 0
        for k, steps in enumerate(g["steps"].tolist()):
            e.step_aos(ps, steps - done)
            done = steps
            assert fnv1a_pos(ps) == int(g["fnv"][k]), f"positions differ from the reference after step {steps}"
    assert ps.tobytes() != g["initial"].tobytes()


def test_parity_mode_dam64k_sha_of_the_reference(golden):
    g = golden("dam64k.npz")
    ps, _ = scenes.dam_break(32, 64, 32, stressed=True)
    assert len(ps) == int(g["n"])
    with sph_b200.Engine(parity_mode=1) as e:
        e.upload(ps)
        e.update_density()
        e.update_forces()
        e.download(ps, F_ALL)
    assert sha(ps["density"]) == str(g["sha_density"])
    assert sha(ps["pressure"]) == str(g["sha_pressure"])
    assert sha(ps["accel"]) == str(g["sha_accel"])
