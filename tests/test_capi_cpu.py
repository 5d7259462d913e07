"""CPU-side checks of the drop-in boundary: the C-ABI library loads, exports every symbol the header
declares, and the ctypes mirror of sphb_params matches the compiled struct.  No compute calls."""
import ctypes
import os
import re

import numpy as np
import pytest

import sph_b200
from sph_b200 import capi, scenes

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    src = open(os.path.join(ROOT, "include", "sphb200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(sphb_[a-z0-9_]+)\s*\(", src)))


def test_library_exports_every_declared_symbol():
    lib = capi.load_library()
    names = declared_symbols()
    assert len(names) >= 25
    for n in names:
        assert hasattr(lib, n), f"{n} declared in include/sphb200.h but not exported"
    assert sorted(capi.EXPORTS) == names, "capi.EXPORTS out of sync with the header"


def test_params_layout_and_defaults():
    lib = capi.load_library()
    assert lib.sphb_params_size() == ctypes.sizeof(capi.Params)
    p = capi.default_params()
    # main.cpp:30-41 and the literals of SphUtils.cpp
    assert (p.h, p.rho0, p.c, p.dt, p.eps) == tuple(np.float32(x) for x in (0.45, 10.0, 100.0, 0.033, 0.01))
    assert p.eos_k == 100.0 and tuple(p.gravity) == (0.0, -1.0, 0.0)
    assert (p.visc_alpha, p.visc_beta, p.visc_eta) == tuple(np.float32(x) for x in (0.1, 0.2, 0.01))
    assert (p.wall_test_off, p.wall_test_lim, p.restitution) == (0.001, 0.01, 0.0)
    assert p.domain_mode == capi.DOMAIN_BBOX


def test_record_dtypes_match_reference_layout():
    assert capi.PARTICLE.itemsize == 52 and capi.WALL.itemsize == 32 and capi.SPHERE.itemsize == 44
    assert [capi.PARTICLE.fields[k][1] for k in ("pos", "vel", "accel", "mass", "density", "pressure", "thermal")] == [0, 12, 24, 36, 40, 44, 48]
    assert [capi.WALL.fields[k][1] for k in ("normal", "center", "xlength", "ylength")] == [0, 12, 24, 28]


def test_no_gpu_means_loud_failure():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    with pytest.raises(sph_b200.SphbError):
        sph_b200.Engine()


def test_product_never_touches_the_oracle():
    bad = []
    for dirpath, _, files in os.walk(os.path.join(ROOT, "sph_b200")):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".hpp", ".h", ".cpp")):
                txt = open(os.path.join(dirpath, f), errors="ignore").read()
                if re.search(r"(from|import)\s+oracle|liborc|libsphref|sph_oracle", txt):
                    bad.append(f)
    assert not bad, bad


def test_scene_generators():
    ps, walls = scenes.dam_break(8, 16, 8, stressed=True)
    assert len(ps) == 1024 and len(walls) == 5
    # hash RNG: a slice of lattice planes equals the same planes of the whole
    whole, _ = scenes.lattice_block(8, 4, 4, 0.1, (0, 0, 0), vel_amp=0.3)
    part, _ = scenes.lattice_block(8, 4, 4, 0.1, (0, 0, 0), vel_amp=0.3, i_range=(2, 5))
    assert whole[2 * 16:5 * 16].tobytes() == part.tobytes()
    # jitter keeps points distinct
    assert len(np.unique(ps["pos"], axis=0)) == len(ps)
    ps2, _ = scenes.block_drop(scale=16)
    assert len(ps2) == 2 ** 24 // 16 ** 3


def test_multi_gpu_slab_plan_balances_owned_plus_ghost_counts():
    """sphb_multi_plan (the C++ restatement of multigpu.balanced_bounds that sphb_multi_upload_aos uses; host arithmetic, no GPU):
    on a lumpy distribution the cuts are increasing, sit on multiples of the support radius from the left-most particle, give
    every rank work -- its own particles plus the 2R ghost layers -- within 15 % of the mean, and do at least as well as plain
    particle-count quantiles."""
    from sph_b200 import multigpu
    rng = np.random.default_rng(4)
    x = np.concatenate([rng.uniform(0, 26, 300_000), rng.uniform(9, 16, 500_000), rng.normal(13, 0.5, 200_000)]).astype(np.float32)
    R = multigpu.support_radius()
    halo = 2 * R * 1.001
    xs = np.sort(x.astype(np.float64))
    load = lambda b: np.array([np.searchsorted(xs, np.float32(b[r + 1] + (halo if r < len(b) - 2 else 1e30))) -
                               np.searchsorted(xs, np.float32(b[r] - (halo if r else 1e30))) for r in range(len(b) - 1)], np.float64)
    for world in (2, 4, 8):
        b = capi.multi_plan(x, world)
        assert len(b) == world + 1 and b[0] == x.min() and b[-1] == x.max()
        assert np.all(np.diff(b) > 0)
        k = (b[1:-1] - b[0]) / R
        assert np.allclose(k, np.round(k), atol=1e-6)
        w = load(b)
        assert w.max() / w.mean() <= 1.15, (world, w)
        q = multigpu.quantile_bounds(x, world, R, float(x.min()), float(x.max()))
        assert w.max() <= load(q).max() * 1.02
    assert list(capi.multi_plan(x, 1)) == [float(x.min()), float(x.max())]
    with pytest.raises(sph_b200.SphbError):
        capi.multi_plan(x, 0)
