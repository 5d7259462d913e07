"""TEST INFRASTRUCTURE ONLY: ctypes loaders for the two CPU checkers.

* ``Oracle``   -- oracle/liborc.so, our plain-C restatement (oracle/sph_oracle.c).
* ``Reference`` -- oracle/_ref/libsphref.so, the UNMODIFIED reference ``SPH/SphUtils.cpp`` behind
  oracle/ref_driver.cpp (exists only where /root/reference was present at build time, or where the
  prebuilt .so travelled with the snapshot).

Only tests/, ``__graft_entry__.smoke()`` and ``bench.py``'s cpu_baseline / ``--impl reference`` legs may
import this module; nothing under ``sph_b200/`` does.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))

# SPH/Particle.hpp:42-50 (52 B), SPH/Wall.hpp:28-31 (32 B), SPH/Sphere.hpp:29-33 (44 B)
PARTICLE = np.dtype(
    [("pos", "<f4", 3), ("vel", "<f4", 3), ("accel", "<f4", 3), ("mass", "<f4"), ("density", "<f4"),
     ("pressure", "<f4"), ("thermal", "<f4")]
)
WALL = np.dtype([("normal", "<f4", 3), ("center", "<f4", 3), ("xlength", "<f4"), ("ylength", "<f4")])
SPHERE = np.dtype([("pos", "<f4", 3), ("vel", "<f4", 3), ("accel", "<f4", 3), ("radius", "<f4"), ("mass", "<f4")])
assert PARTICLE.itemsize == 52 and WALL.itemsize == 32 and SPHERE.itemsize == 44


class Params(C.Structure):
    _fields_ = [("h", C.c_float), ("rho0", C.c_float), ("c", C.c_float), ("dt", C.c_float), ("eps", C.c_float)]


class Grid(C.Structure):
    _fields_ = [("origin", C.c_float * 3), ("cell", C.c_float * 3), ("dims", C.c_int * 3)]


# main.cpp:30-41
DEFAULTS = dict(h=0.45, rho0=10.0, c=100.0, dt=0.033, eps=0.01)


def build(ref: bool = True) -> None:
    """(Re)build liborc.so and, if /root/reference exists, _ref/libsphref.so."""
    subprocess.run(["make", "-C", HERE, "liborc.so"] + (["ref"] if ref else []), check=True,
                   stdout=subprocess.DEVNULL)


def _vp(a: np.ndarray):
    return a.ctypes.data_as(C.c_void_p)


def _f3(a):
    return np.ascontiguousarray(a, dtype=np.float32).ctypes.data_as(C.POINTER(C.c_float))


class Oracle:
    """The C restatement.  Methods mutate the structured ``particles`` array in place, like the reference."""

    def __init__(self, **kw):
        path = os.path.join(HERE, "liborc.so")
        if not os.path.exists(path):
            build(ref=False)
        self.lib = L = C.CDLL(path)
        d = dict(DEFAULTS)
        d.update(kw)
        self.params = Params(**d)
        self.p = C.byref(self.params)
        L.orc_kernel_function.restype = C.c_float
        L.orc_lap_kernel.restype = C.c_float
        L.orc_fnv1a_pos.restype = C.c_uint64
        L.orc_scene_default.restype = C.c_size_t
        L.orc_add_particle.restype = C.c_size_t
        L.orc_ref_update_cells.restype = C.c_long
        L.orc_max_threads.restype = C.c_int

    # --- kernels (SphUtils.cpp:55-95)
    def kernel_function(self, i, j) -> float:
        return float(self.lib.orc_kernel_function(self.p, _f3(i), _f3(j)))

    def grad_kernel(self, i, j) -> np.ndarray:
        out = np.zeros(3, np.float32)
        self.lib.orc_grad_kernel(self.p, _f3(i), _f3(j), _vp(out))
        return out

    def lap_kernel(self, i, j) -> float:
        return float(self.lib.orc_lap_kernel(self.p, _f3(i), _f3(j)))

    # --- phases
    def update_density(self, ps, cells=False, threads=0):
        if cells:
            self.lib.orc_update_density_cells(self.p, _vp(ps), C.c_size_t(len(ps)), C.c_int(threads))
        else:
            self.lib.orc_update_density_allpairs(self.p, _vp(ps), C.c_size_t(len(ps)))

    def update_forces(self, ps, cells=False, threads=0):
        if cells:
            self.lib.orc_update_forces_cells(self.p, _vp(ps), C.c_size_t(len(ps)), C.c_int(threads))
        else:
            self.lib.orc_update_forces_allpairs(self.p, _vp(ps), C.c_size_t(len(ps)))

    def viscosity_forces(self, ps, cells=False, threads=0):
        """The dead viscosity_forces (SphUtils.cpp:244-290) on its full neighbourhood; adds into accel."""
        if cells:
            self.lib.orc_viscosity_forces_cells(self.p, _vp(ps), C.c_size_t(len(ps)), C.c_int(threads))
        else:
            self.lib.orc_viscosity_forces_allpairs(self.p, _vp(ps), C.c_size_t(len(ps)))

    # --- FP64 referee (same formulas, double arithmetic, float inputs): how far the reference's own rounding is from exact
    def density_f64(self, ps, threads=0) -> np.ndarray:
        """Unclamped density sum per particle in double."""
        out = np.zeros(len(ps), np.float64)
        self.lib.orc_density_f64_cells(self.p, _vp(ps), C.c_size_t(len(ps)), _vp(out), C.c_int(threads))
        return out

    def forces_f64(self, ps, threads=0):
        """(accel[n,3], scale[n]) in double from the float state in ps (its density / pressure fields as given);
        scale = 1 + sum_j |temp_ij| |gradW_ij| (SURVEY.md H2 normaliser)."""
        acc = np.zeros((len(ps), 3), np.float64)
        scale = np.zeros(len(ps), np.float64)
        self.lib.orc_forces_f64_cells(self.p, _vp(ps), C.c_size_t(len(ps)), _vp(acc), _vp(scale), C.c_int(threads))
        return acc, scale

    def update_posvel(self, ps, spheres=None):
        ns = 0 if spheres is None else len(spheres)
        self.lib.orc_update_posvel(self.p, _vp(ps), C.c_size_t(len(ps)), _vp(spheres) if ns else None, C.c_size_t(ns))

    def collision(self, ps, walls, spheres=None):
        ns = 0 if spheres is None else len(spheres)
        self.lib.orc_collision(self.p, _vp(ps), C.c_size_t(len(ps)), _vp(walls), C.c_size_t(len(walls)),
                               _vp(spheres) if ns else None, C.c_size_t(ns))

    def step(self, ps, walls, nsteps=1, cells=False, threads=0):
        if cells:
            self.lib.orc_step_cells(self.p, _vp(ps), C.c_size_t(len(ps)), _vp(walls), C.c_size_t(len(walls)),
                                    C.c_int(nsteps), C.c_int(threads))
        else:
            self.lib.orc_step_allpairs(self.p, _vp(ps), C.c_size_t(len(ps)), _vp(walls), C.c_size_t(len(walls)),
                                       C.c_int(nsteps))

    # --- grids
    def ref_update_cells(self, ps):
        bbox = np.zeros(6, np.float32)
        dims = np.zeros(3, np.int32)
        cell_of = np.zeros(len(ps), np.int32)
        ncell = self.lib.orc_ref_update_cells(self.p, _vp(ps), C.c_size_t(len(ps)), _vp(bbox), _vp(dims), _vp(cell_of))
        return int(ncell), bbox, dims, cell_of

    def bbox(self, ps) -> np.ndarray:
        bbox = np.zeros(6, np.float32)
        self.lib.orc_bbox(_vp(ps), C.c_size_t(len(ps)), _vp(bbox))
        return bbox

    def grid_from_bbox(self, bbox, cell, max_cells) -> Grid:
        g = Grid()
        cell = np.broadcast_to(np.asarray(cell, np.float32), (3,))
        self.lib.orc_grid_from_bbox(_f3(bbox), _f3(cell), C.c_longlong(max_cells), C.byref(g))
        return g

    @staticmethod
    def make_grid(origin, cell, dims) -> Grid:
        g = Grid()
        g.origin[:] = [float(x) for x in origin]
        g.cell[:] = [float(x) for x in np.broadcast_to(np.asarray(cell, np.float32), (3,))]
        g.dims[:] = [int(x) for x in dims]
        return g

    def grid_build(self, g: Grid, ps):
        n = len(ps)
        ncell = g.dims[0] * g.dims[1] * g.dims[2]
        xyz = np.zeros((n, 3), np.int32)
        key = np.zeros(n, np.int32)
        perm = np.zeros(n, np.int32)
        start = np.zeros(ncell + 1, np.int32)
        self.lib.orc_grid_build(C.byref(g), _vp(ps), C.c_size_t(n), _vp(xyz), _vp(key), _vp(perm), _vp(start))
        return xyz, key, perm, start

    # --- viewer picture (shader arithmetic restated; parity unpinned: no graphics API on this box)
    def default_mvp(self) -> np.ndarray:
        m = np.zeros(16, np.float32)
        self.lib.orc_default_mvp(_vp(m))
        return m

    def render_depth(self, ps, mvp, width=1080, height=920, point_scale=50.0, sprite_radius=0.1, clear_value=1.0, depth_test=0):
        img = np.zeros((height, width), np.float32)
        self.lib.orc_render_depth2(_vp(np.ascontiguousarray(ps)), C.c_size_t(len(ps)), _f3(mvp), C.c_int(width), C.c_int(height),
                                   C.c_float(point_scale), C.c_float(sprite_radius), C.c_float(clear_value), C.c_int(depth_test), _vp(img))
        return img

    # --- scene (main.cpp:424-540)
    def scene_default(self):
        ps = np.zeros(1300, PARTICLE)
        n = self.lib.orc_scene_default(_vp(ps), C.c_size_t(1300), C.c_float(self.params.h),
                                       C.c_float(self.params.rho0), C.c_float(self.params.eps))
        walls = np.zeros(5, WALL)
        self.lib.orc_walls_default(_vp(walls))
        return ps[:n].copy(), walls

    def fnv1a_pos(self, ps) -> int:
        return int(self.lib.orc_fnv1a_pos(_vp(np.ascontiguousarray(ps)), C.c_size_t(len(ps))))

    def max_threads(self) -> int:
        return int(self.lib.orc_max_threads())


class Reference:
    """The unmodified reference (SphUtils) behind oracle/ref_driver.cpp."""

    PATH = os.path.join(HERE, "_ref", "libsphref.so")

    @classmethod
    def available(cls) -> bool:
        return os.path.exists(cls.PATH)

    def __init__(self, **kw):
        if not self.available():
            raise RuntimeError("oracle/_ref/libsphref.so not built (needs /root/reference at build time)")
        self.lib = L = C.CDLL(self.PATH)
        d = dict(DEFAULTS)
        d.update(kw)
        L.ref_create.restype = C.c_void_p
        L.ref_kernel_function.restype = C.c_float
        L.ref_lap_kernel.restype = C.c_float
        L.ref_sizeof_particle.restype = C.c_size_t
        L.ref_offsets.restype = C.c_size_t
        self.h = C.c_void_p(L.ref_create(C.c_float(d["h"]), C.c_float(d["rho0"]), C.c_float(d["c"]),
                                         C.c_float(d["dt"]), C.c_float(d["eps"])))

    def __del__(self):
        try:
            self.lib.ref_destroy(self.h)
        except Exception:
            pass

    def default_mvp(self) -> np.ndarray:
        """MVP of main.cpp:48-60 computed by the reference's vendored GLM (column-major, 16 floats)."""
        m = np.zeros(16, np.float32)
        self.lib.ref_default_mvp(_vp(m))
        return m

    def kernel_function(self, i, j) -> float:
        return float(self.lib.ref_kernel_function(self.h, _f3(i), _f3(j)))

    def grad_kernel(self, i, j) -> np.ndarray:
        out = np.zeros(3, np.float32)
        self.lib.ref_grad_kernel(self.h, _f3(i), _f3(j), _vp(out))
        return out

    def lap_kernel(self, i, j) -> float:
        return float(self.lib.ref_lap_kernel(self.h, _f3(i), _f3(j)))

    def update_density(self, ps):
        self.lib.ref_update_density(self.h, _vp(ps), C.c_size_t(len(ps)))

    def update_forces(self, ps, spheres=None):
        ns = 0 if spheres is None else len(spheres)
        self.lib.ref_update_forces(self.h, _vp(ps), C.c_size_t(len(ps)), _vp(spheres) if ns else None, C.c_size_t(ns))

    def update_posvel(self, ps, spheres=None):
        ns = 0 if spheres is None else len(spheres)
        self.lib.ref_update_posvel(self.h, _vp(ps), C.c_size_t(len(ps)), _vp(spheres) if ns else None, C.c_size_t(ns))

    def collision(self, ps, walls, spheres=None):
        ns = 0 if spheres is None else len(spheres)
        self.lib.ref_collision(self.h, _vp(ps), C.c_size_t(len(ps)), _vp(walls), C.c_size_t(len(walls)),
                               _vp(spheres) if ns else None, C.c_size_t(ns))

    def viscosity_forces(self, ps):
        self.lib.ref_viscosity_forces(self.h, _vp(ps), C.c_size_t(len(ps)))

    def step(self, ps, walls, nsteps=1):
        self.lib.ref_step(self.h, _vp(ps), C.c_size_t(len(ps)), _vp(walls), C.c_size_t(len(walls)), C.c_int(nsteps))

    def update_cells(self, ps):
        bbox = np.zeros(6, np.float32)
        dims = np.zeros(3, np.int32)
        cell_of = np.full(len(ps), -1, np.int32)
        ncell = self.lib.ref_update_cells(self.h, _vp(ps), C.c_size_t(len(ps)), _vp(bbox), _vp(dims), _vp(cell_of))
        return int(ncell), bbox, dims, cell_of

    def offsets(self):
        return {k: int(self.lib.ref_offsets(C.c_int(k))) for k in (0, 1, 2, 3, 4, 5, 6, 10, 11, 20, 21, 22, 23, 24)}
