"""Python mirror of the reference's step-engine interface (class SphUtils, SPH/SphUtils.hpp:16-44).

Same method names, argument meaning and (absence of) error behaviour as the reference, over numpy
structured arrays laid out like the reference's Particle / Wall / Sphere records, so that a parity
test reads like the reference's own ``step_scene`` (main.cpp:280-303):

    sph = SphUtils(0.45, 10.0, 100.0, 0.033, 0.01)
    sph.update_density(particles)
    sph.update_forces(particles, spheres)
    sph.update_posvel(particles, spheres)
    sph.collision(particles, walls, spheres)

Each call mutates ``particles`` in place exactly where the reference does (density+pressure; accel;
vel+pos; pos+vel).  All compute happens in libsphb200.so on the GPU; this file only moves records.

Two sync policies (the C++ shim sph_b200/host/SphUtils.hpp has the same two):
* ``sync="call"`` (default, strict drop-in): every method uploads the records, runs its phase on the
  device and downloads the fields that phase writes -- host memory always looks like the reference's.
* ``sync="step"``: update_density uploads, collision() downloads everything; the phases in between
  stay on the device.  This is what a viewer loop needs (main.cpp only reads pos after the step).
"""
from __future__ import annotations

import numpy as np

from . import capi
from .capi import F_ACCEL, F_ALL, F_DENSITY, F_POS, F_VEL, PARTICLE


class SphUtils:
    def __init__(self, nsmooth_length=0.45, nrho0=10.0, nc=100.0, ndt=0.033, nepsilon=0.01, *, device=0, sync="call",
                 **param_overrides):
        if sync not in ("call", "step"):
            raise ValueError("sync must be 'call' or 'step'")
        self.sync = sync
        self.params = capi.default_params(h=nsmooth_length, rho0=nrho0, c=nc, dt=ndt, eps=nepsilon, **param_overrides)
        self.engine = capi.Engine(self.params, device=device)
        self._walls_key = None
        self._resident = False  # device state is ahead of the host records (sync="step")

    # --- kernel helpers (SphUtils.cpp:55-95), evaluated on the device
    def kernel_function(self, i, j) -> float:
        return float(self.engine.eval_kernels(i, j)[0][0])

    def grad_kernel(self, i, j) -> np.ndarray:
        return self.engine.eval_kernels(i, j)[1][0]

    def lap_kernel(self, i, j) -> float:
        return float(self.engine.eval_kernels(i, j)[2][0])

    # --- the five public methods
    def _up(self, particles):
        if self.sync == "call" or not self._resident:
            self.engine.upload(particles)

    def _down(self, particles, mask):
        if self.sync == "call":
            self.engine.download(particles, mask)

    def update_cells(self, particles):
        self._up(particles)
        self.engine.update_cells()

    def update_density(self, particles):
        # first call of a step: the host records are the truth (the vector may have grown, main.cpp:441)
        self.engine.upload(particles)
        self._resident = self.sync == "step"
        self.engine.update_density()
        self._down(particles, F_DENSITY)

    def _spheres_in(self, spheres):
        self.engine.set_spheres(spheres)

    def _spheres_out(self, spheres):
        if spheres is not None and len(spheres):
            self.engine.get_spheres(spheres)       # the reference mutates the caller's Sphere records in place

    def update_forces(self, particles, spheres=None):
        self._spheres_in(spheres)
        self._up(particles)
        self.engine.update_forces()
        self._down(particles, F_ACCEL)
        self._spheres_out(spheres)

    def update_posvel(self, particles, spheres=None):
        self._spheres_in(spheres)
        self._up(particles)
        self.engine.update_posvel()
        self._down(particles, F_POS | F_VEL)
        self._spheres_out(spheres)

    def collision(self, particles, walls, spheres=None):
        self._spheres_in(spheres)
        key = walls.tobytes()
        if key != self._walls_key:
            self.engine.set_walls(walls)
            self._walls_key = key
        self._up(particles)
        self.engine.collision()
        if self.sync == "call":
            self.engine.download(particles, F_POS | F_VEL)
        else:
            self.engine.download(particles, F_ALL)
            self._resident = False
        self._spheres_out(spheres)

    # --- fused fast path: step_scene (main.cpp:280-303) x nsteps on host records
    def step_scene(self, particles, walls, nsteps=1, mask=F_ALL, spheres=None):
        key = walls.tobytes()
        if key != self._walls_key:
            self.engine.set_walls(walls)
            self._walls_key = key
        self._spheres_in(spheres)
        self.engine.step_aos(particles, nsteps, mask)
        self._spheres_out(spheres)
