"""Synthetic particle sets for the parity tests and the bench.

* ``default_scene``  -- main.cpp initParticles/initWalls restated (config 1, 972 particles).
* ``dam_break``      -- jittered lattice block in a tank (configs 2 and 3).
* ``block_drop``     -- pool + falling block (config 4).

Jitter and velocities come from a counter-based hash (splitmix64 of seed, particle index, axis), so
any slice of a scene can be generated independently (per rank, per chunk) and is reproducible.
All scenes return (particles[PARTICLE], walls[WALL]) as numpy structured arrays in the reference's
own record layout; they feed the GPU engine and the CPU oracle alike.
"""
from __future__ import annotations

import numpy as np

from .capi import PARTICLE, WALL

H = np.float32(0.45)        # main.cpp:34
RHO0 = np.float32(10.0)     # main.cpp:36
EPS = np.float32(0.01)      # main.cpp:40
DT = np.float32(0.033)      # main.cpp:30


def particle_mass(h=H, rho0=RHO0) -> np.float32:
    """main.cpp:429: .1f*h*h*h*rho0 in float, left to right."""
    h = np.float32(h)
    return np.float32(np.float32(np.float32(np.float32(np.float32(0.1) * h) * h) * h) * np.float32(rho0))


def make_particles(pos, vel=None, h=H, rho0=RHO0, eps=EPS) -> np.ndarray:
    """Records as addParticle builds them (main.cpp:424-442): density rho0+eps, pressure 1, thermal 1."""
    pos = np.asarray(pos, np.float32).reshape(-1, 3)
    ps = np.zeros(len(pos), PARTICLE)
    ps["pos"] = pos
    if vel is not None:
        ps["vel"] = np.asarray(vel, np.float32).reshape(-1, 3)
    ps["mass"] = particle_mass(h, rho0)
    ps["density"] = np.float32(rho0) + np.float32(eps)
    ps["pressure"] = 1.0
    ps["thermal"] = 1.0
    return ps


def make_walls(planes) -> np.ndarray:
    """planes: iterable of (normal xyz, center xyz)."""
    w = np.zeros(len(planes), WALL)
    for i, (n, c) in enumerate(planes):
        w[i]["normal"] = n
        w[i]["center"] = c
        w[i]["xlength"] = 1.0
        w[i]["ylength"] = 1.0
    return w


def default_scene(max_particles: int = 1300):
    """main.cpp:490-505 and :508-540.  The loop counters are floats advanced by the double literal 0.1,
    so each axis stops one short of what the bounds suggest: 18 x 3 x 18 = 972 particles."""
    def axis(lo, hi_cmp):
        vals = []
        v = np.float32(lo)
        while hi_cmp(v):
            vals.append(v)
            v = np.float32(np.float64(v) + 0.1)
        return vals

    fmax = np.float32(0.9)
    xs = axis(-0.9, lambda v: v <= fmax)
    ys = axis(-0.9, lambda v: np.float64(v) <= -0.6)
    zs = axis(-0.9, lambda v: v <= fmax)
    pos = np.array([(i, j, k) for i in xs for j in ys for k in zs], np.float32)[:max_particles]
    walls = make_walls([
        ((1, 0, 0), (-1, 0, 0)),   # w1
        ((-1, 0, 0), (1, 0, 0)),   # w2
        ((0, 0, -1), (0, 0, 1)),   # w4 (front)
        ((0, 0, 1), (0, 0, -1)),   # w5 (back)
        ((0, 1, 0), (0, -1, 0)),   # w3 (bottom) -- pushed last, main.cpp:535-539
    ])
    return make_particles(pos), walls


# --- counter-based RNG -------------------------------------------------------------------------------

def _splitmix64(x: np.ndarray) -> np.ndarray:
    x = (x + np.uint64(0x9E3779B97F4A7C15)).astype(np.uint64)
    z = x
    z = (z ^ (z >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
    z = (z ^ (z >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
    return z ^ (z >> np.uint64(31))


def hash_uniform(seed: int, index: np.ndarray, axis: int, stream: int = 0) -> np.ndarray:
    """U[0,1) float32 from splitmix64(seed ^ index*8 + axis + 4*stream), 24 mantissa bits."""
    with np.errstate(over="ignore"):
        k = (index.astype(np.uint64) * np.uint64(8) + np.uint64(axis + 4 * stream)) ^ np.uint64(seed)
        z = _splitmix64(_splitmix64(k))
    return ((z >> np.uint64(40)).astype(np.float32)) * np.float32(1.0 / (1 << 24))


def lattice_block(nx, ny, nz, spacing, corner, seed=1234, jitter=0.2, vel_amp=0.0, index0=0, i_range=None):
    """nx x ny x nz lattice (x slowest, z fastest, like main.cpp's loop nest), position jitter
    U(-jitter, jitter)*spacing per axis and velocity U(-vel_amp, vel_amp) per axis.
    i_range=(a, b) generates only lattice planes a <= ix < b (particle indices stay global)."""
    a, b = (0, nx) if i_range is None else i_range
    ix, iy, iz = np.meshgrid(np.arange(a, b), np.arange(ny), np.arange(nz), indexing="ij")
    ix, iy, iz = ix.ravel(), iy.ravel(), iz.ravel()
    idx = (index0 + (ix.astype(np.int64) * ny + iy) * nz + iz).astype(np.uint64)
    s = np.float32(spacing)
    pos = np.empty((len(idx), 3), np.float32)
    vel = np.zeros((len(idx), 3), np.float32)
    for ax, ii in enumerate((ix, iy, iz)):
        u = hash_uniform(seed, idx, ax, 0)
        pos[:, ax] = np.float32(corner[ax]) + s * (ii.astype(np.float32) + np.float32(jitter) * (np.float32(2.0) * u - np.float32(1.0)))
        if vel_amp:
            v = hash_uniform(seed, idx, ax, 1)
            vel[:, ax] = np.float32(vel_amp) * (np.float32(2.0) * v - np.float32(1.0))
    return pos, vel


def tank_walls(tx, tz):
    """Five planes of an open-top tank [0,tx] x [0,inf) x [0,tz], in the reference's order
    (left, right, front, back, bottom -- main.cpp:535-539)."""
    return make_walls([
        ((1, 0, 0), (0, 0, 0)),
        ((-1, 0, 0), (tx, 0, 0)),
        ((0, 0, -1), (0, 0, tz)),
        ((0, 0, 1), (0, 0, 0)),
        ((0, 1, 0), (0, 0, 0)),
    ])


def dam_break(nx=32, ny=64, nz=32, stressed=False, seed=1234):
    """block with its corner 0.1 from the walls x=0, y=0, z=0; tank length
    2.5x the block.  rest: spacing 0.1, v=0.  stressed: spacing 0.085, v ~ U(-0.2,0.2)^3 so that rho > rho0,
    P > 0 and the artificial-viscosity branch (v.r < 0) are exercised."""
    s = 0.085 if stressed else 0.1
    pos, vel = lattice_block(nx, ny, nz, s, (0.1, 0.1, 0.1), seed=seed, jitter=0.2, vel_amp=0.2 if stressed else 0.0)
    tx = 2.5 * nx * s + 0.2
    tz = nz * s + 0.2
    return make_particles(pos, vel), tank_walls(tx, tz)


def dam_break_domain(nx=32, ny=64, nz=32, stressed=False):
    """A fixed grid domain that holds the dam-break tank with splash room above."""
    s = 0.085 if stressed else 0.1
    tx = 2.5 * nx * s + 0.2
    tz = nz * s + 0.2
    return (-0.25, -0.25, -0.25), (tx + 0.25, 1.5 * ny * s + 0.5, tz + 0.25)


def block_drop(scale=1, seed=1234):
    """SURVEY.md 8(d) config 4 at scale=1: pool 256x192x256 on the tank floor + block 128x256x128 centred
    above it moving down at 1; 2^24 particles.  scale divides every lattice dimension (scale=4 -> 2^18)."""
    px, py, pz = 256 // scale, 192 // scale, 256 // scale
    bx, by, bz = 128 // scale, 256 // scale, 128 // scale
    s = 0.1
    pool_pos, pool_vel = lattice_block(px, py, pz, s, (0.1, 0.1, 0.1), seed=seed, jitter=0.2)
    top = 0.1 + py * s + 0.3
    off = 0.1 + (px - bx) * s / 2
    blk_pos, blk_vel = lattice_block(bx, by, bz, s, (off, top, off), seed=seed, jitter=0.2, index0=px * py * pz)
    blk_vel[:, 1] = -1.0
    pos = np.concatenate([pool_pos, blk_pos])
    vel = np.concatenate([pool_vel, blk_vel])
    t = px * s + 0.2
    return make_particles(pos, vel), tank_walls(t, t)


def block_drop_section(rank, world, scale=1, seed=1234):
    """Weak-scaling variant of block_drop: `world` tank sections of x-length 25.6/scale side by side in one long
    tank, each with the pool + falling block pattern of config 4; returns the particles of section `rank`
    (positions in GLOBAL coordinates).  rank=0, world=1 is block_drop(scale) itself."""
    px, py, pz = 256 // scale, 192 // scale, 256 // scale
    bx, by, bz = 128 // scale, 256 // scale, 128 // scale
    s = 0.1
    per = px * py * pz + bx * by * bz
    x0 = px * s * rank
    pool_pos, pool_vel = lattice_block(px, py, pz, s, (0.1 + x0, 0.1, 0.1), seed=seed, jitter=0.2, index0=per * rank)
    top = 0.1 + py * s + 0.3
    off = 0.1 + (px - bx) * s / 2
    blk_pos, blk_vel = lattice_block(bx, by, bz, s, (off + x0, top, off), seed=seed, jitter=0.2, index0=per * rank + px * py * pz)
    blk_vel[:, 1] = -1.0
    return make_particles(np.concatenate([pool_pos, blk_pos]), np.concatenate([pool_vel, blk_vel]))


def shallow_pool(nx=96, nz=96, layers=3, seed=1234, jitter=0.2):
    """The reference's own kind of scene (main.cpp:490-505: a 3-layer sheet of particles 0.1 above the floor of a
    box), at any width: the only kind its scheme (dt = 0.033, k = 100, CFL ~ 3.7) sustains for hundreds of steps.
    Used for the long-run conservation checks."""
    pos, vel = lattice_block(nx, layers, nz, 0.1, (0.1, 0.1, 0.1), seed=seed, jitter=jitter)
    return make_particles(pos, vel), tank_walls(nx * 0.1 + 0.2, nz * 0.1 + 0.2)
