// Shared device structures and helpers: grid descriptor, bounding-box accumulator, kernel constants, fast MUFU wrappers, cell coordinates.
// Part of libsphb200.so (sm_100a); included through sphb_device.cuh.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace sphb {


constexpr int kMaxWalls = 16;
constexpr int kMaxRows = 1 << 22; // cap on dims[1]*dims[2] ((y,z) rows of cells): bounds the row-group table

// Device-resident description of the uniform grid (rebuilt every step by k_grid_setup).
// Bins behind the last cell for records the multi-GPU exchange marked dead (mass < 0): the counting sort files them
// after every live particle, which removes them without a compaction pass of its own.  Several bins, so that their
// histogram atomics do not all hit one address.
constexpr int kTrashBins = 1024;

struct GridDesc {
    float origin[3];
    float cell[3]; // per axis: cell / cell_x_div in x, cell in y and z
    int dims[3];
    int ncell;     // dims[0]*dims[1]*dims[2]
    int nbins;     // ncell + kTrashBins: counters / scan length (the extra bins collect records marked dead, see k_key_hist)
    int stencil;   // half width of the stencil in y and z cells: ceil(R / cell) (1 when cell >= R)
    int stencil_x; // half width in x cells: ceil(R / cell_x)
};

// Bounding-box accumulator: floats mapped to order-preserving unsigned ints so atomicMin/Max work.
struct BBoxAcc {
    unsigned int lo[3];
    unsigned int hi[3];
};

// Everything the physics kernels need, passed by value (lives in the kernel parameter space).
struct Consts {
    float h, h2, inv_h, inv_h2; // h2 = h*h (the support radius R), inv_h2 = 1/(h*h)
    float R2;                   // R*R: a pair interacts iff d2 < R2
    float Kw;                   // 315 / (64 pi h^9)
    float Kg;                   // -45 / (pi h^6)
    float Kl;                   // 45 / (pi h^6)
    float rho0, eos_k, dt;
    float gx, gy, gz;
    float alpha, beta, eta_h2;  // eta_h2 = eta*h*h
    float one_plus_e;           // 1 + restitution
    double wall_off, wall_lim;
    int nwalls;
    float wn[kMaxWalls][3];     // wall normal (as given)
    float wc[kMaxWalls][3];     // wall centre
    float wu[kMaxWalls][3];     // normalize(normal), computed on the host with IEEE float ops
    int domain_fixed;
    float dom_lo[3], dom_hi[3];
    long long max_cells;
    float lap_coef; // coefficient of the optional Laplacian viscosity (SphUtils.cpp:277), 0 when off
    // Multi-GPU: the box outside which nobody needs a particle's density -- the slab (and the periodic box) grown by ONE support
    // radius.  Ghost layers are two radii wide so that the inner half can have its density recomputed here; the outer half is
    // only ever a neighbour in that sum, never a target.  (-inf, +inf) on an engine that exchanges no slabs.
    float need_lo[3], need_hi[3];
    float cell;   // cell edge in y and z
    float cell_x; // cell edge in x (cell / cell_x_div)
    int stencil_x;
};

// ---------------------------------------------------------------------------------------------------
// small helpers

__device__ __forceinline__ unsigned int f2ord(float f) {
    unsigned int u = __float_as_uint(f);
    return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
__device__ __forceinline__ float ord2f(unsigned int u) {
    return __uint_as_float((u & 0x80000000u) ? (u & 0x7fffffffu) : ~u);
}

__device__ __forceinline__ float4 ldg4(const float4* p) { return __ldg(p); }

// MUFU approximations without the denormal fix-up code the C library versions carry (magnitudes here are nowhere
// near 1e-38).  <= 2 ulp each; seeds of the pm_* functions below.
__device__ __forceinline__ float fast_rsqrt(float x) {
    float r;
    asm("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
    return r;
}
__device__ __forceinline__ float fast_rcp(float x) {
    float r;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
    return r;
}
__device__ __forceinline__ float fast_sqrt(float x) {
    float r;
    asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
    return r;
}

// The special functions of the pair body (SphUtils.cpp:223-228 divides and takes square roots in IEEE float).
// SPHB_PAIR_MATH selects how close to that they get:
//   0  raw MUFU approximations (<= 2 ulp each; round 1)
//   1  MUFU seed + one Newton step for 1/x and 1/sqrt(x) (<= 1 ulp), correctly rounded sqrt     [default]
//   2  correctly rounded everywhere (__frcp_rn / __frsqrt_rn / __fsqrt_rn)
// Measured on B200 (Numerics): time and per-particle error of all three.
#ifndef SPHB_PAIR_MATH
#define SPHB_PAIR_MATH 1
#endif
__device__ __forceinline__ float pm_rsqrt(float x) {
#if SPHB_PAIR_MATH == 0
    return fast_rsqrt(x);
#elif SPHB_PAIR_MATH == 1
    const float r = fast_rsqrt(x);
    return fmaf(r, fmaf(-0.5f * x * r, r, 0.5f), r); // r + r (1/2 - x r^2 / 2); x = 0 stays non-finite -> NaN downstream, like normalize(0)
#else
    return __frsqrt_rn(x);
#endif
}
__device__ __forceinline__ float pm_rcp(float x) {
#if SPHB_PAIR_MATH == 0
    return fast_rcp(x);
#elif SPHB_PAIR_MATH == 1
    const float r = fast_rcp(x);
    return fmaf(r, fmaf(-x, r, 1.0f), r);
#else
    return __frcp_rn(x);
#endif
}
__device__ __forceinline__ float pm_sqrt(float x) {
#if SPHB_PAIR_MATH == 0
    return fast_sqrt(x);
#else
    return __fsqrt_rn(x);
#endif
}

// aux record (rho, P, sqrt|P/rho|, 1/rho).  The last two are helpers of the pair body (SphUtils.cpp:223,228),
// never returned to the host; they are computed by this ONE function wherever they are needed (density
// kernel, upload, per-neighbour re-derivation) so that every path sees the same bits.
__device__ __forceinline__ float4 make_aux(float rho, float P) {
    const float inv = pm_rcp(rho);
    return make_float4(rho, P, pm_sqrt(fabsf(P * inv)), inv);
}

// Cell coordinate along one axis.  Same expression as SphUtils.cpp:38-40, floor((p - min) / cell), in
// IEEE float (no FMA, true division), then clamped to the grid so that strays stay adjacent to their
// neighbours.  Bit-exact against the CPU grid build the tests use.
__device__ __forceinline__ int grid_coord(float x, float origin, float cell, int dim) {
    float f = floorf(__fdiv_rn(__fsub_rn(x, origin), cell));
    if (!(f >= 0.0f)) return 0;
    if (f >= (float)dim) return dim - 1;
    return (int)f;
}

} // namespace sphb
