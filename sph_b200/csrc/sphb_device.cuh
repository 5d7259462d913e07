// Device-side code of the SPH step engine: grid build (cell keys, stable counting sort, reorder),
// density/pressure, fused force+integrate+collide, AoS<->SoA conversion.  sm_100a only.
//
// Reference restated (not ported: the reference is a scalar all-pairs CPU loop):
//   SPH/SphUtils.cpp:8-47   update_cells      -> k_bbox, k_grid_setup, k_key_hist, k_scan_*, k_scatter, k_rank_gather
//   SPH/SphUtils.cpp:55-95  W, gradW, lapW    -> w_shape(), grad_shape(), k_eval_kernels
//   SPH/SphUtils.cpp:100-149 update_density   -> k_density
//   SPH/SphUtils.cpp:157-239,295-304 forces   -> k_force<...>
//   SPH/SphUtils.cpp:321-335 update_posvel    -> integrate()
//   SPH/SphUtils.cpp:343-372 collision (walls)-> collide_walls()
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace sphb {

constexpr int kMaxWalls = 16;
constexpr int kMaxRows = 1 << 22; // cap on dims[1]*dims[2] ((y,z) rows of cells): bounds the row-group table

// Device-resident description of the uniform grid (rebuilt every step by k_grid_setup).
struct GridDesc {
    float origin[3];
    float cell[3]; // per axis (equal unless an axis is periodic and the cell was stretched to tile it)
    int dims[3];
    int ncell;     // dims[0]*dims[1]*dims[2]
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
    int periodic[3];
    long long max_cells;
    float lap_coef; // coefficient of the optional Laplacian viscosity (SphUtils.cpp:277), 0 when off
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

// MUFU approximations without the denormal fix-up code the C library versions carry (the pair body runs
// ~40 times per particle per step; magnitudes here are nowhere near 1e-38).  <= 2 ulp each.
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

// aux record (rho, P, sqrt|P/rho|, 1/rho).  The last two are helpers of the pair body (SphUtils.cpp:223,228),
// never returned to the host; they are computed by this ONE function wherever they are needed (density
// kernel, upload, per-neighbour re-derivation) so that every path sees the same bits.
__device__ __forceinline__ float4 make_aux(float rho, float P) {
    const float inv = fast_rcp(rho);
    return make_float4(rho, P, fast_sqrt(fabsf(P * inv)), inv);
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

// ---------------------------------------------------------------------------------------------------
// K0: bounding box (SphUtils.cpp:11-23).  Grid-stride, warp shuffle, block reduce, then atomics that
// are skipped when they cannot improve the global value.

__device__ __forceinline__ void bbox_block_commit(float lo[3], float hi[3], BBoxAcc* acc) {
    __shared__ float s_lo[3][32];
    __shared__ float s_hi[3][32];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarp = (blockDim.x + 31) >> 5;
#pragma unroll
    for (int a = 0; a < 3; a++) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            lo[a] = fminf(lo[a], __shfl_xor_sync(0xffffffffu, lo[a], o));
            hi[a] = fmaxf(hi[a], __shfl_xor_sync(0xffffffffu, hi[a], o));
        }
        if (lane == 0) { s_lo[a][warp] = lo[a]; s_hi[a][warp] = hi[a]; }
    }
    __syncthreads();
    if (warp == 0) {
#pragma unroll
        for (int a = 0; a < 3; a++) {
            float l = lane < nwarp ? s_lo[a][lane] : INFINITY;
            float h = lane < nwarp ? s_hi[a][lane] : -INFINITY;
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) {
                l = fminf(l, __shfl_xor_sync(0xffffffffu, l, o));
                h = fmaxf(h, __shfl_xor_sync(0xffffffffu, h, o));
            }
            if (lane == 0) {
                unsigned int ul = f2ord(l), uh = f2ord(h);
                if (l <= h) { // at least one finite sample
                    if (ul < *((volatile unsigned int*)&acc->lo[a])) atomicMin(&acc->lo[a], ul);
                    if (uh > *((volatile unsigned int*)&acc->hi[a])) atomicMax(&acc->hi[a], uh);
                }
            }
        }
    }
}

__global__ void __launch_bounds__(256) k_bbox_reset(BBoxAcc* acc) {
    if (threadIdx.x < 3) {
        acc->lo[threadIdx.x] = f2ord(INFINITY);
        acc->hi[threadIdx.x] = f2ord(-INFINITY);
    }
}

__global__ void __launch_bounds__(256) k_bbox(const float4* __restrict__ posm, int n, BBoxAcc* acc) {
    float lo[3] = {INFINITY, INFINITY, INFINITY}, hi[3] = {-INFINITY, -INFINITY, -INFINITY};
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        float4 p = ldg4(posm + i);
        // fminf/fmaxf drop NaNs, like the reference's `if (x < min)` comparisons
        lo[0] = fminf(lo[0], p.x); hi[0] = fmaxf(hi[0], p.x);
        lo[1] = fminf(lo[1], p.y); hi[1] = fmaxf(hi[1], p.y);
        lo[2] = fminf(lo[2], p.z); hi[2] = fmaxf(hi[2], p.z);
    }
    bbox_block_commit(lo, hi, acc);
}

// ---------------------------------------------------------------------------------------------------
// Grid description from the bounding box (SphUtils.cpp:25-27: n = (int)ceil(extent/cell) + 1) or from
// the fixed domain; then clear the cell counters [0, ncell].  Every thread derives the same dims;
// thread 0 of block 0 publishes them.  The tests check the dims against a CPU restatement of the same rule.

__device__ __forceinline__ void derive_grid(const Consts& c, const BBoxAcc* acc, GridDesc& g) {
    float lo[3], hi[3];
    if (c.domain_fixed) {
#pragma unroll
        for (int a = 0; a < 3; a++) { lo[a] = c.dom_lo[a]; hi[a] = c.dom_hi[a]; }
    } else {
#pragma unroll
        for (int a = 0; a < 3; a++) {
            lo[a] = ord2f(acc->lo[a]);
            hi[a] = ord2f(acc->hi[a]);
            if (!(lo[a] <= hi[a])) { lo[a] = 0.0f; hi[a] = 0.0f; } // no finite particle
        }
    }
#pragma unroll
    for (int a = 0; a < 3; a++) {
        g.origin[a] = lo[a];
        g.cell[a] = a == 0 ? c.cell_x : c.cell;
        float ext = ceilf(__fdiv_rn(__fsub_rn(hi[a], lo[a]), g.cell[a]));
        int d = (ext < 1048575.0f) ? (int)ext + 1 : 1048576; // also catches inf/NaN extents
        g.dims[a] = d < 1 ? 1 : d;
    }
    while ((long long)g.dims[0] * g.dims[1] * g.dims[2] > c.max_cells) {
        int a = 0;
        if (g.dims[1] > g.dims[a]) a = 1;
        if (g.dims[2] > g.dims[a]) a = 2;
        if (g.dims[a] == 1) break;
        g.dims[a] = (g.dims[a] + 1) / 2;
    }
    while ((long long)g.dims[1] * g.dims[2] > kMaxRows) { // same policy for the number of (y,z) rows
        const int a = g.dims[2] > g.dims[1] ? 2 : 1;
        g.dims[a] = (g.dims[a] + 1) / 2;
    }
    g.ncell = g.dims[0] * g.dims[1] * g.dims[2];
    g.stencil = (c.cell >= c.h2) ? 1 : 2;
    g.stencil_x = c.stencil_x;
}

__global__ void __launch_bounds__(256) k_grid_setup(Consts c, const BBoxAcc* acc, GridDesc* out, int* __restrict__ cell_count) {
    GridDesc g;
    derive_grid(c, acc, g);
    if (blockIdx.x == 0 && threadIdx.x == 0) *out = g;
    const int total = g.ncell + 1;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) cell_count[i] = 0;
}

// ---------------------------------------------------------------------------------------------------
// K1: cell key + histogram.  slot = arrival rank inside the cell (not deterministic; k_rank_gather
// replaces it by the stable rank).  Also re-arms the bbox accumulator for this step's force kernel.

__device__ __forceinline__ int cell_key(const GridDesc& g, float x, float y, float z, int& cx, int& cy, int& cz) {
    cx = grid_coord(x, g.origin[0], g.cell[0], g.dims[0]);
    cy = grid_coord(y, g.origin[1], g.cell[1], g.dims[1]);
    cz = grid_coord(z, g.origin[2], g.cell[2], g.dims[2]);
    return cx + g.dims[0] * (cy + g.dims[1] * cz);
}

__global__ void __launch_bounds__(256) k_key_hist(const GridDesc* __restrict__ gd, const float4* __restrict__ posm, int n,
                                                  int* __restrict__ key, int* __restrict__ slot, int* __restrict__ cell_count,
                                                  BBoxAcc* acc, int rearm_bbox) {
    const GridDesc g = *gd;
    if (rearm_bbox && blockIdx.x == 0 && threadIdx.x < 3) {
        acc->lo[threadIdx.x] = f2ord(INFINITY);
        acc->hi[threadIdx.x] = f2ord(-INFINITY);
    }
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    float4 p = ldg4(posm + i);
    int cx, cy, cz;
    int k = cell_key(g, p.x, p.y, p.z, cx, cy, cz);
    key[i] = k;
    slot[i] = atomicAdd(cell_count + k, 1);
}

// ---------------------------------------------------------------------------------------------------
// K2: exclusive scan of the cell counters, in place (count -> cell_start), three passes sized by the
// device-side ncell (the host only knows the capacity).  Tile = 2048 ints per block of 256 threads.

constexpr int kScanThreads = 256;
constexpr int kScanPerThread = 8;
constexpr int kScanTile = kScanThreads * kScanPerThread;

__device__ __forceinline__ int block_exclusive_scan(int v, int* total) {
    // returns the exclusive prefix of v over the block (kScanThreads threads); *total = block sum
    __shared__ int s_warp[kScanThreads / 32];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    int inc = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        int t = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc += t;
    }
    if (lane == 31) s_warp[warp] = inc;
    __syncthreads();
    if (warp == 0) {
        int w = lane < kScanThreads / 32 ? s_warp[lane] : 0;
#pragma unroll
        for (int o = 1; o < kScanThreads / 32; o <<= 1) {
            int t = __shfl_up_sync(0xffffffffu, w, o);
            if (lane >= o) w += t;
        }
        if (lane < kScanThreads / 32) s_warp[lane] = w; // inclusive over warps
    }
    __syncthreads();
    int warp_off = warp ? s_warp[warp - 1] : 0;
    *total = s_warp[kScanThreads / 32 - 1];
    __syncthreads(); // s_warp is reused by the next call
    return warp_off + inc - v;
}

__global__ void __launch_bounds__(kScanThreads) k_scan_reduce(const GridDesc* __restrict__ gd, const int* __restrict__ data, int* __restrict__ block_sums) {
    const int total = gd->ncell + 1;
    const int base = blockIdx.x * kScanTile;
    if (base >= total) return;
    int s = 0;
#pragma unroll
    for (int k = 0; k < kScanPerThread; k++) {
        int i = base + k * kScanThreads + threadIdx.x;
        if (i < total) s += data[i];
    }
    int tot;
    block_exclusive_scan(s, &tot);
    if (threadIdx.x == 0) block_sums[blockIdx.x] = tot;
}

// one block: exclusive scan of the per-tile sums, in place
__global__ void __launch_bounds__(kScanThreads) k_scan_top(const GridDesc* __restrict__ gd, int* __restrict__ block_sums) {
    const int ntile = (gd->ncell + 1 + kScanTile - 1) / kScanTile;
    int carry = 0;
    for (int base = 0; base < ntile; base += kScanThreads) {
        int i = base + threadIdx.x;
        int v = i < ntile ? block_sums[i] : 0;
        int tot;
        int ex = block_exclusive_scan(v, &tot);
        if (i < ntile) block_sums[i] = carry + ex;
        carry += tot;
    }
}

__global__ void __launch_bounds__(kScanThreads) k_scan_apply(const GridDesc* __restrict__ gd, int* __restrict__ data, const int* __restrict__ block_sums) {
    const int total = gd->ncell + 1;
    const int base = blockIdx.x * kScanTile;
    if (base >= total) return;
    // each thread owns kScanPerThread consecutive ints
    int v[kScanPerThread];
    int s = 0;
    const int first = base + threadIdx.x * kScanPerThread;
#pragma unroll
    for (int k = 0; k < kScanPerThread; k++) {
        v[k] = (first + k < total) ? data[first + k] : 0;
        s += v[k];
    }
    int tot;
    int ex = block_exclusive_scan(s, &tot) + block_sums[blockIdx.x];
#pragma unroll
    for (int k = 0; k < kScanPerThread; k++) {
        if (first + k < total) data[first + k] = ex;
        ex += v[k];
    }
}

// ---------------------------------------------------------------------------------------------------
// K3: scatter (original id, current index) pairs into their cell segment in arrival order, then
// k_rank_gather replaces the arrival order by the STABLE order (ascending original id inside a cell,
// i.e. exactly what a CPU stable counting sort of the uploaded order produces) and moves the state.

__global__ void __launch_bounds__(256) k_scatter(const int* __restrict__ key, const int* __restrict__ slot, const int* __restrict__ oid,
                                                 const int* __restrict__ cell_start, int2* __restrict__ pairs, int n) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    pairs[cell_start[key[i]] + slot[i]] = make_int2(oid[i], i);
}

// A cell that holds more than this many particles (a clamped boundary cell full of strays, or everything gone
// NaN and filed under cell 0) keeps its ARRIVAL order instead: counting smaller ids is quadratic in the cell size.
// Still a valid permutation, just not the stable one (and not reproducible run to run) for that cell.
constexpr int kStableRankLimit = 4096;

template <bool kCarryAll>
__global__ void __launch_bounds__(256) k_rank_gather(const int2* __restrict__ pairs, const int* __restrict__ key, const int* __restrict__ cell_start,
                                                     const float4* __restrict__ posm_in, const float4* __restrict__ vel_in,
                                                     const float4* __restrict__ aux_in, const float4* __restrict__ acc_in,
                                                     float4* __restrict__ posm_out, float4* __restrict__ vel_out,
                                                     float4* __restrict__ aux_out, float4* __restrict__ acc_out,
                                                     int* __restrict__ oid_out, int n) {
    int p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= n) return;
    const int2 me = pairs[p];
    const int k = key[me.y];
    const int b = cell_start[k], e = cell_start[k + 1];
    int rank = 0;
    if (e - b <= kStableRankLimit) {
        for (int q = b; q < e; q++) rank += (pairs[q].x < me.x);
    } else {
        rank = p - b; // arrival order
    }
    const int dst = b + rank;
    posm_out[dst] = ldg4(posm_in + me.y);
    vel_out[dst] = ldg4(vel_in + me.y);
    if (kCarryAll) {
        aux_out[dst] = ldg4(aux_in + me.y);
        acc_out[dst] = ldg4(acc_in + me.y);
    }
    oid_out[dst] = me.x;
}

// ---------------------------------------------------------------------------------------------------
// SPH kernel shapes (SphUtils.cpp:55-95).  The reference evaluates the polynomial in double and rounds
// once; here everything is FP32 with the constant factor applied after the neighbour sum
// (tolerance-checked: <= 1e-5 relative on density and acceleration, tests/test_gpu_parity.py).

// (h^2 - q^2)^3 with q^2 = d2/h^2, zero outside the support (q <= h  <=>  d2 <= h^4)
__device__ __forceinline__ float w_shape(const Consts& c, float d2) {
    float t = fmaxf(fmaf(-d2, c.inv_h2, c.h2), 0.0f);
    return t * t * t;
}

// Candidate runs.  Around cell (cx,cy,cz) there are (2S+1)^2 rows (dz, dy); along x the 2Sx+1 cells of a
// row are contiguous in the sorted arrays, so each row is ONE run [jb, je) of sorted slots.  Cells are
// shorter in x than in y,z (cell_x = cell / cell_x_div): the run then spans 2R + 2 cell_x instead of 3R,
// i.e. ~30 % fewer candidates to test for free (same number of runs, same two table look-ups per run).  Rows are
// numbered canonically r = (dz+S)*(2S+1) + (dy+S) so that the density and force kernels agree on them;
// a row outside the grid is the empty run.
__device__ __forceinline__ void row_run(const GridDesc& g, const int* __restrict__ cell_start, int cx, int cy, int cz, int r,
                                        int& jb, int& je) {
    const int S = g.stencil, W = 2 * S + 1, Sx = g.stencil_x;
    const int z = cz + r / W - S, y = cy + r % W - S;
    jb = je = 0;
    if (z < 0 || z >= g.dims[2] || y < 0 || y >= g.dims[1]) return;
    const int x0 = max(cx - Sx, 0), x1 = min(cx + Sx, g.dims[0] - 1);
    const int row = g.dims[0] * (y + g.dims[1] * z);
    jb = __ldg(cell_start + row + x0);
    je = __ldg(cell_start + row + x1 + 1);
}

// Neighbour masks.  While the density kernel tests every candidate of a run against the support radius
// anyway, it records the outcome as one bit per candidate, two 32-bit words per row:
// mask[(2r+w)*n + i] bit k  <=>  sorted slot jb_r + 32w + k is a neighbour of i (d2 < R^2, self excluded).  The force kernel then visits neighbours only --
// no distance test on the ~85 % of candidates that fail it -- and its lanes stay converged on the heavy
// pair body.  Candidates beyond the 64th of a run (crowded cells) are re-tested by the force kernel.
constexpr int kMaskBits = 64;

// ---------------------------------------------------------------------------------------------------
// K4: density + equation of state (SphUtils.cpp:100-149) + neighbour masks.  One thread per sorted
// particle.  aux = (rho, P, sqrt|P/rho|, 1/rho) (SphUtils.cpp:223 evaluates sqrt|P/rho| per pair).

__device__ __forceinline__ float4 finish_density(const Consts& c, float sum) {
    float rho = sum * c.Kw;
    if (rho < c.rho0) rho = c.rho0;                 // SphUtils.cpp:141-142 (clamp UP)
    float P = __fmul_rn(c.eos_k, __fsub_rn(rho, c.rho0)); // :144
    return make_aux(rho, P);
}

// Row groups.  The density kernel gives each thread kDT consecutive sorted particles of ONE (y,z) row of cells,
// so that they always share their candidate runs.  Row rho holds the slots [cell_start[rho*nx],
// cell_start[(rho+1)*nx]) and is cut into ceil(len / kDT) groups; gstart = exclusive scan of the group counts
// (nrows + 1 entries), one block.  Thread t of the density kernel finds its row by binary search in gstart.
constexpr int kDT = 4;
constexpr int kRgThreads = 1024;

__global__ void __launch_bounds__(kRgThreads) k_row_groups(const GridDesc* __restrict__ gd, const int* __restrict__ cell_start,
                                                          int* __restrict__ gstart) {
    __shared__ int s_warp[kRgThreads / 32];
    __shared__ int s_carry;
    const int nx = gd->dims[0], nrows = gd->dims[1] * gd->dims[2];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (threadIdx.x == 0) s_carry = 0;
    __syncthreads();
    for (int base = 0; base < nrows; base += kRgThreads) {
        const int rho = base + threadIdx.x;
        int v = 0;
        if (rho < nrows) v = (cell_start[(rho + 1) * nx] - cell_start[rho * nx] + kDT - 1) / kDT;
        int inc = v;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int t = __shfl_up_sync(0xffffffffu, inc, o);
            if (lane >= o) inc += t;
        }
        if (lane == 31) s_warp[warp] = inc;
        __syncthreads();
        if (warp == 0) {
            int w = s_warp[lane];
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const int t = __shfl_up_sync(0xffffffffu, w, o);
                if (lane >= o) w += t;
            }
            s_warp[lane] = w;
        }
        __syncthreads();
        const int carry = s_carry;
        const int ex = carry + (warp ? s_warp[warp - 1] : 0) + inc - v;
        if (rho < nrows) gstart[rho] = ex;
        __syncthreads();
        if (threadIdx.x == kRgThreads - 1) s_carry = ex + v;
        __syncthreads();
    }
    if (threadIdx.x == 0) gstart[nrows] = s_carry;
}

// The row rho with gstart[rho] <= t < gstart[rho + 1] (t = density-group index).  One 32-ary search per WARP
// for its first thread (three dependent loads for 32 K rows instead of fifteen), then each lane walks on from
// there: a warp covers ~128 particles, i.e. almost always one or two rows.  All 32 lanes must call it.
__device__ __forceinline__ int find_row(const int* __restrict__ gstart, int nrows_grid, int t) {
    const int lane = threadIdx.x & 31;
    const int t0 = __shfl_sync(0xffffffffu, t, 0);
    int lo = 0, hi = nrows_grid; // invariant: gstart[lo] <= t0 < gstart[hi]
    while (hi - lo > 1) {
        const int step = (hi - lo + 31) >> 5;
        const int probe = min(lo + (lane + 1) * step, hi);
        const bool le = probe < hi ? (__ldg(gstart + probe) <= t0) : false;
        const int cnt = __popc(__ballot_sync(0xffffffffu, le)); // probes are ascending: the first cnt satisfy <=
        const int nlo = lo + cnt * step, nhi = min(lo + (cnt + 1) * step, hi);
        lo = cnt ? min(nlo, hi - 1) : lo;
        hi = nhi;
        if (lo >= hi) hi = lo + 1;
    }
    int lo_r = lo;
    // lanes further right: walk forward (bounded), then finish with a plain binary search if rows are sparse
    int steps = 0;
    while (lo_r + 1 < nrows_grid && __ldg(gstart + lo_r + 1) <= t && steps < 8) { ++lo_r; ++steps; }
    if (lo_r + 1 < nrows_grid && __ldg(gstart + lo_r + 1) <= t) {
        int a = lo_r, b = nrows_grid;
        while (b - a > 1) {
            const int mid = (a + b) >> 1;
            if (__ldg(gstart + mid) <= t) a = mid; else b = mid;
        }
        lo_r = a;
    }
    return lo_r;
}

// One candidate run of ONE target, scalar: the general path (crowded runs, targets whose row differs from
// their thread-mates', array tails).  Adds the run's W terms (self included) to `sum` and writes the two mask
// words of row r.
template <bool kAux, bool kMask>
__device__ __forceinline__ void density_row_single(const Consts& c, const float4* __restrict__ posm, unsigned int* __restrict__ mask,
                                                   int n, int i, const float4& pi, int r, int jb, int je, float& sum) {
    const int jm = min(je, jb + kMaskBits);
    unsigned int m0 = 0, m1 = 0; // bits enter at the top and slide down: after len steps bit k = candidate k
    int j = jb;
    const int jh = min(jm, jb + 32);
    for (; j < jh; j++) {
        const float4 pj = ldg4(posm + j);
        const float dx = pi.x - pj.x, dy = pi.y - pj.y, dz = pi.z - pj.z;
        const float d2 = fmaf(dx, dx, fmaf(dy, dy, dz * dz));
        if (kAux) sum = fmaf(pj.w, w_shape(c, d2), sum);
        if (kMask) m0 = (m0 >> 1) | ((d2 < c.R2) ? 0x80000000u : 0u);
    }
    if (kMask) m0 = (jh - jb) ? (m0 >> (32 - (jh - jb))) : 0u;
    for (; j < jm; j++) {
        const float4 pj = ldg4(posm + j);
        const float dx = pi.x - pj.x, dy = pi.y - pj.y, dz = pi.z - pj.z;
        const float d2 = fmaf(dx, dx, fmaf(dy, dy, dz * dz));
        if (kAux) sum = fmaf(pj.w, w_shape(c, d2), sum);
        if (kMask) m1 = (m1 >> 1) | ((d2 < c.R2) ? 0x80000000u : 0u);
    }
    if (kMask) m1 = (jm - jh) > 0 ? (m1 >> (32 - (jm - jh))) : 0u;
    if (kAux)
        for (; j < je; j++) { // beyond the mask: crowded run
            const float4 pj = ldg4(posm + j);
            const float dx = pi.x - pj.x, dy = pi.y - pj.y, dz = pi.z - pj.z;
            sum = fmaf(pj.w, w_shape(c, fmaf(dx, dx, fmaf(dy, dy, dz * dz))), sum);
        }
    if (kMask) {
        const int self = i - jb; // the self term is excluded by index (SphUtils.cpp:132,216)
        if (self >= 0 && i < je) {
            if (self < 32) m0 &= ~(1u << self);
            else if (self < kMaskBits) m1 &= ~(1u << (self - 32));
        }
        mask[(size_t)(2 * r) * n + i] = m0;      // word w of row r covers candidates jb + 32*w ...
        mask[(size_t)(2 * r + 1) * n + i] = m1;
    }
}

// Density kernel proper.  Each thread owns kDT = 4 CONSECUTIVE sorted particles ("targets").  Neighbouring
// slots are neighbouring particles, so the four share their candidate runs up to a cell or two in x: every
// candidate is loaded once and tested against all four -- a quarter of the L1 wavefronts per test -- and the
// four tests run as two packed-FP32 pairs (Blackwell FADD2 / FMUL2 / FFMA2: one issue slot, two lanes of
// math), which is what this issue-bound loop needs (profiles/r1b: issue 83 %, L1 data pipe 75 %).
//   u  = d2 - R^2            sign bit = "inside the support"  -> funnel-shifted straight into the mask word
//   t  = -u / h^2 = h^2 - d2/h^2 ;  W ~ max(t, 0)^3
// Union run of the four targets: [jb of the leftmost, je of the rightmost]; a candidate outside a target's own
// run is at least R away from it in x, so its bit is (at most a rounding-level ~0 term) clear and the
// per-target mask is the union mask shifted down by (own jb - union jb).
// 8 resident blocks of 128 threads per SM (64 registers): measured best on B200 -- the loop is latency-bound,
// so occupancy beats the few spilled words (1: 2.38 ms, 6: 2.41, 7: 2.21, 8: 2.16 ms per 16.7 M particles;
// a register-staged software prefetch of the candidates bought nothing on top).
#ifndef SPHB_DENSITY_MINBLOCKS
#define SPHB_DENSITY_MINBLOCKS 8
#endif
template <bool kAux, bool kMask>
__global__ void __launch_bounds__(128, SPHB_DENSITY_MINBLOCKS) k_density(Consts c, const GridDesc* __restrict__ gd, const int* __restrict__ cell_start,
                                                 const int* __restrict__ gstart,
                                                 const float4* __restrict__ posm, const float4* __restrict__ vel,
                                                 float4* __restrict__ aux, float4* __restrict__ vr,
                                                 unsigned int* __restrict__ mask, int n) {
    const GridDesc g = *gd;
    const int t_raw = blockIdx.x * blockDim.x + threadIdx.x;
    const int nrows_grid = g.dims[1] * g.dims[2];
    const int ngroups = __ldg(gstart + nrows_grid);
    if ((t_raw & ~31) >= ngroups) return;          // whole warp beyond the last group
    const bool live = t_raw < ngroups;
    const int t = live ? t_raw : ngroups - 1;       // spare lanes of the last warp shadow the last group (warp-wide votes below)
    const int lo_r = find_row(gstart, nrows_grid, t);
    if (!live) return;
    const int row_end = __ldg(cell_start + (lo_r + 1) * g.dims[0]);
    const int i0 = __ldg(cell_start + lo_r * g.dims[0]) + (t - __ldg(gstart + lo_r)) * kDT;
    const int nt = min(kDT, row_end - i0); // the tail group of a row is short: extra targets repeat the last one
    float4 pi[kDT];
    int cx[kDT], cy[kDT], cz[kDT];
#pragma unroll
    for (int k = 0; k < kDT; k++) {
        pi[k] = ldg4(posm + i0 + min(k, nt - 1));
        cell_key(g, pi[k].x, pi[k].y, pi[k].z, cx[k], cy[k], cz[k]);
    }
    const int S = g.stencil, Sx = g.stencil_x;
    const int nrows = (2 * S + 1) * (2 * S + 1);
    float sum[kDT] = {0.0f, 0.0f, 0.0f, 0.0f};
    {
        const int cxlo = min(min(cx[0], cx[1]), min(cx[2], cx[3])), cxhi = max(max(cx[0], cx[1]), max(cx[2], cx[3]));
        const float2 nx01 = make_float2(-pi[0].x, -pi[1].x), ny01 = make_float2(-pi[0].y, -pi[1].y), nz01 = make_float2(-pi[0].z, -pi[1].z);
        const float2 nx23 = make_float2(-pi[2].x, -pi[3].x), ny23 = make_float2(-pi[2].y, -pi[3].y), nz23 = make_float2(-pi[2].z, -pi[3].z);
        const float2 nR2 = make_float2(-c.R2, -c.R2), nih2 = make_float2(-c.inv_h2, -c.inv_h2);
        float2 acc01 = make_float2(0.0f, 0.0f), acc23 = make_float2(0.0f, 0.0f);
        const int W = 2 * S + 1;
        for (int r = 0; r < nrows; r++) {
            const int z = cz[0] + r / W - S, y = cy[0] + r % W - S;
            int jbu = 0, jeu = 0, jbk[kDT] = {0, 0, 0, 0};
            const bool inside = !(z < 0 || z >= g.dims[2] || y < 0 || y >= g.dims[1]);
            if (inside) {
                const int row = g.dims[0] * (y + g.dims[1] * z);
                jbu = __ldg(cell_start + row + max(cxlo - Sx, 0));
                jeu = __ldg(cell_start + row + min(cxhi + Sx, g.dims[0] - 1) + 1);
#pragma unroll
                for (int k = 0; k < kDT; k++) jbk[k] = (cx[k] == cxlo) ? jbu : __ldg(cell_start + row + max(cx[k] - Sx, 0));
            }
            if (jeu - jbu > kMaskBits) { // crowded union: per-target general path for this row
#pragma unroll
                for (int k = 0; k < kDT; k++) {
                    if (k >= nt) break;
                    int jb, je;
                    row_run(g, cell_start, cx[k], cy[k], cz[k], r, jb, je);
                    density_row_single<kAux, kMask>(c, posm, mask, n, i0 + k, pi[k], r, jb, je, sum[k]);
                }
                continue;
            }
            unsigned int mw[2][kDT];
#pragma unroll
            for (int w = 0; w < 2; w++) {
                const int ja = min(jeu, jbu + 32 * w), jz = min(jeu, jbu + 32 * (w + 1));
                unsigned int m[kDT] = {0u, 0u, 0u, 0u};
#pragma unroll 2
                for (int j = ja; j < jz; j++) {
                    const float4 pj = ldg4(posm + j);
                    const float2 xx = make_float2(pj.x, pj.x), yy = make_float2(pj.y, pj.y), zz = make_float2(pj.z, pj.z), mm = make_float2(pj.w, pj.w);
                    {
                        const float2 dx = __fadd2_rn(xx, nx01), dy = __fadd2_rn(yy, ny01), dz = __fadd2_rn(zz, nz01);
                        const float2 d2 = __ffma2_rn(dx, dx, __ffma2_rn(dy, dy, __fmul2_rn(dz, dz)));
                        const float2 u = __fadd2_rn(d2, nR2);
                        if (kMask) {
                            m[0] = __funnelshift_l(__float_as_uint(u.x), m[0], 1);
                            m[1] = __funnelshift_l(__float_as_uint(u.y), m[1], 1);
                        }
                        if (kAux) {
                            float2 t = __fmul2_rn(u, nih2);
                            t.x = fmaxf(t.x, 0.0f);
                            t.y = fmaxf(t.y, 0.0f);
                            acc01 = __ffma2_rn(mm, __fmul2_rn(__fmul2_rn(t, t), t), acc01);
                        }
                    }
                    {
                        const float2 dx = __fadd2_rn(xx, nx23), dy = __fadd2_rn(yy, ny23), dz = __fadd2_rn(zz, nz23);
                        const float2 d2 = __ffma2_rn(dx, dx, __ffma2_rn(dy, dy, __fmul2_rn(dz, dz)));
                        const float2 u = __fadd2_rn(d2, nR2);
                        if (kMask) {
                            m[2] = __funnelshift_l(__float_as_uint(u.x), m[2], 1);
                            m[3] = __funnelshift_l(__float_as_uint(u.y), m[3], 1);
                        }
                        if (kAux) {
                            float2 t = __fmul2_rn(u, nih2);
                            t.x = fmaxf(t.x, 0.0f);
                            t.y = fmaxf(t.y, 0.0f);
                            acc23 = __ffma2_rn(mm, __fmul2_rn(__fmul2_rn(t, t), t), acc23);
                        }
                    }
                }
                // the bits went in newest-lowest: candidate q of this word sits at bit (len-1-q); put it at bit q
                const int len = jz - ja;
#pragma unroll
                for (int k = 0; k < kDT; k++) mw[w][k] = len ? (__brev(m[k]) >> (32 - len)) : 0u;
            }
            if (kMask) {
#pragma unroll
                for (int k = 0; k < kDT; k++) {
                    unsigned long long m64 = ((unsigned long long)mw[1][k] << 32) | mw[0][k];
                    m64 >>= (jbk[k] - jbu); // union-relative -> relative to this target's own run
                    const int self = i0 + k - jbk[k];
                    if (r == (nrows >> 1) && self >= 0 && self < kMaskBits) m64 &= ~(1ull << self); // self excluded by index
                    if (k < nt) {
                        mask[(size_t)(2 * r) * n + i0 + k] = (unsigned int)m64;
                        mask[(size_t)(2 * r + 1) * n + i0 + k] = (unsigned int)(m64 >> 32);
                    }
                }
            }
        }
        sum[0] += acc01.x; sum[1] += acc01.y; sum[2] += acc23.x; sum[3] += acc23.y;
    }
    if (kAux) {
#pragma unroll
        for (int k = 0; k < kDT; k++) {
            if (k >= nt) break;
            // self term: the runs above included j == i exactly once (its own cell's row); take it out again
            const float s = sum[k] - pi[k].w * (c.h2 * c.h2 * c.h2);
            const float4 a = finish_density(c, s);
            aux[i0 + k] = a;
            // the force kernel gathers two records per neighbour: posm and (vx, vy, vz, rho); P, sqrt|P/rho| and
            // 1/rho are functions of rho and are recomputed there (two gathers instead of three: the kernel is
            // bound by L1 gather wavefronts, profiles/r1b)
            const float4 v = ldg4(vel + i0 + k);
            vr[i0 + k] = make_float4(v.x, v.y, v.z, a.x);
        }
    }
}

// ---------------------------------------------------------------------------------------------------
// K5: pressure + artificial-viscosity pair force (SphUtils.cpp:181-239), gravity (:295-304), and --
// fused -- symplectic Euler (:321-335) and wall collision (:343-372).

// One pair (i <- j), in range (d2 < R2), accumulates the direction-weighted scalar; the factor
// Kg = -45/(pi h^6) is applied once after the sum.
__device__ __forceinline__ void pair_force(const Consts& c, float dx, float dy, float dz, float d2, const float4& vi, const float4& ai,
                                           float mj, const float4& vj, const float4& aj, float& ax, float& ay, float& az) {
    const float vdotr = fmaf(vi.x - vj.x, dx, fmaf(vi.y - vj.y, dy, (vi.z - vj.z) * dz)); // :220
    const float rinv = fast_rsqrt(d2);
    const float r = d2 * rinv;
    const float s = fmaf(-r, c.inv_h, c.h); // h - q, q = r/h              (:70-74)
    float visc = 0.0f;
    if (vdotr < 0.0f) {                                                                  // :222-227
        const float cbar = 0.5f * (ai.z + aj.z);
        const float mu = c.h * vdotr * fast_rcp(d2 + c.eta_h2);
        visc = 2.0f * mu * fmaf(c.beta, mu, -c.alpha * cbar) * fast_rcp(ai.x + aj.x);
    }
    // :228  -(m_j) * (pi + (P_i + P_j) / (2 rho_j rho_i)); ai.w, aj.w = 1/rho
    const float t = -mj * fmaf(0.5f * (ai.y + aj.y), ai.w * aj.w, visc);
    const float wgt = t * s * s * rinv; // times (x_i - x_j) = temp * |gradW| * normalize(i - j) / Kg
    ax = fmaf(wgt, dx, ax);
    ay = fmaf(wgt, dy, ay);
    az = fmaf(wgt, dz, az);
}

// The neighbour's (P, sqrt|P/rho|, 1/rho) from its rho alone, as finish_density computes them (P with the
// same rounding; the other two through MUFU approximations, <= 2 ulp).
__device__ __forceinline__ float4 aux_from_rho(const Consts& c, float rho) {
    return make_aux(rho, __fmul_rn(c.eos_k, __fsub_rn(rho, c.rho0)));
}

// SphUtils.cpp:326-328 with the reference's rounding (mul then add, no FMA): bit-exact given equal accel
__device__ __forceinline__ void integrate(const Consts& c, float4& p, float4& v, float ax, float ay, float az) {
    v.x = __fadd_rn(v.x, __fmul_rn(c.dt, ax));
    v.y = __fadd_rn(v.y, __fmul_rn(c.dt, ay));
    v.z = __fadd_rn(v.z, __fmul_rn(c.dt, az));
    p.x = __fadd_rn(p.x, __fmul_rn(c.dt, v.x));
    p.y = __fadd_rn(p.y, __fmul_rn(c.dt, v.y));
    p.z = __fadd_rn(p.z, __fmul_rn(c.dt, v.z));
}

// SphUtils.cpp:346-372, walls in vector order, each seeing the previous one's correction; the hit test
// runs in double like the reference's double literals.  Bit-exact given equal inputs.
__device__ __forceinline__ void collide_walls(const Consts& c, float4& p, float4& v) {
    for (int w = 0; w < c.nwalls; w++) {
        const float nx = c.wn[w][0], ny = c.wn[w][1], nz = c.wn[w][2];
        float d = __fadd_rn(__fadd_rn(__fmul_rn(nx, __fsub_rn(p.x, c.wc[w][0])), __fmul_rn(ny, __fsub_rn(p.y, c.wc[w][1]))),
                            __fmul_rn(nz, __fsub_rn(p.z, c.wc[w][2])));
        if ((double)d + c.wall_off <= c.wall_lim || (double)d - c.wall_off <= c.wall_lim) {
            const float ndv = __fadd_rn(__fadd_rn(__fmul_rn(nx, v.x), __fmul_rn(ny, v.y)), __fmul_rn(nz, v.z));
            const float k = __fmul_rn(c.one_plus_e, ndv);
            if (d < 0.0f) d = -d;
            p.x = __fadd_rn(p.x, __fmul_rn(d, nx));
            p.y = __fadd_rn(p.y, __fmul_rn(d, ny));
            p.z = __fadd_rn(p.z, __fmul_rn(d, nz));
            v.x = __fadd_rn(v.x, -__fmul_rn(c.wu[w][0], k));
            v.y = __fadd_rn(v.y, -__fmul_rn(c.wu[w][1], k));
            v.z = __fadd_rn(v.z, -__fmul_rn(c.wu[w][2], k));
        }
    }
}

// kFused: also integrate + collide and write the new state to the other buffer (ping-pong: every
//         thread reads only old buffers, so the fusion is race-free).
// kBBox:  also fold the new positions into the bounding box for the next step's grid.
// kAccel: store the acceleration (parity, viewers).
// kRows:  (2S+1)^2, 9 for cell >= R, 25 for cell >= R/2.
// Each thread walks the set bits of its particle's neighbour masks (written by k_density) as ONE stream
// across the rows, so the lanes of a warp stay together on the pair body however the neighbours are
// spread over the rows.  Masks and run starts are staged per thread in shared memory (column tid: no
// conflicts, no barrier needed).
// kDerive: gather (vx,vy,vz,rho) records written by k_density and derive P etc. from rho (two gathers per
//          neighbour); otherwise gather vel and aux separately (three; needed when rho/P were uploaded and P
//          need not equal k(rho - rho0)).
template <bool kFused, bool kBBox, bool kAccel, int kRows, bool kDerive>
__global__ void __launch_bounds__(128) k_force(Consts c, const GridDesc* __restrict__ gd, const int* __restrict__ cell_start,
                                               const float4* __restrict__ posm, const float4* __restrict__ vel, const float4* __restrict__ aux,
                                               const float4* __restrict__ vr, const unsigned int* __restrict__ mask,
                                               float4* __restrict__ accel, float4* __restrict__ posm_out, float4* __restrict__ vel_out,
                                               BBoxAcc* bbox, int n) {
    __shared__ unsigned int s_word[2 * kRows][128];  // this thread's NON-EMPTY mask words, compacted ...
    __shared__ unsigned char s_idx[2 * kRows][128];  // ... which word each is (2*row + half) ...
    __shared__ int s_jb[kRows][128];                 // ... and where each row's run starts
    const int tid = threadIdx.x;
    const int i = blockIdx.x * blockDim.x + tid;
    float lo[3] = {INFINITY, INFINITY, INFINITY}, hi[3] = {-INFINITY, -INFINITY, -INFINITY};
    if (i < n) {
        const GridDesc g = *gd;
        const float4 pi = ldg4(posm + i);
        const float4 vi = ldg4(vel + i);
        const float4 ai = ldg4(aux + i);
        int cx, cy, cz;
        cell_key(g, pi.x, pi.y, pi.z, cx, cy, cz);
        bool crowded = false;
        int nw = 0;
#pragma unroll
        for (int r = 0; r < kRows; r++) {
            int jb, je;
            row_run(g, cell_start, cx, cy, cz, r, jb, je);
            const unsigned int m0 = mask[(size_t)(2 * r) * n + i], m1 = mask[(size_t)(2 * r + 1) * n + i];
            s_jb[r][tid] = jb;
            if (m0) { s_word[nw][tid] = m0; s_idx[nw][tid] = (unsigned char)(2 * r); nw++; }
            if (m1) { s_word[nw][tid] = m1; s_idx[nw][tid] = (unsigned char)(2 * r + 1); nw++; }
            crowded |= (je - jb) > kMaskBits;
        }
        float ax = 0.0f, ay = 0.0f, az = 0.0f;
        // One neighbour per lane per iteration: a lane whose word ran dry steps to its next (non-empty) word first,
        // so the lanes of a warp stay in lockstep on the pair body until their lists end.
        int k = -1, jb = 0;
        unsigned int m = 0;
        for (;;) {
            if (m == 0) {
                if (++k >= nw) break;
                m = s_word[k][tid];
                const int w = s_idx[k][tid];
                jb = s_jb[w >> 1][tid] + ((w & 1) << 5);
            }
            const int j = jb + __ffs((int)m) - 1;
            m &= m - 1;
            const float4 pj = ldg4(posm + j);
            float4 vj, aj;
            if (kDerive) {
                vj = ldg4(vr + j);
                aj = aux_from_rho(c, vj.w);
            } else {
                vj = ldg4(vel + j);
                aj = ldg4(aux + j);
            }
            const float dx = pi.x - pj.x, dy = pi.y - pj.y, dz = pi.z - pj.z;
            const float d2 = fmaf(dx, dx, fmaf(dy, dy, dz * dz));
            pair_force(c, dx, dy, dz, d2, vi, ai, pj.w, vj, aj, ax, ay, az);
        }
        if (crowded) { // candidates past the 64th of a run: test them here (rare)
            for (int rr = 0; rr < kRows; rr++) {
                int jb2, je2;
                row_run(g, cell_start, cx, cy, cz, rr, jb2, je2);
                for (int j = jb2 + kMaskBits; j < je2; j++) {
                    const float4 pj = ldg4(posm + j);
                    const float dx = pi.x - pj.x, dy = pi.y - pj.y, dz = pi.z - pj.z;
                    const float d2 = fmaf(dx, dx, fmaf(dy, dy, dz * dz));
                    // q < h  <=>  d2 < h^4; NaN-safe form keeps the reference's NaN for coincident points
                    if (!(d2 >= c.R2) && j != i) pair_force(c, dx, dy, dz, d2, vi, ai, pj.w, ldg4(vel + j), ldg4(aux + j), ax, ay, az);
                }
            }
        }
        ax = fmaf(c.Kg, ax, c.gx);
        ay = fmaf(c.Kg, ay, c.gy);
        az = fmaf(c.Kg, az, c.gz);
        if (kAccel) accel[i] = make_float4(ax, ay, az, 0.0f);
        if (kFused) {
            float4 p = pi, v = vi;
            integrate(c, p, v, ax, ay, az);
            collide_walls(c, p, v);
            posm_out[i] = p;
            vel_out[i] = v;
            if (kBBox) {
                lo[0] = hi[0] = p.x; lo[1] = hi[1] = p.y; lo[2] = hi[2] = p.z;
#pragma unroll
                for (int a = 0; a < 3; a++)
                    if (!(lo[a] == lo[a])) { lo[a] = INFINITY; hi[a] = -INFINITY; } // NaN: ignored like the reference's compares
            }
        }
    }
    if (kFused && kBBox) bbox_block_commit(lo, hi, bbox);
}

// Optional: the reference's dead viscosity_forces (SphUtils.cpp:244-290; its call at :175 is commented out), on
// the full neighbourhood:  accel_i += coef * m_j / (rho_i rho_j) * lapW_ij * (v_i - v_j),  lapW = 45/(pi h^6) (h - q)
// (:84-95).  A separate pass over the neighbour masks that adds into accel; row by row, off the hot path.
__global__ void __launch_bounds__(128) k_lap_viscosity(Consts c, const GridDesc* __restrict__ gd, const int* __restrict__ cell_start,
                                                       const float4* __restrict__ posm, const float4* __restrict__ vel, const float4* __restrict__ aux,
                                                       const unsigned int* __restrict__ mask, float4* __restrict__ accel, int n) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const GridDesc g = *gd;
    const float4 pi = ldg4(posm + i), vi = ldg4(vel + i), ai = ldg4(aux + i);
    int cx, cy, cz;
    cell_key(g, pi.x, pi.y, pi.z, cx, cy, cz);
    const int nrows = (2 * g.stencil + 1) * (2 * g.stencil + 1);
    float ax = 0.0f, ay = 0.0f, az = 0.0f;
    auto pair = [&](int j) {
        const float4 pj = ldg4(posm + j), vj = ldg4(vel + j), aj = ldg4(aux + j);
        const float dx = pi.x - pj.x, dy = pi.y - pj.y, dz = pi.z - pj.z;
        const float d2 = fmaf(dx, dx, fmaf(dy, dy, dz * dz));
        const float s = fmaf(-sqrtf(d2), c.inv_h, c.h);      // h - q
        const float t = c.lap_coef * pj.w * ai.w * aj.w * s; // times Kl after the sum
        ax = fmaf(t, vi.x - vj.x, ax);
        ay = fmaf(t, vi.y - vj.y, ay);
        az = fmaf(t, vi.z - vj.z, az);
    };
    for (int r = 0; r < nrows; r++) {
        int jb, je;
        row_run(g, cell_start, cx, cy, cz, r, jb, je);
        for (int w = 0; w < 2; w++) {
            unsigned int m = mask[(size_t)(2 * r + w) * n + i];
            while (m) {
                pair(jb + 32 * w + __ffs((int)m) - 1);
                m &= m - 1;
            }
        }
        for (int j = jb + kMaskBits; j < je; j++) { // crowded run: beyond the mask
            const float4 pj = ldg4(posm + j);
            const float dx = pi.x - pj.x, dy = pi.y - pj.y, dz = pi.z - pj.z;
            if (fmaf(dx, dx, fmaf(dy, dy, dz * dz)) < c.R2 && j != i) pair(j);
        }
    }
    float4 a = accel[i];
    a.x = fmaf(c.Kl, ax, a.x);
    a.y = fmaf(c.Kl, ay, a.y);
    a.z = fmaf(c.Kl, az, a.z);
    accel[i] = a;
}

// Phase kernels for the method-by-method API (update_posvel, collision), in place.
__global__ void __launch_bounds__(256) k_posvel(Consts c, float4* __restrict__ posm, float4* __restrict__ vel, const float4* __restrict__ accel, int n) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    float4 p = posm[i], v = vel[i];
    const float4 a = accel[i];
    integrate(c, p, v, a.x, a.y, a.z);
    posm[i] = p;
    vel[i] = v;
}

__global__ void __launch_bounds__(256) k_collision(Consts c, float4* __restrict__ posm, float4* __restrict__ vel, int n) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    float4 p = posm[i], v = vel[i];
    collide_walls(c, p, v);
    posm[i] = p;
    vel[i] = v;
}

// kernel_function / grad_kernel / lap_kernel for explicit point pairs, same device math as above
__global__ void __launch_bounds__(256) k_eval_kernels(Consts c, const float* __restrict__ xi, const float* __restrict__ xj, int count,
                                                      float* __restrict__ w, float* __restrict__ grad, float* __restrict__ lap) {
    int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= count) return;
    const float dx = xi[3 * t] - xj[3 * t], dy = xi[3 * t + 1] - xj[3 * t + 1], dz = xi[3 * t + 2] - xj[3 * t + 2];
    const float d2 = fmaf(dx, dx, fmaf(dy, dy, dz * dz));
    if (w) w[t] = c.Kw * w_shape(c, d2);
    const float rinv = rsqrtf(d2);
    const float r = d2 * rinv;
    const float s = fmaf(-r, c.inv_h, c.h);
    const bool in = !(d2 >= c.R2);
    if (grad) {
        const float m = in ? c.Kg * s * s : 0.0f;
        grad[3 * t] = dx * rinv * m;      // normalize(i - j) * result; NaN at r = 0 like the reference
        grad[3 * t + 1] = dy * rinv * m;
        grad[3 * t + 2] = dz * rinv * m;
    }
    if (lap) lap[t] = (in && d2 > 0.0f) ? c.Kl * s : (in ? c.Kl * c.h : 0.0f);
}

// ---------------------------------------------------------------------------------------------------
// AoS (the reference's 52-byte Particle) <-> SoA.  Records are staged through shared memory so that
// global traffic is fully coalesced 4-byte words regardless of the odd record size.

constexpr int kAosThreads = 256;

// word offsets inside a record (SPH/Particle.hpp:42-50)
constexpr int kOffPos = 0, kOffVel = 3, kOffAcc = 6, kOffMass = 9, kOffRho = 10, kOffP = 11;

__global__ void __launch_bounds__(kAosThreads) k_aos_to_soa(const float* __restrict__ aos, int stride_w, int n, float4* __restrict__ posm,
                                                            float4* __restrict__ vel, float4* __restrict__ aux, float4* __restrict__ accel,
                                                            int* __restrict__ oid) {
    extern __shared__ float s_rec[];
    const int base = blockIdx.x * kAosThreads;
    const int cnt = min(kAosThreads, n - base);
    const long long wbase = (long long)base * stride_w;
    for (int w = threadIdx.x; w < cnt * stride_w; w += kAosThreads) s_rec[w] = aos[wbase + w];
    __syncthreads();
    if (threadIdx.x >= cnt) return;
    const float* r = s_rec + threadIdx.x * stride_w;
    const int i = base + threadIdx.x;
    posm[i] = make_float4(r[kOffPos], r[kOffPos + 1], r[kOffPos + 2], r[kOffMass]);
    vel[i] = make_float4(r[kOffVel], r[kOffVel + 1], r[kOffVel + 2], 1.0f); // w = 1: owned by this engine (multi-GPU: 0 = ghost)
    const float rho = r[kOffRho], P = r[kOffP];
    aux[i] = make_aux(rho, P);
    accel[i] = make_float4(r[kOffAcc], r[kOffAcc + 1], r[kOffAcc + 2], 0.0f);
    oid[i] = i;
}

// Sorted SoA -> AoS in ORIGINAL order.  inv[o] = sorted slot of original particle o.
__global__ void __launch_bounds__(256) k_invert_perm(const int* __restrict__ oid, int* __restrict__ inv, int n) {
    int p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p < n) inv[oid[p]] = p;
}

__global__ void __launch_bounds__(kAosThreads) k_soa_to_aos(float* __restrict__ aos, int stride_w, int n, const int* __restrict__ inv,
                                                            const float4* __restrict__ posm, const float4* __restrict__ vel,
                                                            const float4* __restrict__ aux, const float4* __restrict__ accel, unsigned mask) {
    extern __shared__ float s_rec[];
    const int base = blockIdx.x * kAosThreads;
    const int cnt = min(kAosThreads, n - base);
    const long long wbase = (long long)base * stride_w;
    // read-modify-write so that fields outside the mask (and thermal, padding) keep their bytes
    for (int w = threadIdx.x; w < cnt * stride_w; w += kAosThreads) s_rec[w] = aos[wbase + w];
    __syncthreads();
    if (threadIdx.x < cnt) {
        float* r = s_rec + threadIdx.x * stride_w;
        const int p = inv[base + threadIdx.x];
        if (mask & 1u) { const float4 q = posm[p]; r[kOffPos] = q.x; r[kOffPos + 1] = q.y; r[kOffPos + 2] = q.z; }
        if (mask & 2u) { const float4 q = vel[p]; r[kOffVel] = q.x; r[kOffVel + 1] = q.y; r[kOffVel + 2] = q.z; }
        if (mask & 4u) { const float4 q = accel[p]; r[kOffAcc] = q.x; r[kOffAcc + 1] = q.y; r[kOffAcc + 2] = q.z; }
        if (mask & 8u) { const float4 q = aux[p]; r[kOffRho] = q.x; r[kOffP] = q.y; }
    }
    __syncthreads();
    for (int w = threadIdx.x; w < cnt * stride_w; w += kAosThreads) aos[wbase + w] = s_rec[w];
}

// positions only, packed, ORIGINAL order (the viewer's vertex buffer, main.cpp:97-110)
__global__ void __launch_bounds__(256) k_positions_packed(const float4* __restrict__ posm, const int* __restrict__ inv, float* __restrict__ xyz, int n) {
    const int o = blockIdx.x * blockDim.x + threadIdx.x;
    if (o >= n) return;
    const float4 q = posm[inv[o]];
    xyz[(size_t)3 * o] = q.x; xyz[(size_t)3 * o + 1] = q.y; xyz[(size_t)3 * o + 2] = q.z;
}

// grid diagnostics for the bit-exact grid test: cell coordinates in ORIGINAL order
__global__ void __launch_bounds__(256) k_cells_xyz(const GridDesc* __restrict__ gd, const float4* __restrict__ posm, const int* __restrict__ oid,
                                                   int* __restrict__ xyz, int n) {
    int p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= n) return;
    const GridDesc g = *gd;
    const float4 q = posm[p];
    int cx, cy, cz;
    cell_key(g, q.x, q.y, q.z, cx, cy, cz);
    const int o = oid[p];
    xyz[(size_t)3 * o] = cx; xyz[(size_t)3 * o + 1] = cy; xyz[(size_t)3 * o + 2] = cz;
}

// ---------------------------------------------------------------------------------------------------
// Rigid spheres (SphUtils.cpp:163-166, 300-303, 331-334, 375-385).  The reference resolves particle-sphere
// contacts INSIDE its serial particle loop: every contact changes the sphere's velocity, which the next
// colliding particle then sees.  Which particles touch a sphere does not depend on that order (positions do
// not change in collision()), so: detect in parallel, sort the few contacts by original particle index, and
// replay the velocity exchanges in that order on one thread with the reference's float operations.

constexpr int kMaxSpheres = 16;

struct SphereDev { // same 44-byte layout as the reference's Sphere (SPH/Sphere.hpp:29-33)
    float pos[3], vel[3], acc[3], radius, mass;
};

__global__ void k_sphere_forces(Consts c, SphereDev* __restrict__ sp, int ns) {
    const int j = threadIdx.x;
    if (j >= ns) return;
    sp[j].acc[0] = __fadd_rn(0.0f, c.gx); // accel = 0 (:165) then += gravity (:302)
    sp[j].acc[1] = __fadd_rn(0.0f, c.gy);
    sp[j].acc[2] = __fadd_rn(0.0f, c.gz);
}

__global__ void k_sphere_posvel(Consts c, SphereDev* __restrict__ sp, int ns) {
    const int j = threadIdx.x;
    if (j >= ns) return;
#pragma unroll
    for (int a = 0; a < 3; a++) {
        sp[j].vel[a] = __fadd_rn(sp[j].vel[a], __fmul_rn(c.dt, sp[j].acc[a])); // :332-333
        sp[j].pos[a] = __fadd_rn(sp[j].pos[a], __fmul_rn(c.dt, sp[j].vel[a]));
    }
}

// glm::length(p - s) <= 0.1 + radius, the comparison in double (SphUtils.cpp:376-377)
__device__ __forceinline__ bool sphere_touch(const float4& p, const SphereDev& s) {
    const float dx = __fsub_rn(p.x, s.pos[0]), dy = __fsub_rn(p.y, s.pos[1]), dz = __fsub_rn(p.z, s.pos[2]);
    const float d = __fsqrt_rn(__fadd_rn(__fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy)), __fmul_rn(dz, dz)));
    return (double)d <= 0.1 + (double)s.radius;
}

__global__ void __launch_bounds__(256) k_sphere_detect(const float4* __restrict__ posm, const int* __restrict__ oid, int n,
                                                       const SphereDev* __restrict__ sp, int ns, int2* __restrict__ list, int* __restrict__ count) {
    const int p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= n) return;
    const float4 q = posm[p];
    bool hit = false;
    for (int j = 0; j < ns; j++) hit |= sphere_touch(q, sp[j]);
    if (hit) list[atomicAdd(count, 1)] = make_int2(oid[p], p);
}

// stable order: rank of a contact = number of contacts with a smaller original particle index
__global__ void __launch_bounds__(256) k_sphere_rank(const int2* __restrict__ list, const int* __restrict__ count, int2* __restrict__ sorted) {
    const int K = *count;
    for (int k = blockIdx.x * blockDim.x + threadIdx.x; k < K; k += gridDim.x * blockDim.x) {
        const int2 me = list[k];
        int r = 0;
        for (int l = 0; l < K; l++) r += list[l].x < me.x;
        sorted[r] = me;
    }
}

// the serial replay (SphUtils.cpp:375-385), particles in original index order, spheres in vector order
__global__ void k_sphere_serial(const float4* __restrict__ posm, float4* __restrict__ vel, const int2* __restrict__ sorted,
                                const int* __restrict__ count, SphereDev* __restrict__ sp, int ns) {
    if (threadIdx.x || blockIdx.x) return;
    const int K = *count;
    for (int k = 0; k < K; k++) {
        const int p = sorted[k].y;
        const float4 q = posm[p];
        float4 v = vel[p];
        const float mp = q.w;
        for (int j = 0; j < ns; j++) {
            if (!sphere_touch(q, sp[j])) continue;
            const float ms = sp[j].mass;
            const float u1[3] = {v.x, v.y, v.z};
            const float u2[3] = {sp[j].vel[0], sp[j].vel[1], sp[j].vel[2]};
            const float dm = __fsub_rn(mp, ms), dn = __fsub_rn(ms, mp), sum = __fadd_rn(mp, ms);
            const float two_ms = __fmul_rn(2.0f, ms), two_mp = __fmul_rn(2.0f, mp);
            float nv[3];
#pragma unroll
            for (int a = 0; a < 3; a++) {
                nv[a] = __fdiv_rn(__fadd_rn(__fmul_rn(u1[a], dm), __fmul_rn(two_ms, u2[a])), sum);           // :380
                sp[j].vel[a] = __fdiv_rn(__fadd_rn(__fmul_rn(u2[a], dn), __fmul_rn(two_mp, u1[a])), sum);     // :382
            }
            v.x = nv[0]; v.y = nv[1]; v.z = nv[2];
        }
        vel[p] = v;
    }
}

// ---------------------------------------------------------------------------------------------------
// Headless point-sprite renderer: the arithmetic of SPH/Shaders/sph.vertexshader + sph.fragmentshader and of GL's
// point rasterisation, into a float image.  Two passes because the viewer has no depth buffer: the LAST particle
// (in original order) that covers a pixel wins.  Pass 1: one warp per particle walks its sprite square and
// atomicMax-es (particle index + 1) into the pixels it really covers; pass 2: one thread per pixel shades the winner.

struct Camera {
    float m[16]; // column-major
    int width, height;
    float point_scale, sprite_radius, clear_value;
};

__device__ __forceinline__ float4 mvp_mul(const Camera& c, float x, float y, float z) {
    // m[0]*x + m[1]*y + m[2]*z + m[3]*1, columns
    return make_float4(fmaf(c.m[0], x, fmaf(c.m[4], y, fmaf(c.m[8], z, c.m[12]))), fmaf(c.m[1], x, fmaf(c.m[5], y, fmaf(c.m[9], z, c.m[13]))),
                       fmaf(c.m[2], x, fmaf(c.m[6], y, fmaf(c.m[10], z, c.m[14]))), fmaf(c.m[3], x, fmaf(c.m[7], y, fmaf(c.m[11], z, c.m[15]))));
}

// sprite of one particle: window centre, edge length; false if culled
__device__ __forceinline__ bool sprite_of(const Camera& c, const float4& p, float& xw, float& yw, float& edge) {
    const float4 clip = mvp_mul(c, p.x, p.y, p.z);
    if (!(clip.w > 0.0f) || fabsf(clip.x) > clip.w || fabsf(clip.y) > clip.w || fabsf(clip.z) > clip.w) return false; // point clipping
    xw = (clip.x / clip.w * 0.5f + 0.5f) * (float)c.width;
    yw = (clip.y / clip.w * 0.5f + 0.5f) * (float)c.height;
    edge = fminf(fmaxf(c.point_scale / clip.w, 1.0f), 2048.0f);
    return true;
}

// fragment shader for pixel (px,py) of that sprite: false = discarded; depth = clamped output value
__device__ __forceinline__ bool sprite_fragment(const Camera& c, const float4& p, float xw, float yw, float edge, int px, int py, float& depth) {
    const float s = 0.5f + ((float)px + 0.5f - xw) / edge, t = 0.5f + ((float)py + 0.5f - yw) / edge;
    if (s < 0.0f || s >= 1.0f || t < 0.0f || t >= 1.0f) return false; // pixel centre outside the sprite square
    const float nx = s * 2.0f - 1.0f, ny = t * 2.0f - 1.0f;
    const float mag = nx * nx + ny * ny;
    if (mag > 1.0f) return false;                                      // sph.fragmentshader:14
    const float nz = sqrtf(1.0f - mag);
    const float4 cp = mvp_mul(c, fmaf(nx, c.sprite_radius, p.x), fmaf(ny, c.sprite_radius, p.y), fmaf(nz, c.sprite_radius, p.z));
    depth = fminf(fmaxf(cp.z / cp.w, 0.0f), 1.0f);                     // colour = vec4(depth), 8-bit target clamps
    return true;
}

__global__ void __launch_bounds__(256) k_splat_winner(Camera c, const float4* __restrict__ posm, const int* __restrict__ inv, int n,
                                                      unsigned int* __restrict__ winner) {
    const int o = (blockIdx.x * blockDim.x + threadIdx.x) >> 5; // one warp per particle, ORIGINAL index
    const int lane = threadIdx.x & 31;
    if (o >= n) return;
    const float4 p = posm[inv[o]];
    float xw, yw, edge;
    if (!sprite_of(c, p, xw, yw, edge)) return;
    const int x0 = max(0, (int)floorf(xw - 0.5f * edge)), x1 = min(c.width - 1, (int)ceilf(xw + 0.5f * edge));
    const int y0 = max(0, (int)floorf(yw - 0.5f * edge)), y1 = min(c.height - 1, (int)ceilf(yw + 0.5f * edge));
    const int wx = x1 - x0 + 1, wy = y1 - y0 + 1;
    if (wx <= 0 || wy <= 0) return;
    for (int k = lane; k < wx * wy; k += 32) {
        const int px = x0 + k % wx, py = y0 + k / wx;
        float d;
        if (sprite_fragment(c, p, xw, yw, edge, px, py, d)) atomicMax(winner + (size_t)py * c.width + px, (unsigned int)o + 1u);
    }
}

__global__ void __launch_bounds__(256) k_splat_shade(Camera c, const float4* __restrict__ posm, const int* __restrict__ inv,
                                                     const unsigned int* __restrict__ winner, float* __restrict__ image) {
    const int pix = blockIdx.x * blockDim.x + threadIdx.x;
    if (pix >= c.width * c.height) return;
    const unsigned int w = winner[pix];
    float v = c.clear_value;
    if (w) {
        const float4 p = posm[inv[w - 1]];
        float xw, yw, edge, d;
        if (sprite_of(c, p, xw, yw, edge) && sprite_fragment(c, p, xw, yw, edge, pix % c.width, pix / c.width, d)) v = d;
    }
    image[pix] = v;
}

// NaN / inf census (debug aid: the reference spreads NaN silently once two particles coincide)
__global__ void __launch_bounds__(256) k_count_nonfinite(const float4* __restrict__ posm, const float4* __restrict__ vel, int n,
                                                         unsigned long long* __restrict__ out) {
    int bad = 0;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const float4 p = posm[i], v = vel[i];
        const float s = p.x + p.y + p.z + v.x + v.y + v.z; // NaN or inf anywhere makes s non-finite (inf - inf gives NaN)
        bad += !(fabsf(s) <= 3.0e38f) || !(fabsf(p.x) <= 3.0e38f) || !(fabsf(v.x) <= 3.0e38f);
    }
    const int c = __syncthreads_count(0) + 0; // keep the block converged
    (void)c;
    bad = __reduce_add_sync(0xffffffffu, bad);
    if ((threadIdx.x & 31) == 0 && bad) atomicAdd(out, (unsigned long long)bad);
}

// device-to-device upload helper: ids default to 0..n-1
__global__ void __launch_bounds__(256) k_fill_iota(int* __restrict__ ids, int n) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) ids[i] = i;
}

// ---------------------------------------------------------------------------------------------------
// Multi-GPU x-slab decomposition (nothing like it in the reference, which is one thread).
// After a step every rank classifies the particles it OWNED (vel.w != 0) by their new x:
//   keep       x in [lo - halo, hi + halo): stays in this rank's arrays, owned if x in [lo, hi) else ghost
//   to left    x <  lo + halo  (owned boundary layer of the left neighbour's ghost zone, and migrants)
//   to right   x >= hi - halo
// Ghosts of the previous step (vel.w == 0) are dropped: they are re-sent every step.  All three outputs are
// compacted in input order with a two-level scan (per-block counts -> scan -> write), so the result is
// deterministic.  A message is a header record {count, capacity, 0...} followed by 32-byte particle
// records {posm float4, vel float4}; the receiver appends them and sets the ownership flag itself.

struct Slab {
    float lo, hi, halo;
    int has_left, has_right;
    float shift_left, shift_right; // added to x of a record sent that way (periodic seam), else 0
    int per_y, per_z;              // periodic in y / z over [ylo, ylo + ylen) / [zlo, zlo + zlen)
    float ylo, ylen, zlo, zlen;
};

constexpr int kMgThreads = 256; // == kScanThreads: block_exclusive_scan is reused

// Periodic y / z are handled HERE and nowhere else: an owned particle is wrapped back into the box, and if it lies
// within `halo` of a periodic face it is emitted once more per face (and once for the edge) shifted by the period,
// as a ghost image.  The density / force kernels never see a periodic boundary -- only more ghosts.
struct MgItem {
    float4 p, v;      // wrapped position+mass, velocity (+flag)
    int nimg;         // 1, 2 or 4 records: (0,0), (dy,0), (0,dz), (dy,dz)
    float dy, dz;     // image shifts (0 if not near that face)
    bool keep, left, right;
};

__device__ __forceinline__ float mg_wrap(float x, float lo, float len) {
    x = x - len * floorf((x - lo) / len);
    return x >= lo + len ? lo : x; // rounding can land exactly on the upper face
}

__device__ __forceinline__ MgItem mg_classify(const Slab& s, float4 p, const float4& v) {
    MgItem it;
    const bool owned = v.w != 0.0f;
    it.dy = it.dz = 0.0f;
    if (s.per_y) {
        p.y = mg_wrap(p.y, s.ylo, s.ylen);
        it.dy = p.y < s.ylo + s.halo ? s.ylen : (p.y >= s.ylo + s.ylen - s.halo ? -s.ylen : 0.0f);
    }
    if (s.per_z) {
        p.z = mg_wrap(p.z, s.zlo, s.zlen);
        it.dz = p.z < s.zlo + s.halo ? s.zlen : (p.z >= s.zlo + s.zlen - s.halo ? -s.zlen : 0.0f);
    }
    it.p = p;
    it.v = v;
    it.nimg = (it.dy != 0.0f ? 2 : 1) * (it.dz != 0.0f ? 2 : 1);
    const float x = p.x;
    it.keep = owned && x >= s.lo - s.halo && x < s.hi + s.halo;
    it.left = owned && s.has_left && x < s.lo + s.halo;
    it.right = owned && s.has_right && x >= s.hi - s.halo;
    return it;
}

__device__ __forceinline__ float4 mg_image(const MgItem& it, int img, float xshift) {
    // image order: base, y-image, z-image, edge image (only those that exist)
    const bool sy = it.dy != 0.0f && (img == 1 || img == 3);
    const bool sz = it.dz != 0.0f && ((it.dy != 0.0f) ? (img >= 2) : (img == 1));
    return make_float4(it.p.x + xshift, sy ? it.p.y + it.dy : it.p.y, sz ? it.p.z + it.dz : it.p.z, it.p.w);
}

__device__ __forceinline__ bool mg_inside(const Slab& s, const float4& p) {
    return p.x >= s.lo && p.x < s.hi && (!s.per_y || (p.y >= s.ylo && p.y < s.ylo + s.ylen)) &&
           (!s.per_z || (p.z >= s.zlo && p.z < s.zlo + s.zlen));
}

// pass 1: per-block record counts of (keep, left, right)
__global__ void __launch_bounds__(kMgThreads) k_mg_count(Slab s, const float4* __restrict__ posm, const float4* __restrict__ vel, int n,
                                                         int* __restrict__ block_counts) {
    const int i = blockIdx.x * kMgThreads + threadIdx.x;
    int c[3] = {0, 0, 0};
    if (i < n) {
        const MgItem it = mg_classify(s, posm[i], vel[i]);
        c[0] = it.keep ? it.nimg : 0;
        c[1] = it.left ? it.nimg : 0;
        c[2] = it.right ? it.nimg : 0;
    }
#pragma unroll
    for (int k = 0; k < 3; k++) {
        int tot;
        block_exclusive_scan(c[k], &tot);
        if (threadIdx.x == 0) block_counts[3 * blockIdx.x + k] = tot;
    }
}

// pass 2: the per-block counts become exclusive offsets (and totals).  One block of 1024 threads per stream
// (keep / left / right), four entries per thread per round: 70 K blocks take 18 rounds (the first version, one
// 256-thread block for all three streams, took 0.34 ms of a 7.4 ms step at 17.8 M particles).
constexpr int kMgScanThreads = 1024;

__global__ void __launch_bounds__(kMgScanThreads) k_mg_scan(int* __restrict__ block_counts, int nblocks, int* __restrict__ totals) {
    __shared__ int s_warp[kMgScanThreads / 32];
    const int k = blockIdx.x; // stream
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    int carry = 0;
    for (int base = 0; base < nblocks; base += 4 * kMgScanThreads) {
        const int b0 = base + 4 * threadIdx.x;
        int v[4];
        int tsum = 0;
#pragma unroll
        for (int q = 0; q < 4; q++) {
            v[q] = (b0 + q < nblocks) ? block_counts[3 * (b0 + q) + k] : 0;
            tsum += v[q];
        }
        int inc = tsum;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int t = __shfl_up_sync(0xffffffffu, inc, o);
            if (lane >= o) inc += t;
        }
        if (lane == 31) s_warp[warp] = inc;
        __syncthreads();
        if (warp == 0) {
            int w = s_warp[lane];
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const int t = __shfl_up_sync(0xffffffffu, w, o);
                if (lane >= o) w += t;
            }
            s_warp[lane] = w; // inclusive over warps
        }
        __syncthreads();
        int ex = carry + (warp ? s_warp[warp - 1] : 0) + inc - tsum;
        const int round_total = s_warp[kMgScanThreads / 32 - 1];
#pragma unroll
        for (int q = 0; q < 4; q++) {
            if (b0 + q < nblocks) block_counts[3 * (b0 + q) + k] = ex;
            ex += v[q];
        }
        carry += round_total;
        __syncthreads(); // s_warp is reused next round
    }
    if (threadIdx.x == 0) totals[k] = carry;
}

// pass 3: write.  keep -> (posm_out, vel_out) with the ownership flag refreshed; left/right -> messages.
__global__ void __launch_bounds__(kMgThreads) k_mg_write(Slab s, const float4* __restrict__ posm, const float4* __restrict__ vel, int n,
                                                         const int* __restrict__ block_offsets, const int* __restrict__ totals,
                                                         float4* __restrict__ posm_out, float4* __restrict__ vel_out, int out_capacity,
                                                         float4* __restrict__ msg_left, float4* __restrict__ msg_right, int capacity) {
    const int i = blockIdx.x * kMgThreads + threadIdx.x;
    MgItem it;
    it.keep = it.left = it.right = false;
    it.nimg = 0;
    if (i < n) it = mg_classify(s, posm[i], vel[i]);
    int off[3];
    const int cnt[3] = {it.keep ? it.nimg : 0, it.left ? it.nimg : 0, it.right ? it.nimg : 0};
#pragma unroll
    for (int k = 0; k < 3; k++) {
        int tot;
        off[k] = block_exclusive_scan(cnt[k], &tot) + block_offsets[3 * blockIdx.x + k];
    }
    for (int img = 0; img < cnt[0]; img++) {
        if (off[0] + img >= out_capacity) break;
        const float4 q = mg_image(it, img, 0.0f);
        posm_out[off[0] + img] = q;
        vel_out[off[0] + img] = make_float4(it.v.x, it.v.y, it.v.z, (img == 0 && mg_inside(s, q)) ? 1.0f : 0.0f);
    }
    for (int img = 0; img < cnt[1]; img++) {
        if (off[1] + img >= capacity) break;
        float4 q = mg_image(it, img, s.shift_left);
        // a migrant through the periodic seam must land INSIDE the far slab even if x + period rounds up onto its face
        if (s.shift_left != 0.0f && it.p.x < s.lo) q.x = fminf(q.x, nextafterf(s.lo + s.shift_left, -INFINITY));
        msg_left[2 + 2 * (off[1] + img)] = q;
        msg_left[3 + 2 * (off[1] + img)] = it.v;
    }
    for (int img = 0; img < cnt[2]; img++) {
        if (off[2] + img >= capacity) break;
        float4 q = mg_image(it, img, s.shift_right);
        if (s.shift_right != 0.0f && it.p.x >= s.hi) q.x = fmaxf(q.x, s.hi + s.shift_right);
        msg_right[2 + 2 * (off[2] + img)] = q;
        msg_right[3 + 2 * (off[2] + img)] = it.v;
    }
    if (blockIdx.x == 0 && threadIdx.x == 0) { // headers: count, capacity
        if (msg_left) msg_left[0] = make_float4(__int_as_float(s.has_left ? totals[1] : 0), __int_as_float(capacity), 0.0f, 0.0f);
        if (msg_right) msg_right[0] = make_float4(__int_as_float(s.has_right ? totals[2] : 0), __int_as_float(capacity), 0.0f, 0.0f);
    }
}

// append a received message behind the kept particles; counts[0] = kept, counts[1] = records already appended
__global__ void __launch_bounds__(256) k_mg_append(Slab s, const float4* __restrict__ msg, const int* __restrict__ totals, int which,
                                                   float4* __restrict__ posm_out, float4* __restrict__ vel_out, int out_capacity,
                                                   int* __restrict__ recv_counts) {
    const int count = min(__float_as_int(msg[0].x), __float_as_int(msg[0].y)); // an overflowing sender truncated its payload
    const int base = totals[0] + (which ? recv_counts[0] : 0);
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        recv_counts[which] = count;
        recv_counts[2 + which] = __float_as_int(msg[0].x); // what the sender wanted to send (overflow check)
    }
    for (int r = blockIdx.x * blockDim.x + threadIdx.x; r < count; r += gridDim.x * blockDim.x) {
        if (base + r >= out_capacity) break;
        const float4 p = msg[2 + 2 * r], v = msg[3 + 2 * r];
        posm_out[base + r] = p;
        vel_out[base + r] = make_float4(v.x, v.y, v.z, mg_inside(s, p) ? 1.0f : 0.0f);
    }
}

} // namespace sphb
