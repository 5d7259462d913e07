// Grid build (update_cells, SPH/SphUtils.cpp:8-47): bounding box, grid description, cell keys + histogram, exclusive scan, scatter and stable rank + reorder.
// Part of libsphb200.so (sm_100a); included through sphb_device.cuh.
#pragma once
#include "sphb_common.cuh"

namespace sphb {

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
    g.nbins = g.ncell + kTrashBins;
    g.stencil = (c.cell >= c.h2) ? 1 : 2;
    g.stencil_x = c.stencil_x;
}

__global__ void __launch_bounds__(256) k_grid_setup(Consts c, const BBoxAcc* acc, GridDesc* out, int* __restrict__ cell_count) {
    GridDesc g;
    derive_grid(c, acc, g);
    if (blockIdx.x == 0 && threadIdx.x == 0) *out = g;
    const int total = g.nbins + 1;
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

__global__ void __launch_bounds__(256) k_key_hist(const GridDesc* __restrict__ gd, const float4* __restrict__ posm, int n, const int* __restrict__ n_dev,
                                                  int* __restrict__ key, int* __restrict__ slot, int* __restrict__ cell_count,
                                                  BBoxAcc* acc, int rearm_bbox, int trash) {
    const GridDesc g = *gd;
    if (n_dev) n = *n_dev; // the count lives on the device (sphb_mg_unpack_async); the launch was sized from an estimate
    if (rearm_bbox && blockIdx.x == 0 && threadIdx.x < 3) {
        acc->lo[threadIdx.x] = f2ord(INFINITY);
        acc->hi[threadIdx.x] = f2ord(-INFINITY);
    }
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        float4 p = ldg4(posm + i);
        int cx, cy, cz;
        int k = cell_key(g, p.x, p.y, p.z, cx, cy, cz);
        if (trash && p.w < 0.0f) k = g.ncell + (i & (kTrashBins - 1)); // marked dead by sphb_mg_pack: filed behind the last cell
        key[i] = k;
        slot[i] = atomicAdd(cell_count + k, 1);
    }
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
    const int total = gd->nbins + 1;
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
    const int ntile = (gd->nbins + 1 + kScanTile - 1) / kScanTile;
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
    const int total = gd->nbins + 1;
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

} // namespace sphb
