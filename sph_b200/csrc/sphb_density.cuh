// Density + equation of state (SPH/SphUtils.cpp:55-63, 100-149): kernel shape, candidate runs, row groups, the packed-FP32 density kernel that also records the neighbour masks.
// Part of libsphb200.so (sm_100a); included through sphb_device.cuh.
#pragma once
#include "sphb_common.cuh"
#include "sphb_grid.cuh"

namespace sphb {

// ---------------------------------------------------------------------------------------------------
// K3: scatter (original id, current index) pairs into their cell segment in arrival order, then
// k_rank_gather replaces the arrival order by the STABLE order (ascending original id inside a cell,
// i.e. exactly what a CPU stable counting sort of the uploaded order produces) and moves the state.

__global__ void __launch_bounds__(256) k_scatter(const int* __restrict__ key, const int* __restrict__ slot, const int* __restrict__ oid,
                                                 const int* __restrict__ cell_start, int2* __restrict__ pairs, int n, const int* __restrict__ n_dev) {
    if (n_dev) n = *n_dev; // the count lives on the device (sphb_mg_unpack_async); the launch was sized from an estimate
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x)
        pairs[cell_start[key[i]] + slot[i]] = make_int2(oid ? oid[i] : i, i); // oid == nullptr: ids are the slot numbers
}

// A cell that holds more than this many particles (a clamped boundary cell full of strays, or everything gone
// NaN and filed under cell 0) keeps its ARRIVAL order instead: counting smaller ids is quadratic in the cell size.
// Still a valid permutation, just not the stable one (and not reproducible run to run) for that cell.
constexpr int kStableRankLimit = 4096;

template <bool kCarryAll>
__device__ __forceinline__ void rank_gather_one(int p, const int2* pairs, const int* key, const int* cell_start,
                                                const float4* __restrict__ posm_in, const float4* __restrict__ vel_in,
                                                const float4* __restrict__ aux_in, const float4* __restrict__ acc_in,
                                                float4* __restrict__ posm_out, float4* __restrict__ vel_out,
                                                float4* __restrict__ aux_out, float4* __restrict__ acc_out,
                                                int* __restrict__ oid_out, int ncell, unsigned int* __restrict__ stats);

template <bool kCarryAll>
__global__ void __launch_bounds__(256) k_rank_gather(const int2* __restrict__ pairs, const int* __restrict__ key, const int* __restrict__ cell_start,
                                                     const float4* __restrict__ posm_in, const float4* __restrict__ vel_in,
                                                     const float4* __restrict__ aux_in, const float4* __restrict__ acc_in,
                                                     float4* __restrict__ posm_out, float4* __restrict__ vel_out,
                                                     float4* __restrict__ aux_out, float4* __restrict__ acc_out,
                                                     int* __restrict__ oid_out, int n, const int* __restrict__ n_dev,
                                                     const GridDesc* __restrict__ gd, unsigned int* __restrict__ stats) {
    if (n_dev) n = *n_dev;
    const int ncell = gd->ncell;
    for (int p = blockIdx.x * blockDim.x + threadIdx.x; p < n; p += gridDim.x * blockDim.x)
        rank_gather_one<kCarryAll>(p, pairs, key, cell_start, posm_in, vel_in, aux_in, acc_in, posm_out, vel_out, aux_out, acc_out, oid_out, ncell, stats);
}

template <bool kCarryAll>
__device__ __forceinline__ void rank_gather_one(int p, const int2* pairs, const int* key, const int* cell_start,
                                                const float4* __restrict__ posm_in, const float4* __restrict__ vel_in,
                                                const float4* __restrict__ aux_in, const float4* __restrict__ acc_in,
                                                float4* __restrict__ posm_out, float4* __restrict__ vel_out,
                                                float4* __restrict__ aux_out, float4* __restrict__ acc_out,
                                                int* __restrict__ oid_out, int ncell, unsigned int* __restrict__ stats) {
    const int2 me = pairs[p];
    const int k = key[me.y];
    if (k >= ncell) return; // a trash bin: records the slab exchange marked dead end here and are not carried over
    const int b = cell_start[k], e = cell_start[k + 1];
    int rank = 0;
    if (e - b <= kStableRankLimit) {
        // ids are unique after an upload; after slab exchanges two records of one cell may carry the same id (ranks number their
        // uploads independently unless told otherwise): the arrival slot breaks the tie so that the result stays a permutation
        for (int q = b; q < e; q++) rank += (pairs[q].x < me.x) || (pairs[q].x == me.x && q < p);
    } else {
        rank = p - b; // arrival order
        if (p == b) atomicAdd(stats, 1u); // one count per such cell and sort (sphb_get_stats: arrival_order_cells)
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

// the work of one block of kRgThreads threads (k_row_groups, and one phase of k_sort_small)
template <int kPer>
__device__ __forceinline__ void row_groups_impl(int nx, int nrows, const int* cell_start, int* gstart, int* counts) {
    // counts[1] = particles in real cells = the live count after this sort (records filed in the trash bins excluded):
    // what the kernels after the sort use as n when the host does not know it (sphb_mg_unpack_async)
    if (threadIdx.x == 0 && counts) counts[1] = cell_start[nx * nrows];
    __shared__ int s_warp[kRgThreads / 32];
    __shared__ int s_carry;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (threadIdx.x == 0) s_carry = 0;
    __syncthreads();
    // kPer consecutive rows per thread and round (all their loads in flight together): the block is alone on the GPU, so the
    // number of dependent rounds is its running time (30 K rows, four per thread: 51 -> 36 us)
    for (int base = 0; base < nrows; base += kRgThreads * kPer) {
        const int rho0 = base + threadIdx.x * kPer;
        int b[kPer + 1];
#pragma unroll
        for (int q = 0; q <= kPer; q++) b[q] = cell_start[min(rho0 + q, nrows) * nx];
        int vq[kPer], v = 0;
#pragma unroll
        for (int q = 0; q < kPer; q++) {
            vq[q] = rho0 + q < nrows ? (b[q + 1] - b[q] + kDT - 1) / kDT : 0;
            v += vq[q];
        }
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
        int ex = carry + (warp ? s_warp[warp - 1] : 0) + inc - v;
#pragma unroll
        for (int q = 0; q < kPer; q++) {
            if (rho0 + q < nrows) gstart[rho0 + q] = ex;
            ex += vq[q];
        }
        __syncthreads();
        if (threadIdx.x == kRgThreads - 1) s_carry = ex;
        __syncthreads();
    }
    if (threadIdx.x == 0) gstart[nrows] = s_carry;
}

__device__ __forceinline__ void row_groups_block(int nx, int nrows, const int* cell_start, int* gstart, int* counts) {
    if (nrows > kRgThreads) row_groups_impl<4>(nx, nrows, cell_start, gstart, counts);
    else row_groups_impl<1>(nx, nrows, cell_start, gstart, counts); // small grids: one round either way, fewer loads
}

__global__ void __launch_bounds__(kRgThreads) k_row_groups(const GridDesc* __restrict__ gd, const int* __restrict__ cell_start,
                                                          int* __restrict__ gstart, int* __restrict__ counts) {
    row_groups_block(gd->dims[0], gd->dims[1] * gd->dims[2], cell_start, gstart, counts);
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
                                                   int mstride, int i, const float4& pi, int r, int jb, int je, float& sum) {
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
        mask[(size_t)(2 * r) * mstride + i] = m0;      // word w of row r covers candidates jb + 32*w ...
        mask[(size_t)(2 * r + 1) * mstride + i] = m1;
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
// SPHB_DENSITY_WFORM 1: the packed loop accumulates m min(d2 - R^2, 0)^3 and the factor -1/h^6 is applied once per
// target after the loop (one packed multiply less per candidate and target pair); 0: m max(h^2 - d2/h^2, 0)^3 per candidate.
#ifndef SPHB_DENSITY_WFORM
#define SPHB_DENSITY_WFORM 1
#endif
#ifndef SPHB_DENSITY_TILE
#define SPHB_DENSITY_TILE 0
#endif
#ifndef SPHB_DENSITY_MINBLOCKS
#define SPHB_DENSITY_MINBLOCKS 8
#endif
// kLanes: 1, or 16 for small scenes: the 9 rows of a group go to 9 of 16 consecutive lanes (partial sums joined by
// shuffles, each lane writes its own row's mask words), because with a few hundred groups the GPU is empty and what
// counts is the length of one thread's chain: 972 particles (main.cpp's scene) 30 us -> see DESIGN.md.
// kTile (experiment, SPHB_DENSITY_TILE): the warp stages the union of its lanes' candidate runs of each stencil row in a
// shared-memory tile (coalesced loads, 16-byte records XOR-swizzled so that lanes four records apart read different bank
// groups) and the candidate loop reads LDS.128 instead of LDG.128.  All 32 lanes then stay in the row loop together.
#if SPHB_DENSITY_TILE == 3
constexpr int kTileRecs = 160; // records per warp tile buffer (2.5 KB; two buffers per warp: the next row is copied while this one is read)
#else
constexpr int kTileRecs = 256; // records per warp tile (4 KB)
#endif
__device__ __forceinline__ int tile_swz(int k) { return k ^ ((k >> 3) & 7); }
__device__ __forceinline__ unsigned int smem_u32(const void* p) { return (unsigned int)__cvta_generic_to_shared(p); }
// bounded wait for the phase of an mbarrier (a lost copy must not hang the GPU: the caller then reads whatever is there and the
// parity tests fail loudly)
__device__ __forceinline__ bool mbar_wait(unsigned int bar, unsigned int parity) {
    for (int it = 0; it < (1 << 22); it++) {
        unsigned int ok;
        asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2; selp.u32 %0, 1, 0, p; }" : "=r"(ok) : "r"(bar), "r"(parity) : "memory");
        if (ok) return true;
    }
    return false;
}

template <bool kAux, bool kMask, int kLanes, bool kTile = false>
__device__ __forceinline__ void density_thread(const Consts& c, const GridDesc& g, const int* __restrict__ cell_start,
                                               const int* __restrict__ gstart,
                                               const float4* __restrict__ posm, const float4* __restrict__ vel,
                                               float4* __restrict__ aux, float4* __restrict__ vr,
                                               unsigned int* __restrict__ mask, const int n /* mask stride */, const int tg,
                                               const int nrows_grid, const int ngroups, float4* tile = nullptr,
                                               unsigned long long* bars = nullptr, int* meta = nullptr) {
    const int t_raw = tg / kLanes, sub = tg % kLanes; // group, and which of its rows this lane takes (kLanes > 1)
    if ((tg & ~31) / kLanes >= ngroups) return;    // whole warp beyond the last group
    const bool live = t_raw < ngroups;
    const int t = live ? t_raw : ngroups - 1;       // spare lanes of the last warp shadow the last group (warp-wide votes below)
    const int lo_r = find_row(gstart, nrows_grid, t);
    bool active = live;                             // kTile: lanes without work stay for the warp-wide staging
    if (!kTile && !live) return;
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
    {   // outer ghosts (more than a support radius outside the slab or the periodic box): nobody reads their density, masks or
        // (vx, vy, vz, rho) record -- the fused force kernel passes ghosts through -- so a group made of them is left at once
        bool needed = false;
#pragma unroll
        for (int k = 0; k < kDT; k++)
            needed |= pi[k].x >= c.need_lo[0] && pi[k].x < c.need_hi[0] && pi[k].y >= c.need_lo[1] && pi[k].y < c.need_hi[1] &&
                      pi[k].z >= c.need_lo[2] && pi[k].z < c.need_hi[2];
        if (!kTile && !needed) return;
        active = active && needed;
    }
    // kTile: the tile holds ONE contiguous run per stencil row, which needs every lane of the warp to work on the same row of cells
    const bool same_row = kTile ? __all_sync(0xffffffffu, lo_r == __shfl_sync(0xffffffffu, lo_r, 0)) : false;
    const int S = g.stencil, Sx = g.stencil_x;
    const int nrows = (2 * S + 1) * (2 * S + 1);
    float sum[kDT] = {0.0f, 0.0f, 0.0f, 0.0f};
    {
        const int cxlo = min(min(cx[0], cx[1]), min(cx[2], cx[3])), cxhi = max(max(cx[0], cx[1]), max(cx[2], cx[3]));
        const float2 nx01 = make_float2(-pi[0].x, -pi[1].x), ny01 = make_float2(-pi[0].y, -pi[1].y), nz01 = make_float2(-pi[0].z, -pi[1].z);
        const float2 nx23 = make_float2(-pi[2].x, -pi[3].x), ny23 = make_float2(-pi[2].y, -pi[3].y), nz23 = make_float2(-pi[2].z, -pi[3].z);
        const float2 nR2 = make_float2(-c.R2, -c.R2), nih2 = make_float2(-c.inv_h2, -c.inv_h2);
        float2 acc01 = make_float2(0.0f, 0.0f), acc23 = make_float2(0.0f, 0.0f);
#if SPHB_DENSITY_TILE == 3
        // TMA variant: the union of the warp's candidate runs of a stencil row is ONE contiguous run of the sorted array between two
        // x-cells that are the same for all nine rows; lane 0 copies it into the warp's tile with one cp.async.bulk (completion on
        // an mbarrier), a row ahead of its use, into the buffer the row before last has just been read from
        int x0W = 0, x1W = -1;
        unsigned int ph = 0; // phase parity of the two barriers
        bool tma_ok = false;
        auto tma_issue = [&](int r, int buf) { // lane 0 only
            const int W = 2 * S + 1, z = cz[0] + r / W - S, y = cy[0] + r % W - S;
            int jb = -1;
            if (!(z < 0 || z >= g.dims[2] || y < 0 || y >= g.dims[1])) {
                const int row = g.dims[0] * (y + g.dims[1] * z);
                const int b = __ldg(cell_start + row + x0W), e = __ldg(cell_start + row + x1W + 1);
                if (e > b && e - b <= kTileRecs) {
                    jb = b;
                    const unsigned int bytes = (unsigned int)(e - b) * 16u, bar = smem_u32(bars + buf), dst = smem_u32(tile + buf * kTileRecs);
                    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
                    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst), "l"(posm + b),
                                 "r"(bytes), "r"(bar)
                                 : "memory");
                }
            }
            meta[buf] = jb;
        };
        if (kTile) {
            x0W = __reduce_min_sync(0xffffffffu, active ? max(cxlo - Sx, 0) : 0x7fffffff);
            x1W = __reduce_max_sync(0xffffffffu, active ? min(cxhi + Sx, g.dims[0] - 1) : -1);
            tma_ok = same_row && x1W >= x0W;
            if (tma_ok && (threadIdx.x & 31) == 0) tma_issue(0, 0);
        }
#endif
        // rows in canonical order r = (dz+S)*(2S+1) + (dy+S), counted without the divisions a runtime stencil would cost
        int dy_r = -S, dz_r = -S;
        unsigned int* mrow = mask + i0 + 2 * (size_t)n * (kLanes > 1 ? sub : 0); // word w of row r of target k: mrow[(2r+w)*n + k]
        for (int r = kLanes > 1 ? sub : 0; r < nrows; r += kLanes, mrow += 2 * (size_t)n * kLanes) {
            if (kLanes > 1) { dz_r = r / (2 * S + 1) - S; dy_r = r % (2 * S + 1) - S; }
            const int z = cz[0] + dz_r, y = cy[0] + dy_r;
            if (kLanes == 1 && ++dy_r > S) { dy_r = -S; ++dz_r; }
            int jbu = 0, jeu = 0, jbk[kDT] = {0, 0, 0, 0};
            const bool inside = (!kTile || active) && !(z < 0 || z >= g.dims[2] || y < 0 || y >= g.dims[1]);
            if (inside) {
                const int row = g.dims[0] * (y + g.dims[1] * z);
                jbu = __ldg(cell_start + row + max(cxlo - Sx, 0));
                jeu = __ldg(cell_start + row + min(cxhi + Sx, g.dims[0] - 1) + 1);
#pragma unroll
                for (int k = 0; k < kDT; k++) jbk[k] = (cx[k] == cxlo) ? jbu : __ldg(cell_start + row + max(cx[k] - Sx, 0));
            }
            int jbW = 0;
            bool use_tile = false;
#if SPHB_DENSITY_TILE == 3
            if (kTile && tma_ok) {
                __syncwarp(); // every lane is done with the row before: its buffer may be overwritten
                if ((threadIdx.x & 31) == 0 && r + 1 < nrows) tma_issue(r + 1, (r + 1) & 1);
                __syncwarp();
                jbW = meta[r & 1];
                if (jbW >= 0) {
                    mbar_wait(smem_u32(bars + (r & 1)), (ph >> (r & 1)) & 1u);
                    ph ^= 1u << (r & 1);
                    use_tile = true;
                }
            }
            if (false) {
#else
            if (kTile) { // stage the union of the lanes' runs of this row
#endif
                const bool plain = inside && jeu > jbu && jeu - jbu <= kMaskBits;
                jbW = __reduce_min_sync(0xffffffffu, plain ? jbu : 0x7fffffff);
                const int jeW = __reduce_max_sync(0xffffffffu, plain ? jeu : 0);
                use_tile = same_row && jeW > jbW && jeW - jbW <= kTileRecs;
                if (use_tile) {
                    __syncwarp(); // the previous row's readers are done
                    for (int k = threadIdx.x & 31; k < jeW - jbW; k += 32) tile[SPHB_DENSITY_TILE == 2 ? k : tile_swz(k)] = ldg4(posm + jbW + k);
                    __syncwarp();
                }
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
                // Candidates two at a time; an odd run repeats its last candidate with mass 0 (adds exactly 0, and its
                // duplicate mask bit is dropped below) instead of paying a separate tail loop that some lane of the
                // warp would take in nearly every row.
                auto test4 = [&](const float4& pj, float mj) {
                    const float2 xx = make_float2(pj.x, pj.x), yy = make_float2(pj.y, pj.y), zz = make_float2(pj.z, pj.z), mm = make_float2(mj, mj);
                    {
                        const float2 dx = __fadd2_rn(xx, nx01), dy = __fadd2_rn(yy, ny01), dz = __fadd2_rn(zz, nz01);
                        const float2 d2 = __ffma2_rn(dx, dx, __ffma2_rn(dy, dy, __fmul2_rn(dz, dz)));
                        const float2 u = __fadd2_rn(d2, nR2);
                        if (kMask) {
                            m[0] = __funnelshift_l(__float_as_uint(u.x), m[0], 1);
                            m[1] = __funnelshift_l(__float_as_uint(u.y), m[1], 1);
                        }
                        if (kAux) {
#if SPHB_DENSITY_WFORM
                            const float2 v = make_float2(fminf(u.x, 0.0f), fminf(u.y, 0.0f)); // min(d2 - R^2, 0) = -h^2 max(t, 0)
                            acc01 = __ffma2_rn(__fmul2_rn(v, v), __fmul2_rn(mm, v), acc01);
#else
                            float2 t = __fmul2_rn(u, nih2);
                            t.x = fmaxf(t.x, 0.0f);
                            t.y = fmaxf(t.y, 0.0f);
                            acc01 = __ffma2_rn(mm, __fmul2_rn(__fmul2_rn(t, t), t), acc01);
#endif
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
#if SPHB_DENSITY_WFORM
                            const float2 v = make_float2(fminf(u.x, 0.0f), fminf(u.y, 0.0f)); // min(d2 - R^2, 0) = -h^2 max(t, 0)
                            acc23 = __ffma2_rn(__fmul2_rn(v, v), __fmul2_rn(mm, v), acc23);
#else
                            float2 t = __fmul2_rn(u, nih2);
                            t.x = fmaxf(t.x, 0.0f);
                            t.y = fmaxf(t.y, 0.0f);
                            acc23 = __ffma2_rn(mm, __fmul2_rn(__fmul2_rn(t, t), t), acc23);
#endif
                        }
                    }
                };
                const int len = jz - ja;
                if (len > 0) {
                    // software pipeline: the next pair's loads are in flight during this pair's arithmetic (2.14 -> 2.08 ms
                    // at 16.7 M; only with all 8 blocks resident -- at 80 or 96 registers it loses more than it gains)
#if SPHB_DENSITY_TILE == 3
                    const float4* src = (kTile && use_tile) ? (const float4*)tile + (r & 1) * kTileRecs - jbW : posm;
                    auto cand = [&](int j) -> float4 { return kTile ? src[j] : ldg4(posm + j); };
#elif SPHB_DENSITY_TILE == 2
                    // plain tile layout, one loop for both sources: a generic pointer that is either the tile shifted by the union's
                    // first slot or the global array (no swizzle arithmetic, no second code path)
                    const float4* src = (kTile && use_tile) ? (const float4*)tile - jbW : posm;
                    auto cand = [&](int j) -> float4 { return kTile ? src[j] : ldg4(posm + j); };
#else
                    auto cand = [&](int j) -> float4 { return (kTile && use_tile) ? tile[tile_swz(j - jbW)] : ldg4(posm + j); };
#endif
                    float4 q0 = cand(ja), q1 = cand(min(ja + 1, jz - 1));
                    for (int j = ja; j < jz; j += 2) {
                        const bool two = j + 1 < jz;
                        const float4 p0 = q0, p1 = q1;
                        q0 = cand(min(j + 2, jz - 1));
                        q1 = cand(min(j + 3, jz - 1));
                        test4(p0, p0.w);
                        test4(p1, two ? p1.w : 0.0f);
                    }
                }
                if (kMask && (len & 1)) {
#pragma unroll
                    for (int k = 0; k < kDT; k++) m[k] >>= 1;
                }
                // the bits went in newest-lowest: candidate q of this word sits at bit (len-1-q); put it at bit q
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
                    if (k < nt && (!kTile || active)) { // kTile: lanes that only help with the staging write nothing
                        mrow[k] = (unsigned int)m64;
                        mrow[(size_t)n + k] = (unsigned int)(m64 >> 32);
                    }
                }
            }
        }
#if SPHB_DENSITY_WFORM
        // the packed loop summed m (d2 - R^2)^3 over the neighbours: W ~ (h^2 - d2/h^2)^3 = -(d2 - R^2)^3 / h^6
        const float k3 = -(c.inv_h2 * c.inv_h2 * c.inv_h2);
        sum[0] = fmaf(k3, acc01.x, sum[0]); sum[1] = fmaf(k3, acc01.y, sum[1]); sum[2] = fmaf(k3, acc23.x, sum[2]); sum[3] = fmaf(k3, acc23.y, sum[3]);
#else
        sum[0] += acc01.x; sum[1] += acc01.y; sum[2] += acc23.x; sum[3] += acc23.y;
#endif
    }
    if (kLanes > 1) { // join the rows' partial sums; lane `sub == 0` of the group finishes
        // groups leave as a whole, so a 16-lane segment is complete or gone: name exactly its lanes (the _sync re-converges them)
        const unsigned int am = kLanes == 16 ? ((threadIdx.x & 16) ? 0xffff0000u : 0x0000ffffu) : 0xffffffffu;
#pragma unroll
        for (int k = 0; k < kDT; k++)
#pragma unroll
            for (int o = kLanes / 2; o > 0; o >>= 1) sum[k] += __shfl_xor_sync(am, sum[k], o);
        if (sub != 0) return;
    }
    if (kTile && !active) return;
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

// The kernel: thread tg takes group tg (of kLanes lanes); the group total is read from the device.  n = stride of the mask
// rows (a capacity, not the particle count).  When the host only has an ESTIMATE of the particle count
// (sphb_mg_unpack_async) the main launch (kLoop = false) covers the estimate and a small second launch with kLoop = true,
// first = threads of the main launch, walks whatever lies beyond in grid strides -- normally nothing.  (One launch that
// always loops cost 12 % here: the loop-carried state pushes the 64-register kernel into more spills.)
template <bool kAux, bool kMask, int kLanes = 1, bool kLoop = false>
__global__ void __launch_bounds__(128, SPHB_DENSITY_MINBLOCKS) k_density(Consts c, const GridDesc* __restrict__ gd, const int* __restrict__ cell_start,
                                                 const int* __restrict__ gstart,
                                                 const float4* __restrict__ posm, const float4* __restrict__ vel,
                                                 float4* __restrict__ aux, float4* __restrict__ vr,
                                                 unsigned int* __restrict__ mask, int n, int first) {
    const GridDesc g = *gd;
    const int nrows_grid = g.dims[1] * g.dims[2];
    const int ngroups = __ldg(gstart + nrows_grid);
#if SPHB_DENSITY_TILE == 3
    if (!kLoop && kLanes == 1) {
        __shared__ __align__(128) float4 s_tile[4][2][kTileRecs];
        __shared__ __align__(8) unsigned long long s_bar[4][2];
        __shared__ int s_meta[4][2];
        const int w = threadIdx.x >> 5;
        if ((threadIdx.x & 31) == 0) {
            asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&s_bar[w][0])) : "memory");
            asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&s_bar[w][1])) : "memory");
            asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
            s_meta[w][0] = s_meta[w][1] = -1;
        }
        __syncwarp();
        density_thread<kAux, kMask, kLanes, true>(c, g, cell_start, gstart, posm, vel, aux, vr, mask, n, first + blockIdx.x * blockDim.x + threadIdx.x,
                                                  nrows_grid, ngroups, &s_tile[w][0][0], s_bar[w], s_meta[w]);
        return;
    }
#elif SPHB_DENSITY_TILE
    if (!kLoop && kLanes == 1) {
        __shared__ float4 s_tile[4][kTileRecs];
        density_thread<kAux, kMask, kLanes, true>(c, g, cell_start, gstart, posm, vel, aux, vr, mask, n, first + blockIdx.x * blockDim.x + threadIdx.x,
                                                  nrows_grid, ngroups, s_tile[threadIdx.x >> 5]);
        return;
    }
#endif
    if (!kLoop) {
        density_thread<kAux, kMask, kLanes>(c, g, cell_start, gstart, posm, vel, aux, vr, mask, n, first + blockIdx.x * blockDim.x + threadIdx.x, nrows_grid, ngroups);
    } else {
        const long long nthreads = (long long)ngroups * kLanes;
        for (long long base = first + (long long)blockIdx.x * blockDim.x; base < nthreads; base += (long long)gridDim.x * blockDim.x)
            density_thread<kAux, kMask, kLanes>(c, g, cell_start, gstart, posm, vel, aux, vr, mask, n, (int)base + (int)threadIdx.x, nrows_grid, ngroups);
    }
}

// ---------------------------------------------------------------------------------------------------
// Small scenes: the whole grid build in ONE launch of one block.  At main.cpp's 972 particles the eight kernels
// above are eight dependent launches of 3-5 us each for a few hundred nanoseconds of work; here their bodies run
// back to back with block barriers in between -- same functions, same arithmetic, same stable order.
__device__ __forceinline__ int scan_block_1024(int v, int* s_warp, int* total) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
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
        s_warp[lane] = w; // inclusive over warps
    }
    __syncthreads();
    const int ex = (warp ? s_warp[warp - 1] : 0) + inc - v;
    *total = s_warp[31];
    __syncthreads(); // s_warp is reused
    return ex;
}

template <bool kCarryAll>
__global__ void __launch_bounds__(kRgThreads) k_sort_small(Consts c, BBoxAcc* acc, GridDesc* gd, int* cell_start, int* gstart,
                                                          int* key, int* slot, int2* pairs, const int* __restrict__ oid_in,
                                                          const float4* __restrict__ posm_in, const float4* __restrict__ vel_in,
                                                          const float4* __restrict__ aux_in, const float4* __restrict__ acc_in,
                                                          float4* __restrict__ posm_out, float4* __restrict__ vel_out,
                                                          float4* __restrict__ aux_out, float4* __restrict__ acc_out,
                                                          int* __restrict__ oid_out, int n, const int* __restrict__ n_dev, int* __restrict__ counts,
                                                          int rearm_bbox, int trash, unsigned int* __restrict__ stats) {
    if (n_dev) n = *n_dev;
    __shared__ GridDesc s_g;
    __shared__ int s_scan[32];
    const int tid = threadIdx.x;
    if (rearm_bbox) { // bounding-box grid mode and no box at hand: take it here (SphUtils.cpp:11-23) rather than in two launches of its own
        if (tid < 3) {
            acc->lo[tid] = f2ord(INFINITY);
            acc->hi[tid] = f2ord(-INFINITY);
        }
        __syncthreads();
        float lo[3] = {INFINITY, INFINITY, INFINITY}, hi[3] = {-INFINITY, -INFINITY, -INFINITY};
        for (int i = tid; i < n; i += kRgThreads) {
            const float4 p = ldg4(posm_in + i);
            lo[0] = fminf(lo[0], p.x); hi[0] = fmaxf(hi[0], p.x);
            lo[1] = fminf(lo[1], p.y); hi[1] = fmaxf(hi[1], p.y);
            lo[2] = fminf(lo[2], p.z); hi[2] = fmaxf(hi[2], p.z);
        }
        bbox_block_commit(lo, hi, acc);
        __threadfence_block();
        __syncthreads();
    }
    if (tid == 0) {
        derive_grid(c, acc, s_g);
        *gd = s_g;
    }
    __syncthreads();
    const GridDesc g = s_g;
    for (int i = tid; i <= g.nbins; i += kRgThreads) cell_start[i] = 0;
    __syncthreads();
    for (int i = tid; i < n; i += kRgThreads) { // K1
        const float4 p = ldg4(posm_in + i);
        int cx, cy, cz;
        int k = cell_key(g, p.x, p.y, p.z, cx, cy, cz);
        if (trash && p.w < 0.0f) k = g.ncell + (i & (kTrashBins - 1));
        key[i] = k;
        slot[i] = atomicAdd(cell_start + k, 1);
    }
    __syncthreads();
    int carry = 0;
    for (int base = 0; base <= g.nbins; base += kRgThreads) { // K2: exclusive scan in place
        const int i = base + tid;
        const int v = i <= g.nbins ? cell_start[i] : 0;
        int tot;
        const int ex = scan_block_1024(v, s_scan, &tot);
        if (i <= g.nbins) cell_start[i] = carry + ex;
        carry += tot;
    }
    __syncthreads();
    row_groups_block(g.dims[0], g.dims[1] * g.dims[2], cell_start, gstart, counts);
    for (int i = tid; i < n; i += kRgThreads) pairs[cell_start[key[i]] + slot[i]] = make_int2(oid_in ? oid_in[i] : i, i); // K3
    __syncthreads();
    for (int p = tid; p < n; p += kRgThreads)
        rank_gather_one<kCarryAll>(p, pairs, key, cell_start, posm_in, vel_in, aux_in, acc_in, posm_out, vel_out, aux_out, acc_out, oid_out, g.ncell, stats);
}


} // namespace sphb
