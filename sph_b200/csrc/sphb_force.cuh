// Pair forces, integration, walls (SPH/SphUtils.cpp:68-79, 157-239, 295-304, 321-372): pair body, fused force + integrate + collide kernel, optional Laplacian viscosity, phase kernels, kernel-function evaluation.
// Part of libsphb200.so (sm_100a); included through sphb_device.cuh.
#pragma once
#include "sphb_common.cuh"
#include "sphb_density.cuh"

namespace sphb {

// ---------------------------------------------------------------------------------------------------
// K5: pressure + artificial-viscosity pair force (SphUtils.cpp:181-239), gravity (:295-304), and --
// fused -- symplectic Euler (:321-335) and wall collision (:343-372).

// One pair (i <- j), in range (d2 < R2), accumulates the direction-weighted scalar; the factor
// Kg = -45/(pi h^6) is applied once after the sum.
__device__ __forceinline__ float pair_weight(const Consts& c, float dx, float dy, float dz, float d2, const float4& vi, const float4& ai,
                                             float mj, const float4& vj, const float4& aj) {
    const float vdotr = fmaf(vi.x - vj.x, dx, fmaf(vi.y - vj.y, dy, (vi.z - vj.z) * dz)); // :220
    const float rinv = pm_rsqrt(d2);
    const float r = d2 * rinv;
    const float s = fmaf(-r, c.inv_h, c.h); // h - q, q = r/h              (:70-74)
    float visc = 0.0f;
    if (vdotr < 0.0f) {                                                                  // :222-227
        const float cbar = 0.5f * (ai.z + aj.z);
        const float mu = c.h * vdotr * pm_rcp(d2 + c.eta_h2);
        visc = 2.0f * mu * fmaf(c.beta, mu, -c.alpha * cbar) * pm_rcp(ai.x + aj.x);
    }
    // :228  -(m_j) * (pi + (P_i + P_j) / (2 rho_j rho_i)); ai.w, aj.w = 1/rho
    const float t = -mj * fmaf(0.5f * (ai.y + aj.y), ai.w * aj.w, visc);
    return t * s * s * rinv; // times (x_i - x_j) = temp * |gradW| * normalize(i - j) / Kg
}

__device__ __forceinline__ void pair_force(const Consts& c, float dx, float dy, float dz, float d2, const float4& vi, const float4& ai,
                                           float mj, const float4& vj, const float4& aj, float& ax, float& ay, float& az) {
    const float wgt = pair_weight(c, dx, dy, dz, d2, vi, ai, mj, vj, aj);
    ax = fmaf(wgt, dx, ax);
    ay = fmaf(wgt, dy, ay);
    az = fmaf(wgt, dz, az);
}

// The neighbour's (P, sqrt|P/rho|, 1/rho) from its rho alone, as finish_density computes them (P with the
// same rounding; the other two through make_aux, i.e. the same bits the density kernel stored).
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

// Resident blocks per SM the main force kernel is compiled for: 9 blocks of 128 threads = 56 registers (measured against 8 blocks /
// 64 registers); the latency-oriented small-scene variant (kIlp = 4) and the 25-row variant take what they need.
#ifndef SPHB_FORCE_MINBLOCKS
#define SPHB_FORCE_MINBLOCKS 9
#endif
#ifndef SPHB_FORCE_WORDS
#define SPHB_FORCE_WORDS 13
#endif
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
// kIlp:   neighbours fetched per iteration.  1 for real workloads (more only cost registers there); 4 for small scenes,
//         where a handful of warps per SM hide nothing and a thread's chain of L2-latency gathers IS the kernel time:
//         four neighbours' records are requested before the first is used.  Same accumulation order, same bits.
template <bool kFused, bool kBBox, bool kAccel, int kRows, bool kDerive, int kIlp, bool kGhosts>
__device__ __forceinline__ void force_thread(const Consts& c, const GridDesc* __restrict__ gd, const int* __restrict__ cell_start,
                                             const float4* __restrict__ posm, const float4* __restrict__ vel, const float4* __restrict__ aux,
                                             const float4* __restrict__ vr, const unsigned int* __restrict__ mask,
                                             float4* __restrict__ accel, float4* __restrict__ posm_out, float4* __restrict__ vel_out,
                                             const int n, const int ms /* stride of the mask rows */, const int i, float lo[3], float hi[3]) {
    // kRows == 9: this thread's NON-EMPTY mask words, compacted, each with the sorted slot of its bit 0 -- one LDS.64 per
    // refill.  kRows == 25 would need 51 KB that way (static shared memory ends at 48 KB): there the words are kept with
    // their index (2*row + half) and the rows' run starts separately.
    constexpr bool kPairs = kRows == 9;
    // Only the first kCapW non-empty words of a thread are staged: every KB of shared memory is a KB less L1 for the
    // gathers this kernel lives on (4 KB more per block cost 8 % in time).  13 words = 13.3 KB per block is the most
    // that keeps the 9 resident blocks (56 registers) inside the 132 KB shared-memory carve-out, i.e. L1 at 124 KB:
    // measured 8 / 10 / 11 / 12 / 13 / 14 / 18 words -> 2.23 / 2.21 / 2.23 / 2.11 / 2.12 / 2.25 / 2.28 ms.  A thread with
    // more non-empty words (rare) reads the rest straight from the mask array after the main walk -- same order.
    constexpr int kCapW = kPairs ? SPHB_FORCE_WORDS : 1;
    __shared__ uint2 s_wb[kCapW][128];
    __shared__ unsigned int s_word[kPairs ? 1 : 2 * kRows][128];
    __shared__ unsigned char s_idx[kPairs ? 1 : 2 * kRows][128];
    __shared__ int s_jb[kPairs ? 1 : kRows][128];
    const int tid = threadIdx.x;
    if (i < n) {
        const GridDesc g = *gd;
        const float4 pi = ldg4(posm + i);
        const float4 vi = ldg4(vel + i);
        if (kFused && kGhosts && vi.w == 0.0f) {
            // a ghost (copy of a neighbouring slab's particle, multi-GPU): it is here to be somebody's neighbour; its own force and
            // new position are its owner's business and the next exchange replaces it anyway.  Pass it through untouched.
            // (kGhosts: only engines that exchange slabs run this variant -- on one GPU the test alone cost 1.5 %.)
            posm_out[i] = pi;
            vel_out[i] = vi;
            if (kAccel) accel[i] = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
            return;
        }
        const float4 ai = ldg4(aux + i);
        int cx, cy, cz;
        cell_key(g, pi.x, pi.y, pi.z, cx, cy, cz);
        bool crowded = false;
        int nw = 0, first_over = 2 * kRows;
        if (kRows == 9) {
            // All 18 mask words and all 18 run bounds are requested before the first is consumed (one round trip to
            // L2 each instead of nine: row_run's early return for rows outside the grid keeps the compiler from
            // hoisting the loads itself).  A row outside the grid reads a clamped row and is emptied afterwards.
            unsigned int mw[18];
            int jbr[9], jer[9];
#pragma unroll
            for (int w = 0; w < 18; w++) mw[w] = __ldg(mask + (size_t)w * ms + i);
            const int x0 = max(cx - g.stencil_x, 0), x1 = min(cx + g.stencil_x, g.dims[0] - 1) + 1;
#pragma unroll
            for (int r = 0; r < 9; r++) {
                const int zc = min(max(cz + r / 3 - 1, 0), g.dims[2] - 1), yc = min(max(cy + r % 3 - 1, 0), g.dims[1] - 1);
                const int row = g.dims[0] * (yc + g.dims[1] * zc);
                jbr[r] = __ldg(cell_start + row + x0);
                jer[r] = __ldg(cell_start + row + x1);
            }
#pragma unroll
            for (int r = 0; r < 9; r++) {
                const int z = cz + r / 3 - 1, y = cy + r % 3 - 1;
                if (z < 0 || z >= g.dims[2] || y < 0 || y >= g.dims[1]) jbr[r] = jer[r] = 0;
                if (mw[2 * r]) {
                    if (nw < kCapW) s_wb[nw][tid] = make_uint2(mw[2 * r], (unsigned int)jbr[r]);
                    else if (nw == kCapW) first_over = 2 * r;
                    nw++;
                }
                if (mw[2 * r + 1]) {
                    if (nw < kCapW) s_wb[nw][tid] = make_uint2(mw[2 * r + 1], (unsigned int)(jbr[r] + 32));
                    else if (nw == kCapW) first_over = 2 * r + 1;
                    nw++;
                }
                crowded |= (jer[r] - jbr[r]) > kMaskBits;
            }
        } else {
#pragma unroll
            for (int r = 0; r < kRows; r++) {
                int jb, je;
                row_run(g, cell_start, cx, cy, cz, r, jb, je);
                const unsigned int m0 = mask[(size_t)(2 * r) * ms + i], m1 = mask[(size_t)(2 * r + 1) * ms + i];
                s_jb[r][tid] = jb;
                if (m0) { s_word[nw][tid] = m0; s_idx[nw][tid] = (unsigned char)(2 * r); nw++; }
                if (m1) { s_word[nw][tid] = m1; s_idx[nw][tid] = (unsigned char)(2 * r + 1); nw++; }
                crowded |= (je - jb) > kMaskBits;
            }
        }
        float ax = 0.0f, ay = 0.0f, az = 0.0f;
        // One neighbour per lane per iteration: a lane whose word ran dry steps to its next (non-empty) word first,
        // so the lanes of a warp stay in lockstep on the pair body until their lists end.
        int k = -1, jb = 0;
        unsigned int m = 0;
        auto visit = [&](int j) {
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
        };
        const int nstaged = kPairs ? min(nw, kCapW) : nw;
        if (kIlp > 1 && kPairs && kDerive) {
            for (;;) {
                int jj[kIlp];
                int got = 0;
#pragma unroll
                for (int u = 0; u < kIlp; u++) {
                    if (m == 0 && k + 1 < nstaged) {
                        const uint2 wb = s_wb[++k][tid];
                        m = wb.x;
                        jb = (int)wb.y;
                    }
                    jj[u] = m ? jb + __ffs((int)m) - 1 : -1;
                    if (m) { m &= m - 1; got++; }
                }
                if (got == 0) break;
                float4 pp[kIlp], vv[kIlp];
#pragma unroll
                for (int u = 0; u < kIlp; u++) {
                    const int j = jj[u] >= 0 ? jj[u] : i;
                    pp[u] = ldg4(posm + j);
                    vv[u] = ldg4(vr + j);
                }
#pragma unroll
                for (int u = 0; u < kIlp; u++) {
                    if (jj[u] < 0) continue;
                    const float dx = pi.x - pp[u].x, dy = pi.y - pp[u].y, dz = pi.z - pp[u].z;
                    const float d2 = fmaf(dx, dx, fmaf(dy, dy, dz * dz));
                    pair_force(c, dx, dy, dz, d2, vi, ai, pp[u].w, vv[u], aux_from_rho(c, vv[u].w), ax, ay, az);
                }
            }
        } else
        for (;;) {
            if (m == 0) {
                if (++k >= nstaged) break;
                if (kPairs) {
                    const uint2 wb = s_wb[k][tid];
                    m = wb.x;
                    jb = (int)wb.y;
                } else {
                    m = s_word[k][tid];
                    const int w = s_idx[k][tid];
                    jb = s_jb[w >> 1][tid] + ((w & 1) << 5);
                }
            }
            const int j = jb + __ffs((int)m) - 1;
            m &= m - 1;
            visit(j);
        }
        if (kPairs && nw > kCapW) { // the words that did not fit the staging area, in the same order
            for (int ww = first_over; ww < 2 * kRows; ww++) {
                unsigned int mo = __ldg(mask + (size_t)ww * ms + i);
                if (!mo) continue;
                int jb2, je2;
                row_run(g, cell_start, cx, cy, cz, ww >> 1, jb2, je2);
                jb2 += (ww & 1) << 5;
                while (mo) {
                    visit(jb2 + __ffs((int)mo) - 1);
                    mo &= mo - 1;
                }
            }
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
            if (kBBox) { // fminf/fmaxf drop NaNs, like the reference's `if (x < min)` comparisons
                lo[0] = fminf(lo[0], p.x); hi[0] = fmaxf(hi[0], p.x);
                lo[1] = fminf(lo[1], p.y); hi[1] = fmaxf(hi[1], p.y);
                lo[2] = fminf(lo[2], p.z); hi[2] = fmaxf(hi[2], p.z);
            }
        }
    }
}

// The kernel: thread i takes sorted particle first + i.  n_dev (may be null) = the live count on the device when the host
// only has an estimate of it (sphb_mg_unpack_async): the main launch (kLoop = false) covers the estimate, and a small second
// launch with kLoop = true, first = threads of the main launch, walks whatever lies beyond in grid strides -- normally
// nothing.  ms = stride of the mask rows (a capacity).  kBBox: every warp leaves the box of its new positions in
// warp_box[global warp id] (8 floats: min xyz, -, max xyz, -); k_bbox_fold joins them at the start of the next sort.  (Folding
// into ONE global box from here -- per block behind a barrier, or per warp -- cost 0.36 / 0.5 ms at 16.7 M: the barrier keeps
// blocks resident for their slowest warp, and half a million warps polling one L2 line queue up behind each other.)
template <bool kFused, bool kBBox, bool kAccel, int kRows, bool kDerive, int kIlp = 1, bool kLoop = false, bool kGhosts = false>
__global__ void __launch_bounds__(128, (kRows == 9 && kIlp == 1) ? SPHB_FORCE_MINBLOCKS : 1) k_force(Consts c, const GridDesc* __restrict__ gd, const int* __restrict__ cell_start,
                                               const float4* __restrict__ posm, const float4* __restrict__ vel, const float4* __restrict__ aux,
                                               const float4* __restrict__ vr, const unsigned int* __restrict__ mask,
                                               float4* __restrict__ accel, float4* __restrict__ posm_out, float4* __restrict__ vel_out,
                                               float4* __restrict__ warp_box, int n, const int* __restrict__ n_dev, int ms, int first) {
    if (n_dev) n = *n_dev;
    float lo[3] = {INFINITY, INFINITY, INFINITY}, hi[3] = {-INFINITY, -INFINITY, -INFINITY};
    if (!kLoop) {
        force_thread<kFused, kBBox, kAccel, kRows, kDerive, kIlp, kGhosts>(c, gd, cell_start, posm, vel, aux, vr, mask, accel, posm_out, vel_out, n, ms,
                                                                  first + blockIdx.x * blockDim.x + threadIdx.x, lo, hi);
    } else {
        for (long long base = first + (long long)blockIdx.x * blockDim.x; base < n; base += (long long)gridDim.x * blockDim.x)
            force_thread<kFused, kBBox, kAccel, kRows, kDerive, kIlp, kGhosts>(c, gd, cell_start, posm, vel, aux, vr, mask, accel, posm_out, vel_out, n, ms,
                                                                      (int)base + (int)threadIdx.x, lo, hi);
    }
    if (kFused && kBBox) {
#pragma unroll
        for (int a = 0; a < 3; a++) {
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) {
                lo[a] = fminf(lo[a], __shfl_xor_sync(0xffffffffu, lo[a], o));
                hi[a] = fmaxf(hi[a], __shfl_xor_sync(0xffffffffu, hi[a], o));
            }
        }
        if ((threadIdx.x & 31) == 0) {
            float4* w = warp_box + 2 * (size_t)((first / 32) + blockIdx.x * (blockDim.x / 32) + (threadIdx.x >> 5));
            w[0] = make_float4(lo[0], lo[1], lo[2], 0.0f);
            w[1] = make_float4(hi[0], hi[1], hi[2], 0.0f);
        }
    }
}

// Joins the per-warp boxes the fused force kernel left (count records of 8 floats) into the global accumulator.
__global__ void __launch_bounds__(256) k_bbox_fold(const float4* __restrict__ warp_box, int count, BBoxAcc* acc) {
    float lo[3] = {INFINITY, INFINITY, INFINITY}, hi[3] = {-INFINITY, -INFINITY, -INFINITY};
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < count; i += gridDim.x * blockDim.x) {
        const float4 l = warp_box[2 * (size_t)i], h = warp_box[2 * (size_t)i + 1];
        lo[0] = fminf(lo[0], l.x); lo[1] = fminf(lo[1], l.y); lo[2] = fminf(lo[2], l.z);
        hi[0] = fmaxf(hi[0], h.x); hi[1] = fmaxf(hi[1], h.y); hi[2] = fmaxf(hi[2], h.z);
    }
    bbox_block_commit(lo, hi, acc);
}

// Optional: the reference's dead viscosity_forces (SphUtils.cpp:244-290; its call at :175 is commented out), on
// the full neighbourhood:  accel_i += coef * m_j / (rho_i rho_j) * lapW_ij * (v_i - v_j),  lapW = 45/(pi h^6) (h - q)
// (:84-95).  A separate pass over the neighbour masks that adds into accel; row by row, off the hot path.
__global__ void __launch_bounds__(128) k_lap_viscosity(Consts c, const GridDesc* __restrict__ gd, const int* __restrict__ cell_start,
                                                       const float4* __restrict__ posm, const float4* __restrict__ vel, const float4* __restrict__ aux,
                                                       const unsigned int* __restrict__ mask, float4* __restrict__ accel, int n, int ms) {
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
            unsigned int m = mask[(size_t)(2 * r + w) * ms + i];
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

} // namespace sphb
