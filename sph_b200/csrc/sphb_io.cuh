// Host exchange: AoS (the reference 52-byte Particle) <-> SoA conversion, inverse permutation, packed positions, grid diagnostics.
// Part of libsphb200.so (sm_100a); included through sphb_device.cuh.
#pragma once
#include "sphb_common.cuh"
#include "sphb_grid.cuh"

namespace sphb {

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

} // namespace sphb
