// Rigid-sphere coupling (SPH/SphUtils.cpp:163-166, 300-303, 331-334, 375-385).
// Part of libsphb200.so (sm_100a); included through sphb_device.cuh.
#pragma once
#include "sphb_common.cuh"

namespace sphb {

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

} // namespace sphb
