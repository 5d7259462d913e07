// Parity mode (sphb_params.parity_mode): density and forces evaluated the way the reference's own loops evaluate them -- neighbours
// in ascending ORIGINAL index, every float operation of SPH/SphUtils.cpp:55-79, 100-149, 181-239 and of GLM's dot / length /
// normalize (deps/include/glm/detail/func_geometric.inl:64-158) with its own rounding and no fused multiply-add, W and gradW
// through double like pow(float, int) makes them -- so that rho, P and accel come out BIT FOR BIT as the
// reference's.  A debugging and refereeing mode: one thread per particle re-scans its ~220 candidates once per neighbour to find
// the next original index (no list, no capacity limit), i.e. it is ~50x slower than the fast kernels and never on the timed path.
// Part of libsphb200.so (sm_100a); included through sphb_device.cuh.
#pragma once
#include "sphb_common.cuh"
#include "sphb_density.cuh"

namespace sphb {

struct ParityConsts {
    double den_w; // (double)(float)(64 * pi_f) * pow((double)h, 9)     SphUtils.cpp:61, computed on the host with the C library's pow
    double den_g; // (double)pi_f * pow((double)h, 6)                    SphUtils.cpp:74
    float eos_k, alpha, beta, eta;
};

// glm::dot: tmp = a * b componentwise; tmp.x + tmp.y + tmp.z, left to right
__device__ __forceinline__ float p_dot(float ax, float ay, float az, float bx, float by, float bz) {
    return __fadd_rn(__fadd_rn(__fmul_rn(ax, bx), __fmul_rn(ay, by)), __fmul_rn(az, bz));
}

// kernel_function (SphUtils.cpp:55-63): distance(i, j) = length(j - i)
__device__ __forceinline__ float p_kernel_function(const Consts& c, const ParityConsts& pc, const float4& pi, const float4& pj) {
    const float dx = __fsub_rn(pj.x, pi.x), dy = __fsub_rn(pj.y, pi.y), dz = __fsub_rn(pj.z, pi.z);
    const float q = __fdiv_rn(__fsqrt_rn(p_dot(dx, dy, dz, dx, dy, dz)), c.h);
    float result = 0.0f;
    if (q <= c.h) {
        const float t = __fsub_rn(__fmul_rn(c.h, c.h), __fmul_rn(q, q));
        const double td = (double)t;
        const double t3 = __dmul_rn(__dmul_rn(td, td), td); // t*t is exact in double (24-bit factors), so this is the correctly rounded cube, as pow(t, 3) returns
        result = (float)__ddiv_rn(__dmul_rn(315.0, t3), pc.den_w);
    }
    return result;
}

// the next neighbour candidate after original index `last`: smallest original index > last among the stencil's candidates that lie
// within a hair of the support (pairs outside contribute exactly +-0 in the reference's sums and are skipped, like the oracle's
// cell-list loops do); returns its sorted slot or -1
__device__ __forceinline__ int p_next(const Consts& c, const GridDesc& g, const int* __restrict__ cell_start, const float4* __restrict__ posm,
                                      const int* __restrict__ oid, const float4& pi, int cx, int cy, int cz, int nrows, int last, int& best_id) {
    const float lim = c.R2 * 1.001f;
    int best = 0x7fffffff, bj = -1;
    for (int r = 0; r < nrows; r++) {
        int jb, je;
        row_run(g, cell_start, cx, cy, cz, r, jb, je);
        for (int j = jb; j < je; j++) {
            const int o = __ldg(oid + j);
            if (o <= last || o >= best) continue;
            const float4 pj = ldg4(posm + j);
            const float dx = pj.x - pi.x, dy = pj.y - pi.y, dz = pj.z - pi.z;
            const float d2 = dx * dx + dy * dy + dz * dz;
            if (!(d2 > lim)) { best = o; bj = j; } // keeps NaN distances too
        }
    }
    best_id = best;
    return bj;
}

__global__ void __launch_bounds__(128) k_density_parity(Consts c, ParityConsts pc, const GridDesc* __restrict__ gd, const int* __restrict__ cell_start,
                                                        const float4* __restrict__ posm, const int* __restrict__ oid, float4* __restrict__ aux, int n) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const GridDesc g = *gd;
    const float4 pi = ldg4(posm + i);
    const int me = oid[i];
    int cx, cy, cz;
    cell_key(g, pi.x, pi.y, pi.z, cx, cy, cz);
    const int nrows = (2 * g.stencil + 1) * (2 * g.stencil + 1);
    float density = 0.0f;
    int last = -1;
    for (;;) {
        int id;
        const int j = p_next(c, g, cell_start, posm, oid, pi, cx, cy, cz, nrows, last, id);
        if (j < 0) break;
        last = id;
        if (id == me) continue; // SphUtils.cpp:132: the self term is excluded by index
        const float4 pj = ldg4(posm + j);
        density = __fadd_rn(density, __fmul_rn(pj.w, p_kernel_function(c, pc, pi, pj))); // :133
    }
    if (density < c.rho0) density = c.rho0;                                    // :141-142
    const float P = __fmul_rn(pc.eos_k, __fsub_rn(density, c.rho0));           // :144
    aux[i] = make_aux(density, P);
}

__global__ void __launch_bounds__(128) k_force_parity(Consts c, ParityConsts pc, const GridDesc* __restrict__ gd, const int* __restrict__ cell_start,
                                                      const float4* __restrict__ posm, const float4* __restrict__ vel, const float4* __restrict__ aux,
                                                      const int* __restrict__ oid, float4* __restrict__ accel, int n) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const GridDesc g = *gd;
    const float4 pi = ldg4(posm + i), vi = ldg4(vel + i), ai = ldg4(aux + i);
    const int me = oid[i];
    int cx, cy, cz;
    cell_key(g, pi.x, pi.y, pi.z, cx, cy, cz);
    const int nrows = (2 * g.stencil + 1) * (2 * g.stencil + 1);
    float ax = __fadd_rn(0.0f, c.gx), ay = __fadd_rn(0.0f, c.gy), az = __fadd_rn(0.0f, c.gz); // accel = 0, then += gravity (:159-168, 295-304)
    const float h = c.h;
    int last = -1;
    for (;;) {
        int id;
        const int j = p_next(c, g, cell_start, posm, oid, pi, cx, cy, cz, nrows, last, id);
        if (j < 0) break;
        last = id;
        if (id == me) continue; // :216
        const float4 pj = ldg4(posm + j), vj = ldg4(vel + j), aj = ldg4(aux + j);
        // pressure_forces pair body, :219-230
        const float dvx = __fsub_rn(vi.x, vj.x), dvy = __fsub_rn(vi.y, vj.y), dvz = __fsub_rn(vi.z, vj.z);
        const float dx = __fsub_rn(pi.x, pj.x), dy = __fsub_rn(pi.y, pj.y), dz = __fsub_rn(pi.z, pj.z);
        const float vdotr = p_dot(dvx, dvy, dvz, dx, dy, dz);
        float pi_visc = 0.0f;
        if (vdotr < 0) {
            const float ci = __fsqrt_rn(__fmul_rn(1.0f, fabsf(__fdiv_rn(ai.y, ai.x)))), cj = __fsqrt_rn(__fmul_rn(1.0f, fabsf(__fdiv_rn(aj.y, aj.x))));
            const float cbar = __fdiv_rn(__fadd_rn(ci, cj), 2.0f);
            const float r = __fsqrt_rn(p_dot(dx, dy, dz, dx, dy, dz));
            const float mu = __fdiv_rn(__fmul_rn(h, vdotr), __fadd_rn(__fmul_rn(r, r), __fmul_rn(__fmul_rn(pc.eta, h), h)));
            const float num = __fadd_rn(__fmul_rn(__fmul_rn(-pc.alpha, cbar), mu), __fmul_rn(__fmul_rn(pc.beta, mu), mu));
            pi_visc = __fdiv_rn(__fmul_rn(2.0f, num), __fadd_rn(ai.x, aj.x));
        }
        const float temp = __fmul_rn(-pj.w, __fadd_rn(pi_visc, __fdiv_rn(__fadd_rn(ai.y, aj.y), __fmul_rn(__fmul_rn(2.0f, aj.x), ai.x))));
        // grad_kernel, :68-79: q = distance(i, j) / h with distance = length(j - i); normalize(i - j) * result
        const float ex = __fsub_rn(pj.x, pi.x), ey = __fsub_rn(pj.y, pi.y), ez = __fsub_rn(pj.z, pi.z);
        const float q = __fdiv_rn(__fsqrt_rn(p_dot(ex, ey, ez, ex, ey, ez)), h);
        float result = 0.0f;
        if (q < h) {
            const float hq = __fsub_rn(h, q);
            const float num = __fmul_rn(__fmul_rn(-45.0f, hq), hq);
            result = (float)__ddiv_rn((double)num, pc.den_g);
        }
        const float inv = __fdiv_rn(1.0f, __fsqrt_rn(p_dot(dx, dy, dz, dx, dy, dz)));
        const float gx = __fmul_rn(__fmul_rn(dx, inv), result), gy = __fmul_rn(__fmul_rn(dy, inv), result), gz = __fmul_rn(__fmul_rn(dz, inv), result);
        ax = __fadd_rn(ax, __fmul_rn(temp, gx));
        ay = __fadd_rn(ay, __fmul_rn(temp, gy));
        az = __fadd_rn(az, __fmul_rn(temp, gz));
    }
    accel[i] = make_float4(ax, ay, az, 0.0f);
}

} // namespace sphb
