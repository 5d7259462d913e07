// Multi-GPU x-slab decomposition: classify / compact / message kernels (no counterpart in the reference).
// Part of libsphb200.so (sm_100a); included through sphb_device.cuh.
#pragma once
#include "sphb_common.cuh"
#include "sphb_grid.cuh"

namespace sphb {

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
