// Headless point-sprite renderer (SPH/Shaders/sph.vertexshader, sph.fragmentshader) and small utility kernels (non-finite census, iota).
// Part of libsphb200.so (sm_100a); included through sphb_device.cuh.
#pragma once
#include "sphb_common.cuh"

namespace sphb {

// ---------------------------------------------------------------------------------------------------
// Headless point-sprite renderer: the arithmetic of SPH/Shaders/sph.vertexshader + sph.fragmentshader and of GL's
// point rasterisation, into a float image.  Two passes because the viewer has no depth buffer: the LAST particle
// (in original order) that covers a pixel wins.  Pass 1: one warp per particle walks its sprite square and
// atomicMax-es (particle index + 1) into the pixels it really covers; pass 2: one thread per pixel shades the winner.

struct Camera {
    float m[16]; // column-major
    int width, height;
    float point_scale, sprite_radius, clear_value;
    int depth_test; // 1: the sprite with the smallest point depth wins a pixel (GL_LESS, first drawn on a tie); 0: the last drawn
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

// depth-tested variant: the key (point depth, particle index) is minimised, so the nearest sprite wins and the lowest index on a tie
__global__ void __launch_bounds__(256) k_splat_winner_depth(Camera c, const float4* __restrict__ posm, const int* __restrict__ inv, int n,
                                                            unsigned long long* __restrict__ winner) {
    const int o = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    if (o >= n) return;
    const float4 p = posm[inv[o]];
    float xw, yw, edge;
    if (!sprite_of(c, p, xw, yw, edge)) return;
    const float4 clip = mvp_mul(c, p.x, p.y, p.z);
    const float zwin = clip.z / clip.w * 0.5f + 0.5f;
    if (!(zwin < 1.0f)) return; // GL_LESS against the cleared depth 1
    const unsigned long long key = ((unsigned long long)f2ord(zwin) << 32) | (unsigned int)o;
    const int x0 = max(0, (int)floorf(xw - 0.5f * edge)), x1 = min(c.width - 1, (int)ceilf(xw + 0.5f * edge));
    const int y0 = max(0, (int)floorf(yw - 0.5f * edge)), y1 = min(c.height - 1, (int)ceilf(yw + 0.5f * edge));
    const int wx = x1 - x0 + 1, wy = y1 - y0 + 1;
    if (wx <= 0 || wy <= 0) return;
    for (int k = lane; k < wx * wy; k += 32) {
        const int px = x0 + k % wx, py = y0 + k / wx;
        float d;
        if (sprite_fragment(c, p, xw, yw, edge, px, py, d)) atomicMin(winner + (size_t)py * c.width + px, key);
    }
}

__global__ void __launch_bounds__(256) k_splat_shade_depth(Camera c, const float4* __restrict__ posm, const int* __restrict__ inv,
                                                           const unsigned long long* __restrict__ winner, float* __restrict__ image) {
    const int pix = blockIdx.x * blockDim.x + threadIdx.x;
    if (pix >= c.width * c.height) return;
    const unsigned long long w = winner[pix];
    float v = c.clear_value;
    if (w != ~0ull) {
        const float4 p = posm[inv[(unsigned int)w]];
        float xw, yw, edge, d;
        if (sprite_of(c, p, xw, yw, edge) && sprite_fragment(c, p, xw, yw, edge, pix % c.width, pix / c.width, d)) v = d;
    }
    image[pix] = v;
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

// device-to-device upload of velocities: the fourth component is the engine's ownership flag, 1 for everything a caller uploads
__global__ void __launch_bounds__(256) k_copy_vel_owned(const float4* __restrict__ src, float4* __restrict__ dst, int n) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float4 v = src[i];
    dst[i] = make_float4(v.x, v.y, v.z, 1.0f);
}

// sphb_set_ids: replace each record's upload index by the caller's label for that index
__global__ void __launch_bounds__(256) k_relabel(int* __restrict__ oid, const int* __restrict__ labels, int n) {
    int p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p < n) oid[p] = labels[oid[p]];
}

// Caller-supplied particle ids must be a permutation of 0..n-1: the downloads invert them (k_invert_perm) and the
// stable rank orders cells by them.  seen[] must be filled with -1; bad[0] counts ids out of range or used twice.
__global__ void __launch_bounds__(256) k_check_ids(const int* __restrict__ ids, int n, int* __restrict__ seen, unsigned int* __restrict__ bad) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const int id = ids[i];
    if (id < 0 || id >= n || atomicExch(seen + id, i) != -1) atomicAdd(bad, 1u);
}

} // namespace sphb
