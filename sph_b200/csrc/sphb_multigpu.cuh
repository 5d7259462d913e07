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
// Ghosts of the previous step (vel.w == 0) are dropped: they are re-sent every step.  Kept particles STAY WHERE THEY
// ARE (only a changed ownership flag or a wrapped coordinate is written back); dropped ones are marked dead (mass
// -1, flag 0) and leave the arrays in the next step's counting sort, which files them behind the last cell
// (kTrashBins) -- no compaction pass over the state.  The records that do move -- the two messages and the periodic
// y/z images appended behind the current particles -- are compacted in input order with a two-level scan
// (per-block counts -> scan -> write), so the result is deterministic.  A message is a header record {count,
// capacity, 0...} followed by 32-byte particle records {posm float4, (vx, vy, vz, particle id as bits)}; the receiver appends
// them, files the id in its id array and sets the ownership flag itself.

struct Slab {
    float lo, hi, halo;
    int has_left, has_right;
    float shift_left, shift_right; // added to x of a record sent that way (periodic seam), else 0
    int per_y, per_z;              // periodic in y / z over [ylo, ylo + ylen) / [zlo, zlo + zlen)
    float ylo, ylen, zlo, zlen;
};

constexpr int kMgThreads = 256;
constexpr int kMgPer = 4;                      // consecutive particles per thread (more loads in flight, a quarter of the barriers)
constexpr int kMgTile = kMgThreads * kMgPer;   // particles per block

// Periodic y / z are handled HERE and nowhere else: an owned particle is wrapped back into the box, and if it lies
// within `halo` of a periodic face it is emitted once more per face (and once for the edge) shifted by the period,
// as a ghost image.  The density / force kernels never see a periodic boundary -- only more ghosts.
struct MgItem {
    float4 p, v;      // wrapped position+mass, velocity (+flag)
    int nimg;         // 1, 2 or 4 records: (0,0), (dy,0), (0,dz), (dy,dz)
    float dy, dz;     // image shifts (0 if not near that face)
    bool keep, left, right;
    bool dies;        // alive so far and not kept: to be marked dead
    bool dead;        // dead after this pack (dies, or marked by an earlier pack)
};

__device__ __forceinline__ float mg_wrap(float x, float lo, float len) {
    x = x - len * floorf((x - lo) / len);
    return x >= lo + len ? lo : x; // rounding can land exactly on the upper face
}

__device__ __forceinline__ MgItem mg_classify(const Slab& s, float4 p, const float4& v) {
    MgItem it;
    const bool alive = !(p.w < 0.0f);      // not already marked dead by an earlier pack
    const bool owned = alive && v.w != 0.0f;
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
    it.dies = alive && !it.keep;
    it.dead = !it.keep;
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

// The four record counts of a thread (each <= 4 per particle, block sums <= 4096) travel as 16-bit fields of one 64-bit word, so
// one block scan serves all four streams (the first version ran one 32-bit block scan per stream: k_mg_count and
// k_mg_write were bound by those barriers, not by memory).
constexpr int kMgStreams = 4; // periodic images of kept particles, to left, to right, dead

__device__ __forceinline__ unsigned long long mg_pack4(int a, int b, int c, int d) {
    return (unsigned long long)a | ((unsigned long long)b << 16) | ((unsigned long long)c << 32) | ((unsigned long long)d << 48);
}
__device__ __forceinline__ int mg_field(unsigned long long v, int k) { return (int)((v >> (16 * k)) & 0xffffull); }

// exclusive prefix of v over the block of kMgThreads threads; *total = block sum.  Two barriers.
__device__ __forceinline__ unsigned long long mg_block_scan(unsigned long long v, unsigned long long* total) {
    __shared__ unsigned long long s_warp[kMgThreads / 32];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    unsigned long long inc = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const unsigned long long t = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc += t;
    }
    if (lane == 31) s_warp[warp] = inc;
    __syncthreads();
    unsigned long long off = 0, tot = 0;
#pragma unroll
    for (int w = 0; w < kMgThreads / 32; w++) { // 8 warps: every thread sums the broadcast partials itself
        const unsigned long long x = s_warp[w];
        if (w < warp) off += x;
        tot += x;
    }
    *total = tot;
    return off + inc - v;
}

// pass 1: per-tile record counts, one array per stream: block_counts[k * nb_layout + tile].  n_dev (may be null) = the particle
// count on the device when the host only has an estimate (sphb_mg_unpack_async): blocks then walk the tiles in grid strides;
// nb_layout = tiles the arrays were laid out for (from the capacity).
__global__ void __launch_bounds__(kMgThreads) k_mg_count(Slab s, const float4* __restrict__ posm, const float4* __restrict__ vel, int n,
                                                         const int* __restrict__ n_dev, int nb_layout,
                                                         int* __restrict__ block_counts, unsigned char* __restrict__ touch) {
    if (n_dev) n = *n_dev;
    const int ntiles = (n + kMgTile - 1) / kMgTile;
    __shared__ unsigned long long s_part[kMgThreads / 32];
    for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
        const int i0 = tile * kMgTile + threadIdx.x * kMgPer;
        unsigned long long c = 0;
        bool work = false; // something for pass 3 to do here: records to emit, a flag / wrap to write back, a particle to mark dead
        float4 p[kMgPer], v[kMgPer];
#pragma unroll
        for (int q = 0; q < kMgPer; q++) {
            const int i = max(min(i0 + q, n - 1), 0);
            p[q] = ldg4(posm + i);
            v[q] = ldg4(vel + i);
        }
#pragma unroll
        for (int q = 0; q < kMgPer; q++) {
            if (i0 + q >= n) break;
            const MgItem it = mg_classify(s, p[q], v[q]);
            c += mg_pack4(it.keep ? it.nimg - 1 : 0, it.left ? it.nimg : 0, it.right ? it.nimg : 0, it.dead ? 1 : 0);
            work |= it.left || it.right || it.dies ||
                    (it.keep && (it.nimg > 1 || it.p.y != p[q].y || it.p.z != p[q].z || (mg_inside(s, it.p) ? 1.0f : 0.0f) != v[q].w));
        }
        // one byte per warp (128 consecutive particles): pass 3 skips the warps with nothing to do -- away from the slab
        // faces that is most of them -- without reading their particles again
        const bool any = __any_sync(0xffffffffu, work);
        if ((threadIdx.x & 31) == 0) touch[tile * (kMgThreads / 32) + (threadIdx.x >> 5)] = any ? 1 : 0;
        // block sum only: warp reduction, then one shared-memory round
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) c += __shfl_xor_sync(0xffffffffu, c, o);
        if ((threadIdx.x & 31) == 0) s_part[threadIdx.x >> 5] = c;
        __syncthreads();
        if (threadIdx.x < kMgStreams) {
            unsigned long long tot = 0;
#pragma unroll
            for (int w = 0; w < kMgThreads / 32; w++) tot += s_part[w];
            block_counts[threadIdx.x * nb_layout + tile] = mg_field(tot, threadIdx.x);
        }
        __syncthreads(); // s_part is reused by the next tile
    }
}

// pass 2: the per-block counts become exclusive offsets (and totals).  One block of 1024 threads per stream
// (images / left / right / dead), four entries per thread per round: 70 K blocks take 18 rounds (the first version, one
// 256-thread block for all three streams, took 0.34 ms of a 7.4 ms step at 17.8 M particles).
constexpr int kMgScanThreads = 1024;

__global__ void __launch_bounds__(kMgScanThreads) k_mg_scan(int* __restrict__ block_counts, int nb_layout, int n, const int* __restrict__ n_dev,
                                                           int* __restrict__ totals) {
    __shared__ int s_warp[kMgScanThreads / 32];
    if (n_dev) n = *n_dev;
    const int nblocks = (n + kMgTile - 1) / kMgTile;
    const int k = blockIdx.x; // stream
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    int carry = 0;
    for (int base = 0; base < nblocks; base += 4 * kMgScanThreads) {
        const int b0 = base + 4 * threadIdx.x;
        int v[4];
        int tsum = 0;
#pragma unroll
        for (int q = 0; q < 4; q++) {
            v[q] = (b0 + q < nblocks) ? block_counts[k * nb_layout + b0 + q] : 0;
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
            if (b0 + q < nblocks) block_counts[k * nb_layout + b0 + q] = ex;
            ex += v[q];
        }
        carry += round_total;
        __syncthreads(); // s_warp is reused next round
    }
    // totals: [0] = records in the arrays after the local images are appended (where received records go),
    // [1], [2] = message sizes, [3] = dead records among the first n
    if (threadIdx.x == 0) totals[k] = k == 0 ? carry + n : carry;
}

// pass 3: write.  A kept particle stays in its slot (flag refreshed, wrapped coordinates written back); its periodic
// images go behind the current particles (slot n + offset); a dropped one is marked dead; left/right -> messages.
// posm/vel are read and written by the thread that owns slot i only.
__global__ void __launch_bounds__(kMgThreads) k_mg_write(Slab s, float4* __restrict__ posm, float4* __restrict__ vel, int* __restrict__ oid, int n,
                                                         const int* __restrict__ n_dev,
                                                         int nb_layout, const int* __restrict__ block_offsets, const int* __restrict__ totals,
                                                         int out_capacity, const unsigned char* __restrict__ touch,
                                                         float4* __restrict__ msg_left, float4* __restrict__ msg_right, int cap_left, int cap_right) {
    if (n_dev) n = *n_dev;
    const int ntiles = (n + kMgTile - 1) / kMgTile;
    for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
        const int i0 = tile * kMgTile + threadIdx.x * kMgPer;
        // a tile none of whose eight warps has anything to do (away from the slab faces: nine tiles in ten) is left at once,
        // before the block scan and its barriers
        static_assert(kMgThreads / 32 == 8, "one 8-byte load covers a tile's touch flags");
        if (*reinterpret_cast<const unsigned long long*>(touch + (size_t)tile * 8) == 0ull) continue;
        const bool active = touch[tile * (kMgThreads / 32) + (threadIdx.x >> 5)] != 0; // warp-uniform
        // phase 1: this thread's record counts (nothing is kept in registers: phase 2 reads the four particles again,
        // from L1 -- holding them cost 107 registers and three quarters of the occupancy)
        unsigned long long mine = 0;
        if (active) {
            float4 pp[kMgPer], vv[kMgPer];
#pragma unroll
            for (int q = 0; q < kMgPer; q++) {
                const int i = max(min(i0 + q, n - 1), 0);
                pp[q] = posm[i];
                vv[q] = vel[i];
            }
#pragma unroll
            for (int q = 0; q < kMgPer; q++) {
                if (i0 + q >= n) break;
                const MgItem it = mg_classify(s, pp[q], vv[q]);
                mine += mg_pack4(it.keep ? it.nimg - 1 : 0, it.left ? it.nimg : 0, it.right ? it.nimg : 0, 0);
            }
        }
        unsigned long long tot;
        const unsigned long long ex = mg_block_scan(mine, &tot);
        int off[3];
#pragma unroll
        for (int k = 0; k < 3; k++) off[k] = mg_field(ex, k) + block_offsets[k * nb_layout + tile];
#pragma unroll 1
        for (int q = 0; q < kMgPer; q++) {
            const int i = i0 + q;
            if (i >= n || !active) break;
            const float4 p_q = posm[i];
            const MgItem it = mg_classify(s, p_q, vel[i]);
            const int cnt[3] = {it.keep ? it.nimg - 1 : 0, it.left ? it.nimg : 0, it.right ? it.nimg : 0};
            // the particle's id travels with it: in a message it rides in the velocity record's fourth word (the ownership
            // flag there means nothing to the receiver, who decides ownership by position), locally it stays in the id array
            const int id = (oid && (cnt[0] | cnt[1] | cnt[2])) ? oid[i] : 0; // oid == null: ids are not tracked across exchanges
            const float4 v_msg = make_float4(it.v.x, it.v.y, it.v.z, __int_as_float(id));
            if (it.keep) {
                const float flag = mg_inside(s, it.p) ? 1.0f : 0.0f;
                if (flag != it.v.w) vel[i].w = flag;
                if (it.p.y != p_q.y || it.p.z != p_q.z) posm[i] = it.p; // wrapped into the periodic box
            } else if (it.dies) {
                posm[i].w = -1.0f;
                vel[i].w = 0.0f;
            }
            for (int img = 0; img < cnt[0]; img++) {
                const int dst = n + off[0] + img;
                if (dst >= out_capacity) break;
                posm[dst] = mg_image(it, img + 1, 0.0f);
                vel[dst] = make_float4(it.v.x, it.v.y, it.v.z, 0.0f);
                if (oid) oid[dst] = id;
            }
            for (int img = 0; img < cnt[1]; img++) {
                if (off[1] + img >= cap_left) break;
                float4 w = mg_image(it, img, s.shift_left);
                // a migrant through the periodic seam must land INSIDE the far slab even if x + period rounds up onto its face
                if (s.shift_left != 0.0f && it.p.x < s.lo) w.x = fminf(w.x, nextafterf(s.lo + s.shift_left, -INFINITY));
                msg_left[2 + 2 * (off[1] + img)] = w;
                msg_left[3 + 2 * (off[1] + img)] = v_msg;
            }
            for (int img = 0; img < cnt[2]; img++) {
                if (off[2] + img >= cap_right) break;
                float4 w = mg_image(it, img, s.shift_right);
                if (s.shift_right != 0.0f && it.p.x >= s.hi) w.x = fmaxf(w.x, s.hi + s.shift_right);
                msg_right[2 + 2 * (off[2] + img)] = w;
                msg_right[3 + 2 * (off[2] + img)] = v_msg;
            }
            off[0] += cnt[0]; off[1] += cnt[1]; off[2] += cnt[2];
        }
        __syncthreads(); // the block scan's shared words are reused by the next tile
    }
    if (blockIdx.x == 0 && threadIdx.x == 0) { // headers: count, capacity
        if (msg_left) msg_left[0] = make_float4(__int_as_float(s.has_left ? totals[1] : 0), __int_as_float(cap_left), 0.0f, 0.0f);
        if (msg_right) msg_right[0] = make_float4(__int_as_float(s.has_right ? totals[2] : 0), __int_as_float(cap_right), 0.0f, 0.0f);
    }
}

// append a received message behind the kept particles; counts[0] = kept, counts[1] = records already appended
__global__ void __launch_bounds__(256) k_mg_append(Slab s, const float4* __restrict__ msg, const int* __restrict__ totals, int which,
                                                   float4* __restrict__ posm_out, float4* __restrict__ vel_out, int* __restrict__ oid_out,
                                                   int out_capacity, int* __restrict__ recv_counts) {
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
        if (oid_out) oid_out[base + r] = __float_as_int(v.w); // the sender's id for this particle (see k_mg_write)
    }
}

// ---------------------------------------------------------------------------------------------------
// Peer-to-peer halo link (sphb_mg_link): messages are written by k_mg_write straight into the NEIGHBOUR's receive buffer over
// NVLink, so the only other traffic is a handful of 4-byte flags.  A flag lives in the memory of the GPU that waits on it and is
// set remotely by the GPU it waits for; values only grow (the step number + 1), so a late or repeated set is harmless.
constexpr int kLinkArrivedLeft = 0, kLinkArrivedRight = 1, kLinkConsumedLeft = 2, kLinkConsumedRight = 3, kLinkError = 4;
constexpr int kLinkFlagInts = 64; // 256-byte flag block in front of the receive buffers

// release: everything this stream did before (the pack kernel's remote stores) is visible before the flag is
__global__ void k_link_set(int* f0, int v0, int* f1, int v1) {
    if (threadIdx.x || blockIdx.x) return;
    __threadfence_system();
    if (f0) *((volatile int*)f0) = v0;
    if (f1) *((volatile int*)f1) = v1;
}

// acquire: spin (one thread) until both flags have reached their values; gives up after ~4 s of GPU time and raises the
// link's error flag instead of hanging the device if a neighbour died
__global__ void k_link_wait(const int* f0, int v0, const int* f1, int v1, int* err) {
    if (threadIdx.x || blockIdx.x) return;
    unsigned long long t0;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
    for (;;) {
        const bool ok0 = !f0 || *((volatile const int*)f0) >= v0;
        const bool ok1 = !f1 || *((volatile const int*)f1) >= v1;
        if (ok0 && ok1) break;
        unsigned long long t;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
        if (t - t0 > 4000000000ull) { *((volatile int*)err) = 1; break; }
        __nanosleep(200);
    }
    __threadfence_system();
}

// After both appends: publish the new record count for the kernels of the next step (counts[0], read by the sort when the host
// does not wait for it) and a status record the host fetches WITHOUT waiting for it (sphb_mg_unpack_async / sphb_mg_poll):
// {records before the messages, from left, from right, total, dead among them, sent left, sent right, flags}
// flags: 1 = the left neighbour overflowed its message, 2 = the right one did, 4 = the arrays are full, 8 = a link wait timed out.
__global__ void k_mg_commit(const int* __restrict__ totals, const int* __restrict__ recv, int capacity, int* __restrict__ counts, int* __restrict__ status,
                            const int* __restrict__ link_err) {
    if (threadIdx.x || blockIdx.x) return;
    const long long kept = totals[0], cl = recv[0], cr = recv[1], total = kept + cl + cr;
    int flags = 0;
    if (link_err && *((volatile const int*)link_err)) flags |= 8; // a wait on the peer-to-peer link timed out
    if (recv[2] > recv[0]) flags |= 1;
    if (recv[3] > recv[1]) flags |= 2;
    if (total > capacity) flags |= 4;
    counts[0] = (int)(total > capacity ? capacity : total);
    status[0] = (int)kept; status[1] = (int)cl; status[2] = (int)cr; status[3] = counts[0];
    status[4] = totals[3]; status[5] = totals[1]; status[6] = totals[2]; status[7] = flags;
}

} // namespace sphb
