// Host side of libsphb200.so: the C ABI declared in include/sphb200.h over the kernels of
// sphb_device.cuh.  One engine = one GPU = one stream.  No CPU compute path exists: every entry
// point either launches CUDA work or fails with an error code.
#include "../../include/sphb200.h"
#include "sphb_device.cuh"

#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <new>

using namespace sphb;

#ifndef SPHB_BBOX_FUSED
#define SPHB_BBOX_FUSED 0 /* 1: track the bounding box of the new positions inside the fused force kernel (measured slower, see sphb_step) */
#endif

namespace {

thread_local char g_err[512] = "";

int fail(int code, const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof g_err, fmt, ap);
    va_end(ap);
    return code;
}

#define CK(call)                                                                                              \
    do {                                                                                                      \
        cudaError_t _e = (call);                                                                              \
        if (_e != cudaSuccess)                                                                                \
            return fail(SPHB_ERR_CUDA, "%s failed at %s:%d: %s", #call, __FILE__, __LINE__, cudaGetErrorString(_e)); \
    } while (0)

inline int cdiv(long long a, int b) { return (int)((a + b - 1) / b); }

template <class T>
struct DevBuf {
    T* p = nullptr;
    size_t cap = 0; // elements
    cudaError_t reserve(size_t n) {
        if (n <= cap) return cudaSuccess;
        if (p) cudaFree(p);
        p = nullptr;
        cap = 0;
        cudaError_t e = cudaMalloc((void**)&p, n * sizeof(T));
        if (e == cudaSuccess) cap = n;
        return e;
    }
    void release() {
        if (p) cudaFree(p);
        p = nullptr;
        cap = 0;
    }
};

} // namespace

struct sphb_engine {
    int device = 0;
    cudaStream_t stream = nullptr;
    sphb_params params;
    Consts c;
    size_t n = 0;
    int cur = 0; // which ping-pong buffer holds the current state

    DevBuf<float4> posm[2], vel[2], aux[2], accel[2];
    DevBuf<int> oid[2];
    DevBuf<int> key, slot, inv;
    DevBuf<int2> pairs;
    DevBuf<unsigned int> mask;       // neighbour masks, 2 words x rows x n
    DevBuf<float4> vr;               // (vx, vy, vz, rho) in sorted order, written by k_density for k_force's gathers
    bool vr_valid = false;           // vr matches the current sorted state and aux came from k_density
    int rows = 9;                    // (2*stencil+1)^2
    DevBuf<int> cell_start; // max_cells + 1 (+ slack)
    DevBuf<int> block_sums;
    DevBuf<int> row_groups; // exclusive scan of ceil(row length / kDT) over the (y,z) rows of cells (kMaxRows + 2)
    DevBuf<float> aos;      // staging for host AoS records
    DevBuf<float> scratch;  // eval_kernels, cells download
    SphereDev* d_spheres = nullptr; // kMaxSpheres records + one int (contact count) behind them
    int nspheres = 0;
    DevBuf<int2> contacts;  // particle-sphere contacts: (original id, slot), unsorted then sorted (2 x capacity)
    DevBuf<int> mg_counts;  // multi-GPU pack: 4 streams x tile counts, then totals[4], recv_counts[4] (+8 spare), one byte per warp
    int mg_blocks = 0;
    size_t mg_sent[2] = {0, 0}; // records the last pack put (wanted to put) into its left / right message, known after unpack
    bool mg_pending = false; // sphb_mg_pack ran, sphb_mg_unpack has not committed yet
    bool mg_used = false;    // this engine has exchanged slabs since its last upload: its arrays may hold ghosts
    size_t reserved = 0;
    GridDesc* d_grid = nullptr;
    BBoxAcc* d_bbox = nullptr;
    DevBuf<float4> warp_box; // per-warp boxes of the new positions, left by the fused force kernel (bbox grid mode)
    int box_records = 0;     // how many of them the last fused step wrote; 0 = none pending
    unsigned int* d_stats = nullptr; // [0] cells whose stable rank fell back to arrival order (cumulative), [1..3] spare
    // Particle counts on the device, for the path that never waits for them (sphb_mg_unpack_async):
    // d_cnt[0] = records in the arrays before the next sort (written by k_mg_commit), d_cnt[1] = live records after the last sort
    // (written by the sort).  While n_deferred, e->n is only the host's ESTIMATE (launch sizing); kernels read the real count.
    int* d_cnt = nullptr;
    int* d_status = nullptr;          // k_mg_commit's status record (8 ints)
    int* h_status = nullptr;          // pinned, 2 x 8 ints: where the status records land
    cudaEvent_t ev_status[2] = {nullptr, nullptr};
    unsigned long long status_issued = 0, status_seen = 0; // async unpacks enqueued / whose status has been read
    int last_status[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    bool n_deferred = false;
    int cnt_slot = 0;                 // which of d_cnt[0] / d_cnt[1] holds the current count
    size_t mstride = 0;               // stride of the neighbour-mask rows (capacity rounded up to 32)
    uint64_t steps_done = 0;
    long long max_cells = 0;

    bool have_state = false;
    bool oid_identity = false; // after sphb_mg_unpack: original id == slot; oid[cur] is only written when somebody needs it
    bool global_ids = false;   // sphb_set_ids: ids are the caller's labels and travel with the particles through the slab exchange
                               // (then downloads in "original order" make no sense for this engine and are refused)
    size_t n_dead = 0;       // records among the n that sphb_mg_pack marked dead (mass -1): the next sort files them last and drops them
    bool grid_valid = false; // cell_start / sorted order correspond to the current positions
    bool bbox_valid = false; // d_bbox holds the bbox of the current positions
    bool mask_valid = false; // neighbour masks correspond to the current sorted positions
    bool aos_valid = false;  // staging holds the bytes of the last AoS upload
    size_t aos_stride_w = 13;

    bool timing = false;
    static constexpr int kMaxTimed = 256;     // steps whose phase events are kept since timing was enabled
    cudaEvent_t* ev = nullptr;                // [kMaxTimed][4]
    int timed_steps = 0;
    uint64_t launches = 0;
};

namespace {

#define LAUNCH(e, kernel, grid, block, smem, ...)                                   \
    do {                                                                            \
        kernel<<<(grid), (block), (smem), (e)->stream>>>(__VA_ARGS__);              \
        (e)->launches++;                                                            \
    } while (0)

int use_device(sphb_engine* e) {
    CK(cudaSetDevice(e->device));
    return SPHB_OK;
}

// oid[cur] as an array (see oid_identity)
int ensure_oid(sphb_engine* e) {
    if (e->oid_identity && e->n) {
        LAUNCH(e, k_fill_iota, cdiv(e->n, 256), 256, 0, e->oid[e->cur].p, (int)e->n);
        CK(cudaGetLastError());
    }
    e->oid_identity = false;
    return SPHB_OK;
}

void build_consts(sphb_engine* e) {
    const sphb_params& p = e->params;
    Consts& c = e->c;
    memset(&c, 0, sizeof c);
    const float pi_f = (float)3.14159265358979323846264338327950288; // glm::pi<float>()
    c.h = p.h;
    c.h2 = p.h * p.h;
    c.inv_h = 1.0f / p.h;
    c.inv_h2 = 1.0f / c.h2;
    c.R2 = c.h2 * c.h2;
    // same double expressions as SphUtils.cpp:61,74,91, rounded once
    c.Kw = (float)(315.0 / ((double)(64 * pi_f) * pow((double)p.h, 9)));
    c.Kg = (float)(-45.0 / ((double)pi_f * pow((double)p.h, 6)));
    c.Kl = (float)(45.0 / ((double)pi_f * pow((double)p.h, 6)));
    c.rho0 = p.rho0;
    c.eos_k = p.eos_k;
    c.dt = p.dt;
    c.gx = p.gravity[0]; c.gy = p.gravity[1]; c.gz = p.gravity[2];
    c.alpha = p.visc_alpha;
    c.beta = p.visc_beta;
    c.eta_h2 = p.visc_eta * p.h * p.h;
    c.one_plus_e = 1.0f + p.restitution;
    c.wall_off = p.wall_test_off;
    c.wall_lim = p.wall_test_lim;
    c.domain_fixed = p.domain_mode == SPHB_DOMAIN_FIXED;
    for (int a = 0; a < 3; a++) {
        c.dom_lo[a] = p.domain_lo[a];
        c.dom_hi[a] = p.domain_hi[a];
    }
    c.lap_coef = p.laplacian_viscosity ? p.lap_visc_coef : 0.0f;
    for (int a = 0; a < 3; a++) { c.need_lo[a] = -INFINITY; c.need_hi[a] = INFINITY; }
    c.cell = p.cell_size > 0.0f ? p.cell_size : c.h2;
    const int div = p.cell_x_div > 0 ? p.cell_x_div : 2;
    c.cell_x = c.cell / (float)div;
    c.stencil_x = (int)ceilf(c.h2 / c.cell_x);
    while ((float)c.stencil_x * c.cell_x < c.h2) c.stencil_x++; // float safety: the x stencil must cover R
}

int ensure_capacity(sphb_engine* e, size_t n) {
    size_t want = n ? n : 1;
    if (want < e->reserved) want = e->reserved;
    for (int b = 0; b < 2; b++) {
        CK(e->posm[b].reserve(want));
        CK(e->vel[b].reserve(want));
        CK(e->aux[b].reserve(want));
        CK(e->accel[b].reserve(want));
        CK(e->oid[b].reserve(want));
    }
    CK(e->key.reserve(want));
    CK(e->slot.reserve(want));
    CK(e->inv.reserve(want));
    CK(e->pairs.reserve(want));
    e->rows = (e->c.cell >= e->c.h2) ? 9 : 25;
    e->mstride = (e->posm[0].cap + 31) & ~(size_t)31; // mask rows 128-byte aligned, and independent of the particle count
    CK(e->mask.reserve(e->mstride * (size_t)e->rows * 2));
    CK(e->vr.reserve(want));
    if (!e->c.domain_fixed) CK(e->warp_box.reserve(2 * (e->posm[0].cap / 32 + 4096)));
    if (e->nspheres) CK(e->contacts.reserve(2 * want));
    const int div = e->params.cell_x_div > 0 ? e->params.cell_x_div : 2;
    long long mc = e->params.max_cells > 0 ? e->params.max_cells : (long long)div * (long long)(2 * want + 65536);
    if (mc > (1ll << 30)) mc = 1ll << 30;
    if (mc > e->max_cells) {
        CK(e->cell_start.reserve((size_t)mc + 2 + kTrashBins));
        CK(e->block_sums.reserve((size_t)cdiv(mc + 2 + kTrashBins, kScanTile) + 1));
        CK(e->row_groups.reserve((size_t)(mc + 2 < kMaxRows + 2 ? mc + 2 : kMaxRows + 2)));
        e->max_cells = mc;
    }
    e->c.max_cells = e->max_cells;
    return SPHB_OK;
}

// bbox of the current positions -> d_bbox
int run_bbox(sphb_engine* e) {
    const int n = (int)e->n;
    LAUNCH(e, k_bbox_reset, 1, 32, 0, e->d_bbox);
    int blocks = cdiv(n, 256);
    if (blocks > 148 * 8) blocks = 148 * 8;
    if (blocks < 1) blocks = 1;
    LAUNCH(e, k_bbox, blocks, 256, 0, e->posm[e->cur].p, n, e->d_bbox);
    e->bbox_valid = true;
    CK(cudaGetLastError());
    return SPHB_OK;
}

// densities are needed inside the slab (and the periodic box) grown by one support radius: see Consts::need_lo
void set_need_box(sphb_engine* e, const sphb_slab* s) {
    const float R = e->c.h2;
    for (int a = 0; a < 3; a++) { e->c.need_lo[a] = -INFINITY; e->c.need_hi[a] = INFINITY; }
    if (!s || getenv("SPHB_DENSITY_ALL_GHOSTS")) return; // the environment switch is for A/B measurements
    e->c.need_lo[0] = s->lo - R;
    e->c.need_hi[0] = s->hi + R;
    if (s->periodic_y) { e->c.need_lo[1] = s->ylo - R; e->c.need_hi[1] = s->ylo + s->ylen + R; }
    if (s->periodic_z) { e->c.need_lo[2] = s->zlo - R; e->c.need_hi[2] = s->zlo + s->zlen + R; }
}

// the device-side particle count the kernels should use instead of their n argument (null: the host's n is exact)
inline const int* n_dev(const sphb_engine* e) { return e->n_deferred ? e->d_cnt + e->cnt_slot : nullptr; }

// d_bbox = box of the current positions: joined from the per-warp boxes of the last fused step if there are any, else one pass
int ensure_bbox(sphb_engine* e) {
    if (e->bbox_valid) return SPHB_OK;
    if (e->box_records > 0) {
        LAUNCH(e, k_bbox_reset, 1, 32, 0, e->d_bbox);
        int blocks = cdiv(e->box_records, 256);
        if (blocks > 148 * 4) blocks = 148 * 4;
        LAUNCH(e, k_bbox_fold, blocks, 256, 0, e->warp_box.p, e->box_records, e->d_bbox);
        CK(cudaGetLastError());
        e->box_records = 0;
        e->bbox_valid = true;
        return SPHB_OK;
    }
    return run_bbox(e);
}

constexpr int kSmallSort = 8192; // particles up to which the grid build is one single-block launch

// bookkeeping after either sort path: the new buffer is current, dead records are gone, derived data is stale
int finish_sort(sphb_engine* e, int nxt) {
    e->cur = nxt;
    e->oid_identity = false; // the sort wrote the new buffer's ids
    e->n -= e->n_dead; // the dead records now sit behind the live ones: [n, n + n_dead) of the new buffers, never read again
    e->n_dead = 0;
    e->cnt_slot = 1;   // d_cnt[1] = live count, written by the sort
    e->grid_valid = true;
    e->mask_valid = false;
    e->vr_valid = false;
    return SPHB_OK;
}

// K0..K3: grid description, keys, stable counting sort, reorder into the other ping-pong buffer.
int run_sort(sphb_engine* e, bool carry_all) {
    const int n = (int)e->n;
    const bool small = n <= kSmallSort;
    bool box_in_sort = false;
    if (!e->c.domain_fixed) {
        if (small && !e->bbox_valid && e->box_records == 0) box_in_sort = true; // the single-block sort takes the box itself
        else if (int rc = ensure_bbox(e)) return rc;
    }
    const int cur = e->cur, nxt = cur ^ 1;
    const int ntile_cap = cdiv(e->max_cells + 1 + kTrashBins, kScanTile);
    int setup_blocks = cdiv(e->max_cells + 1 + kTrashBins, 256 * 8);
    if (setup_blocks > 148 * 8) setup_blocks = 148 * 8;
    if (small) { // small scene: the eight kernels below as phases of one single-block launch (same functions)
        const int* oid_in = e->oid_identity ? (const int*)nullptr : e->oid[cur].p;
        const int rearm = box_in_sort ? 1 : 0, trash = (e->n_dead || e->n_deferred) ? 1 : 0;
#define SMALL_ARGS e->c, e->d_bbox, e->d_grid, e->cell_start.p, e->row_groups.p, e->key.p, e->slot.p, e->pairs.p, oid_in, e->posm[cur].p, \
                   e->vel[cur].p, e->aux[cur].p, e->accel[cur].p, e->posm[nxt].p, e->vel[nxt].p, e->aux[nxt].p, e->accel[nxt].p, e->oid[nxt].p, n, \
                   n_dev(e), e->d_cnt, rearm, trash, e->d_stats
        if (carry_all) LAUNCH(e, k_sort_small<true>, 1, kRgThreads, 0, SMALL_ARGS);
        else LAUNCH(e, k_sort_small<false>, 1, kRgThreads, 0, SMALL_ARGS);
#undef SMALL_ARGS
        CK(cudaGetLastError());
        if (box_in_sort) e->bbox_valid = true;
        return finish_sort(e, nxt);
    }
    LAUNCH(e, k_grid_setup, setup_blocks, 256, 0, e->c, e->d_bbox, e->d_grid, e->cell_start.p);
    const int pb = cdiv(n, 256);
    const int* nd = n_dev(e);
    LAUNCH(e, k_key_hist, pb, 256, 0, e->d_grid, e->posm[cur].p, n, nd, e->key.p, e->slot.p, e->cell_start.p, e->d_bbox,
           0, (e->n_dead || e->n_deferred) ? 1 : 0);
    LAUNCH(e, k_scan_reduce, ntile_cap, kScanThreads, 0, e->d_grid, e->cell_start.p, e->block_sums.p);
    LAUNCH(e, k_scan_top, 1, kScanThreads, 0, e->d_grid, e->block_sums.p);
    LAUNCH(e, k_scan_apply, ntile_cap, kScanThreads, 0, e->d_grid, e->cell_start.p, e->block_sums.p);
    LAUNCH(e, k_row_groups, 1, kRgThreads, 0, e->d_grid, e->cell_start.p, e->row_groups.p, e->d_cnt);
    LAUNCH(e, k_scatter, pb, 256, 0, e->key.p, e->slot.p, e->oid_identity ? (const int*)nullptr : e->oid[cur].p, e->cell_start.p, e->pairs.p, n, nd);
    e->oid_identity = false; // k_rank_gather writes the new buffer's ids
    if (carry_all)
        LAUNCH(e, k_rank_gather<true>, pb, 256, 0, e->pairs.p, e->key.p, e->cell_start.p, e->posm[cur].p, e->vel[cur].p, e->aux[cur].p,
               e->accel[cur].p, e->posm[nxt].p, e->vel[nxt].p, e->aux[nxt].p, e->accel[nxt].p, e->oid[nxt].p, n, nd, e->d_grid, e->d_stats);
    else
        LAUNCH(e, k_rank_gather<false>, pb, 256, 0, e->pairs.p, e->key.p, e->cell_start.p, e->posm[cur].p, e->vel[cur].p, e->aux[cur].p,
               e->accel[cur].p, e->posm[nxt].p, e->vel[nxt].p, e->aux[nxt].p, e->accel[nxt].p, e->oid[nxt].p, n, nd, e->d_grid, e->d_stats);
    CK(cudaGetLastError());
    return finish_sort(e, nxt);
}

// K4.  with_aux = false only rebuilds the neighbour masks (update_forces on an uploaded rho/P).
constexpr int kSmallScene = 32768; // particles up to which the latency-oriented kernel variants are used

// The fused step of an engine that exchanges slabs treats ghosts specially (force kernel: passed through; density kernel: the
// outer half of the ghost layers skipped) -- only in the configuration both kernels have that variant for.
bool ghost_path(const sphb_engine* e, int n) {
    const bool bbox_fused = SPHB_BBOX_FUSED && !e->c.domain_fixed;
    return e->mg_used && e->rows == 9 && !bbox_fused && (n > kSmallScene || e->n_deferred);
}

int run_density(sphb_engine* e, bool with_aux, bool skip_outer_ghosts = false) {
    const int n = (int)e->n;
    if (n == 0) return SPHB_OK;
    Consts c = e->c;
    if (!skip_outer_ghosts)
        for (int a = 0; a < 3; a++) { c.need_lo[a] = -INFINITY; c.need_hi[a] = INFINITY; }
    // groups <= n/kDT + (number of non-empty rows); threads beyond the device-side total exit at once
    long long rows_bound = e->max_cells < kMaxRows ? e->max_cells : kMaxRows;
    if (rows_bound > n) rows_bound = n;
    const long long groups = cdiv(n, kDT) + rows_bound;
    if (n <= kSmallScene && e->rows == 9 && !e->n_deferred) { // small scene: a group's rows side by side on 16 lanes (latency, not throughput)
        auto kfs = with_aux ? k_density<true, true, 16> : k_density<false, true, 16>;
        LAUNCH(e, kfs, cdiv(groups * 16, 128), 128, 0, c, e->d_grid, e->cell_start.p, e->row_groups.p, e->posm[e->cur].p,
               e->vel[e->cur].p, e->aux[e->cur].p, e->vr.p, e->mask.p, (int)e->mstride, 0);
    } else {
        auto kfn = with_aux ? k_density<true, true> : k_density<false, true>;
        const int blocks = cdiv(groups, 128);
        LAUNCH(e, kfn, blocks, 128, 0, c, e->d_grid, e->cell_start.p, e->row_groups.p, e->posm[e->cur].p,
               e->vel[e->cur].p, e->aux[e->cur].p, e->vr.p, e->mask.p, (int)e->mstride, 0);
        if (e->n_deferred) { // n was an estimate: a few blocks walk whatever groups lie beyond the main launch (normally none)
            auto ktail = with_aux ? k_density<true, true, 1, true> : k_density<false, true, 1, true>;
            LAUNCH(e, ktail, 148, 128, 0, c, e->d_grid, e->cell_start.p, e->row_groups.p, e->posm[e->cur].p,
                   e->vel[e->cur].p, e->aux[e->cur].p, e->vr.p, e->mask.p, (int)e->mstride, blocks * 128);
        }
    }
    CK(cudaGetLastError());
    e->mask_valid = true;
    e->vr_valid = with_aux;
    return SPHB_OK;
}

ParityConsts parity_consts(const sphb_engine* e) {
    const sphb_params& p = e->params;
    const float pi_f = (float)3.14159265358979323846264338327950288; // glm::pi<float>()
    ParityConsts pc;
    const float den_f = 64 * pi_f; // int * float -> float (SphUtils.cpp:61)
    pc.den_w = (double)den_f * pow((double)p.h, 9);
    pc.den_g = (double)pi_f * pow((double)p.h, 6);
    pc.eos_k = p.eos_k; pc.alpha = p.visc_alpha; pc.beta = p.visc_beta; pc.eta = p.visc_eta;
    return pc;
}

template <bool kFused, bool kBBox, bool kAccel, bool kDerive>
auto force_kernel(int rows) {
    return rows == 9 ? k_force<kFused, kBBox, kAccel, 9, kDerive> : k_force<kFused, kBBox, kAccel, 25, kDerive>;
}

int* sphere_count_ptr(sphb_engine* e) { return (int*)(e->d_spheres + kMaxSpheres); }

// particle-sphere contacts on the current state (after the wall pass): detect, order, replay
int run_sphere_collision(sphb_engine* e) {
    const int n = (int)e->n, cur = e->cur;
    CK(e->contacts.reserve(2 * (e->posm[cur].cap ? e->posm[cur].cap : 1)));
    int2* list = e->contacts.p;
    int2* sorted = e->contacts.p + e->posm[cur].cap;
    CK(cudaMemsetAsync(sphere_count_ptr(e), 0, sizeof(int), e->stream));
    if (int rco = ensure_oid(e)) return rco;
    LAUNCH(e, k_sphere_detect, cdiv(n, 256), 256, 0, e->posm[cur].p, e->oid[cur].p, n, e->d_spheres, e->nspheres, list, sphere_count_ptr(e));
    LAUNCH(e, k_sphere_rank, 148 * 4, 256, 0, list, sphere_count_ptr(e), sorted);
    LAUNCH(e, k_sphere_serial, 1, 32, 0, e->posm[cur].p, e->vel[cur].p, sorted, sphere_count_ptr(e), e->d_spheres, e->nspheres);
    CK(cudaGetLastError());
    return SPHB_OK;
}

// Status record of the latest sphb_mg_unpack_async, once the GPU has produced it.  wait = block on its event (recorded
// before the kernels of the following step, so by the time the next pack asks it has long completed and nothing drains).
int poll_status(sphb_engine* e, bool wait) {
    if (e->status_seen == e->status_issued) return SPHB_OK;
    const int slot = (int)((e->status_issued - 1) & 1);
    if (wait) CK(cudaEventSynchronize(e->ev_status[slot]));
    else {
        cudaError_t q = cudaEventQuery(e->ev_status[slot]);
        if (q == cudaErrorNotReady) return SPHB_OK;
        CK(q);
    }
    memcpy(e->last_status, e->h_status + 8 * slot, sizeof e->last_status);
    e->status_seen = e->status_issued;
    const int* h = e->last_status;
    e->mg_sent[0] = (size_t)h[5];
    e->mg_sent[1] = (size_t)h[6];
    if (e->n_deferred && e->cnt_slot == 0) e->n = (size_t)h[3]; // still before the sort: this IS the count
    if (h[7] & 3) {
        e->have_state = false;
        return fail(SPHB_ERR_CAPACITY, "a neighbour overflowed its message in the last exchange (flags %d): particles were lost, the state is dropped", h[7]);
    }
    if (h[7] & 4) {
        e->have_state = false;
        return fail(SPHB_ERR_CAPACITY, "the particle arrays overflowed in the last exchange: call sphb_reserve with more room (the state is dropped)");
    }
    if (h[7] & 8) {
        e->have_state = false;
        return fail(SPHB_ERR_STATE, "a neighbour never signalled on the peer-to-peer halo link (4 s): its process or GPU is gone; the state is dropped");
    }
    return SPHB_OK;
}

// Make e->n exact again after sphb_mg_unpack_async left it as an estimate (synchronises).
int resolve_counts(sphb_engine* e) {
    if (!e->n_deferred) return SPHB_OK;
    int h[2] = {0, 0};
    CK(cudaMemcpyAsync(h, e->d_cnt, sizeof h, cudaMemcpyDeviceToHost, e->stream));
    CK(cudaStreamSynchronize(e->stream));
    int rc = poll_status(e, true);
    e->n = (size_t)h[e->cnt_slot];
    e->n_dead = e->cnt_slot == 0 ? (size_t)e->last_status[4] : 0; // before the sort the dead are still among the n
    e->n_deferred = false;
    return rc;
}

int check_ready_hot(sphb_engine* e) { // the entry points of the per-step loop: never wait for the GPU
    if (!e) return fail(SPHB_ERR_INVALID, "null engine");
    if (!e->have_state) return fail(SPHB_ERR_STATE, "no particles uploaded");
    return use_device(e);
}

int check_ready(sphb_engine* e) {
    int rc = check_ready_hot(e);
    if (rc) return rc;
    return resolve_counts(e);
}

} // namespace

extern "C" {

void sphb_default_params(sphb_params* p) {
    if (!p) return;
    memset(p, 0, sizeof *p);
    p->h = 0.45f;      // main.cpp:34
    p->rho0 = 10.0f;   // main.cpp:36
    p->c = 100.0f;     // main.cpp:38
    p->dt = 0.033f;    // main.cpp:30
    p->eps = 0.01f;    // main.cpp:40
    p->eos_k = 100.0f; // SphUtils.cpp:144
    p->gravity[0] = 0.0f; p->gravity[1] = -1.0f; p->gravity[2] = 0.0f; // SphUtils.cpp:298
    p->visc_alpha = 0.1f;  // SphUtils.cpp:226
    p->visc_beta = 0.2f;   // SphUtils.cpp:226
    p->visc_eta = 0.01f;   // SphUtils.cpp:225
    p->restitution = 0.0f; // SphUtils.cpp:352
    p->cell_size = 0.0f;
    p->wall_test_off = 0.001; // SphUtils.cpp:349
    p->wall_test_lim = 0.01;
    p->domain_mode = SPHB_DOMAIN_BBOX;
    p->max_cells = 0;
    p->write_accel = 1;
    p->laplacian_viscosity = 0; // the reference's call is commented out (SphUtils.cpp:175)
    p->lap_visc_coef = 0.1f;    // SphUtils.cpp:277
    p->parity_mode = 0;
}

const char* sphb_last_error(void) { return g_err; }
const char* sphb_version(void) { return "sphb200 0.1 (sm_100a)"; }
size_t sphb_params_size(void) { return sizeof(sphb_params); }

int sphb_create(const sphb_params* p, int device, sphb_engine** out) {
    if (!p || !out) return fail(SPHB_ERR_INVALID, "null argument");
    if (!(p->h > 0.0f) || !(p->rho0 > 0.0f)) return fail(SPHB_ERR_INVALID, "h and rho0 must be positive");
    const float R = p->h * p->h;
    if (p->cell_size != 0.0f && !(p->cell_size >= 0.5f * R)) return fail(SPHB_ERR_INVALID, "cell_size must be 0 or >= h*h/2");
    if (p->domain_mode != SPHB_DOMAIN_BBOX && p->domain_mode != SPHB_DOMAIN_FIXED) return fail(SPHB_ERR_INVALID, "bad domain_mode");
    if (p->periodic_reserved[0] || p->periodic_reserved[1] || p->periodic_reserved[2])
        return fail(SPHB_ERR_INVALID, "sphb_params.periodic_reserved must be 0: periodic boxes are set up through sphb_slab (sphb_mg_pack)");
    if (p->cell_x_div < 0 || p->cell_x_div > 64) return fail(SPHB_ERR_INVALID, "cell_x_div must be in [0, 64]");
    if (p->domain_mode == SPHB_DOMAIN_FIXED)
        for (int a = 0; a < 3; a++)
            if (!(p->domain_hi[a] >= p->domain_lo[a])) return fail(SPHB_ERR_INVALID, "fixed domain needs hi >= lo");
    int ndev = 0;
    CK(cudaGetDeviceCount(&ndev));
    if (device < 0 || device >= ndev) return fail(SPHB_ERR_INVALID, "device %d out of range (%d visible)", device, ndev);
    sphb_engine* e = new (std::nothrow) sphb_engine();
    if (!e) return fail(SPHB_ERR_INVALID, "out of host memory");
    e->device = device;
    e->params = *p;
    build_consts(e);
    cudaError_t err = cudaSetDevice(device);
    if (err == cudaSuccess) err = cudaStreamCreateWithFlags(&e->stream, cudaStreamNonBlocking);
    if (err == cudaSuccess) err = cudaMalloc((void**)&e->d_grid, sizeof(GridDesc));
    if (err == cudaSuccess) err = cudaMalloc((void**)&e->d_bbox, sizeof(BBoxAcc));
    if (err == cudaSuccess) err = cudaMalloc((void**)&e->d_cnt, 8 * sizeof(int));
    if (err == cudaSuccess) err = cudaMalloc((void**)&e->d_status, 8 * sizeof(int));
    if (err == cudaSuccess) err = cudaMemsetAsync(e->d_cnt, 0, 8 * sizeof(int), e->stream);
    if (err == cudaSuccess) err = cudaMalloc((void**)&e->d_stats, 4 * sizeof(unsigned int));
    if (err == cudaSuccess) err = cudaMemsetAsync(e->d_stats, 0, 4 * sizeof(unsigned int), e->stream);
    if (err != cudaSuccess) {
        sphb_destroy(e);
        return fail(SPHB_ERR_CUDA, "engine setup failed: %s", cudaGetErrorString(err));
    }
    *out = e;
    g_err[0] = 0;
    return SPHB_OK;
}

void sphb_destroy(sphb_engine* e) {
    if (!e) return;
    cudaSetDevice(e->device);
    if (e->stream) cudaStreamSynchronize(e->stream);
    for (int b = 0; b < 2; b++) {
        e->posm[b].release(); e->vel[b].release(); e->aux[b].release(); e->accel[b].release(); e->oid[b].release();
    }
    e->key.release(); e->slot.release(); e->inv.release(); e->pairs.release(); e->mask.release(); e->vr.release();
    e->cell_start.release(); e->block_sums.release(); e->row_groups.release(); e->aos.release(); e->scratch.release(); e->mg_counts.release(); e->contacts.release(); e->warp_box.release();
    if (e->d_spheres) cudaFree(e->d_spheres);
    if (e->d_grid) cudaFree(e->d_grid);
    if (e->d_bbox) cudaFree(e->d_bbox);
    if (e->d_stats) cudaFree(e->d_stats);
    if (e->d_cnt) cudaFree(e->d_cnt);
    if (e->d_status) cudaFree(e->d_status);
    if (e->h_status) cudaFreeHost(e->h_status);
    for (int k = 0; k < 2; k++)
        if (e->ev_status[k]) cudaEventDestroy(e->ev_status[k]);
    if (e->ev) {
        for (int i = 0; i < 4 * sphb_engine::kMaxTimed; i++) cudaEventDestroy(e->ev[i]);
        delete[] e->ev;
    }
    if (e->stream) cudaStreamDestroy(e->stream);
    delete e;
}

int sphb_set_walls(sphb_engine* e, const void* walls, size_t count, size_t stride) {
    if (!e) return fail(SPHB_ERR_INVALID, "null engine");
    if (count > SPHB_MAX_WALLS) return fail(SPHB_ERR_CAPACITY, "at most %d walls", SPHB_MAX_WALLS);
    if (count && (!walls || stride < 24)) return fail(SPHB_ERR_INVALID, "walls need >= 24-byte records");
    e->c.nwalls = (int)count;
    for (size_t w = 0; w < count; w++) {
        float f[6];
        memcpy(f, (const char*)walls + w * stride, sizeof f);
        for (int a = 0; a < 3; a++) { e->c.wn[w][a] = f[a]; e->c.wc[w][a] = f[3 + a]; }
        // glm::normalize(n) = n * (1 / sqrt(dot(n, n))), float ops in this order (SphUtils.cpp:351)
        volatile float tx = f[0] * f[0], ty = f[1] * f[1], tz = f[2] * f[2];
        volatile float s = tx + ty;
        s = s + tz;
        volatile float inv = 1.0f / sqrtf(s);
        for (int a = 0; a < 3; a++) e->c.wu[w][a] = f[a] * inv;
    }
    return SPHB_OK;
}

int sphb_set_spheres(sphb_engine* e, const void* spheres, size_t count, size_t stride) {
    if (!e) return fail(SPHB_ERR_INVALID, "null engine");
    if (count > SPHB_MAX_SPHERES) return fail(SPHB_ERR_CAPACITY, "at most %d spheres", SPHB_MAX_SPHERES);
    if (count && (!spheres || stride < sizeof(SphereDev))) return fail(SPHB_ERR_INVALID, "spheres need >= 44-byte records");
    int rc = use_device(e);
    if (rc) return rc;
    if (!e->d_spheres) CK(cudaMalloc((void**)&e->d_spheres, kMaxSpheres * sizeof(SphereDev) + 16));
    SphereDev h[kMaxSpheres];
    for (size_t j = 0; j < count; j++) memcpy(&h[j], (const char*)spheres + j * stride, sizeof(SphereDev));
    if (count) {
        CK(cudaMemcpyAsync(e->d_spheres, h, count * sizeof(SphereDev), cudaMemcpyHostToDevice, e->stream));
        CK(cudaStreamSynchronize(e->stream)); // h lives on this stack frame
    }
    e->nspheres = (int)count;
    return SPHB_OK;
}

int sphb_get_spheres(sphb_engine* e, void* spheres, size_t count, size_t stride) {
    if (!e) return fail(SPHB_ERR_INVALID, "null engine");
    if (count != (size_t)e->nspheres) return fail(SPHB_ERR_INVALID, "the engine holds %d spheres, not %zu", e->nspheres, count);
    if (count && (!spheres || stride < sizeof(SphereDev))) return fail(SPHB_ERR_INVALID, "spheres need >= 44-byte records");
    if (!count) return SPHB_OK;
    int rc = use_device(e);
    if (rc) return rc;
    SphereDev h[kMaxSpheres];
    CK(cudaMemcpyAsync(h, e->d_spheres, count * sizeof(SphereDev), cudaMemcpyDeviceToHost, e->stream));
    CK(cudaStreamSynchronize(e->stream));
    for (size_t j = 0; j < count; j++) memcpy((char*)spheres + j * stride, &h[j], sizeof(SphereDev));
    return SPHB_OK;
}

int sphb_upload_aos(sphb_engine* e, const void* particles, size_t n, size_t stride) {
    if (!e) return fail(SPHB_ERR_INVALID, "null engine");
    if (n && !particles) return fail(SPHB_ERR_INVALID, "null particles");
    if (stride < 48 || stride % 4 || stride > 128) return fail(SPHB_ERR_INVALID, "stride must be a multiple of 4 in [48, 128]");
    if (n > 0x7fffff00u) return fail(SPHB_ERR_CAPACITY, "too many particles");
    int rc = use_device(e);
    if (rc) return rc;
    rc = ensure_capacity(e, n);
    if (rc) return rc;
    const size_t sw = stride / 4;
    CK(e->aos.reserve((n ? n : 1) * sw));
    e->n = n;
    e->cur = 0;
    e->aos_stride_w = sw;
    if (n) {
        CK(cudaMemcpyAsync(e->aos.p, particles, n * stride, cudaMemcpyHostToDevice, e->stream));
        LAUNCH(e, k_aos_to_soa, cdiv(n, kAosThreads), kAosThreads, kAosThreads * stride, e->aos.p, (int)sw, (int)n, e->posm[0].p,
               e->vel[0].p, e->aux[0].p, e->accel[0].p, e->oid[0].p);
        CK(cudaGetLastError());
    }
    e->have_state = true;
    set_need_box(e, nullptr);
    e->mg_used = false;
    e->global_ids = false;
    e->n_dead = 0;
    e->n_deferred = false;
    e->status_seen = e->status_issued;
    e->oid_identity = false;
    e->grid_valid = false;
    e->mask_valid = false;
    e->vr_valid = false;
    e->bbox_valid = false;
    e->box_records = 0;
    e->mg_pending = false;
    e->aos_valid = true;
    return SPHB_OK;
}

int sphb_upload_soa_device(sphb_engine* e, const void* posm, const void* vel, const int32_t* ids, size_t n) {
    if (!e) return fail(SPHB_ERR_INVALID, "null engine");
    if (n && (!posm || !vel)) return fail(SPHB_ERR_INVALID, "null device arrays");
    if (n > 0x7fffff00u) return fail(SPHB_ERR_CAPACITY, "too many particles");
    int rc = use_device(e);
    if (rc) return rc;
    rc = ensure_capacity(e, n);
    if (rc) return rc;
    e->n = n;
    e->cur = 0;
    if (n) {
        CK(cudaMemcpyAsync(e->posm[0].p, posm, n * sizeof(float4), cudaMemcpyDeviceToDevice, e->stream));
        LAUNCH(e, k_copy_vel_owned, cdiv(n, 256), 256, 0, (const float4*)vel, e->vel[0].p, (int)n); // w = 1: owned (0 would mean "ghost of another slab")
        CK(cudaMemsetAsync(e->aux[0].p, 0, n * sizeof(float4), e->stream));
        CK(cudaMemsetAsync(e->accel[0].p, 0, n * sizeof(float4), e->stream));
        if (ids) {
            // the ids index the inverse permutation and order the stable rank: anything but a permutation of 0..n-1 would
            // write out of bounds or lose particles later, so it is refused here (one extra pass and a synchronisation,
            // only on this id-carrying path)
            CK(e->scratch.reserve(4));
            unsigned int* d_bad = (unsigned int*)e->scratch.p;
            CK(cudaMemsetAsync(d_bad, 0, sizeof(unsigned int), e->stream));
            CK(cudaMemsetAsync(e->inv.p, 0xff, n * sizeof(int), e->stream));
            LAUNCH(e, k_check_ids, cdiv(n, 256), 256, 0, ids, (int)n, e->inv.p, d_bad);
            CK(cudaGetLastError());
            unsigned int h_bad = 0;
            CK(cudaMemcpyAsync(&h_bad, d_bad, sizeof h_bad, cudaMemcpyDeviceToHost, e->stream));
            CK(cudaStreamSynchronize(e->stream));
            if (h_bad) {
                e->have_state = false;
                return fail(SPHB_ERR_INVALID, "ids must be a permutation of 0..n-1: %u of %zu are out of range or repeated", h_bad, n);
            }
            CK(cudaMemcpyAsync(e->oid[0].p, ids, n * sizeof(int), cudaMemcpyDeviceToDevice, e->stream));
        } else
            LAUNCH(e, k_fill_iota, cdiv(n, 256), 256, 0, e->oid[0].p, (int)n);
        CK(cudaGetLastError());
    }
    e->have_state = true;
    set_need_box(e, nullptr);
    e->mg_used = false;
    e->global_ids = false;
    e->n_dead = 0;
    e->n_deferred = false;
    e->status_seen = e->status_issued;
    e->oid_identity = false;
    e->grid_valid = false;
    e->mask_valid = false;
    e->vr_valid = false;
    e->bbox_valid = false;
    e->box_records = 0;
    e->mg_pending = false;
    e->aos_valid = false;
    return SPHB_OK;
}

namespace {
int no_global_ids(sphb_engine* e, const char* what) {
    if (e->global_ids) return fail(SPHB_ERR_STATE, "%s returns particles in upload order, which this engine does not have: its ids are the caller's labels "
                                                   "(sphb_set_ids); use sphb_download_soa and place the records by their ids", what);
    return SPHB_OK;
}
} // namespace

int sphb_download_aos(sphb_engine* e, void* particles, size_t n, size_t stride, uint32_t field_mask) {
    int rc = check_ready(e);
    if (rc) return rc;
    if ((rc = no_global_ids(e, "sphb_download_aos"))) return rc;
    if (n > e->n) return fail(SPHB_ERR_INVALID, "download of %zu particles but the engine holds %zu", n, e->n);
    if (n && !particles) return fail(SPHB_ERR_INVALID, "null particles");
    if (stride < 48 || stride % 4 || stride > 128) return fail(SPHB_ERR_INVALID, "stride must be a multiple of 4 in [48, 128]");
    if (n == 0) return SPHB_OK;
    const size_t sw = stride / 4;
    if (!e->aos_valid || sw != e->aos_stride_w) {
        // no faithful copy of the host records on the device: fetch them so unmasked fields survive
        CK(e->aos.reserve(n * sw));
        CK(cudaMemcpyAsync(e->aos.p, particles, n * stride, cudaMemcpyHostToDevice, e->stream));
        e->aos_stride_w = sw;
        e->aos_valid = true;
    }
    const int cur = e->cur;
    if (int rco = ensure_oid(e)) return rco;
    LAUNCH(e, k_invert_perm, cdiv(e->n, 256), 256, 0, e->oid[cur].p, e->inv.p, (int)e->n);
    LAUNCH(e, k_soa_to_aos, cdiv(n, kAosThreads), kAosThreads, kAosThreads * stride, e->aos.p, (int)sw, (int)n, e->inv.p, e->posm[cur].p,
           e->vel[cur].p, e->aux[cur].p, e->accel[cur].p, field_mask);
    CK(cudaGetLastError());
    CK(cudaMemcpyAsync(particles, e->aos.p, n * stride, cudaMemcpyDeviceToHost, e->stream));
    CK(cudaStreamSynchronize(e->stream));
    return SPHB_OK;
}

int sphb_download_positions(sphb_engine* e, float* xyz, size_t n) {
    int rc = check_ready(e);
    if (rc) return rc;
    if ((rc = no_global_ids(e, "sphb_download_positions"))) return rc;
    if (n > e->n) return fail(SPHB_ERR_INVALID, "download of %zu particles but the engine holds %zu", n, e->n);
    if (n && !xyz) return fail(SPHB_ERR_INVALID, "null output");
    if (n == 0) return SPHB_OK;
    const int cur = e->cur;
    CK(e->scratch.reserve(3 * n));
    if (int rco = ensure_oid(e)) return rco;
    LAUNCH(e, k_invert_perm, cdiv(e->n, 256), 256, 0, e->oid[cur].p, e->inv.p, (int)e->n);
    LAUNCH(e, k_positions_packed, cdiv(n, 256), 256, 0, e->posm[cur].p, e->inv.p, e->scratch.p, (int)n);
    CK(cudaGetLastError());
    CK(cudaMemcpyAsync(xyz, e->scratch.p, 3 * n * sizeof(float), cudaMemcpyDeviceToHost, e->stream));
    CK(cudaStreamSynchronize(e->stream));
    return SPHB_OK;
}

void sphb_default_camera(sphb_camera* cam) {
    if (!cam) return;
    memset(cam, 0, sizeof *cam);
    // glm::perspective(45.0f, 4.0f/3.0f, 0.1f, 100.0f): fovy is in RADIANS in GLM 0.9.6 (gtc/matrix_transform.inl:212-232)
    const float fovy = 45.0f, aspect = 4.0f / 3.0f, zn = 0.1f, zf = 100.0f;
    const float th = tanf(fovy / 2.0f);
    float P[16] = {0}, V[16] = {0};
    P[0] = 1.0f / (aspect * th);
    P[5] = 1.0f / th;
    P[10] = -(zf + zn) / (zf - zn);
    P[11] = -1.0f;
    P[14] = -(2.0f * zf * zn) / (zf - zn);
    // glm::lookAt(eye (0,-0.8,1.5), center (0,-0.8,0), up (0,1,0))  (gtc/matrix_transform.inl:388-413)
    const float eye[3] = {0.0f, -0.8f, 1.5f}, ctr[3] = {0.0f, -0.8f, 0.0f}, up[3] = {0.0f, 1.0f, 0.0f};
    float f[3] = {ctr[0] - eye[0], ctr[1] - eye[1], ctr[2] - eye[2]};
    float fl = 1.0f / sqrtf(f[0] * f[0] + f[1] * f[1] + f[2] * f[2]);
    for (int a = 0; a < 3; a++) f[a] *= fl;
    float sv[3] = {f[1] * up[2] - f[2] * up[1], f[2] * up[0] - f[0] * up[2], f[0] * up[1] - f[1] * up[0]};
    float sl = 1.0f / sqrtf(sv[0] * sv[0] + sv[1] * sv[1] + sv[2] * sv[2]);
    for (int a = 0; a < 3; a++) sv[a] *= sl;
    const float u[3] = {sv[1] * f[2] - sv[2] * f[1], sv[2] * f[0] - sv[0] * f[2], sv[0] * f[1] - sv[1] * f[0]};
    V[0] = sv[0]; V[4] = sv[1]; V[8] = sv[2];
    V[1] = u[0];  V[5] = u[1];  V[9] = u[2];
    V[2] = -f[0]; V[6] = -f[1]; V[10] = -f[2];
    V[12] = -(sv[0] * eye[0] + sv[1] * eye[1] + sv[2] * eye[2]);
    V[13] = -(u[0] * eye[0] + u[1] * eye[1] + u[2] * eye[2]);
    V[14] = f[0] * eye[0] + f[1] * eye[1] + f[2] * eye[2];
    V[15] = 1.0f;
    for (int col = 0; col < 4; col++)       // MVP = P * V * I, column-major
        for (int row = 0; row < 4; row++) {
            float acc = 0.0f;
            for (int k = 0; k < 4; k++) acc += P[k * 4 + row] * V[col * 4 + k];
            cam->mvp[col * 4 + row] = acc;
        }
    cam->width = 1080;   // main.cpp:13-14
    cam->height = 920;
    cam->point_scale = 50.0f;
    cam->sprite_radius = 0.1f;
    cam->clear_value = 1.0f;
}

int sphb_render_depth(sphb_engine* e, const sphb_camera* cam, float* image) {
    int rc = check_ready(e);
    if (rc) return rc;
    if ((rc = no_global_ids(e, "sphb_render_depth"))) return rc;
    if (!cam || !image) return fail(SPHB_ERR_INVALID, "null argument");
    if (cam->width <= 0 || cam->height <= 0 || (long long)cam->width * cam->height > (1ll << 28)) return fail(SPHB_ERR_INVALID, "bad image size");
    Camera c;
    memcpy(c.m, cam->mvp, sizeof c.m);
    c.width = cam->width; c.height = cam->height;
    c.point_scale = cam->point_scale; c.sprite_radius = cam->sprite_radius; c.clear_value = cam->clear_value;
    c.depth_test = cam->depth_test != 0;
    const size_t npix = (size_t)c.width * c.height;
    const int n = (int)e->n, cur = e->cur;
    CK(e->scratch.reserve(3 * npix));
    float* d_img = e->scratch.p + 2 * npix;
    if (n) {
        if (int rco = ensure_oid(e)) return rco;
        LAUNCH(e, k_invert_perm, cdiv(n, 256), 256, 0, e->oid[cur].p, e->inv.p, n);
    }
    if (c.depth_test) {
        unsigned long long* winner = (unsigned long long*)e->scratch.p;
        CK(cudaMemsetAsync(winner, 0xff, npix * sizeof(unsigned long long), e->stream));
        if (n) LAUNCH(e, k_splat_winner_depth, cdiv((long long)n * 32, 256), 256, 0, c, e->posm[cur].p, e->inv.p, n, winner);
        LAUNCH(e, k_splat_shade_depth, cdiv(npix, 256), 256, 0, c, e->posm[cur].p, e->inv.p, winner, d_img);
    } else {
        unsigned int* winner = (unsigned int*)e->scratch.p;
        CK(cudaMemsetAsync(winner, 0, npix * sizeof(unsigned int), e->stream));
        if (n) LAUNCH(e, k_splat_winner, cdiv((long long)n * 32, 256), 256, 0, c, e->posm[cur].p, e->inv.p, n, winner);
        LAUNCH(e, k_splat_shade, cdiv(npix, 256), 256, 0, c, e->posm[cur].p, e->inv.p, winner, d_img);
    }
    CK(cudaGetLastError());
    CK(cudaMemcpyAsync(image, d_img, npix * sizeof(float), cudaMemcpyDeviceToHost, e->stream));
    CK(cudaStreamSynchronize(e->stream));
    return SPHB_OK;
}

int sphb_update_cells(sphb_engine* e) {
    int rc = check_ready(e);
    if (rc) return rc;
    if (e->n == 0) return SPHB_OK;
    return run_sort(e, true);
}

int sphb_update_density(sphb_engine* e) {
    int rc = check_ready(e);
    if (rc) return rc;
    if (e->n == 0) return SPHB_OK;
    if (!e->grid_valid && (rc = run_sort(e, true))) return rc;
    if (e->params.parity_mode) {
        if ((rc = ensure_oid(e))) return rc;
        LAUNCH(e, k_density_parity, cdiv(e->n, 128), 128, 0, e->c, parity_consts(e), e->d_grid, e->cell_start.p, e->posm[e->cur].p, e->oid[e->cur].p,
               e->aux[e->cur].p, (int)e->n);
        CK(cudaGetLastError());
        e->vr_valid = false;
        return SPHB_OK;
    }
    return run_density(e, true);
}

int sphb_update_forces(sphb_engine* e) {
    int rc = check_ready(e);
    if (rc) return rc;
    if (e->nspheres) LAUNCH(e, k_sphere_forces, 1, 32, 0, e->c, e->d_spheres, e->nspheres);
    if (e->n == 0) return SPHB_OK;
    if (!e->grid_valid && (rc = run_sort(e, true))) return rc;
    if (e->params.parity_mode) {
        if ((rc = ensure_oid(e))) return rc;
        LAUNCH(e, k_force_parity, cdiv(e->n, 128), 128, 0, e->c, parity_consts(e), e->d_grid, e->cell_start.p, e->posm[e->cur].p, e->vel[e->cur].p,
               e->aux[e->cur].p, e->oid[e->cur].p, e->accel[e->cur].p, (int)e->n);
        CK(cudaGetLastError());
        return SPHB_OK;
    }
    if (!e->mask_valid && (rc = run_density(e, false))) return rc;
    const int n = (int)e->n, cur = e->cur;
    // vr_valid: rho/P in aux are this engine's own (update_density ran on this sorted state) -> two-gather kernel
    auto kfn = e->vr_valid ? force_kernel<false, false, true, true>(e->rows) : force_kernel<false, false, true, false>(e->rows);
    LAUNCH(e, kfn, cdiv(n, 128), 128, 0, e->c, e->d_grid, e->cell_start.p, e->posm[cur].p, e->vel[cur].p,
           e->aux[cur].p, e->vr.p, e->mask.p, e->accel[cur].p, (float4*)nullptr, (float4*)nullptr, (float4*)nullptr, n, (const int*)nullptr, (int)e->mstride, 0);
    if (e->params.laplacian_viscosity)
        LAUNCH(e, k_lap_viscosity, cdiv(n, 128), 128, 0, e->c, e->d_grid, e->cell_start.p, e->posm[cur].p, e->vel[cur].p, e->aux[cur].p,
               e->mask.p, e->accel[cur].p, n, (int)e->mstride);
    CK(cudaGetLastError());
    return SPHB_OK;
}

int sphb_update_posvel(sphb_engine* e) {
    int rc = check_ready(e);
    if (rc) return rc;
    if (e->nspheres) LAUNCH(e, k_sphere_posvel, 1, 32, 0, e->c, e->d_spheres, e->nspheres);
    if (e->n == 0) return SPHB_OK;
    const int n = (int)e->n, cur = e->cur;
    LAUNCH(e, k_posvel, cdiv(n, 256), 256, 0, e->c, e->posm[cur].p, e->vel[cur].p, e->accel[cur].p, n);
    CK(cudaGetLastError());
    e->grid_valid = false;
    e->mask_valid = false;
    e->vr_valid = false;
    e->bbox_valid = false;
    e->box_records = 0;
    return SPHB_OK;
}

int sphb_collision(sphb_engine* e) {
    int rc = check_ready(e);
    if (rc) return rc;
    if (e->n == 0) return SPHB_OK;
    const int n = (int)e->n, cur = e->cur;
    LAUNCH(e, k_collision, cdiv(n, 256), 256, 0, e->c, e->posm[cur].p, e->vel[cur].p, n);
    CK(cudaGetLastError());
    if (e->nspheres && (rc = run_sphere_collision(e))) return rc;
    e->grid_valid = false;
    e->mask_valid = false;
    e->vr_valid = false;
    e->bbox_valid = false;
    e->box_records = 0;
    return SPHB_OK;
}

int sphb_step(sphb_engine* e, int nsteps) {
    int rc = check_ready_hot(e);
    if (!rc && e->n_deferred && (e->nspheres || e->params.laplacian_viscosity || e->params.parity_mode || e->rows != 9)) rc = resolve_counts(e); // those paths size work from the host's n
    if (rc) return rc;
    if (nsteps < 0) return fail(SPHB_ERR_INVALID, "nsteps < 0");
    if (e->n == 0) return SPHB_OK;
    // Bounding-box grid mode: the box of the new positions comes from one streaming pass at the start of the next sort (k_bbox,
    // 16 B/particle: 0.05 ms at 16.7 M).  Tracking it inside the fused force kernel instead (SPHB_BBOX_FUSED=1: per-warp boxes,
    // joined by k_bbox_fold) costs that kernel six live registers and 0.17 ms; earlier forms of the fold cost 0.36 - 0.5 ms.
    const bool bbox = SPHB_BBOX_FUSED && !e->c.domain_fixed;
    const bool acc = e->params.write_accel != 0;
    if (e->params.laplacian_viscosity || e->params.parity_mode) { // optional extra force pass / reference-order arithmetic: run the phases one by one
        for (int s = 0; s < nsteps; s++)
            if ((rc = sphb_update_density(e)) || (rc = sphb_update_forces(e)) || (rc = sphb_update_posvel(e)) || (rc = sphb_collision(e)))
                return rc;
        return SPHB_OK;
    }
    for (int s = 0; s < nsteps; s++) {
        const bool timed = e->timing && e->timed_steps < sphb_engine::kMaxTimed;
        cudaEvent_t* ev = timed ? e->ev + 4 * e->timed_steps : nullptr;
        if (timed) CK(cudaEventRecord(ev[0], e->stream));
        // Always re-sort: positions moved.  (First step after an upload of an already valid grid is rare.)
        if ((rc = run_sort(e, false))) return rc;
        const int n = (int)e->n; // after the sort: records marked dead by the slab exchange are gone
        if (n == 0) return SPHB_OK;
        if (timed) CK(cudaEventRecord(ev[1], e->stream));
        if ((rc = run_density(e, true, ghost_path(e, n)))) return rc;
        if (timed) CK(cudaEventRecord(ev[2], e->stream));
        const int cur = e->cur, nxt = cur ^ 1;
        const int blocks = cdiv(n, 128);
#define FORCE_ARGS e->c, e->d_grid, e->cell_start.p, e->posm[cur].p, e->vel[cur].p, e->aux[cur].p, e->vr.p, e->mask.p, e->accel[nxt].p, e->posm[nxt].p, e->vel[nxt].p, e->warp_box.p, n, n_dev(e), (int)e->mstride
        auto kfn = bbox ? (acc ? force_kernel<true, true, true, true>(e->rows) : force_kernel<true, true, false, true>(e->rows))
                        : (acc ? force_kernel<true, false, true, true>(e->rows) : force_kernel<true, false, false, true>(e->rows));
        if (n <= kSmallScene && e->rows == 9 && !e->n_deferred) // small scene: four neighbours in flight per thread (latency, not throughput)
            kfn = bbox ? (acc ? k_force<true, true, true, 9, true, 4> : k_force<true, true, false, 9, true, 4>)
                       : (acc ? k_force<true, false, true, 9, true, 4> : k_force<true, false, false, 9, true, 4>);
        // an engine that exchanges slabs holds ghosts: the variant that passes them through (big scenes, 9-row stencil, no bbox fold)
        const bool ghosts = ghost_path(e, n);
        if (ghosts) kfn = acc ? k_force<true, false, true, 9, true, 1, false, true> : k_force<true, false, false, 9, true, 1, false, true>;
        LAUNCH(e, kfn, blocks, 128, 0, FORCE_ARGS, 0);
        int box_records = bbox ? blocks * 4 : 0;
        if (e->n_deferred) { // n was an estimate: a few blocks walk whatever particles lie beyond the main launch (normally none)
            constexpr int kTail = 148;
            auto ktail = bbox ? (acc ? k_force<true, true, true, 9, true, 1, true> : k_force<true, true, false, 9, true, 1, true>)
                              : (acc ? k_force<true, false, true, 9, true, 1, true> : k_force<true, false, false, 9, true, 1, true>);
            if (ghosts) ktail = acc ? k_force<true, false, true, 9, true, 1, true, true> : k_force<true, false, false, 9, true, 1, true, true>;
            LAUNCH(e, ktail, kTail, 128, 0, FORCE_ARGS, blocks * 128);
            if (bbox) box_records += kTail * 4;
        }
#undef FORCE_ARGS
        CK(cudaGetLastError());
        // the new state (posm, vel, accel) is in nxt; aux (rho, P) and ids stay where they are: copy-free
        // hand-over by swapping the roles of those two buffers as well.
        {
            DevBuf<float4> t = e->aux[cur]; e->aux[cur] = e->aux[nxt]; e->aux[nxt] = t;
            DevBuf<int> u = e->oid[cur]; e->oid[cur] = e->oid[nxt]; e->oid[nxt] = u;
        }
        e->cur = nxt;
        e->grid_valid = false;
        e->mask_valid = false;
        e->vr_valid = false;
        e->bbox_valid = false;
        e->box_records = box_records; // joined into the box at the start of the next sort (ensure_bbox)
        if (e->nspheres) { // the same order as step_scene: forces, posvel (spheres after particles), collision
            LAUNCH(e, k_sphere_forces, 1, 32, 0, e->c, e->d_spheres, e->nspheres);
            LAUNCH(e, k_sphere_posvel, 1, 32, 0, e->c, e->d_spheres, e->nspheres);
            if ((rc = run_sphere_collision(e))) return rc;
        }
        if (timed) {
            CK(cudaEventRecord(ev[3], e->stream));
            e->timed_steps++;
        }
        e->steps_done++;
    }
    return SPHB_OK;
}

int sphb_step_aos(sphb_engine* e, void* particles, size_t n, size_t stride, int nsteps, uint32_t field_mask) {
    int rc = sphb_upload_aos(e, particles, n, stride);
    if (rc) return rc;
    if ((rc = sphb_step(e, nsteps))) return rc;
    return sphb_download_aos(e, particles, n, stride, field_mask); // synchronises
}

int sphb_synchronize(sphb_engine* e) {
    if (!e) return fail(SPHB_ERR_INVALID, "null engine");
    int rc = use_device(e);
    if (rc) return rc;
    CK(cudaStreamSynchronize(e->stream));
    return SPHB_OK;
}

int sphb_host_register(void* ptr, size_t bytes) {
    if (!ptr || !bytes) return fail(SPHB_ERR_INVALID, "null range");
    CK(cudaHostRegister(ptr, bytes, cudaHostRegisterPortable));
    return SPHB_OK;
}

int sphb_host_unregister(void* ptr) {
    if (!ptr) return fail(SPHB_ERR_INVALID, "null range");
    CK(cudaHostUnregister(ptr));
    return SPHB_OK;
}

int sphb_eval_kernels(sphb_engine* e, const float* xi, const float* xj, size_t count, float* w, float* grad, float* lap) {
    if (!e) return fail(SPHB_ERR_INVALID, "null engine");
    if (count == 0) return SPHB_OK;
    if (!xi || !xj) return fail(SPHB_ERR_INVALID, "null points");
    if (count > (1u << 26)) return fail(SPHB_ERR_CAPACITY, "too many pairs");
    int rc = use_device(e);
    if (rc) return rc;
    CK(e->scratch.reserve(count * 11));
    float* d_xi = e->scratch.p;
    float* d_xj = d_xi + 3 * count;
    float* d_w = d_xj + 3 * count;
    float* d_g = d_w + count;
    float* d_l = d_g + 3 * count;
    CK(cudaMemcpyAsync(d_xi, xi, 3 * count * sizeof(float), cudaMemcpyHostToDevice, e->stream));
    CK(cudaMemcpyAsync(d_xj, xj, 3 * count * sizeof(float), cudaMemcpyHostToDevice, e->stream));
    LAUNCH(e, k_eval_kernels, cdiv(count, 256), 256, 0, e->c, d_xi, d_xj, (int)count, d_w, d_g, d_l);
    CK(cudaGetLastError());
    if (w) CK(cudaMemcpyAsync(w, d_w, count * sizeof(float), cudaMemcpyDeviceToHost, e->stream));
    if (grad) CK(cudaMemcpyAsync(grad, d_g, 3 * count * sizeof(float), cudaMemcpyDeviceToHost, e->stream));
    if (lap) CK(cudaMemcpyAsync(lap, d_l, count * sizeof(float), cudaMemcpyDeviceToHost, e->stream));
    CK(cudaStreamSynchronize(e->stream));
    return SPHB_OK;
}

int sphb_get_cells(sphb_engine* e, int32_t* cell_xyz, int32_t* perm, int32_t* dims, float* origin, float* cell) {
    int rc = check_ready(e);
    if (rc) return rc;
    if (!e->grid_valid) return fail(SPHB_ERR_STATE, "no grid: call sphb_update_cells first");
    if (cell_xyz && (rc = no_global_ids(e, "sphb_get_cells (cell_xyz)"))) return rc;
    const size_t n = e->n;
    GridDesc g;
    CK(cudaMemcpyAsync(&g, e->d_grid, sizeof g, cudaMemcpyDeviceToHost, e->stream));
    if (cell_xyz && n) {
        CK(e->scratch.reserve(3 * n));
        if (int rco = ensure_oid(e)) return rco;
        LAUNCH(e, k_cells_xyz, cdiv(n, 256), 256, 0, e->d_grid, e->posm[e->cur].p, e->oid[e->cur].p, (int*)e->scratch.p, (int)n);
        CK(cudaGetLastError());
        CK(cudaMemcpyAsync(cell_xyz, e->scratch.p, 3 * n * sizeof(int), cudaMemcpyDeviceToHost, e->stream));
    }
    if (perm && n) CK(cudaMemcpyAsync(perm, e->oid[e->cur].p, n * sizeof(int), cudaMemcpyDeviceToHost, e->stream));
    CK(cudaStreamSynchronize(e->stream));
    if (dims) { dims[0] = g.dims[0]; dims[1] = g.dims[1]; dims[2] = g.dims[2]; }
    if (origin) { origin[0] = g.origin[0]; origin[1] = g.origin[1]; origin[2] = g.origin[2]; }
    if (cell) { cell[0] = g.cell[0]; cell[1] = g.cell[1]; cell[2] = g.cell[2]; }
    return SPHB_OK;
}

int sphb_get_bbox(sphb_engine* e, float bbox[6]) {
    int rc = check_ready(e);
    if (rc) return rc;
    if (!bbox) return fail(SPHB_ERR_INVALID, "null bbox");
    if (e->n == 0) return fail(SPHB_ERR_STATE, "no particles");
    if ((rc = ensure_bbox(e))) return rc;
    BBoxAcc acc;
    CK(cudaMemcpyAsync(&acc, e->d_bbox, sizeof acc, cudaMemcpyDeviceToHost, e->stream));
    CK(cudaStreamSynchronize(e->stream));
    for (int a = 0; a < 3; a++) {
        unsigned int lo = acc.lo[a], hi = acc.hi[a];
        unsigned int bl = (lo & 0x80000000u) ? (lo & 0x7fffffffu) : ~lo;
        unsigned int bh = (hi & 0x80000000u) ? (hi & 0x7fffffffu) : ~hi;
        memcpy(&bbox[a], &bl, 4);
        memcpy(&bbox[3 + a], &bh, 4);
    }
    return SPHB_OK;
}

size_t sphb_count(const sphb_engine* e) {
    if (!e) return 0;
    sphb_engine* m = const_cast<sphb_engine*>(e);
    if (m->n_deferred && cudaSetDevice(m->device) == cudaSuccess) resolve_counts(m); // after sphb_mg_unpack_async: fetch the count (synchronises)
    return e->n;
}

int sphb_count_nonfinite(sphb_engine* e, size_t* count) {
    int rc = check_ready(e);
    if (rc) return rc;
    if (!count) return fail(SPHB_ERR_INVALID, "null count");
    *count = 0;
    if (e->n == 0) return SPHB_OK;
    CK(e->scratch.reserve(4));
    unsigned long long* d = (unsigned long long*)e->scratch.p;
    CK(cudaMemsetAsync(d, 0, sizeof(unsigned long long), e->stream));
    int blocks = cdiv(e->n, 256);
    if (blocks > 148 * 8) blocks = 148 * 8;
    LAUNCH(e, k_count_nonfinite, blocks, 256, 0, e->posm[e->cur].p, e->vel[e->cur].p, (int)e->n, d);
    CK(cudaGetLastError());
    unsigned long long h = 0;
    CK(cudaMemcpyAsync(&h, d, sizeof h, cudaMemcpyDeviceToHost, e->stream));
    CK(cudaStreamSynchronize(e->stream));
    *count = (size_t)h;
    return SPHB_OK;
}

int sphb_enable_timing(sphb_engine* e, int on) {
    if (!e) return fail(SPHB_ERR_INVALID, "null engine");
    int rc = use_device(e);
    if (rc) return rc;
    if (on && !e->ev) {
        e->ev = new (std::nothrow) cudaEvent_t[4 * sphb_engine::kMaxTimed]();
        if (!e->ev) return fail(SPHB_ERR_INVALID, "out of host memory");
        for (int i = 0; i < 4 * sphb_engine::kMaxTimed; i++) CK(cudaEventCreate(&e->ev[i]));
    }
    e->timing = on != 0;
    e->timed_steps = 0;
    return SPHB_OK;
}

int sphb_get_timing(sphb_engine* e, float ms[4], int* steps) {
    if (!e || !ms) return fail(SPHB_ERR_INVALID, "null argument");
    int rc = use_device(e);
    if (rc) return rc;
    CK(cudaStreamSynchronize(e->stream));
    double acc[4] = {0, 0, 0, 0};
    for (int s = 0; s < e->timed_steps; s++) {
        cudaEvent_t* ev = e->ev + 4 * s;
        float t;
        CK(cudaEventElapsedTime(&t, ev[0], ev[1])); acc[0] += t;
        CK(cudaEventElapsedTime(&t, ev[1], ev[2])); acc[1] += t;
        CK(cudaEventElapsedTime(&t, ev[2], ev[3])); acc[2] += t;
        CK(cudaEventElapsedTime(&t, ev[0], ev[3])); acc[3] += t;
    }
    for (int k = 0; k < 4; k++) ms[k] = e->timed_steps ? (float)(acc[k] / e->timed_steps) : 0.0f;
    if (steps) *steps = e->timed_steps;
    return SPHB_OK;
}

uint64_t sphb_launch_count(const sphb_engine* e) { return e ? e->launches : 0; }

int sphb_get_stats(sphb_engine* e, sphb_stats* out) {
    if (!e || !out) return fail(SPHB_ERR_INVALID, "null argument");
    int rc = use_device(e);
    if (rc) return rc;
    unsigned int h[4] = {0, 0, 0, 0};
    CK(cudaMemcpyAsync(h, e->d_stats, sizeof h, cudaMemcpyDeviceToHost, e->stream));
    CK(cudaStreamSynchronize(e->stream));
    memset(out, 0, sizeof *out);
    out->arrival_order_cells = h[0];
    out->steps = e->steps_done;
    out->launches = e->launches;
    out->particles = e->n;
    out->capacity = e->posm[0].cap;
    return SPHB_OK;
}

int sphb_get_device_state(sphb_engine* e, void** posm, void** vel, void** aux, void** accel, int32_t** ids) {
    int rc = check_ready(e);
    if (rc) return rc;
    const int cur = e->cur;
    if (posm) *posm = e->posm[cur].p;
    if (vel) *vel = e->vel[cur].p;
    if (aux) *aux = e->aux[cur].p;
    if (accel) *accel = e->accel[cur].p;
    if (int rco = ensure_oid(e)) return rco;
    if (ids) *ids = e->oid[cur].p;
    return SPHB_OK;
}

int sphb_download_soa(sphb_engine* e, float* posm, float* vel, float* aux, float* accel, int32_t* ids) {
    int rc = check_ready(e);
    if (rc) return rc;
    const size_t n = e->n;
    const int cur = e->cur;
    if (n) {
        if (posm) CK(cudaMemcpyAsync(posm, e->posm[cur].p, n * sizeof(float4), cudaMemcpyDeviceToHost, e->stream));
        if (vel) CK(cudaMemcpyAsync(vel, e->vel[cur].p, n * sizeof(float4), cudaMemcpyDeviceToHost, e->stream));
        if (aux) CK(cudaMemcpyAsync(aux, e->aux[cur].p, n * sizeof(float4), cudaMemcpyDeviceToHost, e->stream));
        if (accel) CK(cudaMemcpyAsync(accel, e->accel[cur].p, n * sizeof(float4), cudaMemcpyDeviceToHost, e->stream));
        if (int rco = ensure_oid(e)) return rco;
        if (ids) CK(cudaMemcpyAsync(ids, e->oid[cur].p, n * sizeof(int), cudaMemcpyDeviceToHost, e->stream));
    }
    CK(cudaStreamSynchronize(e->stream));
    return SPHB_OK;
}

void* sphb_stream(sphb_engine* e) { return e ? (void*)e->stream : nullptr; }

} // extern "C"

// =====================================================================================================
// Multi-GPU slab decomposition (include/sphb200.h: sphb_mg_*)

extern "C" size_t sphb_mg_message_bytes(size_t capacity) { return (capacity + 1) * 2 * sizeof(float4); }

extern "C" int sphb_reserve(sphb_engine* e, size_t n_max) {
    if (!e) return fail(SPHB_ERR_INVALID, "null engine");
    if (e->have_state && n_max > e->posm[0].cap) return fail(SPHB_ERR_STATE, "sphb_reserve must come before the first upload");
    if (n_max > 0x7fffff00u) return fail(SPHB_ERR_CAPACITY, "too many particles");
    int rc = use_device(e);
    if (rc) return rc;
    e->reserved = n_max;
    return ensure_capacity(e, n_max);
}

extern "C" int sphb_set_ids(sphb_engine* e, const int32_t* ids, size_t n) {
    int rc = check_ready(e);
    if (rc) return rc;
    if (n != e->n) return fail(SPHB_ERR_INVALID, "sphb_set_ids: %zu ids for %zu particles", n, e->n);
    if (n && !ids) return fail(SPHB_ERR_INVALID, "null ids");
    if ((rc = ensure_oid(e))) return rc;
    // ids arrive in ORIGINAL (upload) order; the state may already be sorted: place each through the current ids
    CK(e->scratch.reserve(n ? n : 1));
    int* d_ids = (int*)e->scratch.p;
    if (n) {
        CK(cudaMemcpyAsync(d_ids, ids, n * sizeof(int), cudaMemcpyHostToDevice, e->stream));
        LAUNCH(e, k_relabel, cdiv(n, 256), 256, 0, e->oid[e->cur].p, d_ids, (int)n);
        CK(cudaGetLastError());
        CK(cudaStreamSynchronize(e->stream)); // ids may live on the caller's stack
    }
    e->global_ids = true;
    return SPHB_OK;
}

extern "C" int sphb_mg_pack(sphb_engine* e, const sphb_slab* slab, void* msg_left, void* msg_right, size_t capacity) {
    return sphb_mg_pack_limits(e, slab, msg_left, capacity, msg_right, capacity);
}

extern "C" int sphb_mg_pack_limits(sphb_engine* e, const sphb_slab* slab, void* msg_left, size_t cap_left, void* msg_right, size_t cap_right) {
    int rc = check_ready_hot(e);
    if (rc) return rc;
    if ((rc = poll_status(e, true))) return rc; // the last asynchronous unpack's verdict (overflow) and counts: long since on the host
    if (!slab) return fail(SPHB_ERR_INVALID, "null slab");
    if ((slab->has_left && !msg_left) || (slab->has_right && !msg_right)) return fail(SPHB_ERR_INVALID, "missing message buffer");
    if (cap_left > 0x3fffffffu || cap_right > 0x3fffffffu) return fail(SPHB_ERR_CAPACITY, "message capacity too large");
    if ((slab->periodic_y && !(slab->ylen >= 2.0f * slab->halo)) || (slab->periodic_z && !(slab->zlen >= 2.0f * slab->halo)))
        return fail(SPHB_ERR_INVALID, "a periodic axis must be at least two halos long");
    Slab s;
    s.lo = slab->lo; s.hi = slab->hi; s.halo = slab->halo;
    s.has_left = slab->has_left; s.has_right = slab->has_right;
    s.shift_left = slab->shift_left; s.shift_right = slab->shift_right;
    s.per_y = slab->periodic_y; s.per_z = slab->periodic_z;
    s.ylo = slab->ylo; s.ylen = slab->ylen; s.zlo = slab->zlo; s.zlen = slab->zlen;
    const int n = (int)e->n, cur = e->cur; // n: an estimate while the count lives on the device (the kernels read the real one)
    const int out_cap = (int)(e->posm[cur].cap < 0x7fffffffu ? e->posm[cur].cap : 0x7fffffffu);
    const int nb_layout = cdiv(out_cap > 0 ? out_cap : 1, kMgTile); // arrays laid out for the capacity: the count may not be known here
    const int nb = n ? cdiv(n, kMgTile) : 1;
    CK(e->mg_counts.reserve((size_t)kMgStreams * nb_layout + 16 + 2 * (size_t)nb_layout)); // counts, totals + received counts, one byte per warp
    e->mg_blocks = nb_layout;
    int* totals = e->mg_counts.p + kMgStreams * nb_layout;
    unsigned char* touch = (unsigned char*)(totals + 16);
    const int* nd = n_dev(e);
    LAUNCH(e, k_mg_count, nb, kMgThreads, 0, s, e->posm[cur].p, e->vel[cur].p, n, nd, nb_layout, e->mg_counts.p, touch);
    LAUNCH(e, k_mg_scan, kMgStreams, kMgScanThreads, 0, e->mg_counts.p, nb_layout, n, nd, totals);
    if (e->global_ids) {
        if (int rco = ensure_oid(e)) return rco;
    }
    LAUNCH(e, k_mg_write, nb, kMgThreads, 0, s, e->posm[cur].p, e->vel[cur].p, e->global_ids ? e->oid[cur].p : (int*)nullptr, n, nd, nb_layout, e->mg_counts.p, totals, out_cap, touch,
           (float4*)msg_left, (float4*)msg_right, (int)cap_left, (int)cap_right);
    CK(cudaGetLastError());
    e->mg_pending = true;
    e->mg_used = true;
    return SPHB_OK;
}

extern "C" int sphb_mg_sent(sphb_engine* e, size_t sent[2]) {
    if (!e || !sent) return fail(SPHB_ERR_INVALID, "null argument");
    sent[0] = e->mg_sent[0];
    sent[1] = e->mg_sent[1];
    return SPHB_OK;
}

namespace {
Slab to_slab(const sphb_slab* slab) {
    Slab s;
    s.lo = slab->lo; s.hi = slab->hi; s.halo = slab->halo;
    s.has_left = slab->has_left; s.has_right = slab->has_right;
    s.shift_left = slab->shift_left; s.shift_right = slab->shift_right;
    s.per_y = slab->periodic_y; s.per_z = slab->periodic_z;
    s.ylo = slab->ylo; s.ylen = slab->ylen; s.zlo = slab->zlo; s.zlen = slab->zlen;
    return s;
}
} // namespace

namespace {
int unpack_async_impl(sphb_engine* e, const sphb_slab* slab, const void* from_left, const void* from_right, const int* link_err);

// With CUDA's lazy module loading the FIRST launch of a kernel loads it, and that can wait for the whole device -- which never
// comes back if a stream of this device (or, with one host thread driving several GPUs, the stream the neighbour is waiting
// for) sits in a flag wait.  So every kernel the exchange -> step loop launches is loaded here, once per device, up front.
int preload_hot_kernels() {
#define PRELOAD(f)                                      \
    do {                                                \
        cudaFuncAttributes a_;                          \
        CK(cudaFuncGetAttributes(&a_, f));              \
    } while (0)
    PRELOAD(k_link_set); PRELOAD(k_link_wait); PRELOAD(k_mg_count); PRELOAD(k_mg_scan); PRELOAD(k_mg_write); PRELOAD(k_mg_append); PRELOAD(k_mg_commit);
    PRELOAD(k_grid_setup); PRELOAD(k_key_hist); PRELOAD(k_scan_reduce); PRELOAD(k_scan_top); PRELOAD(k_scan_apply); PRELOAD(k_row_groups);
    PRELOAD(k_scatter); PRELOAD(k_rank_gather<false>); PRELOAD(k_sort_small<false>); PRELOAD(k_bbox_reset); PRELOAD(k_bbox); PRELOAD(k_fill_iota);
    PRELOAD((k_density<true, true>)); PRELOAD((k_density<true, true, 1, true>));
    PRELOAD((k_force<true, false, true, 9, true>)); PRELOAD((k_force<true, false, false, 9, true>));
    PRELOAD((k_force<true, false, true, 9, true, 1, true>)); PRELOAD((k_force<true, false, false, 9, true, 1, true>));
    PRELOAD((k_force<true, false, true, 9, true, 1, false, true>)); PRELOAD((k_force<true, false, false, 9, true, 1, false, true>));
    PRELOAD((k_force<true, false, true, 9, true, 1, true, true>)); PRELOAD((k_force<true, false, false, 9, true, 1, true, true>));
    PRELOAD((k_force<true, true, true, 9, true>)); PRELOAD((k_force<true, true, false, 9, true>));
    PRELOAD((k_force<true, true, true, 9, true, 1, true>)); PRELOAD((k_force<true, true, false, 9, true, 1, true>));
#undef PRELOAD
    return SPHB_OK;
}

// Everything the exchange allocates, allocated now: cudaMalloc / cudaMallocHost may wait for the whole device, which must not
// happen once a stream of this device sits in a flag wait of the peer-to-peer link.
int prepare_exchange(sphb_engine* e) {
    if (!e->h_status) {
        CK(cudaMallocHost((void**)&e->h_status, 16 * sizeof(int)));
        memset(e->h_status, 0, 16 * sizeof(int));
        for (int k = 0; k < 2; k++) CK(cudaEventCreateWithFlags(&e->ev_status[k], cudaEventDisableTiming));
    }
    const size_t cap = e->posm[0].cap ? e->posm[0].cap : 1;
    const size_t nb_layout = (cap + kMgTile - 1) / kMgTile;
    CK(e->mg_counts.reserve((size_t)kMgStreams * nb_layout + 16 + 2 * nb_layout));
    return SPHB_OK;
}
}

extern "C" int sphb_mg_unpack_async(sphb_engine* e, const sphb_slab* slab, const void* from_left, const void* from_right) {
    return unpack_async_impl(e, slab, from_left, from_right, nullptr);
}

namespace {
int unpack_async_impl(sphb_engine* e, const sphb_slab* slab, const void* from_left, const void* from_right, const int* link_err) {
    int rc = check_ready_hot(e);
    if (rc) return rc;
    if (!slab) return fail(SPHB_ERR_INVALID, "null slab");
    if (!e->mg_pending) return fail(SPHB_ERR_STATE, "sphb_mg_unpack_async without sphb_mg_pack");
    if ((rc = prepare_exchange(e))) return rc;
    const Slab s = to_slab(slab);
    set_need_box(e, slab);
    const int cur = e->cur;
    int* totals = e->mg_counts.p + kMgStreams * e->mg_blocks;
    int* recv = totals + 4;
    const int cap = (int)(e->posm[cur].cap < 0x7fffffffu ? e->posm[cur].cap : 0x7fffffffu);
    CK(cudaMemsetAsync(recv, 0, 4 * sizeof(int), e->stream));
    if (from_left) LAUNCH(e, k_mg_append, 148 * 4, 256, 0, s, (const float4*)from_left, totals, 0, e->posm[cur].p, e->vel[cur].p, e->global_ids ? e->oid[cur].p : (int*)nullptr, cap, recv);
    if (from_right) LAUNCH(e, k_mg_append, 148 * 4, 256, 0, s, (const float4*)from_right, totals, 1, e->posm[cur].p, e->vel[cur].p, e->global_ids ? e->oid[cur].p : (int*)nullptr, cap, recv);
    LAUNCH(e, k_mg_commit, 1, 32, 0, totals, recv, cap, e->d_cnt, e->d_status, link_err);
    CK(cudaGetLastError());
    const int slot = (int)(e->status_issued & 1);
    CK(cudaMemcpyAsync(e->h_status + 8 * slot, e->d_status, 8 * sizeof(int), cudaMemcpyDeviceToHost, e->stream));
    CK(cudaEventRecord(e->ev_status[slot], e->stream));
    e->status_issued++;
    e->mg_pending = false;
    // the host's estimate of the new count (launch sizing only): last exchange's total plus 3 % + 8192, or everything the arrays hold
    const size_t last = (size_t)e->last_status[3];
    size_t est = last ? last + last / 32 + 8192 : (size_t)cap;
    if (est > (size_t)cap) est = (size_t)cap;
    e->n = est;
    e->n_deferred = true;
    e->cnt_slot = 0;
    e->oid_identity = !e->global_ids; // ids restart at the slot numbers (materialised only if somebody asks), unless they are the caller's labels
    e->grid_valid = false;
    e->mask_valid = false;
    e->vr_valid = false;
    e->bbox_valid = false;
    e->box_records = 0;
    e->aos_valid = false;
    return SPHB_OK;
}
} // namespace

extern "C" int sphb_mg_poll(sphb_engine* e, size_t status[8], int wait) {
    if (!e || !status) return fail(SPHB_ERR_INVALID, "null argument");
    int rc = use_device(e);
    if (rc) return rc;
    rc = poll_status(e, wait != 0);
    for (int k = 0; k < 8; k++) status[k] = (size_t)e->last_status[k];
    status[7] = (size_t)e->status_seen;
    return rc;
}

extern "C" int sphb_mg_unpack(sphb_engine* e, const sphb_slab* slab, const void* from_left, const void* from_right, size_t counts[4]) {
    int rc = check_ready_hot(e);
    if (rc) return rc;
    if (!slab) return fail(SPHB_ERR_INVALID, "null slab");
    if (!e->mg_pending) return fail(SPHB_ERR_STATE, "sphb_mg_unpack without sphb_mg_pack");
    Slab s;
    s.lo = slab->lo; s.hi = slab->hi; s.halo = slab->halo;
    s.has_left = slab->has_left; s.has_right = slab->has_right;
    s.shift_left = slab->shift_left; s.shift_right = slab->shift_right;
    s.per_y = slab->periodic_y; s.per_z = slab->periodic_z;
    s.ylo = slab->ylo; s.ylen = slab->ylen; s.zlo = slab->zlo; s.zlen = slab->zlen;
    set_need_box(e, slab);
    const int cur = e->cur; // the kept particles stayed in place: everything is appended behind them
    int* totals = e->mg_counts.p + kMgStreams * e->mg_blocks;
    int* recv = totals + 4;
    const int cap = (int)(e->posm[cur].cap < 0x7fffffffu ? e->posm[cur].cap : 0x7fffffffu);
    CK(cudaMemsetAsync(recv, 0, 4 * sizeof(int), e->stream));
    if (from_left) LAUNCH(e, k_mg_append, 148 * 4, 256, 0, s, (const float4*)from_left, totals, 0, e->posm[cur].p, e->vel[cur].p, e->global_ids ? e->oid[cur].p : (int*)nullptr, cap, recv);
    if (from_right) LAUNCH(e, k_mg_append, 148 * 4, 256, 0, s, (const float4*)from_right, totals, 1, e->posm[cur].p, e->vel[cur].p, e->global_ids ? e->oid[cur].p : (int*)nullptr, cap, recv);
    CK(cudaGetLastError());
    int h[8];
    CK(cudaMemcpyAsync(h, totals, 8 * sizeof(int), cudaMemcpyDeviceToHost, e->stream));
    CK(cudaStreamSynchronize(e->stream));
    // h[0] = records present before the messages (old count + periodic images, the dead among them included), h[3] = dead
    const long long kept = h[0], cl = h[4], cr = h[5], total = kept + cl + cr;
    e->mg_pending = false;
    e->n_dead = (size_t)h[3]; // marked in place by the pack, whatever happens below
    e->mg_sent[0] = (size_t)h[1];
    e->mg_sent[1] = (size_t)h[2];
    // An overflow cannot be repaired here: the senders have already handed ownership over and marked their copies dead, so the
    // records that did not fit are gone.  The engine refuses to go on with a half-updated state (upload again, with more room).
    if (h[6] > cl || h[7] > cr) {
        e->have_state = false;
        return fail(SPHB_ERR_CAPACITY, "a neighbour overflowed its message (wanted %d / %d records): particles were lost, the state is dropped", h[6], h[7]);
    }
    if (total > cap) {
        e->have_state = false;
        return fail(SPHB_ERR_CAPACITY, "%lld particles exceed the reserved %d: call sphb_reserve with more room (the state is dropped)", total, cap);
    }
    e->n = (size_t)total;
    e->n_deferred = false; // the count is exact again
    e->last_status[3] = (int)total;
    e->oid_identity = !e->global_ids; // ids restart at the slot numbers (materialised only if somebody asks), unless they are the caller's labels
    e->grid_valid = false;
    e->mask_valid = false;
    e->vr_valid = false;
    e->bbox_valid = false;
    e->box_records = 0;
    e->aos_valid = false;
    if (counts) { counts[0] = (size_t)kept; counts[1] = (size_t)cl; counts[2] = (size_t)cr; counts[3] = (size_t)total; }
    return SPHB_OK;
}

// =====================================================================================================
// Peer-to-peer halo link (include/sphb200.h: sphb_mg_link_*, sphb_mg_exchange)

struct sphb_mg_link {
    sphb_engine* e = nullptr;
    size_t capacity = 0;
    size_t buf_bytes = 0;      // one receive buffer
    char* block = nullptr;     // [flags 256 B][recv_left[0]][recv_left[1]][recv_right[0]][recv_right[1]] on e's device
    char* left = nullptr;      // the neighbours' blocks as seen from e's device (peer mapping), or this link's own block
    char* right = nullptr;
    bool left_ipc = false, right_ipc = false;
    long long step = 0;
};

namespace {
inline int* link_flags(char* block) { return (int*)block; }
inline char* link_buf(const sphb_mg_link* l, char* block, int side /*0 = recv_left, 1 = recv_right*/, int parity) {
    return block + kLinkFlagInts * sizeof(int) + (size_t)(2 * side + parity) * l->buf_bytes;
}
} // namespace

extern "C" int sphb_mg_link_create(sphb_engine* e, size_t capacity, sphb_mg_link** out) {
    if (!e || !out) return fail(SPHB_ERR_INVALID, "null argument");
    if (capacity == 0 || capacity > 0x3fffffffu) return fail(SPHB_ERR_INVALID, "bad link capacity");
    int rc = use_device(e);
    if (rc) return rc;
    sphb_mg_link* l = new (std::nothrow) sphb_mg_link();
    if (!l) return fail(SPHB_ERR_INVALID, "out of host memory");
    l->e = e;
    l->capacity = capacity;
    l->buf_bytes = (sphb_mg_message_bytes(capacity) + 255) & ~(size_t)255;
    const size_t bytes = kLinkFlagInts * sizeof(int) + 4 * l->buf_bytes;
    cudaError_t err = cudaMalloc((void**)&l->block, bytes);
    if (err == cudaSuccess) err = cudaMemset(l->block, 0, bytes); // flags at 0, headers say "0 records"
    if (err != cudaSuccess) {
        if (l->block) cudaFree(l->block);
        delete l;
        return fail(SPHB_ERR_CUDA, "link allocation of %zu bytes failed: %s", bytes, cudaGetErrorString(err));
    }
    if ((rc = preload_hot_kernels()) || (rc = prepare_exchange(e))) { // with the arrays as reserved / uploaded so far
        cudaFree(l->block);
        delete l;
        return rc;
    }
    *out = l;
    return SPHB_OK;
}

extern "C" void sphb_mg_link_destroy(sphb_mg_link* l) {
    if (!l) return;
    cudaSetDevice(l->e->device);
    cudaStreamSynchronize(l->e->stream);
    if (l->left && l->left_ipc) cudaIpcCloseMemHandle(l->left);
    if (l->right && l->right_ipc && l->right != l->left) cudaIpcCloseMemHandle(l->right);
    if (l->block) cudaFree(l->block);
    delete l;
}

extern "C" int sphb_mg_link_export(sphb_mg_link* l, unsigned char handle[SPHB_MG_EXPORT_BYTES]) {
    if (!l || !handle) return fail(SPHB_ERR_INVALID, "null argument");
    static_assert(sizeof(cudaIpcMemHandle_t) <= SPHB_MG_EXPORT_BYTES, "export record too small");
    CK(cudaSetDevice(l->e->device));
    cudaIpcMemHandle_t h;
    CK(cudaIpcGetMemHandle(&h, l->block));
    memset(handle, 0, SPHB_MG_EXPORT_BYTES);
    memcpy(handle, &h, sizeof h);
    return SPHB_OK;
}

extern "C" int sphb_mg_link_connect_ipc(sphb_mg_link* l, const unsigned char* left, const unsigned char* right) {
    if (!l) return fail(SPHB_ERR_INVALID, "null link");
    CK(cudaSetDevice(l->e->device));
    const unsigned char* hs[2] = {left, right};
    char** dst[2] = {&l->left, &l->right};
    bool* ipc[2] = {&l->left_ipc, &l->right_ipc};
    for (int k = 0; k < 2; k++) {
        if (!hs[k]) continue;
        if (k == 1 && left && memcmp(left, right, sizeof(cudaIpcMemHandle_t)) == 0) { // two ranks in a ring: both neighbours are the same block
            l->right = l->left;
            l->right_ipc = false;
            continue;
        }
        cudaIpcMemHandle_t h;
        memcpy(&h, hs[k], sizeof h);
        void* p = nullptr;
        CK(cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess));
        *dst[k] = (char*)p;
        *ipc[k] = true;
    }
    return SPHB_OK;
}

extern "C" int sphb_mg_link_connect_local(sphb_mg_link* l, sphb_mg_link* left, sphb_mg_link* right) {
    if (!l) return fail(SPHB_ERR_INVALID, "null link");
    CK(cudaSetDevice(l->e->device));
    sphb_mg_link* nb[2] = {left, right};
    for (int k = 0; k < 2; k++) {
        if (!nb[k]) continue;
        if (nb[k]->capacity != l->capacity) return fail(SPHB_ERR_INVALID, "links of one chain must share their capacity");
        const int peer = nb[k]->e->device;
        if (peer != l->e->device) {
            int can = 0;
            CK(cudaDeviceCanAccessPeer(&can, l->e->device, peer));
            if (!can) return fail(SPHB_ERR_CUDA, "device %d cannot access device %d's memory", l->e->device, peer);
            cudaError_t err = cudaDeviceEnablePeerAccess(peer, 0);
            if (err == cudaErrorPeerAccessAlreadyEnabled) cudaGetLastError();
            else CK(err);
        }
    }
    l->left = left ? left->block : nullptr;
    l->right = right ? right->block : nullptr;
    l->left_ipc = l->right_ipc = false;
    return SPHB_OK;
}

extern "C" int sphb_mg_exchange(sphb_engine* e, sphb_mg_link* l, const sphb_slab* slab) {
    int rc = check_ready_hot(e);
    if (rc) return rc;
    if (!l || !slab) return fail(SPHB_ERR_INVALID, "null argument");
    if (l->e != e) return fail(SPHB_ERR_INVALID, "the link belongs to another engine");
    if ((slab->has_left && !l->left) || (slab->has_right && !l->right)) return fail(SPHB_ERR_STATE, "the link is not connected on a side the slab has a neighbour on");
    if ((rc = prepare_exchange(e))) return rc; // nothing below may allocate
    const long long s = l->step;
    const int p = (int)(s & 1);
    int* mine = link_flags(l->block);
    const bool hl = slab->has_left != 0, hr = slab->has_right != 0;
    // 1. the neighbours must have consumed what was written into these buffers two steps ago
    if (s >= 2) LAUNCH(e, k_link_wait, 1, 32, 0, hl ? mine + kLinkConsumedLeft : nullptr, (int)(s - 1), hr ? mine + kLinkConsumedRight : nullptr, (int)(s - 1), mine + kLinkError);
    // 2. pack straight into their receive buffers: my message to the left neighbour is what it receives from its right, and vice versa
    char* to_left = hl ? link_buf(l, l->left, 1, p) : nullptr;
    char* to_right = hr ? link_buf(l, l->right, 0, p) : nullptr;
    if ((rc = sphb_mg_pack_limits(e, slab, to_left, l->capacity, to_right, l->capacity))) return rc;
    // 3. tell them it is there
    LAUNCH(e, k_link_set, 1, 32, 0, hl ? link_flags(l->left) + kLinkArrivedRight : nullptr, (int)(s + 1), hr ? link_flags(l->right) + kLinkArrivedLeft : nullptr, (int)(s + 1));
    // 4. wait for theirs
    LAUNCH(e, k_link_wait, 1, 32, 0, hl ? mine + kLinkArrivedLeft : nullptr, (int)(s + 1), hr ? mine + kLinkArrivedRight : nullptr, (int)(s + 1), mine + kLinkError);
    // 5. append (nothing is read back: counts and verdict arrive one call late, sphb_mg_poll)
    if ((rc = unpack_async_impl(e, slab, hl ? link_buf(l, l->block, 0, p) : nullptr, hr ? link_buf(l, l->block, 1, p) : nullptr, mine + kLinkError))) return rc;
    // 6. their buffers of this parity may be reused
    LAUNCH(e, k_link_set, 1, 32, 0, hl ? link_flags(l->left) + kLinkConsumedRight : nullptr, (int)(s + 1), hr ? link_flags(l->right) + kLinkConsumedLeft : nullptr, (int)(s + 1));
    CK(cudaGetLastError());
    l->step = s + 1;
    return SPHB_OK;
}

// =====================================================================================================
// Allocation-free bounding box (include/sphb200.h: sphb_bbox_noalloc).  Self-contained: it shares no
// code path with the engine above.
//
// Scratch = one file-scope __device__ array.  In HOST code the name `g_noalloc_scratch` is only a
// stand-in for the symbol, not a device address: the address each device gave the array is fetched once
// per device with cudaGetSymbolAddress (that device current) and passed to the kernel as a plain pointer.

namespace {

constexpr int kNaThreads = 256;
constexpr int kNaPerBlock = 1024;                                   // points per block (4 per thread)
constexpr int kNaMaxBlocks = SPHB_NOALLOC_MAX_POINTS / kNaPerBlock; // 4096
constexpr int kNaPartialFloats = 8;                                 // 6 used, padded to 32 B
constexpr int kNaCtrlFloats = 64;                                   // 256-byte control block: word 0 = arrival counter
constexpr int kNaMaxDevices = 64;

static __device__ __align__(256) float g_noalloc_scratch[kNaCtrlFloats + kNaMaxBlocks * kNaPartialFloats];

float* g_noalloc_ptr[kNaMaxDevices]; // per-device address of the array above, 0 until looked up

// One block folds 1024 points into 6 numbers and parks them in the scratch; the block that arrives last
// folds all the parked partials into the caller's output.  `ctrl[0]` (arrival counter) must be zero when
// the kernel starts: the entry point clears it on the same stream before every launch.
__global__ void __launch_bounds__(kNaThreads) k_bbox_noalloc(const float4* __restrict__ posm, int n, float* scratch, float* __restrict__ out) {
    __shared__ float s_lo[3][kNaThreads / 32], s_hi[3][kNaThreads / 32];
    __shared__ bool s_last;
    unsigned int* counter = reinterpret_cast<unsigned int*>(scratch);
    float* partial = scratch + kNaCtrlFloats;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;

    float lo[3] = {INFINITY, INFINITY, INFINITY}, hi[3] = {-INFINITY, -INFINITY, -INFINITY};
    const int base = blockIdx.x * kNaPerBlock;
#pragma unroll
    for (int k = 0; k < kNaPerBlock / kNaThreads; k++) {
        const int i = base + k * kNaThreads + threadIdx.x;
        if (i < n) {
            const float4 p = __ldg(posm + i);
            lo[0] = fminf(lo[0], p.x); hi[0] = fmaxf(hi[0], p.x); // fminf/fmaxf ignore NaN operands
            lo[1] = fminf(lo[1], p.y); hi[1] = fmaxf(hi[1], p.y);
            lo[2] = fminf(lo[2], p.z); hi[2] = fmaxf(hi[2], p.z);
        }
    }
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
    if (threadIdx.x == 0) {
        float* mine = partial + (size_t)blockIdx.x * kNaPartialFloats;
        for (int a = 0; a < 3; a++) {
            float l = INFINITY, h = -INFINITY;
            for (int w = 0; w < kNaThreads / 32; w++) { l = fminf(l, s_lo[a][w]); h = fmaxf(h, s_hi[a][w]); }
            mine[a] = l;
            mine[3 + a] = h;
        }
        __threadfence();                                       // partials visible before the ticket is taken
        s_last = atomicAdd(counter, 1u) == gridDim.x - 1;
    }
    __syncthreads();
    if (!s_last) return;
    __threadfence();
    float l[3] = {INFINITY, INFINITY, INFINITY}, h[3] = {-INFINITY, -INFINITY, -INFINITY};
    for (int b = threadIdx.x; b < (int)gridDim.x; b += kNaThreads) {
        const volatile float* q = partial + (size_t)b * kNaPartialFloats;
#pragma unroll
        for (int a = 0; a < 3; a++) { l[a] = fminf(l[a], q[a]); h[a] = fmaxf(h[a], q[3 + a]); }
    }
#pragma unroll
    for (int a = 0; a < 3; a++) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            l[a] = fminf(l[a], __shfl_xor_sync(0xffffffffu, l[a], o));
            h[a] = fmaxf(h[a], __shfl_xor_sync(0xffffffffu, h[a], o));
        }
        if (lane == 0) { s_lo[a][warp] = l[a]; s_hi[a][warp] = h[a]; }
    }
    __syncthreads();
    if (threadIdx.x < 3) {
        const int a = threadIdx.x;
        float ll = INFINITY, hh = -INFINITY;
        for (int w = 0; w < kNaThreads / 32; w++) { ll = fminf(ll, s_lo[a][w]); hh = fmaxf(hh, s_hi[a][w]); }
        out[a] = ll;
        out[3 + a] = hh;
    }
    if (threadIdx.x == 0) *counter = 0u; // leave the counter clean for the next call as well
}

} // namespace

extern "C" int sphb_bbox_noalloc(const void* posm_device, size_t n, void* stream, float* bbox_device) {
    if (!posm_device || !bbox_device) return fail(SPHB_ERR_INVALID, "sphb_bbox_noalloc: null pointer");
    if (n == 0) return fail(SPHB_ERR_INVALID, "sphb_bbox_noalloc: no points");
    if (n > (size_t)SPHB_NOALLOC_MAX_POINTS)
        return fail(SPHB_ERR_CAPACITY, "sphb_bbox_noalloc: %zu points exceed the fixed scratch (at most %d); use an engine and sphb_get_bbox",
                    n, SPHB_NOALLOC_MAX_POINTS);
    int dev = -1;
    CK(cudaGetDevice(&dev));
    if (dev < 0 || dev >= kNaMaxDevices) return fail(SPHB_ERR_INVALID, "sphb_bbox_noalloc: device ordinal %d unsupported", dev);
    if (!g_noalloc_ptr[dev]) {
        // set-up, once per device, with that device current: ask the runtime where this device keeps the array
        float* ptr = nullptr;
        CK(cudaGetSymbolAddress((void**)&ptr, g_noalloc_scratch));
        if (!ptr) return fail(SPHB_ERR_CUDA, "sphb_bbox_noalloc: cudaGetSymbolAddress returned a null address");
        g_noalloc_ptr[dev] = ptr;
    }
    float* scratch = g_noalloc_ptr[dev];
    cudaStream_t st = (cudaStream_t)stream;
    // the array keeps whatever the previous call left: the arrival counter is the one word that must be zero
    CK(cudaMemsetAsync(scratch, 0, sizeof(unsigned int), st));
    const int blocks = (int)((n + kNaPerBlock - 1) / kNaPerBlock);
    k_bbox_noalloc<<<blocks, kNaThreads, 0, st>>>((const float4*)posm_device, (int)n, scratch, bbox_device);
    CK(cudaGetLastError());
    return SPHB_OK;
}
