// sphb_multi_*: one host thread driving several GPUs of one box (include/sphb200.h).  Nothing like it exists in the reference
// (one thread, one core: SPH/main.cpp:280-303).  Built purely on the library's own C ABI: one sphb_engine per GPU, x-slabs whose
// cuts equalise owned-plus-ghost particle counts, a peer-to-peer halo link between neighbours (sphb_mg_link: the pack kernel
// writes into the neighbour's receive buffer over NVLink), particle ids as labels so that results come back in the caller's
// order.  Per step and GPU: sphb_mg_exchange + sphb_step, all asynchronous -- the host thread only enqueues.
#include "../../include/sphb200.h"

#include <algorithm>
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <limits>
#include <new>
#include <vector>

namespace {

thread_local char g_merr[512] = "";

int mfail(int code, const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_merr, sizeof g_merr, fmt, ap);
    va_end(ap);
    return code;
}

// an error of the underlying engine call: keep its message
int efail(int rc) {
    snprintf(g_merr, sizeof g_merr, "%s", sphb_last_error());
    return rc;
}

} // namespace

struct sphb_multi {
    std::vector<int> devices;
    sphb_params params;
    std::vector<sphb_engine*> eng;
    std::vector<sphb_mg_link*> link;
    std::vector<sphb_slab> slab;
    std::vector<float> walls; // 6 floats each
    size_t nwalls = 0;
    std::vector<double> bounds;
    size_t n_total = 0;
    size_t capacity = 0;
    bool have_state = false;
};

namespace {

void drop_state(sphb_multi* m) {
    for (size_t r = 0; r < m->link.size(); r++)
        if (m->link[r]) sphb_mg_link_destroy(m->link[r]);
    m->link.clear();
    for (size_t r = 0; r < m->eng.size(); r++)
        if (m->eng[r]) sphb_destroy(m->eng[r]);
    m->eng.clear();
    m->have_state = false;
}

// Slab cuts: start at the particle-count quantiles of x and move every inner cut towards the heavier neighbour until what a
// rank WORKS on -- its own particles plus the ghost layers of width `halo` on both sides -- is even; snapped to multiples of
// `cell` from the left-most particle.  (The same rule as sph_b200/multigpu.py balanced_bounds.)
std::vector<double> balanced_bounds(std::vector<float> xs, int world, double cell, double halo) {
    std::sort(xs.begin(), xs.end());
    const size_t n = xs.size();
    const double lo = n ? xs.front() : 0.0, hi = n ? xs.back() : 0.0;
    std::vector<double> b(world + 1);
    b[0] = lo;
    b[world] = hi;
    for (int k = 1; k < world; k++) b[k] = n ? xs[std::min(n - 1, (size_t)((double)k / world * n))] : lo;
    auto count = [&](double a, double c) {
        return (double)(std::lower_bound(xs.begin(), xs.end(), (float)c) - std::lower_bound(xs.begin(), xs.end(), (float)a));
    };
    auto load = [&](const std::vector<double>& bb, int r) {
        return count(r ? bb[r] - halo : -1e300, r < world - 1 ? bb[r + 1] + halo : 1e300);
    };
    std::vector<double> best = b;
    double best_imb = 1e300;
    for (int it = 0; it < 40 && world > 1; it++) {
        std::vector<double> w(world);
        double mx = 0, mean = 0;
        for (int r = 0; r < world; r++) { w[r] = load(b, r); mx = std::max(mx, w[r]); mean += w[r] / world; }
        const double imb = mean > 0 ? mx / mean : 1.0;
        if (imb < best_imb) { best_imb = imb; best = b; }
        for (int k = 1; k < world; k++) {
            const double width = std::min(b[k] - b[k - 1], b[k + 1] - b[k]);
            const double tot = w[k] + w[k - 1];
            if (tot > 0) b[k] += 0.25 * width * (w[k] - w[k - 1]) / tot;
        }
        for (int k = 1; k < world; k++) b[k] = std::max(b[k], b[k - 1]);
    }
    for (int k = 1; k < world; k++) best[k] = lo + std::round((best[k] - lo) / cell) * cell;
    for (int k = 1; k < world; k++) best[k] = std::max(best[k], best[k - 1]);
    return best;
}

} // namespace

extern "C" {

const char* sphb_multi_last_error(void) { return g_merr; }

int sphb_multi_plan(const float* x, size_t n, int ndev, float h, double* bounds) {
    if (!bounds || (n && !x)) return mfail(SPHB_ERR_INVALID, "null argument");
    if (ndev < 1 || ndev > 64) return mfail(SPHB_ERR_INVALID, "ndev must be in [1, 64]");
    if (!(h > 0.0f)) return mfail(SPHB_ERR_INVALID, "h must be positive");
    const double R = (double)(h * h);
    const std::vector<double> b = balanced_bounds(std::vector<float>(x, x + n), ndev, R, 2.0 * R * 1.001);
    for (int k = 0; k <= ndev; k++) bounds[k] = b[k];
    return SPHB_OK;
}

int sphb_multi_create(int ndev, const int* devices, const sphb_params* p, sphb_multi** out) {
    if (!p || !out) return mfail(SPHB_ERR_INVALID, "null argument");
    if (ndev < 1 || ndev > 64) return mfail(SPHB_ERR_INVALID, "ndev must be in [1, 64]");
    sphb_multi* m = new (std::nothrow) sphb_multi();
    if (!m) return mfail(SPHB_ERR_INVALID, "out of host memory");
    m->params = *p;
    for (int r = 0; r < ndev; r++) m->devices.push_back(devices ? devices[r] : r);
    // check the devices now rather than at the first upload
    for (int r = 0; r < ndev; r++) {
        sphb_engine* e = nullptr;
        int rc = sphb_create(p, m->devices[r], &e);
        if (rc) {
            delete m;
            return efail(rc);
        }
        sphb_destroy(e);
    }
    *out = m;
    g_merr[0] = 0;
    return SPHB_OK;
}

void sphb_multi_destroy(sphb_multi* m) {
    if (!m) return;
    drop_state(m);
    delete m;
}

int sphb_multi_set_walls(sphb_multi* m, const void* walls, size_t count, size_t stride) {
    if (!m) return mfail(SPHB_ERR_INVALID, "null handle");
    if (count > SPHB_MAX_WALLS) return mfail(SPHB_ERR_CAPACITY, "at most %d walls", SPHB_MAX_WALLS);
    if (count && (!walls || stride < 24)) return mfail(SPHB_ERR_INVALID, "walls need >= 24-byte records");
    m->walls.resize(6 * count);
    for (size_t w = 0; w < count; w++) memcpy(&m->walls[6 * w], (const char*)walls + w * stride, 24);
    m->nwalls = count;
    for (size_t r = 0; r < m->eng.size(); r++) {
        int rc = sphb_set_walls(m->eng[r], m->walls.data(), m->nwalls, 24);
        if (rc) return efail(rc);
    }
    return SPHB_OK;
}

int sphb_multi_upload_aos(sphb_multi* m, const void* particles, size_t n, size_t stride) {
    if (!m) return mfail(SPHB_ERR_INVALID, "null handle");
    if (n && !particles) return mfail(SPHB_ERR_INVALID, "null particles");
    if (stride < 48 || stride % 4 || stride > 128) return mfail(SPHB_ERR_INVALID, "stride must be a multiple of 4 in [48, 128]");
    if (n > 0x7f000000u) return mfail(SPHB_ERR_CAPACITY, "too many particles");
    drop_state(m);
    const int world = (int)m->devices.size();
    const double R = (double)(m->params.h * m->params.h);
    const double halo = 2.0 * R * 1.001; // two support radii: ghost densities are recomputed locally, one exchange per step
    std::vector<float> xs(n);
    for (size_t i = 0; i < n; i++) memcpy(&xs[i], (const char*)particles + i * stride, sizeof(float));
    m->bounds = balanced_bounds(xs, world, R, halo);
    // which rank owns which particle, and how much may cross a face in one step (message capacity)
    std::vector<std::vector<int32_t> > ids(world);
    std::vector<size_t> near_face(world + 1, 0);
    for (size_t i = 0; i < n; i++) {
        const double x = xs[i];
        int r = 0;
        while (r < world - 1 && !(x < m->bounds[r + 1])) r++; // NaN stays with rank 0
        ids[r].push_back((int32_t)i);
        for (int k = 1; k < world; k++)
            if (std::fabs(x - m->bounds[k]) < 2.0 * halo) near_face[k]++;
    }
    size_t cap = 0;
    for (int k = 1; k < world; k++) cap = std::max(cap, near_face[k]);
    m->capacity = 2 * cap + 65536;
    m->eng.assign(world, nullptr);
    m->link.assign(world, nullptr);
    m->slab.assign(world, sphb_slab());
    std::vector<char> share;
    const float inf = std::numeric_limits<float>::infinity();
    for (int r = 0; r < world; r++) {
        int rc = sphb_create(&m->params, m->devices[r], &m->eng[r]);
        if (rc) return efail(rc);
        if ((rc = sphb_set_walls(m->eng[r], m->walls.data(), m->nwalls, 24))) return efail(rc);
        const size_t nr = ids[r].size();
        // own particles + this step's and last step's ghosts (the dropped ones stay, marked dead, until the next sort) + slack for migration
        if ((rc = sphb_reserve(m->eng[r], nr + nr / 4 + 4 * m->capacity))) return efail(rc);
        share.resize((nr ? nr : 1) * stride);
        for (size_t k = 0; k < nr; k++) memcpy(&share[k * stride], (const char*)particles + (size_t)ids[r][k] * stride, stride);
        if ((rc = sphb_upload_aos(m->eng[r], share.data(), nr, stride))) return efail(rc);
        if ((rc = sphb_synchronize(m->eng[r]))) return efail(rc); // `share` is reused for the next rank
        if ((rc = sphb_set_ids(m->eng[r], ids[r].data(), nr))) return efail(rc);
        sphb_slab& s = m->slab[r];
        memset(&s, 0, sizeof s);
        s.lo = r > 0 ? (float)m->bounds[r] : -inf;
        s.hi = r < world - 1 ? (float)m->bounds[r + 1] : inf;
        s.halo = (float)halo;
        s.has_left = r > 0;
        s.has_right = r < world - 1;
    }
    if (world > 1) {
        for (int r = 0; r < world; r++) {
            int rc = sphb_mg_link_create(m->eng[r], m->capacity, &m->link[r]);
            if (rc) return efail(rc);
        }
        for (int r = 0; r < world; r++) {
            int rc = sphb_mg_link_connect_local(m->link[r], r > 0 ? m->link[r - 1] : nullptr, r < world - 1 ? m->link[r + 1] : nullptr);
            if (rc) return efail(rc);
        }
    }
    m->n_total = n;
    m->have_state = true;
    return SPHB_OK;
}

int sphb_multi_step(sphb_multi* m, int nsteps) {
    if (!m) return mfail(SPHB_ERR_INVALID, "null handle");
    if (!m->have_state) return mfail(SPHB_ERR_STATE, "no particles uploaded");
    if (nsteps < 0) return mfail(SPHB_ERR_INVALID, "nsteps < 0");
    const int world = (int)m->eng.size();
    for (int s = 0; s < nsteps; s++) {
        if (world > 1)
            for (int r = 0; r < world; r++) { // every rank's exchange is enqueued before any rank's step: nobody waits on the host
                int rc = sphb_mg_exchange(m->eng[r], m->link[r], &m->slab[r]);
                if (rc) return efail(rc);
            }
        for (int r = 0; r < world; r++) {
            int rc = sphb_step(m->eng[r], 1);
            if (rc) return efail(rc);
        }
    }
    return SPHB_OK;
}

int sphb_multi_synchronize(sphb_multi* m) {
    if (!m) return mfail(SPHB_ERR_INVALID, "null handle");
    for (size_t r = 0; r < m->eng.size(); r++) {
        int rc = sphb_synchronize(m->eng[r]);
        if (rc) return efail(rc);
    }
    return SPHB_OK;
}

int sphb_multi_download_aos(sphb_multi* m, void* particles, size_t n, size_t stride, uint32_t field_mask) {
    if (!m) return mfail(SPHB_ERR_INVALID, "null handle");
    if (!m->have_state) return mfail(SPHB_ERR_STATE, "no particles uploaded");
    if (n != m->n_total) return mfail(SPHB_ERR_INVALID, "download of %zu particles but %zu were uploaded", n, m->n_total);
    if (n && !particles) return mfail(SPHB_ERR_INVALID, "null particles");
    if (stride < 48 || stride % 4 || stride > 128) return mfail(SPHB_ERR_INVALID, "stride must be a multiple of 4 in [48, 128]");
    std::vector<float> posm, vel, aux, acc;
    std::vector<int32_t> ids;
    size_t seen = 0;
    for (size_t r = 0; r < m->eng.size(); r++) {
        const size_t nr = sphb_count(m->eng[r]); // fetches the count (synchronises that GPU)
        posm.resize(4 * nr + 4); vel.resize(4 * nr + 4); aux.resize(4 * nr + 4); acc.resize(4 * nr + 4); ids.resize(nr + 1);
        int rc = sphb_download_soa(m->eng[r], posm.data(), vel.data(), (field_mask & SPHB_F_DENSITY) ? aux.data() : nullptr,
                                   (field_mask & SPHB_F_ACCEL) ? acc.data() : nullptr, ids.data());
        if (rc) return efail(rc);
        for (size_t k = 0; k < nr; k++) {
            if (vel[4 * k + 3] == 0.0f || posm[4 * k + 3] < 0.0f) continue; // a ghost of another slab, or a dropped record
            const size_t id = (size_t)(uint32_t)ids[k];
            if (id >= n) return mfail(SPHB_ERR_STATE, "rank %zu holds a particle labelled %zu of %zu", r, id, n);
            float* rec = (float*)((char*)particles + id * stride);
            if (field_mask & SPHB_F_POS) memcpy(rec + 0, &posm[4 * k], 12);
            if (field_mask & SPHB_F_VEL) memcpy(rec + 3, &vel[4 * k], 12);
            if (field_mask & SPHB_F_ACCEL) memcpy(rec + 6, &acc[4 * k], 12);
            if (field_mask & SPHB_F_DENSITY) memcpy(rec + 10, &aux[4 * k], 8);
            seen++;
        }
    }
    if (seen != n) return mfail(SPHB_ERR_STATE, "%zu of %zu particles are owned by some rank: the exchange lost or duplicated particles", seen, n);
    return SPHB_OK;
}

int sphb_multi_step_aos(sphb_multi* m, void* particles, size_t n, size_t stride, int nsteps, uint32_t field_mask) {
    int rc = sphb_multi_upload_aos(m, particles, n, stride);
    if (rc) return rc;
    if ((rc = sphb_multi_step(m, nsteps))) return rc;
    return sphb_multi_download_aos(m, particles, n, stride, field_mask);
}

int sphb_multi_info(sphb_multi* m, int* ndev, size_t* owned_plus_ghosts, double* bounds) {
    if (!m) return mfail(SPHB_ERR_INVALID, "null handle");
    if (ndev) *ndev = (int)m->devices.size();
    for (size_t r = 0; r < m->eng.size(); r++)
        if (owned_plus_ghosts) owned_plus_ghosts[r] = sphb_count(m->eng[r]);
    if (bounds)
        for (size_t k = 0; k < m->bounds.size(); k++) bounds[k] = m->bounds[k];
    return SPHB_OK;
}

} // extern "C"
