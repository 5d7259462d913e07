/*
 * TEST INFRASTRUCTURE ONLY.  CPU restatement (plain C) of the reference SPH step, used as the
 * parity checker by tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
 * legs.  Nothing under sph_b200/ (the product) may import, link or call this file.
 *
 * Parity status: PINNED.  Every function below is checked bit-for-bit against the unmodified
 * reference (oracle/_ref/libsphref.so, built by oracle/Makefile from /root/reference/SPH/SphUtils.cpp
 * where it lies) by tests/test_oracle_vs_ref.py, and against the committed golden vectors under
 * tests/golden/ (generated from that same library by tests/golden/make_golden.py).
 *
 * Build: gcc -O2 -ffp-contract=off -fno-fast-math (no -march: the reference is pinned as built by
 * g++ -O2 without FMA contraction).
 *
 * All file:line citations are into /root/reference/ (SPH/..., deps/include/glm/...).
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

/* SPH/Particle.hpp:42-50 -- 52 bytes, all public floats */
typedef struct {
    float pos[3];
    float vel[3];
    float accel[3];
    float mass, density, pressure, thermal;
} orc_particle;

/* SPH/Wall.hpp:28-31 -- normal first, then center; lengths are draw-only */
typedef struct {
    float normal[3];
    float center[3];
    float xlength, ylength;
} orc_wall;

/* SPH/Sphere.hpp:29-33 */
typedef struct {
    float pos[3];
    float vel[3];
    float accel[3];
    float radius, mass;
} orc_sphere;

/* SPH/SphUtils.hpp:20-29 ctor arguments (c and epsilon are stored and never read) */
typedef struct {
    float h, rho0, c, dt, eps;
} orc_params;

/* glm::pi<float>() -- deps/include/glm/gtc/constants.inl: T(3.14159265358979323846264338327950288) */
static const float ORC_PI_F = (float)3.14159265358979323846264338327950288;

/* ---- GLM 0.9.6.1 vec3 arithmetic, restated -------------------------------------------------- */

/* deps/include/glm/detail/func_geometric.inl:64-72: tmp = x*y (componentwise); tmp.x + tmp.y + tmp.z */
static inline float g_dot(const float a[3], const float b[3]) {
    float tx = a[0] * b[0], ty = a[1] * b[1], tz = a[2] * b[2];
    return tx + ty + tz;
}
static inline void g_sub(const float a[3], const float b[3], float o[3]) {
    o[0] = a[0] - b[0]; o[1] = a[1] - b[1]; o[2] = a[2] - b[2];
}
/* func_geometric.inl:95-99 */
static inline float g_length(const float v[3]) { return sqrtf(g_dot(v, v)); }
/* func_geometric.inl:112-114: distance(p0,p1) = length(p1 - p0) */
static inline float g_distance(const float p0[3], const float p1[3]) {
    float d[3];
    g_sub(p1, p0, d);
    return g_length(d);
}
/* func_geometric.inl:154-158 + func_exponential.inl:148-153: x * (1 / sqrt(dot(x,x))) */
static inline void g_normalize(const float v[3], float o[3]) {
    float inv = 1.0f / sqrtf(g_dot(v, v));
    o[0] = v[0] * inv; o[1] = v[1] * inv; o[2] = v[2] * inv;
}

/* ---- kernels: SPH/SphUtils.cpp:55-95 --------------------------------------------------------- */

/* SphUtils.cpp:55-63.  q = |j-i|/h compared against h itself (support radius h*h); the cube and the
 * divisions run in double because pow(float,int) promotes; one rounding to float at the end. */
float orc_kernel_function(const orc_params* p, const float i[3], const float j[3]) {
    float h = p->h;
    float q = g_distance(i, j) / h;
    float result = 0.0f;
    if (q <= h) {
        float t = h * h - q * q;
        float den_f = 64 * ORC_PI_F; /* int * float -> float */
        result = (float)(315 * pow((double)t, 3) / ((double)den_f * pow((double)h, 9)));
    }
    return result;
}

/* SphUtils.cpp:68-79.  strict '<'; numerator in float, division in double; normalize(i-j) is NaN for
 * coincident points and that NaN is kept (0*NaN). */
void orc_grad_kernel(const orc_params* p, const float i[3], const float j[3], float out[3]) {
    float h = p->h;
    float q = g_distance(i, j) / h;
    float result = 0.0f;
    if (q < h) {
        float num = -45.0f * (h - q) * (h - q);
        result = (float)((double)num / ((double)ORC_PI_F * pow((double)h, 6)));
    }
    float d[3], g[3];
    g_sub(i, j, d);
    g_normalize(d, g);
    out[0] = g[0] * result; out[1] = g[1] * result; out[2] = g[2] * result;
}

/* SphUtils.cpp:84-95.  45/((float)pi*pow(h,6)) is int/double -> double; times float -> double -> float */
float orc_lap_kernel(const orc_params* p, const float i[3], const float j[3]) {
    float h = p->h;
    float q = g_distance(i, j) / h;
    float result = 0.0f;
    if (q < h) result = (float)(45 / ((double)ORC_PI_F * pow((double)h, 6)) * (double)(h - q));
    return result;
}

/* ---- pair bodies ----------------------------------------------------------------------------- */

/* SphUtils.cpp:133 */
static inline float density_term(const orc_params* p, const orc_particle* a, const orc_particle* b) {
    return b->mass * orc_kernel_function(p, a->pos, b->pos);
}

/* SphUtils.cpp:141-144 */
static inline void density_finish(const orc_params* p, orc_particle* a, float density) {
    if (density < p->rho0) density = p->rho0;
    a->density = density;
    a->pressure = 100.0f * (a->density - p->rho0);
}

/* SphUtils.cpp:219-230, one (i,j) pair added into acc[3] */
static inline void pressure_pair(const orc_params* p, const orc_particle* a, const orc_particle* b, float acc[3]) {
    float h = p->h;
    float pi = 0.0f;
    float dv[3], dx[3];
    g_sub(a->vel, b->vel, dv);
    g_sub(a->pos, b->pos, dx);
    float vdotr = g_dot(dv, dx);
    if (vdotr < 0) {
        float c = (sqrtf(1.0f * fabsf(a->pressure / a->density)) + sqrtf(1.0f * fabsf(b->pressure / b->density))) / 2.0f;
        float r = g_length(dx);
        float mu = h * vdotr / ((r * r) + 0.01f * h * h);
        pi = 2.0f * (-0.1f * c * mu + 0.2f * mu * mu) / (a->density + b->density);
    }
    float temp = -b->mass * (pi + (a->pressure + b->pressure) / (2.0f * b->density * a->density));
    float g[3];
    orc_grad_kernel(p, a->pos, b->pos, g);
    acc[0] += temp * g[0];
    acc[1] += temp * g[1];
    acc[2] += temp * g[2];
}

/* ---- all-pairs phases, exactly the loops the reference runs ---------------------------------- */

/* SphUtils.cpp:100-149 */
void orc_update_density_allpairs(const orc_params* p, orc_particle* ps, size_t n) {
    for (size_t i = 0; i < n; i++) {
        float density = 0.0f;
        for (size_t j = 0; j < n; j++)
            if (i != j) density += density_term(p, &ps[i], &ps[j]);
        density_finish(p, &ps[i], density);
    }
}

/* SphUtils.cpp:157-176 (zero, gravity :295-304, pressure_forces :181-239; viscosity call is commented out) */
void orc_update_forces_allpairs(const orc_params* p, orc_particle* ps, size_t n) {
    for (size_t i = 0; i < n; i++) {
        float acc[3] = {0.0f, 0.0f, 0.0f};
        acc[0] += 0.0f; acc[1] += -1.0f; acc[2] += 0.0f; /* glm::vec3(0.0, -1, 0.0) */
        for (size_t j = 0; j < n; j++)
            if (i != j) pressure_pair(p, &ps[i], &ps[j], acc);
        ps[i].accel[0] = acc[0]; ps[i].accel[1] = acc[1]; ps[i].accel[2] = acc[2];
    }
}

/* SphUtils.cpp:277-278, the pair body of the dead viscosity_forces: accel_i += 0.1 m_j / (rho_i rho_j) lapW (v_i - v_j).
 * (The sign spreads velocity differences instead of damping them; it is restated as written.) */
static inline void viscosity_pair(const orc_params* p, const orc_particle* a, const orc_particle* b, float acc[3]) {
    float temp = 0.1f * b->mass / (a->density * b->density) * orc_lap_kernel(p, a->pos, b->pos);
    acc[0] += temp * (a->vel[0] - b->vel[0]);
    acc[1] += temp * (a->vel[1] - b->vel[1]);
    acc[2] += temp * (a->vel[2] - b->vel[2]);
}

/* viscosity_forces over ALL pairs in ascending index: what SphUtils.cpp:244-290 computes when every particle lies
 * in one of its cells (pinned against the reference in that situation); adds into accel. */
void orc_viscosity_forces_allpairs(const orc_params* p, orc_particle* ps, size_t n) {
    for (size_t i = 0; i < n; i++) {
        float acc[3] = {ps[i].accel[0], ps[i].accel[1], ps[i].accel[2]};
        for (size_t j = 0; j < n; j++)
            if (i != j) viscosity_pair(p, &ps[i], &ps[j], acc);
        ps[i].accel[0] = acc[0]; ps[i].accel[1] = acc[1]; ps[i].accel[2] = acc[2];
    }
}

/* Sphere side of update_forces (:163-166,300-303): accel = 0 then += (0,-1,0) */
void orc_update_forces_spheres(orc_sphere* s, size_t ns) {
    for (size_t i = 0; i < ns; i++) {
        s[i].accel[0] = 0.0f + 0.0f; s[i].accel[1] = 0.0f + -1.0f; s[i].accel[2] = 0.0f + 0.0f;
    }
}

/* SphUtils.cpp:321-335: vel = vel + dt*accel; pos = pos + dt*vel (the new vel) */
void orc_update_posvel(const orc_params* p, orc_particle* ps, size_t n, orc_sphere* s, size_t ns) {
    float dt = p->dt;
    for (size_t i = 0; i < n; i++)
        for (int a = 0; a < 3; a++) {
            ps[i].vel[a] = ps[i].vel[a] + dt * ps[i].accel[a];
            ps[i].pos[a] = ps[i].pos[a] + dt * ps[i].vel[a];
        }
    for (size_t i = 0; i < ns; i++)
        for (int a = 0; a < 3; a++) {
            s[i].vel[a] = s[i].vel[a] + dt * s[i].accel[a];
            s[i].pos[a] = s[i].pos[a] + dt * s[i].vel[a];
        }
}

/* SphUtils.cpp:343-387.  Walls in vector order, later walls see the corrected pos/vel; the hit test is
 * evaluated in double (0.001 and 0.01 are double literals).  Spheres: serial, sphere velocity mutates
 * inside the particle loop. */
void orc_collision(const orc_params* p, orc_particle* ps, size_t n, const orc_wall* w, size_t nw, orc_sphere* s, size_t ns) {
    (void)p;
    for (size_t i = 0; i < n; i++) {
        for (size_t j = 0; j < nw; j++) {
            float rel[3];
            g_sub(ps[i].pos, w[j].center, rel);
            float dot = g_dot(w[j].normal, rel);
            if ((double)dot + 0.001 <= 0.01 || (double)dot - 0.001 <= 0.01) {
                float imp[3];
                g_normalize(w[j].normal, imp);
                float k = (1.0f + 0.0f) * g_dot(w[j].normal, ps[i].vel);
                imp[0] *= k; imp[1] *= k; imp[2] *= k;
                if (dot < 0) dot = -dot;
                for (int a = 0; a < 3; a++) ps[i].pos[a] += dot * w[j].normal[a];
                for (int a = 0; a < 3; a++) ps[i].vel[a] += -imp[a];
            }
        }
        for (size_t j = 0; j < ns; j++) {
            float rel[3];
            g_sub(ps[i].pos, s[j].pos, rel);
            float dist = g_length(rel);
            if ((double)dist <= 0.1 + (double)s[j].radius) {
                float u1[3], u2[3];
                memcpy(u1, ps[i].vel, sizeof u1);
                memcpy(u2, s[j].vel, sizeof u2);
                float mp = ps[i].mass, ms = s[j].mass;
                for (int a = 0; a < 3; a++)
                    ps[i].vel[a] = (u1[a] * (mp - ms) + (2.0f * ms) * u2[a]) / (mp + ms);
                for (int a = 0; a < 3; a++)
                    s[j].vel[a] = (u2[a] * (ms - mp) + (2.0f * mp) * u1[a]) / (mp + ms);
            }
        }
    }
}

/* main.cpp:293-300 call order */
void orc_step_allpairs(const orc_params* p, orc_particle* ps, size_t n, const orc_wall* w, size_t nw, int nsteps) {
    for (int k = 0; k < nsteps; k++) {
        orc_update_density_allpairs(p, ps, n);
        orc_update_forces_allpairs(p, ps, n);
        orc_update_posvel(p, ps, n, NULL, 0);
        orc_collision(p, ps, n, w, nw, NULL, 0);
    }
}

/* ---- reference update_cells (dead code in the reference, SphUtils.cpp:8-47), restated ------- */

/* bbox (:11-23), dims (:25-27, cell = 2h :14), cell coords (:38-40) and the flat index with its stray
 * "+ny" (:43).  Returns nx*ny*nz; cell_of[m] = flat index, or -1-flat if it falls outside the
 * allocation (where the reference itself would write out of bounds). */
long orc_ref_update_cells(const orc_params* p, const orc_particle* ps, size_t n, float bbox[6], int dims[3], int* cell_of) {
    float minx, miny, minz, maxx, maxy, maxz;
    maxx = minx = ps[0].pos[0];
    maxy = miny = ps[0].pos[1];
    maxz = minz = ps[0].pos[2];
    float cell = 2 * p->h;
    for (size_t i = 0; i < n; i++) {
        if (ps[i].pos[0] < minx) minx = ps[i].pos[0];
        if (ps[i].pos[1] < miny) miny = ps[i].pos[1];
        if (ps[i].pos[2] < minz) minz = ps[i].pos[2];
        if (ps[i].pos[0] > maxx) maxx = ps[i].pos[0];
        if (ps[i].pos[1] > maxy) maxy = ps[i].pos[1];
        if (ps[i].pos[2] > maxz) maxz = ps[i].pos[2];
    }
    int nx = (int)ceilf((maxx - minx) / cell) + 1;
    int ny = (int)ceilf((maxy - miny) / cell) + 1;
    int nz = (int)ceilf((maxz - minz) / cell) + 1;
    bbox[0] = minx; bbox[1] = miny; bbox[2] = minz; bbox[3] = maxx; bbox[4] = maxy; bbox[5] = maxz;
    dims[0] = nx; dims[1] = ny; dims[2] = nz;
    long ncell = (long)nx * ny * nz;
    for (size_t m = 0; m < n; m++) {
        int x = (int)floorf((ps[m].pos[0] - minx) / cell);
        int y = (int)floorf((ps[m].pos[1] - miny) / cell);
        int z = (int)floorf((ps[m].pos[2] - minz) / cell);
        unsigned int i = x + nx * y + (nx * ny + ny) * z;
        cell_of[m] = ((long)i < ncell) ? (int)i : -1 - (int)i;
    }
    return ncell;
}

/* ---- the engine's grid, CPU build ("CPU reference build of the same grid") ------------------- */

/* Grid spec shared with sph_b200/csrc (include/sphb200.h, sphb_get_cells):
 *   coord_a = clamp((int)floorf((pos_a - origin_a) / cell), 0, dims_a - 1)      (IEEE float sub and div,
 *             the same expression as SphUtils.cpp:38-40)
 *   key     = x + nx*(y + ny*z)                        (the standard index, not the reference's :43)
 *   order   = stable counting sort by key: ascending original index inside a cell
 * dims from a bounding box follow SphUtils.cpp:25-27: (int)ceilf((max-min)/cell) + 1, then, if the
 * product exceeds max_cells, the largest dimension (ties: x, then y, then z) is halved (rounding up)
 * until it fits; clamping the coordinates keeps every pair within one cell of each other. */
typedef struct {
    float origin[3];
    float cell[3]; /* edge per axis: the engine's cells are shorter in x (cell_x = cell / cell_x_div) */
    int dims[3];
} orc_grid;

void orc_grid_from_bbox(const float bbox[6], const float cell[3], long long max_cells, orc_grid* g) {
    for (int a = 0; a < 3; a++) {
        g->origin[a] = bbox[a];
        g->cell[a] = cell[a];
        float ext = ceilf((bbox[3 + a] - bbox[a]) / cell[a]);
        g->dims[a] = (ext < 1048575.0f) ? (int)ext + 1 : 1048576; /* cap also catches inf/NaN extents */
        if (g->dims[a] < 1) g->dims[a] = 1;
    }
    while ((long long)g->dims[0] * g->dims[1] * g->dims[2] > max_cells) {
        int a = 0;
        if (g->dims[1] > g->dims[a]) a = 1;
        if (g->dims[2] > g->dims[a]) a = 2;
        if (g->dims[a] == 1) break;
        g->dims[a] = (g->dims[a] + 1) / 2;
    }
    while ((long long)g->dims[1] * g->dims[2] > (1 << 22)) { /* same policy for the number of (y,z) rows */
        int a = g->dims[2] > g->dims[1] ? 2 : 1;
        g->dims[a] = (g->dims[a] + 1) / 2;
    }
}

void orc_bbox(const orc_particle* ps, size_t n, float bbox[6]) {
    for (int a = 0; a < 3; a++) bbox[a] = bbox[3 + a] = ps[0].pos[a];
    for (size_t i = 0; i < n; i++)
        for (int a = 0; a < 3; a++) {
            if (ps[i].pos[a] < bbox[a]) bbox[a] = ps[i].pos[a];
            if (ps[i].pos[a] > bbox[3 + a]) bbox[3 + a] = ps[i].pos[a];
        }
}

static inline int grid_coord(const orc_grid* g, float x, int a) {
    float f = floorf((x - g->origin[a]) / g->cell[a]);
    /* clamp in float first so that far-away strays (or NaN) cannot overflow the int conversion */
    if (!(f >= 0.0f)) return 0;
    if (f >= (float)g->dims[a]) return g->dims[a] - 1;
    return (int)f;
}

/* cell_xyz[3n] (original order), key[n] (original order), perm[n] (sorted slot -> original index),
 * cell_start[ncell+1].  Any output may be NULL. */
void orc_grid_build(const orc_grid* g, const orc_particle* ps, size_t n, int* cell_xyz, int* key, int* perm, int* cell_start) {
    long long ncell = (long long)g->dims[0] * g->dims[1] * g->dims[2];
    int* k = key ? key : (int*)malloc(sizeof(int) * (n ? n : 1));
    int* cs = cell_start ? cell_start : (int*)malloc(sizeof(int) * (ncell + 1));
    memset(cs, 0, sizeof(int) * (ncell + 1));
    for (size_t i = 0; i < n; i++) {
        int x = grid_coord(g, ps[i].pos[0], 0), y = grid_coord(g, ps[i].pos[1], 1), z = grid_coord(g, ps[i].pos[2], 2);
        if (cell_xyz) { cell_xyz[3 * i] = x; cell_xyz[3 * i + 1] = y; cell_xyz[3 * i + 2] = z; }
        k[i] = x + g->dims[0] * (y + g->dims[1] * z);
        cs[k[i] + 1]++;
    }
    for (long long c = 0; c < ncell; c++) cs[c + 1] += cs[c];
    if (perm) {
        int* cur = (int*)malloc(sizeof(int) * (ncell ? ncell : 1));
        memcpy(cur, cs, sizeof(int) * ncell);
        for (size_t i = 0; i < n; i++) perm[cur[k[i]]++] = (int)i;
        free(cur);
    }
    if (!key) free(k);
    if (!cell_start) free(cs);
}

/* ---- cell-list phases: same pair bodies, neighbours visited in ascending ORIGINAL index ------ */
/* Bit-identical to the all-pairs loops: a pair outside the support contributes exactly +0 to the
 * density sum (W = 0.0f) and +-0 to the acceleration sum (grad = n * 0.0f), which leave a float
 * accumulator unchanged, so dropping those pairs and keeping the order of the rest changes no bit.
 * (Pinned against the reference in tests/test_oracle_vs_ref.py.) */

typedef struct {
    orc_grid g;
    int* perm;
    int* cell_start;
} cellctx;

static void cellctx_init(cellctx* c, const orc_params* p, const orc_particle* ps, size_t n) {
    float bbox[6];
    orc_bbox(ps, n, bbox);
    /* a hair wider than the support so float rounding in the coordinate cannot hide a pair */
    float cell1 = p->h * p->h * 1.001f;
    float cell[3] = {cell1, cell1, cell1};
    orc_grid_from_bbox(bbox, cell, (long long)4 * (long long)n + 4096, &c->g);
    long long ncell = (long long)c->g.dims[0] * c->g.dims[1] * c->g.dims[2];
    c->perm = (int*)malloc(sizeof(int) * (n ? n : 1));
    c->cell_start = (int*)malloc(sizeof(int) * (ncell + 1));
    orc_grid_build(&c->g, ps, n, NULL, NULL, c->perm, c->cell_start);
}
static void cellctx_free(cellctx* c) { free(c->perm); free(c->cell_start); }

static int cmp_int(const void* a, const void* b) {
    int x = *(const int*)a, y = *(const int*)b;
    return (x > y) - (x < y);
}

/* candidate neighbours of i (superset of the support, self excluded), ascending original index */
static int gather_sorted(const cellctx* c, const orc_params* p, const orc_particle* ps, size_t i, int** buf, int* cap) {
    const orc_grid* g = &c->g;
    int cx = grid_coord(g, ps[i].pos[0], 0), cy = grid_coord(g, ps[i].pos[1], 1), cz = grid_coord(g, ps[i].pos[2], 2);
    float R = p->h * p->h;
    float lim = R * R * 1.001f;
    int cnt = 0;
    for (int z = cz - 1; z <= cz + 1; z++) {
        if (z < 0 || z >= g->dims[2]) continue;
        for (int y = cy - 1; y <= cy + 1; y++) {
            if (y < 0 || y >= g->dims[1]) continue;
            int x0 = cx - 1 < 0 ? 0 : cx - 1, x1 = cx + 1 >= g->dims[0] ? g->dims[0] - 1 : cx + 1;
            long long row = (long long)g->dims[0] * (y + (long long)g->dims[1] * z);
            int b = c->cell_start[row + x0], e = c->cell_start[row + x1 + 1];
            for (int s = b; s < e; s++) {
                int j = c->perm[s];
                if ((size_t)j == i) continue;
                float d[3];
                g_sub(ps[j].pos, ps[i].pos, d);
                float d2 = g_dot(d, d);
                if (!(d2 > lim)) { /* keeps NaN too */
                    if (cnt == *cap) { *cap *= 2; *buf = (int*)realloc(*buf, sizeof(int) * *cap); }
                    (*buf)[cnt++] = j;
                }
            }
        }
    }
    qsort(*buf, cnt, sizeof(int), cmp_int);
    return cnt;
}

int orc_max_threads(void) {
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}

/* SphUtils.cpp:100-149 over a cell list; threads <= 0 means all */
void orc_update_density_cells(const orc_params* p, orc_particle* ps, size_t n, int threads) {
    if (n == 0) return;
    cellctx c;
    cellctx_init(&c, p, ps, n);
    float* rho = (float*)malloc(sizeof(float) * n);
#ifdef _OPENMP
    if (threads <= 0) threads = omp_get_max_threads();
#pragma omp parallel num_threads(threads)
#endif
    {
        int cap = 512;
        int* buf = (int*)malloc(sizeof(int) * cap);
#ifdef _OPENMP
#pragma omp for schedule(dynamic, 256)
#endif
        for (long long i = 0; i < (long long)n; i++) {
            int cnt = gather_sorted(&c, p, ps, (size_t)i, &buf, &cap);
            float density = 0.0f;
            for (int k = 0; k < cnt; k++) density += density_term(p, &ps[i], &ps[buf[k]]);
            rho[i] = density;
        }
        free(buf);
    }
    for (size_t i = 0; i < n; i++) density_finish(p, &ps[i], rho[i]);
    free(rho);
    cellctx_free(&c);
}

/* SphUtils.cpp:157-239 over a cell list */
void orc_update_forces_cells(const orc_params* p, orc_particle* ps, size_t n, int threads) {
    if (n == 0) return;
    cellctx c;
    cellctx_init(&c, p, ps, n);
    float* acc_all = (float*)malloc(sizeof(float) * 3 * n);
#ifdef _OPENMP
    if (threads <= 0) threads = omp_get_max_threads();
#pragma omp parallel num_threads(threads)
#endif
    {
        int cap = 512;
        int* buf = (int*)malloc(sizeof(int) * cap);
#ifdef _OPENMP
#pragma omp for schedule(dynamic, 256)
#endif
        for (long long i = 0; i < (long long)n; i++) {
            int cnt = gather_sorted(&c, p, ps, (size_t)i, &buf, &cap);
            float acc[3] = {0.0f, 0.0f, 0.0f};
            acc[0] += 0.0f; acc[1] += -1.0f; acc[2] += 0.0f;
            for (int k = 0; k < cnt; k++) pressure_pair(p, &ps[i], &ps[buf[k]], acc);
            acc_all[3 * i] = acc[0]; acc_all[3 * i + 1] = acc[1]; acc_all[3 * i + 2] = acc[2];
        }
        free(buf);
    }
    for (size_t i = 0; i < n; i++) memcpy(ps[i].accel, &acc_all[3 * i], 3 * sizeof(float));
    free(acc_all);
    cellctx_free(&c);
}

/* viscosity_forces over a cell list (ascending original index): bit-identical to the all-pairs loop, because a
 * pair outside the support has lapW = 0.0f and adds +-0 */
void orc_viscosity_forces_cells(const orc_params* p, orc_particle* ps, size_t n, int threads) {
    if (n == 0) return;
    cellctx c;
    cellctx_init(&c, p, ps, n);
    float* acc_all = (float*)malloc(sizeof(float) * 3 * n);
#ifdef _OPENMP
    if (threads <= 0) threads = omp_get_max_threads();
#pragma omp parallel num_threads(threads)
#endif
    {
        int cap = 512;
        int* buf = (int*)malloc(sizeof(int) * cap);
#ifdef _OPENMP
#pragma omp for schedule(dynamic, 256)
#endif
        for (long long i = 0; i < (long long)n; i++) {
            int cnt = gather_sorted(&c, p, ps, (size_t)i, &buf, &cap);
            float acc[3] = {ps[i].accel[0], ps[i].accel[1], ps[i].accel[2]};
            for (int k = 0; k < cnt; k++) viscosity_pair(p, &ps[i], &ps[buf[k]], acc);
            acc_all[3 * i] = acc[0]; acc_all[3 * i + 1] = acc[1]; acc_all[3 * i + 2] = acc[2];
        }
        free(buf);
    }
    for (size_t i = 0; i < n; i++) memcpy(ps[i].accel, &acc_all[3 * i], 3 * sizeof(float));
    free(acc_all);
    cellctx_free(&c);
}

/* ---- FP64 referee: the same formulas evaluated in double on the same float inputs ------ */
/* Not a restatement of reference code: it answers "how far is the reference's own float arithmetic from the
 * exact value of its formulas", so that a reader can see how much of a GPU-vs-reference difference is the
 * reference's rounding.  rho_raw[n]: the unclamped density sum.  Neighbours in any order (double sums). */
void orc_density_f64_cells(const orc_params* p, const orc_particle* ps, size_t n, double* rho_raw, int threads) {
    if (n == 0) return;
    cellctx c;
    cellctx_init(&c, p, ps, n);
    const double h = (double)p->h, kw = 315.0 / (64.0 * (double)ORC_PI_F * pow(h, 9));
#ifdef _OPENMP
    if (threads <= 0) threads = omp_get_max_threads();
#pragma omp parallel num_threads(threads)
#endif
    {
        int cap = 512;
        int* buf = (int*)malloc(sizeof(int) * cap);
#ifdef _OPENMP
#pragma omp for schedule(dynamic, 256)
#endif
        for (long long i = 0; i < (long long)n; i++) {
            int cnt = gather_sorted(&c, p, ps, (size_t)i, &buf, &cap);
            double sum = 0.0;
            for (int k = 0; k < cnt; k++) {
                const orc_particle* b = &ps[buf[k]];
                double dx = (double)ps[i].pos[0] - b->pos[0], dy = (double)ps[i].pos[1] - b->pos[1], dz = (double)ps[i].pos[2] - b->pos[2];
                double q = sqrt(dx * dx + dy * dy + dz * dz) / h;
                if (q <= h) {
                    double t = h * h - q * q;
                    sum += (double)b->mass * kw * t * t * t;
                }
            }
            rho_raw[i] = sum;
        }
        free(buf);
    }
    cellctx_free(&c);
}

/* acc[3n]: gravity + pair sum (SphUtils.cpp:219-230) in double from the float pos, vel, mass, density, pressure in ps;
 * scale[n] = 1 + sum_j |temp_ij| |gradW_ij|: the magnitude against which a cancelling sum is accurate. */
void orc_forces_f64_cells(const orc_params* p, const orc_particle* ps, size_t n, double* acc_out, double* scale, int threads) {
    if (n == 0) return;
    cellctx c;
    cellctx_init(&c, p, ps, n);
    const double h = (double)p->h, kg = -45.0 / ((double)ORC_PI_F * pow(h, 6));
#ifdef _OPENMP
    if (threads <= 0) threads = omp_get_max_threads();
#pragma omp parallel num_threads(threads)
#endif
    {
        int cap = 512;
        int* buf = (int*)malloc(sizeof(int) * cap);
#ifdef _OPENMP
#pragma omp for schedule(dynamic, 256)
#endif
        for (long long i = 0; i < (long long)n; i++) {
            int cnt = gather_sorted(&c, p, ps, (size_t)i, &buf, &cap);
            const orc_particle* a = &ps[i];
            double acc[3] = {0.0, -1.0, 0.0}, sc = 1.0;
            for (int k = 0; k < cnt; k++) {
                const orc_particle* b = &ps[buf[k]];
                double dx[3], dv[3];
                for (int d = 0; d < 3; d++) { dx[d] = (double)a->pos[d] - b->pos[d]; dv[d] = (double)a->vel[d] - b->vel[d]; }
                double r2 = dx[0] * dx[0] + dx[1] * dx[1] + dx[2] * dx[2], r = sqrt(r2), q = r / h;
                if (!(q < h)) continue;
                double vdotr = dv[0] * dx[0] + dv[1] * dx[1] + dv[2] * dx[2], pi = 0.0;
                if (vdotr < 0) {
                    double cb = (sqrt(fabs((double)a->pressure / a->density)) + sqrt(fabs((double)b->pressure / b->density))) / 2.0;
                    double mu = h * vdotr / (r2 + 0.01 * h * h);
                    pi = 2.0 * (-0.1 * cb * mu + 0.2 * mu * mu) / ((double)a->density + b->density);
                }
                double temp = -(double)b->mass * (pi + ((double)a->pressure + b->pressure) / (2.0 * b->density * a->density));
                double g = kg * (h - q) * (h - q);
                for (int d = 0; d < 3; d++) acc[d] += temp * g * dx[d] / r;
                sc += fabs(temp * g);
            }
            acc_out[3 * i] = acc[0]; acc_out[3 * i + 1] = acc[1]; acc_out[3 * i + 2] = acc[2];
            if (scale) scale[i] = sc;
        }
        free(buf);
    }
    cellctx_free(&c);
}

void orc_step_cells(const orc_params* p, orc_particle* ps, size_t n, const orc_wall* w, size_t nw, int nsteps, int threads) {
    for (int k = 0; k < nsteps; k++) {
        orc_update_density_cells(p, ps, n, threads);
        orc_update_forces_cells(p, ps, n, threads);
        orc_update_posvel(p, ps, n, NULL, 0);
        orc_collision(p, ps, n, w, nw, NULL, 0);
    }
}

/* ---- scene: main.cpp initParticles / addParticle / initWalls, headless ----------------------- */

/* main.cpp:424-442 (GL calls at :449-457 dropped).  Returns the new count. */
size_t orc_add_particle(orc_particle* ps, size_t n, size_t max_particles, const float x[3], const float v[3], float h, float rho0, float eps) {
    orc_particle q;
    memset(&q, 0, sizeof q);
    memcpy(q.pos, x, sizeof q.pos);
    memcpy(q.vel, v, sizeof q.vel);
    q.density = rho0 + eps;
    q.mass = .1f * h * h * h * rho0;
    q.pressure = 1.0f;
    q.thermal = 1.0f;
    if (n < max_particles) ps[n++] = q;
    return n;
}

/* main.cpp:490-505: float loop counters advanced by the double literal 0.1 and compared against
 * float max (0.9f) / the double literal -0.6.  MAX_PARTICLES = 1300 (main.cpp:16).  972 particles. */
size_t orc_scene_default(orc_particle* ps, size_t cap, float h, float rho0, float eps) {
    float max = 0.9;
    float min = -0.9;
    size_t n = 0;
    size_t lim = cap < 1300 ? cap : 1300;
    for (float i = min; i <= max; i += 0.1)
        for (float j = -0.9; j <= -0.6; j += 0.1)
            for (float k = min; k <= max; k += 0.1) {
                float x[3] = {i, j, k}, v[3] = {0, 0, 0};
                n = orc_add_particle(ps, n, lim, x, v, h, rho0, eps);
            }
    return n;
}

/* main.cpp:508-540: vector order w1 (x=-1), w2 (x=+1), w4 (z=+1), w5 (z=-1), w3 (y=-1) */
void orc_walls_default(orc_wall w[5]) {
    static const float c[5][3] = {{-1, 0, 0}, {1, 0, 0}, {0, 0, 1}, {0, 0, -1}, {0, -1, 0}};
    static const float nrm[5][3] = {{1, 0, 0}, {-1, 0, 0}, {0, 0, -1}, {0, 0, 1}, {0, 1, 0}};
    for (int i = 0; i < 5; i++) {
        memcpy(w[i].center, c[i], sizeof c[i]);
        memcpy(w[i].normal, nrm[i], sizeof nrm[i]);
        w[i].xlength = 1.0f;
        w[i].ylength = 1.0f;
    }
}

/* FNV-1a 64 over the bit patterns of pos.x,pos.y,pos.z, one 32-bit word per multiply */
uint64_t orc_fnv1a_pos(const orc_particle* ps, size_t n) {
    uint64_t hsh = 0xcbf29ce484222325ULL;
    for (size_t i = 0; i < n; i++)
        for (int a = 0; a < 3; a++) {
            uint32_t w;
            memcpy(&w, &ps[i].pos[a], 4);
            hsh ^= (uint64_t)w;
            hsh *= 0x100000001b3ULL;
        }
    return hsh;
}

/* ---- viewer picture: the shader arithmetic restated on the CPU (parity UNPINNED for this part) -------------
 * SPH/Shaders/sph.vertexshader:8-12, sph.fragmentshader:7-24, main.cpp:48-60,229-252,563-571.  Running the GLSL
 * itself needs a graphics API, which is unavailable on this box, so nothing here can be checked against the
 * reference's own output; it is the checker for the CUDA splat kernel only. */

/* main.cpp:48-60 with GLM 0.9.6.1 gtc/matrix_transform.inl:212-232 (perspective, fovy in radians) and :388-413 (lookAt) */
void orc_default_mvp(float mvp[16]) {
    float fovy = 45.0f, aspect = 4.0f / 3.0f, zn = 0.1f, zf = 100.0f;
    float th = tanf(fovy / 2.0f);
    float P[4][4], V[4][4];
    memset(P, 0, sizeof P);
    memset(V, 0, sizeof V);
    P[0][0] = 1.0f / (aspect * th);
    P[1][1] = 1.0f / th;
    P[2][2] = -(zf + zn) / (zf - zn);
    P[2][3] = -1.0f;
    P[3][2] = -(2.0f * zf * zn) / (zf - zn);
    float eye[3] = {0.0f, -0.8f, 1.5f}, center[3] = {0.0f, -0.8f, 0.0f}, up[3] = {0.0f, 1.0f, 0.0f};
    float d[3], f[3], s0[3], s[3], u[3];
    g_sub(center, eye, d);
    g_normalize(d, f);
    s0[0] = f[1] * up[2] - up[1] * f[2]; s0[1] = f[2] * up[0] - up[2] * f[0]; s0[2] = f[0] * up[1] - up[0] * f[1];
    g_normalize(s0, s);
    u[0] = s[1] * f[2] - f[1] * s[2]; u[1] = s[2] * f[0] - f[2] * s[0]; u[2] = s[0] * f[1] - f[0] * s[1];
    for (int a = 0; a < 4; a++) V[a][a] = 1.0f;
    V[0][0] = s[0]; V[1][0] = s[1]; V[2][0] = s[2];
    V[0][1] = u[0]; V[1][1] = u[1]; V[2][1] = u[2];
    V[0][2] = -f[0]; V[1][2] = -f[1]; V[2][2] = -f[2];
    V[3][0] = -g_dot(s, eye); V[3][1] = -g_dot(u, eye); V[3][2] = g_dot(f, eye);
    for (int c = 0; c < 4; c++)
        for (int r = 0; r < 4; r++) {
            float acc = 0.0f;
            for (int k = 0; k < 4; k++) acc += P[k][r] * V[c][k];
            mvp[4 * c + r] = acc;
        }
}

static void mvp_apply(const float m[16], float x, float y, float z, float out[4]) {
    for (int r = 0; r < 4; r++) out[r] = m[r] * x + m[4 + r] * y + m[8 + r] * z + m[12 + r];
}

/* particles in order, later ones overwrite (no depth buffer, main.cpp:730); image[h][w], row 0 at the bottom */
/* depth_test = 0: no depth buffer, a later particle overwrites an earlier one (what GLUT_DOUBLE | GLUT_RGB asks for,
 * main.cpp:730).  depth_test = 1: the visual happens to carry depth bits, so glEnable(GL_DEPTH_TEST) / glDepthFunc(GL_LESS)
 * (main.cpp:567-571) act: the fragment shader does not write gl_FragDepth, so every fragment of a sprite has the depth of the
 * point itself, (clip.z / clip.w + 1) / 2, and the sprite with the smallest one wins a pixel (the first drawn on a tie). */
void orc_render_depth2(const orc_particle* ps, size_t n, const float mvp[16], int width, int height, float point_scale,
                       float sprite_radius, float clear_value, int depth_test, float* image) {
    float* zbuf = NULL;
    if (depth_test) {
        zbuf = (float*)malloc(sizeof(float) * (size_t)width * height);
        for (long i = 0; i < (long)width * height; i++) zbuf[i] = 1.0f; /* glClear(GL_DEPTH_BUFFER_BIT) */
    }
    for (long i = 0; i < (long)width * height; i++) image[i] = clear_value;
    for (size_t k = 0; k < n; k++) {
        float clip[4];
        mvp_apply(mvp, ps[k].pos[0], ps[k].pos[1], ps[k].pos[2], clip);
        if (!(clip[3] > 0.0f) || fabsf(clip[0]) > clip[3] || fabsf(clip[1]) > clip[3] || fabsf(clip[2]) > clip[3]) continue;
        float xw = (clip[0] / clip[3] * 0.5f + 0.5f) * (float)width, yw = (clip[1] / clip[3] * 0.5f + 0.5f) * (float)height;
        float edge = point_scale / clip[3];
        if (edge < 1.0f) edge = 1.0f;
        if (edge > 2048.0f) edge = 2048.0f;
        int x0 = (int)floorf(xw - 0.5f * edge), x1 = (int)ceilf(xw + 0.5f * edge);
        int y0 = (int)floorf(yw - 0.5f * edge), y1 = (int)ceilf(yw + 0.5f * edge);
        if (x0 < 0) x0 = 0;
        if (y0 < 0) y0 = 0;
        if (x1 > width - 1) x1 = width - 1;
        if (y1 > height - 1) y1 = height - 1;
        for (int py = y0; py <= y1; py++)
            for (int px = x0; px <= x1; px++) {
                float s = 0.5f + ((float)px + 0.5f - xw) / edge, t = 0.5f + ((float)py + 0.5f - yw) / edge;
                if (s < 0.0f || s >= 1.0f || t < 0.0f || t >= 1.0f) continue;
                float nx = s * 2.0f - 1.0f, ny = t * 2.0f - 1.0f;
                float mag = nx * nx + ny * ny;
                if (mag > 1.0f) continue;
                float nz = sqrtf(1.0f - mag);
                float cp[4];
                mvp_apply(mvp, ps[k].pos[0] + nx * sprite_radius, ps[k].pos[1] + ny * sprite_radius, ps[k].pos[2] + nz * sprite_radius, cp);
                float depth = cp[2] / cp[3];
                if (depth < 0.0f) depth = 0.0f;
                if (depth > 1.0f) depth = 1.0f;
                if (zbuf) {
                    float zwin = clip[2] / clip[3] * 0.5f + 0.5f;
                    if (!(zwin < zbuf[(size_t)py * width + px])) continue; /* GL_LESS */
                    zbuf[(size_t)py * width + px] = zwin;
                }
                image[(size_t)py * width + px] = depth;
            }
    }
    free(zbuf);
}

void orc_render_depth(const orc_particle* ps, size_t n, const float mvp[16], int width, int height, float point_scale,
                      float sprite_radius, float clear_value, float* image) {
    orc_render_depth2(ps, n, mvp, width, height, point_scale, sprite_radius, clear_value, 0, image);
}
