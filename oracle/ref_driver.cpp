// TEST INFRASTRUCTURE ONLY -- never linked into, imported by, or called from the product path.
//
// Thin extern "C" wrapper around the UNMODIFIED reference step engine.  This file contains no
// reference code: it #includes the reference's own header from where it lies under
// /root/reference (see oracle/Makefile, target _ref/libsphref.so) and is linked against the
// reference's own SPH/SphUtils.cpp compiled in place.  It exists so that the C restatement in
// sph_oracle.c (and the golden fixtures under tests/golden/) can be pinned against the real thing.
//
// Reference interface wrapped: SPH/SphUtils.hpp:16-44 (ctor + five public methods + three kernel
// helpers), SPH/Particle.hpp:42-50 (52-byte record), SPH/Wall.hpp:28-31 (32-byte record).
#include <cstddef>
#include <cstring>
#include <iostream>
#include <math.h>
#include <vector>

#include <glm/glm.hpp>            // guarded: pulled in before the keyword trick below
#include <glm/gtc/matrix_transform.hpp>  // glm::perspective / glm::lookAt, header-only (deps/include/glm/gtc/matrix_transform.inl)

#define private public            // read bbox / cells for the update_cells pin (SphUtils.hpp:46-57)
#include "SphUtils.hpp"
#undef private

static_assert(sizeof(Particle) == 52, "Particle layout drifted");
static_assert(sizeof(Wall) == 32, "Wall layout drifted");
static_assert(sizeof(Sphere) == 44, "Sphere layout drifted");

namespace {
std::vector<Particle> to_vec(const void* p, size_t n) {
    std::vector<Particle> v(n);
    if (n) std::memcpy((void*)v.data(), p, n * sizeof(Particle));
    return v;
}
void from_vec(const std::vector<Particle>& v, void* p) {
    if (!v.empty()) std::memcpy(p, (const void*)v.data(), v.size() * sizeof(Particle));
}
std::vector<Wall> walls_vec(const void* w, size_t n) {
    std::vector<Wall> v(n);
    if (n) std::memcpy((void*)v.data(), w, n * sizeof(Wall));
    return v;
}
std::vector<Sphere> spheres_vec(const void* s, size_t n) {
    std::vector<Sphere> v(n);
    if (n) std::memcpy((void*)v.data(), s, n * sizeof(Sphere));
    return v;
}
}  // namespace

extern "C" {

void* ref_create(float h, float rho0, float c, float dt, float eps) { return new SphUtils(h, rho0, c, dt, eps); }
void ref_destroy(void* s) { delete static_cast<SphUtils*>(s); }

size_t ref_sizeof_particle() { return sizeof(Particle); }
size_t ref_offsets(int which) {
    switch (which) {
        case 0: return offsetof(Particle, pos);
        case 1: return offsetof(Particle, vel);
        case 2: return offsetof(Particle, accel);
        case 3: return offsetof(Particle, mass);
        case 4: return offsetof(Particle, density);
        case 5: return offsetof(Particle, pressure);
        case 6: return offsetof(Particle, thermal);
        case 10: return offsetof(Wall, normal);
        case 11: return offsetof(Wall, center);
        case 20: return offsetof(Sphere, pos);
        case 21: return offsetof(Sphere, vel);
        case 22: return offsetof(Sphere, accel);
        case 23: return offsetof(Sphere, radius);
        case 24: return offsetof(Sphere, mass);
    }
    return (size_t)-1;
}

float ref_kernel_function(void* s, const float* i, const float* j) {
    return static_cast<SphUtils*>(s)->kernel_function(glm::vec3(i[0], i[1], i[2]), glm::vec3(j[0], j[1], j[2]));
}
void ref_grad_kernel(void* s, const float* i, const float* j, float* out) {
    glm::vec3 g = static_cast<SphUtils*>(s)->grad_kernel(glm::vec3(i[0], i[1], i[2]), glm::vec3(j[0], j[1], j[2]));
    out[0] = g.x; out[1] = g.y; out[2] = g.z;
}
float ref_lap_kernel(void* s, const float* i, const float* j) {
    return static_cast<SphUtils*>(s)->lap_kernel(glm::vec3(i[0], i[1], i[2]), glm::vec3(j[0], j[1], j[2]));
}

void ref_update_density(void* s, void* particles, size_t n) {
    auto v = to_vec(particles, n);
    static_cast<SphUtils*>(s)->update_density(v);
    from_vec(v, particles);
}
void ref_update_forces(void* s, void* particles, size_t n, void* spheres, size_t ns) {
    auto v = to_vec(particles, n);
    auto sp = spheres_vec(spheres, ns);
    static_cast<SphUtils*>(s)->update_forces(v, sp);
    from_vec(v, particles);
    if (ns) std::memcpy(spheres, (const void*)sp.data(), ns * sizeof(Sphere));
}
void ref_update_posvel(void* s, void* particles, size_t n, void* spheres, size_t ns) {
    auto v = to_vec(particles, n);
    auto sp = spheres_vec(spheres, ns);
    static_cast<SphUtils*>(s)->update_posvel(v, sp);
    from_vec(v, particles);
    if (ns) std::memcpy(spheres, (const void*)sp.data(), ns * sizeof(Sphere));
}
void ref_collision(void* s, void* particles, size_t n, const void* walls, size_t nw, void* spheres, size_t ns) {
    auto v = to_vec(particles, n);
    auto w = walls_vec(walls, nw);
    auto sp = spheres_vec(spheres, ns);
    static_cast<SphUtils*>(s)->collision(v, w, sp);
    from_vec(v, particles);
    if (ns) std::memcpy(spheres, (const void*)sp.data(), ns * sizeof(Sphere));
}

// The step in the order main.cpp:293-300 calls it (step_scene), nsteps times, no emitter.
void ref_step(void* s, void* particles, size_t n, const void* walls, size_t nw, int nsteps) {
    auto v = to_vec(particles, n);
    auto w = walls_vec(walls, nw);
    std::vector<Sphere> sp;
    SphUtils* u = static_cast<SphUtils*>(s);
    for (int k = 0; k < nsteps; ++k) {
        u->update_density(v);
        u->update_forces(v, sp);
        u->update_posvel(v, sp);
        u->collision(v, w, sp);
    }
    from_vec(v, particles);
}

// viscosity_forces (SphUtils.cpp:244-290) is private and its only call site is commented out (:175).  Exposed
// so that the oracle's restatement of its pair body can be pinned: on a particle set that fits inside ONE of
// the reference's 2h-wide cells, its half stencil {-1,0}^3 degenerates to "all pairs, ascending index".
// Adds into accel, like the reference would.
void ref_viscosity_forces(void* s, void* particles, size_t n) {
    auto v = to_vec(particles, n);
    SphUtils* u = static_cast<SphUtils*>(s);
    u->update_cells(v);
    u->viscosity_forces(v);
    from_vec(v, particles);
}

// update_cells (SphUtils.cpp:8-47) with its private results exposed.  The caller must make sure the
// reference's flat index (with its stray "+ny", SphUtils.cpp:43) stays inside nx*ny*nz, otherwise
// the reference itself writes out of bounds; ref_cells_index_bound() tells.
// out_bbox = {minx,miny,minz,maxx,maxy,maxz}; out_dims = {nx,ny,nz}; cell_of[n] = flat index the
// reference filed particle m under (found by searching the cells it built).
int ref_update_cells(void* s, void* particles, size_t n, float* out_bbox, int* out_dims, int* cell_of) {
    SphUtils* u = static_cast<SphUtils*>(s);
    auto v = to_vec(particles, n);
    u->update_cells(v);
    out_bbox[0] = u->minx; out_bbox[1] = u->miny; out_bbox[2] = u->minz;
    out_bbox[3] = u->maxx; out_bbox[4] = u->maxy; out_bbox[5] = u->maxz;
    int nx = (int)ceil((u->maxx - u->minx) / u->cell_dimensions) + 1;
    int ny = (int)ceil((u->maxy - u->miny) / u->cell_dimensions) + 1;
    int nz = (int)ceil((u->maxz - u->minz) / u->cell_dimensions) + 1;
    out_dims[0] = nx; out_dims[1] = ny; out_dims[2] = nz;
    for (size_t c = 0; c < u->cells.size(); ++c) {
        int prev = -1;
        for (int m : u->cells[c]) {
            if (m <= prev) return -1;  // not ascending => not a stable fill
            prev = m;
            cell_of[m] = (int)c;
        }
    }
    return (int)u->cells.size();
}


// The viewer's camera: the GLM calls of SPH/main.cpp:48-60 (a file-scope global there; main.cpp itself needs GL
// headers and cannot be compiled here) evaluated by the reference's own vendored GLM 0.9.6.1 with main.cpp's
// arguments.  Column-major, 16 floats, as glUniformMatrix4fv(..., GL_FALSE, &MVP[0][0]) would send it.
void ref_default_mvp(float out[16]) {
    glm::mat4 Projection = glm::perspective(45.0f, 4.0f / 3.0f, 0.1f, 100.0f);                                       // main.cpp:48
    glm::mat4 View = glm::lookAt(glm::vec3(0.0f, -0.8f, 1.5f), glm::vec3(0.0f, -0.8f, 0.0f), glm::vec3(0.0f, 1.0f, 0.0f));  // :50-54
    glm::mat4 Model = glm::mat4(1.0f);                                                                                // :57
    glm::mat4 MVP = Projection * View * Model;                                                                        // :60
    std::memcpy(out, &MVP[0][0], 16 * sizeof(float));
}
}  // extern "C"
