// Drop-in replacement for the reference's SPH/SphUtils.hpp (class SphUtils, /root/reference
// SPH/SphUtils.hpp:16-44): same class name, constructors and public method signatures, implemented as
// inline calls into the C ABI of libsphb200.so (include/sphb200.h).  main.cpp (SPH/main.cpp:62,293-300)
// compiles against this header unchanged; SphUtils.cpp is no longer needed.  See INTEGRATION.md.
//
// Build modes
//   default                  includes the reference's own Particle.hpp / Wall.hpp / Sphere.hpp (and GLM)
//   SPHB_SHIM_STANDALONE     uses sphb_records.hpp + sphb_glm_min.hpp (no reference tree needed)
// Several GPUs
//   use_devices(list)        before the first call: the step runs on those GPUs through sphb_multi_* (x-slabs, peer-to-peer halo
//                            links).  The four methods then act as ONE step executed at collision() -- the only point at which
//                            main.cpp looks at the particles (main.cpp:97-110) -- and the host vector is refreshed there.
// Sync policy
//   default                  every method leaves the host vector exactly as the reference would
//                            (density/pressure after update_density, accel after update_forces, ...)
//   SPHB_SHIM_SYNC_ON_COLLISION   state stays on the GPU from update_density to collision(); the host
//                            vector is refreshed once, at the end of collision() -- all main.cpp needs
#ifndef SPHB_SHIM_SPHUTILS_HPP
#define SPHB_SHIM_SPHUTILS_HPP

#ifdef SPHB_SHIM_STANDALONE
#include "sphb_records.hpp"
#else
#include "Particle.hpp"
#include "Wall.hpp"
#include "Sphere.hpp"
#endif
#include <math.h>
#include <stdio.h>
#include <string.h>
#include <iostream>
#include <vector>

#include "sphb200.h"

class SphUtils {
public:
    SphUtils() : eng_(0), multi_(0), failed_(false), resident_(false), nwalls_(0) { sphb_default_params(&params_); }

    SphUtils(float nsmooth_length, float nrho0, float nc, float ndt, float nepsilon)
        : eng_(0), multi_(0), failed_(false), resident_(false), nwalls_(0) {
        // A global SphUtils is constructed before main() (main.cpp:62): no CUDA call here, only numbers.
        sphb_default_params(&params_);
        params_.h = nsmooth_length;
        params_.rho0 = nrho0;
        params_.c = nc;
        params_.dt = ndt;
        params_.eps = nepsilon;
    }

    ~SphUtils() {
        if (eng_) sphb_destroy(eng_);
        if (multi_) sphb_multi_destroy(multi_);
    }

    // Run the step on several GPUs (CUDA ordinals; the same ordinal twice puts two slabs on one GPU).  Before the first call.
    bool use_devices(const std::vector<int>& devices) {
        if (eng_ || multi_ || devices.empty()) return false;
        devices_ = devices;
        return true;
    }

    // Solver constants beyond the five constructor arguments, before the first call (optional).
    sphb_params& params() { return params_; }
    sphb_engine* engine() { return ensure() ? eng_ : 0; }   // null when several GPUs are in use
    bool ready() { return ensure(); }

    float kernel_function(glm::vec3 i, glm::vec3 j) {
        float w = 0.0f;
        eval(i, j, &w, 0, 0);
        return w;
    }
    glm::vec3 grad_kernel(glm::vec3 i, glm::vec3 j) {
        float g[3] = {0.0f, 0.0f, 0.0f};
        eval(i, j, 0, g, 0);
        return glm::vec3(g[0], g[1], g[2]);
    }
    float lap_kernel(glm::vec3 i, glm::vec3 j) {
        float l = 0.0f;
        eval(i, j, 0, 0, &l);
        return l;
    }

    void update_cells(std::vector<Particle>& particles) {
        if (particles.empty() || !ensure()) return;
        if (!push(particles)) return;
        ok(sphb_update_cells(eng_));
    }

    void update_density(std::vector<Particle>& particles) {
        if (particles.empty() || !ensure()) return;
        if (multi_) return; // several GPUs: the whole step runs at collision()
        // first call of a step: the host vector is the truth (it may have grown, main.cpp:441)
        resident_ = false;
        if (!push(particles)) return;
        if (!ok(sphb_update_density(eng_))) return;
        pull(particles, SPHB_F_DENSITY);
    }

    void update_forces(std::vector<Particle>& particles, std::vector<Sphere>& spheres) {
        if (particles.empty() || !ensure()) return;
        if (multi_) return;
        if (!spheres_in(spheres) || !push(particles)) return;
        if (!ok(sphb_update_forces(eng_))) return;
        pull(particles, SPHB_F_ACCEL);
        spheres_out(spheres);
    }

    void update_posvel(std::vector<Particle>& particles, std::vector<Sphere>& spheres) {
        if (particles.empty() || !ensure()) return;
        if (multi_) return;
        if (!spheres_in(spheres) || !push(particles)) return;
        if (!ok(sphb_update_posvel(eng_))) return;
        pull(particles, SPHB_F_POS | SPHB_F_VEL);
        spheres_out(spheres);
    }

    void collision(std::vector<Particle>& particles, std::vector<Wall>& walls, std::vector<Sphere>& spheres) {
        if (particles.empty() || !ensure()) return;
        if (multi_) {
            step_scene(particles, walls, 1);
            return;
        }
        if (!set_walls(walls) || !spheres_in(spheres)) return;
        if (!push(particles)) return;
        if (!ok(sphb_collision(eng_))) return;
#ifdef SPHB_SHIM_SYNC_ON_COLLISION
        ok(sphb_download_aos(eng_, &particles[0], particles.size(), sizeof(Particle), SPHB_F_ALL));
        resident_ = false;
#else
        pull(particles, SPHB_F_POS | SPHB_F_VEL);
#endif
        spheres_out(spheres);
    }

    // step_scene (main.cpp:280-303) x nsteps through the fused fast path; not part of the reference's class.
    void step_scene(std::vector<Particle>& particles, std::vector<Wall>& walls, int nsteps = 1) {
        if (particles.empty() || !ensure() || !set_walls(walls)) return;
        resident_ = false;
        if (multi_) {
            mok(sphb_multi_step_aos(multi_, &particles[0], particles.size(), sizeof(Particle), nsteps, SPHB_F_ALL));
            return;
        }
        ok(sphb_step_aos(eng_, &particles[0], particles.size(), sizeof(Particle), nsteps, SPHB_F_ALL));
    }

private:
    sphb_params params_;
    sphb_engine* eng_;
    sphb_multi* multi_;
    std::vector<int> devices_;
    bool failed_, resident_;
    size_t nwalls_;
    float walls_[SPHB_MAX_WALLS * 6];

    SphUtils(const SphUtils&);
    SphUtils& operator=(const SphUtils&);

    bool ok(int rc) {
        if (rc == SPHB_OK) return true;
        if (!failed_) fprintf(stderr, "sphb200: %s (further calls are ignored)\n", sphb_last_error());
        failed_ = true;
        return false;
    }
    bool mok(int rc) {
        if (rc == SPHB_OK) return true;
        if (!failed_) fprintf(stderr, "sphb200: %s (further calls are ignored)\n", sphb_multi_last_error());
        failed_ = true;
        return false;
    }
    bool ensure() {
        if (failed_) return false;
        if (eng_ || multi_) return true;
        if (!devices_.empty()) return mok(sphb_multi_create((int)devices_.size(), &devices_[0], &params_, &multi_));
        return ok(sphb_create(&params_, 0, &eng_));
    }
    bool push(std::vector<Particle>& particles) {
#ifdef SPHB_SHIM_SYNC_ON_COLLISION
        if (resident_) return true;
        resident_ = true;
#endif
        return ok(sphb_upload_aos(eng_, &particles[0], particles.size(), sizeof(Particle)));
    }
    void pull(std::vector<Particle>& particles, unsigned mask) {
#ifdef SPHB_SHIM_SYNC_ON_COLLISION
        (void)particles; (void)mask;
#else
        ok(sphb_download_aos(eng_, &particles[0], particles.size(), sizeof(Particle), mask));
#endif
    }
    bool set_walls(std::vector<Wall>& walls) {
        float w[SPHB_MAX_WALLS * 6];
        size_t n = walls.size() < SPHB_MAX_WALLS ? walls.size() : SPHB_MAX_WALLS;
        for (size_t k = 0; k < n; k++) memcpy(w + 6 * k, &walls[k], 6 * sizeof(float));   // normal.xyz, center.xyz lead
        if (n == nwalls_ && memcmp(w, walls_, n * 6 * sizeof(float)) == 0) return true;
        memcpy(walls_, w, n * 6 * sizeof(float));
        nwalls_ = n;
        if (multi_) return mok(sphb_multi_set_walls(multi_, walls_, n, 6 * sizeof(float)));
        return ok(sphb_set_walls(eng_, walls_, n, 6 * sizeof(float)));
    }
    void eval(glm::vec3 i, glm::vec3 j, float* w, float* g, float* l) {
        if (!ensure()) return;
        float a[3] = {i.x, i.y, i.z}, b[3] = {j.x, j.y, j.z};
        ok(sphb_eval_kernels(eng_, a, b, 1, w, g, l));
    }
    // the reference mutates the caller's Sphere records in place: hand them to the engine, take them back
    bool spheres_in(std::vector<Sphere>& spheres) {
        size_t n = spheres.size() < SPHB_MAX_SPHERES ? spheres.size() : SPHB_MAX_SPHERES;
        return ok(sphb_set_spheres(eng_, n ? &spheres[0] : 0, n, sizeof(Sphere)));
    }
    void spheres_out(std::vector<Sphere>& spheres) {
        size_t n = spheres.size() < SPHB_MAX_SPHERES ? spheres.size() : SPHB_MAX_SPHERES;
        if (n) ok(sphb_get_spheres(eng_, &spheres[0], n, sizeof(Sphere)));
    }
};
#endif
