// Headless driver: the reference's scene set-up and step loop (SPH/main.cpp initScene :555-559,
// initParticles :490-505, addParticle :424-442, initWalls :508-540, step_scene :280-303) with every GL /
// GLUT call left out, driving the engine through the drop-in SphUtils class exactly as main.cpp does.
//   sph_headless [steps=20] [fused|methods] [emit|noemit] [picture.pgm] [devices=0,1]
// "devices=a,b,..." runs the step on those GPUs (sphb_multi_*: x-slabs, peer-to-peer halo links); the same ordinal twice puts
// two slabs on one GPU.  The picture needs the single-engine path.
// "fused" uses SphUtils::step_scene (one sphb_step_aos call per step) instead of the four methods.
// "emit" switches the emitter of step_scene on (main.cpp:281-290, the Enter key): one particle per step from
// rand(), until MAX_PARTICLES -- the particle vector grows between steps and the engine follows.
// A last argument ending in .pgm writes the viewer's picture of the final state (sphb_render_depth: the arithmetic
// of the reference's point-sprite shaders as CUDA kernels; 1080 x 920, 8-bit grey, top row first).
// Prints particle count and, per reported step, mean density, kinetic energy and the FNV-1a-64 hash of the
// position bits (the same hash tests/golden/default_scene.npz holds for the compiled reference).
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "SphUtils.hpp"

#define MAX_PARTICLES 1300

const float dt = 0.033f;
const float smooth_length = 0.45;
const float rho0 = 10.0f;
const float c = 100.0f;
const float epsilon = 0.01f;

SphUtils sph(smooth_length, rho0, c, dt, epsilon);   // a global, as at main.cpp:62
std::vector<Particle> particles;
std::vector<Wall> walls;
std::vector<Sphere> spheres;

static void addParticle(glm::vec3 x, glm::vec3 v) {
    float density = rho0 + epsilon;
    float mass = .1f * smooth_length * smooth_length * smooth_length * rho0;
    Particle part(x, v, glm::vec3(0.0f, 0.0f, 0.0f), mass, density, 1.0f, 1.0f);
    if (particles.size() < MAX_PARTICLES) particles.push_back(part);
}

static void initScene() {
    float max = 0.9;
    float min = -0.9;
    for (float i = min; i <= max; i += 0.1)
        for (float j = -0.9; j <= -0.6; j += 0.1)
            for (float k = min; k <= max; k += 0.1) addParticle(glm::vec3(i, j, k), glm::vec3(0, 0, 0));
    // centre, normal; pushed as w1, w2, w4, w5, w3
    walls.push_back(Wall(glm::vec3(-1.0f, 0.0f, 0.0f), glm::vec3(1.0f, 0.0f, 0.0f), 1.0f, 1.0f));
    walls.push_back(Wall(glm::vec3(1.0f, 0.0f, 0.0f), glm::vec3(-1.0f, 0.0f, 0.0f), 1.0f, 1.0f));
    walls.push_back(Wall(glm::vec3(0.0f, 0.0f, 1.0f), glm::vec3(0.0f, 0.0f, -1.0f), 1.0f, 1.0f));
    walls.push_back(Wall(glm::vec3(0.0f, 0.0f, -1.0f), glm::vec3(0.0f, 0.0f, 1.0f), 1.0f, 1.0f));
    walls.push_back(Wall(glm::vec3(0.0f, -1.0f, 0.0f), glm::vec3(0.0f, 1.0f, 0.0f), 1.0f, 1.0f));
}

static bool generateParticles = false;

static void step_scene(bool fused) {
    if (generateParticles) { // main.cpp:281-290
        float x = -.2f + rand() / ((float)RAND_MAX / (-0.2f - 0.2f));
        float y = 1.3f;
        float z = -.2f + rand() / ((float)RAND_MAX / (-0.2f - 0.2f));
        float xvel = -.2f + rand() / ((float)RAND_MAX / (-0.2f - 0.2f));
        float zvel = -.2f + rand() / ((float)RAND_MAX / (-0.2f - 0.2f));
        float yvel = -1.0f;
        addParticle(glm::vec3(x, y, z), glm::vec3(xvel, yvel, zvel));
    }
    if (fused) {
        sph.step_scene(particles, walls, 1);
        return;
    }
    sph.update_density(particles);
    sph.update_forces(particles, spheres);
    sph.update_posvel(particles, spheres);
    sph.collision(particles, walls, spheres);
}

static uint64_t fnv_pos() {
    uint64_t h = 0xcbf29ce484222325ULL;
    for (size_t i = 0; i < particles.size(); i++) {
        float p[3] = {particles[i].pos.x, particles[i].pos.y, particles[i].pos.z};
        for (int a = 0; a < 3; a++) {
            uint32_t w;
            memcpy(&w, &p[a], 4);
            h ^= w;
            h *= 0x100000001b3ULL;
        }
    }
    return h;
}

int main(int argc, char** argv) {
    std::vector<int> devices;
    for (int a = 1; a < argc; a++)
        if (strncmp(argv[a], "devices=", 8) == 0) {
            for (const char* q = argv[a] + 8; *q;) {
                devices.push_back(atoi(q));
                while (*q && *q != ',') q++;
                if (*q == ',') q++;
            }
            for (int b = a; b + 1 < argc; b++) argv[b] = argv[b + 1];
            argc--;
            a--;
        }
    if (!devices.empty()) sph.use_devices(devices);
    int steps = argc > 1 ? atoi(argv[1]) : 20;
    bool fused = argc > 2 && strcmp(argv[2], "fused") == 0;
    generateParticles = argc > 3 && strcmp(argv[3], "emit") == 0;
    initScene();
    printf("particles %zu walls %zu\n", particles.size(), walls.size());
    if (!sph.ready()) return 2;
    for (int s = 1; s <= steps; s++) {
        step_scene(fused);
        if (s == 1 || s == 2 || s == 10 || s == 20 || s == 100 || s == steps) {
            double rho = 0.0, ke = 0.0;
            for (size_t i = 0; i < particles.size(); i++) {
                const Particle& q = particles[i];
                rho += q.density;
                ke += 0.5 * q.mass * ((double)q.vel.x * q.vel.x + (double)q.vel.y * q.vel.y + (double)q.vel.z * q.vel.z);
            }
            printf("step %d mean_rho %.6f ke %.6f fnv %016llx n %zu\n", s, rho / particles.size(), ke, (unsigned long long)fnv_pos(),
                   particles.size());
        }
    }
    const char* pgm = argc > 4 ? argv[4] : 0;
    if (pgm && sph.engine()) {
        sphb_camera cam;
        sphb_default_camera(&cam);
        std::vector<float> img((size_t)cam.width * cam.height);
        // the picture is taken from the engine's resident copy of the particles (it holds the state of the last call)
        if (sphb_render_depth(sph.engine(), &cam, &img[0]) != SPHB_OK) {
            fprintf(stderr, "render failed: %s\n", sphb_last_error());
            return 3;
        }
        FILE* f = fopen(pgm, "wb");
        if (!f) return 4;
        fprintf(f, "P5\n%d %d\n255\n", cam.width, cam.height);
        for (int y = cam.height - 1; y >= 0; y--)          // image rows are bottom-up (GL), PGM is top-down
            for (int x = 0; x < cam.width; x++) fputc((int)(img[(size_t)y * cam.width + x] * 255.0f + 0.5f), f);
        fclose(f);
        printf("wrote %s\n", pgm);
    }
    return 0;
}
