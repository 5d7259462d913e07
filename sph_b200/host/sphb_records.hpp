// Host records for standalone builds (SPHB_SHIM_STANDALONE): same member names and byte layout as the
// reference's Particle (52 B), Wall (32 B) and Sphere (44 B) -- /root/reference SPH/Particle.hpp:42-50,
// Wall.hpp:28-31, Sphere.hpp:29-33 -- so code written against the reference compiles against these too.
// A real drop-in build does NOT use this file: it includes the reference's own three headers.
#ifndef SPHB_RECORDS_HPP
#define SPHB_RECORDS_HPP
#include "sphb_glm_min.hpp"

struct Particle {
    glm::vec3 pos, vel, accel;
    float mass = 0.0f, density = 0.0f, pressure = 0.0f, thermal = 0.0f;
    Particle() {}
    Particle(glm::vec3 p, glm::vec3 v, glm::vec3 a, float m, float rho, float prs, float th)
        : pos(p), vel(v), accel(a), mass(m), density(rho), pressure(prs), thermal(th) {}
};

struct Wall {
    glm::vec3 normal, center;   // normal comes first in memory (Wall.hpp:28-29) although the ctor takes center first
    float xlength = 0.0f, ylength = 0.0f;
    Wall() {}
    Wall(glm::vec3 c, glm::vec3 n, float xl, float yl) : normal(n), center(c), xlength(xl), ylength(yl) {}
};

struct Sphere {
    glm::vec3 pos, vel, accel;
    float radius = 1.0f, mass = 1.0f;
    Sphere() {}
    Sphere(glm::vec3 p, glm::vec3 v, float r, float m) : pos(p), vel(v), radius(r), mass(m) {}
};

static_assert(sizeof(Particle) == 52 && sizeof(Wall) == 32 && sizeof(Sphere) == 44, "record layout");
#endif
