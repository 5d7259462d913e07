/*
 * sphb200.h -- C ABI of the B200-native SPH step engine (libsphb200.so).
 *
 * This is the drop-in boundary for the one hot path of Dimitime/SPH: the per-step
 *   update_density -> update_forces -> update_posvel -> collision
 * sequence that SPH/main.cpp:293-300 (step_scene) runs on the CPU through class SphUtils
 * (SPH/SphUtils.hpp:16-44).  Everything here is extern "C": opaque handle, plain pointers and sizes,
 * int status codes.  All compute runs in hand-written CUDA kernels (sm_100a); there is no CPU path.
 *
 * Host exchange format = the reference's own records, byte for byte:
 *   Particle  52 B  pos@0 vel@12 accel@24 mass@36 density@40 pressure@44 thermal@48   (SPH/Particle.hpp:42-50)
 *   Wall      32 B  normal@0 center@12 xlength@24 ylength@28                           (SPH/Wall.hpp:28-31)
 *   Sphere    44 B  pos@0 vel@12 accel@24 radius@36 mass@40                            (SPH/Sphere.hpp:29-33)
 *
 * Threading: one engine per GPU, calls on one engine from one thread at a time (the reference is
 * single-threaded, main.cpp idle()/keyboard() callbacks).  Different engines are independent.
 */
#ifndef SPHB200_H
#define SPHB200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define SPHB_OK 0
#define SPHB_ERR_INVALID 1   /* bad argument */
#define SPHB_ERR_CUDA 2      /* a CUDA runtime call failed; see sphb_last_error() */
#define SPHB_ERR_STATE 3     /* call out of order (e.g. step before upload) */
#define SPHB_ERR_CAPACITY 4  /* a fixed capacity was exceeded */

#define SPHB_MAX_WALLS 16
#define SPHB_MAX_SPHERES 16

/* field_mask bits for sphb_download_aos / sphb_step_aos */
#define SPHB_F_POS 1u
#define SPHB_F_VEL 2u
#define SPHB_F_ACCEL 4u
#define SPHB_F_DENSITY 8u  /* density and pressure */
#define SPHB_F_ALL 15u

/* domain_mode */
#define SPHB_DOMAIN_BBOX 0  /* grid origin = bounding-box minimum of the particles, recomputed on the device every
                               step: the behaviour of update_cells, SPH/SphUtils.cpp:11-27 */
#define SPHB_DOMAIN_FIXED 1 /* grid covers [domain_lo, domain_hi]; cell coordinates are clamped to it */

typedef struct sphb_engine sphb_engine;

/*
 * Solver constants.  The first five are the SphUtils constructor's arguments (SPH/SphUtils.hpp:20-29;
 * main.cpp:30-41 passes h=0.45, rho0=10, c=100, dt=0.033, eps=0.01).  The rest are literals that the
 * reference hard-codes in SphUtils.cpp; sphb_default_params() sets them to those literals.
 */
typedef struct sphb_params {
    float h;               /* smooth_length.  NOTE the support radius is h*h (SphUtils.cpp:57-61: q=r/h compared to h) */
    float rho0;            /* rest density; density is clamped UP to it (SphUtils.cpp:141-142) */
    float c;               /* stored, never read by the reference */
    float dt;              /* time step (SphUtils.cpp:326-328) */
    float eps;             /* stored, never read by the reference */
    float eos_k;           /* 100.0f: pressure = eos_k*(density-rho0)            (SphUtils.cpp:144) */
    float gravity[3];      /* (0,-1,0)                                           (SphUtils.cpp:298) */
    float visc_alpha;      /* 0.1f  artificial viscosity, linear term            (SphUtils.cpp:226) */
    float visc_beta;       /* 0.2f  quadratic term                               (SphUtils.cpp:226) */
    float visc_eta;        /* 0.01f: mu denominator r*r + eta*h*h                (SphUtils.cpp:225) */
    float restitution;     /* 0.0f: impulse factor (1+restitution)               (SphUtils.cpp:352) */
    float cell_size;       /* grid cell edge; 0 => support radius h*h.  Must be >= h*h/2 */
    double wall_test_off;  /* 0.001: hit if dot - off <= wall_test_lim, compared in double like the reference's
                              double literals                                    (SphUtils.cpp:349) */
    double wall_test_lim;  /* 0.01 */
    int32_t domain_mode;   /* SPHB_DOMAIN_* */
    float domain_lo[3];    /* used when domain_mode == SPHB_DOMAIN_FIXED */
    float domain_hi[3];
    int32_t periodic_reserved[3]; /* must be 0.  Periodic boxes (an extension: the reference has only planes) are set up through
                              sphb_slab -- shift_left/right for x, periodic_y/z for the other two -- and sphb_mg_pack, where they
                              become ghost images; a single engine is then its own x-neighbour.  The kernels carry no
                              minimum-image code */
    int64_t max_cells;     /* capacity of the cell table; 0 => chosen from the particle count at upload */
    int32_t write_accel;   /* 1 => sphb_step also stores per-particle acceleration (parity / viewers); 0 => skipped */
    int32_t cell_x_div;    /* the grid cell is cell_size / cell_x_div long in x (cell_size in y and z): the cells of a
                              (y,z) row are contiguous in memory, so a finer x pitch only trims the candidate run
                              from 3 cells to 2R + 2 cell_x.  0 => 2 (fastest measured on B200: finer pitches cost more in
                              the sort and in L1 wavefronts than they save in tests).  1 => cubic cells */
    int32_t laplacian_viscosity; /* 1 => update_forces also adds the reference's viscosity_forces term
                              accel_i += lap_visc_coef * m_j / (rho_i rho_j) * lapW_ij * (v_i - v_j)   (SphUtils.cpp:244-290,
                              277-278), on the full neighbourhood.  0 (default): off, as in the reference, whose call is
                              commented out (SphUtils.cpp:175).  When on, sphb_step runs the phases unfused */
    float lap_visc_coef;   /* 0.1f (SphUtils.cpp:277) */
    int32_t parity_mode;   /* 1 => update_density / update_forces (and sphb_step, unfused) evaluate every pair the way the reference's own
                              loops do: neighbours in ascending original index, each float operation of SphUtils.cpp:55-79, 100-149,
                              181-239 with its rounding, W and gradW through double -- rho, P and accel then equal the reference's BIT FOR
                              BIT.  ~50x slower than the fast kernels: for refereeing and debugging.  0 (default): the fast kernels,
                              equal to the reference within 1e-5 (tests/test_gpu_parity.py) */
    int32_t reserved[3];
} sphb_params;

/* Fill *p with the reference's constants (main.cpp:30-41 and the literals of SphUtils.cpp). */
void sphb_default_params(sphb_params* p);

/* Last error message of the calling thread ("" if none). */
const char* sphb_last_error(void);

/* Library / build identification, e.g. "sphb200 0.1 sm_100a". */
const char* sphb_version(void);

/* sizeof(sphb_params) as the library was built: lets a foreign-language binding check its struct layout. */
size_t sphb_params_size(void);

/* Replaces `SphUtils sph(h, rho0, c, dt, eps)` (main.cpp:62, SphUtils.hpp:20-29).  device = CUDA ordinal. */
int sphb_create(const sphb_params* p, int device, sphb_engine** out);
void sphb_destroy(sphb_engine* e);

/* Walls as the `walls` vector of collision() (SphUtils.cpp:343-372, main.cpp:508-540), in vector order.
 * `walls` points at `count` records of `stride` bytes whose first six floats are normal.xyz, center.xyz
 * (stride 32 for the reference's Wall, 24 for packed floats).  count <= SPHB_MAX_WALLS. */
int sphb_set_walls(sphb_engine* e, const void* walls, size_t count, size_t stride);

/* Rigid spheres as the `spheres` vector of update_forces / update_posvel / collision (SphUtils.cpp:163-166,
 * 300-303, 331-334, 375-385; main.cpp:542-552).  `spheres` points at `count` records of `stride` bytes laid out
 * like the reference's Sphere (44 B: pos, vel, accel, radius, mass); count <= SPHB_MAX_SPHERES, 0 removes them.
 * The engine keeps them on the device: update_forces sets their acceleration to gravity, update_posvel
 * integrates them, collision replays the particle-sphere velocity exchanges in the reference's serial order
 * (ascending particle index, spheres in vector order).  sphb_get_spheres copies them back (synchronises). */
int sphb_set_spheres(sphb_engine* e, const void* spheres, size_t count, size_t stride);
int sphb_get_spheres(sphb_engine* e, void* spheres, size_t count, size_t stride);

/* Host AoS -> device SoA.  `particles` points at n records of `stride` bytes laid out like the
 * reference's Particle (stride 52).  pos, vel, mass are taken; accel, density, pressure are taken too so
 * that a single phase can be run on a state produced elsewhere.  n may change between calls
 * (main.cpp:441 grows the vector).  Particle order is remembered: downloads return it. */
int sphb_upload_aos(sphb_engine* e, const void* particles, size_t n, size_t stride);

/* Device SoA -> host AoS, original order, only the fields in field_mask are taken from the engine.  n may be smaller than
 * sphb_count: the first n particles of the original order are returned.  The records are rebuilt from the engine's copy of the
 * last upload plus the masked fields, so a field OUTSIDE the mask keeps the value it had at that upload: if the host has edited
 * such a field since, upload again before relying on it (with no upload to go by, the host records are fetched first). */
int sphb_download_aos(sphb_engine* e, void* particles, size_t n, size_t stride, uint32_t field_mask);

/* Positions only, packed x,y,z (12 B per particle), original order: what the viewer's update_positions() copies into
 * its vertex buffer every frame (main.cpp:97-110).  A quarter of the PCIe traffic of a full-record download. */
int sphb_download_positions(sphb_engine* e, float* xyz, size_t n);

/* --- the five public methods of SphUtils (SPH/SphUtils.hpp:40-44), on the device-resident state ---- */

/* update_cells (SphUtils.cpp:8-47): bounding box, cell coordinates, stable counting sort; the state is
 * permuted into cell order.  Called implicitly by sphb_update_density when positions changed. */
int sphb_update_cells(sphb_engine* e);
/* update_density (SphUtils.cpp:100-149): rho_i = sum_{j != i} m_j W_ij, clamp up to rho0, P = k(rho - rho0). */
int sphb_update_density(sphb_engine* e);
/* update_forces (SphUtils.cpp:157-239, 295-304): accel = gravity + pressure/artificial-viscosity pair sum. */
int sphb_update_forces(sphb_engine* e);
/* update_posvel (SphUtils.cpp:321-335): vel += dt*accel; pos += dt*vel. */
int sphb_update_posvel(sphb_engine* e);
/* collision (SphUtils.cpp:343-372), wall part: sequential plane test/response per particle. */
int sphb_collision(sphb_engine* e);

/* step_scene (main.cpp:280-303) x nsteps, fused fast path: sort, density, then ONE kernel doing
 * force + integrate + wall collision.  Asynchronous: returns after enqueueing; use sphb_synchronize. */
int sphb_step(sphb_engine* e, int nsteps);

/* The whole reference-facing call on HOST buffers: upload, nsteps steps, download(field_mask), synchronised. */
int sphb_step_aos(sphb_engine* e, void* particles, size_t n, size_t stride, int nsteps, uint32_t field_mask);

int sphb_synchronize(sphb_engine* e);

/* Page-lock / release a host range (e.g. the std::vector<Particle> buffer the viewer owns) so that
 * uploads and downloads run at full PCIe rate.  Optional: pageable memory works, slower. */
int sphb_host_register(void* ptr, size_t bytes);
int sphb_host_unregister(void* ptr);

/* kernel_function / grad_kernel / lap_kernel (SphUtils.cpp:55-95) evaluated ON THE DEVICE with the same
 * device functions the step kernels use, for `count` point pairs.  xi, xj: 3*count floats (host).
 * Any of w (count), grad (3*count), lap (count) may be NULL. */
int sphb_eval_kernels(sphb_engine* e, const float* xi, const float* xj, size_t count, float* w, float* grad, float* lap);

/* Grid of the last sort, for the bit-exact grid test: cell_xyz[3n] in ORIGINAL particle order,
 * perm[n] = original index held by sorted slot s, dims[3], origin[3], cell[3] (edge per axis).  Any may be NULL. */
int sphb_get_cells(sphb_engine* e, int32_t* cell_xyz, int32_t* perm, int32_t* dims, float* origin, float* cell);

/* Bounding box {minx,miny,minz,maxx,maxy,maxz} of the current positions (SphUtils.cpp:11-23). */
int sphb_get_bbox(sphb_engine* e, float bbox[6]);

size_t sphb_count(const sphb_engine* e);

/* Number of particles whose position or velocity is NaN or infinite (the reference propagates NaN silently once
 * two particles coincide, SphUtils.cpp:75; this lets a driver notice).  Synchronises. */
int sphb_count_nonfinite(sphb_engine* e, size_t* count);

/* Per-phase GPU time, measured with CUDA events on the engine's stream around the kernels of each
 * sphb_step step (no host synchronisation is added).  sphb_enable_timing(e, 1) starts (and resets) the
 * record, up to 256 steps; sphb_get_timing synchronises and returns the AVERAGE ms per step over the
 * recorded steps: [0] grid+sort+reorder, [1] density, [2] force+integrate+collide, [3] whole step;
 * *steps (may be NULL) = how many steps were averaged. */
int sphb_enable_timing(sphb_engine* e, int on);
int sphb_get_timing(sphb_engine* e, float ms[4], int* steps);

/* Kernel launches issued by this engine since creation (for bench.py's gpu_launches). */
uint64_t sphb_launch_count(const sphb_engine* e);

/* Counters since the engine was created (nothing like them in the reference, which has no diagnostics at all).
 * arrival_order_cells: how many times a grid cell held more than 4,096 particles at sort time and therefore kept the
 * (not reproducible) arrival order instead of the stable one -- only pathological states do this: everything clamped
 * into one boundary cell, or a state gone NaN.  0 means every sort so far produced the stable permutation.  Synchronises. */
typedef struct sphb_stats {
    uint64_t arrival_order_cells;
    uint64_t steps;      /* sphb_step steps run */
    uint64_t launches;   /* = sphb_launch_count */
    uint64_t particles;  /* = sphb_count */
    uint64_t capacity;   /* particle records the arrays can hold without growing */
    uint64_t reserved[3];
} sphb_stats;
int sphb_get_stats(sphb_engine* e, sphb_stats* out);

/* --- device-resident access for the headless bench and the multi-GPU driver ----------------------- */

/* Upload from DEVICE SoA arrays (float4 posm = x,y,z,mass; float4 vel = vx,vy,vz,anything: the engine keeps its ownership flag
 * there and sets it itself), n particles,
 * device-to-device on the engine's stream.  ids (may be NULL = 0..n-1) names each particle's place in the "original
 * order" that downloads return: it must be a permutation of 0..n-1.  Anything else is refused with SPHB_ERR_INVALID
 * (checked on the device; that check synchronises) and leaves the engine without a state. */
int sphb_upload_soa_device(sphb_engine* e, const void* posm, const void* vel, const int32_t* ids, size_t n);
/* Replace the particle ids (0..n-1 in upload order after an upload) by the caller's own labels, ids[k] for the k-th uploaded
 * particle: any int32 values.  From then on the labels TRAVEL with the particles through the slab exchange (a message record
 * carries its particle's label), so a driver that gave every particle of a multi-GPU run a unique label can find each one again
 * on whichever rank it ended up (sphb_download_soa returns the labels); sphb_multi_* does exactly that.  In exchange the
 * calls that return particles in upload order (sphb_download_aos, sphb_download_positions, sphb_render_depth) are refused for
 * this engine until its next upload.  Without this call ids restart at the slot numbers at every exchange. */
int sphb_set_ids(sphb_engine* e, const int32_t* ids, size_t n);

/* Current device pointers (sorted order, valid until the next call that steps or uploads):
 * posm float4, vel float4, aux float4 (density, pressure, sqrt|P/rho|, 1/rho), accel float4, ids int32. */
int sphb_get_device_state(sphb_engine* e, void** posm, void** vel, void** aux, void** accel, int32_t** ids);
/* Copy the current state to host arrays in SORTED order (sphb_count records each; any pointer may be NULL):
 * posm/vel/aux/accel 4 floats per particle, ids int32.  Synchronises. */
int sphb_download_soa(sphb_engine* e, float* posm, float* vel, float* aux, float* accel, int32_t* ids);
/* The CUDA stream (cudaStream_t) all of this engine's work is enqueued on. */
void* sphb_stream(sphb_engine* e);

/* --- headless point-sprite renderer (the arithmetic of SPH/Shaders/sph.vertexshader + sph.fragmentshader) ------ */
/* The viewer draws every particle as a GL point sprite whose fragment shader turns it into a sphere impostor and
 * writes its clip-space depth as a grey value (main.cpp:229-252, sph.vertexshader:8-12, sph.fragmentshader:7-24).
 * No graphics API exists on the target machines; this is the same arithmetic as a CUDA kernel into a plain float
 * image.  Per particle: clip = MVP (pos, 1); culled unless -w <= x,y,z <= w; window centre ((x/w + 1)/2 W,
 * (y/w + 1)/2 H); sprite edge = point_scale / clip.w pixels, clamped to [1, 2048]; a pixel whose centre lies in the
 * sprite square gets N.xy = 2 (s,t) - 1 with (s,t) = 1/2 + (pixel centre - window centre) / edge (origin lower
 * left, main.cpp:565), is discarded if |N.xy|^2 > 1, else N.z = sqrt(1 - |N.xy|^2) and
 * value = clamp((MVP (pos + sprite_radius N, 1)).z / .w, 0, 1).  No depth buffer is requested by the viewer
 * (GLUT_DOUBLE | GLUT_RGB, main.cpp:730), so by default a later particle overwrites an earlier one: particle order decides
 * (sphb_camera.depth_test = 1 gives the other possible outcome, see there). */
typedef struct sphb_camera {
    float mvp[16];        /* column-major, like glm::mat4 / glUniformMatrix4fv(..., GL_FALSE, ...) */
    int32_t width, height;
    float point_scale;    /* 50   (sph.vertexshader:10) */
    float sprite_radius;  /* 0.1  (sph.fragmentshader:19) */
    float clear_value;    /* 1.0  (glClearColor(1,1,1,1), main.cpp:230) */
    int32_t depth_test;   /* 0 (default): no depth buffer -- the viewer asks for GLUT_DOUBLE | GLUT_RGB (main.cpp:730), so the particle drawn
                             last wins a pixel.  1: as if the visual carried depth bits after all: glEnable(GL_DEPTH_TEST), GL_LESS
                             (main.cpp:567-571) then make the sprite with the smallest point depth (clip.z / clip.w + 1) / 2 win, the
                             first drawn on a tie (the fragment shader writes no gl_FragDepth).  Which of the two a real GL
                             implementation shows is its choice of visual; both are available here */
} sphb_camera;

/* main.cpp:48-60: perspective(45.0f [radians in GLM 0.9.6], 4/3, 0.1, 100) * lookAt((0,-0.8,1.5), (0,-0.8,0),
 * (0,1,0)) * identity; 1080 x 920 (main.cpp:13-14). */
void sphb_default_camera(sphb_camera* cam);

/* Renders the current particle positions (original order) into image[height][width] (row 0 = bottom, as GL). */
int sphb_render_depth(sphb_engine* e, const sphb_camera* cam, float* image);

/* --- multi-GPU: x-slab decomposition, one engine (one process) per GPU ------------------------------- */
/* Nothing like this exists in the reference (single thread).  Each rank owns the particles
 * whose x lies in [lo, hi) and keeps a ghost layer of width `halo` (2 support radii, so that ghost densities
 * can be recomputed locally and ONE exchange per step suffices) copied from its two x-neighbours.  Ownership
 * is carried in vel.w (1 = owned, 0 = ghost).  Per step:
 *     sphb_mg_pack  ->  exchange the two messages with the neighbours (NCCL send/recv, or peer copies)
 *                   ->  sphb_mg_unpack  ->  sphb_step(e, 1)
 * A message is (capacity + 1) records of 32 bytes: a header {int32 count, int32 capacity, ...} and then
 * {float4 posm, float4 vel} per particle, all in device memory owned by the caller. */
typedef struct sphb_slab {
    float lo, hi;             /* owned x-interval [lo, hi); use -INFINITY / +INFINITY at the ends of the domain */
    float halo;               /* ghost width; >= 2*h*h for the single-exchange scheme */
    int32_t has_left, has_right;
    float shift_left, shift_right; /* added to x of records sent that way: +-period across a periodic seam, else 0 */
    int32_t periodic_y, periodic_z; /* periodic box in y / z (extension: the reference has only planes).  Owned particles
                                  are wrapped into [lo, lo+len) and those within `halo` of a periodic face are copied as
                                  ghost images shifted by the period, so the kernels need no minimum-image logic */
    float ylo, ylen, zlo, zlen;
} sphb_slab;

/* Bytes of a message buffer that can carry `capacity` particles. */
size_t sphb_mg_message_bytes(size_t capacity);

/* Make room for up to n_max particles (owned + TWO generations of ghosts: the last step's stay in the arrays,
 * marked dead, until the next step's sort drops them) before the first upload, so that the arrays need not
 * grow when neighbours send more than this rank started with. */
int sphb_reserve(sphb_engine* e, size_t n_max);

/* Classify the particles this rank owned by their current x: the ones it keeps (as owned or as ghosts) stay in
 * their slots, the ones a neighbour needs (its ghost zone, and migrants) are copied into
 * msg_left / msg_right (either may be NULL when that neighbour does not exist).  Previous ghosts are dropped
 * (marked dead in place; a negative mass means "dead" to an engine that has used sphb_mg_pack).
 * Deterministic (order-preserving compaction of the messages).  Asynchronous on the engine's stream. */
int sphb_mg_pack(sphb_engine* e, const sphb_slab* slab, void* msg_left, void* msg_right, size_t capacity);
/* The same with a capacity per message (see sphb_mg_sent). */
int sphb_mg_pack_limits(sphb_engine* e, const sphb_slab* slab, void* msg_left, size_t cap_left, void* msg_right, size_t cap_right);

/* Append what the neighbours sent (NULL = none), refresh the ownership flags and adopt the new particle
 * set.  Synchronises once to read the counts.  counts[4] (may be NULL) = {records before the messages (the dropped
 * ones stay in the arrays, marked dead with mass -1, until the next step's sort removes them -- reserve room for
 * them), from_left, from_right, total}.  sphb_count() excludes the dead again after the next step.
 * SPHB_ERR_CAPACITY if a sender overflowed its message or the arrays reserved with sphb_reserve are full; that is final (the
 * senders have already given the particles up): the engine drops its state and must be uploaded again. */
int sphb_mg_unpack(sphb_engine* e, const sphb_slab* slab, const void* from_left, const void* from_right, size_t counts[4]);

/* The same without waiting for the GPU: nothing is read back, so the host can enqueue the next step at once and the GPU never
 * idles between steps.  Until some call needs the exact count (sphb_count, downloads, ... -- they synchronise and fetch it),
 * the engine works with an ESTIMATE of its particle count for sizing launches while its kernels read the real count from
 * device memory.  The verdict of the exchange -- its counts, and whether a message or the arrays overflowed -- arrives one call
 * late: sphb_mg_poll returns it, and the next sphb_mg_pack reports an overflow as SPHB_ERR_CAPACITY.  An overflow is final
 * either way (the senders have already given the particles up): the engine drops its state and must be uploaded again. */
int sphb_mg_unpack_async(sphb_engine* e, const sphb_slab* slab, const void* from_left, const void* from_right);

/* Status of the latest sphb_mg_unpack_async whose result has reached the host: status[0..6] = {records before the messages,
 * from_left, from_right, total, dead among them, sent left, sent right}, status[7] = how many asynchronous unpacks have reported so
 * far.  wait != 0 blocks until the latest one has (its event sits before the kernels of the following step, so this does not
 * drain the GPU); wait == 0 returns the newest record available.  SPHB_ERR_CAPACITY if that exchange overflowed. */
int sphb_mg_poll(sphb_engine* e, size_t status[8], int wait);

/* Records the last sphb_mg_pack wanted to put into its left / right message (known once sphb_mg_unpack has read the
 * counts back).  Both ends of a link see the same number -- the sender here, the receiver in counts[1] / counts[2] --
 * so they can agree, without talking, to transfer only a prefix of the message buffers next step (pass the smaller
 * capacity to sphb_mg_pack; sph_b200/multigpu.py does). */
int sphb_mg_sent(sphb_engine* e, size_t sent[2]);

/* --- multi-GPU: peer-to-peer halo link --------------------------------------------------------------- */
/* The exchange without a communication library in the data path: each rank owns a block of device memory with its four
 * receive buffers (from the left / from the right, two of each so that consecutive steps alternate) and a few flags, the
 * neighbours map that block (CUDA IPC between processes, plain peer access inside one process), and sphb_mg_exchange runs
 *     wait until the neighbours have consumed what these buffers held two steps ago
 *     pack -- the kernel writes the two messages STRAIGHT INTO the neighbours' receive buffers over NVLink
 *     raise their "arrived" flags, wait for mine, append what arrived (as sphb_mg_unpack_async: nothing is read back)
 *     raise their "consumed" flags
 * all as kernels on the engine's stream; the host never waits.  Only the records that exist travel (no fixed-size transfer,
 * no size negotiation).  Flags only grow, waits give up after ~4 s of GPU time (reported by sphb_mg_poll / the next
 * exchange as SPHB_ERR_STATE) instead of hanging the device.  NCCL (or any other transport) remains possible through
 * sphb_mg_pack / sphb_mg_unpack; nothing like either exists in the reference. */
typedef struct sphb_mg_link sphb_mg_link;
#define SPHB_MG_EXPORT_BYTES 64

/* Allocate the block for messages of up to `capacity` particles per neighbour and step, on the engine's GPU.  Create the
 * link AFTER sphb_reserve / the first upload: it also allocates the exchange's scratch for the arrays' current capacity, so that
 * sphb_mg_exchange itself allocates nothing (an allocation can wait for the whole device, which a flag wait may be holding). */
int sphb_mg_link_create(sphb_engine* e, size_t capacity, sphb_mg_link** out);
void sphb_mg_link_destroy(sphb_mg_link* l);
/* A handle other PROCESSES can open (cudaIpcMemHandle_t, zero padded): send it to the two neighbour ranks. */
int sphb_mg_link_export(sphb_mg_link* l, unsigned char handle[SPHB_MG_EXPORT_BYTES]);
/* Connect to the neighbours' blocks from their exported handles (NULL = no neighbour on that side). */
int sphb_mg_link_connect_ipc(sphb_mg_link* l, const unsigned char* left, const unsigned char* right);
/* The same inside ONE process (one host driving several GPUs): the neighbours' links themselves; peer access between the
 * devices is enabled here.  A lone rank in a periodic box passes its own link on both sides. */
int sphb_mg_link_connect_local(sphb_mg_link* l, sphb_mg_link* left, sphb_mg_link* right);
/* One halo exchange (see above); follow it with sphb_step(e, 1).  All ranks of a chain must call it the same number of times. */
int sphb_mg_exchange(sphb_engine* e, sphb_mg_link* l, const sphb_slab* slab);

/* --- several GPUs behind one handle ------------------------------------------------------------------- */
/* For a C or C++ host that wants more than one GPU without writing the slab loop itself (the reference's own main.cpp is such a
 * host: one thread, SPH/main.cpp:280-303; nothing like this exists in the reference).  One host thread, `ndev` GPUs of one
 * box: the particles are cut into x-slabs of equal owned-plus-ghost counts at upload, every GPU gets an engine and a
 * peer-to-peer halo link to its neighbours, and sphb_multi_step enqueues, per step, every GPU's exchange and then every GPU's
 * fused step without ever waiting.  Particle order: downloads return the records in the order they were uploaded (every
 * particle carries its upload index as a label through the exchanges).  The cuts stay where the upload put them; spheres
 * and periodic boxes are not supported here.  devices == NULL means 0 .. ndev-1. */
typedef struct sphb_multi sphb_multi;
int sphb_multi_create(int ndev, const int* devices, const sphb_params* p, sphb_multi** out);
void sphb_multi_destroy(sphb_multi* m);
const char* sphb_multi_last_error(void);
int sphb_multi_set_walls(sphb_multi* m, const void* walls, size_t count, size_t stride);
int sphb_multi_upload_aos(sphb_multi* m, const void* particles, size_t n, size_t stride);
int sphb_multi_step(sphb_multi* m, int nsteps);          /* asynchronous */
int sphb_multi_synchronize(sphb_multi* m);
int sphb_multi_download_aos(sphb_multi* m, void* particles, size_t n, size_t stride, uint32_t field_mask);
int sphb_multi_step_aos(sphb_multi* m, void* particles, size_t n, size_t stride, int nsteps, uint32_t field_mask);
/* The slab cuts sphb_multi_upload_aos would choose for these x coordinates (host arithmetic only, no GPU involved): ndev+1
 * increasing values, the first and last being the smallest and largest x; inner cuts sit on multiples of h*h from the first and
 * equalise owned-plus-ghost particle counts (ghost layers 2 h*h wide). */
int sphb_multi_plan(const float* x, size_t n, int ndev, float h, double* bounds);
/* ndev, per-GPU record counts (owned + ghosts; synchronises), and the ndev+1 slab cuts.  Any pointer may be NULL. */
int sphb_multi_info(sphb_multi* m, int* ndev, size_t* owned_plus_ghosts, double* bounds);

/* --- allocation-free bounding box ------------------------------------------------------------------ */

/* Largest point count sphb_bbox_noalloc accepts (its scratch is a fixed array inside the library). */
#define SPHB_NOALLOC_MAX_POINTS 4194304

/*
 * Bounding box of `n` device-resident points without allocating anything.  (No counterpart in the
 * reference: it exists for callers that are not allowed to allocate GPU memory at the moment they need the
 * box -- a render or real-time thread that must not hit the device-wide synchronisation cudaMalloc/cudaFree
 * imply, a plug-in running under a host application's fixed memory budget, code running after the
 * application has handed all remaining memory to another library.  Every other entry point of this library
 * works on an engine, and engines own buffers that they create and grow with cudaMalloc.)
 *
 *   posm_device  n float4 records (x, y, z, anything) in device memory of the CURRENT device, e.g. the
 *                `posm` pointer of sphb_get_device_state, or a mapped vertex buffer
 *   bbox_device  6 floats in device memory, written as {minx, miny, minz, maxx, maxy, maxz}; NaN
 *                coordinates are skipped like sphb_get_bbox does; with no finite point the result is
 *                {+inf x3, -inf x3}
 *   stream       cudaStream_t to enqueue on (NULL = the default stream).  The call returns after
 *                enqueueing; the result is ready when the stream reaches that point.
 *
 * The per-block partial results live in one file-scope __device__ array of the library
 * (SPHB_NOALLOC_MAX_POINTS / 1024 blocks x 32 B = 128 KB, plus a 256-byte control block), which every GPU
 * that loads the library holds until the process exits.  Consequences the caller must respect:
 *   - n > SPHB_NOALLOC_MAX_POINTS (or n == 0) is refused with SPHB_ERR_CAPACITY / SPHB_ERR_INVALID;
 *   - all calls on one GPU share that array: two calls whose kernels may run at the same time (different
 *     streams, different host threads) race and can return wrong boxes -- serialise them;
 *   - the FIRST call on each device looks the array's device address up (cudaGetSymbolAddress, with that
 *     device current) and caches it; make that first call before starting any stream capture.
 * Calls no cudaMalloc / cudaFree / library allocator, creates no stream or event.
 */
int sphb_bbox_noalloc(const void* posm_device, size_t n, void* stream, float* bbox_device);

#ifdef __cplusplus
}
#endif
#endif /* SPHB200_H */
