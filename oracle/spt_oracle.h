/*
 * oracle/spt_oracle.h -- TEST INFRASTRUCTURE ONLY.
 *
 * Plain-C restatement of the path-tracing hot path of ilia-glushchenko/SimplePathTracer
 * (SURVEY.md section 8a).  It exists to CHECK the CUDA path; only tests/,
 * __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may use it.
 * The product (simplepathtracer_b200/, include/) never includes, links or calls it.
 *
 * Pinning: this restatement is pinned bit-for-bit against the UNMODIFIED reference
 * compiled here (oracle/_ref/libspt_ref.so, built by oracle/Makefile from
 * /root/reference/include): first-hit index/t/normal, scene generators, and whole
 * renders of RenderSegment / RenderSegmentTask with the reference's own RNG sequence
 * (tests/test_oracle_vs_ref.py; frozen outputs in tests/golden/).  The reference holds
 * no golden vectors or tests of its own.
 *
 * Every float operation is a single IEEE-754 binary32 operation in the reference's
 * order; build with -ffp-contract=off (no FMA), no -ffast-math.
 */
#ifndef SPT_ORACLE_H
#define SPT_ORACLE_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* Material values: /root/reference/include/Definitions.hpp:7-13 */
enum { SPO_SKYBOX = 0, SPO_REFLECTIVE = 1, SPO_REFRACTIVE = 2, SPO_DIFFUSE = 3 };

/* SoA scene exactly as the reference globals (include/Globals.hpp:31-37): Vec4 = 4 packed floats */
typedef struct spo_scene {
    uint32_t n;
    const float *centers4;    /* g_spheres   n*4 */
    const float *radii;       /* g_radii     n   */
    const uint8_t *materials; /* g_materials n   */
    const float *colors4;     /* g_colors    n*4 */
    const float *fuzz;        /* g_diffuses  n   */
} spo_scene;

typedef struct spo_camera {
    float view[16]; /* viewMatrix.array, row-major (include/Globals.hpp:21, Renderer.hpp:321) */
    float eye[4];   /* eyePos   (include/Globals.hpp:23) */
    float sky[4];   /* initColor (include/Globals.hpp:29) */
} spo_camera;

enum { SPO_RNG_COUNTER = 0, SPO_RNG_REFSEQ = 1 };
enum { SPO_POW_EXACT = 0, SPO_POW_LIBM = 1 };
enum { SPO_TRACER_RECURSIVE = 0, SPO_TRACER_TASK = 1 };

typedef struct spo_params {
    uint32_t width, height;   /* g_width, g_height (include/Globals.hpp:14-15) */
    uint32_t depth;           /* g_bounces (include/Globals.hpp:12) */
    uint32_t pass_cap;        /* 0 = recursive tracer (unbounded specular recursion);
                                 10 = task tracer's literal pass count (TaskBasedPathTracer.hpp:81) */
    uint32_t sample_begin, sample_end; /* sample indices [begin,end) of each pixel (counter RNG mode) */
    uint64_t seed;            /* counter RNG mode: 64-bit seed */
    int pow_mode;             /* SPO_POW_EXACT: x^5 in binary64 rounded once (matches the CUDA path bit for bit);
                                 SPO_POW_LIBM: powf(x,5.f) as the compiled reference calls it */
} spo_params;

typedef struct spo_counters {
    uint64_t segments;     /* FindClosestIntersectionSphere calls */
    uint64_t tests;        /* ray-sphere boolean tests = segments * n */
    uint64_t passed;       /* tests that passed RaySphereIntersection */
    uint64_t samples;      /* camera samples traced */
    uint64_t dropped;      /* samples dropped by pass_cap */
    uint64_t rng_draws;
} spo_counters;

/* ---- RNG (include/Random.hpp:11-46, :86-93, :115-127) ---- */
uint32_t spo_splitmix_next(uint64_t *state);
void spo_splitmix_pair(uint64_t *state, uint32_t *hi, uint32_t *lo); /* counter front end: both halves */
uint64_t spo_splitmix_seed(uint32_t seed32);             /* seed << 31 | seed  (Random.hpp:20) */
float spo_u32_to_unit(uint32_t draw);                    /* libstdc++ generate_canonical<float,24> */
uint64_t spo_stream_key(uint64_t seed, uint32_t pixel, uint32_t sample);
uint64_t spo_event_state(uint64_t key, uint32_t event);
void spo_rng_uniform_seq(uint32_t seed32, float lo, float hi, uint32_t count, float *out);
void spo_rng_shell_seq(uint32_t seed32, uint32_t count, float *out3);

/* ---- math / camera (include/Math.hpp:107-231) ---- */
void spo_camera_basis(const float eye[3], const float look[3], const float up[3], float view16[16]);
float spo_pow5(float x, int pow_mode);

/* ---- collision (include/Collision.hpp:9-109) ---- */
uint32_t spo_closest_hit(const spo_scene *sc, const float d[4], const float o[4]);
void spo_closest_hit_many(const spo_scene *sc, uint32_t count, const float *dir3, const float *org3, uint32_t *idx);
void spo_first_hit(const spo_scene *sc, const spo_camera *cam, uint32_t width, uint32_t height,
                   uint32_t xBegin, uint32_t xEnd, uint32_t yBegin, uint32_t yEnd,
                   uint32_t *idx, float *t, float *normal3, float *dir3);

/* ---- tracers ---- */
/* Counter-RNG render: for every pixel of the half-open segment adds, in sample order, the radiance of
 * samples [sample_begin,sample_end) into accum4[(y*width+x)*4+0..2] and the number of samples that
 * were not dropped into [..+3].  This is the function the CUDA kernels must reproduce bit for bit. */
void spo_render_accum(const spo_scene *sc, const spo_camera *cam, const spo_params *p,
                      uint32_t yBegin, uint32_t yEnd, uint32_t xBegin, uint32_t xEnd, float *accum4,
                      spo_counters *ctr);
/* radiance of one (pixel, sample) in counter mode; returns 0 if the sample was dropped */
int spo_sample_radiance(const spo_scene *sc, const spo_camera *cam, const spo_params *p,
                        uint32_t x, uint32_t y, uint32_t sample, float rgb[4], spo_counters *ctr);

/* Reference-sequence render: replays RenderSegment (tracer 0, SingleThreadPathTracer.hpp:114-137) or
 * RenderSegmentTask (tracer 1, TaskBasedPathTracer.hpp:54-206) in the reference's loop and queue order
 * drawing from ONE splitmix stream seeded like Random.hpp:88-89 with seed32, so that the output equals
 * oracle/_ref's bit for bit.  mean4 (W*H*4) gets the linear colour handed to io::WritePixel at the
 * flipped pixel slot; rgb8 (W*H*3) gets g_data. */
void spo_render_refseq(const spo_scene *sc, const spo_camera *cam, uint32_t width, uint32_t height,
                       uint32_t spp, uint32_t depth, int tracer, int pow_mode,
                       uint32_t yBegin, uint32_t yEnd, uint32_t xBegin, uint32_t xEnd, uint32_t seed32,
                       float *mean4, uint8_t *rgb8, spo_counters *ctr);

/* ---- output (include/IOHelpers.hpp:17-22; index SingleThreadPathTracer.hpp:120) ---- */
void spo_write_pixel(const float rgb[3], uint8_t out[3]);
/* accum4 (unflipped, y*W+x) -> rgb8 at the reference's flipped slot, colour = sum * (1/count) */
void spo_resolve(const float *accum4, uint32_t width, uint32_t height,
                 uint32_t yBegin, uint32_t yEnd, uint32_t xBegin, uint32_t xEnd, uint8_t *rgb8);

/* ---- scene generators (include/SceneGenerators.hpp:6-133) with the reference RNG sequence ---- */
/* Each returns the sphere count and fills caller arrays of capacity cap (returns needed n if cap too small). */
uint32_t spo_init_spheres(uint32_t seed32, uint32_t cap, float *centers4, float *radii, uint8_t *materials,
                          float *colors4, float *fuzz);
uint32_t spo_generate_spheres(uint32_t seed32, uint32_t cap, float *centers4, float *radii, uint8_t *materials,
                              float *colors4, float *fuzz);

#ifdef __cplusplus
}
#endif
#endif
