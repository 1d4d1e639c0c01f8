/*
 * oracle/spt_oracle.c -- TEST INFRASTRUCTURE ONLY (see spt_oracle.h).
 *
 * CPU restatement of the reference's tracing hot path.  Citations are into
 * /root/reference/.  Each float operation below is one IEEE binary32 operation,
 * written in the order the reference's SSE code performs it; compile with
 * -ffp-contract=off.  Vec4 is carried as 4 floats (w is always 0 on this path) so
 * that the dpps / haddps summation order (x+y)+(z+w) is literal.
 */
#include "spt_oracle.h"

#include <float.h>
#include <math.h>
#include <stdlib.h>
#include <string.h>

typedef struct v4 { float v[4]; } v4;

/* ------------------------------------------------------------------ Math.hpp */
/* Vec4::operator- (Math.hpp:16-21) */
static v4 vsub(v4 a, v4 b) { v4 r = {{a.v[0] - b.v[0], a.v[1] - b.v[1], a.v[2] - b.v[2], a.v[3] - b.v[3]}}; return r; }
/* Vec4::operator+ (Math.hpp:33-38) */
static v4 vadd(v4 a, v4 b) { v4 r = {{a.v[0] + b.v[0], a.v[1] + b.v[1], a.v[2] + b.v[2], a.v[3] + b.v[3]}}; return r; }
/* Vec4::operator*(float) (Math.hpp:50-54) */
static v4 vscale(v4 a, float s) { v4 r = {{a.v[0] * s, a.v[1] * s, a.v[2] * s, a.v[3] * s}}; return r; }
/* Vec4::operator-() (Math.hpp:66-70): sign flip */
static v4 vneg(v4 a) { v4 r = {{-a.v[0], -a.v[1], -a.v[2], -a.v[3]}}; return r; }
static v4 v3(float x, float y, float z) { v4 r = {{x, y, z, 0.f}}; return r; }
static v4 vload4(const float *p) { v4 r = {{p[0], p[1], p[2], p[3]}}; return r; }

/* Dot = _mm_dp_ps mask 0xF1 (Math.hpp:107-111): (x*x' + y*y') + (z*z' + w*w') */
static float vdot(v4 a, v4 b)
{
    float const p0 = a.v[0] * b.v[0], p1 = a.v[1] * b.v[1], p2 = a.v[2] * b.v[2], p3 = a.v[3] * b.v[3];
    float const s01 = p0 + p1, s23 = p2 + p3;
    return s01 + s23;
}

/* LengthSquared = mul, hadd, hadd (Math.hpp:122-133): same (x+y)+(z+w) order */
static float vlensq(v4 a) { return vdot(a, a); }

/* Length (Math.hpp:135-138) */
static float vlen(v4 a) { return sqrtf(vlensq(a)); }

/* Normalize = v / sqrt(lensq), all four lanes (Math.hpp:140-154) */
static v4 vnormalize(v4 a)
{
    float const l = sqrtf(vlensq(a));
    v4 r = {{a.v[0] / l, a.v[1] / l, a.v[2] / l, a.v[3] / l}};
    return r;
}

/* Reflect = vec - normal * Dot(vec, normal) * 2 (Math.hpp:156-159) */
static v4 vreflect(v4 vec, v4 n) { return vsub(vec, vscale(vscale(n, vdot(vec, n)), 2.f)); }

/* Mat4::operator*(Vec4): four dpps (Math.hpp:178-186) */
static v4 mat_mul(const float m[16], v4 x)
{
    v4 r;
    for (int i = 0; i < 4; ++i)
        r.v[i] = vdot(vload4(m + 4 * i), x);
    return r;
}

/* Cross as written, INCLUDING its z-lane bug (Math.hpp:113-120) */
static v4 vcross_ref(v4 a, v4 b)
{
    return v3(a.v[1] * b.v[2] - a.v[2] * b.v[1], a.v[2] * b.v[0] - a.v[0] * b.v[2], a.v[0] * b.v[0] - a.v[1] * b.v[0]);
}

/* viewMatrix = Transpose(CreateCameraBasisMatrix(eye, lookAt, up)) (Math.hpp:198-231, Renderer.hpp:321) */
void spo_camera_basis(const float eye[3], const float look[3], const float up[3], float view16[16])
{
    v4 const e = v3(eye[0], eye[1], eye[2]), l = v3(look[0], look[1], look[2]);
    v4 upDir = v3(up[0], up[1], up[2]);
    v4 const viewDir = vnormalize(vsub(l, e));
    v4 const right = vnormalize(vcross_ref(upDir, viewDir));
    upDir = vcross_ref(viewDir, right);
    float m[16];
    memcpy(m + 0, right.v, 16);
    memcpy(m + 4, upDir.v, 16);
    memcpy(m + 8, viewDir.v, 16);
    memset(m + 12, 0, 16);
    for (int r = 0; r < 4; ++r)
        for (int c = 0; c < 4; ++c)
            view16[4 * r + c] = m[4 * c + r];
}

float spo_pow5(float x, int pow_mode)
{
    if (pow_mode == SPO_POW_LIBM)
        return powf(x, 5.f); /* pow(1.f - c, 5.f), SingleThreadPathTracer.hpp:60,74 */
    /* x^5 evaluated in binary64 (3 roundings at 2^-53) and rounded once to binary32: the
     * correctly rounded value except on near-ties; identical on CPU and GPU. */
    double const xd = (double)x, x2 = xd * xd, x4 = x2 * x2;
    return (float)(x4 * xd);
}

/* ---------------------------------------------------------------- Random.hpp */
#define SPO_GOLDEN 0x9E3779B97F4A7C15ULL

/* splitmix::operator() (Random.hpp:35-41): note the 31-bit final shifts */
uint32_t spo_splitmix_next(uint64_t *state)
{
    uint64_t z = (*state += SPO_GOLDEN);
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ULL;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBULL;
    return (uint32_t)((z ^ (z >> 31)) >> 31);
}

/* splitmix(uint64_t seed) : m_seed(seed << 31 | seed) with the clock count cast to unsigned int
 * (Random.hpp:20, :88-89) */
uint64_t spo_splitmix_seed(uint32_t seed32)
{
    uint64_t const s = seed32;
    return (s << 31) | s;
}

/* std::uniform_real_distribution<float> on a 32-bit engine as libstdc++ evaluates it
 * (generate_canonical<float,24>: one draw, float(draw) / 2^32, clamped below 1).  The mapping is
 * library-defined (MSVC's differs); it only has to produce U[0,1). */
float spo_u32_to_unit(uint32_t draw)
{
    float r = (float)draw * 2.3283064365386962890625e-10f;
    if (r >= 1.f)
        r = 0.99999994f; /* nextafterf(1, 0) */
    return r;
}

static uint64_t mix64(uint64_t z)
{
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ULL;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBULL;
    return z ^ (z >> 31);
}

/* Counter-based keying that replaces the per-thread, clock-seeded generator (Random.hpp:88-89):
 * one splitmix stream per (pixel, sample, event).  Event 0 = camera jitter; every scatter
 * (diffuse bounce, reflective, refractive) opens the next event.  The pixel part is mixed (one splitmix
 * finaliser per pixel and launch on the GPU); sample and event enter as Weyl increments of the engine's
 * state -- the engine step itself (state += G; output = finaliser(state), Random.hpp:35-41) is what turns
 * distinct (pixel, sample, event, step) counters into independent-looking draws, so a second finaliser per
 * sample would only repeat its work. */
uint64_t spo_stream_key(uint64_t seed, uint32_t pixel, uint32_t sample)
{
    uint64_t const k = mix64(seed + SPO_GOLDEN * ((uint64_t)pixel + 1ULL));
    return k ^ (0xD1B54A32D192ED03ULL * ((uint64_t)sample + 1ULL));
}

uint64_t spo_event_state(uint64_t key, uint32_t event) { return key + 0xA0761D6478BD642FULL * (uint64_t)event; }

typedef struct rng_t {
    uint64_t state;
    uint64_t key;
    uint32_t event;
    int counter_mode;
    spo_counters *ctr;
} rng_t;

static void rng_open_event(rng_t *r)
{
    if (r->counter_mode) {
        r->state = spo_event_state(r->key, r->event);
        r->event++;
    }
}

/* GenerateUniformRealDist (Random.hpp:86-93): u(rand) = canonical * (max - min) + min.
 * One engine step per draw with the reference's output function -- the REFSEQ front end. */
static float rng_uniform(rng_t *r, float lo, float hi)
{
    uint32_t const d = spo_splitmix_next(&r->state);
    if (r->ctr)
        r->ctr->rng_draws++;
    return spo_u32_to_unit(d) * (hi - lo) + lo;
}

/* COUNTER front end: the same engine step (Random.hpp:37-39), but BOTH 32-bit halves of the mixed word
 * z ^ (z >> 31) are used as draws (the reference keeps bits 31..62 only).  Halves the integer work per
 * draw on the GPU; each half is a full-quality 32-bit output of splitmix64. */
void spo_splitmix_pair(uint64_t *state, uint32_t *hi, uint32_t *lo)
{
    uint64_t z = (*state += SPO_GOLDEN);
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ULL;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBULL;
    z = z ^ (z >> 31);
    *hi = (uint32_t)(z >> 32);
    *lo = (uint32_t)z;
}

static float map_uniform(uint32_t d, float lo, float hi) { return spo_u32_to_unit(d) * (hi - lo) + lo; }

/* one draw; counter mode: one engine step, high half (low half unused) */
static float rng_uniform1(rng_t *r, float lo, float hi)
{
    if (!r->counter_mode)
        return rng_uniform(r, lo, hi);
    uint32_t a, b;
    spo_splitmix_pair(&r->state, &a, &b);
    if (r->ctr)
        r->ctr->rng_draws++;
    return map_uniform(a, lo, hi);
}

/* two draws in call order; counter mode: one engine step (high half first) */
static void rng_uniform2(rng_t *r, float lo, float hi, float *u0, float *u1)
{
    if (!r->counter_mode) {
        *u0 = rng_uniform(r, lo, hi);
        *u1 = rng_uniform(r, lo, hi);
        return;
    }
    uint32_t a, b;
    spo_splitmix_pair(&r->state, &a, &b);
    if (r->ctr)
        r->ctr->rng_draws += 2;
    *u0 = map_uniform(a, lo, hi);
    *u1 = map_uniform(b, lo, hi);
}

/* GenerateUniformDistInsideSphereVector == GenerateNormalDistInsideSphereVector (Random.hpp:115-141):
 * redraw while Length < radius, i.e. KEEP points of the +-0.5 cube that lie OUTSIDE the 0.5 ball.
 * REFSEQ: three sequential draws per attempt, as the reference.
 * COUNTER: attempt k is ONE engine step (word k of the event's stream) whose 64 mixed bits are cut into three
 * fixed-point uniforms -- x = bits 63..43, y = bits 42..22 (21 bits each), z = bits 21..0 (22 bits) -- with
 * canonical = k * 2^-21 (2^-22) and the reference's map canonical * (max - min) + min; all of it exact in float,
 * so any attempt is addressable and nothing needs clamping. */
static v4 rng_shell(rng_t *r)
{
    float const radius = 0.5f;
    v4 res;
    do {
        float x, y, z;
        if (r->counter_mode) {
            uint32_t hi, lo;
            spo_splitmix_pair(&r->state, &hi, &lo);
            if (r->ctr)
                r->ctr->rng_draws += 3;
            uint32_t const kx = hi >> 11, ky = ((hi << 10) | (lo >> 22)) & 0x1FFFFFu, kz = lo & 0x3FFFFFu;
            x = (float)kx * 4.76837158203125e-7f * (radius + radius) + -radius;
            y = (float)ky * 4.76837158203125e-7f * (radius + radius) + -radius;
            z = (float)kz * 2.384185791015625e-7f * (radius + radius) + -radius;
        } else {
            rng_uniform2(r, -radius, radius, &x, &y);
            z = rng_uniform1(r, -radius, radius);
        }
        res = v3(x, y, z);
    } while (vlen(res) < radius);
    return res;
}

void spo_rng_uniform_seq(uint32_t seed32, float lo, float hi, uint32_t count, float *out)
{
    rng_t r = {spo_splitmix_seed(seed32), 0, 0, 0, NULL};
    for (uint32_t i = 0; i < count; ++i)
        out[i] = rng_uniform(&r, lo, hi);
}

void spo_rng_shell_seq(uint32_t seed32, uint32_t count, float *out3)
{
    rng_t r = {spo_splitmix_seed(seed32), 0, 0, 0, NULL};
    for (uint32_t i = 0; i < count; ++i) {
        v4 const s = rng_shell(&r);
        out3[3 * i + 0] = s.v[0];
        out3[3 * i + 1] = s.v[1];
        out3[3 * i + 2] = s.v[2];
    }
}

/* ------------------------------------------------------------- Collision.hpp */
/* RaySphereIntersection (Collision.hpp:9-17) */
static int ray_sphere_test(v4 center, float radius, v4 dir, v4 org)
{
    float const threshold = 1e-3f;
    v4 const L = vsub(center, org);
    float const tCenter = vdot(L, dir);
    float const distanceSquare = vdot(L, L) - tCenter * tCenter;
    return (tCenter > threshold) && (radius * radius - distanceSquare > threshold);
}

/* CalculateRaySphereMinIntersectionFactor (Collision.hpp:19-28) */
static float ray_sphere_tmin(v4 raySphere, float radius, v4 dir)
{
    float const tCenter = vdot(raySphere, dir);
    float const distanceSquare = vdot(raySphere, raySphere) - tCenter * tCenter;
    float const tDelta = sqrtf(radius * radius - distanceSquare);
    return tCenter - tDelta;
}

/* CalculateRaySphereMaxIntersectionFactor (Collision.hpp:30-37) */
static float ray_sphere_tmax(v4 raySphere, float radius, v4 dir)
{
    float const tCenter = vdot(raySphere, dir);
    float const distanceSquare = vdot(raySphere, raySphere) - tCenter * tCenter;
    float const tDelta = sqrtf(radius * radius - distanceSquare);
    return tCenter + tDelta;
}

/* CalculateRaySphereClosestContactPoint (Collision.hpp:49-56) */
static v4 closest_contact(v4 center, float radius, v4 org, v4 dir)
{
    return vadd(org, vscale(dir, ray_sphere_tmin(vsub(center, org), radius, dir)));
}

/* CalculateRaySphereFarthestContactPoint (Collision.hpp:58-65) */
static v4 farthest_contact(v4 center, float radius, v4 org, v4 dir)
{
    return vadd(org, vscale(dir, ray_sphere_tmax(vsub(center, org), radius, dir)));
}

/* CalculateRaySphereContactNormal (Collision.hpp:67-71) */
static v4 contact_normal(v4 point, v4 center) { return vnormalize(vsub(point, center)); }

/* FindClosestIntersectionSphere (Collision.hpp:87-109).  The index is carried as uint32_t: identical
 * to the reference's uint8_t for n <= 255, and the only defined behaviour beyond. */
static uint32_t closest_hit(const spo_scene *sc, v4 dir, v4 org, spo_counters *ctr)
{
    uint32_t minIndex = sc->n;
    float minDistanceSq = FLT_MAX;
    uint64_t passed = 0;
    for (uint32_t i = 0; i < sc->n; ++i) {
        v4 const c = vload4(sc->centers4 + 4 * i);
        float const r = sc->radii[i];
        if (ray_sphere_test(c, r, dir, org)) {
            ++passed;
            v4 const p = closest_contact(c, r, org, dir);
            if (vdot(org, dir) < vdot(p, dir)) {
                float const distanceSq = vlensq(vsub(org, p));
                minIndex = minDistanceSq > distanceSq ? i : minIndex;
                minDistanceSq = minDistanceSq > distanceSq ? distanceSq : minDistanceSq;
            }
        }
    }
    if (ctr) {
        ctr->segments++;
        ctr->tests += sc->n;
        ctr->passed += passed;
    }
    return minIndex;
}

uint32_t spo_closest_hit(const spo_scene *sc, const float d[4], const float o[4])
{
    return closest_hit(sc, vload4(d), vload4(o), NULL);
}

void spo_closest_hit_many(const spo_scene *sc, uint32_t count, const float *dir3, const float *org3, uint32_t *idx)
{
    for (uint32_t k = 0; k < count; ++k)
        idx[k] = closest_hit(sc, v3(dir3[3 * k], dir3[3 * k + 1], dir3[3 * k + 2]),
                             v3(org3[3 * k], org3[3 * k + 1], org3[3 * k + 2]), NULL);
}

/* ----------------------------------------------- camera rays (a2 in SURVEY 8a) */
/* SingleThreadPathTracer.hpp:125-127 == TaskBasedPathTracer.hpp:72-76.  NOTE the swapped divisors:
 * u (built from y) is divided by g_width, v (built from x) by g_height. */
static v4 primary_direction(const spo_camera *cam, uint32_t width, uint32_t height, uint32_t x, uint32_t y,
                            float xiU, float xiV)
{
    float const u = ((float)y + xiU) / (float)width;
    float const v = ((float)x + xiV) / (float)height;
    return vnormalize(mat_mul(cam->view, v3(-1.f + 2.f * v, -1.f + 2.f * u, 1.f)));
}

void spo_first_hit(const spo_scene *sc, const spo_camera *cam, uint32_t width, uint32_t height,
                   uint32_t xBegin, uint32_t xEnd, uint32_t yBegin, uint32_t yEnd,
                   uint32_t *idx, float *t, float *normal3, float *dir3)
{
    v4 const o = vload4(cam->eye);
    for (uint32_t y = yBegin; y < yEnd; ++y) {
        for (uint32_t x = xBegin; x < xEnd; ++x) {
            v4 const d = primary_direction(cam, width, height, x, y, 0.f, 0.f);
            uint32_t const i = closest_hit(sc, d, o, NULL);
            size_t const p = (size_t)(x - xBegin) + (size_t)(y - yBegin) * (xEnd - xBegin);
            idx[p] = i;
            if (dir3) {
                dir3[3 * p + 0] = d.v[0];
                dir3[3 * p + 1] = d.v[1];
                dir3[3 * p + 2] = d.v[2];
            }
            if (i < sc->n) {
                v4 const c = vload4(sc->centers4 + 4 * i);
                t[p] = ray_sphere_tmin(vsub(c, o), sc->radii[i], d);
                v4 const n = contact_normal(closest_contact(c, sc->radii[i], o, d), c);
                normal3[3 * p + 0] = n.v[0];
                normal3[3 * p + 1] = n.v[1];
                normal3[3 * p + 2] = n.v[2];
            } else {
                t[p] = 0.f;
                normal3[3 * p + 0] = normal3[3 * p + 1] = normal3[3 * p + 2] = 0.f;
            }
        }
    }
}

/* ------------------------------------------------------------ shading (a6-a9) */
typedef struct trace_ctx {
    const spo_scene *sc;
    const spo_camera *cam;
    uint32_t depth;
    uint32_t pass_cap;
    int pow_mode;
    spo_counters *ctr;
} trace_ctx;

/* SampleColorSkybox (SingleThreadPathTracer.hpp:11-19; TaskBasedPathTracer.hpp:184) */
static v4 sample_sky(const trace_ctx *c, v4 dir) { return vscale(vscale(vload4(c->cam->sky), dir.v[1] + 1.f), 0.5f); }

/* SampleColorDiffuse (SingleThreadPathTracer.hpp:21-37; TaskBasedPathTracer.hpp:83-106): terminal.
 * Result is the FIRST hit's colour times 0.5^k; later hits' colours/materials are ignored, there is
 * no sky term, and the 2nd+ directions add `origin` into the sum (quirks Q6, Q7). */
static v4 sample_diffuse(const trace_ctx *c, v4 dir, v4 org, uint32_t bounceCount, uint32_t idx, rng_t *rng)
{
    const spo_scene *sc = c->sc;
    v4 sampleColor = vscale(vload4(sc->colors4 + 4 * idx), 0.5f);
    v4 center = vload4(sc->centers4 + 4 * idx);
    org = closest_contact(center, sc->radii[idx], org, dir);
    rng_open_event(rng);
    dir = vnormalize(vadd(contact_normal(org, center), rng_shell(rng)));
    idx = closest_hit(sc, dir, org, c->ctr);
    while (--bounceCount && idx < sc->n) {
        sampleColor = vscale(sampleColor, 0.5f);
        center = vload4(sc->centers4 + 4 * idx);
        org = closest_contact(center, sc->radii[idx], org, dir);
        rng_open_event(rng);
        dir = vnormalize(vadd(vadd(org, contact_normal(org, center)), rng_shell(rng)));
        idx = closest_hit(sc, dir, org, c->ctr);
    }
    return sampleColor;
}

/* SampleColorReflective (SingleThreadPathTracer.hpp:39-46; TaskBasedPathTracer.hpp:108-122):
 * new ray only; no attenuation (quirk Q8); the shell vector is drawn even when fuzz is 0. */
static void scatter_reflective(const trace_ctx *c, v4 *dir, v4 *org, uint32_t idx, rng_t *rng)
{
    const spo_scene *sc = c->sc;
    v4 const center = vload4(sc->centers4 + 4 * idx);
    *org = closest_contact(center, sc->radii[idx], *org, *dir);
    v4 const normal = contact_normal(*org, center);
    rng_open_event(rng);
    *dir = vnormalize(vadd(vreflect(*dir, normal), vscale(rng_shell(rng), sc->fuzz[idx])));
}

/* SampleColorRefractive (SingleThreadPathTracer.hpp:48-92; TaskBasedPathTracer.hpp:124-180).
 * Enters, teleports to the far side of the SAME sphere, leaves (quirk Q9). */
static void scatter_refractive(const trace_ctx *c, v4 *pdir, v4 *porg, uint32_t idx, rng_t *rng)
{
    const spo_scene *sc = c->sc;
    float const nAir = 1.0f, nGlass = 1.5f;
    v4 dir = *pdir, org = *porg;
    v4 const center = vload4(sc->centers4 + 4 * idx);
    float const radius = sc->radii[idx];

    org = closest_contact(center, radius, org, dir);
    v4 n = contact_normal(org, center);
    rng_open_event(rng);

    float cc = vdot(vneg(n), dir);
    float r = nAir / nGlass;
    float const q1 = (nAir - nGlass) / (nAir + nGlass);
    float rSq = q1 * q1; /* pow(q, 2.f): exactly one rounded product (checked against powf in tests) */
    float schlick = rSq + (1.f - rSq) * spo_pow5(1.f - cc, c->pow_mode);

    if (rng_uniform1(rng, 0.f, 1.f) < schlick) {
        dir = vreflect(dir, n);
    } else if (r * sqrtf(1.f - cc * cc) < 1.f) {
        dir = vnormalize(vadd(vscale(dir, r), vscale(n, r * cc - sqrtf(1.f - r * r * (1.f - cc * cc)))));
        org = farthest_contact(center, radius, org, dir);
        n = vneg(contact_normal(org, center));

        cc = vdot(vneg(n), dir);
        r = nGlass / nAir;
        float const q2 = (nGlass - nAir) / (nGlass + nAir);
        rSq = q2 * q2;
        schlick = rSq + (1.f - rSq) * spo_pow5(1.f - cc, c->pow_mode);

        if (rng_uniform1(rng, 0.f, 1.f) < schlick) {
            dir = vreflect(dir, n);
        } else if (r * sqrtf(1.f - cc * cc) < 1.f) {
            dir = vnormalize(vadd(vscale(dir, r), vscale(n, r * cc - sqrtf(1.f - r * r * (1.f - cc * cc)))));
        } else {
            dir = vreflect(dir, n);
        }
    } else {
        dir = vreflect(dir, n);
    }
    *pdir = dir;
    *porg = org;
}

#define SPO_SPECULAR_SAFETY_CAP 65536u

/* TraceAndSampleColor (SingleThreadPathTracer.hpp:94-112), recursion unrolled into a loop (the
 * specular handlers tail-call it with bounceCount unchanged).  pass_cap > 0 adds the task tracer's
 * rule (TaskBasedPathTracer.hpp:81): a path still alive after pass_cap specular events is dropped
 * and its sample is not counted.  Returns 1 if the sample contributes. */
static int trace_path(const trace_ctx *c, v4 dir, v4 org, rng_t *rng, v4 *out)
{
    const spo_scene *sc = c->sc;
    uint32_t specular = 0;
    for (;;) {
        uint32_t const idx = closest_hit(sc, dir, org, c->ctr);
        if (idx >= sc->n) {
            *out = sample_sky(c, dir);
            return 1;
        }
        switch (sc->materials[idx]) {
        case SPO_DIFFUSE:
            *out = sample_diffuse(c, dir, org, c->depth, idx, rng);
            return 1;
        case SPO_REFLECTIVE:
            scatter_reflective(c, &dir, &org, idx, rng);
            break;
        case SPO_REFRACTIVE:
            scatter_refractive(c, &dir, &org, idx, rng);
            break;
        default: /* Material::SKYBOX on a sphere falls out of the switch to the sky return */
            *out = sample_sky(c, dir);
            return 1;
        }
        ++specular;
        if (c->pass_cap && specular >= c->pass_cap)
            return 0;
        if (specular >= SPO_SPECULAR_SAFETY_CAP) { /* the reference would overflow its stack */
            *out = v3(0.f, 0.f, 0.f);
            return 1;
        }
    }
}

int spo_sample_radiance(const spo_scene *sc, const spo_camera *cam, const spo_params *p,
                        uint32_t x, uint32_t y, uint32_t sample, float rgb[4], spo_counters *ctr)
{
    trace_ctx const c = {sc, cam, p->depth, p->pass_cap, p->pow_mode, ctr};
    rng_t rng = {0, spo_stream_key(p->seed, y * p->width + x, sample), 0, 1, ctr};
    rng_open_event(&rng);
    float xiU, xiV;
    rng_uniform2(&rng, -1.f, 1.f, &xiU, &xiV);
    v4 const d = primary_direction(cam, p->width, p->height, x, y, xiU, xiV);
    v4 out = v3(0.f, 0.f, 0.f);
    if (ctr)
        ctr->samples++;
    int const ok = trace_path(&c, d, vload4(cam->eye), &rng, &out);
    if (!ok && ctr)
        ctr->dropped++;
    memcpy(rgb, out.v, 16);
    return ok;
}

void spo_render_accum(const spo_scene *sc, const spo_camera *cam, const spo_params *p,
                      uint32_t yBegin, uint32_t yEnd, uint32_t xBegin, uint32_t xEnd, float *accum4,
                      spo_counters *ctr)
{
    for (uint32_t y = yBegin; y < yEnd; ++y) {
        for (uint32_t x = xBegin; x < xEnd; ++x) {
            /* pixelColor += sample, in sample order (SingleThreadPathTracer.hpp:121-131).  The running sum
             * CONTINUES from what the accumulator holds, so consecutive sample ranges rendered one after another
             * give the very bits of a single render of the whole range (progressive refinement, resume). */
            float *a = accum4 + 4 * ((size_t)y * p->width + x);
            v4 sum = v3(a[0], a[1], a[2]);
            float count = a[3];
            for (uint32_t s = p->sample_begin; s < p->sample_end; ++s) {
                float rgb[4];
                if (spo_sample_radiance(sc, cam, p, x, y, s, rgb, ctr)) {
                    sum = vadd(sum, vload4(rgb));
                    count += 1.f;
                }
            }
            a[0] = sum.v[0];
            a[1] = sum.v[1];
            a[2] = sum.v[2];
            a[3] = count;
        }
    }
}

/* ------------------------------------------------------------- IOHelpers.hpp */
static uint8_t to_u8(float f)
{
    if (!(f == f))
        return 0; /* NaN: UB in the reference; x86 cvttss2si and CUDA cvt.rzi both give low byte 0 */
    if (f >= 2147483648.f)
        return 0xFF;
    if (f <= -2147483648.f)
        return 0;
    return (uint8_t)(int32_t)f;
}

/* io::WritePixel (IOHelpers.hpp:17-22): gamma 2 then round-half-away */
void spo_write_pixel(const float rgb[3], uint8_t out[3])
{
    for (int k = 0; k < 3; ++k)
        out[k] = to_u8(roundf(sqrtf(rgb[k] / 255.f) * 255.f));
}

/* byte offset of pixel (x,y): g_size - ((g_width - x)*g_stride + y*g_width*g_stride)
 * (SingleThreadPathTracer.hpp:120, TaskBasedPathTracer.hpp:200-202) = pixel (x, H-1-y) */
static uint32_t pixel_index(uint32_t width, uint32_t height, uint32_t x, uint32_t y)
{
    uint32_t const size = width * height * 3u;
    return size - ((width - x) * 3u + y * width * 3u);
}

void spo_resolve(const float *accum4, uint32_t width, uint32_t height,
                 uint32_t yBegin, uint32_t yEnd, uint32_t xBegin, uint32_t xEnd, uint8_t *rgb8)
{
    for (uint32_t y = yBegin; y < yEnd; ++y) {
        for (uint32_t x = xBegin; x < xEnd; ++x) {
            const float *a = accum4 + 4 * ((size_t)y * width + x);
            /* colors[i] *= 1.f / samples[i] (TaskBasedPathTracer.hpp:198);
             * pixelColor *= 1.f / g_samples (SingleThreadPathTracer.hpp:133) */
            float const inv = 1.f / a[3];
            float const rgb[3] = {a[0] * inv, a[1] * inv, a[2] * inv};
            spo_write_pixel(rgb, rgb8 + pixel_index(width, height, x, y));
        }
    }
}

/* -------------------------------------------- reference-sequence replays (pinning) */
static void render_refseq_recursive(const trace_ctx *c, uint32_t width, uint32_t height, uint32_t spp,
                                    uint32_t yBegin, uint32_t yEnd, uint32_t xBegin, uint32_t xEnd, rng_t *rng,
                                    float *mean4, uint8_t *rgb8)
{
    /* RenderSegment (SingleThreadPathTracer.hpp:114-137) */
    for (uint32_t y = yBegin; y < yEnd; ++y) {
        for (uint32_t x = xBegin; x < xEnd; ++x) {
            uint32_t const index = pixel_index(width, height, x, y);
            v4 pixelColor = v3(0.f, 0.f, 0.f);
            for (uint32_t s = 0; s < spp; ++s) {
                float const xiU = rng_uniform(rng, -1.f, 1.f);
                float const xiV = rng_uniform(rng, -1.f, 1.f);
                v4 const d = primary_direction(c->cam, width, height, x, y, xiU, xiV);
                v4 out;
                if (c->ctr)
                    c->ctr->samples++;
                trace_path(c, d, vload4(c->cam->eye), rng, &out);
                pixelColor = vadd(pixelColor, out);
            }
            pixelColor = vscale(pixelColor, 1.f / (float)spp);
            if (mean4)
                memcpy(mean4 + 4 * (index / 3), pixelColor.v, 16);
            if (rgb8)
                spo_write_pixel(pixelColor.v, rgb8 + index);
        }
    }
}

typedef struct task_t { /* RaySampleTask (Definitions.hpp:23-31) */
    v4 origin, direction;
    uint32_t sphereIndex, bounceCount, x, y;
} task_t;

typedef struct queue_t { task_t *items; uint32_t size; } queue_t;
typedef struct tasks_t { queue_t diffuse, reflective, refractive, skybox; } tasks_t; /* Tasks (Definitions.hpp:33-39) */

static void q_push(queue_t *q, task_t t) { q->items[q->size++] = t; }

/* TraceAndSampleColor(task, tasks) (TaskBasedPathTracer.hpp:9-30) */
static void enqueue_task(const trace_ctx *c, task_t task, tasks_t *tasks, int count_segment)
{
    task.sphereIndex = closest_hit(c->sc, task.direction, task.origin, count_segment ? c->ctr : NULL);
    if (task.sphereIndex < c->sc->n) {
        switch (c->sc->materials[task.sphereIndex]) {
        case SPO_DIFFUSE: q_push(&tasks->diffuse, task); return;
        case SPO_REFLECTIVE: q_push(&tasks->reflective, task); return;
        case SPO_REFRACTIVE: q_push(&tasks->refractive, task); return;
        default: break;
        }
    }
    q_push(&tasks->skybox, task);
}

static void tasks_alloc(tasks_t *t, uint32_t cap)
{
    t->diffuse.items = (task_t *)malloc(sizeof(task_t) * cap);
    t->reflective.items = (task_t *)malloc(sizeof(task_t) * cap);
    t->refractive.items = (task_t *)malloc(sizeof(task_t) * cap);
    t->skybox.items = (task_t *)malloc(sizeof(task_t) * cap);
    t->diffuse.size = t->reflective.size = t->refractive.size = t->skybox.size = 0;
}
static void tasks_free(tasks_t *t) { free(t->diffuse.items); free(t->reflective.items); free(t->refractive.items); free(t->skybox.items); }
static void tasks_clear(tasks_t *t) { t->diffuse.size = t->reflective.size = t->refractive.size = t->skybox.size = 0; }
static uint32_t tasks_size(const tasks_t *t) { return t->diffuse.size + t->reflective.size + t->refractive.size + t->skybox.size; }

static void render_refseq_task(const trace_ctx *c, uint32_t width, uint32_t height, uint32_t spp,
                               uint32_t yBegin, uint32_t yEnd, uint32_t xBegin, uint32_t xEnd, rng_t *rng,
                               float *mean4, uint8_t *rgb8)
{
    /* RenderSegmentTask (TaskBasedPathTracer.hpp:54-206) */
    uint32_t const segmentWidth = xEnd - xBegin, segmentHeight = yEnd - yBegin;
    uint32_t const npix = segmentWidth * segmentHeight;
    v4 *colors = (v4 *)calloc(npix, sizeof(v4));
    float *samples = (float *)calloc(npix, sizeof(float));
    tasks_t tasks, next;
    tasks_alloc(&tasks, npix);
    tasks_alloc(&next, npix);

    for (uint32_t s = 0; s < spp; ++s) {
        tasks_clear(&tasks);
        for (uint32_t y = yBegin; y < yEnd; ++y) {
            for (uint32_t x = xBegin; x < xEnd; ++x) {
                float const xiU = rng_uniform(rng, -1.f, 1.f);
                float const xiV = rng_uniform(rng, -1.f, 1.f);
                task_t const t = {vload4(c->cam->eye), primary_direction(c->cam, width, height, x, y, xiU, xiV),
                                  c->sc->n, c->depth, x, y};
                if (c->ctr)
                    c->ctr->samples++;
                enqueue_task(c, t, &tasks, 1);
            }
        }
        for (uint32_t pass = 0; pass < 10 && tasks_size(&tasks) > 0; ++pass) { /* literal 10, :81 */
            for (uint32_t i = 0; i < tasks.diffuse.size; ++i) {
                task_t *t = &tasks.diffuse.items[i];
                v4 const sampleColor = sample_diffuse(c, t->direction, t->origin, t->bounceCount, t->sphereIndex, rng);
                /* stride is segmentHeight, as written (:103) -- only right for square segments (quirk Q15) */
                uint32_t const colorIndex = (t->x - xBegin) + (t->y - yBegin) * segmentHeight;
                if (colorIndex < npix) { /* the reference would write out of bounds */
                    colors[colorIndex] = vadd(colors[colorIndex], sampleColor);
                    samples[colorIndex] += 1.f;
                }
            }
            for (uint32_t i = 0; i < tasks.reflective.size; ++i) {
                task_t *t = &tasks.reflective.items[i];
                scatter_reflective(c, &t->direction, &t->origin, t->sphereIndex, rng);
                /* the handler's own FindClosest (:117) is repeated by the enqueue helper (:11);
                 * only the second is counted as an algorithmic segment */
                task_t const nt = {t->origin, t->direction, 0, c->depth, t->x, t->y};
                enqueue_task(c, nt, &next, 1);
            }
            for (uint32_t i = 0; i < tasks.refractive.size; ++i) {
                task_t *t = &tasks.refractive.items[i];
                scatter_refractive(c, &t->direction, &t->origin, t->sphereIndex, rng);
                task_t const nt = {t->origin, t->direction, 0, c->depth, t->x, t->y};
                enqueue_task(c, nt, &next, 1);
            }
            for (uint32_t i = 0; i < tasks.skybox.size; ++i) {
                task_t *t = &tasks.skybox.items[i];
                v4 const sampleColor = sample_sky(c, t->direction);
                uint32_t const colorIndex = (t->x - xBegin) + (t->y - yBegin) * segmentHeight;
                if (colorIndex < npix) {
                    colors[colorIndex] = vadd(colors[colorIndex], sampleColor);
                    samples[colorIndex] += 1.f;
                }
            }
            tasks_t const tmp = tasks; /* SwapTasks + ClearTasks(nextFrameTasks) (:191-192) */
            tasks = next;
            next = tmp;
            tasks_clear(&next);
        }
        if (c->ctr)
            c->ctr->dropped += tasks_size(&tasks);
    }

    for (uint32_t i = 0; i < npix; ++i) {
        colors[i] = vscale(colors[i], 1.f / samples[i]);
        uint32_t const x = i % segmentWidth + xBegin;
        uint32_t const y = i / segmentWidth + yBegin;
        uint32_t const index = pixel_index(width, height, x, y);
        if (mean4)
            memcpy(mean4 + 4 * (index / 3), colors[i].v, 16);
        if (rgb8)
            spo_write_pixel(colors[i].v, rgb8 + index);
    }
    tasks_free(&tasks);
    tasks_free(&next);
    free(colors);
    free(samples);
}

void spo_render_refseq(const spo_scene *sc, const spo_camera *cam, uint32_t width, uint32_t height,
                       uint32_t spp, uint32_t depth, int tracer, int pow_mode,
                       uint32_t yBegin, uint32_t yEnd, uint32_t xBegin, uint32_t xEnd, uint32_t seed32,
                       float *mean4, uint8_t *rgb8, spo_counters *ctr)
{
    trace_ctx const c = {sc, cam, depth, 0 /* the task replay drops through its literal pass loop */, pow_mode, ctr};
    rng_t rng = {spo_splitmix_seed(seed32), 0, 0, 0, ctr};
    if (tracer == SPO_TRACER_RECURSIVE)
        render_refseq_recursive(&c, width, height, spp, yBegin, yEnd, xBegin, xEnd, &rng, mean4, rgb8);
    else
        render_refseq_task(&c, width, height, spp, yBegin, yEnd, xBegin, xEnd, &rng, mean4, rgb8);
}

/* --------------------------------------------------------- SceneGenerators.hpp */
typedef struct scene_out {
    uint32_t cap, n;
    float *centers4, *radii, *colors4, *fuzz;
    uint8_t *materials;
} scene_out;

static void scene_put(scene_out *o, uint32_t i, v4 center, float radius, v4 color, uint8_t material)
{
    if (i < o->cap) {
        memcpy(o->centers4 + 4 * i, center.v, 16);
        memcpy(o->colors4 + 4 * i, color.v, 16);
        o->radii[i] = radius;
        o->materials[i] = material;
    }
}

/* GenerateUnitVector (Random.hpp:95-113): three U[-1,1) draws (x, y, z), normalised; only feeds the
 * unused g_attenuations, but it advances the stream the fuzz values come from. */
static void rng_unit_vector_discard(rng_t *r)
{
    (void)rng_uniform(r, -1.f, 1.f);
    (void)rng_uniform(r, -1.f, 1.f);
    (void)rng_uniform(r, -1.f, 1.f);
}

/* InitSpheres (SceneGenerators.hpp:68-133).  g_sphereNumber keeps its initialiser 10 (Globals.hpp:37). */
uint32_t spo_init_spheres(uint32_t seed32, uint32_t cap, float *centers4, float *radii, uint8_t *materials,
                          float *colors4, float *fuzz)
{
    static const float colors[10][3] = {{30, 144, 255}, {10, 255, 110}, {110, 10, 255}, {255, 100, 230}, {200, 255, 110},
                                        {210, 10, 255}, {255, 100, 150}, {50, 255, 200}, {10, 210, 255}, {255, 100, 220}};
    static const float centers[10][3] = {{0, -1e3f - 0.5f, 0}, {-1, 0, 0}, {0, 0, 0}, {1, 0, 0}, {-1, 1, 0},
                                         {0, 1, 0}, {1, 1, 0}, {-1, 2, 0}, {0, 2, 0}, {1, 2, 0}};
    static const uint8_t mats[10] = {SPO_DIFFUSE, SPO_DIFFUSE, SPO_REFLECTIVE, SPO_DIFFUSE, SPO_DIFFUSE,
                                     SPO_REFRACTIVE, SPO_DIFFUSE, SPO_DIFFUSE, SPO_REFLECTIVE, SPO_DIFFUSE};
    uint32_t const n = 10;
    if (cap < n)
        return n;
    scene_out o = {cap, n, centers4, radii, colors4, fuzz, materials};
    for (uint32_t i = 0; i < n; ++i)
        scene_put(&o, i, v3(centers[i][0], centers[i][1], centers[i][2]), i == 0 ? 1e3f : 0.5f,
                  v3(colors[i][0], colors[i][1], colors[i][2]), mats[i]);
    rng_t r = {spo_splitmix_seed(seed32), 0, 0, 0, NULL};
    for (uint32_t i = 0; i < n; ++i) /* g_attenuations (:122-125) */
        if (rng_uniform(&r, 0.f, 1.f) > 0.3f)
            rng_unit_vector_discard(&r);
    for (uint32_t i = 0; i < n; ++i) { /* g_diffuses (:127-130) */
        fuzz[i] = 0.01f;
        if (rng_uniform(&r, 0.f, 1.f) > 0.3f)
            fuzz[i] = rng_uniform(&r, 0.f, 1.f);
    }
    fuzz[2] = 0.f; /* :132 */
    return n;
}

/* GenerateSpheres (SceneGenerators.hpp:6-66) */
uint32_t spo_generate_spheres(uint32_t seed32, uint32_t cap, float *centers4, float *radii, uint8_t *materials,
                              float *colors4, float *fuzz)
{
    /* worst case: 16 rows of at most ceil((11 + 2*0.85*18.75)/1.25)+1 columns, plus 4 */
    enum { MAXN = 4 + 16 * 40 };
    v4 *C = (v4 *)malloc(sizeof(v4) * MAXN), *K = (v4 *)malloc(sizeof(v4) * MAXN);
    float *R = (float *)malloc(sizeof(float) * MAXN);
    uint8_t *M = (uint8_t *)malloc(MAXN);
    uint32_t n = 0;
    rng_t r = {spo_splitmix_seed(seed32), 0, 0, 0, NULL};

    K[n] = v3(30, 144, 255); C[n] = v3(0, -1e6f, 0); R[n] = 1e6f; M[n] = SPO_DIFFUSE; ++n;
    K[n] = v3(0, 0, 0); C[n] = v3(0, 3, 10); R[n] = 3; M[n] = SPO_REFRACTIVE; ++n;
    K[n] = v3(0, 0, 0); C[n] = v3(5, 3, 5); R[n] = 3; M[n] = SPO_REFLECTIVE; ++n;
    K[n] = v3(223, 55, 132); C[n] = v3(-7, 3, 14); R[n] = 3; M[n] = SPO_DIFFUSE; ++n;

    float const minR = 0.3f, maxR = 0.5f;
    for (float z = 0; z < 20; z += 1.25f) {
        float const bound = fabsf(z) * 0.85f;
        for (float x = -5 - bound; x < 6 + bound; x += 1.25f) {
            if (rng_uniform(&r, 0.f, 1.f) > 0.5f) {
                float const rad = rng_uniform(&r, minR, maxR);
                float const cx = x + rng_uniform(&r, 0.f, minR);
                float const cz = z + rng_uniform(&r, 0.f, minR);
                v4 const c = v3(cx, rad, cz);
                if ((vlen(vsub(c, C[1])) - rad - R[1] < 0.5f) || (vlen(vsub(c, C[2])) - rad - R[2] < 0.5f) ||
                    (vlen(vsub(c, C[3])) - rad - R[3] < 0.5f))
                    continue;
                float const k0 = rng_uniform(&r, 0.f, 255.f);
                float const k1 = rng_uniform(&r, 0.f, 255.f);
                float const k2 = rng_uniform(&r, 0.f, 255.f);
                float const m = fminf(roundf(rng_uniform(&r, 0.5f, 6.0f)), 3.0f);
                if (n < MAXN) {
                    C[n] = c; R[n] = rad; K[n] = v3(k0, k1, k2); M[n] = (uint8_t)m; ++n;
                }
            }
        }
    }
    if (cap >= n) {
        scene_out o = {cap, n, centers4, radii, colors4, fuzz, materials};
        for (uint32_t i = 0; i < n; ++i)
            scene_put(&o, i, C[i], R[i], K[i], M[i]);
        for (uint32_t i = 0; i < n; ++i) /* g_attenuations (:56-59) */
            if (rng_uniform(&r, 0.f, 1.f) > 0.2f)
                rng_unit_vector_discard(&r);
        for (uint32_t i = 0; i < n; ++i) { /* g_diffuses (:61-64) */
            fuzz[i] = 0.0f;
            if (rng_uniform(&r, 0.f, 1.f) > 0.2f)
                fuzz[i] = rng_uniform(&r, 0.f, 1.f);
        }
        fuzz[2] = 0.01f; /* :65 */
    }
    free(C); free(K); free(R); free(M);
    return n;
}
