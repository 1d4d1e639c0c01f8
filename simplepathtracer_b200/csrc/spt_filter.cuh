// spt_filter.cuh -- the closest-hit loop as the B200 runs it: a conservative FMA pre-filter over the
// whole sphere list in packed f32x2 arithmetic, then the reference's exact sequence on the few survivors.
//
// Why: cd::RaySphereIntersection (Collision.hpp:9-17) is 17 separately rounded flops per (ray, sphere).
// Spelled that way it costs ~23 issue slots per test and, having no FMA, can never pass 50 % of the FP32
// roofline.  But only ~2 of N spheres pass per ray, and only those decide the result.  So every sphere first
// goes through a cheaper test that is PROVABLY implied by the exact one (no false negatives), and only
// spheres that pass it are evaluated with the exact, separately rounded sequence (exact_sphere<>), which
// alone decides index, t and normal.  Results are therefore bit-identical to the strict loop.
//
// The pre-filter, per (ray, sphere), with C = centre, O = origin, D = direction:
//     tC = C.D - O.D (+ epsT)                           3 FMA   chain seeded with a per-ray constant
//     na = -(|C|^2 - r^2) + 2 C.O  (+ S_i - 1e-3)       3 FMA   chain seeded with a per-sphere constant
//     v  = tC*tC + na  - |O|^2 (+ S_ray)                1 FMA + 1 ADD
//     candidate  <=>  min(v, tC) >= 0                    i.e.  r^2 - d^2 > 1e-3 - slack  and  tC > -epsT
// v is the discriminant r^2 - (|C-O|^2 - tC^2) of Collision.hpp:14-16 expanded about the world origin so that
// nothing of the form (C - O) has to be formed per test.  Two spheres are tested per instruction
// (fma.rn.f32x2 / add.rn.f32x2: one issue slot, two lanes of the FP32 pipe), the sign of min(v, tC) is shifted
// into a per-lane bit mask (one FMNMX + one SHF per test), and nothing branches inside the loop.
//
// Slack (u = 2^-24; A = |C_i|, B = |O|, |D| <= 1 + 2^-20).  Worst-case rounding analysis of both sides
// ("pre-filter bound"):  |disc_exact_sequence - disc_real| <= u(33.4(A^2+B^2) + 2r^2),
// |v - (disc_real - 1e-3 + S_i + S_ray)| <= u(37.8A^2 + 41.9B^2 + 6r^2), so S_i = 96u(A^2 + r^2) and
// S_ray = 96u B^2 cover the sum with margin.  For tC the exact sequence errs by <= 4.2u(A+B) and the chain
// by <= 6.2u(A+B): with A + B <= 1024 a strict pass (tC_exact > 1e-3) implies tC_chain > 0, so epsT = 0;
// beyond that epsT = 2^-20 (Amax + B) is seeded into the chain and S_ray grows by 2^-18 (Amax + B)^2.
// Spheres too large for this to be useful (r > kBigRadius or |C| > kBigCentre: the ground sphere) are kept
// out of the filter and always evaluated exactly.
#pragma once

#include "spt_device.cuh"

namespace spt
{

#ifndef SPT_FILTER_TC
#define SPT_FILTER_TC 0 // 1: the pre-filter also tests the sign of tC (one more FMNMX per sphere, fewer candidates)
#endif
#ifndef SPT_WORD_UNROLL2
#define SPT_WORD_UNROLL2 1 // the single-ray word loop takes two full words per trip while it can (+0.6 .. 1.0 % on config 3; 0: one)
#endif
constexpr float kBigRadius = 64.f;
constexpr float kBigCentre = 4096.f;
constexpr uint32_t kWordSpheres = 32;                // one mask word
constexpr uint32_t kQuadSpheres = 8;                 // filter slots are padded to a multiple of this (4 records)
constexpr uint32_t kSuperWords = 4;                  // words per candidate-resolution round
constexpr double kSlackU = 96.0 / 16777216.0;        // 96 u

// One record = TWO spheres, component-interleaved so that each 64-bit half is an f32x2 operand:
//   a = (cx0, cx1, cy0, cy1)   b = (cz0, cz1, nk0, nk1),   nk = -(|C|^2 - r^2) + S_i - 1e-3 (rounded up)
struct __align__(16) FilterPair
{
    float cx[2], cy[2], cz[2], nk[2];
};

// Loop shape of a span of slots (a multiple of kQuadSpheres): mask words, words with all 16 records, quads (4 records) in the
// last word.  For the resident list it is computed on the host and read from the kernel's parameter bank: ncu on the bounce kernel
// showed ptxas recomputing it from nSlots in vector registers on every trip of the word loop when registers are short (12 control
// instructions per trip instead of 3).
struct SpanShape
{
    uint32_t nWords, nFull, tailQuads;
};
__host__ __device__ __forceinline__ SpanShape span_shape(uint32_t spanSlots)
{
    SpanShape s;
    s.nWords = (spanSlots + 31u) / 32u;
    s.tailQuads = s.nWords ? (spanSlots - (s.nWords - 1u) * 32u) / 8u : 0u; // 1..4 quads in the last word
    s.nFull = s.tailQuads == 4u ? s.nWords : (s.nWords ? s.nWords - 1u : 0u); // words with all 16 records
    return s;
}

// The filter works on SLOTS: the filterable spheres in index order, compacted (the big spheres take no slot)
// and padded with never-pass records to a multiple of kQuadSpheres.  slotIdx maps a slot back to its sphere
// index (0xFFFFFFFF for padding).  `pairs` and `slotIdx` are ONE device allocation ([records | indices]) so that a
// single TMA bulk copy stages both.
struct FilterScene
{
    const FilterPair *pairs;  // nSlots / 2 records
    const uint32_t *slotIdx;  // nSlots entries
    uint32_t nSlots;          // multiple of kQuadSpheres
    SpanShape shape;          // span_shape(nSlots), filled on the host
    const uint32_t *bigIdx;   // spheres always evaluated exactly
    uint32_t nBig;
    float amax;               // max |C_i| over filtered spheres
    float nearOO;             // origins with |O|^2 <= nearOO have amax + |O| <= 1024: the tC test needs no widening (host, rounded down)
    float epsTCam;            // camera form only: the epsT the launch's records were built with (0 unless the eye is far away)
};

__host__ __device__ __forceinline__ uint32_t filter_blob_bytes(uint32_t nSlots)
{
    return nSlots / 2u * uint32_t(sizeof(FilterPair)) + nSlots * uint32_t(sizeof(uint32_t));
}
// point the filter at a staged copy of the blob
__device__ __forceinline__ void rebase_filter(FilterScene &fs, const void *blob)
{
    fs.pairs = reinterpret_cast<const FilterPair *>(blob);
    fs.slotIdx = reinterpret_cast<const uint32_t *>(fs.pairs + fs.nSlots / 2u);
}

// ---- packed f32x2 helpers (sm_100a: FFMA2 / FADD2) -------------------------------------------------
__device__ __forceinline__ unsigned long long pack2(float lo, float hi)
{
    unsigned long long r;
    asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi));
    return r;
}
__device__ __forceinline__ void unpack2(unsigned long long v, float &lo, float &hi)
{
    asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v));
}
__device__ __forceinline__ unsigned long long fma2(unsigned long long a, unsigned long long b, unsigned long long c)
{
    unsigned long long d;
    asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c));
    return d;
}
__device__ __forceinline__ unsigned long long add2(unsigned long long a, unsigned long long b)
{
    unsigned long long d;
    asm("add.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
    return d;
}

// Per-ray constants of the pre-filter, each rounded in the direction that can only ADD candidates.
struct FilterRay
{
    unsigned long long dx, dy, dz;    // direction, broadcast to both halves
    unsigned long long ox2, oy2, oz2; // 2 * origin
    unsigned long long nod;           // -(O.D) + epsT
    unsigned long long noo;           // -|O|^2 + S_ray
};

__device__ __forceinline__ FilterRay make_filter_ray(V3 dir, V3 org, float od, float amax, float nearOO)
{
    float const oo = __fmaf_ru(org.z, org.z, __fmaf_ru(org.y, org.y, __fmul_ru(org.x, org.x))); // >= |O|^2
    float sRay = __fmul_ru(oo, float(kSlackU) * 1.0001f);
    float epsT = 0.f;
    if (!(oo <= nearOO)) // amax + |O| may exceed 1024 (decided without a square root: nearOO = (1024 - amax)^2, rounded down)
    { // far-away origin: widen the tC test, and pay for shifting tC in the square
        float const m = __fadd_ru(amax, __fsqrt_ru(oo));
        epsT = __fmul_ru(m, 9.5367431640625e-7f);                   // 2^-20 (Amax + B)
        sRay = __fadd_ru(sRay, __fmul_ru(__fmul_ru(m, m), 3.8146973e-6f)); // + 2^-18 (Amax + B)^2
    }
    float const oolo = __fmaf_rd(org.z, org.z, __fmaf_rd(org.y, org.y, __fmul_rd(org.x, org.x))); // <= |O|^2
    float const nod = __fadd_ru(-od, epsT); // od is the exact sequence's Dot(O,D): its <= 3.1u B error is in the bound
    float const noo = __fadd_ru(-oolo, sRay);
    FilterRay r;
    r.dx = pack2(dir.x, dir.x);
    r.dy = pack2(dir.y, dir.y);
    r.dz = pack2(dir.z, dir.z);
    r.ox2 = pack2(org.x + org.x, org.x + org.x);
    r.oy2 = pack2(org.y + org.y, org.y + org.y);
    r.oz2 = pack2(org.z + org.z, org.z + org.z);
    r.nod = pack2(nod, nod);
    r.noo = pack2(noo, noo);
    return r;
}

// one record (two spheres): shifts two FAIL bits into `mask` (sign of v)
__device__ __forceinline__ void filter_pair_regs(ulonglong2 a, ulonglong2 b, const FilterRay &r, uint32_t &mask)
{
    // a = (cx0,cx1) (cy0,cy1), b = (cz0,cz1) (nk0,nk1)
    unsigned long long const tc = fma2(b.x, r.dz, fma2(a.y, r.dy, fma2(a.x, r.dx, r.nod)));
    unsigned long long const na = fma2(b.x, r.oz2, fma2(a.y, r.oy2, fma2(a.x, r.ox2, b.y)));
    unsigned long long const v = add2(fma2(tc, tc, na), r.noo);
    float t0, t1, v0, v1;
    unpack2(tc, t0, t1);
    unpack2(v, v0, v1);
#if SPT_FILTER_TC
    mask = __funnelshift_l(__float_as_uint(fminf(v0, t0)), mask, 1);
    mask = __funnelshift_l(__float_as_uint(fminf(v1, t1)), mask, 1);
#else
    // discriminant only: the tC > 1e-3 half of the test is left to the exact sequence's early-out
    mask = __funnelshift_l(__float_as_uint(v0), mask, 1);
    mask = __funnelshift_l(__float_as_uint(v1), mask, 1);
#endif
}
__device__ __forceinline__ void filter_pair(const FilterPair *__restrict__ rec, const FilterRay &r, uint32_t &mask)
{
    filter_pair_regs(*reinterpret_cast<const ulonglong2 *>(rec->cx), *reinterpret_cast<const ulonglong2 *>(rec->cz), r, mask);
}

// Camera rays all start at the eye E, so everything of v that does not involve the direction is a per-sphere constant of
// the launch: the record's 4th field then holds  nkc = -(|C - E|^2 - r^2) + S_i + S_ray(E) - 1e-3  (binary64 on the host,
// rounded up) and the test is  v = tC^2 + nkc  with the same tC chain: 4 FFMA2 per record instead of 7 FFMA2 + 1 FADD2.
// Fewer rounding steps than the general chain under the same slack, so it is conservative a fortiori.
__device__ __forceinline__ void filter_pair_cam(const FilterPair *__restrict__ rec, const FilterRay &r, uint32_t &mask)
{
    ulonglong2 const a = *reinterpret_cast<const ulonglong2 *>(rec->cx); // (cx0,cx1) (cy0,cy1)
    ulonglong2 const b = *reinterpret_cast<const ulonglong2 *>(rec->cz); // (cz0,cz1) (nkc0,nkc1)
    unsigned long long const tc = fma2(b.x, r.dz, fma2(a.y, r.dy, fma2(a.x, r.dx, r.nod)));
    unsigned long long const v = fma2(tc, tc, b.y);
    float v0, v1;
    unpack2(v, v0, v1);
    mask = __funnelshift_l(__float_as_uint(v0), mask, 1);
    mask = __funnelshift_l(__float_as_uint(v1), mask, 1);
}
// per-ray constants of the camera form: direction and -(E.D) + epsT (epsT is a constant of the launch)
__device__ __forceinline__ FilterRay make_filter_ray_cam(V3 dir, float od, float epsT)
{
    float const nod = __fadd_ru(-od, epsT);
    FilterRay r;
    r.dx = pack2(dir.x, dir.x);
    r.dy = pack2(dir.y, dir.y);
    r.dz = pack2(dir.z, dir.z);
    r.nod = pack2(nod, nod);
    r.ox2 = r.oy2 = r.oz2 = r.noo = 0ull; // unused
    return r;
}

// The pre-filter over a SPAN of slots (a multiple of kQuadSpheres; the whole list, or one staged chunk of it),
// followed by the exact sequence on the span's candidates.  minIndex / minDistanceSq carry the running closest
// hit across spans.  `live` = false marks a lane that only keeps its warp company (finished pixel): it runs the
// convergent filter loop with the rest of the warp but never enters the exact path.
//
// Loop shape: ONE rolled loop over mask words whose body is 16 unrolled records (32 slots), so the hot code is
// ~230 instructions that stay in the instruction cache however many warps are in it; the last word runs
// only the quads (4 records) that exist.  The masks of up to kSuperWords words are parked in registers -- a shift
// register with the newest word in c0, so no select on the word index is needed -- and the divergent candidate
// loop (resolve_candidates) runs once per super-word, so lanes with candidates in different words of the same
// super-word are served in the same trips.
//
// resolve_candidates: the exact sequence on the candidates parked in c0..c4: c_k holds the mask of word (wEnd - 1 - k) of the span.  Leaves all
// five masks zero.  Lanes without candidates idle here (the only divergent part of the closest hit).
template <typename GeomPtr>
__device__ __forceinline__ void resolve_candidates(uint32_t &c0, uint32_t &c1, uint32_t &c2, uint32_t &c3, uint32_t &c4, uint32_t wEnd,
                                                   const uint32_t *slotIdx, GeomPtr geom, uint32_t n, V3 dir, V3 org, float od,
                                                   uint32_t &minIndex, float &minDistanceSq, uint32_t &exactEvals)
{
    for (;;)
    {
        uint32_t k, bits;
        if (c0) { k = 0; bits = c0; }
        else if (c1) { k = 1; bits = c1; }
        else if (c2) { k = 2; bits = c2; }
        else if (c3) { k = 3; bits = c3; }
        else if (c4) { k = 4; bits = c4; }
        else break;
        uint32_t const j = __clz(bits);
        uint32_t const cleared = bits & ~(0x80000000u >> j);
        if (k == 0) c0 = cleared;
        else if (k == 1) c1 = cleared;
        else if (k == 2) c2 = cleared;
        else if (k == 3) c3 = cleared;
        else c4 = cleared;
        uint32_t const i = slotIdx[(wEnd - 1u - k) * kWordSpheres + j];
        if (i < n) // padding records can only show up for NaN rays
        {
            exact_sphere<false>(geom[i], i, dir, org, od, minIndex, minDistanceSq);
            ++exactEvals;
        }
    }
}

// When the parked masks are resolved: after kSuperWords words -- unless exactly ONE word is left in the span, which then
// joins the round (a 149-sphere list is 4 full words + 24 slots: one divergent round per closest hit instead of two; the
// trips of a round are the maximum over the lanes, so one round over five words makes fewer than two rounds over 4 + 1).
#ifndef SPT_RESOLVE_MERGE_TAIL
#define SPT_RESOLVE_MERGE_TAIL 1
#endif
__device__ __forceinline__ bool resolve_now(uint32_t pend, uint32_t w, uint32_t nWords)
{
#if SPT_RESOLVE_MERGE_TAIL
    return w == nWords || (pend >= kSuperWords && w + 1u != nWords);
#else
    return w == nWords || pend >= kSuperWords;
#endif
}

template <bool CAM = false, typename GeomPtr>
__device__ __forceinline__ void filter_span(const FilterPair *rec, const uint32_t *slotIdx, SpanShape const shape, const FilterRay &fr,
                                            GeomPtr geom, uint32_t n, V3 dir, V3 org, float od, uint32_t &minIndex,
                                            float &minDistanceSq, uint32_t &exactEvals, bool live)
{
    uint32_t const nWords = shape.nWords, tailQuads = shape.tailQuads, nFull = shape.nFull;
    // candidate masks of the last (up to kSuperWords + 1) words filtered, newest in c0 (a shift register: no per-word selects)
    uint32_t c0 = 0u, c1 = 0u, c2 = 0u, c3 = 0u, c4 = 0u, pend = 0u;
#pragma unroll 1
    for (uint32_t w = 0; w < nWords;) // everything about w is warp-uniform
    {
        uint32_t mask = 0xFFFFFFFFu; // all "fail"
#if SPT_WORD_UNROLL2
        if (w + 2u <= nFull) // two full words per trip: the loop control is paid once per 64 slots
        {
            uint32_t mask0 = 0xFFFFFFFFu;
#pragma unroll
            for (uint32_t pp = 0; pp < kWordSpheres / 2; ++pp)
            {
                if (CAM)
                    filter_pair_cam(rec + pp, fr, mask0);
                else
                    filter_pair(rec + pp, fr, mask0);
            }
#pragma unroll
            for (uint32_t pp = kWordSpheres / 2; pp < kWordSpheres; ++pp)
            {
                if (CAM)
                    filter_pair_cam(rec + pp, fr, mask);
                else
                    filter_pair(rec + pp, fr, mask);
            }
            rec += kWordSpheres;
            c4 = c2;
            c3 = c1;
            c2 = c0;
            c1 = live ? ~mask0 : 0u;
            ++w;
            ++pend;
        }
        else
#endif
        {
            if (w < nFull)
            {
#pragma unroll
                for (uint32_t pp = 0; pp < kWordSpheres / 2; ++pp)
                {
                    if (CAM)
                        filter_pair_cam(rec + pp, fr, mask);
                    else
                        filter_pair(rec + pp, fr, mask);
                }
                rec += kWordSpheres / 2;
            }
            else // the partial last word: only the quads (4 records) that exist
            {
#pragma unroll 1
                for (uint32_t q = 0; q < tailQuads; ++q)
                {
#pragma unroll
                    for (uint32_t pp = 0; pp < kQuadSpheres / 2; ++pp)
                    {
                        if (CAM)
                            filter_pair_cam(rec + pp, fr, mask);
                        else
                            filter_pair(rec + pp, fr, mask);
                    }
                    rec += kQuadSpheres / 2;
                }
                uint32_t const sh = kWordSpheres - kQuadSpheres * tailQuads; // 8, 16 or 24 missing slots: fail bits
                mask = (mask << sh) | ((1u << sh) - 1u);
            }
            c4 = c3;
            c3 = c2;
            c2 = c1;
            c1 = c0;
        }
        c0 = live ? ~mask : 0u; // bit (31 - j) set <=> slot j of the word is a candidate
        ++w;
        ++pend;
        if (resolve_now(pend, w, nWords))
        {
            resolve_candidates(c0, c1, c2, c3, c4, w, slotIdx, geom, n, dir, org, od, minIndex, minDistanceSq, exactEvals);
            pend = 0u;
        }
    }
}

// the spheres the filter cannot judge (the ground): always through the exact sequence
template <typename GeomPtr>
__device__ __forceinline__ void exact_big_spheres(const FilterScene &fs, GeomPtr geom, V3 dir, V3 org, float od, uint32_t &minIndex,
                                                  float &minDistanceSq, uint32_t &exactEvals, bool live)
{
    if (live)
    {
        for (uint32_t b = 0; b < fs.nBig; ++b)
        {
            uint32_t const i = fs.bigIdx[b];
            exact_sphere<false>(geom[i], i, dir, org, od, minIndex, minDistanceSq);
            ++exactEvals;
        }
    }
}

// FindClosestIntersectionSphere (Collision.hpp:87-109) = big spheres + pre-filter + exact candidates, with the
// whole filter blob addressable (shared memory, or global memory for the probes).  Returns exactly what
// closest_hit_strict returns.  `exactEvals` counts exact_sphere calls (instrumentation).
template <typename GeomPtr>
__device__ __forceinline__ uint32_t closest_hit_filtered(const FilterScene &fs, GeomPtr geom, uint32_t n, V3 dir, V3 org,
                                                         uint32_t &exactEvals, bool live = true)
{
    uint32_t minIndex = n;
    float minDistanceSq = FLT_MAX;
    float const od = vdot(org, dir);
    exact_big_spheres(fs, geom, dir, org, od, minIndex, minDistanceSq, exactEvals, live);
    FilterRay const fr = make_filter_ray(dir, org, od, fs.amax, fs.nearOO);
    filter_span(fs.pairs, fs.slotIdx, fs.shape, fr, geom, n, dir, org, od, minIndex, minDistanceSq, exactEvals, live);
    return minIndex;
}

// ---- camera form, TWO rays per lane -----------------------------------------------------------------
// The camera form is so cheap (4 FFMA2 per record) that the shared-memory pipe, not the issue port, bounds it: two
// broadcast LDS.128 are four LSU wavefronts per record and warp, 16 per SM for every 12 issue cycles of its four
// sub-partitions (ncu: L1/TEX throughput 94 % on primary_hits_kernel with one ray per lane).  Running two camera rays
// of the same pixel per lane through each loaded record halves the wavefronts per test and doubles the independent
// FMA chains.
__device__ __forceinline__ void filter_pair_cam2(const FilterPair *__restrict__ rec, const FilterRay &ra, const FilterRay &rb,
                                                 uint32_t &maskA, uint32_t &maskB)
{
    ulonglong2 const a = *reinterpret_cast<const ulonglong2 *>(rec->cx);
    ulonglong2 const b = *reinterpret_cast<const ulonglong2 *>(rec->cz);
    unsigned long long const tca = fma2(b.x, ra.dz, fma2(a.y, ra.dy, fma2(a.x, ra.dx, ra.nod)));
    unsigned long long const tcb = fma2(b.x, rb.dz, fma2(a.y, rb.dy, fma2(a.x, rb.dx, rb.nod)));
    unsigned long long const va = fma2(tca, tca, b.y);
    unsigned long long const vb = fma2(tcb, tcb, b.y);
    float a0, a1, b0, b1;
    unpack2(va, a0, a1);
    unpack2(vb, b0, b1);
    maskA = __funnelshift_l(__float_as_uint(a0), maskA, 1);
    maskA = __funnelshift_l(__float_as_uint(a1), maskA, 1);
    maskB = __funnelshift_l(__float_as_uint(b0), maskB, 1);
    maskB = __funnelshift_l(__float_as_uint(b1), maskB, 1);
}

// filter_span for two camera rays (A, B) sharing the origin; liveA / liveB as `live` in filter_span
template <typename GeomPtr>
__device__ __forceinline__ void filter_span_cam2(const FilterPair *rec, const uint32_t *slotIdx, SpanShape const shape, const FilterRay &fa,
                                                 const FilterRay &fb, GeomPtr geom, uint32_t n, V3 org, V3 dirA, float odA, V3 dirB,
                                                 float odB, uint32_t &minA, float &distA, uint32_t &minB, float &distB,
                                                 uint32_t &exactEvals, bool liveA, bool liveB)
{
    uint32_t const nWords = shape.nWords, tailQuads = shape.tailQuads, nFull = shape.nFull;
    uint32_t a0 = 0u, a1 = 0u, a2 = 0u, a3 = 0u, a4 = 0u, b0 = 0u, b1 = 0u, b2 = 0u, b3 = 0u, b4 = 0u, pend = 0u;
#pragma unroll 1
    for (uint32_t w = 0; w < nWords;)
    {
        uint32_t maskA = 0xFFFFFFFFu, maskB = 0xFFFFFFFFu;
        if (w < nFull)
        {
#pragma unroll
            for (uint32_t pp = 0; pp < kWordSpheres / 2; ++pp)
                filter_pair_cam2(rec + pp, fa, fb, maskA, maskB);
            rec += kWordSpheres / 2;
        }
        else
        {
#pragma unroll 1
            for (uint32_t q = 0; q < tailQuads; ++q)
            {
#pragma unroll
                for (uint32_t pp = 0; pp < kQuadSpheres / 2; ++pp)
                    filter_pair_cam2(rec + pp, fa, fb, maskA, maskB);
                rec += kQuadSpheres / 2;
            }
            uint32_t const sh = kWordSpheres - kQuadSpheres * tailQuads;
            maskA = (maskA << sh) | ((1u << sh) - 1u);
            maskB = (maskB << sh) | ((1u << sh) - 1u);
        }
        a4 = a3; a3 = a2; a2 = a1; a1 = a0; a0 = liveA ? ~maskA : 0u;
        b4 = b3; b3 = b2; b2 = b1; b1 = b0; b0 = liveB ? ~maskB : 0u;
        ++w;
        ++pend;
        if (resolve_now(pend, w, nWords))
        {
            resolve_candidates(a0, a1, a2, a3, a4, w, slotIdx, geom, n, dirA, org, odA, minA, distA, exactEvals);
            resolve_candidates(b0, b1, b2, b3, b4, w, slotIdx, geom, n, dirB, org, odB, minB, distB, exactEvals);
            pend = 0u;
        }
    }
}

// ---- pixel cone: the spheres ANY camera sample of a pixel can hit, collected once per pixel and launch -----------------
// The camera records carry a radius fattened by |C - E| * delta, delta >= |D - Dc| for every jittered ray D of a pixel with
// central ray Dc (host: pixel_cone_delta / ensure_camera_blob in spt_cuda.cu, where the bound is derived): a sphere that
// passes the reference's exact test for some sample of the pixel passes the camera-form test for Dc.  cone_collect runs that
// test over a span of slots for Dc and appends the sphere indices of the survivors to the lane's list -- a column of a
// shared-memory array, entry k of thread t at list[k * THREADS + t] -- counting past `cap` without storing (overflow: the
// caller falls back to filtering per sample).  Once per pixel and launch, so nothing here is hot; the per-sample work that
// remains is the exact sequence on the listed spheres (cone_closest_hit), which alone decides the hit.
template <int THREADS>
__device__ __forceinline__ void cone_collect(const FilterPair *rec, const uint32_t *slotIdx, uint32_t spanSlots, const FilterRay &fr, uint32_t n,
                                             bool live, uint16_t *list, uint32_t cap, uint32_t &count)
{
    uint32_t const nWords = (spanSlots + kWordSpheres - 1u) / kWordSpheres;
#pragma unroll 1
    for (uint32_t w = 0; w < nWords; ++w)
    {
        uint32_t const slots = min(kWordSpheres, spanSlots - w * kWordSpheres); // a multiple of kQuadSpheres
        uint32_t mask = 0xFFFFFFFFu;
#pragma unroll 1
        for (uint32_t q = 0; q < slots / kQuadSpheres; ++q)
        {
#pragma unroll
            for (uint32_t pp = 0; pp < kQuadSpheres / 2; ++pp)
                filter_pair_cam(rec + pp, fr, mask);
            rec += kQuadSpheres / 2;
        }
        uint32_t const sh = kWordSpheres - slots; // missing slots of a partial word: fail bits
        if (sh)
            mask = (mask << sh) | ((1u << sh) - 1u);
        uint32_t cand = live ? ~mask : 0u;
        while (cand)
        {
            uint32_t const j = __clz(cand);
            cand &= ~(0x80000000u >> j);
            uint32_t const i = slotIdx[w * kWordSpheres + j];
            if (i < n) // padding records can only show up for NaN rays
            {
                if (count < cap)
                    list[count * THREADS + threadIdx.x] = uint16_t(i);
                ++count;
            }
        }
    }
}

// FindClosestIntersectionSphere (Collision.hpp:87-109) for a camera ray of a pixel whose candidate list is complete: the big
// spheres and the listed ones through the exact sequence.  Returns what closest_hit_strict returns.
template <int THREADS, typename GeomPtr>
__device__ __forceinline__ uint32_t cone_closest_hit(const FilterScene &fs, GeomPtr geom, uint32_t n, V3 dir, V3 org, const uint16_t *list,
                                                     uint32_t count, uint32_t &exactEvals, bool live)
{
    uint32_t minIndex = n;
    float minDistanceSq = FLT_MAX;
    float const od = vdot(org, dir);
    exact_big_spheres(fs, geom, dir, org, od, minIndex, minDistanceSq, exactEvals, live);
    if (live)
    {
        for (uint32_t k = 0; k < count; ++k)
        {
            uint32_t const i = list[k * THREADS + threadIdx.x];
            exact_sphere<false>(geom[i], i, dir, org, od, minIndex, minDistanceSq);
        }
        exactEvals += count;
    }
    return minIndex;
}

// the same for a camera ray (org = the eye the records in `fs` were specialised for)
template <typename GeomPtr>
__device__ __forceinline__ uint32_t closest_hit_camera(const FilterScene &fs, GeomPtr geom, uint32_t n, V3 dir, V3 org,
                                                       uint32_t &exactEvals, bool live = true)
{
    uint32_t minIndex = n;
    float minDistanceSq = FLT_MAX;
    float const od = vdot(org, dir);
    exact_big_spheres(fs, geom, dir, org, od, minIndex, minDistanceSq, exactEvals, live);
    FilterRay const fr = make_filter_ray_cam(dir, od, fs.epsTCam);
    filter_span<true>(fs.pairs, fs.slotIdx, fs.shape, fr, geom, n, dir, org, od, minIndex, minDistanceSq, exactEvals, live);
    return minIndex;
}

// two camera rays of one lane (resident blob)
template <typename GeomPtr>
__device__ __forceinline__ void closest_hit_camera2(const FilterScene &fs, GeomPtr geom, uint32_t n, V3 org, V3 dirA, V3 dirB,
                                                    uint32_t &idxA, uint32_t &idxB, uint32_t &exactEvals, bool liveA, bool liveB)
{
    idxA = idxB = n;
    float distA = FLT_MAX, distB = FLT_MAX;
    float const odA = vdot(org, dirA), odB = vdot(org, dirB);
    exact_big_spheres(fs, geom, dirA, org, odA, idxA, distA, exactEvals, liveA);
    exact_big_spheres(fs, geom, dirB, org, odB, idxB, distB, exactEvals, liveB);
    FilterRay const fa = make_filter_ray_cam(dirA, odA, fs.epsTCam), fb = make_filter_ray_cam(dirB, odB, fs.epsTCam);
    filter_span_cam2(fs.pairs, fs.slotIdx, fs.shape, fa, fb, geom, n, org, dirA, odA, dirB, odB, idxA, distA, idxB, distB, exactEvals, liveA,
                     liveB);
}

} // namespace spt
