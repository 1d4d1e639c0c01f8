/*
 * oracle/ref_shim/intrin.h -- TEST INFRASTRUCTURE ONLY (never on the product path).
 *
 * Stand-in for MSVC's <intrin.h>, found first on the include path when the
 * UNMODIFIED reference headers under /root/reference/include are compiled with
 * g++ (see oracle/Makefile, target oracle/_ref/libspt_ref.so).
 *
 * The reference (include/Math.hpp:3, :9-14, :52-53, :178-186) relies on three
 * MSVC-isms that g++ rejects:
 *   - <intrin.h> itself,
 *   - __m128 being a *union* with a `.m128_f32[4]` member, which is what lets
 *     `Vec4{0, 1, 0}` brace-initialise through `union { __m128 m128; float xyzw[4]; }`,
 *   - std::sqrtf (missing from libstdc++ 13).
 * This header re-creates exactly that environment: an aggregate `spt_m128`
 * with the MSVC member name, thin wrappers that forward every SSE intrinsic
 * the reference uses to the real one (same instruction, same IEEE rounding),
 * and `#define __m128 spt_m128` so the reference text binds to it.
 * No arithmetic is altered: each wrapper is a load, the real intrinsic, a store.
 */
#pragma once

#include <immintrin.h>

#include <algorithm>
#include <atomic>
#include <cfloat>
#include <chrono>
#include <cmath>
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <random>
#include <condition_variable>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

namespace std
{
using ::sqrtf; /* Math.hpp:137, Collision.hpp:24,34,44 call std::sqrtf */
}

typedef __m128 spt_real_m128;

struct spt_m128
{
    float m128_f32[4];

    operator spt_real_m128() const { return _mm_loadu_ps(m128_f32); }
};

static inline spt_m128 spt_wrap(spt_real_m128 v)
{
    spt_m128 r;
    _mm_storeu_ps(r.m128_f32, v);
    return r;
}

static inline spt_m128 spt_mm_sub_ps(spt_m128 a, spt_m128 b) { return spt_wrap(_mm_sub_ps(a, b)); }
static inline spt_m128 spt_mm_add_ps(spt_m128 a, spt_m128 b) { return spt_wrap(_mm_add_ps(a, b)); }
static inline spt_m128 spt_mm_mul_ps(spt_m128 a, spt_m128 b) { return spt_wrap(_mm_mul_ps(a, b)); }
static inline spt_m128 spt_mm_div_ps(spt_m128 a, spt_m128 b) { return spt_wrap(_mm_div_ps(a, b)); }
static inline spt_m128 spt_mm_xor_ps(spt_m128 a, spt_m128 b) { return spt_wrap(_mm_xor_ps(a, b)); }
static inline spt_m128 spt_mm_hadd_ps(spt_m128 a, spt_m128 b) { return spt_wrap(_mm_hadd_ps(a, b)); }
static inline spt_m128 spt_mm_sqrt_ps(spt_m128 a) { return spt_wrap(_mm_sqrt_ps(a)); }
static inline spt_m128 spt_mm_set1_ps(float s) { return spt_wrap(_mm_set1_ps(s)); }
static inline spt_m128 spt_mm_load_ps(float const *p) { return spt_wrap(_mm_loadu_ps(p)); }
static inline void spt_mm_store_ps(float *p, spt_m128 a) { _mm_storeu_ps(p, a); }
static inline void spt_mm_store_ss(float *p, spt_m128 a) { _mm_store_ss(p, a); }
/* _mm_dp_ps needs a compile-time immediate: forward through the builtin. */
#define spt_mm_dp_ps(a, b, imm) \
    spt_wrap((spt_real_m128)__builtin_ia32_dpps((__v4sf)(spt_real_m128)(a), (__v4sf)(spt_real_m128)(b), (imm)))

#undef _mm_dp_ps
#define _mm_sub_ps spt_mm_sub_ps
#define _mm_add_ps spt_mm_add_ps
#define _mm_mul_ps spt_mm_mul_ps
#define _mm_div_ps spt_mm_div_ps
#define _mm_xor_ps spt_mm_xor_ps
#define _mm_hadd_ps spt_mm_hadd_ps
#define _mm_sqrt_ps spt_mm_sqrt_ps
#define _mm_set1_ps spt_mm_set1_ps
#define _mm_set_ps1 spt_mm_set1_ps
#define _mm_load_ps spt_mm_load_ps
#define _mm_store_ps spt_mm_store_ps
#define _mm_store_ss spt_mm_store_ss
#define _mm_dp_ps spt_mm_dp_ps
#define __m128 spt_m128
