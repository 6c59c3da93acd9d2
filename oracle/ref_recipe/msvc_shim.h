// TEST INFRASTRUCTURE — part of the recipe that compiles the UNMODIFIED reference
// sources (where they lie under /root/reference) into oracle/_ref/.
//
// Force-included (-include) into every reference translation unit. It supplies
// the MSVC / Win32 / Intel-SVML names the reference uses that g++ + glibc do
// not have. Nothing here changes the reference's arithmetic except that the
// closed-source SVML transcendentals are served by glibc's libmvec / libm
// (a few ulp apart; SURVEY.md §8(c) "parity unpinned" at that boundary).
#pragma once
#ifdef __cplusplus

#include <immintrin.h>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <cmath>
#include <csignal>

// ---- MSVC keywords ---------------------------------------------------------
#ifndef __forceinline
#define __forceinline __attribute__((always_inline))
#endif
#define _declspec(x)
#define __declspec(x)
static inline void __debugbreak() { raise(SIGTRAP); }

// ---- CRT ---------------------------------------------------------------------
static inline void * _aligned_malloc(size_t size, size_t align) {
	void * p = nullptr;
	if (size == 0) size = align;
	if (posix_memalign(&p, align < sizeof(void *) ? sizeof(void *) : align, size) != 0) return nullptr;
	return p;
}
static inline void _aligned_free(void * p) { free(p); }

static inline int fopen_s(FILE ** file, const char * name, const char * mode) {
	*file = fopen(name, mode);
	return *file ? 0 : 1;
}

// ---- SVML -> libmvec (AVX2 variants, 8 lanes) ------------------------------------
extern "C" {
	__m256 _ZGVdN8v_expf (__m256);
	__m256 _ZGVdN8v_sinf (__m256);
	__m256 _ZGVdN8v_cosf (__m256);
	__m256 _ZGVdN8v_tanf (__m256);
	__m256 _ZGVdN8v_asinf(__m256);
	__m256 _ZGVdN8v_acosf(__m256);
	__m256 _ZGVdN8v_atanf(__m256);
	__m256 _ZGVdN8vv_atan2f(__m256, __m256);

	__m128 _ZGVbN4v_expf (__m128);
	__m128 _ZGVbN4v_sinf (__m128);
	__m128 _ZGVbN4v_cosf (__m128);
	__m128 _ZGVbN4v_tanf (__m128);
	__m128 _ZGVbN4v_asinf(__m128);
	__m128 _ZGVbN4v_acosf(__m128);
	__m128 _ZGVbN4v_atanf(__m128);
	__m128 _ZGVbN4vv_atan2f(__m128, __m128);
}

static inline __m256 _mm256_exp_ps  (__m256 x) { return _ZGVdN8v_expf (x); }
static inline __m256 _mm256_sin_ps  (__m256 x) { return _ZGVdN8v_sinf (x); }
static inline __m256 _mm256_cos_ps  (__m256 x) { return _ZGVdN8v_cosf (x); }
static inline __m256 _mm256_tan_ps  (__m256 x) { return _ZGVdN8v_tanf (x); }
static inline __m256 _mm256_asin_ps (__m256 x) { return _ZGVdN8v_asinf(x); }
static inline __m256 _mm256_acos_ps (__m256 x) { return _ZGVdN8v_acosf(x); }
static inline __m256 _mm256_atan_ps (__m256 x) { return _ZGVdN8v_atanf(x); }
static inline __m256 _mm256_atan2_ps(__m256 y, __m256 x) { return _ZGVdN8vv_atan2f(y, x); }
static inline __m256 _mm256_invsqrt_ps(__m256 x) { return _mm256_div_ps(_mm256_set1_ps(1.0f), _mm256_sqrt_ps(x)); }

static inline __m128 _mm_exp_ps  (__m128 x) { return _ZGVbN4v_expf (x); }
static inline __m128 _mm_sin_ps  (__m128 x) { return _ZGVbN4v_sinf (x); }
static inline __m128 _mm_cos_ps  (__m128 x) { return _ZGVbN4v_cosf (x); }
static inline __m128 _mm_tan_ps  (__m128 x) { return _ZGVbN4v_tanf (x); }
static inline __m128 _mm_asin_ps (__m128 x) { return _ZGVbN4v_asinf(x); }
static inline __m128 _mm_acos_ps (__m128 x) { return _ZGVbN4v_acosf(x); }
static inline __m128 _mm_atan_ps (__m128 x) { return _ZGVbN4v_atanf(x); }
static inline __m128 _mm_atan2_ps(__m128 y, __m128 x) { return _ZGVbN4vv_atan2f(y, x); }
static inline __m128 _mm_invsqrt_ps(__m128 x) { return _mm_div_ps(_mm_set1_ps(1.0f), _mm_sqrt_ps(x)); }

// SVML integer division: per-lane loop
static inline __m256i _mm256_div_epi32(__m256i a, __m256i b) {
	alignas(32) int x[8], y[8];
	_mm256_store_si256(reinterpret_cast<__m256i *>(x), a);
	_mm256_store_si256(reinterpret_cast<__m256i *>(y), b);
	for (int i = 0; i < 8; i++) x[i] = y[i] ? x[i] / y[i] : 0;
	return _mm256_load_si256(reinterpret_cast<const __m256i *>(x));
}
static inline __m128i _mm_div_epi32(__m128i a, __m128i b) {
	alignas(16) int x[4], y[4];
	_mm_store_si128(reinterpret_cast<__m128i *>(x), a);
	_mm_store_si128(reinterpret_cast<__m128i *>(y), b);
	for (int i = 0; i < 4; i++) x[i] = y[i] ? x[i] / y[i] : 0;
	return _mm_load_si128(reinterpret_cast<const __m128i *>(x));
}

#endif // __cplusplus
