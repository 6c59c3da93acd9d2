// rt_math.cuh — per-thread FP32 vector math with the reference's association order.
//
// One CUDA thread carries one ray, which is one lane of the reference's SIMD packet
// (SIMD_Vector3.h, SIMD_floats.h). A multiply-add is fused exactly where the reference calls
// madd / msub (SIMD_floats.h:223-224 -> _mm256_fmadd_ps / _mm256_fmsub_ps; used by dot, cross
// and the matrix transforms, SIMD_Vector3.h:53-63, Matrix4.h:39-85) and nowhere else: the file
// is compiled with -fmad=false and spells every fused operation as __fmaf_rn. Division and
// square root are IEEE (nvcc defaults -prec-div=true -prec-sqrt=true; no --use_fast_math).
#pragma once
#include "rt_types.h"
#ifndef RT_EMULATE
#include <math_constants.h>
#endif

// The kernels are compiled once per arithmetic mode (kernels_tu.cu): rt_exact with -fmad=false, rt_fast with
// -fmad=true -DRT_FAST. Same source, two namespaces, so both sets of kernels live in one library.
#ifndef RT_NAMESPACE
#define RT_NAMESPACE rt_exact
#endif
// What the FAST unit changes besides contraction (-fmad=true), one bit each so that every candidate can be built and
// timed alone (Makefile: fast_* variants; DESIGN.md has the table): 1 = slab test as one multiply-add per plane,
// 2 = precomputed-transform triangle test + approximate reciprocal, 4 = RGBA16F texels.
#ifdef RT_FAST
#ifndef RT_FAST_FEATURES
#define RT_FAST_FEATURES 7
#endif
#if RT_FAST_FEATURES & 1
#define RT_FAST_SLAB
#endif
#if RT_FAST_FEATURES & 2
#define RT_FAST_TRI
#endif
#if RT_FAST_FEATURES & 4
#define RT_FAST_TEXEL
#endif
#define RT_FEATURE_BITS RT_FAST_FEATURES
#else
#define RT_FEATURE_BITS 0
#endif
#if defined(RT_FAST_TEXEL) && !defined(RT_EMULATE)
#include <cuda_fp16.h>
#endif

namespace RT_NAMESPACE {
using namespace rt;

#define RT_EPSILON 0.005f               // Ray::EPSILON, Ray.h:5
#define RT_ONE_OVER_PI     0.31830988618f // Util.h:9
#define RT_ONE_OVER_TWO_PI 0.15915494309f // Util.h:12

struct v3 { float x, y, z; };

// Loads of the two read-only streams a ray pulls through L1: BVH nodes + triangles and texels, both ld.global.nc.
// (Cache-policy hints — L1::evict_last / evict_first / no_allocate on either stream — were timed in round 1 and
// changed nothing beyond noise: profiles/r01_variants_cache_hints.log.)
__device__ __forceinline__ float4 load_node(const float4 * p)  { return __ldg(p); }
__device__ __forceinline__ float4 load_texel(const float4 * p) { return __ldg(p); }
// -DRT_PREFETCH_SKIP (experiment, off by default): a threaded record's skip target is known as soon as the record
// is; asking for its line while the box test runs takes the load off the path of a ray that misses the box.
__device__ __forceinline__ void prefetch_record(const float4 * p) {
#if defined(RT_PREFETCH_SKIP) && !defined(RT_EMULATE)
	asm volatile("prefetch.global.L1 [%0];" : : "l"(p));
#else
	(void) p;
#endif
}
// %globaltimer: nanoseconds, one clock for the whole GPU (the kernels time themselves with it)
__device__ __forceinline__ unsigned long long global_timer_ns() {
#ifdef RT_EMULATE
	return emu::global_timer_ns();
#else
	unsigned long long now;
	asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(now));
	return now;
#endif
}

__device__ __forceinline__ v3 V(float x, float y, float z) { v3 r; r.x = x; r.y = y; r.z = z; return r; }
__device__ __forceinline__ v3 V(float4 a) { return V(a.x, a.y, a.z); }
__device__ __forceinline__ v3 operator+(v3 a, v3 b) { return V(a.x + b.x, a.y + b.y, a.z + b.z); }
__device__ __forceinline__ v3 operator-(v3 a, v3 b) { return V(a.x - b.x, a.y - b.y, a.z - b.z); }
__device__ __forceinline__ v3 operator*(v3 a, v3 b) { return V(a.x * b.x, a.y * b.y, a.z * b.z); }
__device__ __forceinline__ v3 operator*(v3 a, float f) { return V(a.x * f, a.y * f, a.z * f); }   // v * f and f * v (SIMD_Vector3.h:150-158)
__device__ __forceinline__ v3 addf(v3 a, float f) { return V(a.x + f, a.y + f, a.z + f); }        // f + v
__device__ __forceinline__ v3 neg(v3 a) { return V(0.0f - a.x, 0.0f - a.y, 0.0f - a.z); }          // SIMD_Vector3.h:138-141

__device__ __forceinline__ float dot(v3 a, v3 b) { return __fmaf_rn(a.x, b.x, __fmaf_rn(a.y, b.y, a.z * b.z)); } // :53-55
__device__ __forceinline__ v3 cross(v3 a, v3 b) {                                                               // :57-63
	return V(__fmaf_rn(a.y, b.z, -(a.z * b.y)), __fmaf_rn(a.z, b.x, -(a.x * b.z)), __fmaf_rn(a.x, b.y, -(a.y * b.x)));
}
__device__ __forceinline__ v3 madd(v3 a, float b, v3 c) { return V(__fmaf_rn(a.x, b, c.x), __fmaf_rn(a.y, b, c.y), __fmaf_rn(a.z, b, c.z)); } // :107-113
__device__ __forceinline__ float inv_sqrt(float x) { return 1.0f / sqrtf(x); }
__device__ __forceinline__ v3 normalize(v3 a) { return a * inv_sqrt(dot(a, a)); }            // :44-51
__device__ __forceinline__ v3 divf(v3 a, float f) { return a * (1.0f / f); }                 // v / f = v * rcp(f), :153

// x86 MINPS/MAXPS semantics: the second operand comes back when either one is NaN.
__device__ __forceinline__ float sse_min(float a, float b) { return a < b ? a : b; }
__device__ __forceinline__ float sse_max(float a, float b) { return a > b ? a : b; }

// std::max / std::min as the scalar texture code uses them (Texture.cpp)
__device__ __forceinline__ float std_max(float a, float b) { return a < b ? b : a; }
__device__ __forceinline__ float std_min(float a, float b) { return b < a ? b : a; }

// Util::float_to_int (cvtss2si, Util.h:39-41) and _mm256_cvtps_epi32: round to nearest even;
// NaN and out-of-range give 0x80000000 (CUDA's cvt.rni saturates and maps NaN to 0 instead).
__device__ __forceinline__ int float_to_int(float x) {
	return (x >= -2147483648.0f && x < 2147483648.0f) ? __float2int_rn(x) : (int)0x80000000u;
}

__device__ __forceinline__ int positive_mod(int value, int modulus) { // Math::mod, Math.h:46-53
	int r = value % modulus;
	return r < 0 ? r + modulus : r;
}
// The same for a modulus that is KNOWN to be a power of two — every level of a mip-mapped texture (Texture.cpp:50): the
// positive modulo is the low bits, for negative values too (two's complement); the same integer without the ~40
// instructions of a division by a run-time divisor (4 % of the kernel's instructions under EWA filtering).
__device__ __forceinline__ int positive_mod_pow2(int value, int modulus) { return value & (modulus - 1); }

__device__ __forceinline__ v3 transform_position(const Affine & m, v3 p) {
	return V(__fmaf_rn(m.r0.x, p.x, __fmaf_rn(m.r0.y, p.y, __fmaf_rn(m.r0.z, p.z, m.r0.w))),
	         __fmaf_rn(m.r1.x, p.x, __fmaf_rn(m.r1.y, p.y, __fmaf_rn(m.r1.z, p.z, m.r1.w))),
	         __fmaf_rn(m.r2.x, p.x, __fmaf_rn(m.r2.y, p.y, __fmaf_rn(m.r2.z, p.z, m.r2.w))));
}
__device__ __forceinline__ v3 transform_direction(const Affine & m, v3 d) {
	return V(__fmaf_rn(m.r0.x, d.x, __fmaf_rn(m.r0.y, d.y, m.r0.z * d.z)),
	         __fmaf_rn(m.r1.x, d.x, __fmaf_rn(m.r1.y, d.y, m.r1.z * d.z)),
	         __fmaf_rn(m.r2.x, d.x, __fmaf_rn(m.r2.y, d.y, m.r2.z * d.z)));
}

} // namespace RT_NAMESPACE
