// cuda_emu.h — TEST INFRASTRUCTURE. Runs the kernels' SOURCE (cpu-raytracer_b200/csrc/*.cuh, *.cu) on the host.
//
// There is no GPU in the build container, and a gpurun call costs minutes of a small budget, so the logic of
// the kernels — the frame-stack walk, the deep-pixel queue with its countdowns and its termination protocol, the
// threaded layouts — is first run here: tests/emu/Makefile compiles csrc/b200rt.cu and csrc/kernels_tu.cu with g++
// (-DRT_EMULATE, -ffp-contract=off) into tests/emu/libb200rt_emu.so, which exports the same C ABI as libb200rt.so,
// and tests/test_kernel_source_on_host.py holds its frames against the oracle.
//
//  * every CUDA thread is a fibre (ucontext); a launch runs its blocks one after the other, or all at once for the
//    persistent kernels whose warps wait for one another; fibres switch at warp collectives, __syncthreads,
//    __nanosleep and — so that interleavings get exercised — at every atomic;
//  * warp collectives (__shfl*_sync, __syncwarp) are rendezvous of the lanes named in the mask;
//  * the CUDA runtime calls b200rt.cu makes are mapped onto malloc / memcpy; streams and events do nothing, because
//    every launch has completed when it returns.
// It is NOT a fallback of the product: nothing in cpu-raytracer_b200/ or bench.py loads it, it proves nothing about
// the GPU's memory model or speed, and libm stands in for CUDA's math functions (the same functions the oracle calls,
// so frames equal the oracle's bit for bit).
#pragma once
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <functional>

#define __device__
#define __host__
#define __global__
#define __forceinline__ inline __attribute__((always_inline))
#define __noinline__ __attribute__((noinline))
#define __launch_bounds__(...)
#define __constant__
#define __shared__ static
#define __align__(n) __attribute__((aligned(n)))
#define CUDART_INF_F (__builtin_inff())

struct __attribute__((aligned(16))) float4 { float x, y, z, w; };
struct __attribute__((aligned(16))) int4 { int x, y, z, w; };
struct uint3 { unsigned x, y, z; };
struct uint2 { unsigned x, y; };
static inline uint2 make_uint2(unsigned x, unsigned y) { uint2 r; r.x = x; r.y = y; return r; }
struct dim3 { unsigned x, y, z; dim3(unsigned x_ = 1, unsigned y_ = 1, unsigned z_ = 1) : x(x_), y(y_), z(z_) {} };
static inline float4 make_float4(float x, float y, float z, float w) { float4 r; r.x = x; r.y = y; r.z = z; r.w = w; return r; }

// ------------------------------------------------------------------------------------ runtime API ---
typedef int cudaError_t;
enum { cudaSuccess = 0, cudaErrorMemoryAllocation = 2, cudaErrorInvalidValue = 1, cudaErrorNotSupported = 801 };
typedef struct emu_stream * cudaStream_t;
typedef struct emu_event * cudaEvent_t;
enum cudaMemcpyKind { cudaMemcpyHostToDevice = 1, cudaMemcpyDeviceToHost = 2, cudaMemcpyDeviceToDevice = 3 };
enum { cudaStreamNonBlocking = 1, cudaEventDisableTiming = 2, cudaHostAllocDefault = 0, cudaMemoryTypeHost = 1, cudaMemoryTypeUnregistered = 0 };
struct cudaDeviceProp { int multiProcessorCount; };
struct cudaPointerAttributes { int type; };

namespace emu {
void * device_alloc(size_t bytes);
void   device_free(void * p);
void * host_alloc(size_t bytes);
void   host_free(void * p);
bool   is_host_alloc(const void * p);
double now_ms();
unsigned long long global_timer_ns();
void launch(dim3 grid, dim3 block, bool blocks_concurrent, const std::function<void()> & body);

struct Thread;                       // one CUDA thread = one fibre
extern Thread * current;
uint3 & thread_idx(); uint3 & block_idx(); dim3 & block_dim(); dim3 & grid_dim();
void yield();
void sync_threads();
unsigned warp_exchange(unsigned mask, unsigned value, int source_lane);   // all lanes of mask deposit `value`, each reads lane source_lane's
long long clock();
}
struct emu_event { double ms; };

static inline cudaError_t cudaGetDeviceCount(int * n) { *n = 1; return cudaSuccess; }
static inline cudaError_t cudaSetDevice(int) { return cudaSuccess; }
static inline cudaError_t cudaGetDeviceProperties(cudaDeviceProp * p, int) { p->multiProcessorCount = 3; return cudaSuccess; }
static inline const char * cudaGetErrorString(cudaError_t e) { return e == cudaSuccess ? "no error" : "emulated CUDA error"; }
static inline cudaError_t cudaGetLastError() { return cudaSuccess; }
static inline cudaError_t cudaMalloc(void ** p, size_t n) { *p = emu::device_alloc(n); return *p ? cudaSuccess : cudaErrorMemoryAllocation; }
template<typename T> static inline cudaError_t cudaMalloc(T ** p, size_t n) { return cudaMalloc(reinterpret_cast<void **>(p), n); }
static inline cudaError_t cudaFree(void * p) { emu::device_free(p); return cudaSuccess; }
static inline cudaError_t cudaMallocHost(void ** p, size_t n) { *p = emu::host_alloc(n); return *p ? cudaSuccess : cudaErrorMemoryAllocation; }
static inline cudaError_t cudaHostAlloc(void ** p, size_t n, unsigned) { return cudaMallocHost(p, n); }
static inline cudaError_t cudaFreeHost(void * p) { emu::host_free(p); return cudaSuccess; }
static inline cudaError_t cudaMemcpy(void * d, const void * s, size_t n, cudaMemcpyKind) { if (n) memcpy(d, s, n); return cudaSuccess; }
static inline cudaError_t cudaMemcpyAsync(void * d, const void * s, size_t n, cudaMemcpyKind, cudaStream_t) { if (n) memcpy(d, s, n); return cudaSuccess; }
static inline cudaError_t cudaMemset(void * d, int v, size_t n) { memset(d, v, n); return cudaSuccess; }
static inline cudaError_t cudaMemsetAsync(void * d, int v, size_t n, cudaStream_t) { memset(d, v, n); return cudaSuccess; }
template<typename T> static inline cudaError_t cudaGetSymbolAddress(void ** p, T & symbol) { *p = &symbol; return cudaSuccess; }
template<typename T> static inline cudaError_t cudaMemcpyToSymbol(T & symbol, const void * s, size_t n) { memcpy(&symbol, s, n); return cudaSuccess; }
static inline cudaError_t cudaStreamCreateWithFlags(cudaStream_t * s, unsigned) { *s = reinterpret_cast<cudaStream_t>(malloc(8)); return cudaSuccess; }
static inline cudaError_t cudaStreamDestroy(cudaStream_t s) { free(s); return cudaSuccess; }
static inline cudaError_t cudaStreamSynchronize(cudaStream_t) { return cudaSuccess; }
static inline cudaError_t cudaStreamWaitEvent(cudaStream_t, cudaEvent_t, unsigned) { return cudaSuccess; }
static inline cudaError_t cudaEventCreate(cudaEvent_t * e) { *e = new emu_event{ 0.0 }; return cudaSuccess; }
static inline cudaError_t cudaEventCreateWithFlags(cudaEvent_t * e, unsigned) { return cudaEventCreate(e); }
static inline cudaError_t cudaEventDestroy(cudaEvent_t e) { delete e; return cudaSuccess; }
static inline cudaError_t cudaEventRecord(cudaEvent_t e, cudaStream_t) { e->ms = emu::now_ms(); return cudaSuccess; }
static inline cudaError_t cudaEventSynchronize(cudaEvent_t) { return cudaSuccess; }
static inline cudaError_t cudaEventQuery(cudaEvent_t) { return cudaSuccess; }   // launches run to completion before they return
static inline cudaError_t cudaEventElapsedTime(float * ms, cudaEvent_t a, cudaEvent_t b) { *ms = float(b->ms - a->ms); return cudaSuccess; }
static inline cudaError_t cudaPointerGetAttributes(cudaPointerAttributes * a, const void * p) { a->type = emu::is_host_alloc(p) ? cudaMemoryTypeHost : cudaMemoryTypeUnregistered; return cudaSuccess; }
struct cudaIpcMemHandle_t { char reserved[64]; };
enum { cudaIpcMemLazyEnablePeerAccess = 1 };
static inline cudaError_t cudaIpcGetMemHandle(cudaIpcMemHandle_t * h, void * p) { memset(h, 0, sizeof(*h)); memcpy(h, &p, sizeof(p)); return cudaSuccess; }   // one process: the handle IS the pointer
static inline cudaError_t cudaIpcOpenMemHandle(void ** p, cudaIpcMemHandle_t h, unsigned) { memcpy(p, &h, sizeof(*p)); return cudaSuccess; }
static inline cudaError_t cudaIpcCloseMemHandle(void *) { return cudaSuccess; }
template<typename K> static inline cudaError_t cudaOccupancyMaxActiveBlocksPerMultiprocessor(int * n, K, int, size_t) { *n = 1; return cudaSuccess; }

// ------------------------------------------------------------------------------------- device side ---
#define threadIdx (emu::thread_idx())
#define blockIdx  (emu::block_idx())
#define blockDim  (emu::block_dim())
#define gridDim   (emu::grid_dim())

template<typename T> static inline T __ldg(const T * p) { return *p; }
template<typename T> static inline T __ldcg(const T * p) { return *reinterpret_cast<const volatile T *>(p); }
static inline float4 __ldcg(const float4 * p) { emu::yield(); return *p; }
template<typename T> static inline void __stcg(T * p, T v) { *p = v; }
static inline float __fmaf_rn(float a, float b, float c) { return fmaf(a, b, c); }
static inline int   __float_as_int(float f) { int i; memcpy(&i, &f, 4); return i; }
static inline float __int_as_float(int i) { float f; memcpy(&f, &i, 4); return f; }
static inline int   __float2int_rn(float f) { return int(lrintf(f)); }   // callers keep the argument in range (rt_math.cuh: float_to_int)
static inline void  __threadfence() {}
static inline void  __threadfence_system() {}
static inline void  __syncthreads() { emu::sync_threads(); }
static inline void  __nanosleep(unsigned) { emu::yield(); }
static inline long long clock64() { return emu::clock(); }
static inline int __popc(unsigned v) { return __builtin_popcount(v); }
using std::min; using std::max;

static inline void __syncwarp(unsigned mask = 0xffffffffu) { emu::warp_exchange(mask, 0, -1); }
static inline unsigned emu_lane() { return (threadIdx.x + threadIdx.y * blockDim.x) & 31u; }
static inline unsigned __ballot_sync(unsigned mask, bool predicate) { return emu::warp_exchange(mask, predicate ? 1u << emu_lane() : 0u, -2); }
static inline int __ffs(unsigned v) { return __builtin_ffs(int(v)); }
template<typename T> static inline T __shfl_sync(unsigned mask, T v, int lane) { unsigned bits; static_assert(sizeof(T) == 4, "32-bit shuffles only"); memcpy(&bits, &v, 4); bits = emu::warp_exchange(mask, bits, lane & 31); memcpy(&v, &bits, 4); return v; }
template<typename T> static inline T __shfl_xor_sync(unsigned mask, T v, int lane_mask) { return __shfl_sync(mask, v, int(emu_lane()) ^ lane_mask); }
template<typename T> static inline T __shfl_up_sync(unsigned mask, T v, unsigned delta) { const int lane = int(emu_lane()); return __shfl_sync(mask, v, lane >= int(delta) ? lane - int(delta) : lane); }

#include <immintrin.h>
static inline float emu_half_to_float(unsigned short h) { return _cvtsh_ss(h); }
static inline unsigned emu_float_to_half(float f) { return _cvtss_sh(f, 0); }   // round to nearest even

// atomics: one OS thread runs all fibres, so plain operations are atomic; the yield lets another fibre run between
// this thread's memory operations, as another SM would on the device
template<typename T> static inline T atomicAdd(T * p, T v) { emu::yield(); T o = *p; *p = o + v; return o; }
static inline int atomicSub(int * p, int v) { emu::yield(); int o = *p; *p = o - v; return o; }
template<typename T> static inline T atomicMax(T * p, T v) { emu::yield(); T o = *p; if (v > o) *p = v; return o; }
static inline unsigned atomicOr(unsigned * p, unsigned v) { emu::yield(); unsigned o = *p; *p = o | v; return o; }
static inline unsigned atomicCAS(unsigned * p, unsigned expect, unsigned v) { emu::yield(); unsigned o = *p; if (o == expect) *p = v; return o; }
