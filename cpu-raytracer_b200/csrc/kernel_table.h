// kernel_table.h — what a kernel translation unit (kernels_tu.cu, compiled once per arithmetic mode) offers the C ABI
// layer (b200rt.cu). Plain function pointers: b200rt.cu holds no device code and no template arguments.
#pragma once
#include "rt_types.h"

namespace rt {

struct KernelTable {
	const char * name;   // "rt_exact" | "rt_fast"
	int fast_features;   // RT_FAST_FEATURES of the unit: 2 = it tests precomputed triangle transforms (SceneParams::triangles_test), 4 = it reads SceneParams::texels_half
	// constant tables of the unit (Texture::ewa_weight_table, Texture.h:20-21,53-62)
	cudaError_t (*init)(const float ewa_weight_table[128]);
	// resident blocks per SM of the render kernel the launch below would pick for these parameters
	cudaError_t (*render_occupancy)(const SceneParams & P, int count, int * blocks_per_sm);
	// Raytracer::render_tile over this context's tiles; instantiation chosen from P (bounces, blas_stack_entries, queue_mode)
	cudaError_t (*render)(const SceneParams & P, unsigned * frame_buffer, float * colour_float, FrameCounters * counters, int count, unsigned blocks, cudaStream_t stream);
	cudaError_t (*primary_hits)(const SceneParams & P, int * ids, float * hits, cudaStream_t stream);
	// orders the tiles for the next frame and clears the next frame's counter block
	cudaError_t (*schedule)(unsigned * tile_cost, unsigned * tile_order, int tile_count, FrameCounters * next_counters, cudaStream_t stream);
	cudaError_t (*fxaa)(const unsigned * frame_buffer, unsigned char * rgb8, const float * gamma_table, int width, int height, int enable_fxaa, int row_begin, int row_end, cudaStream_t stream);
	cudaError_t (*texture_level0)(const unsigned char * rgb8, const float * gamma_lut, float4 * out, long long texel_count, cudaStream_t stream);
	cudaError_t (*texture_mip)(const float4 * previous, float4 * out, int width, int height, int previous_width, cudaStream_t stream);
	cudaError_t (*texel_halves)(const float4 * texels, uint2 * out, long long texel_count, cudaStream_t stream);
	cudaError_t (*barrier)(unsigned * const * flag_blocks, int count, int rank, unsigned epoch, cudaStream_t stream);
};

const KernelTable & kernels_exact();   // -fmad=false: the reference's association order, bit for bit
#ifdef RT_HAVE_FAST
const KernelTable & kernels_fast();    // -fmad=true -DRT_FAST: inside north_star's tolerance (99.9 % ids, 50 dB)
#endif

} // namespace rt
