// kernels_tu.cu — one kernel translation unit. Compiled once per arithmetic mode (Makefile):
//   -DRT_NAMESPACE=rt_exact -fmad=false            the reference's arithmetic, bit for bit (the default mode)
//   -DRT_NAMESPACE=rt_fast  -fmad=true -DRT_FAST   contraction and the FAST-mode traversal arithmetic
// and exports its launchers through a KernelTable (kernel_table.h).
#include "kernel_table.h"
#include "kernels.cuh"

#ifdef RT_EMULATE
#define RT_LAUNCH(kernel, grid, block, stream, concurrent, ...) emu::launch(grid, block, concurrent, [=]() { kernel(__VA_ARGS__); })
#else
#define RT_LAUNCH(kernel, grid, block, stream, concurrent, ...) kernel<<<grid, block, 0, stream>>>(__VA_ARGS__)
#endif

namespace RT_NAMESPACE {

typedef void (*RenderKernel)(const SceneParams, unsigned *, float *, FrameCounters *);

// The instantiation for a frame: frames[] sized by the bounce limit (1 / 3 / 8), BLAS stack by the deepest tree (32 / 64).
static RenderKernel pick_render(const SceneParams & P, int count) {
	if (count) return render_kernel<true, RT_MAX_BOUNCES, 64>;
	const bool deep = P.blas_stack_entries > 32 || P.tlas_stack_entries > RT_TLAS_STACK_SMALL;
	if (P.queue_mode != 0) {
		if (deep) return render_queue_kernel<8, 64>;
		return P.bounces <= 3 ? render_queue_kernel<3, 32> : render_queue_kernel<8, 32>;
	}
	if (deep) return render_kernel<false, 8, 64>;
	return P.bounces <= 1 ? render_kernel<false, 1, 32> : P.bounces <= 3 ? render_kernel<false, 3, 32> : render_kernel<false, 8, 32>;
}

static cudaError_t tu_init(const float ewa[128]) { return cudaMemcpyToSymbol(c_ewa_weight_table, ewa, 128 * sizeof(float)); }

static cudaError_t tu_render_occupancy(const SceneParams & P, int count, int * blocks_per_sm) {
	return cudaOccupancyMaxActiveBlocksPerMultiprocessor(blocks_per_sm, pick_render(P, count), RT_BLOCK_THREADS, 0);
}

static cudaError_t tu_render(const SceneParams & P, unsigned * frame_buffer, float * colour_float, FrameCounters * counters, int count, unsigned blocks, cudaStream_t stream) {
	RenderKernel kernel = pick_render(P, count);
	RT_LAUNCH(kernel, blocks, RT_BLOCK_THREADS, stream, true, P, frame_buffer, colour_float, counters);
	return cudaGetLastError();
}

static cudaError_t tu_primary_hits(const SceneParams & P, int * ids, float * hits, cudaStream_t stream) {
	const dim3 grid((P.width + 31) / 32, (P.height + 7) / 8);
	RT_LAUNCH(primary_hits_kernel, grid, RT_BLOCK_THREADS, stream, false, P, ids, hits);
	return cudaGetLastError();
}

static cudaError_t tu_schedule(unsigned * tile_cost, unsigned * tile_order, int tile_count, FrameCounters * next_counters, cudaStream_t stream) {
	RT_LAUNCH(tile_schedule_kernel, 1, 1024, stream, false, tile_cost, tile_order, tile_count, next_counters);
	return cudaGetLastError();
}

static cudaError_t tu_fxaa(const unsigned * frame_buffer, unsigned char * rgb8, const float * gamma_table, int width, int height, int enable_fxaa, int row_begin, int row_end, cudaStream_t stream) {
	const dim3 block(FXAA_TILE_W, FXAA_TILE_H);
	const dim3 grid((width + FXAA_TILE_W - 1) / FXAA_TILE_W, (row_end - row_begin + FXAA_TILE_H - 1) / FXAA_TILE_H);
	RT_LAUNCH(fxaa_kernel, grid, block, stream, false, frame_buffer, rgb8, gamma_table, width, height, enable_fxaa, row_begin, row_end);
	return cudaGetLastError();
}

static cudaError_t tu_texture_level0(const unsigned char * rgb8, const float * gamma_lut, float4 * out, long long texel_count, cudaStream_t stream) {
	const unsigned blocks = unsigned(texel_count + 255 < 256ll * 148 * 8 ? (texel_count + 255) / 256 : 148 * 8);
	RT_LAUNCH(texture_level0_kernel, blocks ? blocks : 1u, 256, stream, false, rgb8, gamma_lut, out, texel_count);
	return cudaGetLastError();
}

static cudaError_t tu_texture_mip(const float4 * previous, float4 * out, int width, int height, int previous_width, cudaStream_t stream) {
	RT_LAUNCH(texture_mip_kernel, dim3((width + 31) / 32, (height + 7) / 8), 256, stream, false, previous, out, width, height, previous_width);
	return cudaGetLastError();
}

static cudaError_t tu_texel_halves(const float4 * texels, uint2 * out, long long texel_count, cudaStream_t stream) {
	const unsigned blocks = unsigned(texel_count + 255 < 256ll * 148 * 8 ? (texel_count + 255) / 256 : 148 * 8);
	RT_LAUNCH(texel_halves_kernel, blocks ? blocks : 1u, 256, stream, false, texels, out, texel_count);
	return cudaGetLastError();
}

static cudaError_t tu_barrier(unsigned * const * flag_blocks, int count, int rank, unsigned epoch, cudaStream_t stream) {
	BarrierPeers peers;
	for (int q = 0; q < B200RT_MAX_BANDS; q++) peers.block[q] = q < count ? flag_blocks[q] : nullptr;
	RT_LAUNCH(barrier_kernel, 1, 32, stream, false, peers, count, rank, epoch);
	return cudaGetLastError();
}

#define RT_STRINGIFY2(x) #x
#define RT_STRINGIFY(x) RT_STRINGIFY2(x)
static const rt::KernelTable table = { RT_STRINGIFY(RT_NAMESPACE), RT_FEATURE_BITS, tu_init, tu_render_occupancy, tu_render, tu_primary_hits, tu_schedule, tu_fxaa, tu_texture_level0, tu_texture_mip, tu_texel_halves, tu_barrier };

} // namespace RT_NAMESPACE

namespace rt {
#ifdef RT_FAST
const KernelTable & kernels_fast() { return rt_fast::table; }
#else
const KernelTable & kernels_exact() { return rt_exact::table; }
#endif
}
