// kernels.cuh — __global__ entry points.
#pragma once
#include "render.cuh"

namespace rt {

__device__ __forceinline__ unsigned warp_sum(unsigned v) {
#pragma unroll
	for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
	return v;
}

// Maps a work unit to a pixel. The frame is cut into the reference's 32x32 tiles (Window.h:32-33,
// Window.cpp:11-12); this GPU owns the tiles t with t % tile_stride == tile_phase (SURVEY.md
// 8(e)), handed out in the order of P.tile_order (tile_schedule_kernel: expensive tiles first); a
// unit is one warp-sized 8x4 pixel block, 32 units per tile, numbered tile-major so that
// consecutive units (taken by neighbouring warps) stay in one tile.
__device__ __forceinline__ bool unit_pixel(const SceneParams & P, unsigned unit, int lane, int & local_tile, int & px, int & py) {
	local_tile = int(__ldg(P.tile_order + (unit >> 5)));
	const int tile = P.tile_phase + local_tile * P.tile_stride;
	const int sub  = int(unit & 31);
	px = (tile % P.tiles_x) * 32 + (sub & 3) * 8 + (lane & 7);
	py = (tile / P.tiles_x) * 32 + (sub >> 2) * 4 + (lane >> 3);
	return px < P.width && py < P.height; // edge tiles are clipped, WorkerThread.cpp:60-61
}

// Where a pixel of row py goes when the frame is exchanged in bands (include/b200rt.h:
// b200rt_set_band_targets): into the image of the GPU that owns the row's band and, within band_halo
// rows of a band border, into the neighbour's too — cyclically, because the FXAA taps wrap at the
// frame border (GL_REPEAT, Window.cpp:48-50). Plain stores over NVLink peer memory: the exchange
// rides on the kernel's last instruction, every link carries 1/N of the frame, nobody gathers.
__device__ __forceinline__ void store_pixel_banded(const SceneParams & P, size_t p, int py, unsigned value) {
	const int last = P.band_count - 1;
	int b = py / P.band_rows;
	if (b > last) b = last;
	P.band_targets[b][p] = value;
	const int begin = b * P.band_rows, end = b == last ? P.height : begin + P.band_rows;
	if (py - begin < P.band_halo) P.band_targets[b == 0 ? last : b - 1][p] = value;
	if (end - py <= P.band_halo)  P.band_targets[b == last ? 0 : b + 1][p] = value;
}

// Clears the frame's counter block. A kernel rather than a memset: memsets are served by a copy
// engine, and with frames pipelined that engine is busy moving the previous frame's images to the
// host — the memset, and the render kernel behind it, waited 60 us per frame for it.
__global__ void reset_counters_kernel(FrameCounters * counters) {
	unsigned long long * words = reinterpret_cast<unsigned long long *>(counters);
	for (unsigned i = threadIdx.x; i < sizeof(FrameCounters) / sizeof(unsigned long long); i += blockDim.x) words[i] = 0ull;
}

// Persistent warps pull units off one device-side counter — the GPU form of the reference's
// InterlockedDecrement tile countdown (WorkerThread.cpp:53-65).
template<bool COUNT>
__global__ void __launch_bounds__(RT_BLOCK_THREADS, RT_MIN_BLOCKS)
render_kernel(const SceneParams P, unsigned * __restrict__ frame_buffer, float * __restrict__ colour_float, FrameCounters * __restrict__ counters) {
	const int lane = threadIdx.x & 31;
	const unsigned total_units = unsigned(P.tile_count_mine) * 32u;

	Counters cnt = { 0, 0, 0, 0, 0, 0, 0, 0 };
	unsigned primary = 0;

	if (threadIdx.x == 0) { // when the launch started: with last_done_ns the kernel times itself, and the pipelined frames need no timing events
		unsigned long long now;
		asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(now));
		atomicMax(&counters->first_start_ns, ~now);
	}

	for (;;) {
		unsigned unit = 0;
		if (lane == 0) unit = atomicAdd(&counters->next_unit, 1u);
		unit = __shfl_sync(0xffffffffu, unit, 0);
		if (unit >= total_units) break;

		const long long unit_begin = clock64();
		int local_tile, px, py;
		if (unit_pixel(P, unit, lane, local_tile, px, py)) {
			const Ray ray = primary_ray(P, px, py);
			primary++;
			const v3 colour = shade_pixel<COUNT>(P, ray, cnt);

			const size_t p = size_t(px) + size_t(P.width) * py;
			if (P.band_count == 0) frame_buffer[p] = plot_colour(colour);
			else store_pixel_banded(P, p, py, plot_colour(colour));
			if (colour_float) { colour_float[3 * p] = colour.x; colour_float[3 * p + 1] = colour.y; colour_float[3 * p + 2] = colour.z; }
		}
		__syncwarp();
		if (lane == 0) atomicAdd(P.tile_cost + local_tile, unsigned((clock64() - unit_begin) >> 6)); // what this tile cost, for the next frame's order
	}

	if (lane == 0) { // when did this warp run dry? (min / max over warps = how long the launch ran partly empty)
		unsigned long long now;
		asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(now));
		atomicMax(&counters->first_idle_ns, ~now);   // stored complemented: the earliest time wins an atomicMax, and an all-zero block is the initial state
		atomicMax(&counters->last_done_ns, now);
	}

	// one atomic per warp and counter
	const unsigned s_primary = warp_sum(primary), s_shadow = warp_sum(cnt.shadow), s_refl = warp_sum(cnt.reflection), s_refr = warp_sum(cnt.refraction);
	if (lane == 0) {
		atomicAdd(&counters->primary, (unsigned long long)s_primary);
		atomicAdd(&counters->shadow, (unsigned long long)s_shadow);
		atomicAdd(&counters->reflection, (unsigned long long)s_refl);
		atomicAdd(&counters->refraction, (unsigned long long)s_refr);
	}
	if (COUNT) {
		const unsigned s_nodes = warp_sum(cnt.nodes), s_tris = warp_sum(cnt.triangles), s_texels = warp_sum(cnt.texels);
		const unsigned s_accepted = warp_sum(cnt.accepted), s_instances = warp_sum(cnt.instances);
		if (lane == 0) {
			atomicAdd(&counters->accepted, (unsigned long long)s_accepted);
			atomicAdd(&counters->instances, (unsigned long long)s_instances);
			atomicAdd(&counters->nodes, (unsigned long long)s_nodes);
			atomicAdd(&counters->triangles, (unsigned long long)s_tris);
			atomicAdd(&counters->texels, (unsigned long long)s_texels);
		}
	}
}

// Next frame's hand-out order from this frame's measured tile costs: most expensive first (counting
// sort over 256 logarithmic cost classes, one block), so that the launch does not end with a few
// warps inside a costly tile and the rest of the GPU idle. Any order gives the same image; the
// reference's countdown uses index order (WorkerThread.cpp:53-58). Resets the costs.
// Measured (tools/tail_probe.py, bench frame): time with some warps already out of work 0.17 -> 0.08
// ms on one GPU's share of the frame. Splitting expensive tiles into 8-pixel items as well was tried
// and dropped: what bounds a costly unit is the serial recursion of its most expensive PIXEL
// (~0.3 ms), so narrower warps only multiplied the warp-time.
__global__ void __launch_bounds__(1024)
tile_schedule_kernel(unsigned * __restrict__ tile_cost, unsigned * __restrict__ tile_order, int tile_count) {
	__shared__ unsigned histogram[256];
	__shared__ unsigned offset[256];
	for (int i = threadIdx.x; i < 256; i += blockDim.x) histogram[i] = 0;
	__syncthreads();
	auto cost_class = [](unsigned cost) { // 8 classes per octave, expensive = low class number
		const int k = 255 - int(8.0f * log2f(float(cost) + 1.0f));
		return k < 0 ? 0 : k;
	};
	for (int t = threadIdx.x; t < tile_count; t += blockDim.x) atomicAdd(&histogram[cost_class(tile_cost[t])], 1u);
	__syncthreads();
	if (threadIdx.x < 32) { // exclusive prefix sum of the 256 classes by one warp: 8 classes per lane, then a shuffle scan over the lanes
		unsigned local[8], sum = 0;
#pragma unroll
		for (int k = 0; k < 8; k++) { local[k] = sum; sum += histogram[threadIdx.x * 8 + k]; }
		unsigned before = sum;
#pragma unroll
		for (int o = 1; o < 32; o <<= 1) { const unsigned v = __shfl_up_sync(0xffffffffu, before, o); if (int(threadIdx.x) >= o) before += v; }
		before -= sum;
#pragma unroll
		for (int k = 0; k < 8; k++) offset[threadIdx.x * 8 + k] = before + local[k];
	}
	__syncthreads();
	for (int t = threadIdx.x; t < tile_count; t += blockDim.x) tile_order[atomicAdd(&offset[cost_class(tile_cost[t])], 1u)] = unsigned(t);
	__syncthreads();
	for (int t = threadIdx.x; t < tile_count; t += blockDim.x) tile_cost[t] = 0;
}

// Closest-hit record of the primary ray per pixel: (instance, flattened triangle) and
// (distance, material id, u, v) — what the id-instrumented reference build dumps. Test output only.
__global__ void __launch_bounds__(RT_BLOCK_THREADS)
primary_hits_kernel(const SceneParams P, int * __restrict__ ids, float * __restrict__ hits) {
	const int px = blockIdx.x * 32 + (threadIdx.x & 31);
	const int py = blockIdx.y * 8 + (threadIdx.x >> 5);
	if (px >= P.width || py >= P.height) return;

	const Ray ray = primary_ray(P, px, py);
	Counters cnt = { 0, 0, 0, 0, 0, 0, 0, 0 };
	Candidate best;
	trace_closest<false>(P, ray.origin, ray.direction, best, cnt);

	const size_t p = size_t(px) + size_t(P.width) * py;
	const bool is_hit = best.instance != HIT_NONE;
	if (ids) { ids[2 * p] = is_hit ? best.instance : -1; ids[2 * p + 1] = is_hit ? best.primitive : -1; }
	if (hits) {
		float material = -1.0f, u = 0.0f, v = 0.0f;
		if (is_hit) {
			HitRecord hit;
			build_record(P, ray, best, hit);
			material = float(hit.material_id); u = hit.u; v = hit.v;
		}
		hits[4 * p] = best.t; hits[4 * p + 1] = material; hits[4 * p + 2] = u; hits[4 * p + 3] = v;
	}
}

// -------------------------------------------------------------------- texture ingest ---
// Texture::load's two host loops (Texture.cpp:62-117) on the device. Level 0: 8-bit sRGB -> linear
// float through the 256-entry table of gamma_to_linear(i * 0.00392156862f) (the reference converts
// texel by texel; 256 possible inputs give 256 possible outputs). Level k: 2x2 box filter of level
// k-1, summed in the reference's order ((c0 + c1) + c2) + c3, times 0.25. Texels are float4 (pad 0).
__global__ void __launch_bounds__(256)
texture_level0_kernel(const unsigned char * __restrict__ rgb8, const float * __restrict__ gamma_lut, float4 * __restrict__ out, long long texel_count) {
	__shared__ float lut[256];
	lut[threadIdx.x] = gamma_lut[threadIdx.x];
	__syncthreads();
	for (long long i = blockIdx.x * 256ll + threadIdx.x; i < texel_count; i += gridDim.x * 256ll)
		out[i] = make_float4(lut[rgb8[3 * i]], lut[rgb8[3 * i + 1]], lut[rgb8[3 * i + 2]], 0.0f);
}

__global__ void __launch_bounds__(256)
texture_mip_kernel(const float4 * __restrict__ previous, float4 * __restrict__ out, int width, int height, int previous_width) {
	const int i = blockIdx.x * 32 + (threadIdx.x & 31), j = blockIdx.y * 8 + (threadIdx.x >> 5);
	if (i >= width || j >= height) return;
	const float4 c0 = previous[size_t(2 * i)     + size_t(2 * j)     * previous_width];
	const float4 c1 = previous[size_t(2 * i + 1) + size_t(2 * j)     * previous_width];
	const float4 c2 = previous[size_t(2 * i)     + size_t(2 * j + 1) * previous_width];
	const float4 c3 = previous[size_t(2 * i + 1) + size_t(2 * j + 1) * previous_width];
	out[size_t(i) + size_t(j) * width] = make_float4((((c0.x + c1.x) + c2.x) + c3.x) * 0.25f, (((c0.y + c1.y) + c2.y) + c3.y) * 0.25f, (((c0.z + c1.z) + c2.z) + c3.z) * 0.25f, 0.0f);
}

// ------------------------------------------------------------------------------ FXAA ---
// Data/Shaders/fragment_fxaa.glsl:11-74 (+ fragment_identity.glsl) as a compute kernel. Sampler
// state of Window.cpp:48-50: NEAREST, GL_REPEAT, RGBA8; uv of pixel (px, py) = ((px+.5)/W,
// (py+.5)/H) (vertex.glsl:7-11). get_screen() = pow(texel, 1/2.2) of an 8-bit value = a 256-entry
// table, computed on the host with powf and kept in shared memory.
#define FXAA_TILE_W 64
#define FXAA_TILE_H 4

__device__ __forceinline__ v3 fxaa_screen(const unsigned * __restrict__ fb, const float * gamma, int W, int H, float fx, float fy) {
	// GL_REPEAT. The taps reach a few pixels beyond the frame at most: the integer division of positive_mod
	// (~20 instructions, 18 of them per pixel) is only paid by the pixels that do wrap
	int x = int(floorf(fx)), y = int(floorf(fy));
	if (x < 0 || x >= W) x = positive_mod(x, W);
	if (y < 0 || y >= H) y = positive_mod(y, H);
	const unsigned c = __ldg(fb + size_t(x) + size_t(W) * y);
	return V(gamma[(c >> 16) & 255], gamma[(c >> 8) & 255], gamma[c & 255]);
}
__device__ __forceinline__ float fxaa_luma(v3 c) { return (0.299f * c.x + 0.587f * c.y) + 0.114f * c.z; }
__device__ __forceinline__ unsigned fxaa_unorm8(float c) {
	if (!(c > 0.0f)) c = 0.0f;
	if (c > 1.0f) c = 1.0f;
	return unsigned(__float2int_rn(c * 255.0f));
}

__global__ void __launch_bounds__(FXAA_TILE_W * FXAA_TILE_H)
fxaa_kernel(const unsigned * __restrict__ fb, unsigned char * __restrict__ rgb8, const float * __restrict__ gamma_table, int W, int H, int enable_fxaa, int row_begin, int row_end) {
	__shared__ float gamma[256];
	__shared__ __align__(16) unsigned char staged[FXAA_TILE_H][FXAA_TILE_W * 3];

	const int tid = threadIdx.y * FXAA_TILE_W + threadIdx.x;
	gamma[tid] = gamma_table[tid]; // 256 threads, 256 entries
	__syncthreads();

	const int px = blockIdx.x * FXAA_TILE_W + threadIdx.x;
	const int py = row_begin + blockIdx.y * FXAA_TILE_H + threadIdx.y; // rows [row_begin, row_end): the whole frame, or this GPU's band

	if (px < W && py < row_end) {
		const float cx = float(px) + 0.5f, cy = float(py) + 0.5f;
		v3 out;
		if (!enable_fxaa) {
			out = fxaa_screen(fb, gamma, W, H, cx, cy);
		} else {
			// vec2(+-inv.x, +-inv_screen_size) takes .x for both components (fragment_fxaa.glsl:25-28):
			// +-1 px in x, +-H/W px in y
			const float diag_y = float(H) / float(W);
			const float luma_tl = fxaa_luma(fxaa_screen(fb, gamma, W, H, cx - 1.0f, cy - diag_y));
			const float luma_tr = fxaa_luma(fxaa_screen(fb, gamma, W, H, cx + 1.0f, cy - diag_y));
			const float luma_bl = fxaa_luma(fxaa_screen(fb, gamma, W, H, cx - 1.0f, cy + diag_y));
			const float luma_br = fxaa_luma(fxaa_screen(fb, gamma, W, H, cx + 1.0f, cy + diag_y));
			const float luma_m  = fxaa_luma(fxaa_screen(fb, gamma, W, H, cx, cy));

			const float luma_min = fminf(fminf(fminf(luma_tl, luma_tr), fminf(luma_bl, luma_br)), luma_m);
			const float luma_max = fmaxf(fmaxf(fmaxf(luma_tl, luma_tr), fmaxf(luma_bl, luma_br)), luma_m);

			float dir_x = (luma_bl + luma_br) - (luma_tl + luma_tr);
			float dir_y = (luma_tl + luma_bl) - (luma_tr + luma_br);

			const float reduce = fmaxf(1.0f / 128.0f, (luma_tl + luma_tr + luma_bl + luma_br) * 0.25f * (1.0f / 8.0f));
			const float adjust = 1.0f / (fminf(fabsf(dir_x), fabsf(dir_y)) + reduce);

			dir_x = fminf(fmaxf(dir_x * adjust, -8.0f), 8.0f); // in pixels: * inv_screen_size * screen_size
			dir_y = fminf(fmaxf(dir_y * adjust, -8.0f), 8.0f);

			const float k1 = 1.0f / 3.0f - 0.5f, k2 = 2.0f / 3.0f - 0.5f;
			const v3 a0 = fxaa_screen(fb, gamma, W, H, cx + dir_x * k1, cy + dir_y * k1);
			const v3 a1 = fxaa_screen(fb, gamma, W, H, cx + dir_x * k2, cy + dir_y * k2);
			const v3 b0 = fxaa_screen(fb, gamma, W, H, cx + dir_x * -0.5f, cy + dir_y * -0.5f);
			const v3 b1 = fxaa_screen(fb, gamma, W, H, cx + dir_x * 0.5f,  cy + dir_y * 0.5f);

			const v3 result_a = (a0 + a1) * 0.5f;
			const v3 result_b = (b0 + b1) * 0.5f;
			const v3 result   = (result_a + result_b) * 0.5f;

			const float luma_result = fxaa_luma(result);
			out = (luma_result < luma_min || luma_result > luma_max) ? result_a : result;
		}
		staged[threadIdx.y][threadIdx.x * 3]     = (unsigned char)fxaa_unorm8(out.x);
		staged[threadIdx.y][threadIdx.x * 3 + 1] = (unsigned char)fxaa_unorm8(out.y);
		staged[threadIdx.y][threadIdx.x * 3 + 2] = (unsigned char)fxaa_unorm8(out.z);
	}
	__syncthreads();

	// 64 pixels x 3 bytes = 48 aligned 32-bit words per row: coalesced word stores
	const int row_px0 = blockIdx.x * FXAA_TILE_W;
	if (py < row_end && threadIdx.x < 48) {
		const int valid_bytes = min(FXAA_TILE_W, W - row_px0) * 3;
		const size_t row_base = (size_t(py) * W + row_px0) * 3; // multiple of 4 when W % 4 == 0 or row_px0 * 3 aligned
		const int b = threadIdx.x * 4;
		if ((row_base & 3) == 0 && b + 4 <= valid_bytes) {
			*reinterpret_cast<unsigned *>(rgb8 + row_base + b) = *reinterpret_cast<const unsigned *>(&staged[threadIdx.y][b]);
		} else {
			for (int k = b; k < b + 4 && k < valid_bytes; k++) rgb8[row_base + k] = staged[threadIdx.y][k];
		}
	}
}

} // namespace rt
