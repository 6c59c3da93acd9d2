// kernels.cuh — __global__ entry points.
#pragma once
#include "render.cuh"

namespace RT_NAMESPACE {

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

// Clears the NEXT frame's counter block (the context keeps a ring of four: b200rt.cu) and arms its count of
// outstanding work. Runs inside tile_schedule_kernel, the last kernel of every frame, so a frame costs no launch for it.
__device__ __forceinline__ void reset_counters(FrameCounters * counters, unsigned total_units) {
	unsigned long long * words = reinterpret_cast<unsigned long long *>(counters);
	for (unsigned i = threadIdx.x; i < sizeof(FrameCounters) / sizeof(unsigned long long); i += blockDim.x) words[i] = 0ull;
	__syncthreads();
	if (threadIdx.x == 0) counters->q_outstanding = int(total_units);
}

__device__ __forceinline__ void flush_ray_counters(FrameCounters * __restrict__ counters, const Counters & cnt, unsigned primary, bool count) {
	const int lane = threadIdx.x & 31;
	// one atomic per warp and counter
	const unsigned s_primary = warp_sum(primary), s_shadow = warp_sum(cnt.shadow), s_refl = warp_sum(cnt.reflection), s_refr = warp_sum(cnt.refraction);
	if (lane == 0) {
		atomicAdd(&counters->primary, (unsigned long long)s_primary);
		atomicAdd(&counters->shadow, (unsigned long long)s_shadow);
		atomicAdd(&counters->reflection, (unsigned long long)s_refl);
		atomicAdd(&counters->refraction, (unsigned long long)s_refr);
	}
	if (count) {
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

// Persistent warps pull units off one device-side counter — the GPU form of the reference's
// InterlockedDecrement tile countdown (WorkerThread.cpp:53-65). One thread walks a pixel's whole ray tree.
template<bool COUNT, int MAXB, int BSTACK>
__global__ void __launch_bounds__(RT_BLOCK_THREADS, RT_MIN_BLOCKS)
render_kernel(const SceneParams P, unsigned * __restrict__ frame_buffer, float * __restrict__ colour_float, FrameCounters * __restrict__ counters) {
	const int lane = threadIdx.x & 31;
	const unsigned total_units = unsigned(P.tile_count_mine) * 32u;
	const PixelOut out = { frame_buffer, colour_float };

	Counters cnt = { 0, 0, 0, 0, 0, 0, 0, 0 };
	unsigned primary = 0;

	if (threadIdx.x == 0) atomicMax(&counters->first_start_ns, ~global_timer_ns()); // when the launch started: with last_done_ns the kernel times itself, and the pipelined frames need no timing events

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
			float distance;
			write_pixel(P, out, px, py, shade_subtree<COUNT, MAXB, BSTACK>(P, ray, P.bounces, cnt, distance));
		}
		__syncwarp();
		if (lane == 0) atomicAdd(P.tile_cost + local_tile, unsigned((clock64() - unit_begin) >> 6)); // what this tile cost, for the next frame's order
	}

	if (lane == 0) { // when did this warp run dry? (min / max over warps = how long the launch ran partly empty)
		const unsigned long long now = global_timer_ns();
		atomicMax(&counters->first_idle_ns, ~now);   // stored complemented: the earliest time wins an atomicMax, and an all-zero block is the initial state
		atomicMax(&counters->last_done_ns, now);
	}
	flush_ray_counters(counters, cnt, primary, COUNT);
}

// Warp-collective counter update: one atomicAdd per warp.
__device__ __forceinline__ void warp_add(int * counter, int n) {
#pragma unroll
	for (int o = 16; o > 0; o >>= 1) n += __shfl_xor_sync(0xffffffffu, n, o);
	if ((threadIdx.x & 31) == 0 && n != 0) atomicAdd(counter, n);
}

// The same frame with the deep-pixel queue (render.cuh). Every ray — a pixel's primary ray or an item popped from the
// bounce queue — is the same depth-first walk (shade_subtree_body<QUEUE>), whose top node may be exported to the queue.
// A warp works on ONE kind at a time: bounce items first (the deep pixels' rays are the frame's critical path), then
// tile units.
//
// Items are claimed by TICKET: one atomicAdd on the queue's head hands a warp the next n slots — it never fails and is
// never retried. (The first version claimed with compare-and-swap on the head: with thousands of warps contending, one
// wins per L2 round trip and the frame's million items went through at ~20 per microsecond — 150 ms instead of 1;
// profiles/r02_variants_call1.jsonl, r02_variants_call2.jsonl.) A ticket may name a slot that is not published yet, or
// never will be: a lane keeps its ticket across iterations until the slot's ready word shows this frame's serial
// number, and meanwhile the warp takes tile units; tickets die when the frame's count of outstanding work (tile units +
// items not finished) reaches zero. A warp with no ticket that finds nothing looks three more times (2 / 4 / 8 us
// apart) and leaves: whoever publishes an item looks at the queue again right afterwards, so nothing is orphaned, and no
// warp polls for long.
//
// Why this cannot stall. Slots are handed out in order, so the tickets in flight always cover [published .. head): an
// item published below the head has a holder — a warp never leaves while a lane holds a ticket, and looks at its
// tickets' ready words on every turn —, an item at or above the head is drawn by its own publisher at the latest, whose
// next turn starts with a look at the queue. There is ONE queue: the first version of this kernel had a second one for
// shadow rays, and a warp that held early tickets of one queue never looked at the other — with every other warp gone,
// items of the other queue stayed unclaimed and the launch never ended (one run in a few hundred on SCENE_DYNAMIC);
// shadow rays are traced by the thread that shades the node, which was the faster setting anyway (5.5 against 6.9 ms on
// configs[3]). As a last line of defence a warp that has waited on its tickets for RT_QUEUE_WATCHDOG_NS raises
// FrameCounters::q_abort, every warp leaves, and the host renders the frame again depth-first (b200rt.cu).
template<int MAXB, int BSTACK>
__global__ void __launch_bounds__(RT_BLOCK_THREADS, RT_MIN_BLOCKS)
render_queue_kernel(const SceneParams P, unsigned * __restrict__ frame_buffer, float * __restrict__ colour_float, FrameCounters * counters) {
	const int lane = threadIdx.x & 31;
	const unsigned lanes_below = (1u << lane) - 1u;
	const unsigned total_units = unsigned(P.tile_count_mine) * 32u;
	const PixelOut out = { frame_buffer, colour_float };
	enum { WORK_NONE = 0, WORK_UNIT = 1, WORK_BOUNCE = 2, WORK_WAIT = 4, WORK_VOID = 5 };
	const unsigned NO_TICKET = 0xffffffffu;

	Counters cnt = { 0, 0, 0, 0, 0, 0, 0, 0 };
	unsigned primary = 0;
	int idle_looks = 0;
	bool units_left = total_units > 0;   // this warp's view; once false it stays false (the cursor only grows)
	unsigned ticket = NO_TICKET;         // this lane's claimed slot that is not published yet
	unsigned long long waiting_since = 0ull;   // lane 0: when the warp began to wait on tickets with nothing else to do

	if (threadIdx.x == 0) atomicMax(&counters->first_start_ns, ~global_timer_ns());

	for (;;) {
		// ---- lane 0 chooses the warp's next piece of work ------------------------------------------------
		const unsigned holding = __ballot_sync(0xffffffffu, ticket != NO_TICKET);
		int kind = WORK_NONE; unsigned first = 0, n = 0;
		if (lane == 0) {
			const unsigned free_lanes = 32u - unsigned(__popc(holding));
			const unsigned hb = peek(&counters->q_bounce_head), rb = min(peek(&counters->q_bounce_reserved), P.q_bounce_capacity), ab = rb > hb ? rb - hb : 0u;
			if (peek(&counters->q_abort) != 0u) kind = WORK_VOID;
			else if (holding || ab >= unsigned(P.queue_item_batch) || (ab > 0u && !units_left)) kind = WORK_BOUNCE;
			if (kind == WORK_BOUNCE && ab > 0u && free_lanes > 0u) { n = min(ab, free_lanes); first = atomicAdd(&counters->q_bounce_head, n); }
			if (kind == WORK_NONE && units_left) {
				first = atomicAdd(&counters->next_unit, 1u);
				if (first < total_units) kind = WORK_UNIT; else units_left = false;
			}
			if (kind == WORK_NONE && (units_left || ab > 0u || (idle_looks < P.queue_idle_looks && peek(&counters->q_outstanding) > 0))) kind = WORK_WAIT;
		}
		kind = __shfl_sync(0xffffffffu, kind, 0); first = __shfl_sync(0xffffffffu, first, 0); n = __shfl_sync(0xffffffffu, n, 0);
		if (kind == WORK_NONE || kind == WORK_VOID) break;
		if (kind == WORK_WAIT) {
			if (lane == 0 && idle_looks == 0) atomicMax(&counters->first_idle_ns, ~global_timer_ns());
			__nanosleep(2000u << min(idle_looks, P.queue_idle_shift));
			idle_looks++;
			continue;
		}

		// ---- items: new tickets go to the lanes that hold none; whose slot is published? --------------------
		bool active = false;
		if (kind == WORK_BOUNCE) {
			const unsigned rank = unsigned(__popc(~holding & lanes_below));
			if (ticket == NO_TICKET && rank < n) ticket = first + rank;
			if (ticket != NO_TICKET && ticket >= P.q_bounce_capacity) ticket = NO_TICKET;   // beyond the pool: that item was never queued
			if (ticket != NO_TICKET) {
				for (int look = 0; look < 3 && !active; look++) active = __ldcg(P.q_bounce_ready + ticket) == P.frame_serial;   // reserved a moment before it is written
				if (active) __threadfence();
			}
			if (__ballot_sync(0xffffffffu, active) == 0u) { // nothing published yet: the tickets wait, the warp does something else
				kind = WORK_NONE;
				if (lane == 0) {
					if (peek(&counters->q_outstanding) <= 0) kind = WORK_VOID;   // the frame is done: the tickets are void
					else if (units_left) {
						first = atomicAdd(&counters->next_unit, 1u);
						if (first < total_units) kind = WORK_UNIT; else units_left = false;
					}
					if (kind == WORK_NONE) { // nothing but waiting: for how long already?
						const unsigned long long now = global_timer_ns();
						if (waiting_since == 0ull) waiting_since = now;
						else if (now - waiting_since > RT_QUEUE_WATCHDOG_NS) { atomicOr(&counters->q_abort, 1u); kind = WORK_VOID; }
					}
				}
				kind = __shfl_sync(0xffffffffu, kind, 0); first = __shfl_sync(0xffffffffu, first, 0);
				if (kind == WORK_VOID) { ticket = NO_TICKET; continue; }
				if (kind == WORK_NONE) { __nanosleep(500); continue; }
			}
		}
		idle_looks = 0;
		waiting_since = 0ull;

		int settled = 0;   // the item (or the tile unit) this lane finishes: comes off the frame's outstanding work
		{ // a walk: a pixel's (WORK_UNIT) or a popped bounce item's
			const long long unit_begin = clock64();
			Ray ray; int bounces_left = 0, local_tile = 0;
			QueueContext q = { counters, 0, 0 };
			if (kind == WORK_UNIT) {
				int px, py;
				active = unit_pixel(P, first, lane, local_tile, px, py);
				if (active) { ray = primary_ray(P, px, py); primary++; bounces_left = P.bounces; q.parent = -(px + P.width * py + 1); }
			} else if (active) {
				read_bounce_item(P, ticket, ray, q.parent, q.which, bounces_left);
				ticket = NO_TICKET;
				if (q.parent == QUEUE_NULL_ITEM) active = false; else settled = 1;
			}
#ifdef RT_QUEUE_TRACE   // development: when was the last walk of each level picked up / finished (us since the launch began), how many
			const int trace_level = kind == WORK_UNIT ? 9 : bounces_left;
			if (kind != WORK_UNIT) { // how full are the warps that walk items? [10] iterations, [11] lanes with an item, [12..16] iterations by fill (1, 2-3, 4-7, 8-15, 16-32 lanes)
				const unsigned lanes = unsigned(__popc(__ballot_sync(0xffffffffu, active)));
				if (lane == 0 && lanes > 0u) { atomicAdd(&counters->q_pad0[10], 1u); atomicAdd(&counters->q_pad0[11], lanes); atomicAdd(&counters->q_pad0[12 + (lanes >= 16u ? 4 : lanes >= 8u ? 3 : lanes >= 4u ? 2 : lanes >= 2u ? 1 : 0)], 1u); }
			}
			if (active) {
				const unsigned now_us = unsigned((global_timer_ns() - ~counters->first_start_ns) >> 10);
				atomicMax(&counters->q_pad0[trace_level], now_us);
				atomicAdd(&counters->q_pad0[18 + (trace_level > 9 ? 9 : trace_level)], 1u);
			}
#endif
			bool exported = false;
			if (active) {
				float distance;
				const v3 colour = shade_subtree_body<false, MAXB, BSTACK, true>(P, ray, bounces_left, cnt, distance, &q, exported);
#ifdef RT_QUEUE_TRACE
				atomicMax(&counters->q_pad1[trace_level], unsigned((global_timer_ns() - ~counters->first_start_ns) >> 10));
#endif
				// the walk's value goes to its parent: the pixel, or the record the walk last exported (the walk went on below it)
				if (q.parent < 0) { const int p = -(q.parent + 1); write_pixel(P, out, p % P.width, p / P.width, colour); }
				else deliver(P, out, q.parent, q.which, colour, distance);
			}
			if (kind == WORK_UNIT) {
				// What this tile cost, for the next frame's hand-out order. The subtrees this unit exported are walked by
				// others, later: their time is not in this clock, and without it every tile looks cheap and the tiles whose
				// pixels spawn the frame's longest chains are handed out anywhere — at the end, too, and the launch then
				// waits for chains that started late. An exported pixel is booked at what a deep pixel costs depth-first.
				const unsigned exports = unsigned(__popc(__ballot_sync(0xffffffffu, exported)));
				if (lane == 0) { atomicAdd(P.tile_cost + local_tile, unsigned((clock64() - unit_begin) >> 6) + exports * RT_QUEUE_EXPORT_COST); settled = 1; }   // the unit itself
			}
		}
		warp_add(&counters->q_outstanding, -settled);   // what a walk queued was added before it became visible (export_node)
	}

	if (lane == 0) {
		const unsigned long long now = global_timer_ns();
		if (idle_looks == 0) atomicMax(&counters->first_idle_ns, ~now);
		atomicMax(&counters->last_done_ns, now);
	}
	flush_ray_counters(counters, cnt, primary, false);
}

// Next frame's hand-out order from this frame's measured tile costs: most expensive first (counting
// sort over 256 logarithmic cost classes, one block), so that the launch does not end with a few
// warps inside a costly tile and the rest of the GPU idle. Any order gives the same image; the
// reference's countdown uses index order (WorkerThread.cpp:53-58). Resets the costs, and the NEXT frame's counter block.
// Measured (tools/tail_probe.py, bench frame): time with some warps already out of work 0.17 -> 0.08
// ms on one GPU's share of the frame. Splitting expensive tiles into 8-pixel items as well was tried
// and dropped: what bounds a costly unit is the serial recursion of its most expensive PIXEL
// (~0.3 ms), so narrower warps only multiplied the warp-time.
__global__ void __launch_bounds__(1024)
tile_schedule_kernel(unsigned * __restrict__ tile_cost, unsigned * __restrict__ tile_order, int tile_count, FrameCounters * __restrict__ next_counters) {
	__shared__ unsigned histogram[256];
	reset_counters(next_counters, unsigned(tile_count) * 32u);
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
	trace_closest<false, 64>(P, ray.origin, ray.direction, best, cnt);

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

// ------------------------------------------------------------------------ device barrier ---
// b200rt_barrier (include/b200rt.h): thread q tells rank q "rank `rank` has reached barrier `epoch`" with a store into
// q's flag block (peer memory over NVLink) and waits until this rank's own block says the same of rank q. The fences
// order the earlier stores of this GPU (the render kernel's pixels in the peers' images) before the flag, and the
// peers' flags before whatever this stream runs next.
struct BarrierPeers { unsigned * block[B200RT_MAX_BANDS]; };
__global__ void __launch_bounds__(32)
barrier_kernel(BarrierPeers peers, int count, int rank, unsigned epoch) {
	const int q = threadIdx.x;
	if (q < count) {
		__threadfence_system();
		*reinterpret_cast<volatile unsigned *>(peers.block[q] + rank) = epoch;
		while (*reinterpret_cast<volatile unsigned *>(peers.block[rank] + q) < epoch) __nanosleep(100);
		__threadfence_system();
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

// FAST arithmetic, RT_FAST_TEXEL: the float4 pool re-stored as four halves per texel (round to nearest even)
__global__ void __launch_bounds__(256)
texel_halves_kernel(const float4 * __restrict__ texels, uint2 * __restrict__ out, long long texel_count) {
	for (long long i = blockIdx.x * 256ll + threadIdx.x; i < texel_count; i += gridDim.x * 256ll) {
		const float4 t = texels[i];
#ifdef RT_EMULATE
		const unsigned r = emu_float_to_half(t.x), g = emu_float_to_half(t.y), b = emu_float_to_half(t.z);
#elif defined(RT_FAST_TEXEL)
		const unsigned r = __half_as_ushort(__float2half_rn(t.x)), g = __half_as_ushort(__float2half_rn(t.y)), b = __half_as_ushort(__float2half_rn(t.z));
#else
		const unsigned r = 0, g = 0, b = 0; (void) t;
#endif
		out[i] = make_uint2(r | (g << 16), b);
	}
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

// texture2D(screen, uv): texel (floor(u * W) mod W, floor(v * H) mod H) (OpenGL 4.2 spec 3.9.11: NEAREST, GL_REPEAT)
__device__ __forceinline__ v3 fxaa_screen(const unsigned * __restrict__ fb, const float * gamma, int W, int H, float u, float v) {
	// The taps reach a few pixels beyond the frame at most: the integer division of positive_mod
	// (~20 instructions, 18 of them per pixel) is only paid by the pixels that do wrap
	int x = int(floorf(u * float(W))), y = int(floorf(v * float(H)));
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
		// The shader's own coordinates: in_uv = ((px + .5) / W, (py + .5) / H) (vertex.glsl:7-11), inv_screen_size = (1 / W, 1 / H)
		// (Window.cpp:57); taps are formed in uv space like the GLSL's and only fxaa_screen() goes back to texels.
		const float inv_x = 1.0f / float(W), inv_y = 1.0f / float(H);
		const float in_u = (float(px) + 0.5f) / float(W), in_v = (float(py) + 0.5f) / float(H);
		v3 out;
		if (!enable_fxaa) {
			out = fxaa_screen(fb, gamma, W, H, in_u, in_v);
		} else {
			// vec2(+-inv.x, +-inv_screen_size) takes .x for both components (fragment_fxaa.glsl:25-28)
			const float luma_tl = fxaa_luma(fxaa_screen(fb, gamma, W, H, in_u - inv_x, in_v - inv_x));
			const float luma_tr = fxaa_luma(fxaa_screen(fb, gamma, W, H, in_u + inv_x, in_v - inv_x));
			const float luma_bl = fxaa_luma(fxaa_screen(fb, gamma, W, H, in_u - inv_x, in_v + inv_x));
			const float luma_br = fxaa_luma(fxaa_screen(fb, gamma, W, H, in_u + inv_x, in_v + inv_x));
			const float luma_m  = fxaa_luma(fxaa_screen(fb, gamma, W, H, in_u, in_v));

			const float luma_min = fminf(fminf(fminf(luma_tl, luma_tr), fminf(luma_bl, luma_br)), luma_m);
			const float luma_max = fmaxf(fmaxf(fmaxf(luma_tl, luma_tr), fmaxf(luma_bl, luma_br)), luma_m);

			float dir_x = (luma_bl + luma_br) - (luma_tl + luma_tr);
			float dir_y = (luma_tl + luma_bl) - (luma_tr + luma_br);

			const float reduce = fmaxf(1.0f / 128.0f, (luma_tl + luma_tr + luma_bl + luma_br) * 0.25f * (1.0f / 8.0f));
			const float adjust = 1.0f / (fminf(fabsf(dir_x), fabsf(dir_y)) + reduce);

			dir_x = fminf(fmaxf(dir_x * adjust, -8.0f), 8.0f) * inv_x; // :46
			dir_y = fminf(fmaxf(dir_y * adjust, -8.0f), 8.0f) * inv_y;

			const float k1 = 1.0f / 3.0f - 0.5f, k2 = 2.0f / 3.0f - 0.5f;
			const v3 a0 = fxaa_screen(fb, gamma, W, H, in_u + dir_x * k1, in_v + dir_y * k1);
			const v3 a1 = fxaa_screen(fb, gamma, W, H, in_u + dir_x * k2, in_v + dir_y * k2);
			const v3 b0 = fxaa_screen(fb, gamma, W, H, in_u + dir_x * -0.5f, in_v + dir_y * -0.5f);
			const v3 b1 = fxaa_screen(fb, gamma, W, H, in_u + dir_x * 0.5f,  in_v + dir_y * 0.5f);

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

} // namespace RT_NAMESPACE
