// b200rt.cu — the C ABI of include/b200rt.h over the sm_100a kernels (kernels.cuh).
//
// There is no CPU fallback in this library: every entry point either runs the CUDA path or
// returns an error.
#include <cmath>
#include <cstdio>
#include <cstring>
#include <mutex>
#include <string>
#include <vector>

#include "kernel_table.h"      // rt_types.h (cuda_runtime.h, include/b200rt.h) + the launchers of the kernel units
#include "threaded_nodes.h"

using namespace rt;

namespace {

thread_local std::string g_create_error;   // errors of calls that have no context to keep them in (b200rt_create, b200rt_host_alloc)

struct DeviceBuffer {
	void * ptr = nullptr; size_t bytes = 0;
	cudaError_t ensure(size_t n) {
		if (n <= bytes) return cudaSuccess;
		if (ptr) cudaFree(ptr);
		ptr = nullptr; bytes = 0;
		cudaError_t e = cudaMalloc(&ptr, n ? n : 16);
		if (e == cudaSuccess) bytes = n;
		return e;
	}
	void release() { if (ptr) cudaFree(ptr); ptr = nullptr; bytes = 0; }
};

} // namespace

struct b200rt_ctx {
	b200rt_config config;
	std::string   error;
	cudaStream_t  stream = nullptr;      // stream in use (the context's own, or a caller's: b200rt_set_stream)
	cudaStream_t  own_stream = nullptr;
	cudaEvent_t   ev[5] = { nullptr, nullptr, nullptr, nullptr, nullptr }; // total begin, render begin, render end / fxaa begin, fxaa end, total end
	const KernelTable * kernels = nullptr;   // the kernel unit of config.arithmetic

	// static scene in HBM
	DeviceBuffer nodes, shadow_nodes, hot, cold, materials, textures, texels, sky, gamma;
	std::vector<int> shadow_node_offset;  // per BLAS: first node of its threaded copy (build_threaded_nodes)
	std::vector<int> records_per_copy;    // per BLAS: records in one threaded copy (RT_OCTANT_THREADED)
	std::vector<int> wide_node_offset;    // per BLAS: first record of its four-wide copy, -1 if it has none (RT_WIDE_SHADOW)
	DeviceBuffer wide_nodes;
	DeviceBuffer hot_fast, texels_half;   // FAST unit: precomputed triangle transforms, RGBA16F texel pool
	int sky_size = 0;
	bool have_static = false;
	std::vector<b200rt_blas> blas;

	// per-frame data: one pinned staging block + one device block, sub-allocated
	DeviceBuffer frame_block[2];          // alternate with the staging halves; uploaded on upload_stream, ahead of the kernels
	cudaStream_t upload_stream = nullptr;
	cudaEvent_t consumed[2] = { nullptr, nullptr };                // the last kernel that reads each device block has been issued up to here
	int block_in_use = 0;
	void * frame_staging = nullptr; size_t frame_staging_bytes = 0; // two halves, used alternately
	cudaEvent_t staged[2] = { nullptr, nullptr };                  // the host->device copy out of each half
	int staging_half = 0;

	// outputs
	DeviceBuffer frame_buffer, presented, colour_float, ids, hits, counters, l2_scratch;
	// `counters` is a ring of COUNTER_BLOCKS FrameCounters: frame k counts in block k % 4 and its last kernel
	// (tile_schedule_kernel) clears block (k + 1) % 4 — no memset, no extra launch, and a block is cleared three frames
	// after the frame that used it, when its copy to the host (at most two frames are in flight) is long done
	unsigned frame_number = 0;
	// deep-pixel queue (render.cuh): pools, allocated on first use
	DeviceBuffer q_nodes, q_bounce, q_bounce_ready;
	int queue_inline_below = 0, queue_idle_looks = 3, queue_idle_shift = 2, queue_item_batch = 32, queue_keep_refraction = 2;
	bool queue_stalled = false;           // the queue kernel's watchdog fired once: this context renders depth-first from then on
	bool test_stall = false;              // B200RT_TEST_QUEUE_STALL (tests): the first queued frame is marked as given up, to exercise the recovery
	int blas_stack_entries = 1, tlas_stack_entries = 1;
	DeviceBuffer tile_cost, tile_order; // schedule of the render kernel (kernels.cuh: tile_schedule_kernel)
	int tile_count_mine = 0;
	void * pinned_out = nullptr; size_t pinned_out_bytes = 0;
	FrameCounters * pinned_counters = nullptr;

	// pipelined frames (b200rt_submit_frame / b200rt_wait_frame): two slots, each with its own images
	// and counters, so that frame n+1 renders while frame n is still on its way to the host
	struct Slot {
		DeviceBuffer frame_buffer, presented;
		FrameCounters * pinned_counters = nullptr;
		cudaEvent_t render_end = nullptr, fxaa_end = nullptr, copied = nullptr; // ordering only, no timestamps: a timing event costs the stream ~20 us, four of them 7 % of a pipelined frame
		bool busy = false;
	} slots[2];
	cudaStream_t copy_stream = nullptr;
	unsigned submitted = 0, waited = 0;

	uint64_t band_targets[B200RT_MAX_BANDS] = {};   // b200rt_set_band_targets
	unsigned * barrier_blocks[B200RT_MAX_BANDS] = {}; // b200rt_barrier_setup
	int barrier_count = 0, barrier_rank = 0;
	unsigned barrier_epoch = 0;
	int band_count = 0, band_rows = 0, band_halo = 0;

	SceneParams params;      // last frame's parameters (for b200rt_time_resident)
	bool have_frame = false;
	int  num_sms = 0, blocks_per_sm[2][2] = { { 0, 0 }, { 0, 0 } };   // resident blocks per SM of the render kernel, [counting][queue]
	unsigned launch_serial = 0;   // SceneParams::frame_serial of the last queue launch

	// config.deep_queue == B200RT_QUEUE_AUTO. Which of the two walks is faster depends on the view (a hall of opaque
	// walls: depth-first by 45 %; glass filling the frame: queued by 30 %), so the context measures: after WARM frames
	// (the tile order has to settle on measured costs first) PAIRS frames alternate between the two candidates, each
	// frame's own %globaltimer stamps (FrameCounters::first_start_ns / last_done_ns) come back through a 24-byte copy
	// behind the frame — no synchronisation, no timing event —, the candidate with the smaller minimum is kept, and
	// every PERIOD frames the comparison is repeated. Both walks produce the same bits, so switching is invisible.
	struct AutoQueue {
		enum { WARM = 2, PAIRS = 3, SLOTS = 4, PERIOD = 4096 };
		int   frame = 0;            // frames since the comparison (re)started
		int   choice = 0;           // 0 = depth-first, 1 = queued
		bool  decided = false;
		float best_ms[2] = { 1e30f, 1e30f };
		int   samples[2] = { 0, 0 };
		int   issued = -1;          // candidate of the frame being issued if it is one of the timed ones (launch_render -> launch_schedule)
		struct Probe { cudaEvent_t copied = nullptr; int candidate = -1; } probes[SLOTS];
		unsigned long long * stamps = nullptr;   // pinned, SLOTS x { first_idle_ns, last_done_ns, first_start_ns }
	} auto_queue;
};

enum { COUNTER_BLOCKS = 4 };
static FrameCounters * counter_block(b200rt_ctx * ctx, unsigned frame) { return static_cast<FrameCounters *>(ctx->counters.ptr) + frame % COUNTER_BLOCKS; }

#define CHECK_CUDA(ctx, call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) { (ctx)->error = std::string(#call) + ": " + cudaGetErrorString(e_); return -1; } } while (0)

static int fail(b200rt_ctx * ctx, const std::string & message) { ctx->error = message; return -1; }

extern "C" const char * b200rt_last_error(const b200rt_ctx * ctx) { return ctx ? ctx->error.c_str() : g_create_error.c_str(); }

extern "C" uint64_t b200rt_stream(const b200rt_ctx * ctx) { return ctx ? uint64_t(reinterpret_cast<uintptr_t>(ctx->stream)) : 0; }

extern "C" int b200rt_set_stream(b200rt_ctx * ctx, uint64_t stream) {
	if (!ctx) return -1;
	cudaStreamSynchronize(ctx->stream);
	ctx->stream = reinterpret_cast<cudaStream_t>(uintptr_t(stream)); // 0 is a stream too: CUDA's default stream
	return 0;
}

extern "C" int b200rt_reset_stream(b200rt_ctx * ctx) {
	if (!ctx) return -1;
	cudaStreamSynchronize(ctx->stream);
	ctx->stream = ctx->own_stream;
	return 0;
}

extern "C" int b200rt_create(const b200rt_config * config, b200rt_ctx ** out) {
	if (!config || !out) { g_create_error = "null argument"; return -1; }
	*out = nullptr;
	if (config->width <= 0 || config->height <= 0) { g_create_error = "bad resolution"; return -1; }
	if (config->bounces < 0 || config->bounces > RT_MAX_BOUNCES) { g_create_error = "bounces must be in 0..8"; return -1; }
	if (config->mipmap_filter < 0 || config->mipmap_filter > 2) { g_create_error = "bad mipmap_filter"; return -1; }
	if (config->texture_sample_mode < 0 || config->texture_sample_mode > 2) { g_create_error = "bad texture_sample_mode"; return -1; }
	if (config->traversal_naive < 0 || config->traversal_naive > 1) { g_create_error = "bad traversal_naive"; return -1; }
	if (config->tile_stride < 1 || config->tile_phase < 0 || config->tile_phase >= config->tile_stride) { g_create_error = "bad tile partition"; return -1; }
	if (config->arithmetic < 0 || config->arithmetic > 1) { g_create_error = "bad arithmetic mode"; return -1; }
	if (config->deep_queue < 0 || config->deep_queue > B200RT_QUEUE_AUTO) { g_create_error = "bad deep_queue"; return -1; }
#ifndef RT_HAVE_FAST
	if (config->arithmetic == B200RT_ARITHMETIC_FAST) { g_create_error = "this build of the library carries no FAST kernels"; return -1; }
#endif

	int device_count = 0;
	cudaError_t e = cudaGetDeviceCount(&device_count);
	if (e != cudaSuccess || device_count == 0) { g_create_error = std::string("no CUDA device: ") + cudaGetErrorString(e); return -1; }
	if (config->device < 0 || config->device >= device_count) { g_create_error = "bad device ordinal"; return -1; }
	if ((e = cudaSetDevice(config->device)) != cudaSuccess) { g_create_error = cudaGetErrorString(e); return -1; }

	b200rt_ctx * ctx = new b200rt_ctx();
	ctx->config = *config;
	memset(&ctx->params, 0, sizeof(ctx->params));
#ifdef RT_HAVE_FAST
	ctx->kernels = config->arithmetic == B200RT_ARITHMETIC_FAST ? &kernels_fast() : &kernels_exact();
#else
	ctx->kernels = &kernels_exact();
#endif
	if (const char * text = getenv("B200RT_QUEUE_INLINE_BELOW")) ctx->queue_inline_below = atoi(text);   // tuning knobs (tools/): see SceneParams
	if (const char * text = getenv("B200RT_TEST_QUEUE_STALL")) ctx->test_stall = atoi(text) != 0;
	if (const char * text = getenv("B200RT_QUEUE_IDLE_LOOKS")) ctx->queue_idle_looks = atoi(text);
	if (const char * text = getenv("B200RT_QUEUE_KEEP_REFRACTION")) ctx->queue_keep_refraction = std::max(0, std::min(atoi(text), 2));
	if (const char * text = getenv("B200RT_QUEUE_ITEM_BATCH")) ctx->queue_item_batch = std::max(1, atoi(text));
	if (const char * text = getenv("B200RT_QUEUE_IDLE_SHIFT")) ctx->queue_idle_shift = std::max(0, std::min(atoi(text), 8));

	auto bail = [&](const char * what, cudaError_t err) { g_create_error = std::string(what) + ": " + cudaGetErrorString(err); b200rt_destroy(ctx); return -1; };

	if ((e = cudaStreamCreateWithFlags(&ctx->own_stream, cudaStreamNonBlocking)) != cudaSuccess) return bail("cudaStreamCreate", e);
	ctx->stream = ctx->own_stream;
	for (auto & ev : ctx->ev) if ((e = cudaEventCreate(&ev)) != cudaSuccess) return bail("cudaEventCreate", e);
	for (auto & ev : ctx->staged) if ((e = cudaEventCreateWithFlags(&ev, cudaEventDisableTiming)) != cudaSuccess) return bail("cudaEventCreate", e);
	for (auto & ev : ctx->consumed) if ((e = cudaEventCreateWithFlags(&ev, cudaEventDisableTiming)) != cudaSuccess) return bail("cudaEventCreate", e);
	if ((e = cudaStreamCreateWithFlags(&ctx->upload_stream, cudaStreamNonBlocking)) != cudaSuccess) return bail("cudaStreamCreate", e);
	if (config->deep_queue == B200RT_QUEUE_AUTO) {
		for (auto & probe : ctx->auto_queue.probes) if ((e = cudaEventCreateWithFlags(&probe.copied, cudaEventDisableTiming)) != cudaSuccess) return bail("cudaEventCreate", e);
		if ((e = cudaMallocHost(reinterpret_cast<void **>(&ctx->auto_queue.stamps), b200rt_ctx::AutoQueue::SLOTS * 3 * sizeof(unsigned long long))) != cudaSuccess) return bail("cudaMallocHost", e);
	}

	cudaDeviceProp prop;
	if ((e = cudaGetDeviceProperties(&prop, config->device)) != cudaSuccess) return bail("cudaGetDeviceProperties", e);
	ctx->num_sms = prop.multiProcessorCount;

	// Texture::init (Texture.h:53-62) and the gamma table of get_screen() (fragment_fxaa.glsl:16-18)
	float ewa[128];
	{
		const float alpha = 2.0f, denom = 1.0f / 127.0f, exp_neg_alpha = expf(-alpha);
		for (int i = 0; i < 128; i++) ewa[i] = expf(-alpha * (float(i) * denom)) - exp_neg_alpha;
	}
	if ((e = ctx->kernels->init(ewa)) != cudaSuccess) return bail("kernel tables (is this an sm_100a device?)", e);

	float gamma[256];
	for (int i = 0; i < 256; i++) gamma[i] = powf(float(i) / 255.0f, 1.0f / 2.2f);
	if ((e = ctx->gamma.ensure(sizeof(gamma))) != cudaSuccess) return bail("cudaMalloc", e);
	if ((e = cudaMemcpy(ctx->gamma.ptr, gamma, sizeof(gamma), cudaMemcpyHostToDevice)) != cudaSuccess) return bail("cudaMemcpy", e);

	const size_t pixels = size_t(config->width) * config->height;
	if ((e = ctx->frame_buffer.ensure(pixels * 4)) != cudaSuccess) return bail("cudaMalloc", e);
	if ((e = ctx->presented.ensure(pixels * 3 + 16)) != cudaSuccess) return bail("cudaMalloc", e);
	if ((e = ctx->counters.ensure(COUNTER_BLOCKS * sizeof(FrameCounters))) != cudaSuccess) return bail("cudaMalloc", e);
	if ((e = cudaMemset(ctx->frame_buffer.ptr, 0, pixels * 4)) != cudaSuccess) return bail("cudaMemset", e);
	if ((e = cudaMallocHost(reinterpret_cast<void **>(&ctx->pinned_counters), sizeof(FrameCounters))) != cudaSuccess) return bail("cudaMallocHost", e);
	{ // schedule buffers; the first frame has no cost history: tiles in index order
		const int tiles = ((config->width + 31) / 32) * ((config->height + 31) / 32);
		ctx->tile_count_mine = tiles > config->tile_phase ? (tiles - config->tile_phase + config->tile_stride - 1) / config->tile_stride : 0;
		const int mine = std::max(ctx->tile_count_mine, 1);
		std::vector<unsigned> identity(mine);
		for (int i = 0; i < mine; i++) identity[i] = unsigned(i);
		if ((e = ctx->tile_cost.ensure(size_t(mine) * 4)) != cudaSuccess || (e = ctx->tile_order.ensure(size_t(mine) * 4)) != cudaSuccess) return bail("cudaMalloc", e);
		if ((e = cudaMemset(ctx->tile_cost.ptr, 0, size_t(mine) * 4)) != cudaSuccess) return bail("cudaMemset", e);
		if ((e = cudaMemcpy(ctx->tile_order.ptr, identity.data(), size_t(mine) * 4, cudaMemcpyHostToDevice)) != cudaSuccess) return bail("cudaMemcpy", e);
		// the counter ring starts cleared and armed, as tile_schedule_kernel leaves the next frame's block
		std::vector<FrameCounters> blocks(COUNTER_BLOCKS);
		memset(blocks.data(), 0, blocks.size() * sizeof(FrameCounters));
		for (FrameCounters & b : blocks) b.q_outstanding = ctx->tile_count_mine * 32;
		if ((e = cudaMemcpy(ctx->counters.ptr, blocks.data(), blocks.size() * sizeof(FrameCounters), cudaMemcpyHostToDevice)) != cudaSuccess) return bail("cudaMemcpy", e);
	}

	*out = ctx;
	return 0;
}

extern "C" void b200rt_destroy(b200rt_ctx * ctx) {
	if (!ctx) return;
	cudaSetDevice(ctx->config.device);
	if (ctx->stream) cudaStreamSynchronize(ctx->stream);
	if (ctx->upload_stream) cudaStreamSynchronize(ctx->upload_stream);
	if (ctx->copy_stream) cudaStreamSynchronize(ctx->copy_stream);
	for (DeviceBuffer * b : { &ctx->nodes, &ctx->shadow_nodes, &ctx->hot, &ctx->cold, &ctx->materials, &ctx->textures, &ctx->texels, &ctx->sky, &ctx->gamma,
	                          &ctx->frame_block[0], &ctx->frame_block[1], &ctx->frame_buffer, &ctx->presented, &ctx->colour_float, &ctx->ids, &ctx->hits, &ctx->counters, &ctx->l2_scratch, &ctx->tile_cost, &ctx->tile_order,
	                          &ctx->q_nodes, &ctx->q_bounce, &ctx->q_bounce_ready, &ctx->wide_nodes, &ctx->hot_fast, &ctx->texels_half }) b->release();
	for (auto & slot : ctx->slots) {
		slot.frame_buffer.release(); slot.presented.release();
		if (slot.pinned_counters) cudaFreeHost(slot.pinned_counters);
		for (cudaEvent_t e : { slot.render_end, slot.fxaa_end, slot.copied }) if (e) cudaEventDestroy(e);
	}
	if (ctx->copy_stream) cudaStreamDestroy(ctx->copy_stream);
	if (ctx->frame_staging) cudaFreeHost(ctx->frame_staging);
	if (ctx->pinned_out) cudaFreeHost(ctx->pinned_out);
	if (ctx->pinned_counters) cudaFreeHost(ctx->pinned_counters);
	for (auto & probe : ctx->auto_queue.probes) if (probe.copied) cudaEventDestroy(probe.copied);
	if (ctx->auto_queue.stamps) cudaFreeHost(ctx->auto_queue.stamps);
	for (auto & ev : ctx->ev) if (ev) cudaEventDestroy(ev);
	for (auto & ev : ctx->staged) if (ev) cudaEventDestroy(ev);
	for (auto & ev : ctx->consumed) if (ev) cudaEventDestroy(ev);
	if (ctx->upload_stream) cudaStreamDestroy(ctx->upload_stream);
	if (ctx->own_stream) cudaStreamDestroy(ctx->own_stream);
	delete ctx;
}

// Page-locked host memory for output images: b200rt_render_frame copies device -> caller buffer
// directly (one DMA, no staging copy) when the buffer is page-locked.
extern "C" int b200rt_host_alloc(uint64_t bytes, void ** out) {
	if (!out) return -1;
	*out = nullptr;
	cudaError_t e = cudaHostAlloc(out, size_t(bytes ? bytes : 16), cudaHostAllocDefault);
	if (e != cudaSuccess) { g_create_error = std::string("cudaHostAlloc: ") + cudaGetErrorString(e); *out = nullptr; return -1; }
	return 0;
}
extern "C" void b200rt_host_free(void * p) { if (p) cudaFreeHost(p); }

static bool is_page_locked(const void * p) {
	cudaPointerAttributes attr;
	if (cudaPointerGetAttributes(&attr, p) != cudaSuccess) { cudaGetLastError(); return false; }
	return attr.type == cudaMemoryTypeHost;
}

// ------------------------------------------------------------------------ static upload ---
static int upload(b200rt_ctx * ctx, DeviceBuffer & buffer, const void * host, size_t bytes) {
	CHECK_CUDA(ctx, buffer.ensure(bytes));
	if (bytes) CHECK_CUDA(ctx, cudaMemcpy(buffer.ptr, host, bytes, cudaMemcpyHostToDevice));
	return 0;
}

extern "C" int b200rt_upload_static(b200rt_ctx * ctx, const b200rt_static_scene * s) {
	if (!ctx) return -1;
	if (!s) return fail(ctx, "null scene");
	CHECK_CUDA(ctx, cudaSetDevice(ctx->config.device));

	if (s->node_count <= 0 || s->triangle_count <= 0 || s->blas_count <= 0 || !s->nodes || !s->triangles_hot || !s->triangles_cold || !s->blas) return fail(ctx, "static scene needs at least one BLAS with nodes and triangles");
	if (s->material_count <= 0 || s->material_count > 256 || !s->materials) return fail(ctx, "material_count must be in 1..256 (MAX_MATERIALS, Config.h:18)");
	if (s->sky_size <= 0 || !s->sky) return fail(ctx, "a sky probe is required (Sky.cpp:11)");
	const bool gpu_ingest = !s->texels && s->texels_rgb8;
	if (s->texture_count < 0 || (s->texture_count > 0 && (!s->textures || (!s->texels && !s->texels_rgb8)))) return fail(ctx, "bad texture arrays");
	if (gpu_ingest && !s->gamma_lut) return fail(ctx, "texels_rgb8 needs gamma_lut");

	// validate indices: the kernels trust them
	int blas_depth = 0;
	for (int b = 0; b < s->blas_count; b++) {
		const b200rt_blas & d = s->blas[b];
		if (d.node_offset < 0 || d.node_count <= 0 || d.node_offset + d.node_count > s->node_count || d.triangle_offset < 0 || d.triangle_count < 0 || d.triangle_offset + d.triangle_count > s->triangle_count) return fail(ctx, "BLAS range out of bounds");
		int depth = 0;
		if (const char * problem = check_tree(s->nodes + d.node_offset, d.node_count, d.triangle_count, &depth)) return fail(ctx, std::string("BLAS ") + std::to_string(b) + ": " + problem);
		blas_depth = std::max(blas_depth, depth);
		if (depth + 1 > 64) return fail(ctx, "a BLAS is deeper than BVH_TRAVERSAL_STACK_SIZE allows (Config.h:25: 64 entries)");
		for (int t = 0; t < d.triangle_count; t++) {
			const int m = d.material_offset + s->triangles_cold[d.triangle_offset + t].material_id;
			if (m < 0 || m >= s->material_count) return fail(ctx, "triangle material id out of range");
		}
	}
	for (int m = 0; m < s->material_count; m++) if (s->materials[m].texture >= s->texture_count) return fail(ctx, "material texture index out of range");
	for (int t = 0; t < s->texture_count; t++) {
		const b200rt_texture & tex = s->textures[t];
		if (tex.width <= 0 || tex.height <= 0 || tex.mip_levels < 1 || tex.mip_levels > B200RT_MAX_MIP_LEVELS || tex.texel_offset < 0) return fail(ctx, "bad texture descriptor");
		if (tex.mipmapped && ((tex.width & (tex.width - 1)) || (tex.height & (tex.height - 1)))) return fail(ctx, "a mip-mapped texture must have power-of-two sides (Texture.cpp:50): the mip-map filters wrap coordinates by masking");
		const int last = tex.mip_levels - 1;
		const int64_t end = tex.texel_offset + tex.mip_offsets[last] + int64_t(std::max(tex.width >> last, 1)) * std::max(tex.height >> last, 1);
		if (end > s->texel_count) return fail(ctx, "texture pyramid exceeds the texel pool");
	}

	// BVH nodes: same 32 bytes, viewed as two float4
	if (upload(ctx, ctx->nodes, s->nodes, size_t(s->node_count) * sizeof(b200rt_bvh_node))) return -1;
	{ // and the threaded copies the shadow rays walk
		std::vector<b200rt_bvh_node> threaded;
		threaded.reserve(size_t(s->node_count));
		ctx->shadow_node_offset.assign(s->blas_count, 0);
		ctx->records_per_copy.assign(s->blas_count, 0);
		for (int b = 0; b < s->blas_count; b++) {
			ctx->shadow_node_offset[b] = int(threaded.size());
			if (const char * problem = build_threaded_nodes(s->nodes + s->blas[b].node_offset, s->blas[b].node_count, threaded)) return fail(ctx, problem);
#ifdef RT_OCTANT_THREADED
			ctx->records_per_copy[b] = int(threaded.size()) - ctx->shadow_node_offset[b];
			for (int octant = 0; octant < 8; octant++) // copy `octant` lies at shadow_node_offset + (1 + octant) * records_per_copy
				if (const char * problem = build_threaded_nodes(s->nodes + s->blas[b].node_offset, s->blas[b].node_count, threaded, octant)) return fail(ctx, problem);
#endif
		}
		if (upload(ctx, ctx->shadow_nodes, threaded.data(), threaded.size() * sizeof(b200rt_bvh_node))) return -1;
		ctx->wide_node_offset.assign(s->blas_count, -1);
#ifdef RT_WIDE_SHADOW
		std::vector<float> wide;
		for (int b = 0; b < s->blas_count; b++) {
			const size_t first = wide.size() / 32;
			int entries = 0;
			const char * problem = build_wide_nodes(s->nodes + s->blas[b].node_offset, s->blas[b].node_count, wide, &entries);
			if (problem || entries > 48) wide.resize(first * 32);     // RT_WIDE_STACK; this tree keeps the binary walk
			else ctx->wide_node_offset[b] = int(first);
		}
		if (upload(ctx, ctx->wide_nodes, wide.data(), std::max<size_t>(wide.size(), 32) * sizeof(float))) return -1;
#endif
	}

	{ // TriangleHot 36 B -> 48 B (three aligned float4)
		std::vector<float> packed(size_t(s->triangle_count) * 12, 0.0f);
		for (int i = 0; i < s->triangle_count; i++) memcpy(&packed[size_t(i) * 12], &s->triangles_hot[i], sizeof(b200rt_triangle_hot));
		if (upload(ctx, ctx->hot, packed.data(), packed.size() * sizeof(float))) return -1;
	}
	if (upload(ctx, ctx->cold, s->triangles_cold, size_t(s->triangle_count) * sizeof(b200rt_triangle_cold))) return -1;
	if (ctx->kernels->fast_features & 2) {
		// Woop / Baldwin-Weber: rows of [e1 e2 n]^-1 (n = e1 x e2) with the offset -row . p0, so that the transformed
		// ray's third coordinate gives t and the first two give u, v. Computed in double, stored as float: {w row}{u row}{v row}.
		std::vector<float> packed(size_t(s->triangle_count) * 12, 0.0f);
		for (int i = 0; i < s->triangle_count; i++) {
			const b200rt_triangle_hot & t = s->triangles_hot[i];
			const double p0[3] = { t.position_0[0], t.position_0[1], t.position_0[2] }, e1[3] = { t.position_edge_1[0], t.position_edge_1[1], t.position_edge_1[2] }, e2[3] = { t.position_edge_2[0], t.position_edge_2[1], t.position_edge_2[2] };
			auto cross = [](const double a[3], const double b[3], double r[3]) { r[0] = a[1] * b[2] - a[2] * b[1]; r[1] = a[2] * b[0] - a[0] * b[2]; r[2] = a[0] * b[1] - a[1] * b[0]; };
			double n[3], ru[3], rv[3];
			cross(e1, e2, n); cross(e2, n, ru); cross(n, e1, rv);
			const double det = n[0] * n[0] + n[1] * n[1] + n[2] * n[2];   // (e1 x e2) . n
			const double * rows[3] = { n, ru, rv };
			for (int r = 0; r < 3; r++) {
				double row[3], offset = 0.0;
				for (int c = 0; c < 3; c++) { row[c] = det > 0.0 ? rows[r][c] / det : 0.0; offset -= row[c] * p0[c]; }
				float * out = &packed[size_t(i) * 12 + 4 * r];
				out[0] = float(row[0]); out[1] = float(row[1]); out[2] = float(row[2]); out[3] = float(offset);
			}
		}
		if (upload(ctx, ctx->hot_fast, packed.data(), packed.size() * sizeof(float))) return -1;
	}

	{ // Material -> three float4
		std::vector<DeviceMaterial> packed(s->material_count);
		for (int i = 0; i < s->material_count; i++) {
			const b200rt_material & m = s->materials[i];
			int texture = m.texture < 0 ? -1 : m.texture;
			float texture_bits; memcpy(&texture_bits, &texture, 4);
			packed[i].diffuse_texture = make_float4(m.diffuse[0], m.diffuse[1], m.diffuse[2], texture_bits);
			packed[i].reflection_ior  = make_float4(m.reflection[0], m.reflection[1], m.reflection[2], m.index_of_refraction);
			packed[i].transmittance   = make_float4(m.transmittance[0], m.transmittance[1], m.transmittance[2], 0.0f);
		}
		if (upload(ctx, ctx->materials, packed.data(), packed.size() * sizeof(DeviceMaterial))) return -1;
	}
	{
		std::vector<DeviceTexture> packed(std::max(s->texture_count, 1));
		memset(packed.data(), 0, packed.size() * sizeof(DeviceTexture));
		for (int i = 0; i < s->texture_count; i++) {
			const b200rt_texture & t = s->textures[i];
			packed[i].texel_offset = t.texel_offset; packed[i].width = t.width; packed[i].height = t.height;
			packed[i].mipmapped = t.mipmapped; packed[i].mip_levels = t.mip_levels;
			memcpy(packed[i].mip_offsets, t.mip_offsets, sizeof(t.mip_offsets));
		}
		if (upload(ctx, ctx->textures, packed.data(), packed.size() * sizeof(DeviceTexture))) return -1;
	}
	{ // texels and sky: float3 -> float4, in chunks to bound the host staging memory
		auto upload_rgb = [&](DeviceBuffer & buffer, const float * rgb, size_t count) -> int {
			CHECK_CUDA(ctx, buffer.ensure(std::max<size_t>(count, 1) * sizeof(float4)));
			const size_t chunk = size_t(1) << 22;
			std::vector<float4> staging(std::min(count, chunk));
			for (size_t base = 0; base < count; base += chunk) {
				const size_t n = std::min(chunk, count - base);
				for (size_t i = 0; i < n; i++) staging[i] = make_float4(rgb[(base + i) * 3], rgb[(base + i) * 3 + 1], rgb[(base + i) * 3 + 2], 0.0f);
				CHECK_CUDA(ctx, cudaMemcpy(static_cast<float4 *>(buffer.ptr) + base, staging.data(), n * sizeof(float4), cudaMemcpyHostToDevice));
			}
			return 0;
		};
		if (!gpu_ingest) {
			if (upload_rgb(ctx->texels, s->texels, size_t(s->texel_count))) return -1;
		} else {
			// the 8-bit level-0 images go up as they are; conversion and mip chains are built in HBM
			size_t level0_texels = 0;
			for (int t = 0; t < s->texture_count; t++) level0_texels += size_t(s->textures[t].width) * s->textures[t].height;
			DeviceBuffer bytes, lut;
			cudaError_t e = bytes.ensure(level0_texels * 3);
			if (e == cudaSuccess) e = lut.ensure(256 * sizeof(float));
			if (e == cudaSuccess) e = ctx->texels.ensure(std::max<size_t>(size_t(s->texel_count), 1) * sizeof(float4));
			if (e == cudaSuccess) e = cudaMemcpyAsync(bytes.ptr, s->texels_rgb8, level0_texels * 3, cudaMemcpyHostToDevice, ctx->stream);
			if (e == cudaSuccess) e = cudaMemcpyAsync(lut.ptr, s->gamma_lut, 256 * sizeof(float), cudaMemcpyHostToDevice, ctx->stream);
			size_t source = 0;
			for (int t = 0; t < s->texture_count && e == cudaSuccess; t++) {
				const b200rt_texture & tex = s->textures[t];
				float4 * pool = static_cast<float4 *>(ctx->texels.ptr) + tex.texel_offset;
				const long long n = (long long)tex.width * tex.height;
				e = ctx->kernels->texture_level0(static_cast<const unsigned char *>(bytes.ptr) + source * 3, static_cast<const float *>(lut.ptr), pool, n, ctx->stream);
				source += size_t(n);
				int previous_width = tex.width;
				for (int level = 1; level < tex.mip_levels && tex.mipmapped && e == cudaSuccess; level++) {
					const int w = tex.width >> level, h = tex.height >> level;
					e = ctx->kernels->texture_mip(pool + tex.mip_offsets[level - 1], pool + tex.mip_offsets[level], w, h, previous_width, ctx->stream);
					previous_width = w;
				}
			}
			if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
			bytes.release(); lut.release();
			if (e != cudaSuccess) { ctx->error = std::string("texture ingest: ") + cudaGetErrorString(e); return -1; }
		}
		if (upload_rgb(ctx->sky, s->sky, size_t(s->sky_size) * s->sky_size + 1)) return -1; // + the pad texel (Sky.cpp:45)
	}
	if (ctx->kernels->fast_features & 4) {
		CHECK_CUDA(ctx, ctx->texels_half.ensure(std::max<size_t>(size_t(s->texel_count), 1) * sizeof(uint2)));
		CHECK_CUDA(ctx, ctx->kernels->texel_halves(static_cast<const float4 *>(ctx->texels.ptr), static_cast<uint2 *>(ctx->texels_half.ptr), s->texel_count, ctx->stream));
		CHECK_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
	}
	ctx->sky_size = s->sky_size;
	ctx->blas_stack_entries = blas_depth + 1;
	ctx->blas.assign(s->blas, s->blas + s->blas_count);
	ctx->have_static = true;
	ctx->have_frame = false;
	return 0;
}

extern "C" int b200rt_download_texels(b200rt_ctx * ctx, float * texels, int64_t first, int64_t count) {
	if (!ctx) return -1;
	if (!texels || first < 0 || count < 0 || size_t(first + count) * sizeof(float4) > ctx->texels.bytes) return fail(ctx, "texel range out of bounds");
	CHECK_CUDA(ctx, cudaSetDevice(ctx->config.device));
	std::vector<float4> staging(static_cast<size_t>(count));
	CHECK_CUDA(ctx, cudaMemcpy(staging.data(), static_cast<const float4 *>(ctx->texels.ptr) + first, size_t(count) * sizeof(float4), cudaMemcpyDeviceToHost));
	for (int64_t i = 0; i < count; i++) { texels[3 * i] = staging[i].x; texels[3 * i + 1] = staging[i].y; texels[3 * i + 2] = staging[i].z; }
	return 0;
}

// ------------------------------------------------------------------------- per frame ---
static size_t align16(size_t n) { return (n + 15) & ~size_t(15); }

// Validates the frame, packs it into the pinned staging block and issues ONE host->device copy.
static int stage_frame(b200rt_ctx * ctx, const b200rt_frame * f) {
	if (!ctx->have_static) return fail(ctx, "b200rt_upload_static has not been called");
	if (!f) return fail(ctx, "null frame");
	if (f->instance_count <= 0 || !f->instances || !f->tlas_nodes || !f->tlas_indices || f->tlas_node_count <= 0) return fail(ctx, "frame needs at least one instance and a TLAS (TopLevelBVH.cpp:6)");
	if (f->sphere_count < 0 || f->plane_count < 0 || f->light_count < 0) return fail(ctx, "negative count");

	const int material_count = int(ctx->materials.bytes / sizeof(DeviceMaterial));
	for (int i = 0; i < f->instance_count; i++) {
		if (f->instances[i].blas < 0 || f->instances[i].blas >= int(ctx->blas.size())) return fail(ctx, "instance BLAS index out of range");
		if (f->tlas_indices[i] < 0 || f->tlas_indices[i] >= f->instance_count) return fail(ctx, "TLAS primitive index out of range");
	}
	int tlas_depth = 0;
	if (const char * problem = check_tree(f->tlas_nodes, f->tlas_node_count, f->instance_count, &tlas_depth)) return fail(ctx, std::string("TLAS: ") + problem);
	if (tlas_depth + 1 > RT_TLAS_STACK_SIZE) return fail(ctx, "TLAS deeper than the traversal stack (RT_TLAS_STACK_SIZE)");
	ctx->tlas_stack_entries = tlas_depth + 1;
	if ((f->sphere_count > 0 && !f->spheres) || (f->plane_count > 0 && !f->planes) || (f->light_count > 0 && !f->lights)) return fail(ctx, "a count is positive but its array is NULL");
	for (int i = 0; i < f->sphere_count; i++) if (f->spheres[i].material_id < 0 || f->spheres[i].material_id >= material_count) return fail(ctx, "sphere material id out of range");
	for (int i = 0; i < f->plane_count;  i++) if (f->planes[i].material_id  < 0 || f->planes[i].material_id  >= material_count) return fail(ctx, "plane material id out of range");
	for (int i = 0; i < f->light_count; i++) {
		if (f->lights[i].kind < 0 || f->lights[i].kind > 2) return fail(ctx, "bad light kind");
		if (i > 0 && f->lights[i].kind < f->lights[i - 1].kind) return fail(ctx, "lights must be ordered point, spot, directional (Raytracer.cpp:155-198)");
	}

	const size_t off_nodes     = 0;
	const size_t off_indices   = off_nodes     + align16(size_t(f->tlas_node_count) * sizeof(b200rt_bvh_node));
	const size_t off_instances = off_indices   + align16(size_t(f->instance_count) * sizeof(int));
	const size_t off_spheres   = off_instances + size_t(f->instance_count) * sizeof(DeviceInstance);
	const size_t off_planes    = off_spheres   + size_t(f->sphere_count) * sizeof(DeviceSphere);
	const size_t off_lights    = off_planes    + size_t(f->plane_count) * sizeof(DevicePlane);
	const size_t total         = off_lights    + size_t(f->light_count) * sizeof(DeviceLight);

	if (total > ctx->frame_staging_bytes) {
		CHECK_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
		if (ctx->frame_staging) cudaFreeHost(ctx->frame_staging);
		ctx->frame_staging = nullptr; ctx->frame_staging_bytes = 0;
		CHECK_CUDA(ctx, cudaMallocHost(&ctx->frame_staging, total * 4));
		ctx->frame_staging_bytes = total * 2;
	}
	CHECK_CUDA(ctx, ctx->frame_block[ctx->staging_half].ensure(total));

	// Frames may be issued without waiting for the previous one (b200rt_render_frame_device with
	// stats == NULL): the two staging halves alternate, and a half is rewritten only after the
	// copy that last read it has completed.
	const int half = ctx->staging_half;
	ctx->staging_half ^= 1;
	CHECK_CUDA(ctx, cudaEventSynchronize(ctx->staged[half]));
	char * base = static_cast<char *>(ctx->frame_staging) + size_t(half) * ctx->frame_staging_bytes;
	memcpy(base + off_nodes,   f->tlas_nodes,   size_t(f->tlas_node_count) * sizeof(b200rt_bvh_node));
	memcpy(base + off_indices, f->tlas_indices, size_t(f->instance_count) * sizeof(int));

	auto as_float = [](int v) { float r; memcpy(&r, &v, 4); return r; };

	DeviceInstance * instances = reinterpret_cast<DeviceInstance *>(base + off_instances);
	for (int i = 0; i < f->instance_count; i++) {
		const b200rt_instance & in = f->instances[i];
		const b200rt_blas & blas = ctx->blas[in.blas];
		// Matrix4 cells[i + 4j]: the first twelve floats are the three rows the transforms use (Matrix4.h:39-85)
		memcpy(&instances[i].inverse, in.inverse, 12 * sizeof(float));
		memcpy(&instances[i].world,   in.world,   12 * sizeof(float));
		instances[i].node_offset = blas.node_offset; instances[i].triangle_offset = blas.triangle_offset;
		instances[i].material_offset = blas.material_offset; instances[i].shadow_node_offset = ctx->shadow_node_offset[in.blas];
		memset(instances[i].reserved, 0, sizeof(instances[i].reserved));
		instances[i].reserved[0] = ctx->records_per_copy[in.blas];
		instances[i].reserved[1] = ctx->wide_node_offset[in.blas];
	}
	DeviceSphere * spheres = reinterpret_cast<DeviceSphere *>(base + off_spheres);
	for (int i = 0; i < f->sphere_count; i++) {
		const b200rt_sphere & s = f->spheres[i];
		spheres[i].center_r2 = make_float4(s.center[0], s.center[1], s.center[2], s.radius_squared);
		spheres[i].rinv_mat  = make_float4(s.radius_inv, as_float(s.material_id), 0.0f, 0.0f);
	}
	DevicePlane * planes = reinterpret_cast<DevicePlane *>(base + off_planes);
	for (int i = 0; i < f->plane_count; i++) {
		const b200rt_plane & p = f->planes[i];
		planes[i].normal_d   = make_float4(p.normal[0], p.normal[1], p.normal[2], p.distance);
		planes[i].u_axis_mat = make_float4(p.u_axis[0], p.u_axis[1], p.u_axis[2], as_float(p.material_id));
		planes[i].v_axis     = make_float4(p.v_axis[0], p.v_axis[1], p.v_axis[2], 0.0f);
	}
	DeviceLight * lights = reinterpret_cast<DeviceLight *>(base + off_lights);
	for (int i = 0; i < f->light_count; i++) {
		const b200rt_light & l = f->lights[i];
		lights[i].colour_kind    = make_float4(l.colour[0], l.colour[1], l.colour[2], as_float(l.kind));
		lights[i].position_inner = make_float4(l.position[0], l.position[1], l.position[2], l.inner_cutoff);
		lights[i].negdir_outer   = make_float4(l.negative_direction[0], l.negative_direction[1], l.negative_direction[2], l.outer_cutoff);
	}

	// The block travels on its own stream: when frames are pipelined it is in HBM long before the
	// previous frame's kernels are done, and the compute stream carries kernels only. The device block
	// is rewritten only after the kernels of the frame that last read it (two frames ago).
	CHECK_CUDA(ctx, cudaStreamWaitEvent(ctx->upload_stream, ctx->consumed[half], 0));
	CHECK_CUDA(ctx, cudaMemcpyAsync(ctx->frame_block[half].ptr, base, total, cudaMemcpyHostToDevice, ctx->upload_stream));
	CHECK_CUDA(ctx, cudaEventRecord(ctx->staged[half], ctx->upload_stream));
	CHECK_CUDA(ctx, cudaStreamWaitEvent(ctx->stream, ctx->staged[half], 0));
	ctx->block_in_use = half;

	const char * dev = static_cast<const char *>(ctx->frame_block[half].ptr);
	SceneParams & P = ctx->params;
	P.nodes = static_cast<const float4 *>(ctx->nodes.ptr);
	P.shadow_nodes = static_cast<const float4 *>(ctx->shadow_nodes.ptr);
	P.wide_nodes = static_cast<const float4 *>(ctx->wide_nodes.ptr);
	P.triangles_hot = static_cast<const float4 *>(ctx->hot.ptr);
	P.triangles_test = (ctx->kernels->fast_features & 2) ? static_cast<const float4 *>(ctx->hot_fast.ptr) : P.triangles_hot;
	P.texels_half = static_cast<const uint2 *>(ctx->texels_half.ptr);
	P.triangles_cold = static_cast<const float4 *>(ctx->cold.ptr);
	P.materials = static_cast<const DeviceMaterial *>(ctx->materials.ptr);
	P.textures = static_cast<const DeviceTexture *>(ctx->textures.ptr);
	P.texels = static_cast<const float4 *>(ctx->texels.ptr);
	P.sky = static_cast<const float4 *>(ctx->sky.ptr);
	P.sky_size = ctx->sky_size;
	P.tlas_nodes   = reinterpret_cast<const float4 *>(dev + off_nodes);
	P.tlas_indices = reinterpret_cast<const int *>(dev + off_indices);
	P.instances    = reinterpret_cast<const DeviceInstance *>(dev + off_instances);
	P.spheres      = reinterpret_cast<const DeviceSphere *>(dev + off_spheres);
	P.planes       = reinterpret_cast<const DevicePlane *>(dev + off_planes);
	P.lights       = reinterpret_cast<const DeviceLight *>(dev + off_lights);
	P.sphere_count = f->sphere_count; P.plane_count = f->plane_count; P.light_count = f->light_count;
	for (int c = 0; c < 3; c++) {
		P.cam_position[c] = f->camera.position[c]; P.cam_top_left[c] = f->camera.top_left_corner[c];
		P.cam_x_axis[c] = f->camera.x_axis[c];     P.cam_y_axis[c] = f->camera.y_axis[c];
		P.ambient[c] = f->ambient[c];
	}
	const b200rt_config & c = ctx->config;
	P.width = c.width; P.height = c.height; P.bounces = c.bounces; P.mipmap_filter = c.mipmap_filter; P.max_anisotropy = c.max_anisotropy;
	P.traversal_naive = c.traversal_naive; P.texture_sample_mode = c.texture_sample_mode; P.count_shipped_walk = c.count_traversal == 2;
	P.tiles_x = (c.width + 31) / 32; P.tiles_y = (c.height + 31) / 32; // Window.cpp:11-12
	P.tile_stride = c.tile_stride; P.tile_phase = c.tile_phase;
	const int tiles = P.tiles_x * P.tiles_y;
	P.tile_count_mine = tiles > c.tile_phase ? (tiles - c.tile_phase + c.tile_stride - 1) / c.tile_stride : 0;
	for (int b = 0; b < B200RT_MAX_BANDS; b++) P.band_targets[b] = reinterpret_cast<unsigned *>(uintptr_t(ctx->band_targets[b]));
	P.band_count = ctx->band_count; P.band_rows = ctx->band_rows; P.band_halo = ctx->band_halo;
	P.tile_order = static_cast<const unsigned *>(ctx->tile_order.ptr);
	P.tile_cost  = static_cast<unsigned *>(ctx->tile_cost.ptr);

	ctx->have_frame = true;
	return 0;
}

// Pools of the deep-pixel queue (render.cuh), sized by this context's share of the frame; allocated on first use.
// A pool that runs out costs nothing but speed: the producer walks that subtree depth-first (FrameCounters::q_fallbacks).
static int ensure_queue(b200rt_ctx * ctx, SceneParams & P) {
	const size_t pixels = size_t(std::max(ctx->tile_count_mine, 1)) * 1024;
	size_t nodes = std::min<size_t>(4 * pixels, (size_t(1) << 26) - 1), bounce = 2 * nodes;   // two child slots per record
	if (const char * text = getenv("B200RT_QUEUE_CAPACITY")) nodes = bounce = size_t(std::max(atoi(text), 32));   // tests: small pools force the fall-back paths
	if (!ctx->q_nodes.ptr) {
		CHECK_CUDA(ctx, ctx->q_nodes.ensure(nodes * 6 * sizeof(float4)));
		CHECK_CUDA(ctx, ctx->q_bounce.ensure(bounce * 5 * sizeof(float4)));
		CHECK_CUDA(ctx, ctx->q_bounce_ready.ensure(bounce * sizeof(unsigned)));
		// "ready" means "holds this launch's serial number": the arrays must not start with a previous owner's
		CHECK_CUDA(ctx, cudaMemset(ctx->q_bounce_ready.ptr, 0, bounce * sizeof(unsigned)));
	}
	P.q_nodes = static_cast<float4 *>(ctx->q_nodes.ptr); P.q_bounce = static_cast<float4 *>(ctx->q_bounce.ptr);
	P.q_bounce_ready = static_cast<unsigned *>(ctx->q_bounce_ready.ptr);
	P.q_node_capacity = unsigned(nodes); P.q_bounce_capacity = unsigned(bounce);
	return 0;
}

// B200RT_QUEUE_AUTO: collects the stamps that have arrived, decides when both candidates have their samples, and names
// the candidate of the frame about to be issued (see b200rt_ctx::AutoQueue).
static int auto_queue_candidate(b200rt_ctx * ctx) {
	typedef b200rt_ctx::AutoQueue A;
	A & a = ctx->auto_queue;
	for (int s = 0; s < A::SLOTS; s++) {
		A::Probe & probe = a.probes[s];
		if (probe.candidate < 0 || cudaEventQuery(probe.copied) != cudaSuccess) continue;
		const unsigned long long * t = a.stamps + 3 * s;   // first_idle_ns, last_done_ns, first_start_ns (the firsts stored as ~t)
		if (t[2] != 0 && t[1] > ~t[2]) a.best_ms[probe.candidate] = std::min(a.best_ms[probe.candidate], float(double(t[1] - ~t[2]) * 1e-6));
		a.samples[probe.candidate]++;
		probe.candidate = -1;
	}
	if (!a.decided && a.samples[0] >= A::PAIRS && a.samples[1] >= A::PAIRS) {
		a.choice = a.best_ms[1] < 0.97f * a.best_ms[0] ? 1 : 0;   // a tie goes to the walk without pools
		a.decided = true;
	}
	int candidate = a.choice;
	a.issued = -1;
	if (a.frame >= A::WARM && a.frame < A::WARM + 2 * A::PAIRS) a.issued = candidate = (a.frame - A::WARM) & 1;
	if (++a.frame >= A::PERIOD) { a.frame = A::WARM; a.decided = false; a.best_ms[0] = a.best_ms[1] = 1e30f; a.samples[0] = a.samples[1] = 0; }
	return candidate;
}

// Launches the render kernel over this context's tiles; the frame counts in the ring's current block. frame_buffer is
// a device pointer. banded: store through the band targets (b200rt_set_band_targets) — b200rt_render_frame_device
// only; every other entry point renders into the one image it was given.
static int launch_render(b200rt_ctx * ctx, unsigned * frame_buffer, float * colour_float, bool count, bool banded = false) {
	const unsigned total_warps = unsigned(ctx->params.tile_count_mine) * 32u;
	if (total_warps == 0) return 0;
	SceneParams P = ctx->params;
	if (!banded) P.band_count = 0;
	P.blas_stack_entries = ctx->blas_stack_entries; P.tlas_stack_entries = ctx->tlas_stack_entries;
	P.queue_mode = (count || P.bounces == 0 || ctx->queue_stalled) ? 0 : ctx->config.deep_queue;
	P.queue_inline_below = ctx->queue_inline_below; P.queue_idle_looks = ctx->queue_idle_looks; P.queue_idle_shift = ctx->queue_idle_shift; P.queue_item_batch = ctx->queue_item_batch; P.queue_keep_refraction = ctx->queue_keep_refraction;
	ctx->auto_queue.issued = -1;
	if (P.queue_mode == B200RT_QUEUE_AUTO) {
		P.queue_mode = auto_queue_candidate(ctx) ? B200RT_QUEUE_ALL : B200RT_QUEUE_OFF;
	}
	if (P.queue_mode != 0) {
		if (ensure_queue(ctx, P)) return -1;
		if (++ctx->launch_serial == 0) ctx->launch_serial = 1;
		P.frame_serial = ctx->launch_serial;
	}
	// persistent grid: every SM filled to the occupancy limit, never more warps than units
	int & per_sm = ctx->blocks_per_sm[count ? 1 : 0][P.queue_mode != 0 ? 1 : 0];
	if (per_sm == 0) {
		CHECK_CUDA(ctx, ctx->kernels->render_occupancy(P, count, &per_sm));
		if (per_sm < 1) per_sm = 1;
	}
	unsigned blocks = unsigned(ctx->num_sms) * unsigned(per_sm);
	const unsigned needed = (total_warps + 7) / 8;   // 8 warps per 256-thread block
	if (blocks > needed) blocks = needed;
	CHECK_CUDA(ctx, ctx->kernels->render(P, frame_buffer, colour_float, counter_block(ctx, ctx->frame_number), count ? 1 : 0, blocks, ctx->stream));
	if (ctx->test_stall && P.queue_mode != 0) { // as if the watchdog had fired in this launch
		ctx->test_stall = false;
		CHECK_CUDA(ctx, cudaMemsetAsync(&counter_block(ctx, ctx->frame_number)->q_abort, 1, sizeof(unsigned), ctx->stream));
	}
	return 0;
}

// Called after the last kernel of a frame has been issued: from here on the frame's device block may be rewritten (stage_frame).
static int release_frame_block(b200rt_ctx * ctx) {
	CHECK_CUDA(ctx, cudaEventRecord(ctx->consumed[ctx->block_in_use], ctx->stream));
	return 0;
}
// The last kernel of every frame: orders the tiles for the next one by what they cost in this one, and clears the
// next frame's counter block. Returns this frame's block (for the copy to the host) through *counted.
static int launch_schedule(b200rt_ctx * ctx, FrameCounters ** counted = nullptr) {
	if (counted) *counted = counter_block(ctx, ctx->frame_number);
	CHECK_CUDA(ctx, ctx->kernels->schedule(static_cast<unsigned *>(ctx->tile_cost.ptr), static_cast<unsigned *>(ctx->tile_order.ptr), ctx->params.tile_count_mine, counter_block(ctx, ctx->frame_number + 1), ctx->stream));
	if (ctx->auto_queue.issued >= 0) {   // one of AUTO's timed frames: its stamps follow it to the host (the block is cleared three frames from now)
		typedef b200rt_ctx::AutoQueue A;
		A & a = ctx->auto_queue;
		for (int s = 0; s < A::SLOTS; s++) if (a.probes[s].candidate < 0) {
			CHECK_CUDA(ctx, cudaMemcpyAsync(a.stamps + 3 * s, &counter_block(ctx, ctx->frame_number)->first_idle_ns, 3 * sizeof(unsigned long long), cudaMemcpyDeviceToHost, ctx->stream));
			CHECK_CUDA(ctx, cudaEventRecord(a.probes[s].copied, ctx->stream));
			a.probes[s].candidate = a.issued;
			break;
		}
		a.issued = -1;
	}
	ctx->frame_number++;
	return 0;
}

static int launch_fxaa(b200rt_ctx * ctx, const unsigned * frame_buffer, unsigned char * presented, int row_begin = 0, int row_end = -1) {
	if (row_end < 0) row_end = ctx->config.height;
	if (row_end <= row_begin) return 0;
	CHECK_CUDA(ctx, ctx->kernels->fxaa(frame_buffer, presented, static_cast<const float *>(ctx->gamma.ptr), ctx->config.width, ctx->config.height, ctx->config.enable_fxaa, row_begin, row_end, ctx->stream));
	return 0;
}

// The queue kernel's watchdog (kernels.cuh: RT_QUEUE_WATCHDOG_NS) gave the frame up: its image is incomplete. The
// context renders depth-first from then on; a synchronous entry point renders the frame again, an asynchronous one
// reports the frame as failed.
static bool queue_stalled(b200rt_ctx * ctx, const FrameCounters & c) {
	if (c.q_abort == 0 || ctx->queue_stalled) return false;
	ctx->queue_stalled = true;
	fprintf(stderr, "b200rt: the deep-pixel queue made no progress for %.1f s; this context renders depth-first from now on\n", double(RT_QUEUE_WATCHDOG_NS) * 1e-9);
	return true;
}

static void counters_to_stats(const FrameCounters & c, b200rt_stats * stats) {
	stats->primary_rays = c.primary; stats->shadow_rays = c.shadow; stats->reflection_rays = c.reflection; stats->refraction_rays = c.refraction;
	stats->nodes_visited = c.nodes; stats->triangles_tested = c.triangles; stats->texels_fetched = c.texels;
	stats->hits_accepted = c.accepted; stats->instances_entered = c.instances;
	stats->device_ms_render = stats->device_ms_fxaa = stats->device_ms_total = 0.0f;
	stats->device_ms_tail = (c.first_idle_ns != 0 && c.last_done_ns > ~c.first_idle_ns) ? float(double(c.last_done_ns - ~c.first_idle_ns) * 1e-6) : 0.0f;
	stats->queue_nodes = c.q_node_reserved; stats->queue_bounce_items = c.q_bounce_reserved; stats->queue_shadow_items = 0; stats->queue_fallbacks = c.q_fallbacks;
}

static void fill_stats(b200rt_ctx * ctx, b200rt_stats * stats) {
	counters_to_stats(*ctx->pinned_counters, stats);
	cudaEventElapsedTime(&stats->device_ms_render, ctx->ev[1], ctx->ev[2]);
	cudaEventElapsedTime(&stats->device_ms_fxaa,   ctx->ev[2], ctx->ev[3]);
}

extern "C" int b200rt_render_frame(b200rt_ctx * ctx, const b200rt_frame * frame, const b200rt_outputs * out) {
	if (!ctx) return -1;
	if (!out) return fail(ctx, "null outputs");
	CHECK_CUDA(ctx, cudaSetDevice(ctx->config.device));

	const size_t pixels = size_t(ctx->config.width) * ctx->config.height;
	const size_t off_fb = 0, off_rgb = align16(pixels * 4), pinned_total = off_rgb + align16(pixels * 3);
	if (pinned_total > ctx->pinned_out_bytes) {
		if (ctx->pinned_out) cudaFreeHost(ctx->pinned_out);
		ctx->pinned_out = nullptr; ctx->pinned_out_bytes = 0;
		CHECK_CUDA(ctx, cudaMallocHost(&ctx->pinned_out, pinned_total));
		ctx->pinned_out_bytes = pinned_total;
	}
	if (out->colour_float) CHECK_CUDA(ctx, ctx->colour_float.ensure(pixels * 12));
	if (out->primary_ids)  CHECK_CUDA(ctx, ctx->ids.ensure(pixels * 8));
	if (out->primary_hit)  CHECK_CUDA(ctx, ctx->hits.ensure(pixels * 16));

	CHECK_CUDA(ctx, cudaEventRecord(ctx->ev[0], ctx->stream));
	if (stage_frame(ctx, frame)) return -1;

	unsigned * fb = static_cast<unsigned *>(ctx->frame_buffer.ptr);
	render_again:   // (only after the queue kernel's watchdog fired: the same frame once more, depth-first)
	CHECK_CUDA(ctx, cudaEventRecord(ctx->ev[1], ctx->stream));
	if (launch_render(ctx, fb, out->colour_float ? static_cast<float *>(ctx->colour_float.ptr) : nullptr, ctx->config.count_traversal != 0)) return -1;
	CHECK_CUDA(ctx, cudaEventRecord(ctx->ev[2], ctx->stream));
	if (out->presented_rgb8 && launch_fxaa(ctx, fb, static_cast<unsigned char *>(ctx->presented.ptr))) return -1;
	FrameCounters * counted = nullptr;
	if (launch_schedule(ctx, &counted)) return -1;
	CHECK_CUDA(ctx, cudaEventRecord(ctx->ev[3], ctx->stream));

	if (out->primary_ids || out->primary_hit)
		CHECK_CUDA(ctx, ctx->kernels->primary_hits(ctx->params, out->primary_ids ? static_cast<int *>(ctx->ids.ptr) : nullptr, out->primary_hit ? static_cast<float *>(ctx->hits.ptr) : nullptr, ctx->stream));
	if (release_frame_block(ctx)) return -1;

	// device -> pinned host, then one synchronisation
	// page-locked caller buffers (b200rt_host_alloc, cudaHostRegister, ...) are written by the DMA
	// engine directly; pageable ones go through the context's pinned staging block + one memcpy
	char * pinned = static_cast<char *>(ctx->pinned_out);
	const bool direct_fb  = out->frame_buffer   && is_page_locked(out->frame_buffer);
	const bool direct_rgb = out->presented_rgb8 && is_page_locked(out->presented_rgb8);
	if (out->frame_buffer)   CHECK_CUDA(ctx, cudaMemcpyAsync(direct_fb  ? static_cast<void *>(out->frame_buffer)   : pinned + off_fb,  fb, pixels * 4, cudaMemcpyDeviceToHost, ctx->stream));
	if (out->presented_rgb8) CHECK_CUDA(ctx, cudaMemcpyAsync(direct_rgb ? static_cast<void *>(out->presented_rgb8) : pinned + off_rgb, ctx->presented.ptr, pixels * 3, cudaMemcpyDeviceToHost, ctx->stream));
	CHECK_CUDA(ctx, cudaMemcpyAsync(ctx->pinned_counters, counted, sizeof(FrameCounters), cudaMemcpyDeviceToHost, ctx->stream));
	if (out->colour_float) CHECK_CUDA(ctx, cudaMemcpyAsync(out->colour_float, ctx->colour_float.ptr, pixels * 12, cudaMemcpyDeviceToHost, ctx->stream));
	if (out->primary_ids)  CHECK_CUDA(ctx, cudaMemcpyAsync(out->primary_ids, ctx->ids.ptr, pixels * 8, cudaMemcpyDeviceToHost, ctx->stream));
	if (out->primary_hit)  CHECK_CUDA(ctx, cudaMemcpyAsync(out->primary_hit, ctx->hits.ptr, pixels * 16, cudaMemcpyDeviceToHost, ctx->stream));
	CHECK_CUDA(ctx, cudaEventRecord(ctx->ev[4], ctx->stream));
	cudaError_t sync = cudaStreamSynchronize(ctx->stream);
	if (sync != cudaSuccess) { ctx->error = std::string("render failed: ") + cudaGetErrorString(sync); return -1; }
	if (queue_stalled(ctx, *ctx->pinned_counters)) goto render_again;

	if (out->frame_buffer   && !direct_fb)  memcpy(out->frame_buffer,   pinned + off_fb,  pixels * 4);
	if (out->presented_rgb8 && !direct_rgb) memcpy(out->presented_rgb8, pinned + off_rgb, pixels * 3);
	if (out->stats) {
		fill_stats(ctx, out->stats);
		cudaEventElapsedTime(&out->stats->device_ms_total, ctx->ev[0], ctx->ev[4]);
	}
	return 0;
}

extern "C" int b200rt_render_frame_device(b200rt_ctx * ctx, const b200rt_frame * frame, uint64_t frame_buffer_dev, b200rt_stats * stats) {
	if (!ctx) return -1;
	if (!frame_buffer_dev) return fail(ctx, "null device frame buffer");
	CHECK_CUDA(ctx, cudaSetDevice(ctx->config.device));
	// timing events only for the caller who asks for times: each one stalls the stream for ~20 us
	if (stats) CHECK_CUDA(ctx, cudaEventRecord(ctx->ev[0], ctx->stream));
	if (stage_frame(ctx, frame)) return -1;
	if (stats) CHECK_CUDA(ctx, cudaEventRecord(ctx->ev[1], ctx->stream));
	if (launch_render(ctx, reinterpret_cast<unsigned *>(uintptr_t(frame_buffer_dev)), nullptr, false, true)) return -1;   // the one entry point that stores through the band targets
	if (stats) CHECK_CUDA(ctx, cudaEventRecord(ctx->ev[2], ctx->stream));
	FrameCounters * counted = nullptr;
	if (launch_schedule(ctx, &counted)) return -1;
	if (stats) CHECK_CUDA(ctx, cudaEventRecord(ctx->ev[3], ctx->stream));
	if (release_frame_block(ctx)) return -1;
	if (!stats) return 0; // asynchronous: the caller orders later work on the context's stream
	CHECK_CUDA(ctx, cudaMemcpyAsync(ctx->pinned_counters, counted, sizeof(FrameCounters), cudaMemcpyDeviceToHost, ctx->stream));
	CHECK_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
	if (queue_stalled(ctx, *ctx->pinned_counters)) return b200rt_render_frame_device(ctx, frame, frame_buffer_dev, stats);
	fill_stats(ctx, stats); cudaEventElapsedTime(&stats->device_ms_total, ctx->ev[0], ctx->ev[3]);
	return 0;
}

extern "C" int b200rt_last_frame_stats(b200rt_ctx * ctx, b200rt_stats * stats) {
	if (!ctx) return -1;
	if (!stats) return fail(ctx, "null stats");
	if (!ctx->have_frame || ctx->frame_number == 0) return fail(ctx, "no frame has been rendered yet");
	CHECK_CUDA(ctx, cudaSetDevice(ctx->config.device));
	CHECK_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
	CHECK_CUDA(ctx, cudaMemcpy(ctx->pinned_counters, counter_block(ctx, ctx->frame_number - 1), sizeof(FrameCounters), cudaMemcpyDeviceToHost));
	const FrameCounters & c = *ctx->pinned_counters;
#ifdef RT_QUEUE_TRACE
	if (c.q_pad0[10]) fprintf(stderr, "trace items: %u warp iterations, %.1f lanes each on average; iterations with 1 / 2-3 / 4-7 / 8-15 / 16-32 lanes: %u / %u / %u / %u / %u\n", c.q_pad0[10], double(c.q_pad0[11]) / c.q_pad0[10], c.q_pad0[12], c.q_pad0[13], c.q_pad0[14], c.q_pad0[15], c.q_pad0[16]);
	for (int level = 9; level >= 0; level--) if (c.q_pad0[18 + level]) fprintf(stderr, "trace level %d: %u walks, last picked up at %u us, last finished at %u us\n", level, c.q_pad0[18 + level], c.q_pad0[level], c.q_pad1[level]);
#endif
	if (queue_stalled(ctx, c)) return fail(ctx, "the deep-pixel queue stalled in the last frame: its image is incomplete; render it again (the context has switched to the depth-first walk)");
	counters_to_stats(c, stats);
	if (c.first_start_ns != 0 && c.last_done_ns > ~c.first_start_ns) stats->device_ms_render = float(double(c.last_done_ns - ~c.first_start_ns) * 1e-6);
	stats->device_ms_total = stats->device_ms_render;
	return 0;
}

extern "C" int b200rt_present_device(b200rt_ctx * ctx, uint64_t frame_buffer_dev, uint64_t presented_dev, float * device_ms) {
	if (!ctx) return -1;
	if (!frame_buffer_dev || !presented_dev) return fail(ctx, "null device pointer");
	CHECK_CUDA(ctx, cudaSetDevice(ctx->config.device));
	if (device_ms) CHECK_CUDA(ctx, cudaEventRecord(ctx->ev[2], ctx->stream));
	if (launch_fxaa(ctx, reinterpret_cast<const unsigned *>(uintptr_t(frame_buffer_dev)), reinterpret_cast<unsigned char *>(uintptr_t(presented_dev)))) return -1;
	if (device_ms) CHECK_CUDA(ctx, cudaEventRecord(ctx->ev[3], ctx->stream));
	if (!device_ms) return 0; // asynchronous
	CHECK_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
	cudaEventElapsedTime(device_ms, ctx->ev[2], ctx->ev[3]);
	return 0;
}

extern "C" int b200rt_present_rows_device(b200rt_ctx * ctx, uint64_t frame_buffer_dev, uint64_t presented_dev, int32_t row_begin, int32_t row_end, float * device_ms) {
	if (!ctx) return -1;
	if (!frame_buffer_dev || !presented_dev) return fail(ctx, "null device pointer");
	if (row_begin < 0 || row_end > ctx->config.height || row_begin > row_end) return fail(ctx, "bad row range");
	CHECK_CUDA(ctx, cudaSetDevice(ctx->config.device));
	if (device_ms) CHECK_CUDA(ctx, cudaEventRecord(ctx->ev[2], ctx->stream));
	if (launch_fxaa(ctx, reinterpret_cast<const unsigned *>(uintptr_t(frame_buffer_dev)), reinterpret_cast<unsigned char *>(uintptr_t(presented_dev)), row_begin, row_end)) return -1;
	if (device_ms) CHECK_CUDA(ctx, cudaEventRecord(ctx->ev[3], ctx->stream));
	if (!device_ms) return 0; // asynchronous
	CHECK_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
	cudaEventElapsedTime(device_ms, ctx->ev[2], ctx->ev[3]);
	return 0;
}

extern "C" int b200rt_set_band_targets(b200rt_ctx * ctx, const uint64_t * targets, int32_t band_count, int32_t band_rows, int32_t halo_rows) {
	if (!ctx) return -1;
	if (band_count == 0) { ctx->band_count = 0; return 0; }
	if (!targets || band_count < 1 || band_count > B200RT_MAX_BANDS) return fail(ctx, "band_count must be in 0..8");
	const int height = ctx->config.height;
	// the taps reach 4 rows along the edge direction, and the diagonal taps H/W pixels vertically (fragment_fxaa.glsl:25-28)
	const int reach = std::max(4, int(0.5f + float(height) / float(ctx->config.width)));
	if (halo_rows < reach) return fail(ctx, "halo_rows is smaller than the reach of fragment_fxaa.glsl (4 rows, more for a frame over 4.5 times as high as wide)");
	if (band_rows < halo_rows || (band_count - 1) * band_rows >= height || height - (band_count - 1) * band_rows < halo_rows) return fail(ctx, "bands must cover the frame and none may be shorter than halo_rows");
	for (int b = 0; b < band_count; b++) if (!targets[b]) return fail(ctx, "null band target");
	for (int b = 0; b < B200RT_MAX_BANDS; b++) ctx->band_targets[b] = b < band_count ? targets[b] : 0;
	ctx->band_count = band_count; ctx->band_rows = band_rows; ctx->band_halo = halo_rows;
	return 0;
}

// ------------------------------------------------------ peer memory + device barrier ---
extern "C" int b200rt_peer_alloc(b200rt_ctx * ctx, uint64_t bytes, uint64_t * device_ptr, b200rt_peer_handle * handle) {
	if (!ctx) return -1;
	if (!device_ptr || !handle || bytes == 0) return fail(ctx, "bad argument");
	static_assert(sizeof(cudaIpcMemHandle_t) == sizeof(b200rt_peer_handle), "b200rt_peer_handle is a cudaIpcMemHandle_t");
	CHECK_CUDA(ctx, cudaSetDevice(ctx->config.device));
	void * p = nullptr;
	CHECK_CUDA(ctx, cudaMalloc(&p, size_t(bytes)));
	cudaError_t e = cudaMemset(p, 0, size_t(bytes));
	cudaIpcMemHandle_t h;
	if (e == cudaSuccess) e = cudaIpcGetMemHandle(&h, p);
	if (e != cudaSuccess) { cudaFree(p); ctx->error = std::string("peer allocation: ") + cudaGetErrorString(e); return -1; }
	memcpy(handle->bytes, &h, sizeof(h));
	*device_ptr = uint64_t(reinterpret_cast<uintptr_t>(p));
	return 0;
}
extern "C" int b200rt_peer_free(b200rt_ctx * ctx, uint64_t device_ptr) {
	if (!ctx) return -1;
	CHECK_CUDA(ctx, cudaSetDevice(ctx->config.device));
	CHECK_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
	CHECK_CUDA(ctx, cudaFree(reinterpret_cast<void *>(uintptr_t(device_ptr))));
	return 0;
}
extern "C" int b200rt_peer_open(b200rt_ctx * ctx, const b200rt_peer_handle * handle, uint64_t * device_ptr) {
	if (!ctx) return -1;
	if (!handle || !device_ptr) return fail(ctx, "bad argument");
	CHECK_CUDA(ctx, cudaSetDevice(ctx->config.device));
	cudaIpcMemHandle_t h;
	memcpy(&h, handle->bytes, sizeof(h));
	void * p = nullptr;
	CHECK_CUDA(ctx, cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess));
	*device_ptr = uint64_t(reinterpret_cast<uintptr_t>(p));
	return 0;
}
extern "C" int b200rt_peer_close(b200rt_ctx * ctx, uint64_t device_ptr) {
	if (!ctx) return -1;
	CHECK_CUDA(ctx, cudaSetDevice(ctx->config.device));
	CHECK_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
	CHECK_CUDA(ctx, cudaIpcCloseMemHandle(reinterpret_cast<void *>(uintptr_t(device_ptr))));
	return 0;
}
extern "C" int b200rt_peer_read(b200rt_ctx * ctx, uint64_t device_ptr, void * host, uint64_t bytes) {
	if (!ctx) return -1;
	if (!device_ptr || !host) return fail(ctx, "null pointer");
	CHECK_CUDA(ctx, cudaSetDevice(ctx->config.device));
	CHECK_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
	CHECK_CUDA(ctx, cudaMemcpy(host, reinterpret_cast<const void *>(uintptr_t(device_ptr)), size_t(bytes), cudaMemcpyDeviceToHost));
	return 0;
}
extern "C" int b200rt_barrier_setup(b200rt_ctx * ctx, const uint64_t * flag_blocks, int32_t count, int32_t rank) {
	if (!ctx) return -1;
	if (!flag_blocks || count < 1 || count > B200RT_MAX_BANDS || rank < 0 || rank >= count) return fail(ctx, "barrier needs 1..8 flag blocks and a rank among them");
	for (int q = 0; q < count; q++) if (!flag_blocks[q]) return fail(ctx, "null flag block");
	for (int q = 0; q < B200RT_MAX_BANDS; q++) ctx->barrier_blocks[q] = q < count ? reinterpret_cast<unsigned *>(uintptr_t(flag_blocks[q])) : nullptr;
	ctx->barrier_count = count; ctx->barrier_rank = rank; ctx->barrier_epoch = 0;
	return 0;
}
extern "C" int b200rt_barrier(b200rt_ctx * ctx) {
	if (!ctx) return -1;
	if (ctx->barrier_count == 0) return fail(ctx, "b200rt_barrier_setup has not been called");
	CHECK_CUDA(ctx, cudaSetDevice(ctx->config.device));
	CHECK_CUDA(ctx, ctx->kernels->barrier(ctx->barrier_blocks, ctx->barrier_count, ctx->barrier_rank, ++ctx->barrier_epoch, ctx->stream));
	return 0;
}

// ------------------------------------------------------------- the entry point that allocates nothing ---
// Scratch of b200rt_render_resident_noalloc: one FrameCounters block (the unit counter the persistent warps pull work
// from, the ray counters, the kernel's time stamps) followed by the per-tile cost words the kernel adds to. File scope,
// fixed size: every GPU that loads the library holds these 64.5 KB until the process exits. Shared by every call in
// the process — the entry point clears it on the stream before each launch; calls that overlap on one GPU race.
static __device__ __align__(256) unsigned char g_noalloc_scratch[sizeof(FrameCounters) + size_t(B200RT_NOALLOC_MAX_TILES) * sizeof(unsigned)];

// In host code `g_noalloc_scratch` is only a handle for the runtime, not a device address: the address is asked for
// once per device, with that device current, and kept.
static int noalloc_scratch_address(b200rt_ctx * ctx, unsigned char ** out) {
	enum { MAX_DEVICES = 64 };
	static std::mutex lock;
	static unsigned char * address[MAX_DEVICES] = {};
	const int device = ctx->config.device;
	if (device < 0 || device >= MAX_DEVICES) return fail(ctx, "device ordinal beyond the no-alloc table");
	std::lock_guard<std::mutex> guard(lock);
	if (!address[device]) {
		void * p = nullptr;
		CHECK_CUDA(ctx, cudaGetSymbolAddress(&p, g_noalloc_scratch));
		if (!p) return fail(ctx, "cudaGetSymbolAddress returned no address");
		address[device] = static_cast<unsigned char *>(p);
	}
	*out = address[device];
	return 0;
}

extern "C" int b200rt_render_resident_noalloc(b200rt_ctx * ctx, uint64_t frame_buffer_dev) {
	if (!ctx) return -1;
	if (!frame_buffer_dev) return fail(ctx, "null device frame buffer");
	if (ctx->tile_count_mine > B200RT_NOALLOC_MAX_TILES)
		return fail(ctx, "b200rt_render_resident_noalloc: this context renders " + std::to_string(ctx->tile_count_mine) + " tiles, the static scratch holds B200RT_NOALLOC_MAX_TILES = " + std::to_string(int(B200RT_NOALLOC_MAX_TILES)));
	if (!ctx->have_frame) return fail(ctx, "no frame is resident: render one through b200rt_render_frame / b200rt_render_frame_device first");
	const unsigned total_warps = unsigned(ctx->params.tile_count_mine) * 32u;
	if (total_warps == 0) return 0;
	CHECK_CUDA(ctx, cudaSetDevice(ctx->config.device));
	unsigned char * scratch = nullptr;
	if (noalloc_scratch_address(ctx, &scratch)) return -1;

	// the scratch keeps the previous call's counts: the unit counter, the ray counters and the tile costs start at zero
	const size_t used = sizeof(FrameCounters) + size_t(ctx->params.tile_count_mine) * sizeof(unsigned);
	CHECK_CUDA(ctx, cudaMemsetAsync(scratch, 0, used, ctx->stream));

	SceneParams P = ctx->params;
	P.band_count = 0;                           // one target: the caller's image
	P.blas_stack_entries = ctx->blas_stack_entries; P.tlas_stack_entries = ctx->tlas_stack_entries;
	P.queue_mode = 0;                           // the depth-first walk: the queued one needs pools
	P.tile_cost = reinterpret_cast<unsigned *>(scratch + sizeof(FrameCounters));   // the context's own cost history and tile order stay as they are
	int & per_sm = ctx->blocks_per_sm[0][0];
	if (per_sm == 0) {
		CHECK_CUDA(ctx, ctx->kernels->render_occupancy(P, false, &per_sm));    // a query, no allocation
		if (per_sm < 1) per_sm = 1;
	}
	unsigned blocks = unsigned(ctx->num_sms) * unsigned(per_sm);
	const unsigned needed = (total_warps + 7) / 8;
	if (blocks > needed) blocks = needed;
	CHECK_CUDA(ctx, ctx->kernels->render(P, reinterpret_cast<unsigned *>(uintptr_t(frame_buffer_dev)), nullptr, reinterpret_cast<FrameCounters *>(scratch), 0, blocks, ctx->stream));
	return 0;
}

extern "C" int b200rt_time_resident(b200rt_ctx * ctx, int repeat, uint64_t flush_l2_bytes, float * ms_render, float * ms_fxaa) {
	if (!ctx) return -1;
	if (!ctx->have_frame) return fail(ctx, "no frame has been rendered yet");
	if (repeat < 1) repeat = 1;
	CHECK_CUDA(ctx, cudaSetDevice(ctx->config.device));
	unsigned * fb = static_cast<unsigned *>(ctx->frame_buffer.ptr);
	float total_render = 0.0f, total_fxaa = 0.0f;
	if (flush_l2_bytes) CHECK_CUDA(ctx, ctx->l2_scratch.ensure(size_t(flush_l2_bytes)));
	// AUTO's comparison frames are not what a caller wants timed: let it decide first (bounded; same frames, same L2 state)
	const int settle = ctx->config.deep_queue == B200RT_QUEUE_AUTO ? 16 : 0;
	for (int i = -settle; i < repeat; i++) {
		if (i < 0 && ctx->auto_queue.decided) { i = -1; continue; }
		if (flush_l2_bytes) CHECK_CUDA(ctx, cudaMemsetAsync(ctx->l2_scratch.ptr, i & 0xff, size_t(flush_l2_bytes), ctx->stream)); // evict L2, untimed
		CHECK_CUDA(ctx, cudaEventRecord(ctx->ev[1], ctx->stream));
		if (launch_render(ctx, fb, nullptr, false)) return -1;
		CHECK_CUDA(ctx, cudaEventRecord(ctx->ev[2], ctx->stream));
		if (launch_fxaa(ctx, fb, static_cast<unsigned char *>(ctx->presented.ptr))) return -1;
		if (launch_schedule(ctx)) return -1;
		CHECK_CUDA(ctx, cudaEventRecord(ctx->ev[3], ctx->stream));
		CHECK_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
		float a = 0.0f, b = 0.0f;
		cudaEventElapsedTime(&a, ctx->ev[1], ctx->ev[2]);
		cudaEventElapsedTime(&b, ctx->ev[2], ctx->ev[3]);
		if (i >= 0) { total_render += a; total_fxaa += b; }
	}
	if (ms_render) *ms_render = total_render / float(repeat);
	if (ms_fxaa)   *ms_fxaa   = total_fxaa / float(repeat);
	return 0;
}

// ------------------------------------------------------------------- pipelined frames ---
// The reference's presentation is asynchronous too: glTexImage2D + draw + SwapBuffers (Window.cpp:
// 87-119) only queue work for the graphics driver, and the CPU goes on to trace the next frame.
// Here: frame n's images travel to the host on a copy stream while frame n+1 is being traced.
static int ensure_slot(b200rt_ctx * ctx, b200rt_ctx::Slot & slot) {
	const size_t pixels = size_t(ctx->config.width) * ctx->config.height;
	if (slot.copied) return 0;
	CHECK_CUDA(ctx, slot.frame_buffer.ensure(pixels * 4));
	CHECK_CUDA(ctx, slot.presented.ensure(pixels * 3 + 16));
	CHECK_CUDA(ctx, cudaMemset(slot.frame_buffer.ptr, 0, pixels * 4));
	CHECK_CUDA(ctx, cudaMallocHost(reinterpret_cast<void **>(&slot.pinned_counters), sizeof(FrameCounters)));
	for (cudaEvent_t * e : { &slot.render_end, &slot.fxaa_end, &slot.copied }) CHECK_CUDA(ctx, cudaEventCreateWithFlags(e, cudaEventDisableTiming));
	if (!ctx->copy_stream) CHECK_CUDA(ctx, cudaStreamCreateWithFlags(&ctx->copy_stream, cudaStreamNonBlocking));
	return 0;
}

extern "C" int b200rt_submit_frame(b200rt_ctx * ctx, const b200rt_frame * frame, const b200rt_outputs * out) {
	if (!ctx) return -1;
	if (!out) return fail(ctx, "null outputs");
	if (out->colour_float || out->primary_ids || out->primary_hit) return fail(ctx, "b200rt_submit_frame delivers frame_buffer, presented_rgb8 and stats only");
	if ((out->frame_buffer && !is_page_locked(out->frame_buffer)) || (out->presented_rgb8 && !is_page_locked(out->presented_rgb8)))
		return fail(ctx, "b200rt_submit_frame needs page-locked output images (b200rt_host_alloc): the copies are asynchronous");
	CHECK_CUDA(ctx, cudaSetDevice(ctx->config.device));

	b200rt_ctx::Slot & slot = ctx->slots[ctx->submitted & 1];
	if (slot.busy) return fail(ctx, "two frames are in flight: call b200rt_wait_frame first");
	if (ensure_slot(ctx, slot)) return -1;

	const size_t pixels = size_t(ctx->config.width) * ctx->config.height;
	unsigned * fb = static_cast<unsigned *>(slot.frame_buffer.ptr);

	if (stage_frame(ctx, frame)) return -1;
	if (launch_render(ctx, fb, nullptr, ctx->config.count_traversal != 0)) return -1;
	CHECK_CUDA(ctx, cudaEventRecord(slot.render_end, ctx->stream));
	if (out->presented_rgb8 && launch_fxaa(ctx, fb, static_cast<unsigned char *>(slot.presented.ptr))) return -1;
	FrameCounters * counted = nullptr;
	if (launch_schedule(ctx, &counted)) return -1;
	CHECK_CUDA(ctx, cudaEventRecord(slot.fxaa_end, ctx->stream));
	if (release_frame_block(ctx)) return -1;

	// the pre-FXAA image can leave as soon as tracing is done, in parallel with the FXAA kernel
	CHECK_CUDA(ctx, cudaStreamWaitEvent(ctx->copy_stream, slot.render_end, 0));
	if (out->frame_buffer) CHECK_CUDA(ctx, cudaMemcpyAsync(out->frame_buffer, fb, pixels * 4, cudaMemcpyDeviceToHost, ctx->copy_stream));
	CHECK_CUDA(ctx, cudaMemcpyAsync(slot.pinned_counters, counted, sizeof(FrameCounters), cudaMemcpyDeviceToHost, ctx->copy_stream));   // this block is cleared three frames from now
	CHECK_CUDA(ctx, cudaStreamWaitEvent(ctx->copy_stream, slot.fxaa_end, 0));
	if (out->presented_rgb8) CHECK_CUDA(ctx, cudaMemcpyAsync(out->presented_rgb8, slot.presented.ptr, pixels * 3, cudaMemcpyDeviceToHost, ctx->copy_stream));
	CHECK_CUDA(ctx, cudaEventRecord(slot.copied, ctx->copy_stream));

	slot.busy = true;
	ctx->submitted++;
	return 0;
}

extern "C" int b200rt_wait_frame(b200rt_ctx * ctx, b200rt_stats * stats) {
	if (!ctx) return -1;
	b200rt_ctx::Slot & slot = ctx->slots[ctx->waited & 1];
	if (!slot.busy) return fail(ctx, "no frame in flight");
	CHECK_CUDA(ctx, cudaSetDevice(ctx->config.device));
	cudaError_t e = cudaEventSynchronize(slot.copied);
	slot.busy = false;
	ctx->waited++;
	if (e != cudaSuccess) { ctx->error = std::string("frame failed: ") + cudaGetErrorString(e); return -1; }
	if (queue_stalled(ctx, *slot.pinned_counters)) return fail(ctx, "the deep-pixel queue stalled in this frame: its image is incomplete; submit it again (the context has switched to the depth-first walk)");
	if (stats) {
		const FrameCounters & c = *slot.pinned_counters;
		counters_to_stats(c, stats);
		// no timing events on this path: the render kernel stamps its own start and end (globaltimer)
		if (c.first_start_ns != 0 && c.last_done_ns > ~c.first_start_ns) stats->device_ms_render = float(double(c.last_done_ns - ~c.first_start_ns) * 1e-6);
		stats->device_ms_total = stats->device_ms_render;
	}
	return 0;
}
