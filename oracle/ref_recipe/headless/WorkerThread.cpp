// TEST INFRASTRUCTURE — std::thread port of the reference's Win32 WorkerThread.cpp.
// Same scheme: `remaining` counts down atomically, each decrement hands out one 32x32
// tile id, row-major, highest id first, edge tiles clipped (WorkerThread.cpp:53-65);
// per-thread PerformanceStats summed and multiplied by the lane count
// (WorkerThread.cpp:131-148). Threads are created per frame here (the reference parks
// them on events); thread count comes from REF_THREADS or hardware_concurrency, and
// threads are pinned 1:1 to cores as the reference does (WorkerThread.cpp:37-47).
#include "WorkerThread.h"

#include <atomic>
#include <thread>
#include <vector>
#include <pthread.h>
#include <sched.h>

static int thread_count = 1;
static const Raytracer * the_raytracer;
static const Window    * the_window;
static std::vector<PerformanceStats> stats;
static std::vector<std::thread> threads;
static std::atomic<int> remaining;

int WorkerThreads_thread_count() { return thread_count; }

static void run_tiles(int thread_id) {
	const Window & window = *the_window;
	while (remaining.load(std::memory_order_relaxed) > 0) {
		int task = remaining.fetch_sub(1) - 1;
		if (task < 0) break;

		int x = (task % window.tile_count_x) * window.tile_width;
		int y = (task / window.tile_count_x) * window.tile_height;

		int w = x + window.tile_width  < window.width  ? window.tile_width  : window.width  - x;
		int h = y + window.tile_height < window.height ? window.tile_height : window.height - y;

		the_raytracer->render_tile(window, x, y, w, h, stats[thread_id]);
	}
}

void WorkerThreads::init(const Raytracer & raytracer, const Window & window) {
	the_raytracer = &raytracer;
	the_window    = &window;

	const char * env = getenv("REF_THREADS");
	thread_count = env ? atoi(env) : int(std::thread::hardware_concurrency());
	if (thread_count < 1) thread_count = 1;

	stats.assign(thread_count, PerformanceStats { 0, 0, 0, 0 });
}

void WorkerThreads::wake_up_worker_threads(int job_count) {
	remaining = job_count;
	for (auto & s : stats) s = PerformanceStats { 0, 0, 0, 0 };

	threads.clear();
	for (int i = 1; i < thread_count; i++) {
		threads.emplace_back(run_tiles, i);

		cpu_set_t set; CPU_ZERO(&set); CPU_SET(i % int(std::thread::hardware_concurrency()), &set);
		pthread_setaffinity_np(threads.back().native_handle(), sizeof(set), &set);
	}
	run_tiles(0);
}

void WorkerThreads::wait_on_worker_threads() {
	for (auto & t : threads) t.join();
	threads.clear();
}

PerformanceStats WorkerThreads::sum_performance_stats() {
	PerformanceStats result = { 0, 0, 0, 0 };
	for (const auto & s : stats) {
		result.num_primary_rays    += s.num_primary_rays;
		result.num_shadow_rays     += s.num_shadow_rays;
		result.num_reflection_rays += s.num_reflection_rays;
		result.num_refraction_rays += s.num_refraction_rays;
	}
	result.num_primary_rays    *= SIMD_LANE_SIZE;
	result.num_shadow_rays     *= SIMD_LANE_SIZE;
	result.num_reflection_rays *= SIMD_LANE_SIZE;
	result.num_refraction_rays *= SIMD_LANE_SIZE;
	return result;
}
