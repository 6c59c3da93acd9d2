// TEST INFRASTRUCTURE — headless replacement for the reference's Main.cpp (which needs an
// SDL2/OpenGL/ImGui window). Same frame loop as Main.cpp:51-59 without the presentation:
//   scene.update(dt) -> wake worker threads over all tiles -> wait.
// Adds: scene description file input, raw dumps of the frame buffer, timing, and (in the
// ORACLE_IDS build) a primary-ray pass that records the closest-hit
// record per pixel through the reference's own Scene::trace_primitives.
//
// usage: ref_<variant> scene.txt --out PREFIX [--frames N] [--dt SECONDS] [--dump a,b,c]
//                      [--width W --height H] [--quiet] [--per-ray]
// --per-ray: instead of render_tile's 4x2 pixel packets (Raytracer.cpp:28-42) every pixel gets a
//   packet whose lanes all carry the SAME ray (primary ray expressions of Raytracer.cpp:45-59),
//   pushed through the reference's own Raytracer::bounce; lane 0 is plotted. Arithmetic is the
//   SIMD build's (fused madd, SVML->libmvec), but traversal order and every all/any early-out
//   depend on that one ray only — the semantics a one-ray-per-thread GPU renderer has.
// dumps (little-endian raw):  PREFIX.fNNNN.u32   uint32 [H][W]   0x00RRGGBB (Window::plot)
//                             PREFIX.fNNNN.rgbf  float  [H][W][3] un-quantised colour
//                             PREFIX.fNNNN.ids   int32  [H][W][2] (instance, triangle)   ORACLE_IDS
//                             PREFIX.fNNNN.hit   float  [H][W][4] (distance, material id, u, v)   ORACLE_IDS
// --bridge (B200RT_BRIDGE builds, "_bridge" variants): after the CPU has traced a frame, the SAME Scene object is handed
//   to libb200rt.so through bridge/B200Bridge.h — the binding a maintainer of the reference would add — and the GPU's
//   frame is dumped beside the CPU's:  PREFIX.fNNNN.gpu.u32 (Window::plot format), PREFIX.fNNNN.gpu.rgb8 (after FXAA).
#include <chrono>
#include <string>
#include <vector>
#include <algorithm>

#include <thread>
#include <atomic>
#include <map>
#include <unordered_map>
#include <filesystem>
#include <fstream>
#include <sstream>

#include "SceneFile.h"

#ifdef B200RT_BRIDGE
#define private public   // Texture, Sky, Sphere, Plane: see the head of B200Bridge.h
#endif
#include "Scene.h"
#ifdef B200RT_BRIDGE
#undef private
#include "B200Bridge.h"
#endif

// `Raytracer::bounce` is private (Raytracer.h:16-17). The per-ray mode below has to call it with a
// ray of its own, so the access specifier is lifted for this one header — layout and code of the
// reference class are untouched.
#define private public
#include "Raytracer.h"
#undef private
#include "WorkerThread.h"

#ifdef ORACLE_IDS
thread_local int g_oracle_current_instance = -1;
#endif

int WorkerThreads_thread_count();

static void write_file(const std::string & name, const void * data, size_t bytes) {
	FILE * f = fopen(name.c_str(), "wb");
	if (!f) { fprintf(stderr, "cannot write %s\n", name.c_str()); exit(2); }
	fwrite(data, 1, bytes, f);
	fclose(f);
}

// Primary ray for pixel (i, j), all lanes identical: the expressions of Raytracer.cpp:45-59.
static Ray primary_ray(const Scene & scene, int i, int j) {
	Ray ray;
	ray.origin = SIMD_Vector3(scene.camera.position);
#if RAY_DIFFERENTIALS_ENABLED
	ray.dO_dx  = SIMD_Vector3(0.0f);
	ray.dO_dy  = SIMD_Vector3(0.0f);
#endif
	SIMD_float is{float(i)};
	SIMD_float js{float(j)};

	SIMD_Vector3 direction = SIMD_Vector3::madd(scene.camera.rotated_x_axis, is,
		SIMD_Vector3::madd(scene.camera.rotated_y_axis, js, scene.camera.rotated_top_left_corner));

	SIMD_float d_dot_d  = SIMD_Vector3::dot(direction, direction);
	SIMD_float inv_sqrt = SIMD_float::inv_sqrt(d_dot_d);
	SIMD_float denom    = inv_sqrt / d_dot_d;
#if RAY_DIFFERENTIALS_ENABLED
	ray.dD_dx = (d_dot_d * scene.camera.rotated_x_axis - SIMD_Vector3::dot(direction, scene.camera.rotated_x_axis) * direction) * denom;
	ray.dD_dy = (d_dot_d * scene.camera.rotated_y_axis - SIMD_Vector3::dot(direction, scene.camera.rotated_y_axis) * direction) * denom;
#endif
	ray.direction = direction * inv_sqrt;
	return ray;
}

static PerformanceStats g_per_ray_stats = { 0, 0, 0, 0 };

// One identical-lane packet per pixel through the reference's bounce(); rows are handed out
// by an atomic counter to `threads` std::threads.
static void render_per_ray(const Raytracer & raytracer, const Scene & scene, const Window & window, int threads) {
	std::atomic<int> next_row(0);
	std::vector<PerformanceStats> stats(threads, PerformanceStats { 0, 0, 0, 0 });
	auto work = [&](int id) {
		for (int j = next_row.fetch_add(1); j < window.height; j = next_row.fetch_add(1)) {
			for (int i = 0; i < window.width; i++) {
				Ray ray = primary_ray(scene, i, j);
				stats[id].num_primary_rays++;
				SIMD_float distance;
				SIMD_Vector3 colour = raytracer.bounce(ray, NUMBER_OF_BOUNCES, distance, stats[id]);
				window.plot(i, j, Vector3(colour.x[0], colour.y[0], colour.z[0]));
			}
		}
	};
	std::vector<std::thread> pool;
	for (int t = 1; t < threads; t++) pool.emplace_back(work, t);
	work(0);
	for (auto & t : pool) t.join();

	g_per_ray_stats = PerformanceStats { 0, 0, 0, 0 };
	for (const auto & s : stats) {
		g_per_ray_stats.num_primary_rays    += s.num_primary_rays;
		g_per_ray_stats.num_shadow_rays     += s.num_shadow_rays;
		g_per_ray_stats.num_reflection_rays += s.num_reflection_rays;
		g_per_ray_stats.num_refraction_rays += s.num_refraction_rays;
	}
}

// What Scene::update produced for this frame, for checking the host library against the
// reference: PREFIX.fNNNN.frame = int32 instance_count, node_count; per instance float world[16],
// inverse[16], aabb[6]; BVHNode nodes[node_count]; int32 indices[instance_count]; float camera[12].
static void dump_frame_inputs(const Scene & scene, const std::string & prefix) {
	FILE * f = fopen((prefix + ".frame").c_str(), "wb");
	if (!f) return;
	const TopLevelBVH & tlas = scene.top_level_bvh;
	int counts[2] = { tlas.primitive_count, tlas.node_count };
	fwrite(counts, sizeof(int), 2, f);
	for (int i = 0; i < tlas.primitive_count; i++) {
		fwrite(tlas.primitives[i].transform.world_matrix.cells, sizeof(float), 16, f);
		fwrite(tlas.primitives[i].transform_inv.cells,          sizeof(float), 16, f);
		fwrite(&tlas.primitives[i].aabb, sizeof(float), 6, f);
	}
	fwrite(tlas.nodes,   sizeof(BVHNode), tlas.node_count, f);
	fwrite(tlas.indices, sizeof(int), tlas.primitive_count, f);
	float camera[12] = {
		scene.camera.position.x, scene.camera.position.y, scene.camera.position.z,
		scene.camera.rotated_top_left_corner.x[0], scene.camera.rotated_top_left_corner.y[0], scene.camera.rotated_top_left_corner.z[0],
		scene.camera.rotated_x_axis.x[0], scene.camera.rotated_x_axis.y[0], scene.camera.rotated_x_axis.z[0],
		scene.camera.rotated_y_axis.x[0], scene.camera.rotated_y_axis.y[0], scene.camera.rotated_y_axis.z[0] };
	fwrite(camera, sizeof(float), 12, f);
	fclose(f);
}

#ifdef ORACLE_IDS
// Primary-ray closest hits through the reference's own trace_primitives (identical-lane packet,
// lane 0 read back).
static void dump_primary_hits(const Scene & scene, int width, int height, const std::string & prefix) {
	std::vector<int>   ids(size_t(width) * height * 2);
	std::vector<float> hits(size_t(width) * height * 4);

	for (int j = 0; j < height; j++) {
		for (int i = 0; i < width; i++) {
			Ray ray = primary_ray(scene, i, j);

			RayHit hit;
			scene.trace_primitives(ray, hit);

			size_t p = size_t(i) + size_t(width) * j;
			bool is_hit = (SIMD_float::mask(hit.hit) & 1) != 0;

			int instance = -1, triangle = -1;
			if (is_hit) {
				instance = hit.instance_id[0];
				triangle = hit.triangle_id[0];
				if (triangle < 0) { // analytic primitive: identify through its unique material id
					for (int s = 0; s < scene.spheres.primitive_count; s++) if (scene.spheres[s].material_id == hit.material_id[0]) { instance = -2; triangle = s; }
					for (int s = 0; s < scene.planes.primitive_count;  s++) if (scene.planes[s].material_id  == hit.material_id[0]) { instance = -3; triangle = s; }
				}
			}
			ids[2*p]     = instance;
			ids[2*p + 1] = triangle;

			hits[4*p]     = hit.distance[0];
			hits[4*p + 1] = is_hit ? float(hit.material_id[0]) : -1.0f;
			hits[4*p + 2] = is_hit ? hit.u[0] : 0.0f;
			hits[4*p + 3] = is_hit ? hit.v[0] : 0.0f;
		}
	}
	write_file(prefix + ".ids", ids.data(),  ids.size()  * sizeof(int));
	write_file(prefix + ".hit", hits.data(), hits.size() * sizeof(float));
}
#endif

int main(int argc, char ** argv) {
	if (argc < 2) { fprintf(stderr, "usage: %s scene.txt --out PREFIX [--frames N] [--dt S] [--dump a,b] [--width W --height H]\n", argv[0]); return 2; }

	std::string out_prefix;
	int   frames = 1;
	float dt = 0.0f;
	int   width = -1, height = -1;
	bool  quiet = false;
	bool  per_ray = false;
	bool  bridge = false;
	std::vector<int> dump_frames;

	for (int a = 2; a < argc; a++) {
		std::string arg(argv[a]);
		auto next = [&]() -> const char * { if (a + 1 >= argc) { fprintf(stderr, "missing value for %s\n", arg.c_str()); exit(2); } return argv[++a]; };
		if      (arg == "--out")    out_prefix = next();
		else if (arg == "--frames") frames = atoi(next());
		else if (arg == "--dt")     dt = float(atof(next()));
		else if (arg == "--width")  width  = atoi(next());
		else if (arg == "--height") height = atoi(next());
		else if (arg == "--quiet")  quiet = true;
		else if (arg == "--per-ray") per_ray = true;
		else if (arg == "--bridge")  bridge = true;
		else if (arg == "--dump") {
			std::string list(next());
			size_t pos = 0;
			while (pos < list.size()) {
				size_t comma = list.find(',', pos);
				if (comma == std::string::npos) comma = list.size();
				dump_frames.push_back(atoi(list.substr(pos, comma - pos).c_str()));
				pos = comma + 1;
			}
		} else { fprintf(stderr, "unknown argument %s\n", arg.c_str()); return 2; }
	}

	g_scene_file.load(argv[1]);
	if (width  <= 0) width  = g_scene_file.width;
	if (height <= 0) height = g_scene_file.height;

	Window window(width, height);

	// Same initialisation order as Main.cpp:26-44
	Texture::init();
	MaterialBuffer::init();

	Scene * scene = new Scene(); // leaked on purpose, see SceneFromFile.cpp
	scene->camera.resize(width, height);

	Raytracer raytracer;
	raytracer.scene = scene;

	WorkerThreads::init(raytracer, window);

	std::vector<double> ms_total, ms_render, ms_bridge;
#ifdef B200RT_BRIDGE
	B200Bridge gpu;
	std::vector<unsigned> gpu_frame_buffer(size_t(width) * height);
	std::vector<unsigned char> gpu_presented(size_t(width) * height * 3);
	b200rt_stats gpu_stats;
	memset(&gpu_stats, 0, sizeof(gpu_stats));
	if (bridge) gpu.init(*scene, width, height);
#else
	if (bridge) { fprintf(stderr, "--bridge needs a _bridge build\n"); return 2; }
#endif
	PerformanceStats last_stats = { 0, 0, 0, 0 };

	for (int frame = 0; frame < frames; frame++) {
		auto t0 = std::chrono::steady_clock::now();

		scene->update(dt);

		auto t1 = std::chrono::steady_clock::now();

		if (per_ray) {
			render_per_ray(raytracer, *scene, window, WorkerThreads_thread_count());
		} else {
			WorkerThreads::wake_up_worker_threads(window.tile_count_x * window.tile_count_y);
			WorkerThreads::wait_on_worker_threads();
		}

		auto t2 = std::chrono::steady_clock::now();

		ms_total .push_back(std::chrono::duration<double, std::milli>(t2 - t0).count());
		ms_render.push_back(std::chrono::duration<double, std::milli>(t2 - t1).count());

		last_stats = per_ray ? g_per_ray_stats : WorkerThreads::sum_performance_stats();

#ifdef B200RT_BRIDGE
		if (bridge) {   // the same Scene, after the same Scene::update: Main.cpp:56-59 on the GPU
			auto b0 = std::chrono::steady_clock::now();
			gpu.render(*scene, gpu_frame_buffer.data(), gpu_presented.data(), &gpu_stats);
			ms_bridge.push_back(std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - b0).count());
		}
#endif

		if (!out_prefix.empty() && std::find(dump_frames.begin(), dump_frames.end(), frame) != dump_frames.end()) {
			char tag[32]; snprintf(tag, sizeof(tag), ".f%04d", frame);
			std::string prefix = out_prefix + tag;
			write_file(prefix + ".u32",  window.frame_buffer,       size_t(width) * height * sizeof(unsigned));
			write_file(prefix + ".rgbf", window.frame_buffer_float, size_t(width) * height * 3 * sizeof(float));
#ifdef ORACLE_IDS
			dump_primary_hits(*scene, width, height, prefix);
#endif
			dump_frame_inputs(*scene, prefix);
#ifdef B200RT_BRIDGE
			if (bridge) {
				write_file(prefix + ".gpu.u32",  gpu_frame_buffer.data(), gpu_frame_buffer.size() * sizeof(unsigned));
				write_file(prefix + ".gpu.rgb8", gpu_presented.data(),    gpu_presented.size());
			}
#endif
		}
	}

	if (!quiet) {
		printf("{\"impl\": \"reference\", \"mode\": \"%s\", \"lanes\": %d, \"bounces\": %d, \"threads\": %d, \"width\": %d, \"height\": %d, \"frames\": %d, ",
			per_ray ? "per-ray" : "packet", SIMD_LANE_SIZE, NUMBER_OF_BOUNCES, WorkerThreads_thread_count(), width, height, frames);
		printf("\"rays\": {\"primary\": %d, \"shadow\": %d, \"reflection\": %d, \"refraction\": %d}, ",
			last_stats.num_primary_rays, last_stats.num_shadow_rays, last_stats.num_reflection_rays, last_stats.num_refraction_rays);
		printf("\"ms_total\": [");
		for (size_t i = 0; i < ms_total.size(); i++) printf("%s%.4f", i ? ", " : "", ms_total[i]);
		printf("], \"ms_render\": [");
		for (size_t i = 0; i < ms_render.size(); i++) printf("%s%.4f", i ? ", " : "", ms_render[i]);
		printf("]");
#ifdef B200RT_BRIDGE
		if (bridge) {
			printf(", \"gpu_rays\": {\"primary\": %llu, \"shadow\": %llu, \"reflection\": %llu, \"refraction\": %llu}, \"ms_bridge\": [",
				(unsigned long long) gpu_stats.primary_rays, (unsigned long long) gpu_stats.shadow_rays, (unsigned long long) gpu_stats.reflection_rays, (unsigned long long) gpu_stats.refraction_rays);
			for (size_t i = 0; i < ms_bridge.size(); i++) printf("%s%.4f", i ? ", " : "", ms_bridge[i]);
			printf("]");
		}
#endif
		printf("}\n");
	}
	fflush(stdout);
	_Exit(0); // skip static destructors: Scene is leaked, lights were placement-new'd
}
