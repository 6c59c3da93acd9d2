// TEST INFRASTRUCTURE — headless replacement for the reference's Main.cpp (which needs an
// SDL2/OpenGL/ImGui window). Same frame loop as Main.cpp:51-59 without the presentation:
//   scene.update(dt) -> wake worker threads over all tiles -> wait.
// Adds: scene description file input, raw dumps of the frame buffer, timing, and (in the
// ORACLE_IDS build, lane size 1 only) a primary-ray pass that records the closest-hit
// record per pixel through the reference's own Scene::trace_primitives.
//
// usage: ref_<variant> scene.txt --out PREFIX [--frames N] [--dt SECONDS] [--dump a,b,c]
//                      [--width W --height H] [--quiet]
// dumps (little-endian raw):  PREFIX.fNNNN.u32   uint32 [H][W]   0x00RRGGBB (Window::plot)
//                             PREFIX.fNNNN.rgbf  float  [H][W][3] un-quantised colour
//                             PREFIX.fNNNN.ids   int32  [H][W][2] (instance, triangle)   ORACLE_IDS
//                             PREFIX.fNNNN.hit   float  [H][W][4] (distance, material id, u, v)   ORACLE_IDS
#include <chrono>
#include <string>
#include <vector>
#include <algorithm>

#include "SceneFile.h"

#include "Raytracer.h"
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

#ifdef ORACLE_IDS
// Primary-ray closest hits through the reference's own trace_primitives. The ray is built
// with the same SIMD expressions as Raytracer.cpp:45-59 (lane size 1).
static void dump_primary_hits(const Scene & scene, int width, int height, const std::string & prefix) {
	static_assert(SIMD_LANE_SIZE == 1, "the id dump is a lane-1 build");

	std::vector<int>   ids(size_t(width) * height * 2);
	std::vector<float> hits(size_t(width) * height * 4);

	for (int j = 0; j < height; j++) {
		for (int i = 0; i < width; i++) {
			Ray ray;
			ray.origin = SIMD_Vector3(scene.camera.position);
			ray.dO_dx  = SIMD_Vector3(0.0f);
			ray.dO_dy  = SIMD_Vector3(0.0f);

			SIMD_float is{float(i)};
			SIMD_float js{float(j)};

			SIMD_Vector3 direction = SIMD_Vector3::madd(scene.camera.rotated_x_axis, is,
				SIMD_Vector3::madd(scene.camera.rotated_y_axis, js, scene.camera.rotated_top_left_corner));

			SIMD_float d_dot_d  = SIMD_Vector3::dot(direction, direction);
			SIMD_float inv_sqrt = SIMD_float::inv_sqrt(d_dot_d);
			SIMD_float denom    = inv_sqrt / d_dot_d;

			ray.dD_dx = (d_dot_d * scene.camera.rotated_x_axis - SIMD_Vector3::dot(direction, scene.camera.rotated_x_axis) * direction) * denom;
			ray.dD_dy = (d_dot_d * scene.camera.rotated_y_axis - SIMD_Vector3::dot(direction, scene.camera.rotated_y_axis) * direction) * denom;
			ray.direction = direction * inv_sqrt;

			RayHit hit;
			scene.trace_primitives(ray, hit);

			size_t p = size_t(i) + size_t(width) * j;
			bool is_hit = SIMD_float::mask(hit.hit) != 0;

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

	FILE * saved_stdout = nullptr; (void)saved_stdout;

	Scene * scene = new Scene(); // leaked on purpose, see SceneFromFile.cpp
	scene->camera.resize(width, height);

	Raytracer raytracer;
	raytracer.scene = scene;

	WorkerThreads::init(raytracer, window);

	std::vector<double> ms_total, ms_render;
	PerformanceStats last_stats = { 0, 0, 0, 0 };

	for (int frame = 0; frame < frames; frame++) {
		auto t0 = std::chrono::steady_clock::now();

		scene->update(dt);

		auto t1 = std::chrono::steady_clock::now();

		WorkerThreads::wake_up_worker_threads(window.tile_count_x * window.tile_count_y);
		WorkerThreads::wait_on_worker_threads();

		auto t2 = std::chrono::steady_clock::now();

		ms_total .push_back(std::chrono::duration<double, std::milli>(t2 - t0).count());
		ms_render.push_back(std::chrono::duration<double, std::milli>(t2 - t1).count());

		last_stats = WorkerThreads::sum_performance_stats();

		if (!out_prefix.empty() && std::find(dump_frames.begin(), dump_frames.end(), frame) != dump_frames.end()) {
			char tag[32]; snprintf(tag, sizeof(tag), ".f%04d", frame);
			std::string prefix = out_prefix + tag;
			write_file(prefix + ".u32",  window.frame_buffer,       size_t(width) * height * sizeof(unsigned));
			write_file(prefix + ".rgbf", window.frame_buffer_float, size_t(width) * height * 3 * sizeof(float));
#ifdef ORACLE_IDS
			dump_primary_hits(*scene, width, height, prefix);
#endif
		}
	}

	if (!quiet) {
		printf("{\"impl\": \"reference\", \"lanes\": %d, \"bounces\": %d, \"threads\": %d, \"width\": %d, \"height\": %d, \"frames\": %d, ",
			SIMD_LANE_SIZE, NUMBER_OF_BOUNCES, WorkerThreads_thread_count(), width, height, frames);
		printf("\"rays_packet_convention\": {\"primary\": %d, \"shadow\": %d, \"reflection\": %d, \"refraction\": %d}, ",
			last_stats.num_primary_rays, last_stats.num_shadow_rays, last_stats.num_reflection_rays, last_stats.num_refraction_rays);
		printf("\"ms_total\": [");
		for (size_t i = 0; i < ms_total.size(); i++) printf("%s%.4f", i ? ", " : "", ms_total[i]);
		printf("], \"ms_render\": [");
		for (size_t i = 0; i < ms_render.size(); i++) printf("%s%.4f", i ? ", " : "", ms_render[i]);
		printf("]}\n");
	}
	fflush(stdout);
	_Exit(0); // skip static destructors: Scene is leaked, lights were placement-new'd
}
