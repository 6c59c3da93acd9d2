// C ABI of the host library (include/b200rt_host.h) over the C++ Scene.
#include "../../include/b200rt_host.h"

#include <exception>
#include <string>
#include <vector>

#include "Scene.h"

struct b200rt_scene { b200rt::Scene scene; };

static thread_local std::string g_error;

template<typename F> static int guarded(F && f) {
	try { f(); return 0; }
	catch (const std::exception & e) { g_error = e.what(); return -1; }
	catch (...) { g_error = "unknown error"; return -1; }
}

extern "C" {

const char * b200rt_host_last_error(void) { return g_error.c_str(); }

int b200rt_scene_load(const char * scene_file, b200rt_scene ** out) {
	if (!scene_file || !out) { g_error = "null argument"; return -1; }
	*out = nullptr;
	b200rt_scene * s = nullptr;
	int rc = guarded([&] { s = new b200rt_scene(); s->scene.load(scene_file); });
	if (rc != 0) { delete s; return rc; }
	*out = s;
	return 0;
}

void b200rt_scene_free(b200rt_scene * scene) { delete scene; }

int b200rt_scene_resize(b200rt_scene * s, int width, int height) {
	if (!s || width <= 0 || height <= 0) { g_error = "bad argument"; return -1; }
	s->scene.width = width; s->scene.height = height;
	s->scene.camera.resize(width, height);
	return 0;
}

int b200rt_scene_resolution(const b200rt_scene * s, int * width, int * height) {
	if (!s) { g_error = "null scene"; return -1; }
	if (width) *width = s->scene.width;
	if (height) *height = s->scene.height;
	return 0;
}

int b200rt_scene_set_camera(b200rt_scene * s, const float position[3], const float rotation[4]) {
	if (!s) { g_error = "null scene"; return -1; }
	if (position) s->scene.camera.position = b200rt::Vector3(position[0], position[1], position[2]);
	if (rotation) s->scene.camera.rotation = b200rt::Quaternion(rotation[0], rotation[1], rotation[2], rotation[3]);
	return 0;
}

int b200rt_scene_update(b200rt_scene * s, float delta, int animation) {
	if (!s) { g_error = "null scene"; return -1; }
	return guarded([&] { s->scene.update(delta, animation ? b200rt::ANIMATION_DYNAMIC : b200rt::ANIMATION_NONE); });
}

int b200rt_scene_static_view(b200rt_scene * s, b200rt_static_scene * out) {
	if (!s || !out) { g_error = "null argument"; return -1; }
	return guarded([&] { *out = s->scene.static_view(); });
}

int b200rt_scene_frame_view(b200rt_scene * s, b200rt_frame * out) {
	if (!s || !out) { g_error = "null argument"; return -1; }
	return guarded([&] { *out = s->scene.frame_view(); });
}

int b200rt_scene_blas_original_indices(const b200rt_scene * s, int blas, int32_t * out, int capacity) {
	if (!s || blas < 0 || blas >= int(s->scene.blas.size())) { g_error = "bad blas index"; return -1; }
	const std::vector<int> & map = s->scene.blas[blas]->original_index;
	if (out) for (int i = 0; i < capacity && i < int(map.size()); i++) out[i] = map[i];
	return int(map.size());
}

int b200rt_build_sbvh(const float * vertices, int triangle_count, int threads, b200rt_bvh_node * nodes, int node_capacity, int32_t * indices, int index_capacity, int * node_count, int * index_count) {
	if (!vertices || triangle_count <= 0 || !node_count || !index_count) { g_error = "bad argument"; return -1; }
	return guarded([&] {
		std::vector<b200rt::BuildTriangle> triangles(triangle_count);
		for (int i = 0; i < triangle_count; i++) {
			for (int k = 0; k < 3; k++) triangles[i].vertex[k] = b200rt::Vector3(vertices[9 * i + 3 * k], vertices[9 * i + 3 * k + 1], vertices[9 * i + 3 * k + 2]);
			triangles[i].finish();
		}
		b200rt::SBVHResult r = b200rt::build_sbvh(triangles, threads);
		*node_count = int(r.nodes.size()); *index_count = int(r.indices.size());
		if (nodes)   for (int i = 0; i < node_capacity  && i < *node_count;  i++) nodes[i]   = r.nodes[i];
		if (indices) for (int i = 0; i < index_capacity && i < *index_count; i++) indices[i] = r.indices[i];
	});
}

int b200rt_bvh_to_container(const char * legacy_bvh, const char * container) {
	if (!legacy_bvh || !container) { g_error = "null argument"; return -1; }
	return guarded([&] { b200rt::BottomLevelBVH bvh; bvh.load_from_disk(legacy_bvh); bvh.save_container(container); });
}

int b200rt_bvh_container_to_legacy(const char * container, const char * legacy_bvh) {
	if (!legacy_bvh || !container) { g_error = "null argument"; return -1; }
	return guarded([&] { b200rt::BottomLevelBVH bvh; bvh.load_container(container); bvh.save_to_disk(legacy_bvh); });
}

int b200rt_tile_count(int width, int height, int * tiles_x, int * tiles_y) {
	// Window.cpp:11-12: 32x32 tiles, the last row / column may be partial
	int tx = (width + 31) / 32, ty = (height + 31) / 32;
	if (tiles_x) *tiles_x = tx;
	if (tiles_y) *tiles_y = ty;
	return tx * ty;
}

} // extern "C"
