// Host-side scene: the reference's Scene / Mesh / Camera / Sphere / Plane / light surface
// (Scene.h:19-47, Mesh.h:8-26, Camera.h:11-33, Sphere.h, Plane.h, *Light.h) re-hosted in
// portable C++. It owns everything Scene::update (Scene.cpp:139-171) touches, rebuilds the
// TopLevelBVH every frame on the CPU as the reference does, and exposes the result as the
// flat b200rt_static_scene / b200rt_frame views the device path consumes.
#pragma once
#include <memory>
#include <string>
#include <unordered_map>
#include <vector>

#include "BVH.h"
#include "Math.h"
#include "Texture.h"

namespace b200rt {

struct Material { // Material.h:7-14
	Vector3 diffuse = Vector3(1.0f);
	int     texture = -1;
	Vector3 reflection;
	Vector3 transmittance;
	float   index_of_refraction = 1.0f;
};

struct Camera { // Camera.h:11-33
	Vector3    position;
	Quaternion rotation;
	float      fov = deg_to_rad(110.0f);

	Vector3 top_left_corner, x_axis, y_axis;                         // view pyramid, Camera::resize
	Vector3 rotated_top_left_corner, rotated_x_axis, rotated_y_axis; // Camera::update

	void resize(int width, int height);
	void update();
};

struct Mesh { // Mesh.h:8-26
	Transform transform;
	Matrix4   transform_inv;
	AABB      aabb;
	int       bvh = -1; // index into Scene::blas (shared between instances, BottomLevelBVH.cpp:16-22)

	void update(const BottomLevelBVH & blas);
};

struct Sphere { Transform transform; int material_id = 0; float radius_inv = 1.0f, radius_squared = 1.0f; };

struct Plane {
	Transform transform; int material_id = 0;
	Vector3 world_normal, u_axis, v_axis; float world_distance = 0.0f;
	void update(); // Plane.cpp:3-11
};

struct TopLevelBVH { // TopLevelBVH.h:6-33
	std::vector<Mesh>    primitives;
	std::vector<BVHNode> nodes;
	int                  node_count = 0;
	SAHBuilder           builder;   // persistent per-axis index arrays; builder.indices[0] = TopLevelBVH::indices

	void update(const std::vector<std::unique_ptr<BottomLevelBVH>> & blas);
	void build_bvh();
};

enum SceneAnimation { ANIMATION_NONE = 0, ANIMATION_DYNAMIC = 1 }; // Config.h SCENE

struct Scene {
	int width = 1280, height = 720;

	std::vector<Sphere> spheres;
	std::vector<Plane>  planes;
	TopLevelBVH         top_level_bvh;

	std::vector<b200rt_light> lights; // sorted: point, spot, directional
	Vector3 ambient_lighting = Vector3(0.2f);

	std::vector<float> sky; int sky_size = 0;
	Camera camera;

	std::vector<Material> materials;                     // MaterialBuffer::materials
	TexturePool           texture_pool;
	std::vector<std::unique_ptr<BottomLevelBVH>> blas;
	std::unordered_map<std::string, int> blas_cache;     // bvh_cache, BottomLevelBVH.cpp:16

	float time = 0.0f; // `static float time` of Scene.cpp:143

	// Reads a b200rt scene description (DESIGN.md "Scene description"). Throws std::runtime_error.
	void load(const std::string & scene_file);

	// Scene::update, Scene.cpp:139-171. `animation` selects the SCENE_DYNAMIC motion block.
	void update(float delta, SceneAnimation animation);

	// Flat views (valid until the next load/update).
	b200rt_static_scene static_view();
	b200rt_frame        frame_view();

	int load_blas(const std::string & obj_path); // BottomLevelBVH::load, BottomLevelBVH.cpp:18-58

private:
	int reserve_material() { materials.emplace_back(); return int(materials.size()) - 1; } // MaterialBuffer::reserve
	void load_mtl(BottomLevelBVH & bvh, const std::string & obj_path);
	void load_obj(const std::string & obj_path, std::vector<TriangleHot> & hot, std::vector<TriangleCold> & cold, std::vector<Vector3> & vertices, const std::vector<std::string> & material_names);

	// flat staging
	std::vector<BVHNode> flat_nodes; std::vector<TriangleHot> flat_hot; std::vector<TriangleCold> flat_cold;
	std::vector<b200rt_blas> flat_blas; std::vector<b200rt_material> flat_materials;
	std::vector<b200rt_instance> flat_instances; std::vector<b200rt_sphere> flat_spheres; std::vector<b200rt_plane> flat_planes;
	std::vector<std::string> last_material_names;
};

} // namespace b200rt
