#include "Scene.h"

#include <cmath>
#include <cstdlib>
#include <limits>
#include <thread>

#include <algorithm>
#include <cstdio>
#include <fstream>
#include <sstream>
#include <stdexcept>
#include <sys/stat.h>

namespace b200rt {

static bool file_exists(const std::string & path) { struct stat st; return stat(path.c_str(), &st) == 0; }

static std::string directory_of(const std::string & path) {
	size_t slash = path.find_last_of('/');
	return slash == std::string::npos ? std::string("./") : path.substr(0, slash + 1);
}

static std::string resolve(const std::string & dir, const std::string & p) {
	if (p.empty() || p == "-") return "";
	return p[0] == '/' ? p : dir + p;
}

// ------------------------------------------------------------------------------ Camera ---
void Camera::resize(int width, int height) { // Camera.cpp:5-16
	float half_width  = 0.5f * width;
	float half_height = 0.5f * height;
	float d = half_width / tanf(0.5f * fov); // distance to the viewing plane

	top_left_corner = Vector3(-half_width, half_height, d);
	x_axis = Vector3(1.0f,  0.0f, 0.0f);
	y_axis = Vector3(0.0f, -1.0f, 0.0f);
}

void Camera::update() { // Camera.cpp:44-47 (no keyboard in a headless host)
	rotated_top_left_corner = rotation * top_left_corner;
	rotated_x_axis          = rotation * x_axis;
	rotated_y_axis          = rotation * y_axis;
}

// ------------------------------------------------------------------------- primitives ---
void Mesh::update(const BottomLevelBVH & blas) { // Mesh.cpp:9-15
	transform.calc_world_matrix();

	const BVHNode & root = blas.nodes[0];
	AABB object_box;
	object_box.min = Vector3(root.aabb_min[0], root.aabb_min[1], root.aabb_min[2]);
	object_box.max = Vector3(root.aabb_max[0], root.aabb_max[1], root.aabb_max[2]);
	aabb = AABB::transform(object_box, transform.world_matrix);

	transform_inv = Matrix4::invert(transform.world_matrix);
}

void Plane::update() { // Plane.cpp:3-11
	transform.calc_world_matrix();

	world_normal   =  Matrix4::transform_direction(transform.world_matrix, Vector3(0.0f, 1.0f, 0.0f));
	world_distance = -dot(world_normal, transform.position);

	u_axis = Matrix4::transform_direction(transform.world_matrix, Vector3(1.0f, 0.0f, 0.0f));
	v_axis = cross(u_axis, world_normal);
}

void TopLevelBVH::update(const std::vector<std::unique_ptr<BottomLevelBVH>> & blas) { // TopLevelBVH.cpp:47-51
	for (Mesh & mesh : primitives) mesh.update(*blas[mesh.bvh]);
}

void TopLevelBVH::build_bvh() { // TopLevelBVH.cpp:32-45
	std::vector<BuildPrimitive> build(primitives.size());
	for (size_t i = 0; i < primitives.size(); i++) {
		build[i].aabb     = primitives[i].aabb;
		build[i].position = primitives[i].transform.position; // Mesh::get_position
	}
	node_count = builder.build(build, nodes);
}

// ----------------------------------------------------------------------- OBJ / MTL ---
namespace {
struct MtlEntry { std::string name; Material material; std::string texture; };

std::vector<MtlEntry> parse_mtl(const std::string & path) {
	std::vector<MtlEntry> entries;
	std::ifstream in(path);
	if (!in) return entries;

	std::string line;
	while (std::getline(in, line)) {
		while (!line.empty() && (line.back() == '\r' || line.back() == ' ' || line.back() == '\t')) line.pop_back();
		std::istringstream ss(line);
		std::string key;
		if (!(ss >> key) || key[0] == '#') continue;

		if (key == "newmtl") {
			MtlEntry e;
			ss >> e.name;
			// tinyobj's InitMaterial zeroes diffuse/specular/transmittance and sets ior = 1
			e.material.diffuse = Vector3(0.0f);
			entries.push_back(e);
			continue;
		}
		if (entries.empty()) continue;
		Material & m = entries.back().material;

		auto read3 = [&](Vector3 & v) { float x = 0, y = 0, z = 0; ss >> x; if (!(ss >> y)) { y = 0; z = 0; } else if (!(ss >> z)) z = 0; v = Vector3(x, y, z); };

		// OBJLoader.cpp:19-29: Kd -> diffuse, Ks -> reflection (!), Kt/Tf -> transmittance, Ni -> ior
		if      (key == "Kd") read3(m.diffuse);
		else if (key == "Ks") read3(m.reflection);
		else if (key == "Kt" || key == "Tf") read3(m.transmittance);
		else if (key == "Ni") ss >> m.index_of_refraction;
		else if (key == "map_Kd") {
			std::string token, name;
			while (ss >> token) name = token; // options (-bm ...) precede the file name
			std::replace(name.begin(), name.end(), '\\', '/'); // sponza.mtl uses Windows separators
			entries.back().texture = name;
		}
	}
	return entries;
}
}

// load_materials, OBJLoader.cpp:8-41: appends the mesh's materials to the global table
void Scene::load_mtl(BottomLevelBVH & bvh, const std::string & obj_path) {
	bvh.material_offset = int(materials.size());

	std::string mtl_path = obj_path.substr(0, obj_path.size() - 4) + ".mtl";
	std::vector<MtlEntry> entries = parse_mtl(mtl_path);

	last_material_names.clear();
	if (entries.empty()) { // OBJLoader.cpp:32-38: a mesh without materials is magenta
		Material m; m.diffuse = Vector3(1.0f, 0.0f, 1.0f);
		materials.push_back(m);
		return;
	}
	const std::string dir = directory_of(obj_path);
	for (MtlEntry & e : entries) {
		if (!e.texture.empty()) e.material.texture = texture_pool.load(dir + e.texture);
		materials.push_back(e.material);
		last_material_names.push_back(e.name);
	}
	if (materials.size() > 256) throw std::runtime_error("Max Material limit reached (MAX_MATERIALS 256, Config.h:18)");
}

namespace {
struct Corner { int v, t, n; };

// Crossing-number test of (x, y) against a triangle, the way tinyobjloader's pnpoly decides it
// (tiny_obj_loader.h:1316-1326): same edge order, same float expression.
bool inside_triangle(const float tx[3], const float ty[3], float x, float y) {
	bool inside = false;
	for (int i = 0, j = 2; i < 3; j = i++) {
		if ((ty[i] > y) != (ty[j] > y) && x < (tx[j] - tx[i]) * (y - ty[i]) / (ty[j] - ty[i]) + tx[i]) inside = !inside;
	}
	return inside;
}

// The reference loads meshes through tinyobjloader 2.0.0 with triangulate = true (OBJLoader.cpp:97),
// which does NOT fan polygons: it ear-clips them in a 2-D projection (tiny_obj_loader.h:1355-1557).
// The triangle order and the corner rotation of every triangle feed the BVH builder and the
// barycentric interpolation, so they are reproduced exactly:
//   * projection axes from the first corner whose cross product is above FLT_EPSILON — drop x if it
//     holds the strictly largest component, else drop y unless z is strictly largest, then drop z;
//   * signed area of the projected polygon by the shoelace sum;
//   * walk a candidate ear (g, g+1, g+2): skipped when it turns against the polygon or when another
//     remaining corner lies inside it; otherwise emitted and corner g+1 removed, g staying put;
//   * a polygon that cannot be clipped any further in `remaining` tries is abandoned; whatever is
//     left is emitted only when exactly three corners remain.
void ear_clip(const std::vector<Corner> & face, const std::vector<Vector3> & positions, std::vector<Corner> & triangles) {
	const size_t corners = face.size();
	if (corners < 3) return;
	if (corners == 3) { triangles.insert(triangles.end(), face.begin(), face.end()); return; }

	int axis_u = 1, axis_v = 2;
	for (size_t k = 0; k < corners; k++) {
		const Vector3 & a = positions[face[k].v], & b = positions[face[(k + 1) % corners].v], & c = positions[face[(k + 2) % corners].v];
		const float e0[3] = { b.x - a.x, b.y - a.y, b.z - a.z };
		const float e1[3] = { c.x - b.x, c.y - b.y, c.z - b.z };
		const float cx = fabsf(e0[1] * e1[2] - e0[2] * e1[1]);
		const float cy = fabsf(e0[2] * e1[0] - e0[0] * e1[2]);
		const float cz = fabsf(e0[0] * e1[1] - e0[1] * e1[0]);
		const float eps = std::numeric_limits<float>::epsilon();
		if (cx > eps || cy > eps || cz > eps) {
			if (!(cx > cy && cx > cz)) {
				axis_u = 0;
				if (cz > cx && cz > cy) axis_v = 1;
			}
			break;
		}
	}

	float area = 0.0f;
	for (size_t k = 0; k < corners; k++) {
		const Vector3 & a = positions[face[k].v], & b = positions[face[(k + 1) % corners].v];
		area += (a[axis_u] * b[axis_v] - a[axis_v] * b[axis_u]) * 0.5f;
	}

	std::vector<Corner> left = face;
	size_t guess = 0, tries = corners, previous = corners;
	while (left.size() > 3 && tries > 0) {
		const size_t n = left.size();
		if (guess >= n) guess -= n;
		if (previous != n) { previous = n; tries = n; } else tries--;

		Corner ear[3]; float tx[3], ty[3];
		for (int k = 0; k < 3; k++) {
			ear[k] = left[(guess + k) % n];
			tx[k] = positions[ear[k].v][axis_u]; ty[k] = positions[ear[k].v][axis_v];
		}
		const float turn = (tx[1] - tx[0]) * (ty[2] - ty[1]) - (ty[1] - ty[0]) * (tx[2] - tx[1]);
		if (turn * area < 0.0f) { guess++; continue; }

		bool blocked = false;
		for (size_t other = 3; other < n && !blocked; other++) {
			const Vector3 & p = positions[left[(guess + other) % n].v];
			blocked = inside_triangle(tx, ty, p[axis_u], p[axis_v]);
		}
		if (blocked) { guess++; continue; }

		triangles.insert(triangles.end(), ear, ear + 3);
		left.erase(left.begin() + (guess + 1) % n);
	}
	if (left.size() == 3) triangles.insert(triangles.end(), left.begin(), left.end());
}
}

// Triangle soup from a Wavefront OBJ, polygons ear-clipped like tinyobjloader does (above). Needs
// `vn` (the reference indexes normals unconditionally, OBJLoader.cpp:143-145); `vt` is optional and
// V is flipped to 1 - v (OBJLoader.cpp:139-141).
void Scene::load_obj(const std::string & obj_path, std::vector<TriangleHot> & hot, std::vector<TriangleCold> & cold, std::vector<Vector3> & vertices, const std::vector<std::string> & material_names) {
	std::ifstream in(obj_path);
	if (!in) throw std::runtime_error("Failed to load obj file " + obj_path);

	std::vector<Vector3> positions, normals;
	std::vector<float> texcoords;
	int material_id = 0;

	std::string line;
	while (std::getline(in, line)) {
		std::istringstream ss(line);
		std::string key;
		if (!(ss >> key)) continue;
		if (key == "v")  { Vector3 p; ss >> p.x >> p.y >> p.z; positions.push_back(p); }
		else if (key == "vn") { Vector3 n; ss >> n.x >> n.y >> n.z; normals.push_back(n); }
		else if (key == "vt") { float u = 0, v = 0; ss >> u >> v; texcoords.push_back(u); texcoords.push_back(v); }
		else if (key == "usemtl") {
			std::string name; ss >> name;
			auto it = std::find(material_names.begin(), material_names.end(), name);
			material_id = it == material_names.end() ? 0 : int(it - material_names.begin());
		} else if (key == "f") {
			std::vector<Corner> corners;
			std::string token;
			while (ss >> token) {
				Corner c = { 0, 0, 0 };
				int field = 0; size_t start = 0;
				for (size_t i = 0; i <= token.size(); i++) {
					if (i == token.size() || token[i] == '/') {
						if (i > start) {
							int value = atoi(token.substr(start, i - start).c_str());
							if (field == 0) c.v = value; else if (field == 1) c.t = value; else c.n = value;
						}
						field++; start = i + 1;
					}
				}
				auto fix = [](int index, size_t count) { return index > 0 ? index - 1 : (index < 0 ? int(count) + index : -1); };
				c.v = fix(c.v, positions.size()); c.t = fix(c.t, texcoords.size() / 2); c.n = fix(c.n, normals.size());
				corners.push_back(c);
			}
			for (const Corner & c : corners) if (c.v < 0 || c.v >= int(positions.size()) || c.n < 0 || c.n >= int(normals.size())) throw std::runtime_error("OBJ face needs valid v and vn indices: " + obj_path);
			std::vector<Corner> triangles;
			ear_clip(corners, positions, triangles);
			for (size_t k = 0; k + 2 < triangles.size(); k += 3) {
				const Corner c[3] = { triangles[k], triangles[k + 1], triangles[k + 2] };

				TriangleHot  h;
				TriangleCold t;
				Vector3 p[3], n[3]; float uv[3][2];
				for (int i = 0; i < 3; i++) {
					p[i] = positions[c[i].v]; n[i] = normals[c[i].n];
					if (c[i].t >= 0 && c[i].t < int(texcoords.size() / 2)) { uv[i][0] = texcoords[2 * c[i].t]; uv[i][1] = 1.0f - texcoords[2 * c[i].t + 1]; }
					else { uv[i][0] = 0.0f; uv[i][1] = 0.0f; }
				}
				for (int d = 0; d < 3; d++) {
					h.position_0[d] = p[0][d]; h.position_edge_1[d] = p[1][d] - p[0][d]; h.position_edge_2[d] = p[2][d] - p[0][d];
					t.normal_0[d]   = n[0][d]; t.normal_edge_1[d]   = n[1][d] - n[0][d]; t.normal_edge_2[d]   = n[2][d] - n[0][d];
				}
				for (int d = 0; d < 2; d++) {
					t.tex_coord_0[d] = uv[0][d]; t.tex_coord_edge_1[d] = uv[1][d] - uv[0][d]; t.tex_coord_edge_2[d] = uv[2][d] - uv[0][d];
				}
				t.material_id = material_id;
				hot.push_back(h); cold.push_back(t);
				vertices.push_back(p[0]); vertices.push_back(p[1]); vertices.push_back(p[2]);
			}
		}
	}
	if (hot.empty()) throw std::runtime_error("OBJ has no faces: " + obj_path);
}

// BottomLevelBVH::load, BottomLevelBVH.cpp:18-58: one BLAS per file name; `<obj>.bvh` is used
// when present (then only the .mtl is read), else the OBJ is parsed, a BVH built and cached.
int Scene::load_blas(const std::string & obj_path) {
	auto it = blas_cache.find(obj_path);
	if (it != blas_cache.end()) return it->second;

	auto bvh = std::make_unique<BottomLevelBVH>();
	const std::string bvh_path = obj_path + ".bvh";            // the reference's dump: what its binary reads and writes
	const std::string container_path = obj_path + ".b200bvh";   // the same payload behind a header with a checksum

	bool loaded = false;
	if (file_exists(container_path)) {
		try { bvh->load_container(container_path); loaded = true; }
		catch (const std::exception & e) { fprintf(stderr, "b200rt: ignoring %s (%s)\n", container_path.c_str(), e.what()); *bvh = BottomLevelBVH(); }
	}
	if (!loaded && file_exists(bvh_path)) {
		bvh->load_from_disk(bvh_path);
		loaded = true;
		try { bvh->save_container(container_path); } catch (const std::exception &) { }
	}
	if (loaded) {
		load_mtl(*bvh, obj_path);
	} else {
		load_mtl(*bvh, obj_path);
		std::vector<TriangleHot> hot; std::vector<TriangleCold> cold; std::vector<Vector3> vertices;
		load_obj(obj_path, hot, cold, vertices, last_material_names);
		// MESH_ACCELERATOR (Config.h:32-35): the reference's default is the SBVH; B200RT_MESH_ACCELERATOR=bvh selects the plain SAH BVH
		const char * accel = getenv("B200RT_MESH_ACCELERATOR");
		if (accel && std::string(accel) == "bvh") bvh->build(std::move(hot), std::move(cold));
		else bvh->build_sbvh(std::move(hot), std::move(cold), vertices, int(std::max(1u, std::thread::hardware_concurrency())));
		try { bvh->save_to_disk(bvh_path); bvh->save_container(container_path); } catch (const std::exception &) { /* read-only asset dir: keep going uncached */ }
	}

	blas.push_back(std::move(bvh));
	blas_cache[obj_path] = int(blas.size()) - 1;
	return int(blas.size()) - 1;
}

// ------------------------------------------------------------------------ scene file ---
namespace {
struct MaterialDesc { Vector3 diffuse = Vector3(1.0f), reflection, transmittance; float ior = 1.0f; std::string texture; };
struct SphereDesc   { float radius = 1.0f; Vector3 position; MaterialDesc material; };
struct PlaneDesc    { Vector3 position; Quaternion rotation; MaterialDesc material; };
struct MeshDesc     { std::string path; Vector3 position; Quaternion rotation; };

void read_material(std::istringstream & ss, MaterialDesc & m, const std::string & dir) {
	std::string texture;
	ss >> m.diffuse.x >> m.diffuse.y >> m.diffuse.z
	   >> m.reflection.x >> m.reflection.y >> m.reflection.z
	   >> m.transmittance.x >> m.transmittance.y >> m.transmittance.z
	   >> m.ior >> texture;
	m.texture = resolve(dir, texture);
}
}

void Scene::load(const std::string & scene_file) {
	std::ifstream in(scene_file);
	if (!in) throw std::runtime_error("cannot open scene file " + scene_file);
	const std::string dir = directory_of(scene_file);

	std::vector<SphereDesc> sphere_descs; std::vector<PlaneDesc> plane_descs; std::vector<MeshDesc> mesh_descs;
	std::vector<b200rt_light> parsed_lights;
	std::string sky_path;
	float fov_deg = 110.0f;

	std::string line;
	while (std::getline(in, line)) {
		std::istringstream ss(line);
		std::string key;
		if (!(ss >> key) || key[0] == '#') continue;

		if (key == "b200rt_scene") { int version; ss >> version; }
		else if (key == "resolution") ss >> width >> height;
		else if (key == "camera") ss >> fov_deg >> camera.position.x >> camera.position.y >> camera.position.z >> camera.rotation.x >> camera.rotation.y >> camera.rotation.z >> camera.rotation.w;
		else if (key == "ambient") ss >> ambient_lighting.x >> ambient_lighting.y >> ambient_lighting.z;
		else if (key == "sky") { std::string p; ss >> p; sky_path = resolve(dir, p); }
		else if (key == "sphere") { SphereDesc s; ss >> s.radius >> s.position.x >> s.position.y >> s.position.z; read_material(ss, s.material, dir); sphere_descs.push_back(s); }
		else if (key == "plane")  { PlaneDesc p; ss >> p.position.x >> p.position.y >> p.position.z >> p.rotation.x >> p.rotation.y >> p.rotation.z >> p.rotation.w; read_material(ss, p.material, dir); plane_descs.push_back(p); }
		else if (key == "mesh")   { MeshDesc m; std::string p; ss >> p >> m.position.x >> m.position.y >> m.position.z >> m.rotation.x >> m.rotation.y >> m.rotation.z >> m.rotation.w; m.path = resolve(dir, p); mesh_descs.push_back(m); }
		else if (key == "point_light" || key == "spot_light" || key == "directional_light") {
			b200rt_light l = {};
			ss >> l.colour[0] >> l.colour[1] >> l.colour[2];
			if (key != "directional_light") ss >> l.position[0] >> l.position[1] >> l.position[2];
			if (key != "point_light") {
				float d[3]; ss >> d[0] >> d[1] >> d[2];
				for (int i = 0; i < 3; i++) l.negative_direction[i] = -d[i]; // SpotLight.h:12, DirectionalLight.h:7
			}
			l.kind = key == "point_light" ? 0 : (key == "spot_light" ? 1 : 2);
			if (l.kind == 1) { // SpotLight.h:13-14: cosines of the HALF angles, given in degrees
				float inner, outer; ss >> inner >> outer;
				l.inner_cutoff = cosf(deg_to_rad(0.5f * inner));
				l.outer_cutoff = cosf(deg_to_rad(0.5f * outer));
			}
			parsed_lights.push_back(l);
		} else throw std::runtime_error("scene file " + scene_file + ": unknown key '" + key + "'");

		if (ss.fail()) throw std::runtime_error("scene file " + scene_file + ": bad line '" + line + "'");
	}
	if (mesh_descs.empty()) throw std::runtime_error("scene file " + scene_file + ": at least one mesh is required (TopLevelBVH.cpp:6)");

	// Construction order of the reference (Main.cpp:26-27, Scene.h:20-23, Primitive.h:6): material 0 is
	// the black default, then one reserved material per sphere, per plane, then the meshes' materials.
	materials.clear();
	Material default_material; default_material.diffuse = Vector3(0.0f); // MaterialBuffer::init, Material.h:52-60
	materials.push_back(default_material);

	spheres.resize(sphere_descs.size());
	planes .resize(plane_descs.size());
	for (Sphere & s : spheres) s.material_id = reserve_material();
	for (Plane  & p : planes)  p.material_id = reserve_material();

	auto apply = [&](int id, const MaterialDesc & d) {
		Material & m = materials[id];
		m.diffuse = d.diffuse; m.reflection = d.reflection; m.transmittance = d.transmittance; m.index_of_refraction = d.ior;
		if (!d.texture.empty()) m.texture = texture_pool.load(d.texture);
	};
	for (size_t i = 0; i < spheres.size(); i++) {
		spheres[i].radius_inv     = 1.0f / sphere_descs[i].radius; // Sphere::init, Sphere.h:13-16
		spheres[i].radius_squared = sphere_descs[i].radius * sphere_descs[i].radius;
		spheres[i].transform.position = sphere_descs[i].position;
		apply(spheres[i].material_id, sphere_descs[i].material);
	}
	for (size_t i = 0; i < planes.size(); i++) {
		planes[i].transform.position = plane_descs[i].position;
		planes[i].transform.rotation = plane_descs[i].rotation;
		apply(planes[i].material_id, plane_descs[i].material);
	}

	top_level_bvh.primitives.resize(mesh_descs.size());
	top_level_bvh.builder.init(int(mesh_descs.size()));
	for (size_t i = 0; i < mesh_descs.size(); i++) {
		Mesh & mesh = top_level_bvh.primitives[i];
		mesh.transform.position = mesh_descs[i].position;
		mesh.transform.rotation = mesh_descs[i].rotation;
		mesh.bvh = load_blas(mesh_descs[i].path);
	}

	lights.clear();
	for (int kind = 0; kind < 3; kind++) for (const b200rt_light & l : parsed_lights) if (l.kind == kind) lights.push_back(l);

	// Sky, Sky.cpp:8-26: square raw float RGB probe. One texel of padding is appended because the
	// reference clamps the texel index to [0, size*size] INCLUSIVE (Sky.cpp:44-45) and so can read
	// one element past the end; the pad repeats the last texel.
	sky.clear(); sky_size = 0;
	if (!sky_path.empty()) {
		FILE * f = fopen(sky_path.c_str(), "rb");
		if (!f) throw std::runtime_error("cannot open sky probe " + sky_path);
		fseek(f, 0, SEEK_END);
		long bytes = ftell(f);
		rewind(f);
		int size_squared = int(bytes / long(3 * sizeof(float)));
		sky_size = int(sqrtf(float(size_squared)));
		if (sky_size * sky_size != size_squared || size_squared == 0) { fclose(f); throw std::runtime_error("sky probe is not square: " + sky_path); }
		sky.resize(size_t(size_squared + 1) * 3);
		size_t got = fread(sky.data(), 3 * sizeof(float), size_squared, f);
		fclose(f);
		if (got != size_t(size_squared)) throw std::runtime_error("short read on sky probe " + sky_path);
		for (int c = 0; c < 3; c++) sky[size_t(size_squared) * 3 + c] = sky[size_t(size_squared - 1) * 3 + c];
	} else throw std::runtime_error("scene file " + scene_file + ": a sky probe is required (Sky.cpp:11)");

	camera.fov = deg_to_rad(fov_deg);
	camera.resize(width, height);
	time = 0.0f;
}

// ---------------------------------------------------------------------- per-frame prep ---
void Scene::update(float delta, SceneAnimation animation) {
	if (animation == ANIMATION_DYNAMIC) { // Scene.cpp:141-155
		std::vector<Mesh> & p = top_level_bvh.primitives;
		if (p.size() < 6) throw std::runtime_error("the dynamic animation drives 6 mesh instances (Scene.cpp:141-155)");
		const Vector3 X(1.0f, 0.0f, 0.0f), Y(0.0f, 1.0f, 0.0f);

		p[0].transform.rotation = Quaternion::axis_angle(Y, delta) * p[0].transform.rotation;

		time += delta;

		p[1].transform.position.y = 1.0f + 2.0f * sinf(time);
		p[2].transform.position.x -= delta * 0.5f;
		p[3].transform.position = Vector3(6.0f, 4.0f + 2.0f * sinf(time * 0.5f), 4.0f + 2.0f * cosf(time * 0.5f));
		p[3].transform.rotation = Quaternion::axis_angle(Y, delta * 0.5f) * p[3].transform.rotation;
		p[4].transform.rotation = Quaternion::axis_angle(X, delta) * p[4].transform.rotation;
		p[5].transform.rotation = Quaternion::nlerp(Quaternion(), Quaternion::axis_angle(X, deg_to_rad(-90.0f)), 0.5f + 0.5f * sinf(time));
	}

	camera.update();                                      // Scene.cpp:165
	for (Sphere & s : spheres) s.transform.calc_world_matrix(); // Sphere.cpp:5-7
	for (Plane  & p : planes)  p.update();
	top_level_bvh.update(blas);                           // Scene.cpp:169
	top_level_bvh.build_bvh();                            // Scene.cpp:170 — rebuilt EVERY frame
}

// ------------------------------------------------------------------------- flat views ---
b200rt_static_scene Scene::static_view() {
	flat_nodes.clear(); flat_hot.clear(); flat_cold.clear(); flat_blas.clear(); flat_materials.clear();

	for (const auto & b : blas) {
		b200rt_blas d;
		d.node_offset = int(flat_nodes.size());   d.node_count = int(b->nodes.size());
		d.triangle_offset = int(flat_hot.size()); d.triangle_count = int(b->triangles_hot.size());
		d.material_offset = b->material_offset;
		flat_blas.push_back(d);
		flat_nodes.insert(flat_nodes.end(), b->nodes.begin(), b->nodes.end());
		flat_hot  .insert(flat_hot.end(),   b->triangles_hot.begin(),  b->triangles_hot.end());
		flat_cold .insert(flat_cold.end(),  b->triangles_cold.begin(), b->triangles_cold.end());
	}
	for (const Material & m : materials) {
		b200rt_material f = {};
		for (int c = 0; c < 3; c++) { f.diffuse[c] = m.diffuse[c]; f.reflection[c] = m.reflection[c]; f.transmittance[c] = m.transmittance[c]; }
		f.texture = m.texture; f.index_of_refraction = m.index_of_refraction;
		flat_materials.push_back(f);
	}

	b200rt_static_scene s = {};
	s.nodes = flat_nodes.data();         s.node_count = int(flat_nodes.size());
	s.triangles_hot = flat_hot.data();   s.triangles_cold = flat_cold.data(); s.triangle_count = int(flat_hot.size());
	s.blas = flat_blas.data();           s.blas_count = int(flat_blas.size());
	s.materials = flat_materials.data(); s.material_count = int(flat_materials.size());
	s.textures = texture_pool.textures.data(); s.texture_count = int(texture_pool.textures.size());
	s.texels = texture_pool.texels.data();     s.texel_count = int64_t(texture_pool.texels.size() / 3);
	s.texels_rgb8 = texture_pool.rgb8.empty() ? nullptr : texture_pool.rgb8.data();
	s.gamma_lut   = texture_pool.gamma_lut;
	s.sky = sky.data();                  s.sky_size = sky_size;
	return s;
}

b200rt_frame Scene::frame_view() {
	flat_instances.clear(); flat_spheres.clear(); flat_planes.clear();

	for (const Mesh & mesh : top_level_bvh.primitives) {
		b200rt_instance inst = {};
		memcpy(inst.world,   mesh.transform.world_matrix.cells, sizeof(inst.world));
		memcpy(inst.inverse, mesh.transform_inv.cells,          sizeof(inst.inverse));
		inst.blas = mesh.bvh;
		flat_instances.push_back(inst);
	}
	for (const Sphere & sphere : spheres) {
		b200rt_sphere s = {};
		for (int c = 0; c < 3; c++) s.center[c] = sphere.transform.position[c];
		s.radius_squared = sphere.radius_squared; s.radius_inv = sphere.radius_inv; s.material_id = sphere.material_id;
		flat_spheres.push_back(s);
	}
	for (const Plane & plane : planes) {
		b200rt_plane p = {};
		for (int c = 0; c < 3; c++) { p.normal[c] = plane.world_normal[c]; p.u_axis[c] = plane.u_axis[c]; p.v_axis[c] = plane.v_axis[c]; }
		p.distance = plane.world_distance; p.material_id = plane.material_id;
		flat_planes.push_back(p);
	}

	b200rt_frame f = {};
	for (int c = 0; c < 3; c++) {
		f.camera.position[c]        = camera.position[c];
		f.camera.top_left_corner[c] = camera.rotated_top_left_corner[c];
		f.camera.x_axis[c]          = camera.rotated_x_axis[c];
		f.camera.y_axis[c]          = camera.rotated_y_axis[c];
		f.ambient[c]                = ambient_lighting[c];
	}
	f.instances = flat_instances.data(); f.instance_count = int(flat_instances.size());
	f.tlas_nodes = top_level_bvh.nodes.data(); f.tlas_node_count = top_level_bvh.node_count;
	f.tlas_indices = top_level_bvh.builder.indices[0].data();
	f.spheres = flat_spheres.data(); f.sphere_count = int(flat_spheres.size());
	f.planes  = flat_planes.data();  f.plane_count  = int(flat_planes.size());
	f.lights  = lights.data();       f.light_count  = int(lights.size());
	return f;
}

} // namespace b200rt
