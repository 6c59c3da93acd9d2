// TEST INFRASTRUCTURE — builds the reference's `Scene` object from a b200rt scene
// description file instead of the two hard-coded constructors of Scene.cpp:7-131
// (those are stripped from the generated copy of Scene.cpp by build_ref.py; everything
// else in Scene.cpp — update, trace_primitives, intersect_primitives — is the
// reference's own code). Only the reference's public API is used to fill the Scene, in
// the same order as the shipped constructors so that material ids come out the same:
// material 0 = default (Main), spheres, planes (member initialisers, Primitive.h:6),
// then every mesh's materials in load order (OBJLoader.cpp:9).
#include "SceneFile.h"

#include "Scene.h"

#include <fstream>
#include <sstream>
#include <stdexcept>
#include <new>

SceneFile g_scene_file;

static std::string resolve(const std::string & dir, const std::string & p) {
	if (p.empty() || p == "-") return "";
	if (p[0] == '/') return p;
	return dir + p;
}

void SceneFile::load(const char * path) {
	std::ifstream in(path);
	if (!in) { fprintf(stderr, "cannot open scene file %s\n", path); exit(2); }

	std::string dir(path);
	size_t slash = dir.find_last_of('/');
	dir = slash == std::string::npos ? std::string("./") : dir.substr(0, slash + 1);

	std::string line;
	while (std::getline(in, line)) {
		std::istringstream ss(line);
		std::string key;
		if (!(ss >> key) || key[0] == '#') continue;

		if (key == "b200rt_scene") { int v; ss >> v; }
		else if (key == "resolution") ss >> width >> height;
		else if (key == "camera")     ss >> fov_deg >> cam_pos[0] >> cam_pos[1] >> cam_pos[2] >> cam_rot[0] >> cam_rot[1] >> cam_rot[2] >> cam_rot[3];
		else if (key == "ambient")    ss >> ambient[0] >> ambient[1] >> ambient[2];
		else if (key == "sky")        { std::string p; ss >> p; sky = resolve(dir, p); }
		else if (key == "sphere") {
			SphereDesc s; std::string tex;
			ss >> s.radius >> s.pos[0] >> s.pos[1] >> s.pos[2];
			for (int i = 0; i < 3; i++) ss >> s.mat.diffuse[i];
			for (int i = 0; i < 3; i++) ss >> s.mat.reflection[i];
			for (int i = 0; i < 3; i++) ss >> s.mat.transmittance[i];
			ss >> s.mat.ior >> tex;
			s.mat.texture = resolve(dir, tex);
			spheres.push_back(s);
		} else if (key == "plane") {
			PlaneDesc p; std::string tex;
			ss >> p.pos[0] >> p.pos[1] >> p.pos[2] >> p.rot[0] >> p.rot[1] >> p.rot[2] >> p.rot[3];
			for (int i = 0; i < 3; i++) ss >> p.mat.diffuse[i];
			for (int i = 0; i < 3; i++) ss >> p.mat.reflection[i];
			for (int i = 0; i < 3; i++) ss >> p.mat.transmittance[i];
			ss >> p.mat.ior >> tex;
			p.mat.texture = resolve(dir, tex);
			planes.push_back(p);
		} else if (key == "mesh") {
			MeshDesc m; std::string p;
			ss >> p >> m.pos[0] >> m.pos[1] >> m.pos[2] >> m.rot[0] >> m.rot[1] >> m.rot[2] >> m.rot[3];
			m.path = resolve(dir, p);
			meshes.push_back(m);
		} else if (key == "point_light") {
			LightDesc l; l.kind = 0;
			ss >> l.colour[0] >> l.colour[1] >> l.colour[2] >> l.pos[0] >> l.pos[1] >> l.pos[2];
			lights.push_back(l);
		} else if (key == "spot_light") {
			LightDesc l; l.kind = 1;
			ss >> l.colour[0] >> l.colour[1] >> l.colour[2] >> l.pos[0] >> l.pos[1] >> l.pos[2] >> l.dir[0] >> l.dir[1] >> l.dir[2] >> l.inner_deg >> l.outer_deg;
			lights.push_back(l);
		} else if (key == "directional_light") {
			LightDesc l; l.kind = 2;
			ss >> l.colour[0] >> l.colour[1] >> l.colour[2] >> l.dir[0] >> l.dir[1] >> l.dir[2];
			lights.push_back(l);
		} else {
			fprintf(stderr, "scene file %s: unknown key '%s'\n", path, key.c_str());
			exit(2);
		}
		if (ss.fail()) { fprintf(stderr, "scene file %s: bad line '%s'\n", path, line.c_str()); exit(2); }
	}
	if (meshes.empty()) { fprintf(stderr, "scene file %s: the reference needs at least one mesh (TopLevelBVH.cpp:6)\n", path); exit(2); }
}

static void apply_material(int material_id, const MaterialDesc & m) {
	Material & material = MaterialBuffer::materials[material_id];
	material.diffuse       = Vector3(m.diffuse[0], m.diffuse[1], m.diffuse[2]);
	material.reflection    = Vector3(m.reflection[0], m.reflection[1], m.reflection[2]);
	material.transmittance = Vector3(m.transmittance[0], m.transmittance[1], m.transmittance[2]);
	material.index_of_refraction = m.ior;
	if (!m.texture.empty()) material.texture = Texture::load(m.texture.c_str());
}

Scene::Scene() :
	spheres(int(g_scene_file.spheres.size())),
	planes (int(g_scene_file.planes.size())),
	sky(g_scene_file.sky.c_str()),
	camera(DEG_TO_RAD(g_scene_file.fov_deg))
{
	const SceneFile & f = g_scene_file;

	for (size_t i = 0; i < f.spheres.size(); i++) {
		spheres[int(i)].init(f.spheres[i].radius);
		spheres[int(i)].transform.position = Vector3(f.spheres[i].pos[0], f.spheres[i].pos[1], f.spheres[i].pos[2]);
		apply_material(spheres[int(i)].material_id, f.spheres[i].mat);
	}
	for (size_t i = 0; i < f.planes.size(); i++) {
		planes[int(i)].transform.position = Vector3(f.planes[i].pos[0], f.planes[i].pos[1], f.planes[i].pos[2]);
		planes[int(i)].transform.rotation = Quaternion(f.planes[i].rot[0], f.planes[i].rot[1], f.planes[i].rot[2], f.planes[i].rot[3]);
		apply_material(planes[int(i)].material_id, f.planes[i].mat);
	}

	top_level_bvh.init(int(f.meshes.size()));
	for (size_t i = 0; i < f.meshes.size(); i++) {
		Mesh & mesh = top_level_bvh.primitives[i];
		mesh.transform.position = Vector3(f.meshes[i].pos[0], f.meshes[i].pos[1], f.meshes[i].pos[2]);
		mesh.transform.rotation = Quaternion(f.meshes[i].rot[0], f.meshes[i].rot[1], f.meshes[i].rot[2], f.meshes[i].rot[3]);
		mesh.init(f.meshes[i].path.c_str());
	}

	// Lights: placement-new into raw storage because the reference's light types have no
	// default constructor (PointLight.h:7, SpotLight.h:12, DirectionalLight.h:7). 64-byte aligned for
	// the AVX members; the Scene is never destroyed (Main leaks it on purpose).
	int counts[3] = { 0, 0, 0 };
	for (const auto & l : f.lights) counts[l.kind]++;

	point_lights       = counts[0] ? static_cast<PointLight       *>(operator new[](counts[0] * sizeof(PointLight), std::align_val_t(64)))       : nullptr;
	spot_lights        = counts[1] ? static_cast<SpotLight        *>(operator new[](counts[1] * sizeof(SpotLight), std::align_val_t(64)))        : nullptr;
	directional_lights = counts[2] ? static_cast<DirectionalLight *>(operator new[](counts[2] * sizeof(DirectionalLight), std::align_val_t(64))) : nullptr;

	for (const auto & l : f.lights) {
		Vector3 colour(l.colour[0], l.colour[1], l.colour[2]);
		Vector3 pos(l.pos[0], l.pos[1], l.pos[2]);
		Vector3 dir(l.dir[0], l.dir[1], l.dir[2]);
		if (l.kind == 0) new (&point_lights[point_light_count++]) PointLight(colour, pos);
		if (l.kind == 1) new (&spot_lights[spot_light_count++]) SpotLight(colour, pos, dir, l.inner_deg, l.outer_deg);
		if (l.kind == 2) new (&directional_lights[directional_light_count++]) DirectionalLight(colour, dir);
	}

	ambient_lighting = Vector3(f.ambient[0], f.ambient[1], f.ambient[2]);

	camera.position = Vector3(f.cam_pos[0], f.cam_pos[1], f.cam_pos[2]);
	camera.rotation = Quaternion(f.cam_rot[0], f.cam_rot[1], f.cam_rot[2], f.cam_rot[3]);
}
