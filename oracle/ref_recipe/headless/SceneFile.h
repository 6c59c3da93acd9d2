// TEST INFRASTRUCTURE — parsed form of a b200rt scene description file (the same text
// file the CUDA host library reads; format documented in DESIGN.md "Scene description").
#pragma once
#include <string>
#include <vector>

struct MaterialDesc {
	float diffuse[3]       = { 1, 1, 1 };
	float reflection[3]    = { 0, 0, 0 };
	float transmittance[3] = { 0, 0, 0 };
	float ior = 1.0f;
	std::string texture;
};
struct SphereDesc { float radius = 1, pos[3] = { 0, 0, 0 }; MaterialDesc mat; };
struct PlaneDesc  { float pos[3] = { 0, 0, 0 }, rot[4] = { 0, 0, 0, 1 }; MaterialDesc mat; };
struct MeshDesc   { std::string path; float pos[3] = { 0, 0, 0 }, rot[4] = { 0, 0, 0, 1 }; };
struct LightDesc  { int kind = 0; float colour[3] = { 1, 1, 1 }, pos[3] = { 0, 0, 0 }, dir[3] = { 0, -1, 0 }, inner_deg = 0, outer_deg = 0; };

struct SceneFile {
	int   width = 1280, height = 720;
	float fov_deg = 110.0f;
	float cam_pos[3] = { 0, 0, 0 };
	float cam_rot[4] = { 0, 0, 0, 1 };
	float ambient[3] = { 0.2f, 0.2f, 0.2f };
	std::string sky;

	std::vector<SphereDesc> spheres;
	std::vector<PlaneDesc>  planes;
	std::vector<MeshDesc>   meshes;
	std::vector<LightDesc>  lights;

	void load(const char * path);
};

extern SceneFile g_scene_file;
