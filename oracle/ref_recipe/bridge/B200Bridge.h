// B200Bridge.h — the reference-side binding of libb200rt.so: the one header a maintainer of the reference adds.
//
// It is written against the reference's OWN headers (Scene.h, TopLevelBVH.h, BottomLevelBVH.h, Material.h, Texture.h,
// Sky.h, the light and primitive headers) and include/b200rt.h, and it is compiled: oracle/ref_recipe/build_ref.py
// builds the "_bridge" variant of the headless reference with it, and tests/test_reference_bridge.py runs that binary —
// the reference's Scene, Scene::update, BVH builders and loaders, with Main.cpp:56-59 replaced by B200Bridge::render —
// and compares the GPU frame with the frame the same process traces on the CPU.
//
// Four reference classes keep the data this needs in private members (Texture: data, width, height, mipmapped,
// mip_levels, mip_offsets; Sky: size, data; Sphere: radius_inv, radius_squared; Plane: world_normal, world_distance,
// u_axis, v_axis). A maintainer adds `friend struct B200Bridge;` to each. The reference's files cannot be edited from
// here, so the test build includes Scene.h with the access specifier lifted (Main.cpp, B200RT_BRIDGE) — same layout.
//
// Lives under oracle/ because only the test build uses it; it calls nothing in oracle/ and links libb200rt.so only.
#pragma once

#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <map>
#include <vector>

#include "b200rt.h"
#include "Scene.h"

struct B200Bridge {
	b200rt_ctx * ctx = nullptr;

	// static scene: the reference's arrays, concatenated per distinct BottomLevelBVH
	std::vector<b200rt_bvh_node>      nodes;
	std::vector<b200rt_triangle_hot>  hot;
	std::vector<b200rt_triangle_cold> cold;
	std::vector<b200rt_blas>          blas;
	std::vector<b200rt_material>      materials;
	std::vector<b200rt_texture>       textures;
	std::vector<float>                texels, sky;
	std::map<const BottomLevelBVH *, int> blas_index;     // the filename-keyed cache of BottomLevelBVH.cpp:16-22 hands out one object per file
	std::map<const Texture *, int>        texture_index;

	// per-frame
	std::vector<b200rt_instance> instances;
	std::vector<b200rt_sphere>   spheres;
	std::vector<b200rt_plane>    planes;
	std::vector<b200rt_light>    lights;

	static void fail(const char * what, const b200rt_ctx * ctx) { fprintf(stderr, "B200Bridge: %s: %s\n", what, b200rt_last_error(ctx)); abort(); }   // the reference abort()s too

	static void put(float * out, const Vector3 & v) { out[0] = v.x; out[1] = v.y; out[2] = v.z; }
	static void put(float * out, const SIMD_Vector3 & v) { out[0] = v.x[0]; out[1] = v.y[0]; out[2] = v.z[0]; }   // every lane holds the same value

	// After Scene::Scene (Main.cpp:41). width / height: Window's; the other settings are the reference's Config.h macros.
	void init(const Scene & scene, int width, int height, int device = 0) {
		static_assert(sizeof(BVHNode) == sizeof(b200rt_bvh_node), "BVHNode.h:10-16");
		static_assert(sizeof(BottomLevelBVH::TriangleHot) == sizeof(b200rt_triangle_hot), "BottomLevelBVH.h:6-10");
		static_assert(sizeof(BottomLevelBVH::TriangleCold) == sizeof(b200rt_triangle_cold), "BottomLevelBVH.h:12-22");
		static_assert(sizeof(Matrix4) == 16 * sizeof(float), "Matrix4.h:19-28");

		b200rt_config cfg;
		memset(&cfg, 0, sizeof(cfg));
		cfg.width = width; cfg.height = height;
		cfg.bounces = NUMBER_OF_BOUNCES;
		cfg.mipmap_filter = MIPMAP_FILTER;        // MIPMAP_FILTER_TRILINEAR 0, _ANISOTROPIC 1, _EWA 2 = B200RT_FILTER_*
		cfg.max_anisotropy = MAX_ANISOTROPY;
		cfg.enable_fxaa = ENABLE_FXAA ? 1 : 0;
		cfg.device = device;
		cfg.tile_stride = 1; cfg.tile_phase = 0;
		cfg.traversal_naive = BVH_TRAVERSAL_STRATEGY == BVH_TRAVERSE_TREE_NAIVE;
		cfg.texture_sample_mode = TEXTURE_SAMPLE_MODE == TEXTURE_SAMPLE_MODE_NEAREST ? B200RT_TEXTURE_NEAREST : TEXTURE_SAMPLE_MODE == TEXTURE_SAMPLE_MODE_BILINEAR ? B200RT_TEXTURE_BILINEAR : B200RT_TEXTURE_MIPMAP;
		cfg.arithmetic = B200RT_ARITHMETIC_EXACT;
		cfg.deep_queue = B200RT_QUEUE_AUTO;
		if (b200rt_create(&cfg, &ctx)) fail("b200rt_create", nullptr);

		// one b200rt_blas per distinct BottomLevelBVH, in order of first use by the TLAS primitives
		const TopLevelBVH & tlas = scene.top_level_bvh;
		for (int i = 0; i < tlas.primitive_count; i++) {
			const BottomLevelBVH * bvh = tlas.primitives[i].bvh;
			if (blas_index.count(bvh)) continue;
			blas_index[bvh] = int(blas.size());
			b200rt_blas entry = { int(nodes.size()), bvh->node_count, int(hot.size()), bvh->index_count, bvh->material_offset };
			blas.push_back(entry);
			const b200rt_bvh_node * n = reinterpret_cast<const b200rt_bvh_node *>(bvh->nodes);
			nodes.insert(nodes.end(), n, n + bvh->node_count);
			// flattened by BottomLevelBVH::flatten (BottomLevelBVH.cpp:196-212): index_count triangles in leaf order
			const b200rt_triangle_hot  * h = reinterpret_cast<const b200rt_triangle_hot  *>(bvh->triangles_hot);
			const b200rt_triangle_cold * c = reinterpret_cast<const b200rt_triangle_cold *>(bvh->triangles_cold);
			hot.insert(hot.end(), h, h + bvh->index_count);
			cold.insert(cold.end(), c, c + bvh->index_count);
		}

		// Material.h:7-14: the Texture pointer becomes an index into `textures`
		for (int m = 0; m < MaterialBuffer::material_count; m++) {
			const Material & material = MaterialBuffer::materials[m];
			b200rt_material out;
			memset(&out, 0, sizeof(out));
			put(out.diffuse, material.diffuse); put(out.reflection, material.reflection); put(out.transmittance, material.transmittance);
			out.index_of_refraction = material.index_of_refraction;
			out.texture = -1;
			if (const Texture * texture = material.texture) {
				if (!texture_index.count(texture)) {
					texture_index[texture] = int(textures.size());
					b200rt_texture t;
					memset(&t, 0, sizeof(t));
					t.texel_offset = int64_t(texels.size() / 3);
					t.width = texture->width; t.height = texture->height;
					t.mipmapped = texture->mipmapped ? 1 : 0;
					t.mip_levels = texture->mip_levels;                                  // 1 when not mip-mapped (Texture.cpp:115-118)
					if (t.mip_levels > B200RT_MAX_MIP_LEVELS) { fprintf(stderr, "B200Bridge: texture with %d mip levels\n", t.mip_levels); abort(); }
					for (int l = 0; l < t.mip_levels; l++) t.mip_offsets[l] = texture->mip_offsets[l];
					// Texture.cpp:57-61: level 0 plus a third of it again when mip-mapped; Vector3 = three floats
					const size_t count = size_t(t.width) * t.height + (texture->mipmapped ? size_t(t.width) * t.height / 3 : 0);
					const float * data = reinterpret_cast<const float *>(texture->data);
					texels.insert(texels.end(), data, data + 3 * count);
					textures.push_back(t);
				}
				out.texture = texture_index[texture];
			}
			materials.push_back(out);
		}

		// Sky.cpp:13-25 + the one texel past the end that Sky::sample's clamp can reach (Sky.cpp:45)
		const size_t sky_texels = size_t(scene.sky.size) * scene.sky.size;
		const float * sky_data = reinterpret_cast<const float *>(scene.sky.data);
		sky.assign(sky_data, sky_data + 3 * sky_texels);
		sky.insert(sky.end(), sky_data + 3 * (sky_texels - 1), sky_data + 3 * sky_texels);

		b200rt_static_scene s;
		memset(&s, 0, sizeof(s));
		s.nodes = nodes.data(); s.node_count = int(nodes.size());
		s.triangles_hot = hot.data(); s.triangles_cold = cold.data(); s.triangle_count = int(hot.size());
		s.blas = blas.data(); s.blas_count = int(blas.size());
		s.materials = materials.data(); s.material_count = int(materials.size());
		s.textures = textures.data(); s.texture_count = int(textures.size());
		s.texels = texels.data(); s.texel_count = int64_t(texels.size() / 3);
		s.sky = sky.data(); s.sky_size = scene.sky.size;
		if (b200rt_upload_static(ctx, &s)) fail("b200rt_upload_static", ctx);
	}

	// Replaces Main.cpp:56-59 (wake_up_worker_threads, wait_on_worker_threads, draw_quad) after Scene::update.
	void render(const Scene & scene, unsigned * frame_buffer, unsigned char * presented_rgb8, b200rt_stats * stats) {
		const TopLevelBVH & tlas = scene.top_level_bvh;
		instances.resize(tlas.primitive_count);
		for (int i = 0; i < tlas.primitive_count; i++) {               // Mesh.h:8-16 after Mesh::update (Mesh.cpp:9-15)
			const Mesh & mesh = tlas.primitives[i];
			memset(&instances[i], 0, sizeof(b200rt_instance));
			memcpy(instances[i].world,   mesh.transform.world_matrix.cells, 16 * sizeof(float));
			memcpy(instances[i].inverse, mesh.transform_inv.cells,          16 * sizeof(float));
			instances[i].blas = blas_index.at(mesh.bvh);
		}
		spheres.resize(scene.spheres.primitive_count);
		for (int i = 0; i < scene.spheres.primitive_count; i++) {      // Sphere.h:8-22
			const Sphere & sphere = scene.spheres[i];
			memset(&spheres[i], 0, sizeof(b200rt_sphere));
			put(spheres[i].center, sphere.transform.position);
			spheres[i].radius_squared = sphere.radius_squared; spheres[i].radius_inv = sphere.radius_inv; spheres[i].material_id = sphere.material_id;
		}
		planes.resize(scene.planes.primitive_count);
		for (int i = 0; i < scene.planes.primitive_count; i++) {       // Plane.h:8-22 after Plane::update (Plane.cpp:3-11)
			const Plane & plane = scene.planes[i];
			memset(&planes[i], 0, sizeof(b200rt_plane));
			put(planes[i].normal, plane.world_normal); planes[i].distance = plane.world_distance;
			put(planes[i].u_axis, plane.u_axis); put(planes[i].v_axis, plane.v_axis);
			planes[i].material_id = plane.material_id;
		}
		lights.clear();                                                // the order of Raytracer.cpp:155-198: point, spot, directional
		for (int i = 0; i < scene.point_light_count; i++) {
			b200rt_light l; memset(&l, 0, sizeof(l)); l.kind = 0;
			put(l.colour, scene.point_lights[i].colour); put(l.position, scene.point_lights[i].position);
			lights.push_back(l);
		}
		for (int i = 0; i < scene.spot_light_count; i++) {
			const SpotLight & spot = scene.spot_lights[i];
			b200rt_light l; memset(&l, 0, sizeof(l)); l.kind = 1;
			put(l.colour, spot.colour); put(l.position, spot.position); put(l.negative_direction, spot.negative_direction);
			l.inner_cutoff = spot.inner_cutoff; l.outer_cutoff = spot.outer_cutoff;
			lights.push_back(l);
		}
		for (int i = 0; i < scene.directional_light_count; i++) {
			b200rt_light l; memset(&l, 0, sizeof(l)); l.kind = 2;
			put(l.colour, scene.directional_lights[i].colour); put(l.negative_direction, scene.directional_lights[i].negative_direction);
			lights.push_back(l);
		}

		b200rt_frame f;
		memset(&f, 0, sizeof(f));
		put(f.camera.position, scene.camera.position);                 // Camera.h:12, :21-23 after Camera::update
		put(f.camera.top_left_corner, scene.camera.rotated_top_left_corner);
		put(f.camera.x_axis, scene.camera.rotated_x_axis);
		put(f.camera.y_axis, scene.camera.rotated_y_axis);
		put(f.ambient, scene.ambient_lighting);
		f.instances = instances.data(); f.instance_count = tlas.primitive_count;
		f.tlas_nodes = reinterpret_cast<const b200rt_bvh_node *>(tlas.nodes); f.tlas_node_count = tlas.node_count;   // same bytes: passed in place
		f.tlas_indices = tlas.indices;
		f.spheres = spheres.data(); f.sphere_count = int(spheres.size());
		f.planes = planes.data(); f.plane_count = int(planes.size());
		f.lights = lights.data(); f.light_count = int(lights.size());

		b200rt_outputs out;
		memset(&out, 0, sizeof(out));
		out.frame_buffer = frame_buffer; out.presented_rgb8 = presented_rgb8; out.stats = stats;
		if (b200rt_render_frame(ctx, &f, &out)) fail("b200rt_render_frame", ctx);
	}

	void destroy() { b200rt_destroy(ctx); ctx = nullptr; }
};
