// render.cuh — the frame kernels: Whitted ray tracing over all tiles, primary-hit dump, FXAA.
//
// Reference: Raytracer::render_tile (Raytracer.cpp:3-85), Raytracer::bounce (:87-400),
// WorkerThreads tile hand-out (WorkerThread.cpp:49-69), Window::plot (Window.h:56-65),
// Data/Shaders/fragment_fxaa.glsl.
#pragma once
#include "shade.cuh"

namespace rt {

#define RT_MAX_BOUNCES 8
#ifndef RT_BLOCK_THREADS
#define RT_BLOCK_THREADS 256
#endif
#ifndef RT_MIN_BLOCKS
#define RT_MIN_BLOCKS 4   // resident 256-thread blocks per SM the register allocation is sized for: 64 registers, 32 warps/SM.
                          // Measured on the bench frame (gpurun_out/r1_variants_*.log): 2 -> 1.87 ms, 3 -> 1.53, 4 -> 1.46, 5 -> 1.48, 6 -> 1.57, 8 -> 1.69
#endif

// Raytracer::render_tile, Raytracer.cpp:45-59: primary ray + its differentials for pixel (i, j)
__device__ __forceinline__ Ray primary_ray(const SceneParams & P, int i, int j) {
	const v3 x_axis = V(P.cam_x_axis[0], P.cam_x_axis[1], P.cam_x_axis[2]);
	const v3 y_axis = V(P.cam_y_axis[0], P.cam_y_axis[1], P.cam_y_axis[2]);
	const v3 top_left = V(P.cam_top_left[0], P.cam_top_left[1], P.cam_top_left[2]);

	Ray ray;
	ray.origin = V(P.cam_position[0], P.cam_position[1], P.cam_position[2]);
	ray.dO_dx = V(0.0f, 0.0f, 0.0f);
	ray.dO_dy = V(0.0f, 0.0f, 0.0f);

	const v3 direction = madd(x_axis, float(i), madd(y_axis, float(j), top_left));

	const float d_dot_d          = dot(direction, direction);
	const float inv_sqrt_d_dot_d = inv_sqrt(d_dot_d);
	const float denom            = inv_sqrt_d_dot_d / d_dot_d; // d_dot_d ^ -3/2

	ray.dD_dx = (x_axis * d_dot_d - direction * dot(direction, x_axis)) * denom;
	ray.dD_dy = (y_axis * d_dot_d - direction * dot(direction, y_axis)) * denom;

	ray.direction = direction * inv_sqrt_d_dot_d;
	return ray;
}

// Window::plot, Window.h:56-65: linear colour -> 0x00RRGGBB. cvtss2si(NaN) = 0x80000000 is
// reproduced by float_to_int (Math::clamp lets NaN through, Math.h:13-18).
__device__ __forceinline__ unsigned plot_colour(v3 colour) {
	float c[3] = { colour.x, colour.y, colour.z };
	unsigned q[3];
#pragma unroll
	for (int k = 0; k < 3; k++) {
		float v = c[k] * 255.0f;
		if (v < 0.0f) v = 0.0f;
		if (v > 255.0f) v = 255.0f;
		q[k] = unsigned(float_to_int(v - 0.5f));
	}
	return (q[0] << 16) | (q[1] << 8) | q[2];
}

// One pending level of Raytracer::bounce's recursion (:204-397), kept in local memory. The
// recursion is binary (reflection, then refraction) and is unrolled depth-first with the same
// association of every sum as the reference, so the colour is bit-identical to the recursive form.
struct BounceFrame {
	v3    result;            // direct lighting (+ reflection once stage 2 is reached)
	v3    colour_reflection; // material_reflection * child colour
	Ray   refracted;         // prepared at the hit, traced after the reflection child returns
	float distance;          // this node's own hit distance (what bounce() hands to its caller)
	float F_r;               // Schlick reflectance of the refraction branch
	int   material_id;
	int   flags;             // see FRAME_*
};
enum { FRAME_AFTER_REFLECTION = 1, FRAME_AFTER_REFRACTION = 2, FRAME_HAS_REFRACTION = 4, FRAME_TIR = 8, FRAME_ENTERING = 16 };

// Raytracer::bounce for one primary ray, iteratively. Returns the pixel colour.
template<bool COUNT>
__device__ v3 shade_pixel(const SceneParams & P, Ray ray, Counters & cnt) {
	BounceFrame frames[RT_MAX_BOUNCES];
	int depth = 0;                // number of pending frames = recursion depth of the current ray
	int bounces_left = P.bounces;

	v3    returned;               // value of the bounce() call that just finished
	float returned_distance;

	for (;;) {
		// ---- bounce(ray, bounces_left): Raytracer.cpp:87-201 ---------------------------------
		Candidate best;
		trace_closest<COUNT>(P, ray.origin, ray.direction, best, cnt);

		bool descend = false;

		if (best.instance == HIT_NONE) { // :105-111
			returned = sky_sample<COUNT>(P, ray.direction, cnt);
			returned_distance = CUDART_INF_F;
		} else {
			HitRecord hit;
			build_record(P, ray, best, hit);

			const float4 * mp = reinterpret_cast<const float4 *>(P.materials + hit.material_id);
			const v3 material_diffuse = get_albedo<COUNT>(P, __ldg(mp), hit, cnt);

			v3 result = V(0.0f, 0.0f, 0.0f);

			if (dot(material_diffuse, material_diffuse) > 0.0f) { // :143-201
				v3 diffuse = V(P.ambient[0], P.ambient[1], P.ambient[2]);
				const v3 to_camera = normalize(V(P.cam_position[0], P.cam_position[1], P.cam_position[2]) - hit.point);

				for (int i = 0; i < P.light_count; i++) { // point, then spot, then directional (:155-198)
					const float4 l0 = __ldg(&P.lights[i].colour_kind), l1 = __ldg(&P.lights[i].position_inner), l2 = __ldg(&P.lights[i].negdir_outer);
					const int kind = __float_as_int(l0.w);
					cnt.shadow++;

					// one shadow query for all three kinds (a single inlined copy of the traversal)
					v3 to_light; float distance_to_light = CUDART_INF_F, distance_to_light_squared = 0.0f;
					if (kind == 2) { // DirectionalLight.h:9-11
						to_light = V(l2);
					} else {
						to_light = V(l1) - hit.point;
						distance_to_light_squared = dot(to_light, to_light);
						distance_to_light         = sqrtf(distance_to_light_squared);
						to_light = divf(to_light, distance_to_light);
					}
					if (trace_any<COUNT>(P, hit.point, to_light, distance_to_light, cnt)) continue;

					if (kind == 2) {
						diffuse = diffuse + light_base(V(l0), hit.normal, to_light, to_camera);
					} else {
						v3 lighting = divf(light_base(V(l0), hit.normal, to_light, to_camera), distance_to_light_squared); // PointLight.h:9-11
						if (kind == 1) { // SpotLight.h:17-33
							const float d = dot(to_light, V(l2));
							if (!(d > l2.w)) lighting = V(0.0f, 0.0f, 0.0f);
							else {
								float radial_falloff = (d - l2.w) / (l1.w - l2.w);
								if (radial_falloff > 1.0f) radial_falloff = 1.0f;
								lighting = lighting * radial_falloff;
							}
						}
						diffuse = diffuse + lighting;
					}
				}
				result = V(__fmaf_rn(diffuse.x, material_diffuse.x, result.x), __fmaf_rn(diffuse.y, material_diffuse.y, result.y), __fmaf_rn(diffuse.z, material_diffuse.z, result.z)); // :200
			}

			returned = result;
			returned_distance = best.t;

			if (bounces_left > 0) { // :204-247
				const float4 m1 = __ldg(mp + 1), m2 = __ldg(mp + 2);
				const v3 material_reflection = V(m1), material_transmittance = V(m2);

				const bool reflection_mask = dot(material_reflection,    material_reflection)    > 0.0f;
				const bool refraction_mask = dot(material_transmittance, material_transmittance) > 0.0f;

				if (reflection_mask || refraction_mask) {
					BounceFrame & frame = frames[depth];
					frame.result = result;
					frame.colour_reflection = V(0.0f, 0.0f, 0.0f);
					frame.distance = best.t;
					frame.material_id = hit.material_id;
					int flags = 0;

					const float d_dot_n = dot(ray.direction, hit.normal);
					const float dDN_dx = dot(ray.dD_dx, hit.normal) + dot(ray.direction, hit.dN_dx);
					const float dDN_dy = dot(ray.dD_dy, hit.normal) + dot(ray.direction, hit.dN_dy);

					if (refraction_mask) { // :273-342 — everything that only depends on this hit
						flags |= FRAME_HAS_REFRACTION;
						const bool entering = d_dot_n < 0.0f;
						if (entering) flags |= FRAME_ENTERING;

						const float ior = m1.w;
						const float n_1 = entering ? 1.0f : ior;
						const float n_2 = entering ? ior : 1.0f;

						float    cos_theta = entering ? 0.0f - d_dot_n : d_dot_n;
						const v3 normal    = entering ? hit.normal : neg(hit.normal);

						const float eta = n_1 / n_2;
						const float k   = 1.0f - (eta * eta * (1.0f - (cos_theta * cos_theta)));

						if (k < 0.0f) flags |= FRAME_TIR; // :310-314
						else {
							Ray & refracted = frame.refracted;
							refracted.origin    = hit.point;
							refracted.direction = ray.direction * eta + normal * ((eta * cos_theta) - sqrtf(k)); // Math::refract, Math.h:34-36
							refracted.dO_dx = hit.dO_dx;
							refracted.dO_dy = hit.dO_dy;

							const float D_dot_N      = 0.0f - cos_theta;
							const float Dprime_dot_N = 0.0f - sqrtf(k);
							const float mu = 0.0f - (eta * cos_theta + Dprime_dot_N);

							// :341-342, float + vector broadcast and all
							refracted.dD_dx = ray.dD_dx * eta - addf(hit.dN_dx * hit.normal, mu * D_dot_N) * dDN_dx;
							refracted.dD_dy = ray.dD_dy * eta - addf(hit.dN_dy * hit.normal, mu * D_dot_N) * dDN_dy;

							// Schlick, :379-391
							float r_0 = (n_1 - n_2) / (n_1 + n_2);
							r_0 = r_0 * r_0;
							if (n_1 > n_2) cos_theta = 0.0f - dot(refracted.direction, normal);

							const float one_minus_cos         = 1.0f - cos_theta;
							const float one_minus_cos_squared = one_minus_cos * one_minus_cos;
							frame.F_r = r_0 + ((1.0f - r_0) * one_minus_cos_squared) * (one_minus_cos_squared * one_minus_cos);
						}
					}

					if (reflection_mask) { // :249-271
						Ray reflected;
						reflected.origin    = hit.point;
						reflected.direction = ray.direction - hit.normal * (2.0f * d_dot_n); // Math::reflect, Math.h:28-31
						reflected.dO_dx = hit.dO_dx;
						reflected.dO_dy = hit.dO_dy;
						reflected.dD_dx = ray.dD_dx - (hit.dN_dx * d_dot_n + hit.normal * dDN_dx) * 2.0f;
						reflected.dD_dy = ray.dD_dy - (hit.dN_dy * d_dot_n + hit.normal * dDN_dy) * 2.0f;

						cnt.reflection++;
						frame.flags = flags | FRAME_AFTER_REFLECTION;
						ray = reflected;
						depth++; bounces_left--;
						descend = true;
					} else if (flags & FRAME_TIR) {
						// no reflection, total internal reflection: bounce() returns `result` (:312-314)
					} else {
						cnt.refraction++;
						frame.flags = flags | FRAME_AFTER_REFRACTION;
						ray = frame.refracted;
						depth++; bounces_left--;
						descend = true;
					}
				}
			}
		}
		if (descend) continue;

		// ---- a bounce() call returned: unwind pending frames (:264-271, :344-396) ---------------
		while (depth > 0) {
			BounceFrame & frame = frames[depth - 1];

			if (frame.flags & FRAME_AFTER_REFLECTION) {
				const float4 m1 = __ldg(reinterpret_cast<const float4 *>(P.materials + frame.material_id) + 1);
				const v3 colour_reflection = V(m1) * returned;   // :268
				const v3 result = frame.result + colour_reflection; // :270

				if (!(frame.flags & FRAME_HAS_REFRACTION)) {
					returned = result; returned_distance = frame.distance;
					depth--; bounces_left++;
				} else if (frame.flags & FRAME_TIR) {
					returned = result + colour_reflection;           // reflection counted twice, :312-314
					returned_distance = frame.distance;
					depth--; bounces_left++;
				} else {
					frame.result = result;
					frame.colour_reflection = colour_reflection;
					frame.flags = (frame.flags & ~FRAME_AFTER_REFLECTION) | FRAME_AFTER_REFRACTION;
					cnt.refraction++;
					ray = frame.refracted;                           // same depth: sibling of the reflection child
					descend = true;
					break;
				}
			} else { // FRAME_AFTER_REFRACTION
				v3 colour_refraction = returned;
				if (frame.flags & FRAME_ENTERING) { // Beer's law with the child's hit distance, :350-376
					const float4 m2 = __ldg(reinterpret_cast<const float4 *>(P.materials + frame.material_id) + 2);
					colour_refraction.x = colour_refraction.x * expf((m2.x - 1.0f) * returned_distance);
					colour_refraction.y = colour_refraction.y * expf((m2.y - 1.0f) * returned_distance);
					colour_refraction.z = colour_refraction.z * expf((m2.z - 1.0f) * returned_distance);
				}
				const float F_r = frame.F_r, F_t = 1.0f - F_r;
				const v3 blend = frame.colour_reflection * F_r + colour_refraction * F_t; // :393
				returned = frame.result + blend;                                          // :395
				returned_distance = frame.distance;
				depth--; bounces_left++;
			}
		}
		if (!descend) break;
	}
	return returned;
}

} // namespace rt
