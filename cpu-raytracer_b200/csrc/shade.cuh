// shade.cuh — hit record, texture filtering, sky, lights.
//
// Reference: hit records of triangle_trace (BottomLevelBVH.cpp:254-308), Sphere::trace
// (Sphere.cpp:43-89), Plane::trace (Plane.cpp:24-68); Material::get_albedo (Material.h:16-22);
// Texture::sample* (Texture.cpp:131-337); Sky::sample (Sky.cpp:28-68); Light / PointLight /
// SpotLight / DirectionalLight::calc_lighting (Light.h:12-26, PointLight.h:9-11,
// SpotLight.h:17-33, DirectionalLight.h:9-11).
#pragma once
#include "traverse.cuh"

namespace RT_NAMESPACE {

struct HitRecord {         // RayHit.h:4-22
	v3    point, normal;
	int   material_id;
	float u, v;
	float ds_dx, ds_dy, dt_dx, dt_dy;
	v3    dO_dx, dO_dy, dN_dx, dN_dy;
};

// ------------------------------------------------------------------------- hit records ---
__device__ __forceinline__ void triangle_record(const SceneParams & P, const Ray & ray, const Candidate & c, HitRecord & hit) {
	const float4 * ip = reinterpret_cast<const float4 *>(P.instances + c.instance);
	Affine inverse, world;
	inverse.r0 = __ldg(ip);   inverse.r1 = __ldg(ip + 1); inverse.r2 = __ldg(ip + 2);
	world.r0   = __ldg(ip + 3); world.r1 = __ldg(ip + 4); world.r2   = __ldg(ip + 5);
	const int4 offsets = __ldg(reinterpret_cast<const int4 *>(ip + 6));

	// Mesh::trace, Mesh.cpp:17-31: ray and differentials in object space
	const v3 o     = transform_position (inverse, ray.origin);
	const v3 d     = transform_direction(inverse, ray.direction);
	const v3 dO_dx = transform_direction(inverse, ray.dO_dx);
	const v3 dO_dy = transform_direction(inverse, ray.dO_dy);
	const v3 dD_dx = transform_direction(inverse, ray.dD_dx);
	const v3 dD_dy = transform_direction(inverse, ray.dD_dy);

	const float4 * hot  = P.triangles_hot  + 3 * (size_t(offsets.y) + c.primitive);
	const float4 * cold = P.triangles_cold + 4 * (size_t(offsets.y) + c.primitive);
	const float4 a0 = __ldg(hot), a1 = __ldg(hot + 1), a2 = __ldg(hot + 2);
	const float4 c0 = __ldg(cold), c1 = __ldg(cold + 1), c2 = __ldg(cold + 2), c3 = __ldg(cold + 3);

	const v3 edge_1 = V(a0.w, a1.x, a1.y), edge_2 = V(a1.z, a1.w, a2.x);
	// TriangleCold: tex0.xy tex_e1.xy | tex_e2.xy n0.xy | n0.z n_e1.xyz | n_e2.xyz material
	const float t0x = c0.x, t0y = c0.y, te1x = c0.z, te1y = c0.w, te2x = c1.x, te2y = c1.y;
	const v3 n0 = V(c1.z, c1.w, c2.x), n_edge_1 = V(c2.y, c2.z, c2.w), n_edge_2 = V(c3.x, c3.y, c3.z);

	const float t = c.t, u = c.u, v = c.v;

	const v3 n = (n0 + n_edge_1 * u) + n_edge_2 * v; // Math::barycentric, Math.h:22-25

	hit.point  = transform_position (world, o + d * t);
	hit.normal = transform_direction(world, normalize(n));
	hit.material_id = offsets.z + __float_as_int(c3.w);

	hit.u = (t0x + u * te1x) + v * te2x;
	hit.v = (t0y + u * te1y) + v * te2y;

	// Ray Tracing Gems ch. 20, BottomLevelBVH.cpp:278-308
	const float one_over_k = 1.0f / dot(cross(edge_1, edge_2), d);

	const v3 q = madd(dD_dx, t, dO_dx);
	const v3 r = madd(dD_dy, t, dO_dy);

	const v3 c_u = cross(edge_2, d);
	const v3 c_v = cross(d, edge_1);

	const float du_dx = one_over_k * dot(c_u, q), du_dy = one_over_k * dot(c_u, r);
	const float dv_dx = one_over_k * dot(c_v, q), dv_dy = one_over_k * dot(c_v, r);

	hit.dO_dx = edge_1 * du_dx + edge_2 * dv_dx; // stays in object space (reference quirk)
	hit.dO_dy = edge_1 * du_dy + edge_2 * dv_dy;

	const v3 dn_dx = n_edge_1 * du_dx + n_edge_2 * dv_dx;
	const v3 dn_dy = n_edge_1 * du_dy + n_edge_2 * dv_dy;

	const float n_dot_n = dot(n, n);
	const float N_denom = inv_sqrt(n_dot_n) / n_dot_n;

	hit.dN_dx = (dn_dx * n_dot_n - n * dot(n, dn_dx)) * N_denom;
	hit.dN_dy = (dn_dy * n_dot_n - n * dot(n, dn_dy)) * N_denom;

	hit.ds_dx = du_dx * te1x + dv_dx * te2x;
	hit.ds_dy = du_dy * te1x + dv_dy * te2x;
	hit.dt_dx = du_dx * te1y + dv_dx * te2y;
	hit.dt_dy = du_dy * te1y + dv_dy * te2y;
}

__device__ __forceinline__ void sphere_record(const SceneParams & P, const Ray & ray, const Candidate & c, HitRecord & hit) {
	const float4 s0 = __ldg(&P.spheres[c.primitive].center_r2), s1 = __ldg(&P.spheres[c.primitive].rinv_mat);
	const v3 center = V(s0.x, s0.y, s0.z);
	const float one_over_r = s1.x, t = c.t;

	hit.point  = ray.origin + ray.direction * t;
	hit.normal = (hit.point - center) * one_over_r;
	hit.material_id = __float_as_int(s1.y);

	hit.u = __fmaf_rn(atan2f(hit.normal.z, hit.normal.x), RT_ONE_OVER_TWO_PI, 0.5f); // Sphere.cpp:60-61
	hit.v = __fmaf_rn(acosf (hit.normal.y),               RT_ONE_OVER_PI,     0.5f);

	// Igehy transfer, Sphere.cpp:65-78
	const v3 a_x = madd(ray.dD_dx, t, ray.dO_dx);
	const v3 a_y = madd(ray.dD_dy, t, ray.dO_dy);

	const float denom = (0.0f - 1.0f) / dot(ray.direction, hit.normal);
	const float dt_dx = dot(a_x, hit.normal) * denom;
	const float dt_dy = dot(a_y, hit.normal) * denom;

	const v3 dP_dx = madd(ray.direction, dt_dx, a_x);
	const v3 dP_dy = madd(ray.direction, dt_dy, a_y);

	hit.dO_dx = dP_dx; hit.dO_dy = dP_dy;
	hit.dN_dx = dP_dx * one_over_r;
	hit.dN_dy = dP_dy * one_over_r;

	const v3 n = hit.normal; // Sphere.cpp:80-88
	const float ds_denom = RT_ONE_OVER_TWO_PI / (n.x * n.x + n.z * n.z + 1e-8f);
	hit.ds_dx = (n.x * hit.dN_dx.z - n.z * hit.dN_dx.x) * ds_denom;
	hit.ds_dy = (n.x * hit.dN_dy.z - n.z * hit.dN_dy.x) * ds_denom;

	const float dt_denom = (0.0f - RT_ONE_OVER_PI) * inv_sqrt(1.0f - n.y * n.y + 1e-8f);
	hit.dt_dx = hit.dN_dx.y * dt_denom;
	hit.dt_dy = hit.dN_dy.y * dt_denom;
}

__device__ __forceinline__ void plane_record(const SceneParams & P, const Ray & ray, const Candidate & c, HitRecord & hit) {
	const float4 p0 = __ldg(&P.planes[c.primitive].normal_d), p1 = __ldg(&P.planes[c.primitive].u_axis_mat), p2 = __ldg(&P.planes[c.primitive].v_axis);
	const v3 normal = V(p0.x, p0.y, p0.z), u = V(p1.x, p1.y, p1.z), v = V(p2.x, p2.y, p2.z);
	const float t = c.t;

	hit.point  = ray.origin + ray.direction * t;
	hit.normal = normal;
	hit.material_id = __float_as_int(p1.w);

	hit.u = dot(hit.point, u);
	hit.v = dot(hit.point, v);

	const v3 a_x = madd(ray.dD_dx, t, ray.dO_dx); // Plane.cpp:42-68
	const v3 a_y = madd(ray.dD_dy, t, ray.dO_dy);

	const float denom = (0.0f - 1.0f) / (dot(ray.direction, hit.normal) + 1e-8f);
	const float dt_dx = dot(a_x, hit.normal) * denom;
	const float dt_dy = dot(a_y, hit.normal) * denom;

	const v3 dP_dx = madd(ray.direction, dt_dx, a_x);
	const v3 dP_dy = madd(ray.direction, dt_dy, a_y);

	hit.dO_dx = dP_dx; hit.dO_dy = dP_dy;
	hit.dN_dx = V(0.0f, 0.0f, 0.0f);
	hit.dN_dy = V(0.0f, 0.0f, 0.0f);

	hit.ds_dx = dot(dP_dx, u); hit.ds_dy = dot(dP_dy, u);
	hit.dt_dx = dot(dP_dx, v); hit.dt_dy = dot(dP_dy, v);
}

__device__ __forceinline__ void build_record(const SceneParams & P, const Ray & ray, const Candidate & c, HitRecord & hit) {
	if (c.instance >= 0) triangle_record(P, ray, c, hit);
	else if (c.instance == HIT_SPHERE) sphere_record(P, ray, c, hit);
	else plane_record(P, ray, c, hit);
}

// ---------------------------------------------------------------------------- textures ---
__constant__ float c_ewa_weight_table[128]; // Texture::ewa_weight_table, Texture.h:20-21,53-62 (filled by the host)

// A texel in HBM: float4 (linear RGB + pad; the reference keeps 3 floats, Texture.cpp:30-129), or — FAST arithmetic,
// RT_FAST_TEXEL — four halves in 8 bytes (11 significant bits: the source images have 8).
#ifdef RT_FAST_TEXEL
typedef uint2 Texel;
__device__ __forceinline__ float half_bits_to_float(unsigned h) {
#ifdef RT_EMULATE
	return emu_half_to_float((unsigned short)h);
#else
	return __half2float(__ushort_as_half((unsigned short)h));
#endif
}
__device__ __forceinline__ v3 texel_rgb(const Texel * p) { const uint2 h = __ldg(p); return V(half_bits_to_float(h.x & 0xffffu), half_bits_to_float(h.x >> 16), half_bits_to_float(h.y & 0xffffu)); }
__device__ __forceinline__ const Texel * texel_pool(const SceneParams & P) { return P.texels_half; }
#else
typedef float4 Texel;
__device__ __forceinline__ v3 texel_rgb(const Texel * p) { return V(load_texel(p)); }
__device__ __forceinline__ const Texel * texel_pool(const SceneParams & P) { return P.texels; }
#endif

struct TexView { const Texel * texels; int width, height, mip_levels; const int * mip_offsets; };

// Texture::fetch_texel, Texture.cpp:131-147 (repeat wrap by positive modulo). POW2: the caller is one of the mip-map
// filters, which only ever see mip-mapped textures, and those have power-of-two sides (Texture.cpp:50).
template<bool COUNT, bool POW2 = false>
__device__ __forceinline__ v3 fetch_texel(const TexView & tex, int x, int y, int level, Counters & cnt) {
	const int level_width  = tex.width  >> level;
	const int level_height = tex.height >> level;
	x = POW2 ? positive_mod_pow2(x, level_width)  : positive_mod(x, level_width);
	y = POW2 ? positive_mod_pow2(y, level_height) : positive_mod(y, level_height);
	if (COUNT) cnt.texels++;
	return texel_rgb(tex.texels + (__ldg(tex.mip_offsets + level) + x + y * level_width));
}

// Texture::sample_bilinear, Texture.cpp:156-186
template<bool COUNT, bool POW2 = false>
__device__ __forceinline__ v3 sample_bilinear(const TexView & tex, float s, float t, int level, Counters & cnt) {
	const int level_width  = tex.width  >> level;
	const int level_height = tex.height >> level;

	s = s * float(level_width)  - 0.5f;
	t = t * float(level_height) - 0.5f;

	const float fractional_s = s - floorf(s);
	const float fractional_t = t - floorf(t);

	const float one_minus_fractional_s = 1.0f - fractional_s;
	const float one_minus_fractional_t = 1.0f - fractional_t;

	const float w0 = one_minus_fractional_s * one_minus_fractional_t;
	const float w1 =           fractional_s * one_minus_fractional_t;
	const float w2 = one_minus_fractional_s *           fractional_t;
	const float w3 = 1.0f - w0 - w1 - w2;

	const int int_s = float_to_int(s - 0.5f);
	const int int_t = float_to_int(t - 0.5f);

	const v3 c0 = fetch_texel<COUNT, POW2>(tex, int_s,     int_t,     level, cnt) * w0;
	const v3 c1 = fetch_texel<COUNT, POW2>(tex, int_s + 1, int_t,     level, cnt) * w1;
	const v3 c2 = fetch_texel<COUNT, POW2>(tex, int_s,     int_t + 1, level, cnt) * w2;
	const v3 c3 = fetch_texel<COUNT, POW2>(tex, int_s + 1, int_t + 1, level, cnt) * w3;
	return ((c0 + c1) + c2) + c3;
}

// Texture::sample_mipmap_trilinear, Texture.cpp:189-204
template<bool COUNT>
__device__ v3 sample_trilinear(const TexView & tex, float s, float t, float ds_dx, float ds_dy, float dt_dx, float dt_dy, Counters & cnt) {
	const float width = 2.0f * std_max(std_max(fabsf(ds_dx), fabsf(ds_dy)), std_max(fabsf(dt_dx), fabsf(dt_dy)));

	const float lambda = float(tex.mip_levels) - 1.0f + log2f(std_max(width, 1e-8f));
	const int   level  = float_to_int(lambda - 0.5f);

	if (level < 0)                   return sample_bilinear<COUNT, true>(tex, s, t, 0, cnt);
	if (level >= tex.mip_levels - 1) return fetch_texel<COUNT, true>(tex, 0, 0, tex.mip_levels - 1, cnt);

	const float f = lambda - floorf(lambda);
	return sample_bilinear<COUNT, true>(tex, s, t, level, cnt) * (1.0f - f) + sample_bilinear<COUNT, true>(tex, s, t, level + 1, cnt) * f;
}

// Texture::sample_mipmap_anisotropic, Texture.cpp:207-239
template<bool COUNT>
__device__ v3 sample_anisotropic(const TexView & tex, float max_anisotropy, float s, float t, float ds_dx, float ds_dy, float dt_dx, float dt_dy, Counters & cnt) {
	const float p_x = std_max(fabsf(ds_dx), fabsf(dt_dx));
	const float p_y = std_max(fabsf(ds_dy), fabsf(dt_dy));

	const float p_min = std_min(p_x, p_y);
	const float p_max = std_max(p_x, p_y);

	const float N = std_min(ceilf(p_max / p_min), max_anisotropy);
	const float one_over_N = 1.0f / N;

	const float lambda = float(tex.mip_levels) - 1.0f + log2f(p_max * one_over_N);
	const int   level  = float_to_int(lambda);

	if (level < 0)                   return sample_bilinear<COUNT, true>(tex, s, t, 0, cnt);
	if (level >= tex.mip_levels - 1) return fetch_texel<COUNT, true>(tex, 0, 0, tex.mip_levels - 1, cnt);

	const bool  x_major = p_x > p_y;
	const float step_s = x_major ? ds_dx : ds_dy;
	const float step_t = x_major ? dt_dx : dt_dy;

	const float one_over_N_plus_1 = 1.0f / (N + 1.0f);

	v3 sum = V(0.0f, 0.0f, 0.0f);
	for (float i = 1.0f; i <= N + 0.001f; i += 1.0f) {
		const float x = s + step_s * (i * one_over_N_plus_1 - 0.5f);
		const float y = t + step_t * (i * one_over_N_plus_1 - 0.5f);
		sum = sum + sample_bilinear<COUNT, true>(tex, x, y, level, cnt);
	}
	return sum * one_over_N;
}

// Texture::sample_mipmap_ewa, Texture.cpp:242-337
template<bool COUNT>
__device__ v3 sample_ewa(const TexView & tex, float max_anisotropy, float s, float t, float ds_dx, float ds_dy, float dt_dx, float dt_dy, Counters & cnt) {
	float major_x = ds_dx, major_y = dt_dx;
	float minor_x = ds_dy, minor_y = dt_dy;

	float major_length = sqrtf(major_x * major_x + major_y * major_y);
	float minor_length = sqrtf(minor_x * minor_x + minor_y * minor_y);

	if (minor_length > major_length) {
		float tx = minor_x, ty = minor_y, tl = minor_length;
		minor_x = major_x; minor_y = major_y; minor_length = major_length;
		major_x = tx; major_y = ty; major_length = tl;
	}

	if (minor_length < 0.00001f) return sample_bilinear<COUNT, true>(tex, s, t, 0, cnt);
	if (major_length > float(tex.width)) return fetch_texel<COUNT, true>(tex, 0, 0, tex.mip_levels - 1, cnt);

	if (minor_length * max_anisotropy < major_length) { // clamp the eccentricity
		const float scale = major_length / (minor_length * max_anisotropy);
		minor_x *= scale; minor_y *= scale;
		minor_length *= scale;
	}

	const float lambda = std_max(0.0f, float(tex.mip_levels) - 1.0f + log2f(minor_length));
	const int   level  = float_to_int(lambda);

	if (level >= tex.mip_levels - 1) return fetch_texel<COUNT, true>(tex, 0, 0, tex.mip_levels - 1, cnt);

	const float level_width  = float(tex.width  >> level);
	const float level_height = float(tex.height >> level);

	s = s * level_width  - 0.5f;
	t = t * level_height - 0.5f;

	const float major_sx = major_x * level_width, major_sy = major_y * level_height;
	const float minor_sx = minor_x * level_width, minor_sy = minor_y * level_height;

	float a =  1.0f + (major_sy * major_sy + minor_sy * minor_sy);
	float b = -2.0f * (major_sx * major_sy + minor_sx * minor_sy);
	float c =  1.0f + (major_sx * major_sx + minor_sx * minor_sx);

	const float one_over_f = 1.0f / (a * c - b * b * 0.25f);
	a *= one_over_f; b *= one_over_f; c *= one_over_f;

	const float det = -b * b + 4.0f * a * c;

	const float sqrt_u = sqrtf(det * c);
	const float sqrt_v = sqrtf(det * a);

	const float two_inv_det = 2.0f / det;
	const float two_inv_det_sqrt_u = two_inv_det * sqrt_u;
	const float two_inv_det_sqrt_v = two_inv_det * sqrt_v;

	const int s0 = float_to_int(s - two_inv_det_sqrt_u + 0.5f);
	const int s1 = float_to_int(s + two_inv_det_sqrt_u - 0.5f);
	const int t0 = float_to_int(t - two_inv_det_sqrt_v + 0.5f);
	const int t1 = float_to_int(t + two_inv_det_sqrt_v - 0.5f);

	v3    sum = V(0.0f, 0.0f, 0.0f);
	float sum_weights = 0.0f;

	float tf = float(t0);
	for (int ti = t0; ti <= t1; ti++, tf += 1.0f) {
		const float tt = tf - t;
		float sf = float(s0);
		for (int si = s0; si <= s1; si++, sf += 1.0f) {
			const float ss = sf - s;
			const float r2 = a * ss * ss + b * ss * tt + c * tt * tt;
			if (r2 < 1.0f) {
				int index = float_to_int(r2 * 128.0f);
				if (index > 127) index = 127;
				const float weight = c_ewa_weight_table[index];
				sum = sum + fetch_texel<COUNT, true>(tex, si, ti, level, cnt) * weight;
				sum_weights += weight;
			}
		}
	}
	return divf(sum, sum_weights);
}

// Material::get_albedo + Texture::sample, Material.h:16-22, Texture.h:33-49
template<bool COUNT>
__device__ __forceinline__ v3 get_albedo(const SceneParams & P, float4 diffuse_texture, const HitRecord & hit, Counters & cnt) {
	const v3 diffuse = V(diffuse_texture);
	const int texture = __float_as_int(diffuse_texture.w);
	if (texture < 0) return diffuse;

	const DeviceTexture * dt = P.textures + texture;
	TexView tex;
	tex.texels = texel_pool(P) + __ldg(&dt->texel_offset);
	tex.width = __ldg(&dt->width); tex.height = __ldg(&dt->height); tex.mip_levels = __ldg(&dt->mip_levels);
	tex.mip_offsets = dt->mip_offsets;

	v3 sample;   // TEXTURE_SAMPLE_MODE (Config.h), then MIPMAP_FILTER
	if (P.texture_sample_mode == B200RT_TEXTURE_NEAREST)  // Texture::sample_nearest, Texture.cpp:149-154
		sample = fetch_texel<COUNT>(tex, float_to_int(hit.u * float(tex.width)), float_to_int(hit.v * float(tex.height)), 0, cnt);
	else if (P.texture_sample_mode == B200RT_TEXTURE_BILINEAR || !__ldg(&dt->mipmapped)) sample = sample_bilinear<COUNT>(tex, hit.u, hit.v, 0, cnt);
	else if (P.mipmap_filter == B200RT_FILTER_TRILINEAR)  sample = sample_trilinear<COUNT>(tex, hit.u, hit.v, hit.ds_dx, hit.ds_dy, hit.dt_dx, hit.dt_dy, cnt);
	else if (P.mipmap_filter == B200RT_FILTER_EWA)        sample = sample_ewa<COUNT>(tex, P.max_anisotropy, hit.u, hit.v, hit.ds_dx, hit.ds_dy, hit.dt_dx, hit.dt_dy, cnt);
	else                                                  sample = sample_anisotropic<COUNT>(tex, P.max_anisotropy, hit.u, hit.v, hit.ds_dx, hit.ds_dy, hit.dt_dx, hit.dt_dy, cnt);
	return diffuse * sample;
}

// --------------------------------------------------------------------------------- sky ---
// Sky::sample, Sky.cpp:28-68 (AVX build: texel index by round-to-nearest-even, SIMD.h:29; the
// index is clamped to [0, size*size] INCLUSIVE, the host pads the probe by one texel).
template<bool COUNT>
__device__ __forceinline__ v3 sky_sample(const SceneParams & P, v3 d, Counters & cnt) {
	const int   size = P.sky_size;
	const float r = 0.5f * RT_ONE_OVER_PI * acosf(d.z) * inv_sqrt(d.x * d.x + d.y * d.y);

	const float u = __fmaf_rn(d.x, r, 0.5f);
	const float v = __fmaf_rn(d.y, r, 0.5f);

	const int x = float_to_int(u * float(size));
	const int y = float_to_int(v * float(size));

	int index = int(unsigned(y) * unsigned(size) + unsigned(x));
	if (index < 0) index = 0;
	if (index > size * size) index = size * size;

	if (COUNT) cnt.texels++;
	return V(__ldg(P.sky + index)) * RT_ONE_OVER_PI;
}

// ------------------------------------------------------------------------------ lights ---
__device__ __forceinline__ float pow128(float x) { // Math::pow2<128>, Math.h:92-96
	x = x * x; x = x * x; x = x * x; x = x * x; x = x * x; x = x * x; x = x * x;
	return x;
}

// Light::calc_lighting, Light.h:12-26 (Blinn-Phong; even power keeps a negative n.h)
__device__ __forceinline__ v3 light_base(v3 colour, v3 normal, v3 to_light, v3 to_camera) {
	float intensity = dot(normal, to_light);
	if (!(intensity > 0.0f)) return V(0.0f, 0.0f, 0.0f);

	const v3 half_angle = normalize(to_light + to_camera);
	const float specular_factor = dot(normal, half_angle);
	intensity = intensity + pow128(specular_factor);
	return colour * intensity;
}

} // namespace RT_NAMESPACE
