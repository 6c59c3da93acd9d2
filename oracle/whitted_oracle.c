/* whitted_oracle.c — TEST INFRASTRUCTURE. CPU restatement of the reference's per-frame render
 * hot path (clayne/CPU-Raytracer), one ray at a time, in plain C.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
 * load this library, and only as the CHECKER. The product (libb200rt.so) never links or calls it.
 *
 * Parity status: PINNED. The oracle is checked (tests/test_oracle_vs_reference.py, here; and the
 * committed fixtures under tests/golden/, everywhere) against outputs of the real reference built
 * by oracle/ref_recipe/build_ref.py, run in its --per-ray mode (one ray per packet through the
 * reference's own Raytracer::bounce) from the "_ids" variants, which are compiled with
 * -ffp-contract=off. Under that flag a multiply-add is fused exactly where the reference wrote a
 * madd/msub intrinsic (SIMD_floats.h:223-224 -> _mm256_fmadd_ps/_mm256_fmsub_ps) and nowhere
 * else; this file spells those as fmaf() and must itself be compiled with -ffp-contract=off.
 * Not pinned bit-for-bit: SVML transcendentals (exp, acos, atan2), which the reference build gets
 * from glibc libmvec and this file from glibc libm (a few ulp).
 *
 * Semantics are those of the reference's AVX build (SIMD_LANE_SIZE 8) for a packet whose lanes all
 * carry the same ray: traversal order from that ray's direction signs (BVHNode.h:30-40), every
 * all/any early-out decided by that ray alone. Every function cites the reference lines it follows.
 */
#include <math.h>
#include <pthread.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#include "../include/b200rt.h"
#include "whitted_oracle.h"

#define EPSILON 0.005f            /* Ray::EPSILON, Ray.h:5 */
#define ONE_OVER_PI     0.31830988618f /* Util.h:9  */
#define ONE_OVER_TWO_PI 0.15915494309f /* Util.h:12 */
#define STACK_SIZE 64             /* BVH_TRAVERSAL_STACK_SIZE, Config.h:25 */

typedef struct { float x, y, z; } v3;

typedef struct {           /* Ray.h:7-16 */
	v3 origin, direction;
	v3 dO_dx, dO_dy, dD_dx, dD_dy;
} ray_t;

typedef struct {           /* RayHit.h:4-22 (+ ids, as in the instrumented reference build) */
	int   hit;
	float distance;
	v3    point, normal;
	int   material_id;
	float u, v;
	float ds_dx, ds_dy, dt_dx, dt_dy;
	v3    dO_dx, dO_dy, dN_dx, dN_dy;
	int   instance_id, triangle_id;
} hit_t;

typedef struct {           /* per-thread counters */
	uint64_t primary, shadow, reflection, refraction;
	uint64_t nodes, triangles, texels, accepted, instances;
} counters_t;

typedef struct {
	const b200rt_static_scene * s;
	const b200rt_frame        * f;
	const b200rt_config       * c;
	float ewa_weight_table[128];
} ctx_t;

/* ---------------------------------------------------------------- SIMD_Vector3.h ------- */
static inline v3 V(float x, float y, float z) { v3 r = { x, y, z }; return r; }
static inline v3 V3(const float * p) { v3 r = { p[0], p[1], p[2] }; return r; }
static inline v3 add(v3 a, v3 b) { return V(a.x + b.x, a.y + b.y, a.z + b.z); }
static inline v3 sub(v3 a, v3 b) { return V(a.x - b.x, a.y - b.y, a.z - b.z); }
static inline v3 mul(v3 a, v3 b) { return V(a.x * b.x, a.y * b.y, a.z * b.z); }
static inline v3 scale(v3 a, float f) { return V(a.x * f, a.y * f, a.z * f); }          /* v * f and f * v, :150-158 */
static inline v3 addf(v3 a, float f) { return V(a.x + f, a.y + f, a.z + f); }            /* f + v, :155 */
static inline v3 neg(v3 a) { return V(0.0f - a.x, 0.0f - a.y, 0.0f - a.z); }             /* :138-141 */
static inline float dot(v3 a, v3 b) { return fmaf(a.x, b.x, fmaf(a.y, b.y, a.z * b.z)); } /* :53-55 */
static inline v3 cross(v3 a, v3 b) {                                                     /* :57-63: msub(a,b,c) = a*b - c */
	return V(fmaf(a.y, b.z, -(a.z * b.y)), fmaf(a.z, b.x, -(a.x * b.z)), fmaf(a.x, b.y, -(a.y * b.x)));
}
static inline v3 madd(v3 a, float b, v3 c) { return V(fmaf(a.x, b, c.x), fmaf(a.y, b, c.y), fmaf(a.z, b, c.z)); } /* :107-113 */
static inline float inv_sqrt(float x) { return 1.0f / sqrtf(x); }   /* _mm256_invsqrt_ps, served as 1/sqrt by the shim */
static inline v3 normalize(v3 a) { return scale(a, inv_sqrt(dot(a, a))); }               /* :44-51 */
static inline v3 divf(v3 a, float f) { return scale(a, 1.0f / f); }                      /* v / f = v * rcp(f), :153 */

/* x86 min/max: the SECOND operand is returned when either is NaN (_mm256_min_ps) */
static inline float sse_min(float a, float b) { return a < b ? a : b; }
static inline float sse_max(float a, float b) { return a > b ? a : b; }

/* Util::float_to_int / _mm256_cvtps_epi32: round to nearest even, NaN and overflow -> INT_MIN */
static inline int float_to_int(float x) {
	if (!(x >= -2147483648.0f && x < 2147483648.0f)) return (int)0x80000000u;
	return (int)lrintf(x);
}

/* Matrix4.h:39-58 / :69-85 — nested fused multiply-adds, column-major cells[i + 4j] */
static inline v3 transform_position(const float * m, v3 p) {
	return V(fmaf(m[0], p.x, fmaf(m[1], p.y, fmaf(m[2],  p.z, m[3]))),
	         fmaf(m[4], p.x, fmaf(m[5], p.y, fmaf(m[6],  p.z, m[7]))),
	         fmaf(m[8], p.x, fmaf(m[9], p.y, fmaf(m[10], p.z, m[11]))));
}
static inline v3 transform_direction(const float * m, v3 d) {
	return V(fmaf(m[0], d.x, fmaf(m[1], d.y, m[2]  * d.z)),
	         fmaf(m[4], d.x, fmaf(m[5], d.y, m[6]  * d.z)),
	         fmaf(m[8], d.x, fmaf(m[9], d.y, m[10] * d.z)));
}

/* ---------------------------------------------------------------- AABB.cpp:38-52 ------- */
static inline int aabb_intersect(const b200rt_bvh_node * n, v3 o, v3 inv_d, float max_distance) {
	float t0x = (n->aabb_min[0] - o.x) * inv_d.x, t1x = (n->aabb_max[0] - o.x) * inv_d.x;
	float t0y = (n->aabb_min[1] - o.y) * inv_d.y, t1y = (n->aabb_max[1] - o.y) * inv_d.y;
	float t0z = (n->aabb_min[2] - o.z) * inv_d.z, t1z = (n->aabb_max[2] - o.z) * inv_d.z;

	float tminx = sse_min(t0x, t1x), tmaxx = sse_max(t0x, t1x);
	float tminy = sse_min(t0y, t1y), tmaxy = sse_max(t0y, t1y);
	float tminz = sse_min(t0z, t1z), tmaxz = sse_max(t0z, t1z);

	float t_near = sse_max(sse_max(EPSILON,      tminx), sse_max(tminy, tminz));
	float t_far  = sse_min(sse_min(max_distance, tmaxx), sse_min(tmaxy, tmaxz));
	return t_near < t_far;
}

/* BVHNode.h:18-40 */
static inline int node_count(const b200rt_bvh_node * n) { return n->count & 0x3FFFFFFF; }
static inline int visit_left_first(const ctx_t * ctx, const b200rt_bvh_node * n, v3 d) {
	if (ctx->c->traversal_naive) return 1; /* BVH_TRAVERSE_TREE_NAIVE, BVHNode.h:31-32 */
	switch ((uint32_t)n->count >> 30) {
		case 1:  return d.x > 0.0f;
		case 2:  return d.y > 0.0f;
		default: return d.z > 0.0f; /* 3; an axis of 0 cannot occur on an inner node */
	}
}

/* ------------------------------------------------- BottomLevelBVH.cpp:214-309 / :311-346 --- */
/* Moller-Trumbore; all six comparisons strict. Returns 1 and fills t, u, v on acceptance. */
static inline int triangle_test(const b200rt_triangle_hot * tri, v3 o, v3 d, float max_t, float * t_out, float * u_out, float * v_out) {
	v3 edge_1 = V3(tri->position_edge_1), edge_2 = V3(tri->position_edge_2);

	v3    h = cross(d, edge_2);
	float a = dot(edge_1, h);
	float f = 1.0f / a;
	v3    s = sub(o, V3(tri->position_0));
	float u = f * dot(s, h);
	if (!(u > 0.0f && u < 1.0f)) return 0;

	v3    q = cross(s, edge_1);
	float v = f * dot(d, q);
	if (!(v > 0.0f && (u + v) < 1.0f)) return 0;

	float t = f * dot(edge_2, q);
	if (!(t > EPSILON && t < max_t)) return 0;

	*t_out = t; *u_out = u; *v_out = v;
	return 1;
}

/* Hit record of an accepted triangle hit, BottomLevelBVH.cpp:254-308 */
static void triangle_hit_record(const b200rt_triangle_hot * hot, const b200rt_triangle_cold * cold, int material_offset,
                                const ray_t * ray, const float * world, float t, float u, float v, hit_t * hit) {
	v3 edge_1 = V3(hot->position_edge_1), edge_2 = V3(hot->position_edge_2);
	v3 n_edge_1 = V3(cold->normal_edge_1), n_edge_2 = V3(cold->normal_edge_2);

	/* Math::barycentric: base + u * edge1 + v * edge2, Math.h:22-25 */
	v3 n = add(add(V3(cold->normal_0), scale(n_edge_1, u)), scale(n_edge_2, v));

	hit->hit      = 1;
	hit->distance = t;
	hit->point    = transform_position (world, add(ray->origin, scale(ray->direction, t)));
	hit->normal   = transform_direction(world, normalize(n));
	hit->material_id = material_offset + cold->material_id;

	hit->u = (cold->tex_coord_0[0] + u * cold->tex_coord_edge_1[0]) + v * cold->tex_coord_edge_2[0];
	hit->v = (cold->tex_coord_0[1] + u * cold->tex_coord_edge_1[1]) + v * cold->tex_coord_edge_2[1];

	/* Ray Tracing Gems ch. 20 differentials, :278-308 */
	float one_over_k = 1.0f / dot(cross(edge_1, edge_2), ray->direction);

	v3 q = madd(ray->dD_dx, t, ray->dO_dx);
	v3 r = madd(ray->dD_dy, t, ray->dO_dy);

	v3 c_u = cross(edge_2, ray->direction);
	v3 c_v = cross(ray->direction, edge_1);

	float du_dx = one_over_k * dot(c_u, q), du_dy = one_over_k * dot(c_u, r);
	float dv_dx = one_over_k * dot(c_v, q), dv_dy = one_over_k * dot(c_v, r);

	hit->dO_dx = add(scale(edge_1, du_dx), scale(edge_2, dv_dx)); /* object space on purpose, SURVEY App. A.5 */
	hit->dO_dy = add(scale(edge_1, du_dy), scale(edge_2, dv_dy));

	v3 dn_dx = add(scale(n_edge_1, du_dx), scale(n_edge_2, dv_dx));
	v3 dn_dy = add(scale(n_edge_1, du_dy), scale(n_edge_2, dv_dy));

	float n_dot_n = dot(n, n);
	float N_denom = inv_sqrt(n_dot_n) / n_dot_n;

	hit->dN_dx = scale(sub(scale(dn_dx, n_dot_n), scale(n, dot(n, dn_dx))), N_denom);
	hit->dN_dy = scale(sub(scale(dn_dy, n_dot_n), scale(n, dot(n, dn_dy))), N_denom);

	hit->ds_dx = du_dx * cold->tex_coord_edge_1[0] + dv_dx * cold->tex_coord_edge_2[0];
	hit->ds_dy = du_dy * cold->tex_coord_edge_1[0] + dv_dy * cold->tex_coord_edge_2[0];
	hit->dt_dx = du_dx * cold->tex_coord_edge_1[1] + dv_dx * cold->tex_coord_edge_2[1];
	hit->dt_dy = du_dy * cold->tex_coord_edge_1[1] + dv_dy * cold->tex_coord_edge_2[1];
}

/* ---------------------------------------------- BottomLevelBVH::trace, :355-396 ---------- */
static void blas_trace(const ctx_t * ctx, const b200rt_blas * blas, const ray_t * ray, const float * world, int instance, hit_t * hit, counters_t * cnt) {
	const b200rt_bvh_node      * nodes = ctx->s->nodes + blas->node_offset;
	const b200rt_triangle_hot  * hot   = ctx->s->triangles_hot  + blas->triangle_offset;
	const b200rt_triangle_cold * cold  = ctx->s->triangles_cold + blas->triangle_offset;

	int stack[STACK_SIZE];
	int stack_size = 1;
	stack[0] = 0;

	v3 inv_direction = V(1.0f / ray->direction.x, 1.0f / ray->direction.y, 1.0f / ray->direction.z);

	while (stack_size > 0) {
		const b200rt_bvh_node * node = &nodes[stack[--stack_size]];
		cnt->nodes++;

		if (!aabb_intersect(node, ray->origin, inv_direction, hit->distance)) continue;

		if (node_count(node) > 0) {
			for (int i = node->left_or_first; i < node->left_or_first + node->count; i++) {
				float t, u, v;
				cnt->triangles++;
				if (triangle_test(&hot[i], ray->origin, ray->direction, hit->distance, &t, &u, &v)) {
					cnt->accepted++;
					triangle_hit_record(&hot[i], &cold[i], blas->material_offset, ray, world, t, u, v, hit);
					hit->instance_id = instance;
					hit->triangle_id = i;
				}
			}
		} else if (visit_left_first(ctx, node, ray->direction)) {
			stack[stack_size++] = node->left_or_first + 1;
			stack[stack_size++] = node->left_or_first;
		} else {
			stack[stack_size++] = node->left_or_first;
			stack[stack_size++] = node->left_or_first + 1;
		}
	}
}

/* BottomLevelBVH::intersect, :398-437 — any hit closer than max_distance */
static int blas_intersect(const ctx_t * ctx, const b200rt_blas * blas, v3 origin, v3 direction, float max_distance, counters_t * cnt) {
	const b200rt_bvh_node     * nodes = ctx->s->nodes + blas->node_offset;
	const b200rt_triangle_hot * hot   = ctx->s->triangles_hot + blas->triangle_offset;

	int stack[STACK_SIZE];
	int stack_size = 1;
	stack[0] = 0;

	v3 inv_direction = V(1.0f / direction.x, 1.0f / direction.y, 1.0f / direction.z);

	while (stack_size > 0) {
		const b200rt_bvh_node * node = &nodes[stack[--stack_size]];
		cnt->nodes++;

		if (!aabb_intersect(node, origin, inv_direction, max_distance)) continue;

		if (node_count(node) > 0) {
			for (int i = node->left_or_first; i < node->left_or_first + node->count; i++) {
				float t, u, v;
				cnt->triangles++;
				if (triangle_test(&hot[i], origin, direction, max_distance, &t, &u, &v)) return 1;
			}
		} else if (visit_left_first(ctx, node, direction)) {
			stack[stack_size++] = node->left_or_first + 1;
			stack[stack_size++] = node->left_or_first;
		} else {
			stack[stack_size++] = node->left_or_first;
			stack[stack_size++] = node->left_or_first + 1;
		}
	}
	return 0;
}

/* Mesh::trace, Mesh.cpp:17-31: ray (and its four differentials) into object space */
static void mesh_trace(const ctx_t * ctx, int instance, const ray_t * ray, hit_t * hit, counters_t * cnt) {
	const b200rt_instance * inst = &ctx->f->instances[instance];
	cnt->instances++;
	ray_t local;
	local.origin    = transform_position (inst->inverse, ray->origin);
	local.direction = transform_direction(inst->inverse, ray->direction);
	local.dO_dx     = transform_direction(inst->inverse, ray->dO_dx);
	local.dO_dy     = transform_direction(inst->inverse, ray->dO_dy);
	local.dD_dx     = transform_direction(inst->inverse, ray->dD_dx);
	local.dD_dy     = transform_direction(inst->inverse, ray->dD_dy);
	blas_trace(ctx, &ctx->s->blas[inst->blas], &local, inst->world, instance, hit, cnt);
}

/* Mesh::intersect, Mesh.cpp:33-40 */
static int mesh_intersect(const ctx_t * ctx, int instance, v3 origin, v3 direction, float max_distance, counters_t * cnt) {
	const b200rt_instance * inst = &ctx->f->instances[instance];
	cnt->instances++;
	return blas_intersect(ctx, &ctx->s->blas[inst->blas], transform_position(inst->inverse, origin), transform_direction(inst->inverse, direction), max_distance, cnt);
}

/* TopLevelBVH::trace, TopLevelBVH.cpp:53-87 */
static void tlas_trace(const ctx_t * ctx, const ray_t * ray, hit_t * hit, counters_t * cnt) {
	const b200rt_bvh_node * nodes = ctx->f->tlas_nodes;

	int stack[STACK_SIZE];
	int stack_size = 1;
	stack[0] = 0;

	v3 inv_direction = V(1.0f / ray->direction.x, 1.0f / ray->direction.y, 1.0f / ray->direction.z);

	while (stack_size > 0) {
		const b200rt_bvh_node * node = &nodes[stack[--stack_size]];
		cnt->nodes++;

		if (!aabb_intersect(node, ray->origin, inv_direction, hit->distance)) continue;

		if (node_count(node) > 0) {
			for (int i = node->left_or_first; i < node->left_or_first + node->count; i++) mesh_trace(ctx, ctx->f->tlas_indices[i], ray, hit, cnt);
		} else if (visit_left_first(ctx, node, ray->direction)) {
			stack[stack_size++] = node->left_or_first + 1;
			stack[stack_size++] = node->left_or_first;
		} else {
			stack[stack_size++] = node->left_or_first;
			stack[stack_size++] = node->left_or_first + 1;
		}
	}
}

/* TopLevelBVH::intersect, TopLevelBVH.cpp:89-129 */
static int tlas_intersect(const ctx_t * ctx, v3 origin, v3 direction, float max_distance, counters_t * cnt) {
	const b200rt_bvh_node * nodes = ctx->f->tlas_nodes;

	int stack[STACK_SIZE];
	int stack_size = 1;
	stack[0] = 0;

	v3 inv_direction = V(1.0f / direction.x, 1.0f / direction.y, 1.0f / direction.z);

	while (stack_size > 0) {
		const b200rt_bvh_node * node = &nodes[stack[--stack_size]];
		cnt->nodes++;

		if (!aabb_intersect(node, origin, inv_direction, max_distance)) continue;

		if (node_count(node) > 0) {
			for (int i = node->left_or_first; i < node->left_or_first + node->count; i++) {
				if (mesh_intersect(ctx, ctx->f->tlas_indices[i], origin, direction, max_distance, cnt)) return 1;
			}
		} else if (visit_left_first(ctx, node, direction)) {
			stack[stack_size++] = node->left_or_first + 1;
			stack[stack_size++] = node->left_or_first;
		} else {
			stack[stack_size++] = node->left_or_first;
			stack[stack_size++] = node->left_or_first + 1;
		}
	}
	return 0;
}

/* ------------------------------------------------------------ Sphere.cpp:9-90 --------- */
static void sphere_trace(const b200rt_sphere * sphere, int index, const ray_t * ray, hit_t * hit) {
	v3 center = V3(sphere->center);

	v3 oc = sub(ray->origin, center);
	float a = dot(ray->direction, ray->direction);
	float b = 2.0f * dot(oc, ray->direction);
	float c = dot(oc, oc) - sphere->radius_squared;

	float d = b * b - 4.0f * a * c;
	if (!(d >= 0.0f)) return;

	float sqrt_d    = sqrtf(d);
	float inv_denom = 0.0f - 1.0f / (2.0f * a);

	float t0 = (b + sqrt_d) * inv_denom;
	float t1 = (b - sqrt_d) * inv_denom;

	float t = t0 > EPSILON ? t0 : t1; /* the nearer root unless it lies behind the origin, :32-36 */
	if (!(t > EPSILON && t < hit->distance)) return;

	hit->hit      = 1;
	hit->distance = t;
	hit->point    = add(ray->origin, scale(ray->direction, t));
	hit->normal   = scale(sub(hit->point, center), sphere->radius_inv);
	hit->material_id = sphere->material_id;
	hit->instance_id = -2; hit->triangle_id = index;

	/* spherical coordinates of the normal, :60-61 */
	hit->u = fmaf(atan2f(hit->normal.z, hit->normal.x), ONE_OVER_TWO_PI, 0.5f);
	hit->v = fmaf(acosf (hit->normal.y),                ONE_OVER_PI,     0.5f);

	/* Igehy transfer, :65-78 */
	v3 a_x = madd(ray->dD_dx, t, ray->dO_dx);
	v3 a_y = madd(ray->dD_dy, t, ray->dO_dy);

	float denom = (0.0f - 1.0f) / dot(ray->direction, hit->normal);
	float dt_dx = dot(a_x, hit->normal) * denom;
	float dt_dy = dot(a_y, hit->normal) * denom;

	v3 dP_dx = madd(ray->direction, dt_dx, a_x);
	v3 dP_dy = madd(ray->direction, dt_dy, a_y);

	hit->dO_dx = dP_dx;
	hit->dO_dy = dP_dy;
	hit->dN_dx = scale(dP_dx, sphere->radius_inv);
	hit->dN_dy = scale(dP_dy, sphere->radius_inv);

	/* derivatives of (u, v), :80-88 */
	v3 n = hit->normal;
	float ds_denom = ONE_OVER_TWO_PI / (n.x * n.x + n.z * n.z + 1e-8f);
	hit->ds_dx = (n.x * hit->dN_dx.z - n.z * hit->dN_dx.x) * ds_denom;
	hit->ds_dy = (n.x * hit->dN_dy.z - n.z * hit->dN_dy.x) * ds_denom;

	float dt_denom = (0.0f - ONE_OVER_PI) * inv_sqrt(1.0f - n.y * n.y + 1e-8f);
	hit->dt_dx = hit->dN_dx.y * dt_denom;
	hit->dt_dy = hit->dN_dy.y * dt_denom;
}

/* Sphere::intersect, Sphere.cpp:92-112 — geometric form, near root only */
static int sphere_intersect(const b200rt_sphere * sphere, v3 origin, v3 direction, float max_distance) {
	v3    c  = sub(V3(sphere->center), origin);
	float t  = dot(c, direction);
	v3    Q  = sub(c, scale(direction, t));
	float p2 = dot(Q, Q);

	if (!(p2 < sphere->radius_squared)) return 0;

	t = t - sqrtf(sphere->radius_squared - p2);
	return t > EPSILON && t < max_distance;
}

/* ------------------------------------------------------------ Plane.cpp:13-79 --------- */
static void plane_trace(const b200rt_plane * plane, int index, const ray_t * ray, hit_t * hit) {
	v3 normal = V3(plane->normal);

	float t = (0.0f - (dot(normal, ray->origin) + plane->distance)) / dot(normal, ray->direction);
	if (!(t > EPSILON && t < hit->distance)) return;

	hit->hit      = 1;
	hit->distance = t;
	hit->point    = add(ray->origin, scale(ray->direction, t));
	hit->normal   = normal;
	hit->material_id = plane->material_id;
	hit->instance_id = -3; hit->triangle_id = index;

	v3 u = V3(plane->u_axis), v = V3(plane->v_axis);
	hit->u = dot(hit->point, u);
	hit->v = dot(hit->point, v);

	v3 a_x = madd(ray->dD_dx, t, ray->dO_dx);
	v3 a_y = madd(ray->dD_dy, t, ray->dO_dy);

	float denom = (0.0f - 1.0f) / (dot(ray->direction, hit->normal) + 1e-8f);
	float dt_dx = dot(a_x, hit->normal) * denom;
	float dt_dy = dot(a_y, hit->normal) * denom;

	v3 dP_dx = madd(ray->direction, dt_dx, a_x);
	v3 dP_dy = madd(ray->direction, dt_dy, a_y);

	hit->dO_dx = dP_dx;
	hit->dO_dy = dP_dy;
	hit->dN_dx = V(0.0f, 0.0f, 0.0f);
	hit->dN_dy = V(0.0f, 0.0f, 0.0f);

	hit->ds_dx = dot(dP_dx, u); hit->ds_dy = dot(dP_dy, u);
	hit->dt_dx = dot(dP_dx, v); hit->dt_dy = dot(dP_dy, v);
}

static int plane_intersect(const b200rt_plane * plane, v3 origin, v3 direction, float max_distance) {
	v3 normal = V3(plane->normal);
	float t = (0.0f - (dot(normal, origin) + plane->distance)) / dot(normal, direction);
	return t > EPSILON && t < max_distance;
}

/* ------------------------------------------------------ Scene.cpp:173-190 ------------ */
static void trace_primitives(const ctx_t * ctx, const ray_t * ray, hit_t * hit, counters_t * cnt) {
	for (int i = 0; i < ctx->f->sphere_count; i++) sphere_trace(&ctx->f->spheres[i], i, ray, hit);
	for (int i = 0; i < ctx->f->plane_count;  i++) plane_trace (&ctx->f->planes[i],  i, ray, hit);
	tlas_trace(ctx, ray, hit, cnt);
}

static int intersect_primitives(const ctx_t * ctx, v3 origin, v3 direction, float max_distance, counters_t * cnt) {
	for (int i = 0; i < ctx->f->sphere_count; i++) if (sphere_intersect(&ctx->f->spheres[i], origin, direction, max_distance)) return 1;
	for (int i = 0; i < ctx->f->plane_count;  i++) if (plane_intersect (&ctx->f->planes[i],  origin, direction, max_distance)) return 1;
	return tlas_intersect(ctx, origin, direction, max_distance, cnt);
}

/* --------------------------------------------------------------- Texture.cpp ---------- */
static inline int positive_mod(int value, int modulus) { /* Math::mod, Math.h:46-53 */
	int result = value % modulus;
	if (result < 0) result += modulus;
	return result;
}

/* Texture::fetch_texel, Texture.cpp:131-147 — repeat wrap */
static inline v3 fetch_texel(const ctx_t * ctx, const b200rt_texture * tex, int x, int y, int level, counters_t * cnt) {
	int level_width  = tex->width  >> level;
	int level_height = tex->height >> level;
	x = positive_mod(x, level_width);
	y = positive_mod(y, level_height);
	cnt->texels++;
	return V3(ctx->s->texels + 3 * (tex->texel_offset + tex->mip_offsets[level] + x + (int64_t)y * level_width));
}

/* Texture::sample_bilinear, Texture.cpp:156-186 */
static v3 sample_bilinear(const ctx_t * ctx, const b200rt_texture * tex, float s, float t, int level, counters_t * cnt) {
	int level_width  = tex->width  >> level;
	int level_height = tex->height >> level;

	s = s * level_width  - 0.5f;
	t = t * level_height - 0.5f;

	float fractional_s = s - floorf(s);
	float fractional_t = t - floorf(t);

	float one_minus_fractional_s = 1.0f - fractional_s;
	float one_minus_fractional_t = 1.0f - fractional_t;

	float w0 = one_minus_fractional_s * one_minus_fractional_t;
	float w1 =           fractional_s * one_minus_fractional_t;
	float w2 = one_minus_fractional_s *           fractional_t;
	float w3 = 1.0f - w0 - w1 - w2;

	int int_s = float_to_int(s - 0.5f);
	int int_t = float_to_int(t - 0.5f);

	v3 c0 = scale(fetch_texel(ctx, tex, int_s,     int_t,     level, cnt), w0);
	v3 c1 = scale(fetch_texel(ctx, tex, int_s + 1, int_t,     level, cnt), w1);
	v3 c2 = scale(fetch_texel(ctx, tex, int_s,     int_t + 1, level, cnt), w2);
	v3 c3 = scale(fetch_texel(ctx, tex, int_s + 1, int_t + 1, level, cnt), w3);
	return add(add(add(c0, c1), c2), c3);
}

static inline float maxf(float a, float b) { return a < b ? b : a; } /* std::max */
static inline float minf(float a, float b) { return b < a ? b : a; } /* std::min */

/* Texture::sample_mipmap_trilinear, Texture.cpp:189-204 */
static v3 sample_trilinear(const ctx_t * ctx, const b200rt_texture * tex, float s, float t, float ds_dx, float ds_dy, float dt_dx, float dt_dy, counters_t * cnt) {
	float width = 2.0f * maxf(maxf(fabsf(ds_dx), fabsf(ds_dy)), maxf(fabsf(dt_dx), fabsf(dt_dy)));

	float lambda = (float)tex->mip_levels - 1.0f + log2f(maxf(width, 1e-8f));
	int   level  = float_to_int(lambda - 0.5f);

	if (level < 0)                    return sample_bilinear(ctx, tex, s, t, 0, cnt);
	if (level >= tex->mip_levels - 1) return fetch_texel(ctx, tex, 0, 0, tex->mip_levels - 1, cnt);

	float f = lambda - floorf(lambda);
	return add(scale(sample_bilinear(ctx, tex, s, t, level, cnt), 1.0f - f), scale(sample_bilinear(ctx, tex, s, t, level + 1, cnt), f));
}

/* Texture::sample_mipmap_anisotropic, Texture.cpp:207-239 */
static v3 sample_anisotropic(const ctx_t * ctx, const b200rt_texture * tex, float s, float t, float ds_dx, float ds_dy, float dt_dx, float dt_dy, counters_t * cnt) {
	float p_x = maxf(fabsf(ds_dx), fabsf(dt_dx));
	float p_y = maxf(fabsf(ds_dy), fabsf(dt_dy));

	float p_min = minf(p_x, p_y);
	float p_max = maxf(p_x, p_y);

	float N = minf(ceilf(p_max / p_min), ctx->c->max_anisotropy);
	float one_over_N = 1.0f / N;

	float lambda = (float)tex->mip_levels - 1.0f + log2f(p_max * one_over_N);
	int   level  = float_to_int(lambda);

	if (level < 0)                    return sample_bilinear(ctx, tex, s, t, 0, cnt);
	if (level >= tex->mip_levels - 1) return fetch_texel(ctx, tex, 0, 0, tex->mip_levels - 1, cnt);

	int   x_major = p_x > p_y;
	float step_s = x_major ? ds_dx : ds_dy;
	float step_t = x_major ? dt_dx : dt_dy;

	float one_over_N_plus_1 = 1.0f / (N + 1.0f);

	v3 sum = V(0.0f, 0.0f, 0.0f);
	for (float i = 1.0f; i <= N + 0.001f; i += 1.0f) {
		float x = s + step_s * (i * one_over_N_plus_1 - 0.5f);
		float y = t + step_t * (i * one_over_N_plus_1 - 0.5f);
		sum = add(sum, sample_bilinear(ctx, tex, x, y, level, cnt));
	}
	return scale(sum, one_over_N);
}

/* Texture::sample_mipmap_ewa, Texture.cpp:242-337 */
static v3 sample_ewa(const ctx_t * ctx, const b200rt_texture * tex, float s, float t, float ds_dx, float ds_dy, float dt_dx, float dt_dy, counters_t * cnt) {
	float major_x = ds_dx, major_y = dt_dx;
	float minor_x = ds_dy, minor_y = dt_dy;

	float major_length = sqrtf(major_x * major_x + major_y * major_y);
	float minor_length = sqrtf(minor_x * minor_x + minor_y * minor_y);

	if (minor_length > major_length) {
		float tx = minor_x, ty = minor_y, tl = minor_length;
		minor_x = major_x; minor_y = major_y; minor_length = major_length;
		major_x = tx; major_y = ty; major_length = tl;
	}

	if (minor_length < 0.00001f) return sample_bilinear(ctx, tex, s, t, 0, cnt);
	if (major_length > (float)tex->width) return fetch_texel(ctx, tex, 0, 0, tex->mip_levels - 1, cnt);

	if (minor_length * ctx->c->max_anisotropy < major_length) {
		float scale_f = major_length / (minor_length * ctx->c->max_anisotropy);
		minor_x *= scale_f; minor_y *= scale_f;
		minor_length *= scale_f;
	}

	float lambda = maxf(0.0f, (float)tex->mip_levels - 1.0f + log2f(minor_length));
	int   level  = float_to_int(lambda);

	if (level >= tex->mip_levels - 1) return fetch_texel(ctx, tex, 0, 0, tex->mip_levels - 1, cnt);

	float level_width  = (float)(tex->width  >> level);
	float level_height = (float)(tex->height >> level);

	s = s * level_width  - 0.5f;
	t = t * level_height - 0.5f;

	float major_sx = major_x * level_width, major_sy = major_y * level_height;
	float minor_sx = minor_x * level_width, minor_sy = minor_y * level_height;

	float a =  1.0f + (major_sy * major_sy + minor_sy * minor_sy);
	float b = -2.0f * (major_sx * major_sy + minor_sx * minor_sy);
	float c =  1.0f + (major_sx * major_sx + minor_sx * minor_sx);

	float one_over_f = 1.0f / (a * c - b * b * 0.25f);
	a *= one_over_f; b *= one_over_f; c *= one_over_f;

	float det = -b * b + 4.0f * a * c;

	float sqrt_u = sqrtf(det * c);
	float sqrt_v = sqrtf(det * a);

	float two_inv_det = 2.0f / det;
	float two_inv_det_sqrt_u = two_inv_det * sqrt_u;
	float two_inv_det_sqrt_v = two_inv_det * sqrt_v;

	int s0 = float_to_int(s - two_inv_det_sqrt_u + 0.5f);
	int s1 = float_to_int(s + two_inv_det_sqrt_u - 0.5f);
	int t0 = float_to_int(t - two_inv_det_sqrt_v + 0.5f);
	int t1 = float_to_int(t + two_inv_det_sqrt_v - 0.5f);

	v3    sum = V(0.0f, 0.0f, 0.0f);
	float sum_weights = 0.0f;

	float tf = (float)t0;
	for (int ti = t0; ti <= t1; ti++, tf += 1.0f) {
		float tt = tf - t;
		float sf = (float)s0;
		for (int si = s0; si <= s1; si++, sf += 1.0f) {
			float ss = sf - s;
			float r2 = a * ss * ss + b * ss * tt + c * tt * tt;
			if (r2 < 1.0f) {
				int index = float_to_int(r2 * 128.0f);
				if (index > 127) index = 127;
				float weight = ctx->ewa_weight_table[index];
				sum = add(sum, scale(fetch_texel(ctx, tex, si, ti, level, cnt), weight));
				sum_weights += weight;
			}
		}
	}
	return divf(sum, sum_weights);
}

/* Material::get_albedo + Texture::sample, Material.h:16-22, Texture.h:33-49 */
static v3 get_albedo(const ctx_t * ctx, const b200rt_material * m, const hit_t * hit, counters_t * cnt) {
	v3 diffuse = V3(m->diffuse);
	if (m->texture < 0) return diffuse;

	const b200rt_texture * tex = &ctx->s->textures[m->texture];
	v3 sample;
	if (ctx->c->texture_sample_mode == B200RT_TEXTURE_NEAREST) /* Texture::sample_nearest, Texture.cpp:149-154 */
		sample = fetch_texel(ctx, tex, float_to_int(hit->u * (float)tex->width), float_to_int(hit->v * (float)tex->height), 0, cnt);
	else if (ctx->c->texture_sample_mode == B200RT_TEXTURE_BILINEAR || !tex->mipmapped) sample = sample_bilinear(ctx, tex, hit->u, hit->v, 0, cnt);
	else switch (ctx->c->mipmap_filter) {
		case B200RT_FILTER_TRILINEAR: sample = sample_trilinear  (ctx, tex, hit->u, hit->v, hit->ds_dx, hit->ds_dy, hit->dt_dx, hit->dt_dy, cnt); break;
		case B200RT_FILTER_EWA:       sample = sample_ewa        (ctx, tex, hit->u, hit->v, hit->ds_dx, hit->ds_dy, hit->dt_dx, hit->dt_dy, cnt); break;
		default:                      sample = sample_anisotropic(ctx, tex, hit->u, hit->v, hit->ds_dx, hit->ds_dy, hit->dt_dx, hit->dt_dy, cnt); break;
	}
	return mul(diffuse, sample);
}

/* --------------------------------------------------------------- Sky.cpp:28-68 -------- */
static v3 sky_sample(const ctx_t * ctx, v3 d, counters_t * cnt) {
	int   size = ctx->s->sky_size;
	float r = 0.5f * ONE_OVER_PI * acosf(d.z) * inv_sqrt(d.x * d.x + d.y * d.y);

	float u = fmaf(d.x, r, 0.5f);
	float v = fmaf(d.y, r, 0.5f);

	int x = float_to_int(u * (float)size); /* _mm256_cvtps_epi32: round to nearest even (SIMD.h:29) */
	int y = float_to_int(v * (float)size);

	int index = (int)((uint32_t)y * (uint32_t)size + (uint32_t)x); /* _mm256_mullo_epi32 / add wrap */
	if (index < 0) index = 0;
	if (index > size * size) index = size * size; /* INCLUSIVE upper bound, Sky.cpp:45 */

	cnt->texels++;
	return scale(V3(ctx->s->sky + 3 * (size_t)index), ONE_OVER_PI);
}

/* ------------------------------------------------------- Light.h:12-26 and subclasses --- */
static inline float pow128(float x) { /* Math::pow2<128>: seven squarings, Math.h:92-96 */
	for (int i = 0; i < 7; i++) x = x * x;
	return x;
}

static v3 light_base(const b200rt_light * l, v3 normal, v3 to_light, v3 to_camera) {
	float intensity = dot(normal, to_light);
	if (!(intensity > 0.0f)) return V(0.0f, 0.0f, 0.0f);

	v3 half_angle = normalize(add(to_light, to_camera));
	float specular_factor = dot(normal, half_angle);
	intensity = intensity + pow128(specular_factor);

	return scale(V3(l->colour), intensity);
}

/* PointLight.h:9-11 */
static v3 light_point(const b200rt_light * l, v3 normal, v3 to_light, v3 to_camera, float distance_squared) {
	return divf(light_base(l, normal, to_light, to_camera), distance_squared);
}

/* SpotLight.h:17-33 */
static v3 light_spot(const b200rt_light * l, v3 normal, v3 to_light, v3 to_camera, float distance_squared) {
	float d = dot(to_light, V3(l->negative_direction));
	if (!(d > l->outer_cutoff)) return V(0.0f, 0.0f, 0.0f);

	float radial_falloff = (d - l->outer_cutoff) / (l->inner_cutoff - l->outer_cutoff);
	if (radial_falloff > 1.0f) radial_falloff = 1.0f;

	return scale(light_point(l, normal, to_light, to_camera, distance_squared), radial_falloff);
}

/* ----------------------------------------------------- Raytracer::bounce, :87-400 ------ */
static v3 bounce(const ctx_t * ctx, const ray_t * ray, int bounces_left, float * distance, counters_t * cnt) {
	v3 result = V(0.0f, 0.0f, 0.0f);

	hit_t hit;
	memset(&hit, 0, sizeof(hit));
	hit.distance = INFINITY;
	hit.instance_id = -1; hit.triangle_id = -1;

	trace_primitives(ctx, ray, &hit, cnt);

	if (!hit.hit) { /* :105-111 */
		*distance = INFINITY;
		return sky_sample(ctx, ray->direction, cnt);
	}
	*distance = hit.distance;

	const b200rt_material * material = &ctx->s->materials[hit.material_id];
	v3 material_diffuse = get_albedo(ctx, material, &hit, cnt);

	if (dot(material_diffuse, material_diffuse) > 0.0f) { /* :143-201 */
		v3 diffuse = V3(ctx->f->ambient);
		v3 to_camera = normalize(sub(V3(ctx->f->camera.position), hit.point));

		for (int i = 0; i < ctx->f->light_count; i++) {
			const b200rt_light * l = &ctx->f->lights[i];
			cnt->shadow++;
			if (l->kind == 2) { /* directional, :189-198 */
				v3 to_light = V3(l->negative_direction);
				if (intersect_primitives(ctx, hit.point, to_light, INFINITY, cnt)) continue;
				diffuse = add(diffuse, light_base(l, hit.normal, to_light, to_camera));
			} else {            /* point :155-169, spot :172-186 */
				v3    to_light = sub(V3(l->position), hit.point);
				float distance_to_light_squared = dot(to_light, to_light);
				float distance_to_light         = sqrtf(distance_to_light_squared);
				to_light = divf(to_light, distance_to_light);

				if (intersect_primitives(ctx, hit.point, to_light, distance_to_light, cnt)) continue;
				diffuse = add(diffuse, l->kind == 0
					? light_point(l, hit.normal, to_light, to_camera, distance_to_light_squared)
					: light_spot (l, hit.normal, to_light, to_camera, distance_to_light_squared));
			}
		}
		result = V(fmaf(diffuse.x, material_diffuse.x, result.x), fmaf(diffuse.y, material_diffuse.y, result.y), fmaf(diffuse.z, material_diffuse.z, result.z));
	}

	if (bounces_left > 0) { /* :204-397 */
		v3 colour_reflection = V(0.0f, 0.0f, 0.0f);

		v3 material_reflection    = V3(material->reflection);
		v3 material_transmittance = V3(material->transmittance);

		int reflection_mask = dot(material_reflection,    material_reflection)    > 0.0f;
		int refraction_mask = dot(material_transmittance, material_transmittance) > 0.0f;

		if (reflection_mask) { /* :249-271 */
			ray_t reflected;
			reflected.origin = hit.point;
			/* Math::reflect, Math.h:28-31 */
			reflected.direction = sub(ray->direction, scale(hit.normal, 2.0f * dot(ray->direction, hit.normal)));

			reflected.dO_dx = hit.dO_dx;
			reflected.dO_dy = hit.dO_dy;

			float dDN_dx = dot(ray->dD_dx, hit.normal) + dot(ray->direction, hit.dN_dx);
			float dDN_dy = dot(ray->dD_dy, hit.normal) + dot(ray->direction, hit.dN_dy);

			float d_dot_n = dot(ray->direction, hit.normal);
			reflected.dD_dx = sub(ray->dD_dx, scale(add(scale(hit.dN_dx, d_dot_n), scale(hit.normal, dDN_dx)), 2.0f));
			reflected.dD_dy = sub(ray->dD_dy, scale(add(scale(hit.dN_dy, d_dot_n), scale(hit.normal, dDN_dy)), 2.0f));

			cnt->reflection++;

			float reflection_distance;
			colour_reflection = mul(material_reflection, bounce(ctx, &reflected, bounces_left - 1, &reflection_distance, cnt));
			result = add(result, colour_reflection);
		}

		if (refraction_mask) { /* :273-396 */
			float d_dot_n  = dot(ray->direction, hit.normal);
			int   entering = d_dot_n < 0.0f;

			float ior = material->index_of_refraction;
			float n_1 = entering ? 1.0f : ior;
			float n_2 = entering ? ior : 1.0f;

			float cos_theta = entering ? 0.0f - d_dot_n : d_dot_n;
			v3    normal    = entering ? hit.normal : neg(hit.normal);

			float eta = n_1 / n_2;
			float k   = 1.0f - (eta * eta * (1.0f - (cos_theta * cos_theta)));

			if (k < 0.0f) { /* total internal reflection: the reflection is added a second time, :310-314 */
				return reflection_mask ? add(result, colour_reflection) : result;
			}

			ray_t refracted;
			refracted.origin = hit.point;
			/* Math::refract, Math.h:34-36 */
			refracted.direction = add(scale(ray->direction, eta), scale(normal, (eta * cos_theta) - sqrtf(k)));

			cnt->refraction++;

			refracted.dO_dx = hit.dO_dx;
			refracted.dO_dy = hit.dO_dy;

			float dDN_dx = dot(ray->dD_dx, hit.normal) + dot(ray->direction, hit.dN_dx);
			float dDN_dy = dot(ray->dD_dy, hit.normal) + dot(ray->direction, hit.dN_dy);

			float D_dot_N      = 0.0f - cos_theta;
			float Dprime_dot_N = 0.0f - sqrtf(k);

			float mu = 0.0f - (eta * cos_theta + Dprime_dot_N);

			/* literal transcription of :341-342, float + vector broadcast included */
			refracted.dD_dx = sub(scale(ray->dD_dx, eta), scale(addf(mul(hit.dN_dx, hit.normal), mu * D_dot_N), dDN_dx));
			refracted.dD_dy = sub(scale(ray->dD_dy, eta), scale(addf(mul(hit.dN_dy, hit.normal), mu * D_dot_N), dDN_dy));

			float refraction_distance;
			v3 colour_refraction = bounce(ctx, &refracted, bounces_left - 1, &refraction_distance, cnt);

			/* Beer's law with the CHILD's hit distance, only when entering, :350-376 */
			if (entering) {
				colour_refraction.x = colour_refraction.x * expf((material_transmittance.x - 1.0f) * refraction_distance);
				colour_refraction.y = colour_refraction.y * expf((material_transmittance.y - 1.0f) * refraction_distance);
				colour_refraction.z = colour_refraction.z * expf((material_transmittance.z - 1.0f) * refraction_distance);
			}

			/* Schlick, :379-391 */
			float r_0 = (n_1 - n_2) / (n_1 + n_2);
			r_0 = r_0 * r_0;

			if (n_1 > n_2) cos_theta = 0.0f - dot(refracted.direction, normal);

			float one_minus_cos         = 1.0f - cos_theta;
			float one_minus_cos_squared = one_minus_cos * one_minus_cos;

			float F_r = r_0 + ((1.0f - r_0) * one_minus_cos_squared) * (one_minus_cos_squared * one_minus_cos);
			float F_t = 1.0f - F_r;

			v3 blend = add(scale(colour_reflection, F_r), scale(colour_refraction, F_t));
			return add(result, blend);
		}
	}
	return result;
}

/* -------------------------------------------------- Raytracer::render_tile, :3-85 ------ */
static ray_t primary_ray(const ctx_t * ctx, int i, int j) {
	const b200rt_camera * cam = &ctx->f->camera;
	v3 x_axis = V3(cam->x_axis), y_axis = V3(cam->y_axis), top_left = V3(cam->top_left_corner);

	ray_t ray;
	ray.origin = V3(cam->position);
	ray.dO_dx = V(0.0f, 0.0f, 0.0f);
	ray.dO_dy = V(0.0f, 0.0f, 0.0f);

	v3 direction = madd(x_axis, (float)i, madd(y_axis, (float)j, top_left)); /* :45-47 */

	float d_dot_d          = dot(direction, direction);
	float inv_sqrt_d_dot_d = inv_sqrt(d_dot_d);
	float denom            = inv_sqrt_d_dot_d / d_dot_d; /* d_dot_d ^ -3/2 */

	ray.dD_dx = scale(sub(scale(x_axis, d_dot_d), scale(direction, dot(direction, x_axis))), denom); /* :55-56 */
	ray.dD_dy = scale(sub(scale(y_axis, d_dot_d), scale(direction, dot(direction, y_axis))), denom);

	ray.direction = scale(direction, inv_sqrt_d_dot_d);
	return ray;
}

/* Window::plot, Window.h:56-65. cvtss2si(NaN) = 0x80000000, reproduced by float_to_int. */
static inline uint32_t plot_colour(v3 colour) {
	float c[3] = { colour.x, colour.y, colour.z };
	int q[3];
	for (int k = 0; k < 3; k++) {
		float v = c[k] * 255.0f;
		if (v < 0.0f) v = 0.0f;       /* Math::clamp passes NaN through, Math.h:13-18 */
		if (v > 255.0f) v = 255.0f;
		q[k] = float_to_int(v - 0.5f);
	}
	return ((uint32_t)q[0] << 16) | ((uint32_t)q[1] << 8) | (uint32_t)q[2];
}

typedef struct {
	const ctx_t * ctx;
	const b200rt_outputs * out;
	int * next_tile; pthread_mutex_t * lock;
	counters_t cnt;
} worker_t;

static void render_pixel(const ctx_t * ctx, const b200rt_outputs * out, int i, int j, counters_t * cnt) {
	const int width = ctx->c->width;
	const size_t p = (size_t)i + (size_t)width * j;

	ray_t ray = primary_ray(ctx, i, j);
	cnt->primary++;

	float distance;
	v3 colour = bounce(ctx, &ray, ctx->c->bounces, &distance, cnt);

	if (out->frame_buffer) out->frame_buffer[p] = plot_colour(colour);
	if (out->colour_float) { out->colour_float[3*p] = colour.x; out->colour_float[3*p+1] = colour.y; out->colour_float[3*p+2] = colour.z; }

	if (out->primary_ids || out->primary_hit) { /* what the instrumented reference build dumps */
		counters_t scratch; memset(&scratch, 0, sizeof(scratch));
		hit_t hit; memset(&hit, 0, sizeof(hit));
		hit.distance = INFINITY; hit.instance_id = -1; hit.triangle_id = -1;
		trace_primitives(ctx, &ray, &hit, &scratch);
		if (out->primary_ids) { out->primary_ids[2*p] = hit.hit ? hit.instance_id : -1; out->primary_ids[2*p+1] = hit.hit ? hit.triangle_id : -1; }
		if (out->primary_hit) {
			out->primary_hit[4*p]   = hit.distance;
			out->primary_hit[4*p+1] = hit.hit ? (float)hit.material_id : -1.0f;
			out->primary_hit[4*p+2] = hit.hit ? hit.u : 0.0f;
			out->primary_hit[4*p+3] = hit.hit ? hit.v : 0.0f;
		}
	}
}

/* Tile hand-out of WorkerThread.cpp:53-65: 32x32 tiles, countdown, highest id first */
static void * worker(void * arg) {
	worker_t * w = (worker_t *)arg;
	const b200rt_config * c = w->ctx->c;
	const int tiles_x = (c->width + 31) / 32;
	const int stride = c->tile_stride > 0 ? c->tile_stride : 1;

	for (;;) {
		pthread_mutex_lock(w->lock);
		int task = --(*w->next_tile);
		pthread_mutex_unlock(w->lock);
		if (task < 0) break;
		if (task % stride != c->tile_phase) continue; /* multi-GPU interleave, SURVEY.md 8(e) */

		int x0 = (task % tiles_x) * 32, y0 = (task / tiles_x) * 32;
		int x1 = x0 + 32 < c->width  ? x0 + 32 : c->width;
		int y1 = y0 + 32 < c->height ? y0 + 32 : c->height;
		for (int j = y0; j < y1; j++) for (int i = x0; i < x1; i++) render_pixel(w->ctx, w->out, i, j, &w->cnt);
	}
	return NULL;
}

int oracle_render(const b200rt_static_scene * scene, const b200rt_frame * frame, const b200rt_config * config, const b200rt_outputs * outputs, int threads) {
	if (!scene || !frame || !config || !outputs) return -1;
	if (config->width <= 0 || config->height <= 0 || config->bounces < 0) return -1;
	if (threads < 1) threads = 1;
	if (threads > 256) threads = 256;

	ctx_t ctx;
	ctx.s = scene; ctx.f = frame; ctx.c = config;
	{ /* Texture::init, Texture.h:53-62 */
		const float alpha = 2.0f, denom = 1.0f / 127.0f, exp_neg_alpha = expf(-alpha);
		for (int i = 0; i < 128; i++) ctx.ewa_weight_table[i] = expf(-alpha * ((float)i * denom)) - exp_neg_alpha;
	}

	int next_tile = ((config->width + 31) / 32) * ((config->height + 31) / 32);
	pthread_mutex_t lock = PTHREAD_MUTEX_INITIALIZER;

	worker_t * workers = (worker_t *)calloc((size_t)threads, sizeof(worker_t));
	pthread_t * ids = (pthread_t *)calloc((size_t)threads, sizeof(pthread_t));
	for (int t = 0; t < threads; t++) { workers[t].ctx = &ctx; workers[t].out = outputs; workers[t].next_tile = &next_tile; workers[t].lock = &lock; }
	for (int t = 1; t < threads; t++) pthread_create(&ids[t], NULL, worker, &workers[t]);
	worker(&workers[0]);
	for (int t = 1; t < threads; t++) pthread_join(ids[t], NULL);

	if (outputs->stats) {
		b200rt_stats * s = outputs->stats;
		memset(s, 0, sizeof(*s));
		for (int t = 0; t < threads; t++) {
			s->primary_rays += workers[t].cnt.primary; s->shadow_rays += workers[t].cnt.shadow;
			s->reflection_rays += workers[t].cnt.reflection; s->refraction_rays += workers[t].cnt.refraction;
			s->nodes_visited += workers[t].cnt.nodes; s->triangles_tested += workers[t].cnt.triangles; s->texels_fetched += workers[t].cnt.texels;
			s->hits_accepted += workers[t].cnt.accepted; s->instances_entered += workers[t].cnt.instances;
		}
	}
	free(workers); free(ids);

	if (outputs->presented_rgb8 && outputs->frame_buffer) oracle_present(outputs->frame_buffer, config->width, config->height, config->enable_fxaa, outputs->presented_rgb8);
	return 0;
}

/* ------------------------------------ Data/Shaders/fragment_fxaa.glsl:11-74 ------------ */
/* The GLSL cannot be executed (no graphics stack on either box): this is its arithmetic, with the
 * sampler state of Window.cpp:48-50 (NEAREST, default GL_REPEAT wrap, RGBA8 UNORM texture) and the
 * uv mapping of vertex.glsl:7-11 (pixel (px, py) <-> uv ((px + 0.5)/W, (py + 0.5)/H), row 0 = top).
 * get_screen() = pow(texel.rgb, 1/2.2) on 8-bit values, i.e. a 256-entry table.
 * "parity unpinned" for this pass only: no reference execution of the shader is possible. */
typedef struct { const uint32_t * fb; int w, h; const float * gamma; } screen_t;

/* texture2D(screen, uv) on the NEAREST / GL_REPEAT sampler of Window.cpp:48-50: the texel is
 * (floor(u * W) mod W, floor(v * H) mod H) (OpenGL 4.2 spec 3.9.11: unnormalised coordinate, floor, wrap). */
static inline v3 get_screen(const screen_t * s, float u, float v) {
	int x = positive_mod((int)floorf(u * (float)s->w), s->w);
	int y = positive_mod((int)floorf(v * (float)s->h), s->h);
	uint32_t c = s->fb[(size_t)x + (size_t)s->w * y];
	return V(s->gamma[(c >> 16) & 255], s->gamma[(c >> 8) & 255], s->gamma[c & 255]);
}
static inline float luma_of(v3 c) { return (0.299f * c.x + 0.587f * c.y) + 0.114f * c.z; }
static inline uint8_t to_unorm8(float c) {
	if (!(c > 0.0f)) c = 0.0f;
	if (c > 1.0f) c = 1.0f;
	return (uint8_t)lrintf(c * 255.0f);
}

void oracle_gamma_table(float table[256]) {
	for (int i = 0; i < 256; i++) table[i] = powf((float)i / 255.0f, 1.0f / 2.2f);
}

void oracle_present(const uint32_t * frame_buffer, int width, int height, int enable_fxaa, uint8_t * rgb8) {
	float gamma[256];
	oracle_gamma_table(gamma);
	screen_t s = { frame_buffer, width, height, gamma };

	/* The shader's own coordinates: in_uv = ((px + .5) / W, (py + .5) / H) (vertex.glsl:7-11, interpolated to the pixel
	 * centre), inv_screen_size = (1 / W, 1 / H) (Window.cpp:57). Every tap is formed in uv space, as the GLSL forms it, and
	 * only get_screen() goes back to texels: a tap within rounding distance of a texel border picks the texel the shader's
	 * arithmetic picks (round 1 stepped in pixel space: 60 of 2 073 600 pixels came out differently at 1080p). */
	const float inv_x = 1.0f / (float)width, inv_y = 1.0f / (float)height;

	for (int py = 0; py < height; py++) {
		for (int px = 0; px < width; px++) {
			const float in_u = ((float)px + 0.5f) / (float)width, in_v = ((float)py + 0.5f) / (float)height;
			v3 out;
			if (!enable_fxaa) { /* fragment_identity.glsl */
				out = get_screen(&s, in_u, in_v);
			} else {
				/* vec2(+-inv.x, +-inv_screen_size) takes .x for BOTH components (fragment_fxaa.glsl:25-28) */
				float luma_tl = luma_of(get_screen(&s, in_u - inv_x, in_v - inv_x));
				float luma_tr = luma_of(get_screen(&s, in_u + inv_x, in_v - inv_x));
				float luma_bl = luma_of(get_screen(&s, in_u - inv_x, in_v + inv_x));
				float luma_br = luma_of(get_screen(&s, in_u + inv_x, in_v + inv_x));
				float luma_m  = luma_of(get_screen(&s, in_u, in_v));

				float luma_min = fminf(fminf(fminf(luma_tl, luma_tr), fminf(luma_bl, luma_br)), luma_m);
				float luma_max = fmaxf(fmaxf(fmaxf(luma_tl, luma_tr), fmaxf(luma_bl, luma_br)), luma_m);

				float dir_x = (luma_bl + luma_br) - (luma_tl + luma_tr);
				float dir_y = (luma_tl + luma_bl) - (luma_tr + luma_br);

				float reduce = fmaxf(1.0f / 128.0f, (luma_tl + luma_tr + luma_bl + luma_br) * 0.25f * (1.0f / 8.0f));
				float adjust = 1.0f / (fminf(fabsf(dir_x), fabsf(dir_y)) + reduce);

				/* clamp(dir * adjust, -8, 8) * inv_screen_size (:46) */
				dir_x = fminf(fmaxf(dir_x * adjust, -8.0f), 8.0f) * inv_x;
				dir_y = fminf(fmaxf(dir_y * adjust, -8.0f), 8.0f) * inv_y;

				const float k1 = 1.0f / 3.0f - 0.5f, k2 = 2.0f / 3.0f - 0.5f;
				v3 a0 = get_screen(&s, in_u + dir_x * k1, in_v + dir_y * k1);
				v3 a1 = get_screen(&s, in_u + dir_x * k2, in_v + dir_y * k2);
				v3 b0 = get_screen(&s, in_u + dir_x * -0.5f, in_v + dir_y * -0.5f);
				v3 b1 = get_screen(&s, in_u + dir_x * 0.5f,  in_v + dir_y * 0.5f);

				v3 result_a = scale(add(a0, a1), 0.5f);
				v3 result_b = scale(add(b0, b1), 0.5f);
				v3 result   = scale(add(result_a, result_b), 0.5f);

				float luma_result = luma_of(result);
				out = (luma_result < luma_min || luma_result > luma_max) ? result_a : result;
			}
			uint8_t * o = rgb8 + 3 * ((size_t)px + (size_t)width * py);
			o[0] = to_unorm8(out.x); o[1] = to_unorm8(out.y); o[2] = to_unorm8(out.z);
		}
	}
}
