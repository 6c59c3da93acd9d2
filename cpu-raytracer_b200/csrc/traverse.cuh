// traverse.cuh — two-level BVH traversal and primitive intersection, one ray per thread.
//
// Reference: Scene::trace_primitives / intersect_primitives (Scene.cpp:173-190) ->
// PrimitiveList (PrimitiveList.h:29-45) -> Sphere (Sphere.cpp:9-112), Plane (Plane.cpp:13-79),
// TopLevelBVH::trace / intersect (TopLevelBVH.cpp:53-129) -> Mesh (Mesh.cpp:17-40) ->
// BottomLevelBVH::trace / intersect (BottomLevelBVH.cpp:355-437) -> triangle_trace /
// triangle_intersect (:214-346), AABB::intersect (AABB.cpp:38-52), BVHNode.h:30-40.
//
// What is kept from the reference because it decides the RESULT: the slab test's operation
// order and x86 NaN behaviour, strict inequalities of the triangle test, t shared across the
// rigid instance transform, leaf/near-far visiting order (ties between equidistant hits).
// What is restructured for the GPU: a closest-hit query only records (t, u, v, primitive,
// instance); the full hit record is computed once, after traversal, for the winner
// (shade.cuh) — the reference recomputes it on every accepted candidate
// (BottomLevelBVH.cpp:254-308), a pure function of the same inputs.
#pragma once
#include "rt_math.cuh"

namespace RT_NAMESPACE {

// BVH_TRAVERSAL_STACK_SIZE (Config.h:25) is 64 — a capacity, not semantics: a traversal that pushes both children
// and pops one holds at most depth + 1 entries. The host measures every tree's depth (threaded_nodes.h: check_tree)
// and the launch picks the instantiation whose stacks are deep enough: BSTACK = 32 BLAS entries with
// RT_TLAS_STACK_SMALL TLAS entries, or 64 with RT_TLAS_STACK_SIZE (rt_types.h; the TLAS is measured per frame). Smaller arrays = a smaller local-memory frame per
// thread (151 552 resident threads: every 4 bytes are 0.6 MB of L2 footprint).

struct Ray {               // Ray.h:7-16
	v3 origin, direction;
	v3 dO_dx, dO_dy, dD_dx, dD_dy;
};

enum { HIT_NONE = -1, HIT_SPHERE = -2, HIT_PLANE = -3 }; // instance >= 0: mesh instance

struct Candidate {         // closest hit so far
	float t, u, v;
	int instance;          // HIT_* or TopLevelBVH primitive index
	int primitive;         // flattened triangle index within its BLAS, or sphere / plane index
};

struct Counters { unsigned nodes, triangles, texels, shadow, reflection, refraction, accepted, instances; };

// AABB::intersect, AABB.cpp:38-52. MINPS/MAXPS NaN semantics matter: a ray with a zero
// direction component lying exactly on a slab plane produces 0 * inf = NaN.
//
// The slab distances are computed with the reference's operations ((plane - o) * inv_d). How the
// six of them are reduced depends on the ray:
//   regular ray  (origin and 1/direction all finite): no NaN can appear (finite - finite is finite,
//     finite * finite is finite or +-inf), so MINPS/MAXPS equal IEEE min/max and the reduction is
//     6 FMNMX + 2 FMNMX3 instead of 24 FSETP+FSEL — the v1 profile had 31 % of all issued
//     instructions there (profiles/r01_ncu_full_render_v1.csv);
//   irregular ray (a zero direction component, or NaN upstream): the exact x86 emulation.
// `exact` is a per-ray flag, computed once per (ray, space).
__device__ __forceinline__ bool ray_is_regular(v3 o, v3 inv_d) {
	return isfinite(o.x) && isfinite(o.y) && isfinite(o.z) && isfinite(inv_d.x) && isfinite(inv_d.y) && isfinite(inv_d.z);
}
#ifdef RT_FAST_SLAB
// FAST arithmetic (b200rt_config.arithmetic): one fused multiply-add per slab plane, t = plane * (1/d) - o * (1/d),
// against the reference's subtract-then-multiply: 6 instructions instead of 12 in the hottest loop of the kernel.
// `o` here is slab_origin() = -(o * inv_d); NaNs (0 * inf) fall out of fminf / fmaxf, i.e. that slab does not constrain.
// 1/d is kept finite (a zero component gives +-1e30, not inf): inf * plane - inf * o would be NaN or the wrong
// infinity exactly where the reference's (plane - o) * inf is right — a ray parallel to a slab and inside it.
__device__ __forceinline__ v3 inv_direction(v3 d) {
	return V(fminf(fmaxf(1.0f / d.x, -1e30f), 1e30f), fminf(fmaxf(1.0f / d.y, -1e30f), 1e30f), fminf(fmaxf(1.0f / d.z, -1e30f), 1e30f));
}
__device__ __forceinline__ v3 slab_origin(v3 o, v3 inv_d) { return V(0.0f - o.x * inv_d.x, 0.0f - o.y * inv_d.y, 0.0f - o.z * inv_d.z); }
__device__ __forceinline__ bool aabb_intersect(float4 n0, float4 n1, v3 o, v3 inv_d, float max_distance, bool) {
	const float t0x = __fmaf_rn(n0.x, inv_d.x, o.x), t1x = __fmaf_rn(n0.w, inv_d.x, o.x);
	const float t0y = __fmaf_rn(n0.y, inv_d.y, o.y), t1y = __fmaf_rn(n1.x, inv_d.y, o.y);
	const float t0z = __fmaf_rn(n0.z, inv_d.z, o.z), t1z = __fmaf_rn(n1.y, inv_d.z, o.z);
	const float t_near = fmaxf(fmaxf(RT_EPSILON,   fminf(t0x, t1x)), fmaxf(fminf(t0y, t1y), fminf(t0z, t1z)));
	const float t_far  = fminf(fminf(max_distance, fmaxf(t0x, t1x)), fminf(fmaxf(t0y, t1y), fmaxf(t0z, t1z)));
	return t_near < t_far;
}
#else
__device__ __forceinline__ v3 inv_direction(v3 d) { return V(1.0f / d.x, 1.0f / d.y, 1.0f / d.z); }
__device__ __forceinline__ v3 slab_origin(v3 o, v3) { return o; }
__device__ __forceinline__ bool aabb_intersect(float4 n0, float4 n1, v3 o, v3 inv_d, float max_distance, bool exact) {
	float t0x = (n0.x - o.x) * inv_d.x, t1x = (n0.w - o.x) * inv_d.x;
	float t0y = (n0.y - o.y) * inv_d.y, t1y = (n1.x - o.y) * inv_d.y;
	float t0z = (n0.z - o.z) * inv_d.z, t1z = (n1.y - o.z) * inv_d.z;

	if (!exact) {
		const float t_near = fmaxf(fmaxf(RT_EPSILON,   fminf(t0x, t1x)), fmaxf(fminf(t0y, t1y), fminf(t0z, t1z)));
		const float t_far  = fminf(fminf(max_distance, fmaxf(t0x, t1x)), fminf(fmaxf(t0y, t1y), fmaxf(t0z, t1z)));
		return t_near < t_far;
	}
	float tminx = sse_min(t0x, t1x), tmaxx = sse_max(t0x, t1x);
	float tminy = sse_min(t0y, t1y), tmaxy = sse_max(t0y, t1y);
	float tminz = sse_min(t0z, t1z), tmaxz = sse_max(t0z, t1z);

	float t_near = sse_max(sse_max(RT_EPSILON,   tminx), sse_max(tminy, tminz));
	float t_far  = sse_min(sse_min(max_distance, tmaxx), sse_min(tmaxy, tmaxz));
	return t_near < t_far;
}
#endif

// BVHNode::should_visit_left_first, BVHNode.h:30-40 (axis in the top two bits of count)
__device__ __forceinline__ bool visit_left_first(int count, v3 d) {
	unsigned axis = unsigned(count) >> 30;
	float c = axis == 1 ? d.x : (axis == 2 ? d.y : d.z);
	return c > 0.0f;
}
// The same decision for every node of one traversal: bit `axis` of the mask is direction[axis] > 0
// (axis 1 = x, 2 = y, 3 = z; the unused code 0 falls through to z like the chain above), so a node
// costs two shifts and an AND instead of two compares, two selects and a float compare.
__device__ __forceinline__ unsigned direction_order_mask(v3 d) {
	return (d.z > 0.0f ? 9u : 0u) | (d.x > 0.0f ? 2u : 0u) | (d.y > 0.0f ? 4u : 0u);
}
// BVH_TRAVERSE_TREE_NAIVE (BVHNode.h:31-32): the left child first, whatever the ray — every bit of the mask set.
__device__ __forceinline__ unsigned traversal_order_mask(const SceneParams & P, v3 d) {
	return P.traversal_naive ? 0xffffffffu : direction_order_mask(d);
}
__device__ __forceinline__ int left_first_bit(int count, unsigned order_mask) {
	return int((order_mask >> (unsigned(count) >> 30)) & 1u);
}

// Moller-Trumbore with the reference's operation order; all comparisons strict
// (BottomLevelBVH.cpp:214-252 / :311-346).
#ifdef RT_FAST_TRI
// FAST arithmetic: the triangle as a precomputed world-to-barycentric transform (Woop / Baldwin-Weber; b200rt.cu builds
// it at upload, same 48 bytes as TriangleHot): rows of [e1 e2 n]^-1 with the offset -row . p0 in .w. The ray's
// transformed z gives t, then u and v are one multiply-add each — about 25 operations and one MUFU.RCP against
// Moller-Trumbore's 31 and an IEEE division, and the test that fails most often (t) comes first.
__device__ __forceinline__ float fast_rcp(float x) {
#ifdef RT_EMULATE
	return 1.0f / x;
#else
	float r; asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x)); return r;
#endif
}
__device__ __forceinline__ bool triangle_test(const float4 * __restrict__ tri, v3 o, v3 d, float max_t, float & t_out, float & u_out, float & v_out) {
	const float4 rw = load_node(tri), ru = load_node(tri + 1), rv = load_node(tri + 2);
	const float dz = __fmaf_rn(rw.x, d.x, __fmaf_rn(rw.y, d.y, rw.z * d.z));
	const float oz = __fmaf_rn(rw.x, o.x, __fmaf_rn(rw.y, o.y, __fmaf_rn(rw.z, o.z, rw.w)));
	const float t = (0.0f - oz) * fast_rcp(dz);
	if (!(t > RT_EPSILON && t < max_t)) return false;
	const float u = __fmaf_rn(t, __fmaf_rn(ru.x, d.x, __fmaf_rn(ru.y, d.y, ru.z * d.z)), __fmaf_rn(ru.x, o.x, __fmaf_rn(ru.y, o.y, __fmaf_rn(ru.z, o.z, ru.w))));
	if (!(u > 0.0f && u < 1.0f)) return false;
	const float v = __fmaf_rn(t, __fmaf_rn(rv.x, d.x, __fmaf_rn(rv.y, d.y, rv.z * d.z)), __fmaf_rn(rv.x, o.x, __fmaf_rn(rv.y, o.y, __fmaf_rn(rv.z, o.z, rv.w))));
	if (!(v > 0.0f && (u + v) < 1.0f)) return false;
	t_out = t; u_out = u; v_out = v;
	return true;
}
#else
__device__ __forceinline__ bool triangle_test(const float4 * __restrict__ tri, v3 o, v3 d, float max_t, float & t_out, float & u_out, float & v_out) {
	const float4 a0 = load_node(tri), a1 = load_node(tri + 1), a2 = load_node(tri + 2);
	const v3 p0 = V(a0.x, a0.y, a0.z), edge_1 = V(a0.w, a1.x, a1.y), edge_2 = V(a1.z, a1.w, a2.x);

	v3    h = cross(d, edge_2);
	float a = dot(edge_1, h);
	float f = 1.0f / a;
	v3    s = o - p0;
	float u = f * dot(s, h);
	if (!(u > 0.0f && u < 1.0f)) return false;

	v3    q = cross(s, edge_1);
	float v = f * dot(d, q);
	if (!(v > 0.0f && (u + v) < 1.0f)) return false;

	float t = f * dot(edge_2, q);
	if (!(t > RT_EPSILON && t < max_t)) return false;

	t_out = t; u_out = u; v_out = v;
	return true;
}
#endif

__device__ __forceinline__ DeviceInstance load_instance_inverse(const DeviceInstance * __restrict__ inst) {
	const float4 * p = reinterpret_cast<const float4 *>(inst);
	DeviceInstance r;
	r.inverse.r0 = __ldg(p); r.inverse.r1 = __ldg(p + 1); r.inverse.r2 = __ldg(p + 2);
	const int4 o = __ldg(reinterpret_cast<const int4 *>(p + 6));
	r.node_offset = o.x; r.triangle_offset = o.y; r.material_offset = o.z; r.shadow_node_offset = o.w;
	return r;
}

// BottomLevelBVH::trace, BottomLevelBVH.cpp:355-396 (object-space ray, shared t)
template<bool COUNT>
__device__ __forceinline__ void blas_trace(const float4 * __restrict__ nodes, const float4 * __restrict__ hot, v3 o, v3 d, int instance, Candidate & best, int * stack, Counters & cnt, unsigned order_mask) {
	const v3 inv_d = inv_direction(d);
	const bool exact = !ray_is_regular(o, inv_d);
	const v3 so = slab_origin(o, inv_d);

	int stack_size = 1;
	stack[0] = 0;

	while (stack_size > 0) {
		const int index = stack[--stack_size];
		const float4 n0 = load_node(nodes + 2 * index), n1 = load_node(nodes + 2 * index + 1);
		if (COUNT) cnt.nodes++;

		if (!aabb_intersect(n0, n1, so, inv_d, best.t, exact)) continue;

		const int left_or_first = __float_as_int(n1.z), count = __float_as_int(n1.w);
		if ((count & 0x3FFFFFFF) > 0) {
			for (int i = left_or_first; i < left_or_first + count; i++) {
				float t, u, v;
				if (COUNT) cnt.triangles++;
				if (triangle_test(hot + 3 * size_t(i), o, d, best.t, t, u, v)) {
					if (COUNT) cnt.accepted++;
					best.t = t; best.u = u; best.v = v; best.instance = instance; best.primitive = i;
				}
			}
		} else { // push far, then near (BottomLevelBVH.cpp:380-387), without a divergent branch
			const int left_first = left_first_bit(count, order_mask);
			stack[stack_size]     = left_or_first + left_first;
			stack[stack_size + 1] = left_or_first + 1 - left_first;
			stack_size += 2;
		}
	}
}

// BottomLevelBVH::intersect, BottomLevelBVH.cpp:398-437.
// The answer of an any-hit query does not depend on the order the children are visited in: the
// boxes are tested against the fixed max_distance, so the set of leaves the ray can reach is the
// same for every order, and the result is "some triangle in a reachable leaf is hit". The
// reference's front-to-back order (BottomLevelBVH.cpp:421-428) is the slowest of those measured:
// a shadow ray starts ON a surface, and front-to-back first wades through the geometry around its
// own origin. Bench frame, ms per frame (profiles/r01_variants_anyhit_order.log; identical images):
// front-to-back 1.374, leader lane's front-to-back 1.355, left first 1.295, larger-area child first
// 1.255, more-triangles child first 1.199, back-to-front 1.168, RIGHT FIRST 1.153 — also the best
// on the other three scenes (Sibenik 4K 3.15 -> 2.91, SCENE_DYNAMIC 1.07 -> 0.96, dielectrics/EWA
// 7.13 -> 6.50). The counting instantiation keeps the reference's order, so that nodes_visited
// stays the reference algorithm's work.
template<bool COUNT>
__device__ __forceinline__ bool blas_intersect(const float4 * __restrict__ nodes, const float4 * __restrict__ hot, v3 o, v3 d, float max_distance, int * stack, Counters & cnt, unsigned order_mask) {
	const v3 inv_d = inv_direction(d);
	const bool exact = !ray_is_regular(o, inv_d);
	const v3 so = slab_origin(o, inv_d);

	int stack_size = 1;
	stack[0] = 0;

	while (stack_size > 0) {
		const int index = stack[--stack_size];
		const float4 n0 = load_node(nodes + 2 * index), n1 = load_node(nodes + 2 * index + 1);
		if (COUNT) cnt.nodes++;

		if (!aabb_intersect(n0, n1, so, inv_d, max_distance, exact)) continue;

		const int left_or_first = __float_as_int(n1.z), count = __float_as_int(n1.w);
		if ((count & 0x3FFFFFFF) > 0) {
			for (int i = left_or_first; i < left_or_first + count; i++) {
				float t, u, v;
				if (COUNT) cnt.triangles++;
				if (triangle_test(hot + 3 * size_t(i), o, d, max_distance, t, u, v)) return true;
			}
		} else { // the right child is visited first (see above); the counting build pushes far, then near like BottomLevelBVH.cpp:421-428
			const int left_first = COUNT ? left_first_bit(count, order_mask) : 0;
			stack[stack_size]     = left_or_first + left_first;
			stack[stack_size + 1] = left_or_first + 1 - left_first;
			stack_size += 2;
		}
	}
	return false;
}

// The same query on the threaded copy of the tree (threaded_nodes.h: build_threaded_nodes): records lie in
// the order they are visited (right child first), so a hit inner node continues with the next
// record and everything else follows the record's skip link — no stack, no push, no pop.
template<bool COUNT>
__device__ __forceinline__ bool blas_shadow(const float4 * __restrict__ records, const float4 * __restrict__ hot, v3 o, v3 d, float max_distance, Counters & cnt) {
	const v3 inv_d = inv_direction(d);
	const bool exact = !ray_is_regular(o, inv_d);
	const v3 so = slab_origin(o, inv_d);

	int index = 0;
	while (index >= 0) {
		const float4 n0 = load_node(records + 2 * index), n1 = load_node(records + 2 * index + 1);
		int next = __float_as_int(n1.z);                       // skip link
		if (next >= 0) prefetch_record(records + 2 * next);
		if (COUNT) cnt.nodes++;
		if (aabb_intersect(n0, n1, so, inv_d, max_distance, exact)) {
			const int leaf = __float_as_int(n1.w);
			if (leaf != 0) {
				const int first = leaf >> 6, count = leaf & 63;
				for (int i = first; i < first + count; i++) {
					float t, u, v;
					if (COUNT) cnt.triangles++;
					if (triangle_test(hot + 3 * size_t(i), o, d, max_distance, t, u, v)) return true;
				}
			} else {
				next = index + 1;
			}
		}
		index = next;
	}
	return false;
}

#ifdef RT_WIDE_SHADOW
// Experiment (off by default): the any-hit query on the tree collapsed to four children per record (threaded_nodes.h:
// build_wide_nodes) — half the dependent loads per ray. Same boolean as the binary walk for REGULAR rays (the slab
// computation is monotone in the box, and a skipped node's box contains the tested ones below it); irregular rays
// (a zero direction component: NaN arithmetic is not monotone) keep the binary walk.
#define RT_WIDE_STACK 48
__device__ __forceinline__ bool blas_shadow_wide(const float4 * __restrict__ records, const float4 * __restrict__ hot, v3 o, v3 inv_d, v3 d, float max_distance) {
	int stack[RT_WIDE_STACK];
	int size = 1;
	stack[0] = 0;
	while (size > 0) {
		const float4 * r = records + 8 * size_t(stack[--size]);
		const float4 lox = load_node(r), loy = load_node(r + 1), loz = load_node(r + 2), hix = load_node(r + 3), hiy = load_node(r + 4), hiz = load_node(r + 5);
		const float4 links = load_node(r + 6), counts = load_node(r + 7);
		const float lo_x[4] = { lox.x, lox.y, lox.z, lox.w }, lo_y[4] = { loy.x, loy.y, loy.z, loy.w }, lo_z[4] = { loz.x, loz.y, loz.z, loz.w };
		const float hi_x[4] = { hix.x, hix.y, hix.z, hix.w }, hi_y[4] = { hiy.x, hiy.y, hiy.z, hiy.w }, hi_z[4] = { hiz.x, hiz.y, hiz.z, hiz.w };
		const int link[4] = { __float_as_int(links.x), __float_as_int(links.y), __float_as_int(links.z), __float_as_int(links.w) };
		const int count[4] = { __float_as_int(counts.x), __float_as_int(counts.y), __float_as_int(counts.z), __float_as_int(counts.w) };
#pragma unroll
		for (int k = 0; k < 4; k++) {
			if (count[k] < 0) continue;
			// AABB::intersect with the reference's operations (aabb_intersect, regular rays)
			const float t0x = (lo_x[k] - o.x) * inv_d.x, t1x = (hi_x[k] - o.x) * inv_d.x;
			const float t0y = (lo_y[k] - o.y) * inv_d.y, t1y = (hi_y[k] - o.y) * inv_d.y;
			const float t0z = (lo_z[k] - o.z) * inv_d.z, t1z = (hi_z[k] - o.z) * inv_d.z;
			const float t_near = fmaxf(fmaxf(RT_EPSILON,   fminf(t0x, t1x)), fmaxf(fminf(t0y, t1y), fminf(t0z, t1z)));
			const float t_far  = fminf(fminf(max_distance, fmaxf(t0x, t1x)), fminf(fmaxf(t0y, t1y), fmaxf(t0z, t1z)));
			if (!(t_near < t_far)) continue;
			if (count[k] == 0) { stack[size++] = link[k]; continue; }
			for (int i = link[k]; i < link[k] + count[k]; i++) {
				float t, u, v;
				if (triangle_test(hot + 3 * size_t(i), o, d, max_distance, t, u, v)) return true;
			}
		}
	}
	return false;
}
#endif

#ifdef RT_OCTANT_THREADED
// Experiment (off by default): closest hit on the copy of the tree threaded in this ray's visiting order
// (threaded_nodes.h: build_threaded_nodes with an octant). The records come in the sequence the stack traversal pops
// the nodes in, each tested against the running best.t: same tests, same order, no stack.
__device__ __forceinline__ void blas_trace_threaded(const float4 * __restrict__ records, const float4 * __restrict__ hot, v3 o, v3 d, int instance, Candidate & best) {
	const v3 inv_d = inv_direction(d);
	const bool exact = !ray_is_regular(o, inv_d);
	const v3 so = slab_origin(o, inv_d);

	int index = 0;
	while (index >= 0) {
		const float4 n0 = load_node(records + 2 * index), n1 = load_node(records + 2 * index + 1);
		int next = __float_as_int(n1.z);                       // skip link
		if (next >= 0) prefetch_record(records + 2 * next);
		const int leaf = __float_as_int(n1.w);                 // bit 31: continuation of a long leaf, no box test; bit 30: the leaf goes on in the next record
		if (leaf < 0 || aabb_intersect(n0, n1, so, inv_d, best.t, exact)) {
			if (leaf != 0) {
				const int first = (leaf & 0x3FFFFFFF) >> 6, count = leaf & 63;
				if (leaf & 0x40000000) next = index + 1;
				for (int i = first; i < first + count; i++) {
					float t, u, v;
					if (triangle_test(hot + 3 * size_t(i), o, d, best.t, t, u, v)) {
						best.t = t; best.u = u; best.v = v; best.instance = instance; best.primitive = i;
					}
				}
			} else {
				next = index + 1;
			}
		}
		index = next;
	}
}
#endif

// Sphere::trace, Sphere.cpp:9-40 — t only; the record is built in shade.cuh
__device__ __forceinline__ bool sphere_test(const DeviceSphere * __restrict__ sphere, v3 o, v3 d, float max_t, float & t_out) {
	const float4 c = __ldg(&sphere->center_r2);
	v3 oc = o - V(c.x, c.y, c.z);
	float a = dot(d, d);
	float b = 2.0f * dot(oc, d);
	float cc = dot(oc, oc) - c.w;

	float disc = b * b - 4.0f * a * cc;
	if (!(disc >= 0.0f)) return false;

	float sqrt_d    = sqrtf(disc);
	float inv_denom = 0.0f - 1.0f / (2.0f * a);

	float t0 = (b + sqrt_d) * inv_denom;
	float t1 = (b - sqrt_d) * inv_denom;

	float t = t0 > RT_EPSILON ? t0 : t1;
	if (!(t > RT_EPSILON && t < max_t)) return false;
	t_out = t;
	return true;
}

// Sphere::intersect, Sphere.cpp:92-112 (geometric form: not the same arithmetic as trace)
__device__ __forceinline__ bool sphere_intersect(const DeviceSphere * __restrict__ sphere, v3 o, v3 d, float max_distance) {
	const float4 c4 = __ldg(&sphere->center_r2);
	v3    c  = V(c4.x, c4.y, c4.z) - o;
	float t  = dot(c, d);
	v3    Q  = c - d * t;
	float p2 = dot(Q, Q);
	if (!(p2 < c4.w)) return false;
	t = t - sqrtf(c4.w - p2);
	return t > RT_EPSILON && t < max_distance;
}

// Plane::trace / intersect, Plane.cpp:13-22,71-79
__device__ __forceinline__ float plane_t(const DevicePlane * __restrict__ plane, v3 o, v3 d) {
	const float4 n = __ldg(&plane->normal_d);
	v3 normal = V(n.x, n.y, n.z);
	return (0.0f - (dot(normal, o) + n.w)) / dot(normal, d);
}

// Scene::trace_primitives, Scene.cpp:173-177: spheres, planes, then the TLAS, all against one
// running closest distance.
template<bool COUNT, int BSTACK>
__device__ __forceinline__ void trace_closest(const SceneParams & P, v3 o, v3 d, Candidate & best, Counters & cnt) {
	best.t = CUDART_INF_F; best.u = 0.0f; best.v = 0.0f; best.instance = HIT_NONE; best.primitive = -1;

	for (int i = 0; i < P.sphere_count; i++) {
		float t;
		if (sphere_test(P.spheres + i, o, d, best.t, t)) { best.t = t; best.instance = HIT_SPHERE; best.primitive = i; }
	}
	for (int i = 0; i < P.plane_count; i++) {
		float t = plane_t(P.planes + i, o, d);
		if (t > RT_EPSILON && t < best.t) { best.t = t; best.instance = HIT_PLANE; best.primitive = i; }
	}

	// TopLevelBVH::trace, TopLevelBVH.cpp:53-87
	int tlas_stack[BSTACK > 32 ? RT_TLAS_STACK_SIZE : RT_TLAS_STACK_SMALL];   // the deep instantiations carry both full stacks
	int blas_stack[BSTACK];

	const v3 inv_d = inv_direction(d);
	const bool exact = !ray_is_regular(o, inv_d);
	const v3 so = slab_origin(o, inv_d);
	int stack_size = 1;
	tlas_stack[0] = 0;

	while (stack_size > 0) {
		const int index = tlas_stack[--stack_size];
		const float4 n0 = __ldg(P.tlas_nodes + 2 * index), n1 = __ldg(P.tlas_nodes + 2 * index + 1);
		if (COUNT) cnt.nodes++;

		if (!aabb_intersect(n0, n1, so, inv_d, best.t, exact)) continue;

		const int left_or_first = __float_as_int(n1.z), count = __float_as_int(n1.w);
		if ((count & 0x3FFFFFFF) > 0) {
			for (int i = left_or_first; i < left_or_first + count; i++) {
				const int instance = __ldg(P.tlas_indices + i);
				const DeviceInstance inst = load_instance_inverse(P.instances + instance);
				if (COUNT) cnt.instances++;
				// Mesh::trace, Mesh.cpp:17-31 (the differentials are transformed in shade.cuh, for the winner only)
				const v3 local_d = transform_direction(inst.inverse, d);
#ifdef RT_OCTANT_THREADED
				if (!COUNT) { // BVH_TRAVERSE_TREE_NAIVE, the left child first at every node, is the order of octant 7
					const int octant = P.traversal_naive ? 7 : (local_d.x > 0.0f ? 1 : 0) | (local_d.y > 0.0f ? 2 : 0) | (local_d.z > 0.0f ? 4 : 0);
					const int records_per_copy = __ldg(&P.instances[instance].reserved[0]);
					blas_trace_threaded(P.shadow_nodes + 2 * (size_t(inst.shadow_node_offset) + size_t(1 + octant) * size_t(records_per_copy)),
					                    P.triangles_test + 3 * size_t(inst.triangle_offset), transform_position(inst.inverse, o), local_d, instance, best);
				} else
#endif
				blas_trace<COUNT>(P.nodes + 2 * size_t(inst.node_offset), P.triangles_test + 3 * size_t(inst.triangle_offset),
				                  transform_position(inst.inverse, o), local_d, instance, best, blas_stack, cnt, traversal_order_mask(P, local_d));
			}
		} else if (P.traversal_naive || visit_left_first(count, d)) {
			tlas_stack[stack_size++] = left_or_first + 1;
			tlas_stack[stack_size++] = left_or_first;
		} else {
			tlas_stack[stack_size++] = left_or_first;
			tlas_stack[stack_size++] = left_or_first + 1;
		}
	}
}

// Scene::intersect_primitives, Scene.cpp:179-190
template<bool COUNT, int BSTACK>
__device__ __forceinline__ bool trace_any(const SceneParams & P, v3 o, v3 d, float max_distance, Counters & cnt) {
	for (int i = 0; i < P.sphere_count; i++) if (sphere_intersect(P.spheres + i, o, d, max_distance)) return true;
	for (int i = 0; i < P.plane_count; i++) {
		float t = plane_t(P.planes + i, o, d);
		if (t > RT_EPSILON && t < max_distance) return true;
	}

	// TopLevelBVH::intersect, TopLevelBVH.cpp:89-129
	int tlas_stack[BSTACK > 32 ? RT_TLAS_STACK_SIZE : RT_TLAS_STACK_SMALL];   // the deep instantiations carry both full stacks
	int blas_stack[COUNT ? BSTACK : 1];   // the shipped any-hit walk is stackless (blas_shadow); only the counting build pushes

	const v3 inv_d = inv_direction(d);
	const bool exact = !ray_is_regular(o, inv_d);
	const v3 so = slab_origin(o, inv_d);
	int stack_size = 1;
	tlas_stack[0] = 0;

	while (stack_size > 0) {
		const int index = tlas_stack[--stack_size];
		const float4 n0 = __ldg(P.tlas_nodes + 2 * index), n1 = __ldg(P.tlas_nodes + 2 * index + 1);
		if (COUNT) cnt.nodes++;

		if (!aabb_intersect(n0, n1, so, inv_d, max_distance, exact)) continue;

		const int left_or_first = __float_as_int(n1.z), count = __float_as_int(n1.w);
		if ((count & 0x3FFFFFFF) > 0) {
			for (int i = left_or_first; i < left_or_first + count; i++) {
				const int instance = __ldg(P.tlas_indices + i);
				const DeviceInstance inst = load_instance_inverse(P.instances + instance);
				if (COUNT) cnt.instances++;
				const float4 * hot = P.triangles_test + 3 * size_t(inst.triangle_offset);
				const v3 local_o = transform_position(inst.inverse, o), local_d = transform_direction(inst.inverse, d);
#ifdef RT_WIDE_SHADOW
				if (!COUNT) {
					const v3 local_inv_d = inv_direction(local_d);
					const int wide_offset = __ldg(&P.instances[instance].reserved[1]);
					if (wide_offset >= 0 && ray_is_regular(local_o, local_inv_d)) {
						if (blas_shadow_wide(P.wide_nodes + 8 * size_t(wide_offset), hot, local_o, local_inv_d, local_d, max_distance)) return true;
						continue;
					}
				}
#endif
				// the counting build counts the reference's walk, or (count_traversal = 2) the walk that ships
				if (COUNT && !P.count_shipped_walk ? blas_intersect<COUNT>(P.nodes + 2 * size_t(inst.node_offset), hot, local_o, local_d, max_distance, blas_stack, cnt, traversal_order_mask(P, local_d))
				                                   : blas_shadow<COUNT>(P.shadow_nodes + 2 * size_t(inst.shadow_node_offset), hot, local_o, local_d, max_distance, cnt)) return true;
			}
		} else if (P.traversal_naive || visit_left_first(count, d)) {
			tlas_stack[stack_size++] = left_or_first + 1;
			tlas_stack[stack_size++] = left_or_first;
		} else {
			tlas_stack[stack_size++] = left_or_first;
			tlas_stack[stack_size++] = left_or_first + 1;
		}
	}
	return false;
}

} // namespace RT_NAMESPACE
