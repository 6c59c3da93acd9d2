// render.cuh — Raytracer::bounce for one ray per thread: shade_subtree walks the whole ray tree of one call of bounce()
// depth-first with an explicit frame stack. With the deep-pixel queue on, the TOP node of a walk may instead be
// exported: its children (reflection / refraction rays) and its shadow rays go to device-side queues as items of
// their own, each walked by whichever thread pops it, and the node's value is assembled bottom-up when the last of
// them has reported (kernels.cuh has the loop that feeds persistent warps from the queues). Same bits either way.
//
// Reference: Raytracer::render_tile (Raytracer.cpp:3-85), Raytracer::bounce (:87-400), Window::plot (Window.h:56-65).
#pragma once
#include "shade.cuh"

namespace RT_NAMESPACE {

#ifndef RT_BLOCK_THREADS
#define RT_BLOCK_THREADS 256
#endif
#ifndef RT_MIN_BLOCKS
#define RT_MIN_BLOCKS 4   // resident 256-thread blocks per SM the register allocation is sized for: 64 registers, 32 warps/SM.
                          // Measured on the bench frame (profiles/r01_variants_traversal.log): 2 -> 1.87 ms, 3 -> 1.53, 4 -> 1.46, 5 -> 1.48, 6 -> 1.57, 8 -> 1.69
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

// Where a pixel of row py goes when the frame is exchanged in bands (include/b200rt.h:
// b200rt_set_band_targets): into the image of the GPU that owns the row's band and, within band_halo
// rows of a band border, into the neighbour's too — cyclically, because the FXAA taps wrap at the
// frame border (GL_REPEAT, Window.cpp:48-50). Plain stores over NVLink peer memory: the exchange
// rides on the kernel's last instruction, every link carries 1/N of the frame, nobody gathers.
__device__ __forceinline__ void store_pixel_banded(const SceneParams & P, size_t p, int py, unsigned value) {
	const int last = P.band_count - 1;
	int b = py / P.band_rows;
	if (b > last) b = last;
	P.band_targets[b][p] = value;
	const int begin = b * P.band_rows, end = b == last ? P.height : begin + P.band_rows;
	if (py - begin < P.band_halo) P.band_targets[b == 0 ? last : b - 1][p] = value;
	if (end - py <= P.band_halo)  P.band_targets[b == last ? 0 : b + 1][p] = value;
}

// The one place a finished pixel is written (by the thread that traced it, or by whichever thread completed
// the root node of a queued pixel).
struct PixelOut { unsigned * frame_buffer; float * colour_float; };
__device__ __forceinline__ void write_pixel(const SceneParams & P, const PixelOut & out, int px, int py, v3 colour) {
	const size_t p = size_t(px) + size_t(P.width) * py;
	if (P.band_count == 0) out.frame_buffer[p] = plot_colour(colour);
	else store_pixel_banded(P, p, py, plot_colour(colour));
	if (out.colour_float) { out.colour_float[3 * p] = colour.x; out.colour_float[3 * p + 1] = colour.y; out.colour_float[3 * p + 2] = colour.z; }
}

// ---------------------------------------------------------------- pieces of bounce() both forms share ---
enum { FRAME_AFTER_REFLECTION = 1, FRAME_AFTER_REFRACTION = 2, FRAME_HAS_REFRACTION = 4, FRAME_TIR = 8, FRAME_ENTERING = 16, FRAME_HAS_REFLECTION = 32 };

// One light seen from a hit point: direction and length of the shadow ray (Raytracer.cpp:155-198; DirectionalLight.h:9-11)
struct LightSample { float4 l0, l1, l2; v3 to_light; float distance, distance_squared; int kind; };
__device__ __forceinline__ void light_sample(const SceneParams & P, int i, v3 point, LightSample & s) {
	s.l0 = __ldg(&P.lights[i].colour_kind); s.l1 = __ldg(&P.lights[i].position_inner); s.l2 = __ldg(&P.lights[i].negdir_outer);
	s.kind = __float_as_int(s.l0.w);
	s.distance = CUDART_INF_F; s.distance_squared = 0.0f;
	if (s.kind == 2) {
		s.to_light = V(s.l2);
	} else {
		s.to_light = V(s.l1) - point;
		s.distance_squared = dot(s.to_light, s.to_light);
		s.distance         = sqrtf(s.distance_squared);
		s.to_light = divf(s.to_light, s.distance);
	}
}
// diffuse + what an unoccluded light adds (PointLight.h:9-11, SpotLight.h:17-33, DirectionalLight.h:9-11)
__device__ __forceinline__ v3 light_add(v3 diffuse, const LightSample & s, v3 normal, v3 to_camera) {
	if (s.kind == 2) return diffuse + light_base(V(s.l0), normal, s.to_light, to_camera);
	v3 lighting = divf(light_base(V(s.l0), normal, s.to_light, to_camera), s.distance_squared);
	if (s.kind == 1) {
		const float d = dot(s.to_light, V(s.l2));
		if (!(d > s.l2.w)) lighting = V(0.0f, 0.0f, 0.0f);
		else {
			float radial_falloff = (d - s.l2.w) / (s.l1.w - s.l2.w);
			if (radial_falloff > 1.0f) radial_falloff = 1.0f;
			lighting = lighting * radial_falloff;
		}
	}
	return diffuse + lighting;
}
__device__ __forceinline__ v3 lit_result(v3 diffuse, v3 material_diffuse) { // result = madd(diffuse, material_diffuse, result = 0), :200
	return V(__fmaf_rn(diffuse.x, material_diffuse.x, 0.0f), __fmaf_rn(diffuse.y, material_diffuse.y, 0.0f), __fmaf_rn(diffuse.z, material_diffuse.z, 0.0f));
}

// Everything about the refraction branch that only depends on this hit (Raytracer.cpp:273-342, Schlick :379-391).
// Returns the FRAME_* flags of the branch; `refracted` and F_r are filled unless the hit reflects totally.
__device__ __forceinline__ int refraction_setup(const Ray & ray, const HitRecord & hit, float ior, float d_dot_n, float dDN_dx, float dDN_dy, Ray & refracted, float & F_r) {
	int flags = FRAME_HAS_REFRACTION;
	const bool entering = d_dot_n < 0.0f;
	if (entering) flags |= FRAME_ENTERING;

	const float n_1 = entering ? 1.0f : ior;
	const float n_2 = entering ? ior : 1.0f;

	float    cos_theta = entering ? 0.0f - d_dot_n : d_dot_n;
	const v3 normal    = entering ? hit.normal : neg(hit.normal);

	const float eta = n_1 / n_2;
	const float k   = 1.0f - (eta * eta * (1.0f - (cos_theta * cos_theta)));

	if (k < 0.0f) return flags | FRAME_TIR; // :310-314

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

	float r_0 = (n_1 - n_2) / (n_1 + n_2);
	r_0 = r_0 * r_0;
	if (n_1 > n_2) cos_theta = 0.0f - dot(refracted.direction, normal);

	const float one_minus_cos         = 1.0f - cos_theta;
	const float one_minus_cos_squared = one_minus_cos * one_minus_cos;
	F_r = r_0 + ((1.0f - r_0) * one_minus_cos_squared) * (one_minus_cos_squared * one_minus_cos);
	return flags;
}
// :249-271
__device__ __forceinline__ Ray reflection_ray(const Ray & ray, const HitRecord & hit, float d_dot_n, float dDN_dx, float dDN_dy) {
	Ray reflected;
	reflected.origin    = hit.point;
	reflected.direction = ray.direction - hit.normal * (2.0f * d_dot_n); // Math::reflect, Math.h:28-31
	reflected.dO_dx = hit.dO_dx;
	reflected.dO_dy = hit.dO_dy;
	reflected.dD_dx = ray.dD_dx - (hit.dN_dx * d_dot_n + hit.normal * dDN_dx) * 2.0f;
	reflected.dD_dy = ray.dD_dy - (hit.dN_dy * d_dot_n + hit.normal * dDN_dy) * 2.0f;
	return reflected;
}
// The refraction child returned: Beer's law with the child's hit distance (:350-376), Fresnel blend (:393-395)
__device__ __forceinline__ v3 refraction_blend(const SceneParams & P, int material_id, int flags, float F_r, v3 result, v3 colour_reflection, v3 returned, float returned_distance) {
	v3 colour_refraction = returned;
	if (flags & FRAME_ENTERING) {
		const float4 m2 = __ldg(reinterpret_cast<const float4 *>(P.materials + material_id) + 2);
		colour_refraction.x = colour_refraction.x * expf((m2.x - 1.0f) * returned_distance);
		colour_refraction.y = colour_refraction.y * expf((m2.y - 1.0f) * returned_distance);
		colour_refraction.z = colour_refraction.z * expf((m2.z - 1.0f) * returned_distance);
	}
	const float F_t = 1.0f - F_r;
	const v3 blend = colour_reflection * F_r + colour_refraction * F_t; // :393
	return result + blend;                                               // :395
}

// ------------------------------------------------------------------------------ deep-pixel queue ---
// North_star's "wavefront traversal with ray compaction", restricted to where it pays. A deep pixel is not walked by
// one thread: the top node of a walk can be EXPORTED — it becomes a node record in HBM and its reflection / refraction
// rays become items of the bounce queue, which all persistent warps pop (kernels.cuh), so the sibling subtrees of a
// fan-out run side by side on whatever warps are free. Every popped ray is again a walk whose top node may be
// exported, so a deep tree unfolds level by level. A node's value is assembled when the last of its children has
// reported (a countdown in the record), by the reporting thread, with the reference's sums in the reference's order:
// the same bits as the recursive form. Shadow rays stay with the thread that shades the node (a second queue for them
// was built and measured: slower, and two queues can starve each other — kernels.cuh).
//
//   node record  6 x float4   {parent, flags | which << 8, pending, material id} {point, distance} {normal, F_r}
//                             {the node's direct lighting, -}
//                             {reflection child's colour, -} {refraction child's colour, its distance}
//   bounce item  5 x float4   the Ray (18 floats), parent node, which | bounces_left << 1     (parent = QUEUE_NULL_ITEM: skip)
// An item is visible to consumers once its word in the queue's `ready` array holds the frame's serial number (written
// after a __threadfence()). Records and items are bump-allocated per frame; when a pool is exhausted the node is not
// exported and its subtree is walked by the thread that holds it: always correct, never blocked. Records and items
// are read and written with .cg accesses (L2 only): another SM's L1 may hold a stale copy of the line.
#define QUEUE_NULL_ITEM int(0x80000000u)

__device__ __forceinline__ float as_f(int v) { return __int_as_float(v); }
__device__ __forceinline__ int   as_i(float v) { return __float_as_int(v); }
__device__ __forceinline__ unsigned peek(const unsigned * p) { return *reinterpret_cast<const volatile unsigned *>(p); }   // a fresh value, not this SM's L1 copy
__device__ __forceinline__ int      peek(const int * p)      { return *reinterpret_cast<const volatile int *>(p); }

// Value of a node whose children have all reported (the record is complete and no longer written).
__device__ __forceinline__ v3 node_value(const SceneParams & P, const float4 * rec, float & distance_out, int & parent, int & which) {
	const float4 r0 = __ldcg(rec), r1 = __ldcg(rec + 1), r2 = __ldcg(rec + 2), r3 = __ldcg(rec + 3);
	parent = as_i(r0.x);
	const int flags = as_i(r0.y) & 0xff; which = (as_i(r0.y) >> 8) & 1;
	const int material_id = as_i(r0.w);
	distance_out = r1.w;

	v3 result = V(r3);   // the direct lighting, finished by the thread that exported the node
	v3 colour_reflection = V(0.0f, 0.0f, 0.0f);
	if (flags & FRAME_HAS_REFLECTION) {
		const float4 m1 = __ldg(reinterpret_cast<const float4 *>(P.materials + material_id) + 1);
		colour_reflection = V(m1) * V(__ldcg(rec + 4));
		result = result + colour_reflection;
		if (!(flags & FRAME_HAS_REFRACTION)) return result;
		if (flags & FRAME_TIR) return result + colour_reflection;
	} else if (!(flags & FRAME_HAS_REFRACTION) || (flags & FRAME_TIR)) {
		return result;
	}
	const float4 r5 = __ldcg(rec + 5);
	return refraction_blend(P, material_id, flags, r2.w, result, colour_reflection, V(r5), r5.w);
}

// A walk's colour (and hit distance) goes to its parent: the pixel, or a node record — and whoever brings a record's
// countdown to zero computes the node's value and carries it one level up.
__device__ __noinline__ void deliver(const SceneParams & P, const PixelOut & out, int parent, int which, v3 value, float distance) {
	for (;;) {
		if (parent < 0) { // the root's parent is the pixel
			const int p = -(parent + 1);
			write_pixel(P, out, p % P.width, p / P.width, value);
			return;
		}
		float4 * rec = P.q_nodes + 6 * size_t(parent);
		__stcg(rec + 4 + which, make_float4(value.x, value.y, value.z, distance));
		__threadfence();
		if (atomicSub(reinterpret_cast<int *>(rec) + 2, 1) != 1) return;
		__threadfence();
		value = node_value(P, rec, distance, parent, which);
	}
}
__device__ __forceinline__ void write_bounce_item(const SceneParams & P, unsigned slot, const Ray & r, int parent, int which, int bounces_left) {
	float4 * item = P.q_bounce + 5 * size_t(slot);
	__stcg(item,     make_float4(r.origin.x, r.origin.y, r.origin.z, r.direction.x));
	__stcg(item + 1, make_float4(r.direction.y, r.direction.z, r.dO_dx.x, r.dO_dx.y));
	__stcg(item + 2, make_float4(r.dO_dx.z, r.dO_dy.x, r.dO_dy.y, r.dO_dy.z));
	__stcg(item + 3, make_float4(r.dD_dx.x, r.dD_dx.y, r.dD_dx.z, r.dD_dy.x));
	__stcg(item + 4, make_float4(r.dD_dy.y, r.dD_dy.z, as_f(parent), as_f(which | (bounces_left << 1))));
}
__device__ __forceinline__ void read_bounce_item(const SceneParams & P, unsigned slot, Ray & r, int & parent, int & which, int & bounces_left) {
	const float4 * item = P.q_bounce + 5 * size_t(slot);
	const float4 a = __ldcg(item), b = __ldcg(item + 1), c = __ldcg(item + 2), d = __ldcg(item + 3), e = __ldcg(item + 4);
	r.origin = V(a.x, a.y, a.z); r.direction = V(a.w, b.x, b.y);
	r.dO_dx = V(b.z, b.w, c.x); r.dO_dy = V(c.y, c.z, c.w);
	r.dD_dx = V(d.x, d.y, d.z); r.dD_dy = V(d.w, e.x, e.y);
	parent = as_i(e.z); which = as_i(e.w) & 1; bounces_left = as_i(e.w) >> 1;
}

// What a walk needs to export its top node.
struct QueueContext { FrameCounters * counters; int parent, which; };

// One reservation for everything an export needs, for the lanes that happen to stand here together (whoever is converged
// at this point — any grouping is correct, a larger one is cheaper): a node record per lane and an item slot per lane
// that queues a child, taken with ONE 64-bit atomicAdd on the pair of cursors (records in the low word, item slots in
// the high word), and the queued children added to the frame's outstanding work with one more.
__device__ __forceinline__ void group_reserve_export(FrameCounters * counters, bool queues_a_child, unsigned & node_slot, unsigned & item_slot) {
	unsigned long long * cursors = reinterpret_cast<unsigned long long *>(&counters->q_node_reserved);   // {q_node_reserved, q_bounce_reserved}
#ifdef RT_EMULATE
	const unsigned long long old = atomicAdd(cursors, 1ull | (static_cast<unsigned long long>(queues_a_child ? 1 : 0) << 32));
	if (queues_a_child) atomicAdd(&counters->q_outstanding, 1);
	node_slot = unsigned(old); item_slot = unsigned(old >> 32);
#else
	const unsigned mask = __activemask();
	const unsigned queueing = __ballot_sync(mask, queues_a_child);
	const int lane = threadIdx.x & 31, leader = __ffs(mask) - 1;
	unsigned lo = 0, hi = 0;
	if (lane == leader) {
		const unsigned records = unsigned(__popc(mask)), items = unsigned(__popc(queueing));
		const unsigned long long old = atomicAdd(cursors, static_cast<unsigned long long>(records) | (static_cast<unsigned long long>(items) << 32));
		// before anything is visible: the count never runs low. The returned value is USED, so the add has been performed
		// at L2 before the shuffle below hands out the slots — every lane's stores (payload, fence, ready word) are issued
		// after that; a consumer that sees an item of this group cannot bring the count to zero ahead of this add.
		const int before = atomicAdd(&counters->q_outstanding, int(items));
		lo = unsigned(old) + (before == int(0x80000000u) ? 1u : 0u); hi = unsigned(old >> 32);
	}
	lo = __shfl_sync(mask, lo, leader); hi = __shfl_sync(mask, hi, leader);
	__syncwarp(mask);
	const unsigned below = (1u << lane) - 1u;
	node_slot = lo + unsigned(__popc(mask & below));
	// the group's items in lane order: the lanes of a group are neighbouring pixels and the items are all refraction rays,
	// so whoever draws 32 consecutive tickets gets a coherent warp
	item_slot = hi + unsigned(__popc(queueing & below));
#endif
}

// Exports the top node of a walk as a record. The walking thread then goes on with ONE child itself, as a new walk whose
// parent is the record, and only the OTHER child of a node that has two goes to the queue (queued_ray, queued_which): a chain of single-child nodes stays in its thread without a
// single queue round trip, a fan-out splits off exactly one item, and the rays that stay are the coherent ones (the
// warp's neighbouring pixels, one level down). Returns false when a pool is exhausted (nothing has been published then,
// and the caller goes on walking depth-first). result: the node's finished direct lighting; node_out: the record.
__device__ __forceinline__ bool export_node(const SceneParams & P, const QueueContext & q, const HitRecord & hit, float distance, int flags, float F_r, v3 result,
                                            const Ray & queued_ray, int queued_which, int bounces_left, Counters & cnt, int & node_out) {
	const bool has_child[2] = { (flags & FRAME_HAS_REFLECTION) != 0, (flags & FRAME_HAS_REFRACTION) && !(flags & FRAME_TIR) };
	const int n_children = (has_child[0] ? 1 : 0) + (has_child[1] ? 1 : 0);
	const bool queues_a_child = n_children == 2;
	unsigned node_slot, item_slot;
	group_reserve_export(q.counters, queues_a_child, node_slot, item_slot);
	if (node_slot >= P.q_node_capacity || (queues_a_child && item_slot >= P.q_bounce_capacity)) {
		// a pool is exhausted. An item slot inside the pool must still be published (somebody may hold a ticket for it): as a null item
		if (queues_a_child && item_slot < P.q_bounce_capacity) {
			__stcg(P.q_bounce + 5 * size_t(item_slot) + 4, make_float4(0.0f, 0.0f, as_f(QUEUE_NULL_ITEM), 0.0f));
			__threadfence();
			__stcg(P.q_bounce_ready + item_slot, P.frame_serial);
		}
		if (queues_a_child) atomicSub(&q.counters->q_outstanding, 1);   // it will not become work (the walk that queued nothing still holds its own count)
		atomicAdd(&q.counters->q_fallbacks, 1u);
		return false;
	}

	const int node = int(node_slot);
	float4 * rec = P.q_nodes + 6 * size_t(node);
	__stcg(rec,     make_float4(as_f(q.parent), as_f(flags | (q.which << 8)), as_f(n_children), as_f(hit.material_id)));
	__stcg(rec + 1, make_float4(hit.point.x, hit.point.y, hit.point.z, distance));
	__stcg(rec + 2, make_float4(hit.normal.x, hit.normal.y, hit.normal.z, F_r));
	__stcg(rec + 3, make_float4(result.x, result.y, result.z, 0.0f));
	if (queues_a_child) { // payload, ONE fence, then the ready word: record and item are in L2 before any consumer can see the item that names the record
		if (queued_which == 0) cnt.reflection++; else cnt.refraction++;
		write_bounce_item(P, item_slot, queued_ray, node, queued_which, bounces_left - 1);
		__threadfence();
		__stcg(P.q_bounce_ready + item_slot, P.frame_serial);
	} else {
		__threadfence();   // the record is in L2 before this thread's own delivery into it (or a descendant's) can complete it
	}
	node_out = node;
	return true;
}

// ------------------------------------------------------------------------- depth-first, one thread ---
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

// bounce(ray, bounces_left), iteratively: returns the colour; distance_out = the hit distance bounce() hands to its
// caller (Beer's law one level up). MAXB = frames reserved (>= bounces_left).
// QUEUE: the top node of the walk is exported to the deep-pixel queue when the frame's policy says so (q.parent < 0:
// the walk is a pixel's — queue_mode 1: its primary ray hit a transmissive material, 2: it hit anything that spawns rays;
// q.parent >= 0: the walk is a popped item's — more than queue_inline_below bounces left); `exported` then comes back
// true and the returned colour means nothing: the pixel is written by whoever completes the root record.
template<bool COUNT, int MAXB, int BSTACK, bool QUEUE>
__device__ __forceinline__ v3 shade_subtree_body(const SceneParams & P, Ray ray, int bounces_left, Counters & cnt, float & distance_out, QueueContext * q, bool & exported) {
	BounceFrame frames[MAXB];
	int depth = 0;                // number of pending frames = recursion depth of the current ray

	v3    returned;               // value of the bounce() call that just finished
	float returned_distance;
	bool  top = QUEUE;            // the next hit is the walk's top node

	for (;;) {
		// ---- bounce(ray, bounces_left): Raytracer.cpp:87-201 ---------------------------------
		Candidate best;
		trace_closest<COUNT, BSTACK>(P, ray.origin, ray.direction, best, cnt);

		bool descend = false;

		if (best.instance == HIT_NONE) { // :105-111
			returned = sky_sample<COUNT>(P, ray.direction, cnt);
			returned_distance = CUDART_INF_F;
		} else {
			HitRecord hit;
			build_record(P, ray, best, hit);

			const float4 * mp = reinterpret_cast<const float4 *>(P.materials + hit.material_id);
			const v3 material_diffuse = get_albedo<COUNT>(P, __ldg(mp), hit, cnt);
			const bool lit = dot(material_diffuse, material_diffuse) > 0.0f;

			// QUEUE: may this node be exported?
			bool may_export = false;
			if (QUEUE && top && bounces_left > 0) {
				const float4 m1 = __ldg(mp + 1), m2 = __ldg(mp + 2);
				const bool reflects = dot(V(m1), V(m1)) > 0.0f, refracts = dot(V(m2), V(m2)) > 0.0f;
				may_export = q->parent >= 0 ? (bounces_left > P.queue_inline_below && (reflects || refracts))
				                            : (P.queue_mode == 2 ? (reflects || refracts) : refracts);
			}
			v3 result = V(0.0f, 0.0f, 0.0f);

			if (lit) { // :143-201
				v3 diffuse = V(P.ambient[0], P.ambient[1], P.ambient[2]);
				const v3 to_camera = normalize(V(P.cam_position[0], P.cam_position[1], P.cam_position[2]) - hit.point);

				for (int i = 0; i < P.light_count; i++) { // point, then spot, then directional (:155-198)
					LightSample s;
					light_sample(P, i, hit.point, s);
					cnt.shadow++;
					// one shadow query for all three kinds (a single inlined copy of the traversal)
					if (trace_any<COUNT, BSTACK>(P, hit.point, s.to_light, s.distance, cnt)) continue;
					diffuse = light_add(diffuse, s, hit.normal, to_camera);
				}
				result = lit_result(diffuse, material_diffuse); // :200
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
					frame.F_r = 0.0f;
					int flags = 0;

					const float d_dot_n = dot(ray.direction, hit.normal);
					const float dDN_dx = dot(ray.dD_dx, hit.normal) + dot(ray.direction, hit.dN_dx);
					const float dDN_dy = dot(ray.dD_dy, hit.normal) + dot(ray.direction, hit.dN_dy);

					if (refraction_mask) flags |= refraction_setup(ray, hit, m1.w, d_dot_n, dDN_dx, dDN_dy, frame.refracted, frame.F_r);

					Ray reflected;
					if (reflection_mask) reflected = reflection_ray(ray, hit, d_dot_n, dDN_dx, dDN_dy); // :249-271

					if (QUEUE && may_export) {
						const int node_flags = flags | (reflection_mask ? FRAME_HAS_REFLECTION : 0);
						const bool any_child = reflection_mask || ((flags & FRAME_HAS_REFRACTION) && !(flags & FRAME_TIR));
						int node;
						// of two children the thread keeps one and queues the other (P.queue_keep_refraction: which)
						const bool refraction_child = (flags & FRAME_HAS_REFRACTION) && !(flags & FRAME_TIR);
						// 0: the reflection ray stays; 1: the refraction ray; 2: the ray that stays INSIDE the medium (the refraction
						// ray on the way in, the reflection ray on the way out): the deeper subtree, as a rule
						const bool refraction_stays = P.queue_keep_refraction == 2 ? (flags & FRAME_ENTERING) != 0 : P.queue_keep_refraction != 0;
						const bool keep_reflection = reflection_mask && !(refraction_child && refraction_stays);
						if (any_child && export_node(P, *q, hit, best.t, node_flags, frame.F_r, result, keep_reflection ? frame.refracted : reflected, keep_reflection ? 1 : 0, bounces_left, cnt, node)) {
							// the node is a record now; this thread goes on with one of its children, as a walk of its own whose
							// value is delivered into the record (the other child, if there is one, is in the queue)
							exported = true;
							if (keep_reflection) { cnt.reflection++; ray = reflected; q->which = 0; }
							else                 { cnt.refraction++; ray = frame.refracted; q->which = 1; }
							q->parent = node;
							bounces_left--;
							continue;   // depth is still 0 and `top` still true: the next hit is a top node again
						}
					}

					if (reflection_mask) {
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
		top = false;
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
				returned = refraction_blend(P, frame.material_id, frame.flags, frame.F_r, frame.result, frame.colour_reflection, returned, returned_distance);
				returned_distance = frame.distance;
				depth--; bounces_left++;
			}
		}
		if (!descend) break;
	}
	distance_out = returned_distance;
	return returned;
}
template<bool COUNT, int MAXB, int BSTACK>
__device__ __forceinline__ v3 shade_subtree(const SceneParams & P, const Ray & ray, int bounces_left, Counters & cnt, float & distance_out) {
	bool exported = false;
	return shade_subtree_body<COUNT, MAXB, BSTACK, false>(P, ray, bounces_left, cnt, distance_out, nullptr, exported);
}

} // namespace RT_NAMESPACE
