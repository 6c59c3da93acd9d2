// rt_types.h — how the scene lies in HBM: the plain records shared by the C ABI layer (b200rt.cu) and by the kernel
// translation units (kernels_tu.cu, compiled once per arithmetic mode). No device code in this file.
//
// Everything a ray touches is stored as 16-byte aligned records read with LDG.128:
//   BVH node      32 B = 2 x float4   {min.xyz, max.x} {max.yz, left_or_first, count}   (BVHNode.h:10-16, same bytes)
//   TriangleHot   48 B = 3 x float4   {p0.xyz, e1.x} {e1.yz, e2.xy} {e2.z, -, -, -}      (36 B of payload, BottomLevelBVH.h:6-10)
//   TriangleCold  64 B = 4 x float4   same bytes as BottomLevelBVH.h:12-22
//   texel         16 B = float4       linear RGB + pad                                   (12 B of payload, Texture.cpp:146)
//   instance     128 B = 8 x float4   inverse rows 0-2, world rows 0-2, {node_offset, triangle_offset, material_offset, -}
// The child pair of an inner node (left, left + 1) is one aligned 64-byte line, as in the
// reference (root 0, node 1 unused, BottomLevelBVH.cpp:92,132).
#pragma once
#ifdef RT_EMULATE
#include "cuda_emu.h"              // tests/emu: the kernels' source run on the host, fibre per CUDA thread (test infrastructure)
#else
#include <cuda_runtime.h>
#endif
#include <stdint.h>
#include "../../include/b200rt.h"

namespace rt {

// Row-major 3x4 affine matrix (three float4 rows): row i = (m(0,i), m(1,i), m(2,i), m(3,i)) of the
// reference's column-major Matrix4, so the nested fused multiply-adds of Matrix4.h:39-85 read
// x' = fma(r.x, p.x, fma(r.y, p.y, fma(r.z, p.z, r.w))).
struct Affine { float4 r0, r1, r2; };

#define RT_TLAS_STACK_SIZE 32   // entries of the TLAS traversal stack; b200rt.cu rejects a frame whose TLAS is deeper
#define RT_TLAS_STACK_SMALL 8   // ... of the instantiations for shallow trees (a TLAS over up to ~100 instances): the kernels are picked per frame
#define RT_MAX_BOUNCES 8        // NUMBER_OF_BOUNCES the kernels are instantiated up to

struct DeviceInstance {
	Affine inverse;   // Mesh::transform_inv
	Affine world;     // Transform::world_matrix
	int node_offset, triangle_offset, material_offset, shadow_node_offset;
	int reserved[4];    // [0] records per octant copy (RT_OCTANT_THREADED), [1] first wide record or -1 (RT_WIDE_SHADOW); keeps the record at 128 bytes (one L2 line)
};
static_assert(sizeof(DeviceInstance) == 128, "instance record is 128 bytes");

struct DeviceTexture {   // Texture.h:8-18
	long long texel_offset;
	int width, height;
	int mipmapped, mip_levels;
	int mip_offsets[B200RT_MAX_MIP_LEVELS];
};

struct DeviceMaterial {  // Material.h:7-14, 48 bytes
	float4 diffuse_texture;      // diffuse.xyz, as_float(texture index)
	float4 reflection_ior;       // reflection.xyz, index_of_refraction
	float4 transmittance;        // transmittance.xyz, -
};

struct DeviceSphere { float4 center_r2; float4 rinv_mat; };                  // centre.xyz, radius^2 | 1/radius, as_float(material id)
struct DevicePlane  { float4 normal_d; float4 u_axis_mat; float4 v_axis; };  // normal.xyz, distance | u.xyz, as_float(material id) | v.xyz
struct DeviceLight  { float4 colour_kind; float4 position_inner; float4 negdir_outer; }; // colour.xyz, as_float(kind) | position.xyz, inner_cutoff | -direction.xyz, outer_cutoff

// Per-frame + static pointers, passed to the kernels by value (kernel parameter space).
struct SceneParams {
	// static (uploaded once)
	const float4 * __restrict__ nodes;          // all BLAS, concatenated
	const float4 * __restrict__ shadow_nodes;   // the same trees threaded for the any-hit traversal (threaded_nodes.h: build_threaded_nodes)
	const float4 * __restrict__ wide_nodes;     // -DRT_WIDE_SHADOW: the same trees, four children per record (threaded_nodes.h: build_wide_nodes)
	const float4 * __restrict__ triangles_hot;
	const float4 * __restrict__ triangles_test; // what the traversal tests: triangles_hot, or the precomputed transforms of the FAST unit (RT_FAST_TRI)
	const uint2  * __restrict__ texels_half;    // the texel pool as four halves per texel (FAST unit, RT_FAST_TEXEL)
	const float4 * __restrict__ triangles_cold;
	const DeviceMaterial * __restrict__ materials;
	const DeviceTexture  * __restrict__ textures;
	const float4 * __restrict__ texels;
	const float4 * __restrict__ sky;
	int sky_size;
	// per frame
	const float4 * __restrict__ tlas_nodes;
	const int    * __restrict__ tlas_indices;
	const DeviceInstance * __restrict__ instances;
	const DeviceSphere   * __restrict__ spheres;
	const DevicePlane    * __restrict__ planes;
	const DeviceLight    * __restrict__ lights;
	int sphere_count, plane_count, light_count;
	float cam_position[3], cam_top_left[3], cam_x_axis[3], cam_y_axis[3];
	float ambient[3];
	// configuration (Config.h mirror)
	int width, height, bounces, mipmap_filter;
	float max_anisotropy;
	int traversal_naive, texture_sample_mode;   // BVH_TRAVERSAL_STRATEGY, TEXTURE_SAMPLE_MODE
	int count_shipped_walk;                     // counting build: count the any-hit walk that ships (threaded, right child first) instead of the reference's
	int tiles_x, tiles_y, tile_stride, tile_phase, tile_count_mine;
	// schedule (kernels.cuh, tile_schedule_kernel): this GPU's tiles in hand-out order, and where this
	// frame's cost per tile is accumulated for the next frame's order
	const unsigned * __restrict__ tile_order;
	unsigned * __restrict__ tile_cost;
	// multi-GPU exchange fused into the pixel store (b200rt_set_band_targets): band b of the frame
	// (band_rows rows, the last band takes the rest) belongs to the GPU whose image is band_targets[b]
	unsigned * band_targets[B200RT_MAX_BANDS];
	int band_count, band_rows, band_halo;   // band_count 0: one image, the kernel's frame_buffer argument
	// deep-pixel queue (render.cuh): pools in HBM, bump-allocated per frame through FrameCounters::q_*
	float4   * q_nodes;          // 6 float4 per node record
	float4   * q_bounce;         // 5 float4 per bounce item
	unsigned * q_bounce_ready;   // item i is readable once ready[i] == frame_serial
	unsigned q_node_capacity, q_bounce_capacity;
	unsigned frame_serial;       // never 0, different for every launch of a context
	int queue_mode;              // 0 off; 1: pixels whose primary ray hits a transmissive material; 2: every pixel whose primary hit spawns rays
	int queue_inline_below;      // a popped ray with at most this many bounces left is walked to the end by the thread that popped it
	int queue_idle_looks;        // a warp that finds neither item nor unit looks this many times more (2 / 4 / 8 / ... us apart) before it leaves
	int queue_idle_shift;        // ... the pause doubles this many times at most (2: 8 us)
	int queue_keep_refraction;   // of two children the exporting thread keeps the reflection ray (0), the refraction ray (1) or the one that stays inside the medium (2) and queues the other
	int queue_item_batch;        // while tile units are left, a warp turns to the queue only when it holds at least this many items (32: a full warp)
	int blas_stack_entries;      // deepest BLAS + 1: what a push-both-pop-one traversal can hold
	int tlas_stack_entries;      // this frame's TLAS depth + 1
};

#define RT_QUEUE_EXPORT_COST 300u   // tile cost (clock64 >> 6) booked per pixel whose ray tree went to the queue: ~10 us, a deep pixel's depth-first share of its unit
#ifdef RT_EMULATE
#define RT_QUEUE_WATCHDOG_NS 900000000000ull // (tests/emu: an emulated frame takes seconds of real time, and a fibre that waits is not a stall)
#else
#define RT_QUEUE_WATCHDOG_NS 2000000000ull   // kernels.cuh: render_queue_kernel gives a frame up after waiting this long on unpublished items
#endif

struct FrameCounters {   // accumulated with one atomicAdd per warp
	unsigned long long primary, shadow, reflection, refraction;
	unsigned long long nodes, triangles, texels, accepted, instances;
	unsigned int next_unit;      // work-stealing cursor (WorkerThread.cpp:53-54)
	unsigned int pad;
	unsigned long long first_idle_ns, last_done_ns; // globaltimer when the first (stored as ~t) / the last warp ran out of work: the tail of the launch
	unsigned long long first_start_ns, pad2[3];     // globaltimer when the first block started (stored as ~t)
	// deep-pixel queue (render.cuh). Producers' cursors, consumers' heads and the count of outstanding work (tile units
	// + items not finished) each on a 128-byte line of their own: they are hit by different warps at the same time
	unsigned int q_node_reserved, q_bounce_reserved, q_unused0, q_fallbacks, q_pad0[28];
	unsigned int q_bounce_head, q_abort, q_pad1[30];   // q_abort: the queue kernel's watchdog fired (kernels.cuh), the frame is incomplete
	int q_outstanding; unsigned int q_pad2[31];
};
static_assert(sizeof(FrameCounters) == 512, "four 128-byte lines");

} // namespace rt
