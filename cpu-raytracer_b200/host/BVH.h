// Bounding volume hierarchies on the host: the 32-byte node of the reference
// (BVHNode.h:10-16), a full-sweep SAH builder (BVHBuilders.h:8-46, BVHPartitions.h:27-114)
// used every frame for the TopLevelBVH and, when no `<obj>.bvh` cache exists, for a
// BottomLevelBVH, plus the reference's cached-on-disk BLAS format
// (BottomLevelBVH.cpp:149-192).
#pragma once
#include <cstdint>
#include <string>
#include <vector>

#include "Math.h"
#include "../../include/b200rt.h"

namespace b200rt {

using BVHNode      = b200rt_bvh_node;      // same bytes as the reference's BVHNode
using TriangleHot  = b200rt_triangle_hot;  // BottomLevelBVH::TriangleHot
using TriangleCold = b200rt_triangle_cold; // BottomLevelBVH::TriangleCold

constexpr int BVH_AXIS_SHIFT = 30;                       // BVHNode.h:5-8
constexpr int BVH_AXIS_MASK  = int(0xC0000000u);

inline int  node_count(const BVHNode & n) { return n.count & ~BVH_AXIS_MASK; }
inline bool node_is_leaf(const BVHNode & n) { return node_count(n) > 0; }

// What the SAH builder needs to know about a primitive (a triangle or a mesh instance):
// its box and the point it is sorted by. Note the reference sorts instances by
// transform.position, not by box centre (Mesh.h:23-25, TopLevelBVH.cpp:33-35).
struct BuildPrimitive {
	AABB    aabb;
	Vector3 position;
};

// Persistent state of one SAH build: three index arrays kept sorted per axis, reused
// between frames for the TLAS exactly as TopLevelBVH::init/build_bvh do
// (TopLevelBVH.cpp:16-45), so that std::sort sees the same input order the reference's does.
struct SAHBuilder {
	std::vector<int>   indices[3];
	std::vector<float> sah;
	std::vector<int>   temp;

	void init(int primitive_count);

	// Sorts the three index arrays by position and builds into `nodes` (resized to
	// 2 * primitive_count, root at 0, node 1 unused). Returns the node count. After the call
	// indices[0] is the leaf-order primitive index array.
	int build(const std::vector<BuildPrimitive> & primitives, std::vector<BVHNode> & nodes);

private:
	void build_recursive(BVHNode & node, const std::vector<BuildPrimitive> & primitives, std::vector<BVHNode> & nodes, int & node_index, int first_index, int index_count);
};

// A triangle as the SBVH builder sees it (Triangle.h:9-26): the three ORIGINAL vertices (not
// p0 + edge, which rounds differently), their box (fattened twice, Triangle.h:16-21) and the
// centroid the object splits sort by.
struct BuildTriangle {
	Vector3 vertex[3];
	AABB    aabb;
	Vector3 position;

	void finish(); // box + centroid from the vertices
};

// Spatial-split BVH (Stich et al. 2009) the way the reference builds it (BVHBuilders.h:48-329,
// BVHPartitions.h:116-378): full-sweep object split with child boxes clipped to the node box,
// 256-bin spatial split considered when the object split's child overlap exceeds alpha = 1e-4 of
// the root area, reference unsplitting, leaves below 3 references. Output is byte-identical to
// the reference's (same decisions, same node numbering, same leaf order) and independent of
// `threads`: subtrees are built as relocatable pieces by a task pool and stitched in depth-first
// order.
struct SBVHResult {
	std::vector<BVHNode> nodes;    // root 0, node 1 unused, child pairs from 2
	std::vector<int>     indices;  // triangle ids in leaf order; size >= triangle count (duplicated references)
};
SBVHResult build_sbvh(const std::vector<BuildTriangle> & triangles, int threads);

// One mesh's acceleration structure + triangles, shared between instances.
struct BottomLevelBVH {
	std::vector<TriangleHot>  triangles_hot;   // flattened: leaf order (after flatten())
	std::vector<TriangleCold> triangles_cold;
	std::vector<BVHNode>      nodes;
	std::vector<int>          original_index;   // flattened triangle -> triangle in the file / OBJ order
	int material_offset = 0;                     // BottomLevelBVH::material_offset

	// `<obj>.bvh`: int32 triangle_count, hot[], cold[], int32 node_count, nodes[],
	// int32 index_count, indices[] (BottomLevelBVH.cpp:149-192). Throws std::runtime_error.
	void load_from_disk(const std::string & bvh_filename);
	void save_to_disk  (const std::string & bvh_filename) const; // writes the un-flattened form back

	// Versioned container around the same payload: the reference's dump has no
	// magic, version or checksum, so a truncated or foreign file is only noticed as garbage geometry.
	//   char magic[8] = "B200BVH\0"; u32 endian_tag = 0x01020304; u32 version = 1;
	//   u32 triangle_count, node_count, index_count; u32 crc32(payload); then the legacy payload.
	// load_container throws std::runtime_error on a bad magic / version / endianness / size / CRC.
	// `<obj>.bvh` stays the legacy dump (the reference reads it); the container lives beside it as
	// `<obj>.b200bvh` and is preferred when valid.
	void load_container(const std::string & filename);
	void save_container(const std::string & filename) const;

	// Builds a plain SAH BVH over (hot, cold) given in OBJ order (MESH_ACCELERATOR_BVH, Config.h:32).
	void build(std::vector<TriangleHot> hot, std::vector<TriangleCold> cold);

	// Builds the SBVH (MESH_ACCELERATOR_SBVH, the reference's default, Config.h:35) from the original
	// vertices; `vertices` holds 3 per triangle, in (hot, cold) order.
	void build_sbvh(std::vector<TriangleHot> hot, std::vector<TriangleCold> cold, const std::vector<Vector3> & vertices, int threads);

private:
	// Un-flattened form as stored on disk
	std::vector<TriangleHot>  file_hot;
	std::vector<TriangleCold> file_cold;
	std::vector<int>          file_indices;

	void flatten(); // BottomLevelBVH.cpp:196-212
	void parse_legacy(const unsigned char * data, size_t size, const std::string & name);
};

} // namespace b200rt
