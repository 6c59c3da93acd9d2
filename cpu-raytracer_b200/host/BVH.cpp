#include "BVH.h"

#include <algorithm>
#include <cstdio>
#include <cstring>
#include <stdexcept>

#include <zlib.h>

namespace b200rt {

static void set_aabb(BVHNode & node, const AABB & box) {
	for (int d = 0; d < 3; d++) { node.aabb_min[d] = box.min[d]; node.aabb_max[d] = box.max[d]; }
}

void SAHBuilder::init(int primitive_count) {
	for (int d = 0; d < 3; d++) {
		indices[d].resize(primitive_count);
		for (int i = 0; i < primitive_count; i++) indices[d][i] = i;
	}
	sah .assign(primitive_count, 0.0f);
	temp.assign(primitive_count, 0);
}

int SAHBuilder::build(const std::vector<BuildPrimitive> & primitives, std::vector<BVHNode> & nodes) {
	const int n = int(primitives.size());
	if (int(indices[0].size()) != n) init(n);

	// TopLevelBVH.cpp:33-35 / BottomLevelBVH.cpp:77-79: one sort per axis, by primitive position
	for (int d = 0; d < 3; d++) {
		std::sort(indices[d].begin(), indices[d].end(), [&](int a, int b) { return primitives[a].position[d] < primitives[b].position[d]; });
	}

	nodes.assign(size_t(2) * std::max(n, 1), BVHNode {});

	int node_index = 2; // root = 0, node 1 stays unused so that child pairs share a 64-byte line
	build_recursive(nodes[0], primitives, nodes, node_index, 0, n);
	return node_index;
}

// Full-sweep SAH, BVHBuilders.h:8-46. Reference quirks kept on purpose, because the node
// numbering decides traversal order and the node-visit counters:
//  * the two child slots are reserved BEFORE the termination test, so a node that turns out
//    to be a leaf leaves an unused pair behind (BVHBuilders.h:20-33);
//  * the split plane is the position of the first primitive of the right half
//    (BVHBuilders.h:36) and ties are resolved by membership in the left half of the split
//    axis' own ordering (BVHPartitions.h:36-55).
void SAHBuilder::build_recursive(BVHNode & node, const std::vector<BuildPrimitive> & primitives, std::vector<BVHNode> & nodes, int & node_index, int first_index, int index_count) {
	AABB bounds = AABB::empty();
	for (int i = first_index; i < first_index + index_count; i++) bounds.expand(primitives[indices[0][i]].aabb);
	bounds.fix_if_needed();
	set_aabb(node, bounds);

	auto make_leaf = [&]() { node.left_or_first = first_index; node.count = index_count; };

	if (index_count < 3) { make_leaf(); return; }

	const int left_child = node_index;
	node_index += 2;

	// --- evaluate every split position along every axis (BVHPartitions.h:76-114)
	float best_cost  = INFINITY;
	int   best_index = -1;
	int   best_dim   = -1;

	for (int d = 0; d < 3; d++) {
		const int * idx = indices[d].data() + first_index;

		AABB left = AABB::empty();
		for (int i = 0; i < index_count - 1; i++) {
			left.expand(primitives[idx[i]].aabb);
			sah[i] = left.surface_area() * float(i + 1);
		}

		AABB right = AABB::empty();
		for (int i = index_count - 1; i > 0; i--) {
			right.expand(primitives[idx[i]].aabb);
			float cost = sah[i - 1] + right.surface_area() * float(index_count - i);
			if (cost < best_cost) { best_cost = cost; best_index = first_index + i; best_dim = d; }
		}
	}

	float parent_cost = bounds.surface_area() * float(index_count);
	if (best_cost >= parent_cost) { make_leaf(); return; }

	// --- partition the two other axes' orderings consistently (BVHPartitions.h:27-73)
	const float split = primitives[indices[best_dim][best_index]].position[best_dim];
	const int   n_left = best_index - first_index;

	for (int d = 0; d < 3; d++) {
		if (d == best_dim) continue;

		int left = 0, right = n_left;
		for (int i = first_index; i < first_index + index_count; i++) {
			const int   p   = indices[d][i];
			const float pos = primitives[p].position[best_dim];

			bool goes_left = pos < split;
			if (pos == split) {
				for (int j = best_index - 1; j >= first_index && primitives[indices[best_dim][j]].position[best_dim] == split; j--) {
					if (indices[best_dim][j] == p) { goes_left = true; break; }
				}
			}
			temp[goes_left ? left++ : right++] = p;
		}
		memcpy(indices[d].data() + first_index, temp.data(), size_t(index_count) * sizeof(int));
	}

	node.left_or_first = left_child;
	node.count = (best_dim + 1) << BVH_AXIS_SHIFT;

	build_recursive(nodes[left_child    ], primitives, nodes, node_index, first_index,          n_left);
	build_recursive(nodes[left_child + 1], primitives, nodes, node_index, first_index + n_left, index_count - n_left);
}

// ---------------------------------------------------------------------------------------------

namespace {
struct File {
	FILE * f;
	File(const std::string & name, const char * mode) : f(fopen(name.c_str(), mode)) { }
	~File() { if (f) fclose(f); }
};
}

// Parses the reference's headerless dump (BottomLevelBVH.cpp:149-192) from memory and validates
// every index before anything is trusted.
void BottomLevelBVH::parse_legacy(const unsigned char * data, size_t size, const std::string & name) {
	size_t at = 0;
	auto read_count = [&](const char * what) {
		int32_t n = 0;
		if (at + 4 > size) throw std::runtime_error(std::string("truncated BVH cache ") + name);
		memcpy(&n, data + at, 4); at += 4;
		if (n < 0) throw std::runtime_error(std::string("bad ") + what + " in BVH cache " + name);
		return size_t(n);
	};
	auto read_array = [&](auto & v, size_t n) {
		using T = typename std::remove_reference<decltype(v)>::type::value_type;
		if (n > (size - at) / sizeof(T)) throw std::runtime_error("truncated BVH cache " + name);
		v.resize(n);
		if (n) memcpy(v.data(), data + at, n * sizeof(T));
		at += n * sizeof(T);
	};

	const size_t triangle_count = read_count("triangle count");
	read_array(file_hot,  triangle_count);
	read_array(file_cold, triangle_count);
	const size_t n_nodes = read_count("node count");
	read_array(nodes, n_nodes);
	const size_t index_count = read_count("index count");
	read_array(file_indices, index_count);

	for (int i : file_indices) if (i < 0 || size_t(i) >= triangle_count) throw std::runtime_error("index out of range in BVH cache " + name);
	// Node 1 is never written by the reference's builders (root 0, child pairs from 2:
	// BottomLevelBVH.cpp:92,132), so a cache written by the reference holds heap garbage there.
	// It is unreachable; clear it instead of validating it.
	if (n_nodes > 1) memset(&nodes[1], 0, sizeof(BVHNode));
	for (const BVHNode & n : nodes) {
		if (node_is_leaf(n)) {
			if (n.left_or_first < 0 || size_t(n.left_or_first) + size_t(node_count(n)) > index_count) throw std::runtime_error("leaf range out of bounds in BVH cache " + name);
		} else if (n.count != 0 && (n.left_or_first < 0 || size_t(n.left_or_first) + 1 >= n_nodes)) {
			throw std::runtime_error("child index out of bounds in BVH cache " + name);
		}
	}
	flatten();
}

static std::vector<unsigned char> read_whole_file(const std::string & name) {
	File file(name, "rb");
	if (!file.f) throw std::runtime_error("cannot open BVH cache " + name);
	std::vector<unsigned char> bytes;
	unsigned char chunk[1 << 16];
	size_t n;
	while ((n = fread(chunk, 1, sizeof(chunk), file.f)) > 0) bytes.insert(bytes.end(), chunk, chunk + n);
	return bytes;
}

void BottomLevelBVH::load_from_disk(const std::string & name) {
	const std::vector<unsigned char> bytes = read_whole_file(name);
	parse_legacy(bytes.data(), bytes.size(), name);
}

void BottomLevelBVH::save_to_disk(const std::string & name) const {
	File file(name, "wb");
	if (!file.f) throw std::runtime_error("cannot write BVH cache " + name);

	auto write_count = [&](size_t n) { int32_t c = int32_t(n); fwrite(&c, sizeof(c), 1, file.f); };

	write_count(file_hot.size());
	fwrite(file_hot .data(), sizeof(TriangleHot),  file_hot .size(), file.f);
	fwrite(file_cold.data(), sizeof(TriangleCold), file_cold.size(), file.f);
	write_count(nodes.size());
	fwrite(nodes.data(), sizeof(BVHNode), nodes.size(), file.f);
	write_count(file_indices.size());
	fwrite(file_indices.data(), sizeof(int), file_indices.size(), file.f);
}

// ------------------------------------------------------------------ versioned container ---
namespace {
const char     CONTAINER_MAGIC[8] = { 'B', '2', '0', '0', 'B', 'V', 'H', 0 };
const uint32_t CONTAINER_ENDIAN   = 0x01020304u;
const uint32_t CONTAINER_VERSION  = 1;
struct ContainerHeader { char magic[8]; uint32_t endian_tag, version, triangle_count, node_count, index_count, crc; };
static_assert(sizeof(ContainerHeader) == 32, "container header is 32 bytes");

std::vector<unsigned char> legacy_payload(const std::vector<TriangleHot> & hot, const std::vector<TriangleCold> & cold, const std::vector<BVHNode> & nodes, const std::vector<int> & indices) {
	std::vector<unsigned char> out;
	auto put = [&](const void * p, size_t n) { const unsigned char * b = static_cast<const unsigned char *>(p); out.insert(out.end(), b, b + n); };
	int32_t c = int32_t(hot.size());     put(&c, 4); put(hot.data(), hot.size() * sizeof(TriangleHot)); put(cold.data(), cold.size() * sizeof(TriangleCold));
	c = int32_t(nodes.size());           put(&c, 4); put(nodes.data(), nodes.size() * sizeof(BVHNode));
	c = int32_t(indices.size());         put(&c, 4); put(indices.data(), indices.size() * sizeof(int));
	return out;
}
}

void BottomLevelBVH::save_container(const std::string & name) const {
	const std::vector<unsigned char> payload = legacy_payload(file_hot, file_cold, nodes, file_indices);
	ContainerHeader h;
	memcpy(h.magic, CONTAINER_MAGIC, 8);
	h.endian_tag = CONTAINER_ENDIAN; h.version = CONTAINER_VERSION;
	h.triangle_count = uint32_t(file_hot.size()); h.node_count = uint32_t(nodes.size()); h.index_count = uint32_t(file_indices.size());
	h.crc = uint32_t(crc32(0L, payload.data(), uInt(payload.size())));

	const std::string partial = name + ".part"; // written beside, renamed when complete: a reader never sees half a file
	{
		File file(partial, "wb");
		if (!file.f) throw std::runtime_error("cannot write BVH container " + name);
		if (fwrite(&h, sizeof(h), 1, file.f) != 1 || fwrite(payload.data(), 1, payload.size(), file.f) != payload.size()) throw std::runtime_error("short write to BVH container " + name);
	}
	if (rename(partial.c_str(), name.c_str()) != 0) throw std::runtime_error("cannot rename BVH container " + name);
}

void BottomLevelBVH::load_container(const std::string & name) {
	const std::vector<unsigned char> bytes = read_whole_file(name);
	ContainerHeader h;
	if (bytes.size() < sizeof(h))                         throw std::runtime_error("truncated header in BVH container " + name);
	memcpy(&h, bytes.data(), sizeof(h));
	if (memcmp(h.magic, CONTAINER_MAGIC, 8) != 0)         throw std::runtime_error("not a BVH container (bad magic): " + name);
	if (h.endian_tag != CONTAINER_ENDIAN)                 throw std::runtime_error("BVH container written on a machine of the other endianness: " + name);
	if (h.version != CONTAINER_VERSION)                   throw std::runtime_error("unsupported BVH container version in " + name);

	const size_t expect = 12 + size_t(h.triangle_count) * (sizeof(TriangleHot) + sizeof(TriangleCold)) + size_t(h.node_count) * sizeof(BVHNode) + size_t(h.index_count) * sizeof(int);
	if (bytes.size() - sizeof(h) != expect)               throw std::runtime_error("size mismatch in BVH container " + name);
	const unsigned char * payload = bytes.data() + sizeof(h);
	if (uint32_t(crc32(0L, payload, uInt(expect))) != h.crc) throw std::runtime_error("checksum mismatch in BVH container " + name);
	parse_legacy(payload, expect, name);
}

void BottomLevelBVH::build(std::vector<TriangleHot> hot, std::vector<TriangleCold> cold) {
	file_hot  = std::move(hot);
	file_cold = std::move(cold);

	std::vector<BuildPrimitive> primitives(file_hot.size());
	for (size_t i = 0; i < file_hot.size(); i++) {
		const TriangleHot & t = file_hot[i];
		Vector3 p0(t.position_0[0], t.position_0[1], t.position_0[2]);
		Vector3 p1 = p0 + Vector3(t.position_edge_1[0], t.position_edge_1[1], t.position_edge_1[2]);
		Vector3 p2 = p0 + Vector3(t.position_edge_2[0], t.position_edge_2[1], t.position_edge_2[2]);

		AABB box = AABB::empty();
		box.expand(p0); box.expand(p1); box.expand(p2);
		box.fix_if_needed(); // Triangle.h:16-21 applies the fix twice (from_points + calc_aabb)
		box.fix_if_needed();

		primitives[i].aabb     = box;
		primitives[i].position = (p0 + p1 + p2) * (1.0f / 3.0f); // Triangle.h:23-25
	}

	SAHBuilder builder;
	builder.init(int(primitives.size()));
	int n_nodes = builder.build(primitives, nodes);
	nodes.resize(n_nodes);

	file_indices = builder.indices[0];
	flatten();
}

void BottomLevelBVH::build_sbvh(std::vector<TriangleHot> hot, std::vector<TriangleCold> cold, const std::vector<Vector3> & vertices, int threads) {
	if (vertices.size() != hot.size() * 3) throw std::runtime_error("build_sbvh needs three vertices per triangle");
	file_hot  = std::move(hot);
	file_cold = std::move(cold);

	std::vector<BuildTriangle> triangles(file_hot.size());
	for (size_t i = 0; i < triangles.size(); i++) {
		for (int k = 0; k < 3; k++) triangles[i].vertex[k] = vertices[3 * i + k];
		triangles[i].finish();
	}
	SBVHResult result = b200rt::build_sbvh(triangles, threads);
	nodes        = std::move(result.nodes);
	file_indices = std::move(result.indices);
	flatten();
}

void BottomLevelBVH::flatten() {
	triangles_hot .resize(file_indices.size());
	triangles_cold.resize(file_indices.size());
	for (size_t i = 0; i < file_indices.size(); i++) {
		triangles_hot [i] = file_hot [file_indices[i]];
		triangles_cold[i] = file_cold[file_indices[i]];
	}
	original_index = file_indices;
}

} // namespace b200rt
