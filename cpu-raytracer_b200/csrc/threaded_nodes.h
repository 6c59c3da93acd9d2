// threaded_nodes.h — host side, plain C++: lays a BLAS out in the order a traversal visits it.
// Included by b200rt.cu (b200rt_upload_static) and by the g++ harness of tests/test_threaded_nodes.py.
#pragma once
#include <string.h>
#include <algorithm>
#include <utility>
#include <vector>
#include "../../include/b200rt.h"

// Walks the nodes a traversal can reach from the root (node 0) and checks what the kernels trust: a leaf's range
// lies inside [0, leaf_limit) and its count carries no axis bits (the kernels' leaf loops run to left_or_first + count
// with the RAW count, like the reference's, BottomLevelBVH.cpp:368-376), an inner node's children exist — whatever
// its axis code — and the nodes form a tree. Unreachable slots (node 1, never written by the reference's builders:
// BottomLevelBVH.cpp:92,132) may hold anything. *depth_out = the deepest node's level (root = 0): a traversal that
// pushes both children and pops one holds at most depth + 1 entries.
static inline const char * check_tree(const b200rt_bvh_node * nodes, int node_count, int leaf_limit, int * depth_out = nullptr) {
	if (node_count <= 0) return "empty tree";
	std::vector<std::pair<int, int>> dfs(1, std::make_pair(0, 0));
	size_t visited = 0;
	int depth = 0;
	while (!dfs.empty()) {
		const int n = dfs.back().first, level = dfs.back().second; dfs.pop_back();
		if (++visited > size_t(node_count)) return "nodes do not form a tree";
		depth = std::max(depth, level);
		const int count = nodes[n].count & 0x3FFFFFFF, left = nodes[n].left_or_first;
		if (count > 0) {
			if ((unsigned(nodes[n].count) >> 30) != 0) return "leaf count carries axis bits";
			if (left < 0 || left > leaf_limit - count) return "leaf range out of bounds";
		} else {
			if (left <= 0 || left >= node_count - 1) return "child index out of bounds";
			dfs.push_back(std::make_pair(left, level + 1)); dfs.push_back(std::make_pair(left + 1, level + 1));
		}
	}
	if (depth_out) *depth_out = depth;
	return nullptr;
}

// Threaded copy of one BLAS for the any-hit (shadow ray) traversal, traverse.cuh: blas_intersect.
// The answer of an any-hit query does not depend on the order the nodes are visited in, so the
// order is fixed — right child first, the fastest of those measured — which lets the tree be laid
// out in exactly that order with a skip link per node: "hit an inner node" is "next record",
// everything else is "follow the link", and the traversal needs no stack at all.
//   record = BVHNode's 32 bytes: {min[3], max[3], skip, leaf}; skip = record that follows this
//   node's subtree (-1 = done); leaf = 0 for an inner node, else (first_triangle << 6) | count with
//   count in 1..63 — a larger leaf becomes a chain of records with the same box.
// Unreachable records of the reference's array (node 1, reserved pairs) are dropped.
//
// RT_OCTANT_THREADED (experiment, off by default): the same layout for the closest-hit rays. Their order is the
// reference's (BVHNode::should_visit_left_first: the child on the ray's side of the split axis first), which depends
// on the ray only through the signs of its direction: eight orders, eight copies
// (octant = x > 0 | (y > 0) << 1 | (z > 0) << 2). Records are visited in the same sequence, and tested against the
// same running t, as by the stack traversal. A leaf of more than 63 triangles is again a chain, but the reference
// tests the box ONCE per leaf: every record of the chain links past the whole chain (a miss on the first record skips
// it all), bit 30 of `leaf` says "hit: the leaf goes on in the next record", and the records after the first carry bit
// 31, "no box test" (t may have shrunk below the box's entry distance in between). 2^24 triangles per BLAS here.
static inline const char * build_threaded_nodes(const b200rt_bvh_node * nodes, int node_count, std::vector<b200rt_bvh_node> & out, int octant = -1) {
	const size_t base = out.size();
	const unsigned order_mask = octant < 0 ? 0u : ((octant & 4) ? 9u : 0u) | ((octant & 1) ? 2u : 0u) | ((octant & 2) ? 4u : 0u); // traverse.cuh: direction_order_mask
	auto first_child = [&](const b200rt_bvh_node & n) { // the child visited first; shadow rays: the right one
		return n.left_or_first + (octant < 0 ? 1 : 1 - int((order_mask >> (unsigned(n.count) >> 30)) & 1u));
	};
	std::vector<int> order;            // node ids in layout order: preorder, first_child first
	{
		std::vector<int> dfs(1, 0);
		while (!dfs.empty()) {
			const int n = dfs.back(); dfs.pop_back();
			order.push_back(n);
			if ((nodes[n].count & 0x3FFFFFFF) == 0) {
				const int left = nodes[n].left_or_first;
				if (left <= 0 || left + 1 >= node_count) return "BLAS child index out of bounds";
				const int near_child = first_child(nodes[n]);
				dfs.push_back(2 * left + 1 - near_child); dfs.push_back(near_child); // popped first
			}
			if (order.size() > size_t(node_count)) return "BLAS nodes do not form a tree";
		}
	}
	std::vector<int> first_record(order.size());
	std::vector<int> position_of;      // node id -> position in `order`
	{
		int max_node = 0;
		for (int n : order) max_node = std::max(max_node, n);
		position_of.assign(size_t(max_node) + 1, -1);
		for (size_t i = 0; i < order.size(); i++) position_of[order[i]] = int(i);
	}
	int next = 0;
	for (size_t i = 0; i < order.size(); i++) {
		const int count = nodes[order[i]].count & 0x3FFFFFFF;
		first_record[i] = next;
		next += count > 0 ? (count + 62) / 63 : 1;
	}
	const int total = next;
	// subtree extent: the layout is a preorder, so a subtree is a contiguous run; its end = the end of its last (left) child
	std::vector<int> subtree_end(order.size());
	for (size_t i = order.size(); i-- > 0; ) {
		const b200rt_bvh_node & n = nodes[order[i]];
		const int count = n.count & 0x3FFFFFFF;
		subtree_end[i] = count > 0 ? first_record[i] + (count + 62) / 63 : subtree_end[position_of[2 * n.left_or_first + 1 - first_child(n)]];
	}
	out.resize(base + size_t(total));
	for (size_t i = 0; i < order.size(); i++) {
		const b200rt_bvh_node & n = nodes[order[i]];
		const int count = n.count & 0x3FFFFFFF;
		const int after = subtree_end[i] == total ? -1 : subtree_end[i];
		if (count == 0) {
			b200rt_bvh_node r = n;
			r.left_or_first = after; r.count = 0;
			out[base + size_t(first_record[i])] = r;
		} else {
			if (n.left_or_first < 0 || n.left_or_first >= (1 << 25)) return "BLAS too large for the shadow-ray layout (2^25 triangles per BLAS)";
			if (octant >= 0 && n.left_or_first + count > (1 << 24)) return "BLAS too large for the octant-threaded layout (2^24 triangles per BLAS)";
			for (int k = 0, done = 0; done < count; k++, done += 63) {
				b200rt_bvh_node r = n;
				const int part = std::min(63, count - done);
				const bool more = done + part < count;
				r.left_or_first = more && octant < 0 ? first_record[i] + k + 1 : after;
				r.count = ((n.left_or_first + done) << 6) | part;
				if (octant >= 0 && more)  r.count |= 0x40000000;
				if (octant >= 0 && k > 0) r.count |= int(0x80000000u);
				out[base + size_t(first_record[i] + k)] = r;
			}
		}
	}
	return nullptr;
}

// RT_WIDE_SHADOW (experiment): the any-hit tree collapsed to four children per node, so that a shadow ray takes half
// as many dependent steps. One record = 128 bytes = 8 x float4:
//   [0..5]  min.x, min.y, min.z, max.x, max.y, max.z of the four children (one float4 each, component k = child k)
//   [6]     link of child k: first triangle of a leaf, or index of the child's own wide record
//   [7]     count of child k: > 0 leaf of that many triangles, 0 inner, < 0 empty slot
// The boolean "some triangle within max_distance is hit" is unchanged, because a skipped inner node's box contains the
// boxes that are tested below it and the slab computation is monotone in the box (traverse.cuh: aabb_intersect, regular
// rays): if the smaller box is hit, so is the larger — the set of leaves a ray reaches is the binary tree's. A child is
// only pulled up through a parent whose box really contains it (checked on the stored floats; the builder's boxes
// always do); tests/test_wide_nodes.py replays both walks on every fixture tree.
// *stack_entries = the most records a traversal can have pending. Records are appended to `out` (8 float4 each, as floats).
static inline const char * build_wide_nodes(const b200rt_bvh_node * nodes, int node_count, std::vector<float> & out, int * stack_entries, int * not_contained = nullptr) {
	struct Pending { int binary; size_t record; int pending; };   // pending: bound on the stack entries left waiting while this record is looked at
	const size_t base = out.size() / 32;
	auto is_leaf = [&](int n) { return (nodes[n].count & 0x3FFFFFFF) > 0; };
	auto area = [&](int n) {
		const float dx = nodes[n].aabb_max[0] - nodes[n].aabb_min[0], dy = nodes[n].aabb_max[1] - nodes[n].aabb_min[1], dz = nodes[n].aabb_max[2] - nodes[n].aabb_min[2];
		return dx * dy + dy * dz + dz * dx;
	};
	auto contains = [&](int outer, int inner) {
		for (int a = 0; a < 3; a++) if (!(nodes[inner].aabb_min[a] >= nodes[outer].aabb_min[a] && nodes[inner].aabb_max[a] <= nodes[outer].aabb_max[a])) return false;
		return true;
	};
	int deepest = 1, escaped = 0;
	if (!is_leaf(0)) { // the root's own box is never tested by the wide walk: its children must lie inside it
		const int l = nodes[0].left_or_first;
		if (l <= 0 || l + 1 >= node_count) return "BLAS child index out of bounds";
		if (!contains(0, l) || !contains(0, l + 1)) return "the root's box does not contain its children";
	}
	std::vector<Pending> todo;
	out.resize(out.size() + 32, 0.0f);
	todo.push_back({ 0, base, 0 });
	size_t visited = 0;
	while (!todo.empty()) {
		const Pending p = todo.back(); todo.pop_back();
		if (++visited > size_t(node_count)) return "BLAS nodes do not form a tree";
		int slots[4], used = 0;
		if (is_leaf(p.binary)) slots[used++] = p.binary;   // a tree that is one leaf: its box is the one slot
		else { slots[used++] = nodes[p.binary].left_or_first; slots[used++] = nodes[p.binary].left_or_first + 1; }
		for (;;) { // pull the children of the largest inner slot up, while there is room
			int pick = -1;
			for (int k = 0; k < used; k++) {
				if (is_leaf(slots[k])) continue;
				const int l = nodes[slots[k]].left_or_first;
				if (l <= 0 || l + 1 >= node_count) return "BLAS child index out of bounds";
				if (!contains(slots[k], l) || !contains(slots[k], l + 1)) { continue; }
				if (pick < 0 || area(slots[k]) > area(slots[pick])) pick = k;
			}
			if (pick < 0 || used == 4) break;
			const int l = nodes[slots[pick]].left_or_first;
			slots[pick] = l; slots[used++] = l + 1;
		}
		float * rec = &out[p.record * 32];
		int inner = 0;
		for (int k = 0; k < 4; k++) {
			int link = 0, count = -1;
			if (k < used) {
				const b200rt_bvh_node & n = nodes[slots[k]];
				for (int a = 0; a < 3; a++) { rec[4 * a + k] = n.aabb_min[a]; rec[4 * (3 + a) + k] = n.aabb_max[a]; }
				if (is_leaf(slots[k])) { link = n.left_or_first; count = n.count & 0x3FFFFFFF; }
				else {
					const int l = n.left_or_first;
					if (!contains(slots[k], l) || !contains(slots[k], l + 1)) escaped++;
					count = 0; link = int(out.size() / 32 - base);
					out.resize(out.size() + 32, 0.0f);
					rec = &out[p.record * 32];
					todo.push_back({ slots[k], base + size_t(link), 0 });
					inner++;
				}
			}
			memcpy(&rec[24 + k], &link, 4); memcpy(&rec[28 + k], &count, 4);
		}
		// all hit inner children are pushed, then one is popped: while it is walked, the others wait
		deepest = std::max(deepest, p.pending + inner);
		for (int i = 0; i < inner; i++) todo[todo.size() - inner + i].pending = p.pending + inner - 1;
	}
	if (stack_entries) *stack_entries = deepest + 1;
	if (not_contained) *not_contained = escaped;
	return nullptr;
}
