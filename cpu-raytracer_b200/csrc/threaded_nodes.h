// threaded_nodes.h — host side, plain C++: lays a BLAS out in the order a traversal visits it.
// Included by b200rt.cu (b200rt_upload_static) and by the g++ harness of tests/test_threaded_nodes.py.
#pragma once
#include <algorithm>
#include <vector>
#include "../../include/b200rt.h"

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
// same running t, as by the stack traversal. A leaf of more than 63 triangles is again a chain, but the records
// after the first carry bit 31 of `leaf`: "no box test" — the reference tests the box once per leaf, and t may
// have shrunk below the box's entry distance in between.
static const char * build_threaded_nodes(const b200rt_bvh_node * nodes, int node_count, std::vector<b200rt_bvh_node> & out, int octant = -1) {
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
			for (int k = 0, done = 0; done < count; k++, done += 63) {
				b200rt_bvh_node r = n;
				const int part = std::min(63, count - done);
				r.left_or_first = done + part < count ? first_record[i] + k + 1 : after;
				r.count = ((n.left_or_first + done) << 6) | part;
				if (octant >= 0 && k > 0) r.count |= int(0x80000000u);
				out[base + size_t(first_record[i] + k)] = r;
			}
		}
	}
	return nullptr;
}
