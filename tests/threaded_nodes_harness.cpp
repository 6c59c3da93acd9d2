// g++ harness for cpu-raytracer_b200/csrc/threaded_nodes.h (built on the fly by tests/test_threaded_nodes.py).
#include "threaded_nodes.h"

#include <cstring>

// Lays out one BLAS; octant -1 = the shadow rays' order. Returns the number of records written to `out`
// (capacity records), -1 when the builder refuses the tree, -2 when `out` is too small.
extern "C" int threaded_build(const b200rt_bvh_node * nodes, int node_count, int octant, b200rt_bvh_node * out, int capacity) {
	std::vector<b200rt_bvh_node> records;
	if (build_threaded_nodes(nodes, node_count, records, octant) != nullptr) return -1;
	if (int(records.size()) > capacity) return -2;
	memcpy(out, records.data(), records.size() * sizeof(b200rt_bvh_node));
	return int(records.size());
}

// Four-wide collapse of one BLAS (build_wide_nodes). out: capacity records of 32 floats. Returns the number of records,
// -1 when the builder refuses the tree, -2 when `out` is too small; *stack_entries / *not_contained as the builder reports them.
extern "C" int wide_build(const b200rt_bvh_node * nodes, int node_count, float * out, int capacity, int * stack_entries, int * not_contained) {
	std::vector<float> records;
	if (build_wide_nodes(nodes, node_count, records, stack_entries, not_contained) != nullptr) return -1;
	if (int(records.size() / 32) > capacity) return -2;
	memcpy(out, records.data(), records.size() * sizeof(float));
	return int(records.size() / 32);
}

// check_tree: 0 when the tree is sound (depth through *depth), else -1 and the reason in `why` (capacity bytes).
extern "C" int tree_check(const b200rt_bvh_node * nodes, int node_count, int leaf_limit, int * depth, char * why, int capacity) {
	const char * problem = check_tree(nodes, node_count, leaf_limit, depth);
	if (!problem) return 0;
	strncpy(why, problem, size_t(capacity) - 1); why[capacity - 1] = 0;
	return -1;
}
