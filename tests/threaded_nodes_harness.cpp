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
