// Spatial-split BVH builder. Restates the reference's construction (BVHBuilders.h:48-329,
// BVHPartitions.h:116-378, BottomLevelBVH.cpp:108-147) so that the `<obj>.bvh` it produces is the
// reference's byte for byte (tests/test_sbvh.py compares against caches written by the reference's
// own binary), organised differently:
//   * a node's three sorted reference lists travel as a value (`Refs`), not as windows into three
//     shared over-allocated arrays, so nothing is written in place and nothing can overflow;
//   * a subtree is built as a relocatable `Piece` (local node numbering, local leaf offsets) and
//     stitched into its parent in the reference's depth-first order, which makes sibling subtrees
//     independent tasks: big ones go to a small thread pool, the result does not depend on timing.
// All arithmetic is FP32 in the reference's operation order; this file must be compiled without
// FMA contraction (the Makefile targets baseline x86-64: no FMA instructions exist there).
#include "BVH.h"

#include <algorithm>
#include <atomic>
#include <cmath>
#include <future>

namespace b200rt {

namespace {

constexpr int   BIN_COUNT = 256;     // SBVH_BIN_COUNT, BVHPartitions.h:174
constexpr float ALPHA     = 10e-5f;  // BVHBuilders.h:88 (written "10e-5": 1e-4)

inline bool box_is_valid(const AABB & b) { return b.max.x > b.min.x && b.max.y > b.min.y && b.max.z > b.min.z; } // AABB.h:15-17

// AABB::overlap, AABB.cpp:27-37: intersection, or the empty box when it has no volume
inline AABB box_overlap(const AABB & a, const AABB & b) {
	AABB r; r.min = vmax(a.min, b.min); r.max = vmin(a.max, b.max);
	return box_is_valid(r) ? r : AABB::empty();
}

// AABB::from_points, AABB.cpp:12-25 (includes one fix_if_needed)
inline AABB box_from_points(const Vector3 * points, int count) {
	AABB b = AABB::empty();
	for (int i = 0; i < count; i++) b.expand(points[i]);
	b.fix_if_needed();
	return b;
}

struct Refs { std::vector<int> sorted[3]; int count() const { return int(sorted[0].size()); } };

// A built subtree, relocatable: `pairs` holds the nodes BELOW the subtree's root in the order the
// reference numbers them (root's child pair first, then the left subtree's, then the right's);
// child links and leaf offsets are relative to the piece.
struct Piece {
	BVHNode              root;
	std::vector<BVHNode> pairs;
	std::vector<int>     leaf_refs[3]; // the reference's three index arrays over this subtree's leaves; [0] is the one it keeps
};

struct ObjectSplit  { int index = -1, dimension = -1; float cost = INFINITY; AABB left, right; };
struct SpatialSplit { int dimension = -1; float cost = INFINITY, plane = NAN; AABB left, right; int count_left = 0, count_right = 0; };

// BVHPartitions::partition_object, BVHPartitions.h:118-171. Prefix/suffix boxes are clipped to the
// node box at every step and the CLIPPED box is what the next step expands.
ObjectSplit partition_object(const std::vector<BuildTriangle> & tris, const Refs & refs, const AABB & node_box) {
	const int n = refs.count();
	ObjectSplit best;
	std::vector<AABB>  left(n), right(n + 1);
	std::vector<float> sah(n);

	for (int d = 0; d < 3; d++) {
		const std::vector<int> & order = refs.sorted[d];
		left[0] = AABB::empty();
		right[n] = AABB::empty();

		for (int i = 1; i < n; i++) {
			AABB b = left[i - 1];
			b.expand(tris[order[i - 1]].aabb);
			left[i] = box_overlap(b, node_box);
			sah[i] = left[i].surface_area() * float(i);
		}
		for (int i = n - 1; i > 0; i--) {
			AABB b = right[i + 1];
			b.expand(tris[order[i]].aabb);
			right[i] = box_overlap(b, node_box);

			const float cost = sah[i] + right[i].surface_area() * float(n - i);
			if (cost < best.cost) { best.cost = cost; best.index = i; best.dimension = d; best.left = left[i]; best.right = right[i]; }
		}
	}
	return best;
}

// The part of a triangle between the two planes of a bin, as a box (BVHPartitions.h:226-279).
// v[] are the vertices sorted along `d`.
inline AABB clip_to_bin(const Vector3 v[3], const AABB & triangle_box, int d, float plane_left, float plane_right, bool & outside) {
	const float vertex_min = v[0][d], vertex_max = v[2][d];
	outside = false;
	if (vertex_min >= plane_right || vertex_max <= plane_left) { outside = true; return AABB::empty(); }
	if (vertex_min >= plane_left && vertex_max <= plane_right) return triangle_box;

	Vector3 cut[6];
	int cuts = 0;
	for (int i = 0; i < 3; i++) {
		const float vi = v[i][d];
		for (int j = i + 1; j < 3; j++) {
			const float vj = v[j][d];
			const float delta = vj - vi;
			if (vi < plane_left && plane_left <= vj) {
				const float t = (plane_left - vi) / delta;
				cut[cuts++] = (1.0f - t) * v[i] + t * v[j];
			}
			if (vi < plane_right && plane_right <= vj) {
				const float t = (plane_right - vi) / delta;
				cut[cuts++] = (1.0f - t) * v[i] + t * v[j];
			}
		}
	}
	AABB box = box_from_points(cut, cuts);
	if (v[1][d] >= plane_left && v[1][d] < plane_right) box.expand(v[1]);      // the middle vertex lies in the bin
	if (cuts == 2) box.expand(vertex_max < plane_right ? v[2] : v[0]);         // so does one of the outer ones
	box.fix_if_needed();
	return box;
}

// BVHPartitions::partition_spatial, BVHPartitions.h:173-378
SpatialSplit partition_spatial(const std::vector<BuildTriangle> & tris, const Refs & refs, const AABB & bounds) {
	const int n = refs.count();
	SpatialSplit best;

	struct Bin { AABB box; int entries, exits; };
	std::vector<Bin>   bins(BIN_COUNT);
	std::vector<AABB>  left(BIN_COUNT), right(BIN_COUNT + 1);
	std::vector<int>   count_left(BIN_COUNT), count_right(BIN_COUNT + 1);
	std::vector<float> bin_sah(BIN_COUNT);

	for (int d = 0; d < 3; d++) {
		const float bounds_min  = bounds.min[d] - 0.001f;
		const float bounds_max  = bounds.max[d] + 0.001f;
		const float bounds_step = (bounds_max - bounds_min) / float(BIN_COUNT);
		const float inv_bounds_delta = 1.0f / (bounds_max - bounds_min);

		for (Bin & b : bins) { b.box = AABB::empty(); b.entries = 0; b.exits = 0; }

		for (int index : refs.sorted[d]) {
			const BuildTriangle & tri = tris[index];
			Vector3 v[3] = { tri.vertex[0], tri.vertex[1], tri.vertex[2] };
			if (v[0][d] > v[1][d]) std::swap(v[0], v[1]);
			if (v[1][d] > v[2][d]) std::swap(v[1], v[2]);
			if (v[0][d] > v[1][d]) std::swap(v[0], v[1]);

			int bin_min = int(float(BIN_COUNT) * ((tri.aabb.min[d] - bounds_min) * inv_bounds_delta));
			int bin_max = int(float(BIN_COUNT) * ((tri.aabb.max[d] - bounds_min) * inv_bounds_delta));
			bin_min = bin_min < 0 ? 0 : (bin_min > BIN_COUNT - 1 ? BIN_COUNT - 1 : bin_min);
			bin_max = bin_max < 0 ? 0 : (bin_max > BIN_COUNT - 1 ? BIN_COUNT - 1 : bin_max);

			bins[bin_min].entries++;
			bins[bin_max].exits++;

			for (int b = bin_min; b <= bin_max; b++) {
				const float plane_left  = bounds_min + float(b) * bounds_step;
				const float plane_right = plane_left + bounds_step;
				bool outside;
				const AABB piece = clip_to_bin(v, tri.aabb, d, plane_left, plane_right, outside);
				if (outside) continue;
				bins[b].box.expand(piece);
				bins[b].box = box_overlap(bins[b].box, bounds);
			}
		}

		left[0] = AABB::empty();          count_left[0] = 0;
		right[BIN_COUNT] = AABB::empty(); count_right[BIN_COUNT] = 0;

		for (int b = 1; b < BIN_COUNT; b++) {
			left[b] = left[b - 1];
			left[b].expand(bins[b - 1].box);
			count_left[b] = count_left[b - 1] + bins[b - 1].entries;
			bin_sah[b] = count_left[b] < n ? left[b].surface_area() * float(count_left[b]) : INFINITY;
		}
		for (int b = BIN_COUNT - 1; b > 0; b--) {
			right[b] = right[b + 1];
			right[b].expand(bins[b].box);
			count_right[b] = count_right[b + 1] + bins[b].exits;
			if (count_right[b] < n) bin_sah[b] += right[b].surface_area() * float(count_right[b]);
			else bin_sah[b] = INFINITY;
		}
		for (int b = 1; b < BIN_COUNT; b++) {
			if (bin_sah[b] < best.cost) {
				best.cost = bin_sah[b]; best.dimension = d;
				best.left = left[b]; best.right = right[b];
				best.count_left = count_left[b]; best.count_right = count_right[b];
				best.plane = bounds_min + bounds_step * float(b);
			}
		}
	}
	return best;
}

struct Builder {
	const std::vector<BuildTriangle> & tris;
	float            inv_root_area;
	int              max_tasks;
	std::atomic<int> tasks { 0 };

	Builder(const std::vector<BuildTriangle> & t, float inv_area, int threads) : tris(t), inv_root_area(inv_area), max_tasks(threads > 1 ? threads - 1 : 0) { }

	static void make_leaf(Piece & piece, Refs & refs, bool pair_was_reserved) {
		piece.root.left_or_first = 0;           // offset into this piece's leaf_refs
		piece.root.count = refs.count();
		// The reference reserves the child pair BEFORE it decides to stop (BVHBuilders.h:59-60, 103-108):
		// a leaf made by the SAH test leaves two unused node slots behind; the numbering depends on them.
		if (pair_was_reserved) { BVHNode unused; memset(&unused, 0, sizeof(unused)); piece.pairs.assign(2, unused); }
		for (int d = 0; d < 3; d++) piece.leaf_refs[d] = std::move(refs.sorted[d]);
	}

	// Object split, BVHBuilders.h:123-171: a reference goes left iff its centroid is below the split
	// centroid; ties are resolved by looking the reference up among the tied entries left of the split.
	void split_object(const Refs & refs, const ObjectSplit & split, Refs & left, Refs & right) const {
		const std::vector<int> & along = refs.sorted[split.dimension];
		const float plane = tris[along[split.index]].position[split.dimension];
		for (int d = 0; d < 3; d++) {
			for (int index : refs.sorted[d]) {
				const float pos = tris[index].position[split.dimension];
				bool goes_left = pos < plane;
				if (pos == plane) {
					for (int j = split.index - 1; j >= 0 && tris[along[j]].position[split.dimension] == plane; j--) {
						if (along[j] == index) { goes_left = true; break; }
					}
				}
				(goes_left ? left : right).sorted[d].push_back(index);
			}
		}
	}

	// Spatial split with reference unsplitting, BVHBuilders.h:172-297. `split.left/right` grow when a
	// straddling triangle is kept whole on one side, and later triangles are judged against the grown boxes.
	void split_spatial(const Refs & refs, SpatialSplit & split, Refs & left, Refs & right) const {
		thread_local std::vector<unsigned char> to_left, to_right;   // decision per triangle id, valid for this node's references only
		if (to_left.size() < tris.size()) { to_left.resize(tris.size()); to_right.resize(tris.size()); }

		const int d = split.dimension;
		float n_1 = float(split.count_left), n_2 = float(split.count_right);

		for (int index : refs.sorted[d]) {
			const BuildTriangle & tri = tris[index];
			bool goes_left  = tri.vertex[0][d] <  split.plane || tri.vertex[1][d] <  split.plane || tri.vertex[2][d] <  split.plane;
			bool goes_right = tri.vertex[0][d] >= split.plane || tri.vertex[1][d] >= split.plane || tri.vertex[2][d] >= split.plane;

			if (goes_left && goes_right) { // straddler
				const bool valid_left  = box_is_valid(box_overlap(tri.aabb, split.left));
				const bool valid_right = box_is_valid(box_overlap(tri.aabb, split.right));

				if (valid_left && valid_right) {
					AABB grown_left = split.left, grown_right = split.right;
					grown_left.expand(tri.aabb);
					grown_right.expand(tri.aabb);

					const float area_left = split.left.surface_area(), area_right = split.right.surface_area();
					const float c_split = area_left                   *  n_1         + area_right                   *  n_2;
					const float c_1     = grown_left.surface_area()   *  n_1         + area_right                   * (n_2 - 1.0f);
					const float c_2     = area_left                   * (n_1 - 1.0f) + grown_right.surface_area()   *  n_2;

					bool only_right = false, only_left = false;
					if (c_1 < c_split) { if (c_2 < c_1) only_right = true; else only_left = true; }
					else if (c_2 < c_split) only_right = true;

					if (only_right) { goes_left = false;  n_1 -= 1.0f; split.right.expand(tri.aabb); }
					if (only_left)  { goes_right = false; n_2 -= 1.0f; split.left.expand(tri.aabb); }
				} else {
					if (!valid_right) goes_right = false;
					if (!valid_left)  goes_left  = false;
				}
			}
			to_left[index] = goes_left; to_right[index] = goes_right;
		}
		for (int k = 0; k < 3; k++) {
			for (int index : refs.sorted[k]) {
				if (to_left[index])  left.sorted[k].push_back(index);
				if (to_right[index]) right.sorted[k].push_back(index);
			}
		}
	}

	// BVHBuilders::build_sbvh, BVHBuilders.h:48-329
	Piece build(Refs refs, const AABB & node_box) {
		Piece piece;
		memset(&piece.root, 0, sizeof(piece.root));
		for (int k = 0; k < 3; k++) { piece.root.aabb_min[k] = node_box.min[k]; piece.root.aabb_max[k] = node_box.max[k]; }

		const int n = refs.count();
		if (n < 3) { make_leaf(piece, refs, false); return piece; }

		const ObjectSplit object = partition_object(tris, refs, node_box);

		SpatialSplit spatial;
		float lambda = 0.0f;
		const AABB overlap = box_overlap(object.left, object.right);
		if (box_is_valid(overlap)) lambda = overlap.surface_area();
		if (lambda * inv_root_area > ALPHA) spatial = partition_spatial(tris, refs, node_box);

		const float parent_cost = node_box.surface_area() * float(n);
		if (parent_cost <= object.cost && parent_cost <= spatial.cost) { make_leaf(piece, refs, true); return piece; }

		Refs left, right;
		AABB box_left, box_right;
		if (object.cost <= spatial.cost) {
			split_object(refs, object, left, right);
			box_left = object.left; box_right = object.right;
		} else {
			split_spatial(refs, spatial, left, right);
			box_left = spatial.left; box_right = spatial.right;
		}
		for (int k = 0; k < 3; k++) { std::vector<int>().swap(refs.sorted[k]); }

		// children: the left subtree may run as a task; stitching below is in the reference's order either way
		Piece piece_left, piece_right;
		bool spawned = false;
		std::future<Piece> pending;
		if (max_tasks > 0 && left.count() > 4096 && right.count() > 4096) {
			int running = tasks.load();
			while (running < max_tasks && !tasks.compare_exchange_weak(running, running + 1)) { }
			if (running < max_tasks) {
				spawned = true;
				pending = std::async(std::launch::async, [this, l = std::move(left), box_left]() mutable { Piece p = build(std::move(l), box_left); tasks.fetch_sub(1); return p; });
			}
		}
		if (!spawned) piece_left = build(std::move(left), box_left);
		piece_right = build(std::move(right), box_right);
		if (spawned) piece_left = pending.get();

		// stitch: [this node's child pair][left subtree's nodes][right subtree's nodes]; leaves of the left
		// subtree come first in the index arrays (BVHBuilders.h:300-312)
		const int leaves_left = int(piece_left.leaf_refs[0].size());
		const int pairs_left  = int(piece_left.pairs.size());
		auto relocate = [](BVHNode & node, int pair_offset, int leaf_offset) {
			if ((node.count & 0x3FFFFFFF) > 0) node.left_or_first += leaf_offset;
			else if (node.count != 0)          node.left_or_first += pair_offset;
		};
		piece.root.left_or_first = 0;
		piece.root.count = (object.dimension + 1) << BVH_AXIS_SHIFT; // the object split's axis even after a spatial split (BVHBuilders.h:106)
		piece.pairs.reserve(2 + piece_left.pairs.size() + piece_right.pairs.size());
		piece.pairs.push_back(piece_left.root);  relocate(piece.pairs.back(), 2, 0);
		piece.pairs.push_back(piece_right.root); relocate(piece.pairs.back(), 2 + pairs_left, leaves_left);
		for (BVHNode node : piece_left.pairs)  { relocate(node, 2, 0); piece.pairs.push_back(node); }
		for (BVHNode node : piece_right.pairs) { relocate(node, 2 + pairs_left, leaves_left); piece.pairs.push_back(node); }
		for (int k = 0; k < 3; k++) {
			piece.leaf_refs[k] = std::move(piece_left.leaf_refs[k]);
			piece.leaf_refs[k].insert(piece.leaf_refs[k].end(), piece_right.leaf_refs[k].begin(), piece_right.leaf_refs[k].end());
		}
		return piece;
	}
};

} // namespace

void BuildTriangle::finish() {
	aabb = box_from_points(vertex, 3); // Triangle.h:16-21: from_points fixes once, calc_aabb fixes again
	aabb.fix_if_needed();
	position = (vertex[0] + vertex[1] + vertex[2]) * (1.0f / 3.0f);
}

// BottomLevelBVH::build_sbvh, BottomLevelBVH.cpp:108-147
SBVHResult build_sbvh(const std::vector<BuildTriangle> & tris, int threads) {
	SBVHResult result;
	const int n = int(tris.size());
	if (n == 0) return result;

	Refs refs;
	for (int d = 0; d < 3; d++) {
		refs.sorted[d].resize(n);
		for (int i = 0; i < n; i++) refs.sorted[d][i] = i;
		std::sort(refs.sorted[d].begin(), refs.sorted[d].end(), [&](int a, int b) { return tris[a].position[d] < tris[b].position[d]; });
	}

	AABB root_box = AABB::empty(); // BVHPartitions::calculate_bounds, BVHPartitions.h:11-24
	for (int i = 0; i < n; i++) root_box.expand(tris[refs.sorted[0][i]].aabb);
	root_box.fix_if_needed();

	Builder builder(tris, 1.0f / root_box.surface_area(), threads);
	Piece whole = builder.build(std::move(refs), root_box);

	BVHNode unused; memset(&unused, 0, sizeof(unused));
	result.nodes.reserve(2 + whole.pairs.size());
	result.nodes.push_back(whole.root);
	result.nodes.push_back(unused);
	for (const BVHNode & node : whole.pairs) result.nodes.push_back(node);
	for (BVHNode & node : result.nodes) if ((node.count & 0x3FFFFFFF) == 0 && node.count != 0) node.left_or_first += 2; // pieces number their pairs from 0; globally they start at 2
	result.indices = std::move(whole.leaf_refs[0]);
	return result;
}

} // namespace b200rt
