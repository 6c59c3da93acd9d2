#include "Math.h"

namespace b200rt {

// 3x3 determinant of `m` with row `skip_r` and column `skip_c` removed (flat 4x4, index r*4+c).
static float minor3(const float * m, int skip_r, int skip_c) {
	int r[3], c[3];
	for (int i = 0, k = 0; i < 4; i++) if (i != skip_r) r[k++] = i;
	for (int i = 0, k = 0; i < 4; i++) if (i != skip_c) c[k++] = i;
	auto e = [&](int i, int j) { return m[r[i] * 4 + c[j]]; };
	return e(0, 0) * e(1, 1) * e(2, 2) - e(0, 0) * e(1, 2) * e(2, 1)
	     - e(1, 0) * e(0, 1) * e(2, 2) + e(1, 0) * e(0, 2) * e(2, 1)
	     + e(2, 0) * e(0, 1) * e(1, 2) - e(2, 0) * e(0, 2) * e(1, 1);
}

// Adjugate / determinant in FP32. The reference spells the 16 cofactors out by hand
// (Matrix4.h:88-138); this is the same six-term expansion per cofactor, written as a loop.
// A singular matrix yields the identity, as there (Matrix4.h:126-136).
Matrix4 Matrix4::invert(const Matrix4 & matrix) {
	const float * m = matrix.cells;

	float adj[16];
	for (int r = 0; r < 4; r++) {
		for (int c = 0; c < 4; c++) {
			float cofactor = minor3(m, r, c);
			adj[c * 4 + r] = ((r + c) & 1) ? -cofactor : cofactor;
		}
	}

	float det = m[0] * adj[0] + m[1] * adj[4] + m[2] * adj[8] + m[3] * adj[12];

	Matrix4 result;
	if (det != 0.0f) {
		float inv_det = 1.0f / det;
		for (int i = 0; i < 16; i++) result.cells[i] = adj[i] * inv_det;
	}
	return result;
}

} // namespace b200rt
