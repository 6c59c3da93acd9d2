#include "Math.h"

namespace b200rt {

// Signed 3x3 cofactor of `m` with row `skip_r` and column `skip_c` removed (flat 4x4, index r*4+c).
// A negative cofactor is the sum of the NEGATED terms, as the reference writes it
// (Matrix4.h:92-93 "-a*b*c + a*d*e ..."), not the negated sum: the two differ in the sign of a
// zero result (-(+0) = -0, but -0 + 0 = +0), and the inverse's zeros reach the device.
static float cofactor3(const float * m, int skip_r, int skip_c) {
	int r[3], c[3];
	for (int i = 0, k = 0; i < 4; i++) if (i != skip_r) r[k++] = i;
	for (int i = 0, k = 0; i < 4; i++) if (i != skip_c) c[k++] = i;
	auto e = [&](int i, int j) { return m[r[i] * 4 + c[j]]; };
	if (((skip_r + skip_c) & 1) == 0)
		return e(0, 0) * e(1, 1) * e(2, 2) - e(0, 0) * e(1, 2) * e(2, 1)
		     - e(1, 0) * e(0, 1) * e(2, 2) + e(1, 0) * e(0, 2) * e(2, 1)
		     + e(2, 0) * e(0, 1) * e(1, 2) - e(2, 0) * e(0, 2) * e(1, 1);
	return -e(0, 0) * e(1, 1) * e(2, 2) + e(0, 0) * e(1, 2) * e(2, 1)
	     +  e(1, 0) * e(0, 1) * e(2, 2) - e(1, 0) * e(0, 2) * e(2, 1)
	     -  e(2, 0) * e(0, 1) * e(1, 2) + e(2, 0) * e(0, 2) * e(1, 1);
}

// Adjugate / determinant in FP32. The reference spells the 16 cofactors out by hand
// (Matrix4.h:88-138); this is the same six-term expansion per cofactor, written as a loop.
// A singular matrix yields the identity, as there (Matrix4.h:126-136).
Matrix4 Matrix4::invert(const Matrix4 & matrix) {
	const float * m = matrix.cells;

	float adj[16];
	for (int r = 0; r < 4; r++) {
		for (int c = 0; c < 4; c++) {
			adj[c * 4 + r] = cofactor3(m, r, c);
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
