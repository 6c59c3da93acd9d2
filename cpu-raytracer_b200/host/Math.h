// Host-side scalar math for the b200rt host library: Vector3 / Quaternion / Matrix4 /
// Transform with the reference's conventions (Vector3.h, Quaternion.h, Matrix4.h,
// Transform.h of clayne/CPU-Raytracer). Only what Scene::update needs.
//
// Convention that matters for parity: Matrix4 is column-major, element (i, j) lives in
// cells[i + 4*j] (Matrix4.h:19-28); a position transforms as
//   x' = m(0,0) x + m(1,0) y + m(2,0) z + m(3,0)            (Matrix4.h:31-37)
// so the translation sits in cells[3], cells[7], cells[11].
#pragma once
#include <cmath>
#include <cstring>

namespace b200rt {

constexpr float PI = 3.14159265359f; // Util.h:8

inline float deg_to_rad(float angle) { return angle * PI * 0.00555555555f; } // Util.h:14

struct Vector3 {
	float x = 0.0f, y = 0.0f, z = 0.0f;

	Vector3() = default;
	explicit Vector3(float f) : x(f), y(f), z(f) { }
	Vector3(float x, float y, float z) : x(x), y(y), z(z) { }

	float   operator[](int i) const { return (&x)[i]; }
	float & operator[](int i)       { return (&x)[i]; }
};

inline Vector3 operator+(const Vector3 & a, const Vector3 & b) { return { a.x + b.x, a.y + b.y, a.z + b.z }; }
inline Vector3 operator-(const Vector3 & a, const Vector3 & b) { return { a.x - b.x, a.y - b.y, a.z - b.z }; }
inline Vector3 operator*(const Vector3 & a, float s)           { return { a.x * s, a.y * s, a.z * s }; }
inline Vector3 operator*(float s, const Vector3 & a)           { return { a.x * s, a.y * s, a.z * s }; }
inline Vector3 operator-(const Vector3 & a)                    { return { -a.x, -a.y, -a.z }; }

inline float   dot  (const Vector3 & a, const Vector3 & b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
inline Vector3 cross(const Vector3 & a, const Vector3 & b) { return { a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x }; }
inline Vector3 vmin (const Vector3 & a, const Vector3 & b) { return { a.x < b.x ? a.x : b.x, a.y < b.y ? a.y : b.y, a.z < b.z ? a.z : b.z }; }
inline Vector3 vmax (const Vector3 & a, const Vector3 & b) { return { a.x > b.x ? a.x : b.x, a.y > b.y ? a.y : b.y, a.z > b.z ? a.z : b.z }; }

inline Vector3 normalize(const Vector3 & v) { // Vector3.h:25-32: multiply by 1/length
	float inv_length = 1.0f / sqrtf(dot(v, v));
	return { v.x * inv_length, v.y * inv_length, v.z * inv_length };
}

struct Quaternion {
	float x = 0.0f, y = 0.0f, z = 0.0f, w = 1.0f;

	Quaternion() = default;
	Quaternion(float x, float y, float z, float w) : x(x), y(y), z(z), w(w) { }

	static Quaternion axis_angle(const Vector3 & axis, float angle) { // Quaternion.h:28-38
		float half = 0.5f * angle;
		float s = sinf(half);
		return { axis.x * s, axis.y * s, axis.z * s, cosf(half) };
	}

	static Quaternion normalize(const Quaternion & q) { // Quaternion.h:18-21
		float inv = 1.0f / sqrtf(q.x * q.x + q.y * q.y + q.z * q.z + q.w * q.w);
		return { q.x * inv, q.y * inv, q.z * inv, q.w * inv };
	}

	static Quaternion nlerp(const Quaternion & a, const Quaternion & b, float t) { // Quaternion.h:107-116
		float s = 1.0f - t;
		return normalize({ s * a.x + t * b.x, s * a.y + t * b.y, s * a.z + t * b.z, s * a.w + t * b.w });
	}
};

// Hamilton product, Quaternion.h:119-126
inline Quaternion operator*(const Quaternion & l, const Quaternion & r) {
	return {
		l.x * r.w + l.w * r.x + l.y * r.z - l.z * r.y,
		l.y * r.w + l.w * r.y + l.z * r.x - l.x * r.z,
		l.z * r.w + l.w * r.z + l.x * r.y - l.y * r.x,
		l.w * r.w - l.x * r.x - l.y * r.y - l.z * r.z
	};
}

// Rotate a vector, Quaternion.h:128-134:  2(q.v)q + (w^2 - q.q)v + 2w(q x v)
inline Vector3 operator*(const Quaternion & q, const Vector3 & v) {
	Vector3 u(q.x, q.y, q.z);
	return 2.0f * dot(u, v) * u + (q.w * q.w - dot(u, u)) * v + 2.0f * q.w * cross(u, v);
}

struct Matrix4 {
	float cells[16];

	Matrix4() { // identity, Matrix4.h:11-17
		memset(cells, 0, sizeof(cells));
		cells[0] = cells[5] = cells[10] = cells[15] = 1.0f;
	}

	float & operator()(int i, int j)       { return cells[i + (j << 2)]; }
	float   operator()(int i, int j) const { return cells[i + (j << 2)]; }

	static Vector3 transform_position(const Matrix4 & m, const Vector3 & p) { // Matrix4.h:31-37
		return {
			m(0, 0) * p.x + m(1, 0) * p.y + m(2, 0) * p.z + m(3, 0),
			m(0, 1) * p.x + m(1, 1) * p.y + m(2, 1) * p.z + m(3, 1),
			m(0, 2) * p.x + m(1, 2) * p.y + m(2, 2) * p.z + m(3, 2)
		};
	}

	static Vector3 transform_direction(const Matrix4 & m, const Vector3 & d) { // Matrix4.h:61-67
		return {
			m(0, 0) * d.x + m(1, 0) * d.y + m(2, 0) * d.z,
			m(0, 1) * d.x + m(1, 1) * d.y + m(2, 1) * d.z,
			m(0, 2) * d.x + m(1, 2) * d.y + m(2, 2) * d.z
		};
	}

	// General 4x4 inverse by cofactors in FP32 — the reference inverts the rigid world
	// matrix this way (Matrix4.h:88-138), so the instance inverse carries the same rounding.
	static Matrix4 invert(const Matrix4 & matrix);
};

struct Transform { // Transform.h:6-43
	Vector3    position;
	Quaternion rotation;
	Matrix4    world_matrix;

	void calc_world_matrix() {
		const Quaternion & q = rotation;
		float xx = q.x * q.x, yy = q.y * q.y, zz = q.z * q.z;
		float xz = q.x * q.z, xy = q.x * q.y, yz = q.y * q.z;
		float wx = q.w * q.x, wy = q.w * q.y, wz = q.w * q.z;

		Matrix4 & m = world_matrix;
		m(0, 0) = 1.0f - 2.0f * (yy + zz); m(0, 1) = 2.0f * (xy + wz);        m(0, 2) = 2.0f * (xz - wy);
		m(1, 0) = 2.0f * (xy - wz);        m(1, 1) = 1.0f - 2.0f * (xx + zz); m(1, 2) = 2.0f * (yz + wx);
		m(2, 0) = 2.0f * (xz + wy);        m(2, 1) = 2.0f * (yz - wx);        m(2, 2) = 1.0f - 2.0f * (xx + yy);
		m(3, 0) = position.x;              m(3, 1) = position.y;              m(3, 2) = position.z;
	}
};

struct AABB { // AABB.h:7-9
	Vector3 min, max;

	static AABB empty() { AABB b; b.min = Vector3(+INFINITY); b.max = Vector3(-INFINITY); return b; }

	void expand(const Vector3 & p) { min = vmin(min, p);     max = vmax(max, p); }
	void expand(const AABB & b)    { min = vmin(min, b.min); max = vmax(max, b.max); }

	// AABB.h:27-33: boxes thinner than 1e-3 along an axis are fattened by 5e-3 on the max side
	void fix_if_needed() {
		for (int d = 0; d < 3; d++) if (max[d] - min[d] < 0.001f) max[d] += 0.005f;
	}

	float surface_area() const { // AABB.h:35-41
		Vector3 d = max - min;
		return 2.0f * (d.x * d.y + d.y * d.z + d.z * d.x);
	}

	// World-space box of a transformed box: centre/extent form with |M| (AABB.cpp:55-73)
	static AABB transform(const AABB & box, const Matrix4 & m) {
		Vector3 center = 0.5f * (box.min + box.max);
		Vector3 extent = 0.5f * (box.max - box.min);

		Matrix4 abs_m;
		for (int i = 0; i < 16; i++) abs_m.cells[i] = fabsf(m.cells[i]);

		Vector3 c = Matrix4::transform_position (m,     center);
		Vector3 e = Matrix4::transform_direction(abs_m, extent);

		AABB result; result.min = c - e; result.max = c + e;
		return result;
	}
};

} // namespace b200rt
