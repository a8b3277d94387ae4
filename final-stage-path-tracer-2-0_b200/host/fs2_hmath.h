// fs2_hmath.h — host-side float arithmetic used while loading scenes and building the
// reference's octrees.  The flattened scene must be byte-identical to what the reference
// builds, so each helper performs the reference's operations in the reference's order
// (compile with -ffp-contract=off; x86-64 baseline has no FMA):
//   approximate sqrt/normalize  reference source/math/scalar.h:233-245, vector3.h:166-177
//   cross / dot / rotate        reference source/math/vector3.h:179-187,291-293,315-333
//   planes                      reference source/math/plane.h:68-102, normal.h:67-79
//   bounds growth               reference source/math/volume.cpp:119-126,162-228
//   matrix4 TRS                 reference source/math/matrix4.h:248-301,533-568
//   xorshift64* stream          reference source/math/random.cpp:6-49
#ifndef FS2_HMATH_H_
#define FS2_HMATH_H_

#include <math.h>
#include <stdint.h>
#include <string.h>

namespace fs2h {

const float kPi = 3.14159262f;
const float kEpsilon = 1.0e-5f;

struct Vec3 {
  float x = 0, y = 0, z = 0;
  Vec3() {}
  Vec3(float a, float b, float c) : x(a), y(b), z(c) {}
  float operator[](int i) const { return i == 0 ? x : (i == 1 ? y : z); }
};
struct Vec4 {
  float x = 0, y = 0, z = 0, w = 0;
};
typedef Vec4 Plane;

inline Vec3 operator+(Vec3 a, Vec3 b) { return Vec3(a.x + b.x, a.y + b.y, a.z + b.z); }
inline Vec3 operator-(Vec3 a, Vec3 b) { return Vec3(a.x - b.x, a.y - b.y, a.z - b.z); }
inline Vec3 operator*(Vec3 a, float s) { return Vec3(a.x * s, a.y * s, a.z * s); }
inline float dot(Vec3 a, Vec3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
inline Vec3 cross(Vec3 a, Vec3 b) {
  return Vec3((a.y * b.z) - (a.z * b.y), (a.z * b.x) - (a.x * b.z), (a.x * b.y) - (a.y * b.x));
}
inline float inv_sqrt(float f) {
  float half = 0.5f * f;
  int32_t i;
  memcpy(&i, &f, 4);
  i = 0x5f3759df - (i >> 1);
  memcpy(&f, &i, 4);
  return f * (1.5f - half * f * f);
}
inline float approx_sqrt(float f) { return 1.0f / inv_sqrt(f); }
inline float length(Vec3 a) { return approx_sqrt(dot(a, a)); }
inline Vec3 normalize(Vec3 a) {
  float l = length(a);
  if (l == 0.0) l = 1.0f;
  return Vec3(a.x / l, a.y / l, a.z / l);
}
inline bool near_zero(float a) { return a > (0.0f - kEpsilon) && a < (0.0f + kEpsilon); }

// vector3::rotate(angle, axis): axis used as given
inline Vec3 rotate(Vec3 p, float angle, Vec3 a) {
  float c = (float)cos(angle);
  float s = (float)sin(angle);
  Vec3 o;
  o.x = (c + (1 - c) * a.x * a.x) * p.x;
  o.x += ((1 - c) * a.x * a.y - a.z * s) * p.y;
  o.x += ((1 - c) * a.x * a.z + a.y * s) * p.z;
  o.y = ((1 - c) * a.x * a.y + a.z * s) * p.x;
  o.y += (c + (1 - c) * a.y * a.y) * p.y;
  o.y += ((1 - c) * a.y * a.z - a.x * s) * p.z;
  o.z = ((1 - c) * a.x * a.z - a.y * s) * p.x;
  o.z += ((1 - c) * a.y * a.z + a.x * s) * p.y;
  o.z += (c + (1 - c) * a.z * a.z) * p.z;
  return o;
}

inline Plane plane_from(Vec3 n, Vec3 p) {
  Plane o;
  o.x = n.x;
  o.y = n.y;
  o.z = n.z;
  o.w = -dot(n, p);
  return o;
}
inline Vec3 face_normal(Vec3 a, Vec3 b, Vec3 c) { return normalize(cross(b - a, c - a)); }
inline Plane plane_from(Vec3 p1, Vec3 p2, Vec3 p3) { return plane_from(face_normal(p1, p2, p3), p1); }
inline float plane_distance(const Plane& pl, Vec3 p) {
  return (float)pl.x * p.x + pl.y * p.y + pl.z * p.z + pl.w;
}

// bounds with the reference's "first point initialises" rule
struct Box {
  uint32_t count = 0;
  Vec3 mn, mx;
  void clear() {
    count = 0;
    mn = Vec3();
    mx = Vec3();
  }
  void add(Vec3 p) {
    if (count == 0) {
      mn = p;
      mx = p;
      count = 1;
      return;
    }
    count++;
    if (p.x > mx.x) mx.x = p.x;
    if (p.y > mx.y) mx.y = p.y;
    if (p.z > mx.z) mx.z = p.z;
    if (p.x < mn.x) mn.x = p.x;
    if (p.y < mn.y) mn.y = p.y;
    if (p.z < mn.z) mn.z = p.z;
  }
  void add(const Box& b) {
    if (count == 0) {
      mn = b.mn;
      mx = b.mx;
      count = 1;
      return;
    }
    count++;
    if (b.mx.x > mx.x) mx.x = b.mx.x;
    if (b.mx.y > mx.y) mx.y = b.mx.y;
    if (b.mx.z > mx.z) mx.z = b.mx.z;
    if (b.mn.x < mn.x) mn.x = b.mn.x;
    if (b.mn.y < mn.y) mn.y = b.mn.y;
    if (b.mn.z < mn.z) mn.z = b.mn.z;
  }
  Vec3 center() const { return (mn + mx) * 0.5f; }
  float volume() const { return (mx.x - mn.x) * (mx.y - mn.y) * (mx.z - mn.z); }
};

// column-major 4x4 (m[col*4+row])
struct Mat4 {
  float m[16];
  static Mat4 identity() {
    Mat4 o;
    memset(o.m, 0, sizeof o.m);
    o.m[0] = o.m[5] = o.m[10] = o.m[15] = 1;
    return o;
  }
  static Mat4 translation(float tx, float ty, float tz) {
    Mat4 o = identity();
    o.m[12] = tx;
    o.m[13] = ty;
    o.m[14] = tz;
    return o;
  }
  static Mat4 scale(float sx, float sy, float sz) {
    Mat4 o = identity();
    o.m[0] = sx;
    o.m[5] = sy;
    o.m[10] = sz;
    return o;
  }
  static Mat4 rotation(float rad, Vec3 axis) {
    float c = cos(rad);
    float s = sin(rad);
    float t = 1.0f - c;
    Vec3 a = normalize(axis);
    float x = a.x, y = a.y, z = a.z;
    Mat4 o;
    o.m[0] = t * x * x + c;
    o.m[4] = t * x * y - s * z;
    o.m[8] = t * x * z + s * y;
    o.m[12] = 0.0f;
    o.m[1] = t * x * y + s * z;
    o.m[5] = t * y * y + c;
    o.m[9] = t * y * z - s * x;
    o.m[13] = 0.0f;
    o.m[2] = t * x * z - s * y;
    o.m[6] = t * y * z + s * x;
    o.m[10] = t * z * z + c;
    o.m[14] = 0.0f;
    o.m[3] = 0.0f;
    o.m[7] = 0.0f;
    o.m[11] = 0.0f;
    o.m[15] = 1.0f;
    return o;
  }
  Mat4 operator*(const Mat4& r) const {
    Mat4 o;
    for (int k = 0; k < 16; k++) {
      int row = k % 4, col = k / 4;
      o.m[k] = 0.0;
      for (int i = 0; i < 4; i++) o.m[k] += m[row + (i * 4)] * r.m[(col * 4) + i];
    }
    return o;
  }
  // matrix * (v, w = 1), xyz of the result
  Vec3 transform_point(Vec3 v) const {
    float w = 1.0f;
    return Vec3(v.x * m[0] + v.y * m[4] + v.z * m[8] + w * m[12],
                v.x * m[1] + v.y * m[5] + v.z * m[9] + w * m[13],
                v.x * m[2] + v.y * m[6] + v.z * m[10] + w * m[14]);
  }
};

// xorshift64*, the reference's process-wide main-thread stream (default state 521288629)
struct Xorshift {
  uint64_t state = 521288629ull;
  uint64_t next() {
    uint64_t x = state;
    x ^= x << 12;
    x ^= x >> 25;
    x ^= x << 27;
    state = x;
    return x * 0x2545F4914F6CDD1Dull;
  }
  float uniform() { return ((float)(next() % 0x7FFFFFFF)) / 0x7FFFFFFF; }
  float range(float lo, float hi) { return lo + uniform() * (hi - lo); }
};

}  // namespace fs2h
#endif
