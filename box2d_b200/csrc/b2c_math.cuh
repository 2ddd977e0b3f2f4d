// libb2cuda — shared host/device math for the rigid-body step (fp32, no FMA contraction: build with
// -fmad=false and -Xcompiler -ffp-contract=off). Operation order is normative: every expression follows
// the association order of the reference's C# (include/MathFunctions.cs:7-308), because the parity bar
// is bit-exactness against its AVX/scalar paths, which contain no fused multiply-adds.
#pragma once
#include <stdint.h>
#include <math.h>
#include <float.h>

#if defined(__CUDACC__)
#define B2C_HD __host__ __device__ __forceinline__
#else
#define B2C_HD inline
#endif

namespace b2c {

// src/Constants.cs:8-39, include/MathFunctions.cs:311
constexpr float kEps = 1.192092896e-07f;
constexpr float kPi = 3.14159274f;
constexpr float kTau = 6.28318548f;
constexpr float kHuge = 100000.0f;
constexpr float kLinearSlop = 0.005f;
constexpr float kSpeculativeDistance = 4.0f * kLinearSlop;
constexpr float kAABBMargin = 0.05f;
constexpr float kMaxRotation = 0.25f * kPi;
constexpr int kGraphColorCount = 24;
constexpr int kIslandMaxBodies = 192;   // island solver: body rows staged in shared memory (b2c_islands.cuh)
constexpr int kMaxBigQueries = 1024;    // wide broadphase queries answered by a leaf sweep per step (more fall back to a one-thread traversal)
constexpr int kBigQueryVisits = 256;
constexpr int kWideQueryCells = 256;    // a query box covering this many average leaf cells of the dynamic tree is wide without traversing   // node visits after which a one-thread tree query is reclassified as wide
constexpr int kDynamicColorCount = 20;
constexpr int kOverflowIndex = kGraphColorCount - 1;
constexpr int kMaxPolygonVertices = 8;

enum BodyType { kStaticBody = 0, kKinematicBody = 1, kDynamicBody = 2 };
enum ShapeType { kCircle = 0, kCapsule = 1, kSegment = 2, kPolygon = 3, kChainSegment = 4 };
enum BodyFlags : uint32_t {     // src/Body.cs:7-34
  kLockLinearX = 1, kLockLinearY = 2, kLockAngularZ = 4, kIsFast = 8, kIsBullet = 0x10, kIsSpeedCapped = 0x20,
  kHadTimeOfImpact = 0x40, kAllowFastRotation = 0x80, kEnlargeBounds = 0x100, kDynamicFlag = 0x200
};

struct V2 { float x, y; };
struct Rot { float c, s; };
struct Xf { V2 p; Rot q; };
struct Box { V2 lo, hi; };      // AABB

// .NET Math.Max/Min(float,float): NaN-propagating, +0 > -0 (scalar C# paths)
B2C_HD float netMax(float a, float b) {
  if (a != b) { if (!(a != a)) return b < a ? a : b; return a; }
  return signbit(b) ? a : b;
}
B2C_HD float netMin(float a, float b) {
  if (a != b) { if (!(a != a)) return a < b ? a : b; return a; }
  return signbit(a) ? a : b;
}
B2C_HD float netClamp(float v, float lo, float hi) { if (v < lo) return lo; if (v > hi) return hi; return v; }
// vmaxps/vminps semantics of the AVX contact solver: second operand when unordered or equal
B2C_HD float avxMax(float a, float b) { return a > b ? a : b; }
B2C_HD float avxMin(float a, float b) { return a < b ? a : b; }

B2C_HD V2 mk(float x, float y) { V2 v; v.x = x; v.y = y; return v; }
B2C_HD V2 operator+(V2 a, V2 b) { return mk(a.x + b.x, a.y + b.y); }
B2C_HD V2 operator-(V2 a, V2 b) { return mk(a.x - b.x, a.y - b.y); }
B2C_HD V2 operator-(V2 a) { return mk(-a.x, -a.y); }
B2C_HD V2 operator*(float s, V2 v) { return mk(s * v.x, s * v.y); }
B2C_HD float dot(V2 a, V2 b) { return a.x * b.x + a.y * b.y; }
B2C_HD float cross(V2 a, V2 b) { return a.x * b.y - a.y * b.x; }
B2C_HD V2 crossSV(float s, V2 v) { return mk(-s * v.y, s * v.x); }
B2C_HD V2 leftPerp(V2 v) { return mk(-v.y, v.x); }
B2C_HD V2 rightPerp(V2 v) { return mk(v.y, -v.x); }
B2C_HD V2 lerp(V2 a, V2 b, float t) { return mk(a.x + (b.x - a.x) * t, a.y + (b.y - a.y) * t); }
B2C_HD V2 mulAdd(V2 a, float s, V2 b) { return mk(a.x + s * b.x, a.y + s * b.y); }
B2C_HD V2 mulSub(V2 a, float s, V2 b) { return mk(a.x - s * b.x, a.y - s * b.y); }
B2C_HD V2 vmin(V2 a, V2 b) { return mk(a.x < b.x ? a.x : b.x, a.y < b.y ? a.y : b.y); }
B2C_HD V2 vmax(V2 a, V2 b) { return mk(a.x > b.x ? a.x : b.x, a.y > b.y ? a.y : b.y); }
B2C_HD float length(V2 v) { return sqrtf(v.x * v.x + v.y * v.y); }
B2C_HD float lengthSq(V2 v) { return v.x * v.x + v.y * v.y; }
B2C_HD float distSq(V2 a, V2 b) { V2 c = b - a; return c.x * c.x + c.y * c.y; }
B2C_HD float dist(V2 a, V2 b) { float dx = b.x - a.x, dy = b.y - a.y; return sqrtf(dx * dx + dy * dy); }
B2C_HD V2 normalize(V2 v) {
  float len = sqrtf(v.x * v.x + v.y * v.y);
  if (len < kEps) return mk(0.0f, 0.0f);
  float inv = 1.0f / len; return mk(v.x * inv, v.y * inv);
}
B2C_HD V2 lengthAndNormalize(V2 v, float* len) {
  *len = sqrtf(v.x * v.x + v.y * v.y);
  if (*len < kEps) return mk(0.0f, 0.0f);
  float inv = 1.0f / *len; return mk(v.x * inv, v.y * inv);
}
B2C_HD Rot mkRot(float c, float s) { Rot q; q.c = c; q.s = s; return q; }
B2C_HD Rot normalizeRot(Rot q) {
  float mag = sqrtf(q.s * q.s + q.c * q.c), inv = mag > 0.0f ? 1.0f / mag : 0.0f;
  return mkRot(q.c * inv, q.s * inv);
}
B2C_HD Rot nlerpRot(Rot q1, Rot q2, float t) {       // Rotation.NLerp include/MathFunctions.cs:139-144
  Rot q = mkRot(q1.c + (q2.c - q1.c) * t, q1.s + (q2.s - q1.s) * t);
  float mag = sqrtf(q.c * q.c + q.s * q.s), inv = mag > 0.0f ? 1.0f / mag : 0.0f;
  return mkRot(q.c * inv, q.s * inv);
}
B2C_HD Rot integrateRot(Rot q, float deltaAngle) {
  Rot q2 = mkRot(q.c - deltaAngle * q.s, q.s + deltaAngle * q.c);
  float mag = sqrtf(q2.s * q2.s + q2.c * q2.c), inv = mag > 0.0f ? 1.0f / mag : 0.0f;
  return mkRot(q2.c * inv, q2.s * inv);
}
B2C_HD Rot mulRot(Rot q, Rot r) { return mkRot(q.c * r.c - q.s * r.s, q.s * r.c + q.c * r.s); }
B2C_HD Rot invMulRot(Rot a, Rot b) { return mkRot(a.c * b.c + a.s * b.s, a.c * b.s - a.s * b.c); }
B2C_HD V2 rotate(Rot q, V2 v) { return mk(q.c * v.x - q.s * v.y, q.s * v.x + q.c * v.y); }
B2C_HD V2 invRotate(Rot q, V2 v) { return mk(q.c * v.x + q.s * v.y, -q.s * v.x + q.c * v.y); }
B2C_HD Xf mkXf(V2 p, Rot q) { Xf t; t.p = p; t.q = q; return t; }
B2C_HD V2 xfPoint(Xf t, V2 p) { return mk(t.q.c * p.x - t.q.s * p.y + t.p.x, t.q.s * p.x + t.q.c * p.y + t.p.y); }
B2C_HD Xf invMulXf(Xf a, Xf b) { return mkXf(invRotate(a.q, b.p - a.p), invMulRot(a.q, b.q)); }

B2C_HD bool boxContains(Box a, Box b) { return a.lo.x <= b.lo.x && a.lo.y <= b.lo.y && b.hi.x <= a.hi.x && b.hi.y <= a.hi.y; }
B2C_HD bool boxOverlaps(Box a, Box b) { return !(b.lo.x > a.hi.x || b.lo.y > a.hi.y || a.lo.x > b.hi.x || a.lo.y > b.hi.y); }
B2C_HD V2 boxCenter(Box a) { return mk(0.5f * (a.lo.x + a.hi.x), 0.5f * (a.lo.y + a.hi.y)); }
B2C_HD Box boxUnion(Box a, Box b) {
  Box r; r.lo = mk(netMin(a.lo.x, b.lo.x), netMin(a.lo.y, b.lo.y)); r.hi = mk(netMax(a.hi.x, b.hi.x), netMax(a.hi.y, b.hi.y)); return r;
}
B2C_HD float boxPerimeter(Box a) { return 2.0f * (a.hi.x - a.lo.x + a.hi.y - a.lo.y); }
B2C_HD Box boxGrow(Box a, float m) { Box r; r.lo = mk(a.lo.x - m, a.lo.y - m); r.hi = mk(a.hi.x + m, a.hi.y + m); return r; }

// src/Solver.cs:10-26
struct Softness { float biasRate, massScale, impulseScale; };
B2C_HD Softness makeSoft(float hertz, float zeta, float h) {
  Softness s;
  if (hertz == 0.0f) { s.biasRate = 0.0f; s.massScale = 0.0f; s.impulseScale = 0.0f; return s; }
  float omega = kTau * hertz;
  float a1 = 2.0f * zeta + h * omega;
  float a2 = h * omega * a1;
  float a3 = 1.0f / (1.0f + a2);
  s.biasRate = omega / a1; s.massScale = a2 * a3; s.impulseScale = a3;
  return s;
}

// Shape geometry as stored on the device (include/Collision.cs:65,184,376,632).
// circle: v[0]=center; capsule/segment: v[0],v[1]; polygon: count vertices + normals.
struct Geom {
  int type, count;
  float radius;
  float pad;
  V2 v[kMaxPolygonVertices];
  V2 n[kMaxPolygonVertices];
  V2 centroid;
};

B2C_HD Box geomAABB(const Geom& g, Xf xf) {        // include/Collision.cs:73,194,388,639
  Box r;
  if (g.type == kCircle) {
    V2 p = xfPoint(xf, g.v[0]);
    r.lo = mk(p.x - g.radius, p.y - g.radius); r.hi = mk(p.x + g.radius, p.y + g.radius);
  } else if (g.type == kCapsule) {
    V2 v1 = xfPoint(xf, g.v[0]), v2 = xfPoint(xf, g.v[1]);
    V2 rr = mk(g.radius, g.radius);
    r.lo = vmin(v1, v2) - rr; r.hi = vmax(v1, v2) + rr;
  } else if (g.type == kSegment) {
    V2 v1 = xfPoint(xf, g.v[0]), v2 = xfPoint(xf, g.v[1]);
    r.lo = vmin(v1, v2); r.hi = vmax(v1, v2);
  } else {
    V2 lower = xfPoint(xf, g.v[0]), upper = lower;
    for (int i = 1; i < g.count; i++) { V2 v = xfPoint(xf, g.v[i]); lower = vmin(lower, v); upper = vmax(upper, v); }
    V2 rr = mk(g.radius, g.radius);
    r.lo = lower - rr; r.hi = upper + rr;
  }
  return r;
}

}  // namespace b2c
