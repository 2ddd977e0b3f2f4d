// ORACLE — TEST INFRASTRUCTURE ONLY. Not part of the shipped product path.
// CPU restatement of the reference's math layer (MWstudios/Box2D, C#).
// Parity status: UNPINNED by the reference (it ships no tests / golden vectors; no .NET
// runtime exists in this image so the C# cannot be executed). Every function cites the
// reference file:line it follows. Build with -ffp-contract=off: the reference's scalar and
// AVX code paths contain no FMA contraction.
//
// Follows: include/MathFunctions.cs:7-308, .NET BCL Math.Min/Max/Clamp semantics.
#pragma once
#include <cmath>
#include <cstdint>
#include <cfloat>

namespace b2o {

constexpr float kEps = 1.192092896e-07f;   // include/MathFunctions.cs:311
constexpr float kPi = 3.14159274f;         // MathF.PI
constexpr float kTau = 6.28318548f;        // MathF.Tau
constexpr float kHuge = 100000.0f;         // src/Constants.cs:8
constexpr float kLinearSlop = 0.005f;      // src/Constants.cs:18
constexpr float kSpeculativeDistance = 4.0f * kLinearSlop;  // src/Constants.cs:26
constexpr float kAABBMargin = 0.05f;       // src/Constants.cs:31
constexpr float kMaxRotation = 0.25f * kPi;  // src/Constants.cs:22
constexpr float kTimeToSleep = 0.5f;       // src/Constants.cs:33
constexpr int kGraphColorCount = 24;       // src/Constants.cs:13
constexpr int kDynamicColorCount = 20;     // src/Constants.cs:36
constexpr int kOverflowIndex = kGraphColorCount - 1;

// .NET Math.Max(float,float) / Math.Min(float,float): NaN-propagating, +0 > -0.
inline float netMax(float a, float b) {
  if (a != b) { if (!std::isnan(a)) return b < a ? a : b; return a; }
  return std::signbit(b) ? a : b;
}
inline float netMin(float a, float b) {
  if (a != b) { if (!std::isnan(a)) return a < b ? a : b; return a; }
  return std::signbit(a) ? a : b;
}
// .NET Math.Clamp(value,min,max)
inline float netClamp(float v, float lo, float hi) { if (v < lo) return lo; if (v > hi) return hi; return v; }
// vmaxps / vminps: second operand when unordered or equal (src/ContactSolver.cs AVX path)
inline float avxMax(float a, float b) { return a > b ? a : b; }
inline float avxMin(float a, float b) { return a < b ? a : b; }

struct Vec2 { float x, y; };
inline Vec2 operator+(Vec2 a, Vec2 b) { return {a.x + b.x, a.y + b.y}; }
inline Vec2 operator-(Vec2 a, Vec2 b) { return {a.x - b.x, a.y - b.y}; }
inline Vec2 operator-(Vec2 a) { return {-a.x, -a.y}; }
inline Vec2 operator*(float s, Vec2 v) { return {s * v.x, s * v.y}; }
inline Vec2 operator*(Vec2 v, float s) { return {s * v.x, s * v.y}; }
inline bool operator==(Vec2 a, Vec2 b) { return a.x == b.x && a.y == b.y; }
inline float Dot(Vec2 a, Vec2 b) { return a.x * b.x + a.y * b.y; }          // :15
inline float Cross(Vec2 a, Vec2 b) { return a.x * b.y - a.y * b.x; }        // :17
inline Vec2 CrossVS(Vec2 v, float s) { return {s * v.y, -s * v.x}; }        // :19
inline Vec2 CrossSV(float s, Vec2 v) { return {-s * v.y, s * v.x}; }        // :21
inline Vec2 LeftPerp(Vec2 v) { return {-v.y, v.x}; }                        // :23
inline Vec2 RightPerp(Vec2 v) { return {v.y, -v.x}; }                       // :25
inline Vec2 Lerp(Vec2 a, Vec2 b, float t) { return {a.x + (b.x - a.x) * t, a.y + (b.y - a.y) * t}; }  // :32
inline Vec2 MulAdd(Vec2 a, float s, Vec2 b) { return {a.x + s * b.x, a.y + s * b.y}; }  // :45
inline Vec2 MulSub(Vec2 a, float s, Vec2 b) { return {a.x - s * b.x, a.y - s * b.y}; }  // :47
inline Vec2 Min(Vec2 a, Vec2 b) { return {a.x < b.x ? a.x : b.x, a.y < b.y ? a.y : b.y}; }  // :51
inline Vec2 Max(Vec2 a, Vec2 b) { return {a.x > b.x ? a.x : b.x, a.y > b.y ? a.y : b.y}; }  // :53
inline float Length(Vec2 v) { return sqrtf(v.x * v.x + v.y * v.y); }        // :57
inline float LengthSquared(Vec2 v) { return v.x * v.x + v.y * v.y; }        // :78
inline float Distance(Vec2 a, Vec2 b) { float dx = b.x - a.x, dy = b.y - a.y; return sqrtf(dx * dx + dy * dy); }  // :59
inline float DistanceSquared(Vec2 a, Vec2 b) { Vec2 c = b - a; return c.x * c.x + c.y * c.y; }  // :80
inline Vec2 Normalize(Vec2 v) {                                              // :61-66
  float length = sqrtf(v.x * v.x + v.y * v.y);
  if (length < kEps) return {0, 0};
  float inv = 1 / length; return {v.x * inv, v.y * inv};
}
inline Vec2 GetLengthAndNormalize(Vec2 v, float* length) {                   // :71-76
  *length = sqrtf(v.x * v.x + v.y * v.y);
  if (*length < kEps) return {0, 0};
  float inv = 1 / *length; return {v.x * inv, v.y * inv};
}

struct Rot { float c = 1, s = 0; };
inline Rot NormalizeRot(Rot q) {                                             // :121-125
  float mag = sqrtf(q.s * q.s + q.c * q.c), inv = mag > 0 ? 1 / mag : 0;
  return {q.c * inv, q.s * inv};
}
inline Rot IntegrateRot(Rot q, float deltaAngle) {                           // :128-133
  Rot q2 = {q.c - deltaAngle * q.s, q.s + deltaAngle * q.c};
  float mag = sqrtf(q2.s * q2.s + q2.c * q2.c), inv = mag > 0 ? 1 / mag : 0;
  return {q2.c * inv, q2.s * inv};
}
inline Rot NLerp(Rot q1, Rot q2, float t) {                                  // :139-144
  Rot q = {q1.c + (q2.c - q1.c) * t, q1.s + (q2.s - q1.s) * t};
  float mag = sqrtf(q.c * q.c + q.s * q.s), inv = mag > 0 ? 1 / mag : 0;
  return {q.c * inv, q.s * inv};
}
inline Rot MulRot(Rot q, Rot r) { return {q.c * r.c - q.s * r.s, q.s * r.c + q.c * r.s}; }      // :157
inline Rot InvMulRot(Rot a, Rot b) { return {a.c * b.c + a.s * b.s, a.c * b.s - a.s * b.c}; }   // :160
inline Vec2 RotateVector(Rot q, Vec2 v) { return {q.c * v.x - q.s * v.y, q.s * v.x + q.c * v.y}; }   // :166
inline Vec2 InvRotateVector(Rot q, Vec2 v) { return {q.c * v.x + q.s * v.y, -q.s * v.x + q.c * v.y}; }  // :168
inline float RotGetAngle(Rot q) { return atan2f(q.s, q.c); }                 // :151 (CRT dependent)
inline float UnwindAngle(float radians) { return fmodf(radians + kPi, kTau) - kPi; }  // :164

struct Transform { Vec2 p{0, 0}; Rot q; };
inline Vec2 TransformPoint(const Transform& t, Vec2 p) {                     // :180
  return {t.q.c * p.x - t.q.s * p.y + t.p.x, t.q.s * p.x + t.q.c * p.y + t.p.y};
}
inline Vec2 InvTransformPoint(const Transform& t, Vec2 p) {                  // :182
  float vx = p.x - t.p.x, vy = p.y - t.p.y; return {t.q.c * vx + t.q.s * vy, -t.q.s * vx + t.q.c * vy};
}
inline Transform InvMulTransforms(const Transform& a, const Transform& b) {  // :192
  return {InvRotateVector(a.q, b.p - a.p), InvMulRot(a.q, b.q)};
}

struct Mat22 { Vec2 cx, cy; };
inline Vec2 Solve22(const Mat22& m, Vec2 b) {                                // :211-215
  float det = m.cx.x * m.cy.y - m.cy.x * m.cx.y; if (det != 0) det = 1 / det;
  return {det * (m.cy.y * b.x - m.cy.x * b.y), det * (m.cx.x * b.y - m.cx.y * b.x)};
}

struct AABB { Vec2 lo{0, 0}, hi{0, 0}; };
inline bool Contains(const AABB& a, const AABB& b) {                         // :230
  return a.lo.x <= b.lo.x && a.lo.y <= b.lo.y && b.hi.x <= a.hi.x && b.hi.y <= a.hi.y;
}
inline Vec2 Center(const AABB& a) { return {0.5f * (a.lo.x + a.hi.x), 0.5f * (a.lo.y + a.hi.y)}; }  // :232
inline AABB Union(const AABB& a, const AABB& b) {                            // :236-237
  return {{netMin(a.lo.x, b.lo.x), netMin(a.lo.y, b.lo.y)}, {netMax(a.hi.x, b.hi.x), netMax(a.hi.y, b.hi.y)}};
}
inline bool Overlaps(const AABB& a, const AABB& b) {                         // :239
  return !(b.lo.x > a.hi.x || b.lo.y > a.hi.y || a.lo.x > b.hi.x || a.lo.y > b.hi.y);
}
inline float Perimeter(const AABB& a) { return 2 * (a.hi.x - a.lo.x + a.hi.y - a.lo.y); }  // :286
inline bool Enlarge(AABB& a, const AABB& b) {                                // :287-295
  bool changed = false;
  if (b.lo.x < a.lo.x) { a.lo.x = b.lo.x; changed = true; }
  if (b.lo.y < a.lo.y) { a.lo.y = b.lo.y; changed = true; }
  if (a.hi.x < b.hi.x) { a.hi.x = b.hi.x; changed = true; }
  if (a.hi.y < b.hi.y) { a.hi.y = b.hi.y; changed = true; }
  return changed;
}

// src/Solver.cs:10-26
struct Softness {
  float biasRate = 0, massScale = 0, impulseScale = 0;
  Softness() {}
  Softness(float hertz, float zeta, float h) {
    if (hertz == 0) { biasRate = 0; massScale = 0; impulseScale = 0; return; }
    float omega = kTau * hertz;
    float a1 = 2 * zeta + h * omega;
    float a2 = h * omega * a1;
    float a3 = 1 / (1 + a2);
    biasRate = omega / a1; massScale = a2 * a3; impulseScale = a3;
  }
};

}  // namespace b2o
