// ORACLE — TEST INFRASTRUCTURE ONLY (see b2o_math.h header). Parity unpinned by the reference.
// Shape geometry: AABB, mass, extent, centroid. Follows include/Collision.cs:64-92 (Circle),
// :184-228 (Capsule), :376-497 (Polygon), :632-655 (Segment); src/Geometry.cs:88-131 (MakeBox/MakeOffsetBox).
#pragma once
#include "b2o_math.h"

namespace b2o {

enum ShapeType { kCircle = 0, kCapsule = 1, kSegment = 2, kPolygon = 3, kChainSegment = 4 };  // include/Types.cs:271
constexpr int kMaxPolygonVertices = 8;

// circle: v[0]=center; capsule: v[0],v[1]=centers; segment: v[0],v[1]; polygon: count verts + normals
struct Geom {
  int type = kCircle;
  int count = 0;
  float radius = 0;
  Vec2 v[kMaxPolygonVertices] = {};
  Vec2 n[kMaxPolygonVertices] = {};
  Vec2 centroid{0, 0};
};

struct MassData { float mass = 0; Vec2 center{0, 0}; float rotationalInertia = 0; };
struct ShapeExtent { float minExtent = 0, maxExtent = 0; };

inline Geom MakeBox(float hx, float hy) {                 // src/Geometry.cs:88-99
  Geom g; g.type = kPolygon; g.count = 4; g.radius = 0;
  g.v[0] = {-hx, -hy}; g.v[1] = {hx, -hy}; g.v[2] = {hx, hy}; g.v[3] = {-hx, hy};
  g.n[0] = {0, -1}; g.n[1] = {1, 0}; g.n[2] = {0, 1}; g.n[3] = {-1, 0};
  g.centroid = {0, 0};
  return g;
}
inline Geom MakeOffsetBox(float hx, float hy, Vec2 center, Rot rotation) {   // src/Geometry.cs:113-131
  Transform xf; xf.p = center; xf.q = rotation;
  Geom g; g.type = kPolygon; g.count = 4; g.radius = 0;
  g.v[0] = TransformPoint(xf, {-hx, -hy}); g.v[1] = TransformPoint(xf, {hx, -hy});
  g.v[2] = TransformPoint(xf, {hx, hy}); g.v[3] = TransformPoint(xf, {-hx, hy});
  g.n[0] = RotateVector(xf.q, {0, -1}); g.n[1] = RotateVector(xf.q, {1, 0});
  g.n[2] = RotateVector(xf.q, {0, 1}); g.n[3] = RotateVector(xf.q, {-1, 0});
  g.centroid = xf.p;
  return g;
}
// src/Manifold.cs:8-17
inline Geom MakeCapsulePoly(Vec2 p1, Vec2 p2, float radius) {
  Geom g; g.type = kPolygon; g.count = 2; g.radius = radius;
  g.v[0] = p1; g.v[1] = p2; g.centroid = Lerp(p1, p2, 0.5f);
  Vec2 d = p2 - p1;
  Vec2 axis = Normalize(d);
  Vec2 normal = RightPerp(axis);
  g.n[0] = normal; g.n[1] = -normal;
  return g;
}

inline float GeomRadius(const Geom& g) { return g.type == kSegment ? 0.0f : g.radius; }   // IShape.GetRadius default 0 for Segment

inline AABB ComputeAABB(const Geom& g, const Transform& xf) {
  switch (g.type) {
    case kCircle: {                                        // include/Collision.cs:73-77
      Vec2 p = TransformPoint(xf, g.v[0]);
      return {{p.x - g.radius, p.y - g.radius}, {p.x + g.radius, p.y + g.radius}};
    }
    case kCapsule: {                                       // :194-200
      Vec2 v1 = TransformPoint(xf, g.v[0]), v2 = TransformPoint(xf, g.v[1]);
      Vec2 r = {g.radius, g.radius};
      return {Min(v1, v2) - r, Max(v1, v2) + r};
    }
    case kSegment: {                                       // :639-645
      Vec2 v1 = TransformPoint(xf, g.v[0]), v2 = TransformPoint(xf, g.v[1]);
      return {Min(v1, v2), Max(v1, v2)};
    }
    default: {                                             // :388-402
      Vec2 lower = TransformPoint(xf, g.v[0]), upper = lower;
      for (int i = 1; i < g.count; i++) { Vec2 v = TransformPoint(xf, g.v[i]); lower = Min(lower, v); upper = Max(upper, v); }
      Vec2 r = {g.radius, g.radius};
      return {lower - r, upper + r};
    }
  }
}

inline Vec2 GetCentroid(const Geom& g) {
  switch (g.type) {
    case kCircle: return g.v[0];                           // :79
    case kCapsule: return Lerp(g.v[0], g.v[1], 0.5f);      // :201
    case kSegment: return Lerp(g.v[0], g.v[1], 0.5f);      // :646
    default: return g.centroid;                            // :403
  }
}

inline MassData CircleMass(Vec2 center, float radius, float density) {      // :82-91
  float rr = radius * radius, mass = density * kPi * rr;
  MassData m; m.mass = mass; m.center = center; m.rotationalInertia = mass * 0.5f * rr; return m;
}
inline MassData CapsuleMass(Vec2 p1, Vec2 p2, float radius, float density) {   // :205-223
  float rr = radius * radius;
  float ll = DistanceSquared(p1, p2), length = sqrtf(ll);
  float circleMass = density * (kPi * radius * radius);
  float boxMass = density * (2 * radius * radius);
  MassData m; m.mass = circleMass + boxMass; m.center = {0.5f * (p1.x + p2.x), 0.5f * (p1.y + p2.y)};
  float lc = 4 * radius / (3 * kPi);
  float h = 0.5f * length;
  float circleInertia = circleMass * (0.5f * rr + h * h + 2 * h * lc);
  float boxInertia = boxMass * (4 * rr + ll) / 12;
  m.rotationalInertia = circleInertia + boxInertia;
  return m;
}
inline MassData ComputeMass(const Geom& g, float density) {
  if (g.type == kCircle) return CircleMass(g.v[0], g.radius, density);
  if (g.type == kCapsule) return CapsuleMass(g.v[0], g.v[1], g.radius, density);
  if (g.type != kPolygon) return MassData();              // Segment: IShape default (src/Shape.cs) -> zero mass
  // include/Collision.cs:431-484
  if (g.count == 1) return CircleMass(g.v[0], g.radius, density);
  if (g.count == 2) return CapsuleMass(g.v[0], g.v[1], g.radius, density);
  Vec2 vertices[kMaxPolygonVertices] = {};
  if (g.radius > 0) {
    float sqrt2 = 1.4142135623730950488016887242097f;
    for (int i = 0, j = g.count - 1; i < g.count; j = i++) {
      Vec2 n1 = g.n[j], n2 = g.n[i];
      Vec2 mid = Normalize(n1 + n2);
      vertices[i] = MulAdd(g.v[i], sqrt2 * g.radius, mid);
    }
  } else for (int i = 0; i < g.count; i++) vertices[i] = g.v[i];
  Vec2 center = {0, 0};
  float area = 0, rotationalInertia = 0;
  Vec2 r = vertices[0];
  const float inv3 = 1 / 3.0f;
  for (int i = 1; i < g.count - 1; i++) {
    Vec2 e1 = vertices[i] - r, e2 = vertices[i + 1] - r;
    float D = Cross(e1, e2);
    float triangleArea = 0.5f * D;
    area += triangleArea;
    center = MulAdd(center, triangleArea * inv3, e1 + e2);
    float intx2 = e1.x * e1.x + e2.x * e1.x + e2.x * e2.x;
    float inty2 = e1.y * e1.y + e2.y * e1.y + e2.y * e2.y;
    rotationalInertia += 0.25f * inv3 * D * (intx2 + inty2);
  }
  float invArea = 1 / area;
  center.x *= invArea; center.y *= invArea;
  MassData m;
  m.mass = density * area;
  m.center = r + center;
  m.rotationalInertia = density * rotationalInertia - density * area * Dot(center, center);
  return m;
}

inline ShapeExtent ComputeExtent(const Geom& g, Vec2 localCenter) {
  ShapeExtent e;
  switch (g.type) {
    case kCircle: e.minExtent = g.radius; e.maxExtent = Distance(g.v[0], localCenter) + g.radius; break;   // :92
    case kCapsule:                                                                                        // :224-228
      e.minExtent = g.radius;
      e.maxExtent = sqrtf(netMax(DistanceSquared(g.v[0], localCenter), DistanceSquared(g.v[1], localCenter))) + g.radius;
      break;
    case kSegment:                                                                                        // :649-653
      e.minExtent = 0;
      e.maxExtent = sqrtf(netMax(DistanceSquared(g.v[0], localCenter), DistanceSquared(g.v[1], localCenter)));
      break;
    default: {                                                                                            // :485-496
      float minExtent = kHuge, maxExtentSqr = 0;
      for (int i = 0; i < g.count; i++) {
        Vec2 v = g.v[i];
        float planeOffset = Dot(g.n[i], v - g.centroid);
        minExtent = netMin(minExtent, planeOffset);
        maxExtentSqr = netMax(maxExtentSqr, DistanceSquared(v, localCenter));
      }
      e.minExtent = minExtent + g.radius; e.maxExtent = sqrtf(maxExtentSqr) + g.radius;
    }
  }
  return e;
}

}  // namespace b2o
