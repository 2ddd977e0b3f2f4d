// libb2cuda — GJK closest points between two convex point clouds (Distance.ShapeDistance, src/Distance.cs:86-318), fresh cache only
// (cache.count = 0: the port's MakeSimplexFromCache reads the B-side vertices of a warm cache from proxyA, quirk Q19 — no caller here
// passes one). Expression order follows the source; compiled without contraction like everything else.
#pragma once
#include "b2c_world.cuh"

namespace b2c {

struct DistProxy { V2 pts[kMaxPolygonVertices]; int count; float radius; };
struct DistOut { V2 pointA, pointB, normal; float distance; int iterations; };
struct SVertex { V2 wA, wB, w; float a; int indexA, indexB; };
struct Simplex { SVertex v1, v2, v3; int count; };

B2C_HD DistProxy makeDistProxy(const Geom& g) {      // Circle/Capsule/Polygon/Segment.MakeProxy include/Collision.cs:169,340,592,722
  DistProxy p;
  p.count = g.type == kCircle ? 1 : (g.type == kPolygon ? g.count : 2);
  for (int i = 0; i < kMaxPolygonVertices; i++) p.pts[i] = i < p.count ? g.v[i] : mk(0, 0);
  p.radius = g.type == kSegment ? 0.0f : g.radius;
  return p;
}
B2C_HD int findSupport(const DistProxy& p, V2 dir) {  // :86-100
  int best = 0; float bestValue = dot(p.pts[0], dir);
  for (int i = 1; i < p.count; i++) { float v = dot(p.pts[i], dir); if (v > bestValue) { best = i; bestValue = v; } }
  return best;
}
B2C_HD V2 weight2(float a1, V2 w1, float a2, V2 w2) { return mk(a1 * w1.x + a2 * w2.x, a1 * w1.y + a2 * w2.y); }
B2C_HD V2 weight3(float a1, V2 w1, float a2, V2 w2, float a3, V2 w3) { return mk(a1 * w1.x + a2 * w2.x + a3 * w3.x, a1 * w1.y + a2 * w2.y + a3 * w3.y); }
B2C_HD void simplexWitness(V2* a, V2* b, const Simplex& s) {   // :144-153
  if (s.count == 1) { *a = s.v1.wA; *b = s.v1.wB; }
  else if (s.count == 2) { *a = weight2(s.v1.a, s.v1.wA, s.v2.a, s.v2.wA); *b = weight2(s.v1.a, s.v1.wB, s.v2.a, s.v2.wB); }
  else { *a = weight3(s.v1.a, s.v1.wA, s.v2.a, s.v2.wA, s.v3.a, s.v3.wA); *b = *a; }
}
B2C_HD V2 solveSimplex2(Simplex& s) {                 // :154-166
  V2 w1 = s.v1.w, w2 = s.v2.w, e12 = w2 - w1;
  float d12_2 = -dot(w1, e12);
  if (d12_2 <= 0.0f) { s.v1.a = 1; s.count = 1; return mk(-w1.x, -w1.y); }
  float d12_1 = dot(w2, e12);
  if (d12_1 <= 0.0f) { s.v2.a = 1; s.count = 1; s.v1 = s.v2; return mk(-w2.x, -w2.y); }
  float inv = 1.0f / (d12_1 + d12_2);
  s.v1.a = d12_1 * inv; s.v2.a = d12_2 * inv; s.count = 2;
  return crossSV(cross(w1 + w2, e12), e12);
}
B2C_HD V2 solveSimplex3(Simplex& s) {                 // :167-215
  V2 w1 = s.v1.w, w2 = s.v2.w, w3 = s.v3.w, e12 = w2 - w1;
  float w1e12 = dot(w1, e12), w2e12 = dot(w2, e12);
  float d12_1 = w2e12, d12_2 = -w1e12;
  V2 e13 = w3 - w1;
  float w1e13 = dot(w1, e13), w3e13 = dot(w3, e13);
  float d13_1 = w3e13, d13_2 = -w1e13;
  V2 e23 = w3 - w2;
  float w2e23 = dot(w2, e23), w3e23 = dot(w3, e23);
  float d23_1 = w3e23, d23_2 = -w2e23;
  float n123 = cross(e12, e13);
  float d123_1 = n123 * cross(w2, w3), d123_2 = n123 * cross(w3, w1), d123_3 = n123 * cross(w1, w2);
  if (d12_2 <= 0.0f && d13_2 <= 0.0f) { s.v1.a = 1; s.count = 1; return mk(-w1.x, -w1.y); }
  if (d12_1 > 0.0f && d12_2 > 0.0f && d123_3 <= 0.0f) {
    float inv = 1.0f / (d12_1 + d12_2);
    s.v1.a = d12_1 * inv; s.v2.a = d12_2 * inv; s.count = 2;
    return crossSV(cross(w1 + w2, e12), e12);
  }
  if (d13_1 > 0.0f && d13_2 > 0.0f && d123_2 <= 0.0f) {
    float inv = 1.0f / (d13_1 + d13_2);
    s.v1.a = d13_1 * inv; s.v3.a = d13_2 * inv; s.count = 2; s.v2 = s.v3;
    return crossSV(cross(w1 + w3, e13), e13);
  }
  if (d12_1 <= 0.0f && d23_2 <= 0.0f) { s.v1.a = 1; s.count = 1; s.v1 = s.v2; return mk(-w2.x, -w2.y); }   // a = 1 lands in the vertex that is overwritten, as written (:196)
  if (d13_1 <= 0.0f && d23_1 <= 0.0f) { s.v3.a = 1; s.count = 1; s.v1 = s.v3; return mk(-w3.x, -w3.y); }
  if (d23_1 > 0.0f && d23_2 > 0.0f && d123_1 <= 0.0f) {
    float inv = 1.0f / (d23_1 + d23_2);
    s.v2.a = d23_1 * inv; s.v3.a = d23_2 * inv; s.count = 2; s.v1 = s.v3;
    return crossSV(cross(w2 + w3, e23), e23);
  }
  float inv = 1.0f / (d123_1 + d123_2 + d123_3);
  s.v1.a = d123_1 * inv; s.v2.a = d123_2 * inv; s.v3.a = d123_3 * inv; s.count = 3;
  return mk(0, 0);
}
// SimplexCache include/Collision.cs:779-788. count == 0: a fresh start.
struct SimplexCache { int count; int indexA[3], indexB[3]; };
// ShapeDistance :219-318. A warm cache (count > 0) goes through MakeSimplexFromCache as the port writes it (:100-135, quirk Q19): the B-side
// vertex of every cached simplex point is read from proxyA (`wB = proxyA.points[indexB]`), in proxy A's frame. Where indexB is past proxy
// A's last point the C# indexer throws IndexOutOfRangeException: *throws is set and the call returns (the caller reports it; nothing here
// invents a result for a step the reference cannot finish). The cache is written back at the normal exit only (:307), as in the source.
B2C_HD DistOut shapeDistanceCached(const DistProxy& proxyA, const DistProxy& proxyB, Xf xfA, Xf xfB, bool useRadii, SimplexCache& cache, bool* throws) {
  DistOut out; out.pointA = mk(0, 0); out.pointB = mk(0, 0); out.normal = mk(0, 0); out.distance = 0.0f; out.iterations = 0;
  DistProxy localB; localB.count = proxyB.count; localB.radius = 0.0f;
  Xf rel = invMulXf(xfA, xfB);
  for (int i = 0; i < kMaxPolygonVertices; i++) localB.pts[i] = i < proxyB.count ? xfPoint(rel, proxyB.pts[i]) : mk(0, 0);
  Simplex s;
  s.v1.indexA = 0; s.v1.indexB = 0; s.v1.wA = proxyA.pts[0]; s.v1.wB = localB.pts[0]; s.v1.a = 1; s.v1.w = s.v1.wA - s.v1.wB;
  s.v2 = s.v1; s.v3 = s.v1; s.v2.a = 0; s.v3.a = 0;
  s.count = 1;
  if (cache.count > 0) {
    s.count = cache.count;
    SVertex* vs[3] = {&s.v1, &s.v2, &s.v3};
    for (int k = 0; k < cache.count; k++) {
      SVertex& v = *vs[k];
      v.indexA = cache.indexA[k]; v.indexB = cache.indexB[k]; v.a = -1.0f;
      if (v.indexA >= proxyA.count || v.indexB >= proxyA.count) { *throws = true; return out; }
      v.wA = proxyA.pts[v.indexA]; v.wB = proxyA.pts[v.indexB]; v.w = v.wA - v.wB;
    }
  }
  V2 nonUnitNormal = mk(0, 0);
  int saveA1 = 0, saveA2 = 0, saveA3 = 0, saveB1 = 0, saveB2 = 0, saveB3 = 0;
  int iteration = 0;
  while (iteration < 20) {
    int saveCount = s.count;
    if (saveCount > 0) { saveA1 = s.v1.indexA; saveB1 = s.v1.indexB; }
    if (saveCount > 1) { saveA2 = s.v2.indexA; saveB2 = s.v2.indexB; }
    if (saveCount > 2) { saveA3 = s.v3.indexA; saveB3 = s.v3.indexB; }
    V2 d;
    if (s.count == 1) d = mk(-s.v1.w.x, -s.v1.w.y);
    else if (s.count == 2) d = solveSimplex2(s);
    else d = solveSimplex3(s);
    if (s.count == 3 || dot(d, d) < FLT_EPSILON * FLT_EPSILON) {       // overlap: distance stays 0, no normal (:257-270)
      V2 la, lb; simplexWitness(&la, &lb, s);
      out.pointA = xfPoint(xfA, la); out.pointB = xfPoint(xfB, lb);
      return out;
    }
    nonUnitNormal = d;
    SVertex* vertex = s.count == 1 ? &s.v2 : (s.count == 2 ? &s.v3 : &s.v1);
    vertex->indexA = findSupport(proxyA, d); vertex->wA = proxyA.pts[vertex->indexA];
    vertex->indexB = findSupport(localB, mk(-d.x, -d.y)); vertex->wB = localB.pts[vertex->indexB];
    vertex->w = vertex->wA - vertex->wB;
    ++iteration;
    bool duplicate = false;
    if (saveCount > 0) {
      if (vertex->indexA == saveA1 && vertex->indexB == saveB1) duplicate = true;
      else if (saveCount > 1) {
        if (vertex->indexA == saveA2 && vertex->indexB == saveB2) duplicate = true;
        else if (saveCount > 2) { if (vertex->indexA == saveA3 && vertex->indexB == saveB3) duplicate = true; }
      }
    }
    if (duplicate) break;
    s.count++;
  }
  V2 normal = rotate(xfA.q, normalize(nonUnitNormal));
  V2 la, lb; simplexWitness(&la, &lb, s);
  out.normal = normal; out.distance = dist(la, lb);
  out.pointA = xfPoint(xfA, la); out.pointB = xfPoint(xfB, lb);
  out.iterations = iteration;
  cache.count = s.count;                           // MakeSimplexCache :136-143
  if (s.count > 0) { cache.indexA[0] = s.v1.indexA; cache.indexB[0] = s.v1.indexB; }
  if (s.count > 1) { cache.indexA[1] = s.v2.indexA; cache.indexB[1] = s.v2.indexB; }
  if (s.count > 2) { cache.indexA[2] = s.v3.indexA; cache.indexB[2] = s.v3.indexB; }
  if (useRadii && out.distance > 0.1f * kLinearSlop) {
    float rA = proxyA.radius, rB = proxyB.radius;
    out.distance = netMax(0.0f, out.distance - rA - rB);
    out.pointA = mulAdd(out.pointA, rA, normal);
    out.pointB = mulSub(out.pointB, rB, normal);
  }
  return out;
}

B2C_HD DistOut shapeDistance(const DistProxy& proxyA, const DistProxy& proxyB, Xf xfA, Xf xfB, bool useRadii) {   // fresh cache
  SimplexCache cache; cache.count = 0; bool throws = false;
  return shapeDistanceCached(proxyA, proxyB, xfA, xfB, useRadii, cache, &throws);
}

B2C_HD V2 geomCentroid(const Geom& g) {              // GetCentroid include/Collision.cs:78,201,(polygon centroid),(segment lerp)
  if (g.type == kCircle) return g.v[0];
  if (g.type == kPolygon) return g.centroid;
  return lerp(g.v[0], g.v[1], 0.5f);
}
B2C_HD float geomProjectedPerimeter(const Geom& g, V2 line) {   // GetProjectedPerimeter include/Collision.cs:80,203,417-429,648
  if (g.type == kCircle) return 2.0f * g.radius;
  if (g.type == kCapsule) return fabsf(dot(g.v[1] - g.v[0], line)) + 2.0f * g.radius;
  if (g.type == kSegment) return fabsf(dot(g.v[1], line) - dot(g.v[0], line));
  float value = dot(g.v[0], line), lower = value, upper = value;
  for (int i = 1; i < g.count; i++) { value = dot(g.v[i], line); lower = netMin(lower, value); upper = netMax(upper, value); }
  return upper - lower + 2.0f * g.radius;
}

}  // namespace b2c
