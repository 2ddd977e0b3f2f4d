// ORACLE — TEST INFRASTRUCTURE ONLY (see b2o_math.h header). Parity unpinned by the reference.
// Narrowphase manifold functions with the port's quirks preserved.
// Follows src/Manifold.cs:22-568, src/Distance.cs:20-72, include/Collision.cs:916-957.
#pragma once
#include "b2o_geometry.h"

namespace b2o {

struct ManifoldPoint {          // include/Collision.cs:916-945
  Vec2 point{0, 0}, anchorA{0, 0}, anchorB{0, 0};
  float separation = 0, normalImpulse = 0, tangentImpulse = 0, totalNormalImpulse = 0, normalVelocity = 0;
  uint16_t id = 0;
  bool persisted = false;
};
struct Manifold {               // :947-958
  Vec2 normal{0, 0};
  float rollingImpulse = 0;
  ManifoldPoint point0, point1;
  int pointCount = 0;
};

inline uint16_t MakeId(int a, int b) { return (uint16_t)((a << 8) | (uint8_t)b); }   // src/Manifold.cs:22

struct SegmentDistanceResult { Vec2 closest1, closest2; float fraction1 = 0, fraction2 = 0, distanceSquared = 0; };
inline SegmentDistanceResult SegmentDistance(Vec2 p1, Vec2 q1, Vec2 p2, Vec2 q2) {   // src/Distance.cs:20-72
  SegmentDistanceResult result;
  Vec2 d1 = q1 - p1, d2 = q2 - p2, r = p1 - p2;
  float dd1 = Dot(d1, d1), dd2 = Dot(d2, d2);
  float rd1 = Dot(r, d1), rd2 = Dot(r, d2);
  float epsSqr = kEps * kEps;
  if (dd1 < epsSqr || dd2 < epsSqr) {
    if (dd1 >= epsSqr) { result.fraction1 = netClamp(-rd1 / dd1, 0, 1); result.fraction2 = 0; }
    else if (dd2 >= epsSqr) { result.fraction1 = 0; result.fraction2 = netClamp(rd2 / dd2, 0, 1); }
    else { result.fraction1 = 0; result.fraction2 = 0; }
  } else {
    float d12 = Dot(d1, d2);
    float denominator = dd1 * dd2 - d12 * d12;
    float f1 = 0;
    if (denominator != 0) f1 = netClamp((d12 * rd2 - rd1 * dd2) / denominator, 0, 1);
    float f2 = (d12 * f1 + rd2) / dd2;
    if (f2 < 0) { f2 = 0; f1 = netClamp(-rd1 / dd1, 0, 1); }
    else if (f2 > 1) { f2 = 1; f1 = netClamp((d12 - rd1) / dd1, 0, 1); }
    result.fraction1 = f1; result.fraction2 = f2;
  }
  result.closest1 = MulAdd(p1, result.fraction1, d1);
  result.closest2 = MulAdd(p2, result.fraction2, d2);
  result.distanceSquared = DistanceSquared(result.closest1, result.closest2);
  return result;
}

// src/Manifold.cs:24-46
inline Manifold CollideCircles(Vec2 centerA, float radiusA, Transform xfA, Vec2 centerB, float radiusB, Transform xfB) {
  Manifold manifold;
  Transform xf = InvMulTransforms(xfA, xfB);
  Vec2 pointA = centerA, pointB = TransformPoint(xf, centerB);
  float distance;
  Vec2 normal = GetLengthAndNormalize(pointB - pointA, &distance);
  float separation = distance - radiusA - radiusB;
  if (separation > kSpeculativeDistance) return manifold;
  Vec2 cA = MulAdd(pointA, radiusA, normal), cB = MulAdd(pointB, -radiusB, normal);
  Vec2 contactPointA = Lerp(cA, cB, 0.5f);
  manifold.normal = RotateVector(xfA.q, normal);
  ManifoldPoint& mp = manifold.point0;
  mp.anchorA = RotateVector(xfA.q, contactPointA);
  mp.anchorB = mp.anchorA + (xfA.p - xfB.p);
  mp.point = mp.anchorA + xfA.p;
  mp.separation = separation; mp.id = 0;
  manifold.pointCount = 1;
  return manifold;
}
// src/Manifold.cs:49-83 (Q6: mp.point = xfA.p - anchorA)
inline Manifold CollideCapsuleAndCircle(Vec2 p1, Vec2 p2, float radiusA, Transform xfA, Vec2 centerB, float radiusB, Transform xfB) {
  Manifold manifold;
  Transform xf = InvMulTransforms(xfA, xfB);
  Vec2 pB = TransformPoint(xf, centerB);
  Vec2 e = p2 - p1;
  Vec2 pA;
  float s1 = Dot(pB - p1, e), s2 = Dot(p2 - pB, e);
  if (s1 < 0) pA = p1;
  else if (s2 < 0) pA = p2;
  else { float s = s1 / Dot(e, e); pA = MulAdd(p1, s, e); }
  float distance;
  Vec2 normal = GetLengthAndNormalize(pB - pA, &distance);
  float separation = distance - radiusA - radiusB;
  if (separation > kSpeculativeDistance) return manifold;
  Vec2 cA = MulAdd(pA, radiusA, normal), cB = MulAdd(pB, -radiusB, normal);
  Vec2 contactPointA = Lerp(cA, cB, 0.5f);
  manifold.normal = RotateVector(xfA.q, normal);
  ManifoldPoint& mp = manifold.point0;
  mp.anchorA = RotateVector(xfA.q, contactPointA);
  mp.anchorB = mp.anchorA + (xfA.p - xfB.p);
  mp.point = xfA.p - mp.anchorA;
  mp.separation = separation; mp.id = 0;
  manifold.pointCount = 1;
  return manifold;
}
// src/Manifold.cs:91-165 (Q6: `mp.point = xfA.p = mp.anchorA` in the v1 region)
inline Manifold CollidePolygonAndCircle(const Geom& polygonA, Transform xfA, Vec2 centerB, float radiusB, Transform xfB) {
  Manifold manifold;
  Transform xf = InvMulTransforms(xfA, xfB);
  Vec2 center = TransformPoint(xf, centerB);
  float radiusA = polygonA.radius;
  float radius = radiusA + radiusB;
  int normalIndex = 0;
  float separation = -FLT_MAX;
  int vertexCount = polygonA.count;
  const Vec2* vertices = polygonA.v; const Vec2* normals = polygonA.n;
  for (int i = 0; i < vertexCount; i++) {
    float s = Dot(normals[i], center - vertices[i]);
    if (s > separation) { separation = s; normalIndex = i; }
  }
  if (separation > radius + kSpeculativeDistance) return manifold;
  int vertIndex1 = normalIndex;
  int vertIndex2 = vertIndex1 + 1 < vertexCount ? vertIndex1 + 1 : 0;
  Vec2 v1 = vertices[vertIndex1], v2 = vertices[vertIndex2];
  float u1 = Dot(center - v1, v2 - v1), u2 = Dot(center - v2, v1 - v2);
  if (u1 < 0 && separation > kEps) {
    Vec2 normal = Normalize(center - v1);
    separation = Dot(center - v1, normal);
    if (separation > radius + kSpeculativeDistance) return manifold;
    Vec2 cA = MulAdd(v1, radiusA, normal), cB = MulSub(center, radiusB, normal);
    Vec2 contactPointA = Lerp(cA, cB, 0.5f);
    manifold.normal = RotateVector(xfA.q, normal);
    ManifoldPoint& mp = manifold.point0;
    mp.anchorA = RotateVector(xfA.q, contactPointA);
    mp.anchorB = mp.anchorA + (xfA.p - xfB.p);
    mp.point = mp.anchorA;   // quirk: xfA.p assigned, not added
    mp.separation = Dot(cB - cA, normal); mp.id = 0;
    manifold.pointCount = 1;
  } else if (u2 < 0 && separation > kEps) {
    Vec2 normal = Normalize(center - v2);
    separation = Dot(center - v2, normal);
    if (separation > radius + kSpeculativeDistance) return manifold;
    Vec2 cA = MulAdd(v2, radiusA, normal), cB = MulSub(center, radiusB, normal);
    Vec2 contactPointA = Lerp(cA, cB, 0.5f);
    manifold.normal = RotateVector(xfA.q, normal);
    ManifoldPoint& mp = manifold.point0;
    mp.anchorA = RotateVector(xfA.q, contactPointA);
    mp.anchorB = mp.anchorA + (xfA.p - xfB.p);
    mp.point = xfA.p + mp.anchorA;
    mp.separation = Dot(cB - cA, normal); mp.id = 0;
    manifold.pointCount = 1;
  } else {
    Vec2 normal = normals[normalIndex];
    manifold.normal = RotateVector(xfA.q, normal);
    Vec2 cA = MulAdd(center, radiusA - Dot(center - v1, normal), normal);
    Vec2 cB = MulSub(center, radiusB, normal);
    Vec2 contactPointA = Lerp(cA, cB, 0.5f);
    ManifoldPoint& mp = manifold.point0;
    mp.anchorA = RotateVector(xfA.q, contactPointA);
    mp.anchorB = mp.anchorA + (xfA.p - xfB.p);
    mp.point = xfA.p + mp.anchorA;
    mp.separation = separation - radius; mp.id = 0;
    manifold.pointCount = 1;
  }
  return manifold;
}
// src/Manifold.cs:167-315 (Q5 quirks preserved verbatim)
inline Manifold CollideCapsules(Vec2 a1, Vec2 a2, float radiusA, Transform xfA, Vec2 b1, Vec2 b2, float radiusB, Transform xfB) {
  Vec2 origin = a1;
  Transform sfA; sfA.p = xfA.p + RotateVector(xfA.q, origin); sfA.q = xfA.q;
  Transform xf = InvMulTransforms(sfA, xfB);
  Vec2 p1 = {0, 0};
  Vec2 q1 = a2 - origin;
  Vec2 p2 = TransformPoint(xf, b1), q2 = TransformPoint(xf, b2);
  Vec2 d1 = q1 - p1, d2 = q2 - p2;
  float dd1 = Dot(d1, d1), dd2 = Dot(d2, d2);
  const float epsSqr = kEps * kEps;
  Vec2 r = p1 - p2;
  float rd1 = Dot(r, d1), rd2 = Dot(r, d2), d12 = Dot(d1, d2);
  float denom = dd1 * dd2 - d12 * d12;
  float f1 = 0;
  if (denom != 0) f1 = netClamp((d12 * rd2 - rd1 * dd2) / denom, 0, 1);
  float f2 = (d12 * f1 + rd2) / dd2;
  if (f2 < 0) { f2 = 0; f1 = netClamp(-rd1 / dd1, 0, 1); }
  else if (f2 > 1) { f2 = 1; f1 = netClamp((d12 - rd1) / dd1, 0, 1); }
  Vec2 closest1 = MulAdd(p1, f1, d1), closest2 = MulAdd(p2, f2, d2);
  float distanceSquared = DistanceSquared(closest1, closest2);
  Manifold manifold;
  float radius = radiusA + radiusB;
  float maxDistance = radius + kSpeculativeDistance;
  if (distanceSquared > maxDistance * maxDistance) return manifold;
  float distance = sqrtf(distanceSquared);
  float length1, length2;
  Vec2 u1 = GetLengthAndNormalize(d1, &length1), u2 = GetLengthAndNormalize(d2, &length2);
  float fp2 = Dot(p2 - p1, u1), fq2 = Dot(q2 - p1, u1);
  bool outsideA = (fp2 < 0 && fq2 < 0) || (fp2 >= length1 && fq2 >= length1);
  float fp1 = Dot(p1 - p2, u2), fq1 = Dot(q1 - p2, u2);
  bool outsideB = (fp1 < 0 && fq1 < 0) || (fp1 >= length2 && fq1 >= length2);
  if (!outsideA && !outsideB) {
    Vec2 normalA; float separationA;
    {
      normalA = LeftPerp(u1);
      float ss1 = Dot(p2 - p1, normalA), ss2 = Dot(q2 - p1, normalA);
      float s1p = ss1 < ss2 ? ss1 : ss2;
      float s1n = -ss1 < -ss2 ? -ss1 : -ss2;
      if (s1p > s1n) separationA = s1p; else { separationA = s1n; normalA = -normalA; }
    }
    Vec2 normalB; float separationB;
    {
      normalB = LeftPerp(u2);
      float ss1 = Dot(p1 - p2, normalB), ss2 = Dot(q1 - p2, normalB);
      float s1p = ss1 < ss2 ? ss1 : ss2;
      float s1n = -ss1 < -ss2 ? -ss1 : -ss2;
      if (s1p > s1n) separationB = s1p; else { separationB = s1n; normalB = -normalB; }
    }
    if (separationA + 0.1f * kLinearSlop >= separationB) {
      manifold.normal = normalA;
      Vec2 cp = p2, cq = q2;
      if (fp2 < 0 && fq2 > 0) cp = Lerp(p2, q2, -fp2 / (fq2 - fp2));
      else if (fq2 < 0 && fp2 > 0) cq = Lerp(q2, p2, -fq2 / (fq2 - fp2));
      if (fp2 > length1 && fq2 < length1) cp = Lerp(q2, p2, (fp2 - length1) / (fp2 - fq2));
      else if (fq2 > length1 && fp2 < length1) cq = Lerp(q2, p2, (fq2 - length1) / (fq2 - fp2));
      float sp = Dot(cp - p1, normalA), sq = Dot(cq - p1, normalA);
      if (sp <= distance + kLinearSlop || sq <= distance + kLinearSlop) {
        manifold.point0.anchorA = MulAdd(cp, 0.5f * (radiusA - radiusB - sp), normalA);
        manifold.point0.separation = sp - radius;
        manifold.point0.id = MakeId(0, 0);
        manifold.point1.anchorA = MulAdd(cq, 0.5f * (radiusA - radiusB - sq), normalA);
        manifold.point1.separation = sq - radius;
        manifold.point1.id = MakeId(0, 1);
        manifold.pointCount = 2;
      }
    } else {
      manifold.normal = -normalB;
      Vec2 cp = p1, cq = q1;
      if (fp1 < 0 && fq1 > 0) cp = Lerp(p1, q1, -fp1 / (fq1 - fp1));
      else if (fq1 < 0 && fp1 > 0) cq = Lerp(q1, p1, -fq1 / (fp1 - fq1));
      if (fp1 > length2 && fq1 > length2) cp = Lerp(p1, q1, (fp1 - length2) / (fp1 - fq1));
      else if (fq1 > length2 && fp1 > length2) cq = Lerp(q1, p1, (fq1 - length2) / (fq1 - fp1));
      float sp = Dot(cp - p2, normalB), sq = Dot(cq - p2, normalB);
      if (sp <= distance + kLinearSlop || sq <= distance + kLinearSlop) {
        manifold.point0.anchorA = MulAdd(cp, 0.5f * (radiusB - radiusA - sp), normalB);
        manifold.point0.separation = sp - radius;
        manifold.point0.id = MakeId(0, 0);
        manifold.point1.anchorA = MulAdd(cq, 0.5f * (radiusB - radiusA - sp), normalB);
        manifold.point1.separation = sp - radius;
        manifold.point1.id = MakeId(1, 0);
        manifold.pointCount = 2;
      }
    }
  }
  if (manifold.pointCount == 0) {
    Vec2 normal = closest2 - closest1;
    if (Dot(normal, normal) > epsSqr) normal = Normalize(normal); else normal = LeftPerp(u1);
    Vec2 c1 = MulAdd(closest1, radiusA, normal), c2 = MulAdd(closest2, -radiusB, normal);
    int i1 = f1 == 0 ? 0 : 1, i2 = f2 == 0 ? 0 : 1;
    manifold.normal = normal;
    manifold.point0.anchorA = Lerp(c1, c2, 0.5f);
    manifold.point0.separation = sqrtf(distanceSquared) - radius;
    manifold.point0.id = MakeId(i1, i2);
    manifold.pointCount = 1;
  }
  manifold.normal = RotateVector(xfA.q, manifold.normal);
  if (manifold.pointCount > 0) {
    ManifoldPoint& mp = manifold.point0;
    mp.anchorA = RotateVector(xfA.q, mp.anchorA + origin);
    mp.anchorB = mp.anchorA + (xfA.p - xfB.p);
    mp.point = xfA.p + mp.anchorA;
  }
  if (manifold.pointCount > 1) {
    ManifoldPoint& mp = manifold.point1;
    mp.anchorA = RotateVector(xfA.q, mp.anchorA + origin);
    mp.anchorB = mp.anchorA + (xfA.p - xfB.p);
    mp.point = xfA.p + mp.anchorA;
  }
  return manifold;
}

struct LocalPoly { int count; float radius; Vec2 v[kMaxPolygonVertices], n[kMaxPolygonVertices]; };

// src/Manifold.cs:329-388
inline Manifold ClipPolygons(const LocalPoly& polyA, const LocalPoly& polyB, int edgeA, int edgeB, bool flip) {
  Manifold manifold;
  const LocalPoly* poly1 = &polyA; const LocalPoly* poly2 = &polyB;
  int i11 = edgeA, i12 = edgeA + 1 < polyA.count ? edgeA + 1 : 0;
  int i21 = edgeB, i22 = edgeB + 1 < polyB.count ? edgeB + 1 : 0;
  if (flip) { poly1 = &polyB; poly2 = &polyA; int t11 = i11, t12 = i12; i11 = i21; i12 = i22; i21 = t11; i22 = t12; }
  Vec2 normal = poly1->n[i11];
  Vec2 v11 = poly1->v[i11], v12 = poly1->v[i12];
  Vec2 v21 = poly2->v[i21], v22 = poly2->v[i22];
  Vec2 tangent = CrossSV(1, normal);
  float lower1 = 0;
  float upper1 = Dot(v12 - v11, tangent);
  float upper2 = Dot(v21 - v11, tangent);
  float lower2 = Dot(v22 - v11, tangent);
  if (upper2 < lower1 || upper1 < lower2) return manifold;
  Vec2 vLower = (lower2 < lower1 && upper2 - lower2 > kEps) ? Lerp(v22, v21, (lower1 - lower2) / (upper2 - lower2)) : v22;
  Vec2 vUpper = (upper2 > upper1 && upper2 - lower2 > kEps) ? Lerp(v22, v21, (upper1 - lower2) / (upper2 - lower2)) : v21;
  float separationLower = Dot(vLower - v11, normal);
  float separationUpper = Dot(vUpper - v11, normal);
  float r1 = poly1->radius, r2 = poly2->radius;
  vLower = MulAdd(vLower, 0.5f * (r1 - r2 - separationLower), normal);
  vUpper = MulAdd(vUpper, 0.5f * (r1 - r2 - separationUpper), normal);
  float radius = r1 + r2;
  if (!flip) {
    manifold.normal = normal;
    manifold.point0.anchorA = vLower; manifold.point0.separation = separationLower - radius; manifold.point0.id = MakeId(i11, i22);
    manifold.point1.anchorA = vUpper; manifold.point1.separation = separationUpper - radius; manifold.point1.id = MakeId(i12, i21);
    manifold.pointCount = 2;
  } else {
    manifold.normal = -normal;
    manifold.point0.anchorA = vUpper; manifold.point0.separation = separationUpper - radius; manifold.point0.id = MakeId(i21, i12);
    manifold.point1.anchorA = vLower; manifold.point1.separation = separationLower - radius; manifold.point1.id = MakeId(i22, i11);
    manifold.pointCount = 2;
  }
  return manifold;
}
// src/Manifold.cs:389-408
inline float FindMaxSeparation(int* edgeIndex, const LocalPoly& poly1, const LocalPoly& poly2) {
  int bestIndex = 0;
  float maxSeparation = -FLT_MAX;
  for (int i = 0; i < poly1.count; i++) {
    Vec2 n = poly1.n[i], v1 = poly1.v[i];
    float si = FLT_MAX;
    for (int j = 0; j < poly2.count; j++) { float sij = Dot(n, poly2.v[j] - v1); if (sij < si) si = sij; }
    if (si > maxSeparation) { maxSeparation = si; bestIndex = i; }
  }
  *edgeIndex = bestIndex;
  return maxSeparation;
}
// src/Manifold.cs:410-562
inline Manifold CollidePolygons(const Geom& polygonA, Transform xfA, const Geom& polygonB, Transform xfB) {
  Vec2 origin = polygonA.v[0];
  Transform sfA; sfA.p = xfA.p + RotateVector(xfA.q, origin); sfA.q = xfA.q;
  Transform xf = InvMulTransforms(sfA, xfB);
  LocalPoly localPolyA; localPolyA.radius = polygonA.radius; localPolyA.count = polygonA.count;
  localPolyA.v[0] = {0, 0}; localPolyA.n[0] = polygonA.n[0];
  for (int i = 1; i < localPolyA.count; i++) { localPolyA.v[i] = polygonA.v[i] - origin; localPolyA.n[i] = polygonA.n[i]; }
  LocalPoly localPolyB; localPolyB.radius = polygonB.radius; localPolyB.count = polygonB.count;
  for (int i = 0; i < localPolyB.count; i++) { localPolyB.v[i] = TransformPoint(xf, polygonB.v[i]); localPolyB.n[i] = RotateVector(xf.q, polygonB.n[i]); }
  int edgeA, edgeB;
  float separationA = FindMaxSeparation(&edgeA, localPolyA, localPolyB);
  float separationB = FindMaxSeparation(&edgeB, localPolyB, localPolyA);
  float radius = localPolyA.radius + localPolyB.radius;
  if (separationA > kSpeculativeDistance + radius || separationB > kSpeculativeDistance + radius) return Manifold();
  bool flip;
  if (separationA >= separationB) {
    flip = false;
    Vec2 searchDirection = localPolyA.n[edgeA];
    edgeB = 0; float minDot = FLT_MAX;
    for (int i = 0; i < localPolyB.count; i++) { float dot = Dot(searchDirection, localPolyB.n[i]); if (dot < minDot) { minDot = dot; edgeB = i; } }
  } else {
    flip = true;
    Vec2 searchDirection = localPolyB.n[edgeB];
    edgeA = 0; float minDot = FLT_MAX;
    for (int i = 0; i < localPolyA.count; i++) { float dot = Dot(searchDirection, localPolyA.n[i]); if (dot < minDot) { minDot = dot; edgeA = i; } }
  }
  Manifold manifold;
  if (separationA > 0.1f * kLinearSlop || separationB > 0.1f * kLinearSlop) {
    int i11 = edgeA, i12 = edgeA + 1 < localPolyA.count ? edgeA + 1 : 0;
    int i21 = edgeB, i22 = edgeB + 1 < localPolyB.count ? edgeB + 1 : 0;
    Vec2 v11 = localPolyA.v[i11], v12 = localPolyA.v[i12];
    Vec2 v21 = localPolyB.v[i21], v22 = localPolyB.v[i22];
    SegmentDistanceResult result = SegmentDistance(v11, v12, v21, v22);
    float distance = sqrtf(result.distanceSquared);
    float separation = distance - radius;
    if (distance - radius > kSpeculativeDistance) return manifold;
    manifold = ClipPolygons(localPolyA, localPolyB, edgeA, edgeB, flip);
    float minSeparation = FLT_MAX;
    if (manifold.pointCount > 0) minSeparation = netMin(minSeparation, manifold.point0.separation);
    if (manifold.pointCount > 1) minSeparation = netMin(minSeparation, manifold.point1.separation);
    if (separation + 0.1f * kLinearSlop < minSeparation) {
      Vec2 va, vb; int ia = -1, ib = -1;
      if (result.fraction1 == 0 && result.fraction2 == 0) { va = v11; vb = v21; ia = i11; ib = i21; }
      else if (result.fraction1 == 0 && result.fraction2 == 1) { va = v11; vb = v22; ia = i11; ib = i22; }
      else if (result.fraction1 == 1 && result.fraction2 == 0) { va = v12; vb = v21; ia = i12; ib = i21; }
      else if (result.fraction1 == 1 && result.fraction2 == 1) { va = v12; vb = v22; ia = i12; ib = i22; }
      if (ia >= 0) {
        Vec2 normal = vb - va;
        float invDistance = 1 / distance;
        normal = normal * invDistance;
        Vec2 c1 = MulAdd(va, localPolyA.radius, normal), c2 = MulAdd(vb, -localPolyB.radius, normal);
        manifold.normal = normal;
        manifold.point0.anchorA = Lerp(c1, c2, 0.5f);
        manifold.point0.separation = distance - radius;
        manifold.point0.id = MakeId(ia, ib);
        manifold.pointCount = 1;
      }
    }
  } else manifold = ClipPolygons(localPolyA, localPolyB, edgeA, edgeB, flip);
  if (manifold.pointCount > 0) {
    manifold.normal = RotateVector(xfA.q, manifold.normal);
    ManifoldPoint& mp = manifold.point0;
    mp.anchorA = RotateVector(xfA.q, mp.anchorA + origin);
    mp.anchorB = mp.anchorA + (xfA.p - xfB.p);
    mp.point = xfA.p + mp.anchorA;
    if (manifold.pointCount > 1) {
      ManifoldPoint& mq = manifold.point1;
      mq.anchorA = RotateVector(xfA.q, mq.anchorA + origin);
      mq.anchorB = mq.anchorA + (xfA.p - xfB.p);
      mq.point = xfA.p + mq.anchorA;
    }
  }
  return manifold;
}

// src/Contact.cs:420-474 dispatch (primary ordering: Capsule-Circle, Polygon-Circle, Polygon-Capsule, Segment-*)
inline bool PairSupported(int tA, int tB) { return tA != kChainSegment && tB != kChainSegment && !(tA == kSegment && tB == kSegment); }
inline bool PairPrimary(int tA, int tB) {
  // registered (type1,type2) with primary=true
  if (tA == tB) return true;
  if (tA == kCapsule && tB == kCircle) return true;
  if (tA == kPolygon && (tB == kCircle || tB == kCapsule)) return true;
  if (tA == kSegment && (tB == kCircle || tB == kCapsule || tB == kPolygon)) return true;
  return false;
}
inline Manifold ComputeManifold(const Geom& a, Transform xfA, const Geom& b, Transform xfB) {
  switch (a.type) {
    case kCircle: return CollideCircles(a.v[0], a.radius, xfA, b.v[0], b.radius, xfB);
    case kCapsule:
      if (b.type == kCircle) return CollideCapsuleAndCircle(a.v[0], a.v[1], a.radius, xfA, b.v[0], b.radius, xfB);
      return CollideCapsules(a.v[0], a.v[1], a.radius, xfA, b.v[0], b.v[1], b.radius, xfB);
    case kSegment:
      if (b.type == kCircle) return CollideCapsuleAndCircle(a.v[0], a.v[1], 0, xfA, b.v[0], b.radius, xfB);       // :85-89
      if (b.type == kCapsule) return CollideCapsules(a.v[0], a.v[1], 0, xfA, b.v[0], b.v[1], b.radius, xfB);      // :317-321
      return CollidePolygons(MakeCapsulePoly(a.v[0], a.v[1], 0), xfA, b, xfB);                                    // :564-568
    default:
      if (b.type == kCircle) return CollidePolygonAndCircle(a, xfA, b.v[0], b.radius, xfB);
      if (b.type == kCapsule) return CollidePolygons(a, xfA, MakeCapsulePoly(b.v[0], b.v[1], b.radius), xfB);     // :323-327
      return CollidePolygons(a, xfA, b, xfB);
  }
}

}  // namespace b2o
