// libb2cuda — narrowphase manifold functions as device code, one thread per shape pair.
// Semantics follow src/Manifold.cs:22-568 and src/Distance.cs:20-72 of the reference, INCLUDING the port's
// deviations from upstream Box2D, since parity is judged against this C# port. Operation order
// is normative (no FMA); early-outs return pointCount == 0.
#pragma once
#include "b2c_math.cuh"

namespace b2c {

struct MPoint {                 // include/Collision.cs:916-945
  V2 point, anchorA, anchorB;
  float separation, normalImpulse, tangentImpulse, totalNormalImpulse, normalVelocity;
  uint32_t id, persisted;
};
struct Manifold {               // include/Collision.cs:947-958
  V2 normal; float rollingImpulse; int pointCount;
  MPoint p0, p1;
};

__device__ __forceinline__ void clearPoint(MPoint& p) {
  p.point = mk(0, 0); p.anchorA = mk(0, 0); p.anchorB = mk(0, 0);
  p.separation = 0; p.normalImpulse = 0; p.tangentImpulse = 0; p.totalNormalImpulse = 0; p.normalVelocity = 0; p.id = 0; p.persisted = 0;
}
__device__ __forceinline__ Manifold emptyManifold() {
  Manifold m; m.normal = mk(0, 0); m.rollingImpulse = 0; m.pointCount = 0; clearPoint(m.p0); clearPoint(m.p1); return m;
}
__device__ __forceinline__ uint32_t makeId(int a, int b) { return (uint32_t)(((a << 8) | (b & 0xff)) & 0xffff); }   // src/Manifold.cs:22

// src/Manifold.cs:24-46
__device__ inline Manifold collideCircles(V2 centerA, float radiusA, Xf xfA, V2 centerB, float radiusB, Xf xfB) {
  Manifold m = emptyManifold();
  Xf xf = invMulXf(xfA, xfB);
  V2 pointA = centerA, pointB = xfPoint(xf, centerB);
  float distance;
  V2 normal = lengthAndNormalize(pointB - pointA, &distance);
  float separation = distance - radiusA - radiusB;
  if (separation > kSpeculativeDistance) return m;
  V2 cA = mulAdd(pointA, radiusA, normal), cB = mulAdd(pointB, -radiusB, normal);
  V2 contact = lerp(cA, cB, 0.5f);
  m.normal = rotate(xfA.q, normal);
  m.p0.anchorA = rotate(xfA.q, contact);
  m.p0.anchorB = m.p0.anchorA + (xfA.p - xfB.p);
  m.p0.point = m.p0.anchorA + xfA.p;
  m.p0.separation = separation; m.p0.id = 0;
  m.pointCount = 1;
  return m;
}

// src/Manifold.cs:49-83 (also segment-vs-circle :85-89 with radiusA = 0). Q6: point = xfA.p - anchorA.
__device__ inline Manifold collideCapsuleAndCircle(V2 p1, V2 p2, float radiusA, Xf xfA, V2 centerB, float radiusB, Xf xfB) {
  Manifold m = emptyManifold();
  Xf xf = invMulXf(xfA, xfB);
  V2 pB = xfPoint(xf, centerB);
  V2 e = p2 - p1;
  V2 pA;
  float s1 = dot(pB - p1, e), s2 = dot(p2 - pB, e);
  if (s1 < 0.0f) pA = p1;
  else if (s2 < 0.0f) pA = p2;
  else { float s = s1 / dot(e, e); pA = mulAdd(p1, s, e); }
  float distance;
  V2 normal = lengthAndNormalize(pB - pA, &distance);
  float separation = distance - radiusA - radiusB;
  if (separation > kSpeculativeDistance) return m;
  V2 cA = mulAdd(pA, radiusA, normal), cB = mulAdd(pB, -radiusB, normal);
  V2 contact = lerp(cA, cB, 0.5f);
  m.normal = rotate(xfA.q, normal);
  m.p0.anchorA = rotate(xfA.q, contact);
  m.p0.anchorB = m.p0.anchorA + (xfA.p - xfB.p);
  m.p0.point = xfA.p - m.p0.anchorA;
  m.p0.separation = separation; m.p0.id = 0;
  m.pointCount = 1;
  return m;
}

// src/Manifold.cs:91-165. Q6: in the v1 vertex region `mp.point = xfA.p = mp.anchorA`.
__device__ inline Manifold collidePolygonAndCircle(const Geom& polyA, Xf xfA, V2 centerB, float radiusB, Xf xfB) {
  Manifold m = emptyManifold();
  Xf xf = invMulXf(xfA, xfB);
  V2 center = xfPoint(xf, centerB);
  float radiusA = polyA.radius;
  float radius = radiusA + radiusB;
  int normalIndex = 0;
  float separation = -FLT_MAX;
  int count = polyA.count;
  for (int i = 0; i < count; i++) {
    float s = dot(polyA.n[i], center - polyA.v[i]);
    if (s > separation) { separation = s; normalIndex = i; }
  }
  if (separation > radius + kSpeculativeDistance) return m;
  int i1 = normalIndex, i2 = i1 + 1 < count ? i1 + 1 : 0;
  V2 v1 = polyA.v[i1], v2 = polyA.v[i2];
  float u1 = dot(center - v1, v2 - v1), u2 = dot(center - v2, v1 - v2);
  if (u1 < 0.0f && separation > kEps) {
    V2 normal = normalize(center - v1);
    separation = dot(center - v1, normal);
    if (separation > radius + kSpeculativeDistance) return m;
    V2 cA = mulAdd(v1, radiusA, normal), cB = mulSub(center, radiusB, normal);
    V2 contact = lerp(cA, cB, 0.5f);
    m.normal = rotate(xfA.q, normal);
    m.p0.anchorA = rotate(xfA.q, contact);
    m.p0.anchorB = m.p0.anchorA + (xfA.p - xfB.p);
    m.p0.point = m.p0.anchorA;
    m.p0.separation = dot(cB - cA, normal); m.p0.id = 0;
    m.pointCount = 1;
  } else if (u2 < 0.0f && separation > kEps) {
    V2 normal = normalize(center - v2);
    separation = dot(center - v2, normal);
    if (separation > radius + kSpeculativeDistance) return m;
    V2 cA = mulAdd(v2, radiusA, normal), cB = mulSub(center, radiusB, normal);
    V2 contact = lerp(cA, cB, 0.5f);
    m.normal = rotate(xfA.q, normal);
    m.p0.anchorA = rotate(xfA.q, contact);
    m.p0.anchorB = m.p0.anchorA + (xfA.p - xfB.p);
    m.p0.point = xfA.p + m.p0.anchorA;
    m.p0.separation = dot(cB - cA, normal); m.p0.id = 0;
    m.pointCount = 1;
  } else {
    V2 normal = polyA.n[normalIndex];
    m.normal = rotate(xfA.q, normal);
    V2 cA = mulAdd(center, radiusA - dot(center - v1, normal), normal);
    V2 cB = mulSub(center, radiusB, normal);
    V2 contact = lerp(cA, cB, 0.5f);
    m.p0.anchorA = rotate(xfA.q, contact);
    m.p0.anchorB = m.p0.anchorA + (xfA.p - xfB.p);
    m.p0.point = xfA.p + m.p0.anchorA;
    m.p0.separation = separation - radius; m.p0.id = 0;
    m.pointCount = 1;
  }
  return m;
}

// src/Manifold.cs:167-315 with the port's Q5 deviations kept verbatim (also segment-vs-capsule :317-321).
__device__ inline Manifold collideCapsules(V2 a1, V2 a2, float radiusA, Xf xfA, V2 b1, V2 b2, float radiusB, Xf xfB) {
  V2 origin = a1;
  Xf sfA = mkXf(xfA.p + rotate(xfA.q, origin), xfA.q);
  Xf xf = invMulXf(sfA, xfB);
  V2 p1 = mk(0, 0);
  V2 q1 = a2 - origin;
  V2 p2 = xfPoint(xf, b1), q2 = xfPoint(xf, b2);
  V2 d1 = q1 - p1, d2 = q2 - p2;
  float dd1 = dot(d1, d1), dd2 = dot(d2, d2);
  const float epsSqr = kEps * kEps;
  V2 r = p1 - p2;
  float rd1 = dot(r, d1), rd2 = dot(r, d2), d12 = dot(d1, d2);
  float denom = dd1 * dd2 - d12 * d12;
  float f1 = 0.0f;
  if (denom != 0.0f) f1 = netClamp((d12 * rd2 - rd1 * dd2) / denom, 0.0f, 1.0f);
  float f2 = (d12 * f1 + rd2) / dd2;
  if (f2 < 0.0f) { f2 = 0.0f; f1 = netClamp(-rd1 / dd1, 0.0f, 1.0f); }
  else if (f2 > 1.0f) { f2 = 1.0f; f1 = netClamp((d12 - rd1) / dd1, 0.0f, 1.0f); }
  V2 closest1 = mulAdd(p1, f1, d1), closest2 = mulAdd(p2, f2, d2);
  float distanceSquared = distSq(closest1, closest2);
  Manifold m = emptyManifold();
  float radius = radiusA + radiusB;
  float maxDistance = radius + kSpeculativeDistance;
  if (distanceSquared > maxDistance * maxDistance) return m;
  float distance = sqrtf(distanceSquared);
  float length1, length2;
  V2 u1 = lengthAndNormalize(d1, &length1), u2 = lengthAndNormalize(d2, &length2);
  float fp2 = dot(p2 - p1, u1), fq2 = dot(q2 - p1, u1);
  bool outsideA = (fp2 < 0.0f && fq2 < 0.0f) || (fp2 >= length1 && fq2 >= length1);
  float fp1 = dot(p1 - p2, u2), fq1 = dot(q1 - p2, u2);
  bool outsideB = (fp1 < 0.0f && fq1 < 0.0f) || (fp1 >= length2 && fq1 >= length2);
  if (!outsideA && !outsideB) {
    V2 normalA; float separationA;
    {
      normalA = leftPerp(u1);
      float ss1 = dot(p2 - p1, normalA), ss2 = dot(q2 - p1, normalA);
      float s1p = ss1 < ss2 ? ss1 : ss2;
      float s1n = -ss1 < -ss2 ? -ss1 : -ss2;
      if (s1p > s1n) separationA = s1p; else { separationA = s1n; normalA = -normalA; }
    }
    V2 normalB; float separationB;
    {
      normalB = leftPerp(u2);
      float ss1 = dot(p1 - p2, normalB), ss2 = dot(q1 - p2, normalB);
      float s1p = ss1 < ss2 ? ss1 : ss2;
      float s1n = -ss1 < -ss2 ? -ss1 : -ss2;
      if (s1p > s1n) separationB = s1p; else { separationB = s1n; normalB = -normalB; }
    }
    if (separationA + 0.1f * kLinearSlop >= separationB) {
      m.normal = normalA;
      V2 cp = p2, cq = q2;
      if (fp2 < 0.0f && fq2 > 0.0f) cp = lerp(p2, q2, -fp2 / (fq2 - fp2));
      else if (fq2 < 0.0f && fp2 > 0.0f) cq = lerp(q2, p2, -fq2 / (fq2 - fp2));
      if (fp2 > length1 && fq2 < length1) cp = lerp(q2, p2, (fp2 - length1) / (fp2 - fq2));
      else if (fq2 > length1 && fp2 < length1) cq = lerp(q2, p2, (fq2 - length1) / (fq2 - fp2));
      float sp = dot(cp - p1, normalA), sq = dot(cq - p1, normalA);
      if (sp <= distance + kLinearSlop || sq <= distance + kLinearSlop) {
        m.p0.anchorA = mulAdd(cp, 0.5f * (radiusA - radiusB - sp), normalA);
        m.p0.separation = sp - radius; m.p0.id = makeId(0, 0);
        m.p1.anchorA = mulAdd(cq, 0.5f * (radiusA - radiusB - sq), normalA);
        m.p1.separation = sq - radius; m.p1.id = makeId(0, 1);
        m.pointCount = 2;
      }
    } else {
      m.normal = -normalB;
      V2 cp = p1, cq = q1;
      if (fp1 < 0.0f && fq1 > 0.0f) cp = lerp(p1, q1, -fp1 / (fq1 - fp1));
      else if (fq1 < 0.0f && fp1 > 0.0f) cq = lerp(q1, p1, -fq1 / (fp1 - fq1));
      if (fp1 > length2 && fq1 > length2) cp = lerp(p1, q1, (fp1 - length2) / (fp1 - fq1));
      else if (fq1 > length2 && fp1 > length2) cq = lerp(q1, p1, (fq1 - length2) / (fq1 - fp1));
      float sp = dot(cp - p2, normalB), sq = dot(cq - p2, normalB);
      if (sp <= distance + kLinearSlop || sq <= distance + kLinearSlop) {
        m.p0.anchorA = mulAdd(cp, 0.5f * (radiusB - radiusA - sp), normalB);
        m.p0.separation = sp - radius; m.p0.id = makeId(0, 0);
        m.p1.anchorA = mulAdd(cq, 0.5f * (radiusB - radiusA - sp), normalB);
        m.p1.separation = sp - radius; m.p1.id = makeId(1, 0);
        m.pointCount = 2;
      }
    }
  }
  if (m.pointCount == 0) {
    V2 normal = closest2 - closest1;
    if (dot(normal, normal) > epsSqr) normal = normalize(normal); else normal = leftPerp(u1);
    V2 c1 = mulAdd(closest1, radiusA, normal), c2 = mulAdd(closest2, -radiusB, normal);
    int i1 = f1 == 0.0f ? 0 : 1, i2 = f2 == 0.0f ? 0 : 1;
    m.normal = normal;
    m.p0.anchorA = lerp(c1, c2, 0.5f);
    m.p0.separation = sqrtf(distanceSquared) - radius;
    m.p0.id = makeId(i1, i2);
    m.pointCount = 1;
  }
  m.normal = rotate(xfA.q, m.normal);
  {
    m.p0.anchorA = rotate(xfA.q, m.p0.anchorA + origin);
    m.p0.anchorB = m.p0.anchorA + (xfA.p - xfB.p);
    m.p0.point = xfA.p + m.p0.anchorA;
  }
  if (m.pointCount > 1) {
    m.p1.anchorA = rotate(xfA.q, m.p1.anchorA + origin);
    m.p1.anchorB = m.p1.anchorA + (xfA.p - xfB.p);
    m.p1.point = xfA.p + m.p1.anchorA;
  }
  return m;
}

struct LocalPoly { int count; float radius; V2 v[kMaxPolygonVertices], n[kMaxPolygonVertices]; };

// src/Distance.cs:20-72
struct SegDist { float fraction1, fraction2, distanceSquared; };
__device__ inline SegDist segmentDistance(V2 p1, V2 q1, V2 p2, V2 q2) {
  SegDist res;
  V2 d1 = q1 - p1, d2 = q2 - p2, r = p1 - p2;
  float dd1 = dot(d1, d1), dd2 = dot(d2, d2);
  float rd1 = dot(r, d1), rd2 = dot(r, d2);
  const float epsSqr = kEps * kEps;
  float f1, f2;
  if (dd1 < epsSqr || dd2 < epsSqr) {
    if (dd1 >= epsSqr) { f1 = netClamp(-rd1 / dd1, 0.0f, 1.0f); f2 = 0.0f; }
    else if (dd2 >= epsSqr) { f1 = 0.0f; f2 = netClamp(rd2 / dd2, 0.0f, 1.0f); }
    else { f1 = 0.0f; f2 = 0.0f; }
  } else {
    float d12 = dot(d1, d2);
    float denominator = dd1 * dd2 - d12 * d12;
    f1 = 0.0f;
    if (denominator != 0.0f) f1 = netClamp((d12 * rd2 - rd1 * dd2) / denominator, 0.0f, 1.0f);
    f2 = (d12 * f1 + rd2) / dd2;
    if (f2 < 0.0f) { f2 = 0.0f; f1 = netClamp(-rd1 / dd1, 0.0f, 1.0f); }
    else if (f2 > 1.0f) { f2 = 1.0f; f1 = netClamp((d12 - rd1) / dd1, 0.0f, 1.0f); }
  }
  res.fraction1 = f1; res.fraction2 = f2;
  V2 c1 = mulAdd(p1, f1, d1), c2 = mulAdd(p2, f2, d2);
  res.distanceSquared = distSq(c1, c2);
  return res;
}

// src/Manifold.cs:329-388
__device__ inline Manifold clipPolygons(const LocalPoly& polyA, const LocalPoly& polyB, int edgeA, int edgeB, bool flip) {
  Manifold m = emptyManifold();
  const LocalPoly* poly1 = &polyA; const LocalPoly* poly2 = &polyB;
  int i11 = edgeA, i12 = edgeA + 1 < polyA.count ? edgeA + 1 : 0;
  int i21 = edgeB, i22 = edgeB + 1 < polyB.count ? edgeB + 1 : 0;
  if (flip) { poly1 = &polyB; poly2 = &polyA; int t11 = i11, t12 = i12; i11 = i21; i12 = i22; i21 = t11; i22 = t12; }
  V2 normal = poly1->n[i11];
  V2 v11 = poly1->v[i11], v12 = poly1->v[i12];
  V2 v21 = poly2->v[i21], v22 = poly2->v[i22];
  V2 tangent = crossSV(1.0f, normal);
  float lower1 = 0.0f;
  float upper1 = dot(v12 - v11, tangent);
  float upper2 = dot(v21 - v11, tangent);
  float lower2 = dot(v22 - v11, tangent);
  if (upper2 < lower1 || upper1 < lower2) return m;
  V2 vLower = (lower2 < lower1 && upper2 - lower2 > kEps) ? lerp(v22, v21, (lower1 - lower2) / (upper2 - lower2)) : v22;
  V2 vUpper = (upper2 > upper1 && upper2 - lower2 > kEps) ? lerp(v22, v21, (upper1 - lower2) / (upper2 - lower2)) : v21;
  float separationLower = dot(vLower - v11, normal);
  float separationUpper = dot(vUpper - v11, normal);
  float r1 = poly1->radius, r2 = poly2->radius;
  vLower = mulAdd(vLower, 0.5f * (r1 - r2 - separationLower), normal);
  vUpper = mulAdd(vUpper, 0.5f * (r1 - r2 - separationUpper), normal);
  float radius = r1 + r2;
  if (!flip) {
    m.normal = normal;
    m.p0.anchorA = vLower; m.p0.separation = separationLower - radius; m.p0.id = makeId(i11, i22);
    m.p1.anchorA = vUpper; m.p1.separation = separationUpper - radius; m.p1.id = makeId(i12, i21);
  } else {
    m.normal = -normal;
    m.p0.anchorA = vUpper; m.p0.separation = separationUpper - radius; m.p0.id = makeId(i21, i12);
    m.p1.anchorA = vLower; m.p1.separation = separationLower - radius; m.p1.id = makeId(i22, i11);
  }
  m.pointCount = 2;
  return m;
}

// src/Manifold.cs:389-408
__device__ inline float findMaxSeparation(int* edgeIndex, const LocalPoly& poly1, const LocalPoly& poly2) {
  int bestIndex = 0;
  float maxSeparation = -FLT_MAX;
  for (int i = 0; i < poly1.count; i++) {
    V2 n = poly1.n[i], v1 = poly1.v[i];
    float si = FLT_MAX;
    for (int j = 0; j < poly2.count; j++) { float sij = dot(n, poly2.v[j] - v1); if (sij < si) si = sij; }
    if (si > maxSeparation) { maxSeparation = si; bestIndex = i; }
  }
  *edgeIndex = bestIndex;
  return maxSeparation;
}

// src/Manifold.cs:410-562
__device__ inline Manifold collidePolygons(const Geom& polygonA, Xf xfA, const Geom& polygonB, Xf xfB) {
  V2 origin = polygonA.v[0];
  Xf sfA = mkXf(xfA.p + rotate(xfA.q, origin), xfA.q);
  Xf xf = invMulXf(sfA, xfB);
  LocalPoly A; A.radius = polygonA.radius; A.count = polygonA.count;
  A.v[0] = mk(0, 0); A.n[0] = polygonA.n[0];
  for (int i = 1; i < A.count; i++) { A.v[i] = polygonA.v[i] - origin; A.n[i] = polygonA.n[i]; }
  LocalPoly B; B.radius = polygonB.radius; B.count = polygonB.count;
  for (int i = 0; i < B.count; i++) { B.v[i] = xfPoint(xf, polygonB.v[i]); B.n[i] = rotate(xf.q, polygonB.n[i]); }
  int edgeA, edgeB;
  float separationA = findMaxSeparation(&edgeA, A, B);
  float separationB = findMaxSeparation(&edgeB, B, A);
  float radius = A.radius + B.radius;
  if (separationA > kSpeculativeDistance + radius || separationB > kSpeculativeDistance + radius) return emptyManifold();
  bool flip;
  if (separationA >= separationB) {
    flip = false;
    V2 dir = A.n[edgeA];
    edgeB = 0; float minDot = FLT_MAX;
    for (int i = 0; i < B.count; i++) { float dt = dot(dir, B.n[i]); if (dt < minDot) { minDot = dt; edgeB = i; } }
  } else {
    flip = true;
    V2 dir = B.n[edgeB];
    edgeA = 0; float minDot = FLT_MAX;
    for (int i = 0; i < A.count; i++) { float dt = dot(dir, A.n[i]); if (dt < minDot) { minDot = dt; edgeA = i; } }
  }
  Manifold m = emptyManifold();
  if (separationA > 0.1f * kLinearSlop || separationB > 0.1f * kLinearSlop) {
    int i11 = edgeA, i12 = edgeA + 1 < A.count ? edgeA + 1 : 0;
    int i21 = edgeB, i22 = edgeB + 1 < B.count ? edgeB + 1 : 0;
    V2 v11 = A.v[i11], v12 = A.v[i12], v21 = B.v[i21], v22 = B.v[i22];
    SegDist result = segmentDistance(v11, v12, v21, v22);
    float distance = sqrtf(result.distanceSquared);
    float separation = distance - radius;
    if (distance - radius > kSpeculativeDistance) return m;
    m = clipPolygons(A, B, edgeA, edgeB, flip);
    float minSeparation = FLT_MAX;
    if (m.pointCount > 0) minSeparation = netMin(minSeparation, m.p0.separation);
    if (m.pointCount > 1) minSeparation = netMin(minSeparation, m.p1.separation);
    if (separation + 0.1f * kLinearSlop < minSeparation) {
      V2 va = mk(0, 0), vb = mk(0, 0); int ia = -1, ib = -1;
      if (result.fraction1 == 0.0f && result.fraction2 == 0.0f) { va = v11; vb = v21; ia = i11; ib = i21; }
      else if (result.fraction1 == 0.0f && result.fraction2 == 1.0f) { va = v11; vb = v22; ia = i11; ib = i22; }
      else if (result.fraction1 == 1.0f && result.fraction2 == 0.0f) { va = v12; vb = v21; ia = i12; ib = i21; }
      else if (result.fraction1 == 1.0f && result.fraction2 == 1.0f) { va = v12; vb = v22; ia = i12; ib = i22; }
      if (ia >= 0) {
        V2 normal = vb - va;
        float invDistance = 1.0f / distance;
        normal = invDistance * normal;
        V2 c1 = mulAdd(va, A.radius, normal), c2 = mulAdd(vb, -B.radius, normal);
        m.normal = normal;
        m.p0.anchorA = lerp(c1, c2, 0.5f);
        m.p0.separation = distance - radius;
        m.p0.id = makeId(ia, ib);
        m.pointCount = 1;
      }
    }
  } else m = clipPolygons(A, B, edgeA, edgeB, flip);
  if (m.pointCount > 0) {
    m.normal = rotate(xfA.q, m.normal);
    m.p0.anchorA = rotate(xfA.q, m.p0.anchorA + origin);
    m.p0.anchorB = m.p0.anchorA + (xfA.p - xfB.p);
    m.p0.point = xfA.p + m.p0.anchorA;
    if (m.pointCount > 1) {
      m.p1.anchorA = rotate(xfA.q, m.p1.anchorA + origin);
      m.p1.anchorB = m.p1.anchorA + (xfA.p - xfB.p);
      m.p1.point = xfA.p + m.p1.anchorA;
    }
  }
  return m;
}

// ---- the same function with the vertex counts known at compile time (boxes: 4 x 4). Every loop unrolls, the vertex and normal arrays live
// in registers (the few run-time indexed reads go through select chains) and the 32 independent dot products of the two FindMaxSeparation
// passes overlap instead of running as one dependent chain out of local memory. Operation for operation the arithmetic of collidePolygons
// above: each value is computed by the same expression, comparisons run in the same order, nothing is contracted (-fmad=false).
template <int C> struct PolyT { float radius; V2 v[C], n[C]; };
template <int C> __device__ __forceinline__ V2 pickT(const V2 (&arr)[C], int i) {
  V2 r = arr[0];
#pragma unroll
  for (int k = 1; k < C; k++) if (i == k) r = arr[k];
  return r;
}
template <int C1, int C2> __device__ __forceinline__ float findMaxSeparationT(int* edgeIndex, const PolyT<C1>& poly1, const PolyT<C2>& poly2) {
  int bestIndex = 0;
  float maxSeparation = -FLT_MAX;
#pragma unroll
  for (int i = 0; i < C1; i++) {
    V2 n = poly1.n[i], v1 = poly1.v[i];
    float si = FLT_MAX;
#pragma unroll
    for (int j = 0; j < C2; j++) { float sij = dot(n, poly2.v[j] - v1); if (sij < si) si = sij; }
    if (si > maxSeparation) { maxSeparation = si; bestIndex = i; }
  }
  *edgeIndex = bestIndex;
  return maxSeparation;
}
// ClipPolygons (src/Manifold.cs:329-388) on picked vertices: (normal, v11, v12) from the reference polygon, (v21, v22) from the incident one
__device__ __forceinline__ Manifold clipPicked(V2 normal, V2 v11, V2 v12, V2 v21, V2 v22, float r1, float r2, int i11, int i12, int i21, int i22, bool flip) {
  Manifold m = emptyManifold();
  V2 tangent = crossSV(1.0f, normal);
  float lower1 = 0.0f;
  float upper1 = dot(v12 - v11, tangent);
  float upper2 = dot(v21 - v11, tangent);
  float lower2 = dot(v22 - v11, tangent);
  if (upper2 < lower1 || upper1 < lower2) return m;
  V2 vLower = (lower2 < lower1 && upper2 - lower2 > kEps) ? lerp(v22, v21, (lower1 - lower2) / (upper2 - lower2)) : v22;
  V2 vUpper = (upper2 > upper1 && upper2 - lower2 > kEps) ? lerp(v22, v21, (upper1 - lower2) / (upper2 - lower2)) : v21;
  float separationLower = dot(vLower - v11, normal);
  float separationUpper = dot(vUpper - v11, normal);
  vLower = mulAdd(vLower, 0.5f * (r1 - r2 - separationLower), normal);
  vUpper = mulAdd(vUpper, 0.5f * (r1 - r2 - separationUpper), normal);
  float radius = r1 + r2;
  if (!flip) {
    m.normal = normal;
    m.p0.anchorA = vLower; m.p0.separation = separationLower - radius; m.p0.id = makeId(i11, i22);
    m.p1.anchorA = vUpper; m.p1.separation = separationUpper - radius; m.p1.id = makeId(i12, i21);
  } else {
    m.normal = -normal;
    m.p0.anchorA = vUpper; m.p0.separation = separationUpper - radius; m.p0.id = makeId(i21, i12);
    m.p1.anchorA = vLower; m.p1.separation = separationLower - radius; m.p1.id = makeId(i22, i11);
  }
  m.pointCount = 2;
  return m;
}
template <int CA, int CB> __device__ inline Manifold collidePolygonsT(const Geom& polygonA, Xf xfA, const Geom& polygonB, Xf xfB) {
  V2 origin = polygonA.v[0];
  Xf sfA = mkXf(xfA.p + rotate(xfA.q, origin), xfA.q);
  Xf xf = invMulXf(sfA, xfB);
  PolyT<CA> A; A.radius = polygonA.radius;
  A.v[0] = mk(0, 0); A.n[0] = polygonA.n[0];
#pragma unroll
  for (int i = 1; i < CA; i++) { A.v[i] = polygonA.v[i] - origin; A.n[i] = polygonA.n[i]; }
  PolyT<CB> B; B.radius = polygonB.radius;
#pragma unroll
  for (int i = 0; i < CB; i++) { B.v[i] = xfPoint(xf, polygonB.v[i]); B.n[i] = rotate(xf.q, polygonB.n[i]); }
  int edgeA, edgeB;
  float separationA = findMaxSeparationT(&edgeA, A, B);
  float separationB = findMaxSeparationT(&edgeB, B, A);
  float radius = A.radius + B.radius;
  if (separationA > kSpeculativeDistance + radius || separationB > kSpeculativeDistance + radius) return emptyManifold();
  bool flip;
  if (separationA >= separationB) {
    flip = false;
    V2 dir = pickT(A.n, edgeA);
    edgeB = 0; float minDot = FLT_MAX;
#pragma unroll
    for (int i = 0; i < CB; i++) { float dt = dot(dir, B.n[i]); if (dt < minDot) { minDot = dt; edgeB = i; } }
  } else {
    flip = true;
    V2 dir = pickT(B.n, edgeB);
    edgeA = 0; float minDot = FLT_MAX;
#pragma unroll
    for (int i = 0; i < CA; i++) { float dt = dot(dir, A.n[i]); if (dt < minDot) { minDot = dt; edgeA = i; } }
  }
  const int i11 = edgeA, i12 = edgeA + 1 < CA ? edgeA + 1 : 0;
  const int i21 = edgeB, i22 = edgeB + 1 < CB ? edgeB + 1 : 0;
  const V2 v11 = pickT(A.v, i11), v12 = pickT(A.v, i12), v21 = pickT(B.v, i21), v22 = pickT(B.v, i22);
  // ClipPolygons swaps the roles when flip is set: the reference polygon is B, the incident one A
  const V2 clipNormal = flip ? pickT(B.n, i21) : pickT(A.n, i11);
  Manifold m = emptyManifold();
  if (separationA > 0.1f * kLinearSlop || separationB > 0.1f * kLinearSlop) {
    SegDist result = segmentDistance(v11, v12, v21, v22);
    float distance = sqrtf(result.distanceSquared);
    float separation = distance - radius;
    if (distance - radius > kSpeculativeDistance) return m;
    m = flip ? clipPicked(clipNormal, v21, v22, v11, v12, B.radius, A.radius, i21, i22, i11, i12, true)
             : clipPicked(clipNormal, v11, v12, v21, v22, A.radius, B.radius, i11, i12, i21, i22, false);
    float minSeparation = FLT_MAX;
    if (m.pointCount > 0) minSeparation = netMin(minSeparation, m.p0.separation);
    if (m.pointCount > 1) minSeparation = netMin(minSeparation, m.p1.separation);
    if (separation + 0.1f * kLinearSlop < minSeparation) {
      V2 va = mk(0, 0), vb = mk(0, 0); int ia = -1, ib = -1;
      if (result.fraction1 == 0.0f && result.fraction2 == 0.0f) { va = v11; vb = v21; ia = i11; ib = i21; }
      else if (result.fraction1 == 0.0f && result.fraction2 == 1.0f) { va = v11; vb = v22; ia = i11; ib = i22; }
      else if (result.fraction1 == 1.0f && result.fraction2 == 0.0f) { va = v12; vb = v21; ia = i12; ib = i21; }
      else if (result.fraction1 == 1.0f && result.fraction2 == 1.0f) { va = v12; vb = v22; ia = i12; ib = i22; }
      if (ia >= 0) {
        V2 normal = vb - va;
        float invDistance = 1.0f / distance;
        normal = invDistance * normal;
        V2 c1 = mulAdd(va, A.radius, normal), c2 = mulAdd(vb, -B.radius, normal);
        m.normal = normal;
        m.p0.anchorA = lerp(c1, c2, 0.5f);
        m.p0.separation = distance - radius;
        m.p0.id = makeId(ia, ib);
        m.pointCount = 1;
      }
    }
  } else m = flip ? clipPicked(clipNormal, v21, v22, v11, v12, B.radius, A.radius, i21, i22, i11, i12, true)
                  : clipPicked(clipNormal, v11, v12, v21, v22, A.radius, B.radius, i11, i12, i21, i22, false);
  if (m.pointCount > 0) {
    m.normal = rotate(xfA.q, m.normal);
    m.p0.anchorA = rotate(xfA.q, m.p0.anchorA + origin);
    m.p0.anchorB = m.p0.anchorA + (xfA.p - xfB.p);
    m.p0.point = xfA.p + m.p0.anchorA;
    if (m.pointCount > 1) {
      m.p1.anchorA = rotate(xfA.q, m.p1.anchorA + origin);
      m.p1.anchorB = m.p1.anchorA + (xfA.p - xfB.p);
      m.p1.point = xfA.p + m.p1.anchorA;
    }
  }
  return m;
}

// Geometry.MakeCapsule as a 2-gon (src/Manifold.cs:8-17)
__device__ inline void makeCapsulePoly(Geom& g, V2 p1, V2 p2, float radius) {
  g.type = kPolygon; g.count = 2; g.radius = radius;
  g.v[0] = p1; g.v[1] = p2; g.centroid = lerp(p1, p2, 0.5f);
  V2 axis = normalize(p2 - p1);
  V2 normal = rightPerp(axis);
  g.n[0] = normal; g.n[1] = -normal;
}

// src/Contact.cs:420-474 ContactRegister: which (typeA,typeB) orders are registered
__host__ __device__ __forceinline__ bool pairSupported(int tA, int tB) { return tA <= kPolygon && tB <= kPolygon && !(tA == kSegment && tB == kSegment); }
__host__ __device__ __forceinline__ bool pairPrimary(int tA, int tB) {
  if (tA == tB) return true;
  if (tA == kCapsule && tB == kCircle) return true;
  if (tA == kPolygon && (tB == kCircle || tB == kCapsule)) return true;
  if (tA == kSegment && (tB == kCircle || tB == kCapsule || tB == kPolygon)) return true;
  return false;
}

__device__ inline Manifold computeManifold(const Geom& a, Xf xfA, const Geom& b, Xf xfB) {
  switch (a.type) {
    case kCircle: return collideCircles(a.v[0], a.radius, xfA, b.v[0], b.radius, xfB);
    case kCapsule:
      if (b.type == kCircle) return collideCapsuleAndCircle(a.v[0], a.v[1], a.radius, xfA, b.v[0], b.radius, xfB);
      return collideCapsules(a.v[0], a.v[1], a.radius, xfA, b.v[0], b.v[1], b.radius, xfB);
    case kSegment: {
      if (b.type == kCircle) return collideCapsuleAndCircle(a.v[0], a.v[1], 0.0f, xfA, b.v[0], b.radius, xfB);
      if (b.type == kCapsule) return collideCapsules(a.v[0], a.v[1], 0.0f, xfA, b.v[0], b.v[1], b.radius, xfB);
      Geom cap; makeCapsulePoly(cap, a.v[0], a.v[1], 0.0f);
      return collidePolygons(cap, xfA, b, xfB);
    }
    default: {
      if (b.type == kCircle) return collidePolygonAndCircle(a, xfA, b.v[0], b.radius, xfB);
      if (b.type == kCapsule) { Geom cap; makeCapsulePoly(cap, b.v[0], b.v[1], b.radius); return collidePolygons(a, xfA, cap, xfB); }
      if (a.count == 4 && b.count == 4) return collidePolygonsT<4, 4>(a, xfA, b, xfB);   // boxes: the unrolled path (same arithmetic)
      return collidePolygons(a, xfA, b, xfB);
    }
  }
}

}  // namespace b2c
