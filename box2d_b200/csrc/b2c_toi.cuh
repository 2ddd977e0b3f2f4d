// libb2cuda — time of impact (Distance.TimeOfImpact, src/Distance.cs:529-646) with its separation function (:381-527) and the sweep
// (:9-18), as the port writes them. Port behaviour kept on purpose:
//   * the GJK cache is carried from one conservative-advancement iteration to the next, so every ShapeDistance call after the first goes
//     through the port's MakeSimplexFromCache (quirk Q19, see b2c_distance.cuh): a deterministic but different start simplex, or an
//     IndexOutOfRangeException where B's cached index is past proxy A's points (*throws);
//   * SeparationType.FaceA computes `normal = xfA.q * f.localPoint` BEFORE localPoint is assigned (:434-435): the sign test sees a zero
//     vector and never flips the axis.
// Shared host/device code (the host uses it for nothing today; it compiles both ways like the rest of the geometry).
#pragma once
#include "b2c_distance.cuh"

namespace b2c {

struct Sweep { V2 localCenter, c1, c2; Rot q1, q2; };                  // include/Collision.cs:867-879
enum ToiState { kToiUnknown = 0, kToiFailed = 1, kToiOverlapped = 2, kToiHit = 3, kToiSeparated = 4 };
struct ToiOut { int state; float fraction; V2 point, normal; };
enum SepType { kSepPoints = 0, kSepFaceA = 1, kSepFaceB = 2 };
struct SepFn { const DistProxy* proxyA; const DistProxy* proxyB; Sweep sweepA, sweepB; V2 localPoint, axis; int type; };

B2C_HD Xf sweepTransform(const Sweep& sw, float time) {               // GetSweepTransform :9-18
  Xf xf;
  const float u = 1.0f - time;
  xf.p = mk(u * sw.c1.x + time * sw.c2.x, u * sw.c1.y + time * sw.c2.y);
  xf.q = normalizeRot(mkRot(u * sw.q1.c + time * sw.q2.c, u * sw.q1.s + time * sw.q2.s));
  xf.p = xf.p - rotate(xf.q, sw.localCenter);
  return xf;
}
B2C_HD V2 crossVS(V2 v, float s) { return mk(s * v.y, -s * v.x); }     // include/MathFunctions.cs Vector2.CrossVS

B2C_HD SepFn makeSeparationFunction(const SimplexCache& cache, const DistProxy& proxyA, const Sweep& sweepA, const DistProxy& proxyB, const Sweep& sweepB, float t1) {   // :381-443
  SepFn f; f.proxyA = &proxyA; f.proxyB = &proxyB; f.sweepA = sweepA; f.sweepB = sweepB; f.localPoint = mk(0, 0); f.axis = mk(0, 0); f.type = kSepPoints;
  const int count = cache.count;
  Xf xfA = sweepTransform(sweepA, t1), xfB = sweepTransform(sweepB, t1);
  if (count == 1) {
    f.type = kSepPoints;
    V2 pointA = xfPoint(xfA, proxyA.pts[cache.indexA[0]]), pointB = xfPoint(xfB, proxyB.pts[cache.indexB[0]]);
    f.axis = normalize(pointB - pointA);
    f.localPoint = mk(0, 0);
    return f;
  }
  if (cache.indexA[0] == cache.indexA[1]) {
    f.type = kSepFaceB;
    V2 localPointB1 = proxyB.pts[cache.indexB[0]], localPointB2 = proxyB.pts[cache.indexB[1]];
    f.axis = normalize(crossVS(localPointB2 - localPointB1, 1.0f));
    V2 normal = rotate(xfB.q, f.axis);
    f.localPoint = mk(0.5f * (localPointB1.x + localPointB2.x), 0.5f * (localPointB1.y + localPointB2.y));
    V2 pointB = xfPoint(xfB, f.localPoint);
    V2 pointA = xfPoint(xfA, proxyA.pts[cache.indexA[0]]);
    float s = dot(pointA - pointB, normal);
    if (s < 0.0f) f.axis = mk(-f.axis.x, -f.axis.y);
    return f;
  }
  f.type = kSepFaceA;
  V2 localPointA1 = proxyA.pts[cache.indexA[0]], localPointA2 = proxyA.pts[cache.indexA[1]];
  f.axis = normalize(crossVS(localPointA2 - localPointA1, 1.0f));
  V2 normal = rotate(xfA.q, f.localPoint);                             // :434 as written: localPoint is still (0, 0) here
  f.localPoint = mk(0.5f * (localPointA1.x + localPointA2.x), 0.5f * (localPointA1.y + localPointA2.y));
  V2 pointA = xfPoint(xfA, f.localPoint);
  V2 pointB = xfPoint(xfB, proxyB.pts[cache.indexB[0]]);
  float s = dot(pointB - pointA, normal);
  if (s < 0.0f) f.axis = mk(-f.axis.x, -f.axis.y);
  return f;
}

B2C_HD float findMinSeparation(const SepFn& f, int* indexA, int* indexB, float t) {   // :445-490
  Xf xfA = sweepTransform(f.sweepA, t), xfB = sweepTransform(f.sweepB, t);
  if (f.type == kSepPoints) {
    V2 axisA = invRotate(xfA.q, f.axis), axisB = invRotate(xfB.q, mk(-f.axis.x, -f.axis.y));
    *indexA = findSupport(*f.proxyA, axisA); *indexB = findSupport(*f.proxyB, axisB);
    V2 pointA = xfPoint(xfA, f.proxyA->pts[*indexA]), pointB = xfPoint(xfB, f.proxyB->pts[*indexB]);
    return dot(pointB - pointA, f.axis);
  }
  if (f.type == kSepFaceA) {
    V2 normal = rotate(xfA.q, f.axis);
    V2 pointA = xfPoint(xfA, f.localPoint);
    V2 axisB = invRotate(xfB.q, mk(-normal.x, -normal.y));
    *indexA = -1; *indexB = findSupport(*f.proxyB, axisB);
    V2 pointB = xfPoint(xfB, f.proxyB->pts[*indexB]);
    return dot(pointB - pointA, normal);
  }
  V2 normal = rotate(xfB.q, f.axis);
  V2 pointB = xfPoint(xfB, f.localPoint);
  V2 axisA = invRotate(xfA.q, mk(-normal.x, -normal.y));
  *indexB = -1; *indexA = findSupport(*f.proxyA, axisA);
  V2 pointA = xfPoint(xfA, f.proxyA->pts[*indexA]);
  return dot(pointA - pointB, normal);
}

B2C_HD float evaluateSeparation(const SepFn& f, int indexA, int indexB, float t) {    // :491-527
  Xf xfA = sweepTransform(f.sweepA, t), xfB = sweepTransform(f.sweepB, t);
  if (f.type == kSepPoints) {
    V2 pointA = xfPoint(xfA, f.proxyA->pts[indexA]), pointB = xfPoint(xfB, f.proxyB->pts[indexB]);
    return dot(pointB - pointA, f.axis);
  }
  if (f.type == kSepFaceA) {
    V2 normal = rotate(xfA.q, f.axis);
    V2 pointA = xfPoint(xfA, f.localPoint), pointB = xfPoint(xfB, f.proxyB->pts[indexB]);
    return dot(pointB - pointA, normal);
  }
  V2 normal = rotate(xfB.q, f.axis);
  V2 pointB = xfPoint(xfB, f.localPoint), pointA = xfPoint(xfA, f.proxyA->pts[indexA]);
  return dot(pointA - pointB, normal);
}

B2C_HD ToiOut timeOfImpact(const DistProxy& proxyA, const DistProxy& proxyB, const Sweep& sweepA, const Sweep& sweepB, float maxFraction, bool* throws) {   // :529-646
  ToiOut output; output.state = kToiUnknown; output.fraction = maxFraction; output.point = mk(0, 0); output.normal = mk(0, 0);
  const float tMax = maxFraction;
  const float totalRadius = proxyA.radius + proxyB.radius;
  const float target = netMax(kLinearSlop, totalRadius - kLinearSlop);
  const float tolerance = 0.25f * kLinearSlop;
  float t1 = 0.0f;
  const int maxPushBack = proxyA.count > proxyB.count ? proxyA.count : proxyB.count;   // `MAX_POLYGON_VERTICES` :547
  int distanceIterations = 0;
  SimplexCache cache; cache.count = 0;
  for (;;) {
    Xf xfA = sweepTransform(sweepA, t1), xfB = sweepTransform(sweepB, t1);
    DistOut distanceOutput = shapeDistanceCached(proxyA, proxyB, xfA, xfB, false, cache, throws);
    if (*throws) return output;
    distanceIterations++;
    if (distanceOutput.distance <= 0.0f) { output.state = kToiOverlapped; output.fraction = 0.0f; break; }
    if (distanceOutput.distance <= target + tolerance) {
      output.state = kToiHit;
      V2 pA = mulAdd(distanceOutput.pointA, proxyA.radius, distanceOutput.normal), pB = mulAdd(distanceOutput.pointB, -proxyB.radius, distanceOutput.normal);
      output.point = lerp(pA, pB, 0.5f); output.normal = distanceOutput.normal; output.fraction = t1;
      break;
    }
    SepFn fcn = makeSeparationFunction(cache, proxyA, sweepA, proxyB, sweepB, t1);
    bool done = false;
    float t2 = tMax;
    int pushBackIterations = 0;
    for (;;) {
      int indexA, indexB;
      float s2 = findMinSeparation(fcn, &indexA, &indexB, t2);
      if (s2 > target + tolerance) { output.state = kToiSeparated; output.fraction = tMax; done = true; break; }
      if (s2 > target - tolerance) { t1 = t2; break; }
      float s1 = evaluateSeparation(fcn, indexA, indexB, t1);
      if (s1 < target - tolerance) { output.state = kToiFailed; output.fraction = t1; done = true; break; }
      if (s1 <= target + tolerance) {
        output.state = kToiHit;
        V2 pA = mulAdd(distanceOutput.pointA, proxyA.radius, distanceOutput.normal), pB = mulAdd(distanceOutput.pointB, -proxyB.radius, distanceOutput.normal);
        output.point = lerp(pA, pB, 0.5f); output.fraction = t1;
        done = true; break;
      }
      int rootIterationCount = 0;
      float a1 = t1, a2 = t2;
      while (rootIterationCount < 50) {
        float t;
        if ((rootIterationCount & 1) == 1) t = a1 + (target - s1) * (a2 - a1) / (s2 - s1);
        else t = 0.5f * (a1 + a2);
        rootIterationCount++;
        float s = evaluateSeparation(fcn, indexA, indexB, t);
        if (fabsf(s - target) < tolerance) { t2 = t; break; }
        if (s > target) { a1 = t; s1 = s; } else { a2 = t; s2 = s; }
      }
      pushBackIterations++;
      if (pushBackIterations == maxPushBack) break;
    }
    if (done) break;
    if (distanceIterations == 20) {
      output.state = kToiFailed;
      V2 pA = mulAdd(distanceOutput.pointA, proxyA.radius, distanceOutput.normal), pB = mulAdd(distanceOutput.pointB, -proxyB.radius, distanceOutput.normal);
      output.point = lerp(pA, pB, 0.5f); output.normal = distanceOutput.normal; output.fraction = t1;
      break;
    }
  }
  return output;
}

// ShapeExtent.minExtent of a shape about its own centroid (include/Collision.cs:92,224-228,485-496,649-653), for the core-shape retry of
// ContinuousQueryCallback (src/Solver.cs:291-299)
B2C_HD float geomMinExtent(const Geom& g) {
  if (g.type == kCircle || g.type == kCapsule) return g.radius;
  if (g.type == kSegment) return 0.0f;
  float mn = kHuge;
  for (int i = 0; i < g.count; i++) mn = netMin(mn, dot(g.n[i], g.v[i] - g.centroid));
  return mn + g.radius;
}

}  // namespace b2c
