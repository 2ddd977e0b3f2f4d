// libb2cuda — sensor overlaps and sensor begin / end events (World.OverlapSensors, src/Sensor.cs:36-218), run at the end of the step after
// FinalizeBodies and the refit (api/World.cs:97-102). One thread per sensor walks the three trees with the sensor shape's AABB and mask
// (DynamicTree.Query, src/DynamicTree.cs:531-557), filters as SensorQueryCallback does (:36-76, as the port writes it: other sensors and
// static bodies are NOT skipped) and keeps the visitors whose GJK distance with radii is below 10 * FLT_EPSILON. The reference sorts the
// hits by shape id and removes duplicates, so the traversal order does not show; the sorted set is compared with the previous step's and
// the difference becomes events, sensor by sensor in World.sensors order (ascending bit order of eventBits, :143-217), through an
// exclusive scan of the per-sensor event counts. Included by b2c_step.cu (the phase runs at grid scope or, for a batch world, at CTA scope).
#pragma once
#include "b2c_distance.cuh"

namespace b2c {

constexpr int kSensorStack = 128;

// the two overlap lists of sensor i sit side by side ([sensor][parity][kSensorMaxOverlaps]) so that the arrays can grow by plain copies
__device__ __forceinline__ int2* sensorList(const WorldDev& d, int i, int parity) { return d.sensorOverlaps + ((size_t)i * 2 + parity) * kSensorMaxOverlaps; }

// overlaps2 of one sensor into `out` (sorted by shape id); returns the count
__device__ inline int sensorQuery(const WorldDev& d, int sensorIndex, int sensorShapeId, int2* out) {
  Scalars& s = *d.sc;
  const ShapeDev& ss = d.shapes[sensorShapeId];
  int n = 0;
  {                                                      // :90-93 overlaps2 starts with the continuous hits of this step
    const int hits = min(d.sensorHitCount[sensorIndex], kSensorMaxHits);
    const int2* h = d.sensorHits + (size_t)sensorIndex * kSensorMaxHits;
    for (int i = 0; i < hits; i++) {
      int k = 0; while (k < n && out[k].x < h[i].x) k++;
      if (k < n && out[k].x == h[i].x) continue;         // the unique pass (:111-116) keeps the first of equal shape ids
      for (int m = n; m > k; m--) out[m] = out[m - 1];
      out[k] = h[i]; n++;
    }
    d.sensorHitCount[sensorIndex] = 0;
  }
  if (!(ss.flags & SF_SENSOR_EVENTS)) return n;          // :96-101 (the disabled set does not exist here): the hits stay, nothing is queried
  const float4 xa = d.bodyXf[ss.bodyId];
  const Xf xfA = mkXf(mk(xa.x, xa.y), mkRot(xa.z, xa.w));
  const Box bounds = toBox(d.shapeAABB[sensorShapeId]);
  const DistProxy proxyA = makeDistProxy(ss.geom);
  for (int tt = 0; tt < 3; tt++) {
    const TreeDev& t = d.trees[tt];
    int root = t.meta[0];
    if (t.meta[1] == 0 || root == -1) continue;
    int stack[kSensorStack]; int sp = 0;
    stack[sp++] = root;
    while (sp > 0) {
      int id = stack[--sp];
      if (!boxOverlaps(toBox(t.aabb[id]), bounds) || (t.cat[id] & ss.maskBits) == 0) continue;
      int2 c = t.child[id];
      if (c.x != -1) {
        if (sp + 2 > kSensorStack) { atomicOr(&s.errorFlags, (int)ERR_STACK_OVERFLOW); continue; }
        stack[sp++] = c.x; stack[sp++] = c.y;
        continue;
      }
      const int other = t.user[id];
      if (other == sensorShapeId) continue;
      const ShapeDev& so = d.shapes[other];
      if (!(so.flags & SF_SENSOR_EVENTS)) continue;
      if (so.bodyId == ss.bodyId) continue;
      if (!shouldShapesCollide(ss, so)) continue;
      const float4 xb = d.bodyXf[so.bodyId];
      DistOut o = shapeDistance(proxyA, makeDistProxy(so.geom), xfA, mkXf(mk(xb.x, xb.y), mkRot(xb.z, xb.w)), true);
      if (o.distance >= 10.0f * 1.192092896e-07f) continue;
      if (n >= kSensorMaxOverlaps) { atomicOr(&s.errorFlags, (int)ERR_SENSOR_CAPACITY); continue; }
      int k = 0; while (k < n && out[k].x < other) k++;  // insertion by shape id; a shape sits in one tree, but a continuous hit may have brought it already
      if (k < n && out[k].x == other) continue;
      for (int m = n; m > k; m--) out[m] = out[m - 1];
      out[k] = make_int2(other, (int)so.generation); n++;
    }
  }
  return n;
}

// the merge of :155-214: counts the events (WRITE = false) or writes them at the given offsets
template <bool WRITE> __device__ inline int2 sensorMerge(const int2* o1, int count1, const int2* o2, int count2, B2CShapeRef sensor,
                                                         B2CSensorEvent* beginOut, B2CSensorEvent* endOut) {
  int index1 = 0, index2 = 0, nb = 0, ne = 0;
  auto ev = [&](int2 v) { B2CSensorEvent e; e.sensor = sensor; e.visitor.shapeId = v.x; e.visitor.generation = (uint32_t)v.y; return e; };
  while (index1 < count1 && index2 < count2) {
    int2 r1 = o1[index1], r2 = o2[index2];
    if (r1.x == r2.x) {
      if ((uint32_t)r1.y < (uint32_t)r2.y) { if (WRITE) endOut[ne] = ev(r1); ne++; index1++; }
      else if ((uint32_t)r1.y > (uint32_t)r2.y) { if (WRITE) beginOut[nb] = ev(r2); nb++; index2++; }
      else { index1++; index2++; }
    } else if (r1.x < r2.x) { if (WRITE) endOut[ne] = ev(r1); ne++; index1++; }
    else { if (WRITE) beginOut[nb] = ev(r2); nb++; index2++; }
  }
  while (index1 < count1) { if (WRITE) endOut[ne] = ev(o1[index1]); ne++; index1++; }
  while (index2 < count2) { if (WRITE) beginOut[nb] = ev(o2[index2]); nb++; index2++; }
  return make_int2(nb, ne);
}

template <class S> __device__ void phase_sensors(const S& grid, const WorldDev& d, int endIndex) {
  const int n = d.sensorCount;
  if (n == 0) return;
  Scalars& s = *d.sc;
  const int tid = grid.tid(), stride = grid.size();
  const int parity = 1 - s.sensorParity;                 // :88 the lists trade places; every thread reads the old value before the flip below
  grid.sync();
  if (tid == 0) s.sensorParity = parity;
  for (int i = tid; i < n; i += stride) {
    const int shapeId = d.sensorShape[i];
    int2* o2 = sensorList(d, i, parity); const int2* o1 = sensorList(d, i, 1 - parity);
    const int count2 = sensorQuery(d, i, shapeId, o2), count1 = d.sensorOverlapCount[2 * i + 1 - parity];
    d.sensorOverlapCount[2 * i + parity] = count2;
    B2CShapeRef ref; ref.shapeId = shapeId; ref.generation = d.shapes[shapeId].generation;
    int2 c = sensorMerge<false>(o1, count1, o2, count2, ref, nullptr, nullptr);   // equal lists produce no events: the eventBits test (:118-127) is implied
    d.sensorScan[i] = make_int4(c.x, c.y, 0, 0);
  }
  grid.sync();
  gridScan<int4>(grid, d.sensorScan, d.scanBlockSums4, n, d.sensorScan + n);
  const int4 total = d.sensorScan[n];
  const int endBase = s.sensorEndCount[endIndex];        // DestroyShape / DestroyBody of a sensor since the last step appended here already
  const bool fits = total.x <= d.sensorEventCapacity && endBase + total.y <= d.sensorEventCapacity;
  if (fits) for (int i = tid; i < n; i += stride) {
    const int shapeId = d.sensorShape[i];
    const int4 off = addT(d.sensorScan[i], d.scanBlockSums4[i / (int)blockDim.x]);
    B2CShapeRef ref; ref.shapeId = shapeId; ref.generation = d.shapes[shapeId].generation;
    sensorMerge<true>(sensorList(d, i, 1 - parity), d.sensorOverlapCount[2 * i + 1 - parity], sensorList(d, i, parity), d.sensorOverlapCount[2 * i + parity], ref,
                      d.sensorBeginEvents + off.x, d.sensorEndEvents[endIndex] + endBase + off.y);
  }
  grid.sync();
  if (tid == 0) {
    if (fits) { s.sensorBeginCount = total.x; s.sensorEndCount[endIndex] = endBase + total.y; }
    else { s.sensorBeginCount = 0; s.errorFlags |= (int)ERR_SENSOR_CAPACITY; }
  }
}

}  // namespace b2c
