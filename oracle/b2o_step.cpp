// ORACLE — TEST INFRASTRUCTURE ONLY (see b2o_math.h header). Parity unpinned by the reference.
// Part 4: World.Solve driver (src/Solver.cs:739-1322), FinalizeBodies (:402-503),
// SolveContinuous no-hit path (:315-401), WorldAPI.Step (api/World.cs:31-113).
#include "b2o_world.h"
#include <chrono>
static inline double NowMs() { return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now().time_since_epoch()).count(); }

namespace b2o {

static inline int ProxyType(int key) { return key & 3; }
static inline int ProxyId(int key) { return key >> 2; }
static bool ShouldShapesCollideF(const Filter& a, const Filter& b) {
  if (a.groupIndex == b.groupIndex && a.groupIndex != 0) return a.groupIndex > 0;
  return (a.maskBits & b.categoryBits) != 0 && (a.categoryBits & b.maskBits) != 0;
}

// World::SolveContinuous lives in b2o_continuous.inc (it needs the GJK code of b2o_distance.inc).

void World::FinalizeBodies(StepContext& stepContext, BitSet& enlargedSimBitSet) {     // :402-503
  int n = (int)awakeSet.bodySims.size();
  for (int simIndex = 0; simIndex < n; simIndex++) {
    BodyState& state = awakeSet.bodyStates[simIndex];
    BodySim& sim = awakeSet.bodySims[simIndex];
    if (state.flags & kLockLinearX) state.linearVelocity.x = 0;
    if (state.flags & kLockLinearY) state.linearVelocity.y = 0;
    if (state.flags & kLockAngularZ) state.angularVelocity = 0;
    Vec2 v = state.linearVelocity; float w = state.angularVelocity;
    sim.center = sim.center + state.deltaPosition;
    sim.transform.q = NormalizeRot(MulRot(state.deltaRotation, sim.transform.q));
    float maxVelocity = Length(v) + fabsf(w) * sim.maxExtent;
    float maxDeltaPosition = Length(state.deltaPosition) + fabsf(state.deltaRotation.s) * sim.maxExtent;
    float sleepVelocity = netMax(maxVelocity, 0.5f * stepContext.inv_dt * maxDeltaPosition);
    state.deltaPosition = {0, 0}; state.deltaRotation = Rot();
    sim.transform.p = sim.center - RotateVector(sim.transform.q, sim.localCenter);
    Body& body = bodies[sim.bodyId];
    sim.force = {0, 0}; sim.torque = 0;
    body.flags &= ~(kIsFast | kIsSpeedCapped | kHadTimeOfImpact);
    body.flags |= (sim.flags & (kIsSpeedCapped | kHadTimeOfImpact));
    sim.flags &= ~(kIsFast | kIsSpeedCapped | kHadTimeOfImpact);
    if (!enableSleep || !body.enableSleep || sleepVelocity > body.sleepThreshold) {
      body.sleepTime = 0;
      if (body.type == kDynamicBody && enableContinuous && maxVelocity * stepContext.dt > 0.5f * sim.minExtent) {
        sim.flags |= kIsFast;
        if (sim.flags & kIsBullet) bulletBodies.push_back(simIndex);     // :447-449 (one worker: ascending sim index)
        else SolveContinuous(simIndex);
      } else { sim.center0 = sim.center; sim.rotation0 = sim.transform.q; }
    } else { sim.center0 = sim.center; sim.rotation0 = sim.transform.q; body.sleepTime += stepContext.dt; }
    Transform transform = sim.transform;
    bool isFast = (sim.flags & kIsFast) != 0;
    int shapeId = body.headShapeId;
    while (shapeId != -1) {
      Shape& shape = shapes[shapeId];
      if (isFast) enlargedSimBitSet.SetBit(simIndex);
      else {
        AABB aabb = ComputeAABB(shape.geom, transform);
        aabb.lo.x -= kSpeculativeDistance; aabb.lo.y -= kSpeculativeDistance; aabb.hi.x += kSpeculativeDistance; aabb.hi.y += kSpeculativeDistance;
        shape.aabb = aabb;
        if (!Contains(shape.fatAABB, aabb)) {
          shape.fatAABB = {{aabb.lo.x - kAABBMargin, aabb.lo.y - kAABBMargin}, {aabb.hi.x + kAABBMargin, aabb.hi.y + kAABBMargin}};
          shape.enlargedAABB = true;
          enlargedSimBitSet.SetBit(simIndex);
        }
      }
      shapeId = shape.nextShapeId;
    }
  }
}

void World::Solve(StepContext& stepContext) {      // :739-1322
  stepIndex++;
  int awakeBodyCount = (int)awakeSet.bodySims.size();
  if (awakeBodyCount == 0) return;
  // stage preparation :764-871 — active colors, SIMD-padded contact pointer array, joint array
  int c = 0, simdContactCount = 0, awakeJointCount = 0;
  for (int i = 0; i < kGraphColorCount - 1; i++) {
    int colorContactCount = (int)colors[i].contactSims.size(), colorJointCount = (int)colors[i].jointSims.size();
    if (colorContactCount + colorJointCount > 0) {
      stepContext.activeColorIndices[c] = i;
      simdContactCount += colorContactCount > 0 ? ((colorContactCount - 1) >> 3) + 1 : 0;
      awakeJointCount += colorJointCount;
      c++;
    }
  }
  stepContext.activeColorCount = c;
  stepContext.contacts.assign((size_t)8 * simdContactCount, nullptr);
  stepContext.lanes.assign((size_t)8 * simdContactCount, ConstraintLane());
  stepContext.joints.clear(); stepContext.joints.reserve(awakeJointCount);
  {
    int contactBase = 0;
    for (int i = 0; i < c; i++) {
      GraphColor& color = colors[stepContext.activeColorIndices[i]];
      int colorContactCount = (int)color.contactSims.size();
      color.simdBase = 8 * contactBase;
      if (colorContactCount > 0) {
        for (int k = 0; k < colorContactCount; ++k) stepContext.contacts[8 * contactBase + k] = &color.contactSims[k];
        contactBase += ((colorContactCount - 1) >> 3) + 1;
      }
      for (JointSim& j : color.jointSims) stepContext.joints.push_back(&j);
    }
  }
  double ts0 = NowMs();
  SolverStages(stepContext);
  double ts1 = NowMs();
  profileMs[4] += ts1 - ts0;
  BitSet enlargedSimBitSet; enlargedSimBitSet.SetBitCountAndClear(awakeBodyCount);
  FinalizeBodies(stepContext, enlargedSimBitSet);
  // hit events :1151-1190
  for (int i = 0; i < kGraphColorCount; i++) {
    for (const ContactSim& contactSim : colors[i].contactSims) {
      if (!(contactSim.simFlags & kSimEnableHitEvent)) continue;
      ContactHitEvent ev; ev.approachSpeed = hitEventThreshold; ev.point = {0, 0};
      bool hit = false;
      if (contactSim.manifold.pointCount > 0) {
        const ManifoldPoint& mp = contactSim.manifold.point0;
        float approachSpeed = -mp.normalVelocity;
        if (approachSpeed > ev.approachSpeed && mp.totalNormalImpulse > 0) { ev.approachSpeed = approachSpeed; ev.point = mp.point; hit = true; }
      }
      if (contactSim.manifold.pointCount > 1) {
        const ManifoldPoint& mp = contactSim.manifold.point1;
        float approachSpeed = -mp.normalVelocity;
        if (approachSpeed > ev.approachSpeed && mp.totalNormalImpulse > 0) { ev.approachSpeed = approachSpeed; ev.point = mp.point; hit = true; }
      }
      if (hit) { ev.normal = contactSim.manifold.normal; ev.shapeIdA = contactSim.shapeIdA; ev.shapeIdB = contactSim.shapeIdB; contactHitEvents.push_back(ev); }
    }
  }
  // refit :1191-1237 — enlarged bodies in ascending sim index, shapes in body list order
  for (uint32_t k = 0; k < enlargedSimBitSet.blockCount; k++) {
    uint64_t word = enlargedSimBitSet.bits[k];
    while (word != 0) {
      int bodySimIndex = (int)(64 * k + __builtin_ctzll(word));
      BodySim& bodySim = awakeSet.bodySims[bodySimIndex];
      Body& body = bodies[bodySim.bodyId];
      int shapeId = body.headShapeId;
      if ((bodySim.flags & (kIsBullet | kIsFast)) == (kIsBullet | kIsFast)) {     // :1210-1218 a fast bullet: its proxies are enlarged after the bullet stage
        while (shapeId != -1) { Shape& shape = shapes[shapeId]; BufferMove(shape.proxyKey); shapeId = shape.nextShapeId; }
      } else while (shapeId != -1) {
        Shape& shape = shapes[shapeId];
        if (shape.enlargedAABB) {
          trees[ProxyType(shape.proxyKey)].EnlargeProxy(ProxyId(shape.proxyKey), shape.fatAABB);   // src/BroadPhase.cs:96-104
          BufferMove(shape.proxyKey);
          shape.enlargedAABB = false;
        }
        shapeId = shape.nextShapeId;
      }
      word &= word - 1;
    }
  }
  // bullet stage :1238-1272: continuous collision of the fast bullets against all three trees, then their proxies are enlarged (no BufferMove:
  // the refit above put them in the move set)
  for (int simIndex : bulletBodies) SolveContinuous(simIndex);
  for (int simIndex : bulletBodies) {
    BodySim& bulletBodySim = awakeSet.bodySims[simIndex];
    if (!(bulletBodySim.flags & kEnlargeBounds)) continue;
    bulletBodySim.flags &= ~kEnlargeBounds;
    Body& bulletBody = bodies[bulletBodySim.bodyId];
    int shapeId = bulletBody.headShapeId;
    while (shapeId != -1) {
      Shape& shape = shapes[shapeId];
      if (shape.enlargedAABB) { shape.enlargedAABB = false; trees[kDynamicBody].EnlargeProxy(ProxyId(shape.proxyKey), shape.fatAABB); }
      shapeId = shape.nextShapeId;
    }
  }
  bulletBodies.clear();
}

void World::Step(float timeStep, int subStepCount) {      // api/World.cs:31-113
  lastStepDtZero = timeStep == 0.0f;                        // :38 bodyMoveEvents.Clear(); :44-51 early return
  sensorBeginEvents.clear(); contactBeginEvents.clear(); contactHitEvents.clear(); jointEvents.clear();
  if (timeStep == 0) { endEventArrayIndex = 1 - endEventArrayIndex; sensorEndEvents[endEventArrayIndex].clear(); contactEndEvents[endEventArrayIndex].clear(); return; }
  double t0 = NowMs();
  UpdateBroadPhasePairs();
  double t1 = NowMs();
  StepContext context;
  context.dt = timeStep; context.subStepCount = subStepCount > 1 ? subStepCount : 1;
  if (timeStep > 0) { context.inv_dt = 1 / timeStep; context.h = timeStep / context.subStepCount; context.inv_h = context.subStepCount * context.inv_dt; }
  else { context.inv_dt = 0; context.h = 0; context.inv_h = 0; }
  inv_h = context.inv_h; inv_dt = context.inv_dt;
  float hz = netMin(contactHertz, 0.125f * context.inv_h);
  context.contactSoftness = Softness(hz, contactDampingRatio, context.h);
  context.staticSoftness = Softness(2 * hz, contactDampingRatio, context.h);
  context.restitutionThreshold = restitutionThreshold; context.maxLinearVelocity = maxLinearSpeed; context.enableWarmStarting = enableWarmStarting;
  Collide(context);
  double t2 = NowMs();
  if (context.dt > 0) Solve(context);
  double t3 = NowMs();
  profileMs[0] += t3 - t0; profileMs[1] += t1 - t0; profileMs[2] += t2 - t1; profileMs[3] += t3 - t2;
  OverlapSensors();                                         // :97-102
  endEventArrayIndex = 1 - endEventArrayIndex;
  sensorEndEvents[endEventArrayIndex].clear(); contactEndEvents[endEventArrayIndex].clear();
}

}  // namespace b2o
