// ORACLE — TEST INFRASTRUCTURE ONLY (see b2o_math.h header). Parity unpinned by the reference.
// Part 1: world construction, body/shape/joint creation, mass data, setters.
#include "b2o_world.h"
#include <cassert>
#include <cstring>

namespace b2o {

static inline int ProxyType(int key) { return key & 3; }        // src/BroadPhase.cs:30-32
static inline int ProxyId(int key) { return key >> 2; }
static inline int ProxyKey(int id, int type) { return (id << 2) | type; }
static inline uint64_t ShapePairKey(uint64_t k1, uint64_t k2) { return k1 < k2 ? (k1 << 32 | k2) : (k2 << 32 | k1); }  // src/Contact.cs:126

World::World(const WorldDef& def) {      // src/World.cs:183-226, src/ConstraintGraph.cs:73-83
  gravity = def.gravity; hitEventThreshold = def.hitEventThreshold; restitutionThreshold = def.restitutionThreshold;
  maxLinearSpeed = def.maximumLinearSpeed; contactSpeed = def.contactSpeed; contactHertz = def.contactHertz;
  contactDampingRatio = def.contactDampingRatio; enableSleep = def.enableSleep;
  enableContactSoftening = def.enableContactSoftening; enableContinuous = def.enableContinuous;
  for (int i = 0; i < kGraphColorCount - 1; i++) colors[i].bodySet.SetBitCountAndClear(16);
  if (enableSleep) { counters.unsupported++; lastError = "oracle: enableSleep=true (islands/sleep) is not restated"; }
}

int World::CreateBody(const BodyDef& def) {     // api/Body.cs:16-105
  if (!def.isEnabled || !(def.isAwake || !def.enableSleep)) { counters.unsupported++; lastError = "oracle: disabled/asleep bodies not restated"; }
  int setId = def.type == kStaticBody ? kStaticSet : kAwakeSet;
  int bodyId = bodyIdPool.AllocId();
  uint32_t lockFlags = (def.lockLinearX ? kLockLinearX : 0) | (def.lockLinearY ? kLockLinearY : 0) | (def.lockAngularZ ? kLockAngularZ : 0) |
                       (def.isBullet ? kIsBullet : 0) | (def.allowFastRotation ? kAllowFastRotation : 0);
  SolverSet& set = setId == kStaticSet ? staticSet : awakeSet;
  BodySim sim;
  sim.transform.p = def.position; sim.transform.q = def.rotation; sim.center = def.position; sim.rotation0 = def.rotation;
  sim.center0 = def.position; sim.minExtent = kHuge; sim.maxExtent = 0; sim.linearDamping = def.linearDamping;
  sim.angularDamping = def.angularDamping; sim.gravityScale = def.gravityScale; sim.bodyId = bodyId;
  sim.flags = lockFlags | (def.type == kDynamicBody ? kDynamicFlag : 0);
  set.bodySims.push_back(sim);
  if (setId == kAwakeSet) {
    BodyState st; st.linearVelocity = def.linearVelocity; st.angularVelocity = def.angularVelocity; st.flags = sim.flags;
    set.bodyStates.push_back(st);
  }
  if (bodyId == (int)bodies.size()) bodies.emplace_back();
  Body& body = bodies[bodyId];
  uint16_t gen = body.generation;
  body = Body();
  body.generation = (uint16_t)(gen + 1);
  body.userData = def.userData; body.setIndex = setId; body.localIndex = (int)set.bodySims.size() - 1;
  body.id = bodyId; body.sleepThreshold = def.sleepThreshold; body.type = def.type; body.flags = sim.flags; body.enableSleep = def.enableSleep;
  return bodyId;
}

int World::BroadPhaseCreateProxy(int proxyType, const AABB& aabb, uint64_t categoryBits, int shapeIndex, bool forcePairCreation) {
  // src/BroadPhase.cs:71-79
  int proxyId = trees[proxyType].CreateProxy(aabb, categoryBits, (uint64_t)shapeIndex);
  int proxyKey = ProxyKey(proxyId, proxyType);
  if (proxyType != kStaticBody || forcePairCreation) BufferMove(proxyKey);
  return proxyKey;
}

void World::UpdateShapeAABBs(Shape& s, const Transform& xf, int proxyType) {    // src/Shape.cs:64-75
  AABB aabb = ComputeAABB(s.geom, xf);
  aabb.lo.x -= kSpeculativeDistance; aabb.lo.y -= kSpeculativeDistance; aabb.hi.x += kSpeculativeDistance; aabb.hi.y += kSpeculativeDistance;
  s.aabb = aabb;
  float margin = proxyType == kStaticBody ? kSpeculativeDistance : kAABBMargin;
  s.fatAABB = {{aabb.lo.x - margin, aabb.lo.y - margin}, {aabb.hi.x + margin, aabb.hi.y + margin}};
}

int World::CreateShape(int bodyId, const ShapeDef& def, const Geom& geom) {     // api/Shape.cs:8-24, src/Shape.cs:171-218
  if (geom.type == kChainSegment) { counters.unsupported++; lastError = "oracle: chain segments not restated"; }
  Body& body = bodies[bodyId];
  Transform transform = GetBodySim(body).transform;
  int shapeId = shapeIdPool.AllocId();
  if (shapeId == (int)shapes.size()) shapes.emplace_back();
  Shape& shape = shapes[shapeId];
  uint16_t gen = shape.generation;
  shape = Shape();
  shape.generation = (uint16_t)(gen + 1);
  shape.geom = geom; shape.id = shapeId; shape.bodyId = body.id; shape.density = def.density; shape.material = def.material;
  shape.filter = def.filter; shape.userData = def.userData; shape.enableSensorEvents = def.enableSensorEvents;
  shape.enableContactEvents = def.enableContactEvents; shape.enableCustomFiltering = def.enableCustomFiltering;
  shape.enableHitEvents = def.enableHitEvents; shape.enablePreSolveEvents = def.enablePreSolveEvents;
  shape.localCentroid = GetCentroid(geom);
  {
    int proxyType = body.type;     // src/Shape.cs:126-131
    UpdateShapeAABBs(shape, transform, proxyType);
    shape.proxyKey = BroadPhaseCreateProxy(proxyType, shape.fatAABB, shape.filter.categoryBits, shapeId, def.invokeContactCreation || def.isSensor);
  }
  if (body.headShapeId != -1) shapes[body.headShapeId].prevShapeId = shapeId;
  shape.prevShapeId = -1; shape.nextShapeId = body.headShapeId; body.headShapeId = shapeId; body.shapeCount++;
  if (def.isSensor) { shape.sensorIndex = (int)sensors.size(); sensors.emplace_back(); sensors.back().shapeId = shapeId; }   // src/Shape.cs:210-215
  else shape.sensorIndex = -1;
  if (def.updateBodyMass) UpdateBodyMassData(body);
  return shapeId;
}

void World::UpdateBodyMassData(Body& body) {      // src/Body.cs:255-353
  BodySim& bodySim = GetBodySim(body);
  body.mass = 0; body.inertia = 0; bodySim.invMass = 0; bodySim.invInertia = 0; bodySim.localCenter = {0, 0};
  bodySim.minExtent = kHuge; bodySim.maxExtent = 0;
  if (body.type != kDynamicBody) {
    bodySim.center = bodySim.transform.p; bodySim.center0 = bodySim.center;
    if (body.type == kKinematicBody) {
      int shapeId = body.headShapeId;
      while (shapeId != -1) {
        const Shape& s = shapes[shapeId];
        ShapeExtent extent = ComputeExtent(s.geom, {0, 0});
        bodySim.minExtent = netMin(bodySim.minExtent, extent.minExtent);
        bodySim.maxExtent = netMax(bodySim.maxExtent, extent.maxExtent);
        shapeId = s.nextShapeId;
      }
    }
    return;
  }
  int shapeCount = body.shapeCount;
  std::vector<MassData> masses(shapeCount);
  Vec2 localCenter = {0, 0};
  int shapeIndex = 0;
  int shapeId = body.headShapeId;
  while (shapeId != -1) {
    const Shape& s = shapes[shapeId];
    shapeId = s.nextShapeId;
    if (s.density == 0) { masses[shapeIndex] = MassData(); continue; }
    MassData massData = ComputeMass(s.geom, s.density);
    body.mass += massData.mass;
    localCenter = MulAdd(localCenter, massData.mass, massData.center);
    masses[shapeIndex] = massData;
    shapeIndex++;
  }
  if (body.mass > 0) { bodySim.invMass = 1 / body.mass; localCenter = bodySim.invMass * localCenter; }
  for (shapeIndex = 0; shapeIndex < shapeCount; shapeIndex++) {
    MassData massData = masses[shapeIndex];
    if (massData.mass == 0) continue;
    Vec2 offset = localCenter - massData.center;
    float inertia = massData.rotationalInertia + massData.mass * Dot(offset, offset);
    body.inertia += inertia;
  }
  if (body.inertia > 0) bodySim.invInertia = 1 / body.inertia; else { body.inertia = 0; bodySim.invInertia = 0; }
  Vec2 oldCenter = bodySim.center;
  bodySim.localCenter = localCenter;
  bodySim.center = TransformPoint(bodySim.transform, bodySim.localCenter);
  bodySim.center0 = bodySim.center;
  BodyState* state = GetBodyState(body);
  if (state) { Vec2 deltaLinear = CrossSV(state->angularVelocity, bodySim.center - oldCenter); state->linearVelocity = state->linearVelocity + deltaLinear; }
  shapeId = body.headShapeId;
  while (shapeId != -1) {
    const Shape& s = shapes[shapeId];
    ShapeExtent extent = ComputeExtent(s.geom, localCenter);
    bodySim.minExtent = netMin(bodySim.minExtent, extent.minExtent);
    bodySim.maxExtent = netMax(bodySim.maxExtent, extent.maxExtent);
    shapeId = s.nextShapeId;
  }
}

bool World::ShouldBodiesCollide(const Body& bodyA, const Body& bodyB) const {   // src/Body.cs:354-371
  if (bodyA.type != kDynamicBody && bodyB.type != kDynamicBody) return false;
  int jointKey, otherBodyId;
  if (bodyA.jointCount < bodyB.jointCount) { jointKey = bodyA.headJointKey; otherBodyId = bodyB.id; }
  else { jointKey = bodyB.headJointKey; otherBodyId = bodyA.id; }
  while (jointKey != -1) {
    int jointId = jointKey >> 1, edgeIndex = jointKey & 1, otherEdgeIndex = edgeIndex ^ 1;
    const Joint& joint = joints[jointId];
    if (!joint.collideConnected && (otherEdgeIndex == 1 ? joint.edge1.bodyId : joint.edge0.bodyId) == otherBodyId) return false;
    jointKey = edgeIndex == 1 ? joint.edge1.nextKey : joint.edge0.nextKey;
  }
  return true;
}

int World::AssignJointColor(int bodyIdA, int bodyIdB, int typeA, int typeB) {   // src/ConstraintGraph.cs:91-128
  if (typeA != kStaticBody && typeB != kStaticBody) {
    for (int i = 0; i < kDynamicColorCount; i++) {
      GraphColor& color = colors[i];
      if (color.bodySet.GetBit(bodyIdA) || color.bodySet.GetBit(bodyIdB)) continue;
      if (typeA == kDynamicBody) color.bodySet.SetBitGrow(bodyIdA);
      if (typeB == kDynamicBody) color.bodySet.SetBitGrow(bodyIdB);
      return i;
    }
  } else if (typeA == kDynamicBody) {
    for (int i = kGraphColorCount - 2; i >= 1; i--) {
      GraphColor& color = colors[i];
      if (color.bodySet.GetBit(bodyIdA)) continue;
      color.bodySet.SetBitGrow(bodyIdA); return i;
    }
  } else if (typeB == kDynamicBody) {
    for (int i = kGraphColorCount - 2; i >= 1; i--) {
      GraphColor& color = colors[i];
      if (color.bodySet.GetBit(bodyIdB)) continue;
      color.bodySet.SetBitGrow(bodyIdB); return i;
    }
  }
  return kOverflowIndex;
}

int World::CreateRevoluteJoint(const RevoluteJointDef& r) {      // api/Joint.cs:555 -> the generic path below
  JointDef def; def.type = kRevoluteJoint; def.bodyIdA = r.bodyIdA; def.bodyIdB = r.bodyIdB; def.localFrameA = r.localFrameA; def.localFrameB = r.localFrameB;
  def.forceThreshold = r.forceThreshold; def.torqueThreshold = r.torqueThreshold; def.constraintHertz = r.constraintHertz;
  def.constraintDampingRatio = r.constraintDampingRatio; def.collideConnected = r.collideConnected; def.userData = r.userData;
  def.target = r.targetAngle; def.enableSpring = r.enableSpring; def.hertz = r.hertz; def.dampingRatio = r.dampingRatio; def.enableLimit = r.enableLimit;
  def.lower = r.lowerAngle; def.upper = r.upperAngle; def.enableMotor = r.enableMotor; def.maxMotor = r.maxMotorTorque; def.motorSpeed = r.motorSpeed;
  return CreateJoint(def);
}

// World.CreateJoint src/Joint.cs:181-296 followed by the per-type Create: DistanceJoint.cs:35-63, FilterJoint.cs:7-17, MotorJoint.cs:33-52,
// PrismaticJoint.cs:33-53, RevoluteJoint.cs:32-56, WeldJoint.cs:24-39, WheelJoint.cs:33-56
int World::CreateJoint(const JointDef& def) {
  Body& bodyA = bodies[def.bodyIdA]; Body& bodyB = bodies[def.bodyIdB];
  int bodyIdA = bodyA.id, bodyIdB = bodyB.id;
  int jointId = jointIdPool.AllocId();
  if (jointId == (int)joints.size()) joints.emplace_back();
  Joint& joint = joints[jointId];
  uint16_t gen = joint.generation;
  joint = Joint();
  joint.generation = (uint16_t)(gen + 1);
  joint.jointId = jointId; joint.userData = def.userData; joint.type = def.type; joint.collideConnected = def.collideConnected;
  joint.edge0.bodyId = bodyIdA; joint.edge0.prevKey = -1; joint.edge0.nextKey = bodyA.headJointKey;
  int keyA = jointId << 1;
  if (bodyA.headJointKey != -1) {
    Joint& jointA = joints[bodyA.headJointKey >> 1];
    if (bodyA.headJointKey & 1) jointA.edge1.prevKey = keyA; else jointA.edge0.prevKey = keyA;
  }
  bodyA.headJointKey = keyA; bodyA.jointCount++;
  joint.edge1.bodyId = bodyIdB; joint.edge1.prevKey = -1; joint.edge1.nextKey = bodyB.headJointKey;
  int keyB = (jointId << 1) | 1;
  if (bodyB.headJointKey != -1) {
    Joint& jointB = joints[bodyB.headJointKey >> 1];
    if (bodyA.headJointKey & 1) jointB.edge1.prevKey = keyB; else jointB.edge0.prevKey = keyB;   // Q11: tests bodyA's key
  }
  bodyB.headJointKey = keyB; bodyB.jointCount++;
  const bool parked = bodyA.type != kDynamicBody && bodyB.type != kDynamicBody;   // src/Joint.cs:242-248: the static set holds it, nothing simulates it
  joint.setIndex = parked ? kStaticSet : kAwakeSet;
  int colorIndex = parked ? -1 : AssignJointColor(bodyIdA, bodyIdB, bodyA.type, bodyB.type);   // src/ConstraintGraph.cs:242-251
  JointSim sim; sim.jointId = jointId; sim.bodyIdA = bodyIdA; sim.bodyIdB = bodyIdB;
  sim.localFrameA = def.localFrameA; sim.localFrameB = def.localFrameB; sim.type = def.type;
  sim.constraintHertz = def.constraintHertz; sim.constraintDampingRatio = def.constraintDampingRatio;
  sim.constraintSoftness.biasRate = 0; sim.constraintSoftness.massScale = 1; sim.constraintSoftness.impulseScale = 0;
  sim.forceThreshold = def.forceThreshold; sim.torqueThreshold = def.torqueThreshold;
  switch (def.type) {
    case kRevoluteJoint:
      sim.targetAngle = netClamp(def.target, -kPi, kPi); sim.hertz = def.hertz; sim.dampingRatio = def.dampingRatio;
      sim.lowerAngle = def.lower; sim.upperAngle = def.upper; sim.maxMotorTorque = def.maxMotor; sim.motorSpeed = def.motorSpeed;
      sim.enableSpring = def.enableSpring; sim.enableLimit = def.enableLimit; sim.enableMotor = def.enableMotor;
      break;
    case kDistanceJoint:
      sim.length = netMax(def.length, kLinearSlop); sim.hertz = def.hertz; sim.dampingRatio = def.dampingRatio;
      sim.minLength = netMax(def.minLength, kLinearSlop); sim.maxLength = netMax(def.minLength, def.maxLength);
      sim.maxMotorTorque = def.maxMotor; sim.motorSpeed = def.motorSpeed; sim.enableSpring = def.enableSpring;
      sim.lowerSpringForce = def.lowerSpringForce; sim.upperSpringForce = def.upperSpringForce;
      sim.enableLimit = def.enableLimit; sim.enableMotor = def.enableMotor;
      break;
    case kFilterJoint: break;
    case kMotorJoint:
      sim.linearVelocity = def.linearVelocity; sim.maxVelocityForce = def.maxVelocityForce; sim.angularVelocity = def.angularVelocity;
      sim.maxVelocityTorque = def.maxVelocityTorque; sim.hertz = def.hertz; sim.dampingRatio = def.dampingRatio; sim.maxSpringForce = def.maxSpringForce;
      sim.angularHertz = def.angularHertz; sim.angularDampingRatio = def.angularDampingRatio; sim.maxSpringTorque = def.maxSpringTorque;
      break;
    case kPrismaticJoint:
      sim.hertz = def.hertz; sim.dampingRatio = def.dampingRatio; sim.targetAngle = def.target; sim.lowerAngle = def.lower; sim.upperAngle = def.upper;
      sim.maxMotorTorque = def.maxMotor; sim.motorSpeed = def.motorSpeed;
      sim.enableSpring = def.enableSpring; sim.enableLimit = def.enableLimit; sim.enableMotor = def.enableMotor;
      break;
    case kWeldJoint:
      sim.hertz = def.hertz; sim.dampingRatio = def.dampingRatio; sim.angularHertz = def.angularHertz; sim.angularDampingRatio = def.angularDampingRatio;
      break;
    case kWheelJoint:
      sim.lowerAngle = def.lower; sim.upperAngle = def.upper; sim.maxMotorTorque = def.maxMotor; sim.motorSpeed = def.motorSpeed;
      sim.hertz = def.hertz; sim.dampingRatio = def.dampingRatio;
      sim.enableSpring = def.enableSpring; sim.enableLimit = def.enableLimit; sim.enableMotor = def.enableMotor;
      break;
  }
  if (parked) { staticSet.jointSims.push_back(sim); joint.colorIndex = -1; joint.localIndex = (int)staticSet.jointSims.size() - 1; }
  else { colors[colorIndex].jointSims.push_back(sim); joint.colorIndex = colorIndex; joint.localIndex = (int)colors[colorIndex].jointSims.size() - 1; }
  if (!def.collideConnected) DestroyContactsBetweenBodies(bodyA, bodyB);
  return jointId;
}

// ---- joint setters, api/Joint.cs. `param` numbering = B2C_JP_* of include/b2cuda.h (the test harness drives both sides with it). ----
bool World::JointSetParam(int jointId, int param, float v) {
  JointSim& j = GetJointSim(joints[jointId]);
  const int t = j.type;
  const bool slm = t == kRevoluteJoint || t == kPrismaticJoint || t == kWheelJoint || t == kDistanceJoint;
  switch (param) {
    case 0: j.constraintHertz = v; return true;                 // JointAPI.SetConstraintTuning :130-138
    case 1: j.constraintDampingRatio = v; return true;
    case 2: j.forceThreshold = v; return true;                  // SetForceThreshold :148
    case 3: j.torqueThreshold = v; return true;                 // SetTorqueThreshold :159
    case 4: if (!(slm || t == kWeldJoint || t == kMotorJoint)) return false; j.hertz = v; return true;          // *_SetSpringHertz / *_SetLinearHertz
    case 5: if (!(slm || t == kWeldJoint || t == kMotorJoint)) return false; j.dampingRatio = v; return true;
    case 6: if (t != kRevoluteJoint && t != kPrismaticJoint) return false; j.targetAngle = v; return true;      // :444, :585
    case 7: if (!slm) return false; j.motorSpeed = v; return true;
    case 8: if (!slm) return false; j.maxMotorTorque = v; return true;
    case 9: if (t != kDistanceJoint) return false; j.length = netClamp(v, kLinearSlop, kHuge); j.lowerImpulse = 0; j.upperImpulse = 0; return true;   // :187-194
    case 10: if (t != kDistanceJoint) return false; j.lowerSpringForce = v; return true;                         // :204-210
    case 11: if (t != kDistanceJoint) return false; j.upperSpringForce = v; return true;
    case 12: if (t != kWeldJoint && t != kMotorJoint) return false; j.angularHertz = v; return true;
    case 13: if (t != kWeldJoint && t != kMotorJoint) return false; j.angularDampingRatio = v; return true;
    case 14: if (t != kMotorJoint) return false; j.linearVelocity.x = v; return true;                            // MotorJoint_SetLinearVelocity :327
    case 15: if (t != kMotorJoint) return false; j.linearVelocity.y = v; return true;
    case 16: if (t != kMotorJoint) return false; j.angularVelocity = v; return true;
    case 17: if (t != kMotorJoint) return false; j.maxVelocityForce = netMax(0, v); return true;                 // :333
    case 18: if (t != kMotorJoint) return false; j.maxVelocityTorque = v; return true;                           // :336
    case 19: if (t != kMotorJoint) return false; j.maxSpringForce = netMax(0, v); return true;
    case 20: if (t != kMotorJoint) return false; j.maxSpringTorque = netMax(0, v); return true;
  }
  return false;
}
bool World::JointEnable(int jointId, int feature, bool flag) {
  JointSim& j = GetJointSim(joints[jointId]);
  const int t = j.type;
  if (t != kRevoluteJoint && t != kPrismaticJoint && t != kWheelJoint && t != kDistanceJoint) return false;
  if (feature == 0) {           // *_EnableSpring: DistanceJoint :197 stores; Prismatic :432, Revolute :573, Wheel :735 clear springImpulse on a change
    if (t == kDistanceJoint) j.enableSpring = flag;
    else if (flag != j.enableSpring) { j.enableSpring = flag; j.springImpulse = 0; }
  } else if (feature == 1) {    // *_EnableLimit: DistanceJoint :229 stores; Prismatic :452, Revolute :613; Wheel :755 writes false when changed
    if (t == kDistanceJoint) j.enableLimit = flag;
    else if (flag != j.enableLimit) { j.lowerImpulse = 0; j.upperImpulse = 0; j.enableLimit = t == kWheelJoint ? false : flag; }
  } else if (feature == 2) {    // *_EnableMotor :277, :484, :639, :782
    if (flag != j.enableMotor) { j.enableMotor = flag; j.motorImpulse = 0; }
  } else return false;
  return true;
}
bool World::JointSetLimits(int jointId, float lower, float upper) {
  JointSim& j = GetJointSim(joints[jointId]);
  const int t = j.type;
  if (t == kDistanceJoint) {    // DistanceJoint_SetLengthRange :236-247
    float minLength = netClamp(lower, kLinearSlop, kHuge), maxLength = netClamp(upper, kLinearSlop, kHuge);
    j.minLength = netMin(minLength, maxLength); j.maxLength = netMax(minLength, maxLength);
    j.springImpulse = 0; j.lowerImpulse = 0; j.upperImpulse = 0;
    return true;
  }
  if (t != kRevoluteJoint && t != kPrismaticJoint && t != kWheelJoint) return false;
  if (lower != j.lowerAngle || upper != j.upperAngle) {          // :471-481, :626-637, :769-780
    j.lowerAngle = netMin(lower, upper); j.upperAngle = netMax(lower, upper);
    j.lowerImpulse = 0; j.upperImpulse = 0;
  }
  return true;
}

template <class T> static int RemoveSwap(std::vector<T>& a, int index) {   // include/Box2D.cs:9-19
  int movedIndex = -1;
  if (index != (int)a.size() - 1) { movedIndex = (int)a.size() - 1; a[index] = a[movedIndex]; }
  a.pop_back();
  return movedIndex;
}

// DestroyJointInternal src/Joint.cs:297-375 (wakeBodies: no sleeping sets here; islands are not restated)
void World::DestroyJoint(int jointId) {
  Joint& joint = joints[jointId];
  JointEdge edgeA = joint.edge0, edgeB = joint.edge1;
  Body& bodyA = bodies[edgeA.bodyId]; Body& bodyB = bodies[edgeB.bodyId];
  if (edgeA.prevKey != -1) { Joint& prev = joints[edgeA.prevKey >> 1]; if (edgeA.prevKey & 1) prev.edge1.nextKey = edgeA.nextKey; else prev.edge0.nextKey = edgeA.nextKey; }
  if (edgeA.nextKey != -1) { Joint& next = joints[edgeA.nextKey >> 1]; if (edgeA.nextKey & 1) next.edge1.prevKey = edgeA.prevKey; else next.edge0.prevKey = edgeA.prevKey; }
  int edgeKeyA = jointId << 1;
  if (bodyA.headJointKey == edgeKeyA) bodyA.headJointKey = edgeA.nextKey;
  bodyA.jointCount--;
  if (edgeB.prevKey != -1) { Joint& prev = joints[edgeB.prevKey >> 1]; if (edgeB.prevKey & 1) prev.edge1.nextKey = edgeB.nextKey; else prev.edge0.nextKey = edgeB.nextKey; }
  if (edgeB.nextKey != -1) { Joint& next = joints[edgeB.nextKey >> 1]; if (edgeB.nextKey & 1) next.edge1.prevKey = edgeB.prevKey; else next.edge0.prevKey = edgeB.prevKey; }
  int edgeKeyB = (jointId << 1) | 1;
  if (bodyB.headJointKey == edgeKeyB) bodyB.headJointKey = edgeB.nextKey;
  bodyB.jointCount--;
  if (joint.setIndex == kAwakeSet && joint.colorIndex != -1) {     // RemoveJointFromGraph src/ConstraintGraph.cs:256-276
    GraphColor& color = colors[joint.colorIndex];
    if (joint.colorIndex != kOverflowIndex) { color.bodySet.ClearBit(edgeA.bodyId); color.bodySet.ClearBit(edgeB.bodyId); }
    int movedIndex = RemoveSwap(color.jointSims, joint.localIndex);
    if (movedIndex != -1) joints[color.jointSims[joint.localIndex].jointId].localIndex = joint.localIndex;
  }
  else if (joint.setIndex == kStaticSet) {       // src/Joint.cs:356-366 the sim leaves its solver set
    int movedIndex = RemoveSwap(staticSet.jointSims, joint.localIndex);
    if (movedIndex != -1) joints[staticSet.jointSims[joint.localIndex].jointId].localIndex = joint.localIndex;
  }
  joint.setIndex = -1; joint.localIndex = -1; joint.colorIndex = -1; joint.jointId = -1;
  jointIdPool.FreeId(jointId);
}

void World::UnBufferMove(int proxyKey) {        // src/BroadPhase.cs:60-71
  if (moveSet.erase(proxyKey)) {
    for (size_t i = 0; i < moveArray.size(); i++) if (moveArray[i] == proxyKey) { moveArray[i] = moveArray.back(); moveArray.pop_back(); break; }
  }
}

void World::DestroyBody(int bodyId) {           // api/Body.cs:106-160 (awake or static set; no sensors, chains, islands)
  Body& body = bodies[bodyId];
  int edgeKey = body.headJointKey;
  while (edgeKey != -1) {
    int jointId = edgeKey >> 1, edgeIndex = edgeKey & 1;
    Joint& joint = joints[jointId];
    edgeKey = edgeIndex == 1 ? joint.edge1.nextKey : joint.edge0.nextKey;
    DestroyJoint(jointId);
  }
  edgeKey = body.headContactKey;                // DestroyBodyContacts src/Body.cs:232-244
  while (edgeKey != -1) {
    int contactId = edgeKey >> 1, edgeIndex = edgeKey & 1;
    Contact& contact = contacts[contactId];
    edgeKey = edgeIndex == 1 ? contact.edge1.nextKey : contact.edge0.nextKey;
    DestroyContact(contact);
  }
  int shapeId = body.headShapeId;
  while (shapeId != -1) {
    Shape& shape = shapes[shapeId];
    if (shape.sensorIndex != -1) DestroySensor(shape);      // api/Body.cs:127
    if (shape.proxyKey != -1) {                 // Shape.DestroyProxy src/Shape.cs:128-131 -> BroadPhase.DestroyProxy src/BroadPhase.cs:80-86
      UnBufferMove(shape.proxyKey);
      trees[ProxyType(shape.proxyKey)].DestroyProxy(ProxyId(shape.proxyKey));
      shape.proxyKey = -1;
    }
    shapeIdPool.FreeId(shapeId);
    shape.id = -1;
    shapeId = shape.nextShapeId;
  }
  if (body.setIndex == kAwakeSet) {
    int movedIndex = RemoveSwap(awakeSet.bodySims, body.localIndex);
    if (movedIndex != -1) bodies[awakeSet.bodySims[body.localIndex].bodyId].localIndex = body.localIndex;
    RemoveSwap(awakeSet.bodyStates, body.localIndex);
  } else {
    int movedIndex = RemoveSwap(staticSet.bodySims, body.localIndex);
    if (movedIndex != -1) bodies[staticSet.bodySims[body.localIndex].bodyId].localIndex = body.localIndex;
  }
  bodyIdPool.FreeId(body.id);
  body.setIndex = -1; body.localIndex = -1; body.id = -1;
}

void World::DestroyShape(int shapeId, bool updateBodyMass) {      // api/Shape.cs:48-56, DestroyShapeInternal src/Shape.cs:219-272 (no sensors)
  Shape& shape = shapes[shapeId];
  Body& body = bodies[shape.bodyId];
  if (shape.prevShapeId != -1) shapes[shape.prevShapeId].nextShapeId = shape.nextShapeId;
  if (shape.nextShapeId != -1) shapes[shape.nextShapeId].prevShapeId = shape.prevShapeId;
  if (shapeId == body.headShapeId) body.headShapeId = shape.nextShapeId;
  body.shapeCount--;
  if (shape.proxyKey != -1) {                    // Shape.DestroyProxy src/Shape.cs:128-131
    UnBufferMove(shape.proxyKey);
    trees[ProxyType(shape.proxyKey)].DestroyProxy(ProxyId(shape.proxyKey));
    shape.proxyKey = -1;
  }
  int contactKey = body.headContactKey;
  while (contactKey != -1) {
    int contactId = contactKey >> 1, edgeIndex = contactKey & 1;
    Contact& contact = contacts[contactId];
    contactKey = edgeIndex == 1 ? contact.edge1.nextKey : contact.edge0.nextKey;
    if (contact.shapeIdA == shapeId || contact.shapeIdB == shapeId) DestroyContact(contact);
  }
  if (shape.sensorIndex != -1) DestroySensor(shape);       // src/Shape.cs:245-268 (the same lines as DestroySensor)
  shapeIdPool.FreeId(shapeId);
  shape.id = -1;
  if (updateBodyMass) UpdateBodyMassData(body);
}
void World::ResetProxy(Shape& shape, bool destroyProxy) {         // src/Shape.cs:273-301
  Body& body = bodies[shape.bodyId];
  int shapeId = shape.id;
  int contactKey = body.headContactKey;
  while (contactKey != -1) {
    int contactId = contactKey >> 1, edgeIndex = contactKey & 1;
    Contact& contact = contacts[contactId];
    contactKey = edgeIndex == 1 ? contact.edge1.nextKey : contact.edge0.nextKey;
    if (contact.shapeIdA == shapeId || contact.shapeIdB == shapeId) DestroyContact(contact);
  }
  Transform transform = GetBodySim(body).transform;
  if (shape.proxyKey != -1) {
    int proxyType = ProxyType(shape.proxyKey);
    UpdateShapeAABBs(shape, transform, proxyType);
    if (destroyProxy) {
      UnBufferMove(shape.proxyKey);                // BroadPhase.DestroyProxy src/BroadPhase.cs:80-86
      trees[proxyType].DestroyProxy(ProxyId(shape.proxyKey));
      shape.proxyKey = BroadPhaseCreateProxy(proxyType, shape.fatAABB, shape.filter.categoryBits, shapeId, true);
    } else {
      trees[proxyType].MoveProxy(ProxyId(shape.proxyKey), shape.fatAABB); BufferMove(shape.proxyKey);   // src/BroadPhase.cs:89-95
    }
  } else UpdateShapeAABBs(shape, transform, body.type);
}
void World::SetFilter(int shapeId, Filter filter) {               // api/Shape.cs:157-166
  Shape& shape = shapes[shapeId];
  if (filter.maskBits == shape.filter.maskBits && filter.categoryBits == shape.filter.categoryBits && filter.groupIndex == shape.filter.groupIndex) return;
  bool destroyProxy = filter.categoryBits != shape.filter.categoryBits;
  shape.filter = filter;
  ResetProxy(shape, destroyProxy);
}
void World::SetGeometry(int shapeId, const Geom& geom) {          // api/Shape.cs:289-326
  Shape& shape = shapes[shapeId];
  shape.geom = geom;
  ResetProxy(shape, true);
}
void World::TransferJoint(bool toAwake, Joint& joint) {      // src/SolverSet.cs:377-420
  if (toAwake == (joint.setIndex == kAwakeSet)) return;
  int localIndex = joint.localIndex, colorIndex = joint.colorIndex;
  if (toAwake) {                                  // static set -> graph: AddJointToGraph / CreateJointInGraph src/ConstraintGraph.cs:242-255
    JointSim sim = staticSet.jointSims[localIndex];
    int bodyIdA = joint.edge0.bodyId, bodyIdB = joint.edge1.bodyId;
    int c = AssignJointColor(bodyIdA, bodyIdB, bodies[bodyIdA].type, bodies[bodyIdB].type);
    colors[c].jointSims.push_back(sim);
    joint.colorIndex = c; joint.localIndex = (int)colors[c].jointSims.size() - 1; joint.setIndex = kAwakeSet;
    int movedIndex = RemoveSwap(staticSet.jointSims, localIndex);
    if (movedIndex != -1) joints[staticSet.jointSims[localIndex].jointId].localIndex = localIndex;
  } else {                                        // graph -> static set, then RemoveJointFromGraph :256-276
    JointSim sim = colors[colorIndex].jointSims[localIndex];
    joint.setIndex = kStaticSet; joint.localIndex = (int)staticSet.jointSims.size(); joint.colorIndex = -1;
    staticSet.jointSims.push_back(sim);
    GraphColor& color = colors[colorIndex];
    if (colorIndex != kOverflowIndex) { color.bodySet.ClearBit(joint.edge0.bodyId); color.bodySet.ClearBit(joint.edge1.bodyId); }
    int movedIndex = RemoveSwap(color.jointSims, localIndex);
    if (movedIndex != -1) joints[color.jointSims[localIndex].jointId].localIndex = localIndex;
  }
}
int World::SetType(int bodyId, int type) {        // api/Body.cs:185-268
  Body& body = bodies[bodyId];
  int originalType = body.type;
  if (originalType == type) return 0;
  int edgeKey = body.headContactKey;              // DestroyBodyContacts src/Body.cs:232-244
  while (edgeKey != -1) {
    int contactId = edgeKey >> 1, edgeIndex = edgeKey & 1;
    Contact& contact = contacts[contactId];
    edgeKey = edgeIndex == 1 ? contact.edge1.nextKey : contact.edge0.nextKey;
    DestroyContact(contact);
  }
  for (int jointKey = body.headJointKey; jointKey != -1;) {      // :202-216 every joint of the body is parked in the static set
    Joint& joint = joints[jointKey >> 1];
    jointKey = (jointKey & 1) ? joint.edge1.nextKey : joint.edge0.nextKey;
    TransferJoint(false, joint);
  }
  body.type = type;
  if (type == kDynamicBody) body.flags |= kDynamicFlag; else body.flags &= ~(uint32_t)kDynamicFlag;
  SolverSet& sourceSet = body.setIndex == kStaticSet ? staticSet : awakeSet;
  const int targetSetIndex = type == kStaticBody ? kStaticSet : kAwakeSet;
  SolverSet& targetSet = type == kStaticBody ? staticSet : awakeSet;
  if (&targetSet != &sourceSet) {                 // SolverSet.TransferBody src/SolverSet.cs:354-376
    int sourceIndex = body.localIndex;
    BodySim targetSim = sourceSet.bodySims[sourceIndex];
    targetSim.flags &= ~(uint32_t)(kIsFast | kIsSpeedCapped | kHadTimeOfImpact);
    int targetIndex = (int)targetSet.bodySims.size();
    targetSet.bodySims.push_back(targetSim);
    int movedIndex = RemoveSwap(sourceSet.bodySims, sourceIndex);
    if (movedIndex != -1) bodies[sourceSet.bodySims[sourceIndex].bodyId].localIndex = sourceIndex;
    if (body.setIndex == kAwakeSet) RemoveSwap(sourceSet.bodyStates, sourceIndex);
    else { BodyState st; st.flags = body.flags; targetSet.bodyStates.push_back(st); }
    body.setIndex = targetSetIndex; body.localIndex = targetIndex;
  }
  for (int jointKey = body.headJointKey; jointKey != -1;) {      // :226-241 back into the graph when either body is dynamic now
    Joint& joint = joints[jointKey >> 1];
    jointKey = (jointKey & 1) ? joint.edge1.nextKey : joint.edge0.nextKey;
    if (bodies[joint.edge0.bodyId].type == kDynamicBody || bodies[joint.edge1.bodyId].type == kDynamicBody) TransferJoint(true, joint);
  }
  Transform transform = GetBodySim(body).transform;
  int shapeId = body.headShapeId;
  while (shapeId != -1) {
    Shape& shape = shapes[shapeId];
    shapeId = shape.nextShapeId;
    if (shape.proxyKey != -1) {                   // Shape.DestroyProxy src/Shape.cs:128-131
      UnBufferMove(shape.proxyKey);
      trees[ProxyType(shape.proxyKey)].DestroyProxy(ProxyId(shape.proxyKey));
      shape.proxyKey = -1;
    }
    UpdateShapeAABBs(shape, transform, type);     // Shape.CreateProxy src/Shape.cs:121-127
    shape.proxyKey = BroadPhaseCreateProxy(type, shape.fatAABB, shape.filter.categoryBits, shape.id, true);
  }
  UpdateBodyMassData(body);
  if (BodyState* state = GetBodyState(body)) state->flags = body.flags;
  return 0;
}
void World::SetDensity(int shapeId, float density, bool updateBodyMass) {      // api/Shape.cs:95-103
  Shape& shape = shapes[shapeId];
  if (density == shape.density) return;
  shape.density = density;
  if (updateBodyMass) UpdateBodyMassData(bodies[shape.bodyId]);
}
void World::ApplyMassFromShapes(int bodyId) { UpdateBodyMassData(bodies[bodyId]); }   // api/Body.cs:617-621

void World::JointSetLocalFrame(int jointId, int which, const Transform& frame) {   // api/Joint.cs:44,50: a plain store into the joint sim
  JointSim& j = GetJointSim(joints[jointId]);
  if (which == 0) j.localFrameA = frame; else j.localFrameB = frame;
}
void World::JointSetCollideConnected(int jointId, bool shouldCollide) {           // api/Joint.cs:56-75
  Joint& joint = joints[jointId];
  if (joint.collideConnected == shouldCollide) return;
  joint.collideConnected = shouldCollide;
  Body& bodyA = bodies[joint.edge0.bodyId]; Body& bodyB = bodies[joint.edge1.bodyId];
  if (shouldCollide) {
    int shapeId = bodyA.shapeCount < bodyB.shapeCount ? bodyA.headShapeId : bodyB.headShapeId;
    while (shapeId != -1) {
      Shape& shape = shapes[shapeId];
      if (shape.proxyKey != -1) BufferMove(shape.proxyKey);
      shapeId = shape.nextShapeId;
    }
  } else DestroyContactsBetweenBodies(bodyA, bodyB);
}

void World::DestroyContactsBetweenBodies(Body& bodyA, Body& bodyB) {    // src/Joint.cs:155-180
  int contactKey, otherBodyId;
  if (bodyA.contactCount < bodyB.contactCount) { contactKey = bodyA.headContactKey; otherBodyId = bodyB.id; }
  else { contactKey = bodyB.headContactKey; otherBodyId = bodyA.id; }
  while (contactKey != -1) {
    int contactId = contactKey >> 1, edgeIndex = contactKey & 1;
    Contact& contact = contacts[contactId];
    contactKey = edgeIndex == 1 ? contact.edge1.nextKey : contact.edge0.nextKey;
    int otherEdgeIndex = edgeIndex ^ 1;
    if ((otherEdgeIndex == 1 ? contact.edge1.bodyId : contact.edge0.bodyId) == otherBodyId) DestroyContact(contact);
  }
}

void World::SetTransform(int bodyId, Vec2 position, Rot rotation) {      // api/Body.cs:293-325
  Body& body = bodies[bodyId];
  BodySim& bodySim = GetBodySim(body);
  bodySim.transform.p = position; bodySim.transform.q = rotation;
  bodySim.center = TransformPoint(bodySim.transform, bodySim.localCenter);
  bodySim.rotation0 = bodySim.transform.q; bodySim.center0 = bodySim.center;
  Transform transform = bodySim.transform;
  int shapeId = body.headShapeId;
  while (shapeId != -1) {
    Shape& shape = shapes[shapeId];
    AABB aabb = ComputeAABB(shape.geom, transform);
    aabb.lo.x -= kSpeculativeDistance; aabb.lo.y -= kSpeculativeDistance; aabb.hi.x += kSpeculativeDistance; aabb.hi.y += kSpeculativeDistance;
    shape.aabb = aabb;
    if (!Contains(shape.fatAABB, aabb)) {
      AABB fat = {{aabb.lo.x - kAABBMargin, aabb.lo.y - kAABBMargin}, {aabb.hi.x + kAABBMargin, aabb.hi.y + kAABBMargin}};
      shape.fatAABB = fat;
      if (shape.proxyKey != -1) { trees[ProxyType(shape.proxyKey)].MoveProxy(ProxyId(shape.proxyKey), fat); BufferMove(shape.proxyKey); }   // src/BroadPhase.cs:89-95
    }
    shapeId = shape.nextShapeId;
  }
}
void World::SetTargetTransform(int bodyId, const Transform& target, float timeStep) {    // api/Body.cs:378-400 (awake set only: sleep is not restated)
  Body& body = bodies[bodyId];
  if (body.type == kStaticBody || timeStep <= 0) return;
  BodySim& sim = GetBodySim(body);
  Vec2 center1 = sim.center, center2 = TransformPoint(target, sim.localCenter);
  float invTimeStep = 1 / timeStep;
  Vec2 linearVelocity = invTimeStep * (center2 - center1);
  Rot q1 = sim.transform.q, q2 = target.q;
  float deltaAngle = atan2f(q1.c * q2.s - q1.s * q2.c, q1.c * q2.c + q1.s * q2.s);    // Rotation.RelativeAngle include/MathFunctions.cs:162
  float angularVelocity = invTimeStep * deltaAngle;
  if (BodyState* state = GetBodyState(body)) { state->linearVelocity = linearVelocity; state->angularVelocity = angularVelocity; }
}
void World::SetMassData(int bodyId, float mass, float rotationalInertia, Vec2 localCenter) {   // api/Body.cs:585-598
  Body& body = bodies[bodyId];
  BodySim& bodySim = GetBodySim(body);
  body.mass = mass; body.inertia = rotationalInertia;
  bodySim.localCenter = localCenter;
  Vec2 center = TransformPoint(bodySim.transform, localCenter);
  bodySim.center = center; bodySim.center0 = center;
  bodySim.invMass = body.mass > 0 ? 1 / body.mass : 0;
  bodySim.invInertia = body.inertia > 0 ? 1 / body.inertia : 0;
}
void World::SetMotionLocks(int bodyId, bool linearX, bool linearY, bool angularZ) {            // api/Body.cs:796-818
  uint32_t newFlags = (linearX ? kLockLinearX : 0u) | (linearY ? kLockLinearY : 0u) | (angularZ ? kLockAngularZ : 0u);
  Body& body = bodies[bodyId];
  const uint32_t allLocks = kLockLinearX | kLockLinearY | kLockAngularZ;
  if ((body.flags & allLocks) != newFlags) {
    body.flags &= ~allLocks;
    body.flags |= allLocks;                                 // as written (:807): Body.flags ends with every lock bit set
    BodySim& bodySim = GetBodySim(body);
    bodySim.flags &= ~allLocks;
    bodySim.flags |= newFlags;
    if (BodyState* state = GetBodyState(body)) {
      state->flags = bodySim.flags;
      state->linearVelocity = {linearX ? 0 : state->linearVelocity.x, linearY ? 0 : state->linearVelocity.y};
      state->angularVelocity = angularZ ? 0 : state->angularVelocity;
    }
  }
}
void World::SetSurfaceMaterial(int shapeId, float friction, float restitution, float rollingResistance, float tangentSpeed) {   // api/Shape.cs:110-146
  SurfaceMaterial& m = shapes[shapeId].material;
  m.friction = friction; m.restitution = restitution; m.rollingResistance = rollingResistance; m.tangentSpeed = tangentSpeed;
}
void World::SetLinearVelocity(int bodyId, Vec2 v) {       // api/Body.cs:359-367
  Body& body = bodies[bodyId]; if (body.type == kStaticBody) return;
  if (BodyState* s = GetBodyState(body)) s->linearVelocity = v;
}
void World::SetAngularVelocity(int bodyId, float w) {     // api/Body.cs:369-377
  Body& body = bodies[bodyId]; if (body.type == kStaticBody) return;
  if (BodyState* s = GetBodyState(body)) s->angularVelocity = w;
}
void World::ApplyForceToCenter(int bodyId, Vec2 force) {  // api/Body.cs:457-467
  Body& body = bodies[bodyId]; if (body.type != kDynamicBody) return;
  if (body.setIndex == kAwakeSet) { BodySim& sim = GetBodySim(body); sim.force = sim.force + force; }
}
void World::ApplyTorque(int bodyId, float torque) {       // api/Body.cs:469-479 (assigns, port quirk)
  Body& body = bodies[bodyId]; if (body.type != kDynamicBody) return;
  if (body.setIndex == kAwakeSet) GetBodySim(body).torque = torque;
}
void World::ApplyLinearImpulseToCenter(int bodyId, Vec2 impulse) {   // api/Body.cs:531-545, src/Body.cs:131-135
  Body& body = bodies[bodyId]; if (body.type != kDynamicBody) return;
  if (body.setIndex == kAwakeSet) {
    BodyState& st = awakeSet.bodyStates[body.localIndex]; const BodySim& sim = awakeSet.bodySims[body.localIndex];
    st.linearVelocity = MulAdd(st.linearVelocity, sim.invMass, impulse);
    float v2 = LengthSquared(st.linearVelocity);
    if (v2 > maxLinearSpeed * maxLinearSpeed) st.linearVelocity = st.linearVelocity * (maxLinearSpeed / sqrtf(v2));
  }
}

void World::ApplyForce(int bodyId, Vec2 force, Vec2 point) {   // api/Body.cs:439-451
  Body& body = bodies[bodyId]; if (body.type != kDynamicBody) return;
  if (body.setIndex == kAwakeSet) {
    BodySim& sim = GetBodySim(body);
    sim.force = sim.force + force;
    sim.torque += Cross(point - sim.center, force);
  }
}
void World::ClearForces(int bodyId) {                          // api/Body.cs:490-495
  BodySim& sim = GetBodySim(bodies[bodyId]); sim.force = {0, 0}; sim.torque = 0;
}
void World::ApplyLinearImpulse(int bodyId, Vec2 impulse, Vec2 point) {   // api/Body.cs:507-523
  Body& body = bodies[bodyId]; if (body.type != kDynamicBody) return;
  if (body.setIndex == kAwakeSet) {
    BodyState& st = awakeSet.bodyStates[body.localIndex]; const BodySim& sim = awakeSet.bodySims[body.localIndex];
    st.linearVelocity = MulAdd(st.linearVelocity, sim.invMass, impulse);
    st.angularVelocity += sim.invInertia * Cross(point - sim.center, impulse);
    float v2 = LengthSquared(st.linearVelocity);
    if (v2 > maxLinearSpeed * maxLinearSpeed) st.linearVelocity = st.linearVelocity * (maxLinearSpeed / sqrtf(v2));
  }
}
void World::ApplyAngularImpulse(int bodyId, float impulse) {   // api/Body.cs:554-568
  Body& body = bodies[bodyId]; if (body.type != kDynamicBody) return;
  if (body.setIndex == kAwakeSet) {
    BodyState& st = awakeSet.bodyStates[body.localIndex]; const BodySim& sim = awakeSet.bodySims[body.localIndex];
    st.angularVelocity += sim.invInertia * impulse;
  }
}
void World::SetLinearDamping(int bodyId, float v) { GetBodySim(bodies[bodyId]).linearDamping = v; }     // api/Body.cs:624-629
void World::SetAngularDamping(int bodyId, float v) { GetBodySim(bodies[bodyId]).angularDamping = v; }   // api/Body.cs:635-640
void World::SetGravityScale(int bodyId, float v) { GetBodySim(bodies[bodyId]).gravityScale = v; }       // api/Body.cs:647-653

#include "b2o_distance.inc"
#include "b2o_sensor.inc"
#include "b2o_continuous.inc"

}  // namespace b2o
