// ORACLE — TEST INFRASTRUCTURE ONLY (see b2o_math.h header). Parity unpinned by the reference.
// Part 3: solver stage pipeline. The contact solver follows the AVX path of the reference
// (src/ContactSolver.cs:362-1026) lane by lane; SIMD padding and 8-lane grouping are emulated.
#include "b2o_world.h"

namespace b2o {

struct BodyW { Vec2 v; float w; uint32_t flags; Vec2 dp; Rot dq; };
static inline BodyW Gather(const std::vector<BodyState>& states, int index) {    // src/ContactSolver.cs:415-451
  if (index == -1) return {{0, 0}, 0, 0, {0, 0}, {1, 0}};
  const BodyState& s = states[index];
  return {s.linearVelocity, s.angularVelocity, s.flags, s.deltaPosition, s.deltaRotation};
}
static inline void Scatter(std::vector<BodyState>& states, int index, const BodyW& b) {   // :452-479
  if (index != -1 && (states[index].flags & kDynamicFlag)) {
    BodyState& s = states[index];
    s.linearVelocity = b.v; s.angularVelocity = b.w; s.flags = b.flags; s.deltaPosition = b.dp; s.deltaRotation = b.dq;
  }
}

static void PrepareLane(World& world, StepContext& context, ConstraintLane& c, const ContactSim* contactSim) {   // :480-685
  c = ConstraintLane();
  if (!contactSim) return;
  const auto& awakeStates = world.awakeSet.bodyStates;
  float warmStartScale = world.enableWarmStarting ? 1.0f : 0.0f;
  Manifold manifold = contactSim->manifold;
  int indexA = contactSim->bodySimIndexA, indexB = contactSim->bodySimIndexB;
  c.indexA = indexA; c.indexB = indexB;
  Vec2 vA = {0, 0}; float wA = 0, mA = contactSim->invMassA, iA = contactSim->invIA;
  if (indexA != -1) { vA = awakeStates[indexA].linearVelocity; wA = awakeStates[indexA].angularVelocity; }
  Vec2 vB = {0, 0}; float wB = 0, mB = contactSim->invMassB, iB = contactSim->invIB;
  if (indexB != -1) { vB = awakeStates[indexB].linearVelocity; wB = awakeStates[indexB].angularVelocity; }
  c.invMassA = mA; c.invMassB = mB; c.invIA = iA; c.invIB = iB;
  { float k = iA + iB; c.rollingMass = k > 0.0f ? 1.0f / k : 0.0f; }
  Softness soft = context.contactSoftness;
  if (indexA == -1 || indexB == -1) soft = context.staticSoftness;
  else if (world.enableContactSoftening) {
    float contactHertz = netMin(world.contactHertz, 0.125f * context.inv_h);
    float ratio = 1;
    if (mA < mB) ratio = netMax(0.5f, mA / mB); else if (mB < mA) ratio = netMax(0.5f, mB / mA);
    soft = Softness(ratio * contactHertz, ratio * world.contactDampingRatio, context.h);
  }
  Vec2 normal = manifold.normal;
  c.normal = normal;
  c.friction = contactSim->friction; c.tangentSpeed = contactSim->tangentSpeed; c.restitution = contactSim->restitution;
  c.rollingResistance = contactSim->rollingResistance; c.rollingImpulse = warmStartScale * manifold.rollingImpulse;
  c.biasRate = soft.biasRate; c.massScale = soft.massScale; c.impulseScale = soft.impulseScale;
  Vec2 tangent = RightPerp(normal);
  {
    const ManifoldPoint& mp = manifold.point0;
    Vec2 rA = mp.anchorA, rB = mp.anchorB;
    c.anchorA1 = rA; c.anchorB1 = rB;
    c.baseSeparation1 = mp.separation - Dot(rB - rA, normal);
    c.normalImpulse1 = warmStartScale * mp.normalImpulse; c.tangentImpulse1 = warmStartScale * mp.tangentImpulse; c.totalNormalImpulse1 = 0;
    float rnA = Cross(rA, normal), rnB = Cross(rB, normal);
    float kNormal = mA + mB + iA * rnA * rnA + iB * rnB * rnB;
    c.normalMass1 = kNormal > 0.0f ? 1.0f / kNormal : 0.0f;
    float rtA = Cross(rA, tangent), rtB = Cross(rB, tangent);
    float kTangent = mA + mB + iA * rtA * rtA + iB * rtB * rtB;
    c.tangentMass1 = kTangent > 0.0f ? 1.0f / kTangent : 0.0f;
    Vec2 vrA = vA + CrossSV(wA, rA), vrB = vB + CrossSV(wB, rB);
    c.relativeVelocity1 = Dot(normal, vrB - vrA);
  }
  if (manifold.pointCount == 2) {
    const ManifoldPoint& mp = manifold.point1;
    Vec2 rA = mp.anchorA, rB = mp.anchorB;
    c.anchorA2 = rA; c.anchorB2 = rB;
    c.baseSeparation2 = mp.separation - Dot(rB - rA, normal);
    c.normalImpulse2 = warmStartScale * mp.normalImpulse; c.tangentImpulse2 = warmStartScale * mp.tangentImpulse; c.totalNormalImpulse2 = 0;
    float rnA = Cross(rA, normal), rnB = Cross(rB, normal);
    float kNormal = mA + mB + iA * rnA * rnA + iB * rnB * rnB;
    c.normalMass2 = kNormal > 0.0f ? 1.0f / kNormal : 0.0f;
    float rtA = Cross(rA, tangent), rtB = Cross(rB, tangent);
    float kTangent = mA + mB + iA * rtA * rtA + iB * rtB * rtB;
    c.tangentMass2 = kTangent > 0.0f ? 1.0f / kTangent : 0.0f;
    Vec2 vrA = vA + CrossSV(wA, rA), vrB = vB + CrossSV(wB, rB);
    c.relativeVelocity2 = Dot(normal, vrB - vrA);
  }
}

// applies impulse P at anchors (AVX op order, :794-805)
static inline void ApplyP(BodyW& bA, BodyW& bB, const ConstraintLane& c, Vec2 rA, Vec2 rB, float Px, float Py) {
  bA.v.x = bA.v.x - c.invMassA * Px; bA.v.y = bA.v.y - c.invMassA * Py;
  bA.w = bA.w - c.invIA * (rA.x * Py - rA.y * Px);
  bB.v.x = bB.v.x + c.invMassB * Px; bB.v.y = bB.v.y + c.invMassB * Py;
  bB.w = bB.w + c.invIB * (rB.x * Py - rB.y * Px);
}

static void WarmStartLane(std::vector<BodyState>& states, ConstraintLane& c) {      // :686-744
  BodyW bA = Gather(states, c.indexA), bB = Gather(states, c.indexB);
  float tangentX = c.normal.y, tangentY = 0.0f - c.normal.x;
  {
    Vec2 rA = c.anchorA1, rB = c.anchorB1;
    float Px = c.normalImpulse1 * c.normal.x + c.tangentImpulse1 * tangentX;
    float Py = c.normalImpulse1 * c.normal.y + c.tangentImpulse1 * tangentY;
    bA.w = bA.w - c.invIA * (rA.x * Py - rA.y * Px);
    bA.v.x = bA.v.x - c.invMassA * Px; bA.v.y = bA.v.y - c.invMassA * Py;
    bB.w = bB.w + c.invIB * (rB.x * Py - rB.y * Px);
    bB.v.x = bB.v.x + c.invMassB * Px; bB.v.y = bB.v.y + c.invMassB * Py;
  }
  {
    Vec2 rA = c.anchorA2, rB = c.anchorB2;
    float Px = c.normalImpulse2 * c.normal.x + c.tangentImpulse2 * tangentX;
    float Py = c.normalImpulse2 * c.normal.y + c.tangentImpulse2 * tangentY;
    bA.w = bA.w - c.invIA * (rA.x * Py - rA.y * Px);
    bA.v.x = bA.v.x - c.invMassA * Px; bA.v.y = bA.v.y - c.invMassA * Py;
    bB.w = bB.w + c.invIB * (rB.x * Py - rB.y * Px);
    bB.v.x = bB.v.x + c.invMassB * Px; bB.v.y = bB.v.y + c.invMassB * Py;
  }
  bA.w = bA.w - c.invIA * c.rollingImpulse;
  bB.w = bB.w + c.invIB * c.rollingImpulse;
  Scatter(states, c.indexA, bA); Scatter(states, c.indexB, bB);
}

static inline void SolveNormalPoint(BodyW& bA, BodyW& bB, const ConstraintLane& c, Vec2 dp, Vec2 rA, Vec2 rB, float baseSeparation,
                                    float normalMass, float& normalImpulse, float& totalNormalImpulseP, float& totalNormalImpulse,
                                    float inv_h, float contactSpeed, float biasRate, float massScale, float impulseScale) {   // :773-806
  Vec2 rsA = RotateVector(bA.dq, rA), rsB = RotateVector(bB.dq, rB);
  float dsx = dp.x + (rsB.x - rsA.x), dsy = dp.y + (rsB.y - rsA.y);
  float s = (c.normal.x * dsx + c.normal.y * dsy) + baseSeparation;
  bool mask = s > 0.0f;
  float specBias = s * inv_h, softBias = avxMax(biasRate * s, contactSpeed);
  float bias = mask ? specBias : softBias;
  float pointMassScale = mask ? 1.0f : massScale;
  float pointImpulseScale = mask ? 0.0f : impulseScale;
  float dvx = (bB.v.x - bB.w * rB.y) - (bA.v.x - bA.w * rA.y);
  float dvy = (bB.v.y + bB.w * rB.x) - (bA.v.y + bA.w * rA.x);
  float vn = dvx * c.normal.x + dvy * c.normal.y;
  float negImpulse = normalMass * (pointMassScale * vn + bias) + pointImpulseScale * normalImpulse;
  float newImpulse = avxMax(normalImpulse - negImpulse, 0.0f);
  float impulse = newImpulse - normalImpulse;
  normalImpulse = newImpulse;
  totalNormalImpulseP = totalNormalImpulseP + newImpulse;
  totalNormalImpulse = totalNormalImpulse + newImpulse;
  float Px = impulse * c.normal.x, Py = impulse * c.normal.y;
  ApplyP(bA, bB, c, rA, rB, Px, Py);
}
static inline void SolveFrictionPoint(BodyW& bA, BodyW& bB, const ConstraintLane& c, Vec2 rA, Vec2 rB, float tangentX, float tangentY,
                                      float tangentMass, float normalImpulse, float& tangentImpulse) {    // :844-871
  float dvx = (bB.v.x - bB.w * rB.y) - (bA.v.x - bA.w * rA.y);
  float dvy = (bB.v.y + bB.w * rB.x) - (bA.v.y + bA.w * rA.x);
  float vt = dvx * tangentX + dvy * tangentY;
  vt = vt - c.tangentSpeed;
  float negImpulse = tangentMass * vt;
  float maxFriction = c.friction * normalImpulse;
  float newImpulse = tangentImpulse - negImpulse;
  newImpulse = avxMax(0.0f - maxFriction, avxMin(newImpulse, maxFriction));
  float impulse = newImpulse - tangentImpulse;
  tangentImpulse = newImpulse;
  float Px = impulse * tangentX, Py = impulse * tangentY;
  ApplyP(bA, bB, c, rA, rB, Px, Py);
}

static void SolveLane(std::vector<BodyState>& states, ConstraintLane& c, float inv_h, float worldContactSpeed, bool useBias) {   // :745-912
  float contactSpeed = -worldContactSpeed;
  BodyW bA = Gather(states, c.indexA), bB = Gather(states, c.indexB);
  float biasRate, massScale, impulseScale;
  if (useBias) { biasRate = c.massScale * c.biasRate; massScale = c.massScale; impulseScale = c.impulseScale; }
  else { biasRate = 0; massScale = 1; impulseScale = 0; }
  float totalNormalImpulse = 0;
  Vec2 dp = {bB.dp.x - bA.dp.x, bB.dp.y - bA.dp.y};
  SolveNormalPoint(bA, bB, c, dp, c.anchorA1, c.anchorB1, c.baseSeparation1, c.normalMass1, c.normalImpulse1, c.totalNormalImpulse1,
                   totalNormalImpulse, inv_h, contactSpeed, biasRate, massScale, impulseScale);
  SolveNormalPoint(bA, bB, c, dp, c.anchorA2, c.anchorB2, c.baseSeparation2, c.normalMass2, c.normalImpulse2, c.totalNormalImpulse2,
                   totalNormalImpulse, inv_h, contactSpeed, biasRate, massScale, impulseScale);
  float tangentX = c.normal.y, tangentY = 0.0f - c.normal.x;
  SolveFrictionPoint(bA, bB, c, c.anchorA1, c.anchorB1, tangentX, tangentY, c.tangentMass1, c.normalImpulse1, c.tangentImpulse1);
  SolveFrictionPoint(bA, bB, c, c.anchorA2, c.anchorB2, tangentX, tangentY, c.tangentMass2, c.normalImpulse2, c.tangentImpulse2);
  {
    float deltaLambda = c.rollingMass * (bA.w - bB.w);
    float lambda = c.rollingImpulse;
    float maxLambda = c.rollingResistance * totalNormalImpulse;
    c.rollingImpulse = avxMax(0.0f - maxLambda, avxMin(lambda + deltaLambda, maxLambda));   // SymClampW :372
    deltaLambda = c.rollingImpulse - lambda;
    bA.w = bA.w - c.invIA * deltaLambda;
    bB.w = bB.w + c.invIB * deltaLambda;
  }
  Scatter(states, c.indexA, bA); Scatter(states, c.indexB, bB);
}

// :913-989 with quirk Q1: multiplies by the comparison mask (all-ones bits == NaN, or 0.0) instead of the blended mass.
static inline void RestitutionPoint(BodyW& bA, BodyW& bB, const ConstraintLane& c, Vec2 rA, Vec2 rB, float relativeVelocity,
                                    float& normalImpulse, float& totalNormalImpulseP, float threshold, bool restitutionMask) {
  bool mask1 = (relativeVelocity + threshold) > 0.0f;
  bool mask2 = totalNormalImpulseP == 0.0f;
  bool maskb = mask1 || mask2 || restitutionMask;
  float mask = maskb ? NAN : 0.0f;
  float dvx = (bB.v.x - bB.w * rB.y) - (bA.v.x - bA.w * rA.y);
  float dvy = (bB.v.y + bB.w * rB.x) - (bA.v.y + bA.w * rA.x);
  float vn = dvx * c.normal.x + dvy * c.normal.y;
  float negImpulse = mask * (vn + c.restitution * relativeVelocity);
  float newImpulse = avxMax(normalImpulse - negImpulse, 0.0f);
  float deltaImpulse = newImpulse - normalImpulse;
  normalImpulse = newImpulse;
  totalNormalImpulseP = totalNormalImpulseP + deltaImpulse;
  float Px = deltaImpulse * c.normal.x, Py = deltaImpulse * c.normal.y;
  ApplyP(bA, bB, c, rA, rB, Px, Py);
}
static void RestitutionGroup(std::vector<BodyState>& states, ConstraintLane* lanes, float threshold) {
  bool allZero = true;
  for (int j = 0; j < 8; j++) if (!(lanes[j].restitution == 0.0f)) allZero = false;    // AllZeroW :373
  if (allZero) return;
  // the SIMD code gathers all 8 lanes first, then scatters A (8 lanes) then B (8 lanes)
  BodyW bA[8], bB[8];
  for (int j = 0; j < 8; j++) { bA[j] = Gather(states, lanes[j].indexA); bB[j] = Gather(states, lanes[j].indexB); }
  for (int j = 0; j < 8; j++) {
    ConstraintLane& c = lanes[j];
    bool restitutionMask = c.restitution == 0.0f;
    RestitutionPoint(bA[j], bB[j], c, c.anchorA1, c.anchorB1, c.relativeVelocity1, c.normalImpulse1, c.totalNormalImpulse1, threshold, restitutionMask);
    RestitutionPoint(bA[j], bB[j], c, c.anchorA2, c.anchorB2, c.relativeVelocity2, c.normalImpulse2, c.totalNormalImpulse2, threshold, restitutionMask);
  }
  for (int j = 0; j < 8; j++) Scatter(states, lanes[j].indexA, bA[j]);
  for (int j = 0; j < 8; j++) Scatter(states, lanes[j].indexB, bB[j]);
}

#include "b2o_joints.inc"

// ---- JointSim.PrepareJoint src/Joint.cs:125-130, then IJoint.Prepare. The body/frame part is the same text in every joint type
// (RevoluteJoint.cs:72-95, WeldJoint.cs:46-66, MotorJoint.cs:60-81, PrismaticJoint.cs:86-106, WheelJoint.cs:91-111, DistanceJoint.cs:91-113) ----
static void PrepareJoint(World& world, StepContext& context, JointSim& joint) {
  float hertz = netMin(joint.constraintHertz, 0.25f * context.inv_h);
  joint.constraintSoftness = Softness(hertz, joint.constraintDampingRatio, context.h);
  if (joint.type == kFilterJoint) return;
  const Body& bodyA = world.bodies[joint.bodyIdA]; const Body& bodyB = world.bodies[joint.bodyIdB];
  const BodySim& bodySimA = world.GetBodySim(bodyA); const BodySim& bodySimB = world.GetBodySim(bodyB);
  float mA = bodySimA.invMass, iA = bodySimA.invInertia, mB = bodySimB.invMass, iB = bodySimB.invInertia;
  joint.invMassA = mA; joint.invMassB = mB; joint.invIA = iA; joint.invIB = iB;
  joint.indexA = bodyA.setIndex == kAwakeSet ? bodyA.localIndex : -1;
  joint.indexB = bodyB.setIndex == kAwakeSet ? bodyB.localIndex : -1;
  joint.frameA.q = MulRot(bodySimA.transform.q, joint.localFrameA.q);
  joint.frameA.p = RotateVector(bodySimA.transform.q, joint.localFrameA.p - bodySimA.localCenter);
  joint.frameB.q = MulRot(bodySimB.transform.q, joint.localFrameB.q);
  joint.frameB.p = RotateVector(bodySimB.transform.q, joint.localFrameB.p - bodySimB.localCenter);
  joint.deltaCenter = bodySimB.center - bodySimA.center;
  switch (joint.type) {
    case kDistanceJoint: PrepareDistance(world, context, joint, bodySimA, bodySimB); return;
    case kMotorJoint: PrepareMotor(world, context, joint); return;
    case kPrismaticJoint: PreparePrismatic(world, context, joint); return;
    case kWeldJoint: PrepareWeld(world, context, joint); return;
    case kWheelJoint: PrepareWheel(world, context, joint); return;
    default: break;
  }
  // revolute: src/joints/RevoluteJoint.cs:96-104
  float k = iA + iB;
  joint.axialMass = k > 0 ? 1 / k : 0;
  joint.springSoftness = Softness(joint.hertz, joint.dampingRatio, context.h);
  if (!context.enableWarmStarting) { joint.linearImpulse = {0, 0}; joint.springImpulse = 0; joint.motorImpulse = 0; joint.lowerImpulse = 0; joint.upperImpulse = 0; }
}
static void WarmStartRevolute(World& world, JointSim& j) {
  float mA = j.invMassA, mB = j.invMassB, iA = j.invIA, iB = j.invIB;
  BodyState dummy;
  BodyState* stateA = j.indexA != -1 ? &world.awakeSet.bodyStates[j.indexA] : &dummy;
  BodyState* stateB = j.indexB != -1 ? &world.awakeSet.bodyStates[j.indexB] : &dummy;
  Vec2 rA = RotateVector(stateA->deltaRotation, j.frameA.p), rB = RotateVector(stateB->deltaRotation, j.frameB.p);
  float axialImpulse = j.springImpulse + j.motorImpulse + j.lowerImpulse - j.upperImpulse;
  if (stateA->flags & kDynamicFlag) {
    stateA->linearVelocity = MulSub(stateA->linearVelocity, mA, j.linearImpulse);
    stateA->angularVelocity -= iA * (Cross(rA, j.linearImpulse) + axialImpulse);
  }
  if (stateB->flags & kDynamicFlag) {
    stateB->linearVelocity = MulAdd(stateB->linearVelocity, mB, j.linearImpulse);
    stateB->angularVelocity += iB * (Cross(rB, j.linearImpulse) + axialImpulse);
  }
}
static void SolveRevolute(World& world, StepContext& context, JointSim& j, bool useBias) {
  float mA = j.invMassA, mB = j.invMassB, iA = j.invIA, iB = j.invIB;
  BodyState dummy;
  BodyState* stateA = j.indexA != -1 ? &world.awakeSet.bodyStates[j.indexA] : &dummy;
  BodyState* stateB = j.indexB != -1 ? &world.awakeSet.bodyStates[j.indexB] : &dummy;
  Vec2 vA = stateA->linearVelocity, vB = stateB->linearVelocity;
  float wA = stateA->angularVelocity, wB = stateB->angularVelocity;
  Rot qA = MulRot(stateA->deltaRotation, j.frameA.q), qB = MulRot(stateB->deltaRotation, j.frameB.q);
  Rot relQ = InvMulRot(qA, qB);
  bool fixedRotation = iA + iB == 0;
  if (j.enableSpring && !fixedRotation) {
    float jointAngle = RotGetAngle(relQ);
    float jointAngleDelta = UnwindAngle(jointAngle - j.targetAngle);
    float bias = j.springSoftness.biasRate * jointAngleDelta;
    float massScale = j.springSoftness.massScale, impulseScale = j.springSoftness.impulseScale;
    float Cdot = wB - wA;
    float impulse = -massScale * j.axialMass * (Cdot + bias) - impulseScale * j.springImpulse;
    j.springImpulse += impulse;
    wA -= iA * impulse; wB += iB * impulse;
  }
  if (j.enableMotor && !fixedRotation) {
    float Cdot = wB - wA - j.motorSpeed;
    float impulse = -j.axialMass * Cdot;
    float oldImpulse = j.motorImpulse;
    float maxImpulse = context.h * j.maxMotorTorque;
    j.motorImpulse = netClamp(j.motorImpulse + impulse, -maxImpulse, maxImpulse);
    impulse = j.motorImpulse - oldImpulse;
    wA -= iA * impulse; wB += iB * impulse;
  }
  if (j.enableLimit && !fixedRotation) {
    float jointAngle = RotGetAngle(relQ);
    {
      float C = jointAngle - j.lowerAngle;
      float bias = 0, massScale = 1, impulseScale = 0;
      if (C > 0) bias = C * context.inv_h;
      else if (useBias) { bias = j.constraintSoftness.biasRate * C; massScale = j.constraintSoftness.massScale; impulseScale = j.constraintSoftness.impulseScale; }
      float Cdot = wB - wA;
      float oldImpulse = j.lowerImpulse;
      float impulse = -massScale * j.axialMass * (Cdot + bias) - impulseScale * oldImpulse;
      j.lowerImpulse = netMax(oldImpulse + impulse, 0);
      impulse = j.lowerImpulse - oldImpulse;
      wA -= iA * impulse; wB += iB * impulse;
    }
    {
      float C = j.upperAngle - jointAngle;
      float bias = 0, massScale = 1, impulseScale = 0;
      if (C > 0) bias = C * context.inv_h;
      else if (useBias) { bias = j.constraintSoftness.biasRate * C; massScale = j.constraintSoftness.massScale; impulseScale = j.constraintSoftness.impulseScale; }
      float Cdot = wA - wB;
      float oldImpulse = j.upperImpulse;
      float impulse = -massScale * j.axialMass * (Cdot + bias) - impulseScale * oldImpulse;
      j.upperImpulse = netMax(oldImpulse + impulse, 0);
      impulse = j.upperImpulse - oldImpulse;
      wA += iA * impulse; wB -= iB * impulse;
    }
  }
  {
    Vec2 rA = RotateVector(stateA->deltaRotation, j.frameA.p), rB = RotateVector(stateB->deltaRotation, j.frameB.p);
    Vec2 Cdot = (vB + CrossSV(wB, rB)) - (vA + CrossSV(wA, rA));
    Vec2 bias = {0, 0};
    float massScale = 1, impulseScale = 0;
    if (useBias) {
      Vec2 dcA = stateA->deltaPosition, dcB = stateB->deltaPosition;
      Vec2 separation = ((dcB - dcA) + (rB - rA)) + j.deltaCenter;
      bias = j.constraintSoftness.biasRate * separation;
      massScale = j.constraintSoftness.massScale; impulseScale = j.constraintSoftness.impulseScale;
    }
    Mat22 K;
    K.cx.x = mA + mB + rA.y * rA.y * iA + rB.y * rB.y * iB;
    K.cx.y = -rA.y * rA.x * iA - rB.y * rB.x * iB;
    K.cy.x = K.cx.y;
    K.cy.y = mA + mB + rA.x * rA.x * iA + rB.x * rB.x * iB;
    Vec2 b = Solve22(K, Cdot + bias);
    Vec2 impulse = (-massScale * b) - (impulseScale * j.linearImpulse);
    j.linearImpulse = j.linearImpulse + impulse;
    vA = MulSub(vA, mA, impulse); wA -= iA * Cross(rA, impulse);
    vB = MulAdd(vB, mB, impulse); wB += iB * Cross(rB, impulse);
  }
  if (stateA->flags & kDynamicFlag) { stateA->linearVelocity = vA; stateA->angularVelocity = wA; }
  if (stateB->flags & kDynamicFlag) { stateB->linearVelocity = vB; stateB->angularVelocity = wB; }
}

static void WarmStartJoint(World& world, JointSim& j) {        // JointSim.WarmStart src/Joint.cs:131
  switch (j.type) {
    case kDistanceJoint: WarmStartDistance(world, j); break;
    case kMotorJoint: WarmStartMotor(world, j); break;
    case kPrismaticJoint: WarmStartPrismatic(world, j); break;
    case kRevoluteJoint: WarmStartRevolute(world, j); break;
    case kWeldJoint: WarmStartWeld(world, j); break;
    case kWheelJoint: WarmStartWheel(world, j); break;
    default: break;                                              // filter: IJoint default, empty
  }
}
static void SolveJoint(World& world, StepContext& context, JointSim& j, bool useBias) {   // JointSim.Solve src/Joint.cs:132
  switch (j.type) {
    case kDistanceJoint: SolveDistance(world, context, j, useBias); break;
    case kMotorJoint: SolveMotor(world, context, j, useBias); break;
    case kPrismaticJoint: SolvePrismatic(world, context, j, useBias); break;
    case kRevoluteJoint: SolveRevolute(world, context, j, useBias); break;
    case kWeldJoint: SolveWeld(world, context, j, useBias); break;
    case kWheelJoint: SolveWheel(world, context, j, useBias); break;
    default: break;
  }
}

// ---- integrate: src/Solver.cs:121-160,199-212 ----
static void IntegrateVelocities(World& world, StepContext& context) {
  Vec2 gravity = world.gravity; float h = context.h;
  float maxLinearSpeed = context.maxLinearVelocity;
  float maxAngularSpeed = kMaxRotation * context.inv_dt;
  float maxLinearSpeedSquared = maxLinearSpeed * maxLinearSpeed, maxAngularSpeedSquared = maxAngularSpeed * maxAngularSpeed;
  int n = (int)world.awakeSet.bodySims.size();
  for (int i = 0; i < n; i++) {
    BodySim& sim = world.awakeSet.bodySims[i]; BodyState& state = world.awakeSet.bodyStates[i];
    Vec2 v = state.linearVelocity; float w = state.angularVelocity;
    float linearDamping = 1 / (1 + h * sim.linearDamping), angularDamping = 1 / (1 + h * sim.angularDamping);
    float gravityScale = sim.invMass > 0 ? sim.gravityScale : 0;
    Vec2 linearVelocityDelta = ((h * sim.invMass) * sim.force) + ((h * gravityScale) * gravity);
    float angularVelocityDelta = h * sim.invInertia * sim.torque;
    v = MulAdd(linearVelocityDelta, linearDamping, v);
    w = angularVelocityDelta + angularDamping * w;
    if (Dot(v, v) > maxLinearSpeedSquared) { float ratio = maxLinearSpeed / Length(v); v = ratio * v; sim.flags |= kIsSpeedCapped; }
    if (w * w > maxAngularSpeedSquared && !(sim.flags & kAllowFastRotation)) { float ratio = maxAngularSpeed / fabsf(w); w *= ratio; sim.flags |= kIsSpeedCapped; }
    if (state.flags & kLockLinearX) v.x = 0;
    if (state.flags & kLockLinearY) v.y = 0;
    if (state.flags & kLockAngularZ) w = 0;
    state.linearVelocity = v; state.angularVelocity = w;
  }
}
static void IntegratePositions(World& world, StepContext& context) {
  float h = context.h;
  for (BodyState& state : world.awakeSet.bodyStates) {
    if (state.flags & kLockLinearX) state.linearVelocity.x = 0;
    if (state.flags & kLockLinearY) state.linearVelocity.y = 0;
    if (state.flags & kLockAngularZ) state.angularVelocity = 0;
    state.deltaPosition = MulAdd(state.deltaPosition, h, state.linearVelocity);
    state.deltaRotation = IntegrateRot(state.deltaRotation, h * state.angularVelocity);
  }
}

}  // namespace b2o

#include "b2o_solver_stages.inc"
