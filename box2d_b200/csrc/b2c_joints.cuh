// libb2cuda — joint types other than revolute on the grid solver path: one thread per joint within a colour (no two joints of a colour
// share a dynamic body), body rows gathered/scattered as float4. Replaces the IJoint.Prepare/WarmStart/Solve bodies of
// src/joints/DistanceJoint.cs:91-258, WeldJoint.cs:46-153, MotorJoint.cs:60-214, PrismaticJoint.cs:86-286, WheelJoint.cs:91-284
// (FilterJoint.cs: the IJoint defaults src/Joint.cs:52-61 are empty). fp32, no FMA, the reference's association. Field aliases: JointDev.
// Port quirks kept as written: distance takes invMass for iA/iB (:106-107); weld builds K with cy.x = 0 (:136-137); prismatic limits run
// when C > speculativeDistance, bias = C*inv_h in the soft branch, and push P = (accumulated impulse) * axisA component-wise (:214-248);
// wheel's point-to-line Cdot uses axisA (:266).
#pragma once

struct JointBodies { BodyW a, b; V2 rA, rB; };
__device__ __forceinline__ JointBodies jointGather(const WorldDev& d, const JointDev& j) {
  JointBodies s; s.a = gatherBody(d, j.indexA); s.b = gatherBody(d, j.indexB);
  s.rA = rotate(s.a.dq, j.frameA.p); s.rB = rotate(s.b.dq, j.frameB.p);
  return s;
}
__device__ __forceinline__ void jointScatter(const WorldDev& d, const JointDev& j, JointBodies& s, V2 vA, float wA, V2 vB, float wB) {
  s.a.v = vA; s.a.w = wA; s.b.v = vB; s.b.w = wB;
  scatterBody(d, j.indexA, s.a); scatterBody(d, j.indexB, s.b);
}
__device__ __forceinline__ V2 mulM(float4 m, V2 v) { return mk(m.x * v.x + m.z * v.y, m.y * v.x + m.w * v.y); }   // Mat22 * v, include/MathFunctions.cs:202
__device__ __forceinline__ float4 motorLinearMass(float mA, float mB, float iA, float iB, V2 rA, V2 rB) {        // MotorJoint.cs:85-88,164-167 + Mat22.Inverse :204-208
  float cxx = mA + mB + rA.y * rA.y * iA + rB.y * rB.y * iB;
  float cxy = -rA.y * rA.x * iA - rB.y * rB.x * iB;
  float cyx = cxy;
  float cyy = mA + mB + rA.x * rA.x * iA + rB.x * rB.x * iB;
  float det = cxx * cyy - cyx * cxy; if (det != 0.0f) det = 1.0f / det;
  return make_float4(det * cyy, -det * cxy, -det * cyx, det * cxx);
}

// ---------------- distance ----------------
__device__ void prepareDistance(const StepParams& sp, JointDev& j, float invMassA, float invMassB) {
  float mA = invMassA, iA = invMassA, mB = invMassB, iB = invMassB;
  j.invMassA = mA; j.invMassB = mB; j.invIA = iA; j.invIB = iB;
  V2 rA = j.frameA.p, rB = j.frameB.p;
  V2 separation = (rB - rA) + j.deltaCenter;
  V2 axis = normalize(separation);
  float crA = cross(rA, axis), crB = cross(rB, axis);
  float k = mA + mB + iA * crA * crA + iB * crB * crB;
  j.axialMass = k > 0.0f ? 1.0f / k : 0.0f;
  j.springSoftness = makeSoft(j.hertz, j.dampingRatio, sp.h);
  if (!sp.enableWarmStarting) { j.springImpulse = 0; j.lowerImpulse = 0; j.upperImpulse = 0; j.motorImpulse = 0; }
}
__device__ void warmStartDistance(const WorldDev& d, JointDev& j) {
  JointBodies s = jointGather(d, j);
  V2 ds = (s.b.dp - s.a.dp) + (s.rB - s.rA);
  V2 axis = normalize(j.deltaCenter + ds);
  float axialImpulse = j.springImpulse + j.lowerImpulse - j.upperImpulse + j.motorImpulse;
  V2 P = axialImpulse * axis;
  if (s.a.flags & kDynamicFlag) { s.a.v = mulSub(s.a.v, j.invMassA, P); s.a.w -= j.invIA * cross(s.rA, P); }
  if (s.b.flags & kDynamicFlag) { s.b.v = mulAdd(s.b.v, j.invMassB, P); s.b.w += j.invIB * cross(s.rB, P); }
  scatterBody(d, j.indexA, s.a); scatterBody(d, j.indexB, s.b);
}
__device__ void solveDistance(const WorldDev& d, const StepParams& sp, JointDev& j, bool useBias) {
  float mA = j.invMassA, mB = j.invMassB, iA = j.invIA, iB = j.invIB;
  JointBodies s = jointGather(d, j);
  V2 vA = s.a.v, vB = s.b.v; float wA = s.a.w, wB = s.b.w;
  V2 rA = s.rA, rB = s.rB;
  V2 ds = (s.b.dp - s.a.dp) + (rB - rA);
  V2 separation = j.deltaCenter + ds;
  float len; V2 axis = lengthAndNormalize(separation, &len);
  if (j.enableSpring && (j.minLength < j.maxLength || !j.enableLimit)) {
    if (j.hertz > 0.0f) {
      V2 vr = (vB - vA) + (crossSV(wB, rB) - crossSV(wA, rA));
      float Cdot = dot(axis, vr);
      float C = len - j.length;
      float bias = j.springSoftness.biasRate * C;
      float m = j.springSoftness.massScale * j.axialMass;
      float oldImpulse = j.springImpulse;
      float impulse = -m * (Cdot + bias) - j.springSoftness.impulseScale * oldImpulse;
      j.springImpulse = netClamp(j.springImpulse + impulse, j.lowerSpringForce * sp.h, j.upperSpringForce * sp.h);
      impulse = j.springImpulse - oldImpulse;
      V2 P = impulse * axis;
      vA = mulSub(vA, mA, P); wA -= iA * cross(rA, P);
      vB = mulAdd(vB, mB, P); wB += iB * cross(rB, P);
    }
    if (j.enableLimit) {
      {
        V2 vr = (vB - vA) + (crossSV(wB, rB) - crossSV(wA, rA));
        float Cdot = dot(axis, vr);
        float C = len - j.minLength;
        float bias = 0.0f, massCoeff = 1.0f, impulseCoeff = 0.0f;
        if (C > 0.0f) bias = C * sp.inv_h;
        else if (useBias) { bias = j.constraintSoftness.biasRate * C; massCoeff = j.constraintSoftness.massScale; impulseCoeff = j.constraintSoftness.impulseScale; }
        float impulse = -massCoeff * j.axialMass * (Cdot + bias) - impulseCoeff * j.lowerImpulse;
        float newImpulse = netMax(0.0f, j.lowerImpulse + impulse);
        impulse = newImpulse - j.lowerImpulse;
        j.lowerImpulse = newImpulse;
        V2 P = impulse * axis;
        vA = mulSub(vA, mA, P); wA -= iA * cross(rA, P);
        vB = mulAdd(vB, mB, P); wB += iB * cross(rB, P);
      }
      {
        V2 vr = (vA - vB) + (crossSV(wA, rA) - crossSV(wB, rB));
        float Cdot = dot(axis, vr);
        float C = j.maxLength - len;
        float bias = 0.0f, massScale = 1.0f, impulseScale = 0.0f;
        if (C > 0.0f) bias = C * sp.inv_h;
        else if (useBias) { bias = j.constraintSoftness.biasRate * C; massScale = j.constraintSoftness.massScale; impulseScale = j.constraintSoftness.impulseScale; }
        float impulse = -massScale * j.axialMass * (Cdot + bias) - impulseScale * j.upperImpulse;
        float newImpulse = netMax(0.0f, j.upperImpulse + impulse);
        impulse = newImpulse - j.upperImpulse;
        j.upperImpulse = newImpulse;
        V2 P = -impulse * axis;
        vA = mulSub(vA, mA, P); wA -= iA * cross(rA, P);
        vB = mulAdd(vB, mB, P); wB += iB * cross(rB, P);
      }
    }
    if (j.enableMotor) {
      V2 vr = (vB - vA) + (crossSV(wB, rB) - crossSV(wA, rA));
      float Cdot = dot(axis, vr);
      float impulse = j.axialMass * (j.motorSpeed - Cdot);
      float oldImpulse = j.motorImpulse;
      float maxImpulse = sp.h * j.maxMotorTorque;
      j.motorImpulse = netClamp(j.motorImpulse + impulse, -maxImpulse, maxImpulse);
      impulse = j.motorImpulse - oldImpulse;
      V2 P = impulse * axis;
      vA = mulSub(vA, mA, P); wA -= iA * cross(rA, P);
      vB = mulAdd(vB, mB, P); wB += iB * cross(rB, P);
    }
  } else {
    V2 vr = (vB - vA) + (crossSV(wB, rB) - crossSV(wA, rA));
    float Cdot = dot(axis, vr);
    float C = len - j.length;
    float bias = 0.0f, massScale = 1.0f, impulseScale = 0.0f;
    if (useBias) { bias = j.constraintSoftness.biasRate * C; massScale = j.constraintSoftness.massScale; impulseScale = j.constraintSoftness.impulseScale; }
    float impulse = -massScale * j.axialMass * (Cdot + bias) - impulseScale * j.springImpulse;
    j.springImpulse += impulse;
    V2 P = impulse * axis;
    vA = mulSub(vA, mA, P); wA -= iA * cross(rA, P);
    vB = mulAdd(vB, mB, P); wB += iB * cross(rB, P);
  }
  jointScatter(d, j, s, vA, wA, vB, wB);
}

// ---------------- weld ----------------
__device__ void prepareWeld(const StepParams& sp, JointDev& j) {
  float ka = j.invIA + j.invIB;
  j.axialMass = ka > 0.0f ? 1.0f / ka : 0.0f;
  j.springSoftness = j.hertz == 0.0f ? j.constraintSoftness : makeSoft(j.hertz, j.dampingRatio, sp.h);
  j.angularSoftness = j.angularHertz == 0.0f ? j.constraintSoftness : makeSoft(j.angularHertz, j.angularDampingRatio, sp.h);
  if (!sp.enableWarmStarting) { j.linearImpulse = mk(0, 0); j.springImpulse = 0; }
}
__device__ void warmStartLinearAngular(const WorldDev& d, JointDev& j, V2 linearImpulse, float angularImpulse) {   // WeldJoint.cs:77-92, MotorJoint.cs:99-118
  JointBodies s = jointGather(d, j);
  if (s.a.flags & kDynamicFlag) { s.a.v = mulSub(s.a.v, j.invMassA, linearImpulse); s.a.w -= j.invIA * (cross(s.rA, linearImpulse) + angularImpulse); }
  if (s.b.flags & kDynamicFlag) { s.b.v = mulAdd(s.b.v, j.invMassB, linearImpulse); s.b.w += j.invIB * (cross(s.rB, linearImpulse) + angularImpulse); }
  scatterBody(d, j.indexA, s.a); scatterBody(d, j.indexB, s.b);
}
__device__ void solveWeld(const WorldDev& d, const StepParams& sp, JointDev& j, bool useBias) {
  float mA = j.invMassA, mB = j.invMassB, iA = j.invIA, iB = j.invIB;
  JointBodies s = jointGather(d, j);
  V2 vA = s.a.v, vB = s.b.v; float wA = s.a.w, wB = s.b.w;
  {
    Rot qA = mulRot(s.a.dq, j.frameA.q), qB = mulRot(s.b.dq, j.frameB.q);
    Rot relQ = invMulRot(qA, qB);
    float jointAngle = atan2f(relQ.s, relQ.c);
    float bias = 0.0f, massScale = 1.0f, impulseScale = 0.0f;
    if (useBias || j.angularHertz > 0.0f) { bias = j.angularSoftness.biasRate * jointAngle; massScale = j.angularSoftness.massScale; impulseScale = j.angularSoftness.impulseScale; }
    float Cdot = wB - wA;
    float impulse = -massScale * j.axialMass * (Cdot + bias) - impulseScale * j.springImpulse;
    j.springImpulse += impulse;
    wA -= iA * impulse; wB += iB * impulse;
  }
  {
    V2 rA = s.rA, rB = s.rB;
    V2 bias = mk(0, 0);
    float massScale = 1.0f, impulseScale = 0.0f;
    if (useBias || j.hertz > 0.0f) {
      bias = j.springSoftness.biasRate * (((s.b.dp - s.a.dp) + (rB - rA)) + j.deltaCenter);
      massScale = j.springSoftness.massScale; impulseScale = j.springSoftness.impulseScale;
    }
    V2 Cdot = (vB + crossSV(wB, rB)) - (vA + crossSV(wA, rA));
    float cxx = mA + mB + rA.y * rA.y * iA + rB.y * rB.y * iB;
    float cxy = -rA.y * rA.x * iA - rB.y * rB.x * iB;
    float cyx = 0.0f;                                         // as written in the port
    float cyy = mA + mB + rA.x * rA.x * iA + rB.x * rB.x * iB;
    V2 rhs = Cdot + bias;
    float det = cxx * cyy - cyx * cxy; if (det != 0.0f) det = 1.0f / det;       // Mat22.Solve include/MathFunctions.cs:211-215
    V2 b = mk(det * (cyy * rhs.x - cyx * rhs.y), det * (cxx * rhs.y - cxy * rhs.x));
    V2 impulse = mk(-massScale * b.x - impulseScale * j.linearImpulse.x, -massScale * b.y - impulseScale * j.linearImpulse.y);
    j.linearImpulse = j.linearImpulse + impulse;
    vA = mulSub(vA, mA, impulse); wA -= iA * cross(rA, impulse);
    vB = mulAdd(vB, mB, impulse); wB += iB * cross(rB, impulse);
  }
  jointScatter(d, j, s, vA, wA, vB, wB);
}

// ---------------- motor ----------------
__device__ void prepareMotor(const StepParams& sp, JointDev& j) {
  j.springSoftness = makeSoft(j.hertz, j.dampingRatio, sp.h);
  j.angularSoftness = makeSoft(j.angularHertz, j.angularDampingRatio, sp.h);
  j.linearMass = motorLinearMass(j.invMassA, j.invMassB, j.invIA, j.invIB, j.frameA.p, j.frameB.p);
  float ka = j.invIA + j.invIB;
  j.axialMass = ka > 0.0f ? 1.0f / ka : 0.0f;
  if (!sp.enableWarmStarting) { j.linearImpulse = mk(0, 0); j.motorImpulse = 0; j.linearImpulse2 = mk(0, 0); j.springImpulse = 0; }
}
__device__ void solveMotor(const WorldDev& d, const StepParams& sp, JointDev& j) {
  float mA = j.invMassA, mB = j.invMassB, iA = j.invIA, iB = j.invIB;
  JointBodies s = jointGather(d, j);
  V2 vA = s.a.v, vB = s.b.v; float wA = s.a.w, wB = s.b.w;
  if (j.maxSpringTorque > 0.0f && j.angularHertz > 0.0f) {
    Rot qA = mulRot(s.a.dq, j.frameA.q), qB = mulRot(s.b.dq, j.frameB.q);
    Rot relQ = invMulRot(qA, qB);
    float c = atan2f(relQ.s, relQ.c);
    float bias = j.angularSoftness.biasRate * c;
    float cdot = wB - wA;
    float maxImpulse = sp.h * j.maxSpringTorque;
    float oldImpulse = j.springImpulse;
    float impulse = -j.angularSoftness.massScale * j.axialMass * (cdot + bias) - j.angularSoftness.impulseScale * oldImpulse;
    j.springImpulse = netClamp(oldImpulse + impulse, -maxImpulse, maxImpulse);
    impulse = j.springImpulse - oldImpulse;
    wA -= iA * impulse; wB += iB * impulse;
  }
  if (j.maxVelocityTorque > 0.0f) {
    float cdot = wB - wA - j.angularVelocity;
    float impulse = -j.axialMass * cdot;
    float maxImpulse = sp.h * j.maxVelocityTorque;
    float oldImpulse = j.motorImpulse;
    j.motorImpulse = netClamp(oldImpulse + impulse, -maxImpulse, maxImpulse);
    impulse = j.motorImpulse - oldImpulse;
    wA -= iA * impulse; wB += iB * impulse;
  }
  V2 rA = s.rA, rB = s.rB;
  if (j.maxSpringForce > 0.0f && j.hertz > 0.0f) {
    V2 c = ((s.b.dp - s.a.dp) + (rB - rA)) + j.deltaCenter;
    V2 bias = j.springSoftness.biasRate * c;
    V2 cdot = (vB + crossSV(wB, rB)) - (vA + crossSV(wA, rA));
    cdot = cdot + bias;
    j.linearMass = motorLinearMass(mA, mB, iA, iB, rA, rB);
    V2 b = mulM(j.linearMass, cdot);
    V2 oldImpulse = j.linearImpulse2;
    V2 impulse = (-j.springSoftness.massScale * b) - (j.springSoftness.impulseScale * oldImpulse);
    float maxImpulse = sp.h * j.maxSpringForce;
    j.linearImpulse2 = j.linearImpulse2 + impulse;
    if (lengthSq(j.linearImpulse2) > maxImpulse * maxImpulse) {
      j.linearImpulse2 = normalize(j.linearImpulse2);
      j.linearImpulse2 = maxImpulse * j.linearImpulse2;
    }
    impulse = j.linearImpulse2 - oldImpulse;
    vA = mulSub(vA, mA, impulse); wA -= iA * cross(rA, impulse);
    vB = mulAdd(vB, mB, impulse); wB += iB * cross(rB, impulse);
  }
  if (j.maxVelocityForce > 0.0f) {
    V2 cdot = (vB + crossSV(wB, rB)) - (vA + crossSV(wA, rA));
    cdot = cdot - j.linearVelocity;
    V2 b = mulM(j.linearMass, cdot);
    V2 impulse = -b;
    V2 oldImpulse = j.linearImpulse;
    float maxImpulse = sp.h * j.maxVelocityForce;
    j.linearImpulse = j.linearImpulse + impulse;
    if (lengthSq(j.linearImpulse) > maxImpulse * maxImpulse) {
      j.linearImpulse = normalize(j.linearImpulse);
      j.linearImpulse = maxImpulse * j.linearImpulse;
    }
    impulse = j.linearImpulse - oldImpulse;
    vA = mulSub(vA, mA, impulse); wA -= iA * cross(rA, impulse);
    vB = mulAdd(vB, mB, impulse); wB += iB * cross(rB, impulse);
  }
  jointScatter(d, j, s, vA, wA, vB, wB);
}

// ---------------- prismatic / wheel shared geometry ----------------
struct AxisGeom { V2 d, axisA; float a1, a2; };
__device__ __forceinline__ AxisGeom axisGeom(const JointDev& j, const JointBodies& s) {   // PrismaticJoint.cs:124-129,163-169; WheelJoint.cs:131-137,166-171
  AxisGeom g;
  g.d = ((s.b.dp - s.a.dp) + j.deltaCenter) + (s.rB - s.rA);
  V2 axisA = rotate(j.frameA.q, mk(1.0f, 0.0f));
  g.axisA = rotate(s.a.dq, axisA);
  g.a1 = cross(s.rA + g.d, g.axisA);
  g.a2 = cross(s.rB, g.axisA);
  return g;
}
__device__ __forceinline__ void axisWarmStart(const WorldDev& d, JointDev& j, JointBodies& s, V2 P, float LA, float LB) {
  if (s.a.flags & kDynamicFlag) { s.a.v = mulSub(s.a.v, j.invMassA, P); s.a.w -= j.invIA * LA; }
  if (s.b.flags & kDynamicFlag) { s.b.v = mulAdd(s.b.v, j.invMassB, P); s.b.w += j.invIB * LB; }
  scatterBody(d, j.indexA, s.a); scatterBody(d, j.indexB, s.b);
}

// ---------------- prismatic ----------------
__device__ void preparePrismatic(const StepParams& sp, JointDev& j) {
  j.springSoftness = makeSoft(j.hertz, j.dampingRatio, sp.h);
  if (!sp.enableWarmStarting) { j.linearImpulse = mk(0, 0); j.springImpulse = 0; j.motorImpulse = 0; j.lowerImpulse = 0; j.upperImpulse = 0; }
}
__device__ void warmStartPrismatic(const WorldDev& d, JointDev& j) {
  JointBodies s = jointGather(d, j);
  AxisGeom g = axisGeom(j, s);
  float axialImpulse = j.springImpulse + j.motorImpulse + j.lowerImpulse - j.upperImpulse;
  V2 perpA = leftPerp(g.axisA);
  float s1 = cross(s.rA + g.d, perpA);
  float s2 = cross(s.rB, perpA);
  float perpImpulse = j.linearImpulse.x, angleImpulse = j.linearImpulse.y;
  V2 P = (axialImpulse * g.axisA) + (perpImpulse * perpA);
  float LA = axialImpulse * g.a1 + perpImpulse * s1 + angleImpulse;
  float LB = axialImpulse * g.a2 + perpImpulse * s2 + angleImpulse;
  axisWarmStart(d, j, s, P, LA, LB);
}
__device__ void solvePrismatic(const WorldDev& d, const StepParams& sp, JointDev& j, bool useBias) {
  float mA = j.invMassA, mB = j.invMassB, iA = j.invIA, iB = j.invIB;
  JointBodies s = jointGather(d, j);
  V2 vA = s.a.v, vB = s.b.v; float wA = s.a.w, wB = s.b.w;
  Rot qA = mulRot(s.a.dq, j.frameA.q), qB = mulRot(s.b.dq, j.frameB.q);
  Rot relQ = invMulRot(qA, qB);
  V2 rA = s.rA, rB = s.rB;
  AxisGeom g = axisGeom(j, s);
  V2 dd = g.d, axisA = g.axisA; float a1 = g.a1, a2 = g.a2;
  float translation = dot(axisA, dd);
  float k = mA + mB + iA * a1 * a1 + iB * a2 * a2;
  float axialMass = k > 0.0f ? 1.0f / k : 0.0f;
  if (j.enableSpring) {
    float C = translation - j.targetAngle;
    float bias = j.springSoftness.biasRate * C;
    float massScale = j.springSoftness.massScale, impulseScale = j.springSoftness.impulseScale;
    float Cdot = dot(axisA, vB - vA) + a2 * wB - a1 * wA;
    float deltaImpulse = -massScale * axialMass * (Cdot + bias) - impulseScale * j.springImpulse;
    j.springImpulse += deltaImpulse;
    V2 P = deltaImpulse * axisA;
    float LA = deltaImpulse * a1, LB = deltaImpulse * a2;
    vA = mulSub(vA, mA, P); wA -= iA * LA;
    vB = mulAdd(vB, mB, P); wB += iB * LB;
  }
  if (j.enableMotor) {
    float Cdot = dot(axisA, vB - vA) + a2 * wB - a1 * wA;
    float impulse = axialMass * (j.motorSpeed - Cdot);
    float oldImpulse = j.motorImpulse;
    float maxImpulse = sp.h * j.maxMotorTorque;
    j.motorImpulse = netClamp(j.motorImpulse + impulse, -maxImpulse, maxImpulse);
    impulse = j.motorImpulse - oldImpulse;
    V2 P = impulse * axisA;
    float LA = impulse * a1, LB = impulse * a2;
    vA = mulSub(vA, mA, P); wA -= iA * LA;
    vB = mulAdd(vB, mB, P); wB += iB * LB;
  }
  if (j.enableLimit) {
    float speculativeDistance = 0.25f * (j.upperAngle - j.lowerAngle);
    {
      float C = translation - j.lowerAngle;
      if (C > speculativeDistance) {
        float bias = 0.0f, massScale = 1.0f, impulseScale = 0.0f;
        if (C > 0.0f) bias = netMin(C, 1.0f) * sp.inv_h;       // Box2D.LengthUnitsPerMeter = 1
        else if (useBias) { bias = C * sp.inv_h; massScale = j.constraintSoftness.massScale; impulseScale = j.constraintSoftness.impulseScale; }
        float oldImpulse = j.lowerImpulse;
        float Cdot = dot(axisA, vB - vA) + a2 * wB - a1 * wA;
        float deltaImpulse = -axialMass * massScale * (Cdot + bias) - impulseScale * oldImpulse;
        j.lowerImpulse = netMax(oldImpulse + deltaImpulse, 0.0f);
        deltaImpulse = j.lowerImpulse - oldImpulse;
        V2 P = mk(j.linearImpulse.x * axisA.x, j.linearImpulse.y * axisA.y);    // as written: Vector2 * Vector2 with the accumulated impulse
        float LA = deltaImpulse * a1, LB = deltaImpulse * a2;
        vA = mulSub(vA, mA, P); wA -= iA * LA;
        vB = mulAdd(vB, mB, P); wB += iB * LB;
      } else j.lowerImpulse = 0.0f;
    }
    {
      float C = j.upperAngle - translation;
      if (C > speculativeDistance) {
        float bias = 0.0f, massScale = 1.0f, impulseScale = 0.0f;
        if (C > 0.0f) bias = netMin(C, 1.0f) * sp.inv_h;
        else if (useBias) { bias = C * sp.inv_h; massScale = j.constraintSoftness.massScale; impulseScale = j.constraintSoftness.impulseScale; }
        float oldImpulse = j.upperImpulse;
        float Cdot = dot(axisA, vA - vB) + a1 * wA - a2 * wB;
        float deltaImpulse = -axialMass * massScale * (Cdot + bias) - impulseScale * oldImpulse;
        j.upperImpulse = netMax(oldImpulse + deltaImpulse, 0.0f);
        deltaImpulse = j.upperImpulse - oldImpulse;
        V2 P = mk(j.linearImpulse.x * axisA.x, j.linearImpulse.y * axisA.y);
        float LA = deltaImpulse * a1, LB = deltaImpulse * a2;
        vA = mulAdd(vA, mA, P); wA += iA * LA;
        vB = mulSub(vB, mB, P); wB -= iB * LB;
      } else j.upperImpulse = 0.0f;
    }
  }
  {
    V2 perpA = leftPerp(axisA);
    float s1 = cross(dd + rA, perpA);
    float s2 = cross(rB, perpA);
    V2 Cdot = mk(dot(perpA, vB - vA) + s2 * wB - s1 * wA, wB - wA);
    V2 bias = mk(0, 0);
    float massScale = 1.0f, impulseScale = 0.0f;
    if (useBias) {
      V2 C = mk(dot(perpA, dd), atan2f(relQ.s, relQ.c));
      bias = j.constraintSoftness.biasRate * C;
      massScale = j.constraintSoftness.massScale; impulseScale = j.constraintSoftness.impulseScale;
    }
    float k11 = mA + mB + iA * s1 * s1 + iB * s2 * s2;
    float k12 = iA * s1 + iB * s2;
    float k22 = iA + iB;
    if (k22 == 0.0f) k22 = 1.0f;
    V2 rhs = Cdot + bias;
    float det = k11 * k22 - k12 * k12; if (det != 0.0f) det = 1.0f / det;      // Mat22.Solve, cx=(k11,k12), cy=(k12,k22)
    V2 b = mk(det * (k22 * rhs.x - k12 * rhs.y), det * (k11 * rhs.y - k12 * rhs.x));
    V2 deltaImpulse = (-massScale * b) - (impulseScale * j.linearImpulse);
    j.linearImpulse = j.linearImpulse + deltaImpulse;
    V2 P = deltaImpulse.x * perpA;
    float LA = deltaImpulse.x * s1 + deltaImpulse.y, LB = deltaImpulse.x * s2 + deltaImpulse.y;
    vA = mulSub(vA, mA, P); wA -= iA * LA;
    vB = mulAdd(vB, mB, P); wB += iB * LB;
  }
  jointScatter(d, j, s, vA, wA, vB, wB);
}

// ---------------- wheel ----------------
__device__ void prepareWheel(const StepParams& sp, JointDev& j) {
  float mA = j.invMassA, mB = j.invMassB, iA = j.invIA, iB = j.invIB;
  V2 rA = j.frameA.p, rB = j.frameB.p;
  V2 dd = j.deltaCenter + (rB - rA);
  V2 axisA = rotate(j.frameA.q, mk(1.0f, 0.0f));
  V2 perpA = leftPerp(axisA);
  float s1 = cross(dd + rA, perpA);
  float s2 = cross(rB, perpA);
  float kp = mA + mB + iA * s1 * s1 + iB * s2 * s2;
  j.perpMass = kp > 0.0f ? 1.0f / kp : 0.0f;
  float a1 = cross(dd + rA, axisA);
  float a2 = cross(rB, axisA);
  float ka = mA + mB + iA * a1 * a1 + iB * a2 * a2;
  j.axialMass = ka > 0.0f ? 1.0f / ka : 0.0f;
  j.springSoftness = makeSoft(j.hertz, j.dampingRatio, sp.h);
  float km = iA + iB;
  j.motorMass = km > 0.0f ? 1.0f / km : 0.0f;
  if (!sp.enableWarmStarting) { j.linearImpulse.x = 0; j.springImpulse = 0; j.motorImpulse = 0; j.lowerImpulse = 0; j.upperImpulse = 0; }
}
__device__ void warmStartWheel(const WorldDev& d, JointDev& j) {
  JointBodies s = jointGather(d, j);
  AxisGeom g = axisGeom(j, s);
  V2 perpA = leftPerp(g.axisA);
  float a1 = cross(g.d + s.rA, g.axisA);
  float s1 = cross(g.d + s.rA, perpA);
  float s2 = cross(s.rB, perpA);
  float perpImpulse = j.linearImpulse.x;
  float axialImpulse = j.springImpulse + j.lowerImpulse - j.upperImpulse;
  V2 P = (axialImpulse * g.axisA) + (perpImpulse * perpA);
  float LA = axialImpulse * a1 + perpImpulse * s1 + j.motorImpulse;
  float LB = axialImpulse * g.a2 + perpImpulse * s2 + j.motorImpulse;
  axisWarmStart(d, j, s, P, LA, LB);
}
__device__ void solveWheel(const WorldDev& d, const StepParams& sp, JointDev& j, bool useBias) {
  float mA = j.invMassA, mB = j.invMassB, iA = j.invIA, iB = j.invIB;
  JointBodies s = jointGather(d, j);
  V2 vA = s.a.v, vB = s.b.v; float wA = s.a.w, wB = s.b.w;
  bool fixedRotation = iA + iB == 0.0f;
  V2 rA = s.rA, rB = s.rB;
  AxisGeom g = axisGeom(j, s);
  V2 dd = g.d, axisA = g.axisA;
  float translation = dot(axisA, dd);
  float a1 = cross(dd + rA, axisA);
  float a2 = g.a2;
  if (j.enableMotor && !fixedRotation) {
    float Cdot = wB - wA - j.motorSpeed;
    float impulse = -j.motorMass * Cdot;
    float oldImpulse = j.motorImpulse;
    float maxImpulse = sp.h * j.maxMotorTorque;
    j.motorImpulse = netClamp(j.motorImpulse + impulse, -maxImpulse, maxImpulse);
    impulse = j.motorImpulse - oldImpulse;
    wA -= iA * impulse; wB += iB * impulse;
  }
  if (j.enableSpring) {
    float bias = j.springSoftness.biasRate * translation;
    float massScale = j.springSoftness.massScale, impulseScale = j.springSoftness.impulseScale;
    float Cdot = dot(axisA, vB - vA) + a2 * wB - a1 * wA;
    float impulse = -massScale * j.axialMass * (Cdot + bias) - impulseScale * j.springImpulse;
    j.springImpulse += impulse;
    V2 P = impulse * axisA;
    float LA = impulse * a1, LB = impulse * a2;
    vA = mulSub(vA, mA, P); wA -= iA * LA;
    vB = mulAdd(vB, mB, P); wB += iB * LB;
  }
  if (j.enableLimit) {
    {
      float C = translation - j.lowerAngle;
      float bias = 0.0f, massScale = 1.0f, impulseScale = 0.0f;
      if (C > 0.0f) bias = C * sp.inv_h;
      else if (useBias) { bias = j.constraintSoftness.biasRate * C; massScale = j.constraintSoftness.massScale; impulseScale = j.constraintSoftness.impulseScale; }
      float Cdot = dot(axisA, vB - vA) + a2 * wB - a1 * wA;
      float impulse = -massScale * j.axialMass * (Cdot + bias) - impulseScale * j.lowerImpulse;
      float oldImpulse = j.lowerImpulse;
      j.lowerImpulse = netMax(oldImpulse + impulse, 0.0f);
      impulse = j.lowerImpulse - oldImpulse;
      V2 P = impulse * axisA;
      float LA = impulse * a1, LB = impulse * a2;
      vA = mulSub(vA, mA, P); wA -= iA * LA;
      vB = mulAdd(vB, mB, P); wB += iB * LB;
    }
    {
      float C = j.upperAngle - translation;
      float bias = 0.0f, massScale = 1.0f, impulseScale = 0.0f;
      if (C > 0.0f) bias = C * sp.inv_h;
      else if (useBias) { bias = j.constraintSoftness.biasRate * C; massScale = j.constraintSoftness.massScale; impulseScale = j.constraintSoftness.impulseScale; }
      float Cdot = dot(axisA, vA - vB) + a1 * wA - a2 * wB;
      float impulse = -massScale * j.axialMass * (Cdot + bias) - impulseScale * j.upperImpulse;
      float oldImpulse = j.upperImpulse;
      j.upperImpulse = netMax(oldImpulse + impulse, 0.0f);
      impulse = j.upperImpulse - oldImpulse;
      V2 P = impulse * axisA;
      float LA = impulse * a1, LB = impulse * a2;
      vA = mulAdd(vA, mA, P); wA += iA * LA;
      vB = mulSub(vB, mB, P); wB -= iB * LB;
    }
  }
  {
    V2 perpA = leftPerp(axisA);
    float bias = 0.0f, massScale = 1.0f, impulseScale = 0.0f;
    if (useBias) { bias = j.constraintSoftness.biasRate * dot(perpA, dd); massScale = j.constraintSoftness.massScale; impulseScale = j.constraintSoftness.impulseScale; }
    float s1 = cross(dd + rA, perpA);
    float s2 = cross(rB, perpA);
    float Cdot = dot(axisA, vB - vA) + s2 * wB - s1 * wA;     // as written in the port (axisA)
    float impulse = -massScale * j.perpMass * (Cdot + bias) - impulseScale * j.linearImpulse.x;
    j.linearImpulse.x += impulse;
    V2 P = impulse * perpA;
    float LA = impulse * s1, LB = impulse * s2;
    vA = mulSub(vA, mA, P); wA -= iA * LA;
    vB = mulAdd(vB, mB, P); wB += iB * LB;
  }
  jointScatter(d, j, s, vA, wA, vB, wB);
}

// IJoint.GetReaction per type (DistanceJoint.cs:64-68, MotorJoint.cs:53-57, PrismaticJoint.cs:54-60, RevoluteJoint.cs:57-61, WeldJoint.cs:40-41,
// WheelJoint.cs:57-63; filter: zeros) scaled by inv_h as JointSim.GetJointReaction src/Joint.cs:119-124
__device__ void jointReaction(const JointDev& j, float inv_h, float* force, float* torque) {
  float lin = 0.0f, ang = 0.0f;
  switch (j.type) {
    case kDistanceJoint: lin = fabsf(j.springImpulse + j.lowerImpulse - j.upperImpulse + j.motorImpulse); break;
    case kMotorJoint: lin = length(j.linearImpulse + j.linearImpulse2); ang = fabsf(j.motorImpulse + j.springImpulse); break;
    case kPrismaticJoint: { float p = j.linearImpulse.x, a = j.motorImpulse + j.lowerImpulse - j.upperImpulse; lin = sqrtf(p * p + a * a); ang = fabsf(j.linearImpulse.y); } break;
    case kRevoluteJoint: lin = length(j.linearImpulse); ang = fabsf(j.motorImpulse + j.lowerImpulse - j.upperImpulse); break;
    case kWeldJoint: lin = length(j.linearImpulse); ang = fabsf(j.springImpulse); break;
    case kWheelJoint: { float p = j.linearImpulse.x, a = j.springImpulse + j.lowerImpulse - j.upperImpulse; lin = sqrtf(p * p + a * a); ang = fabsf(j.motorImpulse); } break;
    default: break;
  }
  *force = lin * inv_h; *torque = ang * inv_h;
}
