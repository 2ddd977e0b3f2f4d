// libb2cuda — the overflow colour (src/ConstraintGraph.cs:140,:181-186; src/ContactSolver.cs:40-350).
//
// A touching contact that finds no free colour on its bodies goes to colors[GraphColorCount-1]. The reference solves these
// with scalar code, one after another in the order of that colour's contactSims array (append on add, swap-with-last on
// remove), at the start of every stage before the coloured constraints (src/Solver.cs:620-698). They are few and they
// share bodies, so here one thread walks them in that order inside the grid-path solver; islands that own one are sent
// down the grid path. The port's quirks are reproduced: Q2 SolveOverflowContacts/ApplyOverflowRestitution
// read AND write body B through index A, Q3 restitution clamps -normalImpulse + impulse, Q4 the relax pass keeps useBias.
// Included by b2c_step.cu (after the lane functions).
#pragma once

namespace b2c {

// GraphColor.contactSims of EVERY colour in the reference's order. `ops` are this step's graph ops in ascending contact id (the order
// World.Collide applies them, src/World.cs:382-450); opCol[k] = colour | 0x40 (AddContactToGraph appended it) or | 0x80
// (RemoveContactFromGraph swap-removed it). Arrays of different colours are independent, so one warp replays one colour: the lanes
// scan the op list 32 at a time, lane 0 applies the ops of its colour in order. The array order decides which constraints share an
// 8-lane SIMD group in ApplyRestitutionTask (src/ContactSolver.cs:913-923), the order of the hit events (src/Solver.cs:1154-1160) and,
// for the overflow colour, the order of the serial solve.
__device__ void colReplayOps(const WorldDev& d, int n) {
  Scalars& s = *d.sc;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, warps = blockDim.x >> 5;
  const unsigned full = 0xffffffffu;
  for (int c = warp; c < kGraphColorCount; c += warps) {
    int count = s.colLen[c];
    int* arr = d.colList + (size_t)c * d.contactCapacity;
    const int cap = c == kOverflowIndex ? (d.ovCapacity < d.contactCapacity ? d.ovCapacity : d.contactCapacity) : d.contactCapacity;
    for (int base = 0; base < n; base += 32) {
      const int k = base + lane;
      const int oc = k < n ? (int)d.opCol[k] : 0;
      const bool mine = (oc & 0xc0) != 0 && (oc & 0x3f) == c;
      unsigned m = __ballot_sync(full, mine);
      if (m == 0) continue;
      const int id = mine ? d.opList[k] : -1;
      while (m) {
        const int b = __ffs((int)m) - 1; m &= m - 1;
        const int oid = __shfl_sync(full, id, b), ooc = __shfl_sync(full, oc, b);
        if (lane == 0) {
          if (ooc & 0x40) {                          // color.contactSims.Add  (:184-185)
            if (count >= cap) atomicOr(&s.errorFlags, (int)ERR_CONTACT_CAPACITY);
            else { arr[count] = oid; d.cLocal[oid] = count; count++; }
          } else {                                   // RemoveSwap(localIndex)  (:230-240)
            const int li = d.cLocal[oid], last = count - 1;
            if (li >= 0 && li <= last) {
              if (li != last) { int moved = arr[last]; arr[li] = moved; d.cLocal[moved] = li; }
              d.cLocal[oid] = -1; count--;
            }
          }
        }
      }
    }
    if (lane == 0) s.colLen[c] = count;
    __syncwarp();
  }
}

// PrepareOverflowContacts :40-115, one thread per overflow constraint
__device__ void ovPrepare(const WorldDev& d, const StepParams& sp, int k) {
  const int ci = d.colList[(size_t)kOverflowIndex * d.contactCapacity + k];
  OvConstraint& c = d.ovCons[k];
  const float warmStartScale = sp.enableWarmStarting ? 1.0f : 0.0f;
  int2 idx = d.cSim[ci];
  float4 mass = d.cMass[ci], mn = d.mN[ci], mat = d.cMat[ci];
  const int pointCount = (int)fbits(mn.w);
  c.indexA = idx.x; c.indexB = idx.y; c.normal = mk(mn.x, mn.y);
  c.friction = mat.x; c.restitution = mat.y; c.rollingResistance = mat.z; c.tangentSpeed = mat.w;
  c.rollingImpulse = warmStartScale * mn.z; c.pointCount = pointCount;
  V2 vA = mk(0, 0), vB = mk(0, 0); float wA = 0, wB = 0;
  const float mA = mass.x, iA = mass.y, mB = mass.z, iB = mass.w;
  if (idx.x != -1) { float4 st = d.st0[idx.x]; vA = mk(st.x, st.y); wA = st.z; }
  if (idx.y != -1) { float4 st = d.st0[idx.y]; vB = mk(st.x, st.y); wB = st.z; }
  c.softness = (idx.x == -1 || idx.y == -1) ? sp.staticSoftness : sp.contactSoftness;
  c.invMassA = mA; c.invIA = iA; c.invMassB = mB; c.invIB = iB;
  float kk = iA + iB; c.rollingMass = kk > 0.0f ? 1.0f / kk : 0.0f;
  V2 normal = c.normal, tangent = rightPerp(normal);
  for (int j = 0; j < pointCount; j++) {
    float4 a = j == 1 ? d.mA1[ci] : d.mA0[ci], b = j == 1 ? d.mB1[ci] : d.mB0[ci];
    OvPoint& cp = j == 1 ? c.p1 : c.p0;
    cp.normalImpulse = warmStartScale * b.y; cp.tangentImpulse = warmStartScale * b.z; cp.totalNormalImpulse = 0.0f;
    V2 rA = mk(a.x, a.y), rB = mk(a.z, a.w);
    cp.anchorA = rA; cp.anchorB = rB;
    cp.baseSeparation = b.x - dot(rB - rA, normal);
    float rnA = cross(rA, normal), rnB = cross(rB, normal);
    float kNormal = mA + mB + iA * rnA * rnA + iB * rnB * rnB;
    cp.normalMass = kNormal > 0.0f ? 1.0f / kNormal : 0.0f;
    float rtA = cross(rA, tangent), rtB = cross(rB, tangent);
    float kTangent = mA + mB + iA * rtA * rtA + iB * rtB * rtB;
    cp.tangentMass = kTangent > 0.0f ? 1.0f / kTangent : 0.0f;
    V2 vrA = vA + crossSV(wA, rA), vrB = vB + crossSV(wB, rB);
    cp.relativeVelocity = dot(normal, vrB - vrA);
  }
}

struct OvBody { V2 v; float w; V2 dp; Rot dq; bool dynamic; };
__device__ __forceinline__ OvBody ovLoad(const WorldDev& d, int index) {      // &dummyState when index == -1
  OvBody b;
  if (index < 0) { b.v = mk(0, 0); b.w = 0; b.dp = mk(0, 0); b.dq = mkRot(1, 0); b.dynamic = false; return b; }
  float4 a = d.st0[index], c = d.st1[index];
  b.v = mk(a.x, a.y); b.w = a.z; b.dp = mk(c.x, c.y); b.dq = mkRot(c.z, c.w); b.dynamic = (fbits(a.w) & kDynamicFlag) != 0;
  return b;
}
__device__ __forceinline__ void ovStore(const WorldDev& d, int index, bool dynamic, V2 v, float w) {
  if (index >= 0 && dynamic) { float4 a = d.st0[index]; a.x = v.x; a.y = v.y; a.z = w; d.st0[index] = a; }
}

// WarmStartOverflowContacts :116-164 — sequential over the array
__device__ void ovWarmStart(const WorldDev& d, int n) {
  for (int k = 0; k < n; k++) {
    OvConstraint& c = d.ovCons[k];
    OvBody A = ovLoad(d, c.indexA), B = ovLoad(d, c.indexB);
    V2 vA = A.v, vB = B.v; float wA = A.w, wB = B.w;
    const float mA = c.invMassA, iA = c.invIA, mB = c.invMassB, iB = c.invIB;
    V2 normal = c.normal, tangent = rightPerp(normal);
    for (int j = 0; j < c.pointCount; j++) {
      OvPoint& cp = j == 1 ? c.p1 : c.p0;
      V2 rA = cp.anchorA, rB = cp.anchorB;
      V2 P = (cp.normalImpulse * normal) + (cp.tangentImpulse * tangent);
      wA -= iA * cross(rA, P); vA = mulAdd(vA, -mA, P);
      wB += iB * cross(rB, P); vB = mulAdd(vB, mB, P);
    }
    wA -= iA * c.rollingImpulse; wB += iB * c.rollingImpulse;
    ovStore(d, c.indexA, A.dynamic, vA, wA);
    ovStore(d, c.indexB, B.dynamic, vB, wB);
  }
}

// SolveOverflowContacts :165-264 — Q2: "state B" is addressed through indexA
__device__ void ovSolve(const WorldDev& d, const StepParams& sp, int n, bool useBias) {
  const float contactSpeed = sp.contactSpeed, inv_h = sp.inv_h;
  for (int k = 0; k < n; k++) {
    OvConstraint& c = d.ovCons[k];
    const float mA = c.invMassA, iA = c.invIA, mB = c.invMassB, iB = c.invIB;
    const int ixB = (c.indexB != -1 && c.indexA != -1) ? c.indexA : -1;
    OvBody A = ovLoad(d, c.indexA), B = ovLoad(d, ixB);
    V2 vA = A.v, vB = B.v; float wA = A.w, wB = B.w;
    V2 dp = B.dp - A.dp;
    V2 normal = c.normal, tangent = rightPerp(normal);
    float totalNormalImpulse = 0.0f;
    for (int j = 0; j < c.pointCount; j++) {
      OvPoint& cp = j == 1 ? c.p1 : c.p0;
      V2 rA = cp.anchorA, rB = cp.anchorB;
      V2 ds = (dp + rotate(B.dq, rB)) - rotate(A.dq, rA);
      float s = cp.baseSeparation + dot(ds, normal);
      float velocityBias = 0.0f, massScale = 1.0f, impulseScale = 0.0f;
      if (s > 0.0f) velocityBias = s * inv_h;
      else if (useBias) { velocityBias = netMax(c.softness.massScale * c.softness.biasRate * s, -contactSpeed); massScale = c.softness.massScale; impulseScale = c.softness.impulseScale; }
      V2 vrA = vA + crossSV(wA, rA), vrB = vB + crossSV(wB, rB);
      float vn = dot(vrB - vrA, normal);
      float impulse = -cp.normalMass * (massScale * vn + velocityBias) - impulseScale * cp.normalImpulse;
      float newImpulse = netMax(cp.normalImpulse + impulse, 0.0f);
      impulse = newImpulse - cp.normalImpulse;
      cp.normalImpulse = newImpulse; cp.totalNormalImpulse += newImpulse; totalNormalImpulse += newImpulse;
      V2 P = impulse * normal;
      vA = mulSub(vA, mA, P); wA -= iA * cross(rA, P); vB = mulAdd(vB, mB, P); wB += iB * cross(rB, P);
    }
    for (int j = 0; j < c.pointCount; j++) {
      OvPoint& cp = j == 1 ? c.p1 : c.p0;
      V2 rA = cp.anchorA, rB = cp.anchorB;
      V2 vrB = vB + crossSV(wB, rB), vrA = vA + crossSV(wA, rA);
      float vt = dot(vrB - vrA, tangent) - c.tangentSpeed;
      float impulse = cp.tangentMass * -vt;
      float maxFriction = c.friction * cp.normalImpulse;
      float newImpulse = netClamp(cp.tangentImpulse + impulse, -maxFriction, maxFriction);
      impulse = newImpulse - cp.tangentImpulse;
      cp.tangentImpulse = newImpulse;
      V2 P = impulse * tangent;
      vA = mulSub(vA, mA, P); wA -= iA * cross(rA, P); vB = mulAdd(vB, mB, P); wB += iB * cross(rB, P);
    }
    {
      float deltaLambda = -c.rollingMass * (wB - wA);
      float lambda = c.rollingImpulse;
      float maxLambda = c.rollingResistance * totalNormalImpulse;
      c.rollingImpulse = netClamp(lambda + deltaLambda, -maxLambda, maxLambda);
      deltaLambda = c.rollingImpulse - lambda;
      wA -= iA * deltaLambda; wB += iB * deltaLambda;
    }
    ovStore(d, c.indexA, A.dynamic, vA, wA);
    ovStore(d, ixB, B.dynamic, vB, wB);           // Q2: lands on body A again
  }
}

// ApplyOverflowRestitution :265-322 (Q2, Q3)
__device__ void ovRestitution(const WorldDev& d, const StepParams& sp, int n) {
  const float threshold = sp.restitutionThreshold;
  for (int k = 0; k < n; k++) {
    OvConstraint& c = d.ovCons[k];
    const float restitution = c.restitution;
    if (restitution == 0.0f) continue;
    const float mA = c.invMassA, iA = c.invIA, mB = c.invMassB, iB = c.invIB;
    const int ixB = (c.indexB != -1 && c.indexA != -1) ? c.indexA : -1;
    OvBody A = ovLoad(d, c.indexA), B = ovLoad(d, ixB);
    V2 vA = A.v, vB = B.v; float wA = A.w, wB = B.w;
    V2 normal = c.normal;
    for (int j = 0; j < c.pointCount; j++) {
      OvPoint& cp = j == 1 ? c.p1 : c.p0;
      if (cp.relativeVelocity > -threshold || cp.totalNormalImpulse == 0.0f) continue;
      V2 rA = cp.anchorA, rB = cp.anchorB;
      V2 vrB = vB + crossSV(wB, rB), vrA = vA + crossSV(wA, rA);
      float vn = dot(vrB - vrA, normal);
      float impulse = -cp.normalMass * (vn + restitution * cp.relativeVelocity);
      float newImpulse = netMax(-cp.normalImpulse + impulse, 0.0f);     // Q3
      impulse = newImpulse - cp.normalImpulse;
      cp.normalImpulse = newImpulse; cp.totalNormalImpulse += impulse;
      V2 P = impulse * normal;
      vA = mulSub(vA, mA, P); wA -= iA * cross(rA, P); vB = mulAdd(vB, mB, P); wB += iB * cross(rB, P);
    }
    ovStore(d, c.indexA, A.dynamic, vA, wA);
    ovStore(d, ixB, B.dynamic, vB, wB);
  }
}

// StoreOverflowImpulses :323-350, one thread per constraint
__device__ void ovStoreImpulses(const WorldDev& d, int k) {
  const OvConstraint& c = d.ovCons[k];
  const int ci = d.colList[(size_t)kOverflowIndex * d.contactCapacity + k];
  float4 mn = d.mN[ci];
  const int pointCount = (int)fbits(mn.w);
  if (pointCount > 0) { float4 b = d.mB0[ci]; b.y = c.p0.normalImpulse; b.z = c.p0.tangentImpulse; b.w = c.p0.totalNormalImpulse; d.mB0[ci] = b; d.mC0[ci].z = c.p0.relativeVelocity; }
  if (pointCount > 1) { float4 b = d.mB1[ci]; b.y = c.p1.normalImpulse; b.z = c.p1.tangentImpulse; b.w = c.p1.totalNormalImpulse; d.mB1[ci] = b; d.mC1[ci].z = c.p1.relativeVelocity; }
  mn.z = c.rollingImpulse; d.mN[ci] = mn;
}

}  // namespace b2c
