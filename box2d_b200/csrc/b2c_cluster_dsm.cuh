// SolverTask substep loop (src/Solver.cs:599-727) for ONE large island of contacts, resident in a thread-block cluster's shared memory.
//
// The grid path pays a grid barrier (1.2 us) and an L2 round trip for the body rows in every colour stage (large_pyramid: 152 stages x
// 1.9 us). Here the stage touches no L2 at all: the body rows (BodyState, 32 B) of the island are spread over the shared memory of the
// cluster's CTAs (row k lives in CTA k / rowsPerCta: neighbours in creation order share a CTA) and are read and written through distributed shared memory (mapa + ld/st.shared::cluster),
// every thread keeps its constraint lanes (up to kDsmLanes) in its own CTA's shared memory, and the stage barrier is barrier.cluster.
// Eligible (decided on the device, stepFront): contacts only (no joints, no overflow colour, no restitution group), bodies and padded
// constraints within the cluster's rows and lanes. Arithmetic and order are those of phase_solver_grid: same lane functions, same stages.
#pragma once

namespace b2c {

constexpr int kDsmRows = 768;         // body rows per CTA (16 CTAs: 12 288 bodies)
constexpr int kDsmLanes = 2;          // constraint lanes per thread (16 CTAs x 512 threads: 16 384 padded constraints)
constexpr int kDsmPlanes = 12;        // float4 planes of one lane: Lane + impulses as writeLane lays them out, then the lane's copy of its two bodies' (dp, dq) rows
constexpr size_t kDsmSmemBytes = sizeof(float4) * (2 * (size_t)kDsmRows + (size_t)kDsmPlanes * kDsmLanes * kCoopThreads) + sizeof(int) * (size_t)kDsmLanes * kCoopThreads;   // rows, lane planes, lane -> constraint

__device__ __forceinline__ uint32_t smemAddr(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

// BodyRows over the cluster. A lane names a body by (CTA rank << 16 | row), worked out once when the lane is built; -1 = not in the awake set.
// The two row arrays sit at the same shared-memory offset in every CTA.
struct DsmRows {
  uint32_t base0, base1;              // shared-window addresses of this CTA's st0 / st1 rows
  __device__ __forceinline__ uint32_t remote(uint32_t base, int where) const {      // where = CTA rank << 16 | row
    uint32_t local = base + ((uint32_t)where & 0xffffu) * 16u, rank = (uint32_t)where >> 16, out;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(out) : "r"(local), "r"(rank));
    return out;
  }
  __device__ __forceinline__ float4 ld(uint32_t a) const {
    float4 v; asm volatile("ld.shared::cluster.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(a) : "memory");
    return v;
  }
  __device__ __forceinline__ void st(uint32_t a, float4 v) const {
    asm volatile("st.shared::cluster.v4.f32 [%0], {%1, %2, %3, %4};" :: "r"(a), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
  }
  __device__ __forceinline__ BodyW load(int slot) const {
    BodyW b;
    if (slot < 0) { b.v = mk(0, 0); b.w = 0; b.flags = 0; b.dp = mk(0, 0); b.dq = mkRot(1, 0); return b; }
    float4 a = ld(remote(base0, slot)), c = ld(remote(base1, slot));
    b.v = mk(a.x, a.y); b.w = a.z; b.flags = fbits(a.w); b.dp = mk(c.x, c.y); b.dq = mkRot(c.z, c.w);
    return b;
  }
  __device__ __forceinline__ void store(int slot, const BodyW& b) const {
    if (slot >= 0 && (b.flags & kDynamicFlag)) st(remote(base0, slot), make_float4(b.v.x, b.v.y, b.w, __uint_as_float(b.flags)));
  }
};

// The rows as one lane sees them in one kind of stage. Distributed shared memory moves ~17 B per cycle per SM and a stage is bound by it, so
// a lane fetches only what the stage needs: warm start reads no (dp, dq) at all; (dp, dq) change only in IntegratePositions, so the relax
// stage reads them from the rows and leaves a copy in the lane for the solve stage of the next substep.
template <int MODE>     // 0: velocities only, 1: (dp, dq) from the lane's copy, 2: (dp, dq) from the rows, copied to the lane
struct DsmLaneRows {
  DsmRows r; float4* copyA; float4* copyB;
  __device__ __forceinline__ BodyW get(int slot, float4* copy) const {
    BodyW b;
    if (slot < 0) { b.v = mk(0, 0); b.w = 0; b.flags = 0; b.dp = mk(0, 0); b.dq = mkRot(1, 0); return b; }
    const float4 a = r.ld(r.remote(r.base0, slot));
    float4 c = make_float4(0.0f, 0.0f, 1.0f, 0.0f);
    if (MODE == 1) c = *copy;
    if (MODE == 2) { c = r.ld(r.remote(r.base1, slot)); *copy = c; }
    b.v = mk(a.x, a.y); b.w = a.z; b.flags = fbits(a.w); b.dp = mk(c.x, c.y); b.dq = mkRot(c.z, c.w);
    return b;
  }
  __device__ __forceinline__ BodyW loadA(int slot) const { return get(slot, copyA); }
  __device__ __forceinline__ BodyW loadB(int slot) const { return get(slot, copyB); }
  __device__ __forceinline__ void store(int slot, const BodyW& b) const { r.store(slot, b); }
};

// one lane <-> this thread's column of the CTA's lane planes (plane p, lane j: planes[(p * kDsmLanes + j) * blockDim + threadIdx])
struct DsmLanes {
  float4* planes;
  __device__ __forceinline__ float4& at(int p, int j) const { return planes[(p * kDsmLanes + j) * kCoopThreads + threadIdx.x]; }
  __device__ __forceinline__ void write(int j, const Lane& c, float4 I, float4 J) const {
    at(0, j) = make_float4(__int_as_float(c.indexA), __int_as_float(c.indexB), c.mA, c.mB);
    at(1, j) = make_float4(c.iA, c.iB, c.n.x, c.n.y);
    at(2, j) = make_float4(c.friction, c.tangentSpeed, c.rollingResistance, c.rollingMass);
    at(3, j) = make_float4(c.biasRate, c.massScale, c.impulseScale, c.restitution);
    at(4, j) = make_float4(c.rA1.x, c.rA1.y, c.rB1.x, c.rB1.y);
    at(5, j) = make_float4(c.rA2.x, c.rA2.y, c.rB2.x, c.rB2.y);
    at(6, j) = make_float4(c.normalMass1, c.tangentMass1, c.normalMass2, c.tangentMass2);
    at(7, j) = make_float4(c.baseSep1, c.baseSep2, c.relVel1, c.relVel2);
    at(8, j) = I; at(9, j) = J;
  }
  __device__ __forceinline__ void read(int j, Lane& c, bool anchorsOnly) const {
    float4 A = at(0, j), B = at(1, j), E = at(4, j), F = at(5, j);
    c.indexA = (int)fbits(A.x); c.indexB = (int)fbits(A.y); c.mA = A.z; c.mB = A.w; c.iA = B.x; c.iB = B.y; c.n = mk(B.z, B.w);
    c.rA1 = mk(E.x, E.y); c.rB1 = mk(E.z, E.w); c.rA2 = mk(F.x, F.y); c.rB2 = mk(F.z, F.w);
    if (anchorsOnly) return;
    float4 C = at(2, j), D = at(3, j), G = at(6, j), H = at(7, j);
    c.friction = C.x; c.tangentSpeed = C.y; c.rollingResistance = C.z; c.rollingMass = C.w;
    c.biasRate = D.x; c.massScale = D.y; c.impulseScale = D.z; c.restitution = D.w;
    c.normalMass1 = G.x; c.tangentMass1 = G.y; c.normalMass2 = G.z; c.tangentMass2 = G.w;
    c.baseSep1 = H.x; c.baseSep2 = H.y; c.relVel1 = H.z; c.relVel2 = H.w;
  }
};

__device__ __forceinline__ bool dsmEligible(const WorldDev& d, const StepParams& sp) {
  const Scalars& s = *d.sc;
  return sp.clusterDsm != 0 && d.jointColorStart[kGraphColorCount] == 0 && s.colLen[kOverflowIndex] == 0 && s.anyRestitution == 0 &&
         s.smallConStart > 0 && s.smallConStart <= kDsmLanes * sp.clusterThreads && s.bigBodyCount <= kDsmRows * (sp.clusterThreads / kCoopThreads) &&
         d.debugChecksums == 0;
}

// Runs inside k_solver_cluster (one cluster = the whole grid). `smem` is the kernel's dynamic shared memory (kDsmSmemBytes).
// Returns false (on every thread of the cluster, nothing touched) when one CTA would own more constraints than it has lanes: the caller
// then runs the plain cluster path.
__device__ bool phase_solver_dsm(const ClusterScope& cs, const WorldDev& d, const StepParams& sp, unsigned char* smem) {
  Scalars& s = *d.sc;
  __shared__ int sBegin[kGraphColorCount], sEnd[kGraphColorCount], sActiveCount;
  __shared__ int sHist[kGraphColorCount], sOff[kGraphColorCount + 1], sCursor[kGraphColorCount];
  float4* rows0 = reinterpret_cast<float4*>(smem); float4* rows1 = rows0 + kDsmRows;
  const DsmLanes lanes{rows1 + kDsmRows};
  int* laneCon = reinterpret_cast<int*>(lanes.planes + (size_t)kDsmPlanes * kDsmLanes * kCoopThreads);   // constraint of every lane slot of this CTA
  const int nCta = (int)gridDim.x, rank = (int)blockIdx.x;
  const int nBodies = s.bigBodyCount, nCons = s.smallConStart;
  const DsmRows bs{smemAddr(rows0), smemAddr(rows1)};
  __shared__ int sFirst[17], sWarp[kCoopThreads / 32], sTotal;
  if (threadIdx.x == 0) {
    int n = 0;
    for (int c = 0; c < kOverflowIndex; c++) { int cc = s.colorCount[c]; if (cc > 0) { sBegin[n] = s.colorStart[c]; sEnd[n] = s.colorStart[c] + cc; n++; } }
    sActiveCount = n;
  }
  if (threadIdx.x < kGraphColorCount) { sHist[threadIdx.x] = 0; sCursor[threadIdx.x] = 0; }
  if (threadIdx.x <= 16) sFirst[threadIdx.x] = nBodies;
  // ---- who owns what. Body slots are in sim order (phase_islands): neighbours in creation order are neighbours here. The slots are cut into
  // one CONTIGUOUS range per CTA, the cuts placed so that every CTA owns about the same number of constraints (a constraint belongs to the
  // CTA of its first awake body). Most of a constraint's row traffic then stays inside its CTA; only what crosses a cut uses the cluster
  // network (~17 B per cycle per SM: a round-robin placement of rows and lanes was bound by exactly that).
  int* cnt = d.bodyEnlargedCount;                    // [awake + 1] scratch of FinalizeBodies, free until then
  for (int k = cs.tid(); k < nBodies; k += cs.size()) cnt[k] = 0;
  cs.sync();
  for (int t = cs.tid(); t < nCons; t += cs.size()) {
    if (d.sContact[t] < 0) continue;                 // padding of a colour range
    const int2 loc = d.sLocal[t];
    atomicAdd(&cnt[loc.x >= 0 ? loc.x : loc.y], 1);
  }
  cs.sync();
  int* pre = reinterpret_cast<int*>(lanes.planes);   // [nBodies] exclusive prefix of cnt: lives in the lane planes until the lanes are written
  {
    const int chunk = (nBodies + (int)blockDim.x - 1) / (int)blockDim.x, b0 = min((int)threadIdx.x * chunk, nBodies), b1 = min(b0 + chunk, nBodies);
    int sum = 0;
    for (int k = b0; k < b1; k++) sum += __ldcg(&cnt[k]);
    int inc = sum;
    for (int o = 1; o < 32; o <<= 1) { int v = __shfl_up_sync(0xffffffffu, inc, o); if ((int)(threadIdx.x & 31) >= o) inc += v; }
    if ((threadIdx.x & 31) == 31) sWarp[threadIdx.x >> 5] = inc;
    __syncthreads();
    if (threadIdx.x == 0) { int run = 0; for (int w = 0; w < (int)blockDim.x / 32; w++) { int v = sWarp[w]; sWarp[w] = run; run += v; } sTotal = run; }
    __syncthreads();
    int run = sWarp[threadIdx.x >> 5] + inc - sum;
    for (int k = b0; k < b1; k++) { pre[k] = run; run += __ldcg(&cnt[k]); }
  }
  __syncthreads();
  const int per = max(1, (sTotal + nCta - 1) / nCta);
  for (int k = threadIdx.x; k < nBodies; k += blockDim.x) {
    const int r = min(nCta - 1, pre[k] / per), rp = k > 0 ? min(nCta - 1, pre[k - 1] / per) : -1;
    for (int q = rp + 1; q <= r; q++) sFirst[q] = k;
  }
  __syncthreads();
  const int activeCount = sActiveCount;
  const int myFirst = sFirst[rank], myRows = sFirst[rank + 1] - myFirst;
  // pass 1: constraints this CTA owns, per active colour
  for (int t = threadIdx.x; t < nCons; t += blockDim.x) {
    if (d.sContact[t] < 0) continue;
    const int2 loc = d.sLocal[t];
    const int owner = loc.x >= 0 ? loc.x : loc.y;
    if (owner < myFirst || owner >= myFirst + myRows) continue;
    for (int a = 0; a < activeCount; a++) if (t >= sBegin[a] && t < sEnd[a]) { atomicAdd(&sHist[a], 1); break; }
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    int off = 0;
    for (int a = 0; a < activeCount; a++) { sOff[a] = off; off += sHist[a]; }
    sOff[activeCount] = off;
    if (off > kDsmLanes * kCoopThreads || myRows > kDsmRows || sp.clusterDsm == 2) atomicExch(&s.clusterSolve, 3);      // more than this CTA has lanes or rows for (or the test switch)
  }
  cs.sync();
  if (*(volatile int*)&s.clusterSolve == 3) return false;
  // pass 2: lane slots grouped by colour (order inside a colour is free: its constraints share no writable body)
  for (int t = threadIdx.x; t < nCons; t += blockDim.x) {
    if (d.sContact[t] < 0) continue;
    const int2 loc = d.sLocal[t];
    const int owner = loc.x >= 0 ? loc.x : loc.y;
    if (owner < myFirst || owner >= myFirst + myRows) continue;
    for (int a = 0; a < activeCount; a++) if (t >= sBegin[a] && t < sEnd[a]) { laneCon[sOff[a] + atomicAdd(&sCursor[a], 1)] = t; break; }
  }
  // body rows of this CTA; mass, force and flags of the rows a thread integrates stay in its registers
  constexpr int kRowsPerThread = (kDsmRows + kCoopThreads - 1) / kCoopThreads;
  int rowSim[kRowsPerThread]; float4 rowM[kRowsPerThread], rowF[kRowsPerThread]; uint32_t rowSf[kRowsPerThread], rowSf0[kRowsPerThread];
#pragma unroll
  for (int k = 0; k < kRowsPerThread; k++) {
    const int r = threadIdx.x + k * kCoopThreads;
    rowSim[k] = -1;
    if (r < myRows) {
      const int sim = d.bigBodies[myFirst + r];
      rowSim[k] = sim; rows0[r] = d.st0[sim]; rows1[r] = d.st1[sim]; rowM[k] = d.simMass[sim]; rowF[k] = d.simForce[sim]; rowSf[k] = rowSf0[k] = d.simFlags[sim];
    }
  }
  __syncthreads();
  // the lanes of this thread: constraint, colour, and where its two bodies' rows are (needs the prefix, which the lanes are about to overwrite)
  const int nLanes = sOff[activeCount];
  int myA[kDsmLanes], myCon[kDsmLanes], whereA[kDsmLanes], whereB[kDsmLanes];
#pragma unroll
  for (int j = 0; j < kDsmLanes; j++) {
    myA[j] = -1; myCon[j] = -1; whereA[j] = -1; whereB[j] = -1;
    const int ls = j * kCoopThreads + (int)threadIdx.x;
    if (ls >= nLanes) continue;
    const int t = laneCon[ls];
    for (int a = 0; a < activeCount; a++) if (ls >= sOff[a] && ls < sOff[a + 1]) myA[j] = a;
    myCon[j] = d.sContact[t];
    const int2 loc = d.sLocal[t];
    if (loc.x >= 0) { const int r = min(nCta - 1, pre[loc.x] / per); whereA[j] = (r << 16) | (loc.x - sFirst[r]); }
    if (loc.y >= 0) { const int r = min(nCta - 1, pre[loc.y] / per); whereB[j] = (r << 16) | (loc.y - sFirst[r]); }
  }
  __syncthreads();
  // PrepareContacts (reads pre-step velocities from the global rows)
#pragma unroll
  for (int j = 0; j < kDsmLanes; j++) {
    if (myCon[j] < 0) continue;
    Lane c; float4 I, J;
    computeLane(d, sp, myCon[j], c, I, J);
    lanes.at(10, j) = c.indexA >= 0 ? d.st1[c.indexA] : make_float4(0.0f, 0.0f, 1.0f, 0.0f);   // (dp, dq) the first solve stage sees
    lanes.at(11, j) = c.indexB >= 0 ? d.st1[c.indexB] : make_float4(0.0f, 0.0f, 1.0f, 0.0f);
    c.indexA = whereA[j]; c.indexB = whereB[j];
    lanes.write(j, c, I, J);
  }
  cs.sync();
  for (int sub = 0; sub < sp.subStepCount; sub++) {
#pragma unroll
    for (int k = 0; k < kRowsPerThread; k++) if (rowSim[k] >= 0) { const int r = threadIdx.x + k * kCoopThreads; rows0[r] = integrateVelocityRow(rows0[r], rowM[k], rowF[k], sp, &rowSf[k]); }
    cs.sync();
    for (int a = 0; a < activeCount; a++) {
#pragma unroll
      for (int j = 0; j < kDsmLanes; j++) if (myA[j] == a) { Lane c; lanes.read(j, c, true); warmStartLane(DsmLaneRows<0>{bs, nullptr, nullptr}, c, c.indexA, c.indexB, lanes.at(8, j), lanes.at(9, j).z); }
      cs.sync();
    }
    for (int a = 0; a < activeCount; a++) {
#pragma unroll
      for (int j = 0; j < kDsmLanes; j++) if (myA[j] == a) {
        Lane c; lanes.read(j, c, false); float4 I = lanes.at(8, j), J = lanes.at(9, j);
        solveLane<true>(DsmLaneRows<1>{bs, &lanes.at(10, j), &lanes.at(11, j)}, sp, c, c.indexA, c.indexB, I, J);
        lanes.at(8, j) = I; lanes.at(9, j) = J;
      }
      cs.sync();
    }
#pragma unroll
    for (int k = 0; k < kRowsPerThread; k++) if (rowSim[k] >= 0) {
      const int r = threadIdx.x + k * kCoopThreads;
      float4 a = rows0[r], b = rows1[r];
      integratePositionRow(a, b, sp);
      rows0[r] = a; rows1[r] = b;
    }
    cs.sync();
    for (int a = 0; a < activeCount; a++) {
#pragma unroll
      for (int j = 0; j < kDsmLanes; j++) if (myA[j] == a) {
        Lane c; lanes.read(j, c, false); float4 I = lanes.at(8, j), J = lanes.at(9, j);
        solveLane<false>(DsmLaneRows<2>{bs, &lanes.at(10, j), &lanes.at(11, j)}, sp, c, c.indexA, c.indexB, I, J);
        lanes.at(8, j) = I; lanes.at(9, j) = J;
      }
      cs.sync();
    }
  }
  // (no restitution stage: a world with a restitution group is not eligible) StoreImpulses :990-1025, then the rows go home
#pragma unroll
  for (int j = 0; j < kDsmLanes; j++) {
    if (myCon[j] < 0) continue;
    const float4 H = lanes.at(7, j);
    storeImpulsesRegs(d, myCon[j], lanes.at(8, j), lanes.at(9, j), H.z, H.w);
  }
#pragma unroll
  for (int k = 0; k < kRowsPerThread; k++) if (rowSim[k] >= 0) {
    const int r = threadIdx.x + k * kCoopThreads;
    d.st0[rowSim[k]] = rows0[r]; d.st1[rowSim[k]] = rows1[r];
    if (rowSf[k] != rowSf0[k]) d.simFlags[rowSim[k]] = rowSf[k];
  }
  cs.sync();                                          // no CTA may leave while a peer can still address its shared memory
  return true;
}

}  // namespace b2c
