// libb2cuda — island partition + island-local solver (included by b2c_step.cu after the lane functions).
// An island is a connected component of dynamic bodies under touching contacts and joints (the partition src/Island.cs
// maintains for sleeping). Constraints of different islands never share a writable body, so each island may run the
// reference's stage sequence on its own: colors in order, barrier between colors. A small island runs inside ONE CTA
// with its body rows in shared memory and its constraints in registers, so the ~120 barriers of a step are
// __syncthreads() instead of grid-wide syncs; large islands (and islands with joints or kinematic neighbours) keep the
// grid-wide kernel. Per-body and per-constraint arithmetic is identical in both paths, hence so are the results.
#pragma once

namespace b2c {

constexpr int kIslandThreads = 192;
constexpr int kIslandMaxCons = 1024;       // <= kIslandThreads: constraints live in registers; above: re-read per stage
constexpr int kForceBig = 1 << 20;

__device__ __forceinline__ int ufFind(int* parent, int x) {
  int p = parent[x];
  while (p != x) { int gp = parent[p]; if (gp != p) parent[x] = gp; x = p; p = gp; }   // path halving
  return x;
}
// Read-only find for the flatten pass: path halving there would let a slow thread overwrite an already flattened entry
// with an older, higher ancestor (the forest stays valid but is no longer flat, and every later phase assumes flat).
__device__ __forceinline__ int ufFindReadOnly(const int* parent, int x) {
  int p = parent[x];
  while (p != x) { x = p; p = parent[x]; }
  return x;
}
__device__ __forceinline__ uint32_t ufPriority(int x) { return (uint32_t)x * 0x9E3779B1u; }     // odd multiplier: a bijection on 32 bits, no ties
__device__ __forceinline__ void ufUnion(int* parent, int a, int b) {
  while (true) {
    a = ufFind(parent, a); b = ufFind(parent, b);
    if (a == b) return;
    // hook the root of larger PRIORITY under the other. The priority is a bijective scramble of the index, not the index itself: bodies of
    // a lattice are joined to their index neighbours, and hooking by index built parent chains hundreds of links long before any
    // compression (tumbler: 189 us in this loop); with scrambled priorities the trees stay logarithmic. Any total order keeps the forest
    // acyclic, and the root of a component (its minimum-priority member) is still independent of thread timing.
    const bool aHigh = ufPriority(a) > ufPriority(b);
    int hi = aHigh ? a : b, lo = aHigh ? b : a;
    int old = atomicCAS(&parent[hi], hi, lo);
    if (old == hi) return;
  }
}
__device__ __forceinline__ bool simIsDynamic(const WorldDev& d, int sim) { return sim >= 0 && (fbits(d.st0[sim].w) & kDynamicFlag) != 0; }

// grid-wide exclusive scan of int4 rows inside a cooperative kernel (chunk = blockDim.x); consumers add sums[i / blockDim.x]
template <class S> __device__ void gridScan4(const S& grid, int4* data, int4* sums, int n, int4* totalOut) {
  __shared__ int4 warpSums[32];
  const int chunk = blockDim.x;
  int chunks = (n + chunk - 1) / chunk;
  for (int c = grid.block(); c < chunks; c += grid.blocks()) {
    int i = c * chunk + threadIdx.x;
    int4 v = i < n ? data[i] : make_int4(0, 0, 0, 0);
    int4 total;
    int4 inc = blockInclusive(v, warpSums, &total);
    if (i < n) data[i] = addT(inc, make_neg(v));
    if (threadIdx.x == 0) sums[c] = total;
  }
  grid.sync();
  if (grid.block() == 0) {
    int4 carry = make_int4(0, 0, 0, 0);
    for (int base = 0; base < chunks; base += chunk) {
      int i = base + threadIdx.x;
      int4 v = i < chunks ? sums[i] : make_int4(0, 0, 0, 0);
      int4 total;
      int4 inc = blockInclusive(v, warpSums, &total);
      if (i < chunks) sums[i] = addT(carry, addT(inc, make_neg(v)));
      carry = addT(carry, total);
    }
    if (threadIdx.x == 0) *totalOut = carry;
  }
  grid.sync();
}

// Rebuilds the island partition and the constraint order [grid-path colors | island 0 | island 1 | ...] when the graph
// changed this step; a no-op (one uniform branch) otherwise. Order inside one (group, color) bucket is arbitrary:
// constraints of one color never share a writable body, so it does not influence any result.
__device__ __forceinline__ unsigned long long globalTimerI() { unsigned long long t; asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t)); return t; }
template <class S> __device__ void phase_islands(const S& grid, const WorldDev& d) {
  Scalars& s = *d.sc;
  if (!s.graphDirty) return;
  const int tid = grid.tid(), stride = grid.size();
  const int awake = d.awakeCount, nContacts = s.contactNext, nJoints = d.jointColorStart[kGraphColorCount];
  __shared__ int hist[kGraphColorCount];
  if (d.debugChecksums == 3 && tid == 0) d.sc->dbg[6] = globalTimerI();
  // P0 init
  for (int i = tid; i < awake; i += stride) { d.ufParent[i] = i; d.rootBodyCount[i] = 0; d.rootConCount[i] = 0; d.kinematicUsed[i] = 0; }
  for (int i = tid; i < awake * kGraphColorCount; i += stride) d.rootColorCount[i] = 0;
  if (tid < kGraphColorCount) { s.colorTotal[tid] = 0; s.bigCursor[tid] = 0; s.colorCount[tid] = 0; }
  if (tid == 0) { s.freeBodyCount = 0; s.bigBodyCount = 0; }
  grid.sync();
  // P1 union
  for (int i = tid; i < nContacts; i += stride) {
    if (!(d.cFlags[i] & CF_TOUCHING)) continue;
    int2 sim = d.cSim[i];
    if (simIsDynamic(d, sim.x) && simIsDynamic(d, sim.y)) ufUnion(d.ufParent, sim.x, sim.y);
  }
  for (int k = tid; k < nJoints; k += stride) {
    const JointDev& j = d.joints[d.jointSorted[k]];
    int a = d.bodySim[j.bodyIdA], b = d.bodySim[j.bodyIdB];
    if (simIsDynamic(d, a) && simIsDynamic(d, b)) ufUnion(d.ufParent, a, b);
  }
  grid.sync();
  if (d.debugChecksums == 3 && tid == 0) d.sc->dbg[7] = globalTimerI();
  // P2 flatten + bodies per root
  for (int i = tid; i < awake; i += stride) { int r = ufFindReadOnly(d.ufParent, i); d.ufParent[i] = r; atomicAdd(&d.rootBodyCount[r], 1); }
  grid.sync();
  if (d.debugChecksums == 3 && tid == 0) d.sc->dbg[8] = globalTimerI();
  // P3 constraints per root / color
  if (threadIdx.x < kGraphColorCount) hist[threadIdx.x] = 0;
  __syncthreads();
  for (int i = tid; i < nContacts; i += stride) {
    if (!(d.cFlags[i] & CF_TOUCHING)) continue;
    int c = d.cColor[i];
    atomicAdd(&hist[c], 1);
    int2 sim = d.cSim[i];
    bool dynA = simIsDynamic(d, sim.x), dynB = simIsDynamic(d, sim.y);
    int r = d.ufParent[dynA ? sim.x : sim.y];
    if (c >= kOverflowIndex) {                       // overflow constraints are solved serially by the grid path (b2c_overflow.cuh)
      if (sim.x >= 0 && !dynA) d.kinematicUsed[sim.x] = 1;
      if (sim.y >= 0 && !dynB) d.kinematicUsed[sim.y] = 1;
      atomicAdd(&d.rootConCount[r], kForceBig);
      continue;
    }
    int add = 1;
    if (sim.x >= 0 && !dynA) { d.kinematicUsed[sim.x] = 1; add += kForceBig; }
    if (sim.y >= 0 && !dynB) { d.kinematicUsed[sim.y] = 1; add += kForceBig; }
    atomicAdd(&d.rootConCount[r], add);
    atomicAdd(&d.rootColorCount[r * kGraphColorCount + c], 1);
  }
  for (int k = tid; k < nJoints; k += stride) {
    const JointDev& j = d.joints[d.jointSorted[k]];
    int a = d.bodySim[j.bodyIdA], b = d.bodySim[j.bodyIdB];
    bool dynA = simIsDynamic(d, a), dynB = simIsDynamic(d, b);
    if (a >= 0 && !dynA) d.kinematicUsed[a] = 1;
    if (b >= 0 && !dynB) d.kinematicUsed[b] = 1;
    if (dynA || dynB) atomicAdd(&d.rootConCount[d.ufParent[dynA ? a : b]], kForceBig);
  }
  __syncthreads();
  if (threadIdx.x < kGraphColorCount && hist[threadIdx.x]) atomicAdd(&s.colorTotal[threadIdx.x], hist[threadIdx.x]);
  grid.sync();
  if (d.debugChecksums == 3 && tid == 0) d.sc->dbg[9] = globalTimerI();
  // P4 classify roots
  for (int i = tid; i < awake; i += stride) {
    int4 in = make_int4(0, 0, 0, 0);
    if (d.ufParent[i] == i) {
      int bodies = d.rootBodyCount[i], cons = d.rootConCount[i];
      if (cons == 0) d.rootInfo[i] = d.kinematicUsed[i] ? -2 : -1;
      else if (bodies > d.islandMaxBodies || cons > kIslandMaxCons) {
        d.rootInfo[i] = -2;
        for (int c = 0; c < kOverflowIndex; c++) { int n = d.rootColorCount[i * kGraphColorCount + c]; if (n) atomicAdd(&s.colorCount[c], n); }
      } else { d.rootInfo[i] = 0; in = make_int4(1, bodies, cons, 0); }
    }
    d.rootScan[i] = in;
  }
  grid.sync();
  // grid-path bodies take their place in the body list in sim order (the 4th scan lane): neighbours in creation order stay neighbours,
  // which is what lets the cluster solver keep most of a constraint's rows in its own CTA (b2c_cluster_dsm.cuh)
  for (int i = tid; i < awake; i += stride) if (d.rootInfo[d.ufParent[i]] == -2) d.rootScan[i].w = 1;
  grid.sync();
  gridScan4(grid, d.rootScan, d.scanBlockSums4, awake, d.rootScan + awake);
  // P6 layout of the grid-path color ranges (padded to 32) in front of the islands
  if (tid == 0) {
    int start = 0, touching = 0;
    for (int c = 0; c < kGraphColorCount; c++) { s.colorStart[c] = start; start += (s.colorCount[c] + 31) & ~31; touching += s.colorTotal[c]; }
    int4 tot = d.rootScan[awake];
    s.smallConStart = start; s.islandCount = tot.x; s.constraintTotal = start + tot.z; s.touchingCount = touching; s.bigBodyCount = tot.w;
  }
  grid.sync();
  if (d.debugChecksums == 3 && tid == 0) d.sc->dbg[10] = globalTimerI();
  // P7 per island offsets
  const int smallConStart = s.smallConStart;
  for (int t = tid; t < smallConStart; t += stride) d.sContact[t] = -1;
  for (int i = tid; i < awake; i += stride) {
    if (d.ufParent[i] != i || d.rootInfo[i] < 0) continue;
    int4 sc = addT(d.rootScan[i], d.scanBlockSums4[i / (int)blockDim.x]);
    int id = sc.x;
    d.rootInfo[i] = id;
    d.islandBodyStart[id] = sc.y; d.islandBodyStart[id + 1] = sc.y + d.rootBodyCount[i];
    d.islandConStart[id] = smallConStart + sc.z;
    int off = 0;
    for (int c = 0; c < kGraphColorCount; c++) { d.islandColorOff[id * (kGraphColorCount + 1) + c] = off; off += d.rootColorCount[i * kGraphColorCount + c]; d.rootColorCount[i * kGraphColorCount + c] = 0; }
    d.islandColorOff[id * (kGraphColorCount + 1) + kGraphColorCount] = off;
    d.rootBodyCount[i] = 0;
  }
  grid.sync();
  // P8 bodies -> free list / grid-path list / island lists
  for (int i = tid; i < awake; i += stride) {
    int r = d.ufParent[i], info = d.rootInfo[r];
    if (info == -1) { d.freeBodies[atomicAdd(&s.freeBodyCount, 1)] = i; d.bodyLocal[i] = -1; }
    else if (info == -2) { const int k = d.rootScan[i].w + d.scanBlockSums4[i / (int)blockDim.x].w; d.bigBodies[k] = i; d.bodyLocal[i] = k; }   // slot in the grid-path body list, sim order (the cluster's shared-memory row)
    else { int k = atomicAdd(&d.rootBodyCount[r], 1); d.islandBodies[d.islandBodyStart[info] + k] = i; d.bodyLocal[i] = k; }
  }
  grid.sync();
  if (d.debugChecksums == 3 && tid == 0) d.sc->dbg[11] = globalTimerI();
  // P9 constraints -> slots
  for (int i = tid; i < nContacts; i += stride) {
    if (!(d.cFlags[i] & CF_TOUCHING)) { if (d.cFlags[i] & CF_ALIVE) d.cConstraint[i] = -1; continue; }
    int c = d.cColor[i];
    if (c >= kOverflowIndex) { d.cConstraint[i] = -1; continue; }
    int2 sim = d.cSim[i];
    int r = d.ufParent[simIsDynamic(d, sim.x) ? sim.x : sim.y];
    int info = d.rootInfo[r], t;
    if (info < 0) t = s.colorStart[c] + atomicAdd(&s.bigCursor[c], 1);
    else t = d.islandConStart[info] + d.islandColorOff[info * (kGraphColorCount + 1) + c] + atomicAdd(&d.rootColorCount[r * kGraphColorCount + c], 1);
    d.sContact[t] = i; d.cConstraint[i] = t;
    d.sLocal[t] = make_int2(sim.x >= 0 ? d.bodyLocal[sim.x] : -1, sim.y >= 0 ? d.bodyLocal[sim.y] : -1);
  }
}


// Bodies that no constraint touches: the four substeps are a private loop (IntegrateVelocities + IntegratePositions).
__device__ void freeBodies(const WorldDev& d, const StepParams& sp, int tid, int stride) {
  int n = d.sc->freeBodyCount;
  for (int k = tid; k < n; k += stride) {
    int i = d.freeBodies[k];
    float4 s0 = d.st0[i], s1 = d.st1[i], m = d.simMass[i], f = d.simForce[i];
    uint32_t sf = d.simFlags[i], sf0 = sf;
    for (int sub = 0; sub < sp.subStepCount; sub++) { s0 = integrateVelocityRow(s0, m, f, sp, &sf); integratePositionRow(s0, s1, sp); }
    d.st0[i] = s0; d.st1[i] = s1;
    if (sf != sf0) d.simFlags[i] = sf;
  }
}

// One CTA per small island: the whole SolverTask stage sequence (src/Solver.cs:599-727) with __syncthreads() as the barrier.
template <bool CACHED> __device__ void runIsland(const WorldDev& d, const StepParams& sp, int g, float4* s0, float4* s1, int* colorOff) {
  const int tid = threadIdx.x, T = blockDim.x;
  const int bBegin = d.islandBodyStart[g], nb = d.islandBodyStart[g + 1] - bBegin;
  const int cBegin = d.islandConStart[g];
  if (tid <= kGraphColorCount) colorOff[tid] = d.islandColorOff[g * (kGraphColorCount + 1) + tid];
  int sim = -1; float4 m = make_float4(0, 0, 0, 0), f = m; uint32_t sf = 0, sf0 = 0;
  if (tid < nb) { sim = d.islandBodies[bBegin + tid]; s0[tid] = d.st0[sim]; s1[tid] = d.st1[sim]; m = d.simMass[sim]; f = d.simForce[sim]; sf = sf0 = d.simFlags[sim]; }
  __syncthreads();
  const int K = colorOff[kGraphColorCount];
  BodyRows bs{s0, s1};
  // active colours of this island, kept in registers: the stage loops below touch no memory for the loop control
  int nActive = 0, myActive = -1;
  for (int k = 0; k < kOverflowIndex; k++) {
    int b = colorOff[k], e = colorOff[k + 1];
    if (b == e) continue;
    if (tid >= b && tid < e) myActive = nActive;
    if (!CACHED && tid == 0) { colorOff[kGraphColorCount + 1 + 2 * nActive] = b; colorOff[kGraphColorCount + 2 + 2 * nActive] = e; }
    nActive++;
  }
  const int* act = colorOff + kGraphColorCount + 1;          // (begin, end) per active colour, for islands too large for registers
  // PrepareContacts for this island: lanes are built straight into registers (CACHED) or into the planes (large islands).
  // Velocities read here are the pre-step ones: nothing has integrated yet.
  Lane c; float4 I = make_float4(0, 0, 0, 0), J = I; int2 loc = make_int2(-1, -1); int myContact = -1;
  // anyRest: one of this island's constraints sits in an 8-lane group (of the reference's colour arrays) that ApplyRestitution does not skip
  const bool worldRest = d.sc->anyRestitution != 0;
  bool anyRest = false;
  if (CACHED) {
    if (tid < K) {
      myContact = d.sContact[cBegin + tid];
      computeLane(d, sp, myContact, c, I, J); loc = d.sLocal[cBegin + tid];
      anyRest = worldRest && d.cRestGroup[myContact];
    } else myActive = -1;
  } else {
    for (int t = tid; t < K; t += T) { int ci = d.sContact[cBegin + t]; Lane l; float4 li, lj; computeLane(d, sp, ci, l, li, lj); writeLane(d, cBegin + t, l, li, lj); anyRest |= worldRest && d.cRestGroup[ci]; }
  }
  const bool islandRestitution = __syncthreads_or(anyRest);
  for (int sub = 0; sub < sp.subStepCount; sub++) {
    if (tid < nb) s0[tid] = integrateVelocityRow(s0[tid], m, f, sp, &sf);
    __syncthreads();
    for (int a = 0; a < nActive; a++) {
      if (CACHED) { if (myActive == a) warmStartLane(bs, c, loc.x, loc.y, I, J.z); }
      else for (int t = act[2 * a] + tid; t < act[2 * a + 1]; t += T) { Lane l; loadLane(d, cBegin + t, l, true); int2 lc = d.sLocal[cBegin + t]; warmStartLane(bs, l, lc.x, lc.y, d.sI[cBegin + t], d.sJ[cBegin + t].z); }
      __syncthreads();
    }
    for (int a = 0; a < nActive; a++) {
      if (CACHED) { if (myActive == a) solveLane<true>(bs, sp, c, loc.x, loc.y, I, J); }
      else for (int t = act[2 * a] + tid; t < act[2 * a + 1]; t += T) {
        Lane l; loadLane(d, cBegin + t, l, false); int2 lc = d.sLocal[cBegin + t]; float4 li = d.sI[cBegin + t], lj = d.sJ[cBegin + t];
        solveLane<true>(bs, sp, l, lc.x, lc.y, li, lj); d.sI[cBegin + t] = li; d.sJ[cBegin + t] = lj;
      }
      __syncthreads();
    }
    if (tid < nb) { float4 a = s0[tid], b = s1[tid]; integratePositionRow(a, b, sp); s0[tid] = a; s1[tid] = b; }
    __syncthreads();
    for (int a = 0; a < nActive; a++) {
      if (CACHED) { if (myActive == a) solveLane<false>(bs, sp, c, loc.x, loc.y, I, J); }
      else for (int t = act[2 * a] + tid; t < act[2 * a + 1]; t += T) {
        Lane l; loadLane(d, cBegin + t, l, false); int2 lc = d.sLocal[cBegin + t]; float4 li = d.sI[cBegin + t], lj = d.sJ[cBegin + t];
        solveLane<false>(bs, sp, l, lc.x, lc.y, li, lj); d.sI[cBegin + t] = li; d.sJ[cBegin + t] = lj;
      }
      __syncthreads();
    }
  }
  if (islandRestitution) {
    if (CACHED && tid < K) writeLane(d, cBegin + tid, c, I, J);                     // ApplyRestitution per colour; cRestGroup = the lane's 8-lane group is not AllZeroW
    __syncthreads();
    for (int k = 0; k < kOverflowIndex; k++) {
      if (colorOff[k] == colorOff[k + 1]) continue;
      for (int t = colorOff[k] + tid; t < colorOff[k + 1]; t += T) {
        if (!d.cRestGroup[d.sContact[cBegin + t]]) continue;
        Lane l; loadLane(d, cBegin + t, l, false); int2 lc = d.sLocal[cBegin + t]; float4 li = d.sI[cBegin + t], lj = d.sJ[cBegin + t];
        restitutionLane(bs, sp, l, lc.x, lc.y, li, lj); d.sI[cBegin + t] = li; d.sJ[cBegin + t] = lj;
      }
      __syncthreads();
    }
  }
  __syncthreads();
  if (tid < nb) { d.st0[sim] = s0[tid]; d.st1[sim] = s1[tid]; if (sf != sf0) d.simFlags[sim] = sf; }
  if (CACHED && !islandRestitution) { if (tid < K) storeImpulsesRegs(d, myContact, I, J, c.relVel1, c.relVel2); }
  else for (int t = tid; t < K; t += T) storeImpulses(d, cBegin + t);
  __syncthreads();
}

}  // namespace b2c
#include "b2c_islands_warp.cuh"
namespace b2c {

// skipWarpEligible: this step's small islands are solved by k_solver_islands_warp; only the ones that do not fit a slice stay here
__global__ void __launch_bounds__(kIslandThreads, 3) k_solver_islands(const __grid_constant__ WorldDev d, const __grid_constant__ StepParams sp, int skipWarpEligible) {
  __shared__ float4 s0[kIslandMaxBodies], s1[kIslandMaxBodies];
  __shared__ int colorOff[kGraphColorCount + 1 + 2 * kGraphColorCount];   // colour offsets, then (begin, end) of the active colours
  int n = d.sc->islandCount;
  for (int g = blockIdx.x; g < n; g += gridDim.x) {
    if (skipWarpEligible && warpIslandEligible(d, g)) continue;
    int K = d.islandColorOff[g * (kGraphColorCount + 1) + kGraphColorCount];
    if (K <= (int)blockDim.x) runIsland<true>(d, sp, g, s0, s1, colorOff);
    else runIsland<false>(d, sp, g, s0, s1, colorOff);
  }
  freeBodies(d, sp, gtid(), gsize());
}

}  // namespace b2c
