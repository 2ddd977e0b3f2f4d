// libb2cuda — throughput variant of the island solver: ONE WARP per small island.
//
// k_solver_islands (one CTA per island, lanes in registers) minimises the latency of a single island's stage chain, which is
// what a few hundred islands need. With thousands of islands (batched worlds) the machine is throughput-bound instead: a
// 192-thread CTA keeps one warp busy per stage and the register file admits three CTAs per SM, so an SM issues from ~3 warps.
// Here an island is solved by one warp: its constraint lanes live in a shared-memory slice (10 float4 planes), its body rows
// too, the colour barrier is __syncwarp(), and eight islands share an SM (eight warps, ~28 KB of shared memory each).
// Same lane functions, same per-body operation order, hence the same bits as the CTA and grid paths.
// The host picks the variant per step from the island count of the last synchronised step (World::step); either choice is
// correct: islands that do not fit a slice (more than kWarpIslandCons constraints or kWarpIslandBodies bodies) are left to
// k_solver_islands, which skips the ones handled here.
#pragma once

namespace b2c {

constexpr int kWarpIslandCons = 160, kWarpIslandBodies = 64, kWarpIslandsPerCta = 8;
struct WarpIslandSlice {
  float4 plane[10][kWarpIslandCons];     // sA..sJ of b2c_world.cuh, A.x/A.y hold island-local body indices
  float4 s0[kWarpIslandBodies], s1[kWarpIslandBodies];
  int colorOff[kGraphColorCount + 1];
  int pad[3];
};
__device__ __forceinline__ bool warpIslandEligible(const WorldDev& d, int g) {
  int nb = d.islandBodyStart[g + 1] - d.islandBodyStart[g];
  int K = d.islandColorOff[g * (kGraphColorCount + 1) + kGraphColorCount];
  return nb <= kWarpIslandBodies && K <= kWarpIslandCons;
}
__device__ __forceinline__ void sliceLoadLane(const WarpIslandSlice& sl, int t, Lane& c, bool anchorsOnly) {
  float4 A = sl.plane[0][t], B = sl.plane[1][t], E = sl.plane[4][t], F = sl.plane[5][t];
  c.indexA = (int)fbits(A.x); c.indexB = (int)fbits(A.y); c.mA = A.z; c.mB = A.w; c.iA = B.x; c.iB = B.y; c.n = mk(B.z, B.w);
  c.rA1 = mk(E.x, E.y); c.rB1 = mk(E.z, E.w); c.rA2 = mk(F.x, F.y); c.rB2 = mk(F.z, F.w);
  if (anchorsOnly) return;
  float4 C = sl.plane[2][t], D = sl.plane[3][t], G = sl.plane[6][t], H = sl.plane[7][t];
  c.friction = C.x; c.tangentSpeed = C.y; c.rollingResistance = C.z; c.rollingMass = C.w;
  c.biasRate = D.x; c.massScale = D.y; c.impulseScale = D.z; c.restitution = D.w;
  c.normalMass1 = G.x; c.tangentMass1 = G.y; c.normalMass2 = G.z; c.tangentMass2 = G.w;
  c.baseSep1 = H.x; c.baseSep2 = H.y; c.relVel1 = H.z; c.relVel2 = H.w;
}

// one island, one warp: the whole SolverTask stage sequence with __syncwarp() as the colour barrier
__device__ __forceinline__ void runIslandWarp(const WorldDev& d, const StepParams& sp, int g, WarpIslandSlice& sl, int lane) {
  const unsigned full = 0xffffffffu;
  {
    const int bBegin = d.islandBodyStart[g], nb = d.islandBodyStart[g + 1] - bBegin;
    const int cBegin = d.islandConStart[g];
    if (lane <= kGraphColorCount) sl.colorOff[lane] = d.islandColorOff[g * (kGraphColorCount + 1) + lane];
    // bodies: lane l owns rows l and l + 32
    int sim[2] = {-1, -1}; float4 m[2], f[2]; uint32_t sf[2] = {0, 0}, sf0[2] = {0, 0};
    for (int r = 0; r < 2; r++) {
      int b = lane + 32 * r;
      m[r] = make_float4(0, 0, 0, 0); f[r] = m[r];
      if (b < nb) { sim[r] = d.islandBodies[bBegin + b]; sl.s0[b] = d.st0[sim[r]]; sl.s1[b] = d.st1[sim[r]]; m[r] = d.simMass[sim[r]]; f[r] = d.simForce[sim[r]]; sf[r] = sf0[r] = d.simFlags[sim[r]]; }
    }
    __syncwarp();
    const int K = sl.colorOff[kGraphColorCount];
    BodyRows bs{sl.s0, sl.s1};
    // PrepareContacts into the slice (velocities are the pre-step ones)
    const bool worldRest = d.sc->anyRestitution != 0;
    bool anyRest = false;
    for (int t = lane; t < K; t += 32) {
      Lane c; float4 I, J;
      const int ci = d.sContact[cBegin + t];
      computeLane(d, sp, ci, c, I, J);
      int2 loc = d.sLocal[cBegin + t];
      sl.plane[0][t] = make_float4(__int_as_float(loc.x), __int_as_float(loc.y), c.mA, c.mB);
      sl.plane[1][t] = make_float4(c.iA, c.iB, c.n.x, c.n.y);
      sl.plane[2][t] = make_float4(c.friction, c.tangentSpeed, c.rollingResistance, c.rollingMass);
      sl.plane[3][t] = make_float4(c.biasRate, c.massScale, c.impulseScale, c.restitution);
      sl.plane[4][t] = make_float4(c.rA1.x, c.rA1.y, c.rB1.x, c.rB1.y);
      sl.plane[5][t] = make_float4(c.rA2.x, c.rA2.y, c.rB2.x, c.rB2.y);
      sl.plane[6][t] = make_float4(c.normalMass1, c.tangentMass1, c.normalMass2, c.tangentMass2);
      sl.plane[7][t] = make_float4(c.baseSep1, c.baseSep2, c.relVel1, c.relVel2);
      sl.plane[8][t] = I; sl.plane[9][t] = J;
      anyRest |= worldRest && d.cRestGroup[ci];
    }
    const bool islandRestitution = __any_sync(full, anyRest);
    __syncwarp();
    for (int sub = 0; sub < sp.subStepCount; sub++) {
      for (int r = 0; r < 2; r++) { int b = lane + 32 * r; if (b < nb) sl.s0[b] = integrateVelocityRow(sl.s0[b], m[r], f[r], sp, &sf[r]); }
      __syncwarp();
      for (int k = 0; k < kOverflowIndex; k++) {
        const int b = sl.colorOff[k], e = sl.colorOff[k + 1];
        if (b == e) continue;
        for (int t = b + lane; t < e; t += 32) { Lane c; sliceLoadLane(sl, t, c, true); warmStartLane(bs, c, c.indexA, c.indexB, sl.plane[8][t], sl.plane[9][t].z); }
        __syncwarp();
      }
      for (int k = 0; k < kOverflowIndex; k++) {
        const int b = sl.colorOff[k], e = sl.colorOff[k + 1];
        if (b == e) continue;
        for (int t = b + lane; t < e; t += 32) {
          Lane c; sliceLoadLane(sl, t, c, false); float4 I = sl.plane[8][t], J = sl.plane[9][t];
          solveLane<true>(bs, sp, c, c.indexA, c.indexB, I, J); sl.plane[8][t] = I; sl.plane[9][t] = J;
        }
        __syncwarp();
      }
      for (int r = 0; r < 2; r++) { int b = lane + 32 * r; if (b < nb) { float4 a = sl.s0[b], p = sl.s1[b]; integratePositionRow(a, p, sp); sl.s0[b] = a; sl.s1[b] = p; } }
      __syncwarp();
      for (int k = 0; k < kOverflowIndex; k++) {
        const int b = sl.colorOff[k], e = sl.colorOff[k + 1];
        if (b == e) continue;
        for (int t = b + lane; t < e; t += 32) {
          Lane c; sliceLoadLane(sl, t, c, false); float4 I = sl.plane[8][t], J = sl.plane[9][t];
          solveLane<false>(bs, sp, c, c.indexA, c.indexB, I, J); sl.plane[8][t] = I; sl.plane[9][t] = J;
        }
        __syncwarp();
      }
    }
    if (islandRestitution) {                       // ApplyRestitution per colour; cRestGroup = the lane's 8-lane group is not AllZeroW (as in runIsland)
      for (int k = 0; k < kOverflowIndex; k++) {
        const int b = sl.colorOff[k], e = sl.colorOff[k + 1];
        if (b == e) continue;
        for (int t = b + lane; t < e; t += 32) {
          if (!d.cRestGroup[d.sContact[cBegin + t]]) continue;
          Lane c; sliceLoadLane(sl, t, c, false); float4 I = sl.plane[8][t], J = sl.plane[9][t];
          restitutionLane(bs, sp, c, c.indexA, c.indexB, I, J); sl.plane[8][t] = I; sl.plane[9][t] = J;
        }
        __syncwarp();
      }
    }
    for (int r = 0; r < 2; r++) { int b = lane + 32 * r; if (b < nb) { d.st0[sim[r]] = sl.s0[b]; d.st1[sim[r]] = sl.s1[b]; if (sf[r] != sf0[r]) d.simFlags[sim[r]] = sf[r]; } }
    for (int t = lane; t < K; t += 32) { float4 H = sl.plane[7][t]; storeImpulsesRegs(d, d.sContact[cBegin + t], sl.plane[8][t], sl.plane[9][t], H.z, H.w); }
    __syncwarp();
  }
}

__global__ void __launch_bounds__(32 * kWarpIslandsPerCta, 1) k_solver_islands_warp(const __grid_constant__ WorldDev d, const __grid_constant__ StepParams sp) {
  extern __shared__ __align__(16) unsigned char smemRaw[];
  WarpIslandSlice& sl = reinterpret_cast<WarpIslandSlice*>(smemRaw)[threadIdx.x >> 5];
  const int lane = threadIdx.x & 31;
  const int warpsTotal = gridDim.x * kWarpIslandsPerCta, warpId = blockIdx.x * kWarpIslandsPerCta + (threadIdx.x >> 5);
  const int n = d.sc->islandCount;
  for (int g = warpId; g < n; g += warpsTotal) {
    if (!warpIslandEligible(d, g)) continue;
    runIslandWarp(d, sp, g, sl, lane);
  }
}

// Batches of independent worlds (k_batch_front / k_batch_finalize in b2c_step.cu): ONE WARP per world walks the world's islands that fit a
// slice; eight worlds share an SM. The islands that do not fit were solved by the world's CTA in k_batch_front.
__global__ void __launch_bounds__(32 * kWarpIslandsPerCta, 1) k_batch_solve_warp(const WorldDev* __restrict__ worlds, int worldCount, const __grid_constant__ StepParams sp) {
  extern __shared__ __align__(16) unsigned char smemRaw[];
  const int w = threadIdx.x >> 5, lane = threadIdx.x & 31;
  WarpIslandSlice& sl = reinterpret_cast<WarpIslandSlice*>(smemRaw)[w];
  const int world = blockIdx.x * kWarpIslandsPerCta + w;
  if (world >= worldCount) return;
  const WorldDev& d = worlds[world];                 // read through L1 (the slices leave no shared memory for a copy of the record)
  const int n = d.sc->islandCount;
  for (int g = 0; g < n; g++) {
    if (!warpIslandEligible(d, g)) continue;
    runIslandWarp(d, sp, g, sl, lane);
  }
}

}  // namespace b2c
