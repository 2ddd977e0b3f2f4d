// libb2cuda — wide broadphase queries (part of UpdateBroadPhasePairs, src/BroadPhase.cs:197-232).
//
// FindPairsTask gives each moved proxy one DynamicTree.Query. One device thread per query is the right grain while the
// fat box touches a handful of leaves, but a large kinematic/dynamic shape (the tumbler's walls: thousands of hits each)
// would serialise thousands of callbacks on one thread. Such queries are detected by a node-visit budget in the count
// pass (p_find_pairs<false>) and answered here by the whole grid: every leaf of the queried tree is tested in the order
// the reference's DFS would reach it (child2 first, src/DynamicTree.cs:531-557), accepted hits are compacted with a
// prefix sum, so the hit order — hence contact creation order and contact ids — is unchanged. A leaf overlaps the query
// box only if all its ancestors do, so the set of hits equals the pruned traversal's.
//
// Included by b2c_step.cu after b2c_rebuild.cuh (shares its scratch arrays: subCount, visit, leafIdx are free here
// because the tree rebuild runs after pair finding).
#pragma once

namespace b2c {

// DFS (child1-first) rank of every leaf of the trees; t.leafIdx[rank] = leaf node. The static tree never changes after
// the upload, so its ranks are computed once.
template <class S> __device__ void phase_leaf_ranks(const S& grid, const WorldDev& d, bool needStatic) {
  const int tid = grid.tid(), stride = grid.size();
  for (int tt = 0; tt < 3; tt++) {
    if (tt == kStaticBody && !needStatic) continue;
    const TreeDev& t = d.trees[tt];
    for (int n = tid; n < t.capacity; n += stride) { t.visit[n] = 0; t.subCount[n] = 0; }
  }
  grid.sync();
  for (int tt = 0; tt < 3; tt++) {                 // leaves per subtree, bottom-up: the second arrival at a node continues
    if (tt == kStaticBody && !needStatic) continue;
    const TreeDev& t = d.trees[tt];
    for (int n = tid; n < t.capacity; n += stride) {
      uint32_t flags = t.hf[n] >> 16;
      if (!(flags & NODE_ALLOCATED) || !(flags & NODE_LEAF)) continue;
      t.subCount[n] = 1;
      int cur = n;
      while (true) {
        int p = t.parent[cur];
        if (p == -1) break;
        __threadfence();
        if (atomicAdd(&t.visit[p], 1) == 0) break;
        __threadfence();
        int2 c = t.child[p];
        t.subCount[p] = ((volatile int*)t.subCount)[c.x] + ((volatile int*)t.subCount)[c.y];
        cur = p;
      }
    }
  }
  grid.sync();
  for (int tt = 0; tt < 3; tt++) {
    if (tt == kStaticBody && !needStatic) continue;
    const TreeDev& t = d.trees[tt];
    for (int n = tid; n < t.capacity; n += stride) {
      uint32_t flags = t.hf[n] >> 16;
      if (!(flags & NODE_ALLOCATED) || !(flags & NODE_LEAF)) continue;
      int r = 0, cur = n;
      while (true) { int p = t.parent[cur]; if (p == -1) break; int2 c = t.child[p]; if (c.y == cur) r += t.subCount[c.x]; cur = p; }
      t.leafIdx[r] = n;
    }
  }
  if (tid == 0 && needStatic) d.sc->staticRanksValid = 1;
  grid.sync();
}

__device__ __forceinline__ int bigSegLen(const WorldDev& d, int queryType) {
  int nd = d.trees[kDynamicBody].meta[3];
  return queryType == kDynamicBody ? nd + d.trees[kKinematicBody].meta[3] + d.trees[kStaticBody].meta[3] : nd;
}
// largest b in [lo, hi) with arr[b] <= x
__device__ __forceinline__ int lastNotAbove(const int* arr, int lo, int hi, int x) {
  while (hi - lo > 1) { int mid = (lo + hi) >> 1; if (arr[mid] <= x) lo = mid; else hi = mid; }
  return lo;
}
struct BigItem { int queryKey, queryShape, treeType, leaf; Box fat; };
// sweep position -> (wide query, tree, leaf). A dynamic querier visits the kinematic, static, dynamic trees in that
// order (:216-224), others only the dynamic tree; inside a tree the DFS reaches leaves in reverse child1-first rank.
__device__ __forceinline__ BigItem bigDecode(const WorldDev& d, int g0, int g1, int pos) {
  BigItem it;
  int b = lastNotAbove(d.bigSegStart, g0, g1, pos);
  it.queryKey = d.moveArray[d.bigList[b]];
  int type = it.queryKey & 3, id = it.queryKey >> 2;
  it.fat = toBox(d.trees[type].aabb[id]);
  it.queryShape = d.trees[type].user[id];
  int local = pos - d.bigSegStart[b];
  it.treeType = kDynamicBody;
  if (type == kDynamicBody) {
    int nk = d.trees[kKinematicBody].meta[3], ns = d.trees[kStaticBody].meta[3];
    if (local < nk) it.treeType = kKinematicBody;
    else if (local < nk + ns) { it.treeType = kStaticBody; local -= nk; }
    else local -= nk + ns;
  }
  const TreeDev& t = d.trees[it.treeType];
  it.leaf = t.leafIdx[t.meta[3] - 1 - local];
  return it;
}

// Runs between the count pass and the offsets scan. Uniform control flow: every thread of the grid calls it.
template <class S> __device__ void phase_big_queries(const S& grid, const WorldDev& d) {
  const int tid = grid.tid(), stride = grid.size();
  Scalars& s = *d.sc;
  const int nBig = s.bigQueryCount < kMaxBigQueries ? s.bigQueryCount : kMaxBigQueries;
  if (nBig == 0) return;
  phase_leaf_ranks(grid, d, s.staticRanksValid == 0);
  if (tid == 0) {
    int off = 0;
    for (int b = 0; b < nBig; b++) { d.bigSegStart[b] = off; off += bigSegLen(d, d.moveArray[d.bigList[b]] & 3); }
    d.bigSegStart[nBig] = off;
  }
  grid.sync();
  int hitBase = 0, g0 = 0;
  while (g0 < nBig) {                              // groups of queries whose leaf segments fit the sweep scratch
    const int base = d.bigSegStart[g0];
    int g1 = g0 + 1;
    while (g1 < nBig && d.bigSegStart[g1 + 1] - base <= d.bigScratch) g1++;
    const int len = d.bigSegStart[g1] - base;
    for (int e = tid; e < len; e += stride) {
      BigItem it = bigDecode(d, g0, g1, base + e);
      const TreeDev& t = d.trees[it.treeType];
      int a, b2;
      bool f = boxOverlaps(toBox(t.aabb[it.leaf]), it.fat) &&
               acceptPair(d, it.queryKey, it.queryKey & 3, it.queryShape, it.leaf, it.treeType, t.user[it.leaf], &a, &b2);
      d.bigFlag[e] = f ? 1 : 0; d.bigScan[e] = f ? 1 : 0;
    }
    grid.sync();
    gridScan<int>(grid, d.bigScan, d.scanBlockSums, len, &s.bigGroupTotal);
    for (int e = tid; e < len; e += stride) {
      if (!d.bigFlag[e]) continue;
      int pos = hitBase + d.bigScan[e] + d.scanBlockSums[e / (int)blockDim.x];
      if (pos >= d.pairCapacity) { atomicOr(&s.errorFlags, (int)ERR_PAIR_CAPACITY); continue; }
      BigItem it = bigDecode(d, g0, g1, base + e);
      int hitShape = d.trees[it.treeType].user[it.leaf];
      int proxyKey = (it.leaf << 2) | it.treeType;
      d.bigHits[pos] = proxyKey < it.queryKey ? make_int2(hitShape, it.queryShape) : make_int2(it.queryShape, hitShape);
    }
    for (int b = g0 + tid; b < g1; b += stride) {
      int e0 = d.bigSegStart[b] - base, e1 = d.bigSegStart[b + 1] - base;
      int p0 = d.bigScan[e0] + d.scanBlockSums[e0 / (int)blockDim.x];
      int p1 = e1 < len ? d.bigScan[e1] + d.scanBlockSums[e1 / (int)blockDim.x] : s.bigGroupTotal;
      d.bigOff[b] = hitBase + p0;
      d.moveHitCount[d.bigList[b]] = p1 - p0;
    }
    grid.sync();
    hitBase += s.bigGroupTotal;
    g0 = g1;
  }
  if (tid == 0) { s.bigHitTotal = hitBase < d.pairCapacity ? hitBase : d.pairCapacity; d.bigOff[nBig] = hitBase; }
}

// After the offsets scan: hits of wide queries go to their place in creation order (reverse hit order inside a query).
__device__ void p_big_fill(const WorldDev& d, int tid, int stride) {
  Scalars& s = *d.sc;
  const int nBig = s.bigQueryCount < kMaxBigQueries ? s.bigQueryCount : kMaxBigQueries;
  if (nBig == 0) return;
  const int total = s.bigHitTotal, n = s.moveCount;
  for (int e = tid; e < total; e += stride) {
    int b = lastNotAbove(d.bigOff, 0, nBig, e);
    int i = d.bigList[b], h = e - d.bigOff[b];
    int end = (i + 1 < n) ? d.moveHitCount[i + 1] + d.scanBlockSums[(i + 1) / (int)blockDim.x] : s.newPairCount;
    if (end > d.pairCapacity) { atomicOr(&s.errorFlags, (int)ERR_PAIR_CAPACITY); continue; }
    d.movePairs[end - 1 - h] = d.bigHits[e];
  }
}

}  // namespace b2c
