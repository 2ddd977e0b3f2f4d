// libb2cuda — DynamicTree.Rebuild(false) (src/DynamicTree.cs:802-1000) as a grid-wide parallel algorithm that yields the SAME
// topology as the sequential C#: (1) the frontier (leaves and untouched subtrees under enlarged ancestors) is gathered in the
// reference's DFS order (child1 first) with subtree counts + per-node rank walks instead of a stack; (2) BuildTree's top-down
// median split runs level by level over all ranges at once — PartitionMid's two-pointer swap loop has the closed form
// "k-th misplaced element from the left swaps with the k-th misplaced element from the right", evaluated with one prefix sum
// per level; ranges of <= kSerialRange leaves are finished by one warp each, in registers; (3) boxes, heights
// and category bits of the new internal nodes are filled bottom-up (second arrival computes). Which free node id an internal
// node receives is not reproduced (ids of internal nodes are not observable; leaves keep their proxy ids).
#pragma once

namespace b2c {

constexpr int kSerialRange = 32;

__device__ __forceinline__ void rbAtomicMinF(float* addr, float v) {
  v = v + 0.0f;
  if (v >= 0.0f) atomicMin((int*)addr, __float_as_int(v)); else atomicMax((unsigned*)addr, __float_as_uint(v));
}
__device__ __forceinline__ void rbAtomicMaxF(float* addr, float v) {
  v = v + 0.0f;
  if (v >= 0.0f) atomicMax((int*)addr, __float_as_int(v)); else atomicMin((unsigned*)addr, __float_as_uint(v));
}
__device__ __forceinline__ void rbInitNode(const TreeDev& t, int id) {      // AllocateNode :54-72 (node = new())
  t.aabb[id] = make_float4(0, 0, 0, 0); t.child[id] = make_int2(-1, -1); t.parent[id] = -1; t.hf[id] = (uint32_t)NODE_ALLOCATED << 16;
  t.cat[id] = ~0ull; t.user[id] = -1; t.visit[id] = 0;
}
__device__ __forceinline__ int rbAllocNode(const TreeDev& t) { int id = t.freeIds[atomicAdd(&t.rb[2], 1)]; rbInitNode(t, id); return id; }

// bottom-up completion: called once per finished child; the second arrival at a node computes it and continues upward
__device__ void rbWalkUp(const TreeDev& t, int cur) {
  for (int guard = 0; guard < 4096; guard++) {     // a tree is never this deep; the bound turns a corrupted parent chain into a stop, not a hang
    int p = t.parent[cur];
    if (p == -1) return;
    __threadfence();
    if (atomicAdd(&t.visit[p], 1) == 0) return;
    __threadfence();
    refitFromChildren(t, p);
    cur = p;
  }
}

// BuildTree (:853-959) for a range of <= 32 frontier nodes whose root node is already allocated and linked to its parent,
// by ONE WARP with the range in registers (lane p = position p). Every level all ranges split at once: the center bounds
// are a segmented reduction over lanes, PartitionMid's swaps (k-th misplaced-left with k-th misplaced-right, counted
// with ballots) are two shuffles, child ids come from a block of free ids reserved with a single atomic. Boxes, heights
// and category bits of the new nodes are computed after the last level from the final arrangement: the node whose split
// falls before position j is finished by lane j (each internal node of a binary tree over a sequence has its own split).
// No load in here depends on an earlier store, unlike the one-thread version this replaces.
__device__ void rbBuildWarp(const TreeDev& t, int rs, int re, int rootNode) {
  const unsigned lane = threadIdx.x & 31u, full = 0xffffffffu;
  const int count = re - rs;
  const bool in = (int)lane < count;
  int idx = in ? t.leafIdx[rs + lane] : -1;
  float2 c = in ? t.leafCtr[rs + lane] : make_float2(0.0f, 0.0f);
  int base = 0;
  if (lane == 0 && count > 2) base = atomicAdd(&t.rb[2], count - 2);
  base = __shfl_sync(full, base, 0);
  const int myFree = ((int)lane < count - 2) ? t.freeIds[base + lane] : -1;    // lane k holds the k-th reserved id
  int nextFree = 0;                                                            // warp-uniform
  int s = in ? 0 : (int)lane, e = in ? count : (int)lane;                      // current range of this lane (empty for idle lanes)
  int node = rootNode, depth = 0, leafDepth = 0;
  int recS = 0, recE = 0, recId = -1, recDepth = 0;                            // the internal node this lane finishes
  for (int guard = 0; guard < 32 && __ballot_sync(full, e - s > 1); guard++) {     // <= 31 levels for 32 elements
    const int n = e - s;
    // center bounds of the range (:804-809)
    float lox = c.x, loy = c.y, hix = c.x, hiy = c.y;
    for (int j = 0; j < 32; j++) {
      float cx = __shfl_sync(full, c.x, j), cy = __shfl_sync(full, c.y, j);
      if (j >= s && j < e) { lox = lox < cx ? lox : cx; loy = loy < cy ? loy : cy; hix = hix > cx ? hix : cx; hiy = hiy > cy ? hiy : cy; }
    }
    bool f = false;
    if (n > 2) { float dx = hix - lox, dy = hiy - loy; f = dx > dy ? (c.x < 0.5f * (lox + hix)) : (c.y < 0.5f * (loy + hiy)); }
    const unsigned R = (e >= 32 ? full : ((1u << e) - 1u)) & ~((1u << s) - 1u);
    const unsigned F = __ballot_sync(full, f) & R;
    const int m = __popc(F);
    int split = s + n / 2;
    int partner = lane;
    if (n > 2 && m > 0 && m < n) {
      split = s + m;
      const unsigned leftRegion = R & ((1u << split) - 1u);
      const unsigned ML = ~F & leftRegion, MR = F & ~leftRegion;       // misplaced on the left / on the right
      const unsigned me = 1u << lane;
      if (ML & me) { int k = __popc(ML & (me - 1u)); unsigned x = MR; for (int i = 0; i < k; i++) x &= ~(1u << (31 - __clz(x))); partner = 31 - __clz(x); }
      else if (MR & me) { int k = __popc(MR & ~(me | (me - 1u))); unsigned x = ML; for (int i = 0; i < k; i++) x &= x - 1u; partner = __ffs(x) - 1; }
    }
    idx = __shfl_sync(full, idx, partner); c.x = __shfl_sync(full, c.x, partner); c.y = __shfl_sync(full, c.y, partner);
    // children (:905-930): a single element becomes the child itself, a longer range gets one of the reserved ids
    const bool owner = n > 1 && (int)lane == s;
    const bool needL = owner && split - s > 1, needR = owner && e - split > 1;
    const unsigned BL = __ballot_sync(full, needL), BR = __ballot_sync(full, needR);
    const unsigned below = (1u << lane) - 1u;
    const int slotL = nextFree + __popc(BL & below) + __popc(BR & below), slotR = slotL + (needL ? 1 : 0);
    nextFree += __popc(BL) + __popc(BR);
    int newL = -1, newR = -1;
    for (int j = 0; j < 32; j++) {                 // fetch reserved ids from the lanes that hold them
      int wantL = __shfl_sync(full, needL ? slotL : -1, j), wantR = __shfl_sync(full, needR ? slotR : -1, j);
      int gotL = __shfl_sync(full, myFree, wantL < 0 ? 0 : wantL), gotR = __shfl_sync(full, myFree, wantR < 0 ? 0 : wantR);
      if ((int)lane == j) { newL = gotL; newR = gotR; }
    }
    const int idxAtSplit = __shfl_sync(full, idx, split < 32 ? split : 0);
    int leftId = needL ? newL : idx, rightId = needR ? newR : idxAtSplit;
    if (owner) { t.child[node] = make_int2(leftId, rightId); t.parent[leftId] = node; t.parent[rightId] = node; }
    leftId = __shfl_sync(full, leftId, s < 32 ? s : 0); rightId = __shfl_sync(full, rightId, s < 32 ? s : 0);
    if (n > 1) {
      if ((int)lane == split) { recS = s; recE = e; recId = node; recDepth = depth; }
      if ((int)lane < split) { e = split; node = leftId; } else { s = split; node = rightId; }
      depth++;
      if (e - s == 1) leafDepth = depth;
    }
  }
  // finish the new nodes from the final arrangement
  float4 box = make_float4(0, 0, 0, 0); int h = 0; unsigned catLo = 0, catHi = 0;
  if (in) { box = t.aabb[idx]; h = (int)(t.hf[idx] & 0xffffu); unsigned long long ct = t.cat[idx]; catLo = (unsigned)ct; catHi = (unsigned)(ct >> 32); }
  Box acc; acc.lo = mk(0, 0); acc.hi = mk(0, 0); int accH = 0; unsigned accLo = 0, accHi = 0; bool first = true;
  for (int j = 0; j < 32; j++) {
    float4 bj = make_float4(__shfl_sync(full, box.x, j), __shfl_sync(full, box.y, j), __shfl_sync(full, box.z, j), __shfl_sync(full, box.w, j));
    int hj = __shfl_sync(full, h, j) + __shfl_sync(full, leafDepth, j);
    unsigned cl = __shfl_sync(full, catLo, j), ch = __shfl_sync(full, catHi, j);
    if (recId >= 0 && j >= recS && j < recE) {
      Box b = toBox(bj);
      acc = first ? b : boxUnion(acc, b); first = false;
      int hh = hj - recDepth; accH = accH > hh ? accH : hh;
      accLo |= cl; accHi |= ch;
    }
  }
  if (recId >= 0) {
    t.aabb[recId] = fromBox(acc); t.hf[recId] = ((uint32_t)NODE_ALLOCATED << 16) | (uint32_t)accH;
    t.cat[recId] = ((unsigned long long)accHi << 32) | accLo; t.user[recId] = -1;
    if (recId != rootNode) t.visit[recId] = 0;
  }
  if (in) { t.leafIdx[rs + lane] = idx; t.leafCtr[rs + lane] = c; }
  __threadfence();
  __syncwarp();
}

// is `n` a frontier node of the rebuild (a "leaf" of BuildTree)?  :973-978
__device__ __forceinline__ int rbClassify(const TreeDev& t, int n) {     // 0 = not involved, 1 = frontier, 2 = doomed internal node
  uint32_t hf = t.hf[n];
  uint32_t flags = hf >> 16;
  if (!(flags & NODE_ALLOCATED)) return 0;
  bool leaf = (hf & 0xffffu) == 0;
  if (!leaf && (flags & NODE_ENLARGED)) return 2;
  int p = t.parent[n];
  if (p == -1) return 0;                           // an un-enlarged root: nothing to rebuild (handled by the caller)
  return ((t.hf[p] >> 16) & NODE_ENLARGED) ? 1 : 0;
}


// All threads of the cooperative grid call this; every branch around a grid.sync() is uniform.
template <class S> __device__ void phase_rebuild(const S& grid, const WorldDev& d, int treeType) {
  const TreeDev& t = d.trees[treeType];
  const int tid = grid.tid(), stride = grid.size();
  const int root = t.meta[0];
  if (t.meta[3] == 0 || root == -1) return;
  { uint32_t hf = t.hf[root]; if (!((hf >> 16) & NODE_ENLARGED) || (hf & 0xffffu) == 0) { if (tid == 0) d.sc->rebuildLeafCount[treeType] = 0; return; } }
  const int cap = t.capacity;
  // R0
  for (int n = tid; n < cap; n += stride) { t.visit[n] = 0; t.subCount[n] = 0; }
  if (tid < 8) t.rb[tid] = 0;
  if (tid == 0) t.recBox[0] = make_float4(FLT_MAX, FLT_MAX, -FLT_MAX, -FLT_MAX);
  grid.sync();
  // R1: doomed internal nodes -> freeIds; frontier subtree counts bottom-up
  for (int n = tid; n < cap; n += stride) {
    int cls = rbClassify(t, n);
    if (cls == 2) t.freeIds[atomicAdd(&t.rb[1], 1)] = n;
    else if (cls == 1) {
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
  // R2: DFS (child1-first) rank of every frontier node = leaves to its left
  const int L = t.subCount[root];
  for (int n = tid; n < cap; n += stride) {
    if (rbClassify(t, n) != 1) continue;
    int r = 0, cur = n;
    while (true) { int p = t.parent[cur]; if (p == -1) break; int2 c = t.child[p]; if (c.y == cur) r += t.subCount[c.x]; cur = p; }
    t.leafIdx[r] = n;
    float4 a = t.aabb[n];
    float2 ctr = make_float2(0.5f * (a.x + a.z), 0.5f * (a.y + a.w));
    t.leafCtr[r] = ctr;
    t.segStart[r] = 0;
    float* b = (float*)&t.recBox[0];               // center bounds of the root range (first level's pivot)
    rbAtomicMinF(b + 0, ctr.x); rbAtomicMinF(b + 1, ctr.y); rbAtomicMaxF(b + 2, ctr.x); rbAtomicMaxF(b + 3, ctr.y);
  }
  grid.sync();
  if (t.rb[1] != L - 1) { if (tid == 0) atomicOr(&d.sc->errorFlags, (int)ERR_TREE); return; }
  // R3: root range; the doomed nodes become the pool of new internal nodes
  int* const recEndP[2] = {t.recEnd, t.recEnd2}; int* const recNodeP[2] = {t.recNode, t.recNode2}; float4* const recBoxP[2] = {t.recBox, t.recBox2};
  if (tid == 0) {
    d.sc->rebuildLeafCount[treeType] = L;
    int node0 = rbAllocNode(t);
    t.rb[6] = node0;
    t.recEnd[0] = L; t.recNode[0] = node0;
    if (L > kSerialRange) t.rb[3] = 1; else { t.serialList[0] = 0; t.serEnd[0] = L; t.serNode[0] = node0; t.rb[5] = 1; }
  }
  grid.sync();
  // Level loop, 4 grid barriers per level. Range records (end, node, center bounds) are double-buffered by level parity,
  // keyed by the range's first position; segStart[p] is the range an element is in, or -1 once its range is final
  // (a single element, or <= kSerialRange elements that a warp finishes later).
  int level = 0;
  unsigned long long tLoop = 0;
  if (d.debugChecksums && tid == 0) asm volatile("mov.u64 %0, %globaltimer;" : "=l"(tLoop));
  __shared__ int rbWarpSums[32];
  while (t.rb[3 + (level & 1)] > 0) {
    const int par = level & 1, nxt = par ^ 1, nextSlot = 3 + nxt;
    const int* recEnd = recEndP[par]; const int* recNode = recNodeP[par]; const float4* recBox = recBoxP[par];
    // P1: "< pivot" flag of every element (:810-845), fused into the first pass of the prefix sum
    {
      const int chunk = blockDim.x, chunks = (L + chunk - 1) / chunk;
      for (int c = grid.block(); c < chunks; c += grid.blocks()) {
        int i = c * chunk + threadIdx.x, f = 0;
        if (i < L) {
          int sg = t.segStart[i];
          if (sg >= 0) {
            float4 b = recBox[sg]; float2 ctr = t.leafCtr[i];
            float dx = b.z - b.x, dy = b.w - b.y;
            f = dx > dy ? (ctr.x < 0.5f * (b.x + b.z)) : (ctr.y < 0.5f * (b.y + b.w));
          }
          t.flagArr[i] = f;
        }
        int total, inc = blockInclusive(f, rbWarpSums, &total);
        if (i < L) t.flagScan[i] = inc - f;
        if (threadIdx.x == 0) d.scanBlockSums[c] = total;
      }
      if (tid == 0) t.rb[nextSlot] = 0;
      grid.sync();
      if (grid.block() == 0) {
        int carry = 0;
        for (int base = 0; base < chunks; base += chunk) {
          int i = base + threadIdx.x, v = i < chunks ? d.scanBlockSums[i] : 0;
          int total, inc = blockInclusive(v, rbWarpSums, &total);
          if (i < chunks) d.scanBlockSums[i] = carry + inc - v;
          carry += total;
        }
        if (threadIdx.x == 0) t.flagScan[cap] = carry;
      }
      grid.sync();
    }
    const int total = t.flagScan[cap];
    // P2: split of every range; misplaced elements of the right part record their position; the range's owner creates
    //     the two child ranges (:905-930) in the other parity's records
    for (int p = tid; p < L; p += stride) {
      const int sg = t.segStart[p];
      if (sg < 0) continue;
      const int e = recEnd[sg], node = recNode[sg];
      const int Fs = t.flagScan[sg] + d.scanBlockSums[sg / (int)blockDim.x];
      const int Fe = e < L ? t.flagScan[e] + d.scanBlockSums[e / (int)blockDim.x] : total;
      const int Fp = t.flagScan[p] + d.scanBlockSums[p / (int)blockDim.x];
      const int m = Fe - Fs, left = Fp - Fs;
      if (p >= sg + m && t.flagArr[p]) t.pairR[sg + (m - left - 1)] = p;
      if (p != sg) continue;
      const int split = (m > 0 && m < e - sg) ? sg + m : sg + (e - sg) / 2;
      t.recSplit[sg] = split; t.recM[sg] = m;
      for (int side = 0; side < 2; side++) {
        const int cs = side == 0 ? sg : split, ce = side == 0 ? split : e;
        if (ce - cs == 1) continue;                  // a single element: linked in P3, once the swaps have settled who it is
        const int childNode = rbAllocNode(t);
        t.parent[childNode] = node; ((int*)&t.child[node])[side] = childNode;
        recEndP[nxt][cs] = ce; recNodeP[nxt][cs] = childNode; recBoxP[nxt][cs] = make_float4(FLT_MAX, FLT_MAX, -FLT_MAX, -FLT_MAX);
        if (ce - cs > kSerialRange) atomicAdd(&t.rb[nextSlot], 1);
        else { int k = atomicAdd(&t.rb[5], 1); t.serialList[k] = cs; t.serEnd[k] = ce; t.serNode[k] = childNode; }
      }
    }
    grid.sync();
    // P3: k-th misplaced element from the left swaps with the k-th from the right; every element then moves to its child
    //     range: single elements are linked (and completed upward), long ranges collect their center bounds for the next level
    for (int p = tid; p < L; p += stride) {
      const int sg = t.segStart[p];
      if (sg < 0) continue;
      const int e = recEnd[sg], node = recNode[sg], split = t.recSplit[sg], m = t.recM[sg];
      const int left = (t.flagScan[p] + d.scanBlockSums[p / (int)blockDim.x]) - (t.flagScan[sg] + d.scanBlockSums[sg / (int)blockDim.x]);
      const bool f = t.flagArr[p] != 0;
      if (p >= sg + m && f) continue;               // misplaced on the right: its partner on the left handles both positions
      int pos[2] = {p, -1}; int idx[2] = {t.leafIdx[p], -1}; float2 ctr[2] = {t.leafCtr[p], make_float2(0.0f, 0.0f)};
      if (p < sg + m && !f) {
        const int q = t.pairR[sg + ((p - sg) - left)];
        pos[1] = q; idx[1] = idx[0]; ctr[1] = ctr[0];
        idx[0] = t.leafIdx[q]; ctr[0] = t.leafCtr[q];
        t.leafIdx[p] = idx[0]; t.leafCtr[p] = ctr[0]; t.leafIdx[q] = idx[1]; t.leafCtr[q] = ctr[1];
      }
      for (int k = 0; k < 2; k++) {
        if (pos[k] < 0) continue;
        const int side = pos[k] < split ? 0 : 1;
        const int cs = side == 0 ? sg : split, ce = side == 0 ? split : e;
        if (ce - cs == 1) {
          t.parent[idx[k]] = node; ((int*)&t.child[node])[side] = idx[k];
          t.segStart[pos[k]] = -1;
          __threadfence();
          rbWalkUp(t, idx[k]);
        } else if (ce - cs > kSerialRange) {
          t.segStart[pos[k]] = cs;
          float* b = (float*)&recBoxP[nxt][cs];
          rbAtomicMinF(b + 0, ctr[k].x); rbAtomicMinF(b + 1, ctr[k].y); rbAtomicMaxF(b + 2, ctr[k].x); rbAtomicMaxF(b + 3, ctr[k].y);
        } else t.segStart[pos[k]] = -1;
      }
    }
    grid.sync();
    level++;
  }
  if (d.debugChecksums && tid == 0 && treeType == kDynamicBody) {
    unsigned long long now; asm volatile("mov.u64 %0, %globaltimer;" : "=l"(now));
    d.sc->dbg[12] = level; d.sc->dbg[13] = L; d.sc->dbg[14] = now - tLoop; d.sc->dbg[15] = t.rb[5];
  }
  // small ranges: sequential BuildTree, one thread per range, then completion upward
  const int nSerial = t.rb[5];
  for (int k = tid >> 5; k < nSerial; k += stride >> 5) {       // one warp per range
    int node = t.serNode[k];
    rbBuildWarp(t, t.serialList[k], t.serEnd[k], node);
    if ((threadIdx.x & 31) == 0) rbWalkUp(t, node);
  }
  if (tid == 0) t.meta[0] = t.rb[6];
}

}  // namespace b2c
