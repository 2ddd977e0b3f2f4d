// libb2cuda — DynamicTree.Rebuild(false) (src/DynamicTree.cs:802-1000) as a grid-wide parallel algorithm that yields the SAME
// topology as the sequential C#: (1) the frontier (leaves and untouched subtrees under enlarged ancestors) is gathered in the
// reference's DFS order (child1 first) with subtree counts + per-node rank walks instead of a stack; (2) BuildTree's top-down
// median split runs level by level over all ranges at once — PartitionMid's two-pointer swap loop has the closed form
// "k-th misplaced element from the left swaps with the k-th misplaced element from the right", evaluated with one prefix sum
// per level; ranges of <= kSerialRange leaves are finished by one thread each with the sequential code; (3) boxes, heights
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
  while (true) {
    int p = t.parent[cur];
    if (p == -1) return;
    __threadfence();
    if (atomicAdd(&t.visit[p], 1) == 0) return;
    __threadfence();
    refitFromChildren(t, p);
    cur = p;
  }
}

// sequential BuildTree (:853-959) for a small range whose root node is already allocated and linked to its parent
__device__ void rbBuildSerial(const TreeDev& t, int s, int e, int rootNode) {
  int items[5 * (kSerialRange + 2)];
  int li[kSerialRange]; float2 lc[kSerialRange];      // the range is worked on in thread-local (L1-resident) copies
  const int base = s;
  for (int k = 0; k < e - s; k++) { li[k] = t.leafIdx[s + k]; lc[k] = t.leafCtr[s + k]; }
  e -= s; s = 0;
  int top = 0;
  items[0] = rootNode; items[1] = -1; items[2] = s; items[3] = s + partitionMid(li + s, lc + s, e - s); items[4] = e;
  while (true) {
    int* item = items + 5 * top;
    item[1]++;
    if (item[1] == 2) {
      if (top == 0) break;
      int doneNode = item[0];
      top--;
      int* parentItem = items + 5 * top;
      if (parentItem[1] == 0) t.child[parentItem[0]].x = doneNode; else t.child[parentItem[0]].y = doneNode;
      t.parent[doneNode] = parentItem[0];
      refitFromChildren(t, doneNode);
    } else {
      int startIndex, endIndex;
      if (item[1] == 0) { startIndex = item[2]; endIndex = item[3]; } else { startIndex = item[3]; endIndex = item[4]; }
      int count = endIndex - startIndex;
      if (count == 1) {
        int childIndex = li[startIndex];
        if (item[1] == 0) t.child[item[0]].x = childIndex; else t.child[item[0]].y = childIndex;
        t.parent[childIndex] = item[0];
      } else {
        int n = rbAllocNode(t);
        int split = startIndex + partitionMid(li + startIndex, lc + startIndex, count);
        top++;
        int* ni = items + 5 * top;
        ni[0] = n; ni[1] = -1; ni[2] = startIndex; ni[3] = split; ni[4] = endIndex;
      }
    }
  }
  refitFromChildren(t, rootNode);
  for (int k = 0; k < e; k++) { t.leafIdx[base + k] = li[k]; t.leafCtr[base + k] = lc[k]; }
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
__device__ void phase_rebuild(cg::grid_group& grid, const WorldDev& d, int treeType) {
  const TreeDev& t = d.trees[treeType];
  const int tid = gtid(), stride = gsize();
  const int root = t.meta[0];
  if (t.meta[3] == 0 || root == -1) return;
  { uint32_t hf = t.hf[root]; if (!((hf >> 16) & NODE_ENLARGED) || (hf & 0xffffu) == 0) { if (tid == 0) d.sc->rebuildLeafCount[treeType] = 0; return; } }
  const int cap = t.capacity;
  // R0
  for (int n = tid; n < cap; n += stride) { t.visit[n] = 0; t.subCount[n] = 0; }
  if (tid < 8) t.rb[tid] = 0;
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
    t.leafCtr[r] = make_float2(0.5f * (a.x + a.z), 0.5f * (a.y + a.w));
    t.segStart[r] = 0;
  }
  grid.sync();
  if (t.rb[1] != L - 1) { if (tid == 0) atomicOr(&d.sc->errorFlags, (int)ERR_TREE); return; }
  // R3: root range; the doomed nodes become the pool of new internal nodes
  if (tid == 0) {
    d.sc->rebuildLeafCount[treeType] = L;
    int node0 = rbAllocNode(t);
    t.rb[6] = node0;
    t.recEnd[0] = L; t.recNode[0] = node0; t.recBox[0] = make_float4(FLT_MAX, FLT_MAX, -FLT_MAX, -FLT_MAX);
    if (L > kSerialRange) t.rb[3] = 1; else { t.serialList[0] = 0; t.rb[5] = 1; }
  }
  grid.sync();
  int level = 0;
  while (t.rb[3 + (level & 1)] > 0) {
    const int nextSlot = 3 + ((level + 1) & 1);
    // A: elements move to their (possibly new) range; bounds of the centers per active range  :804-809
    for (int p = tid; p < L; p += stride) {
      int s = t.segStart[p], e = t.recEnd[s];
      if (p >= e) { s = e; t.segStart[p] = s; e = t.recEnd[s]; }
      if (e - s > kSerialRange && t.recNode[s] >= 0) {
        float2 c = t.leafCtr[p]; float* b = (float*)&t.recBox[s];
        rbAtomicMinF(b + 0, c.x); rbAtomicMinF(b + 1, c.y); rbAtomicMaxF(b + 2, c.x); rbAtomicMaxF(b + 3, c.y);
      }
    }
    if (tid == 0) t.rb[nextSlot] = 0;
    grid.sync();
    // B1: "< pivot" flags  :810-845
    for (int p = tid; p < L; p += stride) {
      int s = t.segStart[p], e = t.recEnd[s], f = 0;
      if (e - s > kSerialRange && t.recNode[s] >= 0) {
        float4 b = t.recBox[s]; float2 c = t.leafCtr[p];
        float dx = b.z - b.x, dy = b.w - b.y;
        f = dx > dy ? (c.x < 0.5f * (b.x + b.z)) : (c.y < 0.5f * (b.y + b.w));
      }
      t.flagScan[p] = f; t.flagArr[p] = f;
    }
    grid.sync();
    gridScan<int>(grid, t.flagScan, d.scanBlockSums, L, &t.flagScan[cap]);
    // B2: k-th misplaced element from the right records its position
    const int total = t.flagScan[cap];
    for (int p = tid; p < L; p += stride) {
      int s = t.segStart[p], e = t.recEnd[s];
      if (!(e - s > kSerialRange && t.recNode[s] >= 0)) continue;
      int Fs = t.flagScan[s] + d.scanBlockSums[s / kCoopThreads];
      int Fe = e < L ? t.flagScan[e] + d.scanBlockSums[e / kCoopThreads] : total;
      int Fp = t.flagScan[p] + d.scanBlockSums[p / kCoopThreads];
      int m = Fe - Fs, left = Fp - Fs;
      if (p == s) t.recSplit[s] = (m > 0 && m < e - s) ? s + m : s + (e - s) / 2;
      if (p >= s + m && t.flagArr[p]) t.pairR[s + (m - left - 1)] = p;
    }
    grid.sync();
    // B3: k-th misplaced element from the left swaps with it
    for (int p = tid; p < L; p += stride) {
      int s = t.segStart[p], e = t.recEnd[s];
      if (!(e - s > kSerialRange && t.recNode[s] >= 0)) continue;
      int Fs = t.flagScan[s] + d.scanBlockSums[s / kCoopThreads];
      int Fe = e < L ? t.flagScan[e] + d.scanBlockSums[e / kCoopThreads] : total;
      int Fp = t.flagScan[p] + d.scanBlockSums[p / kCoopThreads];
      int m = Fe - Fs, left = Fp - Fs;
      if (p < s + m && !t.flagArr[p]) {
        int q = t.pairR[s + ((p - s) - left)];
        int ti = t.leafIdx[p]; t.leafIdx[p] = t.leafIdx[q]; t.leafIdx[q] = ti;
        float2 tc = t.leafCtr[p]; t.leafCtr[p] = t.leafCtr[q]; t.leafCtr[q] = tc;
      }
    }
    grid.sync();
    // B4: the owner of each active range creates the two child ranges  :905-930
    for (int p = tid; p < L; p += stride) {
      int s = t.segStart[p];
      if (p != s) continue;
      int e = t.recEnd[s], node = t.recNode[s];
      if (!(e - s > kSerialRange && node >= 0)) continue;
      int split = t.recSplit[s];
      int2 ch;
      for (int side = 0; side < 2; side++) {
        int cs = side == 0 ? s : split, ce = side == 0 ? split : e;
        int childNode;
        if (ce - cs == 1) {
          childNode = t.leafIdx[cs];
          t.parent[childNode] = node;
          t.recEnd[cs] = ce; t.recNode[cs] = -1;
        } else {
          childNode = rbAllocNode(t);
          t.parent[childNode] = node;
          t.recEnd[cs] = ce; t.recNode[cs] = childNode; t.recBox[cs] = make_float4(FLT_MAX, FLT_MAX, -FLT_MAX, -FLT_MAX);
          if (ce - cs > kSerialRange) atomicAdd(&t.rb[nextSlot], 1);
          else t.serialList[atomicAdd(&t.rb[5], 1)] = cs;
        }
        if (side == 0) ch.x = childNode; else ch.y = childNode;
      }
      t.child[node] = ch;
      __threadfence();
      if (split - s == 1) rbWalkUp(t, ch.x);
      if (e - split == 1) rbWalkUp(t, ch.y);
    }
    grid.sync();
    level++;
  }
  // small ranges: sequential BuildTree, one thread per range, then completion upward
  const int nSerial = t.rb[5];
  for (int k = tid; k < nSerial; k += stride) {
    int s = t.serialList[k];
    int node = t.recNode[s];
    rbBuildSerial(t, s, t.recEnd[s], node);
    rbWalkUp(t, node);
  }
  if (tid == 0) t.meta[0] = t.rb[6];
}

}  // namespace b2c
