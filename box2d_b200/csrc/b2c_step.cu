// libb2cuda — the rigid-body step on the device (WorldAPI.Step, api/World.cs:31-113).
// Every kernel is grid-strided over a count that lives in device memory (Scalars), so the whole step is a fixed
// sequence of launches with no host read-back: UpdateBroadPhasePairs -> Collide -> Solve -> Finalize/refit.
// Serial, order-defining parts of the reference (contact id allocation, begin/end processing in ascending contactId,
// greedy graph coloring, move-array order) are reproduced with ordered compaction (prefix sums) and priority rounds,
// so ids, colors and event streams are bit-identical to the single-threaded C# step.
#include "b2c_host.cuh"
#include "b2c_manifold.cuh"
#include <cooperative_groups.h>
#include <cstdio>
#include <algorithm>

namespace cg = cooperative_groups;

namespace b2c {

#define CK(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) return fail(B2C_E_CUDA, std::string(#call) + ": " + cudaGetErrorString(e_)); } while (0)

constexpr int kThreads = 256;
constexpr int kCoopThreads = 512;      // block size of the cooperative phase kernels == chunk of gridScan

__device__ __forceinline__ int gtid() { return blockIdx.x * blockDim.x + threadIdx.x; }
__device__ __forceinline__ int gsize() { return gridDim.x * blockDim.x; }
__device__ __forceinline__ Box toBox(float4 a) { Box b; b.lo = mk(a.x, a.y); b.hi = mk(a.z, a.w); return b; }
__device__ __forceinline__ float4 fromBox(Box b) { return make_float4(b.lo.x, b.lo.y, b.hi.x, b.hi.y); }
__device__ __forceinline__ Xf toXf(float4 a) { return mkXf(mk(a.x, a.y), mkRot(a.z, a.w)); }
__device__ __forceinline__ unsigned long long pairKey(int a, int b) {   // src/Contact.cs:126
  unsigned long long k1 = (unsigned long long)(unsigned)a, k2 = (unsigned long long)(unsigned)b;
  return k1 < k2 ? (k1 << 32 | k2) : (k2 << 32 | k1);
}
__device__ __forceinline__ unsigned hashKey(unsigned long long k) {
  k ^= k >> 33; k *= 0xff51afd7ed558ccdULL; k ^= k >> 33; k *= 0xc4ceb9fe1a85ec53ULL; k ^= k >> 33; return (unsigned)k;
}
constexpr unsigned long long kEmptyKey = ~0ull;

// Barrier scopes. The phases below are written once and run at three scales: the whole cooperative grid (one big world: grid.sync),
// one thread-block cluster (the grid-path solver: barrier.cluster) and ONE CTA (a small world of a batch: __syncthreads, see k_batch_step).
// tid()/size() enumerate the threads of the scope, block()/blocks() its CTAs (prefix-sum chunks are handed out per CTA).
struct GridScope {
  cg::grid_group& g;
  __device__ __forceinline__ void sync() const { g.sync(); }
  __device__ __forceinline__ int tid() const { return blockIdx.x * blockDim.x + threadIdx.x; }
  __device__ __forceinline__ int size() const { return gridDim.x * blockDim.x; }
  __device__ __forceinline__ int block() const { return blockIdx.x; }
  __device__ __forceinline__ int blocks() const { return gridDim.x; }
};
struct ClusterScope {       // launched as a one-cluster grid: grid ranks are cluster ranks
  __device__ __forceinline__ void sync() const { cg::this_cluster().sync(); }
  __device__ __forceinline__ int tid() const { return blockIdx.x * blockDim.x + threadIdx.x; }
  __device__ __forceinline__ int size() const { return gridDim.x * blockDim.x; }
  __device__ __forceinline__ int block() const { return blockIdx.x; }
  __device__ __forceinline__ int blocks() const { return gridDim.x; }
};
struct CtaScope {
  __device__ __forceinline__ void sync() const { __syncthreads(); }
  __device__ __forceinline__ int tid() const { return threadIdx.x; }
  __device__ __forceinline__ int size() const { return blockDim.x; }
  __device__ __forceinline__ int block() const { return 0; }
  __device__ __forceinline__ int blocks() const { return 1; }
};

// ------------------------------------------------------------------------------------------------------------
// block / grid exclusive scans used by the ordered compactions
// ------------------------------------------------------------------------------------------------------------
template <class T> __device__ __forceinline__ T zeroOf();
template <> __device__ __forceinline__ int zeroOf<int>() { return 0; }
template <> __device__ __forceinline__ int4 zeroOf<int4>() { return make_int4(0, 0, 0, 0); }
__device__ __forceinline__ int addT(int a, int b) { return a + b; }
__device__ __forceinline__ int4 addT(int4 a, int4 b) { return make_int4(a.x + b.x, a.y + b.y, a.z + b.z, a.w + b.w); }
__device__ __forceinline__ int make_neg(int v) { return -v; }
__device__ __forceinline__ int4 make_neg(int4 v) { return make_int4(-v.x, -v.y, -v.z, -v.w); }
__device__ __forceinline__ int shflUp(int v, int d) { return __shfl_up_sync(0xffffffffu, v, d); }
__device__ __forceinline__ int4 shflUp(int4 v, int d) {
  return make_int4(__shfl_up_sync(0xffffffffu, v.x, d), __shfl_up_sync(0xffffffffu, v.y, d), __shfl_up_sync(0xffffffffu, v.z, d), __shfl_up_sync(0xffffffffu, v.w, d));
}
// block-wide inclusive scan for 1024 threads; returns inclusive value, total via shared
template <class T> __device__ __forceinline__ T blockInclusive(T v, T* warpSums, T* total) {
  int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  for (int d = 1; d < 32; d <<= 1) { T n = shflUp(v, d); if (lane >= d) v = addT(v, n); }
  if (lane == 31) warpSums[warp] = v;
  __syncthreads();
  if (warp == 0) {
    T w = lane < (int)(blockDim.x >> 5) ? warpSums[lane] : zeroOf<T>();
    for (int d = 1; d < 32; d <<= 1) { T n = shflUp(w, d); if (lane >= d) w = addT(w, n); }
    warpSums[lane] = w;
  }
  __syncthreads();
  if (warp > 0) v = addT(v, warpSums[warp - 1]);
  *total = warpSums[(blockDim.x >> 5) - 1];
  __syncthreads();
  return v;
}
// grid-wide exclusive scan inside a cooperative kernel (chunk = blockDim.x); consumers read data[i] + sums[i / blockDim.x]
template <class T, class S> __device__ void gridScan(const S& grid, T* data, T* sums, int n, T* totalOut) {
  __shared__ T warpSums[32];
  const int chunk = blockDim.x;
  int chunks = (n + chunk - 1) / chunk;
  for (int c = grid.block(); c < chunks; c += grid.blocks()) {
    int i = c * chunk + threadIdx.x;
    T v = i < n ? data[i] : zeroOf<T>();
    T total;
    T inc = blockInclusive(v, warpSums, &total);
    if (i < n) data[i] = addT(inc, make_neg(v));
    if (threadIdx.x == 0) sums[c] = total;
  }
  grid.sync();
  if (grid.block() == 0) {
    T carry = zeroOf<T>();
    for (int base = 0; base < chunks; base += chunk) {
      int i = base + threadIdx.x;
      T v = i < chunks ? sums[i] : zeroOf<T>();
      T total;
      T inc = blockInclusive(v, warpSums, &total);
      if (i < chunks) sums[i] = addT(carry, addT(inc, make_neg(v)));
      carry = addT(carry, total);
    }
    if (threadIdx.x == 0) *totalOut = carry;
  }
  grid.sync();
}

// ------------------------------------------------------------------------------------------------------------
// Phase 0: UpdateBroadPhasePairs (src/BroadPhase.cs:137-268)
// ------------------------------------------------------------------------------------------------------------
__device__ void p_pair_insert(const WorldDev& d, int tid, int stride) {
  if (d.sc->moveCount == 0) return;
  int n = d.sc->contactNext;
  unsigned mask = (unsigned)d.pairTableSize - 1u;
  for (int i = tid; i < n; i += stride) {
    if (!(d.cFlags[i] & CF_ALIVE)) continue;
    int2 s = d.cShape[i];
    unsigned long long key = pairKey(s.x, s.y);
    unsigned h = hashKey(key) & mask;
    for (int probe = 0; probe < d.pairTableSize; probe++) {
      unsigned long long old = atomicCAS(&d.pairTable[h], kEmptyKey, key);
      if (old == kEmptyKey || old == key) break;
      h = (h + 1) & mask;
    }
  }
}
__device__ __forceinline__ bool pairSetContains(const WorldDev& d, unsigned long long key) {
  unsigned mask = (unsigned)d.pairTableSize - 1u;
  unsigned h = hashKey(key) & mask;
  for (int probe = 0; probe < d.pairTableSize; probe++) {
    unsigned long long v = d.pairTable[h];
    if (v == key) return true;
    if (v == kEmptyKey) return false;
    h = (h + 1) & mask;
  }
  return false;
}
__device__ __forceinline__ bool shouldShapesCollide(const ShapeDev& a, const ShapeDev& b) {   // src/Shape.cs:56-61
  if (a.groupIndex == b.groupIndex && a.groupIndex != 0) return a.groupIndex > 0;
  return (a.maskBits & b.categoryBits) != 0 && (a.categoryBits & b.maskBits) != 0;
}
__device__ inline bool shouldBodiesCollide(const WorldDev& d, int bodyIdA, int bodyIdB) {     // src/Body.cs:354-371
  if (d.bodyType[bodyIdA] != kDynamicBody && d.bodyType[bodyIdB] != kDynamicBody) return false;
  int jointKey, otherBodyId;
  if (d.bodyJointCount[bodyIdA] < d.bodyJointCount[bodyIdB]) { jointKey = d.bodyHeadJoint[bodyIdA]; otherBodyId = bodyIdB; }
  else { jointKey = d.bodyHeadJoint[bodyIdB]; otherBodyId = bodyIdA; }
  while (jointKey != -1) {
    int jointId = jointKey >> 1, edge = jointKey & 1;
    int4 e = d.jointEdges[jointId];
    int otherBody = edge == 1 ? e.x : e.z;      // the edge opposite to ours
    if (!d.jointCollide[jointId] && otherBody == otherBodyId) return false;
    jointKey = edge == 1 ? e.w : e.y;
  }
  return true;
}

constexpr int kQueryStack = 128;

// PairQueryCallback (:137-196) for one leaf hit. Returns true when the pair must become a contact.
__device__ __forceinline__ bool acceptPair(const WorldDev& d, int queryKey, int queryType, int queryShape, int hitProxyId, int treeType, int hitShape,
                                           int* shapeIdA, int* shapeIdB) {
  int proxyKey = (hitProxyId << 2) | treeType;
  if (proxyKey == queryKey) return false;
  if (queryType == kDynamicBody) {
    if (treeType == kDynamicBody && proxyKey < queryKey && d.trees[treeType].moved[hitProxyId]) return false;
  } else if (d.trees[treeType].moved[hitProxyId]) return false;
  if (pairSetContains(d, pairKey(hitShape, queryShape))) return false;
  int a, b;
  if (proxyKey < queryKey) { a = hitShape; b = queryShape; } else { a = queryShape; b = hitShape; }
  const ShapeDev& sa = d.shapes[a]; const ShapeDev& sb = d.shapes[b];
  if (sa.bodyId == sb.bodyId) return false;
  if (sa.sensorIndex != -1 || sb.sensorIndex != -1) return false;
  if (!shouldShapesCollide(sa, sb)) return false;
  if (!shouldBodiesCollide(d, sa.bodyId, sb.bodyId)) return false;
  if (!pairSupported(sa.geom.type, sb.geom.type)) return false;   // Contact.CreateContact returns early (src/Contact.cs:137-139)
  *shapeIdA = a; *shapeIdB = b;
  return true;
}

// DynamicTree.Query (src/DynamicTree.cs:531-557): explicit stack, child1 pushed first so child2 is visited first.
// BUDGET: give up (return -1) after kBigQueryVisits node visits — the query is then answered by the leaf sweep below.
__device__ __forceinline__ void prefetchL1(const void* p) { asm volatile("prefetch.global.L1 [%0];" :: "l"(p)); }

struct HitSink {               // where accepted hits of one query go
  int2* out; int outLast;      // FILL: movePairs and the query's last slot (hits are written in reverse)
  int2* buf;                   // count pass: first kHitBuf hits in hit order
};
template <bool FILL, bool BUDGET> __device__ inline int queryTree(const WorldDev& d, int treeType, Box fat, int queryKey, int queryType, int queryShape, int hitBase, const HitSink& sink, int& visits) {
  const TreeDev& t = d.trees[treeType];
  int root = t.meta[0];
  if (t.meta[1] == 0 || root == -1) return 0;
  int stack[kQueryStack];
  int sp = 0, hits = 0;
  stack[sp++] = root;
  while (sp > 0) {
    int id = stack[--sp];
    if (BUDGET && ++visits > d.bigQueryVisits) return -1;
    float4 nb = t.aabb[id]; int2 c = t.child[id];      // independent loads; leaves have child (-1,-1)
    if (!boxOverlaps(toBox(nb), fat)) continue;
    // The reference also tests (categoryBits & maskBits) with maskBits = DEFAULT_MASK_BITS (:541): that only rejects
    // category 0, and a category-0 leaf is rejected by shouldShapesCollide below, so the hit list is the same.
    if (c.x == -1) {
      int a, b;
      if (acceptPair(d, queryKey, queryType, queryShape, id, treeType, t.user[id], &a, &b)) {
        int h = hitBase + hits;
        if (FILL) sink.out[sink.outLast - h] = make_int2(a, b);
        else if (h < kHitBuf) sink.buf[h] = make_int2(a, b);
        hits++;
      }
    } else {
      if (sp + 2 > kQueryStack) { atomicOr(&d.sc->errorFlags, (int)ERR_STACK_OVERFLOW); continue; }
      stack[sp++] = c.x; stack[sp++] = c.y;
      // pointer chasing: ask for child1's node now (it is popped after child2's whole subtree), so that pop finds it in L1
      prefetchL1(&t.aabb[c.x]); prefetchL1(&t.child[c.x]);
    }
  }
  return hits;
}
template <bool FILL, bool BUDGET> __device__ inline int queryTrees(const WorldDev& d, Box fat, int queryKey, int type, int queryShape, const HitSink& sink) {
  int count = 0, visits = 0, h;
  if (type == kDynamicBody) {
    h = queryTree<FILL, BUDGET>(d, kKinematicBody, fat, queryKey, type, queryShape, count, sink, visits); if (h < 0) return -1; count += h;
    h = queryTree<FILL, BUDGET>(d, kStaticBody, fat, queryKey, type, queryShape, count, sink, visits); if (h < 0) return -1; count += h;
  }
  h = queryTree<FILL, BUDGET>(d, kDynamicBody, fat, queryKey, type, queryShape, count, sink, visits); if (h < 0) return -1; count += h;
  return count;
}

// FindPairsTask (:197-232). FILL=false counts accepted hits per move entry (keeping the first kHitBuf); FILL=true writes
// them in contact-creation order: move index ascending, within a query the REVERSE of the hit order (the C# result list
// is LIFO, :193-194,:251-262). Accepted hits are NEW pairs only, so most queries have none and few have more than kHitBuf.
template <bool FILL> __device__ void p_find_pairs(const WorldDev& d, int tid, int stride) {
  int n = d.sc->moveCount;
  for (int i = tid; i < n; i += stride) {
    int queryKey = d.moveArray[i];
    int count = 0;
    if (FILL && d.moveBig[i] >= 0) continue;    // wide query: p_big_fill copies its hits
    if (queryKey != -1) {
      int type = queryKey & 3, id = queryKey >> 2;
      HitSink sink; sink.out = d.movePairs; sink.outLast = 0; sink.buf = d.moveHitBuf + (size_t)i * kHitBuf;
      if (FILL) {
        int begin = d.moveHitCount[i] + d.scanBlockSums[i / (int)blockDim.x];
        int end = (i + 1 < n) ? d.moveHitCount[i + 1] + d.scanBlockSums[(i + 1) / (int)blockDim.x] : d.sc->newPairCount;
        if (end == begin) continue;
        if (end > d.pairCapacity) { atomicOr(&d.sc->errorFlags, (int)ERR_PAIR_CAPACITY); continue; }
        sink.outLast = end - 1;
        if (end - begin <= kHitBuf) { for (int h = 0; h < end - begin; h++) d.movePairs[end - 1 - h] = sink.buf[h]; continue; }
        Box fat = toBox(d.trees[type].aabb[id]);
        queryTrees<true, false>(d, fat, queryKey, type, d.trees[type].user[id], sink);
      } else {
        Box fat = toBox(d.trees[type].aabb[id]);
        int queryShape = d.trees[type].user[id];
        int big = -1;
        // A box that covers many average leaf cells of the dynamic tree is wide for certain (the tumbler's walls): do not
        // spend the visit budget on finding that out. The budget stays as the general detector.
        bool wide = false;
        { const TreeDev& dt = d.trees[kDynamicBody]; int root = dt.meta[0], leaves = dt.meta[3];
          if (root != -1 && leaves > 4 * kWideQueryCells) {
            float4 rb = dt.aabb[root];
            float ox = fminf(fat.hi.x, rb.z) - fmaxf(fat.lo.x, rb.x), oy = fminf(fat.hi.y, rb.w) - fmaxf(fat.lo.y, rb.y);
            float cell = (rb.z - rb.x) * (rb.w - rb.y) / (float)leaves;
            wide = ox > 0.0f && oy > 0.0f && ox * oy > (float)kWideQueryCells * cell;
          } }
        count = wide ? -1 : queryTrees<false, true>(d, fat, queryKey, type, queryShape, sink);
        if (count < 0) {
          int slot = atomicAdd(&d.sc->bigQueryCount, 1);
          if (slot < kMaxBigQueries) { d.bigList[slot] = i; big = slot; count = 0; }
          else count = queryTrees<false, false>(d, fat, queryKey, type, queryShape, sink);
        }
        d.moveBig[i] = big;
      }
    } else if (!FILL) d.moveBig[i] = -1;
    if (!FILL) d.moveHitCount[i] = count;
  }
}

// World.CreateContact (src/Contact.cs:134-227), one thread per new pair in creation order. Ids come from the LIFO pool:
// the r-th allocation pops freeArray[freeCount-1-r], then extends nextIndex (src/IDPool.cs:15-28).
__device__ void p_create_contacts(const WorldDev& d, int tid, int stride) {
  int n = d.sc->newPairCount;
  int freeCount = d.sc->contactFreeCount, next = d.sc->contactNext;
  for (int r = tid; r < n; r += stride) {
    int id = r < freeCount ? d.contactFree[freeCount - 1 - r] : next + (r - freeCount);
    if (id >= d.contactCapacity) { atomicOr(&d.sc->errorFlags, (int)ERR_CONTACT_CAPACITY); continue; }
    int2 p = d.movePairs[r];
    int a = p.x, b = p.y;
    if (!pairPrimary(d.shapes[a].geom.type, d.shapes[b].geom.type)) { int t = a; a = b; b = t; }
    const ShapeDev& sa = d.shapes[a]; const ShapeDev& sb = d.shapes[b];
    uint32_t flags = CF_ALIVE;
    if ((sa.flags | sb.flags) & SF_CONTACT_EVENTS) flags |= CF_ENABLE_CONTACT_EVENTS;
    d.cShape[id] = make_int2(a, b);
    d.cBody[id] = make_int2(sa.bodyId, sb.bodyId);
    d.cFlags[id] = flags;
    d.cColor[id] = -1;
    d.cGen[id] = d.cGen[id] + 1;
    d.cStamp[id] = d.sc->contactStamp + (unsigned)r;
    d.cConstraint[id] = -1; d.cLocal[id] = -1;
    float4 ma = sa.material, mb = sb.material;
    d.cMat[id] = make_float4(sqrtf(ma.x * mb.x), netMax(ma.y, mb.y), 0.0f, 0.0f);   // src/World.cs:139-140
    d.cMass[id] = make_float4(0, 0, 0, 0);
    d.cSim[id] = make_int2(-1, -1);
    float4 z = make_float4(0, 0, 0, 0);
    d.mN[id] = z; d.mA0[id] = z; d.mA1[id] = z; d.mB0[id] = z; d.mB1[id] = z; d.mC0[id] = z; d.mC1[id] = z;
  }
}
__device__ void p_clear_moved(const WorldDev& d, int tid, int stride) {
  int n = d.sc->moveCount;
  for (int i = tid; i < n; i += stride) {
    int key = d.moveArray[i];
    if (key != -1) d.trees[key & 3].moved[key >> 2] = 0;
  }
}
__device__ void p_pairs_finish(const WorldDev& d) {
  Scalars& s = *d.sc;
  int n = s.newPairCount;
  int fromFree = n < s.contactFreeCount ? n : s.contactFreeCount;
  s.contactFreeCount -= fromFree;
  s.contactNext += n - fromFree;
  if (s.contactNext > d.contactCapacity) { s.contactNext = d.contactCapacity; s.errorFlags |= (int)ERR_CONTACT_CAPACITY; }
  s.liveContactCount += n; s.contactStamp += (unsigned)n;
  s.moveCount = 0; s.bigQueryCount = 0;
  s.beginCount = 0; s.hitCount = 0; s.jointEventCount = 0; s.anyRestitution = 0; s.anyHitShapes = 0; s.sensorBeginCount = 0;
}

// ------------------------------------------------------------------------------------------------------------
// Phase 1a: BroadPhase.RebuildTrees (src/BroadPhase.cs:112-116) -> DynamicTree.Rebuild(false) (src/DynamicTree.cs:802-1000).
// The topology this produces decides the DFS order of later queries, hence contact ids; it is reproduced step for step.
// The parallel, bit-exact version lives in b2c_rebuild.cuh; PartitionMid and the child refit below are shared with it.
// ------------------------------------------------------------------------------------------------------------
__device__ int partitionMid(int* indices, float2* centers, int count) {   // :802-847
  if (count <= 2) return count / 2;
  float2 lo = centers[0], hi = centers[0];
  for (int i = 1; i < count; i++) {
    float2 c = centers[i];
    lo.x = lo.x < c.x ? lo.x : c.x; lo.y = lo.y < c.y ? lo.y : c.y;
    hi.x = hi.x > c.x ? hi.x : c.x; hi.y = hi.y > c.y ? hi.y : c.y;
  }
  float dx = hi.x - lo.x, dy = hi.y - lo.y;
  float cx = 0.5f * (lo.x + hi.x), cy = 0.5f * (lo.y + hi.y);
  int i1 = 0, i2 = count;
  if (dx > dy) {
    float pivot = cx;
    while (i1 < i2) {
      while (i1 < i2 && centers[i1].x < pivot) i1++;
      while (i1 < i2 && centers[i2 - 1].x >= pivot) i2--;
      if (i1 < i2) { int ti = indices[i1]; indices[i1] = indices[i2 - 1]; indices[i2 - 1] = ti; float2 tc = centers[i1]; centers[i1] = centers[i2 - 1]; centers[i2 - 1] = tc; i1++; i2--; }
    }
  } else {
    float pivot = cy;
    while (i1 < i2) {
      while (i1 < i2 && centers[i1].y < pivot) i1++;
      while (i1 < i2 && centers[i2 - 1].y >= pivot) i2--;
      if (i1 < i2) { int ti = indices[i1]; indices[i1] = indices[i2 - 1]; indices[i2 - 1] = ti; float2 tc = centers[i1]; centers[i1] = centers[i2 - 1]; centers[i2 - 1] = tc; i1++; i2--; }
    }
  }
  if (i1 > 0 && i1 < count) return i1;
  return count / 2;
}
__device__ __forceinline__ void refitFromChildren(const TreeDev& t, int node) {
  int2 c = t.child[node];
  Box u = boxUnion(toBox(t.aabb[c.x]), toBox(t.aabb[c.y]));
  t.aabb[node] = fromBox(u);
  int h1 = (int)(t.hf[c.x] & 0xffffu), h2 = (int)(t.hf[c.y] & 0xffffu);
  t.hf[node] = (t.hf[node] & 0xffff0000u) | (uint32_t)(1 + (h1 > h2 ? h1 : h2));
  t.cat[node] = t.cat[c.x] | t.cat[c.y];
}

// ------------------------------------------------------------------------------------------------------------
// Phase 1b: CollideTask + UpdateContact (src/World.cs:261-310, src/Contact.cs:324-417), one thread per contact slot
// ------------------------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t fbits(float f) { return __float_as_uint(f); }
__device__ __forceinline__ void loadPoint(MPoint& p, float4 a, float4 b, float4 c) {
  p.anchorA = mk(a.x, a.y); p.anchorB = mk(a.z, a.w); p.separation = b.x; p.normalImpulse = b.y; p.tangentImpulse = b.z; p.totalNormalImpulse = b.w;
  p.point = mk(c.x, c.y); p.normalVelocity = c.z; uint32_t w = fbits(c.w); p.id = w & 0xffffu; p.persisted = (w >> 16) & 1u;
}
__device__ __forceinline__ void storePoint(const MPoint& p, float4* a, float4* b, float4* c) {
  *a = make_float4(p.anchorA.x, p.anchorA.y, p.anchorB.x, p.anchorB.y);
  *b = make_float4(p.separation, p.normalImpulse, p.tangentImpulse, p.totalNormalImpulse);
  *c = make_float4(p.point.x, p.point.y, p.normalVelocity, __uint_as_float((p.id & 0xffffu) | (p.persisted << 16)));
}

__device__ __forceinline__ unsigned long long globalTimer() { unsigned long long t; asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t)); return t; }
__device__ __forceinline__ void prefetchL2(const void* p) { asm volatile("prefetch.global.L2 [%0];" :: "l"(p)); }
__device__ void p_collide(const WorldDev& d, const StepParams& sp, int tid, int stride) {
  int n = d.sc->contactNext;
  // A thread that owns several slots (batched worlds: ~16 per thread) walks a chain of dependent loads per slot
  // (slot -> shape ids -> fat boxes / shape rows / body rows). The ids of the NEXT slot are loaded one iteration early and
  // everything they point to is requested from L2 while the current manifold is computed.
  uint32_t nextFlags = 0; int2 nextSh = make_int2(0, 0), nextBd = make_int2(0, 0);
  if (tid < n) { nextFlags = d.cFlags[tid]; nextSh = d.cShape[tid]; nextBd = d.cBody[tid]; }
  for (int i = tid; i < n; i += stride) {
    uint32_t flags = nextFlags;
    const int2 curSh = nextSh, curBd = nextBd;
    const int j = i + stride;
    if (j < n) {
      nextFlags = d.cFlags[j]; nextSh = d.cShape[j]; nextBd = d.cBody[j];
      if (nextFlags & CF_ALIVE) {
        prefetchL2(&d.shapeFat[nextSh.x]); prefetchL2(&d.shapeFat[nextSh.y]);
        prefetchL2(&d.shapes[nextSh.x]); prefetchL2((const char*)&d.shapes[nextSh.x] + 128);
        prefetchL2(&d.shapes[nextSh.y]); prefetchL2((const char*)&d.shapes[nextSh.y] + 128);
        prefetchL2(&d.bodyXf[nextBd.x]); prefetchL2(&d.bodyXf[nextBd.y]); prefetchL2(&d.bodyMass[nextBd.x]); prefetchL2(&d.bodyMass[nextBd.y]);
        prefetchL2(&d.mN[j]); prefetchL2(&d.mA0[j]); prefetchL2(&d.mA1[j]); prefetchL2(&d.mB0[j]); prefetchL2(&d.mB1[j]); prefetchL2(&d.mC0[j]); prefetchL2(&d.mC1[j]);
      }
    }
    int4 fs = make_int4(0, 0, 0, 0);
    if (flags & CF_ALIVE) {
      int2 sh = curSh;
      bool overlap = boxOverlaps(toBox(d.shapeFat[sh.x]), toBox(d.shapeFat[sh.y]));
      bool wasTouching = (flags & CF_TOUCHING) != 0;
      bool events = (flags & CF_ENABLE_CONTACT_EVENTS) != 0;
      if (!overlap) {
        flags |= CF_DISJOINT;
        fs.z = 1;                                   // destroy
        if (wasTouching) { fs.w = 1; if (events) fs.y = 1; }     // graph removal + end event (DestroyContact :247-254)
      } else {
        const ShapeDev& sa = d.shapes[sh.x]; const ShapeDev& sb = d.shapes[sh.y];
        int2 bd = curBd;
        float4 massA = d.bodyMass[bd.x], massB = d.bodyMass[bd.y];
        d.cSim[i] = make_int2(d.bodySim[bd.x], d.bodySim[bd.y]);
        d.cMass[i] = make_float4(massA.x, massA.y, massB.x, massB.y);
        Xf xfA = toXf(d.bodyXf[bd.x]), xfB = toXf(d.bodyXf[bd.y]);
        V2 centerOffsetA = rotate(xfA.q, mk(massA.z, massA.w)), centerOffsetB = rotate(xfB.q, mk(massB.z, massB.w));
        // UpdateContact
        Manifold old; old.pointCount = (int)fbits(d.mN[i].w); old.rollingImpulse = d.mN[i].z;
        loadPoint(old.p0, d.mA0[i], d.mB0[i], d.mC0[i]); loadPoint(old.p1, d.mA1[i], d.mB1[i], d.mC1[i]);
        Manifold m = computeManifold(sa.geom, xfA, sb.geom, xfB);
        float4 ma = sa.material, mb = sb.material;
        float friction = sqrtf(ma.x * mb.x), restitution = netMax(ma.y, mb.y), rolling = 0.0f;
        if (ma.z > 0.0f || mb.z > 0.0f) {
          float rA = sa.geom.type == kSegment ? 0.0f : sa.geom.radius, rB = sb.geom.type == kSegment ? 0.0f : sb.geom.radius;
          rolling = netMax(ma.z, mb.z) * netMax(rA, rB);
        }
        d.cMat[i] = make_float4(friction, restitution, rolling, ma.w + mb.w);
        int pointCount = m.pointCount;
        bool touching = pointCount > 0;
        if (touching && !(restitution == 0.0f)) d.sc->anyRestitution = 1;    // some 8-lane group is not AllZeroW (src/ContactSolver.cs:923): p_rest_groups finds which
        if (!sp.enableSpeculative && pointCount == 2) {            // Q7: tests point0 twice (src/Contact.cs:366,371)
          if (m.p0.separation > 1.5f * kLinearSlop) { m.p0 = m.p1; m.pointCount = 1; }
          pointCount = m.pointCount;
        }
        // sampled by every UpdateContact (src/Contact.cs:375-377), not when the contact is created
        if (touching && ((sa.flags | sb.flags) & SF_HIT_EVENTS)) { flags |= CF_ENABLE_HIT; d.sc->anyHitShapes = 1; } else flags &= ~CF_ENABLE_HIT;
        if (pointCount > 0) m.rollingImpulse = old.rollingImpulse;
        for (int k = 0; k < pointCount; k++) {
          MPoint& mp2 = k == 1 ? m.p1 : m.p0;
          mp2.anchorA = mp2.anchorA - centerOffsetA;
          mp2.anchorB = mp2.anchorB - centerOffsetB;
          mp2.tangentImpulse = 0; mp2.normalImpulse = 0; mp2.totalNormalImpulse = 0; mp2.normalVelocity = 0; mp2.persisted = 0;
          for (int j = 0; j < old.pointCount; j++) {
            MPoint& mp1 = j == 1 ? old.p1 : old.p0;
            if (mp1.id == mp2.id) {
              mp2.normalImpulse = mp1.normalImpulse; mp2.tangentImpulse = mp1.tangentImpulse; mp2.persisted = 1;
              mp1.normalImpulse = 0; mp1.tangentImpulse = 0;
              break;
            }
          }
        }
        d.mN[i] = make_float4(m.normal.x, m.normal.y, m.rollingImpulse, __uint_as_float((uint32_t)m.pointCount));
        storePoint(m.p0, &d.mA0[i], &d.mB0[i], &d.mC0[i]); storePoint(m.p1, &d.mA1[i], &d.mB1[i], &d.mC1[i]);
        if (touching && !wasTouching) { flags |= CF_STARTED; fs.w = 1; if (events) fs.x = 1; }
        else if (!touching && wasTouching) { flags |= CF_STOPPED; fs.w = 1; if (events) fs.y = 1; }
      }
      d.cFlags[i] = flags;
      if (flags & CF_STATE_MASK) d.sc->anyContactFlag = 1;
    }
    d.flagScan[i] = fs;
  }
}

// Serial pass of World.Collide (:382-450) as ordered compaction: events, id recycling and graph ops in ascending contactId.
__device__ void p_contact_events(const WorldDev& d, const StepParams& sp, int tid, int stride) {
  int n = d.sc->contactNext;
  int freeCount = d.sc->contactFreeCount;
  const int endBase = d.sc->endCount[sp.endIndex];     // events appended by DestroyContact between steps come first (src/Contact.cs:236-252)
  for (int i = tid; i < n; i += stride) {
    uint32_t flags = d.cFlags[i];
    if (!(flags & CF_STATE_MASK)) continue;
    int4 rank = addT(d.flagScan[i], d.scanBlockSums4[i / (int)blockDim.x]);
    int2 sh = d.cShape[i];
    B2CContactTouchEvent ev = {sh.x, sh.y, i, d.cGen[i]};
    if (flags & CF_DISJOINT) {
      if ((flags & CF_TOUCHING) && (flags & CF_ENABLE_CONTACT_EVENTS) && endBase + rank.y < d.eventCapacity) d.endEvents[sp.endIndex][endBase + rank.y] = ev;
      d.contactFree[freeCount + rank.z] = i;
      if (flags & CF_TOUCHING) d.opList[rank.w] = i;
      d.cFlags[i] = 0;                             // slot is dead; cBody/cColor stay readable for the graph op
    } else if (flags & CF_STARTED) {
      if (flags & CF_ENABLE_CONTACT_EVENTS) d.beginEvents[rank.x] = ev;
      d.opList[rank.w] = i;
      d.cFlags[i] = (flags & ~CF_STATE_MASK) | CF_TOUCHING;
    } else {
      if ((flags & CF_ENABLE_CONTACT_EVENTS) && endBase + rank.y < d.eventCapacity) d.endEvents[sp.endIndex][endBase + rank.y] = ev;
      d.opList[rank.w] = i;
      d.cFlags[i] = flags & ~(CF_STATE_MASK | CF_TOUCHING);
    }
  }
}
__device__ void p_events_finish(const WorldDev& d, const StepParams& sp, int4* totals) {
  Scalars& s = *d.sc;
  int4 t = *totals;
  if (s.contactNext == 0) t = make_int4(0, 0, 0, 0);
  s.beginCount = t.x; s.endCount[sp.endIndex] += t.y;
  s.contactFreeCount += t.z; s.liveContactCount -= t.z;
  s.opCount = t.w;
  if (t.w > 0) s.graphDirty = 1;
}

}  // namespace b2c
#include "b2c_overflow.cuh"
namespace b2c {

// AddContactToGraph / RemoveContactFromGraph (src/ConstraintGraph.cs:135-241) in ascending contactId order.
// An op may run once it is the lowest pending op on both of its dynamic bodies, which yields exactly the sequential result.
__device__ void p_color_ops(const WorldDev& d) {
  __shared__ int sRemaining;
  int n = d.sc->opCount;
  if (n == 0) return;
  int rounds = 0;
  constexpr int kOwn = 8;                            // ops a thread keeps in registers (n <= kOwn * blockDim: no reload of ids / bodies / types per round)
  if (n <= kOwn * (int)blockDim.x) {
    // Every round costs three passes over the pending ops; with the op's contact id, bodies and body types in registers a pass is ONE
    // memory round trip (the bodies' lowest pending op, then their colour masks) instead of a chain of five dependent loads.
    int opId[kOwn], bx[kOwn], by[kOwn]; uint32_t st[kOwn];     // st: bit0 A dynamic, bit1 B dynamic, bit2 A static, bit3 B static, bit4 pending, bit5 done this round
#pragma unroll
    for (int i = 0; i < kOwn; i++) {
      const int k = threadIdx.x + i * blockDim.x;
      st[i] = 0;
      if (k < n) {
        d.opCol[k] = 0;
        const int id = d.opList[k]; const int2 b = d.cBody[id];
        const int tA = d.bodyType[b.x], tB = d.bodyType[b.y];
        opId[i] = id; bx[i] = b.x; by[i] = b.y;
        st[i] = (tA == kDynamicBody ? 1u : 0u) | (tB == kDynamicBody ? 2u : 0u) | (tA == kStaticBody ? 4u : 0u) | (tB == kStaticBody ? 8u : 0u) | 16u | ((uint32_t)(d.cColor[id] + 1) << 8);   // bits 8..: colour + 1 (0 = not in the graph: an add)
      }
    }
    while (true) {
      if (threadIdx.x == 0) sRemaining = 0;
#pragma unroll
      for (int i = 0; i < kOwn; i++) if (st[i] & 16u) {
        const int k = threadIdx.x + i * blockDim.x;
        if (st[i] & 1u) atomicMin(&d.bodyMinOp[bx[i]], k);
        if (st[i] & 2u) atomicMin(&d.bodyMinOp[by[i]], k);
      }
      __syncthreads();
#pragma unroll
      for (int i = 0; i < kOwn; i++) if (st[i] & 16u) {
        const int k = threadIdx.x + i * blockDim.x;
        const bool dynA = st[i] & 1u, dynB = st[i] & 2u;
        const int minA = dynA ? *(volatile int*)&d.bodyMinOp[bx[i]] : k, minB = dynB ? *(volatile int*)&d.bodyMinOp[by[i]] : k;
        if (minA != k || minB != k) { sRemaining = 1; continue; }
        const int id = opId[i];
        const int color = (int)(st[i] >> 8) - 1;
        if (color == -1) {                           // add
          uint32_t mA = d.bodyColorMask[bx[i]], mB = d.bodyColorMask[by[i]];
          int c = kOverflowIndex;
          if (!(st[i] & 4u) && !(st[i] & 8u)) {
            uint32_t freeBits = ~(mA | mB) & ((1u << kDynamicColorCount) - 1u);
            if (freeBits) c = __ffs((int)freeBits) - 1;
            if (c != kOverflowIndex) { if (dynA) mA |= 1u << c; if (dynB) mB |= 1u << c; }
          } else if (dynA) {
            uint32_t freeBits = ~mA & 0x007ffffeu;    // colors 22..1
            if (freeBits) { c = 31 - __clz((int)freeBits); mA |= 1u << c; }
          } else if (dynB) {
            uint32_t freeBits = ~mB & 0x007ffffeu;
            if (freeBits) { c = 31 - __clz((int)freeBits); mB |= 1u << c; }
          }
          if (dynA) d.bodyColorMask[bx[i]] = mA;
          if (dynB) d.bodyColorMask[by[i]] = mB;
          d.cColor[id] = c;
          d.opCol[k] = (uint8_t)(0x40 | c);
        } else {                                     // remove
          d.opCol[k] = (uint8_t)(0x80 | color);
          if (color != kOverflowIndex) {
            if (dynA) d.bodyColorMask[bx[i]] &= ~(1u << color);
            if (dynB) d.bodyColorMask[by[i]] &= ~(1u << color);
          }
          d.cColor[id] = -1;
        }
        st[i] = (st[i] & 15u) | 32u;
      }
      __syncthreads();
      // every op that was pending in this round put its index on its bodies: take them all back (the ones that ran AND the ones that wait)
#pragma unroll
      for (int i = 0; i < kOwn; i++) if (st[i] & 48u) { d.bodyMinOp[bx[i]] = 0x7f7f7f7f; d.bodyMinOp[by[i]] = 0x7f7f7f7f; st[i] &= ~32u; }
      rounds++;
      __syncthreads();
      if (sRemaining == 0) break;
      __syncthreads();
    }
  } else {
  for (int k = threadIdx.x; k < n; k += blockDim.x) { d.opDone[k] = 0; d.opCol[k] = 0; }
  __syncthreads();
  while (true) {
    if (threadIdx.x == 0) sRemaining = 0;
    for (int k = threadIdx.x; k < n; k += blockDim.x) {
      if (d.opDone[k]) continue;
      int2 b = d.cBody[d.opList[k]];
      if (d.bodyType[b.x] == kDynamicBody) atomicMin(&d.bodyMinOp[b.x], k);
      if (d.bodyType[b.y] == kDynamicBody) atomicMin(&d.bodyMinOp[b.y], k);
    }
    __syncthreads();
    for (int k = threadIdx.x; k < n; k += blockDim.x) {
      if (d.opDone[k]) continue;
      int id = d.opList[k];
      int2 b = d.cBody[id];
      int typeA = d.bodyType[b.x], typeB = d.bodyType[b.y];
      bool ready = (typeA != kDynamicBody || d.bodyMinOp[b.x] == k) && (typeB != kDynamicBody || d.bodyMinOp[b.y] == k);
      if (!ready) { sRemaining = 1; continue; }
      int color = d.cColor[id];
      if (color == -1) {                           // add
        uint32_t mA = d.bodyColorMask[b.x], mB = d.bodyColorMask[b.y];
        int c = kOverflowIndex;
        if (typeA != kStaticBody && typeB != kStaticBody) {
          uint32_t freeBits = ~(mA | mB) & ((1u << kDynamicColorCount) - 1u);
          if (freeBits) c = __ffs((int)freeBits) - 1;
          if (c != kOverflowIndex) { if (typeA == kDynamicBody) mA |= 1u << c; if (typeB == kDynamicBody) mB |= 1u << c; }
        } else if (typeA == kDynamicBody) {
          uint32_t freeBits = ~mA & 0x007ffffeu;    // colors 22..1
          if (freeBits) { c = 31 - __clz((int)freeBits); mA |= 1u << c; }
        } else if (typeB == kDynamicBody) {
          uint32_t freeBits = ~mB & 0x007ffffeu;
          if (freeBits) { c = 31 - __clz((int)freeBits); mB |= 1u << c; }
        }
        if (typeA == kDynamicBody) d.bodyColorMask[b.x] = mA;
        if (typeB == kDynamicBody) d.bodyColorMask[b.y] = mB;
        d.cColor[id] = c;
        d.opCol[k] = (uint8_t)(0x40 | c);
      } else {                                     // remove
        d.opCol[k] = (uint8_t)(0x80 | color);
        if (color != kOverflowIndex) {
          if (typeA == kDynamicBody) d.bodyColorMask[b.x] &= ~(1u << color);
          if (typeB == kDynamicBody) d.bodyColorMask[b.y] &= ~(1u << color);
        }
        d.cColor[id] = -1;
      }
      d.opDone[k] = 2;                             // done this round: still has to reset its bodies below
    }
    __syncthreads();
    for (int k = threadIdx.x; k < n; k += blockDim.x) {
      if (d.opDone[k] == 1) continue;
      int2 b = d.cBody[d.opList[k]];
      d.bodyMinOp[b.x] = 0x7f7f7f7f; d.bodyMinOp[b.y] = 0x7f7f7f7f;
      if (d.opDone[k] == 2) d.opDone[k] = 1;
    }
    rounds++;
    __syncthreads();
    if (sRemaining == 0) break;
    __syncthreads();
  }
  }
  if (threadIdx.x == 0) { d.sc->colorRounds = rounds; if (d.debugChecksums == 2) d.sc->dbg[4] = globalTimer(); }
  __syncthreads();
  colReplayOps(d, n);
}

// ------------------------------------------------------------------------------------------------------------
// Phase 2b: solver stages (src/Solver.cs:599-727, src/ContactSolver.cs:362-1026 AVX lanes, src/joints/RevoluteJoint.cs)
// ------------------------------------------------------------------------------------------------------------
struct BodyW { V2 v; float w; uint32_t flags; V2 dp; Rot dq; };
__device__ __forceinline__ BodyW gatherBody(const WorldDev& d, int index) {        // GatherBodies :415-451
  BodyW b;
  if (index < 0) { b.v = mk(0, 0); b.w = 0; b.flags = 0; b.dp = mk(0, 0); b.dq = mkRot(1, 0); return b; }
  float4 a = d.st0[index], c = d.st1[index];
  b.v = mk(a.x, a.y); b.w = a.z; b.flags = fbits(a.w); b.dp = mk(c.x, c.y); b.dq = mkRot(c.z, c.w);
  return b;
}
__device__ __forceinline__ void scatterBody(const WorldDev& d, int index, const BodyW& b) {   // ScatterBodies :452-479 (v,w only change)
  if (index >= 0 && (b.flags & kDynamicFlag)) d.st0[index] = make_float4(b.v.x, b.v.y, b.w, __uint_as_float(b.flags));
}
// Body-state accessors for the constraint lanes: rows in global memory (grid path) or staged in shared memory (island path).
struct BodyRows {
  float4* st0; float4* st1;
  __device__ __forceinline__ BodyW load(int index) const {
    BodyW b;
    if (index < 0) { b.v = mk(0, 0); b.w = 0; b.flags = 0; b.dp = mk(0, 0); b.dq = mkRot(1, 0); return b; }
    float4 a = st0[index], c = st1[index];
    b.v = mk(a.x, a.y); b.w = a.z; b.flags = fbits(a.w); b.dp = mk(c.x, c.y); b.dq = mkRot(c.z, c.w);
    return b;
  }
  __device__ __forceinline__ BodyW loadA(int index) const { return load(index); }   // (the cluster's rows tell body A from body B: b2c_cluster_dsm.cuh)
  __device__ __forceinline__ BodyW loadB(int index) const { return load(index); }
  __device__ __forceinline__ void store(int index, const BodyW& b) const {
    if (index >= 0 && (b.flags & kDynamicFlag)) st0[index] = make_float4(b.v.x, b.v.y, b.w, __uint_as_float(b.flags));
  }
};

struct Lane {       // one constraint's read-only planes in registers
  int indexA, indexB; float mA, mB, iA, iB; V2 n; float friction, tangentSpeed, rollingResistance, rollingMass;
  float biasRate, massScale, impulseScale, restitution; V2 rA1, rB1, rA2, rB2; float normalMass1, tangentMass1, normalMass2, tangentMass2;
  float baseSep1, baseSep2, relVel1, relVel2;
};
__device__ __forceinline__ void loadLane(const WorldDev& d, int t, Lane& c, bool anchorsOnly) {
  float4 A = d.sA[t], B = d.sB[t], E = d.sE[t], F = d.sF[t];
  c.indexA = (int)fbits(A.x); c.indexB = (int)fbits(A.y); c.mA = A.z; c.mB = A.w; c.iA = B.x; c.iB = B.y; c.n = mk(B.z, B.w);
  c.rA1 = mk(E.x, E.y); c.rB1 = mk(E.z, E.w); c.rA2 = mk(F.x, F.y); c.rB2 = mk(F.z, F.w);
  if (anchorsOnly) return;
  float4 C = d.sC[t], D = d.sD[t], G = d.sG[t], H = d.sH[t];
  c.friction = C.x; c.tangentSpeed = C.y; c.rollingResistance = C.z; c.rollingMass = C.w;
  c.biasRate = D.x; c.massScale = D.y; c.impulseScale = D.z; c.restitution = D.w;
  c.normalMass1 = G.x; c.tangentMass1 = G.y; c.normalMass2 = G.z; c.tangentMass2 = G.w;
  c.baseSep1 = H.x; c.baseSep2 = H.y; c.relVel1 = H.z; c.relVel2 = H.w;
}
__device__ __forceinline__ void applyP(BodyW& bA, BodyW& bB, const Lane& c, V2 rA, V2 rB, float Px, float Py) {   // :794-805
  bA.v.x = bA.v.x - c.mA * Px; bA.v.y = bA.v.y - c.mA * Py;
  bA.w = bA.w - c.iA * (rA.x * Py - rA.y * Px);
  bB.v.x = bB.v.x + c.mB * Px; bB.v.y = bB.v.y + c.mB * Py;
  bB.w = bB.w + c.iB * (rB.x * Py - rB.y * Px);
}

// PrepareContactsTask :480-685 — builds one constraint lane (registers) from the contact slot `ci`
__device__ __forceinline__ void computeLane(const WorldDev& d, const StepParams& sp, int ci, Lane& c, float4& I, float4& J) {
  float warmStartScale = sp.enableWarmStarting ? 1.0f : 0.0f;
  int2 idx = d.cSim[ci];
  float4 mass = d.cMass[ci];
  float mA = mass.x, iA = mass.y, mB = mass.z, iB = mass.w;
  V2 vA = mk(0, 0), vB = mk(0, 0); float wA = 0, wB = 0;
  if (idx.x != -1) { float4 s = d.st0[idx.x]; vA = mk(s.x, s.y); wA = s.z; }
  if (idx.y != -1) { float4 s = d.st0[idx.y]; vB = mk(s.x, s.y); wB = s.z; }
  float k = iA + iB;
  Softness soft = sp.contactSoftness;
  if (idx.x == -1 || idx.y == -1) soft = sp.staticSoftness;
  else if (sp.enableContactSoftening) {
    float contactHertz = netMin(sp.contactHertz, 0.125f * sp.inv_h);
    float ratio = 1.0f;
    if (mA < mB) ratio = netMax(0.5f, mA / mB); else if (mB < mA) ratio = netMax(0.5f, mB / mA);
    soft = makeSoft(ratio * contactHertz, ratio * sp.contactDampingRatio, sp.h);
  }
  float4 mn = d.mN[ci], mat = d.cMat[ci];
  V2 normal = mk(mn.x, mn.y);
  int pointCount = (int)fbits(mn.w);
  V2 tangent = rightPerp(normal);
  c.indexA = idx.x; c.indexB = idx.y; c.mA = mA; c.mB = mB; c.iA = iA; c.iB = iB; c.n = normal;
  c.friction = mat.x; c.tangentSpeed = mat.w; c.rollingResistance = mat.z; c.rollingMass = k > 0.0f ? 1.0f / k : 0.0f;
  c.biasRate = soft.biasRate; c.massScale = soft.massScale; c.impulseScale = soft.impulseScale; c.restitution = mat.y;
  c.rA2 = mk(0, 0); c.rB2 = mk(0, 0); c.normalMass2 = 0; c.tangentMass2 = 0; c.baseSep2 = 0; c.relVel2 = 0;
  I = make_float4(0, 0, 0, 0);
  {
    float4 a = d.mA0[ci], b = d.mB0[ci];
    V2 rA = mk(a.x, a.y), rB = mk(a.z, a.w);
    c.rA1 = rA; c.rB1 = rB;
    c.baseSep1 = b.x - dot(rB - rA, normal);
    I.x = warmStartScale * b.y; I.y = warmStartScale * b.z;
    float rnA = cross(rA, normal), rnB = cross(rB, normal);
    float kNormal = mA + mB + iA * rnA * rnA + iB * rnB * rnB;
    c.normalMass1 = kNormal > 0.0f ? 1.0f / kNormal : 0.0f;
    float rtA = cross(rA, tangent), rtB = cross(rB, tangent);
    float kTangent = mA + mB + iA * rtA * rtA + iB * rtB * rtB;
    c.tangentMass1 = kTangent > 0.0f ? 1.0f / kTangent : 0.0f;
    V2 vrA = vA + crossSV(wA, rA), vrB = vB + crossSV(wB, rB);
    c.relVel1 = dot(normal, vrB - vrA);
  }
  if (pointCount == 2) {
    float4 a = d.mA1[ci], b = d.mB1[ci];
    V2 rA = mk(a.x, a.y), rB = mk(a.z, a.w);
    c.rA2 = rA; c.rB2 = rB;
    c.baseSep2 = b.x - dot(rB - rA, normal);
    I.z = warmStartScale * b.y; I.w = warmStartScale * b.z;
    float rnA = cross(rA, normal), rnB = cross(rB, normal);
    float kNormal = mA + mB + iA * rnA * rnA + iB * rnB * rnB;
    c.normalMass2 = kNormal > 0.0f ? 1.0f / kNormal : 0.0f;
    float rtA = cross(rA, tangent), rtB = cross(rB, tangent);
    float kTangent = mA + mB + iA * rtA * rtA + iB * rtB * rtB;
    c.tangentMass2 = kTangent > 0.0f ? 1.0f / kTangent : 0.0f;
    V2 vrA = vA + crossSV(wA, rA), vrB = vB + crossSV(wB, rB);
    c.relVel2 = dot(normal, vrB - vrA);
  }
  J = make_float4(0.0f, 0.0f, warmStartScale * mn.z, 0.0f);
}
__device__ __forceinline__ void writeLane(const WorldDev& d, int t, const Lane& c, float4 I, float4 J) {
  d.sA[t] = make_float4(__int_as_float(c.indexA), __int_as_float(c.indexB), c.mA, c.mB);
  d.sB[t] = make_float4(c.iA, c.iB, c.n.x, c.n.y);
  d.sC[t] = make_float4(c.friction, c.tangentSpeed, c.rollingResistance, c.rollingMass);
  d.sD[t] = make_float4(c.biasRate, c.massScale, c.impulseScale, c.restitution);
  d.sE[t] = make_float4(c.rA1.x, c.rA1.y, c.rB1.x, c.rB1.y);
  d.sF[t] = make_float4(c.rA2.x, c.rA2.y, c.rB2.x, c.rB2.y);
  d.sG[t] = make_float4(c.normalMass1, c.tangentMass1, c.normalMass2, c.tangentMass2);
  d.sH[t] = make_float4(c.baseSep1, c.baseSep2, c.relVel1, c.relVel2);
  d.sI[t] = I; d.sJ[t] = J;
}
__device__ void prepareContact(const WorldDev& d, const StepParams& sp, int t) {
  Lane c; float4 I, J;
  computeLane(d, sp, d.sContact[t], c, I, J);
  writeLane(d, t, c, I, J);
}

// WarmStartContactsTask :686-744
template <class Rows> __device__ __forceinline__ void warmStartLane(const Rows& bs, const Lane& c, int ia, int ib, float4 I, float rollingImpulse) {
  BodyW bA = bs.loadA(ia), bB = bs.loadB(ib);
  float tangentX = c.n.y, tangentY = 0.0f - c.n.x;
  {
    float Px = I.x * c.n.x + I.y * tangentX, Py = I.x * c.n.y + I.y * tangentY;
    bA.w = bA.w - c.iA * (c.rA1.x * Py - c.rA1.y * Px);
    bA.v.x = bA.v.x - c.mA * Px; bA.v.y = bA.v.y - c.mA * Py;
    bB.w = bB.w + c.iB * (c.rB1.x * Py - c.rB1.y * Px);
    bB.v.x = bB.v.x + c.mB * Px; bB.v.y = bB.v.y + c.mB * Py;
  }
  {
    float Px = I.z * c.n.x + I.w * tangentX, Py = I.z * c.n.y + I.w * tangentY;
    bA.w = bA.w - c.iA * (c.rA2.x * Py - c.rA2.y * Px);
    bA.v.x = bA.v.x - c.mA * Px; bA.v.y = bA.v.y - c.mA * Py;
    bB.w = bB.w + c.iB * (c.rB2.x * Py - c.rB2.y * Px);
    bB.v.x = bB.v.x + c.mB * Px; bB.v.y = bB.v.y + c.mB * Py;
  }
  bA.w = bA.w - c.iA * rollingImpulse;
  bB.w = bB.w + c.iB * rollingImpulse;
  bs.store(ia, bA); bs.store(ib, bB);
}
__device__ __forceinline__ void warmStartContact(const WorldDev& d, int t) {
  Lane c; loadLane(d, t, c, true);
  BodyRows bs{d.st0, d.st1};
  warmStartLane(bs, c, c.indexA, c.indexB, d.sI[t], d.sJ[t].z);
}

__device__ __forceinline__ void solveNormalPoint(BodyW& bA, BodyW& bB, const Lane& c, V2 dp, V2 rA, V2 rB, float baseSeparation, float normalMass,
                                                 float& normalImpulse, float& totalP, float& total, float inv_h, float contactSpeed,
                                                 float biasRate, float massScale, float impulseScale) {   // :773-806
  V2 rsA = rotate(bA.dq, rA), rsB = rotate(bB.dq, rB);
  float dsx = dp.x + (rsB.x - rsA.x), dsy = dp.y + (rsB.y - rsA.y);
  float s = (c.n.x * dsx + c.n.y * dsy) + baseSeparation;
  bool mask = s > 0.0f;
  float specBias = s * inv_h, softBias = avxMax(biasRate * s, contactSpeed);
  float bias = mask ? specBias : softBias;
  float pointMassScale = mask ? 1.0f : massScale;
  float pointImpulseScale = mask ? 0.0f : impulseScale;
  float dvx = (bB.v.x - bB.w * rB.y) - (bA.v.x - bA.w * rA.y);
  float dvy = (bB.v.y + bB.w * rB.x) - (bA.v.y + bA.w * rA.x);
  float vn = dvx * c.n.x + dvy * c.n.y;
  float negImpulse = normalMass * (pointMassScale * vn + bias) + pointImpulseScale * normalImpulse;
  float newImpulse = avxMax(normalImpulse - negImpulse, 0.0f);
  float impulse = newImpulse - normalImpulse;
  normalImpulse = newImpulse;
  totalP = totalP + newImpulse;
  total = total + newImpulse;
  applyP(bA, bB, c, rA, rB, impulse * c.n.x, impulse * c.n.y);
}
__device__ __forceinline__ void solveFrictionPoint(BodyW& bA, BodyW& bB, const Lane& c, V2 rA, V2 rB, float tangentX, float tangentY, float tangentMass,
                                                   float normalImpulse, float& tangentImpulse) {          // :844-871
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
  applyP(bA, bB, c, rA, rB, impulse * tangentX, impulse * tangentY);
}
// SolveContactsTask :745-912 (useBias=true: Solve, false: Relax)
template <bool USE_BIAS, class Rows> __device__ __forceinline__ void solveLane(const Rows& bs, const StepParams& sp, const Lane& c, int ia, int ib, float4& I, float4& J) {
  float contactSpeed = -sp.contactSpeed;
  BodyW bA = bs.loadA(ia), bB = bs.loadB(ib);
  float biasRate, massScale, impulseScale;
  if (USE_BIAS) { biasRate = c.massScale * c.biasRate; massScale = c.massScale; impulseScale = c.impulseScale; }
  else { biasRate = 0.0f; massScale = 1.0f; impulseScale = 0.0f; }
  float totalNormalImpulse = 0.0f;
  V2 dp = mk(bB.dp.x - bA.dp.x, bB.dp.y - bA.dp.y);
  solveNormalPoint(bA, bB, c, dp, c.rA1, c.rB1, c.baseSep1, c.normalMass1, I.x, J.x, totalNormalImpulse, sp.inv_h, contactSpeed, biasRate, massScale, impulseScale);
  solveNormalPoint(bA, bB, c, dp, c.rA2, c.rB2, c.baseSep2, c.normalMass2, I.z, J.y, totalNormalImpulse, sp.inv_h, contactSpeed, biasRate, massScale, impulseScale);
  float tangentX = c.n.y, tangentY = 0.0f - c.n.x;
  solveFrictionPoint(bA, bB, c, c.rA1, c.rB1, tangentX, tangentY, c.tangentMass1, I.x, I.y);
  solveFrictionPoint(bA, bB, c, c.rA2, c.rB2, tangentX, tangentY, c.tangentMass2, I.z, I.w);
  {
    float deltaLambda = c.rollingMass * (bA.w - bB.w);
    float lambda = J.z;
    float maxLambda = c.rollingResistance * totalNormalImpulse;
    J.z = avxMax(0.0f - maxLambda, avxMin(lambda + deltaLambda, maxLambda));
    deltaLambda = J.z - lambda;
    bA.w = bA.w - c.iA * deltaLambda;
    bB.w = bB.w + c.iB * deltaLambda;
  }
  bs.store(ia, bA); bs.store(ib, bB);
}
template <bool USE_BIAS> __device__ __forceinline__ void solveContact(const WorldDev& d, const StepParams& sp, int t) {
  Lane c; loadLane(d, t, c, false);
  float4 I = d.sI[t], J = d.sJ[t];
  BodyRows bs{d.st0, d.st1};
  solveLane<USE_BIAS>(bs, sp, c, c.indexA, c.indexB, I, J);
  d.sI[t] = I; d.sJ[t] = J;
}

// ApplyRestitutionTask :913-989, with quirk Q1 (multiplies by the compare mask, not the blended mass).
__device__ __forceinline__ void restitutionPoint(BodyW& bA, BodyW& bB, const Lane& c, V2 rA, V2 rB, float relativeVelocity, float& normalImpulse,
                                                 float& totalP, float threshold, bool restitutionMask) {
  bool mask1 = (relativeVelocity + threshold) > 0.0f;
  bool mask2 = totalP == 0.0f;
  float mask = (mask1 || mask2 || restitutionMask) ? __uint_as_float(0xffffffffu) : 0.0f;
  float dvx = (bB.v.x - bB.w * rB.y) - (bA.v.x - bA.w * rA.y);
  float dvy = (bB.v.y + bB.w * rB.x) - (bA.v.y + bA.w * rA.x);
  float vn = dvx * c.n.x + dvy * c.n.y;
  float negImpulse = mask * (vn + c.restitution * relativeVelocity);
  float newImpulse = avxMax(normalImpulse - negImpulse, 0.0f);
  float deltaImpulse = newImpulse - normalImpulse;
  normalImpulse = newImpulse;
  totalP = totalP + deltaImpulse;
  applyP(bA, bB, c, rA, rB, deltaImpulse * c.n.x, deltaImpulse * c.n.y);
}
__device__ __forceinline__ void restitutionLane(const BodyRows& bs, const StepParams& sp, const Lane& c, int ia, int ib, float4& I, float4& J) {
  BodyW bA = bs.loadA(ia), bB = bs.loadB(ib);
  bool restitutionMask = c.restitution == 0.0f;
  restitutionPoint(bA, bB, c, c.rA1, c.rB1, c.relVel1, I.x, J.x, sp.restitutionThreshold, restitutionMask);
  restitutionPoint(bA, bB, c, c.rA2, c.rB2, c.relVel2, I.z, J.y, sp.restitutionThreshold, restitutionMask);
  bs.store(ia, bA); bs.store(ib, bB);
}
__device__ __forceinline__ void restitutionContact(const WorldDev& d, const StepParams& sp, int t) {
  const int ci = d.sContact[t];
  if (ci < 0 || !d.cRestGroup[ci]) return;         // AllZeroW group (or padding lane)
  Lane c; loadLane(d, t, c, false);
  float4 I = d.sI[t], J = d.sJ[t];
  BodyRows bs{d.st0, d.st1};
  restitutionLane(bs, sp, c, c.indexA, c.indexB, I, J);
  d.sI[t] = I; d.sJ[t] = J;
}
// The reference skips an 8-lane SIMD group when all its lanes have restitution 0 (AllZeroW, :923); otherwise EVERY lane of the group runs
// the code above, and under quirk Q1 a zero-restitution lane of such a group has its normal impulses wiped. A group is 8 consecutive
// entries of the colour's contactSims array (colList): every touching contact with restitution != 0 flags the members of its group.
__device__ void p_rest_groups(const WorldDev& d, int tid, int stride) {
  const int n = d.sc->contactNext;
  for (int i = tid; i < n; i += stride) {
    if ((d.cFlags[i] & (CF_ALIVE | CF_TOUCHING)) != (CF_ALIVE | CF_TOUCHING)) continue;
    if (d.cMat[i].y == 0.0f) continue;
    const int c = d.cColor[i], li = d.cLocal[i];
    if (c < 0 || c >= kOverflowIndex || li < 0) continue;      // the overflow colour is solved by scalar code without groups (:265-322)
    const int* arr = d.colList + (size_t)c * d.contactCapacity;
    const int g0 = li & ~7, len = d.sc->colLen[c];
    for (int k = g0; k < g0 + 8 && k < len; k++) d.cRestGroup[arr[k]] = 1;
  }
}
// StoreImpulsesTask :990-1025
__device__ __forceinline__ void storeImpulses(const WorldDev& d, int t) {
  int ci = d.sContact[t];
  float4 I = d.sI[t], J = d.sJ[t], H = d.sH[t];
  d.mN[ci].z = J.z;
  float4 b0 = d.mB0[ci], b1 = d.mB1[ci];
  b0.y = I.x; b0.z = I.y; b0.w = J.x; b1.y = I.z; b1.z = I.w; b1.w = J.y;
  d.mB0[ci] = b0; d.mB1[ci] = b1;
  d.mC0[ci].z = H.z; d.mC1[ci].z = H.w;
}

__device__ __forceinline__ void storeImpulsesRegs(const WorldDev& d, int ci, float4 I, float4 J, float relVel1, float relVel2) {
  d.mN[ci].z = J.z;
  float4 b0 = d.mB0[ci], b1 = d.mB1[ci];
  b0.y = I.x; b0.z = I.y; b0.w = J.x; b1.y = I.z; b1.z = I.w; b1.w = J.y;
  d.mB0[ci] = b0; d.mB1[ci] = b1;
  d.mC0[ci].z = relVel1; d.mC1[ci].z = relVel2;
}

#include "b2c_joints.cuh"

__device__ __noinline__ void prepareJointOther(const StepParams& sp, JointDev& j, float invMassA, float invMassB) {
  switch (j.type) {
    case kDistanceJoint: prepareDistance(sp, j, invMassA, invMassB); break;
    case kMotorJoint: prepareMotor(sp, j); break;
    case kPrismaticJoint: preparePrismatic(sp, j); break;
    case kWeldJoint: prepareWeld(sp, j); break;
    case kWheelJoint: prepareWheel(sp, j); break;
    default: break;
  }
}
// ---- JointSim.PrepareJoint src/Joint.cs:125-130 + IJoint.Prepare: the body/frame part is the same text in every joint type
// (RevoluteJoint.cs:72-95, WeldJoint.cs:46-66, MotorJoint.cs:60-81, PrismaticJoint.cs:86-106, WheelJoint.cs:91-111, DistanceJoint.cs:91-113) ----
__device__ void prepareJoint(const WorldDev& d, const StepParams& sp, JointDev& j) {
  float hertz = netMin(j.constraintHertz, 0.25f * sp.inv_h);
  j.constraintSoftness = makeSoft(hertz, j.constraintDampingRatio, sp.h);
  if (j.type == kFilterJoint) return;
  float4 massA = d.bodyMass[j.bodyIdA], massB = d.bodyMass[j.bodyIdB];
  j.invMassA = massA.x; j.invIA = massA.y; j.invMassB = massB.x; j.invIB = massB.y;
  j.indexA = d.bodySim[j.bodyIdA]; j.indexB = d.bodySim[j.bodyIdB];
  Xf xfA = toXf(d.bodyXf[j.bodyIdA]), xfB = toXf(d.bodyXf[j.bodyIdB]);
  j.frameA.q = mulRot(xfA.q, j.localFrameA.q);
  j.frameA.p = rotate(xfA.q, j.localFrameA.p - mk(massA.z, massA.w));
  j.frameB.q = mulRot(xfB.q, j.localFrameB.q);
  j.frameB.p = rotate(xfB.q, j.localFrameB.p - mk(massB.z, massB.w));
  float2 cA = d.bodyCenter[j.bodyIdA], cB = d.bodyCenter[j.bodyIdB];
  j.deltaCenter = mk(cB.x, cB.y) - mk(cA.x, cA.y);
  if (j.type != kRevoluteJoint) { prepareJointOther(sp, j, massA.x, massB.x); return; }
  // revolute: src/joints/RevoluteJoint.cs:96-104
  float k = j.invIA + j.invIB;
  j.axialMass = k > 0.0f ? 1.0f / k : 0.0f;
  j.springSoftness = makeSoft(j.hertz, j.dampingRatio, sp.h);
  if (!sp.enableWarmStarting) { j.linearImpulse = mk(0, 0); j.springImpulse = 0; j.motorImpulse = 0; j.lowerImpulse = 0; j.upperImpulse = 0; }
}
__device__ void warmStartRevolute(const WorldDev& d, JointDev& j) {
  BodyW bA = gatherBody(d, j.indexA), bB = gatherBody(d, j.indexB);
  V2 rA = rotate(bA.dq, j.frameA.p), rB = rotate(bB.dq, j.frameB.p);
  float axialImpulse = j.springImpulse + j.motorImpulse + j.lowerImpulse - j.upperImpulse;
  if (bA.flags & kDynamicFlag) {
    bA.v = mulSub(bA.v, j.invMassA, j.linearImpulse);
    bA.w -= j.invIA * (cross(rA, j.linearImpulse) + axialImpulse);
  }
  if (bB.flags & kDynamicFlag) {
    bB.v = mulAdd(bB.v, j.invMassB, j.linearImpulse);
    bB.w += j.invIB * (cross(rB, j.linearImpulse) + axialImpulse);
  }
  scatterBody(d, j.indexA, bA); scatterBody(d, j.indexB, bB);
}
__device__ void solveRevolute(const WorldDev& d, const StepParams& sp, JointDev& j, bool useBias) {
  float mA = j.invMassA, mB = j.invMassB, iA = j.invIA, iB = j.invIB;
  BodyW bA = gatherBody(d, j.indexA), bB = gatherBody(d, j.indexB);
  V2 vA = bA.v, vB = bB.v; float wA = bA.w, wB = bB.w;
  Rot qA = mulRot(bA.dq, j.frameA.q), qB = mulRot(bB.dq, j.frameB.q);
  Rot relQ = invMulRot(qA, qB);
  bool fixedRotation = iA + iB == 0.0f;
  if (j.enableSpring && !fixedRotation) {
    float jointAngle = atan2f(relQ.s, relQ.c);
    float jointAngleDelta = fmodf((jointAngle - j.targetAngle) + kPi, kTau) - kPi;
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
    float maxImpulse = sp.h * j.maxMotorTorque;
    j.motorImpulse = netClamp(j.motorImpulse + impulse, -maxImpulse, maxImpulse);
    impulse = j.motorImpulse - oldImpulse;
    wA -= iA * impulse; wB += iB * impulse;
  }
  if (j.enableLimit && !fixedRotation) {
    float jointAngle = atan2f(relQ.s, relQ.c);
    {
      float C = jointAngle - j.lowerAngle;
      float bias = 0.0f, massScale = 1.0f, impulseScale = 0.0f;
      if (C > 0.0f) bias = C * sp.inv_h;
      else if (useBias) { bias = j.constraintSoftness.biasRate * C; massScale = j.constraintSoftness.massScale; impulseScale = j.constraintSoftness.impulseScale; }
      float Cdot = wB - wA;
      float oldImpulse = j.lowerImpulse;
      float impulse = -massScale * j.axialMass * (Cdot + bias) - impulseScale * oldImpulse;
      j.lowerImpulse = netMax(oldImpulse + impulse, 0.0f);
      impulse = j.lowerImpulse - oldImpulse;
      wA -= iA * impulse; wB += iB * impulse;
    }
    {
      float C = j.upperAngle - jointAngle;
      float bias = 0.0f, massScale = 1.0f, impulseScale = 0.0f;
      if (C > 0.0f) bias = C * sp.inv_h;
      else if (useBias) { bias = j.constraintSoftness.biasRate * C; massScale = j.constraintSoftness.massScale; impulseScale = j.constraintSoftness.impulseScale; }
      float Cdot = wA - wB;
      float oldImpulse = j.upperImpulse;
      float impulse = -massScale * j.axialMass * (Cdot + bias) - impulseScale * oldImpulse;
      j.upperImpulse = netMax(oldImpulse + impulse, 0.0f);
      impulse = j.upperImpulse - oldImpulse;
      wA += iA * impulse; wB -= iB * impulse;
    }
  }
  {
    V2 rA = rotate(bA.dq, j.frameA.p), rB = rotate(bB.dq, j.frameB.p);
    V2 Cdot = (vB + crossSV(wB, rB)) - (vA + crossSV(wA, rA));
    V2 bias = mk(0, 0);
    float massScale = 1.0f, impulseScale = 0.0f;
    if (useBias) {
      V2 separation = ((bB.dp - bA.dp) + (rB - rA)) + j.deltaCenter;
      bias = j.constraintSoftness.biasRate * separation;
      massScale = j.constraintSoftness.massScale; impulseScale = j.constraintSoftness.impulseScale;
    }
    float Kxx = mA + mB + rA.y * rA.y * iA + rB.y * rB.y * iB;
    float Kxy = -rA.y * rA.x * iA - rB.y * rB.x * iB;
    float Kyy = mA + mB + rA.x * rA.x * iA + rB.x * rB.x * iB;
    V2 rhs = Cdot + bias;
    float det = Kxx * Kyy - Kxy * Kxy;              // Mat22.Solve, include/MathFunctions.cs:211-215 (cx=(Kxx,Kxy), cy=(Kxy,Kyy))
    if (det != 0.0f) det = 1.0f / det;
    V2 b = mk(det * (Kyy * rhs.x - Kxy * rhs.y), det * (Kxx * rhs.y - Kxy * rhs.x));
    V2 impulse = (-massScale * b) - (impulseScale * j.linearImpulse);
    j.linearImpulse = j.linearImpulse + impulse;
    vA = mulSub(vA, mA, impulse); wA -= iA * cross(rA, impulse);
    vB = mulAdd(vB, mB, impulse); wB += iB * cross(rB, impulse);
  }
  bA.v = vA; bA.w = wA; bB.v = vB; bB.w = wB;
  scatterBody(d, j.indexA, bA); scatterBody(d, j.indexB, bB);
}
// Joint types other than revolute are kept out of line (their own frames and registers): the cooperative front kernel is compiled for
// 512 threads x 128 registers and the configs' hot joint is the revolute, which stays inline.
__device__ __noinline__ void warmStartJointOther(const WorldDev& d, JointDev& j) {
  switch (j.type) {
    case kDistanceJoint: warmStartDistance(d, j); break;
    case kMotorJoint: warmStartLinearAngular(d, j, j.linearImpulse + j.linearImpulse2, j.motorImpulse + j.springImpulse); break;
    case kPrismaticJoint: warmStartPrismatic(d, j); break;
    case kWeldJoint: warmStartLinearAngular(d, j, j.linearImpulse, j.springImpulse); break;
    case kWheelJoint: warmStartWheel(d, j); break;
    default: break;                                                    // filter: IJoint default, empty
  }
}
__device__ __noinline__ void solveJointOther(const WorldDev& d, const StepParams& sp, JointDev& j, bool useBias) {
  switch (j.type) {
    case kDistanceJoint: solveDistance(d, sp, j, useBias); break;
    case kMotorJoint: solveMotor(d, sp, j); break;
    case kPrismaticJoint: solvePrismatic(d, sp, j, useBias); break;
    case kWeldJoint: solveWeld(d, sp, j, useBias); break;
    case kWheelJoint: solveWheel(d, sp, j, useBias); break;
    default: break;
  }
}
__device__ __forceinline__ void warmStartJoint(const WorldDev& d, JointDev& j) {      // JointSim.WarmStart src/Joint.cs:131
  if (j.type == kRevoluteJoint) warmStartRevolute(d, j); else warmStartJointOther(d, j);
}
__device__ __forceinline__ void solveJointOnly(const WorldDev& d, const StepParams& sp, JointDev& j, bool useBias) {    // JointSim.Solve src/Joint.cs:132
  if (j.type == kRevoluteJoint) solveRevolute(d, sp, j, useBias); else solveJointOther(d, sp, j, useBias);
}
// Joints of the overflow colour (src/Joint.cs:377-394): one thread, array order, before the overflow contacts of the stage
// (src/Solver.cs:633,648,669). The relax pass keeps useBias (Q4) and no joint event can come from here: only SolveJointsTask tests the thresholds.
__device__ __noinline__ void ovWarmStartAll(const WorldDev& d) {
  for (int k = d.jointColorStart[kOverflowIndex]; k < d.jointColorStart[kGraphColorCount]; k++) warmStartJoint(d, d.joints[d.jointSorted[k]]);
  const int nc = d.sc->colLen[kOverflowIndex];
  if (nc > 0) ovWarmStart(d, nc);
}
__device__ __noinline__ void ovSolveAll(const WorldDev& d, const StepParams& sp) {
  for (int k = d.jointColorStart[kOverflowIndex]; k < d.jointColorStart[kGraphColorCount]; k++) solveJointOnly(d, sp, d.joints[d.jointSorted[k]], true);
  const int nc = d.sc->colLen[kOverflowIndex];
  if (nc > 0) ovSolve(d, sp, nc, true);
}
__device__ __forceinline__ void solveJoint(const WorldDev& d, const StepParams& sp, JointDev& j, bool useBias) {    // JointSim.Solve src/Joint.cs:132 + SolveJointsTask src/Solver.cs:180-198
  solveJointOnly(d, sp, j, useBias);
  if (useBias && (j.forceThreshold < FLT_MAX || j.torqueThreshold < FLT_MAX) && !d.jointEventFlag[j.jointId]) {   // src/Solver.cs:186-196
    float force, torque; jointReaction(j, sp.inv_h, &force, &torque);
    if (force >= j.forceThreshold || torque >= j.torqueThreshold) d.jointEventFlag[j.jointId] = 1;
  }
}

// IntegrateVelocitiesTask src/Solver.cs:121-160 — pure row form; returns the new (v, w, flags) row, ORs IsSpeedCapped into *simFlags
__device__ __forceinline__ float4 integrateVelocityRow(float4 s, float4 m, float4 f, const StepParams& sp, uint32_t* simFlags) {
  V2 v = mk(s.x, s.y); float w = s.z; uint32_t flags = fbits(s.w);
  float h = sp.h;
  float maxAngularSpeed = kMaxRotation * sp.inv_dt;
  float maxLinSq = sp.maxLinearSpeed * sp.maxLinearSpeed, maxAngSq = maxAngularSpeed * maxAngularSpeed;
  float linearDamping = 1.0f / (1.0f + h * m.z), angularDamping = 1.0f / (1.0f + h * m.w);
  float gravityScale = m.x > 0.0f ? f.w : 0.0f;
  V2 delta = ((h * m.x) * mk(f.x, f.y)) + ((h * gravityScale) * sp.gravity);
  float angularDelta = h * m.y * f.z;
  v = mulAdd(delta, linearDamping, v);
  w = angularDelta + angularDamping * w;
  if (dot(v, v) > maxLinSq) { float ratio = sp.maxLinearSpeed / length(v); v = ratio * v; *simFlags |= kIsSpeedCapped; }
  if (w * w > maxAngSq && !(*simFlags & kAllowFastRotation)) { float ratio = maxAngularSpeed / fabsf(w); w *= ratio; *simFlags |= kIsSpeedCapped; }
  if (flags & kLockLinearX) v.x = 0.0f;
  if (flags & kLockLinearY) v.y = 0.0f;
  if (flags & kLockAngularZ) w = 0.0f;
  return make_float4(v.x, v.y, w, s.w);
}
__device__ __forceinline__ void integrateVelocity(const WorldDev& d, const StepParams& sp, int i) {
  uint32_t sf = d.simFlags[i], sf0 = sf;
  d.st0[i] = integrateVelocityRow(d.st0[i], d.simMass[i], d.simForce[i], sp, &sf);
  if (sf != sf0) d.simFlags[i] = sf;
}
// IntegratePositionsTask src/Solver.cs:199-212 — pure row form
__device__ __forceinline__ void integratePositionRow(float4& s, float4& p, const StepParams& sp) {
  uint32_t flags = fbits(s.w);
  if (flags & kLockLinearX) s.x = 0.0f;
  if (flags & kLockLinearY) s.y = 0.0f;
  if (flags & kLockAngularZ) s.z = 0.0f;
  V2 dp = mulAdd(mk(p.x, p.y), sp.h, mk(s.x, s.y));
  Rot dq = integrateRot(mkRot(p.z, p.w), sp.h * s.z);
  p = make_float4(dp.x, dp.y, dq.c, dq.s);
}
__device__ __forceinline__ void integratePosition(const WorldDev& d, const StepParams& sp, int i) {
  float4 s = d.st0[i], p = d.st1[i];
  uint32_t flags = fbits(s.w);
  integratePositionRow(s, p, sp);
  if (flags & (kLockLinearX | kLockLinearY | kLockAngularZ)) d.st0[i] = s;
  d.st1[i] = p;
}

template <class S> __device__ void phase_prepare(const S& grid, const WorldDev& d, const StepParams& sp, bool contactLanes) {
  const Scalars& s = *d.sc;
  int total = contactLanes ? s.smallConStart : 0; // grid-path constraints only (and only when they do not fit one per thread): small islands prepare their own lanes in registers
  for (int t = grid.tid(); t < total; t += grid.size()) if (d.sContact[t] >= 0) prepareContact(d, sp, t);
  int nj = d.jointColorStart[kGraphColorCount];
  for (int k = grid.tid(); k < nj; k += grid.size()) { int id = d.jointSorted[k]; d.jointEventFlag[id] = 0; prepareJoint(d, sp, d.joints[id]); }
  int nov = s.colLen[kOverflowIndex];
  for (int k = grid.tid(); k < nov; k += grid.size()) ovPrepare(d, sp, k);
}

// SolverTask substep loop (src/Solver.cs:599-727) as one persistent cooperative kernel: a grid-wide barrier replaces the
// reference's per-stage spin barrier; empty colors cost nothing.
// CACHED: the grid-path constraints fit one per thread, so each thread builds its lane in registers (PrepareContacts fused in,
// as in the island kernel) and keeps it there for the whole step: a stage then loads only the two body rows after the barrier.
// The stage barrier: the whole cooperative grid (cg grid.sync, ~1.2 us over 148 CTAs) or one thread-block cluster (barrier.cluster,
// ~0.2 us over <= 16 CTAs of one GPC). With a one-cluster grid gtid()/gsize() are the cluster's thread rank and size.

template <bool CACHED, class Scope> __device__ void phase_solver_grid(const Scope& grid, const WorldDev& d, const StepParams& sp) {
  const Scalars& s = *d.sc;
  __shared__ int sBegin[kGraphColorCount], sEnd[kGraphColorCount], sJBegin[kGraphColorCount], sJEnd[kGraphColorCount], sActiveCount;
  if (threadIdx.x == 0) {
    int n = 0;
    for (int c = 0; c < kOverflowIndex; c++) {
      int cc = s.colorCount[c], jc = d.jointColorStart[c + 1] - d.jointColorStart[c];
      if (cc + jc > 0) { sBegin[n] = s.colorStart[c]; sEnd[n] = s.colorStart[c] + cc; sJBegin[n] = d.jointColorStart[c]; sJEnd[n] = d.jointColorStart[c + 1]; n++; }
    }
    sActiveCount = n;
  }
  __syncthreads();
  const int activeCount = sActiveCount;
  const int tid = grid.tid(), stride = grid.size();
  const int awake = s.bigBodyCount;               // bodies of grid-path islands (+ kinematic bodies some constraint reads)
  if (awake == 0 && activeCount == 0) return;     // everything lives in small islands / free bodies: no grid barrier at all
  // overflow colour: one thread, array order, before the colours of each stage (:633-670), its joints before its contacts
  const int nov = s.colLen[kOverflowIndex] + d.jointColorStart[kGraphColorCount] - d.jointColorStart[kOverflowIndex];
  Lane c; float4 I = make_float4(0, 0, 0, 0), J = I; int myA = -1, myContact = -1;
  const BodyRows bs{d.st0, d.st1};
  if (CACHED) {
    if (tid < s.smallConStart) myContact = d.sContact[tid];
    if (myContact >= 0) {
      computeLane(d, sp, myContact, c, I, J);      // reads pre-step velocities: the barrier below keeps IntegrateVelocities behind it
      for (int a = 0; a < activeCount; a++) if (tid >= sBegin[a] && tid < sEnd[a]) myA = a;
    }
    grid.sync();
  }
  for (int sub = 0; sub < sp.subStepCount; sub++) {
    for (int i = tid; i < awake; i += stride) integrateVelocity(d, sp, d.bigBodies[i]);
    grid.sync();
    if (nov > 0) { if (tid == 0) ovWarmStartAll(d); grid.sync(); }
    for (int a = 0; a < activeCount; a++) {
      for (int k = sJBegin[a] + tid; k < sJEnd[a]; k += stride) warmStartJoint(d, d.joints[d.jointSorted[k]]);
      if (CACHED) { if (myA == a) warmStartLane(bs, c, c.indexA, c.indexB, I, J.z); }
      else for (int t = sBegin[a] + tid; t < sEnd[a]; t += stride) warmStartContact(d, t);
      grid.sync();
    }
    if (nov > 0) { if (tid == 0) ovSolveAll(d, sp); grid.sync(); }
    for (int a = 0; a < activeCount; a++) {
      for (int k = sJBegin[a] + tid; k < sJEnd[a]; k += stride) solveJoint(d, sp, d.joints[d.jointSorted[k]], true);
      if (CACHED) { if (myA == a) solveLane<true>(bs, sp, c, c.indexA, c.indexB, I, J); }
      else for (int t = sBegin[a] + tid; t < sEnd[a]; t += stride) solveContact<true>(d, sp, t);
      grid.sync();
    }
    for (int i = tid; i < awake; i += stride) integratePosition(d, sp, d.bigBodies[i]);
    grid.sync();
    if (nov > 0) { if (tid == 0) ovSolveAll(d, sp); grid.sync(); }      // Q4: the relax pass of the overflow colour keeps useBias
    for (int a = 0; a < activeCount; a++) {
      for (int k = sJBegin[a] + tid; k < sJEnd[a]; k += stride) solveJoint(d, sp, d.joints[d.jointSorted[k]], false);
      if (CACHED) { if (myA == a) solveLane<false>(bs, sp, c, c.indexA, c.indexB, I, J); }
      else for (int t = sBegin[a] + tid; t < sEnd[a]; t += stride) solveContact<false>(d, sp, t);
      grid.sync();
    }
  }
  // restitution: skipped entirely (like every AllZeroW group) when no touching contact has restitution (flag set by p_collide)
  if (s.anyRestitution) {
    if (CACHED) { if (myContact >= 0) writeLane(d, tid, c, I, J); grid.sync(); }      // restitution works on 8-lane groups of the planes
    if (nov > 0) { if (tid == 0) ovRestitution(d, sp, s.colLen[kOverflowIndex]); grid.sync(); }
    for (int a = 0; a < activeCount; a++) {
      for (int t = sBegin[a] + tid; t < sEnd[a]; t += stride) restitutionContact(d, sp, t);
      grid.sync();
    }
  }
  if (CACHED && !s.anyRestitution) { if (myContact >= 0) storeImpulsesRegs(d, myContact, I, J, c.relVel1, c.relVel2); }
  else for (int a = 0; a < activeCount; a++)
    for (int t = sBegin[a] + tid; t < sEnd[a]; t += stride) storeImpulses(d, t);
  if (nov > 0) { const int nc = s.colLen[kOverflowIndex]; for (int k = tid; k < nc; k += stride) ovStoreImpulses(d, k); }
}

}  // namespace b2c
#include "b2c_islands.cuh"
#include "b2c_cluster_dsm.cuh"
#include "b2c_toi.cuh"
namespace b2c {

// ------------------------------------------------------------------------------------------------------------
// Phase 3: FinalizeBodiesTask (src/Solver.cs:402-503) + refit (:1191-1237, src/DynamicTree.cs:482-505)
// ------------------------------------------------------------------------------------------------------------
// World.SolveContinuous (src/Solver.cs:315-401): every shape of the fast body sweeps the static tree (a bullet: then the kinematic and the
// dynamic tree, :342-346) with the union of its old and new boxes (DynamicTree.Query order: child2 first), ContinuousQueryCallback
// (:224-314) runs TimeOfImpact on every candidate in that order with the running fraction as the sweep limit, and the smallest positive
// fraction wins. Returns that fraction (1 = no hit). Sensor shapes collect the visitors whose sweep reaches them before the hit (Sensor.hits).
// A query ends at the fast shape's own proxy (the callback returns false, :230): in the dynamic tree a bullet sees only what comes before itself.
constexpr int kContinuousSensorHits = 8;
constexpr int kBulletStageBit = 1 << 30;            // in Scalars::enlargedShapeCount: this step has fast bullets (StepContext.bulletBodyCount > 0, src/Solver.cs:447-449,1238)
__device__ __noinline__ float solveContinuous(const WorldDev& d, int bodyId, Xf xf2, const Sweep& sweep, uint32_t* simFlags, bool isBullet) {
  Scalars& sc = *d.sc;
  float fraction = 1.0f;
  int hitSensor[kContinuousSensorHits], hitVisitor[kContinuousSensorHits]; float hitFraction[kContinuousSensorHits]; int nSensorHits = 0;
  const int treeCount = isBullet ? 3 : 1;              // kStaticBody = 0, kKinematicBody = 1, kDynamicBody = 2: the order of :341-346
  for (int sid = d.bodyHeadShape[bodyId]; sid != -1; sid = d.shapes[sid].nextShapeId) {
    const ShapeDev& fastShape = d.shapes[sid];
    Box box1 = toBox(d.shapeAABB[sid]);
    Box box2 = geomAABB(fastShape.geom, xf2);
    d.shapeAABB[sid] = fromBox(box2);                    // raw AABB, no speculative margin (:337-339)
    if (fastShape.sensorIndex != -1) continue;
    const Box query = boxUnion(box1, box2);
    const DistProxy proxyB = makeDistProxy(fastShape.geom);
    for (int treeIndex = 0; treeIndex < treeCount; treeIndex++) {
    const TreeDev& t = d.trees[treeIndex];
    int root = t.meta[0];
    if (t.meta[1] == 0 || root == -1) continue;
    int stack[kQueryStack]; int sp = 0;
    stack[sp++] = root;
    while (sp > 0) {
      int id = stack[--sp];
      if (!boxOverlaps(toBox(t.aabb[id]), query) || t.cat[id] == 0) continue;      // DEFAULT_MASK_BITS: only category 0 fails the mask test
      if (!((t.hf[id] >> 16) & NODE_LEAF)) {
        if (sp + 2 <= kQueryStack) { int2 c = t.child[id]; stack[sp++] = c.x; stack[sp++] = c.y; } else atomicOr(&sc.errorFlags, (int)ERR_STACK_OVERFLOW);
        continue;
      }
      const int other = t.user[id];
      if (other == sid) break;                           // the callback returns false: this query ends (:230)
      const ShapeDev& shape = d.shapes[other];
      if (shape.bodyId == bodyId) continue;
      const bool isSensor = shape.sensorIndex != -1;
      if (isSensor && (!(shape.flags & SF_SENSOR_EVENTS) || !(fastShape.flags & SF_SENSOR_EVENTS))) continue;
      if (!shouldShapesCollide(fastShape, shape)) continue;
      const int otherSim = d.bodySim[shape.bodyId];
      if ((otherSim < 0 ? d.bodySimFlags[shape.bodyId] : d.simFlags[otherSim]) & kIsBullet) continue;   // bullets do not stop each other (:241)
      if (!shouldBodiesCollide(d, bodyId, shape.bodyId)) continue;
      // (custom filter and chain segments cannot exist: refused at creation)
      Sweep sweepA;                                      // BodySim.MakeSweep src/Body.cs:172
      if (otherSim < 0) {                                // a body at rest in the static set
        const float4 oxf = d.bodyXf[shape.bodyId]; const float2 oc = d.bodyCenter[shape.bodyId]; const float4 om = d.bodyMass[shape.bodyId];
        sweepA.localCenter = mk(om.z, om.w); sweepA.c1 = mk(oc.x, oc.y); sweepA.c2 = sweepA.c1; sweepA.q1 = mkRot(oxf.z, oxf.w); sweepA.q2 = sweepA.q1;
      } else {                                           // an awake body, finalized (and, if fast, stopped) before the bullet stage: center0 == center
        const float4 o0 = d.simPose0[otherSim], o1 = d.simPose[otherSim], ol = d.simLocal[otherSim];
        sweepA.localCenter = mk(ol.x, ol.y); sweepA.c1 = mk(o0.x, o0.y); sweepA.c2 = mk(o1.x, o1.y); sweepA.q1 = mkRot(o0.z, o0.w); sweepA.q2 = mkRot(o1.z, o1.w);
      }
      const DistProxy proxyA = makeDistProxy(shape.geom);
      bool throws = false;
      atomicAdd(&sc.toiCandidates, 1);
      ToiOut output = timeOfImpact(proxyA, proxyB, sweepA, sweep, fraction, &throws);
      if (throws) { atomicOr(&sc.errorFlags, (int)ERR_REFERENCE_THROWS); continue; }
      if (isSensor) {
        if (output.fraction <= fraction) {
          if (nSensorHits < kContinuousSensorHits) { hitSensor[nSensorHits] = other; hitVisitor[nSensorHits] = sid; hitFraction[nSensorHits] = output.fraction; nSensorHits++; }
          else atomicOr(&sc.errorFlags, (int)ERR_SENSOR_CAPACITY);
        }
        continue;
      }
      float hit = fraction; bool didHit = false;
      if (0.0f < output.fraction && output.fraction < fraction) { hit = output.fraction; didHit = true; }
      else if (0.0f == output.fraction) {                // :291-299 fall back to the shape's core: a point of a quarter of the minimum extent at the centroid
        DistProxy core; for (int k = 0; k < kMaxPolygonVertices; k++) core.pts[k] = mk(0, 0);
        core.pts[0] = geomCentroid(fastShape.geom); core.count = 1; core.radius = 0.25f * geomMinExtent(fastShape.geom);
        output = timeOfImpact(proxyA, core, sweepA, sweep, fraction, &throws);
        if (throws) { atomicOr(&sc.errorFlags, (int)ERR_REFERENCE_THROWS); continue; }
        if (0.0f < output.fraction && output.fraction < fraction) { hit = output.fraction; didHit = true; }
      }
      if (didHit) { *simFlags |= kHadTimeOfImpact; fraction = hit; }
    }
    }
  }
  for (int k = 0; k < nSensorHits; k++) {                // :398-400, then src/Solver.cs:1276-1289
    if (!(hitFraction[k] < fraction)) continue;
    const int sensorIndex = d.shapes[hitSensor[k]].sensorIndex;
    int slot = atomicAdd(&d.sensorHitCount[sensorIndex], 1);
    if (slot < kSensorMaxHits) d.sensorHits[(size_t)sensorIndex * kSensorMaxHits + slot] = make_int2(hitVisitor[k], (int)d.shapes[hitVisitor[k]].generation);
    else atomicOr(&sc.errorFlags, (int)ERR_SENSOR_CAPACITY);
  }
  return fraction;
}

// SolveContinuous after the queries (:356-397): a hit stops the body at the time of impact and recomputes its shape boxes; without a hit the
// raw boxes written by the sweep stay. Either way a box that left its fat box grows it (shape flagged, sim gets EnlargeBounds).
// Returns the number of shapes flagged.
__device__ __forceinline__ int continuousOutcome(const WorldDev& d, int i, int bodyId, const Sweep& sweep, float fraction, uint32_t& simFlags, Xf& xf, V2& center, Rot& q) {
  int enlarged = 0;
  if (fraction < 1.0f) {                             // :356-379 the body stops at the time of impact
    q = nlerpRot(sweep.q1, sweep.q2, fraction);
    center = lerp(sweep.c1, sweep.c2, fraction);
    xf = mkXf(center - rotate(q, sweep.localCenter), q);
    d.simPose[i] = make_float4(center.x, center.y, q.c, q.s);
    d.bodyXf[bodyId] = make_float4(xf.p.x, xf.p.y, q.c, q.s);
    d.bodyCenter[bodyId] = make_float2(center.x, center.y);
    d.simPose0[i] = make_float4(center.x, center.y, q.c, q.s);
    for (int sid = d.bodyHeadShape[bodyId]; sid != -1; sid = d.shapes[sid].nextShapeId) {
      Box aabb = boxGrow(geomAABB(d.shapes[sid].geom, xf), kSpeculativeDistance);
      d.shapeAABB[sid] = fromBox(aabb);
      if (!boxContains(toBox(d.shapeFat[sid]), aabb)) {
        d.shapeFat[sid] = fromBox(boxGrow(aabb, kAABBMargin));
        d.shapes[sid].flags |= SF_ENLARGED;
        simFlags |= kEnlargeBounds;
        enlarged++;
      }
    }
  } else {                                           // :380-397
    d.simPose0[i] = make_float4(center.x, center.y, q.c, q.s);
    for (int sid = d.bodyHeadShape[bodyId]; sid != -1; sid = d.shapes[sid].nextShapeId) {
      Box aabb = toBox(d.shapeAABB[sid]);
      if (!boxContains(toBox(d.shapeFat[sid]), aabb)) {
        d.shapeFat[sid] = fromBox(boxGrow(aabb, kAABBMargin));
        d.shapes[sid].flags |= SF_ENLARGED;
        simFlags |= kEnlargeBounds;
        enlarged++;
      }
    }
  }
  return enlarged;
}

// The fast branch of FinalizeBodies (:444-455), out of line so that the common path keeps its registers: a bullet only registers for the
// bullet stage, any other fast body runs SolveContinuous against the static tree here. Returns the shapes to count for the refit.
// Everything goes in and comes back by value: a pointer to a local of p_finalize would pin that local to the stack for the whole common path.
struct FastBodyOut { int enlarged; uint32_t simFlags; Xf xf; V2 center; Rot q; };
__device__ __noinline__ FastBodyOut finalizeFastBody(const WorldDev& d, int i, int bodyId, V2 localCenter, uint32_t simFlags, Xf xf, V2 center, Rot q) {
  int enlarged = 0;
  if (simFlags & kIsBullet) {
    // a fast bullet waits for the bullet stage (:447-449): pose0 and the shape boxes stay, every shape is buffered as a move by the refit (:1210-1218)
    atomicOr(&d.sc->enlargedShapeCount, kBulletStageBit);   // read back with the enlarged-shape count: one load decides both stages
    for (int sid = d.bodyHeadShape[bodyId]; sid != -1; sid = d.shapes[sid].nextShapeId) enlarged++;
    return FastBodyOut{enlarged, simFlags, xf, center, q};
  }
  // SolveContinuous :315-401 (non-bullet body: static tree only)
  float4 pose0 = d.simPose0[i];
  Sweep sweep; sweep.localCenter = localCenter; sweep.c1 = mk(pose0.x, pose0.y); sweep.c2 = center; sweep.q1 = mkRot(pose0.z, pose0.w); sweep.q2 = q;   // BodySim.MakeSweep src/Body.cs:172
  const float fraction = solveContinuous(d, bodyId, xf, sweep, &simFlags, false);
  enlarged = continuousOutcome(d, i, bodyId, sweep, fraction, simFlags, xf, center, q);
  return FastBodyOut{enlarged, simFlags, xf, center, q};
}

__device__ void p_finalize(const WorldDev& d, const StepParams& sp, int tid, int stride) {
  int n = d.awakeCount;
  for (int i = tid; i < n; i += stride) {
    float4 s = d.st0[i], p = d.st1[i], pose = d.simPose[i], loc = d.simLocal[i];
    uint32_t flags = fbits(s.w);
    if (flags & kLockLinearX) s.x = 0.0f;
    if (flags & kLockLinearY) s.y = 0.0f;
    if (flags & kLockAngularZ) s.z = 0.0f;
    V2 v = mk(s.x, s.y); float w = s.z;
    V2 center = mk(pose.x, pose.y) + mk(p.x, p.y);
    Rot q = normalizeRot(mulRot(mkRot(p.z, p.w), mkRot(pose.z, pose.w)));
    float maxVelocity = length(v) + fabsf(w) * loc.w;
    d.st0[i] = s;
    d.st1[i] = make_float4(0.0f, 0.0f, 1.0f, 0.0f);
    V2 localCenter = mk(loc.x, loc.y);
    Xf xf = mkXf(center - rotate(q, localCenter), q);
    int bodyId = d.simBodyId[i];
    d.simPose[i] = make_float4(center.x, center.y, q.c, q.s);
    d.bodyXf[bodyId] = make_float4(xf.p.x, xf.p.y, q.c, q.s);
    d.bodyCenter[bodyId] = make_float2(center.x, center.y);
    float4 f = d.simForce[i];
    d.simForce[i] = make_float4(0.0f, 0.0f, 0.0f, f.w);
    uint32_t simFlags = d.simFlags[i];
    d.bodyFlags[bodyId] = (d.bodyFlags[bodyId] & ~(kIsFast | kIsSpeedCapped | kHadTimeOfImpact)) | (simFlags & (kIsSpeedCapped | kHadTimeOfImpact));
    simFlags &= ~(kIsFast | kIsSpeedCapped | kHadTimeOfImpact);
    bool isFast = false;
    // sleeping is disabled in this build (world creation rejects enableSleep), so the first branch of :444 always runs
    if (d.bodyType[bodyId] == kDynamicBody && sp.enableContinuous && maxVelocity * sp.dt > 0.5f * loc.z) { simFlags |= kIsFast; isFast = true; }
    else d.simPose0[i] = make_float4(center.x, center.y, q.c, q.s);
    int enlarged = 0;
    if (isFast) {
      const FastBodyOut o = finalizeFastBody(d, i, bodyId, localCenter, simFlags, xf, center, q);   // rare: kept out of line
      enlarged = o.enlarged; simFlags = o.simFlags; xf = o.xf; center = o.center; q = o.q;
    } else {
      for (int sid = d.bodyHeadShape[bodyId]; sid != -1; sid = d.shapes[sid].nextShapeId) {
        Box aabb = boxGrow(geomAABB(d.shapes[sid].geom, xf), kSpeculativeDistance);
        d.shapeAABB[sid] = fromBox(aabb);
        if (!boxContains(toBox(d.shapeFat[sid]), aabb)) {
          d.shapeFat[sid] = fromBox(boxGrow(aabb, kAABBMargin));
          d.shapes[sid].flags |= SF_ENLARGED;
          enlarged++;
        }
      }
    }
    d.simFlags[i] = simFlags;
    if (d.moveOut) { B2CBodyMoveEvent ev; ev.transform.p = {xf.p.x, xf.p.y}; ev.transform.q = {q.c, q.s}; ev.bodyId = bodyId; ev.fellAsleep = 0; d.moveOut[i] = ev; }   // src/Solver.cs:432-438
    if (d.bodyOut) {                               // the same per body id, with velocities (bulk BodyAPI getters)
      B2CBodyRead r;
      r.transform.p = {xf.p.x, xf.p.y}; r.transform.q = {q.c, q.s}; r.center = {center.x, center.y};
      r.linearVelocity = {s.x, s.y}; r.angularVelocity = s.z; r.flags = simFlags;
      d.bodyOut[bodyId] = r;
    }
    d.bodyEnlargedCount[i] = enlarged;
    if (enlarged) atomicAdd(&d.sc->enlargedShapeCount, enlarged);
  }
}

__device__ __forceinline__ void atomicMinF(float* addr, float v) {
  v = v + 0.0f;
  if (v >= 0.0f) atomicMin((int*)addr, __float_as_int(v)); else atomicMax((unsigned*)addr, __float_as_uint(v));
}
__device__ __forceinline__ void atomicMaxF(float* addr, float v) {
  v = v + 0.0f;
  if (v >= 0.0f) atomicMax((int*)addr, __float_as_int(v)); else atomicMin((unsigned*)addr, __float_as_uint(v));
}
// DynamicTree.EnlargeProxy (src/DynamicTree.cs:482-505): the leaf takes the fat box, every ancestor grows to contain it and is flagged Enlarged
__device__ __forceinline__ void enlargeLeaf(const TreeDev& t, int id, float4 fat) {
  t.aabb[id] = fat;
  for (int p = t.parent[id]; p != -1; p = t.parent[p]) {
    float* a = (float*)&t.aabb[p];
    const volatile float* av = a;                     // most ancestors already contain the box: test before the atomic
    float4 cur = make_float4(av[0], av[1], av[2], av[3]);
    if (fat.x < cur.x) atomicMinF(a + 0, fat.x);
    if (fat.y < cur.y) atomicMinF(a + 1, fat.y);
    if (fat.z > cur.z) atomicMaxF(a + 2, fat.z);
    if (fat.w > cur.w) atomicMaxF(a + 3, fat.w);
    if (!((*(volatile uint32_t*)&t.hf[p] >> 16) & NODE_ENLARGED)) atomicOr(&t.hf[p], (uint32_t)NODE_ENLARGED << 16);
  }
}
// Enlarged bodies in ascending sim index, shapes in body list order: BroadPhase.EnlargeProxy + BufferMove (:1202-1233).
// EnlargeProxy's result (ancestor boxes = union, every ancestor flagged Enlarged) does not depend on the order of the
// calls, so it is applied with float atomics; the move array order is produced by the prefix sum.
__device__ void p_refit(const WorldDev& d, int tid, int stride) {
  int n = d.awakeCount;
  for (int i = tid; i < n; i += stride) {
    int offset = d.bodyEnlargedCount[i] + d.scanBlockSums[i / (int)blockDim.x];
    int next = (i + 1 < n) ? d.bodyEnlargedCount[i + 1] + d.scanBlockSums[(i + 1) / (int)blockDim.x] : d.sc->moveCount;
    if (next == offset) continue;
    int bodyId = d.simBodyId[i];
    if ((d.simFlags[i] & (kIsBullet | kIsFast)) == (kIsBullet | kIsFast)) {     // :1210-1218 a fast bullet: BufferMove only, the bullet stage enlarges
      for (int sid = d.bodyHeadShape[bodyId]; sid != -1; sid = d.shapes[sid].nextShapeId) {
        int key = d.shapes[sid].proxyKey;
        if (offset < d.moveCapacity) d.moveArray[offset] = key;
        offset++;
        d.trees[key & 3].moved[key >> 2] = 1;
      }
      continue;
    }
    for (int sid = d.bodyHeadShape[bodyId]; sid != -1; sid = d.shapes[sid].nextShapeId) {
      ShapeDev& sh = d.shapes[sid];
      if (!(sh.flags & SF_ENLARGED)) continue;
      sh.flags &= ~SF_ENLARGED;
      int key = sh.proxyKey;
      const TreeDev& t = d.trees[key & 3];
      int id = key >> 2;
      if (offset < d.moveCapacity) d.moveArray[offset] = key;
      offset++;
      t.moved[id] = 1;
      enlargeLeaf(t, id, d.shapeFat[sid]);
    }
  }
}

// BulletBodyTask (src/Solver.cs:728-738): SolveContinuous for this step's fast bullets, after the refit (the dynamic tree holds the enlarged
// boxes of everything else; every other awake body is at its end pose with center0 == center, so the bullets are independent of each other).
__device__ __noinline__ void p_bullets(const WorldDev& d, int tid, int stride) {
  const int n = d.awakeCount;
  for (int i = tid; i < n; i += stride) {
    uint32_t simFlags = d.simFlags[i];
    if ((simFlags & (kIsBullet | kIsFast)) != (kIsBullet | kIsFast)) continue;
    const int bodyId = d.simBodyId[i];
    const float4 pose = d.simPose[i], pose0 = d.simPose0[i], loc = d.simLocal[i];
    V2 center = mk(pose.x, pose.y); Rot q = mkRot(pose.z, pose.w);
    Sweep sweep; sweep.localCenter = mk(loc.x, loc.y); sweep.c1 = mk(pose0.x, pose0.y); sweep.c2 = center; sweep.q1 = mkRot(pose0.z, pose0.w); sweep.q2 = q;
    Xf xf = mkXf(center - rotate(q, sweep.localCenter), q);
    const float fraction = solveContinuous(d, bodyId, xf, sweep, &simFlags, true);
    continuousOutcome(d, i, bodyId, sweep, fraction, simFlags, xf, center, q);
    d.simFlags[i] = simFlags;
    if (fraction < 1.0f && d.moveOut) { d.moveOut[i].transform.p = {xf.p.x, xf.p.y}; d.moveOut[i].transform.q = {q.c, q.s}; }   // :366-367
  }
}
// :1247-1269 the bullets' enlarged proxies grow in the dynamic tree (no BufferMove: the refit put every bullet shape in the move set)
__device__ __noinline__ void p_bullet_enlarge(const WorldDev& d, int tid, int stride) {
  const int n = d.awakeCount;
  for (int i = tid; i < n; i += stride) {
    uint32_t simFlags = d.simFlags[i];
    if ((simFlags & (kIsBullet | kIsFast)) != (kIsBullet | kIsFast)) continue;
    const int bodyId = d.simBodyId[i];
    if (simFlags & kEnlargeBounds) {
      simFlags &= ~(uint32_t)kEnlargeBounds;
      d.simFlags[i] = simFlags;
      const TreeDev& t = d.trees[kDynamicBody];
      for (int sid = d.bodyHeadShape[bodyId]; sid != -1; sid = d.shapes[sid].nextShapeId) {
        ShapeDev& sh = d.shapes[sid];
        if (!(sh.flags & SF_ENLARGED)) continue;
        sh.flags &= ~SF_ENLARGED;
        enlargeLeaf(t, sh.proxyKey >> 2, d.shapeFat[sid]);
      }
    }
    if (d.bodyOut) {                               // the bulk body read-back written by p_finalize, with the pose and flags of the bullet stage
      const float4 pose = d.simPose[i], xf = d.bodyXf[bodyId];
      B2CBodyRead& r = d.bodyOut[bodyId];
      r.transform.p = {xf.x, xf.y}; r.transform.q = {xf.z, xf.w}; r.center = {pose.x, pose.y}; r.flags = simFlags;
    }
  }
}

// contact hit events (src/Solver.cs:1151-1190) in the reference's order: colours 0..23, each colour's contactSims array front to back.
// Position p of the concatenated colour arrays (colList) -> contact slot; FILL=false writes the hit flag of every position (then an
// exclusive prefix sum ranks them), FILL=true writes the event of every flagged position at its rank.
__device__ __forceinline__ bool hitEventOf(const WorldDev& d, const StepParams& sp, int i, B2CContactHitEvent* ev) {
  uint32_t flags = d.cFlags[i];
  if ((flags & (CF_ALIVE | CF_TOUCHING | CF_ENABLE_HIT)) != (CF_ALIVE | CF_TOUCHING | CF_ENABLE_HIT)) return false;
  float4 mn = d.mN[i];
  int pointCount = (int)fbits(mn.w);
  float best = sp.hitEventThreshold; V2 point = mk(0, 0); bool hit = false;
  if (pointCount > 0) { float4 b = d.mB0[i], c = d.mC0[i]; float a = -c.z; if (a > best && b.w > 0.0f) { best = a; point = mk(c.x, c.y); hit = true; } }
  if (pointCount > 1) { float4 b = d.mB1[i], c = d.mC1[i]; float a = -c.z; if (a > best && b.w > 0.0f) { best = a; point = mk(c.x, c.y); hit = true; } }
  if (hit && ev) { int2 sh = d.cShape[i]; B2CContactHitEvent e = {sh.x, sh.y, {point.x, point.y}, {mn.x, mn.y}, best}; *ev = e; }
  return hit;
}
template <bool FILL> __device__ void p_hit_events(const WorldDev& d, const StepParams& sp, int tid, int stride, int* flags, const int* sums) {
  __shared__ int sStart[kGraphColorCount + 1];
  if (threadIdx.x == 0) { int acc = 0; for (int c = 0; c < kGraphColorCount; c++) { sStart[c] = acc; acc += d.sc->colLen[c]; } sStart[kGraphColorCount] = acc; }
  __syncthreads();
  const int total = sStart[kGraphColorCount];
  for (int p = tid; p < total; p += stride) {
    int c = 0;
    while (p >= sStart[c + 1]) c++;
    const int i = d.colList[(size_t)c * d.contactCapacity + (p - sStart[c])];
    if (!FILL) { flags[p] = hitEventOf(d, sp, i, nullptr) ? 1 : 0; continue; }
    const int rank = flags[p] + sums[p / (int)blockDim.x];
    const int next = (p + 1 < total) ? flags[p + 1] + sums[(p + 1) / (int)blockDim.x] : d.sc->hitCount;
    if (next == rank || rank >= d.eventCapacity) continue;
    B2CContactHitEvent ev;
    if (hitEventOf(d, sp, i, &ev)) d.hitEvents[rank] = ev;
  }
}
__device__ void p_step_finish(const WorldDev& d, int newEndIndex) {
  Scalars& s = *d.sc;
  s.graphDirty = 0;
  s.stepCount++;
  if (s.moveCount > d.moveCapacity) { s.moveCount = d.moveCapacity; s.errorFlags |= (int)ERR_PAIR_CAPACITY; }
  s.endCount[newEndIndex] = 0;                     // api/World.cs:106-108: swap, then clear the buffer the next step appends to
  s.sensorEndCount[newEndIndex] = 0;
}
__global__ void k_zero_step(WorldDev d, int newEndIndex) { Scalars& s = *d.sc; s.beginCount = 0; s.hitCount = 0; s.endCount[newEndIndex] = 0; s.sensorBeginCount = 0; s.sensorEndCount[newEndIndex] = 0; }

}  // namespace b2c
#include "b2c_rebuild.cuh"
#include "b2c_bigquery.cuh"
#include "b2c_sensor.cuh"
namespace b2c {

// ------------------------------------------------------------------------------------------------------------
// Fused cooperative phase kernels: one launch per reference phase, grid.sync() between dependent sub-phases, and a
// uniform early-out when the phase has nothing to do (no moved proxies / no contact state change / no enlarged proxy).
// ------------------------------------------------------------------------------------------------------------
template <class S> __device__ void phase_broadphase(const S& grid, const WorldDev& d) {      // UpdateBroadPhasePairs + RebuildTrees
  const int tid = grid.tid(), stride = grid.size();
  Scalars& s = *d.sc;
  if (s.moveCount > 0) {
    if (d.debugChecksums == 2 && tid == 0) d.sc->dbg[6] = globalTimer();
    for (int i = tid; i < d.pairTableSize; i += stride) d.pairTable[i] = kEmptyKey;
    grid.sync();
    p_pair_insert(d, tid, stride);
    grid.sync();
    if (d.debugChecksums == 2 && tid == 0) d.sc->dbg[7] = globalTimer();
    p_find_pairs<false>(d, tid, stride);
    grid.sync();
    if (d.debugChecksums == 2 && tid == 0) { d.sc->dbg[8] = globalTimer(); d.sc->dbg[11] = d.sc->bigQueryCount; }
    phase_big_queries(grid, d);
    if (d.debugChecksums == 2 && tid == 0) d.sc->dbg[9] = globalTimer();
    gridScan<int>(grid, d.moveHitCount, d.scanBlockSums, s.moveCount, &s.newPairCount);
    p_find_pairs<true>(d, tid, stride);
    p_big_fill(d, tid, stride);
    grid.sync();
    if (d.debugChecksums == 2 && tid == 0) d.sc->dbg[10] = globalTimer();
    p_create_contacts(d, tid, stride);
    p_clear_moved(d, tid, stride);
    grid.sync();
  } else if (tid == 0) s.newPairCount = 0;
  if (tid == 0) p_pairs_finish(d);
}

template <class S> __device__ void phase_collide(const S& grid, const WorldDev& d, const StepParams& sp) {   // World.Collide
  const int tid = grid.tid(), stride = grid.size();
  Scalars& s = *d.sc;
  const bool probe = d.debugChecksums == 2 && tid == 0;   // B2CUDA_DEBUG_CHECKSUMS=2: %globaltimer stamps of the sub-phases in dbg[0..5]
  if (probe) s.dbg[0] = globalTimer();
  p_collide(d, sp, tid, stride);
  grid.sync();
  if (probe) s.dbg[1] = globalTimer();
  if (s.anyContactFlag == 0) {                    // no begin / end / destroy anywhere: the serial pass is empty
    if (tid == 0) { s.beginCount = 0; s.opCount = 0; }      // endCount keeps the events DestroyContact appended since the last step
    return;
  }
  gridScan<int4>(grid, d.flagScan, d.scanBlockSums4, s.contactNext, d.flagScan + d.contactCapacity);
  p_contact_events(d, sp, tid, stride);
  grid.sync();
  if (probe) s.dbg[2] = globalTimer();
  if (tid == 0) { p_events_finish(d, sp, d.flagScan + d.contactCapacity); s.anyContactFlag = 0; }
  grid.sync();
  if (grid.block() == 0) p_color_ops(d);
  if (probe) { s.dbg[5] = globalTimer(); s.dbg[12] = (unsigned long long)s.opCount; s.dbg[13] = (unsigned long long)s.colorRounds; }
}


// ------------------------------------------------------------------------------------------------------------
// Debug aid (B2CUDA_DEBUG_CHECKSUMS=1): order-independent checksums of the state between phases, so that two runs of
// the same scene can be compared phase by phase (scripts/gpu_determinism.py). Off by default: one uniform branch.
// ------------------------------------------------------------------------------------------------------------
__device__ __forceinline__ unsigned long long mix64(unsigned long long x) { x ^= x >> 33; x *= 0xff51afd7ed558ccdULL; x ^= x >> 33; x *= 0xc4ceb9fe1a85ec53ULL; x ^= x >> 33; return x; }
__device__ __forceinline__ unsigned long long hf4(unsigned long long h, float4 v) {
  h = mix64(h + (((unsigned long long)fbits(v.x) << 32) | fbits(v.y)));
  return mix64(h + (((unsigned long long)fbits(v.z) << 32) | fbits(v.w)));
}
__device__ void dbgAdd(const WorldDev& d, int slot, unsigned long long sum) {
  for (int o = 16; o > 0; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
  if ((threadIdx.x & 31) == 0 && sum) atomicAdd(&d.sc->dbg[slot], sum);
}
__device__ void dbgBodies(const WorldDev& d, int slot) {
  if (d.debugChecksums != 1) return;
  unsigned long long sum = 0;
  for (int i = gtid(); i < d.awakeCount; i += gsize()) sum += hf4(hf4(mix64(i + 1), d.st0[i]), d.st1[i]);
  dbgAdd(d, slot, sum);
}
__device__ void dbgContacts(const WorldDev& d, int slot) {
  if (d.debugChecksums != 1) return;
  unsigned long long sum = 0;
  int n = d.sc->contactNext;
  for (int i = gtid(); i < n; i += gsize()) {
    if (!(d.cFlags[i] & CF_ALIVE)) continue;
    unsigned long long h = mix64(i + 1);
    int2 sh = d.cShape[i], sm = d.cSim[i];
    h = mix64(h + d.cFlags[i]); h = mix64(h + (unsigned)d.cColor[i]); h = mix64(h + (((unsigned long long)(unsigned)sh.x << 32) | (unsigned)sh.y));
    h = mix64(h + (((unsigned long long)(unsigned)sm.x << 32) | (unsigned)sm.y));
    h = hf4(h, d.cMass[i]); h = hf4(h, d.cMat[i]); h = hf4(h, d.mN[i]); h = hf4(h, d.mA0[i]); h = hf4(h, d.mB0[i]); h = hf4(h, d.mC0[i]);
    if (fbits(d.mN[i].w) > 1) { h = hf4(h, d.mA1[i]); h = hf4(h, d.mB1[i]); h = hf4(h, d.mC1[i]); }
    sum += h;
  }
  dbgAdd(d, slot, sum);
}
__device__ void dbgLanes(const WorldDev& d, int slot, bool all) {
  if (d.debugChecksums != 1) return;
  unsigned long long sum = 0;
  int n = d.sc->smallConStart;
  for (int t = gtid(); t < n; t += gsize()) {
    int ci = d.sContact[t];
    if (ci < 0) continue;
    unsigned long long h = mix64(ci + 1);
    if (all) { h = hf4(h, d.sB[t]); h = hf4(h, d.sC[t]); h = hf4(h, d.sD[t]); h = hf4(h, d.sE[t]); h = hf4(h, d.sF[t]); h = hf4(h, d.sG[t]); h = hf4(h, d.sH[t]);
               float4 A = d.sA[t]; h = hf4(h, make_float4(0, 0, A.z, A.w)); }
    h = hf4(h, d.sI[t]); h = hf4(h, d.sJ[t]);
    sum += h;
  }
  dbgAdd(d, slot, sum);
}

// debug: number of (color, dynamic body) pairs claimed by more than one touching constraint (must be 0), and the same
// test on the solver's own arrangement (slot -> contact -> bodies)
template <class S> __device__ void dbgColorConflicts(const S& grid, const WorldDev& d, int slot) {
  if (d.debugChecksums != 1) return;
  const int awake = d.awakeCount, n = d.sc->contactNext;
  grid.sync();
  for (int i = grid.tid(); i < awake * kGraphColorCount; i += grid.size()) d.rootColorCount[i] = 0;
  grid.sync();
  unsigned long long bad = 0;
  for (int i = grid.tid(); i < n; i += grid.size()) {
    if (!(d.cFlags[i] & CF_TOUCHING)) continue;
    int c = d.cColor[i];
    if (c == kOverflowIndex) continue;             // solved serially, shares bodies by design
    if (c < 0 || c > kOverflowIndex) { bad += 1000000; continue; }
    if (d.cConstraint[i] < 0 || d.sContact[d.cConstraint[i]] != i) bad += 1000;
    int2 sim = d.cSim[i];
    if (sim.x >= 0 && (fbits(d.st0[sim.x].w) & kDynamicFlag) && atomicAdd(&d.rootColorCount[sim.x * kGraphColorCount + c], 1) > 0) bad++;
    if (sim.y >= 0 && (fbits(d.st0[sim.y].w) & kDynamicFlag) && atomicAdd(&d.rootColorCount[sim.y * kGraphColorCount + c], 1) > 0) bad++;
  }
  dbgAdd(d, slot, bad);
  // partition: both dynamic bodies of a touching contact share a root; roots are fixed points; a small island's constraint
  // indexes its own bodies; a body with a constraint is never on the free list
  unsigned long long bad2 = 0;
  for (int i = grid.tid(); i < awake; i += grid.size()) { int r = d.ufParent[i]; if (r < 0 || r >= awake || d.ufParent[r] != r) bad2 += 1; }
  for (int i = grid.tid(); i < n; i += grid.size()) {
    if (!(d.cFlags[i] & CF_TOUCHING)) continue;
    int2 sim = d.cSim[i];
    bool dynA = sim.x >= 0 && (fbits(d.st0[sim.x].w) & kDynamicFlag), dynB = sim.y >= 0 && (fbits(d.st0[sim.y].w) & kDynamicFlag);
    if (dynA && dynB && d.ufParent[sim.x] != d.ufParent[sim.y]) bad2 += 1000;
    int r = d.ufParent[dynA ? sim.x : sim.y], info = d.rootInfo[r];
    if (info == -1) bad2 += 1000000;
    int t = d.cConstraint[i];
    if (info >= 0 && t >= 0) {
      int2 lc = d.sLocal[t];
      int b0 = d.islandBodyStart[info], nb = d.islandBodyStart[info + 1] - b0;
      if (dynA && (lc.x < 0 || lc.x >= nb || d.islandBodies[b0 + lc.x] != sim.x)) bad2 += 1000000000ull;
      if (dynB && (lc.y < 0 || lc.y >= nb || d.islandBodies[b0 + lc.y] != sim.y)) bad2 += 1000000000ull;
      if (t < d.islandConStart[info] || t >= d.islandConStart[info] + d.islandColorOff[info * (kGraphColorCount + 1) + kGraphColorCount]) bad2 += 1000000000000ull;
    }
  }
  dbgAdd(d, slot + 1, bad2);
  grid.sync();
}

__device__ __forceinline__ void stamp(const WorldDev& d, int k) { if (gtid() == 0) d.sc->phaseStamp[k] = globalTimer(); }

// UpdateBroadPhasePairs -> Collide -> (islands) -> PrepareContacts -> grid-path solver. Phases that have nothing to do leave through a
// uniform branch; barriers after phases that wrote nothing are skipped. One launch at grid scope (k_step_front), or the first part of a
// batch world's step at CTA scope (k_batch_step).
template <class S> __device__ void stepFront(const S& grid, const WorldDev& d, const StepParams& sp) {
  Scalars& s = *d.sc;
  if (grid.tid() == 0) s.phaseStamp[0] = globalTimer();
  if (grid.tid() < 16) s.dbg[grid.tid()] = 0;
  phase_broadphase(grid, d);
  grid.sync();
  if (grid.tid() == 0) s.phaseStamp[1] = globalTimer();
  dbgContacts(d, 0); dbgBodies(d, 1);
  phase_rebuild(grid, d, kDynamicBody);           // BroadPhase.RebuildTrees (src/BroadPhase.cs:112-116), part of Collide in the reference
  phase_rebuild(grid, d, kKinematicBody);
  if (grid.tid() == 0) s.phaseStamp[5] = globalTimer();
  phase_collide(grid, d, sp);
  grid.sync();
  if (s.anyRestitution) {                         // rare: which 8-lane groups of the colour arrays ApplyRestitution will not skip
    for (int i = grid.tid(); i < s.contactNext; i += grid.size()) d.cRestGroup[i] = 0;
    grid.sync();
    p_rest_groups(d, grid.tid(), grid.size());
    grid.sync();
  }
  if (grid.tid() == 0) s.phaseStamp[2] = globalTimer();
  dbgContacts(d, 2);
  const bool dirty = s.graphDirty != 0;
  phase_islands(grid, d);
  if (dirty) grid.sync();
  if (d.debugChecksums == 3 && grid.tid() == 0) s.dbg[12] = globalTimer();
  const bool gridWork = s.smallConStart > 0 || d.jointColorStart[kGraphColorCount] > 0 || s.bigBodyCount > 0;
  // Scenes whose grid-path work is one cluster's worth are solved by k_solver_cluster (launched next when sp.clusterThreads > 0):
  // every colour stage is one barrier + one L2 round trip, so the cheaper barrier wins although 16 SMs do the work of 148.
  const int gridJoints = d.jointColorStart[kGraphColorCount];
  // Measured (B200): the cluster wins where stages are many and thin (joints, an active overflow colour) or the lanes fit its registers;
  // a contact-only scene between 8 192 and 75 776 constraints keeps its lanes in the grid's registers, which beats the cheaper barrier.
  const bool clusterFits = s.smallConStart + gridJoints <= 12 * sp.clusterThreads && s.bigBodyCount <= 8 * sp.clusterThreads;
  const bool clusterWins = gridJoints > 0 || s.colLen[kOverflowIndex] > 0 || s.smallConStart <= sp.clusterThreads || s.smallConStart > grid.size();
  // One large island of contacts that fits the cluster's shared memory: body rows in distributed shared memory, lanes in shared memory,
  // no L2 round trip in a stage (b2c_cluster_dsm.cuh) - beats both the grid barrier and the plain cluster path.
  const bool dsm = gridWork && sp.clusterThreads > 0 && dsmEligible(d, sp);
  const bool clusterPath = gridWork && sp.clusterThreads > 0 && ((clusterFits && clusterWins) || dsm);
  if (grid.tid() == 0) s.clusterSolve = dsm ? 2 : (clusterPath ? 1 : 0);
  const bool cachedLanes = dsm || s.smallConStart <= (clusterPath ? sp.clusterThreads : grid.size());   // one grid-path constraint per thread: lanes stay in registers (dsm: built by the cluster kernel)
  if (gridWork) { phase_prepare(grid, d, sp, !cachedLanes); if (!cachedLanes) grid.sync(); }
  if (grid.tid() == 0) s.phaseStamp[3] = globalTimer();
  dbgLanes(d, 3, true); dbgBodies(d, 4); dbgColorConflicts(grid, d, 10);
  if (d.debugChecksums == 1) grid.sync();
  if (gridWork && !clusterPath) { if (cachedLanes) phase_solver_grid<true>(grid, d, sp); else phase_solver_grid<false>(grid, d, sp); }
  if (grid.tid() == 0) s.phaseStamp[4] = globalTimer();
  if (d.debugChecksums == 1) { grid.sync(); dbgBodies(d, 5); dbgLanes(d, 6, false); dbgContacts(d, 7); }
  if (grid.tid() == 0) { for (int k = 0; k < 4; k++) s.phaseAcc[k] += s.phaseStamp[k + 1] - s.phaseStamp[k]; s.phaseAcc[4] += s.phaseStamp[5] - s.phaseStamp[1]; }
}
__global__ void __launch_bounds__(kCoopThreads) k_step_front(const __grid_constant__ WorldDev d, const __grid_constant__ StepParams sp) {
  cg::grid_group g = cg::this_grid();
  GridScope grid{g};
  stepFront(grid, d, sp);
}

// The grid-path solver on one thread-block cluster (see clusterPath in stepFront). All CTAs of the cluster take the same path.
__global__ void __launch_bounds__(kCoopThreads) k_solver_cluster(const __grid_constant__ WorldDev d, const __grid_constant__ StepParams sp) {
  extern __shared__ __align__(16) unsigned char dsmSmem[];
  Scalars& s = *d.sc;
  if (!s.clusterSolve) return;
  ClusterScope cs;
  if (s.clusterSolve == 2) {
    if (phase_solver_dsm(cs, d, sp, dsmSmem)) return;
    // one CTA would own more constraints than it has lanes: the plain cluster path, lanes in the global planes (stepFront skipped PrepareContacts)
    for (int t = cs.tid(); t < s.smallConStart; t += cs.size()) if (d.sContact[t] >= 0) prepareContact(d, sp, t);
    cs.sync();
    phase_solver_grid<false>(cs, d, sp);
    return;
  }
  if (s.smallConStart <= cs.size()) phase_solver_grid<true>(cs, d, sp); else phase_solver_grid<false>(cs, d, sp);
}

template <class S> __device__ void finalizePhase(const S& grid, const WorldDev& d, const StepParams& sp, int newEndIndex) {   // FinalizeBodies + refit
  const int tid = grid.tid(), stride = grid.size();
  Scalars& s = *d.sc;
  unsigned long long t0 = 0;
  if (tid == 0) t0 = globalTimer();
  if (d.debugChecksums == 1) { dbgBodies(d, 8); dbgContacts(d, 9); grid.sync(); }
  p_finalize(d, sp, tid, stride);
  if (s.anyHitShapes) {                           // hit events, ordered by (colour, position in the colour's array)
    int* hitFlags = (int*)d.flagScan; int* hitSums = (int*)d.scanBlockSums4;
    int total = 0; for (int c = 0; c < kGraphColorCount; c++) total += s.colLen[c];
    p_hit_events<false>(d, sp, tid, stride, hitFlags, hitSums);
    grid.sync();
    gridScan<int>(grid, hitFlags, hitSums, total, &s.hitCount);
    p_hit_events<true>(d, sp, tid, stride, hitFlags, hitSums);
  }
  grid.sync();
  const int enlargedShapes = s.enlargedShapeCount;
  if (enlargedShapes > 0) {
    gridScan<int>(grid, d.bodyEnlargedCount, d.scanBlockSums, d.awakeCount, &s.moveCount);
    p_refit(d, tid, stride);
    grid.sync();
    if (enlargedShapes & kBulletStageBit) {         // bullet stage :1238-1272 (uniform: flagged by p_finalize before the barriers above; a fast
      p_bullets(d, tid, stride);                    // bullet counts its shapes as enlarged, so the test lives in this branch and a step
      grid.sync();                                  // without moves pays nothing for it)
      p_bullet_enlarge(d, tid, stride);
      grid.sync();
    }
  } else if (tid == 0) s.moveCount = 0;
  phase_sensors(grid, d, sp.endIndex);              // World.OverlapSensors, api/World.cs:97-102 (before the end-event buffers trade places)
  if (tid < 32) {                                   // the first warp of the scope
    if (tid == 0) {
      s.enlargedShapeCount = 0; p_step_finish(d, newEndIndex);
      const unsigned long long now = globalTimer();   // kernel-level times without CUDA events in the stream (each record costs ~2.5 us of device time)
      s.phaseAcc[5] += t0 - s.phaseStamp[4]; s.phaseAcc[6] += now - t0; s.phaseAcc[7] += now - s.phaseStamp[0];
    }
    if (d.hostSc) {                                 // the step's scalars go to the host's pinned mirror from here: no copy node behind the kernel
      __syncwarp();
      const int4* src = reinterpret_cast<const int4*>(&s); int4* dst = reinterpret_cast<int4*>(d.hostSc);
      for (int k = tid; k < (int)(sizeof(Scalars) / sizeof(int4)); k += 32) dst[k] = __ldcg(src + k);
    }
  }
}
__global__ void __launch_bounds__(kCoopThreads) k_finalize_phase(const __grid_constant__ WorldDev d, const __grid_constant__ StepParams sp, int newEndIndex) {
  cg::grid_group g = cg::this_grid();
  GridScope grid{g};
  finalizePhase(grid, d, sp, newEndIndex);
}

// ------------------------------------------------------------------------------------------------------------
// Batches of independent worlds (api/World.cs:9-10 "Each world is completely independent and may be simulated in parallel"):
// ONE CTA steps ONE world end to end — the same phases as above at CTA scope (every barrier is a __syncthreads(), no grid-wide
// synchronisation anywhere), then its islands with the CTA island solver, then FinalizeBodies/refit. World index = blockIdx.x; each
// world has its own arrays, id pools, trees and colour graph, so world k of a batch is bit-identical to the same world stepped alone.
// ------------------------------------------------------------------------------------------------------------
// MODE 0: the whole step in one launch. MODE 1 / 2: the step split in three launches — front (this kernel, MODE 1: pairs, collide, colouring,
// islands, the islands that do NOT fit a warp slice, free bodies), k_batch_solve_warp (one warp per world, eight worlds per SM, for the
// islands that fit a slice) and finalize (this kernel, MODE 2). The split gives each part the residency it can use: the solver's stage
// chain keeps one warp busy per world, so eight warp-worlds per SM beat four CTA-worlds whose other warps idle at the barrier.
template <int THREADS, int MINB, int MODE>
__global__ void __launch_bounds__(THREADS, MINB) k_batch_step(const WorldDev* __restrict__ worlds, StepParams sp, int newEndIndex, int* __restrict__ maxIslandCons) {
  __shared__ __align__(16) unsigned char sdRaw[sizeof(WorldDev)];
  __shared__ float4 s0[MODE == 2 ? 1 : kIslandMaxBodies], s1[MODE == 2 ? 1 : kIslandMaxBodies];
  __shared__ int colorOff[kGraphColorCount + 1 + 2 * kGraphColorCount];
  {
    const int* src = reinterpret_cast<const int*>(worlds + blockIdx.x);
    int* dst = reinterpret_cast<int*>(sdRaw);
    for (int i = threadIdx.x; i < (int)(sizeof(WorldDev) / sizeof(int)); i += blockDim.x) dst[i] = src[i];
  }
  __syncthreads();
  const WorldDev& d = *reinterpret_cast<const WorldDev*>(sdRaw);
  CtaScope cta;
  if (MODE != 2) {
    stepFront(cta, d, sp);                        // sp.clusterThreads == 0: large / jointed islands are solved here, at CTA scope
    __syncthreads();
    const int n = d.sc->islandCount;
    for (int g = 0; g < n; g++) {
      int K = d.islandColorOff[g * (kGraphColorCount + 1) + kGraphColorCount];
      if (threadIdx.x == 0 && K > *maxIslandCons) atomicMax(maxIslandCons, K);   // the host sizes the next launches' CTAs by it (lanes stay in registers while K <= blockDim)
      if (MODE == 1 && warpIslandEligible(d, g)) continue;                      // k_batch_solve_warp takes it
      if (K <= (int)blockDim.x) runIsland<true>(d, sp, g, s0, s1, colorOff);
      else runIsland<false>(d, sp, g, s0, s1, colorOff);
    }
    freeBodies(d, sp, threadIdx.x, blockDim.x);
    __syncthreads();
  }
  if (MODE != 1) {
    finalizePhase(cta, d, sp, newEndIndex);       // (stamps the island-solver and finalize times itself)
  }
}

// bodyMoveEvents equivalent: one packed record per body id (src/Solver.cs:432-438)
__global__ void __launch_bounds__(kThreads) k_pack_bodies(WorldDev d, B2CBodyRead* out) {
  for (int i = gtid(); i < d.bodyCapacity; i += gsize()) {
    float4 xf = d.bodyXf[i]; float2 c = d.bodyCenter[i];
    B2CBodyRead r;
    r.transform.p = {xf.x, xf.y}; r.transform.q = {xf.z, xf.w}; r.center = {c.x, c.y};
    int sim = d.bodySim[i];
    if (sim >= 0) { float4 s = d.st0[sim]; r.linearVelocity = {s.x, s.y}; r.angularVelocity = s.z; r.flags = d.simFlags[sim]; }
    else { r.linearVelocity = {0.0f, 0.0f}; r.angularVelocity = 0.0f; r.flags = d.bodySimFlags[i]; }
    out[i] = r;
  }
}

// ------------------------------------------------------------------------------------------------------------
// host driver
// ------------------------------------------------------------------------------------------------------------
static int gridFor(int n, int threads, int maxBlocks) {
  int b = (n + threads - 1) / threads;
  if (b < 1) b = 1;
  return b > maxBlocks ? maxBlocks : b;
}

// The launches of one step on the world's stream; captured into a CUDA graph (external = event-record nodes) or issued directly.
int World::enqueueStep(const StepParams& spIn, bool warpMode, bool clusterLaunch, int newEndIndex) {
  WorldDev& d = d_;
  StepParams sp = spIn;
  const dim3 cgrid(coopBlocks_), cblock(kCoopThreads);
  { void* args[] = {(void*)&d, (void*)&sp}; CK(cudaLaunchCooperativeKernel((void*)k_step_front, cgrid, cblock, args, 0, stream_)); }
  if (clusterLaunch) {
    cudaLaunchConfig_t cfg = {}; cfg.gridDim = dim3(clusterCtas_); cfg.blockDim = dim3(kCoopThreads); cfg.stream = stream_;
    cfg.dynamicSmemBytes = sp.clusterDsm ? kDsmSmemBytes : 0;
    cudaLaunchAttribute at; at.id = cudaLaunchAttributeClusterDimension; at.val.clusterDim.x = clusterCtas_; at.val.clusterDim.y = 1; at.val.clusterDim.z = 1;
    cfg.attrs = &at; cfg.numAttrs = 1;
    CK(cudaLaunchKernelEx(&cfg, k_solver_cluster, d, sp));
  }
  if (warpMode) k_solver_islands_warp<<<coopBlocks_, 32 * kWarpIslandsPerCta, kWarpIslandsPerCta * sizeof(WarpIslandSlice), stream_>>>(d, sp);
  k_solver_islands<<<coopBlocks_ * 8, kIslandThreads, 0, stream_>>>(d, sp, (int)warpMode);
  { void* args[] = {(void*)&d, (void*)&sp, (void*)&newEndIndex}; CK(cudaLaunchCooperativeKernel((void*)k_finalize_phase, cgrid, cblock, args, 0, stream_)); }
  if (!d.hostSc) CK(cudaMemcpyAsync(hostScalars_, d.sc, sizeof(Scalars), cudaMemcpyDeviceToHost, stream_));   // (mapped mirror unavailable)
  if (streamMoves_ && !hMoveEventsDev_) CK(cudaMemcpyAsync(hMoveEvents_, dMoveEvents_, sizeof(B2CBodyMoveEvent) * (size_t)d.awakeCount, cudaMemcpyDeviceToHost, stream_));
  if (streamBodies_) CK(cudaMemcpyAsync(hBodyRead_, dBodyRead_, sizeof(B2CBodyRead) * (size_t)d.bodyCapacity, cudaMemcpyDeviceToHost, stream_));   // the caller reads body records every step: the copy rides behind the step
  return B2C_OK;
}
int World::step(float timeStep, int subStepCount) {
  if (batch_) return fail(B2C_E_INVALID, "this world belongs to a batch: step it with b2cuda_batch_step");
  if (enableSleep) return fail(B2C_E_UNSUPPORTED, "enableSleep=true (islands/sleep, src/Island.cs) is not supported in this build; create the world with enableSleep=0");
  if (!onDevice_) { int rc = upload(); if (rc) return rc; }
  CK(cudaSetDevice(device_));
  WorldDev& d = d_;
  eventsFetched_ = false; bodiesFetched_ = false; movesFetched_ = false;
  lastStepDtZero_ = timeStep == 0.0f;              // api/World.cs:38,44-51: the move events are cleared and the step returns early
  if (timeStep == 0.0f) {                          // api/World.cs:44-51
    lastEndIndex_ = endIndex_;
    endIndex_ = 1 - endIndex_;
    k_zero_step<<<1, 1, 0, stream_>>>(d, endIndex_);
    CK(cudaMemcpyAsync(hostScalars_, d.sc, sizeof(Scalars), cudaMemcpyDeviceToHost, stream_));
    return B2C_OK;
  }
  { // contact rows three quarters full at the last step that reached the host: grow them before they overflow (src: List<T> growth)
    const Scalars& hs = *hostScalars_;
    const int used = std::max(std::max(hs.contactNext, hs.newPairCount), std::max(std::max(hs.beginCount, hs.hitCount), std::max(hs.endCount[0], hs.endCount[1])));
    if (4LL * used > 3LL * d.contactCapacity) { int rc = growContacts(2 * d.contactCapacity); if (rc) return rc; }
  }
  StepParams sp;
  memset(&sp, 0, sizeof(sp));
  sp.dt = timeStep; sp.subStepCount = subStepCount > 1 ? subStepCount : 1;
  if (timeStep > 0.0f) { sp.inv_dt = 1.0f / timeStep; sp.h = timeStep / sp.subStepCount; sp.inv_h = sp.subStepCount * sp.inv_dt; }
  float hz = netMin(contactHertz, 0.125f * sp.inv_h);           // api/World.cs:63-83
  sp.contactSoftness = makeSoft(hz, contactDampingRatio, sp.h);
  sp.staticSoftness = makeSoft(2.0f * hz, contactDampingRatio, sp.h);
  sp.restitutionThreshold = restitutionThreshold; sp.maxLinearSpeed = maxLinearSpeed; sp.contactSpeed = contactSpeed; sp.contactHertz = contactHertz;
  sp.contactDampingRatio = contactDampingRatio; sp.hitEventThreshold = hitEventThreshold; sp.gravity = gravity;
  sp.enableWarmStarting = enableWarmStarting; sp.enableContactSoftening = enableContactSoftening; sp.enableContinuous = enableContinuous;
  sp.enableSpeculative = enableSpeculative; sp.enableSleep = 0; sp.endIndex = endIndex_;

  if (coopBlocks_ == 0) {
    int occ[2] = {0, 0};
    CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ[0], k_step_front, kCoopThreads, 0));
    CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ[1], k_finalize_phase, kCoopThreads, 0));
    for (int o : occ) if (o < 1) return fail(B2C_E_CUDA, "a cooperative phase kernel cannot be made resident");
    cudaDeviceProp prop; CK(cudaGetDeviceProperties(&prop, device_));
    coopBlocks_ = prop.multiProcessorCount;       // one 512-thread CTA per SM: the barrier cost grows with CTA count
    CK(cudaFuncSetAttribute(k_solver_islands_warp, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(kWarpIslandsPerCta * sizeof(WarpIslandSlice))));
    if (const char* e = getenv("B2CUDA_ISLAND_MODE")) islandMode_ = !strcmp(e, "cta") ? 1 : !strcmp(e, "warp") ? 2 : 0;   // tests force a variant
    if (coopBlocks_ < 3) return fail(B2C_E_CUDA, "device too small");
    // thread-block cluster for the grid-path solver: 16 CTAs (non-portable size) when the device can co-schedule them, else 8
    clusterCtas_ = 0;
    const char* ce = getenv("B2CUDA_CLUSTER");
    if (!ce || strcmp(ce, "0")) {
      for (int size : {16, 8}) {
        if (size > 8 && cudaFuncSetAttribute(k_solver_cluster, cudaFuncAttributeNonPortableClusterSizeAllowed, 1) != cudaSuccess) { cudaGetLastError(); continue; }
        cudaLaunchConfig_t cfg = {}; cfg.gridDim = dim3(size); cfg.blockDim = dim3(kCoopThreads); cfg.stream = stream_;
        cudaLaunchAttribute at; at.id = cudaLaunchAttributeClusterDimension; at.val.clusterDim.x = size; at.val.clusterDim.y = 1; at.val.clusterDim.z = 1;
        cfg.attrs = &at; cfg.numAttrs = 1;
        int n = 0;
        // first choice: the cluster with the shared memory of the distributed-shared-memory solver (192 KB per CTA)
        const char* de = getenv("B2CUDA_CLUSTER_DSM");
        clusterDsmFallback_ = de && !strcmp(de, "fallback");    // tests: every eligible step takes the solver's way out (a CTA short of lanes)
        if ((!de || strcmp(de, "0")) && cudaFuncSetAttribute(k_solver_cluster, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kDsmSmemBytes) == cudaSuccess) {
          cfg.dynamicSmemBytes = kDsmSmemBytes;
          cudaError_t oe = cudaOccupancyMaxActiveClusters(&n, k_solver_cluster, &cfg);
          if (getenv("B2CUDA_VERBOSE")) fprintf(stderr, "b2cuda: cluster %d CTAs with %zu B of shared memory: %s, %d active clusters\n", size, kDsmSmemBytes, cudaGetErrorString(oe), n);
          if (oe == cudaSuccess && n >= 1) { clusterCtas_ = size; clusterDsm_ = true; break; }
          cfg.dynamicSmemBytes = 0;
        } else if (getenv("B2CUDA_VERBOSE")) fprintf(stderr, "b2cuda: cluster kernel cannot take %zu B of dynamic shared memory: %s\n", kDsmSmemBytes, cudaGetErrorString(cudaGetLastError()));
        cudaGetLastError();
        if (cudaOccupancyMaxActiveClusters(&n, k_solver_cluster, &cfg) == cudaSuccess && n >= 1) { clusterCtas_ = size; break; }
        cudaGetLastError();
      }
    }
  }
  // few islands: one CTA each (shortest stage chain); many (batched worlds): one warp each (b2c_islands_warp.cuh) - measured
  // break-even ~5 islands per SM (scripts/gpu_islandmode.sh). The count comes from the last step whose scalars reached the host;
  // either variant is correct for any count.
  const bool warpMode = islandMode_ == 2 || (islandMode_ == 0 && hostScalars_->islandCount >= 5 * coopBlocks_);
  // the cluster kernel is launched only when the last step that reached the host had grid-path work (large or jointed islands);
  // the front kernel takes the cluster path only when told the kernel follows, so a stale view costs speed, never correctness
  const bool clusterLaunch = clusterCtas_ > 0 && !d.debugChecksums && (hostScalars_->smallConStart > 0 || hostScalars_->bigBodyCount > 0 || d.jointColorStart[kGraphColorCount] > 0);
  sp.clusterThreads = clusterLaunch ? clusterCtas_ * kCoopThreads : 0;
  // The shared-memory solver's 216 KB leave the cluster kernel almost no L1, which the plain cluster path (joint records, lanes in global
  // planes) lives on: ask for them only when the last step that reached the host looked eligible (the device decides for itself whether it is).
  { const Scalars& hs = *hostScalars_;
    sp.clusterDsm = clusterLaunch && clusterDsm_ && d.jointColorStart[kGraphColorCount] == 0 && hs.colLen[kOverflowIndex] == 0 && hs.anyRestitution == 0 &&
                    hs.smallConStart > 0 && hs.smallConStart <= kDsmLanes * clusterCtas_ * kCoopThreads && hs.bigBodyCount <= kDsmRows * clusterCtas_ ? (clusterDsmFallback_ ? 2 : 1) : 0; }
  if (stepCount_ == 100 && getenv("B2CUDA_VERBOSE")) fprintf(stderr, "b2cuda: step 100: cluster %d CTAs, dsm %d, launch %d; last step on the host: clusterSolve %d, grid-path constraints %d, bodies %d, islands %d, overflow %d, restitution %d\n",
      clusterCtas_, (int)clusterDsm_, (int)clusterLaunch, hostScalars_->clusterSolve, hostScalars_->smallConStart, hostScalars_->bigBodyCount, hostScalars_->islandCount, hostScalars_->colLen[kOverflowIndex], hostScalars_->anyRestitution);
  lastEndIndex_ = endIndex_;
  endIndex_ = 1 - endIndex_;
  // Graph launches trade ~7 us of start latency for immunity to host launch jitter: off by default (a caller that synchronises every step is
  // latency-bound), on through b2cuda_world_enable_graph_launch / B2CUDA_GRAPH=1 (several ranks on one host, or steps enqueued back to back).
  if (graphMode_ < 0) { const char* e = getenv("B2CUDA_GRAPH"); graphMode_ = e && !strcmp(e, "1") && !d.debugChecksums ? 1 : 0; }
  bool launched = false;
  if (graphMode_ == 1) {
    // key = everything baked into the graph's nodes
    std::vector<unsigned char> key(sizeof(WorldDev) + sizeof(StepParams) + 5 * sizeof(int));
    { unsigned char* k = key.data(); memcpy(k, &d, sizeof(WorldDev)); k += sizeof(WorldDev); memcpy(k, &sp, sizeof(StepParams)); k += sizeof(StepParams);
      int f[5] = {endIndex_, warpMode, clusterLaunch, streamMoves_, streamBodies_}; memcpy(k, f, sizeof(f)); }
    StepGraph* g = nullptr;
    for (auto& c : graphs_) if (c.key == key) { g = &c; break; }
    if (!g) {
      if (graphs_.size() >= 8) {                    // parameters keep changing: drop the oldest
        StepGraph& o = graphs_.front();
        if (o.exec) cudaGraphExecDestroy(o.exec);
        graphs_.erase(graphs_.begin());
      }
      StepGraph ng; ng.key = key;
      cudaGraph_t graph = nullptr;
      bool ok = cudaStreamBeginCapture(stream_, cudaStreamCaptureModeThreadLocal) == cudaSuccess;
      if (ok) {
        int rc = enqueueStep(sp, warpMode, clusterLaunch, endIndex_);
        ok = cudaStreamEndCapture(stream_, &graph) == cudaSuccess && rc == B2C_OK && graph != nullptr;
      }
      ok = ok && cudaGraphInstantiate(&ng.exec, graph, 0) == cudaSuccess;
      if (graph) cudaGraphDestroy(graph);
      if (ok) { graphs_.push_back(std::move(ng)); g = &graphs_.back(); }
      else {                                          // this driver cannot capture the sequence: direct launches from now on
        cudaGetLastError();
        if (ng.exec) cudaGraphExecDestroy(ng.exec);
        graphMode_ = 0;
      }
    }
    if (g) {
      CK(cudaGraphLaunch(g->exec, stream_));
      launched = true;
    }
  }
  if (!launched) {
    int rc = enqueueStep(sp, warpMode, clusterLaunch, endIndex_); if (rc) return rc;
  }
  if (streamMoves_) movesFetched_ = true;
  if (streamBodies_) bodiesFetched_ = true;
  CK(cudaGetLastError());
  stepCount_++;
  return B2C_OK;
}

__global__ void __launch_bounds__(kThreads) k_pack_moves(WorldDev d, B2CBodyMoveEvent* out) {
  for (int i = gtid(); i < d.awakeCount; i += gsize()) {
    int bodyId = d.simBodyId[i];
    float4 xf = d.bodyXf[bodyId];
    B2CBodyMoveEvent ev; ev.transform.p = {xf.x, xf.y}; ev.transform.q = {xf.z, xf.w}; ev.bodyId = bodyId; ev.fellAsleep = 0;
    out[i] = ev;
  }
}
int World::getBodyMoveEvents(const B2CBodyMoveEvent** events, int* count) {
  *events = nullptr; *count = 0;
  if (!onDevice_) return fail(B2C_E_INVALID, "get_body_move_events before the first step");
  CK(cudaSetDevice(device_));
  { int rc = ensureHostBuffers(); if (rc) return rc; }
  if (!movesFetched_) {
    int n = d_.awakeCount;
    k_pack_moves<<<gridFor(n, kThreads, (coopBlocks_ > 0 ? coopBlocks_ : 148) * 8), kThreads, 0, stream_>>>(d_, dMoveEvents_);
    CK(cudaMemcpyAsync(hMoveEvents_, dMoveEvents_, sizeof(B2CBodyMoveEvent) * (size_t)n, cudaMemcpyDeviceToHost, stream_));
    movesFetched_ = true;
    if (!batch_) { streamMoves_ = true; d_.moveOut = hMoveEventsDev_ ? hMoveEventsDev_ : dMoveEvents_; }   // from now on FinalizeBodies fills the events (mapped host memory, else device array + copy in step())
  }
  int rc = synchronize(); if (rc) return rc;
  *events = hMoveEvents_; *count = lastStepDtZero_ ? 0 : d_.awakeCount;
  return B2C_OK;
}

// One step of every world of a batch (b2c_batch.cu drives it)
// CTA width: 160 threads at 96 registers (4 worlds per SM) while every island's constraints fit 160 register lanes, else 256 threads
// (2 worlds per SM). Measured on 8192 10-row pyramids (145 constraints each): 0.97 ms against 1.49 ms per step; narrower CTAs at
// 80 / 72 / 64 registers spill and lose (1.14 - 1.17 ms). `maxIslandCons` is a mapped host int the kernel keeps a running maximum in.
cudaError_t prepareBatchKernels() {                 // once per device, before the first launch (and outside any stream capture)
  return cudaFuncSetAttribute(k_batch_solve_warp, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(kWarpIslandsPerCta * sizeof(WarpIslandSlice)));
}
// Three launches while every island fits a warp slice-sized CTA (see MODE above): measured on 8192 10-row pyramids against the one-launch
// step (B2CUDA_BATCH_SPLIT=0 keeps one launch, =1 forces the split).
cudaError_t launchBatchStep(const WorldDev* worlds, int worldCount, const StepParams& sp, int newEndIndex, int* maxIslandConsDev, int maxIslandConsSeen, cudaStream_t stream) {
  static int forced = getenv("B2CUDA_BATCH_THREADS") ? atoi(getenv("B2CUDA_BATCH_THREADS")) : 0;
  static int split = getenv("B2CUDA_BATCH_SPLIT") ? atoi(getenv("B2CUDA_BATCH_SPLIT")) : -1;
  const bool narrow = forced ? forced == 160 : (maxIslandConsSeen > 0 && maxIslandConsSeen <= 160);
  const bool useSplit = split < 0 ? narrow : (split != 0 && narrow);
  if (useSplit) {
    k_batch_step<160, 4, 1><<<worldCount, 160, 0, stream>>>(worlds, sp, newEndIndex, maxIslandConsDev);
    k_batch_solve_warp<<<(worldCount + kWarpIslandsPerCta - 1) / kWarpIslandsPerCta, 32 * kWarpIslandsPerCta, kWarpIslandsPerCta * sizeof(WarpIslandSlice), stream>>>(worlds, worldCount, sp);
    k_batch_step<64, 12, 2><<<worldCount, 64, 0, stream>>>(worlds, sp, newEndIndex, maxIslandConsDev);   // FinalizeBodies of a small world: 64 threads at 80 registers, 12 worlds per SM (a dependent body -> shape -> fat-box load chain per thread: more worlds in flight hide it; 8 per SM at 128 registers 0.864-0.898 ms, 12 at 80: 0.847-0.851, 16 at 64 spills: 0.87-0.88)
  } else if (narrow) k_batch_step<160, 4, 0><<<worldCount, 160, 0, stream>>>(worlds, sp, newEndIndex, maxIslandConsDev);
  else k_batch_step<256, 2, 0><<<worldCount, 256, 0, stream>>>(worlds, sp, newEndIndex, maxIslandConsDev);
  return cudaGetLastError();
}
__global__ void k_batch_zero_step(const WorldDev* worlds, int newEndIndex) { Scalars& s = *worlds[blockIdx.x].sc; s.beginCount = 0; s.hitCount = 0; s.endCount[newEndIndex] = 0; s.sensorBeginCount = 0; s.sensorEndCount[newEndIndex] = 0; }
cudaError_t launchBatchZeroStep(const WorldDev* worlds, int worldCount, int newEndIndex, cudaStream_t stream) {
  k_batch_zero_step<<<worldCount, 1, 0, stream>>>(worlds, newEndIndex);
  return cudaGetLastError();
}
// Batch statistics reduced on the device: out[0..4] sums (awake bodies, |v|^2, touching contacts, begin events, end events of the last step),
// out[5] max |v|, out[6] min y, out[7] OR of the worlds' error flags. One CTA per world, one atomic per CTA and statistic.
__device__ __forceinline__ void atomicMaxD(double* a, double v) { unsigned long long* p = (unsigned long long*)a; unsigned long long old = *p; while (__longlong_as_double((long long)old) < v) { unsigned long long prev = atomicCAS(p, old, (unsigned long long)__double_as_longlong(v)); if (prev == old) break; old = prev; } }
__device__ __forceinline__ void atomicMinD(double* a, double v) { unsigned long long* p = (unsigned long long*)a; unsigned long long old = *p; while (__longlong_as_double((long long)old) > v) { unsigned long long prev = atomicCAS(p, old, (unsigned long long)__double_as_longlong(v)); if (prev == old) break; old = prev; } }
__global__ void k_batch_stats_init(double* out) { out[0] = out[1] = out[2] = out[3] = out[4] = 0.0; out[5] = 0.0; out[6] = 1e300; out[7] = 0.0; }
__global__ void __launch_bounds__(128) k_batch_stats(const WorldDev* worlds, int lastEndIndex, double* out, unsigned int* errOr) {
  const WorldDev& d = worlds[blockIdx.x];
  __shared__ double sSum[4], sMax[4], sMin[4];
  double sum = 0.0, mx = 0.0, mn = 1e300;
  for (int i = threadIdx.x; i < d.awakeCount; i += blockDim.x) {
    float4 s = d.st0[i];
    double v2 = (double)s.x * s.x + (double)s.y * s.y;
    sum += v2; if (v2 > mx) mx = v2;
    double y = d.bodyXf[d.simBodyId[i]].y; if (y < mn) mn = y;
  }
  for (int o = 16; o > 0; o >>= 1) { sum += __shfl_xor_sync(0xffffffffu, sum, o); double a = __shfl_xor_sync(0xffffffffu, mx, o), b = __shfl_xor_sync(0xffffffffu, mn, o); if (a > mx) mx = a; if (b < mn) mn = b; }
  if ((threadIdx.x & 31) == 0) { sSum[threadIdx.x >> 5] = sum; sMax[threadIdx.x >> 5] = mx; sMin[threadIdx.x >> 5] = mn; }
  __syncthreads();
  if (threadIdx.x == 0) {
    for (int w = 1; w < (int)(blockDim.x >> 5); w++) { sum += sSum[w]; if (sMax[w] > mx) mx = sMax[w]; if (sMin[w] < mn) mn = sMin[w]; }
    const Scalars& s = *d.sc;
    atomicAdd(&out[0], (double)d.awakeCount); atomicAdd(&out[1], sum); atomicAdd(&out[2], (double)s.touchingCount);
    atomicAdd(&out[3], (double)s.beginCount); atomicAdd(&out[4], (double)s.endCount[lastEndIndex]);
    atomicMaxD(&out[5], sqrt(mx)); if (d.awakeCount > 0) atomicMinD(&out[6], mn);
    if (s.errorFlags) atomicOr(errOr, (unsigned int)s.errorFlags);
  }
}
__global__ void k_batch_stats_finish(double* out, unsigned int* errOr) { out[7] = (double)*errOr; *errOr = 0u; }
cudaError_t launchBatchStats(const WorldDev* worlds, int worldCount, int lastEndIndex, double* out, unsigned int* errOr, cudaStream_t stream) {
  k_batch_stats_init<<<1, 1, 0, stream>>>(out);
  k_batch_stats<<<worldCount, 128, 0, stream>>>(worlds, lastEndIndex, out, errOr);
  k_batch_stats_finish<<<1, 1, 0, stream>>>(out, errOr);
  return cudaGetLastError();
}

int World::getBodyEvents(const B2CBodyRead** records, int* count) {
  *records = nullptr; *count = 0;
  if (!onDevice_) return fail(B2C_E_INVALID, "get_body_events before the first step");
  CK(cudaSetDevice(device_));
  { int rc = ensureHostBuffers(); if (rc) return rc; }
  if (!bodiesFetched_) {
    int n = d_.bodyCapacity;
    k_pack_bodies<<<gridFor(n, kThreads, (coopBlocks_ > 0 ? coopBlocks_ : 148) * 8), kThreads, 0, stream_>>>(d_, dBodyRead_);
    CK(cudaMemcpyAsync(hBodyRead_, dBodyRead_, sizeof(B2CBodyRead) * (size_t)n, cudaMemcpyDeviceToHost, stream_));
    bodiesFetched_ = true;
    if (!batch_) { streamBodies_ = true; d_.bodyOut = dBodyRead_; }   // from now on FinalizeBodies keeps the records current and step() copies them
  }
  int rc = synchronize(); if (rc) return rc;
  *records = hBodyRead_; *count = d_.bodyCapacity;
  return B2C_OK;
}

}  // namespace b2c
