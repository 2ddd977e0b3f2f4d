// libb2cuda — creation-time proxy insertion (host). Reproduces the topology the reference's incremental SAH tree
// would have after the same CreateProxy/MoveProxy sequence (src/DynamicTree.cs:36-480), because the contact creation
// order — and through it contact ids and graph colors — depends on the DFS order of that exact tree.
// The tree operations themselves are shared with the device (b2c_treeops.cuh); this file owns the growable storage.
// Step-time tree work (Query, EnlargeProxy, Rebuild) is device code in b2c_step.cu.
#include "b2c_host.cuh"
#include "b2c_treeops.cuh"

namespace b2c {

HostTree::HostTree() {
  const int n = 16;                              // :40
  aabb.assign(n, make_float4(0, 0, 0, 0)); child.assign(n, make_int2(-1, -1)); parent.resize(n); hf.assign(n, 0);
  cat.assign(n, 0); user.assign(n, -1);
  for (int i = 0; i < n - 1; i++) parent[i] = i + 1;
  parent[n - 1] = -1;
}

// Growth branch of AllocateNode (:56-66): taken before an allocating operation so the fixed-size view never runs dry.
void HostTree::ensureFreeNode() {
  if (freeList != -1) return;
  int old = capacity(), grown = old + (old >> 1);
  aabb.resize(grown, make_float4(0, 0, 0, 0)); child.resize(grown, make_int2(-1, -1)); parent.resize(grown); hf.resize(grown, 0);
  cat.resize(grown, 0); user.resize(grown, -1);
  for (int i = nodeCount; i < grown; i++) parent[i] = i + 1;
  parent[grown - 1] = -1;
  freeList = nodeCount;
}

namespace {
struct Scope {                                   // view over the vectors; the four scalars round-trip through meta[]
  HostTree& h; int meta[4]; TreeView v;
  explicit Scope(HostTree& t) : h(t), meta{t.root, t.nodeCount, t.freeList, t.proxyCount},
    v{t.aabb.data(), t.child.data(), t.parent.data(), t.hf.data(), t.cat.data(), t.user.data(), meta} {}
  ~Scope() { h.root = meta[0]; h.nodeCount = meta[1]; h.freeList = meta[2]; h.proxyCount = meta[3]; }
};
}  // namespace

int HostTree::createProxy(Box b, uint64_t categoryBits, int userData) {   // :443-458
  int id;
  ensureFreeNode();
  { Scope s(*this); id = tvAllocNode(s.v); tvSetBox(s.v, id, b); user[id] = userData; cat[id] = categoryBits; hf[id] = tvPackHF(0, NODE_ALLOCATED | NODE_LEAF); }
  ensureFreeNode();
  { Scope s(*this); tvInsertLeaf(s.v, id, true); s.meta[3]++; }
  return id;
}

void HostTree::moveProxy(int id, Box b) { Scope s(*this); tvMoveProxy(s.v, id, b); }   // :470-480 (frees a node before it allocates one)

}  // namespace b2c
