// libb2cuda — creation-time proxy insertion (host). Reproduces the topology the reference's incremental SAH tree
// would have after the same CreateProxy/MoveProxy sequence (src/DynamicTree.cs:36-480), because the contact creation
// order — and through it contact ids and graph colors — depends on the DFS order of that exact tree.
// Step-time tree work (Query, EnlargeProxy, Rebuild) is device code in b2c_step.cu.
#include "b2c_host.cuh"
#include <algorithm>

namespace b2c {

static inline uint32_t packHF(int height, uint32_t flags) { return (uint32_t)height | (flags << 16); }

HostTree::HostTree() {
  const int n = 16;                              // :40
  aabb.assign(n, make_float4(0, 0, 0, 0)); child.assign(n, make_int2(-1, -1)); parent.resize(n); hf.assign(n, 0);
  cat.assign(n, 0); user.assign(n, -1);
  for (int i = 0; i < n - 1; i++) parent[i] = i + 1;
  parent[n - 1] = -1;
}

int HostTree::allocNode() {                      // :54-72
  if (freeList == -1) {
    int old = capacity(), grown = old + (old >> 1);
    aabb.resize(grown, make_float4(0, 0, 0, 0)); child.resize(grown, make_int2(-1, -1)); parent.resize(grown); hf.resize(grown, 0);
    cat.resize(grown, 0); user.resize(grown, -1);
    for (int i = nodeCount; i < grown; i++) parent[i] = i + 1;
    parent[grown - 1] = -1;
    freeList = nodeCount;
  }
  int id = freeList;
  freeList = parent[id];
  aabb[id] = make_float4(0, 0, 0, 0); child[id] = make_int2(-1, -1); parent[id] = -1; hf[id] = packHF(0, NODE_ALLOCATED);
  cat[id] = ~0ull; user[id] = -1;                // Box2D.DEFAULT_CATEGORY_BITS (include/Box2D.cs:8)
  nodeCount++;
  return id;
}

void HostTree::freeNode(int id) { parent[id] = freeList; hf[id] &= 0xffffu; freeList = id; nodeCount--; }   // :74-82

int HostTree::findBestSibling(Box boxD) {        // :98-184
  V2 centerD = boxCenter(boxD);
  float areaD = boxPerimeter(boxD);
  Box rootBox = box(root);
  float areaBase = boxPerimeter(rootBox);
  float directCost = boxPerimeter(boxUnion(rootBox, boxD));
  float inheritedCost = 0.0f;
  int bestSibling = root;
  float bestCost = directCost;
  int index = root;
  while (height(index) > 0) {
    int c1 = child[index].x, c2 = child[index].y;
    float cost = directCost + inheritedCost;
    if (cost < bestCost) { bestSibling = index; bestCost = cost; }
    inheritedCost += directCost - areaBase;
    bool leaf1 = height(c1) == 0, leaf2 = height(c2) == 0;
    float lower1 = FLT_MAX, lower2 = FLT_MAX, area1 = 0.0f, area2 = 0.0f;
    Box box1 = box(c1), box2 = box(c2);
    float direct1 = boxPerimeter(boxUnion(box1, boxD));
    if (leaf1) { float cost1 = direct1 + inheritedCost; if (cost1 < bestCost) { bestSibling = c1; bestCost = cost1; } }
    else { area1 = boxPerimeter(box1); lower1 = inheritedCost + direct1 + netMin(areaD - area1, 0.0f); }
    float direct2 = boxPerimeter(boxUnion(box2, boxD));
    if (leaf2) { float cost2 = direct2 + inheritedCost; if (cost2 < bestCost) { bestSibling = c2; bestCost = cost2; } }
    else { area2 = boxPerimeter(box2); lower2 = inheritedCost + direct2 + netMin(areaD - area2, 0.0f); }
    if (leaf1 && leaf2) break;
    if (bestCost <= lower1 && bestCost <= lower2) break;
    if (lower1 == lower2 && !leaf1) {
      lower1 = lengthSq(boxCenter(box1) - centerD);
      lower2 = lengthSq(boxCenter(box2) - centerD);
    }
    if (lower1 < lower2 && !leaf1) { index = c1; areaBase = area1; directCost = direct1; }
    else { index = c2; areaBase = area2; directCost = direct2; }
  }
  return bestSibling;
}

// One of the four rotations of :187-364: node `iUp` (a grandchild) takes the slot of A's child `iMove`, which
// becomes a child of the other child of A (`iInner`) next to `iStay`.
void HostTree::applyRot(int iA, int iInner, int iMove, int iUp, int iStay, bool aChild1, bool innerChild1, Box merged) {
  if (aChild1) child[iA].x = iUp; else child[iA].y = iUp;
  if (innerChild1) child[iInner].x = iMove; else child[iInner].y = iMove;
  parent[iMove] = iInner; parent[iUp] = iA;
  setBox(iInner, merged);
  setHeight(iInner, 1 + std::max(height(iMove), height(iStay)));
  setHeight(iA, 1 + std::max(height(iInner), height(iUp)));
  cat[iInner] = cat[iMove] | cat[iStay];
  cat[iA] = cat[iInner] | cat[iUp];
  orFlags(iInner, (flags(iMove) | flags(iStay)) & NODE_ENLARGED);
  orFlags(iA, (flags(iInner) | flags(iUp)) & NODE_ENLARGED);
}

void HostTree::rotate(int iA) {
  if (height(iA) < 2) return;
  int iB = child[iA].x, iC = child[iA].y;
  if (height(iB) == 0) {
    int iF = child[iC].x, iG = child[iC].y;
    float costBase = boxPerimeter(box(iC));
    Box bg = boxUnion(box(iB), box(iG)); float costBG = boxPerimeter(bg);
    Box bf = boxUnion(box(iB), box(iF)); float costBF = boxPerimeter(bf);
    if (costBase < costBF && costBase < costBG) return;
    if (costBF < costBG) applyRot(iA, iC, iB, iF, iG, true, true, bg);
    else applyRot(iA, iC, iB, iG, iF, true, false, bf);
  } else if (height(iC) == 0) {
    int iD = child[iB].x, iE = child[iB].y;
    float costBase = boxPerimeter(box(iB));
    Box ce = boxUnion(box(iC), box(iE)); float costCE = boxPerimeter(ce);
    Box cd = boxUnion(box(iC), box(iD)); float costCD = boxPerimeter(cd);
    if (costBase < costCD && costBase < costCE) return;
    if (costCD < costCE) applyRot(iA, iB, iC, iD, iE, false, true, ce);
    else applyRot(iA, iB, iC, iE, iD, false, false, cd);
  } else {
    int iD = child[iB].x, iE = child[iB].y, iF = child[iC].x, iG = child[iC].y;
    float areaB = boxPerimeter(box(iB)), areaC = boxPerimeter(box(iC));
    int best = 0; float bestCost = areaB + areaC;
    Box bg = boxUnion(box(iB), box(iG)); float costBF = areaB + boxPerimeter(bg);
    if (costBF < bestCost) { best = 1; bestCost = costBF; }
    Box bf = boxUnion(box(iB), box(iF)); float costBG = areaB + boxPerimeter(bf);
    if (costBG < bestCost) { best = 2; bestCost = costBG; }
    Box ce = boxUnion(box(iC), box(iE)); float costCD = areaC + boxPerimeter(ce);
    if (costCD < bestCost) { best = 3; bestCost = costCD; }
    Box cd = boxUnion(box(iC), box(iD)); float costCE = areaC + boxPerimeter(cd);
    if (costCE < bestCost) best = 4;
    if (best == 1) applyRot(iA, iC, iB, iF, iG, true, true, bg);
    else if (best == 2) applyRot(iA, iC, iB, iG, iF, true, false, bf);
    else if (best == 3) applyRot(iA, iB, iC, iD, iE, false, true, ce);
    else if (best == 4) applyRot(iA, iB, iC, iE, iD, false, false, cd);
  }
}

void HostTree::insertLeaf(int leaf, bool shouldRotate) {   // :365-412
  if (root == -1) { root = leaf; parent[root] = -1; return; }
  Box leafBox = box(leaf);
  int sibling = findBestSibling(leafBox);
  int oldParent = parent[sibling];
  int np = allocNode();
  parent[np] = oldParent;
  setBox(np, boxUnion(leafBox, box(sibling)));
  cat[np] = cat[leaf] | cat[sibling];
  setHeight(np, height(sibling) + 1);
  if (oldParent != -1) { if (child[oldParent].x == sibling) child[oldParent].x = np; else child[oldParent].y = np; }
  else root = np;
  child[np] = make_int2(sibling, leaf);
  parent[sibling] = np; parent[leaf] = np;
  int index = parent[leaf];
  while (index != -1) {
    int c1 = child[index].x, c2 = child[index].y;
    setBox(index, boxUnion(box(c1), box(c2)));
    cat[index] = cat[c1] | cat[c2];
    setHeight(index, 1 + std::max(height(c1), height(c2)));
    orFlags(index, (flags(c1) | flags(c2)) & NODE_ENLARGED);
    if (shouldRotate) rotate(index);
    index = parent[index];
  }
}

void HostTree::removeLeaf(int leaf) {            // :413-441
  if (leaf == root) { root = -1; return; }
  int p = parent[leaf], gp = parent[p];
  int sibling = child[p].x == leaf ? child[p].y : child[p].x;
  if (gp != -1) {
    if (child[gp].x == p) child[gp].x = sibling; else child[gp].y = sibling;
    parent[sibling] = gp;
    freeNode(p);
    for (int index = gp; index != -1; index = parent[index]) {
      int c1 = child[index].x, c2 = child[index].y;
      setBox(index, boxUnion(box(c1), box(c2)));
      cat[index] = cat[c1] | cat[c2];
      setHeight(index, 1 + std::max(height(c1), height(c2)));
    }
  } else { root = sibling; parent[sibling] = -1; freeNode(p); }
}

int HostTree::createProxy(Box b, uint64_t categoryBits, int userData) {   // :443-458
  int id = allocNode();
  setBox(id, b); user[id] = userData; cat[id] = categoryBits;
  hf[id] = packHF(0, NODE_ALLOCATED | NODE_LEAF);
  insertLeaf(id, true);
  proxyCount++;
  return id;
}

void HostTree::moveProxy(int id, Box b) { removeLeaf(id); setBox(id, b); insertLeaf(id, false); }   // :470-480

}  // namespace b2c
