// libb2cuda — incremental DynamicTree operations (src/DynamicTree.cs:36-480) over a raw-pointer view of the SoA tree.
//
// One implementation serves both sides of the upload: HostTree (b2c_host_tree.cu) uses it while the world is being
// created, and the device uses it from one-thread kernels for the between-step mutators that move a proxy
// (BodyAPI.SetTransform, api/Body.cs:293-325). The topology must be the reference's exactly: contact creation order —
// and through it contact ids and graph colours — depends on the DFS order of this tree.
// The view never grows: the caller guarantees a free node before an operation that allocates (HostTree grows its
// vectors first; MoveProxy frees a node before it allocates one).
#pragma once
#include "b2c_world.cuh"

namespace b2c {

struct TreeView {
  float4* aabb; int2* child; int* parent; uint32_t* hf; uint64_t* cat; int* user;
  int* meta;           // [0] root [1] nodeCount [2] freeList [3] proxyCount
};
B2C_HD uint32_t tvPackHF(int height, uint32_t flags) { return (uint32_t)height | (flags << 16); }
B2C_HD Box tvBox(const TreeView& t, int i) { float4 a = t.aabb[i]; Box b; b.lo = mk(a.x, a.y); b.hi = mk(a.z, a.w); return b; }
B2C_HD void tvSetBox(const TreeView& t, int i, Box b) { t.aabb[i] = make_float4(b.lo.x, b.lo.y, b.hi.x, b.hi.y); }
B2C_HD int tvHeight(const TreeView& t, int i) { return (int)(t.hf[i] & 0xffffu); }
B2C_HD uint32_t tvFlags(const TreeView& t, int i) { return t.hf[i] >> 16; }
B2C_HD void tvSetHeight(const TreeView& t, int i, int h) { t.hf[i] = (t.hf[i] & 0xffff0000u) | (uint32_t)h; }
B2C_HD void tvOrFlags(const TreeView& t, int i, uint32_t f) { t.hf[i] |= f << 16; }
B2C_HD int tvMax(int a, int b) { return a > b ? a : b; }

B2C_HD int tvAllocNode(const TreeView& t) {       // :54-72 without the growth branch (see header comment)
  int id = t.meta[2];
  if (id == -1) return -1;
  t.meta[2] = t.parent[id];
  t.aabb[id] = make_float4(0, 0, 0, 0); t.child[id] = make_int2(-1, -1); t.parent[id] = -1; t.hf[id] = tvPackHF(0, NODE_ALLOCATED);
  t.cat[id] = ~0ull; t.user[id] = -1;            // Box2D.DEFAULT_CATEGORY_BITS (include/Box2D.cs:8)
  t.meta[1]++;
  return id;
}
B2C_HD void tvFreeNode(const TreeView& t, int id) { t.parent[id] = t.meta[2]; t.hf[id] &= 0xffffu; t.meta[2] = id; t.meta[1]--; }   // :74-82

B2C_HD int tvFindBestSibling(const TreeView& t, Box boxD) {   // :98-184
  const int root = t.meta[0];
  V2 centerD = boxCenter(boxD);
  float areaD = boxPerimeter(boxD);
  Box rootBox = tvBox(t, root);
  float areaBase = boxPerimeter(rootBox);
  float directCost = boxPerimeter(boxUnion(rootBox, boxD));
  float inheritedCost = 0.0f;
  int bestSibling = root;
  float bestCost = directCost;
  int index = root;
  while (tvHeight(t, index) > 0) {
    int c1 = t.child[index].x, c2 = t.child[index].y;
    float cost = directCost + inheritedCost;
    if (cost < bestCost) { bestSibling = index; bestCost = cost; }
    inheritedCost += directCost - areaBase;
    bool leaf1 = tvHeight(t, c1) == 0, leaf2 = tvHeight(t, c2) == 0;
    float lower1 = FLT_MAX, lower2 = FLT_MAX, area1 = 0.0f, area2 = 0.0f;
    Box box1 = tvBox(t, c1), box2 = tvBox(t, c2);
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
B2C_HD void tvApplyRot(const TreeView& t, int iA, int iInner, int iMove, int iUp, int iStay, bool aChild1, bool innerChild1, Box merged) {
  if (aChild1) t.child[iA].x = iUp; else t.child[iA].y = iUp;
  if (innerChild1) t.child[iInner].x = iMove; else t.child[iInner].y = iMove;
  t.parent[iMove] = iInner; t.parent[iUp] = iA;
  tvSetBox(t, iInner, merged);
  tvSetHeight(t, iInner, 1 + tvMax(tvHeight(t, iMove), tvHeight(t, iStay)));
  tvSetHeight(t, iA, 1 + tvMax(tvHeight(t, iInner), tvHeight(t, iUp)));
  t.cat[iInner] = t.cat[iMove] | t.cat[iStay];
  t.cat[iA] = t.cat[iInner] | t.cat[iUp];
  tvOrFlags(t, iInner, (tvFlags(t, iMove) | tvFlags(t, iStay)) & NODE_ENLARGED);
  tvOrFlags(t, iA, (tvFlags(t, iInner) | tvFlags(t, iUp)) & NODE_ENLARGED);
}

B2C_HD void tvRotate(const TreeView& t, int iA) {
  if (tvHeight(t, iA) < 2) return;
  int iB = t.child[iA].x, iC = t.child[iA].y;
  if (tvHeight(t, iB) == 0) {
    int iF = t.child[iC].x, iG = t.child[iC].y;
    float costBase = boxPerimeter(tvBox(t, iC));
    Box bg = boxUnion(tvBox(t, iB), tvBox(t, iG)); float costBG = boxPerimeter(bg);
    Box bf = boxUnion(tvBox(t, iB), tvBox(t, iF)); float costBF = boxPerimeter(bf);
    if (costBase < costBF && costBase < costBG) return;
    if (costBF < costBG) tvApplyRot(t, iA, iC, iB, iF, iG, true, true, bg);
    else tvApplyRot(t, iA, iC, iB, iG, iF, true, false, bf);
  } else if (tvHeight(t, iC) == 0) {
    int iD = t.child[iB].x, iE = t.child[iB].y;
    float costBase = boxPerimeter(tvBox(t, iB));
    Box ce = boxUnion(tvBox(t, iC), tvBox(t, iE)); float costCE = boxPerimeter(ce);
    Box cd = boxUnion(tvBox(t, iC), tvBox(t, iD)); float costCD = boxPerimeter(cd);
    if (costBase < costCD && costBase < costCE) return;
    if (costCD < costCE) tvApplyRot(t, iA, iB, iC, iD, iE, false, true, ce);
    else tvApplyRot(t, iA, iB, iC, iE, iD, false, false, cd);
  } else {
    int iD = t.child[iB].x, iE = t.child[iB].y, iF = t.child[iC].x, iG = t.child[iC].y;
    float areaB = boxPerimeter(tvBox(t, iB)), areaC = boxPerimeter(tvBox(t, iC));
    int best = 0; float bestCost = areaB + areaC;
    Box bg = boxUnion(tvBox(t, iB), tvBox(t, iG)); float costBF = areaB + boxPerimeter(bg);
    if (costBF < bestCost) { best = 1; bestCost = costBF; }
    Box bf = boxUnion(tvBox(t, iB), tvBox(t, iF)); float costBG = areaB + boxPerimeter(bf);
    if (costBG < bestCost) { best = 2; bestCost = costBG; }
    Box ce = boxUnion(tvBox(t, iC), tvBox(t, iE)); float costCD = areaC + boxPerimeter(ce);
    if (costCD < bestCost) { best = 3; bestCost = costCD; }
    Box cd = boxUnion(tvBox(t, iC), tvBox(t, iD)); float costCE = areaC + boxPerimeter(cd);
    if (costCE < bestCost) best = 4;
    if (best == 1) tvApplyRot(t, iA, iC, iB, iF, iG, true, true, bg);
    else if (best == 2) tvApplyRot(t, iA, iC, iB, iG, iF, true, false, bf);
    else if (best == 3) tvApplyRot(t, iA, iB, iC, iD, iE, false, true, ce);
    else if (best == 4) tvApplyRot(t, iA, iB, iC, iE, iD, false, false, cd);
  }
}

// Returns false when no node could be allocated (the caller reports ERR_TREE / grows and retries).
B2C_HD bool tvInsertLeaf(const TreeView& t, int leaf, bool shouldRotate) {   // :365-412
  if (t.meta[0] == -1) { t.meta[0] = leaf; t.parent[leaf] = -1; return true; }
  Box leafBox = tvBox(t, leaf);
  int sibling = tvFindBestSibling(t, leafBox);
  int oldParent = t.parent[sibling];
  int np = tvAllocNode(t);
  if (np == -1) return false;
  t.parent[np] = oldParent;
  tvSetBox(t, np, boxUnion(leafBox, tvBox(t, sibling)));
  t.cat[np] = t.cat[leaf] | t.cat[sibling];
  tvSetHeight(t, np, tvHeight(t, sibling) + 1);
  if (oldParent != -1) { if (t.child[oldParent].x == sibling) t.child[oldParent].x = np; else t.child[oldParent].y = np; }
  else t.meta[0] = np;
  t.child[np] = make_int2(sibling, leaf);
  t.parent[sibling] = np; t.parent[leaf] = np;
  int index = t.parent[leaf];
  while (index != -1) {
    int c1 = t.child[index].x, c2 = t.child[index].y;
    tvSetBox(t, index, boxUnion(tvBox(t, c1), tvBox(t, c2)));
    t.cat[index] = t.cat[c1] | t.cat[c2];
    tvSetHeight(t, index, 1 + tvMax(tvHeight(t, c1), tvHeight(t, c2)));
    tvOrFlags(t, index, (tvFlags(t, c1) | tvFlags(t, c2)) & NODE_ENLARGED);
    if (shouldRotate) tvRotate(t, index);
    index = t.parent[index];
  }
  return true;
}

B2C_HD void tvRemoveLeaf(const TreeView& t, int leaf) {     // :413-441
  if (leaf == t.meta[0]) { t.meta[0] = -1; return; }
  int p = t.parent[leaf], gp = t.parent[p];
  int sibling = t.child[p].x == leaf ? t.child[p].y : t.child[p].x;
  if (gp != -1) {
    if (t.child[gp].x == p) t.child[gp].x = sibling; else t.child[gp].y = sibling;
    t.parent[sibling] = gp;
    tvFreeNode(t, p);
    for (int index = gp; index != -1; index = t.parent[index]) {
      int c1 = t.child[index].x, c2 = t.child[index].y;
      tvSetBox(t, index, boxUnion(tvBox(t, c1), tvBox(t, c2)));
      t.cat[index] = t.cat[c1] | t.cat[c2];
      tvSetHeight(t, index, 1 + tvMax(tvHeight(t, c1), tvHeight(t, c2)));
    }
  } else { t.meta[0] = sibling; t.parent[sibling] = -1; tvFreeNode(t, p); }
}

B2C_HD bool tvMoveProxy(const TreeView& t, int id, Box b) { tvRemoveLeaf(t, id); tvSetBox(t, id, b); return tvInsertLeaf(t, id, false); }   // :470-480

}  // namespace b2c
