// ORACLE — TEST INFRASTRUCTURE ONLY (see b2o_math.h header). Parity unpinned by the reference.
// Restatement of the reference's incremental SAH dynamic AABB tree.
// Follows: src/DynamicTree.cs:11-557 (node pool, FindBestSibling, RotateNodes, Insert/RemoveLeaf,
// Create/Destroy/Move/EnlargeProxy, Query) and :802-1000 (PartitionMid, BuildTree, Rebuild).
#pragma once
#include <vector>
#include <algorithm>
#include "b2o_math.h"

namespace b2o {

enum : uint16_t { kNodeAllocated = 1, kNodeEnlarged = 2, kNodeLeaf = 4 };

struct TreeNode {            // src/DynamicTree.cs:11-32 (explicit layout; children alias userData)
  AABB aabb;
  uint64_t categoryBits = ~0ull;
  int child1 = -1, child2 = -1;
  uint64_t userData = ~0ull;
  int parent = -1;           // aliases `next` in the free list
  uint16_t height = 0;
  uint16_t flags = kNodeAllocated;
};

struct DynamicTree {
  std::vector<TreeNode> nodes;
  int root = -1, nodeCount = 0, freeList = 0, proxyCount = 0;
  std::vector<int> leafIndices;
  std::vector<Vec2> leafCenters;

  DynamicTree() {              // :36-50
    nodes.resize(16);
    for (int i = 0; i < 15; i++) nodes[i].parent = i + 1;
    nodes[15].parent = -1;
  }
  int AllocateNode() {         // :54-72
    if (freeList == -1) {
      int old = (int)nodes.size();
      nodes.resize(old + (old >> 1));
      for (int i = nodeCount; i < (int)nodes.size(); i++) nodes[i].parent = i + 1;
      nodes.back().parent = -1;
      freeList = nodeCount;
    }
    int id = freeList;
    freeList = nodes[id].parent;
    nodes[id] = TreeNode();
    nodeCount++;
    return id;
  }
  void FreeNode(int id) {      // :74-82
    nodes[id].parent = freeList; nodes[id].flags = 0; freeList = id; nodeCount--;
  }
  int FindBestSibling(const AABB& boxD) {   // :98-184
    Vec2 centerD = Center(boxD);
    float areaD = Perimeter(boxD);
    int rootIndex = root;
    AABB rootBox = nodes[rootIndex].aabb;
    float areaBase = Perimeter(rootBox);
    float directCost = Perimeter(Union(rootBox, boxD));
    float inheritedCost = 0;
    int bestSibling = rootIndex;
    float bestCost = directCost;
    int index = rootIndex;
    while (nodes[index].height > 0) {
      int child1 = nodes[index].child1, child2 = nodes[index].child2;
      float cost = directCost + inheritedCost;
      if (cost < bestCost) { bestSibling = index; bestCost = cost; }
      inheritedCost += directCost - areaBase;
      bool leaf1 = nodes[child1].height == 0, leaf2 = nodes[child2].height == 0;
      float lowerCost1 = FLT_MAX;
      AABB box1 = nodes[child1].aabb;
      float directCost1 = Perimeter(Union(box1, boxD));
      float area1 = 0;
      if (leaf1) {
        float cost1 = directCost1 + inheritedCost;
        if (cost1 < bestCost) { bestSibling = child1; bestCost = cost1; }
      } else {
        area1 = Perimeter(box1);
        lowerCost1 = inheritedCost + directCost1 + netMin(areaD - area1, 0);
      }
      float lowerCost2 = FLT_MAX;
      AABB box2 = nodes[child2].aabb;
      float directCost2 = Perimeter(Union(box2, boxD));
      float area2 = 0;
      if (leaf2) {
        float cost2 = directCost2 + inheritedCost;
        if (cost2 < bestCost) { bestSibling = child2; bestCost = cost2; }
      } else {
        area2 = Perimeter(box2);
        lowerCost2 = inheritedCost + directCost2 + netMin(areaD - area2, 0);
      }
      if (leaf1 && leaf2) break;
      if (bestCost <= lowerCost1 && bestCost <= lowerCost2) break;
      if (lowerCost1 == lowerCost2 && !leaf1) {
        Vec2 d1 = Center(box1) - centerD, d2 = Center(box2) - centerD;
        lowerCost1 = LengthSquared(d1); lowerCost2 = LengthSquared(d2);
      }
      if (lowerCost1 < lowerCost2 && !leaf1) { index = child1; areaBase = area1; directCost = directCost1; }
      else { index = child2; areaBase = area2; directCost = directCost2; }
    }
    return bestSibling;
  }
  // one rotation: `up` child moves to A's slot `aSlot`(1|2); B(or C) takes C's (B's) child slot `cSlot`.
  void ApplyRot(int iA, int iInner, int iMove, int iUp, int iStay, bool aChild1, bool innerChild1, const AABB& box) {
    // A.childX = iUp; inner.childY = iMove; move.parent = inner; up.parent = A   (:211-219 etc.)
    TreeNode& A = nodes[iA]; TreeNode& I = nodes[iInner];
    if (aChild1) A.child1 = iUp; else A.child2 = iUp;
    if (innerChild1) I.child1 = iMove; else I.child2 = iMove;
    nodes[iMove].parent = iInner; nodes[iUp].parent = iA;
    I.aabb = box;
    I.height = (uint16_t)(1 + std::max(nodes[iMove].height, nodes[iStay].height));
    A.height = (uint16_t)(1 + std::max(I.height, nodes[iUp].height));
    I.categoryBits = nodes[iMove].categoryBits | nodes[iStay].categoryBits;
    A.categoryBits = I.categoryBits | nodes[iUp].categoryBits;
    I.flags |= (uint16_t)((nodes[iMove].flags | nodes[iStay].flags) & kNodeEnlarged);
    A.flags |= (uint16_t)((I.flags | nodes[iUp].flags) & kNodeEnlarged);
  }
  void RotateNodes(int iA) {   // :187-364
    TreeNode& A = nodes[iA];
    if (A.height < 2) return;
    int iB = A.child1, iC = A.child2;
    TreeNode& B = nodes[iB]; TreeNode& C = nodes[iC];
    if (B.height == 0) {
      int iF = C.child1, iG = C.child2;
      float costBase = Perimeter(C.aabb);
      AABB aabbBG = Union(B.aabb, nodes[iG].aabb); float costBG = Perimeter(aabbBG);
      AABB aabbBF = Union(B.aabb, nodes[iF].aabb); float costBF = Perimeter(aabbBF);
      if (costBase < costBF && costBase < costBG) return;
      if (costBF < costBG) ApplyRot(iA, iC, iB, iF, iG, true, true, aabbBG);   // swap B and F
      else ApplyRot(iA, iC, iB, iG, iF, true, false, aabbBF);                  // swap B and G
    } else if (C.height == 0) {
      int iD = B.child1, iE = B.child2;
      float costBase = Perimeter(B.aabb);
      AABB aabbCE = Union(C.aabb, nodes[iE].aabb); float costCE = Perimeter(aabbCE);
      AABB aabbCD = Union(C.aabb, nodes[iD].aabb); float costCD = Perimeter(aabbCD);
      if (costBase < costCD && costBase < costCE) return;
      if (costCD < costCE) ApplyRot(iA, iB, iC, iD, iE, false, true, aabbCE);  // swap C and D
      else ApplyRot(iA, iB, iC, iE, iD, false, false, aabbCD);                 // swap C and E
    } else {
      int iD = B.child1, iE = B.child2, iF = C.child1, iG = C.child2;
      float areaB = Perimeter(B.aabb), areaC = Perimeter(C.aabb);
      float costBase = areaB + areaC;
      int best = 0; float bestCost = costBase;
      AABB aabbBG = Union(B.aabb, nodes[iG].aabb); float costBF = areaB + Perimeter(aabbBG);
      if (costBF < bestCost) { best = 1; bestCost = costBF; }
      AABB aabbBF = Union(B.aabb, nodes[iF].aabb); float costBG = areaB + Perimeter(aabbBF);
      if (costBG < bestCost) { best = 2; bestCost = costBG; }
      AABB aabbCE = Union(C.aabb, nodes[iE].aabb); float costCD = areaC + Perimeter(aabbCE);
      if (costCD < bestCost) { best = 3; bestCost = costCD; }
      AABB aabbCD = Union(C.aabb, nodes[iD].aabb); float costCE = areaC + Perimeter(aabbCD);
      if (costCE < bestCost) { best = 4; }
      switch (best) {
        case 1: ApplyRot(iA, iC, iB, iF, iG, true, true, aabbBG); break;
        case 2: ApplyRot(iA, iC, iB, iG, iF, true, false, aabbBF); break;
        case 3: ApplyRot(iA, iB, iC, iD, iE, false, true, aabbCE); break;
        case 4: ApplyRot(iA, iB, iC, iE, iD, false, false, aabbCD); break;
        default: break;
      }
    }
  }
  void InsertLeaf(int leaf, bool shouldRotate) {   // :365-412
    if (root == -1) { root = leaf; nodes[root].parent = -1; return; }
    AABB leafAABB = nodes[leaf].aabb;
    int sibling = FindBestSibling(leafAABB);
    int oldParent = nodes[sibling].parent;
    int newParent = AllocateNode();
    nodes[newParent].parent = oldParent;
    nodes[newParent].userData = ~0ull;
    nodes[newParent].aabb = Union(leafAABB, nodes[sibling].aabb);
    nodes[newParent].categoryBits = nodes[leaf].categoryBits | nodes[sibling].categoryBits;
    nodes[newParent].height = (uint16_t)(nodes[sibling].height + 1);
    if (oldParent != -1) {
      if (nodes[oldParent].child1 == sibling) nodes[oldParent].child1 = newParent;
      else nodes[oldParent].child2 = newParent;
    } else root = newParent;
    nodes[newParent].child1 = sibling; nodes[newParent].child2 = leaf;
    nodes[sibling].parent = newParent; nodes[leaf].parent = newParent;
    int index = nodes[leaf].parent;
    while (index != -1) {
      int c1 = nodes[index].child1, c2 = nodes[index].child2;
      nodes[index].aabb = Union(nodes[c1].aabb, nodes[c2].aabb);
      nodes[index].categoryBits = nodes[c1].categoryBits | nodes[c2].categoryBits;
      nodes[index].height = (uint16_t)(1 + std::max(nodes[c1].height, nodes[c2].height));
      nodes[index].flags |= (uint16_t)((nodes[c1].flags | nodes[c2].flags) & kNodeEnlarged);
      if (shouldRotate) RotateNodes(index);
      index = nodes[index].parent;
    }
  }
  void RemoveLeaf(int leaf) {   // :413-441
    if (leaf == root) { root = -1; return; }
    int parent = nodes[leaf].parent, grandParent = nodes[parent].parent;
    int sibling = nodes[parent].child1 == leaf ? nodes[parent].child2 : nodes[parent].child1;
    if (grandParent != -1) {
      if (nodes[grandParent].child1 == parent) nodes[grandParent].child1 = sibling;
      else nodes[grandParent].child2 = sibling;
      nodes[sibling].parent = grandParent;
      FreeNode(parent);
      int index = grandParent;
      while (index != -1) {
        TreeNode& n = nodes[index]; TreeNode& c1 = nodes[n.child1]; TreeNode& c2 = nodes[n.child2];
        n.aabb = Union(c1.aabb, c2.aabb);
        n.categoryBits = c1.categoryBits | c2.categoryBits;
        n.height = (uint16_t)(1 + std::max(c1.height, c2.height));
        index = n.parent;
      }
    } else { root = sibling; nodes[sibling].parent = -1; FreeNode(parent); }
  }
  int CreateProxy(const AABB& aabb, uint64_t categoryBits, uint64_t userData) {   // :443-458
    int id = AllocateNode();
    nodes[id].aabb = aabb; nodes[id].userData = userData; nodes[id].categoryBits = categoryBits;
    nodes[id].height = 0; nodes[id].flags = kNodeAllocated | kNodeLeaf;
    nodes[id].child1 = nodes[id].child2 = -1;   // leaf: children alias userData in the C# layout
    InsertLeaf(id, true);
    proxyCount++;
    return id;
  }
  void DestroyProxy(int id) { RemoveLeaf(id); FreeNode(id); proxyCount--; }   // :460-468
  void MoveProxy(int id, const AABB& aabb) { RemoveLeaf(id); nodes[id].aabb = aabb; InsertLeaf(id, false); }  // :470-480
  void EnlargeProxy(int id, const AABB& aabb) {   // :482-505
    nodes[id].aabb = aabb;
    int p = nodes[id].parent;
    while (p != -1) {
      bool changed = Enlarge(nodes[p].aabb, aabb);
      nodes[p].flags |= kNodeEnlarged;
      p = nodes[p].parent;
      if (!changed) break;
    }
    while (p != -1) {
      if (nodes[p].flags & kNodeEnlarged) break;
      nodes[p].flags |= kNodeEnlarged;
      p = nodes[p].parent;
    }
  }
  // :531-557. Callback returns false to stop. child1 pushed first, child2 popped first.
  template <class F> void Query(const AABB& aabb, uint64_t maskBits, F&& cb) const {
    if (nodeCount == 0) return;
    std::vector<int> stack; stack.reserve(64);
    stack.push_back(root);
    while (!stack.empty()) {
      int id = stack.back(); stack.pop_back();
      const TreeNode& n = nodes[id];
      if (Overlaps(n.aabb, aabb) && (n.categoryBits & maskBits) != 0) {
        if (n.flags & kNodeLeaf) { if (!cb(id, n.userData)) return; }
        else { stack.push_back(n.child1); stack.push_back(n.child2); }
      }
    }
  }
  static int PartitionMid(int* indices, Vec2* centers, int count) {   // :802-847
    if (count <= 2) return count / 2;
    Vec2 lo = centers[0], hi = centers[0];
    for (int i = 1; i < count; i++) { lo = Min(lo, centers[i]); hi = Max(hi, centers[i]); }
    Vec2 d = hi - lo;
    Vec2 c = {0.5f * (lo.x + hi.x), 0.5f * (lo.y + hi.y)};
    int i1 = 0, i2 = count;
    if (d.x > d.y) {
      float pivot = c.x;
      while (i1 < i2) {
        while (i1 < i2 && centers[i1].x < pivot) i1++;
        while (i1 < i2 && centers[i2 - 1].x >= pivot) i2--;
        if (i1 < i2) { std::swap(indices[i1], indices[i2 - 1]); std::swap(centers[i1], centers[i2 - 1]); i1++; i2--; }
      }
    } else {
      float pivot = c.y;
      while (i1 < i2) {
        while (i1 < i2 && centers[i1].y < pivot) i1++;
        while (i1 < i2 && centers[i2 - 1].y >= pivot) i2--;
        if (i1 < i2) { std::swap(indices[i1], indices[i2 - 1]); std::swap(centers[i1], centers[i2 - 1]); i1++; i2--; }
      }
    }
    if (i1 > 0 && i1 < count) return i1;
    return count / 2;
  }
  struct RebuildItem { int nodeIndex, childCount, startIndex, splitIndex, endIndex; };
  int BuildTree(int leafCount) {   // :853-959
    if (leafCount == 1) { nodes[leafIndices[0]].parent = -1; return leafIndices[0]; }
    std::vector<RebuildItem> stack;
    {
      int n = AllocateNode();
      int split = PartitionMid(leafIndices.data(), leafCenters.data(), leafCount);
      stack.push_back({n, -1, 0, split, leafCount});
    }
    while (true) {
      RebuildItem& item = stack.back();
      item.childCount++;
      if (item.childCount == 2) {
        if (stack.size() == 1) break;
        RebuildItem done = item;
        stack.pop_back();
        RebuildItem& parentItem = stack.back();
        TreeNode& parentNode = nodes[parentItem.nodeIndex];
        if (parentItem.childCount == 0) parentNode.child1 = done.nodeIndex;
        else parentNode.child2 = done.nodeIndex;
        TreeNode& node = nodes[done.nodeIndex];
        node.parent = parentItem.nodeIndex;
        TreeNode& c1 = nodes[node.child1]; TreeNode& c2 = nodes[node.child2];
        node.aabb = Union(c1.aabb, c2.aabb);
        node.height = (uint16_t)(1 + std::max(c1.height, c2.height));
        node.categoryBits = c1.categoryBits | c2.categoryBits;
      } else {
        int startIndex, endIndex;
        if (item.childCount == 0) { startIndex = item.startIndex; endIndex = item.splitIndex; }
        else { startIndex = item.splitIndex; endIndex = item.endIndex; }
        int count = endIndex - startIndex;
        if (count == 1) {
          int childIndex = leafIndices[startIndex];
          TreeNode& node = nodes[item.nodeIndex];
          if (item.childCount == 0) node.child1 = childIndex; else node.child2 = childIndex;
          nodes[childIndex].parent = item.nodeIndex;
        } else {
          int n = AllocateNode();   // may reallocate `nodes`; `item` refers to `stack`, still valid until push
          int split = startIndex + PartitionMid(leafIndices.data() + startIndex, leafCenters.data() + startIndex, count);
          stack.push_back({n, -1, startIndex, split, endIndex});
        }
      }
    }
    TreeNode& rootNode = nodes[stack.front().nodeIndex];
    TreeNode& c1 = nodes[rootNode.child1]; TreeNode& c2 = nodes[rootNode.child2];
    rootNode.aabb = Union(c1.aabb, c2.aabb);
    rootNode.height = (uint16_t)(1 + std::max(c1.height, c2.height));
    rootNode.categoryBits = c1.categoryBits | c2.categoryBits;
    return stack.front().nodeIndex;
  }
  int Rebuild(bool fullBuild) {   // :961-1000
    if (proxyCount == 0) return 0;
    if ((int)leafIndices.size() < proxyCount) { leafIndices.resize(proxyCount + proxyCount / 2); leafCenters.resize(leafIndices.size()); }
    int leafCount = 0;
    std::vector<int> stack;
    int nodeIndex = root;
    while (true) {
      TreeNode& node = nodes[nodeIndex];
      if (node.height == 0 || (!fullBuild && !(node.flags & kNodeEnlarged))) {
        leafIndices[leafCount] = nodeIndex;
        leafCenters[leafCount] = Center(node.aabb);
        leafCount++;
        node.parent = -1;
      } else {
        int doomed = nodeIndex;
        nodeIndex = node.child1;
        stack.push_back(node.child2);
        FreeNode(doomed);
        continue;
      }
      if (stack.empty()) break;
      nodeIndex = stack.back(); stack.pop_back();
    }
    root = BuildTree(leafCount);
    return leafCount;
  }
};

}  // namespace b2o
