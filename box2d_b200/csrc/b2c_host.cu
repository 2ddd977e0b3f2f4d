// libb2cuda — host-side creation API, device upload and read-back (see b2c_host.cuh).
// Mirrors api/Body.cs:16-105 (CreateBody), src/Shape.cs:171-218 (CreateShapeInternal), src/Body.cs:255-353
// (UpdateBodyMassData), src/joints/RevoluteJoint.cs:32-56 + src/Joint.cs:181-296 (CreateJoint),
// src/ConstraintGraph.cs:91-128 (AssignJointColor). None of this is on the step path.
#include "b2c_host.cuh"
#include "b2c_jointops.cuh"
#include "b2c_treeops.cuh"
#include "b2c_distance.cuh"
#include <cstring>
#include <cstdio>
#include <algorithm>

namespace b2c {

#define CK(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) return fail(B2C_E_CUDA, std::string(#call) + ": " + cudaGetErrorString(e_)); } while (0)

static inline int proxyKeyOf(int id, int type) { return (id << 2) | type; }

// ---- mass / extent (include/Collision.cs:82-92,205-228,431-497,649-653) ----
struct MassData { float mass; V2 center; float rotationalInertia; };
static MassData circleMass(V2 center, float radius, float density) {
  float rr = radius * radius, mass = density * kPi * rr;
  MassData m; m.mass = mass; m.center = center; m.rotationalInertia = mass * 0.5f * rr; return m;
}
static MassData capsuleMass(V2 p1, V2 p2, float radius, float density) {
  float rr = radius * radius;
  float ll = distSq(p1, p2), len = sqrtf(ll);
  float circleM = density * (kPi * radius * radius);
  float boxM = density * (2.0f * radius * radius);
  MassData m; m.mass = circleM + boxM; m.center = mk(0.5f * (p1.x + p2.x), 0.5f * (p1.y + p2.y));
  float lc = 4.0f * radius / (3.0f * kPi);
  float h = 0.5f * len;
  float circleInertia = circleM * (0.5f * rr + h * h + 2.0f * h * lc);
  float boxInertia = boxM * (4.0f * rr + ll) / 12.0f;
  m.rotationalInertia = circleInertia + boxInertia;
  return m;
}
static MassData geomMass(const Geom& g, float density) {
  MassData zero; zero.mass = 0; zero.center = mk(0, 0); zero.rotationalInertia = 0;
  if (g.type == kCircle) return circleMass(g.v[0], g.radius, density);
  if (g.type == kCapsule) return capsuleMass(g.v[0], g.v[1], g.radius, density);
  if (g.type != kPolygon) return zero;
  if (g.count == 1) return circleMass(g.v[0], g.radius, density);
  if (g.count == 2) return capsuleMass(g.v[0], g.v[1], g.radius, density);
  V2 verts[kMaxPolygonVertices] = {};
  if (g.radius > 0.0f) {
    const float sqrt2 = 1.4142135623730950488016887242097f;
    for (int i = 0, j = g.count - 1; i < g.count; j = i++) {
      V2 mid = normalize(g.n[j] + g.n[i]);
      verts[i] = mulAdd(g.v[i], sqrt2 * g.radius, mid);
    }
  } else for (int i = 0; i < g.count; i++) verts[i] = g.v[i];
  V2 center = mk(0, 0);
  float area = 0, rotationalInertia = 0;
  V2 r = verts[0];
  const float inv3 = 1.0f / 3.0f;
  for (int i = 1; i < g.count - 1; i++) {
    V2 e1 = verts[i] - r, e2 = verts[i + 1] - r;
    float D = cross(e1, e2);
    float triArea = 0.5f * D;
    area += triArea;
    center = mulAdd(center, triArea * inv3, e1 + e2);
    float intx2 = e1.x * e1.x + e2.x * e1.x + e2.x * e2.x;
    float inty2 = e1.y * e1.y + e2.y * e1.y + e2.y * e2.y;
    rotationalInertia += 0.25f * inv3 * D * (intx2 + inty2);
  }
  float invArea = 1.0f / area;
  center.x *= invArea; center.y *= invArea;
  MassData m;
  m.mass = density * area;
  m.center = r + center;
  m.rotationalInertia = density * rotationalInertia - density * area * dot(center, center);
  return m;
}
static void geomExtent(const Geom& g, V2 localCenter, float* minExtent, float* maxExtent) {
  if (g.type == kCircle) { *minExtent = g.radius; *maxExtent = dist(g.v[0], localCenter) + g.radius; }
  else if (g.type == kCapsule) { *minExtent = g.radius; *maxExtent = sqrtf(netMax(distSq(g.v[0], localCenter), distSq(g.v[1], localCenter))) + g.radius; }
  else if (g.type == kSegment) { *minExtent = 0; *maxExtent = sqrtf(netMax(distSq(g.v[0], localCenter), distSq(g.v[1], localCenter))); }
  else {
    float mn = kHuge, mxSq = 0;
    for (int i = 0; i < g.count; i++) {
      float planeOffset = dot(g.n[i], g.v[i] - g.centroid);
      mn = netMin(mn, planeOffset);
      mxSq = netMax(mxSq, distSq(g.v[i], localCenter));
    }
    *minExtent = mn + g.radius; *maxExtent = sqrtf(mxSq) + g.radius;
  }
}

World::World(const B2CWorldDef& def, int device, Batch* batch, int batchIndex) : device_(device), batch_(batch), batchIndex_(batchIndex) {
  gravity = mk(def.gravity.x, def.gravity.y);
  restitutionThreshold = def.restitutionThreshold; hitEventThreshold = def.hitEventThreshold; contactHertz = def.contactHertz;
  contactDampingRatio = def.contactDampingRatio; contactSpeed = def.contactSpeed; maxLinearSpeed = def.maximumLinearSpeed;
  enableSleep = def.enableSleep != 0; enableContinuous = def.enableContinuous != 0; enableContactSoftening = def.enableContactSoftening != 0;
  capHintBody = def.bodyCapacity; capHintShape = def.shapeCapacity; capHintContact = def.contactCapacity; capHintJoint = def.jointCapacity;
  moved_.resize(3);
  colorJoints_.resize(kGraphColorCount);
}

World::~World() { freeDevice(); }

void World::freeDevice() {
  if (device_ >= 0) cudaSetDevice(device_);
  if (!batch_) for (void* p : allocs_) cudaFree(p);      // a batch member's arrays live in the batch's arena
  allocs_.clear(); allocRecs_.clear();
  if (hostScalars_) { if (batch_) free(hostScalars_); else cudaFreeHost(hostScalars_); }
  if (batch_) hJointColor_ = nullptr;                    // the batch's mapped int
  if (hJointColor_) { cudaFreeHost(hJointColor_); hJointColor_ = nullptr; hJointColorDev_ = nullptr; }
  if (hBegin_) cudaFreeHost(hBegin_);                   // (null for a batch member: its events stay in device memory)
  if (hSensorBegin_) { cudaFreeHost(hSensorBegin_); hSensorBegin_ = nullptr; }
  for (auto& e : hSensorEnd_) { if (e) cudaFreeHost(e); e = nullptr; }
  for (auto& e : hEnd_) { if (e) cudaFreeHost(e); e = nullptr; }
  if (hHit_) cudaFreeHost(hHit_);
  if (hJointEvents_) cudaFreeHost(hJointEvents_);
  if (hBodyRead_) cudaFreeHost(hBodyRead_);
  if (hMoveEvents_) cudaFreeHost(hMoveEvents_);
  hMoveEvents_ = nullptr;
  hBodyRead_ = nullptr;
  hostScalars_ = nullptr; hBegin_ = nullptr; hHit_ = nullptr; hJointEvents_ = nullptr;
  for (auto& g : graphs_) if (g.exec) cudaGraphExecDestroy(g.exec);
  graphs_.clear();
  for (auto& e : timer_) if (e) { cudaEventDestroy(e); e = nullptr; }
  if (stream_ && !batch_) cudaStreamDestroy(stream_);
  stream_ = nullptr;
}

void World::bufferMove(int key) {                 // src/BroadPhase.cs:56-59
  int type = key & 3, id = key >> 2;
  auto& m = moved_[type];
  if ((int)m.size() <= id) m.resize(id + 1, 0);
  if (!m[id]) { m[id] = 1; moveArray_.push_back(key); }
}

// ---- creation after the upload (BodyAPI.CreateBody api/Body.cs:16-105, ShapeAPI.CreateShape api/Shape.cs:8-24) -----------------------------
// The host allocates the id from its LIFO pool and fills its mirror as before; one thread then writes the body's rows (appended to the
// awake rows for a moving body) or creates the shape's proxy with the tree code the host used before the upload — free-list order included:
// the step's tree rebuild hands the doomed internal nodes straight to the new internal nodes and never touches the free list, so the
// node a new leaf receives is the one the reference's tree would give it.
struct BodyRows {
  float4 st0, st1, simMass, simForce, simPose, simLocal, simPose0, bodyXf, bodyMass; float2 bodyCenter;
  uint32_t flags; int type, sim; float sleepThreshold;
};
__global__ void k_create_body(WorldDev d, int id, BodyRows r) {
  d.bodyXf[id] = r.bodyXf; d.bodyMass[id] = r.bodyMass; d.bodyCenter[id] = r.bodyCenter; d.bodySim[id] = r.sim; d.bodyType[id] = r.type;
  d.bodyHeadShape[id] = -1; d.bodyHeadJoint[id] = -1; d.bodyJointCount[id] = 0; d.bodyColorMask[id] = 0; d.bodyMinOp[id] = 0x7f7f7f7f;
  d.bodySleepThreshold[id] = r.sleepThreshold; d.bodyFlags[id] = r.flags; d.bodySimFlags[id] = r.flags; d.bodyForce[id] = make_float4(0, 0, 0, 0);
  if (r.sim < 0) return;
  const int i = r.sim;
  d.st0[i] = r.st0; d.st1[i] = r.st1; d.simMass[i] = r.simMass; d.simForce[i] = r.simForce; d.simPose[i] = r.simPose; d.simLocal[i] = r.simLocal;
  d.simPose0[i] = r.simPose0; d.simFlags[i] = r.flags; d.simBodyId[i] = id;
  d.sc->graphDirty = 1;
}
// growFrom/growTo: AllocateNode's growth branch (src/DynamicTree.cs:56-66) decided by the host, which knows node count and capacity
__device__ void treeGrowSerial(const TreeDev& t, int from, int to) {
  if (from >= to) return;
  for (int i = from; i < to - 1; i++) t.parent[i] = i + 1;
  t.parent[to - 1] = -1;
  t.meta[2] = from;
}
__device__ void destroyProxySerial(const WorldDev& d, int sid);
// BodyAPI.SetType api/Body.cs:185-268 for a body without joints. k_body_retype: DestroyBodyContacts, Body.type / Body.flags, TransferBody
// (src/SolverSet.cs:354-376: awake rows appended with an identity state, or closed by swap-remove). k_shape_retype, once per shape in list
// order: DestroyProxy in the old tree, UpdateAABBs with the new type's margin, CreateProxy (forced move) in the new tree.
__global__ void k_shape_retype(WorldDev d, int shapeId, int type, int grow0From, int grow0To, int grow1From, int grow1To) {
  Scalars& s = *d.sc;
  destroyProxySerial(d, shapeId);
  ShapeDev& sh = d.shapes[shapeId];
  float4 xf4 = d.bodyXf[sh.bodyId];
  Xf xf = mkXf(mk(xf4.x, xf4.y), mkRot(xf4.z, xf4.w));
  Box aabb = boxGrow(geomAABB(sh.geom, xf), kSpeculativeDistance);
  Box fat = boxGrow(aabb, type == kStaticBody ? kSpeculativeDistance : kAABBMargin);
  const TreeDev& t = d.trees[type];
  TreeView v{t.aabb, t.child, t.parent, t.hf, t.cat, t.user, t.meta};
  treeGrowSerial(t, grow0From, grow0To);
  int id = tvAllocNode(v);
  if (id < 0) { atomicOr(&s.errorFlags, (int)ERR_TREE); return; }
  tvSetBox(v, id, fat); t.user[id] = shapeId; t.cat[id] = sh.categoryBits; t.hf[id] = tvPackHF(0, NODE_ALLOCATED | NODE_LEAF); t.moved[id] = 0;
  treeGrowSerial(t, grow1From, grow1To);
  tvInsertLeaf(v, id, true); t.meta[3]++;
  const int key = (id << 2) | type;
  if (s.moveCount < d.moveCapacity) { t.moved[id] = 1; d.moveArray[s.moveCount++] = key; } else atomicOr(&s.errorFlags, (int)ERR_PAIR_CAPACITY);
  if (type == kStaticBody) s.staticRanksValid = 0;
  sh.proxyKey = key; sh.flags &= ~SF_ENLARGED;
  d.shapeAABB[shapeId] = make_float4(aabb.lo.x, aabb.lo.y, aabb.hi.x, aabb.hi.y);
  d.shapeFat[shapeId] = make_float4(fat.lo.x, fat.lo.y, fat.hi.x, fat.hi.y);
}
__device__ void destroyContactsOf(const WorldDev& d, int bodyId, int shapeId, int endIndex, int* sCount);
__global__ void k_body_retype(WorldDev d, int bodyId, int type, int transfer, int newSim, uint32_t bodyFlags, uint32_t simFlags, float linearDamping, float angularDamping, float gravityScale, int endIndex) {
  __shared__ int sCount;
  destroyContactsOf(d, bodyId, -1, endIndex, &sCount);
  if (threadIdx.x != 0) return;
  d.bodyType[bodyId] = type; d.bodyFlags[bodyId] = bodyFlags;
  const int sim = d.bodySim[bodyId];
  if (transfer == 1 && sim >= 0) {                 // awake -> static: bodySims / bodyStates RemoveSwap; the host lowers awakeCount after this launch
    d.bodySimFlags[bodyId] = d.simFlags[sim] & ~(uint32_t)(kIsFast | kIsSpeedCapped | kHadTimeOfImpact);
    d.bodyForce[bodyId] = d.simForce[sim];
    int last = d.awakeCount - 1;
    if (sim != last) {
      d.st0[sim] = d.st0[last]; d.st1[sim] = d.st1[last]; d.simMass[sim] = d.simMass[last]; d.simForce[sim] = d.simForce[last];
      d.simPose[sim] = d.simPose[last]; d.simLocal[sim] = d.simLocal[last]; d.simPose0[sim] = d.simPose0[last]; d.simFlags[sim] = d.simFlags[last];
      int movedBody = d.simBodyId[last]; d.simBodyId[sim] = movedBody; d.bodySim[movedBody] = sim;
    }
    d.bodySim[bodyId] = -1;
  } else if (transfer == 2) {                      // static -> awake: the sim is copied, the state starts as identity with Body.flags
    float4 xf4 = d.bodyXf[bodyId], m = d.bodyMass[bodyId]; float2 c = d.bodyCenter[bodyId];
    const int i = newSim;
    d.st0[i] = make_float4(0, 0, 0, __uint_as_float(bodyFlags)); d.st1[i] = make_float4(0, 0, 1, 0);
    d.simMass[i] = make_float4(m.x, m.y, linearDamping, angularDamping); { float4 f = d.bodyForce[bodyId]; d.simForce[i] = make_float4(f.x, f.y, f.z, gravityScale); }
    d.simPose[i] = make_float4(c.x, c.y, xf4.z, xf4.w); d.simPose0[i] = d.simPose[i];
    d.simLocal[i] = make_float4(m.z, m.w, kHuge, 0);
    d.simFlags[i] = d.bodySimFlags[bodyId] & ~(uint32_t)(kIsFast | kIsSpeedCapped | kHadTimeOfImpact); d.simBodyId[i] = bodyId; d.bodySim[bodyId] = i;
  } else if (sim >= 0) {                           // dynamic <-> kinematic: `state->flags = body.flags` (:264-265)
    float4 s0 = d.st0[sim]; s0.w = __uint_as_float(bodyFlags); d.st0[sim] = s0;
  }
  d.sc->graphDirty = 1;
}
__global__ void k_create_shape(WorldDev d, int shapeId, ShapeDev row, int type, int forceMove, int grow0From, int grow0To, int grow1From, int grow1To) {
  Scalars& s = *d.sc;
  const int bodyId = row.bodyId;
  float4 xf4 = d.bodyXf[bodyId];
  Xf xf = mkXf(mk(xf4.x, xf4.y), mkRot(xf4.z, xf4.w));
  Box aabb = boxGrow(geomAABB(row.geom, xf), kSpeculativeDistance);                       // Shape.UpdateShapeAABBs src/Shape.cs:64-75
  Box fat = boxGrow(aabb, type == kStaticBody ? kSpeculativeDistance : kAABBMargin);
  const TreeDev& t = d.trees[type];
  TreeView v{t.aabb, t.child, t.parent, t.hf, t.cat, t.user, t.meta};
  treeGrowSerial(t, grow0From, grow0To);
  int id = tvAllocNode(v);                                                                 // DynamicTree.CreateProxy src/DynamicTree.cs:443-458
  if (id < 0) { atomicOr(&s.errorFlags, (int)ERR_TREE); return; }
  tvSetBox(v, id, fat); t.user[id] = shapeId; t.cat[id] = row.categoryBits; t.hf[id] = tvPackHF(0, NODE_ALLOCATED | NODE_LEAF); t.moved[id] = 0;
  treeGrowSerial(t, grow1From, grow1To);
  tvInsertLeaf(v, id, true); t.meta[3]++;
  const int key = (id << 2) | type;
  if (type != kStaticBody || forceMove) {                                                 // BroadPhase.CreateProxy src/BroadPhase.cs:71-79 -> BufferMove :56-59
    if (s.moveCount < d.moveCapacity) { t.moved[id] = 1; d.moveArray[s.moveCount++] = key; } else atomicOr(&s.errorFlags, (int)ERR_PAIR_CAPACITY);
  }
  if (type == kStaticBody) s.staticRanksValid = 0;
  row.proxyKey = key; row.nextShapeId = d.bodyHeadShape[bodyId];
  d.shapes[shapeId] = row; d.bodyHeadShape[bodyId] = shapeId;
  d.shapeAABB[shapeId] = make_float4(aabb.lo.x, aabb.lo.y, aabb.hi.x, aabb.hi.y);
  d.shapeFat[shapeId] = make_float4(fat.lo.x, fat.lo.y, fat.hi.x, fat.hi.y);
}
// World.ResetProxy src/Shape.cs:273-301 behind ShapeAPI.SetFilter (api/Shape.cs:157-166) and SetCircle/SetCapsule/SetSegment/SetPolygon
// (:289-326): the shape's contacts go (contact-list order), the AABBs are recomputed from the body's current transform, and the proxy is
// either re-created (the leaf and its parent come straight back from the LIFO free list, the move is buffered) or moved.
__device__ void destroyContactsOf(const WorldDev& d, int bodyId, int shapeId, int endIndex, int* sCount);
__global__ void k_reset_proxy(WorldDev d, int shapeId, int setGeom, Geom geom, int setFilter, uint64_t categoryBits, uint64_t maskBits, int groupIndex, int destroyProxy, int endIndex) {
  __shared__ int sCount;
  Scalars& s = *d.sc;
  const int bodyId = d.shapes[shapeId].bodyId;
  destroyContactsOf(d, bodyId, shapeId, endIndex, &sCount);
  if (threadIdx.x != 0) return;
  ShapeDev& sh = d.shapes[shapeId];
  if (setGeom) sh.geom = geom;
  if (setFilter) { sh.categoryBits = categoryBits; sh.maskBits = maskBits; sh.groupIndex = groupIndex; }
  float4 xf4 = d.bodyXf[bodyId];
  Xf xf = mkXf(mk(xf4.x, xf4.y), mkRot(xf4.z, xf4.w));
  const int key = sh.proxyKey;
  const int type = key != -1 ? (key & 3) : d.bodyType[bodyId];
  Box aabb = boxGrow(geomAABB(sh.geom, xf), kSpeculativeDistance);
  Box fat = boxGrow(aabb, type == kStaticBody ? kSpeculativeDistance : kAABBMargin);
  d.shapeAABB[shapeId] = make_float4(aabb.lo.x, aabb.lo.y, aabb.hi.x, aabb.hi.y);
  d.shapeFat[shapeId] = make_float4(fat.lo.x, fat.lo.y, fat.hi.x, fat.hi.y);
  sh.flags &= ~SF_ENLARGED;
  if (key == -1) return;
  const TreeDev& t = d.trees[type];
  TreeView v{t.aabb, t.child, t.parent, t.hf, t.cat, t.user, t.meta};
  int proxyId = key >> 2;
  if (destroyProxy) {
    if (t.moved[proxyId]) {                       // UnBufferMove
      t.moved[proxyId] = 0;
      for (int i = 0; i < s.moveCount; i++) if (d.moveArray[i] == key) { d.moveArray[i] = d.moveArray[s.moveCount - 1]; s.moveCount--; break; }
    }
    tvRemoveLeaf(v, proxyId); tvFreeNode(v, proxyId); t.meta[3]--;
    proxyId = tvAllocNode(v);
    tvSetBox(v, proxyId, fat); t.user[proxyId] = shapeId; t.cat[proxyId] = sh.categoryBits; t.hf[proxyId] = tvPackHF(0, NODE_ALLOCATED | NODE_LEAF); t.moved[proxyId] = 0;
    tvInsertLeaf(v, proxyId, true); t.meta[3]++;
    sh.proxyKey = (proxyId << 2) | type;
  } else tvMoveProxy(v, proxyId, fat);
  if (!t.moved[proxyId]) {                        // BufferMove (forced for the new proxy, src/BroadPhase.cs:76-77; always after MoveProxy :94)
    if (s.moveCount < d.moveCapacity) { t.moved[proxyId] = 1; d.moveArray[s.moveCount++] = sh.proxyKey; } else atomicOr(&s.errorFlags, (int)ERR_PAIR_CAPACITY);
  }
  if (type == kStaticBody) s.staticRanksValid = 0;
}
int World::resetProxy(int shapeId, const Geom* geom, bool setFilter, bool destroyProxy) {
  HShape& sh = shapes_[shapeId];
  Geom g; memset(&g, 0, sizeof(g));
  if (geom) g = *geom;
  CK(cudaSetDevice(device_));
  k_reset_proxy<<<1, 256, 0, stream_>>>(d_, shapeId, geom ? 1 : 0, g, setFilter ? 1 : 0, sh.categoryBits, sh.maskBits, sh.groupIndex, destroyProxy ? 1 : 0, endIndex_);
  CK(cudaGetLastError());
  eventsFetched_ = false;
  return B2C_OK;
}
int World::setShapeFilter(int shapeId, uint64_t categoryBits, uint64_t maskBits, int groupIndex) {   // ShapeAPI.SetFilter api/Shape.cs:157-166
  if (shapeId < 0 || shapeId >= (int)shapes_.size() || shapes_[shapeId].id != shapeId) return fail(B2C_E_INVALID, "bad shape id");
  HShape& sh = shapes_[shapeId];
  if (maskBits == sh.maskBits && categoryBits == sh.categoryBits && groupIndex == sh.groupIndex) return B2C_OK;
  if (!onDevice_) { int rc = upload(); if (rc) return rc; }   // the first call that needs the device state uploads the world (as the first step would)
  const bool destroyProxy = categoryBits != sh.categoryBits;
  sh.categoryBits = categoryBits; sh.maskBits = maskBits; sh.groupIndex = groupIndex;
  return resetProxy(shapeId, nullptr, true, destroyProxy);
}
int World::setShapeGeometry(int shapeId, const B2CGeometry& g) {  // ShapeAPI.SetCircle / SetCapsule / SetSegment / SetPolygon api/Shape.cs:289-326
  if (shapeId < 0 || shapeId >= (int)shapes_.size() || shapes_[shapeId].id != shapeId) return fail(B2C_E_INVALID, "bad shape id");
  if (g.type < 0 || g.type > kPolygon) return fail(B2C_E_UNSUPPORTED, "chain segments are not supported");
  if (g.type == kPolygon && (g.count < 1 || g.count > kMaxPolygonVertices)) return fail(B2C_E_UNSUPPORTED, "polygon vertex count must be 1..8");
  if (!onDevice_) { int rc = upload(); if (rc) return rc; }   // the first call that needs the device state uploads the world (as the first step would)
  HShape& sh = shapes_[shapeId];
  memset(&sh.geom, 0, sizeof(Geom));
  sh.geom.type = g.type; sh.geom.count = g.count; sh.geom.radius = g.radius; sh.geom.centroid = mk(g.centroid.x, g.centroid.y);
  for (int i = 0; i < kMaxPolygonVertices; i++) { sh.geom.v[i] = mk(g.v[i].x, g.v[i].y); sh.geom.n[i] = mk(g.n[i].x, g.n[i].y); }
  return resetProxy(shapeId, &sh.geom, false, true);
}
__global__ void k_body_mass_update(WorldDev d, int bodyId, int type, float invMass, float invInertia, V2 localCenter, float minExtent, float maxExtent);
__global__ void k_joint_color(WorldDev d, int bodyIdA, int bodyIdB, int typeA, int typeB, int* out);
__global__ void k_joint_set_color(WorldDev d, int jointId, int color);
__global__ void k_joint_removed(WorldDev d, int bodyIdA, int bodyIdB, int color);
int World::setBodyType(int bodyId, int type) {   // BodyAPI.SetType api/Body.cs:185-268
  if (bodyId < 0 || bodyId >= (int)bodies_.size() || bodies_[bodyId].id != bodyId) return fail(B2C_E_INVALID, "bad body id");
  if (type < kStaticBody || type > kDynamicBody) return fail(B2C_E_INVALID, "bad body type");
  HBody& b = bodies_[bodyId];
  const int oldType = b.type;
  if (oldType == type) return B2C_OK;
  if (!onDevice_) { int rc = upload(); if (rc) return rc; }   // the first call that needs the device state uploads the world (as the first step would)
  std::vector<int> jids;                           // the body's joints in list order
  for (int key = b.headJointKey; key != -1;) { const HJoint& j = joints_[key >> 1]; jids.push_back(key >> 1); key = (key & 1) ? j.nextKeyB : j.nextKeyA; }
  const int transfer = type == kStaticBody ? 1 : (oldType == kStaticBody ? 2 : 0);
  if (transfer == 2) { int rc = growBodies(0, (int)awakeOrder_.size() + 1); if (rc) return rc; }
  // growth of the target tree, shape by shape, before anything changes (HostTree::createProxy: ensureFreeNode before each allocation)
  std::vector<int> sids; std::vector<int> grows;
  { int cap = d_.trees[type].capacity, nodes = treeNodeCount_[type], proxies = treeProxyCount_[type];
    for (int sid = b.headShapeId; sid != -1; sid = shapes_[sid].nextShapeId) {
      int g[4] = {0, 0, 0, 0};
      if (nodes == cap) { g[0] = cap; cap += cap >> 1; g[1] = cap; }
      if (nodes + 1 == cap) { g[2] = cap; cap += cap >> 1; g[3] = cap; }
      nodes += proxies > 0 ? 2 : 1; proxies++;
      sids.push_back(sid); grows.insert(grows.end(), g, g + 4);
    }
    int rc = growTree(type, cap); if (rc) return rc;
  }
  CK(cudaSetDevice(device_));
  for (int jid : jids) {                            // :202-216 TransferJoint(staticSet <- graph): colour bits and the colour array give the joint up
    HJoint& j = joints_[jid];
    if (!j.inGraph) continue;
    const int color = j.sim.colorIndex;
    std::vector<int>& arr = colorJoints_[color];
    for (size_t k = 0; k < arr.size(); k++) if (arr[k] == jid) { arr[k] = arr.back(); arr.pop_back(); break; }
    k_joint_removed<<<1, 1, 0, stream_>>>(d_, j.bodyIdA, j.bodyIdB, color);
    CK(cudaGetLastError());
    j.inGraph = false; j.sim.colorIndex = -1;
  }
  b.type = type;
  if (type == kDynamicBody) b.bodyFlags |= kDynamicFlag; else b.bodyFlags &= ~(uint32_t)kDynamicFlag;
  const uint32_t simFlags = b.flags & ~(uint32_t)(kIsFast | kIsSpeedCapped | kHadTimeOfImpact);
  const int newSim = transfer == 2 ? (int)awakeOrder_.size() : -1;
  CK(cudaSetDevice(device_));
  k_body_retype<<<1, 256, 0, stream_>>>(d_, bodyId, type, transfer, newSim, b.bodyFlags, simFlags, b.linearDamping, b.angularDamping, b.gravityScale, endIndex_);
  CK(cudaGetLastError());
  if (transfer == 1) {
    int last = (int)awakeOrder_.size() - 1;
    if (b.simIndex != last) { int moved = awakeOrder_[last]; awakeOrder_[b.simIndex] = moved; bodies_[moved].simIndex = b.simIndex; }
    awakeOrder_.pop_back(); b.simIndex = -1;
  } else if (transfer == 2) { b.simIndex = newSim; awakeOrder_.push_back(bodyId); }
  d_.awakeCount = (int)awakeOrder_.size();
  for (int jid : jids) {                            // :226-241 back into the graph when either body is dynamic now (fresh colour, appended)
    HJoint& j = joints_[jid];
    const HBody& bodyA = bodies_[j.bodyIdA]; const HBody& bodyB = bodies_[j.bodyIdB];
    if (bodyA.type != kDynamicBody && bodyB.type != kDynamicBody) continue;
    k_joint_color<<<1, 1, 0, stream_>>>(d_, bodyA.id, bodyB.id, bodyA.type, bodyB.type, hJointColorDev_);
    CK(cudaGetLastError());
    CK(cudaStreamSynchronize(stream_));
    const int color = *hJointColor_;
    colorJoints_[color].push_back(jid);
    j.inGraph = true; j.sim.colorIndex = color;
    k_joint_set_color<<<1, 1, 0, stream_>>>(d_, jid, color);
    CK(cudaGetLastError());
  }
  if (!jids.empty()) { int rc = uploadJointLists(); if (rc) return rc; }
  for (size_t k = 0; k < sids.size(); k++) {
    const int* g = &grows[4 * k];
    k_shape_retype<<<1, 1, 0, stream_>>>(d_, sids[k], type, g[0], g[1], g[2], g[3]);
    CK(cudaGetLastError());
    if (g[1] > d_.trees[type].capacity) d_.trees[type].capacity = g[1];
    if (g[3] > d_.trees[type].capacity) d_.trees[type].capacity = g[3];
    if (shapes_[sids[k]].proxyKey != -1) treeProxyRemoved(oldType);
    treeNodeCount_[type] += treeProxyCount_[type] > 0 ? 2 : 1; treeProxyCount_[type]++;
    shapes_[sids[k]].proxyKey = 0;
  }
  updateBodyMassData(b);                           // UpdateBodyMassData src/Body.cs:255-353 with the new type
  k_body_mass_update<<<1, 1, 0, stream_>>>(d_, bodyId, b.type, b.invMass, b.invInertia, b.localCenter, b.minExtent, b.maxExtent);
  CK(cudaGetLastError());
  eventsFetched_ = false; bodiesFetched_ = false; movesFetched_ = false;
  return B2C_OK;
}
void World::treeProxyRemoved(int type) {          // DestroyProxy: the leaf and, unless it was the root, its parent go back to the free list
  if (!onDevice_) return;
  treeNodeCount_[type] -= treeProxyCount_[type] > 1 ? 2 : 1;
  treeProxyCount_[type]--;
}

int World::createBody(const B2CBodyDef& def, int* bodyId) {
  if (onDevice_) {
    int rc = growBodies(bodyIds_.freeArray.empty() ? (int)bodies_.size() + 1 : 0, def.type != kStaticBody ? (int)awakeOrder_.size() + 1 : 0);
    if (rc) return rc;
  }
  if (!def.isEnabled || !(def.isAwake || !def.enableSleep))
    return fail(B2C_E_UNSUPPORTED, "disabled / initially-asleep bodies are not supported");
  int id = bodyIds_.alloc();
  if (id == (int)bodies_.size()) bodies_.emplace_back();
  HBody& b = bodies_[id];
  uint16_t gen = b.generation;
  b = HBody();
  b.generation = (uint16_t)(gen + 1);
  b.id = id; b.type = def.type; b.userData = def.userData; b.sleepThreshold = def.sleepThreshold; b.enableSleep = def.enableSleep != 0;
  b.flags = (def.lockLinearX ? kLockLinearX : 0) | (def.lockLinearY ? kLockLinearY : 0) | (def.lockAngularZ ? kLockAngularZ : 0) |
            (def.isBullet ? kIsBullet : 0) | (def.allowFastRotation ? kAllowFastRotation : 0) | (def.type == kDynamicBody ? kDynamicFlag : 0);   // api/Body.cs:42-62
  b.xf.p = mk(def.position.x, def.position.y); b.xf.q = mkRot(def.rotation.c, def.rotation.s);
  b.center = b.xf.p; b.center0 = b.xf.p; b.rot0 = b.xf.q; b.localCenter = mk(0, 0); b.force = mk(0, 0);
  b.linearDamping = def.linearDamping; b.angularDamping = def.angularDamping; b.gravityScale = def.gravityScale;
  b.v = mk(def.linearVelocity.x, def.linearVelocity.y); b.w = def.angularVelocity;
  b.bodyFlags = b.flags;
  if (def.type != kStaticBody) { b.simIndex = (int)awakeOrder_.size(); awakeOrder_.push_back(id); }
  if ((int)colorMask_.size() <= id) colorMask_.resize(id + 1, 0);
  colorMask_[id] = 0;
  *bodyId = id;
  if (!onDevice_) return B2C_OK;
  BodyRows r;
  const uint32_t fl = b.flags;
  r.st0 = make_float4(b.v.x, b.v.y, b.w, __builtin_bit_cast(float, fl)); r.st1 = make_float4(0, 0, 1, 0);
  r.simMass = make_float4(b.invMass, b.invInertia, b.linearDamping, b.angularDamping); r.simForce = make_float4(b.force.x, b.force.y, b.torque, b.gravityScale);
  r.simPose = make_float4(b.center.x, b.center.y, b.xf.q.c, b.xf.q.s); r.simLocal = make_float4(b.localCenter.x, b.localCenter.y, b.minExtent, b.maxExtent);
  r.simPose0 = make_float4(b.center0.x, b.center0.y, b.rot0.c, b.rot0.s);
  r.bodyXf = make_float4(b.xf.p.x, b.xf.p.y, b.xf.q.c, b.xf.q.s); r.bodyMass = make_float4(b.invMass, b.invInertia, b.localCenter.x, b.localCenter.y);
  r.bodyCenter = make_float2(b.center.x, b.center.y); r.flags = fl; r.type = b.type; r.sim = b.simIndex; r.sleepThreshold = b.sleepThreshold;
  CK(cudaSetDevice(device_));
  k_create_body<<<1, 1, 0, stream_>>>(d_, id, r);
  CK(cudaGetLastError());
  d_.awakeCount = (int)awakeOrder_.size(); d_.bodyCapacity = (int)bodies_.size();
  bodiesFetched_ = false; movesFetched_ = false;
  return B2C_OK;
}

void World::updateShapeAABBs(HShape& s, Xf xf, int proxyType) {   // src/Shape.cs:64-75
  Box aabb = boxGrow(geomAABB(s.geom, xf), kSpeculativeDistance);
  s.aabb = aabb;
  s.fat = boxGrow(aabb, proxyType == kStaticBody ? kSpeculativeDistance : kAABBMargin);
}

// a sensor created after the upload: its slot in World.sensors starts with two empty overlap lists (src/Shape.cs:212-213)
__global__ void k_sensor_created(WorldDev d, int sensorIndex, int shapeId) {
  d.sensorShape[sensorIndex] = shapeId; d.sensorOverlapCount[2 * sensorIndex] = 0; d.sensorOverlapCount[2 * sensorIndex + 1] = 0; d.sensorHitCount[sensorIndex] = 0;
}
int World::createShape(int bodyId, const B2CShapeDef& def, const B2CGeometry& g, int* shapeId) {
  if (bodyId < 0 || bodyId >= (int)bodies_.size() || bodies_[bodyId].id != bodyId) return fail(B2C_E_INVALID, "bad body id");
  int grow[4] = {0, 0, 0, 0}, grownCap = 0;
  if (onDevice_) {                               // every check first: a refused call leaves both sides untouched
    if (shapeIds_.freeArray.empty()) { int rc = growShapes((int)shapes_.size() + 1); if (rc) return rc; }
    // the contact rows keep the reserve of the first upload (12 per shape) as shapes are added, unless the caller fixed it (contactCapacity)
    if (!batch_ && capHintContact <= 0 && 12 * ((int)shapes_.size() + 1) > d_.contactCapacity) { int rc = growContacts(12 * ((int)shapes_.size() + 1)); if (rc) return rc; }
    const int t = bodies_[bodyId].type;
    int cap = d_.trees[t].capacity, nodes = treeNodeCount_[t];
    if (nodes == cap) { grow[0] = cap; cap += cap >> 1; grow[1] = cap; }                  // HostTree::createProxy: ensureFreeNode before each allocation
    if (nodes + 1 == cap) { grow[2] = cap; cap += cap >> 1; grow[3] = cap; }
    { int rc = growTree(t, cap); if (rc) return rc; }
    grownCap = cap;
  }
  if (def.isSensor && onDevice_) { int rc = growSensors((int)sensorShapes_.size() + 1); if (rc) return rc; }
  if (def.enableCustomFiltering || def.enablePreSolveEvents) return fail(B2C_E_UNSUPPORTED, "managed callbacks (custom filter / preSolve) cannot run on the device");
  if (g.type < 0 || g.type > kPolygon) return fail(B2C_E_UNSUPPORTED, "chain segments are not supported");
  if (g.type == kPolygon && (g.count < 1 || g.count > kMaxPolygonVertices)) return fail(B2C_E_UNSUPPORTED, "polygon vertex count must be 1..8");
  HBody& body = bodies_[bodyId];
  int id = shapeIds_.alloc();
  if (id == (int)shapes_.size()) shapes_.emplace_back();
  HShape& s = shapes_[id];
  uint16_t gen = s.generation;
  s = HShape();
  s.generation = (uint16_t)(gen + 1);
  s.id = id; s.bodyId = bodyId;
  memset(&s.geom, 0, sizeof(Geom));
  s.geom.type = g.type; s.geom.count = g.count; s.geom.radius = g.radius; s.geom.centroid = mk(g.centroid.x, g.centroid.y);
  for (int i = 0; i < kMaxPolygonVertices; i++) { s.geom.v[i] = mk(g.v[i].x, g.v[i].y); s.geom.n[i] = mk(g.n[i].x, g.n[i].y); }
  s.friction = def.friction; s.restitution = def.restitution; s.rollingResistance = def.rollingResistance; s.tangentSpeed = def.tangentSpeed;
  s.density = def.density; s.categoryBits = def.categoryBits; s.maskBits = def.maskBits; s.groupIndex = def.groupIndex; s.userData = def.userData;
  s.flags = (def.enableContactEvents ? SF_CONTACT_EVENTS : 0) | (def.enableHitEvents ? SF_HIT_EVENTS : 0) | (def.enableSensorEvents ? SF_SENSOR_EVENTS : 0);
  if (def.isSensor) { s.sensorIndex = (int)sensorShapes_.size(); sensorShapes_.push_back(id); }      // src/Shape.cs:210-215
  const bool forcePairCreation = def.invokeContactCreation || def.isSensor;                       // src/Shape.cs:199
  int proxyType = body.type;
  if (onDevice_) {
    ShapeDev row; memset(&row, 0, sizeof(row));
    row.geom = s.geom; row.material = make_float4(s.friction, s.restitution, s.rollingResistance, s.tangentSpeed);
    row.categoryBits = s.categoryBits; row.maskBits = s.maskBits; row.groupIndex = s.groupIndex; row.bodyId = bodyId; row.flags = s.flags; row.sensorIndex = s.sensorIndex; row.generation = s.generation;
    CK(cudaSetDevice(device_));
    k_create_shape<<<1, 1, 0, stream_>>>(d_, id, row, proxyType, forcePairCreation ? 1 : 0, grow[0], grow[1], grow[2], grow[3]);
    CK(cudaGetLastError());
    if (def.isSensor) { k_sensor_created<<<1, 1, 0, stream_>>>(d_, s.sensorIndex, id); CK(cudaGetLastError()); d_.sensorCount = (int)sensorShapes_.size(); }
    d_.trees[proxyType].capacity = grownCap;
    treeNodeCount_[proxyType] += treeProxyCount_[proxyType] > 0 ? 2 : 1; treeProxyCount_[proxyType]++;
    s.proxyKey = 0;                               // "has a proxy"; the key itself lives on the device (b2cuda_world_read_shapes)
    d_.shapeCapacity = (int)shapes_.size();
    if (body.headShapeId != -1) shapes_[body.headShapeId].prevShapeId = id;
    s.prevShapeId = -1; s.nextShapeId = body.headShapeId; body.headShapeId = id; body.shapeCount++;
    *shapeId = id;
    return def.updateBodyMass ? applyMassFromShapes(bodyId) : B2C_OK;
  }
  updateShapeAABBs(s, body.xf, proxyType);
  {                                               // src/BroadPhase.cs:71-79
    int proxyId = trees_[proxyType].createProxy(s.fat, s.categoryBits, id);
    s.proxyKey = proxyKeyOf(proxyId, proxyType);
    if (proxyType != kStaticBody || forcePairCreation) bufferMove(s.proxyKey);
  }
  if (body.headShapeId != -1) shapes_[body.headShapeId].prevShapeId = id;
  s.prevShapeId = -1; s.nextShapeId = body.headShapeId; body.headShapeId = id; body.shapeCount++;
  if (def.updateBodyMass) updateBodyMassData(body);
  *shapeId = id;
  return B2C_OK;
}

void World::updateBodyMassData(HBody& body) {     // src/Body.cs:255-353
  body.mass = 0; body.inertia = 0; body.invMass = 0; body.invInertia = 0; body.localCenter = mk(0, 0);
  body.minExtent = kHuge; body.maxExtent = 0;
  if (body.type != kDynamicBody) {
    if (!onDevice_) { body.center = body.xf.p; body.center0 = body.center; }
    if (body.type == kKinematicBody) {
      for (int sid = body.headShapeId; sid != -1; sid = shapes_[sid].nextShapeId) {
        float mn, mx; geomExtent(shapes_[sid].geom, mk(0, 0), &mn, &mx);
        body.minExtent = netMin(body.minExtent, mn); body.maxExtent = netMax(body.maxExtent, mx);
      }
    }
    return;
  }
  std::vector<MassData> masses(body.shapeCount);
  for (auto& m : masses) { m.mass = 0; m.center = mk(0, 0); m.rotationalInertia = 0; }
  V2 localCenter = mk(0, 0);
  int shapeIndex = 0;
  for (int sid = body.headShapeId; sid != -1;) {
    const HShape& s = shapes_[sid];
    sid = s.nextShapeId;
    if (s.density == 0.0f) continue;              // :291 (the C# `continue` skips the index increment too)
    MassData md = geomMass(s.geom, s.density);
    body.mass += md.mass;
    localCenter = mulAdd(localCenter, md.mass, md.center);
    masses[shapeIndex] = md;
    shapeIndex++;
  }
  if (body.mass > 0.0f) { body.invMass = 1.0f / body.mass; localCenter = body.invMass * localCenter; }
  for (shapeIndex = 0; shapeIndex < body.shapeCount; shapeIndex++) {
    MassData md = masses[shapeIndex];
    if (md.mass == 0.0f) continue;
    V2 offset = localCenter - md.center;
    float inertia = md.rotationalInertia + md.mass * dot(offset, offset);
    body.inertia += inertia;
  }
  if (body.inertia > 0.0f) body.invInertia = 1.0f / body.inertia; else { body.inertia = 0; body.invInertia = 0; }
  body.localCenter = localCenter;
  if (!onDevice_) {                               // after the upload pose and velocity live on the device: k_body_mass_update
    V2 oldCenter = body.center;
    body.center = xfPoint(body.xf, localCenter);
    body.center0 = body.center;
    body.v = body.v + crossSV(body.w, body.center - oldCenter);
  }
  for (int sid = body.headShapeId; sid != -1; sid = shapes_[sid].nextShapeId) {
    float mn, mx; geomExtent(shapes_[sid].geom, localCenter, &mn, &mx);
    body.minExtent = netMin(body.minExtent, mn); body.maxExtent = netMax(body.maxExtent, mx);
  }
}

int World::assignJointColor(int bodyIdA, int bodyIdB, int typeA, int typeB) {   // src/ConstraintGraph.cs:91-128
  uint32_t& mA = colorMask_[bodyIdA]; uint32_t& mB = colorMask_[bodyIdB];
  if (typeA != kStaticBody && typeB != kStaticBody) {
    for (int i = 0; i < kDynamicColorCount; i++) {
      if (((mA >> i) & 1u) || ((mB >> i) & 1u)) continue;
      if (typeA == kDynamicBody) mA |= 1u << i;
      if (typeB == kDynamicBody) mB |= 1u << i;
      return i;
    }
  } else if (typeA == kDynamicBody) {
    for (int i = kGraphColorCount - 2; i >= 1; i--) { if ((mA >> i) & 1u) continue; mA |= 1u << i; return i; }
  } else if (typeB == kDynamicBody) {
    for (int i = kGraphColorCount - 2; i >= 1; i--) { if ((mB >> i) & 1u) continue; mB |= 1u << i; return i; }
  }
  return kOverflowIndex;
}

// AssignJointColor (src/ConstraintGraph.cs:91-128) against the colour bitsets the steps maintain on the device; the colour comes back
// through a mapped int so that the host can file the joint under it.
__global__ void k_joint_color(WorldDev d, int bodyIdA, int bodyIdB, int typeA, int typeB, int* out) {
  uint32_t mA = d.bodyColorMask[bodyIdA], mB = d.bodyColorMask[bodyIdB];
  int color = kOverflowIndex;
  if (typeA != kStaticBody && typeB != kStaticBody) {
    for (int i = 0; i < kDynamicColorCount; i++) {
      if (((mA >> i) & 1u) || ((mB >> i) & 1u)) continue;
      if (typeA == kDynamicBody) mA |= 1u << i;
      if (typeB == kDynamicBody) mB |= 1u << i;
      color = i; break;
    }
  } else if (typeA == kDynamicBody) {
    for (int i = kGraphColorCount - 2; i >= 1; i--) { if ((mA >> i) & 1u) continue; mA |= 1u << i; color = i; break; }
  } else if (typeB == kDynamicBody) {
    for (int i = kGraphColorCount - 2; i >= 1; i--) { if ((mB >> i) & 1u) continue; mB |= 1u << i; color = i; break; }
  }
  d.bodyColorMask[bodyIdA] = mA; d.bodyColorMask[bodyIdB] = mB;
  *out = color;
}
__global__ void k_joint_set_color(WorldDev d, int jointId, int color) { d.joints[jointId].colorIndex = color; d.jointEventFlag[jointId] = 0; d.sc->graphDirty = 1; }
__device__ void destroyContactSerial(const WorldDev& d, int id, int endIndex);
// The joint's row, its collide flag and, unless the bodies may collide, DestroyContactsBetweenBodies (src/Joint.cs:155-180): both contact
// lists hold the shared contacts in the same relative order (newest first), so the walk over the shorter list is a descending-stamp walk.
__global__ void k_joint_created(WorldDev d, int jointId, JointDev row, int collide, int endIndex) {
  __shared__ int sCount;
  Scalars& s = *d.sc;
  if (threadIdx.x == 0) { sCount = 0; d.joints[jointId] = row; d.jointCollide[jointId] = collide; d.jointEventFlag[jointId] = 0; s.graphDirty = 1; }
  __syncthreads();
  if (collide) return;
  const int n = s.contactNext, a = row.bodyIdA, b = row.bodyIdB;
  for (int i = threadIdx.x; i < n; i += blockDim.x) {
    if (!(d.cFlags[i] & CF_ALIVE)) continue;
    int2 cb = d.cBody[i];
    if ((cb.x == a && cb.y == b) || (cb.x == b && cb.y == a)) { int k = atomicAdd(&sCount, 1); d.opList[k] = i; }
  }
  __syncthreads();
  if (threadIdx.x != 0) return;
  const int count = sCount;
  for (int i = 1; i < count; i++) {
    int id = d.opList[i]; uint32_t st = d.cStamp[id]; int j = i - 1;
    while (j >= 0 && d.cStamp[d.opList[j]] < st) { d.opList[j + 1] = d.opList[j]; j--; }
    d.opList[j + 1] = id;
  }
  for (int i = 0; i < count; i++) destroyContactSerial(d, d.opList[i], endIndex);
}

int World::createRevoluteJoint(const B2CRevoluteJointDef& r, int* jointId) {     // api/Joint.cs:555 -> the generic path
  B2CJointDef def; memset(&def, 0, sizeof(def));
  def.type = B2C_JOINT_REVOLUTE; def.bodyIdA = r.bodyIdA; def.bodyIdB = r.bodyIdB; def.localFrameA = r.localFrameA; def.localFrameB = r.localFrameB;
  def.forceThreshold = r.forceThreshold; def.torqueThreshold = r.torqueThreshold; def.constraintHertz = r.constraintHertz;
  def.constraintDampingRatio = r.constraintDampingRatio; def.drawScale = r.drawScale; def.collideConnected = r.collideConnected; def.userData = r.userData;
  def.enableSpring = r.enableSpring; def.enableLimit = r.enableLimit; def.enableMotor = r.enableMotor; def.hertz = r.hertz; def.dampingRatio = r.dampingRatio;
  def.target = r.targetAngle; def.lower = r.lowerAngle; def.upper = r.upperAngle; def.maxMotor = r.maxMotorTorque; def.motorSpeed = r.motorSpeed;
  return createJoint(def, jointId);
}

// World.CreateJoint src/Joint.cs:181-296 + the per-type Create (DistanceJoint.cs:35-63, FilterJoint.cs:7-17, MotorJoint.cs:33-52,
// PrismaticJoint.cs:33-53, RevoluteJoint.cs:32-56, WeldJoint.cs:24-39, WheelJoint.cs:33-56)
int World::createJoint(const B2CJointDef& def, int* jointId) {
  if (def.type < B2C_JOINT_DISTANCE || def.type > B2C_JOINT_WHEEL) return fail(B2C_E_INVALID, "bad joint type");
  int nb = (int)bodies_.size();
  if (def.bodyIdA < 0 || def.bodyIdA >= nb || def.bodyIdB < 0 || def.bodyIdB >= nb || bodies_[def.bodyIdA].id != def.bodyIdA || bodies_[def.bodyIdB].id != def.bodyIdB) return fail(B2C_E_INVALID, "bad body id");
  HBody& bodyA = bodies_[def.bodyIdA]; HBody& bodyB = bodies_[def.bodyIdB];
  const bool parked = bodyA.type != kDynamicBody && bodyB.type != kDynamicBody;   // src/Joint.cs:242-248: kept in the static set until SetType makes a body dynamic
  if (def.type == B2C_JOINT_DISTANCE && (!(def.length > 0.0f) || !(def.lowerSpringForce <= def.upperSpringForce)))
    return fail(B2C_E_INVALID, "distance joint: length must be > 0 and lowerSpringForce <= upperSpringForce");
  int deviceColor = -1;
  if (onDevice_ && jointIds_.freeArray.empty()) { int rc = growJoints((int)joints_.size() + 1); if (rc) return rc; }
  if (onDevice_ && !parked) {                    // the colour first: a refused call must leave both sides as they were
    CK(cudaSetDevice(device_));
    k_joint_color<<<1, 1, 0, stream_>>>(d_, bodyA.id, bodyB.id, bodyA.type, bodyB.type, hJointColorDev_);
    CK(cudaGetLastError());
    CK(cudaStreamSynchronize(stream_));
    deviceColor = *hJointColor_;
  }
  // a body with joints in every colour it may use sends the next one to the overflow colour (no bit set, src/ConstraintGraph.cs:91-128):
  // it is filed like any other and solved serially before the overflow contacts (src/Joint.cs:377-394)
  const int color = parked ? -1 : (onDevice_ ? deviceColor : assignJointColor(bodyA.id, bodyB.id, bodyA.type, bodyB.type));
  int id = jointIds_.alloc();
  if (id == (int)joints_.size()) joints_.emplace_back();
  HJoint& j = joints_[id];
  uint16_t gen = j.generation;
  j = HJoint();
  j.generation = (uint16_t)(gen + 1);
  j.id = id; j.collideConnected = def.collideConnected != 0;
  j.bodyIdA = bodyA.id; j.prevKeyA = -1; j.nextKeyA = bodyA.headJointKey;      // src/Joint.cs:208-236
  const int keyA = id << 1, keyB = (id << 1) | 1;
  if (bodyA.headJointKey != -1) { HJoint& h = joints_[bodyA.headJointKey >> 1]; if (bodyA.headJointKey & 1) h.prevKeyB = keyA; else h.prevKeyA = keyA; }
  bodyA.headJointKey = keyA; bodyA.jointCount++;
  j.bodyIdB = bodyB.id; j.prevKeyB = -1; j.nextKeyB = bodyB.headJointKey;
  if (bodyB.headJointKey != -1) { HJoint& h = joints_[bodyB.headJointKey >> 1]; if (bodyA.headJointKey & 1) h.prevKeyB = keyB; else h.prevKeyA = keyB; }   // tests bodyA's key, as written (:229)
  bodyB.headJointKey = keyB; bodyB.jointCount++;
  j.inGraph = !parked;
  JointDev& s = j.sim;
  memset(&s, 0, sizeof(s));
  s.jointId = id; s.bodyIdA = bodyA.id; s.bodyIdB = bodyB.id; s.colorIndex = color; s.type = def.type;
  auto T = [](B2CTransform t) { return mkXf(mk(t.p.x, t.p.y), mkRot(t.q.c, t.q.s)); };
  s.localFrameA = T(def.localFrameA); s.localFrameB = T(def.localFrameB);
  s.constraintHertz = def.constraintHertz; s.constraintDampingRatio = def.constraintDampingRatio;
  s.forceThreshold = def.forceThreshold; s.torqueThreshold = def.torqueThreshold;
  s.constraintSoftness.biasRate = 0; s.constraintSoftness.massScale = 1; s.constraintSoftness.impulseScale = 0;
  const bool spring = def.enableSpring != 0, limit = def.enableLimit != 0, motor = def.enableMotor != 0;
  switch (def.type) {
    case B2C_JOINT_REVOLUTE:
      s.hertz = def.hertz; s.dampingRatio = def.dampingRatio; s.targetAngle = netClamp(def.target, -kPi, kPi);
      s.maxMotorTorque = def.maxMotor; s.motorSpeed = def.motorSpeed; s.lowerAngle = def.lower; s.upperAngle = def.upper;
      s.enableSpring = spring; s.enableMotor = motor; s.enableLimit = limit;
      break;
    case B2C_JOINT_DISTANCE:
      s.length = netMax(def.length, kLinearSlop); s.hertz = def.hertz; s.dampingRatio = def.dampingRatio;
      s.minLength = netMax(def.minLength, kLinearSlop); s.maxLength = netMax(def.minLength, def.maxLength);
      s.maxMotorTorque = def.maxMotor; s.motorSpeed = def.motorSpeed; s.lowerSpringForce = def.lowerSpringForce; s.upperSpringForce = def.upperSpringForce;
      s.enableSpring = spring; s.enableMotor = motor; s.enableLimit = limit;
      break;
    case B2C_JOINT_FILTER: break;
    case B2C_JOINT_MOTOR:
      s.linearVelocity = mk(def.linearVelocity.x, def.linearVelocity.y); s.maxVelocityForce = def.maxVelocityForce; s.angularVelocity = def.angularVelocity;
      s.maxVelocityTorque = def.maxVelocityTorque; s.hertz = def.hertz; s.dampingRatio = def.dampingRatio; s.maxSpringForce = def.maxSpringForce;
      s.angularHertz = def.angularHertz; s.angularDampingRatio = def.angularDampingRatio; s.maxSpringTorque = def.maxSpringTorque;
      break;
    case B2C_JOINT_PRISMATIC:
      s.hertz = def.hertz; s.dampingRatio = def.dampingRatio; s.targetAngle = def.target; s.lowerAngle = def.lower; s.upperAngle = def.upper;
      s.maxMotorTorque = def.maxMotor; s.motorSpeed = def.motorSpeed; s.enableSpring = spring; s.enableMotor = motor; s.enableLimit = limit;
      break;
    case B2C_JOINT_WELD:
      s.hertz = def.hertz; s.dampingRatio = def.dampingRatio; s.angularHertz = def.angularHertz; s.angularDampingRatio = def.angularDampingRatio;
      break;
    case B2C_JOINT_WHEEL:
      s.lowerAngle = def.lower; s.upperAngle = def.upper; s.maxMotorTorque = def.maxMotor; s.motorSpeed = def.motorSpeed;
      s.hertz = def.hertz; s.dampingRatio = def.dampingRatio; s.enableSpring = spring; s.enableMotor = motor; s.enableLimit = limit;
      break;
  }
  if (!parked) colorJoints_[color].push_back(id);
  *jointId = id;
  if (!onDevice_) return B2C_OK;                  // DestroyContactsBetweenBodies (src/Joint.cs:155-180): no contacts exist before the first step
  CK(cudaSetDevice(device_));
  k_joint_created<<<1, 256, 0, stream_>>>(d_, id, s, j.collideConnected ? 1 : 0, endIndex_);
  CK(cudaGetLastError());
  d_.jointCapacity = (int)joints_.size();
  eventsFetched_ = false;
  return uploadJointLists();
}

// ---- JointAPI.DestroyJoint api/Joint.cs:9-17 / BodyAPI.DestroyBody api/Body.cs:106-160 -------------------------------------------------
// The host mirror owns what only creation and destruction change (id pools, body shape/joint lists, joint colour arrays, the awake-set
// order); the device owns what the step changes (contacts, colour bitsets, trees, move array, body rows). A destroy call edits the host
// part as the reference writes it, then patches the device part with a small kernel on the world's stream.
__global__ void k_joint_removed(WorldDev d, int bodyIdA, int bodyIdB, int color) {      // RemoveJointFromGraph src/ConstraintGraph.cs:256-264
  if (color != kOverflowIndex) { d.bodyColorMask[bodyIdA] &= ~(1u << color); d.bodyColorMask[bodyIdB] &= ~(1u << color); }
  d.sc->graphDirty = 1;                           // islands follow the joints
}
__device__ void destroyContactSerial(const WorldDev& d, int id, int endIndex) {          // World.DestroyContact src/Contact.cs:228-312
  Scalars& s = *d.sc;
  uint32_t flags = d.cFlags[id];
  int2 sh = d.cShape[id], b = d.cBody[id];
  bool touching = (flags & CF_TOUCHING) != 0;
  if (touching && (flags & CF_ENABLE_CONTACT_EVENTS)) {
    int k = s.endCount[endIndex];
    if (k < d.eventCapacity) { B2CContactTouchEvent ev = {sh.x, sh.y, id, d.cGen[id]}; d.endEvents[endIndex][k] = ev; s.endCount[endIndex] = k + 1; }
  }
  int color = d.cColor[id];
  if (color != -1) {                              // RemoveContactFromGraph src/ConstraintGraph.cs:221-241
    if (color != kOverflowIndex) { d.bodyColorMask[b.x] &= ~(1u << color); d.bodyColorMask[b.y] &= ~(1u << color); }
    colRemoveSwap(d, color, id);
    d.cColor[id] = -1;
  }
  d.cFlags[id] = 0;
  d.contactFree[s.contactFreeCount++] = id;       // IDPool.FreeId src/IDPool.cs
  s.liveContactCount--;
  s.graphDirty = 1;
}
// One CTA: every thread looks for the body's contacts, thread 0 then walks them in the order of the body's contact list (newest first),
// removes the shapes' proxies and takes the body out of the awake set by swap-remove (api/Body.cs:142-153).
// Collects the live contacts of a body (shapeId < 0) or of one of its shapes and destroys them in the order of the body's contact list.
__device__ void destroyContactsOf(const WorldDev& d, int bodyId, int shapeId, int endIndex, int* sCount) {
  Scalars& s = *d.sc;
  if (threadIdx.x == 0) *sCount = 0;
  __syncthreads();
  const int n = s.contactNext;
  for (int i = threadIdx.x; i < n; i += blockDim.x) {
    if (!(d.cFlags[i] & CF_ALIVE)) continue;
    bool hit;
    if (shapeId < 0) { int2 b = d.cBody[i]; hit = b.x == bodyId || b.y == bodyId; }
    else { int2 sh = d.cShape[i]; hit = sh.x == shapeId || sh.y == shapeId; }
    if (hit) { int k = atomicAdd(sCount, 1); d.opList[k] = i; }
  }
  __syncthreads();
  if (threadIdx.x != 0) return;
  const int count = *sCount;
  for (int i = 1; i < count; i++) {               // insertion sort, descending creation stamp (a body has tens of contacts)
    int id = d.opList[i]; uint32_t st = d.cStamp[id]; int j = i - 1;
    while (j >= 0 && d.cStamp[d.opList[j]] < st) { d.opList[j + 1] = d.opList[j]; j--; }
    d.opList[j + 1] = id;
  }
  for (int i = 0; i < count; i++) destroyContactSerial(d, d.opList[i], endIndex);     // DestroyBodyContacts src/Body.cs:232-244, src/Shape.cs:235-244
}
__device__ void destroyProxySerial(const WorldDev& d, int sid) {      // Shape.DestroyProxy src/Shape.cs:128-131 -> BroadPhase.DestroyProxy src/BroadPhase.cs:80-86
  Scalars& s = *d.sc;
  int key = d.shapes[sid].proxyKey;
  if (key == -1) return;
  const TreeDev& t = d.trees[key & 3];
  const int proxyId = key >> 2;
  if (t.moved[proxyId]) {                         // BroadPhase.UnBufferMove src/BroadPhase.cs:60-71
    t.moved[proxyId] = 0;
    for (int i = 0; i < s.moveCount; i++) if (d.moveArray[i] == key) { d.moveArray[i] = d.moveArray[s.moveCount - 1]; s.moveCount--; break; }
  }
  TreeView v{t.aabb, t.child, t.parent, t.hf, t.cat, t.user, t.meta};
  tvRemoveLeaf(v, proxyId); tvFreeNode(v, proxyId); t.meta[3]--;         // DynamicTree.DestroyProxy src/DynamicTree.cs:460-468
  if ((key & 3) == kStaticBody) s.staticRanksValid = 0;
  d.shapes[sid].proxyKey = -1;
  d.shapeAABB[sid] = make_float4(0, 0, 0, 0); d.shapeFat[sid] = make_float4(0, 0, 0, 0);
}
__global__ void k_destroy_shape(WorldDev d, int shapeId, int bodyId, int prevShapeId, int nextShapeId, int endIndex) {   // DestroyShapeInternal src/Shape.cs:219-272
  __shared__ int sCount;
  destroyContactsOf(d, bodyId, shapeId, endIndex, &sCount);
  if (threadIdx.x != 0) return;
  destroyProxySerial(d, shapeId);
  if (prevShapeId != -1) d.shapes[prevShapeId].nextShapeId = nextShapeId;
  else d.bodyHeadShape[bodyId] = nextShapeId;
}
__global__ void k_destroy_body(WorldDev d, int bodyId, int endIndex) {
  __shared__ int sCount;
  Scalars& s = *d.sc;
  destroyContactsOf(d, bodyId, -1, endIndex, &sCount);
  if (threadIdx.x != 0) return;
  for (int sid = d.bodyHeadShape[bodyId]; sid != -1; sid = d.shapes[sid].nextShapeId) destroyProxySerial(d, sid);
  d.bodyHeadShape[bodyId] = -1;
  int sim = d.bodySim[bodyId];
  if (sim >= 0) {                                 // bodySims / bodyStates RemoveSwap; the host lowers awakeCount (a kernel parameter) after this launch
    int last = d.awakeCount - 1;
    if (sim != last) {
      d.st0[sim] = d.st0[last]; d.st1[sim] = d.st1[last]; d.simMass[sim] = d.simMass[last]; d.simForce[sim] = d.simForce[last];
      d.simPose[sim] = d.simPose[last]; d.simLocal[sim] = d.simLocal[last]; d.simPose0[sim] = d.simPose0[last]; d.simFlags[sim] = d.simFlags[last];
      int movedBody = d.simBodyId[last]; d.simBodyId[sim] = movedBody; d.bodySim[movedBody] = sim;
    }
    d.bodySim[bodyId] = -1;
  }
  d.bodyType[bodyId] = kStaticBody; d.bodyColorMask[bodyId] = 0; d.bodyHeadJoint[bodyId] = -1; d.bodyJointCount[bodyId] = 0;
  d.bodyXf[bodyId] = make_float4(0, 0, 0, 0); d.bodyCenter[bodyId] = make_float2(0, 0); d.bodyFlags[bodyId] = 0;
  s.graphDirty = 1;
}

int World::destroyJoint(int jointId) {           // DestroyJointInternal src/Joint.cs:297-375 (wakeBodies: no sleeping sets in this build)
  if (jointId < 0 || jointId >= (int)joints_.size() || joints_[jointId].id != jointId) return fail(B2C_E_INVALID, "bad joint id");
  HJoint& j = joints_[jointId];
  HBody& bodyA = bodies_[j.bodyIdA]; HBody& bodyB = bodies_[j.bodyIdB];
  auto setNext = [&](int key, int v) { HJoint& o = joints_[key >> 1]; if (key & 1) o.nextKeyB = v; else o.nextKeyA = v; };
  auto setPrev = [&](int key, int v) { HJoint& o = joints_[key >> 1]; if (key & 1) o.prevKeyB = v; else o.prevKeyA = v; };
  if (j.prevKeyA != -1) setNext(j.prevKeyA, j.nextKeyA);
  if (j.nextKeyA != -1) setPrev(j.nextKeyA, j.prevKeyA);
  if (bodyA.headJointKey == (jointId << 1)) bodyA.headJointKey = j.nextKeyA;
  bodyA.jointCount--;
  if (j.prevKeyB != -1) setNext(j.prevKeyB, j.nextKeyB);
  if (j.nextKeyB != -1) setPrev(j.nextKeyB, j.prevKeyB);
  if (bodyB.headJointKey == ((jointId << 1) | 1)) bodyB.headJointKey = j.nextKeyB;
  bodyB.jointCount--;
  const int color = j.sim.colorIndex;
  const bool inGraph = j.inGraph;
  if (inGraph) {
    std::vector<int>& arr = colorJoints_[color];      // RemoveJointFromGraph: RemoveSwap in the colour's array
    for (size_t k = 0; k < arr.size(); k++) if (arr[k] == jointId) { arr[k] = arr.back(); arr.pop_back(); break; }
    if (!onDevice_ && color != kOverflowIndex) { colorMask_[j.bodyIdA] &= ~(1u << color); colorMask_[j.bodyIdB] &= ~(1u << color); }
  }
  const int idA = j.bodyIdA, idB = j.bodyIdB;
  j.id = -1;
  jointIds_.release(jointId);
  if (!onDevice_) return B2C_OK;
  CK(cudaSetDevice(device_));
  if (inGraph) { k_joint_removed<<<1, 1, 0, stream_>>>(d_, idA, idB, color); CK(cudaGetLastError()); }
  return uploadJointLists();
}
int World::uploadJointLists() {                  // joint edges, per-body list heads and the colour-grouped joint order after a destroy
  const int nJoint = (int)joints_.size(), nBody = (int)bodies_.size();
  std::vector<int4> je(nJoint); std::vector<int> jsorted, head(nBody), cnt(nBody);
  for (int i = 0; i < nJoint; i++) { const HJoint& j = joints_[i]; je[i] = make_int4(j.bodyIdA, j.nextKeyA, j.bodyIdB, j.nextKeyB); }
  for (int i = 0; i < nBody; i++) { head[i] = bodies_[i].id == i ? bodies_[i].headJointKey : -1; cnt[i] = bodies_[i].id == i ? bodies_[i].jointCount : 0; }
  d_.jointColorStart[0] = 0;
  for (int c = 0; c < kGraphColorCount; c++) { for (int id : colorJoints_[c]) jsorted.push_back(id); d_.jointColorStart[c + 1] = (int)jsorted.size(); }
  CK(cudaStreamSynchronize(stream_));            // pageable sources: the copies below must not outlive these vectors
  if (nJoint) CK(cudaMemcpy(d_.jointEdges, je.data(), sizeof(int4) * nJoint, cudaMemcpyHostToDevice));
  if (!jsorted.empty()) CK(cudaMemcpy(d_.jointSorted, jsorted.data(), sizeof(int) * jsorted.size(), cudaMemcpyHostToDevice));
  CK(cudaMemcpy(d_.bodyHeadJoint, head.data(), sizeof(int) * nBody, cudaMemcpyHostToDevice));
  CK(cudaMemcpy(d_.bodyJointCount, cnt.data(), sizeof(int) * nBody, cudaMemcpyHostToDevice));
  return B2C_OK;
}
// World.DestroySensor src/Sensor.cs:219-235 (the same lines sit inside DestroyShapeInternal, src/Shape.cs:245-268): an end event for every shape
// in the sensor's current overlap list, appended to the end-event buffer of the running index, then RemoveSwap in World.sensors.
__global__ void k_destroy_sensor(WorldDev d, int sensorIndex, int shapeId, int lastIndex, int endIndex) {
  Scalars& s = *d.sc;
  const int parity = s.sensorParity;
  const int n = d.sensorOverlapCount[2 * sensorIndex + parity];
  const int2* o2 = d.sensorOverlaps + ((size_t)sensorIndex * 2 + parity) * kSensorMaxOverlaps;
  B2CShapeRef ref; ref.shapeId = shapeId; ref.generation = d.shapes[shapeId].generation;
  int base = s.sensorEndCount[endIndex];
  for (int i = 0; i < n; i++) {
    if (base >= d.sensorEventCapacity) { s.errorFlags |= (int)ERR_SENSOR_CAPACITY; break; }
    B2CSensorEvent e; e.sensor = ref; e.visitor.shapeId = o2[i].x; e.visitor.generation = (uint32_t)o2[i].y;
    d.sensorEndEvents[endIndex][base++] = e;
  }
  s.sensorEndCount[endIndex] = base;
  if (sensorIndex != lastIndex) {
    for (int p = 0; p < 2; p++) {
      const int m = d.sensorOverlapCount[2 * lastIndex + p];
      d.sensorOverlapCount[2 * sensorIndex + p] = m;
      const int2* src = d.sensorOverlaps + ((size_t)lastIndex * 2 + p) * kSensorMaxOverlaps; int2* dst = d.sensorOverlaps + ((size_t)sensorIndex * 2 + p) * kSensorMaxOverlaps;
      for (int i = 0; i < m; i++) dst[i] = src[i];
    }
    const int hits = d.sensorHitCount[lastIndex];
    d.sensorHitCount[sensorIndex] = hits;
    for (int i = 0; i < hits; i++) d.sensorHits[(size_t)sensorIndex * kSensorMaxHits + i] = d.sensorHits[(size_t)lastIndex * kSensorMaxHits + i];
    const int moved = d.sensorShape[lastIndex];
    d.sensorShape[sensorIndex] = moved; d.shapes[moved].sensorIndex = sensorIndex;
  }
  d.shapes[shapeId].sensorIndex = -1;
}
int World::destroySensor(int shapeId) {
  HShape& sh = shapes_[shapeId];
  const int index = sh.sensorIndex, last = (int)sensorShapes_.size() - 1;
  k_destroy_sensor<<<1, 1, 0, stream_>>>(d_, index, shapeId, last, endIndex_);
  CK(cudaGetLastError());
  if (index != last) { int moved = sensorShapes_[last]; sensorShapes_[index] = moved; shapes_[moved].sensorIndex = index; }
  sensorShapes_.pop_back(); sh.sensorIndex = -1;
  d_.sensorCount = (int)sensorShapes_.size();
  eventsFetched_ = false;
  return B2C_OK;
}
int World::destroyBody(int bodyId) {             // BodyAPI.DestroyBody api/Body.cs:106-160
  if (bodyId < 0 || bodyId >= (int)bodies_.size() || bodies_[bodyId].id != bodyId) return fail(B2C_E_INVALID, "bad body id");
  if (!onDevice_) { int rc = upload(); if (rc) return rc; }   // the first call that needs the device state uploads the world (as the first step would)
  HBody& b = bodies_[bodyId];
  int edgeKey = b.headJointKey;
  while (edgeKey != -1) {
    int jointId = edgeKey >> 1, edge = edgeKey & 1;
    edgeKey = edge == 1 ? joints_[jointId].nextKeyB : joints_[jointId].nextKeyA;
    int rc = destroyJoint(jointId); if (rc) return rc;
  }
  CK(cudaSetDevice(device_));
  for (int sid = b.headShapeId; sid != -1; sid = shapes_[sid].nextShapeId) if (shapes_[sid].sensorIndex != -1) { int rc = destroySensor(sid); if (rc) return rc; }   // api/Body.cs:127, shape-list order
  k_destroy_body<<<1, 256, 0, stream_>>>(d_, bodyId, endIndex_);      // contacts (list order), proxies, awake-set swap-remove
  CK(cudaGetLastError());
  for (int sid = b.headShapeId; sid != -1;) { HShape& sh = shapes_[sid]; int next = sh.nextShapeId; if (sh.proxyKey != -1) treeProxyRemoved(b.type); sh.proxyKey = -1; sh.id = -1; shapeIds_.release(sid); sid = next; }
  if (b.simIndex >= 0) {
    int last = (int)awakeOrder_.size() - 1;
    if (b.simIndex != last) { int moved = awakeOrder_[last]; awakeOrder_[b.simIndex] = moved; bodies_[moved].simIndex = b.simIndex; }
    awakeOrder_.pop_back();
    d_.awakeCount = (int)awakeOrder_.size();
  }
  bodyIds_.release(bodyId);
  b.id = -1; b.simIndex = -1; b.headShapeId = -1; b.headJointKey = -1; b.jointCount = 0;
  eventsFetched_ = false; bodiesFetched_ = false; movesFetched_ = false;
  return B2C_OK;
}

__global__ void k_body_mass_update(WorldDev d, int bodyId, int type, float invMass, float invInertia, V2 localCenter, float minExtent, float maxExtent);
int World::applyMassFromShapes(int bodyId) {     // BodyAPI.ApplyMassFromShapes api/Body.cs:617-621 -> UpdateBodyMassData src/Body.cs:255-353
  if (bodyId < 0 || bodyId >= (int)bodies_.size() || bodies_[bodyId].id != bodyId) return fail(B2C_E_INVALID, "bad body id");
  HBody& b = bodies_[bodyId];
  updateBodyMassData(b);
  if (!onDevice_) return B2C_OK;
  CK(cudaSetDevice(device_));
  k_body_mass_update<<<1, 1, 0, stream_>>>(d_, bodyId, b.type, b.invMass, b.invInertia, b.localCenter, b.minExtent, b.maxExtent);
  CK(cudaGetLastError());
  bodiesFetched_ = false; movesFetched_ = false;
  return B2C_OK;
}
int World::setShapeDensity(int shapeId, float density, bool updateBodyMass) {   // ShapeAPI.SetDensity api/Shape.cs:95-103
  if (shapeId < 0 || shapeId >= (int)shapes_.size() || shapes_[shapeId].id != shapeId) return fail(B2C_E_INVALID, "bad shape id");
  HShape& s = shapes_[shapeId];
  if (density == s.density) return B2C_OK;
  s.density = density;
  return updateBodyMass ? applyMassFromShapes(s.bodyId) : B2C_OK;
}
int World::destroyShape(int shapeId, bool updateBodyMass) {      // ShapeAPI.DestroyShape api/Shape.cs:48-56, DestroyShapeInternal src/Shape.cs:219-272
  if (shapeId < 0 || shapeId >= (int)shapes_.size() || shapes_[shapeId].id != shapeId) return fail(B2C_E_INVALID, "bad shape id");
  if (!onDevice_) { int rc = upload(); if (rc) return rc; }   // the first call that needs the device state uploads the world (as the first step would)
  HShape& s = shapes_[shapeId];
  HBody& body = bodies_[s.bodyId];
  const int prev = s.prevShapeId, next = s.nextShapeId, bodyId = s.bodyId;
  if (prev != -1) shapes_[prev].nextShapeId = next;
  if (next != -1) shapes_[next].prevShapeId = prev;
  if (shapeId == body.headShapeId) body.headShapeId = next;
  body.shapeCount--;
  CK(cudaSetDevice(device_));
  k_destroy_shape<<<1, 256, 0, stream_>>>(d_, shapeId, bodyId, prev, next, endIndex_);
  CK(cudaGetLastError());
  if (s.sensorIndex != -1) { int rc = destroySensor(shapeId); if (rc) return rc; }      // src/Shape.cs:245-268
  if (s.proxyKey != -1) treeProxyRemoved(body.type);
  s.proxyKey = -1; s.id = -1;
  shapeIds_.release(shapeId);
  eventsFetched_ = false;
  return updateBodyMass ? applyMassFromShapes(bodyId) : B2C_OK;
}

// Joint setters (api/Joint.cs): the host mirror is patched before the upload, the device record after it (stream-ordered, one thread).
__global__ void k_joint_patch(WorldDev d, int jointId, int patch, int which, float a, float b) { jointPatch(d.joints[jointId], patch, which, a, b); }
int World::jointOp(int jointId, int patch, int which, float a, float b) {
  if (jointId < 0 || jointId >= (int)joints_.size() || joints_[jointId].id != jointId) return fail(B2C_E_INVALID, "bad joint id");
  JointDev& s = joints_[jointId].sim;
  if (patch == JPATCH_PARAM && !jointParamValid(s.type, which)) return fail(B2C_E_INVALID, "joint parameter does not exist for this joint type");
  if (patch == JPATCH_ENABLE && !jointHasSpringLimitMotor(s.type, which)) return fail(B2C_E_INVALID, "joint type has no spring/limit/motor switch");
  if (patch == JPATCH_LIMITS && !jointHasSpringLimitMotor(s.type, B2C_JOINT_LIMIT)) return fail(B2C_E_INVALID, "joint type has no limits");
  if (!onDevice_) { jointPatch(s, patch, which, a, b); return B2C_OK; }
  CK(cudaSetDevice(device_));
  k_joint_patch<<<1, 1, 0, stream_>>>(d_, jointId, patch, which, a, b);
  CK(cudaGetLastError());
  return B2C_OK;
}

// JointAPI.SetLocalFrameA/B api/Joint.cs:44,50: a plain store into the joint sim (no wake, no impulse reset)
__global__ void k_joint_frame(WorldDev d, int jointId, int which, Xf f) { if (which) d.joints[jointId].localFrameB = f; else d.joints[jointId].localFrameA = f; }
int World::jointSetLocalFrame(int jointId, int which, Xf frame) {
  if (jointId < 0 || jointId >= (int)joints_.size() || joints_[jointId].id != jointId) return fail(B2C_E_INVALID, "bad joint id");
  if (which != 0 && which != 1) return fail(B2C_E_INVALID, "frame selector must be 0 (A) or 1 (B)");
  JointDev& s = joints_[jointId].sim;
  if (which) s.localFrameB = frame; else s.localFrameA = frame;      // the host record is the upload source and the read-back of parked joints
  if (!onDevice_) return B2C_OK;
  CK(cudaSetDevice(device_));
  k_joint_frame<<<1, 1, 0, stream_>>>(d_, jointId, which, frame);
  CK(cudaGetLastError());
  return B2C_OK;
}
// JointAPI.SetCollideConnected api/Joint.cs:56-75: on -> BufferMove for the shapes of the body with fewer shapes (pairs are found by the
// next UpdateBroadPhasePairs); off -> DestroyContactsBetweenBodies, in contact-list order as at joint creation.
__global__ void k_joint_collide(WorldDev d, int jointId, int flag, int bodyIdA, int bodyIdB, int moveBody, int endIndex) {
  __shared__ int sCount;
  Scalars& s = *d.sc;
  if (threadIdx.x == 0) { sCount = 0; d.jointCollide[jointId] = flag; }
  __syncthreads();
  if (flag) {
    if (threadIdx.x != 0) return;
    for (int sid = d.bodyHeadShape[moveBody]; sid != -1; sid = d.shapes[sid].nextShapeId) {
      int key = d.shapes[sid].proxyKey;
      if (key == -1) continue;
      const TreeDev& t = d.trees[key & 3];
      if (t.moved[key >> 2]) continue;
      if (s.moveCount < d.moveCapacity) { t.moved[key >> 2] = 1; d.moveArray[s.moveCount++] = key; } else s.errorFlags |= (int)ERR_PAIR_CAPACITY;
    }
    return;
  }
  const int n = s.contactNext;
  for (int i = threadIdx.x; i < n; i += blockDim.x) {
    if (!(d.cFlags[i] & CF_ALIVE)) continue;
    int2 cb = d.cBody[i];
    if ((cb.x == bodyIdA && cb.y == bodyIdB) || (cb.x == bodyIdB && cb.y == bodyIdA)) { int k = atomicAdd(&sCount, 1); d.opList[k] = i; }
  }
  __syncthreads();
  if (threadIdx.x != 0) return;
  const int count = sCount;
  for (int i = 1; i < count; i++) {              // newest first = the order of either body's contact list
    int id = d.opList[i]; uint32_t st = d.cStamp[id]; int j = i - 1;
    while (j >= 0 && d.cStamp[d.opList[j]] < st) { d.opList[j + 1] = d.opList[j]; j--; }
    d.opList[j + 1] = id;
  }
  for (int i = 0; i < count; i++) destroyContactSerial(d, d.opList[i], endIndex);
}
int World::jointSetCollideConnected(int jointId, bool shouldCollide) {
  if (jointId < 0 || jointId >= (int)joints_.size() || joints_[jointId].id != jointId) return fail(B2C_E_INVALID, "bad joint id");
  HJoint& j = joints_[jointId];
  if (j.collideConnected == shouldCollide) return B2C_OK;
  j.collideConnected = shouldCollide;
  const HBody& a = bodies_[j.bodyIdA]; const HBody& b = bodies_[j.bodyIdB];
  const int moveBody = a.shapeCount < b.shapeCount ? j.bodyIdA : j.bodyIdB;
  if (!onDevice_) {                               // no contacts exist before the first step
    if (shouldCollide) for (int sid = bodies_[moveBody].headShapeId; sid != -1; sid = shapes_[sid].nextShapeId) if (shapes_[sid].proxyKey != -1) bufferMove(shapes_[sid].proxyKey);
    return B2C_OK;
  }
  CK(cudaSetDevice(device_));
  k_joint_collide<<<1, 256, 0, stream_>>>(d_, jointId, shouldCollide ? 1 : 0, j.bodyIdA, j.bodyIdB, moveBody, endIndex_);
  CK(cudaGetLastError());
  eventsFetched_ = false;
  return B2C_OK;
}

// BodyAPI sleep surface (api/Body.cs:664-716) while islands / sleep are not built: every body outside the static set is awake and stays
// awake, so the calls that only wake (SetAwake(true), WakeTouching, EnableSleep(false)) have nothing to do, the stores are kept for
// read-back, and the one call that would put an island to sleep is refused.
__global__ void k_body_sleep_threshold(WorldDev d, int bodyId, float v) { d.bodySleepThreshold[bodyId] = v; }
int World::bodySetSleepThreshold(int bodyId, float v) {               // api/Body.cs:712
  if (bodyId < 0 || bodyId >= (int)bodies_.size() || bodies_[bodyId].id != bodyId) return fail(B2C_E_INVALID, "bad body id");
  bodies_[bodyId].sleepThreshold = v;
  if (!onDevice_) return B2C_OK;
  CK(cudaSetDevice(device_));
  k_body_sleep_threshold<<<1, 1, 0, stream_>>>(d_, bodyId, v);
  CK(cudaGetLastError());
  return B2C_OK;
}
int World::bodyEnableSleep(int bodyId, bool flag) {                   // api/Body.cs:700-706 (WakeBody finds the body awake)
  if (bodyId < 0 || bodyId >= (int)bodies_.size() || bodies_[bodyId].id != bodyId) return fail(B2C_E_INVALID, "bad body id");
  bodies_[bodyId].enableSleep = flag;
  return B2C_OK;
}
int World::bodySetAwake(int bodyId, bool awake) {                     // api/Body.cs:668-679
  if (bodyId < 0 || bodyId >= (int)bodies_.size() || bodies_[bodyId].id != bodyId) return fail(B2C_E_INVALID, "bad body id");
  if (awake || bodies_[bodyId].simIndex < 0) return B2C_OK;           // already awake / a static-set body: the reference does nothing
  return fail(B2C_E_UNSUPPORTED, "SetAwake(false) puts the body's island to sleep (TrySleepIsland, src/SolverSet.cs:136): islands / sleep are not built");
}
int World::bodyWakeTouching(int bodyId) {                             // api/Body.cs:682-697: WakeBody on bodies that are all awake
  if (bodyId < 0 || bodyId >= (int)bodies_.size() || bodies_[bodyId].id != bodyId) return fail(B2C_E_INVALID, "bad body id");
  return B2C_OK;
}
int World::bodySetEnabled(int bodyId, bool enabled) {                 // api/Body.cs:721-790
  if (bodyId < 0 || bodyId >= (int)bodies_.size() || bodies_[bodyId].id != bodyId) return fail(B2C_E_INVALID, "bad body id");
  if (enabled) return B2C_OK;                                         // Enable on a body that is not in the disabled set returns at once (:757)
  return fail(B2C_E_UNSUPPORTED, "Body.Disable moves the body, its contacts and joints to the disabled solver set (api/Body.cs:721-752): the disabled set is not built");
}
__global__ void k_body_set_bullet(WorldDev d, int bodyId, int flag) {
  const int sim = d.bodySim[bodyId];
  uint32_t* f = sim < 0 ? &d.bodySimFlags[bodyId] : &d.simFlags[sim];
  *f = flag ? (*f | kIsBullet) : (*f & ~(uint32_t)kIsBullet);
  if (d.bodyOut) d.bodyOut[bodyId].flags = *f;      // the streamed read-back row of a static body is not rewritten by the step
}
int World::bodySetBullet(int bodyId, bool flag) {                     // api/Body.cs:835-842
  if (bodyId < 0 || bodyId >= (int)bodies_.size() || bodies_[bodyId].id != bodyId) return fail(B2C_E_INVALID, "bad body id");
  HBody& b = bodies_[bodyId];
  b.flags = flag ? (b.flags | kIsBullet) : (b.flags & ~(uint32_t)kIsBullet);   // BodySim.flags only: Body.flags and BodyState.flags keep their bit
  if (!onDevice_) return B2C_OK;
  CK(cudaSetDevice(device_));
  k_body_set_bullet<<<1, 1, 0, stream_>>>(d_, bodyId, flag ? 1 : 0);
  CK(cudaGetLastError());
  bodiesFetched_ = false;
  return B2C_OK;
}

// BodyAPI.SetTransform after the upload: the same teleport as one device thread, ordered on the world's stream. The fat
// box test needs the device's current fatAABB (enlarged by earlier steps), the proxy moves through the shared tree ops,
// and BufferMove appends behind the moves the last step's FinalizeBodies buffered (src/BroadPhase.cs:56-59, :90-96).
__global__ void k_set_transform(WorldDev d, int bodyId, V2 p, Rot q) {
  Scalars& s = *d.sc;
  Xf xf = mkXf(p, q);
  float4 mass = d.bodyMass[bodyId];
  V2 center = xfPoint(xf, mk(mass.z, mass.w));
  float4 pose = make_float4(center.x, center.y, q.c, q.s);
  d.bodyXf[bodyId] = make_float4(p.x, p.y, q.c, q.s);
  d.bodyCenter[bodyId] = make_float2(center.x, center.y);
  int sim = d.bodySim[bodyId];
  if (sim >= 0) { d.simPose[sim] = pose; d.simPose0[sim] = pose; }
  for (int sid = d.bodyHeadShape[bodyId]; sid != -1; sid = d.shapes[sid].nextShapeId) {
    Box aabb = boxGrow(geomAABB(d.shapes[sid].geom, xf), kSpeculativeDistance);
    d.shapeAABB[sid] = make_float4(aabb.lo.x, aabb.lo.y, aabb.hi.x, aabb.hi.y);
    float4 f4 = d.shapeFat[sid]; Box cur; cur.lo = mk(f4.x, f4.y); cur.hi = mk(f4.z, f4.w);
    if (boxContains(cur, aabb)) continue;
    Box fat = boxGrow(aabb, kAABBMargin);
    d.shapeFat[sid] = make_float4(fat.lo.x, fat.lo.y, fat.hi.x, fat.hi.y);
    int key = d.shapes[sid].proxyKey;
    if (key == -1) continue;
    const TreeDev& t = d.trees[key & 3];
    TreeView v{t.aabb, t.child, t.parent, t.hf, t.cat, t.user, t.meta};
    if (!tvMoveProxy(v, key >> 2, fat)) s.errorFlags |= (int)ERR_TREE;
    if ((key & 3) == kStaticBody) s.staticRanksValid = 0;
    if (!t.moved[key >> 2]) {
      t.moved[key >> 2] = 1;
      if (s.moveCount < d.moveCapacity) d.moveArray[s.moveCount++] = key; else s.errorFlags |= (int)ERR_PAIR_CAPACITY;
    }
  }
}

int World::setTransform(int bodyId, V2 p, Rot q) {          // api/Body.cs:293-325
  if (bodyId < 0 || bodyId >= (int)bodies_.size() || bodies_[bodyId].id != bodyId) return fail(B2C_E_INVALID, "bad body id");
  if (onDevice_) {
    CK(cudaSetDevice(device_));
    k_set_transform<<<1, 1, 0, stream_>>>(d_, bodyId, p, q);
    CK(cudaGetLastError());
    return B2C_OK;
  }
  HBody& b = bodies_[bodyId];
  b.xf.p = p; b.xf.q = q;
  b.center = xfPoint(b.xf, b.localCenter);
  b.rot0 = q; b.center0 = b.center;
  for (int sid = b.headShapeId; sid != -1; sid = shapes_[sid].nextShapeId) {
    HShape& s = shapes_[sid];
    Box aabb = boxGrow(geomAABB(s.geom, b.xf), kSpeculativeDistance);
    s.aabb = aabb;
    if (!boxContains(s.fat, aabb)) {
      s.fat = boxGrow(aabb, kAABBMargin);
      if (s.proxyKey != -1) { trees_[s.proxyKey & 3].moveProxy(s.proxyKey >> 2, s.fat); bufferMove(s.proxyKey); }
    }
  }
  return B2C_OK;
}

template <class T> T* World::dalloc(size_t n) {
  void* p = nullptr;
  if (n == 0) n = 1;
  if (batch_) {                                    // bump allocation inside the batch's arena (freed with the batch)
    p = batchArenaAlloc(batch_, n * sizeof(T));    // arena chunks are zero-filled when they are created
    if (!p) return nullptr;
    arenaBlocks_.push_back({p, n * sizeof(T)});
    allocs_.push_back(p);
    return (T*)p;
  }
  if (cudaMalloc(&p, n * sizeof(T)) != cudaSuccess) return nullptr;
  cudaMemsetAsync(p, 0, n * sizeof(T), stream_);
  allocs_.push_back(p);
  return (T*)p;
}

#define UP(dst, vec) do { if (!(vec).empty()) CK(cudaMemcpyAsync((dst), (vec).data(), (vec).size() * sizeof((vec)[0]), cudaMemcpyHostToDevice, stream_)); } while (0)
#define DA(ptr, T, n) do { (ptr) = dalloc<T>(n); if (!(ptr)) return fail(B2C_E_CUDA, "cudaMalloc failed for " #ptr); allocRecs_.push_back({(void**)&(ptr), sizeof(T), std::max<size_t>((size_t)(n), 1)}); } while (0)

// ---- growth between steps ---------------------------------------------------------------------------------------------------------
// The reference's lists grow on demand (List<T>, the tree's x1.5 node pool). Here a creation call that finds its class of rows full
// re-allocates that class at x1.5 on the world's stream — new array, zero fill, device-to-device copy of the old rows, member pointer
// swapped — before it touches anything else. Kernels take WorldDev by value, so the next launch simply sees the new pointers (a cached
// CUDA graph is keyed by WorldDev and is re-captured). Contact-sized arrays do not grow: their capacity error stays loud.
int World::regrow(void** slot, size_t newCount, int fillByte) {
  for (AllocRec& r : allocRecs_) {
    if (r.slot != slot) continue;
    if (newCount <= r.count) return B2C_OK;
    void* p = nullptr;
    if (batch_) return fail(B2C_E_CAPACITY, "a world of a batch cannot grow its device arrays: reserve rows with B2CWorldDef.bodyCapacity / shapeCapacity / jointCapacity");
    if (cudaMalloc(&p, newCount * r.elem) != cudaSuccess) { cudaGetLastError(); return fail(B2C_E_CUDA, "cudaMalloc failed while growing a device array"); }
    CK(cudaMemsetAsync(p, fillByte, newCount * r.elem, stream_));
    void* old = *slot;
    CK(cudaMemcpyAsync(p, old, r.count * r.elem, cudaMemcpyDeviceToDevice, stream_));
    CK(cudaStreamSynchronize(stream_));
    cudaFree(old);
    for (void*& a : allocs_) if (a == old) a = p;
    *slot = p; r.count = newCount;
    for (auto& g : graphs_) if (g.exec) cudaGraphExecDestroy(g.exec);
    graphs_.clear();                               // captured launches hold the old pointers
    return B2C_OK;
  }
  return fail(B2C_E_INVALID, "regrow: unknown device array");
}
#define RG(ptr, n) do { int rc_ = regrow((void**)&(ptr), (size_t)(n)); if (rc_) return rc_; } while (0)
#define RGF(ptr, n, fill) do { int rc_ = regrow((void**)&(ptr), (size_t)(n), fill); if (rc_) return rc_; } while (0)
int World::growScanScratch() {
  int scanBlocks = std::max(std::max(std::max(d_.contactCapacity, d_.bigScratch), std::max(d_.moveCapacity, capAwake_)), capSensor_) / 64 + 2;   // chunk = block size of the scope, 64 at the narrowest (a batch world's finalize CTA)
  RG(d_.scanBlockSums, scanBlocks); RG(d_.scanBlockSums4, scanBlocks);
  return B2C_OK;
}
int World::growBodies(int minBodies, int minAwake) {
  WorldDev& d = d_;
  CK(cudaSetDevice(device_));
  if (minBodies > capBody_) {
    const int c = std::max(minBodies, capBody_ + capBody_ / 2);
    RG(d.bodyXf, c); RG(d.bodyMass, c); RG(d.bodyCenter, c); RG(d.bodySim, c); RG(d.bodyType, c); RG(d.bodyHeadShape, c); RG(d.bodyHeadJoint, c);
    RG(d.bodyJointCount, c); RG(d.bodyColorMask, c); RGF(d.bodyMinOp, c, 0x7f); RG(d.bodySleepThreshold, c); RG(d.bodyFlags, c); RG(d.bodySimFlags, c); RG(d.bodyForce, c); RG(dBodyRead_, c);
    if (hBodyRead_) cudaFreeHost(hBodyRead_);
    hBodyRead_ = nullptr;
    CK(cudaHostAlloc((void**)&hBodyRead_, sizeof(B2CBodyRead) * (size_t)c, cudaHostAllocDefault));
    if (d.bodyOut) d.bodyOut = dBodyRead_;
    bodiesFetched_ = false;
    capBody_ = c;
  }
  if (minAwake > capAwake_) {
    const int c = std::max(minAwake, capAwake_ + capAwake_ / 2);
    RG(d.st0, c); RG(d.st1, c); RG(d.simMass, c); RG(d.simForce, c); RG(d.simPose, c); RG(d.simLocal, c); RG(d.simPose0, c); RG(d.simFlags, c); RG(d.simBodyId, c);
    RG(d.ufParent, c); RG(d.rootBodyCount, c); RG(d.rootConCount, c); RG(d.rootColorCount, (size_t)c * kGraphColorCount);
    RG(d.rootInfo, c); RG(d.rootScan, c + 1); RG(d.islandBodyStart, c + 1); RG(d.islandConStart, c + 1);
    RG(d.islandColorOff, (size_t)(c + 1) * (kGraphColorCount + 1)); RG(d.islandBodies, c); RG(d.bodyLocal, c);
    RG(d.freeBodies, c); RG(d.bigBodies, c); RG(d.kinematicUsed, c); RG(d.bodyEnlargedCount, c + 1); RG(dMoveEvents_, c);
    if (hMoveEvents_) cudaFreeHost(hMoveEvents_);
    hMoveEvents_ = nullptr;
    CK(cudaHostAlloc((void**)&hMoveEvents_, sizeof(B2CBodyMoveEvent) * (size_t)c, cudaHostAllocMapped));
    if (hMoveEventsDev_ && cudaHostGetDevicePointer((void**)&hMoveEventsDev_, hMoveEvents_, 0) != cudaSuccess) { hMoveEventsDev_ = nullptr; cudaGetLastError(); }
    if (d.moveOut) d.moveOut = hMoveEventsDev_ ? hMoveEventsDev_ : dMoveEvents_;
    movesFetched_ = false;
    capAwake_ = c;
    int rc = growScanScratch(); if (rc) return rc;
  }
  return B2C_OK;
}
// Contact-sized rows (the reference's contact array, colour arrays, pair set and event lists grow without bound). Called between steps
// when the shapes outgrew the reserve of the first upload or when the last step that reached the host left the rows three quarters full.
// The colour arrays are rows of a [colour][capacity] table: they are laid out again at the new stride; the pair set is rebuilt by every
// step that has moves, so it only needs the room; pending end events (DestroyShape between steps appends to them) are carried over.
__global__ void k_relayout_col_list(const int* __restrict__ oldList, int* __restrict__ newList, const Scalars* sc, int oldCap, int newCap) {
  const int c = blockIdx.y, n = sc->colLen[c];
  for (int k = blockIdx.x * blockDim.x + threadIdx.x; k < n; k += gridDim.x * blockDim.x) newList[(size_t)c * newCap + k] = oldList[(size_t)c * oldCap + k];
}
int World::growContacts(int minContacts) {
  WorldDev& d = d_;
  if (minContacts <= d.contactCapacity) return B2C_OK;
  if (batch_) return fail(B2C_E_CAPACITY, "a world of a batch cannot grow its contact rows: reserve them with B2CWorldDef.contactCapacity");
  CK(cudaSetDevice(device_));
  const int oldCap = d.contactCapacity, c = std::max(minContacts, oldCap + oldCap / 2);
  RG(d.cShape, c); RG(d.cBody, c); RG(d.cFlags, c); RG(d.cColor, c); RG(d.cGen, c); RG(d.cStamp, c); RG(d.cMat, c); RG(d.cMass, c); RG(d.cSim, c);
  RG(d.mN, c); RG(d.mA0, c); RG(d.mA1, c); RG(d.mB0, c); RG(d.mB1, c); RG(d.mC0, c); RG(d.mC1, c); RG(d.contactFree, c);
  RG(d.flagScan, c + 1); RG(d.opList, c); RG(d.opDone, c); RG(d.opCol, c); RGF(d.cLocal, c, 0xff); RG(d.cRestGroup, c); RG(d.cConstraint, c);
  RG(d.blockHist, (size_t)(c / 256 + 2) * kGraphColorCount);
  { int* oldList = d.colList; int* newList = nullptr;          // [colour][capacity] rows at the new stride
    if (cudaMalloc((void**)&newList, sizeof(int) * (size_t)kGraphColorCount * c) != cudaSuccess) { cudaGetLastError(); return fail(B2C_E_CUDA, "cudaMalloc failed while growing the colour arrays"); }
    k_relayout_col_list<<<dim3(64, kGraphColorCount), 256, 0, stream_>>>(oldList, newList, d.sc, oldCap, c);
    CK(cudaGetLastError());
    CK(cudaStreamSynchronize(stream_));
    cudaFree(oldList);
    for (void*& a : allocs_) if (a == oldList) a = newList;
    for (AllocRec& r : allocRecs_) if (r.slot == (void**)&d.colList) r.count = (size_t)kGraphColorCount * c;
    d.colList = newList; }
  const int cc = c + 32 * kGraphColorCount;
  RG(d.sA, cc); RG(d.sB, cc); RG(d.sC, cc); RG(d.sD, cc); RG(d.sE, cc); RG(d.sF, cc); RG(d.sG, cc); RG(d.sH, cc); RG(d.sI, cc); RG(d.sJ, cc);
  RG(d.sContact, cc); RG(d.sLocal, cc);
  const int ov = c / 4 + 1024;
  RG(d.ovCons, ov);
  RG(d.movePairs, c); RG(d.bigHits, c);
  int tableSize = d.pairTableSize; while (tableSize < 4 * c) tableSize <<= 1;
  RG(d.pairTable, tableSize);
  // event lists: mapped pinned host memory the kernels write directly
  CK(cudaStreamSynchronize(stream_));
  auto regrowPinned = [&](void** host, void** dev, size_t elem) -> bool {
    void* p = nullptr;
    if (cudaHostAlloc(&p, elem * (size_t)c, cudaHostAllocMapped) != cudaSuccess) { cudaGetLastError(); return false; }
    memcpy(p, *host, elem * (size_t)oldCap);
    cudaFreeHost(*host);
    *host = p;
    if (cudaHostGetDevicePointer(dev, p, 0) != cudaSuccess) { cudaGetLastError(); return false; }
    return true;
  };
  if (!regrowPinned((void**)&hBegin_, (void**)&d.beginEvents, sizeof(B2CContactTouchEvent)) || !regrowPinned((void**)&hEnd_[0], (void**)&d.endEvents[0], sizeof(B2CContactTouchEvent)) ||
      !regrowPinned((void**)&hEnd_[1], (void**)&d.endEvents[1], sizeof(B2CContactTouchEvent)) || !regrowPinned((void**)&hHit_, (void**)&d.hitEvents, sizeof(B2CContactHitEvent)))
    return fail(B2C_E_CUDA, "cudaHostAlloc failed while growing the event lists");
  d.contactCapacity = c; d.pairCapacity = c; d.eventCapacity = c; d.constraintCapacity = cc; d.ovCapacity = ov; d.pairTableSize = tableSize;
  return growScanScratch();
}
int World::growShapes(int minShapes) {
  if (minShapes <= capShape_) return B2C_OK;
  WorldDev& d = d_;
  CK(cudaSetDevice(device_));
  const int c = std::max(minShapes, capShape_ + capShape_ / 2), extra = c - capShape_;
  RG(d.shapes, c); RG(d.shapeAABB, c); RG(d.shapeFat, c);
  const int m = d.moveCapacity + extra;              // every new shape may sit in the move buffer
  RG(d.moveArray, m); RG(d.moveHitCount, m + 1); RG(d.moveBig, m); RG(d.moveHitBuf, (size_t)m * kHitBuf);
  d.moveCapacity = m;
  capShape_ = c;
  return growScanScratch();
}
int World::growSensors(int minSensors) {
  if (minSensors <= capSensor_) return B2C_OK;
  if (batch_) return fail(B2C_E_CAPACITY, "a world of a batch cannot grow its sensor arrays: create its sensor shapes before the first step");
  WorldDev& d = d_;
  CK(cudaSetDevice(device_));
  const int c = std::max(minSensors, capSensor_ + capSensor_ / 2 + 16);
  RG(d.sensorShape, c); RG(d.sensorOverlaps, (size_t)c * 2 * kSensorMaxOverlaps); RG(d.sensorOverlapCount, (size_t)c * 2); RG(d.sensorScan, c + 1); RG(d.sensorHits, (size_t)c * kSensorMaxHits); RG(d.sensorHitCount, c);
  // event buffers are mapped host memory: a larger copy replaces them once the stream has drained (the pending end events move over)
  CK(cudaStreamSynchronize(stream_));
  const int evCap = c * kSensorMaxOverlaps;
  B2CSensorEvent* nb = nullptr; B2CSensorEvent* ne[2] = {nullptr, nullptr};
  CK(cudaHostAlloc((void**)&nb, sizeof(B2CSensorEvent) * (size_t)evCap, cudaHostAllocMapped));
  for (auto& e : ne) CK(cudaHostAlloc((void**)&e, sizeof(B2CSensorEvent) * (size_t)evCap, cudaHostAllocMapped));
  if (hSensorBegin_) {
    memcpy(nb, hSensorBegin_, sizeof(B2CSensorEvent) * (size_t)d.sensorEventCapacity); cudaFreeHost(hSensorBegin_);
    for (int k = 0; k < 2; k++) { memcpy(ne[k], hSensorEnd_[k], sizeof(B2CSensorEvent) * (size_t)d.sensorEventCapacity); cudaFreeHost(hSensorEnd_[k]); }
  }
  hSensorBegin_ = nb; hSensorEnd_[0] = ne[0]; hSensorEnd_[1] = ne[1];
  CK(cudaHostGetDevicePointer((void**)&d.sensorBeginEvents, hSensorBegin_, 0));
  for (int k = 0; k < 2; k++) CK(cudaHostGetDevicePointer((void**)&d.sensorEndEvents[k], hSensorEnd_[k], 0));
  d.sensorEventCapacity = evCap; d.sensorCapacity = c; capSensor_ = c;
  return growScanScratch();
}
int World::growJoints(int minJoints) {
  if (minJoints <= capJoint_) return B2C_OK;
  WorldDev& d = d_;
  CK(cudaSetDevice(device_));
  const int c = std::max(minJoints, capJoint_ + capJoint_ / 2);
  RG(d.joints, c); RG(d.jointEdges, c); RG(d.jointCollide, c); RG(d.jointSorted, c); RG(d.jointEventFlag, c);
  if (hJointEvents_) cudaFreeHost(hJointEvents_);
  hJointEvents_ = nullptr;
  CK(cudaHostAlloc((void**)&hJointEvents_, sizeof(int) * (size_t)c, cudaHostAllocDefault));
  capJoint_ = c;
  return B2C_OK;
}
int World::growTree(int type, int minNodes) {      // the allocation behind the tree's logical capacity (which follows the reference's x1.5 steps)
  if (minNodes <= treeAllocCap_[type]) return B2C_OK;
  TreeDev& td = d_.trees[type];
  CK(cudaSetDevice(device_));
  int cap = treeAllocCap_[type];
  while (cap < minNodes) cap += cap >> 1;
  RG(td.aabb, cap); RG(td.child, cap); RG(td.parent, cap); RG(td.hf, cap); RG(td.cat, cap); RG(td.user, cap); RG(td.moved, cap); RG(td.leafIdx, cap); RG(td.leafCtr, cap);
  RG(td.segStart, cap); RG(td.recEnd, cap + 1); RG(td.recNode, cap + 1); RG(td.recSplit, cap + 1); RG(td.recBox, cap + 1);
  RG(td.flagScan, cap + 1); RG(td.flagArr, cap + 1); RG(td.pairR, cap + 1); RG(td.freeIds, cap + 1); RG(td.subCount, cap);
  RG(td.visit, cap); RG(td.serialList, cap + 1);
  RG(td.recEnd2, cap + 1); RG(td.recNode2, cap + 1); RG(td.recBox2, cap + 1); RG(td.recM, cap + 1); RG(td.serEnd, cap + 1); RG(td.serNode, cap + 1);
  treeAllocCap_[type] = cap;
  int sumCap = 0; for (int t = 0; t < 3; t++) sumCap += treeAllocCap_[t];
  if (sumCap > d_.bigScratch) { RG(d_.bigScan, sumCap); RG(d_.bigFlag, sumCap); d_.bigScratch = sumCap; }
  return growScanScratch();
}

int World::upload() {
  CK(cudaSetDevice(device_));
  const bool member = batch_ != nullptr;           // batch member: the batch's stream and arena, slim reserves, no per-world events or pinned buffers
  if (member) { stream_ = batchStream(batch_); if (!hostScalars_) hostScalars_ = (Scalars*)malloc(sizeof(Scalars)); }
  else {
    CK(cudaStreamCreateWithFlags(&stream_, cudaStreamNonBlocking));
    for (auto& e : timer_) CK(cudaEventCreate(&e));
    CK(cudaHostAlloc((void**)&hostScalars_, sizeof(Scalars), cudaHostAllocMapped));
  }
  memset(hostScalars_, 0, sizeof(Scalars));
  WorldDev& d = d_;
  const int nBody = (int)bodies_.size(), nAwake = (int)awakeOrder_.size(), nShape = (int)shapes_.size(), nJoint = (int)joints_.size();
  d.awakeCount = nAwake; d.bodyCapacity = nBody; d.shapeCapacity = nShape; d.jointCapacity = nJoint;
  // Headroom for bodies and shapes created after this upload (b2cuda_create_body / _create_shape between steps): a quarter more, at least
  // 256, or what the world definition asks for. Past it the create call fails with B2C_E_CAPACITY; nothing here grows behind a running step.
  const int reserve = member ? 8 : 256;            // a batch world is built before its first step; larger reserves come from the world definition
  capBody_ = std::max(nBody + std::max(reserve, nBody / 4), capHintBody); capShape_ = std::max(nShape + std::max(reserve, nShape / 4), capHintShape);
  capAwake_ = nAwake + (capBody_ - nBody);
  capJoint_ = std::max(nJoint + std::max(reserve, nJoint / 4), capHintJoint);
  const int capBody = capBody_, capAwake = capAwake_, capShape = capShape_, capJoint = capJoint_;
  // ---- bodies ----
  std::vector<float4> st0(nAwake), st1(nAwake), simMass(nAwake), simForce(nAwake), simPose(nAwake), simLocal(nAwake), simPose0(nAwake);
  std::vector<uint32_t> simFlags(nAwake); std::vector<int> simBodyId(nAwake);
  for (int i = 0; i < nAwake; i++) {
    const HBody& b = bodies_[awakeOrder_[i]];
    uint32_t fl = b.flags;
    st0[i] = make_float4(b.v.x, b.v.y, b.w, __builtin_bit_cast(float, fl));
    st1[i] = make_float4(0, 0, 1, 0);
    simMass[i] = make_float4(b.invMass, b.invInertia, b.linearDamping, b.angularDamping);
    simForce[i] = make_float4(b.force.x, b.force.y, b.torque, b.gravityScale);
    simPose[i] = make_float4(b.center.x, b.center.y, b.xf.q.c, b.xf.q.s);
    simLocal[i] = make_float4(b.localCenter.x, b.localCenter.y, b.minExtent, b.maxExtent);
    simPose0[i] = make_float4(b.center0.x, b.center0.y, b.rot0.c, b.rot0.s);
    simFlags[i] = fl; simBodyId[i] = b.id;
  }
  DA(d.st0, float4, capAwake); DA(d.st1, float4, capAwake); DA(d.simMass, float4, capAwake); DA(d.simForce, float4, capAwake);
  DA(d.simPose, float4, capAwake); DA(d.simLocal, float4, capAwake); DA(d.simPose0, float4, capAwake); DA(d.simFlags, uint32_t, capAwake); DA(d.simBodyId, int, capAwake);
  UP(d.st0, st0); UP(d.st1, st1); UP(d.simMass, simMass); UP(d.simForce, simForce); UP(d.simPose, simPose); UP(d.simLocal, simLocal);
  UP(d.simPose0, simPose0); UP(d.simFlags, simFlags); UP(d.simBodyId, simBodyId);
  std::vector<float4> bodyXf(nBody), bodyMass(nBody); std::vector<float2> bodyCenter(nBody);
  std::vector<int> bodySim(nBody), bodyType(nBody), bodyHeadShape(nBody), bodyHeadJoint(nBody), bodyJointCount(nBody);
  std::vector<float> bodySleep(nBody); std::vector<uint32_t> bodyFlags(nBody);
  for (int i = 0; i < nBody; i++) {
    const HBody& b = bodies_[i];
    bodyXf[i] = make_float4(b.xf.p.x, b.xf.p.y, b.xf.q.c, b.xf.q.s);
    bodyMass[i] = make_float4(b.invMass, b.invInertia, b.localCenter.x, b.localCenter.y);
    bodyCenter[i] = make_float2(b.center.x, b.center.y);
    bodySim[i] = b.simIndex; bodyType[i] = b.type; bodyHeadShape[i] = b.headShapeId; bodyHeadJoint[i] = b.headJointKey; bodyJointCount[i] = b.jointCount;
    bodySleep[i] = b.sleepThreshold; bodyFlags[i] = b.flags;
  }
  for (int id : lockQuirk_) bodyFlags[id] |= kLockLinearX | kLockLinearY | kLockAngularZ;
  colorMask_.resize(nBody, 0);
  DA(d.bodyXf, float4, capBody); DA(d.bodyMass, float4, capBody); DA(d.bodyCenter, float2, capBody); DA(d.bodySim, int, capBody); DA(d.bodyType, int, capBody);
  DA(d.bodyHeadShape, int, capBody); DA(d.bodyHeadJoint, int, capBody); DA(d.bodyJointCount, int, capBody); DA(d.bodyColorMask, uint32_t, capBody);
  DA(d.bodyMinOp, int, capBody); DA(d.bodySleepThreshold, float, capBody); DA(d.bodyFlags, uint32_t, capBody); DA(d.bodySimFlags, uint32_t, capBody); DA(d.bodyForce, float4, capBody);
  UP(d.bodyXf, bodyXf); UP(d.bodyMass, bodyMass); UP(d.bodyCenter, bodyCenter); UP(d.bodySim, bodySim); UP(d.bodyType, bodyType);
  UP(d.bodyHeadShape, bodyHeadShape); UP(d.bodyHeadJoint, bodyHeadJoint); UP(d.bodyJointCount, bodyJointCount); UP(d.bodyColorMask, colorMask_);
  { std::vector<uint32_t> sf(nBody); for (int i = 0; i < nBody; i++) sf[i] = bodies_[i].flags; UP(d.bodySimFlags, sf); CK(cudaStreamSynchronize(stream_)); }
  UP(d.bodySleepThreshold, bodySleep); UP(d.bodyFlags, bodyFlags);
  CK(cudaMemsetAsync(d.bodyMinOp, 0x7f, sizeof(int) * (size_t)std::max(capBody, 1), stream_));
  // ---- shapes ----
  std::vector<ShapeDev> sh(nShape); std::vector<float4> shAABB(nShape), shFat(nShape);
  for (int i = 0; i < nShape; i++) {
    const HShape& s = shapes_[i]; ShapeDev& o = sh[i];
    memset(&o, 0, sizeof(o));
    o.geom = s.geom; o.material = make_float4(s.friction, s.restitution, s.rollingResistance, s.tangentSpeed);
    o.categoryBits = s.categoryBits; o.maskBits = s.maskBits; o.groupIndex = s.groupIndex; o.bodyId = s.bodyId; o.nextShapeId = s.nextShapeId;
    o.proxyKey = s.proxyKey; o.flags = s.flags; o.sensorIndex = s.sensorIndex; o.generation = s.generation;
    shAABB[i] = make_float4(s.aabb.lo.x, s.aabb.lo.y, s.aabb.hi.x, s.aabb.hi.y);
    shFat[i] = make_float4(s.fat.lo.x, s.fat.lo.y, s.fat.hi.x, s.fat.hi.y);
  }
  DA(d.shapes, ShapeDev, capShape); DA(d.shapeAABB, float4, capShape); DA(d.shapeFat, float4, capShape);
  UP(d.shapes, sh); UP(d.shapeAABB, shAABB); UP(d.shapeFat, shFat);
  // ---- joints ----
  std::vector<JointDev> jd(nJoint); std::vector<int4> je(nJoint); std::vector<int> jc(nJoint), jsorted;
  for (int i = 0; i < nJoint; i++) {
    const HJoint& j = joints_[i];
    jd[i] = j.sim; je[i] = make_int4(j.bodyIdA, j.nextKeyA, j.bodyIdB, j.nextKeyB); jc[i] = j.collideConnected ? 1 : 0;
  }
  d.jointColorStart[0] = 0;
  for (int c = 0; c < kGraphColorCount; c++) {
    for (int id : colorJoints_[c]) jsorted.push_back(id);
    d.jointColorStart[c + 1] = (int)jsorted.size();
  }
  DA(d.joints, JointDev, capJoint); DA(d.jointEdges, int4, capJoint); DA(d.jointCollide, int, capJoint); DA(d.jointSorted, int, capJoint); DA(d.jointEventFlag, uint8_t, capJoint);
  UP(d.joints, jd); UP(d.jointEdges, je); UP(d.jointCollide, jc); UP(d.jointSorted, jsorted);
  // ---- trees ----
  int totalProxies = 0;
  for (int t = 0; t < 3; t++) {
    HostTree& h = trees_[t]; TreeDev& td = d.trees[t];
    const int logicalCap = h.capacity();
    int cap = logicalCap;                          // allocation: the reference's growth steps (x1.5) until the reserved shapes fit
    while (cap - h.nodeCount < 2 * (capShape - nShape) + 2) cap += cap >> 1;
    treeAllocCap_[t] = cap; treeNodeCount_[t] = h.nodeCount; treeProxyCount_[t] = h.proxyCount;
    td.capacity = logicalCap; td.stackCapacity = member ? 1024 : 16384;
    DA(td.aabb, float4, cap); DA(td.child, int2, cap); DA(td.parent, int, cap); DA(td.hf, uint32_t, cap); DA(td.cat, uint64_t, cap);
    DA(td.user, int, cap); DA(td.moved, uint8_t, cap); DA(td.meta, int, 4); DA(td.leafIdx, int, cap); DA(td.leafCtr, float2, cap); DA(td.stack, int, td.stackCapacity);
    DA(td.segStart, int, cap); DA(td.recEnd, int, cap + 1); DA(td.recNode, int, cap + 1); DA(td.recSplit, int, cap + 1); DA(td.recBox, float4, cap + 1);
    DA(td.flagScan, int, cap + 1); DA(td.flagArr, int, cap + 1); DA(td.pairR, int, cap + 1); DA(td.freeIds, int, cap + 1); DA(td.subCount, int, cap);
    DA(td.visit, int, cap); DA(td.serialList, int, cap + 1); DA(td.rb, int, 8);
    DA(td.recEnd2, int, cap + 1); DA(td.recNode2, int, cap + 1); DA(td.recBox2, float4, cap + 1); DA(td.recM, int, cap + 1); DA(td.serEnd, int, cap + 1); DA(td.serNode, int, cap + 1);
    UP(td.aabb, h.aabb); UP(td.child, h.child); UP(td.parent, h.parent); UP(td.hf, h.hf); UP(td.cat, h.cat); UP(td.user, h.user);
    std::vector<uint8_t> mv(cap, 0);
    for (size_t k = 0; k < moved_[t].size() && k < (size_t)cap; k++) mv[k] = moved_[t][k];
    UP(td.moved, mv);
    int meta[4] = {h.root, h.nodeCount, h.freeList, h.proxyCount};
    CK(cudaMemcpyAsync(td.meta, meta, sizeof(meta), cudaMemcpyHostToDevice, stream_));
    CK(cudaStreamSynchronize(stream_));
    totalProxies += h.proxyCount;
  }
  d.moveCapacity = std::max(totalProxies + (capShape - nShape), 16);
  int contactCap = capHintContact > 0 ? capHintContact : std::max(member ? 256 : 4096, 12 * nShape);
  d.contactCapacity = contactCap; d.pairCapacity = contactCap; d.eventCapacity = contactCap;
  d.constraintCapacity = contactCap + 32 * kGraphColorCount;
  int tableSize = 1; while (tableSize < (member ? 2 : 4) * contactCap) tableSize <<= 1;
  d.pairTableSize = tableSize;
  DA(d.moveArray, int, d.moveCapacity); DA(d.moveHitCount, int, d.moveCapacity + 1); DA(d.movePairs, int2, d.pairCapacity);
  DA(d.pairTable, unsigned long long, tableSize);
  { int sumCap = 0; for (int t = 0; t < 3; t++) sumCap += treeAllocCap_[t];
    d.bigScratch = member ? sumCap : std::max(1 << 21, sumCap);
    d.bigQueryVisits = kBigQueryVisits;
    d.islandMaxBodies = kIslandMaxBodies;
    d.debugChecksums = getenv("B2CUDA_DEBUG_CHECKSUMS") ? atoi(getenv("B2CUDA_DEBUG_CHECKSUMS")) : 0;
    if (const char* e = getenv("B2CUDA_ISLAND_MAX_BODIES")) { int v = atoi(e); if (v >= 0 && v <= kIslandMaxBodies) d.islandMaxBodies = v; }
    if (const char* e = getenv("B2CUDA_BIG_QUERY_VISITS")) { int v = atoi(e); if (v > 0) d.bigQueryVisits = v; }
    DA(d.moveBig, int, d.moveCapacity); DA(d.moveHitBuf, int2, (size_t)d.moveCapacity * kHitBuf); DA(d.bigList, int, kMaxBigQueries); DA(d.bigSegStart, int, kMaxBigQueries + 1); DA(d.bigOff, int, kMaxBigQueries + 1);
    DA(d.bigScan, int, d.bigScratch); DA(d.bigFlag, uint8_t, d.bigScratch); DA(d.bigHits, int2, d.pairCapacity); }
  UP(d.moveArray, moveArray_);
  DA(d.cShape, int2, contactCap); DA(d.cBody, int2, contactCap); DA(d.cFlags, uint32_t, contactCap); DA(d.cColor, int, contactCap); DA(d.cGen, uint32_t, contactCap); DA(d.cStamp, uint32_t, contactCap);
  DA(d.cMat, float4, contactCap); DA(d.cMass, float4, contactCap); DA(d.cSim, int2, contactCap);
  DA(d.mN, float4, contactCap); DA(d.mA0, float4, contactCap); DA(d.mA1, float4, contactCap); DA(d.mB0, float4, contactCap); DA(d.mB1, float4, contactCap);
  DA(d.mC0, float4, contactCap); DA(d.mC1, float4, contactCap); DA(d.contactFree, int, contactCap);
  DA(d.flagScan, int4, contactCap + 1); DA(d.opList, int, contactCap); DA(d.opDone, uint8_t, contactCap); DA(d.opCol, uint8_t, contactCap);
  d.ovCapacity = contactCap / 4 + (member ? 16 : 1024);
  DA(d.colList, int, (size_t)kGraphColorCount * contactCap); DA(d.cLocal, int, contactCap); DA(d.cRestGroup, uint8_t, contactCap); DA(d.ovCons, OvConstraint, d.ovCapacity);
  CK(cudaMemsetAsync(d.cLocal, 0xff, sizeof(int) * (size_t)contactCap, stream_));
  int cc = d.constraintCapacity;
  DA(d.sA, float4, cc); DA(d.sB, float4, cc); DA(d.sC, float4, cc); DA(d.sD, float4, cc); DA(d.sE, float4, cc); DA(d.sF, float4, cc);
  DA(d.sG, float4, cc); DA(d.sH, float4, cc); DA(d.sI, float4, cc); DA(d.sJ, float4, cc); DA(d.sContact, int, cc); DA(d.cConstraint, int, contactCap);
  DA(d.blockHist, int, (size_t)(contactCap / 256 + 2) * kGraphColorCount);
  DA(d.ufParent, int, capAwake); DA(d.rootBodyCount, int, capAwake); DA(d.rootConCount, int, capAwake); DA(d.rootColorCount, int, (size_t)capAwake * kGraphColorCount);
  DA(d.rootInfo, int, capAwake); DA(d.rootScan, int4, capAwake + 1); DA(d.islandBodyStart, int, capAwake + 1); DA(d.islandConStart, int, capAwake + 1);
  DA(d.islandColorOff, int, (size_t)(capAwake + 1) * (kGraphColorCount + 1)); DA(d.islandBodies, int, capAwake); DA(d.bodyLocal, int, capAwake);
  DA(d.freeBodies, int, capAwake); DA(d.bigBodies, int, capAwake); DA(d.kinematicUsed, uint8_t, capAwake); DA(d.sLocal, int2, cc);
  DA(d.bodyEnlargedCount, int, capAwake + 1);
  // ---- sensors (src/Sensor.cs): World.sensors order, both overlap lists per sensor (empty at the start), event buffers ----
  { const int nSensor = (int)sensorShapes_.size();
    capSensor_ = nSensor + (member ? 2 : std::max(16, nSensor / 4));
    d.sensorCount = nSensor; d.sensorCapacity = capSensor_; d.sensorEventCapacity = capSensor_ * kSensorMaxOverlaps;
    DA(d.sensorShape, int, capSensor_); DA(d.sensorOverlaps, int2, (size_t)capSensor_ * 2 * kSensorMaxOverlaps); DA(d.sensorOverlapCount, int, (size_t)capSensor_ * 2);
    DA(d.sensorScan, int4, capSensor_ + 1); DA(d.sensorHits, int2, (size_t)capSensor_ * kSensorMaxHits); DA(d.sensorHitCount, int, capSensor_);
    UP(d.sensorShape, sensorShapes_);
    if (member) {
      DA(d.sensorBeginEvents, B2CSensorEvent, d.sensorEventCapacity); DA(d.sensorEndEvents[0], B2CSensorEvent, d.sensorEventCapacity); DA(d.sensorEndEvents[1], B2CSensorEvent, d.sensorEventCapacity);
    } else {
      CK(cudaHostAlloc((void**)&hSensorBegin_, sizeof(B2CSensorEvent) * (size_t)d.sensorEventCapacity, cudaHostAllocMapped));
      for (auto& e : hSensorEnd_) CK(cudaHostAlloc((void**)&e, sizeof(B2CSensorEvent) * (size_t)d.sensorEventCapacity, cudaHostAllocMapped));
      CK(cudaHostGetDevicePointer((void**)&d.sensorBeginEvents, hSensorBegin_, 0));
      for (int k = 0; k < 2; k++) CK(cudaHostGetDevicePointer((void**)&d.sensorEndEvents[k], hSensorEnd_[k], 0));
    } }
  DA(d.sc, Scalars, 1);
  d.hostSc = nullptr;                              // the step's last warp keeps the pinned mirror current (a batch member is fetched on demand)
  if (!member && cudaHostGetDevicePointer((void**)&d.hostSc, hostScalars_, 0) != cudaSuccess) { d.hostSc = nullptr; cudaGetLastError(); }
  int scanBlocks = std::max(std::max(std::max(contactCap, d.bigScratch), std::max(d.moveCapacity, capAwake)), capSensor_) / 64 + 2;   // chunk = block size of the scope, 64 at the narrowest (a batch world's finalize CTA)
  DA(d.scanBlockSums, int, scanBlocks); DA(d.scanBlockSums4, int4, scanBlocks);
  Scalars init; memset(&init, 0, sizeof(init));
  init.moveCount = (int)moveArray_.size(); init.awakeCount = nAwake; init.graphDirty = 1;
  CK(cudaMemcpyAsync(d.sc, &init, sizeof(init), cudaMemcpyHostToDevice, stream_));
  d.bodyOut = nullptr; d.moveOut = nullptr;
  DA(dBodyRead_, B2CBodyRead, capBody);
  DA(dMoveEvents_, B2CBodyMoveEvent, capAwake);
  hMoveEventsDev_ = nullptr;
  if (member) {                                    // events stay in device memory and are copied on demand; pinned read-back buffers come on first use
    DA(d.beginEvents, B2CContactTouchEvent, d.eventCapacity); DA(d.endEvents[0], B2CContactTouchEvent, d.eventCapacity);
    DA(d.endEvents[1], B2CContactTouchEvent, d.eventCapacity); DA(d.hitEvents, B2CContactHitEvent, d.eventCapacity);
    hJointColor_ = batchMappedInt(batch_, &hJointColorDev_);
  } else {
    // contact events are written by the kernels straight into mapped pinned host memory (few per step; no copy, no extra sync)
    CK(cudaHostAlloc((void**)&hBegin_, sizeof(B2CContactTouchEvent) * d.eventCapacity, cudaHostAllocMapped));
    for (auto& e : hEnd_) CK(cudaHostAlloc((void**)&e, sizeof(B2CContactTouchEvent) * d.eventCapacity, cudaHostAllocMapped));
    CK(cudaHostAlloc((void**)&hHit_, sizeof(B2CContactHitEvent) * d.eventCapacity, cudaHostAllocMapped));
    CK(cudaHostGetDevicePointer((void**)&d.beginEvents, hBegin_, 0));
    for (int k = 0; k < 2; k++) CK(cudaHostGetDevicePointer((void**)&d.endEvents[k], hEnd_[k], 0));
    CK(cudaHostGetDevicePointer((void**)&d.hitEvents, hHit_, 0));
    { int rc = ensureHostBuffers(); if (rc) return rc; }
    // B2CUDA_ZEROCOPY_MOVES=1: FinalizeBodies writes the move events straight into this mapped buffer instead of a device array + copy behind
    // the step. Measured slower end to end (many_pyramids 560 vs 582 M body-substeps/s, batched 1663 vs 1779): off by default.
    { const char* e = getenv("B2CUDA_ZEROCOPY_MOVES");
      if (e && !strcmp(e, "1") && cudaHostGetDevicePointer((void**)&hMoveEventsDev_, hMoveEvents_, 0) != cudaSuccess) { hMoveEventsDev_ = nullptr; cudaGetLastError(); } }
    CK(cudaHostAlloc((void**)&hJointColor_, sizeof(int), cudaHostAllocMapped));
    CK(cudaHostGetDevicePointer((void**)&hJointColorDev_, hJointColor_, 0));
  }
  CK(cudaStreamSynchronize(stream_));
  onDevice_ = true;
  return B2C_OK;
}

void World::saveHostSnap(HostSnap& h) const {
  h.bodyIds = bodyIds_; h.shapeIds = shapeIds_; h.jointIds = jointIds_; h.bodies = bodies_; h.shapes = shapes_; h.joints = joints_;
  h.awakeOrder = awakeOrder_; h.sensorShapes = sensorShapes_; h.colorJoints = colorJoints_; h.colorMask = colorMask_; h.d = d_;
  for (int t = 0; t < 3; t++) { h.treeNodeCount[t] = treeNodeCount_[t]; h.treeProxyCount[t] = treeProxyCount_[t]; }
  h.valid = true;
}
void World::restoreHostSnap(const HostSnap& h) {
  bodyIds_ = h.bodyIds; shapeIds_ = h.shapeIds; jointIds_ = h.jointIds; bodies_ = h.bodies; shapes_ = h.shapes; joints_ = h.joints;
  awakeOrder_ = h.awakeOrder; sensorShapes_ = h.sensorShapes; colorJoints_ = h.colorJoints; colorMask_ = h.colorMask; d_ = h.d;
  for (int t = 0; t < 3; t++) { treeNodeCount_[t] = h.treeNodeCount[t]; treeProxyCount_[t] = h.treeProxyCount[t]; }
  eventsFetched_ = false; bodiesFetched_ = false; movesFetched_ = false; lastStepDtZero_ = false;
}

int World::ensureHostBuffers() {                   // pinned read-back buffers (bulk body records, move events, joint events)
  if (!hMoveEvents_) CK(cudaHostAlloc((void**)&hMoveEvents_, sizeof(B2CBodyMoveEvent) * (size_t)std::max(capAwake_, 1), cudaHostAllocMapped));
  if (!hBodyRead_) CK(cudaHostAlloc((void**)&hBodyRead_, sizeof(B2CBodyRead) * (size_t)std::max(capBody_, 1), cudaHostAllocDefault));
  if (!hJointEvents_) CK(cudaHostAlloc((void**)&hJointEvents_, sizeof(int) * (size_t)std::max(capJoint_, 1), cudaHostAllocDefault));
  return B2C_OK;
}

int World::synchronize() {
  if (!onDevice_) return B2C_OK;
  CK(cudaSetDevice(device_));
  CK(cudaStreamSynchronize(stream_));
  if (batch_ && scalarsStale_) { CK(cudaMemcpy(hostScalars_, d_.sc, sizeof(Scalars), cudaMemcpyDeviceToHost)); scalarsStale_ = false; }
  if (hostScalars_->errorFlags) {
    char buf[160]; snprintf(buf, sizeof(buf), "device error flags 0x%x (capacity/stack/overflow-color/TOI/sensor overlaps; see b2c_world.cuh ERR_*)", hostScalars_->errorFlags);
    uint32_t f = (uint32_t)hostScalars_->errorFlags;
    if (f & (ERR_CONTACT_CAPACITY | ERR_PAIR_CAPACITY | ERR_HASH_FULL | ERR_STACK_OVERFLOW)) return fail(B2C_E_CAPACITY, buf);
    if (f & ERR_OVERFLOW_COLOR) return fail(B2C_E_UNSUPPORTED, buf);
    if (f & ERR_REFERENCE_THROWS) return fail(B2C_E_UNSUPPORTED, std::string(buf) + ": the C# reference throws IndexOutOfRangeException in this step (Distance.MakeSimplexFromCache reads B's cached vertex from proxy A, src/Distance.cs:116-132, reached from TimeOfImpact)");
    return fail(B2C_E_CUDA, buf);
  }
  return B2C_OK;
}

// ---- mutators that patch device rows directly (api/Body.cs:359-545) ----
int World::setLinearVelocity(int bodyId, V2 v) {
  if (bodyId < 0 || bodyId >= (int)bodies_.size() || bodies_[bodyId].id != bodyId) return fail(B2C_E_INVALID, "bad body id");
  HBody& b = bodies_[bodyId];
  if (b.type == kStaticBody) return B2C_OK;
  if (!onDevice_) { b.v = v; return B2C_OK; }
  CK(cudaSetDevice(device_));
  float xy[2] = {v.x, v.y};
  CK(cudaMemcpyAsync(&d_.st0[b.simIndex], xy, sizeof(xy), cudaMemcpyHostToDevice, stream_));
  CK(cudaStreamSynchronize(stream_));
  return B2C_OK;
}
int World::setAngularVelocity(int bodyId, float w) {
  if (bodyId < 0 || bodyId >= (int)bodies_.size() || bodies_[bodyId].id != bodyId) return fail(B2C_E_INVALID, "bad body id");
  HBody& b = bodies_[bodyId];
  if (b.type == kStaticBody) return B2C_OK;
  if (!onDevice_) { b.w = w; return B2C_OK; }
  CK(cudaSetDevice(device_));
  CK(cudaMemcpyAsync(&d_.st0[b.simIndex].z, &w, sizeof(float), cudaMemcpyHostToDevice, stream_));
  CK(cudaStreamSynchronize(stream_));
  return B2C_OK;
}
// Force / impulse / per-body tuning mutators run as a one-thread kernel ordered on the world's stream: no read-back, no
// synchronisation, and the arithmetic is the reference's expression by expression (fp32, no contraction).
enum BodyOp { OP_FORCE_CENTER, OP_FORCE_POINT, OP_TORQUE, OP_CLEAR_FORCES, OP_IMPULSE_CENTER, OP_IMPULSE_POINT, OP_ANGULAR_IMPULSE,
              OP_LINEAR_DAMPING, OP_ANGULAR_DAMPING, OP_GRAVITY_SCALE };
__global__ void k_body_op(WorldDev d, int op, int sim, V2 a, V2 point, float maxLinearSpeed) {
  float4 f = d.simForce[sim], m = d.simMass[sim], s0 = d.st0[sim];
  V2 center = mk(d.simPose[sim].x, d.simPose[sim].y);
  switch (op) {
    case OP_FORCE_CENTER: f.x = f.x + a.x; f.y = f.y + a.y; d.simForce[sim] = f; break;                                  // api/Body.cs:457-467
    case OP_FORCE_POINT: f.x = f.x + a.x; f.y = f.y + a.y; f.z += cross(point - center, a); d.simForce[sim] = f; break;  // :439-451
    case OP_TORQUE: f.z = a.x; d.simForce[sim] = f; break;                                                               // :469-479 (assigns — port quirk)
    case OP_CLEAR_FORCES: f.x = 0.0f; f.y = 0.0f; f.z = 0.0f; d.simForce[sim] = f; break;                                // :490-495
    case OP_IMPULSE_CENTER: case OP_IMPULSE_POINT: {                                                                     // :507-523, :531-545
      V2 v = mulAdd(mk(s0.x, s0.y), m.x, a);
      if (op == OP_IMPULSE_POINT) s0.z += m.y * cross(point - center, a);
      float v2 = lengthSq(v);                                                                                            // src/Body.cs:131-135
      if (v2 > maxLinearSpeed * maxLinearSpeed) v = (maxLinearSpeed / sqrtf(v2)) * v;
      s0.x = v.x; s0.y = v.y; d.st0[sim] = s0;
    } break;
    case OP_ANGULAR_IMPULSE: s0.z += m.y * a.x; d.st0[sim] = s0; break;                                                  // :554-568
    case OP_LINEAR_DAMPING: m.z = a.x; d.simMass[sim] = m; break;                                                        // :624-629
    case OP_ANGULAR_DAMPING: m.w = a.x; d.simMass[sim] = m; break;                                                       // :635-640
    case OP_GRAVITY_SCALE: f.w = a.x; d.simForce[sim] = f; break;                                                        // :647-653
  }
}
int World::bodyOp(int bodyId, int op, V2 a, V2 point) {
  if (bodyId < 0 || bodyId >= (int)bodies_.size() || bodies_[bodyId].id != bodyId) return fail(B2C_E_INVALID, "bad body id");
  HBody& b = bodies_[bodyId];
  const bool tuning = op == OP_LINEAR_DAMPING || op == OP_ANGULAR_DAMPING || op == OP_GRAVITY_SCALE;
  if (!tuning && op != OP_CLEAR_FORCES && b.type != kDynamicBody) return B2C_OK;
  if (!onDevice_) {                                 // before the first step the host mirror is the world
    switch (op) {
      case OP_FORCE_CENTER: b.force = b.force + a; break;
      case OP_FORCE_POINT: b.force = b.force + a; b.torque += cross(point - b.center, a); break;
      case OP_TORQUE: b.torque = a.x; break;
      case OP_CLEAR_FORCES: b.force = mk(0, 0); b.torque = 0; break;
      case OP_IMPULSE_CENTER: case OP_IMPULSE_POINT: {
        V2 v = mulAdd(b.v, b.invMass, a);
        if (op == OP_IMPULSE_POINT) b.w += b.invInertia * cross(point - b.center, a);
        float v2 = lengthSq(v);
        if (v2 > maxLinearSpeed * maxLinearSpeed) v = (maxLinearSpeed / sqrtf(v2)) * v;
        b.v = v;
      } break;
      case OP_ANGULAR_IMPULSE: b.w += b.invInertia * a.x; break;
      case OP_LINEAR_DAMPING: b.linearDamping = a.x; break;
      case OP_ANGULAR_DAMPING: b.angularDamping = a.x; break;
      case OP_GRAVITY_SCALE: b.gravityScale = a.x; break;
    }
    return B2C_OK;
  }
  if (op == OP_LINEAR_DAMPING) b.linearDamping = a.x; else if (op == OP_ANGULAR_DAMPING) b.angularDamping = a.x; else if (op == OP_GRAVITY_SCALE) b.gravityScale = a.x;
  if (b.simIndex < 0) {                            // static bodies have no solver rows; ClearForces still reaches a parked sim's force (api/Body.cs:490-495)
    if (op == OP_CLEAR_FORCES) { CK(cudaSetDevice(device_)); CK(cudaMemsetAsync(&d_.bodyForce[bodyId], 0, sizeof(float4), stream_)); }
    return B2C_OK;
  }
  CK(cudaSetDevice(device_));
  k_body_op<<<1, 1, 0, stream_>>>(d_, op, b.simIndex, a, point, maxLinearSpeed);
  CK(cudaGetLastError());
  return B2C_OK;
}
int World::applyForceToCenter(int bodyId, V2 f) { return bodyOp(bodyId, OP_FORCE_CENTER, f, mk(0, 0)); }
int World::applyForce(int bodyId, V2 f, V2 p) { return bodyOp(bodyId, OP_FORCE_POINT, f, p); }
int World::applyTorque(int bodyId, float t) { return bodyOp(bodyId, OP_TORQUE, mk(t, 0), mk(0, 0)); }
int World::clearForces(int bodyId) { return bodyOp(bodyId, OP_CLEAR_FORCES, mk(0, 0), mk(0, 0)); }
int World::applyLinearImpulseToCenter(int bodyId, V2 i) { return bodyOp(bodyId, OP_IMPULSE_CENTER, i, mk(0, 0)); }
int World::applyLinearImpulse(int bodyId, V2 i, V2 p) { return bodyOp(bodyId, OP_IMPULSE_POINT, i, p); }
int World::applyAngularImpulse(int bodyId, float i) { return bodyOp(bodyId, OP_ANGULAR_IMPULSE, mk(i, 0), mk(0, 0)); }
int World::setLinearDamping(int bodyId, float v) { return bodyOp(bodyId, OP_LINEAR_DAMPING, mk(v, 0), mk(0, 0)); }
int World::setAngularDamping(int bodyId, float v) { return bodyOp(bodyId, OP_ANGULAR_DAMPING, mk(v, 0), mk(0, 0)); }
int World::setGravityScale(int bodyId, float v) { return bodyOp(bodyId, OP_GRAVITY_SCALE, mk(v, 0), mk(0, 0)); }

// BodyAPI.SetTargetTransform api/Body.cs:378-400, SetMassData :585-598, SetMotionLocks :796-818; ShapeAPI.SetSurfaceMaterial api/Shape.cs:110-146.
__global__ void k_body_target_transform(WorldDev d, int bodyId, Xf target, float timeStep) {
  int sim = d.bodySim[bodyId];
  if (sim < 0) return;
  float4 pose = d.simPose[sim], loc = d.simLocal[sim], s0 = d.st0[sim];
  V2 center1 = mk(pose.x, pose.y), center2 = xfPoint(target, mk(loc.x, loc.y));
  float invTimeStep = 1.0f / timeStep;
  V2 linearVelocity = invTimeStep * (center2 - center1);
  Rot q1 = mkRot(pose.z, pose.w), q2 = target.q;
  float deltaAngle = atan2f(q1.c * q2.s - q1.s * q2.c, q1.c * q2.c + q1.s * q2.s);     // Rotation.RelativeAngle include/MathFunctions.cs:162
  s0.x = linearVelocity.x; s0.y = linearVelocity.y; s0.z = invTimeStep * deltaAngle;
  d.st0[sim] = s0;
}
__global__ void k_body_mass_data(WorldDev d, int bodyId, float mass, float inertia, V2 localCenter) {
  float4 xf4 = d.bodyXf[bodyId];
  V2 center = xfPoint(mkXf(mk(xf4.x, xf4.y), mkRot(xf4.z, xf4.w)), localCenter);
  float invMass = mass > 0.0f ? 1.0f / mass : 0.0f, invInertia = inertia > 0.0f ? 1.0f / inertia : 0.0f;
  d.bodyMass[bodyId] = make_float4(invMass, invInertia, localCenter.x, localCenter.y);
  d.bodyCenter[bodyId] = make_float2(center.x, center.y);
  int sim = d.bodySim[bodyId];
  if (sim < 0) return;
  float4 m = d.simMass[sim]; m.x = invMass; m.y = invInertia; d.simMass[sim] = m;
  float4 loc = d.simLocal[sim]; loc.x = localCenter.x; loc.y = localCenter.y; d.simLocal[sim] = loc;
  float4 pose = d.simPose[sim]; pose.x = center.x; pose.y = center.y; d.simPose[sim] = pose;
  float4 pose0 = d.simPose0[sim]; pose0.x = center.x; pose0.y = center.y; d.simPose0[sim] = pose0;
}
// World.UpdateBodyMassData src/Body.cs:255-353 after the upload: the host sums the shapes (geometry and densities never change on the
// device), this patch does the part that depends on the body's current pose and velocity.
__global__ void k_body_mass_update(WorldDev d, int bodyId, int type, float invMass, float invInertia, V2 localCenter, float minExtent, float maxExtent) {
  float4 xf4 = d.bodyXf[bodyId];
  int sim = d.bodySim[bodyId];
  if (type != kDynamicBody) {                      // :262-279
    d.bodyMass[bodyId] = make_float4(0, 0, 0, 0);
    d.bodyCenter[bodyId] = make_float2(xf4.x, xf4.y);
    if (sim < 0) return;
    float4 m = d.simMass[sim]; m.x = 0; m.y = 0; d.simMass[sim] = m;
    d.simLocal[sim] = make_float4(0, 0, minExtent, maxExtent);
    float4 pose = d.simPose[sim]; pose.x = xf4.x; pose.y = xf4.y; d.simPose[sim] = pose;
    float4 pose0 = d.simPose0[sim]; pose0.x = xf4.x; pose0.y = xf4.y; d.simPose0[sim] = pose0;
    return;
  }
  V2 center = xfPoint(mkXf(mk(xf4.x, xf4.y), mkRot(xf4.z, xf4.w)), localCenter);
  d.bodyMass[bodyId] = make_float4(invMass, invInertia, localCenter.x, localCenter.y);
  d.bodyCenter[bodyId] = make_float2(center.x, center.y);
  if (sim < 0) return;
  float4 m = d.simMass[sim]; m.x = invMass; m.y = invInertia; d.simMass[sim] = m;
  d.simLocal[sim] = make_float4(localCenter.x, localCenter.y, minExtent, maxExtent);
  float4 pose = d.simPose[sim];
  V2 oldCenter = mk(pose.x, pose.y);
  pose.x = center.x; pose.y = center.y; d.simPose[sim] = pose;
  float4 pose0 = d.simPose0[sim]; pose0.x = center.x; pose0.y = center.y; d.simPose0[sim] = pose0;
  float4 st = d.st0[sim];
  V2 dv = crossSV(st.z, center - oldCenter);      // :338-339
  st.x = st.x + dv.x; st.y = st.y + dv.y; d.st0[sim] = st;
}
__global__ void k_body_motion_locks(WorldDev d, int bodyId, uint32_t newFlags) {
  const uint32_t allLocks = kLockLinearX | kLockLinearY | kLockAngularZ;
  uint32_t bf = d.bodyFlags[bodyId];
  if ((bf & allLocks) == newFlags) return;
  d.bodyFlags[bodyId] = bf | allLocks;                     // `body.flags |= allLocks` as written (api/Body.cs:807)
  int sim = d.bodySim[bodyId];
  if (sim < 0) { d.bodySimFlags[bodyId] = (d.bodySimFlags[bodyId] & ~allLocks) | newFlags; return; }      // the static sim's flags (:809-810)
  uint32_t sf = (d.simFlags[sim] & ~allLocks) | newFlags;
  d.simFlags[sim] = sf;
  float4 s0 = d.st0[sim];
  s0.w = __uint_as_float(sf);
  if (newFlags & kLockLinearX) s0.x = 0.0f;
  if (newFlags & kLockLinearY) s0.y = 0.0f;
  if (newFlags & kLockAngularZ) s0.z = 0.0f;
  d.st0[sim] = s0;
}
__global__ void k_shape_material(WorldDev d, int shapeId, float4 material) { d.shapes[shapeId].material = material; }

int World::setTargetTransform(int bodyId, Xf target, float timeStep) {
  if (bodyId < 0 || bodyId >= (int)bodies_.size() || bodies_[bodyId].id != bodyId) return fail(B2C_E_INVALID, "bad body id");
  HBody& b = bodies_[bodyId];
  if (b.type == kStaticBody || timeStep <= 0.0f) return B2C_OK;
  if (!onDevice_) {
    V2 center2 = xfPoint(target, b.localCenter);
    float invTimeStep = 1.0f / timeStep;
    b.v = invTimeStep * (center2 - b.center);
    Rot q1 = b.xf.q, q2 = target.q;
    b.w = invTimeStep * atan2f(q1.c * q2.s - q1.s * q2.c, q1.c * q2.c + q1.s * q2.s);
    return B2C_OK;
  }
  CK(cudaSetDevice(device_));
  k_body_target_transform<<<1, 1, 0, stream_>>>(d_, bodyId, target, timeStep);
  CK(cudaGetLastError());
  return B2C_OK;
}
int World::setMassData(int bodyId, float mass, float inertia, V2 center) {
  if (bodyId < 0 || bodyId >= (int)bodies_.size() || bodies_[bodyId].id != bodyId) return fail(B2C_E_INVALID, "bad body id");
  HBody& b = bodies_[bodyId];
  b.mass = mass; b.inertia = inertia;
  if (!onDevice_) {
    b.localCenter = center; b.center = xfPoint(b.xf, center); b.center0 = b.center;
    b.invMass = mass > 0.0f ? 1.0f / mass : 0.0f; b.invInertia = inertia > 0.0f ? 1.0f / inertia : 0.0f;
    return B2C_OK;
  }
  CK(cudaSetDevice(device_));
  k_body_mass_data<<<1, 1, 0, stream_>>>(d_, bodyId, mass, inertia, center);
  CK(cudaGetLastError());
  return B2C_OK;
}
int World::setMotionLocks(int bodyId, bool lx, bool ly, bool az) {
  if (bodyId < 0 || bodyId >= (int)bodies_.size() || bodies_[bodyId].id != bodyId) return fail(B2C_E_INVALID, "bad body id");
  HBody& b = bodies_[bodyId];
  const uint32_t allLocks = kLockLinearX | kLockLinearY | kLockAngularZ;
  uint32_t newFlags = (lx ? kLockLinearX : 0u) | (ly ? kLockLinearY : 0u) | (az ? kLockAngularZ : 0u);
  if (!onDevice_) {
    // before the upload HBody::flags stands for both Body.flags and BodySim.flags: the sim's locks are what the solver reads
    if ((b.bodyFlags & allLocks) == newFlags) return B2C_OK;     // tested on Body.flags, which holds all three bits after the first change (:805-807)
    b.flags = (b.flags & ~allLocks) | newFlags; b.bodyFlags |= allLocks;
    if (lx) b.v.x = 0.0f;
    if (ly) b.v.y = 0.0f;
    if (az) b.w = 0.0f;
    lockQuirk_.push_back(bodyId);
    return B2C_OK;
  }
  if ((b.bodyFlags & allLocks) != newFlags) { b.bodyFlags |= allLocks; b.flags = (b.flags & ~allLocks) | newFlags; }   // the kernel's branch, mirrored for SetType
  CK(cudaSetDevice(device_));
  k_body_motion_locks<<<1, 1, 0, stream_>>>(d_, bodyId, newFlags);
  CK(cudaGetLastError());
  return B2C_OK;
}
int World::setSurfaceMaterial(int shapeId, float friction, float restitution, float rolling, float tangentSpeed) {
  if (shapeId < 0 || shapeId >= (int)shapes_.size() || shapes_[shapeId].id != shapeId) return fail(B2C_E_INVALID, "bad shape id");
  HShape& sh = shapes_[shapeId];
  sh.friction = friction; sh.restitution = restitution; sh.rollingResistance = rolling; sh.tangentSpeed = tangentSpeed;
  if (!onDevice_) return B2C_OK;
  CK(cudaSetDevice(device_));
  k_shape_material<<<1, 1, 0, stream_>>>(d_, shapeId, make_float4(friction, restitution, rolling, tangentSpeed));
  CK(cudaGetLastError());
  return B2C_OK;
}

// ShapeAPI.EnableContactEvents api/Shape.cs:182-186 / EnableHitEvents :205-209: the flag is read when a contact is created
// (src/Contact.cs:188) or starts touching (:386), contacts that exist keep what they have.
__global__ void k_shape_flag(WorldDev d, int shapeId, uint32_t bit, int on) { uint32_t f = d.shapes[shapeId].flags; d.shapes[shapeId].flags = on ? (f | bit) : (f & ~bit); }
int World::shapeEnableEvents(int shapeId, int which, bool flag) {
  if (shapeId < 0 || shapeId >= (int)shapes_.size() || shapes_[shapeId].id != shapeId) return fail(B2C_E_INVALID, "bad shape id");
  if (which != B2C_SHAPE_CONTACT_EVENTS && which != B2C_SHAPE_HIT_EVENTS) return fail(B2C_E_INVALID, "bad event selector");
  const uint32_t bit = which == B2C_SHAPE_CONTACT_EVENTS ? SF_CONTACT_EVENTS : SF_HIT_EVENTS;
  HShape& sh = shapes_[shapeId];
  sh.flags = flag ? (sh.flags | bit) : (sh.flags & ~bit);
  if (!onDevice_) return B2C_OK;
  CK(cudaSetDevice(device_));
  k_shape_flag<<<1, 1, 0, stream_>>>(d_, shapeId, bit, flag ? 1 : 0);
  CK(cudaGetLastError());
  return B2C_OK;
}
int World::bodyEnableEvents(int bodyId, int which, bool flag) {     // BodyAPI.EnableContactEvents api/Body.cs:848-858 / EnableHitEvents :862-872
  if (bodyId < 0 || bodyId >= (int)bodies_.size() || bodies_[bodyId].id != bodyId) return fail(B2C_E_INVALID, "bad body id");
  for (int sid = bodies_[bodyId].headShapeId; sid != -1; sid = shapes_[sid].nextShapeId) { int rc = shapeEnableEvents(sid, which, flag); if (rc) return rc; }
  return B2C_OK;
}

// WorldAPI.Explode api/World.cs:703-776: one thread walks the dynamic tree in the reference's Query order (child2 popped first) and applies
// the impulses in that order (a body with several shapes accumulates them in callback order). The point proxy is built with radius 1 as
// the port writes it (`Distance.MakeProxy([position], 1)`, :721 — upstream passes count 1, radius 0).
__global__ void k_explode(WorldDev d, V2 position, float radius, float falloff, float impulsePerLength, uint64_t maskBits) {
  const TreeDev& t = d.trees[kDynamicBody];
  int root = t.meta[0];
  if (t.meta[1] == 0 || root == -1) return;
  const float reach = radius + falloff;
  Box query; query.lo = mk(position.x - reach, position.y - reach); query.hi = mk(position.x + reach, position.y + reach);
  DistProxy point; for (int i = 0; i < kMaxPolygonVertices; i++) point.pts[i] = mk(0, 0);
  point.pts[0] = position; point.count = 1; point.radius = 1.0f;
  const Xf identity = mkXf(mk(0, 0), mkRot(1, 0));
  constexpr int kExplodeStack = 256;
  int stack[kExplodeStack]; int sp = 0;
  stack[sp++] = root;
  while (sp > 0) {
    int id = stack[--sp];
    if (!boxOverlaps(tvBox(TreeView{t.aabb, t.child, t.parent, t.hf, t.cat, t.user, t.meta}, id), query) || (t.cat[id] & maskBits) == 0) continue;
    if (!((t.hf[id] >> 16) & NODE_LEAF)) { if (sp + 2 <= kExplodeStack) { int2 c = t.child[id]; stack[sp++] = c.x; stack[sp++] = c.y; } else atomicOr(&d.sc->errorFlags, (int)ERR_STACK_OVERFLOW); continue; }
    const ShapeDev& shape = d.shapes[t.user[id]];
    const int bodyId = shape.bodyId;
    float4 xf4 = d.bodyXf[bodyId];
    Xf xf = mkXf(mk(xf4.x, xf4.y), mkRot(xf4.z, xf4.w));
    DistOut out = shapeDistance(makeDistProxy(shape.geom), point, xf, identity, true);
    if (out.distance > reach) continue;
    const int sim = d.bodySim[bodyId];
    if (sim < 0) continue;
    V2 closest = out.pointA;
    if (out.distance == 0.0f) closest = xfPoint(xf, geomCentroid(shape.geom));
    V2 direction = closest - position;
    direction = lengthSq(direction) > 100.0f * FLT_EPSILON * FLT_EPSILON ? normalize(direction) : mk(1, 0);
    V2 localLine = invRotate(xf.q, leftPerp(direction));
    float perimeter = geomProjectedPerimeter(shape.geom, localLine);
    float scale = 1.0f;
    if (out.distance > radius && falloff > 0.0f) scale = netClamp((radius + falloff - out.distance) / falloff, 0.0f, 1.0f);
    float magnitude = impulsePerLength * perimeter * scale;
    V2 impulse = magnitude * direction;
    float4 s0 = d.st0[sim], m = d.simMass[sim], pose = d.simPose[sim];
    V2 v = mulAdd(mk(s0.x, s0.y), m.x, impulse);
    s0.x = v.x; s0.y = v.y;
    s0.z += m.y * cross(closest - mk(pose.x, pose.y), impulse);
    d.st0[sim] = s0;
  }
}
int World::explode(V2 position, float radius, float falloff, float impulsePerLength, uint64_t maskBits) {
  if (!(radius >= 0.0f) || !(falloff >= 0.0f)) return fail(B2C_E_INVALID, "explode: radius and falloff must be >= 0");
  if (!onDevice_) { int rc = upload(); if (rc) return rc; }
  CK(cudaSetDevice(device_));
  k_explode<<<1, 1, 0, stream_>>>(d_, position, radius, falloff, impulsePerLength, maskBits);
  CK(cudaGetLastError());
  bodiesFetched_ = false;
  return B2C_OK;
}

// ---- read-back ----
template <class T> static int fetch(std::vector<T>& out, const T* dev, size_t n, cudaStream_t s) {
  out.resize(n);
  if (n == 0) return 0;
  return cudaMemcpyAsync(out.data(), dev, n * sizeof(T), cudaMemcpyDeviceToHost, s) == cudaSuccess ? 0 : 1;
}

int World::readBodies(B2CBodyRead* dst, int capacity, int* count) {
  int n = (int)bodies_.size();
  *count = n;
  if (!onDevice_) {
    for (int i = 0; i < n && i < capacity; i++) {
      const HBody& b = bodies_[i]; B2CBodyRead r; memset(&r, 0, sizeof(r));
      r.transform.p = {b.xf.p.x, b.xf.p.y}; r.transform.q = {b.xf.q.c, b.xf.q.s}; r.center = {b.center.x, b.center.y};
      r.linearVelocity = {b.v.x, b.v.y}; r.angularVelocity = b.w; r.flags = b.flags;
      dst[i] = r;
    }
    return B2C_OK;
  }
  int rc = synchronize(); if (rc) return rc;
  std::vector<float4> xf, st0; std::vector<float2> ctr; std::vector<uint32_t> sf, bsf;
  if (fetch(bsf, d_.bodySimFlags, n, stream_) | fetch(xf, d_.bodyXf, n, stream_) | fetch(ctr, d_.bodyCenter, n, stream_) | fetch(st0, d_.st0, d_.awakeCount, stream_) | fetch(sf, d_.simFlags, d_.awakeCount, stream_))
    return fail(B2C_E_CUDA, "read_bodies copy failed");
  CK(cudaStreamSynchronize(stream_));
  for (int i = 0; i < n && i < capacity; i++) {
    const HBody& b = bodies_[i]; B2CBodyRead r; memset(&r, 0, sizeof(r));
    if (b.id != -1) {
      r.transform.p = {xf[i].x, xf[i].y}; r.transform.q = {xf[i].z, xf[i].w}; r.center = {ctr[i].x, ctr[i].y};
      r.flags = bsf[i];
      if (b.simIndex >= 0) { float4 s = st0[b.simIndex]; r.linearVelocity = {s.x, s.y}; r.angularVelocity = s.z; r.flags = sf[b.simIndex]; }
    }
    dst[i] = r;
  }
  return B2C_OK;
}

// BodyAPI.GetTransform / GetLinearVelocity / GetAngularVelocity / GetWorldCenterOfMass for one body (api/Body.cs:288, :345-357, :580)
int World::readBody(int bodyId, B2CBodyRead* out) {
  if (bodyId < 0 || bodyId >= (int)bodies_.size() || bodies_[bodyId].id == -1) return fail(B2C_E_INVALID, "bad body id");
  const HBody& b = bodies_[bodyId]; B2CBodyRead r; memset(&r, 0, sizeof(r));
  if (!onDevice_) {
    r.transform.p = {b.xf.p.x, b.xf.p.y}; r.transform.q = {b.xf.q.c, b.xf.q.s}; r.center = {b.center.x, b.center.y};
    r.linearVelocity = {b.v.x, b.v.y}; r.angularVelocity = b.w; r.flags = b.flags;
    *out = r; return B2C_OK;
  }
  int rc = synchronize(); if (rc) return rc;
  float4 xf; float2 ctr; float4 s = make_float4(0, 0, 0, 0); uint32_t sf = b.flags;
  CK(cudaMemcpyAsync(&sf, &d_.bodySimFlags[bodyId], sizeof(sf), cudaMemcpyDeviceToHost, stream_));
  CK(cudaMemcpyAsync(&xf, &d_.bodyXf[bodyId], sizeof(xf), cudaMemcpyDeviceToHost, stream_));
  CK(cudaMemcpyAsync(&ctr, &d_.bodyCenter[bodyId], sizeof(ctr), cudaMemcpyDeviceToHost, stream_));
  if (b.simIndex >= 0) {
    CK(cudaMemcpyAsync(&s, &d_.st0[b.simIndex], sizeof(s), cudaMemcpyDeviceToHost, stream_));
    CK(cudaMemcpyAsync(&sf, &d_.simFlags[b.simIndex], sizeof(sf), cudaMemcpyDeviceToHost, stream_));
  }
  CK(cudaStreamSynchronize(stream_));
  r.transform.p = {xf.x, xf.y}; r.transform.q = {xf.z, xf.w}; r.center = {ctr.x, ctr.y};
  r.linearVelocity = {s.x, s.y}; r.angularVelocity = s.z; r.flags = sf;
  *out = r;
  return B2C_OK;
}

int World::readContacts(B2CContactRead* dst, int capacity, int* count) {
  *count = 0;
  if (!onDevice_) return B2C_OK;
  int rc = synchronize(); if (rc) return rc;
  int n = hostScalars_->contactNext;
  std::vector<int2> cShape; std::vector<uint32_t> cFlags; std::vector<int> cColor, cLocal; std::vector<float4> cMat, mN, mA0, mA1, mB0, mB1, mC0, mC1;
  if (fetch(cShape, d_.cShape, n, stream_) | fetch(cFlags, d_.cFlags, n, stream_) | fetch(cColor, d_.cColor, n, stream_) | fetch(cLocal, d_.cLocal, n, stream_) | fetch(cMat, d_.cMat, n, stream_) |
      fetch(mN, d_.mN, n, stream_) | fetch(mA0, d_.mA0, n, stream_) | fetch(mA1, d_.mA1, n, stream_) | fetch(mB0, d_.mB0, n, stream_) |
      fetch(mB1, d_.mB1, n, stream_) | fetch(mC0, d_.mC0, n, stream_) | fetch(mC1, d_.mC1, n, stream_))
    return fail(B2C_E_CUDA, "read_contacts copy failed");
  CK(cudaStreamSynchronize(stream_));
  int k = 0;
  for (int i = 0; i < n; i++) {
    if (!(cFlags[i] & CF_ALIVE)) continue;
    if (k < capacity) {
      B2CContactRead r; memset(&r, 0, sizeof(r));
      r.contactId = i; r.shapeIdA = cShape[i].x; r.shapeIdB = cShape[i].y; r.colorIndex = cColor[i]; r.localIndex = cColor[i] != -1 ? cLocal[i] : -1;
      r.flags = (cFlags[i] & CF_TOUCHING) ? 1u : 0u;
      r.pointCount = __builtin_bit_cast(int, mN[i].w); r.normal = {mN[i].x, mN[i].y}; r.rollingImpulse = mN[i].z;
      r.friction = cMat[i].x; r.restitution = cMat[i].y; r.rollingResistance = cMat[i].z; r.tangentSpeed = cMat[i].w;
      auto fill = [](B2CManifoldPointRead& o, float4 a, float4 b, float4 c) {
        o.anchorA = {a.x, a.y}; o.anchorB = {a.z, a.w}; o.separation = b.x; o.normalImpulse = b.y; o.tangentImpulse = b.z; o.totalNormalImpulse = b.w;
        o.point = {c.x, c.y}; o.normalVelocity = c.z; uint32_t w = __builtin_bit_cast(uint32_t, c.w); o.id = w & 0xffffu; o.persisted = (w >> 16) & 1u;
      };
      fill(r.points[0], mA0[i], mB0[i], mC0[i]); fill(r.points[1], mA1[i], mB1[i], mC1[i]);
      dst[k] = r;
    }
    k++;
  }
  *count = k;
  return B2C_OK;
}

int World::readJoints(B2CJointRead* dst, int capacity, int* count) {
  int n = (int)joints_.size();
  *count = n;
  std::vector<JointDev> jd;
  if (onDevice_) {
    int rc = synchronize(); if (rc) return rc;
    if (fetch(jd, d_.joints, n, stream_)) return fail(B2C_E_CUDA, "read_joints copy failed");
    CK(cudaStreamSynchronize(stream_));
  } else { jd.resize(n); for (int i = 0; i < n; i++) jd[i] = joints_[i].sim; }
  int m = 0;
  for (int i = 0; i < n; i++) {
    if (joints_[i].id == -1 || !joints_[i].inGraph) continue;   // destroyed (the slot waits in the id pool) or parked in the static set
    const JointDev& s = jd[i];
    if (m < capacity) dst[m] = {s.jointId, s.colorIndex, {s.linearImpulse.x, s.linearImpulse.y}, s.springImpulse, s.motorImpulse, s.lowerImpulse, s.upperImpulse, s.type, {s.linearImpulse2.x, s.linearImpulse2.y}};
    m++;
  }
  *count = m;
  return B2C_OK;
}

int World::readShapes(B2CShapeRead* dst, int capacity, int* count) {
  int n = (int)shapes_.size();
  *count = n;
  if (!onDevice_) {
    for (int i = 0; i < n && i < capacity; i++) {
      const HShape& s = shapes_[i];
      if (s.id == -1) { dst[i] = {{0, 0}, {0, 0}, {0, 0}, {0, 0}, -1, 0}; continue; }
      dst[i] = {{s.aabb.lo.x, s.aabb.lo.y}, {s.aabb.hi.x, s.aabb.hi.y}, {s.fat.lo.x, s.fat.lo.y}, {s.fat.hi.x, s.fat.hi.y}, s.proxyKey, 0};
    }
    return B2C_OK;
  }
  int rc = synchronize(); if (rc) return rc;
  std::vector<float4> a, f; std::vector<ShapeDev> sh;
  if (fetch(a, d_.shapeAABB, n, stream_) | fetch(f, d_.shapeFat, n, stream_) | fetch(sh, d_.shapes, n, stream_)) return fail(B2C_E_CUDA, "read_shapes copy failed");
  CK(cudaStreamSynchronize(stream_));
  for (int i = 0; i < n && i < capacity; i++) {
    if (shapes_[i].id == -1) { dst[i] = {{0, 0}, {0, 0}, {0, 0}, {0, 0}, -1, 0}; continue; }      // destroyed with its body
    dst[i] = {{a[i].x, a[i].y}, {a[i].z, a[i].w}, {f[i].x, f[i].y}, {f[i].z, f[i].w}, sh[i].proxyKey, (sh[i].flags & SF_ENLARGED) ? 1 : 0};
  }
  return B2C_OK;
}

int World::getCounters(B2CCounters* out) {
  memset(out, 0, sizeof(*out));
  out->bodyCount = bodyIds_.count(); out->awakeBodyCount = (int)awakeOrder_.size(); out->shapeCount = shapeIds_.count(); out->jointCount = jointIds_.count();
  out->unsupportedCount = unsupported; out->stepCount = stepCount_;
  if (!onDevice_) {
    out->moveCount = (int)moveArray_.size();
    for (int c = 0; c < kGraphColorCount; c++) out->colorCounts[c] = (int)colorJoints_[c].size();
    return B2C_OK;
  }
  int rc = synchronize();
  const Scalars& s = *hostScalars_;
  out->contactCount = s.liveContactCount; out->touchingContactCount = s.touchingCount; out->contactIdCapacity = s.contactNext;
  int active = 0;
  for (int c = 0; c < kGraphColorCount; c++) {
    out->colorCounts[c] = s.colorTotal[c] + (d_.jointColorStart[c + 1] - d_.jointColorStart[c]);
    if (c < kOverflowIndex && out->colorCounts[c] > 0) active++;
  }
  out->activeColorCount = active; out->moveCount = s.moveCount; out->newPairCount = s.newPairCount; out->toiCandidates = s.toiCandidates;
  out->errorFlags = s.errorFlags;
  return rc;
}

int World::getContactEvents(const B2CContactTouchEvent** begin, int* nb, const B2CContactTouchEvent** end, int* ne, const B2CContactHitEvent** hit, int* nh) {
  *begin = nullptr; *end = nullptr; *hit = nullptr; *nb = *ne = *nh = 0;
  if (!onDevice_) return B2C_OK;
  int rc = synchronize(); if (rc) return rc;
  if (!eventsFetched_) {
    const Scalars& s = *hostScalars_;
    hBeginN_ = s.beginCount; hEndN_ = s.endCount[lastEndIndex_]; hHitN_ = s.hitCount;
    if (hBeginN_ > d_.eventCapacity) hBeginN_ = d_.eventCapacity;
    if (hEndN_ > d_.eventCapacity) hEndN_ = d_.eventCapacity;
    if (hHitN_ > d_.eventCapacity) hHitN_ = d_.eventCapacity;
    if (batch_) {                                  // batch member: the event arrays are device memory
      stageBegin_.resize(std::max(hBeginN_, 1)); stageEnd_.resize(std::max(hEndN_, 1)); stageHit_.resize(std::max(hHitN_, 1));
      if (hBeginN_) CK(cudaMemcpy(stageBegin_.data(), d_.beginEvents, sizeof(B2CContactTouchEvent) * hBeginN_, cudaMemcpyDeviceToHost));
      if (hEndN_) CK(cudaMemcpy(stageEnd_.data(), d_.endEvents[lastEndIndex_], sizeof(B2CContactTouchEvent) * hEndN_, cudaMemcpyDeviceToHost));
      if (hHitN_) CK(cudaMemcpy(stageHit_.data(), d_.hitEvents, sizeof(B2CContactHitEvent) * hHitN_, cudaMemcpyDeviceToHost));
    }
    eventsFetched_ = true;
  }
  if (batch_) { *begin = stageBegin_.data(); *nb = hBeginN_; *end = stageEnd_.data(); *ne = hEndN_; *hit = stageHit_.data(); *nh = hHitN_; return B2C_OK; }
  *begin = hBegin_; *nb = hBeginN_; *end = hEnd_[lastEndIndex_]; *ne = hEndN_; *hit = hHit_; *nh = hHitN_;
  return B2C_OK;
}

int World::getSensorEvents(const B2CSensorEvent** begin, int* nb, const B2CSensorEvent** end, int* ne) {   // WorldAPI.GetSensorEvents api/World.cs:289-299
  *begin = nullptr; *end = nullptr; *nb = 0; *ne = 0;
  if (!onDevice_) return B2C_OK;
  int rc = synchronize(); if (rc) return rc;
  int b = hostScalars_->sensorBeginCount, e = hostScalars_->sensorEndCount[lastEndIndex_];
  if (lastStepDtZero_) b = 0;
  b = std::min(b, d_.sensorEventCapacity); e = std::min(e, d_.sensorEventCapacity);
  if (batch_) {                                    // batch member: the event arrays are device memory
    stageSensorBegin_.resize(std::max(b, 1)); stageSensorEnd_.resize(std::max(e, 1));
    if (b) CK(cudaMemcpy(stageSensorBegin_.data(), d_.sensorBeginEvents, sizeof(B2CSensorEvent) * b, cudaMemcpyDeviceToHost));
    if (e) CK(cudaMemcpy(stageSensorEnd_.data(), d_.sensorEndEvents[lastEndIndex_], sizeof(B2CSensorEvent) * e, cudaMemcpyDeviceToHost));
    *begin = stageSensorBegin_.data(); *end = stageSensorEnd_.data();
  } else { *begin = hSensorBegin_; *end = hSensorEnd_[lastEndIndex_]; }
  *nb = b; *ne = e;
  return B2C_OK;
}
__global__ void k_shape_flag(WorldDev d, int shapeId, uint32_t bit, int on);
int World::shapeEnableSensorEvents(int shapeId, bool flag) {          // ShapeAPI.EnableSensorEvents api/Shape.cs:170-174
  if (shapeId < 0 || shapeId >= (int)shapes_.size() || shapes_[shapeId].id != shapeId) return fail(B2C_E_INVALID, "bad shape id");
  HShape& s = shapes_[shapeId];
  s.flags = flag ? (s.flags | SF_SENSOR_EVENTS) : (s.flags & ~(uint32_t)SF_SENSOR_EVENTS);
  if (!onDevice_) return B2C_OK;
  CK(cudaSetDevice(device_));
  k_shape_flag<<<1, 1, 0, stream_>>>(d_, shapeId, SF_SENSOR_EVENTS, flag ? 1 : 0);
  CK(cudaGetLastError());
  return B2C_OK;
}
int World::shapeGetSensorOverlaps(int shapeId, B2CShapeRef* out, int capacity, int* count) {   // ShapeAPI.GetSensorData api/Shape.cs:409-419
  *count = 0;
  if (shapeId < 0 || shapeId >= (int)shapes_.size() || shapes_[shapeId].id != shapeId) return fail(B2C_E_INVALID, "bad shape id");
  const int index = shapes_[shapeId].sensorIndex;
  if (index == -1 || !onDevice_) return B2C_OK;
  int rc = synchronize(); if (rc) return rc;
  CK(cudaStreamSynchronize(stream_));
  Scalars sc; CK(cudaMemcpy(&sc, d_.sc, sizeof(sc), cudaMemcpyDeviceToHost));      // a DestroySensor since the step may have moved lists: read the live parity
  int n = 0; CK(cudaMemcpy(&n, d_.sensorOverlapCount + 2 * index + sc.sensorParity, sizeof(int), cudaMemcpyDeviceToHost));
  n = std::min(n, capacity);
  if (n > 0) {
    std::vector<int2> tmp(n);
    CK(cudaMemcpy(tmp.data(), d_.sensorOverlaps + ((size_t)index * 2 + sc.sensorParity) * kSensorMaxOverlaps, sizeof(int2) * n, cudaMemcpyDeviceToHost));
    for (int i = 0; i < n; i++) { out[i].shapeId = tmp[i].x; out[i].generation = (uint32_t)tmp[i].y; }
  }
  *count = n;
  return B2C_OK;
}
int World::getJointEvents(const int** ids, int* count) {
  *ids = hJointEvents_; *count = 0;
  if (!onDevice_ || joints_.empty()) return B2C_OK;
  int rc = synchronize(); if (rc) return rc;
  rc = ensureHostBuffers(); if (rc) return rc;
  *ids = hJointEvents_;
  std::vector<uint8_t> flags;
  if (fetch(flags, d_.jointEventFlag, joints_.size(), stream_)) return fail(B2C_E_CUDA, "joint event copy failed");
  CK(cudaStreamSynchronize(stream_));
  int n = 0;
  for (size_t i = 0; i < flags.size(); i++) if (flags[i] && joints_[i].id != -1 && joints_[i].inGraph) hJointEvents_[n++] = (int)i;   // a destroyed joint keeps its last flag: prepare only clears live joints
  *count = n;
  return B2C_OK;
}

int World::getProfile(B2CProfile* out) {       // averages over the steps since the previous call
  // Every time is a difference of %globaltimer stamps taken by the step kernels themselves and summed on the device (Scalars::phaseAcc):
  // no CUDA event is recorded in the stream of a step. Island solver = end of the front kernel to the start of the finalize kernel, whole
  // step = start of the front kernel to the end of the finalize kernel (launch gaps between the kernels included, as an event pair would).
  memset(out, 0, sizeof(*out));
  if (!onDevice_) return B2C_OK;
  int rc = synchronize(); if (rc) return rc;
  double ph[8]; const int n = stepCount_ - lastProfStep_ > 0 ? stepCount_ - lastProfStep_ : 1;
  for (int k = 0; k < 8; k++) { ph[k] = (double)(hostScalars_->phaseAcc[k] - lastPhaseAcc_[k]) * 1e-6 / n; lastPhaseAcc_[k] = hostScalars_->phaseAcc[k]; }
  lastProfStep_ = stepCount_;
  out->pairs = (float)ph[0]; out->collide = (float)ph[1]; out->treeRebuild = (float)ph[4]; out->prepareConstraints = (float)ph[2];
  out->solveConstraints = (float)(ph[3] + ph[5]);          // grid-path solver + island solver kernel(s)
  out->transforms = (float)ph[6]; out->refit = out->transforms;
  out->solve = out->prepareConstraints + out->solveConstraints + out->transforms;
  out->step = (float)(batch_ ? ph[0] + ph[1] + ph[2] + ph[3] + ph[5] + ph[6] : ph[7]);   // a batch world's CTA has no stamp 7 of its own meaning: its parts are summed
  return B2C_OK;
}

int World::debugChecksums(unsigned long long* out16) {
  for (int k = 0; k < 16; k++) out16[k] = 0;
  if (!onDevice_) return B2C_OK;
  int rc = synchronize(); if (rc) return rc;
  for (int k = 0; k < 16; k++) out16[k] = hostScalars_->dbg[k];
  return B2C_OK;
}

// The timer brackets device time: the start event sits behind a kernel that holds the stream for kTimerHoldNs, so the launches the caller
// enqueues next are queued by the time it fires and the region does not count the host's enqueue latency of the first of them.
constexpr unsigned long long kTimerHoldNs = 60000;
__global__ void k_timer_hold(unsigned long long ns) {
  unsigned long long t0, t; asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t0));
  do { asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t)); } while (t - t0 < ns);
}
int World::timerStart() {
  if (!onDevice_) { int rc = upload(); if (rc) return rc; }
  CK(cudaSetDevice(device_));
  k_timer_hold<<<1, 1, 0, stream_>>>(kTimerHoldNs);
  CK(cudaGetLastError());
  CK(cudaEventRecord(timer_[0], stream_));
  return B2C_OK;
}
int World::timerStop(float* ms) {
  *ms = 0;
  if (!onDevice_) return fail(B2C_E_INVALID, "timer not started");
  CK(cudaSetDevice(device_));
  CK(cudaEventRecord(timer_[1], stream_));
  CK(cudaEventSynchronize(timer_[1]));
  CK(cudaEventElapsedTime(ms, timer_[0], timer_[1]));
  return B2C_OK;
}

}  // namespace b2c
