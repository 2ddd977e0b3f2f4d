// libb2cuda — host-side creation bookkeeping (the part of Box2D.API that is NOT on the step path:
// id pools, body/shape/joint creation, mass data, proxy insertion into the SAH tree, joint coloring).
// It only builds the arrays that are uploaded to the device; it contains no step code — the step
// runs exclusively on the GPU (b2c_step.cu) and there is no CPU fallback.
#pragma once
#include <vector>
#include <string>
#include "b2c_world.cuh"

namespace b2c {

struct IdPool {                 // src/IDPool.cs (LIFO recycling)
  std::vector<int> freeArray; int nextIndex = 0;
  int alloc() { if (!freeArray.empty()) { int id = freeArray.back(); freeArray.pop_back(); return id; } return nextIndex++; }
  void release(int id) { freeArray.push_back(id); }
  int capacity() const { return nextIndex; }
  int count() const { return nextIndex - (int)freeArray.size(); }
};

// Host replica of the reference's incremental tree, used at creation time only (src/DynamicTree.cs:36-480).
struct HostTree {
  std::vector<float4> aabb; std::vector<int2> child; std::vector<int> parent; std::vector<uint32_t> hf;
  std::vector<uint64_t> cat; std::vector<int> user;
  int root = -1, nodeCount = 0, freeList = 0, proxyCount = 0;
  HostTree();
  int capacity() const { return (int)aabb.size(); }
  int createProxy(Box box, uint64_t categoryBits, int userData);
  void moveProxy(int id, Box box);
  void ensureFreeNode();
};

struct HBody {                  // src/Body.cs:35-93 Body + BodySim + BodyState at creation time
  int id = -1, type = kStaticBody, simIndex = -1;
  int headShapeId = -1, shapeCount = 0, headJointKey = -1, jointCount = 0;
  float mass = 0, inertia = 0, sleepThreshold = 0.05f;
  uint32_t flags = 0, bodyFlags = 0;      // BodySim.flags (persistent bits) / Body.flags: they part ways in SetMotionLocks (api/Body.cs:807) and SetType (:217-218)
  uint16_t generation = 0; uint64_t userData = 0; bool enableSleep = true;
  Xf xf; V2 center, center0, localCenter, force; Rot rot0;
  float torque = 0, invMass = 0, invInertia = 0, minExtent = kHuge, maxExtent = 0, linearDamping = 0, angularDamping = 0, gravityScale = 1;
  V2 v; float w = 0;
};
struct HShape {
  int id = -1, bodyId = -1, prevShapeId = -1, nextShapeId = -1, proxyKey = -1;
  Geom geom; float friction = 0.6f, restitution = 0, rollingResistance = 0, tangentSpeed = 0, density = 1;
  uint64_t categoryBits = ~0ull, maskBits = ~0ull; int groupIndex = 0; uint32_t flags = 0; uint16_t generation = 0;
  Box aabb, fat; uint64_t userData = 0;
  int sensorIndex = -1;           // Shape.sensorIndex: position in World.sensors (sensorShapes_) or -1
};
struct HJoint {
  int id = -1; uint16_t generation = 0; bool collideConnected = false;
  bool inGraph = true;            // false: parked in the static set by SetType (both bodies non-dynamic, SolverSet.TransferJoint) — not simulated, not reported
  int bodyIdA = -1, nextKeyA = -1, prevKeyA = -1, bodyIdB = -1, nextKeyB = -1, prevKeyB = -1;
  JointDev sim;
};

class Batch;
cudaStream_t batchStream(Batch* b);                       // b2c_batch.cu
void* batchArenaAlloc(Batch* b, size_t bytes);
int* batchMappedInt(Batch* b, int** devPtr);
void batchMarkDirty(Batch* b, int index);
class World {
 public:
  explicit World(const B2CWorldDef& def, int device, Batch* batch = nullptr, int batchIndex = -1);
  ~World();
  // --- membership in a batch of independent worlds (b2c_batch.cu): the batch owns the stream and the device memory arena, steps all its
  //     worlds with one launch and keeps a device array of their WorldDev records ---
  bool inBatch() const { return batch_ != nullptr; }
  void markDeviceDirty() { if (batch_ && !devDirty_) { devDirty_ = true; batchMarkDirty(batch_, batchIndex_); } }   // d_ may differ from the batch's device copy of it
  // --- creation API (host) ---
  int createBody(const B2CBodyDef& def, int* bodyId);
  int createShape(int bodyId, const B2CShapeDef& def, const B2CGeometry& g, int* shapeId);
  int createRevoluteJoint(const B2CRevoluteJointDef& def, int* jointId);
  int createJoint(const B2CJointDef& def, int* jointId);
  int destroyJoint(int jointId);
  int destroyBody(int bodyId);
  int destroyShape(int shapeId, bool updateBodyMass);
  int setShapeDensity(int shapeId, float density, bool updateBodyMass);
  int applyMassFromShapes(int bodyId);
  int shapeEnableEvents(int shapeId, int which, bool flag);
  int bodyEnableEvents(int bodyId, int which, bool flag);
  int uploadJointLists();
  void treeProxyRemoved(int type);
  int resetProxy(int shapeId, const Geom* geom, bool setFilter, bool destroyProxy);
  int setShapeFilter(int shapeId, uint64_t categoryBits, uint64_t maskBits, int groupIndex);
  int setShapeGeometry(int shapeId, const B2CGeometry& g);
  int setBodyType(int bodyId, int type);
  int explode(V2 position, float radius, float falloff, float impulsePerLength, uint64_t maskBits);
  int jointOp(int jointId, int patch, int which, float a, float b);
  int jointSetLocalFrame(int jointId, int which, Xf frame);
  int jointSetCollideConnected(int jointId, bool shouldCollide);
  int bodySetSleepThreshold(int bodyId, float v);
  int bodyEnableSleep(int bodyId, bool flag);
  int bodySetAwake(int bodyId, bool awake);
  int bodyWakeTouching(int bodyId);
  int bodySetEnabled(int bodyId, bool enabled);
  int bodySetBullet(int bodyId, bool flag);
  void enableGraphLaunch(bool flag) { graphMode_ = flag ? 1 : 0; }
  int setTransform(int bodyId, V2 p, Rot q);
  int setLinearVelocity(int bodyId, V2 v);
  int setTargetTransform(int bodyId, Xf target, float timeStep);
  int setMassData(int bodyId, float mass, float inertia, V2 center);
  int setMotionLocks(int bodyId, bool lx, bool ly, bool az);
  int setSurfaceMaterial(int shapeId, float friction, float restitution, float rolling, float tangentSpeed);
  int setAngularVelocity(int bodyId, float w);
  int applyForceToCenter(int bodyId, V2 f);
  int applyTorque(int bodyId, float t);
  int applyLinearImpulseToCenter(int bodyId, V2 i);
  int applyForce(int bodyId, V2 f, V2 point);
  int clearForces(int bodyId);
  int applyLinearImpulse(int bodyId, V2 i, V2 point);
  int applyAngularImpulse(int bodyId, float i);
  int setLinearDamping(int bodyId, float v);
  int setAngularDamping(int bodyId, float v);
  int setGravityScale(int bodyId, float v);
  int bodyOp(int bodyId, int op, V2 a, V2 point);
  int readBody(int bodyId, B2CBodyRead* out);
  // --- step + readback (device) ---
  int step(float dt, int subSteps);
  int synchronize();
  int readBodies(B2CBodyRead* dst, int capacity, int* count);
  int getBodyEvents(const B2CBodyRead** records, int* count);
  int readContacts(B2CContactRead* dst, int capacity, int* count);
  int readJoints(B2CJointRead* dst, int capacity, int* count);
  int readShapes(B2CShapeRead* dst, int capacity, int* count);
  int getCounters(B2CCounters* out);
  int getBodyMoveEvents(const B2CBodyMoveEvent** events, int* count);
  int getProfile(B2CProfile* out);
  int debugChecksums(unsigned long long* out16);
  int timerStart();
  int timerStop(float* ms);
  int getContactEvents(const B2CContactTouchEvent** begin, int* nb, const B2CContactTouchEvent** end, int* ne, const B2CContactHitEvent** hit, int* nh);
  int getJointEvents(const int** ids, int* count);
  int getSensorEvents(const B2CSensorEvent** begin, int* nb, const B2CSensorEvent** end, int* ne);
  int shapeEnableSensorEvents(int shapeId, bool flag);
  int shapeGetSensorOverlaps(int shapeId, B2CShapeRef* out, int capacity, int* count);

  // read-only view for b2c_noalloc.cu (b2cuda_world_copy_body_move_events_noalloc): nothing here changes the world
  struct NoAllocView { int device; cudaStream_t stream; bool onDevice, lastStepDtZero; const WorldDev* dev; };
  NoAllocView noAllocView() const { return NoAllocView{device_, stream_, onDevice_, lastStepDtZero_, &d_}; }

  // tunables (api/World.cs setters)
  V2 gravity; float restitutionThreshold, hitEventThreshold, contactHertz, contactDampingRatio, contactSpeed, maxLinearSpeed;
  bool enableSleep, enableContinuous, enableContactSoftening, enableWarmStarting = true, enableSpeculative = true;
  std::string error;
  int unsupported = 0;
  int capHintBody, capHintShape, capHintContact, capHintJoint;
  int capJoint_ = 0; int* hJointColor_ = nullptr; int* hJointColorDev_ = nullptr;   // joint rows; mapped int the colour kernel answers through
  int capBody_ = 0, capAwake_ = 0, capShape_ = 0;                       // device rows allocated at upload (bodies / awake sims / shapes)
  int treeAllocCap_[3] = {0, 0, 0}, treeNodeCount_[3] = {0, 0, 0}, treeProxyCount_[3] = {0, 0, 0};   // allocated nodes per tree; the logical capacity (TreeDev::capacity) grows x1.5 as the reference's does

 private:
  void updateBodyMassData(HBody& body);
  void updateShapeAABBs(HShape& s, Xf xf, int proxyType);
  int assignJointColor(int bodyIdA, int bodyIdB, int typeA, int typeB);
  void bufferMove(int key);
  int upload();                 // host arrays -> device (first step)
  void freeDevice();
  int fail(int code, const std::string& msg) { error = msg; return code; }
  template <class T> T* dalloc(size_t n);

  int device_;
  cudaStream_t stream_ = nullptr;
  bool onDevice_ = false;
  friend class Batch;
  Batch* batch_ = nullptr; int batchIndex_ = -1;
  bool devDirty_ = false;
  bool scalarsStale_ = false;     // batch member: hostScalars_ has not been fetched since the last batch step
  std::vector<std::pair<void*, size_t>> arenaBlocks_;   // batch member: its device arrays inside the batch's arena, in allocation order
  struct HostSnap {               // everything creation / destruction calls change on the host, for Batch::resetWorld
    IdPool bodyIds, shapeIds, jointIds; std::vector<HBody> bodies; std::vector<HShape> shapes; std::vector<HJoint> joints;
    std::vector<int> awakeOrder, sensorShapes; std::vector<std::vector<int>> colorJoints; std::vector<uint32_t> colorMask;
    int treeNodeCount[3], treeProxyCount[3]; WorldDev d; bool valid = false;
  };
  void saveHostSnap(HostSnap& h) const;
  void restoreHostSnap(const HostSnap& h);
  int growSensors(int minSensors);
  int destroySensor(int shapeId);   // World.DestroySensor src/Sensor.cs:219-235
  std::vector<int> sensorShapes_;   // World.sensors: shape id per sensor index
  int capSensor_ = 0;
  B2CSensorEvent* hSensorBegin_ = nullptr; B2CSensorEvent* hSensorEnd_[2] = {nullptr, nullptr};   // mapped pinned (a batch member: device memory + staging)
  std::vector<B2CSensorEvent> stageSensorBegin_, stageSensorEnd_;
  int ensureHostBuffers();        // pinned read-back buffers; a batch member allocates them on first use
  std::vector<B2CContactTouchEvent> stageBegin_, stageEnd_; std::vector<B2CContactHitEvent> stageHit_;   // batch member: events live in device memory and are copied on demand
  // host creation state
  IdPool bodyIds_, shapeIds_, jointIds_;
  std::vector<HBody> bodies_; std::vector<HShape> shapes_; std::vector<HJoint> joints_;
  std::vector<int> awakeOrder_;        // bodyId per awake sim index
  HostTree trees_[3];
  std::vector<int> moveArray_; std::vector<std::vector<uint8_t>> moved_;
  std::vector<uint32_t> colorMask_;    // per bodyId
  std::vector<std::vector<int>> colorJoints_;
  std::vector<int> lockQuirk_;   // bodies whose Body.flags got all three lock bits from SetMotionLocks before the upload (api/Body.cs:807)
  // device state
  WorldDev d_{};
  std::vector<void*> allocs_;
  struct AllocRec { void** slot; size_t elem, count; };
  std::vector<AllocRec> allocRecs_;     // every device array with the member that points at it: lets a class of arrays grow between steps
  int regrow(void** slot, size_t newCount, int fillByte = 0);
  int growBodies(int minBodies, int minAwake);
  int growShapes(int minShapes);
  int growContacts(int minContacts);
  int growJoints(int minJoints);
  int growTree(int type, int minNodes);
  int growScanScratch();
  Scalars* hostScalars_ = nullptr;     // pinned
  int endIndex_ = 0;
  int stepCount_ = 0;
  bool eventsFetched_ = false;
  B2CContactTouchEvent *hBegin_ = nullptr, *hEnd_[2] = {nullptr, nullptr}; B2CContactHitEvent* hHit_ = nullptr; int* hJointEvents_ = nullptr;   // event arrays: mapped pinned host memory the kernels write into
  int hBeginN_ = 0, hEndN_ = 0, hHitN_ = 0, hJointN_ = 0;
  int lastEndIndex_ = 0;
  cudaEvent_t timer_[2] = {};
  int lastProfStep_ = 0;
  // One CUDA graph per distinct step configuration (parameters, end-event buffer parity, solver variant, streamed read-backs): a step is
  // then ONE launch, so the kernels run back to back whatever the host's launch latency is (matters with 8 ranks on one host).
  struct StepGraph { std::vector<unsigned char> key; cudaGraphExec_t exec = nullptr; };
  std::vector<StepGraph> graphs_;
  int graphMode_ = -1;          // -1 undecided (B2CUDA_GRAPH decides at the first step), 0 direct launches (default, or capture failed), 1 graphs
  int enqueueStep(const StepParams& sp, bool warpMode, bool clusterLaunch, int newEndIndex);
  unsigned long long lastPhaseAcc_[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  B2CProfile profile_{};
  bool profileValid_ = false;
  int coopBlocks_ = 0;
  int clusterCtas_ = 0;   // CTAs of the solver cluster (16, 8 or 0 = off)
  bool clusterDsm_ = false;      // the cluster kernel can be launched with kDsmSmemBytes of dynamic shared memory per CTA
  bool clusterDsmFallback_ = false;   // B2CUDA_CLUSTER_DSM=fallback
  B2CBodyRead* dBodyRead_ = nullptr; B2CBodyRead* hBodyRead_ = nullptr; bool bodiesFetched_ = false; bool streamBodies_ = false;
  B2CBodyMoveEvent* dMoveEvents_ = nullptr; B2CBodyMoveEvent* hMoveEvents_ = nullptr; B2CBodyMoveEvent* hMoveEventsDev_ = nullptr;   // device alias of the mapped host buffer
  bool movesFetched_ = false, streamMoves_ = false, lastStepDtZero_ = false;
  int islandMode_ = 0;   // 0 auto, 1 CTA per island, 2 warp per island   // streamBodies_: every step packs + copies the body records (set by the first GetBodyEvents)
  friend struct StepLauncher;
};


// One handle = N independent worlds on one device (api/World.cs:9-10), stepped in lockstep by ONE launch: a CTA per world runs the
// whole step (k_batch_step in b2c_step.cu). The worlds are built through the ordinary creation API on the handles b2cuda_batch_world
// hands out; every world keeps its own id pools, trees, contact ids and colour graph in its own slice of the batch's device arena, so
// world k is bit-identical to the same world stepped alone. Step parameters (gravity, contact tuning, thresholds, switches) are the
// batch's: they are taken from the world definition the batch was created with.
class Batch {
 public:
  Batch(const B2CWorldDef& def, int device, int worldCount);
  ~Batch();
  int init();
  int worldCount() const { return (int)worlds_.size(); }
  World* world(int i) { return i >= 0 && i < (int)worlds_.size() ? worlds_[i] : nullptr; }
  int step(float dt, int subSteps);
  int synchronize();
  int stats(B2CBatchStats* out, double* deviceDst);
  int resetWorld(int index);
  int timerStart();
  int timerStop(float* ms);
  std::string error;
  // used by member worlds
  cudaStream_t stream() { return stream_; }
  void* arenaAlloc(size_t bytes);
  bool isChunkBase(const void* p) const;
  int* mappedInt(int** devPtr) { *devPtr = mappedIntDev_; return mappedInt_; }
  void markDirty(int index) { dirty_.push_back(index); }
 private:
  int fail(int code, const std::string& msg) { error = msg; return code; }
  int ensureUploaded();
  int device_; B2CWorldDef def_;
  cudaStream_t stream_ = nullptr;
  std::vector<World*> worlds_;
  WorldDev* dWorlds_ = nullptr;
  bool uploaded_ = false;
  std::vector<int> dirty_;
  std::vector<void*> chunks_; char* chunkPtr_ = nullptr; size_t chunkLeft_ = 0;   // device arena: chunks of >= 256 MB, zero-filled, bump allocation, freed with the batch
  struct Snap { World::HostSnap host; char* dev = nullptr; size_t bytes = 0; };   // state at the first step, for resetWorld
  std::vector<Snap> snaps_; std::vector<void*> snapChunks_;
  int* mappedInt_ = nullptr; int* mappedIntDev_ = nullptr;
  int* maxCons_ = nullptr; int* maxConsDev_ = nullptr;     // running maximum of an island's constraint count, kept by k_batch_step (mapped host int)
  int endIndex_ = 0, stepCount_ = 0;
  double* dStats_ = nullptr; double* hStats_ = nullptr;
  cudaEvent_t timer_[2] = {nullptr, nullptr};
  struct StepGraphB { StepParams sp; int endIndex, narrow, n; cudaGraphExec_t exec; };   // captured launches of one step
  std::vector<StepGraphB> graphs_;
};

}  // namespace b2c
