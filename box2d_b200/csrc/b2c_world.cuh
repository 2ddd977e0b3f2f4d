// libb2cuda — device-resident world layout (SoA, float4 rows) and the host-side world that owns it.
// The layout is ours (B200-first), the semantics are the reference's: each array cites the C# type it replaces.
#pragma once
#include <cuda_runtime.h>
#include <vector>
#include <string>
#include "b2c_math.cuh"
#include "../../include/b2cuda.h"

namespace b2c {

// ---- contact slot flags (src/Contact.cs:6-16,66-85 merged into one word; slot index == contactId) ----
enum : uint32_t {
  CF_ALIVE = 1u, CF_TOUCHING = 2u, CF_ENABLE_CONTACT_EVENTS = 4u, CF_ENABLE_HIT = 8u, CF_ENABLE_PRESOLVE = 16u,
  CF_SHAPE_HIT = 32u,            // either shape has enableHitEvents (static property of the pair)
  CF_DISJOINT = 0x100u, CF_STARTED = 0x200u, CF_STOPPED = 0x400u, CF_STATE_MASK = 0x700u
};
enum : uint32_t { ERR_CONTACT_CAPACITY = 1u, ERR_STACK_OVERFLOW = 2u, ERR_OVERFLOW_COLOR = 4u, ERR_PAIR_CAPACITY = 8u,
                  ERR_TOI_CANDIDATE = 16u, ERR_HASH_FULL = 32u, ERR_TREE = 64u, ERR_SENSOR_CAPACITY = 128u,
                  ERR_REFERENCE_THROWS = 256u };   // the C# reference throws here (IndexOutOfRangeException in Distance.MakeSimplexFromCache, quirk Q19): the step is reported as failed
enum : uint16_t { NODE_ALLOCATED = 1, NODE_ENLARGED = 2, NODE_LEAF = 4 };   // src/DynamicTree.cs TreeNodeFlags

// ---- dynamic AABB tree replica (src/DynamicTree.cs:11-32 TreeNode, SoA) ----
struct TreeDev {
  float4* aabb;        // lo.x lo.y hi.x hi.y
  int2* child;         // child1, child2 (-1,-1 for leaves)
  int* parent;         // parent, or next-free for nodes on the free list
  uint32_t* hf;        // height | flags << 16
  uint64_t* cat;       // categoryBits
  int* user;           // leaf: shapeId
  uint8_t* moved;      // BroadPhase.moveSet membership, indexed by proxy id
  int* meta;           // [0] root [1] nodeCount [2] freeList [3] proxyCount
  int* leafIdx;        // rebuild scratch (leafIndices)
  float2* leafCtr;     // rebuild scratch (leafCenters)
  int* stack;          // rebuild scratch: DFS stack + RebuildItem stack (5 ints each)
  // parallel rebuild scratch (b2c_rebuild.cuh), all sized `capacity`
  int* segStart; int* recEnd; int* recNode; int* recSplit; float4* recBox; int* flagScan; int* flagArr; int* pairR; int* freeIds;
  int* subCount; int* visit; int* serialList;
  int* recEnd2; int* recNode2; float4* recBox2; int* recM; int* serEnd; int* serNode;   // second parity of the range records, flag count per range
  int* rb;             // [0] L [1] freed internal nodes [2] node cursor [3],[4] active ranges (level parity) [5] serial ranges [6] new root
  int capacity;
  int stackCapacity;
};

constexpr int kSensorMaxHits = 16;       // continuous-collision hits one sensor can collect in one step (Sensor.hits, src/Solver.cs:1276-1289)
constexpr int kSensorMaxOverlaps = 64;   // shapes one sensor can overlap at a time (src/Sensor.cs overlaps2); more sets ERR_SENSOR_CAPACITY and the step fails loudly

struct ShapeDev {      // src/Shape.cs:25-52 (hot subset) — one AoS row per shapeId, 16-byte aligned
  Geom geom;                         // 152 B
  float4 material;                   // friction, restitution, rollingResistance, tangentSpeed
  uint64_t categoryBits, maskBits;
  int groupIndex, bodyId, nextShapeId, proxyKey;
  uint32_t flags;                    // SF_*
  int sensorIndex;                   // Shape.sensorIndex (position in World.sensors) or -1
  uint32_t generation; int pad1;
};
enum : uint32_t { SF_CONTACT_EVENTS = 1u, SF_HIT_EVENTS = 2u, SF_PRESOLVE = 4u, SF_CUSTOM_FILTER = 8u, SF_ENLARGED = 16u, SF_SENSOR_EVENTS = 32u };

struct JointDev {      // src/Joint.cs:74-133 + src/joints/RevoluteJoint.cs:6-31
  int jointId, bodyIdA, bodyIdB, colorIndex;
  Xf localFrameA, localFrameB;
  float constraintHertz, constraintDampingRatio, forceThreshold, torqueThreshold;
  float hertz, dampingRatio, targetAngle, maxMotorTorque, motorSpeed, lowerAngle, upperAngle;
  int enableSpring, enableMotor, enableLimit;
  // prepared each step
  int indexA, indexB;
  float invMassA, invMassB, invIA, invIB;
  Xf frameA, frameB; V2 deltaCenter; float axialMass;
  Softness constraintSoftness, springSoftness;
  // accumulated impulses
  V2 linearImpulse; float springImpulse, motorImpulse, lowerImpulse, upperImpulse;
  // ---- joint types other than revolute (src/joints/*.cs). The slots above are shared where the reference's field plays the same role:
  //   distance  impulse -> springImpulse, maxMotorForce -> maxMotorTorque, distanceSoftness -> springSoftness, anchorA/B -> frameA/B.p
  //   weld      linearHertz/linearDampingRatio -> hertz/dampingRatio, angularImpulse -> springImpulse, linearSpring -> springSoftness
  //   motor     linearVelocityImpulse -> linearImpulse, angularVelocityImpulse -> motorImpulse, linearSpringImpulse -> linearImpulse2,
  //             angularSpringImpulse -> springImpulse, angularMass -> axialMass, linearSpring -> springSoftness
  //   prismatic impulse -> linearImpulse, targetTranslation -> targetAngle, maxMotorForce -> maxMotorTorque, lower/upperTranslation -> lower/upperAngle
  //   wheel     perpImpulse -> linearImpulse.x, lower/upperTranslation -> lower/upperAngle
  int type;            // JointType include/Types.cs:423-426
  float length, minLength, maxLength, lowerSpringForce, upperSpringForce;                              // distance
  float angularHertz, angularDampingRatio;                                                             // weld, motor
  float maxSpringForce, maxSpringTorque, maxVelocityForce, maxVelocityTorque, angularVelocity;         // motor
  V2 linearVelocity, linearImpulse2;
  Softness angularSoftness; float4 linearMass;   // Mat22 as (cx.x, cx.y, cy.x, cy.y)
  float perpMass, motorMass;
};
enum { kDistanceJoint = 0, kFilterJoint = 1, kMotorJoint = 2, kPrismaticJoint = 3, kRevoluteJoint = 4, kWeldJoint = 5, kWheelJoint = 6 };

// scalar overflow constraint (src/ContactSolver.cs:9-37)
struct OvPoint { V2 anchorA, anchorB; float baseSeparation, relativeVelocity, normalImpulse, tangentImpulse, totalNormalImpulse, normalMass, tangentMass; };
struct OvConstraint {
  int indexA, indexB; OvPoint p0, p1; V2 normal; float invMassA, invMassB, invIA, invIB;
  float friction, restitution, tangentSpeed, rollingResistance, rollingMass, rollingImpulse; Softness softness; int pointCount;
};

constexpr int kHitBuf = 8;     // accepted hits a pair query keeps from its count pass; a query with more walks the tree again in the fill pass (a full walk by one thread: ~100 us in a 20 000-leaf tree)

struct Scalars {       // device scalars, mirrored to pinned host memory at the end of each step
  int moveCount, newPairCount, contactNext, contactFreeCount;
  int beginCount, endCount[2], hitCount;
  int colorCount[kGraphColorCount], colorStart[kGraphColorCount];
  int constraintTotal, graphDirty, opCount, enlargedShapeCount;
  int errorFlags, toiCandidates, jointEventCount, stepCount;
  int touchingCount, liveContactCount, rebuildLeafCount[3], colorRounds;
  int awakeCount, anyRestitution, anyContactFlag, anyHitShapes;
  unsigned long long phaseStamp[8], phaseAcc[8];   // %globaltimer stamps inside the step kernels and their per-phase sums (ns): 0 pairs, 1 collide, 2 prepare, 3 grid solver, 4 tree rebuild, 5 island solver kernels, 6 finalize, 7 whole step
  int islandCount, freeBodyCount, bigBodyCount, smallConStart;   // island partition of the awake set (b2c_islands.cuh)
  unsigned int contactStamp;   // contacts created so far (see cStamp)
  int clusterSolve;    // this step's grid-path constraints are solved by k_solver_cluster (one thread-block cluster) instead of the front kernel
  int colorTotal[kGraphColorCount], bigCursor[kGraphColorCount];
  int bigQueryCount, bigHitTotal, bigGroupTotal, staticRanksValid;   // wide broadphase queries (b2c_bigquery.cuh)
  int rebuildLevels[3];
  int colLen[kGraphColorCount];         // GraphColor.contactSims.Count per colour (src/ConstraintGraph.cs:10-24); [kOverflowIndex] = contacts in the overflow colour
  int sensorParity, sensorBeginCount, sensorEndCount[2];   // src/Sensor.cs: which overlap buffer is overlaps2; sensor event counts (end events double-buffered like contact end events)
  unsigned long long dbg[16];           // order-independent state checksums per phase (B2CUDA_DEBUG_CHECKSUMS=1; b2cuda_world_debug_checksums)
};

static_assert(sizeof(Scalars) % 16 == 0, "Scalars is copied to the host mirror in 16-byte words");

struct StepParams {    // src/Solver.cs:57-112 StepContext (scalars only)
  float dt, inv_dt, h, inv_h;
  int subStepCount;
  Softness contactSoftness, staticSoftness;
  float restitutionThreshold, maxLinearSpeed, contactSpeed, contactHertz, contactDampingRatio, hitEventThreshold;
  V2 gravity;
  int enableWarmStarting, enableContactSoftening, enableContinuous, enableSpeculative, enableSleep;
  int endIndex;        // which contactEndEvents buffer this step appends to
  int clusterThreads;  // > 0: k_solver_cluster follows the front kernel with this many threads (one cluster); 0: grid path only
  int clusterDsm;      // the cluster kernel was launched with the shared memory of the distributed-shared-memory solver (b2c_cluster_dsm.cuh)
};

struct WorldDev;
struct WorldDev {
  // --- awake body set, index = awake sim index (src/Body.cs:119-173 BodyState + BodySim) ---
  float4* st0;         // v.x v.y w flags
  float4* st1;         // dp.x dp.y dq.c dq.s
  float4* simMass;     // invMass invInertia linearDamping angularDamping
  float4* simForce;    // force.x force.y torque gravityScale
  float4* simPose;     // center.x center.y q.c q.s
  float4* simLocal;    // localCenter.x localCenter.y minExtent maxExtent
  float4* simPose0;    // center0.x center0.y q0.c q0.s
  uint32_t* simFlags;
  int* simBodyId;
  int awakeCount;
  // --- per bodyId (all sets): what narrowphase / pair filtering / joints need ---
  float4* bodyXf;      // p.x p.y q.c q.s
  float4* bodyMass;    // invMass invInertia localCenter.x localCenter.y
  float2* bodyCenter;  // center (static bodies keep theirs; awake bodies refreshed in finalize)
  int* bodySim;        // awake sim index or -1
  int* bodyType;
  int* bodyHeadShape;
  int* bodyHeadJoint;  // joint key (jointId<<1|edge) or -1
  int* bodyJointCount;
  uint32_t* bodyColorMask;   // bit c <=> colors[c].bodySet.GetBit(bodyId) (src/ConstraintGraph.cs:10-24 transposed)
  int* bodyMinOp;      // coloring scratch
  float* bodySleepThreshold;
  uint32_t* bodyFlags;
  float4* bodyForce;         // force.x force.y torque of a sim parked in the static set by SetType (BodySim is copied whole, src/SolverSet.cs:360)
  uint32_t* bodySimFlags;   // BodySim.flags of a body that has no awake row (static set): kept so that SetType can bring them back (src/SolverSet.cs:360-361)
  int bodyCapacity;
  // --- shapes ---
  ShapeDev* shapes;
  float4* shapeAABB;   // aabb lo/hi
  float4* shapeFat;    // fatAABB lo/hi
  int shapeCapacity;
  // --- joints ---
  JointDev* joints;
  int4* jointEdges;    // per jointId: edge0.bodyId, edge0.nextKey, edge1.bodyId, edge1.nextKey
  int* jointCollide;   // collideConnected
  int* jointSorted;    // joint ids grouped by color
  int jointColorStart[kGraphColorCount + 1];
  uint8_t* jointEventFlag;
  int jointCapacity;
  // --- broadphase ---
  TreeDev trees[3];
  int* moveArray;      // proxy keys, BroadPhase.moveArray order
  int* moveHitCount;   // per move entry: accepted hits, then exclusive offsets
  int2* movePairs;     // (shapeIdA, shapeIdB) in contact-creation order
  int moveCapacity, pairCapacity;
  // wide queries (a moved proxy whose fat box overlaps a large part of a tree) are answered by a grid-wide sweep over the
  // tree's leaves in DFS order instead of a one-thread traversal (b2c_bigquery.cuh)
  int* moveBig;        // per move entry: slot in bigList or -1
  int2* moveHitBuf;    // [moveCapacity * kHitBuf] first accepted hits of each query, kept from the count pass
  int* bigList;        // [kMaxBigQueries] move indices
  int* bigSegStart;    // [kMaxBigQueries+1] start of each wide query's leaf segment in the sweep
  int* bigOff;         // [kMaxBigQueries+1] start of each wide query's hits in bigHits
  int* bigScan;        // [bigScratch] sweep flags -> exclusive ranks
  uint8_t* bigFlag;    // [bigScratch]
  int2* bigHits;       // [pairCapacity] accepted pairs in hit order
  int bigScratch;
  int bigQueryVisits;  // node-visit budget of a one-thread query (kBigQueryVisits; B2CUDA_BIG_QUERY_VISITS overrides it for tests)
  unsigned long long* pairTable;   // open-addressing set of shape pair keys (World.pairSet)
  int pairTableSize;   // power of two
  // --- contacts, slot == contactId (src/Contact.cs Contact + ContactSim) ---
  int2* cShape; int2* cBody; uint32_t* cFlags; int* cColor; uint32_t* cGen;
  uint32_t* cStamp;    // creation sequence number: a body's contact list (src/Contact.cs:196-216, pushed at the head) is its contacts by descending stamp
  float4* cMat; float4* cMass; int2* cSim;
  float4* mN;          // normal.x normal.y rollingImpulse pointCount(bits)
  float4* mA0; float4* mA1;   // anchorA.xy anchorB.xy
  float4* mB0; float4* mB1;   // separation normalImpulse tangentImpulse totalNormalImpulse
  float4* mC0; float4* mC1;   // point.xy normalVelocity (id | persisted<<16)(bits)
  int* contactFree;    // IDPool.freeArray
  int contactCapacity;
  // --- per-step ordered compaction scratch ---
  int4* flagScan;      // per contact slot: (begin, end, destroy, op) flags -> exclusive ranks
  int* opList;         // contact ids with a graph op (add/remove), ascending
  uint8_t* opDone;
  uint8_t* opCol;      // per op: colour | 0x40 (added to it) or | 0x80 (removed from it)
  int* colList;        // [kGraphColorCount][contactCapacity] contact ids of every colour in the reference's array order (GraphColor.contactSims:
                       // append on add, swap-with-last on remove) — decides the 8-lane restitution groups and the hit-event order
  int* cLocal;         // per contact slot: Contact.localIndex inside its colour's array, or -1
  uint8_t* cRestGroup; // per contact slot: its 8-lane group (localIndex >> 3 within the colour) holds a lane with restitution != 0
  OvConstraint* ovCons; int ovCapacity;
  // --- solver constraint planes, index = constraint (grouped by color) ---
  float4* sA; float4* sB; float4* sC; float4* sD; float4* sE; float4* sF; float4* sG; float4* sH; float4* sI; float4* sJ;
  int* sContact;       // contact slot of constraint t
  int* cConstraint;    // constraint index of contact slot (or -1)
  int constraintCapacity;
  int* blockHist;      // color sort scratch [numChunks][24]
  // --- islands: connected components of dynamic bodies under touching contacts + joints (src/Island.cs keeps the same
  //     partition for sleeping; here it only decides which solver kernel a constraint runs in) ---
  int* ufParent;       // [awake] union-find parent, root after flatten
  int* rootBodyCount;  // [awake] bodies per root, then body cursor
  int* rootConCount;   // [awake] constraints per root (+ 1<<20 per joint / kinematic neighbour: forces the grid path)
  int* rootColorCount; // [awake*24] constraints per root and color, then cursors
  int* rootInfo;       // [awake] at roots: -1 free body, -2 grid-path island, >= 0 compact island id
  int4* rootScan;      // [awake+1]
  int* islandBodyStart; int* islandConStart; int* islandColorOff;   // per island id; islandColorOff is [id*25 + color]
  int* islandBodies;   // [awake] sim indices grouped by island
  int* bodyLocal;      // [awake] position inside its island's body list
  int* freeBodies; int* bigBodies;   // [awake]
  uint8_t* kinematicUsed;            // [awake] non-dynamic awake body referenced by a constraint
  int2* sLocal;        // [constraint] island-local body indices (-1 = not in the awake set)
  int debugChecksums;
  int islandMaxBodies; // kIslandMaxBodies; B2CUDA_ISLAND_MAX_BODIES=0 sends every island down the grid path (tests)
  // --- events ---
  B2CContactTouchEvent* beginEvents; B2CContactTouchEvent* endEvents[2]; B2CContactHitEvent* hitEvents;
  int eventCapacity;
  B2CBodyMoveEvent* moveOut;   // when set, FinalizeBodies writes bodyMoveEvents[awake sim index]
  B2CBodyRead* bodyOut;   // when set, FinalizeBodies also writes the packed per-body record (bodyMoveEvents equivalent)
  // --- sensors (src/Sensor.cs:22-27): World.sensors order = sensor index; both overlap lists of every sensor, sorted by shape id ---
  int* sensorShape;          // [sensorCapacity] Sensor.shapeId
  int2* sensorOverlaps;      // [sensorCapacity][2][kSensorMaxOverlaps] (shapeId, generation); list Scalars.sensorParity is overlaps2, the other overlaps1
  int* sensorOverlapCount;   // [sensorCapacity][2]
  int2* sensorHits;          // [sensorCapacity][kSensorMaxHits] Sensor.hits: visitors whose swept core crossed the sensor this step (TOI), any order
  int* sensorHitCount;       // [sensorCapacity]
  int4* sensorScan;          // [sensorCapacity + 1] per sensor (begin events, end events, 0, 0) -> exclusive offsets
  B2CSensorEvent* sensorBeginEvents; B2CSensorEvent* sensorEndEvents[2];
  int sensorCount, sensorCapacity, sensorEventCapacity;
  // --- finalize / refit scratch ---
  int* bodyEnlargedCount;   // per awake sim: enlarged shapes, then exclusive offsets
  Scalars* sc;
  Scalars* hostSc;     // the host's pinned mirror of *sc, mapped: FinalizeBodies' last warp stores the step's scalars there (nullptr: batch member)
  // generic scan scratch
  int* scanBlockSums; int4* scanBlockSums4;
};


#ifdef __CUDACC__
// GraphColor.contactSims as the reference maintains it: Add (src/ConstraintGraph.cs:184-185) and RemoveSwap (:230-240, include/Box2D.cs:9-19)
__device__ __forceinline__ bool colAppend(const WorldDev& d, int c, int id) {
  Scalars& s = *d.sc;
  int n = s.colLen[c];
  if (n >= d.contactCapacity || (c == kOverflowIndex && n >= d.ovCapacity)) return false;
  d.colList[(size_t)c * d.contactCapacity + n] = id; d.cLocal[id] = n; s.colLen[c] = n + 1;
  return true;
}
__device__ __forceinline__ void colRemoveSwap(const WorldDev& d, int c, int id) {
  Scalars& s = *d.sc;
  int li = d.cLocal[id], last = s.colLen[c] - 1;
  if (li < 0 || li > last) return;
  int* arr = d.colList + (size_t)c * d.contactCapacity;
  if (li != last) { int moved = arr[last]; arr[li] = moved; d.cLocal[moved] = li; }
  d.cLocal[id] = -1; s.colLen[c] = last;
}
#endif

}  // namespace b2c
