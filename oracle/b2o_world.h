// ORACLE — TEST INFRASTRUCTURE ONLY (see b2o_math.h header). Parity unpinned by the reference.
// Data model of the reference world restricted to the rigid-body step path:
// src/Body.cs, src/Shape.cs, src/Contact.cs, src/Joint.cs, src/joints/RevoluteJoint.cs,
// src/ConstraintGraph.cs, src/SolverSet.cs (static + awake sets only), src/BroadPhase.cs.
// Sensors: src/Sensor.cs (b2o_sensor.cpp). Continuous collision and bullets: b2o_continuous.inc. Not restated (asserted unused): islands/sleep, chains.
#pragma once
#include <vector>
#include <unordered_set>
#include <string>
#include "b2o_math.h"
#include "b2o_tree.h"
#include "b2o_geometry.h"
#include "b2o_manifold.h"

namespace b2o {

enum BodyType { kStaticBody = 0, kKinematicBody = 1, kDynamicBody = 2 };   // include/Types.cs:147
enum SetType { kStaticSet = 0, kDisabledSet = 1, kAwakeSet = 2 };          // src/World.cs:8-14
enum BodyFlags : uint32_t {                                                 // src/Body.cs:7-34
  kLockLinearX = 1, kLockLinearY = 2, kLockAngularZ = 4, kIsFast = 8, kIsBullet = 0x10, kIsSpeedCapped = 0x20,
  kHadTimeOfImpact = 0x40, kAllowFastRotation = 0x80, kEnlargeBounds = 0x100, kDynamicFlag = 0x200
};
enum ContactFlags : uint32_t { kContactTouching = 1, kContactHitEvent = 2, kContactEnableContactEvents = 4 };   // src/Contact.cs:6-16
enum ContactSimFlags : uint32_t {                                           // src/Contact.cs:66-85
  kSimTouching = 0x10000, kSimDisjoint = 0x20000, kSimStartedTouching = 0x40000, kSimStoppedTouching = 0x80000,
  kSimEnableHitEvent = 0x100000, kSimEnablePreSolveEvents = 0x200000
};

struct IdPool {                 // src/IDPool.cs
  std::vector<int> freeArray; int nextIndex = 0;
  int AllocId() { if (!freeArray.empty()) { int id = freeArray.back(); freeArray.pop_back(); return id; } return nextIndex++; }
  void FreeId(int id) { freeArray.push_back(id); }
  int GetIdCapacity() const { return nextIndex; }
  int GetIdCount() const { return nextIndex - (int)freeArray.size(); }
};

struct BitSet {                 // src/BitSet.cs (set-of-int semantics, ascending iteration)
  std::vector<uint64_t> bits; uint32_t blockCount = 0;
  void SetBitCountAndClear(int bitCount) { blockCount = (uint32_t)((bitCount + 63) >> 6); bits.assign(std::max<size_t>(bits.size(), blockCount), 0); }
  void SetBit(int i) { bits[i >> 6] |= 1ull << (i & 63); }
  void SetBitGrow(int i) { uint32_t b = (uint32_t)(i >> 6); if (b >= blockCount) { if (bits.size() < b + 1) bits.resize(b + 1 + ((b + 1) >> 1), 0); blockCount = b + 1; } bits[b] |= 1ull << (i & 63); }
  void ClearBit(int i) { uint32_t b = (uint32_t)(i >> 6); if (b >= blockCount) return; bits[b] &= ~(1ull << (i & 63)); }
  bool GetBit(int i) const { uint32_t b = (uint32_t)(i >> 6); if (b >= blockCount) return false; return (bits[b] >> (i & 63)) & 1; }
};

struct Body {                   // src/Body.cs:35-93
  uint64_t userData = 0;
  int setIndex = -1, localIndex = -1;
  int headContactKey = -1, contactCount = 0;
  int headShapeId = -1, shapeCount = 0;
  int headJointKey = -1, jointCount = 0;
  float mass = 0, inertia = 0, sleepThreshold = 0.05f, sleepTime = 0;
  int id = -1;
  uint32_t flags = 0;
  int type = kStaticBody;
  uint16_t generation = 0;
  bool enableSleep = true;
};
struct BodyState {              // src/Body.cs:119-136 (32 bytes)
  Vec2 linearVelocity{0, 0}; float angularVelocity = 0; uint32_t flags = 0;
  Vec2 deltaPosition{0, 0}; Rot deltaRotation;
};
struct BodySim {                // src/Body.cs:139-173
  Transform transform; Vec2 center{0, 0}; Rot rotation0; Vec2 center0{0, 0}, localCenter{0, 0}, force{0, 0};
  float torque = 0, invMass = 0, invInertia = 0, minExtent = kHuge, maxExtent = 0;
  float linearDamping = 0, angularDamping = 0, gravityScale = 1;
  int bodyId = -1; uint32_t flags = 0;
};
struct Filter { uint64_t categoryBits = ~0ull, maskBits = ~0ull; int groupIndex = 0; };   // include/Types.cs:226
struct SurfaceMaterial { float friction = 0.6f, restitution = 0, rollingResistance = 0, tangentSpeed = 0; uint64_t userMaterialId = 0; };
struct Shape {                  // src/Shape.cs:25-52
  int id = -1, bodyId = -1, prevShapeId = -1, nextShapeId = -1, sensorIndex = -1;
  Geom geom; SurfaceMaterial material; float density = 1;
  AABB aabb, fatAABB; Vec2 localCentroid{0, 0}; int proxyKey = -1;
  Filter filter; uint64_t userData = 0; uint16_t generation = 0;
  bool enableSensorEvents = false, enableContactEvents = false, enableCustomFiltering = false,
       enableHitEvents = false, enablePreSolveEvents = false, enlargedAABB = false;
};
struct ContactEdge { int bodyId = -1, prevKey = -1, nextKey = -1; };
struct Contact {                // src/Contact.cs:32-63
  int setIndex = -1, colorIndex = -1, localIndex = -1;
  ContactEdge edge0, edge1;
  int shapeIdA = -1, shapeIdB = -1, contactId = -1;
  uint32_t flags = 0; uint32_t generation = 0;
};
struct ContactSim {             // src/Contact.cs:90-123
  int contactId = -1, bodySimIndexA = -1, bodySimIndexB = -1, shapeIdA = -1, shapeIdB = -1;
  float invMassA = 0, invIA = 0, invMassB = 0, invIB = 0;
  Manifold manifold;
  float friction = 0, restitution = 0, rollingResistance = 0, tangentSpeed = 0;
  uint32_t simFlags = 0;
};
struct JointEdge { int bodyId = -1, prevKey = -1, nextKey = -1; };
struct Joint {                  // src/Joint.cs:18-50
  uint64_t userData = 0;
  int setIndex = -1, colorIndex = -1, localIndex = -1;
  JointEdge edge0, edge1;
  int jointId = -1; int type = 4; uint16_t generation = 0; bool collideConnected = false;
};
struct JointSim {               // src/Joint.cs:74-133 + src/joints/RevoluteJoint.cs:6-31
  int jointId = -1, bodyIdA = -1, bodyIdB = -1, type = 4;
  Transform localFrameA, localFrameB;
  float invMassA = 0, invMassB = 0, invIA = 0, invIB = 0;
  float constraintHertz = 60, constraintDampingRatio = 2;
  Softness constraintSoftness;
  float forceThreshold = FLT_MAX, torqueThreshold = FLT_MAX;
  // revolute
  Vec2 linearImpulse{0, 0}; float springImpulse = 0, motorImpulse = 0, lowerImpulse = 0, upperImpulse = 0;
  float hertz = 0, dampingRatio = 0, targetAngle = 0, maxMotorTorque = 0, motorSpeed = 0, lowerAngle = 0, upperAngle = 0;
  int indexA = -1, indexB = -1; Transform frameA, frameB; Vec2 deltaCenter{0, 0}; float axialMass = 0; Softness springSoftness;
  bool enableSpring = false, enableMotor = false, enableLimit = false;
  // Other joint types (src/joints/*.cs) reuse the slots above where the reference's field has the same role:
  //   distance  impulse -> springImpulse, maxMotorForce -> maxMotorTorque, distanceSoftness -> springSoftness, anchorA/B -> frameA/B.p
  //   weld      linearHertz/linearDampingRatio -> hertz/dampingRatio, angularImpulse -> springImpulse, linearSpring -> springSoftness
  //   motor     linearVelocityImpulse -> linearImpulse, angularVelocityImpulse -> motorImpulse, linearSpringImpulse -> linearImpulse2,
  //             angularSpringImpulse -> springImpulse, angularMass -> axialMass, linearSpring -> springSoftness
  //   prismatic impulse -> linearImpulse, targetTranslation -> targetAngle, maxMotorForce -> maxMotorTorque, lower/upperTranslation -> lower/upperAngle
  //   wheel     perpImpulse -> linearImpulse.x, lower/upperTranslation -> lower/upperAngle
  float length = 0, minLength = 0, maxLength = 0, lowerSpringForce = 0, upperSpringForce = 0;                       // distance
  float angularHertz = 0, angularDampingRatio = 0;                                                                   // weld, motor
  float maxSpringForce = 0, maxSpringTorque = 0, maxVelocityForce = 0, maxVelocityTorque = 0, angularVelocity = 0;   // motor
  Vec2 linearVelocity{0, 0}, linearImpulse2{0, 0};
  Softness angularSoftness; Mat22 linearMass{{0, 0}, {0, 0}}; float perpMass = 0, motorMass = 0;
};
enum { kDistanceJoint = 0, kFilterJoint = 1, kMotorJoint = 2, kPrismaticJoint = 3, kRevoluteJoint = 4, kWeldJoint = 5, kWheelJoint = 6 };  // include/Types.cs:423-426

// AVX lane of ContactConstraintSIMD (src/ContactSolver.cs:378-406)
struct ConstraintLane {
  int indexA = -1, indexB = -1;
  float invMassA = 0, invMassB = 0, invIA = 0, invIB = 0;
  Vec2 normal{0, 0};
  float friction = 0, tangentSpeed = 0, rollingResistance = 0, rollingMass = 0, rollingImpulse = 0;
  float biasRate = 0, massScale = 0, impulseScale = 0;
  Vec2 anchorA1{0, 0}, anchorB1{0, 0}; float normalMass1 = 0, tangentMass1 = 0, baseSeparation1 = 0, normalImpulse1 = 0, totalNormalImpulse1 = 0, tangentImpulse1 = 0;
  Vec2 anchorA2{0, 0}, anchorB2{0, 0}; float baseSeparation2 = 0, normalImpulse2 = 0, totalNormalImpulse2 = 0, tangentImpulse2 = 0, normalMass2 = 0, tangentMass2 = 0;
  float restitution = 0, relativeVelocity1 = 0, relativeVelocity2 = 0;
};
// scalar overflow constraint (src/ContactSolver.cs:9-37)
struct OverflowPoint { Vec2 anchorA, anchorB; float baseSeparation, relativeVelocity, normalImpulse, tangentImpulse, totalNormalImpulse, normalMass, tangentMass; };
struct OverflowConstraint {
  int indexA, indexB; OverflowPoint point0, point1; Vec2 normal; float invMassA, invMassB, invIA, invIB;
  float friction, restitution, tangentSpeed, rollingResistance, rollingMass, rollingImpulse; Softness softness; int pointCount;
};

struct GraphColor {             // src/ConstraintGraph.cs:10-24
  BitSet bodySet;
  std::vector<ContactSim> contactSims;
  std::vector<JointSim> jointSims;
  std::vector<OverflowConstraint> overflowConstraints;
  int simdBase = 0;             // first lane (multiple of 8) of this color in StepContext::lanes
};
struct SolverSet {              // src/SolverSet.cs:13-30 (subset)
  std::vector<BodySim> bodySims; std::vector<BodyState> bodyStates; std::vector<ContactSim> contactSims;
  std::vector<JointSim> jointSims;   // static set only: joints whose bodies are both non-dynamic (parked by SetType, SolverSet.TransferJoint)
};

struct Visitor { int shapeId; uint16_t generation; };                      // src/Sensor.cs:12-17
struct Sensor { std::vector<Visitor> hits, overlaps1, overlaps2; int shapeId = -1; };   // src/Sensor.cs:22-27
struct SensorEvent { int sensorShapeId; uint32_t sensorGeneration; int visitorShapeId; uint32_t visitorGeneration; };   // SensorBegin/EndTouchEvent as (ShapeID, ShapeID)
struct ContactTouchEvent { int shapeIdA, shapeIdB, contactId; uint32_t contactGeneration; };
struct ContactHitEvent { int shapeIdA, shapeIdB; Vec2 point, normal; float approachSpeed; };

struct WorldDef {               // include/Types.cs:75-144
  Vec2 gravity{0, -10}; float restitutionThreshold = 1, hitEventThreshold = 1, contactHertz = 30, contactDampingRatio = 10,
    contactSpeed = 3, maximumLinearSpeed = 400;
  bool enableSleep = true, enableContinuous = true, enableContactSoftening = false;
};
struct BodyDef {                // include/Types.cs:170-223
  int type = kStaticBody; Vec2 position{0, 0}; Rot rotation; Vec2 linearVelocity{0, 0}; float angularVelocity = 0,
    linearDamping = 0, angularDamping = 0, gravityScale = 1, sleepThreshold = 0.05f; uint64_t userData = 0;
  bool lockLinearX = false, lockLinearY = false, lockAngularZ = false, enableSleep = true, isAwake = true, isBullet = false,
    isEnabled = true, allowFastRotation = false;
};
struct ShapeDef {               // include/Types.cs:307-344
  uint64_t userData = 0; SurfaceMaterial material; float density = 1; Filter filter;
  bool enableCustomFiltering = false, isSensor = false, enableSensorEvents = false, enableContactEvents = false,
    enableHitEvents = false, enablePreSolveEvents = false, invokeContactCreation = true, updateBodyMass = true;
};
struct JointDef {               // include/Types.cs:431-456 (base) + the per-type defs :460-640, flattened
  int type = kRevoluteJoint;
  int bodyIdA = -1, bodyIdB = -1; Transform localFrameA, localFrameB; float forceThreshold = FLT_MAX, torqueThreshold = FLT_MAX,
    constraintHertz = 60, constraintDampingRatio = 2; bool collideConnected = false; uint64_t userData = 0;
  bool enableSpring = false, enableLimit = false, enableMotor = false;
  float hertz = 0, dampingRatio = 0, target = 0, lower = 0, upper = 0, maxMotor = 0, motorSpeed = 0;
  float length = 1, minLength = 0, maxLength = kHuge, lowerSpringForce = -FLT_MAX, upperSpringForce = FLT_MAX;
  float angularHertz = 0, angularDampingRatio = 0;
  Vec2 linearVelocity{0, 0}; float angularVelocity = 0, maxVelocityForce = 0, maxVelocityTorque = 0, maxSpringForce = 0, maxSpringTorque = 0;
};
struct RevoluteJointDef {       // include/Types.cs:431-456,569-597
  int bodyIdA = -1, bodyIdB = -1; Transform localFrameA, localFrameB; float forceThreshold = FLT_MAX, torqueThreshold = FLT_MAX,
    constraintHertz = 60, constraintDampingRatio = 2; bool collideConnected = false; uint64_t userData = 0;
  float targetAngle = 0; bool enableSpring = false; float hertz = 0, dampingRatio = 0; bool enableLimit = false;
  float lowerAngle = 0, upperAngle = 0; bool enableMotor = false; float maxMotorTorque = 0, motorSpeed = 0;
};

struct StepContext {            // src/Solver.cs:57-112
  float dt = 0, inv_dt = 0, h = 0, inv_h = 0; int subStepCount = 1;
  Softness contactSoftness, staticSoftness; float restitutionThreshold = 0, maxLinearVelocity = 0; bool enableWarmStarting = true;
  std::vector<ContactSim*> contacts;      // padded with nullptr per color to a multiple of 8
  std::vector<ConstraintLane> lanes;
  std::vector<JointSim*> joints;
  int activeColorCount = 0; int activeColorIndices[kGraphColorCount];
};

struct Counters { int toiCandidates = 0, unsupported = 0, referenceThrows = 0; };   // referenceThrows: the C# code raises an exception at that point of the step

class World {
 public:
  explicit World(const WorldDef& def);
  // --- API mirrors (api/Body.cs, api/Shape.cs, api/Joint.cs, api/World.cs) ---
  int CreateBody(const BodyDef& def);
  int CreateShape(int bodyId, const ShapeDef& def, const Geom& geom);
  int CreateRevoluteJoint(const RevoluteJointDef& def);
  int CreateJoint(const JointDef& def);
  void DestroyJoint(int jointId);          // JointAPI.DestroyJoint api/Joint.cs:9-17 -> DestroyJointInternal src/Joint.cs:297-375
  void DestroyBody(int bodyId);            // BodyAPI.DestroyBody api/Body.cs:106-160
  void DestroyShape(int shapeId, bool updateBodyMass);             // ShapeAPI.DestroyShape api/Shape.cs:48-56 -> DestroyShapeInternal src/Shape.cs:219-272
  void SetDensity(int shapeId, float density, bool updateBodyMass);  // ShapeAPI.SetDensity api/Shape.cs:95-103
  void ResetProxy(Shape& shape, bool destroyProxy);                  // src/Shape.cs:273-301
  void SetFilter(int shapeId, Filter filter);                        // ShapeAPI.SetFilter api/Shape.cs:157-166
  void SetGeometry(int shapeId, const Geom& geom);                   // ShapeAPI.SetCircle/SetCapsule/SetSegment/SetPolygon api/Shape.cs:289-326
  void Explode(Vec2 position, float radius, float falloff, float impulsePerLength, uint64_t maskBits);   // WorldAPI.Explode api/World.cs:703-776
  int SetType(int bodyId, int type);                                // BodyAPI.SetType api/Body.cs:185-268
  void ApplyMassFromShapes(int bodyId);                              // BodyAPI.ApplyMassFromShapes api/Body.cs:617-621
  void UnBufferMove(int proxyKey);         // src/BroadPhase.cs:60-71
  void SetTargetTransform(int bodyId, const Transform& target, float timeStep);
  void SetMassData(int bodyId, float mass, float rotationalInertia, Vec2 center);
  void SetMotionLocks(int bodyId, bool linearX, bool linearY, bool angularZ);
  void SetSurfaceMaterial(int shapeId, float friction, float restitution, float rollingResistance, float tangentSpeed);
  bool JointSetParam(int jointId, int param, float value);      // api/Joint.cs setters, see b2o_world.cpp
  bool JointEnable(int jointId, int feature, bool flag);
  bool JointSetLimits(int jointId, float lower, float upper);
  void JointSetLocalFrame(int jointId, int which, const Transform& frame);   // JointAPI.SetLocalFrameA/B api/Joint.cs:44,50
  void JointSetCollideConnected(int jointId, bool shouldCollide);            // JointAPI.SetCollideConnected api/Joint.cs:56-75
  void SetTransform(int bodyId, Vec2 position, Rot rotation);
  void SetLinearVelocity(int bodyId, Vec2 v);
  void SetAngularVelocity(int bodyId, float w);
  void ApplyForceToCenter(int bodyId, Vec2 force);
  void ApplyTorque(int bodyId, float torque);
  void ApplyLinearImpulseToCenter(int bodyId, Vec2 impulse);
  void ApplyForce(int bodyId, Vec2 force, Vec2 point);
  void ClearForces(int bodyId);
  void ApplyLinearImpulse(int bodyId, Vec2 impulse, Vec2 point);
  void ApplyAngularImpulse(int bodyId, float impulse);
  void SetLinearDamping(int bodyId, float v);
  void SetAngularDamping(int bodyId, float v);
  void SetGravityScale(int bodyId, float v);
  void SetBullet(int bodyId, bool flag) { BodySim& s = GetBodySim(bodies[bodyId]); if (flag) s.flags |= kIsBullet; else s.flags &= ~(uint32_t)kIsBullet; }   // api/Body.cs:835-842
  void Step(float timeStep, int subStepCount);

  // --- state ---
  DynamicTree trees[3];
  std::unordered_set<int> moveSet; std::vector<int> moveArray;
  std::unordered_set<uint64_t> pairSet;
  GraphColor colors[kGraphColorCount];
  IdPool bodyIdPool, shapeIdPool, jointIdPool, contactIdPool;
  std::vector<Body> bodies; std::vector<Shape> shapes; std::vector<Joint> joints; std::vector<Contact> contacts;
  SolverSet staticSet, awakeSet;
  bool lastStepDtZero = false;
  std::vector<ContactTouchEvent> contactBeginEvents, contactEndEvents[2]; int endEventArrayIndex = 0;
  std::vector<ContactHitEvent> contactHitEvents;
  std::vector<int> jointEvents;
  std::vector<Sensor> sensors; std::vector<SensorEvent> sensorBeginEvents, sensorEndEvents[2];   // src/World.cs:102-106
  Vec2 gravity; float hitEventThreshold, restitutionThreshold, maxLinearSpeed, contactSpeed, contactHertz, contactDampingRatio;
  float inv_h = 0, inv_dt = 0;
  bool enableSleep, enableWarmStarting = true, enableContactSoftening, enableContinuous, enableSpeculative = true;
  uint64_t stepIndex = 0;
  Counters counters;
  std::vector<int> bulletBodies;           // StepContext.bulletBodies src/Solver.cs:447-449: sim indices of this step's fast bullets
  double profileMs[8] = {0, 0, 0, 0, 0, 0, 0, 0};   // cumulative: step, pairs, collide, solve, solverStages
  std::vector<uint64_t> lastNewPairKeys;   // pair keys created by the last UpdateBroadPhasePairs, in creation order
  std::string lastError;

  // --- internals ---
  BodySim& GetBodySim(const Body& b) { return b.setIndex == kAwakeSet ? awakeSet.bodySims[b.localIndex] : staticSet.bodySims[b.localIndex]; }
  BodyState* GetBodyState(const Body& b) { return b.setIndex == kAwakeSet ? &awakeSet.bodyStates[b.localIndex] : nullptr; }
  ContactSim& GetContactSim(const Contact& c) { return c.colorIndex != -1 ? colors[c.colorIndex].contactSims[c.localIndex] : awakeSet.contactSims[c.localIndex]; }
  JointSim& GetJointSim(const Joint& j) { return j.colorIndex == -1 ? staticSet.jointSims[j.localIndex] : colors[j.colorIndex].jointSims[j.localIndex]; }
  void TransferJoint(bool toAwake, Joint& joint);   // src/SolverSet.cs:377-420 between the static set and the awake set (graph)
  void BufferMove(int key) { if (moveSet.insert(key).second) moveArray.push_back(key); }
  int BroadPhaseCreateProxy(int proxyType, const AABB& aabb, uint64_t categoryBits, int shapeIndex, bool forcePairCreation);
  void UpdateShapeAABBs(Shape& s, const Transform& xf, int proxyType);
  void UpdateBodyMassData(Body& body);
  bool ShouldBodiesCollide(const Body& a, const Body& b) const;
  void CreateContact(Shape& shapeA, Shape& shapeB);
  void DestroyContact(Contact& contact);
  void DestroyContactsBetweenBodies(Body& a, Body& b);
  bool UpdateContact(ContactSim& sim, const Shape& shapeA, const Transform& xfA, Vec2 centerOffsetA, const Shape& shapeB, const Transform& xfB, Vec2 centerOffsetB);
  int AssignJointColor(int bodyIdA, int bodyIdB, int typeA, int typeB);
  void AddContactToGraph(const ContactSim& sim, Contact& contact);
  void RemoveContactFromGraph(int bodyIdA, int bodyIdB, int colorIndex, int localIndex);
  void UpdateBroadPhasePairs();
  void Collide(StepContext& ctx);
  void Solve(StepContext& ctx);
  void SolverStages(StepContext& ctx);
  void FinalizeBodies(StepContext& ctx, BitSet& enlargedSimBitSet);
  void SolveContinuous(int bodySimIndex);
  void OverlapSensors();                   // src/Sensor.cs:130-218
  void DestroySensor(Shape& sensorShape);  // src/Sensor.cs:219-235 (and the same lines inside DestroyShapeInternal src/Shape.cs:245-268)
  void EnableSensorEvents(int shapeId, bool flag) { shapes[shapeId].enableSensorEvents = flag; }   // ShapeAPI.EnableSensorEvents api/Shape.cs:170-174
  int GetSensorOverlaps(int shapeId, Visitor* out, int capacity) const;   // ShapeAPI.GetSensorData api/Shape.cs:409-419
};

}  // namespace b2o
