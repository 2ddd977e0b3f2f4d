/* libb2cuda.so — C-ABI of the B200-native rigid-body step for the Box2D 3 C# port (MWstudios/Box2D).
 *
 * The reference has no FFI seam today (no DllImport/LibraryImport anywhere); the seam this library
 * introduces is the static Box2D.API surface. Each entry point below names the reference method whose
 * body a [LibraryImport("b2cuda")] stub replaces (see INTEGRATION.md and csharp/Box2D.Cuda/).
 *
 * Conventions: every function returns int32 status (0 = B2C_OK, <0 = error); no exceptions cross the
 * ABI; b2cuda_last_error() returns a library-owned UTF-8 string. Ids are the reference's raw ints
 * (BodyID.index1 - 1 etc., include/ID.cs:4-49); the library owns the id pools and reproduces the
 * reference's LIFO recycling (src/IDPool.cs). A world is not re-entrant (api/World.cs:36-37).
 * All state that the step touches lives in device memory; there is no CPU fallback: if no CUDA device
 * is usable b2cuda_world_create fails with B2C_E_NO_DEVICE.
 */
#ifndef B2CUDA_H
#define B2CUDA_H
#include <stdint.h>
#include <stddef.h>
#ifdef __cplusplus
extern "C" {
#endif

#define B2C_OK 0
#define B2C_E_INVALID (-1)       /* bad handle / id / argument */
#define B2C_E_UNSUPPORTED (-2)   /* feature of the reference that this build does not cover (never silent) */
#define B2C_E_NO_DEVICE (-3)
#define B2C_E_CUDA (-4)
#define B2C_E_CAPACITY (-5)      /* a device buffer overflowed (contacts, pairs, events) */
#define B2C_E_LOCKED (-6)

typedef struct B2CVec2 { float x, y; } B2CVec2;
typedef struct B2CRot { float c, s; } B2CRot;
typedef struct B2CTransform { B2CVec2 p; B2CRot q; } B2CTransform;

/* WorldDef POD — include/Types.cs:75-144 (callbacks/task system are not representable: default mixers only) */
typedef struct B2CWorldDef {
  B2CVec2 gravity;
  float restitutionThreshold, hitEventThreshold, contactHertz, contactDampingRatio, contactSpeed, maximumLinearSpeed;
  int32_t enableSleep, enableContinuous, enableContactSoftening;
  int32_t bodyCapacity, shapeCapacity, contactCapacity, jointCapacity;   /* 0 = default. body/shape: rows reserved for creation after the first step; contact: contact slots */
} B2CWorldDef;

/* BodyDef POD — include/Types.cs:170-223 */
typedef struct B2CBodyDef {
  int32_t type;                  /* 0 static, 1 kinematic, 2 dynamic (include/Types.cs:147) */
  B2CVec2 position; B2CRot rotation; B2CVec2 linearVelocity;
  float angularVelocity, linearDamping, angularDamping, gravityScale, sleepThreshold;
  uint64_t userData;
  int32_t lockLinearX, lockLinearY, lockAngularZ, enableSleep, isAwake, isBullet, isEnabled, allowFastRotation;
} B2CBodyDef;

/* ShapeDef POD — include/Types.cs:285-344 */
typedef struct B2CShapeDef {
  uint64_t userData;
  float friction, restitution, rollingResistance, tangentSpeed;
  uint64_t userMaterialId;
  float density;
  int32_t groupIndex;
  uint64_t categoryBits, maskBits;
  int32_t enableCustomFiltering, isSensor, enableSensorEvents, enableContactEvents, enableHitEvents,
          enablePreSolveEvents, invokeContactCreation, updateBodyMass;
} B2CShapeDef;

/* Geometry POD — include/Collision.cs:65,184,376,632. type: 0 circle, 1 capsule, 2 segment, 3 polygon.
 * circle: v[0]=center; capsule: v[0],v[1]=centers; segment: v[0],v[1]; polygon: count<=8 vertices+normals
 * (built host-side by the unchanged C# Geometry.MakeBox/MakePolygon, src/Geometry.cs). */
#define B2C_MAX_POLYGON_VERTICES 8
typedef struct B2CGeometry {
  int32_t type, count;
  float radius;
  B2CVec2 v[B2C_MAX_POLYGON_VERTICES], n[B2C_MAX_POLYGON_VERTICES];
  B2CVec2 centroid;
} B2CGeometry;

/* RevoluteJointDef POD — include/Types.cs:431-456,569-597 */
typedef struct B2CRevoluteJointDef {
  int32_t bodyIdA, bodyIdB;
  B2CTransform localFrameA, localFrameB;
  float forceThreshold, torqueThreshold, constraintHertz, constraintDampingRatio, drawScale;
  int32_t collideConnected;
  uint64_t userData;
  float targetAngle; int32_t enableSpring; float hertz, dampingRatio;
  int32_t enableLimit; float lowerAngle, upperAngle;
  int32_t enableMotor; float maxMotorTorque, motorSpeed;
} B2CRevoluteJointDef;

/* Generic joint definition: JointDef base (include/Types.cs:431-456) + the per-type definitions flattened — DistanceJointDef :460-492,
 * MotorJointDef :496-523, FilterJointDef :524-529, PrismaticJointDef :530-545, RevoluteJointDef :546-561, WeldJointDef :562-570,
 * WheelJointDef :571-586. `type` is the reference's JointType (include/Types.cs:423-426). Fields a type does not have are ignored.
 *   hertz/dampingRatio   distance, prismatic, revolute, wheel spring; weld/motor LINEAR hertz/dampingRatio
 *   target               revolute targetAngle / prismatic targetTranslation
 *   lower/upper          revolute angles / prismatic, wheel translations
 *   maxMotor             revolute, wheel maxMotorTorque / distance, prismatic maxMotorForce */
#define B2C_JOINT_DISTANCE 0
#define B2C_JOINT_FILTER 1
#define B2C_JOINT_MOTOR 2
#define B2C_JOINT_PRISMATIC 3
#define B2C_JOINT_REVOLUTE 4
#define B2C_JOINT_WELD 5
#define B2C_JOINT_WHEEL 6
typedef struct B2CJointDef {
  int32_t type, bodyIdA, bodyIdB;
  B2CTransform localFrameA, localFrameB;
  float forceThreshold, torqueThreshold, constraintHertz, constraintDampingRatio, drawScale;
  int32_t collideConnected;
  uint64_t userData;
  int32_t enableSpring, enableLimit, enableMotor;
  float hertz, dampingRatio, target, lower, upper, maxMotor, motorSpeed;
  float length, minLength, maxLength, lowerSpringForce, upperSpringForce;              /* distance */
  float angularHertz, angularDampingRatio;                                             /* weld, motor */
  B2CVec2 linearVelocity; float angularVelocity, maxVelocityForce, maxVelocityTorque, maxSpringForce, maxSpringTorque;   /* motor */
} B2CJointDef;

/* Read-back records */
typedef struct B2CBodyRead {      /* BodyAPI.GetTransform/GetLinearVelocity/GetAngularVelocity... api/Body.cs:282-357 */
  B2CTransform transform; B2CVec2 center; B2CVec2 linearVelocity; float angularVelocity; uint32_t flags;
} B2CBodyRead;
typedef struct B2CBodyMoveEvent {       /* BodyMoveEvent include/Types.cs:761-767 without the managed userData (the shim maps bodyId -> userData) */
  B2CTransform transform; int32_t bodyId; uint32_t fellAsleep;
} B2CBodyMoveEvent;
typedef struct B2CManifoldPointRead {   /* include/Collision.cs:916-945 */
  B2CVec2 point, anchorA, anchorB;
  float separation, normalImpulse, tangentImpulse, totalNormalImpulse, normalVelocity;
  uint32_t id, persisted;
} B2CManifoldPointRead;
typedef struct B2CContactRead {         /* ContactAPI.GetData api/Contact.cs + graph colour and Contact.localIndex (src/Contact.cs:41-47: the position in
                                           the colour's contactSims array for a touching contact, -1 otherwise) */
  int32_t contactId, shapeIdA, shapeIdB, colorIndex, localIndex;
  uint32_t flags;                        /* bit0 touching */
  int32_t pointCount;
  B2CVec2 normal; float rollingImpulse, friction, restitution, rollingResistance, tangentSpeed;
  B2CManifoldPointRead points[2];
} B2CContactRead;
typedef struct B2CContactTouchEvent { int32_t shapeIdA, shapeIdB, contactId; uint32_t contactGeneration; } B2CContactTouchEvent; /* include/Types.cs:697-730 */
typedef struct B2CContactHitEvent { int32_t shapeIdA, shapeIdB; B2CVec2 point, normal; float approachSpeed; } B2CContactHitEvent;
/* Accumulated joint impulses. Slots by type (reference field -> slot): revolute as named; distance impulse -> springImpulse; weld
 * angularImpulse -> springImpulse; prismatic impulse -> linearImpulse; wheel perpImpulse -> linearImpulse.x; motor linearVelocityImpulse ->
 * linearImpulse, angularVelocityImpulse -> motorImpulse, linearSpringImpulse -> linearImpulse2, angularSpringImpulse -> springImpulse. */
typedef struct B2CJointRead { int32_t jointId, colorIndex; B2CVec2 linearImpulse; float springImpulse, motorImpulse, lowerImpulse, upperImpulse;
                              int32_t type; B2CVec2 linearImpulse2; } B2CJointRead;
typedef struct B2CShapeRead { B2CVec2 aabbLo, aabbHi, fatLo, fatHi; int32_t proxyKey, enlarged; } B2CShapeRead;
typedef struct B2CCounters {            /* WorldAPI.GetCounters api/World.cs:817-834 */
  int32_t bodyCount, awakeBodyCount, shapeCount, contactCount, touchingContactCount, jointCount, contactIdCapacity,
          activeColorCount, moveCount, newPairCount, toiCandidates, unsupportedCount, stepCount, errorFlags;
  int32_t colorCounts[24];
} B2CCounters;
typedef struct B2CProfile {             /* WorldAPI.GetProfile api/World.cs:814; fields of include/Types.cs:378-403 (ms, CUDA events) */
  float step, pairs, collide, solve, prepareStages, solveConstraints, prepareConstraints, integrateVelocities, warmStart,
        solveImpulses, integratePositions, relaxImpulses, applyRestitution, storeImpulses, transforms, refit, readback,
        treeRebuild;        /* BroadPhase.RebuildTrees: inside `collide` as in the reference (src/World.cs:313), also reported alone */
} B2CProfile;

typedef struct B2CWorld B2CWorld;

const char* b2cuda_last_error(void);
int32_t b2cuda_default_world_def(B2CWorldDef* def);       /* include/Types.cs:75-144 defaults */
int32_t b2cuda_default_body_def(B2CBodyDef* def);         /* :170-223 */
int32_t b2cuda_default_shape_def(B2CShapeDef* def);       /* :285-344 */
int32_t b2cuda_default_revolute_joint_def(B2CRevoluteJointDef* def);   /* :431-456,569-597 */

/* WorldAPI.CreateWorld / DestroyWorld — api/World.cs:12-18, src/World.cs:183-260 */
int32_t b2cuda_world_create(const B2CWorldDef* def, int32_t device, B2CWorld** out);
int32_t b2cuda_world_destroy(B2CWorld* world);
/* BodyAPI.CreateBody — api/Body.cs:16-105 */
int32_t b2cuda_create_body(B2CWorld* world, const B2CBodyDef* def, int32_t* bodyId);
/* ShapeAPI.CreateCircleShape/CreateCapsuleShape/CreateSegmentShape/CreatePolygonShape — api/Shape.cs:8-33 */
int32_t b2cuda_create_shape(B2CWorld* world, int32_t bodyId, const B2CShapeDef* def, const B2CGeometry* geometry, int32_t* shapeId);
/* JointAPI.CreateRevoluteJoint — api/Joint.cs:555, src/joints/RevoluteJoint.cs:32-56 */
int32_t b2cuda_create_revolute_joint(B2CWorld* world, const B2CRevoluteJointDef* def, int32_t* jointId);
/* JointAPI.CreateDistanceJoint :182 / CreateMotorJoint :324 / CreateFilterJoint :408 / CreatePrismaticJoint :412 / CreateRevoluteJoint :555 /
 * CreateWeldJoint :672 / CreateWheelJoint :720 — api/Joint.cs; World.CreateJoint src/Joint.cs:181-296 */
int32_t b2cuda_default_joint_def(int32_t type, B2CJointDef* def);
int32_t b2cuda_create_joint(B2CWorld* world, const B2CJointDef* def, int32_t* jointId);
/* JointAPI.DestroyJoint api/Joint.cs:9-17 (DestroyJointInternal src/Joint.cs:297-375) and BodyAPI.DestroyBody api/Body.cs:106-160: the body's joints,
 * then its contacts in contact-list order (end-touch events go to the buffer the NEXT step reports, ids return to the LIFO pool), its shapes'
 * proxies (UnBufferMove + tree RemoveLeaf), then the body leaves the awake set by swap-remove. Allowed between steps once the world has been
 * stepped (destroy_joint also before). b2cuda_create_body / b2cuda_create_shape / b2cuda_create_joint also work between steps after the first
 * one (the joint's colour is assigned against the device's colour bitsets, contacts between its bodies are destroyed unless collideConnected): rows
 * are reserved at the first step (a quarter more than the world holds then, at least 256, or B2CWorldDef.bodyCapacity / shapeCapacity /
 * jointCapacity) and, when a class of rows is full, re-allocated at x1.5 on the world's stream before the call touches anything — as the
 * reference's lists grow. Pointers handed out by the event getters are then stale, as after a step. Contact-sized buffers do not grow
 * (B2CWorldDef.contactCapacity, default 12 per shape present at the first step): their overflow is reported as B2C_E_CAPACITY. */
int32_t b2cuda_destroy_joint(B2CWorld* world, int32_t jointId, int32_t wakeAttached);
int32_t b2cuda_destroy_body(B2CWorld* world, int32_t bodyId);
/* ShapeAPI.DestroyShape api/Shape.cs:48-56 (DestroyShapeInternal src/Shape.cs:219-272: the shape leaves its body's list, its proxy leaves the tree,
 * the body's contacts that involve it are destroyed in contact-list order), ShapeAPI.SetDensity api/Shape.cs:95-103 and
 * BodyAPI.ApplyMassFromShapes api/Body.cs:617-621. updateBodyMass runs World.UpdateBodyMassData src/Body.cs:255-353: mass, inertia, local centre
 * and extents are summed over the remaining shapes, the centre of mass moves and the linear velocity takes w x (newCentre - oldCentre).
 * destroy_shape needs a stepped world (like destroy_body); the other two also work while the world is being assembled. */
int32_t b2cuda_destroy_shape(B2CWorld* world, int32_t shapeId, int32_t updateBodyMass);
int32_t b2cuda_shape_set_density(B2CWorld* world, int32_t shapeId, float density, int32_t updateBodyMass);
int32_t b2cuda_body_apply_mass_from_shapes(B2CWorld* world, int32_t bodyId);
/* ShapeAPI.EnableContactEvents api/Shape.cs:182-186 / EnableHitEvents :205-209 and the per-body loops BodyAPI.EnableContactEvents
 * api/Body.cs:848-858 / EnableHitEvents :862-872. The flag is sampled when a contact is created (contact events) or starts touching
 * (hit events); existing contacts keep theirs, as in the reference. */
#define B2C_SHAPE_CONTACT_EVENTS 0
#define B2C_SHAPE_HIT_EVENTS 1
int32_t b2cuda_shape_enable_events(B2CWorld* world, int32_t shapeId, int32_t which, int32_t flag);
int32_t b2cuda_body_enable_events(B2CWorld* world, int32_t bodyId, int32_t which, int32_t flag);
/* ShapeAPI.SetFilter api/Shape.cs:157-166 and ShapeAPI.SetCircle / SetCapsule / SetSegment / SetPolygon :289-326, both through World.ResetProxy
 * src/Shape.cs:273-301: the shape's contacts are destroyed (end events in the next step), its AABBs are recomputed from the body's current
 * transform and the proxy is re-created (new category bits or new geometry) or moved; contacts come back in the next step. The geometry
 * setters leave the mass alone, as in the reference (call b2cuda_body_apply_mass_from_shapes). Between steps of a stepped world only. */
int32_t b2cuda_shape_set_filter(B2CWorld* world, int32_t shapeId, uint64_t categoryBits, uint64_t maskBits, int32_t groupIndex);
int32_t b2cuda_shape_set_geometry(B2CWorld* world, int32_t shapeId, const B2CGeometry* geometry);
/* BodyAPI.SetType api/Body.cs:185-268 (type: 0 static, 1 kinematic, 2 dynamic) between steps of a stepped world: the body's contacts are
 * destroyed, it moves between the static and the awake set (a body that becomes movable starts at rest, SolverSet.TransferBody
 * src/SolverSet.cs:354-376), its proxies move to the tree of the new type and the mass data is recomputed. Its joints pass through the
 * static set (SolverSet.TransferJoint :377-420): each comes back with a fresh colour when one of its bodies is dynamic afterwards; a joint
 * between two non-dynamic bodies stays parked — not simulated, absent from b2cuda_world_read_joints — until one of them is dynamic again. */
int32_t b2cuda_body_set_type(B2CWorld* world, int32_t bodyId, int32_t type);
/* WorldAPI.Explode api/World.cs:703-776 (ExplosionDef include/Types.cs: maskBits, position, radius, falloff, impulsePerLength): every shape in
 * the dynamic tree within radius + falloff of the position (GJK distance, Distance.ShapeDistance src/Distance.cs:219-318) pushes its body
 * away with impulsePerLength x projected perimeter x falloff scale, in the tree's query order. */
int32_t b2cuda_world_explode(B2CWorld* world, B2CVec2 position, float radius, float falloff, float impulsePerLength, uint64_t maskBits);
/* Joint mutators — api/Joint.cs. One generic entry per shape of setter; `param` selects the field (B2C_JP_*). Wrong joint type for a
 * field returns B2C_E_INVALID (the reference asserts the type, api/Joint.cs:169-179). After the first step each call is a one-thread
 * patch of the device joint record, ordered on the world's stream (no read-back, no synchronisation).
 *   set_param   plain stores: JointAPI.SetConstraintTuning :130, SetForceThreshold :148, SetTorqueThreshold :159, <Type>Joint_SetSpringHertz /
 *               SetSpringDampingRatio / SetMotorSpeed / SetMaxMotorForce|Torque / SetTargetAngle|Translation, DistanceJoint_SetLength :187
 *               (clamped to [linearSlop, huge], clears lower/upper impulse), DistanceJoint_SetSpringForceRange :204, MotorJoint_Set* :327-360
 *               (Max(0, x) where the reference clamps), WeldJoint_Set* :675-700
 *   enable      <Type>Joint_EnableSpring / EnableLimit / EnableMotor: on a change the matching accumulated impulses are cleared
 *               (WheelJoint_EnableLimit :755 stores false whenever the flag changes - kept as written)
 *   set_limits  RevoluteJoint_SetLimits :626, PrismaticJoint_SetLimits :471, WheelJoint_SetLimits :769 (min/max, impulses cleared when
 *               changed), DistanceJoint_SetLengthRange :236 (clamped, all three impulses cleared) */
#define B2C_JP_CONSTRAINT_HERTZ 0
#define B2C_JP_CONSTRAINT_DAMPING_RATIO 1
#define B2C_JP_FORCE_THRESHOLD 2
#define B2C_JP_TORQUE_THRESHOLD 3
#define B2C_JP_SPRING_HERTZ 4            /* weld/motor: linear hertz */
#define B2C_JP_SPRING_DAMPING_RATIO 5    /* weld/motor: linear damping ratio */
#define B2C_JP_TARGET 6                  /* revolute target angle / prismatic target translation */
#define B2C_JP_MOTOR_SPEED 7
#define B2C_JP_MAX_MOTOR 8               /* max motor torque (revolute, wheel) / force (distance, prismatic) */
#define B2C_JP_LENGTH 9
#define B2C_JP_LOWER_SPRING_FORCE 10
#define B2C_JP_UPPER_SPRING_FORCE 11
#define B2C_JP_ANGULAR_HERTZ 12
#define B2C_JP_ANGULAR_DAMPING_RATIO 13
#define B2C_JP_LINEAR_VELOCITY_X 14
#define B2C_JP_LINEAR_VELOCITY_Y 15
#define B2C_JP_ANGULAR_VELOCITY 16
#define B2C_JP_MAX_VELOCITY_FORCE 17
#define B2C_JP_MAX_VELOCITY_TORQUE 18
#define B2C_JP_MAX_SPRING_FORCE 19
#define B2C_JP_MAX_SPRING_TORQUE 20
#define B2C_JP_COUNT 21
#define B2C_JOINT_SPRING 0
#define B2C_JOINT_LIMIT 1
#define B2C_JOINT_MOTOR 2
int32_t b2cuda_joint_set_param(B2CWorld* world, int32_t jointId, int32_t param, float value);
int32_t b2cuda_joint_enable(B2CWorld* world, int32_t jointId, int32_t feature, int32_t flag);
int32_t b2cuda_joint_set_limits(B2CWorld* world, int32_t jointId, float lower, float upper);
/* JointAPI.SetLocalFrameA / SetLocalFrameB api/Joint.cs:44,50 (which = 0 for A, 1 for B): a plain store into the joint sim, picked up by the next
 * PrepareJoints. JointAPI.SetCollideConnected api/Joint.cs:56-75: switching it on buffers a move for the shapes of the body with fewer shapes
 * (the next step's pair update creates the contacts), switching it off destroys the contacts between the two bodies in contact-list order. */
int32_t b2cuda_joint_set_local_frame(B2CWorld* world, int32_t jointId, int32_t which, B2CTransform localFrame);
int32_t b2cuda_joint_set_collide_connected(B2CWorld* world, int32_t jointId, int32_t shouldCollide);
/* BodyAPI sleep / enable / bullet surface (api/Body.cs): SetSleepThreshold :712, EnableSleep :700, SetAwake :668, WakeTouching :682,
 * Disable :721, Enable :754, SetBullet :835. See DESIGN.md for which of them act and which are refused with B2C_E_UNSUPPORTED. */
int32_t b2cuda_body_set_sleep_threshold(B2CWorld* world, int32_t bodyId, float sleepThreshold);
int32_t b2cuda_body_enable_sleep(B2CWorld* world, int32_t bodyId, int32_t enableSleep);
int32_t b2cuda_body_set_awake(B2CWorld* world, int32_t bodyId, int32_t awake);
int32_t b2cuda_body_wake_touching(B2CWorld* world, int32_t bodyId);
int32_t b2cuda_body_disable(B2CWorld* world, int32_t bodyId);
int32_t b2cuda_body_enable(B2CWorld* world, int32_t bodyId);
int32_t b2cuda_body_set_bullet(B2CWorld* world, int32_t bodyId, int32_t flag);
/* BodyAPI.SetTransform :293 / SetLinearVelocity :359 / SetAngularVelocity :369 / ApplyForceToCenter :457 /
 * ApplyTorque :469 / ApplyLinearImpulseToCenter :531 — api/Body.cs */
int32_t b2cuda_body_set_transform(B2CWorld* world, int32_t bodyId, B2CVec2 position, B2CRot rotation);
int32_t b2cuda_body_set_linear_velocity(B2CWorld* world, int32_t bodyId, B2CVec2 v);
int32_t b2cuda_body_set_angular_velocity(B2CWorld* world, int32_t bodyId, float w);
int32_t b2cuda_body_apply_force_to_center(B2CWorld* world, int32_t bodyId, B2CVec2 force, int32_t wake);
int32_t b2cuda_body_apply_torque(B2CWorld* world, int32_t bodyId, float torque, int32_t wake);
int32_t b2cuda_body_apply_linear_impulse_to_center(B2CWorld* world, int32_t bodyId, B2CVec2 impulse, int32_t wake);
/* BodyAPI.ApplyForce :439 / ClearForces :490 / ApplyLinearImpulse :507 / ApplyAngularImpulse :554 / SetLinearDamping :624 /
 * SetAngularDamping :635 / SetGravityScale :647 — api/Body.cs. Stream-ordered device patches: no host synchronisation. */
int32_t b2cuda_body_apply_force(B2CWorld* world, int32_t bodyId, B2CVec2 force, B2CVec2 point, int32_t wake);
int32_t b2cuda_body_clear_forces(B2CWorld* world, int32_t bodyId);
int32_t b2cuda_body_apply_linear_impulse(B2CWorld* world, int32_t bodyId, B2CVec2 impulse, B2CVec2 point, int32_t wake);
int32_t b2cuda_body_apply_angular_impulse(B2CWorld* world, int32_t bodyId, float impulse, int32_t wake);
int32_t b2cuda_body_set_linear_damping(B2CWorld* world, int32_t bodyId, float linearDamping);
int32_t b2cuda_body_set_angular_damping(B2CWorld* world, int32_t bodyId, float angularDamping);
int32_t b2cuda_body_set_gravity_scale(B2CWorld* world, int32_t bodyId, float gravityScale);
/* BodyAPI.SetTargetTransform :378 (velocities that reach `target` in `timeStep`; the angle uses Atan2) / SetMassData :585 /
 * SetMotionLocks :796 — api/Body.cs; ShapeAPI.SetFriction :110 / SetRestitution :122 / SetSurfaceMaterial :144 — api/Shape.cs (the mixed
 * contact friction/restitution is recomputed by every UpdateContact, src/Contact.cs:330-340, so the change shows on the next step). */
int32_t b2cuda_body_set_target_transform(B2CWorld* world, int32_t bodyId, B2CTransform target, float timeStep);
int32_t b2cuda_body_set_mass_data(B2CWorld* world, int32_t bodyId, float mass, float rotationalInertia, B2CVec2 center);
int32_t b2cuda_body_set_motion_locks(B2CWorld* world, int32_t bodyId, int32_t linearX, int32_t linearY, int32_t angularZ);
int32_t b2cuda_shape_set_surface_material(B2CWorld* world, int32_t shapeId, float friction, float restitution, float rollingResistance, float tangentSpeed);
/* BodyAPI.GetTransform :288 / GetLinearVelocity :345 / GetAngularVelocity :352 / GetWorldCenterOfMass :580 for one body
 * (synchronises with the world's stream; use b2cuda_world_read_bodies or the move events for bulk reads). */
int32_t b2cuda_body_get(B2CWorld* world, int32_t bodyId, B2CBodyRead* out);
/* WorldAPI.SetRestitutionThreshold :666 / SetHitEventThreshold :678 / SetMaximumLinearSpeed :789 — api/World.cs */
int32_t b2cuda_world_set_restitution_threshold(B2CWorld* world, float value);
int32_t b2cuda_world_set_hit_event_threshold(B2CWorld* world, float value);
int32_t b2cuda_world_set_maximum_linear_speed(B2CWorld* world, float maximumLinearSpeed);
/* WorldAPI.SetGravity :698 / EnableWarmStarting :801 / EnableSpeculative :941 / EnableContinuous :654 / SetContactTuning :780 */
int32_t b2cuda_world_set_gravity(B2CWorld* world, B2CVec2 gravity);
int32_t b2cuda_world_enable_warm_starting(B2CWorld* world, int32_t flag);
int32_t b2cuda_world_enable_speculative(B2CWorld* world, int32_t flag);
int32_t b2cuda_world_enable_continuous(B2CWorld* world, int32_t flag);
/* WorldAPI.EnableSleeping api/World.cs:631-645: flag 0 is accepted (nothing sleeps in this build, so there is nothing to wake);
 * flag 1 returns B2C_E_UNSUPPORTED and leaves sleeping off (islands/sleep, src/Island.cs, are not built). */
int32_t b2cuda_world_enable_sleeping(B2CWorld* world, int32_t flag);
int32_t b2cuda_world_set_contact_tuning(B2CWorld* world, float hertz, float dampingRatio, float pushSpeed);

/* WorldAPI.Step — api/World.cs:31-113. Asynchronous on the world's stream; read-backs synchronise. */
int32_t b2cuda_world_step(B2CWorld* world, float timeStep, int32_t subStepCount);
int32_t b2cuda_world_synchronize(B2CWorld* world);

/* WorldAPI.GetContactEvents — api/World.cs:302-312. Pointers into library-owned pinned host memory, valid until
 * the next step ("transient, do not store"). End events are the reference's double-buffered previous set. */
int32_t b2cuda_world_get_contact_events(B2CWorld* world, const B2CContactTouchEvent** begin, int32_t* beginCount,
                                        const B2CContactTouchEvent** end, int32_t* endCount,
                                        const B2CContactHitEvent** hit, int32_t* hitCount);
/* WorldAPI.GetJointEvents — api/World.cs:314-319: joint ids whose reaction exceeded a threshold */
int32_t b2cuda_world_get_joint_events(B2CWorld* world, const int32_t** jointIds, int32_t* count);
/* WorldAPI.GetBodyEvents / BodyAPI getters in bulk: record i = bodyId i (free slots zeroed) */
int32_t b2cuda_world_read_bodies(B2CWorld* world, B2CBodyRead* dst, int32_t capacity, int32_t* count);
/* WorldAPI.GetBodyEvents — api/World.cs:283-290 (bodyMoveEvents, src/Solver.cs:432-438): one record per body id, packed on the
 * device and copied once into library-owned pinned host memory; the pointer is valid until the next step ("transient"). */
int32_t b2cuda_world_get_body_events(B2CWorld* world, const B2CBodyRead** records, int32_t* count);
/* WorldAPI.GetBodyEvents exactly as the reference fills it (src/Solver.cs:432-438): one B2CBodyMoveEvent per AWAKE body in
 * awake-set order (24 B each), written by FinalizeBodies; after the first call every step copies the array to pinned host
 * memory behind the step, so the call only synchronises. Pointer valid until the next step. */
int32_t b2cuda_world_get_body_move_events(B2CWorld* world, const B2CBodyMoveEvent** events, int32_t* count);
/* No counterpart in the reference: added for callers that cannot call cudaMalloc (code under
 * CUDA graph capture, tight serving loops, harnesses that forbid it). Same records as b2cuda_world_get_body_move_events — one
 * B2CBodyMoveEvent per AWAKE body in awake-set order — but packed into a file-scope device array of B2C_NOALLOC_MAX_MOVE_EVENTS records
 * (1.5 MiB, held by every GPU that loads the library until exit) and copied into `dst`, which the CALLER owns: host memory (pageable or
 * pinned) or device memory of at least `capacity` records. Calls no cudaMalloc / cudaFree / cudaHostAlloc or any other allocating routine.
 * More awake bodies than the maximum: B2C_E_CAPACITY; `capacity` too small: B2C_E_INVALID; *count = records written (0 after a dt = 0 step).
 * Synchronises the world's stream before returning. Every call on a GPU shares the one scratch array, which keeps the last call's
 * contents (nothing needs clearing: records [0, count) are overwritten before they are copied): calls that overlap on the GPU race. */
#define B2C_NOALLOC_MAX_MOVE_EVENTS 65536
int32_t b2cuda_world_copy_body_move_events_noalloc(B2CWorld* world, B2CBodyMoveEvent* dst, int32_t capacity, int32_t* count);
/* ContactAPI.GetData for every live contact in ascending contactId */
int32_t b2cuda_world_read_contacts(B2CWorld* world, B2CContactRead* dst, int32_t capacity, int32_t* count);
int32_t b2cuda_world_read_joints(B2CWorld* world, B2CJointRead* dst, int32_t capacity, int32_t* count);
int32_t b2cuda_world_read_shapes(B2CWorld* world, B2CShapeRead* dst, int32_t capacity, int32_t* count);
int32_t b2cuda_world_get_counters(B2CWorld* world, B2CCounters* out);
/* Debug aid, no reference counterpart: with B2CUDA_DEBUG_CHECKSUMS=1 in the environment at world upload, out16 receives
   order-independent checksums of the device state between the phases of the last step (scripts/gpu_determinism.py). */
int32_t b2cuda_world_debug_checksums(B2CWorld* world, uint64_t* out16);
int32_t b2cuda_world_get_profile(B2CWorld* world, B2CProfile* out);   /* averages over the steps since the previous call (<= 64) */
/* ---- Sensors (src/Sensor.cs). A shape created with B2CShapeDef.isSensor generates no contacts (src/BroadPhase.cs:168); after every step
 * World.OverlapSensors (src/Sensor.cs:130-218, called from WorldAPI.Step api/World.cs:97-102) queries the three trees with the sensor's AABB,
 * keeps the shapes with enableSensorEvents on another body that pass the filter and whose GJK distance (radii included) is below
 * 10 * FLT_EPSILON, sorted by shape id, and reports the difference to the previous step's set as begin / end events (sensor by sensor in
 * sensor-array order, visitors ascending). End events use the same double buffer as the contact end events (api/World.cs:283-299).
 *   b2cuda_world_get_sensor_events   WorldAPI.GetSensorEvents api/World.cs:289-299; pointers valid until the next step.
 *   b2cuda_shape_enable_sensor_events ShapeAPI.EnableSensorEvents api/Shape.cs:170-174 (a plain flag store).
 *   b2cuda_shape_get_sensor_overlaps  ShapeAPI.GetSensorData api/Shape.cs:409-419: the current overlap set of one sensor shape. */
typedef struct B2CShapeRef { int32_t shapeId; uint32_t generation; } B2CShapeRef;
typedef struct B2CSensorEvent { B2CShapeRef sensor, visitor; } B2CSensorEvent;
int32_t b2cuda_world_get_sensor_events(B2CWorld* world, const B2CSensorEvent** begin, int32_t* beginCount, const B2CSensorEvent** end, int32_t* endCount);
int32_t b2cuda_shape_enable_sensor_events(B2CWorld* world, int32_t shapeId, int32_t flag);
int32_t b2cuda_shape_get_sensor_overlaps(B2CWorld* world, int32_t shapeId, B2CShapeRef* out, int32_t capacity, int32_t* count);
/* ---- Bulk scene upload (new; no reference counterpart: the reference builds scenes call by call). One call replays a recorded
 * sequence of BodyAPI.CreateBody / ShapeAPI.CreateShape / JointAPI.Create*Joint calls in order, so the ids, lists and trees come out exactly as
 * if the calls had been made one by one. Blob layout (all little-endian PODs of this header, tightly packed in this order):
 *   B2CSnapshotHeader, then `opCount` records of { int32 kind; int32 ref; payload }:
 *     kind 0  body   ref unused          payload B2CBodyDef
 *     kind 1  shape  ref = body ordinal  payload B2CShapeDef, B2CGeometry        (ordinal = position among the blob's body records)
 *     kind 2  joint  ref unused          payload B2CJointDef whose bodyIdA / bodyIdB are body ordinals
 * bodyIds / shapeIds / jointIds (each may be NULL) receive the ids in record order. Fails on the first refused record (the world then holds
 * what was created before it). */
#define B2C_SNAPSHOT_MAGIC 0x42324353u   /* "B2CS" */
typedef struct B2CSnapshotHeader { uint32_t magic, version; int32_t opCount, bodyCount, shapeCount, jointCount; } B2CSnapshotHeader;
int32_t b2cuda_world_upload_snapshot(B2CWorld* world, const void* blob, size_t bytes, int32_t* bodyIds, int32_t* shapeIds, int32_t* jointIds);
/* ---- Batches of independent worlds (new; the reference's counterpart is the remark "Each world is completely independent and may be
 * simulated in parallel", api/World.cs:9-10, and a loop over WorldAPI.Step). One handle owns `worldCount` worlds on one device and steps
 * them in lockstep with ONE kernel launch: a thread block per world runs the whole step (pairs, collide, colouring, solver, finalize)
 * with block-local barriers, so world k of a batch is bit-identical to the same world stepped alone with b2cuda_world_step.
 *   b2cuda_batch_world   the handle of world `index`: build it (and read it back, and mutate it between steps) with the ordinary
 *                        b2cuda_create_* / b2cuda_world_read_* / event entry points. b2cuda_world_step and b2cuda_world_destroy are
 *                        refused on such a handle. Device rows do not grow inside a batch: reserve rows for bodies created after the
 *                        first step with B2CWorldDef.bodyCapacity / shapeCapacity / jointCapacity / contactCapacity.
 *   b2cuda_batch_step    WorldAPI.Step for every world; asynchronous on the batch's stream. Step parameters (gravity, contact tuning,
 *                        thresholds, continuous / warm-starting switches) are those of the B2CWorldDef given to b2cuda_batch_create.
 *   b2cuda_batch_reset_world  puts one world back into the state it had at the first b2cuda_batch_step (bodies, contacts, ids, trees).
 *   b2cuda_batch_stats   statistics reduced over all worlds ON THE DEVICE; `deviceDst` (may be NULL) receives the same 8 doubles in device
 *                        memory on the batch's stream, ready for ncclAllReduce: [bodies, sum |v|^2, touching contacts, begin events,
 *                        end events] are sums, [max |v|] a maximum, [min y] a minimum, [error flags] an OR kept as a double. */
typedef struct B2CBatch B2CBatch;
typedef struct B2CBatchStats { double awakeBodies, sumV2, touchingContacts, beginEvents, endEvents, maxSpeed, minY, errorFlags; } B2CBatchStats;
int32_t b2cuda_batch_create(const B2CWorldDef* def, int32_t device, int32_t worldCount, B2CBatch** out);
int32_t b2cuda_batch_destroy(B2CBatch* batch);
int32_t b2cuda_batch_world(B2CBatch* batch, int32_t index, B2CWorld** world);
int32_t b2cuda_batch_step(B2CBatch* batch, float timeStep, int32_t subStepCount);
int32_t b2cuda_batch_synchronize(B2CBatch* batch);
int32_t b2cuda_batch_reset_world(B2CBatch* batch, int32_t index);
int32_t b2cuda_batch_stats(B2CBatch* batch, B2CBatchStats* out, double* deviceDst);
int32_t b2cuda_batch_timer_start(B2CBatch* batch);
int32_t b2cuda_batch_timer_stop(B2CBatch* batch, float* elapsedMs);
/* (new, no reference counterpart) one CUDA-graph launch per step instead of one launch per kernel. The kernels then run back to back whatever
 * the host's launch latency is (many_pyramids: 0.068 against 0.074 ms per step) - use it when steps are enqueued without reading results in
 * between, or when several processes drive GPUs from one host. The first kernel starts later than with direct launches, so a caller that
 * synchronises after every step is better off without: off by default (B2CUDA_GRAPH=1 also enables). */
int32_t b2cuda_world_enable_graph_launch(B2CWorld* world, int32_t flag);
/* (new, no reference counterpart) device-time stopwatch on the world's own stream: CUDA events around whatever was
 * enqueued between the two calls; stop synchronises. Used by bench.py so timing does not depend on torch's stream. start first
 * holds the stream for 60 us (a spinning one-thread kernel) and records its event behind that, so the launches the caller enqueues
 * next are already queued when the region opens: the region is device time and does not contain the host's enqueue latency. */
int32_t b2cuda_world_timer_start(B2CWorld* world);
int32_t b2cuda_world_timer_stop(B2CWorld* world, float* elapsedMs);

#ifdef __cplusplus
}
#endif
#endif
