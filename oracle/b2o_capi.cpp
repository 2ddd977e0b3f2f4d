// ORACLE — TEST INFRASTRUCTURE ONLY (see b2o_math.h header). Parity unpinned by the reference.
// C entry points with the same shapes as include/b2cuda.h (prefix b2o_) so the same scene script can
// drive the oracle and the product. Uses the product header for POD layouts only.
#include "../include/b2cuda.h"
#include "b2o_world.h"
#include <cstring>
#include <algorithm>

using namespace b2o;

static inline Vec2 V(B2CVec2 v) { return {v.x, v.y}; }
static inline B2CVec2 CV(Vec2 v) { return {v.x, v.y}; }
static inline Rot R(B2CRot r) { Rot q; q.c = r.c; q.s = r.s; return q; }
static inline Transform T(B2CTransform t) { Transform x; x.p = V(t.p); x.q = R(t.q); return x; }

extern "C" {

int32_t b2o_world_create(const B2CWorldDef* d, World** out) {
  WorldDef def; def.gravity = V(d->gravity); def.restitutionThreshold = d->restitutionThreshold; def.hitEventThreshold = d->hitEventThreshold;
  def.contactHertz = d->contactHertz; def.contactDampingRatio = d->contactDampingRatio; def.contactSpeed = d->contactSpeed;
  def.maximumLinearSpeed = d->maximumLinearSpeed; def.enableSleep = d->enableSleep != 0; def.enableContinuous = d->enableContinuous != 0;
  def.enableContactSoftening = d->enableContactSoftening != 0;
  *out = new World(def);
  return 0;
}
int32_t b2o_world_destroy(World* w) { delete w; return 0; }
int32_t b2o_create_body(World* w, const B2CBodyDef* d, int32_t* bodyId) {
  BodyDef def; def.type = d->type; def.position = V(d->position); def.rotation = R(d->rotation); def.linearVelocity = V(d->linearVelocity);
  def.angularVelocity = d->angularVelocity; def.linearDamping = d->linearDamping; def.angularDamping = d->angularDamping;
  def.gravityScale = d->gravityScale; def.sleepThreshold = d->sleepThreshold; def.userData = d->userData;
  def.lockLinearX = d->lockLinearX; def.lockLinearY = d->lockLinearY; def.lockAngularZ = d->lockAngularZ; def.enableSleep = d->enableSleep;
  def.isAwake = d->isAwake; def.isBullet = d->isBullet; def.isEnabled = d->isEnabled; def.allowFastRotation = d->allowFastRotation;
  *bodyId = w->CreateBody(def);
  return 0;
}
int32_t b2o_create_shape(World* w, int32_t bodyId, const B2CShapeDef* d, const B2CGeometry* g, int32_t* shapeId) {
  ShapeDef def; def.userData = d->userData; def.material.friction = d->friction; def.material.restitution = d->restitution;
  def.material.rollingResistance = d->rollingResistance; def.material.tangentSpeed = d->tangentSpeed; def.material.userMaterialId = d->userMaterialId;
  def.density = d->density; def.filter.categoryBits = d->categoryBits; def.filter.maskBits = d->maskBits; def.filter.groupIndex = d->groupIndex;
  def.enableCustomFiltering = d->enableCustomFiltering; def.isSensor = d->isSensor; def.enableSensorEvents = d->enableSensorEvents;
  def.enableContactEvents = d->enableContactEvents; def.enableHitEvents = d->enableHitEvents; def.enablePreSolveEvents = d->enablePreSolveEvents;
  def.invokeContactCreation = d->invokeContactCreation; def.updateBodyMass = d->updateBodyMass;
  Geom geom; geom.type = g->type; geom.count = g->count; geom.radius = g->radius; geom.centroid = V(g->centroid);
  for (int i = 0; i < kMaxPolygonVertices; i++) { geom.v[i] = V(g->v[i]); geom.n[i] = V(g->n[i]); }
  *shapeId = w->CreateShape(bodyId, def, geom);
  return 0;
}
int32_t b2o_create_revolute_joint(World* w, const B2CRevoluteJointDef* d, int32_t* jointId) {
  RevoluteJointDef def; def.bodyIdA = d->bodyIdA; def.bodyIdB = d->bodyIdB; def.localFrameA = T(d->localFrameA); def.localFrameB = T(d->localFrameB);
  def.forceThreshold = d->forceThreshold; def.torqueThreshold = d->torqueThreshold; def.constraintHertz = d->constraintHertz;
  def.constraintDampingRatio = d->constraintDampingRatio; def.collideConnected = d->collideConnected; def.userData = d->userData;
  def.targetAngle = d->targetAngle; def.enableSpring = d->enableSpring; def.hertz = d->hertz; def.dampingRatio = d->dampingRatio;
  def.enableLimit = d->enableLimit; def.lowerAngle = d->lowerAngle; def.upperAngle = d->upperAngle; def.enableMotor = d->enableMotor;
  def.maxMotorTorque = d->maxMotorTorque; def.motorSpeed = d->motorSpeed;
  *jointId = w->CreateRevoluteJoint(def);
  return 0;
}
int32_t b2o_create_joint(World* w, const B2CJointDef* d, int32_t* jointId) {
  JointDef def; def.type = d->type; def.bodyIdA = d->bodyIdA; def.bodyIdB = d->bodyIdB; def.localFrameA = T(d->localFrameA); def.localFrameB = T(d->localFrameB);
  def.forceThreshold = d->forceThreshold; def.torqueThreshold = d->torqueThreshold; def.constraintHertz = d->constraintHertz;
  def.constraintDampingRatio = d->constraintDampingRatio; def.collideConnected = d->collideConnected; def.userData = d->userData;
  def.enableSpring = d->enableSpring; def.enableLimit = d->enableLimit; def.enableMotor = d->enableMotor;
  def.hertz = d->hertz; def.dampingRatio = d->dampingRatio; def.target = d->target; def.lower = d->lower; def.upper = d->upper;
  def.maxMotor = d->maxMotor; def.motorSpeed = d->motorSpeed;
  def.length = d->length; def.minLength = d->minLength; def.maxLength = d->maxLength; def.lowerSpringForce = d->lowerSpringForce; def.upperSpringForce = d->upperSpringForce;
  def.angularHertz = d->angularHertz; def.angularDampingRatio = d->angularDampingRatio;
  def.linearVelocity = V(d->linearVelocity); def.angularVelocity = d->angularVelocity; def.maxVelocityForce = d->maxVelocityForce;
  def.maxVelocityTorque = d->maxVelocityTorque; def.maxSpringForce = d->maxSpringForce; def.maxSpringTorque = d->maxSpringTorque;
  *jointId = w->CreateJoint(def);
  return 0;
}
int32_t b2o_joint_set_param(World* w, int32_t id, int32_t param, float v) { return w->JointSetParam(id, param, v) ? 0 : -1; }
int32_t b2o_joint_enable(World* w, int32_t id, int32_t feature, int32_t flag) { return w->JointEnable(id, feature, flag != 0) ? 0 : -1; }
int32_t b2o_body_set_bullet(World* w, int32_t id, int32_t flag) { if (id < 0 || id >= (int)w->bodies.size() || w->bodies[id].id != id) return -1; w->SetBullet(id, flag != 0); return 0; }
int32_t b2o_joint_set_local_frame(World* w, int32_t id, int32_t which, B2CTransform f) { if (id < 0 || id >= (int)w->joints.size() || w->joints[id].jointId != id) return -1; w->JointSetLocalFrame(id, which, T(f)); return 0; }
int32_t b2o_joint_set_collide_connected(World* w, int32_t id, int32_t flag) { if (id < 0 || id >= (int)w->joints.size() || w->joints[id].jointId != id) return -1; w->JointSetCollideConnected(id, flag != 0); return 0; }
int32_t b2o_joint_set_limits(World* w, int32_t id, float lo, float hi) { return w->JointSetLimits(id, lo, hi) ? 0 : -1; }
int32_t b2o_body_set_target_transform(World* w, int32_t id, B2CTransform t, float dt) { w->SetTargetTransform(id, T(t), dt); return 0; }
int32_t b2o_body_set_mass_data(World* w, int32_t id, float mass, float inertia, B2CVec2 c) { w->SetMassData(id, mass, inertia, V(c)); return 0; }
int32_t b2o_body_set_motion_locks(World* w, int32_t id, int32_t lx, int32_t ly, int32_t az) { w->SetMotionLocks(id, lx != 0, ly != 0, az != 0); return 0; }
int32_t b2o_shape_set_surface_material(World* w, int32_t id, float f, float r, float rr, float ts) { w->SetSurfaceMaterial(id, f, r, rr, ts); return 0; }
int32_t b2o_destroy_joint(World* w, int32_t id, int32_t) { if (id < 0 || id >= (int)w->joints.size() || w->joints[id].jointId != id) return -1; w->DestroyJoint(id); return 0; }
int32_t b2o_destroy_body(World* w, int32_t id) { if (id < 0 || id >= (int)w->bodies.size() || w->bodies[id].id != id) return -1; w->DestroyBody(id); return 0; }
int32_t b2o_destroy_shape(World* w, int32_t id, int32_t upd) { if (id < 0 || id >= (int)w->shapes.size() || w->shapes[id].id != id) return -1; w->DestroyShape(id, upd != 0); return 0; }
int32_t b2o_shape_set_density(World* w, int32_t id, float density, int32_t upd) { if (id < 0 || id >= (int)w->shapes.size() || w->shapes[id].id != id) return -1; w->SetDensity(id, density, upd != 0); return 0; }
int32_t b2o_body_apply_mass_from_shapes(World* w, int32_t id) { if (id < 0 || id >= (int)w->bodies.size() || w->bodies[id].id != id) return -1; w->ApplyMassFromShapes(id); return 0; }
int32_t b2o_shape_enable_events(World* w, int32_t id, int32_t which, int32_t flag) {      // api/Shape.cs:182-186, :205-209
  if (id < 0 || id >= (int)w->shapes.size() || w->shapes[id].id != id || which < 0 || which > 1) return -1;
  if (which == 0) w->shapes[id].enableContactEvents = flag != 0; else w->shapes[id].enableHitEvents = flag != 0;
  return 0;
}
int32_t b2o_body_enable_events(World* w, int32_t id, int32_t which, int32_t flag) {       // api/Body.cs:848-872
  if (id < 0 || id >= (int)w->bodies.size() || w->bodies[id].id != id || which < 0 || which > 1) return -1;
  for (int sid = w->bodies[id].headShapeId; sid != -1; sid = w->shapes[sid].nextShapeId) b2o_shape_enable_events(w, sid, which, flag);
  return 0;
}
int32_t b2o_shape_set_filter(World* w, int32_t id, uint64_t categoryBits, uint64_t maskBits, int32_t groupIndex) {
  if (id < 0 || id >= (int)w->shapes.size() || w->shapes[id].id != id) return -1;
  Filter f; f.categoryBits = categoryBits; f.maskBits = maskBits; f.groupIndex = groupIndex;
  w->SetFilter(id, f); return 0;
}
int32_t b2o_shape_set_geometry(World* w, int32_t id, const B2CGeometry* g) {
  if (id < 0 || id >= (int)w->shapes.size() || w->shapes[id].id != id) return -1;
  Geom geom; geom.type = g->type; geom.count = g->count; geom.radius = g->radius; geom.centroid = V(g->centroid);
  for (int i = 0; i < kMaxPolygonVertices; i++) { geom.v[i] = V(g->v[i]); geom.n[i] = V(g->n[i]); }
  w->SetGeometry(id, geom); return 0;
}
int32_t b2o_world_explode(World* w, B2CVec2 position, float radius, float falloff, float impulsePerLength, uint64_t maskBits) { w->Explode(V(position), radius, falloff, impulsePerLength, maskBits); return 0; }
int32_t b2o_body_set_type(World* w, int32_t id, int32_t type) { if (id < 0 || id >= (int)w->bodies.size() || w->bodies[id].id != id || type < 0 || type > 2) return -1; return w->SetType(id, type); }
int32_t b2o_body_set_transform(World* w, int32_t id, B2CVec2 p, B2CRot q) { w->SetTransform(id, V(p), R(q)); return 0; }
int32_t b2o_body_set_linear_velocity(World* w, int32_t id, B2CVec2 v) { w->SetLinearVelocity(id, V(v)); return 0; }
int32_t b2o_body_set_angular_velocity(World* w, int32_t id, float x) { w->SetAngularVelocity(id, x); return 0; }
int32_t b2o_body_apply_force_to_center(World* w, int32_t id, B2CVec2 f, int32_t) { w->ApplyForceToCenter(id, V(f)); return 0; }
int32_t b2o_body_apply_torque(World* w, int32_t id, float t, int32_t) { w->ApplyTorque(id, t); return 0; }
int32_t b2o_body_apply_linear_impulse_to_center(World* w, int32_t id, B2CVec2 i, int32_t) { w->ApplyLinearImpulseToCenter(id, V(i)); return 0; }
int32_t b2o_body_apply_force(World* w, int32_t id, B2CVec2 f, B2CVec2 p, int32_t) { w->ApplyForce(id, V(f), V(p)); return 0; }
int32_t b2o_body_clear_forces(World* w, int32_t id) { w->ClearForces(id); return 0; }
int32_t b2o_body_apply_linear_impulse(World* w, int32_t id, B2CVec2 i, B2CVec2 p, int32_t) { w->ApplyLinearImpulse(id, V(i), V(p)); return 0; }
int32_t b2o_body_apply_angular_impulse(World* w, int32_t id, float i, int32_t) { w->ApplyAngularImpulse(id, i); return 0; }
int32_t b2o_body_set_linear_damping(World* w, int32_t id, float v) { w->SetLinearDamping(id, v); return 0; }
int32_t b2o_body_set_angular_damping(World* w, int32_t id, float v) { w->SetAngularDamping(id, v); return 0; }
int32_t b2o_body_set_gravity_scale(World* w, int32_t id, float v) { w->SetGravityScale(id, v); return 0; }
// api/World.cs:666-670, :678-682, :789-794
int32_t b2o_world_set_restitution_threshold(World* w, float v) { w->restitutionThreshold = v < 0 ? 0 : v; return 0; }
int32_t b2o_world_set_hit_event_threshold(World* w, float v) { w->hitEventThreshold = v < 0 ? 0 : v; return 0; }
int32_t b2o_world_set_maximum_linear_speed(World* w, float v) { w->maxLinearSpeed = v; return 0; }
int32_t b2o_world_set_gravity(World* w, B2CVec2 g) { w->gravity = V(g); return 0; }
int32_t b2o_world_enable_warm_starting(World* w, int32_t f) { w->enableWarmStarting = f != 0; return 0; }
int32_t b2o_world_enable_speculative(World* w, int32_t f) { w->enableSpeculative = f != 0; return 0; }
int32_t b2o_world_enable_sleeping(World* w, int32_t f) { if (f) return -1; w->enableSleep = false; return 0; }   // api/World.cs:631-645 (sleep is not restated)
int32_t b2o_world_enable_continuous(World* w, int32_t f) { w->enableContinuous = f != 0; return 0; }
int32_t b2o_world_set_contact_tuning(World* w, float hertz, float damping, float pushSpeed) {   // api/World.cs:780-787
  w->contactHertz = netClamp(hertz, 0, FLT_MAX); w->contactDampingRatio = netClamp(damping, 0, FLT_MAX); w->contactSpeed = netClamp(pushSpeed, 0, FLT_MAX);
  return 0;
}
int32_t b2o_world_step(World* w, float dt, int32_t sub) {
  int before = w->counters.referenceThrows;
  w->Step(dt, sub);
  if (w->counters.referenceThrows != before) { w->lastError = "the C# reference throws IndexOutOfRangeException in this step (Distance.MakeSimplexFromCache, src/Distance.cs:116-132)"; return B2C_E_UNSUPPORTED; }
  return 0;
}
int32_t b2o_world_synchronize(World*) { return 0; }

static thread_local std::vector<B2CContactTouchEvent> tBegin, tEnd;
static thread_local std::vector<B2CContactHitEvent> tHit;
int32_t b2o_world_get_contact_events(World* w, const B2CContactTouchEvent** begin, int32_t* beginCount, const B2CContactTouchEvent** end,
                                     int32_t* endCount, const B2CContactHitEvent** hit, int32_t* hitCount) {
  tBegin.clear(); tEnd.clear(); tHit.clear();
  for (auto& e : w->contactBeginEvents) tBegin.push_back({e.shapeIdA, e.shapeIdB, e.contactId, e.contactGeneration});
  for (auto& e : w->contactEndEvents[1 - w->endEventArrayIndex]) tEnd.push_back({e.shapeIdA, e.shapeIdB, e.contactId, e.contactGeneration});
  for (auto& e : w->contactHitEvents) tHit.push_back({e.shapeIdA, e.shapeIdB, CV(e.point), CV(e.normal), e.approachSpeed});
  *begin = tBegin.data(); *beginCount = (int)tBegin.size(); *end = tEnd.data(); *endCount = (int)tEnd.size();
  *hit = tHit.data(); *hitCount = (int)tHit.size();
  return 0;
}
static thread_local std::vector<B2CSensorEvent> tSensorBegin, tSensorEnd;
int32_t b2o_world_get_sensor_events(World* w, const B2CSensorEvent** begin, int32_t* beginCount, const B2CSensorEvent** end, int32_t* endCount) {
  tSensorBegin.clear(); tSensorEnd.clear();
  for (auto& e : w->sensorBeginEvents) tSensorBegin.push_back({{e.sensorShapeId, e.sensorGeneration}, {e.visitorShapeId, e.visitorGeneration}});
  for (auto& e : w->sensorEndEvents[1 - w->endEventArrayIndex]) tSensorEnd.push_back({{e.sensorShapeId, e.sensorGeneration}, {e.visitorShapeId, e.visitorGeneration}});
  *begin = tSensorBegin.data(); *beginCount = (int)tSensorBegin.size(); *end = tSensorEnd.data(); *endCount = (int)tSensorEnd.size();
  return 0;
}
int32_t b2o_shape_enable_sensor_events(World* w, int32_t id, int32_t flag) { if (id < 0 || id >= (int)w->shapes.size() || w->shapes[id].id != id) return -1; w->EnableSensorEvents(id, flag != 0); return 0; }
int32_t b2o_shape_get_sensor_overlaps(World* w, int32_t id, B2CShapeRef* out, int32_t capacity, int32_t* count) {
  if (id < 0 || id >= (int)w->shapes.size() || w->shapes[id].id != id) return -1;
  std::vector<Visitor> v(capacity > 0 ? capacity : 0);
  int n = w->GetSensorOverlaps(id, v.data(), capacity);
  for (int i = 0; i < n; i++) out[i] = {v[i].shapeId, v[i].generation};
  *count = n; return 0;
}
int32_t b2o_world_get_joint_events(World* w, const int32_t** ids, int32_t* count) { *ids = w->jointEvents.data(); *count = (int)w->jointEvents.size(); return 0; }
int32_t b2o_world_read_bodies(World* w, B2CBodyRead* dst, int32_t capacity, int32_t* count) {
  int n = (int)w->bodies.size(); *count = n;
  for (int i = 0; i < n && i < capacity; i++) {
    B2CBodyRead r; memset(&r, 0, sizeof(r));
    const Body& b = w->bodies[i];
    if (b.id != -1) {
      const BodySim& s = w->GetBodySim(b);
      r.transform.p = CV(s.transform.p); r.transform.q = {s.transform.q.c, s.transform.q.s}; r.center = CV(s.center); r.flags = s.flags;
      if (const BodyState* st = w->GetBodyState(b)) { r.linearVelocity = CV(st->linearVelocity); r.angularVelocity = st->angularVelocity; }
    }
    dst[i] = r;
  }
  return 0;
}
int32_t b2o_body_get(World* w, int32_t id, B2CBodyRead* out) {      // api/Body.cs:288, :345-357, :580
  if (id < 0 || id >= (int)w->bodies.size() || w->bodies[id].id == -1) return 1;
  B2CBodyRead r; memset(&r, 0, sizeof(r));
  const Body& b = w->bodies[id]; const BodySim& s = w->GetBodySim(b);
  r.transform.p = CV(s.transform.p); r.transform.q = {s.transform.q.c, s.transform.q.s}; r.center = CV(s.center); r.flags = s.flags;
  if (const BodyState* st = w->GetBodyState(b)) { r.linearVelocity = CV(st->linearVelocity); r.angularVelocity = st->angularVelocity; }
  *out = r;
  return 0;
}
static thread_local std::vector<B2CBodyRead> tBodies;
int32_t b2o_world_get_body_events(World* w, const B2CBodyRead** records, int32_t* count) {
  int n = (int)w->bodies.size();
  tBodies.resize(n);
  b2o_world_read_bodies(w, tBodies.data(), n, count);
  *records = tBodies.data();
  return 0;
}
static thread_local std::vector<B2CBodyMoveEvent> tMoves;
int32_t b2o_world_get_body_move_events(World* w, const B2CBodyMoveEvent** events, int32_t* count) {   // src/Solver.cs:432-438
  const auto& sims = w->awakeSet.bodySims;
  tMoves.resize(sims.size());
  for (size_t i = 0; i < sims.size(); i++) {
    B2CBodyMoveEvent e; e.transform.p = CV(sims[i].transform.p); e.transform.q = {sims[i].transform.q.c, sims[i].transform.q.s}; e.bodyId = sims[i].bodyId; e.fellAsleep = 0;
    tMoves[i] = e;
  }
  *events = tMoves.data(); *count = w->lastStepDtZero ? 0 : (int)sims.size();
  return 0;
}
static void FillPoint(B2CManifoldPointRead& o, const ManifoldPoint& p) {
  o.point = CV(p.point); o.anchorA = CV(p.anchorA); o.anchorB = CV(p.anchorB); o.separation = p.separation; o.normalImpulse = p.normalImpulse;
  o.tangentImpulse = p.tangentImpulse; o.totalNormalImpulse = p.totalNormalImpulse; o.normalVelocity = p.normalVelocity; o.id = p.id; o.persisted = p.persisted;
}
int32_t b2o_world_read_contacts(World* w, B2CContactRead* dst, int32_t capacity, int32_t* count) {
  int n = 0;
  for (size_t i = 0; i < w->contacts.size(); i++) {
    const Contact& c = w->contacts[i];
    if (c.contactId == -1) continue;
    if (n < capacity) {
      const ContactSim& s = w->GetContactSim(c);
      B2CContactRead r; memset(&r, 0, sizeof(r));
      r.contactId = c.contactId; r.shapeIdA = c.shapeIdA; r.shapeIdB = c.shapeIdB; r.colorIndex = c.colorIndex; r.localIndex = c.colorIndex != -1 ? c.localIndex : -1;
      r.flags = (c.flags & kContactTouching) ? 1u : 0u; r.pointCount = s.manifold.pointCount; r.normal = CV(s.manifold.normal);
      r.rollingImpulse = s.manifold.rollingImpulse; r.friction = s.friction; r.restitution = s.restitution;
      r.rollingResistance = s.rollingResistance; r.tangentSpeed = s.tangentSpeed;
      FillPoint(r.points[0], s.manifold.point0); FillPoint(r.points[1], s.manifold.point1);
      dst[n] = r;
    }
    n++;
  }
  *count = n;
  return 0;
}
int32_t b2o_world_read_joints(World* w, B2CJointRead* dst, int32_t capacity, int32_t* count) {
  int n = 0;
  for (size_t i = 0; i < w->joints.size(); i++) {
    const Joint& j = w->joints[i];
    if (j.jointId == -1 || j.colorIndex == -1) continue;
    if (n < capacity) {
      const JointSim& s = w->GetJointSim(j);
      dst[n] = {j.jointId, j.colorIndex, CV(s.linearImpulse), s.springImpulse, s.motorImpulse, s.lowerImpulse, s.upperImpulse, s.type, CV(s.linearImpulse2)};
    }
    n++;
  }
  *count = n;
  return 0;
}
int32_t b2o_world_read_shapes(World* w, B2CShapeRead* dst, int32_t capacity, int32_t* count) {
  int n = (int)w->shapes.size(); *count = n;
  for (int i = 0; i < n && i < capacity; i++) {
    const Shape& s = w->shapes[i];
    if (s.id == -1) { dst[i] = {{0, 0}, {0, 0}, {0, 0}, {0, 0}, -1, 0}; continue; }      // destroyed with its body
    dst[i] = {CV(s.aabb.lo), CV(s.aabb.hi), CV(s.fatAABB.lo), CV(s.fatAABB.hi), s.proxyKey, s.enlargedAABB ? 1 : 0};
  }
  return 0;
}
int32_t b2o_world_get_counters(World* w, B2CCounters* out) {
  memset(out, 0, sizeof(*out));
  out->bodyCount = w->bodyIdPool.GetIdCount(); out->awakeBodyCount = (int)w->awakeSet.bodySims.size(); out->shapeCount = w->shapeIdPool.GetIdCount();
  out->contactCount = w->contactIdPool.GetIdCount(); out->jointCount = w->jointIdPool.GetIdCount(); out->contactIdCapacity = w->contactIdPool.GetIdCapacity();
  int touching = 0, active = 0;
  for (int i = 0; i < kGraphColorCount; i++) {
    out->colorCounts[i] = (int)(w->colors[i].contactSims.size() + w->colors[i].jointSims.size()); touching += (int)w->colors[i].contactSims.size();
    if (i < kOverflowIndex && out->colorCounts[i] > 0) active++;
  }
  out->touchingContactCount = touching; out->activeColorCount = active; out->moveCount = (int)w->moveArray.size();
  out->newPairCount = (int)w->lastNewPairKeys.size(); out->toiCandidates = w->counters.toiCandidates; out->unsupportedCount = w->counters.unsupported;
  out->stepCount = (int)w->stepIndex;
  return 0;
}
// pair keys created by the last step, in the reference's creation order (ordering rule A.3-2)
int32_t b2o_world_get_new_pairs(World* w, uint64_t* dst, int32_t capacity, int32_t* count) {
  *count = (int)w->lastNewPairKeys.size();
  for (int i = 0; i < *count && i < capacity; i++) dst[i] = w->lastNewPairKeys[i];
  return 0;
}
// DFS leaf rank (child2 first, src/DynamicTree.cs:539,552) of every proxy of one tree; -1 for non-leaves
int32_t b2o_world_get_tree_ranks(World* w, int32_t treeType, int32_t* dst, int32_t capacity, int32_t* nodeCapacity) {
  const DynamicTree& t = w->trees[treeType];
  *nodeCapacity = (int)t.nodes.size();
  for (int i = 0; i < capacity; i++) dst[i] = -1;
  if (t.root == -1) return 0;
  std::vector<int> stack{t.root}; int rank = 0;
  while (!stack.empty()) {
    int id = stack.back(); stack.pop_back();
    const TreeNode& n = t.nodes[id];
    if (n.flags & kNodeLeaf) { if (id < capacity) dst[id] = rank; rank++; }
    else { stack.push_back(n.child1); stack.push_back(n.child2); }
  }
  return 0;
}
int32_t b2o_world_get_profile_ms(World* w, double* out8) { for (int i = 0; i < 8; i++) out8[i] = w->profileMs[i]; return 0; }
const char* b2o_last_error(World* w) { return w->lastError.c_str(); }

}  // extern "C"
