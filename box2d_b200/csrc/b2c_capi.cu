// libb2cuda — C-ABI entry points declared in include/b2cuda.h. No exceptions cross the ABI; every call returns a status.
#include "b2c_host.cuh"
#include <cstring>
#include <new>

using namespace b2c;

struct B2CWorld { World* w; };
static thread_local std::string g_lastError;
void b2cSetLastError(const std::string& msg) { g_lastError = msg; }   // for entry points that live in other files (b2c_noalloc.cu)
static int done(B2CWorld* h, int rc) {
  if (h && h->w) h->w->markDeviceDirty();        // batch member: the call may have changed counts / capacities the kernels take from WorldDev
  if (rc != B2C_OK && h && h->w) g_lastError = h->w->error;
  return rc;
}
#define CHECK_HANDLE(h) do { if (!(h) || !(h)->w) { g_lastError = "null world handle"; return B2C_E_INVALID; } } while (0)

extern "C" {

const char* b2cuda_last_error(void) { return g_lastError.c_str(); }

int32_t b2cuda_default_world_def(B2CWorldDef* d) {          // include/Types.cs:75-144
  if (!d) return B2C_E_INVALID;
  memset(d, 0, sizeof(*d));
  d->gravity = {0.0f, -10.0f}; d->restitutionThreshold = 1.0f; d->hitEventThreshold = 1.0f; d->contactHertz = 30.0f;
  d->contactDampingRatio = 10.0f; d->contactSpeed = 3.0f; d->maximumLinearSpeed = 400.0f;
  d->enableSleep = 1; d->enableContinuous = 1; d->enableContactSoftening = 0;
  return B2C_OK;
}
int32_t b2cuda_default_body_def(B2CBodyDef* d) {            // include/Types.cs:170-223
  if (!d) return B2C_E_INVALID;
  memset(d, 0, sizeof(*d));
  d->type = 0; d->rotation = {1.0f, 0.0f}; d->gravityScale = 1.0f; d->sleepThreshold = 0.05f; d->enableSleep = 1; d->isAwake = 1; d->isEnabled = 1;
  return B2C_OK;
}
int32_t b2cuda_default_shape_def(B2CShapeDef* d) {          // include/Types.cs:285-344
  if (!d) return B2C_E_INVALID;
  memset(d, 0, sizeof(*d));
  d->friction = 0.6f; d->density = 1.0f; d->categoryBits = ~0ull; d->maskBits = ~0ull; d->invokeContactCreation = 1; d->updateBodyMass = 1;
  return B2C_OK;
}
int32_t b2cuda_default_revolute_joint_def(B2CRevoluteJointDef* d) {   // include/Types.cs:431-456,569-597
  if (!d) return B2C_E_INVALID;
  memset(d, 0, sizeof(*d));
  d->bodyIdA = -1; d->bodyIdB = -1; d->localFrameA.q = {1.0f, 0.0f}; d->localFrameB.q = {1.0f, 0.0f};
  d->forceThreshold = FLT_MAX; d->torqueThreshold = FLT_MAX; d->constraintHertz = 60.0f; d->constraintDampingRatio = 2.0f; d->drawScale = 1.0f;
  return B2C_OK;
}
int32_t b2cuda_default_joint_def(int32_t type, B2CJointDef* d) {   // include/Types.cs:431-456 (base) and :460-586 (per type)
  if (!d || type < B2C_JOINT_DISTANCE || type > B2C_JOINT_WHEEL) return B2C_E_INVALID;
  memset(d, 0, sizeof(*d));
  d->type = type; d->bodyIdA = -1; d->bodyIdB = -1; d->localFrameA.q = {1.0f, 0.0f}; d->localFrameB.q = {1.0f, 0.0f};
  d->forceThreshold = FLT_MAX; d->torqueThreshold = FLT_MAX; d->constraintHertz = 60.0f; d->constraintDampingRatio = 2.0f; d->drawScale = 1.0f;
  d->length = 1.0f; d->maxLength = 100000.0f; d->lowerSpringForce = -FLT_MAX; d->upperSpringForce = FLT_MAX;   // DistanceJointDef :465-482
  return B2C_OK;
}

int32_t b2cuda_world_create(const B2CWorldDef* def, int32_t device, B2CWorld** out) {
  if (!def || !out) { g_lastError = "null argument"; return B2C_E_INVALID; }
  *out = nullptr;
  int count = 0;
  cudaError_t e = cudaGetDeviceCount(&count);
  if (e != cudaSuccess || count == 0) { g_lastError = std::string("no usable CUDA device: ") + cudaGetErrorString(e) + " (libb2cuda has no CPU fallback)"; return B2C_E_NO_DEVICE; }
  if (device < 0 || device >= count) { g_lastError = "device index out of range"; return B2C_E_NO_DEVICE; }
  if (def->enableSleep) { g_lastError = "enableSleep=1 (islands/sleep, src/Island.cs) is not supported in this build; set enableSleep=0"; return B2C_E_UNSUPPORTED; }
  B2CWorld* h = new (std::nothrow) B2CWorld();
  if (!h) return B2C_E_INVALID;
  h->w = new (std::nothrow) World(*def, device);
  if (!h->w) { delete h; return B2C_E_INVALID; }
  *out = h;
  return B2C_OK;
}
int32_t b2cuda_world_destroy(B2CWorld* h) {
  if (!h) return B2C_E_INVALID;
  if (h->w && h->w->inBatch()) { g_lastError = "this world belongs to a batch: b2cuda_batch_destroy frees it"; return B2C_E_INVALID; }
  delete h->w; delete h; return B2C_OK;
}

int32_t b2cuda_create_body(B2CWorld* h, const B2CBodyDef* def, int32_t* bodyId) { CHECK_HANDLE(h); if (!def || !bodyId) return B2C_E_INVALID; return done(h, h->w->createBody(*def, bodyId)); }
int32_t b2cuda_create_shape(B2CWorld* h, int32_t bodyId, const B2CShapeDef* def, const B2CGeometry* g, int32_t* shapeId) {
  CHECK_HANDLE(h); if (!def || !g || !shapeId) return B2C_E_INVALID; return done(h, h->w->createShape(bodyId, *def, *g, shapeId));
}
int32_t b2cuda_create_revolute_joint(B2CWorld* h, const B2CRevoluteJointDef* def, int32_t* jointId) { CHECK_HANDLE(h); if (!def || !jointId) return B2C_E_INVALID; return done(h, h->w->createRevoluteJoint(*def, jointId)); }
int32_t b2cuda_create_joint(B2CWorld* h, const B2CJointDef* def, int32_t* jointId) { CHECK_HANDLE(h); if (!def || !jointId) return B2C_E_INVALID; return done(h, h->w->createJoint(*def, jointId)); }
int32_t b2cuda_destroy_joint(B2CWorld* h, int32_t id, int32_t) { CHECK_HANDLE(h); return done(h, h->w->destroyJoint(id)); }
int32_t b2cuda_destroy_body(B2CWorld* h, int32_t id) { CHECK_HANDLE(h); return done(h, h->w->destroyBody(id)); }
int32_t b2cuda_destroy_shape(B2CWorld* h, int32_t id, int32_t upd) { CHECK_HANDLE(h); return done(h, h->w->destroyShape(id, upd != 0)); }
int32_t b2cuda_shape_set_density(B2CWorld* h, int32_t id, float density, int32_t upd) { CHECK_HANDLE(h); return done(h, h->w->setShapeDensity(id, density, upd != 0)); }
int32_t b2cuda_body_apply_mass_from_shapes(B2CWorld* h, int32_t id) { CHECK_HANDLE(h); return done(h, h->w->applyMassFromShapes(id)); }
int32_t b2cuda_shape_enable_events(B2CWorld* h, int32_t id, int32_t which, int32_t flag) { CHECK_HANDLE(h); return done(h, h->w->shapeEnableEvents(id, which, flag != 0)); }
int32_t b2cuda_body_enable_events(B2CWorld* h, int32_t id, int32_t which, int32_t flag) { CHECK_HANDLE(h); return done(h, h->w->bodyEnableEvents(id, which, flag != 0)); }
int32_t b2cuda_shape_set_filter(B2CWorld* h, int32_t id, uint64_t categoryBits, uint64_t maskBits, int32_t groupIndex) { CHECK_HANDLE(h); return done(h, h->w->setShapeFilter(id, categoryBits, maskBits, groupIndex)); }
int32_t b2cuda_shape_set_geometry(B2CWorld* h, int32_t id, const B2CGeometry* g) { CHECK_HANDLE(h); if (!g) return B2C_E_INVALID; return done(h, h->w->setShapeGeometry(id, *g)); }
int32_t b2cuda_body_set_type(B2CWorld* h, int32_t id, int32_t type) { CHECK_HANDLE(h); return done(h, h->w->setBodyType(id, type)); }
int32_t b2cuda_world_explode(B2CWorld* h, B2CVec2 position, float radius, float falloff, float impulsePerLength, uint64_t maskBits) { CHECK_HANDLE(h); return done(h, h->w->explode(mk(position.x, position.y), radius, falloff, impulsePerLength, maskBits)); }
int32_t b2cuda_joint_set_param(B2CWorld* h, int32_t id, int32_t param, float v) { CHECK_HANDLE(h); return done(h, h->w->jointOp(id, 0, param, v, 0.0f)); }
int32_t b2cuda_joint_enable(B2CWorld* h, int32_t id, int32_t feature, int32_t flag) { CHECK_HANDLE(h); return done(h, h->w->jointOp(id, 1, feature, flag ? 1.0f : 0.0f, 0.0f)); }
int32_t b2cuda_joint_set_limits(B2CWorld* h, int32_t id, float lo, float hi) { CHECK_HANDLE(h); return done(h, h->w->jointOp(id, 2, 0, lo, hi)); }
int32_t b2cuda_body_set_transform(B2CWorld* h, int32_t id, B2CVec2 p, B2CRot q) { CHECK_HANDLE(h); return done(h, h->w->setTransform(id, mk(p.x, p.y), mkRot(q.c, q.s))); }
int32_t b2cuda_body_set_linear_velocity(B2CWorld* h, int32_t id, B2CVec2 v) { CHECK_HANDLE(h); return done(h, h->w->setLinearVelocity(id, mk(v.x, v.y))); }
int32_t b2cuda_body_set_angular_velocity(B2CWorld* h, int32_t id, float w) { CHECK_HANDLE(h); return done(h, h->w->setAngularVelocity(id, w)); }
int32_t b2cuda_body_apply_force_to_center(B2CWorld* h, int32_t id, B2CVec2 f, int32_t) { CHECK_HANDLE(h); return done(h, h->w->applyForceToCenter(id, mk(f.x, f.y))); }
int32_t b2cuda_body_apply_torque(B2CWorld* h, int32_t id, float t, int32_t) { CHECK_HANDLE(h); return done(h, h->w->applyTorque(id, t)); }
int32_t b2cuda_body_apply_linear_impulse_to_center(B2CWorld* h, int32_t id, B2CVec2 i, int32_t) { CHECK_HANDLE(h); return done(h, h->w->applyLinearImpulseToCenter(id, mk(i.x, i.y))); }
int32_t b2cuda_body_apply_force(B2CWorld* h, int32_t id, B2CVec2 f, B2CVec2 p, int32_t) { CHECK_HANDLE(h); return done(h, h->w->applyForce(id, mk(f.x, f.y), mk(p.x, p.y))); }
int32_t b2cuda_body_clear_forces(B2CWorld* h, int32_t id) { CHECK_HANDLE(h); return done(h, h->w->clearForces(id)); }
int32_t b2cuda_body_apply_linear_impulse(B2CWorld* h, int32_t id, B2CVec2 i, B2CVec2 p, int32_t) { CHECK_HANDLE(h); return done(h, h->w->applyLinearImpulse(id, mk(i.x, i.y), mk(p.x, p.y))); }
int32_t b2cuda_body_apply_angular_impulse(B2CWorld* h, int32_t id, float i, int32_t) { CHECK_HANDLE(h); return done(h, h->w->applyAngularImpulse(id, i)); }
int32_t b2cuda_body_set_linear_damping(B2CWorld* h, int32_t id, float v) { CHECK_HANDLE(h); return done(h, h->w->setLinearDamping(id, v)); }
int32_t b2cuda_body_set_angular_damping(B2CWorld* h, int32_t id, float v) { CHECK_HANDLE(h); return done(h, h->w->setAngularDamping(id, v)); }
int32_t b2cuda_body_set_gravity_scale(B2CWorld* h, int32_t id, float v) { CHECK_HANDLE(h); return done(h, h->w->setGravityScale(id, v)); }
int32_t b2cuda_body_set_target_transform(B2CWorld* h, int32_t id, B2CTransform t, float dt) { CHECK_HANDLE(h); return done(h, h->w->setTargetTransform(id, mkXf(mk(t.p.x, t.p.y), mkRot(t.q.c, t.q.s)), dt)); }
int32_t b2cuda_body_set_mass_data(B2CWorld* h, int32_t id, float mass, float inertia, B2CVec2 c) { CHECK_HANDLE(h); return done(h, h->w->setMassData(id, mass, inertia, mk(c.x, c.y))); }
int32_t b2cuda_body_set_motion_locks(B2CWorld* h, int32_t id, int32_t lx, int32_t ly, int32_t az) { CHECK_HANDLE(h); return done(h, h->w->setMotionLocks(id, lx != 0, ly != 0, az != 0)); }
int32_t b2cuda_shape_set_surface_material(B2CWorld* h, int32_t id, float f, float r, float rr, float ts) { CHECK_HANDLE(h); return done(h, h->w->setSurfaceMaterial(id, f, r, rr, ts)); }
int32_t b2cuda_body_get(B2CWorld* h, int32_t id, B2CBodyRead* out) { CHECK_HANDLE(h); if (!out) return B2C_E_INVALID; return done(h, h->w->readBody(id, out)); }
int32_t b2cuda_world_set_restitution_threshold(B2CWorld* h, float v) { CHECK_HANDLE(h); h->w->restitutionThreshold = netClamp(v, 0.0f, FLT_MAX); return B2C_OK; }
int32_t b2cuda_world_set_hit_event_threshold(B2CWorld* h, float v) { CHECK_HANDLE(h); h->w->hitEventThreshold = netClamp(v, 0.0f, FLT_MAX); return B2C_OK; }
int32_t b2cuda_world_set_maximum_linear_speed(B2CWorld* h, float v) { CHECK_HANDLE(h); if (!(v > 0.0f)) return B2C_E_INVALID; h->w->maxLinearSpeed = v; return B2C_OK; }
int32_t b2cuda_world_set_gravity(B2CWorld* h, B2CVec2 g) { CHECK_HANDLE(h); h->w->gravity = mk(g.x, g.y); return B2C_OK; }
int32_t b2cuda_world_enable_warm_starting(B2CWorld* h, int32_t f) { CHECK_HANDLE(h); h->w->enableWarmStarting = f != 0; return B2C_OK; }
int32_t b2cuda_world_enable_speculative(B2CWorld* h, int32_t f) { CHECK_HANDLE(h); h->w->enableSpeculative = f != 0; return B2C_OK; }
int32_t b2cuda_world_enable_sleeping(B2CWorld* h, int32_t f) {
  CHECK_HANDLE(h);
  if (f) { h->w->error = "enable_sleeping(1): islands/sleep (src/Island.cs) are not supported in this build; sleeping stays off"; return done(h, B2C_E_UNSUPPORTED); }
  h->w->enableSleep = false; return B2C_OK;
}
int32_t b2cuda_world_enable_continuous(B2CWorld* h, int32_t f) { CHECK_HANDLE(h); h->w->enableContinuous = f != 0; return B2C_OK; }
int32_t b2cuda_world_set_contact_tuning(B2CWorld* h, float hertz, float damping, float push) {   // api/World.cs:780-787
  CHECK_HANDLE(h);
  h->w->contactHertz = netClamp(hertz, 0.0f, FLT_MAX); h->w->contactDampingRatio = netClamp(damping, 0.0f, FLT_MAX); h->w->contactSpeed = netClamp(push, 0.0f, FLT_MAX);
  return B2C_OK;
}
int32_t b2cuda_world_step(B2CWorld* h, float dt, int32_t sub) { CHECK_HANDLE(h); return done(h, h->w->step(dt, sub)); }
int32_t b2cuda_world_synchronize(B2CWorld* h) { CHECK_HANDLE(h); return done(h, h->w->synchronize()); }
int32_t b2cuda_world_get_contact_events(B2CWorld* h, const B2CContactTouchEvent** b, int32_t* nb, const B2CContactTouchEvent** e, int32_t* ne, const B2CContactHitEvent** hit, int32_t* nh) {
  CHECK_HANDLE(h); return done(h, h->w->getContactEvents(b, nb, e, ne, hit, nh));
}
int32_t b2cuda_world_get_joint_events(B2CWorld* h, const int32_t** ids, int32_t* n) { CHECK_HANDLE(h); return done(h, h->w->getJointEvents(ids, n)); }
int32_t b2cuda_world_read_bodies(B2CWorld* h, B2CBodyRead* dst, int32_t cap, int32_t* n) { CHECK_HANDLE(h); return done(h, h->w->readBodies(dst, cap, n)); }
int32_t b2cuda_world_get_body_move_events(B2CWorld* h, const B2CBodyMoveEvent** e, int32_t* n) { CHECK_HANDLE(h); if (!e || !n) return B2C_E_INVALID; return done(h, h->w->getBodyMoveEvents(e, n)); }
int32_t b2cuda_world_get_body_events(B2CWorld* h, const B2CBodyRead** r, int32_t* n) { CHECK_HANDLE(h); if (!r || !n) return B2C_E_INVALID; return done(h, h->w->getBodyEvents(r, n)); }
int32_t b2cuda_world_read_contacts(B2CWorld* h, B2CContactRead* dst, int32_t cap, int32_t* n) { CHECK_HANDLE(h); return done(h, h->w->readContacts(dst, cap, n)); }
int32_t b2cuda_world_read_joints(B2CWorld* h, B2CJointRead* dst, int32_t cap, int32_t* n) { CHECK_HANDLE(h); return done(h, h->w->readJoints(dst, cap, n)); }
int32_t b2cuda_world_read_shapes(B2CWorld* h, B2CShapeRead* dst, int32_t cap, int32_t* n) { CHECK_HANDLE(h); return done(h, h->w->readShapes(dst, cap, n)); }
int32_t b2cuda_world_get_counters(B2CWorld* h, B2CCounters* out) { CHECK_HANDLE(h); return done(h, h->w->getCounters(out)); }
int32_t b2cuda_world_debug_checksums(B2CWorld* h, uint64_t* out16) { CHECK_HANDLE(h); return done(h, h->w->debugChecksums((unsigned long long*)out16)); }
int32_t b2cuda_world_get_profile(B2CWorld* h, B2CProfile* out) { CHECK_HANDLE(h); return done(h, h->w->getProfile(out)); }

int32_t b2cuda_world_enable_graph_launch(B2CWorld* h, int32_t flag) { CHECK_HANDLE(h); h->w->enableGraphLaunch(flag != 0); return B2C_OK; }
int32_t b2cuda_world_timer_start(B2CWorld* h) { CHECK_HANDLE(h); return done(h, h->w->timerStart()); }
int32_t b2cuda_world_timer_stop(B2CWorld* h, float* ms) { CHECK_HANDLE(h); if (!ms) return B2C_E_INVALID; return done(h, h->w->timerStop(ms)); }

int32_t b2cuda_world_get_sensor_events(B2CWorld* h, const B2CSensorEvent** begin, int32_t* nb, const B2CSensorEvent** end, int32_t* ne) {
  CHECK_HANDLE(h); if (!begin || !nb || !end || !ne) return B2C_E_INVALID;
  return done(h, h->w->getSensorEvents(begin, nb, end, ne));
}
int32_t b2cuda_shape_enable_sensor_events(B2CWorld* h, int32_t id, int32_t flag) { CHECK_HANDLE(h); return done(h, h->w->shapeEnableSensorEvents(id, flag != 0)); }
int32_t b2cuda_shape_get_sensor_overlaps(B2CWorld* h, int32_t id, B2CShapeRef* out, int32_t capacity, int32_t* count) {
  CHECK_HANDLE(h); if (!count || (capacity > 0 && !out)) return B2C_E_INVALID;
  return done(h, h->w->shapeGetSensorOverlaps(id, out, capacity, count));
}
int32_t b2cuda_joint_set_local_frame(B2CWorld* h, int32_t id, int32_t which, B2CTransform f) { CHECK_HANDLE(h); return done(h, h->w->jointSetLocalFrame(id, which, mkXf(mk(f.p.x, f.p.y), mkRot(f.q.c, f.q.s)))); }
int32_t b2cuda_joint_set_collide_connected(B2CWorld* h, int32_t id, int32_t flag) { CHECK_HANDLE(h); return done(h, h->w->jointSetCollideConnected(id, flag != 0)); }
int32_t b2cuda_body_set_sleep_threshold(B2CWorld* h, int32_t id, float v) { CHECK_HANDLE(h); return done(h, h->w->bodySetSleepThreshold(id, v)); }
int32_t b2cuda_body_enable_sleep(B2CWorld* h, int32_t id, int32_t flag) { CHECK_HANDLE(h); return done(h, h->w->bodyEnableSleep(id, flag != 0)); }
int32_t b2cuda_body_set_awake(B2CWorld* h, int32_t id, int32_t awake) { CHECK_HANDLE(h); return done(h, h->w->bodySetAwake(id, awake != 0)); }
int32_t b2cuda_body_wake_touching(B2CWorld* h, int32_t id) { CHECK_HANDLE(h); return done(h, h->w->bodyWakeTouching(id)); }
int32_t b2cuda_body_disable(B2CWorld* h, int32_t id) { CHECK_HANDLE(h); return done(h, h->w->bodySetEnabled(id, false)); }
int32_t b2cuda_body_enable(B2CWorld* h, int32_t id) { CHECK_HANDLE(h); return done(h, h->w->bodySetEnabled(id, true)); }
int32_t b2cuda_body_set_bullet(B2CWorld* h, int32_t id, int32_t flag) { CHECK_HANDLE(h); return done(h, h->w->bodySetBullet(id, flag != 0)); }
int32_t b2cuda_world_upload_snapshot(B2CWorld* h, const void* blob, size_t bytes, int32_t* bodyIds, int32_t* shapeIds, int32_t* jointIds) {
  CHECK_HANDLE(h);
  const unsigned char* p = (const unsigned char*)blob; const unsigned char* end = p + bytes;
  B2CSnapshotHeader hd;
  if (!blob || bytes < sizeof(hd)) { g_lastError = "snapshot blob is null or shorter than its header"; return B2C_E_INVALID; }
  memcpy(&hd, p, sizeof(hd)); p += sizeof(hd);
  if (hd.magic != B2C_SNAPSHOT_MAGIC || hd.version != 1u || hd.opCount < 0) { g_lastError = "snapshot blob: bad magic / version"; return B2C_E_INVALID; }
  std::vector<int32_t> bodyOf; bodyOf.reserve(hd.bodyCount > 0 ? hd.bodyCount : 0);
  int nb = 0, ns = 0, nj = 0;
  for (int i = 0; i < hd.opCount; i++) {
    int32_t rec[2];
    if ((size_t)(end - p) < sizeof(rec)) { g_lastError = "snapshot blob truncated"; return B2C_E_INVALID; }
    memcpy(rec, p, sizeof(rec)); p += sizeof(rec);
    int rc = B2C_OK;
    if (rec[0] == 0) {
      B2CBodyDef def; if ((size_t)(end - p) < sizeof(def)) { g_lastError = "snapshot blob truncated"; return B2C_E_INVALID; }
      memcpy(&def, p, sizeof(def)); p += sizeof(def);
      int32_t id = -1; rc = h->w->createBody(def, &id);
      if (rc == B2C_OK) { bodyOf.push_back(id); if (bodyIds) bodyIds[nb] = id; nb++; }
    } else if (rec[0] == 1) {
      B2CShapeDef def; B2CGeometry g; if ((size_t)(end - p) < sizeof(def) + sizeof(g)) { g_lastError = "snapshot blob truncated"; return B2C_E_INVALID; }
      memcpy(&def, p, sizeof(def)); p += sizeof(def); memcpy(&g, p, sizeof(g)); p += sizeof(g);
      if (rec[1] < 0 || rec[1] >= (int)bodyOf.size()) { g_lastError = "snapshot blob: shape refers to a body ordinal that has not been created"; return B2C_E_INVALID; }
      int32_t id = -1; rc = h->w->createShape(bodyOf[rec[1]], def, g, &id);
      if (rc == B2C_OK) { if (shapeIds) shapeIds[ns] = id; ns++; }
    } else if (rec[0] == 2) {
      B2CJointDef def; if ((size_t)(end - p) < sizeof(def)) { g_lastError = "snapshot blob truncated"; return B2C_E_INVALID; }
      memcpy(&def, p, sizeof(def)); p += sizeof(def);
      if (def.bodyIdA < 0 || def.bodyIdA >= (int)bodyOf.size() || def.bodyIdB < 0 || def.bodyIdB >= (int)bodyOf.size()) { g_lastError = "snapshot blob: joint refers to a body ordinal that has not been created"; return B2C_E_INVALID; }
      def.bodyIdA = bodyOf[def.bodyIdA]; def.bodyIdB = bodyOf[def.bodyIdB];
      int32_t id = -1; rc = h->w->createJoint(def, &id);
      if (rc == B2C_OK) { if (jointIds) jointIds[nj] = id; nj++; }
    } else { g_lastError = "snapshot blob: unknown record kind"; return B2C_E_INVALID; }
    if (rc != B2C_OK) return done(h, rc);
  }
  return done(h, B2C_OK);
}
// ---- batches of independent worlds (b2c_batch.cu) ----
}  // extern "C"
#include <vector>
struct B2CBatch { Batch* b; std::vector<B2CWorld> handles; };
#define CHECK_BATCH(h) do { if (!(h) || !(h)->b) { g_lastError = "null batch handle"; return B2C_E_INVALID; } } while (0)
static int doneB(B2CBatch* h, int rc) { if (rc != B2C_OK) g_lastError = h->b->error; return rc; }
extern "C" {
int32_t b2cuda_batch_create(const B2CWorldDef* def, int32_t device, int32_t worldCount, B2CBatch** out) {
  if (!def || !out || worldCount < 0) { g_lastError = "null argument or negative world count"; return B2C_E_INVALID; }
  *out = nullptr;
  int count = 0;
  cudaError_t e = cudaGetDeviceCount(&count);
  if (e != cudaSuccess || count == 0) { g_lastError = std::string("no usable CUDA device: ") + cudaGetErrorString(e) + " (libb2cuda has no CPU fallback)"; return B2C_E_NO_DEVICE; }
  if (device < 0 || device >= count) { g_lastError = "device index out of range"; return B2C_E_NO_DEVICE; }
  if (def->enableSleep) { g_lastError = "enableSleep=1 (islands/sleep, src/Island.cs) is not supported in this build; set enableSleep=0"; return B2C_E_UNSUPPORTED; }
  B2CBatch* h = new (std::nothrow) B2CBatch();
  if (!h) return B2C_E_INVALID;
  h->b = new (std::nothrow) Batch(*def, device, worldCount);
  if (!h->b) { delete h; return B2C_E_INVALID; }
  int rc = h->b->init();
  if (rc != B2C_OK) { g_lastError = h->b->error; delete h->b; delete h; return rc; }
  h->handles.resize(worldCount);
  for (int i = 0; i < worldCount; i++) h->handles[i].w = h->b->world(i);
  *out = h;
  return B2C_OK;
}
int32_t b2cuda_batch_destroy(B2CBatch* h) { if (!h) return B2C_E_INVALID; delete h->b; delete h; return B2C_OK; }
int32_t b2cuda_batch_world(B2CBatch* h, int32_t index, B2CWorld** world) {
  CHECK_BATCH(h); if (!world) return B2C_E_INVALID;
  if (index < 0 || index >= (int)h->handles.size()) { g_lastError = "world index out of range"; *world = nullptr; return B2C_E_INVALID; }
  *world = &h->handles[index];
  return B2C_OK;
}
int32_t b2cuda_batch_step(B2CBatch* h, float dt, int32_t sub) { CHECK_BATCH(h); return doneB(h, h->b->step(dt, sub)); }
int32_t b2cuda_batch_synchronize(B2CBatch* h) { CHECK_BATCH(h); return doneB(h, h->b->synchronize()); }
int32_t b2cuda_batch_reset_world(B2CBatch* h, int32_t index) { CHECK_BATCH(h); return doneB(h, h->b->resetWorld(index)); }
int32_t b2cuda_batch_stats(B2CBatch* h, B2CBatchStats* out, double* deviceDst) { CHECK_BATCH(h); return doneB(h, h->b->stats(out, deviceDst)); }
int32_t b2cuda_batch_timer_start(B2CBatch* h) { CHECK_BATCH(h); return doneB(h, h->b->timerStart()); }
int32_t b2cuda_batch_timer_stop(B2CBatch* h, float* ms) { CHECK_BATCH(h); if (!ms) return B2C_E_INVALID; return doneB(h, h->b->timerStop(ms)); }
}  // extern "C"
