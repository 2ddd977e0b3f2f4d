// ORACLE — TEST INFRASTRUCTURE ONLY (see b2o_math.h header). Parity unpinned by the reference.
// Part 2: contact lifecycle, broadphase pair update, collide phase, constraint-graph coloring.
#include "b2o_world.h"

namespace b2o {

static inline int ProxyType(int key) { return key & 3; }
static inline int ProxyId(int key) { return key >> 2; }
static inline int ProxyKey(int id, int type) { return (id << 2) | type; }
static inline uint64_t ShapePairKey(uint64_t k1, uint64_t k2) { return k1 < k2 ? (k1 << 32 | k2) : (k2 << 32 | k1); }

template <class T> static int RemoveSwap(std::vector<T>& a, int index) {   // include/Box2D.cs:9-19
  int movedIndex = -1;
  if (index != (int)a.size() - 1) { movedIndex = (int)a.size() - 1; a[index] = a[movedIndex]; }
  a.pop_back();
  return movedIndex;
}

static bool ShouldShapesCollide(const Filter& a, const Filter& b) {       // src/Shape.cs:56-61
  if (a.groupIndex == b.groupIndex && a.groupIndex != 0) return a.groupIndex > 0;
  return (a.maskBits & b.categoryBits) != 0 && (a.categoryBits & b.maskBits) != 0;
}

void World::CreateContact(Shape& shapeA, Shape& shapeB) {      // src/Contact.cs:134-227
  int type1 = shapeA.geom.type, type2 = shapeB.geom.type;
  if (!PairSupported(type1, type2)) return;
  if (!PairPrimary(type1, type2)) { CreateContact(shapeB, shapeA); return; }
  Body& bodyA = bodies[shapeA.bodyId]; Body& bodyB = bodies[shapeB.bodyId];
  int contactId = contactIdPool.AllocId();
  if (contactId == (int)contacts.size()) contacts.emplace_back();
  int shapeIdA = shapeA.id, shapeIdB = shapeB.id;
  Contact& contact = contacts[contactId];
  contact.contactId = contactId; contact.generation++;
  contact.setIndex = kAwakeSet; contact.colorIndex = -1; contact.localIndex = (int)awakeSet.contactSims.size();
  contact.shapeIdA = shapeIdA; contact.shapeIdB = shapeIdB; contact.flags = 0;
  if (shapeA.enableContactEvents || shapeB.enableContactEvents) contact.flags |= kContactEnableContactEvents;
  {
    contact.edge0.bodyId = shapeA.bodyId; contact.edge0.prevKey = -1; contact.edge0.nextKey = bodyA.headContactKey;
    int keyA = contactId << 1, head = bodyA.headContactKey;
    if (head != -1) { Contact& hc = contacts[head >> 1]; if (head & 1) hc.edge1.prevKey = keyA; else hc.edge0.prevKey = keyA; }
    bodyA.headContactKey = keyA; bodyA.contactCount++;
  }
  {
    contact.edge1.bodyId = shapeB.bodyId; contact.edge1.prevKey = -1; contact.edge1.nextKey = bodyB.headContactKey;
    int keyB = (contactId << 1) | 1, head = bodyB.headContactKey;
    if (head != -1) { Contact& hc = contacts[head >> 1]; if (head & 1) hc.edge1.prevKey = keyB; else hc.edge0.prevKey = keyB; }
    bodyB.headContactKey = keyB; bodyB.contactCount++;
  }
  uint64_t pairKey = ShapePairKey((uint64_t)shapeIdA, (uint64_t)shapeIdB);
  pairSet.insert(pairKey);
  lastNewPairKeys.push_back(pairKey);
  ContactSim sim;
  sim.contactId = contactId; sim.shapeIdA = shapeIdA; sim.shapeIdB = shapeIdB;
  sim.friction = sqrtf(shapeA.material.friction * shapeB.material.friction);            // src/World.cs:139
  sim.restitution = netMax(shapeA.material.restitution, shapeB.material.restitution);   // src/World.cs:140
  sim.tangentSpeed = 0; sim.simFlags = 0;
  if (shapeA.enablePreSolveEvents || shapeB.enablePreSolveEvents) sim.simFlags |= kSimEnablePreSolveEvents;
  awakeSet.contactSims.push_back(sim);
}

void World::DestroyContact(Contact& contact) {      // src/Contact.cs:228-312 (wakeBodies irrelevant: no sleeping sets)
  uint64_t pairKey = ShapePairKey((uint64_t)contact.shapeIdA, (uint64_t)contact.shapeIdB);
  pairSet.erase(pairKey);
  int bodyIdA = contact.edge0.bodyId, bodyIdB = contact.edge1.bodyId;
  Body& bodyA = bodies[bodyIdA]; Body& bodyB = bodies[bodyIdB];
  bool touching = (contact.flags & kContactTouching) != 0;
  if (touching && (contact.flags & kContactEnableContactEvents))
    contactEndEvents[endEventArrayIndex].push_back({contact.shapeIdA, contact.shapeIdB, contact.contactId, contact.generation});
  auto fixPrevNext = [&](ContactEdge& e) {
    if (e.prevKey != -1) { Contact& p = contacts[e.prevKey >> 1]; if (e.prevKey & 1) p.edge1.nextKey = e.nextKey; else p.edge0.nextKey = e.nextKey; }
    if (e.nextKey != -1) { Contact& n = contacts[e.nextKey >> 1]; if (e.nextKey & 1) n.edge1.prevKey = e.prevKey; else n.edge0.prevKey = e.prevKey; }
  };
  int contactId = contact.contactId;
  fixPrevNext(contact.edge0);
  if (bodyA.headContactKey == (contactId << 1)) bodyA.headContactKey = contact.edge0.nextKey;
  bodyA.contactCount--;
  fixPrevNext(contact.edge1);
  if (bodyB.headContactKey == ((contactId << 1) | 1)) bodyB.headContactKey = contact.edge1.nextKey;
  bodyB.contactCount--;
  if (contact.colorIndex != -1) RemoveContactFromGraph(bodyIdA, bodyIdB, contact.colorIndex, contact.localIndex);
  else {
    int movedIndex = RemoveSwap(awakeSet.contactSims, contact.localIndex);
    if (movedIndex != -1) contacts[awakeSet.contactSims[contact.localIndex].contactId].localIndex = contact.localIndex;
  }
  contact.contactId = -1; contact.setIndex = -1; contact.colorIndex = -1; contact.localIndex = -1;
  contactIdPool.FreeId(contactId);
}

// src/Contact.cs:324-417 (default friction/restitution mixers, no preSolve callback)
bool World::UpdateContact(ContactSim& contactSim, const Shape& shapeA, const Transform& transformA, Vec2 centerOffsetA,
                          const Shape& shapeB, const Transform& transformB, Vec2 centerOffsetB) {
  Manifold oldManifold = contactSim.manifold;
  contactSim.manifold = ComputeManifold(shapeA.geom, transformA, shapeB.geom, transformB);
  contactSim.friction = sqrtf(shapeA.material.friction * shapeB.material.friction);
  contactSim.restitution = netMax(shapeA.material.restitution, shapeB.material.restitution);
  if (shapeA.material.rollingResistance > 0 || shapeB.material.rollingResistance > 0)
    contactSim.rollingResistance = netMax(shapeA.material.rollingResistance, shapeB.material.rollingResistance) *
                                   netMax(GeomRadius(shapeA.geom), GeomRadius(shapeB.geom));
  else contactSim.rollingResistance = 0;
  contactSim.tangentSpeed = shapeA.material.tangentSpeed + shapeB.material.tangentSpeed;
  int pointCount = contactSim.manifold.pointCount;
  bool touching = pointCount > 0;
  if (!enableSpeculative && pointCount == 2) {      // Q7: tests point0 twice
    if (contactSim.manifold.point0.separation > 1.5f * kLinearSlop) { contactSim.manifold.point0 = contactSim.manifold.point1; contactSim.manifold.pointCount = 1; }
    else if (contactSim.manifold.point0.separation > 1.5f * kLinearSlop) contactSim.manifold.pointCount = 1;
    pointCount = contactSim.manifold.pointCount;
  }
  if (touching && (shapeA.enableHitEvents || shapeB.enableHitEvents)) contactSim.simFlags |= kSimEnableHitEvent;
  else contactSim.simFlags &= ~kSimEnableHitEvent;
  if (pointCount > 0) contactSim.manifold.rollingImpulse = oldManifold.rollingImpulse;
  for (int i = 0; i < pointCount; i++) {
    ManifoldPoint& mp2 = i == 1 ? contactSim.manifold.point1 : contactSim.manifold.point0;
    mp2.anchorA = mp2.anchorA - centerOffsetA;
    mp2.anchorB = mp2.anchorB - centerOffsetB;
    mp2.tangentImpulse = 0; mp2.normalImpulse = 0; mp2.totalNormalImpulse = 0; mp2.normalVelocity = 0; mp2.persisted = false;
    uint16_t id2 = mp2.id;
    for (int j = 0; j < oldManifold.pointCount; j++) {
      ManifoldPoint& mp1 = j == 1 ? oldManifold.point1 : oldManifold.point0;
      if (mp1.id == id2) {
        mp2.normalImpulse = mp1.normalImpulse; mp2.tangentImpulse = mp1.tangentImpulse; mp2.persisted = true;
        mp1.normalImpulse = 0; mp1.tangentImpulse = 0;
        break;
      }
    }
  }
  if (touching) contactSim.simFlags |= kSimTouching; else contactSim.simFlags &= ~kSimTouching;
  return touching;
}

void World::AddContactToGraph(const ContactSim& contactSim, Contact& contact) {     // src/ConstraintGraph.cs:135-220
  int colorIndex = kOverflowIndex;
  int bodyIdA = contact.edge0.bodyId, bodyIdB = contact.edge1.bodyId;
  int typeA = bodies[bodyIdA].type, typeB = bodies[bodyIdB].type;
  if (typeA != kStaticBody && typeB != kStaticBody) {
    for (int i = 0; i < kDynamicColorCount; i++) {
      GraphColor& color = colors[i];
      if (color.bodySet.GetBit(bodyIdA) || color.bodySet.GetBit(bodyIdB)) continue;
      if (typeA == kDynamicBody) color.bodySet.SetBitGrow(bodyIdA);
      if (typeB == kDynamicBody) color.bodySet.SetBitGrow(bodyIdB);
      colorIndex = i; break;
    }
  } else if (typeA == kDynamicBody) {
    for (int i = kGraphColorCount - 2; i >= 1; i--) {
      GraphColor& color = colors[i];
      if (color.bodySet.GetBit(bodyIdA)) continue;
      color.bodySet.SetBitGrow(bodyIdA); colorIndex = i; break;
    }
  } else if (typeB == kDynamicBody) {
    for (int i = kGraphColorCount - 2; i >= 1; i--) {
      GraphColor& color = colors[i];
      if (color.bodySet.GetBit(bodyIdB)) continue;
      color.bodySet.SetBitGrow(bodyIdB); colorIndex = i; break;
    }
  }
  GraphColor& color = colors[colorIndex];
  contact.colorIndex = colorIndex; contact.localIndex = (int)color.contactSims.size();
  color.contactSims.push_back(contactSim);
  ContactSim& newContact = color.contactSims.back();
  if (typeA == kStaticBody) { newContact.bodySimIndexA = -1; newContact.invMassA = 0; newContact.invIA = 0; }
  else {
    int localIndex = bodies[bodyIdA].localIndex; newContact.bodySimIndexA = localIndex;
    const BodySim& s = awakeSet.bodySims[localIndex]; newContact.invMassA = s.invMass; newContact.invIA = s.invInertia;
  }
  if (typeB == kStaticBody) { newContact.bodySimIndexB = -1; newContact.invMassB = 0; newContact.invIB = 0; }
  else {
    int localIndex = bodies[bodyIdB].localIndex; newContact.bodySimIndexB = localIndex;
    const BodySim& s = awakeSet.bodySims[localIndex]; newContact.invMassB = s.invMass; newContact.invIB = s.invInertia;
  }
}

void World::RemoveContactFromGraph(int bodyIdA, int bodyIdB, int colorIndex, int localIndex) {   // src/ConstraintGraph.cs:221-241
  GraphColor& color = colors[colorIndex];
  if (colorIndex != kOverflowIndex) { color.bodySet.ClearBit(bodyIdA); color.bodySet.ClearBit(bodyIdB); }
  int movedIndex = RemoveSwap(color.contactSims, localIndex);
  if (movedIndex != -1) contacts[color.contactSims[localIndex].contactId].localIndex = localIndex;
}

void World::UpdateBroadPhasePairs() {      // src/BroadPhase.cs:137-268
  lastNewPairKeys.clear();
  int moveCount = (int)moveArray.size();
  if (moveCount == 0) return;
  std::vector<std::vector<std::pair<int, int>>> moveResults(moveCount);   // per query, in hit order
  for (int i = 0; i < moveCount; i++) {
    int queryProxyKey = moveArray[i];
    if (queryProxyKey == -1) continue;
    int proxyType = ProxyType(queryProxyKey), proxyId = ProxyId(queryProxyKey);
    AABB fatAABB = trees[proxyType].nodes[proxyId].aabb;
    int queryShapeIndex = (int)trees[proxyType].nodes[proxyId].userData;
    auto& out = moveResults[i];
    auto runQuery = [&](int treeType) {
      trees[treeType].Query(fatAABB, ~0ull, [&](int hitProxyId, uint64_t userData) -> bool {   // PairQueryCallback :137-196
        int shapeId = (int)userData;
        int proxyKey = ProxyKey(hitProxyId, treeType);
        if (proxyKey == queryProxyKey) return true;
        if (proxyType == kDynamicBody) {
          if (treeType == kDynamicBody && proxyKey < queryProxyKey) { if (moveSet.count(proxyKey)) return true; }
        } else { if (moveSet.count(proxyKey)) return true; }
        uint64_t pairKey = ShapePairKey((uint64_t)shapeId, (uint64_t)queryShapeIndex);
        if (pairSet.count(pairKey)) return true;
        int shapeIdA, shapeIdB;
        if (proxyKey < queryProxyKey) { shapeIdA = shapeId; shapeIdB = queryShapeIndex; } else { shapeIdA = queryShapeIndex; shapeIdB = shapeId; }
        const Shape& shapeA = shapes[shapeIdA]; const Shape& shapeB = shapes[shapeIdB];
        int bodyIdA = shapeA.bodyId, bodyIdB = shapeB.bodyId;
        if (bodyIdA == bodyIdB) return true;
        if (shapeA.sensorIndex != -1 || shapeB.sensorIndex != -1) return true;
        if (!ShouldShapesCollide(shapeA.filter, shapeB.filter)) return true;
        if (!ShouldBodiesCollide(bodies[bodyIdA], bodies[bodyIdB])) return true;
        out.push_back({shapeIdA, shapeIdB});
        return true;
      });
    };
    if (proxyType == kDynamicBody) { runQuery(kKinematicBody); runQuery(kStaticBody); }   // :216-226
    runQuery(kDynamicBody);                                                            // :227-230
  }
  for (int i = 0; i < moveCount; i++) {          // :251-262 — list is LIFO: reverse hit order
    auto& r = moveResults[i];
    for (int k = (int)r.size() - 1; k >= 0; k--) CreateContact(shapes[r[k].first], shapes[r[k].second]);
  }
  moveArray.clear(); moveSet.clear();
}

void World::Collide(StepContext& ctx) {      // src/World.cs:261-453
  trees[kDynamicBody].Rebuild(false); trees[kKinematicBody].Rebuild(false);   // UpdateTreesTask inline (:311-314,:341)
  int contactCount = 0;
  for (int i = 0; i < kGraphColorCount; i++) contactCount += (int)colors[i].contactSims.size();
  contactCount += (int)awakeSet.contactSims.size();
  if (contactCount == 0) return;
  BitSet bitSet; bitSet.SetBitCountAndClear(contactIdPool.GetIdCapacity());
  auto collideOne = [&](ContactSim& contactSim) {     // CollideTask :261-310
    int contactId = contactSim.contactId;
    const Shape& shapeA = shapes[contactSim.shapeIdA]; const Shape& shapeB = shapes[contactSim.shapeIdB];
    bool overlap = Overlaps(shapeA.fatAABB, shapeB.fatAABB);
    if (!overlap) {
      contactSim.simFlags |= kSimDisjoint; contactSim.simFlags &= ~kSimTouching; bitSet.SetBit(contactId);
    } else {
      bool wasTouching = (contactSim.simFlags & kSimTouching) != 0;
      const Body& bodyA = bodies[shapeA.bodyId]; const Body& bodyB = bodies[shapeB.bodyId];
      const BodySim& bodySimA = GetBodySim(bodyA); const BodySim& bodySimB = GetBodySim(bodyB);
      contactSim.bodySimIndexA = bodyA.setIndex == kAwakeSet ? bodyA.localIndex : -1;
      contactSim.invMassA = bodySimA.invMass; contactSim.invIA = bodySimA.invInertia;
      contactSim.bodySimIndexB = bodyB.setIndex == kAwakeSet ? bodyB.localIndex : -1;
      contactSim.invMassB = bodySimB.invMass; contactSim.invIB = bodySimB.invInertia;
      Transform transformA = bodySimA.transform, transformB = bodySimB.transform;
      Vec2 centerOffsetA = RotateVector(transformA.q, bodySimA.localCenter);
      Vec2 centerOffsetB = RotateVector(transformB.q, bodySimB.localCenter);
      bool touching = UpdateContact(contactSim, shapeA, transformA, centerOffsetA, shapeB, transformB, centerOffsetB);
      if (touching && !wasTouching) { contactSim.simFlags |= kSimStartedTouching; bitSet.SetBit(contactId); }
      else if (!touching && wasTouching) { contactSim.simFlags |= kSimStoppedTouching; bitSet.SetBit(contactId); }
    }
  };
  for (int i = 0; i < kGraphColorCount; i++) for (auto& cs : colors[i].contactSims) collideOne(cs);
  for (auto& cs : awakeSet.contactSims) collideOne(cs);
  for (uint32_t k = 0; k < bitSet.blockCount; k++) {     // :382-450 ascending contactId
    uint64_t bits = bitSet.bits[k];
    while (bits != 0) {
      int ctz = __builtin_ctzll(bits);
      int contactId = (int)(64 * k + ctz);
      Contact& contact = contacts[contactId];
      int colorIndex = contact.colorIndex, localIndex = contact.localIndex;
      ContactSim& contactSim = colorIndex != -1 ? colors[colorIndex].contactSims[localIndex] : awakeSet.contactSims[localIndex];
      uint32_t flags = contact.flags, simFlags = contactSim.simFlags;
      ContactTouchEvent ev = {contact.shapeIdA, contact.shapeIdB, contactId, contact.generation};
      if (simFlags & kSimDisjoint) DestroyContact(contact);
      else if (simFlags & kSimStartedTouching) {
        if (flags & kContactEnableContactEvents) contactBeginEvents.push_back(ev);
        contact.flags |= kContactTouching;
        ContactSim& cs = awakeSet.contactSims[localIndex];
        cs.simFlags &= ~kSimStartedTouching;
        AddContactToGraph(cs, contact);
        int movedIndex = RemoveSwap(awakeSet.contactSims, localIndex);      // RemoveNonTouchingContact :323-336
        if (movedIndex != -1) contacts[awakeSet.contactSims[localIndex].contactId].localIndex = localIndex;
      } else if (simFlags & kSimStoppedTouching) {
        contactSim.simFlags &= ~kSimStoppedTouching;
        contact.flags &= ~kContactTouching;
        if (contact.flags & kContactEnableContactEvents) contactEndEvents[endEventArrayIndex].push_back(ev);
        int bodyIdA = contact.edge0.bodyId, bodyIdB = contact.edge1.bodyId;
        contact.colorIndex = -1; contact.localIndex = (int)awakeSet.contactSims.size();     // AddNonTouchingContact :315-322
        awakeSet.contactSims.push_back(contactSim);
        RemoveContactFromGraph(bodyIdA, bodyIdB, colorIndex, localIndex);
      }
      bits &= bits - 1;
    }
  }
}

}  // namespace b2o
