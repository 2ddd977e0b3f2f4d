"""Known-answer tests whose expected numbers are derived BY HAND (or by a few lines of numpy) from the C# text of the reference — not from
oracle/ — and checked against BOTH implementations: the CPU oracle (plain run) and the CUDA library (-m gpu). They are what keeps the two
transcriptions honest: each manifold pair type in a configuration whose manifold can be written down from src/Manifold.cs, the integrator
formulas of src/Solver.cs:121-160 replayed in numpy fp32, sensor / hit / touch events in configurations with an obvious answer.

A manifold read back after the FIRST step is the one Collide computed from the initial poses (api/World.cs:31-113: pairs, then Collide, then
Solve), so the expected values below are functions of the scene definition only. Gravity is zero and nothing moves unless the test says so."""
import numpy as np
import pytest
from box2d_b200 import api, scenes, geometry as G

f32 = np.float32


def _backend(name):
    if name == "cuda":
        return api.CudaBackend()
    from oracle_backend import OracleBackend
    return OracleBackend()


BACKENDS = [pytest.param("oracle", id="oracle"), pytest.param("cuda", id="cuda", marks=pytest.mark.gpu)]


def _world(backend, gravity=(0.0, 0.0), **kw):
    wd = scenes.world_def(**kw)
    wd.gravity = api.Vec2(*gravity)
    return api.World(_backend(backend), wd)


def _body(w, x, y, kind=api.DYNAMIC_BODY, **kw):
    bd = api.default_body_def()
    bd.type = kind
    bd.position = api.Vec2(x, y)
    for k, v in kw.items():
        setattr(bd, k, v)
    return w.CreateBody(bd)


def _sd(**kw):
    sd = api.default_shape_def()
    sd.enableContactEvents = 1
    for k, v in kw.items():
        setattr(sd, k, v)
    return sd


def _manifold_of(w, shape_a, shape_b):
    """The one contact between two shapes after the first step, oriented so that the normal points from shape_a to shape_b."""
    w.Step(1.0 / 60.0, 4)
    cs = [c for c in w.ReadContacts() if {int(c["shapeIdA"]), int(c["shapeIdB"])} == {shape_a, shape_b}]
    assert len(cs) == 1, f"expected one contact, got {len(cs)}"
    c = cs[0]
    sign = 1.0 if int(c["shapeIdA"]) == shape_a else -1.0
    n = np.array([c["normal"]["x"], c["normal"]["y"]], dtype=np.float64) * sign
    pts = [(float(p["point"]["x"]), float(p["point"]["y"]), float(p["separation"]), int(p["id"])) for p in c["points"][:int(c["pointCount"])]]
    return int(c["pointCount"]), n, pts


def _close(a, b, tol=2e-6):
    return np.allclose(np.asarray(a, dtype=np.float64), np.asarray(b, dtype=np.float64), atol=tol, rtol=0)


# ---------------------------------------------------------------- manifolds, one per pair type (src/Manifold.cs)

@pytest.mark.parametrize("backend", BACKENDS)
def test_kat_circle_circle_overlap(backend):
    """CollideCircles :24-50. Centres 0.9 apart, radii 0.5: normal (1,0), separation 0.9 - 1.0, point midway between the two surface points
    cA = 0.5, cB = 0.4 -> x = 0.45, id 0."""
    w = _world(backend)
    try:
        a, b = _body(w, 0.0, 0.0), _body(w, 0.9, 0.0)
        sa, sb = w.CreateCircleShape(a, _sd(), G.Circle((0, 0), 0.5)), w.CreateCircleShape(b, _sd(), G.Circle((0, 0), 0.5))
        count, n, pts = _manifold_of(w, sa, sb)
        assert count == 1 and _close(n, (1, 0)) and _close(pts[0][:3], (0.45, 0.0, -0.1)) and pts[0][3] == 0
    finally:
        w.Destroy()


@pytest.mark.parametrize("backend", BACKENDS)
def test_kat_circle_circle_speculative_margin(backend):
    """Same function: a gap of 0.01 is inside the speculative distance (0.02) and yields a point with separation +0.01; a gap of 0.05 yields
    an empty manifold although the fat boxes still overlap and the contact exists."""
    for gap, expect in ((0.01, 1), (0.05, 0)):
        w = _world(backend)
        try:
            a, b = _body(w, 0.0, 0.0), _body(w, 1.0 + gap, 0.0)
            sa, sb = w.CreateCircleShape(a, _sd(), G.Circle((0, 0), 0.5)), w.CreateCircleShape(b, _sd(), G.Circle((0, 0), 0.5))
            count, n, pts = _manifold_of(w, sa, sb)
            assert count == expect
            if expect:
                assert _close(n, (1, 0)) and abs(pts[0][2] - gap) < 2e-6
            assert w.GetCounters().touchingContactCount == expect and len(w.GetContactEvents()[0]) == expect
        finally:
            w.Destroy()


@pytest.mark.parametrize("backend", BACKENDS)
def test_kat_capsule_circle_side_and_cap(backend):
    """CollideCapsuleAndCircle :52-87. Capsule (-1,0)-(1,0) r 0.25. Circle r 0.25 at (0.3, 0.45): closest point (0.3, 0), distance 0.45,
    separation -0.05, normal (0,1), anchor midway between (0.3,0.25) and (0.3,0.2). Circle at (1.3, 0): end cap, closest (1,0), distance
    0.3, separation -0.2, normal (1,0), anchor midway between (1.25,0) and (1.05,0). The port writes the world point of this one function
    as `xfA.p - anchorA` (:80; every other manifold function adds), so with body A at the origin the reported point is the NEGATED anchor."""
    for pos, en, ep in (((0.3, 0.45), (0, 1), (-0.3, -0.225, -0.05)), ((1.3, 0.0), (1, 0), (-1.15, 0.0, -0.2))):
        w = _world(backend)
        try:
            a, b = _body(w, 0.0, 0.0), _body(w, *pos)
            sa = w.CreateCapsuleShape(a, _sd(), G.Capsule((-1.0, 0.0), (1.0, 0.0), 0.25))
            sb = w.CreateCircleShape(b, _sd(), G.Circle((0, 0), 0.25))
            count, n, pts = _manifold_of(w, sa, sb)
            assert count == 1 and _close(n, en) and _close(pts[0][:3], ep)
        finally:
            w.Destroy()


@pytest.mark.parametrize("backend", BACKENDS)
def test_kat_capsules_parallel_two_points(backend):
    """CollideCapsules :89-234, the parallel branch: A (-1,0)-(1,0), B (-0.5,0.45)-(0.5,0.45), radii 0.25. B's segment projects inside A's, so
    the clipped points are B's end points: two manifold points at x = -0.5 and +0.5, separation 0.45 - 0.5, normal (0,1); A's axis is the
    reference face (separationA + 0.1*slop >= separationB, :240), whose branch numbers the points B2_MAKE_ID(0,0) = 0 and B2_MAKE_ID(0,1) = 1
    (:254,:257); each point lies midway between the two surfaces: y = 0.225."""
    w = _world(backend)
    try:
        a, b = _body(w, 0.0, 0.0), _body(w, 0.0, 0.45)
        sa = w.CreateCapsuleShape(a, _sd(), G.Capsule((-1.0, 0.0), (1.0, 0.0), 0.25))
        sb = w.CreateCapsuleShape(b, _sd(), G.Capsule((-0.5, 0.0), (0.5, 0.0), 0.25))
        count, n, pts = _manifold_of(w, sa, sb)
        assert count == 2 and _close(n, (0, 1))
        assert sorted(round(p[0], 5) for p in pts) == [-0.5, 0.5] and all(abs(p[1] - 0.225) < 2e-6 and abs(p[2] + 0.05) < 2e-6 for p in pts)
        assert sorted(p[3] for p in pts) == [0, 1]
    finally:
        w.Destroy()


@pytest.mark.parametrize("backend", BACKENDS)
def test_kat_capsules_crossed_one_point(backend):
    """CollideCapsules, the general branch: A (-1,0)-(1,0), B (0,0.4)-(0,1.4), radii 0.25. Closest points (0,0) and (0,0.4): one point,
    normal (0,1), separation 0.4 - 0.5, midway between (0,0.25) and (0,0.15)."""
    w = _world(backend)
    try:
        a, b = _body(w, 0.0, 0.0), _body(w, 0.0, 0.0)
        sa = w.CreateCapsuleShape(a, _sd(), G.Capsule((-1.0, 0.0), (1.0, 0.0), 0.25))
        sb = w.CreateCapsuleShape(b, _sd(), G.Capsule((0.0, 0.4), (0.0, 1.4), 0.25))
        count, n, pts = _manifold_of(w, sa, sb)
        assert count == 1 and _close(n, (0, 1)) and _close(pts[0][:3], (0.0, 0.2, -0.1))
    finally:
        w.Destroy()


@pytest.mark.parametrize("backend", BACKENDS)
def test_kat_box_circle_face_and_vertex(backend):
    """CollidePolygonAndCircle :258-327. Unit box at the origin. Circle r 0.5 at (0, 0.9): face region, separation 0.9 - 0.5 - 0.5, normal
    (0,1), point midway between (0,0.5) and (0,0.4). Circle at (0.8, 0.8): vertex region, closest vertex (0.5,0.5), distance 0.3*sqrt(2),
    normal (1,1)/sqrt(2), separation distance - 0.5."""
    d = 0.3 * np.sqrt(2.0)
    for pos, en, sep in (((0.0, 0.9), (0, 1), -0.1), ((0.8, 0.8), (np.sqrt(0.5), np.sqrt(0.5)), d - 0.5)):
        w = _world(backend)
        try:
            a, b = _body(w, 0.0, 0.0), _body(w, *pos)
            sa = w.CreatePolygonShape(a, _sd(), G.MakeBox(0.5, 0.5))
            sb = w.CreateCircleShape(b, _sd(), G.Circle((0, 0), 0.5))
            count, n, pts = _manifold_of(w, sa, sb)
            assert count == 1 and _close(n, en) and abs(pts[0][2] - sep) < 2e-6
            # the point halves the gap between the box surface point and the circle surface point along the normal
            surface = np.array((0.0, 0.5)) if pos[0] == 0.0 else np.array((0.5, 0.5))
            assert _close(pts[0][:2], surface + 0.5 * sep * np.array(en))
        finally:
            w.Destroy()


@pytest.mark.parametrize("backend", BACKENDS)
def test_kat_box_on_box_face_contact_ids_and_points(backend):
    """CollidePolygons :410-562 + ClipPolygons :329-388. Unit boxes at (0,0) and (0,0.99). MakeBox orders its vertices (-h,-h),(h,-h),(h,h),
    (-h,h) with normals (0,-1),(1,0),(0,1),(-1,0) (src/Geometry.cs:88-99). Both separations are -0.01, so flip is false; the reference edge
    is A's edge 2 (normal (0,1)), the incident edge is B's edge 0. Clipping keeps B's vertices 1 and 0: points (0.5, 0.495) with id
    (2 << 8 | 1) = 513 and (-0.5, 0.495) with id (3 << 8 | 0) = 768, both with separation -0.01."""
    w = _world(backend)
    try:
        a, b = _body(w, 0.0, 0.0), _body(w, 0.0, 0.99)
        sa, sb = w.CreatePolygonShape(a, _sd(), G.MakeBox(0.5, 0.5)), w.CreatePolygonShape(b, _sd(), G.MakeBox(0.5, 0.5))
        count, n, pts = _manifold_of(w, sa, sb)
        assert count == 2 and _close(n, (0, 1))
        byid = {p[3]: p for p in pts}
        assert set(byid) == {513, 768}
        assert _close(byid[513][:3], (0.5, 0.495, -0.01)) and _close(byid[768][:3], (-0.5, 0.495, -0.01))
    finally:
        w.Destroy()


@pytest.mark.parametrize("backend", BACKENDS)
def test_kat_box_corner_to_corner_single_point(backend):
    """CollidePolygons' vertex-vertex fallback (:498-545): two unit boxes offset diagonally so that only the corners (0.5,0.5) and
    (0.51,0.51) are near: distance 0.01*sqrt(2) > 0.1*linearSlop, both segment fractions clamp to an end, one point along (1,1)/sqrt(2)
    with that separation, halfway between the corners."""
    w = _world(backend)
    try:
        a, b = _body(w, 0.0, 0.0), _body(w, 1.01, 1.01)
        sa, sb = w.CreatePolygonShape(a, _sd(), G.MakeBox(0.5, 0.5)), w.CreatePolygonShape(b, _sd(), G.MakeBox(0.5, 0.5))
        count, n, pts = _manifold_of(w, sa, sb)
        assert count == 1 and _close(n, (np.sqrt(0.5), np.sqrt(0.5)), 1e-5)
        assert abs(pts[0][2] - 0.01 * np.sqrt(2.0)) < 2e-6 and _close(pts[0][:2], (0.505, 0.505), 1e-5)
    finally:
        w.Destroy()


@pytest.mark.parametrize("backend", BACKENDS)
def test_kat_rounded_box_circle_uses_the_radius(backend):
    """CollidePolygonAndCircle with a rounded polygon (MakeRoundedBox: the polygon radius adds to the circle's): box 0.4 + radius 0.1 and a
    circle r 0.3 at (0, 0.75): separation 0.75 - 0.4 - 0.1 - 0.3."""
    w = _world(backend)
    try:
        a, b = _body(w, 0.0, 0.0), _body(w, 0.0, 0.75)
        sa = w.CreatePolygonShape(a, _sd(), G.MakeRoundedBox(0.4, 0.4, 0.1))
        sb = w.CreateCircleShape(b, _sd(), G.Circle((0, 0), 0.3))
        count, n, pts = _manifold_of(w, sa, sb)
        assert count == 1 and _close(n, (0, 1)) and abs(pts[0][2] + 0.05) < 2e-6
    finally:
        w.Destroy()


@pytest.mark.parametrize("backend", BACKENDS)
def test_kat_box_capsule_lying_on_the_face(backend):
    """CollidePolygonAndCapsule :236-256 turns the capsule into a 2-gon with radius and calls CollidePolygons: a capsule (-0.3,0)-(0.3,0)
    r 0.2 lying 0.69 above the centre of a unit box: two points at x = +-0.3, separation 0.69 - 0.5 - 0.2, normal (0,1)."""
    w = _world(backend)
    try:
        a, b = _body(w, 0.0, 0.0), _body(w, 0.0, 0.69)
        sa = w.CreatePolygonShape(a, _sd(), G.MakeBox(0.5, 0.5))
        sb = w.CreateCapsuleShape(b, _sd(), G.Capsule((-0.3, 0.0), (0.3, 0.0), 0.2))
        count, n, pts = _manifold_of(w, sa, sb)
        assert count == 2 and _close(n, (0, 1))
        assert sorted(round(p[0], 5) for p in pts) == [-0.3, 0.3] and all(abs(p[2] + 0.01) < 2e-6 for p in pts)
    finally:
        w.Destroy()


@pytest.mark.parametrize("backend", BACKENDS)
def test_kat_segment_against_circle_capsule_box(backend):
    """CollideSegmentAndCircle / AndCapsule / AndPolygon :536-568 (the segment becomes a capsule of radius 0). A static segment (-2,0)-(2,0):
    a circle r 0.5 at (0.3, 0.45) -> one point, separation -0.05; a capsule (-0.4,0)-(0.4,0) r 0.2 at height 0.19 -> two points at
    x = +-0.4, separation -0.01; a unit box at height 0.49 -> two points at x = +-0.5, separation -0.01. Normal (0,1) from the segment."""
    cases = (("circle", (0.3, 0.45), 1, -0.05, None), ("capsule", (0.0, 0.19), 2, -0.01, [-0.4, 0.4]), ("box", (0.0, 0.49), 2, -0.01, [-0.5, 0.5]))
    for kind, pos, ecount, esep, xs in cases:
        w = _world(backend)
        try:
            g = _body(w, 0.0, 0.0, kind=api.STATIC_BODY)
            sa = w.CreateSegmentShape(g, _sd(), G.Segment((-2.0, 0.0), (2.0, 0.0)))
            b = _body(w, *pos)
            geom = {"circle": G.Circle((0, 0), 0.5), "capsule": G.Capsule((-0.4, 0.0), (0.4, 0.0), 0.2), "box": G.MakeBox(0.5, 0.5)}[kind]
            sb = w.CreateShape(b, _sd(), geom)
            count, n, pts = _manifold_of(w, sa, sb)
            assert count == ecount and _close(n, (0, 1)) and all(abs(p[2] - esep) < 2e-6 for p in pts), (kind, pts)
            if xs:
                assert sorted(round(abs(p[0]), 5) for p in pts) == sorted(abs(x) for x in xs)
        finally:
            w.Destroy()


# ---------------------------------------------------------------- integrator formulas (src/Solver.cs:121-160, :177-211), replayed in numpy fp32

def _integrate(v, steps, sub, dt, damping=0.0, accel=0.0):
    h = f32(f32(dt) / f32(sub))
    d = f32(f32(1.0) / f32(f32(1.0) + f32(h * f32(damping))))
    x = f32(0.0)
    for _ in range(steps):
        dx = f32(0.0)
        for _ in range(sub):
            v = f32(f32(h * f32(accel)) + f32(d * v))        # MulAdd(linearVelocityDelta, linearDamping, v)
            dx = f32(dx + f32(h * v))                         # deltaPosition accumulates over the substeps, added to the centre in FinalizeBodies
        x = f32(x + dx)
    return v, x


@pytest.mark.parametrize("backend", BACKENDS)
def test_kat_linear_damping_and_gravity_scale_exact(backend):
    """v <- h*gravityScale*g + v / (1 + h*damping) per substep; the position gains the substep sum at the end of the step. Bit-exact."""
    w = _world(backend, gravity=(0.0, -10.0))
    try:
        a = _body(w, 0.0, 0.0, linearVelocity=api.Vec2(3.0, 0.0), linearDamping=0.5, gravityScale=0.0)
        b = _body(w, 5.0, 0.0, gravityScale=0.25)
        w.CreateCircleShape(a, _sd(), G.Circle((0, 0), 0.2))
        w.CreateCircleShape(b, _sd(), G.Circle((0, 0), 0.2))
        for _ in range(7):
            w.Step(1.0 / 60.0, 4)
        r = w.ReadBodies()
        va, xa = _integrate(f32(3.0), 7, 4, 1.0 / 60.0, damping=0.5)
        vb, yb = _integrate(f32(0.0), 7, 4, 1.0 / 60.0, accel=f32(0.25) * f32(-10.0))
        assert f32(r[a]["linearVelocity"]["x"]) == va and f32(r[a]["transform"]["p"]["x"]) == xa
        assert f32(r[b]["linearVelocity"]["y"]) == vb and f32(r[b]["transform"]["p"]["y"]) == yb
    finally:
        w.Destroy()


@pytest.mark.parametrize("backend", BACKENDS)
def test_kat_speed_caps(backend):
    """|v| above maximumLinearSpeed (4 m/s here) is scaled to it every substep (ratio = max / |v|), |w|*dt above MaxRotation = pi/4 likewise
    (src/Solver.cs:143-154)."""
    wd = scenes.world_def()
    wd.gravity = api.Vec2(0.0, 0.0)
    wd.maximumLinearSpeed = 4.0
    w = api.World(_backend(backend), wd)
    try:
        a = _body(w, 0.0, 0.0, linearVelocity=api.Vec2(30.0, 40.0), angularVelocity=100.0)
        w.CreateCircleShape(a, _sd(), G.Circle((0, 0), 0.2))
        w.Step(1.0 / 60.0, 4)
        r = w.ReadBodies()[a]
        speed = np.hypot(float(r["linearVelocity"]["x"]), float(r["linearVelocity"]["y"]))
        assert abs(speed - 4.0) < 1e-5 and abs(float(r["linearVelocity"]["y"]) / float(r["linearVelocity"]["x"]) - 4.0 / 3.0) < 1e-5
        assert abs(float(r["angularVelocity"]) - 0.25 * np.pi * 60.0) < 1e-3
    finally:
        w.Destroy()


@pytest.mark.parametrize("backend", BACKENDS)
def test_kat_impulse_force_and_mass_of_a_box(backend):
    """A 1 x 1 box of density 1 has mass 1 and inertia (1/12)(1+1) = 1/6 (Polygon.ComputeMass). ApplyLinearImpulseToCenter adds
    invMass*impulse, ApplyAngularImpulse adds invInertia*impulse = 6*impulse (api/Body.cs:507-554); a force of 12 N held for one step
    adds h*invMass*F per substep = 0.2 m/s and is cleared by FinalizeBodies."""
    w = _world(backend)
    try:
        a = _body(w, 0.0, 0.0)
        w.CreatePolygonShape(a, _sd(), G.MakeBox(0.5, 0.5))
        w.ApplyLinearImpulseToCenter(a, (2.0, -1.0), True)
        w.ApplyAngularImpulse(a, 0.5, True)
        r = w.ReadBodies()[a]
        assert _close((r["linearVelocity"]["x"], r["linearVelocity"]["y"]), (2.0, -1.0), 1e-6) and abs(float(r["angularVelocity"]) - 3.0) < 1e-5
        w.SetLinearVelocity(a, (0.0, 0.0))
        w.SetAngularVelocity(a, 0.0)
        w.ApplyForceToCenter(a, (12.0, 0.0), True)
        w.Step(1.0 / 60.0, 4)
        assert abs(float(w.ReadBodies()[a]["linearVelocity"]["x"]) - 0.2) < 1e-6
        w.Step(1.0 / 60.0, 4)
        assert abs(float(w.ReadBodies()[a]["linearVelocity"]["x"]) - 0.2) < 1e-6            # no force left
    finally:
        w.Destroy()


@pytest.mark.parametrize("backend", BACKENDS)
def test_kat_kinematic_body_and_motion_locks(backend):
    """A kinematic body keeps its velocity (invMass = 0: no gravity term) and advances by the substep sum; a dynamic body with lockLinearX
    and lockAngularZ keeps x and rotation while it falls (src/Solver.cs:155-157, :412-414)."""
    w = _world(backend, gravity=(0.0, -10.0))
    try:
        k = _body(w, 0.0, 5.0, kind=api.KINEMATIC_BODY, linearVelocity=api.Vec2(1.5, 0.0))
        d = _body(w, 4.0, 5.0, linearVelocity=api.Vec2(2.0, 0.0), angularVelocity=1.0, lockLinearX=1, lockAngularZ=1)
        w.CreatePolygonShape(k, _sd(), G.MakeBox(0.5, 0.1))
        w.CreateCircleShape(d, _sd(), G.Circle((0, 0), 0.2))
        for _ in range(5):
            w.Step(1.0 / 60.0, 4)
        r = w.ReadBodies()
        _, xk = _integrate(f32(1.5), 5, 4, 1.0 / 60.0)
        assert f32(r[k]["transform"]["p"]["x"]) == xk and f32(r[k]["transform"]["p"]["y"]) == f32(5.0) and f32(r[k]["linearVelocity"]["x"]) == f32(1.5)
        vy, dy = _integrate(f32(0.0), 5, 4, 1.0 / 60.0, accel=-10.0)
        assert f32(r[d]["transform"]["p"]["x"]) == f32(4.0) and f32(r[d]["linearVelocity"]["x"]) == f32(0.0) and f32(r[d]["angularVelocity"]) == f32(0.0)
        assert f32(r[d]["linearVelocity"]["y"]) == vy and f32(r[d]["transform"]["q"]["s"]) == f32(0.0)
    finally:
        w.Destroy()


@pytest.mark.parametrize("backend", BACKENDS)
def test_kat_friction_decelerates_at_mu_g(backend):
    """A box sliding on the ground at 3 m/s with friction sqrt(0.6*0.6) = 0.6 (geometric mean, include/Types.cs MixFriction) loses
    mu*g*t: after 0.25 s it moves at 3 - 0.6*10*0.25 = 1.5 m/s (soft contact and sub-stepping: within 3 %)."""
    w = _world(backend, gravity=(0.0, -10.0))
    try:
        g = _body(w, 0.0, 0.0, kind=api.STATIC_BODY)
        w.CreatePolygonShape(g, _sd(), G.MakeOffsetBox(50.0, 1.0, (0.0, -1.0)))
        b = _body(w, 0.0, 0.5, linearVelocity=api.Vec2(3.0, 0.0))
        w.CreatePolygonShape(b, _sd(), G.MakeBox(0.5, 0.5))
        for _ in range(15):
            w.Step(1.0 / 60.0, 4)
        assert abs(float(w.ReadBodies()[b]["linearVelocity"]["x"]) - 1.5) < 0.05
    finally:
        w.Destroy()


@pytest.mark.parametrize("backend", BACKENDS)
def test_kat_restitution_is_swallowed_by_the_port_and_hit_event(backend):
    """A ball with restitution 1 thrown at the ground at 5 m/s. Upstream Box2D would send it back at 5 m/s; THIS port does not: its AVX
    ApplyRestitution multiplies the correction by the comparison MASK instead of the blended mass (src/ContactSolver.cs:931,934 -> quirk Q1):
    a lane that qualifies has mask = +0.0, so its restitution impulse is exactly zero and the ball simply stops (never moves upwards).
    The contact still reports a hit event with the approach speed of 5 m/s (> hitEventThreshold = 1) and the ground's normal."""
    w = _world(backend, gravity=(0.0, 0.0))
    try:
        g = _body(w, 0.0, 0.0, kind=api.STATIC_BODY)
        w.CreatePolygonShape(g, _sd(enableHitEvents=1, restitution=1.0), G.MakeOffsetBox(10.0, 1.0, (0.0, -1.0)))
        b = _body(w, 0.0, 0.6, linearVelocity=api.Vec2(0.0, -5.0))
        w.CreateCircleShape(b, _sd(enableHitEvents=1, restitution=1.0), G.Circle((0, 0), 0.25))
        hits, vy = [], []
        for _ in range(12):
            w.Step(1.0 / 60.0, 4)
            hits.extend(w.GetContactEvents()[2].tolist())
            vy.append(float(w.ReadBodies()[b]["linearVelocity"]["y"]))
        assert vy[0] == -5.0 and max(vy) < 1e-6 and abs(vy[-1]) < 1e-6, vy
        assert len(hits) == 1 and abs(hits[0][4] - 5.0) < 0.25 and _close(np.abs(hits[0][3]), (0.0, 1.0), 1e-5), hits
    finally:
        w.Destroy()


@pytest.mark.parametrize("backend", BACKENDS)
def test_kat_touch_events_of_a_passing_circle(backend):
    """A circle flying sideways over a static circle (no gravity, grazing overlap): one begin event when the manifold first has a point, one
    end event when it has none again, both naming the same contact id; the end event is readable after the step that ended it."""
    w = _world(backend)
    try:
        g = _body(w, 0.0, 0.0, kind=api.STATIC_BODY)
        sg = w.CreateCircleShape(g, _sd(), G.Circle((0, 0), 0.5))
        b = _body(w, -2.0, 1.01, linearVelocity=api.Vec2(6.0, 0.0))
        sb = w.CreateCircleShape(b, _sd(), G.Circle((0, 0), 0.5))
        begins, ends = [], []
        for step in range(45):
            w.Step(1.0 / 60.0, 4)
            bgn, end, _ = w.GetContactEvents()
            begins += [(step,) + tuple(e)[:3] for e in bgn.tolist()]
            ends += [(step,) + tuple(e)[:3] for e in end.tolist()]
        assert len(begins) == 1 and len(ends) == 1 and begins[0][0] < ends[0][0], (begins, ends)
        assert {begins[0][1], begins[0][2]} == {sg, sb} and begins[0][3] == ends[0][3]
    finally:
        w.Destroy()


# ---------------------------------------------------------------- sensors (src/Sensor.cs)

@pytest.mark.parametrize("backend", BACKENDS)
def test_kat_sensor_overlap_threshold_and_events(backend):
    """A static circular sensor r 1 at the origin. Visitors r 0.25 at distance 1.2 (inside: 1.2 < 1.25) and 1.3 (outside), a visitor inside
    with enableSensorEvents off, and one inside but filtered out (category 2 against mask 1): exactly one begin event at the first step, the
    overlap list holds that one shape, and moving it away produces exactly one end event."""
    w = _world(backend)
    try:
        z = _body(w, 0.0, 0.0, kind=api.STATIC_BODY)
        s = w.CreateCircleShape(z, _sd(isSensor=1, enableSensorEvents=1, maskBits=1), G.Circle((0, 0), 1.0))
        ids = []
        for pos, ev, cat in (((1.2, 0.0), 1, 1), ((0.0, 1.3), 1, 1), ((-1.0, 0.0), 0, 1), ((0.0, -1.0), 1, 2)):
            b = _body(w, *pos)
            ids.append((b, w.CreateCircleShape(b, _sd(enableSensorEvents=ev, categoryBits=cat), G.Circle((0, 0), 0.25))))
        w.Step(1.0 / 60.0, 4)
        bgn, end = w.GetSensorEvents()
        assert [(int(e["sensor"]["shapeId"]), int(e["visitor"]["shapeId"])) for e in bgn] == [(s, ids[0][1])] and len(end) == 0
        assert [v[0] for v in w.GetSensorOverlaps(s)] == [ids[0][1]]
        w.Step(1.0 / 60.0, 4)
        assert len(w.GetSensorEvents()[0]) == 0
        w.SetTransform(ids[0][0], (3.0, 0.0), (1.0, 0.0))
        w.Step(1.0 / 60.0, 4)
        bgn, end = w.GetSensorEvents()
        assert len(bgn) == 0 and [(int(e["sensor"]["shapeId"]), int(e["visitor"]["shapeId"])) for e in end] == [(s, ids[0][1])]
    finally:
        w.Destroy()


# ---------------------------------------------------------------- joints: geometric invariants (src/joints/*.cs)

@pytest.mark.parametrize("backend", BACKENDS)
def test_kat_revolute_pendulum_keeps_its_length_and_energy(backend):
    """A point mass on a revolute joint 2 m from a static pivot, released horizontally: the distance to the pivot stays 2 m (soft
    constraint: within 2 mm) and at the bottom the speed is sqrt(2 g L) = 6.32 m/s within 3 %."""
    w = _world(backend, gravity=(0.0, -10.0))
    try:
        pivot = _body(w, 0.0, 10.0, kind=api.STATIC_BODY)
        bob = _body(w, 2.0, 10.0)
        w.CreateCircleShape(bob, _sd(), G.Circle((0, 0), 0.1))
        jd = api.default_revolute_joint_def()
        jd.bodyIdA, jd.bodyIdB = pivot, bob
        jd.localFrameB = api.Transform(api.Vec2(-2.0, 0.0), api.Rot(1.0, 0.0))
        w.CreateRevoluteJoint(jd)
        best = 0.0
        for _ in range(60):
            w.Step(1.0 / 60.0, 4)
            r = w.ReadBodies()[bob]
            x, y = float(r["transform"]["p"]["x"]), float(r["transform"]["p"]["y"]) - 10.0
            assert abs(np.hypot(x, y) - 2.0) < 2e-3
            best = max(best, np.hypot(float(r["linearVelocity"]["x"]), float(r["linearVelocity"]["y"])))
        assert abs(best - np.sqrt(2 * 10.0 * 2.0)) < 0.03 * 6.32
    finally:
        w.Destroy()


@pytest.mark.parametrize("backend", BACKENDS)
def test_kat_joint_force_threshold_emits_joint_event(backend):
    """A revolute joint carrying a 1 kg box reports no joint event while forceThreshold is FLT_MAX and reports one every step once the
    threshold (1 N) is below the 10 N it carries (GetJointReaction vs forceThreshold, src/Solver.cs:1125-1150)."""
    w = _world(backend, gravity=(0.0, -10.0))
    try:
        pivot = _body(w, 0.0, 10.0, kind=api.STATIC_BODY)
        box = _body(w, 0.0, 10.0)
        w.CreatePolygonShape(box, _sd(), G.MakeBox(0.5, 0.5))
        jd = api.default_revolute_joint_def()
        jd.bodyIdA, jd.bodyIdB = pivot, box
        j = w.CreateRevoluteJoint(jd)
        for _ in range(5):
            w.Step(1.0 / 60.0, 4)
            assert list(w.GetJointEvents()) == []
        from box2d_b200 import capi
        w.JointSetParam(j, capi.JP_FORCE_THRESHOLD, 1.0)
        for _ in range(3):
            w.Step(1.0 / 60.0, 4)
            assert list(w.GetJointEvents()) == [j]
    finally:
        w.Destroy()


@pytest.mark.parametrize("backend", BACKENDS)
def test_kat_zero_dt_step_only_trades_the_end_buffers(backend):
    """api/World.cs:44-51: Step(0) clears the begin events, flips endEventArrayIndex and returns; nothing moves."""
    w = _world(backend, gravity=(0.0, -10.0))
    try:
        b = _body(w, 0.0, 1.0, linearVelocity=api.Vec2(1.0, 0.0))
        w.CreateCircleShape(b, _sd(), G.Circle((0, 0), 0.2))
        w.Step(1.0 / 60.0, 4)
        before = w.ReadBodies()[b].tobytes()
        w.Step(0.0, 4)
        assert w.ReadBodies()[b].tobytes() == before and all(len(e) == 0 for e in w.GetContactEvents())
    finally:
        w.Destroy()
