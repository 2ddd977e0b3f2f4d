"""Scene builders for BASELINE.json's configs, written against the Box2D.API mirror (api.World).

The reference ships no samples (README.md:4); these definitions are ours and are used
identically for the oracle and the CUDA library. All coordinates are exactly representable or pass
through np.float32 so both sides receive bit-identical inputs.
"""
import numpy as np
from . import api, geometry as geo

f32 = np.float32


def splitmix64(state):
    state = (state + 0x9E3779B97F4A7C15) & 0xFFFFFFFFFFFFFFFF
    z = state
    z = ((z ^ (z >> 30)) * 0xBF58476D1CE4E5B9) & 0xFFFFFFFFFFFFFFFF
    z = ((z ^ (z >> 27)) * 0x94D049BB133111EB) & 0xFFFFFFFFFFFFFFFF
    return state, z ^ (z >> 31)


def _uniform(state, lo, hi):
    state, z = splitmix64(state)
    u = (z >> 11) / float(1 << 53)
    return state, float(f32(lo + (hi - lo) * u))


def world_def(enable_sleep=False, enable_continuous=True, **kw):
    d = api.default_world_def()
    d.enableSleep = int(enable_sleep)
    d.enableContinuous = int(enable_continuous)
    for k, v in kw.items():
        setattr(d, k, v)
    return d


def _shape_def(events=True, **kw):
    sd = api.default_shape_def()
    sd.enableContactEvents = int(events)
    for k, v in kw.items():
        setattr(sd, k, v)
    return sd


def _dynamic_body(world, x, y, **kw):
    bd = api.default_body_def()
    bd.type = api.DYNAMIC_BODY
    bd.position = api.Vec2(float(f32(x)), float(f32(y)))
    for k, v in kw.items():
        setattr(bd, k, v)
    return world.CreateBody(bd)


def add_pyramid(world, rows, origin=(0.0, 0.0), ground_half_width=None, jitter_seed=None, jitter=0.0):
    """Static ground box with its top at origin.y, then `rows` rows of unit boxes (row i has rows-i boxes)."""
    ox, oy = origin
    bd = api.default_body_def()
    bd.position = api.Vec2(float(f32(ox)), float(f32(oy)))
    ground = world.CreateBody(bd)
    ghw = ground_half_width if ground_half_width is not None else float(rows)
    world.CreatePolygonShape(ground, _shape_def(), geo.MakeOffsetBox(ghw, 1.0, (0.0, -1.0)))
    box = geo.MakeBox(0.5, 0.5)
    sd = _shape_def()
    bodies = []
    state = jitter_seed if jitter_seed is not None else 0
    for i in range(rows):
        count = rows - i
        for j in range(count):
            x = ox + (j - (count - 1) * 0.5) * 1.0
            if jitter_seed is not None:
                state, dx = _uniform(state, -jitter, jitter)
                x += dx
            y = oy + 0.5 + i * 1.0
            b = _dynamic_body(world, x, y)
            world.CreatePolygonShape(b, sd, box)
            bodies.append(b)
    return ground, bodies


def large_pyramid(world, rows=100):
    return add_pyramid(world, rows)


def many_pyramids(world, nx=20, ny=20, rows=10, pitch=15.0):
    out = []
    for gy in range(ny):
        for gx in range(nx):
            out.append(add_pyramid(world, rows, origin=(gx * pitch, gy * pitch), ground_half_width=6.0))
    return out


def tumbler(world, n=20000, half=25.0, pitch=0.3, seed=12345, angular_velocity=0.7853981852531433):
    """Kinematic rotating container with n mixed circles/capsules/boxes on a jittered lattice."""
    bd = api.default_body_def()
    bd.type = api.KINEMATIC_BODY
    bd.angularVelocity = angular_velocity      # 0.25*pi as fp32
    container = world.CreateBody(bd)
    sd = _shape_def()
    t = 0.5
    for (hx, hy, c) in [(t, half + t, (half, 0.0)), (t, half + t, (-half, 0.0)), (half + t, t, (0.0, half)), (half + t, t, (0.0, -half))]:
        world.CreatePolygonShape(container, sd, geo.MakeOffsetBox(hx, hy, c))
    cols = int(np.ceil(np.sqrt(n)))
    state = seed
    bodies = []
    circle = geo.Circle((0.0, 0.0), 0.1)
    capsule = geo.Capsule((-0.05, 0.0), (0.05, 0.0), 0.075)
    box = geo.MakeBox(0.1, 0.1)
    for k in range(n):
        i, j = k % cols, k // cols
        state, dx = _uniform(state, -0.01, 0.01)
        state, dy = _uniform(state, -0.01, 0.01)
        x = (i - (cols - 1) * 0.5) * pitch + dx
        y = (j - (cols - 1) * 0.5) * pitch + dy
        b = _dynamic_body(world, x, y)
        world.CreateShape(b, sd, (circle, capsule, box)[k % 3])
        bodies.append(b)
    return container, bodies


def joint_grid(world, n=100, with_ground=True):
    """n x n circles on a 1 m lattice, revolute joints to the right and lower neighbour, top row pinned."""
    bd = api.default_body_def()
    anchor = world.CreateBody(bd)
    sd = _shape_def()
    if with_ground:
        ground = world.CreateBody(bd)
        world.CreatePolygonShape(ground, sd, geo.MakeOffsetBox(float(n), 1.0, (0.5 * (n - 1), -(n - 1) - 0.6 - 1.0)))
    circle = geo.Circle((0.0, 0.0), 0.4)
    ids = [[None] * n for _ in range(n)]
    for j in range(n):          # row j (0 = top), y = -j
        for i in range(n):
            b = _dynamic_body(world, float(i), float(-j))
            world.CreateCircleShape(b, sd, circle)
            ids[j][i] = b
    joints = []
    for j in range(n):
        for i in range(n):
            if j == 0:
                jd = api.default_revolute_joint_def()
                jd.bodyIdA, jd.bodyIdB = anchor, ids[j][i]
                jd.localFrameA.p = api.Vec2(float(i), 0.5)
                jd.localFrameB.p = api.Vec2(0.0, 0.5)
                joints.append(world.CreateRevoluteJoint(jd))
            if i + 1 < n:
                jd = api.default_revolute_joint_def()
                jd.bodyIdA, jd.bodyIdB = ids[j][i], ids[j][i + 1]
                jd.localFrameA.p = api.Vec2(0.5, 0.0)
                jd.localFrameB.p = api.Vec2(-0.5, 0.0)
                joints.append(world.CreateRevoluteJoint(jd))
            if j + 1 < n:
                jd = api.default_revolute_joint_def()
                jd.bodyIdA, jd.bodyIdB = ids[j][i], ids[j + 1][i]
                jd.localFrameA.p = api.Vec2(0.0, -0.5)
                jd.localFrameB.p = api.Vec2(0.0, 0.5)
                joints.append(world.CreateRevoluteJoint(jd))
    return ids, joints


def mixed_drop(world, n=120, seed=7):
    """Small scene touching every manifold pair type the step path supports: static segments + a static rounded
    box as terrain, a kinematic paddle, and dynamic circles / capsules / boxes / rounded boxes / 3-gons dropped on it."""
    bd = api.default_body_def()
    ground = world.CreateBody(bd)
    sd = _shape_def()
    world.CreateSegmentShape(ground, sd, geo.Segment((-12.0, 0.0), (12.0, 0.0)))
    world.CreateSegmentShape(ground, sd, geo.Segment((-12.0, 0.0), (-14.0, 6.0)))
    world.CreateSegmentShape(ground, sd, geo.Segment((12.0, 0.0), (14.0, 6.0)))
    world.CreatePolygonShape(ground, sd, geo.MakeOffsetBox(2.0, 0.25, (0.0, 0.25)))
    kd = api.default_body_def()
    kd.type = api.KINEMATIC_BODY
    kd.position = api.Vec2(-5.0, 1.5)
    kd.angularVelocity = 1.0
    paddle = world.CreateBody(kd)
    world.CreatePolygonShape(paddle, sd, geo.MakeBox(1.5, 0.1))
    world.CreateCapsuleShape(paddle, sd, geo.Capsule((0.0, -0.6), (0.0, 0.6), 0.1))
    shapes = [geo.Circle((0.0, 0.0), 0.2), geo.Capsule((-0.2, 0.0), (0.2, 0.0), 0.12), geo.MakeBox(0.2, 0.15),
              geo.MakeRoundedBox(0.15, 0.15, 0.05), geo.Circle((0.1, 0.0), 0.15)]
    state = seed
    bodies = []
    cols = 12
    for k in range(n):
        i, j = k % cols, k // cols
        state, dx = _uniform(state, -0.05, 0.05)
        b = _dynamic_body(world, (i - 5.5) * 0.9 + dx, 1.2 + j * 0.6)
        world.CreateShape(b, sd, shapes[k % len(shapes)])
        if k % 7 == 0:      # a second shape on some bodies (multi-shape mass + shape-list order)
            world.CreateCircleShape(b, sd, geo.Circle((0.0, 0.2), 0.1))
        bodies.append(b)
    return ground, paddle, bodies


def sensor_field(world, n=60, seed=21):
    """Sensors of every kind (src/Sensor.cs): static sensor boxes / circles / a capsule over a ground, a sensor riding a kinematic
    sweeper, a sensor carried by a dynamic body, a sensor whose own events are off, and visitors (circles, capsules, boxes) of which
    some have enableSensorEvents off and some are filtered out by category. Returns (sensor shape ids, visitor body ids)."""
    sd = _shape_def()
    ground = world.CreateBody(api.default_body_def())
    world.CreatePolygonShape(ground, _shape_def(enableSensorEvents=1), geo.MakeOffsetBox(14.0, 0.5, (0.0, -0.5)))   # a static visitor
    sensors = []
    zone = world.CreateBody(api.default_body_def())
    sen = _shape_def(isSensor=1, enableSensorEvents=1)
    sensors.append(world.CreatePolygonShape(zone, sen, geo.MakeOffsetBox(3.0, 1.0, (-6.0, 2.5))))
    sensors.append(world.CreateCircleShape(zone, sen, geo.Circle((0.0, 3.0), 1.5)))
    sensors.append(world.CreateCapsuleShape(zone, sen, geo.Capsule((4.0, 1.0), (8.0, 2.0), 0.6)))
    sensors.append(world.CreatePolygonShape(zone, _shape_def(isSensor=1, enableSensorEvents=0), geo.MakeOffsetBox(2.0, 2.0, (0.0, 2.0))))   # events off: never reports
    sensors.append(world.CreatePolygonShape(zone, _shape_def(isSensor=1, enableSensorEvents=1, maskBits=2), geo.MakeOffsetBox(14.0, 3.0, (0.0, 3.0))))   # sees category 2 only
    sensors.append(world.CreatePolygonShape(zone, sen, geo.MakeOffsetBox(1.0, 0.4, (-6.0, 2.0))))      # a sensor inside another sensor's box (the port does not skip sensors)
    kd = api.default_body_def()
    kd.type = api.KINEMATIC_BODY
    kd.position = api.Vec2(-10.0, 1.2)
    kd.linearVelocity = api.Vec2(4.0, 0.0)
    sweeper = world.CreateBody(kd)
    world.CreatePolygonShape(sweeper, sd, geo.MakeBox(0.3, 0.3))
    sensors.append(world.CreateCircleShape(sweeper, sen, geo.Circle((0.0, 0.0), 1.2)))
    carrier = _dynamic_body(world, 9.0, 6.0)
    world.CreatePolygonShape(carrier, _shape_def(enableSensorEvents=1), geo.MakeBox(0.4, 0.4))
    sensors.append(world.CreateCircleShape(carrier, sen, geo.Circle((0.0, 0.0), 1.0)))                 # rides a falling body; its own body's box is skipped
    shapes = [geo.Circle((0.0, 0.0), 0.2), geo.Capsule((-0.2, 0.0), (0.2, 0.0), 0.12), geo.MakeBox(0.2, 0.15), geo.MakeRoundedBox(0.15, 0.15, 0.05)]
    state = seed
    bodies = []
    cols = 15
    for k in range(n):
        i, j = k % cols, k // cols
        state, dx = _uniform(state, -0.1, 0.1)
        b = _dynamic_body(world, (i - 7.0) * 1.3 + dx, 4.5 + j * 0.8)
        vd = _shape_def(enableSensorEvents=int(k % 5 != 0), categoryBits=(2 if k % 3 == 0 else 1))
        world.CreateShape(b, vd, shapes[k % len(shapes)])
        bodies.append(b)
    return sensors, bodies, carrier, sweeper


def fast_rain(world, n=48, seed=5, walls="box"):
    """Continuous collision (src/Solver.cs:315-401): small fast bodies fired at thin static walls and a thin floor, so that without a time of
    impact they would tunnel; a static sensor strip in front of the floor collects continuous sensor hits. walls="box": every static shape
    is a polygon (4 points: the port's cached-simplex read stays in range); walls="segment": segment walls, for which the C# code throws."""
    sd = _shape_def(enableSensorEvents=1)
    ground = world.CreateBody(api.default_body_def())
    if walls == "box":
        world.CreatePolygonShape(ground, sd, geo.MakeOffsetBox(20.0, 0.05, (0.0, -0.05)))
        world.CreatePolygonShape(ground, sd, geo.MakeOffsetBox(0.05, 6.0, (12.0, 6.0)))
        world.CreatePolygonShape(ground, sd, geo.MakeOffsetBox(0.05, 6.0, (-12.0, 6.0)))
        world.CreatePolygonShape(ground, sd, geo.MakeOffsetBox(0.6, 0.6, (0.0, 1.0), geo.MakeRot(0.4)))
        world.CreatePolygonShape(ground, sd, geo.MakeOffsetBox(1.5, 0.1, (5.5, 1.0), geo.MakeRot(0.25)))
    elif walls == "round":        # a static circle (1 point) and capsule (2 points): a fast box or capsule indexes past them -> the C# code throws
        world.CreatePolygonShape(ground, sd, geo.MakeOffsetBox(20.0, 0.05, (0.0, -0.05)))
        world.CreateCircleShape(ground, sd, geo.Circle((0.0, 1.0), 0.6))
        world.CreateCapsuleShape(ground, sd, geo.Capsule((4.0, 0.6), (7.0, 1.4), 0.15))
    else:
        world.CreateSegmentShape(ground, sd, geo.Segment((-20.0, 0.0), (20.0, 0.0)))
    strip = world.CreatePolygonShape(ground, _shape_def(isSensor=1, enableSensorEvents=1), geo.MakeOffsetBox(20.0, 0.1, (0.0, 0.5)))
    shapes = [geo.Circle((0.0, 0.0), 0.08), geo.Capsule((-0.1, 0.0), (0.1, 0.0), 0.05), geo.MakeBox(0.1, 0.06), geo.MakeRoundedBox(0.06, 0.06, 0.03),
              geo.Circle((0.05, 0.0), 0.06)]
    state = seed
    bodies = []
    for k in range(n):
        state, x = _uniform(state, -10.0, 10.0)
        state, y = _uniform(state, 3.0, 9.0)
        state, vx = _uniform(state, -60.0, 60.0)
        state, vy = _uniform(state, -150.0, -20.0)
        state, w = _uniform(state, -20.0, 20.0)
        b = _dynamic_body(world, x, y, linearVelocity=api.Vec2(vx, vy), angularVelocity=w)
        world.CreateShape(b, sd, shapes[k % len(shapes)])
        if k % 6 == 0:
            world.CreateCircleShape(b, sd, geo.Circle((0.0, 0.1), 0.04))
        bodies.append(b)
    return ground, strip, bodies


def bullet_range(world, n=24, seed=7, targets="box"):
    """Bullets (BodyDef.isBullet, src/Solver.cs:447-449,728-738,1238-1272): small fast bodies fired at thin DYNAMIC plates and a moving
    kinematic plate, which only the bullet stage's sweep of the kinematic and dynamic trees can stop; every third projectile is not a
    bullet (it stops at static shapes only and tunnels through the plates), two bullets cross each other's path (bullets ignore bullets),
    one bullet is jointed to a plate with collideConnected off. targets="box": every target is a 4-point polygon (the port's cached-simplex
    read stays in range); targets="round": dynamic circles, for which the C# code indexes past the proxy and throws.
    Returns (ground, plates, projectiles)."""
    sd = _shape_def(enableSensorEvents=1)
    ground = world.CreateBody(api.default_body_def())
    world.CreatePolygonShape(ground, sd, geo.MakeOffsetBox(30.0, 0.05, (0.0, -0.05)))
    world.CreatePolygonShape(ground, sd, geo.MakeOffsetBox(0.05, 8.0, (16.0, 8.0)))
    world.CreatePolygonShape(ground, _shape_def(isSensor=1, enableSensorEvents=1), geo.MakeOffsetBox(0.2, 8.0, (14.0, 8.0)))
    plates = []
    for k in range(6):                                  # thin upright dynamic plates, 0.08 m thick, standing on the ground
        b = _dynamic_body(world, 4.0 + 1.5 * k, 1.5)
        if targets == "box":
            world.CreatePolygonShape(b, sd, geo.MakeBox(0.04, 1.5))
        else:
            world.CreateCircleShape(b, sd, geo.Circle((0.0, 0.0), 0.6))
        plates.append(b)
    bd = api.default_body_def()
    bd.type = api.KINEMATIC_BODY
    bd.position = api.Vec2(-3.0, 6.0)
    bd.linearVelocity = api.Vec2(0.5, 0.0)
    mover = world.CreateBody(bd)
    world.CreatePolygonShape(mover, sd, geo.MakeBox(0.04, 2.0))
    plates.append(mover)
    shapes = [geo.MakeBox(0.08, 0.05), geo.Circle((0.0, 0.0), 0.07), geo.Capsule((-0.08, 0.0), (0.08, 0.0), 0.04), geo.MakeRoundedBox(0.05, 0.05, 0.02)]
    state = seed
    projectiles = []
    for k in range(n):
        state, y = _uniform(state, 0.4, 7.5)
        state, vx = _uniform(state, 80.0, 260.0)
        state, vy = _uniform(state, -15.0, 15.0)
        state, w = _uniform(state, -30.0, 30.0)
        b = _dynamic_body(world, -8.0 - 0.3 * k, y, linearVelocity=api.Vec2(vx, vy), angularVelocity=w, isBullet=0 if k % 3 == 2 else 1,
                          gravityScale=0.2)
        # (a projectile that is not a bullet is itself a target of the bullets: polygons only, or the C# code throws)
        world.CreateShape(b, sd, shapes[k % len(shapes)] if k % 3 != 2 else shapes[0 if k % 2 else 3])
        if k % 5 == 0:
            world.CreatePolygonShape(b, sd, geo.MakeOffsetBox(0.03, 0.03, (0.0, 0.09)))
        projectiles.append(b)
    # two bullets on crossing paths
    for sx in (-1.0, 1.0):
        b = _dynamic_body(world, 1.0 + 2.0 * sx, 10.0, linearVelocity=api.Vec2(-120.0 * sx, 0.0), isBullet=1, gravityScale=0.0)
        world.CreatePolygonShape(b, sd, geo.MakeBox(0.06, 0.06))
        projectiles.append(b)
    # a bullet tied to the first plate by a joint that switches their collision off: it flies through that plate and stops at the next
    b = _dynamic_body(world, 2.5, 1.5, linearVelocity=api.Vec2(200.0, 0.0), isBullet=1, gravityScale=0.0)
    world.CreatePolygonShape(b, sd, geo.MakeBox(0.05, 0.05))
    jd = api.default_joint_def(api.capi.JOINT_DISTANCE)
    jd.bodyIdA, jd.bodyIdB = plates[0], b
    jd.length = 1.5
    jd.enableSpring, jd.hertz, jd.dampingRatio = 1, 0.2, 0.0
    jd.collideConnected = 0
    world.CreateJoint(jd)
    projectiles.append(b)
    return ground, plates, projectiles


def joint_hub(world, spokes=28, anchors=25):
    """Joints in the overflow colour (src/ConstraintGraph.cs:91-128, src/Joint.cs:377-394): a dynamic hub with `spokes` joints to dynamic
    satellites (a pair of dynamic bodies may use 20 colours: the rest overflow) and a second hub with `anchors` joints to the static ground
    (22 colours). Plain revolutes, rigid and springy distance joints and velocity motors (the bit-exact types), thresholds on some of
    them: an overflow joint never reports a joint event. The hubs' contacts find every colour taken and overflow as well.
    Returns (hub, satellites, joint ids)."""
    J = api.capi
    sd = _shape_def()
    ground = world.CreateBody(api.default_body_def())
    world.CreateSegmentShape(ground, sd, geo.Segment((-30.0, 0.0), (30.0, 0.0)))

    def jdef(t, a, b, pa, pb):
        d = api.default_joint_def(t)
        d.bodyIdA, d.bodyIdB = a, b
        d.localFrameA.p = api.Vec2(float(f32(pa[0])), float(f32(pa[1])))
        d.localFrameB.p = api.Vec2(float(f32(pb[0])), float(f32(pb[1])))
        return d

    joints = []
    hub = _dynamic_body(world, 0.0, 4.0, angularVelocity=0.7)
    world.CreatePolygonShape(hub, sd, geo.MakeBox(0.8, 0.8))
    satellites = []
    for k in range(spokes):
        a = 2.0 * np.pi * k / spokes
        cx, cy = 3.0 * np.cos(a), 3.0 * np.sin(a)
        b = _dynamic_body(world, cx, 4.0 + cy)
        if k % 2:
            world.CreateCircleShape(b, sd, geo.Circle((0.0, 0.0), 0.25))
        else:
            world.CreatePolygonShape(b, sd, geo.MakeBox(0.2, 0.2))
        satellites.append(b)
        hx, hy = 0.7 * np.cos(a), 0.7 * np.sin(a)
        kind = k % 4
        if kind == 0:
            d = jdef(J.JOINT_REVOLUTE, hub, b, (hx, hy), (hx - cx, hy - cy))
            d.forceThreshold = 2.0
        elif kind == 1:
            d = jdef(J.JOINT_DISTANCE, hub, b, (hx, hy), (0.0, 0.0))
            d.length = 2.3
        elif kind == 2:
            d = jdef(J.JOINT_DISTANCE, hub, b, (hx, hy), (0.0, 0.0))
            d.length = 2.3
            d.enableSpring, d.hertz, d.dampingRatio = 1, 2.0, 0.2
            d.enableLimit, d.minLength, d.maxLength = 1, 1.5, 3.0
        else:
            d = jdef(J.JOINT_MOTOR, hub, b, (cx, cy), (0.0, 0.0))
            d.linearVelocity, d.maxVelocityForce = api.Vec2(0.0, 0.0), 30.0
            d.angularVelocity, d.maxVelocityTorque = 0.5, 2.0
        d.collideConnected = 1 if k % 5 == 0 else 0
        joints.append(world.CreateJoint(d))
    # a second hub tied to the static ground body
    hub2 = _dynamic_body(world, 12.0, 3.0)
    world.CreatePolygonShape(hub2, sd, geo.MakeBox(1.0, 0.5))
    for k in range(anchors):
        x = 12.0 + (k - anchors / 2) * 0.4
        if k % 3 == 0:
            d = jdef(J.JOINT_REVOLUTE, ground, hub2, (12.0, 3.0 + 0.001 * k), (0.0, 0.001 * k))
            d.forceThreshold = 1.0
        else:
            d = jdef(J.JOINT_DISTANCE, ground, hub2, (x, 8.0), ((k % 5 - 2) * 0.2, 0.0))
            d.length = 5.0
            d.enableSpring, d.hertz, d.dampingRatio = 1, 1.0 + 0.1 * k, 0.1
        joints.append(world.CreateJoint(d))
    # boxes dropped on both hubs: their contacts with a hub find its colours taken
    for k in range(6):
        b = _dynamic_body(world, -0.5 + 0.2 * k if k < 3 else 11.5 + 0.3 * (k - 3), 5.5 + 0.6 * (k % 3))
        world.CreatePolygonShape(b, sd, geo.MakeBox(0.25, 0.25))
    return hub, satellites, joints


def overflow_plank(world, n=26):
    """A long dynamic plank resting on n small dynamic boxes: the plank needs more than the 20 dynamic colours, so some
    contacts land in the overflow colour (src/ConstraintGraph.cs:140) and are solved by the scalar overflow path."""
    bd = api.default_body_def()
    ground = world.CreateBody(bd)
    sd = _shape_def()
    half = 0.5 * n * 0.6 + 2.0
    world.CreateSegmentShape(ground, sd, geo.Segment((-half, 0.0), (half, 0.0)))
    bodies = []
    for i in range(n):
        b = _dynamic_body(world, (i - 0.5 * (n - 1)) * 0.6, 0.25)
        world.CreatePolygonShape(b, sd, geo.MakeBox(0.25, 0.25))
        bodies.append(b)
    plank = _dynamic_body(world, 0.0, 0.62)
    world.CreatePolygonShape(plank, sd, geo.MakeBox(0.5 * n * 0.6, 0.1))
    for i in range(0, n, 3):        # a few circles dropped on the plank keep contacts changing
        c = _dynamic_body(world, (i - 0.5 * (n - 1)) * 0.6, 1.2 + 0.05 * (i % 4))
        world.CreateCircleShape(c, sd, geo.Circle((0.0, 0.0), 0.15))
        bodies.append(c)
    return ground, plank, bodies


def joint_zoo(world, copies=1, pitch=40.0, angles=True):
    """Every joint type of the reference (include/Types.cs:423-426) with its options switched on, over a ground segment so that
    contacts share the colouring with the joints: rigid and springy distance joints (limit + motor), rigid and soft welds, a
    prismatic slider (spring, limit, motor), motor joints (velocity and spring drives), a filter joint (overlapping bodies that must not collide) and a revolute pendulum with force/torque thresholds."""
    J = api.capi
    sd = _shape_def()
    out = []
    for c in range(copies):
        ox = c * pitch
        ground = world.CreateBody(api.default_body_def())
        world.CreateSegmentShape(ground, sd, geo.Segment((ox - 18.0, 0.0), (ox + 18.0, 0.0)))

        def dyn(x, y, angle=0.0, w=0.0, v=(0.0, 0.0)):
            bd = api.default_body_def()
            bd.type = api.DYNAMIC_BODY
            bd.position = api.Vec2(ox + x, y)
            bd.rotation = api.Rot(float(np.float32(np.cos(angle))), float(np.float32(np.sin(angle))))
            bd.angularVelocity = w
            bd.linearVelocity = api.Vec2(*v)
            return world.CreateBody(bd)

        def jdef(t, a, b, pa, pb, qa=0.0, qb=0.0):
            d = api.default_joint_def(t)
            d.bodyIdA, d.bodyIdB = a, b
            d.localFrameA = api.Transform(api.Vec2(*pa), api.Rot(float(np.float32(np.cos(qa))), float(np.float32(np.sin(qa)))))
            d.localFrameB = api.Transform(api.Vec2(*pb), api.Rot(float(np.float32(np.cos(qb))), float(np.float32(np.sin(qb)))))
            return d

        ids = {}
        # distance: rigid rope from the ground anchor, then a springy link with limits and a motor between two dynamic bodies
        b0 = dyn(-15.0, 6.0, w=0.5)
        world.CreatePolygonShape(b0, sd, geo.MakeBox(0.4, 0.3))
        d = jdef(J.JOINT_DISTANCE, ground, b0, (ox - 14.0, 9.0), (0.2, 0.1))
        d.length = 3.0
        d.forceThreshold = 20.0
        ids["distance_rigid"] = world.CreateJoint(d)
        b1 = dyn(-15.0, 4.0)
        world.CreateCircleShape(b1, sd, geo.Circle((0.0, 0.0), 0.35))
        d = jdef(J.JOINT_DISTANCE, b0, b1, (0.0, -0.3), (0.0, 0.1))
        d.length = 1.5
        d.enableSpring, d.hertz, d.dampingRatio = 1, 2.0, 0.3
        d.lowerSpringForce, d.upperSpringForce = -50.0, 80.0
        d.enableLimit, d.minLength, d.maxLength = 1, 1.0, 2.2
        d.enableMotor, d.maxMotor, d.motorSpeed = 1, 3.0, 0.5
        ids["distance_spring"] = world.CreateJoint(d)
        # weld: rigid cantilever off the ground body, then a soft weld at its tip
        prev, px = ground, -10.0
        for k in range(3 if angles else 0):
            b = dyn(px + 1.0 * k + 0.5, 5.0)
            world.CreatePolygonShape(b, sd, geo.MakeBox(0.5, 0.125))
            if prev == ground:
                d = jdef(J.JOINT_WELD, prev, b, (ox + px, 5.0), (-0.5, 0.0))
            else:
                d = jdef(J.JOINT_WELD, prev, b, (0.5, 0.0), (-0.5, 0.0))
            if k == 2:
                d.hertz, d.dampingRatio, d.angularHertz, d.angularDampingRatio = 5.0, 0.7, 3.0, 0.5
                d.torqueThreshold = 0.5
            ids[f"weld{k}"] = world.CreateJoint(d)
            prev = b
        # prismatic: slider on a tilted axis with spring, limits and a motor
        if angles:
            b = dyn(-4.0, 3.0, angle=0.2)
            world.CreatePolygonShape(b, sd, geo.MakeBox(0.5, 0.25))
            d = jdef(J.JOINT_PRISMATIC, ground, b, (ox - 4.0, 3.0), (0.0, 0.0), qa=0.35, qb=0.15)
            d.enableSpring, d.hertz, d.dampingRatio, d.target = 1, 1.5, 0.4, 0.5
            d.enableLimit, d.lower, d.upper = 1, -1.0, 2.0
            d.enableMotor, d.maxMotor, d.motorSpeed = 1, 8.0, -1.0
            ids["prismatic"] = world.CreateJoint(d)
            b2 = dyn(-1.0, 2.0)
            world.CreateCircleShape(b2, sd, geo.Circle((0.0, 0.0), 0.3))
            d = jdef(J.JOINT_PRISMATIC, b, b2, (0.5, 0.0), (0.0, 0.0), qa=1.2)
            ids["prismatic_plain"] = world.CreateJoint(d)
        # motor: velocity drive against the ground, spring drive between two dynamic bodies
        m0 = dyn(10.0, 4.0)
        world.CreatePolygonShape(m0, sd, geo.MakeBox(0.4, 0.4))
        d = jdef(J.JOINT_MOTOR, ground, m0, (ox + 10.0, 4.0), (0.0, 0.0))
        d.linearVelocity, d.maxVelocityForce = api.Vec2(0.5, 0.25), 200.0
        d.angularVelocity, d.maxVelocityTorque = 1.0, 20.0
        ids["motor_velocity"] = world.CreateJoint(d)
        m1 = dyn(12.0, 4.5, angle=0.3)
        world.CreateCapsuleShape(m1, sd, geo.Capsule((-0.3, 0.0), (0.3, 0.0), 0.15))
        d = jdef(J.JOINT_MOTOR, m0, m1, (0.4, 0.0), (-0.45, 0.0))
        d.hertz, d.dampingRatio, d.maxSpringForce = 3.0, 0.5, 60.0
        if angles:
            d.angularHertz, d.angularDampingRatio, d.maxSpringTorque = 2.0, 0.8, 5.0
        d.maxVelocityForce, d.maxVelocityTorque = 1.0, 0.5
        ids["motor_spring"] = world.CreateJoint(d)
        # filter: two overlapping boxes that fall together without colliding
        f0, f1 = dyn(15.0, 2.0), dyn(15.2, 2.1)
        world.CreatePolygonShape(f0, sd, geo.MakeBox(0.4, 0.4))
        world.CreatePolygonShape(f1, sd, geo.MakeBox(0.4, 0.4))
        ids["filter"] = world.CreateJoint(jdef(J.JOINT_FILTER, f0, f1, (0.0, 0.0), (0.0, 0.0)))
        # revolute pendulum with thresholds (joint events)
        p0 = dyn(-17.0, 3.0)
        world.CreatePolygonShape(p0, sd, geo.MakeBox(0.2, 1.0))
        d = jdef(J.JOINT_REVOLUTE, ground, p0, (ox - 17.0, 4.0), (0.0, 1.0))
        d.forceThreshold = 5.0
        ids["revolute"] = world.CreateJoint(d)
        out.append(ids)
    return out


def wheel_rig(world, copies=1, pitch=12.0):
    """Wheel joints (spring, limit, motor). The port's WheelJoint.Solve measures the point-to-line velocity along axisA instead of
    perpA (src/joints/WheelJoint.cs:266), so a free chassis on two sprung wheels gains energy within a few steps; the rig hangs the
    wheels from the static ground body on vertical axes instead, where the same code stays bounded, and adds one free two-body pair."""
    J = api.capi
    sd = _shape_def()
    out = []
    for c in range(copies):
        ox = c * pitch
        ground = world.CreateBody(api.default_body_def())
        world.CreateSegmentShape(ground, sd, geo.Segment((ox - 6.0, -20.0), (ox + 6.0, -20.0)))      # out of reach: the rig swings wide
        ids = []
        for k, wx in enumerate((-3.0, 0.0, 3.0)):
            bd = api.default_body_def()
            bd.type = api.DYNAMIC_BODY
            bd.position = api.Vec2(ox + wx, 2.0)
            bd.angularVelocity = 0.5 * k
            wheel = world.CreateBody(bd)
            world.CreateCircleShape(wheel, sd, geo.Circle((0.0, 0.0), 0.4))
            d = api.default_joint_def(J.JOINT_WHEEL)
            d.bodyIdA, d.bodyIdB = ground, wheel
            q = 0.5 * np.pi + 0.1 * k
            d.localFrameA = api.Transform(api.Vec2(ox + wx, 2.0), api.Rot(float(np.float32(np.cos(q))), float(np.float32(np.sin(q)))))
            d.enableSpring, d.hertz, d.dampingRatio = 1, 3.0, 0.7
            d.enableLimit, d.lower, d.upper = int(k != 1), -0.15, 0.25
            d.enableMotor, d.maxMotor, d.motorSpeed = int(k != 2), 4.0, 2.0
            d.torqueThreshold = 1.0
            ids.append(world.CreateJoint(d))
        out.append(ids)
    return out


def bounce_mix(world, n=48, seed=11, hits=True):
    """Restitution != 0 mixed with restitution 0 (ApplyRestitutionTask, src/ContactSolver.cs:913-989, quirk Q1) and hit events
    (src/Solver.cs:1151-1190): many separate islands land on one static segment, so their ground contacts share colours and the
    8-lane SIMD groups of a colour mix zero and non-zero restitution lanes of different islands; bodies that leave the ground again
    swap-remove their contacts from the colour arrays, which reshuffles the groups and the hit-event order."""
    bd = api.default_body_def()
    ground = world.CreateBody(bd)
    sd0 = _shape_def(enableHitEvents=int(hits))
    half = 0.6 * n + 12.0
    world.CreateSegmentShape(ground, sd0, geo.Segment((-half, 0.0), (half, 0.0)))
    rest = (0.0, 0.6, 0.0, 0.0, 0.9, 0.3, 0.0)
    shapes = (geo.Circle((0.0, 0.0), 0.25), geo.MakeBox(0.25, 0.2), geo.Capsule((-0.15, 0.0), (0.15, 0.0), 0.15))
    state = seed
    bodies = []
    for k in range(n):
        state, dx = _uniform(state, -0.05, 0.05)
        x = (k - 0.5 * (n - 1)) * 1.1 + dx
        y = 0.6 + (k % 5) * 0.45
        b = _dynamic_body(world, x, y, linearVelocity=api.Vec2(0.0, -float(2 + (k % 3))))
        sd = _shape_def(enableHitEvents=int(hits and k % 4 != 3))
        sd.restitution = rest[k % len(rest)]
        world.CreateShape(b, sd, shapes[k % 3])
        bodies.append(b)
        if k % 6 == 0:      # a second body dropped on top: body-body contacts in the dynamic colours, restitution from either side
            t = _dynamic_body(world, x, y + 0.7, linearVelocity=api.Vec2(0.0, -3.0))
            sdt = _shape_def(enableHitEvents=int(hits))
            sdt.restitution = rest[(k + 1) % len(rest)]
            world.CreatePolygonShape(t, sdt, geo.MakeBox(0.2, 0.2))
            bodies.append(t)
    # a small resting pyramid (restitution 0 everywhere) on the same ground: its ground contacts share colour 22 with the bouncers
    box = geo.MakeBox(0.5, 0.5)
    px = 0.5 * (n - 1) * 1.1 + 6.0
    for i in range(3):
        for j in range(3 - i):
            b = _dynamic_body(world, px + (j - (3 - i - 1) * 0.5) * 1.0, 0.5 + i * 1.0)
            world.CreatePolygonShape(b, sd0, box)
            bodies.append(b)
    return ground, bodies


def batched_pyramids(world, n_worlds=8192, first=0, rows=10, pitch=15.0, cols=128):
    """BASELINE.json configs[4]: `n_worlds` independent 10-row pyramids (world index first..first+n), each on its own ground,
    x positions jittered per world by splitmix64(seed=world index) so worlds are not bit-copies."""
    for k in range(first, first + n_worlds):
        add_pyramid(world, rows, origin=((k % cols) * pitch, (k // cols) * pitch), ground_half_width=6.0, jitter_seed=k, jitter=0.001)


SCENES = {
    "fast_rain": fast_rain,
    "sensor_field": sensor_field,
    "large_pyramid": large_pyramid,
    "many_pyramids": many_pyramids,
    "tumbler": tumbler,
    "joint_grid": joint_grid,
    "mixed_drop": mixed_drop,
    "overflow_plank": overflow_plank,
    "joint_zoo": joint_zoo,
    "wheel_rig": wheel_rig,
    "batched_pyramids": batched_pyramids,
    "bounce_mix": bounce_mix,
}
