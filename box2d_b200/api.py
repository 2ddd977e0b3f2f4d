"""Host-side mirror of the reference's Box2D.API surface for the step path, over the C-ABI.

Names and argument meaning follow the C# facade (api/World.cs, api/Body.cs, api/Shape.cs, api/Joint.cs):
``WorldAPI.CreateWorld/Step/GetContactEvents``, ``BodyAPI.CreateBody/SetTransform/...``,
``ShapeAPI.CreatePolygonShape/...``, ``JointAPI.CreateRevoluteJoint``. Errors from the library raise
``B2Error`` (the C# shim would throw likewise); unsupported features fail loudly at the call that needs them.

``World`` is backend-agnostic: it is handed a bound function table (see capi.bind). ``CudaBackend`` binds the
product library; the oracle is bound only by tests/ (never from this package).
"""
import os
import ctypes as C
import numpy as np
from . import capi
from .capi import (Vec2, Rot, Transform, WorldDef, BodyDef, ShapeDef, Geometry, RevoluteJointDef, JointDef, BodyRead, ContactRead,
                   ContactTouchEvent, ContactHitEvent, JointRead, ShapeRead, Counters, Profile)

STATIC_BODY, KINEMATIC_BODY, DYNAMIC_BODY = 0, 1, 2
FLT_MAX = 3.4028234663852886e38


class B2Error(RuntimeError):
    pass


_DTYPES = {}


def _dt(ty):
    """numpy dtype of a ctypes Structure, converted once (np.dtype(struct) costs ~25 us a call)."""
    d = _DTYPES.get(ty)
    if d is None:
        d = _DTYPES[ty] = np.dtype(ty)
    return d


def _records(ptr, n, ty, copy):
    """`n` records of ctypes type `ty` at `ptr` as a numpy structured array: a view of the library's buffer, or a copy."""
    d = _dt(ty)
    if n <= 0:
        return np.zeros(0, dtype=d)
    arr = np.frombuffer((C.c_uint8 * (n * d.itemsize)).from_address(C.cast(ptr, C.c_void_p).value), dtype=d)
    return arr.copy() if copy else arr


def default_world_def():
    """include/Types.cs:75-144"""
    d = WorldDef()
    d.gravity = Vec2(0.0, -10.0)
    d.restitutionThreshold = 1.0
    d.hitEventThreshold = 1.0
    d.contactHertz = 30.0
    d.contactDampingRatio = 10.0
    d.contactSpeed = 3.0
    d.maximumLinearSpeed = 400.0
    d.enableSleep = 1
    d.enableContinuous = 1
    d.enableContactSoftening = 0
    return d


def default_body_def():
    """include/Types.cs:170-223"""
    d = BodyDef()
    d.type = STATIC_BODY
    d.rotation = Rot(1.0, 0.0)
    d.gravityScale = 1.0
    d.sleepThreshold = 0.05
    d.enableSleep = 1
    d.isAwake = 1
    d.isEnabled = 1
    return d


def default_shape_def():
    """include/Types.cs:285-344"""
    d = ShapeDef()
    d.friction = 0.6
    d.density = 1.0
    d.categoryBits = 0xFFFFFFFFFFFFFFFF
    d.maskBits = 0xFFFFFFFFFFFFFFFF
    d.invokeContactCreation = 1
    d.updateBodyMass = 1
    return d


def default_revolute_joint_def():
    """include/Types.cs:431-456,569-597"""
    d = RevoluteJointDef()
    d.bodyIdA = -1
    d.bodyIdB = -1
    d.localFrameA = Transform(Vec2(0, 0), Rot(1, 0))
    d.localFrameB = Transform(Vec2(0, 0), Rot(1, 0))
    d.forceThreshold = FLT_MAX
    d.torqueThreshold = FLT_MAX
    d.constraintHertz = 60.0
    d.constraintDampingRatio = 2.0
    d.drawScale = 1.0
    return d


def default_joint_def(joint_type):
    """JointDef base include/Types.cs:431-456 + the per-type defaults :460-586 (Distance/Motor/Filter/Prismatic/Revolute/Weld/Wheel
    JointDef), flattened as B2CJointDef. ``joint_type`` is the reference's JointType value (capi.JOINT_*)."""
    d = JointDef()
    d.type = joint_type
    d.bodyIdA = -1
    d.bodyIdB = -1
    d.localFrameA = Transform(Vec2(0, 0), Rot(1, 0))
    d.localFrameB = Transform(Vec2(0, 0), Rot(1, 0))
    d.forceThreshold = FLT_MAX
    d.torqueThreshold = FLT_MAX
    d.constraintHertz = 60.0
    d.constraintDampingRatio = 2.0
    d.drawScale = 1.0
    d.length = 1.0
    d.maxLength = 100000.0
    d.lowerSpringForce = -FLT_MAX
    d.upperSpringForce = FLT_MAX
    return d


class Backend:
    """A bound C-ABI: fns (common entry points), create(def)->handle, last_error(handle)."""

    def __init__(self, fns, create, last_error, name):
        self.fns, self.create, self.last_error, self.name = fns, create, last_error, name


_cuda_backend = None


def CudaBackend(device=0):
    """Binds libb2cuda.so. Raises if the library (the CUDA extension) is missing: there is no fallback."""
    global _cuda_backend
    path = capi.product_library_path()
    try:
        lib = C.CDLL(path)
    except OSError as e:
        raise B2Error(f"libb2cuda.so not built/loadable at {path}: {e}. Run __graft_entry__.build().") from e
    fns = capi.bind(lib, "b2cuda_")
    extra = {}
    for name, (args, res) in capi.PRODUCT_ONLY.items():
        try:
            f = getattr(lib, "b2cuda_" + name)
        except AttributeError:
            if os.environ.get("B2CUDA_LIB"):      # A/B run against an older build of the library: newer entry points are simply absent
                continue
            raise
        f.argtypes, f.restype = args, res
        extra[name] = f
    fns.update(extra)

    def create(wdef):
        h = C.c_void_p()
        rc = extra["world_create"](C.byref(wdef), device, C.byref(h))
        if rc != 0:
            raise B2Error(f"b2cuda_world_create failed rc={rc}: {extra['last_error']().decode()}")
        return h

    be = Backend(fns, create, lambda h: extra["last_error"]().decode(), "cuda")
    be.device = device
    return be


class World:
    """WorldAPI + BodyAPI + ShapeAPI + JointAPI for one world (api/World.cs:7-113 etc.)."""

    def __init__(self, backend, world_def=None):
        self.backend = backend
        self.fns = backend.fns
        self.handle = backend.create(world_def if world_def is not None else default_world_def())

    def _check(self, rc):
        if rc != 0:
            raise B2Error(f"{self.backend.name} rc={rc}: {self.backend.last_error(self.handle)}")

    # --- BodyAPI ---
    def CreateBody(self, body_def):
        out = C.c_int32(-1)
        self._check(self.fns["create_body"](self.handle, C.byref(body_def), C.byref(out)))
        return out.value

    def SetTransform(self, body_id, position, rotation=(1.0, 0.0)):
        self._check(self.fns["body_set_transform"](self.handle, body_id, Vec2(*position), Rot(*rotation)))

    def SetTargetTransform(self, body_id, position, rotation, time_step):
        t = Transform(Vec2(*position), Rot(*rotation))
        self._check(self.fns["body_set_target_transform"](self.handle, body_id, t, time_step))

    def SetMassData(self, body_id, mass, rotational_inertia, center=(0.0, 0.0)):
        self._check(self.fns["body_set_mass_data"](self.handle, body_id, mass, rotational_inertia, Vec2(*center)))

    def SetMotionLocks(self, body_id, linear_x, linear_y, angular_z):
        self._check(self.fns["body_set_motion_locks"](self.handle, body_id, int(linear_x), int(linear_y), int(angular_z)))

    def SetSurfaceMaterial(self, shape_id, friction, restitution=0.0, rolling_resistance=0.0, tangent_speed=0.0):
        self._check(self.fns["shape_set_surface_material"](self.handle, shape_id, friction, restitution, rolling_resistance, tangent_speed))

    def SetLinearVelocity(self, body_id, v):
        self._check(self.fns["body_set_linear_velocity"](self.handle, body_id, Vec2(*v)))

    def SetAngularVelocity(self, body_id, w):
        self._check(self.fns["body_set_angular_velocity"](self.handle, body_id, float(w)))

    def ApplyForceToCenter(self, body_id, f, wake=True):
        self._check(self.fns["body_apply_force_to_center"](self.handle, body_id, Vec2(*f), int(wake)))

    def ApplyTorque(self, body_id, t, wake=True):
        self._check(self.fns["body_apply_torque"](self.handle, body_id, float(t), int(wake)))

    def ApplyLinearImpulseToCenter(self, body_id, i, wake=True):
        self._check(self.fns["body_apply_linear_impulse_to_center"](self.handle, body_id, Vec2(*i), int(wake)))

    def ApplyForce(self, body_id, f, point, wake=True):
        self._check(self.fns["body_apply_force"](self.handle, body_id, Vec2(*f), Vec2(*point), int(wake)))

    def ClearForces(self, body_id):
        self._check(self.fns["body_clear_forces"](self.handle, body_id))

    def ApplyLinearImpulse(self, body_id, i, point, wake=True):
        self._check(self.fns["body_apply_linear_impulse"](self.handle, body_id, Vec2(*i), Vec2(*point), int(wake)))

    def ApplyAngularImpulse(self, body_id, i, wake=True):
        self._check(self.fns["body_apply_angular_impulse"](self.handle, body_id, float(i), int(wake)))

    def SetLinearDamping(self, body_id, v):
        self._check(self.fns["body_set_linear_damping"](self.handle, body_id, float(v)))

    def SetAngularDamping(self, body_id, v):
        self._check(self.fns["body_set_angular_damping"](self.handle, body_id, float(v)))

    def SetGravityScale(self, body_id, v):
        self._check(self.fns["body_set_gravity_scale"](self.handle, body_id, float(v)))

    def _body(self, body_id):
        out = BodyRead()
        self._check(self.fns["body_get"](self.handle, body_id, C.byref(out)))
        return out

    def GetTransform(self, body_id):
        t = self._body(body_id).transform
        return (t.p.x, t.p.y), (t.q.c, t.q.s)

    def GetPosition(self, body_id):
        return self.GetTransform(body_id)[0]

    def GetLinearVelocity(self, body_id):
        v = self._body(body_id).linearVelocity
        return (v.x, v.y)

    def GetAngularVelocity(self, body_id):
        return self._body(body_id).angularVelocity

    def GetWorldCenterOfMass(self, body_id):
        c = self._body(body_id).center
        return (c.x, c.y)

    # --- ShapeAPI ---
    def CreateShape(self, body_id, shape_def, geometry):
        out = C.c_int32(-1)
        self._check(self.fns["create_shape"](self.handle, body_id, C.byref(shape_def), C.byref(geometry), C.byref(out)))
        return out.value

    CreatePolygonShape = CreateCircleShape = CreateCapsuleShape = CreateSegmentShape = CreateShape

    # --- JointAPI ---
    def CreateRevoluteJoint(self, joint_def):
        out = C.c_int32(-1)
        self._check(self.fns["create_revolute_joint"](self.handle, C.byref(joint_def), C.byref(out)))
        return out.value

    def DestroyJoint(self, joint_id, wake_attached=True):
        """JointAPI.DestroyJoint api/Joint.cs:9-17"""
        self._check(self.fns["destroy_joint"](self.handle, joint_id, int(bool(wake_attached))))

    def DestroyBody(self, body_id):
        """BodyAPI.DestroyBody api/Body.cs:106-160"""
        self._check(self.fns["destroy_body"](self.handle, body_id))

    def DestroyShape(self, shape_id, update_body_mass=True):
        """ShapeAPI.DestroyShape api/Shape.cs:48-56"""
        self._check(self.fns["destroy_shape"](self.handle, shape_id, int(bool(update_body_mass))))

    def SetDensity(self, shape_id, density, update_body_mass=True):
        """ShapeAPI.SetDensity api/Shape.cs:95-103"""
        self._check(self.fns["shape_set_density"](self.handle, shape_id, density, int(bool(update_body_mass))))

    def ApplyMassFromShapes(self, body_id):
        """BodyAPI.ApplyMassFromShapes api/Body.cs:617-621"""
        self._check(self.fns["body_apply_mass_from_shapes"](self.handle, body_id))

    def Shape_EnableContactEvents(self, shape_id, flag):
        """ShapeAPI.EnableContactEvents api/Shape.cs:182-186"""
        self._check(self.fns["shape_enable_events"](self.handle, shape_id, capi.SHAPE_CONTACT_EVENTS, int(bool(flag))))

    def Shape_EnableHitEvents(self, shape_id, flag):
        """ShapeAPI.EnableHitEvents api/Shape.cs:205-209"""
        self._check(self.fns["shape_enable_events"](self.handle, shape_id, capi.SHAPE_HIT_EVENTS, int(bool(flag))))

    def Body_EnableContactEvents(self, body_id, flag):
        """BodyAPI.EnableContactEvents api/Body.cs:848-858"""
        self._check(self.fns["body_enable_events"](self.handle, body_id, capi.SHAPE_CONTACT_EVENTS, int(bool(flag))))

    def Body_EnableHitEvents(self, body_id, flag):
        """BodyAPI.EnableHitEvents api/Body.cs:862-872"""
        self._check(self.fns["body_enable_events"](self.handle, body_id, capi.SHAPE_HIT_EVENTS, int(bool(flag))))

    def Explode(self, position, radius, falloff, impulse_per_length, mask_bits=0xFFFFFFFFFFFFFFFF):
        """WorldAPI.Explode api/World.cs:758-776"""
        self._check(self.fns["world_explode"](self.handle, Vec2(*position), radius, falloff, impulse_per_length, mask_bits))

    def EnableSleeping(self, flag):
        """WorldAPI.EnableSleeping api/World.cs:631-645 (only False is supported)"""
        self._check(self.fns["world_enable_sleeping"](self.handle, int(bool(flag))))

    def SetType(self, body_id, body_type):
        """BodyAPI.SetType api/Body.cs:185-268"""
        self._check(self.fns["body_set_type"](self.handle, body_id, body_type))

    def SetFilter(self, shape_id, category_bits, mask_bits, group_index=0):
        """ShapeAPI.SetFilter api/Shape.cs:157-166"""
        self._check(self.fns["shape_set_filter"](self.handle, shape_id, category_bits, mask_bits, group_index))

    def SetShapeGeometry(self, shape_id, geometry):
        """ShapeAPI.SetCircle / SetCapsule / SetSegment / SetPolygon api/Shape.cs:289-326"""
        self._check(self.fns["shape_set_geometry"](self.handle, shape_id, C.byref(geometry)))

    def CreateJoint(self, joint_def):
        """JointAPI.Create{Distance,Motor,Filter,Prismatic,Revolute,Weld,Wheel}Joint api/Joint.cs:182,324,408,412,555,672,720"""
        out = C.c_int32(-1)
        self._check(self.fns["create_joint"](self.handle, C.byref(joint_def), C.byref(out)))
        return out.value

    def _typed(self, joint_def, joint_type):
        assert joint_def.type == joint_type, "joint definition built for another joint type"
        return self.CreateJoint(joint_def)

    def CreateDistanceJoint(self, joint_def):
        return self._typed(joint_def, capi.JOINT_DISTANCE)

    def CreateMotorJoint(self, joint_def):
        return self._typed(joint_def, capi.JOINT_MOTOR)

    def CreateFilterJoint(self, joint_def):
        return self._typed(joint_def, capi.JOINT_FILTER)

    def CreatePrismaticJoint(self, joint_def):
        return self._typed(joint_def, capi.JOINT_PRISMATIC)

    def CreateWeldJoint(self, joint_def):
        return self._typed(joint_def, capi.JOINT_WELD)

    def CreateWheelJoint(self, joint_def):
        return self._typed(joint_def, capi.JOINT_WHEEL)

    # Joint setters (api/Joint.cs). The C# shim maps each named setter onto one of these three calls, e.g.
    # RevoluteJoint_SetMotorSpeed(j, v) -> JointSetParam(j, JP_MOTOR_SPEED, v); WheelJoint_EnableMotor(j, f) -> JointEnable(j, JOINT_MOTOR_SWITCH, f);
    # PrismaticJoint_SetLimits / DistanceJoint_SetLengthRange -> JointSetLimits(j, lower, upper).
    def JointSetParam(self, joint_id, param, value):
        self._check(self.fns["joint_set_param"](self.handle, joint_id, param, value))

    def JointEnable(self, joint_id, feature, flag):
        self._check(self.fns["joint_enable"](self.handle, joint_id, feature, int(bool(flag))))

    def JointSetLimits(self, joint_id, lower, upper):
        self._check(self.fns["joint_set_limits"](self.handle, joint_id, lower, upper))

    def SetLocalFrameA(self, joint_id, frame):          # JointAPI.SetLocalFrameA api/Joint.cs:44
        self._check(self.fns["joint_set_local_frame"](self.handle, joint_id, 0, frame))

    def SetLocalFrameB(self, joint_id, frame):          # JointAPI.SetLocalFrameB api/Joint.cs:50
        self._check(self.fns["joint_set_local_frame"](self.handle, joint_id, 1, frame))

    def SetCollideConnected(self, joint_id, should_collide):   # JointAPI.SetCollideConnected api/Joint.cs:56
        self._check(self.fns["joint_set_collide_connected"](self.handle, joint_id, int(bool(should_collide))))

    # BodyAPI sleep / enable / bullet surface (api/Body.cs:664-842)
    def SetSleepThreshold(self, body_id, v):
        self._check(self.fns["body_set_sleep_threshold"](self.handle, body_id, v))

    def EnableSleep(self, body_id, flag):
        self._check(self.fns["body_enable_sleep"](self.handle, body_id, int(bool(flag))))

    def SetAwake(self, body_id, awake):
        self._check(self.fns["body_set_awake"](self.handle, body_id, int(bool(awake))))

    def WakeTouching(self, body_id):
        self._check(self.fns["body_wake_touching"](self.handle, body_id))

    def DisableBody(self, body_id):
        self._check(self.fns["body_disable"](self.handle, body_id))

    def EnableBody(self, body_id):
        self._check(self.fns["body_enable"](self.handle, body_id))

    def SetBullet(self, body_id, flag):
        self._check(self.fns["body_set_bullet"](self.handle, body_id, int(bool(flag))))

    def SetConstraintTuning(self, joint_id, hertz, damping_ratio):
        self.JointSetParam(joint_id, capi.JP_CONSTRAINT_HERTZ, hertz)
        self.JointSetParam(joint_id, capi.JP_CONSTRAINT_DAMPING_RATIO, damping_ratio)

    def MotorJoint_SetLinearVelocity(self, joint_id, v):
        self.JointSetParam(joint_id, capi.JP_LINEAR_VELOCITY_X, v[0])
        self.JointSetParam(joint_id, capi.JP_LINEAR_VELOCITY_Y, v[1])

    # --- WorldAPI ---
    def SetGravity(self, g):
        self._check(self.fns["world_set_gravity"](self.handle, Vec2(*g)))

    def EnableWarmStarting(self, flag):
        self._check(self.fns["world_enable_warm_starting"](self.handle, int(flag)))

    def EnableSpeculative(self, flag):
        self._check(self.fns["world_enable_speculative"](self.handle, int(flag)))

    def EnableContinuous(self, flag):
        self._check(self.fns["world_enable_continuous"](self.handle, int(flag)))

    def SetContactTuning(self, hertz, damping, push):
        self._check(self.fns["world_set_contact_tuning"](self.handle, hertz, damping, push))

    def SetRestitutionThreshold(self, value):
        self._check(self.fns["world_set_restitution_threshold"](self.handle, float(value)))

    def SetHitEventThreshold(self, value):
        self._check(self.fns["world_set_hit_event_threshold"](self.handle, float(value)))

    def SetMaximumLinearSpeed(self, value):
        self._check(self.fns["world_set_maximum_linear_speed"](self.handle, float(value)))

    def Step(self, time_step, sub_step_count):
        self._check(self.fns["world_step"](self.handle, float(time_step), int(sub_step_count)))

    def Synchronize(self):
        self._check(self.fns["world_synchronize"](self.handle))

    def Destroy(self):
        if self.handle is not None:
            self.fns["world_destroy"](self.handle)
            self.handle = None

    def GetCounters(self):
        c = Counters()
        self._check(self.fns["world_get_counters"](self.handle, C.byref(c)))
        return c

    def GetProfile(self):
        p = Profile()
        self._check(self.fns["world_get_profile"](self.handle, C.byref(p)))
        return p

    def DebugChecksums(self):
        """Per-phase state checksums of the last step (zeros unless B2CUDA_DEBUG_CHECKSUMS=1 was set at the first step)."""
        out = (C.c_uint64 * 16)()
        self._check(self.fns["world_debug_checksums"](self.handle, out))
        return list(out)

    def EnableGraphLaunch(self, flag=True):
        """(no reference counterpart) one CUDA-graph launch per step; see include/b2cuda.h"""
        self._check(self.fns["world_enable_graph_launch"](self.handle, int(bool(flag))))

    def TimerStart(self):
        self._check(self.fns["world_timer_start"](self.handle))

    def TimerStop(self):
        ms = C.c_float()
        self._check(self.fns["world_timer_stop"](self.handle, C.byref(ms)))
        return ms.value

    def GetContactEvents(self):
        """Returns (begin, end, hit) as numpy structured arrays (copies)."""
        b, e, h = C.POINTER(ContactTouchEvent)(), C.POINTER(ContactTouchEvent)(), C.POINTER(ContactHitEvent)()
        nb, ne, nh = C.c_int32(), C.c_int32(), C.c_int32()
        self._check(self.fns["world_get_contact_events"](self.handle, C.byref(b), C.byref(nb), C.byref(e), C.byref(ne), C.byref(h), C.byref(nh)))

        return _records(b, nb.value, ContactTouchEvent, True), _records(e, ne.value, ContactTouchEvent, True), _records(h, nh.value, ContactHitEvent, True)

    def GetSensorEvents(self):
        """WorldAPI.GetSensorEvents api/World.cs:289-299 -> (begin, end) numpy structured arrays of (sensor, visitor) shape references."""
        b, e = C.POINTER(capi.SensorEvent)(), C.POINTER(capi.SensorEvent)()
        nb, ne = C.c_int32(), C.c_int32()
        self._check(self.fns["world_get_sensor_events"](self.handle, C.byref(b), C.byref(nb), C.byref(e), C.byref(ne)))

        def grab(ptr, n):
            return _records(ptr, n, capi.SensorEvent, True)
        return grab(b, nb.value), grab(e, ne.value)

    def EnableSensorEvents(self, shape_id, flag):       # ShapeAPI.EnableSensorEvents api/Shape.cs:170
        self._check(self.fns["shape_enable_sensor_events"](self.handle, shape_id, int(bool(flag))))

    def GetSensorOverlaps(self, shape_id, capacity=64):  # ShapeAPI.GetSensorData api/Shape.cs:409
        out, n = (capi.ShapeRef * capacity)(), C.c_int32()
        self._check(self.fns["shape_get_sensor_overlaps"](self.handle, shape_id, out, capacity, C.byref(n)))
        return [(out[i].shapeId, out[i].generation) for i in range(n.value)]

    def GetJointEvents(self):
        p, n = C.POINTER(C.c_int32)(), C.c_int32()
        self._check(self.fns["world_get_joint_events"](self.handle, C.byref(p), C.byref(n)))
        return np.array([p[i] for i in range(n.value)], dtype=np.int32)

    def _read(self, name, ty, hint):
        n = C.c_int32()
        cap = max(hint, 16)
        while True:
            buf = (ty * cap)()
            self._check(self.fns[name](self.handle, buf, cap, C.byref(n)))
            if n.value <= cap:
                return np.frombuffer(buf, dtype=_dt(ty), count=n.value).copy()
            cap = n.value

    def GetBodyEvents(self, copy=True):
        """WorldAPI.GetBodyEvents: one record per body id. copy=False returns a view of the library's pinned buffer (valid until the next Step)."""
        p, n = C.POINTER(BodyRead)(), C.c_int32()
        self._check(self.fns["world_get_body_events"](self.handle, C.byref(p), C.byref(n)))
        return _records(p, n.value, BodyRead, copy)

    def GetBodyMoveEvents(self, copy=True):
        """WorldAPI.GetBodyEvents as the reference fills it: one (transform, bodyId, fellAsleep) per awake body, awake-set order."""
        p, n = C.POINTER(capi.BodyMoveEvent)(), C.c_int32()
        self._check(self.fns["world_get_body_move_events"](self.handle, C.byref(p), C.byref(n)))
        return _records(p, n.value, capi.BodyMoveEvent, copy)

    def CopyBodyMoveEventsNoAlloc(self, out):
        """b2cuda_world_copy_body_move_events_noalloc: the move events into the caller's array `out` (numpy, dtype BodyMoveEvent) through
        the library's fixed device scratch, with no allocation anywhere in the library. Returns the number of records written."""
        n = C.c_int32()
        rc = self.fns["world_copy_body_move_events_noalloc"](self.handle, out.ctypes.data_as(C.POINTER(capi.BodyMoveEvent)), len(out), C.byref(n))
        self._check(rc)
        return n.value

    def ReadBodies(self):
        return self._read("world_read_bodies", BodyRead, self.GetCounters().bodyCount + 16)

    def ReadContacts(self):
        return self._read("world_read_contacts", ContactRead, self.GetCounters().contactCount + 16)

    def ReadJoints(self):
        return self._read("world_read_joints", JointRead, self.GetCounters().jointCount + 16)

    def ReadShapes(self):
        return self._read("world_read_shapes", ShapeRead, self.GetCounters().shapeCount + 16)


class Snapshot:
    """Records CreateBody / CreateShape / CreateJoint calls (ids returned are ORDINALS inside the recording) and replays them into a world
    with ONE b2cuda_world_upload_snapshot call; a recording can be uploaded into any number of worlds (batched worlds of one topology)."""

    def __init__(self):
        self.parts, self.n_ops, self.n_bodies, self.n_shapes, self.n_joints = [], 0, 0, 0, 0
        self._blob = None

    def _rec(self, kind, ref, *payload):
        self.parts.append(bytes((C.c_int32 * 2)(kind, ref)) + b"".join(bytes(p) for p in payload))
        self.n_ops += 1
        self._blob = None

    def CreateBody(self, body_def):
        self._rec(0, 0, body_def)
        self.n_bodies += 1
        return self.n_bodies - 1

    def CreateShape(self, body, shape_def, geometry):
        self._rec(1, body, shape_def, geometry)
        self.n_shapes += 1
        return self.n_shapes - 1

    CreatePolygonShape = CreateCircleShape = CreateCapsuleShape = CreateSegmentShape = CreateShape

    def CreateJoint(self, joint_def):
        self._rec(2, 0, joint_def)
        self.n_joints += 1
        return self.n_joints - 1

    def CreateRevoluteJoint(self, r):
        """api/Joint.cs:555: the revolute definition goes down the generic joint path (as b2cuda_create_revolute_joint does)"""
        d = JointDef()
        d.type = capi.JOINT_REVOLUTE
        for f in ("bodyIdA", "bodyIdB", "localFrameA", "localFrameB", "forceThreshold", "torqueThreshold", "constraintHertz", "constraintDampingRatio",
                  "drawScale", "collideConnected", "userData", "enableSpring", "enableLimit", "enableMotor", "hertz", "dampingRatio", "motorSpeed"):
            setattr(d, f, getattr(r, f))
        d.target, d.lower, d.upper, d.maxMotor = r.targetAngle, r.lowerAngle, r.upperAngle, r.maxMotorTorque
        return self.CreateJoint(d)

    def blob(self):
        if self._blob is None:
            hd = capi.SnapshotHeader(capi.SNAPSHOT_MAGIC, 1, self.n_ops, self.n_bodies, self.n_shapes, self.n_joints)
            self._blob = bytes(hd) + b"".join(self.parts)
        return self._blob

    def upload(self, world):
        """Returns (body ids, shape ids, joint ids) as the world assigned them, in recording order."""
        data = self.blob()
        b, s, j = (C.c_int32 * max(self.n_bodies, 1))(), (C.c_int32 * max(self.n_shapes, 1))(), (C.c_int32 * max(self.n_joints, 1))()
        world._check(world.fns["world_upload_snapshot"](world.handle, data, len(data), b, s, j))
        return list(b)[:self.n_bodies], list(s)[:self.n_shapes], list(j)[:self.n_joints]


class Batch:
    """b2cuda_batch_*: N independent worlds on one device, stepped in lockstep by one launch (api/World.cs:9-10).
    `worlds[k]` is an ordinary World facade on the handle of world k (build it, read it back, mutate it between steps)."""

    def __init__(self, backend, n_worlds, world_def=None):
        self.backend, self.fns = backend, backend.fns
        self.handle = C.c_void_p()
        wd = world_def if world_def is not None else default_world_def()
        rc = self.fns["batch_create"](C.byref(wd), getattr(backend, "device", 0), n_worlds, C.byref(self.handle))
        if rc != 0:
            raise B2Error(f"b2cuda_batch_create failed rc={rc}: {self.fns['last_error']().decode()}")
        self.worlds = []
        for k in range(n_worlds):
            h = C.c_void_p()
            self._check(self.fns["batch_world"](self.handle, k, C.byref(h)))
            w = World.__new__(World)
            w.backend, w.fns, w.handle = backend, backend.fns, h
            w.Destroy = lambda: None            # owned by the batch
            self.worlds.append(w)

    def _check(self, rc):
        if rc != 0:
            raise B2Error(f"cuda batch rc={rc}: {self.fns['last_error']().decode()}")

    def Step(self, dt, substeps):
        self._check(self.fns["batch_step"](self.handle, dt, substeps))

    def Synchronize(self):
        self._check(self.fns["batch_synchronize"](self.handle))

    def ResetWorld(self, index):
        self._check(self.fns["batch_reset_world"](self.handle, index))

    def Stats(self, device_ptr=None):
        """Statistics reduced over all worlds on the device; device_ptr (e.g. a torch float64[8] tensor's data_ptr()) also receives them."""
        s = capi.BatchStats()
        self._check(self.fns["batch_stats"](self.handle, C.byref(s), C.c_void_p(device_ptr) if device_ptr else None))
        return s

    def TimerStart(self):
        self._check(self.fns["batch_timer_start"](self.handle))

    def TimerStop(self):
        ms = C.c_float()
        self._check(self.fns["batch_timer_stop"](self.handle, C.byref(ms)))
        return ms.value

    def Destroy(self):
        if self.handle is not None:
            self.fns["batch_destroy"](self.handle)
            self.handle = None
