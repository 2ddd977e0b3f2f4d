"""ctypes mirror of include/b2cuda.h (POD structs + function prototypes).

The same prototypes are bound twice: on libb2cuda.so with prefix ``b2cuda_`` (the product) and — by the
tests only — on the oracle library with prefix ``b2o_``. Nothing in this package loads the oracle.
"""
import ctypes as C
import os

MAX_POLY = 8


class Vec2(C.Structure):
    _fields_ = [("x", C.c_float), ("y", C.c_float)]


class Rot(C.Structure):
    _fields_ = [("c", C.c_float), ("s", C.c_float)]


class Transform(C.Structure):
    _fields_ = [("p", Vec2), ("q", Rot)]


class WorldDef(C.Structure):
    _fields_ = [("gravity", Vec2), ("restitutionThreshold", C.c_float), ("hitEventThreshold", C.c_float),
                ("contactHertz", C.c_float), ("contactDampingRatio", C.c_float), ("contactSpeed", C.c_float),
                ("maximumLinearSpeed", C.c_float), ("enableSleep", C.c_int32), ("enableContinuous", C.c_int32),
                ("enableContactSoftening", C.c_int32), ("bodyCapacity", C.c_int32), ("shapeCapacity", C.c_int32),
                ("contactCapacity", C.c_int32), ("jointCapacity", C.c_int32)]


class BodyDef(C.Structure):
    _fields_ = [("type", C.c_int32), ("position", Vec2), ("rotation", Rot), ("linearVelocity", Vec2),
                ("angularVelocity", C.c_float), ("linearDamping", C.c_float), ("angularDamping", C.c_float),
                ("gravityScale", C.c_float), ("sleepThreshold", C.c_float), ("userData", C.c_uint64),
                ("lockLinearX", C.c_int32), ("lockLinearY", C.c_int32), ("lockAngularZ", C.c_int32),
                ("enableSleep", C.c_int32), ("isAwake", C.c_int32), ("isBullet", C.c_int32), ("isEnabled", C.c_int32),
                ("allowFastRotation", C.c_int32)]


class ShapeDef(C.Structure):
    _fields_ = [("userData", C.c_uint64), ("friction", C.c_float), ("restitution", C.c_float),
                ("rollingResistance", C.c_float), ("tangentSpeed", C.c_float), ("userMaterialId", C.c_uint64),
                ("density", C.c_float), ("groupIndex", C.c_int32), ("categoryBits", C.c_uint64), ("maskBits", C.c_uint64),
                ("enableCustomFiltering", C.c_int32), ("isSensor", C.c_int32), ("enableSensorEvents", C.c_int32),
                ("enableContactEvents", C.c_int32), ("enableHitEvents", C.c_int32), ("enablePreSolveEvents", C.c_int32),
                ("invokeContactCreation", C.c_int32), ("updateBodyMass", C.c_int32)]


class Geometry(C.Structure):
    _fields_ = [("type", C.c_int32), ("count", C.c_int32), ("radius", C.c_float), ("v", Vec2 * MAX_POLY),
                ("n", Vec2 * MAX_POLY), ("centroid", Vec2)]


class ShapeRef(C.Structure):
    _fields_ = [("shapeId", C.c_int32), ("generation", C.c_uint32)]


class SensorEvent(C.Structure):
    _fields_ = [("sensor", ShapeRef), ("visitor", ShapeRef)]


class RevoluteJointDef(C.Structure):
    _fields_ = [("bodyIdA", C.c_int32), ("bodyIdB", C.c_int32), ("localFrameA", Transform), ("localFrameB", Transform),
                ("forceThreshold", C.c_float), ("torqueThreshold", C.c_float), ("constraintHertz", C.c_float),
                ("constraintDampingRatio", C.c_float), ("drawScale", C.c_float), ("collideConnected", C.c_int32),
                ("userData", C.c_uint64), ("targetAngle", C.c_float), ("enableSpring", C.c_int32), ("hertz", C.c_float),
                ("dampingRatio", C.c_float), ("enableLimit", C.c_int32), ("lowerAngle", C.c_float), ("upperAngle", C.c_float),
                ("enableMotor", C.c_int32), ("maxMotorTorque", C.c_float), ("motorSpeed", C.c_float)]


# B2C_JP_* / B2C_JOINT_{SPRING,LIMIT,MOTOR} of include/b2cuda.h
(JP_CONSTRAINT_HERTZ, JP_CONSTRAINT_DAMPING_RATIO, JP_FORCE_THRESHOLD, JP_TORQUE_THRESHOLD, JP_SPRING_HERTZ, JP_SPRING_DAMPING_RATIO, JP_TARGET,
 JP_MOTOR_SPEED, JP_MAX_MOTOR, JP_LENGTH, JP_LOWER_SPRING_FORCE, JP_UPPER_SPRING_FORCE, JP_ANGULAR_HERTZ, JP_ANGULAR_DAMPING_RATIO,
 JP_LINEAR_VELOCITY_X, JP_LINEAR_VELOCITY_Y, JP_ANGULAR_VELOCITY, JP_MAX_VELOCITY_FORCE, JP_MAX_VELOCITY_TORQUE, JP_MAX_SPRING_FORCE,
 JP_MAX_SPRING_TORQUE) = range(21)
SHAPE_CONTACT_EVENTS, SHAPE_HIT_EVENTS = 0, 1
JOINT_SPRING, JOINT_LIMIT, JOINT_MOTOR_SWITCH = range(3)
JOINT_DISTANCE, JOINT_FILTER, JOINT_MOTOR, JOINT_PRISMATIC, JOINT_REVOLUTE, JOINT_WELD, JOINT_WHEEL = range(7)


class JointDef(C.Structure):
    _fields_ = [("type", C.c_int32), ("bodyIdA", C.c_int32), ("bodyIdB", C.c_int32), ("localFrameA", Transform),
                ("localFrameB", Transform), ("forceThreshold", C.c_float), ("torqueThreshold", C.c_float),
                ("constraintHertz", C.c_float), ("constraintDampingRatio", C.c_float), ("drawScale", C.c_float),
                ("collideConnected", C.c_int32), ("userData", C.c_uint64), ("enableSpring", C.c_int32),
                ("enableLimit", C.c_int32), ("enableMotor", C.c_int32), ("hertz", C.c_float), ("dampingRatio", C.c_float),
                ("target", C.c_float), ("lower", C.c_float), ("upper", C.c_float), ("maxMotor", C.c_float),
                ("motorSpeed", C.c_float), ("length", C.c_float), ("minLength", C.c_float), ("maxLength", C.c_float),
                ("lowerSpringForce", C.c_float), ("upperSpringForce", C.c_float), ("angularHertz", C.c_float),
                ("angularDampingRatio", C.c_float), ("linearVelocity", Vec2), ("angularVelocity", C.c_float),
                ("maxVelocityForce", C.c_float), ("maxVelocityTorque", C.c_float), ("maxSpringForce", C.c_float),
                ("maxSpringTorque", C.c_float)]


class BodyRead(C.Structure):
    _fields_ = [("transform", Transform), ("center", Vec2), ("linearVelocity", Vec2), ("angularVelocity", C.c_float),
                ("flags", C.c_uint32)]


class ManifoldPointRead(C.Structure):
    _fields_ = [("point", Vec2), ("anchorA", Vec2), ("anchorB", Vec2), ("separation", C.c_float),
                ("normalImpulse", C.c_float), ("tangentImpulse", C.c_float), ("totalNormalImpulse", C.c_float),
                ("normalVelocity", C.c_float), ("id", C.c_uint32), ("persisted", C.c_uint32)]


class ContactRead(C.Structure):
    _fields_ = [("contactId", C.c_int32), ("shapeIdA", C.c_int32), ("shapeIdB", C.c_int32), ("colorIndex", C.c_int32),
                ("localIndex", C.c_int32), ("flags", C.c_uint32), ("pointCount", C.c_int32), ("normal", Vec2), ("rollingImpulse", C.c_float),
                ("friction", C.c_float), ("restitution", C.c_float), ("rollingResistance", C.c_float),
                ("tangentSpeed", C.c_float), ("points", ManifoldPointRead * 2)]


class ContactTouchEvent(C.Structure):
    _fields_ = [("shapeIdA", C.c_int32), ("shapeIdB", C.c_int32), ("contactId", C.c_int32), ("contactGeneration", C.c_uint32)]


class ContactHitEvent(C.Structure):
    _fields_ = [("shapeIdA", C.c_int32), ("shapeIdB", C.c_int32), ("point", Vec2), ("normal", Vec2), ("approachSpeed", C.c_float)]


class JointRead(C.Structure):
    _fields_ = [("jointId", C.c_int32), ("colorIndex", C.c_int32), ("linearImpulse", Vec2), ("springImpulse", C.c_float),
                ("motorImpulse", C.c_float), ("lowerImpulse", C.c_float), ("upperImpulse", C.c_float),
                ("type", C.c_int32), ("linearImpulse2", Vec2)]


class ShapeRead(C.Structure):
    _fields_ = [("aabbLo", Vec2), ("aabbHi", Vec2), ("fatLo", Vec2), ("fatHi", Vec2), ("proxyKey", C.c_int32), ("enlarged", C.c_int32)]


class Counters(C.Structure):
    _fields_ = [(n, C.c_int32) for n in ("bodyCount", "awakeBodyCount", "shapeCount", "contactCount", "touchingContactCount",
                                         "jointCount", "contactIdCapacity", "activeColorCount", "moveCount", "newPairCount",
                                         "toiCandidates", "unsupportedCount", "stepCount", "errorFlags")] + \
               [("colorCounts", C.c_int32 * 24)]


class BodyMoveEvent(C.Structure):
    _fields_ = [("transform", Transform), ("bodyId", C.c_int32), ("fellAsleep", C.c_uint32)]


class Profile(C.Structure):
    _fields_ = [(n, C.c_float) for n in ("step", "pairs", "collide", "solve", "prepareStages", "solveConstraints",
                                         "prepareConstraints", "integrateVelocities", "warmStart", "solveImpulses",
                                         "integratePositions", "relaxImpulses", "applyRestitution", "storeImpulses",
                                         "transforms", "refit", "readback", "treeRebuild")]


# name -> (argtypes after the world handle, has world handle)
_W = C.c_void_p
PROTOS = {
    "create_body": [_W, C.POINTER(BodyDef), C.POINTER(C.c_int32)],
    "create_shape": [_W, C.c_int32, C.POINTER(ShapeDef), C.POINTER(Geometry), C.POINTER(C.c_int32)],
    "create_revolute_joint": [_W, C.POINTER(RevoluteJointDef), C.POINTER(C.c_int32)],
    "create_joint": [_W, C.POINTER(JointDef), C.POINTER(C.c_int32)],
    "destroy_joint": [_W, C.c_int32, C.c_int32],
    "destroy_body": [_W, C.c_int32],
    "destroy_shape": [_W, C.c_int32, C.c_int32],
    "shape_set_density": [_W, C.c_int32, C.c_float, C.c_int32],
    "body_apply_mass_from_shapes": [_W, C.c_int32],
    "shape_enable_events": [_W, C.c_int32, C.c_int32, C.c_int32],
    "body_enable_events": [_W, C.c_int32, C.c_int32, C.c_int32],
    "body_set_type": [_W, C.c_int32, C.c_int32],
    "world_explode": [_W, Vec2, C.c_float, C.c_float, C.c_float, C.c_uint64],
    "shape_set_filter": [_W, C.c_int32, C.c_uint64, C.c_uint64, C.c_int32],
    "shape_set_geometry": [_W, C.c_int32, C.POINTER(Geometry)],
    "joint_set_param": [_W, C.c_int32, C.c_int32, C.c_float],
    "joint_enable": [_W, C.c_int32, C.c_int32, C.c_int32],
    "joint_set_limits": [_W, C.c_int32, C.c_float, C.c_float],
    "joint_set_local_frame": [_W, C.c_int32, C.c_int32, Transform],
    "joint_set_collide_connected": [_W, C.c_int32, C.c_int32],
    "body_set_transform": [_W, C.c_int32, Vec2, Rot],
    "body_set_linear_velocity": [_W, C.c_int32, Vec2],
    "body_set_target_transform": [_W, C.c_int32, Transform, C.c_float],
    "body_set_mass_data": [_W, C.c_int32, C.c_float, C.c_float, Vec2],
    "body_set_motion_locks": [_W, C.c_int32, C.c_int32, C.c_int32, C.c_int32],
    "shape_set_surface_material": [_W, C.c_int32, C.c_float, C.c_float, C.c_float, C.c_float],
    "body_set_angular_velocity": [_W, C.c_int32, C.c_float],
    "body_apply_force_to_center": [_W, C.c_int32, Vec2, C.c_int32],
    "body_apply_torque": [_W, C.c_int32, C.c_float, C.c_int32],
    "body_apply_linear_impulse_to_center": [_W, C.c_int32, Vec2, C.c_int32],
    "body_apply_force": [_W, C.c_int32, Vec2, Vec2, C.c_int32],
    "body_clear_forces": [_W, C.c_int32],
    "body_apply_linear_impulse": [_W, C.c_int32, Vec2, Vec2, C.c_int32],
    "body_apply_angular_impulse": [_W, C.c_int32, C.c_float, C.c_int32],
    "body_set_linear_damping": [_W, C.c_int32, C.c_float],
    "body_set_angular_damping": [_W, C.c_int32, C.c_float],
    "body_set_gravity_scale": [_W, C.c_int32, C.c_float],
    "body_get": [_W, C.c_int32, C.POINTER(BodyRead)],
    "world_set_restitution_threshold": [_W, C.c_float],
    "world_set_hit_event_threshold": [_W, C.c_float],
    "world_set_maximum_linear_speed": [_W, C.c_float],
    "world_set_gravity": [_W, Vec2],
    "world_enable_warm_starting": [_W, C.c_int32],
    "world_enable_speculative": [_W, C.c_int32],
    "world_enable_continuous": [_W, C.c_int32],
    "world_enable_sleeping": [_W, C.c_int32],
    "world_set_contact_tuning": [_W, C.c_float, C.c_float, C.c_float],
    "body_set_bullet": [_W, C.c_int32, C.c_int32],
    "world_step": [_W, C.c_float, C.c_int32],
    "world_synchronize": [_W],
    "world_destroy": [_W],
    "world_get_contact_events": [_W, C.POINTER(C.POINTER(ContactTouchEvent)), C.POINTER(C.c_int32),
                                 C.POINTER(C.POINTER(ContactTouchEvent)), C.POINTER(C.c_int32),
                                 C.POINTER(C.POINTER(ContactHitEvent)), C.POINTER(C.c_int32)],
    "world_get_sensor_events": [_W, C.POINTER(C.POINTER(SensorEvent)), C.POINTER(C.c_int32), C.POINTER(C.POINTER(SensorEvent)), C.POINTER(C.c_int32)],
    "shape_enable_sensor_events": [_W, C.c_int32, C.c_int32],
    "shape_get_sensor_overlaps": [_W, C.c_int32, C.POINTER(ShapeRef), C.c_int32, C.POINTER(C.c_int32)],
    "world_get_joint_events": [_W, C.POINTER(C.POINTER(C.c_int32)), C.POINTER(C.c_int32)],
    "world_get_body_events": [_W, C.POINTER(C.POINTER(BodyRead)), C.POINTER(C.c_int32)],
    "world_get_body_move_events": [_W, C.POINTER(C.POINTER(BodyMoveEvent)), C.POINTER(C.c_int32)],
    "world_read_bodies": [_W, C.POINTER(BodyRead), C.c_int32, C.POINTER(C.c_int32)],
    "world_read_contacts": [_W, C.POINTER(ContactRead), C.c_int32, C.POINTER(C.c_int32)],
    "world_read_joints": [_W, C.POINTER(JointRead), C.c_int32, C.POINTER(C.c_int32)],
    "world_read_shapes": [_W, C.POINTER(ShapeRead), C.c_int32, C.POINTER(C.c_int32)],
    "world_get_counters": [_W, C.POINTER(Counters)],
}
# symbols only the product exports (declared in include/b2cuda.h)
PRODUCT_ONLY = {
    "last_error": ([], C.c_char_p),
    "default_world_def": ([C.POINTER(WorldDef)], C.c_int32),
    "default_body_def": ([C.POINTER(BodyDef)], C.c_int32),
    "default_shape_def": ([C.POINTER(ShapeDef)], C.c_int32),
    "default_revolute_joint_def": ([C.POINTER(RevoluteJointDef)], C.c_int32),
    "default_joint_def": ([C.c_int32, C.POINTER(JointDef)], C.c_int32),
    "world_create": ([C.POINTER(WorldDef), C.c_int32, C.POINTER(_W)], C.c_int32),
    "world_get_profile": ([_W, C.POINTER(Profile)], C.c_int32),
    "world_enable_graph_launch": ([_W, C.c_int32], C.c_int32),
    "world_timer_start": ([_W], C.c_int32),
    "world_timer_stop": ([_W, C.POINTER(C.c_float)], C.c_int32),
    "world_debug_checksums": ([_W, C.POINTER(C.c_uint64)], C.c_int32),
    "world_copy_body_move_events_noalloc": ([_W, C.POINTER(BodyMoveEvent), C.c_int32, C.POINTER(C.c_int32)], C.c_int32),
}
class BatchStats(C.Structure):
    _fields_ = [(n, C.c_double) for n in ("awakeBodies", "sumV2", "touchingContacts", "beginEvents", "endEvents", "maxSpeed", "minY", "errorFlags")]


_B = C.c_void_p
PRODUCT_ONLY.update({
    "batch_create": ([C.POINTER(WorldDef), C.c_int32, C.c_int32, C.POINTER(_B)], C.c_int32),
    "batch_destroy": ([_B], C.c_int32),
    "batch_world": ([_B, C.c_int32, C.POINTER(_W)], C.c_int32),
    "batch_step": ([_B, C.c_float, C.c_int32], C.c_int32),
    "batch_synchronize": ([_B], C.c_int32),
    "batch_reset_world": ([_B, C.c_int32], C.c_int32),
    "batch_stats": ([_B, C.POINTER(BatchStats), C.c_void_p], C.c_int32),
    "batch_timer_start": ([_B], C.c_int32),
    "batch_timer_stop": ([_B, C.POINTER(C.c_float)], C.c_int32),
})
class SnapshotHeader(C.Structure):
    _fields_ = [("magic", C.c_uint32), ("version", C.c_uint32), ("opCount", C.c_int32), ("bodyCount", C.c_int32), ("shapeCount", C.c_int32), ("jointCount", C.c_int32)]


SNAPSHOT_MAGIC = 0x42324353
PRODUCT_ONLY["world_upload_snapshot"] = ([_W, C.c_void_p, C.c_size_t, C.POINTER(C.c_int32), C.POINTER(C.c_int32), C.POINTER(C.c_int32)], C.c_int32)
PRODUCT_ONLY.update({
    "body_set_sleep_threshold": ([_W, C.c_int32, C.c_float], C.c_int32),
    "body_enable_sleep": ([_W, C.c_int32, C.c_int32], C.c_int32),
    "body_set_awake": ([_W, C.c_int32, C.c_int32], C.c_int32),
    "body_wake_touching": ([_W, C.c_int32], C.c_int32),
    "body_disable": ([_W, C.c_int32], C.c_int32),
    "body_enable": ([_W, C.c_int32], C.c_int32),
})
NOALLOC_MAX_MOVE_EVENTS = 65536     # B2C_NOALLOC_MAX_MOVE_EVENTS


def bind(lib, prefix):
    """Attach argtypes/restype for every common entry point; returns {name: function}."""
    fns = {}
    for name, args in PROTOS.items():
        f = getattr(lib, prefix + name)
        f.argtypes = args
        f.restype = C.c_int32
        fns[name] = f
    return fns


def product_library_path():
    if os.environ.get("B2CUDA_LIB"):              # A/B runs of two builds of the library (scripts/ab_phases.py); still the CUDA library, never a fallback
        return os.environ["B2CUDA_LIB"]
    return os.path.join(os.path.dirname(os.path.abspath(__file__)), "csrc", "libb2cuda.so")


def exported_symbols():
    """Every symbol include/b2cuda.h declares (used by the CPU-side export test)."""
    return ["b2cuda_" + n for n in list(PROTOS) + list(PRODUCT_ONLY)]
