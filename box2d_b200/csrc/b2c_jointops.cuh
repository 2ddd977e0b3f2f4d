// libb2cuda — joint setters of api/Joint.cs as one patch function shared by the host mirror (before the upload) and the one-thread
// device kernel (after it). Returns false when the field does not exist for the joint's type (the reference asserts the type).
#pragma once
#include "b2c_world.cuh"
#include "../../include/b2cuda.h"

namespace b2c {

enum JointPatch { JPATCH_PARAM = 0, JPATCH_ENABLE = 1, JPATCH_LIMITS = 2 };

B2C_HD bool jointHasSpringLimitMotor(int type, int feature) {
  if (type == kRevoluteJoint || type == kPrismaticJoint || type == kWheelJoint || type == kDistanceJoint) return feature >= 0 && feature <= 2;
  return false;
}
B2C_HD bool jointParamValid(int type, int param) {
  const bool slm = type == kRevoluteJoint || type == kPrismaticJoint || type == kWheelJoint || type == kDistanceJoint;
  switch (param) {
    case B2C_JP_CONSTRAINT_HERTZ: case B2C_JP_CONSTRAINT_DAMPING_RATIO: case B2C_JP_FORCE_THRESHOLD: case B2C_JP_TORQUE_THRESHOLD: return true;
    case B2C_JP_SPRING_HERTZ: case B2C_JP_SPRING_DAMPING_RATIO: return slm || type == kWeldJoint || type == kMotorJoint;
    case B2C_JP_TARGET: return type == kRevoluteJoint || type == kPrismaticJoint;
    case B2C_JP_MOTOR_SPEED: case B2C_JP_MAX_MOTOR: return slm;
    case B2C_JP_LENGTH: case B2C_JP_LOWER_SPRING_FORCE: case B2C_JP_UPPER_SPRING_FORCE: return type == kDistanceJoint;
    case B2C_JP_ANGULAR_HERTZ: case B2C_JP_ANGULAR_DAMPING_RATIO: return type == kWeldJoint || type == kMotorJoint;
    case B2C_JP_LINEAR_VELOCITY_X: case B2C_JP_LINEAR_VELOCITY_Y: case B2C_JP_ANGULAR_VELOCITY: case B2C_JP_MAX_VELOCITY_FORCE:
    case B2C_JP_MAX_VELOCITY_TORQUE: case B2C_JP_MAX_SPRING_FORCE: case B2C_JP_MAX_SPRING_TORQUE: return type == kMotorJoint;
    default: return false;
  }
}
B2C_HD void jointPatch(JointDev& j, int patch, int which, float a, float b) {
  if (patch == JPATCH_PARAM) {
    switch (which) {
      case B2C_JP_CONSTRAINT_HERTZ: j.constraintHertz = a; break;                       // api/Joint.cs:130-138
      case B2C_JP_CONSTRAINT_DAMPING_RATIO: j.constraintDampingRatio = a; break;
      case B2C_JP_FORCE_THRESHOLD: j.forceThreshold = a; break;                         // :148-153
      case B2C_JP_TORQUE_THRESHOLD: j.torqueThreshold = a; break;                       // :159-164
      case B2C_JP_SPRING_HERTZ: j.hertz = a; break;
      case B2C_JP_SPRING_DAMPING_RATIO: j.dampingRatio = a; break;
      case B2C_JP_TARGET: j.targetAngle = a; break;                                     // :444, :585 (no clamp in the setter)
      case B2C_JP_MOTOR_SPEED: j.motorSpeed = a; break;
      case B2C_JP_MAX_MOTOR: j.maxMotorTorque = a; break;
      case B2C_JP_LENGTH: j.length = netClamp(a, kLinearSlop, kHuge); j.lowerImpulse = 0.0f; j.upperImpulse = 0.0f; break;   // :187-194
      case B2C_JP_LOWER_SPRING_FORCE: j.lowerSpringForce = a; break;                    // :204-210
      case B2C_JP_UPPER_SPRING_FORCE: j.upperSpringForce = a; break;
      case B2C_JP_ANGULAR_HERTZ: j.angularHertz = a; break;
      case B2C_JP_ANGULAR_DAMPING_RATIO: j.angularDampingRatio = a; break;
      case B2C_JP_LINEAR_VELOCITY_X: j.linearVelocity.x = a; break;                     // :327
      case B2C_JP_LINEAR_VELOCITY_Y: j.linearVelocity.y = a; break;
      case B2C_JP_ANGULAR_VELOCITY: j.angularVelocity = a; break;
      case B2C_JP_MAX_VELOCITY_FORCE: j.maxVelocityForce = netMax(0.0f, a); break;      // :333
      case B2C_JP_MAX_VELOCITY_TORQUE: j.maxVelocityTorque = a; break;                  // :336 (no clamp)
      case B2C_JP_MAX_SPRING_FORCE: j.maxSpringForce = netMax(0.0f, a); break;
      case B2C_JP_MAX_SPRING_TORQUE: j.maxSpringTorque = netMax(0.0f, a); break;
    }
  } else if (patch == JPATCH_ENABLE) {
    const int flag = a != 0.0f;
    if (which == B2C_JOINT_SPRING) {
      if (j.type == kDistanceJoint) j.enableSpring = flag;                              // :197 plain store
      else if (flag != j.enableSpring) { j.enableSpring = flag; j.springImpulse = 0.0f; }
    } else if (which == B2C_JOINT_LIMIT) {
      if (j.type == kDistanceJoint) j.enableLimit = flag;                               // :229 plain store
      else if (flag != j.enableLimit) { j.enableLimit = j.type == kWheelJoint ? 0 : flag; j.lowerImpulse = 0.0f; j.upperImpulse = 0.0f; }   // wheel :755 as written
    } else {
      if (flag != j.enableMotor) { j.enableMotor = flag; j.motorImpulse = 0.0f; }
    }
  } else {
    if (j.type == kDistanceJoint) {                                                     // DistanceJoint_SetLengthRange :236-247
      float minLength = netClamp(a, kLinearSlop, kHuge), maxLength = netClamp(b, kLinearSlop, kHuge);
      j.minLength = netMin(minLength, maxLength); j.maxLength = netMax(minLength, maxLength);
      j.springImpulse = 0.0f; j.lowerImpulse = 0.0f; j.upperImpulse = 0.0f;
    } else if (a != j.lowerAngle || b != j.upperAngle) {                                // :471-481, :626-637, :769-780
      j.lowerAngle = netMin(a, b); j.upperAngle = netMax(a, b);
      j.lowerImpulse = 0.0f; j.upperImpulse = 0.0f;
    }
  }
}

}  // namespace b2c
