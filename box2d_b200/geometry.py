"""Host-side geometry helpers mirroring the reference's Geometry.* / shape constructors (fp32, same op order).

src/Geometry.cs:88-131 (MakeBox, MakeOffsetBox), include/Collision.cs:65 (Circle), :184 (Capsule), :632 (Segment).
These stay host code in the reference too (shape construction is not on the step path).
"""
import numpy as np
from .capi import Geometry, Vec2

CIRCLE, CAPSULE, SEGMENT, POLYGON = 0, 1, 2, 3
f32 = np.float32


def _tp(q, p, v):
    """Transform.TransformPoint (include/MathFunctions.cs:180) in fp32."""
    c, s = f32(q[0]), f32(q[1])
    x, y = f32(v[0]), f32(v[1])
    return (float(f32(f32(c * x) - f32(s * y)) + f32(p[0])), float(f32(f32(s * x) + f32(c * y)) + f32(p[1])))


def _rot(q, v):
    c, s = f32(q[0]), f32(q[1])
    x, y = f32(v[0]), f32(v[1])
    return (float(f32(c * x) - f32(s * y)), float(f32(s * x) + f32(c * y)))


def MakeBox(hx, hy):
    g = Geometry()
    g.type, g.count, g.radius = POLYGON, 4, 0.0
    vs = [(-hx, -hy), (hx, -hy), (hx, hy), (-hx, hy)]
    ns = [(0.0, -1.0), (1.0, 0.0), (0.0, 1.0), (-1.0, 0.0)]
    for i in range(4):
        g.v[i] = Vec2(*vs[i])
        g.n[i] = Vec2(*ns[i])
    g.centroid = Vec2(0.0, 0.0)
    return g


def MakeSquare(h):
    return MakeBox(h, h)


def MakeOffsetBox(hx, hy, center, rotation=(1.0, 0.0)):
    g = Geometry()
    g.type, g.count, g.radius = POLYGON, 4, 0.0
    vs = [(-hx, -hy), (hx, -hy), (hx, hy), (-hx, hy)]
    ns = [(0.0, -1.0), (1.0, 0.0), (0.0, 1.0), (-1.0, 0.0)]
    for i in range(4):
        g.v[i] = Vec2(*_tp(rotation, center, vs[i]))
        g.n[i] = Vec2(*_rot(rotation, ns[i]))
    g.centroid = Vec2(float(f32(center[0])), float(f32(center[1])))
    return g


def MakeRoundedBox(hx, hy, radius):
    g = MakeBox(hx, hy)
    g.radius = radius
    return g


def Circle(center, radius):
    g = Geometry()
    g.type, g.count, g.radius = CIRCLE, 1, radius
    g.v[0] = Vec2(*center)
    return g


def Capsule(c1, c2, radius):
    g = Geometry()
    g.type, g.count, g.radius = CAPSULE, 2, radius
    g.v[0] = Vec2(*c1)
    g.v[1] = Vec2(*c2)
    return g


def Segment(p1, p2):
    g = Geometry()
    g.type, g.count, g.radius = SEGMENT, 2, 0.0
    g.v[0] = Vec2(*p1)
    g.v[1] = Vec2(*p2)
    return g


def MakeRot(radians):
    """Rotation(float radians) (include/MathFunctions.cs:102). MathF.SinCos is CRT dependent (parity unpinned);
    callers that need exact parity pass explicit (c, s) pairs instead."""
    return (float(f32(np.cos(f32(radians)))), float(f32(np.sin(f32(radians)))))
