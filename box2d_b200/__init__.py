"""box2d_b200 — B200-native rigid-body step behind the Box2D.API surface."""
from . import capi, api, geometry, scenes, batch  # noqa: F401
