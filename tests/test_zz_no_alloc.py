"""-m gpu, collected last on purpose (a failure here must not hide other tests under `pytest -x`).

b2cuda_world_copy_body_move_events_noalloc (no counterpart in the reference): the body
move events through the library's fixed file-scope device scratch, with no allocation. Checked through the C-ABI against the existing
entry point b2cuda_world_get_body_move_events, bit for bit (the tolerance every move-event test here uses), on calls in a row."""
import ctypes as C
import numpy as np
import pytest
from box2d_b200 import api, capi, scenes

pytestmark = pytest.mark.gpu

DT, SUB = 1.0 / 60.0, 4
EV = np.dtype(capi.BodyMoveEvent)


def _noalloc(w, capacity):
    out = np.zeros(capacity, dtype=EV)
    out["bodyId"] = -7                      # poison: the call must overwrite exactly `count` records and leave the rest alone
    n = w.CopyBodyMoveEventsNoAlloc(out)
    assert (out["bodyId"][n:] == -7).all()
    return out[:n]


def test_noalloc_move_events_match_existing_entry_point_on_calls_in_a_row():
    w = api.World(api.CudaBackend(), scenes.world_def())
    try:
        scenes.mixed_drop(w)
        cap = w.GetCounters().awakeBodyCount + 8
        for _ in range(5):
            w.Step(DT, SUB)
        # first call: new entry point BEFORE the existing one has ever run on this world (it must not depend on its buffers)
        first = _noalloc(w, cap)
        ref = w.GetBodyMoveEvents()
        assert len(first) == len(ref) > 100 and first.tobytes() == ref.tobytes()
        # second call in a row, same state: the scratch still holds the first call's records
        again = _noalloc(w, cap)
        assert again.tobytes() == ref.tobytes()
        # a later state, then a SMALLER awake set (two bodies destroyed): stale records of the earlier calls must not show through
        w.Step(DT, SUB)
        a, b = _noalloc(w, cap), w.GetBodyMoveEvents()
        assert a.tobytes() == b.tobytes() and a.tobytes() != ref.tobytes()
        w.DestroyBody(5)
        w.DestroyBody(9)
        w.Step(DT, SUB)
        a, b = _noalloc(w, cap), w.GetBodyMoveEvents()
        assert len(a) == len(ref) - 2 and a.tobytes() == b.tobytes()
        c = _noalloc(w, cap)
        assert c.tobytes() == b.tobytes()
        # dt = 0: the reference clears the move events (api/World.cs:38,44-51)
        w.Step(0.0, SUB)
        assert len(_noalloc(w, cap)) == 0 and len(w.GetBodyMoveEvents()) == 0
    finally:
        w.Destroy()


def test_noalloc_second_world_shares_the_scratch_and_rejects_bad_sizes():
    be = api.CudaBackend()
    w1 = api.World(be, scenes.world_def())
    w2 = api.World(be, scenes.world_def())
    try:
        scenes.add_pyramid(w1, 10)
        scenes.add_pyramid(w2, 6, origin=(3.0, 0.0))
        for _ in range(3):
            w1.Step(DT, SUB)
            w2.Step(DT, SUB)
        a1, a2 = _noalloc(w1, 64), _noalloc(w2, 64)        # back to back on one GPU: the same scratch array, one call after the other
        assert a1.tobytes() == w1.GetBodyMoveEvents().tobytes() and a2.tobytes() == w2.GetBodyMoveEvents().tobytes()
        assert len(a1) == 55 and len(a2) == 21
        # destination too small: refused, nothing written
        small = np.zeros(10, dtype=EV)
        n = C.c_int32(123)
        rc = w1.fns["world_copy_body_move_events_noalloc"](w1.handle, small.ctypes.data_as(C.POINTER(capi.BodyMoveEvent)), len(small), C.byref(n))
        assert rc == -1 and n.value == 0 and "needed" in be.last_error(w1.handle)
        # null destination
        rc = w1.fns["world_copy_body_move_events_noalloc"](w1.handle, None, 64, C.byref(n))
        assert rc == -1
        # before the first step there is no device state
        w3 = api.World(be, scenes.world_def())
        try:
            scenes.add_pyramid(w3, 3)
            rc = w3.fns["world_copy_body_move_events_noalloc"](w3.handle, small.ctypes.data_as(C.POINTER(capi.BodyMoveEvent)), len(small), C.byref(n))
            assert rc == -1 and "before the first step" in be.last_error(w3.handle)
        finally:
            w3.Destroy()
    finally:
        w1.Destroy()
        w2.Destroy()


def test_noalloc_rejects_more_awake_bodies_than_the_fixed_maximum():
    """One body more than B2C_NOALLOC_MAX_MOVE_EVENTS: B2C_E_CAPACITY with a message that names the limit; the existing entry point still serves it."""
    w = api.World(api.CudaBackend(), scenes.world_def())
    try:
        bd = api.default_body_def()
        bd.type = api.DYNAMIC_BODY
        sd = scenes._shape_def(events=False)
        from box2d_b200 import geometry as geo
        circle = geo.Circle((0.0, 0.0), 0.2)
        total = capi.NOALLOC_MAX_MOVE_EVENTS + 1
        for k in range(total):
            bd.position = api.Vec2(float(k % 512), float(k // 512))
            b = w.CreateBody(bd)
            if k < 64:
                w.CreateCircleShape(b, sd, circle)
        w.Step(DT, SUB)
        out = np.zeros(total, dtype=EV)
        n = C.c_int32(5)
        rc = w.fns["world_copy_body_move_events_noalloc"](w.handle, out.ctypes.data_as(C.POINTER(capi.BodyMoveEvent)), len(out), C.byref(n))
        assert rc == -5 and n.value == 0
        msg = w.backend.last_error(w.handle)
        assert str(capi.NOALLOC_MAX_MOVE_EVENTS) in msg and "exceed" in msg
        assert len(w.GetBodyMoveEvents()) == total
    finally:
        w.Destroy()
