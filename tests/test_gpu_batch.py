"""-m gpu: b2cuda_batch_* (batches of independent worlds, api/World.cs:9-10; BASELINE.json configs[4]) through the C-ABI.

World k of a batch must be bit-identical to the oracle stepping the same world ALONE (own id pools, trees, contact ids, colour
graph, coordinates at the origin): bodies, contacts, shapes, joints, events and counters are compared after every step."""
import ctypes as C
import numpy as np
import pytest
import parity
from box2d_b200 import api, scenes, geometry
from oracle_backend import OracleBackend

pytestmark = pytest.mark.gpu


def _builders():
    """A mix of small worlds: pyramids of several heights, mixed shapes, joints, restitution + hit events."""
    out = []
    for rows in (1, 2, 3, 5, 8, 10, 10, 12):
        out.append(lambda w, r=rows: scenes.add_pyramid(w, r))
    out.append(lambda w: scenes.mixed_drop(w, n=40, seed=3))
    out.append(lambda w: scenes.mixed_drop(w, n=90, seed=5))
    out.append(lambda w: scenes.joint_zoo(w, copies=1, angles=False))   # Atan2 joints are tolerance-only (test_all_joint_types_within_tolerance)
    out.append(lambda w: scenes.wheel_rig(w, copies=1))
    out.append(lambda w: scenes.bounce_mix(w, n=24, seed=2))
    out.append(lambda w: scenes.joint_grid(w, n=6))
    out.append(lambda w: scenes.fast_rain(w, n=30))             # continuous collision + sensors (a polygon strip) at CTA scope
    out.append(lambda w: scenes.bullet_range(w, n=15))          # the bullet stage at CTA scope
    out.append(lambda w: None)                                  # an empty world inside a batch
    return out


def _compare_world(k, step, cw, ow):
    a, b = parity.snapshot(cw), parity.snapshot(ow)
    rep = []
    for key in ("begin", "end", "hit", "sensor_begin", "sensor_end"):
        assert parity.compare_struct(f"w{k}.step{step}.{key}", a[key], b[key], report=rep) == 0.0, rep[:4]
    assert a["joint_events"].tolist() == b["joint_events"].tolist()
    ca, cb = a["counters"], b["counters"]
    for f in ("bodyCount", "awakeBodyCount", "shapeCount", "jointCount", "contactCount", "touchingContactCount", "contactIdCapacity", "activeColorCount", "moveCount", "newPairCount"):
        assert getattr(ca, f) == getattr(cb, f), f"world {k} step {step} counters.{f}: {getattr(ca, f)} vs {getattr(cb, f)}"
    assert list(ca.colorCounts) == list(cb.colorCounts)
    e = 0.0
    for key, skip in (("contacts", ()), ("bodies", ()), ("shapes", ("enlarged",)), ("joints", ())):
        e = max(e, parity.compare_struct(f"w{k}.step{step}.{key}", a[key], b[key], skip=skip, report=rep))
    assert e == 0.0, f"world {k} step {step}: " + "; ".join(rep[:6])


def test_batch_worlds_match_oracle_worlds_stepped_alone():
    builders = _builders()
    be = api.CudaBackend()
    batch = api.Batch(be, len(builders), scenes.world_def())
    oracles = [api.World(OracleBackend(), scenes.world_def()) for _ in builders]
    try:
        for k, b in enumerate(builders):
            b(batch.worlds[k])
            b(oracles[k])
        for step in range(25):
            batch.Step(1.0 / 60.0, 4)
            for o in oracles:
                o.Step(1.0 / 60.0, 4)
            if step < 10 or step % 5 == 4:
                for k in range(len(builders)):
                    _compare_world(k, step, batch.worlds[k], oracles[k])
    finally:
        batch.Destroy()
        for o in oracles:
            o.Destroy()


def test_batch_reset_world_and_mutation_between_steps():
    """reset_world puts ONE world back to its state at the first step (ids, trees, contacts, impulses); the others run on.
    Bodies created in one world between steps stay in that world."""
    n = 6
    be = api.CudaBackend()
    wd = scenes.world_def()
    wd.bodyCapacity, wd.shapeCapacity = 128, 128
    batch = api.Batch(be, n, wd)
    oracles = [api.World(OracleBackend(), scenes.world_def()) for _ in range(n)]
    try:
        for k in range(n):
            scenes.add_pyramid(batch.worlds[k], 4 + k)
            scenes.add_pyramid(oracles[k], 4 + k)
        for step in range(8):
            batch.Step(1.0 / 60.0, 4)
            for o in oracles:
                o.Step(1.0 / 60.0, 4)
        # world 2 starts over: a fresh oracle world is its reference from here on
        batch.ResetWorld(2)
        oracles[2].Destroy()
        oracles[2] = api.World(OracleBackend(), scenes.world_def())
        scenes.add_pyramid(oracles[2], 6)
        # world 4 gains a body between steps
        for w in (batch.worlds[4], oracles[4]):
            b = scenes._dynamic_body(w, 0.3, 14.0)
            w.CreateCircleShape(b, scenes._shape_def(), geometry.Circle((0.0, 0.0), 0.4))
        for step in range(8, 20):
            batch.Step(1.0 / 60.0, 4)
            for o in oracles:
                o.Step(1.0 / 60.0, 4)
            for k in range(n):
                _compare_world(k, step, batch.worlds[k], oracles[k])
        # a second reset, after the world was reset once already
        batch.ResetWorld(2)
        oracles[2].Destroy()
        oracles[2] = api.World(OracleBackend(), scenes.world_def())
        scenes.add_pyramid(oracles[2], 6)
        for step in range(20, 26):
            batch.Step(1.0 / 60.0, 4)
            for o in oracles:
                o.Step(1.0 / 60.0, 4)
        for k in range(n):
            _compare_world(k, 25, batch.worlds[k], oracles[k])
    finally:
        batch.Destroy()
        for o in oracles:
            o.Destroy()


def test_batch_stats_reduced_on_device_match_read_back():
    import torch
    n = 50
    be = api.CudaBackend()
    batch = api.Batch(be, n, scenes.world_def())
    try:
        for k in range(n):
            scenes.add_pyramid(batch.worlds[k], 1 + k % 10)
        for _ in range(12):
            batch.Step(1.0 / 60.0, 4)
        dev = torch.zeros(8, dtype=torch.float64, device="cuda")
        s = batch.Stats(dev.data_ptr())
        torch.cuda.synchronize()
        bodies = [w.ReadBodies() for w in batch.worlds]
        dyn = [b[b["flags"] != 0] for b in bodies]
        n_awake = sum(w.GetCounters().awakeBodyCount for w in batch.worlds)
        touching = sum(w.GetCounters().touchingContactCount for w in batch.worlds)
        begins = sum(len(w.GetContactEvents()[0]) for w in batch.worlds)
        assert s.awakeBodies == n_awake and s.touchingContacts == touching and s.beginEvents == begins
        v2 = sum(float(np.sum(d["linearVelocity"]["x"].astype(np.float64) ** 2 + d["linearVelocity"]["y"].astype(np.float64) ** 2)) for d in bodies)
        assert abs(s.sumV2 - v2) <= 1e-9 * max(v2, 1.0)
        assert s.errorFlags == 0.0
        assert dev.cpu().tolist() == [s.awakeBodies, s.sumV2, s.touchingContacts, s.beginEvents, s.endEvents, s.maxSpeed, s.minY, s.errorFlags]
    finally:
        batch.Destroy()


def test_batch_refuses_what_it_cannot_do():
    be = api.CudaBackend()
    batch = api.Batch(be, 2, scenes.world_def())
    try:
        scenes.add_pyramid(batch.worlds[0], 3)
        with pytest.raises(api.B2Error):
            batch.worlds[0].Step(1.0 / 60.0, 4)                 # a member is stepped by its batch
        assert be.fns["world_destroy"](batch.worlds[0].handle) != 0
        h = C.c_void_p()
        assert be.fns["batch_world"](batch.handle, 2, C.byref(h)) != 0
        batch.Step(1.0 / 60.0, 4)
        batch.Step(0.0, 4)                                      # dt == 0: event buffers swap only
        assert len(batch.worlds[0].GetContactEvents()[0]) == 0
    finally:
        batch.Destroy()
    with pytest.raises(api.B2Error):
        api.Batch(be, 2, scenes.world_def(enable_sleep=True))


def test_batch_full_size_8192_pyramids_sampled_worlds_bit_exact():
    """configs[4] at the size BASELINE.json names: 8192 independent 10-row pyramids, 12 steps; every world must equal the ONE
    oracle pyramid (they are identical scenes), checked on the body state of all worlds and in full on sampled worlds."""
    n = 8192
    be = api.CudaBackend()
    batch = api.Batch(be, n, scenes.world_def())
    orc = api.World(OracleBackend(), scenes.world_def())
    try:
        for k in range(n):
            scenes.add_pyramid(batch.worlds[k], 10)
        scenes.add_pyramid(orc, 10)
        for step in range(12):
            batch.Step(1.0 / 60.0, 4)
            orc.Step(1.0 / 60.0, 4)
        ref = orc.ReadBodies().tobytes()
        for k in range(0, n, 97):
            assert batch.worlds[k].ReadBodies().tobytes() == ref, f"world {k} differs"
        for k in (0, 1, 4095, 8191):
            _compare_world(k, 11, batch.worlds[k], orc)
        s = batch.Stats()
        assert s.awakeBodies == n * 55 and s.errorFlags == 0.0
    finally:
        batch.Destroy()
        orc.Destroy()


def test_snapshot_upload_equals_call_by_call_build():
    """b2cuda_world_upload_snapshot replays a recorded creation sequence in one call: ids, lists, trees and every later step are those of
    the call-by-call build (compared with the oracle built call by call); one recording fills several worlds of a batch."""
    def scene(w):
        scenes.add_pyramid(w, 7)
        scenes.joint_grid(w, n=5)
        scenes.mixed_drop(w, n=30, seed=9)

    rec = api.Snapshot()
    scene(rec)
    cuda = api.World(api.CudaBackend(), scenes.world_def())
    orc = api.World(OracleBackend(), scenes.world_def())
    try:
        b, s, j = rec.upload(cuda)
        scene(orc)
        assert len(b) == rec.n_bodies and len(s) == rec.n_shapes and len(j) == rec.n_joints and sorted(b) == list(range(len(b)))
        for step in range(15):
            cuda.Step(1.0 / 60.0, 4)
            orc.Step(1.0 / 60.0, 4)
            _compare_world(0, step, cuda, orc)
        # truncated and corrupted blobs are refused, not half-read
        data = rec.blob()
        for bad in (data[:40], b"\0" * 64, data[:len(data) - 7]):
            rc = cuda.fns["world_upload_snapshot"](cuda.handle, bad, len(bad), None, None, None)
            assert rc != 0
    finally:
        cuda.Destroy()
        orc.Destroy()
    pyr = api.Snapshot()
    scenes.add_pyramid(pyr, 10)
    batch = api.Batch(api.CudaBackend(), 64, scenes.world_def())
    orc = api.World(OracleBackend(), scenes.world_def())
    try:
        for w in batch.worlds:
            pyr.upload(w)
        scenes.add_pyramid(orc, 10)
        for step in range(8):
            batch.Step(1.0 / 60.0, 4)
            orc.Step(1.0 / 60.0, 4)
        for k in (0, 31, 63):
            _compare_world(k, 7, batch.worlds[k], orc)
    finally:
        batch.Destroy()
        orc.Destroy()
