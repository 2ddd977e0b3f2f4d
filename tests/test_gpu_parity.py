"""-m gpu: the CUDA step vs the oracle through the C-ABI, bit-exact unless a tolerance is stated."""
import numpy as np
import pytest
import parity
from box2d_b200 import api, scenes, geometry as geo

pytestmark = pytest.mark.gpu


def test_small_pyramid_bit_exact():
    res = parity.run_pair(lambda w: scenes.add_pyramid(w, 10), steps=20, verbose=True)
    assert res["max_rel_err"] == 0.0


def test_mixed_drop_bit_exact():
    res = parity.run_pair(lambda w: scenes.mixed_drop(w), steps=60, verbose=True)
    assert res["max_rel_err"] == 0.0
    assert res["toi"] == res["toi_oracle"]      # fast bodies take the no-hit continuous path on both sides


def test_large_pyramid_10_steps():
    res = parity.run_pair(lambda w: scenes.large_pyramid(w, 40), steps=10, verbose=True)
    assert res["max_rel_err"] == 0.0


def test_joint_grid_small():
    res = parity.run_pair(lambda w: scenes.joint_grid(w, 12), steps=30, verbose=True)
    assert res["max_rel_err"] == 0.0


def test_tumbler_small():
    res = parity.run_pair(lambda w: scenes.tumbler(w, n=600, half=6.0), steps=40, verbose=True)
    assert res["max_rel_err"] == 0.0


@pytest.mark.parametrize("name", ["pyramid10", "mixed_drop", "joint_grid8", "tumbler300", "joint_zoo_exact", "wheel_rig", "bounce_mix"])
def test_cuda_matches_committed_golden(name):
    """The CUDA path against tests/golden/*.npz (bit-exact pair sets, ids, colors, feature ids, events, body states)."""
    from box2d_b200 import api
    from test_oracle_golden import check_against_golden
    w = api.World(api.CudaBackend(), scenes.world_def())
    try:
        check_against_golden(w, name)
    finally:
        w.Destroy()


@pytest.mark.parametrize("scene,steps", [("tumbler", 60), ("mixed", 40), ("pyramid", 12)])
def test_wide_query_sweep_bit_exact(scene, steps, monkeypatch):
    """Every tree query that visits more than 4 nodes is answered by the grid-wide leaf sweep (b2c_bigquery.cuh):
    pair sets, contact ids and creation order must still equal the reference's DFS order."""
    monkeypatch.setenv("B2CUDA_BIG_QUERY_VISITS", "4")
    build = {"tumbler": lambda w: scenes.tumbler(w, n=600, half=6.0),
             "mixed": lambda w: scenes.mixed_drop(w),
             "pyramid": lambda w: scenes.add_pyramid(w, 12)}[scene]
    res = parity.run_pair(build, steps=steps, verbose=True)
    assert res["max_rel_err"] == 0.0


def test_tumbler_large_walls():
    """Wide kinematic walls over a few thousand bodies: exercises the default visit budget and the parallel rebuild."""
    res = parity.run_pair(lambda w: scenes.tumbler(w, n=4000, half=10.5), steps=30, verbose=True)
    assert res["max_rel_err"] == 0.0


def test_island_partition_valid_and_run_to_run_deterministic(monkeypatch):
    """Device-side self checks (B2CUDA_DEBUG_CHECKSUMS): no colour holds two touching constraints of one dynamic body,
    the island partition is flat and consistent with the constraint arrangement, and two runs agree phase by phase."""
    monkeypatch.setenv("B2CUDA_DEBUG_CHECKSUMS", "1")
    from box2d_b200 import api

    def run():
        w = api.World(api.CudaBackend(), scenes.world_def())
        out = []
        try:
            scenes.mixed_drop(w)
            for _ in range(70):
                w.Step(1.0 / 60.0, 4)
                out.append(w.DebugChecksums())
        finally:
            w.Destroy()
        return out

    a, b = run(), run()
    assert any(any(row) for row in a), "debug checksums not enabled"
    for step, (ra, rb) in enumerate(zip(a, b)):
        assert ra[10] == 0 and ra[11] == 0, f"step {step}: colour conflicts {ra[10]} partition errors {ra[11]}"
        assert ra[:12] == rb[:12], f"step {step}: phase checksums differ between two runs: slots {[k for k in range(12) if ra[k] != rb[k]]}"   # 12..15 hold timings


def test_overflow_colour_bit_exact():
    """A body with more than 20 dynamic contacts: some land in the overflow colour and take the scalar, serial solver
    (quirks Q2-Q4 included), then leave it again (swap-remove order of the overflow array)."""
    res = parity.run_pair(lambda w: scenes.overflow_plank(w), steps=45, verbose=True)
    assert res["max_rel_err"] == 0.0
    assert res["max_overflow"] > 0, "scene did not reach the overflow colour"


def test_tumbler_dense_pile_long():
    """Long enough for the pile to compact (contacts churn, colours fill up)."""
    res = parity.run_pair(lambda w: scenes.tumbler(w, n=1500, half=5.0, pitch=0.22), steps=160, verbose=False)
    assert res["max_rel_err"] == 0.0


@pytest.mark.parametrize("scene,steps", [("pyramids", 25), ("mixed", 60), ("tumbler", 40)])
def test_warp_per_island_variant_bit_exact(scene, steps, monkeypatch):
    """Force the throughput variant of the island solver (one warp per island, lanes in shared memory)."""
    monkeypatch.setenv("B2CUDA_ISLAND_MODE", "warp")
    build = {"pyramids": lambda w: scenes.batched_pyramids(w, n_worlds=64),
             "mixed": lambda w: scenes.mixed_drop(w),
             "tumbler": lambda w: scenes.tumbler(w, n=600, half=6.0)}[scene]
    res = parity.run_pair(build, steps=steps, verbose=True)
    assert res["max_rel_err"] == 0.0


def test_body_move_events_match_reference_layout():
    """WorldAPI.GetBodyEvents: one event per awake body in awake-set order, equal to the oracle's, streamed after the first call."""
    from box2d_b200 import api
    from oracle_backend import OracleBackend
    cuda = api.World(api.CudaBackend(), scenes.world_def())
    orc = api.World(OracleBackend(), scenes.world_def())
    try:
        scenes.mixed_drop(cuda)
        scenes.mixed_drop(orc)
        for step in range(12):
            cuda.Step(1.0 / 60.0, 4)
            orc.Step(1.0 / 60.0, 4)
            a, b = cuda.GetBodyMoveEvents(), orc.GetBodyMoveEvents()
            assert len(a) == len(b) == cuda.GetCounters().awakeBodyCount
            assert a.tobytes() == b.tobytes(), f"step {step}"
            bodies = cuda.ReadBodies()
            assert a["transform"].tobytes() == bodies["transform"][a["bodyId"]].tobytes()
        cuda.Step(0.0, 4)
        orc.Step(0.0, 4)
        assert len(cuda.GetBodyMoveEvents()) == 0 and len(orc.GetBodyMoveEvents()) == 0
    finally:
        cuda.Destroy()
        orc.Destroy()


def _poke(world, step):
    """Every body mutator of the boundary, interleaved with steps (api/Body.cs:293-653, api/World.cs:666-794)."""
    if step == 0:                                   # before the first step: the host mirror takes the patch
        world.SetGravityScale(3, 0.5)
        world.SetLinearDamping(4, 0.8)
        world.SetAngularDamping(4, 1.5)
        world.ApplyLinearImpulse(5, (0.7, 1.9), (0.2, 3.1))
        world.ApplyForce(6, (30.0, 12.0), (1.0, 2.0))
        world.ApplyAngularImpulse(7, 0.35)
    if step == 2:
        world.ApplyForce(9, (-40.0, 25.0), (0.3, 4.4))
        world.ApplyForceToCenter(9, (3.0, 1.0))
        world.ApplyTorque(10, 7.5)
        world.ApplyForce(11, (50.0, 50.0), (0.0, 0.0))
        world.ClearForces(11)
    if step == 3:
        world.ApplyLinearImpulse(12, (2.5, 0.25), (-1.2, 2.6))
        world.ApplyLinearImpulseToCenter(13, (-1.0, 3.0))
        world.ApplyAngularImpulse(14, -0.6)
        world.SetGravityScale(15, -0.25)
        world.SetLinearDamping(16, 2.0)
        world.SetAngularDamping(16, 0.4)
        world.SetLinearVelocity(17, (1.5, 6.0))
        world.SetAngularVelocity(18, -3.0)
    if step == 5:
        world.SetRestitutionThreshold(0.25)
        world.SetHitEventThreshold(0.1)
        world.SetMaximumLinearSpeed(5.0)
        world.ApplyLinearImpulse(20, (40.0, 90.0), (0.0, 5.0))     # clipped by the new speed limit
    if step == 7:
        world.SetTransform(21, (6.0, 9.0), (0.8, 0.6))
        world.SetMaximumLinearSpeed(4.0 * 100.0)


def test_body_and_world_mutators_between_steps_bit_exact():
    res = parity.run_pair(lambda w: scenes.add_pyramid(w, 8), steps=24, mutate=_poke)
    assert res["max_rel_err"] == 0.0


def test_single_body_getters_match_bulk_read():
    cuda = api.World(api.CudaBackend(), scenes.world_def())
    try:
        scenes.add_pyramid(cuda, 5)
        for phase in range(2):                      # host mirror, then device rows
            bulk = cuda.ReadBodies()
            for i in (0, 1, 7, 15):
                (p, q), v, w, c = cuda.GetTransform(i), cuda.GetLinearVelocity(i), cuda.GetAngularVelocity(i), cuda.GetWorldCenterOfMass(i)
                r = bulk[i]
                assert p == (r["transform"]["p"]["x"], r["transform"]["p"]["y"]) and q == (r["transform"]["q"]["c"], r["transform"]["q"]["s"])
                assert v == (r["linearVelocity"]["x"], r["linearVelocity"]["y"]) and w == r["angularVelocity"]
                assert c == (r["center"]["x"], r["center"]["y"])
            for _ in range(5):
                cuda.Step(1.0 / 60.0, 4)
        with pytest.raises(Exception):
            cuda.GetTransform(10_000)
    finally:
        cuda.Destroy()


def _teleport(world, step):
    """BodyAPI.SetTransform after the upload (api/Body.cs:293-325): device-side MoveProxy + BufferMove."""
    if step == 3:
        world.SetTransform(55, (7.5, 0.6), (1.0, 0.0))            # the top box, dropped beside the pile
    if step == 5:
        world.SetTransform(0, (0.25, -0.05), (1.0, 0.0))          # the static ground: static tree move, pairs re-found
    if step == 8:
        world.SetTransform(54, (0.1, 3.2), (0.6, 0.8))            # into the middle of the pile
        world.SetTransform(54, (0.2, 3.3), (0.6, 0.8))            # twice: buffered once
        world.SetTransform(12, (-30.0, 40.0), (0.0, 1.0))         # far away: every contact of the box ends
    if step == 9:
        world.SetTransform(12, (-30.0, 40.01), (0.0, 1.0))        # inside the fat box: no proxy move
    if step == 12:
        for b in range(20, 30):
            world.SetTransform(b, (float(b) - 25.0, 14.0 + 0.01 * b), (1.0, 0.0))


def test_set_transform_after_upload_bit_exact():
    res = parity.run_pair(lambda w: scenes.add_pyramid(w, 10), steps=30, mutate=_teleport)
    assert res["max_rel_err"] == 0.0


def test_set_transform_kinematic_and_jointed_bodies():
    def poke(world, step):
        if step in (4, 9):
            world.SetTransform(0, (0.0, 0.5 if step == 4 else -0.5), (0.98006657, 0.19866933))   # the kinematic tumbler frame
            world.SetTransform(40, (0.0, 10.0), (1.0, 0.0))
    res = parity.run_pair(lambda w: scenes.tumbler(w, 400), steps=16, mutate=poke)
    assert res["max_rel_err"] == 0.0
    res = parity.run_pair(lambda w: scenes.joint_grid(w, 8), steps=12,
                          mutate=lambda w, s: w.SetTransform(10, (3.0, -2.0), (0.8, -0.6)) if s == 5 else None)
    assert res["max_rel_err"] == 0.0


def test_all_joint_types_atan2_free_subset_bit_exact():
    """Distance (rigid, spring + limit + motor), motor (velocity and linear spring drives), filter and revolute joints over a ground
    with contacts: none of them evaluates Atan2, so impulses, joint events and body states must be bit-identical."""
    res = parity.run_pair(lambda w: scenes.joint_zoo(w, copies=3, angles=False), steps=120, verbose=True)
    assert res["max_rel_err"] == 0.0


def test_wheel_joints_bit_exact():
    res = parity.run_pair(lambda w: scenes.wheel_rig(w, copies=3), steps=60, verbose=True)
    assert res["max_rel_err"] == 0.0


def test_all_joint_types_within_tolerance():
    """Weld, prismatic and the motor joint's angular spring call Atan2 every solve (MathF.Atan2 is CRT-dependent in .NET too): the
    north-star tolerance for body state (1e-5 relative after a step) is applied; events, ids and colours stay exact."""
    res = parity.run_pair(lambda w: scenes.joint_zoo(w, copies=3), steps=1, tol=1e-5, verbose=True)
    assert res["max_rel_err"] <= 1e-5
    # longer run: ids, colours, contact and joint events stay exact (checked inside run_pair); the float measure is relative per
    # component, so velocity components passing through zero dominate it - reported, not bounded; the 600-step test below is the 1 % check
    res = parity.run_pair(lambda w: scenes.joint_zoo(w, copies=3), steps=60, tol=float("inf"))
    print("joint_zoo 60 steps: max per-component relative difference", res["max_rel_err"])


def test_all_joint_types_600_steps_configuration_within_1_percent():
    """North-star long-run criterion (resting configuration / kinetic energy within 1 % over 600 steps) on the scene with every joint type."""
    import numpy as np
    from oracle_backend import OracleBackend
    out = []
    for be in (api.CudaBackend(), OracleBackend()):
        w = api.World(be, scenes.world_def())
        scenes.joint_zoo(w, copies=2)
        for _ in range(600):
            w.Step(1.0 / 60.0, 4)
        b = w.ReadBodies()
        out.append((np.stack([b["transform"]["p"]["x"], b["transform"]["p"]["y"]], 1).astype(np.float64),
                    np.stack([b["linearVelocity"]["x"], b["linearVelocity"]["y"]], 1).astype(np.float64)))
        w.Destroy()
    (pa, va), (pb, vb) = out
    assert np.abs(pa - pb).max() <= 0.01 * max(1.0, np.abs(pb).max())
    ka, kb = (va ** 2).sum(), (vb ** 2).sum()
    assert abs(ka - kb) <= 0.01 * max(kb, 1e-3)


def test_joint_warm_starting_off_and_substep_variants():
    def mutate(w, step):
        if step == 0:
            w.EnableWarmStarting(False)
        if step == 10:
            w.EnableWarmStarting(True)
    res = parity.run_pair(lambda w: scenes.joint_zoo(w, copies=1, angles=False), steps=30, verbose=True, mutate=mutate)
    assert res["max_rel_err"] == 0.0
    res = parity.run_pair(lambda w: scenes.wheel_rig(w), steps=20, substeps=8, dt=1.0 / 120.0, verbose=True)
    assert res["max_rel_err"] == 0.0


def test_joint_setters_between_steps_bit_exact():
    """api/Joint.cs setters applied between steps on both sides (device: one-thread patches on the world's stream): parameters,
    spring/limit/motor switches with their impulse resets, limits, thresholds, constraint tuning."""
    from box2d_b200 import capi as J
    holder = {}

    def build(w):
        holder["zoo"] = scenes.joint_zoo(w, copies=2, angles=False)
        holder["rig"] = scenes.wheel_rig(w, copies=1, pitch=-30.0)

    def mutate(w, step):
        zoo, rig = holder["zoo"], holder["rig"][0]
        if step == 3:
            for ids in zoo:
                w.JointSetParam(ids["distance_spring"], J.JP_MOTOR_SPEED, -0.75)
                w.JointSetParam(ids["distance_spring"], J.JP_SPRING_HERTZ, 3.5)
                w.JointSetParam(ids["distance_rigid"], J.JP_LENGTH, 2.5)
                w.MotorJoint_SetLinearVelocity(ids["motor_velocity"], (-0.25, 0.5))
                w.JointSetParam(ids["motor_velocity"], J.JP_ANGULAR_VELOCITY, -2.0)
                w.JointSetParam(ids["motor_spring"], J.JP_MAX_SPRING_FORCE, -1.0)      # Max(0, x)
                w.JointSetParam(ids["revolute"], J.JP_FORCE_THRESHOLD, 1.0e9)
            w.JointSetParam(rig[0], J.JP_MOTOR_SPEED, -4.0)
            w.JointSetLimits(rig[0], 0.3, -0.1)                                        # swapped on purpose: min/max
        if step == 6:
            for ids in zoo:
                w.JointEnable(ids["distance_spring"], J.JOINT_MOTOR_SWITCH, False)
                w.JointEnable(ids["distance_spring"], J.JOINT_LIMIT, False)
                w.JointSetLimits(ids["distance_spring"], 0.5, 3.0)
                w.JointEnable(ids["revolute"], J.JOINT_MOTOR_SWITCH, True)
                w.JointSetParam(ids["revolute"], J.JP_MAX_MOTOR, 30.0)
                w.JointSetParam(ids["revolute"], J.JP_MOTOR_SPEED, 1.5)
                w.SetConstraintTuning(ids["revolute"], 30.0, 1.0)
            w.JointEnable(rig[1], J.JOINT_LIMIT, True)      # was off -> the port stores false and clears the impulses
            w.JointEnable(rig[2], J.JOINT_SPRING, False)
        if step == 12:
            for ids in zoo:
                w.JointEnable(ids["distance_spring"], J.JOINT_MOTOR_SWITCH, True)
                w.JointSetLimits(ids["revolute"], -0.2, 0.3)          # limit stays off: the revolute limit evaluates Atan2 (tolerance-only)
            w.JointEnable(rig[0], J.JOINT_MOTOR_SWITCH, False)

    res = parity.run_pair(build, steps=40, verbose=True, mutate=mutate)
    assert res["max_rel_err"] == 0.0


def test_joint_setter_rejects_fields_of_other_types():
    from box2d_b200 import capi as J
    w = api.World(api.CudaBackend(), scenes.world_def())
    try:
        ids = scenes.joint_zoo(w, copies=1, angles=False)[0]
        for jid, param in ((ids["filter"], J.JP_MOTOR_SPEED), (ids["revolute"], J.JP_LENGTH), (ids["distance_rigid"], J.JP_TARGET),
                           (ids["motor_velocity"], J.JP_MAX_MOTOR)):
            with pytest.raises(api.B2Error):
                w.JointSetParam(jid, param, 1.0)
        with pytest.raises(api.B2Error):
            w.JointEnable(ids["filter"], J.JOINT_SPRING, True)
        with pytest.raises(api.B2Error):
            w.JointSetLimits(ids["motor_velocity"], 0.0, 1.0)
        with pytest.raises(api.B2Error):
            w.JointSetParam(999, J.JP_MOTOR_SPEED, 1.0)
    finally:
        w.Destroy()


def test_mass_locks_material_mutators_between_steps_bit_exact():
    """BodyAPI.SetMassData / SetMotionLocks and ShapeAPI.SetSurfaceMaterial before the first step and between steps."""
    def build(w):
        scenes.add_pyramid(w, 8)
        w.SetMotionLocks(5, False, False, True)           # before the upload: host mirror
        w.SetMassData(6, 3.0, 0.75, (0.1, -0.05))
        w.SetSurfaceMaterial(7, 0.1, 0.0, 0.0, 0.5)

    def mutate(w, step):
        if step == 2:
            w.SetMotionLocks(9, True, False, False)
            w.SetMotionLocks(10, False, True, True)
            w.SetMassData(11, 0.5, 0.02, (0.0, 0.2))
            w.SetSurfaceMaterial(0, 0.05, 0.0, 0.0, 0.0)   # the ground becomes slippery
            w.SetSurfaceMaterial(12, 0.9, 0.0, 0.1, -0.3)  # rolling resistance + tangent speed
        if step == 6:
            w.SetMotionLocks(9, False, False, False)
            w.SetMotionLocks(5, False, False, True)        # same locks again: no change
            w.SetMassData(13, 0.0, 0.0, (0.0, 0.0))        # massless: infinite-mass body in the solver
            w.ApplyLinearImpulseToCenter(14, (3.0, 1.0))
    res = parity.run_pair(build, steps=30, verbose=True, mutate=mutate)
    assert res["max_rel_err"] == 0.0


def test_set_target_transform_within_tolerance():
    """BodyAPI.SetTargetTransform: the angular part goes through Atan2 (CRT-dependent), the linear part is exact."""
    def build(w):
        scenes.add_pyramid(w, 4)
        kd = api.default_body_def()
        kd.type = api.KINEMATIC_BODY
        kd.position = api.Vec2(-6.0, 2.0)
        k = w.CreateBody(kd)
        w.CreatePolygonShape(k, scenes._shape_def(), __import__("box2d_b200").geometry.MakeBox(0.5, 0.25))

    def mutate(w, step):
        k = w.GetCounters().bodyCount - 1
        if step % 3 == 0:
            w.SetTargetTransform(k, (-6.0 + 0.25 * step, 2.0), (0.9800666, 0.19866933), 1.0 / 60.0)
    res = parity.run_pair(build, steps=20, tol=1e-5, verbose=True, mutate=mutate)
    assert res["max_rel_err"] <= 1e-5


@pytest.mark.parametrize("scene,steps", [("pyramid", 20), ("joints", 30), ("tumbler", 30), ("zoo", 40)])
def test_grid_barrier_solver_variant_bit_exact(scene, steps, monkeypatch):
    """Large / jointed islands are solved by one thread-block cluster by default (k_solver_cluster); B2CUDA_CLUSTER=0 keeps them in
    the cooperative front kernel with grid-wide barriers. Both must give the reference's floats."""
    monkeypatch.setenv("B2CUDA_CLUSTER", "0")
    build = {"pyramid": lambda w: scenes.large_pyramid(w, 40), "joints": lambda w: scenes.joint_grid(w, 12),
             "tumbler": lambda w: scenes.tumbler(w, n=600, half=6.0), "zoo": lambda w: scenes.joint_zoo(w, copies=2, angles=False)}[scene]
    res = parity.run_pair(build, steps=steps, verbose=True)
    assert res["max_rel_err"] == 0.0


@pytest.mark.parametrize("mode", ["0", "fallback"])
def test_cluster_solver_without_its_shared_memory_variant_bit_exact(mode, monkeypatch):
    """One large island of contacts is solved in the cluster's distributed shared memory by default (b2c_cluster_dsm.cuh; every other large-island
    test runs that way). B2CUDA_CLUSTER_DSM=0 keeps the plain cluster path (lanes in global planes); =fallback makes every eligible step take the
    shared-memory solver's way out (a CTA that would own more constraints than it has lanes: PrepareContacts inside the cluster kernel, then the
    plain path). Both must give the reference's floats."""
    monkeypatch.setenv("B2CUDA_CLUSTER_DSM", mode)
    res = parity.run_pair(lambda w: scenes.large_pyramid(w, 40), steps=20, verbose=True)
    assert res["max_rel_err"] == 0.0
    res = parity.run_pair(lambda w: scenes.tumbler(w, n=600, half=6.0), steps=30)
    assert res["max_rel_err"] == 0.0


@pytest.mark.parametrize("scene", ["mixed", "zoo"])
def test_graph_launch_path_bit_exact(scene, monkeypatch):
    """B2CUDA_GRAPH=1 / b2cuda_world_enable_graph_launch: a step is ONE CUDA-graph launch instead of one launch per kernel. Same results."""
    monkeypatch.setenv("B2CUDA_GRAPH", "1")
    build = {"mixed": lambda w: scenes.mixed_drop(w), "zoo": lambda w: scenes.joint_zoo(w, copies=1, angles=False)}[scene]
    res = parity.run_pair(build, steps=25, verbose=True)
    assert res["max_rel_err"] == 0.0


def test_graph_rekeys_when_step_parameters_change(monkeypatch):
    """Setters that change what is baked into the graph (gravity, tuning, warm starting, dt, substeps) re-key it."""
    monkeypatch.setenv("B2CUDA_GRAPH", "1")

    def mutate(w, step):
        if step == 4:
            w.SetGravity((1.0, -9.0))
        if step == 8:
            w.SetContactTuning(20.0, 5.0, 2.0)
        if step == 12:
            w.EnableWarmStarting(False)
        if step == 16:
            w.EnableWarmStarting(True)
            w.SetGravity((0.0, -10.0))
    res = parity.run_pair(lambda w: scenes.add_pyramid(w, 8), steps=24, verbose=True, mutate=mutate)
    assert res["max_rel_err"] == 0.0
    # varying dt / substeps every step: more than 8 distinct configurations, the oldest graphs are dropped
    cuda = api.World(api.CudaBackend(), scenes.world_def())
    from oracle_backend import OracleBackend
    orc = api.World(OracleBackend(), scenes.world_def())
    try:
        for w in (cuda, orc):
            scenes.add_pyramid(w, 6)
        for k in range(24):
            dt, sub = 1.0 / (50 + (k % 11)), 2 + (k % 3)
            cuda.Step(dt, sub)
            orc.Step(dt, sub)
        a, b = cuda.ReadBodies(), orc.ReadBodies()
        assert a.tobytes() == b.tobytes()
    finally:
        cuda.Destroy()
        orc.Destroy()


def test_destroy_bodies_between_steps_bit_exact():
    """BodyAPI.DestroyBody (api/Body.cs:106-160) between steps: the body's contacts leave in contact-list order (end events reported by the
    next step, ids back to the LIFO pool, colour bits cleared), its proxies leave the trees, the awake set closes the gap by swap-remove.
    Bodies from the bottom row, the middle and the top of a pyramid, then the static ground itself."""
    def mutate(w, step):
        if step == 6:
            w.DestroyBody(5)
        elif step == 10:
            w.DestroyBody(30)
            w.DestroyBody(31)
        elif step == 15:
            w.DestroyBody(55)
            w.DestroyBody(1)
        elif step == 22:
            w.DestroyBody(0)
    res = parity.run_pair(lambda w: scenes.add_pyramid(w, 10), steps=36, mutate=mutate, verbose=True)
    assert res["max_rel_err"] == 0.0
    assert res["bodies"] == 56 - 6


def test_destroy_bodies_in_overflow_colour_and_pile_bit_exact():
    def mutate(w, step):
        if step == 5:
            w.DestroyBody(3)           # a box under the plank: one of its contacts sits in the overflow colour
        elif step == 9:
            w.DestroyBody(27)          # the plank
    res = parity.run_pair(lambda w: scenes.overflow_plank(w, 26), steps=20, mutate=mutate, verbose=True)
    assert res["max_rel_err"] == 0.0

    def cull(w, step):
        if step in (8, 16, 24):
            for b in range(10 + step, 400, 37):
                w.DestroyBody(b)
    res = parity.run_pair(lambda w: scenes.tumbler(w, 400), steps=32, mutate=cull, verbose=True)
    assert res["max_rel_err"] == 0.0


def test_destroy_joints_and_jointed_bodies_bit_exact():
    """JointAPI.DestroyJoint (DestroyJointInternal src/Joint.cs:297-375) before and after the first step, and DestroyBody on bodies that
    carry joints (api/Body.cs:117-129 walks the joint list first)."""
    holder = {}

    def build(w):
        holder["zoo"] = scenes.joint_zoo(w, copies=3, angles=False)
        w.DestroyJoint(holder["zoo"][2]["distance_rigid"])          # host-side path: the world is still being assembled

    def mutate(w, step):
        zoo = holder["zoo"]
        if step == 4:
            w.DestroyJoint(zoo[0]["filter"])                        # the two overlapping boxes start to collide
            w.DestroyJoint(zoo[1]["revolute"])
        elif step == 9:
            w.DestroyJoint(zoo[0]["distance_spring"])
            w.DestroyJoint(zoo[1]["motor_velocity"])
        elif step == 14:
            j = w.ReadJoints()
            assert zoo[0]["filter"] not in j["jointId"].tolist()
            w.DestroyBody(3)                                        # carries joints in the Atan2-free zoo
            w.DestroyBody(1)
        elif step == 20:
            w.DestroyBody(0)                                        # the first copy's ground: every joint anchored to it goes too
    res = parity.run_pair(build, steps=40, mutate=mutate, verbose=True)
    assert res["max_rel_err"] == 0.0

    res = parity.run_pair(lambda w: scenes.joint_grid(w, 8), steps=24, verbose=True,
                          mutate=lambda w, s: (w.DestroyBody(27), w.DestroyJoint(5), w.DestroyJoint(40)) if s == 6 else None)
    assert res["max_rel_err"] == 0.0


def test_destroy_rejects_stale_ids_and_reports_counts():
    cuda = api.World(api.CudaBackend(), scenes.world_def())
    try:
        scenes.add_pyramid(cuda, 4)
        cuda.Step(1.0 / 60.0, 4)
        cuda.EnableSleeping(False)                                  # accepted: nothing sleeps in this build
        with pytest.raises(Exception, match="sleep"):
            cuda.EnableSleeping(True)                               # refused loudly, sleeping stays off
        cuda.DestroyBody(3)
        with pytest.raises(Exception):
            cuda.DestroyBody(3)
        with pytest.raises(Exception):
            cuda.DestroyJoint(0)
        with pytest.raises(Exception):
            cuda.GetTransform(3)
        for call in (lambda: cuda.SetTransform(3, (0.0, 9.0), (1.0, 0.0)), lambda: cuda.SetLinearVelocity(3, (1.0, 0.0)),
                     lambda: cuda.SetMassData(3, 1.0, 1.0, (0.0, 0.0)), lambda: cuda.ApplyMassFromShapes(3), lambda: cuda.DestroyShape(3, True),
                     lambda: cuda.SetSurfaceMaterial(3, 0.5, 0.0, 0.0, 0.0), lambda: cuda.Shape_EnableHitEvents(3, True)):
            with pytest.raises(Exception):
                call()                                              # body 3 and its shape 3 are gone: every mutator says so
        cuda.Step(1.0 / 60.0, 4)
        c = cuda.GetCounters()
        assert c.bodyCount == 10 and c.awakeBodyCount == 9 and c.shapeCount == 10
        assert len(cuda.GetBodyMoveEvents()) == 9
    finally:
        cuda.Destroy()


def test_destroy_shapes_and_mass_updates_bit_exact():
    """ShapeAPI.DestroyShape (api/Shape.cs:48-56), ShapeAPI.SetDensity (:95-103) and BodyAPI.ApplyMassFromShapes (api/Body.cs:617-621) between
    steps on the mixed scene: second and first shapes of compound bodies (with and without the mass update), a static terrain piece, a
    kinematic paddle's capsule, a body left without any shape."""
    def mutate(w, step):
        if step == 5:
            w.DestroyShape(7, True)            # body 2: the extra circle
            w.DestroyShape(14, True)           # body 9: its main shape, the small circle stays
            w.DestroyShape(22, False)          # body 16: mass data left as it was
        elif step == 9:
            w.DestroyShape(3, True)            # the static rounded box
            w.DestroyShape(5, True)            # the paddle's capsule
        elif step == 12:
            w.SetDensity(8, 3.0, True)
            w.SetDensity(16, 0.25, True)
            w.SetDensity(9, 5.0, False)
        elif step == 14:
            w.ApplyMassFromShapes(4)           # picks up shape 9's new density
            w.ApplyMassFromShapes(0)
            w.ApplyMassFromShapes(1)
        elif step == 16:
            w.DestroyShape(8, True)            # body 3 has no shape left: zero mass, it hangs where it is
        elif step == 20:
            w.DestroyBody(2)
            w.DestroyBody(3)
    res = parity.run_pair(lambda w: scenes.mixed_drop(w, 60), steps=40, mutate=mutate, verbose=True)
    assert res["max_rel_err"] == 0.0


def test_event_enable_flags_between_steps_bit_exact():
    """ShapeAPI/BodyAPI.EnableContactEvents and EnableHitEvents (api/Shape.cs:182-209, api/Body.cs:848-872) toggled between steps: contacts
    created afterwards sample the new flag, contacts that exist keep theirs, so begin/end event lists must still agree exactly."""
    wd = scenes.world_def()
    wd.hitEventThreshold = 0.5

    def mutate(w, step):
        if step == 2:
            for sid in range(6, 30, 2):
                w.Shape_EnableContactEvents(sid, False)
            w.Body_EnableContactEvents(0, False)          # the ground: its four shapes
        elif step == 8:
            for sid in range(6, 30, 4):
                w.Shape_EnableContactEvents(sid, True)
                w.Shape_EnableHitEvents(sid, True)
            w.Body_EnableHitEvents(1, True)
        elif step == 20:
            w.Body_EnableContactEvents(0, True)
    res = parity.run_pair(lambda w: scenes.mixed_drop(w, 60), steps=50, world_def=wd, mutate=mutate, verbose=True)
    assert res["max_rel_err"] == 0.0


def _spawn(w, x, y, kind, v=(0.0, 0.0), body_type=None):
    bd = api.default_body_def()
    bd.type = (api.STATIC_BODY if kind == "static" else api.DYNAMIC_BODY) if body_type is None else body_type
    bd.position = api.Vec2(x, y)
    bd.linearVelocity = api.Vec2(*v)
    bd.angularVelocity = 0.5 if kind != "static" else 0.0
    b = w.CreateBody(bd)
    sd = scenes._shape_def()
    if kind == "box":
        w.CreatePolygonShape(b, sd, geo.MakeBox(0.5, 0.5))
    elif kind == "circle":
        w.CreateCircleShape(b, sd, geo.Circle((0.0, 0.0), 0.4))
    elif kind == "compound":
        w.CreatePolygonShape(b, sd, geo.MakeBox(0.4, 0.2))
        w.CreateCircleShape(b, sd, geo.Circle((0.5, 0.0), 0.25))
        w.CreateCapsuleShape(b, sd, geo.Capsule((-0.6, -0.2), (-0.6, 0.2), 0.1))
    elif kind == "static":
        w.CreateSegmentShape(b, sd, geo.Segment((-3.0, 0.0), (3.0, 0.5)))
    elif kind == "paddle":
        w.CreatePolygonShape(b, sd, geo.MakeBox(2.0, 0.1))
    return b


def test_create_bodies_and_shapes_between_steps_bit_exact():
    """BodyAPI.CreateBody (api/Body.cs:16-105) and ShapeAPI.CreateShape (api/Shape.cs:8-24) after the first step: new rows are appended to
    the awake set, proxies enter the device trees through the reference's CreateProxy / InsertLeaf (free-list order included, so proxy keys
    agree), the move buffer and the mass update run as at creation time. Ids freed by DestroyBody come back LIFO."""
    made = {}

    def mutate(w, step):
        if step == 5:
            made[id(w)] = [_spawn(w, -2.0, 14.0, "box", (0.0, -3.0)), _spawn(w, 1.3, 15.0, "circle"), _spawn(w, 0.2, 17.0, "compound", (0.5, -1.0))]
        elif step == 9:
            _spawn(w, 14.0, 3.0, "static")
            _spawn(w, 13.0, 8.0, "box")
            _spawn(w, -12.0, 6.0, "paddle", body_type=api.KINEMATIC_BODY)
        elif step == 14:
            w.DestroyBody(made[id(w)][0])
            w.DestroyBody(7)
            a = _spawn(w, -1.0, 16.0, "circle", (0.0, -6.0))         # takes body id 7 and its shape id back
            b = _spawn(w, 3.0, 16.0, "compound")
            assert a == 7 and b == made[id(w)][0]
        elif step == 20:
            for k in range(12):
                _spawn(w, -6.0 + k, 12.0 + 0.3 * k, ("box", "circle", "compound")[k % 3])
        elif step == 26:
            sd = scenes._shape_def()
            w.CreateCircleShape(3, sd, geo.Circle((0.0, 0.6), 0.3))    # a second shape on a body of the pyramid, mass updated
            sd.updateBodyMass = 0
            w.CreateCircleShape(4, sd, geo.Circle((0.0, 0.6), 0.3))    # and one that leaves the mass alone
    res = parity.run_pair(lambda w: scenes.add_pyramid(w, 8), steps=60, mutate=mutate, verbose=True)
    assert res["max_rel_err"] == 0.0
    assert res["bodies"] == 37 + 3 + 3 + 12


def test_create_many_bodies_grows_the_device_arrays():
    """Spawning far more than the rows reserved at upload (default: a quarter more, at least 256): body, awake, shape, move-buffer and
    tree arrays are re-allocated at x1.5 between steps, the tree's node pool follows the reference's own x1.5 steps inside them, and the
    step stays bit-exact. Also with rows reserved up front through the world definition, and with move/body events streaming."""
    def mutate(w, step):
        if step in (3, 6, 9, 12):
            for k in range(150):
                _spawn(w, -30.0 + 0.41 * k, 6.0 + (k % 5) * 1.1 + step, ("box", "circle")[k % 2])
        if step == 5:
            w.GetBodyMoveEvents()                        # from here on the step streams the events: the buffers must follow the growth
            w.GetBodyEvents()
    res = parity.run_pair(lambda w: scenes.add_pyramid(w, 5), steps=26, mutate=mutate, verbose=True)
    assert res["max_rel_err"] == 0.0 and res["bodies"] == 16 + 600

    wd = scenes.world_def()
    wd.bodyCapacity, wd.shapeCapacity = 2000, 2000
    res = parity.run_pair(lambda w: scenes.add_pyramid(w, 5), steps=16, world_def=wd, mutate=mutate, verbose=True)
    assert res["max_rel_err"] == 0.0

    cuda = api.World(api.CudaBackend(), scenes.world_def())
    try:
        scenes.add_pyramid(cuda, 3)
        cuda.Step(1.0 / 60.0, 4)
        for k in range(700):
            _spawn(cuda, float(k), 30.0, "circle")
        cuda.Step(1.0 / 60.0, 4)
        n = cuda.GetCounters().bodyCount
        assert n == 7 + 700 and len(cuda.GetBodyMoveEvents()) == n - 1 and len(cuda.GetBodyEvents()) == n
        moves = cuda.GetBodyMoveEvents()
        assert sorted(moves["bodyId"].tolist()) == list(range(1, n))
    finally:
        cuda.Destroy()


def test_create_many_joints_grows_the_joint_rows():
    from box2d_b200 import capi as J

    def mutate(w, step):
        if step == 4:
            for k in range(300):                         # chains the boxes of the pile pairwise, far past the 256 reserved joint rows
                d = _jdef(J.JOINT_DISTANCE, 2 + k, 3 + k, (0.0, 0.0), (0.0, 0.0))
                d.length, d.enableSpring, d.hertz, d.dampingRatio, d.collideConnected = 0.6, 1, 4.0, 0.5, 1
                w.CreateJoint(d)
    res = parity.run_pair(lambda w: scenes.tumbler(w, 400), steps=14, mutate=mutate, verbose=True)
    assert res["max_rel_err"] == 0.0


def _jdef(t, a, b, pa, pb):
    d = api.default_joint_def(t)
    d.bodyIdA, d.bodyIdB = a, b
    d.localFrameA = api.Transform(api.Vec2(*pa), api.Rot(1.0, 0.0))
    d.localFrameB = api.Transform(api.Vec2(*pb), api.Rot(1.0, 0.0))
    return d


def test_create_joints_between_steps_bit_exact():
    """World.CreateJoint (src/Joint.cs:181-296) after the first step: the colour is assigned against the device's colour bitsets (contacts
    have taken bits since the upload), the contacts between the two bodies are destroyed when collideConnected is off (:155-180, end events
    in the next step), joint ids freed by DestroyJoint come back LIFO. Atan2-free joint types, so everything is bit-exact."""
    from box2d_b200 import capi as J
    made = {}

    def mutate(w, step):
        if step == 6:
            # bodies 1 and 2 are neighbours in the bottom row, 1 and 9 touch across rows: their contacts go away with the joints
            d = _jdef(J.JOINT_REVOLUTE, 1, 2, (0.5, 0.0), (-0.5, 0.0))
            d.forceThreshold = 1.0
            j0 = w.CreateJoint(d)
            j1 = w.CreateJoint(_jdef(J.JOINT_FILTER, 1, 9, (0.0, 0.0), (0.0, 0.0)))
            d = _jdef(J.JOINT_DISTANCE, 0, 36, (0.0, 14.0), (0.0, 0.0))
            d.length, d.enableSpring, d.hertz, d.dampingRatio = 6.0, 1, 2.0, 0.3
            d.minLength, d.maxLength, d.enableLimit = 4.0, 8.0, 1
            j2 = w.CreateJoint(d)
            made[id(w)] = [j0, j1, j2]
        elif step == 11:
            d = _jdef(J.JOINT_MOTOR, 0, 30, (0.0, 0.0), (0.0, 0.0))
            d.linearVelocity, d.maxVelocityForce, d.angularVelocity, d.maxVelocityTorque = api.Vec2(0.5, 0.2), 40.0, 1.0, 5.0
            d.collideConnected = 1
            w.CreateJoint(d)
            b = _spawn(w, 6.0, 9.0, "box")
            d = _jdef(J.JOINT_DISTANCE, b, 35, (0.0, 0.0), (0.0, 0.0))
            d.length = 4.0
            made[id(w)].append(w.CreateJoint(d))
        elif step == 18:
            w.DestroyJoint(made[id(w)][1])
            w.DestroyJoint(made[id(w)][0])
            d = _jdef(J.JOINT_WHEEL, 20, 21, (0.5, 0.0), (-0.5, 0.0))
            d.enableSpring, d.hertz, d.dampingRatio = 1, 3.0, 0.5
            j = w.CreateJoint(d)
            assert j == made[id(w)][0]                      # freed last, handed out first
        elif step == 25:
            w.DestroyBody(36)                               # takes its distance joint with it
            w.CreateJoint(_jdef(J.JOINT_REVOLUTE, 0, 33, (1.0, 9.0), (0.0, 0.0)))
    res = parity.run_pair(lambda w: scenes.add_pyramid(w, 8), steps=50, mutate=mutate, verbose=True)
    assert res["max_rel_err"] == 0.0

    # a world uploaded without any joint gets its first ones afterwards
    res = parity.run_pair(lambda w: scenes.mixed_drop(w, 40), steps=30, verbose=True,
                          mutate=lambda w, s: (w.CreateJoint(_jdef(J.JOINT_REVOLUTE, 1, 5, (1.0, 0.0), (0.0, 0.0))),
                                               w.CreateJoint(_jdef(J.JOINT_FILTER, 6, 7, (0.0, 0.0), (0.0, 0.0)))) if s == 8 else None)
    assert res["max_rel_err"] == 0.0


def test_set_filter_and_geometry_between_steps_bit_exact():
    """ShapeAPI.SetFilter (api/Shape.cs:157-166) and SetCircle/SetCapsule/SetSegment/SetPolygon (:289-326) through World.ResetProxy
    (src/Shape.cs:273-301): mask-only changes move the proxy, category or geometry changes re-create it (same node back from the LIFO free
    list), the shape's contacts are destroyed first and return in the next step if the filter still lets them."""
    def mutate(w, step):
        if step == 6:
            w.SetFilter(10, 1, 0)                      # mask 0: falls through everything from now on
            w.SetFilter(12, 2, 0xFFFFFFFFFFFFFFFF)     # new category: proxy re-created, still collides
            w.SetFilter(16, 1, 0xFFFFFFFFFFFFFFFF, -3)
            w.SetFilter(17, 1, 0xFFFFFFFFFFFFFFFF, -3)  # same negative group: these two stop colliding with each other
            w.SetFilter(18, 1, 0xFFFFFFFFFFFFFFFF, 0)   # unchanged: early return
        elif step == 11:
            w.SetShapeGeometry(9, geo.Circle((0.0, 0.0), 0.3))
            w.SetShapeGeometry(11, geo.MakeBox(0.3, 0.1))
            w.SetShapeGeometry(13, geo.Capsule((-0.25, 0.0), (0.25, 0.0), 0.1))
            w.ApplyMassFromShapes(5)                   # shape 11's body takes the new mass, the other two keep the old one
        elif step == 15:
            w.SetShapeGeometry(3, geo.Segment((-2.0, 0.5), (2.0, 0.9)))      # the static rounded box becomes a ramp
            w.SetShapeGeometry(5, geo.Capsule((0.0, -1.0), (0.0, 1.0), 0.15))  # the kinematic paddle's capsule grows
            w.SetFilter(0, 1, 0xFFFFFFFFFFFFFFFE)      # the ground segment ignores category 1: only shape 12's body stays on it
        elif step == 24:
            w.SetFilter(0, 1, 0xFFFFFFFFFFFFFFFF)
            w.SetFilter(10, 1, 0xFFFFFFFFFFFFFFFF)
    res = parity.run_pair(lambda w: scenes.mixed_drop(w, 60), steps=44, mutate=mutate, verbose=True)
    assert res["max_rel_err"] == 0.0


def test_set_body_type_between_steps_bit_exact():
    """BodyAPI.SetType (api/Body.cs:185-268) on bodies without joints: a box frozen to static and released again (it restarts at rest with
    an identity state, SolverSet.TransferBody), dynamic <-> kinematic in place (velocity kept, mass data recomputed), a spawned static
    platform turned kinematic, a compound body. Contacts of the body go first (end events), proxies change trees in shape-list order."""
    def mutate(w, step):
        if step == 6:
            w.ApplyForceToCenter(5, (30.0, 50.0))       # stays in the parked sim and acts in the first step after the body is dynamic again
            w.ApplyTorque(5, 4.0)
            w.SetType(5, api.STATIC_BODY)
            w.SetType(5, api.STATIC_BODY)               # same type: early return
        elif step == 10:
            w.SetType(20, api.KINEMATIC_BODY)
            w.SetLinearVelocity(20, (0.4, 0.0))
        elif step == 14:
            w.SetType(5, api.DYNAMIC_BODY)
            made = _spawn(w, 12.0, 4.0, "static")
            w.SetMotionLocks(30, False, False, True)
        elif step == 18:
            w.SetType(20, api.DYNAMIC_BODY)
            w.SetType(30, api.STATIC_BODY)
            w.SetType(37, api.KINEMATIC_BODY)           # the platform spawned at step 14
            w.SetAngularVelocity(37, 0.3)
        elif step == 24:
            w.SetType(30, api.DYNAMIC_BODY)             # comes back with its lock bits (Body.flags carries all three, api/Body.cs:807)
            w.SetType(36, api.STATIC_BODY)              # the top of the pyramid
    res = parity.run_pair(lambda w: scenes.add_pyramid(w, 8), steps=44, mutate=mutate, verbose=True)
    assert res["max_rel_err"] == 0.0

    def mutate2(w, step):
        if step == 7:
            w.SetType(2, api.STATIC_BODY)               # two shapes
            w.SetType(1, api.DYNAMIC_BODY)              # the kinematic paddle (box + capsule) becomes a free body
        elif step == 15:
            w.SetType(2, api.DYNAMIC_BODY)
            w.SetType(0, api.KINEMATIC_BODY)            # the ground itself: four proxies move from the static to the kinematic tree
    res = parity.run_pair(lambda w: scenes.mixed_drop(w, 60), steps=30, mutate=mutate2, verbose=True)
    assert res["max_rel_err"] == 0.0



def test_set_body_type_with_joints_bit_exact():
    """SetType on bodies that carry joints (api/Body.cs:202-241): every joint of the body is parked in the static set
    (SolverSet.TransferJoint src/SolverSet.cs:377-420: colour bits and the colour array give it up) and comes back with a fresh colour,
    appended, when one of its bodies is dynamic afterwards; a joint between two non-dynamic bodies stays parked (not simulated, not
    reported) until one of them turns dynamic again. Accumulated impulses travel with the joint."""
    def mutate(w, step):
        if step == 5:
            w.SetType(27, api.STATIC_BODY)              # an inner node of the revolute grid: four joints re-coloured against a static body
        elif step == 9:
            w.SetType(28, api.KINEMATIC_BODY)           # its neighbour: the joint 27-28 now links two non-dynamic bodies and is parked
            w.SetAngularVelocity(28, 0.5)
        elif step == 14:
            w.SetType(27, api.DYNAMIC_BODY)             # the parked joint returns
        elif step == 19:
            w.SetType(28, api.DYNAMIC_BODY)
            w.SetType(9, api.STATIC_BODY)
        elif step == 23:
            w.DestroyBody(9)                            # a static body with re-coloured joints
            w.SetType(40, api.STATIC_BODY)
            w.SetType(41, api.STATIC_BODY)              # parks the joint 40-41 ...
        elif step == 26:
            w.DestroyBody(41)                           # ... which is destroyed while parked
    res = parity.run_pair(lambda w: scenes.joint_grid(w, 8), steps=40, mutate=mutate, verbose=True)
    assert res["max_rel_err"] == 0.0

    holder = {}

    def build(w):
        holder["zoo"] = scenes.joint_zoo(w, copies=2, angles=False)

    def mutate2(w, step):
        if step == 6:
            w.SetType(1, api.KINEMATIC_BODY)
            w.SetType(3, api.STATIC_BODY)
        elif step == 14:
            w.SetType(1, api.DYNAMIC_BODY)
            w.SetType(3, api.DYNAMIC_BODY)
            w.SetType(0, api.KINEMATIC_BODY)            # the first copy's ground with every joint anchored to it
    res = parity.run_pair(build, steps=30, mutate=mutate2, verbose=True)
    assert res["max_rel_err"] == 0.0


def test_mutators_before_the_first_step_bit_exact():
    """Destroying, re-typing and re-shaping before the world was ever stepped: the first such call uploads the world (as the first step
    would) and runs the device path; creation afterwards goes through the after-upload kernels. Same result as the reference."""
    def build(w):
        scenes.add_pyramid(w, 7)
        w.DestroyBody(4)
        w.DestroyShape(9, True)
        w.SetType(12, api.STATIC_BODY)
        w.SetFilter(15, 2, 0xFFFFFFFFFFFFFFFF)
        w.SetShapeGeometry(16, geo.Circle((0.0, 0.0), 0.5))
        w.ApplyMassFromShapes(16)
        assert _spawn(w, 0.0, 12.0, "compound") == 4
        _spawn(w, 2.0, 13.0, "box")
    res = parity.run_pair(build, steps=30, verbose=True)
    assert res["max_rel_err"] == 0.0


class _Model:
    """Python-side book-keeping of what is alive, kept per backend (both must hand out the same ids)."""
    def __init__(self):
        self.bodies, self.types, self.shapes, self.joints, self.jointed = [], {}, {}, {}, {}


def _churn(seed, steps, n0):
    import random
    from box2d_b200 import capi as J
    models = {}

    def build(w):
        ground, paddle, bodies = scenes.mixed_drop(w, n0)
        m = models[id(w)] = _Model()
        m.bodies = list(bodies)
        for b in bodies:
            m.types[b] = api.DYNAMIC_BODY

    def mutate(w, step):
        m = models[id(getattr(w, 'w', w))]
        rnd = random.Random(seed * 1000 + step)
        for _ in range(rnd.randint(0, 4)):
            op = rnd.choice(("spawn", "spawn", "destroy", "retype", "joint", "unjoint", "filter", "teleport", "impulse", "geometry", "density", "shapekill", "locks", "addshape", "explode"))
            live = m.bodies
            if op == "spawn":
                b = _spawn(w, rnd.uniform(-8.0, 8.0), rnd.uniform(5.0, 12.0), rnd.choice(("box", "circle", "compound")), (rnd.uniform(-2, 2), rnd.uniform(-4, 0)))
                m.bodies.append(b); m.types[b] = api.DYNAMIC_BODY
            elif op == "destroy" and len(live) > 10:
                b = live.pop(rnd.randrange(len(live)))
                for j in [j for j, (a, c) in m.joints.items() if b in (a, c)]:
                    del m.joints[j]
                w.DestroyBody(b)
            elif op == "retype" and live:
                b = rnd.choice(live)                      # jointed or not: its joints are parked / re-coloured
                t = rnd.choice((api.STATIC_BODY, api.KINEMATIC_BODY, api.DYNAMIC_BODY))
                w.SetType(b, t); m.types[b] = t
            elif op == "joint" and len(live) > 2:
                a, c = rnd.sample(live, 2)
                if True:                                          # two non-dynamic bodies: the joint is created parked (src/Joint.cs:242-248)
                    d = _jdef(rnd.choice((J.JOINT_DISTANCE, J.JOINT_FILTER, J.JOINT_REVOLUTE)), a, c, (0.0, 0.0), (0.0, 0.0))
                    if d.type == J.JOINT_DISTANCE:               # the revolute spring evaluates Atan2 (tolerance-only): plain pins here
                        d.length, d.enableSpring, d.hertz, d.dampingRatio = 2.0, 1, 2.0, 0.7
                    d.collideConnected = rnd.randint(0, 1)
                    m.joints[w.CreateJoint(d)] = (a, c)
            elif op == "unjoint" and m.joints:
                j = rnd.choice(sorted(m.joints))
                del m.joints[j]
                w.DestroyJoint(j)
            elif op in ("filter", "geometry"):
                alive = [i for i, k in enumerate(w.ReadShapes()["proxyKey"].tolist()) if k != -1 and i > 5]   # not the terrain / paddle
                if alive:
                    sid = rnd.choice(alive)
                    if op == "filter":
                        w.SetFilter(sid, rnd.choice((1, 2, 4)), rnd.choice((0xFFFFFFFFFFFFFFFF, 0xFFFFFFFFFFFFFFFD, 0xFFFFFFFFFFFFFFFB)), rnd.choice((0, 0, -1)))
                    else:
                        w.SetShapeGeometry(sid, rnd.choice((geo.Circle((0.0, 0.0), 0.25), geo.MakeBox(0.3, 0.2), geo.Capsule((-0.2, 0.0), (0.2, 0.0), 0.1))))
            elif op == "teleport" and live:
                b = rnd.choice(live)
                w.SetTransform(b, (rnd.uniform(-8.0, 8.0), rnd.uniform(3.0, 10.0)), (1.0, 0.0))
            elif op == "impulse" and live:
                b = rnd.choice(live)
                w.ApplyLinearImpulseToCenter(b, (rnd.uniform(-1, 1), rnd.uniform(0, 2)))
            elif op == "density" and live:
                w.ApplyMassFromShapes(rnd.choice(live))
            elif op == "shapekill":
                alive = [i for i, k in enumerate(w.ReadShapes()["proxyKey"].tolist()) if k != -1 and i > 5]
                if alive:
                    w.DestroyShape(rnd.choice(alive), bool(rnd.randint(0, 1)))
            elif op == "explode":
                w.Explode((rnd.uniform(-8.0, 8.0), rnd.uniform(0.0, 6.0)), rnd.uniform(0.5, 3.0), rnd.choice((0.0, 1.0, 2.5)), rnd.uniform(-0.5, 1.0))
            elif op == "locks" and live:
                w.SetMotionLocks(rnd.choice(live), bool(rnd.randint(0, 1)), False, bool(rnd.randint(0, 1)))
            elif op == "addshape" and live:
                sd = scenes._shape_def()
                sd.density = rnd.choice((0.5, 1.0, 2.0))
                w.CreateCircleShape(rnd.choice(live), sd, geo.Circle((rnd.uniform(-0.3, 0.3), rnd.uniform(-0.3, 0.3)), 0.15))
    return build, mutate


@pytest.mark.parametrize("seed", [1, 2, 3, 4, 5, 6])
def test_random_churn_of_creation_and_destruction_bit_exact(seed):
    """A seeded random mix of spawn / DestroyBody / SetType / CreateJoint / DestroyJoint / SetTransform / impulses between steps, applied
    identically to both backends for 120 steps: ids handed out, events, colours and every body state must agree exactly. (This test found
    the sticky EnlargeBounds bit that a body carries into the static set.)"""
    build, mutate = _churn(seed, 120, 40)
    # continuous collision off: explosions make bodies fast next to the terrain segments, where the C# TimeOfImpact throws (see test_continuous_collision_where_the_reference_throws)
    res = parity.run_pair(build, steps=120, mutate=mutate, verbose=True, world_def=scenes.world_def(enable_continuous=False))
    assert res["max_rel_err"] == 0.0


def test_motion_locks_repeated_before_and_after_upload_bit_exact():
    """SetMotionLocks tests Body.flags, which carries all three lock bits after the first change (api/Body.cs:805-807): a repeated call with
    the same locks runs again and zeroes the locked velocity components once more — before the first step (host mirror) and after it."""
    def build(w):
        scenes.add_pyramid(w, 5)
        w.SetMotionLocks(3, True, False, False)
        w.SetLinearVelocity(3, (2.0, 1.0))
        w.SetMotionLocks(3, True, False, False)        # runs again: vx back to 0
        w.SetMotionLocks(4, True, True, True)
        w.SetLinearVelocity(4, (1.0, 1.0))
        w.SetMotionLocks(4, True, True, True)          # all three: Body.flags equals the request, early return, velocity stays

    def mutate(w, step):
        if step in (4, 9):
            w.SetLinearVelocity(3, (1.5, 0.5))
            w.SetMotionLocks(3, True, False, False)
            w.SetLinearVelocity(6, (1.0, 2.0))
            w.SetMotionLocks(6, False, True, False)
    res = parity.run_pair(build, steps=16, mutate=mutate, verbose=True)
    assert res["max_rel_err"] == 0.0


def test_explode_bit_exact():
    """WorldAPI.Explode (api/World.cs:703-776) between steps: GJK distance of every dynamic shape in reach to the centre (circles, capsules,
    boxes, rounded boxes, compound bodies), impulses applied in the tree's query order. Masks and a zero-falloff blast included."""
    def mutate(w, step):
        if step == 6:
            w.Explode((0.0, 1.5), 3.0, 2.0, 1.5)
        elif step == 12:
            w.Explode((-4.0, 0.5), 2.0, 0.0, 0.8)
            w.Explode((3.0, 4.0), 1.0, 4.0, -0.6)              # an implosion
        elif step == 18:
            w.Explode((0.0, 0.0), 50.0, 0.0, 0.2, 0x2)         # nobody carries category 2
            w.Explode((1.0, 0.3), 0.4, 0.3, 2.0)               # small: centre inside some shape (distance 0 -> centroid direction)
    res = parity.run_pair(lambda w: scenes.mixed_drop(w, 120), steps=30, mutate=mutate, verbose=True)
    assert res["max_rel_err"] == 0.0
    res = parity.run_pair(lambda w: scenes.add_pyramid(w, 12), steps=24, verbose=True,
                          mutate=lambda w, s: w.Explode((0.0, 3.0), 4.0, 3.0, 2.0) if s in (8, 16) else None)
    assert res["max_rel_err"] == 0.0


def test_joints_between_non_dynamic_bodies_are_parked_until_one_turns_dynamic():
    """World.CreateJoint puts a joint whose bodies are both non-dynamic into the static set (src/Joint.cs:242-248): it is not simulated and
    not reported; SetType(dynamic) on one of its bodies brings it into the graph (api/Body.cs:226-241). Before and after the first step."""
    from box2d_b200 import capi as J
    made = {}

    def build(w):
        scenes.mixed_drop(w, 40)
        made[id(w)] = [w.CreateJoint(_jdef(J.JOINT_REVOLUTE, 0, 1, (-5.0, 1.5), (0.0, 0.0)))]      # ground (static) - paddle (kinematic)
        assert len(w.ReadJoints()) == 0 and w.GetCounters().jointCount == 1

    def mutate(w, step):
        if step == 5:
            p = _spawn(w, 6.0, 3.0, "static")
            k = _spawn(w, 6.0, 5.0, "paddle", body_type=api.KINEMATIC_BODY)
            d = _jdef(J.JOINT_DISTANCE, p, k, (0.0, 0.0), (0.0, 0.0))
            d.length = 2.0
            made[id(w)] += [w.CreateJoint(d), p, k]
            assert len(w.ReadJoints()) == 0
        elif step == 10:
            w.SetType(1, api.DYNAMIC_BODY)                   # the paddle now swings on its pin
            assert w.ReadJoints()["jointId"].tolist() == [made[id(w)][0]]
        elif step == 16:
            w.SetType(made[id(w)][3], api.DYNAMIC_BODY)
        elif step == 24:
            w.DestroyJoint(made[id(w)][0])
            w.SetType(made[id(w)][3], api.STATIC_BODY)       # parked again
            w.DestroyJoint(made[id(w)][1])                   # destroyed while parked
    res = parity.run_pair(build, steps=34, mutate=mutate, verbose=True)
    assert res["max_rel_err"] == 0.0


def _bounce_mutate(w, step):
    """Kicks some bouncers off the ground (their contacts stop touching: swap-remove in the colour arrays), destroys two bodies and
    turns restitution on/off on live shapes, so the 8-lane restitution groups and the hit-event order keep changing."""
    if step in (24, 40):
        for b in range(3, 40, 5):
            w.ApplyLinearImpulseToCenter(b, (0.3, 2.5 + 0.1 * (b % 3)))
    elif step == 34:
        w.DestroyBody(7)
        w.DestroyBody(20)
    elif step == 29:
        for sid in range(2, 30, 3):
            w.SetSurfaceMaterial(sid, 0.6, 0.4 if sid % 2 else 0.0)


@pytest.mark.parametrize("mode", ["auto", "cta", "warp", "grid", "gridsync"])
def test_restitution_mixed_groups_and_hit_events_bit_exact(mode, monkeypatch):
    """ApplyRestitutionTask (src/ContactSolver.cs:913-989): the 8-lane groups are 8 consecutive entries of a colour's contactSims array
    in the reference's own order (append / swap-remove), the AllZeroW skip and quirk Q1 are per group; hit events (src/Solver.cs:1151-1190)
    come out in (colour, array position) order. Every solver variant must reproduce them bit for bit."""
    if mode in ("cta", "warp"):
        monkeypatch.setenv("B2CUDA_ISLAND_MODE", mode)
    elif mode in ("grid", "gridsync"):
        monkeypatch.setenv("B2CUDA_ISLAND_MAX_BODIES", "0")      # every island down the grid / cluster path
        if mode == "gridsync":
            monkeypatch.setenv("B2CUDA_CLUSTER", "0")
    res = parity.run_pair(lambda w: scenes.bounce_mix(w), steps=70, mutate=_bounce_mutate, verbose=True)
    assert res["max_rel_err"] == 0.0
    assert res["hits"] > 20          # the scene does produce hit events, in the reference's order (compared every step by run_pair)


def test_restitution_in_pyramids_and_overflow_bit_exact():
    """Restitution on a few boxes of a pile that also owns overflow-colour contacts: the scalar overflow restitution (Q2, Q3) and the
    SIMD-group restitution in one step."""
    def build(w):
        scenes.overflow_plank(w)
        scenes.add_pyramid(w, 6, origin=(30.0, 0.0))

    def mutate(w, step):
        if step == 0:
            for sid in range(1, 60, 4):
                w.SetSurfaceMaterial(sid, 0.6, 0.5)
    res = parity.run_pair(build, steps=40, mutate=mutate, verbose=True)
    assert res["max_rel_err"] == 0.0
    assert res["max_overflow"] > 0


def test_joint_local_frames_and_collide_connected_between_steps_bit_exact():
    """JointAPI.SetLocalFrameA/B (api/Joint.cs:44,50) and SetCollideConnected (:56-75) between steps: chains of overlapping boxes hinged
    to each other; switching collision on buffers moves (contacts and begin events appear the next step), switching it off destroys the
    contacts between the two bodies in contact-list order; the frame setters move the hinge."""
    from box2d_b200 import geometry as G
    holder = {}

    def build(w):
        ground = w.CreateBody(api.default_body_def())
        w.CreatePolygonShape(ground, scenes._shape_def(), G.MakeOffsetBox(40.0, 1.0, (0.0, -1.0)))
        joints = []
        for k in range(6):
            prev = None
            for i in range(5):
                b = scenes._dynamic_body(w, -20.0 + 8.0 * k + 0.6 * i, 0.5 + 0.3 * i)
                w.CreatePolygonShape(b, scenes._shape_def(), G.MakeBox(0.5, 0.25))
                if i % 2 == 0:
                    w.CreateCircleShape(b, scenes._shape_def(), G.Circle((0.0, 0.2), 0.3))     # some bodies carry two shapes: shapeCount decides whose shapes move
                if prev is not None:
                    jd = api.default_revolute_joint_def()
                    jd.bodyIdA, jd.bodyIdB = prev, b
                    jd.localFrameA = api.Transform(api.Vec2(0.3, 0.15), api.Rot(1.0, 0.0))
                    jd.localFrameB = api.Transform(api.Vec2(-0.3, -0.15), api.Rot(1.0, 0.0))
                    jd.collideConnected = int(k % 2 == 1)
                    joints.append(w.CreateRevoluteJoint(jd))
                prev = b
        holder["joints"] = joints

    def mutate(w, step):
        js = holder["joints"]
        if step == 4:
            for j in js[0::3]:
                w.SetCollideConnected(j, True)
            for j in js[1::3]:
                w.SetCollideConnected(j, False)
        if step == 9:
            for j in js[0::4]:
                w.SetLocalFrameA(j, api.Transform(api.Vec2(0.25, 0.2), api.Rot(1.0, 0.0)))
            for j in js[2::4]:
                w.SetLocalFrameB(j, api.Transform(api.Vec2(-0.35, -0.1), api.Rot(1.0, 0.0)))
        if step == 14:
            for j in js:
                w.SetCollideConnected(j, False)
        if step == 20:
            for j in js[::2]:
                w.SetCollideConnected(j, True)
                w.SetCollideConnected(j, True)              # a repeated call returns at once

    res = parity.run_pair(build, steps=36, verbose=True, mutate=mutate)
    assert res["max_rel_err"] == 0.0
    # before the first step the call only edits the host mirror (moves are buffered, no contacts exist yet)
    def build2(w):
        build(w)
        for j in holder["joints"][::2]:
            w.SetCollideConnected(j, True)
        w.SetLocalFrameA(holder["joints"][1], api.Transform(api.Vec2(0.2, 0.1), api.Rot(1.0, 0.0)))
    res = parity.run_pair(build2, steps=12, verbose=True)
    assert res["max_rel_err"] == 0.0


def test_body_sleep_enable_bullet_surface_while_sleep_is_not_built():
    """api/Body.cs:664-842: the calls that only wake or store succeed, the calls that need islands / the disabled set are refused
    loudly (B2C_E_UNSUPPORTED = -2), and none of them disturbs the simulation (SetBullet: test_bullets_*)."""
    cuda = api.World(api.CudaBackend(), scenes.world_def())
    try:
        scenes.add_pyramid(cuda, 4)
        cuda.Step(1.0 / 60.0, 4)
        before = cuda.ReadBodies().tobytes()
        b = 3
        cuda.SetSleepThreshold(b, 0.2)
        cuda.EnableSleep(b, False)
        cuda.SetAwake(b, True)
        cuda.WakeTouching(b)
        cuda.WakeTouching(0)                 # the static ground
        cuda.SetAwake(0, False)              # a static body: nothing to do
        cuda.EnableBody(b)
        cuda.SetBullet(b, False)
        for call in (lambda: cuda.SetAwake(b, False), lambda: cuda.DisableBody(b)):
            with pytest.raises(api.B2Error, match="rc=-2"):
                call()
        with pytest.raises(api.B2Error):
            cuda.SetSleepThreshold(9999, 0.1)
        assert cuda.ReadBodies().tobytes() == before
    finally:
        cuda.Destroy()


def test_sensor_events_bit_exact():
    """src/Sensor.cs: static, kinematic-borne and dynamic-borne sensors, visitors with events off / filtered out, a sensor with its own
    events off; sensor begin / end events (sensor order, visitors ascending) and the overlap sets must equal the oracle's every step."""
    holder = {}

    def build(w):
        holder[w.backend.name] = scenes.sensor_field(w)


    # continuous collision off: the falling boxes and capsules become fast next to the static circle / capsule sensors, where the C#
    # TimeOfImpact indexes past proxy A and throws (test_continuous_collision_where_the_reference_throws); fast_rain covers continuous sensor hits
    wd = scenes.world_def(enable_continuous=False)
    cuda = api.World(api.CudaBackend(), wd)
    from oracle_backend import OracleBackend
    orc = api.World(OracleBackend(), wd)
    try:
        build(cuda), build(orc)
        total = 0
        for step in range(90):
            cuda.Step(1.0 / 60.0, 4), orc.Step(1.0 / 60.0, 4)
            a, b = cuda.GetSensorEvents(), orc.GetSensorEvents()
            for x, y, name in ((a[0], b[0], "begin"), (a[1], b[1], "end")):
                assert x.tobytes() == y.tobytes(), f"step {step}: sensor {name} events differ: {x.tolist()[:6]} vs {y.tolist()[:6]}"
            total += len(a[0]) + len(a[1])
            if step % 10 == 9:
                for sid in holder["cuda"][0]:
                    assert cuda.GetSensorOverlaps(sid) == orc.GetSensorOverlaps(sid), f"step {step}: overlaps of sensor shape {sid}"
                assert cuda.ReadBodies().tobytes() == orc.ReadBodies().tobytes()
        assert total > 60, f"scene produced only {total} sensor events"
    finally:
        cuda.Destroy(), orc.Destroy()


def test_sensor_mutations_between_steps_bit_exact():
    """DestroyShape of a sensor with visitors (end events go to the running end buffer, World.sensors swap-remove), EnableSensorEvents on
    sensors and visitors, a sensor created after the first step, DestroyBody of a body that carries a sensor, visitor destroyed inside a sensor."""
    from box2d_b200 import geometry as G
    holder = {}

    def build(w):
        holder[id(w)] = scenes.sensor_field(w)

    def mutate(w, step):
        sensors, bodies, carrier, sweeper = holder[id(w)]
        if step == 30:
            w.DestroyShape(sensors[1])                        # the circle zone, with visitors inside
        if step == 36:
            w.EnableSensorEvents(sensors[0], False)           # the sensor stops reporting: end events for everything inside
            w.EnableSensorEvents(sensors[3], True)            # the silent one starts
        if step == 42:
            zone = w.CreateBody(api.default_body_def())
            holder[(id(w), "late")] = w.CreatePolygonShape(zone, scenes._shape_def(isSensor=1, enableSensorEvents=1), G.MakeOffsetBox(6.0, 1.0, (0.0, 1.0)))
        if step == 48:
            w.DestroyBody(carrier)                            # takes its sensor with it
            w.DestroyBody(bodies[7])
            w.EnableSensorEvents(sensors[0], True)
        if step == 55:
            w.DestroyShape(holder[(id(w), "late")])
            w.DestroyBody(sweeper)

    res = parity.run_pair(build, steps=70, verbose=True, mutate=mutate, world_def=scenes.world_def(enable_continuous=False))   # see test_sensor_events_bit_exact
    assert res["max_rel_err"] == 0.0 and res["sensor_events"] > 40


def test_continuous_collision_bit_exact():
    """World.SolveContinuous + Distance.TimeOfImpact as the port behaves (src/Solver.cs:224-401, src/Distance.cs:529-646): fast bodies
    against thin static polygons; bodies stopped at a time of impact (HadTimeOfImpact), bodies the port lets tunnel (its cached-simplex
    read, quirk Q19, derails most sweeps after the first GJK call) and continuous sensor hits must all equal the oracle's."""
    res = parity.run_pair(lambda w: scenes.fast_rain(w, n=64), steps=45, verbose=True)
    assert res["max_rel_err"] == 0.0
    assert res["toi"] > 100 and res["toi"] == res["toi_oracle"], res
    res = parity.run_pair(lambda w: scenes.fast_rain(w, n=40, seed=77), steps=30, substeps=2, dt=1.0 / 30.0)
    assert res["max_rel_err"] == 0.0 and res["toi"] == res["toi_oracle"]


@pytest.mark.parametrize("walls", ["round", "segment"])
def test_continuous_collision_where_the_reference_throws(walls):
    """A fast polygon or capsule swept against a static shape with fewer points makes the C# code index past proxy A
    (Distance.MakeSimplexFromCache, src/Distance.cs:116-132): IndexOutOfRangeException out of WorldAPI.Step. Both sides must report it,
    in the same step, instead of inventing a result."""
    from oracle_backend import OracleBackend
    failed = {}
    for name, be in (("cuda", api.CudaBackend()), ("oracle", OracleBackend())):
        w = api.World(be, scenes.world_def())
        try:
            scenes.fast_rain(w, walls=walls)
            failed[name] = None
            for step in range(60):
                try:
                    w.Step(1.0 / 60.0, 4)
                    w.GetCounters()                    # the device reports at the next synchronising call
                except api.B2Error as e:
                    assert "throws" in str(e)
                    failed[name] = step
                    break
        finally:
            w.Destroy()
    assert failed["cuda"] == failed["oracle"], failed
    if walls == "round":
        assert failed["cuda"] is not None


def test_continuous_parallel_motion_is_not_stopped():
    """Known answer: a fast circle flying parallel to a thin static box, 0.1 m above it (distance 0.1 > target 0.075 + tolerance: TimeOfImpact returns Separated), is a TOI candidate every step (the swept boxes
    overlap the fat box) and is never stopped: x advances by v*dt per step exactly as free flight, on both backends."""
    from box2d_b200 import geometry as G

    def build(w):
        ground = w.CreateBody(api.default_body_def())
        w.CreatePolygonShape(ground, scenes._shape_def(), G.MakeOffsetBox(50.0, 0.05, (0.0, -0.05)))
        b = scenes._dynamic_body(w, -20.0, 0.1, linearVelocity=api.Vec2(120.0, 0.0), gravityScale=0.0)
        w.CreateCircleShape(b, scenes._shape_def(), G.Circle((0.0, 0.0), 0.08))

    res = parity.run_pair(build, steps=12)
    assert res["max_rel_err"] == 0.0 and res["toi"] >= 10
    w = api.World(api.CudaBackend(), scenes.world_def())
    try:
        build(w)
        h = np.float32(np.float32(1.0 / 60.0) / np.float32(4))
        x = np.float32(-20.0)
        for _ in range(12):
            w.Step(1.0 / 60.0, 4)
            dx = np.float32(0.0)
            for _ in range(4):
                dx = np.float32(dx + np.float32(h * np.float32(120.0)))
            x = np.float32(x + dx)
        assert np.float32(w.ReadBodies()[1]["transform"]["p"]["x"]) == x
    finally:
        w.Destroy()


def test_bullets_bit_exact():
    """BodyDef.isBullet (src/Solver.cs:447-449, 728-738, 1210-1218, 1238-1272): fast bullets wait for the bullet stage, sweep the static,
    kinematic and dynamic trees after the refit, stop at thin dynamic and kinematic plates (or tunnel where the port's cached-simplex read
    derails the sweep), ignore each other and jointed bodies with collideConnected off; their shapes enter the move set in the refit and
    their proxies grow after the stage. State, events, move counts and TOI candidate counts must equal the oracle's every step."""
    res = parity.run_pair(lambda w: scenes.bullet_range(w, n=24), steps=45, verbose=True)
    assert res["max_rel_err"] == 0.0
    assert res["toi"] > 300 and res["toi"] == res["toi_oracle"], res
    res = parity.run_pair(lambda w: scenes.bullet_range(w, n=40, seed=31), steps=30, substeps=2, dt=1.0 / 30.0)
    assert res["max_rel_err"] == 0.0 and res["toi"] == res["toi_oracle"]


def test_bullet_move_events_and_set_bullet_between_steps():
    """BodyAPI.SetBullet (api/Body.cs:835-842) flips BodySim.flags between steps; the body move events carry the pose the bullet stage
    left (src/Solver.cs:366-367). Bullets are switched off and on while they fly; a static body takes the call without effect."""
    from oracle_backend import OracleBackend
    cuda = api.World(api.CudaBackend(), scenes.world_def())
    orc = api.World(OracleBackend(), scenes.world_def())
    try:
        made = [scenes.bullet_range(w, n=24, seed=11) for w in (cuda, orc)]
        assert made[0] == made[1]
        ground, plates, proj = made[0]
        had_toi = 0
        for step in range(40):
            for w in (cuda, orc):
                if step == 2:
                    for b in proj[::2]:
                        w.SetBullet(b, False)
                    w.SetBullet(ground, True)
                if step == 5:
                    for b in proj[::2]:
                        w.SetBullet(b, True)
                    w.SetBullet(proj[2], True)           # one that was never a bullet
                    w.SetBullet(ground, False)
                w.Step(1.0 / 60.0, 4)
            a, b = cuda.GetBodyMoveEvents(), orc.GetBodyMoveEvents()
            assert a.tobytes() == b.tobytes(), f"step {step}: body move events differ"
            ra, rb = cuda.GetBodyEvents(), orc.GetBodyEvents()
            assert ra.tobytes() == rb.tobytes(), f"step {step}: body read-back differs"
            assert cuda.ReadBodies().tobytes() == orc.ReadBodies().tobytes(), f"step {step}"
            assert cuda.ReadShapes().tobytes() == orc.ReadShapes().tobytes(), f"step {step}"
            had_toi += int(np.count_nonzero(ra["flags"] & 0x40))
            ca, co = cuda.GetCounters(), orc.GetCounters()
            assert (ca.moveCount, ca.contactCount, ca.toiCandidates) == (co.moveCount, co.contactCount, co.toiCandidates)
        assert had_toi > 0
    finally:
        cuda.Destroy()
        orc.Destroy()


def test_bullets_where_the_reference_throws():
    """A bullet swept against a dynamic circle indexes past proxy A in the C# code (Distance.MakeSimplexFromCache): both sides report the
    step as failed, in the same step."""
    from oracle_backend import OracleBackend
    failed = {}
    for name, be in (("cuda", api.CudaBackend()), ("oracle", OracleBackend())):
        w = api.World(be, scenes.world_def())
        try:
            scenes.bullet_range(w, targets="round")
            failed[name] = None
            for step in range(30):
                try:
                    w.Step(1.0 / 60.0, 4)
                    w.GetCounters()
                except api.B2Error as e:
                    assert "throws" in str(e)
                    failed[name] = step
                    break
        finally:
            w.Destroy()
    assert failed["cuda"] == failed["oracle"] and failed["cuda"] is not None, failed


def test_overflow_joints_bit_exact():
    """A body with more joints than colours (src/ConstraintGraph.cs:91-128): the surplus joints sit in the overflow colour and are prepared,
    warm-started and solved serially before the overflow contacts of each stage (src/Joint.cs:377-394, src/Solver.cs:620,633,648,669),
    relax keeps useBias, and they report no joint events. Joints are destroyed and created between steps, and a hub changes type and back."""
    holder = {}

    def build(w):
        holder[w.backend.name] = scenes.joint_hub(w)

    def mutate(w, step):
        hub, sats, joints = holder[w.backend.name]
        if step == 6:
            w.DestroyJoint(joints[22])                      # an overflow joint
            w.DestroyJoint(joints[3])                       # frees a colour of the hub: the next joint takes it
        if step == 9:
            for k in (3, 22):
                d = api.default_joint_def(api.capi.JOINT_DISTANCE)
                d.bodyIdA, d.bodyIdB = hub, sats[k]
                d.length = 2.5
                d.enableSpring, d.hertz, d.dampingRatio = 1, 3.0, 0.3
                joints.append(w.CreateJoint(d))
        if step == 14:
            w.SetType(hub, api.KINEMATIC_BODY)              # hub joints stay in the graph (the satellites are dynamic) with fresh colours
        if step == 20:
            w.SetType(hub, api.DYNAMIC_BODY)

    res = parity.run_pair(build, steps=60, mutate=mutate, verbose=True)
    assert res["max_rel_err"] == 0.0 and res["max_overflow"] >= 11, res


def test_contact_rows_grow_between_steps():
    """The reference's contact array, colour arrays, pair set and event lists grow without bound. Here the contact-sized device rows grow
    (a) when shapes created between steps outgrow the first upload's reserve and (b) when a step leaves them three quarters full; ids, colour
    array order, pending end events and every later step stay bit-exact."""
    holder = {}

    def build(w):
        holder[w.backend.name] = scenes.add_pyramid(w, 3)

    def mutate(w, step):
        if step == 3:
            scenes.add_pyramid(w, 30, origin=(60.0, 0.0))          # 465 bodies: 12 per shape passes the 4096 rows of the first upload
        if step == 6:
            w.DestroyBody(holder[w.backend.name][1][0])            # end events pending while ...
            scenes.add_pyramid(w, 24, origin=(-60.0, 0.0))         # ... the rows grow again
    res = parity.run_pair(build, steps=14, mutate=mutate, verbose=True)
    assert res["max_rel_err"] == 0.0 and res["contacts"] > 1500, res

    # (b): a fixed small reserve, contacts pile up gradually
    wd = scenes.world_def(contactCapacity=160)
    res = parity.run_pair(lambda w: scenes.mixed_drop(w, n=90, seed=5), steps=120, world_def=wd, full_every=10)
    assert res["max_rel_err"] == 0.0 and res["contacts"] > 160, res
