"""Parity harness: drives the CUDA library and the CPU oracle through the same Box2D.API mirror and compares the
reference-visible state after every step. TEST INFRASTRUCTURE (imports the oracle)."""
import numpy as np
from box2d_b200 import api, scenes
from oracle_backend import OracleBackend


def _flat(a):
    """structured array -> dict of flat leaf arrays"""
    out = {}

    def rec(prefix, arr):
        if arr.dtype.names:
            for n in arr.dtype.names:
                rec(prefix + "." + n if prefix else n, arr[n])
        else:
            out[prefix] = np.asarray(arr)
    rec("", a)
    return out


def compare_struct(name, a, b, exact_fields=None, skip=(), report=None):
    """Bit-exact comparison of two structured arrays; returns max relative error over float leaves."""
    assert len(a) == len(b), f"{name}: length {len(a)} vs {len(b)}"
    fa, fb = _flat(a), _flat(b)
    worst = 0.0
    for k in fa:
        if k in skip:
            continue
        x, y = fa[k], fb[k]
        if x.dtype.kind == "f":
            same = (x.view(np.uint32) == y.view(np.uint32)) | ((x == 0) & (y == 0))
            if not same.all():
                denom = np.maximum(np.abs(y), 1e-6)
                err = float(np.max(np.abs(x.astype(np.float64) - y) / denom))
                worst = max(worst, err)
                if report is not None:
                    idx = np.argwhere(~same)[:3].tolist()
                    report.append(f"{name}.{k}: {int((~same).sum())} mismatches, max rel {err:.3e}, first at {idx}")
        else:
            if not np.array_equal(x, y):
                bad = np.argwhere(x != y)[:3].tolist()
                raise AssertionError(f"{name}.{k}: integer mismatch at {bad}: {x[tuple(bad[0])]} vs {y[tuple(bad[0])]}")
    return worst


def snapshot(w, full=True):
    ev = w.GetContactEvents()
    out = dict(begin=ev[0], end=ev[1], hit=ev[2], counters=w.GetCounters(), joint_events=np.asarray(w.GetJointEvents(), dtype=np.int32))
    out["sensor_begin"], out["sensor_end"] = w.GetSensorEvents()
    if full:
        out.update(bodies=w.ReadBodies(), contacts=w.ReadContacts(), shapes=w.ReadShapes(), joints=w.ReadJoints())
    return out


def run_pair(build, steps=10, dt=1.0 / 60.0, substeps=4, world_def=None, tol=0.0, verbose=False, mutate=None, full_every=1):
    """Builds the same scene on both backends, steps both, compares after every step. Returns summary dict.
    full_every > 1 (long full-size runs): events, joint events and counters are still compared after EVERY step; the bulk state
    (contacts, bodies, shapes, joints) after the first 10 steps, then every `full_every`-th step and after the last one."""
    cuda = api.World(api.CudaBackend(), world_def if world_def is not None else scenes.world_def())
    orc = api.World(OracleBackend(), world_def if world_def is not None else scenes.world_def())
    try:
        build(cuda)
        build(orc)
        worst = 0.0
        notes = []
        max_overflow = 0
        hits_total = 0
        sensor_total = 0
        for step in range(steps):
            if mutate is not None:
                mutate(cuda, step)
                mutate(orc, step)
            cuda.Step(dt, substeps)
            orc.Step(dt, substeps)
            full = step < 10 or step % full_every == 0 or step == steps - 1
            a, b = snapshot(cuda, full), snapshot(orc, full)
            rep = []
            # exact: pair set / ids / colors / manifold ids / events
            for key in ("begin", "end", "hit", "sensor_begin", "sensor_end"):   # hit events: reference order = colour, then position in the colour's array; sensor events: sensor order, visitors ascending
                err = compare_struct(f"step{step}.{key}", a[key], b[key], report=rep)
                assert err == 0.0, f"step{step} {key} events differ: " + "; ".join(rep[:4])
            assert a["joint_events"].tolist() == b["joint_events"].tolist(), f"step{step} joint events {a['joint_events']} vs {b['joint_events']}"
            hits_total += len(a["hit"])
            sensor_total += len(a["sensor_begin"]) + len(a["sensor_end"])
            ca, cb = a["counters"], b["counters"]
            max_overflow = max(max_overflow, ca.colorCounts[23])
            for f in ("bodyCount", "awakeBodyCount", "shapeCount", "jointCount", "contactCount", "touchingContactCount", "contactIdCapacity", "activeColorCount", "moveCount", "newPairCount"):
                assert getattr(ca, f) == getattr(cb, f), f"step{step} counters.{f}: {getattr(ca, f)} vs {getattr(cb, f)}"
            assert list(ca.colorCounts) == list(cb.colorCounts), f"step{step} colorCounts {list(ca.colorCounts)} vs {list(cb.colorCounts)}"
            e = 0.0
            if full:
                e = compare_struct(f"step{step}.contacts", a["contacts"], b["contacts"], report=rep)
                e = max(e, compare_struct(f"step{step}.bodies", a["bodies"], b["bodies"], report=rep))
                e = max(e, compare_struct(f"step{step}.shapes", a["shapes"], b["shapes"], skip=("enlarged",), report=rep))
                e = max(e, compare_struct(f"step{step}.joints", a["joints"], b["joints"], report=rep))
            if rep and verbose:
                print("\n".join(rep[:12]))
            if e > tol:
                raise AssertionError(f"step {step}: max rel err {e:.3e} > {tol}\n" + "\n".join(rep[:12]))
            worst = max(worst, e)
            notes.extend(rep)
        c = cuda.GetCounters()
        co = orc.GetCounters()
        return dict(max_rel_err=worst, steps=steps, contacts=c.contactCount, touching=c.touchingContactCount, bodies=c.bodyCount,
                    colors=c.activeColorCount, toi=c.toiCandidates, toi_oracle=co.toiCandidates, max_overflow=max_overflow, hits=hits_total, sensor_events=sensor_total)
    finally:
        cuda.Destroy()
        orc.Destroy()
