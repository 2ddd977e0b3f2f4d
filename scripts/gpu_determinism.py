"""Run the CUDA step several times on the same scene and report the first step / field where two runs differ.
Debug aid for races (the step must be run-to-run deterministic); not part of the test suite."""
import os
import sys
import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
from box2d_b200 import api, scenes  # noqa: E402


def run(build, steps):
    w = api.World(api.CudaBackend(), scenes.world_def())
    out = []
    try:
        build(w)
        for _ in range(steps):
            w.Step(1.0 / 60.0, 4)
            if os.environ.get("DET_SLEEP"):
                import time
                time.sleep(float(os.environ["DET_SLEEP"]))
            if os.environ.get("DET_FULL"):
                w.GetContactEvents(); w.ReadShapes(); w.ReadJoints()
            b = w.ReadBodies()
            c = w.ReadContacts()
            cnt = w.GetCounters()
            out.append((b, c, cnt.moveCount, cnt.contactCount, cnt.touchingContactCount, w.DebugChecksums()))
    finally:
        w.Destroy()
    return out


def fields(x):
    if isinstance(x, dict):
        return x.items()
    if hasattr(x, "dtype") and x.dtype.names:
        return [(n, x[n]) for n in x.dtype.names]
    return [("v", np.asarray(x))]


def first_diff(a, b):
    for step, (sa, sb) in enumerate(zip(a, b)):
        msgs = []
        for k, (name, idx) in enumerate((("bodies", 0), ("contacts", 1))):
            for f, va in fields(sa[idx]):
                vb = dict(fields(sb[idx]))[f]
                va, vb = np.ascontiguousarray(va), np.ascontiguousarray(vb)
                if va.shape != vb.shape:
                    msgs.append(f"{name}.{f} shape {va.shape} vs {vb.shape}")
                    continue
                ne = va.view(np.uint8).reshape(va.shape[0], -1) != vb.view(np.uint8).reshape(vb.shape[0], -1) if va.size else np.zeros((0, 1), bool)
                rows = np.nonzero(ne.any(axis=1))[0]
                if rows.size:
                    r = rows[0]
                    msgs.append(f"{name}.{f}: {rows.size} rows differ {rows[:12].tolist()}, first row {r}:\n        {va[r]}\n     vs {vb[r]}")
        for k, nm in ((2, "moveCount"), (3, "contactCount"), (4, "touching")):
            if sa[k] != sb[k]:
                msgs.append(f"{nm} {sa[k]} vs {sb[k]}")
        if sa[5] != sb[5]:
            msgs.append("checksum slots differing: " + str([k for k in range(16) if sa[5][k] != sb[5][k]]) + f" slot10 {sa[5][10]} vs {sb[5][10]} slot11 {sa[5][11]} vs {sb[5][11]}")
        if msgs:
            return step, msgs
    return None


def main():
    scene = sys.argv[1] if len(sys.argv) > 1 else "mixed"
    steps = int(sys.argv[2]) if len(sys.argv) > 2 else 60
    runs = int(sys.argv[3]) if len(sys.argv) > 3 else 6
    build = {"mixed": lambda w: scenes.mixed_drop(w), "tumbler": lambda w: scenes.tumbler(w, n=600, half=6.0),
             "tumbler4k": lambda w: scenes.tumbler(w, n=4000, half=10.5), "pyramid": lambda w: scenes.add_pyramid(w, 20)}[scene]
    ref = run(build, steps)
    bad = 0
    for r in range(runs):
        cur = run(build, steps)
        d = first_diff(ref, cur)
        if d:
            bad += 1
            print(f"[{scene}] run {r}: first difference at step {d[0]}")
            for m in d[1][:12]:
                print("    ", m)
    print('debug: color conflicts (slot 10):', [x[5][10] for x in ref if x[5][10]][:10], 'partition errors (slot 11):', [(k, x[5][11]) for k, x in enumerate(ref) if x[5][11]][:10])
    import hashlib
    print("hash", hashlib.md5(b"".join(x[0].tobytes() + x[1].tobytes() for x in ref)).hexdigest()[:10])
    print(f"[{scene}] env={ {k: v for k, v in os.environ.items() if k.startswith('B2CUDA_')} } nondeterministic runs: {bad}/{runs}")


if __name__ == "__main__":
    main()
