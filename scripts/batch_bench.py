#!/usr/bin/env python
"""Times b2cuda_batch_step on N independent 10-row pyramids (BASELINE.json configs[4]) on one GPU: world-steps/s and body-substeps/s."""
import argparse, json, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from box2d_b200 import api, scenes

ap = argparse.ArgumentParser()
ap.add_argument("--worlds", type=int, default=8192)
ap.add_argument("--rows", type=int, default=10)
ap.add_argument("--steps", type=int, default=100)
ap.add_argument("--warmup", type=int, default=30)
a = ap.parse_args()
be = api.CudaBackend()
t0 = time.perf_counter()
b = api.Batch(be, a.worlds, scenes.world_def())
for w in b.worlds:
    scenes.add_pyramid(w, a.rows)
build_s = time.perf_counter() - t0
t0 = time.perf_counter()
b.Step(1 / 60, 4); b.Synchronize()
first_s = time.perf_counter() - t0
for _ in range(a.warmup):
    b.Step(1 / 60, 4)
b.Synchronize()
b.worlds[0].GetProfile(); b.worlds[a.worlds // 2].GetProfile()
b.TimerStart()
for _ in range(a.steps):
    b.Step(1 / 60, 4)
ms = b.TimerStop()
prof = []
for k in (0, a.worlds // 2):
    p = b.worlds[k].GetProfile()
    prof.append({f: round(getattr(p, f) * 1000, 2) for f in ("pairs", "collide", "treeRebuild", "prepareConstraints", "solveConstraints", "transforms", "step")})
s = b.Stats()
bodies = s.awakeBodies
print(json.dumps({"worlds": a.worlds, "rows": a.rows, "ms_per_step": ms / a.steps, "world_steps_per_s": a.worlds * a.steps / (ms * 1e-3),
                  "body_substeps_per_s": bodies * 4 * a.steps / (ms * 1e-3), "build_s": build_s, "first_step_s": first_s,
                  "touching": s.touchingContacts, "cta_phase_us": prof, "errorFlags": s.errorFlags}))
b.Destroy()
