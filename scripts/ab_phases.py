#!/usr/bin/env python
"""Phase times of one workload on the library B2CUDA_LIB points at (A/B runs of two builds on one box)."""
import argparse, json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from box2d_b200 import api, scenes
ap = argparse.ArgumentParser()
ap.add_argument("--workload", default="many_pyramids")
ap.add_argument("--warmup", type=int, default=60)
ap.add_argument("--steps", type=int, default=200)
a = ap.parse_args()
w = api.World(api.CudaBackend(), scenes.world_def())
scenes.SCENES[a.workload](w)
for _ in range(a.warmup):
    w.Step(1 / 60, 4)
w.Synchronize(); w.GetProfile()
acc = {}
done = 0
while done < a.steps:
    n = min(32, a.steps - done)
    for _ in range(n):
        w.Step(1 / 60, 4); w.Synchronize()
    p = w.GetProfile()
    for f in ("pairs", "collide", "treeRebuild", "prepareConstraints", "solveConstraints", "transforms", "step"):
        acc[f] = acc.get(f, 0.0) + getattr(p, f) * n
    done += n
w.TimerStart()
for _ in range(a.steps):
    w.Step(1 / 60, 4)
ms = w.TimerStop()
print(json.dumps({"lib": os.environ.get("B2CUDA_LIB", "in-tree"), "workload": a.workload, "back_to_back_ms": round(ms / a.steps, 4), **{k: round(v / a.steps * 1000, 1) for k, v in acc.items()}}))
w.Destroy()
