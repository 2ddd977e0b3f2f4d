#!/usr/bin/env python
"""Back-to-back device time per step of one workload (no profile reads): A/B runs of launch-path switches."""
import argparse, json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from box2d_b200 import api, scenes
ap = argparse.ArgumentParser()
ap.add_argument("--workload", default="many_pyramids")
ap.add_argument("--warmup", type=int, default=60)
ap.add_argument("--steps", type=int, default=300)
ap.add_argument("--repeats", type=int, default=5)
a = ap.parse_args()
w = api.World(api.CudaBackend(), scenes.world_def())
scenes.SCENES[a.workload](w)
for _ in range(a.warmup):
    w.Step(1 / 60, 4)
w.Synchronize()
out = []
for _ in range(a.repeats):
    w.TimerStart()
    for _ in range(a.steps):
        w.Step(1 / 60, 4)
    out.append(round(w.TimerStop() / a.steps, 5))
print(json.dumps({"env": {k: v for k, v in os.environ.items() if k.startswith("B2CUDA_")}, "workload": a.workload, "ms_per_step": sorted(out)}))
w.Destroy()
