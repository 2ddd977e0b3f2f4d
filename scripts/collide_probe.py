#!/usr/bin/env python
"""Sub-phase times of World.Collide on the device (B2CUDA_DEBUG_CHECKSUMS=2: %globaltimer stamps in dbg[0..5], op count and colouring rounds)."""
import os, sys
os.environ["B2CUDA_DEBUG_CHECKSUMS"] = "2"
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from box2d_b200 import api, scenes
wl = sys.argv[1] if len(sys.argv) > 1 else "tumbler"
w = api.World(api.CudaBackend(), scenes.world_def())
scenes.SCENES[wl](w)
for _ in range(120):
    w.Step(1 / 60, 4)
acc = [0.0] * 5; ops = rounds = 0; n = 40
for _ in range(n):
    w.Step(1 / 60, 4)
    d = w.DebugChecksums()
    if d[5]:
        acc[0] += d[1] - d[0]; acc[1] += d[2] - d[1]; acc[2] += d[4] - d[2]; acc[3] += d[5] - d[4]; acc[4] += d[5] - d[0]
        ops += d[12]; rounds += d[13]
print(wl, "us per step: manifolds %.1f, scan+events %.1f, finish+colouring rounds %.1f, colour array replay %.1f, total %.1f; ops %.0f, rounds %.1f" %
      tuple([a / n / 1000 for a in acc] + [ops / n, rounds / n]))
w.Destroy()
