#!/usr/bin/env python
"""Sub-phase times of the island partition (B2CUDA_DEBUG_CHECKSUMS=3: %globaltimer stamps in dbg[6..12])."""
import os, sys
os.environ["B2CUDA_DEBUG_CHECKSUMS"] = "3"
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from box2d_b200 import api, scenes
wl = sys.argv[1] if len(sys.argv) > 1 else "tumbler"
w = api.World(api.CudaBackend(), scenes.world_def())
scenes.SCENES[wl](w)
for _ in range(120):
    w.Step(1 / 60, 4)
names = ("init+union", "flatten", "count", "classify+scan+layout", "offsets+bodies", "constraints->slots")
acc = [0.0] * 6; n = 0
for _ in range(40):
    w.Step(1 / 60, 4)
    d = w.DebugChecksums()
    if d[12] and d[6]:
        n += 1
        for i in range(6):
            acc[i] += d[7 + i] - d[6 + i]
print(wl, "island partition, us per dirty step:", ", ".join("%s %.1f" % (names[i], acc[i] / max(n, 1) / 1000) for i in range(6)), "(%d dirty steps of 40)" % n)
w.Destroy()
