#!/usr/bin/env python
"""Sub-phase times of UpdateBroadPhasePairs on the device (B2CUDA_DEBUG_CHECKSUMS=2: %globaltimer stamps in dbg[6..10])."""
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
names = ("pair set rebuild", "tree queries (count pass)", "wide queries", "scan + fill")
acc = [0.0] * 4; n = 0; big = 0; moves = 0; pairs = 0
for _ in range(40):
    w.Step(1 / 60, 4)
    d = w.DebugChecksums()
    c = w.GetCounters()
    if d[10] and d[6]:
        n += 1; big += d[11]; moves += c.moveCount; pairs += c.newPairCount
        for i in range(4):
            acc[i] += d[7 + i] - d[6 + i]
print(wl, "pairs, us per step:", ", ".join("%s %.1f" % (names[i], acc[i] / max(n, 1) / 1000) for i in range(4)), "| wide queries %.1f, moves %.0f, new pairs %.0f (%d steps)" % (big / max(n, 1), moves / max(n, 1), pairs / max(n, 1), n))
w.Destroy()
