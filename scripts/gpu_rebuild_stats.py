"""Debug: rebuild statistics (levels, frontier size, level-loop time) per scene. Needs B2CUDA_DEBUG_CHECKSUMS=1."""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from box2d_b200 import api, scenes
for name in sys.argv[1:]:
    w = api.World(api.CudaBackend(), scenes.world_def())
    scenes.SCENES[name](w)
    rows = []
    for k in range(60):
        w.Step(1.0 / 60.0, 4)
        c = w.DebugChecksums()
        rows.append((c[12], c[13], c[14] / 1000.0, c[15]))
    p = w.GetProfile()
    w.Destroy()
    a = np.array(rows[20:], dtype=float)
    print(name, "levels mean/max %.1f/%d  L mean %.0f  level-loop us mean %.1f  serial ranges %.0f | profile pairs %.3f rebuild %.3f collide %.3f" % (
        a[:, 0].mean(), a[:, 0].max(), a[:, 1].mean(), a[:, 2].mean(), a[:, 3].mean(), p.pairs, p.treeRebuild, p.collide))
