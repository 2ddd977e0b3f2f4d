"""Host-side cost of one Step() call: a world so small that the GPU is never the limiter (wall time per step = launch path)."""
import sys, time
sys.path.insert(0, '.')
from box2d_b200 import api, scenes
for name, build in (("pyramid3", lambda w: scenes.add_pyramid(w, 3)), ("joint_grid4", lambda w: scenes.joint_grid(w, 4))):
    w = api.World(api.CudaBackend(), scenes.world_def())
    build(w)
    for _ in range(50):
        w.Step(1 / 60, 4)
    w.Synchronize()
    n = 100
    t0 = time.perf_counter()
    for _ in range(n):
        w.Step(1 / 60, 4)
    t1 = time.perf_counter()
    w.Synchronize()
    t2 = time.perf_counter()
    print(name, "host us/step (enqueue only)", round((t1 - t0) / n * 1e6, 2), "incl. drain", round((t2 - t0) / n * 1e6, 2))
    w.Destroy()
