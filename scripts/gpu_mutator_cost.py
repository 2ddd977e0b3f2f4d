"""Host cost of the between-step mutators (wall clock around the C-ABI calls, stream synchronised at the end)."""
import sys, time
sys.path.insert(0, '/root/repo'); sys.path.insert(0, '/root/repo/tests')
from box2d_b200 import api, scenes, geometry as geo
wd = scenes.world_def()
w = api.World(api.CudaBackend(), wd)
scenes.many_pyramids(w, 10, 10, 10)
for _ in range(10):
    w.Step(1 / 60, 4)
w.Synchronize()
sd = scenes._shape_def()
N = 2000
t0 = time.perf_counter()
ids = []
for k in range(N):
    bd = api.default_body_def(); bd.type = api.DYNAMIC_BODY; bd.position = api.Vec2(-50.0 + 0.05 * k, 40.0 + (k % 7))
    b = w.CreateBody(bd); w.CreateCircleShape(b, sd, geo.Circle((0.0, 0.0), 0.2)); ids.append(b)
w.Synchronize(); t1 = time.perf_counter()
w.Step(1 / 60, 4); w.Synchronize(); t2 = time.perf_counter()
for b in ids:
    w.DestroyBody(b)
w.Synchronize(); t3 = time.perf_counter()
for b in range(1, 1001):
    w.SetTransform(b, (float(b), 80.0), (1.0, 0.0))
w.Synchronize(); t4 = time.perf_counter()
w.Step(1 / 60, 4); w.Synchronize()
c = w.GetCounters()
print(f"create body+shape: {1e6 * (t1 - t0) / N:.1f} us each ({N} bodies past the reserved rows: includes the array growth); step after: {1e3 * (t2 - t1):.2f} ms")
print(f"destroy body: {1e6 * (t3 - t2) / N:.1f} us each; set transform: {1e6 * (t4 - t3) / 1000:.1f} us each; bodies now {c.bodyCount}, errorFlags {c.errorFlags}")
w.Destroy()
