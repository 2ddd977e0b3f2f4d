"""Where the end-to-end step time goes (host side): Step+Synchronize, + body read-back, + contact events."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from box2d_b200 import api, scenes
w = api.World(api.CudaBackend(), scenes.world_def())
scenes.many_pyramids(w)
for _ in range(40):
    w.Step(1 / 60, 4)
w.Synchronize()
N = 300
def timed(f):
    w.Synchronize(); t = time.perf_counter()
    for _ in range(N): f()
    w.Synchronize(); return (time.perf_counter() - t) / N * 1e6
def a(): w.Step(1 / 60, 4)
def b(): w.Step(1 / 60, 4); w.Synchronize()
def c(): w.Step(1 / 60, 4); w.GetBodyEvents(copy=False)
def d(): w.Step(1 / 60, 4); w.GetBodyEvents(copy=False); w.GetContactEvents()
def e(): w.Step(1 / 60, 4); w.GetContactEvents()
for name, f in (("step back-to-back", a), ("step+sync", b), ("step+bodies", c), ("step+bodies+events", d), ("step+events", e)):
    print(f"{name:22s} {timed(f):8.1f} us/step")
w.Destroy()
