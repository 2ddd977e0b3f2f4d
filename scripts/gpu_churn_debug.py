import sys
sys.path.insert(0, '/root/repo'); sys.path.insert(0, '/root/repo/tests')
import parity, test_gpu_parity as T
from box2d_b200 import api, scenes
import os
for seed in range(int(os.environ.get("CHURN_FIRST", "1")), int(os.environ.get("CHURN_LAST", "12")) + 1):
    STEPS = int(os.environ.get("CHURN_STEPS", "120")); build, mutate = T._churn(seed, STEPS, int(os.environ.get("CHURN_BODIES", "40")))
    calls = []
    class Spy:
        def __init__(self, w): self.w = w
        def __getattr__(self, n):
            f = getattr(self.w, n)
            if not callable(f) or n.startswith("Read") or n.startswith("Get"): return f
            def g(*a, **k):
                r = f(*a, **k)
                if self.w.backend.name != "oracle": calls.append((cur[0], n, a[:2] if n != "CreateJoint" else (a[0].type, a[0].bodyIdA, a[0].bodyIdB), r))
                return r
            return g
    cur = [0]
    def mut(w, step):
        cur[0] = step
        mutate(Spy(w), step)
    try:
        res = parity.run_pair(build, steps=STEPS, mutate=mut, verbose=False)
        print("seed", seed, "ok", res["max_rel_err"])
    except AssertionError as e:
        print("seed", seed, "FAIL", str(e)[:400])
        for c in calls[-60:]: print("   ", c)
