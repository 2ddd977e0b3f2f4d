import os, sys
import numpy as np
sys.path.insert(0, '/root/repo')
from box2d_b200 import api, scenes
w = api.World(api.CudaBackend(), scenes.world_def()); scenes.tumbler(w)
rows=[]
for k in range(80):
    w.Step(1/60, 4); c = w.DebugChecksums(); cn = w.GetCounters()
    rows.append([c[7]-c[6], c[8]-c[7], c[9]-c[8], c[10]-c[9], cn.moveCount, cn.newPairCount, c[11]])
a=np.array(rows[30:],dtype=float)
print("tumbler us: clear+insert %.1f  count %.1f  bigqueries %.1f  scan+fill %.1f | moves %.0f newpairs %.0f bigqueries %.1f max %d" % tuple(list(a[:,:4].mean(0)/1000)+[a[:,4].mean(), a[:,5].mean(), a[:,6].mean(), a[:,6].max()]))
