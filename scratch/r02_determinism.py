"""Are two identically driven handles bit-equal on the small body scene? (EE contact solves use float atomics.)"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from xrtailor_b200 import scene
sc, body, bind = scene.body_scene(0.05)
for cluster in (0, 1):
    outs = []
    for rep in range(3):
        s = scene.make_solver(sc, bindings=bind, persist_cluster=cluster)
        for f in range(1, 5):
            s.UpdateObstacleAnimationFrame(body.pose(f))
            s.SimulateSwift()
        outs.append(s.GetPositions(0))
        s.close()
    print("cluster", cluster, "max diff between identical runs:", [float(np.abs(outs[0] - o).max()) for o in outs[1:]])
