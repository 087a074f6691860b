"""ms / frame of ONE garment in each launch shape (C2 and C1m)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import xrtailor_b200 as X
from xrtailor_b200 import scene

shapes = [int(a) for a in sys.argv[1:]] or [0, -2, 16]
sc2, body, bind = scene.body_scene()
sc1 = scene.drop_scene(224)
for name, sc, b, motion in (("C2 30k", sc2, bind, lambda f: body.pose(f + 1)),
                            ("C1m 50k", sc1, None, lambda f: scene.sphere_motion(sc1.obs_pos, f + 1))):
    for sh in shapes:
        s = scene.make_solver(sc, persist_cluster=sh, bindings=b)
        s.SetObstacleSequence(np.stack([motion(f) for f in range(60)]))
        t = [s.SimulateSwift() for _ in range(40)]
        print(f"{name} shape {sh:3d}: {np.median(t[10:]):.3f} ms/frame  ({s.stats()['launches_per_frame']} launches)")
        s.close()
