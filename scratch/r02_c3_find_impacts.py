"""C3 dress (50k vertices, Quality): stage times of FindImpacts on the state after three frames (QualityProfile)."""
import os, sys, types
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import xrtailor_b200 as X
from xrtailor_b200 import scene
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
from r02_c3_c4 import qparams
sc, body, bind = scene.dress_scene(1.0)
V = len(sc.cloth_pos)
cs, _ = X.greedy_color(sc.stretch_idx, V); cb, _ = X.greedy_color(sc.bend_idx, V); cf, _ = X.greedy_color(sc.cloth_tri, V)
s = scene.make_solver(sc, colors=(cs, cb), bindings=bind, quality=qparams(), strain_colors=cf, num_substeps=1, num_iterations=200,
                      enable_self_collision=0, persist_cluster=0)
for f in range(3):
    s.UpdateObstacleAnimationFrame(body.pose(f + 1, amp_deg=25.0, hz=1.2))
    print("frame", f, round(s.SimulateQuality(), 1), "ms")
names = ["tree_build", "tree_update", "ccd_tree_build", "pairs", "ccd_narrow", "find_impacts"]
for what, name in enumerate(names):
    ms, cnt = s.QualityProfile(what, reps=5)
    print(f"{name}: {ms:.3f} ms, count {int(cnt)}")
