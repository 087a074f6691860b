"""Per-phase wall time of the persistent substep kernel (XRT_PERSIST_PROF=1)."""
import os, sys
os.environ["XRT_PERSIST_PROF"] = "1"
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import xrtailor_b200 as X
from xrtailor_b200 import scene

NAMES = ["predict", "collide", "apply1", "interp+normals", "refit", "generate", "stretch", "bend", "lra",
         "solve_vertex", "solve_edge", "apply2", "finalize"]
n = int(sys.argv[1]) if len(sys.argv) > 1 else 224
for cl in [int(a) for a in sys.argv[2:]] or [8, 16]:
    sc = scene.drop_scene(n)
    s = scene.make_solver(sc, persist_cluster=cl)
    frames = 20
    for f in range(5):
        s.UpdateObstacleAnimationFrame(scene.sphere_motion(sc.obs_pos, f + 1)); s.SimulateSwift()
    p0 = s.debug("PERSIST_PROF", np.uint64).copy()
    ms = []
    for f in range(5, 5 + frames):
        s.UpdateObstacleAnimationFrame(scene.sphere_motion(sc.obs_pos, f + 1)); ms.append(s.SimulateSwift())
    p = (s.debug("PERSIST_PROF", np.uint64) - p0).astype(np.float64) / frames / 1e3
    print(f"cluster {cl}: frame {np.median(ms):.3f} ms; per-phase-kind us/frame (sum {p[:13].sum():.0f}):")
    for nm, v in zip(NAMES, p[:13]):
        print(f"   {nm:16s} {v:9.1f}")
    s.close()
