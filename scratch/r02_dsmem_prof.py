"""Round 2: in-kernel phase times of the distributed-shared-memory shape (XRT_PERSIST_PROF=1), GPU box only."""
import os
import sys
os.environ["XRT_PERSIST_PROF"] = "1"
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np  # noqa: E402
from xrtailor_b200 import scene  # noqa: E402

names = ["load/predict", "collide", "apply", "interp+normals", "refit", "generate", "stretch", "bend", "LRA", "vertex solve",
         "edge solve", "apply2", "finalize+writeback"]
for n in (224, 317):
    sc = scene.drop_scene(n)
    for shape in (-3, 16):
        s = scene.make_solver(sc, persist_cluster=shape)
        ms = []
        for f in range(12):
            s.UpdateObstacleAnimationFrame(scene.sphere_motion(sc.obs_pos, f + 1))
            ms.append(s.SimulateSwift())
        prof = s.debug("PERSIST_PROF", np.uint64)[:13].astype(np.float64) / 12 / 1e6
        print(f"n={n} shape {shape}: {np.median(ms[3:]):.3f} ms/frame; in-kernel ms/frame by phase kind (thread 0 of CTA 0):")
        print("   " + ", ".join(f"{a} {b:.3f}" for a, b in zip(names, prof) if b > 0), f"| sum {prof.sum():.3f}")
        st = s.stats()
        print(f"   colours {st['n_stretch_colors']}+{st['n_bend_colors']}, launches/frame {st['launches_per_frame']}")
        s.close()
