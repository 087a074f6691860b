# scratch timing script for the first GPU run (not part of the product)
import sys, time
import numpy as np
import xrtailor_b200 as X
from xrtailor_b200 import scene
n = int(sys.argv[1]) if len(sys.argv) > 1 else 224
frames = int(sys.argv[2]) if len(sys.argv) > 2 else 30
moving = int(sys.argv[3]) if len(sys.argv) > 3 else 1
sc = scene.drop_scene(n)
s = scene.make_solver(sc)
ms = []
for f in range(frames):
    if moving:
        s.UpdateObstacleAnimationFrame(scene.sphere_motion(sc.obs_pos, f + 1))
    ms.append(s.SimulateSwift())
ms = np.array(ms[5:])
st = s.stats()
print(f"n={n} V={len(sc.cloth_pos)} median {np.median(ms):.3f} ms/frame p10 {np.percentile(ms,10):.3f} p90 {np.percentile(ms,90):.3f} "
      f"launches {st['launches_per_frame']} colors {st['n_stretch_colors']}/{st['n_bend_colors']} bytes/frame {st['bytes_per_frame']/1e6:.1f} MB "
      f"-> {st['bytes_per_frame']/np.median(ms)/1e6:.1f} GB/s stretch_rms {st['stretch_rms']:.4f}")
