import faulthandler, sys, os
faulthandler.dump_traceback_later(15, exit=True)
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from xrtailor_b200 import scene
small = scene.drop_scene(64, sphere=(24, 48))
s = scene.make_solver(small, persist_cluster=0)
rng = np.random.default_rng(0)
for scale in (1e3, 1e6, 1e8, 1e12, 1e30):
    x = (small.cloth_pos * scale).astype(np.float32)
    s.SetClothPredicted(x)
    print("hash at scale", scale, flush=True)
    s.HashParticles()
    print("  ok", flush=True)
x[::7] = np.inf; x[3::11] = np.nan
s.SetClothPredicted(x); print("hash with inf/nan", flush=True); s.HashParticles(); print("  ok", flush=True)
