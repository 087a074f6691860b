import faulthandler, sys, os
faulthandler.dump_traceback_later(15, exit=True)
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from xrtailor_b200 import scene
small = scene.drop_scene(64, sphere=(24, 48))
cs = np.zeros(len(small.stretch_rest), np.uint32); cb = np.zeros(len(small.bend_theta), np.uint32)
cl = int(sys.argv[1])
print("make", flush=True)
kw = eval(sys.argv[2]) if len(sys.argv) > 2 else {}
s = scene.make_solver(small, colors=(cs, cb), persist_cluster=cl, particle_diameter=small.particle_diameter(), **kw)
print("made", flush=True)
for f in range(3):
    ms = s.SimulateSwift(); x = s.GetPositions(0)
    print("frame", f, ms, np.isfinite(x).all(), np.abs(x[np.isfinite(x)]).max(), flush=True)
print(s.stats())
