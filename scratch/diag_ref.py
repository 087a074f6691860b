import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import xrtailor_b200 as X
from oracle.ref_harness import ref as R
from xrtailor_b200 import scene
from tests.common import rel_rms, rms
from tests.test_ref_parity_gpu import _improper
S = int(sys.argv[1]) if len(sys.argv) > 1 else 1
selfc = int(sys.argv[2]) if len(sys.argv) > 2 else 1
settle = int(sys.argv[3]) if len(sys.argv) > 3 else 30
sc = scene.drop_scene(64, sphere=(24, 48))
p = X.default_params(); p.particle_diameter = sc.particle_diameter(); p.num_substeps = S; p.enable_self_collision = selfc
r2 = R.RefSolver(sc, p, seed=7)
cs, cb, nsc, nbc = r2.colors()
out = ["S=%d self=%d settle=%d colours %d/%d conflicts %d/%d" % (S, selfc, settle, nsc, nbc, _improper(sc.stretch_idx, cs), _improper(sc.bend_idx, cb))]
for _ in range(settle): r2.step_frame()
st = r2.state(0)
s = scene.make_solver(sc, colors=(cs, cb), morton_fast_div=1, num_substeps=S, enable_self_collision=selfc)
s.SetClothState(st["x0"], st["v"])
s.SimulateSwift(); r2.step_frame()
a, b = s.GetPositions(0), r2.state(0)["x0"]
rc = r2.contacts(); vf, rt, ee = s.contacts()
d = np.sqrt(((a - b) ** 2).sum(1))
disp = np.sqrt(((b - st["x0"]) ** 2).sum(1).mean())
out.append("rms %.3e rel %.3e disp %.3e max %.3e n>1e-5: %d n>1e-6: %d" % (rms(a, b), rel_rms(a, b), disp, d.max(), (d > 1e-5).sum(), (d > 1e-6).sum()))
out.append("vf contacts ref %d ours %d same idx %.4f; rt ref %d ours %d same %.4f; ee live ref %d ours %d" % ((rc["vf_idx"] >= 0).sum(), (vf["index"] >= 0).sum(), (vf["index"] == rc["vf_idx"]).mean(), (rc["rt_idx"] >= 0).sum(), (rt["index"] >= 0).sum(), (rt["index"] == rc["rt_idx"]).mean(), (rc["ee"][:, 0] < 1e6).sum(), (ee[:, 0] < 1e6).sum()))
bad = np.argsort(-d)[:8]
out.append("worst verts %s err %s vf ref %s ours %s" % (bad.tolist(), ["%.1e" % e for e in d[bad]], rc["vf_idx"][bad].tolist(), vf["index"][bad].tolist()))
# is the reference deterministic? run a second reference from the same state
print("\nDIAG " + "\nDIAG ".join(out), flush=True)
