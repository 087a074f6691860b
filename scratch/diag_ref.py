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
# (1) the reference against itself from the same state: its own run-to-run spread
r3 = R.RefSolver(sc, p, seed=7)
c3 = r3.colors()
r3.set_cloth_state(st["x0"], st["x0"], st["v"])
r3.step_frame()
b3 = r3.state(0)["x0"]
out.append("REF vs REF (same colours: %s): rms %.3e rel %.3e max %.3e" % (bool((c3[0] == cs).all() and (c3[1] == cb).all()), rms(b3, b), rel_rms(b3, b), np.sqrt(((b3 - b) ** 2).sum(1)).max()))
# (2) both with OUR proper colouring
gs, _ = X.greedy_color(sc.stretch_idx, len(sc.cloth_pos)); gb, _ = X.greedy_color(sc.bend_idx, len(sc.cloth_pos))
r3.set_colors(gs, gb)
r3.set_cloth_state(st["x0"], st["x0"], st["v"])
r3.step_frame()
s2 = scene.make_solver(sc, colors=(gs, gb), morton_fast_div=1, num_substeps=S, enable_self_collision=selfc)
s2.SetClothState(st["x0"], st["v"]); s2.SimulateSwift()
a2, b2 = s2.GetPositions(0), r3.state(0)["x0"]
d2 = np.sqrt(((a2 - b2) ** 2).sum(1))
out.append("PROPER colouring: rms %.3e rel %.3e max %.3e n>1e-5: %d n>1e-6: %d" % (rms(a2, b2), rel_rms(a2, b2), d2.max(), (d2 > 1e-5).sum(), (d2 > 1e-6).sum()))
bad = np.argsort(-d2)[:6]
rc = r3.contacts(); vf, rt, ee = s2.contacts()
out.append("PROPER worst %s err %s vf ref %s ours %s" % (bad.tolist(), ["%.1e" % e for e in d2[bad]], rc["vf_idx"][bad].tolist(), vf["index"][bad].tolist()))
# run a second reference from the same state
print("\nDIAG " + "\nDIAG ".join(out), flush=True)
