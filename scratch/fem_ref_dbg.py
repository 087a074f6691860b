import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import xrtailor_b200 as X
from oracle import oracle as O
from oracle.ref_harness import ref as R
from xrtailor_b200 import scene
sc = scene.drop_scene(64, sphere=(24, 48))
p = X.default_params(); p.particle_diameter = sc.particle_diameter()
r = R.RefSolver(sc, p)
pos, tri, bend_idx = sc.cloth_pos, sc.cloth_tri, sc.bend_idx
cs, _ = X.greedy_color(tri, len(pos)); cb, _ = X.greedy_color(bend_idx, len(pos))
rng = np.random.default_rng(5)
x = pos.copy(); x[:, 0] *= 1.08; x[:, 1] += 0.02 * np.sin(7 * x[:, 0]).astype(np.float32); x += rng.normal(0, 5e-4, x.shape).astype(np.float32)
w = sc.inv_mass.astype(np.float32)
for name, m, b, cbb in [("strain only", [5000, 5000, 1000, 0.3, 0.25, 0.5], np.zeros((0, 4), np.uint32), np.zeros(0, np.uint32)),
                        ("strain + bend", [5000, 5000, 1000, 0.3, 0.25, 0.5], bend_idx, cb),
                        ("iso strain", [5000, 5000, 1000, 0.3, 0.3, 0.5], np.zeros((0, 4), np.uint32), np.zeros(0, np.uint32))]:
    m = np.array(m, np.float32)
    for it in (1, 3):
        ref = r.fem_solve(x, m, b, cs, cbb, it, 1 / 60)
        cpu = O.fem_solve(pos, x, w, tri, b, m, cs, cbb, it, 1 / 60)
        dr, dc = ref - x, cpu - x
        print(f"{name} it={it}: |disp| rms ref {np.sqrt((dr**2).sum(1).mean()):.3e} cpu {np.sqrt((dc**2).sum(1).mean()):.3e} "
              f"diff rms {np.sqrt(((dr-dc)**2).sum(1).mean()):.3e} max {np.abs(dr-dc).max():.3e} "
              f"at vertex {np.abs(dr-dc).max(1).argmax()}")
print("--- bending only (strain stiffness ~0), crumpled sheets")
m = np.array([1e-6, 1e-6, 1e-6, 0.3, 0.3, 0.5], np.float32)
for amp in (0.0, 0.002, 0.01, 0.03):
    x = pos.copy()
    x[:, 1] += (amp * np.sin(40 * x[:, 0]) * np.cos(33 * x[:, 2])).astype(np.float32)
    for it in (1, 3):
        ref = r.fem_solve(x, m, bend_idx, cs, cb, it, 1 / 60)
        cpu = O.fem_solve(pos, x, w, tri, bend_idx, m, cs, cb, it, 1 / 60)
        dr, dc = ref - x, cpu - x
        print(f"amp {amp} it={it}: |disp| rms ref {np.sqrt((dr**2).sum(1).mean()):.3e} cpu {np.sqrt((dc**2).sum(1).mean()):.3e} "
              f"diff rms {np.sqrt(((dr-dc)**2).sum(1).mean()):.3e} rel {np.sqrt(((dr-dc)**2).sum()/max((dr**2).sum(),1e-30)):.3e}")
