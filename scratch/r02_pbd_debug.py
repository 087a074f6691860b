"""Round 2 debug: where does one PBD repulsion solve differ between the reference's CUDA and the product / oracle?
Isolates cloth-cloth from cloth-obstacle pairs and prints the worst vertices.  Test infrastructure (GPU box only)."""
import os
import sys
import types

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import xrtailor_b200 as X  # noqa: E402
from oracle import oracle as O  # noqa: E402
from oracle.ref_harness import ref as R  # noqa: E402
from xrtailor_b200 import scene  # noqa: E402


def qparams(**kw):
    q = types.SimpleNamespace(xx_stiffness=1.0, yy_stiffness=1.0, xy_stiffness=1.0, xy_poisson_ratio=0.3,
                              yx_poisson_ratio=0.3, bending_stiffness=1e-2, enable_imminent_repulsion=1,
                              imminent_thickness=1e-3, enable_pbd_repulsion=1, pbd_thickness=1e-4, relaxation_rate=0.75,
                              obstacle_mass=1e3, impact_zone_thickness=1e-4)
    for k, v in kw.items():
        setattr(q, k, v)
    return q


def run(tag, n, layers, gap, y, wave, h):
    sc = scene.layered_scene(n, layers, gap=gap, sphere=(12, 24), wave=wave, y=y)
    q = qparams(pbd_thickness=h)
    p = X.default_params()
    p.num_substeps, p.num_iterations, p.enable_self_collision = 1, 20, 0
    p.particle_diameter = sc.particle_diameter()
    cs, _ = X.greedy_color(sc.stretch_idx, len(sc.cloth_pos))
    cb, _ = X.greedy_color(sc.bend_idx, len(sc.cloth_pos))
    cf, _ = X.greedy_color(sc.cloth_tri, len(sc.cloth_pos))
    s = scene.make_solver(sc, colors=(cs, cb), quality=q, strain_colors=cf, num_substeps=1, num_iterations=20,
                          enable_self_collision=0, morton_fast_div=1, persist_cluster=0)
    r = R.RefSolver(sc, p)
    r.setup_quality(sc, q, cf, cb)
    o = O.Scene(sc, p, (cs, cb))
    o.set_quality(sc, q, cf)
    x0 = sc.cloth_pos.copy()
    v = np.zeros_like(x0)
    rng = np.random.default_rng(5)
    x1 = (x0 + rng.uniform(-2e-4, 2e-4, x0.shape)).astype(np.float32)
    s.SetClothState(x0, v); s.SetClothPredicted(x1)
    r.set_cloth_state(x0, x1, v)
    o.set_state(x0, v); o.set_predicted(x1)
    s.QualityBeginFrame()
    o.quality_begin_frame()
    r.imminent_generate()          # builds the reference's cloth tree
    op = o.rep_update_pairs()
    gp = s.RepUpdatePairs()
    rp = r.rep_update_pairs()
    F = len(sc.cloth_tri)
    print(f"[{tag}] pairs product {len(gp)} reference {len(rp)} equal {np.array_equal(gp, rp)}; obstacle pairs {(gp[:, 1] >= F).sum()}")
    print(f"[{tag}] rtri equal: {np.array_equal(s.ClothRTri(F), r.cloth_rtri())}")
    worst_rows = {}
    s.RepSolve()
    # contributions BEFORE the reference moves (its own functions on its own pair list)
    pre = {}
    for i in range(len(x0)):
        pass
    r_contrib = lambda i: r.rep_contributions(i)
    o_contrib = lambda i: o.rep_contributions(i)
    cache = {}
    probe = list(range(len(x0)))
    for i in probe:
        cache[i] = (r_contrib(i), o_contrib(i))
    ed, ec = r.rep_evaluate_only()
    r.rep_solve(float(p.frame_dt))
    g = s.debug("PREDICTED", np.float32).reshape(-1, 3)
    rx = r.state(0)["x"]
    dg, dr = g - x1, rx - x1
    err = np.sqrt(((g - rx) ** 2).sum(1))
    print(f"[{tag}] h={h}: product moved {int((np.abs(dg).max(1) > 0).sum())} rms {np.sqrt((dg ** 2).sum(1).mean()):.3e}; reference moved "
          f"{int((np.abs(dr).max(1) > 0).sum())} rms {np.sqrt((dr ** 2).sum(1).mean()):.3e}; diff rms {np.sqrt((err ** 2).mean()):.3e} max {err.max():.3e}; "
          f"vertices off by > 1e-6: {int((err > 1e-6).sum())}")
    per = len(x0) // layers
    for i in np.argsort(-err)[:8]:
        print(f"   v{i} layer {i // per} grid ({(i % per) // n},{(i % per) % n}) pos {x1[i]} product d {dg[i]} reference d {dr[i]}")
    for i in np.argsort(-err)[:4]:
        (rr, rc), (orow, oc) = cache[i]
        print(f"   v{i}: EvaluateRep_Kernel left sum {ed[i]} count {ec[i]}")
        print(f"   v{i}: reference-function terms {len(rr)}: sum/count*rate = {rc.sum(0) / max(len(rc), 1) * 0.75}; oracle terms {len(orow)}: {oc.sum(0) / max(len(oc), 1) * 0.75}")
        for a, cc in zip(rr, rc):
            print(f"      pair {a[0]} faces {rp[a[0]]} slot {a[1]} nodes {a[2:]} corr {cc}")
    shown = 0
    for i in probe:
        (rr, rc), (orow, oc) = cache[i]
        rk = {(int(a[0]), int(a[1])) for a in rr}
        ok = {(int(a[0]), int(a[1])) for a in orow}
        if rk != ok and shown < 6:
            shown += 1
            print(f"   v{i}: reference terms {len(rr)}, oracle terms {len(orow)}; only oracle {sorted(ok - rk)[:6]}, only reference {sorted(rk - ok)[:6]}")
            for a, cc in zip(orow, oc):
                if (int(a[0]), int(a[1])) in ok - rk:
                    print(f"      oracle-only pair {a[0]} = faces {op[a[0]]} slot {a[1]} nodes {a[2:]} corr {cc}")
            for a, cc in zip(rr, rc):
                if (int(a[0]), int(a[1])) in rk - ok:
                    print(f"      reference-only pair {a[0]} = faces {rp[a[0]]} slot {a[1]} nodes {a[2:]} corr {cc}")
    print(f"[{tag}] probed {len(probe)} vertices: term sets differ on {sum(1 for i in probe if {(int(a[0]), int(a[1])) for a in cache[i][0][0]} != {(int(a[0]), int(a[1])) for a in cache[i][1][0]})}")
    s.close(); r.close()


if __name__ == "__main__":
    run("cloth-cloth", 16, 2, 1e-3, 1.0, 1e-3, 2e-3)
