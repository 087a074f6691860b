"""Round 2: run the reference's Quality-mode pieces (through oracle/ref_harness) on small layered scenes,
compare the CPU oracle with them, and write golden fixtures (tests/golden/quality_*.npz are copies of
gpurun_out/quality_*.npz made by this script on the GPU box).  Test infrastructure."""
import os
import sys
import types

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import xrtailor_b200 as X  # noqa: E402
from oracle import oracle as O  # noqa: E402
from oracle.ref_harness import ref as R  # noqa: E402
from xrtailor_b200 import scene  # noqa: E402

OUT = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "gpurun_out")
os.makedirs(OUT, exist_ok=True)


def qparams(**kw):
    q = types.SimpleNamespace(xx_stiffness=1.0, yy_stiffness=1.0, xy_stiffness=1.0, xy_poisson_ratio=0.3,
                              yx_poisson_ratio=0.3, bending_stiffness=1e-2, enable_imminent_repulsion=1,
                              imminent_thickness=1e-3, enable_pbd_repulsion=1, pbd_thickness=1e-4, relaxation_rate=0.75,
                              obstacle_mass=1e3, impact_zone_thickness=1e-4)
    for k, v in kw.items():
        setattr(q, k, v)
    return q


def setup(n, layers, gap, sphere, N=20, q=None, S=1, wave=0.0):
    sc = scene.layered_scene(n, layers, gap=gap, sphere=sphere, wave=wave)
    p = X.default_params()
    p.num_substeps, p.num_iterations, p.enable_self_collision = S, N, 0
    p.particle_diameter = sc.particle_diameter()
    cs, _ = X.greedy_color(sc.stretch_idx, len(sc.cloth_pos))
    cb, _ = X.greedy_color(sc.bend_idx, len(sc.cloth_pos))
    cf, _ = X.greedy_color(sc.cloth_tri, len(sc.cloth_pos))
    q = q or qparams()
    r = R.RefSolver(sc, p)
    r.setup_quality(sc, q, cf, cb)
    o = O.Scene(sc, p, (cs, cb))
    o.set_quality(sc, q, cf)
    return sc, p, q, r, o, (cs, cb, cf)


def velocities(sc, layers, v_lo=-1.0, v_hi=-2.5):
    V = len(sc.cloth_pos)
    per = V // layers
    v = np.zeros((V, 3), np.float32)
    for k in range(layers):
        v[k * per:(k + 1) * per, 1] = v_lo + (v_hi - v_lo) * k / max(layers - 1, 1)
    return v


def key_rows(nodes, twn):
    """canonical key per impact: kind + node quad (VF: vertex | sorted face; EE: sorted edges, sorted pair)"""
    out = []
    for nd, t in zip(nodes, twn):
        if t[8] > 0.5:
            out.append((1, int(nd[0])) + tuple(sorted(int(x) for x in nd[1:])))
        else:
            a, b = tuple(sorted((int(nd[0]), int(nd[1])))), tuple(sorted((int(nd[2]), int(nd[3]))))
            out.append((0,) + tuple(sorted((a, b))[0]) + tuple(sorted((a, b))[1]))
    return out


def compare_impacts(tag, rn, rt, on, ot):
    rk, ok = key_rows(rn, rt), key_rows(on, ot)
    rs, os_ = set(rk), set(ok)
    both = rs & os_
    print(f"[{tag}] impacts: reference {len(rk)} ({len(rs)} distinct), oracle {len(ok)} ({len(os_)} distinct), "
          f"common {len(both)}, only reference {len(rs - os_)}, only oracle {len(os_ - rs)}")
    # t of common keys: min over duplicates
    def tmin(keys, twn):
        d = {}
        for k, t in zip(keys, twn):
            d[k] = min(d.get(k, 9.0), float(t[0]))
        return d
    a, b = tmin(rk, rt), tmin(ok, ot)
    if both:
        dt = np.array([abs(a[k] - b[k]) for k in both])
        print(f"[{tag}] |t_ref - t_oracle| over common impacts: max {dt.max():.2e}, mean {dt.mean():.2e}")
    same_order = len(rk) == len(ok) and all(x == y for x, y in zip(rk, ok))
    print(f"[{tag}] same rows in the same order: {same_order}")
    return len(both), len(rs | os_)


def main():
    res = {}
    # ---------------- A: FindImpacts / independentImpacts / CollisionStep from one predicted state ----------------
    for tag, n, layers, sphere in (("s16", 16, 2, (8, 16)), ("s32", 32, 3, (12, 24))):
        sc, p, q, r, o, cols = setup(n, layers, 4e-3, sphere)
        dt = np.float32(p.frame_dt / p.num_substeps)
        x0 = sc.cloth_pos.copy()
        v = velocities(sc, layers)
        x1 = (x0 + v * dt).astype(np.float32)
        r.set_cloth_state(x0, x1, v)
        o.set_state(x0, v)
        o.set_predicted(x1)
        rn, rt, rp = r.find_impacts(q.impact_zone_thickness, float(dt))
        on, ot, op = o.find_impacts(q.impact_zone_thickness)
        print(f"[{tag}] V={len(x0)} pairs: reference {rp}, oracle {op}")
        common, union = compare_impacts(tag + " find", rn, rt, on, ot)
        res[tag + "_find"] = (common, union)
        # independent impacts: the reference's own selection, and the oracle's selection on the REFERENCE's list
        rin, rit = r.independent_impacts(0)
        oin, oit = o.independent_impacts(rn, rt, 0)
        compare_impacts(tag + " independent(ref list)", rin, rit, oin, oit)
        np.savez_compressed(os.path.join(OUT, f"quality_impacts_{tag}.npz"), cloth_pos=sc.cloth_pos, cloth_tri=sc.cloth_tri,
                            cloth_edges=sc.cloth_edges, obs_pos=sc.obs_pos, obs_tri=sc.obs_tri, obs_edges=sc.obs_edges,
                            inv_mass=sc.inv_mass, x0=x0, x1=x1, v=v, thickness=q.impact_zone_thickness, dt=dt,
                            ref_nodes=rn, ref_twn=rt, ref_pairs=rp, ref_ind_nodes=rin, ref_ind_twn=rit)
        # CollisionStep on both
        r.collision_step(float(dt))
        o.collision_step()
        rx, ox = r.state(0)["x"], o.predicted()
        d = np.sqrt(((rx - ox) ** 2).sum(1))
        moved = np.sqrt(((rx - x1) ** 2).sum(1))
        print(f"[{tag}] CollisionStep: oracle {o.collision_stats()}, rms(ref, oracle) {np.sqrt((d**2).mean()):.3e} m, "
              f"max {d.max():.3e}; the step moved {int((moved > 1e-7).sum())} vertices, rms {np.sqrt((moved**2).mean()):.3e}")
        # penetration statistics: impacts left after the step, on both (each with its own positions)
        rn2, rt2, _ = r.find_impacts(q.impact_zone_thickness, float(dt))
        on2, ot2, _ = o.find_impacts(q.impact_zone_thickness)
        print(f"[{tag}] impacts left after CollisionStep: reference {len(rn2)}, oracle {len(on2)}")
        np.savez_compressed(os.path.join(OUT, f"quality_step_{tag}.npz"), ref_x=rx, oracle_x=ox, x1=x1)
        r.close()

    # ---------------- B: imminent repulsion + PBD repulsion from one state ----------------
    sc, p, q, r, o, cols = setup(24, 2, 1e-3, (12, 24), q=qparams(pbd_thickness=2e-3), wave=1e-3)
    dt = np.float32(p.frame_dt / p.num_substeps)
    x0 = sc.cloth_pos.copy()
    v = velocities(sc, 2, -0.2, -0.6)
    r.set_cloth_state(x0, x0, v)
    o.set_state(x0, v)
    rqn, rqb = r.imminent_generate()
    o.quality_begin_frame()
    oqn, oqb = o.imminent_generate()
    print(f"[imminent] quadratures: reference {len(rqn)}, oracle {len(oqn)}; same rows "
          f"{rqn.shape == oqn.shape and bool((rqn == oqn).all())}")
    if len(rqn) and rqn.shape == oqn.shape and (rqn == oqn).all():
        print(f"[imminent] max |bary diff| {np.abs(rqb - oqb).max():.2e}")
    else:
        rs, os_ = set(map(tuple, rqn.tolist())), set(map(tuple, oqn.tolist()))
        print(f"[imminent] distinct quads: reference {len(rs)}, oracle {len(os_)}, common {len(rs & os_)}")
    x1 = (x0 + v * dt).astype(np.float32)
    r.set_cloth_state(x0, x1, v)
    o.set_predicted(x1)
    r.imminent_solve(float(dt))
    o.imminent_solve()
    rx, ox = r.state(0)["x"], o.predicted()
    print(f"[imminent] one Solve: moved rms {np.sqrt(((rx - x1) ** 2).sum(1).mean()):.3e} m "
          f"({int((np.abs(rx - x1).max(1) > 0).sum())} vertices), oracle vs reference rms "
          f"{np.sqrt(((rx - ox) ** 2).sum(1).mean()):.3e}, max {np.abs(rx - ox).max():.3e}")
    np.savez_compressed(os.path.join(OUT, "quality_imminent.npz"), cloth_pos=sc.cloth_pos, cloth_tri=sc.cloth_tri,
                        cloth_edges=sc.cloth_edges, obs_pos=sc.obs_pos, obs_tri=sc.obs_tri, obs_edges=sc.obs_edges,
                        x0=x0, x1=x1, v=v, dt=dt, ref_quad_nodes=rqn, ref_quad_bary=rqb, ref_x_after_solve=rx)
    # PBD repulsion from the state after the imminent solve
    x2 = rx.copy()
    r.set_cloth_state(x0, x2, v)
    o.set_predicted(x2)
    rpairs = r.rep_update_pairs()
    opairs = o.rep_update_pairs()
    print(f"[pbd] pairs: reference {len(rpairs)}, oracle {len(opairs)}, equal in order "
          f"{rpairs.shape == opairs.shape and bool((rpairs == opairs).all())}")
    r.rep_solve(float(dt))
    o.rep_solve()
    rx, ox = r.state(0)["x"], o.predicted()
    print(f"[pbd] one SolveNoProximity (h = {q.pbd_thickness}): moved rms {np.sqrt(((rx - x2) ** 2).sum(1).mean()):.3e} m "
          f"({int((np.abs(rx - x2).max(1) > 0).sum())} vertices), oracle vs reference rms "
          f"{np.sqrt(((rx - ox) ** 2).sum(1).mean()):.3e}, max {np.abs(rx - ox).max():.3e}")
    np.savez_compressed(os.path.join(OUT, "quality_pbd.npz"), x0=x0, x2=x2, v=v, ref_pairs=rpairs, ref_x_after_solve=rx,
                        pbd_thickness=q.pbd_thickness)
    r.close()

    # ---------------- C: whole Quality frames ----------------
    for tag, n, layers, sphere, N in (("f16", 16, 2, (8, 16), 20), ("f32", 32, 2, (12, 24), 20)):
        sc, p, q, r, o, cols = setup(n, layers, 4e-3, sphere, N=N)
        x0 = sc.cloth_pos.copy()
        v = velocities(sc, layers)
        r.set_cloth_state(x0, x0, v)
        o.set_state(x0, v)
        errs, ms = [], []
        for f in range(10):
            ms.append(r.step_frame_quality())
            o.step_frame_quality()
            errs.append(float(np.sqrt(((r.state(0)["x0"] - o.positions()) ** 2).sum(1).mean())))
        print(f"[{tag}] 10 Quality frames: rms oracle vs reference per frame {['%.1e' % e for e in errs]}; "
              f"reference ms/frame {np.mean(ms):.1f}; oracle {o.collision_stats()}")
        np.savez_compressed(os.path.join(OUT, f"quality_frames_{tag}.npz"), ref_x=r.state(0)["x0"], oracle_x=o.positions())
        r.close()


if __name__ == "__main__":
    main()
