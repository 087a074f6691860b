"""Round 2: BASELINE.json configs[2] (C3, Quality-mode dress on the walking body) and configs[3] (C4, 200k-vertex
layered tubes: BVH build / refit, pair lists, CCD narrow phase) on the product, with the reference's own CUDA (through
oracle/ref_harness) beside it where it finishes.  Writes gpurun_out/r02_c3_c4.json.  GPU box only."""
import json
import os
import subprocess
import sys
import time
import types

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
OUT = os.path.join(ROOT, "gpurun_out")
os.makedirs(OUT, exist_ok=True)


def qparams(**kw):
    q = types.SimpleNamespace(xx_stiffness=1.0, yy_stiffness=1.0, xy_stiffness=1.0, xy_poisson_ratio=0.3,
                              yx_poisson_ratio=0.3, bending_stiffness=1.0, enable_imminent_repulsion=1,
                              imminent_thickness=1e-3, enable_pbd_repulsion=1, pbd_thickness=1e-4, relaxation_rate=0.25,
                              obstacle_mass=1e3, impact_zone_thickness=1e-4)
    for k, v in kw.items():
        setattr(q, k, v)
    return q


def c4_ours(n, layers):
    import xrtailor_b200 as X
    from xrtailor_b200 import scene
    res = {}
    t = time.perf_counter()
    sc = scene.layered_tubes_scene(n, layers)
    res["scene_numpy_s"] = round(time.perf_counter() - t, 3)
    V, F = len(sc.cloth_pos), len(sc.cloth_tri)
    t = time.perf_counter()
    c = X.build_constraints(sc.cloth_pos, sc.cloth_tri)
    res["build_constraints_s"] = round(time.perf_counter() - t, 3)
    t = time.perf_counter()
    cs, ncs = X.greedy_color(c["edges"], V)
    cb, ncb = X.greedy_color(c["bend_idx"], V)
    cf, ncf = X.greedy_color(sc.cloth_tri, V)
    res["greedy_colouring_s"] = round(time.perf_counter() - t, 3)
    res["colours"] = [int(ncs), int(ncb), int(ncf)]
    res.update(verts=V, faces=F, edges=len(c["edges"]), bend=len(c["bend_idx"]))
    t = time.perf_counter()
    s = scene.make_solver(sc, colors=(cs, cb), quality=qparams(max_pairs=min(400 * F, 2 ** 28 - 1), max_candidates=min(4000 * F, 2 ** 31), max_impacts=50 * F), strain_colors=cf, num_substeps=1, num_iterations=20,
                          enable_self_collision=0, persist_cluster=0)
    res["finalize_setup_s"] = round(time.perf_counter() - t, 3)
    # falling state: layer k moves a little faster than layer k - 1, so the layers close their 2 mm gaps inside the frame
    x0 = sc.cloth_pos.copy()
    v = np.zeros_like(x0)
    per = V // layers
    for k in range(layers):
        v[k * per:(k + 1) * per, 1] = -0.2
        rad = x0[k * per:(k + 1) * per][:, [0, 2]]
        v[k * per:(k + 1) * per, 0] = -0.12 * k * rad[:, 0] / np.linalg.norm(rad, axis=1)
        v[k * per:(k + 1) * per, 2] = -0.12 * k * rad[:, 1] / np.linalg.norm(rad, axis=1)
    dt = np.float32(1.0 / 60.0)
    x1 = (x0 + v * dt).astype(np.float32)
    s.SetClothState(x0, v)
    s.SetClothPredicted(x1)
    s.QualityBeginFrame()
    names = ["tree_build_ms", "tree_update_ms", "ccd_tree_build_ms", "pairs_ms", "ccd_narrow_ms", "find_impacts_ms"]
    for what, name in enumerate(names):
        ms, cnt = s.QualityProfile(what, reps=5)
        res[name] = round(ms, 4)
        if what == 3:
            res["pairs"] = int(cnt)
            res["pairs_per_s"] = round(cnt / (ms * 1e-3))
        if what == 4:
            res["ccd_candidates"] = int(cnt)
            res["ccd_tests_per_s"] = round(15 * res["pairs"] / (ms * 1e-3))
            res["ccd_cubic_solves_per_s"] = round(cnt / (ms * 1e-3))
        if what == 5:
            res["impacts"] = int(cnt)
    res["build_faces_per_s"] = round(F / (res["tree_build_ms"] * 1e-3))
    # Swift frames on the same mesh
    s.close()
    s2 = scene.make_solver(sc, colors=(cs, cb), persist_cluster=0)
    ms = [s2.SimulateSwift() for _ in range(8)]
    res["swift_ms_per_frame_per_stage_graph"] = round(float(np.median(ms[3:])), 3)
    s2.close()
    return res, (x0, x1, v)


def c4_reference(n, layers, state_file, timeout=1500):
    """in a child process with a time limit: the reference's setup at 200k builds O(n^2) adjacency tables"""
    code = f"""
import sys, time, json, types, numpy as np
sys.path.insert(0, {ROOT!r})
from workloads import scene
from oracle.ref_harness import ref as R
sc = scene.layered_tubes_scene({n}, {layers})
p = scene.swift_params(num_substeps=1, num_iterations=20, enable_self_collision=0)
p.particle_diameter = sc.particle_diameter()
st = np.load({state_file!r})
t = time.perf_counter()
r = R.RefSolver(sc, p)
setup = time.perf_counter() - t
r.set_cloth_state(st['x0'], st['x1'], st['v'])
res = r.quality_timing(1e-4, 1.0 / 60.0, reps=3)
res['setup_s'] = round(setup, 2)
ms = [r.step_frame() for _ in range(6)]
res['swift_ms_per_frame'] = float(np.median(ms[2:]))
print('JSON' + json.dumps(res))
"""
    try:
        r = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, timeout=timeout)
        line = [ln for ln in r.stdout.splitlines() if ln.startswith("JSON")]
        if not line:
            return {"error": (r.stderr or r.stdout)[-300:]}
        return json.loads(line[-1][4:])
    except subprocess.TimeoutExpired:
        return {"error": f"did not finish in {timeout} s"}


def c3(scale, frames, N, with_ref=True):
    import xrtailor_b200 as X
    from xrtailor_b200 import scene
    sc, body, bind = scene.dress_scene(scale)
    V = len(sc.cloth_pos)
    q = qparams()
    cs, _ = X.greedy_color(sc.stretch_idx, V)
    cb, _ = X.greedy_color(sc.bend_idx, V)
    cf, _ = X.greedy_color(sc.cloth_tri, V)
    s = scene.make_solver(sc, colors=(cs, cb), bindings=bind, quality=q, strain_colors=cf, num_substeps=1, num_iterations=N,
                          enable_self_collision=0, persist_cluster=0)
    ms, stats = [], []
    for f in range(frames):
        s.UpdateObstacleAnimationFrame(body.pose(f + 1, amp_deg=25.0, hz=1.2))
        ms.append(s.SimulateQuality())
    g = s.GetPositions(0)
    res = {"verts": V, "N": N, "frames": frames, "ms_per_frame": [round(m, 2) for m in ms], "finite": bool(np.isfinite(g).all()),
           "launches_per_frame": int(s.stats()["launches_per_frame"])}
    s.close()
    if with_ref:
        code = f"""
import sys, time, json, types, numpy as np
sys.path.insert(0, {ROOT!r})
from workloads import scene
from oracle.ref_harness import ref as R
import xrtailor_b200 as X
sc, body, bind = scene.dress_scene({scale})
V = len(sc.cloth_pos)
p = scene.swift_params(num_substeps=1, num_iterations={N}, enable_self_collision=0)
p.particle_diameter = sc.particle_diameter()
q = types.SimpleNamespace(**{vars(q)!r})
cs, _ = X.greedy_color(sc.stretch_idx, V); cb, _ = X.greedy_color(sc.bend_idx, V); cf, _ = X.greedy_color(sc.cloth_tri, V)
r = R.RefSolver(sc, p)
r.set_bindings(*bind)
r.setup_quality(sc, q, cf, cb)
ms = []
for f in range({frames}):
    r.set_obstacle_target(body.pose(f + 1, amp_deg=25.0, hz=1.2))
    ms.append(r.step_frame_quality())
print('JSON' + json.dumps([round(m, 2) for m in ms]))
"""
        try:
            r = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, timeout=900)
            line = [ln for ln in r.stdout.splitlines() if ln.startswith("JSON")]
            res["reference_ms_per_frame"] = json.loads(line[-1][4:]) if line else {"error": (r.stderr or r.stdout)[-300:]}
        except subprocess.TimeoutExpired:
            res["reference_ms_per_frame"] = {"error": "did not finish in 900 s"}
        a = np.array(res["ms_per_frame"][1:])
        if isinstance(res["reference_ms_per_frame"], list):
            b = np.array(res["reference_ms_per_frame"][1:])
            res["speedup"] = round(float(b.mean() / a.mean()), 2)
    return res


if __name__ == "__main__":
    out = {}
    which = sys.argv[1:] or ["c4small", "c4", "c3small", "c3"]
    if "c4small" in which:
        r, st = c4_ours(64, 4)
        f = os.path.join(OUT, "c4small_state.npz")
        np.savez(f, x0=st[0], x1=st[1], v=st[2])
        r["reference"] = c4_reference(64, 4, f, timeout=300)
        out["C4_small_4x64x64"] = r
        print(json.dumps({"C4_small_4x64x64": r}), flush=True)
    if "c4" in which:
        r, st = c4_ours(224, 4)
        f = os.path.join(OUT, "c4_state.npz")
        np.savez(f, x0=st[0], x1=st[1], v=st[2])
        r["reference"] = c4_reference(224, 4, f, timeout=1200)
        os.remove(f)
        out["C4_4x224x224"] = r
        print(json.dumps({"C4_4x224x224": r}), flush=True)
    if "c3small" in which:
        out["C3_dress_scale0.05_N50"] = c3(0.05, 6, 50)
        print(json.dumps({"C3small": out["C3_dress_scale0.05_N50"]}), flush=True)
    if "c3" in which:
        out["C3_dress_50k_N200"] = c3(1.0, 4, 200, with_ref="noref" not in which)
        print(json.dumps({"C3": out["C3_dress_50k_N200"]}), flush=True)
    json.dump(out, open(os.path.join(OUT, "r02_c3_c4.json"), "w"), indent=1)
