"""C2 scene sanity on the CPU oracle: finite, garments stay outside the body."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import xrtailor_b200 as X
from oracle import oracle as O
from xrtailor_b200 import scene

def seg_dist(p, a, b):
    ab = b - a
    t = np.clip(((p - a) @ ab) / (ab @ ab), 0, 1)
    return np.linalg.norm(p - (a + t[:, None] * ab), axis=1)

def penetration(body, pose, x):
    """max depth of a cloth vertex inside any capsule (m)."""
    worst = 0.0
    for (name, r, h, c, joint, *_), (_, lo, hi, _) in zip(scene._BODY_PARTS, body.parts):
        top, bot = pose[lo], pose[hi - 1]            # poles
        ax = (top - bot) / np.linalg.norm(top - bot)
        d = seg_dist(x, top - ax * r, bot + ax * r)
        worst = max(worst, float((r - d).max()))
    return worst

if __name__ == "__main__":
    scale = float(sys.argv[1]) if len(sys.argv) > 1 else 0.05
    frames = int(sys.argv[2]) if len(sys.argv) > 2 else 30
    amp = float(sys.argv[3]) if len(sys.argv) > 3 else 30.0
    sc, body, bind = scene.body_scene(scale)
    p = X.default_params(); p.particle_diameter = sc.particle_diameter()
    cs, _ = X.greedy_color(sc.stretch_idx, len(sc.cloth_pos)); cb, _ = X.greedy_color(sc.bend_idx, len(sc.cloth_pos))
    o = O.Scene(sc, p, (cs, cb))
    o.set_binding(*bind)
    t0 = time.time()
    for f in range(frames):
        pose = body.pose(f + 1, amp_deg=amp)
        o.set_obstacle_target(pose); o.step_frame()
        x = o.positions()
        if f % 5 == 4 or f == frames - 1:
            print(f"frame {f+1}: finite {np.isfinite(x).all()}  max penetration {penetration(body, pose, x)*1e3:.2f} mm  "
                  f"y range {x[:,1].min():.3f}..{x[:,1].max():.3f}")
    d = np.linalg.norm(x[sc.stretch_idx[:, 0]] - x[sc.stretch_idx[:, 1]], axis=1) / sc.stretch_rest
    print("stretch ratio max %.3f mean %.4f" % (d.max(), d.mean()))
    print("verts", len(x), "s/frame", (time.time() - t0) / frames)
