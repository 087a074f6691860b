"""Run-to-run spread of ten Quality frames (two 32 x 32 sheets on the sphere): the reference's own CUDA against itself,
the product against itself, and product against reference, each run in a fresh process.  Both implementations
accumulate with float atomics, so none of them is reproducible bit for bit; this says how far apart two runs of the
SAME code end up, which is the floor of any cross-implementation gate."""
import os, subprocess, sys, json
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np

CHILD = r'''
import sys, numpy as np
sys.path.insert(0, %r); sys.path.insert(0, %r + "/tests")
import test_quality_gpu as T
n, frames, out = int(sys.argv[1]), int(sys.argv[2]), sys.argv[3]
c = T.Case(n, 2, 4e-3, (8, 16) if n == 16 else (12, 24), N=20, fast_div=0)
x0 = c.sc.cloth_pos.copy(); v = c.velocities()
c.s.SetClothState(x0, v); c.r.set_cloth_state(x0, x0, v)
P, Rr = [], []
for f in range(frames):
    c.s.SimulateQuality(); P.append(c.s.GetPositions(0).copy())
    c.r.step_frame_quality(); Rr.append(c.r.state(0)["x0"].copy())
np.savez(out, P=np.array(P), R=np.array(Rr))
''' % (ROOT, ROOT)

def rms(a, b):
    return float(np.sqrt(((a.astype(np.float64) - b) ** 2).sum(-1).mean()))

res = {}
for n, frames in ((16, 5), (32, 10)):
    runs = []
    for k in range(4):
        out = f"/tmp/qspread_{n}_{k}.npz"
        subprocess.run([sys.executable, "-c", CHILD, str(n), str(frames), out], check=True, capture_output=True)
        runs.append(np.load(out))
    pairs = [(a, b) for a in range(4) for b in range(a + 1, 4)]
    last = frames - 1
    res[f"n{n}"] = {
        "reference_vs_reference_last_frame": [round(rms(runs[a]["R"][last], runs[b]["R"][last]), 6) for a, b in pairs],
        "product_vs_product_last_frame": [round(rms(runs[a]["P"][last], runs[b]["P"][last]), 6) for a, b in pairs],
        "product_vs_reference_last_frame": [round(rms(runs[a]["P"][last], runs[b]["R"][last]), 6) for a in range(4) for b in range(4)],
        "product_vs_reference_max_over_frames": [round(max(rms(runs[a]["P"][f], runs[a]["R"][f]) for f in range(frames)), 6) for a in range(4)],
    }
print(json.dumps(res, indent=1))
json.dump(res, open(os.path.join(ROOT, "gpurun_out", "r02_quality_spread.json"), "w"), indent=1)
