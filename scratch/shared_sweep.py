"""C2: shared-launch batch throughput vs cluster size (one garment per resident cluster)."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import xrtailor_b200 as X
from xrtailor_b200 import scene
sc, body, bindings = scene.body_scene()
seq = np.stack([body.pose(f + 1, amp_deg=30.0) for f in range(60)])
for arg in sys.argv[1:] or ["1", "2", "4", "8"]:
    parts = arg.split(":")
    cl = int(parts[0])
    G = int(parts[1]) if len(parts) > 1 else 1
    M = int(parts[2]) if len(parts) > 2 and parts[2] else X.persist_info(cl)[1]
    W = int(parts[3]) if len(parts) > 3 else max(0, 32 // G - 1)
    t0 = time.time()
    solvers = []
    for i in range(M):
        s = scene.make_solver(sc, persist_cluster=cl, bindings=bindings)
        s.SetObstacleSequence(seq)
        solvers.append(s)
    t1 = time.time()
    b = X.Batch(solvers, groups=G, workers=W)
    b.run(3)
    frames = 10
    ms = b.run(frames)
    kms, kb = b.profile(4)
    print(f"cluster {cl} x {M} garments in {G} groups x {W} workers: {ms / frames:.2f} ms per round = {M * frames / ms * 1e3:.0f} frames/s; "
          f"substep launch {kms * 1e3:.0f} us ({kb / kms / 1e6:.0f} GB/s algorithmic); setup {t1 - t0:.1f} s", flush=True)
    x = solvers[-1].GetPositions(0)
    assert np.isfinite(x).all()
    b.close()
    for s in solvers:
        s.close()
