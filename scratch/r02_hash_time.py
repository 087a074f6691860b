"""Time of SpatialHashGPU::Hash (hash + sort + cell ranges + neighbour cache) on one C2 garment, device-wide kernels."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from xrtailor_b200 import scene
sc, body, bind = scene.body_scene()
s = scene.make_solver(sc, bindings=bind, persist_cluster=0)
for f in range(3):
    s.UpdateObstacleAnimationFrame(body.pose(f + 1)); s.SimulateSwift()
s.PredictPositions()
for _ in range(20): s.HashParticles()
s.Sync()
t = time.perf_counter()
n = 300
for _ in range(n): s.HashParticles()
s.Sync()
print(f"hash of {len(sc.cloth_pos)} particles: {(time.perf_counter() - t) / n * 1e6:.1f} us per call (5 launches + sync)")
nb = s.debug("NEIGHBORS", np.uint32)
print("neighbour table checksum", int(nb.astype(np.uint64).sum()), "mean neighbours", float((nb.reshape(-1, len(sc.cloth_pos)) <= len(sc.cloth_pos)).sum(0).mean()))
