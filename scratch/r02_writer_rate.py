"""Throughput of the streaming OBJ writer on this host's cores (C2 garment, positions + normals, 2.97 MB per file)."""
import os, sys, tempfile, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import xrtailor_b200 as X
from workloads import scene
sc, body, bind = scene.body_scene()
V, tri = len(sc.cloth_pos), sc.cloth_tri
x = sc.cloth_pos.astype(np.float32)
n = np.random.default_rng(0).standard_normal((V, 3)).astype(np.float32)
print("host cores:", len(os.sched_getaffinity(0)))
for th in (1, 2, 4, 8, 16):
    d = tempfile.mkdtemp(dir="/dev/shm")
    w = X.ObjSequenceWriter(d, tri, V, with_normals=True, queue_frames=32, threads=th)
    t = time.perf_counter()
    for f in range(128):
        w.push(x, n)
    fr, by = w.close()
    dt = time.perf_counter() - t
    print(f"threads {th}: {fr / dt:.0f} files/s, {by / dt / 1e6:.0f} MB/s")
    for f in os.listdir(d):
        os.remove(os.path.join(d, f))
    os.rmdir(d)
