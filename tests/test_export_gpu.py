"""The output path end to end: frames leave a shared-launch batch through the pinned read-back ring
(xrt_batch_set_readback) while the next frame computes and go straight to the streaming OBJ writers; the files hold
the frames an immediate GetPositions returns, six decimals each."""
import os

import numpy as np
import pytest

import xrtailor_b200 as X
from xrtailor_b200 import scene

pytestmark = pytest.mark.gpu


def parse_obj(path):
    v, vn = [], []
    for ln in open(path):
        if ln.startswith("v "):
            v.append([float(t) for t in ln.split()[1:]])
        elif ln.startswith("vn "):
            vn.append([float(t) for t in ln.split()[1:]])
    return np.array(v), np.array(vn)


def test_batch_frames_stream_to_obj_files(tmp_path):
    sc = scene.drop_scene(24, sphere=(8, 16), y=0.6)
    n, frames = 2, 4
    ss = [scene.make_solver(sc, persist_cluster=1, enable_self_collision=0) for _ in range(n)]
    V = len(sc.cloth_pos)
    b = X.Batch(ss, groups=1)
    b.SetReadback(True)
    ws = [X.ObjSequenceWriter(tmp_path / f"g{i}", sc.cloth_tri, V, with_normals=True) for i in range(n)]
    kept = []
    tg = lambda f: np.concatenate([scene.sphere_motion(sc.obs_pos, f + 1 + 5 * i) for i in range(n)])
    b.UpdateObstacleAnimationFrames(tg(0))
    b.SimulateSwiftAsync()
    for f in range(1, frames + 1):
        if f < frames:
            b.UpdateObstacleAnimationFrames(tg(f))
            b.SimulateSwiftAsync()  # frame f computes while frame f - 1 is read back and written
        x, nr = b.GetPositions(0, normals=True)
        x, nr = x.reshape(n, V, 3), nr.reshape(n, V, 3)
        kept.append((x.copy(), nr.copy()))
        for i, w in enumerate(ws):
            w.push(x[i], nr[i])
    for w in ws:
        assert w.close()[0] == frames
    for i in range(n):
        assert sorted(os.listdir(tmp_path / f"g{i}")) == ["%06d.obj" % f for f in range(frames)]
        for f in range(frames):
            v, vn = parse_obj(tmp_path / f"g{i}" / ("%06d.obj" % f))
            assert np.abs(v - kept[f][0][i]).max() <= 0.5e-6 and np.abs(vn - kept[f][1][i]).max() <= 0.5e-6
    assert np.abs(kept[-1][0][0] - kept[0][0][0]).max() > 1e-4  # the frames differ: the cloth is falling
    b.close()
    for s in ss:
        s.close()
