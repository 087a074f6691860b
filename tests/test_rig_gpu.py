"""Body-animation producer on the device (include/xrt_b200.h: xrt_set_obstacle_rig / _pose /
_pose_sequence) through the C ABI against the CPU oracle (oracle/xrt_oracle_rig.c).  The library's default build is
IEEE without FMA contraction and accumulates in the reference kernel's order, so the skinned targets are compared bit
for bit; the joint transforms are host arithmetic on both sides (gate 1e-6, measured 0)."""
import numpy as np
import pytest

import xrtailor_b200 as X
from oracle import oracle as O
from xrtailor_b200 import scene

pytestmark = pytest.mark.gpu


def rms(a, b):
    return float(np.sqrt(((np.asarray(a, np.float64) - b) ** 2).sum(-1).mean()))


def small_body_scene():
    sc, body, bind = scene.body_scene(0.05)
    return sc, body, bind, scene.body_rig(body)


def set_rig(s, rig, **kw):
    s.SetObstacleRig(rig["rest"], rig["w_outer"], rig["w_inner"], rig["w_val"], rig["joints"], rig["parent"], **kw)


def oracle_target(rig, pose, trans=None, M=None, B=None):
    T = O.rig_joint_transforms(pose, rig["joints"], rig["parent"], trans)
    par = O.rig_pose_blend_params(pose) if B is not None else None
    return T, O.rig_lbs(rig["rest"], rig["w_outer"], rig["w_inner"], rig["w_val"], T, model_matrix=M, pose_blend=B,
                        blend_params=par)


def test_skinned_target_equals_the_oracle_bit_for_bit():
    sc, body, bind, rig = small_body_scene()
    s = scene.make_solver(sc, bindings=bind, persist_cluster=0)
    rng = np.random.default_rng(3)
    set_rig(s, rig)
    for f in (1, 9):
        pose = scene.body_pose_params(body, f) + rng.uniform(-0.2, 0.2, (len(rig["joints"]), 3)).astype(np.float32)
        trans = rng.uniform(-0.1, 0.1, 3).astype(np.float32)
        s.SetObstaclePose(pose, trans)
        T, want = oracle_target(rig, pose, trans)
        assert np.abs(s.ObstacleJointTransforms() - T).max() <= 1e-6
        np.testing.assert_array_equal(s.ObstacleTarget(), want)
    # model matrix (UpdateObstaclePredicted's model * align) and pose blend shapes
    ang = np.float32(0.3)
    M = np.array([[1, 0, 0, 0.05], [0, np.cos(ang), -np.sin(ang), 0.0], [0, np.sin(ang), np.cos(ang), -0.02], [0, 0, 0, 1]], np.float32)
    B = (rng.standard_normal((3 * len(rig["rest"]), 9 * (len(rig["joints"]) - 1))) * 1e-3).astype(np.float32)
    set_rig(s, rig, model_matrix=M, pose_blend=B)
    pose = scene.body_pose_params(body, 5)
    s.SetObstaclePose(pose)
    _, want = oracle_target(rig, pose, M=M, B=B)
    np.testing.assert_array_equal(s.ObstacleTarget(), want)
    s.close()


def test_one_hot_rig_reproduces_the_generator_pose():
    sc, body, bind, _ = small_body_scene()
    rig = scene.body_rig(body, blend=0.0)
    s = scene.make_solver(sc, bindings=bind, persist_cluster=0)
    set_rig(s, rig)
    for f in (1, 14):
        s.SetObstaclePose(scene.body_pose_params(body, f))
        assert np.abs(s.ObstacleTarget() - body.pose(f)).max() <= 2e-6
    s.close()


@pytest.mark.parametrize("cluster", [0, 1])
def test_frames_driven_by_poses_equal_frames_driven_by_uploaded_targets(cluster):
    """A frame whose obstacle target was skinned on the device is the frame the reference's path gives: posed vertices
    computed off-line (here: the oracle) and uploaded with xrt_set_obstacle_target."""
    sc, body, bind, rig = small_body_scene()
    a = scene.make_solver(sc, bindings=bind, persist_cluster=cluster)
    b = scene.make_solver(sc, bindings=bind, persist_cluster=cluster)
    set_rig(a, rig)
    for f in range(1, 5):
        pose = scene.body_pose_params(body, f)
        a.SetObstaclePose(pose)
        b.UpdateObstacleAnimationFrame(oracle_target(rig, pose)[1])
        a.SimulateSwift()
        b.SimulateSwift()
    # the obstacle is bit-equal; the cloth is in contact with it and the edge-contact solve accumulates with float
    # atomics (like the reference's), so two identical runs differ by ~1e-3 m after four frames (scratch/r02_determinism.py)
    np.testing.assert_array_equal(a.GetPositions(1), b.GetPositions(1))
    assert rms(a.GetPositions(0), b.GetPositions(0)) <= 1e-3
    a.close()
    b.close()


def test_resident_pose_sequence_in_a_shared_launch_batch():
    """xrt_set_obstacle_pose_sequence: no per-frame upload; single handles and a shared-launch batch take the same frames."""
    sc, body, bind, rig = small_body_scene()
    n = 3
    seqs = [np.stack([scene.body_pose_params(body, f + 1, phase=0.7 * i) for f in range(4)]) for i in range(n)]

    def run(batched):
        ss = [scene.make_solver(sc, bindings=bind, persist_cluster=1) for _ in range(n)]
        for s, q in zip(ss, seqs):
            set_rig(s, rig)
            s.SetObstaclePoseSequence(q)
        if batched:
            bt = X.Batch(ss, groups=2)
            bt.run(5)  # wraps after four frames
            bt.close()
        else:
            for _ in range(5):
                for s in ss:
                    s.SimulateSwift()
        out = [(s.GetPositions(0), s.GetPositions(1)) for s in ss]
        for s in ss:
            s.close()
        return out

    single, batch = run(False), run(True)
    for (x, ox), (y, oy) in zip(single, batch):
        np.testing.assert_array_equal(ox, oy)  # the same frames were taken, wrap included
        assert rms(x, y) <= 5e-3  # five frames in contact: run-to-run divergence of the atomics, see above
    # and the sequence is what explicit poses give
    s = scene.make_solver(sc, bindings=bind, persist_cluster=1)
    set_rig(s, rig)
    for f in range(5):
        s.SetObstaclePose(seqs[0][f % 4])
        s.SimulateSwift()
    np.testing.assert_array_equal(s.GetPositions(1), single[0][1])
    assert rms(s.GetPositions(0), single[0][0]) <= 5e-3
    s.close()


def test_rig_argument_checks():
    sc, body, bind, rig = small_body_scene()
    s = scene.make_solver(sc, bindings=bind, persist_cluster=0)
    with pytest.raises(X.XrtError):
        s._n_joints = len(rig["joints"])
        s.SetObstaclePose(np.zeros((len(rig["joints"]), 3), np.float32))  # no rig yet
    bad = dict(rig, w_inner=rig["w_inner"].copy())
    bad["w_inner"][5] = 99
    with pytest.raises(X.XrtError):
        set_rig(s, bad)
    with pytest.raises(X.XrtError):
        set_rig(s, dict(rig, rest=rig["rest"][:-1]))
    s.close()
