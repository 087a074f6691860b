"""Body-animation producer, CPU side: the oracle's restatement of Rodrigues, the joint hierarchy
and linear blend skinning (oracle/xrt_oracle_rig.c) against a dense numpy formulation in float64 (4x4 matrices, full
weight matrix), and the capsule-body rig of the workloads against Body.pose."""
import numpy as np

from oracle import oracle as O
from workloads import scene


def rodrigues64(v):
    v = np.asarray(v, np.float64)
    th = np.linalg.norm(v)
    if th < 1e-5:
        return np.eye(3)
    r = v / th
    K = np.array([[0, -r[2], r[1]], [r[2], 0, -r[0]], [-r[1], r[0], 0]])
    return np.cos(th) * np.eye(3) + (1 - np.cos(th)) * np.outer(r, r) + np.sin(th) * K


def dense_skinning(rest, W, joints, parent, pose, trans, M=None, B=None):
    """textbook SMPL: G_j = G_parent * [R_j | J_j - J_parent], vertex = sum_j w_vj G_j [I | -J_j] v"""
    J = len(joints)
    G = np.zeros((J, 4, 4))
    for j in range(J):
        L = np.eye(4)
        L[:3, :3] = rodrigues64(pose[j])
        L[:3, 3] = joints[j] + trans if j == 0 else joints[j] - joints[parent[j]]
        G[j] = L if j == 0 else G[parent[j]] @ L
    A = G.copy()
    for j in range(J):
        A[j, :3, 3] -= G[j, :3, :3] @ joints[j]
    v = rest.astype(np.float64)
    if B is not None:
        par = np.concatenate([(rodrigues64(pose[j]) - np.eye(3)).ravel() for j in range(1, J)])
        v = v + (B.astype(np.float64) @ par).reshape(-1, 3)
    vh = np.concatenate([v, np.ones((len(v), 1))], 1)
    out = np.einsum("vj,jab,vb->va", W.astype(np.float64), A, vh)[:, :3]
    if M is not None:
        M = np.asarray(M, np.float64).reshape(4, 4)
        out = out @ M[:3, :3].T + M[:3, 3]
    return out, A


def random_rig(rng, V=500, J=9):
    joints = rng.uniform(-1, 1, (J, 3)).astype(np.float32)
    parent = np.array([-1] + [int(rng.integers(0, j)) for j in range(1, J)], np.int32)
    rest = rng.uniform(-1, 1, (V, 3)).astype(np.float32)
    W = np.zeros((V, J), np.float32)
    outer, inner, val = [0], [], []
    for v in range(V):
        k = int(rng.integers(1, 5))
        js = np.sort(rng.choice(J, k, replace=False))
        w = rng.uniform(0.1, 1, k).astype(np.float32)
        w /= w.sum()
        W[v, js] = w
        inner += list(js)
        val += list(w)
        outer.append(len(inner))
    return rest, W, np.array(outer, np.uint32), np.array(inner, np.uint32), np.array(val, np.float32), joints, parent


def test_rodrigues_and_joint_hierarchy_against_dense_float64():
    rng = np.random.default_rng(20261104)
    rest, W, wo, wi, wv, joints, parent = random_rig(rng)
    for trial in range(4):
        pose = rng.uniform(-1.2, 1.2, (len(joints), 3)).astype(np.float32)
        pose[2] = 0.0  # the identity branch of Rodrigues
        trans = rng.uniform(-0.5, 0.5, 3).astype(np.float32)
        T = O.rig_joint_transforms(pose, joints, parent, trans)
        ref, A = dense_skinning(rest, W, joints, parent, pose, trans)
        assert np.abs(T.reshape(-1, 3, 4) - A[:, :3, :]).max() <= 2e-6
        got = O.rig_lbs(rest, wo, wi, wv, T)
        assert np.abs(got - ref).max() <= 5e-6


def test_model_matrix_and_pose_blend_shapes_against_dense_float64():
    rng = np.random.default_rng(7)
    rest, W, wo, wi, wv, joints, parent = random_rig(rng, V=200, J=6)
    P = 9 * (len(joints) - 1)
    B = (rng.standard_normal((3 * len(rest), P)) * 0.01).astype(np.float32)
    ang = 0.5
    M = np.array([[1, 0, 0, 0.1], [0, np.cos(ang), -np.sin(ang), -0.2], [0, np.sin(ang), np.cos(ang), 0.3], [0, 0, 0, 1]], np.float32)
    pose = rng.uniform(-1, 1, (len(joints), 3)).astype(np.float32)
    T = O.rig_joint_transforms(pose, joints, parent)
    par = O.rig_pose_blend_params(pose)
    got = O.rig_lbs(rest, wo, wi, wv, T, model_matrix=M, pose_blend=B, blend_params=par)
    ref, _ = dense_skinning(rest, W, joints, parent, pose, np.zeros(3), M=M, B=B)
    assert np.abs(got - ref).max() <= 5e-6


def test_capsule_body_rig_reproduces_body_pose():
    body = scene.Body()
    rig = scene.body_rig(body, blend=0.0)
    assert rig["w_outer"][-1] == len(body.rest) and (rig["parent"][1:] == 0).all()
    for f in (1, 7, 23):
        pose = scene.body_pose_params(body, f)
        T = O.rig_joint_transforms(pose, rig["joints"], rig["parent"])
        got = O.rig_lbs(rig["rest"], rig["w_outer"], rig["w_inner"], rig["w_val"], T)
        assert np.abs(got - body.pose(f)).max() <= 2e-6
    blended = scene.body_rig(body)  # limbs blend into the torso below their joints: weights still sum to one
    s = np.add.reduceat(blended["w_val"], blended["w_outer"][:-1].astype(np.int64))
    assert np.allclose(s, 1.0) and len(blended["w_val"]) > len(body.rest)
