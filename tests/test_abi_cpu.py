"""CPU-side checks of the boundary: the C-ABI library loads, exports every
symbol include/xrt_b200.h declares, fails loudly without a GPU, and the host
logic (colouring, scene generation, sharding) behaves."""
import ctypes
import os
import re

import numpy as np
import pytest

import xrtailor_b200 as X
from xrtailor_b200 import _abi, batch, scene

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _has_gpu():
    try:
        import torch
        return torch.cuda.is_available()
    except Exception:
        return False


def test_library_exports_every_declared_symbol():
    hdr = open(os.path.join(ROOT, "include", "xrt_b200.h")).read()
    declared = set(re.findall(r"\b(xrt_[a-z_0-9]+)\s*\(", hdr))
    assert len(declared) >= 30
    lib = ctypes.CDLL(_abi.LIB_PATH)
    missing = [s for s in sorted(declared) if not hasattr(lib, s)]
    assert not missing, missing
    assert declared == set(_abi.SYMBOLS)


def test_params_struct_matches_header_defaults():
    p = X.default_params()
    assert (p.num_substeps, p.num_iterations, p.num_collision_passes, p.max_num_neighbors) == (4, 10, 10, 64)
    assert abs(p.gravity[1] + 9.8) < 1e-6 and abs(p.frame_dt - 1 / 60) < 1e-9
    assert p.bend_compliance == 100.0 and abs(p.relaxation_factor - 0.9) < 1e-7
    assert p.inter_leaved_hash == 3 and p.device == 0
    assert ctypes.sizeof(_abi.XrtParams) == 4 * 24


@pytest.mark.skipif(_has_gpu(), reason="checks the no-GPU failure mode")
def test_no_cpu_fallback():
    with pytest.raises(X.XrtError) as e:
        X.ClothSolver()
    assert e.value.code == 8000
    pos, tri = scene.uv_sphere(0.3, 4, 6)
    with pytest.raises(X.XrtError) as e:
        X.bvh_build_standalone(pos, tri)
    assert e.value.code == 8000


def test_product_never_imports_oracle():
    for dirpath, _, files in os.walk(os.path.join(ROOT, "xrtailor_b200")):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                src = open(os.path.join(dirpath, f), errors="ignore").read()
                assert "oracle" not in src.replace("CPU oracle", "").replace("the oracle", ""), f


def test_scene_sizes_match_baseline_md():
    sc = scene.drop_scene(224)
    assert (len(sc.cloth_pos), len(sc.cloth_tri), len(sc.cloth_edges)) == (50176, 99458, 149633)
    assert (len(sc.obs_pos), len(sc.obs_tri)) == (8066, 16128)
    assert len(scene.grid_cloth(317)[0]) == 100489
    # closed sphere: every edge is shared by exactly two faces, normals point outward
    assert len(sc.obs_edges) * 2 == len(sc.obs_tri) * 3
    p, t = sc.obs_pos, sc.obs_tri
    fn = np.cross(p[t[:, 1]] - p[t[:, 0]], p[t[:, 2]] - p[t[:, 0]])
    assert ((p[t].mean(1) * fn).sum(1) > 0).all()


def test_bend_elements_follow_reference_convention():
    # two triangles sharing edge (0,1): flat -> theta 0 with the reference's setup normals
    pos = np.array([[0, 0, 0], [1, 0, 0], [0.5, 0, 1], [0.5, 0, -1]], np.float32)
    tri = np.array([[0, 1, 3], [1, 0, 2]], np.uint32)
    idx, theta = scene.bend_constraints(pos, tri)
    assert idx.tolist() == [[0, 1, 3, 2]] and abs(theta[0]) < 1e-6
    pos[2] = [0.5, 1, 0]  # fold by 90 degrees
    _, theta = scene.bend_constraints(pos, tri)
    assert abs(theta[0] - np.pi / 2) < 1e-6
    sc = scene.drop_scene(16, sphere=(4, 6))
    n = 16
    assert len(sc.bend_idx) == len(sc.cloth_edges) - 4 * (n - 1)  # one per interior edge


@pytest.mark.parametrize("n", [5, 40])
def test_greedy_colouring_is_proper(n):
    sc = scene.drop_scene(n, sphere=(4, 6))
    for idx in (sc.stretch_idx, sc.bend_idx):
        col, nc = X.greedy_color(idx, len(sc.cloth_pos))
        assert nc == col.max() + 1
        for c in range(nc):
            v = idx[col == c].ravel()
            assert len(np.unique(v)) == len(v)  # no vertex twice inside one colour
    assert X.greedy_color(sc.stretch_idx, len(sc.cloth_pos))[1] <= 8


def test_greedy_colouring_rejects_bad_index():
    with pytest.raises(X.XrtError):
        X.greedy_color(np.array([[0, 99]], np.uint32), 4)


def test_shard_partition():
    for n, g in ((512, 8), (7, 2), (3, 4), (0, 2)):
        parts = [batch.shard(n, g, r) for r in range(g)]
        assert sorted(sum(parts, [])) == list(range(n))
        assert max(map(len, parts)) - min(map(len, parts)) <= 1
    with pytest.raises(ValueError):
        batch.shard(4, 2, 2)
    assert batch.frames_per_second(100, [10.0, 20.0]) == 5000.0


def test_build_constraints_equals_the_numpy_generators():
    """xrt_build_constraints (edges, face -> edge, edge -> face, stretch, bend in O(F log F) on the
    host) against the numpy generators the scenes use, and on a non-manifold fin (three faces on one edge: three bend
    elements, like GenerateBendingElements' i < j loop over the edge's face star)."""
    for sc in (scene.drop_scene(24, sphere=(6, 12)), scene.body_scene(0.05)[0], scene.dress_scene(0.05)[0]):
        c = X.build_constraints(sc.cloth_pos, sc.cloth_tri)
        assert np.array_equal(c["edges"], sc.cloth_edges)
        np.testing.assert_allclose(c["stretch_rest"], sc.stretch_rest, rtol=0, atol=1e-7)
        assert np.array_equal(c["bend_idx"], sc.bend_idx)
        np.testing.assert_allclose(c["bend_theta"], sc.bend_theta, rtol=0, atol=1e-6)
        # face -> edge: edge j of a face joins its vertices j and j + 1
        tri, fe = sc.cloth_tri.astype(np.int64), c["face_edges"]
        for j in range(3):
            e = c["edges"][fe[:, j]].astype(np.int64)
            a, b = tri[:, j], tri[:, (j + 1) % 3]
            assert np.array_equal(e[:, 0], np.minimum(a, b)) and np.array_equal(e[:, 1], np.maximum(a, b))
        # edge -> faces: ascending, and exactly the faces that list the edge
        off, ef = c["ef_offsets"], c["ef_faces"]
        assert off[-1] == 3 * len(tri) and (np.diff(off) >= 1).all() and (np.diff(off) <= 2).all()
        for e in (0, len(off) // 2, len(off) - 2):
            faces = ef[off[e]:off[e + 1]]
            assert list(faces) == sorted(faces) and all(e in fe[f] for f in faces)
    pos = np.array([[0, 0, 0], [1, 0, 0], [0, 1, 0], [0, -1, 0], [0, 0, 1]], np.float32)
    tri = np.array([[0, 1, 2], [1, 0, 3], [0, 1, 4]], np.uint32)
    c = X.build_constraints(pos, tri)
    assert len(c["edges"]) == 7 and len(c["bend_idx"]) == 3
    assert sorted(map(tuple, c["bend_idx"][:, 2:].tolist())) == [(2, 3), (2, 4), (3, 4)]
    assert list(c["ef_faces"][c["ef_offsets"][0]:c["ef_offsets"][1]]) == [0, 1, 2]     # edge (0, 1)
    with pytest.raises(X.XrtError):
        X.build_constraints(pos, np.array([[0, 1, 9]], np.uint32))


def test_workload_params_mirror_the_library_defaults():
    """workloads.scene.swift_params (what the reference arm of bench.py uses without loading the product) must stay
    equal to xrt_default_params."""
    p, w = X.default_params(), scene.swift_params()
    for name, _ in _abi.XrtParams._fields_:
        a, b = getattr(p, name), getattr(w, name)
        if name == "gravity":
            assert list(a) == pytest.approx(list(b))
        else:
            assert a == pytest.approx(b), name
