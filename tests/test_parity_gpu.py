"""Parity of the CUDA path (through the C ABI) against the CPU oracle.
Integer / index / box results: bit exact.  Positions: relative RMS <= 1e-5
after one substep, <= 1e-3 m RMS after 10 frames (BASELINE.json north_star)."""
import numpy as np
import pytest

import xrtailor_b200 as X
from oracle import oracle as O
from xrtailor_b200 import scene
from tests.common import assert_bvh_equal, pair, rel_rms, rms, stretch_stats

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def small():
    return scene.drop_scene(64, sphere=(24, 48))


def test_bvh_build_bit_exact(small):
    s, o = pair(small)
    assert_bvh_equal(s.bvh(), o.bvh())
    np.testing.assert_array_equal(s.debug("RTRI", np.uint32), o.rtri())
    s.close()


@pytest.mark.parametrize("n,sph", [(8, (4, 6)), (33, (9, 17)), (120, (64, 128))])
def test_bvh_standalone_sizes(n, sph):
    pos, tri = scene.uv_sphere(0.3, *sph)
    g = X.bvh_build_standalone(pos, tri, fast_div=0)
    assert_bvh_equal(g, O.Bvh(pos, tri).arrays())
    pos, tri = scene.grid_cloth(n, jitter=1e-4)
    g = X.bvh_build_standalone(pos, tri, fast_div=0)
    assert_bvh_equal(g, O.Bvh(pos, tri).arrays())


def test_bvh_duplicate_morton_codes():
    # no jitter, flat in y: many identical codes -> the reference's 64-bit key path
    pos, tri = scene.grid_cloth(40, jitter=0.0)
    tri = np.concatenate([tri, tri])  # every face twice
    o = O.Bvh(pos, tri).arrays()
    assert o["used64"] == 1
    assert_bvh_equal(X.bvh_build_standalone(pos, tri, fast_div=0), o)


def test_bvh_tiny():
    pos = np.array([[0, 0, 0], [1, 0, 0], [0, 1, 0], [1, 1, 1]], np.float32)
    for tri in ([[0, 1, 2]], [[0, 1, 2], [1, 2, 3]], [[0, 1, 2], [1, 2, 3], [0, 2, 3]]):
        assert_bvh_equal(X.bvh_build_standalone(pos, np.array(tri, np.uint32), fast_div=0),
                         O.Bvh(pos, np.array(tri, np.uint32)).arrays())


@pytest.mark.parametrize("n,sph,th", [(8, (4, 6), 1e-3), (65, (16, 32), 0.0), (224, (64, 128), 2e-3)])
def test_pair_traversal_bit_exact_and_ordered(n, sph, th):
    # BVH::Traverse = CountPairs + inclusive_scan + FindPairs (BVH.cu:1149-1238): same rows in the same order
    pl, tl = scene.grid_cloth(n, 1.0, 0.3, 1e-4)
    pr, tr = scene.uv_sphere(0.3, *sph)
    want = O.Bvh(pl, tl).pairs(O.Bvh(pr, tr), th)
    got = X.traverse_pairs(pl, tl, pr, tr, thickness=th, fast_div=0)
    assert len(want) > 0
    np.testing.assert_array_equal(got, want)


@pytest.mark.parametrize("n", [3, 40, 150])
def test_pair_traversal_self_bit_exact(n):
    # TraverseSelf (BVH.cu:1240-1340)
    pl, tl = scene.grid_cloth(n, 1.0, 0.0, 1e-4)
    np.testing.assert_array_equal(X.traverse_pairs(pl, tl, thickness=1e-3, fast_div=0), O.Bvh(pl, tl).pairs(None, 1e-3))


def test_pair_traversal_edge_cases():
    pl, tl = scene.grid_cloth(5, 0.5, 5.0, 0.0)
    pr, tr = scene.uv_sphere(0.3, 6, 8)
    assert X.traverse_pairs(pl, tl, pr, tr, thickness=1e-3, fast_div=0).shape == (0, 2)  # disjoint
    one = np.array([[0, 1, 2]], np.uint32)
    p3 = np.array([[0, 0, 0], [1, 0, 0], [0, 1, 0]], np.float32)
    np.testing.assert_array_equal(X.traverse_pairs(p3, one, fast_div=0), [[0, 0]])  # one-leaf tree against itself
    np.testing.assert_array_equal(X.traverse_pairs(p3, one, pr, tr, thickness=2.0, fast_div=0),
                                  O.Bvh(p3, one).pairs(O.Bvh(pr, tr), 2.0))  # everything overlaps
    # duplicate Morton codes (64-bit key path) on both sides
    pd, td = scene.grid_cloth(20, jitter=0.0)
    td = np.concatenate([td, td])
    np.testing.assert_array_equal(X.traverse_pairs(pd, td, thickness=1e-3, fast_div=0), O.Bvh(pd, td).pairs(None, 1e-3))


FEM_MATERIAL = np.array([5000.0, 5000.0, 1000.0, 0.3, 0.25, 0.5], np.float32)


def _fem_case(n, seed=11):
    pos, tri = scene.grid_cloth(n, 1.0, 0.0, 1e-4)
    bend_idx, _ = scene.bend_constraints(pos, tri)
    cs, _ = X.greedy_color(tri, len(pos))
    cb, _ = X.greedy_color(bend_idx, len(pos))
    rng = np.random.default_rng(seed)
    x = pos.copy()
    x[:, 0] *= 1.1                                                     # stretched
    x[:, 1] += 0.03 * np.sin(6 * x[:, 0]).astype(np.float32)           # wrinkled
    x += rng.normal(0, 1e-3, x.shape).astype(np.float32)
    w = np.ones(len(pos), np.float32)
    w[:n] = 0.0                                                        # one pinned border row
    return pos, tri, bend_idx, cs, cb, x, w


@pytest.mark.parametrize("n,iters", [(6, 1), (40, 5), (120, 10)])
def test_fem_strain_and_isometric_bending_match_the_oracle(n, iters):
    # row a25: FEMHelper.cu:94-212 (strain) and :317-388 (bending), colour by colour as SimulateQuality runs them
    pos, tri, bend_idx, cs, cb, x, w = _fem_case(n)
    want = O.fem_solve(pos, x, w, tri, bend_idx, FEM_MATERIAL, cs, cb, iters, 1.0 / 60)
    got = X.fem_solve(pos, x, w, tri, bend_idx, FEM_MATERIAL, cs, cb, iters, 1.0 / 60)
    assert np.isfinite(got).all()
    assert rel_rms(got, want) <= 1e-6, rel_rms(got, want)   # IEEE arithmetic in the same order: rounding only
    assert np.abs(got - x).max() > 1e-4                     # the solve did move the sheet
    np.testing.assert_array_equal(got[:n], x[:n])           # pinned row untouched


def test_fem_improper_colouring_uses_atomics_and_edge_cases():
    pos, tri, bend_idx, cs, cb, x, w = _fem_case(20)
    # everything in one colour: races resolved by float atomics -> same answer to summation-order rounding only
    # when the colouring is proper; here only finiteness and the pins are gated
    got = X.fem_solve(pos, x, w, tri, bend_idx, FEM_MATERIAL, np.zeros_like(cs), np.zeros_like(cb), 3, 1.0 / 60)
    assert np.isfinite(got).all()
    np.testing.assert_array_equal(got[:20], x[:20])
    # no hinges, zero iterations, out-of-range index
    got = X.fem_solve(pos, x, w, tri, np.zeros((0, 4), np.uint32), FEM_MATERIAL, cs, np.zeros(0, np.uint32), 2, 1.0 / 60)
    want = O.fem_solve(pos, x, w, tri, np.zeros((0, 4), np.uint32), FEM_MATERIAL, cs, np.zeros(0, np.uint32), 2, 1.0 / 60)
    assert rel_rms(got, want) <= 1e-6
    np.testing.assert_array_equal(X.fem_solve(pos, x, w, tri, bend_idx, FEM_MATERIAL, cs, cb, 0, 1.0 / 60), x)
    bad = tri.copy()
    bad[0, 0] = len(pos)
    with pytest.raises(X.XrtError):
        X.fem_solve(pos, x, w, bad, bend_idx, FEM_MATERIAL, cs, cb, 1, 1.0 / 60)


def test_refit_midstep_bit_exact_boxes(small):
    s, o = pair(small)
    tgt = scene.sphere_motion(small.obs_pos, 7)
    s.UpdateObstacleAnimationFrame(tgt)
    o.set_obstacle_target(tgt)
    for step in range(3):
        s.BvhRefitObstacle(step)
        o.collision_pass(step)
        g, c = s.bvh(), o.bvh()
        np.testing.assert_array_equal(g["nodes"], c["nodes"])
        # pred = x1 + tol*n1 goes through normals (different summation order): boxes to 1e-6
        np.testing.assert_allclose(g["aabb"], c["aabb"], rtol=0, atol=2e-6)
        np.testing.assert_allclose(s.debug("OBSTACLE_PRED", np.float32).reshape(-1, 9), c["pred"], rtol=0, atol=2e-6)
    s.close()


def _settle(s, o, frames):
    for _ in range(frames):
        s.SimulateSwift()
        o.step_frame()


def test_candidate_sets_and_contacts(small):
    s, o = pair(small, enable_self_collision=0)
    _settle(s, o, 6)
    # same state on both sides, then one generation pass on each
    x0, v = o.positions(), o.velocity()
    s.SetClothState(x0, v)
    s.PredictPositions()
    o.predict()
    # IEEE build: predict is bit exact; pin the inputs of the queries anyway
    np.testing.assert_allclose(s.debug("PREDICTED", np.float32).reshape(-1, 3), o.predicted(), rtol=0, atol=1e-7)
    s.SetClothPredicted(o.predicted())
    s.GenerateContacts()
    o.generate_contacts()
    oc = o.contacts()
    vf, rt, ee = s.contacts()
    for kind, key in ((0, "cnt_vf"), (1, "cnt_rt"), (2, "cnt_ee")):
        off, items = s.candidates(kind)
        np.testing.assert_array_equal(np.diff(off), oc[key])
    # candidate lists in traversal order against the oracle's query
    ob = O.Bvh(small.obs_pos, small.obs_tri)
    off, items = s.candidates(0)
    pred = o.predicted()
    hit = np.nonzero(np.diff(off))[0]
    assert len(hit) > 0
    for i in hit[:200]:
        np.testing.assert_array_equal(items[off[i]:off[i + 1]], ob.query(0, pred[i]))
    off, items = s.candidates(2)
    hit = np.nonzero(np.diff(off))[0]
    for i in hit[:200]:
        e = small.cloth_edges[i]
        np.testing.assert_array_equal(items[off[i]:off[i + 1]], ob.query(1, pred[e[0]], pred[e[1]]))
    np.testing.assert_array_equal(vf["index"], oc["vf_idx"])
    np.testing.assert_allclose(vf["q"], oc["vf_q"], rtol=0, atol=1e-6)
    np.testing.assert_array_equal(rt["index"], oc["rt_idx"])
    np.testing.assert_allclose(rt["q"], oc["rt_q"], rtol=0, atol=1e-6)
    live = oc["ee"][:, 0] < 1e6
    assert live.sum() > 0
    np.testing.assert_array_equal(ee[:, 0] < 1e6, live)
    np.testing.assert_allclose(ee[live], oc["ee"][live, :12], rtol=0, atol=2e-6)
    s.close()


def test_neighbor_lists(small):
    s, o = pair(small)
    s.PredictPositions()
    o.predict()
    s.HashParticles()
    o.hash()
    gh = s.debug("PARTICLE_HASH", np.uint32)
    oh, oi = o.particle_hash()
    V = len(small.cloth_pos)
    np.testing.assert_array_equal(gh[:V], oh)
    np.testing.assert_array_equal(gh[V:], oi)
    np.testing.assert_array_equal(s.debug("NEIGHBORS", np.uint32), o.neighbors())
    s.close()


@pytest.mark.parametrize("self_collision", [0, 1])
def test_one_substep_positions(small, self_collision):
    s, o = pair(small, num_substeps=1, enable_self_collision=self_collision)
    s.SimulateSwift()
    o.step_frame()
    g, c = s.GetPositions(0), o.positions()
    assert rel_rms(g, c) <= 1e-5, rel_rms(g, c)
    # displacement parity too (positions are dominated by the rest pose)
    d0 = small.cloth_pos.astype(np.float64)
    assert rel_rms(g - d0, c - d0) <= 1e-3
    s.close()


def test_ten_frames_positions_moving_obstacle(small):
    s, o = pair(small)
    for f in range(10):
        tgt = scene.sphere_motion(small.obs_pos, f + 1)
        s.UpdateObstacleAnimationFrame(tgt)
        o.set_obstacle_target(tgt)
        s.SimulateSwift()
        o.step_frame()
    g, gn = s.GetPositions(0, normals=True)
    assert np.isfinite(g).all()
    assert rms(g, o.positions()) <= 1e-3, rms(g, o.positions())
    np.testing.assert_allclose(s.GetPositions(1), scene.sphere_motion(small.obs_pos, 10), atol=1e-7)
    # matching constraint-residual and penetration statistics (north_star's third gate)
    st = s.stats()
    o_rms, o_max = stretch_stats(small, o.positions())
    assert abs(st["stretch_rms"] - o_rms) <= 1e-3 and abs(st["stretch_max"] - o_max) <= 2e-2
    oc = o.contacts()
    # statistic, not a set: after 10 frames the EE solve's float atomics have
    # reordered last bits (as they do between two runs of the reference itself)
    n_g, n_o = int((s.contacts()[0]["index"] >= 0).sum()), int((oc["vf_idx"] >= 0).sum())
    assert abs(n_g - n_o) <= max(4, 0.02 * n_o), (n_g, n_o)
    assert st["launches_per_frame"] > 0
    s.close()


def test_graph_replay_equals_eager_building_blocks(small):
    """xrt_step_frame (CUDA graph) == the same frame driven call by call."""
    s1, _ = pair(small, num_substeps=1, num_iterations=4, num_collision_passes=1, enable_self_collision=0,
                 persist_cluster=0)
    s2, _ = pair(small, num_substeps=1, num_iterations=4, num_collision_passes=1, enable_self_collision=0)
    s1.SimulateSwift()
    s2.PredictPositions()
    s2.GenerateContacts()
    for it in range(4):
        s2.SolveIteration(it)
    s2.Finalize()
    np.testing.assert_array_equal(s1.GetPositions(0), s2.GetPositions(0))
    s1.close()
    s2.close()


def test_full_size_properties():
    """C1 at full size (50k verts): properties that need no oracle run."""
    sc = scene.drop_scene(224)
    s = scene.make_solver(sc)
    for _ in range(3):
        s.SimulateSwift()
    x = s.GetPositions(0)
    assert np.isfinite(x).all()
    b = s.bvh()
    F = len(sc.obs_tri)
    assert (np.diff(b["morton"].astype(np.int64)) >= 0).all()          # sortedness
    assert sorted(b["object"].tolist()) == list(range(F))               # permutation
    # every node's box contains its children's boxes
    n, a = b["nodes"], b["aabb"]
    for i in range(F - 1):
        for ch in (n[i, 1], n[i, 2]):
            assert (a[i, :3] <= a[ch, :3]).all() and (a[i, 3:] >= a[ch, 3:]).all()
            assert n[ch, 0] == i
    st = s.stats()
    g_rms, g_max = stretch_stats(sc, x)
    assert abs(st["stretch_rms"] - g_rms) < 1e-5
    # nothing fell through the sphere
    r = np.linalg.norm(x, axis=1)
    assert r.min() > 0.29
    s.close()


def test_error_codes():
    sc = scene.drop_scene(8, sphere=(4, 6))
    s = X.ClothSolver()
    with pytest.raises(X.XrtError) as e:
        s.SimulateSwift()  # not finalized
    assert e.value.code == 8100
    with pytest.raises(X.XrtError) as e:
        s.AddCloth(sc.cloth_pos, np.array([[0, 1, 999999]], np.uint32), sc.cloth_edges)
    assert e.value.code == 8100
    s.close()


# ---- persistent cluster kernel (xrt_persist.cu) vs the per-stage launch shape ----

def _run_frames(sc, frames, **kw):
    s, _ = pair(sc, **kw)
    for f in range(frames):
        s.UpdateObstacleAnimationFrame(scene.sphere_motion(sc.obs_pos, f + 1))
        s.SimulateSwift()
    x = s.GetPositions(0)
    st = s.stats()
    vf = s.contacts()[0]["index"].copy()
    s.close()
    return x, st, vf


@pytest.mark.parametrize("cluster", [1, 2, 8, 16, -2, -3])
def test_persistent_bit_equal_before_contact(cluster):
    """Same device bodies, same phase order: with a proper colouring and no EE
    contact (whose Jacobi sum uses float atomics) the two launch shapes agree
    bit for bit, refit + contact generation + self-collision included."""
    sc = scene.drop_scene(48, sphere=(16, 32), y=0.6)
    x0, st0, vf0 = _run_frames(sc, 3, persist_cluster=0)
    x1, st1, vf1 = _run_frames(sc, 3, persist_cluster=cluster)
    np.testing.assert_array_equal(x0, x1)
    np.testing.assert_array_equal(vf0, vf1)
    assert st1["launches_per_frame"] < st0["launches_per_frame"] / (5 if cluster == -3 else 10)


def test_dsmem_shape_with_bindings_tethers_and_limits():
    """persist_cluster -3 (positions in the distributed shared memory of a 16-CTA cluster, collision passes as
    device-wide kernels in between): the per-stage shape's result on the C2 body scene (bindings, moving body, 4
    substeps, hash every third), bit-equal with long-range attachments before contact; improper colourings are refused."""
    sc, body, bind = scene.body_scene(0.05)
    out = []
    for shape in (0, -3):
        s = scene.make_solver(sc, bindings=bind, persist_cluster=shape)
        for f in range(4):
            s.UpdateObstacleAnimationFrame(body.pose(f + 1))
            s.SimulateSwift()
        out.append(s.GetPositions(0))
        s.close()
    # edge contacts sum their Jacobi terms with float atomics, so in contact two shapes agree to rounding, not in bits
    assert rms(out[0], out[1]) <= 1e-3
    n = 40
    sc = scene.drop_scene(n, sphere=(12, 24), y=0.6, fixed=(0, n - 1))
    lra = scene.corner_tethers(sc, (0, n - 1))
    out = []
    for shape in (0, -3):
        s = scene.make_solver(sc, lra=lra, persist_cluster=shape, geodesic_lra=1, long_range_stretchiness=1.0)
        for f in range(3):
            s.SimulateSwift()
        out.append(s.GetPositions(0))
        s.close()
    np.testing.assert_array_equal(out[0], out[1])
    bad = np.zeros(len(sc.stretch_rest), np.uint32)          # every stretch constraint in one colour: not proper
    with pytest.raises(X.XrtError) as e:
        scene.make_solver(sc, colors=(bad, X.greedy_color(sc.bend_idx, len(sc.cloth_pos))[0]), persist_cluster=-3)
    assert e.value.code == 8100


@pytest.mark.parametrize("cluster", [1, 8, -1, -2, -3])
def test_persistent_in_contact_vs_oracle(small, cluster):
    s, o = pair(small, persist_cluster=cluster)
    for f in range(6):
        tgt = scene.sphere_motion(small.obs_pos, f + 1)
        s.UpdateObstacleAnimationFrame(tgt)
        o.set_obstacle_target(tgt)
        s.SimulateSwift()
        o.step_frame()
    g = s.GetPositions(0)
    assert np.isfinite(g).all()
    assert rms(g, o.positions()) <= 1e-3, rms(g, o.positions())
    assert_bvh_equal(s.bvh(), o.bvh())   # refit inside the persistent kernel: boxes bit exact
    s.close()


@pytest.mark.parametrize("cluster,groups", [(1, 1), (2, 2), (8, 3), (1, 2)])
def test_shared_launch_batch_bit_equal_to_single_handles(cluster, groups):
    """xrt_batch_*: one launch per substep over all garments (one CTA / cluster each) gives every garment
    exactly what it gets stepped on its own stream: different garments, different obstacle motion, a handle
    whose obstacle does not move, hash + BVH build per garment between the shared launches."""
    scs = [scene.drop_scene(40, sphere=(12, 24), y=0.6), scene.drop_scene(33, sphere=(16, 32), y=0.55),
           scene.drop_scene(48, sphere=(8, 16), y=0.6)]
    amps = [0.05, 0.0, 0.03]

    def run(batched):
        ss = [scene.make_solver(sc, persist_cluster=cluster) for sc in scs]
        b = X.Batch(ss, groups=groups) if batched else None
        for f in range(3):
            for s, sc, a in zip(ss, scs, amps):
                if a:
                    s.UpdateObstacleAnimationFrame(scene.sphere_motion(sc.obs_pos, f + 1, amp=a))
            if batched:
                b.SimulateSwiftAsync()
                b.Sync()
            else:
                for s in ss:
                    s.SimulateSwift()
        out = [(s.GetPositions(0), s.debug("VF_CONTACTS", np.uint32)) for s in ss]
        if b:
            b.close()
        for s in ss:
            s.close()
        return out

    for (xa, ca), (xb, cb) in zip(run(False), run(True)):
        np.testing.assert_array_equal(xa, xb)
        np.testing.assert_array_equal(ca, cb)


def test_shared_launch_batch_in_contact_vs_oracle(small):
    ss, oo = zip(*[pair(small, persist_cluster=1) for _ in range(3)])
    b = X.Batch(ss)
    for f in range(6):
        for i, (s, o) in enumerate(zip(ss, oo)):
            tgt = scene.sphere_motion(small.obs_pos, f + 1 + 7 * i)
            s.UpdateObstacleAnimationFrame(tgt)
            o.set_obstacle_target(tgt)
        b.run(1)
        for o in oo:
            o.step_frame()
    for s, o in zip(ss, oo):
        g = s.GetPositions(0)
        assert np.isfinite(g).all()
        assert rms(g, o.positions()) <= 1e-3, rms(g, o.positions())
    b.close()
    for s in ss:
        s.close()


def test_shared_launch_readback_returns_frames_in_order(small):
    """Read-back mode: frame f can be fetched after frame f+1 was enqueued, and is exactly what an immediate
    GetPositions after frame f returns; a third outstanding frame is refused."""
    def run(readback):
        ss = [scene.make_solver(small, persist_cluster=1) for _ in range(3)]
        b = X.Batch(ss, groups=2, workers=1)
        frames = []
        tg = lambda f: np.concatenate([scene.sphere_motion(small.obs_pos, f + 1 + 3 * i) for i in range(3)])
        if readback:
            b.SetReadback(True)
            b.UpdateObstacleAnimationFrames(tg(0))
            b.SimulateSwiftAsync()
            for f in range(1, 5):
                b.UpdateObstacleAnimationFrames(tg(f))
                b.SimulateSwiftAsync()
                frames.append(tuple(a.copy() for a in b.GetPositions(0, normals=True)))
            with pytest.raises(X.XrtError):
                b.SimulateSwiftAsync()
                b.SimulateSwiftAsync()
        else:
            for f in range(4):
                b.UpdateObstacleAnimationFrames(tg(f))
                b.SimulateSwiftAsync()
                frames.append(tuple(a.copy() for a in b.GetPositions(0, normals=True)))
        b.close()
        for s in ss:
            s.close()
        return frames

    for (xa, na), (xb, nb) in zip(run(False), run(True)):
        np.testing.assert_array_equal(xa, xb)
        np.testing.assert_array_equal(na, nb)


def test_shared_launch_batch_rejects_mixed_shapes(small):
    a = scene.make_solver(small, persist_cluster=8)
    c = scene.make_solver(small, persist_cluster=4)
    d = scene.make_solver(small, persist_cluster=0)
    for hs in ([a, c], [d, d]):
        with pytest.raises(X.XrtError):
            X.Batch(hs)
    for s in (a, c, d):
        s.close()


def test_persistent_one_substep_positions(small):
    s, o = pair(small, num_substeps=1, persist_cluster=8)
    s.SimulateSwift()
    o.step_frame()
    assert rel_rms(s.GetPositions(0), o.positions()) <= 1e-5
    s.close()


def test_per_stage_shape_still_matches_oracle(small):
    s, o = pair(small, num_substeps=1, persist_cluster=0)
    s.SimulateSwift()
    o.step_frame()
    assert rel_rms(s.GetPositions(0), o.positions()) <= 1e-5
    s.close()


def _waist_bindings(sc, n=64, seed=3):
    """Tether n cloth vertices to the obstacle triangle under them (a belt)."""
    rng = np.random.default_rng(seed)
    verts = rng.choice(len(sc.cloth_pos), n, replace=False).astype(np.uint32)
    tris = sc.obs_tri[rng.integers(0, len(sc.obs_tri), n)].astype(np.uint32)
    bary = rng.dirichlet(np.ones(3), n).astype(np.float32)
    stiff = rng.uniform(0.2, 1.0, n).astype(np.float32)
    dist = np.where(rng.random(n) < 0.7, 0.0, rng.uniform(0.05, 0.5, n)).astype(np.float32)  # 0: length at setup
    return verts, tris, bary, stiff, dist


@pytest.mark.parametrize("cluster", [0, 8])
def test_binding_constraints_vs_oracle(small, cluster):
    """SolveBinding_Kernel (row a8): tethers pull while the obstacle moves away."""
    b = _waist_bindings(small)
    s, o = pair(small, bindings=b, num_substeps=1, persist_cluster=cluster)
    s0, o0 = pair(small, num_substeps=1, persist_cluster=cluster)
    tgt = scene.sphere_motion(small.obs_pos, 45, amp=0.05)  # sphere 5 cm down: the tethers above it go taut
    for obj in (s, s0):
        obj.UpdateObstacleAnimationFrame(tgt)
    for obj in (o, o0):
        obj.set_obstacle_target(tgt)
    for obj in (s, s0):
        obj.SimulateSwift()
    for obj in (o, o0):
        obj.step_frame()
    g, c = s.GetPositions(0), o.positions()
    assert rel_rms(g, c) <= 1e-5, rel_rms(g, c)
    # the tethers did something, and the same thing on both sides
    moved_g = np.linalg.norm(g - s0.GetPositions(0), axis=1)[b[0]]
    moved_o = np.linalg.norm(c - o0.positions(), axis=1)[b[0]]
    assert (moved_o > 1e-4).sum() >= 8
    np.testing.assert_allclose(moved_g, moved_o, atol=2e-5)
    for obj in (s, s0):
        obj.close()


@pytest.mark.parametrize("cluster", [0, 8, -3])
def test_c2_body_scene_vs_oracle(cluster):
    """BASELINE configs[1] at oracle size: capsule body with a pose sequence, shirt + trousers
    tethered by binding constraints; moving-obstacle contact on every limb."""
    sc, body, bind = scene.body_scene(0.05)
    # one substep from rest (gravity against the tethers, body still): the strict gate
    s, o = pair(sc, bindings=bind, persist_cluster=cluster, num_substeps=1)
    s.SimulateSwift()
    o.step_frame()
    assert rel_rms(s.GetPositions(0), o.positions()) <= 1e-5
    s.close()
    # ten frames of the pose sequence: every limb pushes cloth from the first frame on
    s, o = pair(sc, bindings=bind, persist_cluster=cluster)
    for f in range(10):
        tgt = body.pose(f + 1)
        s.UpdateObstacleAnimationFrame(tgt)
        o.set_obstacle_target(tgt)
        s.SimulateSwift()
        o.step_frame()
    g = s.GetPositions(0)
    assert np.isfinite(g).all()
    assert rms(g, o.positions()) <= 1e-3, rms(g, o.positions())
    assert_bvh_equal(s.bvh(), o.bvh())
    # nothing inside a limb: distance to each capsule's axis segment >= its radius
    pose = body.pose(10)
    for (name, r, *_), (_, lo, hi, _) in zip(scene._BODY_PARTS, body.parts):
        top, bot = pose[lo].astype(np.float64), pose[hi - 1].astype(np.float64)
        ax = (top - bot) / np.linalg.norm(top - bot)
        a, b = top - ax * r, bot + ax * r
        t = np.clip(((g - a) @ (b - a)) / ((b - a) @ (b - a)), 0, 1)
        d = np.linalg.norm(g - (a + t[:, None] * (b - a)), axis=1)
        assert d.min() >= r - 1e-3, (name, d.min(), r)
    s.close()


# ---- edge cases: missing pieces, tiny and ragged inputs ----

@pytest.mark.parametrize("cluster", [0, 8])
def test_no_obstacle_free_fall(cluster):
    sc0 = scene.drop_scene(24, sphere=(6, 8))
    sc = scene.Scene(sc0.cloth_pos, sc0.cloth_tri, fixed=(0, 23))   # two pinned corners, nothing to collide with
    s, o = pair(sc, persist_cluster=cluster)
    for _ in range(3):
        s.SimulateSwift()
        o.step_frame()
    g = s.GetPositions(0)
    assert rel_rms(g, o.positions()) <= 1e-5
    np.testing.assert_array_equal(g[[0, 23]], sc.cloth_pos[[0, 23]])   # pinned vertices do not move
    assert g[:, 1].mean() < sc.cloth_pos[:, 1].mean()                  # the rest falls
    s.close()


@pytest.mark.parametrize("cluster", [0, 1, 16])
def test_switches_off(small, cluster):
    """solve_bending = 0, self collision off, one iteration, one collision pass."""
    s, o = pair(small, persist_cluster=cluster, solve_bending=0, enable_self_collision=0, num_iterations=1,
                num_collision_passes=1, num_substeps=2)
    s.SimulateSwift()
    o.step_frame()
    assert rel_rms(s.GetPositions(0), o.positions()) <= 1e-5
    s.close()


@pytest.mark.parametrize("cluster", [0, 8])
def test_tiny_meshes(cluster):
    """3 x 3 cloth (8 faces, fewer items than one warp) over a 4-triangle obstacle."""
    cp, ct = scene.grid_cloth(3, 0.2, 0.25, 1e-4)
    op = np.array([[0, 0.2, 0], [0.3, -0.1, -0.2], [-0.3, -0.1, -0.2], [0, -0.1, 0.3]], np.float32)
    ot = np.array([[0, 2, 1], [0, 1, 3], [0, 3, 2], [1, 2, 3]], np.uint32)
    sc = scene.Scene(cp, ct, op, ot)
    s, o = pair(sc, persist_cluster=cluster)
    assert_bvh_equal(s.bvh(), o.bvh())
    for _ in range(3):      # (a 9-vertex cloth sliding off a peak is chaotic after that)
        s.SimulateSwift()
        o.step_frame()
    g = s.GetPositions(0)
    assert np.isfinite(g).all()
    assert rms(g, o.positions()) <= 1e-3, rms(g, o.positions())
    s.close()


@pytest.mark.parametrize("cluster", [0, 8])
def test_improper_colouring_uses_atomics(small, cluster):
    """Every constraint in colour 0 (what the reference's randomized colouring degenerates to
    when it conflicts): the projection must fall back to float atomics.  Six un-averaged
    corrections per vertex overshoot, so the garment blows up within a frame (the reference's
    does too); what is checked is that neither launch shape hangs or faults on the way
    (coordinates of 1e12 m once hung the 3x3x3 cell loops of the spatial hash)."""
    cs = np.zeros(len(small.stretch_rest), np.uint32)
    cb = np.zeros(len(small.bend_theta), np.uint32)
    out = []
    for cl in (0, cluster if cluster else 1):
        s = scene.make_solver(small, colors=(cs, cb), persist_cluster=cl, particle_diameter=small.particle_diameter())
        for _ in range(3):
            s.SimulateSwift()
        out.append(s.GetPositions(0))
        assert s.stats()["n_stretch_colors"] == 1
        s.close()
    assert out[0].shape == out[1].shape


def test_hash_survives_a_blown_up_garment(small):
    s, _ = pair(small, persist_cluster=0)
    x = (small.cloth_pos * np.float32(1e12)).astype(np.float32)
    x[::7] = np.inf
    x[3::11] = np.nan
    s.SetClothPredicted(x)
    s.HashParticles()          # returns (it used to spin forever in the cell loops)
    assert s.debug("NEIGHBORS", np.uint32).shape[0] > 0
    s.close()


def test_batch_launch_count_counts_at_the_launch_sites(small):
    """xrt_batch_launch_count: kernels of the library per frame of a uniform shared-launch batch = 2 substep launches
    (substeps 0-2 share one: the hash is due at substeps 0 and 3) + 6 build + 2 x 4 hash + 2 end-of-frame stages per
    group, whatever the number of garments; cub's segmented sort is counted as calls (1 build + 2 hashes)."""
    ss = [scene.make_solver(small, persist_cluster=1) for _ in range(4)]
    b = X.Batch(ss, groups=2)
    for s in ss:
        s.UpdateObstacleAnimationFrame(scene.sphere_motion(small.obs_pos, 1))
    a0, c0 = b.launch_count()
    b.run(1)
    a1, c1 = b.launch_count()
    assert (a1 - a0, c1 - c0) == (2 * (2 + 6 + 8 + 2), 2 * 3)
    b.run(1)  # the obstacle did not move: no tree build
    a2, c2 = b.launch_count()
    assert (a2 - a1, c2 - c1) == (2 * (2 + 8 + 2), 2 * 2)
    b.close()
    for s in ss:
        s.close()
