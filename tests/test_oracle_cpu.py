"""The CPU oracle against known answers and brute-force restatements (numpy),
so the checker itself is checked before the GPU parity tests trust it."""
import numpy as np
import pytest

import xrtailor_b200 as X
from oracle import oracle as O
from xrtailor_b200 import scene


def test_expand_bits_and_morton_known_answers():
    # expand_bits inserts two zeros after each of the 10 input bits (BVH.cu:16-24)
    assert O.L.oracle_expand_bits(0) == 0
    assert O.L.oracle_expand_bits(1) == 1
    assert O.L.oracle_expand_bits(0b11) == 0b1001
    assert O.L.oracle_expand_bits(0x3FF) == 0x09249249
    for v in (5, 77, 512, 1023):
        ref = sum(((v >> b) & 1) << (3 * b) for b in range(10))
        assert O.L.oracle_expand_bits(v) == ref
    # morton = xx*4 + yy*2 + zz with 10-bit quantisation, clamped to [0, 1023]
    assert O.L.oracle_morton_code(0.0, 0.0, 0.0) == 0
    assert O.L.oracle_morton_code(1.0, 1.0, 1.0) == 0x3FFFFFFF
    assert O.L.oracle_morton_code(2.0, -1.0, 0.5) == (0x09249249 << 2) + O.L.oracle_expand_bits(512)
    assert O.L.oracle_morton_code(float("nan"), 0.0, 0.0) == 0
    assert O.L.oracle_morton_code(1 / 1024, 0.0, 0.0) == 4


def _brute_tree_checks(b, F):
    n, aabb = b["nodes"], b["aabb"]
    NI = F - 1
    assert n[0, 0] == 0xFFFFFFFF
    # every non-root node has exactly one parent that lists it as a child
    for i in range(1, 2 * F - 1):
        p = n[i, 0]
        assert i in (n[p, 1], n[p, 2])
    # leaves hold the sorted objects, boxes are the exact union of children
    np.testing.assert_array_equal(n[NI:, 3], b["object"])
    for i in range(NI):
        l, r = n[i, 1], n[i, 2]
        np.testing.assert_array_equal(aabb[i, :3], np.minimum(aabb[l, :3], aabb[r, :3]))
        np.testing.assert_array_equal(aabb[i, 3:], np.maximum(aabb[l, 3:], aabb[r, 3:]))
    assert (np.diff(b["morton"].astype(np.int64)) >= 0).all()


def _dfs_escape(n, F):
    """Escape links from a pre-order walk: the node visited after a subtree."""
    NI = F - 1
    order, stack = [], [0]
    size = {}
    def walk(i):
        order.append(i)
        if n[i, 3] != 0xFFFFFFFF:
            size[i] = 1
            return 1
        s = 1 + walk(int(n[i, 1])) + walk(int(n[i, 2]))
        size[i] = s
        return s
    import sys
    sys.setrecursionlimit(10000)
    walk(0)
    pos = {v: k for k, v in enumerate(order)}
    esc = {}
    for v in order:
        k = pos[v] + size[v]
        esc[v] = order[k] if k < len(order) else -1
    return esc


@pytest.mark.parametrize("mesh", ["sphere", "grid", "dups"])
def test_bvh_structure_and_escape_tables(mesh):
    if mesh == "sphere":
        pos, tri = scene.uv_sphere(0.3, 10, 16)
    elif mesh == "grid":
        pos, tri = scene.grid_cloth(12, jitter=1e-4)
    else:
        pos, tri = scene.grid_cloth(9, jitter=0.0)
        tri = np.concatenate([tri, tri[:17]])
    F = len(tri)
    b = O.Bvh(pos, tri).arrays()
    assert b["used64"] == (1 if mesh == "dups" else 0) or mesh == "grid"
    _brute_tree_checks(b, F)
    esc = _dfs_escape(b["nodes"], F)
    NI = F - 1
    assert [esc[i] for i in range(NI)] == b["esc_int"].tolist()
    assert [esc[NI + i] for i in range(F)] == b["esc_ext"].tolist()
    # the O(1) formula the CUDA build uses (xrt_bvh.cu escape_after) gives the same links
    first = {}
    def rng(i):
        if b["nodes"][i, 3] != 0xFFFFFFFF:
            k = i - NI
            return k, k
        l = rng(int(b["nodes"][i, 1])); r = rng(int(b["nodes"][i, 2]))
        first[i] = (l[0], r[1])
        return first[i]
    rng(0)
    def after(last):
        if last >= F - 1:
            return -1
        k = last + 1
        return k if (k <= F - 2 and first[k][0] == k) else NI + k
    assert [after(first[i][1]) for i in range(NI)] == b["esc_int"].tolist()
    assert [after(i) for i in range(F)] == b["esc_ext"].tolist()


def test_stackless_queries_equal_brute_force():
    pos, tri = scene.uv_sphere(0.3, 12, 20)
    F = len(tri)
    bv = O.Bvh(pos, tri)
    b = bv.arrays()
    leaf = b["aabb"][F - 1:]
    obj = b["object"]
    rng = np.random.default_rng(scene.SEED)
    pts = rng.uniform(-0.35, 0.35, (300, 3)).astype(np.float32)
    hits_total = 0
    for p in pts:
        inside = np.all((p > leaf[:, :3]) & (p < leaf[:, 3:]), axis=1)
        got = bv.query(0, p)
        np.testing.assert_array_equal(got, obj[inside])  # traversal order == leaf order
        hits_total += len(got)
    assert hits_total > 20
    for _ in range(200):
        p1 = rng.uniform(-0.4, 0.4, 3).astype(np.float32)
        p2 = (p1 + rng.uniform(-0.05, 0.05, 3)).astype(np.float32)
        brute = [obj[i] for i in range(F) if O.L.oracle_overlap_segment_f(O._pf(leaf[i].copy()), O._pf(p1), O._pf(p2))]
        np.testing.assert_array_equal(bv.query(1, p1, p2), np.array(brute, np.uint32))


def test_nearest_point_against_dense_sampling():
    rng = np.random.default_rng(7)
    for _ in range(100):
        t = rng.normal(size=(3, 3)).astype(np.float32)
        p = rng.normal(size=3).astype(np.float32) * 2
        qs = np.zeros(3, np.float32)
        d = O.L.oracle_nearest_point_f(O._pf(p), O._pf(t.ravel().copy()), O._pf(qs))
        u, v = np.meshgrid(np.linspace(0, 1, 120), np.linspace(0, 1, 120))
        m = u + v <= 1
        cand = t[0] + u[m][:, None] * (t[1] - t[0]) + v[m][:, None] * (t[2] - t[0])
        brute = np.linalg.norm(cand - p, axis=1).min()
        assert d <= brute + 1e-5 and d >= brute - 0.03
        assert abs(np.linalg.norm(p - qs) - d) < 1e-5


def test_ray_and_segment_known_answers():
    tri = np.array([0, 0, 0, 1, 0, 0, 0, 1, 0], np.float32)
    o = np.array([0.25, 0.25, 1.0], np.float32)
    d = np.array([0, 0, -1], np.float32)
    assert abs(O.L.oracle_ray_intersect_f(O._pf(o), O._pf(d), 2.0, O._pf(tri)) - 1.0) < 1e-6
    assert O.L.oracle_ray_intersect_f(O._pf(o), O._pf(d), 0.5, O._pf(tri)) == np.inf       # too short: "no hit"
    o2 = np.array([2, 2, 1], np.float32)
    assert O.L.oracle_ray_intersect_f(O._pf(o2), O._pf(d), 2.0, O._pf(tri)) == np.float32(3.4028235e38)  # outside
    par = np.array([1, 0, 0], np.float32)
    assert O.L.oracle_ray_intersect_f(O._pf(o), O._pf(par), 2.0, O._pf(tri)) == np.float32(3.4028235e38)  # parallel
    # crossing segments at distance 1 (Ericson 5.1.9)
    seg = np.array([-1, 0, 0, 1, 0, 0, 0, -1, 1, 0, 1, 1], np.float32)
    st, cc = np.zeros(2, np.float32), np.zeros(6, np.float32)
    dist = O.L.oracle_closest_seg_seg_f(O._pf(seg), O._pf(st), O._pf(cc))
    assert abs(dist - 1) < 1e-6 and np.allclose(st, [0.5, 0.5]) and np.allclose(cc, [0, 0, 0, 0, 0, 1])
    # both degenerate
    seg = np.array([0, 0, 0, 0, 0, 0, 3, 4, 0, 3, 4, 0], np.float32)
    assert abs(O.L.oracle_closest_seg_seg_f(O._pf(seg), O._pf(st), O._pf(cc)) - 5) < 1e-6


def test_rtriangle_owns_every_edge_and_vertex_once():
    pos, tri = scene.uv_sphere(0.3, 8, 12)
    info = O.rtri(tri, len(pos))
    verts = np.zeros(len(pos), int)
    edges = {}
    for f, t in enumerate(tri):
        for i in range(3):
            if (info[f] >> i) & 1:
                verts[t[i]] += 1
            if (info[f] >> (i + 3)) & 1:
                e = tuple(sorted((int(t[i]), int(t[(i + 1) % 3]))))
                edges[e] = edges.get(e, 0) + 1
    assert (verts == 1).all()
    assert len(edges) == len(scene.unique_edges(tri)) and set(edges.values()) == {1}


def test_free_fall_and_rest_are_exact():
    """No obstacle, no constraints violated: positions follow x += v dt with v += g dt (damped)."""
    sc = scene.Scene(*scene.grid_cloth(6, jitter=0.0))
    p = X.default_params()
    p.particle_diameter = sc.particle_diameter()
    p.enable_self_collision = 0
    cs, _ = X.greedy_color(sc.stretch_idx, len(sc.cloth_pos))
    cb, _ = X.greedy_color(sc.bend_idx, len(sc.cloth_pos))
    o = O.Scene(sc, p, (cs, cb))
    o.step_frame()
    dt = np.float32(p.frame_dt / p.num_substeps)
    y, v = np.float32(sc.cloth_pos[0, 1]), np.float32(0)
    for _ in range(p.num_substeps):
        v = np.float32(v + np.float32(-9.8) * dt)
        yn = np.float32(y + v * dt)
        v = np.float32(np.float32((yn - y) / dt) * np.float32(1 - np.float32(p.damping) * dt))
        y = yn
    # rigid translation: stretch keeps rest lengths to rounding, bend adds an in-plane push of ~1e-7
    assert np.abs(o.positions()[:, 1] - y).max() < 2e-5
    assert np.abs(o.velocity()[:, 1] - v).max() < 2e-3
    # pinned cloth stays put
    sc2 = scene.Scene(*scene.grid_cloth(6, jitter=0.0), fixed=range(36))
    o2 = O.Scene(sc2, p, (cs, cb))
    o2.step_frame()
    np.testing.assert_array_equal(o2.positions(), sc2.cloth_pos)


def test_contact_keeps_cloth_outside_sphere():
    sc = scene.drop_scene(24, sphere=(12, 24))
    p = X.default_params()
    p.particle_diameter = sc.particle_diameter()
    cs, _ = X.greedy_color(sc.stretch_idx, len(sc.cloth_pos))
    cb, _ = X.greedy_color(sc.bend_idx, len(sc.cloth_pos))
    o = O.Scene(sc, p, (cs, cb))
    for _ in range(20):
        o.step_frame()
    x = o.positions()
    assert np.isfinite(x).all()
    r = np.linalg.norm(x, axis=1)
    assert r.min() > 0.295          # nothing tunnels into the r = 0.3 sphere (faceted: ~0.297 at face centres)
    c = o.contacts()
    assert (c["vf_idx"] >= 0).sum() > 10 and (c["ee"][:, 0] < 1e6).sum() > 10


def test_binding_known_answer():
    """SolveBinding_Kernel by hand: vertex 1 m above its anchor, rest length 0.25, stiffness 0.5
    -> C = 0.75, moved down by 0.5 * 0.75; slack tethers and pinned vertices do not move."""
    from xrtailor_b200 import scene
    sc = scene.drop_scene(4, sphere=(4, 6))
    p = X.default_params()
    p.particle_diameter = sc.particle_diameter()
    cs, _ = X.greedy_color(sc.stretch_idx, len(sc.cloth_pos))
    cb, _ = X.greedy_color(sc.bend_idx, len(sc.cloth_pos))
    o = O.Scene(sc, p, (cs, cb))
    tri = sc.obs_tri[:2].astype(np.uint32)
    bary = np.array([[1, 0, 0], [0.25, 0.25, 0.5]], np.float32)
    anchor = np.einsum("nk,nkd->nd", bary, sc.obs_pos[tri])
    o.set_binding(np.array([0, 1], np.uint32), tri, bary, np.array([0.5, 1.0], np.float32),
                  np.array([0.25, 100.0], np.float32))
    x = o.predicted().copy()
    x[0] = anchor[0] + [0, 1.0, 0]
    x[1] = anchor[1] + [0, 1.0, 0]          # slack: rest length 100
    o.set_predicted(x)
    o.solve_binding()
    y = o.predicted()
    np.testing.assert_allclose(y[0], anchor[0] + [0, 1.0 - 0.5 * 0.75, 0], atol=1e-6)
    np.testing.assert_array_equal(y[1], x[1])


def _leaf_boxes(pos, tri):
    p = pos[tri]
    return p.min(1), p.max(1)


def _brute_pairs(pl, tl, pr, tr, th):
    lo_l, hi_l = _leaf_boxes(pl, tl)
    lo_r, hi_r = _leaf_boxes(pr, tr)
    lo_q, hi_q = lo_l - np.float32(th), hi_l + np.float32(th)
    hit = np.all((lo_r[None] <= hi_q[:, None]) & (hi_r[None] >= lo_q[:, None]), axis=2)
    return {(int(a), int(b)) for a, b in zip(*np.nonzero(hit))}


@pytest.mark.parametrize("th", [0.0, 2e-3])
def test_pair_traversal_equals_brute_force_and_keeps_the_reference_order(th):
    # BVH.cu:1149-1238: every overlapping leaf pair exactly once, grouped by lhs leaf in sorted order
    pl, tl = scene.grid_cloth(21, 0.5, 0.29, 1e-4)
    pr, tr = scene.uv_sphere(0.3, 12, 20)
    bl, br = O.Bvh(pl, tl), O.Bvh(pr, tr)
    got = bl.pairs(br, th)
    want = _brute_pairs(pl, tl, pr, tr, th)
    assert len(got) == len(want) > 0
    assert {(int(a), int(b)) for a, b in got} == want
    order = {int(o): i for i, o in enumerate(bl.arrays()["object"])}
    rank = np.array([order[int(a)] for a in got[:, 0]])
    assert np.all(np.diff(rank) >= 0)
    # within one lhs leaf the rhs leaves come out in reverse traversal (= descending sorted-leaf) order
    r_order = {int(o): i for i, o in enumerate(br.arrays()["object"])}
    for leaf in np.unique(rank)[:50]:
        rr = [r_order[int(b)] for b in got[rank == leaf, 1]]
        assert rr == sorted(rr, reverse=True)


def test_pair_traversal_self_has_the_diagonal_and_is_symmetric():
    # CountPairsSelf/FindPairsSelf, BVH.cu:1240-1340: lhs == rhs, nothing filtered at this stage
    pl, tl = scene.grid_cloth(17, 0.5, 0.0, 1e-4)
    b = O.Bvh(pl, tl)
    got = {(int(a), int(c)) for a, c in b.pairs(None, 1e-3)}
    assert got == _brute_pairs(pl, tl, pl, tl, 1e-3)
    assert all((a, a) in got for a in range(len(tl)))
    assert all((c, a) in got for a, c in got)


def test_pair_traversal_disjoint_and_single_face():
    pl, tl = scene.grid_cloth(5, 0.5, 5.0, 0.0)
    pr, tr = scene.uv_sphere(0.3, 6, 8)
    assert len(O.Bvh(pl, tl).pairs(O.Bvh(pr, tr), 1e-3)) == 0
    one = np.array([[0, 1, 2]], np.uint32)
    p3 = np.array([[0, 0, 0], [1, 0, 0], [0, 1, 0]], np.float32)
    np.testing.assert_array_equal(O.Bvh(p3, one).pairs(None, 0.0), [[0, 0]])


# ---- Quality-mode cloth model (row a25): FEM strain + isometric bending --------------------------
MATERIAL = np.array([5000.0, 5000.0, 1000.0, 0.3, 0.3, 1e-3], np.float32)  # xx, yy, xy, nu_xy, nu_yx, bending


def _fem_case(n=9):
    pos, tri = scene.grid_cloth(n, 0.5, 0.0, 1e-4)
    bend_idx, _ = scene.bend_constraints(pos, tri)
    cs, _ = X.greedy_color(tri, len(pos))
    cb, _ = X.greedy_color(bend_idx, len(pos))
    return pos, tri, bend_idx, cs, cb


def _strain_energy(rest, x, tri):
    # Green strain energy with unit isotropic stiffness: a scale-free progress measure for the tests
    e = 0.0
    for t in tri:
        R = np.stack([rest[t[1]] - rest[t[0]], rest[t[2]] - rest[t[0]]], 1).astype(np.float64)
        D = np.stack([x[t[1]] - x[t[0]], x[t[2]] - x[t[0]]], 1).astype(np.float64)
        q, r = np.linalg.qr(R)
        Fm = D @ np.linalg.inv(r)
        G = 0.5 * (Fm.T @ Fm - np.eye(2))
        e += 0.5 * abs(np.linalg.det(r)) * (G * G).sum()
    return e


def test_fem_rest_state_is_a_fixed_point():
    # FEMHelper.cu:94-175: zero Green strain -> zero energy -> zero correction; a flat sheet has zero
    # isometric bending energy (Q x = 0), so the early exit at :343 is taken
    pos, tri, bend_idx, cs, cb = _fem_case()
    w = np.ones(len(pos), np.float32)
    out = O.fem_solve(pos, pos, w, tri, bend_idx, MATERIAL, cs, cb, 3, 1.0 / 60)
    np.testing.assert_allclose(out, pos, rtol=0, atol=2e-6)


def test_fem_strain_pulls_a_stretched_sheet_back_and_respects_pins():
    pos, tri, bend_idx, cs, cb = _fem_case()
    x = pos.copy()
    x[:, 0] *= 1.2                      # 20 % stretch along x
    w = np.ones(len(pos), np.float32)
    w[:3] = 0.0                          # pinned vertices
    e0 = _strain_energy(pos, x, tri)
    mid = O.fem_solve(pos, x, w, tri, bend_idx, MATERIAL, cs, cb, 5, 1.0 / 60)
    out = O.fem_solve(pos, x, w, tri, bend_idx, MATERIAL, cs, cb, 20, 1.0 / 60)
    assert np.isfinite(out).all()
    # the constraint is the element's energy itself (C = A psi, step = C / |grad C|^2): monotone, not fast
    assert _strain_energy(pos, out, tri) < _strain_energy(pos, mid, tri) < e0
    assert _strain_energy(pos, out, tri) < 0.5 * e0
    np.testing.assert_array_equal(out[:3], x[:3])


def test_fem_is_invariant_under_rigid_motion():
    pos, tri, bend_idx, cs, cb = _fem_case(7)
    rng = np.random.default_rng(3)
    x = pos + rng.normal(0, 5e-3, pos.shape).astype(np.float32)
    w = np.ones(len(pos), np.float32)
    a = 0.7
    R = np.array([[np.cos(a), -np.sin(a), 0], [np.sin(a), np.cos(a), 0], [0, 0, 1]], np.float32)
    t = np.array([0.3, -0.2, 0.5], np.float32)
    out = O.fem_solve(pos, x, w, tri, bend_idx, MATERIAL, cs, cb, 4, 1.0 / 60)
    out_r = O.fem_solve(pos, x @ R.T + t, w, tri, bend_idx, MATERIAL, cs, cb, 4, 1.0 / 60)
    np.testing.assert_allclose(out_r, out @ R.T + t, rtol=0, atol=5e-5)


def test_fem_bending_flattens_a_fold():
    pos, tri, bend_idx, cs, cb = _fem_case()
    x = pos.copy()
    right = x[:, 0] > x[:, 0].mean()
    x[right, 1] += (x[right, 0] - x[:, 0].mean()) * 0.8   # fold the right half up (an isometry of each half)
    x[right, 0] = x[:, 0].mean() + (x[right, 0] - x[:, 0].mean()) * 0.6
    w = np.ones(len(pos), np.float32)
    stiff = MATERIAL.copy()
    stiff[5] = 10.0
    def fold_angle(p):
        # largest dihedral angle over the hinges ([0],[1] shared edge, [2],[3] opposite vertices)
        a, b, c, d = (p[bend_idx[:, k]].astype(np.float64) for k in range(4))
        n1, n2 = np.cross(b - a, c - a), np.cross(d - a, b - a)
        cosang = (n1 * n2).sum(1) / np.linalg.norm(n1, axis=1) / np.linalg.norm(n2, axis=1)
        return np.arccos(np.clip(cosang, -1, 1)).max()

    one = O.fem_solve(pos, x, w, tri, bend_idx, stiff, cs, cb, 1, 1.0 / 60)
    out = O.fem_solve(pos, x, w, tri, bend_idx, stiff, cs, cb, 30, 1.0 / 60)
    assert np.isfinite(out).all()
    assert fold_angle(out) < fold_angle(one) < fold_angle(x)
    # one iteration moves the vertices of the hinges across the fold, the rest only by the 1e-4 jitter's worth
    # (each half is an isometry of its rest shape)
    far = np.abs(pos[:, 0] - pos[:, 0].mean()) > 0.15
    move = np.linalg.norm(one - x, axis=1)
    assert move[far].max() < 0.1 * move[~far].max(), (move[far].max(), move[~far].max())


# ---- Quality mode (oracle/xrt_oracle_quality.c) ----

def _quality_oracle(sc, N=20):
    import types
    q = types.SimpleNamespace(xx_stiffness=1.0, yy_stiffness=1.0, xy_stiffness=1.0, xy_poisson_ratio=0.3,
                              yx_poisson_ratio=0.3, bending_stiffness=1e-2, enable_imminent_repulsion=1,
                              imminent_thickness=1e-3, enable_pbd_repulsion=1, pbd_thickness=1e-4, relaxation_rate=0.75,
                              obstacle_mass=1e3, impact_zone_thickness=1e-4)
    p = X.default_params()
    p.num_substeps, p.num_iterations, p.enable_self_collision = 1, N, 0
    p.particle_diameter = sc.particle_diameter()
    cs, _ = X.greedy_color(sc.stretch_idx, len(sc.cloth_pos))
    cb, _ = X.greedy_color(sc.bend_idx, len(sc.cloth_pos))
    cf, _ = X.greedy_color(sc.cloth_tri, len(sc.cloth_pos))
    o = O.Scene(sc, p, (cs, cb))
    o.set_quality(sc, q, cf)
    return o, q


def test_solve_cubic_and_check_impact_known_answers():
    """MathFunctions::SolveCubic (MathFunctions.cu:122-146) on cubics with known roots, and the continuous vertex-face
    test (ImpactHelper.cu:34-87) on a vertex falling through the middle of a static triangle."""
    r = sorted(O.solve_cubic(1.0, -6.0, 11.0, -6.0))            # (t - 1)(t - 2)(t - 3)
    np.testing.assert_allclose(r, [1, 2, 3], atol=1e-4)
    r = O.solve_cubic(1.0, 0.0, 1.0, -2.0)                      # t^3 + t - 2: one real root, t = 1
    assert len(r) == 1 and abs(r[0] - 1.0) < 1e-5
    r = sorted(O.solve_cubic(0.0, 1.0, -3.0, 2.0))              # degenerates to (t - 1)(t - 2)
    np.testing.assert_allclose(r, [1, 2], atol=1e-5)
    # a scene of one triangle pair is enough to drive FindImpacts: vertex 3 of the cloth falls through face (0, 1, 2)
    pos = np.array([[0, 0, 0], [1, 0, 0], [0, 1, 0], [0.25, 0.25, 0.5], [2, 2, 0.5], [2, 3, 0.5]], np.float32)
    tri = np.array([[0, 1, 2], [3, 4, 5]], np.uint32)
    sc = scene.Scene(pos, tri, None, None)
    o, q = _quality_oracle(sc)
    x1 = pos.copy()
    x1[3:, 2] -= 1.0                                             # the second triangle moves 1 m down in the step
    o.set_state(pos, np.zeros_like(pos))
    o.set_predicted(x1)
    nodes, twn, npairs = o.find_impacts(1e-4)
    vf = [(n, t) for n, t in zip(nodes, twn) if t[8] > 0.5 and n[0] == 3]
    assert vf, (nodes, twn)
    n, t = vf[0]
    assert sorted(n[1:]) == [0, 1, 2]
    assert abs(t[0] - 0.5) < 1e-5                                # reaches z = 0 at half the step
    np.testing.assert_allclose(sorted(-t[2:5]), sorted([0.5, 0.25, 0.25]), atol=1e-5)   # barycentric weights
    assert abs(abs(t[7]) - 1.0) < 1e-5                           # normal along z


def test_quality_oracle_against_golden_reference_outputs():
    """tests/golden/quality_*_s16.npz hold what the REFERENCE's own CUDA returned on a B200 for a two-sheet scene
    (scratch/r02_quality_ref.py wrote them from oracle/_ref through the harness): FindImpacts rows, the
    independentImpacts selection, and the positions after CollisionStep.  The CPU oracle must reproduce them here."""
    import os
    g = np.load(os.path.join(os.path.dirname(__file__), "golden", "quality_impacts_s16.npz"))
    sc = scene.layered_scene(16, 2, gap=4e-3, sphere=(8, 16))
    np.testing.assert_array_equal(sc.cloth_pos, g["cloth_pos"])  # the generator still makes the scene of the fixture
    np.testing.assert_array_equal(sc.cloth_tri, g["cloth_tri"])
    o, q = _quality_oracle(sc)
    o.set_state(g["x0"], g["v"])
    o.set_predicted(g["x1"])
    nodes, twn, npairs = o.find_impacts(float(g["thickness"]))
    assert npairs == int(g["ref_pairs"]) == 29608
    rn, rt = g["ref_nodes"], g["ref_twn"]
    assert len(nodes) == len(rn) == 19260
    np.testing.assert_array_equal(nodes, rn)                      # the reference's rows, in the reference's order
    np.testing.assert_array_equal(twn[:, 8], rt[:, 8])
    assert np.abs(twn[:, 0] - rt[:, 0]).max() <= 5e-5 and np.median(np.abs(twn[:, 0] - rt[:, 0])) <= 1e-6
    inn, itw = o.independent_impacts(rn, rt, 0)
    np.testing.assert_array_equal(inn, g["ref_ind_nodes"])       # 102 impacts, none sharing a movable node
    np.testing.assert_array_equal(itw[:, 0], g["ref_ind_twn"][:, 0])
    assert len(inn) == 102
    # CollisionStep from that state: every impact resolved, within 1e-3 m RMS of the reference's result
    o.collision_step()
    st = o.collision_stats()
    assert st["success"] == 1
    gs = np.load(os.path.join(os.path.dirname(__file__), "golden", "quality_step_s16.npz"))
    d = np.sqrt(((o.predicted() - gs["ref_x"]) ** 2).sum(1))
    assert np.sqrt((d ** 2).mean()) <= 1e-3
    left, _, _ = o.find_impacts(float(g["thickness"]))
    assert len(left) == 0
