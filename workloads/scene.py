"""Synthetic scenes of BASELINE.md (C1 / C100k / C2-like) and the setup-time
derivations the reference performs on the host before the solver starts:
unique edges, stretch constraints (PrepareBasicStretchData,
xrtailor/src/physics/ClothSolverHelper.cpp:90-109), bend elements
(GenerateBendingElements, :4-88).  numpy only; every RNG seeded 20261104."""
import numpy as np

SEED = 20261104


def unique_edges(tri):
    tri = np.asarray(tri, np.int64).reshape(-1, 3)
    e = np.concatenate([tri[:, [0, 1]], tri[:, [1, 2]], tri[:, [2, 0]]])
    e.sort(axis=1)
    return np.unique(e, axis=0).astype(np.uint32)


def stretch_constraints(pos, edges):
    """(idx01, rest) — one distance constraint per mesh edge."""
    pos = np.asarray(pos, np.float32)
    d = pos[edges[:, 0]] - pos[edges[:, 1]]
    rest = np.sqrt((d * d).sum(1, dtype=np.float32)).astype(np.float32)
    return edges.astype(np.uint32), rest


def bend_constraints(pos, tri):
    """(idx0123, theta): [0],[1] = shared edge in the winding of the first
    (lower-index) face, [2] = its opposite vertex, [3] = the other face's
    opposite vertex; theta = acos(clamp(n1.n2)) with the reference's normals."""
    pos = np.asarray(pos, np.float32)
    tri = np.asarray(tri, np.int64).reshape(-1, 3)
    F = len(tri)
    a = np.concatenate([tri[:, 0], tri[:, 1], tri[:, 2]])
    b = np.concatenate([tri[:, 1], tri[:, 2], tri[:, 0]])
    opp = np.concatenate([tri[:, 2], tri[:, 0], tri[:, 1]])
    face = np.tile(np.arange(F), 3)
    lo, hi = np.minimum(a, b), np.maximum(a, b)
    nv = int(tri.max()) + 1
    key = lo * nv + hi
    order = np.lexsort((face, key))
    key, a, b, opp = key[order], a[order], b[order], opp[order]
    same = key[1:] == key[:-1]
    # manifold interior edges: exactly two half-edges share the key
    first = np.nonzero(same)[0]
    if len(first):
        run_prev = np.concatenate([[False], same[:-1]])[first]
        run_next = np.concatenate([same[1:], [False]])[first]
        first = first[~run_prev & ~run_next]
    i1, i2, i3, i4 = a[first], b[first], opp[first], opp[first + 1]
    v1, v2, v3, v4 = pos[i1], pos[i2], pos[i3], pos[i4]

    def nrm(v):
        return v / np.sqrt((v * v).sum(1, keepdims=True, dtype=np.float32))

    n1 = nrm(np.cross(v1 - v3, v2 - v3))
    n2 = nrm(np.cross(v2 - v4, v1 - v4))
    d = np.clip((n1 * n2).sum(1, dtype=np.float32), -1.0, 1.0)
    theta = np.arccos(d).astype(np.float32)
    idx = np.stack([i1, i2, i3, i4], 1).astype(np.uint32)
    return idx, theta


def grid_cloth(n, side=1.0, y=1.0, jitter=1e-4, seed=SEED):
    """n x n regular grid in the xz plane at height y (C1: n=224, C100k: n=317)."""
    rng = np.random.default_rng(seed)
    u = np.linspace(-side / 2, side / 2, n, dtype=np.float32)
    xx, zz = np.meshgrid(u, u, indexing="ij")
    pos = np.stack([xx.ravel(), np.full(n * n, y, np.float32), zz.ravel()], 1).astype(np.float32)
    pos += rng.uniform(-jitter, jitter, pos.shape).astype(np.float32)
    i, j = np.meshgrid(np.arange(n - 1), np.arange(n - 1), indexing="ij")
    v00 = (i * n + j).ravel()
    v01, v10, v11 = v00 + 1, v00 + n, v00 + n + 1
    tri = np.concatenate([np.stack([v00, v01, v11], 1), np.stack([v00, v11, v10], 1)]).astype(np.uint32)
    return pos, tri


def uv_sphere(r=0.3, stacks=64, slices=128, center=(0.0, 0.0, 0.0)):
    """UV sphere with outward normals: (stacks-1)*slices + 2 verts, 2*slices*(stacks-1) faces."""
    th = np.pi * np.arange(1, stacks, dtype=np.float64) / stacks
    ph = 2 * np.pi * np.arange(slices, dtype=np.float64) / slices
    T, P = np.meshgrid(th, ph, indexing="ij")
    ring = np.stack([np.sin(T) * np.cos(P), np.cos(T), np.sin(T) * np.sin(P)], -1).reshape(-1, 3)
    pos = np.concatenate([[[0, 1, 0]], ring, [[0, -1, 0]]]) * r + np.asarray(center)
    top, bot = 0, 1 + (stacks - 1) * slices
    tris = []
    s = np.arange(slices)
    s1 = (s + 1) % slices
    tris.append(np.stack([np.full(slices, top), 1 + s1, 1 + s], 1))
    for k in range(stacks - 2):
        a, b = 1 + k * slices, 1 + (k + 1) * slices
        tris.append(np.stack([a + s, a + s1, b + s1], 1))
        tris.append(np.stack([a + s, b + s1, b + s], 1))
    a = 1 + (stacks - 2) * slices
    tris.append(np.stack([np.full(slices, bot), a + s, a + s1], 1))
    tri = np.concatenate(tris).astype(np.uint32)
    pos = pos.astype(np.float32)
    # orient outward
    c = np.asarray(center, np.float32)
    fn = np.cross(pos[tri[:, 1]] - pos[tri[:, 0]], pos[tri[:, 2]] - pos[tri[:, 0]])
    flip = ((pos[tri].mean(1) - c) * fn).sum(1) < 0
    tri[flip] = tri[flip][:, [0, 2, 1]]
    return pos, tri


def lathe(rows, slices, center=(0.0, 0.0, 0.0)):
    """Closed surface of revolution about y: rows = [(y, radius), ...] from top to bottom
    (poles added above / below at the first / last row's y + / - its radius... see capsule)."""
    rows = np.asarray(rows, np.float64)
    ph = 2 * np.pi * np.arange(slices, dtype=np.float64) / slices
    ring = np.stack([np.outer(rows[:, 1], np.cos(ph)), np.repeat(rows[:, :1], slices, 1),
                     np.outer(rows[:, 1], np.sin(ph))], -1).reshape(-1, 3)
    return ring, len(rows)


def capsule(r, h, slices=32, hemi=8, cyl=20, center=(0.0, 0.0, 0.0)):
    """Capsule with its axis on y: cylinder of height h (cyl rows) capped by two hemispheres
    (hemi rows each), outward normals.  (2*hemi + cyl - 1) * slices + 2 vertices."""
    th = 0.5 * np.pi * np.arange(1, hemi + 1, dtype=np.float64) / hemi           # (0, pi/2]
    top = [(h / 2 + r * np.cos(t), r * np.sin(t)) for t in th]
    mid = [(h / 2 - h * k / cyl, r) for k in range(1, cyl)]
    bot = [(-h / 2 - r * np.cos(t), r * np.sin(t)) for t in th[::-1]]
    ring, nrows = lathe(top + mid + bot, slices)
    pos = np.concatenate([[[0, h / 2 + r, 0]], ring, [[0, -h / 2 - r, 0]]])
    s = np.arange(slices)
    s1 = (s + 1) % slices
    tris = [np.stack([np.zeros(slices, np.int64), 1 + s1, 1 + s], 1)]
    for k in range(nrows - 1):
        a, b = 1 + k * slices, 1 + (k + 1) * slices
        tris.append(np.stack([a + s, a + s1, b + s1], 1))
        tris.append(np.stack([a + s, b + s1, b + s], 1))
    a = 1 + (nrows - 1) * slices
    tris.append(np.stack([np.full(slices, a + slices), a + s, a + s1], 1))
    tri = np.concatenate(tris).astype(np.uint32)
    pos = pos.astype(np.float32)
    fn = np.cross(pos[tri[:, 1]] - pos[tri[:, 0]], pos[tri[:, 2]] - pos[tri[:, 0]])
    flip = (pos[tri].mean(1) * fn).sum(1) < 0
    tri[flip] = tri[flip][:, [0, 2, 1]]
    return pos + np.asarray(center, np.float32), tri


def tube_cloth(r, y_top, y_bot, n_around, n_high, center_xz=(0.0, 0.0), jitter=1e-4, rng=None, r_bot=None):
    """Open cylinder (r_bot: cone frustum) of cloth about a vertical axis (C2 / C3 garments); row 0 is the top ring."""
    ph = 2 * np.pi * np.arange(n_around, dtype=np.float64) / n_around
    ys = np.linspace(y_top, y_bot, n_high)
    rr = np.repeat(np.linspace(r, r if r_bot is None else r_bot, n_high), n_around)
    pos = np.stack([center_xz[0] + rr * np.tile(np.cos(ph), n_high), np.repeat(ys, n_around),
                    center_xz[1] + rr * np.tile(np.sin(ph), n_high)], 1).astype(np.float32)
    if rng is not None and jitter > 0:
        pos += rng.uniform(-jitter, jitter, pos.shape).astype(np.float32)
    i, j = np.meshgrid(np.arange(n_high - 1), np.arange(n_around), indexing="ij")
    v00 = (i * n_around + j).ravel()
    v01 = (i * n_around + (j + 1) % n_around).ravel()
    v10, v11 = v00 + n_around, v01 + n_around
    tri = np.concatenate([np.stack([v00, v01, v11], 1), np.stack([v00, v11, v10], 1)]).astype(np.uint32)
    return pos, tri


class Scene:
    """Everything AddCloth/AddObstacle + the constraint generators hand to the solver."""

    def __init__(self, cloth_pos, cloth_tri, obs_pos=None, obs_tri=None, fixed=()):
        self.cloth_pos = np.ascontiguousarray(cloth_pos, np.float32)
        self.cloth_tri = np.ascontiguousarray(cloth_tri, np.uint32)
        self.cloth_edges = unique_edges(self.cloth_tri)
        self.inv_mass = np.ones(len(self.cloth_pos), np.float32)
        self.inv_mass[list(fixed)] = 0.0
        self.stretch_idx, self.stretch_rest = stretch_constraints(self.cloth_pos, self.cloth_edges)
        self.bend_idx, self.bend_theta = bend_constraints(self.cloth_pos, self.cloth_tri)
        self.obs_pos = None if obs_pos is None else np.ascontiguousarray(obs_pos, np.float32)
        self.obs_tri = None if obs_tri is None else np.ascontiguousarray(obs_tri, np.uint32)
        self.obs_edges = None if obs_tri is None else unique_edges(self.obs_tri)

    def particle_diameter(self, scale=1.3):  # PARTICLE_DIAMETER of examples/smpl/simulation_conf_smpl.json
        # ClothObjectUniversal::Start, impl/universal/ClothObject.hpp:36-37
        t = self.cloth_tri[0]
        return float(np.linalg.norm(self.cloth_pos[t[0]] - self.cloth_pos[t[1]])) * scale


def drop_scene(n=224, sphere=(64, 128), y=None, seed=SEED, fixed=()):
    """C1 / C100k: n x n cloth dropped on a UV sphere r = 0.3 m at the origin."""
    r = 0.3
    if y is None:
        y = r + 0.02
    cp, ct = grid_cloth(n, 1.0, y, 1e-4, seed)
    op, ot = uv_sphere(r, sphere[0], sphere[1])
    return Scene(cp, ct, op, ot, fixed)


def layered_scene(n=24, layers=2, gap=4e-3, sphere=(12, 24), y=None, side=0.5, seed=SEED, fixed=(), wave=0.0):
    """Quality-mode test scene: `layers` n x n sheets stacked `gap` apart above a UV sphere, as ONE cloth mesh
    (the reference merges garments into one PhysicsMesh).  With a downward initial velocity the sheets hit the
    sphere and each other inside one frame, which exercises the VF / EE CCD of both kinds of pairs.  wave > 0
    ripples every odd layer by wave * sin(6 pi x / side), so neighbouring layers come within `gap - wave` of each
    other in stripes (proximity pairs for the repulsion stages without any motion)."""
    r = 0.3
    if y is None:
        y = r + 0.01
    pos, tri = [], []
    for k in range(layers):
        p, t = grid_cloth(n, side, y + k * gap, 1e-4, seed + k)
        if wave and k % 2:
            p[:, 1] += np.float32(wave) * np.sin(6.0 * np.pi * p[:, 0] / side).astype(np.float32)
        tri.append(t + len(p) * k)
        pos.append(p)
    op, ot = uv_sphere(r, sphere[0], sphere[1])
    return Scene(np.concatenate(pos), np.concatenate(tri), op, ot, fixed)


def corner_tethers(scene, pins):
    """Geodesic long-range attachments (ClothObjectBase::GenerateLongRangeConstraints,
    pipeline/base/ClothObjectBase.cpp:70-94) with ONE slot per vertex: every free vertex is tethered to the
    nearest of the pinned vertices `pins`, rest length = straight-line distance at setup (on a flat sheet that is
    the geodesic distance).  Returns (src, tgt, rest)."""
    pins = np.asarray(pins, np.int64)
    pos = scene.cloth_pos
    d = np.sqrt(((pos[:, None, :] - pos[pins][None, :, :]) ** 2).sum(-1, dtype=np.float32))
    k = d.argmin(1)
    tgt = np.setdiff1d(np.arange(len(pos)), pins)
    return pins[k[tgt]].astype(np.uint32), tgt.astype(np.uint32), d[tgt, k[tgt]].astype(np.float32)


def swift_params(**kw):
    """xrt_default_params (include/xrt_b200.h; the reference's Swift defaults of examples/smpl/simulation_conf_smpl.json
    + Common.hpp HOST_INITs) as a plain namespace, for callers that must not load the product library (the reference
    arm of bench.py).  tests/test_abi_cpu.py checks it against the library's own defaults."""
    import types
    p = types.SimpleNamespace(
        num_substeps=4, num_iterations=10, num_collision_passes=10, max_num_neighbors=64, gravity=[0.0, -9.8, 0.0],
        damping=0.98, relaxation_factor=0.9, long_range_stretchiness=1.1, sdf_friction=0.1, bvh_tolerance=0.003,
        stretch_compliance=0.0, bend_compliance=100.0, particle_diameter=0.0, hash_cell_size=1.5, inter_leaved_hash=3,
        enable_self_collision=1, solve_bending=1, geodesic_lra=1, frame_dt=1.0 / 60.0, morton_fast_div=1,
        persist_cluster=8, device=0)
    for k, v in kw.items():
        setattr(p, k, v)
    return p


def sphere_motion(obs_pos, frame, amp=0.05, hz=1.0, fps=60.0):
    """Synthetic body motion: the obstacle bobs vertically and breathes."""
    t = frame / fps
    s = 1.0 + 0.03 * np.sin(2 * np.pi * hz * t)
    out = obs_pos * np.float32(s)
    out[:, 1] += np.float32(amp * np.sin(2 * np.pi * hz * t))
    return out.astype(np.float32)


# ---- C2: capsule body with a procedural pose sequence, shirt + trousers ----------
# (SMPL's own topology is a licensed download and not in the archive, so
# the body is a union of six capsules of SMPL's size: ~7k vertices / ~14k faces.)
_BODY_PARTS = (
    # name, radius, cylinder height, centre at rest, joint the part swings about, slices, hemi, cyl rows
    ("torso", 0.150, 0.43, (0.00, 1.235, 0.0), None, 48, 12, 20),
    ("head", 0.095, 0.05, (0.00, 1.72, 0.0), None, 32, 10, 2),
    ("arm_l", 0.045, 0.52, (0.235, 1.15, 0.0), (0.235, 1.41, 0.0), 20, 5, 30),
    ("arm_r", 0.045, 0.52, (-0.235, 1.15, 0.0), (-0.235, 1.41, 0.0), 20, 5, 30),
    ("leg_l", 0.070, 0.74, (0.085, 0.49, 0.0), (0.085, 0.90, 0.0), 32, 8, 42),
    ("leg_r", 0.070, 0.74, (-0.085, 0.49, 0.0), (-0.085, 0.90, 0.0), 32, 8, 42),
)


class Body:
    """Obstacle mesh of the C2 scene and its pose function."""

    def __init__(self):
        pos, tri, self.parts = [], [], []
        base = 0
        for name, r, h, c, joint, sl, hemi, cyl in _BODY_PARTS:
            p, t = capsule(r, h, sl, hemi, cyl, center=c)
            pos.append(p)
            tri.append(t + base)
            self.parts.append((name, base, base + len(p), joint))
            base += len(p)
        self.rest = np.concatenate(pos).astype(np.float32)
        self.tri = np.concatenate(tri).astype(np.uint32)

    def pose(self, frame, amp_deg=30.0, hz=1.0, fps=60.0, phase=0.0):
        """Sinusoidal shoulder / hip swing about the x axis: arms and legs in opposite phase."""
        t = frame / fps
        a = np.deg2rad(amp_deg) * np.sin(2 * np.pi * hz * t + phase)
        out = self.rest.copy()
        for name, lo, hi, joint in self.parts:
            if joint is None:
                continue
            ang = a if name in ("arm_l", "leg_r") else -a
            c, s = np.float32(np.cos(ang)), np.float32(np.sin(ang))
            j = np.asarray(joint, np.float32)
            d = self.rest[lo:hi] - j
            out[lo:hi, 1] = j[1] + c * d[:, 1] - s * d[:, 2]
            out[lo:hi, 2] = j[2] + s * d[:, 1] + c * d[:, 2]
        return out


def body_rig(body, blend=0.04):
    """Skinning rig of the capsule body for the device-side animation producer: joint 0 = torso (root, at its centre),
    1 = head, 2..5 = the shoulder / hip joints of the limbs, all children of the root.  Weights are one-hot on the
    part except within `blend` metres below a limb's joint plane, where the limb blends linearly into the torso (what
    a skinned body has around a shoulder); blend = 0 reproduces Body.pose exactly.
    Returns dict(rest, w_outer, w_inner, w_val, joints, parent)."""
    joints, parent = [], []
    for name, r, h, c, joint, *_ in _BODY_PARTS:
        joints.append(c if joint is None else joint)
        parent.append(-1 if not parent else 0)
    joints = np.asarray(joints, np.float32)
    outer, inner, val = [0], [], []
    for k, (name, lo, hi, joint) in enumerate(body.parts):
        for v in range(lo, hi):
            w = 1.0
            if joint is not None and blend > 0.0:
                w = float(np.clip((joint[1] - body.rest[v, 1]) / blend, 0.0, 1.0)) if body.rest[v, 1] < joint[1] else 0.0
            if w >= 1.0 or k == 0:
                inner.append(k); val.append(1.0)
            elif w <= 0.0:
                inner.append(0); val.append(1.0)
            else:
                inner += [0, k]; val += [1.0 - w, w]
            outer.append(len(inner))
    return dict(rest=body.rest.copy(), w_outer=np.asarray(outer, np.uint32), w_inner=np.asarray(inner, np.uint32),
                w_val=np.asarray(val, np.float32), joints=joints, parent=np.asarray(parent, np.int32))


def body_pose_params(body, frame, amp_deg=30.0, hz=1.0, fps=60.0, phase=0.0):
    """Angle-axis pose (J x 3) of Body.pose's motion: limbs swing about the x axis, arms and legs in opposite phase."""
    t = frame / fps
    a = np.deg2rad(amp_deg) * np.sin(2 * np.pi * hz * t + phase)
    pose = np.zeros((len(body.parts), 3), np.float32)
    for k, (name, lo, hi, joint) in enumerate(body.parts):
        if joint is not None:
            pose[k, 0] = a if name in ("arm_l", "leg_r") else -a
    return pose


def body_scene(scale=1.0, seed=SEED, gap=0.012):
    """C2: shirt (~18k verts) around the torso and two trouser legs (~6k verts each) around the
    legs, `gap` metres off the body.  The top ring of each garment is tethered to the body with
    binding constraints (what the reference's SMPL pipeline does with its bind_* lists,
    BindingConstraint.cu:11-55): anchor = nearest body vertex, rest length = distance at setup.
    scale < 1 coarsens the garments for the oracle-sized tests.
    Returns (Scene, Body, bindings) with bindings = (cloth_idx, tri, bary, stiffness, distance)."""
    rng = np.random.default_rng(seed)
    body = Body()
    k = float(np.sqrt(scale))
    na_s, nh_s = max(24, int(190 * k)), max(12, int(95 * k))
    na_p, nh_p = max(16, int(80 * k)), max(12, int(75 * k))
    sp, st = tube_cloth(0.150 + gap, 1.44, 1.04, na_s, nh_s, rng=rng)
    lp, lt = tube_cloth(0.070 + gap, 0.84, 0.22, na_p, nh_p, center_xz=(0.085, 0.0), rng=rng)
    rp, rt = tube_cloth(0.070 + gap, 0.84, 0.22, na_p, nh_p, center_xz=(-0.085, 0.0), rng=rng)
    pos = np.concatenate([sp, lp, rp])
    tri = np.concatenate([st, lt + len(sp), rt + len(sp) + len(lp)])
    o1, o2 = len(sp), len(sp) + len(lp)
    ring = np.array(list(range(na_s)) + list(range(o1, o1 + na_p)) + list(range(o2, o2 + na_p)), np.uint32)
    sc = Scene(pos, tri, body.rest, body.tri)
    # anchor of a ring vertex: the nearest body vertex, as corner 0 of one of its triangles
    d = ((pos[ring][:, None, :] - body.rest[None, :, :]) ** 2).sum(-1)
    near = d.argmin(1)
    first_tri = np.full(len(body.rest), -1, np.int64)
    for c in range(3):
        first_tri[body.tri[::-1, c]] = np.arange(len(body.tri))[::-1]
    t = body.tri[first_tri[near]]
    corner = (t == near[:, None]).argmax(1)
    bary = np.zeros((len(ring), 3), np.float32)
    bary[np.arange(len(ring)), corner] = 1.0
    bindings = (ring, t.astype(np.uint32), bary, np.ones(len(ring), np.float32), np.zeros(len(ring), np.float32))
    return sc, body, bindings


def _ring_bindings(pos, ring, body):
    """tether every vertex of `ring` to the nearest body vertex (corner 0 of one of its triangles)"""
    d = ((pos[ring][:, None, :] - body.rest[None, :, :]) ** 2).sum(-1)
    near = d.argmin(1)
    first_tri = np.full(len(body.rest), -1, np.int64)
    for c in range(3):
        first_tri[body.tri[::-1, c]] = np.arange(len(body.tri))[::-1]
    t = body.tri[first_tri[near]]
    corner = (t == near[:, None]).argmax(1)
    bary = np.zeros((len(ring), 3), np.float32)
    bary[np.arange(len(ring)), corner] = 1.0
    return (ring, t.astype(np.uint32), bary, np.ones(len(ring), np.float32), np.zeros(len(ring), np.float32))


def dress_scene(scale=1.0, seed=SEED, gap=0.012):
    """C3 (BASELINE.json configs[2]): a cone-frustum dress, 224 x 224 grid wrapped around the C2 body from the chest
    (r = body + gap) to a 0.42 m hem at knee height, 50 176 vertices at scale 1; the top ring is tethered to the body.
    Motion: Body.pose(frame, amp_deg=25, hz=1.2), the walking swing SURVEY 8(d) names.  Quality mode.
    Returns (Scene, Body, bindings)."""
    rng = np.random.default_rng(seed)
    body = Body()
    k = float(np.sqrt(scale))
    na, nh = max(24, int(224 * k)), max(12, int(224 * k))
    pos, tri = tube_cloth(0.150 + gap, 1.40, 0.45, na, nh, rng=rng, r_bot=0.42)
    sc = Scene(pos, tri, body.rest, body.tri)
    return sc, body, _ring_bindings(pos, np.arange(na, dtype=np.uint32), body)


def layered_tubes_scene(n=224, layers=4, gap=2e-3, sphere=(64, 128), seed=SEED, r=0.2):
    """C4 (BASELINE.json configs[3]): `layers` concentric vertical tubes of n x n vertices, `gap` apart, hanging just
    above the UV sphere (r = 0.3 m) they fall onto: 4 x 224^2 = 200 704 vertices, every layer within 2 mm of the
    next -- the self-contact stress case for BVH build / refit, pair lists and the CCD narrow phase."""
    rng = np.random.default_rng(seed)
    pos, tri, base = [], [], 0
    for k in range(layers):
        p, t = tube_cloth(r + k * gap, 0.85, 0.35, n, n, rng=rng)
        pos.append(p)
        tri.append(t + base)
        base += len(p)
    op, ot = uv_sphere(0.3, sphere[0], sphere[1])
    return Scene(np.concatenate(pos), np.concatenate(tri), op, ot)
