"""Host-side mirror of the reference's solver component for the Swift path.

Method names follow ``ClothSolverBase`` / ``ClothSolverUniversal``
(xrtailor/include/xrtailor/pipeline/base/ClothSolverBase.hpp:55-189,
xrtailor/src/pipeline/impl/universal/ClothSolver.cpp:189-350): ``AddCloth``,
``AddObstacle``, ``AddGeodesic``, ``Start``, ``SimulateSwift``, ``SaveCache``.
All numerics happen in libxrt_b200.so; this file only marshals numpy arrays
across the C ABI.
"""
import ctypes as C

import numpy as np

from ._abi import (XrtRigDesc, XrtCollisionStats, XrtConstraints, XrtError, XrtMeshDesc, XrtParams, XrtQualityParams, XrtStats, default_params,
                   default_quality_params, lib)

DBG = dict(MORTON=0, SORTED_OBJECT=1, NODES=2, AABBS=3, ESCAPE_INT=4, ESCAPE_EXT=5, VF_CONTACTS=6, RT_CONTACTS=7,
           EE_CONTACTS=8, PREDICTED=9, VELOCITY=10, COLORS_STRETCH=11, COLORS_BEND=12, NEIGHBORS=13, RTRI=14,
           OBSTACLE_PRED=15, STATS=16, PARTICLE_HASH=17, VF_CAND_COUNT=18, RT_CAND_COUNT=19, EE_CAND_COUNT=20, PERSIST_PROF=21)

_f32p = C.POINTER(C.c_float)
_u32p = C.POINTER(C.c_uint32)
_i32p = C.POINTER(C.c_int32)


def _f32(a):
    return np.ascontiguousarray(a, dtype=np.float32)


def _u32(a):
    return np.ascontiguousarray(a, dtype=np.uint32)


def _pf(a):
    return a.ctypes.data_as(_f32p) if a is not None else None


def _pu(a):
    return a.ctypes.data_as(_u32p) if a is not None else None


class ClothSolver:
    """One solver handle = one CUDA stream on one device."""

    def __init__(self, params: XrtParams = None, **overrides):
        self.params = params if params is not None else default_params()
        for k, v in overrides.items():
            if k == "gravity":
                for i in range(3):
                    self.params.gravity[i] = float(v[i])
            else:
                setattr(self.params, k, v)
        self._h = C.c_void_p()
        rc = lib.xrt_create(C.byref(self.params), C.byref(self._h))
        if rc:
            raise XrtError(rc, "xrt_create failed (no usable CUDA device? there is no CPU fallback)")
        self.n_cloth_verts = self.n_cloth_edges = self.n_obstacle_verts = self.n_obstacle_faces = 0
        self.solver_time = 0.0
        # host frame cache like ClothSolverBase::h_cloth_positions_cache (ClothSolverBase.hpp:107-112)
        self.h_cloth_positions_cache, self.h_cloth_normals_cache = [], []

    # -- lifetime ---------------------------------------------------------
    def OnDestroy(self):
        if self._h:
            lib.xrt_destroy(self._h)
            self._h = C.c_void_p()

    close = OnDestroy

    def __del__(self):
        try:
            self.OnDestroy()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.OnDestroy()

    def _ck(self, rc):
        if rc:
            raise XrtError(rc, lib.xrt_last_error(self._h).decode(errors="replace"))

    # -- setup ------------------------------------------------------------
    def _mesh(self, positions, indices, edges, inv_mass):
        pos, idx, edg = _f32(positions).reshape(-1, 3), _u32(indices).reshape(-1, 3), _u32(edges).reshape(-1, 2)
        inv = _f32(inv_mass) if inv_mass is not None else None
        d = XrtMeshDesc(_pf(pos), _pu(idx), _pu(edg), _pf(inv), len(pos), len(idx), len(edg))
        return d, (pos, idx, edg, inv)

    def AddCloth(self, positions, indices, edges, inv_mass=None):
        d, keep = self._mesh(positions, indices, edges, inv_mass)
        off = C.c_int32()
        self._ck(lib.xrt_add_cloth(self._h, C.byref(d), C.byref(off)))
        self.n_cloth_verts, self.n_cloth_edges = d.n_verts, d.n_edges
        return off.value

    def AddObstacle(self, positions, indices, edges=None, inv_mass=None):
        if edges is None:
            edges = np.zeros((0, 2), np.uint32)
        d, keep = self._mesh(positions, indices, edges, inv_mass)
        off = C.c_int32()
        self._ck(lib.xrt_add_obstacle(self._h, C.byref(d), C.byref(off)))
        self.n_obstacle_verts, self.n_obstacle_faces = d.n_verts, d.n_faces
        return off.value

    def SetStretch(self, idx01, rest):
        idx, r = _u32(idx01).reshape(-1, 2), _f32(rest)
        self._ck(lib.xrt_set_stretch(self._h, _pu(idx), _pf(r), len(r)))

    def SetBend(self, idx0123, theta):
        idx, t = _u32(idx0123).reshape(-1, 4), _f32(theta)
        self._ck(lib.xrt_set_bend(self._h, _pu(idx), _pf(t), len(t)))

    def AddGeodesic(self, src, tgt, rest):
        s, t, r = _u32(src), _u32(tgt), _f32(rest)
        self._ck(lib.xrt_add_geodesic(self._h, _pu(s), _pu(t), _pf(r), len(r)))

    def AddBinding(self, cloth_idx, tri, bary, stiffness, distance):
        """BindingConstraint::Add (BindingConstraint.cu:11-20) with the SkinParam anchor passed in."""
        v, t, b = _u32(cloth_idx), _u32(tri).reshape(-1, 3), _f32(bary).reshape(-1, 3)
        k, d = _f32(stiffness), _f32(distance)
        self._ck(lib.xrt_add_binding(self._h, _pu(v), _pu(t), _pf(b), _pf(k), _pf(d), len(v)))

    def SetColors(self, kind, colors):
        c = _u32(colors)
        self._ck(lib.xrt_set_colors(self._h, int(kind), _pu(c), len(c)))

    def Start(self):
        self._ck(lib.xrt_finalize_setup(self._h))

    # -- per frame --------------------------------------------------------
    def UpdateObstacleAnimationFrame(self, target_xyz):
        t = _f32(target_xyz).reshape(-1, 3)
        self._ck(lib.xrt_set_obstacle_target(self._h, _pf(t), len(t)))

    def SetObstacleSequence(self, frames_xyz):
        t = _f32(frames_xyz).reshape(-1, self.n_obstacle_verts, 3)
        self._ck(lib.xrt_set_obstacle_sequence(self._h, _pf(t), len(t)))

    def SetObstacleRig(self, rest, w_outer, w_inner, w_val, joints, parent, model_matrix=None, pose_blend=None):
        """Body animation on the device: the skinning rig of the obstacle (smplx::Body, runtime/rag_doll/smpl)."""
        keep = dict(rest=_f32(rest).reshape(-1, 3), wo=_u32(w_outer), wi=_u32(w_inner), wv=_f32(w_val),
                    j=_f32(joints).reshape(-1, 3), par=np.ascontiguousarray(parent, np.int32),
                    M=None if model_matrix is None else _f32(model_matrix).reshape(16),
                    pb=None if pose_blend is None else _f32(pose_blend))
        d = XrtRigDesc()
        d.rest_positions, d.weight_outer, d.weight_inner, d.weight_values = _pf(keep["rest"]), _pu(keep["wo"]), _pu(keep["wi"]), _pf(keep["wv"])
        d.joints, d.parent = _pf(keep["j"]), keep["par"].ctypes.data_as(_i32p)
        d.model_matrix = _pf(keep["M"]) if keep["M"] is not None else None
        d.pose_blend = _pf(keep["pb"]) if keep["pb"] is not None else None
        d.n_verts, d.n_joints = len(keep["rest"]), len(keep["j"])
        self._n_joints = d.n_joints
        self._ck(lib.xrt_set_obstacle_rig(self._h, C.byref(d)))

    def SetObstaclePose(self, axis_angle, trans=None):
        """Sequence::set_pose + Body::Update + UpdateObstaclePredicted for one frame (impl/smpl/ClothSolver.cpp:244-249)."""
        a = _f32(axis_angle).reshape(self._n_joints, 3)
        t = None if trans is None else _f32(trans).reshape(3)
        self._ck(lib.xrt_set_obstacle_pose(self._h, _pf(a), None if t is None else _pf(t)))

    def SetObstaclePoseSequence(self, axis_angle, trans=None):
        a = _f32(axis_angle).reshape(-1, self._n_joints, 3)
        t = None if trans is None else _f32(trans).reshape(len(a), 3)
        self._ck(lib.xrt_set_obstacle_pose_sequence(self._h, _pf(a), None if t is None else _pf(t), len(a)))

    def ObstacleJointTransforms(self):
        out = np.zeros((self._n_joints, 12), np.float32)
        self._ck(lib.xrt_get_obstacle_joint_transforms(self._h, _pf(out)))
        return out

    def ObstacleTarget(self):
        out = np.zeros((self.n_obstacle_verts, 3), np.float32)
        self._ck(lib.xrt_get_obstacle_target(self._h, _pf(out)))
        return out

    def ProfileKernel(self, which, reps=20):
        ms, b = C.c_float(), C.c_double()
        self._ck(lib.xrt_profile_kernel(self._h, int(which), int(reps), C.byref(ms), C.byref(b)))
        return ms.value, b.value

    def SimulateSwift(self):
        ms = C.c_float()
        self._ck(lib.xrt_step_frame(self._h, C.byref(ms)))
        self.solver_time = ms.value
        return ms.value

    def SimulateSwiftAsync(self):
        self._ck(lib.xrt_step_frame_async(self._h))

    def Sync(self):
        self._ck(lib.xrt_sync(self._h))

    def GetPositions(self, which=0, normals=False):
        n = self.n_cloth_verts if which == 0 else self.n_obstacle_verts
        xyz = np.empty((n, 3), np.float32)
        nrm = np.empty((n, 3), np.float32) if normals else None
        self._ck(lib.xrt_get_positions(self._h, which, _pf(xyz), _pf(nrm)))
        return (xyz, nrm) if normals else xyz

    def SaveCache(self):
        xyz, nrm = self.GetPositions(0, normals=True)
        self.h_cloth_positions_cache.append(xyz)
        self.h_cloth_normals_cache.append(nrm)

    def SetClothState(self, xyz=None, vel=None):
        x = _f32(xyz).reshape(-1, 3) if xyz is not None else None
        v = _f32(vel).reshape(-1, 3) if vel is not None else None
        self._ck(lib.xrt_set_cloth_state(self._h, _pf(x), _pf(v)))

    def SetClothPredicted(self, xyz):
        x = _f32(xyz).reshape(-1, 3)
        self._ck(lib.xrt_set_cloth_predicted(self._h, _pf(x)))

    # -- Quality mode (ClothSolverBase::SimulateQuality, ClothSolverBase.cpp:263-362) ---------------
    def SetQuality(self, q=None, strain_colors=None, **overrides):
        """Before Start: the handle steps Quality frames.  q: XrtQualityParams or any object with its fields."""
        qp = default_quality_params()
        if q is not None:
            for name, _ in XrtQualityParams._fields_:
                if hasattr(q, name):
                    setattr(qp, name, getattr(q, name))
        for k, v in overrides.items():
            setattr(qp, k, v)
        self.quality = qp
        c = _u32(strain_colors) if strain_colors is not None else None
        self._ck(lib.xrt_set_quality(self._h, C.byref(qp), _pu(c)))

    def SimulateQuality(self):
        return self.SimulateSwift()  # xrt_step_frame runs the mode the handle was set up for

    def QualityBeginFrame(self):
        self._ck(lib.xrt_quality_begin_frame(self._h))

    def ImminentGenerate(self):
        n = C.c_size_t(0)
        self._ck(lib.xrt_quality_imminent_generate(self._h, None, None, 0, C.byref(n)))
        nodes, bary = np.zeros((n.value, 4), np.uint32), np.zeros((n.value, 4), np.float32)
        if n.value:
            self._ck(lib.xrt_quality_imminent_generate(self._h, _pu(nodes), _pf(bary), n.value, C.byref(n)))
        return nodes, bary

    def ImminentSolve(self):
        self._ck(lib.xrt_quality_imminent_solve(self._h))

    def RepUpdatePairs(self):
        n = C.c_size_t(0)
        self._ck(lib.xrt_quality_rep_update_pairs(self._h, None, 0, C.byref(n)))
        pairs = np.zeros((n.value, 2), np.uint32)
        if n.value:
            # the second call refreshes the same pairs (the cloth tree only grows by positions it already holds)
            self._ck(lib.xrt_quality_rep_update_pairs(self._h, _pu(pairs), n.value, C.byref(n)))
        return pairs

    def RepSolve(self):
        self._ck(lib.xrt_quality_rep_solve(self._h))

    def ObstacleBvhUpdate(self):
        self._ck(lib.xrt_quality_obstacle_bvh_update(self._h))

    def QualityIteration(self, iteration):
        self._ck(lib.xrt_quality_iteration(self._h, int(iteration)))

    def FindImpacts(self, thickness):
        """(nodes[n,4], twn[n,9] = t, w[4], n[3], is_vf, n_pairs, n_candidates)"""
        n, npairs, ncand = C.c_size_t(0), C.c_size_t(0), C.c_size_t(0)
        self._ck(lib.xrt_quality_find_impacts(self._h, thickness, None, None, 0, C.byref(n), C.byref(npairs), C.byref(ncand)))
        nodes, twn = np.zeros((n.value, 4), np.uint32), np.zeros((n.value, 9), np.float32)
        if n.value:
            self._ck(lib.xrt_quality_find_impacts(self._h, thickness, _pu(nodes), _pf(twn), n.value, C.byref(n),
                                                  C.byref(npairs), C.byref(ncand)))
        return nodes, twn, int(npairs.value), int(ncand.value)

    def IndependentImpacts(self, nodes, twn, deform=0):
        nodes, twn = _u32(nodes).reshape(-1, 4), _f32(twn).reshape(-1, 9)
        on, ot = np.zeros_like(nodes), np.zeros_like(twn)
        m = C.c_size_t(0)
        self._ck(lib.xrt_quality_independent_impacts(self._h, _pu(nodes), _pf(twn), len(nodes), int(deform), _pu(on), _pf(ot),
                                                     C.byref(m)))
        return on[:m.value].copy(), ot[:m.value].copy()

    def IsolateZones(self, nodes):
        nodes = _u32(nodes).reshape(-1, 4)
        col = np.zeros(len(nodes), np.int32)
        nz = C.c_int32(0)
        self._ck(lib.xrt_quality_isolate_zones(self._h, _pu(nodes), len(nodes), col.ctypes.data_as(_i32p), C.byref(nz)))
        return col, int(nz.value)

    def OptimizeZones(self, nodes, twn, colors, deform, thickness, obstacle_mass):
        nodes, twn = _u32(nodes).reshape(-1, 4), _f32(twn).reshape(-1, 9)
        colors = np.ascontiguousarray(colors, np.int32)
        st = np.zeros(3, np.int32)
        nz = int(colors.max()) + 1 if len(colors) else 0
        self._ck(lib.xrt_quality_optimize_zones(self._h, _pu(nodes), _pf(twn), colors.ctypes.data_as(_i32p), len(nodes), nz,
                                                int(deform), float(thickness), float(obstacle_mass), st.ctypes.data_as(_i32p)))
        return dict(outer=int(st[0]), line_search=int(st[1]), converged=int(st[2]))

    def CollisionStep(self):
        st = XrtCollisionStats()
        self._ck(lib.xrt_quality_collision_step(self._h, C.byref(st)))
        return {name: int(getattr(st, name)) for name, _ in XrtCollisionStats._fields_}

    def QualityProfile(self, what, reps=5):
        ms, n = C.c_float(), C.c_double()
        self._ck(lib.xrt_quality_profile(self._h, int(what), int(reps), C.byref(ms), C.byref(n)))
        return ms.value, n.value

    def ClothRTri(self, n_faces):
        out = np.zeros(n_faces, np.uint32)
        self._ck(lib.xrt_quality_cloth_rtri(self._h, _pu(out), n_faces))
        return out

    def ObstaclePredicted(self):
        out = np.zeros((self.n_obstacle_verts, 3), np.float32)
        self._ck(lib.xrt_quality_get_obstacle_predicted(self._h, _pf(out)))
        return out

    # -- building blocks (parity tests) ------------------------------------
    def BvhBuildObstacle(self):
        self._ck(lib.xrt_bvh_build_obstacle(self._h))

    def BvhRefitObstacle(self, step):
        self._ck(lib.xrt_bvh_refit_obstacle(self._h, int(step)))

    def GenerateContacts(self):
        self._ck(lib.xrt_generate_contacts(self._h))

    def HashParticles(self):
        self._ck(lib.xrt_hash_particles(self._h))

    def PredictPositions(self):
        self._ck(lib.xrt_predict(self._h))

    def SolveIteration(self, iteration):
        self._ck(lib.xrt_solve_iteration(self._h, int(iteration)))

    def Finalize(self):
        self._ck(lib.xrt_finalize(self._h))

    # -- debug ------------------------------------------------------------
    def debug(self, what, dtype=np.uint8):
        code = DBG[what] if isinstance(what, str) else int(what)
        n = C.c_size_t(0)
        self._ck(lib.xrt_debug_get(self._h, code, None, C.byref(n)))
        buf = np.empty(n.value, np.uint8)
        self._ck(lib.xrt_debug_get(self._h, code, buf.ctypes.data_as(C.c_void_p), C.byref(n)))
        return buf.view(dtype)

    def stats(self):
        raw = self.debug("STATS")
        st = XrtStats.from_buffer_copy(raw.tobytes())
        return {f: getattr(st, f) for f, _ in XrtStats._fields_}

    def bvh(self):
        """Obstacle BVH as the reference's host accessors return it (BVH.cu:579-602)."""
        return dict(
            morton=self.debug("MORTON", np.uint32), object=self.debug("SORTED_OBJECT", np.uint32),
            nodes=self.debug("NODES", np.uint32).reshape(-1, 4), aabb=self.debug("AABBS", np.float32).reshape(-1, 6),
            esc_int=self.debug("ESCAPE_INT", np.int32), esc_ext=self.debug("ESCAPE_EXT", np.int32))

    def contacts(self):
        vf = self.debug("VF_CONTACTS", np.dtype([("index", np.int32), ("q", np.float32, 3)]))
        rt = self.debug("RT_CONTACTS", np.dtype([("index", np.int32), ("q", np.float32, 3)]))
        ee = self.debug("EE_CONTACTS", np.float32).reshape(-1, 12)  # dist, p2, q2, s, t, fn
        return vf, rt, ee

    def candidates(self, kind):
        n = self.n_cloth_edges if kind == 2 else self.n_cloth_verts
        off = np.zeros(n + 1, np.uint32)
        cnt = C.c_size_t(0)
        self._ck(lib.xrt_debug_candidates(self._h, kind, _pu(off), None, C.byref(cnt)))
        items = np.zeros(max(cnt.value, 1), np.uint32)
        cnt2 = C.c_size_t(cnt.value)
        self._ck(lib.xrt_debug_candidates(self._h, kind, _pu(off), _pu(items), C.byref(cnt2)))
        return off, items[:cnt.value]


def batch_run(solvers, frames):
    """Run `frames` frames on every solver concurrently; returns device ms."""
    arr = (C.c_void_p * len(solvers))(*[s._h for s in solvers])
    ms = C.c_float()
    rc = lib.xrt_batch_run(arr, len(solvers), int(frames), C.byref(ms))
    if rc:
        raise XrtError(rc, lib.xrt_last_error(solvers[0]._h).decode(errors="replace"))
    return ms.value


class Batch:
    """Shared-launch batch over solvers with the same persist_cluster >= 1: every substep of all of them is one
    kernel launch (xrt_batch_*).  The solvers stay usable on their own between frames.  groups > 1: that many
    independently stepping sets, so one set's BVH builds / hashes overlap the other sets' substep kernels."""

    def __init__(self, solvers, groups=1, workers=3):
        self.solvers = list(solvers)
        arr = (C.c_void_p * len(self.solvers))(*[s._h for s in self.solvers])
        h = C.c_void_p()
        rc = lib.xrt_batch_create(arr, len(self.solvers), int(groups), int(workers), C.byref(h))
        if rc:
            raise XrtError(rc, lib.xrt_last_error(self.solvers[0]._h).decode(errors="replace"))
        self._h = h

    def _ck(self, rc):
        if rc:
            raise XrtError(rc, lib.xrt_last_error(self.solvers[0]._h).decode(errors="replace"))

    def SimulateSwiftAsync(self):
        self._ck(lib.xrt_batch_step_frame_async(self._h))

    def Sync(self):
        self._ck(lib.xrt_batch_sync(self._h))

    def run(self, frames):
        """`frames` frames of every garment back to back; returns device ms."""
        ms = C.c_float()
        self._ck(lib.xrt_batch_step_frames(self._h, int(frames), C.byref(ms)))
        return ms.value

    def profile(self, reps=8):
        ms, b = C.c_float(), C.c_double()
        self._ck(lib.xrt_batch_profile(self._h, int(reps), C.byref(ms), C.byref(b)))
        return ms.value, b.value

    def GetPositions(self, which=0, xyz=None, nrm=None, normals=False):
        """Positions (and normals) of every garment back to back, into caller arrays when given."""
        n = sum((s.n_cloth_verts if which == 0 else s.n_obstacle_verts) for s in self.solvers)
        if xyz is None:
            xyz = np.empty((n, 3), np.float32)
        if nrm is None and normals:
            nrm = np.empty((n, 3), np.float32)
        assert xyz.shape == (n, 3) and xyz.dtype == np.float32 and xyz.flags.c_contiguous
        self._ck(lib.xrt_batch_get_positions(self._h, int(which), _pf(xyz), _pf(nrm)))
        return (xyz, nrm) if nrm is not None else xyz

    def UpdateObstacleAnimationFrames(self, xyz):
        t = _f32(xyz).reshape(-1, 3)
        self._ck(lib.xrt_batch_set_obstacle_targets(self._h, _pf(t), len(t)))

    def SetReadback(self, on=True):
        """Frames end with a snapshot that is copied to the host while the next frame computes; GetPositions(0)
        then returns the oldest unread frame (at most two outstanding)."""
        self._ck(lib.xrt_batch_set_readback(self._h, int(bool(on))))

    def launch_count(self):
        """(kernels of this library enqueued so far, counted at the launch sites; calls into cub's segmented sort)"""
        a, b = C.c_uint64(), C.c_uint64()
        self._ck(lib.xrt_batch_launch_count(self._h, C.byref(a), C.byref(b)))
        return a.value, b.value

    def launches_per_frame(self, moved=True):
        n = C.c_uint32()
        self._ck(lib.xrt_batch_launches_per_frame(self._h, int(bool(moved)), C.byref(n)))
        return n.value

    def close(self):
        if getattr(self, "_h", None):
            lib.xrt_batch_destroy(self._h)
            self._h = None

    def __del__(self):
        self.close()


def batch_profile_substep(solvers, reps=20):
    """(ms of one round of the persistent substep kernel on every solver side by side, bytes of that round)."""
    arr = (C.c_void_p * len(solvers))(*[s._h for s in solvers])
    ms, b = C.c_float(), C.c_double()
    rc = lib.xrt_batch_profile_substep(arr, len(solvers), int(reps), C.byref(ms), C.byref(b))
    if rc:
        raise XrtError(rc, lib.xrt_last_error(solvers[0]._h).decode(errors="replace"))
    return ms.value, b.value


def persist_info(cluster=-1):
    """(largest cluster the device schedules, clusters of `cluster` CTAs resident at once)."""
    mx, res = C.c_int32(), C.c_int32()
    rc = lib.xrt_persist_info(int(cluster), C.byref(mx), C.byref(res))
    if rc:
        raise XrtError(rc, "xrt_persist_info failed (no CUDA device)")
    return mx.value, res.value


def build_constraints(positions, indices):
    """xrt_build_constraints: dict(edges[E,2], face_edges[F,3], stretch_rest[E], bend_idx[B,4], bend_theta[B],
    ef_offsets[E+1], ef_faces) from a vertex array and an index buffer (host work, no GPU needed)."""
    pos, tri = _f32(positions).reshape(-1, 3), _u32(indices).reshape(-1, 3)
    c = XrtConstraints()
    rc = lib.xrt_build_constraints(_pf(pos), len(pos), _pu(tri), len(tri), C.byref(c))
    if rc:
        raise XrtError(rc, "xrt_build_constraints")
    E, B, F = c.n_edges, c.n_bend, len(tri)
    arr = lambda p, n, shape: np.ctypeslib.as_array(p, shape=(max(n, 1),))[:n].copy().reshape(shape)  # noqa: E731
    out = dict(edges=arr(c.edges, 2 * E, (-1, 2)), face_edges=arr(c.face_edges, 3 * F, (-1, 3)),
               stretch_rest=arr(c.stretch_rest, E, (-1,)), bend_idx=arr(c.bend_idx, 4 * B, (-1, 4)),
               bend_theta=arr(c.bend_theta, B, (-1,)), ef_offsets=arr(c.ef_offsets, E + 1, (-1,)))
    out["ef_faces"] = arr(c.ef_faces, int(out["ef_offsets"][-1]), (-1,))
    lib.xrt_free_constraints(C.byref(c))
    return out


def greedy_color(idx, n_verts):
    idx = _u32(idx)
    arity = idx.shape[1]
    colors = np.zeros(len(idx), np.uint32)
    nc = C.c_uint32()
    rc = lib.xrt_greedy_color(_pu(idx), len(idx), arity, n_verts, _pu(colors), C.byref(nc))
    if rc:
        raise XrtError(rc, "greedy colouring failed")
    return colors, nc.value


def bvh_build_standalone(positions, indices, fast_div=1, device=0):
    pos, tri = _f32(positions).reshape(-1, 3), _u32(indices).reshape(-1, 3)
    F = len(tri)
    out = dict(morton=np.zeros(F, np.uint32), object=np.zeros(F, np.uint32), nodes=np.zeros((2 * F - 1, 4), np.uint32),
               aabb=np.zeros((2 * F - 1, 6), np.float32), esc_int=np.zeros(max(F - 1, 1), np.int32),
               esc_ext=np.zeros(F, np.int32))
    ms = C.c_float()
    rc = lib.xrt_bvh_build_standalone(_pf(pos), len(pos), _pu(tri), F, fast_div, device, _pu(out["morton"]),
                                      _pu(out["object"]), _pu(out["nodes"]), _pf(out["aabb"]),
                                      out["esc_int"].ctypes.data_as(_i32p), out["esc_ext"].ctypes.data_as(_i32p),
                                      C.byref(ms))
    if rc:
        raise XrtError(rc, "standalone BVH build failed (no CUDA device? there is no CPU fallback)")
    out["esc_int"] = out["esc_int"][:F - 1]
    out["build_ms"] = ms.value
    return out


class ObjSequenceWriter:
    """xrt_obj_writer_*: frames are written as <dir>/%06d.obj by background threads while the solver runs
    (ObjSequenceExporter, runtime/export/ObjSequenceExporter.cpp)."""

    def __init__(self, directory, indices, n_verts, with_normals=False, queue_frames=4, threads=2):
        tri = _u32(indices).reshape(-1, 3)
        h = C.c_void_p()
        rc = lib.xrt_obj_writer_open(str(directory).encode(), _pu(tri), len(tri), int(n_verts), int(with_normals),
                                     int(queue_frames), int(threads), C.byref(h))
        if rc:
            raise XrtError(rc, "obj writer: bad directory or indices")
        self._h, self._n, self._normals = h, int(n_verts), bool(with_normals)

    def push(self, xyz, normals=None):
        x = _f32(xyz).reshape(self._n, 3)
        n = _f32(normals).reshape(self._n, 3) if self._normals else None
        rc = lib.xrt_obj_writer_push(self._h, _pf(x), None if n is None else _pf(n))
        if rc:
            raise XrtError(rc, "obj writer: push failed (closed, or a file could not be written)")

    def close(self):
        if self._h is None:
            return 0, 0
        frames, nbytes = C.c_uint64(), C.c_uint64()
        rc = lib.xrt_obj_writer_close(self._h, C.byref(frames), C.byref(nbytes))
        self._h = None
        if rc:
            raise XrtError(rc, "obj writer: not every frame was written")
        return frames.value, nbytes.value


def morton_codes_noalloc(d_positions_xyzw, n_verts, d_indices, n_faces, d_morton, d_scene_box=0, fast_div=1, stream=0):
    """xrt_morton_codes_noalloc: Morton codes in face order through the library's static device scratch, no allocation.
    All d_* arguments are raw DEVICE addresses (ints) of the current CUDA device; the call only enqueues on `stream`."""
    err = C.create_string_buffer(512)
    rc = lib.xrt_morton_codes_noalloc(C.c_void_p(d_positions_xyzw), n_verts, C.c_void_p(d_indices), n_faces, fast_div,
                                      C.c_void_p(stream), C.c_void_p(d_morton), C.c_void_p(d_scene_box or None), err, len(err))
    if rc:
        raise XrtError(rc, err.value.decode() or "xrt_morton_codes_noalloc failed")


def traverse_pairs(positions, indices, other_positions=None, other_indices=None, thickness=0.0, fast_div=1, device=0,
                   timing=False):
    """BVH::Traverse (two meshes) / TraverseSelf (other_* = None), BVH.cu:1342-1400: face-face candidate rows
    (left face, right face) in the reference's order, on the GPU."""
    pl, tl = _f32(positions).reshape(-1, 3), _u32(indices).reshape(-1, 3)
    self_mode = other_indices is None
    pr = None if self_mode else _f32(other_positions).reshape(-1, 3)
    tr = None if self_mode else _u32(other_indices).reshape(-1, 3)
    n, ms = C.c_size_t(0), C.c_float()
    args = (_pf(pl), len(pl), _pu(tl), len(tl), _pf(pr), 0 if self_mode else len(pr), _pu(tr),
            0 if self_mode else len(tr), thickness, fast_div, device)
    rc = lib.xrt_traverse_pairs_standalone(*args, None, 0, C.byref(n), None)
    if rc:
        raise XrtError(rc, "pair traversal failed (no CUDA device? there is no CPU fallback)")
    out = np.zeros((n.value, 2), np.uint32)
    rc = lib.xrt_traverse_pairs_standalone(*args, _pu(out), n.value, C.byref(n), C.byref(ms))
    if rc:
        raise XrtError(rc, "pair traversal failed")
    return (out, ms.value) if timing else out


def fem_solve(rest, pos, inv_mass, tri, bend_idx, material, colors_strain, colors_bend, iterations, dt, device=0,
              timing=False):
    """FEM::StrainSolver + FEM::IsometricBendingSolver (FEMSolver.cu) for `iterations` Quality-mode iterations on
    the GPU; returns the positions."""
    rest, x, w = _f32(rest).reshape(-1, 3), _f32(pos).reshape(-1, 3), _f32(inv_mass)
    tri, bend = _u32(tri).reshape(-1, 3), _u32(bend_idx).reshape(-1, 4)
    m, cs, cb = _f32(material), _u32(colors_strain), _u32(colors_bend)
    out, ms = np.empty_like(x), C.c_float()
    rc = lib.xrt_fem_solve_standalone(_pf(rest), _pf(x), _pf(w), len(x), _pu(tri), len(tri), _pu(bend), len(bend),
                                      _pf(m), _pu(cs), _pu(cb), int(iterations), float(dt), device, _pf(out),
                                      C.byref(ms))
    if rc:
        raise XrtError(rc, "FEM solve failed (no CUDA device? there is no CPU fallback)")
    return (out, ms.value) if timing else out
