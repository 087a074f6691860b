"""ctypes wrapper of the CPU oracle (oracle/xrt_oracle.c).

TEST INFRASTRUCTURE ONLY: imported by tests/, __graft_entry__.smoke() and
bench.py's cpu_baseline / --impl reference legs.  Never imported by
xrtailor_b200/.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "liboracle.so")


def build(force=False):
    srcs = [os.path.join(_HERE, f) for f in ("xrt_oracle.c", "xrt_oracle_quality.c", "xrt_oracle_rig.c")]
    if force or not os.path.exists(_SO) or os.path.getmtime(_SO) < max(os.path.getmtime(f) for f in srcs):
        subprocess.check_call(["make", "-C", _HERE, "-s", "liboracle.so"])
    return _SO


def _load():
    L = C.CDLL(build())
    vp, f32p, u32p, i32p = C.c_void_p, C.POINTER(C.c_float), C.POINTER(C.c_uint32), C.POINTER(C.c_int32)
    L.oracle_expand_bits.argtypes = [C.c_uint32]; L.oracle_expand_bits.restype = C.c_uint32
    L.oracle_morton_code.argtypes = [C.c_float] * 3; L.oracle_morton_code.restype = C.c_uint32
    L.oracle_bvh_build.argtypes = [f32p, f32p, f32p, u32p, C.c_uint32, C.c_float, C.c_int]
    L.oracle_bvh_build.restype = vp
    L.oracle_bvh_free.argtypes = [vp]; L.oracle_bvh_free.restype = None
    L.oracle_bvh_refit_midstep.argtypes = [vp, f32p, f32p, u32p, C.c_float]; L.oracle_bvh_refit_midstep.restype = None
    for n, t in [("morton", u32p), ("object", u32p), ("nodes", u32p), ("aabb", f32p), ("esc_int", i32p),
                 ("esc_ext", i32p), ("pred", f32p)]:
        f = getattr(L, "oracle_bvh_" + n); f.argtypes = [vp]; f.restype = t
    L.oracle_bvh_used64.argtypes = [vp]; L.oracle_bvh_used64.restype = C.c_int
    L.oracle_fem_solve.argtypes = [f32p, f32p, f32p, C.c_uint32, u32p, C.c_uint32, u32p, C.c_uint32, f32p, u32p, u32p,
                                   C.c_int, C.c_float]
    L.oracle_fem_solve.restype = None
    L.oracle_rodrigues.argtypes = [f32p, f32p]; L.oracle_rodrigues.restype = None
    L.oracle_local_to_global.argtypes = [f32p, f32p, f32p, i32p, C.c_uint32, f32p]; L.oracle_local_to_global.restype = None
    L.oracle_pose_blend_params.argtypes = [f32p, C.c_uint32, f32p]; L.oracle_pose_blend_params.restype = None
    L.oracle_lbs.argtypes = [f32p, C.c_uint32, u32p, u32p, f32p, f32p, f32p, f32p, f32p, C.c_uint32, f32p]
    L.oracle_lbs.restype = None
    L.oracle_traverse_pairs.argtypes = [vp, vp, C.c_float, u32p, C.c_size_t]; L.oracle_traverse_pairs.restype = C.c_size_t
    L.oracle_nearest_point_f.argtypes = [f32p, f32p, f32p]; L.oracle_nearest_point_f.restype = C.c_float
    L.oracle_ray_intersect_f.argtypes = [f32p, f32p, C.c_float, f32p]; L.oracle_ray_intersect_f.restype = C.c_float
    L.oracle_closest_seg_seg_f.argtypes = [f32p, f32p, f32p]; L.oracle_closest_seg_seg_f.restype = C.c_float
    L.oracle_overlap_point_f.argtypes = [f32p, f32p]; L.oracle_overlap_point_f.restype = C.c_int
    L.oracle_overlap_segment_f.argtypes = [f32p, f32p, f32p]; L.oracle_overlap_segment_f.restype = C.c_int
    L.oracle_query_f.argtypes = [vp, C.c_int, f32p, f32p, u32p]; L.oracle_query_f.restype = C.c_uint32
    L.oracle_rtri.argtypes = [u32p, u32p, u32p, C.c_uint32, u32p]; L.oracle_rtri.restype = None
    L.oracle_scene_new.argtypes = [i32p, f32p]; L.oracle_scene_new.restype = vp
    L.oracle_scene_set_cloth.argtypes = [vp, f32p, f32p, C.c_uint32, u32p, C.c_uint32, u32p, C.c_uint32]
    L.oracle_scene_set_stretch.argtypes = [vp, u32p, f32p, u32p, C.c_uint32]
    L.oracle_scene_set_bend.argtypes = [vp, u32p, f32p, u32p, C.c_uint32]
    L.oracle_scene_set_lra.argtypes = [vp, u32p, u32p, f32p, C.c_uint32]
    L.oracle_scene_set_binding.argtypes = [vp, u32p, u32p, f32p, f32p, f32p, C.c_uint32]
    L.oracle_scene_set_binding.restype = None
    L.oracle_scene_set_obstacle.argtypes = [vp, f32p, C.c_uint32, u32p, C.c_uint32]
    L.oracle_scene_set_obstacle_target.argtypes = [vp, f32p]
    L.oracle_scene_set_state.argtypes = [vp, f32p, f32p]
    for n in ["set_cloth", "set_stretch", "set_bend", "set_lra", "set_obstacle", "set_obstacle_target", "set_state",
              "free"]:
        getattr(L, "oracle_scene_" + n).restype = None
    L.oracle_scene_free.argtypes = [vp]
    for n in ["oracle_step_frame", "oracle_generate_contacts", "oracle_hash", "oracle_collide_particles",
              "oracle_solve_lra", "oracle_apply_deltas", "oracle_solve_vf_ccd", "oracle_solve_vf_dcd",
              "oracle_solve_ee_dcd", "oracle_solve_binding"]:
        f = getattr(L, n); f.argtypes = [vp]; f.restype = None
    L.oracle_predict.argtypes = [vp, C.c_float]; L.oracle_predict.restype = None
    L.oracle_finalize.argtypes = [vp, C.c_float]; L.oracle_finalize.restype = None
    L.oracle_iteration.argtypes = [vp, C.c_float, C.c_int]; L.oracle_iteration.restype = None
    L.oracle_collision_pass.argtypes = [vp, C.c_int]; L.oracle_collision_pass.restype = None
    L.oracle_solve_stretch.argtypes = [vp, C.c_uint32, C.c_float, C.c_int]; L.oracle_solve_stretch.restype = None
    L.oracle_solve_bend.argtypes = [vp, C.c_uint32, C.c_float]; L.oracle_solve_bend.restype = None
    for n, t in [("x0", f32p), ("x", f32p), ("v", f32p), ("n", f32p), ("ox0", f32p), ("on", f32p), ("bvh", vp),
                 ("rtri", u32p), ("vf_idx", i32p), ("vf_q", f32p), ("rt_idx", i32p), ("rt_q", f32p), ("ee", f32p),
                 ("neighbors", u32p), ("p_hash", u32p), ("p_index", u32p)]:
        f = getattr(L, "oracle_scene_" + n); f.argtypes = [vp]; f.restype = t
    L.oracle_scene_cnt.argtypes = [vp, C.c_int]; L.oracle_scene_cnt.restype = u32p
    # ---- Quality mode (xrt_oracle_quality.c) ----
    szp = C.POINTER(C.c_size_t)
    L.oracle_scene_set_quality.argtypes = [vp, i32p, f32p, u32p, C.c_uint32, u32p]; L.oracle_scene_set_quality.restype = None
    for n in ["oracle_step_frame_quality", "oracle_imminent_generate", "oracle_rep_update_pairs", "oracle_rep_solve",
              "oracle_quality_begin_frame"]:
        f = getattr(L, n); f.argtypes = [vp]; f.restype = None
    L.oracle_imminent_solve.argtypes = [vp, C.c_float]; L.oracle_imminent_solve.restype = None
    L.oracle_collision_step.argtypes = [vp, C.c_float]; L.oracle_collision_step.restype = None
    L.oracle_quality_iteration.argtypes = [vp, C.c_float, C.c_int]; L.oracle_quality_iteration.restype = None
    L.oracle_find_impacts.argtypes = [vp, C.c_float, u32p, f32p, C.c_size_t, szp]; L.oracle_find_impacts.restype = C.c_size_t
    L.oracle_independent_impacts.argtypes = [vp, u32p, f32p, C.c_size_t, C.c_int, u32p, f32p]
    L.oracle_independent_impacts.restype = C.c_size_t
    L.oracle_isolate_zones.argtypes = [vp, u32p, C.c_size_t, i32p]; L.oracle_isolate_zones.restype = C.c_int
    L.oracle_optimize_zones.argtypes = [vp, u32p, f32p, i32p, C.c_size_t, C.c_int, C.c_int, C.c_float, C.c_float, i32p]
    L.oracle_optimize_zones.restype = None
    L.oracle_rep_contributions.argtypes = [vp, C.c_int, i32p, f32p, C.c_int]; L.oracle_rep_contributions.restype = C.c_int
    L.oracle_quality_num_quads.argtypes = [vp]; L.oracle_quality_num_quads.restype = C.c_size_t
    L.oracle_quality_get_quads.argtypes = [vp, u32p, f32p]; L.oracle_quality_get_quads.restype = None
    L.oracle_quality_num_rep_pairs.argtypes = [vp]; L.oracle_quality_num_rep_pairs.restype = C.c_size_t
    L.oracle_quality_rep_pairs.argtypes = [vp]; L.oracle_quality_rep_pairs.restype = u32p
    L.oracle_quality_stats.argtypes = [vp, i32p]; L.oracle_quality_stats.restype = None
    L.oracle_quality_cloth_rtri.argtypes = [vp]; L.oracle_quality_cloth_rtri.restype = u32p
    L.oracle_scene_ox.argtypes = [vp]; L.oracle_scene_ox.restype = f32p
    L.oracle_bvh_update.argtypes = [vp, f32p, f32p, u32p, C.c_int]; L.oracle_bvh_update.restype = None
    L.oracle_solve_cubic.argtypes = [C.c_float] * 4 + [f32p]; L.oracle_solve_cubic.restype = C.c_int
    return L


L = _load()
_f32p, _u32p, _i32p = C.POINTER(C.c_float), C.POINTER(C.c_uint32), C.POINTER(C.c_int32)


def _f(a):
    return np.ascontiguousarray(a, np.float32)


def _u(a):
    return np.ascontiguousarray(a, np.uint32)


def _pf(a):
    return a.ctypes.data_as(_f32p) if a is not None else None


def _pu(a):
    return a.ctypes.data_as(_u32p) if a is not None else None


def _arr(ptr, n, dtype):
    if n == 0:
        return np.zeros(0, dtype)
    return np.ctypeslib.as_array(ptr, shape=(n,)).astype(dtype, copy=True)


def _bvh_dict(h, F):
    NN = 2 * F - 1
    return dict(morton=_arr(L.oracle_bvh_morton(h), F, np.uint32), object=_arr(L.oracle_bvh_object(h), F, np.uint32),
                nodes=_arr(L.oracle_bvh_nodes(h), 4 * NN, np.uint32).reshape(-1, 4),
                aabb=_arr(L.oracle_bvh_aabb(h), 6 * NN, np.float32).reshape(-1, 6),
                esc_int=_arr(L.oracle_bvh_esc_int(h), F - 1, np.int32),
                esc_ext=_arr(L.oracle_bvh_esc_ext(h), F, np.int32),
                pred=_arr(L.oracle_bvh_pred(h), 9 * F, np.float32).reshape(-1, 9), used64=L.oracle_bvh_used64(h))


class Bvh:
    """BVH::UpdateData + Construct on the CPU."""

    def __init__(self, x0, tri, x=None, normals=None, tol=0.0, ccd=False):
        self.x0, self.tri = _f(x0).reshape(-1, 3), _u(tri).reshape(-1, 3)
        self.x = _f(x).reshape(-1, 3) if x is not None else self.x0
        self.nrm = _f(normals).reshape(-1, 3) if normals is not None else None
        self.F = len(self.tri)
        self.h = L.oracle_bvh_build(_pf(self.x0), _pf(self.x), _pf(self.nrm), _pu(self.tri), self.F, tol, int(ccd))

    def arrays(self):
        return _bvh_dict(self.h, self.F)

    def refit_midstep(self, x1, n1, tol):
        x1, n1 = _f(x1), _f(n1)
        L.oracle_bvh_refit_midstep(self.h, _pf(x1), _pf(n1), _pu(self.tri), tol)

    def query(self, kind, p1, p2=None):
        out = np.zeros(256, np.uint32)
        p1 = _f(p1)
        p2 = _f(p2) if p2 is not None else None
        n = L.oracle_query_f(self.h, kind, _pf(p1), _pf(p2), _pu(out))
        return out[:n].copy()

    def pairs(self, other=None, thickness=0.0):
        """CountPairs + scan + FindPairs (BVH.cu:1149-1340): (lhs face, rhs face) rows in the reference's order;
        other=None is the self traversal."""
        o = other if other is not None else self
        n = L.oracle_traverse_pairs(self.h, o.h, thickness, None, 0)
        out = np.zeros((n, 2), np.uint32)
        if n:
            L.oracle_traverse_pairs(self.h, o.h, thickness, _pu(out), n)
        return out

    def __del__(self):
        if getattr(self, "h", None):
            L.oracle_bvh_free(self.h)
            self.h = None


class Scene:
    """The Swift frame on the CPU, from a xrtailor_b200.scene.Scene + params + colours."""

    def __init__(self, sc, params, colors):
        p = params
        ip = np.array([p.num_substeps, p.num_iterations, p.num_collision_passes, p.max_num_neighbors,
                       p.inter_leaved_hash, p.enable_self_collision, p.solve_bending, p.geodesic_lra], np.int32)
        fp = np.array([p.gravity[0], p.gravity[1], p.gravity[2], p.damping, p.relaxation_factor,
                       p.long_range_stretchiness, p.sdf_friction, p.bvh_tolerance, p.stretch_compliance,
                       p.bend_compliance, p.particle_diameter, p.hash_cell_size, p.frame_dt], np.float32)
        self.h = L.oracle_scene_new(ip.ctypes.data_as(_i32p), _pf(fp))
        self.V, self.E = len(sc.cloth_pos), len(sc.cloth_edges)
        self.Vo = 0 if sc.obs_pos is None else len(sc.obs_pos)
        self.Fo = 0 if sc.obs_tri is None else len(sc.obs_tri)
        self.max_nb = p.max_num_neighbors
        self.dt = p.frame_dt / p.num_substeps
        L.oracle_scene_set_cloth(self.h, _pf(_f(sc.cloth_pos)), _pf(_f(sc.inv_mass)), self.V, _pu(_u(sc.cloth_tri)),
                                 len(sc.cloth_tri), _pu(_u(sc.cloth_edges)), self.E)
        L.oracle_scene_set_stretch(self.h, _pu(_u(sc.stretch_idx)), _pf(_f(sc.stretch_rest)), _pu(_u(colors[0])),
                                   len(sc.stretch_rest))
        L.oracle_scene_set_bend(self.h, _pu(_u(sc.bend_idx)), _pf(_f(sc.bend_theta)), _pu(_u(colors[1])),
                                len(sc.bend_theta))
        if self.Vo:
            L.oracle_scene_set_obstacle(self.h, _pf(_f(sc.obs_pos)), self.Vo, _pu(_u(sc.obs_tri)), self.Fo)

    def set_lra(self, src, tgt, rest):
        L.oracle_scene_set_lra(self.h, _pu(_u(src)), _pu(_u(tgt)), _pf(_f(rest)), len(rest))

    def set_binding(self, vert, tri3, bary3, stiffness, distance):
        """BindingConstraint::Add + Generate (BindingConstraint.cu:11-55)."""
        L.oracle_scene_set_binding(self.h, _pu(_u(vert)), _pu(_u(tri3)), _pf(_f(bary3)), _pf(_f(stiffness)),
                                   _pf(_f(distance)), len(vert))

    def solve_binding(self):
        L.oracle_solve_binding(self.h)

    def set_obstacle_target(self, xyz):
        L.oracle_scene_set_obstacle_target(self.h, _pf(_f(xyz)))

    def set_state(self, x0=None, v=None):
        L.oracle_scene_set_state(self.h, _pf(_f(x0)) if x0 is not None else None, _pf(_f(v)) if v is not None else None)

    def step_frame(self):
        L.oracle_step_frame(self.h)

    def predict(self):
        L.oracle_predict(self.h, self.dt)

    def finalize(self):
        L.oracle_finalize(self.h, self.dt)

    def iteration(self, it, self_collision=False):
        if self_collision:
            L.oracle_collide_particles(self.h)
        L.oracle_iteration(self.h, self.dt, it)

    def collision_pass(self, cnt=0):
        L.oracle_collision_pass(self.h, cnt)

    def generate_contacts(self):
        L.oracle_generate_contacts(self.h)

    def hash(self):
        L.oracle_hash(self.h)

    def positions(self):
        return _arr(L.oracle_scene_x0(self.h), 3 * self.V, np.float32).reshape(-1, 3)

    def predicted(self):
        return _arr(L.oracle_scene_x(self.h), 3 * self.V, np.float32).reshape(-1, 3)

    def set_predicted(self, x):
        """Overwrite Node::x (test helper)."""
        np.ctypeslib.as_array(L.oracle_scene_x(self.h), shape=(3 * self.V,))[:] = _f(x).ravel()

    def velocity(self):
        return _arr(L.oracle_scene_v(self.h), 3 * self.V, np.float32).reshape(-1, 3)

    def normals(self):
        return _arr(L.oracle_scene_n(self.h), 3 * self.V, np.float32).reshape(-1, 3)

    def bvh(self):
        return _bvh_dict(L.oracle_scene_bvh(self.h), self.Fo)

    def rtri(self):
        return _arr(L.oracle_scene_rtri(self.h), self.Fo, np.uint32)

    def contacts(self):
        return dict(vf_idx=_arr(L.oracle_scene_vf_idx(self.h), self.V, np.int32),
                    vf_q=_arr(L.oracle_scene_vf_q(self.h), 3 * self.V, np.float32).reshape(-1, 3),
                    rt_idx=_arr(L.oracle_scene_rt_idx(self.h), self.V, np.int32),
                    rt_q=_arr(L.oracle_scene_rt_q(self.h), 3 * self.V, np.float32).reshape(-1, 3),
                    ee=_arr(L.oracle_scene_ee(self.h), 14 * self.E, np.float32).reshape(-1, 14),
                    cnt_vf=_arr(L.oracle_scene_cnt(self.h, 0), self.V, np.uint32),
                    cnt_rt=_arr(L.oracle_scene_cnt(self.h, 1), self.V, np.uint32),
                    cnt_ee=_arr(L.oracle_scene_cnt(self.h, 2), self.E, np.uint32))

    def neighbors(self):
        return _arr(L.oracle_scene_neighbors(self.h), self.V * self.max_nb, np.uint32)

    def particle_hash(self):
        return (_arr(L.oracle_scene_p_hash(self.h), self.V, np.uint32),
                _arr(L.oracle_scene_p_index(self.h), self.V, np.uint32))

    # ---- Quality mode (SimulateQuality, ClothSolverBase.cpp:263-362) ----
    def set_quality(self, sc, q, strain_colors):
        """q: any object with the xrt_quality_params fields; strain_colors: one colour per cloth face."""
        qp = np.array([q.enable_imminent_repulsion, q.enable_pbd_repulsion], np.int32)
        qf = np.array([q.xx_stiffness, q.yy_stiffness, q.xy_stiffness, q.xy_poisson_ratio, q.yx_poisson_ratio,
                       q.bending_stiffness, q.imminent_thickness, q.pbd_thickness, q.relaxation_rate, q.obstacle_mass,
                       q.impact_zone_thickness], np.float32)
        oe = _u(sc.obs_edges) if self.Vo else None
        self.F = len(sc.cloth_tri)
        L.oracle_scene_set_quality(self.h, qp.ctypes.data_as(_i32p), _pf(qf), _pu(oe), 0 if oe is None else len(oe),
                                   _pu(_u(strain_colors)))

    def step_frame_quality(self):
        L.oracle_step_frame_quality(self.h)

    def quality_begin_frame(self):
        L.oracle_quality_begin_frame(self.h)

    def quality_iteration(self, it):
        L.oracle_quality_iteration(self.h, self.dt, it)

    def collision_step(self):
        L.oracle_collision_step(self.h, self.dt)

    def collision_stats(self):
        out = np.zeros(4, np.int32)
        L.oracle_quality_stats(self.h, out.ctypes.data_as(_i32p))
        return dict(rounds=int(out[0]), success=int(out[1]), impacts=int(out[2]), deform=int(out[3]))

    def find_impacts(self, thickness):
        """FindImpacts with fresh ccd trees: (nodes[n,4] global ids, twn[n,9] = t, w[4], n[3], is_vf, n_pairs)."""
        npairs = C.c_size_t(0)
        n = L.oracle_find_impacts(self.h, thickness, None, None, 0, C.byref(npairs))
        nodes, twn = np.zeros((n, 4), np.uint32), np.zeros((n, 9), np.float32)
        if n:
            L.oracle_find_impacts(self.h, thickness, _pu(nodes), _pf(twn), n, C.byref(npairs))
        return nodes, twn, int(npairs.value)

    def independent_impacts(self, nodes, twn, deform=0):
        nodes, twn = _u(nodes).reshape(-1, 4), _f(twn).reshape(-1, 9)
        on, ot = np.zeros_like(nodes), np.zeros_like(twn)
        m = L.oracle_independent_impacts(self.h, _pu(nodes), _pf(twn), len(nodes), deform, _pu(on), _pf(ot))
        return on[:m].copy(), ot[:m].copy()

    def isolate_zones(self, nodes):
        nodes = _u(nodes).reshape(-1, 4)
        col = np.zeros(len(nodes), np.int32)
        nc = L.oracle_isolate_zones(self.h, _pu(nodes), len(nodes), col.ctypes.data_as(_i32p))
        return col, nc

    def optimize_zones(self, nodes, twn, colors, deform, thickness, obstacle_mass):
        nodes, twn, colors = _u(nodes).reshape(-1, 4), _f(twn).reshape(-1, 9), np.ascontiguousarray(colors, np.int32)
        st = np.zeros(3, np.int32)
        L.oracle_optimize_zones(self.h, _pu(nodes), _pf(twn), colors.ctypes.data_as(_i32p), len(nodes),
                                int(colors.max()) + 1 if len(colors) else 0, deform, thickness, obstacle_mass,
                                st.ctypes.data_as(_i32p))
        return dict(outer=int(st[0]), line_search=int(st[1]), converged=int(st[2]))

    def imminent_generate(self):
        L.oracle_imminent_generate(self.h)
        n = L.oracle_quality_num_quads(self.h)
        nodes, bary = np.zeros((n, 4), np.uint32), np.zeros((n, 4), np.float32)
        if n:
            L.oracle_quality_get_quads(self.h, _pu(nodes), _pf(bary))
        return nodes, bary

    def imminent_solve(self):
        L.oracle_imminent_solve(self.h, self.dt)

    def rep_update_pairs(self):
        L.oracle_rep_update_pairs(self.h)
        n = L.oracle_quality_num_rep_pairs(self.h)
        return _arr(L.oracle_quality_rep_pairs(self.h), 2 * n, np.uint32).reshape(-1, 2)

    def rep_solve(self):
        L.oracle_rep_solve(self.h)

    def rep_contributions(self, vertex, cap=256):
        rows, corr = np.zeros((cap, 6), np.int32), np.zeros((cap, 3), np.float32)
        n = L.oracle_rep_contributions(self.h, int(vertex), rows.ctypes.data_as(_i32p), _pf(corr), cap)
        return rows[:min(n, cap)], corr[:min(n, cap)]

    def cloth_rtri(self):
        return _arr(L.oracle_quality_cloth_rtri(self.h), self.F, np.uint32)

    def obstacle_predicted(self):
        return _arr(L.oracle_scene_ox(self.h), 3 * self.Vo, np.float32).reshape(-1, 3)

    def __del__(self):
        if getattr(self, "h", None):
            L.oracle_scene_free(self.h)
            self.h = None


def solve_cubic(a, b, c, d):
    """MathFunctions::SolveCubic (math/MathFunctions.cu:122-146): roots of a t^3 + b t^2 + c t + d."""
    x = np.zeros(3, np.float32)
    n = L.oracle_solve_cubic(a, b, c, d, _pf(x))
    return x[:n].copy()


def rtri(indices, n_verts):
    tri = _u(indices).reshape(-1, 3)
    F = len(tri)
    prefix = np.zeros(n_verts + 1, np.uint32)
    np.add.at(prefix, tri.ravel() + 1, 1)
    prefix = np.cumsum(prefix).astype(np.uint32)
    order = np.argsort(tri.ravel(), kind="stable")
    nb = (order // 3).astype(np.uint32)  # faces ascending per vertex (stable sort of face-major list)
    out = np.zeros(F, np.uint32)
    L.oracle_rtri(_pu(tri), _pu(nb), _pu(prefix), F, _pu(out))
    return out


def fem_solve(rest, pos, inv_mass, tri, bend_idx, material, colors_strain, colors_bend, iterations, dt):
    """FEM strain + isometric bending (FEMHelper.cu) for `iterations` Quality-mode iterations; returns positions."""
    rest, x, w = _f(rest).reshape(-1, 3), _f(pos).reshape(-1, 3).copy(), _f(inv_mass)
    tri, bend = _u(tri).reshape(-1, 3), _u(bend_idx).reshape(-1, 4)
    m, cs, cb = _f(material), _u(colors_strain), _u(colors_bend)
    L.oracle_fem_solve(_pf(rest), _pf(x), _pf(w), len(x), _pu(tri), len(tri), _pu(bend), len(bend), _pf(m), _pu(cs),
                       _pu(cb), int(iterations), float(dt))
    return x


def rig_joint_transforms(pose, joints, parent, trans=None):
    """Body<>::Update + _local_to_global (runtime/rag_doll/smpl/body.cpp:308-325,386-415): J x 12 joint transforms."""
    pose, joints = _f(pose).reshape(-1, 3), _f(joints).reshape(-1, 3)
    par = np.ascontiguousarray(parent, np.int32)
    T = np.zeros((len(joints), 12), np.float32)
    t = None if trans is None else _f(trans)
    L.oracle_local_to_global(_pf(pose), None if t is None else _pf(t), _pf(joints), par.ctypes.data_as(C.POINTER(C.c_int32)),
                              len(joints), _pf(T))
    return T


def rig_pose_blend_params(pose):
    pose = _f(pose).reshape(-1, 3)
    out = np.zeros(9 * (len(pose) - 1), np.float32)
    L.oracle_pose_blend_params(_pf(pose), len(pose), _pf(out))
    return out


def rig_lbs(rest, w_outer, w_inner, w_val, T, model_matrix=None, pose_blend=None, blend_params=None):
    """device::lbs (runtime/rag_doll/smpl/cuda/body.cu:39-63) + UpdateObstaclePredicted_Kernel (ClothSolverHelper.cu:70-79)."""
    rest = _f(rest).reshape(-1, 3)
    wo, wi, wv, T = _u(w_outer), _u(w_inner), _f(w_val), _f(T)
    M = None if model_matrix is None else _f(model_matrix)
    pb = None if pose_blend is None else _f(pose_blend)
    bp = None if blend_params is None else _f(blend_params)
    out = np.zeros_like(rest)
    L.oracle_lbs(_pf(rest), len(rest), _pu(wo), _pu(wi), _pf(wv), _pf(T), None if M is None else _pf(M),
                  None if pb is None else _pf(pb), None if bp is None else _pf(bp), 0 if bp is None else len(bp), _pf(out))
    return out
