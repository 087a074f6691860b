"""ctypes wrapper of oracle/_ref/libxrt_ref.so — the reference's own CUDA
physics (compiled from /root/reference by oracle/ref_harness/Makefile) behind a
GL-free harness.  TEST INFRASTRUCTURE ONLY: used by tests/ as the oracle on the
GPU box and by bench.py's reference arm.  Needs a GPU; nothing here reads
/root/reference at run time."""
import ctypes as C
import os
import time

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
# XRT_REF_NO_TIMERS=1 selects the build with the reference's per-kernel GPU timers stubbed (Makefile NT=1)
SO = os.path.join(_HERE, "..", "_ref",
                  "libxrt_ref_nt.so" if os.environ.get("XRT_REF_NO_TIMERS") == "1" else "libxrt_ref.so")
_lib = None


def available():
    if not os.path.exists(SO):
        return False
    try:
        import torch
        return torch.cuda.is_available()
    except Exception:
        return False


def lib():
    global _lib
    if _lib is None:
        L = C.CDLL(SO)
        vp, f32p, u32p, i32p = C.c_void_p, C.POINTER(C.c_float), C.POINTER(C.c_uint32), C.POINTER(C.c_int32)
        L.ref_create.argtypes = [i32p, f32p, f32p, f32p, C.c_int, u32p, C.c_int, u32p, C.c_int, u32p, f32p, C.c_int,
                                 u32p, f32p, C.c_int, f32p, C.c_int, u32p, C.c_int, u32p, C.c_int]
        L.ref_create.restype = vp
        L.ref_set_obstacle_target.argtypes = [vp, f32p]; L.ref_set_obstacle_target.restype = None
        L.ref_set_bindings.argtypes = [vp, u32p, u32p, f32p, f32p, f32p, C.c_int]; L.ref_set_bindings.restype = None
        L.ref_step_frame.argtypes = [vp]; L.ref_step_frame.restype = C.c_float
        L.ref_get_state.argtypes = [vp, C.c_int, f32p, f32p, f32p, f32p]; L.ref_get_state.restype = None
        L.ref_set_cloth_state.argtypes = [vp, f32p, f32p, f32p]; L.ref_set_cloth_state.restype = None
        L.ref_num_colors.argtypes = [vp, C.c_int]; L.ref_num_colors.restype = C.c_int
        L.ref_get_colors.argtypes = [vp, C.c_int, u32p, C.c_int]; L.ref_get_colors.restype = None
        L.ref_set_colors.argtypes = [vp, C.c_int, u32p, C.c_int, C.c_int]; L.ref_set_colors.restype = None
        L.ref_get_bvh.argtypes = [vp, u32p, u32p, f32p, i32p, i32p, f32p]; L.ref_get_bvh.restype = None
        L.ref_get_contacts.argtypes = [vp, i32p, f32p, i32p, f32p, f32p]; L.ref_get_contacts.restype = None
        L.ref_generate_contacts.argtypes = [vp]; L.ref_generate_contacts.restype = None
        L.ref_get_rtri.argtypes = [vp, u32p, C.c_int]; L.ref_get_rtri.restype = None
        L.ref_fem_solve.argtypes = [vp, f32p, f32p, u32p, C.c_int, u32p, C.c_int, u32p, C.c_int, C.c_int, C.c_float, f32p]
        L.ref_fem_solve.restype = C.c_int
        L.ref_traverse_pairs.argtypes = [vp, C.c_int, C.c_float, u32p, C.c_longlong]
        L.ref_traverse_pairs.restype = C.c_longlong
        L.ref_destroy.argtypes = [vp]; L.ref_destroy.restype = None
        ll, llp = C.c_longlong, C.POINTER(C.c_longlong)
        L.ref_setup_quality.argtypes = [vp, i32p, f32p, u32p, u32p, u32p, u32p, C.c_int, u32p, C.c_int, u32p, u32p]
        L.ref_setup_quality.restype = None
        L.ref_step_frame_quality.argtypes = [vp]; L.ref_step_frame_quality.restype = C.c_float
        L.ref_find_impacts.argtypes = [vp, C.c_float, C.c_float, u32p, f32p, ll, llp]; L.ref_find_impacts.restype = ll
        L.ref_independent_impacts.argtypes = [vp, C.c_int, u32p, f32p, ll]; L.ref_independent_impacts.restype = ll
        L.ref_collision_step.argtypes = [vp, C.c_float]; L.ref_collision_step.restype = None
        L.ref_quality_timing.argtypes = [vp, C.c_float, C.c_float, C.c_int, f32p, C.POINTER(C.c_longlong)]; L.ref_quality_timing.restype = None
        L.ref_rep_evaluate_only.argtypes = [vp, f32p, i32p]; L.ref_rep_evaluate_only.restype = None
        L.ref_rep_contributions.argtypes = [vp, C.c_int, i32p, f32p, C.c_int]; L.ref_rep_contributions.restype = C.c_int
        L.ref_imminent_generate.argtypes = [vp, u32p, f32p, ll]; L.ref_imminent_generate.restype = ll
        L.ref_imminent_solve.argtypes = [vp, C.c_float, C.c_int]; L.ref_imminent_solve.restype = None
        L.ref_rep_update_pairs.argtypes = [vp, u32p, ll]; L.ref_rep_update_pairs.restype = ll
        L.ref_rep_solve.argtypes = [vp, C.c_float, C.c_int]; L.ref_rep_solve.restype = None
        L.ref_obstacle_bvh_update.argtypes = [vp]; L.ref_obstacle_bvh_update.restype = None
        L.ref_get_cloth_rtri.argtypes = [vp, u32p, C.c_int]; L.ref_get_cloth_rtri.restype = None
        L.ref_set_obstacle_state.argtypes = [vp, f32p, f32p, f32p]; L.ref_set_obstacle_state.restype = None
        L.ref_build_obstacle_bvh.argtypes = [vp]; L.ref_build_obstacle_bvh.restype = None
        L.ref_get_mortons.argtypes = [vp, u32p]; L.ref_get_mortons.restype = None
        L.ref_add_geodesics.argtypes = [vp, u32p, u32p, f32p, C.c_int]; L.ref_add_geodesics.restype = None
        _lib = L
    return _lib


def _f(a):
    return np.ascontiguousarray(a, np.float32)


def _u(a):
    return np.ascontiguousarray(a, np.uint32)


def _pf(a):
    return a.ctypes.data_as(C.POINTER(C.c_float)) if a is not None else None


def _pu(a):
    return a.ctypes.data_as(C.POINTER(C.c_uint32)) if a is not None else None


def _pi(a):
    return a.ctypes.data_as(C.POINTER(C.c_int32)) if a is not None else None


class RefSolver:
    """The reference's Swift solver on a xrtailor_b200.scene.Scene with xrt_params p."""

    def __init__(self, sc, p, seed=20261104, device=0):
        L = lib()
        ip = np.array([p.num_substeps, p.num_iterations, p.num_collision_passes, p.max_num_neighbors,
                       p.inter_leaved_hash, p.enable_self_collision, p.solve_bending, p.geodesic_lra, device,
                       seed & 0x7FFFFFFF], np.int32)
        fp = np.array([p.gravity[0], p.gravity[1], p.gravity[2], p.damping, p.relaxation_factor,
                       p.long_range_stretchiness, p.sdf_friction, p.bvh_tolerance, p.stretch_compliance,
                       p.bend_compliance, p.particle_diameter, p.hash_cell_size], np.float32)
        self.V, self.E, self.F = len(sc.cloth_pos), len(sc.cloth_edges), len(sc.cloth_tri)
        self.Vo, self.Fo = len(sc.obs_pos), len(sc.obs_tri)
        self.ns, self.nb = len(sc.stretch_rest), len(sc.bend_theta)
        keep = [_f(sc.cloth_pos), _f(sc.inv_mass), _u(sc.cloth_tri), _u(sc.cloth_edges), _u(sc.stretch_idx),
                _f(sc.stretch_rest), _u(sc.bend_idx), _f(sc.bend_theta), _f(sc.obs_pos), _u(sc.obs_tri),
                _u(sc.obs_edges)]
        self.h = L.ref_create(_pi(ip), _pf(fp), _pf(keep[0]), _pf(keep[1]), self.V, _pu(keep[2]), self.F,
                              _pu(keep[3]), self.E, _pu(keep[4]), _pf(keep[5]), self.ns, _pu(keep[6]), _pf(keep[7]),
                              self.nb, _pf(keep[8]), self.Vo, _pu(keep[9]), self.Fo, _pu(keep[10]),
                              len(sc.obs_edges))

    def colors(self):
        cs, cb = np.zeros(self.ns, np.uint32), np.zeros(self.nb, np.uint32)
        lib().ref_get_colors(self.h, 0, _pu(cs), self.ns)
        lib().ref_get_colors(self.h, 1, _pu(cb), self.nb)
        return cs, cb, lib().ref_num_colors(self.h, 0), lib().ref_num_colors(self.h, 1)

    def set_colors(self, cs, cb):
        cs, cb = _u(cs), _u(cb)
        lib().ref_set_colors(self.h, 0, _pu(cs), len(cs), int(cs.max()) + 1)
        lib().ref_set_colors(self.h, 1, _pu(cb), len(cb), int(cb.max()) + 1)

    def set_bindings(self, idx, tri, bary, stiffness, distance):
        k = [_u(idx), _u(tri), _f(bary), _f(stiffness), _f(distance)]
        lib().ref_set_bindings(self.h, _pu(k[0]), _pu(k[1]), _pf(k[2]), _pf(k[3]), _pf(k[4]), len(k[0]))

    def set_obstacle_target(self, xyz):
        lib().ref_set_obstacle_target(self.h, _pf(_f(xyz)))

    def step_frame(self):
        return lib().ref_step_frame(self.h)

    def state(self, which=0):
        n = self.V if which == 0 else self.Vo
        out = [np.zeros((n, 3), np.float32) for _ in range(4)]
        lib().ref_get_state(self.h, which, *[_pf(a) for a in out])
        return dict(x0=out[0], x=out[1], v=out[2], n=out[3])

    def set_cloth_state(self, x0=None, x=None, v=None):
        a = [None if t is None else _f(t) for t in (x0, x, v)]
        lib().ref_set_cloth_state(self.h, *[_pf(t) for t in a])

    def bvh(self):
        F = self.Fo
        nodes, aabb = np.zeros((2 * F - 1, 4), np.uint32), np.zeros((2 * F - 1, 6), np.float32)
        ei, ee, pred = np.zeros(F - 1, np.int32), np.zeros(F, np.int32), np.zeros((F, 9), np.float32)
        lib().ref_get_bvh(self.h, None, _pu(nodes), _pf(aabb), _pi(ei), _pi(ee), _pf(pred))
        return dict(nodes=nodes, aabb=aabb, esc_int=ei, esc_ext=ee, pred=pred, object=nodes[F - 1:, 3].copy())

    def generate_contacts(self):
        lib().ref_generate_contacts(self.h)

    def contacts(self):
        vi, ri = np.zeros(self.V, np.int32), np.zeros(self.V, np.int32)
        vq, rq = np.zeros((self.V, 3), np.float32), np.zeros((self.V, 3), np.float32)
        ee = np.zeros((self.E, 12), np.float32)
        lib().ref_get_contacts(self.h, _pi(vi), _pf(vq), _pi(ri), _pf(rq), _pf(ee))
        return dict(vf_idx=vi, vf_q=vq, rt_idx=ri, rt_q=rq, ee=ee)

    def fem_solve(self, pos, material, bend_idx, colors_strain, colors_bend, iterations, dt):
        """The reference's FEM::StrainSolver + FEM::IsometricBendingSolver on this solver's cloth, from predicted
        positions `pos`, with the given colours injected (row a25)."""
        pos, m = _f(pos).reshape(-1, 3), _f(material)
        bend, cs, cb = _u(bend_idx).reshape(-1, 4), _u(colors_strain), _u(colors_bend)
        out = np.zeros_like(pos)
        rc = lib().ref_fem_solve(self.h, _pf(pos), _pf(m), _pu(bend), len(bend), _pu(cs), int(cs.max()) + 1, _pu(cb),
                                 (int(cb.max()) + 1) if len(cb) else 0, int(iterations), float(dt), _pf(out))
        if rc:
            raise RuntimeError("ref_fem_solve failed")
        return out

    def traverse_pairs(self, self_mode=False, thickness=0.0):
        """BVH::UpdateData + Construct + Traverse on the current cloth / obstacle positions (row a16): rows
        (cloth face, other face) in the reference's order."""
        n = lib().ref_traverse_pairs(self.h, int(self_mode), float(thickness), None, 0)
        out = np.zeros((n, 2), np.uint32)
        if n:
            lib().ref_traverse_pairs(self.h, int(self_mode), float(thickness), _pu(out), n)
        return out

    def rtri(self):
        out = np.zeros(self.Fo, np.uint32)
        lib().ref_get_rtri(self.h, _pu(out), self.Fo)
        return out

    def build_obstacle_bvh(self):
        lib().ref_build_obstacle_bvh(self.h)

    def mortons(self):
        """Sorted Morton codes of the obstacle tree, from the reference's own ComputeMortons_Kernel."""
        out = np.zeros(self.Fo, np.uint32)
        lib().ref_get_mortons(self.h, _pu(out))
        return out

    def add_geodesics(self, src, tgt, rest):
        k = [_u(src), _u(tgt), _f(rest)]
        lib().ref_add_geodesics(self.h, _pu(k[0]), _pu(k[1]), _pf(k[2]), len(k[0]))

    def set_obstacle_state(self, x0=None, x=None, v=None):
        a = [None if t is None else _f(t) for t in (x0, x, v)]
        lib().ref_set_obstacle_state(self.h, *[_pf(t) for t in a])

    # ---- Quality mode ----
    def setup_quality(self, sc, q, strain_colors, bend_colors):
        """q: xrt_quality_params-like object.  Colours are injected into the reference's FEM solvers."""
        cs, cb = _u(strain_colors), _u(bend_colors)
        qp = np.array([q.enable_imminent_repulsion, q.enable_pbd_repulsion, int(cs.max()) + 1,
                       (int(cb.max()) + 1) if len(cb) else 0], np.int32)
        qf = np.array([q.xx_stiffness, q.yy_stiffness, q.xy_stiffness, q.xy_poisson_ratio, q.yx_poisson_ratio,
                       q.bending_stiffness, q.imminent_thickness, q.pbd_thickness, q.relaxation_rate, q.obstacle_mass,
                       q.impact_zone_thickness], np.float32)
        keep = [_u(sc.cloth_tri), _u(sc.cloth_edges), _u(sc.obs_tri), _u(sc.obs_edges), _u(sc.bend_idx)]
        lib().ref_setup_quality(self.h, _pi(qp), _pf(qf), _pu(keep[0]), _pu(keep[1]), _pu(keep[2]), _pu(keep[3]),
                                len(keep[3]), _pu(keep[4]), len(keep[4]), _pu(cs), _pu(cb))

    def step_frame_quality(self):
        return lib().ref_step_frame_quality(self.h)

    def find_impacts(self, thickness, dt):
        npairs = C.c_longlong(0)
        n = lib().ref_find_impacts(self.h, thickness, dt, None, None, 0, C.byref(npairs))
        nodes, twn = np.zeros((n, 4), np.uint32), np.zeros((n, 9), np.float32)
        if n:
            lib().ref_find_impacts(self.h, thickness, dt, _pu(nodes), _pf(twn), n, C.byref(npairs))
        return nodes, twn, int(npairs.value)

    def independent_impacts(self, deform=0):
        n = lib().ref_independent_impacts(self.h, deform, None, None, 0)
        nodes, twn = np.zeros((n, 4), np.uint32), np.zeros((n, 9), np.float32)
        if n:
            lib().ref_independent_impacts(self.h, deform, _pu(nodes), _pf(twn), n)
        return nodes, twn

    def collision_step(self, dt):
        lib().ref_collision_step(self.h, dt)

    def imminent_generate(self):
        n = lib().ref_imminent_generate(self.h, None, None, 0)
        nodes, bary = np.zeros((n, 4), np.uint32), np.zeros((n, 4), np.float32)
        if n:
            lib().ref_imminent_generate(self.h, _pu(nodes), _pf(bary), n)
        return nodes, bary

    def imminent_solve(self, dt, iteration=0):
        lib().ref_imminent_solve(self.h, dt, iteration)

    def rep_update_pairs(self):
        n = lib().ref_rep_update_pairs(self.h, None, 0)
        out = np.zeros((n, 2), np.uint32)
        if n:
            lib().ref_rep_update_pairs(self.h, _pu(out), n)
        return out

    def rep_solve(self, dt, iteration=0):
        lib().ref_rep_solve(self.h, dt, iteration)

    def quality_timing(self, thickness, dt, reps=3):
        """ms of the reference's ccd tree build, BVH::Update, the two Traverses and FindImpacts; pair and impact counts"""
        out, cnt = np.zeros(4, np.float32), np.zeros(2, np.int64)
        lib().ref_quality_timing(self.h, float(thickness), float(dt), int(reps), _pf(out), cnt.ctypes.data_as(C.POINTER(C.c_longlong)))
        return dict(tree_build_ms=float(out[0]), tree_update_ms=float(out[1]), traverse_ms=float(out[2]), find_impacts_ms=float(out[3]),
                    pairs=int(cnt[0]), impacts=int(cnt[1]))

    def rep_evaluate_only(self):
        V = self.V if hasattr(self, "V") else len(self.state(0)["x"])
        d, c = np.zeros((V, 3), np.float32), np.zeros(V, np.int32)
        lib().ref_rep_evaluate_only(self.h, _pf(d), _pi(c))
        return d, c

    def rep_contributions(self, vertex, cap=256):
        """(rows[n,6] = pair, slot, 4 node ids; corr[n,3]) of every PBD repulsion term on one cloth vertex (debug)"""
        rows, corr = np.zeros((cap, 6), np.int32), np.zeros((cap, 3), np.float32)
        n = lib().ref_rep_contributions(self.h, int(vertex), _pi(rows), _pf(corr), cap)
        return rows[:min(n, cap)], corr[:min(n, cap)]

    def obstacle_bvh_update(self):
        lib().ref_obstacle_bvh_update(self.h)

    def cloth_rtri(self):
        out = np.zeros(self.F, np.uint32)
        lib().ref_get_cloth_rtri(self.h, _pu(out), self.F)
        return out

    def close(self):
        if self.h:
            lib().ref_destroy(self.h)
            self.h = None


def time_frames(sc, steps, warmup, device=0, params=None, motion=None, bindings=None, ready=None):
    """Reference arm of bench.py: the reference's own Solver_Total (CUDA events around SimulateSwift) over `steps`
    frames after `warmup`, same workload.  Nothing of the product is loaded: scenes and defaults come from the
    `workloads` package.  ready: optional callable run after setup + warm-up (the start line of a concurrent run)."""
    from workloads import scene
    p = params or scene.swift_params()
    if p.particle_diameter <= 0:
        p.particle_diameter = sc.particle_diameter()
    r = RefSolver(sc, p, device=device)
    if bindings is not None:
        r.set_bindings(*bindings)
    if motion is None:
        motion = lambda f: scene.sphere_motion(sc.obs_pos, f + 1)  # noqa: E731
    ms = []
    wall0 = None
    t_warm = time.perf_counter()
    extra = 0
    while time.perf_counter() - t_warm < 1.0 and extra < 400:  # an idle GPU needs ~1 s of work to reach its boost clock
        r.set_obstacle_target(motion(extra))
        r.step_frame()
        extra += 1
    for f in range(warmup + steps):
        if f == warmup:
            if ready is not None:
                ready()
            wall0 = time.perf_counter()
        r.set_obstacle_target(motion(f))
        ms.append(r.step_frame())
    wall = time.perf_counter() - wall0
    ms = np.array(ms[warmup:])
    cs, cb, nsc, nbc = r.colors()
    r.close()
    timers = "stubbed" if os.environ.get("XRT_REF_NO_TIMERS") == "1" else "as shipped"
    return {"frames_per_s": steps / (ms.sum() * 1e-3), "ms_per_frame": float(ms.mean()),
            "ms_median": float(np.median(ms)), "ms_p10": float(np.percentile(ms, 10)), "ms_p90": float(np.percentile(ms, 90)),
            "wall_s": wall, "wall_frames_per_s": steps / wall, "cores": 1,
            "colors": [int(nsc), int(nbc)],
            "sample": f"{steps} frames, reference CUDA physics (sm_100a, --use_fast_math, per-kernel GPU timers {timers}) "
                      f"through the GL-free harness, Solver_Total event time; {nsc}/{nbc} stretch/bend colours"}
