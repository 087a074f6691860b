"""ctypes binding of include/xrt_b200.h.  Loads the in-tree libxrt_b200.so and
fails loudly when it is missing — there is no other implementation to fall
back to."""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
# XRT_B200_LIB=fast selects the --use_fast_math build (A/B timing only)
LIB_PATH = os.path.join(_HERE, "libxrt_b200_fast.so" if os.environ.get("XRT_B200_LIB") == "fast" else "libxrt_b200.so")


class XrtError(RuntimeError):
    def __init__(self, code, msg=""):
        super().__init__(f"xrt_b200 error {code}: {msg}")
        self.code = code


class XrtRigDesc(C.Structure):
    _fields_ = [
        ("rest_positions", C.POINTER(C.c_float)), ("weight_outer", C.POINTER(C.c_uint32)),
        ("weight_inner", C.POINTER(C.c_uint32)), ("weight_values", C.POINTER(C.c_float)),
        ("joints", C.POINTER(C.c_float)), ("parent", C.POINTER(C.c_int32)),
        ("model_matrix", C.POINTER(C.c_float)), ("pose_blend", C.POINTER(C.c_float)),
        ("n_verts", C.c_uint32), ("n_joints", C.c_uint32),
    ]


class XrtParams(C.Structure):
    _fields_ = [
        ("num_substeps", C.c_int32), ("num_iterations", C.c_int32),
        ("num_collision_passes", C.c_int32), ("max_num_neighbors", C.c_int32),
        ("gravity", C.c_float * 3), ("damping", C.c_float), ("relaxation_factor", C.c_float),
        ("long_range_stretchiness", C.c_float), ("sdf_friction", C.c_float), ("bvh_tolerance", C.c_float),
        ("stretch_compliance", C.c_float), ("bend_compliance", C.c_float), ("particle_diameter", C.c_float),
        ("hash_cell_size", C.c_float), ("inter_leaved_hash", C.c_int32), ("enable_self_collision", C.c_int32),
        ("solve_bending", C.c_int32), ("geodesic_lra", C.c_int32), ("frame_dt", C.c_float),
        ("morton_fast_div", C.c_int32), ("persist_cluster", C.c_int32), ("device", C.c_int32),
    ]


class XrtMeshDesc(C.Structure):
    _fields_ = [
        ("positions", C.POINTER(C.c_float)), ("indices", C.POINTER(C.c_uint32)),
        ("edges", C.POINTER(C.c_uint32)), ("inv_mass", C.POINTER(C.c_float)),
        ("n_verts", C.c_uint32), ("n_faces", C.c_uint32), ("n_edges", C.c_uint32),
    ]


class XrtStats(C.Structure):
    _fields_ = [
        ("stretch_rms", C.c_double), ("stretch_max", C.c_double), ("bend_rms", C.c_double),
        ("vf_penetrations", C.c_uint32), ("n_stretch_colors", C.c_uint32), ("n_bend_colors", C.c_uint32),
        ("launches_per_frame", C.c_uint32), ("bytes_per_frame", C.c_double),
    ]


class XrtQualityParams(C.Structure):
    _fields_ = [
        ("xx_stiffness", C.c_float), ("yy_stiffness", C.c_float), ("xy_stiffness", C.c_float),
        ("xy_poisson_ratio", C.c_float), ("yx_poisson_ratio", C.c_float), ("bending_stiffness", C.c_float),
        ("enable_imminent_repulsion", C.c_int32), ("imminent_thickness", C.c_float),
        ("enable_pbd_repulsion", C.c_int32), ("pbd_thickness", C.c_float), ("relaxation_rate", C.c_float),
        ("obstacle_mass", C.c_float), ("impact_zone_thickness", C.c_float),
        ("max_pairs", C.c_uint32), ("max_candidates", C.c_uint32), ("max_impacts", C.c_uint32),
    ]


class XrtConstraints(C.Structure):
    _fields_ = [
        ("n_edges", C.c_uint32), ("n_bend", C.c_uint32), ("edges", C.POINTER(C.c_uint32)),
        ("face_edges", C.POINTER(C.c_uint32)), ("stretch_rest", C.POINTER(C.c_float)),
        ("bend_idx", C.POINTER(C.c_uint32)), ("bend_theta", C.POINTER(C.c_float)),
        ("ef_offsets", C.POINTER(C.c_uint32)), ("ef_faces", C.POINTER(C.c_uint32)),
    ]


class XrtCollisionStats(C.Structure):
    _fields_ = [
        ("rounds", C.c_int32), ("success", C.c_int32), ("total_impacts", C.c_int32), ("deform", C.c_int32),
        ("n_pairs", C.c_int32), ("n_candidates", C.c_int32), ("outer_iterations", C.c_int32),
        ("line_search_steps", C.c_int32),
    ]


# every symbol include/xrt_b200.h declares
SYMBOLS = [
    "xrt_default_params", "xrt_create", "xrt_destroy", "xrt_last_error", "xrt_add_cloth", "xrt_add_obstacle",
    "xrt_set_stretch", "xrt_set_bend", "xrt_add_geodesic", "xrt_set_colors", "xrt_finalize_setup",
    "xrt_set_obstacle_target", "xrt_set_obstacle_sequence", "xrt_batch_run", "xrt_profile_kernel", "xrt_step_frame", "xrt_step_frame_async", "xrt_sync", "xrt_get_positions",
    "xrt_set_cloth_state", "xrt_set_cloth_predicted", "xrt_bvh_build_obstacle", "xrt_bvh_refit_obstacle", "xrt_generate_contacts",
    "xrt_hash_particles", "xrt_predict", "xrt_solve_iteration", "xrt_finalize", "xrt_debug_get",
    "xrt_debug_candidates", "xrt_greedy_color", "xrt_bvh_build_standalone", "xrt_persist_info", "xrt_batch_profile_substep", "xrt_add_binding",
    "xrt_traverse_pairs_standalone", "xrt_fem_solve_standalone", "xrt_batch_create", "xrt_batch_step_frame_async", "xrt_batch_sync",
    "xrt_batch_step_frames", "xrt_batch_profile", "xrt_batch_destroy", "xrt_batch_get_positions",
    "xrt_batch_set_obstacle_targets", "xrt_batch_launches_per_frame", "xrt_batch_launch_count", "xrt_batch_set_readback",
    "xrt_build_constraints", "xrt_free_constraints", "xrt_global_init", "xrt_morton_codes_noalloc", "xrt_obj_writer_open", "xrt_obj_writer_push", "xrt_obj_writer_close", "xrt_set_obstacle_rig", "xrt_set_obstacle_pose",
    "xrt_set_obstacle_pose_sequence", "xrt_get_obstacle_joint_transforms", "xrt_get_obstacle_target",
    "xrt_default_quality_params", "xrt_set_quality", "xrt_quality_begin_frame", "xrt_quality_imminent_generate",
    "xrt_quality_imminent_solve", "xrt_quality_rep_update_pairs", "xrt_quality_rep_solve",
    "xrt_quality_obstacle_bvh_update", "xrt_quality_iteration", "xrt_quality_find_impacts",
    "xrt_quality_independent_impacts", "xrt_quality_isolate_zones", "xrt_quality_optimize_zones",
    "xrt_quality_collision_step", "xrt_quality_profile", "xrt_quality_cloth_rtri", "xrt_quality_get_obstacle_predicted",
]


def _load():
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            "(nvcc, sm_100a).  xrtailor_b200 has no CPU or PyTorch fallback.")
    L = C.CDLL(LIB_PATH)
    vp, u32p, f32p, i32p = C.c_void_p, C.POINTER(C.c_uint32), C.POINTER(C.c_float), C.POINTER(C.c_int32)
    sig = {
        "xrt_default_params": [C.POINTER(XrtParams)],
        "xrt_create": [C.POINTER(XrtParams), C.POINTER(vp)],
        "xrt_add_cloth": [vp, C.POINTER(XrtMeshDesc), i32p],
        "xrt_add_obstacle": [vp, C.POINTER(XrtMeshDesc), i32p],
        "xrt_set_stretch": [vp, u32p, f32p, C.c_size_t],
        "xrt_set_bend": [vp, u32p, f32p, C.c_size_t],
        "xrt_add_geodesic": [vp, u32p, u32p, f32p, C.c_size_t],
        "xrt_set_colors": [vp, C.c_int, u32p, C.c_size_t],
        "xrt_finalize_setup": [vp],
        "xrt_set_obstacle_target": [vp, f32p, C.c_size_t],
        "xrt_set_obstacle_sequence": [vp, f32p, C.c_size_t],
        "xrt_batch_run": [C.POINTER(vp), C.c_int, C.c_int, f32p],
        "xrt_profile_kernel": [vp, C.c_int, C.c_int, f32p, C.POINTER(C.c_double)],
        "xrt_step_frame": [vp, f32p],
        "xrt_step_frame_async": [vp],
        "xrt_sync": [vp],
        "xrt_get_positions": [vp, C.c_int, f32p, f32p],
        "xrt_set_cloth_state": [vp, f32p, f32p],
        "xrt_set_cloth_predicted": [vp, f32p],
        "xrt_bvh_build_obstacle": [vp],
        "xrt_bvh_refit_obstacle": [vp, C.c_int],
        "xrt_generate_contacts": [vp],
        "xrt_hash_particles": [vp],
        "xrt_predict": [vp],
        "xrt_solve_iteration": [vp, C.c_int],
        "xrt_finalize": [vp],
        "xrt_debug_get": [vp, C.c_int, vp, C.POINTER(C.c_size_t)],
        "xrt_debug_candidates": [vp, C.c_int, u32p, u32p, C.POINTER(C.c_size_t)],
        "xrt_greedy_color": [u32p, C.c_size_t, C.c_int, C.c_uint32, u32p, u32p],
        "xrt_batch_profile_substep": [C.POINTER(vp), C.c_int, C.c_int, f32p, C.POINTER(C.c_double)],
        "xrt_add_binding": [vp, u32p, u32p, f32p, f32p, f32p, C.c_size_t],
        "xrt_persist_info": [C.c_int, i32p, i32p],
        "xrt_bvh_build_standalone": [f32p, C.c_uint32, u32p, C.c_uint32, C.c_int, C.c_int, u32p, u32p, u32p, f32p,
                                     i32p, i32p, f32p],
        "xrt_set_obstacle_rig": [vp, C.POINTER(XrtRigDesc)],
        "xrt_set_obstacle_pose": [vp, f32p, f32p],
        "xrt_set_obstacle_pose_sequence": [vp, f32p, f32p, C.c_size_t],
        "xrt_get_obstacle_joint_transforms": [vp, f32p],
        "xrt_get_obstacle_target": [vp, f32p],
        "xrt_obj_writer_open": [C.c_char_p, u32p, C.c_uint32, C.c_uint32, C.c_int, C.c_int, C.c_int, C.POINTER(vp)],
        "xrt_obj_writer_push": [vp, f32p, f32p],
        "xrt_obj_writer_close": [vp, C.POINTER(C.c_uint64), C.POINTER(C.c_uint64)],
        "xrt_morton_codes_noalloc": [vp, C.c_uint32, vp, C.c_uint32, C.c_int, vp, vp, vp, C.c_char_p, C.c_size_t],
        "xrt_traverse_pairs_standalone": [f32p, C.c_uint32, u32p, C.c_uint32, f32p, C.c_uint32, u32p, C.c_uint32,
                                          C.c_float, C.c_int, C.c_int, u32p, C.c_size_t, C.POINTER(C.c_size_t), f32p],
        "xrt_fem_solve_standalone": [f32p, f32p, f32p, C.c_uint32, u32p, C.c_uint32, u32p, C.c_uint32, f32p, u32p, u32p,
                                     C.c_int, C.c_float, C.c_int, f32p, f32p],
        "xrt_batch_create": [C.POINTER(vp), C.c_int, C.c_int, C.c_int, C.POINTER(vp)],
        "xrt_batch_step_frame_async": [vp],
        "xrt_batch_sync": [vp],
        "xrt_batch_step_frames": [vp, C.c_int, f32p],
        "xrt_batch_profile": [vp, C.c_int, f32p, C.POINTER(C.c_double)],
        "xrt_batch_get_positions": [vp, C.c_int, f32p, f32p],
        "xrt_batch_set_obstacle_targets": [vp, f32p, C.c_size_t],
        "xrt_batch_launch_count": [vp, C.POINTER(C.c_uint64), C.POINTER(C.c_uint64)],
        "xrt_batch_launches_per_frame": [vp, C.c_int, u32p],
        "xrt_batch_set_readback": [vp, C.c_int],
        "xrt_build_constraints": [f32p, C.c_uint32, u32p, C.c_uint32, C.POINTER(XrtConstraints)],
        "xrt_default_quality_params": [C.POINTER(XrtQualityParams)],
        "xrt_set_quality": [vp, C.POINTER(XrtQualityParams), u32p],
        "xrt_quality_begin_frame": [vp],
        "xrt_quality_imminent_generate": [vp, u32p, f32p, C.c_size_t, C.POINTER(C.c_size_t)],
        "xrt_quality_imminent_solve": [vp],
        "xrt_quality_rep_update_pairs": [vp, u32p, C.c_size_t, C.POINTER(C.c_size_t)],
        "xrt_quality_rep_solve": [vp],
        "xrt_quality_obstacle_bvh_update": [vp],
        "xrt_quality_iteration": [vp, C.c_int],
        "xrt_quality_find_impacts": [vp, C.c_float, u32p, f32p, C.c_size_t, C.POINTER(C.c_size_t), C.POINTER(C.c_size_t),
                                     C.POINTER(C.c_size_t)],
        "xrt_quality_independent_impacts": [vp, u32p, f32p, C.c_size_t, C.c_int, u32p, f32p, C.POINTER(C.c_size_t)],
        "xrt_quality_isolate_zones": [vp, u32p, C.c_size_t, i32p, i32p],
        "xrt_quality_optimize_zones": [vp, u32p, f32p, i32p, C.c_size_t, C.c_int, C.c_int, C.c_float, C.c_float, i32p],
        "xrt_quality_collision_step": [vp, C.POINTER(XrtCollisionStats)],
        "xrt_quality_profile": [vp, C.c_int, C.c_int, f32p, C.POINTER(C.c_double)],
        "xrt_quality_cloth_rtri": [vp, u32p, C.c_size_t],
        "xrt_quality_get_obstacle_predicted": [vp, f32p],
    }
    for name, args in sig.items():
        fn = getattr(L, name)
        fn.argtypes = args
        fn.restype = C.c_int
    L.xrt_destroy.argtypes = [vp]
    L.xrt_destroy.restype = None
    L.xrt_batch_destroy.argtypes = [vp]
    L.xrt_batch_destroy.restype = None
    L.xrt_free_constraints.argtypes = [C.POINTER(XrtConstraints)]
    L.xrt_free_constraints.restype = None
    L.xrt_last_error.argtypes = [vp]
    L.xrt_last_error.restype = C.c_char_p
    L.xrt_global_init.argtypes = []
    L.xrt_global_init.restype = C.c_int
    L.xrt_global_init()  # the tests and the bench step batches of up to 32 streams: ask for the queues before any CUDA call
    return L


lib = _load()


def default_quality_params() -> XrtQualityParams:
    q = XrtQualityParams()
    rc = lib.xrt_default_quality_params(C.byref(q))
    if rc:
        raise XrtError(rc)
    return q


def default_params() -> XrtParams:
    p = XrtParams()
    rc = lib.xrt_default_params(C.byref(p))
    if rc:
        raise XrtError(rc)
    return p
