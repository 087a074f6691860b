/*
 * xrt_b200.h — C ABI of the B200-native XRTailor Swift-mode cloth hot path.
 *
 * Drop-in boundary for the reference's solver component
 *   ClothSolverBase (xrtailor/include/xrtailor/pipeline/base/ClothSolverBase.hpp:55-189)
 * Every entry point names the reference interface it replaces.  Plain pointers
 * and sizes only; the caller owns every host array, the library copies.
 * Every call returns 0 on success, else a TAILOR_EXIT-compatible code
 * (xrtailor/include/xrtailor/core/ErrorDefs.hpp:3-35).  Nothing here calls
 * exit() or throws across the boundary.  There is NO CPU fallback: when no
 * CUDA device is usable xrt_create fails with XRT_ERR_CUDA.
 */
#ifndef XRT_B200_H_
#define XRT_B200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define XRT_OK 0
#define XRT_ERR_CUDA 8000          /* TAILOR_EXIT CUDA error                */
#define XRT_ERR_NODE_POOL 8001     /* node pool exhausted                   */
#define XRT_ERR_EDGE_POOL 8002
#define XRT_ERR_FACE_POOL 8003
#define XRT_ERR_INVALID 8100       /* bad argument / wrong call order       */
#define XRT_ERR_FRAME_RANGE 9000

typedef struct xrt_solver xrt_solver;

/* Mirrors the fields of SimParams the Swift path reads
 * (xrtailor/include/xrtailor/core/Common.hpp:31-116) plus the Swift config
 * keys of SimulationConfigParser (hash_cell_size, inter_leaved_hash). */
typedef struct xrt_params {
  int32_t num_substeps;          /* SimParams::num_substeps            */
  int32_t num_iterations;        /* SimParams::num_iterations          */
  int32_t num_collision_passes;  /* SimParams::num_collision_passes    */
  int32_t max_num_neighbors;     /* SimParams::max_num_neighbors (64)  */
  float gravity[3];              /* SimParams::gravity                 */
  float damping;                 /* SimParams::damping                 */
  float relaxation_factor;       /* SimParams::relaxation_factor       */
  float long_range_stretchiness; /* SimParams::long_range_stretchiness */
  float sdf_friction;            /* SimParams::sdf_friction            */
  float bvh_tolerance;           /* SimParams::bvh_tolerance           */
  float stretch_compliance;      /* per-cloth in ref, AddCloth arg      */
  float bend_compliance;
  float particle_diameter;       /* SimParams::particle_diameter       */
  float hash_cell_size;          /* swift.collision.self_contact       */
  int32_t inter_leaved_hash;     /* swift.collision.self_contact       */
  int32_t enable_self_collision; /* SimParams::enable_self_collision   */
  int32_t solve_bending;         /* SimParams::solve_bending           */
  int32_t geodesic_lra;          /* SimParams::geodesic_LRA            */
  float frame_dt;                /* Timer::FixedDeltaTime(), 1/60      */
  /* 0: IEEE division in the Morton quantiser (matches the CPU oracle bit
   * for bit); 1: div.approx like the reference's --use_fast_math build
   * (CMakeLists.txt:217), for bit parity against the reference's CUDA. */
  int32_t morton_fast_div;
  /* Launch shape of the substep.  0: one launch per stage and colour,
   * replayed as a CUDA graph.  C >= 1: one persistent kernel per substep on a
   * thread-block cluster of C CTAs (1, 2, 4, 8 or 16; one CTA per SM) with
   * hardware cluster barriers between the phases; 148 / C garments run
   * concurrently on a B200.  -1: the largest cluster the device schedules.
   * -2: every SM of the device on ONE garment (cooperative launch, grid-wide
   * software barrier), no concurrency.
   * -3: lowest latency for a single garment: a 16-CTA cluster keeps the predicted positions of
   * the whole garment in its distributed shared memory and runs the Gauss-Seidel phases between two collision
   * passes as ONE kernel (gathers are 215-cycle ld.shared::cluster, the phase boundary is the hardware cluster
   * barrier); the collision passes themselves (BVH refit + queries) run device-wide in between.  Garments up to
   * 131 072 vertices, proper colourings only.
   * All shapes call the same device bodies and give identical results. */
  int32_t persist_cluster;
  int32_t device;                /* CUDA device ordinal                */
} xrt_params;

/* Fills *p with the reference's Swift defaults
 * (examples/smpl/simulation_conf_smpl.json + Common.hpp HOST_INITs). */
int xrt_default_params(xrt_params* p);

/* Mesh as the reference's AddCloth/AddObstacle receive it
 * (pipeline/impl/universal/ClothSolver.cpp:189-350): world-space positions,
 * triangle indices, unique edges.  inv_mass may be NULL (cloth: 1, obstacle:
 * 1e-3, PhysicsMeshHelper.cu:139). */
typedef struct xrt_mesh_desc {
  const float* positions;   /* 3*n_verts  */
  const uint32_t* indices;  /* 3*n_faces  */
  const uint32_t* edges;    /* 2*n_edges (unique, undirected) */
  const float* inv_mass;    /* n_verts or NULL */
  uint32_t n_verts, n_faces, n_edges;
} xrt_mesh_desc;

/* Optional, once per process, before the host's first CUDA call: asks the driver for 32 hardware queues
 * (CUDA_DEVICE_MAX_CONNECTIONS, unless the host set it), which batches of more than 8 streams need to run side by
 * side.  The library changes nothing in the process environment unless this is called. */
int xrt_global_init(void);

/* ClothSolverBase lifetime: constructor / OnDestroy. */
int xrt_create(const xrt_params* p, xrt_solver** out);
void xrt_destroy(xrt_solver* s);
const char* xrt_last_error(const xrt_solver* s);

/* ClothSolverUniversal::AddCloth / AddObstacle
 * (impl/universal/ClothSolver.cpp:189,247).  Call once per garment / collider:
 * like the reference, the library merges them into one cloth and one obstacle
 * mesh with running index offsets.  *index_offset receives the first vertex
 * index of this mesh in the merged numbering (the reference's return value);
 * constraints, tethers and bindings added afterwards use merged indices. */
int xrt_add_cloth(xrt_solver* s, const xrt_mesh_desc* m, int32_t* index_offset);
int xrt_add_obstacle(xrt_solver* s, const xrt_mesh_desc* m, int32_t* index_offset);

/* StretchSolver::AddCloth / BendSolver::AddCloth
 * (physics/dynamics/BasicSolver.cu:36-62,101-117).  Constraints as produced by
 * PrepareBasicStretchData / GenerateBendingElements
 * (physics/ClothSolverHelper.cpp:4-109), or by xrt_build_constraints.  Each call
 * APPENDS one garment's constraints (indices in the merged numbering). */
int xrt_set_stretch(xrt_solver* s, const uint32_t* idx01, const float* rest, size_t n);
int xrt_set_bend(xrt_solver* s, const uint32_t* idx0123, const float* theta, size_t n);

/* ClothSolverBase::AddGeodesic (pipeline/base/ClothSolverBase.cpp:527-532). */
int xrt_add_geodesic(xrt_solver* s, const uint32_t* src, const uint32_t* tgt,
                     const float* rest, size_t n);

/* BindingConstraint::Add + Generate (physics/dynamics/BindingConstraint.cu:11-55,
 * called from ClothSolverSMPL::Start, pipeline/impl/smpl/ClothSolver.cpp): tether
 * cloth vertex cloth_idx[i] to the point bary[3i..] of the obstacle triangle with
 * vertex ids tri[3i..] (the reference's SkinParam).  distance[i] < 1e-6 means
 * "the distance at setup".  Solved one-sided after the contact solves of every
 * iteration (SolveBinding_Kernel, BindingConstraint.cu:56-92).  Before finalize. */
int xrt_add_binding(xrt_solver* s, const uint32_t* cloth_idx, const uint32_t* tri, const float* bary,
                    const float* stiffness, const float* distance, size_t n);

/* Parity injection of the reference's GraphColoring::colors
 * (physics/graph_coloring/Base.cuh:36).  kind 0 = stretch, 1 = bend.
 * Without it xrt_finalize_setup colours deterministically (greedy). */
int xrt_set_colors(xrt_solver* s, int kind, const uint32_t* colors, size_t n);

/* ClothSolverUniversal::Start (impl/universal/ClothSolver.cpp:46-93): builds
 * the obstacle BVH, RTriangle masks, spatial hash, colour-compacted
 * constraint lists; allocates every device buffer the step path needs. */
int xrt_finalize_setup(xrt_solver* s);

/* UpdateObstaclePredicted (physics/ClothSolverHelper.cu:81-97): target
 * obstacle vertex positions of the coming frame (host pointer, 3*n floats). */
int xrt_set_obstacle_target(xrt_solver* s, const float* xyz, size_t n_verts);

/* Pre-stage a whole obstacle animation in device memory (n_frames x n_verts
 * positions).  With a sequence set, every xrt_step_frame that has no explicit
 * xrt_set_obstacle_target pending takes the next frame of the sequence
 * (wrapping) with a device-to-device copy: the reference's SMPL pipeline keeps
 * its pose sequence resident the same way (impl/smpl/ClothSolver.cpp:239-249). */
int xrt_set_obstacle_sequence(xrt_solver* s, const float* xyz, size_t n_frames);

/* Body-animation producer on the device.  Replaces, for a skinned obstacle, the reference's
 * per-frame chain smplx::Body<>::Update (runtime/rag_doll/smpl/body.cpp:270-383, cuda/body.cu:212-313) ->
 * Body::verts() device-to-host -> UpdateObstaclePredicted's cudaMalloc + host-to-device + kernel + cudaFree
 * (physics/ClothSolverHelper.cu:81-97; call site impl/smpl/ClothSolver.cpp:244-249).  The rig stays resident; a
 * frame uploads 12 floats per joint (or nothing, with a resident pose sequence) and one kernel writes Node::x of the
 * obstacle: x = M * (sum_j w_vj (R_j v + t_j)), v = rest [+ pose_blend * (R - I)].  Call after xrt_finalize_setup.
 *   rest_positions : 3*n_verts, the SHAPED rest mesh (shape blend shapes are static and applied by the caller)
 *   weight_*       : skinning weights in CSR by vertex (outer n_verts+1, inner = joint index, values)
 *   joints         : 3*n_joints shaped joint positions (joint regressor * rest); parent[j] < j, parent[0] ignored
 *   model_matrix   : 16 floats row-major = model_matrix * align_matrix of UpdateObstaclePredicted, or NULL
 *   pose_blend     : optional [3*n_verts][9*(n_joints-1)] row-major pose blend shapes (body.cu:243-249), or NULL */
typedef struct xrt_rig_desc {
  const float* rest_positions;
  const uint32_t* weight_outer;
  const uint32_t* weight_inner;
  const float* weight_values;
  const float* joints;
  const int32_t* parent;
  const float* model_matrix;
  const float* pose_blend;
  uint32_t n_verts, n_joints;
} xrt_rig_desc;
int xrt_set_obstacle_rig(xrt_solver* s, const xrt_rig_desc* rig);
/* One frame: angle-axis pose (3*n_joints, joint 0 = root orientation) and root translation (3 or NULL), the
 * arguments of Sequence::set_pose + Body::Update.  Same effect as xrt_set_obstacle_target with the skinned mesh. */
int xrt_set_obstacle_pose(xrt_solver* s, const float* axis_angle, const float* trans);
/* A whole pose sequence resident on the device (n_frames x 3*n_joints, trans n_frames x 3 or NULL): every frame
 * without an explicit target or pose pending takes the next one (wrapping); no per-frame host-to-device traffic. */
int xrt_set_obstacle_pose_sequence(xrt_solver* s, const float* axis_angle, const float* trans, size_t n_frames);
/* n_joints x 12 joint transforms ([R | t] row-major) of the last xrt_set_obstacle_pose (Body::_local_to_global). */
int xrt_get_obstacle_joint_transforms(xrt_solver* s, float* out);
/* The obstacle's pending target positions Node::x (3*n_verts), whoever wrote them. */
int xrt_get_obstacle_target(xrt_solver* s, float* xyz);

/* ClothSolverBase::SimulateSwift (pipeline/base/ClothSolverBase.cpp:171-261).
 * solver_ms (may be NULL) receives the CUDA-event time of the frame, the
 * reference's "Solver_Total". */
int xrt_step_frame(xrt_solver* s, float* solver_ms);
/* Same, but enqueue only (no sync, no timing); pair with xrt_sync. */
int xrt_step_frame_async(xrt_solver* s);
int xrt_sync(xrt_solver* s);

/* Batch data generation: run `frames` frames on each of n independent
 * handles (same device), all streams released by one start event and joined
 * into one end event; *ms = device time between the two.  No host sync
 * between frames. */
int xrt_batch_run(xrt_solver** handles, int n, int frames, float* ms);

/* Persistent launch shape on the current device: *max_cluster = largest
 * cluster (CTAs) the device co-schedules, *resident = how many clusters of
 * `cluster` CTAs fit at once (= garments that step concurrently).  Either
 * out pointer may be NULL.  Fails with XRT_ERR_CUDA without a device. */
int xrt_persist_info(int cluster, int* max_cluster, int* resident);

/* Launch the named step kernel `reps` times on the handle's stream between
 * two CUDA events; *avg_ms = mean duration of ONE launch, *bytes = its
 * algorithmic bytes.  which: 0 bend colour pass, 1 stretch colour
 * pass, 2 collide particles, 3 contact generation (vertex), 4 predict, 5 the persistent substep kernel
 * (advances the state by one substep per launch). */
int xrt_profile_kernel(xrt_solver* s, int which, int reps, float* avg_ms, double* bytes);
/* The persistent substep kernel as it runs in batch mode: every handle
 * launches it `reps` times on its own stream between one start and one end
 * event.  *avg_ms = duration of one round (n launches running side by side on
 * disjoint clusters), *bytes = algorithmic bytes of those n launches. */
int xrt_batch_profile_substep(xrt_solver** handles, int n, int reps, float* avg_ms, double* bytes);

/* Shared-launch batch ("instances batched into shared launches").
 * n finalized handles on one device, all with the same persist_cluster >= 1
 * and substep count, step in lockstep: each substep of the whole batch is ONE
 * launch with one CTA (persist_cluster 1) or one cluster per garment, so the
 * batch is as wide as the device (148 garments at one SM each) instead of as
 * wide as the hardware queue count.  The handles stay usable on their own
 * (targets, sequences, getters); the batch keeps no state besides launch
 * tables.  Destroy the batch before its handles.  `groups` >= 1 splits the
 * garments into that many independently stepping sets (own stream and shared
 * launches each), so one set's per-garment work between substeps (obstacle
 * BVH build, predict + spatial hash) overlaps the other sets' substep kernels;
 * that per-garment work is spread over `workers` extra streams per set (0: the
 * set's own stream).  Keep groups * (1 + workers) <= 32, the hardware queues of
 * a device.  Everything a batch enqueues runs on the batch's streams; a
 * handle's own stream joins it on the next per-handle call.
 *   xrt_batch_step_frame_async  enqueue one SimulateSwift frame of every handle
 *   xrt_batch_sync              wait for all of it
 *   xrt_batch_step_frames       `frames` frames back to back; *ms = device time
 *   xrt_batch_profile           `reps` shared launches of one substep between two
 *                               events: *avg_ms per launch, *bytes algorithmic */
typedef struct xrt_batch xrt_batch;
int xrt_batch_create(xrt_solver** handles, int n, int groups, int workers, xrt_batch** out);
int xrt_batch_step_frame_async(xrt_batch* b);
int xrt_batch_sync(xrt_batch* b);
int xrt_batch_step_frames(xrt_batch* b, int frames, float* ms);
int xrt_batch_profile(xrt_batch* b, int reps, float* avg_ms, double* bytes);
/* xrt_get_positions / xrt_set_obstacle_target for every handle of the batch in
 * one call: rows of the handles back to back in batch order (xyz / nrm:
 * 3 * sum(n_verts) floats; n_verts_total is checked).  One garment's copy
 * overlaps the previous garment's memcpy into the caller's buffer. */
int xrt_batch_get_positions(xrt_batch* b, int which, float* xyz, float* nrm);
int xrt_batch_set_obstacle_targets(xrt_batch* b, const float* xyz, size_t n_verts_total);
/* Read-back mode.  on != 0: every frame ends with a device-side snapshot of each
 * garment's positions + normals that a copy stream brings to pinned host memory
 * while the next frame computes.  xrt_batch_get_positions(b, 0, ...) then returns
 * the OLDEST unread snapshot (frame f may be fetched after frame f+1 has been
 * enqueued); at most two frames may be outstanding - a third
 * xrt_batch_step_frame_async fails with XRT_ERR_INVALID until one is read.
 * xrt_batch_step_frames cannot be used in this mode for more than two frames. */
int xrt_batch_set_readback(xrt_batch* b, int on);
/* Kernels one frame of the batch launches (moved: with a new obstacle target). */
int xrt_batch_launches_per_frame(xrt_batch* b, int moved, uint32_t* out);
/* Kernels of this library the batch has enqueued since xrt_batch_create, counted where they are launched (substep
 * kernels, stage kernels, target unpack and read-back pack kernels), and the number of calls into cub's segmented sort. */
int xrt_batch_launch_count(xrt_batch* b, uint64_t* own_kernels, uint64_t* sort_calls);
void xrt_batch_destroy(xrt_batch* b);

/* SaveCache (impl/universal/ClothSolver.cpp:107-144): which 0 = cloth,
 * 1 = obstacle.  xyz / nrm: 3*n_verts floats each, either may be NULL. */
int xrt_get_positions(xrt_solver* s, int which, float* xyz, float* nrm);
/* Overwrite cloth state (x0 and v), for tests and batch seeding. */
int xrt_set_cloth_state(xrt_solver* s, const float* xyz, const float* vel);
/* Overwrite the predicted positions (Node::x) only, for parity tests. */
int xrt_set_cloth_predicted(xrt_solver* s, const float* xyz);

/* Building blocks exposed for parity tests; each one is the same device
 * code xrt_step_frame runs. */
int xrt_bvh_build_obstacle(xrt_solver* s);                 /* BVH::UpdateData+Construct, BVH.cu:565,618 */
int xrt_bvh_refit_obstacle(xrt_solver* s, int step);       /* Interpolate+midstep normals+RefitMidstep  */
int xrt_generate_contacts(xrt_solver* s);                  /* PBDCollision.cu:17,100,236                */
int xrt_hash_particles(xrt_solver* s);                     /* SpatialHashGPU::Hash                      */
int xrt_predict(xrt_solver* s);                            /* PredictPositions                          */
int xrt_solve_iteration(xrt_solver* s, int iteration);     /* one pass of the inner loop, no generate   */
int xrt_finalize(xrt_solver* s);                           /* Finalize                                  */

/* Debug getters (the reference's BVH::*Host / PhysicsMesh::Host* accessors,
 * BVH.cu:579-602).  *bytes in: capacity of dst, out: bytes written/needed.
 * dst == NULL only queries the size. */
enum xrt_debug_what {
  XRT_DBG_MORTON = 0,        /* uint32[F]  sorted Morton codes                 */
  XRT_DBG_SORTED_OBJECT = 1, /* uint32[F]  object index of leaf i              */
  XRT_DBG_NODES = 2,         /* uint32[4*(2F-1)] parent,left,right,object      */
  XRT_DBG_AABBS = 3,         /* float[6*(2F-1)]  lower xyz, upper xyz          */
  XRT_DBG_ESCAPE_INT = 4,    /* int32[F-1]                                     */
  XRT_DBG_ESCAPE_EXT = 5,    /* int32[F]                                       */
  XRT_DBG_VF_CONTACTS = 6,   /* {int32 index; float qs[3]}[V]                  */
  XRT_DBG_RT_CONTACTS = 7,   /* {int32 index; float qc[3]}[V]                  */
  XRT_DBG_EE_CONTACTS = 8,   /* {float dist,p2[3],q2[3],s,t,fn[3]}[E]          */
  XRT_DBG_PREDICTED = 9,     /* float[3V] cloth x                              */
  XRT_DBG_VELOCITY = 10,     /* float[3V] cloth v                              */
  XRT_DBG_COLORS_STRETCH = 11,
  XRT_DBG_COLORS_BEND = 12,
  XRT_DBG_NEIGHBORS = 13,    /* uint32[max_nb*V] strided by V like the ref     */
  XRT_DBG_RTRI = 14,         /* uint32[Fo] representative-triangle masks       */
  XRT_DBG_OBSTACLE_PRED = 15,/* float[9*Fo] primitive pred1..3                 */
  XRT_DBG_STATS = 16,        /* xrt_stats                                      */
  XRT_DBG_PARTICLE_HASH = 17,/* uint32[V] sorted hash; then uint32[V] index    */
  XRT_DBG_VF_CAND_COUNT = 18,/* uint32[V] #leaf hits of the VF point query     */
  XRT_DBG_RT_CAND_COUNT = 19,/* uint32[V] #leaf hits of the vertex-ray query   */
  XRT_DBG_EE_CAND_COUNT = 20,/* uint32[E] #leaf hits of the edge query         */
  /* uint64[16] ns per phase kind of the persistent kernel since setup (empty
   * unless XRT_PERSIST_PROF=1 was set at xrt_finalize_setup): 0 predict,
   * 1 collide, 2 apply, 3 interpolate+normals, 4 refit, 5 generate, 6 stretch,
   * 7 bend, 8 LRA, 9 vertex solve, 10 edge solve, 11 apply, 12 finalize */
  XRT_DBG_PERSIST_PROF = 21
};
int xrt_debug_get(xrt_solver* s, int what, void* dst, size_t* bytes);

/* Per-vertex / per-edge candidate sets of the three broad-phase queries in
 * traversal order (the content of the reference's per-thread buffer[256],
 * BVH.cu:957-1055).  kind 0 = VF point, 1 = vertex ray, 2 = edge segment.
 * offsets: uint32[n+1] (CSR), items: uint32[offsets[n]].  Two-call protocol:
 * call with items == NULL to get *n_items. */
int xrt_debug_candidates(xrt_solver* s, int kind, uint32_t* offsets,
                         uint32_t* items, size_t* n_items);

/* Constraint-residual and penetration statistics (north_star's third gate). */
typedef struct xrt_stats {
  double stretch_rms;       /* RMS of (|xi-xj|-rest)/rest over stretch constraints */
  double stretch_max;
  double bend_rms;          /* RMS of (angle-theta0)                               */
  uint32_t vf_penetrations; /* vertices with n.(x-qs) < 0 on their VF contact      */
  uint32_t n_stretch_colors, n_bend_colors;
  uint32_t launches_per_frame; /* kernels enqueued by one xrt_step_frame         */
  double bytes_per_frame;      /* algorithmic bytes of one frame     */
} xrt_stats;

/* Deterministic greedy colouring (replaces GraphColoring::Paint's random
 * palette, graph_coloring/Base.cu:163-317, in O(n)).  arity 2 (stretch), 3 (FEM strain: triangles) or 4 (bend). */
int xrt_greedy_color(const uint32_t* idx, size_t n, int arity, uint32_t n_verts,
                     uint32_t* colors, uint32_t* n_colors);

/* Setup-time derivations of the reference on an index buffer: the unique undirected edges
 * (sorted by (lower vertex, higher vertex)), the face -> edge table (edge j of a face joins its vertices j and j+1;
 * what the reference's mesh loader hands to Face::SetEdges), the edge -> face adjacency (BuildEFAdjacency,
 * physics/ClothSolverHelper.cpp:111-125), one stretch constraint per edge (PrepareBasicStretchData, :90-109) and
 * one bend element per pair of faces on an edge (GenerateBendingElements, :4-88: [0], [1] the edge in the winding
 * of the lower-index face, [2] its opposite vertex, [3] the other face's; theta = acos(clamp(n1 . n2))).
 * O(F log F) on the host without vcglib's half-edge mesh; the arrays are allocated by the library and released
 * with xrt_free_constraints.  Feed them to xrt_add_cloth (edges), xrt_set_stretch and xrt_set_bend. */
typedef struct xrt_constraints {
  uint32_t n_edges, n_bend;
  uint32_t* edges;       /* 2 n_edges                       */
  uint32_t* face_edges;  /* 3 n_faces                       */
  float* stretch_rest;   /* n_edges                         */
  uint32_t* bend_idx;    /* 4 n_bend                        */
  float* bend_theta;     /* n_bend                          */
  uint32_t* ef_offsets;  /* n_edges + 1 (CSR)               */
  uint32_t* ef_faces;    /* ef_offsets[n_edges], ascending  */
} xrt_constraints;
int xrt_build_constraints(const float* positions, uint32_t n_verts, const uint32_t* indices, uint32_t n_faces,
                          xrt_constraints* out);
void xrt_free_constraints(xrt_constraints* c);

/* Standalone LBVH build over triangles (BVH::UpdateData(ccd=false)+Construct)
 * for parity tests and the 200k-face scaling config.  Outputs as above. */
int xrt_bvh_build_standalone(const float* positions, uint32_t n_verts,
                             const uint32_t* indices, uint32_t n_faces,
                             int morton_fast_div, int device,
                             uint32_t* morton, uint32_t* sorted_object,
                             uint32_t* nodes, float* aabbs,
                             int32_t* esc_int, int32_t* esc_ext, float* build_ms);

/* Streaming OBJ-sequence writer, host threads only.  Replaces
 * ObjSequenceExporter::ExportObjSequence (src/runtime/export/ObjSequenceExporter.cpp:155-250), which runs after the
 * simulation over frames the solver kept in host memory (ClothSolver.cpp:107-144).  Frames are pushed as they come
 * out of xrt_get_positions / xrt_batch_get_positions and written by `threads` workers (default 2) while the GPU
 * computes the next ones; push copies the frame and returns, and blocks only when `queue_frames` (default 4) frames
 * are waiting.  Files are <dir>/%06d.obj, byte for byte the reference's SaveObj output (:21-46 with normals,
 * :83-103 without): header, "o mesh", "v" lines with 6 decimals, ["vn" lines], "s 1", 1-based faces.
 * indices: 3*n_faces (each < n_verts, checked); xyz / nrm: 3*n_verts per frame (nrm required iff with_normals).
 * close drains the queue, joins the workers and reports how many files and bytes were written. */
typedef struct xrt_obj_writer xrt_obj_writer;
int xrt_obj_writer_open(const char* dir, const uint32_t* indices, uint32_t n_faces, uint32_t n_verts,
                        int with_normals, int queue_frames, int threads, xrt_obj_writer** out);
int xrt_obj_writer_push(xrt_obj_writer* w, const float* xyz, const float* nrm);
int xrt_obj_writer_close(xrt_obj_writer* w, uint64_t* frames_written, uint64_t* bytes_written);

/* Morton codes of a triangle mesh (leaf boxes -> scene box -> 30-bit codes, the first stage of the tree build)
 * WITHOUT ANY ALLOCATION.  No counterpart in the reference: added for callers
 * that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
 * The stage's scratch (scene box, leaf boxes, object indices: 28 bytes per face) is a file-scope
 * `static __device__` array of 256 + 28 * XRT_NOALLOC_MAX_FACES bytes (7.0 MiB) that every GPU loading the library
 * holds until exit.  The call allocates and frees nothing, does not synchronise, and only enqueues 3-4 kernels on
 * `stream` (a cudaStream_t, NULL = default stream) of the CURRENT device.
 *   d_positions_xyzw : DEVICE, n_verts * 4 floats (x, y, z, w ignored -- the library's float4 layout)
 *   d_indices        : DEVICE, 3 * n_faces vertex indices (each < n_verts; not checked)
 *   d_morton         : DEVICE out, n_faces codes in object (face) order; equal to xrt_bvh_build_standalone's
 *                      `morton` after that call's own sort: morton_sorted[i] == d_morton[sorted_object[i]]
 *   d_scene_box      : DEVICE out or NULL, 6 floats (min xyz, max xyz over all leaf boxes)
 *   err, err_cap     : optional buffer for a message when the call fails
 * n_faces > XRT_NOALLOC_MAX_FACES fails with XRT_ERR_INVALID and a message; so do host pointers.
 * Every call shares the one array per device, which keeps the last call's contents (the call itself resets what
 * must start fresh): calls that overlap on the GPU -- two streams, two host threads -- race on it; serialise them.
 * The first call on a device looks up the array's device address (cudaGetSymbolAddress, once per device); make that
 * first call outside stream capture. */
#define XRT_NOALLOC_MAX_FACES 262144u
int xrt_morton_codes_noalloc(const float* d_positions_xyzw, uint32_t n_verts,
                             const uint32_t* d_indices, uint32_t n_faces,
                             int morton_fast_div, void* stream,
                             uint32_t* d_morton, float* d_scene_box,
                             char* err, size_t err_cap);

/* Quality-mode cloth model on the GPU: orthotropic StVK strain
 * per triangle and isometric bending per hinge as coloured Gauss-Seidel passes,
 * `iterations` times [every strain colour, then every bending colour], the inner
 * loop of ClothSolverBase::SimulateQuality (ClothSolverBase.cpp:315-324).
 * Replaces FEM::StrainSolver::Init/Solve and FEM::IsometricBendingSolver::Init/
 * Solve (physics/dynamics/FEMSolver.cu:28-83, FEMHelper.cu:10-388).
 * rest_pos / pos / out_pos: 3*n_verts floats; inv_mass: n_verts; tri: 3*n_faces;
 * bend_idx: 4*n_bend (may be NULL with n_bend 0); material: xx, yy, xy
 * stiffness, xy and yx Poisson ratio, bending stiffness; colors_*: one colour
 * per constraint (any colouring: an improper one falls back to float atomics). */
int xrt_fem_solve_standalone(const float* rest_pos, const float* pos, const float* inv_mass,
                             uint32_t n_verts, const uint32_t* tri, uint32_t n_faces,
                             const uint32_t* bend_idx, uint32_t n_bend, const float* material,
                             const uint32_t* colors_strain, const uint32_t* colors_bend,
                             int iterations, float dt, int device, float* out_pos,
                             float* ms_per_iteration);

/* Face-face broad phase between two triangle meshes (F_r == 0: the left mesh
 * against itself).  Replaces BVH::Traverse / TraverseSelf with
 * CountPairs_Kernel + thrust::inclusive_scan + FindPairs_Kernel
 * (xrtailor/src/physics/broad_phase/lbvh/BVH.cu:1149-1340, callers
 * :1342-1400): both trees are built here (UpdateData(ccd=false)+Construct),
 * each left leaf box is dilated by `thickness` (Bounds.cu:95-128) and walked
 * through the right tree.  pairs: uint32[2*capacity] rows (left face, right
 * face) in the reference's order.  *n_pairs is always the full count; call
 * with pairs == NULL to size the buffer (XRT_ERR_INVALID if capacity is too
 * small).  traverse_ms times count + scan + fill on the device. */
int xrt_traverse_pairs_standalone(const float* pos_l, uint32_t nv_l,
                                  const uint32_t* tri_l, uint32_t F_l,
                                  const float* pos_r, uint32_t nv_r,
                                  const uint32_t* tri_r, uint32_t F_r,
                                  float thickness, int morton_fast_div,
                                  int device, uint32_t* pairs, size_t capacity,
                                  size_t* n_pairs, float* traverse_ms);

/* ------------------------------------------------------------------------------------------------
 * Quality mode: ClothSolverBase::SimulateQuality (pipeline/base/ClothSolverBase.cpp:263-362).
 * A handle on which xrt_set_quality was called before xrt_finalize_setup steps Quality frames:
 * xrt_step_frame then runs   cloth BVH build -> imminent-repulsion proximity + quadratures ->
 * obstacle BVH Update(ccd) -> per substep [predict; num_iterations x (imminent solve, PBD
 * repulsion with pairs refreshed every 10th iteration, FEM strain colours, isometric-bending
 * colours, LRA, binding); CollisionStep (cubic CCD, independent impacts, impact zones, augmented-
 * Lagrangian zone optimiser); finalize].  num_substeps / num_iterations / gravity / damping come
 * from the xrt_params struct, which stands for quality.solver of the reference's config (1 substep, 200 iterations).
 * ---------------------------------------------------------------------------------------------- */

/* QualityModeSettings (include/xrtailor/config/Base.hpp:47-57,108-127) */
typedef struct xrt_quality_params {
  float xx_stiffness, yy_stiffness, xy_stiffness;  /* FEMFabricSettings                        */
  float xy_poisson_ratio, yx_poisson_ratio;
  float bending_stiffness;
  int32_t enable_imminent_repulsion;               /* RepulsionSettings                        */
  float imminent_thickness;
  int32_t enable_pbd_repulsion;
  float pbd_thickness;
  float relaxation_rate;
  float obstacle_mass;                             /* ImpactZoneSettings                       */
  float impact_zone_thickness;
  /* arena sizes (every list of the collision path lives in preallocated device memory; an overflow
   * makes the call fail with XRT_ERR_FACE_POOL and names the arena).  0 = default.               */
  uint32_t max_pairs;       /* face-face pairs of one broad phase; default max(1 << 20, 256 F); quadratures share
                             * this bound, repulsion items get four times it                                  */
  uint32_t max_candidates;  /* narrow-phase candidates (<= 15 per pair); default 4 * max_pairs    */
  uint32_t max_impacts;     /* impacts of one FindImpacts; default max(1 << 20, max_pairs / 2)    */
} xrt_quality_params;
int xrt_default_quality_params(xrt_quality_params* q);

/* Turns the handle into a Quality-mode solver.  Before xrt_finalize_setup, after xrt_add_cloth and
 * xrt_set_bend (the hinges double as the isometric-bending elements, like the reference builds both
 * from GenerateBendingElements; bend colours are the ones of xrt_set_colors(kind 1)).
 * strain_colors: one colour per cloth face for FEM::StrainSolver (graph_coloring/impl/
 * FEMStrainColoring.cu), or NULL for the deterministic greedy colouring. */
int xrt_set_quality(xrt_solver* s, const xrt_quality_params* q, const uint32_t* strain_colors);

/* Result of the last CollisionStep (impact_zone/Entry.cu:140-256) */
typedef struct xrt_collision_stats {
  int32_t rounds;          /* FindImpacts rounds that found something                            */
  int32_t success;         /* ended with no impact left                                          */
  int32_t total_impacts;   /* independent impacts handed to the zones                            */
  int32_t deform;          /* 1: the rigid pass failed and the obstacle was allowed to deform    */
  int32_t n_pairs;         /* face pairs of the last broad phase                                 */
  int32_t n_candidates;    /* narrow-phase tests that passed the box filters, last round         */
  int32_t outer_iterations, line_search_steps; /* of the last zone solve (slowest zone)          */
} xrt_collision_stats;

/* The stages of a Quality frame one by one, for parity tests; each is the device code xrt_step_frame
 * runs.  Impacts travel as rows: nodes uint32[4n] (cloth node i = i, obstacle node j = V + j) and
 * twn float[9n] = t, w[4], n[3], is_vf. */
int xrt_quality_begin_frame(xrt_solver* s);  /* cloth BVH UpdateData(ccd=false)+Construct, ClothSolverBase.cpp:277-278 */
/* ImminentRepulsion::UpdateProximity + Generate (repulsion/ImminentRepulsion.cu:21-98).  nodes/bary may be
 * NULL; *n_quads is always the full count. */
int xrt_quality_imminent_generate(xrt_solver* s, uint32_t* nodes, float* bary, size_t capacity, size_t* n_quads);
int xrt_quality_imminent_solve(xrt_solver* s);                      /* ImminentRepulsion::Solve, :153-169   */
/* PBDRepulsionSolver::UpdatePairs (repulsion/PBDRepulsion.cu:22-33): pairs uint32[2n] (cloth face, other
 * face; obstacle faces offset by F) */
int xrt_quality_rep_update_pairs(xrt_solver* s, uint32_t* pairs, size_t capacity, size_t* n_pairs);
int xrt_quality_rep_solve(xrt_solver* s);                           /* SolveNoProximity, :470-516           */
int xrt_quality_obstacle_bvh_update(xrt_solver* s);                 /* obstacle_bvh_->Update(ccd), BVH.cu:836 */
int xrt_quality_iteration(xrt_solver* s, int iteration);            /* ClothSolverBase.cpp:296-328          */
/* FindImpacts on fresh ccd trees (impact_zone/ImpactHelper.cu:216-243, Entry.cu:146-155) */
int xrt_quality_find_impacts(xrt_solver* s, float thickness, uint32_t* nodes, float* twn, size_t capacity,
                             size_t* n_impacts, size_t* n_pairs, size_t* n_candidates);
/* independentImpacts (Entry.cu:25-70) on a caller-supplied list; returns the kept rows in sorted order */
int xrt_quality_independent_impacts(xrt_solver* s, const uint32_t* nodes, const float* twn, size_t n, int deform,
                                    uint32_t* out_nodes, float* out_twn, size_t* n_out);
/* IsolateImpacts (impact_zone/ImpactZoneHelper.cu:227-334): zone colour per impact, -1 beyond 100 zones */
int xrt_quality_isolate_zones(xrt_solver* s, const uint32_t* nodes, size_t n, int32_t* colors, int32_t* n_zones);
/* ParallelImpactZoneOptimizer ctor + ParallelOptimizer::Solve (ParallelOptimizer.cu:513-890) on a caller-
 * supplied impact list with ascending zone colours; moves the predicted positions.  stats: outer iterations
 * and line-search steps of the slowest zone, all zones converged. */
int xrt_quality_optimize_zones(xrt_solver* s, const uint32_t* nodes, const float* twn, const int32_t* colors,
                               size_t n, int n_zones, int deform, float thickness, float obstacle_mass,
                               int32_t* stats3);
int xrt_quality_collision_step(xrt_solver* s, xrt_collision_stats* stats);  /* CollisionStep, Entry.cu:140-256 */
/* CUDA-event time of one stage on the handle's current state, mean of `reps` runs.  what: 0 cloth BVH build
 * (UpdateData + Construct), 1 BVH::Update(ccd), 2 ccd tree build, 3 face-pair broad phase (cloth x obstacle + cloth x
 * cloth on the ccd trees), 4 CCD narrow phase on those pairs (filter, expand, cubic tests, compaction), 5 FindImpacts
 * as a whole.  *count: faces (0-2), pairs (3), candidates that reached the cubic solver (4), impacts (5). */
int xrt_quality_profile(xrt_solver* s, int what, int reps, float* avg_ms, double* count);
/* uint32[F] representative-triangle masks of the cloth (ClothSolverBase::SetupRTriangle) */
int xrt_quality_cloth_rtri(xrt_solver* s, uint32_t* out, size_t n_faces);
/* obstacle predicted positions (Node::x), float[3 Vo]: the deform pass of CollisionStep moves them */
int xrt_quality_get_obstacle_predicted(xrt_solver* s, float* xyz);

#ifdef __cplusplus
}
#endif
#endif /* XRT_B200_H_ */
