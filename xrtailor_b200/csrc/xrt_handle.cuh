// The solver handle behind the C ABI (include/xrt_b200.h) and the small host helpers every
// translation unit of the library shares.  One handle = one CUDA stream on one device, no globals
// (the reference keeps its state in ClothSolverBase members and process-wide Global::sim_params,
// xrtailor/include/xrtailor/pipeline/base/ClothSolverBase.hpp:55-189, core/Global.hpp:22-25).
#pragma once
#include "../../include/xrt_b200.h"
#include "xrt_persist.cuh"

#include <cstdarg>
#include <cstdio>
#include <vector>

struct RigState;      // body-animation producer (xrt_rig.cu); null without xrt_set_obstacle_rig
struct QualityState;  // Quality-mode state (xrt_quality.cu); null for a Swift-mode handle

struct xrt_solver {
  xrt_params p{};
  char err[512] = {0};
  cudaStream_t st = nullptr;
  cudaStream_t st2 = nullptr;  // side branch of the per-stage frame graph (edge queries, next pass's obstacle interpolation)
  cudaEvent_t ev_fork = nullptr, ev_prep = nullptr, ev_side = nullptr, ev_gv = nullptr;
  float4* pred_alt = nullptr;  // second copy of the obstacle primitives: pass k + 1 is refitted while pass k's contacts are solved
  DevBVH frame_bvh{};          // the tree as the pass being enqueued sees it (pred = this pass's copy)
  int frame_passes = 0;
  bool side_open = false;      // the side stream has joined the capture of this frame
  cudaEvent_t ev0 = nullptr, ev1 = nullptr;
  bool finalized = false;
  // host staging (caller data is copied, never retained)
  std::vector<float> c_pos, c_invm, o_pos, o_invm;
  std::vector<uint32_t> c_tri, c_edge, o_tri, o_edge;
  std::vector<uint32_t> s_idx, b_idx, s_color, b_color, l_src, l_tgt;
  std::vector<float> s_rest, b_theta, l_rest;
  std::vector<uint32_t> bd_vert, bd_tri;
  std::vector<float> bd_bary, bd_k, bd_dist;
  bool s_color_set = false, b_color_set = false;
  // colour ranges over the permuted constraint arrays
  std::vector<uint32_t> s_off, b_off;
  uint32_t *d_s_off = nullptr, *d_b_off = nullptr;
  uint32_t *d_s_cta_off = nullptr, *d_b_cta_off = nullptr;  // persist_cluster -3: [colours][17] per-CTA ranges
  int dsmem_shift = -1;
  bool atomic_s = false, atomic_b = false;
  // device state
  ClothDev c;
  ObstacleDev o;
  DevBVH bvh{};
  BvhWork bw;
  StepConst k{};
  bool target_pending = false;  // xrt_set_obstacle_target called since the last frame
  cudaGraphExec_t graph[2] = {nullptr, nullptr};  // [obstacle moved?]
  // shared-launch batches: this garment's own work before substep `sub` ([2 * sub + moved]; null = nothing to do)
  // and after the last substep ([moved]), each one graph launch
  std::vector<cudaGraphExec_t> pre_graph;
  cudaGraphExec_t post_graph[2] = {nullptr, nullptr};
  cudaEvent_t batch_ev = nullptr;  // a batch enqueued work on this garment: the handle's stream joins it on next use
  bool touched = true;             // the handle's own stream has work the batch has not joined yet
  uint32_t launches[2] = {0, 0};
  uint32_t launch_counter = 0;
  float4* stage = nullptr;  // pinned host staging for obstacle targets / readback
  size_t stage_elems = 0;
  float4* d_stage = nullptr;
  // per-frame transfers: raw float3 arrays through pinned memory, converted on the device
  float *h_up = nullptr, *d_up = nullptr;  // [3 Vo] obstacle target upload
  float *h_dn = nullptr, *d_dn = nullptr;  // [6 max(V, Vo)] positions + normals download
  cudaEvent_t ev_up = nullptr;             // upload buffer free again
  float4* d_seq = nullptr;  // pre-staged obstacle animation [seq_frames][Vo]
  size_t seq_frames = 0, seq_cursor = 0;
  cudaEvent_t ev_join = nullptr;
  int cluster = 0;  // resolved persist_cluster (0: per-stage launches)
  unsigned long long* d_grid_bar = nullptr;  // grid-wide barrier counter (persist_cluster -2)
  unsigned long long* d_prof = nullptr;  // [16] per-phase-kind ns of the persistent kernel (XRT_PERSIST_PROF=1)
  QualityState* q = nullptr;             // set by xrt_set_quality: the handle runs SimulateQuality
  RigState* rig = nullptr;               // set by xrt_set_obstacle_rig: the obstacle target is skinned on the device
};

static inline void xrt_set_error(xrt_solver* s, const char* fmt, ...) {
  if (!s) return;
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(s->err, sizeof(s->err), fmt, ap);
  va_end(ap);
}

#define XRT_REQUIRE(cond, msg)          \
  do {                                  \
    if (!(cond)) {                      \
      xrt_set_error(s, "%s", msg);      \
      return XRT_ERR_INVALID;           \
    }                                   \
  } while (0)

template <class T>
static inline cudaError_t dev_alloc(T** p, size_t n) {
  return cudaMalloc((void**)p, (n ? n : 1) * sizeof(T));
}
template <class T>
static inline cudaError_t dev_upload(T** p, const std::vector<T>& h) {
  cudaError_t e = dev_alloc(p, h.size());
  if (e == cudaSuccess && !h.empty()) e = cudaMemcpy(*p, h.data(), h.size() * sizeof(T), cudaMemcpyHostToDevice);
  return e;
}

// Work a shared-launch batch enqueued for this garment lives on the batch's streams; the
// handle's own stream joins it lazily, on the first per-handle call (a wait parked in every
// handle stream would sit in the shared hardware queues and couple unrelated batches).
static inline cudaError_t join_batch(xrt_solver* s) {
  s->touched = true;
  if (!s->batch_ev) return cudaSuccess;
  cudaEvent_t e = s->batch_ev;
  s->batch_ev = nullptr;
  return cudaStreamWaitEvent(s->st, e, 0);
}
#define XRT_READY()                                            \
  if (!s) return XRT_ERR_INVALID;                              \
  XRT_REQUIRE(s->finalized, "call xrt_finalize_setup first");  \
  XRT_CUDA_OK(cudaSetDevice(s->p.device));                     \
  XRT_CUDA_OK(join_batch(s))


// setup-time host helpers (xrt_setup.cu)
void xrt_color_ranges(const std::vector<uint32_t>& idx, int arity, const std::vector<uint32_t>& color, uint32_t n_verts,
                      std::vector<uint32_t>& perm, std::vector<uint32_t>& off, bool& improper);
void xrt_build_vf(const std::vector<uint32_t>& tri, uint32_t V, std::vector<uint32_t>& off, std::vector<uint32_t>& idx);
void xrt_build_rtri(const std::vector<uint32_t>& tri, const std::vector<uint32_t>& vf_off,
                    const std::vector<uint32_t>& vf_idx, std::vector<uint32_t>& rtri);
// frame path pieces of xrt_api.cu that the batch code (xrt_batch.cu) shares
PersistArgs xrt_make_persist_args(const xrt_solver* s, int substep, bool moved, bool do_predict);
int xrt_take_sequence_frame(xrt_solver* s, cudaStream_t st);
double xrt_frame_bytes(const xrt_solver* s, bool moved);  // algorithmic bytes of one frame, SURVEY 8(d)
double xrt_bvh_build_bytes(const xrt_solver* s);
// body-animation producer (xrt_rig.cu)
int xrt_rig_take_sequence_frame(xrt_solver* s, cudaStream_t st);
void xrt_rig_destroy(xrt_solver* s);
// Quality mode (xrt_quality.cu)
int xrt_quality_finalize(xrt_solver* s);           // called by xrt_finalize_setup when xrt_set_quality was called
int xrt_quality_step_frame(xrt_solver* s, float* solver_ms);  // ClothSolverBase::SimulateQuality
void xrt_quality_destroy(xrt_solver* s);
