// C ABI of the B200 Swift-mode cloth path (include/xrt_b200.h) and the host
// driver that restates ClothSolverBase::SimulateSwift
// (xrtailor/src/pipeline/base/ClothSolverBase.cpp:171-261) as one CUDA graph
// per frame.  No CPU fallback lives here: without a CUDA device xrt_create fails.
#include "../../include/xrt_b200.h"
#include "xrt_solver.cuh"

#include <algorithm>
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <vector>

struct xrt_solver {
  xrt_params p{};
  char err[512] = {0};
  cudaStream_t st = nullptr;
  cudaEvent_t ev0 = nullptr, ev1 = nullptr;
  bool finalized = false;
  // host staging (caller data is copied, never retained)
  std::vector<float> c_pos, c_invm, o_pos, o_invm;
  std::vector<uint32_t> c_tri, c_edge, o_tri, o_edge;
  std::vector<uint32_t> s_idx, b_idx, s_color, b_color, l_src, l_tgt;
  std::vector<float> s_rest, b_theta, l_rest;
  bool s_color_set = false, b_color_set = false;
  // colour ranges over the permuted constraint arrays
  std::vector<uint32_t> s_off, b_off;
  uint32_t *d_s_off = nullptr, *d_b_off = nullptr;
  bool atomic_s = false, atomic_b = false;
  // device state
  ClothDev c;
  ObstacleDev o;
  DevBVH bvh{};
  BvhWork bw;
  StepConst k{};
  bool target_pending = false;  // xrt_set_obstacle_target called since the last frame
  cudaGraphExec_t graph[2] = {nullptr, nullptr};  // [obstacle moved?]
  uint32_t launches[2] = {0, 0};
  uint32_t launch_counter = 0;
  float4* stage = nullptr;  // pinned host staging for obstacle targets / readback
  size_t stage_elems = 0;
  float4* d_stage = nullptr;
};

static void xrt_set_error(xrt_solver* s, const char* fmt, ...) {
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
static cudaError_t dev_alloc(T** p, size_t n) {
  return cudaMalloc((void**)p, (n ? n : 1) * sizeof(T));
}
template <class T>
static cudaError_t dev_upload(T** p, const std::vector<T>& h) {
  cudaError_t e = dev_alloc(p, h.size());
  if (e == cudaSuccess && !h.empty()) e = cudaMemcpy(*p, h.data(), h.size() * sizeof(T), cudaMemcpyHostToDevice);
  return e;
}

extern "C" int xrt_default_params(xrt_params* p) {
  if (!p) return XRT_ERR_INVALID;
  memset(p, 0, sizeof(*p));
  p->num_substeps = 4;           // examples/smpl/simulation_conf_smpl.json
  p->num_iterations = 10;
  p->num_collision_passes = 10;  // Common.hpp:101
  p->max_num_neighbors = 64;     // Common.hpp:36
  p->gravity[0] = 0.f; p->gravity[1] = -9.8f; p->gravity[2] = 0.f;
  p->damping = 0.98f;
  p->relaxation_factor = 0.25f;
  p->long_range_stretchiness = 1.0f;
  p->sdf_friction = 0.1f;
  p->bvh_tolerance = 0.003f;     // config/Base.hpp:131 default (F10 in SURVEY.md)
  p->stretch_compliance = 0.0f;
  p->bend_compliance = 10.0f;
  p->particle_diameter = 0.0f;   // 0: derive from the first edge like ClothObject.hpp:36-37
  p->hash_cell_size = 1.5f;
  p->inter_leaved_hash = 3;
  p->enable_self_collision = 1;
  p->solve_bending = 1;
  p->geodesic_lra = 1;
  p->frame_dt = 1.0f / 60.0f;
  p->morton_fast_div = 1;
  p->fused_iterations = 0;
  p->device = 0;
  return XRT_OK;
}

extern "C" int xrt_create(const xrt_params* p, xrt_solver** out) {
  if (!p || !out) return XRT_ERR_INVALID;
  *out = nullptr;
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev <= 0 || p->device < 0 || p->device >= ndev) {
    (void)cudaGetLastError();
    return XRT_ERR_CUDA;  // no CPU fallback
  }
  if (cudaSetDevice(p->device) != cudaSuccess) return XRT_ERR_CUDA;
  xrt_solver* s = new xrt_solver();
  s->p = *p;
  if (cudaStreamCreateWithFlags(&s->st, cudaStreamNonBlocking) != cudaSuccess ||
      cudaEventCreate(&s->ev0) != cudaSuccess || cudaEventCreate(&s->ev1) != cudaSuccess) {
    delete s;
    return XRT_ERR_CUDA;
  }
  *out = s;
  return XRT_OK;
}

extern "C" const char* xrt_last_error(const xrt_solver* s) { return s ? s->err : "null solver"; }

static int copy_mesh(xrt_solver* s, const xrt_mesh_desc* m, std::vector<float>& pos, std::vector<float>& invm,
                     std::vector<uint32_t>& tri, std::vector<uint32_t>& edge, float default_inv_mass) {
  XRT_REQUIRE(m && m->positions && m->n_verts > 0, "mesh: positions missing");
  XRT_REQUIRE(m->n_faces == 0 || m->indices, "mesh: indices missing");
  XRT_REQUIRE(m->n_edges == 0 || m->edges, "mesh: edges missing");
  if (m->n_verts > 1000000u) {  // MemoryPool.cuh:12-14 pool sizes
    xrt_set_error(s, "node pool exhausted (%u > 1000000)", m->n_verts);
    return XRT_ERR_NODE_POOL;
  }
  if (m->n_edges > 3000000u) { xrt_set_error(s, "edge pool exhausted"); return XRT_ERR_EDGE_POOL; }
  if (m->n_faces > 2000000u) { xrt_set_error(s, "face pool exhausted"); return XRT_ERR_FACE_POOL; }
  for (size_t i = 0; i < 3 * (size_t)m->n_faces; i++) XRT_REQUIRE(m->indices[i] < m->n_verts, "mesh: face index out of range");
  for (size_t i = 0; i < 2 * (size_t)m->n_edges; i++) XRT_REQUIRE(m->edges[i] < m->n_verts, "mesh: edge index out of range");
  pos.assign(m->positions, m->positions + 3 * (size_t)m->n_verts);
  if (m->inv_mass) invm.assign(m->inv_mass, m->inv_mass + m->n_verts);
  else invm.assign(m->n_verts, default_inv_mass);
  tri.assign(m->indices, m->indices + 3 * (size_t)m->n_faces);
  edge.assign(m->edges, m->edges + 2 * (size_t)m->n_edges);
  return XRT_OK;
}

extern "C" int xrt_add_cloth(xrt_solver* s, const xrt_mesh_desc* m, int32_t* index_offset) {
  if (!s) return XRT_ERR_INVALID;
  XRT_REQUIRE(!s->finalized, "add_cloth after finalize_setup");
  XRT_REQUIRE(s->c_pos.empty(), "one cloth mesh per handle (merge meshes on the caller side)");
  int rc = copy_mesh(s, m, s->c_pos, s->c_invm, s->c_tri, s->c_edge, 1.0f);
  if (rc == XRT_OK && index_offset) *index_offset = 0;
  return rc;
}

extern "C" int xrt_add_obstacle(xrt_solver* s, const xrt_mesh_desc* m, int32_t* index_offset) {
  if (!s) return XRT_ERR_INVALID;
  XRT_REQUIRE(!s->finalized, "add_obstacle after finalize_setup");
  XRT_REQUIRE(s->o_pos.empty(), "one obstacle mesh per handle (merge meshes on the caller side)");
  int rc = copy_mesh(s, m, s->o_pos, s->o_invm, s->o_tri, s->o_edge, 1.0f / 1e3f);  // PhysicsMeshHelper.cu:139
  if (rc == XRT_OK && index_offset) *index_offset = 0;
  return rc;
}

extern "C" int xrt_set_stretch(xrt_solver* s, const uint32_t* idx01, const float* rest, size_t n) {
  if (!s) return XRT_ERR_INVALID;
  XRT_REQUIRE(!s->finalized, "set_stretch after finalize_setup");
  XRT_REQUIRE(n == 0 || (idx01 && rest), "set_stretch: null array");
  s->s_idx.assign(idx01, idx01 + 2 * n);
  s->s_rest.assign(rest, rest + n);
  s->s_color_set = false;
  return XRT_OK;
}

extern "C" int xrt_set_bend(xrt_solver* s, const uint32_t* idx0123, const float* theta, size_t n) {
  if (!s) return XRT_ERR_INVALID;
  XRT_REQUIRE(!s->finalized, "set_bend after finalize_setup");
  XRT_REQUIRE(n == 0 || (idx0123 && theta), "set_bend: null array");
  s->b_idx.assign(idx0123, idx0123 + 4 * n);
  s->b_theta.assign(theta, theta + n);
  s->b_color_set = false;
  return XRT_OK;
}

extern "C" int xrt_add_geodesic(xrt_solver* s, const uint32_t* src, const uint32_t* tgt, const float* rest, size_t n) {
  if (!s) return XRT_ERR_INVALID;
  XRT_REQUIRE(!s->finalized, "add_geodesic after finalize_setup");
  XRT_REQUIRE(n == 0 || (src && tgt && rest), "add_geodesic: null array");
  s->l_src.insert(s->l_src.end(), src, src + n);
  s->l_tgt.insert(s->l_tgt.end(), tgt, tgt + n);
  s->l_rest.insert(s->l_rest.end(), rest, rest + n);
  return XRT_OK;
}

extern "C" int xrt_set_colors(xrt_solver* s, int kind, const uint32_t* colors, size_t n) {
  if (!s) return XRT_ERR_INVALID;
  XRT_REQUIRE(!s->finalized, "set_colors after finalize_setup");
  XRT_REQUIRE(kind == 0 || kind == 1, "set_colors: kind must be 0 (stretch) or 1 (bend)");
  XRT_REQUIRE(colors || n == 0, "set_colors: null array");
  if (kind == 0) {
    XRT_REQUIRE(n == s->s_rest.size(), "set_colors: size differs from the stretch constraint count");
    s->s_color.assign(colors, colors + n);
    s->s_color_set = true;
  } else {
    XRT_REQUIRE(n == s->b_theta.size(), "set_colors: size differs from the bend constraint count");
    s->b_color.assign(colors, colors + n);
    s->b_color_set = true;
  }
  return XRT_OK;
}
