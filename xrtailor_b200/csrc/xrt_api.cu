// C ABI of the B200 Swift-mode cloth path (include/xrt_b200.h) and the host
// driver that restates ClothSolverBase::SimulateSwift
// (xrtailor/src/pipeline/base/ClothSolverBase.cpp:171-261) as one CUDA graph
// per frame.  No CPU fallback lives here: without a CUDA device xrt_create fails.
#include "../../include/xrt_b200.h"
#include "xrt_handle.cuh"

#include <algorithm>
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

// Batch mode steps many handles side by side, every stream a chain of dependent launches.  Streams map onto
// CUDA_DEVICE_MAX_CONNECTIONS hardware queues (default 8); two chains in one queue serialise behind each other's
// dependencies, which capped a batch at 8 garments in flight however many clusters were free
// (profiles/r01h_hardware_queue_finding.txt).  The variable is read when the CUDA context is created, so a host that
// wants 32 queues calls this once before its first CUDA call; a value the host already set wins.  Nothing happens at
// load time any more (round 1 did this in a library constructor).
extern "C" int xrt_global_init(void) {
  return setenv("CUDA_DEVICE_MAX_CONNECTIONS", "32", 0) == 0 ? XRT_OK : XRT_ERR_INVALID;
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
  p->relaxation_factor = 0.9f;      // examples/smpl/simulation_conf_smpl.json FABRIC
  p->long_range_stretchiness = 1.1f;
  p->sdf_friction = 0.1f;
  p->bvh_tolerance = 0.003f;     // config/Base.hpp:131 default (F10 in SURVEY.md)
  p->stretch_compliance = 0.0f;
  p->bend_compliance = 100.0f;      // config/Base.hpp:40
  p->particle_diameter = 0.0f;   // 0: derive from the first edge like ClothObject.hpp:36-37
  p->hash_cell_size = 1.5f;
  p->inter_leaved_hash = 3;
  p->enable_self_collision = 1;
  p->solve_bending = 1;
  p->geodesic_lra = 1;
  p->frame_dt = 1.0f / 60.0f;
  p->morton_fast_div = 1;
  p->persist_cluster = 8;
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
      cudaStreamCreateWithFlags(&s->st2, cudaStreamNonBlocking) != cudaSuccess ||
      cudaEventCreateWithFlags(&s->ev_fork, cudaEventDisableTiming) != cudaSuccess ||
      cudaEventCreateWithFlags(&s->ev_prep, cudaEventDisableTiming) != cudaSuccess ||
      cudaEventCreateWithFlags(&s->ev_side, cudaEventDisableTiming) != cudaSuccess ||
      cudaEventCreateWithFlags(&s->ev_gv, cudaEventDisableTiming) != cudaSuccess ||
      cudaEventCreate(&s->ev0) != cudaSuccess || cudaEventCreate(&s->ev1) != cudaSuccess) {
    delete s;
    return XRT_ERR_CUDA;
  }
  *out = s;
  return XRT_OK;
}

extern "C" const char* xrt_last_error(const xrt_solver* s) { return s ? s->err : "null solver"; }

// Appends mesh m to the merged mesh (pos / invm / tri / edge), its indices shifted by the vertices already there:
// the reference merges every garment into one PhysicsMesh the same way (RegisterNewMesh with a running offset,
// impl/universal/ClothSolver.cpp:247-350) and returns that offset to the caller.
static int copy_mesh(xrt_solver* s, const xrt_mesh_desc* m, std::vector<float>& pos, std::vector<float>& invm,
                     std::vector<uint32_t>& tri, std::vector<uint32_t>& edge, float default_inv_mass, int32_t* index_offset) {
  XRT_REQUIRE(m && m->positions && m->n_verts > 0, "mesh: positions missing");
  XRT_REQUIRE(m->n_faces == 0 || m->indices, "mesh: indices missing");
  XRT_REQUIRE(m->n_edges == 0 || m->edges, "mesh: edges missing");
  const size_t v0 = invm.size();
  if (v0 + m->n_verts > 1000000u) {  // MemoryPool.cuh:12-14 pool sizes
    xrt_set_error(s, "node pool exhausted (%zu > 1000000)", v0 + m->n_verts);
    return XRT_ERR_NODE_POOL;
  }
  if (edge.size() / 2 + m->n_edges > 3000000u) { xrt_set_error(s, "edge pool exhausted"); return XRT_ERR_EDGE_POOL; }
  if (tri.size() / 3 + m->n_faces > 2000000u) { xrt_set_error(s, "face pool exhausted"); return XRT_ERR_FACE_POOL; }
  for (size_t i = 0; i < 3 * (size_t)m->n_faces; i++) XRT_REQUIRE(m->indices[i] < m->n_verts, "mesh: face index out of range");
  for (size_t i = 0; i < 2 * (size_t)m->n_edges; i++) XRT_REQUIRE(m->edges[i] < m->n_verts, "mesh: edge index out of range");
  pos.insert(pos.end(), m->positions, m->positions + 3 * (size_t)m->n_verts);
  if (m->inv_mass) invm.insert(invm.end(), m->inv_mass, m->inv_mass + m->n_verts);
  else invm.insert(invm.end(), m->n_verts, default_inv_mass);
  for (size_t i = 0; i < 3 * (size_t)m->n_faces; i++) tri.push_back(m->indices[i] + (uint32_t)v0);
  for (size_t i = 0; i < 2 * (size_t)m->n_edges; i++) edge.push_back(m->edges[i] + (uint32_t)v0);
  if (index_offset) *index_offset = (int32_t)v0;
  return XRT_OK;
}

extern "C" int xrt_add_cloth(xrt_solver* s, const xrt_mesh_desc* m, int32_t* index_offset) {
  if (!s) return XRT_ERR_INVALID;
  XRT_REQUIRE(!s->finalized, "add_cloth after finalize_setup");
  return copy_mesh(s, m, s->c_pos, s->c_invm, s->c_tri, s->c_edge, 1.0f, index_offset);
}

extern "C" int xrt_add_obstacle(xrt_solver* s, const xrt_mesh_desc* m, int32_t* index_offset) {
  if (!s) return XRT_ERR_INVALID;
  XRT_REQUIRE(!s->finalized, "add_obstacle after finalize_setup");
  return copy_mesh(s, m, s->o_pos, s->o_invm, s->o_tri, s->o_edge, 1.0f / 1e3f, index_offset);  // PhysicsMeshHelper.cu:139
}

extern "C" int xrt_set_stretch(xrt_solver* s, const uint32_t* idx01, const float* rest, size_t n) {
  if (!s) return XRT_ERR_INVALID;
  XRT_REQUIRE(!s->finalized, "set_stretch after finalize_setup");
  XRT_REQUIRE(n == 0 || (idx01 && rest), "set_stretch: null array");
  s->s_idx.insert(s->s_idx.end(), idx01, idx01 + 2 * n);  // one call per garment, indices already in the merged numbering
  s->s_rest.insert(s->s_rest.end(), rest, rest + n);
  s->s_color_set = false;
  return XRT_OK;
}

extern "C" int xrt_set_bend(xrt_solver* s, const uint32_t* idx0123, const float* theta, size_t n) {
  if (!s) return XRT_ERR_INVALID;
  XRT_REQUIRE(!s->finalized, "set_bend after finalize_setup");
  XRT_REQUIRE(n == 0 || (idx0123 && theta), "set_bend: null array");
  s->b_idx.insert(s->b_idx.end(), idx0123, idx0123 + 4 * n);
  s->b_theta.insert(s->b_theta.end(), theta, theta + n);
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

extern "C" int xrt_add_binding(xrt_solver* s, const uint32_t* cloth_idx, const uint32_t* tri, const float* bary,
                               const float* stiffness, const float* distance, size_t n) {
  if (!s) return XRT_ERR_INVALID;
  XRT_REQUIRE(!s->finalized, "add_binding after finalize_setup");
  XRT_REQUIRE(n == 0 || (cloth_idx && tri && bary && stiffness && distance), "add_binding: null array");
  s->bd_vert.insert(s->bd_vert.end(), cloth_idx, cloth_idx + n);
  s->bd_tri.insert(s->bd_tri.end(), tri, tri + 3 * n);
  s->bd_bary.insert(s->bd_bary.end(), bary, bary + 3 * n);
  s->bd_k.insert(s->bd_k.end(), stiffness, stiffness + n);
  s->bd_dist.insert(s->bd_dist.end(), distance, distance + n);
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

static std::vector<float4> to_float4(const std::vector<float>& pos, const std::vector<float>& w) {
  std::vector<float4> out(w.size());
  for (size_t i = 0; i < w.size(); i++) out[i] = make_float4(pos[3 * i], pos[3 * i + 1], pos[3 * i + 2], w[i]);
  return out;
}

static void fill_step_const(xrt_solver* s) {
  const xrt_params& p = s->p;
  StepConst& k = s->k;
  k.dt = p.frame_dt / (float)p.num_substeps;
  k.gx = p.gravity[0]; k.gy = p.gravity[1]; k.gz = p.gravity[2];
  k.damping = p.damping;
  k.relaxation = p.relaxation_factor;
  k.lra_stretchiness = p.long_range_stretchiness;
  k.friction = p.sdf_friction;
  k.stretch_compliance = p.stretch_compliance;
  k.bend_compliance = p.bend_compliance;
  k.particle_diameter = p.particle_diameter;
  k.cell_spacing = p.particle_diameter * p.hash_cell_size;  // SpatialHashGPU ctor, SpatialHashGPU.cu:161-167
  k.table_size = 2 * (int)s->c.V;
  k.max_neighbors = p.max_num_neighbors;
}

extern "C" int xrt_finalize_setup(xrt_solver* s) {
  if (!s) return XRT_ERR_INVALID;
  XRT_REQUIRE(!s->finalized, "finalize_setup called twice");
  XRT_REQUIRE(!s->c_pos.empty(), "finalize_setup: no cloth");
  XRT_REQUIRE(s->p.num_substeps > 0 && s->p.num_iterations > 0 && s->p.num_collision_passes > 0,
              "finalize_setup: substeps, iterations and collision passes must be positive");
  XRT_CUDA_OK(cudaSetDevice(s->p.device));
  ClothDev& c = s->c;
  ObstacleDev& o = s->o;
  c.V = (uint32_t)s->c_invm.size();
  c.F = (uint32_t)(s->c_tri.size() / 3);
  c.E = (uint32_t)(s->c_edge.size() / 2);
  for (uint32_t v : s->s_idx) XRT_REQUIRE(v < c.V, "stretch index out of range");
  for (uint32_t v : s->b_idx) XRT_REQUIRE(v < c.V, "bend index out of range");
  for (uint32_t v : s->l_src) XRT_REQUIRE(v < c.V, "geodesic src out of range");
  for (uint32_t v : s->l_tgt) XRT_REQUIRE(v < c.V, "geodesic tgt out of range");
  // particle diameter like ClothObjectUniversal::Start (impl/universal/ClothObject.hpp:36-37):
  // |p[idx0]-p[idx1]| of the first triangle; PARTICLE_DIAMETER scaling is the caller's
  if (s->p.particle_diameter <= 0.f && c.F) {
    const float* a = &s->c_pos[3 * s->c_tri[0]];
    const float* b = &s->c_pos[3 * s->c_tri[1]];
    s->p.particle_diameter = sqrtf((a[0] - b[0]) * (a[0] - b[0]) + (a[1] - b[1]) * (a[1] - b[1]) + (a[2] - b[2]) * (a[2] - b[2]));
  }
  // --- colours ---
  const uint32_t ns = (uint32_t)s->s_rest.size(), nb = (uint32_t)s->b_theta.size();
  if (!s->s_color_set) {
    s->s_color.resize(ns);
    if (xrt_greedy_color(s->s_idx.data(), ns, 2, c.V, s->s_color.data(), nullptr)) {
      xrt_set_error(s, "stretch colouring failed");
      return XRT_ERR_INVALID;
    }
  }
  if (!s->b_color_set) {
    s->b_color.resize(nb);
    if (xrt_greedy_color(s->b_idx.data(), nb, 4, c.V, s->b_color.data(), nullptr)) {
      xrt_set_error(s, "bend colouring failed");
      return XRT_ERR_INVALID;
    }
  }
  std::vector<uint32_t> s_perm, b_perm;
  xrt_color_ranges(s->s_idx, 2, s->s_color, c.V, s_perm, s->s_off, s->atomic_s);
  xrt_color_ranges(s->b_idx, 4, s->b_color, c.V, b_perm, s->b_off, s->atomic_b);
  std::vector<uint32_t> s_cta, b_cta;
  if (s->p.persist_cluster == -3) {
    // distributed-shared-memory shape: inside each colour the constraints are ordered by the CTA that owns their
    // lowest vertex (slice = vertices [r << shift, (r + 1) << shift)), and every CTA gets its own range of the colour
    s->dsmem_shift = persist_dsmem_shift(c.V);
    XRT_REQUIRE(s->dsmem_shift >= 0, "persist_cluster -3: needs 16-CTA clusters and a garment of at most 131072 vertices");
    XRT_REQUIRE(!s->atomic_s && !s->atomic_b, "persist_cluster -3: needs proper stretch and bend colourings");
    auto by_owner = [&](const std::vector<uint32_t>& idx, int arity, std::vector<uint32_t>& perm, const std::vector<uint32_t>& off,
                        std::vector<uint32_t>& cta) {
      auto owner = [&](uint32_t k) {
        uint32_t m = idx[(size_t)k * arity];
        for (int a = 1; a < arity; a++) m = std::min(m, idx[(size_t)k * arity + a]);
        return m >> s->dsmem_shift;
      };
      cta.assign((off.size() - 1) * (XRT_DSMEM_CLUSTER + 1), 0);
      for (size_t col = 0; col + 1 < off.size(); col++) {
        std::stable_sort(perm.begin() + off[col], perm.begin() + off[col + 1], [&](uint32_t x, uint32_t y) { return owner(x) < owner(y); });
        uint32_t* o = &cta[col * (XRT_DSMEM_CLUSTER + 1)];
        uint32_t at = off[col];
        for (int r = 0; r <= XRT_DSMEM_CLUSTER; r++) {
          while (at < off[col + 1] && (int)owner(perm[at]) < r) at++;
          o[r] = at;
        }
        o[XRT_DSMEM_CLUSTER] = off[col + 1];
      }
    };
    by_owner(s->s_idx, 2, s_perm, s->s_off, s_cta);
    by_owner(s->b_idx, 4, b_perm, s->b_off, b_cta);
  }
  c.n_stretch = ns; c.n_sc = (uint32_t)s->s_off.size() - 1;
  c.n_bend = nb;    c.n_bc = (uint32_t)s->b_off.size() - 1;
  std::vector<uint2> hs(ns);
  std::vector<float> hr(ns), hl(ns, 0.f);
  for (uint32_t i = 0; i < ns; i++) {
    hs[i] = make_uint2(s->s_idx[2 * s_perm[i]], s->s_idx[2 * s_perm[i] + 1]);
    hr[i] = s->s_rest[s_perm[i]];
  }
  std::vector<uint4> hb(nb);
  std::vector<float> ht(nb);
  for (uint32_t i = 0; i < nb; i++) {
    const uint32_t* q = &s->b_idx[4 * (size_t)b_perm[i]];
    hb[i] = make_uint4(q[0], q[1], q[2], q[3]);
    ht[i] = s->b_theta[b_perm[i]];
  }
  // --- cloth device state ---
  std::vector<float4> cx = to_float4(s->c_pos, s->c_invm);
  XRT_CUDA_OK(dev_upload(&c.x0, cx));
  XRT_CUDA_OK(dev_upload(&c.x, cx));
  XRT_CUDA_OK(dev_upload(&c.rest, cx));
  XRT_CUDA_OK(dev_alloc(&c.v, c.V));
  XRT_CUDA_OK(cudaMemset(c.v, 0, c.V * sizeof(float4)));
  XRT_CUDA_OK(dev_alloc(&c.n, c.V));
  XRT_CUDA_OK(dev_alloc(&c.delta, c.V));
  XRT_CUDA_OK(cudaMemset(c.delta, 0, c.V * sizeof(float4)));
  XRT_CUDA_OK(dev_upload(&c.tri, s->c_tri));
  std::vector<uint2> he(c.E);
  for (uint32_t i = 0; i < c.E; i++) he[i] = make_uint2(s->c_edge[2 * i], s->c_edge[2 * i + 1]);
  XRT_CUDA_OK(dev_upload(&c.edge, he));
  std::vector<uint32_t> vfo, vfi;
  xrt_build_vf(s->c_tri, c.V, vfo, vfi);
  XRT_CUDA_OK(dev_upload(&c.vf_off, vfo));
  XRT_CUDA_OK(dev_upload(&c.vf_idx, vfi));
  XRT_CUDA_OK(dev_upload(&c.s_idx, hs));
  XRT_CUDA_OK(dev_upload(&c.s_rest, hr));
  XRT_CUDA_OK(dev_upload(&c.s_lambda, hl));
  XRT_CUDA_OK(dev_upload(&c.s_perm, s_perm));
  XRT_CUDA_OK(dev_upload(&c.b_idx, hb));
  XRT_CUDA_OK(dev_upload(&c.b_theta, ht));
  XRT_CUDA_OK(dev_upload(&s->d_s_off, s->s_off));
  XRT_CUDA_OK(dev_upload(&s->d_b_off, s->b_off));
  if (s->dsmem_shift >= 0) {
    XRT_CUDA_OK(dev_upload(&s->d_s_cta_off, s_cta));
    XRT_CUDA_OK(dev_upload(&s->d_b_cta_off, b_cta));
  }
  c.n_lra = (uint32_t)s->l_rest.size();
  XRT_CUDA_OK(dev_upload(&c.l_src, s->l_src));
  XRT_CUDA_OK(dev_upload(&c.l_tgt, s->l_tgt));
  XRT_CUDA_OK(dev_upload(&c.l_rest, s->l_rest));
  XRT_CUDA_OK(dev_alloc(&c.vf, c.V));
  XRT_CUDA_OK(dev_alloc(&c.rt, c.V));
  XRT_CUDA_OK(dev_alloc(&c.ee, c.E));
  XRT_CUDA_OK(cudaMemset(c.vf, 0xFF, c.V * sizeof(VFContact)));  // index = -1: no contact
  XRT_CUDA_OK(cudaMemset(c.rt, 0xFF, c.V * sizeof(VFContact)));
  {
    std::vector<EEContact> e0(c.E);
    for (auto& e : e0) { memset(&e, 0, sizeof(e)); e.dist = XRT_SCALAR_MAX; }
    if (c.E) XRT_CUDA_OK(cudaMemcpy(c.ee, e0.data(), c.E * sizeof(EEContact), cudaMemcpyHostToDevice));
  }
  XRT_CUDA_OK(dev_alloc(&c.v_active, (c.V + 31) / 32));
  XRT_CUDA_OK(cudaMemset(c.v_active, 0, std::max<size_t>((c.V + 31) / 32, 1) * 4));
  XRT_CUDA_OK(dev_alloc(&c.e_active, (c.E + 31) / 32));
  XRT_CUDA_OK(cudaMemset(c.e_active, 0, std::max<size_t>((c.E + 31) / 32, 1) * 4));
  XRT_CUDA_OK(dev_alloc(&c.cnt_vf, c.V));
  XRT_CUDA_OK(dev_alloc(&c.cnt_rt, c.V));
  XRT_CUDA_OK(dev_alloc(&c.cnt_ee, c.E));
  fill_step_const(s);
  if (hash_alloc(&c, s->k) != 0) {
    xrt_set_error(s, "spatial hash allocation failed");
    return XRT_ERR_CUDA;
  }
  // --- obstacle ---
  o.V = (uint32_t)s->o_invm.size();
  o.F = (uint32_t)(s->o_tri.size() / 3);
  if (o.V) {
    std::vector<float4> ox = to_float4(s->o_pos, s->o_invm);
    XRT_CUDA_OK(dev_upload(&o.x0, ox));
    XRT_CUDA_OK(dev_upload(&o.x, ox));
    XRT_CUDA_OK(dev_upload(&o.x1, ox));
    XRT_CUDA_OK(dev_alloc(&o.n, o.V));
    XRT_CUDA_OK(dev_alloc(&o.n1, o.V));
    XRT_CUDA_OK(dev_upload(&o.tri, s->o_tri));
    std::vector<uint32_t> ovo, ovi, rt;
    xrt_build_vf(s->o_tri, o.V, ovo, ovi);
    xrt_build_rtri(s->o_tri, ovo, ovi, rt);  // ClothSolverBase::SetupRTriangle
    XRT_CUDA_OK(dev_upload(&o.vf_off, ovo));
    XRT_CUDA_OK(dev_upload(&o.vf_idx, ovi));
    XRT_CUDA_OK(dev_upload(&o.rtri, rt));
    if (bvh_alloc(&s->bvh, &s->bw, o.F) != 0) {
      xrt_set_error(s, "BVH allocation failed");
      return XRT_ERR_CUDA;
    }
    XRT_CUDA_OK(dev_alloc(&s->pred_alt, 3 * (size_t)o.F));
  }
  // --- bindings: GenerateBinding_Kernel (BindingConstraint.cu:22-46) on the host ---
  c.n_bind = (uint32_t)s->bd_vert.size();
  if (c.n_bind) {
    XRT_REQUIRE(o.V, "bindings need an obstacle");
    std::vector<uint4> bi(c.n_bind);
    std::vector<float4> bp(c.n_bind);
    for (uint32_t i = 0; i < c.n_bind; i++) {
      const uint32_t cv = s->bd_vert[i];
      const uint32_t* t = &s->bd_tri[3 * (size_t)i];
      const float* w = &s->bd_bary[3 * (size_t)i];
      XRT_REQUIRE(cv < c.V && t[0] < o.V && t[1] < o.V && t[2] < o.V, "binding index out of range");
      float L = s->bd_dist[i];
      if (L < XRT_EPS) {
        float d2 = 0.f;
        for (int k = 0; k < 3; k++) {
          const float pb = w[0] * s->o_pos[3 * t[0] + k] + w[1] * s->o_pos[3 * t[1] + k] + w[2] * s->o_pos[3 * t[2] + k];
          const float d = s->c_pos[3 * cv + k] - pb;
          d2 += d * d;
        }
        L = sqrtf(d2);
      }
      bi[i] = make_uint4(cv, t[0], t[1], t[2]);
      bp[i] = make_float4(w[0], w[1], w[2], L);
    }
    XRT_CUDA_OK(dev_upload(&c.bd_idx, bi));
    XRT_CUDA_OK(dev_upload(&c.bd_par, bp));
    XRT_CUDA_OK(dev_upload(&c.bd_k, s->bd_k));
  }
  // launch shape (xrt_params.persist_cluster)
  {
    int cl = s->p.persist_cluster;
    const int mx = persist_max_cluster();
    if (cl == -2) {
      XRT_REQUIRE(persist_grid_ctas() > 0, "persist_cluster -2: the substep kernel cannot be co-resident on this device");
      cl = XRT_PERSIST_GRID;
      XRT_CUDA_OK(dev_alloc(&s->d_grid_bar, 1));
      XRT_CUDA_OK(cudaMemset(s->d_grid_bar, 0, sizeof(unsigned long long)));
    } else if (cl == -3) {
      cl = XRT_PERSIST_DSMEM;  // checked where the constraints were ordered, above
    } else if (cl < 0 || cl > mx) {
      cl = mx;
    }
    XRT_REQUIRE(cl == 0 || cl == XRT_PERSIST_GRID || cl == XRT_PERSIST_DSMEM || (cl & (cl - 1)) == 0,
                "persist_cluster must be 0, -1, -2, -3 or a power of two up to 16");
    s->cluster = cl;
    const char* pe = getenv("XRT_PERSIST_PROF");
    if (cl > 0 && pe && pe[0] == '1') {
      XRT_CUDA_OK(dev_alloc(&s->d_prof, 32));
      XRT_CUDA_OK(cudaMemset(s->d_prof, 0, 32 * sizeof(unsigned long long)));
    }
  }
  const size_t stage = std::max<size_t>(std::max(c.V, o.V), 1);
  XRT_CUDA_OK(cudaMallocHost((void**)&s->stage, stage * sizeof(float4)));
  XRT_CUDA_OK(dev_alloc(&s->d_stage, stage));
  s->stage_elems = stage;
  XRT_CUDA_OK(cudaMallocHost((void**)&s->h_up, std::max<size_t>(o.V, 1) * 3 * sizeof(float)));
  XRT_CUDA_OK(dev_alloc(&s->d_up, std::max<size_t>(o.V, 1) * 3));
  XRT_CUDA_OK(cudaMallocHost((void**)&s->h_dn, stage * 6 * sizeof(float)));
  XRT_CUDA_OK(dev_alloc(&s->d_dn, stage * 6));
  XRT_CUDA_OK(cudaEventCreateWithFlags(&s->ev_up, cudaEventDisableTiming));
  // ClothSolverUniversal::Start: obstacle normals, BVH UpdateData(0, false) + Construct
  if (o.F) {
    launch_normals(o.x0, o.n, o.tri, o.vf_off, o.vf_idx, o.V, s->st);
    bvh_build(s->bvh, s->bw, o.x0, o.tri, o.F, s->p.morton_fast_div, s->st);
  }
  launch_normals(c.x0, c.n, c.tri, c.vf_off, c.vf_idx, c.V, s->st);
  XRT_CUDA_OK(cudaStreamSynchronize(s->st));
  XRT_CUDA_OK(cudaGetLastError());
  if (s->q)
    if (int rc = xrt_quality_finalize(s)) return rc;
  s->finalized = true;
  return XRT_OK;
}

// ---------------------------------------------------------------------------
// the step path
// ---------------------------------------------------------------------------

// one pass of the inner loop after contact generation
// (ClothSolverBase.cpp:215-237; binding->Solve is a no-op without bindings)
static void enqueue_iteration(xrt_solver* s, int iteration, uint32_t* launches) {
  const ClothDev& c = s->c;
  if (!s->frame_bvh.tnodes) s->frame_bvh = s->bvh;  // eager use outside a frame
  cudaStream_t st = s->st;
  for (uint32_t col = 0; col < c.n_sc; col++) {
    launch_stretch_color(c, s->k, s->s_off[col], s->s_off[col + 1], iteration, s->atomic_s, st);
    (*launches)++;
  }
  if (s->p.solve_bending)
    for (uint32_t col = 0; col < c.n_bc; col++) {
      launch_bend_color(c, s->k, s->b_off[col], s->b_off[col + 1], s->atomic_b, st);
      (*launches)++;
    }
  if (s->p.geodesic_lra && c.n_lra) {
    launch_lra(c, s->k, st);
    (*launches)++;
  }
  if (s->o.F) {
    launch_solve_contacts(c, s->frame_bvh, s->k, st);  // RT solve, VF solve, EE solve
    (*launches) += c.E ? 2 : 1;
    launch_apply_deltas(c, s->k, st);
    (*launches)++;
    if (c.n_bind) {
      launch_binding(c, s->o, st);  // binding_->Solve, ClothSolverBase.cpp:236
      (*launches)++;
    }
  }
}

// The tree as collision pass `cnt` of this frame sees it.  The primitives (pred = x1 + tol n1) exist twice: the contact
// solves after pass k still read pass k's copy while the side branch already refits the tree for pass k + 1 into the
// other one.  The last pass of a frame always lands in s->bvh.pred (what the getters and the eager entry points read).
static DevBVH pass_bvh(const xrt_solver* s, int cnt) {
  DevBVH b = s->bvh;
  if (((s->frame_passes - 1 - cnt) & 1) && s->pred_alt) b.pred = s->pred_alt;
  return b;
}

// Obstacle side of collision pass `cnt`: Interpolate + UpdateMidstepNormals + RefitMidstep (ClothSolverBase.cpp:
// 205-209).  It needs nothing from the cloth, so in the frame graph it runs on the side branch while the main branch
// is still busy with the Gauss-Seidel phases before the pass.  It waits for the vertex queries of the previous pass
// (they read the tree this refit grows); the edge queries of that pass precede it on the side branch itself.
static void enqueue_obstacle_prep(xrt_solver* s, int cnt, uint32_t* launches) {
  cudaStream_t st2 = s->st2;
  if (!s->side_open) {
    cudaEventRecord(s->ev_fork, s->st);
    cudaStreamWaitEvent(st2, s->ev_fork, 0);
    s->side_open = true;
  }
  if (cnt > 0) cudaStreamWaitEvent(st2, s->ev_gv, 0);
  launch_interpolate(s->o, cnt, s->p.num_collision_passes, st2);
  launch_normals(s->o.x1, s->o.n1, s->o.tri, s->o.vf_off, s->o.vf_idx, s->o.V, st2);
  bvh_refit_midstep(pass_bvh(s, cnt), s->bw, s->o.x1, s->o.n1, s->o.tri, s->p.bvh_tolerance, st2);
  cudaEventRecord(s->ev_prep, st2);
  (*launches) += 2 + bvh_refit_launches();
}

// One collision pass of the per-stage frame graph.  Main branch: vertex queries; side branch: edge queries
// (independent of the vertex queries: both only read the tree and the positions), then the obstacle side of the
// next pass.  last: no pass follows in this frame.
static void enqueue_collision_pass(xrt_solver* s, bool moved, int cnt, bool last, uint32_t* launches) {
  if (!s->o.F) return;
  cudaStream_t st = s->st, st2 = s->st2;
  if (moved) {
    cudaStreamWaitEvent(st, s->ev_prep, 0);  // the tree of this pass (enqueue_obstacle_prep)
    s->frame_bvh = pass_bvh(s, cnt);
  }
  if (s->c.E) {
    cudaEventRecord(s->ev_fork, st);
    cudaStreamWaitEvent(st2, s->ev_fork, 0);
    s->side_open = true;
  }
  launch_generate_vertex_contacts(s->c, s->frame_bvh, st);
  if (moved && !last) cudaEventRecord(s->ev_gv, st);
  if (s->c.E) {
    launch_generate_edge_contacts(s->c, s->frame_bvh, s->o.rtri, st2);
    cudaEventRecord(s->ev_side, st2);
    cudaStreamWaitEvent(st, s->ev_side, 0);
  }
  (*launches) += s->c.E ? 2 : 1;
  if (moved && !last) enqueue_obstacle_prep(s, cnt + 1, launches);
}

// the side branch must have rejoined the main one before the capture (or the frame) ends
static void close_side_branch(xrt_solver* s) {
  if (!s->side_open) return;
  cudaEventRecord(s->ev_side, s->st2);
  cudaStreamWaitEvent(s->st, s->ev_side, 0);
  s->side_open = false;
}

// arguments of the persistent kernel for one substep of the frame
PersistArgs xrt_make_persist_args(const xrt_solver* s, int substep, bool moved, bool do_predict) {
  const xrt_params& p = s->p;
  uint32_t frequency = (uint32_t)(p.num_substeps * p.num_iterations) / (uint32_t)p.num_collision_passes;
  if (frequency == 0) frequency = 1;
  PersistArgs a{};
  a.c = s->c; a.o = s->o; a.bvh = s->bvh; a.k = s->k;
  a.s_off = s->d_s_off; a.b_off = s->d_b_off;
  a.refit_flags = s->bw.flags;
  a.tol = p.bvh_tolerance;
  a.substep = substep;
  a.n_iter = p.num_iterations; a.n_passes = p.num_collision_passes;
  a.frequency = frequency;
  a.do_predict = do_predict ? 1 : 0; a.do_finalize = 1;
  a.self_collision = p.enable_self_collision; a.solve_bending = p.solve_bending;
  a.lra = p.geodesic_lra; a.moved = moved ? 1 : 0;
  a.atomic_s = s->atomic_s ? 1 : 0; a.atomic_b = s->atomic_b ? 1 : 0;
  a.prof = s->d_prof;
  a.grid_bar = s->d_grid_bar;
  return a;
}

// SimulateSwift, ClothSolverBase.cpp:171-247
static uint32_t enqueue_frame(xrt_solver* s, bool moved) {
  const xrt_params& p = s->p;
  const ClothDev& c = s->c;
  cudaStream_t st = s->st;
  uint32_t launches = 0;
  uint32_t frequency = (uint32_t)(p.num_substeps * p.num_iterations) / (uint32_t)p.num_collision_passes;
  if (frequency == 0) frequency = 1;  // the reference divides by zero here
  int n_pass = 0;  // collision passes of this frame
  for (int step = 0; step < p.num_substeps * p.num_iterations; step++) n_pass += (uint32_t)step % frequency == 0;
  s->frame_passes = n_pass;
  s->frame_bvh = s->bvh;
  const bool staged = s->cluster == 0 || s->cluster == XRT_PERSIST_DSMEM;  // collision passes are separate kernels
  if (moved && s->o.F) {
    // UpdateObstacleAnimationFrame: BVH UpdateData(0, ccd=false) + Construct on the frame-start positions
    // (impl/smpl/ClothSolver.cpp:257-261).  In the staged shapes the build and the obstacle side of the first pass
    // run on the side branch, beside predict, the spatial hash and the first self-contact solve.
    cudaStream_t bst = st;
    if (staged) {
      cudaEventRecord(s->ev_fork, st);
      cudaStreamWaitEvent(s->st2, s->ev_fork, 0);
      s->side_open = true;
      bst = s->st2;
    }
    bvh_build(s->bvh, s->bw, s->o.x0, s->o.tri, s->o.F, p.morton_fast_div, bst);
    launches += bvh_build_launches();
    if (staged) enqueue_obstacle_prep(s, 0, &launches);
  }
  int cnt = 0;

  for (int substep = 0; substep < p.num_substeps; substep++) {
    const bool hash = p.enable_self_collision && substep % std::max(1, p.inter_leaved_hash) == 0;
    if (s->cluster == XRT_PERSIST_DSMEM) {
      // distributed-shared-memory shape: the substep is cut at its collision passes; those run as device-wide
      // per-stage kernels, everything between them as one segment on the 16-CTA cluster (xrt_persist.cu)
      if (hash) {
        launch_predict(c, s->k, st);
        launch_hash(c, s->k, st);
        launches += 6;
      }
      PersistArgs a = xrt_make_persist_args(s, substep, moved, !hash);
      a.x_shift = (uint32_t)s->dsmem_shift;
      a.s_cta_off = s->d_s_cta_off;
      a.b_cta_off = s->d_b_cta_off;
      auto due = [&](int it) { return s->o.F && (uint32_t)(substep * p.num_iterations + it) % frequency == 0; };
      int it = 0;
      bool resume = false;
      while (true) {
        int g = resume ? it + 1 : it;
        while (g < p.num_iterations && !due(g)) g++;
        a.bvh = s->frame_bvh;
        a.it_begin = it;
        a.stop_at = g;  // == num_iterations: run to the end of the substep (and finalize)
        a.resume = resume ? 1 : 0;
        launch_swift_segment(a, st);
        launches++;
        a.do_predict = 0;
        if (g >= p.num_iterations) break;
        enqueue_collision_pass(s, moved, cnt, cnt + 1 >= n_pass, &launches);
        cnt++;
        it = g;
        resume = true;
      }
      continue;
    }
    if (s->cluster > 0) {
      // persistent shape: [predict + hash when due] + ONE kernel for the rest of the substep
      if (hash) {
        launch_predict(c, s->k, st);
        launch_hash(c, s->k, st);
        launches += 6;
      }
      const PersistArgs a = xrt_make_persist_args(s, substep, moved, !hash);
      launch_swift_substep(a, s->cluster, st);
      launches++;
      continue;
    }
    launch_predict(c, s->k, st);
    launches++;
    if (hash) {
      launch_hash(c, s->k, st);
      launches += 5;
    }
    for (int it = 0; it < p.num_iterations; it++) {
      if (p.enable_self_collision) {
        launch_collide_particles(c, s->k, st);
        launches += 2;
      }
      const uint32_t step = (uint32_t)(substep * p.num_iterations + it);
      if (step % frequency == 0) {
        enqueue_collision_pass(s, moved, cnt, cnt + 1 >= n_pass, &launches);
        cnt++;
      }
      enqueue_iteration(s, it, &launches);
    }
    launch_finalize(c, s->k, st);
    launches++;
  }
  // obstacle_->PredictedToPositions(); cloth_/obstacle_->UpdateNormals()
  if (s->o.V) {
    if (moved) cudaMemcpyAsync(s->o.x0, s->o.x, s->o.V * sizeof(float4), cudaMemcpyDeviceToDevice, st);
    launch_normals(s->o.x0, s->o.n, s->o.tri, s->o.vf_off, s->o.vf_idx, s->o.V, st);
    launches++;
  }
  launch_normals(c.x0, c.n, c.tri, c.vf_off, c.vf_idx, c.V, st);
  launches++;
  close_side_branch(s);
  return launches;
}

static int ensure_graph(xrt_solver* s, int moved) {
  if (s->graph[moved]) return XRT_OK;
  cudaGraph_t g = nullptr;
  XRT_CUDA_OK(cudaStreamBeginCapture(s->st, cudaStreamCaptureModeThreadLocal));
  s->launches[moved] = enqueue_frame(s, moved != 0);
  // a launch that failed inside the capture shows up here at the latest: the capture is always ended, so the
  // stream never stays in capture mode, and the first error is the one reported
  const cudaError_t launch_err = cudaGetLastError();
  const cudaError_t end_err = cudaStreamEndCapture(s->st, &g);
  s->side_open = false;
  if (launch_err != cudaSuccess || end_err != cudaSuccess || !g) {
    if (g) cudaGraphDestroy(g);
    const cudaError_t e = launch_err != cudaSuccess ? launch_err : end_err;
    xrt_set_error(s, "frame graph capture failed: %s", cudaGetErrorString(e));
    (void)cudaGetLastError();
    return XRT_ERR_CUDA;
  }
  const cudaError_t inst = cudaGraphInstantiate(&s->graph[moved], g, 0);
  cudaGraphDestroy(g);
  if (inst != cudaSuccess) {
    s->graph[moved] = nullptr;
    xrt_set_error(s, "frame graph instantiation failed: %s", cudaGetErrorString(inst));
    return XRT_ERR_CUDA;
  }
  return XRT_OK;
}

extern "C" int xrt_set_obstacle_target(xrt_solver* s, const float* xyz, size_t n_verts) {
  XRT_READY();
  XRT_REQUIRE(xyz && n_verts == s->o.V, "set_obstacle_target: vertex count differs from the obstacle");
  // no stream sync: only the previous upload must have left the pinned buffer
  XRT_CUDA_OK(cudaEventSynchronize(s->ev_up));
  memcpy(s->h_up, xyz, n_verts * 3 * sizeof(float));
  XRT_CUDA_OK(cudaMemcpyAsync(s->d_up, s->h_up, n_verts * 3 * sizeof(float), cudaMemcpyHostToDevice, s->st));
  XRT_CUDA_OK(cudaEventRecord(s->ev_up, s->st));
  launch_unpack_xyz(s->d_up, s->o.x, (uint32_t)n_verts, s->st);  // keeps o.x.w = inverse mass
  s->target_pending = true;
  return XRT_OK;
}

extern "C" int xrt_set_obstacle_sequence(xrt_solver* s, const float* xyz, size_t n_frames) {
  XRT_READY();
  XRT_REQUIRE(s->o.V && xyz && n_frames, "set_obstacle_sequence: no obstacle or empty sequence");
  const size_t Vo = s->o.V;
  std::vector<float4> h(n_frames * Vo);
  for (size_t f = 0; f < n_frames; f++)
    for (size_t i = 0; i < Vo; i++) {
      const float* p = xyz + 3 * (f * Vo + i);
      h[f * Vo + i] = make_float4(p[0], p[1], p[2], s->o_invm[i]);
    }
  XRT_CUDA_OK(cudaStreamSynchronize(s->st));
  cudaFree(s->d_seq);
  s->d_seq = nullptr;
  XRT_CUDA_OK(dev_upload(&s->d_seq, h));
  s->seq_frames = n_frames;
  s->seq_cursor = 0;
  return XRT_OK;
}

// next frame of the resident sequence -> obstacle target (device to device)
int xrt_take_sequence_frame(xrt_solver* s, cudaStream_t st) {
  if (s->rig)  // a resident pose sequence skins the frame on the device (xrt_rig.cu)
    if (int rc = xrt_rig_take_sequence_frame(s, st)) return rc;
  if (s->target_pending || !s->seq_frames) return XRT_OK;
  const float4* src = s->d_seq + (s->seq_cursor % s->seq_frames) * s->o.V;
  XRT_CUDA_OK(cudaMemcpyAsync(s->o.x, src, s->o.V * sizeof(float4), cudaMemcpyDeviceToDevice, st));
  s->seq_cursor++;
  s->target_pending = true;
  return XRT_OK;
}

extern "C" int xrt_step_frame_async(xrt_solver* s) {
  XRT_READY();
  if (s->q) {  // a Quality frame synchronises inside CollisionStep: there is no enqueue-only form
    if (int rc0 = xrt_take_sequence_frame(s, s->st)) return rc0;
    return xrt_quality_step_frame(s, nullptr);
  }
  if (int rc0 = xrt_take_sequence_frame(s, s->st)) return rc0;
  const int moved = s->target_pending ? 1 : 0;
  int rc = ensure_graph(s, moved);
  if (rc) return rc;
  XRT_CUDA_OK(cudaGraphLaunch(s->graph[moved], s->st));
  s->target_pending = false;
  return XRT_OK;
}

extern "C" int xrt_sync(xrt_solver* s) {
  XRT_READY();
  XRT_CUDA_OK(cudaStreamSynchronize(s->st));
  return XRT_OK;
}

extern "C" int xrt_step_frame(xrt_solver* s, float* solver_ms) {
  XRT_READY();
  if (int rc0 = xrt_take_sequence_frame(s, s->st)) return rc0;
  if (s->q) return xrt_quality_step_frame(s, solver_ms);  // ClothSolverBase::SimulateQuality
  const int moved = s->target_pending ? 1 : 0;
  int rc = ensure_graph(s, moved);
  if (rc) return rc;
  XRT_CUDA_OK(cudaEventRecord(s->ev0, s->st));  // Timer::StartTimerGPU("Solver_Total")
  XRT_CUDA_OK(cudaGraphLaunch(s->graph[moved], s->st));
  XRT_CUDA_OK(cudaEventRecord(s->ev1, s->st));
  s->target_pending = false;
  XRT_CUDA_OK(cudaStreamSynchronize(s->st));
  if (solver_ms) XRT_CUDA_OK(cudaEventElapsedTime(solver_ms, s->ev0, s->ev1));
  return XRT_OK;
}

extern "C" int xrt_batch_run(xrt_solver** handles, int n, int frames, float* ms) {
  if (!handles || n <= 0 || frames < 0) return XRT_ERR_INVALID;
  xrt_solver* s = handles[0];
  XRT_READY();
  for (int i = 0; i < n; i++) {
    if (!handles[i] || !handles[i]->finalized || handles[i]->p.device != s->p.device) {
      xrt_set_error(s, "batch_run: handle %d is null, not finalized or on another device", i);
      return XRT_ERR_INVALID;
    }
    if (!handles[i]->ev_join) XRT_CUDA_OK(cudaEventCreateWithFlags(&handles[i]->ev_join, cudaEventDisableTiming));
  }
  XRT_CUDA_OK(cudaEventRecord(s->ev0, s->st));
  for (int i = 1; i < n; i++) XRT_CUDA_OK(cudaStreamWaitEvent(handles[i]->st, s->ev0, 0));
  for (int f = 0; f < frames; f++)
    for (int i = 0; i < n; i++) {
      int rc = xrt_step_frame_async(handles[i]);
      if (rc) {
        if (i) xrt_set_error(s, "batch_run: handle %d: %s", i, handles[i]->err);
        return rc;
      }
    }
  for (int i = 1; i < n; i++) {
    XRT_CUDA_OK(cudaEventRecord(handles[i]->ev_join, handles[i]->st));
    XRT_CUDA_OK(cudaStreamWaitEvent(s->st, handles[i]->ev_join, 0));
  }
  XRT_CUDA_OK(cudaEventRecord(s->ev1, s->st));
  XRT_CUDA_OK(cudaStreamSynchronize(s->st));
  if (ms) XRT_CUDA_OK(cudaEventElapsedTime(ms, s->ev0, s->ev1));
  return XRT_OK;
}

// boxes in the reference's node numbering (BVH::AABBsHost) from the sibling-pair slots
static cudaError_t download_aabbs(const DevBVH& b, size_t F, float* o) {
  if (!F) return cudaSuccess;
  std::vector<TNode> t(2 * F);
  std::vector<BVNode> nd(2 * F - 1);
  cudaError_t e = cudaMemcpy(t.data(), b.tnodes, t.size() * sizeof(TNode), cudaMemcpyDeviceToHost);
  if (e == cudaSuccess) e = cudaMemcpy(nd.data(), b.nodes, nd.size() * sizeof(BVNode), cudaMemcpyDeviceToHost);
  if (e != cudaSuccess) return e;
  for (size_t i = 0; i < nd.size(); i++) {
    const TNode& n = t[tnode_slot(nd.data(), (uint32_t)i)];
    o[6 * i] = n.lx; o[6 * i + 1] = n.ly; o[6 * i + 2] = n.lz;
    o[6 * i + 3] = n.ux; o[6 * i + 4] = n.uy; o[6 * i + 5] = n.uz;
  }
  return cudaSuccess;
}

double xrt_bvh_build_bytes(const xrt_solver* s) {
  const double Fo = s->o.F;
  return 84.0 * Fo + 24.0 * Fo + 8.0 * Fo * 8.0 + 40.0 * (2 * Fo - 1) + 8.0 * Fo;
}

extern "C" int xrt_persist_info(int cluster, int* max_cluster, int* resident) {
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev <= 0) {
    (void)cudaGetLastError();
    return XRT_ERR_CUDA;
  }
  const int mx = persist_max_cluster();
  if (max_cluster) *max_cluster = mx;
  if (cluster < 0 || cluster > mx) cluster = mx;
  if (resident) *resident = persist_active_clusters(cluster < 1 ? 1 : cluster);
  return XRT_OK;
}

extern "C" int xrt_profile_kernel(xrt_solver* s, int which, int reps, float* avg_ms, double* bytes) {
  XRT_READY();
  XRT_REQUIRE(reps > 0 && avg_ms && bytes, "profile_kernel: bad argument");
  const ClothDev& c = s->c;
  int launches = 0;
  double b = 0;
  for (int pass = 0; pass < 2; pass++) {  // pass 0 warms up, pass 1 is timed
    if (pass == 1) XRT_CUDA_OK(cudaEventRecord(s->ev0, s->st));
    launches = 0;
    for (int r = 0; r < (pass ? reps : 3); r++) {
      switch (which) {
        case 0:
          for (uint32_t col = 0; col < c.n_bc; col++, launches++)
            launch_bend_color(c, s->k, s->b_off[col], s->b_off[col + 1], s->atomic_b, s->st);
          b = 160.0 * c.n_bend / std::max(1u, c.n_bc);
          break;
        case 1:
          for (uint32_t col = 0; col < c.n_sc; col++, launches++)
            launch_stretch_color(c, s->k, s->s_off[col], s->s_off[col + 1], 1, s->atomic_s, s->st);
          b = 84.0 * c.n_stretch / std::max(1u, c.n_sc);
          break;
        case 2:
          launch_collide_particles(c, s->k, s->st);
          launches += 2;
          b = ((4.0 * 8 + 16.0 * 8 + 36.0) * c.V + 52.0 * c.V) / 2;
          break;
        case 3:
          XRT_REQUIRE(s->o.F, "no obstacle");
          launch_generate_contacts(c, s->bvh, s->o.rtri, false, s->st);
          launches += c.E ? 2 : 1;
          b = (96.0 * c.V + 88.0 * c.E) / (c.E ? 2 : 1);
          break;
        case 4:
          launch_predict(c, s->k, s->st);
          launches++;
          b = 64.0 * c.V;
          break;
        case 5: {
          // the persistent substep kernel (substep 1: predict inside, no hash before it);
          // algorithmic bytes = the frame's minus the BVH build, per substep
          XRT_REQUIRE(s->cluster > 0 && s->cluster != XRT_PERSIST_DSMEM, "profile_kernel: no persistent substep kernel in this launch shape");
          const PersistArgs a = xrt_make_persist_args(s, 1 % s->p.num_substeps, s->o.F != 0, true);
          XRT_CUDA_OK(launch_swift_substep(a, s->cluster, s->st));
          launches++;
          b = (xrt_frame_bytes(s, s->o.F != 0) - (s->o.F ? xrt_bvh_build_bytes(s) : 0.0)) / s->p.num_substeps;
          break;
        }
        default: xrt_set_error(s, "profile_kernel: unknown kernel %d", which); return XRT_ERR_INVALID;
      }
    }
  }
  XRT_CUDA_OK(cudaEventRecord(s->ev1, s->st));
  XRT_CUDA_OK(cudaStreamSynchronize(s->st));
  float ms = 0;
  XRT_CUDA_OK(cudaEventElapsedTime(&ms, s->ev0, s->ev1));
  *avg_ms = ms / (float)std::max(1, launches);
  *bytes = b;
  return XRT_OK;
}

extern "C" int xrt_batch_profile_substep(xrt_solver** handles, int n, int reps, float* avg_ms, double* bytes) {
  if (!handles || n <= 0 || reps <= 0 || !avg_ms || !bytes) return XRT_ERR_INVALID;
  xrt_solver* s = handles[0];
  XRT_READY();
  for (int i = 0; i < n; i++) {
    if (!handles[i] || !handles[i]->finalized || handles[i]->p.device != s->p.device || handles[i]->cluster <= 0 ||
        handles[i]->cluster == XRT_PERSIST_DSMEM) {
      xrt_set_error(s, "batch_profile_substep: handle %d is null, not finalized, on another device or per-stage", i);
      return XRT_ERR_INVALID;
    }
    if (!handles[i]->ev_join) XRT_CUDA_OK(cudaEventCreateWithFlags(&handles[i]->ev_join, cudaEventDisableTiming));
  }
  double b = 0;
  for (int pass = 0; pass < 2; pass++) {  // pass 0 warms up
    XRT_CUDA_OK(cudaEventRecord(s->ev0, s->st));
    for (int i = 1; i < n; i++) XRT_CUDA_OK(cudaStreamWaitEvent(handles[i]->st, s->ev0, 0));
    b = 0;
    for (int i = 0; i < n; i++) {
      xrt_solver* h = handles[i];
      const PersistArgs a = xrt_make_persist_args(h, 1 % h->p.num_substeps, h->o.F != 0, true);
      for (int r = 0; r < (pass ? reps : 2); r++) XRT_CUDA_OK(launch_swift_substep(a, h->cluster, h->st));
      b += (xrt_frame_bytes(h, h->o.F != 0) - (h->o.F ? xrt_bvh_build_bytes(h) : 0.0)) / h->p.num_substeps;
    }
    for (int i = 1; i < n; i++) {
      XRT_CUDA_OK(cudaEventRecord(handles[i]->ev_join, handles[i]->st));
      XRT_CUDA_OK(cudaStreamWaitEvent(s->st, handles[i]->ev_join, 0));
    }
    XRT_CUDA_OK(cudaEventRecord(s->ev1, s->st));
    XRT_CUDA_OK(cudaStreamSynchronize(s->st));
  }
  float ms = 0;
  XRT_CUDA_OK(cudaEventElapsedTime(&ms, s->ev0, s->ev1));
  *avg_ms = ms / (float)reps;
  *bytes = b;
  return XRT_OK;
}


// building blocks for the parity tests (eager launches of the same kernels)
extern "C" int xrt_bvh_build_obstacle(xrt_solver* s) {
  XRT_READY();
  XRT_REQUIRE(s->o.F, "no obstacle");
  bvh_build(s->bvh, s->bw, s->o.x0, s->o.tri, s->o.F, s->p.morton_fast_div, s->st);
  XRT_CUDA_OK(cudaStreamSynchronize(s->st));
  return XRT_OK;
}
extern "C" int xrt_bvh_refit_obstacle(xrt_solver* s, int step) {
  XRT_READY();
  XRT_REQUIRE(s->o.F, "no obstacle");
  launch_interpolate(s->o, step, s->p.num_collision_passes, s->st);
  launch_normals(s->o.x1, s->o.n1, s->o.tri, s->o.vf_off, s->o.vf_idx, s->o.V, s->st);
  bvh_refit_midstep(s->bvh, s->bw, s->o.x1, s->o.n1, s->o.tri, s->p.bvh_tolerance, s->st);
  XRT_CUDA_OK(cudaStreamSynchronize(s->st));
  return XRT_OK;
}
extern "C" int xrt_generate_contacts(xrt_solver* s) {
  XRT_READY();
  XRT_REQUIRE(s->o.F, "no obstacle");
  launch_generate_contacts(s->c, s->bvh, s->o.rtri, true, s->st);
  XRT_CUDA_OK(cudaStreamSynchronize(s->st));
  return XRT_OK;
}
extern "C" int xrt_hash_particles(xrt_solver* s) {
  XRT_READY();
  launch_hash(s->c, s->k, s->st);
  XRT_CUDA_OK(cudaStreamSynchronize(s->st));
  return XRT_OK;
}
extern "C" int xrt_predict(xrt_solver* s) {
  XRT_READY();
  launch_predict(s->c, s->k, s->st);
  XRT_CUDA_OK(cudaStreamSynchronize(s->st));
  return XRT_OK;
}
extern "C" int xrt_solve_iteration(xrt_solver* s, int iteration) {
  XRT_READY();
  uint32_t l = 0;
  s->frame_bvh = s->bvh;
  if (s->p.enable_self_collision) launch_collide_particles(s->c, s->k, s->st);
  enqueue_iteration(s, iteration, &l);
  XRT_CUDA_OK(cudaStreamSynchronize(s->st));
  return XRT_OK;
}
extern "C" int xrt_finalize(xrt_solver* s) {
  XRT_READY();
  launch_finalize(s->c, s->k, s->st);
  XRT_CUDA_OK(cudaStreamSynchronize(s->st));
  return XRT_OK;
}

// ---------------------------------------------------------------------------
// read-back, debug, teardown
// ---------------------------------------------------------------------------

static int download_xyz(xrt_solver* s, const float4* d, size_t n, float* out) {
  if (!out || !n) return XRT_OK;
  XRT_CUDA_OK(cudaMemcpyAsync(s->stage, d, n * sizeof(float4), cudaMemcpyDeviceToHost, s->st));
  XRT_CUDA_OK(cudaStreamSynchronize(s->st));
  for (size_t i = 0; i < n; i++) {
    out[3 * i] = s->stage[i].x;
    out[3 * i + 1] = s->stage[i].y;
    out[3 * i + 2] = s->stage[i].z;
  }
  return XRT_OK;
}

extern "C" int xrt_get_positions(xrt_solver* s, int which, float* xyz, float* nrm) {
  XRT_READY();
  XRT_REQUIRE(which == 0 || which == 1, "get_positions: which must be 0 (cloth) or 1 (obstacle)");
  const size_t n = which == 0 ? s->c.V : s->o.V;
  if (!n || (!xyz && !nrm)) return XRT_OK;
  const float4* a = which == 0 ? s->c.x0 : s->o.x0;
  const float4* b = which == 0 ? s->c.n : s->o.n;
  // pack to float3 on the device, one D2H into pinned memory, memcpy out
  launch_pack_xyz(xyz ? a : nullptr, nrm ? b : nullptr, s->d_dn, (uint32_t)n, s->st);
  XRT_CUDA_OK(cudaMemcpyAsync(s->h_dn, s->d_dn, n * 6 * sizeof(float), cudaMemcpyDeviceToHost, s->st));
  XRT_CUDA_OK(cudaStreamSynchronize(s->st));
  if (xyz) memcpy(xyz, s->h_dn, n * 3 * sizeof(float));
  if (nrm) memcpy(nrm, s->h_dn + 3 * n, n * 3 * sizeof(float));
  return XRT_OK;
}

extern "C" int xrt_set_cloth_state(xrt_solver* s, const float* xyz, const float* vel) {
  XRT_READY();
  const size_t V = s->c.V;
  XRT_CUDA_OK(cudaStreamSynchronize(s->st));
  if (xyz) {
    for (size_t i = 0; i < V; i++) s->stage[i] = make_float4(xyz[3 * i], xyz[3 * i + 1], xyz[3 * i + 2], s->c_invm[i]);
    XRT_CUDA_OK(cudaMemcpyAsync(s->c.x0, s->stage, V * sizeof(float4), cudaMemcpyHostToDevice, s->st));
    XRT_CUDA_OK(cudaMemcpyAsync(s->c.x, s->stage, V * sizeof(float4), cudaMemcpyHostToDevice, s->st));
    XRT_CUDA_OK(cudaStreamSynchronize(s->st));
  }
  if (vel) {
    for (size_t i = 0; i < V; i++) s->stage[i] = make_float4(vel[3 * i], vel[3 * i + 1], vel[3 * i + 2], 0.f);
    XRT_CUDA_OK(cudaMemcpyAsync(s->c.v, s->stage, V * sizeof(float4), cudaMemcpyHostToDevice, s->st));
    XRT_CUDA_OK(cudaStreamSynchronize(s->st));
  }
  return XRT_OK;
}

extern "C" int xrt_set_cloth_predicted(xrt_solver* s, const float* xyz) {
  XRT_READY();
  XRT_REQUIRE(xyz, "set_cloth_predicted: null array");
  const size_t V = s->c.V;
  XRT_CUDA_OK(cudaStreamSynchronize(s->st));
  for (size_t i = 0; i < V; i++) s->stage[i] = make_float4(xyz[3 * i], xyz[3 * i + 1], xyz[3 * i + 2], s->c_invm[i]);
  XRT_CUDA_OK(cudaMemcpyAsync(s->c.x, s->stage, V * sizeof(float4), cudaMemcpyHostToDevice, s->st));
  XRT_CUDA_OK(cudaStreamSynchronize(s->st));
  return XRT_OK;
}

// algorithmic bytes of one frame, SoA-ideal, SURVEY.md section 8(d) with the
// exact counts of this scene (see DESIGN.md "Roofline accounting")
double xrt_frame_bytes(const xrt_solver* s, bool moved) {
  const xrt_params& p = s->p;
  const double V = s->c.V, E = s->c.E, Fo = s->o.F, Vo = s->o.V;
  const double ns = s->c.n_stretch, nb = p.solve_bending ? s->c.n_bend : 0, nl = p.geodesic_lra ? s->c.n_lra : 0;
  const double S = p.num_substeps, N = p.num_iterations, C = p.num_collision_passes;
  double per_iter = 84.0 * ns + 160.0 * nb + 60.0 * nl;
  if (Fo) per_iter += 96.0 * V + 96.0 * V + 120.0 * E + 52.0 * V;  // VF-CCD, VF-DCD, EE solves, apply deltas
  double b = S * N * per_iter + S * (64.0 + 64.0) * V;              // + predict + finalize
  if (Fo) b += C * (48.0 * V + 48.0 * V + 88.0 * E);                // contact generation, compulsory traffic only
  if (moved && Fo) {
    b += xrt_bvh_build_bytes(s);  // BVH build
    b += C * (36.0 * Fo + 24.0 * (2 * Fo - 1) + 32.0 * Vo);                          // refit + interpolate
  }
  if (p.enable_self_collision) {
    const double K = 8.0;  // mean neighbour count assumed by SURVEY.md 8(d)
    b += S * N * ((4.0 * K + 16.0 * K + 36.0) * V + 52.0 * V);
  }
  return b;
}

extern "C" int xrt_debug_get(xrt_solver* s, int what, void* dst, size_t* bytes) {
  XRT_READY();
  XRT_REQUIRE(bytes, "debug_get: bytes is null");
  XRT_CUDA_OK(cudaStreamSynchronize(s->st));
  const size_t F = s->bvh.n_obj, V = s->c.V, E = s->c.E;
  const void* src = nullptr;
  size_t need = 0;
  std::vector<float> tmp;
  std::vector<uint32_t> tmpu;
  xrt_stats stats;
  switch (what) {
    case XRT_DBG_MORTON: src = s->bvh.morton; need = F * 4; break;
    case XRT_DBG_SORTED_OBJECT: src = s->bvh.sorted_obj; need = F * 4; break;
    case XRT_DBG_NODES: src = s->bvh.nodes; need = F ? (2 * F - 1) * sizeof(BVNode) : 0; break;
    case XRT_DBG_ESCAPE_INT: src = s->bvh.esc_int; need = F ? (F - 1) * 4 : 0; break;
    case XRT_DBG_ESCAPE_EXT: src = s->bvh.esc_ext; need = F * 4; break;
    case XRT_DBG_VF_CONTACTS: src = s->c.vf; need = V * sizeof(VFContact); break;
    case XRT_DBG_RT_CONTACTS: src = s->c.rt; need = V * sizeof(VFContact); break;
    case XRT_DBG_EE_CONTACTS: need = E * sizeof(EEContact); break;
    case XRT_DBG_NEIGHBORS: src = s->c.neighbors; need = V * (size_t)s->p.max_num_neighbors * 4; break;
    case XRT_DBG_RTRI: src = s->o.rtri; need = (size_t)s->o.F * 4; break;
    case XRT_DBG_VF_CAND_COUNT: src = s->c.cnt_vf; need = V * 4; break;
    case XRT_DBG_RT_CAND_COUNT: src = s->c.cnt_rt; need = V * 4; break;
    case XRT_DBG_EE_CAND_COUNT: src = s->c.cnt_ee; need = E * 4; break;
    case XRT_DBG_PERSIST_PROF: src = s->d_prof; need = s->d_prof ? 32 * sizeof(unsigned long long) : 0; break;
    case XRT_DBG_COLORS_STRETCH: need = s->s_color.size() * 4; break;
    case XRT_DBG_COLORS_BEND: need = s->b_color.size() * 4; break;
    case XRT_DBG_AABBS: need = F ? (2 * F - 1) * 6 * sizeof(float) : 0; break;
    case XRT_DBG_PREDICTED: case XRT_DBG_VELOCITY: need = V * 3 * sizeof(float); break;
    case XRT_DBG_OBSTACLE_PRED: need = (size_t)s->o.F * 9 * sizeof(float); break;
    case XRT_DBG_PARTICLE_HASH: need = V * 8; break;
    case XRT_DBG_STATS: need = sizeof(xrt_stats); break;
    default: xrt_set_error(s, "debug_get: unknown item %d", what); return XRT_ERR_INVALID;
  }
  const size_t cap = *bytes;
  *bytes = need;
  if (!dst) return XRT_OK;
  XRT_REQUIRE(cap >= need, "debug_get: destination too small");
  if (src) {
    if (need) XRT_CUDA_OK(cudaMemcpy(dst, src, need, cudaMemcpyDeviceToHost));
    return XRT_OK;
  }
  switch (what) {
    case XRT_DBG_EE_CONTACTS: {
      // records exist only where the active bit is set; the rest reads as "no contact"
      // like the reference's initial value (dist = SCALAR_MAX, PBDCollision.cu:105)
      std::vector<uint32_t> bits((E + 31) / 32);
      if (need) XRT_CUDA_OK(cudaMemcpy(dst, s->c.ee, need, cudaMemcpyDeviceToHost));
      if (E) XRT_CUDA_OK(cudaMemcpy(bits.data(), s->c.e_active, bits.size() * 4, cudaMemcpyDeviceToHost));
      EEContact* e = (EEContact*)dst;
      for (size_t i = 0; i < E; i++)
        if (!((bits[i >> 5] >> (i & 31)) & 1u)) {
          memset(&e[i], 0, sizeof(EEContact));
          e[i].dist = XRT_SCALAR_MAX;
        }
      break;
    }
    case XRT_DBG_COLORS_STRETCH: memcpy(dst, s->s_color.data(), need); break;
    case XRT_DBG_COLORS_BEND: memcpy(dst, s->b_color.data(), need); break;
    case XRT_DBG_AABBS: XRT_CUDA_OK(download_aabbs(s->bvh, F, (float*)dst)); break;
    case XRT_DBG_PREDICTED: return download_xyz(s, s->c.x, V, (float*)dst);
    case XRT_DBG_VELOCITY: return download_xyz(s, s->c.v, V, (float*)dst);
    case XRT_DBG_OBSTACLE_PRED: {
      std::vector<float4> t(3 * (size_t)s->o.F);
      XRT_CUDA_OK(cudaMemcpy(t.data(), s->bvh.pred, t.size() * sizeof(float4), cudaMemcpyDeviceToHost));
      float* o = (float*)dst;
      for (size_t i = 0; i < t.size(); i++) { o[3 * i] = t[i].x; o[3 * i + 1] = t[i].y; o[3 * i + 2] = t[i].z; }
      break;
    }
    case XRT_DBG_PARTICLE_HASH:
      XRT_CUDA_OK(cudaMemcpy(dst, s->c.p_hash, V * 4, cudaMemcpyDeviceToHost));
      XRT_CUDA_OK(cudaMemcpy((char*)dst + V * 4, s->c.p_index, V * 4, cudaMemcpyDeviceToHost));
      break;
    case XRT_DBG_STATS: {
      memset(&stats, 0, sizeof(stats));
      std::vector<float> x(3 * V);
      int rc = download_xyz(s, s->c.x0, V, x.data());
      if (rc) return rc;
      auto P = [&](uint32_t i) { return f3{x[3 * i], x[3 * i + 1], x[3 * i + 2]}; };
      double acc = 0, mx = 0;
      const size_t ns = s->s_rest.size();
      for (size_t i = 0; i < ns; i++) {
        const double r = s->s_rest[i];
        const double e = r > 0 ? (len3(P(s->s_idx[2 * i]) - P(s->s_idx[2 * i + 1])) - r) / r : 0.0;
        acc += e * e;
        mx = std::max(mx, std::fabs(e));
      }
      stats.stretch_rms = ns ? std::sqrt(acc / ns) : 0;
      stats.stretch_max = mx;
      acc = 0;
      const size_t nb = s->b_theta.size();
      for (size_t i = 0; i < nb; i++) {
        const uint32_t* q = &s->b_idx[4 * i];
        const f3 p1 = P(q[0]), p2 = P(q[1]) - p1, p3 = P(q[2]) - p1, p4 = P(q[3]) - p1;
        const f3 n1 = normalize3(cross3(p2, p3)), n2 = normalize3(cross3(p2, p4));
        const float d = clampf(dot3(n1, n2), 0.f, 1.f);
        const double e = std::isnan(d) ? 0.0 : std::acos((double)d) - s->b_theta[i];
        acc += e * e;
      }
      stats.bend_rms = nb ? std::sqrt(acc / nb) : 0;
      if (s->o.F) {
        std::vector<VFContact> vf(V);
        std::vector<float4> pr(3 * (size_t)s->o.F);
        XRT_CUDA_OK(cudaMemcpy(vf.data(), s->c.vf, V * sizeof(VFContact), cudaMemcpyDeviceToHost));
        XRT_CUDA_OK(cudaMemcpy(pr.data(), s->bvh.pred, pr.size() * sizeof(float4), cudaMemcpyDeviceToHost));
        for (size_t i = 0; i < V; i++) {
          if (vf[i].index < 0) continue;
          const size_t o = 3 * (size_t)vf[i].index;
          const f3 nn = face_normal(mk3(pr[o]), mk3(pr[o + 1]), mk3(pr[o + 2]));
          if (dot3(nn, P((uint32_t)i) - f3{vf[i].q[0], vf[i].q[1], vf[i].q[2]}) < 0.f) stats.vf_penetrations++;
        }
      }
      stats.n_stretch_colors = s->c.n_sc;
      stats.n_bend_colors = s->c.n_bc;
      const int moved = s->graph[1] && !s->graph[0] ? 1 : 0;
      stats.launches_per_frame = s->launches[moved];
      stats.bytes_per_frame = xrt_frame_bytes(s, moved != 0);
      memcpy(dst, &stats, sizeof(stats));
      break;
    }
    default: break;
  }
  return XRT_OK;
}

extern "C" int xrt_debug_candidates(xrt_solver* s, int kind, uint32_t* offsets, uint32_t* items, size_t* n_items) {
  XRT_READY();
  XRT_REQUIRE(kind >= 0 && kind <= 2 && offsets && n_items, "debug_candidates: bad argument");
  XRT_REQUIRE(s->o.F, "no obstacle");
  const uint32_t n = kind == 2 ? s->c.E : s->c.V;
  uint32_t* d_counts = kind == 0 ? s->c.cnt_vf : (kind == 1 ? s->c.cnt_rt : s->c.cnt_ee);
  launch_candidates(s->c, s->bvh, kind, nullptr, nullptr, d_counts, s->st);
  std::vector<uint32_t> cnt(n);
  XRT_CUDA_OK(cudaStreamSynchronize(s->st));
  if (n) XRT_CUDA_OK(cudaMemcpy(cnt.data(), d_counts, n * 4, cudaMemcpyDeviceToHost));
  offsets[0] = 0;
  for (uint32_t i = 0; i < n; i++) offsets[i + 1] = offsets[i] + cnt[i];
  const size_t total = offsets[n];
  const size_t cap = *n_items;
  *n_items = total;
  if (!items) return XRT_OK;
  XRT_REQUIRE(cap >= total, "debug_candidates: items too small");
  uint32_t *d_off = nullptr, *d_items = nullptr;
  XRT_CUDA_OK(cudaMalloc((void**)&d_off, (n + 1) * 4));
  XRT_CUDA_OK(cudaMalloc((void**)&d_items, (total ? total : 1) * 4));
  XRT_CUDA_OK(cudaMemcpy(d_off, offsets, (n + 1) * 4, cudaMemcpyHostToDevice));
  launch_candidates(s->c, s->bvh, kind, d_off, d_items, nullptr, s->st);
  XRT_CUDA_OK(cudaStreamSynchronize(s->st));
  if (total) XRT_CUDA_OK(cudaMemcpy(items, d_items, total * 4, cudaMemcpyDeviceToHost));
  cudaFree(d_off);
  cudaFree(d_items);
  return XRT_OK;
}

extern "C" void xrt_destroy(xrt_solver* s) {
  if (!s) return;
  cudaSetDevice(s->p.device);
  if (s->st) cudaStreamSynchronize(s->st);
  for (int i = 0; i < 2; i++) {
    if (s->graph[i]) cudaGraphExecDestroy(s->graph[i]);
    if (s->post_graph[i]) cudaGraphExecDestroy(s->post_graph[i]);
  }
  for (cudaGraphExec_t g : s->pre_graph)
    if (g) cudaGraphExecDestroy(g);
  xrt_quality_destroy(s);
  ClothDev& c = s->c;
  void* ptrs[] = {c.x0, c.x, c.v, c.n, c.rest, c.tri, c.edge, c.vf_off, c.vf_idx, c.s_idx, c.s_rest, c.s_lambda,
                  c.s_perm, c.b_idx, c.b_theta, c.l_src, c.l_tgt, c.l_rest, c.delta, c.vf, c.rt, c.ee, c.p_hash,
                  c.p_index, c.p_hash_b, c.p_index_b, c.cell_start, c.cell_end, c.neighbors, c.hash_tmp, c.cnt_vf, c.v_active, c.e_active, s->d_prof, s->d_grid_bar, c.bd_idx, c.bd_par, c.bd_k,
                  c.cnt_rt, c.cnt_ee, s->o.x0, s->o.x, s->o.x1, s->o.n, s->o.n1, s->o.tri, s->o.vf_off,
                  s->o.vf_idx, s->o.rtri, s->d_s_off, s->d_b_off, s->d_s_cta_off, s->d_b_cta_off, s->d_stage, s->d_seq, s->d_up, s->d_dn};
  for (void* p : ptrs) cudaFree(p);
  if (s->bvh.n_obj) bvh_free(&s->bvh, &s->bw);
  if (s->stage) cudaFreeHost(s->stage);
  if (s->h_up) cudaFreeHost(s->h_up);
  if (s->h_dn) cudaFreeHost(s->h_dn);
  if (s->ev_up) cudaEventDestroy(s->ev_up);
  if (s->ev0) cudaEventDestroy(s->ev0);
  if (s->ev1) cudaEventDestroy(s->ev1);
  if (s->ev_join) cudaEventDestroy(s->ev_join);
  if (s->st2) cudaStreamDestroy(s->st2);
  if (s->ev_fork) cudaEventDestroy(s->ev_fork);
  if (s->ev_prep) cudaEventDestroy(s->ev_prep);
  if (s->ev_side) cudaEventDestroy(s->ev_side);
  if (s->ev_gv) cudaEventDestroy(s->ev_gv);
  cudaFree(s->pred_alt);
  xrt_rig_destroy(s);
  if (s->st) cudaStreamDestroy(s->st);
  delete s;
}

extern "C" int xrt_bvh_build_standalone(const float* positions, uint32_t n_verts, const uint32_t* indices,
                                        uint32_t n_faces, int morton_fast_div, int device, uint32_t* morton,
                                        uint32_t* sorted_object, uint32_t* nodes, float* aabbs, int32_t* esc_int,
                                        int32_t* esc_ext, float* build_ms) {
  if (!positions || !indices || !n_verts || !n_faces) return XRT_ERR_INVALID;
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || device < 0 || device >= ndev) return XRT_ERR_CUDA;
  if (cudaSetDevice(device) != cudaSuccess) return XRT_ERR_CUDA;
  std::vector<float4> hx(n_verts);
  for (uint32_t i = 0; i < n_verts; i++) hx[i] = make_float4(positions[3 * i], positions[3 * i + 1], positions[3 * i + 2], 0.f);
  float4* dx = nullptr;
  uint32_t* dt = nullptr;
  DevBVH b{};
  BvhWork w;
  cudaEvent_t e0, e1;
  int rc = XRT_OK;
  const size_t F = n_faces;
  if (cudaMalloc((void**)&dx, n_verts * sizeof(float4)) != cudaSuccess ||
      cudaMalloc((void**)&dt, 3 * F * 4) != cudaSuccess || bvh_alloc(&b, &w, n_faces) != 0)
    rc = XRT_ERR_CUDA;
  if (rc == XRT_OK) {
    cudaMemcpy(dx, hx.data(), n_verts * sizeof(float4), cudaMemcpyHostToDevice);
    cudaMemcpy(dt, indices, 3 * F * 4, cudaMemcpyHostToDevice);
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    bvh_build(b, w, dx, dt, n_faces, morton_fast_div, 0);  // warm-up (also sizes cub state)
    cudaEventRecord(e0, 0);
    bvh_build(b, w, dx, dt, n_faces, morton_fast_div, 0);
    cudaEventRecord(e1, 0);
    if (cudaDeviceSynchronize() != cudaSuccess) rc = XRT_ERR_CUDA;
    if (rc == XRT_OK && build_ms) cudaEventElapsedTime(build_ms, e0, e1);
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
  }
  if (rc == XRT_OK) {
    if (morton) cudaMemcpy(morton, b.morton, F * 4, cudaMemcpyDeviceToHost);
    if (sorted_object) cudaMemcpy(sorted_object, b.sorted_obj, F * 4, cudaMemcpyDeviceToHost);
    if (nodes) cudaMemcpy(nodes, b.nodes, (2 * F - 1) * sizeof(BVNode), cudaMemcpyDeviceToHost);
    if (esc_int && F > 1) cudaMemcpy(esc_int, b.esc_int, (F - 1) * 4, cudaMemcpyDeviceToHost);
    if (esc_ext) cudaMemcpy(esc_ext, b.esc_ext, F * 4, cudaMemcpyDeviceToHost);
    if (aabbs && download_aabbs(b, F, aabbs) != cudaSuccess) rc = XRT_ERR_CUDA;
    if (cudaGetLastError() != cudaSuccess) rc = XRT_ERR_CUDA;
  }
  cudaFree(dx);
  cudaFree(dt);
  if (b.n_obj) bvh_free(&b, &w);
  return rc;
}
