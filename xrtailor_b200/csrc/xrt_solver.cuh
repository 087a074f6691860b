// Device-side view of one solver instance (all SoA, all preallocated at
// xrt_finalize_setup) and the launch wrappers of the step kernels.
#pragma once
#include "xrt_bvh.cuh"

struct ClothDev {
  uint32_t V = 0, E = 0, F = 0;
  float4* x0 = nullptr;   // xyz = position at substep start, w = inv_mass   (Node::x0, inv_mass)
  float4* x = nullptr;    // xyz = predicted position,        w = inv_mass   (Node::x)
  float4* v = nullptr;    // xyz = velocity                                  (Node::v)
  float4* n = nullptr;    // vertex normals                                  (Node::n)
  float4* rest = nullptr; // initial positions for the self-contact filter (SpatialHashGPU::initial_positions)
  uint32_t* tri = nullptr;   // [3F]
  uint2* edge = nullptr;     // [E]  mesh edges (EE contact queries)
  // vertex -> incident faces CSR, for atomic-free normals
  uint32_t *vf_off = nullptr, *vf_idx = nullptr;
  // stretch constraints, permuted so each colour is one contiguous range
  uint32_t n_stretch = 0, n_sc = 0;
  uint2* s_idx = nullptr;
  float* s_rest = nullptr;
  float* s_lambda = nullptr;
  uint32_t* s_perm = nullptr;  // permuted slot -> original constraint id
  // bend constraints, same
  uint32_t n_bend = 0, n_bc = 0;
  uint4* b_idx = nullptr;
  float* b_theta = nullptr;
  // geodesic long-range attachments
  uint32_t n_lra = 0;
  uint32_t *l_src = nullptr, *l_tgt = nullptr;
  float* l_rest = nullptr;
  // bindings (BindingItem): idx = cloth vertex + 3 obstacle vertices, par = barycentric u,v,w + rest length
  uint32_t n_bind = 0;
  uint4* bd_idx = nullptr;
  float4* bd_par = nullptr;
  float* bd_k = nullptr;
  // Jacobi accumulators (MemoryPool::deltas / delta_counts)
  float4* delta = nullptr;  // xyz = sum, w = count
  // contacts
  VFContact* vf = nullptr;  // [V]
  VFContact* rt = nullptr;  // [V]
  EEContact* ee = nullptr;  // [E]  (record valid only where the e_active bit is set)
  uint32_t* v_active = nullptr;  // [ceil(V/32)] bit i: vertex i has a VF or RT contact
  uint32_t* e_active = nullptr;  // [ceil(E/32)] bit i: edge i has an EE contact
  // spatial hash (SpatialHashGPU members)
  uint32_t *p_hash = nullptr, *p_index = nullptr, *p_hash_b = nullptr, *p_index_b = nullptr;
  uint32_t *cell_start = nullptr, *cell_end = nullptr, *neighbors = nullptr;
  void* hash_tmp = nullptr;
  size_t hash_tmp_bytes = 0;
  // debug candidate counts
  uint32_t *cnt_vf = nullptr, *cnt_rt = nullptr, *cnt_ee = nullptr;
};

struct ObstacleDev {
  uint32_t V = 0, F = 0;
  float4* x0 = nullptr;  // frame-start positions (Node::x0)
  float4* x = nullptr;   // frame-end target      (Node::x)
  float4* x1 = nullptr;  // interpolated midstep  (Node::x1)
  float4* n = nullptr;   // normals               (Node::n)
  float4* n1 = nullptr;  // midstep normals       (Node::n1)
  uint32_t* tri = nullptr;
  uint32_t *vf_off = nullptr, *vf_idx = nullptr;
  uint32_t* rtri = nullptr;  // representative-triangle masks
};

struct StepConst {
  float dt;  // substep time
  float gx, gy, gz;
  float damping, relaxation, lra_stretchiness, friction;
  float stretch_compliance, bend_compliance;
  float particle_diameter;
  float cell_spacing;
  int table_size, max_neighbors;
};

// ---- dynamics (xrt_dynamics.cu) ----
void launch_predict(const ClothDev& c, const StepConst& k, cudaStream_t st);
void launch_finalize(const ClothDev& c, const StepConst& k, cudaStream_t st);
void launch_stretch_color(const ClothDev& c, const StepConst& k, uint32_t begin, uint32_t end, int iter,
                          bool atomic, cudaStream_t st);
void launch_bend_color(const ClothDev& c, const StepConst& k, uint32_t begin, uint32_t end, bool atomic,
                       cudaStream_t st);
void launch_lra(const ClothDev& c, const StepConst& k, cudaStream_t st);
void launch_binding(const ClothDev& c, const ObstacleDev& o, cudaStream_t st);
void launch_apply_deltas(const ClothDev& c, const StepConst& k, cudaStream_t st);
void launch_normals(const float4* x, float4* n, const uint32_t* tri, const uint32_t* vf_off,
                    const uint32_t* vf_idx, uint32_t V, cudaStream_t st);
// ABI boundary conversions: float[3n] -> float4[n] (w kept) and two float4[n] -> float[3n] | float[3n]
void launch_unpack_xyz(const float* src, float4* dst, uint32_t n, cudaStream_t st);
void launch_pack_xyz(const float4* a, const float4* b, float* out, uint32_t n, cudaStream_t st);
void launch_interpolate(const ObstacleDev& o, int step, int num_steps, cudaStream_t st);
// end of frame for a whole batch: obstacle_->PredictedToPositions() (x = copy_src where the garment's target moved), then
// UpdateNormals of every mesh; two launches (blockIdx.y = mesh)
struct NormalsBatchItem {
  float4* x;               // positions the normals are taken on (x0)
  const float4* copy_src;  // obstacle: Node::x, copied into x first; cloth: nullptr
  float4* n;
  const uint32_t *tri, *vf_off, *vf_idx;
  uint32_t V;
  uint32_t garment;        // index into the moved flags
};
void launch_post_batched(const NormalsBatchItem* d_items, uint32_t n_items, uint32_t max_v, const int* d_moved, cudaStream_t st);
// ---- contacts (xrt_contact.cu) ----
// counts: also record the per-query leaf-hit counts (debug getters); the frame path passes false
void launch_generate_contacts(const ClothDev& c, const DevBVH& bvh, const uint32_t* rtri, bool counts, cudaStream_t st);
// the two halves of launch_generate_contacts, for the frame graph that runs them on two branches
void launch_generate_vertex_contacts(const ClothDev& c, const DevBVH& bvh, cudaStream_t st);
void launch_generate_edge_contacts(const ClothDev& c, const DevBVH& bvh, const uint32_t* rtri, cudaStream_t st);
void launch_solve_contacts(const ClothDev& c, const DevBVH& bvh, const StepConst& k, cudaStream_t st);
void launch_candidates(const ClothDev& c, const DevBVH& bvh, int kind, const uint32_t* offsets,
                       uint32_t* items, uint32_t* counts, cudaStream_t st);

// ---- self contact (xrt_hash.cu) ----
// predict + spatial hash of every garment of a shared-launch batch in five launches (blockIdx.y = garment)
struct HashBatchItem {
  ClothDev c;
  StepConst k;
  uint32_t seg;  // where this garment's keys start in the batch's key / value buffers
};
struct HashBatch {
  uint32_t n = 0, total = 0, max_v = 0, max_table = 0;
  int bits = 0;
  HashBatchItem* items = nullptr;  // device
  uint32_t* seg_off = nullptr;
  uint32_t *key_in = nullptr, *val_in = nullptr, *key_out = nullptr, *val_out = nullptr;
  void* cub_tmp = nullptr;
  size_t cub_bytes = 0;
};
#include <vector>
int hash_batch_alloc(HashBatch* hb, const std::vector<HashBatchItem>& items);
void hash_batch_free(HashBatch* hb);
void launch_predict_hash_batched(const HashBatch& hb, cudaStream_t st);
int hash_alloc(ClothDev* c, const StepConst& k);
void launch_hash(const ClothDev& c, const StepConst& k, cudaStream_t st);
void launch_collide_particles(const ClothDev& c, const StepConst& k, cudaStream_t st);
