// LBVH build / refit for the B200 path.  Interface only; see xrt_bvh.cu.
#pragma once
#include "xrt_common.cuh"

struct BvhWork {
  uint32_t cap = 0;  // faces this scratch was sized for
  uint32_t *key_a = nullptr, *key_b = nullptr, *val_a = nullptr, *val_b = nullptr;
  void* cub_tmp = nullptr;
  size_t cub_bytes = 0;
  int* flags = nullptr;     // [F-1] arrival counters of the bottom-up pass
  float* leaf_box = nullptr;  // [6F] leaf boxes in object order (not carried through the sort)
  int* scene = nullptr;     // [6] scene box as order-preserving ints
  int2* ranges = nullptr;   // [F-1] Karras range of each internal node
};

int bvh_alloc(DevBVH* b, BvhWork* w, uint32_t n_faces);
void bvh_free(DevBVH* b, BvhWork* w);

// BVH::UpdateData(tol=0, ccd=false) + BVH::Construct(scr=0)
// (xrtailor/src/physics/broad_phase/lbvh/BVH.cu:565-573,618-722).
// x: vertex positions (float4, w ignored); tri: uint32[3F].
void bvh_build(const DevBVH& b, const BvhWork& w, const float4* x, const uint32_t* tri,
               uint32_t n_faces, int fast_div, cudaStream_t st);

// BVH::UpdateData(tol=0, ccd=true) + Construct: leaf boxes over x (Node::x0) and xp (Node::x), pred = xp
// (what CollisionStep builds, impact_zone/Entry.cu:146-155).  xp == nullptr is bvh_build.
void bvh_build_ccd(const DevBVH& b, const BvhWork& w, const float4* x, const float4* xp, const uint32_t* tri,
                   uint32_t n_faces, int fast_div, cudaStream_t st);

// BVH::Update (BVH.cu:820-852): leaf boxes grow by x (and x0 when ccd), pred = x, internal boxes follow.
void bvh_update(const DevBVH& b, const BvhWork& w, const float4* x0, const float4* x, const uint32_t* tri, int ccd,
                cudaStream_t st);

// BVH::RefitMidstep (BVH.cu:795-818): pred = x1 + tol*n1, leaf boxes grow,
// internal boxes = exact merge of children.
void bvh_refit_midstep(const DevBVH& b, const BvhWork& w, const float4* x1, const float4* n1,
                       const uint32_t* tri, float tol, cudaStream_t st);

// number of kernels one call enqueues (for the launch count the bench reports)
int bvh_build_launches();
int bvh_refit_launches();

// ---- batched build (shared-launch batches): every tree of a group in seven launches ----
struct BvhBatchItem {
  DevBVH b;
  BvhWork w;
  const float4* x;      // vertex positions the tree is built on (obstacle x0)
  const uint32_t* tri;
  uint32_t F;
  uint32_t seg;         // where this tree's keys start in the batch's key / value buffers
};
struct BvhBatch {
  uint32_t n = 0, total = 0, max_f = 0;
  BvhBatchItem* items = nullptr;   // device
  uint32_t* seg_off = nullptr;     // device [n + 1]
  uint32_t *key_in = nullptr, *val_in = nullptr, *key_out = nullptr, *val_out = nullptr;  // [total]
  void* cub_tmp = nullptr;
  size_t cub_bytes = 0;
};
#ifdef __cplusplus
#include <vector>
int bvh_batch_alloc(BvhBatch* bb, const std::vector<BvhBatchItem>& items);
#endif
void bvh_batch_free(BvhBatch* bb);
// d_moved: device int[n] or nullptr (all); trees with moved == 0 are left as they are
void bvh_build_batched(const BvhBatch& bb, const int* d_moved, int fast_div, cudaStream_t st);

// ---- face-face pair lists (xrt_pairs.cu): BVH::Traverse / TraverseSelf (BVH.cu:1149-1400) on preallocated scratch ----
struct PairWork {
  uint32_t cap = 0;  // left faces this scratch was sized for
  int* num = nullptr;
  void* scan_tmp = nullptr;
  size_t scan_bytes = 0;
};
int pairs_alloc(PairWork* w, uint32_t F_left);
void pairs_free(PairWork* w);
// Appends the (left face, right face + right_off) rows of Traverse(L, R, thickness) to `pairs` at *count (a
// device-side counter) in the reference's order and advances *count; rows beyond `cap` are dropped and bit 0 of
// *overflow is set.  Three kernels and a scan, no allocation, no host synchronisation.
void pairs_traverse(const DevBVH& L, const DevBVH& R, float thickness, const PairWork& w, uint2* pairs, uint32_t cap,
                    uint32_t* count, uint32_t right_off, uint32_t* overflow, cudaStream_t st);
