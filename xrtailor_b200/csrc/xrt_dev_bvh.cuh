// Device bodies of the LBVH bottom-up box pass and the midstep refit, shared by
// xrt_bvh.cu (per-stage kernels) and xrt_persist.cu (persistent kernel).
#pragma once
#include "xrt_common.cuh"

namespace xrtdev {

__device__ __forceinline__ void tnode_store_box(TNode* t, const float lo[3], const float hi[3]) {
  // box lives in the first 24 bytes: one float4 + one float2 store
  reinterpret_cast<float4*>(t)[0] = make_float4(lo[0], lo[1], lo[2], hi[0]);
  reinterpret_cast<float2*>(t)[2] = make_float2(hi[1], hi[2]);
}
__device__ __forceinline__ void tnode_load_box(const TNode* t, float lo[3], float hi[3]) {
  // L2 loads: sibling boxes written by other SMs must not come from a stale L1 line
  float4 a = __ldcg(reinterpret_cast<const float4*>(t));
  float2 b = __ldcg(reinterpret_cast<const float2*>(t) + 2);
  lo[0] = a.x; lo[1] = a.y; lo[2] = a.z;
  hi[0] = a.w; hi[1] = b.x; hi[2] = b.y;
}

// Bottom-up boxes in one pass (replaces AssignAABBAtomic_Kernel's
// O(F log F) CAS walk, BVH.cu:540-556; min/max are exact so boxes are
// identical).  The second thread to arrive at a parent merges its children.
// PARITY: the counters are never cleared; a parent is reached exactly twice per
// pass, so "first arrival" is "old value even" (saves the memset the per-stage
// path issues before each pass).
// ONE_CTA: every thread of the pass is in one CTA (one SM per garment), so the box store only
// has to be ordered before the counter at CTA scope -- no MEMBAR.GPU per level.
template <bool PARITY, bool ONE_CTA = false>
__device__ __forceinline__ void walk_up(TNode* tn, const BVNode* nodes, int* flags, uint32_t leaf_node) {
  uint32_t parent = nodes[leaf_node].parent;
  while (parent != 0xFFFFFFFFu) {
    if (ONE_CTA)
      __threadfence_block();
    else
      __threadfence();
    const int old = atomicAdd(flags + parent, 1);
    if (PARITY ? ((old & 1) == 0) : (old == 0)) return;
    float la[3], ha[3], lb[3], hb[3];
    tnode_load_box(tn + tnode_child_slot(parent, false), la, ha);
    tnode_load_box(tn + tnode_child_slot(parent, true), lb, hb);
#pragma unroll
    for (int k = 0; k < 3; k++) {
      la[k] = fminf(la[k], lb[k]);
      ha[k] = fmaxf(ha[k], hb[k]);
    }
    const uint32_t gp = nodes[parent].parent;
    const uint32_t slot = gp == 0xFFFFFFFFu ? 0u : tnode_child_slot(gp, nodes[gp].right == parent);
    tnode_store_box(tn + slot, la, ha);
    parent = gp;
  }
}

// Interpolated obstacle -> primitive pred + grow-only leaf boxes + exact
// internal merge (RefitExternalNodeMidstep_Kernel + AssignAABBCAS_Kernel,
// BVH.cu:770-793,515-538) for leaf i.
template <bool PARITY, bool ONE_CTA = false>
__device__ __forceinline__ void refit_item(TNode* tn, const BVNode* nodes, float4* pred, const float4* x1,
                                           const float4* n1, const uint32_t* tri, float tol, int* flags,
                                           uint32_t F, uint32_t i) {
  const uint32_t node = F - 1 + i;
  const uint32_t o = nodes[node].object;
  TNode* leaf = tn + tnode_slot(nodes, node);
  float lo[3], hi[3];
  tnode_load_box(leaf, lo, hi);
#pragma unroll
  for (int k = 0; k < 3; k++) {
    const uint32_t v = tri[3 * o + k];
    const float4 p = x1[v], nn = n1[v];
    const float4 q = make_float4(p.x + tol * nn.x, p.y + tol * nn.y, p.z + tol * nn.z, 0.f);
    pred[3 * o + k] = q;
    lo[0] = fminf(lo[0], q.x); lo[1] = fminf(lo[1], q.y); lo[2] = fminf(lo[2], q.z);
    hi[0] = fmaxf(hi[0], q.x); hi[1] = fmaxf(hi[1], q.y); hi[2] = fmaxf(hi[2], q.z);
  }
  tnode_store_box(leaf, lo, hi);
  walk_up<PARITY, ONE_CTA>(tn, nodes, flags, node);
}

}  // namespace xrtdev
