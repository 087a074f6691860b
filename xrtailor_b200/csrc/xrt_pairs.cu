// Face-face candidate pairs between two LBVHs (or one with itself): the
// broad phase that feeds the reference's Quality-mode repulsion and impact
// stages.  Restates Traverse / CountPairs_Kernel / FindPairs_Kernel and the
// *Self variants (xrtailor/src/physics/broad_phase/lbvh/BVH.cu:1149-1340):
// every leaf of the left tree queries the right tree with its own box dilated
// by `thickness` (Bounds::Overlap(b, thickness), Bounds.cu:115-128); a count
// pass, an inclusive scan and a fill pass give the pairs in the reference's
// exact order (leaf by leaf in sorted-leaf order, each leaf's hits reversed).
// Here the right tree is walked on the packed 32-byte traversal nodes (box +
// links in one sector, siblings in one 64-byte pair) and the query box is
// dilated once per leaf, not once per visited node.
#include "../../include/xrt_b200.h"
#include "xrt_bvh.cuh"
#include "xrt_dev_contact.cuh"

#include <cub/device/device_scan.cuh>
#include <algorithm>
#include <vector>

namespace {
using namespace xrtdev;

// rhs.Overlap(lhs.Dilate(thickness)) with lhs already dilated (Bounds.cu:115-128): closed boxes
__device__ __forceinline__ bool overlap_box(const Box& r, const Box& q) {
  return !(r.lx > q.ux) && !(r.ux < q.lx) && !(r.ly > q.uy) && !(r.uy < q.ly) && !(r.lz > q.uz) && !(r.uz < q.lz);
}

// base: where this traversal's pairs start in the arena (a device-side count, so two traversals chain without a
// host round trip); right_off is added to the right face id (global face numbering: cloth first); pairs beyond
// `cap` are dropped and reported by pairs_commit_kernel.
template <bool FILL>
__global__ void pairs_kernel(const TNode* __restrict__ tl, const BVNode* __restrict__ nl, uint32_t Fl,
                             const TNode* __restrict__ tr, float thickness, int* __restrict__ num,
                             uint2* __restrict__ pairs, const uint32_t* __restrict__ base, uint32_t right_off,
                             uint32_t cap) {
  const uint32_t id = blockIdx.x * blockDim.x + threadIdx.x;
  if (id >= Fl) return;
  const NodeView leaf = load_node<true>(tl, (int)tnode_slot(nl, Fl - 1 + id));
  const uint32_t obj_l = (uint32_t)~leaf.hit;
  Box q = leaf.b;  // Bounds::Dilate, Bounds.cu:95-104
  q.lx -= thickness; q.ly -= thickness; q.lz -= thickness;
  q.ux += thickness; q.uy += thickness; q.uz += thickness;
  // inclusive scan: this leaf's pairs end at num[id]-1 (the reference fills each leaf's range back to front)
  long long index = FILL ? (long long)(base ? *base : 0u) + num[id] - 1 : 0;
  int n = 0;
  int node = 0;
  while (node != -1) {
    const NodeView r = load_node<true>(tr, node);
    if (overlap_box(r.b, q)) {
      if (r.hit < 0) {  // external node
        if (FILL) {
          if (index < (long long)cap) pairs[index] = make_uint2(obj_l, (uint32_t)~r.hit + right_off);
          index--;
        }
        n++;
        node = r.miss;
      } else {
        node = r.hit;
      }
    } else {
      node = r.miss;
    }
  }
  if (!FILL) num[id] = n;
}

__global__ void pairs_commit_kernel(const int* __restrict__ num, uint32_t Fl, uint32_t* count, uint32_t cap,
                                    uint32_t* overflow) {
  const unsigned long long total = (unsigned long long)*count + (unsigned long long)num[Fl - 1];
  if (total > cap) {
    atomicOr(overflow, 1u);
    *count = cap;
  } else {
    *count = (uint32_t)total;
  }
}

struct Tree {
  float4* x = nullptr;
  uint32_t* tri = nullptr;
  DevBVH b{};
  BvhWork w;
  bool ok = false;
};

bool make_tree(Tree& t, const float* pos, uint32_t nv, const uint32_t* idx, uint32_t F, int fast_div) {
  std::vector<float4> hx(nv);
  for (uint32_t i = 0; i < nv; i++) hx[i] = make_float4(pos[3 * i], pos[3 * i + 1], pos[3 * i + 2], 0.f);
  if (cudaMalloc((void**)&t.x, nv * sizeof(float4)) != cudaSuccess) return false;
  if (cudaMalloc((void**)&t.tri, 3 * (size_t)F * 4) != cudaSuccess) return false;
  if (bvh_alloc(&t.b, &t.w, F) != 0) return false;
  cudaMemcpy(t.x, hx.data(), nv * sizeof(float4), cudaMemcpyHostToDevice);
  cudaMemcpy(t.tri, idx, 3 * (size_t)F * 4, cudaMemcpyHostToDevice);
  bvh_build(t.b, t.w, t.x, t.tri, F, fast_div, 0);
  t.ok = true;
  return true;
}
void free_tree(Tree& t) {
  cudaFree(t.x);
  cudaFree(t.tri);
  if (t.b.n_obj) bvh_free(&t.b, &t.w);
}

}  // namespace

int pairs_alloc(PairWork* w, uint32_t F_left) {
  w->cap = F_left;
  w->scan_bytes = 0;
  cub::DeviceScan::InclusiveSum(nullptr, w->scan_bytes, w->num, w->num, (int)F_left, 0);
  cudaError_t e = cudaMalloc((void**)&w->num, (size_t)(F_left ? F_left : 1) * sizeof(int));
  if (e == cudaSuccess) e = cudaMalloc(&w->scan_tmp, w->scan_bytes ? w->scan_bytes : 16);
  return (int)e;
}
void pairs_free(PairWork* w) {
  cudaFree(w->num);
  cudaFree(w->scan_tmp);
  *w = PairWork{};
}
void pairs_traverse(const DevBVH& L, const DevBVH& R, float thickness, const PairWork& w, uint2* pairs, uint32_t cap,
                    uint32_t* count, uint32_t right_off, uint32_t* overflow, cudaStream_t st) {
  const uint32_t Fl = L.n_obj;
  if (!Fl || !R.n_obj) return;
  pairs_kernel<false><<<xrt_grid(Fl), XRT_BLOCK, 0, st>>>(L.tnodes, L.nodes, Fl, R.tnodes, thickness, w.num, nullptr, nullptr,
                                                         0u, 0u);
  size_t bytes = w.scan_bytes;
  cub::DeviceScan::InclusiveSum(w.scan_tmp, bytes, w.num, w.num, (int)Fl, st);
  pairs_kernel<true><<<xrt_grid(Fl), XRT_BLOCK, 0, st>>>(L.tnodes, L.nodes, Fl, R.tnodes, thickness, w.num, pairs, count,
                                                        right_off, cap);
  pairs_commit_kernel<<<1, 1, 0, st>>>(w.num, Fl, count, cap, overflow);
}

extern "C" int xrt_traverse_pairs_standalone(const float* pos_l, uint32_t nv_l, const uint32_t* tri_l, uint32_t F_l,
                                             const float* pos_r, uint32_t nv_r, const uint32_t* tri_r, uint32_t F_r,
                                             float thickness, int morton_fast_div, int device, uint32_t* pairs,
                                             size_t capacity, size_t* n_pairs, float* traverse_ms) {
  if (!pos_l || !tri_l || !nv_l || !F_l || !n_pairs) return XRT_ERR_INVALID;
  const bool self = (F_r == 0);
  if (!self && (!pos_r || !tri_r || !nv_r)) return XRT_ERR_INVALID;
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || device < 0 || device >= ndev) {
    (void)cudaGetLastError();
    return XRT_ERR_CUDA;  // no CPU fallback
  }
  if (cudaSetDevice(device) != cudaSuccess) return XRT_ERR_CUDA;
  Tree L, R;
  int* num = nullptr;
  uint2* d_pairs = nullptr;
  void* scan_tmp = nullptr;
  int rc = XRT_OK;
  if (!make_tree(L, pos_l, nv_l, tri_l, F_l, morton_fast_div)) rc = XRT_ERR_CUDA;
  if (rc == XRT_OK && !self && !make_tree(R, pos_r, nv_r, tri_r, F_r, morton_fast_div)) rc = XRT_ERR_CUDA;
  const Tree& Rt = self ? L : R;
  size_t scan_bytes = 0;
  if (rc == XRT_OK) {
    cub::DeviceScan::InclusiveSum(nullptr, scan_bytes, num, num, (int)F_l, 0);
    if (cudaMalloc((void**)&num, (size_t)F_l * sizeof(int)) != cudaSuccess ||
        cudaMalloc(&scan_tmp, scan_bytes ? scan_bytes : 16) != cudaSuccess)
      rc = XRT_ERR_CUDA;
  }
  cudaEvent_t e0 = nullptr, e1 = nullptr;
  size_t total = 0;
  if (rc == XRT_OK) {
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    cudaEventRecord(e0, 0);
    pairs_kernel<false><<<xrt_grid(F_l), XRT_BLOCK>>>(L.b.tnodes, L.b.nodes, F_l, Rt.b.tnodes, thickness, num, nullptr, nullptr, 0u, 0u);
    cub::DeviceScan::InclusiveSum(scan_tmp, scan_bytes, num, num, (int)F_l, 0);
    int last = 0;
    if (cudaMemcpy(&last, num + (F_l - 1), sizeof(int), cudaMemcpyDeviceToHost) != cudaSuccess) rc = XRT_ERR_CUDA;
    total = (size_t)(last < 0 ? 0 : last);
    if (rc == XRT_OK && total && cudaMalloc((void**)&d_pairs, total * sizeof(uint2)) != cudaSuccess) rc = XRT_ERR_CUDA;
    if (rc == XRT_OK && total)
      pairs_kernel<true><<<xrt_grid(F_l), XRT_BLOCK>>>(L.b.tnodes, L.b.nodes, F_l, Rt.b.tnodes, thickness, num, d_pairs, nullptr, 0u,
                                                         (uint32_t)std::min<size_t>(total, 0xFFFFFFFFu));
    cudaEventRecord(e1, 0);
    if (cudaDeviceSynchronize() != cudaSuccess) rc = XRT_ERR_CUDA;
    if (rc == XRT_OK && traverse_ms) cudaEventElapsedTime(traverse_ms, e0, e1);
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
  }
  if (rc == XRT_OK) {
    *n_pairs = total;
    if (pairs && total) {
      if (capacity < total)
        rc = XRT_ERR_INVALID;  // *n_pairs tells the caller how much to allocate
      else if (cudaMemcpy(pairs, d_pairs, total * sizeof(uint2), cudaMemcpyDeviceToHost) != cudaSuccess)
        rc = XRT_ERR_CUDA;
    }
  }
  cudaFree(num);
  cudaFree(d_pairs);
  cudaFree(scan_tmp);
  free_tree(L);
  if (!self) free_tree(R);
  (void)cudaGetLastError();
  return rc;
}
