// LBVH for sm_100a: fused leaf-box + scene-box kernel, Morton quantiser,
// cub radix sort of (key,index) only, Karras hierarchy on (morton,object)
// pairs, single-pass bottom-up boxes with arrival counters, O(1) escape links
// from the Karras ranges, nodes packed to one 32-byte sector for traversal.
// Behaviour follows xrtailor/src/physics/broad_phase/lbvh/BVH.cu:22-46
// (morton), :57-210 (determine_range / find_split), :212-264 (escape tables),
// :423-462 (leaf boxes), :618-722 (Construct), :770-818 (RefitMidstep).
#include "xrt_bvh.cuh"
#include "xrt_dev_bvh.cuh"
#include <cub/device/device_radix_sort.cuh>
#include <cub/device/device_segmented_radix_sort.cuh>
#include <algorithm>
#include <vector>

namespace {
using namespace xrtdev;

// order-preserving float <-> int map so atomicMin/Max work on floats
__device__ __forceinline__ int f2ord(float f) {
  int i = __float_as_int(f);
  return i >= 0 ? i : i ^ 0x7FFFFFFF;
}
__device__ __forceinline__ float ord2f(int i) { return __int_as_float(i >= 0 ? i : i ^ 0x7FFFFFFF); }

__device__ __forceinline__ void scene_init_body(int* scene) {
  if (threadIdx.x < 3) scene[threadIdx.x] = f2ord(INFINITY);
  else if (threadIdx.x < 6) scene[threadIdx.x] = f2ord(-INFINITY);
}
__global__ void scene_init_kernel(int* scene) { scene_init_body(scene); }

// UpdateObjects_Kernel + InitAABBs_Kernel + thrust::reduce (BVH.cu:382-462,642)
// in one pass over the faces.
__device__ __forceinline__ void leaf_box_body(const float4* __restrict__ x, const float4* __restrict__ xp,
                                              const uint32_t* __restrict__ tri, float4* __restrict__ pred,
                                              float* __restrict__ leaf_box, int* __restrict__ scene, uint32_t F, uint32_t i) {
  float lo[3] = {INFINITY, INFINITY, INFINITY}, hi[3] = {-INFINITY, -INFINITY, -INFINITY};
  if (i < F) {
    const uint32_t i0 = tri[3 * i], i1 = tri[3 * i + 1], i2 = tri[3 * i + 2];
    float4 a = x[i0], b = x[i1], c = x[i2];
    lo[0] = fminf(a.x, fminf(b.x, c.x));
    lo[1] = fminf(a.y, fminf(b.y, c.y));
    lo[2] = fminf(a.z, fminf(b.z, c.z));
    hi[0] = fmaxf(a.x, fmaxf(b.x, c.x));
    hi[1] = fmaxf(a.y, fmaxf(b.y, c.y));
    hi[2] = fmaxf(a.z, fmaxf(b.z, c.z));
    if (xp) {  // ccd: pred = Node::x, box over v and pred (UpdateObjects_Kernel / InitAABBs_Kernel, BVH.cu:382-462)
      a = xp[i0]; b = xp[i1]; c = xp[i2];
      lo[0] = fminf(lo[0], fminf(a.x, fminf(b.x, c.x)));
      lo[1] = fminf(lo[1], fminf(a.y, fminf(b.y, c.y)));
      lo[2] = fminf(lo[2], fminf(a.z, fminf(b.z, c.z)));
      hi[0] = fmaxf(hi[0], fmaxf(a.x, fmaxf(b.x, c.x)));
      hi[1] = fmaxf(hi[1], fmaxf(a.y, fmaxf(b.y, c.y)));
      hi[2] = fmaxf(hi[2], fmaxf(a.z, fmaxf(b.z, c.z)));
    }
    pred[3 * i] = a;
    pred[3 * i + 1] = b;
    pred[3 * i + 2] = c;
    float* o = leaf_box + 6 * (size_t)i;
    o[0] = lo[0]; o[1] = lo[1]; o[2] = lo[2];
    o[3] = hi[0]; o[4] = hi[1]; o[5] = hi[2];
  }
  // warp shuffle reduce, then one atomic per warp per component
#pragma unroll
  for (int k = 0; k < 3; k++) {
    float l = lo[k], h = hi[k];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      l = fminf(l, __shfl_xor_sync(0xFFFFFFFFu, l, o));
      h = fmaxf(h, __shfl_xor_sync(0xFFFFFFFFu, h, o));
    }
    if ((threadIdx.x & 31) == 0) {
      atomicMin(scene + k, f2ord(l));
      atomicMax(scene + 3 + k, f2ord(h));
    }
  }
}
__global__ void leaf_box_kernel(const float4* __restrict__ x, const float4* __restrict__ xp,
                                const uint32_t* __restrict__ tri, float4* __restrict__ pred,
                                float* __restrict__ leaf_box, int* __restrict__ scene, uint32_t F) {
  leaf_box_body(x, xp, tri, pred, leaf_box, scene, F, blockIdx.x * blockDim.x + threadIdx.x);
}

__device__ __forceinline__ uint32_t expand_bits10(uint32_t v) {  // BVH.cu:18-24
  v = (v * 0x00010001u) & 0xFF0000FFu;
  v = (v * 0x00000101u) & 0x0F00F00Fu;
  v = (v * 0x00000011u) & 0xC30C30C3u;
  v = (v * 0x00000005u) & 0x49249249u;
  return v;
}
__device__ __forceinline__ uint32_t quant10(float t) {
  t = t * 1024.0f;
  t = t > 0.0f ? t : 0.0f;        // MathFunctions::max(a,b) = a > b ? a : b  (NaN -> 0)
  t = t < 1023.0f ? t : 1023.0f;  // MathFunctions::min
  return (uint32_t)t;
}

// ComputeMortons_Kernel + InitObjectIndices_Kernel (BVH.cu:464-487)
template <bool FAST_DIV>
__device__ __forceinline__ void morton_body(const float* __restrict__ leaf_box, const int* __restrict__ scene,
                                            uint32_t* __restrict__ key, uint32_t* __restrict__ val, uint32_t F, uint32_t i) {
  if (i >= F) return;
  const float* b = leaf_box + 6 * (size_t)i;
  float wl[3] = {ord2f(scene[0]), ord2f(scene[1]), ord2f(scene[2])};
  float wu[3] = {ord2f(scene[3]), ord2f(scene[4]), ord2f(scene[5])};
  uint32_t q[3];
#pragma unroll
  for (int k = 0; k < 3; k++) {
    float p = 0.5f * (b[k] + b[3 + k]);
    p = __fsub_rn(p, wl[k]);
    float d = __fsub_rn(wu[k], wl[k]);
    p = FAST_DIV ? __fdividef(p, d) : __fdiv_rn(p, d);
    q[k] = quant10(p);
  }
  key[i] = expand_bits10(q[0]) * 4 + expand_bits10(q[1]) * 2 + expand_bits10(q[2]);
  val[i] = i;
}
template <bool FAST_DIV>
__global__ void morton_kernel(const float* __restrict__ leaf_box, const int* __restrict__ scene,
                              uint32_t* __restrict__ key, uint32_t* __restrict__ val, uint32_t F) {
  morton_body<FAST_DIV>(leaf_box, scene, key, val, F, blockIdx.x * blockDim.x + threadIdx.x);
}

// common_upper_bits on the reference's 64-bit key (morton << 32 | object),
// evaluated from the two halves.  For distinct Morton codes it equals the
// 32-bit __clz the reference uses when all codes are unique (BVH.cu:47-55), so
// one code path gives the reference's tree in both of its modes.
__device__ __forceinline__ int delta_pair(uint32_t m0, uint32_t o0, uint32_t m1, uint32_t o1) {
  return m0 != m1 ? __clz(m0 ^ m1) : 32 + __clz(o0 ^ o1);
}

// ConstructLeafNodes_Kernel + ConstructInternalNodes64Bits_Kernel (BVH.cu:297-318,489-500)
__device__ __forceinline__ void karras_body(const uint32_t* __restrict__ morton, const uint32_t* __restrict__ obj,
                                            BVNode* __restrict__ nodes, int2* __restrict__ ranges, uint32_t F, uint32_t tid) {
  if (tid >= F) return;
  const int n = (int)F;
  // leaf tid
  nodes[F - 1 + tid].left = 0xFFFFFFFFu;
  nodes[F - 1 + tid].right = 0xFFFFFFFFu;
  nodes[F - 1 + tid].object = obj[tid];
  if (F == 1) nodes[0].parent = 0xFFFFFFFFu;
  if (tid >= F - 1) return;
  const int idx = (int)tid;
  const uint32_t mc = morton[idx], oc = obj[idx];
  auto delta = [&](int j) -> int {
    return (j < 0 || j >= n) ? -1 : delta_pair(mc, oc, morton[j], obj[j]);
  };
  int first, last;
  if (idx == 0) {
    first = 0;
    last = n - 1;
    nodes[0].parent = 0xFFFFFFFFu;
  } else {
    const int dl = delta(idx - 1), dr = delta(idx + 1);
    const int d = dr > dl ? 1 : -1;
    const int dmin = min(dl, dr);
    int lmax = 2;
    while (delta(idx + d * lmax) > dmin) lmax <<= 1;
    int l = 0;
    for (int t = lmax >> 1; t > 0; t >>= 1)
      if (delta(idx + (l + t) * d) > dmin) l += t;
    int j = idx + l * d;
    first = min(idx, j);
    last = max(idx, j);
  }
  // find_split (keys are unique by construction, so no equal-code shortcut)
  const uint32_t mf = morton[first], of = obj[first];
  const int dnode = delta_pair(mf, of, morton[last], obj[last]);
  int split = first, stride = last - first;
  do {
    stride = (stride + 1) >> 1;
    const int mid = split + stride;
    if (mid < last && delta_pair(mf, of, morton[mid], obj[mid]) > dnode) split = mid;
  } while (stride > 1);
  uint32_t L = split, R = split + 1;
  if (first == split) L += F - 1;
  if (last == split + 1) R += F - 1;
  nodes[idx].left = L;
  nodes[idx].right = R;
  nodes[idx].object = 0xFFFFFFFFu;
  nodes[L].parent = idx;
  nodes[R].parent = idx;
  ranges[idx] = make_int2(first, last);
}
__global__ void karras_kernel(const uint32_t* __restrict__ morton, const uint32_t* __restrict__ obj,
                              BVNode* __restrict__ nodes, int2* __restrict__ ranges, uint32_t F) {
  karras_body(morton, obj, nodes, ranges, F, blockIdx.x * blockDim.x + threadIdx.x);
}

__device__ __forceinline__ void build_boxes_body(TNode* tn, const BVNode* __restrict__ nodes,
                                                 const float* __restrict__ leaf_box,
                                                 const uint32_t* __restrict__ obj, int* flags, uint32_t F, uint32_t i) {
  if (i >= F) return;
  const float* b = leaf_box + 6 * (size_t)obj[i];
  float lo[3] = {b[0], b[1], b[2]}, hi[3] = {b[3], b[4], b[5]};
  tnode_store_box(tn + tnode_slot(nodes, F - 1 + i), lo, hi);
  walk_up<false>(tn, nodes, flags, F - 1 + i);
}
__global__ void build_boxes_kernel(TNode* tn, const BVNode* __restrict__ nodes,
                                   const float* __restrict__ leaf_box,
                                   const uint32_t* __restrict__ obj, int* flags, uint32_t F) {
  build_boxes_body(tn, nodes, leaf_box, obj, flags, F, blockIdx.x * blockDim.x + threadIdx.x);
}

// Escape links in O(1) from the Karras ranges: the subtree that follows a
// range ending at leaf `last` is rooted at internal node last+1 when that node
// starts at last+1, else at leaf last+1.  Equals the parent walk of
// Build{Internal,External}EscapeIndexTable_Kernel (BVH.cu:212-264).
__device__ __forceinline__ int escape_after(int last, const int2* ranges, int n) {
  if (last >= n - 1) return -1;
  const int k = last + 1;
  if (k <= n - 2 && ranges[k].x == k) return k;
  return n - 1 + k;
}

__device__ __forceinline__ void link_body(TNode* tn, const BVNode* __restrict__ nodes, const int2* __restrict__ ranges,
                                          int32_t* __restrict__ esc_int, int32_t* __restrict__ esc_ext, uint32_t F, uint32_t i) {
  const int n = (int)F;
  if (i >= 2 * F - 1) return;
  TNode* me = tn + tnode_slot(nodes, i);
  if (i < F - 1) {
    int e = escape_after(ranges[i].y, ranges, n);
    esc_int[i] = e;
    me->link_hit = (int)tnode_child_slot(i, false);
    me->link_miss = e < 0 ? -1 : (int)tnode_slot(nodes, (uint32_t)e);
  } else {
    int leaf = (int)(i - (F - 1));
    int e = escape_after(leaf, ranges, n);
    esc_ext[leaf] = e;
    me->link_hit = ~(int)nodes[i].object;
    me->link_miss = e < 0 ? -1 : (int)tnode_slot(nodes, (uint32_t)e);
  }
}
__global__ void link_kernel(TNode* tn, const BVNode* __restrict__ nodes, const int2* __restrict__ ranges,
                            int32_t* __restrict__ esc_int, int32_t* __restrict__ esc_ext, uint32_t F) {
  link_body(tn, nodes, ranges, esc_int, esc_ext, F, blockIdx.x * blockDim.x + threadIdx.x);
}

// ---- the same seven stages for a whole batch of trees: blockIdx.y = tree, one launch per stage ----
// (a shared-launch batch rebuilds 148 obstacle trees per frame; as 148 x 7 launches of less than one wave each they
// took 10 % of the frame).  Trees whose `moved` flag is 0 keep last frame's tree.
__global__ void batch_stage0_kernel(const BvhBatchItem* __restrict__ items, const int* __restrict__ moved) {
  const BvhBatchItem& it = items[blockIdx.y];
  if ((moved && !moved[blockIdx.y]) || !it.F) return;
  if (blockIdx.x == 0) scene_init_body(it.w.scene);
  // arrival counters of the bottom-up box pass
  for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < it.F; i += gridDim.x * blockDim.x) it.w.flags[i] = 0;
}
__global__ void batch_leaf_box_kernel(const BvhBatchItem* __restrict__ items, const int* __restrict__ moved) {
  const BvhBatchItem& it = items[blockIdx.y];
  if (moved && !moved[blockIdx.y]) return;
  if (blockIdx.x * blockDim.x >= it.F) return;
  leaf_box_body(it.x, nullptr, it.tri, it.b.pred, it.w.leaf_box, it.w.scene, it.F, blockIdx.x * blockDim.x + threadIdx.x);
}
template <bool FAST_DIV>
__global__ void batch_morton_kernel(const BvhBatchItem* __restrict__ items, const int* __restrict__ moved,
                                    uint32_t* __restrict__ key, uint32_t* __restrict__ val) {
  const BvhBatchItem& it = items[blockIdx.y];
  if (moved && !moved[blockIdx.y]) return;
  morton_body<FAST_DIV>(it.w.leaf_box, it.w.scene, key + it.seg, val + it.seg, it.F, blockIdx.x * blockDim.x + threadIdx.x);
}
__global__ void batch_karras_kernel(const BvhBatchItem* __restrict__ items, const int* __restrict__ moved,
                                    const uint32_t* __restrict__ key, const uint32_t* __restrict__ val) {
  const BvhBatchItem& it = items[blockIdx.y];
  if (moved && !moved[blockIdx.y]) return;
  karras_body(key + it.seg, val + it.seg, it.b.nodes, it.w.ranges, it.F, blockIdx.x * blockDim.x + threadIdx.x);
}
__global__ void batch_boxes_kernel(const BvhBatchItem* __restrict__ items, const int* __restrict__ moved,
                                   const uint32_t* __restrict__ val) {
  const BvhBatchItem& it = items[blockIdx.y];
  if (moved && !moved[blockIdx.y]) return;
  build_boxes_body(it.b.tnodes, it.b.nodes, it.w.leaf_box, val + it.seg, it.w.flags, it.F, blockIdx.x * blockDim.x + threadIdx.x);
}
__global__ void batch_link_kernel(const BvhBatchItem* __restrict__ items, const int* __restrict__ moved) {
  const BvhBatchItem& it = items[blockIdx.y];
  if (moved && !moved[blockIdx.y]) return;
  link_body(it.b.tnodes, it.b.nodes, it.w.ranges, it.b.esc_int, it.b.esc_ext, it.F, blockIdx.x * blockDim.x + threadIdx.x);
}

// RefitMidstep, one kernel (body in xrt_dev_bvh.cuh)
__global__ void refit_kernel(TNode* tn, const BVNode* __restrict__ nodes, float4* __restrict__ pred,
                             const float4* __restrict__ x1, const float4* __restrict__ n1,
                             const uint32_t* __restrict__ tri, float tol, int* flags, uint32_t F) {
  uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < F) refit_item<false>(tn, nodes, pred, x1, n1, tri, tol, flags, F, i);
}

// BVH::Update (BVH.cu:820-852): Update_Kernel grows every leaf box by Face::ComputeBounds(ccd) (x, and x0 when
// ccd) and sets pred = x; AssignAABBAtomic then merges upwards.  Leaves only grow, so "old box merged with the
// children" equals the exact merge of the children, which is what the single bottom-up pass writes.
__global__ void update_kernel(TNode* tn, const BVNode* __restrict__ nodes, float4* __restrict__ pred,
                              const float4* __restrict__ x0, const float4* __restrict__ x,
                              const uint32_t* __restrict__ tri, int ccd, int* flags, uint32_t F) {
  uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= F) return;
  const uint32_t node = F - 1 + i;
  const uint32_t o = nodes[node].object;
  TNode* leaf = tn + tnode_slot(nodes, node);
  float lo[3], hi[3];
  tnode_load_box(leaf, lo, hi);
#pragma unroll
  for (int k = 0; k < 3; k++) {
    const uint32_t v = tri[3 * o + k];
    const float4 p = x[v];
    pred[3 * o + k] = p;
    lo[0] = fminf(lo[0], p.x); lo[1] = fminf(lo[1], p.y); lo[2] = fminf(lo[2], p.z);
    hi[0] = fmaxf(hi[0], p.x); hi[1] = fmaxf(hi[1], p.y); hi[2] = fmaxf(hi[2], p.z);
    if (ccd) {
      const float4 q = x0[v];
      lo[0] = fminf(lo[0], q.x); lo[1] = fminf(lo[1], q.y); lo[2] = fminf(lo[2], q.z);
      hi[0] = fmaxf(hi[0], q.x); hi[1] = fmaxf(hi[1], q.y); hi[2] = fmaxf(hi[2], q.z);
    }
  }
  tnode_store_box(leaf, lo, hi);
  walk_up<false>(tn, nodes, flags, node);
}

}  // namespace

int bvh_alloc(DevBVH* b, BvhWork* w, uint32_t F) {
  if (F == 0) return 0;
  const size_t nn = 2 * (size_t)F - 1;
  b->n_obj = F;
  cudaError_t e = cudaSuccess;
  auto A = [&](void** p, size_t bytes) {
    if (e == cudaSuccess) e = cudaMalloc(p, bytes ? bytes : 16);
  };
  A((void**)&b->tnodes, (nn + 1) * sizeof(TNode));  // 2F sibling-pair slots
  A((void**)&b->nodes, nn * sizeof(BVNode));
  A((void**)&b->esc_int, F * sizeof(int32_t));
  A((void**)&b->esc_ext, F * sizeof(int32_t));
  A((void**)&b->pred, 3 * (size_t)F * sizeof(float4));
  A((void**)&w->key_a, F * 4); A((void**)&w->key_b, F * 4);
  A((void**)&w->val_a, F * 4); A((void**)&w->val_b, F * 4);
  A((void**)&w->flags, F * sizeof(int));
  A((void**)&w->leaf_box, 6 * (size_t)F * sizeof(float));
  A((void**)&w->scene, 8 * sizeof(int));
  A((void**)&w->ranges, F * sizeof(int2));
  if (e != cudaSuccess) return (int)e;
  cub::DoubleBuffer<uint32_t> k(w->key_a, w->key_b), v(w->val_a, w->val_b);
  w->cub_bytes = 0;
  e = cub::DeviceRadixSort::SortPairs(nullptr, w->cub_bytes, k, v, (int)F, 0, 30);
  A(&w->cub_tmp, w->cub_bytes);
  w->cap = F;
  b->morton = nullptr;
  b->sorted_obj = nullptr;
  b->aabbs = nullptr;
  return (int)e;
}

void bvh_free(DevBVH* b, BvhWork* w) {
  cudaFree(b->tnodes); cudaFree(b->nodes); cudaFree(b->esc_int); cudaFree(b->esc_ext); cudaFree(b->pred);
  cudaFree(w->key_a); cudaFree(w->key_b); cudaFree(w->val_a); cudaFree(w->val_b);
  cudaFree(w->flags); cudaFree(w->leaf_box); cudaFree(w->scene); cudaFree(w->ranges); cudaFree(w->cub_tmp);
  *b = DevBVH{};
  *w = BvhWork{};
}

int bvh_build_launches() { return 7; }   // scene_init, leaf_box, morton, sort(~2: histogram+onesweep counted as 2), karras, boxes, link  (memset excluded)
int bvh_refit_launches() { return 1; }

int bvh_batch_alloc(BvhBatch* bb, const std::vector<BvhBatchItem>& items) {
  bb->n = (uint32_t)items.size();
  bb->total = 0;
  bb->max_f = 0;
  std::vector<BvhBatchItem> h = items;
  std::vector<uint32_t> off(items.size() + 1, 0);
  for (size_t i = 0; i < h.size(); i++) {
    h[i].seg = bb->total;
    off[i] = bb->total;
    bb->total += h[i].F;
    bb->max_f = std::max(bb->max_f, h[i].F);
  }
  off[h.size()] = bb->total;
  cudaError_t e = cudaSuccess;
  auto A = [&](void** p, size_t bytes) {
    if (e == cudaSuccess) e = cudaMalloc(p, bytes ? bytes : 16);
  };
  A((void**)&bb->items, h.size() * sizeof(BvhBatchItem));
  A((void**)&bb->seg_off, off.size() * 4);
  A((void**)&bb->key_in, (size_t)bb->total * 4); A((void**)&bb->val_in, (size_t)bb->total * 4);
  A((void**)&bb->key_out, (size_t)bb->total * 4); A((void**)&bb->val_out, (size_t)bb->total * 4);
  if (e != cudaSuccess) return (int)e;
  cudaMemcpy(bb->items, h.data(), h.size() * sizeof(BvhBatchItem), cudaMemcpyHostToDevice);
  cudaMemcpy(bb->seg_off, off.data(), off.size() * 4, cudaMemcpyHostToDevice);
  bb->cub_bytes = 0;
  e = cub::DeviceSegmentedRadixSort::SortPairs(nullptr, bb->cub_bytes, bb->key_in, bb->key_out, bb->val_in, bb->val_out, (int)bb->total,
                                               (int)bb->n, bb->seg_off, bb->seg_off + 1, 0, 30);
  A(&bb->cub_tmp, bb->cub_bytes);
  return (int)e;
}
void bvh_batch_free(BvhBatch* bb) {
  cudaFree(bb->items); cudaFree(bb->seg_off); cudaFree(bb->key_in); cudaFree(bb->val_in); cudaFree(bb->key_out);
  cudaFree(bb->val_out); cudaFree(bb->cub_tmp);
  *bb = BvhBatch{};
}
// bvh_build of every tree of the batch in seven launches (the segmented sort is cub's, like the per-tree sort)
void bvh_build_batched(const BvhBatch& bb, const int* d_moved, int fast_div, cudaStream_t st) {
  if (!bb.n || !bb.total) return;
  const dim3 g(xrt_grid(bb.max_f), bb.n), g2(xrt_grid(2 * (size_t)bb.max_f), bb.n);
  batch_stage0_kernel<<<dim3(std::min<unsigned>(g.x, 8u), bb.n), XRT_BLOCK, 0, st>>>(bb.items, d_moved);
  batch_leaf_box_kernel<<<g, XRT_BLOCK, 0, st>>>(bb.items, d_moved);
  if (fast_div) batch_morton_kernel<true><<<g, XRT_BLOCK, 0, st>>>(bb.items, d_moved, bb.key_in, bb.val_in);
  else batch_morton_kernel<false><<<g, XRT_BLOCK, 0, st>>>(bb.items, d_moved, bb.key_in, bb.val_in);
  size_t bytes = bb.cub_bytes;
  // a tree that did not move re-sorts last frame's keys (key_in is kept): same order, its other stages are skipped
  cub::DeviceSegmentedRadixSort::SortPairs(bb.cub_tmp, bytes, bb.key_in, bb.key_out, bb.val_in, bb.val_out, (int)bb.total, (int)bb.n,
                                           bb.seg_off, bb.seg_off + 1, 0, 30, st);
  batch_karras_kernel<<<g, XRT_BLOCK, 0, st>>>(bb.items, d_moved, bb.key_out, bb.val_out);
  batch_boxes_kernel<<<g, XRT_BLOCK, 0, st>>>(bb.items, d_moved, bb.val_out);
  batch_link_kernel<<<g2, XRT_BLOCK, 0, st>>>(bb.items, d_moved);
}

void bvh_build(const DevBVH& b_, const BvhWork& w, const float4* x, const uint32_t* tri, uint32_t F,
               int fast_div, cudaStream_t st) {
  bvh_build_ccd(b_, w, x, nullptr, tri, F, fast_div, st);
}

void bvh_update(const DevBVH& b, const BvhWork& w, const float4* x0, const float4* x, const uint32_t* tri, int ccd,
                cudaStream_t st) {
  const uint32_t F = b.n_obj;
  if (F == 0) return;
  cudaMemsetAsync(w.flags, 0, F * sizeof(int), st);
  update_kernel<<<xrt_grid(F), XRT_BLOCK, 0, st>>>(b.tnodes, b.nodes, b.pred, x0, x, tri, ccd, w.flags, F);
}

void bvh_build_ccd(const DevBVH& b_, const BvhWork& w, const float4* x, const float4* xp, const uint32_t* tri,
                   uint32_t F, int fast_div, cudaStream_t st) {
  DevBVH& b = const_cast<DevBVH&>(b_);
  if (F == 0) return;
  const unsigned g = xrt_grid(F);
  scene_init_kernel<<<1, 32, 0, st>>>(w.scene);
  leaf_box_kernel<<<g, XRT_BLOCK, 0, st>>>(x, xp, tri, b.pred, w.leaf_box, w.scene, F);
  if (fast_div)
    morton_kernel<true><<<g, XRT_BLOCK, 0, st>>>(w.leaf_box, w.scene, w.key_a, w.val_a, F);
  else
    morton_kernel<false><<<g, XRT_BLOCK, 0, st>>>(w.leaf_box, w.scene, w.key_a, w.val_a, F);
  // stable LSD radix sort over the 30 key bits == thrust::stable_sort_by_key (BVH.cu:666)
  cub::DoubleBuffer<uint32_t> k(w.key_a, w.key_b), v(w.val_a, w.val_b);
  size_t bytes = w.cub_bytes;
  cub::DeviceRadixSort::SortPairs(w.cub_tmp, bytes, k, v, (int)F, 0, 30, st);
  b.morton = k.Current();
  b.sorted_obj = v.Current();
  cudaMemsetAsync(w.flags, 0, F * sizeof(int), st);
  karras_kernel<<<g, XRT_BLOCK, 0, st>>>(b.morton, b.sorted_obj, b.nodes, w.ranges, F);
  build_boxes_kernel<<<g, XRT_BLOCK, 0, st>>>(b.tnodes, b.nodes, w.leaf_box, b.sorted_obj, w.flags, F);
  link_kernel<<<xrt_grid(2 * (size_t)F - 1), XRT_BLOCK, 0, st>>>(b.tnodes, b.nodes, w.ranges, b.esc_int,
                                                                  b.esc_ext, F);
}

void bvh_refit_midstep(const DevBVH& b, const BvhWork& w, const float4* x1, const float4* n1,
                       const uint32_t* tri, float tol, cudaStream_t st) {
  const uint32_t F = b.n_obj;
  if (F == 0) return;
  cudaMemsetAsync(w.flags, 0, F * sizeof(int), st);
  refit_kernel<<<xrt_grid(F), XRT_BLOCK, 0, st>>>(b.tnodes, b.nodes, b.pred, x1, n1, tri, tol, w.flags, F);
}

// ------------------------------------------------------------------------------------------------
// xrt_morton_codes_noalloc: the first stage of the tree build (leaf boxes -> scene box -> Morton codes,
// BVH.cu:423-487,642) for callers that must not allocate (CUDA graph capture, serving loops).  No counterpart in the
// reference.  The stage's scratch -- the scene box the leaf boxes are
// reduced into, the leaf boxes themselves and the object indices the Morton kernel writes beside its keys -- lives in
// ONE file-scope device array that every GPU loading the library holds until exit.  Every call shares it and it keeps
// the last call's contents: the scene box is re-initialised by the first kernel of each call, leaf boxes and indices
// are fully overwritten for [0, n_faces) before they are read.  Calls that overlap on the GPU (different streams,
// different host threads) race on it.
// ------------------------------------------------------------------------------------------------
#include "../../include/xrt_b200.h"
#include <cstdarg>
#include <cstdio>
#include <mutex>

namespace {
constexpr size_t NOALLOC_HEAD = 256;  // scene box (6 ints), padded so the leaf boxes start on a 256-byte boundary
constexpr size_t NOALLOC_BYTES = NOALLOC_HEAD + (size_t)XRT_NOALLOC_MAX_FACES * (6 * sizeof(float) + sizeof(uint32_t));
static __device__ __align__(256) unsigned char xrt_noalloc_scratch[NOALLOC_BYTES];

// In host code `xrt_noalloc_scratch` is a host-side stand-in, not a device address: the address each device gave its
// copy is asked for once per device, with that device current, and only that pointer is handed to kernels.
constexpr int NOALLOC_MAX_DEVICES = 64;
unsigned char* g_noalloc_ptr[NOALLOC_MAX_DEVICES] = {};
std::mutex g_noalloc_mu;

cudaError_t noalloc_scratch_for(int dev, unsigned char** out) {
  std::lock_guard<std::mutex> lock(g_noalloc_mu);
  if (!g_noalloc_ptr[dev]) {
    void* p = nullptr;
    const cudaError_t e = cudaGetSymbolAddress(&p, xrt_noalloc_scratch);  // current device == dev (checked by the caller)
    if (e != cudaSuccess) return e;
    g_noalloc_ptr[dev] = static_cast<unsigned char*>(p);
  }
  *out = g_noalloc_ptr[dev];
  return cudaSuccess;
}

// leaf boxes of ccd = false without the primitive copy of leaf_box_kernel (same min / max, so the same boxes)
__global__ void noalloc_leaf_box_kernel(const float4* __restrict__ x, const uint32_t* __restrict__ tri,
                                        float* __restrict__ leaf_box, int* __restrict__ scene, uint32_t F) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  float lo[3] = {INFINITY, INFINITY, INFINITY}, hi[3] = {-INFINITY, -INFINITY, -INFINITY};
  if (i < F) {
    const float4 a = x[tri[3 * i]], b = x[tri[3 * i + 1]], c = x[tri[3 * i + 2]];
    lo[0] = fminf(a.x, fminf(b.x, c.x)); hi[0] = fmaxf(a.x, fmaxf(b.x, c.x));
    lo[1] = fminf(a.y, fminf(b.y, c.y)); hi[1] = fmaxf(a.y, fmaxf(b.y, c.y));
    lo[2] = fminf(a.z, fminf(b.z, c.z)); hi[2] = fmaxf(a.z, fmaxf(b.z, c.z));
    float* o = leaf_box + 6 * (size_t)i;
    o[0] = lo[0]; o[1] = lo[1]; o[2] = lo[2];
    o[3] = hi[0]; o[4] = hi[1]; o[5] = hi[2];
  }
#pragma unroll
  for (int k = 0; k < 3; k++) {
    float l = lo[k], h = hi[k];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      l = fminf(l, __shfl_xor_sync(0xFFFFFFFFu, l, o));
      h = fmaxf(h, __shfl_xor_sync(0xFFFFFFFFu, h, o));
    }
    if ((threadIdx.x & 31) == 0) {
      atomicMin(scene + k, f2ord(l));
      atomicMax(scene + 3 + k, f2ord(h));
    }
  }
}
__global__ void noalloc_scene_out_kernel(const int* __restrict__ scene, float* __restrict__ out) {
  if (threadIdx.x < 6) out[threadIdx.x] = ord2f(scene[threadIdx.x]);
}

int noalloc_fail(char* err, size_t cap, int code, const char* fmt, ...) {
  if (err && cap) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(err, cap, fmt, ap);
    va_end(ap);
  }
  return code;
}
bool noalloc_on_device(const void* p, int dev) {
  cudaPointerAttributes a{};
  if (cudaPointerGetAttributes(&a, p) != cudaSuccess) {
    cudaGetLastError();
    return false;
  }
  return (a.type == cudaMemoryTypeDevice && a.device == dev) || a.type == cudaMemoryTypeManaged;
}
}  // namespace

extern "C" int xrt_morton_codes_noalloc(const float* d_positions_xyzw, uint32_t n_verts, const uint32_t* d_indices,
                                        uint32_t n_faces, int morton_fast_div, void* stream, uint32_t* d_morton,
                                        float* d_scene_box, char* err, size_t err_cap) {
  if (err && err_cap) err[0] = 0;
  if (!d_positions_xyzw || !d_indices || !d_morton || !n_verts || !n_faces)
    return noalloc_fail(err, err_cap, XRT_ERR_INVALID,
                        "xrt_morton_codes_noalloc: positions, indices and the morton output are required and must not be empty");
  if (n_faces > XRT_NOALLOC_MAX_FACES)
    return noalloc_fail(err, err_cap, XRT_ERR_INVALID,
                        "xrt_morton_codes_noalloc: %u faces exceed the static scratch of XRT_NOALLOC_MAX_FACES = %u faces "
                        "(use xrt_bvh_build_standalone or a solver handle, which allocate)",
                        n_faces, (unsigned)XRT_NOALLOC_MAX_FACES);
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess)
    return noalloc_fail(err, err_cap, XRT_ERR_CUDA, "xrt_morton_codes_noalloc: no current CUDA device (%s)", cudaGetErrorString(e));
  if (dev < 0 || dev >= NOALLOC_MAX_DEVICES)
    return noalloc_fail(err, err_cap, XRT_ERR_INVALID, "xrt_morton_codes_noalloc: device ordinal %d is not supported", dev);
  if (!noalloc_on_device(d_positions_xyzw, dev) || !noalloc_on_device(d_indices, dev) || !noalloc_on_device(d_morton, dev) ||
      (d_scene_box && !noalloc_on_device(d_scene_box, dev)))
    return noalloc_fail(err, err_cap, XRT_ERR_INVALID,
                        "xrt_morton_codes_noalloc: inputs and outputs must be device memory of the current device %d", dev);
  unsigned char* base = nullptr;
  e = noalloc_scratch_for(dev, &base);
  if (e != cudaSuccess)
    return noalloc_fail(err, err_cap, XRT_ERR_CUDA, "xrt_morton_codes_noalloc: cudaGetSymbolAddress failed on device %d (%s)", dev,
                        cudaGetErrorString(e));
  int* scene = reinterpret_cast<int*>(base);
  float* leaf_box = reinterpret_cast<float*>(base + NOALLOC_HEAD);
  uint32_t* val = reinterpret_cast<uint32_t*>(base + NOALLOC_HEAD + (size_t)XRT_NOALLOC_MAX_FACES * 6 * sizeof(float));
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const float4* x = reinterpret_cast<const float4*>(d_positions_xyzw);
  const unsigned g = xrt_grid(n_faces);
  scene_init_kernel<<<1, 32, 0, st>>>(scene);  // the array keeps the last call's scene box: start from +-inf again
  noalloc_leaf_box_kernel<<<g, XRT_BLOCK, 0, st>>>(x, d_indices, leaf_box, scene, n_faces);
  if (morton_fast_div)
    morton_kernel<true><<<g, XRT_BLOCK, 0, st>>>(leaf_box, scene, d_morton, val, n_faces);
  else
    morton_kernel<false><<<g, XRT_BLOCK, 0, st>>>(leaf_box, scene, d_morton, val, n_faces);
  if (d_scene_box) noalloc_scene_out_kernel<<<1, 32, 0, st>>>(scene, d_scene_box);
  e = cudaGetLastError();
  if (e != cudaSuccess)
    return noalloc_fail(err, err_cap, XRT_ERR_CUDA, "xrt_morton_codes_noalloc: launch failed (%s)", cudaGetErrorString(e));
  return XRT_OK;
}
