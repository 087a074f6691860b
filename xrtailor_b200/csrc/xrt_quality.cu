// Quality mode for sm_100a: ClothSolverBase::SimulateQuality
// (xrtailor/src/pipeline/base/ClothSolverBase.cpp:263-362) and the collision pipeline it drives --
// cubic vertex-face / edge-edge CCD on compacted candidate lists (impact_zone/ImpactHelper.cu),
// independent impacts (impact_zone/Entry.cu:25-70), impact zones (ImpactZoneHelper.cu, ImpactZone.cu),
// the augmented-Lagrangian zone optimiser (ParallelOptimizer.cu, ParallelImpactZoneOptimizer.cu),
// imminent repulsion (repulsion/ImminentRepulsion*.cu) and PBD repulsion (repulsion/PBDRepulsion.cu).
//
// Shape of the B200 path (the reference's shape in brackets):
//  * every list (face pairs, narrow-phase candidates, impacts, quadratures, repulsion items) lives in a
//    preallocated arena with a DEVICE-side count; kernels are grid-stride over "148 x k" CTAs and read
//    the count from memory, so a whole repulsion iteration block and a whole FindImpacts enqueue
//    without a host round trip  [thrust::device_vector per call, .size() on the host].
//  * order-preserving compaction in three launches (per-CTA contiguous chunks: count, one-CTA scan of
//    the CTA sums, fill) instead of writing 15 Impact slots (1.3 KB) per pair and remove_if over them.
//  * the narrow phase runs one candidate per lane on a dense list: the 15 tests of a pair are first
//    filtered by shared nodes and ccd boxes from 12 position loads per pair, only survivors reach the
//    cubic solver  [one thread walks all 15 tests of a pair].
//  * PBD repulsion expands its pairs once per refresh into (vertex, face) / (edge, edge) items that
//    pass the static representative-triangle and shared-node filters; the 10 iterations until the
//    next refresh evaluate one item per lane  [12 tests per pair per iteration, serial in a thread].
//  * independent-impact selection and zone flood fill are single-CTA kernels with the loop on the
//    device  [host loop of thrust sorts; a <<<1,1>>> flood fill with a cudaMemcpyFromSymbol per zone].
//  * the zone optimiser is ONE launch, one CTA per zone, outer iterations and the backtracking line
//    search inside the kernel  [~15 launches + 2 sorts + a D2H per line-search step].
// Discrete decisions are IEEE single precision in the reference's evaluation order (xrt_dev_quality.cuh).
#include "xrt_handle.cuh"
#include "xrt_dev_quality.cuh"
#include "xrt_fem.cuh"

#include <cub/block/block_reduce.cuh>
#include <cub/block/block_scan.cuh>
#include <cub/device/device_radix_sort.cuh>

#include <algorithm>
#include <cstring>
#include <set>
#include <chrono>
#include <vector>

using namespace xrtq;

namespace {

constexpr int QB = 256;       // threads per CTA of the list kernels
constexpr int QG_MAX = 1024;  // CTAs a chunked compaction may use (one-CTA scan of the CTA sums)

enum Counter : int {
  N_PAIRS = 0, N_CAND, N_IMP, OVERFLOW, N_QUAD_EE, N_QUAD, N_REP, N_IND, N_ZONES, OPT_OUTER, OPT_LS, OPT_CONV, N_COUNTERS = 16
};
enum OverflowBit : uint32_t { OV_PAIRS = 1u, OV_CAND = 2u, OV_IMP = 4u, OV_QUAD = 8u, OV_REP = 16u };

}  // namespace

// host-side impact zone (impact_zone/ImpactZone.cuh): the impacts of one connected patch
struct HostZone {
  std::vector<DImpact> imp;
  std::vector<int> node_idx;  // 4 per impact: Node::index, LOCAL to its mesh (FlattenNode_Kernel)
  bool active = true, sorted = false, enabled = true;
  bool prev_active = true;  // grew or was new in the previous round (its tests of that round may predate the growth)
};

struct QualityState {
  xrt_quality_params qp{};
  std::vector<uint32_t> strain_colors;
  bool strain_colors_set = false;
  // topology beyond ClothDev / ObstacleDev
  uint32_t Eo = 0;
  uint32_t *cfe = nullptr, *ofe = nullptr, *cef_off = nullptr, *cef_idx = nullptr, *oef_off = nullptr, *oef_idx = nullptr;
  uint32_t* crtri = nullptr;
  uint2* oedge = nullptr;
  float4 *cfn0 = nullptr, *ofn0 = nullptr;
  float4* ov = nullptr;               // obstacle Node::v
  float4 *cx1 = nullptr, *ox1 = nullptr;  // Node::x1 snapshots of CollisionStep (the optimiser's objective)
  // trees: the frame's cloth tree (cloth_bvh_) and the two ccd trees CollisionStep builds
  DevBVH cbvh{}, ccd_c{}, ccd_o{};
  BvhWork cbw, ccd_cw, ccd_ow;
  PairWork pw;
  // arenas
  uint32_t pair_cap = 0, cand_cap = 0, imp_cap = 0, quad_cap = 0, rep_cap = 0;
  uint2* pairs = nullptr;
  uint16_t* masks = nullptr;   // per pair: which of its tests pass the static filters
  uint32_t* cand = nullptr;    // pair << 4 | slot
  float* cand_t = nullptr;     // time of impact, -1: none
  DImpact *impacts = nullptr, *imp_sorted = nullptr, *imp_ind = nullptr;
  uint32_t *sort_key_a = nullptr, *sort_key_b = nullptr, *sort_val_a = nullptr, *sort_val_b = nullptr;
  void* sort_tmp = nullptr;
  size_t sort_bytes = 0;
  uint4* quad_n = nullptr;
  float4* quad_b = nullptr;
  uint2* rep_items = nullptr;
  uint32_t *bsum = nullptr, *boff = nullptr;  // [2 * QG_MAX] CTA sums / offsets of a compaction
  uint32_t* cnt = nullptr;                    // device counters
  uint32_t* h_cnt = nullptr;                  // pinned copy
  int grid = 148;                             // CTAs of the list kernels (a multiple of the SM count)
  // independent impacts / zones
  int *removed = nullptr, *min_index = nullptr, *ncol = nullptr;  // [V + Vo]
  uint8_t* flags = nullptr;                                       // [3 * imp_cap] taken / avail / win
  uint32_t* frontier = nullptr;                                   // [2 * (V + Vo + 8)]
  int* imp_color = nullptr;                                       // [imp_cap]
  // optimiser
  size_t opt_cap_imp = 0, opt_cap_node = 0;
  DImpact* o_imp = nullptr;
  int* o_idx = nullptr;       // [4 n] variable index of each impact node or -1
  uint32_t* o_node = nullptr; // [nn] global node id
  uint32_t *o_inc_off = nullptr, *o_inc_slot = nullptr;
  uint8_t* o_skip = nullptr;
  float *o_X = nullptr, *o_Xn = nullptr, *o_G = nullptr, *o_obj = nullptr, *o_con = nullptr, *o_lam = nullptr;
  int4* o_zone = nullptr;     // node_off, n_nodes, imp_off, n_imp
  int* o_stat = nullptr;      // [3 nz] outer, line search, converged
  FemDev fem;
  xrt_collision_stats stats{};
  bool frame_begun = false;
};

namespace {

// ------------------------------------------------------------------------------------------------
// list plumbing
// ------------------------------------------------------------------------------------------------

// contiguous chunk of [0, n) this CTA owns (a multiple of the CTA size, so tiles never straddle chunks)
__device__ __forceinline__ void chunk_range(uint32_t n, uint32_t& lo, uint32_t& hi) {
  unsigned long long per = ((unsigned long long)n + gridDim.x - 1) / gridDim.x;
  per = (per + QB - 1) / QB * QB;
  const unsigned long long l = (unsigned long long)blockIdx.x * per;
  lo = l < n ? (uint32_t)l : n;
  hi = (l + per) < n ? (uint32_t)(l + per) : n;
}
__device__ __forceinline__ uint32_t block_sum(uint32_t v) {
  typedef cub::BlockReduce<uint32_t, QB> BR;
  __shared__ typename BR::TempStorage tmp;
  const uint32_t r = BR(tmp).Sum(v);
  __syncthreads();
  return r;
}
// exclusive offset of this thread's `c` items inside the tile; *tile_total = items of the tile (all threads)
__device__ __forceinline__ uint32_t tile_scan(uint32_t c, uint32_t* tile_total) {
  typedef cub::BlockScan<uint32_t, QB> BS;
  __shared__ typename BS::TempStorage tmp;
  uint32_t off, tot;
  BS(tmp).ExclusiveSum(c, off, tot);
  __syncthreads();
  *tile_total = tot;
  return off;
}

// One CTA: exclusive scan of the per-CTA sums of one or two streams (the second stream starts where the first
// ends), total -> *count_b (and the first stream's total -> *count_a), clamped to cap.
__global__ void __launch_bounds__(QG_MAX) scan_sums_kernel(const uint32_t* __restrict__ sum_a, uint32_t* __restrict__ off_a,
                                 const uint32_t* __restrict__ sum_b, uint32_t* __restrict__ off_b, int G,
                                 uint32_t* count_a, uint32_t* count_b, uint32_t cap, uint32_t* overflow, uint32_t bit) {
  typedef cub::BlockScan<unsigned long long, QG_MAX> BS;
  __shared__ typename BS::TempStorage tmp;
  __shared__ unsigned long long total_a;
  const int t = threadIdx.x;
  unsigned long long v = t < G ? sum_a[t] : 0ull, off, tot;
  BS(tmp).ExclusiveSum(v, off, tot);
  if (t < G) off_a[t] = (uint32_t)(off < cap ? off : cap);
  if (t == 0) total_a = tot;
  __syncthreads();
  unsigned long long total = tot;
  if (sum_b) {
    v = t < G ? sum_b[t] : 0ull;
    BS(tmp).ExclusiveSum(v, off, tot);
    off += total_a;
    if (t < G) off_b[t] = (uint32_t)(off < cap ? off : cap);
    total = total_a + tot;
  }
  if (t == 0) {
    if (count_a) *count_a = (uint32_t)(total_a < cap ? total_a : cap);
    *count_b = (uint32_t)(total < cap ? total : cap);
    if (total > cap) atomicOr(overflow, bit);
  }
}

// ------------------------------------------------------------------------------------------------
// CCD narrow phase on compacted lists (FindImpacts, ImpactHelper.cu:216-243)
// ------------------------------------------------------------------------------------------------

// Pass 1 over the face pairs: which of the 15 tests of a pair survive the shared-node test and the ccd boxes
// (CheckVertexFaceImpact / CheckEdgeEdgeImpact, ImpactHelper.cu:89-114).  12 position loads per pair: the six
// node boxes are built once and merged into the two face boxes and six edge boxes (min / max are exact, so the
// merged boxes are the ones Node/Edge/Face::ComputeBounds(true) return).
__global__ void __launch_bounds__(QB) ccd_filter_kernel(QMesh m, const uint2* __restrict__ pairs, const uint32_t* __restrict__ n_pairs,
                                                        float th, uint16_t* __restrict__ masks, uint32_t* __restrict__ bsum) {
  uint32_t lo, hi;
  chunk_range(*n_pairs, lo, hi);
  uint32_t local = 0;
  for (uint32_t p = lo + threadIdx.x; p < hi; p += QB) {
    const uint2 pr = pairs[p];
    uint32_t nd[6];
    QBox nb[6];
#pragma unroll
    for (int k = 0; k < 6; k++) {
      nd[k] = face_node(m, k < 3 ? pr.x : pr.y, k % 3);
      box_init(nb[k]);
      box_add(nb[k], x_of(m, nd[k]));
      box_add(nb[k], x0_of(m, nd[k]));
    }
    auto merge = [](const QBox& a, const QBox& b) {
      QBox r;
#pragma unroll
      for (int k = 0; k < 3; k++) { r.lo[k] = qmin(a.lo[k], b.lo[k]); r.hi[k] = qmax(a.hi[k], b.hi[k]); }
      return r;
    };
    const QBox fb0 = merge(merge(nb[0], nb[1]), nb[2]), fb1 = merge(merge(nb[3], nb[4]), nb[5]);
    uint32_t mask = 0;
#pragma unroll
    for (int i = 0; i < 3; i++) {
      if (nd[i] != nd[3] && nd[i] != nd[4] && nd[i] != nd[5] && box_overlap(nb[i], fb1, th)) mask |= 1u << i;
      if (nd[3 + i] != nd[0] && nd[3 + i] != nd[1] && nd[3 + i] != nd[2] && box_overlap(nb[3 + i], fb0, th)) mask |= 1u << (3 + i);
    }
    // edge j of a face joins its vertices j and j+1 (the face -> edge table is built that way at setup)
#pragma unroll
    for (int i = 0; i < 3; i++) {
      const uint32_t a0 = nd[i], a1 = nd[(i + 1) % 3];
      const QBox ea = merge(nb[i], nb[(i + 1) % 3]);
#pragma unroll
      for (int j = 0; j < 3; j++) {
        const uint32_t b0 = nd[3 + j], b1 = nd[3 + (j + 1) % 3];
        if (a0 == b0 || a0 == b1 || a1 == b0 || a1 == b1) continue;
        const QBox eb = merge(nb[3 + j], nb[3 + (j + 1) % 3]);
        if (box_overlap(ea, eb, th)) mask |= 1u << (6 + 3 * i + j);
      }
    }
    masks[p] = (uint16_t)mask;
    local += __popc(mask);
  }
  const uint32_t tot = block_sum(local);
  if (threadIdx.x == 0) bsum[blockIdx.x] = tot;
}

// Pass 3: the surviving tests as a dense list, in pair order then slot order
__global__ void __launch_bounds__(QB) ccd_expand_kernel(const uint32_t* __restrict__ n_pairs, const uint16_t* __restrict__ masks,
                                                        const uint32_t* __restrict__ boff, uint32_t* __restrict__ cand, uint32_t cap) {
  uint32_t lo, hi;
  chunk_range(*n_pairs, lo, hi);
  uint32_t running = boff[blockIdx.x];
  for (uint32_t tile = lo; tile < hi; tile += QB) {
    const uint32_t p = tile + threadIdx.x;
    uint32_t mask = p < hi ? masks[p] : 0u;
    uint32_t tot;
    uint32_t at = running + tile_scan(__popc(mask), &tot);
    while (mask) {
      const int slot = __ffs(mask) - 1;
      mask &= mask - 1;
      if (at < cap) cand[at] = (p << 4) | (uint32_t)slot;
      at++;
    }
    running += tot;
  }
}

__device__ __forceinline__ bool run_candidate(const QMesh& m, const uint2* pairs, uint32_t c, DImpact* out, float* t_only) {
  const uint2 pr = pairs[c >> 4];
  uint32_t n[4];
  const bool is_vf = slot_nodes(m, pr.x, pr.y, (int)(c & 15u), n);
  f3 p0[4], p1[4];
#pragma unroll
  for (int k = 0; k < 4; k++) { p0[k] = x0_of(m, n[k]); p1[k] = x_of(m, n[k]); }
  float t = -1.0f, w[4] = {0.f, 0.f, 0.f, 0.f};
  f3 nrm = mk3(0.f, 0.f, 0.f);
  const bool hit = check_impact(is_vf, p0, p1, t, w, nrm);
  if (t_only) *t_only = hit ? t : -1.0f;
  if (hit && out) {
#pragma unroll
    for (int k = 0; k < 4; k++) { out->n[k] = n[k]; out->w[k] = w[k]; }
    out->t = t;
    out->nrm[0] = nrm.x; out->nrm[1] = nrm.y; out->nrm[2] = nrm.z;
    out->is_vf = is_vf ? 1u : 0u;
  }
  return hit;
}

// Pass 1 over the candidates: the cubic test (CheckImpact), one candidate per lane
__global__ void __launch_bounds__(QB) ccd_test_kernel(QMesh m, const uint2* __restrict__ pairs, const uint32_t* __restrict__ cand,
                                                      const uint32_t* __restrict__ n_cand, float* __restrict__ cand_t,
                                                      uint32_t* __restrict__ bsum) {
  uint32_t lo, hi;
  chunk_range(*n_cand, lo, hi);
  uint32_t local = 0;
  for (uint32_t i = lo + threadIdx.x; i < hi; i += QB) {
    float t;
    local += run_candidate(m, pairs, cand[i], nullptr, &t) ? 1u : 0u;
    cand_t[i] = t;
  }
  const uint32_t tot = block_sum(local);
  if (threadIdx.x == 0) bsum[blockIdx.x] = tot;
}

// Pass 3: the hits as Impact records in candidate order (thrust::remove_if keeps that order, ImpactHelper.cu:238-240).
// A hit is evaluated a second time here to produce w and n: hits are a small part of the candidates, and the
// alternative is 52 bytes of scratch per candidate.
__global__ void __launch_bounds__(QB) ccd_compact_kernel(QMesh m, const uint2* __restrict__ pairs, const uint32_t* __restrict__ cand,
                                                         const uint32_t* __restrict__ n_cand, const float* __restrict__ cand_t,
                                                         const uint32_t* __restrict__ boff, DImpact* __restrict__ out, uint32_t cap) {
  uint32_t lo, hi;
  chunk_range(*n_cand, lo, hi);
  uint32_t running = boff[blockIdx.x];
  for (uint32_t tile = lo; tile < hi; tile += QB) {
    const uint32_t i = tile + threadIdx.x;
    const bool hit = i < hi && cand_t[i] >= 0.0f;
    uint32_t tot;
    const uint32_t at = running + tile_scan(hit ? 1u : 0u, &tot);
    if (hit && at < cap) run_candidate(m, pairs, cand[i], out + at, nullptr);
    running += tot;
  }
}

// ------------------------------------------------------------------------------------------------
// independent impacts (Entry.cu:25-70, ImpactHelper.cu:245-341), one CTA, loop on the device
// ------------------------------------------------------------------------------------------------
constexpr int IB = 1024;

__global__ void impact_keys_kernel(const DImpact* __restrict__ imp, uint32_t n, uint32_t* __restrict__ key, uint32_t* __restrict__ val) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  key[i] = __float_as_uint(imp[i].t + 0.0f);  // t in [0, 1]: the bit pattern orders like the float; -0 -> +0
  val[i] = i;
}
__global__ void impact_gather_kernel(const DImpact* __restrict__ imp, const uint32_t* __restrict__ order, uint32_t n,
                                     DImpact* __restrict__ out) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) out[i] = imp[order[i]];
}

__device__ __forceinline__ int block_sum_i(int v, int* sh) {
  typedef cub::BlockReduce<int, IB> BR;
  __shared__ typename BR::TempStorage tmp;
  const int r = BR(tmp).Sum(v);
  if (threadIdx.x == 0) *sh = r;
  __syncthreads();
  return *sh;
}

// imp: impacts sorted by t (stable).  A node is eligible when it may move: free cloth nodes, or every node in the
// deform pass.  Rounds of: impacts none of whose eligible nodes is taken are available; every eligible node
// learns the earliest available impact that lists it; an impact that is the earliest on all its eligible nodes
// wins and takes them.  Stops when a round selects nothing.
__global__ void __launch_bounds__(IB) independent_kernel(QMesh m, const DImpact* __restrict__ imp, int n, int deform,
                                                         int* __restrict__ removed, int* __restrict__ min_index,
                                                         uint8_t* __restrict__ taken, uint8_t* __restrict__ avail,
                                                         uint8_t* __restrict__ win, DImpact* __restrict__ out, uint32_t* n_out) {
  __shared__ int sh;
  const int tid = threadIdx.x;
  auto elig = [&](uint32_t g) { return deform == 1 || free_of(m, g); };
  for (int i = tid; i < n; i += IB) {  // InitializeImpactNode_Kernel
    taken[i] = 0;
    for (int j = 0; j < 4; j++) {
      const uint32_t g = imp[i].n[j];
      if (elig(g)) { removed[g] = 0; min_index[g] = n; }
    }
  }
  __syncthreads();
  int num, new_num = n;
  do {
    num = new_num;
    for (int i = tid; i < n; i += IB) {
      bool a = true;
      for (int j = 0; j < 4; j++) {
        const uint32_t g = imp[i].n[j];
        if (elig(g) && removed[g]) { a = false; break; }
      }
      avail[i] = a;
    }
    __syncthreads();
    for (int i = tid; i < n; i += IB)
      if (avail[i])
        for (int j = 0; j < 4; j++) {
          const uint32_t g = imp[i].n[j];
          if (elig(g)) min_index[g] = 0x7FFFFFFF;
        }
    __syncthreads();
    for (int i = tid; i < n; i += IB)
      if (avail[i])
        for (int j = 0; j < 4; j++) {
          const uint32_t g = imp[i].n[j];
          if (elig(g)) atomicMin(&min_index[g], i);
        }
    __syncthreads();
    for (int i = tid; i < n; i += IB) {  // CheckIndependentImpacts_Kernel
      bool f = true;
      for (int j = 0; j < 4; j++) {
        const uint32_t g = imp[i].n[j];
        if (elig(g) && (removed[g] || min_index[g] != i)) { f = false; break; }
      }
      win[i] = f;
    }
    __syncthreads();
    int left = 0;
    for (int i = tid; i < n; i += IB) {
      if (win[i]) {
        taken[i] = 1;
        for (int j = 0; j < 4; j++) {
          const uint32_t g = imp[i].n[j];
          if (elig(g)) removed[g] = 1;
        }
      }
      left += taken[i] ? 0 : 1;
    }
    new_num = block_sum_i(left, &sh);
    __syncthreads();
  } while (num > new_num);
  // the selected impacts, in sorted order
  typedef cub::BlockScan<int, IB> BS;
  __shared__ typename BS::TempStorage stmp;
  int running = 0;
  for (int tile = 0; tile < n; tile += IB) {
    const int i = tile + tid;
    const int c = (i < n && taken[i]) ? 1 : 0;
    int off, tot;
    BS(stmp).ExclusiveSum(c, off, tot);
    __syncthreads();
    if (c) out[running + off] = imp[i];
    running += tot;
  }
  if (tid == 0) *n_out = (uint32_t)running;
}

// ------------------------------------------------------------------------------------------------
// zone isolation (IsolateImpacts, ImpactZoneHelper.cu:227-334), one CTA, frontier-parallel flood fill
// ------------------------------------------------------------------------------------------------
#define Q_NODE_OUTSIDE -2
#define Q_NODE_INSIDE -1

// A zone grows from the first impact that has no zone yet: its four nodes are expanded through the mesh (a node's
// face neighbours), stepping only onto nodes that belong to some impact and have no zone yet; pinned cloth nodes
// take a zone but are not expanded.  Impacts then take the zone of their first zoned node.  At most 100 zones.
// The reference walks each zone with one thread and a 1024-entry stack (FloodFillColoring_Kernel :38-95); the
// zones are the same because a fill is a reachability set, whatever the visiting order.
__global__ void __launch_bounds__(IB) isolate_kernel(QMesh m, const uint32_t* __restrict__ cvf_off, const uint32_t* __restrict__ cvf_idx,
                                                     const uint32_t* __restrict__ ovf_off, const uint32_t* __restrict__ ovf_idx,
                                                     const DImpact* __restrict__ imp, int n, int* __restrict__ ncol,
                                                     int* __restrict__ colors, uint32_t* __restrict__ frontier, uint32_t fcap,
                                                     uint32_t* n_zones) {
  __shared__ int s_pos;
  __shared__ uint32_t s_n[2];
  const int tid = threadIdx.x;
  for (int i = tid; i < n; i += IB) {
    colors[i] = -1;
    for (int j = 0; j < 4; j++) ncol[imp[i].n[j]] = Q_NODE_INSIDE;
  }
  __syncthreads();
  int n_colors = 0;
  for (;;) {
    if (n_colors > 99) break;  // "too many attempts, exit"
    if (tid == 0) s_pos = 0x7FFFFFFF;
    __syncthreads();
    int best = 0x7FFFFFFF;
    for (int i = tid; i < n; i += IB)
      if (colors[i] == -1) { best = i; break; }
    if (best != 0x7FFFFFFF) atomicMin(&s_pos, best);
    __syncthreads();
    const int pos = s_pos;
    if (pos == 0x7FFFFFFF) break;
    uint32_t* cur = frontier;
    uint32_t* nxt = frontier + fcap;
    if (tid < 4) cur[tid] = imp[pos].n[tid];
    if (tid == 0) { s_n[0] = 4; s_n[1] = 0; }
    __syncthreads();
    int which = 0;
    while (s_n[which] > 0) {
      const uint32_t nf = s_n[which];
      for (uint32_t k = tid; k < nf; k += IB) {
        const uint32_t node = cur[k];
        const bool cl = is_cloth(m, node);
        if (cl && !free_of(m, node)) continue;  // pinned cloth nodes are not expanded
        const uint32_t l = cl ? node : node - m.V;
        const uint32_t* off = cl ? cvf_off : ovf_off;
        const uint32_t* idx = cl ? cvf_idx : ovf_idx;
        for (uint32_t a = off[l]; a < off[l + 1]; a++) {
          const uint32_t gf = (cl ? 0u : m.F) + (idx[a] >> 2);
          for (int j = 0; j < 3; j++) {
            const uint32_t g = face_node(m, gf, j);
            if (atomicCAS(&ncol[g], Q_NODE_INSIDE, n_colors) == Q_NODE_INSIDE) {
              const uint32_t at = atomicAdd(&s_n[which ^ 1], 1u);
              if (at < fcap) nxt[at] = g;
            }
          }
        }
      }
      __syncthreads();
      if (tid == 0) s_n[which] = 0;
      which ^= 1;
      uint32_t* t = cur; cur = nxt; nxt = t;
      __syncthreads();
    }
    for (int i = tid; i < n; i += IB) {
      if (colors[i] != -1) continue;
      for (int j = 0; j < 4; j++) {
        const int c = ncol[imp[i].n[j]];
        if (c != Q_NODE_OUTSIDE && c != Q_NODE_INSIDE) { colors[i] = c; break; }
      }
    }
    n_colors++;
    __syncthreads();
  }
  __syncthreads();
  for (int i = tid; i < n; i += IB)  // leave the node table clean for the next call
    for (int j = 0; j < 4; j++) ncol[imp[i].n[j]] = Q_NODE_OUTSIDE;
  if (tid == 0) *n_zones = (uint32_t)n_colors;
}

// ------------------------------------------------------------------------------------------------
// the zone optimiser (ParallelOptimizer::Solve, ParallelOptimizer.cu:513-890; objective, constraint and
// gradients ParallelImpactZoneOptimizer.cu:14-160,383-468): one CTA per zone, everything in the kernel
// ------------------------------------------------------------------------------------------------
#define IZO_EPS_S 1e-12f
#define IZO_EPS_F 1e-10f
#define IZO_ALPHA 0.5f

struct OptArgs {
  QMesh m;
  const float4 *cx1, *ox1;  // Node::x1
  const DImpact* imp;
  const int* idx;           // [4 n]
  const uint32_t* node;     // [nn]
  const uint8_t* skip;      // [nn] 1: a later variable holds the same node and writes it back
  const uint32_t *inc_off, *inc_slot;
  float *X, *Xn, *G, *obj, *con, *lam;
  const int4* zone;
  int* stat;
  float4 *out_cx, *out_ox;  // Node::x of both meshes (Finalize_Kernel writes the result back)
  float thickness, obstacle_mass, inv_mass, mu;
  int deform;
};

__device__ __forceinline__ f3 ld3f(const float* a, uint32_t i) { return mk3(a[3 * i], a[3 * i + 1], a[3 * i + 2]); }
__device__ __forceinline__ void st3f(float* a, uint32_t i, f3 v) { a[3 * i] = v.x; a[3 * i + 1] = v.y; a[3 * i + 2] = v.z; }

// ReduceBaselineLocal_Kernel + ReduceBaselineGlobal_Kernel (ParallelOptimizer.cu:330-420): blocks of 256 summed by
// a stride-doubling tree, then the (at most 256) block sums by the same tree -- the same association order, so
// the sums (and with them the line-search decisions) are the reference's.
// The tree over 256 values is evaluated with warp shuffles for strides 1..16 (value t takes value t + stride, exactly
// the additions s[t] += s[t + stride] of the reference's loop survive in lane t % (2 stride) == 0) and by warp 0 over
// the eight warp sums for strides 32..128: the same binary tree, two block barriers instead of ten.
__device__ __forceinline__ float block_tree_sum(float v, float* sbuf) {
  const int t = threadIdx.x;
#pragma unroll
  for (int stride = 1; stride < 32; stride *= 2) v += __shfl_down_sync(0xFFFFFFFFu, v, stride);
  if ((t & 31) == 0) sbuf[t >> 5] = v;
  __syncthreads();
  if (t < 32) {
    float w = t < QB / 32 ? sbuf[t] : 0.0f;
#pragma unroll
    for (int stride = 1; stride < QB / 32; stride *= 2) w += __shfl_down_sync(0xFFFFFFFFu, w, stride);
    if (t == 0) sbuf[QB / 32] = w;
  }
  __syncthreads();
  const float r = sbuf[QB / 32];
  __syncthreads();  // sbuf is free again
  return r;
}
template <class Get>
__device__ __forceinline__ float zone_reduce(float* sbuf, float* part, uint32_t n, Get get) {
  const int t = threadIdx.x;
  uint32_t nb = (n + QB - 1) / QB;
  if (nb > QB) nb = QB;
  if (nb <= 1) {  // one block: its sum, then the tree over (sum, 0, 0, ...) which adds exact zeros
    const float r = block_tree_sum((uint32_t)t < n ? get((uint32_t)t) : 0.0f, sbuf);
    return r;
  }
  for (uint32_t b = 0; b < nb; b++) {
    const uint32_t i = b * QB + t;
    const float r = block_tree_sum(i < n ? get(i) : 0.0f, sbuf);
    if (t == 0) part[b] = r;
  }
  __syncthreads();
  return block_tree_sum((uint32_t)t < nb ? part[t] : 0.0f, sbuf);
}

// SM: the zone's working set -- variables, next variables, gradient, objective terms, constraint values, multipliers and
// the per-variable / per-impact constants they are computed from -- lives in shared memory (the host picks this variant
// when the largest zone of the launch fits, zone_smem_bytes).  The optimiser is ~5 000 tiny dependent phases per zone
// (load a few values, a handful of flops, store, barrier), so its time is the latency of those loads and stores; the
// arithmetic and its order are the same in both variants.
struct ZoneImp {  // what the phases read of an impact, 48 bytes
  float nrm[3];
  float w[4];
  int idx[4];
  uint32_t pad;
};
__host__ __device__ inline size_t zone_smem_bytes(size_t nn, size_t ni) {
  return (14 * nn + 2 * ni) * sizeof(float) + ni * sizeof(ZoneImp);
}

template <bool SM>
__global__ void __launch_bounds__(QB) zone_optimize_kernel(OptArgs a) {
  __shared__ float sbuf[QB], part[QB];
  extern __shared__ __align__(16) unsigned char zone_dyn[];
  const int4 z = a.zone[blockIdx.x];
  const uint32_t no = (uint32_t)z.x, nn = (uint32_t)z.y, io = (uint32_t)z.z, ni = (uint32_t)z.w;
  const int t = threadIdx.x;
  const QMesh& m = a.m;
  const float mu = a.mu, inv_mass = a.inv_mass;
  // storage, indexed like the global arrays (k in [no, no + nn), i in [io, io + ni))
  float* dyn = reinterpret_cast<float*>(zone_dyn);
  float* X = SM ? dyn - 3 * (size_t)no : a.X;
  float* Xn = SM ? dyn + 3 * (size_t)nn - 3 * (size_t)no : a.Xn;
  float* G = SM ? dyn + 6 * (size_t)nn - 3 * (size_t)no : a.G;
  float* X1 = dyn + 9 * (size_t)nn - 3 * (size_t)no;      // SM only: Node::x1 of the variable
  float* obj = SM ? dyn + 12 * (size_t)nn - (size_t)no : a.obj;
  float* M = dyn + 13 * (size_t)nn - (size_t)no;           // SM only: mass of the variable
  float* con = SM ? dyn + 14 * (size_t)nn - (size_t)io : a.con;
  float* lam = SM ? dyn + 14 * (size_t)nn + (size_t)ni - (size_t)io : a.lam;
  ZoneImp* zi = reinterpret_cast<ZoneImp*>(dyn + 14 * (size_t)nn + 2 * (size_t)ni) - (size_t)io;  // SM only
  auto mass_g = [&](uint32_t k) { const uint32_t g = a.node[k]; return is_cloth(m, g) ? (1.0f / invm_of(m, g)) : a.obstacle_mass; };
  auto x1_g = [&](uint32_t k) { const uint32_t g = a.node[k]; return mk3(is_cloth(m, g) ? a.cx1[g] : a.ox1[g - m.V]); };
  auto mass = [&](uint32_t k) { return SM ? M[k] : mass_g(k); };
  auto x1 = [&](uint32_t k) { return SM ? ld3f(X1, k) : x1_g(k); };
  auto imp_nrm = [&](uint32_t i) { return SM ? mk3(zi[i].nrm[0], zi[i].nrm[1], zi[i].nrm[2]) : mk3(a.imp[i].nrm[0], a.imp[i].nrm[1], a.imp[i].nrm[2]); };
  auto imp_w = [&](uint32_t i, int j) { return SM ? zi[i].w[j] : a.imp[i].w[j]; };
  auto imp_idx = [&](uint32_t i, int j) { return SM ? zi[i].idx[j] : a.idx[4 * i + j]; };
  // CollisionConstraint_Kernel (:77-93): c = -thickness + sum_j w_j n . x_j, variables from XX, the rest from the mesh
  auto constraint = [&](const float* XX) {
    for (uint32_t i = io + t; i < io + ni; i += QB) {
      const f3 nrm = imp_nrm(i);
      float c = -a.thickness;
      for (int j = 0; j < 4; j++) {
        const int k = imp_idx(i, j);
        c += imp_w(i, j) * dot3(nrm, k > -1 ? ld3f(XX, (uint32_t)k) : x_of(m, a.imp[i].n[j]));
      }
      con[i] = c;
    }
    __syncthreads();
  };
  auto coeff = [&]() {  // ClampViolation
    for (uint32_t i = io + t; i < io + ni; i += QB) con[i] = qmin(con[i] + lam[i] / mu, 0.0f);
    __syncthreads();
  };
  auto objective = [&](const float* XX) {  // CollisionObjective_Kernel (:57-65)
    for (uint32_t k = no + t; k < no + nn; k += QB) {
      const f3 d = ld3f(XX, k) - x1(k);
      obj[k] = mass(k) * dot3(d, d);
    }
    __syncthreads();
  };
  for (uint32_t k = no + t; k < no + nn; k += QB) {  // Initialize_Kernel; next_X starts at zero
    st3f(X, k, x_of(m, a.node[k]));
    st3f(Xn, k, mk3(0.f, 0.f, 0.f));
    if (SM) {
      st3f(X1, k, x1_g(k));
      M[k] = mass_g(k);
    }
  }
  for (uint32_t i = io + t; i < io + ni; i += QB) {
    lam[i] = 0.0f;
    if (SM) {
      const DImpact& im = a.imp[i];
      ZoneImp q;
      for (int j = 0; j < 3; j++) q.nrm[j] = im.nrm[j];
      for (int j = 0; j < 4; j++) { q.w[j] = im.w[j]; q.idx[j] = a.idx[4 * i + j]; }
      q.pad = 0;
      zi[i] = q;
    }
  }
  __syncthreads();
  float s = 1e-3f * (float)nn;  // InitStepLength_Kernel
  bool converged = false;
  int iter, ls_total = 0;
  for (iter = 0; iter < 100; iter++) {
    constraint(X);
    // TotalGradient (:213-224)
    for (uint32_t k = no + t; k < no + nn; k += QB) st3f(G, k, (ld3f(X, k) - x1(k)) * (inv_mass * mass(k)));
    __syncthreads();
    coeff();
    for (uint32_t k = no + t; k < no + nn; k += QB) {  // a node gathers its impacts in slot order: no atomics
      f3 g = ld3f(G, k);
      for (uint32_t q = a.inc_off[k]; q < a.inc_off[k + 1]; q++) {
        const uint32_t slot = a.inc_slot[q], i = slot >> 2, j = slot & 3u;
        g = g + imp_nrm(i) * (mu * con[i] * imp_w(i, (int)j));
      }
      st3f(G, k, g);
    }
    __syncthreads();
    const float G2 = zone_reduce(sbuf, part, nn, [&](uint32_t i) { const f3 g = ld3f(G, no + i); return dot3(g, g); });
    objective(X);
    const float C2 = zone_reduce(sbuf, part, ni, [&](uint32_t i) { return con[io + i] * con[io + i]; });
    const float L2 = zone_reduce(sbuf, part, ni, [&](uint32_t i) { return lam[io + i] * lam[io + i]; });
    const float O = zone_reduce(sbuf, part, nn, [&](uint32_t i) { return obj[no + i]; });
    s /= 0.7f;  // PreconditionStepLength_Kernel
    bool wolfe = false;
    do {
      s *= 0.7f;  // AdjustStepLength_Kernel
      for (uint32_t k = no + t; k < no + nn; k += QB) st3f(Xn, k, ld3f(X, k) - ld3f(G, k) * s);  // NextX_Kernel
      __syncthreads();
      objective(Xn);
      constraint(Xn);
      coeff();
      const float C2t = zone_reduce(sbuf, part, ni, [&](uint32_t i) { return con[io + i] * con[io + i]; });
      const float Ot = zone_reduce(sbuf, part, nn, [&](uint32_t i) { return obj[no + i]; });
      // BacktrackingLineSearchAdvanced_Kernel (ParallelImpactZoneOptimizer.cu:383-468)
      const float f = 0.5f * inv_mass * O + 0.5f * mu * C2 - 0.5f * L2 / mu;
      const float ft = 0.5f * inv_mass * Ot + 0.5f * mu * C2t - 0.5f * L2 / mu;
      if (s < IZO_EPS_S || ft < (f - IZO_ALPHA * s * G2) || qabs(f - ft) < IZO_EPS_F) wolfe = true;
      ls_total++;
    } while (!wolfe);
    if (s < IZO_EPS_S) converged = true;  // UpdateConvergency_Kernel
    if (converged) break;
    for (uint32_t k = no + t; k < no + nn; k += QB) st3f(X, k, ld3f(Xn, k));
    __syncthreads();
    constraint(X);  // UpdateMultiplier (:226-235)
    for (uint32_t i = io + t; i < io + ni; i += QB) lam[i] = qmin(lam[i] + mu * con[i], 0.0f);
    __syncthreads();
  }
  // Finalize_Kernel.  A node two zones list (an impact takes the zone of its FIRST zoned node, its other nodes may
  // sit in another zone) is a variable of both; the reference lets the zones race for it, here the zone with the
  // higher index writes it (what a serial pass over the variables would leave behind).
  for (uint32_t k = no + t; k < no + nn; k += QB) {
    if (a.skip[k]) continue;
    const uint32_t g = a.node[k];
    const f3 p = ld3f(X, k);
    if (is_cloth(m, g)) { float4 q = a.out_cx[g]; q.x = p.x; q.y = p.y; q.z = p.z; a.out_cx[g] = q; }
    else { float4 q = a.out_ox[g - m.V]; q.x = p.x; q.y = p.y; q.z = p.z; a.out_ox[g - m.V] = q; }
  }
  if (t == 0) { a.stat[3 * blockIdx.x] = iter; a.stat[3 * blockIdx.x + 1] = ls_total; a.stat[3 * blockIdx.x + 2] = converged ? 1 : 0; }
}

// ------------------------------------------------------------------------------------------------
// imminent repulsion (ImminentRepulsion.cu:21-169, ImminentRepulsionHelper.cu:10-208)
// ------------------------------------------------------------------------------------------------

// which of the 9 edge-edge (bits 0-8, slot 3 i + j) and 3 vertex-face (bits 9-11) quadratures of a pair exist
__device__ __forceinline__ uint32_t quad_mask(const QMesh& m, uint2 pr) {
  if (pr.x == pr.y) return 0u;
  const uint32_t r0 = face_rtri(m, pr.x), r1 = face_rtri(m, pr.y);
  uint32_t mask = 0, n[4];
  float b[4];
  for (int i = 0; i < 3; i++) {
    if (!rtri_edge(r0, i)) continue;
    for (int j = 0; j < 3; j++) {
      if (!rtri_edge(r1, j)) continue;
      if (ee_quadrature(m, face_edge(m, pr.x, i), face_edge(m, pr.y, j), n, b)) mask |= 1u << (3 * i + j);
    }
  }
  for (int i = 0; i < 3; i++) {
    if (!rtri_vertex(r0, i)) continue;
    if (vf_quadrature(m, face_node(m, pr.x, i), pr.y, n, b)) mask |= 1u << (9 + i);
  }
  return mask;
}
__global__ void __launch_bounds__(QB) quad_count_kernel(QMesh m, const uint2* __restrict__ pairs, const uint32_t* __restrict__ n_pairs,
                                                        uint16_t* __restrict__ masks, uint32_t* __restrict__ bsum_ee,
                                                        uint32_t* __restrict__ bsum_vf) {
  uint32_t lo, hi;
  chunk_range(*n_pairs, lo, hi);
  uint32_t ee = 0, vf = 0;
  for (uint32_t p = lo + threadIdx.x; p < hi; p += QB) {
    const uint32_t mask = quad_mask(m, pairs[p]);
    masks[p] = (uint16_t)mask;
    ee += __popc(mask & 0x1FFu);
    vf += __popc(mask >> 9);
  }
  const uint32_t te = block_sum(ee), tv = block_sum(vf);
  if (threadIdx.x == 0) { bsum_ee[blockIdx.x] = te; bsum_vf[blockIdx.x] = tv; }
}
// quadratures in the reference's order: every edge-edge one (pair-major, slot order), then every vertex-face one
__global__ void __launch_bounds__(QB) quad_fill_kernel(QMesh m, const uint2* __restrict__ pairs, const uint32_t* __restrict__ n_pairs,
                                                       const uint16_t* __restrict__ masks, const uint32_t* __restrict__ boff_ee,
                                                       const uint32_t* __restrict__ boff_vf, uint4* __restrict__ quad_n,
                                                       float4* __restrict__ quad_b, uint32_t cap) {
  uint32_t lo, hi;
  chunk_range(*n_pairs, lo, hi);
  uint32_t run_ee = boff_ee[blockIdx.x], run_vf = boff_vf[blockIdx.x];
  for (uint32_t tile = lo; tile < hi; tile += QB) {
    const uint32_t p = tile + threadIdx.x;
    const uint32_t mask = p < hi ? masks[p] : 0u;
    uint32_t tot_ee, tot_vf;
    uint32_t at_ee = run_ee + tile_scan(__popc(mask & 0x1FFu), &tot_ee);
    uint32_t at_vf = run_vf + tile_scan(__popc(mask >> 9), &tot_vf);
    if (mask) {
      const uint2 pr = pairs[p];
      uint32_t n[4];
      float b[4];
      for (int sl = 0; sl < 12; sl++) {
        if (!((mask >> sl) & 1u)) continue;
        uint32_t at;
        if (sl < 9) { ee_quadrature(m, face_edge(m, pr.x, sl / 3), face_edge(m, pr.y, sl % 3), n, b); at = at_ee++; }
        else { vf_quadrature(m, face_node(m, pr.x, sl - 9), pr.y, n, b); at = at_vf++; }
        if (at < cap) { quad_n[at] = make_uint4(n[0], n[1], n[2], n[3]); quad_b[at] = make_float4(b[0], b[1], b[2], b[3]); }
      }
    }
    run_ee += tot_ee;
    run_vf += tot_vf;
  }
}

// AccumulateImpulses / AccumulateRepulsions (ImminentRepulsionHelper.cu:192-206, PBDRepulsion.cu:283-297):
// Jacobi sums on free (cloth) nodes; delta.xyz = sum, delta.w = count
__device__ __forceinline__ void accumulate(const QMesh& m, float4* delta, const uint32_t nodes[4], const f3 corr[4]) {
#pragma unroll
  for (int i = 0; i < 4; i++) {
    if (len3(corr[i]) < XRT_EPS) continue;
    if (!free_of(m, nodes[i])) continue;
    float* d = reinterpret_cast<float*>(delta + nodes[i]);
    atomicAdd(d, corr[i].x); atomicAdd(d + 1, corr[i].y); atomicAdd(d + 2, corr[i].z); atomicAdd(d + 3, 1.0f);
  }
}

// UpdateVelocity_Kernel + ComputeImminentImpulse_Kernel (ImminentRepulsion.cu:104-119, EvaluateImminentImpulse
// ImminentRepulsionHelper.cu:82-150; add_repulsion = true, friction_coeff = 0).  The cloth velocity the reference
// first stores into Node::v, (x - x0) / dt, is formed where it is used.
__global__ void __launch_bounds__(QB) imminent_impulse_kernel(QMesh m, const uint4* __restrict__ quad_n, const float4* __restrict__ quad_b,
                                                              const uint32_t* __restrict__ n_quads, float thickness, float dt,
                                                              float4* __restrict__ delta) {
  const uint32_t n = *n_quads;
  for (uint32_t c = blockIdx.x * QB + threadIdx.x; c < n; c += gridDim.x * QB) {
    const uint4 qn = quad_n[c];
    const float4 qb = quad_b[c];
    const uint32_t nd[4] = {qn.x, qn.y, qn.z, qn.w};
    const float bary[4] = {qb.x, qb.y, qb.z, qb.w};
    f3 pr = mk3(0.f, 0.f, 0.f), vr = mk3(0.f, 0.f, 0.f);
    float invm[4];
#pragma unroll
    for (int i = 0; i < 4; i++) {
      const f3 x0 = x0_of(m, nd[i]);
      const f3 v = is_cloth(m, nd[i]) ? (x_of(m, nd[i]) - x0) / dt : mk3(m.ov[nd[i] - m.V]);
      pr = pr + x0 * bary[i];
      vr = vr + v * bary[i];
      invm[i] = invm_of(m, nd[i]);
    }
    const float npr = len3(pr);
    if (npr > thickness) continue;
    if (npr < 1e-5f) continue;
    pr = normalize3(pr);
    const float vr_nrm = dot3(vr, pr);
    const float target = (thickness - npr) * 0.1f;
    if (vr_nrm > target) continue;
    f3 I = pr * (vr_nrm * dt);
    const f3 vr_tan = vr - pr * vr_nrm;
    if (len3(vr_tan) > thickness * 0.01f) {
      const float imp_norm = len3(I);
      const f3 dir = normalize3(vr_tan);
      const float friction_coeff = 0.0f, friction_scale = 1.0f;
      f3 friction;
      if (imp_norm * friction_coeff > len3(vr_tan)) friction = -vr_tan;
      else friction = ((-dir) * imp_norm) * friction_coeff;
      I = I + friction * friction_scale;
    }
    float inv_cm = 0.0f;
#pragma unroll
    for (int i = 0; i < 4; i++) inv_cm += bary[i] * bary[i] * invm[i];
    if (inv_cm < 1e-5f) continue;
    f3 corr[4];
#pragma unroll
    for (int i = 0; i < 4; i++) corr[i] = I * (-invm[i] * bary[i] / inv_cm);
    accumulate(m, delta, nd, corr);
  }
}
// ApplyImminentImpulse_Kernel / ApplyPBDImpulse_Kernel (ImminentRepulsion.cu:121-136, PBDRepulsion.cu:299-314)
__global__ void jacobi_apply_kernel(float4* __restrict__ x, float4* __restrict__ delta, uint32_t V, float rate) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= V) return;
  const float4 d = delta[i];
  if (d.w <= XRT_EPS) return;
  float4 p = x[i];
  p.x += d.x / d.w * rate; p.y += d.y / d.w * rate; p.z += d.z / d.w * rate;
  x[i] = p;
  delta[i] = make_float4(0.f, 0.f, 0.f, 0.f);
}

// ------------------------------------------------------------------------------------------------
// PBD repulsion (PBDRepulsion.cu:22-33,283-516)
// ------------------------------------------------------------------------------------------------

// which of the 3 vertex-face (bits 0-2) and 9 edge-edge (bits 3-11) tests of a pair can ever fire: ownership by
// the representative triangles and no shared node.  These do not depend on the positions, so they are decided
// once per pair refresh, not once per iteration.
__device__ __forceinline__ uint32_t rep_mask(const QMesh& m, uint2 pr) {
  const uint32_t r0 = face_rtri(m, pr.x), r1 = face_rtri(m, pr.y);
  uint32_t mask = 0;
  const uint32_t g0 = face_node(m, pr.y, 0), g1 = face_node(m, pr.y, 1), g2 = face_node(m, pr.y, 2);
  for (int i = 0; i < 3; i++) {
    if (!rtri_vertex(r0, i)) continue;
    const uint32_t nd = face_node(m, pr.x, i);
    if (nd == g0 || nd == g1 || nd == g2) continue;
    mask |= 1u << i;
  }
  for (int i = 0; i < 3; i++) {
    if (!rtri_edge(r0, i)) continue;
    const uint2 a = edge_nodes(m, face_edge(m, pr.x, i));
    for (int j = 0; j < 3; j++) {
      if (!rtri_edge(r1, j)) continue;
      const uint2 b = edge_nodes(m, face_edge(m, pr.y, j));
      if (a.x == b.x || a.x == b.y || a.y == b.x || a.y == b.y) continue;
      mask |= 1u << (3 + 3 * i + j);
    }
  }
  return mask;
}
__global__ void __launch_bounds__(QB) rep_count_kernel(QMesh m, const uint2* __restrict__ pairs, const uint32_t* __restrict__ n_pairs,
                                                       uint16_t* __restrict__ masks, uint32_t* __restrict__ bsum) {
  uint32_t lo, hi;
  chunk_range(*n_pairs, lo, hi);
  uint32_t local = 0;
  for (uint32_t p = lo + threadIdx.x; p < hi; p += QB) {
    const uint32_t mask = rep_mask(m, pairs[p]);
    masks[p] = (uint16_t)mask;
    local += __popc(mask);
  }
  const uint32_t tot = block_sum(local);
  if (threadIdx.x == 0) bsum[blockIdx.x] = tot;
}
// items: VF (node | 1 << 31, face), EE (edge, edge); pair order, VF before EE inside a pair (the order
// EvaluateRep_Kernel visits them, PBDRepulsion.cu:470-500)
__global__ void __launch_bounds__(QB) rep_fill_kernel(QMesh m, const uint2* __restrict__ pairs, const uint32_t* __restrict__ n_pairs,
                                                      const uint16_t* __restrict__ masks, const uint32_t* __restrict__ boff,
                                                      uint2* __restrict__ items, uint32_t cap) {
  uint32_t lo, hi;
  chunk_range(*n_pairs, lo, hi);
  uint32_t running = boff[blockIdx.x];
  for (uint32_t tile = lo; tile < hi; tile += QB) {
    const uint32_t p = tile + threadIdx.x;
    uint32_t mask = p < hi ? masks[p] : 0u;
    uint32_t tot;
    uint32_t at = running + tile_scan(__popc(mask), &tot);
    if (mask) {
      const uint2 pr = pairs[p];
      while (mask) {
        const int sl = __ffs(mask) - 1;
        mask &= mask - 1;
        uint2 it;
        if (sl < 3) it = make_uint2(face_node(m, pr.x, sl) | 0x80000000u, pr.y);
        else it = make_uint2(face_edge(m, pr.x, (sl - 3) / 3), face_edge(m, pr.y, (sl - 3) % 3));
        if (at < cap) items[at] = it;
        at++;
      }
    }
    running += tot;
  }
}
// EvaluateRep_Kernel (PBDRepulsion.cu:470-500), one item per lane
__global__ void __launch_bounds__(QB) rep_solve_kernel(QMesh m, const uint2* __restrict__ items, const uint32_t* __restrict__ n_items,
                                                       float h, float4* __restrict__ delta) {
  const uint32_t n = *n_items;
  for (uint32_t c = blockIdx.x * QB + threadIdx.x; c < n; c += gridDim.x * QB) {
    const uint2 it = items[c];
    uint32_t nodes[4];
    f3 corr[4];
    const bool hit = (it.x & 0x80000000u) ? rep_vf(m, it.x & 0x7FFFFFFFu, it.y, h, nodes, corr) : rep_ee(m, it.x, it.y, h, nodes, corr);
    if (hit) accumulate(m, delta, nodes, corr);
  }
}

// SetupNode_Kernel / SetupObstacleNode_Kernel (Entry.cu:110-124)
__global__ void snapshot_kernel(const float4* __restrict__ x, const float4* __restrict__ x0, float4* __restrict__ x1,
                                float4* __restrict__ v, float dt, uint32_t n) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const float4 p = x[i];
  x1[i] = p;
  if (v) {
    const float4 q = x0[i];
    v[i] = make_float4((p.x - q.x) / dt, (p.y - q.y) / dt, (p.z - q.z) / dt, 0.f);
  }
}
__global__ void face_normal_kernel(const float4* __restrict__ x, const uint32_t* __restrict__ tri, float4* __restrict__ fn, uint32_t F) {
  const uint32_t f = blockIdx.x * blockDim.x + threadIdx.x;
  if (f >= F) return;  // Face::Update, Face.cu:56-60
  const f3 a = mk3(x[tri[3 * f]]), b = mk3(x[tri[3 * f + 1]]), c = mk3(x[tri[3 * f + 2]]);
  const f3 n = normalize3(cross3(b - a, c - a));
  fn[f] = make_float4(n.x, n.y, n.z, 0.f);
}

// ------------------------------------------------------------------------------------------------
// host side
// ------------------------------------------------------------------------------------------------

QMesh make_mesh(const xrt_solver* s) {
  const QualityState* q = s->q;
  QMesh m{};
  m.V = s->c.V; m.Vo = s->o.V; m.F = s->c.F; m.Fo = s->o.F; m.E = s->c.E; m.Eo = q->Eo;
  m.cx0 = s->c.x0; m.cx = s->c.x; m.ox0 = s->o.x0; m.ox = s->o.x; m.ov = q->ov;
  m.ctri = s->c.tri; m.otri = s->o.tri; m.cfe = q->cfe; m.ofe = q->ofe;
  m.cedge = s->c.edge; m.oedge = q->oedge; m.crtri = q->crtri; m.ortri = s->o.rtri;
  m.cef_off = q->cef_off; m.cef_idx = q->cef_idx; m.oef_off = q->oef_off; m.oef_idx = q->oef_idx;
  m.cfn0 = q->cfn0; m.ofn0 = q->ofn0;
  m.obstacle_inv_mass = 1.0f / 1e3f;
  return m;
}

// face -> edge ids (edge j joins vertices j and j+1; Face::SetEdges, PhysicsMeshHelper.cu:47-50) and edge -> faces
// ascending (BuildEFAdjacency, ClothSolverHelper.cpp:111-125), O(F) through a vertex -> incident-edges table
bool build_edge_topology(const std::vector<uint32_t>& tri, const std::vector<uint32_t>& edge, uint32_t nv,
                         std::vector<uint32_t>& fe, std::vector<uint32_t>& ef_off, std::vector<uint32_t>& ef_idx) {
  const size_t F = tri.size() / 3, E = edge.size() / 2;
  std::vector<uint32_t> off(nv + 1, 0);
  for (size_t e = 0; e < E; e++) { off[edge[2 * e] + 1]++; off[edge[2 * e + 1] + 1]++; }
  for (uint32_t v = 0; v < nv; v++) off[v + 1] += off[v];
  std::vector<uint32_t> cur(off.begin(), off.end() - 1), inc(2 * E);
  for (size_t e = 0; e < E; e++) { inc[cur[edge[2 * e]]++] = (uint32_t)e; inc[cur[edge[2 * e + 1]]++] = (uint32_t)e; }
  fe.assign(3 * F, 0);
  ef_off.assign(E + 1, 0);
  for (size_t f = 0; f < F; f++)
    for (int j = 0; j < 3; j++) {
      const uint32_t a = tri[3 * f + j], b = tri[3 * f + (j + 1) % 3];
      uint32_t found = 0xFFFFFFFFu;
      for (uint32_t i = off[a]; i < off[a + 1]; i++) {
        const uint32_t e = inc[i];
        if ((edge[2 * e] == a && edge[2 * e + 1] == b) || (edge[2 * e] == b && edge[2 * e + 1] == a)) { found = e; break; }
      }
      if (found == 0xFFFFFFFFu) return false;  // the edge list does not cover the faces
      fe[3 * f + j] = found;
      ef_off[found + 1]++;
    }
  for (size_t e = 0; e < E; e++) ef_off[e + 1] += ef_off[e];
  ef_idx.assign(ef_off[E], 0);
  std::vector<uint32_t> c2(ef_off.begin(), ef_off.end() - 1);
  for (size_t f = 0; f < F; f++)
    for (int j = 0; j < 3; j++) ef_idx[c2[fe[3 * f + j]]++] = (uint32_t)f;
  return true;
}

#define Q_READY()                                                            \
  XRT_READY();                                                               \
  XRT_REQUIRE(s->q, "not a Quality-mode handle: call xrt_set_quality before xrt_finalize_setup"); \
  [[maybe_unused]] QualityState* q = s->q

int read_counters(xrt_solver* s) {
  QualityState* q = s->q;
  XRT_CUDA_OK(cudaGetLastError());  // a launch that was refused (not a fault inside a kernel) shows up only here
  XRT_CUDA_OK(cudaMemcpyAsync(q->h_cnt, q->cnt, N_COUNTERS * sizeof(uint32_t), cudaMemcpyDeviceToHost, s->st));
  XRT_CUDA_OK(cudaStreamSynchronize(s->st));
  const uint32_t ov = q->h_cnt[OVERFLOW];
  if (ov) {
    xrt_set_error(s, "Quality-mode arena overflow:%s%s%s%s%s (raise xrt_quality_params.max_pairs / max_candidates / max_impacts)",
                  ov & OV_PAIRS ? " face pairs" : "", ov & OV_CAND ? " narrow-phase candidates" : "", ov & OV_IMP ? " impacts" : "",
                  ov & OV_QUAD ? " quadratures" : "", ov & OV_REP ? " repulsion items" : "");
    cudaMemsetAsync(q->cnt + OVERFLOW, 0, sizeof(uint32_t), s->st);
    return XRT_ERR_FACE_POOL;
  }
  return XRT_OK;
}

// pairs = [cloth x obstacle, cloth x cloth] (FindImpacts ImpactHelper.cu:226-230, UpdatePairs PBDRepulsion.cu:22-33,
// UpdateProximity ImminentRepulsion.cu:21-28), rows as global face ids
void enqueue_pairs(xrt_solver* s, const DevBVH& cloth, const DevBVH* obstacle, float thickness) {
  QualityState* q = s->q;
  cudaMemsetAsync(q->cnt + N_PAIRS, 0, sizeof(uint32_t), s->st);
  if (obstacle && obstacle->n_obj)
    pairs_traverse(cloth, *obstacle, thickness, q->pw, q->pairs, q->pair_cap, q->cnt + N_PAIRS, s->c.F, q->cnt + OVERFLOW, s->st);
  pairs_traverse(cloth, cloth, thickness, q->pw, q->pairs, q->pair_cap, q->cnt + N_PAIRS, 0u, q->cnt + OVERFLOW, s->st);
}

// FindImpacts on the two ccd trees: 2 traversals + 6 list kernels, impacts -> q->impacts, count -> cnt[N_IMP]
void enqueue_find_impacts(xrt_solver* s, float thickness) {
  QualityState* q = s->q;
  const QMesh m = make_mesh(s);
  cudaStream_t st = s->st;
  enqueue_pairs(s, q->ccd_c, s->o.F ? &q->ccd_o : nullptr, thickness);
  const int G = q->grid;
  ccd_filter_kernel<<<G, QB, 0, st>>>(m, q->pairs, q->cnt + N_PAIRS, thickness, q->masks, q->bsum);
  scan_sums_kernel<<<1, QG_MAX, 0, st>>>(q->bsum, q->boff, nullptr, nullptr, G, nullptr, q->cnt + N_CAND, q->cand_cap, q->cnt + OVERFLOW, OV_CAND);
  ccd_expand_kernel<<<G, QB, 0, st>>>(q->cnt + N_PAIRS, q->masks, q->boff, q->cand, q->cand_cap);
  ccd_test_kernel<<<G, QB, 0, st>>>(m, q->pairs, q->cand, q->cnt + N_CAND, q->cand_t, q->bsum);
  scan_sums_kernel<<<1, QG_MAX, 0, st>>>(q->bsum, q->boff, nullptr, nullptr, G, nullptr, q->cnt + N_IMP, q->imp_cap, q->cnt + OVERFLOW, OV_IMP);
  ccd_compact_kernel<<<G, QB, 0, st>>>(m, q->pairs, q->cand, q->cnt + N_CAND, q->cand_t, q->boff, q->impacts, q->imp_cap);
}

// CollisionStep's setup: x1 snapshots, obstacle velocity, two fresh ccd trees (Entry.cu:143-155)
void enqueue_collision_setup(xrt_solver* s) {
  QualityState* q = s->q;
  cudaStream_t st = s->st;
  snapshot_kernel<<<xrt_grid(s->c.V), XRT_BLOCK, 0, st>>>(s->c.x, s->c.x0, q->cx1, nullptr, s->k.dt, s->c.V);
  if (s->o.V) snapshot_kernel<<<xrt_grid(s->o.V), XRT_BLOCK, 0, st>>>(s->o.x, s->o.x0, q->ox1, q->ov, s->k.dt, s->o.V);
  bvh_build_ccd(q->ccd_c, q->ccd_cw, s->c.x0, s->c.x, s->c.tri, s->c.F, s->p.morton_fast_div, st);
  if (s->o.F) bvh_build_ccd(q->ccd_o, q->ccd_ow, s->o.x0, s->o.x, s->o.tri, s->o.F, s->p.morton_fast_div, st);
}

// independentImpacts on n impacts in `src` -> q->imp_ind, count in cnt[N_IND]
int enqueue_independent(xrt_solver* s, const DImpact* src, uint32_t n, int deform) {
  QualityState* q = s->q;
  cudaStream_t st = s->st;
  impact_keys_kernel<<<xrt_grid(n), XRT_BLOCK, 0, st>>>(src, n, q->sort_key_a, q->sort_val_a);
  cub::DoubleBuffer<uint32_t> k(q->sort_key_a, q->sort_key_b), v(q->sort_val_a, q->sort_val_b);
  size_t bytes = q->sort_bytes;
  // thrust::sort by t (Entry.cu:29); a stable sort here, so equal times keep their list order
  XRT_CUDA_OK(cub::DeviceRadixSort::SortPairs(q->sort_tmp, bytes, k, v, (int)n, 0, 32, st));
  impact_gather_kernel<<<xrt_grid(n), XRT_BLOCK, 0, st>>>(src, v.Current(), n, q->imp_sorted);
  independent_kernel<<<1, IB, 0, st>>>(make_mesh(s), q->imp_sorted, (int)n, deform, q->removed, q->min_index, q->flags,
                                       q->flags + q->imp_cap, q->flags + 2 * (size_t)q->imp_cap, q->imp_ind, q->cnt + N_IND);
  return XRT_OK;
}

void enqueue_isolate(xrt_solver* s, const DImpact* imp, uint32_t n) {
  QualityState* q = s->q;
  isolate_kernel<<<1, IB, 0, s->st>>>(make_mesh(s), s->c.vf_off, s->c.vf_idx, s->o.vf_off, s->o.vf_idx, imp, (int)n, q->ncol,
                                      q->imp_color, q->frontier, s->c.V + s->o.V + 8, q->cnt + N_ZONES);
}

int ensure_opt_capacity(xrt_solver* s, size_t n_imp, size_t n_node, size_t n_zone) {
  QualityState* q = s->q;
  if (n_imp > q->opt_cap_imp) {
    const size_t cap = std::max<size_t>(n_imp * 2, 1024);
    cudaFree(q->o_imp); cudaFree(q->o_idx); cudaFree(q->o_con); cudaFree(q->o_lam); cudaFree(q->o_inc_slot);
    cudaFree(q->o_zone); cudaFree(q->o_stat);
    XRT_CUDA_OK(dev_alloc(&q->o_imp, cap));
    XRT_CUDA_OK(dev_alloc(&q->o_idx, 4 * cap));
    XRT_CUDA_OK(dev_alloc(&q->o_con, cap));
    XRT_CUDA_OK(dev_alloc(&q->o_lam, cap));
    XRT_CUDA_OK(dev_alloc(&q->o_inc_slot, 4 * cap));
    XRT_CUDA_OK(dev_alloc(&q->o_zone, cap));  // zones <= impacts
    XRT_CUDA_OK(dev_alloc(&q->o_stat, 3 * cap));
    q->opt_cap_imp = cap;
  }
  if (n_node > q->opt_cap_node) {
    const size_t cap = std::max<size_t>(n_node * 2, 1024);
    cudaFree(q->o_node); cudaFree(q->o_inc_off); cudaFree(q->o_X); cudaFree(q->o_Xn); cudaFree(q->o_G); cudaFree(q->o_obj);
    cudaFree(q->o_skip);
    XRT_CUDA_OK(dev_alloc(&q->o_skip, cap));
    XRT_CUDA_OK(dev_alloc(&q->o_node, cap));
    XRT_CUDA_OK(dev_alloc(&q->o_inc_off, cap + 1));
    XRT_CUDA_OK(dev_alloc(&q->o_X, 3 * cap));
    XRT_CUDA_OK(dev_alloc(&q->o_Xn, 3 * cap));
    XRT_CUDA_OK(dev_alloc(&q->o_G, 3 * cap));
    XRT_CUDA_OK(dev_alloc(&q->o_obj, cap));
    q->opt_cap_node = cap;
  }
  (void)n_zone;
  return XRT_OK;
}

// ParallelImpactZoneOptimizer ctor (ParallelImpactZoneOptimizer.cu:167-282) on the host -- the impacts of the
// active zones are a few thousand rows at most -- then the one-launch solve.  imp: impacts of the zones back to
// back, color[i] = zone of impact i (ascending).  h_invm: the cloth's inverse masses.
int solve_zones(xrt_solver* s, const std::vector<DImpact>& imp, const std::vector<int>& color, int nz, int deform,
                float thickness, float obstacle_mass, int* stats3) {
  QualityState* q = s->q;
  const size_t n = imp.size();
  if (stats3) stats3[0] = stats3[1] = stats3[2] = 0;
  if (n == 0 || nz == 0) return XRT_OK;
  const uint32_t V = s->c.V;
  auto is_free = [&](uint32_t g) { return g < V && s->c_invm[g] != 0.0f; };
  // CollectCollisionNodes (:14-34): the variables are the free cloth nodes, or every node in the deform pass
  struct Ent { int color; uint32_t node; uint32_t slot; };
  std::vector<Ent> ent;
  ent.reserve(4 * n);
  for (size_t i = 0; i < n; i++)
    for (int j = 0; j < 4; j++) {
      const uint32_t g = imp[i].n[j];
      if (deform == 1 || is_free(g)) ent.push_back(Ent{color[i], g, (uint32_t)(4 * i + j)});
    }
  // sort_by_key((colour, node)) :203-207; the slot as the last key makes the order total
  std::sort(ent.begin(), ent.end(), [](const Ent& a, const Ent& b) {
    if (a.color != b.color) return a.color < b.color;
    if (a.node != b.node) return a.node < b.node;
    return a.slot < b.slot;
  });
  // SetDiff on the node only + scan + SetIndices + unique_by_key (:36-49,209-227)
  std::vector<int> indices(4 * n, -1);
  std::vector<uint32_t> nodes, inc_off, inc_slot;
  std::vector<int> ncolor;
  int run = 0;
  for (size_t i = 0; i < ent.size(); i++) {
    const bool first = i == 0 || ent[i].node != ent[i - 1].node;
    if (i > 0 && first) run++;
    indices[ent[i].slot] = run;
    if (first) { nodes.push_back(ent[i].node); ncolor.push_back(ent[i].color); inc_off.push_back((uint32_t)inc_slot.size()); }
    inc_slot.push_back(ent[i].slot);
  }
  inc_off.push_back((uint32_t)inc_slot.size());
  const size_t nn = nodes.size();
  if (nn == 0) return XRT_OK;
  std::vector<uint8_t> skip(nn, 0);
  {
    std::vector<std::pair<uint32_t, uint32_t>> byn(nn);
    for (size_t i = 0; i < nn; i++) byn[i] = {nodes[i], (uint32_t)i};
    std::sort(byn.begin(), byn.end());
    for (size_t i = 0; i + 1 < nn; i++)
      if (byn[i].first == byn[i + 1].first) skip[byn[i].second] = 1;
  }
  std::vector<int4> zone(nz);
  {
    size_t no = 0, io = 0;
    std::vector<size_t> cn((size_t)nz, 0), ci((size_t)nz, 0);
    for (size_t i = 0; i < nn; i++)
      if (ncolor[i] >= 0 && ncolor[i] < nz) cn[(size_t)ncolor[i]]++;
    for (size_t i = 0; i < n; i++)
      if (color[i] >= 0 && color[i] < nz) ci[(size_t)color[i]]++;
    for (int c = 0; c < nz; c++) {
      zone[c] = make_int4((int)no, (int)cn[c], (int)io, (int)ci[c]);
      no += cn[c]; io += ci[c];
    }
  }
  // inv_mass = mean of the per-node inverse masses (thrust::reduce :237-243)
  float inv_mass = 0.0f;
  for (size_t i = 0; i < nn; i++) inv_mass += nodes[i] < V ? s->c_invm[nodes[i]] : 1.0f / obstacle_mass;
  inv_mass = inv_mass / (float)nn;
  if (int rc = ensure_opt_capacity(s, n, nn, (size_t)nz)) return rc;
  cudaStream_t st = s->st;
  XRT_CUDA_OK(cudaMemcpyAsync(q->o_imp, imp.data(), n * sizeof(DImpact), cudaMemcpyHostToDevice, st));
  XRT_CUDA_OK(cudaMemcpyAsync(q->o_idx, indices.data(), 4 * n * sizeof(int), cudaMemcpyHostToDevice, st));
  XRT_CUDA_OK(cudaMemcpyAsync(q->o_node, nodes.data(), nn * sizeof(uint32_t), cudaMemcpyHostToDevice, st));
  XRT_CUDA_OK(cudaMemcpyAsync(q->o_skip, skip.data(), nn * sizeof(uint8_t), cudaMemcpyHostToDevice, st));
  XRT_CUDA_OK(cudaMemcpyAsync(q->o_inc_off, inc_off.data(), (nn + 1) * sizeof(uint32_t), cudaMemcpyHostToDevice, st));
  XRT_CUDA_OK(cudaMemcpyAsync(q->o_inc_slot, inc_slot.data(), inc_slot.size() * sizeof(uint32_t), cudaMemcpyHostToDevice, st));
  XRT_CUDA_OK(cudaMemcpyAsync(q->o_zone, zone.data(), nz * sizeof(int4), cudaMemcpyHostToDevice, st));
  OptArgs a{};
  a.m = make_mesh(s);
  a.cx1 = q->cx1; a.ox1 = q->ox1;
  a.imp = q->o_imp; a.idx = q->o_idx; a.node = q->o_node; a.skip = q->o_skip; a.inc_off = q->o_inc_off; a.inc_slot = q->o_inc_slot;
  a.X = q->o_X; a.Xn = q->o_Xn; a.G = q->o_G; a.obj = q->o_obj; a.con = q->o_con; a.lam = q->o_lam;
  a.zone = q->o_zone; a.stat = q->o_stat;
  a.out_cx = s->c.x; a.out_ox = s->o.x;
  a.thickness = thickness; a.obstacle_mass = obstacle_mass; a.inv_mass = inv_mass; a.mu = 1e3f;
  a.deform = deform;
  {
    // shared-memory variant when the largest zone of this launch fits (opt-in dynamic shared memory, once per device)
    size_t need = 0;
    for (int c = 0; c < nz; c++) need = std::max(need, zone_smem_bytes((size_t)zone[c].y, (size_t)zone[c].w));
    static const size_t limit = [] {
      const char* e = getenv("XRT_ZONE_SMEM");  // measurement switch: 0 keeps the optimiser's state in global memory
      if (e && atoi(e) == 0) return (size_t)0;
      return (size_t)200 * 1024;
    }();
    if (need <= limit) {
      static thread_local int configured_dev = -1;
      int dev = -1;
      cudaGetDevice(&dev);
      if (configured_dev != dev) {
        XRT_CUDA_OK(cudaFuncSetAttribute(zone_optimize_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)limit));
        configured_dev = dev;
      }
      zone_optimize_kernel<true><<<nz, QB, need, st>>>(a);
    } else {
      zone_optimize_kernel<false><<<nz, QB, 0, st>>>(a);
    }
  }
  XRT_CUDA_OK(cudaGetLastError());
  // the copies above read pageable host vectors: wait before they go out of scope; the stats ride along
  std::vector<int> hs(3 * (size_t)nz);
  XRT_CUDA_OK(cudaMemcpyAsync(hs.data(), q->o_stat, hs.size() * sizeof(int), cudaMemcpyDeviceToHost, st));
  XRT_CUDA_OK(cudaStreamSynchronize(st));
  int outer = 0, ls = 0, conv = 1;
  for (int c = 0; c < nz; c++) { outer = std::max(outer, hs[3 * c]); ls = std::max(ls, hs[3 * c + 1]); conv &= hs[3 * c + 2]; }
  q->stats.outer_iterations = outer;
  q->stats.line_search_steps = ls;
  if (stats3) { stats3[0] = outer; stats3[1] = ls; stats3[2] = conv; }
  return XRT_OK;
}

// ImpactZone::Intersects (ImpactZone.cu:74-84): shared Node::index, which is LOCAL to the mesh, so cloth node 5
// and obstacle node 5 count as shared (kept as in the reference)
bool zone_intersects(HostZone& a, HostZone& b) {
  if (!a.sorted) { std::sort(a.node_idx.begin(), a.node_idx.end()); a.sorted = true; }
  if (!b.sorted) { std::sort(b.node_idx.begin(), b.node_idx.end()); b.sorted = true; }
  size_t i = 0, j = 0;
  while (i < a.node_idx.size() && j < b.node_idx.size()) {
    if (a.node_idx[i] < b.node_idx[j]) i++;
    else if (a.node_idx[i] > b.node_idx[j]) j++;
    else return true;
  }
  return false;
}
// ImpactZone::Merge (ImpactZone.cu:57-72)
bool zone_merge(HostZone& a, HostZone& b) {
  if (!b.enabled) return false;
  if (!zone_intersects(a, b)) return false;
  a.imp.insert(a.imp.end(), b.imp.begin(), b.imp.end());
  const size_t mid = a.node_idx.size();  // both lists are sorted (zone_intersects): merge instead of sorting again
  a.node_idx.insert(a.node_idx.end(), b.node_idx.begin(), b.node_idx.end());
  std::inplace_merge(a.node_idx.begin(), a.node_idx.begin() + (std::ptrdiff_t)mid, a.node_idx.end());
  a.active = true; a.sorted = true; a.enabled = true;
  b.enabled = false;
  return true;
}
// AddImpacts (ImpactZoneHelper.cu:344-384) with the zone colours of the new impacts already known
void add_impacts(uint32_t V, std::vector<HostZone>& zones, const std::vector<DImpact>& nimp, const std::vector<int>& col, int nc) {
  const size_t n = nimp.size();
  // stable_sort_by_key(colours) + ranges from the per-zone counts (:305-331): impacts left at -1 by the 100-zone
  // cap sort to the front and shift every range, as in the reference
  std::vector<size_t> order(n);
  std::vector<size_t> cnt((size_t)nc + 1, 0);
  {
    std::vector<size_t> at((size_t)nc + 2, 0);  // bucket c + 1 (colour -1 first), a stable counting sort
    for (size_t i = 0; i < n; i++) at[(size_t)(col[i] + 1) + 1]++;
    for (size_t c = 1; c < at.size(); c++) at[c] += at[c - 1];
    for (size_t i = 0; i < n; i++) {
      order[at[(size_t)(col[i] + 1)]++] = i;
      if (col[i] >= 0) cnt[(size_t)col[i]]++;
    }
  }
  std::vector<HostZone> nz((size_t)nc);
  size_t lhs = 0;
  for (int c = 0; c < nc; c++) {
    for (size_t i = lhs; i < lhs + cnt[c] && i < n; i++) {
      const DImpact& im = nimp[order[i]];
      nz[c].imp.push_back(im);
      for (int j = 0; j < 4; j++) nz[c].node_idx.push_back((int)(im.n[j] < V ? im.n[j] : im.n[j] - V));  // FlattenNode_Kernel
    }
    lhs += cnt[c];
  }
  const size_t n_zone = zones.size();
  for (size_t i = 0; i < n_zone; i++) { zones[i].prev_active = zones[i].active; zones[i].active = false; }
  for (size_t i = 0; i < n_zone; i++)
    for (size_t j = 0; j < nz.size(); j++) zone_merge(zones[i], nz[j]);
  for (size_t j = 0; j < nz.size(); j++)
    if (nz[j].enabled) zones.push_back(std::move(nz[j]));
  // All pairs, in the reference's order.  A pair of zones that both stayed as they were through the whole previous
  // round and this one so far has been tested with exactly these contents before and did not intersect (a zone that
  // grew DURING the previous round's pass may have been tested against its earlier neighbours before it grew, hence
  // prev_active), so that test is skipped: the zones and their order come out the same.
  for (size_t i = 0; i < zones.size(); i++)
    for (size_t j = i + 1; j < zones.size(); j++)
      if (zones[i].active || zones[j].active || zones[i].prev_active || zones[j].prev_active) zone_merge(zones[i], zones[j]);
  std::vector<HostZone> kept;
  for (auto& z : zones)
    if (z.enabled) kept.push_back(std::move(z));
  zones.swap(kept);
}

// CollisionStep, impact_zone/Entry.cu:140-256
int collision_step(xrt_solver* s) {
  QualityState* q = s->q;
  cudaStream_t st = s->st;
  enqueue_collision_setup(s);
  float obstacle_mass = q->qp.obstacle_mass;
  const float thickness = q->qp.impact_zone_thickness;
  q->stats = xrt_collision_stats{};
  bool success = false;
  // XRT_QUALITY_PROF=2: host-clock split of the rounds (every stage ends in a stream sync here)
  static const bool prof = [] { const char* e = getenv("XRT_QUALITY_PROF"); return e && atoi(e) >= 2; }();
  double t_find = 0, t_ind = 0, t_iso = 0, t_host = 0, t_solve = 0;
  auto now = [] { return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now().time_since_epoch()).count(); };
  double t0 = 0;
  std::vector<DImpact> ind, flat;
  std::vector<int> col, fcol;
  for (int deform = 0; deform < 2 && !success; deform++) {
    std::vector<HostZone> zones;
    q->stats.deform = deform;
    for (int round = 0; round < 100; round++) {
      if (prof) { cudaStreamSynchronize(st); t0 = now(); }
      enqueue_find_impacts(s, thickness);
      if (int rc = read_counters(s)) return rc;
      if (prof) { t_find += now() - t0; t0 = now(); }
      const uint32_t n_new = q->h_cnt[N_IMP];
      q->stats.n_pairs = (int32_t)q->h_cnt[N_PAIRS];
      q->stats.n_candidates = (int32_t)q->h_cnt[N_CAND];
      if (n_new == 0) { success = true; break; }
      if (int rc = enqueue_independent(s, q->impacts, n_new, deform)) return rc;
      // the zone kernel reads the count of independent impacts from the host: one more small round trip
      XRT_CUDA_OK(cudaMemcpyAsync(q->h_cnt + N_IND, q->cnt + N_IND, sizeof(uint32_t), cudaMemcpyDeviceToHost, st));
      XRT_CUDA_OK(cudaStreamSynchronize(st));
      const uint32_t n_ind = q->h_cnt[N_IND];
      if (prof) { t_ind += now() - t0; t0 = now(); }
      q->stats.total_impacts += (int32_t)n_ind;
      q->stats.rounds++;
      ind.resize(n_ind);
      col.resize(n_ind);
      int nc = 0;
      if (n_ind) {
        enqueue_isolate(s, q->imp_ind, n_ind);
        XRT_CUDA_OK(cudaMemcpyAsync(ind.data(), q->imp_ind, n_ind * sizeof(DImpact), cudaMemcpyDeviceToHost, st));
        XRT_CUDA_OK(cudaMemcpyAsync(col.data(), q->imp_color, n_ind * sizeof(int), cudaMemcpyDeviceToHost, st));
        XRT_CUDA_OK(cudaMemcpyAsync(q->h_cnt + N_ZONES, q->cnt + N_ZONES, sizeof(uint32_t), cudaMemcpyDeviceToHost, st));
        XRT_CUDA_OK(cudaStreamSynchronize(st));
        nc = (int)q->h_cnt[N_ZONES];
      }
      if (prof) { t_iso += now() - t0; t0 = now(); }
      add_impacts(s->c.V, zones, ind, col, nc);
      // FlattenImpactZones (ImpactZoneHelper.cu:386-416): the impacts of the active zones back to back
      flat.clear();
      fcol.clear();
      int nz = 0;
      for (auto& z : zones) {
        if (!z.active) continue;
        for (auto& im : z.imp) { flat.push_back(im); fcol.push_back(nz); }
        nz++;
      }
      if (prof) { t_host += now() - t0; t0 = now(); }
      if (int rc = solve_zones(s, flat, fcol, nz, deform, thickness, obstacle_mass, nullptr)) return rc;
      if (prof) { cudaStreamSynchronize(st); t_solve += now() - t0; }
      bvh_update(q->ccd_c, q->ccd_cw, s->c.x0, s->c.x, s->c.tri, 1, st);
      if (deform == 1) {
        if (s->o.F) bvh_update(q->ccd_o, q->ccd_ow, s->o.x0, s->o.x, s->o.tri, 1, st);
        obstacle_mass *= 0.5f;
      }
    }
  }
  q->stats.success = success ? 1 : 0;
  if (prof)
    fprintf(stderr, "[xrt quality] CollisionStep rounds %d: find_impacts %.2f | independent %.2f | isolate + copies %.2f | host zones %.2f | "
            "zone optimiser %.2f ms\n", q->stats.rounds, t_find, t_ind, t_iso, t_host, t_solve);
  return XRT_OK;
}

// cloth_bvh_->UpdateData(faces, 0, false) + Construct (ClothSolverBase.cpp:277-278)
void enqueue_begin_frame(xrt_solver* s) {
  QualityState* q = s->q;
  bvh_build(q->cbvh, q->cbw, s->c.x0, s->c.tri, s->c.F, s->p.morton_fast_div, s->st);
  q->frame_begun = true;
}

// ImminentRepulsion::UpdateProximity + Generate
void enqueue_imminent_generate(xrt_solver* s) {
  QualityState* q = s->q;
  const QMesh m = make_mesh(s);
  cudaStream_t st = s->st;
  const int G = q->grid;
  enqueue_pairs(s, q->cbvh, s->o.F ? &s->bvh : nullptr, 1e-6f);
  quad_count_kernel<<<G, QB, 0, st>>>(m, q->pairs, q->cnt + N_PAIRS, q->masks, q->bsum, q->bsum + QG_MAX);
  scan_sums_kernel<<<1, QG_MAX, 0, st>>>(q->bsum, q->boff, q->bsum + QG_MAX, q->boff + QG_MAX, G, q->cnt + N_QUAD_EE, q->cnt + N_QUAD,
                                        q->quad_cap, q->cnt + OVERFLOW, OV_QUAD);
  quad_fill_kernel<<<G, QB, 0, st>>>(m, q->pairs, q->cnt + N_PAIRS, q->masks, q->boff, q->boff + QG_MAX, q->quad_n, q->quad_b, q->quad_cap);
}
int enqueue_imminent_solve(xrt_solver* s) {
  QualityState* q = s->q;
  imminent_impulse_kernel<<<q->grid, QB, 0, s->st>>>(make_mesh(s), q->quad_n, q->quad_b, q->cnt + N_QUAD, q->qp.imminent_thickness, s->k.dt,
                                                     s->c.delta);
  jacobi_apply_kernel<<<xrt_grid(s->c.V), XRT_BLOCK, 0, s->st>>>(s->c.x, s->c.delta, s->c.V, q->qp.relaxation_rate);
  return 2;
}
// PBDRepulsionSolver::UpdatePairs: cloth tree Update(ccd = true), pairs, and the items the next iterations evaluate
int enqueue_rep_update_pairs(xrt_solver* s) {
  QualityState* q = s->q;
  const QMesh m = make_mesh(s);
  cudaStream_t st = s->st;
  const int G = q->grid;
  bvh_update(q->cbvh, q->cbw, s->c.x0, s->c.x, s->c.tri, 1, st);
  enqueue_pairs(s, q->cbvh, s->o.F ? &s->bvh : nullptr, 1e-6f);
  rep_count_kernel<<<G, QB, 0, st>>>(m, q->pairs, q->cnt + N_PAIRS, q->masks, q->bsum);
  scan_sums_kernel<<<1, QG_MAX, 0, st>>>(q->bsum, q->boff, nullptr, nullptr, G, nullptr, q->cnt + N_REP, q->rep_cap, q->cnt + OVERFLOW, OV_REP);
  rep_fill_kernel<<<G, QB, 0, st>>>(m, q->pairs, q->cnt + N_PAIRS, q->masks, q->boff, q->rep_items, q->rep_cap);
  return 2 + 8 + 3;
}
int enqueue_rep_solve(xrt_solver* s) {
  QualityState* q = s->q;
  rep_solve_kernel<<<q->grid, QB, 0, s->st>>>(make_mesh(s), q->rep_items, q->cnt + N_REP, q->qp.pbd_thickness, s->c.delta);
  jacobi_apply_kernel<<<xrt_grid(s->c.V), XRT_BLOCK, 0, s->st>>>(s->c.x, s->c.delta, s->c.V, q->qp.relaxation_rate);
  return 2;
}
// one pass of the Quality inner loop, ClothSolverBase.cpp:296-328
int enqueue_iteration(xrt_solver* s, int iteration) {
  QualityState* q = s->q;
  cudaStream_t st = s->st;
  int launches = 0;
  if (q->qp.enable_imminent_repulsion) launches += enqueue_imminent_solve(s);
  if (q->qp.enable_pbd_repulsion) {
    if (iteration % 10 == 0) launches += enqueue_rep_update_pairs(s);
    launches += enqueue_rep_solve(s);
  }
  if (iteration == 0) fem_reset_lambda(q->fem, st);
  launches += fem_iteration(q->fem, s->c.x, s->k.dt, s->p.solve_bending != 0, st);
  if (s->p.geodesic_lra && s->c.n_lra) { launch_lra(s->c, s->k, st); launches++; }
  if (s->o.F && s->c.n_bind) { launch_binding(s->c, s->o, st); launches++; }
  return launches;
}

void impacts_to_rows(const std::vector<DImpact>& imp, size_t n, uint32_t* nodes, float* twn) {
  for (size_t i = 0; i < n; i++) {
    memcpy(nodes + 4 * i, imp[i].n, 16);
    float* o = twn + 9 * i;
    o[0] = imp[i].t;
    memcpy(o + 1, imp[i].w, 16);
    memcpy(o + 5, imp[i].nrm, 12);
    o[8] = imp[i].is_vf ? 1.0f : 0.0f;
  }
}
void rows_to_impacts(const uint32_t* nodes, const float* twn, size_t n, std::vector<DImpact>& imp) {
  imp.resize(n);
  for (size_t i = 0; i < n; i++) {
    memcpy(imp[i].n, nodes + 4 * i, 16);
    if (twn) {
      const float* o = twn + 9 * i;
      imp[i].t = o[0];
      memcpy(imp[i].w, o + 1, 16);
      memcpy(imp[i].nrm, o + 5, 12);
      imp[i].is_vf = o[8] > 0.5f ? 1u : 0u;
    } else {
      imp[i].t = 0.f;
      memset(imp[i].w, 0, 16);
      memset(imp[i].nrm, 0, 12);
      imp[i].is_vf = 0;
    }
  }
}

}  // namespace

// ------------------------------------------------------------------------------------------------
// ABI
// ------------------------------------------------------------------------------------------------

extern "C" int xrt_default_quality_params(xrt_quality_params* p) {
  if (!p) return XRT_ERR_INVALID;
  memset(p, 0, sizeof(*p));
  p->xx_stiffness = p->yy_stiffness = p->xy_stiffness = 1.0f;  // FEMFabricSettings, config/Base.hpp:47-57
  p->xy_poisson_ratio = p->yx_poisson_ratio = 0.3f;
  p->bending_stiffness = 1.0f;
  p->enable_imminent_repulsion = 1;  // RepulsionSettings, :108-114
  p->imminent_thickness = 1e-3f;
  p->enable_pbd_repulsion = 1;
  p->pbd_thickness = 1e-4f;
  p->relaxation_rate = 0.25f;
  p->obstacle_mass = 1e3f;  // ImpactZoneSettings, :116-119
  p->impact_zone_thickness = 1e-4f;
  return XRT_OK;
}

extern "C" int xrt_set_quality(xrt_solver* s, const xrt_quality_params* qp, const uint32_t* strain_colors) {
  if (!s || !qp) return XRT_ERR_INVALID;
  XRT_REQUIRE(!s->finalized, "set_quality: call before finalize_setup");
  XRT_REQUIRE(!s->c_tri.empty(), "set_quality: add the cloth first");
  if (!s->q) s->q = new QualityState();
  s->q->qp = *qp;
  if (strain_colors) {
    s->q->strain_colors.assign(strain_colors, strain_colors + s->c_tri.size() / 3);
    s->q->strain_colors_set = true;
  }
  return XRT_OK;
}

int xrt_quality_finalize(xrt_solver* s) {
  QualityState* q = s->q;
  const ClothDev& c = s->c;
  const ObstacleDev& o = s->o;
  const uint32_t NV = c.V + o.V;
  // --- topology ---
  std::vector<uint32_t> fe, efo, efi;
  XRT_REQUIRE(build_edge_topology(s->c_tri, s->c_edge, c.V, fe, efo, efi), "set_quality: the cloth edge list does not cover its faces");
  XRT_CUDA_OK(dev_upload(&q->cfe, fe));
  XRT_CUDA_OK(dev_upload(&q->cef_off, efo));
  XRT_CUDA_OK(dev_upload(&q->cef_idx, efi));
  {
    std::vector<uint32_t> vfo, vfi, rt;
    xrt_build_vf(s->c_tri, c.V, vfo, vfi);
    xrt_build_rtri(s->c_tri, vfo, vfi, rt);  // ClothSolverBase::SetupRTriangle, cloth side
    XRT_CUDA_OK(dev_upload(&q->crtri, rt));
  }
  XRT_CUDA_OK(dev_alloc(&q->cfn0, c.F));
  face_normal_kernel<<<xrt_grid(c.F), XRT_BLOCK, 0, s->st>>>(c.x0, c.tri, q->cfn0, c.F);
  XRT_CUDA_OK(dev_alloc(&q->cx1, c.V));
  if (o.V) {
    q->Eo = (uint32_t)(s->o_edge.size() / 2);
    XRT_REQUIRE(build_edge_topology(s->o_tri, s->o_edge, o.V, fe, efo, efi), "set_quality: the obstacle edge list does not cover its faces");
    XRT_CUDA_OK(dev_upload(&q->ofe, fe));
    XRT_CUDA_OK(dev_upload(&q->oef_off, efo));
    XRT_CUDA_OK(dev_upload(&q->oef_idx, efi));
    std::vector<uint2> he(q->Eo);
    for (uint32_t i = 0; i < q->Eo; i++) he[i] = make_uint2(s->o_edge[2 * i], s->o_edge[2 * i + 1]);
    XRT_CUDA_OK(dev_upload(&q->oedge, he));
    XRT_CUDA_OK(dev_alloc(&q->ofn0, o.F));
    face_normal_kernel<<<xrt_grid(o.F), XRT_BLOCK, 0, s->st>>>(o.x0, o.tri, q->ofn0, o.F);
    XRT_CUDA_OK(dev_alloc(&q->ov, o.V));
    XRT_CUDA_OK(cudaMemsetAsync(q->ov, 0, o.V * sizeof(float4), s->st));
    XRT_CUDA_OK(dev_alloc(&q->ox1, o.V));
  }
  // --- trees ---
  if (bvh_alloc(&q->cbvh, &q->cbw, c.F) || bvh_alloc(&q->ccd_c, &q->ccd_cw, c.F) || (o.F && bvh_alloc(&q->ccd_o, &q->ccd_ow, o.F)) ||
      pairs_alloc(&q->pw, c.F)) {
    xrt_set_error(s, "Quality mode: BVH allocation failed");
    return XRT_ERR_CUDA;
  }
  // --- arenas ---
  q->pair_cap = q->qp.max_pairs ? q->qp.max_pairs : std::max<uint32_t>(1u << 20, 256u * c.F);
  XRT_REQUIRE(q->pair_cap < (1u << 28), "max_pairs must stay below 2^28");
  q->cand_cap = q->qp.max_candidates ? q->qp.max_candidates : 4u * q->pair_cap;
  q->imp_cap = q->qp.max_impacts ? q->qp.max_impacts : std::max<uint32_t>(1u << 20, q->pair_cap / 2u);
  q->quad_cap = q->pair_cap;       // <= 12 quadratures per pair, in practice a few per hundred pairs
  q->rep_cap = 4u * q->pair_cap;   // <= 12 repulsion items per pair; ownership leaves about 3
  XRT_CUDA_OK(dev_alloc(&q->pairs, q->pair_cap));
  XRT_CUDA_OK(dev_alloc(&q->masks, q->pair_cap));
  XRT_CUDA_OK(dev_alloc(&q->cand, q->cand_cap));
  XRT_CUDA_OK(dev_alloc(&q->cand_t, q->cand_cap));
  XRT_CUDA_OK(dev_alloc(&q->impacts, q->imp_cap));
  XRT_CUDA_OK(dev_alloc(&q->imp_sorted, q->imp_cap));
  XRT_CUDA_OK(dev_alloc(&q->imp_ind, q->imp_cap));
  XRT_CUDA_OK(dev_alloc(&q->sort_key_a, q->imp_cap));
  XRT_CUDA_OK(dev_alloc(&q->sort_key_b, q->imp_cap));
  XRT_CUDA_OK(dev_alloc(&q->sort_val_a, q->imp_cap));
  XRT_CUDA_OK(dev_alloc(&q->sort_val_b, q->imp_cap));
  {
    cub::DoubleBuffer<uint32_t> k(q->sort_key_a, q->sort_key_b), v(q->sort_val_a, q->sort_val_b);
    XRT_CUDA_OK(cub::DeviceRadixSort::SortPairs(nullptr, q->sort_bytes, k, v, (int)q->imp_cap, 0, 32, s->st));
    XRT_CUDA_OK(cudaMalloc(&q->sort_tmp, q->sort_bytes ? q->sort_bytes : 16));
  }
  XRT_CUDA_OK(dev_alloc(&q->quad_n, q->quad_cap));
  XRT_CUDA_OK(dev_alloc(&q->quad_b, q->quad_cap));
  XRT_CUDA_OK(dev_alloc(&q->rep_items, q->rep_cap));
  XRT_CUDA_OK(dev_alloc(&q->bsum, 2 * QG_MAX));
  XRT_CUDA_OK(dev_alloc(&q->boff, 2 * QG_MAX));
  XRT_CUDA_OK(dev_alloc(&q->cnt, N_COUNTERS));
  XRT_CUDA_OK(cudaMemsetAsync(q->cnt, 0, N_COUNTERS * sizeof(uint32_t), s->st));
  XRT_CUDA_OK(cudaMallocHost((void**)&q->h_cnt, N_COUNTERS * sizeof(uint32_t)));
  memset(q->h_cnt, 0, N_COUNTERS * sizeof(uint32_t));
  {
    cudaDeviceProp prop;
    XRT_CUDA_OK(cudaGetDeviceProperties(&prop, s->p.device));
    q->grid = std::min(QG_MAX, prop.multiProcessorCount * 4);  // 148 x 4 on a B200: every SM holds 4 CTAs of 256
  }
  XRT_CUDA_OK(dev_alloc(&q->removed, NV));
  XRT_CUDA_OK(dev_alloc(&q->min_index, NV));
  XRT_CUDA_OK(dev_alloc(&q->ncol, NV));
  {
    std::vector<int> outside(NV, Q_NODE_OUTSIDE);
    XRT_CUDA_OK(cudaMemcpy(q->ncol, outside.data(), NV * sizeof(int), cudaMemcpyHostToDevice));
  }
  XRT_CUDA_OK(dev_alloc(&q->flags, 3 * (size_t)q->imp_cap));
  XRT_CUDA_OK(dev_alloc(&q->frontier, 2 * ((size_t)NV + 8)));
  XRT_CUDA_OK(dev_alloc(&q->imp_color, q->imp_cap));
  // --- FEM (SetupInternalDynamics, ClothSolverBase.cpp:468-486): strain per face, isometric bending per hinge ---
  if (!q->strain_colors_set) {
    q->strain_colors.resize(c.F);
    if (xrt_greedy_color(s->c_tri.data(), c.F, 3, c.V, q->strain_colors.data(), nullptr)) {
      xrt_set_error(s, "strain colouring failed");
      return XRT_ERR_INVALID;
    }
  }
  const FemMaterial mat{q->qp.xx_stiffness, q->qp.yy_stiffness, q->qp.xy_stiffness, q->qp.xy_poisson_ratio, q->qp.yx_poisson_ratio,
                        q->qp.bending_stiffness};
  const uint32_t nb = (uint32_t)s->b_theta.size();
  if (int e = fem_setup(&q->fem, c.rest, c.V, c.tri, s->c_tri.data(), c.F, s->b_idx.data(), nb, q->strain_colors.data(),
                        s->b_color.data(), mat, s->st)) {
    xrt_set_error(s, "FEM setup failed: %s", cudaGetErrorString((cudaError_t)e));
    return XRT_ERR_CUDA;
  }
  XRT_CUDA_OK(cudaStreamSynchronize(s->st));
  return XRT_OK;
}

void xrt_quality_destroy(xrt_solver* s) {
  QualityState* q = s->q;
  if (!q) return;
  cudaFree(q->cfe); cudaFree(q->ofe); cudaFree(q->cef_off); cudaFree(q->cef_idx); cudaFree(q->oef_off); cudaFree(q->oef_idx);
  cudaFree(q->crtri); cudaFree(q->oedge); cudaFree(q->cfn0); cudaFree(q->ofn0); cudaFree(q->ov); cudaFree(q->cx1); cudaFree(q->ox1);
  if (q->cbvh.n_obj) bvh_free(&q->cbvh, &q->cbw);
  if (q->ccd_c.n_obj) bvh_free(&q->ccd_c, &q->ccd_cw);
  if (q->ccd_o.n_obj) bvh_free(&q->ccd_o, &q->ccd_ow);
  pairs_free(&q->pw);
  cudaFree(q->pairs); cudaFree(q->masks); cudaFree(q->cand); cudaFree(q->cand_t);
  cudaFree(q->impacts); cudaFree(q->imp_sorted); cudaFree(q->imp_ind);
  cudaFree(q->sort_key_a); cudaFree(q->sort_key_b); cudaFree(q->sort_val_a); cudaFree(q->sort_val_b); cudaFree(q->sort_tmp);
  cudaFree(q->quad_n); cudaFree(q->quad_b); cudaFree(q->rep_items); cudaFree(q->bsum); cudaFree(q->boff); cudaFree(q->cnt);
  if (q->h_cnt) cudaFreeHost(q->h_cnt);
  cudaFree(q->removed); cudaFree(q->min_index); cudaFree(q->ncol); cudaFree(q->flags); cudaFree(q->frontier); cudaFree(q->imp_color);
  cudaFree(q->o_imp); cudaFree(q->o_idx); cudaFree(q->o_node); cudaFree(q->o_inc_off); cudaFree(q->o_inc_slot);
  cudaFree(q->o_X); cudaFree(q->o_Xn); cudaFree(q->o_G); cudaFree(q->o_obj); cudaFree(q->o_con); cudaFree(q->o_lam);
  cudaFree(q->o_zone); cudaFree(q->o_stat); cudaFree(q->o_skip);
  fem_free(&q->fem);
  delete q;
  s->q = nullptr;
}

// ClothSolverBase::SimulateQuality, ClothSolverBase.cpp:263-362
int xrt_quality_step_frame(xrt_solver* s, float* solver_ms) {
  QualityState* q = s->q;
  cudaStream_t st = s->st;
  const xrt_params& p = s->p;
  XRT_CUDA_OK(cudaEventRecord(s->ev0, st));  // Timer::StartTimerGPU("Solver_Total")
  if (s->target_pending && s->o.F)          // UpdateObstacleAnimationFrame (impl/smpl/ClothSolver.cpp:257-261)
    bvh_build(s->bvh, s->bw, s->o.x0, s->o.tri, s->o.F, p.morton_fast_div, st);
  enqueue_begin_frame(s);
  if (q->qp.enable_imminent_repulsion) enqueue_imminent_generate(s);
  if (s->o.F) bvh_update(s->bvh, s->bw, s->o.x0, s->o.x, s->o.tri, 1, st);  // obstacle_bvh_->Update(faces, true)
  uint32_t launches = 0;
  // XRT_QUALITY_PROF=1: where a Quality frame's time goes (setup | iterations | CollisionStep), one line per frame
  static const bool prof = [] { const char* e = getenv("XRT_QUALITY_PROF"); return e && atoi(e) != 0; }();
  cudaEvent_t pe[3] = {nullptr, nullptr, nullptr};
  if (prof) {
    for (auto& e : pe) cudaEventCreate(&e);
    cudaEventRecord(pe[0], st);
  }
  for (int substep = 0; substep < p.num_substeps; substep++) {
    launch_predict(s->c, s->k, st);
    for (int it = 0; it < p.num_iterations; it++) launches += (uint32_t)enqueue_iteration(s, it);
    if (prof && substep == 0) cudaEventRecord(pe[1], st);
    if (int rc = collision_step(s)) return rc;
    if (prof && substep == 0) cudaEventRecord(pe[2], st);
    launch_finalize(s->c, s->k, st);
  }
  if (s->o.V) {  // obstacle_->PredictedToPositions(); UpdateNormals
    XRT_CUDA_OK(cudaMemcpyAsync(s->o.x0, s->o.x, s->o.V * sizeof(float4), cudaMemcpyDeviceToDevice, st));
    launch_normals(s->o.x0, s->o.n, s->o.tri, s->o.vf_off, s->o.vf_idx, s->o.V, st);
  }
  launch_normals(s->c.x0, s->c.n, s->c.tri, s->c.vf_off, s->c.vf_idx, s->c.V, st);
  XRT_CUDA_OK(cudaEventRecord(s->ev1, st));
  s->target_pending = false;
  s->launches[0] = s->launches[1] = launches;
  if (int rc = read_counters(s)) return rc;
  XRT_CUDA_OK(cudaGetLastError());
  if (solver_ms) XRT_CUDA_OK(cudaEventElapsedTime(solver_ms, s->ev0, s->ev1));
  if (prof) {
    float a = 0, b = 0, c = 0, t = 0;
    cudaEventElapsedTime(&a, s->ev0, pe[0]);
    cudaEventElapsedTime(&b, pe[0], pe[1]);
    cudaEventElapsedTime(&c, pe[1], pe[2]);
    cudaEventElapsedTime(&t, s->ev0, s->ev1);
    fprintf(stderr, "[xrt quality] frame %.2f ms: setup %.2f | %d iterations %.2f | CollisionStep %.2f (rounds %d, impacts %d, pairs %d, "
            "candidates %d, deform %d) | quads %u rep items %u\n", t, a, p.num_iterations, b, c, q->stats.rounds, q->stats.total_impacts,
            q->stats.n_pairs, q->stats.n_candidates, q->stats.deform, q->h_cnt[N_QUAD], q->h_cnt[N_REP]);
    for (auto& e : pe) cudaEventDestroy(e);
  }
  return XRT_OK;
}

extern "C" int xrt_quality_begin_frame(xrt_solver* s) {
  Q_READY();
  enqueue_begin_frame(s);
  XRT_CUDA_OK(cudaStreamSynchronize(s->st));
  return XRT_OK;
}

extern "C" int xrt_quality_imminent_generate(xrt_solver* s, uint32_t* nodes, float* bary, size_t capacity, size_t* n_quads) {
  Q_READY();
  XRT_REQUIRE(q->frame_begun, "imminent_generate: call xrt_quality_begin_frame (or step a frame) first");
  enqueue_imminent_generate(s);
  if (int rc = read_counters(s)) return rc;
  const size_t n = q->h_cnt[N_QUAD];
  if (n_quads) *n_quads = n;
  if (nodes && bary && n) {
    XRT_REQUIRE(capacity >= n, "imminent_generate: capacity too small");
    XRT_CUDA_OK(cudaMemcpy(nodes, q->quad_n, n * sizeof(uint4), cudaMemcpyDeviceToHost));
    XRT_CUDA_OK(cudaMemcpy(bary, q->quad_b, n * sizeof(float4), cudaMemcpyDeviceToHost));
  }
  return XRT_OK;
}

extern "C" int xrt_quality_imminent_solve(xrt_solver* s) {
  Q_READY();
  enqueue_imminent_solve(s);
  XRT_CUDA_OK(cudaStreamSynchronize(s->st));
  return XRT_OK;
}

extern "C" int xrt_quality_rep_update_pairs(xrt_solver* s, uint32_t* pairs, size_t capacity, size_t* n_pairs) {
  Q_READY();
  XRT_REQUIRE(q->frame_begun, "rep_update_pairs: call xrt_quality_begin_frame (or step a frame) first");
  enqueue_rep_update_pairs(s);
  if (int rc = read_counters(s)) return rc;
  const size_t n = q->h_cnt[N_PAIRS];
  if (n_pairs) *n_pairs = n;
  if (pairs && n) {
    XRT_REQUIRE(capacity >= n, "rep_update_pairs: capacity too small");
    XRT_CUDA_OK(cudaMemcpy(pairs, q->pairs, n * sizeof(uint2), cudaMemcpyDeviceToHost));
  }
  return XRT_OK;
}

extern "C" int xrt_quality_rep_solve(xrt_solver* s) {
  Q_READY();
  enqueue_rep_solve(s);
  XRT_CUDA_OK(cudaStreamSynchronize(s->st));
  return XRT_OK;
}

extern "C" int xrt_quality_obstacle_bvh_update(xrt_solver* s) {
  Q_READY();
  if (s->o.F) bvh_update(s->bvh, s->bw, s->o.x0, s->o.x, s->o.tri, 1, s->st);
  XRT_CUDA_OK(cudaStreamSynchronize(s->st));
  return XRT_OK;
}

extern "C" int xrt_quality_iteration(xrt_solver* s, int iteration) {
  Q_READY();
  XRT_REQUIRE(q->frame_begun, "quality_iteration: call xrt_quality_begin_frame (or step a frame) first");
  enqueue_iteration(s, iteration);
  if (int rc = read_counters(s)) return rc;
  return XRT_OK;
}

extern "C" int xrt_quality_find_impacts(xrt_solver* s, float thickness, uint32_t* nodes, float* twn, size_t capacity,
                                        size_t* n_impacts, size_t* n_pairs, size_t* n_candidates) {
  Q_READY();
  enqueue_collision_setup(s);
  enqueue_find_impacts(s, thickness);
  if (int rc = read_counters(s)) return rc;
  const size_t n = q->h_cnt[N_IMP];
  if (n_impacts) *n_impacts = n;
  if (n_pairs) *n_pairs = q->h_cnt[N_PAIRS];
  if (n_candidates) *n_candidates = q->h_cnt[N_CAND];
  if (nodes && twn && n) {
    XRT_REQUIRE(capacity >= n, "find_impacts: capacity too small");
    std::vector<DImpact> h(n);
    XRT_CUDA_OK(cudaMemcpy(h.data(), q->impacts, n * sizeof(DImpact), cudaMemcpyDeviceToHost));
    impacts_to_rows(h, n, nodes, twn);
  }
  return XRT_OK;
}

extern "C" int xrt_quality_independent_impacts(xrt_solver* s, const uint32_t* nodes, const float* twn, size_t n, int deform,
                                               uint32_t* out_nodes, float* out_twn, size_t* n_out) {
  Q_READY();
  XRT_REQUIRE(n_out && (n == 0 || (nodes && twn)), "independent_impacts: null argument");
  XRT_REQUIRE(n <= q->imp_cap, "independent_impacts: more impacts than max_impacts");
  *n_out = 0;
  if (n == 0) return XRT_OK;
  for (size_t i = 0; i < 4 * n; i++) XRT_REQUIRE(nodes[i] < s->c.V + s->o.V, "independent_impacts: node out of range");
  std::vector<DImpact> h;
  rows_to_impacts(nodes, twn, n, h);
  XRT_CUDA_OK(cudaMemcpy(q->impacts, h.data(), n * sizeof(DImpact), cudaMemcpyHostToDevice));
  if (int rc = enqueue_independent(s, q->impacts, (uint32_t)n, deform)) return rc;
  if (int rc = read_counters(s)) return rc;
  const size_t m = q->h_cnt[N_IND];
  *n_out = m;
  if (m && out_nodes && out_twn) {
    h.resize(m);
    XRT_CUDA_OK(cudaMemcpy(h.data(), q->imp_ind, m * sizeof(DImpact), cudaMemcpyDeviceToHost));
    impacts_to_rows(h, m, out_nodes, out_twn);
  }
  return XRT_OK;
}

extern "C" int xrt_quality_isolate_zones(xrt_solver* s, const uint32_t* nodes, size_t n, int32_t* colors, int32_t* n_zones) {
  Q_READY();
  XRT_REQUIRE(colors && n_zones && (n == 0 || nodes), "isolate_zones: null argument");
  XRT_REQUIRE(n <= q->imp_cap, "isolate_zones: more impacts than max_impacts");
  *n_zones = 0;
  if (n == 0) return XRT_OK;
  for (size_t i = 0; i < 4 * n; i++) XRT_REQUIRE(nodes[i] < s->c.V + s->o.V, "isolate_zones: node out of range");
  std::vector<DImpact> h;
  rows_to_impacts(nodes, nullptr, n, h);
  XRT_CUDA_OK(cudaMemcpy(q->imp_ind, h.data(), n * sizeof(DImpact), cudaMemcpyHostToDevice));
  enqueue_isolate(s, q->imp_ind, (uint32_t)n);
  if (int rc = read_counters(s)) return rc;
  XRT_CUDA_OK(cudaMemcpy(colors, q->imp_color, n * sizeof(int), cudaMemcpyDeviceToHost));
  *n_zones = (int32_t)q->h_cnt[N_ZONES];
  return XRT_OK;
}

extern "C" int xrt_quality_optimize_zones(xrt_solver* s, const uint32_t* nodes, const float* twn, const int32_t* colors, size_t n,
                                          int n_zones, int deform, float thickness, float obstacle_mass, int32_t* stats3) {
  Q_READY();
  XRT_REQUIRE(n == 0 || (nodes && twn && colors), "optimize_zones: null argument");
  for (size_t i = 0; i < 4 * n; i++) XRT_REQUIRE(nodes[i] < s->c.V + s->o.V, "optimize_zones: node out of range");
  for (size_t i = 0; i < n; i++)
    XRT_REQUIRE(colors[i] >= 0 && colors[i] < n_zones && (i == 0 || colors[i] >= colors[i - 1]), "optimize_zones: colours must ascend within [0, n_zones)");
  // Node::x1 = the positions the optimiser is asked to stay close to
  snapshot_kernel<<<xrt_grid(s->c.V), XRT_BLOCK, 0, s->st>>>(s->c.x, s->c.x0, q->cx1, nullptr, s->k.dt, s->c.V);
  if (s->o.V) snapshot_kernel<<<xrt_grid(s->o.V), XRT_BLOCK, 0, s->st>>>(s->o.x, s->o.x0, q->ox1, nullptr, s->k.dt, s->o.V);
  std::vector<DImpact> h;
  rows_to_impacts(nodes, twn, n, h);
  std::vector<int> col(colors, colors + n);
  return solve_zones(s, h, col, n_zones, deform, thickness, obstacle_mass, stats3);
}

extern "C" int xrt_quality_collision_step(xrt_solver* s, xrt_collision_stats* stats) {
  Q_READY();
  const int rc = collision_step(s);
  if (rc == XRT_OK) XRT_CUDA_OK(cudaStreamSynchronize(s->st));
  if (stats) *stats = q->stats;
  return rc;
}

// CUDA-event time of one stage of the Quality path, averaged over `reps` launches on the handle's stream
extern "C" int xrt_quality_profile(xrt_solver* s, int what, int reps, float* avg_ms, double* count) {
  Q_READY();
  XRT_REQUIRE(avg_ms && reps > 0 && what >= 0 && what <= 5, "quality_profile: bad argument");
  cudaStream_t st = s->st;
  const float th = q->qp.impact_zone_thickness;
  if (what >= 2) {  // the stages below work on the ccd trees
    enqueue_collision_setup(s);
    if (what >= 4) enqueue_pairs(s, q->ccd_c, s->o.F ? &q->ccd_o : nullptr, th);
  }
  const QMesh m = make_mesh(s);
  const int G = q->grid;
  XRT_CUDA_OK(cudaStreamSynchronize(st));
  XRT_CUDA_OK(cudaEventRecord(s->ev0, st));
  for (int r = 0; r < reps; r++) {
    switch (what) {
      case 0: bvh_build(q->cbvh, q->cbw, s->c.x0, s->c.tri, s->c.F, s->p.morton_fast_div, st); break;
      case 1: bvh_update(q->cbvh, q->cbw, s->c.x0, s->c.x, s->c.tri, 1, st); break;
      case 2: bvh_build_ccd(q->ccd_c, q->ccd_cw, s->c.x0, s->c.x, s->c.tri, s->c.F, s->p.morton_fast_div, st); break;
      case 3: enqueue_pairs(s, q->ccd_c, s->o.F ? &q->ccd_o : nullptr, th); break;
      case 4:  // the narrow phase on the pairs in the arena: filter, expand, cubic test, compaction
        ccd_filter_kernel<<<G, QB, 0, st>>>(m, q->pairs, q->cnt + N_PAIRS, th, q->masks, q->bsum);
        scan_sums_kernel<<<1, QG_MAX, 0, st>>>(q->bsum, q->boff, nullptr, nullptr, G, nullptr, q->cnt + N_CAND, q->cand_cap, q->cnt + OVERFLOW, OV_CAND);
        ccd_expand_kernel<<<G, QB, 0, st>>>(q->cnt + N_PAIRS, q->masks, q->boff, q->cand, q->cand_cap);
        ccd_test_kernel<<<G, QB, 0, st>>>(m, q->pairs, q->cand, q->cnt + N_CAND, q->cand_t, q->bsum);
        scan_sums_kernel<<<1, QG_MAX, 0, st>>>(q->bsum, q->boff, nullptr, nullptr, G, nullptr, q->cnt + N_IMP, q->imp_cap, q->cnt + OVERFLOW, OV_IMP);
        ccd_compact_kernel<<<G, QB, 0, st>>>(m, q->pairs, q->cand, q->cnt + N_CAND, q->cand_t, q->boff, q->impacts, q->imp_cap);
        break;
      default: enqueue_find_impacts(s, th); break;
    }
  }
  XRT_CUDA_OK(cudaEventRecord(s->ev1, st));
  if (int rc = read_counters(s)) return rc;
  float ms = 0.f;
  XRT_CUDA_OK(cudaEventElapsedTime(&ms, s->ev0, s->ev1));
  *avg_ms = ms / (float)reps;
  if (count) *count = what == 3 ? (double)q->h_cnt[N_PAIRS] : (what == 4 ? (double)q->h_cnt[N_CAND] : (what == 5 ? (double)q->h_cnt[N_IMP] : (double)s->c.F));
  return XRT_OK;
}

extern "C" int xrt_quality_cloth_rtri(xrt_solver* s, uint32_t* out, size_t n_faces) {
  Q_READY();
  XRT_REQUIRE(out && n_faces == s->c.F, "cloth_rtri: face count differs from the cloth");
  XRT_CUDA_OK(cudaMemcpy(out, q->crtri, n_faces * sizeof(uint32_t), cudaMemcpyDeviceToHost));
  return XRT_OK;
}

extern "C" int xrt_quality_get_obstacle_predicted(xrt_solver* s, float* xyz) {
  Q_READY();
  XRT_REQUIRE(xyz && s->o.V, "get_obstacle_predicted: no obstacle");
  XRT_CUDA_OK(cudaStreamSynchronize(s->st));
  std::vector<float4> h(s->o.V);
  XRT_CUDA_OK(cudaMemcpy(h.data(), s->o.x, h.size() * sizeof(float4), cudaMemcpyDeviceToHost));
  for (size_t i = 0; i < h.size(); i++) { xyz[3 * i] = h[i].x; xyz[3 * i + 1] = h[i].y; xyz[3 * i + 2] = h[i].z; }
  return XRT_OK;
}
