// Swift obstacle contact, per-stage launch shape (bodies in xrt_dev_contact.cuh).
#include "xrt_dev_contact.cuh"

namespace {
using namespace xrtdev;

__global__ void generate_vertex_contacts_kernel(const TNode* __restrict__ t, const float4* __restrict__ pred,
                                                const float4* __restrict__ x0, const float4* __restrict__ x,
                                                VFContact* __restrict__ vf, VFContact* __restrict__ rt,
                                                uint32_t* __restrict__ cnt_vf, uint32_t* __restrict__ cnt_rt,
                                                uint32_t* __restrict__ active, uint32_t V) {
  uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  bool live = false;
  if (i < V) live = gen_vertex_item<true>(t, pred, x0, x, vf, rt, cnt_vf, cnt_rt, i);
  store_active_bits(active, i, V, live);
}
__global__ void generate_edge_contacts_kernel(const TNode* __restrict__ t, const float4* __restrict__ pred,
                                              const uint32_t* __restrict__ rtri, const float4* __restrict__ x,
                                              const uint2* __restrict__ edge, EEContact* __restrict__ ee,
                                              uint32_t* __restrict__ cnt_ee, uint32_t* __restrict__ active,
                                              uint32_t E) {
  uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  bool live = false;
  if (i < E) live = gen_edge_item<true>(t, pred, rtri, x, edge, ee, cnt_ee, i);
  store_active_bits(active, i, E, live);
}
__global__ void solve_vertex_contacts_kernel(const float4* __restrict__ pred, float4* __restrict__ x,
                                             const VFContact* __restrict__ vf, const VFContact* __restrict__ rt,
                                             const uint32_t* __restrict__ active, uint32_t V) {
  uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < V && load_active_bit(active, i)) solve_vertex_item(pred, x, vf, rt, i);
}
__global__ void solve_edge_contacts_kernel(const float4* __restrict__ x, const uint2* __restrict__ edge,
                                           const EEContact* __restrict__ ee, const uint32_t* __restrict__ active,
                                           float4* delta, uint32_t E) {
  uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < E && load_active_bit(active, i)) solve_edge_item(x, edge, ee, delta, i);
}

// candidate lists in traversal order for the parity tests
__global__ void candidates_kernel(const TNode* __restrict__ t, const float4* __restrict__ x0,
                                  const float4* __restrict__ x, const uint2* __restrict__ edge, int kind,
                                  const uint32_t* __restrict__ offsets, uint32_t* __restrict__ items,
                                  uint32_t* __restrict__ counts, uint32_t n) {
  uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  uint32_t* out = items ? items + offsets[i] : nullptr;
  uint32_t k = 0;
  auto visit = [&](uint32_t o) {
    if (out) out[k] = o;
    k++;
  };
  uint32_t nf = 0;
  if (kind == 0) {
    const f3 p = mk3(x[i]);
    nf = traverse<true>(t, [&](const Box& b) { return overlap_point(b, p); }, visit);
  } else if (kind == 1) {
    const f3 p1 = mk3(x0[i]), p2 = mk3(x[i]);
    if (!(len3(p2 - p1) < XRT_EPS)) nf = traverse<true>(t, [&](const Box& b) { return overlap_segment(b, p1, p2); }, visit);
  } else {
    const uint2 e = edge[i];
    const f3 p1 = mk3(x[e.x]), p2 = mk3(x[e.y]);
    nf = traverse<true>(t, [&](const Box& b) { return overlap_segment(b, p1, p2); }, visit);
  }
  if (counts) counts[i] = nf;
}

}  // namespace

void launch_generate_contacts(const ClothDev& c, const DevBVH& bvh, const uint32_t* rtri, bool counts,
                              cudaStream_t st) {
  if (!bvh.n_obj) return;
  generate_vertex_contacts_kernel<<<xrt_grid(c.V), XRT_BLOCK, 0, st>>>(
      bvh.tnodes, bvh.pred, c.x0, c.x, c.vf, c.rt, counts ? c.cnt_vf : nullptr, counts ? c.cnt_rt : nullptr,
      c.v_active, c.V);
  if (c.E)
    generate_edge_contacts_kernel<<<xrt_grid(c.E), XRT_BLOCK, 0, st>>>(
        bvh.tnodes, bvh.pred, rtri, c.x, c.edge, c.ee, counts ? c.cnt_ee : nullptr, c.e_active, c.E);
}

void launch_generate_vertex_contacts(const ClothDev& c, const DevBVH& bvh, cudaStream_t st) {
  if (!bvh.n_obj) return;
  generate_vertex_contacts_kernel<<<xrt_grid(c.V), XRT_BLOCK, 0, st>>>(bvh.tnodes, bvh.pred, c.x0, c.x, c.vf, c.rt, nullptr, nullptr,
                                                                        c.v_active, c.V);
}
void launch_generate_edge_contacts(const ClothDev& c, const DevBVH& bvh, const uint32_t* rtri, cudaStream_t st) {
  if (!bvh.n_obj || !c.E) return;
  generate_edge_contacts_kernel<<<xrt_grid(c.E), XRT_BLOCK, 0, st>>>(bvh.tnodes, bvh.pred, rtri, c.x, c.edge, c.ee, nullptr, c.e_active,
                                                                      c.E);
}

void launch_solve_contacts(const ClothDev& c, const DevBVH& bvh, const StepConst& k, cudaStream_t st) {
  if (!bvh.n_obj) return;
  solve_vertex_contacts_kernel<<<xrt_grid(c.V), XRT_BLOCK, 0, st>>>(bvh.pred, c.x, c.vf, c.rt, c.v_active, c.V);
  if (c.E)
    solve_edge_contacts_kernel<<<xrt_grid(c.E), XRT_BLOCK, 0, st>>>(c.x, c.edge, c.ee, c.e_active, c.delta, c.E);
}

void launch_candidates(const ClothDev& c, const DevBVH& bvh, int kind, const uint32_t* offsets, uint32_t* items,
                       uint32_t* counts, cudaStream_t st) {
  const uint32_t n = kind == 2 ? c.E : c.V;
  if (!n || !bvh.n_obj) return;
  candidates_kernel<<<xrt_grid(n), XRT_BLOCK, 0, st>>>(bvh.tnodes, c.x0, c.x, c.edge, kind, offsets, items, counts, n);
}
