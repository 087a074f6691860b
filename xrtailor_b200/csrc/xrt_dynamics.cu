// XPBD dynamics kernels of the Swift step on SoA float4 state (per-stage launch
// shape; bodies in xrt_dev_dynamics.cuh are shared with the persistent kernel).
// Per-colour kernels run over that colour's contiguous constraint range only
// (the reference launches over ALL constraints per colour and exits on
// colour mismatch, BasicSolverHelper.cu:12-14,86-88).
#include "xrt_dev_dynamics.cuh"

namespace {
using namespace xrtdev;

__global__ void predict_kernel(const float4* x0, float4* x, float4* v, StepConst k, uint32_t V) {
  uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < V) predict_item(x0, x, v, k, i);
}
__global__ void finalize_kernel(float4* x0, const float4* x, float4* v, StepConst k, uint32_t V) {
  uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < V) finalize_item(x0, x, v, k, i);
}

template <bool ATOMIC>
__global__ void stretch_color_kernel(float4* x, const uint2* __restrict__ s_idx, const float* __restrict__ s_rest,
                                     float* __restrict__ s_lambda, uint32_t begin, uint32_t end,
                                     float alpha_tilde, int iter) {
  uint32_t id = begin + blockIdx.x * blockDim.x + threadIdx.x;
  if (id >= end) return;
  stretch_one(x, s_idx, s_rest, s_lambda, id, alpha_tilde, iter, ATOMIC);
}

template <bool ATOMIC>
__global__ void bend_color_kernel(float4* x, const uint4* __restrict__ b_idx, const float* __restrict__ b_theta,
                                  uint32_t begin, uint32_t end, float xpbd_bend) {
  uint32_t id = begin + blockIdx.x * blockDim.x + threadIdx.x;
  if (id >= end) return;
  bend_one(x, b_idx, b_theta, id, xpbd_bend, ATOMIC);
}

__global__ void lra_kernel(float4* x, const uint32_t* __restrict__ src, const uint32_t* __restrict__ tgt,
                           const float* __restrict__ rest, uint32_t n, float stretchiness) {
  uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) lra_item(x, src, tgt, rest, stretchiness, i);
}
__global__ void binding_kernel(float4* x, const float4* __restrict__ ox, const uint4* __restrict__ bd_idx,
                               const float4* __restrict__ bd_par, const float* __restrict__ bd_k, uint32_t n) {
  uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) binding_item(x, ox, bd_idx, bd_par, bd_k, i);
}
__global__ void apply_deltas_kernel(float4* x, float4* delta, float relaxation, uint32_t V) {
  uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < V) apply_delta_item(x, delta, relaxation, i);
}
__global__ void normals_kernel(const float4* __restrict__ x, float4* __restrict__ n,
                               const uint32_t* __restrict__ tri, const uint32_t* __restrict__ vf_off,
                               const uint32_t* __restrict__ vf_idx, uint32_t V) {
  uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < V) normals_item(x, n, tri, vf_off, vf_idx, i);
}
// end-of-frame work of a whole batch (blockIdx.y = mesh): PredictedToPositions, then vertex normals
__global__ void batch_copy_kernel(const NormalsBatchItem* __restrict__ items, const int* __restrict__ moved) {
  const NormalsBatchItem& it = items[blockIdx.y];
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (!it.copy_src || i >= it.V || (moved && !moved[it.garment])) return;
  it.x[i] = it.copy_src[i];
}
__global__ void batch_normals_kernel(const NormalsBatchItem* __restrict__ items) {
  const NormalsBatchItem& it = items[blockIdx.y];
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < it.V) normals_item(it.x, it.n, it.tri, it.vf_off, it.vf_idx, i);
}
__global__ void interpolate_kernel(const float4* __restrict__ x0, const float4* __restrict__ x,
                                   float4* __restrict__ x1, float alpha, uint32_t V) {
  uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < V) interpolate_item(x0, x, x1, alpha, i);
}

// host float[3n] <-> device float4[n] at the ABI boundary (dst.w = inverse mass is kept)
__global__ void unpack_xyz_kernel(const float* __restrict__ src, float4* __restrict__ dst, uint32_t n) {
  uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) dst[i] = make_float4(src[3 * i], src[3 * i + 1], src[3 * i + 2], dst[i].w);
}
__global__ void pack_xyz_kernel(const float4* __restrict__ a, const float4* __restrict__ b, float* __restrict__ out,
                                uint32_t n) {
  uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  if (a) {
    const float4 p = a[i];
    out[3 * i] = p.x; out[3 * i + 1] = p.y; out[3 * i + 2] = p.z;
  }
  if (b) {
    const float4 q = b[i];
    float* o2 = out + 3 * (size_t)n;
    o2[3 * i] = q.x; o2[3 * i + 1] = q.y; o2[3 * i + 2] = q.z;
  }
}

}  // namespace

void launch_predict(const ClothDev& c, const StepConst& k, cudaStream_t st) {
  predict_kernel<<<xrt_grid(c.V), XRT_BLOCK, 0, st>>>(c.x0, c.x, c.v, k, c.V);
}
void launch_finalize(const ClothDev& c, const StepConst& k, cudaStream_t st) {
  finalize_kernel<<<xrt_grid(c.V), XRT_BLOCK, 0, st>>>(c.x0, c.x, c.v, k, c.V);
}
void launch_stretch_color(const ClothDev& c, const StepConst& k, uint32_t begin, uint32_t end, int iter,
                          bool atomic, cudaStream_t st) {
  if (end <= begin) return;
  const float at = k.stretch_compliance / (k.dt * k.dt);
  if (atomic)
    stretch_color_kernel<true><<<xrt_grid(end - begin), XRT_BLOCK, 0, st>>>(c.x, c.s_idx, c.s_rest, c.s_lambda,
                                                                           begin, end, at, iter);
  else
    stretch_color_kernel<false><<<xrt_grid(end - begin), XRT_BLOCK, 0, st>>>(c.x, c.s_idx, c.s_rest, c.s_lambda,
                                                                            begin, end, at, iter);
}
void launch_bend_color(const ClothDev& c, const StepConst& k, uint32_t begin, uint32_t end, bool atomic,
                       cudaStream_t st) {
  if (end <= begin) return;
  const float xb = k.bend_compliance / k.dt / k.dt;
  if (atomic)
    bend_color_kernel<true><<<xrt_grid(end - begin), XRT_BLOCK, 0, st>>>(c.x, c.b_idx, c.b_theta, begin, end, xb);
  else
    bend_color_kernel<false><<<xrt_grid(end - begin), XRT_BLOCK, 0, st>>>(c.x, c.b_idx, c.b_theta, begin, end, xb);
}
void launch_lra(const ClothDev& c, const StepConst& k, cudaStream_t st) {
  if (!c.n_lra) return;
  lra_kernel<<<xrt_grid(c.n_lra), XRT_BLOCK, 0, st>>>(c.x, c.l_src, c.l_tgt, c.l_rest, c.n_lra, k.lra_stretchiness);
}
void launch_binding(const ClothDev& c, const ObstacleDev& o, cudaStream_t st) {
  if (c.n_bind && o.V)
    binding_kernel<<<xrt_grid(c.n_bind), XRT_BLOCK, 0, st>>>(c.x, o.x, c.bd_idx, c.bd_par, c.bd_k, c.n_bind);
}
void launch_apply_deltas(const ClothDev& c, const StepConst& k, cudaStream_t st) {
  apply_deltas_kernel<<<xrt_grid(c.V), XRT_BLOCK, 0, st>>>(c.x, c.delta, k.relaxation, c.V);
}
void launch_normals(const float4* x, float4* n, const uint32_t* tri, const uint32_t* vf_off,
                    const uint32_t* vf_idx, uint32_t V, cudaStream_t st) {
  if (!V) return;
  normals_kernel<<<xrt_grid(V), XRT_BLOCK, 0, st>>>(x, n, tri, vf_off, vf_idx, V);
}
void launch_interpolate(const ObstacleDev& o, int step, int num_steps, cudaStream_t st) {
  if (!o.V) return;
  const float alpha = ((float)step + 1.0f) / (float)num_steps;
  interpolate_kernel<<<xrt_grid(o.V), XRT_BLOCK, 0, st>>>(o.x0, o.x, o.x1, alpha, o.V);
}
void launch_unpack_xyz(const float* src, float4* dst, uint32_t n, cudaStream_t st) {
  if (n) unpack_xyz_kernel<<<xrt_grid(n), XRT_BLOCK, 0, st>>>(src, dst, n);
}
void launch_pack_xyz(const float4* a, const float4* b, float* out, uint32_t n, cudaStream_t st) {
  if (n) pack_xyz_kernel<<<xrt_grid(n), XRT_BLOCK, 0, st>>>(a, b, out, n);
}

void launch_post_batched(const NormalsBatchItem* d_items, uint32_t n_items, uint32_t max_v, const int* d_moved, cudaStream_t st) {
  if (!n_items || !max_v) return;
  const dim3 g(xrt_grid(max_v), n_items);
  batch_copy_kernel<<<g, XRT_BLOCK, 0, st>>>(d_items, d_moved);
  batch_normals_kernel<<<g, XRT_BLOCK, 0, st>>>(d_items);
}
