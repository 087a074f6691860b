// XPBD dynamics kernels of the Swift step on SoA float4 state.
// Per-colour kernels run over that colour's contiguous constraint range only
// (the reference launches over ALL constraints per colour and exits on
// colour mismatch, BasicSolverHelper.cu:12-14,86-88).
#include "xrt_solver.cuh"

namespace {

// PredictPositions_Kernel, ClothSolverHelper.cu:125-131
__global__ void predict_kernel(const float4* __restrict__ x0, float4* __restrict__ x, float4* __restrict__ v,
                               StepConst k, uint32_t V) {
  uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= V) return;
  float4 p = x0[i], vel = v[i];
  if (p.w > XRT_EPS) {
    vel.x += k.gx * k.dt;
    vel.y += k.gy * k.dt;
    vel.z += k.gz * k.dt;
    v[i] = vel;
  }
  x[i] = make_float4(p.x + vel.x * k.dt, p.y + vel.y * k.dt, p.z + vel.z * k.dt, p.w);
}

// Finalize_Kernel, ClothSolverHelper.cu:222-229
__global__ void finalize_kernel(float4* __restrict__ x0, const float4* __restrict__ x, float4* __restrict__ v,
                                StepConst k, uint32_t V) {
  uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= V) return;
  float4 p = x[i], q = x0[i];
  const float s = 1.0f - k.damping * k.dt;
  v[i] = make_float4((p.x - q.x) / k.dt * s, (p.y - q.y) / k.dt * s, (p.z - q.z) / k.dt * s, 0.f);
  x0[i] = p;
}

__device__ __forceinline__ void add_pos(float4* x, uint32_t i, f3 c, bool atomic) {
  if (atomic) {
    atomicAdd(&x[i].x, c.x);
    atomicAdd(&x[i].y, c.y);
    atomicAdd(&x[i].z, c.z);
  } else {
    float4 p = x[i];
    p.x += c.x; p.y += c.y; p.z += c.z;
    x[i] = p;
  }
}

// SolveStretch_Kernel, BasicSolverHelper.cu:80-127
__device__ __forceinline__ void stretch_one(float4* x, const uint2* s_idx, const float* s_rest, float* s_lambda,
                                            uint32_t id, float alpha_tilde, int iter, bool atomic) {
  const uint2 ij = s_idx[id];
  const float rest = s_rest[id];
  float lambda = iter == 0 ? 0.f : s_lambda[id];
  const float4 a = x[ij.x], b = x[ij.y];
  const f3 diff = mk3(a) - mk3(b);
  const float dist = len3(diff);
  const float w1 = a.w, w2 = b.w;
  if (dist != rest && w1 + w2 > 0.f) {
    const f3 grad = diff / (dist + XRT_EPS);
    const float C = dist - rest;
    const float dl = (-C - alpha_tilde * lambda) / (w1 + w2 + alpha_tilde);
    const f3 corr = dl * grad;
    lambda += dl;
    if (atomic) {
      add_pos(x, ij.x, corr * w1, true);
      add_pos(x, ij.y, (-corr) * w2, true);
    } else {
      x[ij.x] = make_float4(a.x + corr.x * w1, a.y + corr.y * w1, a.z + corr.z * w1, a.w);
      x[ij.y] = make_float4(b.x - corr.x * w2, b.y - corr.y * w2, b.z - corr.z * w2, b.w);
    }
  }
  s_lambda[id] = lambda;
}

// SolveBend_Kernel, BasicSolverHelper.cu:8-66
__device__ __forceinline__ void bend_one(float4* x, const uint4* b_idx, const float* b_theta, uint32_t id,
                                         float xpbd_bend, bool atomic) {
  const uint4 q = b_idx[id];
  const float theta0 = b_theta[id];
  const float4 a1 = x[q.x], a2 = x[q.y], a3 = x[q.z], a4 = x[q.w];
  const float w1 = a1.w, w2 = a2.w, w3 = a3.w, w4 = a4.w;
  const f3 p1 = mk3(a1);
  const f3 p2 = mk3(a2) - p1, p3 = mk3(a3) - p1, p4 = mk3(a4) - p1;
  const f3 c23 = cross3(p2, p3), c24 = cross3(p2, p4);
  const f3 n1 = normalize3(c23), n2 = normalize3(c24);
  const float dd = dot3(n1, n2);
  const float d = dd < 0.f ? 0.f : (dd > 1.f ? 1.f : dd);  // MathFunctions::Clamp
  const float angle = acosf(d);
  if (angle < XRT_EPS || isnan(d)) return;
  const float l23 = len3(c23) + XRT_EPS, l24 = len3(c24) + XRT_EPS;
  const f3 q3 = (cross3(p2, n2) + cross3(n1, p2) * d) / l23;
  const f3 q4 = (cross3(p2, n1) + cross3(n2, p2) * d) / l24;
  const f3 q2 = -((cross3(p3, n2) + cross3(n1, p3) * d) / l23) - (cross3(p4, n1) + cross3(n2, p4) * d) / l24;
  const f3 q1 = -q2 - q3 - q4;
  const float denom = xpbd_bend + (w1 * dot3(q1, q1) + w2 * dot3(q2, q2) + w3 * dot3(q3, q3) + w4 * dot3(q4, q4));
  if (denom < XRT_EPS) return;
  const float lambda = sqrtf(1.0f - d * d) * (angle - theta0) / denom;
  add_pos(x, q.x, (w1 * lambda) * q1, atomic);
  add_pos(x, q.y, (w2 * lambda) * q2, atomic);
  add_pos(x, q.z, (w3 * lambda) * q3, atomic);
  add_pos(x, q.w, (w4 * lambda) * q4, atomic);
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

// SolveGeodesicLRA_Kernel, LongRangeAttachments.cu:5-26 (the reference's
// non-atomic += on the target is kept: one slot per vertex is race free)
__global__ void lra_kernel(float4* x, const uint32_t* __restrict__ src, const uint32_t* __restrict__ tgt,
                           const float* __restrict__ rest, uint32_t n, float stretchiness) {
  uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const uint32_t s = src[i], t = tgt[i];
  const float target = rest[i] * stretchiness;
  float4 pt = x[t];
  if (pt.w < XRT_EPS && target > 0.f) return;
  const f3 diff = mk3(pt) - mk3(x[s]);
  const float dist = len3(diff);
  if (dist > target) {
    const f3 c = -diff + diff / dist * target;
    pt.x += c.x; pt.y += c.y; pt.z += c.z;
    x[t] = pt;
  }
}

// ApplyDeltas_Kernel, ClothSolverHelper.cu:139-148
__global__ void apply_deltas_kernel(float4* __restrict__ x, float4* __restrict__ delta, float relaxation,
                                    uint32_t V) {
  uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= V) return;
  const float4 d = delta[i];
  if (d.w > 0.f) {
    float4 p = x[i];
    p.x += d.x / d.w * relaxation;
    p.y += d.y / d.w * relaxation;
    p.z += d.z / d.w * relaxation;
    x[i] = p;
    delta[i] = make_float4(0.f, 0.f, 0.f, 0.f);
  }
}

// Vertex normals as a gather over incident faces (deterministic; the
// reference scatters with float atomicAdd, PhysicsMeshHelper.cu:200-214,282-310).
// vf_idx entries are (face << 2) | corner.
__global__ void normals_kernel(const float4* __restrict__ x, float4* __restrict__ n,
                               const uint32_t* __restrict__ tri, const uint32_t* __restrict__ vf_off,
                               const uint32_t* __restrict__ vf_idx, uint32_t V) {
  uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= V) return;
  const f3 p = mk3(x[i]);
  f3 acc = mk3(0.f, 0.f, 0.f);
  for (uint32_t k = vf_off[i]; k < vf_off[i + 1]; k++) {
    const uint32_t fc = vf_idx[k], f = fc >> 2, j = fc & 3u;
    const f3 e0 = mk3(x[tri[3 * f + (j + 1) % 3]]) - p;
    const f3 e1 = mk3(x[tri[3 * f + (j + 2) % 3]]) - p;
    acc = acc + cross3(e0, e1) / (dot3(e0, e0) * dot3(e1, e1));
  }
  const f3 r = normalize3(acc);
  n[i] = make_float4(r.x, r.y, r.z, 0.f);
}

// Interpolate_Kernel, PhysicsMesh.cu:610-614
__global__ void interpolate_kernel(const float4* __restrict__ x0, const float4* __restrict__ x,
                                   float4* __restrict__ x1, float alpha, uint32_t V) {
  uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= V) return;
  const float4 a = x0[i], b = x[i];
  const float om = 1.0f - alpha;
  x1[i] = make_float4(om * a.x + alpha * b.x, om * a.y + alpha * b.y, om * a.z + alpha * b.z, a.w);
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
