// Per-item device bodies of the XPBD dynamics stages.  One definition serves
// both launch shapes: the per-stage kernels (xrt_dynamics.cu, building blocks
// of the parity tests) and the persistent cluster kernel (xrt_persist.cu).
// State arrays are read and written inside one launch by the persistent
// kernel, so nothing here may use the non-coherent load path (no __ldg, no
// const __restrict__ on mutable state).
#pragma once
#include "xrt_solver.cuh"

namespace xrtdev {

// PredictPositions_Kernel, ClothSolverHelper.cu:125-131
__device__ __forceinline__ void predict_item(const float4* x0, float4* x, float4* v, const StepConst& k, uint32_t i) {
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
__device__ __forceinline__ void finalize_item(float4* x0, const float4* x, float4* v, const StepConst& k, uint32_t i) {
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

// SolveStretch_Kernel, BasicSolverHelper.cu:80-127.  The record (ij, rest) is
// passed in so a caller can fetch it ahead of a barrier.
__device__ __forceinline__ void stretch_rec(float4* x, float* s_lambda, uint32_t id, uint2 ij, float rest,
                                            float alpha_tilde, int iter, bool atomic) {
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
__device__ __forceinline__ void stretch_one(float4* x, const uint2* s_idx, const float* s_rest, float* s_lambda,
                                            uint32_t id, float alpha_tilde, int iter, bool atomic) {
  stretch_rec(x, s_lambda, id, s_idx[id], s_rest[id], alpha_tilde, iter, atomic);
}

// SolveBend_Kernel, BasicSolverHelper.cu:8-66
__device__ __forceinline__ void bend_rec(float4* x, uint4 q, float theta0, float xpbd_bend, bool atomic) {
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
__device__ __forceinline__ void bend_one(float4* x, const uint4* b_idx, const float* b_theta, uint32_t id,
                                         float xpbd_bend, bool atomic) {
  bend_rec(x, b_idx[id], b_theta[id], xpbd_bend, atomic);
}

// SolveGeodesicLRA_Kernel, LongRangeAttachments.cu:5-26 (the reference's
// non-atomic += on the target is kept: one slot per vertex is race free)
__device__ __forceinline__ void lra_item(float4* x, const uint32_t* src, const uint32_t* tgt, const float* rest,
                                         float stretchiness, uint32_t i) {
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
__device__ __forceinline__ void apply_delta_item(float4* x, float4* delta, float relaxation, uint32_t i) {
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
__device__ __forceinline__ void normals_item(const float4* x, float4* n, const uint32_t* tri, const uint32_t* vf_off,
                                             const uint32_t* vf_idx, uint32_t i) {
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
__device__ __forceinline__ void interpolate_item(const float4* x0, const float4* x, float4* x1, float alpha,
                                                 uint32_t i) {
  const float4 a = x0[i], b = x[i];
  const float om = 1.0f - alpha;
  x1[i] = make_float4(om * a.x + alpha * b.x, om * a.y + alpha * b.y, om * a.z + alpha * b.z, a.w);
}

}  // namespace xrtdev
