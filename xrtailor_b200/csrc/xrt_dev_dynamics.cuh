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
template <class X>
__device__ __forceinline__ void predict_item(const float4* x0, X x, float4* v, const StepConst& k, uint32_t i) {
  float4 p = x0[i], vel = v[i];
  if (p.w > XRT_EPS) {
    vel.x += k.gx * k.dt;
    vel.y += k.gy * k.dt;
    vel.z += k.gz * k.dt;
    v[i] = vel;
  }
  x[i] = make_float4(fmaf(vel.x, k.dt, p.x), fmaf(vel.y, k.dt, p.y), fmaf(vel.z, k.dt, p.z), p.w);
}

// Finalize_Kernel, ClothSolverHelper.cu:222-229
template <class X>
__device__ __forceinline__ void finalize_item(float4* x0, X x, float4* v, const StepConst& k, uint32_t i) {
  float4 p = x[i], q = x0[i];
  const float s = xf_rcp(k.dt) * (1.0f - k.damping * k.dt);
  v[i] = make_float4((p.x - q.x) * s, (p.y - q.y) * s, (p.z - q.z) * s, 0.f);
  x0[i] = p;
}

// float atomics on a position (improper colouring only); the distributed-shared-memory accessor has no such path
__device__ __forceinline__ void pos_atomic_add(float4* x, uint32_t i, f3 c) {
  atomicAdd(&x[i].x, c.x);
  atomicAdd(&x[i].y, c.y);
  atomicAdd(&x[i].z, c.z);
}
template <class X>
__device__ __forceinline__ void add_pos(X x, uint32_t i, f3 c, bool atomic) {
  if (atomic) {
    pos_atomic_add(x, i, c);
  } else {
    float4 p = x[i];
    p.x += c.x; p.y += c.y; p.z += c.z;
    x[i] = p;
  }
}

// SolveStretch_Kernel, BasicSolverHelper.cu:80-127.  The record (ij, rest) is
// passed in so a caller can fetch it ahead of a barrier.
// `lambda` is the multiplier carried from the previous iteration (0 at iteration 0).
// Projection of one stretch constraint on positions a, b that the caller has already gathered (proper colouring:
// this thread owns both vertices until the phase barrier), so a caller can issue the gathers of its next item
// before the arithmetic of this one.
template <class X>
__device__ __forceinline__ void stretch_apply(X x, float* s_lambda, uint32_t id, uint2 ij, float rest, float lambda,
                                              float alpha_tilde, float4 a, float4 b) {
  const f3 diff = mk3(a) - mk3(b);
  const float dist = xf_len(diff);
  const float w1 = a.w, w2 = b.w;
  if (dist != rest && w1 + w2 > 0.f) {
    const f3 grad = diff * xf_rcp(dist + XRT_EPS);
    const float C = dist - rest;
    const float dl = xf_div(-C - alpha_tilde * lambda, w1 + w2 + alpha_tilde);
    const f3 corr = dl * grad;
    lambda += dl;
    x[ij.x] = make_float4(fmaf(corr.x, w1, a.x), fmaf(corr.y, w1, a.y), fmaf(corr.z, w1, a.z), a.w);
    x[ij.y] = make_float4(fmaf(-corr.x, w2, b.x), fmaf(-corr.y, w2, b.y), fmaf(-corr.z, w2, b.z), b.w);
  }
  s_lambda[id] = lambda;
}
template <class X>
__device__ __forceinline__ void stretch_rec(X x, float* s_lambda, uint32_t id, uint2 ij, float rest,
                                            float lambda, float alpha_tilde, bool atomic) {
  const float4 a = x[ij.x], b = x[ij.y];
  if (!atomic) {
    stretch_apply(x, s_lambda, id, ij, rest, lambda, alpha_tilde, a, b);
    return;
  }
  const f3 diff = mk3(a) - mk3(b);
  const float dist = xf_len(diff);
  const float w1 = a.w, w2 = b.w;
  if (dist != rest && w1 + w2 > 0.f) {
    const f3 grad = diff * xf_rcp(dist + XRT_EPS);
    const float C = dist - rest;
    const float dl = xf_div(-C - alpha_tilde * lambda, w1 + w2 + alpha_tilde);
    const f3 corr = dl * grad;
    lambda += dl;
    add_pos(x, ij.x, corr * w1, true);
    add_pos(x, ij.y, (-corr) * w2, true);
  }
  s_lambda[id] = lambda;
}
template <class X>
__device__ __forceinline__ void stretch_one(X x, const uint2* s_idx, const float* s_rest, float* s_lambda,
                                            uint32_t id, float alpha_tilde, int iter, bool atomic) {
  stretch_rec(x, s_lambda, id, s_idx[id], s_rest[id], iter == 0 ? 0.f : s_lambda[id], alpha_tilde, atomic);
}

// SolveBend_Kernel, BasicSolverHelper.cu:8-66.  Without atomics the four
// positions gathered at the top are the ones updated at the bottom (a proper
// colouring gives this thread exclusive ownership), so x is read once.
template <class X, bool ATOMIC>
__device__ __forceinline__ void bend_project(X x, uint4 q, float theta0, float xpbd_bend, float4 a1, float4 a2, float4 a3,
                                             float4 a4) {
  const float w1 = a1.w, w2 = a2.w, w3 = a3.w, w4 = a4.w;
  const f3 p1 = mk3(a1);
  const f3 p2 = mk3(a2) - p1, p3 = mk3(a3) - p1, p4 = mk3(a4) - p1;
  const f3 c23 = xf_cross(p2, p3), c24 = xf_cross(p2, p4);
  const float s23 = xf_dot(c23, c23), s24 = xf_dot(c24, c24);
  const f3 n1 = c23 * xf_rsqrt(s23), n2 = c24 * xf_rsqrt(s24);
  const float dd = xf_dot(n1, n2);
  const float d = dd < 0.f ? 0.f : (dd > 1.f ? 1.f : dd);  // MathFunctions::Clamp
  const float angle = acosf(d);
  if (angle < XRT_EPS || isnan(d)) return;
  const float i23 = xf_rcp(xf_sqrt(s23) + XRT_EPS), i24 = xf_rcp(xf_sqrt(s24) + XRT_EPS);
  const f3 q3 = xf_axpy(d, xf_cross(n1, p2), xf_cross(p2, n2)) * i23;
  const f3 q4 = xf_axpy(d, xf_cross(n2, p2), xf_cross(p2, n1)) * i24;
  const f3 q2 = -(xf_axpy(d, xf_cross(n1, p3), xf_cross(p3, n2)) * i23) -
                xf_axpy(d, xf_cross(n2, p4), xf_cross(p4, n1)) * i24;
  const f3 q1 = -q2 - q3 - q4;
  const float denom = xpbd_bend + fmaf(w1, xf_dot(q1, q1), fmaf(w2, xf_dot(q2, q2), fmaf(w3, xf_dot(q3, q3), w4 * xf_dot(q4, q4))));
  if (denom < XRT_EPS) return;
  const float lambda = xf_div(xf_sqrt(fmaf(-d, d, 1.0f)) * (angle - theta0), denom);
  if (ATOMIC) {
    add_pos(x, q.x, (w1 * lambda) * q1, true);
    add_pos(x, q.y, (w2 * lambda) * q2, true);
    add_pos(x, q.z, (w3 * lambda) * q3, true);
    add_pos(x, q.w, (w4 * lambda) * q4, true);
  } else {
    const f3 r1 = xf_axpy(w1 * lambda, q1, p1), r2 = xf_axpy(w2 * lambda, q2, mk3(a2));
    const f3 r3 = xf_axpy(w3 * lambda, q3, mk3(a3)), r4 = xf_axpy(w4 * lambda, q4, mk3(a4));
    x[q.x] = make_float4(r1.x, r1.y, r1.z, w1);
    x[q.y] = make_float4(r2.x, r2.y, r2.z, w2);
    x[q.z] = make_float4(r3.x, r3.y, r3.z, w3);
    x[q.w] = make_float4(r4.x, r4.y, r4.z, w4);
  }
}
template <class X>
__device__ __forceinline__ void bend_rec(X x, uint4 q, float theta0, float xpbd_bend, bool atomic) {
  const float4 a1 = x[q.x], a2 = x[q.y], a3 = x[q.z], a4 = x[q.w];
  if (atomic)
    bend_project<X, true>(x, q, theta0, xpbd_bend, a1, a2, a3, a4);
  else
    bend_project<X, false>(x, q, theta0, xpbd_bend, a1, a2, a3, a4);
}
template <class X>
__device__ __forceinline__ void bend_one(X x, const uint4* b_idx, const float* b_theta, uint32_t id,
                                         float xpbd_bend, bool atomic) {
  bend_rec(x, b_idx[id], b_theta[id], xpbd_bend, atomic);
}

// SolveGeodesicLRA_Kernel, LongRangeAttachments.cu:5-26 (the reference's
// non-atomic += on the target is kept: one slot per vertex is race free)
template <class X>
__device__ __forceinline__ void lra_item(X x, const uint32_t* src, const uint32_t* tgt, const float* rest,
                                         float stretchiness, uint32_t i) {
  const uint32_t s = src[i], t = tgt[i];
  const float target = rest[i] * stretchiness;
  float4 pt = x[t];
  if (pt.w < XRT_EPS && target > 0.f) return;
  const f3 diff = mk3(pt) - mk3(x[s]);
  const float dist = xf_len(diff);
  if (dist > target) {
    const f3 c = -diff + diff * (xf_rcp(dist) * target);
    pt.x += c.x; pt.y += c.y; pt.z += c.z;
    x[t] = pt;
  }
}

// SolveBinding_Kernel, BindingConstraint.cu:56-92: one-sided tether of a cloth
// vertex to a barycentric point of an obstacle triangle at its frame-end
// position (Node::x); the kernel's `s = C / inv_mass ... * inv_mass` is kept as written.
template <class X>
__device__ __forceinline__ void binding_item(X x, const float4* ox, const uint4* bd_idx, const float4* bd_par,
                                             const float* bd_k, uint32_t i) {
  const uint4 id = bd_idx[i];
  const float4 par = bd_par[i];
  float4 pc = x[id.x];
  const float inv_mass = pc.w;
  if (inv_mass < XRT_EPS) return;
  const f3 po = par.x * mk3(ox[id.y]) + par.y * mk3(ox[id.z]) + par.z * mk3(ox[id.w]);
  const f3 d = mk3(pc) - po;
  const float dl = xf_len(d);
  const float C = dl - par.w;
  if (C < 0.f) return;
  const f3 n = d * xf_rcp(dl);
  const float s = xf_div(C, inv_mass) * bd_k[i];
  pc.x = fmaf(-s * inv_mass, n.x, pc.x);
  pc.y = fmaf(-s * inv_mass, n.y, pc.y);
  pc.z = fmaf(-s * inv_mass, n.z, pc.z);
  x[id.x] = pc;
}

// ApplyDeltas_Kernel, ClothSolverHelper.cu:139-148
template <class X>
__device__ __forceinline__ void apply_delta_item(X x, float4* delta, float relaxation, uint32_t i) {
  const float4 d = delta[i];
  if (d.w > 0.f) {
    float4 p = x[i];
    const float sc = xf_rcp(d.w) * relaxation;
    p.x = fmaf(d.x, sc, p.x);
    p.y = fmaf(d.y, sc, p.y);
    p.z = fmaf(d.z, sc, p.z);
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
  // incident faces in batches of 4: the index -> corner -> position chains of a
  // batch are issued together (same summation order as one face at a time)
  const uint32_t kb = vf_off[i], ke = vf_off[i + 1];
  for (uint32_t k0 = kb; k0 < ke; k0 += 4) {
    uint32_t ia[4], ib[4];
#pragma unroll
    for (int u = 0; u < 4; u++) {
      const uint32_t fc = k0 + u < ke ? vf_idx[k0 + u] : 0xFFFFFFFFu;
      const uint32_t f = fc >> 2, j = fc & 3u;
      ia[u] = fc != 0xFFFFFFFFu ? tri[3 * f + (j + 1) % 3] : 0xFFFFFFFFu;
      ib[u] = fc != 0xFFFFFFFFu ? tri[3 * f + (j + 2) % 3] : 0xFFFFFFFFu;
    }
    float4 pa[4], pb[4];
#pragma unroll
    for (int u = 0; u < 4; u++) {
      pa[u] = ia[u] != 0xFFFFFFFFu ? x[ia[u]] : make_float4(0.f, 0.f, 0.f, 0.f);
      pb[u] = ib[u] != 0xFFFFFFFFu ? x[ib[u]] : make_float4(0.f, 0.f, 0.f, 0.f);
    }
#pragma unroll
    for (int u = 0; u < 4; u++)
      if (ia[u] != 0xFFFFFFFFu) {
        const f3 e0 = mk3(pa[u]) - p, e1 = mk3(pb[u]) - p;
        acc = acc + cross3(e0, e1) / (dot3(e0, e0) * dot3(e1, e1));
      }
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
