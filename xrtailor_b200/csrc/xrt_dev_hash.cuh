// Device body of the Swift self-contact solve, shared by the per-stage kernel
// (xrt_hash.cu) and the persistent kernel (xrt_persist.cu).
#pragma once
#include "xrt_solver.cuh"

namespace xrtdev {

// ComputeFriction + CollideParticles_Kernel, ClothSolverHelper.cu:150-211
template <class X>
__device__ __forceinline__ void collide_item(float4* delta, X x, const float4* x0,
                                             const uint32_t* neighbors, const StepConst& k, uint32_t V, uint32_t id) {
  const float4 pi4 = x[id];
  const f3 pi = mk3(pi4);
  const float wi = pi4.w;
  f3 acc = mk3(0, 0, 0);
  int count = 0;
  f3 vi = mk3(0, 0, 0);
  bool have_vi = false;
  uint32_t j = neighbors[id];
  for (int s = 0; s < k.max_neighbors && j <= V; s++) {
    // next slot's index is fetched while this neighbour is processed
    const uint32_t jn = s + 1 < k.max_neighbors ? neighbors[(size_t)(s + 1) * V + id] : 0xFFFFFFFFu;
    const float4 pj4 = x[j];
    const float denom = wi + pj4.w;
    const f3 pj = mk3(pj4);
    const f3 diff = pi - pj;
    const float dist = xf_len(diff);
    if (denom > 0 && dist < k.particle_diameter) {
      const f3 grad = diff * xf_rcp(dist + XRT_EPS);
      const float lambda = xf_div(dist - k.particle_diameter, denom);
      const f3 common = lambda * grad;
      count++;
      acc = acc - wi * common;
      if (!have_vi) {
        vi = pi - mk3(x0[id]);
        have_vi = true;
      }
      const f3 rel = vi - (pj - mk3(x0[j]));
      // ComputeFriction(common, rel)
      const float cl = xf_len(common);
      if (k.friction > 0.f && cl > 0.f) {
        const f3 nrm = common * xf_rcp(cl);
        const f3 tan = rel - nrm * xf_dot(rel, nrm);
        const float tl = xf_len(tan);
        const float r = xf_div(cl * k.friction, tl);
        const f3 fr = -tan * (r < 1.0f ? r : 1.0f);
        acc = acc + wi * fr;
      }
    }
    j = jn;
  }
  delta[id] = make_float4(acc.x, acc.y, acc.z, (float)count);
}

}  // namespace xrtdev
