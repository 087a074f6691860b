// Device body of the Swift self-contact solve, shared by the per-stage kernel
// (xrt_hash.cu) and the persistent kernel (xrt_persist.cu).
#pragma once
#include "xrt_solver.cuh"

namespace xrtdev {

// ComputeFriction + CollideParticles_Kernel, ClothSolverHelper.cu:150-211
__device__ __forceinline__ void collide_item(float4* delta, const float4* x, const float4* x0,
                                             const uint32_t* neighbors, const StepConst& k, uint32_t V, uint32_t id) {
  const float4 pi4 = x[id];
  const f3 pi = mk3(pi4);
  const f3 vi = pi - mk3(x0[id]);
  const float wi = pi4.w;
  f3 acc = mk3(0, 0, 0);
  int count = 0;
  for (int s = 0; s < k.max_neighbors; s++) {
    const uint32_t j = neighbors[(size_t)s * V + id];
    if (j > V) break;
    const float4 pj4 = x[j];
    const float denom = wi + pj4.w;
    if (denom <= 0) continue;
    const f3 pj = mk3(pj4);
    const f3 diff = pi - pj;
    const float dist = len3(diff);
    if (dist >= k.particle_diameter) continue;
    const f3 grad = diff / (dist + XRT_EPS);
    const float lambda = (dist - k.particle_diameter) / denom;
    const f3 common = lambda * grad;
    count++;
    acc = acc - wi * common;
    const f3 rel = vi - (pj - mk3(x0[j]));
    // ComputeFriction(common, rel)
    const float cl = len3(common);
    if (k.friction > 0.f && cl > 0.f) {
      const f3 nrm = common / cl;
      const f3 tan = rel - nrm * dot3(rel, nrm);
      const float tl = len3(tan);
      const float r = (cl * k.friction) / tl;
      const f3 fr = -tan * (r < 1.0f ? r : 1.0f);
      acc = acc + wi * fr;
    }
  }
  delta[id] = make_float4(acc.x, acc.y, acc.z, (float)count);
}

}  // namespace xrtdev
