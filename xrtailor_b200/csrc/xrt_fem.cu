// Quality-mode cloth model: orthotropic StVK strain per triangle and
// isometric (quadratic) bending per hinge, as coloured Gauss-Seidel passes.
// Restates InitStrain_Kernel / ComputeStrainMeta / ComputeStrainCorrection / SolveStrain_Kernel
// (xrtailor/src/physics/dynamics/FEMHelper.cu:10-212) and ComputeHessianEnergy /
// InitIsometricBending_Kernel / SolveIsometricBending_Kernel (:221-388), in the order
// ClothSolverBase::SimulateQuality runs them (ClothSolverBase.cpp:315-324): per iteration every
// strain colour, then every bending colour.  A colour touches only its own constraints (ranges over
// a colour-sorted id list; the reference sweeps all constraints per colour), and with a proper
// colouring the position update is a plain 16-byte read-modify-write instead of float atomics.
// IEEE arithmetic (the file is built with -fmad=false) in the reference's evaluation order, so the
// result matches the CPU oracle to rounding.
#include "../../include/xrt_b200.h"
#include "xrt_fem.cuh"

#include <algorithm>
#include <cstdlib>
#include <cstring>
#include <vector>

using Strain = FemStrain;
using IsoBend = FemIsoBend;
using Material = FemMaterial;

namespace {

__device__ __forceinline__ f3 ld3(const float4* x, uint32_t i) { return mk3(x[i]); }

// ComputeStrainMeta + InitStrain_Kernel, FEMHelper.cu:10-84
__global__ void init_strain_kernel(Strain* out, const float4* __restrict__ rest, const uint32_t* __restrict__ tri,
                                   uint32_t F) {
  const uint32_t id = blockIdx.x * blockDim.x + threadIdx.x;
  if (id >= F) return;
  Strain c;
  c.i0 = tri[3 * id]; c.i1 = tri[3 * id + 1]; c.i2 = tri[3 * id + 2];
  const f3 p0 = ld3(rest, c.i0), p1 = ld3(rest, c.i1), p2 = ld3(rest, c.i2);
  const f3 n0 = cross3(p1 - p0, p2 - p0);
  c.area = 0.5f * len3(n0);
  const f3 a1 = normalize3(p1 - p0);
  const f3 a2 = normalize3(cross3(n0, a1));
  const float u0 = dot3(p0, a2), v0 = dot3(p0, a1), u1 = dot3(p1, a2), v1 = dot3(p1, a1), u2 = dot3(p2, a2),
              v2 = dot3(p2, a1);
  float P[2][2];
  P[0][0] = u0 - u2; P[1][0] = v0 - v2; P[0][1] = u1 - u2; P[1][1] = v1 - v2;
  const float det = P[0][0] * P[1][1] - P[1][0] * P[0][1];
  c.inv[0][0] = c.inv[0][1] = c.inv[1][0] = c.inv[1][1] = 0.f;  // glm default when the element is degenerate
  if (fabsf(det) > XRT_EPS) {  // glm::inverse(mat2)
    const float ood = 1.0f / det;
    c.inv[0][0] = P[1][1] * ood; c.inv[0][1] = -P[0][1] * ood;
    c.inv[1][0] = -P[1][0] * ood; c.inv[1][1] = P[0][0] * ood;
  }
  out[id] = c;
}

__device__ __forceinline__ float cot_theta(f3 v, f3 w) { return dot3(v, w) / len3(cross3(v, w)); }

// ComputeHessianEnergy + InitIsometricBending_Kernel, FEMHelper.cu:221-305
__global__ void init_bend_kernel(IsoBend* out, const float4* __restrict__ rest, const uint32_t* __restrict__ idx,
                                 uint32_t B) {
  const uint32_t id = blockIdx.x * blockDim.x + threadIdx.x;
  if (id >= B) return;
  IsoBend c;
  for (int k = 0; k < 4; k++) c.i[k] = idx[4 * id + k];
  const f3 p0 = ld3(rest, c.i[0]), p1 = ld3(rest, c.i[1]), p2 = ld3(rest, c.i[2]), p3 = ld3(rest, c.i[3]);
  const f3 e0 = p1 - p0, e1 = p2 - p0, e2 = p3 - p0, e3 = p2 - p1, e4 = p3 - p1;
  const f3 me0 = mk3(-e0.x, -e0.y, -e0.z);
  const float c01 = cot_theta(e0, e1), c02 = cot_theta(e0, e2), c03 = cot_theta(me0, e3), c04 = cot_theta(me0, e4);
  const float A0 = 0.5f * len3(cross3(p1 - p0, p2 - p0)), A1 = 0.5f * len3(cross3(p0 - p1, p3 - p1));
  const float coeff = 3.0f / (A0 + A1);
  const float K[4] = {c03 + c04, c01 + c02, -c01 - c03, -c02 - c04};
  for (int i = 0; i < 4; i++)
    for (int j = 0; j < 4; j++) c.Q[j][i] = coeff * (K[i] * K[j]);
  out[id] = c;
}

__device__ __forceinline__ void add_x(float4* x, uint32_t i, f3 d, bool atomic) {
  if (atomic) {
    atomicAdd(&x[i].x, d.x); atomicAdd(&x[i].y, d.y); atomicAdd(&x[i].z, d.z);
  } else {
    float4 p = x[i];
    p.x += d.x; p.y += d.y; p.z += d.z;
    x[i] = p;
  }
}

// ComputeStrainCorrection + SolveStrain_Kernel, FEMHelper.cu:94-212; x.w = inverse mass
__global__ void solve_strain_kernel(float4* x, const Strain* __restrict__ cs, const uint32_t* __restrict__ ids,
                                    uint32_t begin, uint32_t end, Material m, bool atomic) {
  const uint32_t t = begin + blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= end) return;
  const Strain c = cs[ids[t]];
  const float4 q0 = x[c.i0], q1 = x[c.i1], q2 = x[c.i2];
  const f3 p0 = mk3(q0), p1 = mk3(q1), p2 = mk3(q2);
  const float w0 = q0.w, w1 = q1.w, w2 = q2.w;
  const float den = 1.0f - m.nu_xy * m.nu_yx;
  const float C00 = m.xx / den, C01 = m.xx * m.nu_yx / den, C11 = m.yy / den, C10 = m.yy * m.nu_xy / den, C22 = m.xy;
  const f3 p13 = p0 - p2, p23 = p1 - p2;
  const float a[3] = {p13.x, p13.y, p13.z}, b[3] = {p23.x, p23.y, p23.z};
  float Fm[3][2];
  for (int i = 0; i < 3; i++) {
    Fm[i][0] = a[i] * c.inv[0][0] + b[i] * c.inv[1][0];
    Fm[i][1] = a[i] * c.inv[0][1] + b[i] * c.inv[1][1];
  }
  const float e00 = 0.5f * (Fm[0][0] * Fm[0][0] + Fm[1][0] * Fm[1][0] + Fm[2][0] * Fm[2][0] - 1.0f);
  const float e11 = 0.5f * (Fm[0][1] * Fm[0][1] + Fm[1][1] * Fm[1][1] + Fm[2][1] * Fm[2][1] - 1.0f);
  const float e01 = 0.5f * (Fm[0][0] * Fm[0][1] + Fm[1][0] * Fm[1][1] + Fm[2][0] * Fm[2][1]);
  const float s00 = C00 * e00 + C01 * e11 + 0.f * e01;
  const float s11 = C10 * e00 + C11 * e11 + 0.f * e01;
  const float s01 = 0.f * e00 + 0.f * e11 + C22 * e01;
  float pk[3][2];  // stress * F (glm mat2 * mat3x2)
  for (int i = 0; i < 3; i++) {
    pk[i][0] = s00 * Fm[i][0] + s01 * Fm[i][1];
    pk[i][1] = s01 * Fm[i][0] + s11 * Fm[i][1];
  }
  float psi = 0.f;
  psi += e00 * s00; psi += e01 * s01; psi += e01 * s01; psi += e11 * s11;
  psi = 0.5f * psi;
  const float energy = c.area * psi;
  float g0[3], g1[3];  // H = transpose(inv) * pk * area; grad_c[0][j] = H[j][0], grad_c[1][j] = H[j][1]
  for (int i = 0; i < 3; i++) {
    g0[i] = (c.inv[0][0] * pk[i][0] + c.inv[0][1] * pk[i][1]) * c.area;
    g1[i] = (c.inv[1][0] * pk[i][0] + c.inv[1][1] * pk[i][1]) * c.area;
  }
  const f3 G0 = mk3(g0[0], g0[1], g0[2]), G1 = mk3(g1[0], g1[1], g1[2]);
  const f3 G2 = mk3(-G0.x - G1.x, -G0.y - G1.y, -G0.z - G1.z);
  float sum = w0 * dot3(G0, G0);
  sum += w1 * dot3(G1, G1);
  sum += w2 * dot3(G2, G2);
  if (!(fabsf(sum) > XRT_EPS)) return;
  const float s = energy / sum;
  if (w0 > XRT_EPS) add_x(x, c.i0, (-(s * w0)) * G0, atomic);
  if (w1 > XRT_EPS) add_x(x, c.i1, (-(s * w1)) * G1, atomic);
  if (w2 > XRT_EPS) add_x(x, c.i2, (-(s * w2)) * G2, atomic);
}

// SolveIsometricBending_Kernel, FEMHelper.cu:317-388
__global__ void solve_bend_kernel(float4* x, const IsoBend* __restrict__ cs, float* lambda,
                                  const uint32_t* __restrict__ ids, uint32_t begin, uint32_t end, float stiffness,
                                  float dt, int iter, bool atomic) {
  const uint32_t t = begin + blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= end) return;
  const uint32_t id = ids[t];
  if (iter == 0) lambda[id] = 0.f;
  const IsoBend c = cs[id];
  f3 p[4];
  float w[4];
  for (int k = 0; k < 4; k++) {
    const float4 q = x[c.i[k]];
    p[k] = mk3(q);
    w[k] = q.w;
  }
  float energy = 0.f;
  for (int k = 0; k < 4; k++)
    for (int j = 0; j < 4; j++) energy += c.Q[j][k] * dot3(p[k], p[j]);
  energy *= 0.5f;
  if (fabsf(energy) < XRT_EPS) return;
  f3 d[4] = {mk3(0, 0, 0), mk3(0, 0, 0), mk3(0, 0, 0), mk3(0, 0, 0)};
  for (int i = 0; i < 4; i++)
    for (int j = 0; j < 4; j++) d[j] = d[j] + c.Q[j][i] * p[i];
  const float n0 = dot3(d[0], d[0]), n1 = dot3(d[1], d[1]), n2 = dot3(d[2], d[2]), n3 = dot3(d[3], d[3]);
  const float division = (w[0] * n0 + w[1] * n1) + (w[2] * n2 + w[3] * n3);  // glm::dot(vec4, vec4)
  const float alpha = 1.0f / (stiffness * dt * dt);
  const float lam = lambda[id];
  const float dl = (-energy - alpha * lam) / (division + alpha);
  for (int k = 0; k < 4; k++) add_x(x, c.i[k], (dl * w[k]) * d[k], atomic);
  lambda[id] = lam + dl;
}

// colour-sorted id list + ranges; returns whether the colouring is proper (no vertex twice in a colour)
bool sort_by_color(const uint32_t* colors, const uint32_t* idx, int arity, uint32_t n, uint32_t n_verts,
                   std::vector<uint32_t>& ids, std::vector<uint32_t>& off) {
  uint32_t nc = 0;
  for (uint32_t i = 0; i < n; i++) nc = std::max(nc, colors[i] + 1);
  off.assign(nc + 1, 0);
  for (uint32_t i = 0; i < n; i++) off[colors[i] + 1]++;
  for (uint32_t c = 0; c < nc; c++) off[c + 1] += off[c];
  ids.resize(n);
  std::vector<uint32_t> cur(off.begin(), off.end() - 1);
  for (uint32_t i = 0; i < n; i++) ids[cur[colors[i]]++] = i;
  std::vector<uint32_t> seen(n_verts, 0xFFFFFFFFu);
  bool proper = true;
  for (uint32_t c = 0; c < nc && proper; c++)
    for (uint32_t t = off[c]; t < off[c + 1] && proper; t++)
      for (int k = 0; k < arity; k++) {
        const uint32_t v = idx[(size_t)arity * ids[t] + k];
        if (v >= n_verts || seen[v] == c) proper = false;
        else seen[v] = c;
      }
  return proper;
}

template <class T>
cudaError_t up(T** d, const T* h, size_t n) {
  cudaError_t e = cudaMalloc((void**)d, (n ? n : 1) * sizeof(T));
  if (e == cudaSuccess && n) e = cudaMemcpy(*d, h, n * sizeof(T), cudaMemcpyHostToDevice);
  return e;
}

}  // namespace

int fem_setup(FemDev* f, const float4* rest, uint32_t n_verts, const uint32_t* d_tri, const uint32_t* h_tri,
              uint32_t n_faces, const uint32_t* h_bend_idx, uint32_t n_bend, const uint32_t* colors_strain,
              const uint32_t* colors_bend, const FemMaterial& m, cudaStream_t st) {
  std::vector<uint32_t> ids_s, off_s, ids_b, off_b;
  f->atomic_s = !sort_by_color(colors_strain, h_tri, 3, n_faces, n_verts, ids_s, off_s);
  f->atomic_b = n_bend ? !sort_by_color(colors_bend, h_bend_idx, 4, n_bend, n_verts, ids_b, off_b) : false;
  f->n_strain = n_faces;
  f->n_bend = n_bend;
  f->m = m;
  f->nc_s = (uint32_t)off_s.size() - 1;
  f->nc_b = n_bend ? (uint32_t)off_b.size() - 1 : 0;
  f->off_s = (uint32_t*)malloc(sizeof(uint32_t) * off_s.size());
  memcpy(f->off_s, off_s.data(), sizeof(uint32_t) * off_s.size());
  if (n_bend) {
    f->off_b = (uint32_t*)malloc(sizeof(uint32_t) * off_b.size());
    memcpy(f->off_b, off_b.data(), sizeof(uint32_t) * off_b.size());
  }
  cudaError_t e = up(&f->ids_s, ids_s.data(), ids_s.size());
  if (e == cudaSuccess) e = up(&f->ids_b, ids_b.data(), ids_b.size());
  if (e == cudaSuccess) e = up(&f->bend_idx, h_bend_idx, 4 * (size_t)n_bend);
  if (e == cudaSuccess) e = cudaMalloc((void**)&f->strain, (size_t)(n_faces ? n_faces : 1) * sizeof(FemStrain));
  if (e == cudaSuccess) e = cudaMalloc((void**)&f->bend, (size_t)(n_bend ? n_bend : 1) * sizeof(FemIsoBend));
  if (e == cudaSuccess) e = cudaMalloc((void**)&f->lambda, (size_t)(n_bend ? n_bend : 1) * sizeof(float));
  if (e == cudaSuccess) e = cudaMemsetAsync(f->lambda, 0, (size_t)(n_bend ? n_bend : 1) * sizeof(float), st);
  if (e != cudaSuccess) return (int)e;
  if (n_faces) init_strain_kernel<<<xrt_grid(n_faces), XRT_BLOCK, 0, st>>>(f->strain, rest, d_tri, n_faces);
  if (n_bend) init_bend_kernel<<<xrt_grid(n_bend), XRT_BLOCK, 0, st>>>(f->bend, rest, f->bend_idx, n_bend);
  return (int)cudaGetLastError();
}

void fem_free(FemDev* f) {
  cudaFree(f->strain); cudaFree(f->bend); cudaFree(f->lambda);
  cudaFree(f->ids_s); cudaFree(f->ids_b); cudaFree(f->bend_idx);
  free(f->off_s); free(f->off_b);
  *f = FemDev{};
}

void fem_reset_lambda(const FemDev& f, cudaStream_t st) {
  if (f.n_bend) cudaMemsetAsync(f.lambda, 0, (size_t)f.n_bend * sizeof(float), st);
}

int fem_iteration(const FemDev& f, float4* x, float dt, bool bending, cudaStream_t st) {
  int launches = 0;
  for (uint32_t c = 0; c < f.nc_s; c++)
    if (f.off_s[c + 1] > f.off_s[c]) {
      solve_strain_kernel<<<xrt_grid(f.off_s[c + 1] - f.off_s[c]), XRT_BLOCK, 0, st>>>(x, f.strain, f.ids_s, f.off_s[c],
                                                                                       f.off_s[c + 1], f.m, f.atomic_s);
      launches++;
    }
  if (bending)
    for (uint32_t c = 0; c < f.nc_b; c++)
      if (f.off_b[c + 1] > f.off_b[c]) {
        // iteration 1: lambda is reset by fem_reset_lambda at the start of the substep, not inside the kernel
        solve_bend_kernel<<<xrt_grid(f.off_b[c + 1] - f.off_b[c]), XRT_BLOCK, 0, st>>>(x, f.bend, f.lambda, f.ids_b, f.off_b[c],
                                                                                       f.off_b[c + 1], f.m.bend, dt, 1, f.atomic_b);
        launches++;
      }
  return launches;
}

extern "C" int xrt_fem_solve_standalone(const float* rest_pos, const float* pos, const float* inv_mass,
                                        uint32_t n_verts, const uint32_t* tri, uint32_t n_faces,
                                        const uint32_t* bend_idx, uint32_t n_bend, const float* material,
                                        const uint32_t* colors_strain, const uint32_t* colors_bend, int iterations,
                                        float dt, int device, float* out_pos, float* ms_per_iteration) {
  if (!rest_pos || !pos || !inv_mass || !n_verts || !tri || !n_faces || !material || !colors_strain || !out_pos ||
      iterations < 0 || (n_bend && (!bend_idx || !colors_bend)))
    return XRT_ERR_INVALID;
  for (uint32_t i = 0; i < 3 * n_faces; i++)
    if (tri[i] >= n_verts) return XRT_ERR_INVALID;
  for (uint32_t i = 0; i < 4 * n_bend; i++)
    if (bend_idx[i] >= n_verts) return XRT_ERR_INVALID;
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || device < 0 || device >= ndev) {
    (void)cudaGetLastError();
    return XRT_ERR_CUDA;  // no CPU fallback
  }
  if (cudaSetDevice(device) != cudaSuccess) return XRT_ERR_CUDA;
  std::vector<float4> hx(n_verts), hr(n_verts);
  for (uint32_t i = 0; i < n_verts; i++) {
    hx[i] = make_float4(pos[3 * i], pos[3 * i + 1], pos[3 * i + 2], inv_mass[i]);
    hr[i] = make_float4(rest_pos[3 * i], rest_pos[3 * i + 1], rest_pos[3 * i + 2], 0.f);
  }
  const FemMaterial m{material[0], material[1], material[2], material[3], material[4], material[5]};
  float4 *dx = nullptr, *dr = nullptr;
  uint32_t* dtri = nullptr;
  FemDev fem;
  cudaEvent_t e0 = nullptr, e1 = nullptr;
  cudaError_t e = up(&dx, hx.data(), hx.size());
  if (e == cudaSuccess) e = up(&dr, hr.data(), hr.size());
  if (e == cudaSuccess) e = up(&dtri, tri, 3 * (size_t)n_faces);
  if (e == cudaSuccess) e = (cudaError_t)fem_setup(&fem, dr, n_verts, dtri, tri, n_faces, bend_idx, n_bend, colors_strain, colors_bend, m, 0);
  if (e == cudaSuccess) e = cudaEventCreate(&e0);
  if (e == cudaSuccess) e = cudaEventCreate(&e1);
  if (e == cudaSuccess) {
    cudaEventRecord(e0, 0);
    for (int it = 0; it < iterations; it++) fem_iteration(fem, dx, dt, n_bend != 0, 0);
    cudaEventRecord(e1, 0);
    e = cudaDeviceSynchronize();
  }
  if (e == cudaSuccess) e = cudaMemcpy(hx.data(), dx, hx.size() * sizeof(float4), cudaMemcpyDeviceToHost);
  if (e == cudaSuccess) {
    for (uint32_t i = 0; i < n_verts; i++) {
      out_pos[3 * i] = hx[i].x; out_pos[3 * i + 1] = hx[i].y; out_pos[3 * i + 2] = hx[i].z;
    }
    if (ms_per_iteration) {
      float ms = 0;
      cudaEventElapsedTime(&ms, e0, e1);
      *ms_per_iteration = iterations ? ms / (float)iterations : 0.f;
    }
  }
  if (e0) cudaEventDestroy(e0);
  if (e1) cudaEventDestroy(e1);
  cudaFree(dx); cudaFree(dr); cudaFree(dtri);
  fem_free(&fem);
  (void)cudaGetLastError();
  return e == cudaSuccess ? XRT_OK : XRT_ERR_CUDA;
}
