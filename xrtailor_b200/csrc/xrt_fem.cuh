// Quality-mode cloth model on preallocated device state (xrt_fem.cu): orthotropic StVK strain per
// triangle + isometric bending per hinge as coloured Gauss-Seidel passes
// (xrtailor/src/physics/dynamics/FEMHelper.cu:10-388, FEMSolver.cu:28-83).
#pragma once
#include "xrt_common.cuh"

struct FemStrain {  // FEM::StrainConstraint: rest area + inverse rest matrix (glm column-major: m[c][r])
  uint32_t i0, i1, i2;
  float area;
  float inv[2][2];
};
struct FemIsoBend {  // FEM::IsometricBendingConstraint
  uint32_t i[4];
  float Q[4][4];
};
struct FemMaterial {
  float xx, yy, xy, nu_xy, nu_yx, bend;
};

struct FemDev {
  uint32_t n_strain = 0, n_bend = 0;
  FemStrain* strain = nullptr;
  FemIsoBend* bend = nullptr;
  float* lambda = nullptr;                 // [n_bend]
  uint32_t *ids_s = nullptr, *ids_b = nullptr;  // constraint ids sorted by colour
  uint32_t* bend_idx = nullptr;            // [4 n_bend] (setup only)
  uint32_t *off_s = nullptr, *off_b = nullptr;  // host arrays: colour ranges [n_colors + 1]
  uint32_t nc_s = 0, nc_b = 0;
  bool atomic_s = false, atomic_b = false;      // improper colouring: float atomics
  FemMaterial m{};
};

// rest: float4[V] rest positions on the device; tri: device uint32[3F]; bend_idx / colours: host arrays.
// Returns 0 or a cudaError_t.
int fem_setup(FemDev* f, const float4* rest, uint32_t n_verts, const uint32_t* d_tri, const uint32_t* h_tri,
              uint32_t n_faces, const uint32_t* h_bend_idx, uint32_t n_bend, const uint32_t* colors_strain,
              const uint32_t* colors_bend, const FemMaterial& m, cudaStream_t st);
void fem_free(FemDev* f);
// FEM::IsometricBendingSolver::Solve resets lambda when iteration == 0 (FEMHelper.cu:325); here once per substep
void fem_reset_lambda(const FemDev& f, cudaStream_t st);
// every strain colour, then (bending) every bending colour: ClothSolverBase.cpp:315-323.  Returns the launches.
int fem_iteration(const FemDev& f, float4* x, float dt, bool bending, cudaStream_t st);
