// Persistent-cluster launch shape of one Swift substep (xrt_persist.cu).
#pragma once
#include "xrt_solver.cuh"

#define XRT_PERSIST_THREADS 1024

struct PersistArgs {
  ClothDev c;
  ObstacleDev o;
  DevBVH bvh;
  StepConst k;
  const uint32_t* s_off;  // device colour ranges over the permuted stretch / bend arrays
  const uint32_t* b_off;
  int* refit_flags;       // [F-1] arrival counters (parity mode: never cleared)
  float tol;              // bvh_tolerance
  int substep;            // index of this substep in the frame
  int n_iter, n_passes;   // num_iterations, num_collision_passes
  uint32_t frequency;     // collision pass every `frequency` (substep, iteration) steps
  int do_predict, do_finalize;
  int self_collision, solve_bending, lra, moved;
  unsigned long long* prof;  // optional [16] per-phase-kind ns accumulators (XRT_PERSIST_PROF=1)
  int atomic_s, atomic_b;  // colouring not proper -> float atomics in the projection
};

// largest cluster size this device can co-schedule (16 on B200 with the non-portable opt-in)
int persist_max_cluster();
// how many clusters of that size can be resident at once
int persist_active_clusters(int cluster);
cudaError_t launch_swift_substep(const PersistArgs& a, int cluster, cudaStream_t st);
