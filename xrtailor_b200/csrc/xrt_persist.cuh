// Persistent-cluster launch shape of one Swift substep (xrt_persist.cu).
#pragma once
#include "xrt_solver.cuh"

#define XRT_PERSIST_THREADS 1024
#define XRT_PERSIST_GRID 1000  // resolved launch shape "every SM of the device on one garment"

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
  unsigned long long* grid_bar;  // ticket counter of the grid-wide barrier (XRT_PERSIST_GRID shape)
  unsigned long long* prof;  // optional [16] per-phase-kind ns accumulators (XRT_PERSIST_PROF=1)
  int atomic_s, atomic_b;  // colouring not proper -> float atomics in the projection
};

// largest cluster size this device can co-schedule (16 on B200 with the non-portable opt-in)
int persist_max_cluster();
// how many clusters of that size can be resident at once
int persist_active_clusters(int cluster);
// CTAs of the grid-wide shape (= SMs), 0 if the kernel cannot be co-resident
int persist_grid_ctas();
cudaError_t launch_swift_substep(const PersistArgs& a, int cluster, cudaStream_t st);
// shared launch over a batch: table = [S][2][n] argument blocks (middle index 1: obstacle target moved),
// moved = [n] flags; garment g runs substeps [sub_begin, sub_end) on CTA g (cluster <= 1) or on cluster g
cudaError_t launch_swift_substep_batch(const PersistArgs* table, const int* moved, uint32_t n, int sub_begin,
                                       int sub_end, int cluster, cudaStream_t st);
