// Persistent-cluster launch shape of one Swift substep (xrt_persist.cu).
#pragma once
#include "xrt_solver.cuh"

#define XRT_PERSIST_THREADS 1024
#define XRT_PERSIST_GRID 1000  // resolved launch shape "every SM of the device on one garment"
#define XRT_PERSIST_DSMEM 2000  // resolved launch shape "positions in the distributed shared memory of a 16-CTA cluster"
#define XRT_DSMEM_CLUSTER 16
#define XRT_DSMEM_MAX_SHIFT 13  // 8192 vertices x 16 B = 128 KB of shared memory per CTA -> garments up to 131 072 vertices

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
  // XRT_PERSIST_DSMEM only: the segment of the substep this launch runs (see swift_segment_kernel)
  int it_begin, stop_at, resume;
  uint32_t x_shift;  // vertices per CTA = 1 << x_shift
  // [n_colours][17]: where the constraints of colour c owned by CTA r start (owner = CTA of the constraint's lowest
  // vertex); the permuted constraint arrays are sorted by (colour, owner) for this shape
  const uint32_t *s_cta_off, *b_cta_off;
};

// largest cluster size this device can co-schedule (16 on B200 with the non-portable opt-in)
int persist_max_cluster();
// how many clusters of that size can be resident at once
int persist_active_clusters(int cluster);
// CTAs of the grid-wide shape (= SMs), 0 if the kernel cannot be co-resident
int persist_grid_ctas();
cudaError_t launch_swift_substep(const PersistArgs& a, int cluster, cudaStream_t st);
// XRT_PERSIST_DSMEM: one segment of a substep (a.it_begin, a.stop_at, a.resume) on a 16-CTA cluster that keeps the
// garment's predicted positions in distributed shared memory
cudaError_t launch_swift_segment(const PersistArgs& a, cudaStream_t st);
// smallest x_shift whose 16 slices hold V vertices, or -1 if the garment does not fit / the device has no 16-clusters
int persist_dsmem_shift(uint32_t V);
// shared launch over a batch: table = [S][2][n] argument blocks (middle index 1: obstacle target moved),
// moved = [n] flags; garment g runs substeps [sub_begin, sub_end) on CTA g (cluster <= 1) or on cluster g
cudaError_t launch_swift_substep_batch(const PersistArgs* table, const int* moved, uint32_t n, int sub_begin,
                                       int sub_end, int cluster, cudaStream_t st);
