// One Swift substep as ONE persistent kernel on a thread-block cluster.
//
// The per-stage launch shape (xrt_api.cu: enqueue_frame with persist == 0)
// replays ~270 dependent 3-5 us kernels per substep; each touches ~1.7 MB that
// lives in L2, so the step is bound by launch + drain latency, not by HBM
// (profiles/r01b_launches_percolour_summary.csv).  Here a cluster of C CTAs
// (C x 1024 threads, one CTA per SM) owns one garment for the whole substep and
// separates the ~270 phases with the hardware cluster barrier
// (barrier.cluster, ~0.2 us) instead of a kernel boundary.  Independent
// garments run on other clusters at the same time (one stream per handle), so
// a B200 carries 148 / C garments concurrently.
//
// Phase order is exactly ClothSolverBase::SimulateSwift's inner loop
// (pipeline/base/ClothSolverBase.cpp:189-243); the bodies are the same device
// functions the per-stage kernels call, so results are bit-identical to that
// path whenever the colouring is proper (no float atomics in the projection).
#include "xrt_dev_bvh.cuh"
#include "xrt_dev_contact.cuh"
#include "xrt_dev_dynamics.cuh"
#include "xrt_dev_hash.cuh"
#include "xrt_persist.cuh"

namespace {
using namespace xrtdev;

constexpr int PT = XRT_PERSIST_THREADS;

__device__ __forceinline__ uint32_t cluster_ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ uint32_t cluster_nctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_nctarank;" : "=r"(r));
  return r;
}

// Barrier between two phases, by launch shape:
//  SYNC_BLOCK    one CTA: bar.sync (one L1, nothing to invalidate);
//  SYNC_CLUSTER  hardware cluster barrier; release/acquire at cluster scope orders
//                the global-memory writes of every CTA before the reads that follow
//                (ptxas emits MEMBAR.GPU before the arrive and CCTL.IVALL after the wait);
//  SYNC_GRID     every SM of the device on ONE garment (cooperative launch): bar.sync,
//                then thread 0 publishes with a gpu-scope fence, takes a ticket on a
//                monotonically increasing counter and spins until all CTAs of this
//                round have arrived; its acquiring fence invalidates the SM's L1.
enum { SYNC_BLOCK = 0, SYNC_CLUSTER = 1, SYNC_GRID = 2 };
template <int MODE>
__device__ __forceinline__ void phase_sync(unsigned long long* grid_bar = nullptr) {
  if (MODE == SYNC_CLUSTER) {
    asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
  } else if (MODE == SYNC_GRID) {
    __syncthreads();
    if (threadIdx.x == 0) {
      __threadfence();
      const unsigned long long n = gridDim.x;
      const unsigned long long ticket = atomicAdd(grid_bar, 1ull);
      const unsigned long long target = (ticket / n + 1ull) * n;
      unsigned long long seen;
      do {
        asm volatile("ld.acquire.gpu.global.u64 %0, [%1];" : "=l"(seen) : "l"(grid_bar) : "memory");
      } while (seen < target);
      __threadfence();
    }
    __syncthreads();
  } else {
    __syncthreads();
  }
}

__device__ __forceinline__ unsigned long long gtimer() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}
// optional per-phase-kind wall time (ns, barrier included), thread 0 of CTA 0 only
struct PhaseClock {
  unsigned long long* acc;
  unsigned long long t;
  __device__ __forceinline__ void start(unsigned long long* a, uint32_t tid) {
    acc = tid == 0 ? a : nullptr;
    if (acc) t = gtimer();
  }
  __device__ __forceinline__ void lap(int kind) {
    if (acc) {
      const unsigned long long n = gtimer();
      acc[kind] += n - t;
      t = n;
    }
  }
};

// first item of colour phase `ph` for this thread (stretch colours, then bend colours)
struct ColorRec {
  uint4 q;
  float f, lambda;
};
__device__ __forceinline__ ColorRec fetch_color_rec(const PersistArgs& a, uint32_t ph, uint32_t n_sc, uint32_t n_ph,
                                                    uint32_t tid, int it) {
  ColorRec r;
  r.q = make_uint4(0, 0, 0, 0);
  r.f = 0.f;
  r.lambda = 0.f;
  if (ph < n_sc) {
    const uint32_t i = a.s_off[ph] + tid;
    if (i < a.s_off[ph + 1]) {
      const uint2 ij = a.c.s_idx[i];
      r.q.x = ij.x;
      r.q.y = ij.y;
      r.f = a.c.s_rest[i];
      if (it != 0) r.lambda = a.c.s_lambda[i];
    }
  } else if (ph < n_ph) {
    const uint32_t col = ph - n_sc, i = a.b_off[col] + tid;
    if (i < a.b_off[col + 1]) {
      r.q = a.c.b_idx[i];
      r.f = a.c.b_theta[i];
    }
  }
  return r;
}

// the same for the distributed-shared-memory shape: first item of this thread in CTA `rank`'s share of the phase
__device__ __forceinline__ ColorRec fetch_color_rec_cta(const PersistArgs& a, uint32_t ph, uint32_t n_sc, uint32_t n_ph,
                                                        uint32_t rank, int it) {
  ColorRec r;
  r.q = make_uint4(0, 0, 0, 0);
  r.f = 0.f;
  r.lambda = 0.f;
  if (ph < n_sc) {
    const uint32_t* o = a.s_cta_off + ph * (XRT_DSMEM_CLUSTER + 1) + rank;
    const uint32_t i = o[0] + threadIdx.x;
    if (i < o[1]) {
      const uint2 ij = a.c.s_idx[i];
      r.q.x = ij.x;
      r.q.y = ij.y;
      r.f = a.c.s_rest[i];
      if (it != 0) r.lambda = a.c.s_lambda[i];
    }
  } else if (ph < n_ph) {
    const uint32_t* o = a.b_cta_off + (ph - n_sc) * (XRT_DSMEM_CLUSTER + 1) + rank;
    const uint32_t i = o[0] + threadIdx.x;
    if (i < o[1]) {
      r.q = a.c.b_idx[i];
      r.f = a.c.b_theta[i];
    }
  }
  return r;
}

// A colour phase of one thread as a software pipeline.  The items of a colour are independent (proper colouring: no
// vertex is shared inside the phase), so the record of item k+2 and the positions of item k+1 are in flight while item k
// is projected: per thread the chain record -> gather -> arithmetic -> scatter overlaps across items instead of paying
// two dependent L2 latencies per item.  Same arithmetic per item, so positions are bit-identical.  Stretch only (-5 % on
// the stretch phases at one SM per garment): the four gathered positions of a bend element leave no registers for a
// second item in flight (measured +55 % with spills), and running only its record ahead changes nothing.
__device__ __forceinline__ void stretch_pipelined(const ClothDev& c, uint32_t i, uint32_t end, uint32_t nth, const ColorRec& first,
                                                  float alpha_s, int it) {
  uint2 ij0 = make_uint2(first.q.x, first.q.y);
  float r0 = first.f, l0 = first.lambda;
  float4 a0 = c.x[ij0.x], b0 = c.x[ij0.y];
  uint32_t i1 = i + nth;
  bool h1 = i1 < end;
  uint2 ij1 = make_uint2(0, 0);
  float r1 = 0.f, l1 = 0.f;
  if (h1) {
    ij1 = c.s_idx[i1];
    r1 = c.s_rest[i1];
    if (it != 0) l1 = c.s_lambda[i1];
  }
  for (;;) {
    const uint32_t i2 = i1 + nth;
    const bool h2 = h1 && i2 < end;
    float4 a1 = a0, b1 = b0;
    uint2 ij2 = make_uint2(0, 0);
    float r2 = 0.f, l2 = 0.f;
    if (h1) {
      a1 = c.x[ij1.x];
      b1 = c.x[ij1.y];
    }
    if (h2) {
      ij2 = c.s_idx[i2];
      r2 = c.s_rest[i2];
      if (it != 0) l2 = c.s_lambda[i2];
    }
    stretch_apply(c.x, c.s_lambda, i, ij0, r0, l0, alpha_s, a0, b0);
    if (!h1) break;
    i = i1; ij0 = ij1; r0 = r1; l0 = l1; a0 = a1; b0 = b1;
    i1 = i2; h1 = h2; ij1 = ij2; r1 = r2; l1 = l2;
  }
}
template <int MODE>
__device__ __forceinline__ void swift_substep_body(const PersistArgs& a) {
  const uint32_t nth = (MODE == SYNC_CLUSTER ? cluster_nctarank() : (MODE == SYNC_GRID ? gridDim.x : 1u)) * PT;
  const uint32_t tid =
      (MODE == SYNC_CLUSTER ? cluster_ctarank() : (MODE == SYNC_GRID ? blockIdx.x : 0u)) * PT + threadIdx.x;
  const ClothDev& c = a.c;
  const ObstacleDev& o = a.o;
  const StepConst& k = a.k;
  const uint32_t V = c.V, E = c.E;
  const float alpha_s = k.stretch_compliance / (k.dt * k.dt);
  const float xpbd_b = k.bend_compliance / k.dt / k.dt;
  PhaseClock pc;
  pc.start(a.prof, tid);

  if (a.do_predict) {
    for (uint32_t i = tid; i < V; i += nth) predict_item(c.x0, c.x, c.v, k, i);
    phase_sync<MODE>(a.grid_bar);
    pc.lap(0);
  }
  for (int it = 0; it < a.n_iter; it++) {
    if (a.self_collision) {
      for (uint32_t i = tid; i < V; i += nth) collide_item(c.delta, c.x, c.x0, c.neighbors, k, V, i);
      phase_sync<MODE>(a.grid_bar);
    pc.lap(1);
      for (uint32_t i = tid; i < V; i += nth) apply_delta_item(c.x, c.delta, k.relaxation, i);
      phase_sync<MODE>(a.grid_bar);
    pc.lap(2);
    }
    const uint32_t step = (uint32_t)(a.substep * a.n_iter + it);
    if (o.F && step % a.frequency == 0) {
      if (a.moved) {
        const int cnt = (int)(step / a.frequency);
        const float alpha = ((float)cnt + 1.0f) / (float)a.n_passes;
        for (uint32_t i = tid; i < o.V; i += nth) interpolate_item(o.x0, o.x, o.x1, alpha, i);
        phase_sync<MODE>(a.grid_bar);
        for (uint32_t i = tid; i < o.V; i += nth) normals_item(o.x1, o.n1, o.tri, o.vf_off, o.vf_idx, i);
        phase_sync<MODE>(a.grid_bar);
    pc.lap(3);
        for (uint32_t i = tid; i < o.F; i += nth)
          refit_item<true, MODE == SYNC_BLOCK>(a.bvh.tnodes, a.bvh.nodes, a.bvh.pred, o.x1, o.n1, o.tri, a.tol, a.refit_flags, o.F, i);
        phase_sync<MODE>(a.grid_bar);
    pc.lap(4);
      }
      // 32-item groups stay warp-aligned so each warp ballots its own word of the active masks
      for (uint32_t i = tid; (i & ~31u) < V; i += nth) {
        const bool live =
            i < V && gen_vertex_item<false>(a.bvh.tnodes, a.bvh.pred, c.x0, c.x, c.vf, c.rt, nullptr, nullptr, i);
        store_active_bits(c.v_active, i, V, live);
      }
      for (uint32_t i = tid; (i & ~31u) < E; i += nth) {
        const bool live =
            i < E && gen_edge_item<false>(a.bvh.tnodes, a.bvh.pred, o.rtri, c.x, c.edge, c.ee, nullptr, i);
        store_active_bits(c.e_active, i, E, live);
      }
      phase_sync<MODE>(a.grid_bar);
      pc.lap(5);
    }
    // Colour phases of stretch then bend.  The record of this thread's first
    // item of the NEXT phase is static data, so it is fetched before the barrier
    // and its latency overlaps the store drain + barrier.
    {
      const uint32_t n_sc = c.n_sc, n_ph = c.n_sc + (a.solve_bending ? c.n_bc : 0u);
      ColorRec cur = fetch_color_rec(a, 0, n_sc, n_ph, tid, it);
      for (uint32_t ph = 0; ph < n_ph; ph++) {
        if (ph < n_sc) {
          const uint32_t end = a.s_off[ph + 1];
          uint32_t i = a.s_off[ph] + tid;
          if (i < end) {
            if (a.atomic_s) {
              stretch_rec(c.x, c.s_lambda, i, make_uint2(cur.q.x, cur.q.y), cur.f, cur.lambda, alpha_s, true);
              for (i += nth; i < end; i += nth) stretch_one(c.x, c.s_idx, c.s_rest, c.s_lambda, i, alpha_s, it, true);
            } else {
              stretch_pipelined(c, i, end, nth, cur, alpha_s, it);
            }
          }
        } else {
          const uint32_t col = ph - n_sc, end = a.b_off[col + 1];
          uint32_t i = a.b_off[col] + tid;
          if (i < end) {
            bend_rec(c.x, cur.q, cur.f, xpbd_b, a.atomic_b != 0);
            for (i += nth; i < end; i += nth) bend_one(c.x, c.b_idx, c.b_theta, i, xpbd_b, a.atomic_b != 0);
          }
        }
        cur = fetch_color_rec(a, ph + 1, n_sc, n_ph, tid, it);
        phase_sync<MODE>(a.grid_bar);
        pc.lap(ph < n_sc ? 6 : 7);
      }
    }
    if (a.lra && c.n_lra) {
      for (uint32_t i = tid; i < c.n_lra; i += nth) lra_item(c.x, c.l_src, c.l_tgt, c.l_rest, k.lra_stretchiness, i);
      phase_sync<MODE>(a.grid_bar);
    pc.lap(8);
    }
    if (o.F) {
      for (uint32_t i = tid; i < V; i += nth)
        if (load_active_bit(c.v_active, i)) solve_vertex_item(a.bvh.pred, c.x, c.vf, c.rt, i);
      phase_sync<MODE>(a.grid_bar);
    pc.lap(9);
      if (E) {
        for (uint32_t i = tid; i < E; i += nth)
          if (load_active_bit(c.e_active, i)) solve_edge_item(c.x, c.edge, c.ee, c.delta, i);
        phase_sync<MODE>(a.grid_bar);
    pc.lap(10);
      }
      for (uint32_t i = tid; i < V; i += nth) apply_delta_item(c.x, c.delta, k.relaxation, i);
      phase_sync<MODE>(a.grid_bar);
    pc.lap(11);
      if (c.n_bind) {  // binding_->Solve, ClothSolverBase.cpp:236
        for (uint32_t i = tid; i < c.n_bind; i += nth) binding_item(c.x, o.x, c.bd_idx, c.bd_par, c.bd_k, i);
        phase_sync<MODE>(a.grid_bar);
      }
    }
  }
  if (a.do_finalize)
    for (uint32_t i = tid; i < V; i += nth) finalize_item(c.x0, c.x, c.v, k, i);
  pc.lap(12);
}

template <int MODE>
__global__ void __launch_bounds__(PT, 1) swift_substep_kernel(const __grid_constant__ PersistArgs a) {
  swift_substep_body<MODE>(a);
}

// Shared launch: ONE kernel steps a whole batch of garments, one CTA (SYNC_BLOCK) or one
// cluster (SYNC_CLUSTER) per garment.  The garment's argument block is copied from the batch's
// device table into shared memory; `moved` selects the variant with / without a new obstacle
// target.  No stream (and no hardware queue) per garment: the batch is as wide as the device.
template <int MODE>
__global__ void __launch_bounds__(PT, 1) swift_substep_batch_kernel(const PersistArgs* __restrict__ table,
                                                                    const int* __restrict__ moved, uint32_t n,
                                                                    int sub_begin, int sub_end) {
  __shared__ PersistArgs sa;
  const uint32_t g = MODE == SYNC_CLUSTER ? blockIdx.x / cluster_nctarank() : blockIdx.x;
  static_assert(sizeof(PersistArgs) % 4 == 0, "argument block is copied word by word");
  const uint32_t mv = moved[g] ? 1u : 0u;
  // XRT_PERSIST_PROF: [13] time inside shared launches, [14] time between them (stage kernels, launch latency, waiting
  // for the slowest garment of the group), [15] end of this garment's previous launch
  unsigned long long* span = nullptr;
  unsigned long long t_in = 0;
  if (threadIdx.x == 0 && (MODE != SYNC_CLUSTER || cluster_ctarank() == 0)) {
    span = table[((size_t)sub_begin * 2u + mv) * n + g].prof;
    if (span) {
      t_in = gtimer();
      if (span[15]) span[14] += t_in - span[15];
    }
  }
  // consecutive substeps that need nothing from outside (no hash, no build) run in one launch:
  // a garment does not wait for the slowest garment of the batch between them
  for (int sub = sub_begin; sub < sub_end; sub++) {
    const uint32_t* src = reinterpret_cast<const uint32_t*>(table + ((size_t)sub * 2u + mv) * n + g);
    uint32_t* dst = reinterpret_cast<uint32_t*>(&sa);
    for (uint32_t w = threadIdx.x; w < sizeof(PersistArgs) / 4; w += PT) dst[w] = src[w];
    __syncthreads();
    swift_substep_body<MODE>(sa);
    phase_sync<MODE>();  // the substep's last phase is complete (and sa is free) before the next one starts
  }
  if (span) {
    const unsigned long long t_out = gtimer();
    span[13] += t_out - t_in;
    span[15] = t_out;
  }
}

// ------------------------------------------------------------------------------------------------
// Distributed-shared-memory shape (single-garment latency).
//
// What bounds ONE garment is not bandwidth but ~1 100 dependent Gauss-Seidel phases per frame, each a short
// gather -> project -> scatter over the predicted positions.  In the cluster shape above those positions travel
// through L2: every phase pays a drain of the scatter to L2 (MEMBAR.GPU), an L1 invalidate and ~700-cycle gathers.
// Here a 16-CTA cluster keeps Node::x of the whole garment in its shared memory (50 176 vertices x 16 B = 0.8 MB
// over 16 x 64 KB; 100 489 over 16 x 128 KB): a gather is a 215-cycle ld.shared::cluster, a scatter a
// st.shared::cluster, the phase boundary the hardware cluster barrier.  Per-vertex phases run on the CTA that owns
// the vertex, so their own position is a local shared-memory access.
// The BVH queries (contact generation) are the opposite kind of work -- wide and independent -- so they stay
// device-wide kernels: the host cuts each substep into SEGMENTS at the iterations where a collision pass is due
// (every `frequency` steps, ClothSolverBase.cpp:204), runs interpolate / refit / generate as per-stage kernels on all
// 148 SMs, and the cluster resumes.  A frame is ~10 generation passes + ~14 segments in one CUDA graph.
// Same device bodies as every other shape, so positions are bit-identical to the per-stage path.
// ------------------------------------------------------------------------------------------------
}  // namespace
namespace xrtdev {
struct XDsm {
  uint32_t base;   // shared-window address of the slice (the same offset in every CTA of the cluster)
  uint32_t shift;  // vertex i lives in CTA i >> shift, slot i & mask
  uint32_t mask;
  struct Ref {
    uint32_t addr;
    __device__ __forceinline__ operator float4() const {
      float4 v;
      asm volatile("ld.shared::cluster.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(addr));
      return v;
    }
    __device__ __forceinline__ void operator=(float4 v) const {
      asm volatile("st.shared::cluster.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
    }
  };
  __device__ __forceinline__ uint32_t addr(uint32_t i) const {
    uint32_t a;
    asm("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(a) : "r"(base + ((i & mask) << 4)), "r"(i >> shift));
    return a;
  }
  __device__ __forceinline__ Ref operator[](uint32_t i) const { return Ref{addr(i)}; }
};
// improper colourings are rejected for this shape at setup; kept so add_pos<XDsm> instantiates
__device__ __forceinline__ void pos_atomic_add(XDsm x, uint32_t i, f3 c) {
  const uint32_t a = x.addr(i);
  asm volatile("red.shared::cluster.add.f32 [%0], %1;" ::"r"(a), "f"(c.x) : "memory");
  asm volatile("red.shared::cluster.add.f32 [%0], %1;" ::"r"(a + 4), "f"(c.y) : "memory");
  asm volatile("red.shared::cluster.add.f32 [%0], %1;" ::"r"(a + 8), "f"(c.z) : "memory");
}
}  // namespace xrtdev
namespace {

// One segment of a substep: [predict] then, from iteration it_begin on,
//   head of an iteration = particle self-contact (collide + apply), tail = colours, LRA, contact solves, binding.
// resume: iteration it_begin starts at its tail (its head and its collision pass ran before this launch).
// The segment ends after the head of iteration stop_at (a collision pass is due there: the host runs it next), or
// after the last iteration, with finalize.
__global__ void __launch_bounds__(PT, 1) swift_segment_kernel(const __grid_constant__ PersistArgs a) {
  extern __shared__ float4 xs[];
  const uint32_t rank = cluster_ctarank();
  const uint32_t nth = cluster_nctarank() * PT, tid = rank * PT + threadIdx.x;
  const ClothDev& c = a.c;
  const ObstacleDev& o = a.o;
  const StepConst& k = a.k;
  const uint32_t V = c.V, E = c.E;
  const float alpha_s = k.stretch_compliance / (k.dt * k.dt);
  const float xpbd_b = k.bend_compliance / k.dt / k.dt;
  const XDsm x{(uint32_t)__cvta_generic_to_shared(xs), a.x_shift, (1u << a.x_shift) - 1u};
  const uint32_t lo = rank << a.x_shift;                                // this CTA owns vertices [lo, hi)
  const uint32_t hi = min(V, lo + (1u << a.x_shift));
  const bool fresh = a.do_predict && a.it_begin == 0 && !a.resume;
  PhaseClock pc;
  pc.start(a.prof, tid);
  if (fresh) {
    for (uint32_t i = lo + threadIdx.x; i < hi; i += PT) predict_item(c.x0, x, c.v, k, i);
  } else {
    for (uint32_t i = lo + threadIdx.x; i < hi; i += PT) xs[i - lo] = c.x[i];
  }
  phase_sync<SYNC_CLUSTER>();
  pc.lap(0);
  int it = a.it_begin;
  bool skip_head = a.resume != 0;
  while (it < a.n_iter) {
    if (!skip_head) {
      if (a.self_collision) {
        for (uint32_t i = lo + threadIdx.x; i < hi; i += PT) collide_item(c.delta, x, c.x0, c.neighbors, k, V, i);
        phase_sync<SYNC_CLUSTER>();
        pc.lap(1);
        for (uint32_t i = lo + threadIdx.x; i < hi; i += PT) apply_delta_item(x, c.delta, k.relaxation, i);
        phase_sync<SYNC_CLUSTER>();
        pc.lap(2);
      }
      if (it == a.stop_at) break;
    }
    skip_head = false;
    {
      // a colour phase: every CTA projects the constraints it owns (those whose lowest vertex lives in its slice), so
      // on a mesh numbered row by row all but the constraints across a slice boundary touch local shared memory only
      // (the cluster's remote path moves 17 B per cycle and SM; a colour with every endpoint remote is bound by it)
      const uint32_t n_sc = c.n_sc, n_ph = c.n_sc + (a.solve_bending ? c.n_bc : 0u);
      ColorRec cur = fetch_color_rec_cta(a, 0, n_sc, n_ph, rank, it);
      for (uint32_t ph = 0; ph < n_ph; ph++) {
        if (ph < n_sc) {
          const uint32_t end = a.s_cta_off[ph * (XRT_DSMEM_CLUSTER + 1) + rank + 1];
          uint32_t i = a.s_cta_off[ph * (XRT_DSMEM_CLUSTER + 1) + rank] + threadIdx.x;
          if (i < end) {
            stretch_rec(x, c.s_lambda, i, make_uint2(cur.q.x, cur.q.y), cur.f, cur.lambda, alpha_s, false);
            for (i += PT; i < end; i += PT) stretch_one(x, c.s_idx, c.s_rest, c.s_lambda, i, alpha_s, it, false);
          }
        } else {
          const uint32_t col = ph - n_sc, end = a.b_cta_off[col * (XRT_DSMEM_CLUSTER + 1) + rank + 1];
          uint32_t i = a.b_cta_off[col * (XRT_DSMEM_CLUSTER + 1) + rank] + threadIdx.x;
          if (i < end) {
            bend_rec(x, cur.q, cur.f, xpbd_b, false);
            for (i += PT; i < end; i += PT) bend_one(x, c.b_idx, c.b_theta, i, xpbd_b, false);
          }
        }
        cur = fetch_color_rec_cta(a, ph + 1, n_sc, n_ph, rank, it);
        phase_sync<SYNC_CLUSTER>();
        pc.lap(ph < n_sc ? 6 : 7);
      }
    }
    if (a.lra && c.n_lra) {
      for (uint32_t i = tid; i < c.n_lra; i += nth) lra_item(x, c.l_src, c.l_tgt, c.l_rest, k.lra_stretchiness, i);
      phase_sync<SYNC_CLUSTER>();
    }
    if (o.F) {
      for (uint32_t i = lo + threadIdx.x; i < hi; i += PT)
        if (load_active_bit(c.v_active, i)) solve_vertex_item(a.bvh.pred, x, c.vf, c.rt, i);
      phase_sync<SYNC_CLUSTER>();
      pc.lap(9);
      if (E) {
        for (uint32_t i = tid; i < E; i += nth)
          if (load_active_bit(c.e_active, i)) solve_edge_item(x, c.edge, c.ee, c.delta, i);
        phase_sync<SYNC_CLUSTER>();
        pc.lap(10);
      }
      for (uint32_t i = lo + threadIdx.x; i < hi; i += PT) apply_delta_item(x, c.delta, k.relaxation, i);
      phase_sync<SYNC_CLUSTER>();
      pc.lap(11);
      if (c.n_bind) {
        for (uint32_t i = tid; i < c.n_bind; i += nth) binding_item(x, o.x, c.bd_idx, c.bd_par, c.bd_k, i);
        phase_sync<SYNC_CLUSTER>();
      }
    }
    it++;
  }
  if (it >= a.n_iter && a.do_finalize)
    for (uint32_t i = lo + threadIdx.x; i < hi; i += PT) finalize_item(c.x0, x, c.v, k, i);
  // Node::x back to global memory for the kernels that run between segments (BVH queries, spatial hash, getters);
  // every CTA's slice is final here: the last phase above ended with a cluster barrier
  for (uint32_t i = lo + threadIdx.x; i < hi; i += PT) c.x[i] = xs[i - lo];
  phase_sync<SYNC_CLUSTER>();  // no CTA leaves while another one may still read its shared memory
  pc.lap(12);
}

}  // namespace

// per-device caches: the non-portable cluster opt-in and the occupancy answers belong to the current device
static int current_device_slot() {
  int dev = 0;
  cudaGetDevice(&dev);
  return dev >= 0 && dev < 64 ? dev : 0;
}

int persist_max_cluster() {
  static int cache[64];
  static bool have[64] = {false};
  const int slot = current_device_slot();
  if (have[slot]) return cache[slot];
  int& cached = cache[slot];
  // clusters above 8 CTAs need the non-portable opt-in
  cudaFuncSetAttribute(swift_substep_kernel<SYNC_CLUSTER>, cudaFuncAttributeNonPortableClusterSizeAllowed, 1);
  cudaFuncSetAttribute(swift_substep_batch_kernel<SYNC_CLUSTER>, cudaFuncAttributeNonPortableClusterSizeAllowed, 1);
  int best = 1;
  for (int cs = 2; cs <= 16; cs *= 2) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(cs);
    cfg.blockDim = dim3(PT);
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeClusterDimension;
    at[0].val.clusterDim.x = cs;
    at[0].val.clusterDim.y = 1;
    at[0].val.clusterDim.z = 1;
    cfg.attrs = at;
    cfg.numAttrs = 1;
    int n = 0;
    if (cudaOccupancyMaxActiveClusters(&n, swift_substep_kernel<SYNC_CLUSTER>, &cfg) == cudaSuccess && n > 0) best = cs;
  }
  (void)cudaGetLastError();
  cached = best;
  have[slot] = true;
  return cached;
}

int persist_active_clusters(int cluster) {
  if (cluster <= 1) {
    int dev = 0, sms = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    return sms;
  }
  persist_max_cluster();
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(cluster);
  cfg.blockDim = dim3(PT);
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeClusterDimension;
  at[0].val.clusterDim.x = cluster;
  at[0].val.clusterDim.y = 1;
  at[0].val.clusterDim.z = 1;
  cfg.attrs = at;
  cfg.numAttrs = 1;
  int n = 0;
  if (cudaOccupancyMaxActiveClusters(&n, swift_substep_kernel<SYNC_CLUSTER>, &cfg) != cudaSuccess) {
    (void)cudaGetLastError();
    return 0;
  }
  return n;
}

int persist_grid_ctas() {
  static int cache[64];
  static bool have[64] = {false};
  const int slot = current_device_slot();
  if (have[slot]) return cache[slot];
  int& cached = cache[slot];
  int dev = 0, sms = 0, per_sm = 0;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, swift_substep_kernel<SYNC_GRID>, PT, 0);
  cached = per_sm > 0 ? sms : 0;  // one CTA per SM, all co-resident
  have[slot] = true;
  return cached;
}

cudaError_t launch_swift_substep(const PersistArgs& a, int cluster, cudaStream_t st) {
  if (cluster == XRT_PERSIST_GRID) {
    // whole device on one garment; the cooperative attribute guarantees co-residency
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(persist_grid_ctas());
    cfg.blockDim = dim3(PT);
    cfg.stream = st;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeCooperative;
    at[0].val.cooperative = 1;
    cfg.attrs = at;
    cfg.numAttrs = 1;
    return cudaLaunchKernelEx(&cfg, swift_substep_kernel<SYNC_GRID>, a);
  }
  if (cluster <= 1) {
    swift_substep_kernel<SYNC_BLOCK><<<1, PT, 0, st>>>(a);
    return cudaGetLastError();
  }
  persist_max_cluster();  // sets the non-portable attribute once
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(cluster);
  cfg.blockDim = dim3(PT);
  cfg.stream = st;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeClusterDimension;
  at[0].val.clusterDim.x = cluster;
  at[0].val.clusterDim.y = 1;
  at[0].val.clusterDim.z = 1;
  cfg.attrs = at;
  cfg.numAttrs = 1;
  return cudaLaunchKernelEx(&cfg, swift_substep_kernel<SYNC_CLUSTER>, a);
}

int persist_dsmem_shift(uint32_t V) {
  if (persist_max_cluster() < XRT_DSMEM_CLUSTER) return -1;
  for (int sh = 6; sh <= XRT_DSMEM_MAX_SHIFT; sh++)
    if (((uint64_t)XRT_DSMEM_CLUSTER << sh) >= V) return sh;
  return -1;
}

cudaError_t launch_swift_segment(const PersistArgs& a, cudaStream_t st) {
  static bool once[64] = {false};
  const int slot = current_device_slot();
  if (!once[slot]) {
    cudaFuncSetAttribute(swift_segment_kernel, cudaFuncAttributeNonPortableClusterSizeAllowed, 1);
    cudaFuncSetAttribute(swift_segment_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (16 << XRT_DSMEM_MAX_SHIFT));
    once[slot] = true;
  }
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(XRT_DSMEM_CLUSTER);
  cfg.blockDim = dim3(PT);
  cfg.dynamicSmemBytes = (size_t)16 << a.x_shift;
  cfg.stream = st;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeClusterDimension;
  at[0].val.clusterDim.x = XRT_DSMEM_CLUSTER;
  at[0].val.clusterDim.y = 1;
  at[0].val.clusterDim.z = 1;
  cfg.attrs = at;
  cfg.numAttrs = 1;
  return cudaLaunchKernelEx(&cfg, swift_segment_kernel, a);
}

cudaError_t launch_swift_substep_batch(const PersistArgs* table, const int* moved, uint32_t n, int sub_begin,
                                       int sub_end, int cluster, cudaStream_t st) {
  if (cluster <= 1) {
    swift_substep_batch_kernel<SYNC_BLOCK><<<n, PT, 0, st>>>(table, moved, n, sub_begin, sub_end);
    return cudaGetLastError();
  }
  persist_max_cluster();
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(n * (uint32_t)cluster);
  cfg.blockDim = dim3(PT);
  cfg.stream = st;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeClusterDimension;
  at[0].val.clusterDim.x = cluster;
  at[0].val.clusterDim.y = 1;
  at[0].val.clusterDim.z = 1;
  cfg.attrs = at;
  cfg.numAttrs = 1;
  return cudaLaunchKernelEx(&cfg, swift_substep_batch_kernel<SYNC_CLUSTER>, table, moved, n, sub_begin, sub_end);
}
