// Shared-launch batches (xrt_batch_*): n garments of one device step in lockstep, each substep of the whole batch
// ONE launch of swift_substep_batch_kernel (one CTA or cluster per garment), the per-garment work in between (obstacle
// BVH build, predict + spatial hash, end-of-frame normals) as small CUDA graphs fanned out over worker streams;
// read-back mode snapshots every frame on the device and brings it to pinned memory on a copy stream.
// "instances batched into shared launches".  Split out of xrt_api.cu in round 2.
#include "xrt_handle.cuh"

#include <algorithm>
#include <cstring>
#include <vector>

// ---- shared-launch batch ------------------------------------------------------------------
// A batch steps many independent garments in lockstep: every substep is ONE launch of
// swift_substep_batch_kernel with one CTA / cluster per garment, so the batch is not limited by
// the number of hardware queues (32) and small clusters (down to one SM per garment, where a phase
// boundary is a bar.sync) fill the device.  What stays per garment and per stream is the work
// before a substep: the obstacle BVH build of the frame and predict + spatial hash when due.
struct BatchGroup {
  int first = 0, n = 0, seg_begin = 0;
  cudaStream_t st = nullptr;
  std::vector<cudaStream_t> wk;   // worker streams for the garments' own graphs between shared launches
  std::vector<cudaEvent_t> ev_wk;
  cudaEvent_t ev_fork = nullptr, ev_flags = nullptr, ev_done = nullptr;
  cudaStream_t st_copy = nullptr;  // read-back mode: D2H of the frame snapshots
  cudaEvent_t ev_snap = nullptr;
  PersistArgs* d_table = nullptr;  // [S][2][n]
  int *d_moved = nullptr, *h_moved = nullptr;
  // the per-garment stages between two shared launches as ONE launch per stage over the whole group (blockIdx.y =
  // garment): obstacle BVH builds, predict + spatial hash, end-of-frame copy + normals.  Used when every garment of
  // the group hashes on the same substeps (`uniform`); otherwise the per-garment graphs on the worker streams.
  bool uniform = false;
  BvhBatch bb;
  HashBatch hb;
  NormalsBatchItem* d_post = nullptr;
  uint32_t n_post = 0, post_max_v = 0;
};
// Garments are split into `groups` that step independently (own stream, own shared launches):
// while one group's substep kernel holds its SMs, another group's builds and hashes run.
// Every stream of the batch should own a hardware queue (32 per device): groups * (1 + workers) <= 32.
struct xrt_batch {
  std::vector<xrt_solver*> h;
  std::vector<BatchGroup> g;
  int cluster = 1, S = 1;
  cudaEvent_t ev0 = nullptr, ev1 = nullptr;
  std::vector<cudaEvent_t> ev_pre;  // per garment: handle stream -> batch join, pipelined read-back
  // read-back mode (xrt_batch_set_readback): every frame ends with a device-side snapshot of the cloth
  // (positions + normals, packed float3) that a copy stream brings to pinned host memory while the next
  // frame computes; two frames may be outstanding
  bool readback = false;
  std::vector<float*> d_snap, h_snap[2];
  std::vector<cudaEvent_t> ev_dn[2];
  unsigned long long snap_written = 0, snap_read = 0;
  // kernels of this library enqueued so far, counted where they are launched (xrt_batch_launch_count); cub's sort
  // kernels are library code and counted as calls
  unsigned long long own_launches = 0, sort_calls = 0;
};

extern "C" void xrt_batch_destroy(xrt_batch* b) {
  if (!b) return;
  if (!b->h.empty()) cudaSetDevice(b->h[0]->p.device);
  for (BatchGroup& g : b->g) {
    if (g.st) cudaStreamSynchronize(g.st);
    for (cudaStream_t w : g.wk)
      if (w) cudaStreamDestroy(w);
    for (cudaEvent_t e : g.ev_wk)
      if (e) cudaEventDestroy(e);
  }
  for (BatchGroup& g : b->g) {
    if (g.st_copy) {
      cudaStreamSynchronize(g.st_copy);
      cudaStreamDestroy(g.st_copy);
    }
    if (g.ev_snap) cudaEventDestroy(g.ev_snap);
  }
  for (float* p : b->d_snap) cudaFree(p);
  for (int k = 0; k < 2; k++) {
    for (float* p : b->h_snap[k]) cudaFreeHost(p);
    for (cudaEvent_t e : b->ev_dn[k])
      if (e) cudaEventDestroy(e);
  }
  for (xrt_solver* h : b->h) h->batch_ev = nullptr;  // everything it pointed at has completed
  for (BatchGroup& g : b->g) {
    for (cudaEvent_t e : {g.ev_fork, g.ev_flags, g.ev_done})
      if (e) cudaEventDestroy(e);
    cudaFree(g.d_table);
    cudaFree(g.d_moved);
    bvh_batch_free(&g.bb);
    hash_batch_free(&g.hb);
    cudaFree(g.d_post);
    cudaFreeHost(g.h_moved);
    if (g.st) cudaStreamDestroy(g.st);
  }
  for (cudaEvent_t e : b->ev_pre)
    if (e) cudaEventDestroy(e);
  for (cudaEvent_t e : {b->ev0, b->ev1})
    if (e) cudaEventDestroy(e);
  delete b;
}

extern "C" int xrt_batch_create(xrt_solver** handles, int n, int groups, int workers, xrt_batch** out) {
  if (!handles || n <= 0 || !out) return XRT_ERR_INVALID;
  xrt_solver* s = handles[0];
  XRT_READY();
  for (int i = 0; i < n; i++) {
    xrt_solver* h = handles[i];
    if (!h || !h->finalized || h->p.device != s->p.device || h->q) {
      xrt_set_error(s, "batch_create: handle %d is null, not finalized, on another device or a Quality-mode handle", i);
      return XRT_ERR_INVALID;
    }
    if (h->cluster < 1 || h->cluster == XRT_PERSIST_GRID || h->cluster == XRT_PERSIST_DSMEM || h->cluster != s->cluster ||
        h->p.num_substeps != s->p.num_substeps) {
      xrt_set_error(s, "batch_create: handle %d: a shared launch needs persist_cluster >= 1 and the same cluster size "
                       "and substep count on every handle", i);
      return XRT_ERR_INVALID;
    }
  }
  groups = std::max(1, std::min(groups, n));
  workers = std::max(0, std::min(workers, 31));
  xrt_batch* b = new xrt_batch;
  b->h.assign(handles, handles + n);
  b->cluster = s->cluster;
  b->S = s->p.num_substeps;
  b->g.resize(groups);
  b->ev_pre.assign(n, nullptr);
  cudaError_t e = cudaEventCreate(&b->ev0);
  if (e == cudaSuccess) e = cudaEventCreate(&b->ev1);
  for (int i = 0; i < n && e == cudaSuccess; i++) e = cudaEventCreateWithFlags(&b->ev_pre[i], cudaEventDisableTiming);
  for (int gi = 0; gi < groups && e == cudaSuccess; gi++) {
    BatchGroup& g = b->g[gi];
    g.first = (int)((long long)n * gi / groups);
    g.n = (int)((long long)n * (gi + 1) / groups) - g.first;
    std::vector<PersistArgs> tab((size_t)b->S * 2 * g.n);
    for (int sub = 0; sub < b->S; sub++)
      for (int mv = 0; mv < 2; mv++)
        for (int i = 0; i < g.n; i++) {
          const xrt_solver* h = handles[g.first + i];
          const bool hash = h->p.enable_self_collision && sub % std::max(1, h->p.inter_leaved_hash) == 0;
          tab[((size_t)sub * 2 + mv) * g.n + i] = xrt_make_persist_args(h, sub, mv != 0, !hash);
        }
    e = cudaStreamCreateWithFlags(&g.st, cudaStreamNonBlocking);
    g.wk.assign(std::min(workers, g.n), nullptr);
    g.ev_wk.assign(g.wk.size(), nullptr);
    for (size_t w = 0; w < g.wk.size() && e == cudaSuccess; w++) {
      e = cudaStreamCreateWithFlags(&g.wk[w], cudaStreamNonBlocking);
      if (e == cudaSuccess) e = cudaEventCreateWithFlags(&g.ev_wk[w], cudaEventDisableTiming);
    }
    if (e == cudaSuccess) e = cudaMalloc((void**)&g.d_table, tab.size() * sizeof(PersistArgs));
    if (e == cudaSuccess)
      e = cudaMemcpy(g.d_table, tab.data(), tab.size() * sizeof(PersistArgs), cudaMemcpyHostToDevice);
    if (e == cudaSuccess) e = cudaMalloc((void**)&g.d_moved, g.n * sizeof(int));
    if (e == cudaSuccess) e = cudaMallocHost((void**)&g.h_moved, g.n * sizeof(int));
    if (e == cudaSuccess) e = cudaEventCreateWithFlags(&g.ev_fork, cudaEventDisableTiming);
    if (e == cudaSuccess) e = cudaEventCreateWithFlags(&g.ev_flags, cudaEventDisableTiming);
    if (e == cudaSuccess) e = cudaEventCreateWithFlags(&g.ev_done, cudaEventDisableTiming);
    // group-wide stages
    if (e == cudaSuccess) {
      const xrt_solver* h0 = handles[g.first];
      g.uniform = true;
      std::vector<BvhBatchItem> bi;
      std::vector<HashBatchItem> hi;
      std::vector<NormalsBatchItem> ni;
      for (int i = 0; i < g.n; i++) {
        xrt_solver* h = handles[g.first + i];
        if (h->p.enable_self_collision != h0->p.enable_self_collision || h->p.inter_leaved_hash != h0->p.inter_leaved_hash ||
            h->p.morton_fast_div != h0->p.morton_fast_div)
          g.uniform = false;
        BvhBatchItem b1{};
        b1.b = h->bvh; b1.w = h->bw; b1.x = h->o.x0; b1.tri = h->o.tri; b1.F = h->o.F;
        bi.push_back(b1);
        HashBatchItem h1{};
        h1.c = h->c; h1.k = h->k;
        hi.push_back(h1);
        if (h->o.V) ni.push_back(NormalsBatchItem{h->o.x0, h->o.x, h->o.n, h->o.tri, h->o.vf_off, h->o.vf_idx, h->o.V, (uint32_t)i});
        ni.push_back(NormalsBatchItem{h->c.x0, nullptr, h->c.n, h->c.tri, h->c.vf_off, h->c.vf_idx, h->c.V, (uint32_t)i});
        g.post_max_v = std::max(g.post_max_v, std::max(h->o.V, h->c.V));
      }
      if (g.uniform) {
        if (bvh_batch_alloc(&g.bb, bi) || hash_batch_alloc(&g.hb, hi)) e = cudaErrorMemoryAllocation;
        g.n_post = (uint32_t)ni.size();
        if (e == cudaSuccess) e = cudaMalloc((void**)&g.d_post, ni.size() * sizeof(NormalsBatchItem));
        if (e == cudaSuccess) e = cudaMemcpy(g.d_post, ni.data(), ni.size() * sizeof(NormalsBatchItem), cudaMemcpyHostToDevice);
        // the sorted keys of a tree now live in the group's buffers (what the BVH getters read)
        uint32_t seg = 0;
        for (int i = 0; i < g.n && e == cudaSuccess; i++) {
          xrt_solver* h = handles[g.first + i];
          if (h->o.F) {
            h->bvh.morton = g.bb.key_out + seg;
            h->bvh.sorted_obj = g.bb.val_out + seg;
          }
          seg += h->o.F;
        }
      }
    }
  }
  if (e != cudaSuccess) {
    xrt_set_error(s, "batch_create: %s", cudaGetErrorString(e));
    xrt_batch_destroy(b);
    return XRT_ERR_CUDA;
  }
  *out = b;
  return XRT_OK;
}

// A garment's own launches between two shared launches, captured once per (substep, moved) so that
// a frame costs the host a handful of calls per garment (eager: ~40 per garment, which made 148
// garments host-bound at 68 ms per frame).
static int batch_pre_graph(xrt_solver* s, int sub, int moved, cudaGraphExec_t* out) {
  const bool hash = s->p.enable_self_collision && sub % std::max(1, s->p.inter_leaved_hash) == 0;
  const bool build = sub == 0 && moved && s->o.F;
  *out = nullptr;
  if (!hash && !build) return XRT_OK;
  if (s->pre_graph.empty()) s->pre_graph.assign((size_t)2 * s->p.num_substeps, nullptr);
  cudaGraphExec_t& slot = s->pre_graph[(size_t)2 * sub + moved];
  if (!slot) {
    cudaGraph_t g = nullptr;
    XRT_CUDA_OK(cudaStreamBeginCapture(s->st, cudaStreamCaptureModeThreadLocal));
    if (build) bvh_build(s->bvh, s->bw, s->o.x0, s->o.tri, s->o.F, s->p.morton_fast_div, s->st);
    if (hash) {
      launch_predict(s->c, s->k, s->st);
      launch_hash(s->c, s->k, s->st);
    }
    XRT_CUDA_OK(cudaStreamEndCapture(s->st, &g));
    XRT_CUDA_OK(cudaGraphInstantiate(&slot, g, 0));
    cudaGraphDestroy(g);
  }
  *out = slot;
  return XRT_OK;
}
// obstacle_->PredictedToPositions(); cloth_/obstacle_->UpdateNormals()
static int batch_post_graph(xrt_solver* s, int moved, cudaGraphExec_t* out) {
  cudaGraphExec_t& slot = s->post_graph[moved];
  if (!slot) {
    cudaGraph_t g = nullptr;
    XRT_CUDA_OK(cudaStreamBeginCapture(s->st, cudaStreamCaptureModeThreadLocal));
    if (s->o.V) {
      if (moved) cudaMemcpyAsync(s->o.x0, s->o.x, s->o.V * sizeof(float4), cudaMemcpyDeviceToDevice, s->st);
      launch_normals(s->o.x0, s->o.n, s->o.tri, s->o.vf_off, s->o.vf_idx, s->o.V, s->st);
    }
    launch_normals(s->c.x0, s->c.n, s->c.tri, s->c.vf_off, s->c.vf_idx, s->c.V, s->st);
    XRT_CUDA_OK(cudaStreamEndCapture(s->st, &g));
    XRT_CUDA_OK(cudaGraphInstantiate(&slot, g, 0));
    cudaGraphDestroy(g);
  }
  *out = slot;
  return XRT_OK;
}

// the garments' own graphs of one stage, spread over the group's worker streams and joined back
static int batch_fan_out(xrt_batch* b, BatchGroup& g, const std::vector<cudaGraphExec_t>& gr) {
  xrt_solver* s = b->h[0];
  const size_t W = g.wk.size();
  if (!W) {
    for (cudaGraphExec_t e : gr)
      if (e) XRT_CUDA_OK(cudaGraphLaunch(e, g.st));
    return XRT_OK;
  }
  bool any = false;
  for (cudaGraphExec_t e : gr) any = any || e;
  if (!any) return XRT_OK;
  XRT_CUDA_OK(cudaEventRecord(g.ev_fork, g.st));
  std::vector<char> used(W, 0);
  size_t k = 0;
  for (cudaGraphExec_t e : gr) {
    if (!e) continue;
    const size_t w = k++ % W;
    if (!used[w]) {
      XRT_CUDA_OK(cudaStreamWaitEvent(g.wk[w], g.ev_fork, 0));
      used[w] = 1;
    }
    XRT_CUDA_OK(cudaGraphLaunch(e, g.wk[w]));
  }
  for (size_t w = 0; w < W; w++)
    if (used[w]) {
      XRT_CUDA_OK(cudaEventRecord(g.ev_wk[w], g.wk[w]));
      XRT_CUDA_OK(cudaStreamWaitEvent(g.st, g.ev_wk[w], 0));
    }
  return XRT_OK;
}

extern "C" int xrt_batch_set_readback(xrt_batch* b, int on) {
  if (!b || b->h.empty()) return XRT_ERR_INVALID;
  xrt_solver* s = b->h[0];
  XRT_CUDA_OK(cudaSetDevice(s->p.device));
  XRT_REQUIRE(b->snap_read == b->snap_written, "batch_set_readback: read the outstanding frames back first");
  if (on && b->d_snap.empty()) {
    const size_t n = b->h.size();
    b->d_snap.assign(n, nullptr);
    for (int k = 0; k < 2; k++) {
      b->h_snap[k].assign(n, nullptr);
      b->ev_dn[k].assign(n, nullptr);
    }
    for (size_t i = 0; i < n; i++) {
      const size_t bytes = (size_t)b->h[i]->c.V * 6 * sizeof(float);
      XRT_CUDA_OK(cudaMalloc((void**)&b->d_snap[i], bytes ? bytes : 16));
      for (int k = 0; k < 2; k++) {
        XRT_CUDA_OK(cudaMallocHost((void**)&b->h_snap[k][i], bytes ? bytes : 16));
        XRT_CUDA_OK(cudaEventCreateWithFlags(&b->ev_dn[k][i], cudaEventDisableTiming));
      }
    }
    for (BatchGroup& g : b->g) {
      XRT_CUDA_OK(cudaStreamCreateWithFlags(&g.st_copy, cudaStreamNonBlocking));
      XRT_CUDA_OK(cudaEventCreateWithFlags(&g.ev_snap, cudaEventDisableTiming));
    }
  }
  b->readback = on != 0;
  return XRT_OK;
}

// end of frame in read-back mode: snapshot on the group's stream, D2H on its copy stream
static int batch_snapshot(xrt_batch* b, BatchGroup& g) {
  xrt_solver* s = b->h[0];
  const int par = (int)(b->snap_written & 1ull);
  for (int i = 0; i < g.n; i++) {
    xrt_solver* h = b->h[g.first + i];
    if (!h->c.V) continue;
    // the copy of the previous frame has left the device staging buffer
    XRT_CUDA_OK(cudaStreamWaitEvent(g.st, b->ev_dn[par ^ 1][g.first + i], 0));
    launch_pack_xyz(h->c.x0, h->c.n, b->d_snap[g.first + i], (uint32_t)h->c.V, g.st);
    b->own_launches++;
  }
  XRT_CUDA_OK(cudaEventRecord(g.ev_snap, g.st));
  XRT_CUDA_OK(cudaStreamWaitEvent(g.st_copy, g.ev_snap, 0));
  for (int i = 0; i < g.n; i++) {
    xrt_solver* h = b->h[g.first + i];
    if (!h->c.V) continue;
    XRT_CUDA_OK(cudaMemcpyAsync(b->h_snap[par][g.first + i], b->d_snap[g.first + i], (size_t)h->c.V * 6 * sizeof(float),
                                cudaMemcpyDeviceToHost, g.st_copy));
    XRT_CUDA_OK(cudaEventRecord(b->ev_dn[par][g.first + i], g.st_copy));
  }
  return XRT_OK;
}

// one frame of every garment (SimulateSwift, ClothSolverBase.cpp:171-247), each group in lockstep
static int batch_enqueue_frame(xrt_batch* b) {
  xrt_solver* s = b->h[0];
  XRT_REQUIRE(!b->readback || b->snap_written - b->snap_read < 2,
              "batch: two frames are waiting to be read back (xrt_batch_get_positions)");
  std::vector<cudaGraphExec_t> gr;
  for (BatchGroup& g : b->g) {
    XRT_CUDA_OK(cudaEventSynchronize(g.ev_flags));  // previous frame's flags have left the pinned buffer
    for (int i = 0; i < g.n; i++) {
      xrt_solver* h = b->h[g.first + i];
      if (h->touched) {  // per-handle calls since the last batch frame (targets, state): order after them
        XRT_CUDA_OK(cudaEventRecord(b->ev_pre[g.first + i], h->st));
        XRT_CUDA_OK(cudaStreamWaitEvent(g.st, b->ev_pre[g.first + i], 0));
        h->touched = false;
      }
      if (int rc = xrt_take_sequence_frame(h, g.st)) return rc;
      g.h_moved[i] = h->target_pending ? 1 : 0;
    }
    XRT_CUDA_OK(cudaMemcpyAsync(g.d_moved, g.h_moved, g.n * sizeof(int), cudaMemcpyHostToDevice, g.st));
    XRT_CUDA_OK(cudaEventRecord(g.ev_flags, g.st));
  }
  for (BatchGroup& g : b->g) g.seg_begin = 0;
  for (int sub = 0; sub <= b->S; sub++)  // sub == S: end of frame
    for (BatchGroup& g : b->g) {
      gr.assign(g.n, nullptr);
      bool any = sub == 0 || sub == b->S;
      if (g.uniform) {
        const xrt_solver* h0 = b->h[g.first];
        const bool hash = sub < b->S && h0->p.enable_self_collision && sub % std::max(1, h0->p.inter_leaved_hash) == 0;
        bool build = false;
        if (sub == 0)
          for (int i = 0; i < g.n; i++) build = build || (g.h_moved[i] && b->h[g.first + i]->o.F);
        if (!(hash || build || sub == b->S)) continue;  // this substep joins the running segment
        if (sub > g.seg_begin) {
          XRT_CUDA_OK(launch_swift_substep_batch(g.d_table, g.d_moved, (uint32_t)g.n, g.seg_begin, sub, b->cluster, g.st));
          b->own_launches++;
        }
        g.seg_begin = sub;
        if (build) {
          bvh_build_batched(g.bb, g.d_moved, h0->p.morton_fast_div, g.st);
          b->own_launches += 6;  // stage0, leaf boxes, morton, karras, boxes, links
          b->sort_calls++;
        }
        if (hash) {
          launch_predict_hash_batched(g.hb, g.st);
          b->own_launches += 4;  // predict + hash, cell clear, cell ranges, neighbour cache
          b->sort_calls++;
        }
        if (sub == b->S) {
          launch_post_batched(g.d_post, g.n_post, g.post_max_v, g.d_moved, g.st);
          b->own_launches += 2;  // copy, normals
          if (b->readback)
            if (int rc = batch_snapshot(b, g)) return rc;
          XRT_CUDA_OK(cudaEventRecord(g.ev_done, g.st));
          for (int i = 0; i < g.n; i++) {
            b->h[g.first + i]->target_pending = false;
            b->h[g.first + i]->batch_ev = g.ev_done;
          }
        }
        XRT_CUDA_OK(cudaGetLastError());
        continue;
      }
      for (int i = 0; i < g.n; i++) {
        xrt_solver* h = b->h[g.first + i];
        const int rc = sub < b->S ? batch_pre_graph(h, sub, g.h_moved[i], &gr[i]) : batch_post_graph(h, g.h_moved[i], &gr[i]);
        if (rc) {
          if (h != s) xrt_set_error(s, "batch: handle %d: %s", g.first + i, h->err);
          return rc;
        }
        any = any || gr[i];
      }
      if (!any) continue;  // this substep joins the running segment
      // substeps [seg_begin, sub) need nothing from outside: one launch
      if (sub > g.seg_begin) {
        XRT_CUDA_OK(launch_swift_substep_batch(g.d_table, g.d_moved, (uint32_t)g.n, g.seg_begin, sub, b->cluster, g.st));
        b->own_launches++;
      }
      g.seg_begin = sub;
      for (int i = 0; i < g.n; i++)  // per-garment graphs of a non-uniform group: the kernels their capture recorded
        if (gr[i]) {
          const xrt_solver* h = b->h[g.first + i];
          const bool hash_due = sub < b->S && h->p.enable_self_collision && sub % std::max(1, h->p.inter_leaved_hash) == 0;
          b->own_launches += sub == b->S ? (h->o.V ? 2u : 1u)
                                         : (hash_due ? 5u : 0u) + ((sub == 0 && g.h_moved[i] && h->o.F) ? 6u : 0u);
        }
      if (int rc = batch_fan_out(b, g, gr)) return rc;
      if (sub == b->S) {
        if (b->readback)
          if (int rc = batch_snapshot(b, g)) return rc;
        XRT_CUDA_OK(cudaEventRecord(g.ev_done, g.st));
        for (int i = 0; i < g.n; i++) {
          b->h[g.first + i]->target_pending = false;
          b->h[g.first + i]->batch_ev = g.ev_done;
        }
      }
    }
  if (b->readback) b->snap_written++;
  return XRT_OK;
}

extern "C" int xrt_batch_step_frame_async(xrt_batch* b) {
  if (!b || b->h.empty()) return XRT_ERR_INVALID;
  xrt_solver* s = b->h[0];
  XRT_CUDA_OK(cudaSetDevice(s->p.device));
  return batch_enqueue_frame(b);
}

// waits for every garment's end-of-frame work
extern "C" int xrt_batch_sync(xrt_batch* b) {
  if (!b || b->h.empty()) return XRT_ERR_INVALID;
  xrt_solver* s = b->h[0];
  XRT_CUDA_OK(cudaSetDevice(s->p.device));
  for (BatchGroup& g : b->g) {
    XRT_CUDA_OK(cudaStreamSynchronize(g.st));
    if (g.st_copy) XRT_CUDA_OK(cudaStreamSynchronize(g.st_copy));
  }
  return XRT_OK;
}

extern "C" int xrt_batch_step_frames(xrt_batch* b, int frames, float* ms) {
  if (!b || b->h.empty() || frames < 0) return XRT_ERR_INVALID;
  xrt_solver* s = b->h[0];
  XRT_CUDA_OK(cudaSetDevice(s->p.device));
  if (int rc = xrt_batch_sync(b)) return rc;
  for (xrt_solver* h : b->h)
    if (h->touched) XRT_CUDA_OK(cudaStreamSynchronize(h->st));  // keep per-handle leftovers out of the timed region
  cudaStream_t st0 = b->g[0].st;
  XRT_CUDA_OK(cudaEventRecord(b->ev0, st0));
  for (size_t gi = 1; gi < b->g.size(); gi++) XRT_CUDA_OK(cudaStreamWaitEvent(b->g[gi].st, b->ev0, 0));
  for (int f = 0; f < frames; f++)
    if (int rc = batch_enqueue_frame(b)) return rc;
  for (size_t gi = 1; gi < b->g.size(); gi++) {
    XRT_CUDA_OK(cudaEventRecord(b->g[gi].ev_fork, b->g[gi].st));
    XRT_CUDA_OK(cudaStreamWaitEvent(st0, b->g[gi].ev_fork, 0));
  }
  XRT_CUDA_OK(cudaEventRecord(b->ev1, st0));
  XRT_CUDA_OK(cudaStreamSynchronize(st0));
  if (ms) XRT_CUDA_OK(cudaEventElapsedTime(ms, b->ev0, b->ev1));
  return XRT_OK;
}

// SaveCache for the whole batch: rows of every handle back to back in batch order.  All packs and
// D2H copies are enqueued first, so the copy of one garment overlaps the memcpy-out of the previous.
extern "C" int xrt_batch_get_positions(xrt_batch* b, int which, float* xyz, float* nrm) {
  if (!b || b->h.empty() || (which != 0 && which != 1)) return XRT_ERR_INVALID;
  xrt_solver* s = b->h[0];
  XRT_CUDA_OK(cudaSetDevice(s->p.device));
  if (!xyz && !nrm) return XRT_OK;
  if (which == 0 && b->readback && b->snap_read < b->snap_written) {
    // oldest unread snapshot: wait for its copy garment by garment, memcpy out
    const int par = (int)(b->snap_read & 1ull);
    size_t row = 0;
    for (size_t i = 0; i < b->h.size(); i++) {
      const size_t n = b->h[i]->c.V;
      if (!n) continue;
      XRT_CUDA_OK(cudaEventSynchronize(b->ev_dn[par][i]));
      if (xyz) memcpy(xyz + 3 * row, b->h_snap[par][i], n * 3 * sizeof(float));
      if (nrm) memcpy(nrm + 3 * row, b->h_snap[par][i] + 3 * n, n * 3 * sizeof(float));
      row += n;
    }
    b->snap_read++;
    return XRT_OK;
  }
  for (BatchGroup& g : b->g)
    for (int i = 0; i < g.n; i++) {
      xrt_solver* h = b->h[g.first + i];
      const size_t n = which == 0 ? h->c.V : h->o.V;
      if (!n) continue;
      if (h->touched) {
        XRT_CUDA_OK(cudaEventRecord(b->ev_pre[g.first + i], h->st));
        XRT_CUDA_OK(cudaStreamWaitEvent(g.st, b->ev_pre[g.first + i], 0));
        h->touched = false;
      }
      launch_pack_xyz(xyz ? (which == 0 ? h->c.x0 : h->o.x0) : nullptr, nrm ? (which == 0 ? h->c.n : h->o.n) : nullptr,
                      h->d_dn, (uint32_t)n, g.st);
      XRT_CUDA_OK(cudaMemcpyAsync(h->h_dn, h->d_dn, n * 6 * sizeof(float), cudaMemcpyDeviceToHost, g.st));
      XRT_CUDA_OK(cudaEventRecord(b->ev_pre[g.first + i], g.st));
    }
  size_t row = 0;
  for (size_t i = 0; i < b->h.size(); i++) {
    xrt_solver* h = b->h[i];
    const size_t n = which == 0 ? h->c.V : h->o.V;
    if (!n) continue;
    XRT_CUDA_OK(cudaEventSynchronize(b->ev_pre[i]));
    if (xyz) memcpy(xyz + 3 * row, h->h_dn, n * 3 * sizeof(float));
    if (nrm) memcpy(nrm + 3 * row, h->h_dn + 3 * n, n * 3 * sizeof(float));
    row += n;
  }
  return XRT_OK;
}

// UpdateObstacleAnimationFrame for the whole batch: xyz = the handles' obstacle targets back to back
extern "C" int xrt_batch_set_obstacle_targets(xrt_batch* b, const float* xyz, size_t n_verts_total) {
  if (!b || b->h.empty() || !xyz) return XRT_ERR_INVALID;
  xrt_solver* s = b->h[0];
  XRT_CUDA_OK(cudaSetDevice(s->p.device));
  size_t total = 0;
  for (xrt_solver* h : b->h) total += h->o.V;
  XRT_REQUIRE(total == n_verts_total, "batch_set_obstacle_targets: vertex count differs from the batch's obstacles");
  size_t row = 0;
  for (BatchGroup& g : b->g) {
    for (int i = 0; i < g.n; i++) {
      xrt_solver* h = b->h[g.first + i];
      const size_t n = h->o.V;
      if (!n) continue;
      if (h->touched) {
        XRT_CUDA_OK(cudaEventRecord(b->ev_pre[g.first + i], h->st));
        XRT_CUDA_OK(cudaStreamWaitEvent(g.st, b->ev_pre[g.first + i], 0));
        h->touched = false;
      }
      XRT_CUDA_OK(cudaEventSynchronize(h->ev_up));  // the previous upload has left the pinned buffer
      memcpy(h->h_up, xyz + 3 * row, n * 3 * sizeof(float));
      XRT_CUDA_OK(cudaMemcpyAsync(h->d_up, h->h_up, n * 3 * sizeof(float), cudaMemcpyHostToDevice, g.st));
      XRT_CUDA_OK(cudaEventRecord(h->ev_up, g.st));
      launch_unpack_xyz(h->d_up, h->o.x, (uint32_t)n, g.st);
      b->own_launches++;
      h->target_pending = true;
      row += n;
    }
    XRT_CUDA_OK(cudaEventRecord(g.ev_done, g.st));
    for (int i = 0; i < g.n; i++) b->h[g.first + i]->batch_ev = g.ev_done;
  }
  return XRT_OK;
}

// kernels of this library the batch has enqueued since it was created, counted at the launch sites, and calls into cub's
// segmented sort (library kernels)
extern "C" int xrt_batch_launch_count(xrt_batch* b, uint64_t* own_kernels, uint64_t* sort_calls) {
  if (!b) return XRT_ERR_INVALID;
  if (own_kernels) *own_kernels = b->own_launches;
  if (sort_calls) *sort_calls = b->sort_calls;
  return XRT_OK;
}

// kernels one frame of the batch launches, from the batch's configuration (xrt_batch_launch_count counts them)
extern "C" int xrt_batch_launches_per_frame(xrt_batch* b, int moved, uint32_t* out) {
  if (!b || b->h.empty() || !out) return XRT_ERR_INVALID;
  uint32_t n = 0;
  for (const BatchGroup& g : b->g)  // shared launches: one per run of substeps without per-garment work in between
    for (int sub = 0; sub < b->S; sub++) {
      bool any = sub == 0;
      for (int i = 0; i < g.n && !any; i++) {
        const xrt_solver* h = b->h[g.first + i];
        any = h->p.enable_self_collision && sub % std::max(1, h->p.inter_leaved_hash) == 0;
      }
      n += any ? 1u : 0u;
    }
  for (const BatchGroup& g : b->g) {
    if (g.uniform) {  // group-wide stages: one launch per stage whatever the number of garments
      const xrt_solver* h0 = b->h[g.first];
      bool obstacle = false;
      for (int i = 0; i < g.n; i++) obstacle = obstacle || b->h[g.first + i]->o.F;
      if (moved && obstacle) n += (uint32_t)bvh_build_launches();
      for (int sub = 0; sub < b->S; sub++)
        if (h0->p.enable_self_collision && sub % std::max(1, h0->p.inter_leaved_hash) == 0) n += 5;  // predict+hash, sort (1), clear, ranges, neighbours
      n += 2;  // end of frame: copy, normals
      continue;
    }
    for (int i = 0; i < g.n; i++) {
      const xrt_solver* h = b->h[g.first + i];
      if (moved && h->o.F) n += (uint32_t)bvh_build_launches();
      for (int sub = 0; sub < b->S; sub++)
        if (h->p.enable_self_collision && sub % std::max(1, h->p.inter_leaved_hash) == 0) n += 6;  // predict + hash
      n += h->o.V ? 2 : 1;  // end-of-frame normals
    }
  }
  *out = n;
  return XRT_OK;
}

// `reps` rounds of shared launches of one substep (no hash, no build), every group side by side:
// the dominant kernel as the batch runs it
extern "C" int xrt_batch_profile(xrt_batch* b, int reps, float* avg_ms, double* bytes) {
  if (!b || b->h.empty() || reps <= 0 || !avg_ms || !bytes) return XRT_ERR_INVALID;
  xrt_solver* s = b->h[0];
  XRT_CUDA_OK(cudaSetDevice(s->p.device));
  if (int rc = xrt_batch_sync(b)) return rc;
  for (xrt_solver* h : b->h) XRT_CUDA_OK(cudaStreamSynchronize(h->st));
  const int sub = 1 % b->S;
  double by = 0;
  std::vector<PersistArgs*> d_tabs(b->g.size(), nullptr);
  int rc = XRT_OK;
  for (size_t gi = 0; gi < b->g.size() && rc == XRT_OK; gi++) {
    BatchGroup& g = b->g[gi];
    std::vector<PersistArgs> tab((size_t)2 * g.n);
    for (int i = 0; i < g.n; i++) {
      const xrt_solver* h = b->h[g.first + i];
      tab[i] = tab[(size_t)g.n + i] = xrt_make_persist_args(h, sub, h->o.F != 0, true);
      by += (xrt_frame_bytes(h, h->o.F != 0) - (h->o.F ? xrt_bvh_build_bytes(h) : 0.0)) / h->p.num_substeps;
    }
    if (dev_upload(&d_tabs[gi], tab) != cudaSuccess) rc = XRT_ERR_CUDA;
  }
  cudaStream_t st0 = b->g[0].st;
  for (int pass = 0; pass < 2 && rc == XRT_OK; pass++) {  // pass 0 warms up
    XRT_CUDA_OK(cudaEventRecord(b->ev0, st0));
    for (size_t gi = 0; gi < b->g.size(); gi++) {
      BatchGroup& g = b->g[gi];
      if (gi) XRT_CUDA_OK(cudaStreamWaitEvent(g.st, b->ev0, 0));
      cudaMemsetAsync(g.d_moved, 0, g.n * sizeof(int), g.st);
      for (int r = 0; r < (pass ? reps : 2); r++)
        launch_swift_substep_batch(d_tabs[gi], g.d_moved, (uint32_t)g.n, 0, 1, b->cluster, g.st);
      if (gi) {
        XRT_CUDA_OK(cudaEventRecord(g.ev_fork, g.st));
        XRT_CUDA_OK(cudaStreamWaitEvent(st0, g.ev_fork, 0));
      }
    }
    XRT_CUDA_OK(cudaEventRecord(b->ev1, st0));
    if (cudaStreamSynchronize(st0) != cudaSuccess) rc = XRT_ERR_CUDA;
  }
  for (PersistArgs* p : d_tabs) cudaFree(p);
  if (rc != XRT_OK) {
    xrt_set_error(s, "batch_profile: %s", cudaGetErrorString(cudaGetLastError()));
    return rc;
  }
  float ms = 0;
  XRT_CUDA_OK(cudaEventElapsedTime(&ms, b->ev0, b->ev1));
  *avg_ms = ms / (float)reps;
  *bytes = by;
  return XRT_OK;
}
