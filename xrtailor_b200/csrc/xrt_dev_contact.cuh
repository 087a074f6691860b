// Device bodies of the Swift obstacle-contact stages: stackless BVH queries
// fused with their narrow phase, and the contact solves.  Shared by the
// per-stage kernels (xrt_contact.cu) and the persistent kernel (xrt_persist.cu).
// The reference fills a 256-entry per-thread local buffer and then loops over
// it (PBDCollision.cu:17-47,100-153,236-268; queries BVH.cu:957-1055); here each
// leaf hit is consumed as the traversal reaches it, in the same order and with
// the same 256-hit cap, so the selected contact is the same and no local memory
// is touched.  One 32-byte TNode load per visited node.
#pragma once
#include "xrt_solver.cuh"

namespace xrtdev {

struct Box {
  float lx, ly, lz, ux, uy, uz;
};
struct NodeView {
  Box b;
  int hit, miss;
};
// NC: non-coherent (read-only cache) loads, legal only when no thread of the
// launch writes the nodes; the persistent kernel refits them and passes false.
template <bool NC>
__device__ __forceinline__ NodeView load_node(const TNode* t, int i) {
  const float4* q = reinterpret_cast<const float4*>(t + i);
  const float4 a = NC ? __ldg(q) : q[0];
  const float4 c = NC ? __ldg(q + 1) : q[1];
  NodeView n;
  n.b = Box{a.x, a.y, a.z, a.w, c.x, c.y};
  n.hit = __float_as_int(c.z);
  n.miss = __float_as_int(c.w);
  return n;
}

// Bounds::Overlap(const Vector3&), Bounds.cu:126-131 (strict)
__device__ __forceinline__ bool overlap_point(const Box& b, f3 p) {
  return p.x > b.lx && p.x < b.ux && p.y > b.ly && p.y < b.uy && p.z > b.lz && p.z < b.uz;
}
// GetIntersection / InBox, Bounds.cu:132-150
__device__ __forceinline__ bool get_isect(float d1, float d2, f3 p1, f3 p2, f3& hit) {
  if (d1 * d2 >= 0.f) return false;
  if (d1 == d2) return false;
  hit = p1 + (p2 - p1) * (-d1 / (d2 - d1));
  return true;
}
__device__ __forceinline__ bool in_box(f3 h, const Box& b, int axis) {
  if (axis == 1 && h.z > b.lz && h.z < b.uz && h.y > b.ly && h.y < b.uy) return true;
  if (axis == 2 && h.z > b.lz && h.z < b.uz && h.x > b.lx && h.x < b.ux) return true;
  if (axis == 3 && h.x > b.lx && h.x < b.ux && h.y > b.ly && h.y < b.uy) return true;
  return false;
}
// Bounds::Overlap(p1, p2, hit), Bounds.cu:151-183
__device__ __forceinline__ bool overlap_segment(const Box& b, f3 p1, f3 p2) {
  if (p2.x < b.lx && p1.x < b.lx) return false;
  if (p2.x > b.ux && p1.x > b.ux) return false;
  if (p2.y < b.ly && p1.y < b.ly) return false;
  if (p2.y > b.uy && p1.y > b.uy) return false;
  if (p2.z < b.lz && p1.z < b.lz) return false;
  if (p2.z > b.uz && p1.z > b.uz) return false;
  if (overlap_point(b, p1)) return true;
  f3 h;
  return (get_isect(p1.x - b.lx, p2.x - b.lx, p1, p2, h) && in_box(h, b, 1)) ||
         (get_isect(p1.y - b.ly, p2.y - b.ly, p1, p2, h) && in_box(h, b, 2)) ||
         (get_isect(p1.z - b.lz, p2.z - b.lz, p1, p2, h) && in_box(h, b, 3)) ||
         (get_isect(p1.x - b.ux, p2.x - b.ux, p1, p2, h) && in_box(h, b, 1)) ||
         (get_isect(p1.y - b.uy, p2.y - b.uy, p1, p2, h) && in_box(h, b, 2)) ||
         (get_isect(p1.z - b.uz, p2.z - b.uz, p1, p2, h) && in_box(h, b, 3));
}

// The same test with the segment's own min/max corners precomputed once per
// query: "both endpoints below lo" is "the larger endpoint below lo", so the six
// slab rejections cost six compares per node instead of twelve.
struct SegQuery {
  f3 p1, p2, lo, hi;
  __device__ __forceinline__ SegQuery(f3 a, f3 b) : p1(a), p2(b) {
    lo = mk3(fminf(a.x, b.x), fminf(a.y, b.y), fminf(a.z, b.z));
    hi = mk3(fmaxf(a.x, b.x), fmaxf(a.y, b.y), fmaxf(a.z, b.z));
  }
  __device__ __forceinline__ bool overlaps(const Box& b) const {
    if (hi.x < b.lx || lo.x > b.ux || hi.y < b.ly || lo.y > b.uy || hi.z < b.lz || lo.z > b.uz) return false;
    if (overlap_point(b, p1)) return true;
    f3 h;
    return (get_isect(p1.x - b.lx, p2.x - b.lx, p1, p2, h) && in_box(h, b, 1)) ||
           (get_isect(p1.y - b.ly, p2.y - b.ly, p1, p2, h) && in_box(h, b, 2)) ||
           (get_isect(p1.z - b.lz, p2.z - b.lz, p1, p2, h) && in_box(h, b, 3)) ||
           (get_isect(p1.x - b.ux, p2.x - b.ux, p1, p2, h) && in_box(h, b, 1)) ||
           (get_isect(p1.y - b.uy, p2.y - b.uy, p1, p2, h) && in_box(h, b, 2)) ||
           (get_isect(p1.z - b.uz, p2.z - b.uz, p1, p2, h) && in_box(h, b, 3));
  }
};

// QueryDeviceStackless (BVH.cu:957-1055) with the leaf visitor inlined.
// Returns the number of leaf hits (capped at XRT_MAX_HITS like the reference).
// "while-while" form: every lane first walks to its next leaf hit (or the end), then the
// warp runs the leaf visitor together -- the heavy primitive tests execute with most lanes
// active instead of only the lanes that happen to stand on a leaf in this iteration.
// Same visits in the same order per lane as the reference's single loop.
template <bool NC, class Pred, class Visit>
__device__ __forceinline__ uint32_t traverse(const TNode* t, Pred pred, Visit visit) {
  int node = 0;
  uint32_t found = 0;
  while (node != -1) {
    int leaf = -1;
    while (node != -1) {
      const NodeView n = load_node<NC>(t, node);
      if (found >= XRT_MAX_HITS) {
        node = -1;
        break;
      }
      if (pred(n.b)) {
        if (n.hit < 0) {
          leaf = ~n.hit;
          node = n.miss;
          break;
        }
        node = n.hit;
      } else {
        node = n.miss;
      }
    }
    if (leaf >= 0) {
      visit((uint32_t)leaf);
      found++;
    }
  }
  return found;
}

// BasicPrimitiveTests::NearestPoint, math/BasicPrimitiveTests.cu:6-84
__device__ __forceinline__ float nearest_point(f3 p, f3 a, f3 b, f3 c, f3& qs) {
  const f3 e1 = b - a, e2 = c - a, e3 = c - b;
  const f3 v1 = p - a;
  const float d1 = dot3(e1, v1), d2 = dot3(e2, v1);
  if (d1 <= 0 && d2 <= 0) { qs = a; return len3(p - qs); }
  const f3 v2 = p - b;
  const float d3 = dot3(e1, v2), d4 = dot3(e2, v2);
  if (d3 >= 0 && d4 <= d3) { qs = b; return len3(p - qs); }
  const float vc = d1 * d4 - d3 * d2;
  if (vc <= 0 && d1 >= 0 && d3 <= 0) {
    const float v = d1 / (d1 - d3);
    qs = a + v * e1;
    return len3(p - qs);
  }
  const f3 v3 = p - c;
  const float d5 = dot3(e1, v3), d6 = dot3(e2, v3);
  if (d6 >= 0 && d5 <= d6) { qs = c; return len3(p - qs); }
  const float vb = d5 * d2 - d1 * d6;
  if (vb <= 0 && d2 >= 0 && d6 <= 0) {
    const float w = d2 / (d2 - d6);
    qs = a + w * e2;
    return len3(p - qs);
  }
  const float va = d3 * d6 - d5 * d4;
  if (va <= 0 && (d4 - d3) >= 0 && (d5 - d6) >= 0) {
    const float w = (d4 - d3) / ((d4 - d3) + (d5 - d6));
    qs = b + w * e3;
    return len3(p - qs);
  }
  const float d = 1 / (va + vb + vc);
  const float v = vb * d, w = vc * d;
  qs = a + e1 * v + e2 * w;
  return len3(p - qs);
}

// BasicPrimitiveTests::RayIntersect(o, d, len, Primitive), BasicPrimitiveTests.cu:242-284
__device__ __forceinline__ float ray_intersect(f3 o, f3 d, float ray_length, f3 a, f3 b, f3 c) {
  const f3 e1 = b - a, e2 = c - a;
  const f3 n = cross3(d, e2);
  const float det = dot3(e1, n);
  if (fabsf(det) < XRT_EPS) return XRT_SCALAR_MAX;
  const float inv = 1 / det;
  const f3 t = o - a;
  const float u = dot3(t, n) * inv;
  if (u < 0 || u > 1) return XRT_SCALAR_MAX;
  const f3 q = cross3(t, e1);
  const float v = dot3(d, q) * inv;
  if (v < 0 || v + u > 1) return XRT_SCALAR_MAX;
  const float s = dot3(e2, q) * inv;
  if (s > 0 && s < ray_length) return s;
  return INFINITY;
}

// BasicPrimitiveTests::ClosestPtSegmentSegment, BasicPrimitiveTests.cu:303-361
__device__ __forceinline__ float closest_seg_seg(f3 p1, f3 q1, f3 p2, f3 q2, float& s, float& t, f3& c1, f3& c2) {
  const f3 d1 = q1 - p1, d2 = q2 - p2, r = p1 - p2;
  const float a = dot3(d1, d1), e = dot3(d2, d2), f = dot3(d2, r);
  if (a <= XRT_EPS && e <= XRT_EPS) {
    s = t = 0;
    c1 = p1;
    c2 = p2;
    return len3(c1 - c2);
  }
  if (a <= XRT_EPS) {
    s = 0;
    t = clampf(f / e, 0.f, 1.f);
  } else {
    const float c = dot3(d1, r);
    if (e <= XRT_EPS) {
      t = 0;
      s = clampf(-c / a, 0.f, 1.f);
    } else {
      const float b = dot3(d1, d2);
      const float denom = a * e - b * b;
      s = denom > XRT_EPS ? clampf((b * f - c * e) / denom, 0.f, 1.f) : 0.f;
      t = (b * s + f) / e;
      if (t < XRT_EPS) {
        t = 0;
        s = clampf(-c / a, 0.f, 1.f);
      } else if (t > 1) {
        t = 1;
        s = clampf((b - c) / a, 0.f, 1.f);
      }
    }
  }
  c1 = p1 + d1 * s;
  c2 = p2 + d2 * t;
  return len3(c1 - c2);
}

// GenerateVF_DCD_Kernel + GenerateVFCCD_Kernel (PBDCollision.cu:17-47,236-268):
// both per-vertex queries for one vertex (x and x0 are loaded once).  Each leaf
// hit goes through its primitive test as the traversal reaches it.  (Measured
// alternatives that lost on B200, see DESIGN.md section 6: staging hits in shared
// memory and testing them after the traversal; two traversals per thread in lock
// step; compacting the items that pass the root box into a dense list.)
// Returns whether the vertex has any contact (either query).
template <bool NC>
__device__ __forceinline__ bool gen_vertex_item(const TNode* t, const float4* pred, const float4* x0, const float4* x,
                                                VFContact* vf, VFContact* rt, uint32_t* cnt_vf, uint32_t* cnt_rt,
                                                uint32_t i) {
  const f3 p = mk3(x[i]);
  bool live_vf;
  {
    float best = XRT_SCALAR_MAX;
    int best_idx = -1;
    f3 best_q = mk3(0, 0, 0);
    const uint32_t nf = traverse<NC>(
        t, [&](const Box& b) { return overlap_point(b, p); },
        [&](uint32_t o) {
          f3 qs;
          const float d = nearest_point(p, mk3(pred[3 * o]), mk3(pred[3 * o + 1]), mk3(pred[3 * o + 2]), qs);
          if (d < best) {
            best = d;
            best_q = qs;
            best_idx = (int)o;
          }
        });
    vf[i] = VFContact{best_idx, {best_q.x, best_q.y, best_q.z}};
    live_vf = best_idx != -1;
    if (cnt_vf) cnt_vf[i] = nf;
  }
  {
    const f3 p1 = mk3(x0[i]);
    f3 d = p - p1;
    const float length = len3(d);
    VFContact out{-1, {0.f, 0.f, 0.f}};
    uint32_t nf = 0;
    if (!(length < XRT_EPS)) {
      d = d / length;
      float best = XRT_SCALAR_MAX;
      const SegQuery seg(p1, p);
      nf = traverse<NC>(
          t, [&](const Box& b) { return seg.overlaps(b); },
          [&](uint32_t o) {
            const float dist = ray_intersect(p1, d, length, mk3(pred[3 * o]), mk3(pred[3 * o + 1]), mk3(pred[3 * o + 2]));
            if (dist < best) {
              const f3 qc = p1 + dist * d;
              out = VFContact{(int)o, {qc.x, qc.y, qc.z}};
              best = dist;
            }
          });
    }
    rt[i] = out;
    if (cnt_rt) cnt_rt[i] = nf;
    return live_vf || out.index != -1;
  }
}

// GenerateEE_DCD_Kernel, PBDCollision.cu:100-153.  Returns whether a contact was
// found; only then is ee[i] written (the active bit, not the record, says "none").
template <bool NC>
__device__ __forceinline__ bool gen_edge_item(const TNode* t, const float4* pred, const uint32_t* rtri,
                                              const float4* x, const uint2* edge, EEContact* ee, uint32_t* cnt_ee,
                                              uint32_t i) {
  const uint2 e = edge[i];
  const f3 pc = mk3(x[e.x]), qc = mk3(x[e.y]);
  float s = 0.f, tt = 0.f, min_dist = XRT_SCALAR_MAX;
  uint32_t best_ok = 0xFFFFFFFFu;  // (face << 2) | edge k of the current minimum
  const SegQuery seg(pc, qc);
  const uint32_t nf = traverse<NC>(
      t, [&](const Box& b) { return seg.overlaps(b); },
      [&](uint32_t o) {
        const uint32_t info = rtri[o] >> 3;  // RTriEdge bits
        if (!info) return;
        const f3 ov[3] = {mk3(pred[3 * o]), mk3(pred[3 * o + 1]), mk3(pred[3 * o + 2])};
#pragma unroll
        for (uint32_t k = 0; k < 3; k++) {
          if ((info >> k) & 1u) {
            float _s, _t;
            f3 c1, c2;
            const uint32_t j = (k + 1u) % 3u;
            // the reference tests the returned distance as a bool: non-zero distance only
            if (closest_seg_seg(pc, qc, ov[k], ov[j], _s, _t, c1, c2) != 0.f) {
              const float dist = len3(c1 - c2);
              if (dist < min_dist) {
                min_dist = dist;
                s = _s;
                tt = _t;
                best_ok = (o << 2) | k;
              }
            }
          }
        }
      });
  if (cnt_ee) cnt_ee[i] = nf;
  if (!(min_dist < 1e6f)) return false;
  const uint32_t o = best_ok >> 2, k = best_ok & 3u, j = (k + 1u) % 3u;
  const float4 a0 = pred[3 * o], a1 = pred[3 * o + 1], a2 = pred[3 * o + 2];
  // the winning face's normal, once (the reference recomputes it on every improvement)
  const f3 fn = face_normal(mk3(a0), mk3(a1), mk3(a2));
  const float4 p2 = k == 0 ? a0 : (k == 1 ? a1 : a2), q2 = j == 0 ? a0 : (j == 1 ? a1 : a2);
  EEContact c;
  c.dist = min_dist;
  c.p2[0] = p2.x; c.p2[1] = p2.y; c.p2[2] = p2.z;
  c.q2[0] = q2.x; c.q2[1] = q2.y; c.q2[2] = q2.z;
  c.s = s; c.t = tt;
  c.fn[0] = fn.x; c.fn[1] = fn.y; c.fn[2] = fn.z;
  ee[i] = c;
  return true;
}

// SolveVFCCD_Kernel then SolveVFDCD_Kernel (PBDCollision.cu:279-297,57-75) on
// the same vertex, in the reference's order, one load/store of x.
template <class X>
__device__ __forceinline__ void solve_vertex_item(const float4* pred, X x, const VFContact* vf,
                                                  const VFContact* rt, uint32_t i) {
  const VFContact c = rt[i], d = vf[i];
  if (c.index == -1 && d.index == -1) return;
  float4 p = x[i];
  bool dirty = false;
  if (c.index != -1) {
    const f3 nc = xf_face_normal(mk3(pred[3 * c.index]), mk3(pred[3 * c.index + 1]), mk3(pred[3 * c.index + 2]));
    const float lambda = xf_dot(nc, mk3(p) - mk3(c.q[0], c.q[1], c.q[2]));
    if (lambda < 0) {
      p.x += -lambda * nc.x; p.y += -lambda * nc.y; p.z += -lambda * nc.z;
      dirty = true;
    }
  }
  if (d.index != -1) {
    const f3 ns = xf_face_normal(mk3(pred[3 * d.index]), mk3(pred[3 * d.index + 1]), mk3(pred[3 * d.index + 2]));
    const float lambda = xf_dot(ns, mk3(p) - mk3(d.q[0], d.q[1], d.q[2])) - 1e-3f;
    if (lambda < 0) {
      p.x += -lambda * ns.x; p.y += -lambda * ns.y; p.z += -lambda * ns.z;
      dirty = true;
    }
  }
  if (dirty) x[i] = p;
}

// SolveEEDCD_Kernel, PBDCollision.cu:163-207.  `active` is the bit the
// generation pass recorded for this edge (dist < 1e6); inactive records are
// not stored, so they must not be read.
template <class X>
__device__ __forceinline__ void solve_edge_item(X x, const uint2* edge, const EEContact* ee,
                                                float4* delta, uint32_t i) {
  const EEContact c = ee[i];
  const uint2 e = edge[i];
  const f3 p1 = mk3(x[e.x]), q1 = mk3(x[e.y]);
  const f3 p2 = mk3(c.p2[0], c.p2[1], c.p2[2]), q2 = mk3(c.q2[0], c.q2[1], c.q2[2]);
  const f3 fn = mk3(c.fn[0], c.fn[1], c.fn[2]);
  const f3 c1 = (1 - c.s) * p1 + c.s * q1;
  const f3 c2 = (1 - c.t) * p2 + c.t * q2;
  const f3 d = c1 - c2;
  if (xf_len(d) < XRT_EPS) return;
  const float lambda = xf_dot(d, fn) - 1e-3f;
  if (lambda > 0) return;
  const f3 k1 = (-(1 - c.s) * lambda) * fn, k2 = (-(c.s) * lambda) * fn;
  atomicAdd(&delta[e.x].x, k1.x); atomicAdd(&delta[e.x].y, k1.y); atomicAdd(&delta[e.x].z, k1.z);
  atomicAdd(&delta[e.x].w, 1.0f);
  atomicAdd(&delta[e.y].x, k2.x); atomicAdd(&delta[e.y].y, k2.y); atomicAdd(&delta[e.y].z, k2.z);
  atomicAdd(&delta[e.y].w, 1.0f);
}

// Bit i of these masks = "item i has a live contact"; one word per 32
// consecutive items, written by lane 0 from a warp ballot in the generation
// pass (callers keep 32-item groups warp-aligned).  The solves read one
// broadcast word per warp instead of a 48-byte record per edge.
// Must be called by all 32 lanes of the warp (lanes past the end pass live = false);
// i = this lane's item, a multiple of 32 on lane 0.
__device__ __forceinline__ void store_active_bits(uint32_t* mask, uint32_t i, uint32_t n, bool live) {
  const uint32_t b = __ballot_sync(0xFFFFFFFFu, live);
  if ((threadIdx.x & 31u) == 0 && i < n) mask[i >> 5] = b;
}
__device__ __forceinline__ bool load_active_bit(const uint32_t* mask, uint32_t i) {
  return (mask[i >> 5] >> (i & 31u)) & 1u;
}

}  // namespace xrtdev
