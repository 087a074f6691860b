// Device math and mesh accessors of the Quality-mode collision path: signed vertex-face / edge-edge
// distances, the cubic coplanarity solver, the continuous impact test, the imminent-repulsion
// quadratures and the PBD repulsion constraints.  IEEE arithmetic in the reference's evaluation
// order (this library is built with -fmad=false and without --use_fast_math), because every
// function here ends in a discrete decision (a root inside [0,1], a barycentric sign, a proximity
// threshold) that must come out the same as in the CPU oracle.
// Follows xrtailor/src/math/BasicPrimitiveTests.cu:286-301,369-438, math/MathFunctions.cu:85-146,
// physics/impact_zone/ImpactHelper.cu:34-114, physics/repulsion/ImminentRepulsionHelper.cu:10-150,
// physics/repulsion/PBDRepulsion.cu:329-468, physics/repulsion/Proximity.cu:10-36.
#pragma once
#include "xrt_common.cuh"

namespace xrtq {

// Both meshes of a handle behind one numbering: cloth node i is node i, obstacle node j is V + j;
// the same for faces (F) and edges (E).  The reference orders Node* / Face* by pool address, cloth
// first (MemoryPool, memory/MemoryPool.cuh:18), which is this order.
struct QMesh {
  uint32_t V, Vo, F, Fo, E, Eo;
  const float4 *cx0, *cx, *ox0, *ox;   // x0 = Node::x0, x = Node::x; .w = inverse mass (cloth)
  const float4* ov;                    // obstacle Node::v (SetupObstacleNode_Kernel, Entry.cu:117-124)
  const uint32_t *ctri, *otri;         // [3F] local vertex ids
  const uint32_t *cfe, *ofe;           // [3F] face -> its three edges (local edge ids), edge j = (v_j, v_j+1)
  const uint2 *cedge, *oedge;          // [E]
  const uint32_t *crtri, *ortri;       // representative-triangle masks
  const uint32_t *cef_off, *cef_idx, *oef_off, *oef_idx;  // edge -> adjacent faces (ascending), Edge::adjacents
  const float4 *cfn0, *ofn0;           // Face::n as set at registration (PhysicsMeshHelper.cu:118-123)
  float obstacle_inv_mass;             // 1 / 1e3, PhysicsMeshHelper.cu:139
};

__device__ __forceinline__ bool is_cloth(const QMesh& m, uint32_t g) { return g < m.V; }
__device__ __forceinline__ float4 ld_x0(const QMesh& m, uint32_t g) { return g < m.V ? m.cx0[g] : m.ox0[g - m.V]; }
__device__ __forceinline__ float4 ld_x(const QMesh& m, uint32_t g) { return g < m.V ? m.cx[g] : m.ox[g - m.V]; }
__device__ __forceinline__ f3 x0_of(const QMesh& m, uint32_t g) { return mk3(ld_x0(m, g)); }
__device__ __forceinline__ f3 x_of(const QMesh& m, uint32_t g) { return mk3(ld_x(m, g)); }
__device__ __forceinline__ float invm_of(const QMesh& m, uint32_t g) { return g < m.V ? m.cx0[g].w : m.obstacle_inv_mass; }
// Node::is_free: cloth nodes unless pinned (inv_mass 0), obstacle nodes never
__device__ __forceinline__ bool free_of(const QMesh& m, uint32_t g) { return g < m.V ? (m.cx0[g].w != 0.0f) : false; }
__device__ __forceinline__ uint32_t face_node(const QMesh& m, uint32_t gf, int k) {
  return gf < m.F ? m.ctri[3 * gf + k] : m.V + m.otri[3 * (gf - m.F) + k];
}
__device__ __forceinline__ uint32_t face_edge(const QMesh& m, uint32_t gf, int k) {
  return gf < m.F ? m.cfe[3 * gf + k] : m.E + m.ofe[3 * (gf - m.F) + k];
}
__device__ __forceinline__ uint2 edge_nodes(const QMesh& m, uint32_t ge) {
  if (ge < m.E) return m.cedge[ge];
  const uint2 e = m.oedge[ge - m.E];
  return make_uint2(m.V + e.x, m.V + e.y);
}
__device__ __forceinline__ uint32_t face_rtri(const QMesh& m, uint32_t gf) { return gf < m.F ? m.crtri[gf] : m.ortri[gf - m.F]; }
__device__ __forceinline__ bool rtri_vertex(uint32_t info, int n) { return (info >> n) & 1u; }      // RTriangle.cu:12-14
__device__ __forceinline__ bool rtri_edge(uint32_t info, int n) { return (info >> (n + 3)) & 1u; }  // RTriangle.cu:15-17
__device__ __forceinline__ bool face_free(const QMesh& m, uint32_t gf) {
  return free_of(m, face_node(m, gf, 0)) && free_of(m, face_node(m, gf, 1)) && free_of(m, face_node(m, gf, 2));
}
__device__ __forceinline__ bool edge_free(const QMesh& m, uint32_t ge) {
  const uint2 e = edge_nodes(m, ge);
  return free_of(m, e.x) && free_of(m, e.y);
}

__device__ __forceinline__ float mixed3(f3 a, f3 b, f3 c) { return dot3(a, cross3(b, c)); }  // MathFunctions.cuh:62-64
__device__ __forceinline__ float qmin(float a, float b) { return a < b ? a : b; }           // MathFunctions.cuh:24
__device__ __forceinline__ float qmax(float a, float b) { return a > b ? a : b; }
__device__ __forceinline__ float qabs(float x) { return x < 0.0f ? -x : x; }
__device__ __forceinline__ float qmin3(float a, float b, float c) { return qmin(a, qmin(b, c)); }
__device__ __forceinline__ float qmin4(float a, float b, float c, float d) { return qmin(qmin(a, b), qmin(c, d)); }

// SignedVertexFaceDistance, BasicPrimitiveTests.cu:369-388.  A degenerate face returns infinity with
// n = the raw cross product and w untouched.
__device__ __forceinline__ float signed_vf_distance(f3 x, f3 y0, f3 y1, f3 y2, f3& n, float w[4]) {
  n = cross3(normalize3(y1 - y0), normalize3(y2 - y0));
  if (dot3(n, n) < 1e-6f) return INFINITY;
  n = normalize3(n);
  const float h = dot3(x - y0, n);
  const float b0 = mixed3(y1 - x, y2 - x, n);
  const float b1 = mixed3(y2 - x, y0 - x, n);
  const float b2 = mixed3(y0 - x, y1 - x, n);
  w[0] = 1.0f;
  w[1] = -b0 / (b0 + b1 + b2);
  w[2] = -b1 / (b0 + b1 + b2);
  w[3] = -b2 / (b0 + b1 + b2);
  return h;
}

// SignedEdgeEdgeDistance, BasicPrimitiveTests.cu:422-438 (n is left un-normalised, as there)
__device__ __forceinline__ float signed_ee_distance(f3 x0, f3 x1, f3 y0, f3 y1, f3& n, float w[4]) {
  n = cross3(normalize3(x1 - x0), normalize3(y1 - y0));
  if (dot3(n, n) < 1e-6f) return INFINITY;
  const float h = dot3(x0 - y0, n);
  const float a0 = mixed3(y1 - x1, y0 - x1, n);
  const float a1 = mixed3(y0 - x0, y1 - x0, n);
  const float b0 = mixed3(x0 - y1, x1 - y1, n);
  const float b1 = mixed3(x1 - y0, x0 - y0, n);
  w[0] = a0 / (a0 + a1);
  w[1] = a1 / (a0 + a1);
  w[2] = -b0 / (b0 + b1);
  w[3] = -b1 / (b0 + b1);
  return h;
}

// NewtonsMethod, MathFunctions.cu:85-103
__device__ __forceinline__ float newtons_method(float a, float b, float c, float d, float x0, int dir) {
  if (dir != 0) {
    const float y0 = d + x0 * (c + x0 * (b + x0 * a));
    const float ddy0 = 2.0f * b + x0 * (6.0f * a);
    x0 += (float)dir * sqrtf(qabs(2.0f * y0 / ddy0));
  }
  for (int iter = 0; iter < 100; iter++) {
    const float y = d + x0 * (c + x0 * (b + x0 * a));
    const float dy = c + x0 * (2.0f * b + 3.0f * x0 * a);
    if (dy == 0) return x0;
    const float x1 = x0 - y / dy;
    if (qabs(x0 - x1) < 1e-6f) return x0;
    x0 = x1;
  }
  return x0;
}
// SolveQuadratic, MathFunctions.cu:105-120
__device__ __forceinline__ int solve_quadratic(float a, float b, float c, float x[2]) {
  const float d = b * b - 4.0f * a * c;
  if (d < 0) {
    x[0] = -b / (2.0f * a);
    return 0;
  }
  const float q = -(b + (float)(b < 0.0f ? -1 : 1) * sqrtf(d)) * 0.5f;
  int i = 0;
  if (qabs(a) > 1e-12f * qabs(q)) x[i++] = q / a;
  if (qabs(q) > 1e-12f * qabs(c)) x[i++] = c / q;
  if (i == 2 && x[0] > x[1]) {
    const float t = x[0];
    x[0] = x[1];
    x[1] = t;
  }
  return i;
}
// SolveCubic, MathFunctions.cu:122-146
__device__ __forceinline__ int solve_cubic(float a, float b, float c, float d, float x[3]) {
  float xc[2];
  const int n = solve_quadratic(3.0f * a, 2.0f * b, c, xc);
  if (n == 0) {
    x[0] = newtons_method(a, b, c, d, xc[0], 0);
    return 1;
  } else if (n == 1) {
    return solve_quadratic(b, c, d, x);
  }
  const float yc0 = d + xc[0] * (c + xc[0] * (b + xc[0] * a)), yc1 = d + xc[1] * (c + xc[1] * (b + xc[1] * a));
  int i = 0;
  if (yc0 * a >= 0) x[i++] = newtons_method(a, b, c, d, xc[0], -1);
  if (yc0 * yc1 <= 0) {
    const int closer = qabs(yc0) < qabs(yc1) ? 0 : 1;
    x[i++] = newtons_method(a, b, c, d, xc[closer], closer == 0 ? 1 : -1);
  }
  if (yc1 * a <= 0) x[i++] = newtons_method(a, b, c, d, xc[1], 1);
  return i;
}

// One continuous-collision record (impact_zone/Impact.cuh:8-18): global node ids, time of impact in
// [0,1], weights (VF: 1, -b1, -b2, -b3; EE: a0, a1, -b0, -b1) and the normal.
struct DImpact {
  uint32_t n[4];
  float t;
  float w[4];
  float nrm[3];
  uint32_t is_vf;
};

// CheckImpact, ImpactHelper.cu:34-87.  p0 / p1: Node::x0 / Node::x of the four nodes.  w and nrm keep what an
// earlier root wrote (the reference works on the Impact in place), starting from zero.
__device__ __forceinline__ bool check_impact(bool is_vf, const f3 p0[4], const f3 p1[4], float& t_out, float w[4],
                                             f3& nrm) {
  const f3 x0 = p0[0];
  const f3 v0 = p1[0] - x0;
  const f3 x1 = p0[1] - x0, x2 = p0[2] - x0, x3 = p0[3] - x0;
  const f3 v1 = (p1[1] - p0[1]) - v0, v2 = (p1[2] - p0[2]) - v0, v3 = (p1[3] - p0[3]) - v0;
  const float a0 = mixed3(x1, x2, x3);
  const float a1 = mixed3(v1, x2, x3) + mixed3(x1, v2, x3) + mixed3(x1, x2, v3);
  const float a2 = mixed3(x1, v2, v3) + mixed3(v1, x2, v3) + mixed3(v1, v2, x3);
  const float a3 = mixed3(v1, v2, v3);
  float t[3];
  const int ns = solve_cubic(a3, a2, a1, a0, t);
  for (int i = 0; i < ns; i++) {
    if (t[i] < 0 || t[i] > 1) continue;
    t_out = t[i];
    f3 q[4];
#pragma unroll
    for (int k = 0; k < 4; k++) q[k] = p0[k] + (p1[k] - p0[k]) * t[i];  // Node::Position, Node.cu:17-19
    float d;
    bool inside;
    if (is_vf) {
      d = signed_vf_distance(q[0], q[1], q[2], q[3], nrm, w);
      inside = qmin3(-w[1], -w[2], -w[3]) >= -1e-6f;
    } else {
      d = signed_ee_distance(q[0], q[1], q[2], q[3], nrm, w);
      inside = qmin4(w[0], w[1], -w[2], -w[3]) >= -1e-6f;
    }
    const f3 rel = (v1 * w[1] + v2 * w[2]) + v3 * w[3];
    if (dot3(nrm, rel) > 0) nrm = -nrm;
    if (qabs(d) < 1e-6f && inside) return true;
  }
  return false;
}

// Node/Edge/Face::ComputeBounds(true) (Node.cu:9-15): box over x and x0 of the listed nodes
struct QBox {
  float lo[3], hi[3];
};
__device__ __forceinline__ void box_init(QBox& b) {
  b.lo[0] = b.lo[1] = b.lo[2] = INFINITY;
  b.hi[0] = b.hi[1] = b.hi[2] = -INFINITY;
}
__device__ __forceinline__ void box_add(QBox& b, f3 p) {
  b.lo[0] = qmin(b.lo[0], p.x); b.lo[1] = qmin(b.lo[1], p.y); b.lo[2] = qmin(b.lo[2], p.z);
  b.hi[0] = qmax(b.hi[0], p.x); b.hi[1] = qmax(b.hi[1], p.y); b.hi[2] = qmax(b.hi[2], p.z);
}
// a.Overlap(b, thickness) = a.Overlap(b.Dilate(thickness)), Bounds.cu:95-128
__device__ __forceinline__ bool box_overlap(const QBox& a, const QBox& b, float th) {
#pragma unroll
  for (int k = 0; k < 3; k++) {
    if (a.lo[k] > b.hi[k] + th) return false;
    if (a.hi[k] < b.lo[k] - th) return false;
  }
  return true;
}

// The 15 tests of one face pair (CheckImpacts_Kernel, ImpactHelper.cu:116-171): slots 0-2 vertex i of f0 against
// f1, 3-5 vertex i of f1 against f0, 6-14 edge (s-6)/3 of f0 against edge (s-6)%3 of f1.  Writes the four node ids
// and says whether it is a VF test.
__device__ __forceinline__ bool slot_nodes(const QMesh& m, uint32_t f0, uint32_t f1, int slot, uint32_t n[4]) {
  if (slot < 6) {
    const uint32_t fv = slot < 3 ? f0 : f1, ff = slot < 3 ? f1 : f0;
    n[0] = face_node(m, fv, slot < 3 ? slot : slot - 3);
    n[1] = face_node(m, ff, 0); n[2] = face_node(m, ff, 1); n[3] = face_node(m, ff, 2);
    return true;
  }
  const uint2 a = edge_nodes(m, face_edge(m, f0, (slot - 6) / 3)), b = edge_nodes(m, face_edge(m, f1, (slot - 6) % 3));
  n[0] = a.x; n[1] = a.y; n[2] = b.x; n[3] = b.y;
  return false;
}
// The filters in front of CheckImpact: shared nodes and the ccd boxes of the two primitives
// (CheckVertexFaceImpact / CheckEdgeEdgeImpact, ImpactHelper.cu:89-114)
__device__ __forceinline__ bool slot_passes(const QMesh& m, bool is_vf, const uint32_t n[4], float th) {
  QBox a, b;
  box_init(a);
  box_init(b);
  if (is_vf) {
    if (n[0] == n[1] || n[0] == n[2] || n[0] == n[3]) return false;
    box_add(a, x_of(m, n[0])); box_add(a, x0_of(m, n[0]));
#pragma unroll
    for (int k = 1; k < 4; k++) { box_add(b, x_of(m, n[k])); box_add(b, x0_of(m, n[k])); }
  } else {
    if (n[0] == n[2] || n[0] == n[3] || n[1] == n[2] || n[1] == n[3]) return false;
#pragma unroll
    for (int k = 0; k < 2; k++) { box_add(a, x_of(m, n[k])); box_add(a, x0_of(m, n[k])); }
#pragma unroll
    for (int k = 2; k < 4; k++) { box_add(b, x_of(m, n[k])); box_add(b, x0_of(m, n[k])); }
  }
  return box_overlap(a, b, th);
}

// Edge::adjacents as global face ids (at most 4 kept, memory/Edge.cuh:30)
__device__ __forceinline__ int edge_adjacent(const QMesh& m, uint32_t ge, uint32_t out[4]) {
  const bool cl = ge < m.E;
  const uint32_t e = cl ? ge : ge - m.E;
  const uint32_t* off = cl ? m.cef_off : m.oef_off;
  const uint32_t* idx = cl ? m.cef_idx : m.oef_idx;
  int n = 0;
  for (uint32_t i = off[e]; i < off[e + 1] && n < 4; i++) out[n++] = (cl ? 0u : m.F) + idx[i];
  return n;
}
__device__ __forceinline__ f3 face_n0(const QMesh& m, uint32_t gf) { return mk3(gf < m.F ? m.cfn0[gf] : m.ofn0[gf - m.F]); }

// InEdge, BasicPrimitiveTests.cu:286-301 (current positions, registration-time face normals)
__device__ __forceinline__ bool in_edge(const QMesh& m, float w, uint32_t e0, uint32_t e1) {
  const uint2 a = edge_nodes(m, e0), b = edge_nodes(m, e1);
  const f3 x = x_of(m, a.x) * (1.0f - w) + x_of(m, a.y) * w;
  uint32_t adj[4];
  const int na = edge_adjacent(m, e1, adj);
  bool in = true;
  for (int i = 0; i < na; i++) {
    const f3 p0 = x_of(m, b.x), p1 = x_of(m, b.y);
    const f3 e = p1 - p0, r = x - p0;
    in &= (mixed3(e, face_n0(m, adj[i]), r) >= 0);
  }
  return in;
}

// CheckImminentVFQuadrature, ImminentRepulsionHelper.cu:10-47 (frame-start positions).  Output node order:
// face nodes, then the vertex; bary = w[1], w[2], w[3], w[0].
__device__ __forceinline__ bool vf_quadrature(const QMesh& m, uint32_t node, uint32_t gf, uint32_t out_n[4], float out_b[4]) {
  const uint32_t f[3] = {face_node(m, gf, 0), face_node(m, gf, 1), face_node(m, gf, 2)};
  if (node == f[0] || node == f[1] || node == f[2]) return false;
  f3 fn;
  float w[4] = {0, 0, 0, 0};
  const float d = signed_vf_distance(x0_of(m, node), x0_of(m, f[0]), x0_of(m, f[1]), x0_of(m, f[2]), fn, w);
  if (!(d < INFINITY)) return false;
  if (!(qmin3(-w[1], -w[2], -w[3]) >= 1e-6f)) return false;
  out_n[0] = f[0]; out_n[1] = f[1]; out_n[2] = f[2]; out_n[3] = node;
  out_b[0] = w[1]; out_b[1] = w[2]; out_b[2] = w[3]; out_b[3] = w[0];
  return true;
}
// CheckImminentEEQuadrature, ImminentRepulsionHelper.cu:49-80
__device__ __forceinline__ bool ee_quadrature(const QMesh& m, uint32_t e0, uint32_t e1, uint32_t out_n[4], float out_b[4]) {
  const uint2 a = edge_nodes(m, e0), b = edge_nodes(m, e1);
  const uint32_t n[4] = {a.x, a.y, b.x, b.y};
  if (n[0] == n[2] || n[0] == n[3] || n[1] == n[2] || n[1] == n[3]) return false;
  float w[4] = {0, 0, 0, 0};
  f3 fn;
  const float d = signed_ee_distance(x0_of(m, n[0]), x0_of(m, n[1]), x0_of(m, n[2]), x0_of(m, n[3]), fn, w);
  if (!(d < INFINITY)) return false;
  const bool inside = qmin4(w[0], w[1], -w[2], -w[3]) >= 1e-6f && in_edge(m, w[1], e0, e1) && in_edge(m, -w[3], e1, e0);
  if (!inside) return false;
#pragma unroll
  for (int i = 0; i < 4; i++) { out_n[i] = n[i]; out_b[i] = w[i]; }
  return true;
}

// Edge::ComputeNormal, Edge.cu:43-58 (current positions)
__device__ __forceinline__ f3 edge_normal(const QMesh& m, uint32_t ge) {
  uint32_t adj[4];
  const int na = edge_adjacent(m, ge, adj);
  f3 ans = mk3(0, 0, 0);
  for (int i = 0; i < na; i++)
    ans = ans + face_normal(x_of(m, face_node(m, adj[i], 0)), x_of(m, face_node(m, adj[i], 1)), x_of(m, face_node(m, adj[i], 2)));
  if (na == 0) return ans;
  return ans / (float)na;
}

// EvaluateVFProximity, PBDRepulsion.cu:329-409: nodes = vertex, face nodes; corr = position corrections
__device__ __forceinline__ bool rep_vf(const QMesh& m, uint32_t node, uint32_t gf, float h, uint32_t nodes[4], f3 corr[4]) {
  const uint32_t f[3] = {face_node(m, gf, 0), face_node(m, gf, 1), face_node(m, gf, 2)};
  if (node == f[0] || node == f[1] || node == f[2]) return false;
  nodes[0] = node; nodes[1] = f[0]; nodes[2] = f[1]; nodes[3] = f[2];
  // Proximity(node, face, 1), Proximity.cu:10-20: the side is decided on the frame-start positions
  f3 pn;
  float w0[4] = {0, 0, 0, 0};
  const float d0 = signed_vf_distance(x0_of(m, nodes[0]), x0_of(m, nodes[1]), x0_of(m, nodes[2]), x0_of(m, nodes[3]), pn, w0);
  if (d0 < 0.0f && face_free(m, gf)) pn = -pn;
  f3 fn;
  float w[4] = {0, 0, 0, 0};
  f3 q = x_of(m, nodes[0]), p1 = x_of(m, nodes[1]), p2 = x_of(m, nodes[2]), p3 = x_of(m, nodes[3]);
  const float dd = signed_vf_distance(q, p1, p2, p3, fn, w);
  if (!(dd < INFINITY)) return false;
  if (!(qmin3(-w[1], -w[2], -w[3]) >= 1e-6f)) return false;
  f3 n = face_normal(p1, p2, p3);
  const float side = dot3(n, pn) > 0.0f ? 1.0f : -1.0f;
  q = q - p1; p2 = p2 - p1; p3 = p3 - p1;
  n = cross3(p2, p3) * side;
  const float c23 = len3(n);
  n = normalize3(n);
  const float C = dot3(q, n) - h;
  if (C > 0) return false;
  const f3 dcq = n;
  const f3 dcp2 = (cross3(p3, q) + cross3(n, p3) * dot3(n, q)) / c23;
  const f3 dcp3 = -((cross3(p2, q) + cross3(n, p2) * dot3(n, q)) / c23);
  const f3 dcp1 = ((-dcq) - dcp2) - dcp3;
  const float w_[4] = {invm_of(m, nodes[0]), invm_of(m, nodes[1]), invm_of(m, nodes[2]), invm_of(m, nodes[3])};
  const float denom = w_[0] * dot3(dcq, dcq) + w_[1] * dot3(dcp1, dcp1) + w_[2] * dot3(dcp2, dcp2) + w_[3] * dot3(dcp3, dcp3);
  if (denom < XRT_EPS) return false;
  const float sc = C / denom;
  corr[0] = dcq * (-w_[0] * sc);
  corr[1] = dcp1 * (-w_[1] * sc);
  corr[2] = dcp2 * (-w_[2] * sc);
  corr[3] = dcp3 * (-w_[3] * sc);
  return true;
}
// EvaluateEEProximity, PBDRepulsion.cu:411-468
__device__ __forceinline__ bool rep_ee(const QMesh& m, uint32_t e0, uint32_t e1, float h, uint32_t nodes[4], f3 corr[4]) {
  const uint2 a = edge_nodes(m, e0), bq = edge_nodes(m, e1);
  nodes[0] = a.x; nodes[1] = a.y; nodes[2] = bq.x; nodes[3] = bq.y;
  if (nodes[0] == nodes[2] || nodes[0] == nodes[3] || nodes[1] == nodes[2] || nodes[1] == nodes[3]) return false;
  // Proximity(edge0, edge1, 1), Proximity.cu:22-36
  f3 pn;
  float w0[4] = {0, 0, 0, 0};
  const float d0 = signed_ee_distance(x0_of(m, nodes[0]), x0_of(m, nodes[1]), x0_of(m, nodes[2]), x0_of(m, nodes[3]), pn, w0);
  if (!edge_free(m, e1)) pn = edge_normal(m, e1);
  else if (d0 < 0.0f) pn = -pn;
  float b[4] = {0, 0, 0, 0};
  f3 n;
  const f3 px[4] = {x_of(m, nodes[0]), x_of(m, nodes[1]), x_of(m, nodes[2]), x_of(m, nodes[3])};
  const float dd = signed_ee_distance(px[0], px[1], px[2], px[3], n, b);
  if (!(dd < INFINITY)) return false;
  const bool inside = qmin4(b[0], b[1], -b[2], -b[3]) >= 1e-6f && in_edge(m, b[1], e0, e1) && in_edge(m, -b[3], e1, e0);
  if (!inside) return false;
  f3 d = mk3(0, 0, 0);
#pragma unroll
  for (int i = 0; i < 4; i++) d = d + px[i] * b[i];
  n = pn;
  const float C = dot3(d, n) - h;
  if (C > 0) return false;
  b[2] *= -1; b[3] *= -1;
  const f3 grads[4] = {n * b[0], n * b[1], n * (-b[2]), n * (-b[3])};
  float denom = 0;
#pragma unroll
  for (int i = 0; i < 4; i++) denom += invm_of(m, nodes[i]) * b[i] * b[i];
  if (denom < XRT_EPS) return false;
  const float sc = C / denom;
#pragma unroll
  for (int i = 0; i < 4; i++) corr[i] = grads[i] * (-sc * invm_of(m, nodes[i]));
  return true;
}

}  // namespace xrtq
