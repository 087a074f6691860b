// Setup-time host work behind the C ABI: everything the reference derives from an index buffer before the
// solver starts, in O(n) or O(n log n) -- unique edges, face -> edge table, edge -> face adjacency, stretch and
// bend constraints (GenerateBendingElements / PrepareBasicStretchData / BuildEFAdjacency,
// xrtailor/src/physics/ClothSolverHelper.cpp:4-125, which walk vcglib's half-edge structure), deterministic
// constraint colouring (graph_coloring/Base.cu:163-317 is a random palette over an O(n^2) adjacency table),
// vertex -> face adjacency (ClothSolverHelper.cu:377-401) and the representative-triangle masks
// (representative_triangle/RTriangle.cu:19-74).
#include "xrt_handle.cuh"

#include <algorithm>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <vector>

// ---------------------------------------------------------------------------
// setup-time host work
// ---------------------------------------------------------------------------

// Deterministic greedy colouring in O(n): a 128-bit mask per vertex holds the
// colours already used by constraints on that vertex.  Replaces the random
// palette colouring + O(n^2) adjacency build of graph_coloring/Base.cu:163-317
// and impl/StretchColoring.cu:6-24; any proper colouring is a valid
// Gauss-Seidel order, and the reference's own order is random per run (F9).
extern "C" int xrt_greedy_color(const uint32_t* idx, size_t n, int arity, uint32_t n_verts, uint32_t* colors,
                                uint32_t* n_colors) {
  if ((n && (!idx || !colors)) || (arity < 2 || arity > 4)) return XRT_ERR_INVALID;
  std::vector<uint64_t> lo(n_verts, 0), hi(n_verts, 0);
  uint32_t used = 0;
  for (size_t c = 0; c < n; c++) {
    uint64_t ml = 0, mh = 0;
    for (int a = 0; a < arity; a++) {
      const uint32_t v = idx[c * arity + a];
      if (v >= n_verts) return XRT_ERR_INVALID;
      ml |= lo[v];
      mh |= hi[v];
    }
    uint32_t col;
    if (~ml) col = (uint32_t)__builtin_ctzll(~ml);
    else if (~mh) col = 64u + (uint32_t)__builtin_ctzll(~mh);
    else return XRT_ERR_INVALID;  // > 128 colours: vertex degree beyond anything a triangle mesh produces
    for (int a = 0; a < arity; a++) {
      const uint32_t v = idx[c * arity + a];
      if (col < 64) lo[v] |= 1ull << col;
      else hi[v] |= 1ull << (col - 64);
    }
    colors[c] = col;
    used = std::max(used, col + 1);
  }
  if (n_colors) *n_colors = used;
  return XRT_OK;
}

// stable counting sort of constraints by colour; returns offsets[n_colors+1]
// and whether any colour touches a vertex twice (then the kernels use atomics
// like the reference does unconditionally)
void xrt_color_ranges(const std::vector<uint32_t>& idx, int arity, const std::vector<uint32_t>& color,
                         uint32_t n_verts, std::vector<uint32_t>& perm, std::vector<uint32_t>& off, bool& improper) {
  const size_t n = color.size();
  uint32_t nc = 0;
  for (uint32_t c : color) nc = std::max(nc, c + 1);
  off.assign(nc + 1, 0);
  for (uint32_t c : color) off[c + 1]++;
  for (uint32_t c = 0; c < nc; c++) off[c + 1] += off[c];
  perm.resize(n);
  std::vector<uint32_t> cur(off.begin(), off.end() - 1);
  for (size_t i = 0; i < n; i++) perm[cur[color[i]]++] = (uint32_t)i;
  improper = false;
  std::vector<uint32_t> stamp(n_verts, 0xFFFFFFFFu);
  for (uint32_t c = 0; c < nc && !improper; c++)
    for (uint32_t s = off[c]; s < off[c + 1] && !improper; s++)
      for (int a = 0; a < arity; a++) {
        const uint32_t v = idx[(size_t)perm[s] * arity + a];
        if (stamp[v] == c) improper = true;
        stamp[v] = c;
      }
}

// vertex -> (face << 2 | corner) CSR, faces ascending per vertex
// (BuildVFAdjacency, physics/ClothSolverHelper.cu:377-401 uses std::set<uint>)
void xrt_build_vf(const std::vector<uint32_t>& tri, uint32_t V, std::vector<uint32_t>& off,
                     std::vector<uint32_t>& idx) {
  const size_t F = tri.size() / 3;
  off.assign(V + 1, 0);
  for (size_t i = 0; i < 3 * F; i++) off[tri[i] + 1]++;
  for (uint32_t v = 0; v < V; v++) off[v + 1] += off[v];
  idx.resize(3 * F);
  std::vector<uint32_t> cur(off.begin(), off.end() - 1);
  for (size_t f = 0; f < F; f++)
    for (uint32_t j = 0; j < 3; j++) idx[cur[tri[3 * f + j]]++] = (uint32_t)(f << 2) | j;
}

// RTriBuild_Kernel (physics/representative_triangle/RTriangle.cu:19-74) on the
// host: bit i   = this face is the lowest-index face on its vertex i,
//       bit 3+i = this face owns edge (i, i+1): no other face shares it, or
//                 the lowest-index other face sharing it has a larger index.
void xrt_build_rtri(const std::vector<uint32_t>& tri, const std::vector<uint32_t>& vf_off,
                       const std::vector<uint32_t>& vf_idx, std::vector<uint32_t>& rtri) {
  const size_t F = tri.size() / 3;
  rtri.assign(F, 0);
  for (size_t f = 0; f < F; f++) {
    uint32_t info = 0;
    for (uint32_t i = 0; i < 3; i++) {
      const uint32_t vi = tri[3 * f + i], vj = tri[3 * f + (i + 1) % 3];
      if (vf_off[vi] < vf_off[vi + 1] && (vf_idx[vf_off[vi]] >> 2) == f) info |= 1u << i;
      uint32_t of = 0xFFFFFFFFu;
      for (uint32_t a = vf_off[vi]; a < vf_off[vi + 1] && of == 0xFFFFFFFFu; a++) {
        const uint32_t g = vf_idx[a] >> 2;
        if (g == f) continue;
        for (uint32_t b = vf_off[vj]; b < vf_off[vj + 1]; b++)
          if ((vf_idx[b] >> 2) == g) { of = g; break; }
      }
      if (f < of) info |= 1u << (i + 3u);
    }
    rtri[f] = info;
  }
}


// ---------------------------------------------------------------------------
// constraints from an index buffer
// ---------------------------------------------------------------------------
namespace {
struct HalfEdge {
  uint64_t key;  // lo * n_verts + hi
  uint32_t a, b, opp, face;
};
template <class T>
T* dup(const std::vector<T>& v) {
  T* p = (T*)malloc((v.size() ? v.size() : 1) * sizeof(T));
  if (p && !v.empty()) memcpy(p, v.data(), v.size() * sizeof(T));
  return p;
}
}  // namespace

extern "C" int xrt_build_constraints(const float* positions, uint32_t n_verts, const uint32_t* indices, uint32_t n_faces,
                                     xrt_constraints* out) {
  if (!positions || !indices || !out || !n_verts || !n_faces) return XRT_ERR_INVALID;
  memset(out, 0, sizeof(*out));
  for (size_t i = 0; i < 3 * (size_t)n_faces; i++)
    if (indices[i] >= n_verts) return XRT_ERR_INVALID;
  // the three half-edges of every face, sorted by their undirected edge, then by face
  std::vector<HalfEdge> he(3 * (size_t)n_faces);
  for (uint32_t f = 0; f < n_faces; f++)
    for (int j = 0; j < 3; j++) {
      const uint32_t a = indices[3 * f + j], b = indices[3 * f + (j + 1) % 3], o = indices[3 * f + (j + 2) % 3];
      he[3 * (size_t)f + j] = HalfEdge{(uint64_t)std::min(a, b) * n_verts + std::max(a, b), a, b, o, f};
    }
  std::vector<uint32_t> order(he.size());
  for (size_t i = 0; i < order.size(); i++) order[i] = (uint32_t)i;
  std::sort(order.begin(), order.end(), [&](uint32_t x, uint32_t y) {
    return he[x].key != he[y].key ? he[x].key < he[y].key : he[x].face < he[y].face;
  });
  auto P = [&](uint32_t v, int k) { return positions[3 * (size_t)v + k]; };
  std::vector<uint32_t> edges, face_edges(3 * (size_t)n_faces), bend, ef_off(1, 0), ef_faces;
  std::vector<float> rest, theta;
  for (size_t r = 0; r < order.size();) {
    size_t e = r;
    while (e < order.size() && he[order[e]].key == he[order[r]].key) e++;
    const uint32_t eid = (uint32_t)(edges.size() / 2);
    const uint32_t lo = (uint32_t)(he[order[r]].key / n_verts), hi = (uint32_t)(he[order[r]].key % n_verts);
    edges.push_back(lo);
    edges.push_back(hi);
    {  // PrepareBasicStretchData: rest length = |p[lo] - p[hi]|
      const float dx = P(lo, 0) - P(hi, 0), dy = P(lo, 1) - P(hi, 1), dz = P(lo, 2) - P(hi, 2);
      rest.push_back(sqrtf(dx * dx + dy * dy + dz * dz));
    }
    for (size_t q = r; q < e; q++) {
      const HalfEdge& h = he[order[q]];
      const uint32_t slot = order[q] % 3u;  // half-edge j of face f joins its vertices j and j+1
      face_edges[3 * (size_t)h.face + slot] = eid;
      ef_faces.push_back(h.face);  // ascending: the sort's second key (BuildEFAdjacency's std::set order)
    }
    ef_off.push_back((uint32_t)ef_faces.size());
    // GenerateBendingElements: every pair of faces on a non-boundary edge; [0], [1] = the edge in the winding of
    // the first (lower-index) face, [2] = its opposite vertex, [3] = the other face's opposite vertex
    for (size_t x = r; x < e; x++)
      for (size_t y = x + 1; y < e; y++) {
        const HalfEdge& h1 = he[order[x]];
        const HalfEdge& h2 = he[order[y]];
        const uint32_t i1 = h1.a, i2 = h1.b, i3 = h1.opp, i4 = h2.opp;
        bend.push_back(i1); bend.push_back(i2); bend.push_back(i3); bend.push_back(i4);
        auto sub = [&](uint32_t u, uint32_t v, float o[3]) { for (int k = 0; k < 3; k++) o[k] = P(u, k) - P(v, k); };
        auto cross = [](const float a[3], const float b[3], float o[3]) {
          o[0] = a[1] * b[2] - b[1] * a[2]; o[1] = a[2] * b[0] - b[2] * a[0]; o[2] = a[0] * b[1] - b[0] * a[1];
        };
        float u[3], v[3], n1[3], n2[3];
        sub(i1, i3, u); sub(i2, i3, v); cross(u, v, n1);
        sub(i2, i4, u); sub(i1, i4, v); cross(u, v, n2);
        const float l1 = sqrtf(n1[0] * n1[0] + n1[1] * n1[1] + n1[2] * n1[2]), l2 = sqrtf(n2[0] * n2[0] + n2[1] * n2[1] + n2[2] * n2[2]);
        float d = (n1[0] / l1) * (n2[0] / l2) + (n1[1] / l1) * (n2[1] / l2) + (n1[2] / l1) * (n2[2] / l2);
        d = d < -1.f ? -1.f : (d > 1.f ? 1.f : d);
        theta.push_back(acosf(d));
      }
    r = e;
  }
  out->n_edges = (uint32_t)(edges.size() / 2);
  out->n_bend = (uint32_t)theta.size();
  out->edges = dup(edges);
  out->face_edges = dup(face_edges);
  out->stretch_rest = dup(rest);
  out->bend_idx = dup(bend);
  out->bend_theta = dup(theta);
  out->ef_offsets = dup(ef_off);
  out->ef_faces = dup(ef_faces);
  if (!out->edges || !out->face_edges || !out->stretch_rest || !out->bend_idx || !out->bend_theta || !out->ef_offsets ||
      !out->ef_faces) {
    xrt_free_constraints(out);
    return XRT_ERR_INVALID;
  }
  return XRT_OK;
}

extern "C" void xrt_free_constraints(xrt_constraints* c) {
  if (!c) return;
  free(c->edges); free(c->face_edges); free(c->stretch_rest); free(c->bend_idx); free(c->bend_theta);
  free(c->ef_offsets); free(c->ef_faces);
  memset(c, 0, sizeof(*c));
}
