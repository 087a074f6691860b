/*
 * xrt_oracle_quality.c — CPU restatement of XRTailor's Quality-mode frame
 * (ClothSolverBase::SimulateQuality, pipeline/base/ClothSolverBase.cpp:263-362):
 * cubic vertex-face / edge-edge CCD (impact_zone/ImpactHelper.cu), independent
 * impacts (impact_zone/Entry.cu:25-70), impact-zone isolation and merging
 * (impact_zone/ImpactZoneHelper.cu, ImpactZone.cu), the augmented-Lagrangian zone
 * optimiser (impact_zone/ParallelOptimizer.cu, ParallelImpactZoneOptimizer.cu),
 * imminent repulsion (repulsion/ImminentRepulsion*.cu), PBD repulsion
 * (repulsion/PBDRepulsion.cu) and BVH::Update (broad_phase/lbvh/BVH.cu:820-852).
 *
 * TEST INFRASTRUCTURE ONLY (see xrt_oracle.c, which #includes this file at its
 * end so the two share the scene struct and the static helpers).  Plain C, one
 * thread, IEEE arithmetic, no FMA contraction.  Every function cites the
 * reference file:line it follows, paths relative to /root/reference/xrtailor/src/.
 *
 * Node numbering: cloth node i is global node i, obstacle node j is V + j; the
 * reference orders nodes by pool address, which is the same order (cloth is
 * registered first).  Node::index is LOCAL to its mesh (PhysicsMeshHelper.cu:17)
 * and the reference merges zones on that local index, so cloth node 5 and
 * obstacle node 5 count as shared: kept (zone_intersects).  Faces and edges are
 * numbered the same way (cloth first).
 */

/* ------------------------------------------------------------------------- */
/* per-scene Quality state                                                    */
/* ------------------------------------------------------------------------- */

typedef struct {
  uint32_t n[4]; /* global node ids */
  float t, w[4];
  v3 nrm;
  int is_vf;
} OImpact;

typedef struct {
  uint32_t n[4];
  float bary[4];
} OQuad;

typedef struct {
  OImpact* imp;
  int* node_idx; /* 4 per impact, Node::index (local) */
  size_t n_imp, cap;
  int active, sorted, enabled;
} OZone;

struct OQuality {
  /* QualityModeSettings, config/Base.hpp:121-160 */
  float material[6]; /* xx, yy, xy stiffness, xy / yx Poisson, bending stiffness */
  float imminent_thickness, pbd_thickness, relaxation_rate, obstacle_mass, iz_thickness;
  int imminent_on, pbd_on;
  /* cloth topology */
  uint32_t *fe, *ef_off, *ef_idx, *crtri;
  uint32_t *vf_off, *vf_idx, *ovf_off, *ovf_idx; /* vertex -> faces, ascending (MemoryPool::*_nb_vf) */
  float* fn0; /* Face::n, set once at registration (PhysicsMeshHelper.cu:118-123) */
  /* obstacle topology */
  uint32_t Eo;
  uint32_t *oedge, *ofe, *oef_off, *oef_idx;
  float *ofn0, *ov;
  float *x1, *oxq1; /* Node::x1 snapshots of CollisionStep */
  /* FEM */
  OStrain* strain;
  OIsoBend* iso;
  float* iso_lambda;
  uint32_t n_iso, n_strain_colors, n_iso_colors;
  uint32_t *strain_color, *iso_color, *iso_idx;
  /* BVHs */
  OBvh* cloth_bvh;
  /* imminent repulsion */
  OQuad* quads;
  size_t n_quads;
  float* im_delta;
  int* im_count;
  /* pbd repulsion */
  uint32_t* rep_pairs;
  size_t n_rep_pairs;
  float* rep_delta;
  int* rep_count;
  /* statistics of the last CollisionStep */
  int cs_rounds, cs_success, cs_total_impacts, cs_deform;
  /* scratch: removed / min_index of Node (memory/Node.cuh:9,13) */
  int *removed, *min_index;
};

/* ---- node / face / edge accessors over the two meshes ---- */
static int q_cloth(const OScene* s, uint32_t g) { return g < s->V; }
static uint32_t q_local(const OScene* s, uint32_t g) { return g < s->V ? g : g - s->V; }
static v3 q_x0(const OScene* s, uint32_t g) { return g < s->V ? ld3(s->x0, g) : ld3(s->ox0, g - s->V); }
static v3 q_x(const OScene* s, uint32_t g) { return g < s->V ? ld3(s->x, g) : ld3(s->ox, g - s->V); }
static v3 q_v(const OScene* s, uint32_t g) { return g < s->V ? ld3(s->v, g) : ld3(s->q->ov, g - s->V); }
static v3 q_x1(const OScene* s, uint32_t g) { return g < s->V ? ld3(s->q->x1, g) : ld3(s->q->oxq1, g - s->V); }
static void q_set_x(OScene* s, uint32_t g, v3 p) { if (g < s->V) st3(s->x, g, p); else st3(s->ox, g - s->V, p); }
/* inv_mass: cloth as given; obstacle 1/1e3 (PhysicsMeshHelper.cu:139) */
static float q_invm(const OScene* s, uint32_t g) { return g < s->V ? s->invm[g] : 1.0f / 1e3f; }
/* is_free: cloth nodes are free unless pinned (inv_mass 0 -> SetFree(false)); obstacle nodes never */
static int q_free(const OScene* s, uint32_t g) { return g < s->V ? (s->invm[g] != 0.0f) : 0; }
static uint32_t q_face_node(const OScene* s, uint32_t gf, int k) {
  return gf < s->F ? s->tri[3 * gf + k] : s->V + s->otri[3 * (gf - s->F) + k];
}
static uint32_t q_face_edge(const OScene* s, uint32_t gf, int k) {
  return gf < s->F ? s->q->fe[3 * gf + k] : s->E + s->q->ofe[3 * (gf - s->F) + k];
}
static uint32_t q_edge_node(const OScene* s, uint32_t ge, int k) {
  return ge < s->E ? s->edge[2 * ge + k] : s->V + s->q->oedge[2 * (ge - s->E) + k];
}
static uint32_t q_face_rtri(const OScene* s, uint32_t gf) { return gf < s->F ? s->q->crtri[gf] : s->rtri[gf - s->F]; }
static v3 q_face_n0(const OScene* s, uint32_t gf) { return gf < s->F ? ld3(s->q->fn0, gf) : ld3(s->q->ofn0, gf - s->F); }
static int q_face_free(const OScene* s, uint32_t gf) {
  return q_free(s, q_face_node(s, gf, 0)) && q_free(s, q_face_node(s, gf, 1)) && q_free(s, q_face_node(s, gf, 2));
}
static int q_edge_free(const OScene* s, uint32_t ge) { return q_free(s, q_edge_node(s, ge, 0)) && q_free(s, q_edge_node(s, ge, 1)); }
/* Edge::adjacents (memory/Edge.cuh:30, RegisterEFAdjacency_Kernel PhysicsMesh.cu:659-671): global face ids, <= 4 */
static int q_edge_adj(const OScene* s, uint32_t ge, uint32_t out[4]) {
  const int cl = ge < s->E;
  const uint32_t e = cl ? ge : ge - s->E;
  const uint32_t* off = cl ? s->q->ef_off : s->q->oef_off;
  const uint32_t* idx = cl ? s->q->ef_idx : s->q->oef_idx;
  int n = 0;
  for (uint32_t i = off[e]; i < off[e + 1] && n < 4; i++) out[n++] = (cl ? 0 : s->F) + idx[i];
  return n;
}
static int rtri_vertex(uint32_t info, int n) { return (info >> n) & 1u; }        /* RTriangle.cu:12-14 */
static int rtri_edge(uint32_t info, int n) { return (info >> (n + 3)) & 1u; }    /* RTriangle.cu:15-17 */

static float mixed(v3 a, v3 b, v3 c) { return dot(a, cross(b, c)); } /* MathFunctions.cuh:62-64 */
static float min3f(float a, float b, float c) { return fmin2(a, fmin2(b, c)); }               /* :28-30 */
static float min4f(float a, float b, float c, float d) { return fmin2(fmin2(a, b), fmin2(c, d)); } /* :33-35 */
static float absf(float x) { return x < 0.0f ? -x : x; }                                       /* :58-60 */

/* ------------------------------------------------------------------------- */
/* math/BasicPrimitiveTests.cu, math/MathFunctions.cu                         */
/* ------------------------------------------------------------------------- */

/* SignedVertexFaceDistance, BasicPrimitiveTests.cu:369-388.  On a degenerate face n is the raw cross
 * product and w is left untouched. */
float oracle_signed_vf_distance(v3 x, v3 y0, v3 y1, v3 y2, v3* n, float* w) {
  *n = cross(normalize(sub(y1, y0)), normalize(sub(y2, y0)));
  if (dot(*n, *n) < 1e-6f) return INFINITY;
  *n = normalize(*n);
  const float h = dot(sub(x, y0), *n);
  const float b0 = mixed(sub(y1, x), sub(y2, x), *n);
  const float b1 = mixed(sub(y2, x), sub(y0, x), *n);
  const float b2 = mixed(sub(y0, x), sub(y1, x), *n);
  w[0] = 1.0f;
  w[1] = -b0 / (b0 + b1 + b2);
  w[2] = -b1 / (b0 + b1 + b2);
  w[3] = -b2 / (b0 + b1 + b2);
  return h;
}

/* SignedEdgeEdgeDistance, BasicPrimitiveTests.cu:422-438.  `glm::normalize(n);` discards its result,
 * so n stays cross(normalize(e0), normalize(e1)) (length = sin of the angle): kept. */
float oracle_signed_ee_distance(v3 x0, v3 x1, v3 y0, v3 y1, v3* n, float* w) {
  *n = cross(normalize(sub(x1, x0)), normalize(sub(y1, y0)));
  if (dot(*n, *n) < 1e-6f) return INFINITY;
  const float h = dot(sub(x0, y0), *n);
  const float a0 = mixed(sub(y1, x1), sub(y0, x1), *n);
  const float a1 = mixed(sub(y0, x0), sub(y1, x0), *n);
  const float b0 = mixed(sub(x0, y1), sub(x1, y1), *n);
  const float b1 = mixed(sub(x1, y0), sub(x0, y0), *n);
  w[0] = a0 / (a0 + a1);
  w[1] = a1 / (a0 + a1);
  w[2] = -b0 / (b0 + b1);
  w[3] = -b1 / (b0 + b1);
  return h;
}

/* NewtonsMethod, MathFunctions.cu:85-103 */
static float newtons_method(float a, float b, float c, float d, float x0, int dir) {
  if (dir != 0) {
    const float y0 = d + x0 * (c + x0 * (b + x0 * a));
    const float ddy0 = 2.0f * b + x0 * (6.0f * a);
    x0 += (float)dir * sqrtf(absf(2.0f * y0 / ddy0));
  }
  for (int iter = 0; iter < 100; iter++) {
    const float y = d + x0 * (c + x0 * (b + x0 * a));
    const float dy = c + x0 * (2.0f * b + 3.0f * x0 * a);
    if (dy == 0) return x0;
    const float x1 = x0 - y / dy;
    if (absf(x0 - x1) < 1e-6f) return x0;
    x0 = x1;
  }
  return x0;
}
/* SolveQuadratic, MathFunctions.cu:105-120 */
static int solve_quadratic(float a, float b, float c, float x[2]) {
  const float d = b * b - 4.0f * a * c;
  if (d < 0) {
    x[0] = -b / (2.0f * a);
    return 0;
  }
  const float q = -(b + (float)(b < 0.0f ? -1 : 1) * sqrtf(d)) * 0.5f;
  int i = 0;
  if (absf(a) > 1e-12f * absf(q)) x[i++] = q / a;
  if (absf(q) > 1e-12f * absf(c)) x[i++] = c / q;
  if (i == 2 && x[0] > x[1]) { const float t = x[0]; x[0] = x[1]; x[1] = t; }
  return i;
}
/* SolveCubic, MathFunctions.cu:122-146 */
int oracle_solve_cubic(float a, float b, float c, float d, float x[3]) {
  float xc[2];
  const int n = solve_quadratic(3.0f * a, 2.0f * b, c, xc);
  if (n == 0) {
    x[0] = newtons_method(a, b, c, d, xc[0], 0);
    return 1;
  } else if (n == 1) {
    return solve_quadratic(b, c, d, x);
  } else {
    const float yc[2] = {d + xc[0] * (c + xc[0] * (b + xc[0] * a)), d + xc[1] * (c + xc[1] * (b + xc[1] * a))};
    int i = 0;
    if (yc[0] * a >= 0) x[i++] = newtons_method(a, b, c, d, xc[0], -1);
    if (yc[0] * yc[1] <= 0) {
      const int closer = absf(yc[0]) < absf(yc[1]) ? 0 : 1;
      x[i++] = newtons_method(a, b, c, d, xc[closer], closer == 0 ? 1 : -1);
    }
    if (yc[1] * a <= 0) x[i++] = newtons_method(a, b, c, d, xc[1], 1);
    return i;
  }
}

/* ------------------------------------------------------------------------- */
/* impact_zone/ImpactHelper.cu                                                */
/* ------------------------------------------------------------------------- */

/* CheckImpact, ImpactHelper.cu:34-87.  p0[k] / p1[k]: Node::x0 / Node::x of the four nodes.
 * imp->w / imp->nrm keep whatever an earlier root wrote, like the reference's in-place Impact. */
int oracle_check_impact(int is_vf, const v3 p0[4], const v3 p1[4], OImpact* imp) {
  const v3 x0 = p0[0];
  const v3 v0 = sub(p1[0], x0);
  const v3 x1 = sub(p0[1], x0), x2 = sub(p0[2], x0), x3 = sub(p0[3], x0);
  const v3 v1 = sub(sub(p1[1], p0[1]), v0), v2 = sub(sub(p1[2], p0[2]), v0), v3_ = sub(sub(p1[3], p0[3]), v0);
  const float a0 = mixed(x1, x2, x3);
  const float a1 = mixed(v1, x2, x3) + mixed(x1, v2, x3) + mixed(x1, x2, v3_);
  const float a2 = mixed(x1, v2, v3_) + mixed(v1, x2, v3_) + mixed(v1, v2, x3);
  const float a3 = mixed(v1, v2, v3_);
  float t[3];
  const int ns = oracle_solve_cubic(a3, a2, a1, a0, t);
  for (int i = 0; i < ns; i++) {
    if (t[i] < 0 || t[i] > 1) continue;
    imp->t = t[i];
    v3 q[4];
    for (int k = 0; k < 4; k++) q[k] = add(p0[k], mul(sub(p1[k], p0[k]), t[i])); /* Node::Position, Node.cu:17-19 */
    float d;
    int inside;
    if (is_vf) {
      d = oracle_signed_vf_distance(q[0], q[1], q[2], q[3], &imp->nrm, imp->w);
      inside = min3f(-imp->w[1], -imp->w[2], -imp->w[3]) >= -1e-6f;
    } else {
      d = oracle_signed_ee_distance(q[0], q[1], q[2], q[3], &imp->nrm, imp->w);
      inside = min4f(imp->w[0], imp->w[1], -imp->w[2], -imp->w[3]) >= -1e-6f;
    }
    const v3 rel = add(add(mul(v1, imp->w[1]), mul(v2, imp->w[2])), mul(v3_, imp->w[3]));
    if (dot(imp->nrm, rel) > 0) imp->nrm = neg(imp->nrm);
    if (absf(d) < 1e-6f && inside) return 1;
  }
  return 0;
}

/* Node/Edge/Face::ComputeBounds(true) + Bounds::Overlap(b, thickness), Node.cu:9-15, Bounds.cu:95-128 */
static void q_bounds(const OScene* s, const uint32_t* nodes, int n, float bb[6]) {
  for (int k = 0; k < 3; k++) { bb[k] = INFINITY; bb[3 + k] = -INFINITY; } /* Bounds(), Bounds.cu:5-8 */
  for (int i = 0; i < n; i++) {
    const v3 a = q_x(s, nodes[i]), b = q_x0(s, nodes[i]);
    const float c[6] = {a.x, a.y, a.z, b.x, b.y, b.z};
    for (int k = 0; k < 6; k++) {
      bb[k % 3] = fmin2(bb[k % 3], c[k]);
      bb[3 + k % 3] = fmax2(bb[3 + k % 3], c[k]);
    }
  }
}
static int q_overlap(const float a[6], const float b[6], float th) { /* a.Overlap(b.Dilate(th)) */
  for (int k = 0; k < 3; k++) {
    if (a[k] > b[3 + k] + th) return 0;
    if (a[3 + k] < b[k] - th) return 0;
  }
  return 1;
}

static void impact_reset(OImpact* im, int is_vf, uint32_t a, uint32_t b, uint32_t c, uint32_t d) {
  im->is_vf = is_vf;
  im->n[0] = a; im->n[1] = b; im->n[2] = c; im->n[3] = d;
  im->w[0] = im->w[1] = im->w[2] = im->w[3] = 0.0f;
  im->nrm = V(0, 0, 0);
  im->t = -1.0f;
}
static int q_check(const OScene* s, OImpact* im) {
  v3 p0[4], p1[4];
  for (int k = 0; k < 4; k++) { p0[k] = q_x0(s, im->n[k]); p1[k] = q_x(s, im->n[k]); }
  if (!oracle_check_impact(im->is_vf, p0, p1, im)) { im->t = -1.0f; return 0; }
  return 1;
}
/* CheckVertexFaceImpact, ImpactHelper.cu:89-100 */
static int q_check_vf(const OScene* s, uint32_t node, uint32_t gf, float th, OImpact* im) {
  const uint32_t f[3] = {q_face_node(s, gf, 0), q_face_node(s, gf, 1), q_face_node(s, gf, 2)};
  impact_reset(im, 1, node, f[0], f[1], f[2]);
  if (node == f[0] || node == f[1] || node == f[2]) return 0;
  float bn[6], bf[6];
  q_bounds(s, &node, 1, bn);
  q_bounds(s, f, 3, bf);
  if (!q_overlap(bn, bf, th)) return 0;
  return q_check(s, im);
}
/* CheckEdgeEdgeImpact, ImpactHelper.cu:102-114 */
static int q_check_ee(const OScene* s, uint32_t e0, uint32_t e1, float th, OImpact* im) {
  const uint32_t a[2] = {q_edge_node(s, e0, 0), q_edge_node(s, e0, 1)}, b[2] = {q_edge_node(s, e1, 0), q_edge_node(s, e1, 1)};
  impact_reset(im, 0, a[0], a[1], b[0], b[1]);
  if (a[0] == b[0] || a[0] == b[1] || a[1] == b[0] || a[1] == b[1]) return 0;
  float ba[6], bb[6];
  q_bounds(s, a, 2, ba);
  q_bounds(s, b, 2, bb);
  if (!q_overlap(ba, bb, th)) return 0;
  return q_check(s, im);
}

/* CheckImpacts_Kernel + remove_if, ImpactHelper.cu:116-171,238-240: 6 VF + 9 EE tests per pair, hits kept in
 * pair order then slot order.  pairs: (global face, global face) rows.  *tests counts the narrow-phase calls
 * that passed the node / box filters.  Returns a malloc'd array. */
static OImpact* q_check_impacts(const OScene* s, const uint32_t* pairs, size_t np, float th, size_t* n_out) {
  size_t cap = 64, n = 0;
  OImpact* out = (OImpact*)malloc(cap * sizeof(OImpact));
  for (size_t p = 0; p < np; p++) {
    const uint32_t f0 = pairs[2 * p], f1 = pairs[2 * p + 1];
    OImpact im;
    for (int slot = 0; slot < 15; slot++) {
      int hit;
      if (slot < 3) hit = q_check_vf(s, q_face_node(s, f0, slot), f1, th, &im);
      else if (slot < 6) hit = q_check_vf(s, q_face_node(s, f1, slot - 3), f0, th, &im);
      else hit = q_check_ee(s, q_face_edge(s, f0, (slot - 6) / 3), q_face_edge(s, f1, (slot - 6) % 3), th, &im);
      if (!hit) continue;
      if (n == cap) { cap *= 2; out = (OImpact*)realloc(out, cap * sizeof(OImpact)); }
      out[n++] = im;
    }
  }
  *n_out = n;
  return out;
}

/* BVH::Update, BVH.cu:820-852: leaf boxes grow by Face::ComputeBounds(ccd), pred = x; internal boxes grow
 * (AssignAABBAtomic merges into the old box; leaves only grow, so this equals the merge of the children). */
void oracle_bvh_update(OBvh* b, const float* x0, const float* x, const uint32_t* tri, int ccd) {
  const uint32_t NI = b->F - 1;
  for (uint32_t i = 0; i < b->F; i++) {
    const uint32_t o = b->nodes[4 * (NI + i) + 3];
    float* bb = b->aabb + 6 * (size_t)(NI + i);
    for (int k = 0; k < 3; k++) {
      const uint32_t vi = tri[3 * o + k];
      const v3 p = ld3(x, vi);
      st3(b->pred, 3 * o + k, p);
      bb[0] = fmin2(bb[0], p.x); bb[1] = fmin2(bb[1], p.y); bb[2] = fmin2(bb[2], p.z);
      bb[3] = fmax2(bb[3], p.x); bb[4] = fmax2(bb[4], p.y); bb[5] = fmax2(bb[5], p.z);
      if (ccd) {
        const v3 q = ld3(x0, vi);
        bb[0] = fmin2(bb[0], q.x); bb[1] = fmin2(bb[1], q.y); bb[2] = fmin2(bb[2], q.z);
        bb[3] = fmax2(bb[3], q.x); bb[4] = fmax2(bb[4], q.y); bb[5] = fmax2(bb[5], q.z);
      }
    }
  }
  if (b->F > 1) refit_internal(b, 0);
}

/* pairs of two Traverse calls, obstacle pairs first then self pairs (FindImpacts ImpactHelper.cu:226-230,
 * PBDRepulsionSolver::UpdatePairs PBDRepulsion.cu:22-33); or self first (ImminentRepulsion::UpdateProximity,
 * ImminentRepulsion.cu:21-28: `pairs.insert(end, self_pairs)` also appends self after obstacle).
 * Rows are global face ids. */
static uint32_t* q_pairs(const OScene* s, const OBvh* cloth, const OBvh* obs, float th, size_t* n_out) {
  const size_t no = obs ? oracle_traverse_pairs(cloth, obs, th, NULL, 0) : 0;
  const size_t ns = oracle_traverse_pairs(cloth, cloth, th, NULL, 0);
  uint32_t* out = (uint32_t*)malloc(8 * (no + ns) + 8);
  if (no) {
    oracle_traverse_pairs(cloth, obs, th, out, no);
    for (size_t i = 0; i < no; i++) out[2 * i + 1] += s->F;
  }
  if (ns) oracle_traverse_pairs(cloth, cloth, th, out + 2 * no, ns);
  *n_out = no + ns;
  return out;
}

/* FindImpacts, ImpactHelper.cu:216-243 */
static OImpact* q_find_impacts(const OScene* s, const OBvh* cloth, const OBvh* obs, float th, size_t* n_out, size_t* n_pairs) {
  size_t np;
  uint32_t* pairs = q_pairs(s, cloth, obs, th, &np);
  OImpact* r = q_check_impacts(s, pairs, np, th, n_out);
  if (n_pairs) *n_pairs = np;
  free(pairs);
  return r;
}

/* ------------------------------------------------------------------------- */
/* independentImpacts, impact_zone/Entry.cu:25-70 + ImpactHelper.cu:245-341   */
/* ------------------------------------------------------------------------- */
static int cmp_impact_t(const void* a, const void* b) { /* Impact::operator<, Impact.cu:7-9; index as tie-break */
  const OImpact* x = *(const OImpact* const*)a;
  const OImpact* y = *(const OImpact* const*)b;
  if (x->t < y->t) return -1;
  if (x->t > y->t) return 1;
  return x < y ? -1 : (x > y ? 1 : 0);
}
static OImpact* q_independent(OScene* s, const OImpact* in, size_t n, int deform, size_t* n_out) {
  /* thrust::sort by t (not stable in the reference; a stable order is used here and in the CUDA path) */
  const OImpact** ord = (const OImpact**)malloc(sizeof(void*) * (n ? n : 1));
  for (size_t i = 0; i < n; i++) ord[i] = in + i;
  qsort(ord, n, sizeof(void*), cmp_impact_t);
  OImpact* sorted = (OImpact*)malloc(sizeof(OImpact) * (n ? n : 1));
  for (size_t i = 0; i < n; i++) sorted[i] = *ord[i];
  free(ord);
  int* removed = s->q->removed;
  int* min_index = s->q->min_index;
  char* taken = (char*)calloc(n ? n : 1, 1);
#define ELIG(g) (deform == 1 || q_free(s, (g)))
  for (size_t i = 0; i < n; i++) /* InitializeImpactNode_Kernel, ImpactHelper.cu:245-258 */
    for (int j = 0; j < 4; j++)
      if (ELIG(sorted[i].n[j])) { removed[sorted[i].n[j]] = 0; min_index[sorted[i].n[j]] = (int)n; }
  size_t num, new_num = n;
  do {
    num = new_num;
    /* CollectRelativeImpacts + sort_by_key + reduce_by_key(min) + SetImpactMinIndices, ImpactHelper.cu:266-306:
     * only nodes listed by an available impact get a new min_index; a first pass must not see the minima of
     * this round, so available flags are computed before any write */
    char* avail = (char*)malloc(n ? n : 1);
    for (size_t i = 0; i < n; i++) {
      avail[i] = 1;
      for (int j = 0; j < 4; j++)
        if (ELIG(sorted[i].n[j]) && removed[sorted[i].n[j]]) { avail[i] = 0; break; }
    }
    for (size_t i = 0; i < n; i++)
      if (avail[i])
        for (int j = 0; j < 4; j++)
          if (ELIG(sorted[i].n[j])) min_index[sorted[i].n[j]] = -1; /* mark: gets a fresh minimum */
    for (size_t i = 0; i < n; i++)
      if (avail[i])
        for (int j = 0; j < 4; j++) {
          const uint32_t g = sorted[i].n[j];
          if (ELIG(g) && (min_index[g] < 0 || (int)i < min_index[g])) min_index[g] = (int)i;
        }
    free(avail);
    /* CheckIndependentImpacts_Kernel, ImpactHelper.cu:308-333 (flags read before this round's removals:
     * an impact that shares a node with a winner fails on min_index anyway) */
    char* win = (char*)calloc(n ? n : 1, 1);
    for (size_t i = 0; i < n; i++) {
      int flag = 1;
      for (int j = 0; j < 4; j++) {
        const uint32_t g = sorted[i].n[j];
        if (ELIG(g) && (removed[g] || min_index[g] != (int)i)) { flag = 0; break; }
      }
      win[i] = (char)flag;
    }
    for (size_t i = 0; i < n; i++)
      if (win[i]) {
        taken[i] = 1;
        for (int j = 0; j < 4; j++)
          if (ELIG(sorted[i].n[j])) removed[sorted[i].n[j]] = 1;
      }
    free(win);
    new_num = 0;
    for (size_t i = 0; i < n; i++) new_num += !taken[i];
  } while (num > new_num);
#undef ELIG
  size_t m = 0;
  OImpact* out = (OImpact*)malloc(sizeof(OImpact) * (n ? n : 1));
  for (size_t i = 0; i < n; i++)
    if (taken[i]) out[m++] = sorted[i];
  free(taken);
  free(sorted);
  *n_out = m;
  return out;
}

/* ------------------------------------------------------------------------- */
/* zones: impact_zone/ImpactZoneHelper.cu, ImpactZone.cu                      */
/* ------------------------------------------------------------------------- */
#define NODE_OUTSIDE -2
#define NODE_INSIDE -1

/* IsolateImpacts, ImpactZoneHelper.cu:227-334, with FloodFillColoring_Kernel :38-95 and
 * WriteBackNodeColor_Kernel :97-117.  colors[i] receives the zone of impact i (or -1 after 100 zones). */
static int q_isolate(const OScene* s, const OImpact* imp, size_t n, int* colors) {
  const uint32_t NV = s->V + s->Vo;
  int* ncol = (int*)malloc(sizeof(int) * (NV ? NV : 1));
  for (uint32_t i = 0; i < NV; i++) ncol[i] = NODE_OUTSIDE;
  for (size_t i = 0; i < n; i++) {
    colors[i] = -1;
    for (int j = 0; j < 4; j++) ncol[imp[i].n[j]] = NODE_INSIDE;
  }
  /* vertex -> faces CSR for both meshes (MemoryPool::cloth_nb_vf / obstacle_nb_vf, ascending face order) */
  uint32_t* stack = (uint32_t*)malloc(sizeof(uint32_t) * ((size_t)NV + 8)); /* a node is pushed once, when it is coloured */
  int n_colors = 0;
  for (;;) {
    if (n_colors > 99) break; /* "too many attempts, exit" */
    size_t pos = 0;
    while (pos < n && colors[pos] != -1) pos++;
    if (pos == n) break;
    size_t sp = 0;
    for (int j = 0; j < 4; j++) stack[sp++] = imp[pos].n[j];
    do {
      const uint32_t node = stack[--sp];
      const int cl = q_cloth(s, node);
      if (cl && !q_free(s, node)) continue; /* pinned cloth nodes are not expanded */
      const uint32_t l = q_local(s, node);
      const uint32_t* off = cl ? s->q->vf_off : s->q->ovf_off;
      const uint32_t* idx = cl ? s->q->vf_idx : s->q->ovf_idx;
      for (uint32_t i = off[l]; i < off[l + 1]; i++) {
        const uint32_t gf = (cl ? 0 : s->F) + idx[i];
        for (int j = 0; j < 3; j++) {
          const uint32_t g = q_face_node(s, gf, j);
          if (ncol[g] == NODE_INSIDE) {
            ncol[g] = n_colors;
            stack[sp++] = g;
          }
        }
      }
    } while (sp > 0);
    for (size_t i = 0; i < n; i++) {
      if (colors[i] != -1) continue;
      for (int j = 0; j < 4; j++) {
        const int c = ncol[imp[i].n[j]];
        if (c != NODE_OUTSIDE && c != NODE_INSIDE) { colors[i] = c; break; }
      }
    }
    n_colors++;
  }
  free(stack);
  free(ncol);
  return n_colors;
}

static void zone_push(OZone* z, const OImpact* im, const OScene* s) {
  if (z->n_imp == z->cap) {
    z->cap = z->cap ? 2 * z->cap : 8;
    z->imp = (OImpact*)realloc(z->imp, z->cap * sizeof(OImpact));
    z->node_idx = (int*)realloc(z->node_idx, z->cap * 4 * sizeof(int));
  }
  z->imp[z->n_imp] = *im;
  for (int j = 0; j < 4; j++) z->node_idx[4 * z->n_imp + j] = (int)q_local(s, im->n[j]); /* FlattenNode_Kernel */
  z->n_imp++;
}
static int cmp_int(const void* a, const void* b) { return (*(const int*)a > *(const int*)b) - (*(const int*)a < *(const int*)b); }
/* ImpactZone::Intersects, ImpactZone.cu:74-84 (sorts both node lists in place) */
static int zone_intersects(OZone* a, OZone* b) {
  if (!a->sorted) { qsort(a->node_idx, 4 * a->n_imp, sizeof(int), cmp_int); a->sorted = 1; }
  if (!b->sorted) { qsort(b->node_idx, 4 * b->n_imp, sizeof(int), cmp_int); b->sorted = 1; }
  size_t i = 0, j = 0;
  while (i < 4 * a->n_imp && j < 4 * b->n_imp) {
    if (a->node_idx[i] < b->node_idx[j]) i++;
    else if (a->node_idx[i] > b->node_idx[j]) j++;
    else return 1;
  }
  return 0;
}
/* ImpactZone::Merge, ImpactZone.cu:57-72 */
static int zone_merge(OZone* a, OZone* b) {
  if (!b->enabled) return 0;
  if (!zone_intersects(a, b)) return 0;
  for (size_t i = 0; i < b->n_imp; i++) {
    if (a->n_imp == a->cap) {
      a->cap = a->cap ? 2 * a->cap : 8;
      a->imp = (OImpact*)realloc(a->imp, a->cap * sizeof(OImpact));
      a->node_idx = (int*)realloc(a->node_idx, a->cap * 4 * sizeof(int));
    }
    a->imp[a->n_imp] = b->imp[i];
    memcpy(a->node_idx + 4 * a->n_imp, b->node_idx + 4 * i, 4 * sizeof(int));
    a->n_imp++;
  }
  a->active = 1;
  a->sorted = 0;
  a->enabled = 1;
  b->enabled = 0;
  return 1;
}

typedef struct {
  OZone* z;
  size_t n, cap;
} OZoneList;
static void zl_push(OZoneList* l, OZone z) {
  if (l->n == l->cap) { l->cap = l->cap ? 2 * l->cap : 8; l->z = (OZone*)realloc(l->z, l->cap * sizeof(OZone)); }
  l->z[l->n++] = z;
}
static void zl_free(OZoneList* l) {
  for (size_t i = 0; i < l->n; i++) { free(l->z[i].imp); free(l->z[i].node_idx); }
  free(l->z);
  l->z = NULL; l->n = l->cap = 0;
}

/* AddImpacts, ImpactZoneHelper.cu:344-384.  A disabled zone that absorbs a later one is enabled again
 * (ImpactZone::Merge calls Enable() on itself): kept. */
static void q_add_impacts(const OScene* s, OZoneList* zones, const OImpact* nimp, size_t n) {
  int* col = (int*)malloc(sizeof(int) * (n ? n : 1));
  const int nc = n ? q_isolate(s, nimp, n, col) : 0;
  OZoneList nz = {0};
  /* stable_sort_by_key(colors) + offsets from the per-flood-fill counts (:305-331); impacts left at -1 after the
   * 100-zone cap sort to the front and shift every range: the ranges are taken from the counts like the reference */
  size_t* cnt = (size_t*)calloc((size_t)nc + 1, sizeof(size_t));
  size_t* order = (size_t*)malloc(sizeof(size_t) * (n ? n : 1));
  size_t m = 0;
  for (int c = -1; c < nc; c++)
    for (size_t i = 0; i < n; i++)
      if (col[i] == c) { order[m++] = i; if (c >= 0) cnt[c]++; }
  size_t lhs = 0;
  for (int c = 0; c < nc; c++) {
    OZone z = {0};
    z.active = 1; z.enabled = 1;
    for (size_t i = lhs; i < lhs + cnt[c] && i < n; i++) zone_push(&z, &nimp[order[i]], s);
    lhs += cnt[c];
    zl_push(&nz, z);
  }
  free(cnt); free(order); free(col);
  const size_t n_zone = zones->n;
  for (size_t i = 0; i < n_zone; i++) zones->z[i].active = 0;
  for (size_t i = 0; i < n_zone; i++)
    for (size_t j = 0; j < nz.n; j++) zone_merge(&zones->z[i], &nz.z[j]);
  for (size_t j = 0; j < nz.n; j++) {
    if (nz.z[j].enabled) {
      zl_push(zones, nz.z[j]);
      nz.z[j].imp = NULL; nz.z[j].node_idx = NULL;
    }
  }
  zl_free(&nz);
  for (size_t i = 0; i < zones->n; i++)
    for (size_t j = i + 1; j < zones->n; j++) zone_merge(&zones->z[i], &zones->z[j]);
  size_t k = 0;
  for (size_t i = 0; i < zones->n; i++) {
    if (zones->z[i].enabled) zones->z[k++] = zones->z[i];
    else { free(zones->z[i].imp); free(zones->z[i].node_idx); }
  }
  zones->n = k;
}

/* ------------------------------------------------------------------------- */
/* the zone optimiser: ParallelImpactZoneOptimizer.cu, ParallelOptimizer.cu   */
/* ------------------------------------------------------------------------- */
#define IZO_BLOCK 256
#define IZO_EPS_S 1e-12f
#define IZO_EPS_F 1e-10f
#define IZO_ALPHA 0.5f

/* ReduceBaselineLocal_Kernel + ReduceBaselineGlobal_Kernel, ParallelOptimizer.cu:330-420: blocks of 256 summed
 * by a stride-doubling tree, then the (<= 256) block sums by the same tree */
static float tree256(const float* v, size_t n) {
  float sbuf[IZO_BLOCK];
  for (int i = 0; i < IZO_BLOCK; i++) sbuf[i] = (size_t)i < n ? v[i] : 0.0f;
  for (int stride = 1; stride < IZO_BLOCK; stride *= 2)
    for (int t = 0; t + stride < IZO_BLOCK; t += 2 * stride) sbuf[t] += sbuf[t + stride];
  return sbuf[0];
}
static float zone_reduce(const float* v, size_t n) {
  float part[IZO_BLOCK];
  size_t nb = (n + IZO_BLOCK - 1) / IZO_BLOCK;
  if (nb > IZO_BLOCK) nb = IZO_BLOCK; /* the reference reads only the first 256 block sums */
  for (size_t b = 0; b < nb; b++) part[b] = tree256(v + b * IZO_BLOCK, n - b * IZO_BLOCK < IZO_BLOCK ? n - b * IZO_BLOCK : IZO_BLOCK);
  return tree256(part, nb);
}

typedef struct {
  size_t imp_off, n_imp, node_off, n_nodes;
  float s, f, ft, O, Ot, C2, C2t, G2, lambda2;
  int converged, wolfe;
} OZoneAttr;

typedef struct {
  int outer_iterations, line_search_steps, all_converged;
} OSolveStats;

typedef struct { int color; uint32_t node; int slot; } Ent;
static int cmp_ent(const void* a, const void* b) {
  const Ent* x = (const Ent*)a; const Ent* y = (const Ent*)b;
  if (x->color != y->color) return x->color < y->color ? -1 : 1;
  if (x->node != y->node) return x->node < y->node ? -1 : 1;
  return (x->slot > y->slot) - (x->slot < y->slot);
}

/* ParallelImpactZoneOptimizer ctor (:167-282) + ParallelOptimizer::Solve (ParallelOptimizer.cu:513-890).
 * imp: impacts of the active zones back to back, color[i] = zone of impact i (ascending). */
static void q_optimize(OScene* s, const OImpact* imp, const int* color, size_t n, int nz, int deform, float thickness,
                       float obstacle_mass, OSolveStats* st) {
  if (n == 0 || nz == 0) return;
  /* CollectCollisionNodes (:14-34): variables are the free cloth nodes, or every node when deform == 1 */
  Ent* ent = (Ent*)malloc(sizeof(Ent) * 4 * n);
  size_t ne = 0;
  for (size_t i = 0; i < n; i++)
    for (int j = 0; j < 4; j++) {
      const uint32_t g = imp[i].n[j];
      if (deform == 1 || (q_cloth(s, g) && q_free(s, g))) { ent[ne].color = color[i]; ent[ne].node = g; ent[ne].slot = (int)(4 * i + j); ne++; }
    }
  /* sort_by_key((color, node ptr)) :203-207; the slot as last key makes it a stable order */
  qsort(ent, ne, sizeof(Ent), cmp_ent);
  /* SetDiff (node pointer only) + inclusive_scan + SetIndices + unique_by_key, :36-49,209-227 */
  int* indices = (int*)malloc(sizeof(int) * 4 * n);
  for (size_t i = 0; i < 4 * n; i++) indices[i] = -1;
  uint32_t* nodes = (uint32_t*)malloc(sizeof(uint32_t) * (ne ? ne : 1));
  int* ncolor = (int*)malloc(sizeof(int) * (ne ? ne : 1));
  size_t nn = 0;
  int run = 0;
  for (size_t i = 0; i < ne; i++) {
    if (i > 0 && ent[i].node != ent[i - 1].node) run++;
    indices[ent[i].slot] = run;
    if (i == 0 || ent[i].node != ent[i - 1].node) { nodes[nn] = ent[i].node; ncolor[nn] = ent[i].color; nn++; }
  }
  free(ent);
  OZoneAttr* z = (OZoneAttr*)calloc((size_t)nz, sizeof(OZoneAttr));
  { /* node_offsets from thrust::count per colour (:229-235), impact offsets from the zone sizes */
    size_t no = 0, io = 0;
    for (int c = 0; c < nz; c++) {
      size_t cn = 0, ci = 0;
      for (size_t i = 0; i < nn; i++) cn += ncolor[i] == c;
      for (size_t i = 0; i < n; i++) ci += color[i] == c;
      z[c].node_off = no; z[c].n_nodes = cn; z[c].imp_off = io; z[c].n_imp = ci;
      no += cn; io += ci;
    }
  }
  /* inv_mass = mean of the per-node inverse masses, thrust::reduce (:237-243) */
  float inv_mass = 0.0f;
  for (size_t i = 0; i < nn; i++) inv_mass += q_cloth(s, nodes[i]) ? q_invm(s, nodes[i]) : 1.0f / obstacle_mass;
  inv_mass = inv_mass / (float)nn;

  const float mu = 1e3f;
  float* X = (float*)calloc(3 * nn + 3, 4);    /* current_X */
  float* Xn = (float*)calloc(3 * nn + 3, 4);   /* next_X (zero-initialised device_vector) */
  float* G = (float*)calloc(3 * nn + 3, 4);
  float* obj = (float*)calloc(nn + 1, 4);
  float* con = (float*)calloc(n + 1, 4);
  float* lam = (float*)calloc(n + 1, 4);
  float* tmp = (float*)calloc((nn > n ? nn : n) + 1, 4);
  for (size_t i = 0; i < nn; i++) st3(X, (uint32_t)i, q_x(s, nodes[i])); /* Initialize_Kernel */
  for (int c = 0; c < nz; c++) z[c].s = 1e-3f * (float)z[c].n_nodes;     /* InitStepLength_Kernel */

#define MASS(i) (q_cloth(s, nodes[i]) ? (1.0f / q_invm(s, nodes[i])) : obstacle_mass)
#define CONSTRAINT(XX)                                                                    \
  for (size_t i_ = 0; i_ < n; i_++) { /* CollisionConstraint_Kernel :77-93 */              \
    float c_ = -thickness;                                                                 \
    for (int j_ = 0; j_ < 4; j_++) {                                                       \
      const int k_ = indices[4 * i_ + j_];                                                 \
      if (k_ > -1) c_ += imp[i_].w[j_] * dot(imp[i_].nrm, ld3(XX, (uint32_t)k_));          \
      else c_ += imp[i_].w[j_] * dot(imp[i_].nrm, q_x(s, imp[i_].n[j_]));                  \
    }                                                                                      \
    con[i_] = c_;                                                                          \
  }
#define COEFF() for (size_t i_ = 0; i_ < n; i_++) con[i_] = fmin2(con[i_] + lam[i_] / mu, 0.0f); /* ClampViolation, sign +1 */
#define OBJECTIVE(XX)                                                                     \
  for (size_t i_ = 0; i_ < nn; i_++) { /* CollisionObjective_Kernel :57-65 */              \
    const v3 d_ = sub(ld3(XX, (uint32_t)i_), q_x1(s, nodes[i_]));                          \
    obj[i_] = MASS(i_) * dot(d_, d_);                                                      \
  }
#define REDUCE_SQ(SRC, OFF, CNT, DST) do { for (size_t i_ = 0; i_ < (CNT); i_++) tmp[i_] = (SRC)[(OFF) + i_] * (SRC)[(OFF) + i_]; DST = zone_reduce(tmp, (CNT)); } while (0)

  int iter;
  int all_conv = 0;
  int ls_total = 0;
  for (iter = 0; iter < 100; iter++) {
    for (int c = 0; c < nz; c++) z[c].wolfe = z[c].converged; /* ResetWolfeCondition_Kernel */
    CONSTRAINT(X);
    /* TotalGradient (:213-224): objective gradient, coefficients, constraint gradient */
    for (size_t i = 0; i < nn; i++) st3(G, (uint32_t)i, mul(sub(ld3(X, (uint32_t)i), q_x1(s, nodes[i])), inv_mass * MASS(i)));
    COEFF();
    for (size_t i = 0; i < n; i++)
      for (int j = 0; j < 4; j++) {
        const int k = indices[4 * i + j];
        if (k > -1) st3(G, (uint32_t)k, add(ld3(G, (uint32_t)k), mul(imp[i].nrm, mu * con[i] * imp[i].w[j])));
      }
    for (int c = 0; c < nz; c++) {
      if (z[c].converged) continue;
      for (size_t i = 0; i < z[c].n_nodes; i++) { const v3 g = ld3(G, (uint32_t)(z[c].node_off + i)); tmp[i] = dot(g, g); }
      z[c].G2 = zone_reduce(tmp, z[c].n_nodes);
    }
    OBJECTIVE(X);
    for (int c = 0; c < nz; c++) {
      if (z[c].converged) continue;
      REDUCE_SQ(con, z[c].imp_off, z[c].n_imp, z[c].C2);
      REDUCE_SQ(lam, z[c].imp_off, z[c].n_imp, z[c].lambda2);
      z[c].O = zone_reduce(obj + z[c].node_off, z[c].n_nodes);
      z[c].s /= 0.7f; /* PreconditionStepLength_Kernel */
    }
    int pending;
    do {
      for (int c = 0; c < nz; c++) {
        if (z[c].converged || z[c].wolfe) continue;
        z[c].s *= 0.7f; /* AdjustStepLength_Kernel */
        for (size_t i = 0; i < z[c].n_nodes; i++) { /* NextX_Kernel */
          const uint32_t k = (uint32_t)(z[c].node_off + i);
          st3(Xn, k, sub(ld3(X, k), mul(ld3(G, k), z[c].s)));
        }
      }
      OBJECTIVE(Xn);
      CONSTRAINT(Xn);
      COEFF();
      pending = 0;
      for (int c = 0; c < nz; c++) {
        if (z[c].converged) continue;
        REDUCE_SQ(con, z[c].imp_off, z[c].n_imp, z[c].C2t);
        const float Ot = zone_reduce(obj + z[c].node_off, z[c].n_nodes);
        /* BacktrackingLineSearchAdvanced_Kernel, ParallelImpactZoneOptimizer.cu:383-468 */
        const float f = 0.5f * inv_mass * z[c].O + 0.5f * mu * z[c].C2 - 0.5f * z[c].lambda2 / mu;
        const float ft = 0.5f * inv_mass * Ot + 0.5f * mu * z[c].C2t - 0.5f * z[c].lambda2 / mu;
        if (z[c].s < IZO_EPS_S || ft < (f - IZO_ALPHA * z[c].s * z[c].G2) || absf(f - ft) < IZO_EPS_F) z[c].wolfe = 1;
        z[c].Ot = Ot; z[c].f = f; z[c].ft = ft;
        if (!z[c].wolfe) pending = 1;
      }
      ls_total++;
    } while (pending);
    all_conv = 1;
    for (int c = 0; c < nz; c++) {
      if (z[c].s < IZO_EPS_S) z[c].converged = 1; /* UpdateConvergency_Kernel */
      if (!z[c].converged) all_conv = 0;
    }
    if (all_conv) break;
    memcpy(X, Xn, 12 * nn); /* previous_X = current_X; current_X = next_X */
    CONSTRAINT(X);          /* UpdateMultiplier (:226-235) */
    for (size_t i = 0; i < n; i++) lam[i] = fmin2(lam[i] + mu * con[i], 0.0f);
  }
  for (size_t i = 0; i < nn; i++) q_set_x(s, nodes[i], ld3(X, (uint32_t)i)); /* Finalize_Kernel */
  if (st) { st->outer_iterations = iter; st->line_search_steps = ls_total; st->all_converged = all_conv; }
#undef MASS
#undef CONSTRAINT
#undef COEFF
#undef OBJECTIVE
#undef REDUCE_SQ
  free(X); free(Xn); free(G); free(obj); free(con); free(lam); free(tmp);
  free(indices); free(nodes); free(ncolor); free(z);
}

/* FlattenImpactZones (ImpactZoneHelper.cu:386-416) + the optimiser over the active zones */
static void q_solve_zones(OScene* s, OZoneList* zones, int deform, float thickness, float obstacle_mass, OSolveStats* st) {
  size_t n = 0;
  int nz = 0;
  for (size_t i = 0; i < zones->n; i++)
    if (zones->z[i].active) { n += zones->z[i].n_imp; nz++; }
  OImpact* imp = (OImpact*)malloc(sizeof(OImpact) * (n ? n : 1));
  int* col = (int*)malloc(sizeof(int) * (n ? n : 1));
  size_t m = 0;
  int c = 0;
  for (size_t i = 0; i < zones->n; i++) {
    if (!zones->z[i].active) continue;
    for (size_t k = 0; k < zones->z[i].n_imp; k++) { imp[m] = zones->z[i].imp[k]; col[m] = c; m++; }
    c++;
  }
  q_optimize(s, imp, col, n, nz, deform, thickness, obstacle_mass, st);
  free(imp); free(col);
}

/* CollisionStep, impact_zone/Entry.cu:140-256 */
void oracle_collision_step(OScene* s, float dt) {
  OQuality* q = s->q;
  memcpy(q->x1, s->x, 12 * (size_t)s->V); /* SetupNode_Kernel */
  for (uint32_t i = 0; i < s->Vo; i++) {  /* SetupObstacleNode_Kernel */
    st3(q->oxq1, i, ld3(s->ox, i));
    st3(q->ov, i, divs(sub(ld3(s->ox, i), ld3(s->ox0, i)), dt));
  }
  OBvh* cb = oracle_bvh_build(s->x0, s->x, NULL, s->tri, s->F, 0.0f, 1);
  OBvh* ob = s->Fo ? oracle_bvh_build(s->ox0, s->ox, NULL, s->otri, s->Fo, 0.0f, 1) : NULL;
  float obstacle_mass = q->obstacle_mass;
  int success = 0;
  q->cs_rounds = 0; q->cs_total_impacts = 0; q->cs_deform = 0;
  for (int deform = 0; deform < 2 && !success; deform++) {
    OZoneList zones = {0};
    q->cs_deform = deform;
    for (int i = 0; i < 100; i++) {
      size_t n_new;
      OImpact* found = q_find_impacts(s, cb, ob, q->iz_thickness, &n_new, NULL);
      if (n_new == 0) { free(found); success = 1; break; }
      size_t n_ind;
      OImpact* ind = q_independent(s, found, n_new, deform, &n_ind);
      free(found);
      q->cs_total_impacts += (int)n_ind;
      q->cs_rounds++;
      q_add_impacts(s, &zones, ind, n_ind);
      free(ind);
      q_solve_zones(s, &zones, deform, q->iz_thickness, obstacle_mass, NULL);
      oracle_bvh_update(cb, s->x0, s->x, s->tri, 1);
      if (deform == 1) {
        if (ob) oracle_bvh_update(ob, s->ox0, s->ox, s->otri, 1);
        obstacle_mass *= 0.5f;
      }
    }
    zl_free(&zones);
  }
  q->cs_success = success;
  oracle_bvh_free(cb);
  if (ob) oracle_bvh_free(ob);
}

/* ------------------------------------------------------------------------- */
/* imminent repulsion: repulsion/ImminentRepulsion.cu, ImminentRepulsionHelper.cu */
/* ------------------------------------------------------------------------- */

/* InEdge, BasicPrimitiveTests.cu:286-301: uses Face::n as set at registration */
static int q_in_edge(const OScene* s, float w, uint32_t e0, uint32_t e1) {
  const v3 x = add(mul(q_x(s, q_edge_node(s, e0, 0)), 1.0f - w), mul(q_x(s, q_edge_node(s, e0, 1)), w));
  uint32_t adj[4];
  const int na = q_edge_adj(s, e1, adj);
  int in = 1;
  for (int i = 0; i < na; i++) {
    const v3 p0 = q_x(s, q_edge_node(s, e1, 0)), p1 = q_x(s, q_edge_node(s, e1, 1));
    const v3 e = sub(p1, p0), r = sub(x, p0);
    in &= (mixed(e, q_face_n0(s, adj[i]), r) >= 0);
  }
  return in;
}

/* CheckImminentVFQuadrature, ImminentRepulsionHelper.cu:10-47 */
static int q_vf_quad(const OScene* s, uint32_t node, uint32_t gf, OQuad* out) {
  const uint32_t f[3] = {q_face_node(s, gf, 0), q_face_node(s, gf, 1), q_face_node(s, gf, 2)};
  if (node == f[0] || node == f[1] || node == f[2]) return 0;
  const uint32_t nodes[4] = {f[0], f[1], f[2], node};
  v3 fn;
  float w[4] = {0, 0, 0, 0}; /* the reference leaves w uninitialised on a degenerate face; such a quad is rejected below */
  const float d = oracle_signed_vf_distance(q_x0(s, node), q_x0(s, f[0]), q_x0(s, f[1]), q_x0(s, f[2]), &fn, w);
  if (!(d < INFINITY)) return 0;
  if (!(min3f(-w[1], -w[2], -w[3]) >= 1e-6f)) return 0;
  for (int i = 0; i < 4; i++) { out->bary[i] = w[(i + 1) % 4]; out->n[i] = nodes[i]; }
  return 1;
}
/* CheckImminentEEQuadrature, ImminentRepulsionHelper.cu:49-80 */
static int q_ee_quad(const OScene* s, uint32_t e0, uint32_t e1, OQuad* out) {
  const uint32_t nodes[4] = {q_edge_node(s, e0, 0), q_edge_node(s, e0, 1), q_edge_node(s, e1, 0), q_edge_node(s, e1, 1)};
  if (nodes[0] == nodes[2] || nodes[0] == nodes[3] || nodes[1] == nodes[2] || nodes[1] == nodes[3]) return 0;
  float w[4] = {0, 0, 0, 0};
  v3 fn;
  const float d = oracle_signed_ee_distance(q_x0(s, nodes[0]), q_x0(s, nodes[1]), q_x0(s, nodes[2]), q_x0(s, nodes[3]), &fn, w);
  if (!(d < INFINITY)) return 0;
  const int inside = min4f(w[0], w[1], -w[2], -w[3]) >= 1e-6f && q_in_edge(s, w[1], e0, e1) && q_in_edge(s, -w[3], e1, e0);
  if (!inside) return 0;
  for (int i = 0; i < 4; i++) { out->bary[i] = w[i]; out->n[i] = nodes[i]; }
  return 1;
}

/* ImminentRepulsion::UpdateProximity + Generate, ImminentRepulsion.cu:21-98: pairs = [cloth x obstacle, self];
 * quads = [EE quads (pair-major, 3x3 slot order), then VF quads] */
void oracle_imminent_generate(OScene* s) {
  OQuality* q = s->q;
  size_t np;
  uint32_t* pairs = q_pairs(s, q->cloth_bvh, s->bvh, 1e-6f, &np);
  free(q->quads);
  q->quads = (OQuad*)malloc(sizeof(OQuad) * (12 * np + 1));
  size_t n = 0;
  for (size_t p = 0; p < np; p++) {
    const uint32_t f0 = pairs[2 * p], f1 = pairs[2 * p + 1];
    if (f0 == f1) continue;
    for (int i = 0; i < 3; i++) {
      if (!rtri_edge(q_face_rtri(s, f0), i)) continue;
      for (int j = 0; j < 3; j++) {
        if (!rtri_edge(q_face_rtri(s, f1), j)) continue;
        if (q_ee_quad(s, q_face_edge(s, f0, i), q_face_edge(s, f1, j), &q->quads[n])) n++;
      }
    }
  }
  for (size_t p = 0; p < np; p++) {
    const uint32_t f0 = pairs[2 * p], f1 = pairs[2 * p + 1];
    if (f0 == f1) continue;
    for (int i = 0; i < 3; i++) {
      if (!rtri_vertex(q_face_rtri(s, f0), i)) continue;
      if (q_vf_quad(s, q_face_node(s, f0, i), f1, &q->quads[n])) n++;
    }
  }
  q->n_quads = n;
  free(pairs);
}

/* Jacobi apply shared by both repulsion solvers: ApplyImminentImpulse_Kernel (ImminentRepulsion.cu:121-136) /
 * ApplyPBDImpulse_Kernel (PBDRepulsion.cu:299-314) */
static void q_apply(OScene* s, float* delta, int* count, float rate) {
  for (uint32_t i = 0; i < s->V; i++) {
    const float c = (float)count[i];
    if (c <= EPS) continue;
    const v3 dx = mul(divs(ld3(delta, i), c), rate);
    st3(s->x, i, add(ld3(s->x, i), dx));
    st3(delta, i, V(0, 0, 0));
    count[i] = 0;
  }
}
/* AccumulateImpulses / AccumulateRepulsions, ImminentRepulsionHelper.cu:192-206, PBDRepulsion.cu:283-297 */
static void q_accumulate(const OScene* s, float* delta, int* count, const uint32_t nodes[4], const v3 corr[4]) {
  for (int i = 0; i < 4; i++) {
    if (len(corr[i]) < EPS) continue;
    if (!q_free(s, nodes[i])) continue;
    const uint32_t l = q_local(s, nodes[i]);
    st3(delta, l, add(ld3(delta, l), corr[i]));
    count[l] += 1;
  }
}

/* ImminentRepulsion::Solve, ImminentRepulsion.cu:153-169, with EvaluateImminentImpulse
 * (ImminentRepulsionHelper.cu:82-150; friction_coeff is 0 there, so the tangential branch adds nothing) */
void oracle_imminent_solve(OScene* s, float dt) {
  OQuality* q = s->q;
  for (uint32_t i = 0; i < s->V; i++) st3(s->v, i, divs(sub(ld3(s->x, i), ld3(s->x0, i)), dt)); /* UpdateVelocity_Kernel */
  const float thickness = q->imminent_thickness;
  for (size_t c = 0; c < q->n_quads; c++) {
    const OQuad* qd = &q->quads[c];
    v3 pr = V(0, 0, 0), vr = V(0, 0, 0);
    for (int i = 0; i < 4; i++) {
      pr = add(pr, mul(q_x0(s, qd->n[i]), qd->bary[i]));
      vr = add(vr, mul(q_v(s, qd->n[i]), qd->bary[i]));
    }
    const float npr = len(pr);
    if (npr > thickness) continue;
    if (npr < 1e-5f) continue;
    pr = normalize(pr);
    const float vr_nrm = dot(vr, pr);
    const float target = (thickness - npr) * 0.1f; /* add_repulsion = true */
    if (vr_nrm > target) continue;
    v3 I = mul(pr, vr_nrm * dt);
    const v3 vr_tan = sub(vr, mul(pr, vr_nrm));
    if (len(vr_tan) > thickness * 0.01f) {
      const float imp_norm = len(I);
      const v3 dir = normalize(vr_tan);
      const float friction_coeff = 0.0f, friction_scale = 1.0f;
      v3 friction;
      if (imp_norm * friction_coeff > len(vr_tan)) friction = neg(vr_tan);
      else friction = mul(mul(neg(dir), imp_norm), friction_coeff);
      I = add(I, mul(friction, friction_scale));
    }
    float inv_cm = 0.0f;
    for (int i = 0; i < 4; i++) inv_cm += qd->bary[i] * qd->bary[i] * q_invm(s, qd->n[i]);
    if (inv_cm < 1e-5f) continue;
    v3 corr[4];
    for (int i = 0; i < 4; i++) corr[i] = mul(I, -q_invm(s, qd->n[i]) * qd->bary[i] / inv_cm);
    q_accumulate(s, q->im_delta, q->im_count, qd->n, corr);
  }
  q_apply(s, q->im_delta, q->im_count, q->relaxation_rate);
}

/* ------------------------------------------------------------------------- */
/* PBD repulsion: repulsion/PBDRepulsion.cu, Proximity.cu                     */
/* ------------------------------------------------------------------------- */

/* Edge::ComputeNormal, Edge.cu:43-58 */
static v3 q_edge_normal(const OScene* s, uint32_t ge) {
  uint32_t adj[4];
  const int na = q_edge_adj(s, ge, adj);
  v3 ans = V(0, 0, 0);
  for (int i = 0; i < na; i++)
    ans = add(ans, face_normal(q_x(s, q_face_node(s, adj[i], 0)), q_x(s, q_face_node(s, adj[i], 1)), q_x(s, q_face_node(s, adj[i], 2))));
  if (na == 0) return ans;
  return divs(ans, (float)na);
}

/* EvaluateVFProximity, PBDRepulsion.cu:329-409 */
static int q_rep_vf(const OScene* s, uint32_t node, uint32_t gf, float h, uint32_t nodes[4], v3 corr[4]) {
  const uint32_t f[3] = {q_face_node(s, gf, 0), q_face_node(s, gf, 1), q_face_node(s, gf, 2)};
  if (node == f[0] || node == f[1] || node == f[2]) return 0;
  nodes[0] = node; nodes[1] = f[0]; nodes[2] = f[1]; nodes[3] = f[2];
  /* Proximity(node, face, 1), Proximity.cu:10-20: side from the frame-start positions */
  v3 pn;
  float w0[4] = {0, 0, 0, 0};
  const float d0 = oracle_signed_vf_distance(q_x0(s, nodes[0]), q_x0(s, nodes[1]), q_x0(s, nodes[2]), q_x0(s, nodes[3]), &pn, w0);
  if (d0 < 0.0f && q_face_free(s, gf)) pn = neg(pn);
  v3 fn;
  float w[4] = {0, 0, 0, 0};
  const float dd = oracle_signed_vf_distance(q_x(s, nodes[0]), q_x(s, nodes[1]), q_x(s, nodes[2]), q_x(s, nodes[3]), &fn, w);
  if (!(dd < INFINITY)) return 0; /* degenerate: w is uninitialised in the reference */
  if (!(min3f(-w[1], -w[2], -w[3]) >= 1e-6f)) return 0;
  v3 n = face_normal(q_x(s, nodes[1]), q_x(s, nodes[2]), q_x(s, nodes[3]));
  const float side = dot(n, pn) > 0.0f ? 1.0f : -1.0f;
  v3 qq = q_x(s, nodes[0]), p1 = q_x(s, nodes[1]), p2 = q_x(s, nodes[2]), p3 = q_x(s, nodes[3]);
  qq = sub(qq, p1); p2 = sub(p2, p1); p3 = sub(p3, p1);
  n = mul(cross(p2, p3), side);
  const float c23 = len(n);
  n = normalize(n);
  const float C = dot(qq, n) - h;
  if (C > 0) return 0;
  const v3 dcq = n;
  const v3 dcp2 = divs(add(cross(p3, qq), mul(cross(n, p3), dot(n, qq))), c23);
  const v3 dcp3 = neg(divs(add(cross(p2, qq), mul(cross(n, p2), dot(n, qq))), c23));
  const v3 dcp1 = sub(sub(neg(dcq), dcp2), dcp3);
  const float w_[4] = {q_invm(s, nodes[0]), q_invm(s, nodes[1]), q_invm(s, nodes[2]), q_invm(s, nodes[3])};
  const float denom = w_[0] * dot(dcq, dcq) + w_[1] * dot(dcp1, dcp1) + w_[2] * dot(dcp2, dcp2) + w_[3] * dot(dcp3, dcp3);
  if (denom < EPS) return 0;
  const float sc = C / denom;
  corr[0] = mul(dcq, -w_[0] * sc);
  corr[1] = mul(dcp1, -w_[1] * sc);
  corr[2] = mul(dcp2, -w_[2] * sc);
  corr[3] = mul(dcp3, -w_[3] * sc);
  return 1;
}
/* EvaluateEEProximity, PBDRepulsion.cu:411-468 */
static int q_rep_ee(const OScene* s, uint32_t e0, uint32_t e1, float h, uint32_t nodes[4], v3 corr[4]) {
  nodes[0] = q_edge_node(s, e0, 0); nodes[1] = q_edge_node(s, e0, 1); nodes[2] = q_edge_node(s, e1, 0); nodes[3] = q_edge_node(s, e1, 1);
  if (nodes[0] == nodes[2] || nodes[0] == nodes[3] || nodes[1] == nodes[2] || nodes[1] == nodes[3]) return 0;
  /* Proximity(edge0, edge1, 1), Proximity.cu:22-36 */
  v3 pn;
  float w0[4] = {0, 0, 0, 0};
  const float d0 = oracle_signed_ee_distance(q_x0(s, nodes[0]), q_x0(s, nodes[1]), q_x0(s, nodes[2]), q_x0(s, nodes[3]), &pn, w0);
  if (!q_edge_free(s, e1)) pn = q_edge_normal(s, e1);
  else if (d0 < 0.0f) pn = neg(pn);
  float b[4] = {0, 0, 0, 0};
  v3 n;
  const float dd = oracle_signed_ee_distance(q_x(s, nodes[0]), q_x(s, nodes[1]), q_x(s, nodes[2]), q_x(s, nodes[3]), &n, b);
  if (!(dd < INFINITY)) return 0;
  const int inside = min4f(b[0], b[1], -b[2], -b[3]) >= 1e-6f && q_in_edge(s, b[1], e0, e1) && q_in_edge(s, -b[3], e1, e0);
  if (!inside) return 0;
  v3 d = V(0, 0, 0);
  for (int i = 0; i < 4; i++) d = add(d, mul(q_x(s, nodes[i]), b[i]));
  n = pn;
  const float C = dot(d, n) - h;
  if (C > 0) return 0;
  b[2] *= -1; b[3] *= -1;
  const v3 grads[4] = {mul(n, b[0]), mul(n, b[1]), mul(n, -b[2]), mul(n, -b[3])};
  float denom = 0;
  for (int i = 0; i < 4; i++) denom += q_invm(s, nodes[i]) * b[i] * b[i];
  if (denom < EPS) return 0;
  const float sc = C / denom;
  for (int i = 0; i < 4; i++) corr[i] = mul(grads[i], -sc * q_invm(s, nodes[i]));
  return 1;
}

/* PBDRepulsionSolver::UpdatePairs, PBDRepulsion.cu:22-33 */
void oracle_rep_update_pairs(OScene* s) {
  OQuality* q = s->q;
  oracle_bvh_update(q->cloth_bvh, s->x0, s->x, s->tri, 1);
  free(q->rep_pairs);
  q->rep_pairs = q_pairs(s, q->cloth_bvh, s->bvh, 1e-6f, &q->n_rep_pairs);
}
/* PBDRepulsionSolver::SolveNoProximity, PBDRepulsion.cu:470-516 */
void oracle_rep_solve(OScene* s) {
  OQuality* q = s->q;
  const float h = q->pbd_thickness;
  for (size_t p = 0; p < q->n_rep_pairs; p++) {
    const uint32_t f0 = q->rep_pairs[2 * p], f1 = q->rep_pairs[2 * p + 1];
    uint32_t nodes[4];
    v3 corr[4];
    for (int i = 0; i < 3; i++) {
      if (!rtri_vertex(q_face_rtri(s, f0), i)) continue;
      if (!q_rep_vf(s, q_face_node(s, f0, i), f1, h, nodes, corr)) continue;
      q_accumulate(s, q->rep_delta, q->rep_count, nodes, corr);
    }
    for (int i = 0; i < 3; i++) {
      if (!rtri_edge(q_face_rtri(s, f0), i)) continue;
      for (int j = 0; j < 3; j++) {
        if (!rtri_edge(q_face_rtri(s, f1), j)) continue;
        if (!q_rep_ee(s, q_face_edge(s, f0, i), q_face_edge(s, f1, j), h, nodes, corr)) continue;
        q_accumulate(s, q->rep_delta, q->rep_count, nodes, corr);
      }
    }
  }
  q_apply(s, q->rep_delta, q->rep_count, q->relaxation_rate);
}

/* ------------------------------------------------------------------------- */
/* setup + the Quality frame                                                  */
/* ------------------------------------------------------------------------- */

static void q_face_normals(const float* x, const uint32_t* tri, uint32_t F, float* fn) { /* Face::Update, Face.cu:56-60 */
  for (uint32_t f = 0; f < F; f++) {
    const v3 d1 = sub(ld3(x, tri[3 * f + 1]), ld3(x, tri[3 * f])), d2 = sub(ld3(x, tri[3 * f + 2]), ld3(x, tri[3 * f]));
    st3(fn, f, normalize(cross(d1, d2)));
  }
}
/* face -> edge ids (the reference receives them from the mesh loader; Face::SetEdges, PhysicsMeshHelper.cu:47-50)
 * and edge -> faces ascending (BuildEFAdjacency, ClothSolverHelper.cpp:111-125) */
static void q_topology(const uint32_t* tri, uint32_t F, const uint32_t* edge, uint32_t E, uint32_t nv, uint32_t** fe_out,
                       uint32_t** ef_off_out, uint32_t** ef_idx_out) {
  /* vertex -> incident edges CSR to look an edge up by its end points */
  uint32_t* off = (uint32_t*)calloc(nv + 2, 4);
  for (uint32_t e = 0; e < E; e++) { off[edge[2 * e] + 1]++; off[edge[2 * e + 1] + 1]++; }
  for (uint32_t v = 0; v < nv; v++) off[v + 1] += off[v];
  uint32_t* cur = (uint32_t*)dup_mem(off, 4 * (size_t)(nv + 1));
  uint32_t* inc = (uint32_t*)malloc(8 * (size_t)E + 4);
  for (uint32_t e = 0; e < E; e++) { inc[cur[edge[2 * e]]++] = e; inc[cur[edge[2 * e + 1]]++] = e; }
  uint32_t* fe = (uint32_t*)malloc(12 * (size_t)F + 4);
  uint32_t* cnt = (uint32_t*)calloc(E + 2, 4);
  for (uint32_t f = 0; f < F; f++)
    for (int j = 0; j < 3; j++) {
      const uint32_t a = tri[3 * f + j], b = tri[3 * f + (j + 1) % 3];
      uint32_t found = 0xFFFFFFFFu;
      for (uint32_t i = off[a]; i < off[a + 1]; i++) {
        const uint32_t e = inc[i];
        if ((edge[2 * e] == a && edge[2 * e + 1] == b) || (edge[2 * e] == b && edge[2 * e + 1] == a)) { found = e; break; }
      }
      fe[3 * f + j] = found;
      if (found != 0xFFFFFFFFu) cnt[found + 1]++;
    }
  for (uint32_t e = 0; e < E; e++) cnt[e + 1] += cnt[e];
  uint32_t* ef_idx = (uint32_t*)malloc(4 * (size_t)cnt[E] + 4);
  uint32_t* c2 = (uint32_t*)dup_mem(cnt, 4 * (size_t)(E + 1));
  for (uint32_t f = 0; f < F; f++) /* faces ascending per edge (std::set order) */
    for (int j = 0; j < 3; j++)
      if (fe[3 * f + j] != 0xFFFFFFFFu) ef_idx[c2[fe[3 * f + j]]++] = f;
  free(off); free(cur); free(inc); free(c2);
  *fe_out = fe; *ef_off_out = cnt; *ef_idx_out = ef_idx;
}
static void q_vf_csr(const uint32_t* tri, uint32_t F, uint32_t nv, uint32_t** off_out, uint32_t** idx_out) {
  uint32_t* off = (uint32_t*)calloc(nv + 2, 4);
  for (uint32_t i = 0; i < 3 * F; i++) off[tri[i] + 1]++;
  for (uint32_t v = 0; v < nv; v++) off[v + 1] += off[v];
  uint32_t* cur = (uint32_t*)dup_mem(off, 4 * (size_t)(nv + 1));
  uint32_t* idx = (uint32_t*)malloc(12 * (size_t)F + 4);
  for (uint32_t f = 0; f < F; f++)
    for (int j = 0; j < 3; j++) idx[cur[tri[3 * f + j]]++] = f;
  free(cur);
  *off_out = off; *idx_out = idx;
}

/* qp: imminent_on, pbd_on; qf: material[6], imminent_thickness, pbd_thickness, relaxation_rate, obstacle_mass,
 * impact-zone thickness.  Call after set_cloth / set_obstacle (obstacle optional).  oedge may be NULL without
 * an obstacle.  Strain colours: one per face; bending uses the scene's bend constraints (set_bend) and colours. */
void oracle_scene_set_quality(OScene* s, const int* qp, const float* qf, const uint32_t* oedge, uint32_t Eo,
                              const uint32_t* strain_color) {
  OQuality* q = (OQuality*)calloc(1, sizeof(OQuality));
  s->q = q;
  q->imminent_on = qp[0]; q->pbd_on = qp[1];
  memcpy(q->material, qf, 24);
  q->imminent_thickness = qf[6]; q->pbd_thickness = qf[7]; q->relaxation_rate = qf[8];
  q->obstacle_mass = qf[9]; q->iz_thickness = qf[10];
  q_topology(s->tri, s->F, s->edge, s->E, s->V, &q->fe, &q->ef_off, &q->ef_idx);
  q_vf_csr(s->tri, s->F, s->V, &s->q->vf_off, &s->q->vf_idx);
  q->crtri = (uint32_t*)malloc(4 * (size_t)s->F + 4);
  oracle_rtri(s->tri, s->q->vf_idx, s->q->vf_off, s->F, q->crtri);
  q->fn0 = (float*)malloc(12 * (size_t)s->F + 4);
  q_face_normals(s->x0, s->tri, s->F, q->fn0);
  q->x1 = (float*)dup_mem(s->x0, 12 * (size_t)s->V);
  if (s->Vo) {
    q->Eo = Eo;
    q->oedge = (uint32_t*)dup_mem(oedge, 8 * (size_t)Eo);
    q_topology(s->otri, s->Fo, q->oedge, Eo, s->Vo, &q->ofe, &q->oef_off, &q->oef_idx);
    q_vf_csr(s->otri, s->Fo, s->Vo, &s->q->ovf_off, &s->q->ovf_idx);
    q->ofn0 = (float*)malloc(12 * (size_t)s->Fo + 4);
    q_face_normals(s->ox0, s->otri, s->Fo, q->ofn0);
    q->ov = (float*)calloc(3 * (size_t)s->Vo + 1, 4);
    q->oxq1 = (float*)dup_mem(s->ox0, 12 * (size_t)s->Vo);
  }
  /* FEM (SetupInternalDynamics, ClothSolverBase.cpp:468-486): strain per face, isometric bending per hinge */
  q->strain = (OStrain*)malloc(sizeof(OStrain) * (s->F ? s->F : 1));
  q->strain_color = (uint32_t*)dup_mem(strain_color, 4 * (size_t)s->F);
  for (uint32_t i = 0; i < s->F; i++) {
    q->strain[i] = fem_init_strain(s->rest, s->tri, i);
    if (strain_color[i] + 1 > q->n_strain_colors) q->n_strain_colors = strain_color[i] + 1;
  }
  q->n_iso = s->nb;
  q->iso = (OIsoBend*)malloc(sizeof(OIsoBend) * (s->nb ? s->nb : 1));
  q->iso_lambda = (float*)calloc(s->nb + 1, 4);
  for (uint32_t i = 0; i < s->nb; i++) q->iso[i] = fem_init_bend(s->rest, s->b_idx, i);
  q->n_iso_colors = s->n_bc;
  q->im_delta = (float*)calloc(3 * (size_t)s->V + 1, 4); q->im_count = (int*)calloc(s->V + 1, 4);
  q->rep_delta = (float*)calloc(3 * (size_t)s->V + 1, 4); q->rep_count = (int*)calloc(s->V + 1, 4);
  q->removed = (int*)calloc(s->V + s->Vo + 1, 4); q->min_index = (int*)calloc(s->V + s->Vo + 1, 4);
}

/* one iteration of the Quality inner loop, ClothSolverBase.cpp:296-328 */
void oracle_quality_iteration(OScene* s, float dt, int iteration) {
  OQuality* q = s->q;
  if (q->imminent_on) oracle_imminent_solve(s, dt);
  if (q->pbd_on) {
    if (iteration % 10 == 0) oracle_rep_update_pairs(s);
    oracle_rep_solve(s);
  }
  for (uint32_t c = 0; c < q->n_strain_colors; c++)
    for (uint32_t i = 0; i < s->F; i++)
      if (q->strain_color[i] == c) fem_solve_strain(s->x, s->invm, &q->strain[i], q->material);
  if (s->bending)
    for (uint32_t c = 0; c < q->n_iso_colors; c++)
      for (uint32_t i = 0; i < q->n_iso; i++)
        if (s->b_color[i] == c) fem_solve_bend(s->x, s->invm, &q->iso[i], &q->iso_lambda[i], q->material[5], dt, iteration);
  if (s->lra_on) oracle_solve_lra(s);
  if (s->bvh) oracle_solve_binding(s);
}

/* ClothSolverBase::SimulateQuality, ClothSolverBase.cpp:263-362 */
void oracle_step_frame_quality(OScene* s) {
  OQuality* q = s->q;
  const float dt = s->frame_dt / (float)s->S;
  if (s->moved && s->bvh) { /* UpdateObstacleAnimationFrame (impl/smpl/ClothSolver.cpp:257-261) */
    oracle_bvh_free(s->bvh);
    s->bvh = oracle_bvh_build(s->ox0, s->ox, s->on, s->otri, s->Fo, 0.0f, 0);
  }
  if (q->cloth_bvh) oracle_bvh_free(q->cloth_bvh);
  q->cloth_bvh = oracle_bvh_build(s->x0, s->x, NULL, s->tri, s->F, 0.0f, 0);
  if (q->imminent_on) oracle_imminent_generate(s);
  if (s->bvh) oracle_bvh_update(s->bvh, s->ox0, s->ox, s->otri, 1);
  for (int substep = 0; substep < s->S; substep++) {
    oracle_predict(s, dt);
    for (int it = 0; it < s->N; it++) oracle_quality_iteration(s, dt, it);
    oracle_collision_step(s, dt);
    oracle_finalize(s, dt);
  }
  if (s->Vo) {
    memcpy(s->ox0, s->ox, 12 * (size_t)s->Vo); /* PredictedToPositions */
    vertex_normals(s->ox0, s->otri, s->Vo, s->Fo, s->on);
  }
  vertex_normals(s->x0, s->tri, s->V, s->F, s->n);
  s->moved = 0;
}

/* ---- test entry points: the pieces on caller-supplied state ---- */

/* FindImpacts on the scene's current x0 / x with fresh ccd BVHs (as CollisionStep builds them, Entry.cu:149-155).
 * out: rows of 13 floats: node0..3 (global ids as floats are avoided: written to out_nodes), t, w[4], n[3], is_vf. */
size_t oracle_find_impacts(OScene* s, float thickness, uint32_t* out_nodes, float* out_twn, size_t cap, size_t* n_pairs) {
  OBvh* cb = oracle_bvh_build(s->x0, s->x, NULL, s->tri, s->F, 0.0f, 1);
  OBvh* ob = s->Fo ? oracle_bvh_build(s->ox0, s->ox, NULL, s->otri, s->Fo, 0.0f, 1) : NULL;
  size_t n;
  OImpact* imp = q_find_impacts(s, cb, ob, thickness, &n, n_pairs);
  for (size_t i = 0; i < n && i < cap; i++) {
    memcpy(out_nodes + 4 * i, imp[i].n, 16);
    float* o = out_twn + 9 * i;
    o[0] = imp[i].t; memcpy(o + 1, imp[i].w, 16);
    o[5] = imp[i].nrm.x; o[6] = imp[i].nrm.y; o[7] = imp[i].nrm.z; o[8] = (float)imp[i].is_vf;
  }
  free(imp);
  oracle_bvh_free(cb);
  if (ob) oracle_bvh_free(ob);
  return n;
}
/* independentImpacts on a caller-supplied list (same row layout); returns the kept count, rows in sorted order */
size_t oracle_independent_impacts(OScene* s, const uint32_t* nodes, const float* twn, size_t n, int deform, uint32_t* out_nodes,
                                  float* out_twn) {
  OImpact* in = (OImpact*)malloc(sizeof(OImpact) * (n ? n : 1));
  for (size_t i = 0; i < n; i++) {
    memcpy(in[i].n, nodes + 4 * i, 16);
    const float* o = twn + 9 * i;
    in[i].t = o[0]; memcpy(in[i].w, o + 1, 16);
    in[i].nrm = V(o[5], o[6], o[7]); in[i].is_vf = (int)o[8];
  }
  size_t m;
  OImpact* out = q_independent(s, in, n, deform, &m);
  for (size_t i = 0; i < m; i++) {
    memcpy(out_nodes + 4 * i, out[i].n, 16);
    float* o = out_twn + 9 * i;
    o[0] = out[i].t; memcpy(o + 1, out[i].w, 16);
    o[5] = out[i].nrm.x; o[6] = out[i].nrm.y; o[7] = out[i].nrm.z; o[8] = (float)out[i].is_vf;
  }
  free(in); free(out);
  return m;
}
/* IsolateImpacts colours of a caller-supplied impact list */
int oracle_isolate_zones(OScene* s, const uint32_t* nodes, size_t n, int* colors) {
  OImpact* in = (OImpact*)calloc(n ? n : 1, sizeof(OImpact));
  for (size_t i = 0; i < n; i++) memcpy(in[i].n, nodes + 4 * i, 16);
  const int nc = q_isolate(s, in, n, colors);
  free(in);
  return nc;
}
/* the zone optimiser on a caller-supplied impact list with zone colours (ascending); moves the scene's x.
 * stats: outer iterations, line-search steps, all zones converged */
void oracle_optimize_zones(OScene* s, const uint32_t* nodes, const float* twn, const int* colors, size_t n, int n_zones,
                           int deform, float thickness, float obstacle_mass, int* stats) {
  OImpact* in = (OImpact*)malloc(sizeof(OImpact) * (n ? n : 1));
  for (size_t i = 0; i < n; i++) {
    memcpy(in[i].n, nodes + 4 * i, 16);
    const float* o = twn + 9 * i;
    in[i].t = o[0]; memcpy(in[i].w, o + 1, 16);
    in[i].nrm = V(o[5], o[6], o[7]); in[i].is_vf = (int)o[8];
  }
  memcpy(s->q->x1, s->x, 12 * (size_t)s->V);
  if (s->Vo) memcpy(s->q->oxq1, s->ox, 12 * (size_t)s->Vo);
  OSolveStats st = {0};
  q_optimize(s, in, colors, n, n_zones, deform, thickness, obstacle_mass, &st);
  if (stats) { stats[0] = st.outer_iterations; stats[1] = st.line_search_steps; stats[2] = st.all_converged; }
  free(in);
}

/* debug twin of ref_rep_contributions (oracle/ref_harness/harness.cu): every term of one oracle_rep_solve pass on
 * cloth vertex `vertex`; rows: pair, slot (0-2 VF, 3-11 EE), 4 global node ids | corr xyz */
int oracle_rep_contributions(OScene* s, int vertex, int* rows, float* corr_out, int cap) {
  OQuality* q = s->q;
  const float h = q->pbd_thickness;
  int n = 0;
  for (size_t p = 0; p < q->n_rep_pairs; p++) {
    const uint32_t f0 = q->rep_pairs[2 * p], f1 = q->rep_pairs[2 * p + 1];
    for (int slot = 0; slot < 12; slot++) {
      uint32_t nodes[4];
      v3 corr[4];
      int hit;
      if (slot < 3) {
        if (!rtri_vertex(q_face_rtri(s, f0), slot)) continue;
        hit = q_rep_vf(s, q_face_node(s, f0, slot), f1, h, nodes, corr);
      } else {
        const int i = (slot - 3) / 3, j = (slot - 3) % 3;
        if (!rtri_edge(q_face_rtri(s, f0), i) || !rtri_edge(q_face_rtri(s, f1), j)) continue;
        hit = q_rep_ee(s, q_face_edge(s, f0, i), q_face_edge(s, f1, j), h, nodes, corr);
      }
      if (!hit) continue;
      for (int k = 0; k < 4; k++) {
        if ((int)nodes[k] != vertex || len(corr[k]) < EPS || !q_free(s, nodes[k])) continue;
        if (n < cap) {
          rows[6 * n] = (int)p; rows[6 * n + 1] = slot;
          for (int t = 0; t < 4; t++) rows[6 * n + 2 + t] = (int)nodes[t];
          corr_out[3 * n] = corr[k].x; corr_out[3 * n + 1] = corr[k].y; corr_out[3 * n + 2] = corr[k].z;
        }
        n++;
      }
    }
  }
  return n;
}

size_t oracle_quality_num_quads(OScene* s) { return s->q->n_quads; }
void oracle_quality_get_quads(OScene* s, uint32_t* nodes, float* bary) {
  for (size_t i = 0; i < s->q->n_quads; i++) { memcpy(nodes + 4 * i, s->q->quads[i].n, 16); memcpy(bary + 4 * i, s->q->quads[i].bary, 16); }
}
size_t oracle_quality_num_rep_pairs(OScene* s) { return s->q->n_rep_pairs; }
uint32_t* oracle_quality_rep_pairs(OScene* s) { return s->q->rep_pairs; }
void oracle_quality_stats(OScene* s, int* out) {
  out[0] = s->q->cs_rounds; out[1] = s->q->cs_success; out[2] = s->q->cs_total_impacts; out[3] = s->q->cs_deform;
}
uint32_t* oracle_quality_cloth_rtri(OScene* s) { return s->q->crtri; }
float* oracle_scene_ox(OScene* s) { return s->ox; }
/* build the frame's cloth BVH without stepping (test helper for the repulsion entry points) */
void oracle_quality_begin_frame(OScene* s) {
  OQuality* q = s->q;
  if (q->cloth_bvh) oracle_bvh_free(q->cloth_bvh);
  q->cloth_bvh = oracle_bvh_build(s->x0, s->x, NULL, s->tri, s->F, 0.0f, 0);
}

static void oracle_quality_free(OScene* s) {
  OQuality* q = s->q;
  if (!q) return;
  free(q->fe); free(q->ef_off); free(q->ef_idx); free(q->crtri); free(q->fn0);
  free(q->oedge); free(q->ofe); free(q->oef_off); free(q->oef_idx); free(q->ofn0); free(q->ov);
  free(q->x1); free(q->oxq1); free(q->strain); free(q->iso); free(q->iso_lambda); free(q->strain_color);
  if (q->cloth_bvh) oracle_bvh_free(q->cloth_bvh);
  free(q->quads); free(q->im_delta); free(q->im_count); free(q->rep_pairs); free(q->rep_delta); free(q->rep_count);
  free(q->removed); free(q->min_index);
  free(s->q->vf_off); free(s->q->vf_idx); free(s->q->ovf_off); free(s->q->ovf_idx);
  free(q);
  s->q = NULL;
}
