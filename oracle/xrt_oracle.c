/*
 * xrt_oracle.c — CPU restatement of XRTailor's Swift-mode hot path.
 *
 * TEST INFRASTRUCTURE ONLY.  Imported by tests/, __graft_entry__.smoke() and
 * bench.py's cpu_baseline / --impl reference legs as the checker; the product
 * (xrtailor_b200/) never links or calls it.
 *
 * Parity status: pinned.  The reference (openxrlab/xrtailor @ 4eaf5c3) has no tests, golden vectors or CPU path
 * (SURVEY.md section 4); this restatement is pinned against the reference's own CUDA physics, compiled from
 * /root/reference by oracle/ref_harness/Makefile into oracle/_ref/libxrt_ref.so and run on the GPU box:
 * (a) live, in tests/test_ref_parity_gpu.py and tests/test_quality_gpu.py (BVH bit for bit, pair lists row for
 * row, contacts index for index, impacts, positions to the north star's tolerances), and (b) offline, through the
 * fixtures under tests/golden/ that scratch/r02_quality_ref.py wrote from the reference's outputs on that box
 * (tests/test_oracle_cpu.py::test_quality_oracle_against_golden_reference_outputs runs here, without a GPU).
 * Brute-force numpy restatements and hand-derived known answers (tests/test_oracle_cpu.py) cover the rest.
 *
 * Every function cites the reference file:line it follows (paths relative to
 * /root/reference/xrtailor/).  Plain C, scalar, one thread, IEEE arithmetic
 * (the reference's CUDA build uses --use_fast_math; differences are a few ulp).
 */
#include <float.h>
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#define EPS 1e-6f            /* core/Scalar.hpp:39 */
#define SCALAR_MAX FLT_MAX   /* core/Scalar.cuh:10 */
#define MAX_HITS 256         /* core/Common.cuh:23 LBVH_MAX_BUFFER_SIZE */

typedef struct { float x, y, z; } v3;
static v3 V(float x, float y, float z) { v3 r = {x, y, z}; return r; }
static v3 add(v3 a, v3 b) { return V(a.x + b.x, a.y + b.y, a.z + b.z); }
static v3 sub(v3 a, v3 b) { return V(a.x - b.x, a.y - b.y, a.z - b.z); }
static v3 mul(v3 a, float s) { return V(a.x * s, a.y * s, a.z * s); }
static v3 divs(v3 a, float s) { return V(a.x / s, a.y / s, a.z / s); }
static v3 neg(v3 a) { return V(-a.x, -a.y, -a.z); }
static float dot(v3 a, v3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
static v3 cross(v3 a, v3 b) { return V(a.y * b.z - b.y * a.z, a.z * b.x - b.z * a.x, a.x * b.y - b.x * a.y); }
static float len(v3 a) { return sqrtf(dot(a, a)); }
static v3 normalize(v3 a) { return mul(a, 1.0f / sqrtf(dot(a, a))); } /* glm::normalize */
static float fmin2(float a, float b) { return a < b ? a : b; }        /* math/MathFunctions.cuh:24 */
static float fmax2(float a, float b) { return a > b ? a : b; }        /* math/MathFunctions.cuh:39 */
static float clampf(float x, float a, float b) { return x < a ? a : (x > b ? b : x); } /* :75 */
static v3 ld3(const float* p, uint32_t i) { return V(p[3 * i], p[3 * i + 1], p[3 * i + 2]); }
static void st3(float* p, uint32_t i, v3 a) { p[3 * i] = a.x; p[3 * i + 1] = a.y; p[3 * i + 2] = a.z; }
static v3 face_normal(v3 a, v3 b, v3 c) { return normalize(cross(sub(b, a), sub(c, a))); } /* math/MathFunctions.cu:14 */

/* ------------------------------------------------------------------------- */
/* LBVH: physics/broad_phase/lbvh/BVH.cu                                      */
/* ------------------------------------------------------------------------- */

/* expand_bits, BVH.cu:18-24 */
uint32_t oracle_expand_bits(uint32_t v) {
  v = (v * 0x00010001u) & 0xFF0000FFu;
  v = (v * 0x00000101u) & 0x0F00F00Fu;
  v = (v * 0x00000011u) & 0xC30C30C3u;
  v = (v * 0x00000005u) & 0x49249249u;
  return v;
}

static uint32_t quant(float t) {
  t = fmin2(fmax2(t * 1024.0f, 0.0f), 1023.0f);
  if (t != t) return 0; /* NaN (flat scene axis): the GPU cast gives 0 */
  return (uint32_t)t;
}
/* morton_code, BVH.cu:27-40 */
uint32_t oracle_morton_code(float x, float y, float z) {
  return oracle_expand_bits(quant(x)) * 4 + oracle_expand_bits(quant(y)) * 2 + oracle_expand_bits(quant(z));
}

typedef struct {
  uint32_t F;
  uint32_t* morton;  /* [F] sorted */
  uint32_t* object;  /* [F] */
  uint32_t* nodes;   /* [4*(2F-1)] parent,left,right,object */
  float* aabb;       /* [6*(2F-1)] */
  int32_t* esc_int;  /* [F-1] */
  int32_t* esc_ext;  /* [F] */
  float* pred;       /* [9*F] primitive pred1..3 (object order) */
  float* v;          /* [9*F] primitive v1..3 */
  int used64;        /* duplicate Morton codes -> 64-bit keys, BVH.cu:671-683 */
} OBvh;

static int clz32(uint32_t x) { return x ? __builtin_clz(x) : 32; }
static int clz64(uint64_t x) { return x ? __builtin_clzll(x) : 64; }

/* determine_range, BVH.cu:57-105 (32-bit) and :106-156 (64-bit) through one key array */
static int key_delta(const uint64_t* key, int wide, int n, int i, int j) {
  if (j < 0 || j >= n) return -1;
  return wide ? clz64(key[i] ^ key[j]) : clz32((uint32_t)key[i] ^ (uint32_t)key[j]);
}
static void determine_range(const uint64_t* key, int wide, int n, int idx, int* first, int* last) {
  if (idx == 0) { *first = 0; *last = n - 1; return; }
  const int L = key_delta(key, wide, n, idx, idx - 1), R = key_delta(key, wide, n, idx, idx + 1);
  const int d = R > L ? 1 : -1;
  const int dmin = L < R ? L : R;
  int lmax = 2;
  int delta = key_delta(key, wide, n, idx, idx + d * lmax);
  while (delta > dmin) {
    lmax <<= 1;
    delta = key_delta(key, wide, n, idx, idx + d * lmax);
  }
  int l = 0;
  for (int t = lmax >> 1; t > 0; t >>= 1) {
    delta = key_delta(key, wide, n, idx, idx + (l + t) * d);
    if (delta > dmin) l += t;
  }
  int j = idx + l * d;
  if (d < 0) { int t = idx; idx = j; j = t; }
  *first = idx; *last = j;
}
/* find_split, BVH.cu:158-210 */
static int find_split(const uint64_t* key, int wide, int first, int last) {
  if (key[first] == key[last]) return (first + last) >> 1;
  const int dnode = wide ? clz64(key[first] ^ key[last]) : clz32((uint32_t)key[first] ^ (uint32_t)key[last]);
  int split = first, stride = last - first;
  do {
    stride = (stride + 1) >> 1;
    const int mid = split + stride;
    if (mid < last) {
      const int d = wide ? clz64(key[first] ^ key[mid]) : clz32((uint32_t)key[first] ^ (uint32_t)key[mid]);
      if (d > dnode) split = mid;
    }
  } while (stride > 1);
  return split;
}

typedef struct { uint32_t key, idx; } SortItem;
static int cmp_item(const void* a, const void* b) {
  const SortItem *x = (const SortItem*)a, *y = (const SortItem*)b;
  if (x->key != y->key) return x->key < y->key ? -1 : 1;
  return x->idx < y->idx ? -1 : (x->idx > y->idx); /* stable: initial indices ascend */
}

void oracle_bvh_free(OBvh* b) {
  if (!b) return;
  free(b->morton); free(b->object); free(b->nodes); free(b->aabb);
  free(b->esc_int); free(b->esc_ext); free(b->pred); free(b->v);
  free(b);
}

/* BVH::UpdateData (BVH.cu:565-573, UpdateObjects_Kernel :382-421) + BVH::Construct (:618-722).
 * x0 / x / n: float[3*V]; tol = torlence; ccd selects pred = x + tol*n. */
OBvh* oracle_bvh_build(const float* x0, const float* x, const float* nrm, const uint32_t* tri, uint32_t F,
                       float tol, int ccd) {
  OBvh* b = (OBvh*)calloc(1, sizeof(OBvh));
  const uint32_t NN = 2 * F - 1, NI = F - 1;
  b->F = F;
  b->morton = (uint32_t*)malloc(4 * F);
  b->object = (uint32_t*)malloc(4 * F);
  b->nodes = (uint32_t*)malloc(16 * (size_t)NN);
  b->aabb = (float*)malloc(24 * (size_t)NN);
  b->esc_int = (int32_t*)malloc(4 * (F > 1 ? NI : 1));
  b->esc_ext = (int32_t*)malloc(4 * F);
  b->pred = (float*)malloc(36 * (size_t)F);
  b->v = (float*)malloc(36 * (size_t)F);
  float* box = (float*)malloc(24 * (size_t)F);
  float whole[6] = {INFINITY, INFINITY, INFINITY, -INFINITY, -INFINITY, -INFINITY};
  for (uint32_t i = 0; i < F; i++) {
    float lo[3] = {INFINITY, INFINITY, INFINITY}, hi[3] = {-INFINITY, -INFINITY, -INFINITY};
    for (int k = 0; k < 3; k++) {
      const uint32_t vi = tri[3 * i + k];
      v3 nn = nrm ? ld3(nrm, vi) : V(0, 0, 0);
      v3 a = tol != 0.0f ? add(ld3(x0, vi), mul(nn, tol)) : ld3(x0, vi);
      v3 p = ccd ? (tol != 0.0f ? add(ld3(x, vi), mul(nn, tol)) : ld3(x, vi)) : a;
      st3(b->v, 3 * i + k, a);
      st3(b->pred, 3 * i + k, p);
      const float c[6] = {a.x, a.y, a.z, p.x, p.y, p.z};
      for (int q = 0; q < 6; q++) {  /* InitAABBs_Kernel, BVH.cu:423-462 with scr = 0 */
        lo[q % 3] = fmin2(lo[q % 3], c[q]);
        hi[q % 3] = fmax2(hi[q % 3], c[q]);
      }
    }
    for (int k = 0; k < 3; k++) {
      box[6 * i + k] = lo[k]; box[6 * i + 3 + k] = hi[k];
      whole[k] = fmin2(whole[k], lo[k]); whole[3 + k] = fmax2(whole[3 + k], hi[k]); /* thrust::reduce, :642 */
    }
  }
  /* ComputeMortons_Kernel, BVH.cu:464-481 */
  SortItem* items = (SortItem*)malloc(sizeof(SortItem) * F);
  for (uint32_t i = 0; i < F; i++) {
    float p[3];
    for (int k = 0; k < 3; k++) {
      p[k] = 0.5f * (box[6 * i + k] + box[6 * i + 3 + k]);
      p[k] -= whole[k];
      p[k] /= (whole[3 + k] - whole[k]);
    }
    items[i].key = oracle_morton_code(p[0], p[1], p[2]);
    items[i].idx = i;
  }
  qsort(items, F, sizeof(SortItem), cmp_item); /* stable_sort_by_key, :666 */
  int unique = 1;
  for (uint32_t i = 0; i < F; i++) {
    b->morton[i] = items[i].key;
    b->object[i] = items[i].idx;
    if (i && items[i].key == items[i - 1].key) unique = 0;
  }
  b->used64 = !unique;
  uint64_t* key = (uint64_t*)malloc(8 * (size_t)F);
  for (uint32_t i = 0; i < F; i++)
    key[i] = unique ? (uint64_t)b->morton[i] : (((uint64_t)b->morton[i]) << 32) | b->object[i]; /* :675-682 */
  free(items);
  /* default nodes / boxes, leaf nodes: :684-693 */
  for (uint32_t i = 0; i < NN; i++) {
    b->nodes[4 * i] = b->nodes[4 * i + 1] = b->nodes[4 * i + 2] = b->nodes[4 * i + 3] = 0xFFFFFFFFu;
    for (int k = 0; k < 3; k++) { b->aabb[6 * i + k] = INFINITY; b->aabb[6 * i + 3 + k] = -INFINITY; }
  }
  for (uint32_t i = 0; i < F; i++) {
    b->nodes[4 * (NI + i) + 3] = b->object[i];
    memcpy(b->aabb + 6 * (size_t)(NI + i), box + 6 * (size_t)b->object[i], 24); /* AABBs travel with the sort */
  }
  /* ConstructInternalNodes{32,64}Bits_Kernel, :275-318 */
  for (uint32_t idx = 0; idx < NI; idx++) {
    int first, last;
    determine_range(key, !unique, (int)F, (int)idx, &first, &last);
    const int gamma = find_split(key, !unique, first, last);
    uint32_t L = (uint32_t)gamma, R = (uint32_t)gamma + 1;
    if ((first < last ? first : last) == gamma) L += NI;
    if ((first > last ? first : last) == gamma + 1) R += NI;
    b->nodes[4 * idx + 1] = L;
    b->nodes[4 * idx + 2] = R;
    b->nodes[4 * L] = idx;
    b->nodes[4 * R] = idx;
  }
  /* AssignAABBAtomic_Kernel, :540-556: every node pushes its box into all ancestors */
  for (uint32_t i = NI; i < NN; i++) {
    uint32_t parent = b->nodes[4 * i];
    while (parent != 0xFFFFFFFFu) {
      for (int k = 0; k < 3; k++) {
        b->aabb[6 * parent + k] = fmin2(b->aabb[6 * parent + k], b->aabb[6 * i + k]);
        b->aabb[6 * parent + 3 + k] = fmax2(b->aabb[6 * parent + 3 + k], b->aabb[6 * i + 3 + k]);
      }
      parent = b->nodes[4 * parent];
    }
  }
  /* BuildInternalEscapeIndexTable_Kernel, :212-249 */
  for (uint32_t idx = 0; idx < NI; idx++) {
    const uint32_t parent = b->nodes[4 * idx];
    int32_t e = -1;
    if (parent != 0xFFFFFFFFu) {
      if (idx == b->nodes[4 * parent + 1]) e = (int32_t)b->nodes[4 * parent + 2];
      else {
        uint32_t cur = parent;
        for (;;) {
          const uint32_t pp = b->nodes[4 * cur];
          if (pp == 0xFFFFFFFFu) { e = -1; break; }
          if (cur == b->nodes[4 * pp + 2]) cur = pp;
          else { e = (int32_t)b->nodes[4 * pp + 2]; break; }
        }
      }
    }
    b->esc_int[idx] = e;
  }
  /* BuildExternalEscapeIndexTable_Kernel, :251-264 */
  for (uint32_t i = 0; i < F; i++) {
    const uint32_t parent = b->nodes[4 * (NI + i)];
    if (parent == 0xFFFFFFFFu) { b->esc_ext[i] = -1; continue; } /* F == 1 */
    if (NI + i == b->nodes[4 * parent + 1]) b->esc_ext[i] = (int32_t)b->nodes[4 * parent + 2];
    else b->esc_ext[i] = b->esc_int[parent];
  }
  free(key);
  free(box);
  return b;
}

/* BVH::RefitMidstep, BVH.cu:770-818: pred = x1 + tol*n1; leaf boxes only grow;
 * internal boxes are recomputed as the merge of their children (AssignAABBCAS). */
static void refit_internal(OBvh* b, uint32_t node) {
  if (b->nodes[4 * node + 3] != 0xFFFFFFFFu) return;
  const uint32_t L = b->nodes[4 * node + 1], R = b->nodes[4 * node + 2];
  refit_internal(b, L);
  refit_internal(b, R);
  for (int k = 0; k < 3; k++) {
    b->aabb[6 * node + k] = fminf(b->aabb[6 * L + k], b->aabb[6 * R + k]);
    b->aabb[6 * node + 3 + k] = fmaxf(b->aabb[6 * L + 3 + k], b->aabb[6 * R + 3 + k]);
  }
}
void oracle_bvh_refit_midstep(OBvh* b, const float* x1, const float* n1, const uint32_t* tri, float tol) {
  const uint32_t NI = b->F - 1;
  for (uint32_t i = 0; i < b->F; i++) {
    const uint32_t o = b->nodes[4 * (NI + i) + 3];
    for (int k = 0; k < 3; k++) {
      const uint32_t vi = tri[3 * o + k];
      const v3 p = add(ld3(x1, vi), mul(ld3(n1, vi), tol));
      st3(b->pred, 3 * o + k, p);
      float* bb = b->aabb + 6 * (size_t)(NI + i);
      bb[0] = fmin2(bb[0], p.x); bb[1] = fmin2(bb[1], p.y); bb[2] = fmin2(bb[2], p.z);
      bb[3] = fmax2(bb[3], p.x); bb[4] = fmax2(bb[4], p.y); bb[5] = fmax2(bb[5], p.z);
    }
  }
  if (b->F > 1) refit_internal(b, 0);
}

/* CountPairs_Kernel + thrust::inclusive_scan + FindPairs_Kernel, BVH.cu:1149-1238 (self variants :1240-1340
 * are the same walk with lhs == rhs).  Overlap test: Bounds::Overlap(b, thickness) = Overlap(b.Dilate(thickness)),
 * Bounds.cu:95-128.  Pairs are (object of the lhs leaf, object of the rhs leaf); leaf `id` owns
 * pairs[scan[id-1] .. scan[id]-1] and fills it from the top down.  Returns the total; writes at most cap pairs. */
static size_t pairs_walk(const OBvh* l, const OBvh* r, float th, uint32_t id, uint32_t* out /* NULL: count */, long index) {
  const uint32_t nl = l->F - 1 + id, ni = r->F - 1;
  const float* a = l->aabb + 6 * (size_t)nl;
  const float q[6] = {a[0] - th, a[1] - th, a[2] - th, a[3] + th, a[4] + th, a[5] + th};
  const uint32_t obj_l = l->nodes[4 * (size_t)nl + 3];
  size_t n = 0;
  int node = 0;
  while (node != -1) {
    const float* b = r->aabb + 6 * (size_t)node;
    const uint32_t obj = r->nodes[4 * (size_t)node + 3];
    const int leaf = obj != 0xFFFFFFFFu;
    int hit = 1;
    for (int k = 0; k < 3; k++)
      if (b[k] > q[3 + k] || b[3 + k] < q[k]) hit = 0;
    if (hit && leaf) {
      if (out) { out[2 * index] = obj_l; out[2 * index + 1] = obj; index--; }
      n++;
    }
    if (hit && !leaf) node = (int)r->nodes[4 * (size_t)node + 1];
    else node = leaf ? r->esc_ext[(uint32_t)node - ni] : r->esc_int[node];
  }
  return n;
}
size_t oracle_traverse_pairs(const OBvh* l, const OBvh* r, float thickness, uint32_t* pairs, size_t cap) {
  size_t* scan = (size_t*)malloc(sizeof(size_t) * l->F);
  size_t total = 0;
  for (uint32_t id = 0; id < l->F; id++) {
    total += pairs_walk(l, r, thickness, id, NULL, 0);
    scan[id] = total;
  }
  if (pairs && cap >= total)
    for (uint32_t id = 0; id < l->F; id++) pairs_walk(l, r, thickness, id, pairs, (long)scan[id] - 1);
  free(scan);
  return total;
}

/* ---- Quality-mode cloth model ------------------------------------------------
 * FEM strain + isometric bending, physics/dynamics/FEMHelper.cu, in the order of
 * ClothSolverBase::SimulateQuality (ClothSolverBase.cpp:315-324): per iteration every strain colour,
 * then every bending colour; inside a colour in constraint order (a proper colouring makes that order
 * irrelevant).  glm matrices are column-major: m[c][r]. */
typedef struct { uint32_t i0, i1, i2; float area; float inv[2][2]; } OStrain;
typedef struct { uint32_t i[4]; float Q[4][4]; } OIsoBend;

/* ComputeStrainMeta + InitStrain_Kernel, FEMHelper.cu:10-84 */
static OStrain fem_init_strain(const float* rest, const uint32_t* tri, uint32_t id) {
  OStrain c;
  c.i0 = tri[3 * id]; c.i1 = tri[3 * id + 1]; c.i2 = tri[3 * id + 2];
  const v3 p0 = ld3(rest, c.i0), p1 = ld3(rest, c.i1), p2 = ld3(rest, c.i2);
  const v3 n0 = cross(sub(p1, p0), sub(p2, p0));
  c.area = 0.5f * len(n0);
  const v3 a1 = normalize(sub(p1, p0));
  const v3 a2 = normalize(cross(n0, a1));
  const float u0 = dot(p0, a2), v0 = dot(p0, a1), u1 = dot(p1, a2), v1 = dot(p1, a1), u2 = dot(p2, a2), v2 = dot(p2, a1);
  float P[2][2];
  P[0][0] = u0 - u2; P[1][0] = v0 - v2; P[0][1] = u1 - u2; P[1][1] = v1 - v2;
  const float det = P[0][0] * P[1][1] - P[1][0] * P[0][1]; /* glm::determinant(mat2) */
  c.inv[0][0] = c.inv[0][1] = c.inv[1][0] = c.inv[1][1] = 0.0f;
  if (fabsf(det) > EPS) { /* glm::inverse(mat2) */
    const float ood = 1.0f / det;
    c.inv[0][0] = P[1][1] * ood; c.inv[0][1] = -P[0][1] * ood;
    c.inv[1][0] = -P[1][0] * ood; c.inv[1][1] = P[0][0] * ood;
  }
  return c;
}
static float cot_theta(v3 v, v3 w) { return dot(v, w) / len(cross(v, w)); } /* MathFunctions.cu:7-12 */
/* ComputeHessianEnergy + InitIsometricBending_Kernel, FEMHelper.cu:221-305 */
static OIsoBend fem_init_bend(const float* rest, const uint32_t* idx, uint32_t id) {
  OIsoBend c;
  for (int k = 0; k < 4; k++) c.i[k] = idx[4 * id + k];
  const v3 p0 = ld3(rest, c.i[0]), p1 = ld3(rest, c.i[1]), p2 = ld3(rest, c.i[2]), p3 = ld3(rest, c.i[3]);
  const v3 e0 = sub(p1, p0), e1 = sub(p2, p0), e2 = sub(p3, p0), e3 = sub(p2, p1), e4 = sub(p3, p1);
  const float c01 = cot_theta(e0, e1), c02 = cot_theta(e0, e2), c03 = cot_theta(neg(e0), e3), c04 = cot_theta(neg(e0), e4);
  const float A0 = 0.5f * len(cross(sub(p1, p0), sub(p2, p0))), A1 = 0.5f * len(cross(sub(p0, p1), sub(p3, p1)));
  const float coeff = 3.0f / (A0 + A1);
  const float K[4] = {c03 + c04, c01 + c02, -c01 - c03, -c02 - c04};
  for (int i = 0; i < 4; i++)
    for (int j = 0; j < 4; j++) c.Q[j][i] = coeff * (K[i] * K[j]);
  return c;
}
/* ComputeStrainCorrection + SolveStrain_Kernel, FEMHelper.cu:94-212 */
static void fem_solve_strain(float* x, const float* w, const OStrain* c, const float* m) {
  const v3 p0 = ld3(x, c->i0), p1 = ld3(x, c->i1), p2 = ld3(x, c->i2);
  const float w0 = w[c->i0], w1 = w[c->i1], w2 = w[c->i2];
  const float den = 1.0f - m[3] * m[4];
  const float C00 = m[0] / den, C01 = m[0] * m[4] / den, C11 = m[1] / den, C10 = m[1] * m[3] / den, C22 = m[2];
  const v3 p13 = sub(p0, p2), p23 = sub(p1, p2);
  const float a[3] = {p13.x, p13.y, p13.z}, b[3] = {p23.x, p23.y, p23.z};
  float F[3][2];
  for (int i = 0; i < 3; i++) {
    F[i][0] = a[i] * c->inv[0][0] + b[i] * c->inv[1][0];
    F[i][1] = a[i] * c->inv[0][1] + b[i] * c->inv[1][1];
  }
  const float e00 = 0.5f * (F[0][0] * F[0][0] + F[1][0] * F[1][0] + F[2][0] * F[2][0] - 1.0f);
  const float e11 = 0.5f * (F[0][1] * F[0][1] + F[1][1] * F[1][1] + F[2][1] * F[2][1] - 1.0f);
  const float e01 = 0.5f * (F[0][0] * F[0][1] + F[1][0] * F[1][1] + F[2][0] * F[2][1]);
  const float s00 = C00 * e00 + C01 * e11 + 0.0f * e01;
  const float s11 = C10 * e00 + C11 * e11 + 0.0f * e01;
  const float s01 = 0.0f * e00 + 0.0f * e11 + C22 * e01;
  float pk[3][2];
  for (int i = 0; i < 3; i++) {
    pk[i][0] = s00 * F[i][0] + s01 * F[i][1];
    pk[i][1] = s01 * F[i][0] + s11 * F[i][1];
  }
  float psi = 0.0f;
  psi += e00 * s00; psi += e01 * s01; psi += e01 * s01; psi += e11 * s11;
  psi = 0.5f * psi;
  const float energy = c->area * psi;
  float g0[3], g1[3];
  for (int i = 0; i < 3; i++) {
    g0[i] = (c->inv[0][0] * pk[i][0] + c->inv[0][1] * pk[i][1]) * c->area;
    g1[i] = (c->inv[1][0] * pk[i][0] + c->inv[1][1] * pk[i][1]) * c->area;
  }
  const v3 G0 = V(g0[0], g0[1], g0[2]), G1 = V(g1[0], g1[1], g1[2]);
  const v3 G2 = V(-G0.x - G1.x, -G0.y - G1.y, -G0.z - G1.z);
  float sum = w0 * dot(G0, G0);
  sum += w1 * dot(G1, G1);
  sum += w2 * dot(G2, G2);
  if (!(fabsf(sum) > EPS)) return;
  const float s = energy / sum;
  if (w0 > EPS) st3(x, c->i0, add(ld3(x, c->i0), mul(G0, -(s * w0))));
  if (w1 > EPS) st3(x, c->i1, add(ld3(x, c->i1), mul(G1, -(s * w1))));
  if (w2 > EPS) st3(x, c->i2, add(ld3(x, c->i2), mul(G2, -(s * w2))));
}
/* SolveIsometricBending_Kernel, FEMHelper.cu:317-388 */
static void fem_solve_bend(float* x, const float* w, const OIsoBend* c, float* lambda, float stiffness, float dt, int iter) {
  if (iter == 0) *lambda = 0.0f;
  v3 p[4];
  float wi[4];
  for (int k = 0; k < 4; k++) { p[k] = ld3(x, c->i[k]); wi[k] = w[c->i[k]]; }
  float energy = 0.0f;
  for (int k = 0; k < 4; k++)
    for (int j = 0; j < 4; j++) energy += c->Q[j][k] * dot(p[k], p[j]);
  energy *= 0.5f;
  if (fabsf(energy) < EPS) return;
  v3 d[4] = {V(0, 0, 0), V(0, 0, 0), V(0, 0, 0), V(0, 0, 0)};
  for (int i = 0; i < 4; i++)
    for (int j = 0; j < 4; j++) d[j] = add(d[j], mul(p[i], c->Q[j][i]));
  const float n0 = dot(d[0], d[0]), n1 = dot(d[1], d[1]), n2 = dot(d[2], d[2]), n3 = dot(d[3], d[3]);
  const float division = (wi[0] * n0 + wi[1] * n1) + (wi[2] * n2 + wi[3] * n3); /* glm::dot(vec4, vec4) */
  const float alpha = 1.0f / (stiffness * dt * dt);
  const float dl = (-energy - alpha * *lambda) / (division + alpha);
  for (int k = 0; k < 4; k++) st3(x, c->i[k], add(ld3(x, c->i[k]), mul(d[k], dl * wi[k])));
  *lambda += dl;
}
/* x: [3 nv] in/out.  material: xx, yy, xy stiffness, xy / yx Poisson ratio, bending stiffness. */
void oracle_fem_solve(const float* rest, float* x, const float* inv_mass, uint32_t nv, const uint32_t* tri, uint32_t F,
                      const uint32_t* bend, uint32_t B, const float* material, const uint32_t* col_s,
                      const uint32_t* col_b, int iterations, float dt) {
  (void)nv;
  OStrain* cs = (OStrain*)malloc(sizeof(OStrain) * (F ? F : 1));
  OIsoBend* cb = (OIsoBend*)malloc(sizeof(OIsoBend) * (B ? B : 1));
  float* lam = (float*)calloc(B ? B : 1, sizeof(float));
  uint32_t ncs = 0, ncb = 0;
  for (uint32_t i = 0; i < F; i++) { cs[i] = fem_init_strain(rest, tri, i); if (col_s[i] + 1 > ncs) ncs = col_s[i] + 1; }
  for (uint32_t i = 0; i < B; i++) { cb[i] = fem_init_bend(rest, bend, i); if (col_b[i] + 1 > ncb) ncb = col_b[i] + 1; }
  for (int it = 0; it < iterations; it++) {
    for (uint32_t c = 0; c < ncs; c++)
      for (uint32_t i = 0; i < F; i++)
        if (col_s[i] == c) fem_solve_strain(x, inv_mass, &cs[i], material);
    for (uint32_t c = 0; c < ncb; c++)
      for (uint32_t i = 0; i < B; i++)
        if (col_b[i] == c) fem_solve_bend(x, inv_mass, &cb[i], &lam[i], material[5], dt, it);
  }
  free(cs); free(cb); free(lam);
}

/* accessors for ctypes */
uint32_t* oracle_bvh_morton(OBvh* b) { return b->morton; }
uint32_t* oracle_bvh_object(OBvh* b) { return b->object; }
uint32_t* oracle_bvh_nodes(OBvh* b) { return b->nodes; }
float* oracle_bvh_aabb(OBvh* b) { return b->aabb; }
int32_t* oracle_bvh_esc_int(OBvh* b) { return b->esc_int; }
int32_t* oracle_bvh_esc_ext(OBvh* b) { return b->esc_ext; }
float* oracle_bvh_pred(OBvh* b) { return b->pred; }
int oracle_bvh_used64(OBvh* b) { return b->used64; }

/* ------------------------------------------------------------------------- */
/* Bounds tests: physics/broad_phase/Bounds.cu                                */
/* ------------------------------------------------------------------------- */

/* Bounds::Overlap(const Vector3&), Bounds.cu:126-131 */
int oracle_overlap_point(const float* bb, v3 p) {
  return p.x > bb[0] && p.x < bb[3] && p.y > bb[1] && p.y < bb[4] && p.z > bb[2] && p.z < bb[5];
}
/* GetIntersection, Bounds.cu:132-140 */
static int get_isect(float d1, float d2, v3 p1, v3 p2, v3* hit) {
  if ((d1 * d2) >= 0.0f) return 0;
  if (d1 == d2) return 0;
  *hit = add(p1, mul(sub(p2, p1), (-d1 / (d2 - d1))));
  return 1;
}
/* InBox, Bounds.cu:141-150 */
static int in_box(v3 h, const float* bb, int axis) {
  if (axis == 1 && h.z > bb[2] && h.z < bb[5] && h.y > bb[1] && h.y < bb[4]) return 1;
  if (axis == 2 && h.z > bb[2] && h.z < bb[5] && h.x > bb[0] && h.x < bb[3]) return 1;
  if (axis == 3 && h.x > bb[0] && h.x < bb[3] && h.y > bb[1] && h.y < bb[4]) return 1;
  return 0;
}
/* Bounds::Overlap(p1, p2, hit), Bounds.cu:151-183 */
int oracle_overlap_segment(const float* bb, v3 p1, v3 p2) {
  v3 hit;
  if (p2.x < bb[0] && p1.x < bb[0]) return 0;
  if (p2.x > bb[3] && p1.x > bb[3]) return 0;
  if (p2.y < bb[1] && p1.y < bb[1]) return 0;
  if (p2.y > bb[4] && p1.y > bb[4]) return 0;
  if (p2.z < bb[2] && p1.z < bb[2]) return 0;
  if (p2.z > bb[5] && p1.z > bb[5]) return 0;
  if (oracle_overlap_point(bb, p1)) return 1;
  if ((get_isect(p1.x - bb[0], p2.x - bb[0], p1, p2, &hit) && in_box(hit, bb, 1)) ||
      (get_isect(p1.y - bb[1], p2.y - bb[1], p1, p2, &hit) && in_box(hit, bb, 2)) ||
      (get_isect(p1.z - bb[2], p2.z - bb[2], p1, p2, &hit) && in_box(hit, bb, 3)) ||
      (get_isect(p1.x - bb[3], p2.x - bb[3], p1, p2, &hit) && in_box(hit, bb, 1)) ||
      (get_isect(p1.y - bb[4], p2.y - bb[4], p1, p2, &hit) && in_box(hit, bb, 2)) ||
      (get_isect(p1.z - bb[5], p2.z - bb[5], p1, p2, &hit) && in_box(hit, bb, 3)))
    return 1;
  return 0;
}

/* QueryDeviceStackless (point: BVH.cu:957-988, segment: :989-1022).
 * kind 0 = point p1, 1 = segment p1-p2.  Returns hits in traversal order. */
uint32_t oracle_query(const OBvh* b, int kind, v3 p1, v3 p2, uint32_t* out) {
  int node = 0;
  uint32_t found = 0;
  const uint32_t NI = b->F - 1;
  while (node != -1) {
    const uint32_t obj = b->nodes[4 * node + 3];
    const float* bb = b->aabb + 6 * (size_t)node;
    if (found >= MAX_HITS) break;
    const int hit = kind == 0 ? oracle_overlap_point(bb, p1) : oracle_overlap_segment(bb, p1, p2);
    if (hit) {
      if (obj != 0xFFFFFFFFu) {
        out[found++] = obj;
        node = b->esc_ext[node - (int)NI];
      } else
        node = (int)b->nodes[4 * node + 1];
    } else {
      if (obj != 0xFFFFFFFFu) node = b->esc_ext[node - (int)NI];
      else node = b->esc_int[node];
    }
  }
  return found;
}

/* ------------------------------------------------------------------------- */
/* Primitive tests: math/BasicPrimitiveTests.cu                               */
/* ------------------------------------------------------------------------- */

/* NearestPoint(p, Primitive, qs, u, v, w), BasicPrimitiveTests.cu:6-84 */
float oracle_nearest_point(v3 p, v3 a, v3 b, v3 c, v3* qs) {
  const v3 e1 = sub(b, a), e2 = sub(c, a), e3 = sub(c, b);
  const v3 v1 = sub(p, a);
  const float d1 = dot(e1, v1), d2 = dot(e2, v1);
  if (d1 <= 0 && d2 <= 0) { *qs = a; return len(sub(p, *qs)); }
  const v3 v2 = sub(p, b);
  const float d3 = dot(e1, v2), d4 = dot(e2, v2);
  if (d3 >= 0 && d4 <= d3) { *qs = b; return len(sub(p, *qs)); }
  const float vc = d1 * d4 - d3 * d2;
  if (vc <= 0 && d1 >= 0 && d3 <= 0) {
    const float v = d1 / (d1 - d3);
    *qs = add(a, mul(e1, v));
    return len(sub(p, *qs));
  }
  const v3 v3_ = sub(p, c);
  const float d5 = dot(e1, v3_), d6 = dot(e2, v3_);
  if (d6 >= 0 && d5 <= d6) { *qs = c; return len(sub(p, *qs)); }
  const float vb = d5 * d2 - d1 * d6;
  if (vb <= 0 && d2 >= 0 && d6 <= 0) {
    const float w = d2 / (d2 - d6);
    *qs = add(a, mul(e2, w));
    return len(sub(p, *qs));
  }
  const float va = d3 * d6 - d5 * d4;
  if (va <= 0 && (d4 - d3) >= 0 && (d5 - d6) >= 0) {
    const float w = (d4 - d3) / ((d4 - d3) + (d5 - d6));
    *qs = add(b, mul(e3, w));
    return len(sub(p, *qs));
  }
  const float d = 1 / (va + vb + vc);
  const float v = vb * d, w = vc * d;
  *qs = add(add(a, mul(e1, v)), mul(e2, w));
  return len(sub(p, *qs));
}

/* RayIntersect(o, d, ray_length, Primitive), BasicPrimitiveTests.cu:242-284 */
float oracle_ray_intersect(v3 o, v3 d, float ray_length, v3 a, v3 b, v3 c) {
  const v3 e1 = sub(b, a), e2 = sub(c, a);
  const v3 n = cross(d, e2);
  const float det = dot(e1, n);
  if (fabsf(det) < EPS) return SCALAR_MAX;
  const float inv = 1 / det;
  const v3 t = sub(o, a);
  const float u = dot(t, n) * inv;
  if (u < 0 || u > 1) return SCALAR_MAX;
  const v3 q = cross(t, e1);
  const float v = dot(d, q) * inv;
  if (v < 0 || v + u > 1) return SCALAR_MAX;
  const float s = dot(e2, q) * inv;
  if (s > 0 && s < ray_length) return s;
  return INFINITY;
}

/* ClosestPtSegmentSegment, BasicPrimitiveTests.cu:303-361 */
float oracle_closest_seg_seg(v3 p1, v3 q1, v3 p2, v3 q2, float* s, float* t, v3* c1, v3* c2) {
  const v3 d1 = sub(q1, p1), d2 = sub(q2, p2), r = sub(p1, p2);
  const float a = dot(d1, d1), e = dot(d2, d2), f = dot(d2, r);
  if (a <= EPS && e <= EPS) {
    *s = *t = 0;
    *c1 = p1; *c2 = p2;
    return len(sub(*c1, *c2));
  }
  if (a <= EPS) {
    *s = 0;
    *t = clampf(f / e, 0.0f, 1.0f);
  } else {
    const float c = dot(d1, r);
    if (e <= EPS) {
      *t = 0;
      *s = clampf(-c / a, 0.0f, 1.0f);
    } else {
      const float b = dot(d1, d2);
      const float denom = a * e - b * b;
      if (denom > EPS) *s = clampf((b * f - c * e) / denom, 0.0f, 1.0f);
      else *s = 0;
      *t = (b * *s + f) / e;
      if (*t < EPS) {
        *t = 0;
        *s = clampf(-c / a, 0.0f, 1.0f);
      } else if (*t > 1) {
        *t = 1;
        *s = clampf((b - c) / a, 0.0f, 1.0f);
      }
    }
  }
  *c1 = add(p1, mul(d1, *s));
  *c2 = add(p2, mul(d2, *t));
  return len(sub(*c1, *c2));
}

/* flat-array wrappers so Python can call the primitive tests directly */
float oracle_nearest_point_f(const float* p, const float* tri9, float* qs) {
  v3 q;
  float d = oracle_nearest_point(ld3(p, 0), ld3(tri9, 0), ld3(tri9, 1), ld3(tri9, 2), &q);
  st3(qs, 0, q);
  return d;
}
float oracle_ray_intersect_f(const float* o, const float* d, float ray_length, const float* tri9) {
  return oracle_ray_intersect(ld3(o, 0), ld3(d, 0), ray_length, ld3(tri9, 0), ld3(tri9, 1), ld3(tri9, 2));
}
float oracle_closest_seg_seg_f(const float* p1q1p2q2, float* st, float* c1c2) {
  v3 c1, c2;
  float d = oracle_closest_seg_seg(ld3(p1q1p2q2, 0), ld3(p1q1p2q2, 1), ld3(p1q1p2q2, 2), ld3(p1q1p2q2, 3), st,
                                   st + 1, &c1, &c2);
  st3(c1c2, 0, c1);
  st3(c1c2, 1, c2);
  return d;
}
int oracle_overlap_point_f(const float* bb, const float* p) { return oracle_overlap_point(bb, ld3(p, 0)); }
int oracle_overlap_segment_f(const float* bb, const float* p1, const float* p2) {
  return oracle_overlap_segment(bb, ld3(p1, 0), ld3(p2, 0));
}
uint32_t oracle_query_f(const OBvh* b, int kind, const float* p1, const float* p2, uint32_t* out) {
  return oracle_query(b, kind, ld3(p1, 0), p2 ? ld3(p2, 0) : V(0, 0, 0), out);
}

/* RTriBuild_Kernel, physics/representative_triangle/RTriangle.cu:19-74.
 * nb_vf / prefix: vertex -> incident faces, ascending (std::set order). */
void oracle_rtri(const uint32_t* indices, const uint32_t* nb_vf, const uint32_t* nb_vf_prefix, uint32_t F,
                 uint32_t* out) {
  for (uint32_t fid = 0; fid < F; fid++) {
    uint32_t fv[3] = {indices[3 * fid], indices[3 * fid + 1], indices[3 * fid + 2]};
    uint32_t info = 0;
    uint32_t j_start = nb_vf_prefix[fv[0]], j_end = nb_vf_prefix[fv[0] + 1];
    for (uint32_t i_v = 0; i_v < 3; i_v++) {
      uint32_t off = j_start, i_end = j_end;
      uint32_t i_f = nb_vf[off++];
      if (i_f == fid) info |= (1u << i_v);
      const uint32_t j_v = (i_v + 1) % 3;
      j_start = nb_vf_prefix[fv[j_v]];
      j_end = nb_vf_prefix[fv[j_v] + 1];
      uint32_t of = 0xFFFFFFFFu;
      for (uint32_t jno = j_start; jno < j_end; jno++)
        if (nb_vf[jno] == i_f && fid != i_f) { of = i_f; break; }
      for (; off < i_end && of == 0xFFFFFFFFu; off++) {
        i_f = nb_vf[off];
        for (uint32_t jno = j_start; jno < j_end; jno++)
          if (nb_vf[jno] == i_f && fid != i_f) { of = i_f; break; }
      }
      if (fid < of) info |= (1u << (i_v + 3u));
    }
    out[fid] = info;
  }
}

/* ------------------------------------------------------------------------- */
/* Scene state + the Swift frame                                              */
/* ------------------------------------------------------------------------- */

typedef struct OQuality OQuality; /* Quality-mode state, xrt_oracle_quality.c */
typedef struct {
  /* params: SimParams (core/Common.hpp:31-116) + swift self-contact config */
  int S, N, C, max_nb, interleaved, self_collision, bending, lra_on;
  float g[3], damping, relax, lra_stretch, friction, tol, s_compliance, b_compliance, diameter, cell_size, frame_dt;
  /* cloth */
  uint32_t V, E, F;
  float *x0, *x, *v, *invm, *n, *rest;
  uint32_t *tri, *edge;
  /* constraints in the caller's order, colour per constraint like GraphColoring::colors */
  uint32_t ns, nb, nl, n_sc, n_bc;
  uint32_t *s_idx, *s_color, *b_idx, *b_color, *l_src, *l_tgt;
  float *s_rest, *s_lambda, *b_theta, *l_rest;
  /* bindings (BindingItem, dynamics/BindingConstraint.cuh:11-21) */
  uint32_t nbind;
  uint32_t *bd_vert, *bd_tri;
  float *bd_bary, *bd_L, *bd_k;
  /* obstacle */
  uint32_t Vo, Fo;
  float *ox0, *ox, *ox1, *on, *on1;
  uint32_t *otri, *rtri;
  OBvh* bvh;
  int moved;
  /* contacts */
  int32_t *vf_idx, *rt_idx;
  float *vf_q, *rt_q, *ee; /* ee: 14 floats per edge: dist,p2,q2,s,t,fn */
  uint32_t *cnt_vf, *cnt_rt, *cnt_ee;
  /* Jacobi accumulators */
  float* delta;
  int* delta_count;
  /* spatial hash */
  uint32_t *p_hash, *p_index, *cell_start, *cell_end, *neighbors;
  int table_size;
  OQuality* q; /* NULL in Swift mode */
} OScene;
static void oracle_quality_free(OScene* s);

static void* dup_mem(const void* p, size_t bytes) {
  void* r = malloc(bytes ? bytes : 1);
  if (p && bytes) memcpy(r, p, bytes);
  return r;
}

OScene* oracle_scene_new(const int* ip, const float* fp) {
  OScene* s = (OScene*)calloc(1, sizeof(OScene));
  s->S = ip[0]; s->N = ip[1]; s->C = ip[2]; s->max_nb = ip[3]; s->interleaved = ip[4];
  s->self_collision = ip[5]; s->bending = ip[6]; s->lra_on = ip[7];
  s->g[0] = fp[0]; s->g[1] = fp[1]; s->g[2] = fp[2];
  s->damping = fp[3]; s->relax = fp[4]; s->lra_stretch = fp[5]; s->friction = fp[6]; s->tol = fp[7];
  s->s_compliance = fp[8]; s->b_compliance = fp[9]; s->diameter = fp[10]; s->cell_size = fp[11]; s->frame_dt = fp[12];
  return s;
}

/* area-weighted vertex normals: updateNodeNormals_Kernel / UpdateNodeMidstepNormals_Kernel,
 * physics/PhysicsMeshHelper.cu:200-230, then FinalizeNodeNormals (normalize) */
static void vertex_normals(const float* x, const uint32_t* tri, uint32_t V_, uint32_t F, float* n) {
  memset(n, 0, 12 * (size_t)V_);
  for (uint32_t f = 0; f < F; f++)
    for (int j = 0; j < 3; j++) {
      const uint32_t a = tri[3 * f + j];
      const v3 e0 = sub(ld3(x, tri[3 * f + (j + 1) % 3]), ld3(x, a));
      const v3 e1 = sub(ld3(x, tri[3 * f + (j + 2) % 3]), ld3(x, a));
      const v3 w = divs(cross(e0, e1), dot(e0, e0) * dot(e1, e1));
      st3(n, a, add(ld3(n, a), w));
    }
  for (uint32_t i = 0; i < V_; i++) st3(n, i, normalize(ld3(n, i)));
}

void oracle_scene_set_cloth(OScene* s, const float* pos, const float* invm, uint32_t V_, const uint32_t* tri,
                            uint32_t F, const uint32_t* edge, uint32_t E) {
  s->V = V_; s->F = F; s->E = E;
  s->x0 = (float*)dup_mem(pos, 12 * (size_t)V_);
  s->x = (float*)dup_mem(pos, 12 * (size_t)V_);
  s->rest = (float*)dup_mem(pos, 12 * (size_t)V_);
  s->v = (float*)calloc(3 * (size_t)V_ + 1, 4);
  s->n = (float*)calloc(3 * (size_t)V_ + 1, 4);
  s->invm = (float*)dup_mem(invm, 4 * (size_t)V_);
  s->tri = (uint32_t*)dup_mem(tri, 12 * (size_t)F);
  s->edge = (uint32_t*)dup_mem(edge, 8 * (size_t)E);
  s->vf_idx = (int32_t*)malloc(4 * (size_t)V_ + 4); s->rt_idx = (int32_t*)malloc(4 * (size_t)V_ + 4);
  for (uint32_t i = 0; i < V_; i++) s->vf_idx[i] = s->rt_idx[i] = -1;
  s->vf_q = (float*)calloc(3 * (size_t)V_ + 1, 4); s->rt_q = (float*)calloc(3 * (size_t)V_ + 1, 4);
  s->ee = (float*)calloc(14 * (size_t)E + 1, 4);
  for (uint32_t i = 0; i < E; i++) s->ee[14 * i] = SCALAR_MAX;
  s->cnt_vf = (uint32_t*)calloc(V_ + 1, 4); s->cnt_rt = (uint32_t*)calloc(V_ + 1, 4); s->cnt_ee = (uint32_t*)calloc(E + 1, 4);
  s->delta = (float*)calloc(3 * (size_t)V_ + 1, 4);
  s->delta_count = (int*)calloc(V_ + 1, 4);
  s->table_size = 2 * (int)V_; /* SpatialHashGPU ctor, SpatialHashGPU.cu:161-175 */
  s->p_hash = (uint32_t*)calloc(V_ + 1, 4); s->p_index = (uint32_t*)calloc(V_ + 1, 4);
  s->cell_start = (uint32_t*)calloc((size_t)s->table_size + 1, 4);
  s->cell_end = (uint32_t*)calloc((size_t)s->table_size + 1, 4);
  s->neighbors = (uint32_t*)malloc(4 * (size_t)V_ * s->max_nb + 4);
  memset(s->neighbors, 0xFF, 4 * (size_t)V_ * s->max_nb);
  vertex_normals(s->x0, s->tri, V_, F, s->n);
}

void oracle_scene_set_stretch(OScene* s, const uint32_t* idx, const float* rest, const uint32_t* color, uint32_t n) {
  s->ns = n;
  s->s_idx = (uint32_t*)dup_mem(idx, 8 * (size_t)n);
  s->s_rest = (float*)dup_mem(rest, 4 * (size_t)n);
  s->s_color = (uint32_t*)dup_mem(color, 4 * (size_t)n);
  s->s_lambda = (float*)calloc(n + 1, 4);
  s->n_sc = 0;
  for (uint32_t i = 0; i < n; i++) if (color[i] + 1 > s->n_sc) s->n_sc = color[i] + 1;
}
void oracle_scene_set_bend(OScene* s, const uint32_t* idx, const float* theta, const uint32_t* color, uint32_t n) {
  s->nb = n;
  s->b_idx = (uint32_t*)dup_mem(idx, 16 * (size_t)n);
  s->b_theta = (float*)dup_mem(theta, 4 * (size_t)n);
  s->b_color = (uint32_t*)dup_mem(color, 4 * (size_t)n);
  s->n_bc = 0;
  for (uint32_t i = 0; i < n; i++) if (color[i] + 1 > s->n_bc) s->n_bc = color[i] + 1;
}
void oracle_scene_set_lra(OScene* s, const uint32_t* src, const uint32_t* tgt, const float* rest, uint32_t n) {
  s->nl = n;
  s->l_src = (uint32_t*)dup_mem(src, 4 * (size_t)n);
  s->l_tgt = (uint32_t*)dup_mem(tgt, 4 * (size_t)n);
  s->l_rest = (float*)dup_mem(rest, 4 * (size_t)n);
}

/* AddObstacle + ClothSolverUniversal::Start (impl/universal/ClothSolver.cpp:46-66):
 * normals, BVH UpdateData(0, false) + Construct, RTriangle */
void oracle_scene_set_obstacle(OScene* s, const float* pos, uint32_t Vo, const uint32_t* tri, uint32_t Fo) {
  s->Vo = Vo; s->Fo = Fo;
  s->ox0 = (float*)dup_mem(pos, 12 * (size_t)Vo);
  s->ox = (float*)dup_mem(pos, 12 * (size_t)Vo);
  s->ox1 = (float*)dup_mem(pos, 12 * (size_t)Vo);
  s->on = (float*)calloc(3 * (size_t)Vo + 1, 4);
  s->on1 = (float*)calloc(3 * (size_t)Vo + 1, 4);
  s->otri = (uint32_t*)dup_mem(tri, 12 * (size_t)Fo);
  vertex_normals(s->ox0, s->otri, Vo, Fo, s->on);
  s->bvh = oracle_bvh_build(s->ox0, s->ox, s->on, s->otri, Fo, 0.0f, 0);
  /* VF adjacency (BuildVFAdjacency, ClothSolverHelper.cu:377-401) -> RTriangle::Init */
  uint32_t* prefix = (uint32_t*)calloc(Vo + 2, 4);
  for (uint32_t i = 0; i < 3 * Fo; i++) prefix[tri[i] + 1]++;
  for (uint32_t v = 0; v < Vo; v++) prefix[v + 1] += prefix[v];
  uint32_t* nb = (uint32_t*)malloc(12 * (size_t)Fo + 4);
  uint32_t* cur = (uint32_t*)dup_mem(prefix, 4 * (size_t)(Vo + 1));
  for (uint32_t f = 0; f < Fo; f++)
    for (int j = 0; j < 3; j++) nb[cur[tri[3 * f + j]]++] = f;
  s->rtri = (uint32_t*)malloc(4 * (size_t)Fo + 4);
  oracle_rtri(tri, nb, prefix, Fo, s->rtri);
  free(prefix); free(nb); free(cur);
}

void oracle_scene_set_obstacle_target(OScene* s, const float* xyz) {
  memcpy(s->ox, xyz, 12 * (size_t)s->Vo);
  s->moved = 1;
}

/* PredictPositions_Kernel, ClothSolverHelper.cu:125-131 */
void oracle_predict(OScene* s, float dt) {
  for (uint32_t i = 0; i < s->V; i++) {
    if (s->invm[i] > EPS) st3(s->v, i, add(ld3(s->v, i), mul(V(s->g[0], s->g[1], s->g[2]), dt)));
    st3(s->x, i, add(ld3(s->x0, i), mul(ld3(s->v, i), dt)));
  }
}
/* Finalize_Kernel, ClothSolverHelper.cu:222-229 */
void oracle_finalize(OScene* s, float dt) {
  for (uint32_t i = 0; i < s->V; i++) {
    const v3 np = ld3(s->x, i);
    const v3 raw = divs(sub(np, ld3(s->x0, i)), dt);
    st3(s->v, i, mul(raw, (1 - s->damping * dt)));
    st3(s->x0, i, np);
  }
}
/* SolveStretch_Kernel, BasicSolverHelper.cu:80-127: full sweep, colour filter */
void oracle_solve_stretch(OScene* s, uint32_t color, float dt, int iter) {
  for (uint32_t id = 0; id < s->ns; id++) {
    if (s->s_color[id] != color) continue;
    if (iter == 0) s->s_lambda[id] = 0;
    const uint32_t i1 = s->s_idx[2 * id], i2 = s->s_idx[2 * id + 1];
    const float rest = s->s_rest[id];
    const v3 diff = sub(ld3(s->x, i1), ld3(s->x, i2));
    const float distance = len(diff);
    const float w1 = s->invm[i1], w2 = s->invm[i2];
    if (distance != rest && w1 + w2 > 0.0f) {
      const v3 gradient = divs(diff, distance + EPS);
      const float alpha_tilde = s->s_compliance / (dt * dt);
      const float constraint = distance - rest;
      const float dl = (-constraint - alpha_tilde * s->s_lambda[id]) / (w1 + w2 + alpha_tilde);
      const v3 corr = mul(gradient, dl);
      s->s_lambda[id] += dl;
      st3(s->x, i1, add(ld3(s->x, i1), mul(corr, w1)));
      st3(s->x, i2, add(ld3(s->x, i2), mul(neg(corr), w2)));
    }
  }
}
/* SolveBend_Kernel, BasicSolverHelper.cu:8-66 */
void oracle_solve_bend(OScene* s, uint32_t color, float dt) {
  for (uint32_t id = 0; id < s->nb; id++) {
    if (s->b_color[id] != color) continue;
    const uint32_t* q = s->b_idx + 4 * (size_t)id;
    const float w1 = s->invm[q[0]], w2 = s->invm[q[1]], w3 = s->invm[q[2]], w4 = s->invm[q[3]];
    const v3 p1 = ld3(s->x, q[0]);
    const v3 p2 = sub(ld3(s->x, q[1]), p1), p3 = sub(ld3(s->x, q[2]), p1), p4 = sub(ld3(s->x, q[3]), p1);
    const v3 n1 = normalize(cross(p2, p3)), n2 = normalize(cross(p2, p4));
    const float d = clampf(dot(n1, n2), 0.0f, 1.0f);
    const float angle = acosf(d);
    if (angle < EPS || d != d) continue;
    const float l23 = len(cross(p2, p3)) + EPS, l24 = len(cross(p2, p4)) + EPS;
    const v3 q3 = divs(add(cross(p2, n2), mul(cross(n1, p2), d)), l23);
    const v3 q4 = divs(add(cross(p2, n1), mul(cross(n2, p2), d)), l24);
    const v3 q2 = sub(neg(divs(add(cross(p3, n2), mul(cross(n1, p3), d)), l23)),
                      divs(add(cross(p4, n1), mul(cross(n2, p4), d)), l24));
    const v3 q1 = sub(sub(neg(q2), q3), q4);
    const float xpbd = s->b_compliance / dt / dt;
    const float denom = xpbd + (w1 * dot(q1, q1) + w2 * dot(q2, q2) + w3 * dot(q3, q3) + w4 * dot(q4, q4));
    if (denom < EPS) continue;
    const float lambda = sqrtf(1.0f - d * d) * (angle - s->b_theta[id]) / denom;
    st3(s->x, q[0], add(ld3(s->x, q[0]), mul(q1, w1 * lambda)));
    st3(s->x, q[1], add(ld3(s->x, q[1]), mul(q2, w2 * lambda)));
    st3(s->x, q[2], add(ld3(s->x, q[2]), mul(q3, w3 * lambda)));
    st3(s->x, q[3], add(ld3(s->x, q[3]), mul(q4, w4 * lambda)));
  }
}
/* SolveGeodesicLRA_Kernel, LongRangeAttachments.cu:5-26 */
void oracle_solve_lra(OScene* s) {
  for (uint32_t i = 0; i < s->nl; i++) {
    const uint32_t src = s->l_src[i], tgt = s->l_tgt[i];
    const float target = s->l_rest[i] * s->lra_stretch;
    if (s->invm[tgt] < EPS && target > 0) continue;
    const v3 diff = sub(ld3(s->x, tgt), ld3(s->x, src));
    const float dist = len(diff);
    if (dist > target) st3(s->x, tgt, add(ld3(s->x, tgt), add(neg(diff), mul(divs(diff, dist), target))));
  }
}
/* BindingConstraint::Add + GenerateBinding_Kernel (BindingConstraint.cu:11-48): the
 * barycentric anchor (skin_params) is passed in; rest length = given distance, or the
 * setup-time distance to the anchor when that is < EPSILON.  Call after set_cloth and
 * set_obstacle. */
void oracle_scene_set_binding(OScene* s, const uint32_t* vert, const uint32_t* tri3, const float* bary3,
                              const float* stiffness, const float* distance, uint32_t n) {
  s->nbind = n;
  s->bd_vert = (uint32_t*)dup_mem(vert, 4 * (size_t)n);
  s->bd_tri = (uint32_t*)dup_mem(tri3, 12 * (size_t)n);
  s->bd_bary = (float*)dup_mem(bary3, 12 * (size_t)n);
  s->bd_k = (float*)dup_mem(stiffness, 4 * (size_t)n);
  s->bd_L = (float*)dup_mem(distance, 4 * (size_t)n);
  for (uint32_t i = 0; i < n; i++) {
    const v3 pc = ld3(s->x0, vert[i]);
    const v3 pb = add(add(mul(ld3(s->ox0, tri3[3 * i]), bary3[3 * i]), mul(ld3(s->ox0, tri3[3 * i + 1]), bary3[3 * i + 1])),
                      mul(ld3(s->ox0, tri3[3 * i + 2]), bary3[3 * i + 2]));
    if (s->bd_L[i] < EPS) s->bd_L[i] = len(sub(pc, pb));
  }
}
/* SolveBinding_Kernel, BindingConstraint.cu:56-92: one-sided tether of a cloth vertex to
 * a point of an obstacle triangle at its frame-end position (Node::x) */
void oracle_solve_binding(OScene* s) {
  for (uint32_t i = 0; i < s->nbind; i++) {
    const uint32_t cv = s->bd_vert[i];
    const float inv_mass = s->invm[cv];
    if (inv_mass < EPS) continue;
    const float* b = s->bd_bary + 3 * i;
    const uint32_t* t = s->bd_tri + 3 * i;
    const v3 po = add(add(mul(ld3(s->ox, t[0]), b[0]), mul(ld3(s->ox, t[1]), b[1])), mul(ld3(s->ox, t[2]), b[2]));
    const v3 d = sub(ld3(s->x, cv), po);
    const float C = len(d) - s->bd_L[i];
    const v3 n = normalize(d);
    float sc = C / inv_mass;
    if (C < 0.0f) continue;
    sc *= s->bd_k[i];
    st3(s->x, cv, add(ld3(s->x, cv), mul(n, -sc * inv_mass)));
  }
}
/* ApplyDeltas_Kernel, ClothSolverHelper.cu:139-148 */
void oracle_apply_deltas(OScene* s) {
  for (uint32_t i = 0; i < s->V; i++) {
    const float count = (float)s->delta_count[i];
    if (count > 0.0f) {
      st3(s->x, i, add(ld3(s->x, i), mul(divs(ld3(s->delta, i), count), s->relax)));
      st3(s->delta, i, V(0, 0, 0));
      s->delta_count[i] = 0;
    }
  }
}

/* ------------------------------------------------------------------------- */
/* Obstacle contact: physics/pbd_collision/PBDCollision.cu                    */
/* ------------------------------------------------------------------------- */

/* GenerateVF_DCD_Kernel :17-47, GenerateVFCCD_Kernel :236-268, GenerateEE_DCD_Kernel :100-153 */
void oracle_generate_contacts(OScene* s) {
  const OBvh* b = s->bvh;
  if (!b) return;
  uint32_t buffer[MAX_HITS];
  for (uint32_t i = 0; i < s->V; i++) {
    const v3 pred = ld3(s->x, i);
    uint32_t nf = oracle_query(b, 0, pred, pred, buffer);
    float hit = SCALAR_MAX;
    int hit_idx = -1;
    v3 hit_qs = V(0, 0, 0);
    for (uint32_t k = 0; k < nf; k++) {
      v3 qs;
      const uint32_t o = buffer[k];
      const float d = oracle_nearest_point(pred, ld3(b->pred, 3 * o), ld3(b->pred, 3 * o + 1), ld3(b->pred, 3 * o + 2), &qs);
      if (d < hit) { hit = d; hit_qs = qs; hit_idx = (int)o; }
    }
    s->vf_idx[i] = hit_idx;
    st3(s->vf_q, i, hit_qs);
    s->cnt_vf[i] = nf;
    /* vertex ray */
    s->rt_idx[i] = -1;
    st3(s->rt_q, i, V(0, 0, 0));
    s->cnt_rt[i] = 0;
    const v3 p1 = ld3(s->x0, i), p2 = pred;
    v3 d = sub(p2, p1);
    const float length = len(d);
    if (length < EPS) continue;
    d = divs(d, length);
    nf = oracle_query(b, 1, p1, p2, buffer);
    s->cnt_rt[i] = nf;
    hit = SCALAR_MAX;
    hit_idx = -1;
    v3 qc = V(0, 0, 0);
    for (uint32_t k = 0; k < nf; k++) {
      const uint32_t o = buffer[k];
      const float dist = oracle_ray_intersect(p1, d, length, ld3(b->pred, 3 * o), ld3(b->pred, 3 * o + 1), ld3(b->pred, 3 * o + 2));
      if (dist < hit) { qc = add(p1, mul(d, dist)); hit = dist; hit_idx = (int)o; }
    }
    s->rt_idx[i] = hit_idx;
    st3(s->rt_q, i, qc);
  }
  for (uint32_t i = 0; i < s->E; i++) {
    const v3 pc = ld3(s->x, s->edge[2 * i]), qc = ld3(s->x, s->edge[2 * i + 1]);
    const uint32_t nf = oracle_query(b, 1, pc, qc, buffer);
    s->cnt_ee[i] = nf;
    float ss = 0, tt = 0, min_dist = SCALAR_MAX;
    v3 p2 = V(0, 0, 0), q2 = V(0, 0, 0), fn = V(0, 0, 0);
    for (uint32_t k = 0; k < nf; k++) {
      const uint32_t f = buffer[k];
      const uint32_t info = s->rtri[f];
      const v3 ov[3] = {ld3(b->pred, 3 * f), ld3(b->pred, 3 * f + 1), ld3(b->pred, 3 * f + 2)};
      for (uint32_t e = 0; e < 3; e++) {
        if ((info >> (e + 3u)) & 1u) { /* RTriEdge */
          float _s, _t;
          v3 c1, c2;
          const uint32_t j = (e + 1u) % 3u;
          if (oracle_closest_seg_seg(pc, qc, ov[e], ov[j], &_s, &_t, &c1, &c2)) {
            const float dist = len(sub(c1, c2));
            if (dist < min_dist) {
              min_dist = dist; p2 = ov[e]; q2 = ov[j]; ss = _s; tt = _t;
              fn = face_normal(ov[0], ov[1], ov[2]);
            }
          }
        }
      }
    }
    float* c = s->ee + 14 * (size_t)i;
    c[0] = min_dist; st3(c + 1, 0, p2); st3(c + 4, 0, q2); c[7] = ss; c[8] = tt; st3(c + 9, 0, fn);
  }
}

/* SolveVFCCD_Kernel :279-297 */
void oracle_solve_vf_ccd(OScene* s) {
  if (!s->bvh) return;
  for (uint32_t i = 0; i < s->V; i++) {
    const int o = s->rt_idx[i];
    if (o == -1) continue;
    const v3 nc = face_normal(ld3(s->bvh->pred, 3 * o), ld3(s->bvh->pred, 3 * o + 1), ld3(s->bvh->pred, 3 * o + 2));
    const float lambda = dot(nc, sub(ld3(s->x, i), ld3(s->rt_q, i)));
    if (lambda < 0) st3(s->x, i, add(ld3(s->x, i), mul(nc, -lambda)));
  }
}
/* SolveVFDCD_Kernel :57-75 */
void oracle_solve_vf_dcd(OScene* s) {
  if (!s->bvh) return;
  for (uint32_t i = 0; i < s->V; i++) {
    const int o = s->vf_idx[i];
    if (o == -1) continue;
    const v3 ns = face_normal(ld3(s->bvh->pred, 3 * o), ld3(s->bvh->pred, 3 * o + 1), ld3(s->bvh->pred, 3 * o + 2));
    const float lambda = dot(ns, sub(ld3(s->x, i), ld3(s->vf_q, i))) - 1e-3f;
    if (lambda < 0) st3(s->x, i, add(ld3(s->x, i), mul(ns, -lambda)));
  }
}
/* SolveEEDCD_Kernel :163-207 */
void oracle_solve_ee_dcd(OScene* s) {
  if (!s->bvh) return;
  for (uint32_t i = 0; i < s->E; i++) {
    const float* c = s->ee + 14 * (size_t)i;
    if (c[0] > 1e6f) continue;
    const uint32_t i1 = s->edge[2 * i], i2 = s->edge[2 * i + 1];
    const v3 p1 = ld3(s->x, i1), q1 = ld3(s->x, i2);
    const v3 p2 = ld3(c + 1, 0), q2 = ld3(c + 4, 0), fn = ld3(c + 9, 0);
    const float ss = c[7], tt = c[8];
    const v3 c1 = add(mul(p1, 1 - ss), mul(q1, ss));
    const v3 c2 = add(mul(p2, 1 - tt), mul(q2, tt));
    const v3 d = sub(c1, c2);
    if (len(d) < EPS) continue;
    const float lambda = dot(d, fn) - 1e-3f;
    if (lambda > 0) continue;
    st3(s->delta, i1, add(ld3(s->delta, i1), mul(fn, -(1 - ss) * lambda)));
    st3(s->delta, i2, add(ld3(s->delta, i2), mul(fn, -(ss)*lambda)));
    s->delta_count[i1]++;
    s->delta_count[i2]++;
  }
}

/* ------------------------------------------------------------------------- */
/* Self contact: broad_phase/spatial_hash/SpatialHashGPU.cu + CollideParticles */
/* ------------------------------------------------------------------------- */

static int int_coord(float v, float cell) {  /* :6-8; clamped like the CUDA path (no effect below 2^30 cells) */
  float c = floorf(v / cell);
  c = c > -1073741824.0f ? c : -1073741824.0f;
  c = c < 1073741824.0f ? c : 1073741824.0f;
  return (int)c;
}
static int hash_coords(int x, int y, int z, int table) {                    /* :9-12 */
  const int h = (int)(((uint32_t)x * 92837111u) ^ ((uint32_t)y * 689287499u) ^ ((uint32_t)z * 283923481u));
  return abs(h % table);
}
static int cmp_hash(const void* a, const void* b) {
  const SortItem *x = (const SortItem*)a, *y = (const SortItem*)b;
  if (x->key != y->key) return x->key < y->key ? -1 : 1;
  return x->idx < y->idx ? -1 : (x->idx > y->idx); /* radix sort is stable */
}
/* HashObjects, SpatialHashGPU.cu:128-160 */
void oracle_hash(OScene* s) {
  const uint32_t n = s->V;
  const float cell = s->diameter * s->cell_size;
  SortItem* it = (SortItem*)malloc(sizeof(SortItem) * (n + 1));
  for (uint32_t i = 0; i < n; i++) {
    const v3 p = ld3(s->x, i);
    it[i].key = (uint32_t)hash_coords(int_coord(p.x, cell), int_coord(p.y, cell), int_coord(p.z, cell), s->table_size);
    it[i].idx = i;
  }
  qsort(it, n, sizeof(SortItem), cmp_hash);
  for (uint32_t i = 0; i < n; i++) { s->p_hash[i] = it[i].key; s->p_index[i] = it[i].idx; }
  free(it);
  memset(s->cell_start, 0xFF, 4 * (size_t)s->table_size);
  for (uint32_t i = 0; i < n; i++) { /* FindCellStart_Kernel :27-55 */
    const uint32_t h = s->p_hash[i];
    if (i == 0 || h != s->p_hash[i - 1]) {
      s->cell_start[h] = i;
      if (i > 0) s->cell_end[s->p_hash[i - 1]] = i;
    }
    if (i == n - 1) s->cell_end[h] = i + 1;
  }
  const float cell2 = cell * cell, diam2 = s->diameter * s->diameter;
  const long long limit = (long long)n * s->max_nb;
  for (uint32_t t = 0; t < n; t++) { /* CacheNeighbors_Kernel :56-100 */
    const uint32_t id = s->p_index[t];
    const v3 p = ld3(s->x, id), r = ld3(s->rest, id);
    const int ix = int_coord(p.x, cell), iy = int_coord(p.y, cell), iz = int_coord(p.z, cell);
    long long ni = id;
    int full = 0;
    for (int x = ix - 1; x <= ix + 1 && !full; x++)
      for (int y = iy - 1; y <= iy + 1 && !full; y++)
        for (int z = iz - 1; z <= iz + 1 && !full; z++) {
          const int h = hash_coords(x, y, z, s->table_size);
          const int start = (int)s->cell_start[h];
          if (start == -1) continue;
          int end = (int)s->cell_end[h];
          if (start + s->max_nb < end) end = start + s->max_nb;
          for (int k = start; k < end; k++) {
            const uint32_t nb = s->p_index[k];
            const v3 d = sub(p, ld3(s->x, nb)), dr = sub(r, ld3(s->rest, nb));
            if (nb != id && dot(d, d) < cell2 && dot(dr, dr) > diam2) {
              s->neighbors[ni] = nb;
              ni += n;
              if (ni >= limit) { full = 1; break; }
            }
          }
        }
    if (!full && ni < limit) s->neighbors[ni] = 0xFFFFFFFFu;
  }
}

/* ComputeFriction + CollideParticles_Kernel, ClothSolverHelper.cu:150-211, then ApplyDeltas */
void oracle_collide_particles(OScene* s) {
  const uint32_t n = s->V;
  for (uint32_t id = 0; id < n; id++) {
    v3 acc = V(0, 0, 0);
    int count = 0;
    const v3 pi = ld3(s->x, id);
    const v3 vi = sub(pi, ld3(s->x0, id));
    const float wi = s->invm[id];
    for (long long k = id; k < (long long)n * s->max_nb; k += n) {
      const uint32_t j = s->neighbors[k];
      if (j > n) break;
      const float denom = wi + s->invm[j];
      if (denom <= 0) continue;
      const v3 pj = ld3(s->x, j);
      const v3 diff = sub(pi, pj);
      const float distance = len(diff);
      if (distance >= s->diameter) continue;
      const v3 grad = divs(diff, distance + EPS);
      const float lambda = (distance - s->diameter) / denom;
      const v3 common = mul(grad, lambda);
      count++;
      acc = sub(acc, mul(common, wi));
      const v3 rel = sub(vi, sub(pj, ld3(s->x0, j)));
      v3 fr = V(0, 0, 0);
      const float cl = len(common);
      if (s->friction > 0 && cl > 0) {
        const v3 nrm = divs(common, cl);
        const v3 tan = sub(rel, mul(nrm, dot(rel, nrm)));
        const float tl = len(tan);
        fr = mul(neg(tan), fmin2(cl * s->friction / tl, 1.0f));
      }
      acc = add(acc, mul(fr, wi));
    }
    st3(s->delta, id, acc);
    s->delta_count[id] = count;
  }
  oracle_apply_deltas(s);
}

/* ------------------------------------------------------------------------- */
/* The frame: ClothSolverBase::SimulateSwift, pipeline/base/ClothSolverBase.cpp:171-247 */
/* ------------------------------------------------------------------------- */

void oracle_collision_pass(OScene* s, int cnt) {
  if (!s->bvh) return;
  if (s->moved) {
    const float alpha = ((float)cnt + 1.0f) / (float)s->C; /* Interpolate_Kernel, PhysicsMesh.cu:610-614 */
    for (uint32_t i = 0; i < s->Vo; i++)
      st3(s->ox1, i, add(mul(ld3(s->ox0, i), 1.0f - alpha), mul(ld3(s->ox, i), alpha)));
    vertex_normals(s->ox1, s->otri, s->Vo, s->Fo, s->on1);
    oracle_bvh_refit_midstep(s->bvh, s->ox1, s->on1, s->otri, s->tol);
  }
  oracle_generate_contacts(s);
}

void oracle_iteration(OScene* s, float dt, int iteration) {
  for (uint32_t c = 0; c < s->n_sc; c++) oracle_solve_stretch(s, c, dt, iteration);
  if (s->bending)
    for (uint32_t c = 0; c < s->n_bc; c++) oracle_solve_bend(s, c, dt);
  if (s->lra_on) oracle_solve_lra(s);
  oracle_solve_vf_ccd(s);
  oracle_solve_vf_dcd(s);
  oracle_solve_ee_dcd(s);
  if (s->bvh) oracle_apply_deltas(s);
  if (s->bvh) oracle_solve_binding(s); /* binding_->Solve, ClothSolverBase.cpp:236 (no-op without an obstacle) */
}

void oracle_step_frame(OScene* s) {
  const float dt = s->frame_dt / (float)s->S;
  if (s->moved && s->bvh) { /* UpdateObstacleAnimationFrame: UpdateData(0,false) + Construct */
    oracle_bvh_free(s->bvh);
    s->bvh = oracle_bvh_build(s->ox0, s->ox, s->on, s->otri, s->Fo, 0.0f, 0);
  }
  uint32_t frequency = (uint32_t)(s->S * s->N) / (uint32_t)s->C;
  if (frequency == 0) frequency = 1;
  int cnt = 0;
  for (int substep = 0; substep < s->S; substep++) {
    oracle_predict(s, dt);
    if (s->self_collision && substep % (s->interleaved > 0 ? s->interleaved : 1) == 0) oracle_hash(s);
    for (int it = 0; it < s->N; it++) {
      if (s->self_collision) oracle_collide_particles(s);
      const uint32_t step = (uint32_t)(substep * s->N + it);
      if (step % frequency == 0) { oracle_collision_pass(s, cnt); cnt++; }
      oracle_iteration(s, dt, it);
    }
    oracle_finalize(s, dt);
  }
  if (s->Vo) {
    if (s->moved) memcpy(s->ox0, s->ox, 12 * (size_t)s->Vo); /* PredictedToPositions */
    vertex_normals(s->ox0, s->otri, s->Vo, s->Fo, s->on);
  }
  vertex_normals(s->x0, s->tri, s->V, s->F, s->n);
  s->moved = 0;
}

/* accessors */
float* oracle_scene_x0(OScene* s) { return s->x0; }
float* oracle_scene_x(OScene* s) { return s->x; }
float* oracle_scene_v(OScene* s) { return s->v; }
float* oracle_scene_n(OScene* s) { return s->n; }
float* oracle_scene_ox0(OScene* s) { return s->ox0; }
float* oracle_scene_on(OScene* s) { return s->on; }
OBvh* oracle_scene_bvh(OScene* s) { return s->bvh; }
uint32_t* oracle_scene_rtri(OScene* s) { return s->rtri; }
int32_t* oracle_scene_vf_idx(OScene* s) { return s->vf_idx; }
float* oracle_scene_vf_q(OScene* s) { return s->vf_q; }
int32_t* oracle_scene_rt_idx(OScene* s) { return s->rt_idx; }
float* oracle_scene_rt_q(OScene* s) { return s->rt_q; }
float* oracle_scene_ee(OScene* s) { return s->ee; }
uint32_t* oracle_scene_cnt(OScene* s, int kind) { return kind == 0 ? s->cnt_vf : (kind == 1 ? s->cnt_rt : s->cnt_ee); }
uint32_t* oracle_scene_neighbors(OScene* s) { return s->neighbors; }
uint32_t* oracle_scene_p_hash(OScene* s) { return s->p_hash; }
uint32_t* oracle_scene_p_index(OScene* s) { return s->p_index; }
void oracle_scene_set_state(OScene* s, const float* x0, const float* v) {
  if (x0) { memcpy(s->x0, x0, 12 * (size_t)s->V); memcpy(s->x, x0, 12 * (size_t)s->V); }
  if (v) memcpy(s->v, v, 12 * (size_t)s->V);
}

void oracle_scene_free(OScene* s) {
  if (!s) return;
  oracle_quality_free(s);
  free(s->x0); free(s->x); free(s->v); free(s->invm); free(s->n); free(s->rest); free(s->tri); free(s->edge);
  free(s->s_idx); free(s->s_color); free(s->b_idx); free(s->b_color); free(s->l_src); free(s->l_tgt);
  free(s->s_rest); free(s->s_lambda); free(s->b_theta); free(s->l_rest);
  free(s->bd_vert); free(s->bd_tri); free(s->bd_bary); free(s->bd_L); free(s->bd_k);
  free(s->ox0); free(s->ox); free(s->ox1); free(s->on); free(s->on1); free(s->otri); free(s->rtri);
  oracle_bvh_free(s->bvh);
  free(s->vf_idx); free(s->rt_idx); free(s->vf_q); free(s->rt_q); free(s->ee);
  free(s->cnt_vf); free(s->cnt_rt); free(s->cnt_ee); free(s->delta); free(s->delta_count);
  free(s->p_hash); free(s->p_index); free(s->cell_start); free(s->cell_end); free(s->neighbors);
  free(s);
}

#include "xrt_oracle_quality.c"
#include "xrt_oracle_rig.c"
