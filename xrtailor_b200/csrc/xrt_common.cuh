// Shared device types and math for the B200 Swift-mode cloth path.
// State is SoA: every per-vertex quantity is a float4 so one 16-byte load
// brings position + inverse mass (the reference chases Node** -> 96-byte Node,
// xrtailor/include/xrtailor/memory/Node.cuh:7-22).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <float.h>
#include <math.h>

#define XRT_EPS 1e-6f           // EPSILON, core/Scalar.hpp:39
#define XRT_SCALAR_MAX FLT_MAX  // SCALAR_MAX, core/Scalar.cuh:10
#define XRT_MAX_HITS 256        // LBVH_MAX_BUFFER_SIZE, core/Common.cuh:23
#define XRT_BLOCK 256

struct f3 {
  float x, y, z;
};
__host__ __device__ __forceinline__ f3 mk3(float x, float y, float z) { return f3{x, y, z}; }
__host__ __device__ __forceinline__ f3 mk3(float4 a) { return f3{a.x, a.y, a.z}; }
__host__ __device__ __forceinline__ f3 operator+(f3 a, f3 b) { return f3{a.x + b.x, a.y + b.y, a.z + b.z}; }
__host__ __device__ __forceinline__ f3 operator-(f3 a, f3 b) { return f3{a.x - b.x, a.y - b.y, a.z - b.z}; }
__host__ __device__ __forceinline__ f3 operator-(f3 a) { return f3{-a.x, -a.y, -a.z}; }
__host__ __device__ __forceinline__ f3 operator*(f3 a, float s) { return f3{a.x * s, a.y * s, a.z * s}; }
__host__ __device__ __forceinline__ f3 operator*(float s, f3 a) { return f3{a.x * s, a.y * s, a.z * s}; }
__host__ __device__ __forceinline__ f3 operator/(f3 a, float s) { return f3{a.x / s, a.y / s, a.z / s}; }
__host__ __device__ __forceinline__ float dot3(f3 a, f3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
__host__ __device__ __forceinline__ f3 cross3(f3 a, f3 b) {
  return f3{a.y * b.z - b.y * a.z, a.z * b.x - b.z * a.x, a.x * b.y - b.x * a.y};
}
__host__ __device__ __forceinline__ float len3(f3 a) { return sqrtf(dot3(a, a)); }
// glm::normalize = v * inversesqrt(dot(v,v))
__host__ __device__ __forceinline__ f3 normalize3(f3 a) {
  return a * (1.0f / sqrtf(dot3(a, a)));
}
__host__ __device__ __forceinline__ float clampf(float x, float lo, float hi) { return fminf(fmaxf(x, lo), hi); }
// MathFunctions::FaceNormal, math/MathFunctions.cu:14-20
__host__ __device__ __forceinline__ f3 face_normal(f3 a, f3 b, f3 c) { return normalize3(cross3(b - a, c - a)); }

// Approximate single-precision helpers of the projection math (stretch, bend,
// self-contact, contact solves).  These are the instructions the reference's
// own build selects with --use_fast_math (CMakeLists.txt:217): MUFU rcp / rsq /
// sqrt and fused multiply-add.  Discrete decisions (Morton bucket, box tests,
// closest feature, hash cell) stay in IEEE arithmetic so candidate sets and
// contact indices are bit-identical to the oracle; only float-valued position
// updates go through here (gate: 1e-5 relative RMS after one substep).
#ifdef __CUDACC__
__device__ __forceinline__ float xf_div(float a, float b) { return __fdividef(a, b); }
__device__ __forceinline__ float xf_rcp(float a) {
  float r;
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(a));
  return r;
}
__device__ __forceinline__ float xf_rsqrt(float a) {
  float r;
  asm("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(a));
  return r;
}
__device__ __forceinline__ float xf_sqrt(float a) {
  float r;
  asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(a));
  return r;
}
__device__ __forceinline__ float xf_dot(f3 a, f3 b) { return fmaf(a.z, b.z, fmaf(a.y, b.y, a.x * b.x)); }
__device__ __forceinline__ f3 xf_cross(f3 a, f3 b) {
  return f3{fmaf(a.y, b.z, -(b.y * a.z)), fmaf(a.z, b.x, -(b.z * a.x)), fmaf(a.x, b.y, -(b.x * a.y))};
}
__device__ __forceinline__ float xf_len(f3 a) { return xf_sqrt(xf_dot(a, a)); }
__device__ __forceinline__ f3 xf_normalize(f3 a) { return a * xf_rsqrt(xf_dot(a, a)); }
__device__ __forceinline__ f3 xf_face_normal(f3 a, f3 b, f3 c) { return xf_normalize(xf_cross(b - a, c - a)); }
// a + s*b
__device__ __forceinline__ f3 xf_axpy(float s, f3 b, f3 a) { return f3{fmaf(s, b.x, a.x), fmaf(s, b.y, a.y), fmaf(s, b.z, a.z)}; }
#endif

// One BVH node packed for traversal: 32 bytes = one DRAM sector.
//   lo.xyz, hi.xyz : box
//   link_hit       : internal: left child index; leaf: ~object_idx (negative)
//   link_miss      : escape index (-1 = done); a leaf goes here hit or miss
struct __align__(16) TNode {
  float lx, ly, lz, ux;
  float uy, uz;
  int32_t link_hit, link_miss;
};

struct BVNode {  // broad_phase/lbvh/BVH.cuh:27-32
  uint32_t parent, left, right, object;
};

// Where node c lives in the traversal array: the root at slot 0, the two
// children of internal node p at slots 2p+2 and 2p+3 (slot 1 unused, 2F slots).
// Siblings are one aligned 64-byte pair, so the load of a left child brings its
// right sibling -- which the stackless traversal always visits next after the
// left subtree -- into the same L1 line.  links in TNode are slots.
__host__ __device__ __forceinline__ uint32_t tnode_child_slot(uint32_t parent, bool right) {
  return 2u + 2u * parent + (right ? 1u : 0u);
}
__host__ __device__ __forceinline__ uint32_t tnode_slot(const BVNode* nodes, uint32_t c) {
  const uint32_t p = nodes[c].parent;
  return p == 0xFFFFFFFFu ? 0u : tnode_child_slot(p, nodes[p].right == c);
}

struct __align__(16) VFContact {  // pbd_collision/Contact.cuh VFContact / RTContact (one 16-byte access)
  int32_t index;
  float q[3];
};
struct __align__(16) EEContact {  // pbd_collision/Contact.cuh EEContact (48 bytes = three 16-byte accesses)
  float dist;
  float p2[3], q2[3];
  float s, t;
  float fn[3];
};

struct DevBVH {
  uint32_t n_obj;       // F
  TNode* tnodes;        // [2F] traversal nodes in sibling-pair slots (tnode_slot)
  BVNode* nodes;        // [2F-1]
  float* aabbs;         // [6*(2F-1)]
  int32_t* esc_int;     // [F-1]
  int32_t* esc_ext;     // [F]
  uint32_t* morton;     // [F] sorted
  uint32_t* sorted_obj; // [F]
  float4* pred;         // [3*F] primitive pred1..3 in object order
};

#define XRT_CUDA_OK(expr)                                                     \
  do {                                                                        \
    cudaError_t e__ = (expr);                                                 \
    if (e__ != cudaSuccess) {                                                 \
      xrt_set_error(s, "%s:%d %s: %s", __FILE__, __LINE__, #expr,             \
                    cudaGetErrorString(e__));                                 \
      return XRT_ERR_CUDA;                                                    \
    }                                                                         \
  } while (0)

static inline unsigned xrt_grid(size_t n) { return (unsigned)((n + XRT_BLOCK - 1) / XRT_BLOCK); }
