// Body-animation producer on the device.
//
// The reference poses its SMPL body every frame with smplx::Body<>::Update (runtime/rag_doll/smpl/body.cpp:270-383,
// cuda/body.cu:212-313): Rodrigues and the joint hierarchy on the host, pose blend shapes and linear blend skinning as
// kernels, then the posed vertices go device -> host (Body::verts()) -> cudaMalloc + host -> device + a kernel that
// applies the model matrix + cudaFree (UpdateObstaclePredicted, physics/ClothSolverHelper.cu:81-97;
// impl/smpl/ClothSolver.cpp:244-249).  Here the rig (shaped rest vertices, CSR skinning weights, optional pose blend
// shapes) stays resident, a frame uploads J x 12 floats of joint transforms (or nothing: a whole pose sequence can be
// made resident), and ONE kernel writes the obstacle's target positions Node::x straight into the solver's float4
// SoA: no posed vertex ever crosses PCIe and nothing is allocated per frame.
#include "xrt_handle.cuh"
#include <cmath>
#include <cstring>
#include <vector>

struct RigState {
  uint32_t J = 0, P = 0;  // joints, pose blend parameters (9 (J - 1) or 0)
  std::vector<float> joints;
  std::vector<int32_t> parent;
  float M[16];
  bool has_M = false;
  float* d_rest = nullptr;       // [3 Vo]
  uint32_t* d_w_outer = nullptr;  // [Vo + 1]
  uint32_t* d_w_inner = nullptr;
  float* d_w_val = nullptr;
  float* d_blend = nullptr;       // [3 Vo][P] row-major or null
  // per frame: [12 J joint transforms | P blend parameters]
  size_t stride = 0;
  float *h_frame = nullptr, *d_frame = nullptr;  // pinned / device, one frame
  cudaEvent_t ev_up = nullptr;
  std::vector<float> last;  // the transforms of the last pose (debug getter)
  float* d_seq = nullptr;   // [n_frames][stride]
  size_t seq_frames = 0, seq_cursor = 0;
};

namespace {

// util::Rodrigues, include/xrtailor/runtime/rag_doll/smpl/util.hpp:79-94 (row-major 3x3)
void rodrigues(const float* v, float* R) {
  const float theta = sqrtf(v[0] * v[0] + v[1] * v[1] + v[2] * v[2]);
  if (fabsf(theta) < 1e-5f) {
    for (int i = 0; i < 9; i++) R[i] = (i % 4 == 0) ? 1.f : 0.f;
    return;
  }
  const float c = cosf(theta), s = sinf(theta);
  const float r[3] = {v[0] / theta, v[1] / theta, v[2] / theta};
  const float skew[9] = {0.f, -r[2], r[1], r[2], 0.f, -r[0], -r[1], r[0], 0.f};
  for (int i = 0; i < 3; i++)
    for (int j = 0; j < 3; j++) R[3 * i + j] = (i == j ? c : 0.f) + ((1.f - c) * r[i]) * r[j] + s * skew[3 * i + j];
}

// Body<>::Update (body.cpp:308-325) + _local_to_global (:386-415): frame = [J x 12 transforms | 9 (J - 1) blend params].
// Host work like the reference's (J <= 55 joints; the reference calls the device version "horribly slow").
void pose_to_frame(const RigState& r, const float* pose, const float* trans, float* frame) {
  const uint32_t J = r.J;
  float* T = frame;
  for (uint32_t i = 0; i < J; i++) {
    float R[9];
    rodrigues(pose + 3 * i, R);
    float* t = T + 12 * i;
    for (int a = 0; a < 3; a++) {
      t[4 * a] = R[3 * a]; t[4 * a + 1] = R[3 * a + 1]; t[4 * a + 2] = R[3 * a + 2]; t[4 * a + 3] = 0.f;
    }
    if (r.P && i > 0)
      for (int k = 0; k < 9; k++) frame[12 * J + 9 * (i - 1) + k] = R[k] - (k % 4 == 0 ? 1.f : 0.f);
  }
  const float* jt = r.joints.data();
  for (int a = 0; a < 3; a++) T[4 * a + 3] = jt[a] + (trans ? trans[a] : 0.f);
  for (uint32_t i = 1; i < J; i++) {
    float* b = T + 12 * i;
    const uint32_t p = (uint32_t)r.parent[i];
    const float* a = T + 12 * p;
    for (int k = 0; k < 3; k++) b[4 * k + 3] = jt[3 * i + k] - jt[3 * p + k];
    float o[12];
    for (int k = 0; k < 3; k++) {  // util::MulAffine, util.hpp:107-112
      for (int c = 0; c < 3; c++) o[4 * k + c] = a[4 * k] * b[c] + a[4 * k + 1] * b[4 + c] + a[4 * k + 2] * b[8 + c];
      o[4 * k + 3] = a[4 * k + 3] + (a[4 * k] * b[3] + a[4 * k + 1] * b[7] + a[4 * k + 2] * b[11]);
    }
    memcpy(b, o, sizeof o);
  }
  for (uint32_t i = 0; i < J; i++) {
    float* t = T + 12 * i;
    const float* j = jt + 3 * i;
    for (int k = 0; k < 3; k++) t[4 * k + 3] -= j[0] * t[4 * k] + j[1] * t[4 * k + 1] + j[2] * t[4 * k + 2];
  }
}

struct RigMat {
  float m[12];  // rows 0..2 of the model matrix (row-major 3x4)
  int on;
};

// mmv_block (pose blend shapes, body.cu:243-249) + device::lbs (body.cu:39-63, the kernel's accumulation order) +
// UpdateObstaclePredicted_Kernel (ClothSolverHelper.cu:70-79) for one vertex; the joint transforms of the frame are
// staged in shared memory (J x 48 bytes).  Writes Node::x of the obstacle, w (inverse mass) untouched.
__global__ void rig_lbs_kernel(const float* __restrict__ rest, const uint32_t* __restrict__ w_outer,
                               const uint32_t* __restrict__ w_inner, const float* __restrict__ w_val,
                               const float* __restrict__ blend, const float* __restrict__ frame, uint32_t J, uint32_t P,
                               RigMat M, float4* __restrict__ x, uint32_t V) {
  extern __shared__ float sh[];  // [12 J + P]
  for (uint32_t i = threadIdx.x; i < 12 * J + P; i += blockDim.x) sh[i] = frame[i];
  __syncthreads();
  const uint32_t v = blockIdx.x * blockDim.x + threadIdx.x;
  if (v >= V) return;
  float p[3] = {rest[3 * v], rest[3 * v + 1], rest[3 * v + 2]};
  if (P) {
    const float* par = sh + 12 * J;
#pragma unroll
    for (int i = 0; i < 3; i++) {
      const float* row = blend + ((size_t)3 * v + i) * P;
      float acc = 0.f;
      for (uint32_t k = 0; k < P; k++) acc += row[k] * par[k];
      p[i] += acc;
    }
  }
  float q[3];
  const uint32_t lo = w_outer[v], hi = w_outer[v + 1];
#pragma unroll
  for (int i = 0; i < 3; i++) {
    float acc = 0.f;
    for (uint32_t it = lo; it < hi; it++) {
      const float w = w_val[it];
      const float* row = sh + 12 * w_inner[it] + 4 * i;
      acc += w * row[0] * p[0];
      acc += w * row[1] * p[1];
      acc += w * row[2] * p[2];
      acc += w * row[3];
    }
    q[i] = acc;
  }
  float4 o = x[v];
  if (M.on) {
    o.x = M.m[0] * q[0] + M.m[1] * q[1] + M.m[2] * q[2] + M.m[3];
    o.y = M.m[4] * q[0] + M.m[5] * q[1] + M.m[6] * q[2] + M.m[7];
    o.z = M.m[8] * q[0] + M.m[9] * q[1] + M.m[10] * q[2] + M.m[11];
  } else {
    o.x = q[0]; o.y = q[1]; o.z = q[2];
  }
  x[v] = o;
}

cudaError_t launch_rig(const xrt_solver* s, const float* d_frame, cudaStream_t st) {
  const RigState& r = *s->rig;
  RigMat M{};
  M.on = r.has_M ? 1 : 0;
  if (r.has_M) memcpy(M.m, r.M, sizeof M.m);
  const uint32_t V = s->o.V;
  rig_lbs_kernel<<<xrt_grid(V), XRT_BLOCK, (12 * r.J + r.P) * sizeof(float), st>>>(
      r.d_rest, r.d_w_outer, r.d_w_inner, r.d_w_val, r.d_blend, d_frame, r.J, r.P, M, s->o.x, V);
  return cudaGetLastError();
}

}  // namespace

void xrt_rig_destroy(xrt_solver* s) {
  RigState* r = s->rig;
  if (!r) return;
  cudaFree(r->d_rest); cudaFree(r->d_w_outer); cudaFree(r->d_w_inner); cudaFree(r->d_w_val); cudaFree(r->d_blend);
  cudaFree(r->d_frame); cudaFree(r->d_seq);
  if (r->h_frame) cudaFreeHost(r->h_frame);
  if (r->ev_up) cudaEventDestroy(r->ev_up);
  delete r;
  s->rig = nullptr;
}

// next frame of a resident pose sequence -> obstacle target (one kernel, nothing crosses PCIe)
int xrt_rig_take_sequence_frame(xrt_solver* s, cudaStream_t st) {
  RigState* r = s->rig;
  if (!r || !r->seq_frames || s->target_pending) return XRT_OK;
  XRT_CUDA_OK(launch_rig(s, r->d_seq + (r->seq_cursor % r->seq_frames) * r->stride, st));
  r->seq_cursor++;
  s->target_pending = true;
  return XRT_OK;
}

extern "C" int xrt_set_obstacle_rig(xrt_solver* s, const xrt_rig_desc* d) {
  XRT_READY();
  XRT_REQUIRE(d && d->rest_positions && d->weight_outer && d->weight_inner && d->weight_values && d->joints && d->parent,
              "set_obstacle_rig: rest positions, CSR weights, joints and parents are required");
  XRT_REQUIRE(s->o.V && d->n_verts == s->o.V, "set_obstacle_rig: vertex count differs from the obstacle");
  XRT_REQUIRE(d->n_joints >= 1 && d->n_joints <= 1024, "set_obstacle_rig: 1..1024 joints");
  const uint32_t V = s->o.V, J = d->n_joints;
  const size_t nnz = d->weight_outer[V];
  XRT_REQUIRE(d->weight_outer[0] == 0, "set_obstacle_rig: weight_outer[0] must be 0");
  for (uint32_t v = 0; v < V; v++)
    XRT_REQUIRE(d->weight_outer[v] <= d->weight_outer[v + 1], "set_obstacle_rig: weight_outer must not decrease");
  for (size_t i = 0; i < nnz; i++) XRT_REQUIRE(d->weight_inner[i] < J, "set_obstacle_rig: joint index out of range");
  for (uint32_t j = 1; j < J; j++)
    XRT_REQUIRE(d->parent[j] >= 0 && (uint32_t)d->parent[j] < j, "set_obstacle_rig: parent[j] must be a joint before j");
  XRT_CUDA_OK(cudaStreamSynchronize(s->st));
  xrt_rig_destroy(s);
  RigState* r = new RigState();
  s->rig = r;
  r->J = J;
  r->P = d->pose_blend ? 9 * (J - 1) : 0;
  r->stride = 12 * (size_t)J + r->P;
  r->joints.assign(d->joints, d->joints + 3 * (size_t)J);
  r->parent.assign(d->parent, d->parent + J);
  r->has_M = d->model_matrix != nullptr;
  if (r->has_M) memcpy(r->M, d->model_matrix, sizeof r->M);
  auto up = [&](auto** dst, const auto* src, size_t n) {
    cudaError_t e = cudaMalloc((void**)dst, (n ? n : 1) * sizeof(**dst));
    if (e == cudaSuccess && n) e = cudaMemcpy(*dst, src, n * sizeof(**dst), cudaMemcpyHostToDevice);
    return e;
  };
  cudaError_t e = up(&r->d_rest, d->rest_positions, 3 * (size_t)V);
  if (e == cudaSuccess) e = up(&r->d_w_outer, d->weight_outer, (size_t)V + 1);
  if (e == cudaSuccess) e = up(&r->d_w_inner, d->weight_inner, nnz);
  if (e == cudaSuccess) e = up(&r->d_w_val, d->weight_values, nnz);
  if (e == cudaSuccess && r->P) e = up(&r->d_blend, d->pose_blend, 3 * (size_t)V * r->P);
  if (e == cudaSuccess) e = cudaMalloc((void**)&r->d_frame, r->stride * sizeof(float));
  if (e == cudaSuccess) e = cudaMallocHost((void**)&r->h_frame, r->stride * sizeof(float));
  if (e == cudaSuccess) e = cudaEventCreateWithFlags(&r->ev_up, cudaEventDisableTiming);
  if (e != cudaSuccess) {
    xrt_set_error(s, "set_obstacle_rig: %s", cudaGetErrorString(e));
    xrt_rig_destroy(s);
    return XRT_ERR_CUDA;
  }
  r->last.assign(12 * (size_t)J, 0.f);
  return XRT_OK;
}

// sequence_->set_pose + body_->Update + UpdateObstaclePredicted (impl/smpl/ClothSolver.cpp:244-249) for one frame
extern "C" int xrt_set_obstacle_pose(xrt_solver* s, const float* axis_angle, const float* trans) {
  XRT_READY();
  RigState* r = s->rig;
  XRT_REQUIRE(r, "set_obstacle_pose: call xrt_set_obstacle_rig first");
  XRT_REQUIRE(axis_angle, "set_obstacle_pose: axis_angle is required");
  XRT_CUDA_OK(cudaEventSynchronize(r->ev_up));  // the previous upload has left the pinned buffer
  pose_to_frame(*r, axis_angle, trans, r->h_frame);
  memcpy(r->last.data(), r->h_frame, 12 * (size_t)r->J * sizeof(float));
  XRT_CUDA_OK(cudaMemcpyAsync(r->d_frame, r->h_frame, r->stride * sizeof(float), cudaMemcpyHostToDevice, s->st));
  XRT_CUDA_OK(cudaEventRecord(r->ev_up, s->st));
  XRT_CUDA_OK(launch_rig(s, r->d_frame, s->st));
  s->target_pending = true;
  return XRT_OK;
}

extern "C" int xrt_set_obstacle_pose_sequence(xrt_solver* s, const float* axis_angle, const float* trans, size_t n_frames) {
  XRT_READY();
  RigState* r = s->rig;
  XRT_REQUIRE(r, "set_obstacle_pose_sequence: call xrt_set_obstacle_rig first");
  XRT_REQUIRE(axis_angle && n_frames, "set_obstacle_pose_sequence: empty sequence");
  std::vector<float> h(n_frames * r->stride);
  for (size_t f = 0; f < n_frames; f++)
    pose_to_frame(*r, axis_angle + f * 3 * r->J, trans ? trans + 3 * f : nullptr, h.data() + f * r->stride);
  XRT_CUDA_OK(cudaStreamSynchronize(s->st));
  cudaFree(r->d_seq);
  r->d_seq = nullptr;
  r->seq_frames = 0;
  XRT_CUDA_OK(cudaMalloc((void**)&r->d_seq, h.size() * sizeof(float)));
  XRT_CUDA_OK(cudaMemcpy(r->d_seq, h.data(), h.size() * sizeof(float), cudaMemcpyHostToDevice));
  r->seq_frames = n_frames;
  r->seq_cursor = 0;
  return XRT_OK;
}

extern "C" int xrt_get_obstacle_joint_transforms(xrt_solver* s, float* out) {
  if (!s || !out) return XRT_ERR_INVALID;
  XRT_REQUIRE(s->rig, "get_obstacle_joint_transforms: no rig");
  memcpy(out, s->rig->last.data(), s->rig->last.size() * sizeof(float));
  return XRT_OK;
}

extern "C" int xrt_get_obstacle_target(xrt_solver* s, float* xyz) {
  XRT_READY();
  XRT_REQUIRE(xyz && s->o.V, "get_obstacle_target: no obstacle or null output");
  std::vector<float4> h(s->o.V);
  XRT_CUDA_OK(cudaStreamSynchronize(s->st));
  XRT_CUDA_OK(cudaMemcpy(h.data(), s->o.x, h.size() * sizeof(float4), cudaMemcpyDeviceToHost));
  for (size_t i = 0; i < h.size(); i++) {
    xyz[3 * i] = h[i].x; xyz[3 * i + 1] = h[i].y; xyz[3 * i + 2] = h[i].z;
  }
  return XRT_OK;
}
