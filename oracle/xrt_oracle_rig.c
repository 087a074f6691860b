/*
 * xrt_oracle_rig.c — CPU restatement of the reference's body-animation producer
 * (SURVEY.md section 8(f) rank 3): angle-axis pose -> joint transforms -> linear blend
 * skinning -> obstacle target positions.
 *
 * TEST INFRASTRUCTURE ONLY (xrt_oracle.c #includes this file at its end).  Plain C, one
 * thread, IEEE arithmetic, no FMA contraction.  Paths relative to /root/reference/xrtailor/.
 *
 * Parity status: UNPINNED against the reference's own code.  The reference's kernels for this
 * step live in an anonymous namespace of src/runtime/rag_doll/smpl/cuda/body.cu and are only
 * reachable through smplx::Body<>::Update, which needs an SMPL model file (a licensed download
 * that is not in the archive) and Eigen; there are no tests or golden vectors for it.  The
 * restatement is checked against a dense numpy formulation (4x4 matrices, full weight matrix)
 * in tests/test_oracle_cpu.py instead.
 */

/* util::Rodrigues, include/xrtailor/runtime/rag_doll/smpl/util.hpp:79-94.
 * R row-major 3x3 = c*I + (1-c)*r r^T + s*skew(r); identity below 1e-5 rad. */
void oracle_rodrigues(const float* v, float* R) {
  const float theta = sqrtf(v[0] * v[0] + v[1] * v[1] + v[2] * v[2]);
  if (fabsf(theta) < 1e-5f) {
    for (int i = 0; i < 9; i++) R[i] = (i % 4 == 0) ? 1.f : 0.f;
    return;
  }
  const float c = cosf(theta), s = sinf(theta);
  const float r[3] = {v[0] / theta, v[1] / theta, v[2] / theta};
  const float skew[9] = {0.f, -r[2], r[1], r[2], 0.f, -r[0], -r[1], r[0], 0.f};
  for (int i = 0; i < 3; i++)
    for (int j = 0; j < 3; j++)
      R[3 * i + j] = (i == j ? c : 0.f) + ((1.f - c) * r[i]) * r[j] + s * skew[3 * i + j];
}

/* Body<>::Update (src/runtime/rag_doll/smpl/body.cpp:308-325) + Body<>::_local_to_global
 * (:386-415).  pose: 3*J angle-axis; trans: 3 (or NULL); joints: 3*J shaped joint positions;
 * parent[0] is ignored.  T: J x 12, row-major 3x4 [R | t] per joint, such that a rest-space
 * point p of the shaped mesh goes to R p + t. */
void oracle_local_to_global(const float* pose, const float* trans, const float* joints, const int32_t* parent,
                            uint32_t J, float* T) {
  for (uint32_t i = 0; i < J; i++) {
    float R[9];
    oracle_rodrigues(pose + 3 * i, R);
    float* t = T + 12 * i;
    for (int r = 0; r < 3; r++) {
      t[4 * r] = R[3 * r];
      t[4 * r + 1] = R[3 * r + 1];
      t[4 * r + 2] = R[3 * r + 2];
      t[4 * r + 3] = 0.f;
    }
  }
  /* root: translation = joint 0 + trans (:391-393) */
  for (int r = 0; r < 3; r++) T[4 * r + 3] = joints[r] + (trans ? trans[r] : 0.f);
  /* children in index order: relative translation, then compose with the parent (util::MulAffine, util.hpp:107-112) */
  for (uint32_t i = 1; i < J; i++) {
    float* b = T + 12 * i;
    const float* a = T + 12 * (uint32_t)parent[i];
    for (int r = 0; r < 3; r++) b[4 * r + 3] = joints[3 * i + r] - joints[3 * (uint32_t)parent[i] + r];
    float o[12];
    for (int r = 0; r < 3; r++) {
      for (int c = 0; c < 3; c++) o[4 * r + c] = a[4 * r] * b[c] + a[4 * r + 1] * b[4 + c] + a[4 * r + 2] * b[8 + c];
      o[4 * r + 3] = a[4 * r + 3] + (a[4 * r] * b[3] + a[4 * r + 1] * b[7] + a[4 * r + 2] * b[11]);
    }
    memcpy(b, o, sizeof o);
  }
  /* centre every transform at the global origin: t -= R * joint (:410-414) */
  for (uint32_t i = 0; i < J; i++) {
    float* t = T + 12 * i;
    const float* j = joints + 3 * i;
    for (int r = 0; r < 3; r++) t[4 * r + 3] -= j[0] * t[4 * r] + j[1] * t[4 * r + 1] + j[2] * t[4 * r + 2];
  }
}

/* Pose blend parameters: R_i - I of joints 1..J-1, row-major, 9 per joint (body.cpp:318-324). */
void oracle_pose_blend_params(const float* pose, uint32_t J, float* out) {
  for (uint32_t i = 1; i < J; i++) {
    float R[9];
    oracle_rodrigues(pose + 3 * i, R);
    for (int k = 0; k < 9; k++) out[9 * (i - 1) + k] = R[k] - (k % 4 == 0 ? 1.f : 0.f);
  }
}

/* device::lbs (src/runtime/rag_doll/smpl/cuda/body.cu:39-63) in the kernel's own accumulation order, optionally
 * preceded by the pose blend shapes (mmv_block, body.cu:243-249: verts_shaped += B * params, B row-major
 * [3V][P]), followed by UpdateObstaclePredicted_Kernel (src/physics/ClothSolverHelper.cu:70-79):
 * x = M * vec4(v, 1) with M = model_matrix * align_matrix given row-major (NULL = identity).
 * rest: 3*V; CSR weights (outer V+1, inner, values); T: J x 12; out: 3*V. */
void oracle_lbs(const float* rest, uint32_t V, const uint32_t* w_outer, const uint32_t* w_inner, const float* w_val,
                const float* T, const float* M, const float* pose_blend, const float* blend_params, uint32_t P,
                float* out) {
  for (uint32_t v = 0; v < V; v++) {
    float p[3] = {rest[3 * v], rest[3 * v + 1], rest[3 * v + 2]};
    if (pose_blend)
      for (int i = 0; i < 3; i++) {
        const float* row = pose_blend + ((size_t)3 * v + i) * P;
        float acc = 0.f;
        for (uint32_t k = 0; k < P; k++) acc += row[k] * blend_params[k];
        p[i] += acc;
      }
    float q[3];
    for (int i = 0; i < 3; i++) {
      float acc = 0.f;
      for (uint32_t it = w_outer[v]; it < w_outer[v + 1]; it++) {
        const float* row = T + 12 * (size_t)w_inner[it] + 4 * i;
        for (int j = 0; j < 3; j++) acc += w_val[it] * row[j] * p[j];
        acc += w_val[it] * row[3];
      }
      q[i] = acc;
    }
    if (M) {
      for (int i = 0; i < 3; i++) out[3 * v + i] = M[4 * i] * q[0] + M[4 * i + 1] * q[1] + M[4 * i + 2] * q[2] + M[4 * i + 3];
    } else {
      out[3 * v] = q[0]; out[3 * v + 1] = q[1]; out[3 * v + 2] = q[2];
    }
  }
}
