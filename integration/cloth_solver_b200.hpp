// ClothSolverB200 -- the host-side binding of libxrt_b200.so in C++, member for member after the reference's
// ClothSolverBase / ClothSolverUniversal (xrtailor/include/xrtailor/pipeline/base/ClothSolverBase.hpp:55-189,
// xrtailor/src/pipeline/impl/universal/ClothSolver.cpp:189-350).  It depends on include/xrt_b200.h only: a
// maintainer of the reference derives it from ClothSolverBase and feeds AddCloth / AddObstacle from `Mesh`
// (INTEGRATION.md shows that adapter); tests/test_integration_cpp.py compiles this header as it stands and runs
// integration/run_one_frame.cpp through it on the GPU box.
#pragma once
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <stdexcept>
#include <string>
#include <vector>

extern "C" {
#include "xrt_b200.h"
}

namespace xrt {

class ClothSolverB200 {
 public:
  explicit ClothSolverB200(const xrt_params& p) : params_(p) { Check(xrt_create(&params_, &h_)); }
  ~ClothSolverB200() { OnDestroy(); }
  ClothSolverB200(const ClothSolverB200&) = delete;
  ClothSolverB200& operator=(const ClothSolverB200&) = delete;

  // ClothSolverUniversal::AddCloth (ClothSolver.cpp:247-350): world-space positions, triangle indices, pinned vertices.
  // Edges, stretch constraints and bend elements come from xrt_build_constraints (what PrepareBasicStretchData /
  // GenerateBendingElements derive from vcglib's half-edge mesh).  Returns the garment's first vertex in the merged mesh.
  int AddCloth(const std::vector<float>& positions, const std::vector<uint32_t>& indices,
               const std::vector<uint32_t>& fixed_indices = {}) {
    const uint32_t nv = (uint32_t)(positions.size() / 3), nf = (uint32_t)(indices.size() / 3);
    xrt_constraints c{};
    Check(xrt_build_constraints(positions.data(), nv, indices.data(), nf, &c));
    std::vector<float> inv_mass(nv, 1.0f);
    for (uint32_t i : fixed_indices) inv_mass.at(i) = 0.0f;  // SetInvMass(0), ClothSolver.cpp:296-299
    xrt_mesh_desc d{positions.data(), indices.data(), c.edges, inv_mass.data(), nv, nf, c.n_edges};
    int32_t offset = 0;
    int rc = xrt_add_cloth(h_, &d, &offset);
    if (rc == XRT_OK) {
      std::vector<uint32_t> idx(c.edges, c.edges + 2 * (size_t)c.n_edges), bend(c.bend_idx, c.bend_idx + 4 * (size_t)c.n_bend);
      for (auto& v : idx) v += (uint32_t)offset;  // merged numbering, like `idx + prevNumParticles` in the reference
      for (auto& v : bend) v += (uint32_t)offset;
      rc = xrt_set_stretch(h_, idx.data(), c.stretch_rest, c.n_edges);
      if (rc == XRT_OK) rc = xrt_set_bend(h_, bend.data(), c.bend_theta, c.n_bend);
    }
    xrt_free_constraints(&c);
    Check(rc);
    n_cloth_ += nv;
    return offset;
  }

  // ClothSolverUniversal::AddObstacle (ClothSolver.cpp:189-245)
  int AddObstacle(const std::vector<float>& positions, const std::vector<uint32_t>& indices) {
    const uint32_t nv = (uint32_t)(positions.size() / 3), nf = (uint32_t)(indices.size() / 3);
    xrt_constraints c{};
    Check(xrt_build_constraints(positions.data(), nv, indices.data(), nf, &c));
    xrt_mesh_desc d{positions.data(), indices.data(), c.edges, nullptr, nv, nf, c.n_edges};
    int32_t offset = 0;
    const int rc = xrt_add_obstacle(h_, &d, &offset);
    xrt_free_constraints(&c);
    Check(rc);
    n_obstacle_ += nv;
    return offset;
  }

  // ClothSolverBase::AddGeodesic (ClothSolverBase.cpp:527-532)
  void AddGeodesic(uint32_t src, uint32_t tgt, float rest) { Check(xrt_add_geodesic(h_, &src, &tgt, &rest, 1)); }
  // BindingConstraint::Add + Generate with the SkinParam anchor (BindingConstraint.cu:11-55)
  void AddBinding(const std::vector<uint32_t>& idx, const std::vector<uint32_t>& tri, const std::vector<float>& bary,
                  const std::vector<float>& stiffness, const std::vector<float>& distance) {
    Check(xrt_add_binding(h_, idx.data(), tri.data(), bary.data(), stiffness.data(), distance.data(), idx.size()));
  }
  // SOLVER_MODE::QUALITY (Global::sim_config.quality): before Start
  void SetQuality(const xrt_quality_params& q) { Check(xrt_set_quality(h_, &q, nullptr)); quality_ = true; }

  void Start() { Check(xrt_finalize_setup(h_)); }  // ClothSolverUniversal::Start, ClothSolver.cpp:46-93

  // UpdateObstacleAnimationFrame: the posed body of the coming frame
  // ClothSolverSMPL: the skinned body as a resident rig; per frame the pose replaces sequence_->set_pose + body_->Update +
  // UpdateObstaclePredicted (impl/smpl/ClothSolver.cpp:244-249).  Call SetObstacleRig after Start().
  void SetObstacleRig(const xrt_rig_desc& rig) { Check(xrt_set_obstacle_rig(h_, &rig)); }
  void UpdateObstaclePose(const std::vector<float>& axis_angle, const float* trans = nullptr) {
    Check(xrt_set_obstacle_pose(h_, axis_angle.data(), trans));
  }
  // ObjSequenceExporter (runtime/export/ObjSequenceExporter.cpp): frames stream to <dir>/%06d.obj while the run goes on
  void OpenObjSequence(const char* dir, const std::vector<uint32_t>& indices, uint32_t n_verts, bool normals = true) {
    if (xrt_obj_writer_open(dir, indices.data(), (uint32_t)(indices.size() / 3), n_verts, normals, 0, 0, &writer_))
      throw std::runtime_error("xrt_b200: cannot open the OBJ sequence directory");
  }
  void ExportFrame() {  // after SaveCache(): hands the newest cached frame to the writer threads
    if (writer_ && xrt_obj_writer_push(writer_, h_cloth_positions_cache.back().data(), h_cloth_normals_cache.back().data()))
      throw std::runtime_error("xrt_b200: OBJ writer failed");
  }
  void CloseObjSequence() {
    if (writer_) xrt_obj_writer_close(writer_, nullptr, nullptr);
    writer_ = nullptr;
  }
  void UpdateObstacleAnimationFrame(const std::vector<float>& xyz) { Check(xrt_set_obstacle_target(h_, xyz.data(), xyz.size() / 3)); }

  // ClothSolverBase::SimulateSwift / SimulateQuality (ClothSolverBase.cpp:171-362); solver_time = Solver_Total in ms
  void Simulate() {
    float ms = 0.f;
    Check(xrt_step_frame(h_, &ms));
    solver_time = ms;
    SaveCache();
  }
  // SaveCache (ClothSolver.cpp:107-144): the exporters read these host frames
  void SaveCache() {
    h_cloth_positions_cache.emplace_back(3 * (size_t)n_cloth_);
    h_cloth_normals_cache.emplace_back(3 * (size_t)n_cloth_);
    Check(xrt_get_positions(h_, 0, h_cloth_positions_cache.back().data(), h_cloth_normals_cache.back().data()));
  }
  void OnDestroy() {
    CloseObjSequence();
    if (h_) xrt_destroy(h_);
    h_ = nullptr;
  }

  double solver_time = 0.0;
  std::vector<std::vector<float>> h_cloth_positions_cache, h_cloth_normals_cache;
  size_t NumClothVertices() const { return n_cloth_; }
  xrt_solver* handle() { return h_; }

 private:
  void Check(int rc) {  // the reference exits with a TAILOR_EXIT code; a library user gets an exception with the same code
    if (rc) throw std::runtime_error("xrt_b200 error " + std::to_string(rc) + ": " + (h_ ? xrt_last_error(h_) : "no handle"));
  }
  xrt_params params_;
  xrt_solver* h_ = nullptr;
  xrt_obj_writer* writer_ = nullptr;
  size_t n_cloth_ = 0, n_obstacle_ = 0;
  bool quality_ = false;
};

}  // namespace xrt
