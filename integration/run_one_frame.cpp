// Drives libxrt_b200.so from C++ through integration/cloth_solver_b200.hpp: two garments added one after the other
// (merged by the library with a running index offset), one obstacle, two Swift frames with a moving obstacle, then a
// Quality-mode solver on the same meshes.  Prints one line per solver; exit code 0 = every position finite and the
// cloth fell.  Built and run by tests/test_integration_cpp.py.
#include <cmath>
#include <cstdio>

#include "cloth_solver_b200.hpp"

static void grid(int n, float side, float y, std::vector<float>& pos, std::vector<uint32_t>& tri) {
  pos.clear();
  tri.clear();
  for (int i = 0; i < n; i++)
    for (int j = 0; j < n; j++) {
      pos.push_back(-side / 2 + side * i / (n - 1) + 1e-4f * (float)((i * 7 + j * 13) % 5 - 2));
      pos.push_back(y);
      pos.push_back(-side / 2 + side * j / (n - 1) + 1e-4f * (float)((i * 11 + j * 3) % 5 - 2));
    }
  for (int i = 0; i + 1 < n; i++)
    for (int j = 0; j + 1 < n; j++) {
      const uint32_t a = i * n + j, b = a + 1, c = a + n, d = c + 1;
      tri.insert(tri.end(), {a, b, d, a, d, c});
    }
}
static void sphere(float r, int stacks, int slices, std::vector<float>& pos, std::vector<uint32_t>& tri) {
  pos = {0, r, 0};
  for (int i = 1; i < stacks; i++)
    for (int j = 0; j < slices; j++) {
      const float th = 3.14159265f * i / stacks, ph = 6.2831853f * j / slices;
      pos.insert(pos.end(), {r * sinf(th) * cosf(ph), r * cosf(th), r * sinf(th) * sinf(ph)});
    }
  pos.insert(pos.end(), {0, -r, 0});
  const uint32_t south = (uint32_t)(pos.size() / 3 - 1);
  auto ring = [&](int i, int j) { return (uint32_t)(1 + (i - 1) * slices + (j % slices)); };
  for (int j = 0; j < slices; j++) tri.insert(tri.end(), {0u, ring(1, j + 1), ring(1, j)});
  for (int i = 1; i + 1 < stacks; i++)
    for (int j = 0; j < slices; j++) tri.insert(tri.end(), {ring(i, j), ring(i, j + 1), ring(i + 1, j + 1), ring(i, j), ring(i + 1, j + 1), ring(i + 1, j)});
  for (int j = 0; j < slices; j++) tri.insert(tri.end(), {south, ring(stacks - 1, j), ring(stacks - 1, j + 1)});
}

static int run(bool quality) {
  std::vector<float> g0, g1, ob;
  std::vector<uint32_t> t0, t1, to;
  grid(24, 0.5f, 0.34f, g0, t0);
  grid(16, 0.4f, 0.36f, g1, t1);
  sphere(0.3f, 12, 24, ob, to);
  xrt_params p;
  xrt_default_params(&p);
  p.persist_cluster = 0;
  if (quality) { p.num_substeps = 1; p.num_iterations = 20; p.enable_self_collision = 0; }
  xrt::ClothSolverB200 solver(p);
  const int off0 = solver.AddCloth(g0, t0), off1 = solver.AddCloth(g1, t1, {0u});
  solver.AddObstacle(ob, to);
  if (quality) {
    xrt_quality_params q;
    xrt_default_quality_params(&q);
    solver.SetQuality(q);
  }
  solver.Start();
  for (int f = 0; f < 2; f++) {
    std::vector<float> tgt = ob;
    for (size_t i = 1; i < tgt.size(); i += 3) tgt[i] += 0.002f * (f + 1);
    solver.UpdateObstacleAnimationFrame(tgt);
    solver.Simulate();
  }
  if (!quality) {
    // the obstacle as a one-joint rig posed on the device, and the frames streamed to an OBJ sequence
    const uint32_t nv = (uint32_t)(ob.size() / 3);
    std::vector<uint32_t> outer(nv + 1), inner(nv, 0u);
    std::vector<float> w(nv, 1.0f);
    for (uint32_t i = 0; i <= nv; i++) outer[i] = i;
    const float joint[3] = {0.f, 0.f, 0.f};
    const int32_t parent[1] = {-1};
    xrt_rig_desc rig{ob.data(), outer.data(), inner.data(), w.data(), joint, parent, nullptr, nullptr, nv, 1u};
    solver.SetObstacleRig(rig);
    std::vector<uint32_t> all = t0;
    for (uint32_t v : t1) all.push_back(v + (uint32_t)off1);
    solver.OpenObjSequence("integration/build/obj_sequence", all, (uint32_t)solver.NumClothVertices());
    for (int f = 0; f < 2; f++) {
      const float trans[3] = {0.f, 0.006f + 0.002f * f, 0.f};
      solver.UpdateObstaclePose({0.f, 0.02f * (f + 1), 0.f}, trans);
      solver.Simulate();
      solver.ExportFrame();
    }
    solver.CloseObjSequence();
  }
  const auto& x = solver.h_cloth_positions_cache.back();
  double sum = 0;
  bool finite = true;
  for (float v : x) { finite = finite && std::isfinite(v); sum += v; }
  const size_t pinned = (size_t)off1;  // vertex 0 of the second garment is pinned: it must not have moved
  const bool pin_ok = std::fabs(x[3 * pinned + 1] - 0.36f) < 1e-6f;
  const bool fell = x[3 * (size_t)(off0 + 5) + 1] < 0.34f;
  std::printf("%s: %zu cloth vertices (garments at %d and %d), checksum %.6f, last frame %.3f ms, finite %d pinned %d fell %d\n",
              quality ? "quality" : "swift", solver.NumClothVertices(), off0, off1, sum, solver.solver_time, (int)finite, (int)pin_ok, (int)fell);
  return finite && pin_ok && fell ? 0 : 1;
}

int main() {
  try {
    return run(false) | run(true);
  } catch (const std::exception& e) {
    std::fprintf(stderr, "%s\n", e.what());
    return 2;
  }
}
