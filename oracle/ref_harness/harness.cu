// GL-free harness around the reference's own CUDA physics translation units
// (openxrlab/xrtailor @ 4eaf5c3, compiled in place from /root/reference by
// oracle/ref_harness/Makefile).  TEST INFRASTRUCTURE ONLY: it is the oracle the
// CUDA product is checked against on the GPU box and the reference arm of
// bench.py.  This file contains no physics: it builds the reference's objects
// (MemoryPool, PhysicsMesh, BVH, StretchSolver, ...) and restates the call
// sequence of ClothSolverUniversal::Start
// (src/pipeline/impl/universal/ClothSolver.cpp:46-93) and
// ClothSolverBase::SimulateSwift (src/pipeline/base/ClothSolverBase.cpp:171-261)
// call for call, with ClothSolverSMPL::UpdateObstacleAnimationFrame's BVH
// rebuild (src/pipeline/impl/smpl/ClothSolver.cpp:257-261) for a moving obstacle.
#include <map>
#include <memory>
#include <set>
#include <vector>

#include <xrtailor/core/Global.hpp>
#include <xrtailor/memory/MemoryPool.cuh>
#include <xrtailor/physics/ClothSolverHelper.cuh>
#include <xrtailor/physics/PhysicsMesh.cuh>
#include <xrtailor/physics/broad_phase/lbvh/BVH.cuh>
#include <xrtailor/physics/broad_phase/spatial_hash/SpatialHashGPU.cuh>
#include <xrtailor/physics/dynamics/BasicSolver.cuh>
#include <xrtailor/physics/dynamics/BindingConstraint.cuh>
#include <xrtailor/physics/dynamics/FEMSolver.cuh>
#include <xrtailor/physics/dynamics/LongRangeAttachments.cuh>
#include <xrtailor/physics/pbd_collision/Constraint.cuh>
#include <xrtailor/physics/representative_triangle/RTriangle.cuh>
#include <xrtailor/physics/impact_zone/Entry.cuh>
#include <xrtailor/physics/impact_zone/ImpactHelper.cuh>
#include <xrtailor/physics/impact_zone/ImpactZoneHelper.cuh>
#include <xrtailor/physics/impact_zone/ParallelImpactZoneOptimizer.cuh>
// the harness reads the solvers' quadrature / pair lists for the parity tests (no accessor exists)
#define private public
#include <xrtailor/physics/repulsion/ImminentRepulsion.cuh>
#include <xrtailor/physics/repulsion/PBDRepulsion.cuh>
#undef private
#include <xrtailor/physics/repulsion/Impulse.cuh>
#include <xrtailor/utils/Timer.hpp>

namespace XRTailor {
// the reference's own Morton kernel (broad_phase/lbvh/BVH.cu:464-481), launched by ref_get_mortons
__global__ void ComputeMortons_Kernel(Primitive* objects, Bounds* aabbs, Bounds aabb_whole, uint* mortons,
                                      uint num_internal_nodes, uint num_objects);
// defined in repulsion/PBDRepulsion.cu:329,411, not declared in a header (ref_rep_contributions calls them)
__host__ __device__ bool EvaluateVFProximity(Node* node, Face* face, const Scalar& h, Impulse* impulse);
__host__ __device__ bool EvaluateEEProximity(Edge* edge0, Edge* edge1, const Scalar& h, Impulse* impulse);
__global__ void EvaluateRep_Kernel(PairFF* pairs, Vector3* deltas, int* delta_counts, Scalar thickness, int n_pairs);
namespace ImpactZoneOptimization {
// defined in impact_zone/Entry.cu:25, not declared in a header
thrust::device_vector<Impact> independentImpacts(const thrust::device_vector<Impact>& impacts, int deform);
}  // namespace ImpactZoneOptimization
}  // namespace XRTailor

using namespace XRTailor;

namespace {
__global__ void SetTarget_Kernel(Node** nodes, const float* xyz, int n) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  nodes[i]->x = Vector3(xyz[3 * i], xyz[3 * i + 1], xyz[3 * i + 2]);  // UpdateObstaclePredicted_Kernel with identity matrices
}
__global__ void SetInvMass_Kernel(Node** nodes, const float* w, int n) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  nodes[i]->inv_mass = w[i];
  if (w[i] == 0.0f) nodes[i]->is_free = false;
}
__global__ void GetState_Kernel(Node** nodes, float* x0, float* x, float* v, float* nrm, int n) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  for (int k = 0; k < 3; k++) {
    if (x0) x0[3 * i + k] = nodes[i]->x0[k];
    if (x) x[3 * i + k] = nodes[i]->x[k];
    if (v) v[3 * i + k] = nodes[i]->v[k];
    if (nrm) nrm[3 * i + k] = nodes[i]->n[k];
  }
}
__global__ void SetState_Kernel(Node** nodes, const float* x0, const float* x, const float* v, int n) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  for (int k = 0; k < 3; k++) {
    if (x0) nodes[i]->x0[k] = x0[3 * i + k];
    if (x) nodes[i]->x[k] = x[3 * i + k];
    if (v) nodes[i]->v[k] = v[3 * i + k];
  }
}

std::vector<uint> FaceEdgeIndices(const uint* tri, int F, const uint* edges, int E) {
  std::map<std::pair<uint, uint>, uint> m;
  for (int e = 0; e < E; e++) m[{std::min(edges[2 * e], edges[2 * e + 1]), std::max(edges[2 * e], edges[2 * e + 1])}] = e;
  std::vector<uint> fe(3 * (size_t)F);
  for (int f = 0; f < F; f++)
    for (int j = 0; j < 3; j++) {
      uint a = tri[3 * f + j], b = tri[3 * f + (j + 1) % 3];
      fe[3 * f + j] = m[{std::min(a, b), std::max(a, b)}];
    }
  return fe;
}

struct RefSolver {
  std::shared_ptr<MemoryPool> pool;
  std::shared_ptr<PhysicsMesh> cloth, obstacle;
  std::shared_ptr<XRTailor::BVH> obstacle_bvh;
  std::shared_ptr<BasicConstraint::StretchSolver> stretch;
  std::shared_ptr<BasicConstraint::BendSolver> bend;
  std::shared_ptr<BindingConstraint> binding;
  std::shared_ptr<PBDCollision::DCD::VFConstraint> vf_dcd;
  std::shared_ptr<PBDCollision::CCD::RTConstraint> vf_ccd;
  std::shared_ptr<PBDCollision::DCD::EEConstraint> ee_dcd;
  std::shared_ptr<RTriangle> obstacle_rtri;
  std::shared_ptr<SpatialHashGPU> hash;
  int interleaved_hash = 3;
  bool target_pending = false;
  // Quality mode (ClothSolverBase members, pipeline/base/ClothSolverBase.hpp)
  std::shared_ptr<XRTailor::BVH> cloth_bvh;
  std::shared_ptr<RTriangle> cloth_rtri;
  std::shared_ptr<ImminentRepulsion> imminent;
  std::shared_ptr<PBDRepulsionSolver> pbd;
  std::shared_ptr<FEM::StrainSolver> fem_strain;
  std::shared_ptr<FEM::IsometricBendingSolver> fem_bend;
  bool imminent_on = true, pbd_on = true;
  float imminent_thickness = 1e-3f, pbd_thickness = 1e-4f, relaxation_rate = 0.75f, obstacle_mass = 1e3f,
        iz_thickness = 1e-4f;
  int frame_index = 0;
  thrust::device_vector<Impact> last_impacts;
  SimParams sp;  // this solver's copy of the process-wide Global::sim_params, re-installed by use()
  float* d_tmp = nullptr;
  size_t tmp_floats = 0;
};

Timer* g_timer = nullptr;

// The reference keeps its parameters in process globals (Global::sim_params and the __constant__ copy uploaded
// by SetSimulationParams), so two solvers alive in one process overwrite each other's.  Every entry point
// re-installs the handle's own copy first.
RefSolver* use(void* h) {
  auto* s = (RefSolver*)h;
  Global::sim_params = s->sp;
  SetSimulationParams(&Global::sim_params);
  return s;
}
}  // namespace

// Timer::CurrentTime() needs it (utils/Timer.hpp:64); no GLFW on this box
extern "C" double glfwGetTime(void) {
  static auto t0 = std::chrono::steady_clock::now();
  return std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
}

extern "C" {

// ip: S, N, C, max_neighbors, interleaved_hash, self_collision, solve_bending, geodesic_lra, device, seed
// fp: gx, gy, gz, damping, relaxation, lra_stretchiness, friction, bvh_tolerance, stretch_compliance,
//     bend_compliance, particle_diameter, hash_cell_size
void* ref_create(const int* ip, const float* fp, const float* cpos, const float* cinvm, int V, const uint* ctri, int F,
                 const uint* cedge, int E, const uint* sidx, const float* srest, int ns, const uint* bidx,
                 const float* btheta, int nb, const float* opos, int Vo, const uint* otri, int Fo, const uint* oedge,
                 int Eo) {
  cudaSetDevice(ip[8]);
  if (!g_timer) g_timer = new Timer();
  SimParams& sp = Global::sim_params;
  sp = SimParams();
  sp.num_substeps = ip[0]; sp.num_iterations = ip[1]; sp.num_collision_passes = ip[2];
  sp.max_num_neighbors = ip[3];
  sp.enable_self_collision = ip[5] != 0; sp.solve_bending = ip[6] != 0; sp.geodesic_LRA = ip[7] != 0;
  sp.gravity = Vector3(fp[0], fp[1], fp[2]);
  sp.damping = fp[3]; sp.relaxation_factor = fp[4]; sp.long_range_stretchiness = fp[5];
  sp.sdf_friction = fp[6]; sp.bvh_tolerance = fp[7]; sp.particle_diameter = fp[10];
  sp.solver_mode = SOLVER_MODE::SWIFT;
  sp.num_particles = V; sp.num_edges = E; sp.num_faces = F; sp.num_overall_particles = V + Vo;
  sp.num_obstacle_vertices = Vo; sp.num_obstacle_edges = Eo; sp.num_obstacle_faces = Fo;
  sp.delta_time = Timer::FixedDeltaTime();

  auto* s = new RefSolver();
  s->interleaved_hash = ip[4] > 0 ? ip[4] : 1;
  s->pool = std::make_shared<MemoryPool>();
  s->cloth = std::make_shared<PhysicsMesh>(s->pool);
  s->obstacle = std::make_shared<PhysicsMesh>(s->pool);
  s->obstacle_bvh = std::make_shared<XRTailor::BVH>();
  s->stretch = std::make_shared<BasicConstraint::StretchSolver>();
  s->bend = std::make_shared<BasicConstraint::BendSolver>();
  s->binding = std::make_shared<BindingConstraint>();

  // --- AddCloth (impl/universal/ClothSolver.cpp:247-350) ---
  std::vector<Vector3> cp(V);
  for (int i = 0; i < V; i++) cp[i] = Vector3(cpos[3 * i], cpos[3 * i + 1], cpos[3 * i + 2]);
  std::vector<uint> cidx(ctri, ctri + 3 * (size_t)F);
  thrust::host_vector<uint> h_edges(cedge, cedge + 2 * (size_t)E);
  std::vector<uint> fe = FaceEdgeIndices(ctri, F, cedge, E);
  thrust::host_vector<uint> h_fe(fe.begin(), fe.end());
  s->pool->SetupInitialPositions(cp);
  s->cloth->RegisterNewMesh(cp, cidx, h_edges, h_fe, 0);
  s->pool->SetupTmpPositions(V);
  s->pool->SetupDeltas(V);
  s->tmp_floats = 3 * (size_t)std::max(V, Vo) + 16;
  cudaMalloc(&s->d_tmp, 4 * s->tmp_floats * sizeof(float));
  cudaMemcpy(s->d_tmp, cinvm, V * sizeof(float), cudaMemcpyHostToDevice);
  SetInvMass_Kernel<<<(V + 255) / 256, 256>>>(pointer(s->cloth->nodes), s->d_tmp, V);  // SetInvMass/SetFree per fixed index
  cudaDeviceSynchronize();
  std::vector<int> ev0(ns), ev1(ns);
  std::vector<Scalar> len(srest, srest + ns);
  for (int i = 0; i < ns; i++) { ev0[i] = sidx[2 * i]; ev1[i] = sidx[2 * i + 1]; }
  s->stretch->AddCloth(ev0, ev1, len, fp[8]);
  if (sp.solve_bending) {
    std::vector<uint> bi(bidx, bidx + 4 * (size_t)nb);
    std::vector<Scalar> ba(btheta, btheta + nb);
    s->bend->AddCloth(bi, ba, fp[9]);
  }
  // --- AddObstacle (impl/universal/ClothSolver.cpp:189-245) ---
  if (Vo) {
    std::vector<Vector3> op(Vo);
    for (int i = 0; i < Vo; i++) op[i] = Vector3(opos[3 * i], opos[3 * i + 1], opos[3 * i + 2]);
    std::vector<uint> oidx(otri, otri + 3 * (size_t)Fo);
    thrust::host_vector<uint> ho_edges(oedge, oedge + 2 * (size_t)Eo);
    std::vector<uint> ofe = FaceEdgeIndices(otri, Fo, oedge, Eo);
    thrust::host_vector<uint> ho_fe(ofe.begin(), ofe.end());
    s->obstacle->SetPtrOffset(s->cloth->NumNodes(), s->cloth->NumEdges(), s->cloth->NumFaces());
    s->obstacle->RegisterNewMesh(op, oidx, ho_edges, ho_fe, 0, false);
    s->obstacle->PredictedToPositions();
    s->obstacle->UpdateNodeGeometries();
  }
  // --- ClothSolverUniversal::Start (impl/universal/ClothSolver.cpp:46-93) ---
  if (Vo) {
    s->obstacle->UpdateNormals();
    s->obstacle_bvh->UpdateData(pointer(s->obstacle->faces), 0.0f, false, s->obstacle->NumFaces());
    s->obstacle_bvh->Construct();
    // SetupRTriangle (pipeline/base/ClothSolverBase.cpp:487-500)
    BuildVFAdjacency(s->obstacle, s->pool->obstacle_nb_vf, s->pool->obstacle_nb_vf_prefix);
    s->obstacle_rtri = std::make_shared<RTriangle>();
    s->obstacle_rtri->Init(s->obstacle->DeviceIndices(), s->pool->obstacle_nb_vf, s->pool->obstacle_nb_vf_prefix);
  } else {
    s->obstacle_rtri = std::make_shared<RTriangle>();
  }
  s->vf_dcd = std::make_shared<PBDCollision::DCD::VFConstraint>(V);
  s->vf_ccd = std::make_shared<PBDCollision::CCD::RTConstraint>(V);
  s->ee_dcd = std::make_shared<PBDCollision::DCD::EEConstraint>(E, s->obstacle_rtri);
  s->hash = std::make_shared<SpatialHashGPU>(sp.particle_diameter, sp.num_particles, fp[11], sp.max_num_neighbors);
  s->hash->SetInitialPositions(pointer(s->cloth->nodes), s->cloth->NumNodes());
  // SetupBinding (pipeline/base/ClothSolverBase.cpp:448-451, called by the SMPL pipeline's Start): sets
  // n_bindings = 0; without it BindingConstraint::n_bindings is uninitialised (BindingConstraint.cu:4)
  s->binding->Generate(pointer(s->cloth->nodes), pointer(s->obstacle->nodes), pointer(s->pool->skin_params));
  // SetupInternalDynamics (pipeline/base/ClothSolverBase.cpp:453-468)
  unsigned int seed = (unsigned int)ip[9];
  s->stretch->SetupGraphColoring(pointer(s->cloth->edges), s->cloth->NumEdges(), seed);
  if (sp.solve_bending) s->bend->SetupGraphColoring(seed);
  cudaDeviceSynchronize();
  s->sp = sp;
  return s;
}

// BindingConstraint::Add + Generate with the SkinParam table filled from the caller's
// anchors (what ClothSolverSMPL::Start does with the garment's skin weights and bind_* lists)
void ref_set_bindings(void* h, const uint* idx, const uint* tri3, const float* bary3, const float* stiff,
                      const float* dist, int n) {
  auto* s = use(h);
  const int V = s->cloth->NumNodes();
  thrust::host_vector<SkinParam> sk(V);
  for (int i = 0; i < V; i++) sk[i] = SkinParam{0, 0, 0, 0.f, 0.f, 0.f};
  std::vector<uint> bi(n);
  std::vector<Scalar> bs(n), bd(n);
  for (int i = 0; i < n; i++) {
    sk[idx[i]] = SkinParam{(int)tri3[3 * i], (int)tri3[3 * i + 1], (int)tri3[3 * i + 2], bary3[3 * i],
                           bary3[3 * i + 1], bary3[3 * i + 2]};
    bi[i] = idx[i]; bs[i] = stiff[i]; bd[i] = dist[i];
  }
  s->pool->skin_params = sk;
  s->binding = std::make_shared<BindingConstraint>();
  s->binding->Add(bi, bs, bd, 0);
  s->binding->Generate(pointer(s->cloth->nodes), pointer(s->obstacle->nodes), pointer(s->pool->skin_params));
  cudaDeviceSynchronize();
}

void ref_set_obstacle_target(void* h, const float* xyz) {
  auto* s = use(h);
  int n = s->obstacle->NumNodes();
  cudaMemcpy(s->d_tmp, xyz, 3 * (size_t)n * sizeof(float), cudaMemcpyHostToDevice);
  SetTarget_Kernel<<<(n + 255) / 256, 256>>>(pointer(s->obstacle->nodes), s->d_tmp, n);
  s->target_pending = true;
}

// ClothSolverBase::SimulateSwift, pipeline/base/ClothSolverBase.cpp:171-261; returns Solver_Total in ms
float ref_step_frame(void* h) {
  auto* s = use(h);
  SimParams& sp = Global::sim_params;
  Timer::StartTimerGPU("Solver_Total");
  Scalar substep_time = Timer::FixedDeltaTime() / sp.num_substeps;
  SetSimulationParams(&sp);
  bool b_obstacle_updated = false;
  if (s->target_pending) {  // UpdateObstacleAnimationFrame, impl/smpl/ClothSolver.cpp:257-261
    ScopedTimerGPU timer("Solver_UpdateBVH");
    s->obstacle_bvh->UpdateData(pointer(s->obstacle->faces), 0.0f, false, s->obstacle->NumFaces());
    s->obstacle_bvh->Construct();
    b_obstacle_updated = true;
    s->target_pending = false;
  }
  uint current_step = 0u;
  uint frequency = (sp.num_substeps * sp.num_iterations) / sp.num_collision_passes;
  uint cnt = 0;
  for (int substep = 0; substep < sp.num_substeps; substep++) {
    PredictPositions(pointer(s->cloth->nodes), substep_time, s->cloth->NumNodes());
    if (sp.enable_self_collision) {
      if (substep % s->interleaved_hash == 0) s->hash->Hash(pointer(s->cloth->nodes), s->cloth->NumNodes());
    }
    for (int iteration = 0; iteration < sp.num_iterations; iteration++) {
      if (sp.enable_self_collision) {
        CollideParticles(pointer(s->pool->deltas), pointer(s->pool->delta_counts), pointer(s->cloth->nodes),
                         pointer(s->hash->neighbors));
      }
      current_step = substep * sp.num_iterations + iteration;
      if (current_step % frequency == 0) {
        if (b_obstacle_updated) {
          s->obstacle->Interpolate(cnt, sp.num_collision_passes);
          s->obstacle->UpdateMidstepNormals();
          s->obstacle_bvh->RefitMidstep(pointer(s->obstacle->faces), sp.bvh_tolerance);
        }
        s->vf_dcd->Generate(pointer(s->cloth->nodes), s->obstacle_bvh);
        s->vf_ccd->Generate(pointer(s->cloth->nodes), s->obstacle_bvh);
        s->ee_dcd->Generate(pointer(s->cloth->nodes), pointer(s->cloth->edges), s->obstacle_bvh);
        cnt++;
      }
      for (int c = 0; c < s->stretch->Coloring()->n_colors; c++)
        s->stretch->Solve(pointer(s->cloth->nodes), pointer(s->pool->tmp_positions), c, substep_time, iteration);
      if (sp.solve_bending)
        for (int c = 0; c < s->bend->Coloring()->n_colors; c++)
          s->bend->Solve(pointer(s->cloth->nodes), c, substep_time, iteration);
      {  // SolveLongRangeAttachments, pipeline/base/ClothSolverBase.cpp:558-566
        ScopedTimerGPU timer("Solver_SolveLRA");
        if (sp.geodesic_LRA)
          SolveGeodesicLRA(pointer(s->cloth->nodes), pointer(s->pool->geodesic_src_indices),
                           pointer(s->pool->geodesic_tgt_indices), pointer(s->pool->geodesic_rest_length),
                           s->pool->NumGeodesics(), sp.long_range_stretchiness);
      }
      s->vf_ccd->Solve(pointer(s->cloth->nodes), s->obstacle_bvh);
      s->vf_dcd->Solve(pointer(s->cloth->nodes), s->obstacle_bvh);
      s->ee_dcd->Solve(pointer(s->cloth->edges), pointer(s->pool->deltas), pointer(s->pool->delta_counts),
                       s->obstacle_bvh);
      ApplyDeltas(pointer(s->cloth->nodes), pointer(s->pool->deltas), pointer(s->pool->delta_counts));
      s->binding->Solve(pointer(s->cloth->nodes), pointer(s->obstacle->nodes));
    }
    Finalize(pointer(s->cloth->nodes), substep_time);
  }
  s->obstacle->PredictedToPositions();
  s->cloth->UpdateNormals();
  s->obstacle->UpdateNormals();
  Timer::EndTimerGPU("Solver_Total");
  checkCudaErrors(cudaDeviceSynchronize());
  Timer::NextFrame();  // GameInstance::MainLoop: timers of the previous frame are collected on the next Start
  return (float)Timer::GetTimerGPU("Solver_Total");
}

// which: 0 cloth, 1 obstacle; any pointer may be null
void ref_get_state(void* h, int which, float* x0, float* x, float* v, float* nrm) {
  auto* s = use(h);
  auto& mesh = which == 0 ? s->cloth : s->obstacle;
  int n = mesh->NumNodes();
  float* d = s->d_tmp;
  size_t st = s->tmp_floats;
  GetState_Kernel<<<(n + 255) / 256, 256>>>(pointer(mesh->nodes), d, d + st, d + 2 * st, d + 3 * st, n);
  cudaDeviceSynchronize();
  if (x0) cudaMemcpy(x0, d, 3 * (size_t)n * 4, cudaMemcpyDeviceToHost);
  if (x) cudaMemcpy(x, d + st, 3 * (size_t)n * 4, cudaMemcpyDeviceToHost);
  if (v) cudaMemcpy(v, d + 2 * st, 3 * (size_t)n * 4, cudaMemcpyDeviceToHost);
  if (nrm) cudaMemcpy(nrm, d + 3 * st, 3 * (size_t)n * 4, cudaMemcpyDeviceToHost);
}

void ref_set_cloth_state(void* h, const float* x0, const float* x, const float* v) {
  auto* s = use(h);
  int n = s->cloth->NumNodes();
  float* d = s->d_tmp;
  size_t st = s->tmp_floats;
  if (x0) cudaMemcpy(d, x0, 3 * (size_t)n * 4, cudaMemcpyHostToDevice);
  if (x) cudaMemcpy(d + st, x, 3 * (size_t)n * 4, cudaMemcpyHostToDevice);
  if (v) cudaMemcpy(d + 2 * st, v, 3 * (size_t)n * 4, cudaMemcpyHostToDevice);
  SetState_Kernel<<<(n + 255) / 256, 256>>>(pointer(s->cloth->nodes), x0 ? d : nullptr, x ? d + st : nullptr,
                                            v ? d + 2 * st : nullptr, n);
  cudaDeviceSynchronize();
}

int ref_num_colors(void* h, int kind) {
  auto* s = use(h);
  return kind == 0 ? s->stretch->Coloring()->n_colors : s->bend->Coloring()->n_colors;
}
// GraphColoring::colors (graph_coloring/Base.cuh:36), for injection into the product (F9)
void ref_get_colors(void* h, int kind, uint* out, int n) {
  auto* s = use(h);
  if (kind == 0) thrust::copy_n(s->stretch->Coloring()->colors.begin(), n, out);
  else thrust::copy_n(s->bend->Coloring()->colors.begin(), n, out);
}

// overwrite GraphColoring::colors / n_colors (public members, graph_coloring/Base.cuh:36) so that both
// implementations can be driven with the same PROPER colouring (the reference's own randomized colouring
// leaves vertex conflicts inside colours, which makes its result depend on atomic timing)
void ref_set_colors(void* h, int kind, const uint* colors, int n, int n_colors) {
  auto* s = use(h);
  if (kind == 0) {
    thrust::copy_n(colors, n, s->stretch->Coloring()->colors.begin());
    s->stretch->Coloring()->n_colors = n_colors;
  } else {
    thrust::copy_n(colors, n, s->bend->Coloring()->colors.begin());
    s->bend->Coloring()->n_colors = n_colors;
  }
}

// the reference's own host accessors, BVH.cu:579-602
void ref_get_bvh(void* h, uint* morton_unused, uint* nodes, float* aabbs, int* esc_int, int* esc_ext, float* pred9) {
  auto* s = use(h);
  (void)morton_unused;
  auto hn = s->obstacle_bvh->NodesHost();
  auto hb = s->obstacle_bvh->AABBsHost();
  auto hi = s->obstacle_bvh->InternalEscapeIndexTableHost();
  auto he = s->obstacle_bvh->ExternalEscapeIndexTableHost();
  auto ho = s->obstacle_bvh->ObjectsHost();
  const size_t F = ho.size();
  for (size_t i = 0; i < hn.size(); i++) {
    nodes[4 * i] = hn[i].parent_idx; nodes[4 * i + 1] = hn[i].left_idx;
    nodes[4 * i + 2] = hn[i].right_idx; nodes[4 * i + 3] = hn[i].object_idx;
    for (int k = 0; k < 3; k++) { aabbs[6 * i + k] = hb[i].lower[k]; aabbs[6 * i + 3 + k] = hb[i].upper[k]; }
  }
  for (size_t i = 0; i + 1 < F; i++) esc_int[i] = hi[i];
  for (size_t i = 0; i < F; i++) {
    esc_ext[i] = he[i];
    if (pred9)
      for (int k = 0; k < 3; k++) {
        pred9[9 * i + k] = ho[i].pred1[k]; pred9[9 * i + 3 + k] = ho[i].pred2[k]; pred9[9 * i + 6 + k] = ho[i].pred3[k];
      }
  }
}

// contacts as the reference stores them (pbd_collision/Contact.cuh)
void ref_get_contacts(void* h, int* vf_idx, float* vf_q, int* rt_idx, float* rt_q, float* ee12) {
  auto* s = use(h);
  thrust::host_vector<VFContact> a = s->vf_dcd->contacts;
  thrust::host_vector<RTContact> b = s->vf_ccd->contacts;
  thrust::host_vector<EEContact> c = s->ee_dcd->contacts;
  for (size_t i = 0; i < a.size(); i++) {
    vf_idx[i] = a[i].index; rt_idx[i] = b[i].index;
    for (int k = 0; k < 3; k++) { vf_q[3 * i + k] = a[i].qs[k]; rt_q[3 * i + k] = b[i].qc[k]; }
  }
  for (size_t i = 0; i < c.size(); i++) {
    float* o = ee12 + 12 * i;
    o[0] = c[i].dist;
    for (int k = 0; k < 3; k++) { o[1 + k] = c[i].p2[k]; o[4 + k] = c[i].q2[k]; o[9 + k] = c[i].fn[k]; }
    o[7] = c[i].s; o[8] = c[i].t;
  }
}

void ref_generate_contacts(void* h) {
  auto* s = use(h);
  SetSimulationParams(&Global::sim_params);
  s->vf_dcd->Generate(pointer(s->cloth->nodes), s->obstacle_bvh);
  s->vf_ccd->Generate(pointer(s->cloth->nodes), s->obstacle_bvh);
  s->ee_dcd->Generate(pointer(s->cloth->nodes), pointer(s->cloth->edges), s->obstacle_bvh);
  cudaDeviceSynchronize();
}

// BVH::UpdateData(0, false) + Construct on the obstacle's current x0 (what ref_step_frame does for a moved obstacle)
void ref_build_obstacle_bvh(void* h) {
  auto* s = use(h);
  s->obstacle_bvh->UpdateData(pointer(s->obstacle->faces), 0.0f, false, s->obstacle->NumFaces());
  s->obstacle_bvh->Construct();
  cudaDeviceSynchronize();
}
// Morton codes of the obstacle tree's leaves in sorted order: the reference's own ComputeMortons_Kernel run on
// the tree's leaf boxes (sorted with the keys, BVH.cu:666) and root box, i.e. the keys Construct sorted.  Valid
// right after a Construct (a refit grows the leaf boxes).
void ref_get_mortons(void* h, uint* out) {
  auto* s = use(h);
  auto dev = s->obstacle_bvh->GetDeviceRepr();
  const uint F = (uint)s->obstacle->NumFaces();
  thrust::device_vector<uint> m(F);
  ComputeMortons_Kernel<<<(F + 255) / 256, 256>>>(dev.objects, dev.aabbs, s->obstacle_bvh->GetRootAABB(), pointer(m), F - 1, F);
  cudaDeviceSynchronize();
  thrust::copy(m.begin(), m.end(), out);
}
// ClothSolverBase::AddGeodesic (ClothSolverBase.cpp:527-532)
void ref_add_geodesics(void* h, const uint* src, const uint* tgt, const float* rest, int n) {
  auto* s = use(h);
  for (int i = 0; i < n; i++) s->pool->AddGeodesic(src[i], tgt[i], rest[i]);
}

void ref_get_rtri(void* h, uint* out, int n) {
  auto* s = use(h);
  thrust::copy_n(s->obstacle_rtri->r_tris.begin(), n, out);
}

// Quality-mode cloth model on the handle's cloth (row a25): the reference's own FEM::StrainSolver and
// FEM::IsometricBendingSolver, set up as ClothSolverBase::SetupInternalDynamics does
// (pipeline/base/ClothSolverBase.cpp:468-486) and iterated as SimulateQuality does (:315-324), from the
// predicted positions `pos`.  The solvers colour themselves with a random palette; the given colours
// overwrite GraphColoring::colors / n_colors so both sides sweep in the same order.
// material: xx, yy, xy stiffness, xy / yx Poisson ratio, bending stiffness.
int ref_fem_solve(void* h, const float* pos, const float* material, const uint* bend_idx, int nb, const uint* col_s,
                  int ncs, const uint* col_b, int ncb, int iterations, float dt, float* out_pos) {
  auto* s = use(h);
  const int F = s->cloth->NumFaces();
  ref_set_cloth_state(h, nullptr, pos, nullptr);
  FEM::StrainSolver strain(pointer(s->cloth->faces), F, 1234u);
  strain.Init(pointer(s->pool->initial_positions), pointer(s->cloth->faces), material[0], material[1], material[2],
              material[3], material[4]);
  thrust::copy_n(col_s, F, strain.Coloring()->colors.begin());
  strain.Coloring()->n_colors = ncs;
  FEM::IsometricBendingSolver bend;
  if (nb) {
    std::vector<uint> bi(bend_idx, bend_idx + 4 * (size_t)nb);
    std::vector<Scalar> ba(nb, 0.0f);
    bend.AddCloth(bi, ba);
    bend.SetupGraphColoring(1234u);
    bend.Init(pointer(s->pool->initial_positions), material[5]);
    thrust::copy_n(col_b, nb, bend.Coloring()->colors.begin());
    bend.Coloring()->n_colors = ncb;
  }
  for (int it = 0; it < iterations; it++) {
    for (int c = 0; c < ncs; c++) strain.Solve(pointer(s->cloth->nodes), c);
    if (nb)
      for (int c = 0; c < ncb; c++) bend.Solve(pointer(s->cloth->nodes), c, it, dt);
  }
  if (cudaDeviceSynchronize() != cudaSuccess) return 1;
  ref_get_state(h, 0, nullptr, out_pos, nullptr, nullptr);
  return 0;
}

// Face-face broad phase (row a16) by the reference's own BVH::UpdateData + Construct + Traverse
// (BVH.cu:565-573,618-722,1315-1340) on the cloth's and the obstacle's current positions (x0).
// self != 0: the cloth tree against itself.  Writes at most cap pairs (face indices local to each
// mesh); returns the number found.
__global__ void PairIndex_Kernel(const PairFF* p, uint* out, int n, int off_l, int off_r) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  out[2 * i] = (uint)(p[i].first->index - off_l);
  out[2 * i + 1] = (uint)(p[i].second->index - off_r);
}
long long ref_traverse_pairs(void* h, int self, float thickness, uint* out, long long cap) {
  auto* s = use(h);
  auto cloth_bvh = std::make_shared<XRTailor::BVH>();
  cloth_bvh->UpdateData(pointer(s->cloth->faces), 0.0f, false, s->cloth->NumFaces());
  cloth_bvh->Construct();
  auto rhs = cloth_bvh;
  if (!self) {
    rhs = std::make_shared<XRTailor::BVH>();
    rhs->UpdateData(pointer(s->obstacle->faces), 0.0f, false, s->obstacle->NumFaces());
    rhs->Construct();
  }
  Scalar th = thickness;
  thrust::device_vector<PairFF> pairs =
      Traverse(cloth_bvh, rhs, pointer(s->cloth->faces), pointer(self ? s->cloth->faces : s->obstacle->faces), th);
  const long long n = (long long)pairs.size();
  if (n && n <= cap) {
    thrust::device_vector<uint> idx(2 * (size_t)n);
    // Face::index is local to its mesh (PhysicsMeshHelper.cu:52)
    PairIndex_Kernel<<<(unsigned)((n + 255) / 256), 256>>>(pointer(pairs), pointer(idx), (int)n, 0, 0);
    cudaDeviceSynchronize();
    thrust::copy(idx.begin(), idx.end(), out);
  }
  return n;
}

// ---------------------------------------------------------------------------------------------
// Quality mode.  ref_setup_quality restates what ClothSolverBase::Start (ClothSolverBase.cpp:129-139),
// ClothSolverUniversal::Start (impl/universal/ClothSolver.cpp:66-79: SetupRTriangle, EF adjacency) and
// SetupInternalDynamics (ClothSolverBase.cpp:468-486) do for SOLVER_MODE::QUALITY.
// qp: imminent_on, pbd_on, n strain colours, n bend colours; qf: material xx, yy, xy, poisson xy, yx, bending,
// imminent_thickness, pbd_thickness, relaxation_rate, obstacle_mass, impact-zone thickness.
static void EFAdjacency(const uint* tri, int F, const uint* edges, int E, std::vector<std::set<uint>>& out) {
  std::vector<uint> fe = FaceEdgeIndices(tri, F, edges, E);
  out.assign(E, std::set<uint>());
  for (int f = 0; f < F; f++)
    for (int j = 0; j < 3; j++) out[fe[3 * f + j]].insert((uint)f);
}
void ref_setup_quality(void* h, const int* qp, const float* qf, const uint* ctri, const uint* cedge, const uint* otri,
                       const uint* oedge, int Eo, const uint* bidx, int nb, const uint* col_s, const uint* col_b) {
  auto* s = use(h);
  SimParams& sp = Global::sim_params;
  sp.solver_mode = SOLVER_MODE::QUALITY;
  sp.imminent_repulsion = qp[0] != 0;
  sp.pbd_repulsion = qp[1] != 0;
  s->imminent_on = qp[0] != 0; s->pbd_on = qp[1] != 0;
  s->imminent_thickness = qf[6]; s->pbd_thickness = qf[7]; s->relaxation_rate = qf[8];
  s->obstacle_mass = qf[9]; s->iz_thickness = qf[10];
  const int F = s->cloth->NumFaces(), E = s->cloth->NumEdges(), Fo = s->obstacle->NumFaces();
  // SetupRTriangle (ClothSolverBase.cpp:487-521): both meshes, registered on the faces
  if (Fo) s->obstacle->RegisterRTris(s->obstacle_rtri->r_tris);
  BuildVFAdjacency(s->cloth, s->pool->cloth_nb_vf, s->pool->cloth_nb_vf_prefix);
  s->cloth_rtri = std::make_shared<RTriangle>();
  s->cloth_rtri->Init(s->cloth->DeviceIndices(), s->pool->cloth_nb_vf, s->pool->cloth_nb_vf_prefix);
  s->cloth->RegisterRTris(s->cloth_rtri->r_tris);
  // EF adjacency (BuildEFAdjacency + BuildDeviceEFAdjacency + RegisterEFAdjacency)
  std::vector<std::set<uint>> ef;
  EFAdjacency(ctri, F, cedge, E, ef);
  BuildDeviceEFAdjacency(ef, s->pool->cloth_nb_ef, s->pool->cloth_nb_ef_prefix);
  s->cloth->RegisterEFAdjacency(s->pool->cloth_nb_ef, s->pool->cloth_nb_ef_prefix);
  if (Fo) {
    EFAdjacency(otri, Fo, oedge, Eo, ef);
    BuildDeviceEFAdjacency(ef, s->pool->obstacle_nb_ef, s->pool->obstacle_nb_ef_prefix);
    s->obstacle->RegisterEFAdjacency(s->pool->obstacle_nb_ef, s->pool->obstacle_nb_ef_prefix);
  }
  s->cloth_bvh = std::make_shared<XRTailor::BVH>();
  s->imminent = std::make_shared<ImminentRepulsion>();
  s->imminent->InitializeDeltas(s->cloth->NumNodes());
  s->pbd = std::make_shared<PBDRepulsionSolver>(s->cloth.get());
  // FEM solvers with the caller's colours (the reference colours them randomly)
  s->fem_strain = std::make_shared<FEM::StrainSolver>(pointer(s->cloth->faces), F, 1234u);
  s->fem_strain->Init(pointer(s->pool->initial_positions), pointer(s->cloth->faces), qf[0], qf[1], qf[2], qf[3], qf[4]);
  thrust::copy_n(col_s, F, s->fem_strain->Coloring()->colors.begin());
  s->fem_strain->Coloring()->n_colors = qp[2];
  s->fem_bend = std::make_shared<FEM::IsometricBendingSolver>();
  if (nb && sp.solve_bending) {
    std::vector<uint> bi(bidx, bidx + 4 * (size_t)nb);
    std::vector<Scalar> ba(nb, 0.0f);
    s->fem_bend->AddCloth(bi, ba);
    s->fem_bend->SetupGraphColoring(1234u);
    s->fem_bend->Init(pointer(s->pool->initial_positions), qf[5]);
    thrust::copy_n(col_b, nb, s->fem_bend->Coloring()->colors.begin());
    s->fem_bend->Coloring()->n_colors = qp[3];
  }
  cudaDeviceSynchronize();
  s->sp = sp;
}

// ClothSolverBase::SimulateQuality, pipeline/base/ClothSolverBase.cpp:263-362 (icm_enable is off: the ICM object
// the reference allocates per frame and GIADetangleStep are left out); returns Solver_Total in ms
float ref_step_frame_quality(void* h) {
  auto* s = use(h);
  SimParams& sp = Global::sim_params;
  Timer::StartTimerGPU("Solver_Total");
  Scalar substep_time = Timer::FixedDeltaTime() / sp.num_substeps;
  SetSimulationParams(&sp);
  if (s->target_pending) {  // UpdateObstacleAnimationFrame, impl/smpl/ClothSolver.cpp:257-261
    ScopedTimerGPU timer("Solver_UpdateBVH");
    s->obstacle_bvh->UpdateData(pointer(s->obstacle->faces), 0.0f, false, s->obstacle->NumFaces());
    s->obstacle_bvh->Construct();
    s->target_pending = false;
  }
  s->cloth_bvh->UpdateData(pointer(s->cloth->faces), 0.0f, false, s->cloth->NumFaces());
  s->cloth_bvh->Construct();
  if (s->imminent_on) {
    s->imminent->UpdateProximity(s->cloth.get(), s->obstacle.get(), s->cloth_bvh.get(), s->obstacle_bvh.get(), 1e-6f);
    s->imminent->Generate();
  }
  s->obstacle_bvh->Update(pointer(s->obstacle->faces), true);
  for (int substep = 0; substep < sp.num_substeps; substep++) {
    PredictPositions(pointer(s->cloth->nodes), substep_time, s->cloth->NumNodes());
    for (int iteration = 0; iteration < sp.num_iterations; iteration++) {
      if (s->imminent_on)
        s->imminent->Solve(s->cloth.get(), s->imminent_thickness, true, s->relaxation_rate, substep_time, s->frame_index,
                           iteration);
      if (s->pbd_on) {
        if (iteration % 10 == 0)
          s->pbd->UpdatePairs(s->cloth.get(), s->obstacle.get(), s->cloth_bvh.get(), s->obstacle_bvh.get(), 1e-6f);
        s->pbd->SolveNoProximity(s->cloth.get(), s->pbd_thickness, s->relaxation_rate, substep_time, s->frame_index,
                                 iteration);
      }
      for (int c = 0; c < s->fem_strain->Coloring()->n_colors; c++) s->fem_strain->Solve(pointer(s->cloth->nodes), c);
      if (sp.solve_bending && s->fem_bend->Coloring())
        for (int c = 0; c < s->fem_bend->Coloring()->n_colors; c++)
          s->fem_bend->Solve(pointer(s->cloth->nodes), c, iteration, substep_time);
      {
        ScopedTimerGPU timer("Solver_SolveLRA");
        if (sp.geodesic_LRA)
          SolveGeodesicLRA(pointer(s->cloth->nodes), pointer(s->pool->geodesic_src_indices),
                           pointer(s->pool->geodesic_tgt_indices), pointer(s->pool->geodesic_rest_length),
                           s->pool->NumGeodesics(), sp.long_range_stretchiness);
      }
      s->binding->Solve(pointer(s->cloth->nodes), pointer(s->obstacle->nodes));
    }
    Timer::StartTimerGPU("Solver_CollisionStep");
    ImpactZoneOptimization::CollisionStep(s->cloth, s->obstacle, s->pool, s->frame_index, substep_time, s->obstacle_mass,
                                          s->iz_thickness);
    Timer::EndTimerGPU("Solver_CollisionStep");
    Finalize(pointer(s->cloth->nodes), substep_time);
  }
  s->obstacle->PredictedToPositions();
  s->cloth->UpdateNormals();
  s->obstacle->UpdateNormals();
  Timer::EndTimerGPU("Solver_Total");
  checkCudaErrors(cudaDeviceSynchronize());
  Timer::NextFrame();
  s->frame_index++;
  return (float)Timer::GetTimerGPU("Solver_Total");
}

// -- pieces, for the parity tests ------------------------------------------------------------------
__global__ void SnapshotX1_Kernel(Node** nodes, int n, float dt, int obstacle) {  // SetupNode / SetupObstacleNode, Entry.cu:110-124
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  nodes[i]->x1 = nodes[i]->x;
  if (obstacle) {
    nodes[i]->is_free = false;
    nodes[i]->v = (nodes[i]->x - nodes[i]->x0) / dt;
    nodes[i]->removed = false;
  }
}
__global__ void ImpactRows_Kernel(const Impact* imp, int n, int V, uint* nodes, float* twn) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  for (int j = 0; j < 4; j++) nodes[4 * i + j] = (uint)(imp[i].nodes[j]->index + (imp[i].nodes[j]->is_cloth ? 0 : V));
  float* o = twn + 9 * i;
  o[0] = imp[i].t;
  for (int j = 0; j < 4; j++) o[1 + j] = imp[i].w[j];
  for (int j = 0; j < 3; j++) o[5 + j] = imp[i].n[j];
  o[8] = imp[i].is_vf ? 1.0f : 0.0f;
}
static long long ImpactsOut(RefSolver* s, const thrust::device_vector<Impact>& imp, uint* nodes, float* twn, long long cap) {
  const long long n = (long long)imp.size();
  if (n && n <= cap && nodes) {
    thrust::device_vector<uint> dn(4 * (size_t)n);
    thrust::device_vector<float> dt(9 * (size_t)n);
    ImpactRows_Kernel<<<(unsigned)((n + 255) / 256), 256>>>(pointer(imp), (int)n, s->cloth->NumNodes(), pointer(dn), pointer(dt));
    cudaDeviceSynchronize();
    thrust::copy(dn.begin(), dn.end(), nodes);
    thrust::copy(dt.begin(), dt.end(), twn);
  }
  return n;
}
// FindImpacts (ImpactHelper.cu:216-243) on the current x0 / x with the two fresh ccd trees CollisionStep builds
// (Entry.cu:146-155).  Rows: 4 global node ids (obstacle ids offset by V) | t, w[4], n[3], is_vf.  The list is
// kept for ref_independent_impacts.
long long ref_find_impacts(void* h, float thickness, float dt, uint* nodes, float* twn, long long cap, long long* n_pairs) {
  auto* s = use(h);
  const int V = s->cloth->NumNodes(), Vo = s->obstacle->NumNodes();
  SnapshotX1_Kernel<<<(V + 255) / 256, 256>>>(pointer(s->cloth->nodes), V, dt, 0);
  if (Vo) SnapshotX1_Kernel<<<(Vo + 255) / 256, 256>>>(pointer(s->obstacle->nodes), Vo, dt, 1);
  auto cb = std::make_shared<XRTailor::BVH>();
  auto ob = std::make_shared<XRTailor::BVH>();
  cb->UpdateData(pointer(s->cloth->faces), 0.0f, true, s->cloth->NumFaces());
  cb->Construct(0.0f);
  ob->UpdateData(pointer(s->obstacle->faces), 0.0f, true, s->obstacle->NumFaces());
  ob->Construct(0.0f);
  if (n_pairs) {
    Scalar th = thickness;
    *n_pairs = (long long)Traverse(cb, ob, pointer(s->cloth->faces), pointer(s->obstacle->faces), th).size() +
               (long long)Traverse(cb, cb, pointer(s->cloth->faces), pointer(s->cloth->faces), th).size();
  }
  s->last_impacts = FindImpacts(cb, ob, pointer(s->cloth->faces), pointer(s->obstacle->faces), thickness, 0, 0);
  cudaDeviceSynchronize();
  return ImpactsOut(s, s->last_impacts, nodes, twn, cap);
}
// CUDA-event times (ms, mean of reps) of the reference's own stages on the current state, for the C4 scaling table:
// out[0] cloth BVH UpdateData(ccd) + Construct, [1] BVH::Update(ccd), [2] the two Traverse calls of FindImpacts,
// [3] FindImpacts as a whole (its Traverses + CheckImpacts_Kernel + remove_if); counts: pairs, impacts
void ref_quality_timing(void* h, float thickness, float dt, int reps, float* out, long long* counts) {
  auto* s = use(h);
  const int V = s->cloth->NumNodes(), Vo = s->obstacle->NumNodes();
  SnapshotX1_Kernel<<<(V + 255) / 256, 256>>>(pointer(s->cloth->nodes), V, dt, 0);
  if (Vo) SnapshotX1_Kernel<<<(Vo + 255) / 256, 256>>>(pointer(s->obstacle->nodes), Vo, dt, 1);
  auto cb = std::make_shared<XRTailor::BVH>();
  auto ob = std::make_shared<XRTailor::BVH>();
  ob->UpdateData(pointer(s->obstacle->faces), 0.0f, true, s->obstacle->NumFaces());
  ob->Construct(0.0f);
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0);
  cudaEventCreate(&e1);
  auto timed = [&](auto&& fn) {
    fn();  // warm-up (first-touch allocations of thrust)
    cudaDeviceSynchronize();
    cudaEventRecord(e0);
    for (int r = 0; r < reps; r++) fn();
    cudaEventRecord(e1);
    cudaEventSynchronize(e1);
    float ms = 0.f;
    cudaEventElapsedTime(&ms, e0, e1);
    return ms / (float)reps;
  };
  out[0] = timed([&] {
    cb->UpdateData(pointer(s->cloth->faces), 0.0f, true, s->cloth->NumFaces());
    cb->Construct(0.0f);
  });
  out[1] = timed([&] { cb->Update(pointer(s->cloth->faces), true); });
  long long np = 0;
  Scalar th = thickness;
  out[2] = timed([&] {
    np = (long long)Traverse(cb, ob, pointer(s->cloth->faces), pointer(s->obstacle->faces), th).size() +
         (long long)Traverse(cb, cb, pointer(s->cloth->faces), pointer(s->cloth->faces), th).size();
  });
  long long ni = 0;
  out[3] = timed([&] { ni = (long long)FindImpacts(cb, ob, pointer(s->cloth->faces), pointer(s->obstacle->faces), thickness, 0, 0).size(); });
  counts[0] = np;
  counts[1] = ni;
  cudaEventDestroy(e0);
  cudaEventDestroy(e1);
}
// independentImpacts (Entry.cu:25-70) on the list ref_find_impacts kept
long long ref_independent_impacts(void* h, int deform, uint* nodes, float* twn, long long cap) {
  auto* s = use(h);
  thrust::device_vector<Impact> ind = ImpactZoneOptimization::independentImpacts(s->last_impacts, deform);
  cudaDeviceSynchronize();
  return ImpactsOut(s, ind, nodes, twn, cap);
}
// CollisionStep (Entry.cu:140-256) on the current state
void ref_collision_step(void* h, float dt) {
  auto* s = use(h);
  ImpactZoneOptimization::CollisionStep(s->cloth, s->obstacle, s->pool, s->frame_index, dt, s->obstacle_mass, s->iz_thickness);
  cudaDeviceSynchronize();
}
// cloth_bvh UpdateData+Construct, ImminentRepulsion::UpdateProximity + Generate (ClothSolverBase.cpp:277-284);
// returns the number of quadratures; rows: 4 global node ids | 4 barycentric weights
__global__ void QuadRows_Kernel(const Quadrature* q, int n, int V, uint* nodes, float* bary) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  for (int j = 0; j < 4; j++) {
    nodes[4 * i + j] = (uint)(q[i].nodes[j]->index + (q[i].nodes[j]->is_cloth ? 0 : V));
    bary[4 * i + j] = q[i].bary[j];
  }
}
long long ref_imminent_generate(void* h, uint* nodes, float* bary, long long cap) {
  auto* s = use(h);
  SetSimulationParams(&Global::sim_params);
  s->cloth_bvh->UpdateData(pointer(s->cloth->faces), 0.0f, false, s->cloth->NumFaces());
  s->cloth_bvh->Construct();
  s->imminent->UpdateProximity(s->cloth.get(), s->obstacle.get(), s->cloth_bvh.get(), s->obstacle_bvh.get(), 1e-6f);
  s->imminent->Generate();
  cudaDeviceSynchronize();
  const long long n = (long long)s->imminent->quads_.size();
  if (n && n <= cap && nodes) {
    thrust::device_vector<uint> dn(4 * (size_t)n);
    thrust::device_vector<float> db(4 * (size_t)n);
    QuadRows_Kernel<<<(unsigned)((n + 255) / 256), 256>>>(pointer(s->imminent->quads_), (int)n, s->cloth->NumNodes(), pointer(dn), pointer(db));
    cudaDeviceSynchronize();
    thrust::copy(dn.begin(), dn.end(), nodes);
    thrust::copy(db.begin(), db.end(), bary);
  }
  return n;
}
void ref_imminent_solve(void* h, float dt, int iteration) {
  auto* s = use(h);
  s->imminent->Solve(s->cloth.get(), s->imminent_thickness, true, s->relaxation_rate, dt, s->frame_index, iteration);
  cudaDeviceSynchronize();
}
// PBDRepulsionSolver::UpdatePairs (PBDRepulsion.cu:22-33) -> number of pairs; rows (cloth face, other face; obstacle
// faces offset by F)
__global__ void PairRows_Kernel(const PairFF* p, int n, int F, uint* out) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  out[2 * i] = (uint)p[i].first->index;
  const bool cloth = p[i].second->nodes[0]->is_cloth;
  out[2 * i + 1] = (uint)(p[i].second->index + (cloth ? 0 : F));
}
long long ref_rep_update_pairs(void* h, uint* out, long long cap) {
  auto* s = use(h);
  s->pbd->UpdatePairs(s->cloth.get(), s->obstacle.get(), s->cloth_bvh.get(), s->obstacle_bvh.get(), 1e-6f);
  cudaDeviceSynchronize();
  const long long n = (long long)s->pbd->pairs_.size();
  if (n && n <= cap && out) {
    thrust::device_vector<uint> d(2 * (size_t)n);
    PairRows_Kernel<<<(unsigned)((n + 255) / 256), 256>>>(pointer(s->pbd->pairs_), (int)n, s->cloth->NumFaces(), pointer(d));
    cudaDeviceSynchronize();
    thrust::copy(d.begin(), d.end(), out);
  }
  return n;
}
void ref_rep_solve(void* h, float dt, int iteration) {
  auto* s = use(h);
  s->pbd->SolveNoProximity(s->cloth.get(), s->pbd_thickness, s->relaxation_rate, dt, s->frame_index, iteration);
  cudaDeviceSynchronize();
}
// Every contribution of one SolveNoProximity pass to cloth vertex `vertex`, evaluated by the reference's own
// EvaluateVFProximity / EvaluateEEProximity over the solver's pair list in EvaluateRep_Kernel's order
// (PBDRepulsion.cu:470-500).  rows: pair, slot (0-2 VF, 3-11 EE), 4 global node ids | corr xyz of that vertex.
__global__ void RepContrib_Kernel(const PairFF* pairs, int n_pairs, Scalar h, int vertex, int V, int* n_out, int* rows, float* corr, int cap) {
  for (int pid = 0; pid < n_pairs; pid++) {
    Face* face0 = pairs[pid].first;
    Face* face1 = pairs[pid].second;
    for (int slot = 0; slot < 12; slot++) {
      Impulse impulse;
      bool hit = false;
      if (slot < 3) {
        if (!RTriVertex(face0->r_tri, slot)) continue;
        hit = EvaluateVFProximity(face0->nodes[slot], face1, h, &impulse);
      } else {
        const int i = (slot - 3) / 3, j = (slot - 3) % 3;
        if (!RTriEdge(face0->r_tri, i) || !RTriEdge(face1->r_tri, j)) continue;
        hit = EvaluateEEProximity(face0->edges[i], face1->edges[j], h, &impulse);
      }
      if (!hit) continue;
      for (int k = 0; k < 4; k++) {
        Node* nd = impulse.nodes[k];
        if (!nd->is_cloth || nd->index != vertex) continue;
        if (glm::length(impulse.corrs[k]) < EPSILON || !nd->is_free) continue;
        const int at = (*n_out)++;
        if (at < cap) {
          rows[6 * at] = pid; rows[6 * at + 1] = slot;
          for (int q = 0; q < 4; q++) rows[6 * at + 2 + q] = impulse.nodes[q]->index + (impulse.nodes[q]->is_cloth ? 0 : V);
          for (int q = 0; q < 3; q++) corr[3 * at + q] = impulse.corrs[k][q];
        }
      }
    }
  }
}
// the reference's EvaluateRep_Kernel alone (no apply): the Jacobi sums and counts it leaves behind (debug)
void ref_rep_evaluate_only(void* h, float* deltas, int* counts) {
  auto* s = use(h);
  const int V = s->cloth->NumNodes();
  const int n_pairs = (int)s->pbd->pairs_.size();
  thrust::device_vector<Vector3> d(V, Vector3(0));
  thrust::device_vector<int> c(V, 0);
  EvaluateRep_Kernel<<<(n_pairs + 255) / 256, 256>>>(pointer(s->pbd->pairs_), pointer(d), pointer(c), s->pbd_thickness, n_pairs);
  cudaDeviceSynchronize();
  thrust::host_vector<Vector3> hd = d;
  thrust::host_vector<int> hc = c;
  for (int i = 0; i < V; i++) {
    for (int k = 0; k < 3; k++) deltas[3 * i + k] = hd[i][k];
    counts[i] = hc[i];
  }
}
int ref_rep_contributions(void* h, int vertex, int* rows, float* corr, int cap) {
  auto* s = use(h);
  thrust::device_vector<int> n(1, 0), dr(6 * (size_t)cap);
  thrust::device_vector<float> dc(3 * (size_t)cap);
  RepContrib_Kernel<<<1, 1>>>(pointer(s->pbd->pairs_), (int)s->pbd->pairs_.size(), s->pbd_thickness, vertex, s->cloth->NumNodes(),
                              pointer(n), pointer(dr), pointer(dc), cap);
  cudaDeviceSynchronize();
  thrust::copy(dr.begin(), dr.end(), rows);
  thrust::copy(dc.begin(), dc.end(), corr);
  return n[0];
}
// BVH::Update(ccd = true) on the obstacle tree (ClothSolverBase.cpp:291)
void ref_obstacle_bvh_update(void* h) {
  auto* s = use(h);
  s->obstacle_bvh->Update(pointer(s->obstacle->faces), true);
  cudaDeviceSynchronize();
}
void ref_get_cloth_rtri(void* h, uint* out, int n) {
  auto* s = use(h);
  thrust::copy_n(s->cloth_rtri->r_tris.begin(), n, out);
}
// obstacle state (x0, x, v), for the tests that load one state into three implementations
void ref_set_obstacle_state(void* h, const float* x0, const float* x, const float* v) {
  auto* s = use(h);
  int n = s->obstacle->NumNodes();
  float* d = s->d_tmp;
  size_t st = s->tmp_floats;
  if (x0) cudaMemcpy(d, x0, 3 * (size_t)n * 4, cudaMemcpyHostToDevice);
  if (x) cudaMemcpy(d + st, x, 3 * (size_t)n * 4, cudaMemcpyHostToDevice);
  if (v) cudaMemcpy(d + 2 * st, v, 3 * (size_t)n * 4, cudaMemcpyHostToDevice);
  SetState_Kernel<<<(n + 255) / 256, 256>>>(pointer(s->obstacle->nodes), x0 ? d : nullptr, x ? d + st : nullptr,
                                            v ? d + 2 * st : nullptr, n);
  cudaDeviceSynchronize();
}

void ref_destroy(void* h) {
  auto* s = use(h);
  cudaFree(s->d_tmp);
  s->cloth->Destroy();
  s->obstacle->Destroy();
  delete s;
}
}  // extern "C"
