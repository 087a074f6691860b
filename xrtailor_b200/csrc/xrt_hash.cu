// Swift self-contact: spatial hash -> cub radix sort -> cell ranges ->
// neighbour cache -> particle push-out with friction.
// Follows xrtailor/src/physics/broad_phase/spatial_hash/SpatialHashGPU.cu:9-160
// and CollideParticles_Kernel (physics/ClothSolverHelper.cu:150-211).
#include "xrt_dev_hash.cuh"
#include <cub/device/device_radix_sort.cuh>

namespace {

// floor(v / cell) as the reference computes it (SpatialHashGPU.cu:6-8), clamped to +-2^30 so
// the 3x3x3 cell loops cannot overflow (the reference's are undefined there): a garment that
// has blown up to 1e12 m must not hang the device.  NaN maps to the lower clamp.
__device__ __forceinline__ int int_coord(float v, float cell) {
  return (int)fminf(fmaxf(floorf(v / cell), -1073741824.0f), 1073741824.0f);
}
__device__ __forceinline__ int hash_coords(int x, int y, int z, int table) {
  const int h = (int)(((uint32_t)x * 92837111u) ^ ((uint32_t)y * 689287499u) ^ ((uint32_t)z * 283923481u));
  return abs(h % table);
}

__global__ void particle_hash_kernel(const float4* __restrict__ x, uint32_t* __restrict__ hash,
                                     uint32_t* __restrict__ index, float cell, int table, uint32_t V) {
  uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= V) return;
  const float4 p = x[i];
  hash[i] = (uint32_t)hash_coords(int_coord(p.x, cell), int_coord(p.y, cell), int_coord(p.z, cell), table);
  index[i] = i;
}

// FindCellStart_Kernel (SpatialHashGPU.cu:27-55); neighbours' hashes come
// through a shuffle + one shared word per warp boundary instead of a
// (block+1)-word shared array.
__global__ void cell_range_kernel(uint32_t* __restrict__ cell_start, uint32_t* __restrict__ cell_end,
                                  const uint32_t* __restrict__ hash, uint32_t V) {
  uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= V) return;
  const uint32_t h = hash[i];
  const uint32_t prev = i > 0 ? hash[i - 1] : 0xFFFFFFFFu;
  if (i == 0 || h != prev) {
    cell_start[h] = i;
    if (i > 0) cell_end[prev] = i;
  }
  if (i == V - 1) cell_end[h] = i + 1;
}

// CacheNeighbors_Kernel, SpatialHashGPU.cu:56-100
__global__ void cache_neighbors_kernel(uint32_t* __restrict__ neighbors, const uint32_t* __restrict__ index,
                                       const uint32_t* __restrict__ cell_start,
                                       const uint32_t* __restrict__ cell_end, const float4* __restrict__ x,
                                       const float4* __restrict__ rest, StepConst k, uint32_t V) {
  uint32_t tid = blockIdx.x * blockDim.x + threadIdx.x;
  if (tid >= V) return;
  const uint32_t id = index[tid];
  const f3 p = mk3(x[id]), r = mk3(rest[id]);
  const float cell2 = k.cell_spacing * k.cell_spacing, diam2 = k.particle_diameter * k.particle_diameter;
  const int ix = int_coord(p.x, k.cell_spacing), iy = int_coord(p.y, k.cell_spacing),
            iz = int_coord(p.z, k.cell_spacing);
  const long long limit = (long long)V * k.max_neighbors;
  long long ni = id;
  // same cell order as the reference; the 9 cell ranges of a plane are fetched together so their
  // latencies overlap (the kernel is a chain of dependent loads: hash -> range -> index -> position)
  for (int cx = ix - 1; cx <= ix + 1; cx++) {
    int st[9], en[9];
#pragma unroll
    for (int j = 0; j < 9; j++) {
      const int h = hash_coords(cx, iy - 1 + j / 3, iz - 1 + j % 3, k.table_size);
      st[j] = (int)cell_start[h];
      en[j] = (int)cell_end[h];  // meaningless where the cell is empty; not used then
    }
#pragma unroll
    for (int j = 0; j < 9; j++) {
      const int start = st[j];
      if (start == -1) continue;
      const int end = min(en[j], start + k.max_neighbors);
      for (int s = start; s < end; s++) {
        const uint32_t nb = index[s];
        if (nb == id) continue;
        const f3 d = p - mk3(x[nb]);
        if (!(dot3(d, d) < cell2)) continue;
        const f3 dr = r - mk3(rest[nb]);
        if (!(dot3(dr, dr) > diam2)) continue;
        neighbors[ni] = nb;
        ni += V;
        if (ni >= limit) return;
      }
    }
  }
  if (ni < limit) neighbors[ni] = 0xFFFFFFFFu;
}

// ComputeFriction + CollideParticles_Kernel, ClothSolverHelper.cu:150-211
__global__ void collide_particles_kernel(float4* __restrict__ delta, const float4* __restrict__ x,
                                         const float4* __restrict__ x0, const uint32_t* __restrict__ neighbors,
                                         StepConst k, uint32_t V) {
  uint32_t id = blockIdx.x * blockDim.x + threadIdx.x;
  if (id < V) xrtdev::collide_item(delta, x, x0, neighbors, k, V, id);
}

}  // namespace

static int hash_max_bit(int table) {
  int b = 0;
  while ((1ll << b) < (long long)table) b++;  // ceil(log2(table_size)), SpatialHashGPU.cu:139
  return b;
}

int hash_alloc(ClothDev* c, const StepConst& k) {
  const size_t V = c->V;
  cudaError_t e = cudaSuccess;
  auto A = [&](void** p, size_t bytes) {
    if (e == cudaSuccess) e = cudaMalloc(p, bytes ? bytes : 16);
  };
  A((void**)&c->p_hash, V * 4); A((void**)&c->p_index, V * 4);
  A((void**)&c->p_hash_b, V * 4); A((void**)&c->p_index_b, V * 4);
  A((void**)&c->cell_start, (size_t)k.table_size * 4);
  A((void**)&c->cell_end, (size_t)k.table_size * 4);
  A((void**)&c->neighbors, V * (size_t)k.max_neighbors * 4);
  if (e != cudaSuccess) return (int)e;
  cudaMemset(c->cell_end, 0, (size_t)k.table_size * 4);
  cudaMemset(c->neighbors, 0xFF, V * (size_t)k.max_neighbors * 4);
  c->hash_tmp_bytes = 0;
  e = cub::DeviceRadixSort::SortPairs(nullptr, c->hash_tmp_bytes, c->p_hash, c->p_hash_b, c->p_index, c->p_index_b,
                                      (int)V, 0, hash_max_bit(k.table_size));
  A(&c->hash_tmp, c->hash_tmp_bytes);
  return (int)e;
}

// SpatialHashGPU::Hash -> HashObjects, SpatialHashGPU.cu:128-160
void launch_hash(const ClothDev& c, const StepConst& k, cudaStream_t st) {
  const uint32_t V = c.V;
  if (!V) return;
  particle_hash_kernel<<<xrt_grid(V), XRT_BLOCK, 0, st>>>(c.x, c.p_hash_b, c.p_index_b, k.cell_spacing,
                                                          k.table_size, V);
  size_t bytes = c.hash_tmp_bytes;
  cub::DeviceRadixSort::SortPairs(c.hash_tmp, bytes, c.p_hash_b, c.p_hash, c.p_index_b, c.p_index, (int)V, 0,
                                  hash_max_bit(k.table_size), st);
  cudaMemsetAsync(c.cell_start, 0xFF, (size_t)k.table_size * 4, st);
  cell_range_kernel<<<xrt_grid(V), XRT_BLOCK, 0, st>>>(c.cell_start, c.cell_end, c.p_hash, V);
  cache_neighbors_kernel<<<xrt_grid(V), XRT_BLOCK, 0, st>>>(c.neighbors, c.p_index, c.cell_start, c.cell_end, c.x,
                                                            c.rest, k, V);
}

void launch_collide_particles(const ClothDev& c, const StepConst& k, cudaStream_t st) {
  if (!c.V) return;
  collide_particles_kernel<<<xrt_grid(c.V), XRT_BLOCK, 0, st>>>(c.delta, c.x, c.x0, c.neighbors, k, c.V);
  launch_apply_deltas(c, k, st);
}
