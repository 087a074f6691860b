// Swift self-contact: spatial hash -> cub radix sort -> cell ranges ->
// neighbour cache -> particle push-out with friction.
// Follows xrtailor/src/physics/broad_phase/spatial_hash/SpatialHashGPU.cu:9-160
// and CollideParticles_Kernel (physics/ClothSolverHelper.cu:150-211).
#include "xrt_dev_hash.cuh"
#include "xrt_dev_dynamics.cuh"
#include <cub/device/device_radix_sort.cuh>
#include <cub/device/device_segmented_radix_sort.cuh>
#include <algorithm>
#include <vector>

static int hash_max_bit(int table) {
  int b = 0;
  while ((1ll << b) < (long long)table) b++;  // ceil(log2(table_size)), SpatialHashGPU.cu:139
  return b;
}

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

__device__ __forceinline__ void particle_hash_body(const float4* __restrict__ x, uint32_t* __restrict__ hash,
                                                   uint32_t* __restrict__ index, float cell, int table, uint32_t V, uint32_t i) {
  if (i >= V) return;
  const float4 p = x[i];
  hash[i] = (uint32_t)hash_coords(int_coord(p.x, cell), int_coord(p.y, cell), int_coord(p.z, cell), table);
  index[i] = i;
}
__global__ void particle_hash_kernel(const float4* __restrict__ x, uint32_t* __restrict__ hash,
                                     uint32_t* __restrict__ index, float cell, int table, uint32_t V) {
  particle_hash_body(x, hash, index, cell, table, V, blockIdx.x * blockDim.x + threadIdx.x);
}

// FindCellStart_Kernel (SpatialHashGPU.cu:27-55); neighbours' hashes come
// through a shuffle + one shared word per warp boundary instead of a
// (block+1)-word shared array.
__device__ __forceinline__ void cell_range_body(uint32_t* __restrict__ cell_start, uint32_t* __restrict__ cell_end,
                                                const uint32_t* __restrict__ hash, uint32_t V, uint32_t i) {
  if (i >= V) return;
  const uint32_t h = hash[i];
  const uint32_t prev = i > 0 ? hash[i - 1] : 0xFFFFFFFFu;
  if (i == 0 || h != prev) {
    cell_start[h] = i;
    if (i > 0) cell_end[prev] = i;
  }
  if (i == V - 1) cell_end[h] = i + 1;
}
__global__ void cell_range_kernel(uint32_t* __restrict__ cell_start, uint32_t* __restrict__ cell_end,
                                  const uint32_t* __restrict__ hash, uint32_t V) {
  cell_range_body(cell_start, cell_end, hash, V, blockIdx.x * blockDim.x + threadIdx.x);
}

// CacheNeighbors_Kernel, SpatialHashGPU.cu:56-100
__device__ __forceinline__ void cache_neighbors_body(uint32_t* __restrict__ neighbors, const uint32_t* __restrict__ index,
                                                     const uint32_t* __restrict__ cell_start,
                                                     const uint32_t* __restrict__ cell_end, const float4* __restrict__ x,
                                                     const float4* __restrict__ rest, const StepConst& k, uint32_t V, uint32_t tid) {
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
__global__ void cache_neighbors_kernel(uint32_t* __restrict__ neighbors, const uint32_t* __restrict__ index,
                                       const uint32_t* __restrict__ cell_start,
                                       const uint32_t* __restrict__ cell_end, const float4* __restrict__ x,
                                       const float4* __restrict__ rest, StepConst k, uint32_t V) {
  cache_neighbors_body(neighbors, index, cell_start, cell_end, x, rest, k, V, blockIdx.x * blockDim.x + threadIdx.x);
}

// ---- predict + hash of a whole batch of garments: blockIdx.y = garment, one launch per stage ----
__global__ void batch_predict_hash_kernel(const HashBatchItem* __restrict__ items, uint32_t* __restrict__ key, uint32_t* __restrict__ val) {
  const HashBatchItem& it = items[blockIdx.y];
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= it.c.V) return;
  xrtdev::predict_item(it.c.x0, it.c.x, it.c.v, it.k, i);  // PredictPositions, then the hash of the predicted position
  particle_hash_body(it.c.x, key + it.seg, val + it.seg, it.k.cell_spacing, it.k.table_size, it.c.V, i);
}
__global__ void batch_cell_clear_kernel(const HashBatchItem* __restrict__ items) {
  const HashBatchItem& it = items[blockIdx.y];
  for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < (uint32_t)it.k.table_size; i += gridDim.x * blockDim.x)
    it.c.cell_start[i] = 0xFFFFFFFFu;
}
__global__ void batch_cell_range_kernel(const HashBatchItem* __restrict__ items, const uint32_t* __restrict__ key) {
  const HashBatchItem& it = items[blockIdx.y];
  cell_range_body(it.c.cell_start, it.c.cell_end, key + it.seg, it.c.V, blockIdx.x * blockDim.x + threadIdx.x);
}
__global__ void batch_cache_neighbors_kernel(const HashBatchItem* __restrict__ items, const uint32_t* __restrict__ val) {
  const HashBatchItem& it = items[blockIdx.y];
  cache_neighbors_body(it.c.neighbors, val + it.seg, it.c.cell_start, it.c.cell_end, it.c.x, it.c.rest, it.k, it.c.V,
                       blockIdx.x * blockDim.x + threadIdx.x);
}

// ComputeFriction + CollideParticles_Kernel, ClothSolverHelper.cu:150-211
__global__ void collide_particles_kernel(float4* __restrict__ delta, const float4* __restrict__ x,
                                         const float4* __restrict__ x0, const uint32_t* __restrict__ neighbors,
                                         StepConst k, uint32_t V) {
  uint32_t id = blockIdx.x * blockDim.x + threadIdx.x;
  if (id < V) xrtdev::collide_item(delta, x, x0, neighbors, k, V, id);
}

}  // namespace


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

int hash_batch_alloc(HashBatch* hb, const std::vector<HashBatchItem>& items) {
  hb->n = (uint32_t)items.size();
  hb->total = 0;
  hb->max_v = 0;
  hb->max_table = 0;
  hb->bits = 0;
  std::vector<HashBatchItem> h = items;
  std::vector<uint32_t> off(items.size() + 1, 0);
  for (size_t i = 0; i < h.size(); i++) {
    h[i].seg = hb->total;
    off[i] = hb->total;
    hb->total += h[i].c.V;
    hb->max_v = std::max(hb->max_v, h[i].c.V);
    hb->max_table = std::max(hb->max_table, (uint32_t)h[i].k.table_size);
    hb->bits = std::max(hb->bits, hash_max_bit(h[i].k.table_size));
  }
  off[h.size()] = hb->total;
  cudaError_t e = cudaSuccess;
  auto A = [&](void** p, size_t bytes) {
    if (e == cudaSuccess) e = cudaMalloc(p, bytes ? bytes : 16);
  };
  A((void**)&hb->items, h.size() * sizeof(HashBatchItem));
  A((void**)&hb->seg_off, off.size() * 4);
  A((void**)&hb->key_in, (size_t)hb->total * 4); A((void**)&hb->val_in, (size_t)hb->total * 4);
  A((void**)&hb->key_out, (size_t)hb->total * 4); A((void**)&hb->val_out, (size_t)hb->total * 4);
  if (e != cudaSuccess) return (int)e;
  cudaMemcpy(hb->items, h.data(), h.size() * sizeof(HashBatchItem), cudaMemcpyHostToDevice);
  cudaMemcpy(hb->seg_off, off.data(), off.size() * 4, cudaMemcpyHostToDevice);
  hb->cub_bytes = 0;
  e = cub::DeviceSegmentedRadixSort::SortPairs(nullptr, hb->cub_bytes, hb->key_in, hb->key_out, hb->val_in, hb->val_out, (int)hb->total,
                                               (int)hb->n, hb->seg_off, hb->seg_off + 1, 0, hb->bits);
  A(&hb->cub_tmp, hb->cub_bytes);
  return (int)e;
}
void hash_batch_free(HashBatch* hb) {
  cudaFree(hb->items); cudaFree(hb->seg_off); cudaFree(hb->key_in); cudaFree(hb->val_in); cudaFree(hb->key_out);
  cudaFree(hb->val_out); cudaFree(hb->cub_tmp);
  *hb = HashBatch{};
}
// PredictPositions + SpatialHashGPU::Hash of every garment of the batch in five launches
void launch_predict_hash_batched(const HashBatch& hb, cudaStream_t st) {
  if (!hb.n || !hb.total) return;
  const dim3 g(xrt_grid(hb.max_v), hb.n);
  batch_predict_hash_kernel<<<g, XRT_BLOCK, 0, st>>>(hb.items, hb.key_in, hb.val_in);
  size_t bytes = hb.cub_bytes;
  cub::DeviceSegmentedRadixSort::SortPairs(hb.cub_tmp, bytes, hb.key_in, hb.key_out, hb.val_in, hb.val_out, (int)hb.total, (int)hb.n,
                                           hb.seg_off, hb.seg_off + 1, 0, hb.bits, st);
  batch_cell_clear_kernel<<<dim3(std::min<unsigned>(xrt_grid(hb.max_table), 64u), hb.n), XRT_BLOCK, 0, st>>>(hb.items);
  batch_cell_range_kernel<<<g, XRT_BLOCK, 0, st>>>(hb.items, hb.key_out);
  batch_cache_neighbors_kernel<<<g, XRT_BLOCK, 0, st>>>(hb.items, hb.val_out);
}

void launch_collide_particles(const ClothDev& c, const StepConst& k, cudaStream_t st) {
  if (!c.V) return;
  collide_particles_kernel<<<xrt_grid(c.V), XRT_BLOCK, 0, st>>>(c.delta, c.x, c.x0, c.neighbors, k, c.V);
  launch_apply_deltas(c, k, st);
}
