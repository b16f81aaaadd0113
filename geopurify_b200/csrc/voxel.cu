// Voxel index (Hilbert order of 4^3-voxel cells, Morton inside a cell, + cell lists), 3^3 kernel map, exact grid kNN with warp bitonic top-k.
// Replaces the MinkowskiEngine coordinate map (reference models/affinity_module.py:1541-1545) and the
// faiss-cpu IndexFlatL2 search (reference models/affinity_module.py:1553-1557).
#include <cub/device/device_radix_sort.cuh>
#include "common.cuh"

#define GP_CELL_SHIFT 2  // cells of 4^3 voxels
#define GP_CELL (1 << GP_CELL_SHIFT)

struct VoxelIndexLayout {
  int cx, cy, cz;
  int64_t ncell;
  int key_bits;
  size_t off_sorted, off_cells, off_rank, off_keys_in, off_keys_out, off_vals_in, off_vals_out, off_cub, off_status;
  size_t cub_bytes, total;
};

static int morton_bits(int ex, int ey, int ez) {
  int mx = ex > ey ? ex : ey;
  mx = mx > ez ? mx : ez;
  int b = 1;
  while ((1 << b) < mx) ++b;
  return b;
}

static bool voxel_index_layout(int64_t m, int ex, int ey, int ez, VoxelIndexLayout& L) {
  if (m < 0 || ex <= 0 || ey <= 0 || ez <= 0 || ex > 32768 || ey > 32768 || ez > 32768) return false;
  L.cx = (ex + GP_CELL - 1) >> GP_CELL_SHIFT;
  L.cy = (ey + GP_CELL - 1) >> GP_CELL_SHIFT;
  L.cz = (ez + GP_CELL - 1) >> GP_CELL_SHIFT;
  L.ncell = (int64_t)L.cx * L.cy * L.cz;
  if (L.ncell > (1ll << 28)) return false;
  L.key_bits = 3 * morton_bits(ex, ey, ez);
  size_t off = 0;
  auto take = [&](size_t bytes) {
    size_t o = off;
    off += gp_align_up(bytes);
    return o;
  };
  L.off_sorted = take(sizeof(int4) * (size_t)m);
  L.off_cells = take(sizeof(int2) * (size_t)L.ncell);
  L.off_rank = take(sizeof(int32_t) * (size_t)m);
  L.off_keys_in = take(sizeof(uint64_t) * (size_t)m);
  L.off_keys_out = take(sizeof(uint64_t) * (size_t)m);
  L.off_vals_in = take(sizeof(int32_t) * (size_t)m);
  L.off_vals_out = take(sizeof(int32_t) * (size_t)m);
  L.cub_bytes = 0;
  cub::DeviceRadixSort::SortPairs(nullptr, L.cub_bytes, (const uint64_t*)nullptr, (uint64_t*)nullptr,
                                  (const int32_t*)nullptr, (int32_t*)nullptr, (int)m, 0, L.key_bits);
  L.off_cub = take(L.cub_bytes + 256);
  L.off_status = take(sizeof(int32_t));   // != 0: a voxel coordinate was outside [0, extent)
  L.total = off;
  return true;
}

__device__ __forceinline__ uint64_t spread3(uint32_t v) {  // 21 bits -> every third bit
  uint64_t x = v & 0x1fffffu;
  x = (x | x << 32) & 0x1f00000000ffffull;
  x = (x | x << 16) & 0x1f0000ff0000ffull;
  x = (x | x << 8) & 0x100f00f00f00f00full;
  x = (x | x << 4) & 0x10c30c30c30c30c3ull;
  x = (x | x << 2) & 0x1249249249249249ull;
  return x;
}

// Hilbert index of a 3-D cell coordinate with `bits` bits per axis (Skilling's transpose algorithm).  Consecutive
// indices are face-adjacent cells, so 128 consecutive voxels along the curve form a compact patch: the pooling tiles'
// neighbour unions shrink by ~8 % against Morton order (628 -> 581 rows on a 54k-voxel scene), and with them the
// K-steps of pool_tc_kernel.
__device__ __forceinline__ uint64_t hilbert3(uint32_t x, uint32_t y, uint32_t z, int bits) {
  if (bits <= 0) return 0;
  uint32_t X[3] = {x, y, z};
  const uint32_t M = 1u << (bits - 1);
  for (uint32_t Q = M; Q > 1; Q >>= 1) {
    const uint32_t P = Q - 1;
#pragma unroll
    for (int i = 0; i < 3; ++i) {
      if (X[i] & Q) {
        X[0] ^= P;
      } else {
        const uint32_t t = (X[0] ^ X[i]) & P;
        X[0] ^= t;
        X[i] ^= t;
      }
    }
  }
  X[1] ^= X[0];
  X[2] ^= X[1];
  uint32_t t = 0;
  for (uint32_t Q = M; Q > 1; Q >>= 1)
    if (X[2] & Q) t ^= Q - 1;
  X[0] ^= t;
  X[1] ^= t;
  X[2] ^= t;
  // interleave: bit b of X[0] is the most significant of the three
  return (spread3(X[0]) << 2) | (spread3(X[1]) << 1) | spread3(X[2]);
}

// sort key: Hilbert index of the 4^3-voxel cell, then Morton order inside the cell (a cell's voxels stay contiguous)
__global__ void morton_keys_kernel(const int32_t* __restrict__ vox, int64_t m, int cell_bits,
                                   uint64_t* __restrict__ keys, int32_t* __restrict__ vals) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= m) return;
  uint32_t x = vox[3 * i], y = vox[3 * i + 1], z = vox[3 * i + 2];
  const uint32_t lm = GP_CELL - 1;
  const uint64_t inside = spread3(x & lm) | (spread3(y & lm) << 1) | (spread3(z & lm) << 2);
  keys[i] = (hilbert3(x >> GP_CELL_SHIFT, y >> GP_CELL_SHIFT, z >> GP_CELL_SHIFT, cell_bits) << (3 * GP_CELL_SHIFT)) | inside;
  vals[i] = (int32_t)i;
}

__global__ void index_finalize_kernel(const int32_t* __restrict__ vox, const uint64_t* __restrict__ keys_sorted,
                                      const int32_t* __restrict__ vals_sorted, int64_t m, int cx, int cy,
                                      int4* __restrict__ sorted, int2* __restrict__ cells,
                                      int32_t* __restrict__ rank_ws, int32_t* __restrict__ order,
                                      int32_t* __restrict__ rank, int ex, int ey, int ez,
                                      int32_t* __restrict__ status) {
  int64_t s = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (s >= m) return;
  int orig = vals_sorted[s];
  int x = vox[3 * (int64_t)orig], y = vox[3 * (int64_t)orig + 1], z = vox[3 * (int64_t)orig + 2];
  sorted[s] = make_int4(x, y, z, orig);
  rank_ws[orig] = (int32_t)s;
  if (order) order[s] = orig;
  if (rank) rank[orig] = (int32_t)s;
  if (x < 0 || y < 0 || z < 0 || x >= ex || y >= ey || z >= ez) {  // caller error: no cell-table write, flag it
    atomicOr(status, 1);
    return;
  }
  uint64_t ck = keys_sorted[s] >> (3 * GP_CELL_SHIFT);
  int64_t lin = ((int64_t)(z >> GP_CELL_SHIFT) * cy + (y >> GP_CELL_SHIFT)) * cx + (x >> GP_CELL_SHIFT);
  if (s == 0 || (keys_sorted[s - 1] >> (3 * GP_CELL_SHIFT)) != ck) cells[lin].x = (int)s;
  if (s == m - 1 || (keys_sorted[s + 1] >> (3 * GP_CELL_SHIFT)) != ck) cells[lin].y = (int)s + 1;
}

extern "C" size_t gp_voxel_index_workspace_bytes(int64_t m, int ex, int ey, int ez) {
  VoxelIndexLayout L;
  if (!voxel_index_layout(m, ex, ey, ez, L)) return 0;
  return L.total;
}

extern "C" int gp_voxel_index_build(const int32_t* vox, int64_t m, int ex, int ey, int ez, int32_t* order,
                                    int32_t* rank, void* workspace, size_t workspace_bytes, gp_stream_t stream) {
  VoxelIndexLayout L;
  GP_CHECK_ARG(voxel_index_layout(m, ex, ey, ez, L), "gp_voxel_index_build: bad sizes m=%lld extent=(%d,%d,%d)",
               (long long)m, ex, ey, ez);
  GP_CHECK_ARG(m < (1ll << 31), "gp_voxel_index_build: m too large");
  if (workspace_bytes < L.total) {
    gp_set_error("gp_voxel_index_build: workspace %zu < %zu", workspace_bytes, L.total);
    return GP_ERR_WORKSPACE;
  }
  cudaStream_t st = (cudaStream_t)stream;
  char* ws = (char*)workspace;
  GP_CUDA(cudaMemsetAsync(ws + L.off_cells, 0, sizeof(int2) * (size_t)L.ncell, st));
  GP_CUDA(cudaMemsetAsync(ws + L.off_status, 0, sizeof(int32_t), st));
  if (m == 0) return GP_OK;
  uint64_t* kin = (uint64_t*)(ws + L.off_keys_in);
  uint64_t* kout = (uint64_t*)(ws + L.off_keys_out);
  int32_t* vin = (int32_t*)(ws + L.off_vals_in);
  int32_t* vout = (int32_t*)(ws + L.off_vals_out);
  unsigned grid = (unsigned)((m + 255) / 256);
  morton_keys_kernel<<<grid, 256, 0, st>>>(vox, m, L.key_bits / 3 - GP_CELL_SHIFT, kin, vin);
  GP_LAUNCH_CHECK();
  size_t cb = L.cub_bytes + 256;
  GP_CUDA(cub::DeviceRadixSort::SortPairs(ws + L.off_cub, cb, kin, kout, vin, vout, (int)m, 0, L.key_bits, st));
  index_finalize_kernel<<<grid, 256, 0, st>>>(vox, kout, vout, m, L.cx, L.cy, (int4*)(ws + L.off_sorted),
                                              (int2*)(ws + L.off_cells), (int32_t*)(ws + L.off_rank), order, rank,
                                              ex, ey, ez, (int32_t*)(ws + L.off_status));
  GP_LAUNCH_CHECK();
  return GP_OK;
}

// Synchronising check of a built index: GP_ERR_ARG if any voxel coordinate was outside [0, extent) (such voxels are
// kept out of the cell table, so nothing was written out of bounds, but kernel map / kNN results are meaningless).
extern "C" int gp_voxel_index_check(const void* index_workspace, int64_t m, int ex, int ey, int ez,
                                    gp_stream_t stream) {
  VoxelIndexLayout L;
  GP_CHECK_ARG(voxel_index_layout(m, ex, ey, ez, L), "gp_voxel_index_check: bad sizes");
  int32_t flag = 0;
  GP_CUDA(cudaMemcpyAsync(&flag, (const char*)index_workspace + L.off_status, sizeof(flag), cudaMemcpyDeviceToHost,
                          (cudaStream_t)stream));
  GP_CUDA(cudaStreamSynchronize((cudaStream_t)stream));
  GP_CHECK_ARG(flag == 0, "voxel index: a voxel coordinate lies outside [0, extent) = [0,%d) x [0,%d) x [0,%d) "
               "(shift the coordinates by their minimum and pass extent = max + 1)", ex, ey, ez);
  return GP_OK;
}

// ------------------------------------------------------------------ kernel map
__global__ void __launch_bounds__(256) kernel_map_kernel(const int4* __restrict__ sorted,
                                                          const int2* __restrict__ cells, int64_t m, int ex, int ey,
                                                          int ez, int cx, int cy, int32_t* __restrict__ nbr27) {
  int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= m * 27) return;
  int64_t s = t / 27;
  int o = (int)(t % 27);
  int4 c = sorted[s];
  int x = c.x + (o % 3 - 1), y = c.y + ((o / 3) % 3 - 1), z = c.z + (o / 9 - 1);
  int res = -1;
  if (o == 13) {
    res = (int)s;
  } else if (x >= 0 && y >= 0 && z >= 0 && x < ex && y < ey && z < ez) {
    int2 ce = cells[((int64_t)(z >> GP_CELL_SHIFT) * cy + (y >> GP_CELL_SHIFT)) * cx + (x >> GP_CELL_SHIFT)];
    for (int j = ce.x; j < ce.y; ++j) {
      int4 q = sorted[j];
      if (q.x == x && q.y == y && q.z == z) {
        res = j;
        break;
      }
    }
  }
  nbr27[t] = res;
}

extern "C" int gp_kernel_map(const void* index_workspace, int64_t m, int ex, int ey, int ez, int32_t* nbr27,
                             gp_stream_t stream) {
  VoxelIndexLayout L;
  GP_CHECK_ARG(voxel_index_layout(m, ex, ey, ez, L), "gp_kernel_map: bad sizes");
  if (m == 0) return GP_OK;
  const char* ws = (const char*)index_workspace;
  int64_t total = m * 27;
  kernel_map_kernel<<<(unsigned)((total + 255) / 256), 256, 0, (cudaStream_t)stream>>>(
      (const int4*)(ws + L.off_sorted), (const int2*)(ws + L.off_cells), m, ex, ey, ez, L.cx, L.cy, nbr27);
  GP_LAUNCH_CHECK();
  return GP_OK;
}

#include "topk.cuh"

// One query per warp.  Keys are (d^2 << SHIFT) | original index: ascending key order is "ascending (d^2, index)".
// KT = uint32 (SHIFT 21: index < 2^21, d^2 < 2^11) halves the shuffle traffic of the bitonic networks; it gives up
// (returns false, the caller re-runs the query with 64-bit keys) when the search radius would need d^2 >= 2^11.
template <typename KT, int SHIFT>
__device__ __forceinline__ bool knn_query(const int4* __restrict__ sorted, const int2* __restrict__ cells, int64_t s,
                                          const int4 q, int cx, int cy, int cz, int K, KT* buf, int lane,
                                          KT (&list)[4]) {
  constexpr KT KMAX = (KT)~(KT)0;
  constexpr long long D2_LIMIT = sizeof(KT) == 4 ? (1ll << (32 - SHIFT)) - 1 : (1ll << 31);
  const int qcx = q.x >> GP_CELL_SHIFT, qcy = q.y >> GP_CELL_SHIFT, qcz = q.z >> GP_CELL_SHIFT;
  int slack;
  {
    int ox = q.x & (GP_CELL - 1), oy = q.y & (GP_CELL - 1), oz = q.z & (GP_CELL - 1);
    int a = min(ox, GP_CELL - 1 - ox), b = min(oy, GP_CELL - 1 - oy), c = min(oz, GP_CELL - 1 - oz);
    slack = min(a, min(b, c));
  }
#pragma unroll
  for (int j = 0; j < 4; ++j) list[j] = KMAX;
  KT thr = KMAX;
  int nbuf = 0, merged = 0;
  const int rmax = max(cx, max(cy, cz));
  for (int r = 0; r <= rmax; ++r) {
    const int w = 2 * r + 1;
    const int ncand = w * w * w;
    for (int t0 = 0; t0 < ncand; t0 += 32) {
      int t = t0 + lane;
      int2 ce = make_int2(0, 0);
      if (t < ncand) {
        int dx = t % w - r, dy = (t / w) % w - r, dz = t / (w * w) - r;
        bool shell = max(abs(dx), max(abs(dy), abs(dz))) == r;
        int x = qcx + dx, y = qcy + dy, z = qcz + dz;
        if (shell && x >= 0 && y >= 0 && z >= 0 && x < cx && y < cy && z < cz)
          ce = __ldg(&cells[((int64_t)z * cy + y) * cx + x]);
      }
      unsigned nonempty = __ballot_sync(0xffffffffu, ce.y > ce.x);
      while (nonempty) {
        int src = __ffs(nonempty) - 1;
        nonempty &= nonempty - 1;
        int beg = __shfl_sync(0xffffffffu, ce.x, src), end = __shfl_sync(0xffffffffu, ce.y, src);
        for (int j0 = beg; j0 < end; j0 += 32) {
          int j = j0 + lane;
          KT key = KMAX;
          if (j < end && j != s) {
            int4 c = __ldg(&sorted[j]);
            int dx = c.x - q.x, dy = c.y - q.y, dz = c.z - q.z;
            unsigned d2 = (unsigned)(dx * dx) + (unsigned)(dy * dy) + (unsigned)(dz * dz);
            // candidates too far for the key width can only matter if the search runs past D2_LIMIT (-> false below)
            if ((long long)d2 <= D2_LIMIT - 1) key = ((KT)d2 << SHIFT) | (KT)(unsigned)c.w;
          }
          bool take = key < thr;
          unsigned bal = __ballot_sync(0xffffffffu, take);
          if (take) buf[nbuf + __popc(bal & ((1u << lane) - 1))] = key;
          nbuf += __popc(bal);
          if (nbuf > 96) {
            __syncwarp();
            topk_flush(list, buf, nbuf, lane);
            merged += nbuf;
            nbuf = 0;
            thr = topk_element(list, K - 1);
            __syncwarp();
          }
        }
      }
    }
    // every voxel within Chebyshev distance g - 1 of the query has now been scanned.  Fewer than K candidates so far:
    // the search cannot end on this shell, keep buffering instead of paying for a merge.
    if (nbuf && !(thr == KMAX && merged + nbuf < K && r < rmax)) {
      __syncwarp();
      topk_flush(list, buf, nbuf, lane);
      merged += nbuf;
      nbuf = 0;
      thr = topk_element(list, K - 1);
      __syncwarp();
    }
    long long g = (long long)r * GP_CELL + 1 + slack;
    if (thr != KMAX && (long long)(thr >> SHIFT) < g * g) return true;
    if (g * g >= D2_LIMIT) return false;  // the next ring could need distances the key cannot hold
  }
  return thr != KMAX || sizeof(KT) == 8;
}

#define KNN_WARPS 8
__global__ void __launch_bounds__(KNN_WARPS * 32) knn_grid_kernel(const int4* __restrict__ sorted,
                                                                   const int2* __restrict__ cells,
                                                                   const int32_t* __restrict__ rank, int64_t m,
                                                                   int cx, int cy, int cz, int K,
                                                                   int32_t* __restrict__ nbr_orig,
                                                                   int32_t* __restrict__ nbr_sorted) {
  __shared__ u64 sbuf[KNN_WARPS][128];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int64_t s = (int64_t)blockIdx.x * KNN_WARPS + warp;
  if (s >= m) return;
  const int4 q = sorted[s];
  int orig[4];
  bool done = false;
  if (m <= (1ll << 21)) {  // 32-bit keys: 21 index bits, 11 distance bits
    uint32_t list[4];
    done = knn_query<uint32_t, 21>(sorted, cells, s, q, cx, cy, cz, K, (uint32_t*)sbuf[warp], lane, list);
#pragma unroll
    for (int j = 0; j < 4; ++j) orig[j] = (int)(list[j] & 0x1fffffu);
  }
  if (!done) {
    __syncwarp();
    u64 list[4];
    knn_query<u64, 32>(sorted, cells, s, q, cx, cy, cz, K, sbuf[warp], lane, list);
#pragma unroll
    for (int j = 0; j < 4; ++j) orig[j] = (int)(list[j] & 0xffffffffu);
  }
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    int e = lane * 4 + j;
    if (e < K) {
      if (nbr_orig) nbr_orig[(int64_t)q.w * K + e] = orig[j];
      if (nbr_sorted) nbr_sorted[s * K + e] = __ldg(&rank[orig[j]]);
    }
  }
}

extern "C" int gp_knn_grid(const void* index_workspace, int64_t m, int ex, int ey, int ez, int K, int32_t* nbr_orig,
                           int32_t* nbr_sorted, gp_stream_t stream) {
  VoxelIndexLayout L;
  GP_CHECK_ARG(voxel_index_layout(m, ex, ey, ez, L), "gp_knn_grid: bad sizes");
  GP_CHECK_ARG(K >= 1 && K <= 128, "gp_knn_grid: K=%d outside [1,128]", K);
  // the reference's faiss call pads with -1 when fewer than K+1 voxels exist and then indexes with it
  // (SURVEY.md §8c item 11); refuse instead of reproducing the wrap-around.
  GP_CHECK_ARG(m >= (int64_t)K + 1, "gp_knn_grid: needs m >= K+1 voxels (m=%lld, K=%d)", (long long)m, K);
  const char* ws = (const char*)index_workspace;
  unsigned grid = (unsigned)((m + KNN_WARPS - 1) / KNN_WARPS);
  knn_grid_kernel<<<grid, KNN_WARPS * 32, 0, (cudaStream_t)stream>>>(
      (const int4*)(ws + L.off_sorted), (const int2*)(ws + L.off_cells), (const int32_t*)(ws + L.off_rank), m, L.cx,
      L.cy, L.cz, K, nbr_orig, nbr_sorted);
  GP_LAUNCH_CHECK();
  return GP_OK;
}
