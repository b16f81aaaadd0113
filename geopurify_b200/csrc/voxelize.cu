// F1: GPU voxeliser.  Replaces Voxelizer.voxelize + sparse_quantize of the reference
// (dataset/voxelizer.py:104-121; dataset/voxelization_utils.py:6-18 fnv_hash_vec, :78-96 np.unique on the keys).
//
//   c  = floor(M [x y z 1]^T)            float64, M = diag(1/voxel_size) (optionally rotated/scaled: M_r M_v)
//   v  = floor(c - min_over_points(c))   integer-valued, non-negative
//   key = FNV-1-style uint64 of v: h = 14695981039346656037; per axis h *= 1099511628211; h ^= v_axis
//   np.unique(key, return_index, return_inverse): voxels are numbered in ascending KEY order (i.e. hash-random in
//   space), the representative of a voxel is its FIRST point, inds_reconstruct maps points to voxels.
// Distinct cells whose keys collide are merged, exactly as the reference merges them (it only ever looks at the key).
#include <cub/device/device_radix_sort.cuh>
#include <cub/device/device_scan.cuh>
#include "common.cuh"

namespace {

struct VoxHeader {
  long long mn[3];
  long long mx[3];
  long long m;
};

__global__ void vox_init_kernel(VoxHeader* h) {
  for (int a = 0; a < 3; ++a) {
    h->mn[a] = 0x7fffffffffffffffll;
    h->mx[a] = -0x7fffffffffffffffll;
  }
  h->m = 0;
}

template <typename T>
__global__ void __launch_bounds__(256) vox_floor_kernel(const T* __restrict__ coords, int64_t n,
                                                        const double* __restrict__ xform, long long* __restrict__ cell,
                                                        VoxHeader* h) {
  __shared__ double M[12];
  if (threadIdx.x < 12) M[threadIdx.x] = xform[threadIdx.x];
  __syncthreads();
  long long mn[3] = {0x7fffffffffffffffll, 0x7fffffffffffffffll, 0x7fffffffffffffffll};
  long long mx[3] = {-0x7fffffffffffffffll, -0x7fffffffffffffffll, -0x7fffffffffffffffll};
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const double x = (double)coords[3 * i], y = (double)coords[3 * i + 1], z = (double)coords[3 * i + 2];
#pragma unroll
    for (int a = 0; a < 3; ++a) {
      // [x y z 1] @ M^T[:, a]: the same k = 0..3 FMA chain numpy's dgemm runs (see project.cu); with a diagonal M
      // every other term is an exact zero, so the result is the single product x * (1 / voxel_size).
      double r = x * M[4 * a];
      r = fma(y, M[4 * a + 1], r);
      r = fma(z, M[4 * a + 2], r);
      r = fma(1.0, M[4 * a + 3], r);
      const long long c = (long long)floor(r);
      cell[3 * i + a] = c;
      mn[a] = min(mn[a], c);
      mx[a] = max(mx[a], c);
    }
  }
#pragma unroll
  for (int a = 0; a < 3; ++a) {
    for (int o = 16; o > 0; o >>= 1) {
      mn[a] = min(mn[a], __shfl_xor_sync(0xffffffffu, mn[a], o));
      mx[a] = max(mx[a], __shfl_xor_sync(0xffffffffu, mx[a], o));
    }
    if ((threadIdx.x & 31) == 0) {
      atomicMin(&h->mn[a], mn[a]);
      atomicMax(&h->mx[a], mx[a]);
    }
  }
}

__global__ void vox_key_kernel(const long long* __restrict__ cell, int64_t n, const VoxHeader* __restrict__ h,
                               uint64_t* __restrict__ keys, int32_t* __restrict__ vals) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  uint64_t k = 14695981039346656037ull;
#pragma unroll
  for (int a = 0; a < 3; ++a) {
    k *= 1099511628211ull;
    k ^= (uint64_t)(cell[3 * i + a] - h->mn[a]);
  }
  keys[i] = k;
  vals[i] = (int32_t)i;
}

__global__ void vox_heads_kernel(const uint64_t* __restrict__ keys_sorted, int64_t n, int32_t* __restrict__ flags) {
  const int64_t s = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (s >= n) return;
  flags[s] = s > 0 && keys_sorted[s] != keys_sorted[s - 1];  // inclusive scan of this = voxel id of sorted entry s
}

__global__ void vox_emit_kernel(const uint64_t* __restrict__ keys_sorted, const int32_t* __restrict__ vals_sorted,
                                const int32_t* __restrict__ vid, int64_t n, const long long* __restrict__ cell,
                                VoxHeader* h, int32_t* __restrict__ vox, int32_t* __restrict__ inds,
                                int32_t* __restrict__ inds_reconstruct, int64_t* __restrict__ m_out,
                                int32_t* __restrict__ extent) {
  const int64_t s = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (s >= n) return;
  const int v = vid[s];
  const int p = vals_sorted[s];
  inds_reconstruct[p] = v;
  if (s == 0 || keys_sorted[s] != keys_sorted[s - 1]) {  // stable sort: the head of a run is the voxel's first point
    inds[v] = p;
#pragma unroll
    for (int a = 0; a < 3; ++a) vox[3 * (int64_t)v + a] = (int32_t)(cell[3 * (int64_t)p + a] - h->mn[a]);
  }
  if (s == n - 1) {
    *m_out = (int64_t)v + 1;
    for (int a = 0; a < 3; ++a) extent[a] = (int32_t)(h->mx[a] - h->mn[a] + 1);
  }
}

struct VoxLayout {
  size_t off_hdr, off_cell, off_kin, off_kout, off_vin, off_vout, off_flags, off_vid, off_cub, cub_bytes, total;
};

void vox_layout(int64_t n, VoxLayout& L) {
  size_t off = 0;
  auto take = [&](size_t bytes) {
    size_t o = off;
    off += gp_align_up(bytes);
    return o;
  };
  L.off_hdr = take(sizeof(VoxHeader));
  L.off_cell = take(sizeof(long long) * 3 * (size_t)n);
  L.off_kin = take(sizeof(uint64_t) * (size_t)n);
  L.off_kout = take(sizeof(uint64_t) * (size_t)n);
  L.off_vin = take(sizeof(int32_t) * (size_t)n);
  L.off_vout = take(sizeof(int32_t) * (size_t)n);
  L.off_flags = take(sizeof(int32_t) * (size_t)n);
  L.off_vid = take(sizeof(int32_t) * (size_t)n);
  size_t a = 0, b = 0;
  cub::DeviceRadixSort::SortPairs(nullptr, a, (const uint64_t*)nullptr, (uint64_t*)nullptr, (const int32_t*)nullptr,
                                  (int32_t*)nullptr, (int)n, 0, 64);
  cub::DeviceScan::InclusiveSum(nullptr, b, (const int32_t*)nullptr, (int32_t*)nullptr, (int)n);
  L.cub_bytes = (a > b ? a : b) + 256;
  L.off_cub = take(L.cub_bytes);
  L.total = off;
}

}  // namespace

extern "C" size_t gp_voxelize_workspace_bytes(int64_t n) {
  if (n < 0) return 0;
  VoxLayout L;
  vox_layout(n, L);
  return L.total;
}

extern "C" int gp_voxelize(const void* coords, int coords_dtype, int64_t n, const double* xform, int32_t* vox,
                           int32_t* inds, int32_t* inds_reconstruct, int64_t* m_out, int32_t* extent, void* workspace,
                           size_t workspace_bytes, gp_stream_t stream) {
  GP_CHECK_ARG(n >= 1 && n < (1ll << 31), "gp_voxelize: n=%lld outside [1, 2^31) (the reference asserts a non-empty cloud)",
               (long long)n);
  GP_CHECK_ARG(coords_dtype == GP_F32 || coords_dtype == GP_F64, "gp_voxelize: coords must be float32 or float64");
  GP_CHECK_ARG(xform != nullptr, "gp_voxelize: xform [3,4] float64 (device) is required");
  VoxLayout L;
  vox_layout(n, L);
  if (workspace_bytes < L.total) {
    gp_set_error("gp_voxelize: workspace %zu < %zu", workspace_bytes, L.total);
    return GP_ERR_WORKSPACE;
  }
  cudaStream_t st = (cudaStream_t)stream;
  char* ws = (char*)workspace;
  VoxHeader* h = (VoxHeader*)(ws + L.off_hdr);
  long long* cell = (long long*)(ws + L.off_cell);
  uint64_t* kin = (uint64_t*)(ws + L.off_kin);
  uint64_t* kout = (uint64_t*)(ws + L.off_kout);
  int32_t* vin = (int32_t*)(ws + L.off_vin);
  int32_t* vout = (int32_t*)(ws + L.off_vout);
  int32_t* flags = (int32_t*)(ws + L.off_flags);
  int32_t* vid = (int32_t*)(ws + L.off_vid);
  const unsigned grid = (unsigned)((n + 255) / 256);
  vox_init_kernel<<<1, 1, 0, st>>>(h);
  const unsigned g2 = grid < 4 * GP_SM_COUNT ? grid : 4 * GP_SM_COUNT;
  if (coords_dtype == GP_F64) vox_floor_kernel<double><<<g2, 256, 0, st>>>((const double*)coords, n, xform, cell, h);
  else vox_floor_kernel<float><<<g2, 256, 0, st>>>((const float*)coords, n, xform, cell, h);
  GP_LAUNCH_CHECK();
  vox_key_kernel<<<grid, 256, 0, st>>>(cell, n, h, kin, vin);
  GP_LAUNCH_CHECK();
  size_t cb = L.cub_bytes;
  GP_CUDA(cub::DeviceRadixSort::SortPairs(ws + L.off_cub, cb, kin, kout, vin, vout, (int)n, 0, 64, st));
  vox_heads_kernel<<<grid, 256, 0, st>>>(kout, n, flags);
  GP_LAUNCH_CHECK();
  cb = L.cub_bytes;
  GP_CUDA(cub::DeviceScan::InclusiveSum(ws + L.off_cub, cb, flags, vid, (int)n, st));
  vox_emit_kernel<<<grid, 256, 0, st>>>(kout, vout, vid, n, cell, h, vox, inds, inds_reconstruct, m_out, extent);
  GP_LAUNCH_CHECK();
  return GP_OK;
}
