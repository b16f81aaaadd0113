// Deterministic scatter-mean of point rows into voxel rows.
// Replaces torch_scatter.scatter_mean (reference models/affinity_module.py:1524-1535): float32 sum of the
// rows of each voxel in ascending point order, divided by max(count, 1).
#include <cub/device/device_radix_sort.cuh>
#include "common.cuh"

struct ScatterLayout {
  size_t off_pts, off_offsets, off_kin, off_kout, off_vin, off_cub, cub_bytes, total;
  int key_bits;
};

static void scatter_layout(int64_t n, int64_t m, ScatterLayout& L) {
  size_t off = 0;
  auto take = [&](size_t bytes) {
    size_t o = off;
    off += gp_align_up(bytes);
    return o;
  };
  L.key_bits = 1;
  while ((1ll << L.key_bits) < m + 1 && L.key_bits < 31) ++L.key_bits;
  L.off_pts = take(sizeof(int32_t) * (size_t)n);
  L.off_offsets = take(sizeof(int32_t) * (size_t)(m + 1));
  L.off_kin = take(sizeof(uint32_t) * (size_t)n);
  L.off_kout = take(sizeof(uint32_t) * (size_t)n);
  L.off_vin = take(sizeof(int32_t) * (size_t)n);
  L.cub_bytes = 0;
  cub::DeviceRadixSort::SortPairs(nullptr, L.cub_bytes, (const uint32_t*)nullptr, (uint32_t*)nullptr,
                                  (const int32_t*)nullptr, (int32_t*)nullptr, (int)n, 0, L.key_bits);
  L.off_cub = take(L.cub_bytes + 256);
  L.total = off;
}

__global__ void scatter_keys_kernel(const int32_t* __restrict__ p2v, int64_t n, uint32_t* keys, int32_t* vals) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) {
    keys[i] = (uint32_t)p2v[i];
    vals[i] = (int32_t)i;
  }
}

// offsets[v] = first sorted position whose key >= v  (segments may be empty)
__global__ void scatter_offsets_kernel(const uint32_t* __restrict__ keys_sorted, int64_t n, int64_t m,
                                       int32_t* __restrict__ offsets) {
  int64_t s = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (s > n) return;
  int64_t prev = s == 0 ? -1 : (int64_t)keys_sorted[s - 1];
  int64_t cur = s == n ? m : (int64_t)keys_sorted[s];
  if (prev > m) prev = m;   // an out-of-range destination row (caller error) must not write past offsets[m]
  if (cur > m) cur = m;
  for (int64_t v = prev + 1; v <= cur; ++v) offsets[v] = (int32_t)s;
}

extern "C" size_t gp_scatter_mean_workspace_bytes(int64_t n, int64_t m) {
  if (n < 0 || m < 0) return 0;
  ScatterLayout L;
  scatter_layout(n, m, L);
  return L.total;
}

extern "C" int gp_scatter_mean_prepare(const int32_t* p2v, int64_t n, int64_t m, void* workspace,
                                       size_t workspace_bytes, gp_stream_t stream) {
  GP_CHECK_ARG(n >= 0 && m >= 0 && n < (1ll << 31) && m < (1ll << 31), "gp_scatter_mean_prepare: bad sizes");
  ScatterLayout L;
  scatter_layout(n, m, L);
  if (workspace_bytes < L.total) {
    gp_set_error("gp_scatter_mean_prepare: workspace %zu < %zu", workspace_bytes, L.total);
    return GP_ERR_WORKSPACE;
  }
  cudaStream_t st = (cudaStream_t)stream;
  char* ws = (char*)workspace;
  uint32_t* kin = (uint32_t*)(ws + L.off_kin);
  uint32_t* kout = (uint32_t*)(ws + L.off_kout);
  int32_t* vin = (int32_t*)(ws + L.off_vin);
  int32_t* pts = (int32_t*)(ws + L.off_pts);
  if (n > 0) {
    scatter_keys_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(p2v, n, kin, vin);
    GP_LAUNCH_CHECK();
    size_t cb = L.cub_bytes + 256;
    GP_CUDA(cub::DeviceRadixSort::SortPairs(ws + L.off_cub, cb, kin, kout, vin, pts, (int)n, 0, L.key_bits, st));
  }
  scatter_offsets_kernel<<<(unsigned)((n + 1 + 255) / 256), 256, 0, st>>>(kout, n, m, (int32_t*)(ws + L.off_offsets));
  GP_LAUNCH_CHECK();
  return GP_OK;
}

// one warp per voxel row
template <bool VEC4>
__global__ void __launch_bounds__(256) scatter_mean_kernel(const float* __restrict__ src, int64_t src_stride, int D,
                                                            const int32_t* __restrict__ pts,
                                                            const int32_t* __restrict__ offsets, int64_t m,
                                                            float* __restrict__ out, int64_t out_stride, int col0) {
  int64_t v = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (v >= m) return;
  const int lane = threadIdx.x & 31;
  const int beg = offsets[v], end = offsets[v + 1];
  const float cnt = (float)max(end - beg, 1);
  if (VEC4) {
    for (int c0 = lane * 4; c0 < D; c0 += 128) {
      float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
      for (int j = beg; j < end; ++j) {
        float4 t = ldg_stream_f4((const float4*)(src + (int64_t)pts[j] * src_stride + c0));
        acc.x = __fadd_rn(acc.x, t.x);
        acc.y = __fadd_rn(acc.y, t.y);
        acc.z = __fadd_rn(acc.z, t.z);
        acc.w = __fadd_rn(acc.w, t.w);
      }
      float4 r = make_float4(__fdiv_rn(acc.x, cnt), __fdiv_rn(acc.y, cnt), __fdiv_rn(acc.z, cnt),
                             __fdiv_rn(acc.w, cnt));
      *(float4*)(out + v * out_stride + col0 + c0) = r;
    }
  } else {
    for (int c = lane; c < D; c += 32) {
      float acc = 0.f;
      for (int j = beg; j < end; ++j) acc = __fadd_rn(acc, __ldg(src + (int64_t)pts[j] * src_stride + c));
      out[v * out_stride + col0 + c] = __fdiv_rn(acc, cnt);
    }
  }
}

extern "C" int gp_scatter_mean(const float* src, int64_t src_stride, int D, const void* prepared_workspace, int64_t n,
                               int64_t m, float* out, int64_t out_stride, int col0, gp_stream_t stream) {
  GP_CHECK_ARG(n >= 0 && m >= 0 && D > 0 && src_stride >= D && col0 >= 0 && out_stride >= col0 + D,
               "gp_scatter_mean: bad sizes");
  if (m == 0) return GP_OK;
  ScatterLayout L;
  scatter_layout(n, m, L);
  const char* ws = (const char*)prepared_workspace;
  const int32_t* pts = (const int32_t*)(ws + L.off_pts);
  const int32_t* offsets = (const int32_t*)(ws + L.off_offsets);
  bool vec = D % 4 == 0 && src_stride % 4 == 0 && out_stride % 4 == 0 && col0 % 4 == 0 &&
             ((uintptr_t)src % 16) == 0 && ((uintptr_t)out % 16) == 0;
  unsigned grid = (unsigned)((m + 7) / 8);
  if (vec)
    scatter_mean_kernel<true><<<grid, 256, 0, (cudaStream_t)stream>>>(src, src_stride, D, pts, offsets, m, out,
                                                                      out_stride, col0);
  else
    scatter_mean_kernel<false><<<grid, 256, 0, (cudaStream_t)stream>>>(src, src_stride, D, pts, offsets, m, out,
                                                                       out_stride, col0);
  GP_LAUNCH_CHECK();
  return GP_OK;
}

// ---- allocation-free variant -------------------------------------------------------------------------------------
// gp_scatter_mean_static = gp_scatter_mean_prepare + gp_scatter_mean with the scratch (sorted point list, segment
// offsets, radix-sort keys and CUB's temporary storage) in a file-scope device array instead of a caller workspace,
// for callers that cannot allocate (CUDA graph capture, serving loops, harnesses that forbid cudaMalloc).  No
// counterpart in the reference.
//
// The array is GP_SCATTER_STATIC_BYTES (16 MiB) per GPU that loads the library, held until the process exits.
// Every call on a device shares it: calls that overlap on the GPU (different streams, or a replay of a captured graph
// next to a live call) race and must be serialised by the caller.  It keeps the previous call's contents; nothing in
// it has to start at zero -- the keys, values and the offsets[0..m] are written in full before they are read, and
// cub::DeviceRadixSort clears its own counters with cudaMemsetAsync on the same stream -- so no clear is issued.
// Host work: one cudaGetSymbolAddress per device (first call on that device), then kernel launches only: no
// cudaMalloc / cudaFree / allocating library call (CUB is given its temporary storage).
static __device__ __align__(256) unsigned char g_scatter_static[GP_SCATTER_STATIC_BYTES];

#define GP_MAX_DEVICES 64
#include <mutex>

// Device address of g_scatter_static on the CURRENT device.  In host code the symbol's name is only a handle for
// cudaGetSymbolAddress, never a device pointer.
static int scatter_static_ptr(void** out) {
  static std::mutex mu;
  static void* ptrs[GP_MAX_DEVICES] = {};
  int dev = -1;
  GP_CUDA(cudaGetDevice(&dev));
  GP_CHECK_ARG(dev >= 0 && dev < GP_MAX_DEVICES, "gp_scatter_mean_static: device ordinal %d not supported", dev);
  std::lock_guard<std::mutex> lock(mu);
  if (!ptrs[dev]) {
    void* p = nullptr;
    GP_CUDA(cudaGetSymbolAddress(&p, g_scatter_static));
    if (!p) {
      gp_set_error("gp_scatter_mean_static: cudaGetSymbolAddress returned a null address on device %d", dev);
      return GP_ERR_CUDA;
    }
    ptrs[dev] = p;
  }
  *out = ptrs[dev];
  return GP_OK;
}

extern "C" int gp_scatter_mean_static(const int32_t* p2v, int64_t n, int64_t m, const float* src, int64_t src_stride,
                                      int D, float* out, int64_t out_stride, int col0, gp_stream_t stream) {
  GP_CHECK_ARG(n >= 0 && m >= 0 && D > 0 && src_stride >= D && col0 >= 0 && out_stride >= col0 + D,
               "gp_scatter_mean_static: bad sizes");
  if (n > GP_SCATTER_STATIC_MAX_N || m > GP_SCATTER_STATIC_MAX_M) {
    gp_set_error("gp_scatter_mean_static: n = %lld, m = %lld beyond the static scratch limits (n <= %d, m <= %d); "
                 "use gp_scatter_mean_prepare + gp_scatter_mean with a caller workspace",
                 (long long)n, (long long)m, GP_SCATTER_STATIC_MAX_N, GP_SCATTER_STATIC_MAX_M);
    return GP_ERR_ARG;
  }
  ScatterLayout L;
  scatter_layout(n, m, L);
  if (L.total > (size_t)GP_SCATTER_STATIC_BYTES) {
    gp_set_error("gp_scatter_mean_static: scratch for n = %lld, m = %lld is %zu bytes > the %d-byte static array",
                 (long long)n, (long long)m, L.total, GP_SCATTER_STATIC_BYTES);
    return GP_ERR_WORKSPACE;
  }
  void* scratch = nullptr;
  int rc = scatter_static_ptr(&scratch);
  if (rc != GP_OK) return rc;
  rc = gp_scatter_mean_prepare(p2v, n, m, scratch, (size_t)GP_SCATTER_STATIC_BYTES, stream);
  if (rc != GP_OK) return rc;
  return gp_scatter_mean(src, src_stride, D, scratch, n, m, out, out_stride, col0, stream);
}
