// F4 (training-side sampling): exact K nearest neighbours on RAW float points.
// Replaces the faiss IndexFlatL2 search of the training step (models/affinity_module.py:1159-1165: index.add(coords),
// index.search(coords, K + 1), first column dropped) for the hybrid contrastive sampler.
//
// Contract: neighbours of point i in ascending (d2, index) order with i itself excluded, where
//   d2 = fl32(fl32(fl32(dx*dx) + fl32(dy*dy)) + fl32(dz*dz)),  dx = fl32(x_j - x_i)        (each operation rounded once)
// -- the float32 arithmetic of a direct squared-L2 evaluation, so the oracle (numpy float32, same order) is matched bit
// for bit.  faiss itself switches between this form and a BLAS ||x||^2 + ||y||^2 - 2<x,y> form by batch size; the two
// agree except where the K-th and (K+1)-th distances differ by float rounding.
//
// Uniform grid over the bounding box (<= 128 cells per axis), points sorted by cell, one warp per query walking
// Chebyshev shells of cells; candidates are kept in the warp-distributed bitonic top-k list of topk.cuh with keys
// (float bits of d2 << 32 | index) -- non-negative floats order like their bit patterns.
#include <cub/device/device_radix_sort.cuh>
#include "common.cuh"
#define GP_TOPK_SMALL  // the one-key-per-lane path for small batches pays off here (single key type; in voxel.cu,
                        // with two key widths inlined, the extra code costs more than it saves: 0.67 vs 0.58 ms)
#include "topk.cuh"

namespace {

constexpr int KP_DIM = 128;
constexpr int KP_NCELLS = KP_DIM * KP_DIM * KP_DIM;
constexpr int KP_WARPS = 8;

struct KpHeader {
  float mn[3];
  float cell, inv_cell;
  int dim[3];
  unsigned mn_bits[3], mx_bits[3];
};

struct KpLayout {
  size_t off_hdr, off_cells, off_sorted, off_kin, off_kout, off_vin, off_vout, off_cub, cub_bytes, total;
};

void kp_layout(int64_t n, KpLayout& L) {
  size_t off = 0;
  auto take = [&](size_t bytes) {
    size_t o = off;
    off += gp_align_up(bytes);
    return o;
  };
  L.off_hdr = take(sizeof(KpHeader));
  L.off_cells = take(sizeof(int2) * (size_t)KP_NCELLS);
  L.off_sorted = take(sizeof(float4) * (size_t)n);
  L.off_kin = take(sizeof(uint32_t) * (size_t)n);
  L.off_kout = take(sizeof(uint32_t) * (size_t)n);
  L.off_vin = take(sizeof(int32_t) * (size_t)n);
  L.off_vout = take(sizeof(int32_t) * (size_t)n);
  L.cub_bytes = 0;
  cub::DeviceRadixSort::SortPairs(nullptr, L.cub_bytes, (const uint32_t*)nullptr, (uint32_t*)nullptr,
                                  (const int32_t*)nullptr, (int32_t*)nullptr, (int)n, 0, 21);
  L.off_cub = take(L.cub_bytes + 256);
  L.total = off;
}

__device__ __forceinline__ unsigned kp_f2ord(float f) {
  unsigned u = __float_as_uint(f);
  return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
__device__ __forceinline__ float kp_ord2f(unsigned u) {
  return __uint_as_float((u & 0x80000000u) ? (u & 0x7fffffffu) : ~u);
}

__global__ void kp_init_kernel(KpHeader* h) {
  for (int a = 0; a < 3; ++a) {
    h->mn_bits[a] = 0xffffffffu;
    h->mx_bits[a] = 0u;
  }
}

__global__ void kp_bbox_kernel(const float* __restrict__ xyz, int64_t n, KpHeader* h) {
  unsigned mn[3] = {0xffffffffu, 0xffffffffu, 0xffffffffu}, mx[3] = {0, 0, 0};
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
    for (int a = 0; a < 3; ++a) {
      unsigned o = kp_f2ord(xyz[3 * i + a]);
      mn[a] = min(mn[a], o);
      mx[a] = max(mx[a], o);
    }
  for (int a = 0; a < 3; ++a) {
    for (int o = 16; o > 0; o >>= 1) {
      mn[a] = min(mn[a], __shfl_xor_sync(0xffffffffu, mn[a], o));
      mx[a] = max(mx[a], __shfl_xor_sync(0xffffffffu, mx[a], o));
    }
    if ((threadIdx.x & 31) == 0) {
      atomicMin(&h->mn_bits[a], mn[a]);
      atomicMax(&h->mx_bits[a], mx[a]);
    }
  }
}

// cell edge: the larger of ext / 127 and the edge that puts about `per_cell` points of a SURFACE sampling in a cell
__global__ void kp_grid_kernel(KpHeader* h, int64_t n, float min_cell) {
  float ext = 0.f;
  for (int a = 0; a < 3; ++a) {
    h->mn[a] = kp_ord2f(h->mn_bits[a]);
    ext = fmaxf(ext, kp_ord2f(h->mx_bits[a]) - h->mn[a]);
  }
  float e3[3];
  for (int a = 0; a < 3; ++a) e3[a] = kp_ord2f(h->mx_bits[a]) - h->mn[a];
  const float area = 2.f * (e3[0] * e3[1] + e3[1] * e3[2] + e3[2] * e3[0]);
  float c = fmaxf(min_cell, ext / (float)(KP_DIM - 1));
  if (n > 0 && area > 0.f) c = fmaxf(c, sqrtf(32.f * area / (float)n));  // ~32 points per cell at bounding-box density
  if (!(c > 0.f)) c = 1.f;
  h->cell = c;
  h->inv_cell = 1.f / c;
  for (int a = 0; a < 3; ++a) {
    int d = (int)floorf((kp_ord2f(h->mx_bits[a]) - h->mn[a]) * h->inv_cell) + 1;
    h->dim[a] = max(1, min(d, KP_DIM));
  }
}

__device__ __forceinline__ void kp_cell_of(const KpHeader* h, float x, float y, float z, int& cx, int& cy, int& cz) {
  cx = min(max((int)floorf((x - h->mn[0]) * h->inv_cell), 0), h->dim[0] - 1);
  cy = min(max((int)floorf((y - h->mn[1]) * h->inv_cell), 0), h->dim[1] - 1);
  cz = min(max((int)floorf((z - h->mn[2]) * h->inv_cell), 0), h->dim[2] - 1);
}

__global__ void kp_keys_kernel(const float* __restrict__ xyz, int64_t n, const KpHeader* __restrict__ h,
                               uint32_t* __restrict__ keys, int32_t* __restrict__ vals) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  int cx, cy, cz;
  kp_cell_of(h, xyz[3 * i], xyz[3 * i + 1], xyz[3 * i + 2], cx, cy, cz);
  keys[i] = (uint32_t)((cz * KP_DIM + cy) * KP_DIM + cx);
  vals[i] = (int32_t)i;
}

__global__ void kp_cells_kernel(const float* __restrict__ xyz, const uint32_t* __restrict__ keys_sorted,
                                const int32_t* __restrict__ vals_sorted, int64_t n, float4* __restrict__ sorted,
                                int2* __restrict__ cells) {
  const int64_t s = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (s >= n) return;
  const uint32_t k = keys_sorted[s];
  const int i = vals_sorted[s];
  sorted[s] = make_float4(xyz[3 * (int64_t)i], xyz[3 * (int64_t)i + 1], xyz[3 * (int64_t)i + 2], __int_as_float(i));
  if (s == 0 || keys_sorted[s - 1] != k) cells[k].x = (int)s;
  if (s == n - 1 || keys_sorted[s + 1] != k) cells[k].y = (int)s + 1;
}

__global__ void __launch_bounds__(KP_WARPS * 32) kp_query_kernel(const KpHeader* __restrict__ h,
                                                                  const float4* __restrict__ sorted,
                                                                  const int2* __restrict__ cells, int64_t n, int K,
                                                                  int32_t* __restrict__ out) {
  __shared__ u64 sbuf[KP_WARPS][128];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int64_t s = (int64_t)blockIdx.x * KP_WARPS + warp;  // queries in sorted (cell) order: neighbouring warps share cells
  if (s >= n) return;
  u64* buf = sbuf[warp];
  const float4 q = sorted[s];
  const int self = __float_as_int(q.w);
  int qx, qy, qz;
  kp_cell_of(h, q.x, q.y, q.z, qx, qy, qz);
  const int dx_ = h->dim[0], dy_ = h->dim[1], dz_ = h->dim[2];
  const int rmax = max(dx_, max(dy_, dz_));
  const double cell = (double)h->cell;
  double fx = ((double)q.x - (double)h->mn[0]) / cell - qx, fy = ((double)q.y - (double)h->mn[1]) / cell - qy,
         fz = ((double)q.z - (double)h->mn[2]) / cell - qz;
  double slack = fmin(fmin(fmin(fx, 1.0 - fx), fmin(fy, 1.0 - fy)), fmin(fz, 1.0 - fz));
  slack = fmax(0.0, fmin(slack, 1.0) - 1e-3) * cell;  // margin for the float rounding of the cell assignment
  constexpr u64 KMAX = ~0ull;
  u64 list[4] = {KMAX, KMAX, KMAX, KMAX};
  u64 thr = KMAX;
  int nbuf = 0, merged = 0;
  for (int r = 0; r <= rmax; ++r) {
    const int w = 2 * r + 1;
    const int ncand = w * w * w;
    for (int t0 = 0; t0 < ncand; t0 += 32) {
      const int t = t0 + lane;
      int2 ce = make_int2(0, 0);
      if (t < ncand) {
        const int ddx = t % w - r, ddy = (t / w) % w - r, ddz = t / (w * w) - r;
        const bool shell = max(abs(ddx), max(abs(ddy), abs(ddz))) == r;
        const int x = qx + ddx, y = qy + ddy, z = qz + ddz;
        if (shell && x >= 0 && y >= 0 && z >= 0 && x < dx_ && y < dy_ && z < dz_)
          ce = __ldg(&cells[(z * KP_DIM + y) * KP_DIM + x]);
      }
      unsigned nonempty = __ballot_sync(0xffffffffu, ce.y > ce.x);
      while (nonempty) {
        const int src = __ffs(nonempty) - 1;
        nonempty &= nonempty - 1;
        const int beg = __shfl_sync(0xffffffffu, ce.x, src), end = __shfl_sync(0xffffffffu, ce.y, src);
        for (int j0 = beg; j0 < end; j0 += 32) {
          const int j = j0 + lane;
          u64 key = KMAX;
          if (j < end) {
            const float4 c = __ldg(&sorted[j]);
            const int ci = __float_as_int(c.w);
            if (ci != self) {
              const float ex = __fsub_rn(c.x, q.x), ey = __fsub_rn(c.y, q.y), ez = __fsub_rn(c.z, q.z);
              const float d2 = __fadd_rn(__fadd_rn(__fmul_rn(ex, ex), __fmul_rn(ey, ey)), __fmul_rn(ez, ez));
              key = ((u64)__float_as_uint(d2) << 32) | (u64)(unsigned)ci;
            }
          }
          const bool take = key < thr;
          const unsigned bal = __ballot_sync(0xffffffffu, take);
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
    if (nbuf && !(thr == KMAX && merged + nbuf < K && r < rmax)) {  // fewer than K candidates so far: keep buffering
      __syncwarp();
      topk_flush(list, buf, nbuf, lane);
      merged += nbuf;
      nbuf = 0;
      thr = topk_element(list, K - 1);
      __syncwarp();
    }
    // every point closer than g has been scanned (cells within Chebyshev distance r of the query's cell)
    const double g = fmax(0.0, r * cell + slack - 1e-3 * cell * (r > 0));
    if (thr != KMAX) {
      const double kth = (double)__uint_as_float((unsigned)(thr >> 32));
      // float32 rounding of d2 is below 1e-6 relative: a point outside radius g cannot round below g^2 (1 - 1e-5)
      if (kth < g * g * (1.0 - 1e-5)) break;
    }
  }
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    const int e = lane * 4 + j;
    if (e < K) out[(int64_t)self * K + e] = list[j] == KMAX ? -1 : (int)(list[j] & 0xffffffffu);
  }
}

}  // namespace

extern "C" size_t gp_knn_points_workspace_bytes(int64_t n) {
  if (n < 0) return 0;
  KpLayout L;
  kp_layout(n, L);
  return L.total;
}

extern "C" int gp_knn_points(const float* xyz, int64_t n, int K, float min_cell, int32_t* nbr, void* workspace,
                             size_t workspace_bytes, gp_stream_t stream) {
  GP_CHECK_ARG(n >= 0 && n < (1ll << 31), "gp_knn_points: bad n");
  GP_CHECK_ARG(K >= 1 && K <= 128, "gp_knn_points: K=%d outside [1,128]", K);
  GP_CHECK_ARG(n == 0 || n >= (int64_t)K + 1, "gp_knn_points: needs n >= K+1 points (n=%lld, K=%d)", (long long)n, K);
  GP_CHECK_ARG(min_cell > 0.f, "gp_knn_points: min_cell must be > 0");
  if (n == 0) return GP_OK;
  KpLayout L;
  kp_layout(n, L);
  if (workspace_bytes < L.total) {
    gp_set_error("gp_knn_points: workspace %zu < %zu", workspace_bytes, L.total);
    return GP_ERR_WORKSPACE;
  }
  cudaStream_t st = (cudaStream_t)stream;
  char* ws = (char*)workspace;
  KpHeader* h = (KpHeader*)(ws + L.off_hdr);
  int2* cells = (int2*)(ws + L.off_cells);
  uint32_t* kin = (uint32_t*)(ws + L.off_kin);
  uint32_t* kout = (uint32_t*)(ws + L.off_kout);
  int32_t* vin = (int32_t*)(ws + L.off_vin);
  int32_t* vout = (int32_t*)(ws + L.off_vout);
  float4* sorted = (float4*)(ws + L.off_sorted);
  GP_CUDA(cudaMemsetAsync(cells, 0, sizeof(int2) * (size_t)KP_NCELLS, st));
  kp_init_kernel<<<1, 1, 0, st>>>(h);
  const unsigned grid = (unsigned)((n + 255) / 256);
  kp_bbox_kernel<<<grid < 1184 ? grid : 1184, 256, 0, st>>>(xyz, n, h);
  kp_grid_kernel<<<1, 1, 0, st>>>(h, n, min_cell);
  kp_keys_kernel<<<grid, 256, 0, st>>>(xyz, n, h, kin, vin);
  GP_LAUNCH_CHECK();
  size_t cb = L.cub_bytes + 256;
  GP_CUDA(cub::DeviceRadixSort::SortPairs(ws + L.off_cub, cb, kin, kout, vin, vout, (int)n, 0, 21, st));
  kp_cells_kernel<<<grid, 256, 0, st>>>(xyz, kout, vout, n, sorted, cells);
  GP_LAUNCH_CHECK();
  kp_query_kernel<<<(unsigned)((n + KP_WARPS - 1) / KP_WARPS), KP_WARPS * 32, 0, st>>>(h, sorted, cells, n, K, nbr);
  GP_LAUNCH_CHECK();
  return GP_OK;
}
