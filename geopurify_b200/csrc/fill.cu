// 1-NN hole fill over float coordinates + row copy / gather helpers.
// Replaces the sklearn KDTree k=1 queries of the reference (models/affinity_module.py:445-449, 693-696;
// run/validation.py:425-430) and the index-select of models/affinity_module.py:1589.
#include <cub/device/device_radix_sort.cuh>
#include "common.cuh"

#define NN1_MAX_DIM 128
#define NN1_COARSE 8                                  // fine cells per coarse cell and axis
#define NN1_CDIM (NN1_MAX_DIM / NN1_COARSE)           // coarse cells per axis
#define NN1_NCELLS (NN1_MAX_DIM * NN1_MAX_DIM * NN1_MAX_DIM)
#define NN1_NCOARSE (NN1_CDIM * NN1_CDIM * NN1_CDIM)

// cell key, coarse-major: all points of one coarse cell (8^3 fine cells) are contiguous after the sort
__host__ __device__ __forceinline__ uint32_t nn1_key(int cx, int cy, int cz) {
  const uint32_t coarse = (uint32_t)(((cz >> 3) * NN1_CDIM + (cy >> 3)) * NN1_CDIM + (cx >> 3));
  const uint32_t fine = (uint32_t)((((cz & 7) << 3) | (cy & 7)) << 3 | (cx & 7));
  return (coarse << 9) | fine;
}

struct Nn1Header {
  float mn[3];
  float inv_cell;
  float cell;
  int dim[3];
  unsigned mn_bits[3], mx_bits[3];  // order-preserving encodings for atomicMin/Max
};

struct Nn1Layout {
  size_t off_hdr, off_cells, off_sorted, off_kin, off_kout, off_vin, off_vout, off_cub, cub_bytes, total;
};

static void nn1_layout(int64_t n, Nn1Layout& L) {
  size_t off = 0;
  auto take = [&](size_t bytes) {
    size_t o = off;
    off += gp_align_up(bytes);
    return o;
  };
  L.off_hdr = take(sizeof(Nn1Header));
  L.off_cells = take(sizeof(int2) * ((size_t)NN1_NCELLS + NN1_NCOARSE));  // fine table, then the coarse table
  L.off_sorted = take(sizeof(float4) * (size_t)n);
  L.off_kin = take(sizeof(uint32_t) * (size_t)n);
  L.off_kout = take(sizeof(uint32_t) * (size_t)n);
  L.off_vin = take(sizeof(int32_t) * (size_t)n);
  L.off_vout = take(sizeof(int32_t) * (size_t)n);
  L.cub_bytes = 0;
  cub::DeviceRadixSort::SortPairs(nullptr, L.cub_bytes, (const uint32_t*)nullptr, (uint32_t*)nullptr,
                                  (const int32_t*)nullptr, (int32_t*)nullptr, (int)n, 0, 23);
  L.off_cub = take(L.cub_bytes + 256);
  L.total = off;
}

__device__ __forceinline__ unsigned f2ord(float f) {
  unsigned u = __float_as_uint(f);
  return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
__device__ __forceinline__ float ord2f(unsigned u) {
  return __uint_as_float((u & 0x80000000u) ? (u & 0x7fffffffu) : ~u);
}

__global__ void nn1_init_kernel(Nn1Header* h) {
  for (int a = 0; a < 3; ++a) {
    h->mn_bits[a] = 0xffffffffu;
    h->mx_bits[a] = 0u;
  }
}

__global__ void nn1_bbox_kernel(const float* __restrict__ xyz, int64_t n, Nn1Header* h) {
  unsigned mn[3] = {0xffffffffu, 0xffffffffu, 0xffffffffu}, mx[3] = {0, 0, 0};
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
    for (int a = 0; a < 3; ++a) {
      unsigned o = f2ord(xyz[3 * i + a]);
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

__global__ void nn1_grid_kernel(Nn1Header* h, float cell) {
  float ext = 0.f;
  for (int a = 0; a < 3; ++a) {
    h->mn[a] = ord2f(h->mn_bits[a]);
    ext = fmaxf(ext, ord2f(h->mx_bits[a]) - h->mn[a]);
  }
  float c = fmaxf(cell, ext / (float)(NN1_MAX_DIM - 1));
  if (!(c > 0.f)) c = 1.f;
  h->cell = c;
  h->inv_cell = 1.f / c;
  for (int a = 0; a < 3; ++a) {
    int d = (int)floorf((ord2f(h->mx_bits[a]) - h->mn[a]) * h->inv_cell) + 1;
    h->dim[a] = max(1, min(d, NN1_MAX_DIM));
  }
}

__device__ __forceinline__ void nn1_cell_of(const Nn1Header* h, float x, float y, float z, int& cx, int& cy, int& cz) {
  cx = min(max((int)floorf((x - h->mn[0]) * h->inv_cell), 0), h->dim[0] - 1);
  cy = min(max((int)floorf((y - h->mn[1]) * h->inv_cell), 0), h->dim[1] - 1);
  cz = min(max((int)floorf((z - h->mn[2]) * h->inv_cell), 0), h->dim[2] - 1);
}

__global__ void nn1_keys_kernel(const float* __restrict__ xyz, const uint8_t* __restrict__ seen, int64_t n,
                                const Nn1Header* __restrict__ h, uint32_t* __restrict__ keys,
                                int32_t* __restrict__ vals, int32_t* __restrict__ nn) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  int cx, cy, cz;
  nn1_cell_of(h, xyz[3 * i], xyz[3 * i + 1], xyz[3 * i + 2], cx, cy, cz);
  uint32_t key = nn1_key(cx, cy, cz);
  const uint8_t st = seen[i];
  if (st == 1) nn[i] = (int32_t)i;
  else if (st == 0) key |= 1u << 21;  // unseen points sort after every seen one, in cell order
  else key |= 1u << 22;               // state 2: neither a source nor a query (nn[i] left untouched)
  keys[i] = key;
  vals[i] = (int32_t)i;
}

__global__ void nn1_cells_kernel(const float* __restrict__ xyz, const uint32_t* __restrict__ keys_sorted,
                                 const int32_t* __restrict__ vals_sorted, int64_t n, float4* __restrict__ sorted,
                                 int2* __restrict__ cells) {
  int64_t s = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (s >= n) return;
  uint32_t k = keys_sorted[s];
  if (k >> 21) return;
  int i = vals_sorted[s];
  sorted[s] = make_float4(xyz[3 * (int64_t)i], xyz[3 * (int64_t)i + 1], xyz[3 * (int64_t)i + 2], __int_as_float(i));
  if (s == 0 || keys_sorted[s - 1] != k) cells[k].x = (int)s;
  if (s == n - 1 || keys_sorted[s + 1] != k) cells[k].y = (int)s + 1;
  int2* coarse = cells + NN1_NCELLS;
  if (s == 0 || (keys_sorted[s - 1] >> 9) != (k >> 9)) coarse[k >> 9].x = (int)s;
  if (s == n - 1 || (keys_sorted[s + 1] >> 9) != (k >> 9)) coarse[k >> 9].y = (int)s + 1;
}

// One warp per unseen point.  Unseen points carry sort keys (1 << 21 | own cell), so they sit at the tail of the
// sorted arrays in cell order (neighbouring warps search neighbouring cells); warp w takes sorted position n-1-w and
// quits when it reaches a seen point.  Lanes enumerate the cells of Chebyshev shell r in parallel, each lane scans
// its cell's members; distances in float64 on the float32 coordinates (as the KD-tree does), ties -> lowest index.
__device__ __forceinline__ bool shell_cell(int r, int t, int& dx, int& dy, int& dz) {
  if (r == 0) {
    dx = dy = dz = 0;
    return t == 0;
  }
  const int w = 2 * r + 1, face = w * w, ring = 8 * r;
  if (t < face) {
    dz = -r;
    dx = t % w - r;
    dy = t / w - r;
    return true;
  }
  t -= face;
  if (t < (w - 2) * ring) {
    dz = t / ring - r + 1;
    int u = t % ring, side = u / (2 * r), j = u % (2 * r);
    dx = side == 0 ? -r + j : (side == 1 ? r : (side == 2 ? r - j : -r));
    dy = side == 0 ? -r : (side == 1 ? -r + j : (side == 2 ? r : r - j));
    return true;
  }
  t -= (w - 2) * ring;
  if (t < face) {
    dz = r;
    dx = t % w - r;
    dy = t / w - r;
    return true;
  }
  return false;
}

__global__ void __launch_bounds__(256) nn1_query_kernel(const float* __restrict__ xyz, int64_t n,
                                                         const Nn1Header* __restrict__ h,
                                                         const uint32_t* __restrict__ keys_sorted,
                                                         const int32_t* __restrict__ vals_sorted,
                                                         const float4* __restrict__ sorted,
                                                         const int2* __restrict__ cells, int32_t* __restrict__ nn) {
  const int lane = threadIdx.x & 31;
  const int64_t w = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (w >= n) return;
  const int64_t spos = n - 1 - w;
  if ((keys_sorted[spos] >> 21) != 1) return;  // only unseen points (state 0) are queries
  const int i = vals_sorted[spos];
  const float x = xyz[3 * (int64_t)i], y = xyz[3 * (int64_t)i + 1], z = xyz[3 * (int64_t)i + 2];
  int qx, qy, qz;
  nn1_cell_of(h, x, y, z, qx, qy, qz);
  const int dx_ = h->dim[0], dy_ = h->dim[1], dz_ = h->dim[2];
  const int rmax = max(dx_, max(dy_, dz_));
  const double cell = (double)h->cell;
  // distance from the query to the faces of its own cell bounds what ring r guarantees
  double fx = ((double)x - (double)h->mn[0]) / cell - qx, fy = ((double)y - (double)h->mn[1]) / cell - qy,
         fz = ((double)z - (double)h->mn[2]) / cell - qz;
  double slack = fmin(fmin(fmin(fx, 1.0 - fx), fmin(fy, 1.0 - fy)), fmin(fz, 1.0 - fz));
  slack = fmax(0.0, fmin(slack, 1.0) - 1e-3) * cell;  // margin for float rounding of the cell assignment
  double best = 1e300;
  int besti = 0x7fffffff;
  bool done = false;
  // ---- phase 1: fine shells around the query's own cell (holes inside seen regions resolve here)
  const int r1 = min(rmax, 2);
  for (int r = 0; r <= r1 && !done; ++r) {
    const int wd = 2 * r + 1;
    const int total = r == 0 ? 1 : 2 * wd * wd + (wd - 2) * 8 * r;
    for (int t = lane; t < total; t += 32) {
      int ddx, ddy, ddz;
      if (!shell_cell(r, t, ddx, ddy, ddz)) continue;
      int cx = qx + ddx, cy = qy + ddy, cz = qz + ddz;
      if (cx < 0 || cy < 0 || cz < 0 || cx >= dx_ || cy >= dy_ || cz >= dz_) continue;
      int2 ce = __ldg(&cells[nn1_key(cx, cy, cz)]);
      for (int j = ce.x; j < ce.y; ++j) {
        float4 p = __ldg(&sorted[j]);
        double ex = (double)p.x - (double)x, ey = (double)p.y - (double)y, ez = (double)p.z - (double)z;
        double d2 = ex * ex + ey * ey + ez * ez;
        int pi = __float_as_int(p.w);
        if (d2 < best || (d2 == best && pi < besti)) {
          best = d2;
          besti = pi;
        }
      }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      double ob = __shfl_xor_sync(0xffffffffu, best, o);
      int oi = __shfl_xor_sync(0xffffffffu, besti, o);
      if (ob < best || (ob == best && oi < besti)) {
        best = ob;
        besti = oi;
      }
    }
    double g = fmax(0.0, r * cell + slack - 1e-3 * cell * (r > 0));
    if (besti != 0x7fffffff && best <= g * g) done = true;
  }
  // ---- phase 2: coarse shells (8^3 fine cells each).  Lanes test the shell's coarse cells against the current best
  // through the distance to the cell's box; the warp then scans the surviving cells' points together (contiguous).
  if (!done) {
    const int2* coarse = cells + NN1_NCELLS;
    const int qX = qx >> 3, qY = qy >> 3, qZ = qz >> 3;
    const int cX = (dx_ + 7) >> 3, cY = (dy_ + 7) >> 3, cZ = (dz_ + 7) >> 3;
    const int rmaxc = max(cX, max(cY, cZ));
    const double ccell = cell * NN1_COARSE, eps = 1e-3 * cell;
    const double ox = (double)h->mn[0], oy = (double)h->mn[1], oz = (double)h->mn[2];
    double gx = ((double)x - ox) / ccell - qX, gy = ((double)y - oy) / ccell - qY, gz = ((double)z - oz) / ccell - qZ;
    double cslack = fmin(fmin(fmin(gx, 1.0 - gx), fmin(gy, 1.0 - gy)), fmin(gz, 1.0 - gz));
    cslack = fmax(0.0, fmin(cslack, 1.0)) * ccell;
    for (int r = 0; r <= rmaxc; ++r) {
      const int wd = 2 * r + 1;
      const int total = r == 0 ? 1 : 2 * wd * wd + (wd - 2) * 8 * r;
      for (int t0 = 0; t0 < total; t0 += 32) {
        const int t = t0 + lane;
        int2 ce = make_int2(0, 0);
        double lb2 = 1e300;
        int ddx, ddy, ddz;
        if (t < total && shell_cell(r, t, ddx, ddy, ddz)) {
          const int cx = qX + ddx, cy = qY + ddy, cz = qZ + ddz;
          if (cx >= 0 && cy >= 0 && cz >= 0 && cx < cX && cy < cY && cz < cZ) {
            ce = __ldg(&coarse[(cz * NN1_CDIM + cy) * NN1_CDIM + cx]);
            const double lx = ox + cx * ccell, ly = oy + cy * ccell, lz = oz + cz * ccell;
            const double ax = fmax(0.0, fmax(lx - (double)x, (double)x - (lx + ccell)) - eps);
            const double ay = fmax(0.0, fmax(ly - (double)y, (double)y - (ly + ccell)) - eps);
            const double az = fmax(0.0, fmax(lz - (double)z, (double)z - (lz + ccell)) - eps);
            lb2 = ax * ax + ay * ay + az * az;
          }
        }
        unsigned cand = __ballot_sync(0xffffffffu, ce.y > ce.x && lb2 <= best);
        while (cand) {
          const int l = __ffs(cand) - 1;
          cand &= cand - 1;
          const int j0 = __shfl_sync(0xffffffffu, ce.x, l), j1 = __shfl_sync(0xffffffffu, ce.y, l);
          const double clb = __shfl_sync(0xffffffffu, lb2, l);
          if (clb > best) continue;  // pruned by a cell scanned since the ballot (best is warp-uniform)
          for (int j = j0 + lane; j < j1; j += 32) {
            float4 p = __ldg(&sorted[j]);
            double ex = (double)p.x - (double)x, ey = (double)p.y - (double)y, ez = (double)p.z - (double)z;
            double d2 = ex * ex + ey * ey + ez * ez;
            int pi = __float_as_int(p.w);
            if (d2 < best || (d2 == best && pi < besti)) {
              best = d2;
              besti = pi;
            }
          }
#pragma unroll
          for (int o = 16; o > 0; o >>= 1) {
            double ob = __shfl_xor_sync(0xffffffffu, best, o);
            int oi = __shfl_xor_sync(0xffffffffu, besti, o);
            if (ob < best || (ob == best && oi < besti)) {
              best = ob;
              besti = oi;
            }
          }
        }
      }
      double g = fmax(0.0, r * ccell + cslack - eps * (r > 0) - eps);
      if (besti != 0x7fffffff && best <= g * g) break;
    }
  }
  if (lane == 0) nn[i] = besti == 0x7fffffff ? -1 : besti;
}

extern "C" size_t gp_nn1_fill_workspace_bytes(int64_t n) {
  if (n < 0) return 0;
  Nn1Layout L;
  nn1_layout(n, L);
  return L.total;
}

extern "C" int gp_nn1_fill(const float* xyz, int64_t n, const uint8_t* seen, float cell, int32_t* nn, void* workspace,
                           size_t workspace_bytes, gp_stream_t stream) {
  GP_CHECK_ARG(n >= 0 && n < (1ll << 31), "gp_nn1_fill: bad n");
  GP_CHECK_ARG(cell > 0.f, "gp_nn1_fill: cell must be > 0");
  if (n == 0) return GP_OK;
  Nn1Layout L;
  nn1_layout(n, L);
  if (workspace_bytes < L.total) {
    gp_set_error("gp_nn1_fill: workspace %zu < %zu", workspace_bytes, L.total);
    return GP_ERR_WORKSPACE;
  }
  cudaStream_t st = (cudaStream_t)stream;
  char* ws = (char*)workspace;
  Nn1Header* h = (Nn1Header*)(ws + L.off_hdr);
  int2* cells = (int2*)(ws + L.off_cells);
  GP_CUDA(cudaMemsetAsync(cells, 0, sizeof(int2) * ((size_t)NN1_NCELLS + NN1_NCOARSE), st));
  nn1_init_kernel<<<1, 1, 0, st>>>(h);
  unsigned grid = (unsigned)((n + 255) / 256);
  nn1_bbox_kernel<<<grid < 1184 ? grid : 1184, 256, 0, st>>>(xyz, n, h);
  nn1_grid_kernel<<<1, 1, 0, st>>>(h, cell);
  uint32_t* kin = (uint32_t*)(ws + L.off_kin);
  uint32_t* kout = (uint32_t*)(ws + L.off_kout);
  int32_t* vin = (int32_t*)(ws + L.off_vin);
  int32_t* vout = (int32_t*)(ws + L.off_vout);
  nn1_keys_kernel<<<grid, 256, 0, st>>>(xyz, seen, n, h, kin, vin, nn);
  GP_LAUNCH_CHECK();
  size_t cb = L.cub_bytes + 256;
  GP_CUDA(cub::DeviceRadixSort::SortPairs(ws + L.off_cub, cb, kin, kout, vin, vout, (int)n, 0, 23, st));
  nn1_cells_kernel<<<grid, 256, 0, st>>>(xyz, kout, vout, n, (float4*)(ws + L.off_sorted), cells);
  nn1_query_kernel<<<(unsigned)((n + 7) / 8), 256, 0, st>>>(xyz, n, h, kout, vout, (const float4*)(ws + L.off_sorted),
                                                           cells, nn);
  GP_LAUNCH_CHECK();
  return GP_OK;
}

// ------------------------------------------------------------------ row helpers
__global__ void copy_rows_kernel(float* __restrict__ rows, int64_t n, int D4, int64_t stride4,
                                 const int32_t* __restrict__ nn) {
  int64_t i = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (i >= n) return;
  int src = nn[i];
  if (src < 0 || src == i) return;
  const float4* s = (const float4*)rows + (int64_t)src * stride4;
  float4* d = (float4*)rows + i * stride4;
  for (int j = threadIdx.x & 31; j < D4; j += 32) d[j] = s[j];
}

extern "C" int gp_copy_rows(float* rows, int64_t n, int D, int64_t stride, const int32_t* nn, gp_stream_t stream) {
  GP_CHECK_ARG(n >= 0 && D > 0 && D % 4 == 0 && stride % 4 == 0 && stride >= D, "gp_copy_rows: bad sizes");
  if (n == 0) return GP_OK;
  copy_rows_kernel<<<(unsigned)((n + 7) / 8), 256, 0, (cudaStream_t)stream>>>(rows, n, D / 4, stride / 4, nn);
  GP_LAUNCH_CHECK();
  return GP_OK;
}

__global__ void gather_rows_kernel(const float* __restrict__ rows, int64_t stride4, int D4,
                                   const int32_t* __restrict__ index, int64_t n, float* __restrict__ out) {
  int64_t i = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (i >= n) return;
  const float4* s = (const float4*)rows + (int64_t)index[i] * stride4;
  float4* d = (float4*)out + i * D4;
  for (int j = threadIdx.x & 31; j < D4; j += 32) d[j] = __ldg(&s[j]);
}

extern "C" int gp_gather_rows(const float* rows, int64_t row_stride, int D, const int32_t* index, int64_t n,
                              float* out, gp_stream_t stream) {
  GP_CHECK_ARG(n >= 0 && D > 0 && D % 4 == 0 && row_stride % 4 == 0 && row_stride >= D, "gp_gather_rows: bad sizes");
  if (n == 0) return GP_OK;
  gather_rows_kernel<<<(unsigned)((n + 7) / 8), 256, 0, (cudaStream_t)stream>>>(rows, row_stride / 4, D / 4, index, n,
                                                                              out);
  GP_LAUNCH_CHECK();
  return GP_OK;
}
