// Stage 1: point -> view projection, occlusion test, z-buffer depth, fused gather-mean lifting.
// Replaces PointCloudToImageMapper.compute_mapping (reference models/utils/fusion_util.py:99-147, :45-82)
// and the gather + mean of lift_lseg_features (reference models/affinity_module.py:404-438).
#include "common.cuh"

#define GP_MAX_VIEWS_SMEM 256

struct Cam {  // 16 doubles per view
  double m[12];  // rows 0..2 of W2C
  double fx, fy, cx, cy;
};

static_assert(sizeof(Cam) == 128, "Cam is 16 doubles");

__device__ __forceinline__ void load_cams(Cam* s, const double* w2c, const double* intr, int v0, int nv) {
  for (int i = threadIdx.x; i < nv * 16; i += blockDim.x) {
    int v = i >> 4, j = i & 15;
    double val = j < 12 ? w2c[(size_t)(v0 + v) * 16 + j] : intr[(size_t)(v0 + v) * 4 + (j - 12)];
    ((double*)s)[i] = val;
  }
}

// Field-major copy for kernels whose LANES walk the views (lift_mean_kernel): s[j * nv + v].  With the 128-byte Cam
// records above, lane = view puts all 32 lanes on the same two banks (a 32-way conflict on each of the 16 loads of a
// projection); field-major, consecutive lanes read consecutive doubles.
__device__ __forceinline__ void load_cams_soa(double* s, const double* w2c, const double* intr, int nv) {
  for (int i = threadIdx.x; i < nv * 16; i += blockDim.x) {
    int v = i >> 4, j = i & 15;
    s[j * nv + v] = j < 12 ? w2c[(size_t)v * 16 + j] : intr[(size_t)v * 4 + (j - 12)];
  }
}
__device__ __forceinline__ Cam cam_from_soa(const double* s, int nv, int v) {
  Cam c;
#pragma unroll
  for (int j = 0; j < 12; ++j) c.m[j] = s[j * nv + v];
  c.fx = s[12 * nv + v];
  c.fy = s[13 * nv + v];
  c.cx = s[14 * nv + v];
  c.cy = s[15 * nv + v];
  return c;
}

// numpy's 4x4 @ 4xN float64 product runs through OpenBLAS dgemm: an FMA chain over k = 0..3 starting from
// a0*x (verified bit-for-bit against numpy); the rest is separate IEEE ops (numpy ufuncs).
__device__ __forceinline__ bool project_point(const Cam& c, double x, double y, double z, int H, int W, int cut,
                                              double& pz, int& row, int& col) {
  double p0 = __fma_rn(c.m[3], 1.0, __fma_rn(c.m[2], z, __fma_rn(c.m[1], y, __dmul_rn(c.m[0], x))));
  double p1 = __fma_rn(c.m[7], 1.0, __fma_rn(c.m[6], z, __fma_rn(c.m[5], y, __dmul_rn(c.m[4], x))));
  double p2 = __fma_rn(c.m[11], 1.0, __fma_rn(c.m[10], z, __fma_rn(c.m[9], y, __dmul_rn(c.m[8], x))));
  double u = __dadd_rn(__ddiv_rn(__dmul_rn(p0, c.fx), p2), c.cx);
  double v = __dadd_rn(__ddiv_rn(__dmul_rn(p1, c.fy), p2), c.cy);
  double ru = rint(u), rv = rint(v);  // np.round: half to even
  pz = p2;
  // NaN / inf / out-of-int64 values become INT64_MIN in numpy's astype(int) and fail `>= cut`; the double
  // comparisons below reject exactly the same inputs.
  bool inside = (ru >= (double)cut) && (rv >= (double)cut) && (ru < (double)(W - cut)) && (rv < (double)(H - cut));
  if (inside) {
    col = (int)ru;
    row = (int)rv;
  }
  return inside;
}

template <int DEPTH_DTYPE>
__device__ __forceinline__ bool depth_test(const void* depth, size_t off, double pz, double thr, float thr32) {
  if (DEPTH_DTYPE == 0) return pz > 0.0;
  if (DEPTH_DTYPE == GP_F64) {
    double d = __ldg((const double*)depth + off);
    return fabs(__dsub_rn(d, pz)) <= __dmul_rn(thr, d);
  } else {
    float d = __ldg((const float*)depth + off);
    // numpy: float32 array * python float stays float32; float32 array - float64 array is float64
    return fabs(__dsub_rn((double)d, pz)) <= (double)__fmul_rn(thr32, d);
  }
}

template <typename T>
__device__ __forceinline__ void load_xyz(const void* coords, int64_t i, double& x, double& y, double& z) {
  const T* c = (const T*)coords + 3 * i;
  x = (double)c[0];
  y = (double)c[1];
  z = (double)c[2];
}

// ------------------------------------------------------------------ gp_project_views
template <typename CT, int DEPTH_DTYPE>
__global__ void __launch_bounds__(256) project_views_kernel(const void* __restrict__ coords, int64_t n,
                                                             const double* __restrict__ w2c,
                                                             const double* __restrict__ intr,
                                                             const void* __restrict__ depth, int V, int H, int W,
                                                             int cut, double thr, int32_t* __restrict__ pix,
                                                             int32_t* __restrict__ counts) {
  extern __shared__ double smem_d[];
  Cam* cams = (Cam*)smem_d;
  int* cnt = (int*)(cams + V);
  load_cams(cams, w2c, intr, 0, V);
  for (int i = threadIdx.x; i < V; i += blockDim.x) cnt[i] = 0;
  __syncthreads();
  const float thr32 = (float)thr;
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  bool live = i < n;
  double x = 0, y = 0, z = 0;
  if (live) load_xyz<CT>(coords, i, x, y, z);
  for (int v = 0; v < V; ++v) {
    int row = 0, col = 0;
    double pz;
    bool vis = false;
    if (live) {
      vis = project_point(cams[v], x, y, z, H, W, cut, pz, row, col);
      if (vis) vis = depth_test<DEPTH_DTYPE>(depth, ((size_t)v * H + row) * W + col, pz, thr, thr32);
      pix[(size_t)v * n + i] = vis ? row * W + col : -1;
    }
    unsigned b = __ballot_sync(0xffffffffu, vis);
    if ((threadIdx.x & 31) == 0 && b) atomicAdd(&cnt[v], __popc(b));
  }
  __syncthreads();
  for (int v = threadIdx.x; v < V; v += blockDim.x)
    if (cnt[v]) atomicAdd(&counts[v], cnt[v]);
}

extern "C" int gp_project_views(const void* coords, int coords_dtype, int64_t n, const double* w2c,
                                const double* intr, const void* depth, int depth_dtype, int V, int H, int W,
                                int cut, double vis_thres, int32_t* pix, int32_t* counts, gp_stream_t stream) {
  GP_CHECK_ARG(n >= 0 && V >= 0 && H > 0 && W > 0 && cut >= 0, "gp_project_views: bad sizes");
  GP_CHECK_ARG(V <= GP_MAX_VIEWS_SMEM, "gp_project_views: V=%d exceeds %d views per call", V, GP_MAX_VIEWS_SMEM);
  GP_CHECK_ARG(coords_dtype == GP_F32 || coords_dtype == GP_F64, "gp_project_views: coords dtype");
  GP_CHECK_ARG((depth == nullptr) == (depth_dtype == 0), "gp_project_views: depth pointer / dtype mismatch");
  GP_CHECK_ARG(depth_dtype == 0 || depth_dtype == GP_F32 || depth_dtype == GP_F64, "gp_project_views: depth dtype");
  GP_CHECK_ARG((int64_t)H * W < (1ll << 31), "gp_project_views: image too large");
  cudaStream_t st = (cudaStream_t)stream;
  if (V == 0) return GP_OK;
  GP_CUDA(cudaMemsetAsync(counts, 0, sizeof(int32_t) * V, st));
  if (n == 0) return GP_OK;
  size_t smem = sizeof(Cam) * V + sizeof(int) * V;
  unsigned grid = (unsigned)((n + 255) / 256);
#define LAUNCH(CT, DD)                                                                                          \
  project_views_kernel<CT, DD><<<grid, 256, smem, st>>>(coords, n, w2c, intr, depth, V, H, W, cut, vis_thres, \
                                                        pix, counts)
  if (coords_dtype == GP_F64) {
    if (depth_dtype == 0) LAUNCH(double, 0);
    else if (depth_dtype == GP_F64) LAUNCH(double, GP_F64);
    else LAUNCH(double, GP_F32);
  } else {
    if (depth_dtype == 0) LAUNCH(float, 0);
    else if (depth_dtype == GP_F64) LAUNCH(float, GP_F64);
    else LAUNCH(float, GP_F32);
  }
#undef LAUNCH
  GP_LAUNCH_CHECK();
  return GP_OK;
}

// ------------------------------------------------------------------ gp_pix_to_mapping
__global__ void pix_to_mapping_kernel(const int32_t* __restrict__ pix, int64_t total, int W,
                                      int64_t* __restrict__ mapping) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= total) return;
  int p = pix[i];
  int64_t r = 0, c = 0, m = 0;
  if (p >= 0) {
    r = p / W;
    c = p % W;
    m = 1;
  }
  mapping[3 * i] = r;
  mapping[3 * i + 1] = c;
  mapping[3 * i + 2] = m;
}

extern "C" int gp_pix_to_mapping(const int32_t* pix, int64_t total, int W, int64_t* mapping, gp_stream_t stream) {
  GP_CHECK_ARG(total >= 0 && W > 0, "gp_pix_to_mapping: bad sizes");
  if (total == 0) return GP_OK;
  pix_to_mapping_kernel<<<(unsigned)((total + 255) / 256), 256, 0, (cudaStream_t)stream>>>(pix, total, W, mapping);
  GP_LAUNCH_CHECK();
  return GP_OK;
}

// ------------------------------------------------------------------ gp_splat_depth
// Positive doubles order like their bit patterns, so the z-buffer min is an atomicMin on uint64.
template <typename CT>
__global__ void __launch_bounds__(256) splat_depth_kernel(const void* __restrict__ coords, int64_t n,
                                                           const double* __restrict__ w2c,
                                                           const double* __restrict__ intr, int V, int H, int W,
                                                           int cut, unsigned long long* __restrict__ depth_bits) {
  extern __shared__ double smem_d[];
  Cam* cams = (Cam*)smem_d;
  load_cams(cams, w2c, intr, 0, V);
  __syncthreads();
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  double x, y, z;
  load_xyz<CT>(coords, i, x, y, z);
  for (int v = 0; v < V; ++v) {
    int row, col;
    double pz;
    if (project_point(cams[v], x, y, z, H, W, cut, pz, row, col) && pz > 0.2)
      atomicMin(&depth_bits[((size_t)v * H + row) * W + col], (unsigned long long)__double_as_longlong(pz));
  }
}

__global__ void fill_f64_kernel(double* p, int64_t n, double val) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) p[i] = val;
}

extern "C" int gp_splat_depth(const void* coords, int coords_dtype, int64_t n, const double* w2c, const double* intr,
                              int V, int H, int W, int cut, double* depth, gp_stream_t stream) {
  GP_CHECK_ARG(n >= 0 && V >= 0 && H > 0 && W > 0 && cut >= 0, "gp_splat_depth: bad sizes");
  GP_CHECK_ARG(V <= GP_MAX_VIEWS_SMEM, "gp_splat_depth: too many views");
  GP_CHECK_ARG(coords_dtype == GP_F32 || coords_dtype == GP_F64, "gp_splat_depth: coords dtype");
  cudaStream_t st = (cudaStream_t)stream;
  int64_t total = (int64_t)V * H * W;
  if (total == 0) return GP_OK;
  fill_f64_kernel<<<(unsigned)((total + 255) / 256), 256, 0, st>>>(depth, total, 999999.0);
  GP_LAUNCH_CHECK();
  if (n == 0) return GP_OK;
  unsigned grid = (unsigned)((n + 255) / 256);
  size_t smem = sizeof(Cam) * V;
  if (coords_dtype == GP_F64)
    splat_depth_kernel<double><<<grid, 256, smem, st>>>(coords, n, w2c, intr, V, H, W, cut,
                                                        (unsigned long long*)depth);
  else
    splat_depth_kernel<float><<<grid, 256, smem, st>>>(coords, n, w2c, intr, V, H, W, cut,
                                                       (unsigned long long*)depth);
  GP_LAUNCH_CHECK();
  return GP_OK;
}

// ------------------------------------------------------------------ gp_accept_views
__global__ void accept_views_kernel(const int32_t* counts, int V, int lo, int hi, uint8_t* keep) {
  int v = blockIdx.x * blockDim.x + threadIdx.x;
  if (v < V) {
    int c = counts[v];
    keep[v] = (c != 0 && c >= lo && c <= hi) ? 1 : 0;
  }
}

extern "C" int gp_accept_views(const int32_t* counts, int V, int min_visible, int max_visible, uint8_t* keep,
                               gp_stream_t stream) {
  GP_CHECK_ARG(V >= 0, "gp_accept_views: bad V");
  if (V == 0) return GP_OK;
  accept_views_kernel<<<(V + 127) / 128, 128, 0, (cudaStream_t)stream>>>(counts, V, min_visible, max_visible, keep);
  GP_LAUNCH_CHECK();
  return GP_OK;
}

// ------------------------------------------------------------------ fused lifting
// One warp per point.  Lane l projects views l, l+32 of a 64-view chunk (cameras in shared memory; the two depth
// samples of a lane are independent loads), the warp ballots the visible set and then walks it in ascending view
// order in BATCHES of up to RB rows: all RB feature rows are requested before the first is added, so a lane has
// RB x (row bytes / 512) independent 128-bit loads in flight instead of one row's worth (round 1: 28 % of HBM peak,
// latency-bound on the depth -> pixel-row chain).  The adds stay in view order: the sum is bit-identical.
// NV4 = float4 accumulators per lane: D = 128 * NV4.
template <int NV4, bool BF16>
struct RowRegs {
  static constexpr int NL = BF16 ? (NV4 / 2 > 0 ? NV4 / 2 : 1) : NV4;   // 128-bit loads per lane per row
  uint4 t[NL];
};

template <int NV4, bool BF16>
__device__ __forceinline__ void load_row(const void* feat, size_t row_index, int D, int lane, RowRegs<NV4, BF16>& r) {
  if constexpr (!BF16) {
    const uint4* p = (const uint4*)((const float*)feat + row_index * (size_t)D);
#pragma unroll
    for (int j = 0; j < NV4; ++j) r.t[j] = ldg_stream_u4(p + j * 32 + lane);
  } else {
    // lane owns 8 consecutive bf16 per 256-element chunk -> two float4 accumulators per uint4
    const uint4* p = (const uint4*)((const __nv_bfloat16*)feat + row_index * (size_t)D);
#pragma unroll
    for (int j = 0; j < NV4 / 2; ++j) r.t[j] = ldg_stream_u4(p + j * 32 + lane);
  }
}

template <int NV4, bool BF16>
__device__ __forceinline__ void add_row(const RowRegs<NV4, BF16>& r, float4 (&acc)[NV4]) {
  if constexpr (!BF16) {
#pragma unroll
    for (int j = 0; j < NV4; ++j) {
      acc[j].x = __fadd_rn(acc[j].x, __uint_as_float(r.t[j].x));
      acc[j].y = __fadd_rn(acc[j].y, __uint_as_float(r.t[j].y));
      acc[j].z = __fadd_rn(acc[j].z, __uint_as_float(r.t[j].z));
      acc[j].w = __fadd_rn(acc[j].w, __uint_as_float(r.t[j].w));
    }
  } else {
#pragma unroll
    for (int j = 0; j < NV4 / 2; ++j) {
      acc[2 * j].x = __fadd_rn(acc[2 * j].x, __uint_as_float(r.t[j].x << 16));
      acc[2 * j].y = __fadd_rn(acc[2 * j].y, __uint_as_float(r.t[j].x & 0xffff0000u));
      acc[2 * j].z = __fadd_rn(acc[2 * j].z, __uint_as_float(r.t[j].y << 16));
      acc[2 * j].w = __fadd_rn(acc[2 * j].w, __uint_as_float(r.t[j].y & 0xffff0000u));
      acc[2 * j + 1].x = __fadd_rn(acc[2 * j + 1].x, __uint_as_float(r.t[j].z << 16));
      acc[2 * j + 1].y = __fadd_rn(acc[2 * j + 1].y, __uint_as_float(r.t[j].z & 0xffff0000u));
      acc[2 * j + 1].z = __fadd_rn(acc[2 * j + 1].z, __uint_as_float(r.t[j].w << 16));
      acc[2 * j + 1].w = __fadd_rn(acc[2 * j + 1].w, __uint_as_float(r.t[j].w & 0xffff0000u));
    }
  }
}

template <int NV4, bool BF16>
__device__ __forceinline__ void store_mean(float* out, int64_t out_stride, int64_t i, int lane, float4 (&acc)[NV4],
                                           int count) {
  // reference: counter[counter == 0] = 1e-6; sum / counter (models/affinity_module.py:435-436)
  float c = count ? (float)count : 1e-6f;
  float* o = out + i * out_stride;
  if constexpr (!BF16) {
#pragma unroll
    for (int j = 0; j < NV4; ++j) {
      float4 r = make_float4(__fdiv_rn(acc[j].x, c), __fdiv_rn(acc[j].y, c), __fdiv_rn(acc[j].z, c),
                             __fdiv_rn(acc[j].w, c));
      ((float4*)o)[j * 32 + lane] = r;
    }
  } else {
#pragma unroll
    for (int j = 0; j < NV4 / 2; ++j) {
      float4 r0 = make_float4(__fdiv_rn(acc[2 * j].x, c), __fdiv_rn(acc[2 * j].y, c), __fdiv_rn(acc[2 * j].z, c),
                              __fdiv_rn(acc[2 * j].w, c));
      float4 r1 = make_float4(__fdiv_rn(acc[2 * j + 1].x, c), __fdiv_rn(acc[2 * j + 1].y, c),
                              __fdiv_rn(acc[2 * j + 1].z, c), __fdiv_rn(acc[2 * j + 1].w, c));
      ((float4*)o)[(j * 32 + lane) * 2] = r0;
      ((float4*)o)[(j * 32 + lane) * 2 + 1] = r1;
    }
  }
}

template <typename CT, int DEPTH_DTYPE, int NV4, bool BF16, bool FROM_PIX>
__global__ void __launch_bounds__(128, 5) lift_mean_kernel(const void* __restrict__ coords, int64_t n,
                                                            const double* __restrict__ w2c,
                                                            const double* __restrict__ intr,
                                                            const void* __restrict__ depth, int V, int H, int W, int cut,
                                                            double thr, const uint8_t* __restrict__ keep,
                                                            const int32_t* __restrict__ pix_in,
                                                            const void* __restrict__ feat, int D, float* __restrict__ out,
                                                            int64_t out_stride, uint8_t* __restrict__ seen,
                                                            const uint8_t* __restrict__ idmap, int Q) {
  extern __shared__ double smem_d[];
  double* cams = smem_d;   // field-major: cams[j * V + v]
  uint8_t* skeep = (uint8_t*)(cams + (FROM_PIX ? 0 : 16 * V));
  if (!FROM_PIX) load_cams_soa(cams, w2c, intr, V);
  for (int i = threadIdx.x; i < V; i += blockDim.x) skeep[i] = keep ? keep[i] : 1;
  __syncthreads();
  const float thr32 = (float)thr;
  const int lane = threadIdx.x & 31;
  const int64_t warps_total = (int64_t)gridDim.x * (blockDim.x >> 5);
  const size_t hw = (size_t)H * W;
  // rows in flight per batch: 32 registers' worth of 128-bit loads per lane (bf16 D = 512: 4 rows x 2 loads)
  constexpr int NL = RowRegs<NV4, BF16>::NL;
  constexpr int RB = 8 / NL > 4 ? 4 : (8 / NL < 1 ? 1 : 8 / NL);
  for (int64_t i = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); i < n; i += warps_total) {
    double x = 0, y = 0, z = 0;
    if (!FROM_PIX) load_xyz<CT>(coords, i, x, y, z);
    float4 acc[NV4];
#pragma unroll
    for (int j = 0; j < NV4; ++j) acc[j] = make_float4(0.f, 0.f, 0.f, 0.f);
    int count = 0;
    for (int v0 = 0; v0 < V; v0 += 64) {
      // ---- which of the chunk's 64 views see the point, and which feature row each one reads
      int p[2];
      uint32_t ri[2];
#pragma unroll
      for (int h = 0; h < 2; ++h) {
        const int v = v0 + 32 * h + lane;
        p[h] = -1;
        if (v < V && skeep[v]) {
          if (FROM_PIX) {
            p[h] = __ldg(pix_in + (size_t)v * n + i);
          } else {
            int row, col;
            double pz;
            if (project_point(cam_from_soa(cams, V, v), x, y, z, H, W, cut, pz, row, col) &&
                depth_test<DEPTH_DTYPE>(depth, (size_t)v * hw + (size_t)row * W + col, pz, thr, thr32))
              p[h] = row * W + col;
          }
        }
      }
#pragma unroll
      for (int h = 0; h < 2; ++h) {
        const int v = v0 + 32 * h + lane;
        // compact 2-D form: the pixel's row is table[v, idmap[v, pixel]] instead of feat[v, pixel]
        ri[h] = 0;
        if (p[h] >= 0)
          ri[h] = idmap ? (uint32_t)(v * Q + min((int)__ldg(idmap + (size_t)v * hw + p[h]), Q - 1))
                        : (uint32_t)((size_t)v * hw + (size_t)p[h]);
      }
      unsigned long long mask = (unsigned long long)__ballot_sync(0xffffffffu, p[0] >= 0) |
                                ((unsigned long long)__ballot_sync(0xffffffffu, p[1] >= 0) << 32);
      count += __popcll(mask);
      // ---- ascending view order, RB rows in flight per batch
      while (mask) {
        RowRegs<NV4, BF16> rows[RB];
        bool ok[RB];
#pragma unroll
        for (int k = 0; k < RB; ++k) {
          ok[k] = mask != 0;
          const int src = ok[k] ? __ffsll((long long)mask) - 1 : 0;
          if (ok[k]) mask &= mask - 1;
          const uint32_t a = __shfl_sync(0xffffffffu, ri[0], src & 31);
          const uint32_t b2 = __shfl_sync(0xffffffffu, ri[1], src & 31);
          if (ok[k]) load_row<NV4, BF16>(feat, (size_t)(src < 32 ? a : b2), D, lane, rows[k]);
        }
#pragma unroll
        for (int k = 0; k < RB; ++k)
          if (ok[k]) add_row<NV4, BF16>(rows[k], acc);
      }
    }
    store_mean<NV4, BF16>(out, out_stride, i, lane, acc, count);
    if (lane == 0) seen[i] = count > 0;
  }
}

template <typename CT, int DD, bool FROM_PIX>
static int launch_lift(const void* coords, int64_t n, const double* w2c, const double* intr, const void* depth, int V,
                       int H, int W, int cut, double thr, const uint8_t* keep, const int32_t* pix, const void* feat,
                       int feat_dtype, int D, float* out, int64_t out_stride, uint8_t* seen, cudaStream_t st,
                       const uint8_t* idmap = nullptr, int Q = 0) {
  size_t smem = (FROM_PIX ? 0 : sizeof(Cam) * V) + V;
  int64_t warps = n;
  int64_t blocks = (warps + 3) / 4;
  int64_t cap = (int64_t)GP_SM_COUNT * 5 * 16;  // grid-stride beyond 16 waves of 5 CTAs (4 warps, ~96 registers) per SM
  unsigned grid = (unsigned)(blocks < cap ? blocks : cap);
  int nv4 = D / 128;
#define LIFT(NV4, BF)                                                                                        \
  lift_mean_kernel<CT, DD, NV4, BF, FROM_PIX><<<grid, 128, smem, st>>>(coords, n, w2c, intr, depth, V, H, W, \
                                                                       cut, thr, keep, pix, feat, D, out,    \
                                                                       out_stride, seen, idmap, Q)
  if (feat_dtype == GP_F32) {
    switch (nv4) {
      case 1: LIFT(1, false); break;
      case 2: LIFT(2, false); break;
      case 4: LIFT(4, false); break;
      case 6: LIFT(6, false); break;
      case 8: LIFT(8, false); break;
      default: gp_set_error("lift: unsupported D=%d (fp32 needs D in {128,256,512,768,1024})", D); return GP_ERR_ARG;
    }
  } else {
    switch (nv4) {
      case 2: LIFT(2, true); break;
      case 4: LIFT(4, true); break;
      case 6: LIFT(6, true); break;
      case 8: LIFT(8, true); break;
      default: gp_set_error("lift: unsupported D=%d (bf16 needs D in {256,512,768,1024})", D); return GP_ERR_ARG;
    }
  }
#undef LIFT
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) {
    gp_set_error("lift kernel launch failed: %s", cudaGetErrorString(e));
    return GP_ERR_CUDA;
  }
  return GP_OK;
}

static int check_lift_common(int64_t n, int V, int H, int W, int feat_dtype, int D, int64_t out_stride,
                             const float* out) {
  GP_CHECK_ARG(n >= 0 && V >= 0 && H > 0 && W > 0, "lift: bad sizes");
  GP_CHECK_ARG(V <= GP_MAX_VIEWS_SMEM, "lift: too many views");
  GP_CHECK_ARG(feat_dtype == GP_F32 || feat_dtype == GP_BF16, "lift: feature dtype must be f32 or bf16");
  GP_CHECK_ARG(D > 0 && D % 128 == 0 && D <= 1024, "lift: D must be a multiple of 128, <= 1024");
  GP_CHECK_ARG(out_stride >= D && out_stride % 4 == 0 && ((uintptr_t)out % 16) == 0,
               "lift: out must be 16-byte aligned with stride %% 4 == 0");
  return GP_OK;
}

extern "C" int gp_lift_mean_fused(const void* coords, int coords_dtype, int64_t n, const double* w2c,
                                  const double* intr, const void* depth, int depth_dtype, int V, int H, int W,
                                  int cut, double vis_thres, const uint8_t* keep, const void* feat, int feat_dtype,
                                  int D, float* out, int64_t out_stride, uint8_t* seen, gp_stream_t stream) {
  int rc = check_lift_common(n, V, H, W, feat_dtype, D, out_stride, out);
  if (rc) return rc;
  GP_CHECK_ARG(coords_dtype == GP_F32 || coords_dtype == GP_F64, "lift: coords dtype");
  GP_CHECK_ARG((depth == nullptr) == (depth_dtype == 0), "lift: depth pointer / dtype mismatch");
  if (n == 0) return GP_OK;
  cudaStream_t st = (cudaStream_t)stream;
#define GO(CT, DD)                                                                                              \
  return launch_lift<CT, DD, false>(coords, n, w2c, intr, depth, V, H, W, cut, vis_thres, keep, nullptr, feat, \
                                    feat_dtype, D, out, out_stride, seen, st)
  if (coords_dtype == GP_F64) {
    if (depth_dtype == 0) GO(double, 0);
    if (depth_dtype == GP_F64) GO(double, GP_F64);
    if (depth_dtype == GP_F32) GO(double, GP_F32);
  } else {
    if (depth_dtype == 0) GO(float, 0);
    if (depth_dtype == GP_F64) GO(float, GP_F64);
    if (depth_dtype == GP_F32) GO(float, GP_F32);
  }
#undef GO
  gp_set_error("lift: depth dtype");
  return GP_ERR_ARG;
}

extern "C" int gp_lift_mean_idmap(const void* coords, int coords_dtype, int64_t n, const double* w2c,
                                  const double* intr, const void* depth, int depth_dtype, int V, int H, int W,
                                  int cut, double vis_thres, const uint8_t* keep, const uint8_t* idmap,
                                  const void* table, int table_dtype, int Q, int D, float* out, int64_t out_stride,
                                  uint8_t* seen, gp_stream_t stream) {
  int rc = check_lift_common(n, V, H, W, table_dtype, D, out_stride, out);
  if (rc) return rc;
  GP_CHECK_ARG(coords_dtype == GP_F32 || coords_dtype == GP_F64, "lift: coords dtype");
  GP_CHECK_ARG((depth == nullptr) == (depth_dtype == 0), "lift: depth pointer / dtype mismatch");
  GP_CHECK_ARG(idmap != nullptr && Q >= 1 && Q <= 256, "gp_lift_mean_idmap: idmap / Q");
  if (n == 0) return GP_OK;
  cudaStream_t st = (cudaStream_t)stream;
#define GO(CT, DD)                                                                                               \
  return launch_lift<CT, DD, false>(coords, n, w2c, intr, depth, V, H, W, cut, vis_thres, keep, nullptr, table, \
                                    table_dtype, D, out, out_stride, seen, st, idmap, Q)
  if (coords_dtype == GP_F64) {
    if (depth_dtype == 0) GO(double, 0);
    if (depth_dtype == GP_F64) GO(double, GP_F64);
    if (depth_dtype == GP_F32) GO(double, GP_F32);
  } else {
    if (depth_dtype == 0) GO(float, 0);
    if (depth_dtype == GP_F64) GO(float, GP_F64);
    if (depth_dtype == GP_F32) GO(float, GP_F32);
  }
#undef GO
  gp_set_error("lift: depth dtype");
  return GP_ERR_ARG;
}

extern "C" int gp_lift_mean_pix(const int32_t* pix, int64_t n, int V, int H, int W, const uint8_t* keep,
                                const void* feat, int feat_dtype, int D, float* out, int64_t out_stride,
                                uint8_t* seen, gp_stream_t stream) {
  int rc = check_lift_common(n, V, H, W, feat_dtype, D, out_stride, out);
  if (rc) return rc;
  if (n == 0) return GP_OK;
  return launch_lift<float, 0, true>(nullptr, n, nullptr, nullptr, nullptr, V, H, W, 0, 0.0, keep, pix, feat,
                                     feat_dtype, D, out, out_stride, seen, (cudaStream_t)stream);
}

// ------------------------------------------------------------------ gp_expand_idmap
// Dense channels-last map from the (id map, embedding table) form: feat[v,h,w,:] = table[v, id[v,h,w], :].
// Stand-in for the 2-D VLM's per-pixel output (out of scope); also the A5 "mask embedding" lookup
// (reference models/affinity_module.py:584-592) materialised per pixel.
template <bool BF16>
__global__ void __launch_bounds__(256) expand_idmap_kernel(const uint8_t* __restrict__ idmap,
                                                            const float* __restrict__ table, int64_t pixels_per_view,
                                                            int64_t total_pixels, int Q, int D,
                                                            void* __restrict__ feat) {
  const int chunks = D >> 3;  // 8 channels per thread
  const int64_t total = total_pixels * chunks;
  for (int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; t < total; t += (int64_t)gridDim.x * blockDim.x) {
    int64_t p = t / chunks;
    int c = (int)(t - p * chunks) << 3;
    int v = (int)(p / pixels_per_view);
    int q = idmap[p];
    const float4* src = (const float4*)(table + ((int64_t)v * Q + q) * D + c);
    float4 a = __ldg(src), b = __ldg(src + 1);
    if (BF16) {
      __nv_bfloat162 h0 = __floats2bfloat162_rn(a.x, a.y), h1 = __floats2bfloat162_rn(a.z, a.w);
      __nv_bfloat162 h2 = __floats2bfloat162_rn(b.x, b.y), h3 = __floats2bfloat162_rn(b.z, b.w);
      uint4 o;
      o.x = *(unsigned*)&h0;
      o.y = *(unsigned*)&h1;
      o.z = *(unsigned*)&h2;
      o.w = *(unsigned*)&h3;
      *(uint4*)((__nv_bfloat16*)feat + p * D + c) = o;
    } else {
      float4* dst = (float4*)((float*)feat + p * D + c);
      dst[0] = a;
      dst[1] = b;
    }
  }
}

extern "C" int gp_expand_idmap(const uint8_t* idmap, const float* table, int V, int H, int W, int Q, int D,
                               void* feat, int feat_dtype, gp_stream_t stream) {
  GP_CHECK_ARG(V >= 0 && H > 0 && W > 0 && Q > 0 && Q <= 256 && D > 0 && D % 8 == 0, "gp_expand_idmap: bad sizes");
  GP_CHECK_ARG(feat_dtype == GP_F32 || feat_dtype == GP_BF16, "gp_expand_idmap: feature dtype");
  GP_CHECK_ARG(((uintptr_t)table % 16) == 0 && ((uintptr_t)feat % 16) == 0, "gp_expand_idmap: alignment");
  int64_t ppv = (int64_t)H * W, total = ppv * V;
  if (total == 0) return GP_OK;
  int64_t blocks = (total * (D / 8) + 255) / 256;
  int64_t cap = (int64_t)GP_SM_COUNT * 8 * 8;
  unsigned grid = (unsigned)(blocks < cap ? blocks : cap);
  if (feat_dtype == GP_BF16)
    expand_idmap_kernel<true><<<grid, 256, 0, (cudaStream_t)stream>>>(idmap, table, ppv, total, Q, D, feat);
  else
    expand_idmap_kernel<false><<<grid, 256, 0, (cudaStream_t)stream>>>(idmap, table, ppv, total, Q, D, feat);
  GP_LAUNCH_CHECK();
  return GP_OK;
}
