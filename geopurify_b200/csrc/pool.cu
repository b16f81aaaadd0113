// Affinity softmax and Geometry-Guided Pooling (v1: one warp per voxel row, gather through L1/L2).
// Replaces reference models/affinity_module.py:1559-1572 (cosine affinities, softmax(20 x)) and
// :1575-1587 (sparse COO A, F <- A F repeated T times through cuSPARSE).
#include "common.cuh"

// lane l ends up with the sum over lanes of v[l]  (31 shuffles for 32 values)
__device__ __forceinline__ float warp_reduce_scatter32(float (&v)[32], int lane) {
#pragma unroll
  for (int step = 16; step >= 1; step >>= 1) {
    const bool hi = (lane & step) != 0;
#pragma unroll
    for (int i = 0; i < step; ++i) {
      float send = hi ? v[i] : v[i + step];
      float keep = hi ? v[i + step] : v[i];
      v[i] = keep + __shfl_xor_sync(0xffffffffu, send, step);
    }
  }
  return v[0];
}

// E = 128 * NE4 ... generic: lane holds NE4 float4 of the centre embedding
template <int NE4>
__global__ void __launch_bounds__(256) affinity_softmax_kernel(const float* __restrict__ embed,
                                                                const int32_t* __restrict__ nbr, int64_t m, int K,
                                                                float sharpen, float* __restrict__ weights) {
  const int lane = threadIdx.x & 31;
  const int64_t s = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (s >= m) return;
  const int E4 = NE4 * 32;
  float4 c[NE4];
#pragma unroll
  for (int j = 0; j < NE4; ++j) c[j] = __ldg((const float4*)embed + s * E4 + j * 32 + lane);
  float a[4];
  const int groups = (K + 31) >> 5;
#pragma unroll
  for (int g = 0; g < 4; ++g) {
    a[g] = -INFINITY;
    if (g < groups) {
      int k = g * 32 + lane;
      int my = k < K ? __ldg(nbr + s * K + k) : -1;
      float part[32];
#pragma unroll
      for (int kk = 0; kk < 32; ++kk) {
        int idx = __shfl_sync(0xffffffffu, my, kk);
        float d = 0.f;
        if (idx >= 0) {
#pragma unroll
          for (int j = 0; j < NE4; ++j) {
            float4 e = __ldg((const float4*)embed + (int64_t)idx * E4 + j * 32 + lane);
            d = fmaf(c[j].x, e.x, d);
            d = fmaf(c[j].y, e.y, d);
            d = fmaf(c[j].z, e.z, d);
            d = fmaf(c[j].w, e.w, d);
          }
        }
        part[kk] = d;
      }
      float dot = warp_reduce_scatter32(part, lane);
      if (k < K) a[g] = dot * sharpen;
    }
  }
  float mx = warp_max(fmaxf(fmaxf(a[0], a[1]), fmaxf(a[2], a[3])));
  float ex[4], sum = 0.f;
#pragma unroll
  for (int g = 0; g < 4; ++g) {
    ex[g] = (g * 32 + lane < K) ? expf(a[g] - mx) : 0.f;
    sum += ex[g];
  }
  sum = warp_sum(sum);
#pragma unroll
  for (int g = 0; g < 4; ++g)
    if (g * 32 + lane < K) weights[s * K + g * 32 + lane] = ex[g] / sum;
}

extern "C" int gp_affinity_softmax(const float* embed, int E, const int32_t* nbr_sorted, int64_t m, int K,
                                   float sharpen, float* weights, gp_stream_t stream) {
  GP_CHECK_ARG(m >= 0 && K >= 1 && K <= 128, "gp_affinity_softmax: K=%d outside [1,128]", K);
  GP_CHECK_ARG(E == 128 || E == 256, "gp_affinity_softmax: E=%d must be 128 or 256", E);
  GP_CHECK_ARG(((uintptr_t)embed % 16) == 0, "gp_affinity_softmax: embed must be 16-byte aligned");
  if (m == 0) return GP_OK;
  unsigned grid = (unsigned)((m + 7) / 8);
  cudaStream_t st = (cudaStream_t)stream;
  if (E == 128)
    affinity_softmax_kernel<1><<<grid, 256, 0, st>>>(embed, nbr_sorted, m, K, sharpen, weights);
  else
    affinity_softmax_kernel<2><<<grid, 256, 0, st>>>(embed, nbr_sorted, m, K, sharpen, weights);
  GP_LAUNCH_CHECK();
  return GP_OK;
}

// ------------------------------------------------------------------ pooling, one application
template <int NV4>
__global__ void __launch_bounds__(256) pool_step_kernel(const float* __restrict__ fin, int64_t in_stride4,
                                                         const int32_t* __restrict__ nbr,
                                                         const float* __restrict__ weights, int64_t m, int K,
                                                         float* __restrict__ fout) {
  const int lane = threadIdx.x & 31;
  const int64_t s = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (s >= m) return;
  float4 acc[NV4];
#pragma unroll
  for (int j = 0; j < NV4; ++j) acc[j] = make_float4(0.f, 0.f, 0.f, 0.f);
  for (int k0 = 0; k0 < K; k0 += 32) {
    int k = k0 + lane;
    int my_idx = k < K ? __ldg(nbr + s * K + k) : 0;
    float my_w = k < K ? __ldg(weights + s * K + k) : 0.f;
    const int kn = min(32, K - k0);
#pragma unroll 4
    for (int kk = 0; kk < kn; ++kk) {
      int idx = __shfl_sync(0xffffffffu, my_idx, kk);
      float w = __shfl_sync(0xffffffffu, my_w, kk);
      const float4* row = (const float4*)fin + (int64_t)idx * in_stride4;
#pragma unroll
      for (int j = 0; j < NV4; ++j) {
        float4 t = __ldg(row + j * 32 + lane);
        acc[j].x = fmaf(w, t.x, acc[j].x);
        acc[j].y = fmaf(w, t.y, acc[j].y);
        acc[j].z = fmaf(w, t.z, acc[j].z);
        acc[j].w = fmaf(w, t.w, acc[j].w);
      }
    }
  }
  float4* o = (float4*)fout + s * (NV4 * 32);
#pragma unroll
  for (int j = 0; j < NV4; ++j) o[j * 32 + lane] = acc[j];
}

extern "C" int gp_pool(const float* f0, int64_t stride, int D, const int32_t* nbr_sorted, const float* weights,
                       int64_t m, int K, int T, float* buf_a, float* buf_b, float** out_host, gp_stream_t stream) {
  GP_CHECK_ARG(m >= 0 && K >= 1 && K <= 128 && T >= 0, "gp_pool: bad sizes");
  GP_CHECK_ARG(D > 0 && D % 128 == 0 && D <= 1024, "gp_pool: D must be a multiple of 128, <= 1024");
  GP_CHECK_ARG(stride % 4 == 0 && stride >= D, "gp_pool: stride must be a multiple of 4 and >= D");
  GP_CHECK_ARG(((uintptr_t)f0 % 16) == 0 && ((uintptr_t)buf_a % 16) == 0 && ((uintptr_t)buf_b % 16) == 0,
               "gp_pool: buffers must be 16-byte aligned");
  GP_CHECK_ARG(T == 0 || (buf_a && buf_b), "gp_pool: ping-pong buffers required");
  cudaStream_t st = (cudaStream_t)stream;
  const float* cur = f0;
  int64_t cur_stride4 = stride / 4;
  float* bufs[2] = {buf_a, buf_b};
  if (T == 0 || m == 0) {
    if (out_host) *out_host = (float*)f0;
    return GP_OK;
  }
  unsigned grid = (unsigned)((m + 7) / 8);
  for (int t = 0; t < T; ++t) {
    float* dst = bufs[t & 1];
    switch (D / 128) {
      case 1: pool_step_kernel<1><<<grid, 256, 0, st>>>(cur, cur_stride4, nbr_sorted, weights, m, K, dst); break;
      case 2: pool_step_kernel<2><<<grid, 256, 0, st>>>(cur, cur_stride4, nbr_sorted, weights, m, K, dst); break;
      case 4: pool_step_kernel<4><<<grid, 256, 0, st>>>(cur, cur_stride4, nbr_sorted, weights, m, K, dst); break;
      case 8: pool_step_kernel<8><<<grid, 256, 0, st>>>(cur, cur_stride4, nbr_sorted, weights, m, K, dst); break;
      default: gp_set_error("gp_pool: unsupported D=%d", D); return GP_ERR_ARG;
    }
    GP_LAUNCH_CHECK();
    cur = dst;
    cur_stride4 = D / 4;
  }
  if (out_host) *out_host = (float*)cur;
  return GP_OK;
}
