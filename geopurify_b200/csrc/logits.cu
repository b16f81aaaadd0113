// Text-embedding classification (v1: fp32 SIMT, one warp per row) and IoU histograms.
// Replaces reference run/validation.py:413-417 (normalise, logits = s f t^T, argmax) and
// util/util.py:160-177 (intersectionAndUnionGPU: ignore overwrite + three histc histograms).
#include "common.cuh"

__global__ void normalize_text_kernel(const float* __restrict__ text, int C, int D, float* __restrict__ out) {
  int c = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (c >= C) return;
  int lane = threadIdx.x & 31;
  float ss = 0.f;
  for (int j = lane; j < D; j += 32) {
    float v = text[(int64_t)c * D + j];
    ss = fmaf(v, v, ss);
  }
  ss = warp_sum(ss);
  float den = fmaxf(sqrtf(ss), 1e-12f);
  for (int j = lane; j < D; j += 32) out[(int64_t)c * D + j] = text[(int64_t)c * D + j] / den;
}

__device__ __forceinline__ float reduce_scatter32(float (&v)[32], int lane) {
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

template <int NV4>
__global__ void __launch_bounds__(256) logits_argmax_kernel(const float* __restrict__ feat, int64_t stride4,
                                                             int64_t rows, const float* __restrict__ text_n, int C,
                                                             float scale, int32_t* __restrict__ label,
                                                             uint8_t* __restrict__ zero, float* __restrict__ logits,
                                                             const uint8_t* __restrict__ only) {
  const int lane = threadIdx.x & 31;
  const int64_t r = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (r >= rows) return;
  if (only && !only[r]) return;  // fix-up pass of the tensor-core path: only the flagged (near-tie / zero) rows
  const int D4 = NV4 * 32;
  float4 f[NV4];
  float ss = 0.f;
#pragma unroll
  for (int j = 0; j < NV4; ++j) {
    f[j] = ldg_stream_f4((const float4*)feat + r * stride4 + j * 32 + lane);
    ss = fmaf(f[j].x, f[j].x, ss);
    ss = fmaf(f[j].y, f[j].y, ss);
    ss = fmaf(f[j].z, f[j].z, ss);
    ss = fmaf(f[j].w, f[j].w, ss);
  }
  ss = warp_sum(ss);
  const float den = fmaxf(sqrtf(ss), 1e-12f);
  float asum = 0.f;
#pragma unroll
  for (int j = 0; j < NV4; ++j) {
    f[j].x /= den;
    f[j].y /= den;
    f[j].z /= den;
    f[j].w /= den;
    asum += fabsf(f[j].x) + fabsf(f[j].y) + fabsf(f[j].z) + fabsf(f[j].w);
  }
  asum = warp_sum(asum);
  float best = -INFINITY;
  int besti = 0x7fffffff;
  for (int c0 = 0; c0 < C; c0 += 32) {
    float part[32];
#pragma unroll
    for (int cc = 0; cc < 32; ++cc) {
      float d = 0.f;
      if (c0 + cc < C) {
        const float4* t = (const float4*)text_n + (int64_t)(c0 + cc) * D4;
#pragma unroll
        for (int j = 0; j < NV4; ++j) {
          float4 tv = __ldg(t + j * 32 + lane);
          d = fmaf(f[j].x, tv.x, d);
          d = fmaf(f[j].y, tv.y, d);
          d = fmaf(f[j].z, tv.z, d);
          d = fmaf(f[j].w, tv.w, d);
        }
      }
      part[cc] = d;
    }
    float dot = reduce_scatter32(part, lane);
    int c = c0 + lane;
    if (c < C) {
      float lg = scale * dot;
      if (logits) logits[r * C + c] = lg;
      if (lg > best) {  // ascending c per lane: strict > keeps the lowest index
        best = lg;
        besti = c;
      }
    }
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    float ob = __shfl_xor_sync(0xffffffffu, best, o);
    int oi = __shfl_xor_sync(0xffffffffu, besti, o);
    if (ob > best || (ob == best && oi < besti)) {
      best = ob;
      besti = oi;
    }
  }
  if (lane == 0) {
    // all-NaN rows: torch.max returns the first NaN position; an all-zero row gives logits 0 -> class 0
    label[r] = besti == 0x7fffffff ? 0 : besti;
    if (zero) zero[r] = asum == 0.f;
  }
}

extern "C" size_t gp_logits_workspace_bytes(int C, int D) {
  if (C <= 0 || D <= 0) return 0;
  return gp_align_up(sizeof(float) * (size_t)C * D);
}

static int logits_argmax_impl(const float* feat, int64_t stride, int D, int64_t rows, const float* text, int C,
                              float logit_scale, int32_t* label, uint8_t* zero, float* logits_or_null,
                              void* workspace, size_t workspace_bytes, const uint8_t* only, gp_stream_t stream);

extern "C" int gp_logits_argmax(const float* feat, int64_t stride, int D, int64_t rows, const float* text, int C,
                                float logit_scale, int32_t* label, uint8_t* zero, float* logits_or_null,
                                void* workspace, size_t workspace_bytes, gp_stream_t stream) {
  return logits_argmax_impl(feat, stride, D, rows, text, C, logit_scale, label, zero, logits_or_null, workspace,
                            workspace_bytes, nullptr, stream);
}

// the same fp32 kernel restricted to the rows with only[r] != 0 (labels / zero flags of the other rows untouched)
int gp_logits_fix_rows(const float* feat, int64_t stride, int D, int64_t rows, const float* text, int C,
                       float logit_scale, int32_t* label, uint8_t* zero, const uint8_t* only, void* workspace,
                       size_t workspace_bytes, cudaStream_t st) {
  return logits_argmax_impl(feat, stride, D, rows, text, C, logit_scale, label, zero, nullptr, workspace,
                            workspace_bytes, only, (gp_stream_t)st);
}

static int logits_argmax_impl(const float* feat, int64_t stride, int D, int64_t rows, const float* text, int C,
                              float logit_scale, int32_t* label, uint8_t* zero, float* logits_or_null,
                              void* workspace, size_t workspace_bytes, const uint8_t* only, gp_stream_t stream) {
  GP_CHECK_ARG(rows >= 0 && C >= 1, "gp_logits_argmax: bad sizes");
  GP_CHECK_ARG(D > 0 && D % 128 == 0 && D <= 1024, "gp_logits_argmax: D must be a multiple of 128, <= 1024");
  GP_CHECK_ARG(stride % 4 == 0 && stride >= D && ((uintptr_t)feat % 16) == 0, "gp_logits_argmax: feat alignment");
  if (workspace_bytes < gp_logits_workspace_bytes(C, D)) {
    gp_set_error("gp_logits_argmax: workspace too small");
    return GP_ERR_WORKSPACE;
  }
  cudaStream_t st = (cudaStream_t)stream;
  float* tn = (float*)workspace;
  normalize_text_kernel<<<(C + 7) / 8, 256, 0, st>>>(text, C, D, tn);
  GP_LAUNCH_CHECK();
  if (rows == 0) return GP_OK;
  unsigned grid = (unsigned)((rows + 7) / 8);
#define GO(NV4) \
  logits_argmax_kernel<NV4><<<grid, 256, 0, st>>>(feat, stride / 4, rows, tn, C, logit_scale, label, zero, logits_or_null, only)
  switch (D / 128) {
    case 1: GO(1); break;
    case 2: GO(2); break;
    case 4: GO(4); break;
    case 8: GO(8); break;
    default: gp_set_error("gp_logits_argmax: unsupported D=%d", D); return GP_ERR_ARG;
  }
#undef GO
  GP_LAUNCH_CHECK();
  return GP_OK;
}

// ------------------------------------------------------------------ IoU histograms
// histc(bins=C, min=0, max=C-1) puts integer label j in bin j for 0 <= j <= C-1 and drops everything else.
__global__ void __launch_bounds__(256) iou_hist_kernel(const int32_t* __restrict__ label_rows,
                                                        const int32_t* __restrict__ index,
                                                        const int64_t* __restrict__ target, int64_t n,
                                                        const int64_t* __restrict__ ignore, int n_ignore, int C,
                                                        int64_t* __restrict__ pred_out,
                                                        unsigned long long* __restrict__ hist) {
  extern __shared__ int sh[];  // [3][C]: intersection, output, target
  for (int i = threadIdx.x; i < 3 * C; i += blockDim.x) sh[i] = 0;
  __syncthreads();
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    int64_t t = target[i];
    int64_t p = label_rows[index ? index[i] : i];
    for (int g = 0; g < n_ignore; ++g)
      if (t == ignore[g]) p = ignore[g];
    if (pred_out) pred_out[i] = p;
    if (p >= 0 && p < C) {
      atomicAdd(&sh[C + (int)p], 1);
      if (p == t) atomicAdd(&sh[(int)p], 1);
    }
    if (t >= 0 && t < C) atomicAdd(&sh[2 * C + (int)t], 1);
  }
  __syncthreads();
  for (int c = threadIdx.x; c < C; c += blockDim.x) {
    int in = sh[c], ou = sh[C + c], tg = sh[2 * C + c];
    if (in) atomicAdd(&hist[c], (unsigned long long)in);
    long long un = (long long)ou + tg - in;  // union = output + target - intersection
    if (un) atomicAdd(&hist[C + c], (unsigned long long)un);
    if (tg) atomicAdd(&hist[2 * C + c], (unsigned long long)tg);
  }
}

extern "C" int gp_iou_hist(const int32_t* label_rows, const int32_t* index, const int64_t* target, int64_t n,
                           const int64_t* ignore, int n_ignore, int C, int64_t* pred_out, int64_t* hist,
                           gp_stream_t stream) {
  GP_CHECK_ARG(n >= 0 && C >= 1 && C <= 4096 && n_ignore >= 0, "gp_iou_hist: bad sizes");
  if (n == 0) return GP_OK;
  int64_t blocks = (n + 255) / 256;
  unsigned grid = (unsigned)(blocks < GP_SM_COUNT * 8 ? blocks : GP_SM_COUNT * 8);
  iou_hist_kernel<<<grid, 256, sizeof(int) * 3 * C, (cudaStream_t)stream>>>(label_rows, index, target, n, ignore,
                                                                            n_ignore, C, pred_out,
                                                                            (unsigned long long*)hist);
  GP_LAUNCH_CHECK();
  return GP_OK;
}
