// Student affinity network forward (v1: fp32 SIMT gather-GEMM per kernel offset).
// Replaces AffinityPredictor.forward on MinkowskiEngine (reference models/affinity_module.py:33-72, 1541-1547):
// out[i] = sum over occupied offsets o (ascending) of W_o^T in[nbr(i,o)]; BN(eval) / ReLU / residual epilogues.
#include <cub/device/device_scan.cuh>
#include "common.cuh"

#define ST_TM 128
#define ST_TN 128
#define ST_TK 16

struct StudentLayout {
  size_t off_flags, off_pos, off_pair_out, off_pair_in, off_act[3], off_cub, cub_bytes, total;
  int hidden;
};

static bool student_layout(const gp_student_desc* d, int64_t m, StudentLayout& L) {
  if (!d || d->n_layers < 1 || d->n_layers > 16 || m < 0) return false;
  int hidden = 0;
  for (int l = 0; l + 1 < d->n_layers; ++l) hidden = d->cout[l] > hidden ? d->cout[l] : hidden;
  L.hidden = hidden;
  size_t off = 0;
  auto take = [&](size_t bytes) {
    size_t o = off;
    off += gp_align_up(bytes);
    return o;
  };
  size_t tot = (size_t)27 * m;
  L.off_flags = take(sizeof(int32_t) * (tot + 1));
  L.off_pos = take(sizeof(int32_t) * (tot + 1));
  L.off_pair_out = take(sizeof(int32_t) * tot);
  L.off_pair_in = take(sizeof(int32_t) * tot);
  for (int i = 0; i < 3; ++i) L.off_act[i] = take(sizeof(float) * (size_t)m * (hidden > 0 ? hidden : 1));
  L.cub_bytes = 0;
  cub::DeviceScan::ExclusiveSum(nullptr, L.cub_bytes, (const int32_t*)nullptr, (int32_t*)nullptr, (int)(tot + 1));
  L.off_cub = take(L.cub_bytes + 256);
  L.total = off;
  return true;
}

extern "C" size_t gp_student_workspace_bytes(const gp_student_desc* desc_host, int64_t m) {
  StudentLayout L;
  if (!student_layout(desc_host, m, L)) return 0;
  return L.total;
}

// flags in offset-major order: flags[o * m + s]
__global__ void pair_flags_kernel(const int32_t* __restrict__ nbr27, int64_t m, int32_t* __restrict__ flags) {
  int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t > 27 * m) return;
  if (t == 27 * m) {
    flags[t] = 0;
    return;
  }
  int o = (int)(t / m);
  int64_t s = t % m;
  flags[t] = nbr27[s * 27 + o] >= 0;
}

__global__ void pair_fill_kernel(const int32_t* __restrict__ nbr27, const int32_t* __restrict__ pos, int64_t m,
                                 int32_t* __restrict__ pair_out, int32_t* __restrict__ pair_in) {
  int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= 27 * m) return;
  int o = (int)(t / m);
  int64_t s = t % m;
  int nb = nbr27[s * 27 + o];
  if (nb >= 0) {
    int p = pos[t];
    pair_out[p] = (int32_t)s;
    pair_in[p] = nb;
  }
}

// C[pair_out[p], n0:n0+128] += A[pair_in[p], :] @ W[:, n0:n0+128] for the pairs of one offset.
// pos_lo / pos_hi point at the scan entries that bracket the offset's pair range (device side, no host sync).
__global__ void __launch_bounds__(256) gather_gemm_kernel(const float* __restrict__ A, int64_t lda, int cin,
                                                           const float* __restrict__ W, int cout,
                                                           const int32_t* __restrict__ pair_out,
                                                           const int32_t* __restrict__ pair_in,
                                                           const int32_t* __restrict__ pos_lo,
                                                           const int32_t* __restrict__ pos_hi, int64_t m_identity,
                                                           float* __restrict__ C, int64_t ldc, int accumulate) {
  __shared__ __align__(16) float As[ST_TK][ST_TM + 4];
  __shared__ __align__(16) float Bs[ST_TK][ST_TN];
  __shared__ int s_in[ST_TM], s_out[ST_TM];
  int64_t p_begin, p_end;
  if (pair_out) {
    p_begin = *pos_lo;
    p_end = *pos_hi;
  } else {
    p_begin = 0;
    p_end = m_identity;
  }
  const int64_t p0 = p_begin + (int64_t)blockIdx.x * ST_TM;
  if (p0 >= p_end) return;
  const int n0 = blockIdx.y * ST_TN;
  const int tid = threadIdx.x;
  if (tid < ST_TM) {
    int64_t p = p0 + tid;
    bool ok = p < p_end;
    s_in[tid] = ok ? (pair_in ? pair_in[p] : (int)p) : -1;
    s_out[tid] = ok ? (pair_out ? pair_out[p] : (int)p) : -1;
  }
  __syncthreads();
  const int tx = tid & 15, ty = tid >> 4;  // 16 x 16 threads, 8 x 8 outputs each
  float acc[8][8];
#pragma unroll
  for (int i = 0; i < 8; ++i)
#pragma unroll
    for (int j = 0; j < 8; ++j) acc[i][j] = 0.f;

  for (int k0 = 0; k0 < cin; k0 += ST_TK) {
#pragma unroll
    for (int r = 0; r < 2; ++r) {
      int idx = tid + r * 256;
      int row = idx >> 2, kq = idx & 3;
      int src = s_in[row];
      float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
      if (src >= 0) v = __ldg((const float4*)(A + (int64_t)src * lda + k0 + kq * 4));
      As[kq * 4 + 0][row] = v.x;
      As[kq * 4 + 1][row] = v.y;
      As[kq * 4 + 2][row] = v.z;
      As[kq * 4 + 3][row] = v.w;
    }
#pragma unroll
    for (int r = 0; r < 2; ++r) {
      int idx = tid + r * 256;
      int kr = idx >> 5, c4 = idx & 31;
      float4 v = __ldg((const float4*)(W + (int64_t)(k0 + kr) * cout + n0 + c4 * 4));
      *(float4*)&Bs[kr][c4 * 4] = v;
    }
    __syncthreads();
#pragma unroll
    for (int k = 0; k < ST_TK; ++k) {
      float a[8], b[8];
      *(float4*)&a[0] = *(const float4*)&As[k][ty * 4];
      *(float4*)&a[4] = *(const float4*)&As[k][64 + ty * 4];
      *(float4*)&b[0] = *(const float4*)&Bs[k][tx * 4];
      *(float4*)&b[4] = *(const float4*)&Bs[k][64 + tx * 4];
#pragma unroll
      for (int i = 0; i < 8; ++i)
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[i][j] = fmaf(a[i], b[j], acc[i][j]);
    }
    __syncthreads();
  }
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    int row = (i < 4) ? ty * 4 + i : 64 + ty * 4 + (i - 4);
    int dst = s_out[row];
    if (dst < 0) continue;
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      float4* cp = (float4*)(C + (int64_t)dst * ldc + n0 + h * 64 + tx * 4);
      float4 r = make_float4(acc[i][h * 4 + 0], acc[i][h * 4 + 1], acc[i][h * 4 + 2], acc[i][h * 4 + 3]);
      if (accumulate) {
        float4 old = *cp;
        r.x += old.x;
        r.y += old.y;
        r.z += old.z;
        r.w += old.w;
      }
      *cp = r;
    }
  }
}

// y = y * scale + shift (+ residual) (ReLU)
__global__ void epilogue_kernel(float* __restrict__ y, int64_t rows, int cout, const float* __restrict__ scale,
                                const float* __restrict__ shift, const float* __restrict__ residual, int relu) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  int c4n = cout >> 2;
  if (i >= rows * c4n) return;
  int c = (int)(i % c4n) * 4;
  float4 v = ((float4*)y)[i];
  if (scale) {
    float4 sc = *(const float4*)(scale + c), sh = *(const float4*)(shift + c);
    v.x = fmaf(v.x, sc.x, sh.x);
    v.y = fmaf(v.y, sc.y, sh.y);
    v.z = fmaf(v.z, sc.z, sh.z);
    v.w = fmaf(v.w, sc.w, sh.w);
  }
  if (residual) {
    float4 r = ((const float4*)residual)[i];
    v.x += r.x;
    v.y += r.y;
    v.z += r.z;
    v.w += r.w;
  }
  if (relu) {
    v.x = fmaxf(v.x, 0.f);
    v.y = fmaxf(v.y, 0.f);
    v.z = fmaxf(v.z, 0.f);
    v.w = fmaxf(v.w, 0.f);
  }
  ((float4*)y)[i] = v;
}

// F.normalize(p=2, dim=1): x / max(||x||, 1e-12), one warp per row
__global__ void l2_normalize_kernel(float* __restrict__ x, int64_t rows, int E) {
  int64_t r = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (r >= rows) return;
  int lane = threadIdx.x & 31;
  float ss = 0.f;
  for (int c = lane; c < E; c += 32) {
    float v = x[r * E + c];
    ss = fmaf(v, v, ss);
  }
  ss = warp_sum(ss);
  float inv = 1.f / fmaxf(sqrtf(ss), 1e-12f);
  for (int c = lane; c < E; c += 32) x[r * E + c] *= inv;
}

// shared with student_tc.cu
int gp_l2_normalize_rows(float* x, int64_t rows, int E, cudaStream_t st) {
  if (rows == 0) return GP_OK;
  l2_normalize_kernel<<<(unsigned)((rows + 7) / 8), 256, 0, st>>>(x, rows, E);
  GP_LAUNCH_CHECK();
  return GP_OK;
}

extern "C" int gp_student_forward(const gp_student_desc* d, const float* weights, const float* x, int64_t cin_stride,
                                  const int32_t* nbr27, int64_t m, float* embed, int normalize, void* workspace,
                                  size_t workspace_bytes, gp_stream_t stream) {
  StudentLayout L;
  GP_CHECK_ARG(student_layout(d, m, L), "gp_student_forward: bad descriptor");
  for (int l = 0; l < d->n_layers; ++l) {
    GP_CHECK_ARG(d->kv[l] == 27 || d->kv[l] == 1, "gp_student_forward: layer %d kernel volume %d", l, d->kv[l]);
    GP_CHECK_ARG(d->cin[l] % ST_TK == 0 && d->cout[l] % ST_TN == 0, "gp_student_forward: layer %d needs cin %% 16 == 0, cout %% 128 == 0", l);
    GP_CHECK_ARG(d->residual_from[l] == -1 || d->residual_from[l] == l - 2, "gp_student_forward: layer %d residual must come from layer l-2", l);
    if (l > 0) GP_CHECK_ARG(d->cin[l] == d->cout[l - 1], "gp_student_forward: layer %d cin != previous cout", l);
  }
  GP_CHECK_ARG(cin_stride >= d->cin[0] && cin_stride % 4 == 0, "gp_student_forward: input stride %lld < padded cin %d", (long long)cin_stride, d->cin[0]);
  if (workspace_bytes < L.total) {
    gp_set_error("gp_student_forward: workspace %zu < %zu", workspace_bytes, L.total);
    return GP_ERR_WORKSPACE;
  }
  if (m == 0) return GP_OK;
  cudaStream_t st = (cudaStream_t)stream;
  char* ws = (char*)workspace;
  int32_t* flags = (int32_t*)(ws + L.off_flags);
  int32_t* pos = (int32_t*)(ws + L.off_pos);
  int32_t* pair_out = (int32_t*)(ws + L.off_pair_out);
  int32_t* pair_in = (int32_t*)(ws + L.off_pair_in);
  const int64_t tot = 27 * m;
  pair_flags_kernel<<<(unsigned)((tot + 1 + 255) / 256), 256, 0, st>>>(nbr27, m, flags);
  GP_LAUNCH_CHECK();
  size_t cb = L.cub_bytes + 256;
  GP_CUDA(cub::DeviceScan::ExclusiveSum(ws + L.off_cub, cb, flags, pos, (int)(tot + 1), st));
  pair_fill_kernel<<<(unsigned)((tot + 255) / 256), 256, 0, st>>>(nbr27, pos, m, pair_out, pair_in);
  GP_LAUNCH_CHECK();

  const float* in = x;
  int64_t ld_in = cin_stride;
  for (int l = 0; l < d->n_layers; ++l) {
    const bool last = l == d->n_layers - 1;
    float* out = last ? embed : (float*)(ws + L.off_act[l % 3]);
    const int cout = d->cout[l], cin = d->cin[l];
    const float* Wl = weights + d->w_off[l];
    dim3 grid((unsigned)((m + ST_TM - 1) / ST_TM), (unsigned)(cout / ST_TN));
    if (d->kv[l] == 27) {
      GP_CUDA(cudaMemsetAsync(out, 0, sizeof(float) * (size_t)m * cout, st));
      for (int o = 0; o < 27; ++o) {
        gather_gemm_kernel<<<grid, 256, 0, st>>>(in, ld_in, cin, Wl + (int64_t)o * cin * cout, cout, pair_out,
                                                 pair_in, pos + (int64_t)o * m, pos + (int64_t)(o + 1) * m, 0, out,
                                                 cout, 1);
      }
    } else {
      gather_gemm_kernel<<<grid, 256, 0, st>>>(in, ld_in, cin, Wl, cout, nullptr, nullptr, nullptr, nullptr, m, out,
                                               cout, 0);
    }
    GP_LAUNCH_CHECK();
    const float* res = d->residual_from[l] >= 0 ? (const float*)(ws + L.off_act[(l - 2) % 3]) : nullptr;
    if (d->bn[l] || res || d->relu[l]) {
      int64_t n4 = m * (cout / 4);
      const float* sc = d->bn[l] ? weights + d->bn_off[l] : nullptr;
      epilogue_kernel<<<(unsigned)((n4 + 255) / 256), 256, 0, st>>>(out, m, cout, sc, sc ? sc + cout : nullptr, res,
                                                                    d->relu[l]);
      GP_LAUNCH_CHECK();
    }
    in = out;
    ld_in = cout;
  }
  if (normalize) {
    l2_normalize_kernel<<<(unsigned)((m + 7) / 8), 256, 0, st>>>(embed, m, d->cout[d->n_layers - 1]);
    GP_LAUNCH_CHECK();
  }
  return GP_OK;
}
