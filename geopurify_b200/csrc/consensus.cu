// A5: X-Decoder-style lifting -- mask-embedding gather, per-view hole fill, consensus top-3 fusion.
// Replaces reference models/affinity_module.py:584-592 (point feature = embedding of the mask covering its pixel),
// :601-625 (per-view KDTree fill of zero rows), :627-630 (normalise, per-view logits), :633-685 (regroup per point,
// mean logits -> consensus class, top-min(3, #views) views by their logit of that class, softmax-weighted sum).
//
// The 2-D side arrives in its compact form: idmap[v,h,w] = index of the mask covering the pixel, table[v,q,:] = that
// mask's embedding (an all-zero row = "no mask").  A point's per-view (feature, logits) pair is therefore one of
// V*Q per-MASK pairs, so normalisation and the logits GEMM run once per mask (V*Q*D*C MACs instead of P*D*C) and the
// per-point state that the reference keeps in a Python dict of tensors collapses to one mask id per (view, point).
#include "common.cuh"

namespace {

// one warp per (view, mask): zero-row flag, L2 normalisation, logits against the normalised text rows
__global__ void __launch_bounds__(256) mask_prepare_kernel(const float* __restrict__ table, int64_t rows, int D,
                                                           const float* __restrict__ textn, int C, float scale,
                                                           float* __restrict__ tn, float* __restrict__ mlog,
                                                           uint8_t* __restrict__ nonzero) {
  const int lane = threadIdx.x & 31;
  const int64_t r = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (r >= rows) return;
  const float* src = table + r * D;
  float s = 0.f, ss = 0.f;
  for (int j = lane; j < D; j += 32) {
    float v = src[j];
    s += v;
    ss += v * v;
  }
  s = warp_sum(s);
  ss = warp_sum(ss);
  const float inv = 1.f / fmaxf(sqrtf(ss), 1e-12f);  // F.normalize eps
  for (int j = lane; j < D; j += 32) tn[r * D + j] = src[j] * inv;
  if (lane == 0) nonzero[r] = s != 0.f;               // the reference tests sum(dim=1) == 0 (:606)
  for (int c = 0; c < C; ++c) {
    float a = 0.f;
    for (int j = lane; j < D; j += 32) a = fmaf(src[j] * inv, __ldg(textn + (int64_t)c * D + j), a);
    a = warp_sum(a);
    if (lane == 0) mlog[r * C + c] = scale * a;
  }
}

__global__ void text_normalize_kernel(const float* __restrict__ text, int C, int D, float* __restrict__ out) {
  const int lane = threadIdx.x & 31;
  const int c = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (c >= C) return;
  float ss = 0.f;
  for (int j = lane; j < D; j += 32) ss += text[(int64_t)c * D + j] * text[(int64_t)c * D + j];
  ss = warp_sum(ss);
  const float inv = 1.f / fmaxf(sqrtf(ss), 1e-12f);
  for (int j = lane; j < D; j += 32) out[(int64_t)c * D + j] = text[(int64_t)c * D + j] * inv;
}

// qid[v,p] = mask id under the point's pixel (-1: not visible / view dropped); state: 1 source, 0 query, 2 neither
__global__ void mask_id_kernel(const int32_t* __restrict__ pix, int64_t n, int V, int HW, const uint8_t* __restrict__ keep,
                               const uint8_t* __restrict__ idmap, const uint8_t* __restrict__ nonzero, int Q,
                               int32_t* __restrict__ qid, uint8_t* __restrict__ state) {
  const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= (int64_t)V * n) return;
  const int v = (int)(t / n);
  const int px = pix[t];
  int q = -1;
  uint8_t st = 2;
  if (px >= 0 && (!keep || keep[v])) {
    q = idmap[(int64_t)v * HW + px];
    if (q >= Q) q = Q - 1;
    st = nonzero[(int64_t)v * Q + q] ? 1 : 0;
  }
  qid[t] = q;
  state[t] = st;
}

__global__ void qid_fill_kernel(int32_t* __restrict__ qid_v, const uint8_t* __restrict__ state_v,
                                const int32_t* __restrict__ nn, int64_t n) {
  const int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= n || state_v[p] != 0) return;
  const int src = nn[p];
  if (src >= 0) qid_v[p] = qid_v[src];  // sources are never rewritten: in place is safe
}

// One warp per point.  NC = ceil(C / 32) classes per lane.
template <int NC>
__global__ void __launch_bounds__(256) consensus_fuse_kernel(const int32_t* __restrict__ qid, int64_t n, int V, int Q,
                                                             int C, int D, const float* __restrict__ tn,
                                                             const float* __restrict__ mlog, float* __restrict__ out,
                                                             int64_t out_stride, uint8_t* __restrict__ seen) {
  const int lane = threadIdx.x & 31;
  const int64_t p = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (p >= n) return;
  // ---- pass 1: sum of the per-view logits in ascending view order, number of views (:673-675)
  float sum[NC];
#pragma unroll
  for (int j = 0; j < NC; ++j) sum[j] = 0.f;
  int count = 0;
  for (int v0 = 0; v0 < V; v0 += 32) {
    const int v = v0 + lane;
    const int q = v < V ? __ldg(qid + (int64_t)v * n + p) : -1;
    unsigned live = __ballot_sync(0xffffffffu, q >= 0);
    count += __popc(live);
    while (live) {
      const int l = __ffs(live) - 1;
      live &= live - 1;
      const int qq = __shfl_sync(0xffffffffu, q, l);
      const float* row = mlog + ((int64_t)(v0 + l) * Q + qq) * C;
#pragma unroll
      for (int j = 0; j < NC; ++j)
        if (lane + 32 * j < C) sum[j] += __ldg(row + lane + 32 * j);
    }
  }
  if (lane == 0) seen[p] = count > 0;
  float* dst = out + p * out_stride;
  if (count == 0) {
    for (int j = lane * 4; j < D; j += 128) *(float4*)(dst + j) = make_float4(0.f, 0.f, 0.f, 0.f);
    return;
  }
  // ---- consensus class: argmax of the mean logits, lowest class index on ties (:676)
  float best = -INFINITY;
  int bestc = 0x7fffffff;
#pragma unroll
  for (int j = 0; j < NC; ++j) {
    const int c = lane + 32 * j;
    if (c < C) {
      const float a = sum[j] / (float)count;
      if (a > best || (a == best && c < bestc)) {
        best = a;
        bestc = c;
      }
    }
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const float ob = __shfl_xor_sync(0xffffffffu, best, o);
    const int oc = __shfl_xor_sync(0xffffffffu, bestc, o);
    if (ob > best || (ob == best && oc < bestc)) {
      best = ob;
      bestc = oc;
    }
  }
  // ---- top-min(3, count) views by their logit of the consensus class, earliest view on ties (:677-679)
  const int k = count < 3 ? count : 3;
  int sel_v[3] = {-1, -1, -1}, sel_q[3] = {0, 0, 0};
  float sel_s[3] = {0.f, 0.f, 0.f};
  for (int r = 0; r < k; ++r) {
    float bs = -INFINITY;
    int bv = 0x7fffffff, bq = 0;
    for (int v = lane; v < V; v += 32) {
      if (v == sel_v[0] || v == sel_v[1]) continue;
      const int q = __ldg(qid + (int64_t)v * n + p);
      if (q < 0) continue;
      const float s = __ldg(mlog + ((int64_t)v * Q + q) * C + bestc);
      if (bv == 0x7fffffff || s > bs) {  // ascending v inside a lane: ties keep the earlier view
        bs = s;
        bv = v;
        bq = q;
      }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      const float os = __shfl_xor_sync(0xffffffffu, bs, o);
      const int ov = __shfl_xor_sync(0xffffffffu, bv, o);
      const int oq = __shfl_xor_sync(0xffffffffu, bq, o);
      if (ov != 0x7fffffff && (bv == 0x7fffffff || os > bs || (os == bs && ov < bv))) {
        bs = os;
        bv = ov;
        bq = oq;
      }
    }
    sel_v[r] = bv;
    sel_q[r] = bq;
    sel_s[r] = bs;
  }
  // ---- softmax over the selected scores, weighted sum of the normalised embeddings in rank order (:682-683)
  float w[3] = {0.f, 0.f, 0.f};
  float den = 0.f;
  for (int r = 0; r < k; ++r) {
    w[r] = expf(sel_s[r] - sel_s[0]);  // sel_s[0] is the maximum
    den += w[r];
  }
  for (int r = 0; r < k; ++r) w[r] /= den;
  for (int j = lane * 4; j < D; j += 128) {
    float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
    for (int r = 0; r < k; ++r) {
      const float4 f = __ldg((const float4*)(tn + ((int64_t)sel_v[r] * Q + sel_q[r]) * D + j));
      acc.x += f.x * w[r];
      acc.y += f.y * w[r];
      acc.z += f.z * w[r];
      acc.w += f.w * w[r];
    }
    *(float4*)(dst + j) = acc;
  }
}

struct ConsLayout {
  size_t off_textn, off_tn, off_mlog, off_nz, off_qid, off_state, off_nn, off_nn1, nn1_bytes, total;
};

void cons_layout(int64_t n, int V, int Q, int D, int C, ConsLayout& L) {
  size_t off = 0;
  auto take = [&](size_t bytes) {
    size_t o = off;
    off += gp_align_up(bytes);
    return o;
  };
  L.off_textn = take(sizeof(float) * (size_t)C * D);
  L.off_tn = take(sizeof(float) * (size_t)V * Q * D);
  L.off_mlog = take(sizeof(float) * (size_t)V * Q * C);
  L.off_nz = take((size_t)V * Q);
  L.off_qid = take(sizeof(int32_t) * (size_t)V * n);
  L.off_state = take((size_t)V * n);
  L.off_nn = take(sizeof(int32_t) * (size_t)n);
  L.nn1_bytes = gp_nn1_fill_workspace_bytes(n);
  L.off_nn1 = take(L.nn1_bytes);
  L.total = off;
}

}  // namespace

extern "C" size_t gp_lift_consensus_workspace_bytes(int64_t n, int V, int Q, int D, int C) {
  if (n < 0 || V < 0 || Q < 1 || D < 1 || C < 1) return 0;
  ConsLayout L;
  cons_layout(n, V, Q, D, C, L);
  return L.total;
}

extern "C" int gp_lift_consensus(const int32_t* pix, int64_t n, int V, int H, int W, const uint8_t* keep,
                                 const uint8_t* idmap, const float* table, int Q, int D, const float* text, int C,
                                 float logit_scale, const float* xyz, float cell, float* out, int64_t out_stride,
                                 uint8_t* seen, void* workspace, size_t workspace_bytes, gp_stream_t stream) {
  GP_CHECK_ARG(n >= 0 && n < (1ll << 31) && V >= 0 && H > 0 && W > 0, "gp_lift_consensus: bad sizes");
  GP_CHECK_ARG(Q >= 1 && Q <= 256, "gp_lift_consensus: Q=%d outside [1,256] (idmap is uint8)", Q);
  GP_CHECK_ARG(D >= 4 && D % 4 == 0 && out_stride >= D && out_stride % 4 == 0, "gp_lift_consensus: D / stride");
  GP_CHECK_ARG(C >= 1 && C <= 256, "gp_lift_consensus: C=%d outside [1,256]", C);
  GP_CHECK_ARG(cell > 0.f, "gp_lift_consensus: cell must be > 0");
  ConsLayout L;
  cons_layout(n, V, Q, D, C, L);
  if (workspace_bytes < L.total) {
    gp_set_error("gp_lift_consensus: workspace %zu < %zu", workspace_bytes, L.total);
    return GP_ERR_WORKSPACE;
  }
  if (n == 0) return GP_OK;
  cudaStream_t st = (cudaStream_t)stream;
  char* ws = (char*)workspace;
  float* textn = (float*)(ws + L.off_textn);
  float* tn = (float*)(ws + L.off_tn);
  float* mlog = (float*)(ws + L.off_mlog);
  uint8_t* nz = (uint8_t*)(ws + L.off_nz);
  int32_t* qid = (int32_t*)(ws + L.off_qid);
  uint8_t* state = (uint8_t*)(ws + L.off_state);
  int32_t* nn = (int32_t*)(ws + L.off_nn);
  text_normalize_kernel<<<(C + 7) / 8, 256, 0, st>>>(text, C, D, textn);
  GP_LAUNCH_CHECK();
  if (V > 0) {
    const int64_t rows = (int64_t)V * Q;
    mask_prepare_kernel<<<(unsigned)((rows + 7) / 8), 256, 0, st>>>(table, rows, D, textn, C, logit_scale, tn, mlog, nz);
    GP_LAUNCH_CHECK();
    const int64_t tot = (int64_t)V * n;
    mask_id_kernel<<<(unsigned)((tot + 255) / 256), 256, 0, st>>>(pix, n, V, H * W, keep, idmap, nz, Q, qid, state);
    GP_LAUNCH_CHECK();
    for (int v = 0; v < V; ++v) {  // per-view hole fill among the points this view sees (:601-625)
      int rc = gp_nn1_fill(xyz, n, state + (int64_t)v * n, cell, nn, ws + L.off_nn1, L.nn1_bytes, stream);
      if (rc) return rc;
      qid_fill_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(qid + (int64_t)v * n, state + (int64_t)v * n, nn, n);
      GP_LAUNCH_CHECK();
    }
  }
  const unsigned grid = (unsigned)((n + 7) / 8);
#define FUSE(NC) consensus_fuse_kernel<NC><<<grid, 256, 0, st>>>(qid, n, V, Q, C, D, tn, mlog, out, out_stride, seen)
  if (C <= 32) FUSE(1);
  else if (C <= 64) FUSE(2);
  else if (C <= 128) FUSE(4);
  else FUSE(8);
#undef FUSE
  GP_LAUNCH_CHECK();
  return GP_OK;
}
