// A5: X-Decoder-style lifting -- mask-embedding gather, per-view hole fill, consensus top-3 fusion.
// Replaces reference models/affinity_module.py:584-592 (point feature = embedding of the mask covering its pixel),
// :601-625 (per-view KDTree fill of zero rows), :627-630 (normalise, per-view logits), :633-685 (regroup per point,
// mean logits -> consensus class, top-min(3, #views) views by their logit of that class, softmax-weighted sum).
//
// The 2-D side arrives in its compact form: idmap[v,h,w] = index of the mask covering the pixel, table[v,q,:] = that
// mask's embedding (an all-zero row = "no mask").  A point's per-view (feature, logits) pair is therefore one of
// V*Q per-MASK pairs, so normalisation and the logits GEMM run once per mask (V*Q*D*C MACs instead of P*D*C) and the
// per-point state that the reference keeps in a Python dict of tensors collapses to one mask id per (view, point).
#include <cub/device/device_scan.cuh>
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

// ---------------------------------------------------------------- per-view hole fill, all views in one pass
// Every view fills its zero-mask points from the nearest point of the SAME view that lies under a real mask
// (:601-625, a KDTree per view in the reference).  All V searches share one uniform grid over the scene (VG cells
// per axis); the sources of every view are binned by (view, cell) with a counting sort (order inside a cell does not
// matter: the search keeps the minimum of (d2, index)), then one kernel walks the [V,n] state array, and each warp
// serves the queries among its 32 entries one after the other: Chebyshev shells of cells around the query, cells
// pruned by their box distance, members scanned by the whole warp; float64 distances on the float32 coordinates,
// ties -> lowest point index, exactly as gp_nn1_fill.
#define VG 32
#define VG3 (VG * VG * VG)

struct ViewGrid {
  float mn[3];
  float cell, inv_cell;
  int dim[3];
  unsigned mn_bits[3], mx_bits[3];
};

__device__ __forceinline__ unsigned vg_f2ord(float f) {
  unsigned u = __float_as_uint(f);
  return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
__device__ __forceinline__ float vg_ord2f(unsigned u) {
  return __uint_as_float((u & 0x80000000u) ? (u & 0x7fffffffu) : ~u);
}

__global__ void vg_init_kernel(ViewGrid* h) {
  for (int a = 0; a < 3; ++a) {
    h->mn_bits[a] = 0xffffffffu;
    h->mx_bits[a] = 0u;
  }
}

__global__ void vg_bbox_kernel(const float* __restrict__ xyz, int64_t n, ViewGrid* h) {
  unsigned mn[3] = {0xffffffffu, 0xffffffffu, 0xffffffffu}, mx[3] = {0, 0, 0};
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
    for (int a = 0; a < 3; ++a) {
      unsigned o = vg_f2ord(xyz[3 * i + a]);
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

__global__ void vg_grid_kernel(ViewGrid* h, float cell) {
  float ext = 0.f;
  for (int a = 0; a < 3; ++a) {
    h->mn[a] = vg_ord2f(h->mn_bits[a]);
    ext = fmaxf(ext, vg_ord2f(h->mx_bits[a]) - h->mn[a]);
  }
  float c = fmaxf(cell, ext / (float)(VG - 1));
  if (!(c > 0.f)) c = 1.f;
  h->cell = c;
  h->inv_cell = 1.f / c;
  for (int a = 0; a < 3; ++a) {
    int d = (int)floorf((vg_ord2f(h->mx_bits[a]) - h->mn[a]) * h->inv_cell) + 1;
    h->dim[a] = max(1, min(d, VG));
  }
}

__device__ __forceinline__ int vg_cell_of(const ViewGrid* h, const float* __restrict__ xyz, int64_t p, int& cx, int& cy,
                                          int& cz) {
  cx = min(max((int)floorf((xyz[3 * p] - h->mn[0]) * h->inv_cell), 0), h->dim[0] - 1);
  cy = min(max((int)floorf((xyz[3 * p + 1] - h->mn[1]) * h->inv_cell), 0), h->dim[1] - 1);
  cz = min(max((int)floorf((xyz[3 * p + 2] - h->mn[2]) * h->inv_cell), 0), h->dim[2] - 1);
  return (cz * VG + cy) * VG + cx;
}

__global__ void vg_count_kernel(const float* __restrict__ xyz, int64_t n, int V, const ViewGrid* __restrict__ h,
                                const uint8_t* __restrict__ state, int32_t* __restrict__ count) {
  const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= (int64_t)V * n || state[t] != 1) return;
  const int v = (int)(t / n);
  int cx, cy, cz;
  const int c = vg_cell_of(h, xyz, t - (int64_t)v * n, cx, cy, cz);
  atomicAdd(&count[(int64_t)v * VG3 + c], 1);
}

__global__ void vg_scatter_kernel(const float* __restrict__ xyz, int64_t n, int V, const ViewGrid* __restrict__ h,
                                  const uint8_t* __restrict__ state, const int32_t* __restrict__ start,
                                  int32_t* __restrict__ cursor, int32_t* __restrict__ members) {
  const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= (int64_t)V * n || state[t] != 1) return;
  const int v = (int)(t / n);
  const int64_t p = t - (int64_t)v * n;
  int cx, cy, cz;
  const int64_t c = (int64_t)v * VG3 + vg_cell_of(h, xyz, p, cx, cy, cz);
  members[start[c] + atomicAdd(&cursor[c], 1)] = (int32_t)p;
}

// t-th cell of Chebyshev shell r (bottom face, side rings, top face); false past the end
__device__ __forceinline__ bool vg_shell_cell(int r, int t, int& dx, int& dy, int& dz) {
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
    const int u = t % ring, side = u / (2 * r), j = u % (2 * r);
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

__global__ void __launch_bounds__(256) vg_query_kernel(const float* __restrict__ xyz, int64_t n, int V,
                                                        const ViewGrid* __restrict__ h,
                                                        const uint8_t* __restrict__ state,
                                                        const int32_t* __restrict__ start,
                                                        const int32_t* __restrict__ members, int32_t* qid) {
  const int lane = threadIdx.x & 31;
  const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int64_t tot = (int64_t)V * n;
  unsigned queries = __ballot_sync(0xffffffffu, t < tot && state[t] == 0);
  if (!queries) return;
  const int64_t tbase = t - lane;
  const int dx_ = h->dim[0], dy_ = h->dim[1], dz_ = h->dim[2];
  const int rmax = max(dx_, max(dy_, dz_));
  const double cell = (double)h->cell, eps = 1e-3 * cell;
  const double ox = (double)h->mn[0], oy = (double)h->mn[1], oz = (double)h->mn[2];
  while (queries) {
    const int64_t tq = tbase + (__ffs(queries) - 1);
    queries &= queries - 1;
    const int v = (int)(tq / n);
    const int64_t p = tq - (int64_t)v * n;
    const int32_t* vstart = start + (int64_t)v * VG3;
    if (vstart[VG3] == vstart[0]) continue;  // this view has no point under a real mask: nothing to copy (:604)
    const float x = xyz[3 * p], y = xyz[3 * p + 1], z = xyz[3 * p + 2];
    int qx, qy, qz;
    vg_cell_of(h, xyz, p, qx, qy, qz);
    const double gx = ((double)x - ox) / cell - qx, gy = ((double)y - oy) / cell - qy, gz = ((double)z - oz) / cell - qz;
    double slack = fmin(fmin(fmin(gx, 1.0 - gx), fmin(gy, 1.0 - gy)), fmin(gz, 1.0 - gz));
    slack = fmax(0.0, fmin(slack, 1.0)) * cell;
    double best = 1e300;
    int besti = 0x7fffffff;
    for (int r = 0; r <= rmax; ++r) {
      const int wd = 2 * r + 1;
      const int total = r == 0 ? 1 : 2 * wd * wd + (wd - 2) * 8 * r;
      for (int t0 = 0; t0 < total; t0 += 32) {
        const int tt = t0 + lane;
        int j0 = 0, j1 = 0;
        double lb2 = 1e300;
        int ddx, ddy, ddz;
        if (tt < total && vg_shell_cell(r, tt, ddx, ddy, ddz)) {
          const int cx = qx + ddx, cy = qy + ddy, cz = qz + ddz;
          if (cx >= 0 && cy >= 0 && cz >= 0 && cx < dx_ && cy < dy_ && cz < dz_) {
            const int c = (cz * VG + cy) * VG + cx;
            j0 = __ldg(vstart + c);
            j1 = __ldg(vstart + c + 1);
            const double lx = ox + cx * cell, ly = oy + cy * cell, lz = oz + cz * cell;
            const double ax = fmax(0.0, fmax(lx - (double)x, (double)x - (lx + cell)) - eps);
            const double ay = fmax(0.0, fmax(ly - (double)y, (double)y - (ly + cell)) - eps);
            const double az = fmax(0.0, fmax(lz - (double)z, (double)z - (lz + cell)) - eps);
            lb2 = ax * ax + ay * ay + az * az;
          }
        }
        unsigned cand = __ballot_sync(0xffffffffu, j1 > j0 && lb2 <= best);
        while (cand) {
          const int l = __ffs(cand) - 1;
          cand &= cand - 1;
          const int a0 = __shfl_sync(0xffffffffu, j0, l), a1 = __shfl_sync(0xffffffffu, j1, l);
          const double clb = __shfl_sync(0xffffffffu, lb2, l);
          if (clb > best) continue;  // pruned by a cell scanned since the ballot (best is warp-uniform)
          for (int j = a0 + lane; j < a1; j += 32) {
            const int pi = __ldg(members + j);
            const double ex = (double)__ldg(xyz + 3 * (int64_t)pi) - (double)x,
                         ey = (double)__ldg(xyz + 3 * (int64_t)pi + 1) - (double)y,
                         ez = (double)__ldg(xyz + 3 * (int64_t)pi + 2) - (double)z;
            const double d2 = ex * ex + ey * ey + ez * ez;
            if (d2 < best || (d2 == best && pi < besti)) {
              best = d2;
              besti = pi;
            }
          }
#pragma unroll
          for (int o = 16; o > 0; o >>= 1) {
            const double ob = __shfl_xor_sync(0xffffffffu, best, o);
            const int oi = __shfl_xor_sync(0xffffffffu, besti, o);
            if (ob < best || (ob == best && oi < besti)) {
              best = ob;
              besti = oi;
            }
          }
        }
      }
      const double g = fmax(0.0, r * cell + slack - eps * (r > 0) - eps);
      if (besti != 0x7fffffff && best <= g * g) break;
    }
    // sources are never rewritten, queries only by their own warp: in place is safe
    if (lane == 0 && besti != 0x7fffffff) qid[(int64_t)v * n + p] = qid[(int64_t)v * n + besti];
  }
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
  size_t off_textn, off_tn, off_mlog, off_nz, off_qid, off_state, off_hdr, off_start, off_cursor, off_members, off_cub,
      cub_bytes, total;
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
  L.off_hdr = take(sizeof(ViewGrid));
  L.off_start = take(sizeof(int32_t) * ((size_t)V * VG3 + 1));   // counts, scanned in place into cell starts
  L.off_cursor = take(sizeof(int32_t) * ((size_t)V * VG3 + 1));
  L.off_members = take(sizeof(int32_t) * (size_t)V * n);
  L.cub_bytes = 0;
  cub::DeviceScan::ExclusiveSum(nullptr, L.cub_bytes, (const int32_t*)nullptr, (int32_t*)nullptr, V * VG3 + 1);
  L.off_cub = take(L.cub_bytes + 256);
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
  GP_CHECK_ARG((int64_t)V * n < (1ll << 31) && V < 32768, "gp_lift_consensus: V*n=%lld must stay below 2^31 (V < 32768)",
               (long long)V * (long long)n);
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
  text_normalize_kernel<<<(C + 7) / 8, 256, 0, st>>>(text, C, D, textn);
  GP_LAUNCH_CHECK();
  if (V > 0) {
    const int64_t rows = (int64_t)V * Q;
    mask_prepare_kernel<<<(unsigned)((rows + 7) / 8), 256, 0, st>>>(table, rows, D, textn, C, logit_scale, tn, mlog, nz);
    GP_LAUNCH_CHECK();
    const int64_t tot = (int64_t)V * n;
    mask_id_kernel<<<(unsigned)((tot + 255) / 256), 256, 0, st>>>(pix, n, V, H * W, keep, idmap, nz, Q, qid, state);
    GP_LAUNCH_CHECK();
    // per-view hole fill among the points each view sees (:601-625), every view in the same launches
    ViewGrid* vg = (ViewGrid*)(ws + L.off_hdr);
    int32_t* start = (int32_t*)(ws + L.off_start);
    int32_t* cursor = (int32_t*)(ws + L.off_cursor);
    int32_t* members = (int32_t*)(ws + L.off_members);
    const size_t tab = sizeof(int32_t) * ((size_t)V * VG3 + 1);
    GP_CUDA(cudaMemsetAsync(start, 0, tab, st));
    GP_CUDA(cudaMemsetAsync(cursor, 0, tab, st));
    vg_init_kernel<<<1, 1, 0, st>>>(vg);
    const unsigned gn = (unsigned)((n + 255) / 256);
    vg_bbox_kernel<<<gn < 1184 ? gn : 1184, 256, 0, st>>>(xyz, n, vg);
    vg_grid_kernel<<<1, 1, 0, st>>>(vg, cell);
    const unsigned gt = (unsigned)((tot + 255) / 256);
    vg_count_kernel<<<gt, 256, 0, st>>>(xyz, n, V, vg, state, start);
    GP_LAUNCH_CHECK();
    size_t cb = L.cub_bytes + 256;
    GP_CUDA(cub::DeviceScan::ExclusiveSum(ws + L.off_cub, cb, start, start, V * VG3 + 1, st));
    vg_scatter_kernel<<<gt, 256, 0, st>>>(xyz, n, V, vg, state, start, cursor, members);
    GP_LAUNCH_CHECK();
    vg_query_kernel<<<gt, 256, 0, st>>>(xyz, n, V, vg, state, start, members, qid);
    GP_LAUNCH_CHECK();
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
