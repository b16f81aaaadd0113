// F2: scene batch assembly on the device.  Replaces the per-view index packing of the reference's loader and collate
// (dataset/data_loader_ablation.py:254-288 view acceptance, :348-353 x_label / y_label of the visible points in point
// order, collate :446-459 concatenation over the kept views with renumbered view ids) and makes the V*N-row
// mask_2ds / unique_maps tensors unnecessary: a CSR (kept view -> visible point list) comes straight from the
// pix [V,n] tensor of gp_project_views.
#include <cub/device/device_scan.cuh>
#include "common.cuh"

namespace {

__global__ void csr_flags_kernel(const int32_t* __restrict__ pix, const uint8_t* __restrict__ keep, int64_t n, int V,
                                 int32_t* __restrict__ flags) {
  const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int64_t total = (int64_t)V * n;
  if (t > total) return;
  int f = 0;
  if (t < total) {
    const int v = (int)(t / n);
    f = pix[t] >= 0 && (!keep || keep[v]);
  }
  flags[t] = f;  // flags[total] = 0: the scan's last entry is the number of pairs
}

__global__ void csr_fill_kernel(const int32_t* __restrict__ pix, const int32_t* __restrict__ pos, int64_t n, int V,
                                int W, int32_t* __restrict__ point_idx, int32_t* __restrict__ view_idx,
                                int32_t* __restrict__ rows, int32_t* __restrict__ cols) {
  const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= (int64_t)V * n) return;
  const int p = pos[t];
  if (pos[t + 1] == p) return;
  const int v = (int)(t / n);
  const int px = pix[t];
  point_idx[p] = (int32_t)(t - (int64_t)v * n);
  if (view_idx) view_idx[p] = v;
  if (rows) rows[p] = px / W;
  if (cols) cols[p] = px % W;
}

// offsets[v] = first pair of view v (original view numbering; dropped views are empty), offsets[V] = P
__global__ void csr_offsets_kernel(const int32_t* __restrict__ pos, int64_t n, int V, int32_t* __restrict__ offsets) {
  const int v = blockIdx.x * blockDim.x + threadIdx.x;
  if (v <= V) offsets[v] = pos[(int64_t)v * n];
}

}  // namespace

extern "C" size_t gp_visible_csr_workspace_bytes(int64_t n, int V) {
  if (n < 0 || V < 0) return 0;
  const size_t total = (size_t)V * n + 1;
  size_t cb = 0;
  cub::DeviceScan::ExclusiveSum(nullptr, cb, (const int32_t*)nullptr, (int32_t*)nullptr, (int)total);
  return gp_align_up(sizeof(int32_t) * total) * 2 + gp_align_up(cb + 256);
}

extern "C" int gp_visible_csr(const int32_t* pix, const uint8_t* keep, int64_t n, int V, int W, int32_t* offsets,
                              int32_t* point_idx, int32_t* view_idx, int32_t* rows, int32_t* cols, void* workspace,
                              size_t workspace_bytes, gp_stream_t stream) {
  GP_CHECK_ARG(n >= 0 && V >= 0 && W > 0 && (int64_t)V * n < (1ll << 31) - 1, "gp_visible_csr: bad sizes");
  GP_CHECK_ARG(offsets && point_idx, "gp_visible_csr: offsets and point_idx are required");
  const size_t need = gp_visible_csr_workspace_bytes(n, V);
  if (workspace_bytes < need) {
    gp_set_error("gp_visible_csr: workspace %zu < %zu", workspace_bytes, need);
    return GP_ERR_WORKSPACE;
  }
  cudaStream_t st = (cudaStream_t)stream;
  const int64_t total = (int64_t)V * n;
  char* ws = (char*)workspace;
  int32_t* flags = (int32_t*)ws;
  int32_t* pos = (int32_t*)(ws + gp_align_up(sizeof(int32_t) * (total + 1)));
  void* cubws = ws + 2 * gp_align_up(sizeof(int32_t) * (total + 1));
  size_t cb = workspace_bytes - 2 * gp_align_up(sizeof(int32_t) * (total + 1));
  const unsigned grid = (unsigned)((total + 1 + 255) / 256);
  csr_flags_kernel<<<grid, 256, 0, st>>>(pix, keep, n, V, flags);
  GP_LAUNCH_CHECK();
  GP_CUDA(cub::DeviceScan::ExclusiveSum(cubws, cb, flags, pos, (int)(total + 1), st));
  if (total > 0) {
    csr_fill_kernel<<<grid, 256, 0, st>>>(pix, pos, n, V, W, point_idx, view_idx, rows, cols);
    GP_LAUNCH_CHECK();
  }
  csr_offsets_kernel<<<(V + 1 + 127) / 128, 128, 0, st>>>(pos, n, V, offsets);
  GP_LAUNCH_CHECK();
  return GP_OK;
}
