// F3: mask post-processing of the 2-D VLM output -- the panoptic-style partition of the reference
// (models/affinity_module.py:544-570): per query score = max softmax class probability (background column dropped),
// keep score > threshold, prob = score * sigmoid(mask logit), every pixel goes to the arg-max query, and stays
// covered only where that query's own sigmoid is >= 0.5.  The surviving masks are disjoint, so the per-point
// "mean of the covering masks' embeddings" (:584-592) is a single table lookup: the output is the uint8 id map
// that gp_lift_consensus / gp_lift_mean_idmap read (value Q = no mask -> the all-zero row appended to the table).
// Replaces a Python loop over <= 200 queries with an .item() sync per query and three [H,W] reductions each.
#include "common.cuh"

namespace {

// one warp per (view, query): softmax over the C1 class logits, max over the first C1-1
__global__ void __launch_bounds__(256) mask_scores_kernel(const float* __restrict__ cls, int64_t rows, int C1,
                                                          float thresh, float* __restrict__ score) {
  const int lane = threadIdx.x & 31;
  const int64_t r = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (r >= rows) return;
  const float* x = cls + r * C1;
  float mx = -INFINITY;
  for (int c = lane; c < C1; c += 32) mx = fmaxf(mx, x[c]);
  mx = warp_max(mx);
  float den = 0.f, best = -INFINITY;
  for (int c = lane; c < C1; c += 32) {
    const float e = expf(x[c] - mx);
    den += e;
    if (c < C1 - 1) best = fmaxf(best, e);
  }
  den = warp_sum(den);
  best = warp_max(best);
  const float s = best / den;
  if (lane == 0) score[r] = s > thresh ? s : -1.f;  // -1: query dropped (scores are probabilities, >= 0)
}

__global__ void __launch_bounds__(256) mask_partition_kernel(const float* __restrict__ logits,
                                                             const float* __restrict__ score, int Q, int64_t hw,
                                                             uint8_t* __restrict__ idmap) {
  extern __shared__ float s_score[];
  const int v = blockIdx.y;
  for (int q = threadIdx.x; q < Q; q += blockDim.x) s_score[q] = score[(int64_t)v * Q + q];
  __syncthreads();
  const int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= hw) return;
  const float* lp = logits + (int64_t)v * Q * hw + p;
  float best = -INFINITY, best_logit = 0.f;
  int bq = -1;
  for (int q = 0; q < Q; ++q) {
    const float sc = s_score[q];
    if (sc < 0.f) continue;
    const float x = __ldg(lp + (int64_t)q * hw);
    const float pr = sc * (1.f / (1.f + expf(-x)));  // cur_scores * cur_masks.sigmoid()
    if (pr > best) {                                 // argmax(0): first maximum among the kept queries
      best = pr;
      best_logit = x;
      bq = q;
    }
  }
  const bool covered = bq >= 0 && (1.f / (1.f + expf(-best_logit))) >= 0.5f;  // cur_masks[k] >= 0.5
  idmap[(int64_t)v * hw + p] = (uint8_t)(covered ? bq : Q);
}

}  // namespace

extern "C" int gp_mask_partition(const float* mask_logits, const float* cls_logits, int V, int Q, int C1, int H, int W,
                                 float score_thresh, uint8_t* idmap, float* scores, gp_stream_t stream) {
  GP_CHECK_ARG(V >= 0 && Q >= 1 && Q <= 255 && C1 >= 2 && H > 0 && W > 0, "gp_mask_partition: bad sizes (Q <= 255)");
  GP_CHECK_ARG(scores != nullptr, "gp_mask_partition: scores [V,Q] buffer is required");
  if (V == 0) return GP_OK;
  cudaStream_t st = (cudaStream_t)stream;
  const int64_t rows = (int64_t)V * Q, hw = (int64_t)H * W;
  mask_scores_kernel<<<(unsigned)((rows + 7) / 8), 256, 0, st>>>(cls_logits, rows, C1, score_thresh, scores);
  GP_LAUNCH_CHECK();
  dim3 grid((unsigned)((hw + 255) / 256), (unsigned)V);
  mask_partition_kernel<<<grid, 256, Q * sizeof(float), st>>>(mask_logits, scores, Q, hw, idmap);
  GP_LAUNCH_CHECK();
  return GP_OK;
}
