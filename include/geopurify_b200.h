/* geopurify_b200 -- C ABI of libgeopurify_b200.so (sm_100a).
 *
 * Drop-in boundary for the inference hot path of tj12323/GeoPurify.  The reference
 * has no FFI of its own: its hot path is Python calling numpy / torch / third-party CUDA libraries.
 * Each entry point below replaces one of those call sites (cited as reference file:line) and is what
 * a ctypes binding in the reference tree would load.
 *
 * Conventions
 *   - every pointer is a DEVICE pointer unless the name ends in _host; plain pointers and sizes only;
 *   - the caller owns all buffers, outputs included; functions never allocate what they return;
 *     scratch comes from a caller-provided workspace whose size a *_workspace_bytes() query returns;
 *   - all work is enqueued on `stream` (a cudaStream_t) and is asynchronous w.r.t. the host;
 *   - return value: 0 = ok, GP_ERR_* otherwise; gp_last_error() gives the text for this thread;
 *   - no global state, re-entrant, one device per process.
 */
#ifndef GEOPURIFY_B200_H
#define GEOPURIFY_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef void* gp_stream_t; /* cudaStream_t */

enum {
  GP_OK = 0,
  GP_ERR_ARG = 1,       /* bad argument (shape, alignment, unsupported size) */
  GP_ERR_WORKSPACE = 2, /* workspace too small */
  GP_ERR_CUDA = 3       /* CUDA runtime error, text in gp_last_error() */
};

enum { GP_F32 = 1, GP_F64 = 2, GP_BF16 = 3 };

int gp_version(void);
const char* gp_last_error(void);

/* Optional measurement hook (off by default, not part of the reference surface): when enabled, CUDA events are
 * recorded on the launching stream around every launch of the dominant kernels; gp_timing_read synchronises them
 * and returns the summed duration and the launch count since the last gp_timing_enable.
 * which: 0 student gather-GEMM, 1 student reduce, 2 pooling application (tcgen05), 3 reserved. */
int gp_timing_enable(int on);
int gp_timing_read(int which, double* total_ms, int* launches);

/* ---- A1 + A3: PointCloudToImageMapper.compute_mapping for every view of a scene in one launch
 * (models/utils/fusion_util.py:99-147, Matterport twin :45-82; callers dataset/data_loader_ablation.py:242-253,
 * dataset/data_loader_matterport.py:194-205).  float64 arithmetic in the reference's operation order.
 *   coords      [n,3]  float32 or float64 (coords_dtype)
 *   w2c         [V,16] float64 row-major world-to-camera (the float32 matrix promoted; NOT transposed)
 *   intr        [V,4]  float64 fx, fy, cx, cy (after the mapper ctor's rescale)
 *   depth       [V,H,W] float32/float64 metres, or NULL with depth_dtype 0 (then the p.z > 0 test applies)
 *   pix  (out)  [V,n]  int32 row*W+col of a visible point, -1 otherwise
 *   counts(out) [V]    int32 visible points per view (zeroed by the call) */
int gp_project_views(const void* coords, int coords_dtype, int64_t n, const double* w2c, const double* intr,
                     const void* depth, int depth_dtype, int V, int H, int W, int cut, double vis_thres,
                     int32_t* pix, int32_t* counts, gp_stream_t stream);

/* pix -> the reference's mapping[V,n,3] int64 (row, col, mask), zeros where invisible (fusion_util.py:142-144). */
int gp_pix_to_mapping(const int32_t* pix, int64_t total, int W, int64_t* mapping, gp_stream_t stream);

/* ---- A2: z-buffer depth synthesis, the string-depth branch (fusion_util.py:126-130).
 *   depth (out) [V,H,W] float64: min p.z over points with p.z > 0.2 inside the cut bound, 999999 elsewhere */
int gp_splat_depth(const void* coords, int coords_dtype, int64_t n, const double* w2c, const double* intr, int V,
                   int H, int W, int cut, double* depth, gp_stream_t stream);

/* ---- A3: keep[v] = counts[v] != 0 && counts[v] >= min_visible && counts[v] <= max_visible
 * (dataset/data_loader_ablation.py:254-288). */
int gp_accept_views(const int32_t* counts, int V, int min_visible, int max_visible, uint8_t* keep,
                    gp_stream_t stream);

/* ---- F3: mask post-processing of the 2-D VLM output (models/affinity_module.py:544-570): panoptic-style partition.
 *   mask_logits [V,Q,H,W] float32 (already resized to mask_shape), cls_logits [V,Q,C1] float32 (last class =
 *   background), score_thresh = the reference's scores_keep_thresh (0.0)
 *   idmap (out) [V,H,W] uint8: arg-max query of the pixel where that query's sigmoid >= 0.5, else Q ("no mask")
 *   scores (out) [V,Q] float32: max softmax probability over the C1-1 real classes, -1 for dropped queries
 * The surviving masks are disjoint, so idmap + the query embeddings (with an all-zero row appended at index Q) is
 * exactly the input of gp_lift_consensus. */
int gp_mask_partition(const float* mask_logits, const float* cls_logits, int V, int Q, int C1, int H, int W,
                      float score_thresh, uint8_t* idmap, float* scores, gp_stream_t stream);

/* ---- F2: scene batch assembly (dataset/data_loader_ablation.py:254-288, 348-353; collate :446-459): CSR of the
 * visible (view, point) pairs in (view, point) order, straight from pix [V,n].
 *   offsets   (out) [V+1] int32 first pair of every view (views with keep == 0 are empty), offsets[V] = P
 *   point_idx (out) [>= P] int32 point of every pair;  view_idx / rows / cols (out, may be NULL) [>= P] int32:
 *   the reference's x_label (row) and y_label (col).  Capacity V*n is always enough; P is offsets[V]. */
size_t gp_visible_csr_workspace_bytes(int64_t n, int V);
int gp_visible_csr(const int32_t* pix, const uint8_t* keep, int64_t n, int V, int W, int32_t* offsets,
                   int32_t* point_idx, int32_t* view_idx, int32_t* rows, int32_t* cols, void* workspace,
                   size_t workspace_bytes, gp_stream_t stream);

/* ---- A1+A4 fused: project, occlusion-test and gather-mean in one pass (models/affinity_module.py:404-438).
 * One warp per point; cameras in shared memory; 128-bit loads of channels-last feature rows; float32 sum in
 * ascending view order (the reference's index_add_ order).
 *   feat        [V,H,W,D] float32 or bf16 (feat_dtype), D % 128 == 0 (fp32) / D % 256 == 0 (bf16), D <= 1024
 *   keep        [V] uint8 or NULL (all views kept)
 *   out  (out)  [n, out_stride] float32 mean feature in columns [0,D); other columns untouched
 *   seen (out)  [n] uint8, 1 if at least one kept view saw the point */
int gp_lift_mean_fused(const void* coords, int coords_dtype, int64_t n, const double* w2c, const double* intr,
                       const void* depth, int depth_dtype, int V, int H, int W, int cut, double vis_thres,
                       const uint8_t* keep, const void* feat, int feat_dtype, int D, float* out,
                       int64_t out_stride, uint8_t* seen, gp_stream_t stream);

/* ---- A1+A4 fused on the COMPACT 2-D form: the feature of pixel (v, p) is table[v, idmap[v, p], :], so the 3.9 GB
 * dense maps are never built (the X-Decoder path of the reference never builds them either: it adds mask embeddings
 * per point, models/affinity_module.py:584-592).  Result is bit-identical to gp_expand_idmap + gp_lift_mean_fused
 * with float32 maps.   idmap [V,H,W] uint8, table [V,Q,D] float32 or bf16 (table_dtype), Q <= 256 */
int gp_lift_mean_idmap(const void* coords, int coords_dtype, int64_t n, const double* w2c, const double* intr,
                       const void* depth, int depth_dtype, int V, int H, int W, int cut, double vis_thres,
                       const uint8_t* keep, const uint8_t* idmap, const void* table, int table_dtype, int Q, int D,
                       float* out, int64_t out_stride, uint8_t* seen, gp_stream_t stream);

/* ---- 2-D feature form conversion: feat[v,h,w,:] = table[v, idmap[v,h,w], :] (the per-pixel view of the mask
 * embedding lookup at models/affinity_module.py:584-592; also the synthetic stand-in for the 2-D VLM output).
 *   idmap [V,H,W] uint8, table [V,Q,D] float32, feat (out) [V,H,W,D] float32 or bf16, D % 8 == 0 */
int gp_expand_idmap(const uint8_t* idmap, const float* table, int V, int H, int W, int Q, int D, void* feat,
                    int feat_dtype, gp_stream_t stream);

/* ---- A4 from precomputed pixels (the reference's x_label / y_label / mask_2d form): same gather-mean, pix from
 * gp_project_views or from the reference data loader. */
int gp_lift_mean_pix(const int32_t* pix, int64_t n, int V, int H, int W, const uint8_t* keep, const void* feat,
                     int feat_dtype, int D, float* out, int64_t out_stride, uint8_t* seen, gp_stream_t stream);

/* ---- A5: X-Decoder-style lifting = mask-embedding gather + per-view hole fill + consensus top-3 fusion
 * (models/affinity_module.py:584-592, 601-630, 633-685).  The 2-D side is in its compact form:
 *   pix    [V,n]   int32 from gp_project_views (or the loader's x_label/y_label/mask_2d)
 *   idmap  [V,H,W] uint8 index of the mask covering each pixel;  table [V,Q,D] float32 mask embeddings, an all-zero
 *          row meaning "no mask" (the reference's zero feature rows, detected by sum(dim=1) == 0)
 *   text   [C,D] float32 raw (normalised inside), C <= 256;  xyz [n,3] float32 coordinates for the per-view 1-NN fill
 *   out (out) [n, out_stride] float32 fused feature in columns [0,D), zeros for never-seen points (fill them with
 *          gp_nn1_fill + gp_copy_rows, :687-696);  seen (out) [n] uint8
 * Per view: a visible point under a zero row takes the mask of the nearest visible point under a non-zero row
 * (a view whose rows are all zero contributes zero features); rows are L2-normalised; logits = scale * f t^T.
 * Per point: mean logits over its views -> consensus class (lowest index on ties) -> top-min(3, #views) views by
 * their logit of that class (earliest view on ties) -> softmax weights -> weighted sum. */
size_t gp_lift_consensus_workspace_bytes(int64_t n, int V, int Q, int D, int C);
int gp_lift_consensus(const int32_t* pix, int64_t n, int V, int H, int W, const uint8_t* keep, const uint8_t* idmap,
                      const float* table, int Q, int D, const float* text, int C, float logit_scale,
                      const float* xyz, float cell, float* out, int64_t out_stride, uint8_t* seen, void* workspace,
                      size_t workspace_bytes, gp_stream_t stream);

/* ---- hole fill: 1-NN over float xyz (sklearn KDTree k=1 at models/affinity_module.py:445-449, 693-696;
 * run/validation.py:425-430).  For every point with seen==0 finds the nearest seen point (float64 distance on
 * the float32 coordinates, ties -> lowest index) and writes nn[i] = that index; nn[i] = i for seen points,
 * -1 everywhere if no point is seen.  `cell` is the search-grid cell size in coordinate units.
 * seen[i] == 2 marks a point that is neither a source nor a query (nn[i] is left untouched). */
size_t gp_nn1_fill_workspace_bytes(int64_t n);
int gp_nn1_fill(const float* xyz, int64_t n, const uint8_t* seen, float cell, int32_t* nn, void* workspace,
                size_t workspace_bytes, gp_stream_t stream);
/* rows[i, 0:D) = rows[nn[i], 0:D) for nn[i] != i (out-of-place safe: sources are seen rows, never rewritten). */
int gp_copy_rows(float* rows, int64_t n, int D, int64_t stride, const int32_t* nn, gp_stream_t stream);

/* ---- F1: Voxelizer.voxelize + sparse_quantize (dataset/voxelizer.py:104-121; dataset/voxelization_utils.py:6-18,
 * 78-96): c = floor(M [x y z 1]^T) in float64, v = c - min(c), FNV-1-style uint64 key per voxel, np.unique on the
 * keys: voxels numbered in ascending key order, representative = first point of the voxel.
 *   coords [n,3] float32/float64;  xform [3,4] float64 row-major (device): diag(1/voxel_size) | 0 for the
 *   identity-augmentation case (or M_r M_v with augmentation)
 *   vox (out) [n,3] int32 (first m rows valid), inds (out) [n] int32 (first m valid: representative point),
 *   inds_reconstruct (out) [n] int32 point -> voxel, m_out (out) device int64, extent (out) device int32[3] */
size_t gp_voxelize_workspace_bytes(int64_t n);
int gp_voxelize(const void* coords, int coords_dtype, int64_t n, const double* xform, int32_t* vox, int32_t* inds,
                int32_t* inds_reconstruct, int64_t* m_out, int32_t* extent, void* workspace, size_t workspace_bytes,
                gp_stream_t stream);

/* ---- voxel index: Morton-sorted order + cell lists over integer voxel coordinates (replaces the MinkowskiEngine
 * coordinate map of models/affinity_module.py:1541-1545 and the faiss index of :1553-1554).
 *   vox         [m,3] int32 non-negative voxel coordinates, unique rows
 *   order (out) [m] int32  sorted position -> original voxel
 *   rank  (out) [m] int32  original voxel -> sorted position
 * (ex,ey,ez) = exclusive upper bound of the coordinates (<= 32768 each).  The built index lives in `workspace`
 * and is consumed by gp_kernel_map / gp_knn_grid, which must be given the same (m, ex, ey, ez). */
size_t gp_voxel_index_workspace_bytes(int64_t m, int ex, int ey, int ez);
int gp_voxel_index_build(const int32_t* vox, int64_t m, int ex, int ey, int ez, int32_t* order, int32_t* rank,
                         void* workspace, size_t workspace_bytes, gp_stream_t stream);
/* Synchronising validity check of a built index: GP_ERR_ARG if a voxel coordinate was outside [0, extent).  The build
 * itself never writes out of bounds for such input (the voxel is left out of the cell table and a flag is raised). */
int gp_voxel_index_check(const void* index_workspace, int64_t m, int ex, int ey, int ez, gp_stream_t stream);

/* ---- A7: 3^3 kernel map in SORTED space: nbr27[s, o] = sorted position of the voxel at coord(s) + delta_o,
 * o = (dx+1) + 3(dy+1) + 9(dz+1), or -1. */
int gp_kernel_map(const void* index_workspace, int64_t m, int ex, int ey, int ez, int32_t* nbr27,
                  gp_stream_t stream);

/* ---- A9: exact K nearest voxels, ascending (d^2, original index), self dropped, K <= 128
 * (faiss IndexFlatL2 search at models/affinity_module.py:1553-1557).
 *   nbr_orig   (out, may be NULL) [m,K] int32 rows/entries in ORIGINAL voxel numbering (the reference's
 *                                  neighbor_indices)
 *   nbr_sorted (out, may be NULL) [m,K] int32 rows/entries in SORTED numbering (input of gp_affinity / gp_pool) */
int gp_knn_grid(const void* index_workspace, int64_t m, int ex, int ey, int ez, int K, int32_t* nbr_orig,
                int32_t* nbr_sorted, gp_stream_t stream);

/* ---- F4 (training-side sampling): exact K nearest neighbours on RAW float points, self excluded, ascending
 * (d2, index) with d2 evaluated in float32 as ((dx*dx + dy*dy) + dz*dz), every operation rounded once
 * (faiss IndexFlatL2 add/search on scene_coords at models/affinity_module.py:1159-1165, first column dropped).
 *   xyz [n,3] float32;  nbr (out) [n,K] int32;  K <= 128, n >= K+1;  min_cell = lower bound of the grid cell edge, which is
 *   otherwise chosen for about 32 points per cell at the bounding box's surface density */
size_t gp_knn_points_workspace_bytes(int64_t n);
int gp_knn_points(const float* xyz, int64_t n, int K, float min_cell, int32_t* nbr, void* workspace,
                  size_t workspace_bytes, gp_stream_t stream);

/* ---- F4 (training-side sampling): hybrid contrastive mining for a given anchor set
 * (SonataXAffinityTrainer.sample_contrastive_pairs_hybrid, models/affinity_module.py:1099-1136, after its randperm).
 *   feat      [n, stride] float32, D columns used (the 3-D teacher features); rows are L2-normalised internally
 *   anchors   [A] int32 distinct point indices;  nbr [n,K] int32 spatial neighbours (gp_knn_points)
 *   positive  (out) [A] int32   arg-max cosine similarity over all points except the anchor
 *   negatives (out) [A, n_macro + n_micro] int32: the n_macro least similar points (anchor and positive excluded,
 *             ascending similarity) followed by the n_micro least similar of the anchor's K neighbours (positive
 *             excluded).  Ties: lowest index (macro) / earliest list position (micro) first.
 * Similarities are an fp32-faithful tcgen05 GEMM (bf16 hi/lo split, error <= 2^-15). n_macro <= 63, n_micro <= 32. */
size_t gp_mine_workspace_bytes(int64_t n, int D, int A, int n_macro);
int gp_mine_contrastive(const float* feat, int64_t stride, int D, int64_t n, const int32_t* anchors, int A,
                        const int32_t* nbr, int K, int n_macro, int n_micro, int32_t* positive, int32_t* negatives,
                        void* workspace, size_t workspace_bytes, gp_stream_t stream);

/* ---- A6: torch_scatter.scatter_mean (models/affinity_module.py:1524-1535), deterministic: float32 sum in
 * ascending point order, divided by the count.
 *   src  [n, src_stride] float32, D columns used;  p2v [n] int32 point -> destination row (already composed
 *   with `rank` when rows are wanted in sorted order);  out [m, out_stride], columns [col0, col0+D) written */
size_t gp_scatter_mean_workspace_bytes(int64_t n, int64_t m);
int gp_scatter_mean_prepare(const int32_t* p2v, int64_t n, int64_t m, void* workspace, size_t workspace_bytes,
                            gp_stream_t stream);
int gp_scatter_mean(const float* src, int64_t src_stride, int D, const void* prepared_workspace, int64_t n,
                    int64_t m, float* out, int64_t out_stride, int col0, gp_stream_t stream);

/* ---- A6 without any allocation.  NO COUNTERPART IN THE REFERENCE: added for
 * callers that cannot call cudaMalloc (CUDA graph capture, tight serving loops, harnesses that forbid it).
 * Same arguments and bit-identical result as gp_scatter_mean_prepare + gp_scatter_mean, but the scratch lives in a
 * 16 MiB `static __device__` array inside the library (held by every GPU that loads it, until process exit) and the
 * call makes no cudaMalloc / cudaFree / allocating library call.
 *   limits   n <= GP_SCATTER_STATIC_MAX_N and m <= GP_SCATTER_STATIC_MAX_M, else GP_ERR_ARG with a clear message
 *            (GP_ERR_WORKSPACE if the layout for (n, m) would still not fit the array);
 *   sharing  all calls on one GPU use the same array: calls that overlap on the GPU (two streams, or a captured graph
 *            replayed beside a live call) RACE -- serialise them.  The array keeps the last call's contents; every word
 *            the call reads it has written first, so there is nothing to clear;
 *   device   the stream's device must be current (as for every entry point); the first call on a device resolves the
 *            array's address with cudaGetSymbolAddress, so make that call outside stream capture. */
#define GP_SCATTER_STATIC_BYTES (16 * 1024 * 1024)
#define GP_SCATTER_STATIC_MAX_N 262144
#define GP_SCATTER_STATIC_MAX_M 262144
int gp_scatter_mean_static(const int32_t* p2v, int64_t n, int64_t m, const float* src, int64_t src_stride, int D,
                           float* out, int64_t out_stride, int col0, gp_stream_t stream);

/* ---- A8: AffinityPredictor.forward + F.normalize (models/affinity_module.py:33-72, 1546-1547).
 * Sparse 3^3 convolutions as gather-GEMM over the kernel map, BN (eval) + ReLU + residual epilogues.
 *   x        [m, cin_stride] float32 rows in sorted order, cin columns valid
 *   weights  packed by gp_student_pack_bytes()/the Python shim: per layer [kv, cin_pad, cout] float32,
 *            then folded BN scale/shift [cout] pairs (see geopurify_b200/student.py)
 *   embed (out) [m, E] float32 unit rows */
typedef struct {
  int n_layers;          /* conv layers, <= 16 */
  int kv[16];            /* 27 or 1 */
  int cin[16];           /* padded to a multiple of 32 */
  int cout[16];
  int bn[16];            /* 1: y = y*scale + shift */
  int relu[16];          /* 1: ReLU after (BN, residual add) */
  int residual_from[16]; /* -1, or index of the layer whose OUTPUT is added before the ReLU (-2: network input) */
  int64_t w_off[16];     /* float offset of the layer's weights inside `weights` */
  int64_t bn_off[16];    /* float offset of scale[cout] followed by shift[cout] */
} gp_student_desc;
size_t gp_student_workspace_bytes(const gp_student_desc* desc_host, int64_t m);
int gp_student_forward(const gp_student_desc* desc_host, const float* weights, const float* x, int64_t cin_stride,
                       const int32_t* nbr27, int64_t m, float* embed, int normalize, void* workspace,
                       size_t workspace_bytes, gp_stream_t stream);

/* ---- A8 on the tensor cores (tcgen05 + TMEM, sm_100a): same contract as gp_student_forward, fp32-faithful through a
 * bf16 hi/lo split of activations and weights (3 MMAs per product term, fp32 accumulate in TMEM).
 *   packed    bf16 K-major repack of the conv weights made once by gp_student_tc_pack (128-byte aligned)
 *   weights   the same float32 buffer as gp_student_forward (the folded BN scale/shift are read from it)
 *   workspace 256-byte aligned */
size_t gp_student_tc_pack_bytes(const gp_student_desc* desc_host);
int gp_student_tc_pack(const gp_student_desc* desc_host, const float* weights, void* packed, gp_stream_t stream);
size_t gp_student_tc_workspace_bytes(const gp_student_desc* desc_host, int64_t m);
int gp_student_forward_tc(const gp_student_desc* desc_host, const float* weights, const void* packed, const float* x,
                          int64_t cin_stride, const int32_t* nbr27, int64_t m, float* embed, int normalize,
                          void* workspace, size_t workspace_bytes, gp_stream_t stream);

/* ---- A10: w[s,k] = softmax_k(sharpen * <e_s, e_nbr(s,k)>) (models/affinity_module.py:1559-1572). */
int gp_affinity_softmax(const float* embed, int E, const int32_t* nbr_sorted, int64_t m, int K, float sharpen,
                        float* weights, gp_stream_t stream);

/* ---- A11: F <- A F, T times (models/affinity_module.py:1575-1587); float32 accumulate.
 *   f0 [m, stride] float32 (D columns pooled), buf_a/buf_b [m, D] ping-pong; result pointer returned in *out */
int gp_pool(const float* f0, int64_t stride, int D, const int32_t* nbr_sorted, const float* weights, int64_t m,
            int K, int T, float* buf_a, float* buf_b, float** out_host, gp_stream_t stream);

/* ---- A10+A11 on the tensor cores (tcgen05 + TMEM): F <- A F as a block-local dense GEMM.  A per-scene plan groups
 * 128 consecutive (Hilbert-sorted) voxels with the sorted union of their neighbour lists and keeps the tile's weights
 * in a compact form (6 bytes per (voxel, neighbour) entry, grouped by 16-row K-step); every application expands the
 * 128 x 16 weight blocks in shared memory, gathers each union row once per (tile, 256-column half) and accumulates
 * 128 x 256 fp32 outputs in TMEM (3 bf16 MMAs per product term), two CTAs per SM.  F moves between applications as
 * bf16 hi/lo planes; the last application writes float32.  D must be 256 or 512, T >= 1; `nbr_sorted` / `weights`
 * serve tiles whose union exceeds 2048 rows.  Precondition: the K ids of one row are distinct (true for kNN lists;
 * gp_pool sums duplicates, this path keeps one of them).
 *   out (out) [m, D] float32;  plan / workspace 256-byte aligned */
size_t gp_pool_plan_bytes(int64_t m, int K);
int gp_pool_plan(const int32_t* nbr_sorted, const float* weights, int64_t m, int K, void* plan, size_t plan_bytes,
                 gp_stream_t stream);
size_t gp_pool_tc_workspace_bytes(int64_t m);
int gp_pool_tc(const float* f0, int64_t stride, int D, const void* plan, const int32_t* nbr_sorted,
               const float* weights, int64_t m, int K, int T, float* out, void* workspace, size_t workspace_bytes,
               gp_stream_t stream);

/* ---- A11 tail: scene_features[n, 0:D) = pooled[rank[p2v_orig[n]], 0:D) (models/affinity_module.py:1589). */
int gp_gather_rows(const float* rows, int64_t row_stride, int D, const int32_t* index, int64_t n, float* out,
                   gp_stream_t stream);

/* ---- A12: normalise, logits = s * f t^T, argmax (run/validation.py:413-417).
 *   feat [rows, stride] float32 (D columns), text [C,D] float32 (raw; normalised inside)
 *   label (out) [rows] int32 lowest-index argmax;  zero (out, may be NULL) [rows] uint8 1 if sum|f^| == 0 */
size_t gp_logits_workspace_bytes(int C, int D);
int gp_logits_argmax(const float* feat, int64_t stride, int D, int64_t rows, const float* text, int C,
                     float logit_scale, int32_t* label, uint8_t* zero, float* logits_or_null, void* workspace,
                     size_t workspace_bytes, gp_stream_t stream);

/* ---- A12 on the tensor cores: the same labels from a tcgen05/TMEM GEMM F[rows,D] x T^[D,C] (bf16 hi/lo split, fp32
 * accumulate) with a fused arg-max epilogue -- no logits in HBM.  Rows whose top-2 margin is below 2e-4 |f| (near ties,
 * all-zero rows) are re-evaluated by the fp32 kernel of gp_logits_argmax, so both entry points return identical labels.
 * C <= 256, logit_scale > 0, workspace 256-byte aligned.  zero (may be NULL) as in gp_logits_argmax. */
size_t gp_logits_tc_workspace_bytes(int64_t rows, int C, int D);
int gp_logits_argmax_tc(const float* feat, int64_t stride, int D, int64_t rows, const float* text, int C,
                        float logit_scale, int32_t* label, uint8_t* zero, void* workspace, size_t workspace_bytes,
                        gp_stream_t stream);

/* ---- A12: intersectionAndUnionGPU (util/util.py:160-177): pred[target in ignore] = ignore, three histograms.
 *   label_rows [rows] int32, index [n] int32 or NULL (pred[i] = label_rows[index[i]] else label_rows[i])
 *   target [n] int64, ignore [n_ignore] int64 (device)
 *   pred_out (may be NULL) [n] int64 prediction after the ignore overwrite
 *   hist (out) [3,C] int64 intersection, union, target; ACCUMULATED into (caller zeroes once per run) */
int gp_iou_hist(const int32_t* label_rows, const int32_t* index, const int64_t* target, int64_t n,
                const int64_t* ignore, int n_ignore, int C, int64_t* pred_out, int64_t* hist, gp_stream_t stream);

#ifdef __cplusplus
}
#endif
#endif
