/* CPU oracle for SURVEY.md §8 row A11 -- TEST INFRASTRUCTURE, NOT PRODUCT.
 *
 * Restates the reference's pooling loop (models/affinity_module.py:1575-1587)
 *     A = torch.sparse_coo_tensor([rows; nbr], w, (M, M));  for _ in range(T): F = torch.sparse.mm(A, F)
 * as a fixed-K CSR SpMM.  torch's CPU kernel (torch 2.11, the version in this image) keeps the entries of a row in
 * the order they were given (rows are already ascending, so its COO -> CSR conversion is a stable no-op) and adds
 * val * F[col, :] to the output row entry by entry with fused multiply-adds (axpy); this file does the same:
 * neighbour order k = 0..K-1, one fmaf per element.  tests/test_oracle_cpu.py::test_pool_c_matches_torch_sparse_mm
 * checks it BIT FOR BIT against torch.sparse.mm.
 *
 * Why it exists: torch.sparse.mm on the host takes ~7 s per application at M = 130k, 140 s for the
 * reference's 19 applications; the full-size parity test (tests/test_fullsize_gpu.py) needs the CPU side in well
 * under a minute.  Parity: unpinned by the reference (it ships no tests); pinned here against torch.sparse.mm itself.
 */
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#include <math.h>

/* out += w * src with one fused multiply-add per element (what torch's axpy does on an FMA machine) */
__attribute__((target("avx2,fma")))
static void axpy_row_fma(float *restrict out, const float *restrict src, float w, int64_t d) {
    for (int64_t j = 0; j < d; ++j) out[j] = __builtin_fmaf(w, src[j], out[j]);
}
static void axpy_row_soft(float *restrict out, const float *restrict src, float w, int64_t d) {
    for (int64_t j = 0; j < d; ++j) out[j] = fmaf(w, src[j], out[j]);   /* libm: exact but slow */
}
static void axpy_row(float *restrict out, const float *restrict src, float w, int64_t d) {
    static int have = -1;
    if (have < 0) have = __builtin_cpu_supports("avx2") && __builtin_cpu_supports("fma");
    if (have) axpy_row_fma(out, src, w, d);
    else axpy_row_soft(out, src, w, d);
}

/* idx [m,k] int64, w [m,k] f32, fin [m,d] f32 -> fout [m,d] f32 (one application).  Returns 0, or 1 on a bad index.
 * `order` (may be NULL) is only the order in which the independent output rows are VISITED (e.g. a Morton order of the
 * voxels, so that consecutive rows share neighbours and their input rows are still in cache: 4x faster at M = 130k);
 * every row is still accumulated k = 0..K-1, so the result does not depend on it. */
int gp_oracle_spmm(const int64_t *idx, const float *w, int64_t m, int k, const float *fin, int64_t d, float *fout,
                   const int64_t *order) {
    int bad = 0;
#pragma omp parallel for schedule(dynamic, 256)
    for (int64_t i = 0; i < m; ++i) {
        const int64_t r = order ? order[i] : i;
        float *o = fout + r * d;
        memset(o, 0, sizeof(float) * (size_t)d);
        for (int j = 0; j < k; ++j) {
            const int64_t c = idx[r * k + j];
            if (c < 0 || c >= m) {
                bad = 1;
                continue;
            }
            axpy_row(o, fin + c * d, w[r * k + j], d);
        }
    }
    return bad;
}
