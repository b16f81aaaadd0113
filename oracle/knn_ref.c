/* CPU oracle for SURVEY.md §8 row A9 -- TEST INFRASTRUCTURE, NOT PRODUCT.
 *
 * Exact brute-force k-nearest-voxel search restating
 *   index = faiss.IndexFlatL2(3); index.add(x); index.search(x, K+1)[:, 1:]
 * (reference models/affinity_module.py:1551-1557; faiss-cpu is a pip/conda dependency that is
 * neither vendored nor pinned, docs/Install.md:126-129).  IndexFlatL2 is an exhaustive scan
 * returning ascending squared-L2 distance; the order among equal distances is unspecified by
 * faiss, so the north star's rule is applied: ascending (d^2, index).  Column 0 is the query
 * itself (d^2 = 0, voxel coordinates are unique) and is dropped.
 *
 * Coordinates are integer voxel indices, so d^2 is an exact integer.  Parity: unpinned by the
 * reference (no tests); cross-checked against the numpy restatement in oracle.py.
 */
#include <stdint.h>
#include <stdlib.h>

static void sift_down(uint64_t *h, int n, int i) {
    for (;;) {
        int l = 2 * i + 1, r = l + 1, b = i;
        if (l < n && h[l] > h[b]) b = l;
        if (r < n && h[r] > h[b]) b = r;
        if (b == i) return;
        uint64_t t = h[i]; h[i] = h[b]; h[b] = t;
        i = b;
    }
}

static int cmp_u64(const void *a, const void *b) {
    uint64_t x = *(const uint64_t *)a, y = *(const uint64_t *)b;
    return (x > y) - (x < y);
}

/* coords: [m,3] int32; out: [m,k] int64.  Returns 0, or 1 if m < k+1. */
int gp_oracle_knn(const int32_t *coords, int64_t m, int k, int64_t *out) {
    if (m < (int64_t)k + 1) return 1;
    const int kk = k + 1;
#pragma omp parallel
    {
        uint64_t *heap = (uint64_t *)malloc(sizeof(uint64_t) * (size_t)kk);
#pragma omp for schedule(dynamic, 64)
        for (int64_t q = 0; q < m; ++q) {
            const int32_t qx = coords[3 * q], qy = coords[3 * q + 1], qz = coords[3 * q + 2];
            int n = 0;
            uint64_t worst = ~(uint64_t)0;
            for (int64_t j = 0; j < m; ++j) {
                int64_t dx = coords[3 * j] - qx, dy = coords[3 * j + 1] - qy, dz = coords[3 * j + 2] - qz;
                uint64_t key = ((uint64_t)(dx * dx + dy * dy + dz * dz) << 32) | (uint64_t)j;
                if (n < kk) {
                    heap[n++] = key;
                    if (n == kk) {
                        for (int i = kk / 2 - 1; i >= 0; --i) sift_down(heap, kk, i);
                        worst = heap[0];
                    }
                } else if (key < worst) {
                    heap[0] = key;
                    sift_down(heap, kk, 0);
                    worst = heap[0];
                }
            }
            qsort(heap, (size_t)kk, sizeof(uint64_t), cmp_u64);
            for (int i = 1; i < kk; ++i) out[q * k + (i - 1)] = (int64_t)(heap[i] & 0xffffffffu);
        }
        free(heap);
    }
    return 0;
}
