// Geometry-Guided Pooling on the tensor cores: F <- A F as a block-local dense GEMM (tcgen05 + TMEM, bf16x3 split).
// Replaces the cuSPARSE SpMM loop of the reference (models/affinity_module.py:1575-1587).
//
// In Hilbert-cell order the K-neighbour lists of consecutive voxels overlap heavily: 128 consecutive voxels reference
// only ~4.3 |tile| distinct rows instead of 128 K.  A per-scene PLAN stores, for every tile of 128 output voxels, the
// sorted union U of its neighbour lists and the tile's weights in a COMPACT form: the 128 K (weight, position) entries
// grouped by K-step (16 union rows), 8 bytes each -- bf16 hi | bf16 lo | byte offset inside the K-step's operand block.
// One application is
//   out_tile[128 x D] = W_tile[128 x |U|] * F[U, :]
// run as D/256 independent PASSES of 256 columns (work item = (tile, column half)), TWO CTAs per SM with 256 TMEM
// columns each, so that one CTA's epilogue (TMEM drain, bf16 split, stores) runs under the other CTA's MMAs.  Per
// K-step the dense 128 x 16 weight block (zeros where a voxel does not use a neighbour) is EXPANDED IN SHARED MEMORY
// from the compact entries by three otherwise idle warps (zero 8 KB, scatter ~360 entries; the entries of a warp's next
// K-step are loaded into registers while the current one waits for its stage); the rows F[U] are gathered once per work
// item by cp.async into the MN-major operand layout, every thread's copies arriving on the stage's mbarrier by
// themselves (cp.async.mbarrier.arrive.noinc), so all free stages have their rows in flight.  F travels between applications as bf16 hi/lo
// planes (16 mantissa bits, relative error 2^-17 per application); products use Whi*Fhi + Whi*Flo + Wlo*Fhi with fp32
// accumulation in TMEM.  HBM traffic per application: F planes in + out (2 x 4 m D bytes), the compact plan (8 m K
// bytes) and the union lists -- no dense weight block and no M x K fp32 tensor is re-read.
// Two plan formats, chosen at build time: GP_POOL_DENSE=1 (default) stores the expanded 128 x 16 bf16 hi/lo blocks and
// streams them with 8 KB bulk copies (300 MB of plan per scene, re-read by every application; faster: the shared-memory
// port, which the MMAs' operand reads already load to ~60 %, is spared the zero-fill + scatter of the expansion);
// GP_POOL_DENSE=0 is the compact format described above (100 MB of plan, 26 % less DRAM traffic per application,
// 17 % slower on B200 -- kept for A/B runs and for memory-tight scenes).
#include "common.cuh"
#include "tc.cuh"

#ifndef GP_POOL_DENSE
#define GP_POOL_DENSE 1   // measured (profiles/r02_*): dense plan 0.233 ms per application at M = 130k, compact 0.272 ms
#endif

#ifndef GP_POOL_EPI_DIRECT
#define GP_POOL_EPI_DIRECT 0
#endif

namespace {

constexpr bool COMPACT = GP_POOL_DENSE == 0;
constexpr int PT = 128;          // output voxels per tile = UMMA M = TMEM lanes
constexpr int PH = 256;          // feature columns per pass = UMMA N = TMEM columns of one CTA
constexpr int PK = 16;           // union rows per K-step = UMMA K (bf16)
constexpr int KSTEP_CAP = 128;   // K-steps a tile may use (2048 union rows); larger unions take the row-wise path
constexpr int A_PLANE = PT * PK * 2;     // bytes of one plane of the expanded weight block per K-step: 4 KB
constexpr int A_STEP = 2 * A_PLANE;      // hi + lo: 8 KB
constexpr int B_PLANE = PK * PH * 2;     // bytes of one plane of gathered rows per K-step: 8 KB
constexpr int PSTAGE = A_STEP + 2 * B_PLANE;  // 24 KB
constexpr int PNSTAGE = 4;
constexpr int P_EPI_LD = 20;     // floats per staged row of the 16-column transpose buffer (16-byte aligned)
constexpr int PTHREADS = 384;    // warps 0, 2, 3 weight expansion (dense build: warp 0 bulk copies), 1 MMA, 2 also TMEM
                                 // alloc, 4-7 epilogue, 8-11 row gather
constexpr int PSMEM = PNSTAGE * PSTAGE + 4 * 32 * P_EPI_LD * 4 + 256 + 1024;   // 107.3 KB: two CTAs per SM
constexpr int KOFS_LD = KSTEP_CAP + 4;   // per-tile K-step offset table (int32), KSTEP_CAP + 1 used

struct PlanLayout {
  int64_t tiles;
  size_t off_count, off_uidx, off_kofs, off_went, off_w, bytes;
};

inline PlanLayout plan_layout(int64_t m, int K) {
  PlanLayout L;
  L.tiles = (m + PT - 1) / PT;
  size_t off = 0;
  auto take = [&](size_t b) {
    size_t o = off;
    off += gp_align_up(b, 1024);
    return o;
  };
  L.off_count = take(sizeof(int32_t) * L.tiles);
  L.off_uidx = take(sizeof(int32_t) * L.tiles * KSTEP_CAP * PK);
  L.off_kofs = L.off_went = L.off_w = 0;
  if (COMPACT) {
    L.off_kofs = take(sizeof(int32_t) * L.tiles * KOFS_LD);
    L.off_went = take(sizeof(uint2) * L.tiles * PT * K);
  } else {
    L.off_w = take((size_t)L.tiles * KSTEP_CAP * A_STEP);
  }
  L.bytes = off;
  return L;
}

__device__ __forceinline__ void split_bf16(float v, __nv_bfloat16& hi, __nv_bfloat16& lo) {
  hi = __float2bfloat16_rn(v);
  lo = __float2bfloat16_rn(v - __bfloat162float(hi));
}

// byte offset of element (row r, k kk) inside one plane of a K-step's weight block: K-major, 32-byte rows,
// SWIZZLE_32B (16-byte chunk c of row r is stored at chunk c ^ ((r >> 2) & 1))
__device__ __forceinline__ uint32_t a_block_offset(int r, int kk) {
  return (uint32_t)(r * 32 + (((kk >> 3) ^ ((r >> 2) & 1)) << 4) + (kk & 7) * 2);
}

// ------------------------------------------------------------------ plan: one CTA per tile
// The union of the tile's 128 K neighbour ids is built in a shared-memory hash set (only ~5 % of the entries are
// distinct), compacted, and only the distinct ids are sorted (bitonic over <= 2048 keys instead of 16384).  The
// (voxel, neighbour) weights are then grouped by the K-step their union position falls into (counting sort in shared
// memory).  Precondition (true for kNN lists): the ids of one row are distinct -- a repeated id would be a second
// entry for the same (row, position) cell and only one of them would survive the expansion.
constexpr int PLAN_HASH = 4096;                 // slots; the union is capped at KSTEP_CAP * PK = 2048 ids
constexpr int PLAN_UCAP = KSTEP_CAP * PK;

__global__ void __launch_bounds__(256) pool_tc_plan_kernel(const int32_t* __restrict__ nbr,
                                                            const float* __restrict__ w, int64_t m, int K,
                                                            int32_t* __restrict__ ucount, int32_t* __restrict__ uidx,
                                                            int32_t* __restrict__ kofs_out, uint2* __restrict__ went,
                                                            uint8_t* __restrict__ wplan) {
  __shared__ uint32_t table[PLAN_HASH];
  __shared__ uint32_t keys[PLAN_UCAP];
  __shared__ int s_nu, s_over;
  const int64_t t = blockIdx.x;
  const int64_t row0 = t * PT;
  const int rows = (int)min((int64_t)PT, m - row0);
  const int nent = rows * K;
  for (int e = threadIdx.x; e < PLAN_HASH; e += 256) table[e] = 0xffffffffu;
  if (threadIdx.x == 0) {
    s_nu = 0;
    s_over = 0;
  }
  __syncthreads();
  for (int e = threadIdx.x; e < nent; e += 256) {
    const uint32_t v = (uint32_t)__ldg(nbr + row0 * K + e);
    uint32_t h = (v * 2654435761u) >> 20;  // 12 bits
    for (int probe = 0; probe < PLAN_HASH; ++probe) {
      const uint32_t old = atomicCAS(&table[h], 0xffffffffu, v);
      if (old == v) break;
      if (old == 0xffffffffu) {
        if (atomicAdd(&s_nu, 1) >= PLAN_UCAP) s_over = 1;
        break;
      }
      if (s_over) break;
      h = (h + 1) & (PLAN_HASH - 1);
    }
  }
  __syncthreads();
  if (s_over) {  // union larger than the tensor-core path takes: the row-wise kernel serves this tile
    if (threadIdx.x == 0) ucount[t] = PLAN_UCAP + 1;
    return;
  }
  const int nu = s_nu;
  __syncthreads();
  if (threadIdx.x == 0) s_nu = 0;
  int np2 = 32;
  while (np2 < nu) np2 <<= 1;
  for (int e = threadIdx.x; e < np2; e += 256) keys[e] = 0xffffffffu;
  __syncthreads();
  for (int e = threadIdx.x; e < PLAN_HASH; e += 256) {
    const uint32_t v = table[e];
    if (v != 0xffffffffu) keys[atomicAdd(&s_nu, 1)] = v;
  }
  __syncthreads();
  for (int k = 2; k <= np2; k <<= 1)
    for (int j = k >> 1; j > 0; j >>= 1) {
      for (int i = threadIdx.x; i < (np2 >> 1); i += 256) {
        int lo = ((i & ~(j - 1)) << 1) | (i & (j - 1));
        int hi = lo | j;
        uint32_t a = keys[lo], b = keys[hi];
        bool up = (lo & k) == 0;
        if ((a > b) == up) {
          keys[lo] = b;
          keys[hi] = a;
        }
      }
      __syncthreads();
    }
  if (threadIdx.x == 0) ucount[t] = nu;
  const int ksteps = (nu + PK - 1) / PK;
  if (ksteps > KSTEP_CAP) return;  // this tile is served by the row-wise kernel
  int32_t* ui = uidx + t * (int64_t)(KSTEP_CAP * PK);
  for (int e = threadIdx.x; e < ksteps * PK; e += 256) ui[e] = (int32_t)keys[e < nu ? e : nu - 1];  // pad: weight 0
  auto position = [&](uint32_t v) {
    int lo = 0, hi = nu - 1;
    while (lo < hi) {
      int mid = (lo + hi) >> 1;
      if (keys[mid] < v) lo = mid + 1;
      else hi = mid;
    }
    return lo;
  };
  if (COMPACT) {
    // counting sort of the entries by K-step; the hash table's storage is free again
    int* cnt = (int*)table;              // [KSTEP_CAP] entries per K-step, then running fill
    int* start = cnt + KSTEP_CAP;        // [KSTEP_CAP + 1] exclusive offsets
    for (int e = threadIdx.x; e < KSTEP_CAP; e += 256) cnt[e] = 0;
    __syncthreads();
    for (int e = threadIdx.x; e < nent; e += 256)
      atomicAdd(&cnt[position((uint32_t)__ldg(nbr + row0 * K + e)) >> 4], 1);
    __syncthreads();
    if (threadIdx.x < 32) {  // exclusive scan of <= 128 counts by one warp
      int carry = 0;
      for (int base = 0; base < KSTEP_CAP; base += 32) {
        const int c = cnt[base + threadIdx.x];
        int x = c;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
          const int y = __shfl_up_sync(0xffffffffu, x, o);
          if ((int)threadIdx.x >= o) x += y;
        }
        start[base + threadIdx.x] = carry + x - c;
        carry += __shfl_sync(0xffffffffu, x, 31);
      }
      if (threadIdx.x == 0) start[KSTEP_CAP] = carry;
    }
    __syncthreads();
    int32_t* ko = kofs_out + t * (int64_t)KOFS_LD;
    for (int e = threadIdx.x; e <= KSTEP_CAP; e += 256) ko[e] = start[e];
    for (int e = threadIdx.x; e < KSTEP_CAP; e += 256) cnt[e] = 0;
    __syncthreads();
    uint2* te = went + t * (int64_t)PT * K;
    for (int e = threadIdx.x; e < nent; e += 256) {
      const int pos = position((uint32_t)__ldg(nbr + row0 * K + e));
      const int ks = pos >> 4;
      const int slot = start[ks] + atomicAdd(&cnt[ks], 1);
      __nv_bfloat16 h, l;
      split_bf16(__ldg(w + row0 * K + e), h, l);
      te[slot] = make_uint2(((uint32_t)__bfloat16_as_ushort(h) << 16) | (uint32_t)__bfloat16_as_ushort(l),
                            a_block_offset(e / K, pos & 15));
    }
  } else {
    uint8_t* wp_base = wplan + (size_t)t * KSTEP_CAP * A_STEP;
    for (int e = threadIdx.x; e < ksteps * (A_STEP / 16); e += 256) ((uint4*)wp_base)[e] = make_uint4(0, 0, 0, 0);
    __syncthreads();
    for (int e = threadIdx.x; e < nent; e += 256) {
      const int pos = position((uint32_t)__ldg(nbr + row0 * K + e));
      __nv_bfloat16 h, l;
      split_bf16(__ldg(w + row0 * K + e), h, l);
      uint8_t* p = wp_base + (size_t)(pos >> 4) * A_STEP + a_block_offset(e / K, pos & 15);
      *(__nv_bfloat16*)p = h;
      *(__nv_bfloat16*)(p + A_PLANE) = l;
    }
  }
}

// fp32 rows -> bf16 hi/lo planes [m][2][D]
__global__ void pool_tc_split_kernel(const float* __restrict__ f, int64_t stride, int64_t m, int D,
                                     __nv_bfloat16* __restrict__ planes) {
  const int dq = D / 4;
  int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= m * dq) return;
  int64_t r = t / dq;
  int c = (int)(t - r * dq) * 4;
  float4 v = *(const float4*)(f + r * stride + c);
  __nv_bfloat16 h[4], l[4];
  split_bf16(v.x, h[0], l[0]);
  split_bf16(v.y, h[1], l[1]);
  split_bf16(v.z, h[2], l[2]);
  split_bf16(v.w, h[3], l[3]);
  *(uint2*)(planes + r * 2 * D + c) = *(const uint2*)h;
  *(uint2*)(planes + r * 2 * D + D + c) = *(const uint2*)l;
}


// ------------------------------------------------------------------ one application: out = W_plan * F[U]
struct PoolParams {
  const __nv_bfloat16* fin;   // [m][2][D] hi | lo planes
  const int32_t* ucount;      // [tiles]
  const int32_t* uidx;        // [tiles][KSTEP_CAP * PK]
  const int32_t* kofs;        // [tiles][KOFS_LD]           (compact plan)
  const uint2* went;          // [tiles][PT * K]  x = hi << 16 | lo, y = byte offset inside the K-step's plane
  const uint8_t* wplan;       // [tiles][KSTEP_CAP][A_STEP] (dense plan)
  int64_t m;
  int tiles, K, D, halves;
  __nv_bfloat16* out_planes;  // next application's input, or nullptr
  float* out_f32;             // [m][D] final result, or nullptr
};

__global__ void __launch_bounds__(PTHREADS, 2) pool_tc_kernel(const PoolParams P) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  float* epi = (float*)(smem + PNSTAGE * PSTAGE);
  uint64_t* bars = (uint64_t*)(smem + PNSTAGE * PSTAGE + 4 * 32 * P_EPI_LD * 4);
  uint64_t* full = bars;
  uint64_t* empty = bars + PNSTAGE;
  uint64_t* tmem_full = bars + 2 * PNSTAGE;
  uint64_t* tmem_empty = tmem_full + 1;
  uint32_t* tmem_slot = (uint32_t*)(tmem_empty + 1);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int items = P.tiles * P.halves;   // work item = (tile, 256-column half)

  if (threadIdx.x == 0) {
    for (int s = 0; s < PNSTAGE; ++s) {
      tc::mbar_init(&full[s], 128 + 1);  // every gather thread's cp.async completion + the weight block (expander
                                         // warp's lane 0 / bulk copy expect_tx)
      tc::mbar_init(&empty[s], 1);
    }
    tc::mbar_init(tmem_full, 1);
    tc::mbar_init(tmem_empty, 4);
    tc::fence_barrier_init();
  }
  if (warp == 2) tc::tmem_alloc<PH>(tmem_slot);
  tc::tc_fence_before();
  __syncthreads();
  tc::tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  auto ksteps_of = [&](int t) -> int {
    int ks = (__ldg(P.ucount + t) + PK - 1) / PK;
    return ks > KSTEP_CAP ? 0 : ks;  // overflow tiles are skipped here (row-wise kernel)
  };

  if (COMPACT && (warp == 0 || warp == 2 || warp == 3)) {
    // ===================== weight blocks: expanded in shared memory from the compact entries =====================
    // three warps take the K-steps round robin; each zeroes the 8 KB block, then scatters the step's entries.  The
    // entries of the warp's NEXT K-step are loaded (registers, PRE per lane) right after the current block is handed
    // over, so their latency runs under the two K-steps the other warps serve.
    const int eid = warp == 0 ? 0 : warp - 1;
    constexpr int PRE = 14;  // 448 entries per K-step in registers (mean ~360 at K = 96); the rest is loaded late
    int wi = blockIdx.x, t = 0, ks = -1, ks_n = 0, e0 = 0, e1 = 0;
    uint32_t it = 0xffffffffu;   // K-steps of this CTA so far, minus one
    const uint2* te = nullptr;
    uint2 pe[PRE];
    bool first = true;
    auto advance = [&]() -> bool {   // next K-step with it % 3 == eid; loads its first PRE * 32 entries
      for (;;) {
        if (first) {
          first = false;
          if (wi >= items) return false;
          t = wi / P.halves;
          ks_n = ksteps_of(t);
          ks = -1;
        }
        ++ks;
        while (ks >= ks_n) {
          wi += gridDim.x;
          if (wi >= items) return false;
          t = wi / P.halves;
          ks_n = ksteps_of(t);
          ks = 0;
        }
        ++it;
        if ((int)(it % 3) == eid) break;
      }
      const int32_t* ko = P.kofs + (int64_t)t * KOFS_LD;
      te = P.went + (int64_t)t * PT * P.K;
      e0 = __ldg(ko + ks);
      e1 = __ldg(ko + ks + 1);
#pragma unroll
      for (int j = 0; j < PRE; ++j) {
        const int e = e0 + j * 32 + lane;
        pe[j] = e < e1 ? __ldg(te + e) : make_uint2(0u, 0xffffffffu);
      }
      return true;
    };
    bool have = advance();
    while (have) {
      const int s = it % PNSTAGE;
      tc::mbar_wait(&empty[s], ((it / PNSTAGE) & 1) ^ 1);
      uint8_t* a = smem + s * PSTAGE;
#pragma unroll
      for (int j = 0; j < A_STEP / 16 / 32; ++j) ((uint4*)a)[j * 32 + lane] = make_uint4(0, 0, 0, 0);
      __syncwarp();
#pragma unroll
      for (int j = 0; j < PRE; ++j)
        if (pe[j].y != 0xffffffffu) {
          *(uint16_t*)(a + pe[j].y) = (uint16_t)(pe[j].x >> 16);
          *(uint16_t*)(a + A_PLANE + pe[j].y) = (uint16_t)(pe[j].x & 0xffffu);
        }
      for (int e = e0 + PRE * 32 + lane; e < e1; e += 32) {
        const uint2 v = __ldg(te + e);
        *(uint16_t*)(a + v.y) = (uint16_t)(v.x >> 16);
        *(uint16_t*)(a + A_PLANE + v.y) = (uint16_t)(v.x & 0xffffu);
      }
      tc::fence_proxy_async();
      __syncwarp();
      if (lane == 0) tc::mbar_arrive(&full[s]);
      have = advance();
    }
  } else if (!COMPACT && warp == 0) {
    // ===================== dense plan: one 8 KB bulk copy per K-step =====================
    if (lane == 0) {
      uint32_t it = 0;
      for (int wi = blockIdx.x; wi < items; wi += gridDim.x) {
        const int t = wi / P.halves;
        const int ks_n = ksteps_of(t);
        const uint8_t* src = P.wplan + (size_t)t * KSTEP_CAP * A_STEP;
        for (int ks = 0; ks < ks_n; ++ks, ++it) {
          const int s = it % PNSTAGE;
          tc::mbar_wait(&empty[s], ((it / PNSTAGE) & 1) ^ 1);
          tc::mbar_arrive_expect_tx(&full[s], A_STEP);
          tc::bulk_load(tc::smem_u32(smem + s * PSTAGE), src + (size_t)ks * A_STEP, A_STEP, &full[s]);
        }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer =====================
    constexpr uint64_t ADESC = tc::umma_desc_kmajor_base<32>();
    constexpr uint64_t BDESC = tc::umma_desc_mnmajor_sw128_base(2048, 1024);
    constexpr uint32_t IDESC = tc::umma_idesc_bf16(PT, PH, true);
    uint32_t it = 0, tcount = 0;
    for (int wi = blockIdx.x; wi < items; wi += gridDim.x) {
      const int ks_n = ksteps_of(wi / P.halves);
      if (ks_n == 0) continue;
      tc::mbar_wait(tmem_empty, (tcount & 1) ^ 1);
      tc::tc_fence_after();
      for (int ks = 0; ks < ks_n; ++ks, ++it) {
        const int s = it % PNSTAGE;
        tc::mbar_wait(&full[s], (it / PNSTAGE) & 1);
        tc::fence_proxy_async();   // the gather threads' cp.async writes (generic proxy) before the async-proxy reads
        tc::tc_fence_after();
        if (lane == 0) {
          const uint32_t a_hi = tc::smem_u32(smem + s * PSTAGE);
          const uint32_t a_lo = a_hi + A_PLANE;
          const uint32_t b_hi = a_hi + A_STEP;
          const uint32_t b_lo = b_hi + B_PLANE;
          // small terms first, then the leading product
          tc::umma_bf16(tmem_base, tc::umma_desc(ADESC, a_lo), tc::umma_desc(BDESC, b_hi), IDESC, ks > 0);
          tc::umma_bf16(tmem_base, tc::umma_desc(ADESC, a_hi), tc::umma_desc(BDESC, b_lo), IDESC, 1);
          tc::umma_bf16(tmem_base, tc::umma_desc(ADESC, a_hi), tc::umma_desc(BDESC, b_hi), IDESC, 1);
          tc::umma_commit(&empty[s]);
          if (ks == ks_n - 1) tc::umma_commit(tmem_full);
        }
        __syncwarp();
      }
      ++tcount;
    }
  } else if (warp >= 4 && warp < 8) {
    // ===================== epilogue: TMEM -> registers -> 16-column smem transpose -> coalesced rows =====================
    const int q = warp & 3;
    [[maybe_unused]] float* buf = epi + q * 32 * P_EPI_LD;
    uint32_t tcount = 0;
    for (int wi = blockIdx.x; wi < items; wi += gridDim.x) {
      const int t = wi / P.halves, half = wi - t * P.halves;
      if (ksteps_of(t) == 0) continue;
      tc::mbar_wait(tmem_full, tcount & 1);
      tc::tc_fence_after();
      const int64_t row_base = (int64_t)t * PT + q * 32;
      const int rows_here = (int)min((int64_t)32, P.m - row_base);
#if GP_POOL_EPI_DIRECT
      // accumulator-fragment loads (tcgen05.ld.16x256b: four threads hold one 32-byte sector of a row), stored straight
      // from the registers: no shared-memory transpose, no __syncwarp between TMEM reads, so the accumulator is handed
      // back to the MMA warp sooner
#pragma unroll 1
      for (int hb = 0; hb < 2; ++hb) {
        const int ra = hb * 16 + (lane >> 2), rb = ra + 8;
        const bool oka = ra < rows_here, okb = rb < rows_here;
        const int c0 = half * PH + 2 * (lane & 3);
#pragma unroll 1
        for (int cg = 0; cg < PH / 64; ++cg) {
          uint32_t v[32];
          tc::tmem_ld_16x256b_x8(tmem_base + ((uint32_t)(q * 32 + hb * 16) << 16) + cg * 64, v);
          tc::tmem_ld_wait();
          if (hb == 1 && cg == PH / 64 - 1) {
            tc::tc_fence_before();
            __syncwarp();
            if (lane == 0) tc::mbar_arrive(tmem_empty);
          }
#pragma unroll
          for (int g = 0; g < 8; ++g) {
            const int col = c0 + cg * 64 + g * 8;
#pragma unroll
            for (int r2 = 0; r2 < 2; ++r2) {
              if (r2 == 0 ? oka : okb) {
                const int64_t row = row_base + (r2 == 0 ? ra : rb);
                const float x = __uint_as_float(v[4 * g + 2 * r2]), y = __uint_as_float(v[4 * g + 2 * r2 + 1]);
                if (P.out_f32) *(float2*)(P.out_f32 + row * P.D + col) = make_float2(x, y);
                if (P.out_planes) {
                  __nv_bfloat16 h[2], l[2];
                  split_bf16(x, h[0], l[0]);
                  split_bf16(y, h[1], l[1]);
                  __nv_bfloat16* dst = P.out_planes + row * 2 * P.D + col;
                  *(uint32_t*)dst = *(const uint32_t*)h;
                  *(uint32_t*)(dst + P.D) = *(const uint32_t*)l;
                }
              }
            }
          }
        }
      }
      ++tcount;
      continue;
#endif
      for (int ch = 0; ch < PH / 32; ++ch) {
        uint32_t v[32];
        tc::tmem_ld32(tmem_base + ((uint32_t)(q * 32) << 16) + ch * 32, v);
        tc::tmem_ld_wait();
        if (ch == PH / 32 - 1) {
          tc::tc_fence_before();
          __syncwarp();
          if (lane == 0) tc::mbar_arrive(tmem_empty);
        }
#pragma unroll
        for (int hh = 0; hh < 2; ++hh) {
#pragma unroll
          for (int j = 0; j < 4; ++j)
            *(float4*)(buf + lane * P_EPI_LD + j * 4) =
                make_float4(__uint_as_float(v[16 * hh + 4 * j]), __uint_as_float(v[16 * hh + 4 * j + 1]),
                            __uint_as_float(v[16 * hh + 4 * j + 2]), __uint_as_float(v[16 * hh + 4 * j + 3]));
          __syncwarp();
#pragma unroll
          for (int rr = 0; rr < 4; ++rr) {
            const int row = rr * 8 + (lane >> 2), c4 = lane & 3;
            if (row < rows_here) {
              float4 o = *(const float4*)(buf + row * P_EPI_LD + c4 * 4);
              const int col = half * PH + ch * 32 + hh * 16 + c4 * 4;
              if (P.out_f32) *(float4*)(P.out_f32 + (row_base + row) * P.D + col) = o;
              if (P.out_planes) {
                __nv_bfloat16 h[4], l[4];
                split_bf16(o.x, h[0], l[0]);
                split_bf16(o.y, h[1], l[1]);
                split_bf16(o.z, h[2], l[2]);
                split_bf16(o.w, h[3], l[3]);
                __nv_bfloat16* dst = P.out_planes + (row_base + row) * 2 * P.D + col;
                *(uint2*)dst = *(const uint2*)h;
                *(uint2*)(dst + P.D) = *(const uint2*)l;
              }
            }
          }
          __syncwarp();
        }
      }
      ++tcount;
    }
  } else if (warp >= 8) {
    // ===================== gather of the union rows into the MN-major operand layout =====================
    // warp w owns union rows k = 4w .. 4w+3 of the K-step.  One warp-wide cp.async moves the 512 contiguous bytes of
    // one row plane's column half (4 cache lines).  16-byte chunk q (0..31) = column block nb = q >> 3, chunk c = q & 7
    // inside the 128-byte atom row: atom (nb, k >> 3) at ((nb * 2 + (k >> 3)) * 1024), row (k & 7) of 128 bytes, chunk
    // c stored at c ^ (k & 7).  Completion is asynchronous (cp.async.mbarrier.arrive.noinc per thread).
    const int w = warp - 8;
    uint32_t it = 0;
    for (int wi = blockIdx.x; wi < items; wi += gridDim.x) {
      const int t = wi / P.halves, half = wi - t * P.halves;
      const int ks_n = ksteps_of(t);
      const int32_t* ui = P.uidx + (int64_t)t * (KSTEP_CAP * PK) + 4 * w + (lane & 3);
      int u_next = ks_n > 0 ? __ldg(ui) : 0;
      const __nv_bfloat16* fbase = P.fin + half * PH + lane * 8;
      for (int ks = 0; ks < ks_n; ++ks, ++it) {
        const int s = it % PNSTAGE;
        const int u_cur = u_next;
        if (ks + 1 < ks_n) u_next = __ldg(ui + (ks + 1) * PK);  // the next K-step's row ids, one iteration ahead
        tc::mbar_wait(&empty[s], ((it / PNSTAGE) & 1) ^ 1);
        const uint32_t stage = tc::smem_u32(smem + s * PSTAGE) + A_STEP;
#pragma unroll
        for (int kk = 0; kk < 4; ++kk) {
          const int k = 4 * w + kk;
          const int u = __shfl_sync(0xffffffffu, u_cur, kk);
          const __nv_bfloat16* src = fbase + (int64_t)u * 2 * P.D;
          const uint32_t dst = stage + (lane >> 3) * 2048 + (k >> 3) * 1024 + (k & 7) * 128 +
                               ((((lane & 7) ^ (k & 7))) << 4);
          tc::cp_async16(dst, src, 16);
          tc::cp_async16(dst + B_PLANE, src + P.D, 16);
        }
        tc::cp_async_arrive_noinc(&full[s]);   // arrives when this thread's copies above have landed
      }
    }
    tc::cp_async_wait<0>();
  }

  tc::tc_fence_before();
  __syncthreads();
  if (warp == 2) tc::tmem_dealloc<PH>(tmem_base);
}

// tiles whose union exceeds KSTEP_CAP K-steps: one warp per output row straight from the neighbour lists
template <int NV>   // D = 32 * NV
__global__ void __launch_bounds__(256) pool_rows_planes_kernel(const __nv_bfloat16* __restrict__ fin,
                                                                const int32_t* __restrict__ ucount,
                                                                const int32_t* __restrict__ nbr,
                                                                const float* __restrict__ w, int64_t m, int K,
                                                                __nv_bfloat16* __restrict__ out_planes,
                                                                float* __restrict__ out_f32) {
  constexpr int D = 32 * NV;
  const int t = blockIdx.x;
  if ((__ldg(ucount + t) + PK - 1) / PK <= KSTEP_CAP) return;
  const int lane = threadIdx.x & 31;
  for (int rr = threadIdx.x >> 5; rr < PT; rr += 8) {
    const int64_t s = (int64_t)t * PT + rr;
    if (s >= m) break;
    float acc[NV];
#pragma unroll
    for (int j = 0; j < NV; ++j) acc[j] = 0.f;
    for (int k = 0; k < K; ++k) {
      const int u = __ldg(nbr + s * K + k);
      const float wk = __ldg(w + s * K + k);
      const __nv_bfloat16* row = fin + (int64_t)u * 2 * D;
#pragma unroll
      for (int j = 0; j < NV; ++j) {
        int c = j * 32 + lane;
        acc[j] = fmaf(wk, __bfloat162float(row[c]) + __bfloat162float(row[D + c]), acc[j]);
      }
    }
#pragma unroll
    for (int j = 0; j < NV; ++j) {
      int c = j * 32 + lane;
      if (out_f32) out_f32[s * D + c] = acc[j];
      if (out_planes) {
        __nv_bfloat16 h, l;
        split_bf16(acc[j], h, l);
        out_planes[s * 2 * D + c] = h;
        out_planes[s * 2 * D + D + c] = l;
      }
    }
  }
}

}  // namespace

// ------------------------------------------------------------------ host side
extern "C" size_t gp_pool_plan_bytes(int64_t m, int K) {
  if (m < 0 || K < 1 || K > 128) return 0;
  return plan_layout(m, K).bytes;
}

extern "C" int gp_pool_plan(const int32_t* nbr_sorted, const float* weights, int64_t m, int K, void* plan,
                            size_t plan_bytes, gp_stream_t stream) {
  GP_CHECK_ARG(m >= 0 && K >= 1 && K <= 128, "gp_pool_plan: bad sizes");
  GP_CHECK_ARG(((uintptr_t)plan % 256) == 0, "gp_pool_plan: plan must be 256-byte aligned");
  PlanLayout L = plan_layout(m, K);
  if (plan_bytes < L.bytes) {
    gp_set_error("gp_pool_plan: plan buffer %zu < %zu", plan_bytes, L.bytes);
    return GP_ERR_WORKSPACE;
  }
  if (m == 0) return GP_OK;
  char* p = (char*)plan;
  const bool tg = gp_timing_begin(GP_T_POOL_PLAN, (cudaStream_t)stream);
  pool_tc_plan_kernel<<<(unsigned)L.tiles, 256, 0, (cudaStream_t)stream>>>(
      nbr_sorted, weights, m, K, (int32_t*)(p + L.off_count), (int32_t*)(p + L.off_uidx), (int32_t*)(p + L.off_kofs),
      (uint2*)(p + L.off_went), (uint8_t*)(p + L.off_w));
  if (tg) gp_timing_end(GP_T_POOL_PLAN, (cudaStream_t)stream);
  GP_LAUNCH_CHECK();
  return GP_OK;
}

extern "C" size_t gp_pool_tc_workspace_bytes(int64_t m) {
  if (m < 0) return 0;
  return 2 * gp_align_up(sizeof(__nv_bfloat16) * 2 * 512 * (size_t)m, 1024);
}

extern "C" int gp_pool_tc(const float* f0, int64_t stride, int D, const void* plan, const int32_t* nbr_sorted,
                          const float* weights, int64_t m, int K, int T, float* out, void* workspace,
                          size_t workspace_bytes, gp_stream_t stream) {
  GP_CHECK_ARG(m >= 0 && K >= 1 && K <= 128 && T >= 1, "gp_pool_tc: bad sizes (T >= 1)");
  GP_CHECK_ARG(D == 512 || D == 256, "gp_pool_tc: D must be 256 or 512 (got %d)", D);
  GP_CHECK_ARG(stride % 4 == 0 && stride >= D && ((uintptr_t)f0 % 16) == 0 && ((uintptr_t)out % 16) == 0,
               "gp_pool_tc: alignment");
  GP_CHECK_ARG(((uintptr_t)workspace % 256) == 0 && ((uintptr_t)plan % 256) == 0, "gp_pool_tc: 256-byte alignment");
  if (workspace_bytes < gp_pool_tc_workspace_bytes(m)) {
    gp_set_error("gp_pool_tc: workspace %zu < %zu", workspace_bytes, gp_pool_tc_workspace_bytes(m));
    return GP_ERR_WORKSPACE;
  }
  if (m == 0) return GP_OK;
  cudaStream_t st = (cudaStream_t)stream;
  static int sm_count = 0;
  if (!sm_count) {
    int dev = 0;
    GP_CUDA(cudaGetDevice(&dev));
    GP_CUDA(cudaDeviceGetAttribute(&sm_count, cudaDevAttrMultiProcessorCount, dev));
    GP_CUDA(cudaFuncSetAttribute(pool_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, PSMEM));
    GP_CUDA(cudaFuncSetAttribute(pool_tc_kernel, cudaFuncAttributePreferredSharedMemoryCarveout,
                                 cudaSharedmemCarveoutMaxShared));
  }
  PlanLayout L = plan_layout(m, K);
  const char* p = (const char*)plan;
  __nv_bfloat16* planes[2];
  planes[0] = (__nv_bfloat16*)workspace;
  planes[1] = (__nv_bfloat16*)((char*)workspace + gp_align_up(sizeof(__nv_bfloat16) * 2 * 512 * (size_t)m, 1024));
  {
    int64_t tot = m * (D / 4);
    pool_tc_split_kernel<<<(unsigned)((tot + 255) / 256), 256, 0, st>>>(f0, stride, m, D, planes[0]);
    GP_LAUNCH_CHECK();
  }
  PoolParams P;
  P.ucount = (const int32_t*)(p + L.off_count);
  P.uidx = (const int32_t*)(p + L.off_uidx);
  P.kofs = (const int32_t*)(p + L.off_kofs);
  P.went = (const uint2*)(p + L.off_went);
  P.wplan = (const uint8_t*)(p + L.off_w);
  P.m = m;
  P.tiles = (int)L.tiles;
  P.K = K;
  P.D = D;
  P.halves = D / PH;
  const int items = P.tiles * P.halves;
  int grid = items < 2 * sm_count ? items : 2 * sm_count;   // two CTAs (256 TMEM columns each) per SM
  for (int t = 0; t < T; ++t) {
    const bool last = t == T - 1;
    P.fin = planes[t & 1];
    P.out_planes = last ? nullptr : planes[(t + 1) & 1];
    P.out_f32 = last ? out : nullptr;
    const bool tg = gp_timing_begin(GP_T_POOL_APPLY, st);
    pool_tc_kernel<<<grid, PTHREADS, PSMEM, st>>>(P);
    if (tg) gp_timing_end(GP_T_POOL_APPLY, st);
    GP_LAUNCH_CHECK();
    if (D == 512)
      pool_rows_planes_kernel<16><<<P.tiles, 256, 0, st>>>(P.fin, P.ucount, nbr_sorted, weights, m, K, P.out_planes,
                                                           P.out_f32);
    else
      pool_rows_planes_kernel<8><<<P.tiles, 256, 0, st>>>(P.fin, P.ucount, nbr_sorted, weights, m, K, P.out_planes,
                                                          P.out_f32);
    GP_LAUNCH_CHECK();
  }
  return GP_OK;
}
