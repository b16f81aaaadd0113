// Student affinity network on the 5th-gen tensor cores (tcgen05 + TMEM), fp32-faithful through a bf16x3 split.
// Replaces AffinityPredictor.forward on MinkowskiEngine (reference models/affinity_module.py:33-72, 1541-1547).
//
// A sparse 3^3 convolution is  out[i] = sum_o W_o^T in[nbr(i,o)]  over the occupied offsets only (~4.9 of 27 per
// voxel on ScanNet-density surfaces), so the work is organised by (output, input) PAIRS, not by output tiles:
//   1. pair lists: for every slab of SLAB output voxels and every offset o, the pairs (i, nbr(i,o)) in voxel order
//      (flags -> exclusive scan -> fill); 128-pair tiles never straddle an offset.
//   2. gather-GEMM (this file's main kernel): C[p, :] = in[pair_in[p], :] W_o for every pair p, one 128 x N tile per
//      CTA iteration.  A rows are gathered with cp.async (zero-filled tails) from activations stored as bf16 hi/lo
//      planes, in K=64 chunks laid out as 128-byte-swizzled K-major rows: 8 consecutive lanes fetch one 128-byte
//      line of one row, so a warp-wide cp.async touches 4 cache lines (the first version fetched 16 bytes from 32
//      different rows per instruction and was bound by the L1TEX tag stage at 85 %).  W_o^T slices arrive by TMA
//      (32-byte swizzle) from a pre-tiled weight image in which every K=16 box is contiguous; a stage holds GP_STU_UK = 4
//      K-steps of one 256-column half, so the MMA issuer pays one wait + fence + commit per 12 MMAs.  Each K=16 step
//      issues  Alo*Bhi + Ahi*Blo + Ahi*Bhi  into a 128 x N fp32 accumulator in TMEM (two 256-column halves with their
//      own full/empty barriers); the epilogue warps drain TMEM to a pair-major staging buffer straight from the
//      accumulator fragments (tcgen05.ld.16x256b, float2 sector stores).
//   3. reduce (student_reduce_kernel): out[i] = sum of its staged pair rows in ascending offset order (the oracle's
//      order, deterministic, no atomics), BN scale/shift, residual, ReLU, then the bf16 hi/lo split for the next layer.
#include <cub/device/device_scan.cuh>
#include "common.cuh"
#include "tc.cuh"
#include "topk.cuh"

namespace {

constexpr int TM = 128;        // pairs (rows) per tile = UMMA M
constexpr int BK = 16;         // K per stage = UMMA K for bf16; 32-byte operand rows, SWIZZLE_32B
constexpr int ROWB = BK * 2;   // bytes per operand row per stage
#ifndef GP_STU_CL
#define GP_STU_CL 2
#endif
#ifndef GP_STU_NSTAGE
#define GP_STU_NSTAGE 3
#define GP_STU_AK 64
#define GP_STU_NA (GP_STU_CL == 2 ? 4 : 2)
#endif
#ifndef GP_STU_UK
#define GP_STU_UK 4  // K=16 steps per B stage = per barrier hand-off of the MMA issuer: 12 MMAs per wait + fence + commit.
                     // Measured (student forward, N = 150k scene): UK 1 / 2 / 4 = 11.55 / 11.06 / 10.78 ms
#endif
#ifndef GP_STU_EDGE
#define GP_STU_EDGE 0  // A chunks at each END of a tile's K loop that run as two passes (half 0, then half 1); see for_units.
                       // 0 = both halves per unit throughout.  Measured: 0 / 2 / 3 chunks = 10.78 / 10.79 / 11.05 ms
#endif
#ifndef GP_STU_EPI_DIRECT
#define GP_STU_EPI_DIRECT 1
#endif
#ifndef GP_STU_PROF
#define GP_STU_PROF 0  // 1: diagnostic build; CTA 0 prints where its MMA issuer, epilogue warp 4 and gather warp 8 waited
#endif
#if GP_STU_PROF
#define PROF_T0() const long long pt0_ = clock64()
#define PROF_ADD(acc) acc += clock64() - pt0_
#else
#define PROF_T0()
#define PROF_ADD(acc)
#endif
#ifndef GP_STU_EPI_QUAD
#define GP_STU_EPI_QUAD 0  // direct epilogue: quad transpose by shuffles, 64 contiguous bytes per quad and STG.128.
                           // Measured: 11.7 vs 10.8 ms -- the shuffle latency chains cost more than the wider stores give
#endif
#ifndef GP_STU_EPI_PIPE
#define GP_STU_EPI_PIPE 1  // direct epilogue: next TMEM fragment load in flight while the current one is stored
#endif
#ifndef GP_STU_ABLATE
#define GP_STU_ABLATE 0  // 1 / 2: diagnostic builds (wrong results) whose epilogue skips the stores / the TMEM loads too
#endif
constexpr int NSTAGE = GP_STU_NSTAGE;  // B (weight) stages: UK K-steps of ONE column half (<= 256 outputs), hi + lo planes
constexpr int UK = GP_STU_UK;          // K=16 steps per stage (a divisor of AKS, so a stage never straddles an A chunk)
constexpr int AK = GP_STU_AK;  // K per A chunk: 64- or 128-byte operand rows, SWIZZLE_64B / SWIZZLE_128B
constexpr int AKS = AK / BK;   // K=16 steps per A chunk
constexpr int NA = GP_STU_NA;  // A chunk buffers
constexpr int EDGE = GP_STU_EDGE < NA ? GP_STU_EDGE : NA - 1;  // an edge block keeps EDGE (+1 when K is not a multiple of
                                                               // the chunk) buffers resident; the gather hands over its
                                                               // lagged chunk before it blocks, so NA buffers suffice

// Order of the (column half h, K-steps) units of one tile, shared by the TMA producer and the MMA issuer.
// N <= 256 (NH = 1): K ascending.  N = 512 (NH = 2 halves of 256 columns, each with its own full/empty barriers):
//   EDGE = 0 (default): both halves per unit position, K ascending -- the two accumulators alternate every unit;
//   EDGE > 0 (build option): the first EDGE chunks run as two passes (all their K-steps into half 0, then the same,
//   still resident, A chunks into half 1), the middle alternates, the last EDGE chunks run as two passes again, so
//   that half 0 is complete one pass before half 1 and drains under half 1's last pass, and half 1 drains under the
//   next tile's first pass.  Why it exists: an SM stores at most 32 B/clk (scripts/probes/store_bw.cu; ~17 B/clk inside
//   this kernel, where the gathers share the path), so draining a 128 x 512 fp32 accumulator takes 8-15k clk of a
//   ~40k-clk tile and the tensor pipe idles meanwhile (builds without the stores: -2.1 ms of a 10.9 ms forward).  Why it
//   is off: hiding a 7.5k-clk half drain needs passes of >= 4 chunks, i.e. >= 6 resident A buffers (192 KB) -- with the
//   2-3 chunks that fit, the issuer still waits ~9k clk per tile on tmem_empty (GP_STU_PROF=1) and nothing is gained.
// A unit is UK consecutive K-steps (the last one of a tile may be shorter): f(h, first K-step, number of K-steps).
template <int NH, typename F>
__device__ __forceinline__ void for_units(int ksteps, F&& f) {
  static_assert(AKS % UK == 0, "a B stage must not straddle an A chunk");
  auto nk = [&](int ks, int end) { return end - ks < UK ? end - ks : UK; };
  if (NH == 1 || EDGE <= 0) {
    for (int ks = 0; ks < ksteps; ks += UK)
      for (int h = 0; h < NH; ++h) f(h, ks, nk(ks, ksteps));
    return;
  }
  int ka = EDGE * AKS < ksteps ? EDGE * AKS : ksteps;                       // end of the first edge block
  int kb = ksteps - EDGE * AKS > ka ? (ksteps - EDGE * AKS) / AKS * AKS : ka;  // start of the last one (chunk-aligned)
  for (int h = 0; h < NH; ++h)
    for (int ks = 0; ks < ka; ks += UK) f(h, ks, nk(ks, ka));
  for (int ks = ka; ks < kb; ks += UK)
    for (int h = 0; h < NH; ++h) f(h, ks, nk(ks, kb));
  for (int h = 0; h < NH; ++h)
    for (int ks = kb; ks < ksteps; ks += UK) f(h, ks, nk(ks, ksteps));
}
constexpr int CL = GP_STU_CL;  // 2: CTA pairs (cta_group::2).  Two tiles of one kernel offset form one M = 256 MMA over
                               // two SMs; each SM gathers its own 128 A rows and ingests only HALF of every W_o slice
constexpr bool PAIR = CL == 2;
constexpr int ALAG = 1;        // A chunks in flight per gather thread besides the one being issued
constexpr int EPI_LD = 36;     // floats per staged row in the epilogue transpose buffer (16-byte aligned, conflict-free)
#ifndef GP_STU_EPI2
#define GP_STU_EPI2 0  // 1: a second set of four epilogue warps (12-15).  Set e drains accumulator e (N = 512: column half e;
                       // N <= 256: the tiles that use accumulator e), so both halves are stored at the same time.
                       // Measured: 10.78-10.83 vs 10.79-10.83 ms, no difference: each set then drains its half in 19k
                       // clk -- the store path, not the number of warps, is the limit
#endif
constexpr bool EPI2 = GP_STU_EPI2 != 0;
static_assert(!EPI2 || GP_STU_EPI_DIRECT, "the second epilogue set has no transpose buffer: direct epilogue only");
constexpr int NTHREADS = EPI2 ? 512 : 384;  // warp 0 TMA, 1 MMA, 2 TMEM alloc, 3 idle, 4-7 epilogue (set 0), 8-11 A gather,
                                            // 12-15 epilogue set 1

struct TileInfo {  // one 128-pair tile
  int offset;      // kernel offset o (selects W_o)
  int p0;          // first pair (index into pair_in); identity mode: first input row
  int nrows;       // valid rows (<= 128)
  int c_row0;      // first row of the tile in C
};

template <int N>
struct Smem {
  static constexpr int UN = N < 256 ? N : 256;      // UMMA N = columns of one accumulator
  static constexpr int NH = N / UN;                 // column halves per tile (2 for N = 512)
  static constexpr int TCOLS = NH == 2 ? N : 2 * N; // TMEM columns: two halves, or two alternating accumulators
  static constexpr int A_BYTES = TM * AK * 2;       // per plane per chunk (16 KB)
  static constexpr int A_CHUNK = 2 * A_BYTES;       // hi + lo
  static constexpr int B_BYTES = (PAIR ? UN / 2 : UN) * ROWB;  // per plane per K-step (a CTA of a pair holds half of the rows)
  static constexpr int STAGE = UK * 2 * B_BYTES;               // [K-step][plane][rows][16]
  static constexpr int A_OFF = 0;
  static constexpr int B_OFF = NA * A_CHUNK;
  static constexpr int EPI_OFF = B_OFF + NSTAGE * STAGE;
  static constexpr int EPI = GP_STU_EPI_DIRECT ? 0 : 4 * 32 * EPI_LD * 4;  // transpose buffer (non-direct epilogue only)
  static constexpr int BAR_OFF = EPI_OFF + EPI;
  static constexpr int TOTAL = BAR_OFF + 512;       // + barriers and the TMEM slot
  static_assert((2 * NSTAGE + 2 * NA + 4) * 8 + 4 <= 512, "barrier block too small");
  static_assert(TOTAL <= 232448, "more shared memory than a CTA can have");
  static_assert(TCOLS <= 512, "TMEM has 512 columns");
};

struct GemmParams {
  const __nv_bfloat16* a;  // activations [rows][2][kp]: hi plane then lo plane
  int64_t lda;             // elements between rows (= 2 * kp)
  int kp;                  // padded K (multiple of 16)
  const int32_t* pair_in;  // gathered input row per pair, or nullptr (identity)
  const TileInfo* tiles;   // tile table, or nullptr (identity: tile t covers rows [128 t, 128 t + 128))
  const int32_t* ntiles_dev;  // number of tiles (device), or nullptr -> ntiles_host
  int ntiles_host;
  int64_t m_rows;          // identity mode: number of rows
  int cout_total;          // rows per (offset, plane, k-step) block of the pre-tiled weights (= layer cout)
  float* c;                // [c_rows][ldc]
  int64_t ldc;
  // fused arg-max epilogue (classification): when label_out != nullptr nothing is stored to c
  int32_t* label_out;      // [m_rows] lowest-index arg-max over the first ncols accumulator columns
  uint8_t* flag_out;       // [m_rows] 1 if the top-2 margin is below margin_rel * rnorm[row] (fp32 re-evaluation)
  const float* rnorm;      // [m_rows] row norms of A
  int ncols;
  float margin_rel;
};

template <int N>
__global__ void __launch_bounds__(NTHREADS, 1) gather_gemm_tc_kernel(const __grid_constant__ CUtensorMap wmap,
                                                                      const GemmParams P) {
  extern __shared__ uint8_t smem_raw[];
  using S = Smem<N>;
  uint8_t* smem = smem_raw;  // dynamic shared memory starts 1024-byte aligned (no static smem in this kernel); checked below
  float* epi = (float*)(smem + S::EPI_OFF);
  uint64_t* bars = (uint64_t*)(smem + S::BAR_OFF);
  uint64_t* b_full = bars;                   // [NSTAGE] weight slice landed (TMA)
  uint64_t* b_empty = b_full + NSTAGE;       // [NSTAGE] MMAs reading the stage have completed
  uint64_t* a_full = b_empty + NSTAGE;       // [NA] A chunk gathered (4 gather warps)
  uint64_t* a_empty = a_full + NA;           // [NA] MMAs reading the chunk have completed
  uint64_t* tmem_full = a_empty + NA;        // [2] accumulator (half) complete
  uint64_t* tmem_empty = tmem_full + 2;      // [2] accumulator (half) drained
  uint32_t* tmem_slot = (uint32_t*)(tmem_empty + 2);
  // TMEM holds two accumulators of UN <= 256 columns.  N = 512: the two column halves of one tile, each with its own
  // full/empty barrier pair, so that a half drains while the MMAs run on the other one (below).  N <= 256: two whole
  // accumulators used by alternate tiles.
  constexpr int UN = S::UN, NH = S::NH;
  auto part_bar = [&](uint32_t tcount, int h) -> int { return NH == 2 ? h : (int)(tcount & 1); };
  auto part_phase = [&](uint32_t tcount) -> uint32_t { return NH == 2 ? (tcount & 1) : ((tcount >> 1) & 1); };
  auto part_col = [&](uint32_t tcount, int h) -> uint32_t { return NH == 2 ? h * UN : (tcount & 1) * N; };

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int ntiles_real = P.ntiles_dev ? *P.ntiles_dev : P.ntiles_host;
  const int ntiles = (ntiles_real + CL - 1) / CL * CL;  // every CTA of a cluster runs the same number of tiles
  const int ksteps = P.kp / BK;
  const uint32_t crank = CL > 1 ? tc::cluster_ctarank() : 0;
  [[maybe_unused]] constexpr uint16_t CMASK = (uint16_t)((1u << CL) - 1);
  const int nchunks = (ksteps + AKS - 1) / AKS;

  if (threadIdx.x == 0) {
    if (tc::smem_u32(smem) & 1023u) __trap();
    tc::tma_prefetch_desc(&wmap);
    for (int s = 0; s < NSTAGE; ++s) {
      tc::mbar_init(&b_full[s], 1);
      tc::mbar_init(&b_empty[s], 1);
    }
    for (int s = 0; s < NA; ++s) {
      tc::mbar_init(&a_full[s], PAIR ? 8 : 4);  // pair: the leader's barrier collects both CTAs' gather warps
      tc::mbar_init(&a_empty[s], 1);
    }
    for (int s = 0; s < 2; ++s) {
      tc::mbar_init(&tmem_full[s], 1);
      tc::mbar_init(&tmem_empty[s], PAIR ? 8 : 4);
    }
    tc::fence_barrier_init();
  }
  if (warp == 2) {
    if (PAIR) tc::tmem_alloc2<S::TCOLS>(tmem_slot);
    else tc::tmem_alloc<S::TCOLS>(tmem_slot);
  }
  tc::tc_fence_before();
  __syncthreads();
  if (CL > 1) tc::cluster_sync_all();  // peers' barriers are initialised before anything is multicast to them
  tc::tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  auto tile_of = [&](int t) -> TileInfo {
    if (P.tiles) {
      TileInfo ti = P.tiles[t < ntiles_real ? t : ntiles_real - 1];
      if (t >= ntiles_real) ti.nrows = 0;
      return ti;
    }
    TileInfo ti;
    ti.offset = 0;
    ti.p0 = t * TM;
    int64_t left = P.m_rows - (int64_t)t * TM;
    ti.nrows = left < TM ? (left > 0 ? (int)left : 0) : TM;
    ti.c_row0 = t * TM;
    return ti;
  };

  if (warp == 0) {
    // ===================== TMA producer: W_o^T[k0:k0+16] hi + lo of ONE column half per stage =====================
    // stage order = the MMA issuer's: K-blocks of KBS steps, inside a block first half 0 then half 1
    if (lane == 0) {
      uint32_t it = 0;
      for (int t = blockIdx.x; t < ntiles; t += gridDim.x) {
        const TileInfo ti = tile_of(t);
        for_units<NH>(ksteps, [&](int h, int ks, int nk) {
          const int s = it % NSTAGE;
          const uint32_t ph = (it / NSTAGE) & 1;
          ++it;
          tc::mbar_wait(&b_empty[s], ph ^ 1);
          uint8_t* st = smem + S::B_OFF + s * S::STAGE;
          // leader of a pair: expect both CTAs' halves; every CTA fetches the UN/2 rows it feeds to the M = 256 instruction
          if (!PAIR || crank == 0) tc::mbar_arrive_expect_tx(&b_full[s], (PAIR ? 2 : 1) * nk * 2 * S::B_BYTES);
          for (int j = 0; j < nk; ++j) {
            const int row0 = (ti.offset * 2 * ksteps + ks + j) * P.cout_total + h * UN;  // plane 0; plane 1: ksteps blocks on
#pragma unroll
            for (int plane = 0; plane < 2; ++plane) {
              const uint32_t dst = tc::smem_u32(st + (j * 2 + plane) * S::B_BYTES);
              if constexpr (PAIR)
                tc::tma_load_2d_2sm(dst, &wmap, 0, row0 + plane * ksteps * P.cout_total + (int)crank * (UN / 2), &b_full[s]);
              else
                tc::tma_load_2d(dst, &wmap, 0, row0 + plane * ksteps * P.cout_total, &b_full[s]);
            }
          }
        });
      }
    }
  } else if (warp == 1 && (!PAIR || crank == 0)) {
    // ===================== MMA issuer (pair mode: the leader CTA issues for both SMs) =====================
    // unit order: for_units above.  Every unit is three MMAs (M = 128 or 256 over the pair, N = UN, K = 16) into the
    // unit's accumulator; a chunk buffer is released after its last K-step in the LAST half that reads it.
    constexpr uint64_t DB = tc::umma_desc_kmajor_base<ROWB>();
    constexpr uint64_t DA = tc::umma_desc_kmajor_base<AK * 2>();
    constexpr uint32_t IDESC = tc::umma_idesc_bf16(PAIR ? 2 * TM : TM, UN);
    uint32_t it = 0, ac = 0, tcount = 0;
    [[maybe_unused]] long long w_te = 0, w_af = 0, w_bf = 0;
    [[maybe_unused]] const long long prof_start = clock64();
    for (int t = blockIdx.x; t < ntiles; t += gridDim.x, ++tcount, ac += nchunks) {
      for_units<NH>(ksteps, [&](int h, int ks, int nk) {
        const int pb = part_bar(tcount, h);
        if (ks == 0) {  // first unit of this half in the tile: the previous contents must have been drained
          PROF_T0();
          tc::mbar_wait(&tmem_empty[pb], part_phase(tcount) ^ 1);
          PROF_ADD(w_te);
        }
        const uint32_t d = tmem_base + part_col(tcount, h);
        const int s = it % NSTAGE;
        const uint32_t ph = (it / NSTAGE) & 1;
        ++it;
        const uint32_t aci = ac + ks / AKS;  // running index of this tile's chunk (a unit never straddles chunks)
        const int ab = aci % NA;
        if (h == 0 && (ks & (AKS - 1)) == 0) {
          PROF_T0();
          tc::mbar_wait(&a_full[ab], (aci / NA) & 1);
          PROF_ADD(w_af);
        }
        {
          PROF_T0();
          tc::mbar_wait(&b_full[s], ph);
          PROF_ADD(w_bf);
        }
        tc::tc_fence_after();
        const int kl = ks + nk - 1;  // last K-step of the unit
        const bool chunk_end = (kl & (AKS - 1)) == AKS - 1 || kl == ksteps - 1;
        if (lane == 0) {
          for (int j = 0; j < nk; ++j) {
            // K=16 slice j of a 128-byte swizzled row starts 32 j bytes into the (1024-byte aligned) tile
            const uint32_t a_hi = tc::smem_u32(smem + S::A_OFF + ab * S::A_CHUNK) + ((ks + j) & (AKS - 1)) * ROWB;
            const uint32_t a_lo = a_hi + S::A_BYTES;
            const uint32_t b_hi = tc::smem_u32(smem + S::B_OFF + s * S::STAGE) + j * 2 * S::B_BYTES;
            const uint32_t b_lo = b_hi + S::B_BYTES;
            // small terms first, then the leading product
            if constexpr (PAIR) {
              tc::umma_bf16_2sm(d, tc::umma_desc(DA, a_lo), tc::umma_desc(DB, b_hi), IDESC, ks + j > 0);
              tc::umma_bf16_2sm(d, tc::umma_desc(DA, a_hi), tc::umma_desc(DB, b_lo), IDESC, 1);
              tc::umma_bf16_2sm(d, tc::umma_desc(DA, a_hi), tc::umma_desc(DB, b_hi), IDESC, 1);
            } else {
              tc::umma_bf16(d, tc::umma_desc(DA, a_lo), tc::umma_desc(DB, b_hi), IDESC, ks + j > 0);
              tc::umma_bf16(d, tc::umma_desc(DA, a_hi), tc::umma_desc(DB, b_lo), IDESC, 1);
              tc::umma_bf16(d, tc::umma_desc(DA, a_hi), tc::umma_desc(DB, b_hi), IDESC, 1);
            }
          }
          if constexpr (PAIR) {  // the same barrier offset in both CTAs
            tc::umma_commit_2sm(&b_empty[s], CMASK);
            if (h == NH - 1 && chunk_end) tc::umma_commit_2sm(&a_empty[ab], CMASK);
            if (kl == ksteps - 1) tc::umma_commit_2sm(&tmem_full[pb], CMASK);
          } else {
            tc::umma_commit(&b_empty[s]);
            if (h == NH - 1 && chunk_end) tc::umma_commit(&a_empty[ab]);
            if (kl == ksteps - 1) tc::umma_commit(&tmem_full[pb]);
          }
        }
        __syncwarp();
      });
    }
#if GP_STU_PROF
    if (blockIdx.x == 0 && lane == 0 && N == 512)
      printf("PROF mma: total %lld clk, tiles %u, wait tmem_empty %lld a_full %lld b_full %lld\n", clock64() - prof_start,
             tcount, w_te, w_af, w_bf);
#endif
  } else if ((warp >= 4 && warp < 8) || warp >= 12) {
    // ===================== epilogue: TMEM -> registers -> global rows, one accumulator (half) at a time ==============
    const int q = warp & 3;  // TMEM lane quadrant this warp may read
    [[maybe_unused]] float* buf = epi + q * 32 * EPI_LD;
    uint32_t tcount = 0;
    [[maybe_unused]] long long w_tf = 0;
    [[maybe_unused]] const long long prof_start = clock64();
    for (int t = blockIdx.x; t < ntiles; t += gridDim.x, ++tcount) {
      const TileInfo ti = tile_of(t);
      const int rows_here = ti.nrows - q * 32;  // valid rows in this warp's quadrant
      for (int h = 0; h < NH; ++h) {
        const int pb = part_bar(tcount, h);
        if (EPI2 && pb != (warp >= 12)) continue;  // the other set's accumulator
        const uint32_t tacc = tmem_base + part_col(tcount, h);
        {
          PROF_T0();
          tc::mbar_wait(&tmem_full[pb], part_phase(tcount));
          PROF_ADD(w_tf);
        }
        tc::tc_fence_after();
        auto release = [&]() {  // this warp has read its quadrant of the accumulator: the MMAs may overwrite it
          tc::tc_fence_before();
          __syncwarp();
          if (lane == 0) {
            if (PAIR && crank != 0) tc::mbar_arrive_cluster(&tmem_empty[pb], 0);
            else tc::mbar_arrive(&tmem_empty[pb]);
          }
        };
        if (P.label_out) {
          // classification (N <= 256, one accumulator per tile): thread = accumulator row, so the arg-max over the
          // columns is thread-local -- no transpose, no logits in HBM.  Strict > in ascending column order keeps the
          // lowest class index on ties.
          float best = -INFINITY, second = -INFINITY;
          int besti = 0;
          const int nch = (P.ncols + 31) / 32;
          for (int ch = 0; ch < UN / 32; ++ch) {
            if (ch < nch) {
              uint32_t v[32];
              tc::tmem_ld32(tacc + ((uint32_t)(q * 32) << 16) + ch * 32, v);
              tc::tmem_ld_wait();
#pragma unroll
              for (int j = 0; j < 32; ++j) {
                const float x = __uint_as_float(v[j]);
                if (ch * 32 + j < P.ncols) {
                  if (x > best) {
                    second = best;
                    best = x;
                    besti = ch * 32 + j;
                  } else if (x > second) {
                    second = x;
                  }
                }
              }
            }
          }
          release();
          if (lane < rows_here) {
            const int64_t row = (int64_t)ti.c_row0 + q * 32 + lane;
            P.label_out[row] = besti;
            // !(a > b) also catches NaN accumulators and all-zero rows (margin 0, norm 0)
            P.flag_out[row] = !((best - second) > P.margin_rel * __ldg(P.rnorm + row)) || !(best > -INFINITY);
          }
          continue;
        }
#if GP_STU_ABLATE == 2
        release();
        continue;
#endif
        float* crow = P.c + ((int64_t)ti.c_row0 + q * 32) * P.ldc + h * UN;
#if GP_STU_EPI_DIRECT
        // accumulator-fragment loads: four threads hold one 32-byte sector of a row, stored as float2 straight from
        // the registers -- no shared-memory transpose (the smem port is shared with the MMAs' operand reads).
        // Software-pipelined over two register sets (GP_STU_EPI_PIPE): the TMEM load of fragment i + 1 is in flight
        // while fragment i is being stored.
        constexpr int NI = 2 * (UN / 64);  // fragments per accumulator: 16 rows x 64 columns each
        auto ldv = [&](int i, uint32_t(&v)[32]) {
          const int hb = i / (UN / 64), cg = i % (UN / 64);
          tc::tmem_ld_16x256b_x8(tacc + ((uint32_t)(q * 32 + hb * 16) << 16) + cg * 64, v);
        };
        auto stv = [&](int i, const uint32_t(&v)[32]) {
#if GP_STU_ABLATE == 1
          if (v[0] == 0x7fc12345u && v[31] == 0x7fc54321u) crow[0] = 1.f;
          return;
#endif
          const int hb = i / (UN / 64), cg = i % (UN / 64);
          const int ra = hb * 16 + (lane >> 2), rb = ra + 8;
#if GP_STU_EPI_QUAD
          // 4 x 4 transpose inside every quad (two shfl.xor rounds): thread j ends up with the 16 CONSECUTIVE columns
          // 16 j .. 16 j + 15 of its row, so a quad stores 64 contiguous bytes per STG.128 instead of four 8-byte pieces
          // of one 32-byte sector per STG.64 (all SMs storing at once: 22 vs 19.4 B/clk/SM, scripts/probes/store_bw.cu)
          const bool o = lane & 1, hq = (lane >> 1) & 1;
#pragma unroll
          for (int rs = 0; rs < 2; ++rs) {
            // X[a][e]: block a = columns 16 a + {2 j, 2 j + 1, 8 + 2 j, 9 + 2 j}
            uint32_t T[2][2][4], Y[4][4];
#pragma unroll
            for (int r = 0; r < 2; ++r)
#pragma unroll
              for (int e = 0; e < 4; ++e) {
                const uint32_t x0 = v[4 * (2 * (2 * r) + (e >> 1)) + rs * 2 + (e & 1)];      // block 2 r
                const uint32_t x1 = v[4 * (2 * (2 * r + 1) + (e >> 1)) + rs * 2 + (e & 1)];  // block 2 r + 1
                const uint32_t send = o ? x0 : x1, keep = o ? x1 : x0;
                const uint32_t recv = __shfl_xor_sync(0xffffffffu, send, 1);
                T[0][r][e] = o ? recv : keep;  // from the quad thread with bit 0 clear
                T[1][r][e] = o ? keep : recv;  // from the one with bit 0 set
              }
#pragma unroll
            for (int sbit = 0; sbit < 2; ++sbit)
#pragma unroll
              for (int e = 0; e < 4; ++e) {
                const uint32_t send = hq ? T[sbit][0][e] : T[sbit][1][e], keep = hq ? T[sbit][1][e] : T[sbit][0][e];
                const uint32_t recv = __shfl_xor_sync(0xffffffffu, send, 2);
                Y[sbit][e] = hq ? recv : keep;      // source thread with bit 1 clear
                Y[2 + sbit][e] = hq ? keep : recv;  // source thread with bit 1 set
              }
            const int row = rs ? rb : ra;
            if (row < rows_here) {
              float4* dst = (float4*)(crow + (int64_t)row * P.ldc + cg * 64 + 16 * (lane & 3));
              dst[0] = make_float4(__uint_as_float(Y[0][0]), __uint_as_float(Y[0][1]), __uint_as_float(Y[1][0]), __uint_as_float(Y[1][1]));
              dst[1] = make_float4(__uint_as_float(Y[2][0]), __uint_as_float(Y[2][1]), __uint_as_float(Y[3][0]), __uint_as_float(Y[3][1]));
              dst[2] = make_float4(__uint_as_float(Y[0][2]), __uint_as_float(Y[0][3]), __uint_as_float(Y[1][2]), __uint_as_float(Y[1][3]));
              dst[3] = make_float4(__uint_as_float(Y[2][2]), __uint_as_float(Y[2][3]), __uint_as_float(Y[3][2]), __uint_as_float(Y[3][3]));
            }
          }
#else
          float* pa = crow + (int64_t)ra * P.ldc + 2 * (lane & 3) + cg * 64;
          float* pb2 = crow + (int64_t)rb * P.ldc + 2 * (lane & 3) + cg * 64;
#pragma unroll
          for (int g = 0; g < 8; ++g) {
            if (ra < rows_here)
              *(float2*)(pa + g * 8) = make_float2(__uint_as_float(v[4 * g]), __uint_as_float(v[4 * g + 1]));
            if (rb < rows_here)
              *(float2*)(pb2 + g * 8) = make_float2(__uint_as_float(v[4 * g + 2]), __uint_as_float(v[4 * g + 3]));
          }
#endif
        };
#if GP_STU_EPI_PIPE
        static_assert(NI % 2 == 0, "two fragments per pipeline round");
        uint32_t va[32], vb[32];
        ldv(0, va);
#pragma unroll
        for (int i = 0; i < NI; i += 2) {
          tc::tmem_ld_wait();
          ldv(i + 1, vb);
          stv(i, va);
          tc::tmem_ld_wait();
          if (i + 2 < NI) ldv(i + 2, va);
          else release();  // every TMEM read of this warp has completed
          stv(i + 1, vb);
        }
#else
#pragma unroll 1
        for (int i = 0; i < NI; ++i) {
          uint32_t v[32];
          ldv(i, v);
          tc::tmem_ld_wait();
          if (i == NI - 1) release();
          stv(i, v);
        }
#endif
#else
        for (int ch = 0; ch < UN / 32; ++ch) {
          uint32_t v[32];
          tc::tmem_ld32(tacc + ((uint32_t)(q * 32) << 16) + ch * 32, v);
          tc::tmem_ld_wait();
          if (ch == UN / 32 - 1) release();
#pragma unroll
          for (int j = 0; j < 8; ++j)
            *(float4*)(buf + lane * EPI_LD + j * 4) =
                make_float4(__uint_as_float(v[4 * j]), __uint_as_float(v[4 * j + 1]), __uint_as_float(v[4 * j + 2]),
                            __uint_as_float(v[4 * j + 3]));
          __syncwarp();
#pragma unroll
          for (int rr = 0; rr < 8; ++rr) {
            const int row = rr * 4 + (lane >> 3), c4 = lane & 7;
            if (row < rows_here) {
              float4 o = *(const float4*)(buf + row * EPI_LD + c4 * 4);
              *(float4*)(crow + (int64_t)row * P.ldc + ch * 32 + c4 * 4) = o;
            }
          }
          __syncwarp();
        }
#endif
      }
    }
#if GP_STU_PROF
    if (blockIdx.x == 0 && threadIdx.x == 128 && N == 512)
      printf("PROF epi: total %lld clk, wait tmem_full %lld (the rest is draining)\n", clock64() - prof_start, w_tf);
#endif
  } else if (warp >= 8 && warp < 12) {
    // ===================== A gather: warp w owns tile rows [32 w, 32 w + 32); 8 lanes fetch one 128-byte row slice ====
    const int w = warp - 8;
    constexpr int LPR = AK * 2 / 16;   // lanes per row slice (16-byte chunks in AK bf16)
    constexpr int RPI = 32 / LPR;      // rows per warp-wide cp.async
    const int rsub = lane / LPR, c = lane % LPR;
    uint32_t ac = 0, signalled = 0;
    [[maybe_unused]] long long w_ae = 0, w_land = 0;
    [[maybe_unused]] const long long prof_start = clock64();
    for (int t = blockIdx.x; t < ntiles; t += gridDim.x) {
      const TileInfo ti = tile_of(t);
      const int r = w * 32 + lane;
      int my_row = -1;
      if (r < ti.nrows) my_row = P.pair_in ? __ldg(P.pair_in + ti.p0 + r) : ti.p0 + r;
      for (int ci = 0; ci < nchunks; ++ci, ++ac) {
        const int ab = ac % NA;
        if (signalled < ac && !__all_sync(0xffffffffu, tc::mbar_test(&a_empty[ab], ((ac / NA) & 1) ^ 1))) {
          // about to block on a buffer the MMAs still read: hand over the chunk already issued first (the issuer may be
          // waiting for exactly that chunk before it can release this buffer)
          tc::cp_async_wait<0>();
          tc::fence_proxy_async();
          __syncwarp();
          for (; signalled < ac; ++signalled)
            if (lane == 0) {
              if (PAIR && crank != 0) tc::mbar_arrive_cluster(&a_full[signalled % NA], 0);
              else tc::mbar_arrive(&a_full[signalled % NA]);
            }
        }
        {
          PROF_T0();
          tc::mbar_wait(&a_empty[ab], ((ac / NA) & 1) ^ 1);
          PROF_ADD(w_ae);
        }
        const int kvalid = min(AK, P.kp - ci * AK);  // elements of this chunk that exist
        const uint32_t abase = tc::smem_u32(smem + S::A_OFF + ab * S::A_CHUNK);
#pragma unroll
        for (int itr = 0; itr < 32 / RPI; ++itr) {
          const int rl = itr * RPI + rsub;           // row inside this warp's 32
          const int in_row = __shfl_sync(0xffffffffu, my_row, rl);
          const int row = w * 32 + rl;
          const uint32_t nbytes = (in_row >= 0 && c * 8 < kvalid) ? 16u : 0u;
          const __nv_bfloat16* g = P.a + (int64_t)(in_row < 0 ? 0 : in_row) * P.lda + ci * AK + (nbytes ? c * 8 : 0);
          const uint32_t dst = abase + row * (AK * 2) + (tc::swizzled_chunk<AK * 2>(row, c) << 4);
          tc::cp_async16(dst, g, nbytes);
          tc::cp_async16(dst + S::A_BYTES, g + P.kp, nbytes);
        }
        tc::cp_async_commit();
        if (ac >= ALAG) {  // the chunk issued ALAG iterations ago has landed
          PROF_T0();
          tc::cp_async_wait<ALAG>();
          PROF_ADD(w_land);
          tc::fence_proxy_async();
          __syncwarp();
          if (lane == 0) {
            if (PAIR && crank != 0) tc::mbar_arrive_cluster(&a_full[signalled % NA], 0);
            else tc::mbar_arrive(&a_full[signalled % NA]);
          }
          ++signalled;
        }
      }
    }
    tc::cp_async_wait<0>();
    tc::fence_proxy_async();
    __syncwarp();
    for (; signalled < ac; ++signalled)
      if (lane == 0) {
        if (PAIR && crank != 0) tc::mbar_arrive_cluster(&a_full[signalled % NA], 0);
        else tc::mbar_arrive(&a_full[signalled % NA]);
      }
#if GP_STU_PROF
    if (blockIdx.x == 0 && threadIdx.x == 256 && N == 512)
      printf("PROF gather: total %lld clk, chunks %u, wait a_empty %lld, wait landing %lld\n", clock64() - prof_start, ac,
             w_ae, w_land);
#endif
  }

  tc::tc_fence_before();
  __syncthreads();
  if (CL > 1) tc::cluster_sync_all();  // no CTA leaves while its peer may still signal it or read its smem
  if (warp == 2) {
    if (PAIR) tc::tmem_dealloc2<S::TCOLS>(tmem_base);
    else tc::tmem_dealloc<S::TCOLS>(tmem_base);
  }
}


// ------------------------------------------------------------------ pair lists and tile tables
// flat index of (slab, offset, row-in-slab): all pairs of one (slab, offset) group are contiguous after the scan
__device__ __forceinline__ int64_t group_index(int slab, int o, int sl, int S) { return ((int64_t)slab * 27 + o) * S + sl; }

__global__ void tc_pair_flags_kernel(const int32_t* __restrict__ nbr27, int64_t m, int S, int64_t total,
                                     int32_t* __restrict__ flags) {
  int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t > total) return;
  int f = 0;
  if (t < total) {
    int sl = (int)(t % S);
    int64_t g = t / S;
    int o = (int)(g % 27);
    int64_t s = (g / 27) * S + sl;
    f = (s < m) && (nbr27[s * 27 + o] >= 0);
  }
  flags[t] = f;
}

__global__ void tc_pair_fill_kernel(const int32_t* __restrict__ nbr27, const int32_t* __restrict__ pos, int64_t m,
                                    int S, int64_t total, int32_t* __restrict__ pair_in) {
  int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= total) return;
  int sl = (int)(t % S);
  int64_t g = t / S;
  int o = (int)(g % 27);
  int64_t s = (g / 27) * S + sl;
  if (s < m) {
    int nb = nbr27[s * 27 + o];
    if (nb >= 0) pair_in[pos[t]] = nb;
  }
}

// one block per slab: tile table + first tile of every (slab, offset) group
__global__ void tc_build_tiles_kernel(const int32_t* __restrict__ pos, int S, int max_tiles,
                                      TileInfo* __restrict__ tiles, int32_t* __restrict__ ntiles,
                                      int32_t* __restrict__ gtile0) {
  __shared__ int s_t0[28], s_p0[28];
  const int slab = blockIdx.x;
  if (threadIdx.x == 0) {
    int run = 0;
    for (int o = 0; o < 27; ++o) {
      int p0 = pos[group_index(slab, o, 0, S)], p1 = pos[group_index(slab, o, 0, S) + S];
      s_t0[o] = run;
      s_p0[o] = p0;
      gtile0[slab * 27 + o] = run;
      run += ((p1 - p0 + TM - 1) / TM + CL - 1) / CL * CL;  // whole clusters of tiles per kernel offset
    }
    s_t0[27] = run;
    s_p0[27] = pos[group_index(slab, 26, 0, S) + S];
    ntiles[slab] = run;
  }
  __syncthreads();
  TileInfo* tt = tiles + (int64_t)slab * max_tiles;
  for (int o = 0; o < 27; ++o) {
    int nt = s_t0[o + 1] - s_t0[o];
    int cnt = s_p0[o + 1] - s_p0[o];
    for (int i = threadIdx.x; i < nt; i += blockDim.x) {
      TileInfo ti;
      ti.offset = o;
      ti.p0 = s_p0[o] + i * TM;
      ti.nrows = max(0, min(TM, cnt - i * TM));  // 0: padding tile of the cluster group
      ti.c_row0 = (s_t0[o] + i) * TM;
      tt[s_t0[o] + i] = ti;
    }
  }
}

// ------------------------------------------------------------------ fp32 -> bf16 hi/lo planes
__device__ __forceinline__ void split_bf16(float v, __nv_bfloat16& hi, __nv_bfloat16& lo) {
  hi = __float2bfloat16_rn(v);
  lo = __float2bfloat16_rn(v - __bfloat162float(hi));
}

__global__ void tc_split_input_kernel(const float* __restrict__ x, int64_t stride, int cin, int kp, int64_t m,
                                      __nv_bfloat16* __restrict__ out) {
  // four columns per thread: one 16-byte load, two 8-byte stores (kp and the row stride are multiples of 4)
  const int kq = kp >> 2;
  int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= m * kq) return;
  int64_t r = t / kq;
  int c = (int)(t - r * kq) * 4;
  float v[4];
  if (c + 3 < cin && (stride & 3) == 0 && ((uintptr_t)x & 15) == 0) {
    const float4 q = *(const float4*)(x + r * stride + c);
    v[0] = q.x; v[1] = q.y; v[2] = q.z; v[3] = q.w;
  } else {
#pragma unroll
    for (int j = 0; j < 4; ++j) v[j] = c + j < cin ? x[r * stride + c + j] : 0.f;
  }
  __nv_bfloat16 hi[4], lo[4];
#pragma unroll
  for (int j = 0; j < 4; ++j) split_bf16(v[j], hi[j], lo[j]);
  *(uint2*)(out + r * 2 * kp + c) = *(const uint2*)hi;
  *(uint2*)(out + r * 2 * kp + kp + c) = *(const uint2*)lo;
}

// weights [kv][cin_pad][cout] fp32 -> Wt [kv][2][kp/16][cout][16] bf16: hi plane then lo plane, and inside a plane one
// contiguous [cout][16] block per K=16 step, so that the TMA box of a stage is one contiguous run of 32-byte rows
__global__ void tc_pack_weights_kernel(const float* __restrict__ w, int kv, int cin_pad, int cout, int kp,
                                       __nv_bfloat16* __restrict__ out) {
  int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  int64_t total = (int64_t)kv * cout * kp;
  if (t >= total) return;
  int k = (int)(t % kp);
  int n = (int)((t / kp) % cout);
  int o = (int)(t / ((int64_t)kp * cout));
  float v = k < cin_pad ? w[((int64_t)o * cin_pad + k) * cout + n] : 0.f;
  __nv_bfloat16 hi, lo;
  split_bf16(v, hi, lo);
  const int ks = k / BK, kk = k % BK, nks = kp / BK;
  out[((((int64_t)o * 2 + 0) * nks + ks) * cout + n) * BK + kk] = hi;
  out[((((int64_t)o * 2 + 1) * nks + ks) * cout + n) * BK + kk] = lo;
}

// ------------------------------------------------------------------ reduce: pair rows -> output rows + epilogue
// one warp per output voxel of the slab; COUT = 128 * NV4
template <int NV4>
__global__ void __launch_bounds__(256) student_reduce_kernel(const float* __restrict__ staging,
                                                              const int32_t* __restrict__ nbr27,
                                                              const int32_t* __restrict__ pos,
                                                              const int32_t* __restrict__ gtile0, int slab, int S,
                                                              int64_t m, const float* __restrict__ scale,
                                                              const float* __restrict__ shift,
                                                              const float* __restrict__ residual, int relu,
                                                              __nv_bfloat16* __restrict__ out_hilo,
                                                              float* __restrict__ out_f32) {
  constexpr int COUT = 128 * NV4;
  const int lane = threadIdx.x & 31;
  const int sl = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  const int64_t s = (int64_t)slab * S + sl;
  if (sl >= S || s >= m) return;
  int row = -1;
  if (lane < 27 && __ldg(nbr27 + s * 27 + lane) >= 0) {
    int64_t gi = group_index(slab, lane, 0, S);
    int rel = __ldg(pos + gi + sl) - __ldg(pos + gi);
    row = (__ldg(gtile0 + slab * 27 + lane) + rel / TM) * TM + rel % TM;
  }
  unsigned live = __ballot_sync(0xffffffffu, row >= 0);
  float4 acc[NV4];
#pragma unroll
  for (int j = 0; j < NV4; ++j) acc[j] = make_float4(0.f, 0.f, 0.f, 0.f);
  while (live) {  // ascending offset order, as the oracle's per-offset index_add_
    int o = __ffs(live) - 1;
    live &= live - 1;
    int rr = __shfl_sync(0xffffffffu, row, o);
    const float4* src = (const float4*)(staging + (int64_t)rr * COUT);
    float4 t[NV4];
#pragma unroll
    for (int j = 0; j < NV4; ++j) t[j] = ldg_stream_f4(src + j * 32 + lane);
#pragma unroll
    for (int j = 0; j < NV4; ++j) {
      acc[j].x += t[j].x;
      acc[j].y += t[j].y;
      acc[j].z += t[j].z;
      acc[j].w += t[j].w;
    }
  }
#pragma unroll
  for (int j = 0; j < NV4; ++j) {
    const int c = (j * 32 + lane) * 4;
    float4 v = acc[j];
    if (scale) {
      float4 sc = __ldg((const float4*)(scale + c)), sh = __ldg((const float4*)(shift + c));
      v.x = fmaf(v.x, sc.x, sh.x);
      v.y = fmaf(v.y, sc.y, sh.y);
      v.z = fmaf(v.z, sc.z, sh.z);
      v.w = fmaf(v.w, sc.w, sh.w);
    }
    if (residual) {
      float4 r = __ldg((const float4*)(residual + s * COUT + c));
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
    if (out_f32) *(float4*)(out_f32 + s * COUT + c) = v;
    if (out_hilo) {
      __nv_bfloat16 h[4], l[4];
      split_bf16(v.x, h[0], l[0]);
      split_bf16(v.y, h[1], l[1]);
      split_bf16(v.z, h[2], l[2]);
      split_bf16(v.w, h[3], l[3]);
      *(uint2*)(out_hilo + s * 2 * COUT + c) = *(const uint2*)h;
      *(uint2*)(out_hilo + s * 2 * COUT + COUT + c) = *(const uint2*)l;
    }
  }
}

}  // namespace

// ------------------------------------------------------------------ host side
#ifndef GP_TC_SLAB
#define GP_TC_SLAB 131072
#endif

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn encode_tiled_fn() {
  static EncodeTiledFn fn = nullptr;
  if (!fn) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = (EncodeTiledFn)p;
  }
  return fn;
}

// 2-D bf16 tensor [rows][16] (the pre-tiled weight image: row = ((offset*2 + plane) * ksteps + kstep) * cout + n),
// box 16 x box_rows, 32-byte swizzle
static bool make_weight_map(CUtensorMap* map, const void* base, int64_t rows, int box_rows) {
  EncodeTiledFn fn = encode_tiled_fn();
  if (!fn) return false;
  cuuint64_t dims[2] = {(cuuint64_t)BK, (cuuint64_t)rows};
  cuuint64_t strides[1] = {(cuuint64_t)BK * 2};
  cuuint32_t box[2] = {BK, (cuuint32_t)box_rows};
  cuuint32_t estr[2] = {1, 1};
  return fn(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(base), dims, strides, box, estr,
            CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_32B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
            CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

struct TcLayout {
  int S, nslab, max_tiles, hidden, kp0;
  int64_t total;  // flag entries
  size_t off_flags, off_pos, off_pair_in, off_tiles, off_ntiles, off_gtile0, off_xin, off_act[3], off_res[2],
      off_staging, off_cub, cub_bytes, bytes;
};

static bool tc_layout(const gp_student_desc* d, int64_t m, TcLayout& L) {
  if (!d || d->n_layers < 2 || d->n_layers > 16 || m < 0 || m >= (1ll << 31) / 27) return false;
  int hidden = d->cout[0];
  for (int l = 0; l + 1 < d->n_layers; ++l)
    if (d->cout[l] != hidden || d->kv[l] != 27) return false;
  if (d->kv[d->n_layers - 1] != 1) return false;
  L.hidden = hidden;
  L.kp0 = d->cin[0];
  int64_t m128 = (m + TM - 1) / TM * TM;
  L.S = (int)(m128 < GP_TC_SLAB ? (m128 > 0 ? m128 : TM) : GP_TC_SLAB);
  L.nslab = (int)((m + L.S - 1) / L.S);
  if (L.nslab < 1) L.nslab = 1;
  L.max_tiles = 27 * (L.S / TM) + 27 * CL;
  L.total = (int64_t)L.nslab * 27 * L.S;
  size_t off = 0;
  auto take = [&](size_t bytes) {
    size_t o = off;
    off += gp_align_up(bytes, 1024);
    return o;
  };
  L.off_flags = take(sizeof(int32_t) * (L.total + 1));
  L.off_pos = take(sizeof(int32_t) * (L.total + 1));
  L.off_pair_in = take(sizeof(int32_t) * L.total);
  L.off_tiles = take(sizeof(TileInfo) * (size_t)L.nslab * L.max_tiles);
  L.off_ntiles = take(sizeof(int32_t) * L.nslab);
  L.off_gtile0 = take(sizeof(int32_t) * L.nslab * 27);
  L.off_xin = take(sizeof(__nv_bfloat16) * 2 * (size_t)m * L.kp0);
  for (int i = 0; i < 3; ++i) L.off_act[i] = take(sizeof(__nv_bfloat16) * 2 * (size_t)m * hidden);
  for (int i = 0; i < 2; ++i) L.off_res[i] = take(sizeof(float) * (size_t)m * hidden);
  L.off_staging = take(sizeof(float) * (size_t)L.max_tiles * TM * hidden);
  L.cub_bytes = 0;
  cub::DeviceScan::ExclusiveSum(nullptr, L.cub_bytes, (const int32_t*)nullptr, (int32_t*)nullptr, (int)(L.total + 1));
  L.off_cub = take(L.cub_bytes + 256);
  L.bytes = off;
  return true;
}

extern "C" size_t gp_student_tc_workspace_bytes(const gp_student_desc* desc_host, int64_t m) {
  TcLayout L;
  if (!tc_layout(desc_host, m, L)) return 0;
  return L.bytes;
}

// bf16 hi/lo K-major repack of every layer's weights: layer l at element offset sum_{j<l} kv*2*cout*cin
extern "C" size_t gp_student_tc_pack_bytes(const gp_student_desc* d) {
  if (!d || d->n_layers < 1 || d->n_layers > 16) return 0;
  size_t n = 0;
  for (int l = 0; l < d->n_layers; ++l) n += (size_t)d->kv[l] * 2 * d->cout[l] * d->cin[l];
  return n * sizeof(__nv_bfloat16);
}

extern "C" int gp_student_tc_pack(const gp_student_desc* d, const float* weights, void* packed, gp_stream_t stream) {
  GP_CHECK_ARG(d && d->n_layers >= 1 && d->n_layers <= 16, "gp_student_tc_pack: bad descriptor");
  GP_CHECK_ARG(((uintptr_t)packed % 128) == 0, "gp_student_tc_pack: packed buffer must be 128-byte aligned");
  size_t off = 0;
  for (int l = 0; l < d->n_layers; ++l) {
    GP_CHECK_ARG(d->cin[l] % BK == 0, "gp_student_tc_pack: layer %d cin must be a multiple of 16", l);
    int64_t total = (int64_t)d->kv[l] * d->cout[l] * d->cin[l];
    tc_pack_weights_kernel<<<(unsigned)((total + 255) / 256), 256, 0, (cudaStream_t)stream>>>(
        weights + d->w_off[l], d->kv[l], d->cin[l], d->cout[l], d->cin[l], (__nv_bfloat16*)packed + off);
    GP_LAUNCH_CHECK();
    off += (size_t)total * 2;
  }
  return GP_OK;
}

template <int N>
static int launch_gemm(const CUtensorMap& map, const GemmParams& P, int grid, cudaStream_t st) {
  static bool configured = false;
  if (!configured) {
    GP_CUDA(cudaFuncSetAttribute(gather_gemm_tc_kernel<N>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                 Smem<N>::TOTAL));
    configured = true;
  }
  grid = grid / CL * CL;
  if (grid < CL) grid = CL;
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3((unsigned)grid, 1, 1);
  cfg.blockDim = dim3(NTHREADS, 1, 1);
  cfg.dynamicSmemBytes = Smem<N>::TOTAL;
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = CL;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  GP_CUDA(cudaLaunchKernelEx(&cfg, gather_gemm_tc_kernel<N>, map, P));
  return GP_OK;
}

extern "C" int gp_student_forward_tc(const gp_student_desc* d, const float* weights, const void* packed,
                                     const float* x, int64_t cin_stride, const int32_t* nbr27, int64_t m,
                                     float* embed, int normalize, void* workspace, size_t workspace_bytes,
                                     gp_stream_t stream) {
  TcLayout L;
  GP_CHECK_ARG(tc_layout(d, m, L),
               "gp_student_forward_tc: needs 3^3 layers of one hidden width followed by a 1^3 output layer");
  const int H = L.hidden, E = d->cout[d->n_layers - 1];
  GP_CHECK_ARG(H == 512 || H == 256 || H == 128, "gp_student_forward_tc: hidden width %d not in {128,256,512}", H);
  GP_CHECK_ARG(E == 128, "gp_student_forward_tc: embed width %d must be 128", E);
  for (int l = 0; l < d->n_layers; ++l) {
    GP_CHECK_ARG(d->cin[l] % BK == 0, "gp_student_forward_tc: layer %d cin %% 16", l);
    GP_CHECK_ARG(d->residual_from[l] == -1 || d->residual_from[l] == l - 2,
                 "gp_student_forward_tc: layer %d residual must come from layer l-2", l);
    if (l > 0) GP_CHECK_ARG(d->cin[l] == d->cout[l - 1], "gp_student_forward_tc: layer %d cin != previous cout", l);
  }
  GP_CHECK_ARG(cin_stride >= d->cin[0], "gp_student_forward_tc: input stride %lld < padded cin %d",
               (long long)cin_stride, d->cin[0]);
  GP_CHECK_ARG(((uintptr_t)workspace % 256) == 0 || m == 0, "gp_student_forward_tc: workspace must be 256-byte aligned");
  if (workspace_bytes < L.bytes) {
    gp_set_error("gp_student_forward_tc: workspace %zu < %zu", workspace_bytes, L.bytes);
    return GP_ERR_WORKSPACE;
  }
  GP_CHECK_ARG(encode_tiled_fn() != nullptr, "gp_student_forward_tc: cuTensorMapEncodeTiled unavailable");
  if (m == 0) return GP_OK;
  cudaStream_t st = (cudaStream_t)stream;
  static int sm_count = 0;
  if (!sm_count) {
    int dev = 0;
    GP_CUDA(cudaGetDevice(&dev));
    GP_CUDA(cudaDeviceGetAttribute(&sm_count, cudaDevAttrMultiProcessorCount, dev));
  }
  char* ws = (char*)workspace;
  int32_t* flags = (int32_t*)(ws + L.off_flags);
  int32_t* pos = (int32_t*)(ws + L.off_pos);
  int32_t* pair_in = (int32_t*)(ws + L.off_pair_in);
  TileInfo* tiles = (TileInfo*)(ws + L.off_tiles);
  int32_t* ntiles = (int32_t*)(ws + L.off_ntiles);
  int32_t* gtile0 = (int32_t*)(ws + L.off_gtile0);
  __nv_bfloat16* xin = (__nv_bfloat16*)(ws + L.off_xin);
  float* staging = (float*)(ws + L.off_staging);

  // pair lists + tile tables (shared by every 3^3 layer)
  tc_pair_flags_kernel<<<(unsigned)((L.total + 1 + 255) / 256), 256, 0, st>>>(nbr27, m, L.S, L.total, flags);
  GP_LAUNCH_CHECK();
  size_t cb = L.cub_bytes + 256;
  GP_CUDA(cub::DeviceScan::ExclusiveSum(ws + L.off_cub, cb, flags, pos, (int)(L.total + 1), st));
  tc_pair_fill_kernel<<<(unsigned)((L.total + 255) / 256), 256, 0, st>>>(nbr27, pos, m, L.S, L.total, pair_in);
  GP_LAUNCH_CHECK();
  tc_build_tiles_kernel<<<L.nslab, 256, 0, st>>>(pos, L.S, L.max_tiles, tiles, ntiles, gtile0);
  GP_LAUNCH_CHECK();
  {
    int64_t tot = m * (L.kp0 / 4);
    tc_split_input_kernel<<<(unsigned)((tot + 255) / 256), 256, 0, st>>>(x, cin_stride, d->cin[0], L.kp0, m, xin);
    GP_LAUNCH_CHECK();
  }

  const __nv_bfloat16* in = xin;
  int in_kp = L.kp0;
  size_t woff = 0;  // element offset into `packed`
  int act_i = 0, res_i = 0;
  const float* res_src[16];
  for (int l = 0; l < 16; ++l) res_src[l] = nullptr;
  for (int l = 0; l + 1 < d->n_layers; ++l) {
    CUtensorMap map;
    const __nv_bfloat16* wl = (const __nv_bfloat16*)packed + woff;
    GP_CHECK_ARG(make_weight_map(&map, wl, (int64_t)27 * 2 * H * (in_kp / BK), (H < 256 ? H : 256) / (PAIR ? 2 : 1)),
                 "gp_student_forward_tc: cuTensorMapEncodeTiled failed (layer %d)", l);
    bool is_res_src = false;
    for (int j = l + 1; j < d->n_layers; ++j) is_res_src |= d->residual_from[j] == l;
    __nv_bfloat16* out_hilo = (__nv_bfloat16*)(ws + L.off_act[act_i]);
    float* out_f32 = nullptr;
    if (is_res_src) {
      out_f32 = (float*)(ws + L.off_res[res_i]);
      res_src[l] = out_f32;
      res_i ^= 1;
    }
    const float* residual = d->residual_from[l] >= 0 ? res_src[d->residual_from[l]] : nullptr;
    const float* sc = d->bn[l] ? weights + d->bn_off[l] : nullptr;
    for (int slab = 0; slab < L.nslab; ++slab) {
      GemmParams P = {};
      P.a = in;
      P.lda = 2 * (int64_t)in_kp;
      P.kp = in_kp;
      P.pair_in = pair_in;
      P.tiles = tiles + (int64_t)slab * L.max_tiles;
      P.ntiles_dev = ntiles + slab;
      P.ntiles_host = 0;
      P.m_rows = m;
      P.cout_total = H;
      P.c = staging;
      P.ldc = H;
      const bool tg = gp_timing_begin(GP_T_STUDENT_GEMM, st);
      int rc = H == 512 ? launch_gemm<512>(map, P, sm_count, st)
                        : (H == 256 ? launch_gemm<256>(map, P, sm_count, st) : launch_gemm<128>(map, P, sm_count, st));
      if (tg) gp_timing_end(GP_T_STUDENT_GEMM, st);
      if (rc) return rc;
      const bool tr = gp_timing_begin(GP_T_STUDENT_REDUCE, st);
      int64_t rows = m - (int64_t)slab * L.S < L.S ? m - (int64_t)slab * L.S : L.S;
      unsigned grid = (unsigned)((rows + 7) / 8);
#define REDUCE(NV4)                                                                                              \
  student_reduce_kernel<NV4><<<grid, 256, 0, st>>>(staging, nbr27, pos, gtile0, slab, L.S, m, sc, sc ? sc + H : nullptr, \
                                                   residual, d->relu[l], out_hilo, out_f32)
      if (H == 512) REDUCE(4);
      else if (H == 256) REDUCE(2);
      else REDUCE(1);
#undef REDUCE
      if (tr) gp_timing_end(GP_T_STUDENT_REDUCE, st);
      GP_LAUNCH_CHECK();
    }
    woff += (size_t)27 * 2 * H * in_kp;
    in = out_hilo;
    in_kp = H;
    act_i = (act_i + 1) % 3;
  }
  {  // 1^3 output layer: plain GEMM [m, H] x [H, E] straight into `embed`
    const int l = d->n_layers - 1;
    CUtensorMap map;
    GP_CHECK_ARG(make_weight_map(&map, (const __nv_bfloat16*)packed + woff, (int64_t)2 * E * (H / BK), E / (PAIR ? 2 : 1)),
                 "gp_student_forward_tc: cuTensorMapEncodeTiled failed (output layer)");
    GP_CHECK_ARG(!d->bn[l] && !d->relu[l] && d->residual_from[l] < 0, "gp_student_forward_tc: output layer is linear");
    GemmParams P = {};
    P.a = in;
    P.lda = 2 * (int64_t)H;
    P.kp = H;
    P.pair_in = nullptr;
    P.tiles = nullptr;
    P.ntiles_dev = nullptr;
    P.ntiles_host = (int)((m + TM - 1) / TM);
    P.m_rows = m;
    P.cout_total = E;
    P.c = embed;
    P.ldc = E;
    int grid = P.ntiles_host < sm_count ? P.ntiles_host : sm_count;
    int rc = launch_gemm<128>(map, P, grid, st);
    if (rc) return rc;
  }
  if (normalize) {
    int rc = gp_l2_normalize_rows(embed, m, E, st);
    if (rc) return rc;
  }
  return GP_OK;
}

// ------------------------------------------------------------------ A12 on the tensor cores
// Text-embedding classification as a tcgen05/TMEM GEMM with a fused arg-max epilogue (run/validation.py:413-417).
// arg-max_c s * <f/|f|, t_c/|t_c|> = arg-max_c <f, t_c/|t_c|> for s > 0, so the rows are not normalised: the kernel
// computes F[rows, D] x T^[D, C] with the fp32-faithful bf16 hi/lo split (3 MMAs per product, error <= 2^-15 |f|),
// each epilogue thread owns one accumulator row and keeps (best, second, arg best) over the C columns -- no logits
// ever reach HBM.  Rows whose top-2 margin is below 2e-4 |f| (near ties, all-zero rows, NaNs) are flagged and
// re-evaluated by the fp32 SIMT kernel, so the labels are identical to gp_logits_argmax's.
namespace {

__global__ void __launch_bounds__(256) logits_split_kernel(const float* __restrict__ f, int64_t stride, int D,
                                                           int64_t rows, __nv_bfloat16* __restrict__ planes,
                                                           float* __restrict__ rnorm) {
  const int lane = threadIdx.x & 31;
  const int64_t r = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (r >= rows) return;
  float ss = 0.f;
  for (int c = lane * 4; c < D; c += 128) {
    const float4 v = ldg_stream_f4((const float4*)(f + r * stride + c));
    ss = fmaf(v.x, v.x, ss);
    ss = fmaf(v.y, v.y, ss);
    ss = fmaf(v.z, v.z, ss);
    ss = fmaf(v.w, v.w, ss);
    __nv_bfloat16 h[4], l[4];
    split_bf16(v.x, h[0], l[0]);
    split_bf16(v.y, h[1], l[1]);
    split_bf16(v.z, h[2], l[2]);
    split_bf16(v.w, h[3], l[3]);
    *(uint2*)(planes + r * 2 * D + c) = *(const uint2*)h;
    *(uint2*)(planes + r * 2 * D + D + c) = *(const uint2*)l;
  }
  ss = warp_sum(ss);
  if (lane == 0) rnorm[r] = sqrtf(ss);
}

// W[k][c] = text[c][k] / max(|text[c]|, 1e-12) for c < C, 0 for the padding columns; one warp per class
__global__ void __launch_bounds__(256) text_weights_kernel(const float* __restrict__ text, int C, int D, int cpad,
                                                           float* __restrict__ w) {
  const int lane = threadIdx.x & 31;
  const int c = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (c >= cpad) return;
  float ss = 0.f;
  if (c < C)
    for (int k = lane; k < D; k += 32) ss = fmaf(text[(int64_t)c * D + k], text[(int64_t)c * D + k], ss);
  ss = warp_sum(ss);
  const float den = fmaxf(sqrtf(ss), 1e-12f);
  for (int k = lane; k < D; k += 32) w[(int64_t)k * cpad + c] = c < C ? text[(int64_t)c * D + k] / den : 0.f;
}

struct LogitsTcLayout {
  int cpad;
  size_t off_simt, simt_bytes, off_w, off_packed, off_planes, off_rnorm, off_flags, bytes;
};

bool logits_tc_layout(int64_t rows, int C, int D, LogitsTcLayout& L) {
  if (rows < 0 || C < 1 || C > 256 || D < 128 || D % 128 != 0 || D > 1024) return false;
  L.cpad = C <= 128 ? 128 : 256;
  size_t off = 0;
  auto take = [&](size_t b) {
    size_t o = off;
    off += gp_align_up(b, 1024);
    return o;
  };
  L.simt_bytes = gp_logits_workspace_bytes(C, D);
  L.off_simt = take(L.simt_bytes);
  L.off_w = take(sizeof(float) * (size_t)D * L.cpad);
  L.off_packed = take(sizeof(__nv_bfloat16) * 2 * (size_t)L.cpad * D);
  L.off_planes = take(sizeof(__nv_bfloat16) * 2 * (size_t)rows * D);
  L.off_rnorm = take(sizeof(float) * (size_t)rows);
  L.off_flags = take((size_t)rows);
  L.bytes = off;
  return true;
}

}  // namespace

extern "C" size_t gp_logits_tc_workspace_bytes(int64_t rows, int C, int D) {
  LogitsTcLayout L;
  return logits_tc_layout(rows, C, D, L) ? L.bytes : 0;
}

extern "C" int gp_logits_argmax_tc(const float* feat, int64_t stride, int D, int64_t rows, const float* text, int C,
                                   float logit_scale, int32_t* label, uint8_t* zero, void* workspace,
                                   size_t workspace_bytes, gp_stream_t stream) {
  LogitsTcLayout L;
  GP_CHECK_ARG(logits_tc_layout(rows, C, D, L), "gp_logits_argmax_tc: needs 1 <= C <= 256, D a multiple of 128 <= 1024");
  GP_CHECK_ARG(logit_scale > 0.f, "gp_logits_argmax_tc: logit_scale must be > 0 (the arg-max drops it)");
  GP_CHECK_ARG(stride % 4 == 0 && stride >= D && ((uintptr_t)feat % 16) == 0, "gp_logits_argmax_tc: feat alignment");
  GP_CHECK_ARG(((uintptr_t)workspace % 256) == 0 || rows == 0, "gp_logits_argmax_tc: workspace must be 256-byte aligned");
  if (workspace_bytes < L.bytes) {
    gp_set_error("gp_logits_argmax_tc: workspace %zu < %zu", workspace_bytes, L.bytes);
    return GP_ERR_WORKSPACE;
  }
  GP_CHECK_ARG(encode_tiled_fn() != nullptr, "gp_logits_argmax_tc: cuTensorMapEncodeTiled unavailable");
  if (rows == 0) return GP_OK;
  cudaStream_t st = (cudaStream_t)stream;
  static int sm_count = 0;
  if (!sm_count) {
    int dev = 0;
    GP_CUDA(cudaGetDevice(&dev));
    GP_CUDA(cudaDeviceGetAttribute(&sm_count, cudaDevAttrMultiProcessorCount, dev));
  }
  char* ws = (char*)workspace;
  float* w = (float*)(ws + L.off_w);
  __nv_bfloat16* packed = (__nv_bfloat16*)(ws + L.off_packed);
  __nv_bfloat16* planes = (__nv_bfloat16*)(ws + L.off_planes);
  float* rnorm = (float*)(ws + L.off_rnorm);
  uint8_t* flags = (uint8_t*)(ws + L.off_flags);
  text_weights_kernel<<<(L.cpad + 7) / 8, 256, 0, st>>>(text, C, D, L.cpad, w);
  GP_LAUNCH_CHECK();
  {
    int64_t total = (int64_t)L.cpad * D;
    tc_pack_weights_kernel<<<(unsigned)((total + 255) / 256), 256, 0, st>>>(w, 1, D, L.cpad, D, packed);
    GP_LAUNCH_CHECK();
  }
  logits_split_kernel<<<(unsigned)((rows + 7) / 8), 256, 0, st>>>(feat, stride, D, rows, planes, rnorm);
  GP_LAUNCH_CHECK();
  CUtensorMap map;
  GP_CHECK_ARG(make_weight_map(&map, packed, (int64_t)2 * L.cpad * (D / BK), (L.cpad < 256 ? L.cpad : 256) / (PAIR ? 2 : 1)),
               "gp_logits_argmax_tc: cuTensorMapEncodeTiled failed");
  GemmParams P = {};
  P.a = planes;
  P.lda = 2 * (int64_t)D;
  P.kp = D;
  P.ntiles_host = (int)((rows + TM - 1) / TM);
  P.m_rows = rows;
  P.cout_total = L.cpad;
  P.label_out = label;
  P.flag_out = flags;
  P.rnorm = rnorm;
  P.ncols = C;
  P.margin_rel = 2e-4f;
  const int grid = P.ntiles_host < sm_count ? P.ntiles_host : sm_count;
  int rc = L.cpad == 128 ? launch_gemm<128>(map, P, grid, st) : launch_gemm<256>(map, P, grid, st);
  if (rc) return rc;
  if (zero) GP_CUDA(cudaMemsetAsync(zero, 0, (size_t)rows, st));
  // near ties, all-zero rows and NaNs: the fp32 kernel decides (and sets their zero flags)
  return gp_logits_fix_rows(feat, stride, D, rows, text, C, logit_scale, label, zero, flags, ws + L.off_simt,
                            L.simt_bytes, st);
}

// ------------------------------------------------------------------ F4: hybrid contrastive mining on the tensor cores
// Replaces SonataXAffinityTrainer.sample_contrastive_pairs_hybrid (models/affinity_module.py:1099-1136) for a given
// anchor set: cosine similarities of A anchors against all n points, positive = arg-max (anchor excluded), 48 "macro"
// negatives = the smallest similarities (anchor and positive excluded), 15 "micro" negatives = the least similar of
// the anchor's K spatial neighbours (positive excluded).  Ties: lowest index first.
// S = Fn[n, D] x Fn[anchors]^T runs through gather_gemm_tc_kernel<512> unchanged: rows = points (identity gather),
// the 512-anchor chunks play the role of the kernel offsets (tile.offset selects the chunk's weight image), the fp32
// result is stored chunk-major [chunk][n][512]; the selection passes then read it with lane = anchor (coalesced).
namespace {

constexpr int MINE_SEG = 16;     // row segments per anchor in the partial selection
constexpr int MINE_LMAX = 64;    // longest macro list (n_macro + 1 <= MINE_LMAX)

// rows -> unit rows (F.normalize, eps 1e-12) -> bf16 hi/lo planes [n][2][kp]; one warp per row
__global__ void __launch_bounds__(256) mine_normalize_split_kernel(const float* __restrict__ f, int64_t stride, int D,
                                                                    int kp, int64_t n, __nv_bfloat16* __restrict__ planes) {
  const int lane = threadIdx.x & 31;
  const int64_t r = (int64_t)blockIdx.x * 8 + (threadIdx.x >> 5);
  if (r >= n) return;
  float ss = 0.f;
  for (int c = lane; c < D; c += 32) {
    const float v = f[r * stride + c];
    ss = fmaf(v, v, ss);
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) ss += __shfl_xor_sync(0xffffffffu, ss, o);
  const float inv = 1.f / fmaxf(sqrtf(ss), 1e-12f);
  for (int c = lane; c < kp; c += 32) {
    const float v = c < D ? f[r * stride + c] * inv : 0.f;
    __nv_bfloat16 hi, lo;
    split_bf16(v, hi, lo);
    planes[r * 2 * kp + c] = hi;
    planes[r * 2 * kp + kp + c] = lo;
  }
}

// anchor rows of the planes -> pre-tiled weight image [chunk][plane][kstep][512][16]
__global__ void mine_anchor_weights_kernel(const __nv_bfloat16* __restrict__ planes, const int32_t* __restrict__ anchors,
                                           int A, int chunks, int kp, __nv_bfloat16* __restrict__ out) {
  const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int64_t total = (int64_t)chunks * 2 * kp * 512;
  if (t >= total) return;
  const int kk = (int)(t % BK);
  const int a = (int)((t / BK) % 512);
  const int ks = (int)((t / (BK * 512)) % (kp / BK));
  const int plane = (int)((t / ((int64_t)BK * 512 * (kp / BK))) % 2);
  const int chunk = (int)(t / ((int64_t)BK * 512 * (kp / BK) * 2));
  const int ai = chunk * 512 + a;
  __nv_bfloat16 v = __float2bfloat16_rn(0.f);
  if (ai < A) v = planes[(int64_t)anchors[ai] * 2 * kp + plane * kp + ks * BK + kk];
  out[t] = v;
}

__global__ void mine_tiles_kernel(int64_t n, int chunks, int64_t n_pad, TileInfo* __restrict__ tiles) {
  const int rt = (int)((n + TM - 1) / TM);
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= rt * chunks) return;
  const int chunk = t / rt, r = t - chunk * rt;
  TileInfo ti;
  ti.offset = chunk;
  ti.p0 = r * TM;
  ti.nrows = (int)min((int64_t)TM, n - (int64_t)r * TM);
  ti.c_row0 = (int)((int64_t)chunk * n_pad + (int64_t)r * TM);
  tiles[t] = ti;
}

// (value, index) ascending with lowest index first on ties
__device__ __forceinline__ bool mine_less(float v, int i, float w, int j) { return v < w || (v == w && i < j); }

// Partial selection: a CTA owns 64 anchors x one row segment.  Row blocks of 32 x 64 similarities are staged in shared
// memory (coalesced 256-byte rows), then every warp serves 8 of the anchors with lane = row: one compare against the
// anchor's current threshold per element, and the rare survivors are inserted WARP-COOPERATIVELY into the anchor's list
// of the 64 smallest keys (2 per lane, registers; key = orderable float bits << 32 | row) by replacing its maximum.
__device__ __forceinline__ unsigned mine_f2ord(float f) {
  const unsigned u = __float_as_uint(f);
  return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
__device__ __forceinline__ float mine_ord2f(unsigned u) {
  return __uint_as_float((u & 0x80000000u) ? (u & 0x7fffffffu) : ~u);
}

__global__ void __launch_bounds__(256) mine_partial_kernel(const float* __restrict__ S, int64_t n, int64_t n_pad, int A,
                                                            const int32_t* __restrict__ anchors, int L,
                                                            float* __restrict__ pbest_v, int32_t* __restrict__ pbest_i,
                                                            float* __restrict__ plist_v, int32_t* __restrict__ plist_i) {
  __shared__ float tile[32][65];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int a0 = blockIdx.x * 64;  // 64 consecutive anchors: one 512-anchor chunk, 256 contiguous bytes per row
  const int seg = blockIdx.y;
  const int chunk = a0 >> 9, col0 = a0 & 511;
  const float* base = S + ((int64_t)chunk * n_pad) * 512 + col0;
  const int64_t rows_per = (n + MINE_SEG - 1) / MINE_SEG;
  const int64_t r0 = seg * rows_per, r1 = min(n, r0 + rows_per);
  u64 key[8][2], thr[8];
  int self[8];
  float best[8];
  int besti[8];
#pragma unroll
  for (int k = 0; k < 8; ++k) {
    const int a = a0 + warp * 8 + k;
    self[k] = a < A ? __ldg(anchors + a) : -1;
    key[k][0] = ~0ull - (u64)(lane);        // 64 distinct "empty" keys; the largest is replaced first
    key[k][1] = ~0ull - (u64)(lane + 32);
    thr[k] = ~0ull;
    best[k] = -INFINITY;
    besti[k] = -1;
  }
  for (int64_t rb = r0; rb < r1; rb += 32) {
    __syncthreads();
    {  // 32 rows x 64 anchors, 16 threads x float4 per row, 16 rows per pass
      const int c4 = (threadIdx.x & 15) * 4;
#pragma unroll
      for (int pass = 0; pass < 2; ++pass) {
        const int rr = pass * 16 + (threadIdx.x >> 4);
        float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
        if (rb + rr < r1) v = __ldg((const float4*)(base + (rb + rr) * 512 + c4));
        tile[rr][c4] = v.x;
        tile[rr][c4 + 1] = v.y;
        tile[rr][c4 + 2] = v.z;
        tile[rr][c4 + 3] = v.w;
      }
    }
    __syncthreads();
    const int64_t r = rb + lane;
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      const float v = tile[lane][warp * 8 + k];
      const bool ok = r < r1 && (int)r != self[k];
      if (ok && (v > best[k] || besti[k] < 0)) {  // ascending rows inside a lane: strict > keeps the lowest row
        best[k] = v;
        besti[k] = (int)r;
      }
      const u64 cand = ((u64)mine_f2ord(v) << 32) | (u64)(unsigned)r;
      unsigned pend = __ballot_sync(0xffffffffu, ok && cand < thr[k]);
      while (pend) {
        const int src = __ffs(pend) - 1;
        const u64 ck = __shfl_sync(0xffffffffu, cand, src);
        pend &= pend - 1;
        if (ck >= thr[k]) continue;  // an earlier insertion lowered the threshold below this candidate
        // replace the list's maximum, then find the new maximum (= threshold)
        u64 lm = key[k][0] > key[k][1] ? key[k][0] : key[k][1];
        u64 wm = lm;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
          const u64 other = __shfl_xor_sync(0xffffffffu, wm, o);
          wm = other > wm ? other : wm;
        }
        if (key[k][0] == wm) key[k][0] = ck;
        else if (key[k][1] == wm) key[k][1] = ck;
        lm = key[k][0] > key[k][1] ? key[k][0] : key[k][1];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
          const u64 other = __shfl_xor_sync(0xffffffffu, lm, o);
          lm = other > lm ? other : lm;
        }
        thr[k] = lm;
      }
    }
  }
  // ---- write the segment's results: arg-max and the L smallest in ascending (value, row) order
#pragma unroll
  for (int k = 0; k < 8; ++k) {
    const int a = a0 + warp * 8 + k;
    float bv = best[k];
    int bi = besti[k];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      const float ov = __shfl_xor_sync(0xffffffffu, bv, o);
      const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
      if (oi >= 0 && (bi < 0 || ov > bv || (ov == bv && oi < bi))) {
        bv = ov;
        bi = oi;
      }
    }
    u64 srt[4] = {key[k][0], key[k][1], ~0ull, ~0ull};
    bitonic_sort128(srt, lane);
    if (a < A) {
      if (lane == 0) {
        pbest_v[(int64_t)seg * A + a] = bv;
        pbest_i[(int64_t)seg * A + a] = bi;
      }
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const int e = lane * 4 + j;
        if (e < L) {
          const bool real = (srt[j] >> 32) != 0xffffffffull;  // "empty" keys have all-ones upper words
          plist_v[((int64_t)seg * L + e) * A + a] = real ? mine_ord2f((unsigned)(srt[j] >> 32)) : INFINITY;
          plist_i[((int64_t)seg * L + e) * A + a] = real ? (int)(srt[j] & 0xffffffffu) : -1;
        }
      }
    }
  }
}

// thread = anchor: merge the segments, then pick the micro negatives among the spatial neighbours
__global__ void mine_merge_kernel(const float* __restrict__ S, int64_t n_pad, int A, const int32_t* __restrict__ anchors,
                                  const float* __restrict__ pbest_v, const int32_t* __restrict__ pbest_i,
                                  const float* __restrict__ plist_v, const int32_t* __restrict__ plist_i, int L,
                                  const int32_t* __restrict__ nbr, int K, int n_macro, int n_micro,
                                  int32_t* __restrict__ positive, int32_t* __restrict__ negatives) {
  const int a = blockIdx.x * blockDim.x + threadIdx.x;
  if (a >= A) return;
  float best = -INFINITY;
  int pos = -1;
  for (int s = 0; s < MINE_SEG; ++s) {  // ascending segments = ascending rows: strict > keeps the lowest index
    const float v = pbest_v[(int64_t)s * A + a];
    const int i = pbest_i[(int64_t)s * A + a];
    if (i >= 0 && (pos < 0 || v > best)) {
      best = v;
      pos = i;
    }
  }
  positive[a] = pos;
  int32_t* neg = negatives + (int64_t)a * (n_macro + n_micro);
  int head[MINE_SEG];
  for (int s = 0; s < MINE_SEG; ++s) head[s] = 0;
  for (int j = 0; j < n_macro; ++j) {
    int bs = -1, bi = -1;
    float bv = INFINITY;
    for (int s = 0; s < MINE_SEG; ++s) {
      while (head[s] < L && plist_i[((int64_t)s * L + head[s]) * A + a] == pos) ++head[s];  // the positive is excluded
      if (head[s] >= L) continue;
      const int i = plist_i[((int64_t)s * L + head[s]) * A + a];
      if (i < 0) continue;
      const float v = plist_v[((int64_t)s * L + head[s]) * A + a];
      if (bs < 0 || mine_less(v, i, bv, bi)) {
        bs = s;
        bv = v;
        bi = i;
      }
    }
    neg[j] = bi;
    if (bs >= 0) ++head[bs];
  }
  // micro negatives: the n_micro smallest similarities among the K spatial neighbours, (value, list position) order
  const int chunk = a >> 9, col = a & 511;
  const float* base = S + ((int64_t)chunk * n_pad) * 512 + col;
  const int32_t* nb = nbr + (int64_t)anchors[a] * K;
  float mv[32];
  int mk[32];
  int cnt = 0;
  for (int k = 0; k < K; ++k) {
    const int p = nb[k];
    const float v = p == pos ? INFINITY : __ldg(base + (int64_t)p * 512);
    if (cnt < n_micro || mine_less(v, k, mv[cnt - 1], mk[cnt - 1])) {
      int q = cnt < n_micro ? cnt : n_micro - 1;
      while (q > 0 && mine_less(v, k, mv[q - 1], mk[q - 1])) {
        mv[q] = mv[q - 1];
        mk[q] = mk[q - 1];
        --q;
      }
      mv[q] = v;
      mk[q] = k;
      if (cnt < n_micro) ++cnt;
    }
  }
  for (int j = 0; j < n_micro; ++j) neg[n_macro + j] = j < cnt ? nb[mk[j]] : -1;
}

struct MineLayout {
  int kp, chunks, L;
  int64_t n_pad, ntiles;
  size_t off_planes, off_w, off_tiles, off_S, off_pbv, off_pbi, off_plv, off_pli, bytes;
};

bool mine_layout(int64_t n, int D, int A, int n_macro, MineLayout& M) {
  if (n < 1 || D < 1 || A < 1 || n >= (1ll << 31) / 8) return false;
  M.kp = (D + BK - 1) / BK * BK;
  M.chunks = (A + 511) / 512;
  M.L = n_macro + 1;
  M.n_pad = (n + TM - 1) / TM * TM;
  M.ntiles = (M.n_pad / TM) * M.chunks;
  if (M.n_pad * M.chunks >= (1ll << 31)) return false;  // TileInfo.c_row0 is an int
  size_t off = 0;
  auto take = [&](size_t b) {
    size_t o = off;
    off += gp_align_up(b, 1024);
    return o;
  };
  M.off_planes = take(sizeof(__nv_bfloat16) * 2 * (size_t)n * M.kp);
  M.off_w = take(sizeof(__nv_bfloat16) * 2 * (size_t)M.chunks * 512 * M.kp);
  M.off_tiles = take(sizeof(TileInfo) * (size_t)M.ntiles);
  M.off_S = take(sizeof(float) * (size_t)M.chunks * M.n_pad * 512);
  M.off_pbv = take(sizeof(float) * (size_t)MINE_SEG * A);
  M.off_pbi = take(sizeof(int32_t) * (size_t)MINE_SEG * A);
  M.off_plv = take(sizeof(float) * (size_t)MINE_SEG * M.L * A);
  M.off_pli = take(sizeof(int32_t) * (size_t)MINE_SEG * M.L * A);
  M.bytes = off;
  return true;
}

}  // namespace

extern "C" size_t gp_mine_workspace_bytes(int64_t n, int D, int A, int n_macro) {
  MineLayout M;
  if (!mine_layout(n, D, A, n_macro, M)) return 0;
  return M.bytes;
}

extern "C" int gp_mine_contrastive(const float* feat, int64_t stride, int D, int64_t n, const int32_t* anchors, int A,
                                   const int32_t* nbr, int K, int n_macro, int n_micro, int32_t* positive,
                                   int32_t* negatives, void* workspace, size_t workspace_bytes, gp_stream_t stream) {
  GP_CHECK_ARG(feat && anchors && nbr && positive && negatives, "gp_mine_contrastive: null argument");
  GP_CHECK_ARG(n_macro >= 1 && n_macro + 1 <= MINE_LMAX, "gp_mine_contrastive: n_macro=%d outside [1,%d]", n_macro,
               MINE_LMAX - 1);
  GP_CHECK_ARG(n_micro >= 0 && n_micro <= 32 && n_micro <= K, "gp_mine_contrastive: n_micro=%d outside [0,min(32,K)]",
               n_micro);
  GP_CHECK_ARG(n >= (int64_t)n_macro + 2, "gp_mine_contrastive: needs n >= n_macro + 2 points");
  GP_CHECK_ARG(stride >= D, "gp_mine_contrastive: stride < D");
  MineLayout M;
  GP_CHECK_ARG(mine_layout(n, D, A, n_macro, M), "gp_mine_contrastive: bad sizes (n=%lld, D=%d, A=%d)", (long long)n, D, A);
  GP_CHECK_ARG(((uintptr_t)workspace % 256) == 0, "gp_mine_contrastive: workspace must be 256-byte aligned");
  if (workspace_bytes < M.bytes) {
    gp_set_error("gp_mine_contrastive: workspace %zu < %zu", workspace_bytes, M.bytes);
    return GP_ERR_WORKSPACE;
  }
  GP_CHECK_ARG(encode_tiled_fn() != nullptr, "gp_mine_contrastive: cuTensorMapEncodeTiled unavailable");
  cudaStream_t st = (cudaStream_t)stream;
  static int sm_count = 0;
  if (!sm_count) {
    int dev = 0;
    GP_CUDA(cudaGetDevice(&dev));
    GP_CUDA(cudaDeviceGetAttribute(&sm_count, cudaDevAttrMultiProcessorCount, dev));
  }
  char* ws = (char*)workspace;
  __nv_bfloat16* planes = (__nv_bfloat16*)(ws + M.off_planes);
  __nv_bfloat16* wimg = (__nv_bfloat16*)(ws + M.off_w);
  TileInfo* tiles = (TileInfo*)(ws + M.off_tiles);
  float* S = (float*)(ws + M.off_S);
  mine_normalize_split_kernel<<<(unsigned)((n + 7) / 8), 256, 0, st>>>(feat, stride, D, M.kp, n, planes);
  GP_LAUNCH_CHECK();
  {
    const int64_t tot = (int64_t)M.chunks * 2 * M.kp * 512;
    mine_anchor_weights_kernel<<<(unsigned)((tot + 255) / 256), 256, 0, st>>>(planes, anchors, A, M.chunks, M.kp, wimg);
    GP_LAUNCH_CHECK();
  }
  mine_tiles_kernel<<<(unsigned)((M.ntiles + 255) / 256), 256, 0, st>>>(n, M.chunks, M.n_pad, tiles);
  GP_LAUNCH_CHECK();
  CUtensorMap map;
  GP_CHECK_ARG(make_weight_map(&map, wimg, (int64_t)M.chunks * 2 * 512 * (M.kp / BK), 256 / (PAIR ? 2 : 1)),
               "gp_mine_contrastive: cuTensorMapEncodeTiled failed");
  GemmParams P = {};
  P.a = planes;
  P.lda = 2 * (int64_t)M.kp;
  P.kp = M.kp;
  P.pair_in = nullptr;
  P.tiles = tiles;
  P.ntiles_dev = nullptr;
  P.ntiles_host = (int)M.ntiles;
  P.m_rows = n;
  P.cout_total = 512;
  P.c = S;
  P.ldc = 512;
  int grid = P.ntiles_host < sm_count ? P.ntiles_host : sm_count;
  int rc = launch_gemm<512>(map, P, grid, st);
  if (rc) return rc;
  float* pbv = (float*)(ws + M.off_pbv);
  int32_t* pbi = (int32_t*)(ws + M.off_pbi);
  float* plv = (float*)(ws + M.off_plv);
  int32_t* pli = (int32_t*)(ws + M.off_pli);
  dim3 pg((unsigned)((A + 63) / 64), MINE_SEG);
  mine_partial_kernel<<<pg, 256, 0, st>>>(S, n, M.n_pad, A, anchors, M.L, pbv, pbi, plv, pli);
  GP_LAUNCH_CHECK();
  mine_merge_kernel<<<(unsigned)((A + 127) / 128), 128, 0, st>>>(S, M.n_pad, A, anchors, pbv, pbi, plv, pli, M.L, nbr, K,
                                                               n_macro, n_micro, positive, negatives);
  GP_LAUNCH_CHECK();
  return GP_OK;
}
