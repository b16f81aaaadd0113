// Calibration probe (not product): how fast can ONE SM push a 128 x 512 fp32 tile (256 KB, rows contiguous) to global
// memory, as a function of the per-instruction store pattern and the number of storing warps?  The student GEMM's
// epilogue drains such a tile from TMEM while the tensor pipe waits (the accumulator fills TMEM, so it cannot be
// double-buffered); the `GP_STU_ABLATE` builds put the cost of the stores at ~2.1 ms of a 10.9 ms forward.
// One CTA per SM, each CTA writes its own sequence of tiles.  Patterns (rows x bytes per warp instruction):
//   0: 8 rows x 32 B  (STG.64, four lanes per sector -- the tcgen05.ld.16x256b fragment stored as is)
//   1: 8 rows x 64 B  (STG.128, a quad writes 64 contiguous bytes)
//   2: 4 rows x 128 B (STG.128, eight lanes per row -- the round-1 smem-transposed epilogue)
//   3: 2 rows x 256 B (STG.128)
//   4: 1 row  x 512 B (STG.128, fully coalesced)
//   5: cp.async.bulk shared -> global, 16 KB per copy, issued by one thread (the TMA engine instead of the LSU)
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o scripts/probes/store_bw scripts/probes/store_bw.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

constexpr int ROWS = 128, COLS = 512;

template <int PAT>
__global__ void __launch_bounds__(256) store_kernel(float* out, int tiles_per_cta, long long* clk) {
  const int nwarps = blockDim.x >> 5, warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int q = warp & 3;                                  // row quadrant, as the TMEM lane restriction forces
  const int chalf = warp >> 2, ncol = COLS / (nwarps / 4); // 8 warps: second set takes the upper column half
  const float v = (float)threadIdx.x;
  __syncthreads();
  const long long t0 = clock64();
  for (int t = 0; t < tiles_per_cta; ++t) {
    float* tile = out + ((size_t)blockIdx.x * tiles_per_cta + t) * ROWS * COLS + (size_t)q * 32 * COLS + chalf * ncol;
    if (PAT == 0) {
      for (int hb = 0; hb < 2; ++hb)
        for (int c0 = 0; c0 < ncol; c0 += 8) {
          float* p = tile + (size_t)(hb * 16 + (lane >> 2)) * COLS + c0 + 2 * (lane & 3);
          *(float2*)p = make_float2(v, v);
          *(float2*)(p + 8 * COLS) = make_float2(v, v);
        }
    } else {
      constexpr int LPR = PAT == 1 ? 4 : PAT == 2 ? 8 : PAT == 3 ? 16 : 32;  // lanes per row
      constexpr int RPI = 32 / LPR;                                           // rows per instruction
      for (int r0 = 0; r0 < 32; r0 += RPI)
        for (int c0 = 0; c0 < ncol; c0 += LPR * 4) {
          float* p = tile + (size_t)(r0 + lane / LPR) * COLS + c0 + (lane % LPR) * 4;
          *(float4*)p = make_float4(v, v, v, v);
        }
    }
  }
  __syncthreads();
  const long long t1 = clock64();
  if (threadIdx.x == 0) clk[blockIdx.x] = t1 - t0;
}

// pattern 0 again, with four MORE warps of the CTA gathering 128-byte lines by cp.async at full rate meanwhile (what the
// GEMM's gather warps do at the start of a drain): does load traffic through the LSU slow the stores down?
__global__ void __launch_bounds__(256) store_with_gather_kernel(float* out, const uint4* src, size_t src_lines,
                                                                int tiles_per_cta, long long* clk, volatile int* stop) {
  __shared__ __align__(128) unsigned char ring[32768];
  __shared__ int done;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == 0) done = 0;
  __syncthreads();
  if (warp < 4) {
    const int q = warp;
    const float v = (float)threadIdx.x;
    const long long t0 = clock64();
    for (int t = 0; t < tiles_per_cta; ++t) {
      float* tile = out + ((size_t)blockIdx.x * tiles_per_cta + t) * ROWS * COLS + (size_t)q * 32 * COLS;
      for (int hb = 0; hb < 2; ++hb)
        for (int c0 = 0; c0 < COLS; c0 += 8) {
          float* p = tile + (size_t)(hb * 16 + (lane >> 2)) * COLS + c0 + 2 * (lane & 3);
          *(float2*)p = make_float2(v, v);
          *(float2*)(p + 8 * COLS) = make_float2(v, v);
        }
    }
    const long long t1 = clock64();
    if (lane == 0) {
      atomicMax((unsigned long long*)&clk[blockIdx.x], (unsigned long long)(t1 - t0));
      atomicAdd(&done, 1);
    }
  } else {
    unsigned long long x = blockIdx.x * 977 + threadIdx.x / 8;
    const unsigned s0 = (unsigned)__cvta_generic_to_shared(ring);
    int it = 0;
    while (*(volatile int*)&done < 4) {
      x = x * 6364136223846793005ull + 1442695040888963407ull;
      const uint4* g = src + ((x >> 20) % src_lines) * 8 + (lane & 7);   // 8 lanes share one 128-byte line
      asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(s0 + ((it & 15) * 2048 + threadIdx.x % 128 * 16)), "l"(g));
      asm volatile("cp.async.commit_group;");
      asm volatile("cp.async.wait_group 8;");
      ++it;
    }
    asm volatile("cp.async.wait_group 0;");
  }
}

// pattern 5: the TMA engine instead of the LSU -- one thread issues cp.async.bulk shared -> global copies of 16 KB
__global__ void __launch_bounds__(128) bulk_store_kernel(float* out, int tiles_per_cta, long long* clk) {
  extern __shared__ __align__(128) unsigned char sm[];
  for (int i = threadIdx.x; i < 32768 / 4; i += blockDim.x) ((float*)sm)[i] = (float)i;
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  __syncthreads();
  const long long t0 = clock64();
  if (threadIdx.x == 0) {
    const unsigned s0 = (unsigned)__cvta_generic_to_shared(sm);
    for (int t = 0; t < tiles_per_cta; ++t) {
      char* tile = (char*)(out + ((size_t)blockIdx.x * tiles_per_cta + t) * ROWS * COLS);
      for (int c = 0; c < 16; ++c) {
        asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], 16384;" ::"l"(tile + c * 16384),
                     "r"(s0 + (c & 1) * 16384)
                     : "memory");
        asm volatile("cp.async.bulk.commit_group;" ::: "memory");
        asm volatile("cp.async.bulk.wait_group.read 4;" ::: "memory");
      }
    }
    asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
  }
  __syncthreads();
  const long long t1 = clock64();
  if (threadIdx.x == 0) clk[blockIdx.x] = t1 - t0;
}

void run_bulk(float* buf, long long* clk, int grid, int tiles) {
  cudaFuncSetAttribute(bulk_store_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 32768);
  bulk_store_kernel<<<grid, 128, 32768>>>(buf, tiles, clk);
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0);
  cudaEventCreate(&e1);
  cudaEventRecord(e0);
  bulk_store_kernel<<<grid, 128, 32768>>>(buf, tiles, clk);
  cudaEventRecord(e1);
  cudaDeviceSynchronize();
  float ms = 0;
  cudaEventElapsedTime(&ms, e0, e1);
  long long h[256], mx = 0;
  cudaMemcpy(h, clk, sizeof(long long) * grid, cudaMemcpyDeviceToHost);
  for (int i = 0; i < grid; ++i) mx = h[i] > mx ? h[i] : mx;
  const double bytes = (double)tiles * ROWS * COLS * 4;
  printf("pattern 5 (cp.async.bulk smem->global)  grid %3d : %7.1f B/clk/SM  (%.0f clk per 256 KB tile)  aggregate %6.0f GB/s  err=%s\n",
         grid, bytes / (double)mx, (double)mx / tiles, bytes * grid / ms / 1e6, cudaGetErrorString(cudaGetLastError()));
}

template <int PAT>
void run(float* buf, long long* clk, int grid, int threads, int tiles) {
  store_kernel<PAT><<<grid, threads>>>(buf, tiles, clk);  // warm-up
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0);
  cudaEventCreate(&e1);
  cudaEventRecord(e0);
  store_kernel<PAT><<<grid, threads>>>(buf, tiles, clk);
  cudaEventRecord(e1);
  cudaDeviceSynchronize();
  float ms = 0;
  cudaEventElapsedTime(&ms, e0, e1);
  long long h[256], mx = 0;
  cudaMemcpy(h, clk, sizeof(long long) * grid, cudaMemcpyDeviceToHost);
  for (int i = 0; i < grid; ++i) mx = h[i] > mx ? h[i] : mx;
  const double bytes = (double)tiles * ROWS * COLS * 4;
  printf("pattern %d  warps %d  grid %3d : %7.1f B/clk/SM  (%.0f clk per 256 KB tile)  aggregate %6.0f GB/s  err=%s\n", PAT,
         threads / 32, grid, bytes / (double)mx, (double)mx / tiles, bytes * grid / ms / 1e6,
         cudaGetErrorString(cudaGetLastError()));
}

int main() {
  const int tiles = 32;  // 8 MB per CTA, 1.2 GB in all: far beyond L2, like the staging buffer
  float* buf;
  long long* clk;
  cudaMalloc(&buf, (size_t)148 * tiles * ROWS * COLS * 4);
  cudaMalloc(&clk, 256 * sizeof(long long));
  for (int grid : {148, 1}) {
    {
      cudaMemset(clk, 0, 256 * sizeof(long long));
      store_with_gather_kernel<<<grid, 256>>>(buf, (const uint4*)buf, (size_t)148 * tiles * ROWS * COLS * 4 / 128, tiles, clk, nullptr);
      cudaDeviceSynchronize();
      long long h[256], mx = 0;
      cudaMemcpy(h, clk, sizeof(long long) * grid, cudaMemcpyDeviceToHost);
      for (int i = 0; i < grid; ++i) mx = h[i] > mx ? h[i] : mx;
      printf("pattern 0 + 4 gather warps (cp.async)  grid %3d : %7.1f B/clk/SM  (%.0f clk per 256 KB tile)  err=%s\n", grid,
             (double)tiles * ROWS * COLS * 4 / (double)mx, (double)mx / tiles, cudaGetErrorString(cudaGetLastError()));
    }
    run_bulk(buf, clk, grid, tiles);
    for (int threads : {128}) {
      run<0>(buf, clk, grid, threads, tiles);
      run<1>(buf, clk, grid, threads, tiles);
      run<2>(buf, clk, grid, threads, tiles);
      run<3>(buf, clk, grid, threads, tiles);
      run<4>(buf, clk, grid, threads, tiles);
    }
  }
  return 0;
}
