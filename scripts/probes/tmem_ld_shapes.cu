// Calibration probe (not product): which TMEM (lane, column) does every register of the tcgen05.ld shapes
// .16x256b / .16x128b / .16x64b hold?  TMEM is filled through .32x32b (thread t of warp w <-> lane 32 w + t, register j
// <-> column j) with value = lane * 1000 + column, then read back with the other shapes and printed for warp 1.
//   nvcc -gencode arch=compute_100a,code=sm_100a -o scripts/probes/tmem_ld_shapes scripts/probes/tmem_ld_shapes.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__global__ void probe(int* out) {
  __shared__ uint32_t slot;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 64;" ::"r"(smem_u32(&slot)) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t base = slot;
  const uint32_t wbase = base + ((uint32_t)(warp * 32) << 16);
  uint32_t v[32];
#pragma unroll
  for (int j = 0; j < 32; ++j) v[j] = (uint32_t)((warp * 32 + lane) * 1000 + j);
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,"
      "%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31,%32};" ::"r"(wbase),
      "r"(v[0]), "r"(v[1]), "r"(v[2]), "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7]), "r"(v[8]), "r"(v[9]),
      "r"(v[10]), "r"(v[11]), "r"(v[12]), "r"(v[13]), "r"(v[14]), "r"(v[15]), "r"(v[16]), "r"(v[17]), "r"(v[18]),
      "r"(v[19]), "r"(v[20]), "r"(v[21]), "r"(v[22]), "r"(v[23]), "r"(v[24]), "r"(v[25]), "r"(v[26]), "r"(v[27]),
      "r"(v[28]), "r"(v[29]), "r"(v[30]), "r"(v[31])
      : "memory");
  asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  // record layout: out[shape][warp][lane][reg]  (16 regs per shape slot)
  uint32_t r[16];
  int* o;
  // shape 0: 16x256b.x1 at lane offset 0 (4 regs)
  asm volatile("tcgen05.ld.sync.aligned.16x256b.x1.b32 {%0,%1,%2,%3}, [%4];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(wbase) : "memory");
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
  o = out + ((0 * 4 + warp) * 32 + lane) * 16;
  for (int j = 0; j < 4; ++j) o[j] = (int)r[j];
  // shape 1: 16x256b.x1 at lane offset 16
  asm volatile("tcgen05.ld.sync.aligned.16x256b.x1.b32 {%0,%1,%2,%3}, [%4];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(wbase + (16u << 16)) : "memory");
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
  o = out + ((1 * 4 + warp) * 32 + lane) * 16;
  for (int j = 0; j < 4; ++j) o[j] = (int)r[j];
  // shape 2: 16x256b.x4 (16 regs) at column 0
  asm volatile("tcgen05.ld.sync.aligned.16x256b.x4.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
                 "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
               : "r"(wbase) : "memory");
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
  o = out + ((2 * 4 + warp) * 32 + lane) * 16;
  for (int j = 0; j < 16; ++j) o[j] = (int)r[j];
  // shape 3: 16x128b.x2 (4 regs)
  asm volatile("tcgen05.ld.sync.aligned.16x128b.x2.b32 {%0,%1,%2,%3}, [%4];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(wbase) : "memory");
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
  o = out + ((3 * 4 + warp) * 32 + lane) * 16;
  for (int j = 0; j < 4; ++j) o[j] = (int)r[j];
  // shape 4: 16x64b.x4 (4 regs)
  asm volatile("tcgen05.ld.sync.aligned.16x64b.x4.b32 {%0,%1,%2,%3}, [%4];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(wbase) : "memory");
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
  o = out + ((4 * 4 + warp) * 32 + lane) * 16;
  for (int j = 0; j < 4; ++j) o[j] = (int)r[j];
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 64;" ::"r"(base) : "memory");
}

int main() {
  int* d;
  const int n = 5 * 4 * 32 * 16;
  cudaMalloc(&d, n * sizeof(int));
  cudaMemset(d, 0xff, n * sizeof(int));
  probe<<<1, 128>>>(d);
  cudaError_t e = cudaDeviceSynchronize();
  if (e != cudaSuccess) { printf("error: %s\n", cudaGetErrorString(e)); return 1; }
  static int h[5 * 4 * 32 * 16];
  cudaMemcpy(h, d, sizeof(h), cudaMemcpyDeviceToHost);
  const char* names[5] = {"16x256b.x1 @lane+0", "16x256b.x1 @lane+16", "16x256b.x4", "16x128b.x2", "16x64b.x4"};
  const int nreg[5] = {4, 4, 16, 4, 4};
  for (int s = 0; s < 5; ++s) {
    printf("== %s (warp 1: value = lane*1000 + column, lane in [32,64))\n", names[s]);
    for (int t = 0; t < 32; ++t) {
      printf("t%02d:", t);
      for (int j = 0; j < nreg[s]; ++j) {
        int v = h[((s * 4 + 1) * 32 + t) * 16 + j];
        printf(" (%d,%d)", v / 1000, v % 1000);
      }
      printf("\n");
    }
  }
  return 0;
}
