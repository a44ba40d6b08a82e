// micro-benchmark: cost of row gathers with 16-byte cp.async from an L2-resident matrix
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
template <int MODE>
__global__ void k(const uint8_t* mat, int nrows, int rowbytes, int rows_per_group, int groups, long long* out) {
  extern __shared__ __align__(128) uint8_t sm[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int ppr = rowbytes / 16;
  uint8_t* my = sm + (size_t)warp * rows_per_group * rowbytes * 2;   // two groups of buffers
  uint32_t seed = (blockIdx.x * 64 + warp) * 2654435761u + 12345u;
  long long t0 = clock64();
  for (int g = 0; g < groups; ++g) {
    uint32_t base = smem_u32(my + (size_t)(g & 1) * rows_per_group * rowbytes) + lane * 16;
    for (int r = 0; r < rows_per_group; ++r) {
      seed = seed * 1664525u + 1013904223u;
      const int row = (seed >> 8) % nrows;
      const uint8_t* src = mat + (size_t)row * rowbytes + lane * 16;
      if (MODE == 0) {
        if (lane < ppr) asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(base + r * rowbytes), "l"(src) : "memory");
        if (lane + 32 < ppr) asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(base + r * rowbytes + 512), "l"(src + 512) : "memory");
      } else {
        // plain LDG.128 + STS.128
        uint4 v0, v1;
        if (lane < ppr) v0 = *reinterpret_cast<const uint4*>(src);
        if (lane + 32 < ppr) v1 = *reinterpret_cast<const uint4*>(src + 512);
        if (lane < ppr) *reinterpret_cast<uint4*>(my + (size_t)(g & 1) * rows_per_group * rowbytes + r * rowbytes + lane * 16) = v0;
        if (lane + 32 < ppr) *reinterpret_cast<uint4*>(my + (size_t)(g & 1) * rows_per_group * rowbytes + r * rowbytes + lane * 16 + 512) = v1;
      }
    }
    asm volatile("cp.async.commit_group;" ::: "memory");
    asm volatile("cp.async.wait_group 1;" ::: "memory");   // one group stays in flight
  }
  asm volatile("cp.async.wait_group 0;" ::: "memory");
  long long t1 = clock64();
  if (lane == 0 && blockIdx.x == 0) out[warp] = t1 - t0;
}
int main() {
  const int nrows = 61443;
  for (int rowbytes : {704, 320}) {
    uint8_t* mat; cudaMalloc(&mat, (size_t)nrows * rowbytes); cudaMemset(mat, 1, (size_t)nrows * rowbytes);
    long long* out; cudaMalloc(&out, 64 * 8);
    for (int mode = 0; mode < 2; ++mode)
    for (int warps : {1, 2, 4, 8}) for (int grid : {1, 148}) {
      const int rpg = 16, groups = 64;
      size_t smem = (size_t)warps * rpg * rowbytes * 2;
      auto kern = mode == 0 ? k<0> : k<1>;
      cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
      for (int rep = 0; rep < 3; ++rep) kern<<<grid, warps * 32, smem>>>(mat, nrows, rowbytes, rpg, groups, out);
      cudaError_t e = cudaDeviceSynchronize();
      long long h[64]; cudaMemcpy(h, out, 64 * 8, cudaMemcpyDeviceToHost);
      double clk_row = (double)h[0] / (rpg * groups);
      printf("%s rowbytes %d warps %d grid %3d : %7.1f clk/row/warp -> %6.1f B/clk/SM  (%s)\n", mode ? "LDG+STS" : "LDGSTS ", rowbytes, warps, grid, clk_row,
             rowbytes * warps / clk_row, cudaGetErrorString(e));
    }
    cudaFree(mat); cudaFree(out);
  }
  return 0;
}
