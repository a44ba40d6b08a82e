// micro-test: TMA tile::gather4 - semantics (box rows, OOB row index) and throughput
#include <cstdio>
#include <cstdint>
#include <vector>
#include <cuda.h>
#include <cuda_runtime.h>
#include <cuda_fp16.h>
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* b, uint32_t c) { asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(b)), "r"(c)); }
__device__ __forceinline__ void mbar_expect(uint64_t* b, uint32_t bytes) { asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(b)), "r"(bytes) : "memory"); }
__device__ __forceinline__ void mbar_wait(uint64_t* b, uint32_t parity) {
  uint32_t ok = 0; int spins = 0;
  while (!ok) { asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(ok) : "r"(smem_u32(b)), "r"(parity) : "memory"); if (++spins > (1 << 22)) __trap(); }
}
__device__ __forceinline__ void gather4(uint32_t dst, const CUtensorMap* map, uint64_t* bar, int col, int r0, int r1, int r2, int r3) {
  asm volatile("cp.async.bulk.tensor.2d.shared::cta.global.tile::gather4.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, %6, %7}], [%2];"
               ::"r"(dst), "l"(map), "r"(smem_u32(bar)), "r"(col), "r"(r0), "r"(r1), "r"(r2), "r"(r3) : "memory");
}
// semantics: one op, dump smem
__global__ void sem(const __grid_constant__ CUtensorMap map, int box_cols, int r0, int r1, int r2, int r3, __half* out) {
  extern __shared__ __align__(1024) uint8_t sm[];
  __shared__ uint64_t bar;
  for (int i = threadIdx.x; i < 4 * box_cols; i += blockDim.x) reinterpret_cast<__half*>(sm)[i] = __float2half(-7.f);
  if (threadIdx.x == 0) { mbar_init(&bar, 1); asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
  __syncthreads();
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  if (threadIdx.x == 0) { mbar_expect(&bar, 4 * box_cols * 2); gather4(smem_u32(sm), &map, &bar, 0, r0, r1, r2, r3); }
  mbar_wait(&bar, 0);
  for (int i = threadIdx.x; i < 4 * box_cols; i += blockDim.x) out[i] = reinterpret_cast<__half*>(sm)[i];
}
// throughput: W warps, each warp's lane 0 issues `ops` gather4 per round into its own buffer, waits, repeats
__global__ void thr(const __grid_constant__ CUtensorMap map, int box_cols, int nrows, int ops, int rounds, long long* out) {
  extern __shared__ __align__(1024) uint8_t sm[];
  __shared__ uint64_t bars[32];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x < 32) { mbar_init(&bars[threadIdx.x], 1); }
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  __syncthreads();
  const uint32_t opbytes = 4 * box_cols * 2;
  uint8_t* my = sm + (size_t)warp * ops * opbytes;
  uint32_t seed = (blockIdx.x * 64 + warp) * 2654435761u + 99u;
  long long t0 = clock64(), tissue = 0;
  for (int r = 0; r < rounds; ++r) {
    if (lane == 0) {
      long long a = clock64();
      mbar_expect(&bars[warp], ops * opbytes);
      for (int o = 0; o < ops; ++o) {
        seed = seed * 1664525u + 1013904223u; int r0 = (seed >> 8) % nrows;
        seed = seed * 1664525u + 1013904223u; int r1 = (seed >> 8) % nrows;
        seed = seed * 1664525u + 1013904223u; int r2 = (seed >> 8) % nrows;
        seed = seed * 1664525u + 1013904223u; int r3 = (seed >> 8) % nrows;
        gather4(smem_u32(my + (size_t)o * opbytes), &map, &bars[warp], 0, r0, r1, r2, r3);
      }
      tissue += clock64() - a;
    }
    mbar_wait(&bars[warp], r & 1);
  }
  long long t1 = clock64();
  if (lane == 0 && blockIdx.x == 0) { out[warp] = t1 - t0; out[32 + warp] = tissue; }
}
typedef CUresult (*EncodeFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
int main() {
  void* fn = nullptr; cudaDriverEntryPointQueryResult q;
  cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &q);
  EncodeFn encode = (EncodeFn)fn;
  const int nrows = 61443, cols = 352;
  std::vector<__half> h((size_t)nrows * cols);
  for (int r = 0; r < nrows; ++r) for (int c = 0; c < cols; ++c) h[(size_t)r * cols + c] = __float2half((float)((r % 500) + c * 0.001f * 0 + (c == 0 ? 0.f : 0.f)));
  __half* mat; cudaMalloc(&mat, h.size() * 2); cudaMemcpy(mat, h.data(), h.size() * 2, cudaMemcpyHostToDevice);
  __half* out; cudaMalloc(&out, 4 * 256 * 2 * 4);
  long long* tout; cudaMalloc(&tout, 64 * 8);
  for (int box_rows : {1, 4}) {
    const int box_cols = 176;
    CUtensorMap map; cuuint64_t dims[2] = {(cuuint64_t)cols, (cuuint64_t)nrows}; cuuint64_t strides[1] = {(cuuint64_t)cols * 2};
    cuuint32_t box[2] = {(cuuint32_t)box_cols, (cuuint32_t)box_rows}; cuuint32_t estr[2] = {1, 1};
    CUresult r = encode(&map, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, mat, dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    printf("box_rows %d: encode -> %d\n", box_rows, (int)r);
    if (r != CUDA_SUCCESS) continue;
    sem<<<1, 128, 4 * box_cols * 2 + 1024>>>(map, box_cols, 5, 300, -1, 77, out);
    cudaError_t e = cudaDeviceSynchronize();
    printf("  sem kernel: %s\n", cudaGetErrorString(e));
    if (e != cudaSuccess) return 1;
    std::vector<__half> o(4 * box_cols); cudaMemcpy(o.data(), out, o.size() * 2, cudaMemcpyDeviceToHost);
    for (int i = 0; i < 4; ++i) printf("  smem row %d: first %g last %g (expect %d)\n", i, __half2float(o[i * box_cols]), __half2float(o[i * box_cols + box_cols - 1]), i == 0 ? 5 : i == 1 ? 300 : i == 2 ? 0 : 77);
    if (box_rows != 1 && box_rows != 4) continue;
    for (int warps : {1, 2, 4, 8}) for (int ops : {4, 16}) for (int grid : {1, 148}) {
      size_t smem = (size_t)warps * ops * 4 * box_cols * 2 + 1024;
      if (smem > 200 * 1024) continue;
      cudaFuncSetAttribute(thr, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
      const int rounds = 200;
      thr<<<grid, warps * 32, smem>>>(map, box_cols, nrows, ops, rounds, tout);
      thr<<<grid, warps * 32, smem>>>(map, box_cols, nrows, ops, rounds, tout);
      e = cudaDeviceSynchronize();
      long long t[64]; cudaMemcpy(t, tout, 64 * 8, cudaMemcpyDeviceToHost);
      double per_op = (double)t[0] / (rounds * ops);
      printf("  warps %d ops/round %2d grid %3d: %7.1f clk/op/warp (issue %6.1f) -> %6.1f B/clk/SM (%s)\n", warps, ops, grid, per_op, (double)t[32] / (rounds * ops),
             4.0 * box_cols * 2 * warps / per_op, cudaGetErrorString(e));
      if (e != cudaSuccess) return 1;
    }
  }
  return 0;
}
