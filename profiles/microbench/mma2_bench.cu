// micro-benchmark: tcgen05.mma.cta_group::2 (M = 256 over a CTA pair, A in TMEM, B = N/2 rows of
// shared memory in each CTA) - cycles per instruction for N = 64 / 128 / 256, kind::f16.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -I graphnet_b200/csrc mma2_bench.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
#include <cuda.h>
#include "../../graphnet_b200/csrc/tc_common.cuh"
using namespace gn::tc;
__device__ __forceinline__ void mma2_ts(uint32_t d, uint32_t a, uint64_t b, uint32_t idesc, uint32_t acc) {
  asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::2.kind::f16 [%0], [%1], %2, %3, p;\n\t}" ::"r"(d), "r"(a), "l"(b), "r"(idesc), "r"(acc) : "memory");
}
__device__ __forceinline__ void mma2_ss(uint32_t d, uint64_t a, uint64_t b, uint32_t idesc, uint32_t acc) {
  asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(d), "l"(a), "l"(b), "r"(idesc), "r"(acc) : "memory");
}
__device__ __forceinline__ void commit2(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;" ::"r"(smem_u32(bar)), "h"((uint16_t)3) : "memory");
}
__device__ __forceinline__ void csync() { asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory"); }
// mode 0: TS (A tmem), 1: SS
template <int mode>
__global__ void __cluster_dims__(2, 1, 1) k(int N, int count, long long* out) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  __shared__ uint64_t bar; __shared__ uint32_t tbase;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  uint32_t rank; asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(rank));
  for (int i = threadIdx.x; i < 65536 / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(smem)[i] = 0;
  if (threadIdx.x == 0) { mbar_init(&bar, 1); fence_mbar_init(); }
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"(smem_u32(&tbase)) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
  }
  fence_proxy_async_smem();
  tc_fence_before(); __syncthreads(); csync(); tc_fence_after();
  const uint32_t tb = tbase;
  const uint32_t idesc = make_idesc_f16(256, N);
  if (warp == 0 && rank == 0) {
    long long t0 = 0, t1 = 0;
    for (int rep = 0; rep < 3; ++rep) {
      t0 = clock64();
      if (elect_one()) {
        const uint64_t b128 = make_kmajor_desc<128>(smem_u32(smem));
        const uint64_t a128 = make_kmajor_desc<128>(smem_u32(smem) + 32768);
#pragma unroll 1
        for (int i = 0; i < count; i += 8) {
#pragma unroll
          for (int u = 0; u < 8; ++u) {
            const uint64_t boff = (uint64_t)((u & 3) * 2);
            if (mode == 0) mma2_ts(tb + 256, tb + (u & 7) * 8, b128 + boff, idesc, 1);
            else mma2_ss(tb + 256, a128 + boff, b128 + boff, idesc, 1);
          }
        }
        commit2(&bar);
      }
      __syncwarp();
      mbar_wait(&bar, rep & 1);
      t1 = clock64();
    }
    if (lane == 0) out[0] = t1 - t0;
  }
  tc_fence_before(); __syncthreads(); csync();
  if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, 512;" ::"r"(tb) : "memory");
}
int main() {
  long long* out; cudaMalloc(&out, 64);
  cudaFuncSetAttribute(k<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, 66560 + 1024);
  cudaFuncSetAttribute(k<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, 66560 + 1024);
  for (int mode = 0; mode < 2; ++mode) for (int N : {64, 128, 256}) {
    const int count = 512;
    auto kern = mode == 0 ? k<0> : k<1>;
    kern<<<2, 128, 66560 + 1024>>>(N, count, out);
    cudaError_t e = cudaDeviceSynchronize();
    long long h[2]; cudaMemcpy(h, out, 16, cudaMemcpyDeviceToHost);
    printf("cta_group::2 %s M=256 N=%3d: %6.1f clk/MMA  (nominal %d)  (%s)\n", mode ? "SS" : "TS", N, (double)h[0] / count, 256 * N / 512, cudaGetErrorString(e));
    if (e != cudaSuccess) return 1;
  }
  return 0;
}
