// micro-benchmark: tcgen05.mma issue-to-complete cost per instruction by operand source / shape
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
#include <cuda.h>
#include "../../graphnet_b200/csrc/tc_common.cuh"
using namespace gn::tc;
__device__ __forceinline__ void mma_ts(uint32_t d, uint32_t a, uint64_t b, uint32_t idesc, uint32_t acc) {
  asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t}" ::"r"(d), "r"(a), "l"(b), "r"(idesc), "r"(acc) : "memory");
}
__device__ __forceinline__ void mma_ss(uint32_t d, uint64_t a, uint64_t b, uint32_t idesc, uint32_t acc) {
  asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(d), "l"(a), "l"(b), "r"(idesc), "r"(acc) : "memory");
}
template <int FILL>
__device__ __forceinline__ void mma_ts_ws(uint32_t d, uint32_t a, uint64_t b, uint32_t idesc, uint32_t acc) {
  if (FILL) asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.ws.cta_group::1.kind::f16.collector::b0::fill [%0], [%1], %2, %3, p;\n\t}" ::"r"(d), "r"(a), "l"(b), "r"(idesc), "r"(acc) : "memory");
  else asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.ws.cta_group::1.kind::f16.collector::b0::lastuse [%0], [%1], %2, %3, p;\n\t}" ::"r"(d), "r"(a), "l"(b), "r"(idesc), "r"(acc) : "memory");
}
// mode 0: TS SW64 B; 1: TS SW128 B; 2: TS ws pairs (SW128); 3: SS (A smem SW128, B SW128); 4: TS two D blocks alternating, non-ws (SW128)
template <int mode>
__global__ void k(int N, int count, int with_ld, long long* out) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  __shared__ uint64_t bar; __shared__ uint32_t tbase;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int i = threadIdx.x; i < 65536 / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(smem)[i] = 0;
  if (threadIdx.x == 0) { mbar_init(&bar, 1); fence_mbar_init(); }
  if (warp == 0) tmem_alloc<512>(&tbase);
  tc_fence_before(); __syncthreads(); tc_fence_after();
  fence_proxy_async_smem();
  const uint32_t tb = tbase;
  const uint32_t idesc = make_idesc_f16(128, N);
  if (warp == 0) {
    long long t0 = 0, t1 = 0;
    for (int rep = 0; rep < 3; ++rep) {
      t0 = clock64();
      if (elect_one()) {
        const uint64_t b64 = make_kmajor_desc<64>(smem_u32(smem));
        const uint64_t b128 = make_kmajor_desc<128>(smem_u32(smem));
        const uint64_t a128 = make_kmajor_desc<128>(smem_u32(smem) + 32768);
        // tight issue loop: 8 MMAs per iteration, constant operands (A columns 0..63, B K-steps 0..3)
#pragma unroll 1
        for (int i = 0; i < count; i += 8) {
#pragma unroll
          for (int u = 0; u < 8; ++u) {
            const uint32_t a = tb + (u & 7) * 8;
            const uint64_t boff = (uint64_t)((u & 3) * 2);
            if (mode == 0) mma_ts(tb + 256, a, b64 + (u & 1) * 2, idesc, 1);
            else if (mode == 1) mma_ts(tb + 256, a, b128 + boff, idesc, 1);
            else if (mode == 2) { if (u & 1) mma_ts_ws<0>(tb + 256 + N, a + 168, b128 + (uint64_t)(((u >> 1) & 3) * 2), idesc, 1); else mma_ts_ws<1>(tb + 256, a, b128 + (uint64_t)(((u >> 1) & 3) * 2), idesc, 1); }
            else if (mode == 3) mma_ss(tb + 256, a128 + boff, b128 + boff, idesc, 1);
            else mma_ts((u & 1) ? tb + 256 + N : tb + 256, a + ((u & 1) ? 168 : 0), b128 + (uint64_t)(((u >> 1) & 3) * 2), idesc, 1);
          }
        }
        mma_commit(&bar);
      }
      __syncwarp();
      mbar_wait(&bar, rep & 1);
      t1 = clock64();
    }
    if (lane == 0) out[0] = t1 - t0;
  } else if (with_ld && warp >= 4 && warp < 8) {
    // concurrent TMEM reads (an epilogue's worth), unsynchronised: only the load on the read port matters
    float v[32]; float acc = 0;
    for (int i = 0; i < with_ld; ++i) { tmem_ld_32x32(tb + ((uint32_t)((warp & 3) * 32) << 16) + 256 + (i & 1) * 32, v); acc += v[0]; }
    if (acc == 123.f) out[1] = 1;
  }
  tc_fence_before(); __syncthreads();
  if (warp == 0) tmem_dealloc<512>(tb);
}
int main() {
  long long* out; cudaMalloc(&out, 64);
  cudaFuncSetAttribute(k<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, 66560 + 1024);
  cudaFuncSetAttribute(k<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, 66560 + 1024);
  cudaFuncSetAttribute(k<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, 66560 + 1024);
  cudaFuncSetAttribute(k<3>, cudaFuncAttributeMaxDynamicSharedMemorySize, 66560 + 1024);
  cudaFuncSetAttribute(k<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, 66560 + 1024);
  const char* names[] = {"TS  B=SW64 ", "TS  B=SW128", "TS.ws pairs", "SS  SW128  ", "TS 2 D blks"};
  for (int with_ld : {0, 4000}) for (int mode = 0; mode < 5; ++mode) for (int N : {64, 128, 256}) {
    if ((mode == 2 || mode == 4) && N == 256) continue;
    const int count = 512;
    auto kern = mode == 0 ? k<0> : mode == 1 ? k<1> : mode == 2 ? k<2> : mode == 3 ? k<3> : k<4>;
    kern<<<1, 256, 66560 + 1024>>>(N, count, with_ld, out);
    cudaError_t e = cudaDeviceSynchronize();
    long long h[2]; cudaMemcpy(h, out, 16, cudaMemcpyDeviceToHost);
    printf("%s N=%3d %s: %6.1f clk/MMA  (nominal %d)  (%s)\n", names[mode], N, with_ld ? "+LDTM" : "     ", (double)h[0] / count, 128 * N / 256, cudaGetErrorString(e));
    if (e != cudaSuccess) return 1;
  }
  return 0;
}
