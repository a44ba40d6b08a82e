// micro-benchmark: tcgen05.ld (TMEM -> registers) throughput per SM by number of reading warps
// and columns per instruction, alone and under a concurrent tcgen05.mma stream (A in TMEM, M=128,
// N=128, accumulating into columns the readers do not touch).
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -I graphnet_b200/csrc tmem_ld_bench.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
#include <cuda.h>
#include "../../graphnet_b200/csrc/tc_common.cuh"
using namespace gn::tc;
__device__ __forceinline__ void mma_ts(uint32_t d, uint32_t a, uint64_t b, uint32_t idesc, uint32_t acc) {
  asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t}" ::"r"(d), "r"(a), "l"(b), "r"(idesc), "r"(acc) : "memory");
}
template <int X> struct Ld;
template <> struct Ld<16> { static __device__ __forceinline__ float go(uint32_t a) {
  uint32_t r[16];
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
    : "=r"(r[0]),"=r"(r[1]),"=r"(r[2]),"=r"(r[3]),"=r"(r[4]),"=r"(r[5]),"=r"(r[6]),"=r"(r[7]),"=r"(r[8]),"=r"(r[9]),"=r"(r[10]),"=r"(r[11]),"=r"(r[12]),"=r"(r[13]),"=r"(r[14]),"=r"(r[15]) : "r"(a) : "memory");
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
  return __uint_as_float(r[0] ^ r[15]); } };
template <> struct Ld<32> { static __device__ __forceinline__ float go(uint32_t a) {
  float v[32]; tmem_ld_32x32(a, v); return v[0] + v[31]; } };
template <> struct Ld<64> { static __device__ __forceinline__ float go(uint32_t a) {
  uint32_t r[64];
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x64.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31,%32,%33,%34,%35,%36,%37,%38,%39,%40,%41,%42,%43,%44,%45,%46,%47,%48,%49,%50,%51,%52,%53,%54,%55,%56,%57,%58,%59,%60,%61,%62,%63}, [%64];"
    : "=r"(r[0]),"=r"(r[1]),"=r"(r[2]),"=r"(r[3]),"=r"(r[4]),"=r"(r[5]),"=r"(r[6]),"=r"(r[7]),"=r"(r[8]),"=r"(r[9]),"=r"(r[10]),"=r"(r[11]),"=r"(r[12]),"=r"(r[13]),"=r"(r[14]),"=r"(r[15]),
      "=r"(r[16]),"=r"(r[17]),"=r"(r[18]),"=r"(r[19]),"=r"(r[20]),"=r"(r[21]),"=r"(r[22]),"=r"(r[23]),"=r"(r[24]),"=r"(r[25]),"=r"(r[26]),"=r"(r[27]),"=r"(r[28]),"=r"(r[29]),"=r"(r[30]),"=r"(r[31]),
      "=r"(r[32]),"=r"(r[33]),"=r"(r[34]),"=r"(r[35]),"=r"(r[36]),"=r"(r[37]),"=r"(r[38]),"=r"(r[39]),"=r"(r[40]),"=r"(r[41]),"=r"(r[42]),"=r"(r[43]),"=r"(r[44]),"=r"(r[45]),"=r"(r[46]),"=r"(r[47]),
      "=r"(r[48]),"=r"(r[49]),"=r"(r[50]),"=r"(r[51]),"=r"(r[52]),"=r"(r[53]),"=r"(r[54]),"=r"(r[55]),"=r"(r[56]),"=r"(r[57]),"=r"(r[58]),"=r"(r[59]),"=r"(r[60]),"=r"(r[61]),"=r"(r[62]),"=r"(r[63]) : "r"(a) : "memory");
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
  return __uint_as_float(r[0] ^ r[63]); } };

// warp 0: MMA issuer (if with_mma); warps 1..nld: readers (TMEM lane quadrant = warp & 3)
template <int X>
__global__ void k(int nld, int loads, int with_mma, long long* out) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  __shared__ uint64_t bar; __shared__ uint32_t tbase; __shared__ int stop;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int i = threadIdx.x; i < 32768 / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(smem)[i] = 0;
  if (threadIdx.x == 0) { mbar_init(&bar, 1); fence_mbar_init(); stop = 0; }
  if (warp == 0) tmem_alloc<512>(&tbase);
  tc_fence_before(); __syncthreads(); tc_fence_after();
  fence_proxy_async_smem();
  const uint32_t tb = tbase;
  if (warp == 0) {
    if (with_mma) {
      const uint32_t idesc = make_idesc_f16(128, 128);
      const uint64_t b128 = make_kmajor_desc<128>(smem_u32(smem));
      long long t0 = clock64(); int count = 0;
      // keep issuing until the readers are done
      while (*(volatile int*)&stop < nld) {
        if (elect_one()) {
#pragma unroll
          for (int u = 0; u < 8; ++u) mma_ts(tb + 384, tb + (u & 7) * 8, b128 + (uint64_t)((u & 3) * 2), idesc, 1);
          mma_commit(&bar);
        }
        __syncwarp();
        mbar_wait(&bar, count & 1);
        ++count;
      }
      long long t1 = clock64();
      if (lane == 0) { out[2] = t1 - t0; out[3] = count * 8; }
    }
  } else if (warp <= nld) {
    const uint32_t a = tb + ((uint32_t)((warp & 3) * 32) << 16);
    float acc = 0;
    __syncwarp();
    long long t0 = clock64();
    for (int i = 0; i < loads; ++i) acc += Ld<X>::go(a + ((i * X) & 127));   // columns 0..255 (+X)
    long long t1 = clock64();
    if (acc == 123.f) out[5] = 1;
    if (lane == 0) { atomicMax((unsigned long long*)&out[0], (unsigned long long)(t1 - t0)); atomicAdd(&stop, 1); }
  }
  tc_fence_before(); __syncthreads();
  if (warp == 0) tmem_dealloc<512>(tb);
}
template <int X> void run(long long* out, int nld, int with_mma) {
  const int loads = 2048;
  cudaMemset(out, 0, 64);
  cudaFuncSetAttribute(k<X>, cudaFuncAttributeMaxDynamicSharedMemorySize, 32768 + 1024);
  k<X><<<1, 32 * 17, 32768 + 1024>>>(nld, loads, with_mma, out);
  cudaError_t e = cudaDeviceSynchronize();
  long long h[4]; cudaMemcpy(h, out, 32, cudaMemcpyDeviceToHost);
  const double bytes = (double)nld * loads * 32 * X * 4;
  printf("x%-3d readers=%2d %s: %7.1f clk/load/warp  %6.1f B/clk/SM", X, nld, with_mma ? "+MMA" : "    ",
         (double)h[0] / loads, bytes / (double)h[0]);
  if (with_mma) printf("   MMA %5.1f clk each (nominal 64)", (double)h[2] / (double)h[3]);
  printf("  (%s)\n", cudaGetErrorString(e));
}
int main() {
  long long* out; cudaMalloc(&out, 64);
  for (int with_mma : {0, 1}) for (int nld : {1, 4, 8, 16}) { run<16>(out, nld, with_mma); run<32>(out, nld, with_mma); run<64>(out, nld, with_mma); }
  return 0;
}
