// micro-benchmark: shared-memory pipe throughput for the builder pattern (2 x LDS.128 + STS.128)
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
__device__ __forceinline__ uint4 lds128(uint32_t a) { uint4 r; asm volatile("ld.shared.v4.b32 {%0,%1,%2,%3}, [%4];" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "r"(a)); return r; }
__device__ __forceinline__ uint2 lds64(uint32_t a) { uint2 r; asm volatile("ld.shared.v2.b32 {%0,%1}, [%2];" : "=r"(r.x), "=r"(r.y) : "r"(a)); return r; }
__device__ __forceinline__ void sts128(uint32_t a, uint4 v) { asm volatile("st.shared.v4.b32 [%0], {%1,%2,%3,%4};" :: "r"(a), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory"); }
__device__ __forceinline__ void sts64(uint32_t a, uint2 v) { asm volatile("st.shared.v2.b32 [%0], {%1,%2};" :: "r"(a), "r"(v.x), "r"(v.y) : "memory"); }
// MODE 0: LDS.128 only; 1: 2xLDS.128 + STS.128 (builder); 2: same with 64-bit ops (twice as many); 3: STS.128 only
template <int MODE>
__global__ void k(int iters, long long* out, uint32_t* sink) {
  extern __shared__ __align__(128) uint8_t sm[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t base = (uint32_t)__cvta_generic_to_shared(sm) + warp * 8192;
  const uint32_t a16 = base + lane * 16, a8 = base + lane * 8;
  uint32_t acc = 0;
  __syncthreads();
  long long t0 = clock64();
  for (int i = 0; i < iters; ++i) {
#pragma unroll
    for (int u = 0; u < 8; ++u) {
      if (MODE == 0) { uint4 v = lds128(a16 + u * 512); acc ^= v.x ^ v.y ^ v.z ^ v.w; }
      else if (MODE == 1) { uint4 q = lds128(a16 + u * 512), p = lds128(a16 + 4096 + (u & 3) * 512); q.x += p.x; q.y += p.y; q.z += p.z; q.w += p.w; sts128(a16 + 4096 + 2048 + (u & 3) * 512, q); }
      else if (MODE == 2) {
        for (int h = 0; h < 2; ++h) { uint2 q = lds64(a8 + u * 512 + h * 256), p = lds64(a8 + 4096 + (u & 3) * 512 + h * 256); q.x += p.x; q.y += p.y; sts64(a8 + 4096 + 2048 + (u & 3) * 512 + h * 256, q); }
      } else { sts128(a16 + u * 512, make_uint4(i, u, lane, warp)); }
    }
  }
  long long t1 = clock64();
  if (lane == 0) out[warp] = t1 - t0;
  if (acc == 0x12345) sink[0] = acc;
}
int main() {
  long long* out; cudaMalloc(&out, 64 * 8); uint32_t* sink; cudaMalloc(&sink, 4);
  const char* names[] = {"LDS.128 only", "2xLDS.128+STS.128", "2x(2xLDS.64+STS.64)", "STS.128 only"};
  for (int mode = 0; mode < 4; ++mode) for (int warps : {1, 4, 8, 16}) {
    const int iters = 2000; size_t smem = warps * 8192;
    auto kern = mode == 0 ? k<0> : mode == 1 ? k<1> : mode == 2 ? k<2> : k<3>;
    cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    kern<<<1, warps * 32, smem>>>(iters, out, sink); kern<<<1, warps * 32, smem>>>(iters, out, sink);
    cudaError_t e = cudaDeviceSynchronize();
    long long h[64]; cudaMemcpy(h, out, 64 * 8, cudaMemcpyDeviceToHost);
    double per_unit = (double)h[0] / (iters * 8);   // clk per (512 B unit) per warp
    double bytes = (mode == 0 || mode == 3) ? 512.0 : 1536.0;
    printf("%-22s warps %2d: %6.1f clk/unit/warp -> %6.1f B/clk/SM (%s)\n", names[mode], warps, per_unit, bytes * warps / per_unit, cudaGetErrorString(e));
  }
  return 0;
}
