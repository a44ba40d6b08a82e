// Backward-pass kernels of the 16-bit DynEdge training path (BASELINE config 3).
//
// The reference has no backward code of its own: gradients come from torch autograd through
// PyG's EdgeConv (components/layers.py:62-76), torch_scatter's pooling (gnn/dynedge.py:252-265)
// and nn.Linear/ReLU (dynedge.py:193-232).  The edge MLP's backward is run UNFUSED here: the
// per-slot hidden activations are re-materialised (bf16) so that all three contractions
//     Z  = H  W2^T          (recomputed pre-activations, for the ReLU mask)
//     dH = G2 W2            (G2 = dOut[target] * [Z > 0])
//     dW2 = G2^T H          (K = all edge slots: split-K, fp32 reductions)
// are plain calls of the tcgen05 GEMM (tc_linear.cu).  The kernels in this file are the byte
// movers between them; all are HBM bound and written for coalesced 16-byte accesses.
#include "common.cuh"

#include <cuda_fp16.h>

#include <cstdlib>

namespace gn {

template <typename T> __device__ __forceinline__ float to_f32(T v);
template <> __device__ __forceinline__ float to_f32<float>(float v) { return v; }
template <> __device__ __forceinline__ float to_f32<__nv_bfloat16>(__nv_bfloat16 v) { return __bfloat162float(v); }
template <> __device__ __forceinline__ float to_f32<__half>(__half v) { return __half2float(v); }

template <typename T> __device__ __forceinline__ T from_f32(float v);
template <> __device__ __forceinline__ float from_f32<float>(float v) { return v; }
template <> __device__ __forceinline__ __nv_bfloat16 from_f32<__nv_bfloat16>(float v) { return __float2bfloat16_rn(v); }
template <> __device__ __forceinline__ __half from_f32<__half>(float v) {
  return __float2half_rn(fminf(fmaxf(v, -65504.f), 65504.f));
}

// ------------------------------------------------------------------ transpose (16-bit elements)
// dst[c, r] = src[r, c]; 64 x 64 tiles through shared memory, both sides coalesced.
__global__ void __launch_bounds__(256) transpose16_kernel(const uint16_t* __restrict__ src,
                                                          int64_t lds_, int64_t rows, int cols,
                                                          uint16_t* __restrict__ dst,
                                                          int64_t ldd) {
  __shared__ uint16_t tile[64][66];
  const int64_t r0 = (int64_t)blockIdx.x * 64;
  const int c0 = blockIdx.y * 64;
  const int tx = threadIdx.x & 63, ty = threadIdx.x >> 6;  // 64 x 4
#pragma unroll
  for (int i = 0; i < 16; ++i) {
    const int r = ty + 4 * i;
    tile[r][tx] = (r0 + r < rows && c0 + tx < cols) ? src[(r0 + r) * lds_ + c0 + tx] : (uint16_t)0;
  }
  __syncthreads();
#pragma unroll
  for (int i = 0; i < 16; ++i) {
    const int c = ty + 4 * i;
    if (c0 + c < cols && r0 + tx < rows) dst[(int64_t)(c0 + c) * ldd + r0 + tx] = tile[tx][c];
  }
}

// ------------------------------------------------------------------ column sums (bias gradients)
// out[n] += sum_m X[m, n].  A CTA takes 256 rows x 64 columns: 8 row groups x 32 column pairs,
// coalesced 128-byte (bf16: 4-byte per lane) row segments, a shared-memory reduction over the
// row groups and one atomic per column and CTA.
template <typename T>
__global__ void __launch_bounds__(256) colsum_kernel(const T* __restrict__ X, int64_t ldx,
                                                     int64_t M, int N, float* __restrict__ out) {
  __shared__ float part[8][64];
  const int cp = threadIdx.x & 31, rg = threadIdx.x >> 5;
  const int n = blockIdx.y * 64 + cp * 2;
  const int64_t m0 = (int64_t)blockIdx.x * 256;
  const int64_t m1 = m0 + 256 < M ? m0 + 256 : M;
  float a0 = 0.f, a1 = 0.f;
  if (n < N) {
    const bool two = n + 1 < N;
    for (int64_t m = m0 + rg; m < m1; m += 8) {
      a0 += to_f32<T>(X[m * ldx + n]);
      if (two) a1 += to_f32<T>(X[m * ldx + n + 1]);
    }
  }
  part[rg][cp * 2] = a0;
  part[rg][cp * 2 + 1] = a1;
  __syncthreads();
  if (threadIdx.x < 64) {
    const int c = blockIdx.y * 64 + threadIdx.x;
    if (c < N) {
      float acc = 0.f;
#pragma unroll
      for (int g = 0; g < 8; ++g) acc += part[g][threadIdx.x];
      atomicAdd(out + c, acc);
    }
  }
}

// ------------------------------------------------------------------ dZ = dY * [Y > 0], converted
// Y == nullptr: plain conversion.  Four elements per thread.
template <typename TG, typename TY, typename TO>
__global__ void __launch_bounds__(256) mask_convert_kernel(const TG* __restrict__ dY, int64_t ldg,
                                                           const TY* __restrict__ Y, int64_t ldy,
                                                           TO* __restrict__ out, int64_t ldo,
                                                           int64_t M, int N) {
  const int nq = (N + 3) / 4;
  const int64_t idx = (int64_t)blockIdx.x * 256 + threadIdx.x;
  if (idx >= M * nq) return;
  const int64_t m = idx / nq;
  const int n0 = (int)(idx % nq) * 4;
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    const int n = n0 + j;
    if (n < N) {
      float g = to_f32<TG>(dY[m * ldg + n]);
      if (Y != nullptr && !(to_f32<TY>(Y[m * ldy + n]) > 0.f)) g = 0.f;
      out[m * ldo + n] = from_f32<TO>(g);
    }
  }
}

// ------------------------------------------------------------------ fp32 -> 16-bit with a scale
template <typename TO>
__global__ void __launch_bounds__(256) convert_scale_kernel(const float* __restrict__ X, int64_t ldx,
                                                            TO* __restrict__ out, int64_t ldo,
                                                            int64_t M, int N, float scale) {
  const int nq = (N + 3) / 4;
  const int64_t idx = (int64_t)blockIdx.x * 256 + threadIdx.x;
  if (idx >= M * nq) return;
  const int64_t m = idx / nq;
  const int n0 = (int)(idx % nq) * 4;
#pragma unroll
  for (int j = 0; j < 4; ++j)
    if (n0 + j < N) out[m * ldo + n0 + j] = from_f32<TO>(X[m * ldx + n0 + j] * scale);
}

// ------------------------------------------------------------------ edge slots: H and H^T
// Row e = i * KP + t of the slot table (ELL, stride KP): H[e, :] = ReLU(P_i + Q_j) for a filled
// slot (j = nbr[e] >= 0), 0 for an empty one; bf16.  A CTA takes 32 slot rows: the row-major copy
// is written directly, the transposed copy HT[k, e] through a shared-memory tile.  Everything
// moves in 16-byte pieces of eight elements per lane (these kernels are bound by instruction
// issue, not by bytes, when they touch memory 4 bytes at a time).
//
// Shared tile [EB_ROWS][W] of 16-bit elements, W a multiple of 32: the 16-byte piece p of row r
// is stored at piece p ^ ((r >> 3) & 3), so that the transposing read below - four lanes on rows
// r, r+8, r+16, r+24 of the same column - touches four different bank groups.
// (128-row tiles, i.e. 256-byte runs per transposed row, were measured: 1.5-2x SLOWER - two
// CTAs of 98 KB per SM do not keep enough loads in flight.)
constexpr int EB_ROWS = 32;
__host__ __device__ __forceinline__ int tile_width(int cols) { return (cols + 31) & ~31; }
__device__ __forceinline__ uint4* tile_piece(uint16_t* tile, int W, int r, int p) {
  return reinterpret_cast<uint4*>(tile + r * W) + (p ^ ((r >> 3) & 3));
}
__device__ __forceinline__ uint16_t tile_at(const uint16_t* tile, int W, int r, int k) {
  return tile[r * W + ((((k >> 3) ^ ((r >> 3) & 3)) << 3) | (k & 7))];
}
// dstT[k, e0 + r] = tile[r, k] for k < cols, r < rvalid: a thread gathers eight rows of one
// column and stores them as one 16-byte piece of the transposed row.
__device__ __forceinline__ void tile_store_transposed(const uint16_t* tile, int W, int cols,
                                                      int rvalid, uint16_t* dstT, int64_t ldt,
                                                      int64_t e0) {
  constexpr int G = EB_ROWS / 8;  // 16-byte pieces per transposed row and CTA
  for (int idx = threadIdx.x; idx < cols * G; idx += blockDim.x) {
    const int k = idx / G, r8 = (idx % G) * 8;
    uint16_t v[8];
#pragma unroll
    for (int t = 0; t < 8; ++t) v[t] = tile_at(tile, W, r8 + t, k);
    uint16_t* dst = dstT + (int64_t)k * ldt + e0 + r8;
    if (r8 + 8 <= rvalid && ((reinterpret_cast<uintptr_t>(dst) & 15) == 0)) {
      *reinterpret_cast<uint4*>(dst) = *reinterpret_cast<const uint4*>(v);
    } else {
#pragma unroll
      for (int t = 0; t < 8; ++t)
        if (r8 + t < rvalid) dst[t] = v[t];
    }
  }
}

__device__ __forceinline__ uint4 pack8_bf16(const float (&f)[8]) {
  __nv_bfloat162 o[4];
#pragma unroll
  for (int x = 0; x < 4; ++x) o[x] = __floats2bfloat162_rn(f[2 * x], f[2 * x + 1]);
  return *reinterpret_cast<const uint4*>(o);
}

__global__ void __launch_bounds__(256) edge_build_h_kernel(
    const __half* __restrict__ P, int64_t ldp, const __half* __restrict__ Q, int64_t ldq, int Kp,
    const int32_t* __restrict__ nbr, int KP, int64_t E, __nv_bfloat16* __restrict__ H,
    int64_t ldh, __nv_bfloat16* __restrict__ HT, int64_t ldht) {
  extern __shared__ __align__(16) uint16_t tile[];  // [EB_ROWS][W]
  const int W = tile_width(Kp);
  const int pieces = Kp >> 3;
  const int64_t e0 = (int64_t)blockIdx.x * EB_ROWS;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int r = warp; r < EB_ROWS; r += 8) {
    const int64_t e = e0 + r;
    int32_t j = -1;
    int64_t i = 0;
    if (e < E) {
      j = nbr[e];
      i = (uint32_t)e / (uint32_t)KP;   // E < 2^31 (checked by the caller): 32-bit division
    }
    for (int pc = lane; pc < pieces; pc += 32) {
      float h[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
      if (j >= 0) {
        const uint4 pv = __ldg(reinterpret_cast<const uint4*>(P + i * ldp) + pc);
        const uint4 qv = __ldg(reinterpret_cast<const uint4*>(Q + (int64_t)j * ldq) + pc);
        const __half2* p2 = reinterpret_cast<const __half2*>(&pv);
        const __half2* q2 = reinterpret_cast<const __half2*>(&qv);
#pragma unroll
        for (int x = 0; x < 4; ++x) {
          const float2 pf = __half22float2(p2[x]), qf = __half22float2(q2[x]);
          h[2 * x] = fmaxf(pf.x + qf.x, 0.f);
          h[2 * x + 1] = fmaxf(pf.y + qf.y, 0.f);
        }
      }
      const uint4 hb = pack8_bf16(h);
      *tile_piece(tile, W, r, pc) = hb;
      if (H != nullptr && e < E) *(reinterpret_cast<uint4*>(H + e * ldh) + pc) = hb;
    }
  }
  if (HT == nullptr) return;
  __syncthreads();
  const int rvalid = (int)(E - e0 < EB_ROWS ? E - e0 : EB_ROWS);
  tile_store_transposed(tile, W, Kp, rvalid, reinterpret_cast<uint16_t*>(HT), ldht, e0);
}

// ------------------------------------------------------------------ G2 = dOut[target] * [Z > 0]
// Z [E, b] recomputed pre-activations (bf16), dOut [N, b] fp32.  Writes G2 [E, ldg] (all
// round_up(b, 8) columns; the pad columns as zeros) and G2T [b, E] (bf16).  Empty slots give
// zero rows.  MASK: the ReLU pattern comes from the bits the training forward recorded
// (gn_tc_edge_mlp_aggregate_mask) instead of from Z: one word per (32-channel group, tile row),
// bit c = [z(row, 32 g + c) > 0]; slot e sits in tile e / used at row e % used (used = slots of
// whole nodes in a 64-row tile).
template <bool MASK>
__global__ void __launch_bounds__(256) edge_grad_z_kernel(
    const __nv_bfloat16* __restrict__ Z, int64_t ldz, const uint32_t* __restrict__ mask,
    int64_t mask_ld, int used, const float* __restrict__ dOut, int64_t ldd, int b,
    const int32_t* __restrict__ nbr, int KP, int64_t E, __nv_bfloat16* __restrict__ G2,
    int64_t ldg, __nv_bfloat16* __restrict__ G2T, int64_t ldgt) {
  extern __shared__ __align__(16) uint16_t tile[];  // [EB_ROWS][W]
  const int W = tile_width(b);
  const int pieces = (b + 7) >> 3;
  const int64_t e0 = (int64_t)blockIdx.x * EB_ROWS;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int r = warp; r < EB_ROWS; r += 8) {
    const int64_t e = e0 + r;
    const bool filled = e < E && nbr[e] >= 0;
    const uint32_t e32 = (uint32_t)e;   // E < 2^31 (checked by the caller): 32-bit divisions
    const int64_t i = e < E ? e32 / (uint32_t)KP : 0;
    const int64_t mrow =
        (MASK && e < E) ? (int64_t)(e32 / (uint32_t)used) * 64 + (e32 % (uint32_t)used) : 0;
    for (int pc = lane; pc < pieces; pc += 32) {
      const int c = pc * 8;
      float g[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
      if (filled) {
        uint32_t bits;
        if (MASK) {
          bits = (__ldg(mask + (int64_t)(c >> 5) * mask_ld + mrow) >> (c & 31)) & 0xffu;
        } else {
          const uint4 zv = __ldg(reinterpret_cast<const uint4*>(Z + e * ldz) + pc);
          const __nv_bfloat162* z2 = reinterpret_cast<const __nv_bfloat162*>(&zv);
          bits = 0;
#pragma unroll
          for (int x = 0; x < 4; ++x) {
            const float2 zf = __bfloat1622float2(z2[x]);
            bits |= (zf.x > 0.f ? 1u : 0u) << (2 * x);
            bits |= (zf.y > 0.f ? 1u : 0u) << (2 * x + 1);
          }
        }
        const float4 d0 = __ldg(reinterpret_cast<const float4*>(dOut + i * ldd + c));
        const float4 d1 = __ldg(reinterpret_cast<const float4*>(dOut + i * ldd + c + 4));
        const float d[8] = {d0.x, d0.y, d0.z, d0.w, d1.x, d1.y, d1.z, d1.w};
#pragma unroll
        for (int x = 0; x < 8; ++x) g[x] = (c + x < b && ((bits >> x) & 1u)) ? d[x] : 0.f;
      }
      const uint4 gb = pack8_bf16(g);
      *tile_piece(tile, W, r, pc) = gb;
      if (e < E) *(reinterpret_cast<uint4*>(G2 + e * ldg) + pc) = gb;
    }
  }
  __syncthreads();
  const int rvalid = (int)(E - e0 < EB_ROWS ? E - e0 : EB_ROWS);
  tile_store_transposed(tile, W, b, rvalid, reinterpret_cast<uint16_t*>(G2T), ldgt, e0);
}

// ------------------------------------------------------------------ dP, dQ from dH
// m = dH * [H > 0]; dP_i = sum over the node's slots (one write), dQ_j += m (fp32 red.add).
// One warp per node, lanes stride the hidden columns in pairs.
__global__ void __launch_bounds__(256) edge_grad_pq_kernel(
    const __nv_bfloat16* __restrict__ dH, int64_t lddh, const __nv_bfloat16* __restrict__ H,
    int64_t ldh, int Kp, const int32_t* __restrict__ nbr, int KP, int64_t N,
    __nv_bfloat16* __restrict__ dP, int64_t ldp, float* __restrict__ dQ, int64_t ldq,
    float* __restrict__ dP32, int64_t ldp32) {
  const int64_t i = (int64_t)blockIdx.x * 8 + (threadIdx.x >> 5);
  const int lane = threadIdx.x & 31;
  if (i >= N) return;
  for (int k = lane * 2; k < Kp; k += 64) {
    float2 acc = make_float2(0.f, 0.f);
    for (int t = 0; t < KP; ++t) {
      const int64_t e = i * KP + t;
      const int32_t j = nbr[e];
      if (j < 0) continue;
      const float2 h = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(H + e * ldh + k));
      float2 g = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(dH + e * lddh + k));
      g.x = h.x > 0.f ? g.x : 0.f;
      g.y = h.y > 0.f ? g.y : 0.f;
      acc.x += g.x;
      acc.y += g.y;
      if (g.x != 0.f) atomicAdd(dQ + (int64_t)j * ldq + k, g.x);
      if (g.y != 0.f) atomicAdd(dQ + (int64_t)j * ldq + k + 1, g.y);
    }
    *reinterpret_cast<__nv_bfloat162*>(dP + i * ldp + k) = __floats2bfloat162_rn(acc.x, acc.y);
    if (dP32 != nullptr) *reinterpret_cast<float2*>(dP32 + i * ldp32 + k) = acc;
  }
}

// ------------------------------------------------------------------ slots grouped by SOURCE node
// For every node j of an ELL graph: the slots e = i*KP + t with nbr[e] == j, in ascending e
// (deterministic), as rslot[rstart[j] .. rstart[j] + rcount[j]).  Edges never leave an event, so
// the lists of event [lo, hi) live in rslot[lo*KP, hi*KP) and no global scan is needed.  One
// warp per event walks its slots three times (count, scan, fill); the fill takes 32 slots at a
// time and ranks equal sources inside the warp with match_any, so list order is slot order.
// rcount must be zero on entry.
__global__ void __launch_bounds__(128) ell_reverse_kernel(const int32_t* __restrict__ nbr, int KP,
                                                          const int32_t* __restrict__ ptr,
                                                          int64_t B, int32_t* __restrict__ rstart,
                                                          int32_t* __restrict__ rcount,
                                                          int32_t* __restrict__ rslot) {
  const int64_t ev = (int64_t)blockIdx.x * 4 + (threadIdx.x >> 5);
  const int lane = threadIdx.x & 31;
  if (ev >= B) return;
  const int lo = ptr[ev], hi = ptr[ev + 1];
  const int64_t s0 = (int64_t)lo * KP, s1 = (int64_t)hi * KP;
  for (int64_t e = s0 + lane; e < s1; e += 32) {
    const int32_t j = nbr[e];
    if (j >= 0) atomicAdd(rcount + j, 1);
  }
  __threadfence();
  __syncwarp();
  int32_t carry = (int32_t)s0;
  for (int base = lo; base < hi; base += 32) {
    const int j = base + lane;
    const int32_t c = j < hi ? __ldcg(rcount + j) : 0;
    int32_t incl = c;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int32_t t = __shfl_up_sync(0xffffffffu, incl, o);
      if (lane >= o) incl += t;
    }
    if (j < hi) {
      rstart[j] = carry + incl - c;
      rcount[j] = 0;  // re-used as the fill cursor; ends at the count again
    }
    carry += __shfl_sync(0xffffffffu, incl, 31);
  }
  __threadfence();
  __syncwarp();
  for (int64_t eb = s0; eb < s1; eb += 32) {
    const int64_t e = eb + lane;
    const int32_t j = e < s1 ? nbr[e] : -1;
    const uint32_t peers = __match_any_sync(0xffffffffu, j);
    const int rank = __popc(peers & ((1u << lane) - 1u));
    const int leader = __ffs(peers) - 1;
    int32_t base = 0;
    if (j >= 0 && lane == leader) base = atomicAdd(rcount + j, __popc(peers));
    base = __shfl_sync(0xffffffffu, base, leader);
    if (j >= 0) rslot[__ldcg(rstart + j) + base + rank] = (int32_t)e;
  }
}

// ------------------------------------------------------------------ dP, dQ from dH by gathering
// m[e] = dH[e] * [P_i + Q_j > 0] (the fp16 sum the forward kernel clamps: the same sign).
//   dP_i = sum of m over the node's own slots (contiguous rows of dH),
//   dQ_j = sum of m over the slots that name j as neighbour (ell_reverse_kernel's lists),
// both as plain sums in registers: no atomics, one write per output element, a fixed order.
// A CTA takes one event and `gch` 16-byte column pieces of every row: the event's rows are read
// from HBM by the first loop and found in L2 by the second.
__device__ __forceinline__ void pq_accumulate(float (&acc)[8], const uint4& g, const uint4& p,
                                              const uint4& q) {
  const __half2* p2 = reinterpret_cast<const __half2*>(&p);
  const __half2* q2 = reinterpret_cast<const __half2*>(&q);
  const __nv_bfloat162* g2 = reinterpret_cast<const __nv_bfloat162*>(&g);
#pragma unroll
  for (int x = 0; x < 4; ++x) {
    const float2 s = __half22float2(__hadd2(p2[x], q2[x]));
    const float2 gv = __bfloat1622float2(g2[x]);
    acc[2 * x] += s.x > 0.f ? gv.x : 0.f;
    acc[2 * x + 1] += s.y > 0.f ? gv.y : 0.f;
  }
}

__device__ __forceinline__ void pq_store(const float (&acc)[8], __nv_bfloat16* d16, float* d32) {
  __nv_bfloat162 o[4];
#pragma unroll
  for (int x = 0; x < 4; ++x) o[x] = __floats2bfloat162_rn(acc[2 * x], acc[2 * x + 1]);
  *reinterpret_cast<uint4*>(d16) = *reinterpret_cast<const uint4*>(o);
  if (d32 != nullptr) {
    *reinterpret_cast<float4*>(d32) = make_float4(acc[0], acc[1], acc[2], acc[3]);
    *reinterpret_cast<float4*>(d32 + 4) = make_float4(acc[4], acc[5], acc[6], acc[7]);
  }
}

__global__ void __launch_bounds__(256) edge_grad_pq_gather_kernel(
    const __nv_bfloat16* __restrict__ dH, int64_t lddh, const __half* __restrict__ P, int64_t ldp,
    const __half* __restrict__ Q, int64_t ldq, int Kp, const int32_t* __restrict__ nbr, int KP,
    const int32_t* __restrict__ ptr, const int32_t* __restrict__ rstart,
    const int32_t* __restrict__ rcount, const int32_t* __restrict__ rslot, int gch,
    __nv_bfloat16* __restrict__ dP, int64_t lddp, __nv_bfloat16* __restrict__ dQ, int64_t lddq,
    float* __restrict__ dP32, int64_t ldp32, float* __restrict__ dQ32, int64_t ldq32) {
  const int lo = ptr[blockIdx.x], n = ptr[blockIdx.x + 1] - lo;
  const int pieces = Kp / 8;
  const int c_first = blockIdx.y * gch;
  const int nc = min(gch, pieces - c_first);
  if (n <= 0 || nc <= 0) return;
  const int items = n * nc;
  // four slots per step: ids first, then all eight 16-byte loads, then the arithmetic - a
  // thread's loads are dependent chains (id -> row), so they must overlap each other
  for (int it = threadIdx.x; it < items; it += blockDim.x) {  // own slots -> dP
    const int64_t i = lo + it / nc;
    const int col = (c_first + it % nc) * 8;
    const uint4 p = __ldg(reinterpret_cast<const uint4*>(P + i * ldp + col));
    float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    for (int t0 = 0; t0 < KP; t0 += 4) {
      int32_t j[4];
      uint4 g[4], q[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) j[u] = t0 + u < KP ? __ldg(nbr + i * KP + t0 + u) : -1;
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        if (j[u] >= 0) {
          g[u] = *reinterpret_cast<const uint4*>(dH + (i * KP + t0 + u) * lddh + col);
          q[u] = __ldg(reinterpret_cast<const uint4*>(Q + (int64_t)j[u] * ldq + col));
        }
      }
#pragma unroll
      for (int u = 0; u < 4; ++u)
        if (j[u] >= 0) pq_accumulate(acc, g[u], p, q[u]);
    }
    pq_store(acc, dP + i * lddp + col, dP32 ? dP32 + i * ldp32 + col : nullptr);
  }
  for (int it = threadIdx.x; it < items; it += blockDim.x) {  // slots naming this node -> dQ
    const int64_t j = lo + it / nc;
    const int col = (c_first + it % nc) * 8;
    const uint4 q = __ldg(reinterpret_cast<const uint4*>(Q + j * ldq + col));
    float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    const int32_t* list = rslot + rstart[j];
    const int cnt = rcount[j];
    for (int r0 = 0; r0 < cnt; r0 += 4) {
      int64_t e[4];
      uint4 g[4], p[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) e[u] = r0 + u < cnt ? (int64_t)__ldg(list + r0 + u) : -1;
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        if (e[u] >= 0) {
          g[u] = *reinterpret_cast<const uint4*>(dH + e[u] * lddh + col);
          p[u] = __ldg(reinterpret_cast<const uint4*>(P + (e[u] / KP) * ldp + col));
        }
      }
#pragma unroll
      for (int u = 0; u < 4; ++u)
        if (e[u] >= 0) pq_accumulate(acc, g[u], p[u], q);
    }
    pq_store(acc, dQ + j * lddq + col, dQ32 ? dQ32 + j * ldq32 + col : nullptr);
  }
}

// ------------------------------------------------------------------ dP, dQ in one pass over dH
// The same sums from ONE streaming read of dH: a thread takes all the slots of one node for one
// 16-byte piece, keeps dP in registers and sends m to dQ32[j] with two red.global.add.v4.f32 per
// slot (four fp32 additions per L2 operation; the scalar form of this kernel spent its time on
// 97 M single-word atomics per layer, and shared-memory float atomics compile to a CAS loop).
// dQ32 must be zero on entry; the order of the additions is not fixed.
__device__ __forceinline__ void red_add_v4(float* addr, float a, float b, float c, float d) {
  asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(addr), "f"(a), "f"(b),
               "f"(c), "f"(d)
               : "memory");
}

__global__ void __launch_bounds__(256) edge_grad_pq_scatter_kernel(
    const __nv_bfloat16* __restrict__ dH, int64_t lddh, const __half* __restrict__ P, int64_t ldp,
    const __half* __restrict__ Q, int64_t ldq, int Kp, const int32_t* __restrict__ nbr, int KP,
    int64_t N, __nv_bfloat16* __restrict__ dP, int64_t lddp, float* __restrict__ dP32,
    int64_t ldp32, float* __restrict__ dQ32, int64_t ldq32) {
  const int pieces = Kp / 8;
  const int64_t it = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (it >= N * pieces) return;
  const int64_t i = it / pieces;
  const int col = (int)(it - i * pieces) * 8;
  const uint4 p = __ldg(reinterpret_cast<const uint4*>(P + i * ldp + col));
  const __half2* p2 = reinterpret_cast<const __half2*>(&p);
  float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
  for (int t0 = 0; t0 < KP; t0 += 4) {
    int32_t j[4];
    uint4 g[4], q[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) j[u] = t0 + u < KP ? __ldg(nbr + i * KP + t0 + u) : -1;
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      if (j[u] >= 0) {
        g[u] = __ldg(reinterpret_cast<const uint4*>(dH + (i * KP + t0 + u) * lddh + col));
        q[u] = __ldg(reinterpret_cast<const uint4*>(Q + (int64_t)j[u] * ldq + col));
      }
    }
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      if (j[u] < 0) continue;
      const __half2* q2 = reinterpret_cast<const __half2*>(&q[u]);
      const __nv_bfloat162* g2 = reinterpret_cast<const __nv_bfloat162*>(&g[u]);
      float m[8];
#pragma unroll
      for (int x = 0; x < 4; ++x) {
        const float2 sum = __half22float2(__hadd2(p2[x], q2[x]));
        const float2 gv = __bfloat1622float2(g2[x]);
        m[2 * x] = sum.x > 0.f ? gv.x : 0.f;
        m[2 * x + 1] = sum.y > 0.f ? gv.y : 0.f;
        acc[2 * x] += m[2 * x];
        acc[2 * x + 1] += m[2 * x + 1];
      }
      float* dst = dQ32 + (int64_t)j[u] * ldq32 + col;
      red_add_v4(dst, m[0], m[1], m[2], m[3]);
      red_add_v4(dst + 4, m[4], m[5], m[6], m[7]);
    }
  }
  pq_store(acc, dP + i * lddp + col, dP32 ? dP32 + i * ldp32 + col : nullptr);
}

// ------------------------------------------------------------------ EdgeConv weight repacking
// One launch per layer and parameter version (in training: per step) instead of two dozen
// small tensor ops on the host.  From nn.Linear's W1 [a, 2d], b1 [a], W2 [b, a], b2 [b]
// (reference gnn/dynedge.py:193-204) to the layouts the tensor-core kernels read:
//   Wpq  [2Kp, Kd]  rows 0..a-1 = W1[:, :d] - W1[:, d:] (acts on x_i), rows Kp..Kp+a-1 = W1[:, d:]
//                   (acts on x_j); fp32 (first layer) or 16-bit
//   bias [2Kp]      b1, then 1, 1 (the "ones" columns that carry b2), zeros
//   W2p  [b16, Kp]  fp16: W2, column a = hi(b2), a+1 = lo(b2)
// and, for the backward pass (pointers may be null): W2b = bf16(W2p), W2bT [Kp, b16] its
// transpose, WpqT [Kd, 2Kp] the transpose of the 16-bit Wpq.
struct PackParams {
  const float *W1, *b1, *W2, *b2;
  int d, a, b, Kp, Kd, b16;
  void* Wpq;
  int wpq_dtype;  // GN_F32 / GN_BF16 / GN_F16
  float* bias;
  __half* W2p;
  __nv_bfloat16 *W2b, *W2bT;
  void* WpqT;  // same dtype as Wpq (16-bit only)
  float ones_act;   // act(1.0) as the h builder computes it (fp16)
};

__device__ __forceinline__ void store_as(void* base, int64_t idx, int dtype, float v) {
  if (dtype == GN_F32) reinterpret_cast<float*>(base)[idx] = v;
  else if (dtype == GN_BF16) reinterpret_cast<__nv_bfloat16*>(base)[idx] = __float2bfloat16_rn(v);
  else reinterpret_cast<__half*>(base)[idx] = __float2half_rn(v);
}

__global__ void __launch_bounds__(256) edgeconv_pack_tc_kernel(PackParams p) {
  const int64_t n_pq = (int64_t)2 * p.Kp * p.Kd;
  const int64_t n_w2 = (int64_t)p.b16 * p.Kp;
  for (int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; idx < n_pq + n_w2 + 2 * p.Kp;
       idx += (int64_t)gridDim.x * blockDim.x) {
    if (idx < n_pq) {
      const int r = (int)(idx / p.Kd), c = (int)(idx % p.Kd);
      const int row = r < p.Kp ? r : r - p.Kp;
      float v = 0.f;
      if (row < p.a && c < p.d) {
        const float wa = p.W1[(int64_t)row * 2 * p.d + c];
        const float wb = p.W1[(int64_t)row * 2 * p.d + p.d + c];
        v = r < p.Kp ? wa - wb : wb;
      }
      store_as(p.Wpq, idx, p.wpq_dtype, v);
      if (p.WpqT != nullptr) store_as(p.WpqT, (int64_t)c * 2 * p.Kp + r, p.wpq_dtype, v);
    } else if (idx < n_pq + n_w2) {
      const int64_t t = idx - n_pq;
      const int r = (int)(t / p.Kp), c = (int)(t % p.Kp);
      __half h = __float2half_rn(0.f);
      if (r < p.b) {
        if (c < p.a) {
          h = __float2half_rn(p.W2[(int64_t)r * p.a + c]);
        } else if (p.b2 != nullptr && c < p.a + 2) {
          // h's "ones" columns hold act(1) (1 for ReLU / LeakyReLU, 0.8413 for GELU): the bias
          // is divided by exactly that fp16 value, so that act(1) * (hi + lo) = b2
          const float bv = p.b2[r] / p.ones_act;
          const __half hi = __float2half_rn(bv);
          h = c == p.a ? hi : __float2half_rn(bv - __half2float(hi));
        }
      }
      p.W2p[t] = h;
      if (p.W2b != nullptr) {
        const __nv_bfloat16 hb = __float2bfloat16_rn(__half2float(h));
        p.W2b[t] = hb;
        p.W2bT[(int64_t)c * p.b16 + r] = hb;
      }
    } else {
      const int c = (int)(idx - n_pq - n_w2);
      float v = 0.f;
      if (c < p.a) v = p.b1 != nullptr ? p.b1[c] : 0.f;
      else if (c < p.a + 2) v = 1.f;
      p.bias[c] = v;
    }
  }
}

// ------------------------------------------------------------------ out += G^T X, X narrow, fp32
// Weight gradient of the FIRST EdgeConv Linear (K = 2F+5 <= 32 raw input features): out[c, k] +=
// sum_n G[n, c] X[n, k], all fp32 - the two halves of that gradient are subtracted from each
// other (W1 acts on x_j - x_i), which amplifies any rounding of G or X.  A CTA takes 128 rows;
// a thread owns output rows c, c+256, ... with KMAX accumulators each; X rows come from smem.
template <int KMAX>
__global__ void __launch_bounds__(256) outer_accumulate_kernel(const float* __restrict__ G,
                                                               int64_t ldg, int cols,
                                                               const float* __restrict__ X,
                                                               int64_t ldx, int K, int64_t N,
                                                               float* __restrict__ out,
                                                               int64_t ldo) {
  __shared__ float xs[128][KMAX];
  const int64_t n0 = (int64_t)blockIdx.x * 128;
  const int rows = (int)(N - n0 < 128 ? N - n0 : 128);
  for (int idx = threadIdx.x; idx < 128 * KMAX; idx += 256) {
    const int r = idx / KMAX, k = idx % KMAX;
    xs[r][k] = (r < rows && k < K) ? X[(n0 + r) * ldx + k] : 0.f;
  }
  __syncthreads();
  for (int c = threadIdx.x; c < cols; c += 256) {
    float acc[KMAX];
#pragma unroll
    for (int k = 0; k < KMAX; ++k) acc[k] = 0.f;
    for (int r = 0; r < rows; ++r) {
      const float g = G[(n0 + r) * ldg + c];
#pragma unroll
      for (int k = 0; k < KMAX; ++k) acc[k] = fmaf(g, xs[r][k], acc[k]);
    }
#pragma unroll
    for (int k = 0; k < KMAX; ++k)
      if (k < K) atomicAdd(out + (int64_t)c * ldo + k, acc[k]);
  }
}

// ------------------------------------------------------------------ pooling backward
// dX[n, c] = sum over schemes of the gradient that reaches row n: sum -> g, mean -> g / count,
// min / max -> g at the FIRST row of the event that attains the pooled value (torch_scatter's
// CPU arg convention), 0 elsewhere.  One CTA per (event, 128 columns), one thread per column.
struct PoolSchemesB { int n; int s[8]; };
template <typename T>
__global__ void __launch_bounds__(128) segment_pool_backward_kernel(
    const T* __restrict__ x, int64_t ldx, int C, const int32_t* __restrict__ ptr,
    PoolSchemesB sch, const float* __restrict__ pooled, int64_t ldpool,
    const float* __restrict__ dpooled, int64_t lddp, float* __restrict__ dX, int64_t lddx) {
  const int64_t ev = blockIdx.x;
  const int c = blockIdx.y * 128 + threadIdx.x;
  if (c >= C) return;
  const int32_t n0 = ptr[ev], n1 = ptr[ev + 1];
  if (n1 <= n0) return;
  float g_all = 0.f;                      // same for every row: sum and mean
  float vmin = 0.f, vmax = 0.f, gmin = 0.f, gmax = 0.f;
  bool has_min = false, has_max = false;
  for (int k = 0; k < sch.n; ++k) {
    const float g = dpooled[ev * lddp + (int64_t)k * C + c];
    const float v = pooled[ev * ldpool + (int64_t)k * C + c];
    if (sch.s[k] == GN_POOL_SUM) g_all += g;
    else if (sch.s[k] == GN_POOL_MEAN) g_all += g / (float)(n1 - n0);
    else if (sch.s[k] == GN_POOL_MIN) { has_min = true; vmin = v; gmin += g; }
    else { has_max = true; vmax = v; gmax += g; }
  }
  bool min_done = !has_min, max_done = !has_max;
  for (int32_t n = n0; n < n1; ++n) {
    const float v = to_f32<T>(x[(int64_t)n * ldx + c]);
    float g = g_all;
    if (!min_done && v == vmin) { g += gmin; min_done = true; }
    if (!max_done && v == vmax) { g += gmax; max_done = true; }
    dX[(int64_t)n * lddx + c] = g;
  }
}

// ------------------------------------------------------------------ CSR -> slot table
// nbr[i*stride + t] = t-th in-edge of node i (order kept), -1 beyond its degree; flags |=
// GN_FLAG_DEGREE_OVERFLOW if a node has more than `stride` in-edges (those are dropped).
__global__ void __launch_bounds__(256) csr_to_ell_kernel(const int32_t* __restrict__ rowptr,
                                                         const int32_t* __restrict__ col,
                                                         int64_t N, int stride,
                                                         int32_t* __restrict__ nbr,
                                                         int32_t* __restrict__ deg,
                                                         int32_t* __restrict__ flags) {
  const int64_t idx = (int64_t)blockIdx.x * 256 + threadIdx.x;
  if (idx >= N * stride) return;
  const int64_t i = idx / stride;
  const int t = (int)(idx % stride);
  const int32_t a = rowptr[i], b = rowptr[i + 1];
  nbr[idx] = t < b - a ? col[a + t] : -1;
  if (t == 0) {
    deg[i] = b - a < stride ? b - a : stride;
    if (b - a > stride && flags) atomicOr(flags, GN_FLAG_DEGREE_OVERFLOW);
  }
}

}  // namespace gn

using namespace gn;

extern "C" int gn_transpose_16(const void* src, int64_t ld_src, int64_t rows, int cols, void* dst,
                               int64_t ld_dst, void* stream) {
  GN_CHECK_ARG(rows >= 0 && cols >= 1 && ld_src >= cols && ld_dst >= rows,
               "gn_transpose_16: bad sizes");
  if (rows == 0) return GN_OK;
  GN_CHECK_ARG(src && dst, "gn_transpose_16: null pointer");
  dim3 grid((unsigned)ceil_div(rows, 64), (unsigned)ceil_div(cols, 64));
  transpose16_kernel<<<grid, 256, 0, as_stream(stream)>>>(
      reinterpret_cast<const uint16_t*>(src), ld_src, rows, cols, reinterpret_cast<uint16_t*>(dst),
      ld_dst);
  GN_LAUNCH_CHECK();
  return GN_OK;
}

extern "C" int gn_colsum_accumulate(const void* X, int x_dtype, int64_t ldx, int64_t M, int N,
                                    float* out, void* stream) {
  GN_CHECK_ARG(M >= 0 && N >= 1 && ldx >= N, "gn_colsum_accumulate: bad sizes");
  if (M == 0) return GN_OK;
  GN_CHECK_ARG(X && out, "gn_colsum_accumulate: null pointer");
  dim3 grid((unsigned)ceil_div(M, 256), (unsigned)ceil_div(N, 64));
  cudaStream_t st = as_stream(stream);
  if (x_dtype == GN_F32)
    colsum_kernel<float><<<grid, 256, 0, st>>>(reinterpret_cast<const float*>(X), ldx, M, N, out);
  else if (x_dtype == GN_BF16)
    colsum_kernel<__nv_bfloat16><<<grid, 256, 0, st>>>(reinterpret_cast<const __nv_bfloat16*>(X),
                                                       ldx, M, N, out);
  else
    GN_CHECK_ARG(false, "gn_colsum_accumulate: bad x_dtype %d", x_dtype);
  GN_LAUNCH_CHECK();
  return GN_OK;
}

extern "C" int gn_relu_backward(const void* dY, int dy_dtype, int64_t ld_dy, const void* Y,
                                int y_dtype, int64_t ld_y, void* out, int64_t ld_out, int64_t M,
                                int N, void* stream) {
  GN_CHECK_ARG(M >= 0 && N >= 1 && ld_dy >= N && ld_out >= N && (Y == nullptr || ld_y >= N),
               "gn_relu_backward: bad sizes");
  if (M == 0) return GN_OK;
  GN_CHECK_ARG(dY && out, "gn_relu_backward: null pointer");
  GN_CHECK_ARG(dy_dtype == GN_F32 || dy_dtype == GN_BF16, "gn_relu_backward: bad dy_dtype");
  GN_CHECK_ARG(Y == nullptr || y_dtype == GN_F32 || y_dtype == GN_BF16,
               "gn_relu_backward: bad y_dtype");
  const int64_t work = M * ((N + 3) / 4);
  const unsigned grid = (unsigned)ceil_div(work, 256);
  cudaStream_t st = as_stream(stream);
  __nv_bfloat16* o = reinterpret_cast<__nv_bfloat16*>(out);
#define GN_MC(TG, TY)                                                                        \
  mask_convert_kernel<TG, TY, __nv_bfloat16><<<grid, 256, 0, st>>>(                          \
      reinterpret_cast<const TG*>(dY), ld_dy, reinterpret_cast<const TY*>(Y), ld_y, o, ld_out, M, N)
  const bool yf = (Y == nullptr) || y_dtype == GN_F32;
  if (dy_dtype == GN_F32) {
    if (yf) GN_MC(float, float); else GN_MC(float, __nv_bfloat16);
  } else {
    if (yf) GN_MC(__nv_bfloat16, float); else GN_MC(__nv_bfloat16, __nv_bfloat16);
  }
#undef GN_MC
  GN_LAUNCH_CHECK();
  return GN_OK;
}

static int check_ell(const gn_graph* graph, const char* who) {
  if (!graph || !graph->col || graph->ell_stride <= 0) {
    set_error("%s: the graph must be in ELL layout (slot table)", who);
    return GN_ERR_INVALID_ARGUMENT;
  }
  return GN_OK;
}

extern "C" int gn_edgeconv_pack_tc(const float* W1, const float* b1, const float* W2,
                                   const float* b2, int d, int a, int b, int Kp, int Kd,
                                   void* Wpq, int wpq_dtype, float* bias_pq, void* W2p, void* W2b,
                                   void* W2bT, void* WpqT, int act, void* stream) {
  GN_CHECK_ARG(act == GN_ACT_RELU || act == GN_ACT_LEAKY_RELU || act == GN_ACT_GELU,
               "gn_edgeconv_pack_tc: activation %d (ReLU, LeakyReLU or GELU)", act);
  GN_CHECK_ARG(d >= 1 && a >= 1 && b >= 1 && Kp >= a + 2 && Kd >= d, "gn_edgeconv_pack_tc: bad sizes");
  GN_CHECK_ARG(W1 && W2 && Wpq && bias_pq && W2p, "gn_edgeconv_pack_tc: null pointer");
  GN_CHECK_ARG(wpq_dtype == GN_F32 || wpq_dtype == GN_BF16 || wpq_dtype == GN_F16,
               "gn_edgeconv_pack_tc: unknown dtype");
  GN_CHECK_ARG((W2b == nullptr) == (W2bT == nullptr), "gn_edgeconv_pack_tc: W2b and W2bT go together");
  GN_CHECK_ARG(WpqT == nullptr || wpq_dtype != GN_F32, "gn_edgeconv_pack_tc: WpqT is 16-bit only");
  PackParams p;
  p.W1 = W1; p.b1 = b1; p.W2 = W2; p.b2 = b2;
  p.d = d; p.a = a; p.b = b; p.Kp = Kp; p.Kd = Kd; p.b16 = (b + 15) & ~15;
  p.Wpq = Wpq; p.wpq_dtype = wpq_dtype; p.bias = bias_pq;
  p.W2p = reinterpret_cast<__half*>(W2p);
  p.W2b = reinterpret_cast<__nv_bfloat16*>(W2b);
  p.W2bT = reinterpret_cast<__nv_bfloat16*>(W2bT);
  p.WpqT = WpqT;
  // what the edge kernel's h builder makes of the "ones" columns: act(1.0), rounded to fp16
  p.ones_act = act == GN_ACT_GELU
                   ? __half2float(__float2half_rn(0.5f * (1.0f + erff(0.70710678118654752440f))))
                   : 1.0f;
  const int64_t total = (int64_t)2 * Kp * Kd + (int64_t)p.b16 * Kp + 2 * Kp;
  edgeconv_pack_tc_kernel<<<(unsigned)ceil_div(total, 256), 256, 0, as_stream(stream)>>>(p);
  GN_LAUNCH_CHECK();
  return GN_OK;
}

extern "C" int gn_edge_build_h(const void* P, int64_t ldp, const void* Q, int64_t ldq, int Kp,
                               const gn_graph* graph, int64_t N, void* H, int64_t ldh, void* HT,
                               int64_t ldht, void* stream) {
  int rc = check_ell(graph, "gn_edge_build_h");
  if (rc != GN_OK) return rc;
  const int KP = graph->ell_stride;
  const int64_t E = N * KP;
  GN_CHECK_ARG(E < INT32_MAX, "gn_edge_build_h: N * stride must fit int32");
  GN_CHECK_ARG(Kp >= 8 && Kp % 8 == 0 && ldp >= Kp && ldq >= Kp && ldp % 8 == 0 && ldq % 8 == 0 &&
                   (H == nullptr || (ldh >= Kp && ldh % 8 == 0)) && (HT == nullptr || ldht >= E),
               "gn_edge_build_h: Kp and the row strides must be multiples of 8");
  if (N == 0) return GN_OK;
  GN_CHECK_ARG(P && Q && (H || HT), "gn_edge_build_h: null pointer");
  GN_CHECK_ARG(((reinterpret_cast<uintptr_t>(P) | reinterpret_cast<uintptr_t>(Q) |
                 reinterpret_cast<uintptr_t>(H)) & 15) == 0,
               "gn_edge_build_h: P, Q, H must be 16-byte aligned");
  const size_t smem = (size_t)EB_ROWS * tile_width(Kp) * 2;
  GN_CHECK_ARG(smem <= 200 * 1024, "gn_edge_build_h: Kp=%d too wide", Kp);
  GN_CUDA(cudaFuncSetAttribute(edge_build_h_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                               (int)smem));
  edge_build_h_kernel<<<(unsigned)ceil_div(E, EB_ROWS), 256, smem, as_stream(stream)>>>(
      reinterpret_cast<const __half*>(P), ldp, reinterpret_cast<const __half*>(Q), ldq, Kp,
      graph->col, KP, E, reinterpret_cast<__nv_bfloat16*>(H), ldh,
      reinterpret_cast<__nv_bfloat16*>(HT), ldht);
  GN_LAUNCH_CHECK();
  return GN_OK;
}

extern "C" int gn_edge_grad_z(const void* Z, int64_t ldz, const float* dOut, int64_t ldd, int b,
                              const gn_graph* graph, int64_t N, void* G2, int64_t ldg, void* G2T,
                              int64_t ldgt, void* stream) {
  int rc = check_ell(graph, "gn_edge_grad_z");
  if (rc != GN_OK) return rc;
  const int KP = graph->ell_stride;
  const int64_t E = N * KP;
  const int bw = (b + 7) & ~7;
  GN_CHECK_ARG(E < INT32_MAX, "gn_edge_grad_z: N * stride must fit int32");
  GN_CHECK_ARG(b >= 2 && b % 2 == 0 && b <= 1024 && ldz >= bw && ldd >= bw && ldg >= bw &&
                   ldgt >= E && ldz % 8 == 0 && ldd % 4 == 0 && ldg % 8 == 0,
               "gn_edge_grad_z: rows must hold round_up(b, 8) columns; strides multiples of 8");
  if (N == 0) return GN_OK;
  GN_CHECK_ARG(Z && dOut && G2 && G2T, "gn_edge_grad_z: null pointer");
  GN_CHECK_ARG(((reinterpret_cast<uintptr_t>(dOut) | reinterpret_cast<uintptr_t>(Z) |
                 reinterpret_cast<uintptr_t>(G2)) & 15) == 0,
               "gn_edge_grad_z: Z, dOut, G2 must be 16-byte aligned");
  const size_t smem = (size_t)EB_ROWS * tile_width(b) * 2;
  GN_CHECK_ARG(smem <= 200 * 1024, "gn_edge_grad_z: b=%d too wide", b);
  GN_CUDA(cudaFuncSetAttribute(edge_grad_z_kernel<false>,
                               cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  edge_grad_z_kernel<false><<<(unsigned)ceil_div(E, EB_ROWS), 256, smem, as_stream(stream)>>>(
      reinterpret_cast<const __nv_bfloat16*>(Z), ldz, nullptr, 0, 1, dOut, ldd, b, graph->col, KP,
      E, reinterpret_cast<__nv_bfloat16*>(G2), ldg, reinterpret_cast<__nv_bfloat16*>(G2T), ldgt);
  GN_LAUNCH_CHECK();
  return GN_OK;
}

extern "C" int gn_edge_grad_z_mask(const uint32_t* relu_mask, int64_t mask_ld, const float* dOut,
                                   int64_t ldd, int b, const gn_graph* graph, int64_t N, void* G2,
                                   int64_t ldg, void* G2T, int64_t ldgt, void* stream) {
  int rc = check_ell(graph, "gn_edge_grad_z_mask");
  if (rc != GN_OK) return rc;
  const int KP = graph->ell_stride;
  const int64_t E = N * KP;
  GN_CHECK_ARG(KP <= 64, "gn_edge_grad_z_mask: %d slots per node (max 64)", KP);
  const int used = (64 / KP) * KP;
  const int bw = (b + 7) & ~7;
  GN_CHECK_ARG(E < INT32_MAX, "gn_edge_grad_z_mask: N * stride must fit int32");
  GN_CHECK_ARG(b >= 2 && b % 2 == 0 && b <= 256 && ldd >= bw && ldg >= bw && ldgt >= E &&
                   ldd % 4 == 0 && ldg % 8 == 0 && mask_ld >= ceil_div(E, (int64_t)used) * 64,
               "gn_edge_grad_z_mask: rows must hold round_up(b, 8) columns; strides multiples of 8");
  if (N == 0) return GN_OK;
  GN_CHECK_ARG(relu_mask && dOut && G2 && G2T, "gn_edge_grad_z_mask: null pointer");
  GN_CHECK_ARG(((reinterpret_cast<uintptr_t>(dOut) | reinterpret_cast<uintptr_t>(G2)) & 15) == 0,
               "gn_edge_grad_z_mask: dOut, G2 must be 16-byte aligned");
  const size_t smem = (size_t)EB_ROWS * tile_width(b) * 2;
  GN_CUDA(cudaFuncSetAttribute(edge_grad_z_kernel<true>,
                               cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  edge_grad_z_kernel<true><<<(unsigned)ceil_div(E, EB_ROWS), 256, smem, as_stream(stream)>>>(
      nullptr, 0, relu_mask, mask_ld, used, dOut, ldd, b, graph->col, KP, E,
      reinterpret_cast<__nv_bfloat16*>(G2), ldg, reinterpret_cast<__nv_bfloat16*>(G2T), ldgt);
  GN_LAUNCH_CHECK();
  return GN_OK;
}

extern "C" int gn_ell_reverse(const gn_graph* graph, const int32_t* ptr, int64_t N, int64_t B,
                              int32_t* rstart, int32_t* rcount, int32_t* rslot, void* stream) {
  int rc = check_ell(graph, "gn_ell_reverse");
  if (rc != GN_OK) return rc;
  GN_CHECK_ARG(N >= 0 && B >= 0 && N * (int64_t)graph->ell_stride < INT32_MAX,
               "gn_ell_reverse: bad sizes");
  if (N == 0 || B == 0) return GN_OK;
  GN_CHECK_ARG(ptr && rstart && rcount && rslot, "gn_ell_reverse: null pointer");
  cudaStream_t st = as_stream(stream);
  GN_CUDA(cudaMemsetAsync(rcount, 0, (size_t)N * sizeof(int32_t), st));
  ell_reverse_kernel<<<(unsigned)ceil_div(B, 4), 128, 0, st>>>(graph->col, graph->ell_stride, ptr,
                                                               B, rstart, rcount, rslot);
  GN_LAUNCH_CHECK();
  return GN_OK;
}

extern "C" int gn_edge_grad_pq_gather(const void* dH, int64_t lddh, const void* P, int64_t ldp,
                                      const void* Q, int64_t ldq, int Kp, const gn_graph* graph,
                                      const int32_t* ptr, int64_t N, int64_t B,
                                      const int32_t* rstart, const int32_t* rcount,
                                      const int32_t* rslot, void* dP, int64_t lddp, void* dQ,
                                      int64_t lddq, float* dP32, int64_t ldp32, float* dQ32,
                                      int64_t ldq32, void* stream) {
  int rc = check_ell(graph, "gn_edge_grad_pq_gather");
  if (rc != GN_OK) return rc;
  GN_CHECK_ARG(Kp >= 8 && Kp % 8 == 0 && lddh >= Kp && ldp >= Kp && ldq >= Kp && lddp >= Kp &&
                   lddq >= Kp && lddh % 8 == 0 && ldp % 8 == 0 && ldq % 8 == 0 && lddp % 8 == 0 &&
                   lddq % 8 == 0,
               "gn_edge_grad_pq_gather: widths and row strides must be multiples of 8");
  if (N == 0 || B == 0) return GN_OK;
  GN_CHECK_ARG(dH && P && Q && ptr && dP && dQ, "gn_edge_grad_pq_gather: null pointer");
  GN_CHECK_ARG(rstart == nullptr || (rcount && rslot),
               "gn_edge_grad_pq_gather: the list form needs all three arrays of gn_ell_reverse");
  GN_CHECK_ARG(((reinterpret_cast<uintptr_t>(dH) | reinterpret_cast<uintptr_t>(P) |
                 reinterpret_cast<uintptr_t>(Q) | reinterpret_cast<uintptr_t>(dP) |
                 reinterpret_cast<uintptr_t>(dQ)) & 15) == 0,
               "gn_edge_grad_pq_gather: 16-byte alignment");
  GN_CHECK_ARG((dP32 == nullptr || (ldp32 >= Kp && ldp32 % 4 == 0 &&
                                    (reinterpret_cast<uintptr_t>(dP32) & 15) == 0)) &&
                   (dQ32 == nullptr || (ldq32 >= Kp && ldq32 % 4 == 0 &&
                                        (reinterpret_cast<uintptr_t>(dQ32) & 15) == 0)),
               "gn_edge_grad_pq_gather: bad fp32 outputs");
  GN_CHECK_ARG(B <= 2147483647LL, "gn_edge_grad_pq_gather: too many events");
  static const int gch_env = getenv("GN_PQ_CHUNKS") ? atoi(getenv("GN_PQ_CHUNKS")) : 0;
  const bool use_lists = rstart != nullptr;
  const int pieces = Kp / 8;
  if (!use_lists) {
    GN_CHECK_ARG(dQ32 != nullptr,
                 "gn_edge_grad_pq_gather: the one-pass form accumulates in dQ32 (fp32), pass it");
    cudaStream_t st = as_stream(stream);
    GN_CUDA(cudaMemset2DAsync(dQ32, (size_t)ldq32 * sizeof(float), 0, (size_t)Kp * sizeof(float),
                              (size_t)N, st));
    const int64_t items = N * pieces;
    edge_grad_pq_scatter_kernel<<<(unsigned)ceil_div(items, 256), 256, 0, st>>>(
        reinterpret_cast<const __nv_bfloat16*>(dH), lddh, reinterpret_cast<const __half*>(P), ldp,
        reinterpret_cast<const __half*>(Q), ldq, Kp, graph->col, graph->ell_stride, N,
        reinterpret_cast<__nv_bfloat16*>(dP), lddp, dP32, ldp32, dQ32, ldq32);
    GN_LAUNCH_CHECK();
    const int64_t work = N * (int64_t)((Kp + 3) / 4);
    mask_convert_kernel<float, float, __nv_bfloat16><<<(unsigned)ceil_div(work, 256), 256, 0, st>>>(
        dQ32, ldq32, nullptr, 0, reinterpret_cast<__nv_bfloat16*>(dQ), lddq, N, Kp);
    GN_LAUNCH_CHECK();
    return GN_OK;
  }
  const int gch = gch_env > 0 ? gch_env : 4;  // 64-byte row pieces: the event's block stays in L2
  dim3 grid((unsigned)B, (unsigned)ceil_div(pieces, gch));
  edge_grad_pq_gather_kernel<<<grid, 256, 0, as_stream(stream)>>>(
      reinterpret_cast<const __nv_bfloat16*>(dH), lddh, reinterpret_cast<const __half*>(P), ldp,
      reinterpret_cast<const __half*>(Q), ldq, Kp, graph->col, graph->ell_stride, ptr, rstart,
      rcount, rslot, gch, reinterpret_cast<__nv_bfloat16*>(dP), lddp,
      reinterpret_cast<__nv_bfloat16*>(dQ), lddq, dP32, ldp32, dQ32, ldq32);
  GN_LAUNCH_CHECK();
  return GN_OK;
}

extern "C" int gn_edge_grad_pq(const void* dH, int64_t lddh, const void* H, int64_t ldh, int Kp,
                               const gn_graph* graph, int64_t N, void* dP, int64_t ldp, float* dQ,
                               int64_t ldq, float* dP32, int64_t ldp32, void* stream) {
  int rc = check_ell(graph, "gn_edge_grad_pq");
  if (rc != GN_OK) return rc;
  GN_CHECK_ARG(Kp >= 2 && Kp % 2 == 0 && lddh >= Kp && ldh >= Kp && ldp >= Kp && ldq >= Kp,
               "gn_edge_grad_pq: bad sizes");
  if (N == 0) return GN_OK;
  GN_CHECK_ARG(dH && H && dP && dQ, "gn_edge_grad_pq: null pointer");
  GN_CHECK_ARG(dP32 == nullptr || (ldp32 >= Kp && ldp32 % 2 == 0 &&
                                   (reinterpret_cast<uintptr_t>(dP32) & 7) == 0),
               "gn_edge_grad_pq: bad dP32");
  edge_grad_pq_kernel<<<(unsigned)ceil_div(N, 8), 256, 0, as_stream(stream)>>>(
      reinterpret_cast<const __nv_bfloat16*>(dH), lddh, reinterpret_cast<const __nv_bfloat16*>(H),
      ldh, Kp, graph->col, graph->ell_stride, N, reinterpret_cast<__nv_bfloat16*>(dP), ldp, dQ,
      ldq, dP32, ldp32);
  GN_LAUNCH_CHECK();
  return GN_OK;
}

extern "C" int gn_segment_pool_backward(const void* x, int x_dtype, int64_t ldx, int C,
                                        const int32_t* ptr, int64_t B, const int32_t* schemes,
                                        int n_schemes, const float* pooled, int64_t ld_pooled,
                                        const float* d_pooled, int64_t ld_dpooled, float* dX,
                                        int64_t ld_dx, void* stream) {
  GN_CHECK_ARG(C >= 1 && n_schemes >= 1 && n_schemes <= 8, "gn_segment_pool_backward: bad sizes");
  GN_CHECK_ARG(ld_pooled >= (int64_t)C * n_schemes && ld_dpooled >= (int64_t)C * n_schemes &&
                   ld_dx >= C && ldx >= C,
               "gn_segment_pool_backward: leading dimension too small");
  if (B == 0) return GN_OK;
  GN_CHECK_ARG(x && ptr && schemes && pooled && d_pooled && dX,
               "gn_segment_pool_backward: null pointer");
  PoolSchemesB sch;
  sch.n = n_schemes;
  for (int k = 0; k < n_schemes; ++k) {
    GN_CHECK_ARG(schemes[k] >= GN_POOL_MIN && schemes[k] <= GN_POOL_MEAN,
                 "gn_segment_pool_backward: unknown scheme %d", schemes[k]);
    sch.s[k] = schemes[k];
  }
  dim3 grid((unsigned)B, (unsigned)ceil_div(C, 128));
  cudaStream_t st = as_stream(stream);
  if (x_dtype == GN_F32)
    segment_pool_backward_kernel<float><<<grid, 128, 0, st>>>(
        reinterpret_cast<const float*>(x), ldx, C, ptr, sch, pooled, ld_pooled, d_pooled,
        ld_dpooled, dX, ld_dx);
  else if (x_dtype == GN_BF16)
    segment_pool_backward_kernel<__nv_bfloat16><<<grid, 128, 0, st>>>(
        reinterpret_cast<const __nv_bfloat16*>(x), ldx, C, ptr, sch, pooled, ld_pooled, d_pooled,
        ld_dpooled, dX, ld_dx);
  else
    GN_CHECK_ARG(false, "gn_segment_pool_backward: bad x_dtype %d", x_dtype);
  GN_LAUNCH_CHECK();
  return GN_OK;
}

extern "C" int gn_csr_to_ell(const int32_t* rowptr, const int32_t* col, int64_t N, int stride,
                             int32_t* nbr, int32_t* deg, int32_t* flags, void* stream) {
  GN_CHECK_ARG(N >= 0 && stride >= 1, "gn_csr_to_ell: bad sizes");
  if (N == 0) return GN_OK;
  GN_CHECK_ARG(rowptr && col && nbr && deg, "gn_csr_to_ell: null pointer");
  csr_to_ell_kernel<<<(unsigned)ceil_div(N * stride, 256), 256, 0, as_stream(stream)>>>(
      rowptr, col, N, stride, nbr, deg, flags);
  GN_LAUNCH_CHECK();
  return GN_OK;
}

extern "C" int gn_outer_accumulate_f32(const float* G, int64_t ldg, int cols, const float* X,
                                       int64_t ldx, int K, int64_t N, float* out, int64_t ldo,
                                       void* stream) {
  GN_CHECK_ARG(N >= 0 && cols >= 1 && K >= 1 && K <= 64 && ldg >= cols && ldx >= K && ldo >= K,
               "gn_outer_accumulate_f32: bad sizes (K must be <= 64)");
  if (N == 0) return GN_OK;
  GN_CHECK_ARG(G && X && out, "gn_outer_accumulate_f32: null pointer");
  const unsigned grid = (unsigned)ceil_div(N, 128);
  cudaStream_t st = as_stream(stream);
  if (K <= 8) outer_accumulate_kernel<8><<<grid, 256, 0, st>>>(G, ldg, cols, X, ldx, K, N, out, ldo);
  else if (K <= 20) outer_accumulate_kernel<20><<<grid, 256, 0, st>>>(G, ldg, cols, X, ldx, K, N, out, ldo);
  else if (K <= 32) outer_accumulate_kernel<32><<<grid, 256, 0, st>>>(G, ldg, cols, X, ldx, K, N, out, ldo);
  else if (K <= 40) outer_accumulate_kernel<40><<<grid, 256, 0, st>>>(G, ldg, cols, X, ldx, K, N, out, ldo);
  else outer_accumulate_kernel<64><<<grid, 256, 0, st>>>(G, ldg, cols, X, ldx, K, N, out, ldo);
  GN_LAUNCH_CHECK();
  return GN_OK;
}

extern "C" int gn_zero(void* p, size_t bytes, void* stream) {
  if (bytes == 0) return GN_OK;
  GN_CHECK_ARG(p != nullptr, "gn_zero: null pointer");
  GN_CUDA(cudaMemsetAsync(p, 0, bytes, as_stream(stream)));
  return GN_OK;
}

extern "C" int gn_convert_scale_16(const float* X, int64_t ldx, void* out, int out_dtype,
                                   int64_t ldo, int64_t M, int N, float scale, void* stream) {
  GN_CHECK_ARG(M >= 0 && N >= 1 && ldx >= N && ldo >= N, "gn_convert_scale_16: bad sizes");
  if (M == 0) return GN_OK;
  GN_CHECK_ARG(X && out, "gn_convert_scale_16: null pointer");
  const unsigned grid = (unsigned)ceil_div(M * ((N + 3) / 4), 256);
  cudaStream_t st = as_stream(stream);
  if (out_dtype == GN_F16)
    convert_scale_kernel<__half><<<grid, 256, 0, st>>>(X, ldx, reinterpret_cast<__half*>(out), ldo,
                                                       M, N, scale);
  else if (out_dtype == GN_BF16)
    convert_scale_kernel<__nv_bfloat16><<<grid, 256, 0, st>>>(
        X, ldx, reinterpret_cast<__nv_bfloat16*>(out), ldo, M, N, scale);
  else
    GN_CHECK_ARG(false, "gn_convert_scale_16: bad out_dtype %d", out_dtype);
  GN_LAUNCH_CHECK();
  return GN_OK;
}
