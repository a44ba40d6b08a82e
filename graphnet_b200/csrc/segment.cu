// Per-event (segmented) reductions: global variables + homophily + "distribute",
// and global min/max/sum/mean pooling.
//
// Replaces torch_scatter.scatter_{min,max,sum,mean} and
// torch_geometric.utils.homophily as reached from the reference at
// models/gnn/dynedge.py:252-265, :267-294, :309-320 and models/utils.py:17-33.
// Both kernels are HBM-bound streaming reductions: rows of one event are
// contiguous, threads map to feature columns (coalesced, 128-bit where aligned),
// every input element is read exactly once and all requested statistics are
// produced in the same pass.
#include <cuda_fp16.h>
#include <math.h>

#include <algorithm>
#include <atomic>

#include "common.cuh"

namespace gn {

// --------------------------------------------------------- global variables
constexpr int GV_THREADS = 128;
constexpr int GV_MAX_THREADS = 512;
constexpr int GV_MAX_F = 64;

struct X0Out {
  void* ptr;      // [N, ld] or null
  int64_t ld;
  int cols;       // columns written: [x | gv[batch] | zeros]
  int dtype;      // GN_F32 / GN_BF16 / GN_F16
};

__device__ __forceinline__ void store_x0(const X0Out& o, int64_t row, int c, float v) {
  if (o.dtype == GN_F32) reinterpret_cast<float*>(o.ptr)[row * o.ld + c] = v;
  else if (o.dtype == GN_BF16X3) {
    // v = hi + lo to 16 mantissa bits; the row is [hi | lo | hi], each `cols` wide: against
    // weights laid out as [W_hi | W_hi | W_lo] one bf16 GEMM gives hi W_hi + lo W_hi + hi W_lo
    __nv_bfloat16* p = reinterpret_cast<__nv_bfloat16*>(o.ptr) + row * o.ld + c;
    const __nv_bfloat16 hi = __float2bfloat16_rn(v);
    p[0] = hi;
    p[o.cols] = __float2bfloat16_rn(v - __bfloat162float(hi));
    p[2 * o.cols] = hi;
  } else if (o.dtype == GN_BF16)
    reinterpret_cast<__nv_bfloat16*>(o.ptr)[row * o.ld + c] = __float2bfloat16_rn(v);
  else
    reinterpret_cast<__half*>(o.ptr)[row * o.ld + c] =
        __float2half_rn(fminf(fmaxf(v, -65504.f), 65504.f));
}

__global__ void __launch_bounds__(GV_MAX_THREADS) global_variables_kernel(
    const float* __restrict__ x, int64_t ldx, int F, GraphView g, const int32_t* __restrict__ ptr,
    const int32_t* __restrict__ n_pulses, float* __restrict__ gv, X0Out o0, X0Out o1) {
  __shared__ float s_part[GV_MAX_THREADS];
  const int T = blockDim.x;  // 128, or 512 for large events (chosen by the host from N / B)
  __shared__ float s_gv[GV_MAX_F + 5];
  __shared__ int s_cnt[5];
  const int e = blockIdx.x;
  const int lo = ptr[e];
  const int n = ptr[e + 1] - lo;
  const int tid = threadIdx.x;
  const int G = F + 5;
  if (tid < 5) s_cnt[tid] = 0;

  // column sums: thread (r, c) walks rows r, r + rpp, ...
  const int rpp = T / F;
  const int r0 = tid / F, c0 = tid % F;
  float s = 0.f;
  if (r0 < rpp)
    for (int r = r0; r < n; r += rpp) s += x[(int64_t)(lo + r) * ldx + c0];
  s_part[tid] = s;
  __syncthreads();
  if (tid < F) {
    float tot = 0.f;
    for (int r = 0; r < rpp; ++r) tot += s_part[r * F + tid];
    float cnt = (float)max(n, 1);
    s_gv[tid] = tot / cnt;
  }

  // homophily of columns 0..3 over the in-edges of this event's nodes
  int eq[4] = {0, 0, 0, 0};
  int ne = 0;
  for (int r = tid; r < n; r += T) {
    const int64_t i = lo + r;
    int64_t start;
    int cnt;
    graph_row(g, i, start, cnt);
    const float xi0 = x[i * ldx + 0], xi1 = x[i * ldx + 1], xi2 = x[i * ldx + 2],
                xi3 = x[i * ldx + 3];
    for (int t = 0; t < cnt; ++t) {
      const int64_t j = g.col[start + t];
      eq[0] += (x[j * ldx + 0] == xi0);
      eq[1] += (x[j * ldx + 1] == xi1);
      eq[2] += (x[j * ldx + 2] == xi2);
      eq[3] += (x[j * ldx + 3] == xi3);
    }
    ne += cnt;
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
#pragma unroll
    for (int c = 0; c < 4; ++c) eq[c] += __shfl_xor_sync(0xffffffffu, eq[c], o);
    ne += __shfl_xor_sync(0xffffffffu, ne, o);
  }
  if ((tid & 31) == 0) {
#pragma unroll
    for (int c = 0; c < 4; ++c) atomicAdd(&s_cnt[c], eq[c]);
    atomicAdd(&s_cnt[4], ne);
  }
  __syncthreads();
  if (tid < 4) {
    const int tot = s_cnt[4];
    s_gv[F + tid] = tot > 0 ? (float)s_cnt[tid] / (float)tot : 0.f;
  }
  if (tid == 4) s_gv[F + 4] = log10f((float)n_pulses[e]);
  __syncthreads();
  for (int c = tid; c < G; c += T) gv[(int64_t)e * G + c] = s_gv[c];

#pragma unroll
  for (int w = 0; w < 2; ++w) {
    const X0Out& o = w == 0 ? o0 : o1;
    if (o.ptr == nullptr) continue;
    // (row, column) advance incrementally: no division per element
    int r = tid / o.cols, c = tid % o.cols;
    const int dr = T / o.cols, dc = T % o.cols;
    while (r < n) {
      float v = 0.f;
      if (c < F) v = x[(int64_t)(lo + r) * ldx + c];
      else if (c < F + G) v = s_gv[c - F];
      store_x0(o, lo + r, c, v);
      r += dr;
      c += dc;
      if (c >= o.cols) { c -= o.cols; ++r; }
    }
  }
}

// ------------------------------------------------------------------ pooling
constexpr int POOL_COLS = 32;   // column-threads per CTA (x4 columns each when vectorised)
constexpr int POOL_ROWG = 4;    // row groups per CTA

struct PoolSchemes {
  int n;
  int s[8];
};

template <typename T, int VEC>
__device__ __forceinline__ void load_vec(const T* p, float (&v)[VEC]);
template <>
__device__ __forceinline__ void load_vec<float, 4>(const float* p, float (&v)[4]) {
  float4 t = *reinterpret_cast<const float4*>(p);
  v[0] = t.x; v[1] = t.y; v[2] = t.z; v[3] = t.w;
}
template <>
__device__ __forceinline__ void load_vec<float, 1>(const float* p, float (&v)[1]) { v[0] = *p; }
template <>
__device__ __forceinline__ void load_vec<__nv_bfloat16, 4>(const __nv_bfloat16* p, float (&v)[4]) {
  uint2 t = *reinterpret_cast<const uint2*>(p);
  v[0] = __uint_as_float(t.x << 16); v[1] = __uint_as_float(t.x & 0xffff0000u);
  v[2] = __uint_as_float(t.y << 16); v[3] = __uint_as_float(t.y & 0xffff0000u);
}
template <>
__device__ __forceinline__ void load_vec<__nv_bfloat16, 1>(const __nv_bfloat16* p, float (&v)[1]) {
  v[0] = __bfloat162float(*p);
}

template <typename T, int VEC>
__global__ void __launch_bounds__(POOL_COLS * POOL_ROWG) segment_pool_kernel(
    const T* __restrict__ x, int64_t ldx, int C, const int32_t* __restrict__ ptr,
    PoolSchemes sch, float* __restrict__ out, int64_t ldo) {
  __shared__ float s_min[POOL_ROWG][POOL_COLS * VEC];
  __shared__ float s_max[POOL_ROWG][POOL_COLS * VEC];
  __shared__ float s_sum[POOL_ROWG][POOL_COLS * VEC];
  const int e = blockIdx.x;
  const int lo = ptr[e];
  const int n = ptr[e + 1] - lo;
  const int ct = threadIdx.x % POOL_COLS;
  const int rg = threadIdx.x / POOL_COLS;
  const int c = (blockIdx.y * POOL_COLS + ct) * VEC;
  const bool col_ok = c < C;
  float mn[VEC], mx[VEC], sm[VEC];
#pragma unroll
  for (int v = 0; v < VEC; ++v) {
    mn[v] = __int_as_float(0x7f800000);
    mx[v] = __int_as_float(0xff800000);
    sm[v] = 0.f;
  }
  if (col_ok) {
    const T* base = x + (int64_t)lo * ldx + c;
    int r = rg;
    // sixteen rows of this row group in flight per pass: an event of up to 64 pulses is read
    // in ONE round trip to HBM (eight per pass needed two dependent ones)
    constexpr int UNR = VEC == 4 ? 16 : 8;
    for (; r + (UNR - 1) * POOL_ROWG < n; r += UNR * POOL_ROWG) {
      float a[UNR][VEC];
#pragma unroll
      for (int u = 0; u < UNR; ++u) load_vec<T, VEC>(base + (int64_t)(r + u * POOL_ROWG) * ldx, a[u]);
#pragma unroll
      for (int u = 0; u < UNR; ++u) {
#pragma unroll
        for (int v = 0; v < VEC; ++v) {
          mn[v] = fminf(mn[v], a[u][v]);
          mx[v] = fmaxf(mx[v], a[u][v]);
          sm[v] += a[u][v];
        }
      }
    }
    for (; r + 1 * POOL_ROWG < n; r += 2 * POOL_ROWG) {
      float a0[VEC], a1[VEC];
      load_vec<T, VEC>(base + (int64_t)r * ldx, a0);
      load_vec<T, VEC>(base + (int64_t)(r + POOL_ROWG) * ldx, a1);
#pragma unroll
      for (int v = 0; v < VEC; ++v) {
        mn[v] = fminf(mn[v], fminf(a0[v], a1[v]));
        mx[v] = fmaxf(mx[v], fmaxf(a0[v], a1[v]));
        sm[v] += a0[v]; sm[v] += a1[v];
      }
    }
    for (; r < n; r += POOL_ROWG) {
      float a0[VEC];
      load_vec<T, VEC>(base + (int64_t)r * ldx, a0);
#pragma unroll
      for (int v = 0; v < VEC; ++v) {
        mn[v] = fminf(mn[v], a0[v]);
        mx[v] = fmaxf(mx[v], a0[v]);
        sm[v] += a0[v];
      }
    }
  }
#pragma unroll
  for (int v = 0; v < VEC; ++v) {
    s_min[rg][ct * VEC + v] = mn[v];
    s_max[rg][ct * VEC + v] = mx[v];
    s_sum[rg][ct * VEC + v] = sm[v];
  }
  __syncthreads();
  if (rg == 0 && col_ok) {
    const float cnt = (float)max(n, 1);
#pragma unroll
    for (int v = 0; v < VEC; ++v) {
      float a = s_min[0][ct * VEC + v], b = s_max[0][ct * VEC + v], s = s_sum[0][ct * VEC + v];
#pragma unroll
      for (int g2 = 1; g2 < POOL_ROWG; ++g2) {
        a = fminf(a, s_min[g2][ct * VEC + v]);
        b = fmaxf(b, s_max[g2][ct * VEC + v]);
        s += s_sum[g2][ct * VEC + v];
      }
      if (n == 0) { a = 0.f; b = 0.f; }  // torch_scatter: empty segment -> 0
      if (c + v < C) {
        for (int k = 0; k < sch.n; ++k) {
          float r = sch.s[k] == GN_POOL_MIN ? a
                  : sch.s[k] == GN_POOL_MAX ? b
                  : sch.s[k] == GN_POOL_SUM ? s : s / cnt;
          out[(int64_t)e * ldo + (int64_t)k * C + c + v] = r;
        }
      }
    }
  }
}

// acc[4][B][C] (min bits / max bits / 64-bit fixed-point sum in units of 2^-32, as
// gn_tc_linear_pool leaves them) -> the pooled row
// of every event in the order of `sch`, plus (optionally) the scaled 16-bit copy the readout GEMM
// reads.  An event without rows pools to 0 (torch_scatter).
__global__ void pool_finalize_kernel(const float* __restrict__ acc, const int32_t* __restrict__ ptr,
                                     int64_t B, int C, PoolSchemes sch, float* __restrict__ out,
                                     int64_t ldo, __half* __restrict__ out16, int64_t ld16,
                                     float scale16) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= B * C) return;
  const int64_t e = i / C;
  const int c = (int)(i - e * C);
  const int n = ptr[e + 1] - ptr[e];
  const int64_t plane = B * (int64_t)C;
  float a = acc[i], b = acc[plane + i];
  const float s = (float)((double)reinterpret_cast<const long long*>(acc + 2 * plane)[i] *
                          (1.0 / 4294967296.0));
  if (n == 0) { a = 0.f; b = 0.f; }
  const float cnt = (float)max(n, 1);
  for (int k = 0; k < sch.n; ++k) {
    const float r = sch.s[k] == GN_POOL_MIN ? a
                  : sch.s[k] == GN_POOL_MAX ? b
                  : sch.s[k] == GN_POOL_SUM ? s : s / cnt;
    out[e * ldo + (int64_t)k * C + c] = r;
    if (out16 != nullptr)
      out16[e * ld16 + (int64_t)k * C + c] =
          __float2half_rn(fmaxf(fminf(r * scale16, 65504.f), -65504.f));
  }
}

}  // namespace gn

using namespace gn;

extern "C" int gn_pool_finalize(const float* acc, const int32_t* ptr, int64_t B, int C,
                                const int32_t* schemes, int num_schemes, float* out, int64_t ldo,
                                void* out16, int64_t ld16, float scale16, void* stream) {
  GN_CHECK_ARG(B >= 0 && C >= 1 && num_schemes >= 1 && num_schemes <= 8 && ldo >= (int64_t)C * num_schemes,
               "gn_pool_finalize: bad sizes");
  if (B == 0) return GN_OK;
  GN_CHECK_ARG(acc && ptr && schemes && out, "gn_pool_finalize: null pointer");
  GN_CHECK_ARG(out16 == nullptr || ld16 >= (int64_t)C * num_schemes, "gn_pool_finalize: bad ld16");
  PoolSchemes sch;
  sch.n = num_schemes;
  for (int k = 0; k < num_schemes; ++k) {
    GN_CHECK_ARG(schemes[k] >= 0 && schemes[k] <= 3, "gn_pool_finalize: unknown scheme %d", schemes[k]);
    sch.s[k] = schemes[k];
  }
  const int64_t total = B * (int64_t)C;
  pool_finalize_kernel<<<(unsigned)ceil_div(total, 256), 256, 0, as_stream(stream)>>>(
      acc, ptr, B, C, sch, out, ldo, reinterpret_cast<__half*>(out16), ld16, scale16);
  GN_LAUNCH_CHECK();
  return GN_OK;
}

extern "C" int gn_global_variables(const float* x, int64_t ldx, int F, const gn_graph* graph,
                                   const int32_t* ptr, const int32_t* n_pulses, int64_t N,
                                   int64_t B, float* gv, void* x0, int64_t ldx0, int x0_cols,
                                   int x0_dtype, void* x0b, int64_t ldx0b, int x0b_cols,
                                   int x0b_dtype, void* stream) {
  GN_CHECK_ARG(F >= 4 && F <= GV_MAX_F,
               "gn_global_variables: F=%d not in [4,%d] (columns 0..3 must be x,y,z,t)", F,
               GV_MAX_F);
  GN_CHECK_ARG(N >= 0 && B >= 0 && ldx >= F, "gn_global_variables: bad sizes");
  if (B == 0) return GN_OK;
  GN_CHECK_ARG(x && ptr && n_pulses && gv, "gn_global_variables: null pointer");
  GraphView g;
  int rc = make_graph_view(graph, &g);
  if (rc != GN_OK) return rc;
  X0Out o0{x0, ldx0, x0_cols, x0_dtype}, o1{x0b, ldx0b, x0b_cols, x0b_dtype};
  for (const X0Out* o : {&o0, &o1}) {
    if (!o->ptr) continue;
    GN_CHECK_ARG(o->cols >= F + F + 5 && o->ld >= (o->dtype == GN_BF16X3 ? 3 : 1) * o->cols,
                 "gn_global_variables: x0 needs >= 2F+5 columns (three times that for GN_BF16X3)");
    GN_CHECK_ARG(o->dtype == GN_F32 || o->dtype == GN_BF16 || o->dtype == GN_F16 ||
                     o->dtype == GN_BF16X3,
                 "gn_global_variables: bad x0 dtype");
  }
  const int threads = (N / (B > 0 ? B : 1)) > 128 ? GV_MAX_THREADS : GV_THREADS;
  global_variables_kernel<<<(unsigned)B, threads, 0, as_stream(stream)>>>(
      x, ldx, F, g, ptr, n_pulses, gv, o0, o1);
  GN_LAUNCH_CHECK();
  return GN_OK;
}

// fp32 [M, cols] -> bf16 pieces of x = x1 + x2 + x3 (x1 = bf16(x), x2 = bf16(x - x1), ...), laid
// out for ONE bf16 GEMM to evaluate x W^T with more mantissa bits than a bf16 operand has:
//   terms = 3: [x1 | x2 | x1]               against [w1 | w1 | w2]              (16 bits)
//   terms = 6: [x1 | x1 | x1 | x2 | x2 | x3] against [w1 | w2 | w3 | w1 | w2 | w1] (24 bits: every
//              product term down to 2^-16 of the leading one)
// Each block is kc columns wide, columns cols..kc-1 zero.
// TERMS = 9: both layouts from one read (out: 3 terms, out6: 6 terms)
template <int TERMS>
__global__ void __launch_bounds__(256) split_bf16_kernel(const float* __restrict__ X, int64_t ldx,
                                                         int64_t M, int cols, int kc,
                                                         __nv_bfloat16* __restrict__ out,
                                                         int64_t ldo,
                                                         __nv_bfloat16* __restrict__ out6 = nullptr,
                                                         int64_t ldo6 = 0) {
  const int q = kc >> 2;   // groups of 4 columns per row
  const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i >= M * q) return;
  const int64_t r = i / q;
  const int c = (int)(i - r * q) * 4;
  float v[4];
#pragma unroll
  for (int t = 0; t < 4; ++t) v[t] = c + t < cols ? X[r * ldx + c + t] : 0.f;
  __nv_bfloat16 x1[4], x2[4], x3[4];
#pragma unroll
  for (int t = 0; t < 4; ++t) {
    x1[t] = __float2bfloat16_rn(v[t]);
    const float r1 = v[t] - __bfloat162float(x1[t]);
    x2[t] = __float2bfloat16_rn(r1);
    x3[t] = __float2bfloat16_rn(r1 - __bfloat162float(x2[t]));
  }
  __nv_bfloat16* o = out + r * ldo + c;
  const uint2 u1 = *reinterpret_cast<const uint2*>(x1), u2 = *reinterpret_cast<const uint2*>(x2);
  if (TERMS == 3 || TERMS == 9) {
    *reinterpret_cast<uint2*>(o) = u1;
    *reinterpret_cast<uint2*>(o + kc) = u2;
    *reinterpret_cast<uint2*>(o + 2 * kc) = u1;
    if (TERMS == 9) o = out6 + r * ldo6 + c;
  }
  if (TERMS == 6 || TERMS == 9) {
    *reinterpret_cast<uint2*>(o) = u1;
    *reinterpret_cast<uint2*>(o + kc) = u1;
    *reinterpret_cast<uint2*>(o + 2 * kc) = u1;
    *reinterpret_cast<uint2*>(o + 3 * kc) = u2;
    *reinterpret_cast<uint2*>(o + 4 * kc) = u2;
    *reinterpret_cast<uint2*>(o + 5 * kc) = *reinterpret_cast<const uint2*>(x3);
  }
}

extern "C" int gn_split_bf16_both(const float* X, int64_t ldx, int64_t M, int cols, int kc,
                                  void* out3, int64_t ldo3, void* out6, int64_t ldo6,
                                  void* stream) {
  GN_CHECK_ARG(M >= 0 && cols >= 1 && kc >= cols && kc % 4 == 0 && ldx >= cols &&
                   ldo3 >= 3 * (int64_t)kc && ldo3 % 4 == 0 && ldo6 >= 6 * (int64_t)kc &&
                   ldo6 % 4 == 0,
               "gn_split_bf16_both: bad sizes (kc and the leading dimensions must be multiples of 4)");
  if (M == 0) return GN_OK;
  GN_CHECK_ARG(X && out3 && out6 && (reinterpret_cast<uintptr_t>(out3) & 7) == 0 &&
                   (reinterpret_cast<uintptr_t>(out6) & 7) == 0,
               "gn_split_bf16_both: null or misaligned pointer");
  const int64_t n = M * (kc / 4);
  split_bf16_kernel<9><<<(unsigned)ceil_div(n, 256), 256, 0, as_stream(stream)>>>(
      X, ldx, M, cols, kc, reinterpret_cast<__nv_bfloat16*>(out3), ldo3,
      reinterpret_cast<__nv_bfloat16*>(out6), ldo6);
  GN_LAUNCH_CHECK();
  return GN_OK;
}

extern "C" int gn_split_bf16(const float* X, int64_t ldx, int64_t M, int cols, int kc, int terms,
                             void* out, int64_t ldo, void* stream) {
  GN_CHECK_ARG(terms == 3 || terms == 6, "gn_split_bf16: terms must be 3 or 6");
  GN_CHECK_ARG(M >= 0 && cols >= 1 && kc >= cols && kc % 4 == 0 && ldx >= cols &&
                   ldo >= (int64_t)terms * kc && ldo % 4 == 0,
               "gn_split_bf16: bad sizes (kc and ldo must be multiples of 4)");
  if (M == 0) return GN_OK;
  GN_CHECK_ARG(X && out && (reinterpret_cast<uintptr_t>(out) & 7) == 0,
               "gn_split_bf16: null or misaligned pointer");
  const int64_t n = M * (kc / 4);
  const unsigned grid = (unsigned)ceil_div(n, 256);
  __nv_bfloat16* o = reinterpret_cast<__nv_bfloat16*>(out);
  if (terms == 3)
    split_bf16_kernel<3><<<grid, 256, 0, as_stream(stream)>>>(X, ldx, M, cols, kc, o, ldo);
  else
    split_bf16_kernel<6><<<grid, 256, 0, as_stream(stream)>>>(X, ldx, M, cols, kc, o, ldo);
  GN_LAUNCH_CHECK();
  return GN_OK;
}

extern "C" int gn_segment_pool(const void* x, int x_dtype, int64_t ldx, int C, const int32_t* ptr,
                               int64_t B, const int32_t* schemes, int n_schemes, float* out,
                               int64_t ldo, void* stream) {
  GN_CHECK_ARG(C >= 1 && n_schemes >= 1 && n_schemes <= 8, "gn_segment_pool: bad sizes");
  GN_CHECK_ARG(ldo >= (int64_t)C * n_schemes, "gn_segment_pool: ldo too small");
  if (B == 0) return GN_OK;
  GN_CHECK_ARG(x && ptr && schemes && out, "gn_segment_pool: null pointer");
  PoolSchemes sch;
  sch.n = n_schemes;
  for (int k = 0; k < n_schemes; ++k) {
    GN_CHECK_ARG(schemes[k] >= GN_POOL_MIN && schemes[k] <= GN_POOL_MEAN,
                 "gn_segment_pool: unknown scheme %d", schemes[k]);
    sch.s[k] = schemes[k];
  }
  cudaStream_t st = as_stream(stream);
  const int threads = POOL_COLS * POOL_ROWG;
  if (x_dtype == GN_F32) {
    const float* xf = reinterpret_cast<const float*>(x);
    bool vec = (C % 4 == 0) && (ldx % 4 == 0) && ((reinterpret_cast<uintptr_t>(x) & 15) == 0);
    if (vec) {
      dim3 grid((unsigned)B, (unsigned)ceil_div(C, POOL_COLS * 4));
      segment_pool_kernel<float, 4><<<grid, threads, 0, st>>>(xf, ldx, C, ptr, sch, out, ldo);
    } else {
      dim3 grid((unsigned)B, (unsigned)ceil_div(C, POOL_COLS));
      segment_pool_kernel<float, 1><<<grid, threads, 0, st>>>(xf, ldx, C, ptr, sch, out, ldo);
    }
  } else if (x_dtype == GN_BF16) {
    const __nv_bfloat16* xb = reinterpret_cast<const __nv_bfloat16*>(x);
    bool vec = (C % 4 == 0) && (ldx % 4 == 0) && ((reinterpret_cast<uintptr_t>(x) & 7) == 0);
    if (vec) {
      dim3 grid((unsigned)B, (unsigned)ceil_div(C, POOL_COLS * 4));
      segment_pool_kernel<__nv_bfloat16, 4><<<grid, threads, 0, st>>>(xb, ldx, C, ptr, sch, out, ldo);
    } else {
      dim3 grid((unsigned)B, (unsigned)ceil_div(C, POOL_COLS));
      segment_pool_kernel<__nv_bfloat16, 1><<<grid, threads, 0, st>>>(xb, ldx, C, ptr, sch, out, ldo);
    }
  } else {
    GN_CHECK_ARG(false, "gn_segment_pool: bad x_dtype %d", x_dtype);
  }
  GN_LAUNCH_CHECK();
  return GN_OK;
}
