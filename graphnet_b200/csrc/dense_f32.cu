// fp32 (FFMA) dense layers: Linear(+activation) and the fused EdgeConv.
//
// This is the accuracy-anchored path (outputs within 1e-4 relative of the CPU
// reference): plain fp32 fused multiply-adds, no tensor cores.  The bf16
// tcgen05 path in tc_*.cu is the throughput path.
//
// EdgeConv (reference: components/layers.py:67 -> torch_geometric EdgeConv with the
// MLP of models/gnn/dynedge.py:193-204) is computed without ever materialising the
// [E, 2d] / [E, a] / [E, b] edge tensors:
//   1. node GEMM   PQ[i] = [ (W1a - W1b) x_i + b1 ,  W1b x_i ]           ([N, 2a])
//   2. per tile of 64 target nodes, for each neighbour slot s:
//        H[r]  = act(P[i_r] + Q[nbr(i_r, s)])       (built in shared memory)
//        D     = H W2^T                              (register-tiled FFMA)
//        acc  (+)= act(D + b2)  for rows with s < deg(i_r)
//   3. out[i] = acc   (sum / mean / max / min over the node's in-edges)
#include <cuda_fp16.h>

#include <algorithm>

#include "common.cuh"

namespace gn {

// ------------------------------------------------------------------- linear
constexpr int LIN_TK = 16, LIN_THREADS = 256;

template <typename OutT>
__device__ __forceinline__ OutT cvt_out(float v);
template <>
__device__ __forceinline__ float cvt_out<float>(float v) { return v; }
template <>
__device__ __forceinline__ __nv_bfloat16 cvt_out<__nv_bfloat16>(float v) {
  return __float2bfloat16_rn(v);
}
template <>
__device__ __forceinline__ __half cvt_out<__half>(float v) {  // saturating
  return __float2half_rn(fminf(fmaxf(v, -65504.f), 65504.f));
}

// TM x TN output tile per CTA, (TM/16) x (TN/16) outputs per thread.
template <int ACT, int TM, int TN, typename OutT>
__global__ void __launch_bounds__(LIN_THREADS) linear_f32_kernel(
    const float* __restrict__ A, int64_t lda, const float* __restrict__ W,
    const float* __restrict__ bias, OutT* __restrict__ C, int64_t ldc, int64_t M, int N, int K) {
  constexpr int RM = TM / 16, RN = TN / 16;
  __shared__ __align__(16) float As[LIN_TK][TM + 4];
  __shared__ __align__(16) float Ws[LIN_TK][TN + 4];
  const int tid = threadIdx.x;
  const int64_t m0 = (int64_t)blockIdx.x * TM;
  const int n0 = blockIdx.y * TN;
  const int ty = tid / 16, tx = tid % 16;
  float acc[RM][RN];
#pragma unroll
  for (int i = 0; i < RM; ++i)
#pragma unroll
    for (int j = 0; j < RN; ++j) acc[i][j] = 0.f;

  for (int k0 = 0; k0 < K; k0 += LIN_TK) {
    // TM x 16 and TN x 16 elements, k fastest within a row
#pragma unroll
    for (int t = 0; t < RM; ++t) {
      const int idx = tid + t * LIN_THREADS;
      const int r = idx / LIN_TK, kk = idx % LIN_TK;
      const int64_t m = m0 + r;
      As[kk][r] = (m < M && k0 + kk < K) ? A[m * lda + k0 + kk] : 0.f;
    }
#pragma unroll
    for (int t = 0; t < RN; ++t) {
      const int idx = tid + t * LIN_THREADS;
      const int r = idx / LIN_TK, kk = idx % LIN_TK;
      const int n = n0 + r;
      Ws[kk][r] = (n < N && k0 + kk < K) ? W[(int64_t)n * K + k0 + kk] : 0.f;
    }
    __syncthreads();
#pragma unroll
    for (int k = 0; k < LIN_TK; ++k) {
      float av[RM], wv[RN];
#pragma unroll
      for (int i = 0; i < RM; ++i) av[i] = As[k][ty * RM + i];
#pragma unroll
      for (int j = 0; j < RN; ++j) wv[j] = Ws[k][tx * RN + j];
#pragma unroll
      for (int i = 0; i < RM; ++i)
#pragma unroll
        for (int j = 0; j < RN; ++j) acc[i][j] = fmaf(av[i], wv[j], acc[i][j]);
    }
    __syncthreads();
  }
#pragma unroll
  for (int i = 0; i < RM; ++i) {
    const int64_t m = m0 + ty * RM + i;
    if (m >= M) continue;
    const int nb = n0 + tx * RN;
    OutT* dst = C + m * ldc + nb;
    if (RN == 4 && sizeof(OutT) == 2 && nb + 4 <= N &&
        (reinterpret_cast<uintptr_t>(dst) & 7) == 0) {
      // four consecutive 16-bit outputs: one 8-byte store
      OutT o[4];
#pragma unroll
      for (int j = 0; j < 4; ++j)
        o[j] = cvt_out<OutT>(activate<ACT>(acc[i][j] + (bias ? bias[nb + j] : 0.f)));
      *reinterpret_cast<uint2*>(dst) = *reinterpret_cast<const uint2*>(o);
    } else {
#pragma unroll
      for (int j = 0; j < RN; ++j) {
        const int n = nb + j;
        if (n < N) {
          float v = acc[i][j] + (bias ? bias[n] : 0.f);
          C[m * ldc + n] = cvt_out<OutT>(activate<ACT>(v));
        }
      }
    }
  }
}

// Small-K variant (K <= 36, 16-bit output): the first EdgeConv's node GEMM has K = 2F+5 = 19
// and is store bound (N*Kp*2 bytes out).  A thread owns two adjacent output columns - its two
// weight rows (2K values) and biases stay in registers - and a CTA walks SK_ROWS input rows that
// are staged in shared memory and read with warp-uniform addresses (16-byte broadcasts).  A
// warp therefore stores 128 contiguous bytes of one output row per instruction; the previous
// shape of this kernel (a thread per row) wrote 16 bytes per lane 2*N bytes apart.
// The k order of the sum (bias first, k ascending) is that of the generic kernel above.
constexpr int SK_MAXK = 36;  // 2F+5 = 33 for the 14 IceCube-Upgrade features
constexpr int SK_ROWS = 64;
template <typename OutT, int KMAX>
__global__ void __launch_bounds__(512) linear_smallk_kernel(
    const float* __restrict__ A, int64_t lda, const float* __restrict__ W,
    const float* __restrict__ bias, OutT* __restrict__ C, int64_t ldc, int64_t M, int N, int K) {
  static_assert(KMAX % 4 == 0, "rows are read as float4");
  __shared__ __align__(16) float s_a[SK_ROWS][KMAX];
  const int n0 = threadIdx.x * 2;
  const bool col_ok = n0 < N;  // N is even
  float w0[KMAX], w1[KMAX];
#pragma unroll
  for (int k = 0; k < KMAX; ++k) {
    w0[k] = (col_ok && k < K) ? __ldg(W + (int64_t)n0 * K + k) : 0.f;
    w1[k] = (col_ok && k < K) ? __ldg(W + (int64_t)(n0 + 1) * K + k) : 0.f;
  }
  const float b0 = (col_ok && bias) ? bias[n0] : 0.f;
  const float b1 = (col_ok && bias) ? bias[n0 + 1] : 0.f;
  const int64_t m0 = (int64_t)blockIdx.x * SK_ROWS;
  const int rows = (int)(M - m0 < SK_ROWS ? M - m0 : SK_ROWS);
  for (int i = threadIdx.x; i < SK_ROWS * KMAX; i += blockDim.x) {
    const int r = i / KMAX, k = i - r * KMAX;
    s_a[r][k] = (r < rows && k < K) ? __ldg(A + (m0 + r) * lda + k) : 0.f;
  }
  __syncthreads();
  if (!col_ok) return;
  for (int r = 0; r < rows; r += 2) {  // two rows per step: four independent FMA chains
    float c00 = b0, c01 = b1, c10 = b0, c11 = b1;
#pragma unroll
    for (int k4 = 0; k4 < KMAX; k4 += 4) {
      const float4 a0 = *reinterpret_cast<const float4*>(&s_a[r][k4]);
      const float4 a1 = *reinterpret_cast<const float4*>(&s_a[r + 1][k4]);
      const float x0[4] = {a0.x, a0.y, a0.z, a0.w};
      const float x1[4] = {a1.x, a1.y, a1.z, a1.w};
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        c00 = fmaf(x0[u], w0[k4 + u], c00);
        c01 = fmaf(x0[u], w1[k4 + u], c01);
        c10 = fmaf(x1[u], w0[k4 + u], c10);
        c11 = fmaf(x1[u], w1[k4 + u], c11);
      }
    }
    OutT o0[2] = {cvt_out<OutT>(c00), cvt_out<OutT>(c01)};
    *reinterpret_cast<uint32_t*>(C + (m0 + r) * ldc + n0) = *reinterpret_cast<const uint32_t*>(o0);
    if (r + 1 < rows) {
      OutT o1[2] = {cvt_out<OutT>(c10), cvt_out<OutT>(c11)};
      *reinterpret_cast<uint32_t*>(C + (m0 + r + 1) * ldc + n0) =
          *reinterpret_cast<const uint32_t*>(o1);
    }
  }
}

// --------------------------------------------------------- weight repacking
__global__ void edgeconv_prepare_kernel(const float* __restrict__ W1, const float* __restrict__ W2,
                                        int d, int a, int b, float* __restrict__ Wpq,
                                        float* __restrict__ W2T) {
  const int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int64_t n1 = (int64_t)a * d;
  if (idx < n1) {
    const int r = (int)(idx / d), c = (int)(idx % d);
    const float wa = W1[(int64_t)r * 2 * d + c];
    const float wb = W1[(int64_t)r * 2 * d + d + c];
    Wpq[idx] = wa - wb;        // rows 0..a-1 : acts on x_i
    Wpq[n1 + idx] = wb;        // rows a..2a-1: acts on x_j
  }
  if (W2 != nullptr) {
    const int64_t n2 = (int64_t)a * b;
    if (idx < n2) {
      const int k = (int)(idx / b), n = (int)(idx % b);  // W2T[k][n] = W2[n][k]
      W2T[idx] = W2[(int64_t)n * a + k];
    }
  }
}

// ------------------------------------------------------------ fused EdgeConv
constexpr int EC_TM = 64;        // target nodes per CTA
constexpr int EC_TN = 128;       // output columns per CTA
constexpr int EC_TK = 16;        // K chunk of W2T staged per step
constexpr int EC_THREADS = 256;  // 16 row-groups (4 rows) x 16 column-groups (8 cols)
constexpr int EC_HS = EC_TM + 4; // padded row length of the transposed H tile

template <int AGGR>
__device__ __forceinline__ float aggr_init() {
  if (AGGR == GN_AGGR_MAX) return __int_as_float(0xff800000);
  if (AGGR == GN_AGGR_MIN) return __int_as_float(0x7f800000);
  return 0.f;
}
template <int AGGR>
__device__ __forceinline__ float aggr_combine(float acc, float v) {
  if (AGGR == GN_AGGR_MAX) return fmaxf(acc, v);
  if (AGGR == GN_AGGR_MIN) return fminf(acc, v);
  return acc + v;
}

// Two-layer MLP variant: dynamic smem = a * EC_HS floats (H^T) + 2 * EC_TK * EC_TN (W2T chunks)
template <int ACT, int AGGR>
__global__ void __launch_bounds__(EC_THREADS) edgeconv_f32_kernel(
    const float* __restrict__ pq, int a, GraphView g, int64_t N, const float* __restrict__ W2T,
    const float* __restrict__ b2, int b, float* __restrict__ out, int64_t ldo) {
  extern __shared__ __align__(16) float smem[];
  float* Hs = smem;                       // [a][EC_HS]  (k-major, node-minor)
  float* Wsm = smem + (size_t)a * EC_HS;  // [2][EC_TK][EC_TN]
  __shared__ int s_deg[EC_TM];
  __shared__ int64_t s_start[EC_TM];
  __shared__ int s_maxdeg;

  const int tid = threadIdx.x;
  const int64_t i0 = (int64_t)blockIdx.x * EC_TM;
  const int n0 = blockIdx.y * EC_TN;
  const int ty = tid / 16, tx = tid % 16;
  const int ld2a = 2 * a;

  if (tid == 0) s_maxdeg = 0;
  __syncthreads();
  if (tid < EC_TM) {
    int64_t i = i0 + tid, start = 0;
    int cnt = 0;
    if (i < N) graph_row(g, i, start, cnt);
    s_deg[tid] = cnt;
    s_start[tid] = start;
    atomicMax(&s_maxdeg, cnt);
  }
  __syncthreads();
  const int maxdeg = s_maxdeg;

  float acc[4][8];
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 8; ++j) acc[i][j] = aggr_init<AGGR>();
  float bias[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    const int n = n0 + (j < 4 ? tx * 4 + j : 64 + tx * 4 + (j - 4));
    bias[j] = (n < b) ? b2[n] : 0.f;
  }
  int mydeg[4];
#pragma unroll
  for (int i = 0; i < 4; ++i) mydeg[i] = s_deg[ty * 4 + i];

  const int nchunks = (a + EC_TK - 1) / EC_TK;
  for (int s = 0; s < maxdeg; ++s) {
    // ---- build H^T for this slot: Hs[k][r] = act(P[i_r][k] + Q[j_r][k]) ----
    __syncthreads();  // previous slot's GEMM has finished reading Hs
    for (int r = tid / 32; r < EC_TM; r += EC_THREADS / 32) {
      const int lane = tid & 31;
      const bool valid = s < s_deg[r];
      const int64_t i = i0 + r;
      int64_t j = 0;
      if (valid) j = g.col[s_start[r] + s];
      const float* Pi = pq + i * ld2a;
      const float* Qj = pq + j * ld2a + a;
      for (int k = lane; k < a; k += 32) {
        float v = 0.f;
        if (valid) v = activate<ACT>(Pi[k] + Qj[k]);
        Hs[k * EC_HS + r] = v;
      }
    }
    // ---- D = H W2^T over K = a, W2T streamed through smem in chunks ----
    float D[4][8];
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
      for (int j = 0; j < 8; ++j) D[i][j] = 0.f;

    auto load_chunk = [&](int ch, int buf) {
      // EC_TK x EC_TN floats = 2048 -> 8 per thread, rows of 128 contiguous columns
      float* dst = Wsm + buf * EC_TK * EC_TN;
#pragma unroll
      for (int t = 0; t < 8; ++t) {
        const int idx = tid + t * EC_THREADS;
        const int kk = idx / EC_TN, nn = idx % EC_TN;
        const int k = ch * EC_TK + kk, n = n0 + nn;
        dst[idx] = (k < a && n < b) ? W2T[(int64_t)k * b + n] : 0.f;
      }
    };
    load_chunk(0, 0);
    __syncthreads();
    for (int ch = 0; ch < nchunks; ++ch) {
      const int buf = ch & 1;
      if (ch + 1 < nchunks) load_chunk(ch + 1, buf ^ 1);
      const float* wbuf = Wsm + buf * EC_TK * EC_TN;
      const int kbase = ch * EC_TK;
#pragma unroll
      for (int kk = 0; kk < EC_TK; ++kk) {
        const int k = kbase + kk;
        if (k < a) {
          const float4 h = *reinterpret_cast<const float4*>(&Hs[k * EC_HS + ty * 4]);
          const float4 w0 = *reinterpret_cast<const float4*>(&wbuf[kk * EC_TN + tx * 4]);
          const float4 w1 = *reinterpret_cast<const float4*>(&wbuf[kk * EC_TN + 64 + tx * 4]);
          const float hv[4] = {h.x, h.y, h.z, h.w};
          const float wv[8] = {w0.x, w0.y, w0.z, w0.w, w1.x, w1.y, w1.z, w1.w};
#pragma unroll
          for (int i = 0; i < 4; ++i)
#pragma unroll
            for (int j = 0; j < 8; ++j) D[i][j] = fmaf(hv[i], wv[j], D[i][j]);
        }
      }
      __syncthreads();
    }
    // ---- message = act(D + b2), aggregate rows whose slot s is a real edge ----
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      if (s < mydeg[i]) {
#pragma unroll
        for (int j = 0; j < 8; ++j)
          acc[i][j] = aggr_combine<AGGR>(acc[i][j], activate<ACT>(D[i][j] + bias[j]));
      }
    }
  }

#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int64_t node = i0 + ty * 4 + i;
    if (node >= N) continue;
    const int dg = mydeg[i];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const int n = n0 + (j < 4 ? tx * 4 + j : 64 + tx * 4 + (j - 4));
      if (n < b) {
        float v = acc[i][j];
        if (dg == 0) v = 0.f;
        else if (AGGR == GN_AGGR_MEAN) v = v / (float)dg;
        out[node * ldo + n] = v;
      }
    }
  }
}

// One-layer MLP variant: out[i] = aggr_j act(P_i + Q_j); one warp per node.
template <int ACT, int AGGR>
__global__ void edgeconv1_f32_kernel(const float* __restrict__ pq, int a, GraphView g, int64_t N,
                                     float* __restrict__ out, int64_t ldo) {
  const int64_t i = (int64_t)blockIdx.x * (blockDim.x / 32) + threadIdx.x / 32;
  if (i >= N) return;
  const int lane = threadIdx.x & 31;
  int64_t start;
  int cnt;
  graph_row(g, i, start, cnt);
  const float* Pi = pq + i * 2 * a;
  for (int k = lane; k < a; k += 32) {
    float acc = aggr_init<AGGR>();
    const float p = Pi[k];
    for (int s = 0; s < cnt; ++s) {
      const int64_t j = g.col[start + s];
      acc = aggr_combine<AGGR>(acc, activate<ACT>(p + pq[j * 2 * a + a + k]));
    }
    if (cnt == 0) acc = 0.f;
    else if (AGGR == GN_AGGR_MEAN) acc = acc / (float)cnt;
    out[i * ldo + k] = acc;
  }
}

template <int ACT, typename OutT>
static int launch_linear(const float* A, int64_t lda, const float* W, const float* bias, void* C,
                         int64_t ldc, int64_t M, int N, int K, cudaStream_t st) {
  // small problems (readout, heads): 32x32 tiles so that the grid still covers the SMs
  const int64_t big_ctas = ceil_div(M, 64) * ceil_div(N, 64);
  if (big_ctas >= 256) {
    dim3 grid((unsigned)ceil_div(M, 64), (unsigned)ceil_div(N, 64));
    linear_f32_kernel<ACT, 64, 64, OutT><<<grid, LIN_THREADS, 0, st>>>(
        A, lda, W, bias, reinterpret_cast<OutT*>(C), ldc, M, N, K);
  } else {
    dim3 grid((unsigned)ceil_div(M, 32), (unsigned)ceil_div(N, 32));
    linear_f32_kernel<ACT, 32, 32, OutT><<<grid, LIN_THREADS, 0, st>>>(
        A, lda, W, bias, reinterpret_cast<OutT*>(C), ldc, M, N, K);
  }
  return GN_OK;
}

template <typename OutT>
static int linear_f32_t(const float* A, int64_t lda, const float* W, const float* bias, void* C,
                        int64_t ldc, int64_t M, int N, int K, int act, cudaStream_t st) {
  switch (act) {
    case GN_ACT_NONE: launch_linear<ACT_NONE, OutT>(A, lda, W, bias, C, ldc, M, N, K, st); break;
    case GN_ACT_RELU: launch_linear<ACT_RELU, OutT>(A, lda, W, bias, C, ldc, M, N, K, st); break;
    case GN_ACT_GELU: launch_linear<ACT_GELU, OutT>(A, lda, W, bias, C, ldc, M, N, K, st); break;
    case GN_ACT_LEAKY_RELU:
      launch_linear<ACT_LEAKY_RELU, OutT>(A, lda, W, bias, C, ldc, M, N, K, st);
      break;
    default: set_error("unknown activation %d", act); return GN_ERR_INVALID_ARGUMENT;
  }
  GN_LAUNCH_CHECK();
  return GN_OK;
}

int linear_f32(const float* A, int64_t lda, const float* W, const float* bias, void* C,
               int64_t ldc, int c_dtype, int64_t M, int N, int K, int act, cudaStream_t st) {
  if (M == 0) return GN_OK;
  if (K <= SK_MAXK && N % 2 == 0 && N <= 1024 && ldc % 2 == 0 && act == GN_ACT_NONE &&
      c_dtype != GN_F32 && (reinterpret_cast<uintptr_t>(C) & 3) == 0 && M * (int64_t)N >= (1 << 16)) {
    const unsigned threads = (unsigned)((N / 2 + 31) / 32 * 32);
    const unsigned grid = (unsigned)ceil_div(M, SK_ROWS);
#define GN_SK_CASE(T_, KM_)                                                                      \
  {                                                                                              \
    linear_smallk_kernel<T_, KM_><<<grid, threads, 0, st>>>(                                     \
        A, lda, W, bias, reinterpret_cast<T_*>(C), ldc, M, N, K);                                \
  }
    if (c_dtype == GN_F16) {
      if (K <= 20) GN_SK_CASE(__half, 20) else if (K <= 32) GN_SK_CASE(__half, 32) else GN_SK_CASE(__half, 36)
    } else {
      if (K <= 20) GN_SK_CASE(__nv_bfloat16, 20)
      else if (K <= 32) GN_SK_CASE(__nv_bfloat16, 32)
      else GN_SK_CASE(__nv_bfloat16, 36)
    }
#undef GN_SK_CASE
    GN_LAUNCH_CHECK();
    return GN_OK;
  }
  if (c_dtype == GN_F32) return linear_f32_t<float>(A, lda, W, bias, C, ldc, M, N, K, act, st);
  if (c_dtype == GN_BF16)
    return linear_f32_t<__nv_bfloat16>(A, lda, W, bias, C, ldc, M, N, K, act, st);
  if (c_dtype == GN_F16) return linear_f32_t<__half>(A, lda, W, bias, C, ldc, M, N, K, act, st);
  set_error("unknown output dtype %d", c_dtype);
  return GN_ERR_INVALID_ARGUMENT;
}

template <int ACT, int AGGR>
static int launch_edgeconv(const float* pq, int a, const GraphView& g, int64_t N, const float* W2T,
                           const float* b2, int b, float* out, int64_t ldo, cudaStream_t st) {
  if (W2T == nullptr) {
    const int warps = 8;
    edgeconv1_f32_kernel<ACT, AGGR><<<(unsigned)ceil_div(N, warps), warps * 32, 0, st>>>(
        pq, a, g, N, out, ldo);
    GN_LAUNCH_CHECK();
    return GN_OK;
  }
  const size_t smem = ((size_t)a * EC_HS + 2 * EC_TK * EC_TN) * sizeof(float);
  if (smem > 220 * 1024) {
    set_error("gn_edgeconv_f32: hidden width a=%d needs %zu B of shared memory (max 220 KB)", a,
              smem);
    return GN_ERR_UNSUPPORTED;
  }
  auto kern = edgeconv_f32_kernel<ACT, AGGR>;
  GN_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  dim3 grid((unsigned)ceil_div(N, EC_TM), (unsigned)ceil_div(b, EC_TN));
  kern<<<grid, EC_THREADS, smem, st>>>(pq, a, g, N, W2T, b2, b, out, ldo);
  GN_LAUNCH_CHECK();
  return GN_OK;
}

template <int ACT>
static int dispatch_aggr(int aggr, const float* pq, int a, const GraphView& g, int64_t N,
                         const float* W2T, const float* b2, int b, float* out, int64_t ldo,
                         cudaStream_t st) {
  switch (aggr) {
    case GN_AGGR_ADD: return launch_edgeconv<ACT, GN_AGGR_ADD>(pq, a, g, N, W2T, b2, b, out, ldo, st);
    case GN_AGGR_MEAN: return launch_edgeconv<ACT, GN_AGGR_MEAN>(pq, a, g, N, W2T, b2, b, out, ldo, st);
    case GN_AGGR_MAX: return launch_edgeconv<ACT, GN_AGGR_MAX>(pq, a, g, N, W2T, b2, b, out, ldo, st);
    case GN_AGGR_MIN: return launch_edgeconv<ACT, GN_AGGR_MIN>(pq, a, g, N, W2T, b2, b, out, ldo, st);
  }
  set_error("unknown aggregation %d", aggr);
  return GN_ERR_INVALID_ARGUMENT;
}

}  // namespace gn

using namespace gn;

extern "C" int gn_linear_f32(const float* A, int64_t lda, const float* W, const float* bias,
                             void* C, int64_t ldc, int c_dtype, int64_t M, int N, int K, int act,
                             void* stream) {
  GN_CHECK_ARG(M >= 0 && N >= 1 && K >= 1 && lda >= K && ldc >= N, "gn_linear_f32: bad sizes");
  GN_CHECK_ARG(M == 0 || (A && W && C), "gn_linear_f32: null pointer");
  return linear_f32(A, lda, W, bias, C, ldc, c_dtype, M, N, K, act, as_stream(stream));
}

extern "C" int gn_edgeconv_prepare_f32(const float* W1, const float* W2, int d, int a, int b,
                                       float* Wpq, float* W2T, void* stream) {
  GN_CHECK_ARG(d >= 1 && a >= 1 && W1 && Wpq, "gn_edgeconv_prepare_f32: bad arguments");
  GN_CHECK_ARG((W2 == nullptr) == (W2T == nullptr), "gn_edgeconv_prepare_f32: W2/W2T mismatch");
  GN_CHECK_ARG(W2 == nullptr || b >= 1, "gn_edgeconv_prepare_f32: bad b");
  int64_t n = (int64_t)a * d;
  if (W2) n = n > (int64_t)a * b ? n : (int64_t)a * b;
  edgeconv_prepare_kernel<<<(unsigned)ceil_div(n, 256), 256, 0, as_stream(stream)>>>(W1, W2, d, a, b,
                                                                                    Wpq, W2T);
  GN_LAUNCH_CHECK();
  return GN_OK;
}

extern "C" int gn_edgeconv_f32(const float* x, int64_t ldx, int d, const gn_graph* graph, int64_t N,
                               const float* Wpq, const float* b1, int a, const float* W2T,
                               const float* b2, int b, int act, int aggr, float* pq, float* out,
                               int64_t ldo, void* stream) {
  GN_CHECK_ARG(N >= 0 && d >= 1 && a >= 1 && b >= 1 && ldx >= d && ldo >= b,
               "gn_edgeconv_f32: bad sizes");
  if (N == 0) return GN_OK;
  GN_CHECK_ARG(x && Wpq && pq && out, "gn_edgeconv_f32: null pointer");
  GN_CHECK_ARG(W2T != nullptr || b == a, "gn_edgeconv_f32: one-layer MLP needs b == a");
  GraphView g;
  int rc = make_graph_view(graph, &g);
  if (rc != GN_OK) return rc;
  cudaStream_t st = as_stream(stream);
  // 1. node GEMM: pq[:, 0:a] = x (W1a-W1b)^T + b1 ; pq[:, a:2a] = x W1b^T
  rc = linear_f32(x, ldx, Wpq, b1, pq, 2 * a, GN_F32, N, a, d, GN_ACT_NONE, st);
  if (rc != GN_OK) return rc;
  rc = linear_f32(x, ldx, Wpq + (int64_t)a * d, nullptr, pq + a, 2 * a, GN_F32, N, a, d,
                  GN_ACT_NONE, st);
  if (rc != GN_OK) return rc;
  // 2./3. gather + second layer + aggregation
  switch (act) {
    case GN_ACT_NONE: return dispatch_aggr<ACT_NONE>(aggr, pq, a, g, N, W2T, b2, b, out, ldo, st);
    case GN_ACT_RELU: return dispatch_aggr<ACT_RELU>(aggr, pq, a, g, N, W2T, b2, b, out, ldo, st);
    case GN_ACT_GELU: return dispatch_aggr<ACT_GELU>(aggr, pq, a, g, N, W2T, b2, b, out, ldo, st);
    case GN_ACT_LEAKY_RELU:
      return dispatch_aggr<ACT_LEAKY_RELU>(aggr, pq, a, g, N, W2T, b2, b, out, ldo, st);
  }
  set_error("unknown activation %d", act);
  return GN_ERR_INVALID_ARGUMENT;
}
