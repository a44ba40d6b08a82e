// add_norm_layer=True (reference gnn/dynedge.py:202-203, :224-226): a torch.nn.LayerNorm sits
// between every Linear and its activation in the EdgeConv MLPs and the post-processing MLP.
// The normalisation runs over the full feature row of every EDGE, which the fused edge kernels
// (column-tiled) cannot see, so this option takes the reference's own route - per-edge rows are
// materialised (fp32) - with three small HBM-bound kernels around the dense layers.
#include "common.cuh"

namespace gn {

// X[m, :] <- act(LayerNorm(X[m, :]) * gamma + beta), in place; one warp per row, values kept
// in registers between the passes (N <= 32 * LN_MAX_PER_LANE)
constexpr int LN_MAX_PER_LANE = 16;
template <int ACT>
__global__ void __launch_bounds__(256) layernorm_act_kernel(float* __restrict__ X, int64_t ldx,
                                                            int64_t M, int N,
                                                            const float* __restrict__ gamma,
                                                            const float* __restrict__ beta,
                                                            float eps) {
  const int64_t m = (int64_t)blockIdx.x * 8 + (threadIdx.x >> 5);
  const int lane = threadIdx.x & 31;
  if (m >= M) return;
  float* row = X + m * ldx;
  float v[LN_MAX_PER_LANE];
  float sum = 0.f;
#pragma unroll
  for (int t = 0; t < LN_MAX_PER_LANE; ++t) {
    const int n = lane + 32 * t;
    v[t] = n < N ? row[n] : 0.f;
    sum += v[t];
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
  const float mean = sum / (float)N;
  float sq = 0.f;
#pragma unroll
  for (int t = 0; t < LN_MAX_PER_LANE; ++t) {
    const int n = lane + 32 * t;
    const float d = n < N ? v[t] - mean : 0.f;
    sq += d * d;
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) sq += __shfl_xor_sync(0xffffffffu, sq, o);
  const float rstd = rsqrtf(sq / (float)N + eps);   // biased variance, as torch.nn.LayerNorm
#pragma unroll
  for (int t = 0; t < LN_MAX_PER_LANE; ++t) {
    const int n = lane + 32 * t;
    if (n < N) {
      float y = (v[t] - mean) * rstd;
      y = y * (gamma ? gamma[n] : 1.f) + (beta ? beta[n] : 0.f);
      row[n] = activate<ACT>(y);
    }
  }
}

// out[e, :] = P[i] + Q[j] for slot e = i*S + t of an ELL graph (P = pq[:, :a], Q = pq[:, a:]);
// empty slots give zero rows
__global__ void __launch_bounds__(256) edge_preact_kernel(const float* __restrict__ pq, int a,
                                                          const int32_t* __restrict__ nbr, int S,
                                                          int64_t E, float* __restrict__ out,
                                                          int64_t ldo) {
  const int64_t e = (int64_t)blockIdx.x * 8 + (threadIdx.x >> 5);
  const int lane = threadIdx.x & 31;
  if (e >= E) return;
  const int32_t j = nbr[e];
  const int64_t i = e / S;
  for (int c = lane; c < a; c += 32)
    out[e * ldo + c] = j >= 0 ? pq[i * 2 * a + c] + pq[(int64_t)j * 2 * a + a + c] : 0.f;
}

// out[i, :] = sum over the FILLED slots of node i of Z[i*S + t, :]
__global__ void __launch_bounds__(256) slot_sum_kernel(const float* __restrict__ Z, int64_t ldz,
                                                       int b, const int32_t* __restrict__ nbr,
                                                       int S, int64_t N, float* __restrict__ out,
                                                       int64_t ldo) {
  const int64_t i = (int64_t)blockIdx.x * 8 + (threadIdx.x >> 5);
  const int lane = threadIdx.x & 31;
  if (i >= N) return;
  for (int c = lane; c < b; c += 32) {
    float acc = 0.f;
    for (int t = 0; t < S; ++t)
      if (nbr[i * S + t] >= 0) acc += Z[(i * S + t) * ldz + c];
    out[i * ldo + c] = acc;
  }
}

// out[i, :] = aggr over the FILLED slots of node i of Z[i*S + t, :]; aggr as GN_AGGR_* (PyG's
// MessagePassing aggregation, components/layers.py:27-76: add / mean / max / min); a node
// without in-edges gives 0.
__global__ void __launch_bounds__(256) slot_aggregate_kernel(const float* __restrict__ Z,
                                                             int64_t ldz, int b,
                                                             const int32_t* __restrict__ nbr, int S,
                                                             int64_t N, int aggr,
                                                             float* __restrict__ out, int64_t ldo) {
  const int64_t i = (int64_t)blockIdx.x * 8 + (threadIdx.x >> 5);
  const int lane = threadIdx.x & 31;
  if (i >= N) return;
  int cnt = 0;
  for (int t = 0; t < S; ++t) cnt += nbr[i * S + t] >= 0;
  for (int c = lane; c < b; c += 32) {
    float acc = aggr == GN_AGGR_MAX ? -INFINITY : (aggr == GN_AGGR_MIN ? INFINITY : 0.f);
    for (int t = 0; t < S; ++t) {
      if (nbr[i * S + t] < 0) continue;
      const float v = Z[(i * S + t) * ldz + c];
      if (aggr == GN_AGGR_MAX) acc = fmaxf(acc, v);
      else if (aggr == GN_AGGR_MIN) acc = fminf(acc, v);
      else acc += v;
    }
    if (cnt == 0) acc = 0.f;
    else if (aggr == GN_AGGR_MEAN) acc /= (float)cnt;
    out[i * ldo + c] = acc;
  }
}

}  // namespace gn

using namespace gn;

extern "C" int gn_layernorm_act_f32(float* X, int64_t ldx, int64_t M, int N, const float* gamma,
                                    const float* beta, float eps, int act, void* stream) {
  GN_CHECK_ARG(M >= 0 && N >= 1 && ldx >= N, "gn_layernorm_act_f32: bad sizes");
  GN_CHECK_ARG(N <= 32 * LN_MAX_PER_LANE, "gn_layernorm_act_f32: N=%d exceeds %d", N,
               32 * LN_MAX_PER_LANE);
  if (M == 0) return GN_OK;
  GN_CHECK_ARG(X != nullptr, "gn_layernorm_act_f32: null pointer");
  const unsigned grid = (unsigned)ceil_div(M, 8);
  cudaStream_t st = as_stream(stream);
  switch (act) {
    case GN_ACT_NONE: layernorm_act_kernel<ACT_NONE><<<grid, 256, 0, st>>>(X, ldx, M, N, gamma, beta, eps); break;
    case GN_ACT_RELU: layernorm_act_kernel<ACT_RELU><<<grid, 256, 0, st>>>(X, ldx, M, N, gamma, beta, eps); break;
    case GN_ACT_GELU: layernorm_act_kernel<ACT_GELU><<<grid, 256, 0, st>>>(X, ldx, M, N, gamma, beta, eps); break;
    case GN_ACT_LEAKY_RELU: layernorm_act_kernel<ACT_LEAKY_RELU><<<grid, 256, 0, st>>>(X, ldx, M, N, gamma, beta, eps); break;
    default: set_error("unknown activation %d", act); return GN_ERR_INVALID_ARGUMENT;
  }
  GN_LAUNCH_CHECK();
  return GN_OK;
}

extern "C" int gn_edge_preact_f32(const float* pq, int a, const gn_graph* graph, int64_t N,
                                  float* out, int64_t ldo, void* stream) {
  GN_CHECK_ARG(graph && graph->col && graph->ell_stride > 0,
               "gn_edge_preact_f32: the graph must be in ELL layout");
  GN_CHECK_ARG(a >= 1 && ldo >= a, "gn_edge_preact_f32: bad sizes");
  if (N == 0) return GN_OK;
  GN_CHECK_ARG(pq && out, "gn_edge_preact_f32: null pointer");
  const int64_t E = N * graph->ell_stride;
  edge_preact_kernel<<<(unsigned)ceil_div(E, 8), 256, 0, as_stream(stream)>>>(
      pq, a, graph->col, graph->ell_stride, E, out, ldo);
  GN_LAUNCH_CHECK();
  return GN_OK;
}

extern "C" int gn_slot_sum_f32(const float* Z, int64_t ldz, int b, const gn_graph* graph, int64_t N,
                               float* out, int64_t ldo, void* stream) {
  GN_CHECK_ARG(graph && graph->col && graph->ell_stride > 0,
               "gn_slot_sum_f32: the graph must be in ELL layout");
  GN_CHECK_ARG(b >= 1 && ldz >= b && ldo >= b, "gn_slot_sum_f32: bad sizes");
  if (N == 0) return GN_OK;
  GN_CHECK_ARG(Z && out, "gn_slot_sum_f32: null pointer");
  slot_sum_kernel<<<(unsigned)ceil_div(N, 8), 256, 0, as_stream(stream)>>>(
      Z, ldz, b, graph->col, graph->ell_stride, N, out, ldo);
  GN_LAUNCH_CHECK();
  return GN_OK;
}

extern "C" int gn_slot_aggregate_f32(const float* Z, int64_t ldz, int b, const gn_graph* graph,
                                     int64_t N, int aggr, float* out, int64_t ldo, void* stream) {
  GN_CHECK_ARG(graph && graph->col && graph->ell_stride > 0,
               "gn_slot_aggregate_f32: the graph must be in ELL layout");
  GN_CHECK_ARG(b >= 1 && ldz >= b && ldo >= b, "gn_slot_aggregate_f32: bad sizes");
  GN_CHECK_ARG(aggr == GN_AGGR_ADD || aggr == GN_AGGR_MEAN || aggr == GN_AGGR_MAX ||
                   aggr == GN_AGGR_MIN,
               "gn_slot_aggregate_f32: unknown aggregation %d", aggr);
  if (N == 0) return GN_OK;
  GN_CHECK_ARG(Z && out, "gn_slot_aggregate_f32: null pointer");
  slot_aggregate_kernel<<<(unsigned)ceil_div(N, 8), 256, 0, as_stream(stream)>>>(
      Z, ldz, b, graph->col, graph->ell_stride, N, aggr, out, ldo);
  GN_LAUNCH_CHECK();
  return GN_OK;
}
