// Fused EdgeConv message + aggregation on tcgen05 / TMEM (fp16 hidden layer, fp32 accumulate).
//
//   out[i] = aggr_{j in N(i)} act( W2 act(P_i + Q_j) + b2 )
//
// with P = x (W1a-W1b)^T + b1 and Q = x W1b^T produced by the node GEMM (tc_linear.cu), so
// that W1 [x_i, x_j - x_i] + b1 = P_i + Q_j.  Replaces torch_geometric EdgeConv.propagate +
// torch_scatter as reached from components/layers.py:67 (MLP of dynedge.py:193-204,
// aggr="add" dynedge.py:208).  The [E, a] hidden tensor and the [E, b] messages never reach
// HBM: they live in shared memory / TMEM for the lifetime of one tile.  The kernel is in
// tc_edge_pair.cuh (a CTA pair per TPC, tcgen05.mma.cta_group::2); this file holds its
// parameter block, the layout helper for graphs that are not slot tables, and the C ABI.
//
// Tile rows are ELL slots: row r of tile t is (node t*NPT + r/KP, slot r%KP), KP slots per node;
// empty slots are all-zero h rows.  b2 is folded into the GEMM: P carries 1.0 in the two pad
// columns a, a+1 and W2's columns a, a+1 hold b2 split into fp16 hi + lo parts, so an empty
// slot contributes exactly 0 and the epilogue needs no bias add.
//
// The hidden layer (P, Q, h, W2) is IEEE fp16 rather than bf16: same tensor-core rate, 3 more
// mantissa bits, which matters because P_i + Q_j cancels (W1b (x_j - x_i) is small for near
// neighbours); pre-activations are bounded by the weights, so fp16's range is ample (the node
// GEMM saturates at +-65504).  Layer inputs/outputs stay bf16.
//
// In-edges beyond the KP slots of a node (a node keeps k+1 edges when >= k+1 coincident pulses
// precede it; user graphs may have any degree) are handled by a second pass of the SAME kernel
// over a compacted list of (node, slot base) groups, accumulating with red.global.add.
#include <cuda_fp16.h>
#include <stdlib.h>

#include <atomic>

#include "tc_common.cuh"

namespace gn {
namespace tc {

constexpr int EK_KCH = 32;                       // hidden width is padded to a multiple of this
constexpr int EK_MAX_CHUNKS = 11;                // hidden width (padded) <= 352

struct EdgeParams {
  const __half* pmat;       // [N, ldp]: P = x (W1a-W1b)^T + b1  (+ the two "ones" columns)
  int64_t ldp;
  const __half* qmat;       // [N, ldq]: Q = x W1b^T   (the gathered operand, kept L2 resident)
  int64_t ldq;
  int Kp;                   // padded hidden width, multiple of 32
  int nch;                  // Kp / 32
  const int32_t* ell;       // [groups, ell_ld] neighbour ids (the first KP of a row are used), -1 = empty slot
  int64_t ell_ld;
  const int32_t* group_node;  // target node of each group, or null: group g is node g
  const int32_t* group_count; // device int: number of groups, or null: N groups
  int64_t group_cap;          // capacity of the group lists (a malformed graph can count past it)
  int accumulate;           // 0: store, 1: red.add into out / coords (overflow pass)
  int64_t N;
  int bnp;                  // output width padded to a multiple of 16 (UMMA N)
  int b;                    // valid output columns
  void* out;                // [N, ldo] bf16 or fp16
  int out_f16;              // 0: bf16, 1: fp16 (saturated)
  int64_t ldo;
  float* coords;            // [N, ncoord] fp32 copy of the columns selected by coord_mask, or null
  int ncoord;
  uint32_t coord_mask[8];
  const void* w2;           // W2p [bnp, Kp] fp16 row-major
  void* dbg;                // profiling builds: wait-time counters + timeline (GN_EDGE_DEBUG), else null
  uint32_t* relu_mask;      // training: [8][mask_ld] words, bit c of word (g, row) = [z(row, 32g+c) > 0]
  int64_t mask_ld;          // words per 32-channel group: >= tiles * 64
  int act;                  // CTA-pair kernel: activation of both Linear layers (gn::Act)
  int aggr;                 // CTA-pair kernel: 0 add, 1 mean, 2 max, 3 min
  const int32_t* deg;       // slots in use per node (ELL graphs; mean / max / min need it)
  const void* neg_row;      // CTA-pair kernel: 1 KB of fp16 -inf, read in place of Q_j by empty slots
  const void* w2lo;         // fp32-accurate kernel: fp16(W2 - fp16(W2)), laid out like w2
  const void* neg_row32;    // fp32-accurate kernel: a row of fp32 -inf
};

__device__ __forceinline__ uint32_t relu_add_f16x2(uint32_t a, uint32_t b) {
  uint32_t r;  // one HFMA2.RELU: max(a * 1 + b, 0)
  asm("fma.rn.relu.f16x2 %0, %1, %2, %3;" : "=r"(r) : "r"(a), "r"(0x3C003C00u), "r"(b));
  return r;
}

__device__ __forceinline__ uint16_t pack1_16(float a, bool f16) {
  if (f16) {
    __half h = __float2half_rn(fmaxf(fminf(a, 65504.f), -65504.f));   // saturate, never inf
    return *reinterpret_cast<uint16_t*>(&h);
  }
  __nv_bfloat16 h = __float2bfloat16_rn(a);
  return *reinterpret_cast<uint16_t*>(&h);
}
__device__ __forceinline__ void red_add1_16(void* dst, float a, bool f16) {
  if (f16) atomicAdd(reinterpret_cast<__half*>(dst),
                     __float2half_rn(fmaxf(fminf(a, 65504.f), -65504.f)));
  else atomicAdd(reinterpret_cast<__nv_bfloat16*>(dst), __float2bfloat16_rn(a));
}

}  // namespace tc
}  // namespace gn
#include "tc_edge_pair.cuh"
#include "tc_edge_x3.cuh"
namespace gn {
namespace tc {

// Graph (CSR or ELL) -> the group lists the main kernel consumes:
//   ell[N, KP]            slots 0..KP-1 of every node (-1 padded)          (pass 1, group = node)
//   ovf_node[], ovf_ell[] one entry per further block of KP in-edges of a node  (pass 2)
// ovf_count is a device counter (zeroed by the caller); entries are appended with atomicAdd,
// their order does not affect the result (pass 2 accumulates with red.add).
__global__ void graph_to_groups_kernel(GraphView g, int64_t N, int KP, int32_t* __restrict__ ell,
                                       int32_t* __restrict__ ovf_node,
                                       int32_t* __restrict__ ovf_ell,
                                       int32_t* __restrict__ ovf_count, int64_t ovf_cap) {
  const int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= N * KP) return;
  const int64_t i = idx / KP;
  const int s = (int)(idx % KP);
  int64_t start;
  int cnt;
  graph_row(g, i, start, cnt);
  ell[idx] = s < cnt ? g.col[start + s] : -1;
  if (s == 0 && cnt > KP) {
    const int extra = (cnt - 1) / KP;  // further blocks of KP edges
    const int base = atomicAdd(ovf_count, extra);
    for (int e = 0; e < extra; ++e) {
      if (base + e >= ovf_cap) break;  // cannot happen when the caller sized the scratch right
      ovf_node[base + e] = (int32_t)i;
      for (int t = 0; t < KP; ++t) {
        const int slot = (e + 1) * KP + t;
        ovf_ell[(int64_t)(base + e) * KP + t] = slot < cnt ? g.col[start + slot] : -1;
      }
    }
  }
}

// Slot table with MORE entries per row than the kernel's slot count (a kNN graph, stride k + 1, run
// with KP = k slots: a row holds k + 1 neighbours only where more than k points coincide with the
// query, so nearly every tile row carries an edge instead of 1 in k + 1 staying empty): the table is
// read in place, and the entries beyond KP go to the overflow group list.
__global__ void ell_overflow_kernel(const int32_t* __restrict__ nbr, int32_t stride,
                                    const int32_t* __restrict__ deg, int64_t N, int KP,
                                    int32_t* __restrict__ ovf_node, int32_t* __restrict__ ovf_ell,
                                    int32_t* __restrict__ ovf_count, int64_t ovf_cap) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= N) return;
  const int cnt = min(max(deg[i], 0), stride);
  if (cnt <= KP) return;
  const int extra = (cnt - 1) / KP;
  const int base = atomicAdd(ovf_count, extra);
  for (int e = 0; e < extra; ++e) {
    if (base + e >= ovf_cap) break;
    ovf_node[base + e] = (int32_t)i;
    for (int t = 0; t < KP; ++t) {
      const int slot = (e + 1) * KP + t;
      ovf_ell[(int64_t)(base + e) * KP + t] = slot < cnt ? nbr[i * stride + slot] : -1;
    }
  }
}

}  // namespace tc
}  // namespace gn

using namespace gn;

static int64_t edge_overflow_cap(const gn_graph* graph, int64_t N, int slots) {
  if (graph->ell_stride > 0) return N * ((graph->ell_stride + slots - 1) / slots - 1);
  return graph->num_edges / slots + 1;
}

#ifdef GN_DEBUG_BUILD
// profiling builds only (declared in include/graphnet_b200_debug.h): 8 wait-time counters, then a
// timeline of cluster 0 - [CTA rank 0..1][warp 0..31][step 0..63][4 stamps], cycles since the
// kernel's set-up barrier
constexpr int EDGE_DBG_WORDS = 8 + 2 * 32 * 64 * 4;
static __device__ unsigned long long g_edge_dbg[EDGE_DBG_WORDS];

// copy out and clear the counters
extern "C" int gn_debug_edge_counters(unsigned long long* host8) {
  GN_CUDA(cudaDeviceSynchronize());
  GN_CUDA(cudaMemcpyFromSymbol(host8, g_edge_dbg, 8 * sizeof(unsigned long long)));
  unsigned long long zero[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  GN_CUDA(cudaMemcpyToSymbol(g_edge_dbg, zero, sizeof(zero)));
  return GN_OK;
}
// the per-warp timeline of the last launch
extern "C" int gn_debug_edge_trace(unsigned long long* host, int words) {
  GN_CUDA(cudaDeviceSynchronize());
  if (words > EDGE_DBG_WORDS - 8) words = EDGE_DBG_WORDS - 8;
  GN_CUDA(cudaMemcpyFromSymbol(host, g_edge_dbg, words * sizeof(unsigned long long),
                               8 * sizeof(unsigned long long)));
  return GN_OK;
}
#endif

// 1 KB of fp16 -inf (what an empty slot gathers instead of a neighbour row)
#define GN_NI4 0xFC00FC00u, 0xFC00FC00u, 0xFC00FC00u, 0xFC00FC00u
#define GN_NI16 GN_NI4, GN_NI4, GN_NI4, GN_NI4
#define GN_NI64 GN_NI16, GN_NI16, GN_NI16, GN_NI16
static __device__ __align__(16) const uint32_t g_neg_inf_row[256] = {GN_NI64, GN_NI64, GN_NI64,
                                                                      GN_NI64};
#undef GN_NI4
#undef GN_NI16
#undef GN_NI64
// 352 fp32 -inf (the fp32-accurate kernel's rows are up to 352 columns wide)
#define GN_NF4 0xFF800000u, 0xFF800000u, 0xFF800000u, 0xFF800000u
#define GN_NF16 GN_NF4, GN_NF4, GN_NF4, GN_NF4
#define GN_NF64 GN_NF16, GN_NF16, GN_NF16, GN_NF16
static __device__ __align__(16) const uint32_t g_neg_inf_row32[352] = {
    GN_NF64, GN_NF64, GN_NF64, GN_NF64, GN_NF64, GN_NF16, GN_NF16};
#undef GN_NF4
#undef GN_NF16
#undef GN_NF64

// How the kernels get at a graph's in-edges:
//   EDGE_DIRECT   an ELL graph whose stride is a supported slot count, run with that many slots:
//                 consumed in place, no lists, no second pass;
//   EDGE_INPLACE  an ELL graph run with FEWER slots than its stride: the first `slots` entries of
//                 every row in place, the rest through the overflow list (ell_overflow_kernel);
//   EDGE_GROUPS   anything else (CSR, odd strides): graph_to_groups_kernel copies the rows.
enum EdgeMode { EDGE_DIRECT, EDGE_INPLACE, EDGE_GROUPS };
static bool edge_supported_slots(int s) { return s == 8 || s == 9 || s == 16 || s == 17 || s == 32; }
static EdgeMode edge_mode(const gn_graph* graph, int slots) {
  if (graph->ell_stride > slots && edge_supported_slots(slots)) return EDGE_INPLACE;
  if (graph->ell_stride > 0 && edge_supported_slots(graph->ell_stride)) return EDGE_DIRECT;
  return EDGE_GROUPS;
}

extern "C" size_t gn_tc_edge_scratch_bytes(const gn_graph* graph, int64_t N, int slots) {
  if (!graph || slots <= 0) return 0;
  const EdgeMode mode = edge_mode(graph, slots);
  if (mode == EDGE_DIRECT) return 16;
  const int64_t n = N > 0 ? N : 1;
  const int64_t cap = edge_overflow_cap(graph, n, slots);
  // [ell [N, slots]] + ovf_ell [cap, slots] + ovf_node [cap] + counter (16-byte padded)
  return (size_t)((mode == EDGE_GROUPS ? n * slots : 0) + cap * slots + cap + 4) * sizeof(int32_t);
}

// the group lists of one call, carved from the caller's scratch and filled on `st`
struct EdgeLists {
  EdgeMode mode;
  int slots;                 // slots the kernels run with
  const int32_t* ell;        // pass 1: [N, ell_ld]
  int64_t ell_ld;
  int32_t *ovf_ell, *ovf_node, *ovf_count;
  int64_t cap;
};
static int edge_prepare_lists(const gn_graph* graph, const GraphView& g, int64_t N, int slots,
                              void* scratch, cudaStream_t st, EdgeLists* L) {
  L->mode = edge_mode(graph, slots);
  L->slots = L->mode == EDGE_DIRECT ? graph->ell_stride : slots;
  L->cap = L->mode == EDGE_DIRECT ? 0 : edge_overflow_cap(graph, N, slots);
  int32_t* base = reinterpret_cast<int32_t*>(scratch);
  int32_t* ell_copy = base;
  L->ovf_ell = base + (L->mode == EDGE_GROUPS ? N * slots : 0);
  L->ovf_node = L->ovf_ell + L->cap * slots;
  L->ovf_count = L->ovf_node + L->cap;
  if (L->mode == EDGE_GROUPS) {
    L->ell = ell_copy;
    L->ell_ld = slots;
    GN_CUDA(cudaMemsetAsync(L->ovf_count, 0, sizeof(int32_t), st));
    tc::graph_to_groups_kernel<<<(unsigned)ceil_div(N * slots, 256), 256, 0, st>>>(
        g, N, slots, ell_copy, L->ovf_node, L->ovf_ell, L->ovf_count, L->cap);
    GN_LAUNCH_CHECK();
  } else {
    L->ell = graph->col;
    L->ell_ld = graph->ell_stride;
    if (L->mode == EDGE_INPLACE) {
      GN_CHECK_ARG(graph->deg != nullptr, "gn_tc_edge_mlp_aggregate: ELL graph without deg");
      GN_CUDA(cudaMemsetAsync(L->ovf_count, 0, sizeof(int32_t), st));
      tc::ell_overflow_kernel<<<(unsigned)ceil_div(N, 256), 256, 0, st>>>(
          graph->col, graph->ell_stride, graph->deg, N, slots, L->ovf_node, L->ovf_ell,
          L->ovf_count, L->cap);
      GN_LAUNCH_CHECK();
    }
  }
  return GN_OK;
}

extern "C" int64_t gn_tc_edge_relu_mask_ld(int64_t N, int slots) {
  if (N <= 0 || !edge_supported_slots(slots)) return 0;
  return ceil_div(N, (int64_t)(tc::E4_ROWS / slots)) * tc::E4_ROWS;
}

static int tc_edge_mlp_aggregate_impl(const void* P, int64_t ldp, const void* Q, int64_t ldq,
                                      int Kp, const gn_graph* graph, int64_t N, int slots,
                                      const void* W2p, int b, void* out, int64_t ldo,
                                      int out_dtype, float* coords,
                                      const int32_t* coord_cols, int ncoord, void* scratch,
                                      size_t scratch_bytes, uint32_t* relu_mask, int64_t mask_ld,
                                      int act, int aggr, void* stream) {
  GN_CHECK_ARG(act == GN_ACT_RELU || act == GN_ACT_LEAKY_RELU || act == GN_ACT_GELU,
               "gn_tc_edge_mlp_aggregate: activation %d (ReLU, LeakyReLU or GELU)", act);
  GN_CHECK_ARG(aggr >= GN_AGGR_ADD && aggr <= GN_AGGR_MIN, "gn_tc_edge_mlp_aggregate: aggr %d", aggr);
  GN_CHECK_ARG(out_dtype == GN_BF16 || out_dtype == GN_F16,
               "gn_tc_edge_mlp_aggregate: out_dtype must be GN_BF16 or GN_F16");
  GN_CHECK_ARG(N >= 0 && b >= 1 && b <= 256, "gn_tc_edge_mlp_aggregate: b=%d not in [1,256]", b);
  GN_CHECK_ARG(Kp >= 32 && Kp % 32 == 0 && Kp <= tc::EK_MAX_CHUNKS * tc::EK_KCH,
               "gn_tc_edge_mlp_aggregate: padded hidden width %d not a multiple of 32 in [32,%d]",
               Kp, tc::EK_MAX_CHUNKS * tc::EK_KCH);
  GN_CHECK_ARG(edge_supported_slots(slots),
               "gn_tc_edge_mlp_aggregate: slots=%d (must be 8, 9, 16, 17 or 32)", slots);
  GN_CHECK_ARG(ldp >= Kp && ldq >= Kp && ldp % 8 == 0 && ldq % 8 == 0 && ldo % 4 == 0 && ldo >= b,
               "gn_tc_edge_mlp_aggregate: bad leading dimensions");
  GN_CHECK_ARG(ncoord >= 0 && ncoord <= 16, "gn_tc_edge_mlp_aggregate: ncoord=%d", ncoord);
  if (N == 0) return GN_OK;
  GN_CHECK_ARG(P && Q && W2p && out && scratch && graph,
               "gn_tc_edge_mlp_aggregate: null pointer");
  GN_CHECK_ARG(scratch_bytes >= gn_tc_edge_scratch_bytes(graph, N, slots),
               "gn_tc_edge_mlp_aggregate: scratch too small");
  GN_CHECK_ARG((reinterpret_cast<uintptr_t>(P) & 15) == 0 &&
               (reinterpret_cast<uintptr_t>(Q) & 15) == 0 &&
               (reinterpret_cast<uintptr_t>(out) & 7) == 0 &&
               (reinterpret_cast<uintptr_t>(scratch) & 3) == 0,
               "gn_tc_edge_mlp_aggregate: P, Q must be 16-byte and out 8-byte aligned");
  GN_CHECK_ARG(N * 33 < (int64_t)1 << 31, "gn_tc_edge_mlp_aggregate: N too large (32-bit slot offsets)");
  GraphView g;
  int rc = make_graph_view(graph, &g);
  if (rc != GN_OK) return rc;
  cudaStream_t st = as_stream(stream);
  const int bnp = (b + 15) & ~15;

  tc::EdgeParams p;
  memset(&p, 0, sizeof(p));
  for (int t = 0; t < ncoord; ++t) {
    GN_CHECK_ARG(coord_cols && coord_cols[t] >= 0 && coord_cols[t] < b,
                 "gn_tc_edge_mlp_aggregate: coordinate column out of range");
    GN_CHECK_ARG(t == 0 || coord_cols[t] > coord_cols[t - 1],
                 "gn_tc_edge_mlp_aggregate: coordinate columns must be strictly increasing");
    p.coord_mask[coord_cols[t] >> 5] |= 1u << (coord_cols[t] & 31);
  }
  if (relu_mask != nullptr || aggr != GN_AGGR_ADD) {
    // every in-edge of a node in ONE row of slots: run with the table's own stride
    if (graph->ell_stride > 0 && edge_supported_slots(graph->ell_stride)) slots = graph->ell_stride;
  }
  const bool direct = edge_mode(graph, slots) == EDGE_DIRECT;
  if (direct) slots = graph->ell_stride;
  if (relu_mask != nullptr) {
    if (!direct) {
      set_error("gn_tc_edge_mlp_aggregate_mask: needs an ELL graph with 8, 9, 16, 17 or 32 slots "
                "per node");
      return GN_ERR_UNSUPPORTED;
    }
    GN_CHECK_ARG(mask_ld >= gn_tc_edge_relu_mask_ld(N, slots),
                 "gn_tc_edge_mlp_aggregate_mask: mask_ld=%lld too small", (long long)mask_ld);
  }
  EdgeLists L;
  rc = edge_prepare_lists(graph, g, N, slots, scratch, st, &L);
  if (rc != GN_OK) return rc;
  const int64_t cap = L.cap;

  p.w2 = W2p;
  p.act = act;
  p.aggr = aggr;
  p.deg = direct ? graph->deg : nullptr;
  if (aggr != GN_AGGR_ADD && !direct) {
    // mean / max / min need every in-edge of a node in ONE group of slots (an overflow group
    // could only be combined with the first by a second reduction)
    set_error("gn_tc_edge_mlp_aggregate: aggr != add needs an ELL graph with 8, 9, 16, 17 or 32 "
              "slots per node (every kNN graph of this library)");
    return GN_ERR_UNSUPPORTED;
  }
  if ((act != GN_ACT_RELU || aggr != GN_AGGR_ADD) && relu_mask != nullptr) {
    set_error("gn_tc_edge_mlp_aggregate_mask: the training forward is ReLU + add only");
    return GN_ERR_UNSUPPORTED;
  }
  const bool generic = act != GN_ACT_RELU || aggr != GN_AGGR_ADD;
  p.relu_mask = relu_mask;
  p.mask_ld = mask_ld;
  p.dbg = nullptr;
#ifdef GN_DEBUG_BUILD   // GN_EDGE_DEBUG=1: wait-time counters and a timeline (graphnet_b200_debug.h)
  static const bool dbg_on = getenv("GN_EDGE_DEBUG") != nullptr;
  if (dbg_on) GN_CUDA(cudaGetSymbolAddress(&p.dbg, g_edge_dbg));
#endif
  p.pmat = reinterpret_cast<const __half*>(P);
  p.ldp = ldp;
  p.qmat = reinterpret_cast<const __half*>(Q);
  p.ldq = ldq;
  p.Kp = Kp;
  p.nch = Kp / tc::EK_KCH;
  p.N = N;
  p.bnp = bnp;
  p.b = b;
  p.out = out;
  p.out_f16 = out_dtype == GN_F16;
  p.ldo = ldo;
  p.coords = ncoord > 0 ? coords : nullptr;
  p.ncoord = ncoord;
  const int npt = tc::E4_ROWS / slots;

  int dev = 0, sms = 0;
  GN_CUDA(cudaGetDevice(&dev));
  {
    static std::atomic<int> sm_count[64];   // per device, fetched once
    sms = sm_count[dev & 63].load(std::memory_order_relaxed);
    if (sms == 0) {
      GN_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
      sm_count[dev & 63].store(sms, std::memory_order_relaxed);
    }
  }
  GN_CHECK_ARG(sms >= 2, "gn_tc_edge_mlp_aggregate: needs a CTA pair (>= 2 SMs)");
  {   // a __device__ symbol has one address per device: looked up once
    static std::atomic<void*> neg_row_addr[64];
    void* nr = neg_row_addr[dev & 63].load(std::memory_order_relaxed);
    if (nr == nullptr) {
      GN_CUDA(cudaGetSymbolAddress(&nr, g_neg_inf_row));
      neg_row_addr[dev & 63].store(nr, std::memory_order_relaxed);
    }
    p.neg_row = nr;
  }
  const size_t smem4 = tc::edge4_smem_bytes(Kp);
  for (int pass = 0; pass < 2; ++pass) {
    int64_t max_tiles;
    if (pass == 0) {
      p.ell = L.ell;
      p.ell_ld = L.ell_ld;
      p.group_node = nullptr;
      p.group_count = nullptr;
      p.accumulate = 0;
      max_tiles = ceil_div(N, npt);
    } else {
      if (cap <= 0) break;
      p.ell = L.ovf_ell;
      p.ell_ld = slots;
      p.group_node = L.ovf_node;
      p.group_count = L.ovf_count;
      p.group_cap = cap;
      p.accumulate = 1;
      max_tiles = ceil_div(cap, npt);
    }
    // CTA pairs: one cluster of 2 per TPC, a step = two 64-row tiles
    const int64_t max_steps = (max_tiles + 1) / 2;
    const unsigned grid4 = 2u * (unsigned)(max_steps < sms / 2 ? max_steps : sms / 2);
  /* the opt-in to > 48 KB of dynamic shared memory is per function and device: once */
#define GN_EDGE_LAUNCH4(KP_, ACC_, F16_)                                                         \
  {                                                                                              \
    const int var_ = generic ? 2 : (!(ACC_) && p.relu_mask != nullptr) ? 1 : 0;                  \
    auto kern = var_ == 2 ? tc::edge_mlp_aggregate_pair_kernel<KP_, ACC_, F16_, true, false>     \
              : var_ == 1 ? tc::edge_mlp_aggregate_pair_kernel<KP_, false, F16_, false, true>    \
                          : tc::edge_mlp_aggregate_pair_kernel<KP_, ACC_, F16_, false, false>;   \
    static std::atomic<unsigned long long> ready_[3];                                            \
    if (!((ready_[var_].load(std::memory_order_relaxed) >> (dev & 63)) & 1ull)) {                \
      GN_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize,            \
                                   (int)tc::edge4_smem_bytes(tc::EK_MAX_CHUNKS * tc::EK_KCH)));  \
      ready_[var_].fetch_or(1ull << (dev & 63), std::memory_order_relaxed);                      \
    }                                                                                            \
    kern<<<grid4, tc::E4_THREADS, smem4, st>>>(p);                                               \
  }
#define GN_EDGE_CASE4(KP_)                                                                       \
  {                                                                                              \
    if (p.accumulate) {                                                                          \
      if (p.out_f16) GN_EDGE_LAUNCH4(KP_, true, true) else GN_EDGE_LAUNCH4(KP_, true, false)     \
    } else {                                                                                     \
      if (p.out_f16) GN_EDGE_LAUNCH4(KP_, false, true) else GN_EDGE_LAUNCH4(KP_, false, false)   \
    }                                                                                            \
  }
    switch (slots) {
      case 8: GN_EDGE_CASE4(8) break;
      case 9: GN_EDGE_CASE4(9) break;
      case 16: GN_EDGE_CASE4(16) break;
      case 17: GN_EDGE_CASE4(17) break;
      default: GN_EDGE_CASE4(32) break;
    }
#undef GN_EDGE_CASE4
#undef GN_EDGE_LAUNCH4
    GN_LAUNCH_CHECK();
  }
  return GN_OK;
}

// fp32 in, fp32 out, fp16 (hi, lo) operand pairs on the tensor cores: see tc_edge_x3.cuh
extern "C" int gn_tc_edge_mlp_aggregate_f32(const float* P, int64_t ldp, const float* Q,
                                            int64_t ldq, int Kp, const gn_graph* graph, int64_t N,
                                            int slots, const void* W2hi, const void* W2lo, int b,
                                            float* out, int64_t ldo, float* coords,
                                            const int32_t* coord_cols, int ncoord, void* scratch,
                                            size_t scratch_bytes, void* stream) {
  GN_CHECK_ARG(N >= 0 && b >= 1 && b <= 256, "gn_tc_edge_mlp_aggregate_f32: b=%d not in [1,256]", b);
  GN_CHECK_ARG(Kp >= 32 && Kp % 32 == 0 && Kp <= 352,
               "gn_tc_edge_mlp_aggregate_f32: padded hidden width %d not a multiple of 32 in "
               "[32,352]", Kp);
  GN_CHECK_ARG(edge_supported_slots(slots),
               "gn_tc_edge_mlp_aggregate_f32: slots=%d (must be 8, 9, 16, 17 or 32)", slots);
  GN_CHECK_ARG(ldp >= Kp && ldq >= Kp && ldp % 4 == 0 && ldq % 4 == 0 && ldo >= b,
               "gn_tc_edge_mlp_aggregate_f32: bad leading dimensions");
  GN_CHECK_ARG(ncoord >= 0 && ncoord <= 16, "gn_tc_edge_mlp_aggregate_f32: ncoord=%d", ncoord);
  if (N == 0) return GN_OK;
  GN_CHECK_ARG(P && Q && W2hi && W2lo && out && scratch && graph,
               "gn_tc_edge_mlp_aggregate_f32: null pointer");
  GN_CHECK_ARG(scratch_bytes >= gn_tc_edge_scratch_bytes(graph, N, slots),
               "gn_tc_edge_mlp_aggregate_f32: scratch too small");
  GN_CHECK_ARG((reinterpret_cast<uintptr_t>(P) & 15) == 0 &&
               (reinterpret_cast<uintptr_t>(Q) & 15) == 0 &&
               (reinterpret_cast<uintptr_t>(W2hi) & 15) == 0 &&
               (reinterpret_cast<uintptr_t>(W2lo) & 15) == 0 &&
               (reinterpret_cast<uintptr_t>(scratch) & 3) == 0,
               "gn_tc_edge_mlp_aggregate_f32: P, Q, W2 must be 16-byte aligned");
  GN_CHECK_ARG(N * 33 < (int64_t)1 << 31,
               "gn_tc_edge_mlp_aggregate_f32: N too large (32-bit slot offsets)");
  GraphView g;
  int rc = make_graph_view(graph, &g);
  if (rc != GN_OK) return rc;
  cudaStream_t st = as_stream(stream);
  tc::EdgeParams p;
  memset(&p, 0, sizeof(p));
  for (int t = 0; t < ncoord; ++t) {
    GN_CHECK_ARG(coord_cols && coord_cols[t] >= 0 && coord_cols[t] < b,
                 "gn_tc_edge_mlp_aggregate_f32: coordinate column out of range");
    GN_CHECK_ARG(t == 0 || coord_cols[t] > coord_cols[t - 1],
                 "gn_tc_edge_mlp_aggregate_f32: coordinate columns must be strictly increasing");
    p.coord_mask[coord_cols[t] >> 5] |= 1u << (coord_cols[t] & 31);
  }
  EdgeLists L;
  rc = edge_prepare_lists(graph, g, N, slots, scratch, st, &L);
  if (rc != GN_OK) return rc;
  slots = L.slots;
  const int64_t cap = L.cap;
  p.w2 = W2hi;
  p.w2lo = W2lo;
  p.act = GN_ACT_RELU;
  p.aggr = GN_AGGR_ADD;
  p.pmat = reinterpret_cast<const __half*>(P);   // (fp32 here: the kernel reads them as such)
  p.ldp = ldp;
  p.qmat = reinterpret_cast<const __half*>(Q);
  p.ldq = ldq;
  p.Kp = Kp;
  p.nch = Kp / tc::EK_KCH;
  p.N = N;
  p.bnp = (b + 15) & ~15;
  p.b = b;
  p.out = out;
  p.ldo = ldo;
  p.coords = ncoord > 0 ? coords : nullptr;
  p.ncoord = ncoord;
  const int npt = tc::E4_ROWS / slots;
  int dev = 0, sms = 0;
  GN_CUDA(cudaGetDevice(&dev));
  GN_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
  GN_CHECK_ARG(sms >= 2, "gn_tc_edge_mlp_aggregate_f32: needs a CTA pair (>= 2 SMs)");
  {
    static std::atomic<void*> neg_row_addr[64];
    void* nr = neg_row_addr[dev & 63].load(std::memory_order_relaxed);
    if (nr == nullptr) {
      GN_CUDA(cudaGetSymbolAddress(&nr, g_neg_inf_row32));
      neg_row_addr[dev & 63].store(nr, std::memory_order_relaxed);
    }
    p.neg_row32 = nr;
  }
  const size_t smem = tc::edge4_smem_bytes(Kp);   // four h tiles, as the 16-bit kernel
  for (int pass = 0; pass < 2; ++pass) {
    int64_t max_tiles;
    if (pass == 0) {
      p.ell = L.ell;
      p.ell_ld = L.ell_ld;
      p.group_node = nullptr;
      p.group_count = nullptr;
      p.accumulate = 0;
      max_tiles = ceil_div(N, npt);
    } else {
      if (cap <= 0) break;
      p.ell = L.ovf_ell;
      p.ell_ld = slots;
      p.group_node = L.ovf_node;
      p.group_count = L.ovf_count;
      p.group_cap = cap;
      p.accumulate = 1;
      max_tiles = ceil_div(cap, npt);
    }
    const int64_t max_steps = (max_tiles + 1) / 2;
    const unsigned grid = 2u * (unsigned)(max_steps < sms / 2 ? max_steps : sms / 2);
#define GN_X3_LAUNCH(KP_, ACC_)                                                                  \
  {                                                                                              \
    auto kern = tc::edge_mlp_aggregate_x3_kernel<KP_, ACC_>;                                     \
    static std::atomic<unsigned long long> ready_;                                               \
    if (!((ready_.load(std::memory_order_relaxed) >> (dev & 63)) & 1ull)) {                      \
      GN_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize,            \
                                   (int)tc::edge4_smem_bytes(352)));                             \
      ready_.fetch_or(1ull << (dev & 63), std::memory_order_relaxed);                            \
    }                                                                                            \
    kern<<<grid, tc::E4_THREADS, smem, st>>>(p);                                                 \
  }
#define GN_X3_CASE(KP_)                                                  \
  {                                                                      \
    if (p.accumulate) GN_X3_LAUNCH(KP_, true) else GN_X3_LAUNCH(KP_, false) \
  }
    switch (slots) {
      case 8: GN_X3_CASE(8) break;
      case 9: GN_X3_CASE(9) break;
      case 16: GN_X3_CASE(16) break;
      case 17: GN_X3_CASE(17) break;
      default: GN_X3_CASE(32) break;
    }
#undef GN_X3_CASE
#undef GN_X3_LAUNCH
    GN_LAUNCH_CHECK();
  }
  return GN_OK;
}

extern "C" int gn_tc_edge_mlp_aggregate(const void* P, int64_t ldp, const void* Q, int64_t ldq,
                                        int Kp, const gn_graph* graph, int64_t N, int slots,
                                        const void* W2p, int b, void* out, int64_t ldo,
                                        int out_dtype, float* coords,
                                        const int32_t* coord_cols, int ncoord, void* scratch,
                                        size_t scratch_bytes, int act, int aggr, void* stream) {
  return tc_edge_mlp_aggregate_impl(P, ldp, Q, ldq, Kp, graph, N, slots, W2p, b, out, ldo,
                                    out_dtype, coords, coord_cols, ncoord, scratch, scratch_bytes,
                                    nullptr, 0, act, aggr, stream);
}

extern "C" int gn_tc_edge_mlp_aggregate_mask(const void* P, int64_t ldp, const void* Q,
                                             int64_t ldq, int Kp, const gn_graph* graph, int64_t N,
                                             int slots, const void* W2p, int b, void* out,
                                             int64_t ldo, int out_dtype, float* coords,
                                             const int32_t* coord_cols, int ncoord, void* scratch,
                                             size_t scratch_bytes, uint32_t* relu_mask,
                                             int64_t mask_ld, void* stream) {
  GN_CHECK_ARG(relu_mask != nullptr, "gn_tc_edge_mlp_aggregate_mask: relu_mask is null");
  return tc_edge_mlp_aggregate_impl(P, ldp, Q, ldq, Kp, graph, N, slots, W2p, b, out, ldo,
                                    out_dtype, coords, coord_cols, ncoord, scratch, scratch_bytes,
                                    relu_mask, mask_ld, GN_ACT_RELU, GN_AGGR_ADD, stream);
}

