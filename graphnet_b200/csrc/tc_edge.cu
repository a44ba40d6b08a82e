// Fused EdgeConv message + aggregation on tcgen05 / TMEM (fp16 hidden layer, fp32 accumulate).
//
//   out[i] = sum_{j in N(i)} ReLU( W2 ReLU(P_i + Q_j) + b2 )
//
// with P = x (W1a-W1b)^T + b1 and Q = x W1b^T produced by the node GEMM (tc_linear.cu), so
// that W1 [x_i, x_j - x_i] + b1 = P_i + Q_j.  Replaces torch_geometric EdgeConv.propagate +
// torch_scatter sum as reached from components/layers.py:67 (MLP of dynedge.py:193-204,
// aggr="add" dynedge.py:208).  The [E, a] hidden tensor and the [E, b] messages never
// reach HBM: they live in shared memory / TMEM for the lifetime of one 128-edge tile.
//
// Persistent kernel, one CTA per SM, 19 warps:
//   warp 0       tcgen05.mma issuer; owns TMEM (2 x 256 fp32 columns: double-buffered D) and
//                loads the whole W2 (bf16, K padded, 64B-swizzled) into shared memory once by TMA
//   warps 1-10   producers: gather P_i / Q_j rows (16-byte chunks, L2 resident), h = ReLU(P+Q)
//                in packed fp16, written into the 64B-swizzled K-major A stage ring; one warp
//                pair per stage, so five K-chunks are being gathered at any time
//   warps 11-18  epilogue: TMEM -> registers, ReLU, reduce-scatter over the KP edge rows of a
//                node with warp shuffles, store bf16 row slice (+ fp32 kNN coordinates)
// Tile rows are ELL slots: row r of tile t is (node t*128/KP + r/KP, slot r%KP); empty slots are
// all-zero A rows.  b2 is folded into the GEMM: P carries 1.0 in the two pad columns a, a+1 and
// W2's columns a, a+1 hold b2 split into bf16 hi + lo parts, so an empty slot contributes
// exactly 0 and the epilogue needs no bias add.
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

#include "tc_common.cuh"

namespace gn {
namespace tc {

constexpr int EK_ROWS = 128;
constexpr int EK_KCH = 32;                       // K elements per stage (64-byte rows)
constexpr int EK_STAGE_BYTES = EK_ROWS * 64;     // 8 KB
constexpr int EK_STAGES = 5;
constexpr int EK_MAX_CHUNKS = 11;                // hidden width (padded) <= 352
constexpr int EK_PROD_WARPS = 2 * EK_STAGES;     // a pair of producer warps owns each stage
constexpr int EK_EPI_WARPS = 8;
constexpr int EK_THREADS = 32 * (1 + EK_PROD_WARPS + EK_EPI_WARPS);  // 544

struct __align__(8) EdgeBarriers {
  uint64_t full[EK_STAGES];
  uint64_t empty[EK_STAGES];
  uint64_t tmem_full[2];
  uint64_t tmem_empty[2];
  uint64_t w2_full;
  uint32_t tmem_base;
};

struct EdgeParams {
  const __half* pmat;       // [N, ldp]: P = x (W1a-W1b)^T + b1  (+ the two "ones" columns)
  int64_t ldp;
  const __half* qmat;       // [N, ldq]: Q = x W1b^T   (the gathered operand, kept L2 resident)
  int64_t ldq;
  int Kp;                   // padded hidden width, multiple of 32
  int nch;                  // Kp / 32
  const int32_t* ell;       // [groups, KP] neighbour ids, -1 = empty slot
  const int32_t* group_node;  // target node of each group, or null: group g is node g
  const int32_t* group_count; // device int: number of groups, or null: N groups
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
  const void* w2;           // W2p [bnp, Kp] fp16 row-major (second-generation kernel reads it directly)
  int use_ws;               // weight-stationary MMA form (collector re-use of the h chunk)
  void* dbg;                // optional 8 x uint64 wait-time counters (GN_EDGE_DEBUG), else null
  uint32_t* relu_mask;      // training: [8][mask_ld] words, bit c of word (g, row) = [z(row, 32g+c) > 0]
  int64_t mask_ld;          // words per 32-channel group: >= tiles * 64
};

__device__ __forceinline__ uint4 ldg16(const __half* p) {
  return __ldg(reinterpret_cast<const uint4*>(p));
}
// Q rows are gathered ~k times each: ask L2 to keep them (evict_last) while the streamed
// operands (P, outputs) pass through with normal priority.
__device__ __forceinline__ uint64_t policy_evict_last() {
  uint64_t pol;
  asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(pol));
  return pol;
}
__device__ __forceinline__ uint4 ldg16_keep(const __half* p, uint64_t pol) {
  uint4 r;
  asm volatile("ld.global.nc.L2::cache_hint.v4.b32 {%0, %1, %2, %3}, [%4], %5;"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
               : "l"(p), "l"(pol));
  return r;
}

__device__ __forceinline__ uint32_t relu_add_f16x2(uint32_t a, uint32_t b) {
  uint32_t r;  // one HFMA2.RELU: max(a * 1 + b, 0)
  asm("fma.rn.relu.f16x2 %0, %1, %2, %3;" : "=r"(r) : "r"(a), "r"(0x3C003C00u), "r"(b));
  return r;
}

__device__ __forceinline__ uint32_t pack2_16(float a, float b, bool f16) {
  if (f16) {
    __half2 h = __floats2half2_rn(fminf(a, 65504.f), fminf(b, 65504.f));  // inputs are >= 0
    return *reinterpret_cast<uint32_t*>(&h);
  }
  __nv_bfloat162 h = __floats2bfloat162_rn(a, b);
  return *reinterpret_cast<uint32_t*>(&h);
}
__device__ __forceinline__ uint16_t pack1_16(float a, bool f16) {
  if (f16) {
    __half h = __float2half_rn(fminf(a, 65504.f));
    return *reinterpret_cast<uint16_t*>(&h);
  }
  __nv_bfloat16 h = __float2bfloat16_rn(a);
  return *reinterpret_cast<uint16_t*>(&h);
}
__device__ __forceinline__ void red_add2_16(void* dst, float a, float b, bool f16) {
  if (f16) atomicAdd(reinterpret_cast<__half2*>(dst), __floats2half2_rn(fminf(a, 65504.f), fminf(b, 65504.f)));
  else atomicAdd(reinterpret_cast<__nv_bfloat162*>(dst), __floats2bfloat162_rn(a, b));
}
__device__ __forceinline__ void red_add1_16(void* dst, float a, bool f16) {
  if (f16) atomicAdd(reinterpret_cast<__half*>(dst), __float2half_rn(fminf(a, 65504.f)));
  else atomicAdd(reinterpret_cast<__nv_bfloat16*>(dst), __float2bfloat16_rn(a));
}

template <int KP>
__global__ void __launch_bounds__(EK_THREADS, 1) edge_mlp_aggregate_kernel(
    const __grid_constant__ CUtensorMap tmW2, const EdgeParams p) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) &
                                             ~uintptr_t(1023));
  const int w2_chunk_bytes = p.bnp * 64;
  uint8_t* w2_smem = smem;
  uint8_t* a_smem = smem + (size_t)p.nch * w2_chunk_bytes;
  EdgeBarriers* bars = reinterpret_cast<EdgeBarriers*>(a_smem + EK_STAGES * EK_STAGE_BYTES);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  constexpr int NPT = EK_ROWS / KP;  // groups (nodes) per tile
  const int nch = p.nch;
  const int64_t num_groups = p.group_count ? (int64_t)__ldg(p.group_count) : p.N;
  const int64_t num_tiles = (num_groups + NPT - 1) / NPT;
  if ((int64_t)blockIdx.x >= num_tiles) return;  // uniform per CTA, before any barrier

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmW2);
#pragma unroll
    for (int s = 0; s < EK_STAGES; ++s) {
      mbar_init(&bars->full[s], 2);  // the pair of producer warps that owns the stage
      mbar_init(&bars->empty[s], 1);
    }
#pragma unroll
    for (int b = 0; b < 2; ++b) {
      mbar_init(&bars->tmem_full[b], 1);
      mbar_init(&bars->tmem_empty[b], EK_EPI_WARPS);
    }
    mbar_init(&bars->w2_full, 1);
    fence_mbar_init();
  }
  if (warp == 0) tmem_alloc<512>(&bars->tmem_base);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = bars->tmem_base;

  if (warp == 0) {
    // ================================================================ MMA issuer
    if (elect_one()) {
      mbar_arrive_expect_tx(&bars->w2_full, (uint32_t)(nch * w2_chunk_bytes));
      for (int kc = 0; kc < nch; ++kc)
        tma_load_2d(w2_smem + (size_t)kc * w2_chunk_bytes, &tmW2, &bars->w2_full, kc * EK_KCH, 0);
    }
    __syncwarp();
    mbar_wait(&bars->w2_full, 0);
    const uint32_t idesc = make_idesc_f16(EK_ROWS, p.bnp);
    const uint32_t w2_addr = smem_u32(w2_smem);
    const uint32_t a_addr = smem_u32(a_smem);
    uint32_t g = 0;  // running stage counter
    uint32_t it = 0;
    for (int64_t tile = blockIdx.x; tile < num_tiles; tile += gridDim.x, ++it) {
      const uint32_t buf = it & 1;
      mbar_wait(&bars->tmem_empty[buf], ((it >> 1) & 1) ^ 1);
      tc_fence_after();
      const uint32_t tmem_d = tmem_base + buf * 256;
      for (int kc = 0; kc < nch; ++kc, ++g) {
        const uint32_t s = g % EK_STAGES;
        mbar_wait(&bars->full[s], (g / EK_STAGES) & 1);
        tc_fence_after();
        if (elect_one()) {
          const uint64_t adesc = make_kmajor_desc<64>(a_addr + s * EK_STAGE_BYTES);
          const uint64_t bdesc = make_kmajor_desc<64>(w2_addr + kc * w2_chunk_bytes);
#pragma unroll
          for (int k2 = 0; k2 < EK_KCH / 16; ++k2)
            mma_bf16_ss(tmem_d, adesc + (uint64_t)(k2 * 2), bdesc + (uint64_t)(k2 * 2), idesc,
                        (kc | k2) != 0);
          mma_commit(&bars->empty[s]);
          if (kc == nch - 1) mma_commit(&bars->tmem_full[buf]);
        }
        __syncwarp();
      }
    }
  } else if (warp <= EK_PROD_WARPS) {
    // ================================================================= producers
    // Stage s of the A ring is owned by the producer warp pair (2s, 2s+1): each warp builds 64
    // of its 128 rows, and the five pairs work on five consecutive K-chunks concurrently.
    // (fence.proxy.async waits for the issuing thread's outstanding global loads, so a warp
    // cannot prefetch across its own fence - the overlap that hides the gather latency comes
    // from the other pairs.  A pair only ever waits for the consumption of its own previous
    // item, which keeps the mbarrier parity protocol one phase deep.)  Lane l owns the 16-byte
    // chunk c = l & 3 of rows 64 half + (l >> 2) + 8 t, t = 0..7; rows are processed two at a
    // time with the next pair's loads in flight.  The 8 rows of one load instruction belong to
    // one node (KP = 8), so the P chunk is fetched once per instruction; Q rows are the gather.
    const int pw = warp - 1;          // 0..9
    const int owner = pw >> 1;        // stage owned by this warp's pair
    const int half = pw & 1;
    const int c = lane & 3;
    const int rq = half * 64 + (lane >> 2);
    const uint32_t a_addr = smem_u32(a_smem);
    const int64_t my_tiles = (num_tiles - blockIdx.x + gridDim.x - 1) / gridDim.x;
    const uint32_t total = (uint32_t)(my_tiles * nch);
    const uint4 zero4 = make_uint4(0, 0, 0, 0);
    const uint64_t pol = policy_evict_last();
    const int64_t ell_len = num_groups * KP;
    const uint32_t base = a_addr + owner * EK_STAGE_BYTES;

    int32_t ids[8];       // neighbour ids of my 8 rows for the current item's tile
    int32_t ids_nx[8];    // ... for my next item (prefetched)
    auto fetch_ids = [&](uint32_t g, int32_t (&dst)[8]) {
      const int64_t tile = blockIdx.x + (int64_t)(g / nch) * gridDim.x;
#pragma unroll
      for (int t = 0; t < 8; ++t) {
        const int64_t e = tile * EK_ROWS + rq + 8 * t;
        dst[t] = (g < total && e < ell_len) ? __ldg(p.ell + e) : -1;
      }
    };
    fetch_ids(owner, ids_nx);

    for (uint32_t g = owner; g < total; g += EK_STAGES) {
#pragma unroll
      for (int t = 0; t < 8; ++t) ids[t] = ids_nx[t];
      fetch_ids(g + EK_STAGES, ids_nx);
      const int64_t tile = blockIdx.x + (int64_t)(g / nch) * gridDim.x;
      const int kc = g % nch;
      const __half* pbase = p.pmat + kc * EK_KCH + c * 8;
      const __half* qbase = p.qmat + kc * EK_KCH + c * 8;

      uint4 Pb[2][2], Qb[2][2];
#define GN_EK_LOAD(BUF, S)                                                                 \
  {                                                                                        \
    _Pragma("unroll") for (int tt = 0; tt < 2; ++tt) {                                     \
      const int t = 2 * (S) + tt;                                                          \
      const int32_t j = ids[t];                                                            \
      if (j >= 0) {                                                                        \
        const int64_t grp = tile * NPT + (rq + 8 * t) / KP;                                \
        const int64_t node = p.group_node ? (int64_t)__ldg(p.group_node + grp) : grp;      \
        Pb[BUF][tt] = ldg16(pbase + node * p.ldp);                                         \
        Qb[BUF][tt] = ldg16_keep(qbase + (int64_t)j * p.ldq, pol);                         \
      } else {                                                                             \
        Pb[BUF][tt] = zero4;                                                               \
        Qb[BUF][tt] = zero4;                                                               \
      }                                                                                    \
    }                                                                                      \
  }
      GN_EK_LOAD(0, 0)
      mbar_wait(&bars->empty[owner], ((g / EK_STAGES) & 1) ^ 1);
#pragma unroll
      for (int sb = 0; sb < 4; ++sb) {
        if (sb + 1 < 4) {
          if ((sb & 1) == 0) GN_EK_LOAD(1, sb + 1) else GN_EK_LOAD(0, sb + 1)
        }
#pragma unroll
        for (int tt = 0; tt < 2; ++tt) {
          const uint4 P = Pb[sb & 1][tt], Q = Qb[sb & 1][tt];
          const int r = rq + 8 * (2 * sb + tt);
          const uint32_t dst = base + (uint32_t)(r * 64 + ((c ^ ((r >> 1) & 3)) << 4));
          asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(dst),
                       "r"(relu_add_f16x2(P.x, Q.x)), "r"(relu_add_f16x2(P.y, Q.y)),
                       "r"(relu_add_f16x2(P.z, Q.z)), "r"(relu_add_f16x2(P.w, Q.w))
                       : "memory");
        }
      }
#undef GN_EK_LOAD
      fence_proxy_async_smem();
      __syncwarp();
      if (lane == 0) mbar_arrive(&bars->full[owner]);
    }
  } else {
    // ================================================================== epilogue
    const int e = warp - (1 + EK_PROD_WARPS);  // 0..7
    const int q = warp & 3;                    // TMEM lane quadrant of this warp
    const int chalf = e >> 2;                  // which 128-column half
    constexpr int CPL = 32 / KP;               // columns per lane after the reduce-scatter
    const int coloff = (lane & (KP - 1)) * CPL;
    uint32_t it = 0;
    for (int64_t tile = blockIdx.x; tile < num_tiles; tile += gridDim.x, ++it) {
      const uint32_t buf = it & 1;
      mbar_wait(&bars->tmem_full[buf], (it >> 1) & 1);
      tc_fence_after();
      const int64_t grp = tile * NPT + (q * 32 + lane) / KP;
      const bool grp_ok = grp < num_groups;
      const int64_t node = grp_ok ? (p.group_node ? (int64_t)__ldg(p.group_node + grp) : grp) : 0;
#pragma unroll 1
      for (int cc = 0; cc < 4; ++cc) {
        const int c0 = chalf * 128 + cc * 32;
        if (c0 >= p.bnp) break;
        float v[32];
        tmem_ld_32x32(tmem_base + buf * 256 + ((uint32_t)(q * 32) << 16) + (uint32_t)c0, v);
#pragma unroll
        for (int i = 0; i < 32; ++i) v[i] = fmaxf(v[i], 0.f);
        // reduce-scatter over the KP lanes (edge rows) of each node
        int n = 32;
#pragma unroll
        for (int mask = KP / 2; mask >= 1; mask >>= 1) {
          n >>= 1;
          const bool up = (lane & mask) != 0;
#pragma unroll
          for (int i = 0; i < 16; ++i) {
            if (i < n) {
              const float send = up ? v[i] : v[i + n];
              const float keep = up ? v[i + n] : v[i];
              v[i] = keep + __shfl_xor_sync(0xffffffffu, send, mask);
            }
          }
        }
        const bool f16 = p.out_f16 != 0;
        if (grp_ok && p.accumulate) {
          const int col = c0 + coloff;
          uint16_t* dst = reinterpret_cast<uint16_t*>(p.out) + node * p.ldo + col;
#pragma unroll
          for (int i = 0; i < CPL; i += 2) {
            if (CPL >= 2 && col + i + 2 <= p.b) {
              red_add2_16(dst + i, v[i], v[i + 1 < CPL ? i + 1 : i], f16);
            } else {
              if (col + i < p.b) red_add1_16(dst + i, v[i], f16);
              if (i + 1 < CPL && col + i + 1 < p.b)
                red_add1_16(dst + i + 1, v[i + 1 < CPL ? i + 1 : i], f16);
            }
          }
          if (p.coords != nullptr) {
#pragma unroll
            for (int i = 0; i < CPL; ++i) {
              const int cidx = col + i;
              const uint32_t word = p.coord_mask[cidx >> 5];
              if (cidx < p.b && ((word >> (cidx & 31)) & 1u)) {
                int rank = __popc(word & ((1u << (cidx & 31)) - 1u));
                for (int w = 0; w < (cidx >> 5); ++w) rank += __popc(p.coord_mask[w]);
                atomicAdd(p.coords + node * p.ncoord + rank, v[i]);
              }
            }
          }
        } else if (grp_ok) {
          const int col = c0 + coloff;
          uint16_t* dst = reinterpret_cast<uint16_t*>(p.out) + node * p.ldo + col;
          if (CPL == 4 && col + 4 <= p.b) {
            uint2 u;
            u.x = pack2_16(v[0], v[1], f16);
            u.y = pack2_16(v[2], v[3], f16);
            *reinterpret_cast<uint2*>(dst) = u;
          } else {
#pragma unroll
            for (int i = 0; i < CPL; ++i)
              if (col + i < p.b) dst[i] = pack1_16(v[i], f16);
          }
          if (p.coords != nullptr) {
#pragma unroll
            for (int i = 0; i < CPL; ++i) {
              const int cidx = col + i;
              const uint32_t word = p.coord_mask[cidx >> 5];
              if (cidx < p.b && ((word >> (cidx & 31)) & 1u)) {
                int rank = __popc(word & ((1u << (cidx & 31)) - 1u));
                for (int w = 0; w < (cidx >> 5); ++w) rank += __popc(p.coord_mask[w]);
                p.coords[node * p.ncoord + rank] = v[i];
              }
            }
          }
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&bars->tmem_empty[buf]);
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc<512>(tmem_base);
}

}  // namespace tc
}  // namespace gn
#include "tc_edge2.cuh"
#include "tc_edge3.cuh"
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

}  // namespace tc
}  // namespace gn

using namespace gn;

static int64_t edge_overflow_cap(const gn_graph* graph, int64_t N, int slots) {
  if (graph->ell_stride > 0) return N * ((graph->ell_stride + slots - 1) / slots - 1);
  return graph->num_edges / slots + 1;
}

// 8 wait-time counters, then a timeline of CTA 0: [warp 0..31][iteration 0..31][4 stamps]
constexpr int EDGE_DBG_WORDS = 8 + 32 * 32 * 4;
static __device__ unsigned long long g_edge_dbg[EDGE_DBG_WORDS];

// diagnostic only (not part of include/graphnet_b200.h): copy out and clear the counters
extern "C" int gn_debug_edge_counters(unsigned long long* host8) {
  GN_CUDA(cudaDeviceSynchronize());
  GN_CUDA(cudaMemcpyFromSymbol(host8, g_edge_dbg, 8 * sizeof(unsigned long long)));
  unsigned long long zero[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  GN_CUDA(cudaMemcpyToSymbol(g_edge_dbg, zero, sizeof(zero)));
  return GN_OK;
}
// diagnostic only: the per-warp timeline of CTA 0 of the last launch (clock64 stamps)
extern "C" int gn_debug_edge_trace(unsigned long long* host, int words) {
  GN_CUDA(cudaDeviceSynchronize());
  if (words > EDGE_DBG_WORDS - 8) words = EDGE_DBG_WORDS - 8;
  GN_CUDA(cudaMemcpyFromSymbol(host, g_edge_dbg, words * sizeof(unsigned long long),
                               8 * sizeof(unsigned long long)));
  return GN_OK;
}

static bool edge_use_v2() {
  // the third-generation kernel (tc_edge3.cuh) is opt-in until it beats the second
  static const bool v3 = getenv("GN_EDGE_V3") != nullptr;
  return !v3;
}
static bool edge_use_v1() {
  static const bool v = getenv("GN_EDGE_V1") != nullptr;  // first-generation kernel (A/B runs)
  return v;
}
// an ELL graph whose stride is a supported slot count is consumed in place: no lists, no 2nd pass
static bool edge_direct_ell(const gn_graph* graph) {
  return !edge_use_v1() && graph->ell_stride > 0 && tc::edge2_supported_slots(graph->ell_stride);
}

extern "C" size_t gn_tc_edge_scratch_bytes(const gn_graph* graph, int64_t N, int slots) {
  if (!graph || slots <= 0) return 0;
  if (edge_direct_ell(graph)) return 16;
  const int64_t n = N > 0 ? N : 1;
  const int64_t cap = edge_overflow_cap(graph, n, slots);
  // ell [N, slots] + ovf_ell [cap, slots] + ovf_node [cap] + counter (16-byte padded)
  return (size_t)(n * slots + cap * slots + cap + 4) * sizeof(int32_t);
}

extern "C" int64_t gn_tc_edge_relu_mask_ld(int64_t N, int slots) {
  if (N <= 0 || !tc::edge2_supported_slots(slots)) return 0;
  return ceil_div(N, (int64_t)(tc::E2_ROWS / slots)) * tc::E2_ROWS;
}

static int tc_edge_mlp_aggregate_impl(const void* P, int64_t ldp, const void* Q, int64_t ldq,
                                      int Kp, const gn_graph* graph, int64_t N, int slots,
                                      const void* W2p, int b, void* out, int64_t ldo,
                                      int out_dtype, float* coords,
                                      const int32_t* coord_cols, int ncoord, void* scratch,
                                      size_t scratch_bytes, uint32_t* relu_mask, int64_t mask_ld,
                                      void* stream) {
  GN_CHECK_ARG(out_dtype == GN_BF16 || out_dtype == GN_F16,
               "gn_tc_edge_mlp_aggregate: out_dtype must be GN_BF16 or GN_F16");
  GN_CHECK_ARG(N >= 0 && b >= 1 && b <= 256, "gn_tc_edge_mlp_aggregate: b=%d not in [1,256]", b);
  GN_CHECK_ARG(Kp >= 32 && Kp % 32 == 0 && Kp <= tc::EK_MAX_CHUNKS * tc::EK_KCH,
               "gn_tc_edge_mlp_aggregate: padded hidden width %d not a multiple of 32 in [32,%d]",
               Kp, tc::EK_MAX_CHUNKS * tc::EK_KCH);
  GN_CHECK_ARG(tc::edge2_supported_slots(slots),
               "gn_tc_edge_mlp_aggregate: slots=%d (must be 8, 9, 16, 17 or 32)", slots);
  GN_CHECK_ARG(!edge_use_v1() || slots == 8 || slots == 16 || slots == 32,
               "gn_tc_edge_mlp_aggregate: the first-generation kernel takes 8, 16 or 32 slots");
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
  const bool direct = edge_direct_ell(graph);
  if (direct) slots = graph->ell_stride;
  if (relu_mask != nullptr) {
    if (!direct || !edge_use_v2()) {
      set_error("gn_tc_edge_mlp_aggregate_mask: needs an ELL graph with 8, 9, 16, 17 or 32 slots "
                "per node and the default (second-generation) kernel");
      return GN_ERR_UNSUPPORTED;
    }
    GN_CHECK_ARG(mask_ld >= gn_tc_edge_relu_mask_ld(N, slots),
                 "gn_tc_edge_mlp_aggregate_mask: mask_ld=%lld too small", (long long)mask_ld);
  }
  const int64_t cap = direct ? 0 : edge_overflow_cap(graph, N, slots);
  int32_t* ell = reinterpret_cast<int32_t*>(scratch);
  int32_t* ovf_ell = ell + N * slots;
  int32_t* ovf_node = ovf_ell + cap * slots;
  int32_t* ovf_count = ovf_node + cap;
  if (!direct) {
    GN_CUDA(cudaMemsetAsync(ovf_count, 0, sizeof(int32_t), st));
    tc::graph_to_groups_kernel<<<(unsigned)ceil_div(N * slots, 256), 256, 0, st>>>(
        g, N, slots, ell, ovf_node, ovf_ell, ovf_count, cap);
    GN_LAUNCH_CHECK();
  }

  CUtensorMap tmW2;
  if (edge_use_v1()) {
    rc = make_tmap_f16_2d(&tmW2, W2p, (uint64_t)bnp, (uint64_t)Kp, (uint64_t)Kp, (uint32_t)bnp,
                          tc::EK_KCH, 64);
    if (rc != GN_OK) return rc;
  }

  p.w2 = W2p;
  p.relu_mask = relu_mask;
  p.mask_ld = mask_ld;
  // GN_EDGE_DEBUG=1: the second-generation kernel accumulates wait-time counters (tc_edge2.cuh)
  // into g_edge_dbg, read back with the undeclared diagnostic export gn_debug_edge_counters()
  static const bool dbg_on = getenv("GN_EDGE_DEBUG") != nullptr;
  p.dbg = nullptr;
  static const bool ws_off = getenv("GN_EDGE_NO_WS") != nullptr;
  p.use_ws = ws_off ? 0 : 1;
  if (dbg_on) GN_CUDA(cudaGetSymbolAddress(&p.dbg, g_edge_dbg));
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
  const bool use_v1 = edge_use_v1();
  const int npt = (use_v1 ? tc::EK_ROWS : tc::E2_ROWS) / slots;

  int dev = 0, sms = 0;
  GN_CUDA(cudaGetDevice(&dev));
  GN_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
  const size_t smem1 = (size_t)p.nch * bnp * 64 + tc::EK_STAGES * tc::EK_STAGE_BYTES +
                       sizeof(tc::EdgeBarriers) + 1024;
  const bool use_v2 = edge_use_v2();
  const size_t smem2 = use_v2 ? tc::edge2_smem_bytes(Kp, slots) : tc::edge3_smem_bytes(Kp);
  for (int pass = 0; pass < 2; ++pass) {
    int64_t max_tiles;
    if (pass == 0) {
      p.ell = direct ? graph->col : ell;
      p.group_node = nullptr;
      p.group_count = nullptr;
      p.accumulate = 0;
      max_tiles = ceil_div(N, npt);
    } else {
      if (cap <= 0) break;
      p.ell = ovf_ell;
      p.group_node = ovf_node;
      p.group_count = ovf_count;
      p.accumulate = 1;
      max_tiles = ceil_div(cap, npt);
    }
    const unsigned grid = (unsigned)(max_tiles < sms ? max_tiles : sms);
#define GN_EDGE_CASE1(KP_)                                                                       \
  {                                                                                              \
    auto kern = tc::edge_mlp_aggregate_kernel<KP_>;                                              \
    GN_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem1)); \
    kern<<<grid, tc::EK_THREADS, smem1, st>>>(tmW2, p);                                          \
  }
#define GN_EDGE_LAUNCH2(KP_, ACC_, F16_)                                                         \
  {                                                                                              \
    auto kern = use_v2 ? tc::edge_mlp_aggregate_v2_kernel<KP_, ACC_, F16_>                       \
                       : tc::edge_mlp_aggregate_v3_kernel<KP_, ACC_, F16_>;                      \
    GN_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem2)); \
    kern<<<grid, use_v2 ? tc::E2_THREADS : tc::E3_THREADS, smem2, st>>>(p);                      \
  }
#define GN_EDGE_CASE2(KP_)                                                                       \
  {                                                                                              \
    if (p.accumulate) {                                                                          \
      if (p.out_f16) GN_EDGE_LAUNCH2(KP_, true, true) else GN_EDGE_LAUNCH2(KP_, true, false)     \
    } else {                                                                                     \
      if (p.out_f16) GN_EDGE_LAUNCH2(KP_, false, true) else GN_EDGE_LAUNCH2(KP_, false, false)   \
    }                                                                                            \
  }
    if (use_v1) {
      if (slots == 8) GN_EDGE_CASE1(8)
      else if (slots == 16) GN_EDGE_CASE1(16)
      else GN_EDGE_CASE1(32)
    } else {
      switch (slots) {
        case 8: GN_EDGE_CASE2(8) break;
        case 9: GN_EDGE_CASE2(9) break;
        case 16: GN_EDGE_CASE2(16) break;
        case 17: GN_EDGE_CASE2(17) break;
        default: GN_EDGE_CASE2(32) break;
      }
    }
#undef GN_EDGE_CASE1
#undef GN_EDGE_CASE2
#undef GN_EDGE_LAUNCH2
#define GN_EDGE_CASE(x)
#undef GN_EDGE_CASE
    GN_LAUNCH_CHECK();
  }
  return GN_OK;
}

extern "C" int gn_tc_edge_mlp_aggregate(const void* P, int64_t ldp, const void* Q, int64_t ldq,
                                        int Kp, const gn_graph* graph, int64_t N, int slots,
                                        const void* W2p, int b, void* out, int64_t ldo,
                                        int out_dtype, float* coords,
                                        const int32_t* coord_cols, int ncoord, void* scratch,
                                        size_t scratch_bytes, void* stream) {
  return tc_edge_mlp_aggregate_impl(P, ldp, Q, ldq, Kp, graph, N, slots, W2p, b, out, ldo,
                                    out_dtype, coords, coord_cols, ncoord, scratch, scratch_bytes,
                                    nullptr, 0, stream);
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
                                    relu_mask, mask_ld, stream);
}

