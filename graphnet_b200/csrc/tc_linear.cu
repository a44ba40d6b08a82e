// bf16 tensor-core Linear (+bias, +activation) on tcgen05 / TMEM, operands staged by TMA.
//
// C[M,N] = act(A[M,K] W[N,K]^T + bias)   A, W bf16 (K contiguous), fp32 accumulate in TMEM.
//
// Used for the node-level GEMMs of the DynEdge path: the factored first layer of every
// EdgeConv MLP (PQ = x [W1a-W1b ; W1b]^T, reference dynedge.py:193-204 via layers.py:67) and
// the post-processing MLP (dynedge.py:222-232, 332).
//
// Persistent kernel, one CTA per SM walking 128 x BN output tiles.  Warp roles: warp 0 = TMA
// producer, warp 1 = tcgen05.mma issuer (owns the TMEM allocation: two 256-column fp32
// accumulators, so the epilogue of tile i overlaps the main loop of tile i+1), warps 2-5 =
// epilogue (one TMEM lane quadrant each).  A/W tiles travel as 128B-swizzled 64-column boxes
// through a 4-stage mbarrier ring; rows/columns beyond M, N, K are zero-filled by the TMA unit.
#include <cuda_fp16.h>
#include <stdlib.h>

#include <atomic>

#include "tc_common.cuh"

namespace gn {

static int make_tmap_16bit_2d(CUtensorMap* map, CUtensorMapDataType dt, const void* base,
                              uint64_t rows, uint64_t cols, uint64_t ld, uint32_t box_rows,
                              uint32_t box_cols, int swizzle_bytes) {
  typedef CUresult (*EncodeFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*,
                               const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                               const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                               CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
  static EncodeFn encode = nullptr;
  if (!encode) {
    void* fn = nullptr;
    cudaDriverEntryPointQueryResult q;
    cudaError_t e = cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &q);
    if (e != cudaSuccess || q != cudaDriverEntryPointSuccess || !fn) {
      set_error("cuTensorMapEncodeTiled is not available from the driver");
      return GN_ERR_CUDA;
    }
    encode = reinterpret_cast<EncodeFn>(fn);
  }
  if ((reinterpret_cast<uintptr_t>(base) & 15) || ((ld * 2) & 15)) {
    set_error("TMA operand must be 16-byte aligned with a row stride that is a multiple of 8 "
              "bf16 elements (base=%p, ld=%llu)", base, (unsigned long long)ld);
    return GN_ERR_INVALID_ARGUMENT;
  }
  cuuint64_t dims[2] = {cols, rows};
  cuuint64_t strides[1] = {ld * 2};
  cuuint32_t box[2] = {box_cols, box_rows};
  cuuint32_t estr[2] = {1, 1};
  CUtensorMapSwizzle sw = swizzle_bytes == 128 ? CU_TENSOR_MAP_SWIZZLE_128B
                        : swizzle_bytes == 64  ? CU_TENSOR_MAP_SWIZZLE_64B
                        : swizzle_bytes == 32  ? CU_TENSOR_MAP_SWIZZLE_32B
                                               : CU_TENSOR_MAP_SWIZZLE_NONE;
  CUresult r = encode(map, dt, 2, const_cast<void*>(base), dims,
                      strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, sw,
                      CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    set_error("cuTensorMapEncodeTiled failed with CUresult %d (rows=%llu cols=%llu ld=%llu box=%ux%u)",
              (int)r, (unsigned long long)rows, (unsigned long long)cols, (unsigned long long)ld,
              box_rows, box_cols);
    return GN_ERR_CUDA;
  }
  return GN_OK;
}

int make_tmap_bf16_2d(CUtensorMap* map, const void* base, uint64_t rows, uint64_t cols,
                      uint64_t ld, uint32_t box_rows, uint32_t box_cols, int swizzle_bytes) {
  return make_tmap_16bit_2d(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, base, rows, cols, ld, box_rows,
                            box_cols, swizzle_bytes);
}
int make_tmap_f16_2d(CUtensorMap* map, const void* base, uint64_t rows, uint64_t cols,
                     uint64_t ld, uint32_t box_rows, uint32_t box_cols, int swizzle_bytes) {
  return make_tmap_16bit_2d(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, base, rows, cols, ld, box_rows,
                            box_cols, swizzle_bytes);
}

namespace tc {

constexpr int LIN_BM = 128;
constexpr int LIN_BK = 64;        // 64 bf16 = one 128-byte swizzle row
constexpr int LIN_STAGES = 8;     // capacity of the barrier arrays; the ring uses `stages` of them
// Epilogue warps: 16 for 16-bit outputs, 8 for fp32.  The epilogue of a 128 x BN tile (TMEM ->
// registers -> bias / activation / pack -> scratch -> coalesced stores) is a long dependent
// chain per warp; with eight warps (two per scheduler) it took ~3900 cycles per tile of the node
// GEMMs against ~1400 of tensor work and paced the kernel (timeline of CTA 0, profiling build).
// Sixteen warps halve the chunks per warp and give every scheduler four warps to interleave.
// fp32 outputs keep eight: their scratch rows are twice as wide, and the shared memory is better
// spent on the resident W tile.
template <typename OutT>
__host__ __device__ constexpr int lin_epi_warps() { return sizeof(OutT) == 4 ? 8 : 16; }
template <typename OutT>
__host__ __device__ constexpr int lin_threads() { return 32 * (2 + lin_epi_warps<OutT>()); }
constexpr int LIN_MAX_N = 2048;   // bias staged in shared memory
constexpr int LIN_A_BYTES = LIN_BM * LIN_BK * 2;  // 16 KB
// per epilogue warp: a 32-row x 32-column chunk on its way from "lane = row" (tcgen05.ld) to
// "lanes = consecutive 16-byte pieces of a row" (coalesced global stores); rows padded by 16 B
// fp32: rows of 128 B padded by 16 B.  16-bit: rows of 64 B, not padded, the 16-byte piece index
// XOR-ed with (row >> 1) & 3 - conflict free both for the row-wise writes and for the reads
// (the padded layout had 2-way conflicts on the 16-bit reads: 41 % of the kernel's wavefronts)
template <typename OutT>
__host__ __device__ constexpr int lin_scr_pitch() { return sizeof(OutT) == 4 ? 32 * 4 + 16 : 64; }
template <typename OutT>
__host__ __device__ constexpr int lin_scr_bytes() { return 32 * lin_scr_pitch<OutT>(); }
constexpr int LIN_SCR_PITCH = 32 * 4 + 16;       // fp32 (the pooling epilogue reads it by column)

struct __align__(16) LinBarriers {
  uint64_t full[LIN_STAGES];
  uint64_t empty[LIN_STAGES];
  uint64_t tmem_full[2];
  uint64_t tmem_empty[2];
  uint64_t w_full;
  uint32_t tmem_base;
};

template <typename OutT>
__device__ __forceinline__ uint32_t pack2(float a, float b);
template <>
__device__ __forceinline__ uint32_t pack2<__nv_bfloat16>(float a, float b) {
  __nv_bfloat162 p = __floats2bfloat162_rn(a, b);
  return *reinterpret_cast<uint32_t*>(&p);
}
template <>
__device__ __forceinline__ uint32_t pack2<__half>(float a, float b) {  // saturating, one CVT
  uint32_t r;
  asm("cvt.rn.satfinite.f16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(b), "f"(a));  // hi <- b, lo <- a
  return r;
}

// one thread stores 32 consecutive columns of its row; 16-byte vector stores for every full
// group of 8 (16-bit) / 4 (fp32) columns, scalar stores only for a ragged tail
template <typename OutT>
__device__ __forceinline__ void store_row32(OutT* dst, const float (&v)[32], int nvalid) {
  const bool aligned = (reinterpret_cast<uintptr_t>(dst) & 15) == 0;
  if (aligned && nvalid >= 32) {
    // common case, branch-free body: 32 columns as 16-byte stores
    if constexpr (sizeof(OutT) == 4) {
#pragma unroll
      for (int j = 0; j < 32; j += 4)
        *reinterpret_cast<float4*>(dst + j) = make_float4(v[j], v[j + 1], v[j + 2], v[j + 3]);
    } else {
#pragma unroll
      for (int j = 0; j < 32; j += 8) {
        uint4 u;
        u.x = pack2<OutT>(v[j], v[j + 1]);
        u.y = pack2<OutT>(v[j + 2], v[j + 3]);
        u.z = pack2<OutT>(v[j + 4], v[j + 5]);
        u.w = pack2<OutT>(v[j + 6], v[j + 7]);
        *reinterpret_cast<uint4*>(dst + j) = u;
      }
    }
    return;
  }
  // ragged tail of a tile / unaligned rows: 16-byte stores for whole groups, scalars after
  if constexpr (sizeof(OutT) == 4) {
#pragma unroll
    for (int j = 0; j < 32; j += 4) {
      if (aligned && j + 4 <= nvalid) {
        *reinterpret_cast<float4*>(dst + j) = make_float4(v[j], v[j + 1], v[j + 2], v[j + 3]);
      } else {
#pragma unroll
        for (int t = 0; t < 4; ++t)
          if (j + t < nvalid) dst[j + t] = v[j + t];
      }
    }
  } else {
    uint16_t* d16 = reinterpret_cast<uint16_t*>(dst);
#pragma unroll
    for (int j = 0; j < 32; j += 8) {
      if (aligned && j + 8 <= nvalid) {
        uint4 u;
        u.x = pack2<OutT>(v[j], v[j + 1]);
        u.y = pack2<OutT>(v[j + 2], v[j + 3]);
        u.z = pack2<OutT>(v[j + 4], v[j + 5]);
        u.w = pack2<OutT>(v[j + 6], v[j + 7]);
        *reinterpret_cast<uint4*>(dst + j) = u;
      } else {
#pragma unroll
        for (int t = 0; t < 8; t += 2) {
          const uint32_t pr = pack2<OutT>(v[j + t], v[j + t + 1]);
          if (j + t < nvalid) d16[j + t] = (uint16_t)(pr & 0xffffu);
          if (j + t + 1 < nvalid) d16[j + t + 1] = (uint16_t)(pr >> 16);
        }
      }
    }
  }
}

#ifdef GN_DEBUG_BUILD
// profiling builds (GN_TC_DEBUG=1): clock64 stamps of CTA 0, [warp 0..9][item 0..15][4]
__device__ unsigned long long g_lin_trace[10 * 16 * 4];
#endif

// Segmented-pooling epilogue (POOL): instead of storing the [M, N] output, every column's min, max
// and sum over the rows of one event (rows = nodes, sorted by event) go straight into
// acc[3][B][N] - the [N, 256] post-processing output and the pooling kernel that re-read it
// (63 MB each way at C2) disappear.  Values are >= 0 (ReLU), so their bit patterns order like
// unsigned integers: min / max are plain atomicMin / atomicMax on uint32 (acc_min starts as
// 0xFFFFFFFF, acc_max as 0).  The sum is accumulated in 64-bit FIXED POINT (units of 2^-32, every
// element converted before it is added): integer addition is associative, so the result does not
// depend on the order in which tiles, warps and atomics happen to run, nor on where an event's
// rows fall relative to the 32-row chunks - the same event pools to the same bits in any batch.
struct PoolEpilogue {
  const int64_t* batch;   // [M] event id of every row, non-decreasing
  const int32_t* ptr;     // [B + 1] first row of every event
  float* acc;             // [4][B][N] words: min bits, max bits, sum (int64 fixed point: 2 planes)
  int64_t B;
};
constexpr float POOL_FIX_SCALE = 4294967296.f;        // 2^32
constexpr float POOL_FIX_MAX = 1073741824.f;          // elements are clamped to 2^30

template <typename OutT, int ACT, bool POOL = false>
__global__ void __launch_bounds__(lin_threads<OutT>(), 1) tc_linear_kernel(
    const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmW,
    const float* __restrict__ bias, OutT* __restrict__ C, int64_t ldc, int64_t M, int N, int K,
    int BN, int num_n_tiles, int64_t num_tiles, uint32_t idesc, int stages, int splits,
    int accumulate, int coalesce_in, int wres, const PoolEpilogue pool) {
  constexpr int LIN_EPI_WARPS = lin_epi_warps<OutT>();
  constexpr int LIN_THREADS = lin_threads<OutT>();
  constexpr int NPAR = LIN_EPI_WARPS / 4;          // warps sharing a TMEM lane quadrant
  constexpr int SCR_PITCH = lin_scr_pitch<OutT>();
  constexpr int SCR_BYTES = lin_scr_bytes<OutT>();
  const bool coalesce = coalesce_in != 0 && accumulate == 0;
#ifdef GN_DEBUG_BUILD
  const bool traced = (coalesce_in & 4) && blockIdx.x == 0 && (threadIdx.x & 31) == 0;
  auto stamp = [&](uint32_t i, int k) {
    if (traced && i < 16 && (threadIdx.x >> 5) < 10)
      g_lin_trace[((threadIdx.x >> 5) * 16 + i) * 4 + k] = clock64();
  };
#else
  auto stamp = [](uint32_t, int) {};
#endif
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  // 1024-byte alignment for the 128B swizzle atoms
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) &
                                             ~uintptr_t(1023));
  const int b_bytes = BN * LIN_BK * 2;
  const int nk_all = (K + LIN_BK - 1) / LIN_BK;
  // wres ("weights resident"): this CTA works on ONE column tile of W for its whole life and
  // keeps that tile (all K blocks) in shared memory; only A streams through the ring.  Without
  // it every 128-row tile re-reads its W tile from L2, which is more bytes than A itself
  // (ncu, node GEMM of C2: 157 MB of TMA loads for 63 MB of A) and the kernel sits at the
  // L2->SM limit with the tensor pipe 27 % busy.
  uint8_t* w_res = smem;                                   // [nk_all][BN x 128 B] when wres
  uint8_t* ring = smem + (wres ? nk_all * b_bytes : 0);
  const int stage_bytes = LIN_A_BYTES + (wres ? 0 : b_bytes);
  LinBarriers* bars = reinterpret_cast<LinBarriers*>(ring + stages * stage_bytes);
  float* s_bias = reinterpret_cast<float*>(bars + 1);  // [num_n_tiles * BN]
  uint8_t* s_scr = reinterpret_cast<uint8_t*>(s_bias + num_n_tiles * BN);  // one scratch per epilogue warp
  s_scr = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(s_scr) + 15) & ~uintptr_t(15));

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int nk = (K + LIN_BK - 1) / LIN_BK;
  for (int n = threadIdx.x; n < num_n_tiles * BN; n += LIN_THREADS)
    s_bias[n] = (bias != nullptr && n < N) ? bias[n] : 0.f;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmW);
#pragma unroll
    for (int s = 0; s < LIN_STAGES; ++s) {
      mbar_init(&bars->full[s], 1);
      mbar_init(&bars->empty[s], 1);
    }
#pragma unroll
    for (int b = 0; b < 2; ++b) {
      mbar_init(&bars->tmem_full[b], 1);
      mbar_init(&bars->tmem_empty[b], LIN_EPI_WARPS);
    }
    mbar_init(&bars->w_full, 1);
    fence_mbar_init();
  }
  if (warp == 1) tmem_alloc<512>(&bars->tmem_base);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = bars->tmem_base;

  // Persistent: this CTA walks work items blockIdx.x, +gridDim.x, ...; item = (tile, K split),
  // split fastest; consecutive tile ids share the A row block (n fastest), so concurrently
  // running CTAs hit the same A tile in L2.  splits > 1 (weight gradients: few tiles, K = rows of
  // the batch) gives every item the K blocks [kb0, kb1) and the epilogue adds with red.add.
  const int64_t num_items = num_tiles * splits;
  // i-th work item of this CTA -> (tile t, K split sp); false when the CTA is done
  const int cta_n = blockIdx.x % num_n_tiles, cta_m0 = blockIdx.x / num_n_tiles;
  const int cta_mstep = gridDim.x / num_n_tiles;           // wres: grid is a multiple of n tiles
  const int64_t num_m_tiles = num_tiles / num_n_tiles;
  auto work_item = [&](uint32_t i, int64_t& t, int& sp) -> bool {
    if (wres) {
      const int64_t mt = cta_m0 + (int64_t)i * cta_mstep;
      if (mt >= num_m_tiles) return false;
      t = mt * num_n_tiles + cta_n;
      sp = 0;
      return true;
    }
    const int64_t w = blockIdx.x + (int64_t)i * gridDim.x;
    if (w >= num_items) return false;
    t = w / splits;
    sp = (int)(w % splits);
    return true;
  };
  if (warp == 0) {
    // ---------------------------------------------------------- TMA producer
    if (elect_one()) {
      uint32_t g = 0;
      if (wres) {  // the CTA's W tile, once
        mbar_arrive_expect_tx(&bars->w_full, (uint32_t)(nk_all * b_bytes));
        for (int kb = 0; kb < nk_all; ++kb)
          tma_load_2d(w_res + kb * b_bytes, &tmW, &bars->w_full, kb * LIN_BK, cta_n * BN);
      }
      int64_t t;
      int sp;
      for (uint32_t i = 0; work_item(i, t, sp); ++i) {
        const int kb0 = (int)((int64_t)sp * nk / splits), kb1 = (int)((int64_t)(sp + 1) * nk / splits);
        const int32_t m0 = (int32_t)((t / num_n_tiles) * LIN_BM);
        const int32_t n0 = (int32_t)(t % num_n_tiles) * BN;
        stamp(i, 0);
        for (int kb = kb0; kb < kb1; ++kb, ++g) {
          const uint32_t s = g % stages;
          mbar_wait(&bars->empty[s], ((g / stages) & 1) ^ 1);
          if (kb == kb0) stamp(i, 1);
          if (kb == kb1 - 1) stamp(i, 2);
          uint8_t* a_dst = ring + s * stage_bytes;
          mbar_arrive_expect_tx(&bars->full[s], (uint32_t)stage_bytes);
          tma_load_2d(a_dst, &tmA, &bars->full[s], kb * LIN_BK, m0);
          if (!wres) tma_load_2d(a_dst + LIN_A_BYTES, &tmW, &bars->full[s], kb * LIN_BK, n0);
        }
      }
    }
  } else if (warp == 1) {
    // ------------------------------------------------------------ MMA issuer
    uint32_t g = 0;
    if (wres) {
      mbar_wait(&bars->w_full, 0);
      tc_fence_after();
    }
    int64_t t;
    int sp;
    for (uint32_t it = 0; work_item(it, t, sp); ++it) {
      const int kb0 = (int)((int64_t)sp * nk / splits), kb1 = (int)((int64_t)(sp + 1) * nk / splits);
      const uint32_t buf = it & 1;
      stamp(it, 0);
      mbar_wait(&bars->tmem_empty[buf], ((it >> 1) & 1) ^ 1);
      stamp(it, 1);
      tc_fence_after();
      const uint32_t tmem_d = tmem_base + buf * 256;
      for (int kb = kb0; kb < kb1; ++kb, ++g) {
        const uint32_t s = g % stages;
        mbar_wait(&bars->full[s], (g / stages) & 1);
        if (kb == kb0) stamp(it, 2);
        if (kb == kb1 - 1) stamp(it, 3);
        tc_fence_after();
        if (elect_one()) {
          const uint32_t a_addr = smem_u32(ring + s * stage_bytes);
          const uint32_t b_addr = wres ? smem_u32(w_res + kb * b_bytes) : a_addr + LIN_A_BYTES;
          const uint64_t adesc = make_kmajor_desc<128>(a_addr);
          const uint64_t bdesc = make_kmajor_desc<128>(b_addr);
#pragma unroll
          for (int k4 = 0; k4 < LIN_BK / 16; ++k4) {
            if (coalesce_in & 8) break;   // profiling: no tensor work, data path only
            // advance 16 bf16 = 32 bytes along K inside the swizzle row: +2 in (addr >> 4) units
            mma_bf16_ss(tmem_d, adesc + (uint64_t)(k4 * 2), bdesc + (uint64_t)(k4 * 2), idesc,
                        ((kb - kb0) | k4) != 0);
          }
          mma_commit(&bars->empty[s]);
          if (kb == kb1 - 1) mma_commit(&bars->tmem_full[buf]);
        }
        __syncwarp();
      }
    }
  } else {
    // -------------------------------------------------------------- epilogue
    const int q = warp & 3;  // TMEM lane quadrant this warp may read
    const int par = (warp - 2) >> 2;  // the two warps of a quadrant take alternate 32-col chunks
    int64_t t;
    int sp;
    for (uint32_t it = 0; work_item(it, t, sp); ++it) {
      const uint32_t buf = it & 1;
      const int64_t m0 = (t / num_n_tiles) * LIN_BM;
      const int n0 = (int)(t % num_n_tiles) * BN;
      stamp(it, 0);
      mbar_wait(&bars->tmem_full[buf], (it >> 1) & 1);
      stamp(it, 1);
      tc_fence_after();
      const int64_t row = m0 + q * 32 + lane;
      // The warp's chunks are c = par*32, +64, ...; the tcgen05.ld of chunk i+1 is issued before
      // chunk i is processed (two register sets), so TMEM latency overlaps the arithmetic,
      // the transpose and the stores.
      const uint32_t t_addr = tmem_base + buf * 256 + ((uint32_t)(q * 32) << 16);
      auto chunk_ok = [&](int c) { return c < BN && n0 + c < N; };
      uint32_t ra[32];
      [[maybe_unused]] uint32_t rb[32];
      auto process = [&](uint32_t (&raw)[32], const int c) {
        const int n = n0 + c;
        float v[32];
#pragma unroll
        for (int j = 0; j < 32; ++j) v[j] = __uint_as_float(raw[j]);
        const uint32_t b4 = smem_u32(s_bias + n);
#pragma unroll
        for (int j = 0; j < 32; j += 4) {
          const uint4 b = lds128(b4 + j * 4);  // same address in every lane: one broadcast wavefront
          v[j] = activate<ACT>(v[j] + __uint_as_float(b.x));
          v[j + 1] = activate<ACT>(v[j + 1] + __uint_as_float(b.y));
          v[j + 2] = activate<ACT>(v[j + 2] + __uint_as_float(b.z));
          v[j + 3] = activate<ACT>(v[j + 3] + __uint_as_float(b.w));
        }
        if constexpr (POOL) {
          // registers (lane = row) -> this warp's scratch; then lane = COLUMN walks the chunk's
          // 32 rows (consecutive nodes) keeping min / max / sum of the current event and
          // flushes them at every event boundary and at the end of the chunk
          uint8_t* scr = s_scr + (warp - 2) * SCR_BYTES;
          const uint32_t mine = smem_u32(scr) + (uint32_t)lane * SCR_PITCH;
#pragma unroll
          for (int j = 0; j < 32; j += 4)
            sts128(mine + j * 4, make_uint4(__float_as_uint(v[j]), __float_as_uint(v[j + 1]),
                                            __float_as_uint(v[j + 2]), __float_as_uint(v[j + 3])));
          __syncwarp();
          const int64_t node0 = m0 + q * 32;
          const int rows = (int)min((int64_t)32, M - node0);
          const int col = n + lane;
          if (rows > 0) {
            int64_t e = __ldg(pool.batch + node0);
            const int64_t plane = pool.B * (int64_t)N;
            const uint32_t col_addr = smem_u32(scr) + (uint32_t)lane * 4u;
            auto ld_row = [&](int r) {   // this lane's column of scratch row r
              float f;
              asm volatile("ld.shared.f32 %0, [%1];" : "=f"(f) : "r"(col_addr + (uint32_t)r * LIN_SCR_PITCH));
              return f;
            };
            auto fix = [](float f) {     // fixed point, units of 2^-32
              return (unsigned long long)__float2ll_rn(fminf(f, POOL_FIX_MAX) * POOL_FIX_SCALE);
            };
            int r = 0;
            while (r < rows) {
              // rows [r, rend) of the chunk belong to event e (events without rows are skipped)
              int64_t end = __ldg(pool.ptr + e + 1);
              while (node0 + r >= end && e + 1 < pool.B) { ++e; end = __ldg(pool.ptr + e + 1); }
              const int rend = (int)min((int64_t)rows, end - node0);
              // four independent chains (the row loop is latency bound otherwise)
              float mn[4], mx[4];
              unsigned long long sm[4];
#pragma unroll
              for (int k = 0; k < 4; ++k) { mn[k] = __int_as_float(0x7f800000); mx[k] = 0.f; sm[k] = 0ull; }
              for (; r + 4 <= rend; r += 4) {
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                  const float xv = ld_row(r + k);
                  mn[k] = fminf(mn[k], xv);
                  mx[k] = fmaxf(mx[k], xv);
                  sm[k] += fix(xv);
                }
              }
              for (; r < rend; ++r) {
                const float xv = ld_row(r);
                mn[0] = fminf(mn[0], xv);
                mx[0] = fmaxf(mx[0], xv);
                sm[0] += fix(xv);
              }
              if (col < N) {
                float* a = pool.acc + e * N + col;
                atomicMin(reinterpret_cast<unsigned int*>(a),
                          __float_as_uint(fminf(fminf(mn[0], mn[1]), fminf(mn[2], mn[3]))));
                atomicMax(reinterpret_cast<unsigned int*>(a + plane),
                          __float_as_uint(fmaxf(fmaxf(mx[0], mx[1]), fmaxf(mx[2], mx[3]))));
                atomicAdd(reinterpret_cast<unsigned long long*>(pool.acc + 2 * plane) + e * N + col,
                          (sm[0] + sm[1]) + (sm[2] + sm[3]));
              }
              if (r < rows) ++e;
            }
          }
          __syncwarp();
          return;
        }
        if (row < M) {
          if constexpr (sizeof(OutT) == 4) {
            if (accumulate) {  // C += : fp32 reductions in L2 (split K, or adding onto a gradient)
              const int nv = min(N - n, BN - c);
              float* dst = reinterpret_cast<float*>(C) + row * ldc + n;
#pragma unroll
              for (int j = 0; j < 32; ++j)
                if (j < nv) atomicAdd(dst + j, v[j]);
              return;
            }
          }
          if (!coalesce) store_row32<OutT>(C + row * ldc + n, v, min(N - n, BN - c));
        }
        if (coalesce) {
          // registers (lane = row) -> this warp's scratch -> 16-byte pieces, lanes along the row:
          // one store instruction covers 32/P rows x (P x 16) contiguous bytes instead of 32 rows
          // x 16 bytes
          constexpr int P = (32 * (int)sizeof(OutT)) / 16;   // pieces per 32-column row: 4 or 8
          constexpr int EPP = 16 / (int)sizeof(OutT);        // elements per piece
          const uint32_t scr = smem_u32(s_scr + (warp - 2) * SCR_BYTES);
          const uint32_t mine = scr + (uint32_t)lane * SCR_PITCH;
          if constexpr (sizeof(OutT) == 4) {
#pragma unroll
            for (int j = 0; j < 32; j += 4)
              sts128(mine + j * 4, make_uint4(__float_as_uint(v[j]), __float_as_uint(v[j + 1]),
                                              __float_as_uint(v[j + 2]), __float_as_uint(v[j + 3])));
          } else {
            const int wsw = (lane >> 1) & 3;   // swizzle of this lane's row
#pragma unroll
            for (int j = 0; j < 32; j += 8) {
              uint4 u;
              u.x = pack2<OutT>(v[j], v[j + 1]);
              u.y = pack2<OutT>(v[j + 2], v[j + 3]);
              u.z = pack2<OutT>(v[j + 4], v[j + 5]);
              u.w = pack2<OutT>(v[j + 6], v[j + 7]);
              sts128(mine + (uint32_t)(((j >> 3) ^ wsw) << 4), u);
            }
          }
          __syncwarp();
          const int nv = min(N - n, BN - c);                 // valid columns of this chunk
          const int piece = lane % P, rsub = lane / P;
#pragma unroll
          for (int i = 0; i < P; ++i) {
            const int r = i * (32 / P) + rsub;
            const int64_t grow = m0 + q * 32 + r;
            if (grow < M && piece * EPP < nv && !(coalesce_in & 2)) {
              const int psw = sizeof(OutT) == 4 ? piece : (piece ^ ((r >> 1) & 3));
              const uint4 u = lds128(scr + (uint32_t)(r * SCR_PITCH + psw * 16));
              OutT* dst = C + grow * ldc + n + piece * EPP;
              if ((piece + 1) * EPP <= nv) {
                *reinterpret_cast<uint4*>(dst) = u;
              } else {
                const OutT* e = reinterpret_cast<const OutT*>(&u);
#pragma unroll
                for (int t = 0; t < EPP; ++t)
                  if (piece * EPP + t < nv) dst[t] = e[t];
              }
            }
          }
          __syncwarp();
        }
      };
      {
        // this warp's chunks: c = par*32, + 32*NPAR, ...
        constexpr int CSTEP = 32 * NPAR;
        int c = par * 32;
        if (chunk_ok(c) && !(coalesce_in & 16)) {   // (16: profiling without the epilogue body)
          if constexpr (NPAR <= 2) {
            // two register sets: the tcgen05.ld of the next chunk is in flight while this one
            // is processed
            tmem_ld_32x32_issue(t_addr + (uint32_t)c, ra);
            for (;;) {
              tmem_ld_wait(ra);
              const bool more1 = chunk_ok(c + CSTEP);
              if (more1) tmem_ld_32x32_issue(t_addr + (uint32_t)(c + CSTEP), rb);
              process(ra, c);
              if (!more1) break;
              tmem_ld_wait(rb);
              const bool more2 = chunk_ok(c + 2 * CSTEP);
              if (more2) tmem_ld_32x32_issue(t_addr + (uint32_t)(c + 2 * CSTEP), ra);
              process(rb, c + CSTEP);
              if (!more2) break;
              c += 2 * CSTEP;
            }
          } else {
            // four warps per quadrant: the other warps of the scheduler cover the TMEM latency
            for (; chunk_ok(c); c += CSTEP) {
              tmem_ld_32x32_issue(t_addr + (uint32_t)c, ra);
              tmem_ld_wait(ra);
              process(ra, c);
            }
          }
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&bars->tmem_empty[buf]);
      stamp(it, 2);
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc<512>(tmem_base);
}

template <typename OutT>
static int launch_tc_linear(const CUtensorMap& tmA, const CUtensorMap& tmW, const float* bias,
                            void* C, int64_t ldc, int64_t M, int N, int K, int BN, int act,
                            int ab_dtype, cudaStream_t st, int splits = 1, int accumulate = 0,
                            const PoolEpilogue* pool_in = nullptr) {
  PoolEpilogue pool;
  memset(&pool, 0, sizeof(pool));
  if (pool_in) pool = *pool_in;
  const uint32_t idesc =
      ab_dtype == GN_F16 ? make_idesc_f16(LIN_BM, BN) : make_idesc_bf16(LIN_BM, BN);
  const int num_n_tiles = (int)ceil_div(N, BN);
  const size_t stage_bytes = (size_t)LIN_A_BYTES + (size_t)BN * LIN_BK * 2;
  const size_t fixed = sizeof(LinBarriers) + (size_t)num_n_tiles * BN * sizeof(float) + 32 + 1024 +
                       (size_t)lin_epi_warps<OutT>() * lin_scr_bytes<OutT>() + 16;
  // 16-byte stores need C rows and the tile's first column on 16-byte boundaries
  static const bool no_coalesce = getenv("GN_TC_NO_COALESCE") != nullptr;
#ifdef GN_DEBUG_BUILD   // ablation / trace switches exist in profiling builds only
  static const bool hack_nostore = getenv("GN_TC_HACK_NOSTORE") != nullptr;
  static const bool dbg_trace = getenv("GN_TC_DEBUG") != nullptr;
  static const int hack_bits = getenv("GN_TC_HACK") ? atoi(getenv("GN_TC_HACK")) : 0;
#else
  constexpr bool hack_nostore = false, dbg_trace = false;
  constexpr int hack_bits = 0;
#endif
  const int coalesce = ((hack_nostore ? 3 : 1) | (dbg_trace ? 4 : 0) | (hack_bits & 24)) * (int)(!no_coalesce && (reinterpret_cast<uintptr_t>(C) % 16 == 0) &&
                       ((ldc * (int64_t)sizeof(OutT)) % 16 == 0) &&
                       ((BN * (int)sizeof(OutT)) % 16 == 0));
  int stages = (int)((227 * 1024 - fixed) / stage_bytes);  // as deep as shared memory allows
  stages = stages > LIN_STAGES ? LIN_STAGES : stages;
  size_t smem = (size_t)stages * stage_bytes + fixed;
  // weights-resident variant: one K-complete W tile per CTA + a ring of A-only stages (>= 4)
  int dev0 = 0, sms0 = 0;
  GN_CUDA(cudaGetDevice(&dev0));
  {
    static std::atomic<int> sm_count[64];   // per device, fetched once
    sms0 = sm_count[dev0 & 63].load(std::memory_order_relaxed);
    if (sms0 == 0) {
      GN_CUDA(cudaDeviceGetAttribute(&sms0, cudaDevAttrMultiProcessorCount, dev0));
      sm_count[dev0 & 63].store(sms0, std::memory_order_relaxed);
    }
  }
  static const bool no_wres = getenv("GN_TC_NO_WRES") != nullptr;
  const size_t w_bytes = (size_t)ceil_div(K, LIN_BK) * BN * LIN_BK * 2;
  int wres = 0;
  if (!no_wres && splits == 1 && !accumulate && num_n_tiles <= sms0 &&
      w_bytes + 4 * (size_t)LIN_A_BYTES + fixed <= 227 * 1024 &&
      ceil_div(M, LIN_BM) >= 2 * (sms0 / num_n_tiles)) {
    wres = 1;
    stages = (int)((227 * 1024 - fixed - w_bytes) / LIN_A_BYTES);
    stages = stages > LIN_STAGES ? LIN_STAGES : stages;
    smem = w_bytes + (size_t)stages * LIN_A_BYTES + fixed;
  }
  const int64_t num_tiles = ceil_div(M, LIN_BM) * num_n_tiles;
  const int dev = dev0, sms = sms0;
  const int nkb = (int)ceil_div(K, LIN_BK);
  if (splits == 0) {  // automatic: enough items to fill the machine, at least 4 K blocks each
    splits = (int)(sms / (num_tiles < sms ? num_tiles : sms));
    if (splits > nkb / 4) splits = nkb / 4;
  }
  if (splits > nkb) splits = nkb;
  if (splits < 1) splits = 1;
  const int64_t num_items = num_tiles * splits;
  unsigned grid = (unsigned)(num_items < sms ? num_items : sms);
  if (wres) grid = (unsigned)((sms / num_n_tiles) * num_n_tiles);
  /* the opt-in to > 48 KB of dynamic shared memory is per function and device: once */        \
#define GN_TC_LIN_LAUNCH(KERN_)                                                                 \
  {                                                                                             \
    auto kern = KERN_;                                                                          \
    static std::atomic<unsigned long long> ready_;                                              \
    if (!((ready_.load(std::memory_order_relaxed) >> (dev & 63)) & 1ull)) {                     \
      GN_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize,           \
                                   227 * 1024));                                                \
      ready_.fetch_or(1ull << (dev & 63), std::memory_order_relaxed);                           \
    }                                                                                           \
    kern<<<grid, lin_threads<OutT>(), smem, st>>>(tmA, tmW, bias, reinterpret_cast<OutT*>(C), ldc, M, N, \
                                          K, BN, num_n_tiles, num_tiles, idesc, stages, splits,  \
                                          accumulate, coalesce, wres, pool);                    \
  }
#define GN_TC_LIN_CASE(A_) GN_TC_LIN_LAUNCH((tc_linear_kernel<OutT, A_, false>))
  if (pool.acc != nullptr) {
    if constexpr (sizeof(OutT) == 4) {
      if (act != GN_ACT_RELU) {
        set_error("gn_tc_linear_pool: the pooling epilogue needs ReLU outputs");
        return GN_ERR_UNSUPPORTED;
      }
      GN_TC_LIN_LAUNCH((tc_linear_kernel<float, ACT_RELU, true>))
      GN_LAUNCH_CHECK();
      return GN_OK;
    }
  }
  switch (act) {
    case GN_ACT_NONE: GN_TC_LIN_CASE(ACT_NONE) break;
    case GN_ACT_RELU: GN_TC_LIN_CASE(ACT_RELU) break;
    case GN_ACT_GELU: GN_TC_LIN_CASE(ACT_GELU) break;
    default: set_error("unknown activation %d", act); return GN_ERR_INVALID_ARGUMENT;
  }
#undef GN_TC_LIN_CASE
#undef GN_TC_LIN_LAUNCH
  GN_LAUNCH_CHECK();
  return GN_OK;
}

}  // namespace tc

static size_t tc_fixed_smem(int N, int BN, int out_bytes) {
  const int num_n_tiles = (N + BN - 1) / BN;
  const size_t scr = out_bytes == 4
                         ? (size_t)tc::lin_epi_warps<float>() * tc::lin_scr_bytes<float>()
                         : (size_t)tc::lin_epi_warps<__half>() * tc::lin_scr_bytes<__half>();
  return sizeof(tc::LinBarriers) + (size_t)num_n_tiles * BN * sizeof(float) + 32 + 1024 + scr + 16;
}
// does a K-complete W tile of width BN fit next to a ring of at least four A stages?
static bool tc_wres_fits(int N, int K, int BN, int out_bytes) {
  const size_t w_bytes = (size_t)ceil_div(K, tc::LIN_BK) * BN * tc::LIN_BK * 2;
  return w_bytes + 4 * (size_t)tc::LIN_A_BYTES + tc_fixed_smem(N, BN, out_bytes) <= 227 * 1024;
}

int tc_pick_bn(int N, int bn_max_arg = 256) {
  // largest tile width <= 256 (multiple of 16) that minimises padding
  static const char* env = getenv("GN_TC_BN_MAX");
  const int bn_max = env ? atoi(env) : bn_max_arg;
  int best = bn_max, best_waste = 1 << 30;
  for (int bn = bn_max; bn >= 64; bn -= 16) {
    int tiles = (N + bn - 1) / bn;
    int waste = tiles * bn - N;
    if (waste < best_waste) { best = bn; best_waste = waste; }
  }
  return best;
}

}  // namespace gn

using namespace gn;

extern "C" int gn_tc_linear(const void* A, int64_t lda, const void* W, int64_t ldw, int ab_dtype,
                            const float* bias, void* C, int64_t ldc, int c_dtype, int64_t M,
                            int N, int K, int act, void* stream) {
  GN_CHECK_ARG(M >= 0 && N >= 1 && K >= 1 && lda >= K && ldw >= K && ldc >= N,
               "gn_tc_linear: bad sizes");
  if (M == 0) return GN_OK;
  GN_CHECK_ARG(A && W && C, "gn_tc_linear: null pointer");
  GN_CHECK_ARG(ab_dtype == GN_BF16 || ab_dtype == GN_F16, "gn_tc_linear: bad ab_dtype");
  GN_CHECK_ARG(N <= tc::LIN_MAX_N, "gn_tc_linear: N=%d exceeds %d", N, tc::LIN_MAX_N);
  GN_CHECK_ARG(c_dtype == GN_F32 || c_dtype == GN_BF16 || c_dtype == GN_F16,
               "gn_tc_linear: bad c_dtype");
  int BN = tc_pick_bn(N);
  const int out_bytes = c_dtype == GN_F32 ? 4 : 2;
  // a narrower tile whose K-complete W fits in shared memory beats a wide one that streams W
  if (!tc_wres_fits(N, K, BN, out_bytes) && M >= 4096) {
    const int bn2 = tc_pick_bn(N, 128);
    if (tc_wres_fits(N, K, bn2, out_bytes)) BN = bn2;
  }
  CUtensorMap tmA, tmW;
  auto mk = ab_dtype == GN_F16 ? make_tmap_f16_2d : make_tmap_bf16_2d;
  int rc = mk(&tmA, A, (uint64_t)M, (uint64_t)K, (uint64_t)lda, tc::LIN_BM, tc::LIN_BK, 128);
  if (rc != GN_OK) return rc;
  rc = mk(&tmW, W, (uint64_t)N, (uint64_t)K, (uint64_t)ldw, (uint32_t)BN, tc::LIN_BK, 128);
  if (rc != GN_OK) return rc;
  cudaStream_t st = as_stream(stream);
  if (c_dtype == GN_F32)
    return tc::launch_tc_linear<float>(tmA, tmW, bias, C, ldc, M, N, K, BN, act, ab_dtype, st);
  if (c_dtype == GN_F16)
    return tc::launch_tc_linear<__half>(tmA, tmW, bias, C, ldc, M, N, K, BN, act, ab_dtype, st);
  return tc::launch_tc_linear<__nv_bfloat16>(tmA, tmW, bias, C, ldc, M, N, K, BN, act, ab_dtype,
                                             st);
}

extern "C" int gn_tc_linear_pool(const void* A, int64_t lda, const void* W, int64_t ldw,
                                 int ab_dtype, const float* bias, int64_t M, int N, int K,
                                 const int64_t* batch, const int32_t* ptr, int64_t B, float* acc,
                                 void* stream) {
  GN_CHECK_ARG(M >= 0 && N >= 1 && K >= 1 && lda >= K && ldw >= K && B >= 0,
               "gn_tc_linear_pool: bad sizes");
  GN_CHECK_ARG(ab_dtype == GN_BF16 || ab_dtype == GN_F16, "gn_tc_linear_pool: bad ab_dtype");
  GN_CHECK_ARG(N <= tc::LIN_MAX_N, "gn_tc_linear_pool: N=%d exceeds %d", N, tc::LIN_MAX_N);
  if (B == 0) return GN_OK;
  GN_CHECK_ARG(acc, "gn_tc_linear_pool: null pointer");
  cudaStream_t st = as_stream(stream);
  // identities of the three reductions: min over bit patterns of non-negative floats starts at
  // 0xFFFFFFFF, max and sum at 0
  const size_t plane = (size_t)B * (size_t)N * sizeof(float);
  GN_CHECK_ARG((reinterpret_cast<uintptr_t>(acc) & 7) == 0 && (plane & 7) == 0,
               "gn_tc_linear_pool: acc must be 8-byte aligned and B*N even");
  GN_CUDA(cudaMemsetAsync(acc, 0xFF, plane, st));
  GN_CUDA(cudaMemsetAsync(reinterpret_cast<uint8_t*>(acc) + plane, 0, 3 * plane, st));
  if (M == 0) return GN_OK;
  GN_CHECK_ARG(A && W && batch && ptr, "gn_tc_linear_pool: null pointer");
  int BN = tc_pick_bn(N);
  if (!tc_wres_fits(N, K, BN, 4) && M >= 4096) {
    const int bn2 = tc_pick_bn(N, 128);
    if (tc_wres_fits(N, K, bn2, 4)) BN = bn2;
  }
  CUtensorMap tmA, tmW;
  auto mk = ab_dtype == GN_F16 ? make_tmap_f16_2d : make_tmap_bf16_2d;
  int rc = mk(&tmA, A, (uint64_t)M, (uint64_t)K, (uint64_t)lda, tc::LIN_BM, tc::LIN_BK, 128);
  if (rc != GN_OK) return rc;
  rc = mk(&tmW, W, (uint64_t)N, (uint64_t)K, (uint64_t)ldw, (uint32_t)BN, tc::LIN_BK, 128);
  if (rc != GN_OK) return rc;
  tc::PoolEpilogue pe;
  pe.batch = batch;
  pe.ptr = ptr;
  pe.acc = acc;
  pe.B = B;
  return tc::launch_tc_linear<float>(tmA, tmW, bias, acc, N, M, N, K, BN, GN_ACT_RELU, ab_dtype,
                                     st, 1, 0, &pe);
}

extern "C" int gn_tc_linear_accumulate(const void* A, int64_t lda, const void* W, int64_t ldw,
                                       int ab_dtype, float* C, int64_t ldc, int64_t M, int N,
                                       int K, int splits, void* stream) {
  GN_CHECK_ARG(M >= 0 && N >= 1 && K >= 1 && lda >= K && ldw >= K && ldc >= N && splits >= 0,
               "gn_tc_linear_accumulate: bad sizes");
  if (M == 0) return GN_OK;
  GN_CHECK_ARG(A && W && C, "gn_tc_linear_accumulate: null pointer");
  GN_CHECK_ARG(ab_dtype == GN_BF16 || ab_dtype == GN_F16, "gn_tc_linear_accumulate: bad ab_dtype");
  GN_CHECK_ARG(N <= tc::LIN_MAX_N, "gn_tc_linear_accumulate: N=%d exceeds %d", N, tc::LIN_MAX_N);
  const int BN = tc_pick_bn(N);
  CUtensorMap tmA, tmW;
  auto mk = ab_dtype == GN_F16 ? make_tmap_f16_2d : make_tmap_bf16_2d;
  int rc = mk(&tmA, A, (uint64_t)M, (uint64_t)K, (uint64_t)lda, tc::LIN_BM, tc::LIN_BK, 128);
  if (rc != GN_OK) return rc;
  rc = mk(&tmW, W, (uint64_t)N, (uint64_t)K, (uint64_t)ldw, (uint32_t)BN, tc::LIN_BK, 128);
  if (rc != GN_OK) return rc;
  return tc::launch_tc_linear<float>(tmA, tmW, nullptr, C, ldc, M, N, K, BN, GN_ACT_NONE, ab_dtype,
                                     as_stream(stream), splits, 1);
}

#ifdef GN_DEBUG_BUILD
// profiling builds only (declared in include/graphnet_b200_debug.h): timeline of CTA 0 of the
// last tc_linear launch
extern "C" int gn_debug_linear_trace(unsigned long long* host, int words) {
  GN_CUDA(cudaDeviceSynchronize());
  if (words > 10 * 16 * 4) words = 10 * 16 * 4;
  GN_CUDA(cudaMemcpyFromSymbol(host, tc::g_lin_trace, words * sizeof(unsigned long long)));
  return GN_OK;
}
#endif
