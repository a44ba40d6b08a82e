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
// per ring stage g < 64 of CTA 0: [0] TMA issued, [1] MMA thread saw `full`, [2] commit issued
__device__ unsigned long long g_lin_stage[3 * 64];
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
    const __grid_constant__ CUtensorMap tmC, const float* __restrict__ bias, OutT* __restrict__ C, int64_t ldc, int64_t M, int N, int K,
    int BN, int num_n_tiles, int64_t num_tiles, uint32_t idesc, int stages, int splits,
    int accumulate, int coalesce_in, int wres, int cshape, const PoolEpilogue pool) {
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
  uint32_t gstage = 0;
  auto stage_stamp = [&](int k) {
    if (traced && gstage < 64) g_lin_stage[k * 64 + gstage] = clock64();
  };
#else
  auto stamp = [](uint32_t, int) {};
  [[maybe_unused]] uint32_t gstage = 0;
  auto stage_stamp = [](int) {};
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
  // K blocks (64 columns) per ring stage: two when W is resident - half as many barrier
  // hand-shakes per tile for the two single threads that pace the kernel
  const int kpb = (int)(((uint32_t)cshape >> 16) & 0xffu);
  const int kb_bytes = LIN_A_BYTES + (wres ? 0 : b_bytes);   // one K block of a stage
  const int stage_bytes = kpb * kb_bytes;
  LinBarriers* bars = reinterpret_cast<LinBarriers*>(ring + stages * stage_bytes);
  float* s_bias = reinterpret_cast<float*>(bars + 1);  // [num_n_tiles * BN]
  uint8_t* s_scr = reinterpret_cast<uint8_t*>(s_bias + num_n_tiles * BN);  // one scratch per epilogue warp
  // (512-byte aligned: a 16-bit scratch is a 64B-swizzled TMA store box)
  s_scr = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(s_scr) + 511) & ~uintptr_t(511));

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int nk = (K + LIN_BK - 1) / LIN_BK;
  for (int n = threadIdx.x; n < num_n_tiles * BN; n += LIN_THREADS)
    s_bias[n] = (bias != nullptr && n < N) ? bias[n] : 0.f;

  // cluster of cn x cm CTAs (see LinPlan): rank = m_idx * cn + n_idx
  const uint32_t cn = (uint32_t)cshape & 0xffu, cm = ((uint32_t)cshape >> 8) & 0xffu, csize = cn * cm;
  const uint32_t crank = blockIdx.x % csize;   // == %cluster_ctarank of a 1-D cluster
  const uint32_t n_idx = crank % cn, m_idx = crank / cn;
  // CTAs that receive the A parts this CTA loads (its cluster row) and the W parts (its column);
  // the same CTAs write into this one's ring, so a slot is free once all of them released it
  const uint16_t mask_a = (uint16_t)(((1u << cn) - 1u) << (m_idx * cn));
  uint16_t mask_w = 0;
  for (uint32_t i = 0; i < cm; ++i) mask_w |= (uint16_t)(1u << (i * cn + n_idx));
  const uint16_t mask_rel = (uint16_t)(mask_a | (wres ? 0 : mask_w));
  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmW);
#pragma unroll
    for (int s = 0; s < LIN_STAGES; ++s) {
      mbar_init(&bars->full[s], 1);
      mbar_init(&bars->empty[s], (uint32_t)__popc((unsigned)mask_rel));
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
  if (csize > 1) cl_sync();   // every CTA's barriers exist before a peer signals them
  tc_fence_after();
  const uint32_t tmem_base = bars->tmem_base;

  // Persistent: this CTA walks work items blockIdx.x, +gridDim.x, ...; item = (tile, K split),
  // split fastest; consecutive tile ids share the A row block (n fastest), so concurrently
  // running CTAs hit the same A tile in L2.  splits > 1 (weight gradients: few tiles, K = rows of
  // the batch) gives every item the K blocks [kb0, kb1) and the epilogue adds with red.add.
  //
  // The three roles walk the same sequence, each in ONE thread (or one warp) of dependent
  // instructions, so the walk must not divide: the item -> (m tile, n tile, split) map is advanced
  // incrementally (the 64-bit divisions and the g % stages ring index of the first version cost
  // ~1100 cycles per item and ~700 per stage in the producer and MMA threads - more than the
  // tensor work of a node-GEMM tile - and paced the whole kernel).
  struct Walk {
    uint32_t mt, nt, sp;          // current item
    uint32_t dm, dn, ds;          // gridDim.x in mixed radix (m tiles, n tiles, splits)
    uint32_t nnt, nsp, nmt, off;
    bool wres;
    // cluster: cluster c takes the row-tile groups c, c + clusters, ...; CTA (m_idx, n_idx) of it
    // the tile (group * cm + m_idx, n_idx).  Every CTA of a cluster walks the same number of items
    // (a row tile past the end loads zeros and stores nothing).
    __device__ __forceinline__ void init_cluster(uint32_t cta, uint32_t grid, uint32_t cn, uint32_t cm,
                                                 uint32_t n_idx, uint32_t m_idx, uint32_t nmt_) {
      const uint32_t csize = cn * cm;
      nnt = cn; nsp = 1; nmt = nmt_; wres = false; off = m_idx;
      nt = n_idx; mt = (cta / csize) * cm + m_idx; sp = 0;
      dm = (grid / csize) * cm; dn = 0; ds = 0;
    }
    __device__ __forceinline__ void init(uint32_t cta, uint32_t grid, uint32_t nnt_, uint32_t nsp_,
                                         uint32_t nmt_, bool wres_) {
      nnt = nnt_; nsp = nsp_; nmt = nmt_; wres = wres_; off = 0;
      if (wres) {                 // one n tile per CTA; the CTAs of an n tile stride over m
        nt = cta % nnt; mt = cta / nnt; sp = 0;
        dm = grid / nnt; dn = 0; ds = 0;
      } else {
        sp = cta % nsp; const uint32_t t = cta / nsp; nt = t % nnt; mt = t / nnt;
        ds = grid % nsp; const uint32_t g = grid / nsp; dn = g % nnt; dm = g / nnt;
      }
    }
    __device__ __forceinline__ bool valid() const { return mt - off < nmt; }
    __device__ __forceinline__ void next() {
      sp += ds;
      if (sp >= nsp) { sp -= nsp; ++nt; }
      nt += dn;
      if (nt >= nnt) { nt -= nnt; ++mt; }
      if (nt >= nnt) { nt -= nnt; ++mt; }   // (carry + dn can pass nnt twice)
      mt += dm;
    }
  };
  const uint32_t num_m_tiles = (uint32_t)(num_tiles / num_n_tiles);
  Walk walk;
  if (csize > 1)
    walk.init_cluster(blockIdx.x, gridDim.x, cn, cm, n_idx, m_idx, num_m_tiles);
  else
    walk.init(blockIdx.x, gridDim.x, (uint32_t)num_n_tiles, (uint32_t)splits, num_m_tiles, wres != 0);
  // K blocks [kb0, kb1) of split sp (splits == 1: all of them)
  auto k_range = [&](uint32_t sp, int& kb0, int& kb1) {
    if (splits == 1) { kb0 = 0; kb1 = nk; return; }
    kb0 = (int)(sp * (uint32_t)nk / (uint32_t)splits);
    kb1 = (int)((sp + 1) * (uint32_t)nk / (uint32_t)splits);
  };
  const int cta_n = (int)walk.nt;   // (wres: the CTA's one column tile)
  if (warp == 0) {
    // ---------------------------------------------------------- TMA producer
    if (elect_one()) {
      if (wres) {  // the CTA's W tile, once
        mbar_arrive_expect_tx(&bars->w_full, (uint32_t)(nk_all * b_bytes));
        for (int kb = 0; kb < nk_all; ++kb)
          tma_load_2d(w_res + kb * b_bytes, &tmW, &bars->w_full, kb * LIN_BK, cta_n * BN);
      }
      uint32_t s = 0, ph = 1;   // ring slot and the parity its `empty` barrier is waited with
      const uint32_t ring_a = smem_u32(ring);
      const uint32_t a_part_rows = LIN_BM / cn, a_part_bytes = a_part_rows * (LIN_BK * 2);
      const uint32_t w_part_rows = (uint32_t)BN / cm, w_part_bytes = w_part_rows * (LIN_BK * 2);
      for (uint32_t i = 0; walk.valid(); walk.next(), ++i) {
        int kb0, kb1;
        k_range(walk.sp, kb0, kb1);
        const int32_t m0 = (int32_t)(walk.mt * LIN_BM);
        const int32_t n0 = (int32_t)walk.nt * BN;
        stamp(i, 0);
        for (int kb = kb0; kb < kb1; kb += kpb) {
          const int nb = min(kpb, kb1 - kb);
          mbar_wait(&bars->empty[s], ph);
          if (kb == kb0) stamp(i, 1);
          if (kb + nb >= kb1) stamp(i, 2);
          mbar_arrive_expect_tx(&bars->full[s], (uint32_t)(nb * kb_bytes));
          stage_stamp(0);
          ++gstage;
          for (int j = 0; j < nb; ++j) {
            const uint32_t a_dst = ring_a + s * (uint32_t)stage_bytes + (uint32_t)(j * kb_bytes);
            const int32_t k0 = (kb + j) * LIN_BK;
            if (cn > 1)   // this CTA's 128 / cn rows of the A tile, into every CTA of its row
              tma_load_2d_mc(a_dst + n_idx * a_part_bytes, &tmA, &bars->full[s], k0,
                             m0 + (int32_t)(n_idx * a_part_rows), mask_a);
            else
              tma_load_2d_a(a_dst, &tmA, &bars->full[s], k0, m0);
            if (!wres) {
              if (cm > 1)
                tma_load_2d_mc(a_dst + LIN_A_BYTES + m_idx * w_part_bytes, &tmW, &bars->full[s],
                               k0, n0 + (int32_t)(m_idx * w_part_rows), mask_w);
              else
                tma_load_2d_a(a_dst + LIN_A_BYTES, &tmW, &bars->full[s], k0, n0);
            }
          }
          if (++s == (uint32_t)stages) { s = 0; ph ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    // ------------------------------------------------------------ MMA issuer
    // The whole warp walks the loop (every value in it is warp-uniform) and ONE elected lane
    // issues: ptxas then keeps descriptors and addresses in uniform registers.  Under
    // `if (lane == 0)` it cannot know that a single thread is active and wraps every tcgen05.mma
    // in an ELECT / 5 x R2UR / branch loop (~70 cycles per instruction, measured).
    if (wres) {
      mbar_wait(&bars->w_full, 0);
      tc_fence_after();
    }
    uint32_t s = 0, ph = 0;
    const uint32_t ring_a = smem_u32(ring), wres_a = smem_u32(w_res);
    for (uint32_t it = 0; walk.valid(); walk.next(), ++it) {
      int kb0, kb1;
      k_range(walk.sp, kb0, kb1);
      const uint32_t buf = it & 1;
      stamp(it, 0);
      mbar_wait(&bars->tmem_empty[buf], ((it >> 1) & 1) ^ 1);
      stamp(it, 1);
      tc_fence_after();
      const uint32_t tmem_d = tmem_base + buf * 256;
      for (int kb = kb0; kb < kb1; kb += kpb) {
        const int nb = min(kpb, kb1 - kb);
        mbar_wait(&bars->full[s], ph);
        stage_stamp(1);
        if (kb == kb0) stamp(it, 2);
        if (kb + nb >= kb1) stamp(it, 3);
        tc_fence_after();
        if (elect_one()) {
          for (int j = 0; j < nb; ++j) {
            const uint32_t a_addr = ring_a + s * (uint32_t)stage_bytes + (uint32_t)(j * kb_bytes);
            const uint32_t b_addr =
                wres ? wres_a + (uint32_t)((kb + j) * b_bytes) : a_addr + LIN_A_BYTES;
            const uint64_t adesc = make_kmajor_desc<128>(a_addr);
            const uint64_t bdesc = make_kmajor_desc<128>(b_addr);
            if (!(coalesce_in & 8)) {   // (8: profiling without tensor work, data path only)
              // advance 16 bf16 = 32 bytes along K inside the swizzle row: +2 in (addr >> 4) units
              mma_bf16_ss(tmem_d, adesc, bdesc, idesc, kb + j != kb0);
              mma_bf16_ss(tmem_d, adesc + 2, bdesc + 2, idesc, 1);
              mma_bf16_ss(tmem_d, adesc + 4, bdesc + 4, idesc, 1);
              mma_bf16_ss(tmem_d, adesc + 6, bdesc + 6, idesc, 1);
            }
          }
          if (csize > 1) mma_commit_mc(&bars->empty[s], mask_rel);
          else mma_commit(&bars->empty[s]);
          if (kb + nb >= kb1) mma_commit(&bars->tmem_full[buf]);
        }
        __syncwarp();
        stage_stamp(2);
        ++gstage;
        if (++s == (uint32_t)stages) { s = 0; ph ^= 1; }
      }
    }
  } else {
    // -------------------------------------------------------------- epilogue
    const int q = warp & 3;  // TMEM lane quadrant this warp may read
    const int par = (warp - 2) >> 2;  // the two warps of a quadrant take alternate 32-col chunks
    for (uint32_t it = 0; walk.valid(); walk.next(), ++it) {
      const uint32_t buf = it & 1;
      const int64_t m0 = (int64_t)walk.mt * LIN_BM;
      const int n0 = (int)walk.nt * BN;
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
          const int nv = min(N - n, BN - c);                 // valid columns of this chunk
          if constexpr (sizeof(OutT) == 2) {
            // the scratch IS a 64B-swizzled [32 rows x 64 bytes] box: a full chunk leaves as one
            // TMA store (rows beyond M are clipped by the unit) - no read-back, no per-lane stores
            if ((coalesce_in & 32) && BN - c >= 32) {
              fence_proxy_async_smem();
              __syncwarp();
              if (lane == 0) {
                if (!(coalesce_in & 2)) tma_store_2d(&tmC, scr, n, (int32_t)(m0 + q * 32));
                tma_store_wait_read();     // (the next chunk rewrites the scratch)
              }
              __syncwarp();
              return;
            }
          }
          __syncwarp();
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
  if (warp >= 2 && lane == 0) tma_store_wait_all();
  tc_fence_before();
  __syncthreads();
  if (csize > 1) cl_sync();   // no CTA leaves while a peer may still signal its barriers
  if (warp == 1) tmem_dealloc<512>(tmem_base);
}

// largest tile width <= bn_max (multiple of 16) that minimises padding
static int tc_pick_bn(int N, int bn_max_arg = 256) {
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

template <typename OutT>
static size_t lin_fixed_smem(int N, int BN) {
  const int num_n_tiles = (N + BN - 1) / BN;
  return sizeof(LinBarriers) + (size_t)num_n_tiles * BN * sizeof(float) + 32 + 1024 +
         (size_t)lin_epi_warps<OutT>() * lin_scr_bytes<OutT>() + 512;
}
// does a K-complete W tile of width BN fit next to a ring of at least four A stages?
template <typename OutT>
static bool lin_wres_fits(int N, int K, int BN) {
  const size_t w_bytes = (size_t)ceil_div(K, LIN_BK) * BN * LIN_BK * 2;
  return w_bytes + 4 * (size_t)LIN_A_BYTES + lin_fixed_smem<OutT>(N, BN) <= 227 * 1024;
}

// How one GEMM runs: tile width, whether the W tile stays in shared memory, and the cluster.
//
// Cluster (cn x cm CTAs = the cn column tiles of cm consecutive row tiles): every stage of A is
// fetched ONCE per cluster row - each of the cn CTAs loads 128 / cn of its rows and the TMA unit
// multicasts them into all cn - and every streamed stage of W once per cluster column.  Without it
// each CTA pulls whole A and W tiles through L2 for itself: 4 x 31 MB of A for the node GEMMs of
// C2, 266 MB of A + 351 MB of W for the first post-processing GEMM, and the kernels ran at the
// L2 -> SM throughput limit (the loads alone, no tensor work, no epilogue, took 2/3 of their time).
struct LinPlan {
  int BN, wres, cn, cm, kpb, stages, num_n_tiles;
  size_t smem;
};

template <typename OutT>
static LinPlan lin_plan(int64_t M, int N, int K, int splits, int accumulate, int sms) {
  static const bool no_wres = getenv("GN_TC_NO_WRES") != nullptr;
  static const int max_cluster = getenv("GN_TC_MAX_CLUSTER") ? atoi(getenv("GN_TC_MAX_CLUSTER")) : 4;
  LinPlan p;
  memset(&p, 0, sizeof(p));
  p.cn = p.cm = 1;
  int BN = tc_pick_bn(N);
  const bool plain = splits == 1 && !accumulate;
  // a narrower tile whose K-complete W fits in shared memory beats a wide one that streams W
  if (plain && !lin_wres_fits<OutT>(N, K, BN) && M >= 4096) {
    const int bn2 = tc_pick_bn(N, 128);
    if (lin_wres_fits<OutT>(N, K, bn2)) BN = bn2;
  }
  int nnt = (int)ceil_div(N, BN);
  const int64_t m_tiles = ceil_div(M, LIN_BM);
  if (plain && !no_wres && nnt <= sms && lin_wres_fits<OutT>(N, K, BN) &&
      m_tiles >= 2 * (sms / nnt)) {
    p.wres = 1;
    // (no cluster here: with W resident only A streams, and multicasting it between the column
    // tiles of a row made the node GEMMs of C2 slower, 38 vs 34 us in the running step - their
    // pace is set by the MMA / epilogue contention for shared memory, not by L2 traffic)
    static const bool wres_cluster = getenv("GN_TC_WRES_CLUSTER") != nullptr;
    if (wres_cluster && (nnt == 2 || nnt == 4) && nnt <= max_cluster) p.cn = nnt;
  } else if (plain && M >= 4096 && max_cluster >= 2) {
    // W streams: share it between cm row tiles, and A between the column tiles.  Parts of a W
    // stage must start on a swizzle atom (8 rows): pick the tile width accordingly.
    for (int bn = 256; bn >= 64; bn -= 16) {
      const int t = (N + bn - 1) / bn;
      if (t != 1 && t != 2 && t != 4) continue;
      if (t > max_cluster) continue;
      const int cm = max_cluster / t;          // 1, 2 or 4
      if ((bn / cm) % 8) continue;
      const int waste = t * bn - N;
      if (waste * 16 > N) continue;            // at most ~6 % padded columns
      BN = bn; nnt = t; p.cn = t; p.cm = cm;
      break;
    }
  }
  p.BN = BN;
  p.num_n_tiles = nnt;
  const size_t fixed = lin_fixed_smem<OutT>(N, BN);
  const size_t b_bytes = (size_t)BN * LIN_BK * 2;
  const size_t w_bytes = (size_t)ceil_div(K, LIN_BK) * b_bytes;
  const size_t avail = 227 * 1024 - fixed - (p.wres ? w_bytes : 0);
  static const bool no_kpb2 = getenv("GN_TC_KPB1") != nullptr;
  p.kpb = (p.wres && !no_kpb2 && ceil_div(K, LIN_BK) >= 2 && avail >= 3 * 2 * (size_t)LIN_A_BYTES) ? 2 : 1;
  const size_t stage_bytes = (size_t)p.kpb * ((size_t)LIN_A_BYTES + (p.wres ? 0 : b_bytes));
  int stages = (int)(avail / stage_bytes);   // as deep as shared memory allows
  p.stages = stages > LIN_STAGES ? LIN_STAGES : stages;
  p.smem = (p.wres ? w_bytes : 0) + (size_t)p.stages * stage_bytes + fixed;
  return p;
}

template <typename OutT>
static int launch_tc_linear(const void* A, int64_t lda, const void* W, int64_t ldw,
                            const float* bias, void* C, int64_t ldc, int64_t M, int N, int K,
                            int act, int ab_dtype, cudaStream_t st, int splits = 1,
                            int accumulate = 0, const PoolEpilogue* pool_in = nullptr) {
  PoolEpilogue pool;
  memset(&pool, 0, sizeof(pool));
  if (pool_in) pool = *pool_in;
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
  const int nkb = (int)ceil_div(K, LIN_BK);
  LinPlan pl = lin_plan<OutT>(M, N, K, splits == 1 ? 1 : 2, accumulate, sms);
  const int BN = pl.BN, num_n_tiles = pl.num_n_tiles;
  const int64_t num_m_tiles = ceil_div(M, LIN_BM);
  const int64_t num_tiles = num_m_tiles * num_n_tiles;
  if (splits == 0) {  // automatic: enough items to fill the machine, at least 4 K blocks each
    splits = (int)(sms / (num_tiles < sms ? num_tiles : sms));
    if (splits > nkb / 4) splits = nkb / 4;
  }
  if (splits > nkb) splits = nkb;
  if (splits < 1) splits = 1;
#ifdef GN_DEBUG_BUILD
  {
    static const int cap = getenv("GN_TC_STAGES") ? atoi(getenv("GN_TC_STAGES")) : 0;
    if (cap > 0 && cap < pl.stages) pl.stages = cap;
  }
#endif
  CUtensorMap tmA, tmW;
  auto mk = ab_dtype == GN_F16 ? make_tmap_f16_2d : make_tmap_bf16_2d;
  int rc = mk(&tmA, A, (uint64_t)M, (uint64_t)K, (uint64_t)lda, LIN_BM / pl.cn, LIN_BK, 128);
  if (rc != GN_OK) return rc;
  rc = mk(&tmW, W, (uint64_t)N, (uint64_t)K, (uint64_t)ldw, (uint32_t)(BN / pl.cm), LIN_BK, 128);
  if (rc != GN_OK) return rc;
  const uint32_t idesc =
      ab_dtype == GN_F16 ? make_idesc_f16(LIN_BM, BN) : make_idesc_bf16(LIN_BM, BN);
  CUtensorMap tmC = tmA;      // (only read by 16-bit outputs with TMA stores)
  bool tma_store = false;
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
  const int coalesce = ((hack_nostore ? 3 : 1) | (dbg_trace ? 4 : 0) | (hack_bits & 24)) *
                       (int)(!no_coalesce && (reinterpret_cast<uintptr_t>(C) % 16 == 0) &&
                             ((ldc * (int64_t)sizeof(OutT)) % 16 == 0) &&
                             ((BN * (int)sizeof(OutT)) % 16 == 0));
  if (sizeof(OutT) == 2 && coalesce != 0 && splits == 1 && !accumulate) {
    static const bool no_tma_store = getenv("GN_TC_NO_TMASTORE") != nullptr;
    if (!no_tma_store) {
      // C as a [M, N] tensor of 16-bit elements, boxes of 32 rows x 32 columns, 64B swizzle
      rc = make_tmap_f16_2d(&tmC, C, (uint64_t)M, (uint64_t)N, (uint64_t)ldc, 32, 32, 64);
      if (rc != GN_OK) return rc;
      tma_store = true;
    }
  }
  const int csize = pl.cn * pl.cm;
  const int64_t num_items = num_tiles * splits;
  const int wres = pl.wres, stages = pl.stages, cshape = pl.cn | (pl.cm << 8) | (pl.kpb << 16);
  const size_t smem = pl.smem;
  /* per kernel instantiation and device, once: the opt-in to > 48 KB of dynamic shared memory   \
     and how many clusters of each size the device holds at a time (a persistent kernel must not  \
     launch more: GPCs whose SM count is not a multiple of the cluster leave SMs out) */          \
#define GN_TC_LIN_LAUNCH(KERN_)                                                                 \
  {                                                                                             \
    auto kern = KERN_;                                                                          \
    static std::atomic<unsigned long long> ready_;                                              \
    static std::atomic<int> max_clusters_[64][9];                                               \
    if (!((ready_.load(std::memory_order_relaxed) >> (dev & 63)) & 1ull)) {                     \
      GN_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize,           \
                                   227 * 1024));                                                \
      ready_.fetch_or(1ull << (dev & 63), std::memory_order_relaxed);                           \
    }                                                                                           \
    cudaLaunchConfig_t cfg;                                                                     \
    memset(&cfg, 0, sizeof(cfg));                                                               \
    cudaLaunchAttribute attr[1];                                                                \
    attr[0].id = cudaLaunchAttributeClusterDimension;                                           \
    attr[0].val.clusterDim.x = (unsigned)csize;                                                 \
    attr[0].val.clusterDim.y = 1;                                                               \
    attr[0].val.clusterDim.z = 1;                                                               \
    cfg.blockDim = dim3(lin_threads<OutT>());                                                   \
    cfg.dynamicSmemBytes = smem;                                                                \
    cfg.stream = st;                                                                            \
    cfg.attrs = attr;                                                                           \
    cfg.numAttrs = 1;                                                                           \
    unsigned grid;                                                                              \
    if (csize > 1) {                                                                            \
      int ncl = max_clusters_[dev & 63][csize].load(std::memory_order_relaxed);                 \
      if (ncl == 0) {                                                                           \
        cfg.gridDim = dim3((unsigned)(sms / csize * csize));                                    \
        cfg.dynamicSmemBytes = 227 * 1024;                                                      \
        GN_CUDA(cudaOccupancyMaxActiveClusters(&ncl, kern, &cfg));                              \
        cfg.dynamicSmemBytes = smem;                                                            \
        if (ncl < 1) {                                                                          \
          set_error("gn_tc_linear: no cluster of %d CTAs fits on this device", csize);          \
          return GN_ERR_UNSUPPORTED;                                                            \
        }                                                                                       \
        max_clusters_[dev & 63][csize].store(ncl, std::memory_order_relaxed);                   \
      }                                                                                         \
      const int64_t supers = ceil_div(num_m_tiles, pl.cm);                                      \
      grid = (unsigned)((supers < ncl ? supers : ncl) * csize);                                 \
    } else if (wres) {                                                                          \
      grid = (unsigned)((sms / num_n_tiles) * num_n_tiles);                                     \
    } else {                                                                                    \
      grid = (unsigned)(num_items < sms ? num_items : sms);                                     \
    }                                                                                           \
    cfg.gridDim = dim3(grid);                                                                   \
    GN_CUDA(cudaLaunchKernelEx(&cfg, kern, tmA, tmW, tmC, bias, reinterpret_cast<OutT*>(C), ldc, M, N, \
                               K, BN, num_n_tiles, num_tiles, idesc, stages, splits,             \
                               accumulate, coalesce | (tma_store ? 32 : 0), wres, cshape, pool)); \
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
    case GN_ACT_LEAKY_RELU: GN_TC_LIN_CASE(ACT_LEAKY_RELU) break;
    default: set_error("unknown activation %d", act); return GN_ERR_INVALID_ARGUMENT;
  }
#undef GN_TC_LIN_CASE
#undef GN_TC_LIN_LAUNCH
  GN_LAUNCH_CHECK();
  return GN_OK;
}

}  // namespace tc
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
  cudaStream_t st = as_stream(stream);
  if (c_dtype == GN_F32)
    return tc::launch_tc_linear<float>(A, lda, W, ldw, bias, C, ldc, M, N, K, act, ab_dtype, st);
  if (c_dtype == GN_F16)
    return tc::launch_tc_linear<__half>(A, lda, W, ldw, bias, C, ldc, M, N, K, act, ab_dtype, st);
  return tc::launch_tc_linear<__nv_bfloat16>(A, lda, W, ldw, bias, C, ldc, M, N, K, act, ab_dtype,
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
  tc::PoolEpilogue pe;
  pe.batch = batch;
  pe.ptr = ptr;
  pe.acc = acc;
  pe.B = B;
  return tc::launch_tc_linear<float>(A, lda, W, ldw, bias, acc, N, M, N, K, GN_ACT_RELU, ab_dtype,
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
  return tc::launch_tc_linear<float>(A, lda, W, ldw, nullptr, C, ldc, M, N, K, GN_ACT_NONE,
                                     ab_dtype, as_stream(stream), splits, 1);
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
extern "C" int gn_debug_linear_stages(unsigned long long* host, int words) {
  GN_CUDA(cudaDeviceSynchronize());
  if (words > 3 * 64) words = 3 * 64;
  GN_CUDA(cudaMemcpyFromSymbol(host, tc::g_lin_stage, words * sizeof(unsigned long long)));
  return GN_OK;
}
#endif
