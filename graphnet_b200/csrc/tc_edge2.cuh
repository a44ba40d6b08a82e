// Fused EdgeConv message + aggregation, second generation ("transposed" formulation).
//
//   D^T[ch, e] = sum_k W2[ch, k] * h[e, k],   h[e] = ReLU(P_i(e) + Q_j(e))
//   out[i, ch] = sum_{e in N(i)} ReLU(D^T[ch, e])
//
// W2 is the A operand and lives in TENSOR MEMORY for the whole kernel (written once with
// tcgen05.st: lane = output channel, 32-bit column = two consecutive k), so shared memory holds
// no weights at all.  The 64-edge tile h is the B operand (N = 64) in the 64B-swizzled K-major
// layout.  What this buys over the first generation (tc_edge.cu, W2 resident in smem):
//   * shared memory is free for a DEEP ASYNCHRONOUS GATHER: eight gather warps copy the neighbour
//     rows Q_j (and node rows P_i) with 16-byte cp.async into a 4-deep ring of 32-row stages that
//     complete on mbarriers; the gather latency (L2 / HBM) is hidden by the ring, not by registers
//     (one cp.async.bulk per 704-byte row was tried first: the per-copy cost of the TMA unit made
//     it slower than the first-generation kernel);
//   * edges are COLUMNS of the accumulator: a thread owns one output channel, so the per-node
//     sum over its KP edges is a plain in-register reduction (no shuffles), and the 32 lanes of
//     a warp write 32 consecutive channels of a node (coalesced);
//   * tensor-core operand traffic from shared memory drops from 96 to 64 B/clk.
//     - and KP need not be a power of two: with KP = k+1 = 9 slots per node a kNN graph needs
//     no overflow pass at all;
// Tile = 64 ELL rows = floor(64/KP) nodes.  Warp roles (17 warps): 0-7 epilogue (warp>>2 =
// channel block of 128, warp&3 = TMEM lane quadrant), 8-15 h builders (8 rows each), 16-19
// gather warps (one per ring stage), 20 tcgen05.mma issuer.  TMEM columns: [0,Kp/2) W2 block 0, [176,176+Kp/2) W2 block 1,
// [352,416) D block 0, [416,480) D block 1 - the epilogue of one channel block overlaps the
// MMAs of the other.
#pragma once

namespace gn {
namespace tc {

constexpr int E2_ROWS = 64;          // edge rows (accumulator columns) per tile
constexpr int E2_HALF = 32;          // rows per gather stage
constexpr int E2_QSTAGES = 4;        // gather ring depth (half tiles): 2 tiles of lookahead
constexpr int E2_HBUFS = 2;          // h tiles (B operand) in shared memory
constexpr int E2_EPI_WARPS = 8;      // (channel block, TMEM lane quadrant)
constexpr int E2_BLD_WARPS = 8;      // h builders: 8 tile rows per warp (one row per lane quad)
constexpr int E2_GAT_SPLIT = 1;     // gather warps per ring stage (2: even / odd rows - tried, the 72-register cap of 25 warps costs the epilogue more than the gather gains)
constexpr int E2_GAT_WARPS = E2_QSTAGES * E2_GAT_SPLIT;
constexpr int E2_MMA_WARP = E2_EPI_WARPS + E2_BLD_WARPS + E2_GAT_WARPS;  // warp 16
constexpr int E2_THREADS = 32 * (E2_MMA_WARP + 1);                       // 544
constexpr int E2_TMEM_A1 = 176;      // column offset of W2 block 1
constexpr int E2_TMEM_D0 = 352;
constexpr int E2_TMEM_D1 = 416;

__host__ __device__ constexpr int e2_pmax(int KP) { return E2_HALF / KP + 2; }  // P rows per stage

struct __align__(16) Edge2Shared {
  uint64_t q_full[E2_QSTAGES];
  uint64_t q_empty[E2_QSTAGES];
  uint64_t h_full[2];
  uint64_t h_empty[2];
  uint64_t d_full[2];
  uint64_t d_empty[2];
  uint32_t tmem_base;
  uint32_t pad_;
  uint8_t row_valid[E2_QSTAGES][E2_HALF];
};

// 16-byte asynchronous global -> shared copy (LDGSTS), L2 only
__device__ __forceinline__ void cp_async16(uint32_t smem_dst, const void* gsrc) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(smem_dst), "l"(gsrc) : "memory");
}
// the same under a predicate (no branch around the copy: the gather loop stays straight-line)
__device__ __forceinline__ void cp_async16_if(uint32_t smem_dst, const void* gsrc, uint32_t on) {
  asm volatile(
      "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %2, 0;\n\t"
      "@p cp.async.cg.shared.global [%0], [%1], 16;\n\t}" ::"r"(smem_dst), "l"(gsrc), "r"(on)
      : "memory");
}
// the mbarrier receives one arrival when all cp.async issued so far by this thread have landed
__device__ __forceinline__ void cp_async_arrive_noinc(uint64_t* bar) {
  asm volatile("cp.async.mbarrier.arrive.noinc.shared::cta.b64 [%0];" ::"r"(smem_u32(bar))
               : "memory");
}

__device__ __forceinline__ void tmem_st_32x32_x16(uint32_t taddr, const uint32_t (&r)[16]) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], "
      "{%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16};" ::"r"(taddr),
      "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]),
      "r"(r[8]), "r"(r[9]), "r"(r[10]), "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15])
      : "memory");
}
__device__ __forceinline__ void tmem_st_wait() {
  asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
}
// D[tmem] (+)= A[tmem] * B[smem]^T  (A: lane = row m, 32-bit column = 2 consecutive k)
__device__ __forceinline__ void mma_f16_ts(uint32_t tmem_d, uint32_t tmem_a, uint64_t bdesc,
                                           uint32_t idesc, bool accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t}" ::"r"(tmem_d),
      "r"(tmem_a), "l"(bdesc), "r"(idesc), "r"((uint32_t)accumulate)
      : "memory");
}

// the same with the accumulate flag fixed at compile time (no predicate set-up per MMA)
template <bool ACC>
__device__ __forceinline__ void mma_f16_ts_c(uint32_t tmem_d, uint32_t tmem_a, uint64_t bdesc,
                                             uint32_t idesc) {
  if (ACC) {
    asm volatile(
        "{\n\t.reg .pred p;\n\tsetp.eq.u32 p, 0, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t}" ::"r"(tmem_d),
        "r"(tmem_a), "l"(bdesc), "r"(idesc)
        : "memory");
  } else {
    asm volatile(
        "{\n\t.reg .pred p;\n\tsetp.ne.u32 p, 0, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t}" ::"r"(tmem_d),
        "r"(tmem_a), "l"(bdesc), "r"(idesc)
        : "memory");
  }
}

// Weight-stationary form: the B operand (the h chunk) is latched in collector buffer b<BUF> by
// the FILL instruction and re-used by the LASTUSE one, so the two channel blocks of a K step read
// the chunk from shared memory once instead of twice.
template <int BUF, bool FILL, bool ACC>
__device__ __forceinline__ void mma_f16_ts_ws(uint32_t tmem_d, uint32_t tmem_a, uint64_t bdesc,
                                              uint32_t idesc) {
#define GN_WS_MMA(BUFS, OP)                                                                      \
  asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.u32 p, %4, 0;\n\t"                               \
               "tcgen05.mma.ws.cta_group::1.kind::f16.collector::" BUFS "::" OP                  \
               " [%0], [%1], %2, %3, p;\n\t}" ::"r"(tmem_d),                                     \
               "r"(tmem_a), "l"(bdesc), "r"(idesc), "r"((uint32_t)ACC)                           \
               : "memory")
  if (BUF == 0) { if (FILL) GN_WS_MMA("b0", "fill"); else GN_WS_MMA("b0", "lastuse"); }
  else { if (FILL) GN_WS_MMA("b1", "fill"); else GN_WS_MMA("b1", "lastuse"); }
#undef GN_WS_MMA
}

// the same with a run-time accumulate flag (first K step of a tile)
template <int BUF, bool FILL>
__device__ __forceinline__ void mma_f16_ts_ws_p(uint32_t tmem_d, uint32_t tmem_a, uint64_t bdesc,
                                                uint32_t idesc, uint32_t acc) {
#define GN_WS_MMA(BUFS, OP)                                                                      \
  asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.u32 p, %4, 0;\n\t"                               \
               "tcgen05.mma.ws.cta_group::1.kind::f16.collector::" BUFS "::" OP                  \
               " [%0], [%1], %2, %3, p;\n\t}" ::"r"(tmem_d),                                     \
               "r"(tmem_a), "l"(bdesc), "r"(idesc), "r"(acc)                                     \
               : "memory")
  if (BUF == 0) { if (FILL) GN_WS_MMA("b0", "fill"); else GN_WS_MMA("b0", "lastuse"); }
  else { if (FILL) GN_WS_MMA("b1", "fill"); else GN_WS_MMA("b1", "lastuse"); }
#undef GN_WS_MMA
}

__device__ __forceinline__ uint4 lds128(uint32_t addr) {
  uint4 r;
  asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
               : "r"(addr));
  return r;
}

// mbar_wait that also accumulates the cycles spent waiting (profiling builds only)
template <bool RELAXED = false>
__device__ __forceinline__ void mbar_wait_t(uint64_t* bar, uint32_t parity, long long& acc,
                                            bool timed) {
  const long long t0 = timed ? clock64() : 0;
  if (RELAXED) mbar_wait_relaxed<40>(bar, parity);
  else mbar_wait(bar, parity);
  if (timed) acc += clock64() - t0;
}

// ACC: second pass (group lists, red.add into out); F16: output element type (else bf16)
template <int KP, bool ACC, bool F16>
__global__ void __launch_bounds__(E2_THREADS, 1) edge_mlp_aggregate_v2_kernel(const EdgeParams p) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) &
                                             ~uintptr_t(1023));
  constexpr int NPT = E2_ROWS / KP;    // nodes (groups) per tile; KP need not be a power of two
  constexpr int USED = NPT * KP;       // tile rows that carry slots (the rest stay zero)
  constexpr int PMAX = e2_pmax(KP);    // node rows a 32-row stage can touch
  const int nch = p.nch;
  const int rowbytes = p.Kp * 2;
  const int htile_bytes = nch * (E2_ROWS * 64);           // [kc][64 rows][64 B]
  const int qstage_bytes = (E2_HALF + PMAX) * rowbytes;   // 32 Q rows then PMAX P rows
  uint8_t* h_smem = smem;                                 // 2 tiles
  uint8_t* q_smem = smem + E2_HBUFS * htile_bytes;        // E2_QSTAGES stages
  Edge2Shared* sh = reinterpret_cast<Edge2Shared*>(q_smem + E2_QSTAGES * qstage_bytes);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int64_t num_groups = p.group_count ? (int64_t)__ldg(p.group_count) : p.N;
  const int64_t num_tiles = (num_groups + NPT - 1) / NPT;
  if ((int64_t)blockIdx.x >= num_tiles) return;  // uniform per CTA, before any barrier
  const int nblk = p.bnp > 128 ? 2 : 1;

  if (threadIdx.x == 0) {
#pragma unroll
    for (int s = 0; s < E2_QSTAGES; ++s) {
      mbar_init(&sh->q_full[s], E2_GAT_SPLIT * (32 + 1));  // the stage's gather warps
      mbar_init(&sh->q_empty[s], E2_BLD_WARPS / 2);
    }
#pragma unroll
    for (int b = 0; b < 2; ++b) {
      mbar_init(&sh->h_full[b], E2_BLD_WARPS);
      mbar_init(&sh->h_empty[b], 1);
      mbar_init(&sh->d_full[b], 1);
      mbar_init(&sh->d_empty[b], E2_EPI_WARPS / 2);
    }
    fence_mbar_init();
  }
  if (warp == E2_MMA_WARP) tmem_alloc<512>(&sh->tmem_base);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = sh->tmem_base;

  // ---- W2 -> tensor memory (once): epilogue warp (blk, q) writes lanes 32q.. of block blk;
  // each lane reads its own weight row with 16-byte loads
  if (warp < E2_EPI_WARPS) {
    const int blk = warp >> 2, q = warp & 3;
    if (blk < nblk) {
      const int ch = blk * 128 + q * 32 + lane;
      const uint4* wrow = reinterpret_cast<const uint4*>(
          reinterpret_cast<const __half*>(p.w2) + (int64_t)ch * p.Kp);
      const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + (blk ? E2_TMEM_A1 : 0);
      for (int cc = 0; cc < nch; ++cc) {  // 4 x 16 bytes = 32 fp16 = one K chunk = 16 columns
        uint32_t r[16];
#pragma unroll
        for (int t = 0; t < 4; ++t) {
          const uint4 w = ch < p.bnp ? __ldg(wrow + cc * 4 + t) : make_uint4(0, 0, 0, 0);
          r[4 * t] = w.x; r[4 * t + 1] = w.y; r[4 * t + 2] = w.z; r[4 * t + 3] = w.w;
        }
        tmem_st_32x32_x16(taddr + cc * 16, r);
      }
      tmem_st_wait();
    }
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();

  const int64_t my_tiles = (num_tiles - blockIdx.x + gridDim.x - 1) / gridDim.x;
  const uint32_t total_halves = (uint32_t)(2 * my_tiles);
  // optional wait-time accounting: dbg[0..7] = cycles waited on q_empty (gather), q_full and
  // h_empty (builders), h_full and d_empty (MMA), d_full (epilogue), total cycles, tiles
  const bool timed = p.dbg != nullptr;
  long long w0 = 0, w1 = 0;
  const long long t_start = clock64();

  if (warp >= E2_EPI_WARPS + E2_BLD_WARPS && warp < E2_MMA_WARP) {
    // ================================================================ gather warps
    // Stage s of the ring belongs to gather warp s; half-stage g2 (= 2*tile_iter + half) lives
    // in stage g2 % 4, so a warp only ever waits for the release of its own previous use.  A stage holds the 32 Q rows of the half tile, then the P rows of the nodes it
    // touches, Kp fp16 each, filled with 16-byte cp.async pieces (one row per warp iteration,
    // lanes = consecutive pieces).  Every thread makes the stage's mbarrier count one arrival
    // when its copies have landed; one plain arrive per warp publishes the row-valid flags.
    const int gw = warp - (E2_EPI_WARPS + E2_BLD_WARPS);
    const int so = gw / E2_GAT_SPLIT;                     // the stage this warp helps to fill
    const int sub = gw % E2_GAT_SPLIT;                    // ... rows r = sub (mod E2_GAT_SPLIT)
    const int ppr = rowbytes / 16;                        // 16-byte pieces per row (<= 44)
    const uint32_t on1 = lane < ppr, on2 = lane + 32 < ppr;
    const uint32_t stage = smem_u32(q_smem) + so * qstage_bytes + lane * 16;
    // base pointers and row pitches live in registers for the whole loop (the per-row address is
    // one 64-bit multiply-add); "+l" keeps the compiler from re-deriving them from the
    // parameter bank in front of every copy
    uint64_t qbase = reinterpret_cast<uint64_t>(p.qmat) + (uint64_t)lane * 16;
    uint64_t pbase = reinterpret_cast<uint64_t>(p.pmat) + (uint64_t)lane * 16;
    uint32_t qpitch = (uint32_t)p.ldq * 2, ppitch = (uint32_t)p.ldp * 2;
    asm volatile("" : "+l"(qbase), "+l"(pbase), "+r"(qpitch), "+r"(ppitch));
    int32_t j_nx = -1, pn_nx = -1;
    auto fetch_ids = [&](uint32_t g2) {
      j_nx = -1;
      pn_nx = -1;
      if (g2 >= total_halves) return;
      const int64_t tile = blockIdx.x + (int64_t)(g2 >> 1) * gridDim.x;
      const int hs = g2 & 1;
      const int rt = hs * E2_HALF + lane;                 // row in tile
      const int64_t grp = tile * NPT + rt / KP;
      if (rt < USED && grp < num_groups) j_nx = __ldg(p.ell + tile * USED + rt);
      const int nfirst = (hs * E2_HALF) / KP;
      const int64_t pg = tile * NPT + nfirst + lane;
      if (lane < PMAX && nfirst + lane < NPT && pg < num_groups)
        pn_nx = p.group_node ? __ldg(p.group_node + pg) : (int32_t)pg;
    };
    fetch_ids(so);
    for (uint32_t g2 = so; g2 < total_halves; g2 += E2_QSTAGES) {
      const int32_t j = j_nx, pn = pn_nx;
      fetch_ids(g2 + E2_QSTAGES);  // ids of this pair's next half-stage, in flight meanwhile
      mbar_wait_t<true>(&sh->q_empty[so], ((g2 / E2_QSTAGES) & 1) ^ 1, w0, timed);
      if (sub == 0) sh->row_valid[so][lane] = j >= 0 ? 1 : 0;
      // four rows per step (independent address chains); lanes = consecutive 16-byte pieces
#pragma unroll 8
      for (int r = sub; r < E2_HALF; r += E2_GAT_SPLIT) {
        const int32_t src = __shfl_sync(0xffffffffu, j, r);
        const uint32_t ok = src >= 0;
        const uint8_t* g = reinterpret_cast<const uint8_t*>(qbase + (uint64_t)(uint32_t)max(src, 0) * qpitch);
        const uint32_t d = stage + (uint32_t)(r * rowbytes);
        cp_async16_if(d, g, ok & on1);
        cp_async16_if(d + 512, g + 512, ok & on2);
      }
#pragma unroll
      for (int r = 0; r < PMAX; ++r) {
        const int32_t src = __shfl_sync(0xffffffffu, pn, r);
        const uint32_t ok = src >= 0 && (r % E2_GAT_SPLIT) == sub;
        const uint8_t* g = reinterpret_cast<const uint8_t*>(pbase + (uint64_t)(uint32_t)max(src, 0) * ppitch);
        const uint32_t d = stage + (uint32_t)((E2_HALF + r) * rowbytes);
        cp_async16_if(d, g, ok & on1);
        cp_async16_if(d + 512, g + 512, ok & on2);
      }
      cp_async_arrive_noinc(&sh->q_full[so]);
      __syncwarp();
      if (lane == 0) mbar_arrive(&sh->q_full[so]);
    }
  } else if (warp == E2_MMA_WARP) {
    // =================================================================== MMA issuer
    const uint32_t idesc = make_idesc_f16(128, E2_ROWS);
    const uint32_t h_addr = smem_u32(h_smem);
    for (int64_t it = 0; it < my_tiles; ++it) {
      const uint32_t hb = (uint32_t)(it % E2_HBUFS);
      mbar_wait_t(&sh->h_full[hb], (uint32_t)((it / E2_HBUFS) & 1), w0, timed);
      if (p.use_ws && nblk == 2) {
        // both channel blocks advance together, K step by K step: the second block takes the h
        // chunk from the collector buffer the first one filled (one shared-memory read, not two)
        mbar_wait_t(&sh->d_empty[0], (uint32_t)(it & 1) ^ 1, w1, timed);
        mbar_wait_t(&sh->d_empty[1], (uint32_t)(it & 1) ^ 1, w1, timed);
        tc_fence_after();
        if (elect_one()) {
          const uint32_t d0 = tmem_base + E2_TMEM_D0, d1 = tmem_base + E2_TMEM_D1;
          uint64_t bdesc = make_kmajor_desc<64>(h_addr + hb * htile_bytes);
          uint32_t a = tmem_base;
          mma_f16_ts_ws<0, true, false>(d0, a, bdesc, idesc);
          mma_f16_ts_ws<0, false, false>(d1, a + E2_TMEM_A1, bdesc, idesc);
          mma_f16_ts_ws<1, true, true>(d0, a + 8, bdesc + 2, idesc);
          mma_f16_ts_ws<1, false, true>(d1, a + E2_TMEM_A1 + 8, bdesc + 2, idesc);
#pragma unroll 1
          for (int kc = 1; kc < nch; ++kc) {
            bdesc += (E2_ROWS * 64) >> 4;
            a += 16;
            mma_f16_ts_ws<0, true, true>(d0, a, bdesc, idesc);
            mma_f16_ts_ws<0, false, true>(d1, a + E2_TMEM_A1, bdesc, idesc);
            mma_f16_ts_ws<1, true, true>(d0, a + 8, bdesc + 2, idesc);
            mma_f16_ts_ws<1, false, true>(d1, a + E2_TMEM_A1 + 8, bdesc + 2, idesc);
          }
          mma_commit(&sh->d_full[0]);
          mma_commit(&sh->d_full[1]);
          mma_commit(&sh->h_empty[hb]);
        }
        __syncwarp();
        continue;
      }
      for (int blk = 0; blk < nblk; ++blk) {
        mbar_wait_t(&sh->d_empty[blk], (uint32_t)(it & 1) ^ 1, w1, timed);
        tc_fence_after();
        if (elect_one()) {
          const uint32_t d_addr = tmem_base + (blk ? E2_TMEM_D1 : E2_TMEM_D0);
          const uint32_t a_addr = tmem_base + (blk ? E2_TMEM_A1 : 0);
          // tight issue loop: one descriptor per tile, then constant increments
          // (K chunk = 4096 B of B -> +256 in descriptor units, = 16 TMEM columns of A;
          //  half chunk (16 k) = 32 B -> +2, = 8 columns)
          uint64_t bdesc = make_kmajor_desc<64>(h_addr + hb * htile_bytes);
          uint32_t a = a_addr;
          mma_f16_ts_c<false>(d_addr, a, bdesc, idesc);
          mma_f16_ts_c<true>(d_addr, a + 8, bdesc + 2, idesc);
#pragma unroll 1
          for (int kc = 1; kc < nch; ++kc) {
            bdesc += (E2_ROWS * 64) >> 4;
            a += 16;
            mma_f16_ts_c<true>(d_addr, a, bdesc, idesc);
            mma_f16_ts_c<true>(d_addr, a + 8, bdesc + 2, idesc);
          }
          mma_commit(&sh->d_full[blk]);
          if (blk == nblk - 1) mma_commit(&sh->h_empty[hb]);
        }
        __syncwarp();
      }
    }
  } else if (warp >= E2_EPI_WARPS) {
    // ==================================================================== h builders
    // warp bw owns rows 16 bw .. 16 bw + 15 of the tile; lane = (row rr = lane >> 2, 16-byte
    // piece c4 = lane & 3 of the 64-byte K chunk), two rows per lane (rr and rr + 8).
    // Quarter-warps touch 2 rows x 64 B: conflict free.
    constexpr int RW = E2_ROWS / E2_BLD_WARPS;   // tile rows per builder warp
    constexpr int U = RW / 8;                    // rows per lane
    const int bw = warp - E2_EPI_WARPS;
    const int hs = (bw * RW) / E2_HALF;          // which gather stage of the tile
    const int rr = lane >> 2, c4 = lane & 3;
    const uint32_t q_addr = smem_u32(q_smem);
    const uint32_t h_addr = smem_u32(h_smem);
    const int nfirst = (hs * E2_HALF) / KP;
    int srow[U];
    uint32_t dst_off[U], q_off[U], p_off[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const int row = bw * RW + rr + 8 * u;           // row in tile
      srow[u] = row - hs * E2_HALF;                   // row in stage
      dst_off[u] = (uint32_t)(row * 64 + ((c4 ^ ((row >> 1) & 3)) << 4));
      q_off[u] = (uint32_t)(srow[u] * rowbytes + c4 * 16);
      const int pslot = row < USED ? row / KP - nfirst : 0;
      p_off[u] = (uint32_t)((E2_HALF + pslot) * rowbytes + c4 * 16);
    }
    for (int64_t it = 0; it < my_tiles; ++it) {
      const uint32_t g2 = (uint32_t)(2 * it + hs);
      const uint32_t s = g2 % E2_QSTAGES;
      const uint32_t hb = (uint32_t)(it % E2_HBUFS);
      mbar_wait_t(&sh->q_full[s], (g2 / E2_QSTAGES) & 1, w0, timed);
      bool valid[U];
#pragma unroll
      for (int u = 0; u < U; ++u) valid[u] = sh->row_valid[s][srow[u]] != 0;
      mbar_wait_t(&sh->h_empty[hb], (uint32_t)((it / E2_HBUFS) & 1) ^ 1, w1, timed);
      const uint32_t sbase = q_addr + s * qstage_bytes;
      const uint32_t dbase = h_addr + hb * htile_bytes;
      // four K chunks per step: all eight loads are issued before the first use
#pragma unroll 1
      for (int kc0 = 0; kc0 < nch; kc0 += 4) {
        uint4 Q[4][U], P[4][U];
#pragma unroll
        for (int i = 0; i < 4; ++i) {
#pragma unroll
          for (int u = 0; u < U; ++u) {
            const bool ld = valid[u] && kc0 + i < nch;
            Q[i][u] = ld ? lds128(sbase + q_off[u] + (kc0 + i) * 64) : make_uint4(0, 0, 0, 0);
            P[i][u] = ld ? lds128(sbase + p_off[u] + (kc0 + i) * 64) : make_uint4(0, 0, 0, 0);
          }
        }
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          if (kc0 + i < nch) {
#pragma unroll
            for (int u = 0; u < U; ++u) {
              // empty slots: P = Q = 0 -> h = 0 (the "ones" columns included)
              asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(
                               dbase + dst_off[u] + (kc0 + i) * (E2_ROWS * 64)),
                           "r"(relu_add_f16x2(P[i][u].x, Q[i][u].x)),
                           "r"(relu_add_f16x2(P[i][u].y, Q[i][u].y)),
                           "r"(relu_add_f16x2(P[i][u].z, Q[i][u].z)),
                           "r"(relu_add_f16x2(P[i][u].w, Q[i][u].w))
                           : "memory");
            }
          }
        }
      }
      fence_proxy_async_smem();
      __syncwarp();
      if (lane == 0) {
        mbar_arrive(&sh->h_full[hb]);
        mbar_arrive(&sh->q_empty[s]);
      }
    }
  } else {
    // ====================================================================== epilogue
    // warp (blk, q) reads TMEM lanes 32q.. of channel block blk: lane = output channel, the 64
    // columns are the tile's edge rows, KP consecutive columns per node -> plain in-register
    // sums; the 32 lanes store 32 consecutive channels of a node (coalesced)
    const int blk = warp >> 2, q = warp & 3;
    if (blk < nblk) {
      const int ch = blk * 128 + q * 32 + lane;
      const bool ch_ok = ch < p.b;
      int crank = -1;  // position of this channel among the kNN coordinate columns
      if (p.coords != nullptr && ch_ok && ((p.coord_mask[ch >> 5] >> (ch & 31)) & 1u)) {
        crank = __popc(p.coord_mask[ch >> 5] & ((1u << (ch & 31)) - 1u));
        for (int w = 0; w < (ch >> 5); ++w) crank += __popc(p.coord_mask[w]);
      }
      const uint32_t d_addr =
          tmem_base + ((uint32_t)(q * 32) << 16) + (blk ? E2_TMEM_D1 : E2_TMEM_D0);
      uint16_t* const out_ch = reinterpret_cast<uint16_t*>(p.out) + ch;
      for (int64_t it = 0; it < my_tiles; ++it) {
        const int64_t tile = blockIdx.x + it * gridDim.x;
        const int64_t grp0 = tile * NPT;
        const int nvalid = (int)min((int64_t)NPT, num_groups - grp0);  // groups of this tile
        mbar_wait_t<true>(&sh->d_full[blk], (uint32_t)(it & 1), w0, timed);
        tc_fence_after();
        float sums[NPT];
#pragma unroll
        for (int n = 0; n < NPT; ++n) sums[n] = 0.f;
#pragma unroll
        for (int half = 0; half < 2; ++half) {   // 32 columns at a time: half the registers
          float v[32];
          tmem_ld_32x32(d_addr + 32 * half, v);
#pragma unroll
          for (int c = 0; c < 32; ++c) {
            const int col = 32 * half + c;
            if (col < USED) sums[col / KP] += fmaxf(v[c], 0.f);
          }
          if (!ACC && p.relu_mask != nullptr) {
            // training: the ReLU pattern of the 32 channels of this warp, one word per tile row
            // (the backward pass masks dOut with it instead of recomputing Z = H W2^T)
            uint32_t mword = 0;
#pragma unroll
            for (int c = 0; c < 32; ++c) {
              const uint32_t w = __ballot_sync(0xffffffffu, v[c] > 0.f);
              if (lane == c) mword = w;
            }
            p.relu_mask[(int64_t)(blk * 4 + q) * p.mask_ld + tile * E2_ROWS + 32 * half + lane] =
                mword;
          }
        }
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(&sh->d_empty[blk]);  // values are in registers now
        if (ch_ok) {
          if (!ACC) {
            // pass 1: group g is node g -> consecutive rows of `out`
            uint16_t* dst = out_ch + grp0 * p.ldo;
#pragma unroll
            for (int n = 0; n < NPT; ++n) {
              if (n < nvalid) dst[(int64_t)n * p.ldo] = pack1_16(sums[n], F16);
            }
            if (crank >= 0) {
#pragma unroll
              for (int n = 0; n < NPT; ++n)
                if (n < nvalid) p.coords[(grp0 + n) * p.ncoord + crank] = sums[n];
            }
          } else {
#pragma unroll
            for (int n = 0; n < NPT; ++n) {
              if (n < nvalid) {
                const int64_t node = __ldg(p.group_node + grp0 + n);
                red_add1_16(out_ch + node * p.ldo, sums[n], F16);
                if (crank >= 0) atomicAdd(p.coords + node * p.ncoord + crank, sums[n]);
              }
            }
          }
        }
      }
    }
  }
  if (timed && lane == 0) {
    unsigned long long* d = reinterpret_cast<unsigned long long*>(p.dbg);
    if (warp >= E2_EPI_WARPS + E2_BLD_WARPS && warp < E2_MMA_WARP) {
      atomicAdd(d + 0, (unsigned long long)w0);
    } else if (warp == E2_MMA_WARP) {
      atomicAdd(d + 3, (unsigned long long)w0);
      atomicAdd(d + 4, (unsigned long long)w1);
      atomicAdd(d + 6, (unsigned long long)(clock64() - t_start));
      atomicAdd(d + 7, (unsigned long long)my_tiles);
    } else if (warp >= E2_EPI_WARPS) {
      atomicAdd(d + 1, (unsigned long long)w0);
      atomicAdd(d + 2, (unsigned long long)w1);
    } else {
      atomicAdd(d + 5, (unsigned long long)w0);
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == E2_MMA_WARP) tmem_dealloc<512>(tmem_base);
}

inline size_t edge2_smem_bytes(int Kp, int slots) {
  const int nch = Kp / 32;
  return (size_t)E2_HBUFS * nch * (E2_ROWS * 64) +
         (size_t)E2_QSTAGES * (E2_HALF + e2_pmax(slots)) * (Kp * 2) + sizeof(Edge2Shared) + 1024;
}
inline bool edge2_supported_slots(int s) { return s == 8 || s == 9 || s == 16 || s == 17 || s == 32; }

}  // namespace tc
}  // namespace gn
