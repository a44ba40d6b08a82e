// Fused EdgeConv message + aggregation, third generation: the second-generation formulation
// (tc_edge2.cuh: W2 resident in tensor memory as the A operand, edges as accumulator columns)
// with the shared-memory traffic cut to what the tensor core itself needs.
//
// What the second generation was bound by (timeline of one CTA + micro-benchmarks):
// shared memory moves ~125 B/clk per SM in total, and a tile cost 276 KB of it - gather writes
// (51 KB), builder reads of Q and P (90 KB), builder writes of h (45 KB), and the tensor core
// reading the h tile once per channel block (90 KB): 2200 clk of shared-memory time per tile,
// 3000 measured.  Here
//   * the gather goes global/L2 -> REGISTERS -> h tile: no staging ring, no gather warps.  Sixteen
//     builder warps in two sets alternate tiles, so one set's loads are in flight while the other
//     set adds, clamps and stores; a builder never has loads outstanding when it executes
//     fence.proxy.async (that fence waits for them), so nothing is software-pipelined inside a
//     warp - the second set is the pipelining;
//   * the two channel blocks of a K step share one read of the h chunk (tcgen05.mma.ws, the B
//     operand latched in a collector buffer by the first block and re-used by the second);
//   * h is a 128B-swizzled K-major tile: a quarter-warp stores one row's 128 contiguous bytes
//     (conflict free), and a warp-wide 16-byte load touches four rows x 128 B = four L1 lines.
// Shared-memory traffic per tile: 45 KB (h written) + 45 KB (h read once) = 90 KB.
//
// STATUS (round 1): OPT-IN (GN_EDGE_V3=1), parity-green, not yet faster than the second
// generation (L2 conv of C2: 120-128 us against 112 us).  The shared-memory pressure is gone,
// but with both channel blocks finishing together (the .ws pairing) the single-buffered
// accumulators serialise tensor core and epilogue - MMA 1670 clk, then ~1000 clk until the
// eight epilogue warps have drained 64 KB of TMEM at 64 B/clk and released it - and TMEM has no
// room for a second accumulator pair (W2 352 + D 2x128 columns > 512).  See DESIGN.md.
//
// Warp roles (25 warps): 0-7 epilogue (warp>>2 = channel block of 128, warp&3 = TMEM lane
// quadrant), 8-23 builders (set = (warp-8)>>3 owns tiles of that parity and h buffer `set`; warp
// (warp-8)&7 of the set owns tile rows 8w..8w+7), 24 tcgen05.mma issuer.  TMEM columns as in the
// second generation: [0,Kp/2) W2 block 0, [176,176+Kp/2) W2 block 1, [352,416) D block 0,
// [416,480) D block 1.
#pragma once

namespace gn {
namespace tc {

constexpr int E3_ROWS = 64;          // edge rows (accumulator columns) per tile
constexpr int E3_EPI_WARPS = 8;
constexpr int E3_SET_WARPS = 8;      // builder warps per set: 8 tile rows each
constexpr int E3_SETS = 2;           // = h buffers
constexpr int E3_BLD_WARPS = E3_SET_WARPS * E3_SETS;
constexpr int E3_MMA_WARP = E3_EPI_WARPS + E3_BLD_WARPS;  // warp 24
constexpr int E3_THREADS = 32 * (E3_MMA_WARP + 1);        // 800
constexpr int E3_CHUNK_BYTES = E3_ROWS * 128;             // one 64-column K chunk of the h tile
constexpr int E3_KHALF = 3;          // 128-byte chunks a builder keeps in flight at a time

struct __align__(16) Edge3Shared {
  uint64_t h_full[E3_SETS];
  uint64_t h_empty[E3_SETS];
  uint64_t d_full[2];
  uint64_t d_empty[2];
  uint32_t tmem_base;
  uint32_t pad_;
};

__device__ __forceinline__ uint4 ldg_stream128(const void* p) {   // read once: keep out of L1
  uint4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.b32 {%0, %1, %2, %3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
               : "l"(p));
  return r;
}
__device__ __forceinline__ int32_t ldg_now32(const int32_t* p) {
  int32_t r;
  asm volatile("ld.global.nc.b32 %0, [%1];" : "=r"(r) : "l"(p));
  return r;
}
__device__ __forceinline__ uint4 ldg_keep128(const void* p) {     // re-read by neighbours: L1
  uint4 r;
  asm volatile("ld.global.nc.v4.b32 {%0, %1, %2, %3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
               : "l"(p));
  return r;
}

// ACC: second pass (group lists, red.add into out); F16: output element type (else bf16)
template <int KP, bool ACC, bool F16>
__global__ void __launch_bounds__(E3_THREADS, 1) edge_mlp_aggregate_v3_kernel(const EdgeParams p) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) &
                                             ~uintptr_t(1023));
  constexpr int NPT = E3_ROWS / KP;    // nodes (groups) per tile; KP need not be a power of two
  constexpr int USED = NPT * KP;       // tile rows that carry slots (the rest stay zero)
  const int ksteps = p.Kp / 16;                           // MMAs per channel block per tile
  const int ppr = p.Kp / 8;                               // 16-byte pieces per row
  const int nchunk = (p.Kp + 63) / 64;                    // 128-byte K chunks of the h tile
  const int htile_bytes = nchunk * E3_CHUNK_BYTES;
  uint8_t* h_smem = smem;                                 // E3_SETS tiles
  Edge3Shared* sh = reinterpret_cast<Edge3Shared*>(h_smem + E3_SETS * htile_bytes);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int64_t num_groups = p.group_count ? (int64_t)__ldg(p.group_count) : p.N;
  const int64_t num_tiles = (num_groups + NPT - 1) / NPT;
  // A CTA takes a CONTIGUOUS range of tiles: consecutive tiles are consecutive nodes of the same
  // event, whose neighbours are the same few dozen rows of Q - every Q row is wanted ~k times
  // within a few tiles, and with the range on one SM all but the first request hit its L1
  // (an interleaved assignment sent those k requests to k different SMs, i.e. all to L2).
  const int64_t tiles_per_cta = (num_tiles + gridDim.x - 1) / gridDim.x;
  const int64_t tile0 = (int64_t)blockIdx.x * tiles_per_cta;
  if (tile0 >= num_tiles) return;  // uniform per CTA, before any barrier
  const int nblk = p.bnp > 128 ? 2 : 1;

  if (threadIdx.x == 0) {
#pragma unroll
    for (int b = 0; b < 2; ++b) {
      mbar_init(&sh->h_full[b], E3_SET_WARPS);
      mbar_init(&sh->h_empty[b], 1);
      mbar_init(&sh->d_full[b], 1);
      mbar_init(&sh->d_empty[b], E3_EPI_WARPS / 2);
    }
    fence_mbar_init();
  }
  if (warp == E3_MMA_WARP) tmem_alloc<512>(&sh->tmem_base);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = sh->tmem_base;

  // ---- W2 -> tensor memory (once): epilogue warp (blk, q) writes lanes 32q.. of block blk;
  // each lane reads its own weight row with 16-byte loads
  if (warp < E3_EPI_WARPS) {
    const int blk = warp >> 2, q = warp & 3;
    if (blk < nblk) {
      const int ch = blk * 128 + q * 32 + lane;
      const uint4* wrow = reinterpret_cast<const uint4*>(
          reinterpret_cast<const __half*>(p.w2) + (int64_t)ch * p.Kp);
      const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + (blk ? E2_TMEM_A1 : 0);
      for (int cc = 0; cc < p.nch; ++cc) {  // 4 x 16 bytes = 32 fp16 = 16 columns
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

  const int64_t my_tiles = min(tiles_per_cta, num_tiles - tile0);
  // optional wait-time accounting (profiling runs): dbg[1] builders on h_empty, dbg[3]/[4] MMA on
  // h_full / d_empty, dbg[5] epilogue on d_full, dbg[6] total cycles, dbg[7] tiles
  const bool timed = p.dbg != nullptr;
  long long w0 = 0, w1 = 0;
  const long long t_start = clock64();
  // timeline of CTA 0 (profiling runs): stamp k of this warp's iteration i
  unsigned long long* const trace =
      (timed && blockIdx.x == 0 && lane == 0)
          ? reinterpret_cast<unsigned long long*>(p.dbg) + 8 + warp * (32 * 4) : nullptr;
  auto stamp = [&](int64_t i, int k) {
    if (trace != nullptr && i < 32) trace[i * 4 + k] = (unsigned long long)clock64();
  };

  if (warp == E3_MMA_WARP) {
    // =================================================================== MMA issuer
    const uint32_t idesc = make_idesc_f16(128, E3_ROWS);
    const uint32_t h_addr = smem_u32(h_smem);
    const uint32_t d0 = tmem_base + E2_TMEM_D0, d1 = tmem_base + E2_TMEM_D1;
    for (int64_t it = 0; it < my_tiles; ++it) {
      const uint32_t hb = (uint32_t)(it % E3_SETS);
      stamp(it, 0);
      mbar_wait_t(&sh->h_full[hb], (uint32_t)((it / E3_SETS) & 1), w0, timed);
      stamp(it, 1);
      const bool ws = nblk == 2 && p.use_ws;
      mbar_wait_t(&sh->d_empty[0], (uint32_t)(it & 1) ^ 1, w1, timed);
      if (ws) mbar_wait_t(&sh->d_empty[1], (uint32_t)(it & 1) ^ 1, w1, timed);
      tc_fence_after();
      stamp(it, 2);
      if (elect_one()) {
        // K step s: B = 32 bytes at offset (s&3)*32 of the 128-byte rows of chunk s>>2 (descriptor
        // units of 16 B: +2 per step, +512 per chunk); A = TMEM columns 8s..8s+7 of each block.
        // Both channel blocks take the step back to back, the second from the collector buffer.
        // The issue loop is the tensor core's pacemaker (one thread, ~32 clk per N=64 MMA): it is
        // unrolled by K chunk (four steps) so that an MMA costs two or three integer instructions.
        const uint64_t bdesc0 = make_kmajor_desc<128>(h_addr + hb * htile_bytes);
        if (ws) {
          uint64_t bd = bdesc0;
          uint32_t a = tmem_base;
#pragma unroll 1
          for (int s0 = 0; s0 < ksteps; s0 += 4) {
            const uint32_t acc0 = s0 > 0;
            mma_f16_ts_ws_p<0, true>(d0, a, bd, idesc, acc0);
            mma_f16_ts_ws_p<0, false>(d1, a + E2_TMEM_A1, bd, idesc, acc0);
            mma_f16_ts_ws<1, true, true>(d0, a + 8, bd + 2, idesc);
            mma_f16_ts_ws<1, false, true>(d1, a + E2_TMEM_A1 + 8, bd + 2, idesc);
            if (s0 + 2 < ksteps) {   // K is a multiple of 32: steps come in pairs
              mma_f16_ts_ws<0, true, true>(d0, a + 16, bd + 4, idesc);
              mma_f16_ts_ws<0, false, true>(d1, a + E2_TMEM_A1 + 16, bd + 4, idesc);
              mma_f16_ts_ws<1, true, true>(d0, a + 24, bd + 6, idesc);
              mma_f16_ts_ws<1, false, true>(d1, a + E2_TMEM_A1 + 24, bd + 6, idesc);
            }
            bd += E3_CHUNK_BYTES >> 4;
            a += 32;
          }
          mma_commit(&sh->d_full[0]);
          mma_commit(&sh->d_full[1]);
        } else {
          for (int blk = 0; blk < nblk; ++blk) {
            if (blk == 1) {  // block 1's accumulator: drained while block 0 was being issued
              mbar_wait(&sh->d_empty[1], (uint32_t)(it & 1) ^ 1);
              tc_fence_after();
            }
            const uint32_t d = blk ? d1 : d0;
            uint64_t bd = bdesc0;
            uint32_t a = tmem_base + (blk ? E2_TMEM_A1 : 0);
#pragma unroll 1
            for (int s0 = 0; s0 < ksteps; s0 += 4) {
              mma_f16_ts(d, a, bd, idesc, s0 > 0);
              mma_f16_ts_c<true>(d, a + 8, bd + 2, idesc);
              if (s0 + 2 < ksteps) {
                mma_f16_ts_c<true>(d, a + 16, bd + 4, idesc);
                mma_f16_ts_c<true>(d, a + 24, bd + 6, idesc);
              }
              bd += E3_CHUNK_BYTES >> 4;
              a += 32;
            }
            mma_commit(&sh->d_full[blk]);
          }
        }
        mma_commit(&sh->h_empty[hb]);
      }
      __syncwarp();
      stamp(it, 3);
    }
  } else if (warp >= E3_EPI_WARPS) {
    // ==================================================================== h builders
    // lane = (r4 = lane >> 3, c8 = lane & 7): rows rowA = 8w + r4 and rowB = rowA + 4 of the tile,
    // 16-byte piece c8 of every 128-byte K chunk.  One warp-wide load = 4 rows x 128 B.
    const int bw = warp - E3_EPI_WARPS;
    const int set = bw / E3_SET_WARPS, w8 = bw % E3_SET_WARPS;
    const int r4 = lane >> 3, c8 = lane & 7;
    int row[2];
    uint32_t dst_off[2];
    bool used[2];
#pragma unroll
    for (int u = 0; u < 2; ++u) {
      row[u] = w8 * 8 + r4 + 4 * u;
      used[u] = row[u] < USED;
      dst_off[u] = (uint32_t)(row[u] * 128 + ((c8 ^ (row[u] & 7)) << 4));
    }
    // bases and pitches pinned in registers (one 64-bit multiply-add per row address)
    uint64_t qbase = reinterpret_cast<uint64_t>(p.qmat) + (uint64_t)c8 * 16;
    uint64_t pbase = reinterpret_cast<uint64_t>(p.pmat) + (uint64_t)c8 * 16;
    uint32_t qpitch = (uint32_t)p.ldq * 2, ppitch = (uint32_t)p.ldp * 2;
    asm volatile("" : "+l"(qbase), "+l"(pbase), "+r"(qpitch), "+r"(ppitch));
    const uint32_t h_base = smem_u32(h_smem) + set * htile_bytes;

    int32_t j_nx[2] = {-1, -1}, n_nx[2] = {-1, -1};
    auto fetch_ids = [&](int64_t it) {
#pragma unroll
      for (int u = 0; u < 2; ++u) {
        j_nx[u] = -1;
        n_nx[u] = -1;
        if (it < my_tiles && used[u]) {
          const int64_t tile = tile0 + it;
          const int64_t grp = tile * NPT + row[u] / KP;
          if (grp < num_groups) {
            // volatile: issued HERE, a tile ahead of their use (the compiler would sink a plain
            // load to its first use and expose the whole latency there)
            j_nx[u] = ldg_now32(p.ell + tile * USED + row[u]);
            n_nx[u] = p.group_node ? ldg_now32(p.group_node + grp) : (int32_t)grp;
          }
        }
      }
    };
    fetch_ids(set);
    for (int64_t it = set; it < my_tiles; it += E3_SETS) {
      const uint8_t* qrow[2];
      const uint8_t* prow[2];
      bool valid[2];
#pragma unroll
      for (int u = 0; u < 2; ++u) {
        valid[u] = j_nx[u] >= 0;
        qrow[u] = reinterpret_cast<const uint8_t*>(qbase + (uint64_t)(uint32_t)max(j_nx[u], 0) * qpitch);
        prow[u] = reinterpret_cast<const uint8_t*>(pbase + (uint64_t)(uint32_t)max(n_nx[u], 0) * ppitch);
      }
      fetch_ids(it + E3_SETS);  // next tile's ids: landed long before this tile's fence
      stamp(it / E3_SETS, 0);
#pragma unroll 1
      for (int c0 = 0; c0 < nchunk; c0 += E3_KHALF) {
        uint4 Q[E3_KHALF][2], P[E3_KHALF][2];
#pragma unroll
        for (int i = 0; i < E3_KHALF; ++i) {
#pragma unroll
          for (int u = 0; u < 2; ++u) {
            // empty slots: Q = -inf, P = 0 -> ReLU(P + Q) = 0 (the "ones" columns included)
            Q[i][u] = make_uint4(0xFC00FC00u, 0xFC00FC00u, 0xFC00FC00u, 0xFC00FC00u);
            P[i][u] = make_uint4(0, 0, 0, 0);
            if (valid[u] && (c0 + i) * 8 + c8 < ppr) {
              Q[i][u] = ldg_keep128(qrow[u] + (c0 + i) * 128);
              P[i][u] = ldg_keep128(prow[u] + (c0 + i) * 128);
            }
          }
        }
        if (c0 == 0)  // the tensor core has finished with this buffer's previous tile
          mbar_wait_t(&sh->h_empty[set], (uint32_t)((it / E3_SETS) & 1) ^ 1, w1, timed);
#pragma unroll
        for (int i = 0; i < E3_KHALF; ++i) {
          if ((c0 + i) * 8 + c8 < ppr) {
#pragma unroll
            for (int u = 0; u < 2; ++u) {
              asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(
                               h_base + dst_off[u] + (c0 + i) * E3_CHUNK_BYTES),
                           "r"(relu_add_f16x2(P[i][u].x, Q[i][u].x)),
                           "r"(relu_add_f16x2(P[i][u].y, Q[i][u].y)),
                           "r"(relu_add_f16x2(P[i][u].z, Q[i][u].z)),
                           "r"(relu_add_f16x2(P[i][u].w, Q[i][u].w))
                           : "memory");
            }
          }
        }
        stamp(it / E3_SETS, c0 == 0 ? 1 : 2);
      }
      fence_proxy_async_smem();
      __syncwarp();
      if (lane == 0) mbar_arrive(&sh->h_full[set]);
      stamp(it / E3_SETS, 3);
    }
  } else {
    // ====================================================================== epilogue
    // warp (blk, q) reads TMEM lanes 32q.. of channel block blk: lane = output channel, the 64
    // columns are the tile's edge rows, KP consecutive columns per node -> plain in-register
    // sums, 32 columns at a time; the 32 lanes store 32 consecutive channels of a node
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
        const int64_t tile = tile0 + it;
        const int64_t grp0 = tile * NPT;
        const int nvalid = (int)min((int64_t)NPT, num_groups - grp0);  // groups of this tile
        stamp(it, 0);
        mbar_wait_t<true>(&sh->d_full[blk], (uint32_t)(it & 1), w0, timed);
        stamp(it, 1);
        tc_fence_after();
        float sums[NPT];
#pragma unroll
        for (int n = 0; n < NPT; ++n) sums[n] = 0.f;
#pragma unroll
        for (int half = 0; half < 2; ++half) {
          float v[32];
          tmem_ld_32x32(d_addr + 32 * half, v);
#pragma unroll
          for (int c = 0; c < 32; ++c) {
            const int col = 32 * half + c;
            if (col < USED) sums[col / KP] += fmaxf(v[c], 0.f);
          }
        }
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(&sh->d_empty[blk]);  // values are in registers now
        stamp(it, 2);
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
        stamp(it, 3);
      }
    }
  }
  if (timed && lane == 0) {
    unsigned long long* d = reinterpret_cast<unsigned long long*>(p.dbg);
    if (warp == E3_MMA_WARP) {
      atomicAdd(d + 3, (unsigned long long)w0);
      atomicAdd(d + 4, (unsigned long long)w1);
      atomicAdd(d + 6, (unsigned long long)(clock64() - t_start));
      atomicAdd(d + 7, (unsigned long long)my_tiles);
    } else if (warp >= E3_EPI_WARPS) {
      atomicAdd(d + 2, (unsigned long long)w1);
    } else {
      atomicAdd(d + 5, (unsigned long long)w0);
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == E3_MMA_WARP) tmem_dealloc<512>(tmem_base);
}

inline size_t edge3_smem_bytes(int Kp) {
  const int nchunk = (Kp + 63) / 64;
  return (size_t)E3_SETS * nchunk * E3_CHUNK_BYTES + sizeof(Edge3Shared) + 1024;
}

}  // namespace tc
}  // namespace gn
