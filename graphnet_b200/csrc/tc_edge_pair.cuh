// Fused EdgeConv message + aggregation, CTA-pair generation.
//
//   D^T[ch, e] = sum_k W2[ch, k] * h[e, k],   h[e] = act(P_i(e) + Q_j(e))
//   out[i, ch] = aggr_{e in N(i)} act(D^T[ch, e])
//
// Two CTAs on the two SMs of a TPC (a cluster of 2) run ONE tcgen05.mma.cta_group::2 stream:
// M = 256 output channels, 128 per CTA, so each SM keeps only ITS half of W2 in tensor memory
// (Kp/2 <= 176 columns instead of 352) and the 512 columns have room for TWO 128-column
// accumulators: the epilogue of one step drains while the MMAs of the next are running.  That
// was the limit of the single-CTA kernels (W2 352 + D 128 columns, single-buffered: tensor core
// and epilogue took turns, tensor pipe 46 %).  N = 128 edge rows per step: each CTA builds the h
// tile of its own 64 ELL rows (the 64-row tile of the ReLU-mask / slot layout: CTA r of the
// pair takes tile 2*T + r) and the hardware hands both halves to both tensor cores, so every h
// byte is written once and read once per SM pair.
//
// Shared-memory traffic is what the tensor core itself needs: the gather goes global/L2 ->
// REGISTERS -> h tile.  A builder warp takes one target node at a time: lanes = consecutive
// 16-byte pieces of the Kp-wide row (512 contiguous bytes per warp-wide load), the node's P
// piece stays in registers for all of its KP slots (P is read once per node, not once per
// edge) and the KP neighbour rows Q_j are all in flight before the first one is used.
//
// Warp roles (26 warps per CTA): 0-7 epilogue (warp & 3 = TMEM lane quadrant = 32 channels,
// warp >> 2 = which 64 columns = which CTA's tile), 8-23 builders (one queue of tasks over all
// steps, four h tiles in flight), 25 = TMEM allocation, and in the leader CTA the tcgen05.mma
// issuer.  TMEM columns: [0, Kp/2) this CTA's half of W2, [192, 320) D buffer 0,
// [320, 448) D buffer 1.
//
// Barriers: h_full / d_empty live in the LEADER's shared memory and count arrivals of both
// CTAs (remote mbarrier.arrive through the cluster address); h_empty / d_full are signalled in
// both CTAs at once by tcgen05.commit ... .multicast::cluster.
#pragma once

namespace gn {
namespace tc {

constexpr int E4_ROWS = 64;          // ELL rows (accumulator columns) per CTA per step
constexpr int E4_EPI_WARPS = 8;
constexpr int E4_SET_WARPS = 8;      // builder warps per set
constexpr int E4_SETS = 2;           // builder sets (steps of one parity each) = accumulator buffers
constexpr int E4_HBUFS = 4;          // h tiles in flight: a set fills buffer it & 3 for step it
constexpr int E4_BLD_WARPS = E4_SET_WARPS * E4_SETS;
// The MMA issuer is warp 25, not 24: warp w is scheduled by sub-partition w & 3, and epilogue warps
// 0 and 4 (channels 0..31, the ones that also write the kNN coordinates) are the busiest of
// their role - the issuer stays off their sub-partition (3 us per launch at C2).  Warp 24 only
// helps with the W2 upload.
constexpr int E4_MMA_WARP = E4_EPI_WARPS + E4_BLD_WARPS + 1;
constexpr int E4_THREADS = 32 * (E4_MMA_WARP + 1);        // 832
constexpr int E4_CHUNK_BYTES = E4_ROWS * 128;             // one 64-column K chunk of the h tile
constexpr int E4_TMEM_D0 = 192;
constexpr int E4_TMEM_D1 = 320;
constexpr int E4_SG = 9;             // neighbour rows a builder keeps in flight at a time

enum : int { E4_AGGR_ADD = 0, E4_AGGR_MEAN = 1, E4_AGGR_MAX = 2, E4_AGGR_MIN = 3 };

struct __align__(16) Edge4Shared {
  uint64_t h_full[E4_HBUFS];   // used in the leader CTA only
  uint64_t h_empty[E4_HBUFS];
  uint64_t d_full[E4_SETS];
  uint64_t d_empty[E4_SETS];   // used in the leader CTA only
  uint32_t tmem_base;
  uint32_t pad_;
};

// ------------------------------------------------------------ cluster / cta_group::2 PTX
__device__ __forceinline__ uint32_t cluster_ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ void cluster_sync_all() {   // every thread of both CTAs
  asm volatile("barrier.cluster.arrive.release.aligned;\n\t"
               "barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
// address of the same shared-memory variable in CTA `rank` of the cluster
__device__ __forceinline__ uint32_t map_to_cta(uint32_t smem_addr, uint32_t rank) {
  uint32_t r;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(smem_addr), "r"(rank));
  return r;
}
// (the default .cta-scoped forms on both sides: .release.cluster makes ptxas put MEMBAR.ALL.GPU in
// front of the arrive - it waits for every outstanding global access of the warp, the epilogue's
// stores, the builders' prefetches, on the critical path of every step - and .acquire.cluster on
// the waiting side adds CCTL.IVALL, which empties the L1 the builders gather through.  What an
// arrival publishes here is ordered by other means: shared-memory writes by fence.proxy.async,
// TMEM reads by tcgen05.wait::ld + tcgen05.fence::before_thread_sync.)
__device__ __forceinline__ void mbar_arrive_cluster(uint32_t cluster_addr) {
  asm volatile("mbarrier.arrive.shared::cluster.b64 _, [%0];" ::"r"(cluster_addr) : "memory");
}
template <int NCOLS>
__device__ __forceinline__ void tmem_alloc_pair(uint32_t* smem_dst) {  // one warp in EACH CTA
  asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(
                   smem_u32(smem_dst)),
               "n"(NCOLS)
               : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
}
template <int NCOLS>
__device__ __forceinline__ void tmem_dealloc_pair(uint32_t taddr) {
  asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(taddr), "n"(NCOLS)
               : "memory");
}
// D[tmem, both CTAs] (+)= A[tmem, both CTAs] * B[smem, half the rows in each CTA]^T
__device__ __forceinline__ void mma_f16_ts_pair(uint32_t tmem_d, uint32_t tmem_a, uint64_t bdesc,
                                                uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::2.kind::f16 [%0], [%1], %2, %3, p;\n\t}" ::"r"(tmem_d),
      "r"(tmem_a), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}
// one arrival on the mbarrier at this shared-memory offset in BOTH CTAs once every MMA issued so
// far has completed
__device__ __forceinline__ void mma_commit_pair(uint64_t* bar) {
  asm volatile(
      "tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 "
      "[%0], %1;" ::"r"(smem_u32(bar)),
      "h"((uint16_t)3)
      : "memory");
}

__device__ __forceinline__ uint4 e4_ldg128(const void* p) {   // re-read by neighbours: keep in L1
  uint4 r;
  asm volatile("ld.global.nc.v4.b32 {%0, %1, %2, %3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
               : "l"(p));
  return r;
}
__device__ __forceinline__ int32_t e4_ldg32(const int32_t* p) {  // volatile: issued where written
  int32_t r;
  asm volatile("ld.global.nc.b32 %0, [%1];" : "=r"(r) : "l"(p));
  return r;
}

// h = act(P + Q) on two packed fp16 values (`act` is uniform over the grid)
__device__ __forceinline__ uint32_t e4_act_add(uint32_t pv, uint32_t qv, int act) {
  if (act == ACT_RELU) return relu_add_f16x2(pv, qv);
  __half2 s = __hadd2(*reinterpret_cast<__half2*>(&pv), *reinterpret_cast<__half2*>(&qv));
  if (act == ACT_LEAKY_RELU) {
    s = __hmax2(s, __hmul2(s, __float2half2_rn(0.01f)));
  } else if (act == ACT_GELU) {
    float2 f = __half22float2(s);
    f.x = activate<ACT_GELU>(f.x);
    f.y = activate<ACT_GELU>(f.y);
    s = __floats2half2_rn(f.x, f.y);
  }
  return *reinterpret_cast<uint32_t*>(&s);
}
__device__ __forceinline__ float e4_act(float v, int act) {
  if (act == ACT_RELU) return fmaxf(v, 0.f);
  if (act == ACT_LEAKY_RELU) return activate<ACT_LEAKY_RELU>(v);
  if (act == ACT_GELU) return activate<ACT_GELU>(v);
  return v;
}

// ACC: second pass (group lists, red.add into out); F16: output element type (else bf16);
// GEN: activation and aggregation taken from the parameters (p.act: ReLU / LeakyReLU / GELU for
// both Linear layers of the edge MLP, p.aggr: add / mean / max / min); the default instance
// (GEN = false) is ReLU + add with nothing to decide per element
// MASK: training - also record the ReLU pattern of the second Linear (p.relu_mask)
template <int KP, bool ACC, bool F16, bool GEN, bool MASK>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(E4_THREADS, 1)
    edge_mlp_aggregate_pair_kernel(const EdgeParams p) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) &
                                             ~uintptr_t(1023));
  constexpr int NPT = E4_ROWS / KP;    // nodes (groups) per 64-row tile
  constexpr int USED = NPT * KP;       // tile rows that carry slots (the rest stay zero)
  const int ksteps = p.Kp / 16;                           // MMAs per step
  const int ppr = p.Kp / 8;                               // 16-byte pieces per row
  const int nchunk = (p.Kp + 63) / 64;                    // 128-byte K chunks of the h tile
  const int htile_bytes = nchunk * E4_CHUNK_BYTES;
  uint8_t* h_smem = smem;                                 // E4_SETS tiles
  Edge4Shared* sh = reinterpret_cast<Edge4Shared*>(h_smem + E4_HBUFS * htile_bytes);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t rank = cluster_ctarank();                // 0 = leader (issues the MMAs)
#ifdef GN_DEBUG_BUILD
  const long long t_entry = clock64();
#endif
  const int64_t num_groups =
      p.group_count ? min((int64_t)__ldg(p.group_count), p.group_cap) : p.N;
  const int64_t num_tiles = (num_groups + NPT - 1) / NPT; // 64-row tiles
  const int64_t num_steps = (num_tiles + 1) / 2;          // a step = two tiles, one per CTA
  // a cluster takes a CONTIGUOUS range of steps: consecutive tiles are consecutive nodes of the
  // same event, whose neighbours are the same few dozen rows of Q - all but the first request
  // for a row hit this SM's L1
  const int64_t clusters = gridDim.x >> 1, cid = blockIdx.x >> 1;
  const int64_t steps_per = (num_steps + clusters - 1) / clusters;
  const int64_t step0 = cid * steps_per;
  if (step0 >= num_steps) return;      // uniform per cluster, before any barrier
  const int64_t my_steps = min(steps_per, num_steps - step0);

  if (threadIdx.x == 0) {
    // builder tasks per tile (see the builder section): every task arrives once, in both CTAs
    const int tasks_per_tile = NPT + (p.Kp > 256 ? (KP <= 16 ? (NPT + 1) / 2 : NPT) : 0);
#pragma unroll
    for (int b = 0; b < E4_HBUFS; ++b) {
      mbar_init(&sh->h_full[b], 2 * tasks_per_tile);
      mbar_init(&sh->h_empty[b], 1);
    }
#pragma unroll
    for (int b = 0; b < E4_SETS; ++b) {
      mbar_init(&sh->d_full[b], 1);
      mbar_init(&sh->d_empty[b], 2 * E4_EPI_WARPS);
    }
    fence_mbar_init();
  }
  if (warp == E4_MMA_WARP) tmem_alloc_pair<512>(&sh->tmem_base);
  // ---- this CTA's half of W2 -> tensor memory (once).  EVERY warp takes part: warp w may write
  // TMEM lanes 32 (w & 3) .. +31 (lane = output channel), and the K chunks of a lane quadrant
  // are dealt over the warps that share it, so a warp moves one or two 64-byte chunks per lane
  // and the first one is in flight while shared memory is being zeroed.  (Eight warps walking
  // eleven chunks each, one dependent DRAM round trip after the other, cost 7 us of set-up.)
  const int wq = warp & 3, wi = warp >> 2;
  const int nwq = (E4_THREADS / 32 - wq + 3) >> 2;        // warps of this lane quadrant
  const int wch = (int)rank * 128 + wq * 32 + lane;        // output channel of this lane
  const uint4* const wrow = reinterpret_cast<const uint4*>(
      reinterpret_cast<const __half*>(p.w2) + (int64_t)wch * p.Kp);
  uint32_t wreg[16];
  auto load_w2_chunk = [&](int cc) {                       // 4 x 16 bytes = 32 fp16 = 16 columns
#pragma unroll
    for (int t = 0; t < 4; ++t) {
      const uint4 w = (wch < p.bnp && cc < p.nch) ? __ldg(wrow + cc * 4 + t) : make_uint4(0, 0, 0, 0);
      wreg[4 * t] = w.x; wreg[4 * t + 1] = w.y; wreg[4 * t + 2] = w.z; wreg[4 * t + 3] = w.w;
    }
  };
  load_w2_chunk(wi);
  // rows that carry no slot (and the K padding of the last chunk) are never written: zero once
  for (int i = threadIdx.x * 16; i < E4_HBUFS * htile_bytes; i += E4_THREADS * 16)
    *reinterpret_cast<uint4*>(h_smem + i) = make_uint4(0, 0, 0, 0);
  fence_proxy_async_smem();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = sh->tmem_base;
  {
    const uint32_t taddr = tmem_base + ((uint32_t)(wq * 32) << 16);
    for (int cc = wi; cc < p.nch; cc += nwq) {
      if (cc != wi) load_w2_chunk(cc);
      tmem_st_32x32_x16(taddr + cc * 16, wreg);
    }
    tmem_st_wait();
  }
  tc_fence_before();
  cluster_sync_all();   // barriers initialised, W2 resident and h tiles zeroed in BOTH CTAs
  tc_fence_after();

  // optional wait-time accounting (profiling runs): dbg[1] builders on h_empty, dbg[3]/[4] MMA on
  // h_full / d_empty, dbg[5] epilogue on d_full, dbg[6] total cycles, dbg[7] steps
#ifdef GN_DEBUG_BUILD
  const bool timed = p.dbg != nullptr;
#else
  constexpr bool timed = false;
#endif
  long long w0 = 0, w1 = 0;
  const long long t_start = clock64();
  // timeline of cluster 0 (profiling builds): stamp k of this warp's step i, cycles since the
  // set-up barrier (the two CTAs leave it within a few cycles of each other)
  unsigned long long* const trace =
      (timed && blockIdx.x < 2 && lane == 0)
          ? reinterpret_cast<unsigned long long*>(p.dbg) + 8 + ((int)rank * 32 + warp) * (64 * 4)
          : nullptr;
  auto stamp = [&](int64_t i, int k) {
    if (trace != nullptr && i < 64) trace[i * 4 + k] = (unsigned long long)(clock64() - t_start);
  };

  if (warp == E4_MMA_WARP) {
    // =================================================================== MMA issuer
    if (rank == 0) {
      const uint32_t idesc = make_idesc_f16(256, 2 * E4_ROWS);
      const uint32_t h_addr = smem_u32(h_smem);
      for (int64_t it = 0; it < my_steps; ++it) {
        const uint32_t hb = (uint32_t)(it % E4_HBUFS), db = (uint32_t)(it & 1);
        long long t0 = timed ? clock64() : 0;
        stamp(it, 0);
        mbar_wait(&sh->h_full[hb], (uint32_t)((it / E4_HBUFS) & 1));
        if (timed) { const long long t1 = clock64(); w0 += t1 - t0; t0 = t1; }
        stamp(it, 1);
        mbar_wait(&sh->d_empty[db], (uint32_t)((it >> 1) & 1) ^ 1);
        if (timed) w1 += clock64() - t0;
        stamp(it, 2);
        tc_fence_after();
        if (elect_one()) {
          // K step s: B = 32 bytes at offset (s&3)*32 of the 128-byte rows of chunk s>>2
          // (descriptor units of 16 B: +2 per step, +512 per chunk); A = TMEM columns 8s..8s+7
          const uint32_t d = tmem_base + (db ? E4_TMEM_D1 : E4_TMEM_D0);
          uint64_t bd = make_kmajor_desc<128>(h_addr + hb * htile_bytes);
          uint32_t a = tmem_base;
#pragma unroll 1
          for (int s0 = 0; s0 < ksteps; s0 += 4) {
            mma_f16_ts_pair(d, a, bd, idesc, s0 > 0);
            mma_f16_ts_pair(d, a + 8, bd + 2, idesc, 1);
            if (s0 + 2 < ksteps) {   // K is a multiple of 32: steps come in pairs
              mma_f16_ts_pair(d, a + 16, bd + 4, idesc, 1);
              mma_f16_ts_pair(d, a + 24, bd + 6, idesc, 1);
            }
            bd += E4_CHUNK_BYTES >> 4;
            a += 32;
          }
          mma_commit_pair(&sh->d_full[db]);
          mma_commit_pair(&sh->h_empty[hb]);
        }
        __syncwarp();
        stamp(it, 3);
      }
    }
  } else if (warp >= E4_EPI_WARPS + E4_BLD_WARPS) {
    // warp 24: nothing to do in the main loop (see E4_MMA_WARP)
  } else if (warp >= E4_EPI_WARPS) {
    // ==================================================================== h builders
    // Tasks of a tile:
    //   A(n): node n, 16-byte pieces 0..31 of its rows (lane = piece);
    //   B(m): pieces 32.. (rows wider than 512 bytes: at most 12 more pieces) - of nodes 2m and
    //         2m+1 at once when the slot ids of a node fit 16 lanes (lane >> 4 = which node),
    //         else of node m alone.
    // Nothing in a task is conditional: an empty slot (or a node past the end) reads a row of
    // -inf instead of Q_j, so ReLU(P + Q) = 0 falls out of the same HFMA2.RELU, and lanes past
    // the row's last piece redo piece 0 and skip the store.  (Predicated loads made ptxas
    // consume each batch of four before issuing the next: three exposed latencies per task.)
    constexpr bool PAIRB = KP <= 16;
    const int bw = warp - E4_EPI_WARPS;                 // 0..15
    const int act = GEN ? p.act : (int)ACT_RELU;
    const int nb_tasks = ppr > 32 ? (PAIRB ? (NPT + 1) / 2 : NPT) : 0;
    const int ntasks = NPT + nb_tasks;                  // = tasks_per_tile above
    // bases and pitches pinned in registers (one 64-bit multiply-add per row address)
    uint64_t qbase = reinterpret_cast<uint64_t>(p.qmat);
    uint64_t pbase = reinterpret_cast<uint64_t>(p.pmat);
    uint64_t negrow = reinterpret_cast<uint64_t>(p.neg_row);
    uint32_t qpitch = (uint32_t)p.ldq * 2, ppitch = (uint32_t)p.ldp * 2;
    asm volatile("" : "+l"(qbase), "+l"(pbase), "+l"(negrow), "+r"(qpitch), "+r"(ppitch));
    const uint32_t h_smem_addr = smem_u32(h_smem);
    const uint32_t hfull_leader0 = map_to_cta(smem_u32(&sh->h_full[0]), 0);

    // ONE queue of tasks over all steps, dealt round robin over the 16 builder warps: task g is
    // task g % ntasks of step g / ntasks.  (Two sets of 8 warps taking alternate tiles left
    // warps with two tasks of a tile next to warps with one: the tile took two task times.)
    // (it, t) of this warp's current task; advancing by 16 tasks
    // A warp may run at most E4_HBUFS steps ahead of the tile it has just filled: the parity wait
    // on h_empty tells "released" from "not yet" only one use of the buffer apart.  With few
    // tasks per tile that caps the number of warps in the queue.
    const int nbw = min(E4_BLD_WARPS, E4_HBUFS * ntasks);
    int64_t it = bw < nbw ? bw / ntasks : my_steps;
    int t = bw % ntasks;
    const int adv_it = nbw / ntasks, adv_t = nbw % ntasks;

    // per-lane ids of a task (slot id of the lane's node, the node itself), fetched one task
    // ahead of their use
    int32_t j_nx = -1, n_nx = -1;
    auto fetch_ids = [&](int64_t it_, int t_) {
      j_nx = -1;
      n_nx = -1;
      if (it_ >= my_steps) return;
      const int64_t tile = 2 * (step0 + it_) + rank;
      int nn, sl;   // node within the tile, slot this lane fetches
      if (t_ < NPT) { nn = t_; sl = lane; }
      else if (PAIRB) { nn = 2 * (t_ - NPT) + (lane >> 4); sl = lane & 15; }
      else { nn = t_ - NPT; sl = lane; }
      const int64_t grp = tile * NPT + nn;
      if (nn < NPT && grp < num_groups) {
        // (32-bit offset: a table has fewer than 2^31 entries)
        if (sl < KP) j_nx = e4_ldg32(p.ell + ((int32_t)grp * (int32_t)p.ell_ld + sl));
        n_nx = p.group_node ? e4_ldg32(p.group_node + grp) : (int32_t)grp;
      }
    };
    fetch_ids(it, t);
    while (it < my_steps) {
      const uint32_t hb = (uint32_t)(it % E4_HBUFS);
      const uint32_t hpar = (uint32_t)((it / E4_HBUFS) & 1) ^ 1;
      const uint32_t h_base = h_smem_addr + hb * htile_bytes;
      const int32_t jv = j_nx, node = n_nx;
      stamp(it, 0);
      // the next task of this warp
      int64_t it_n = it + adv_it;
      int t_n = t + adv_t;
      if (t_n >= ntasks) { t_n -= ntasks; ++it_n; }
      fetch_ids(it_n, t_n);
      {
        const bool isb = t >= NPT;
        const int nn = !isb ? t : PAIRB ? 2 * (t - NPT) + (lane >> 4) : t - NPT;
        const int piece = !isb ? lane : 32 + (PAIRB ? (lane & 15) : lane);
        const int src0 = (isb && PAIRB) ? (lane & 16) : 0;   // lane holding slot 0 of my node
        const bool pok = piece < ppr && nn < NPT;
        const uint32_t pofs = pok ? (uint32_t)piece * 16u : 0u;
        const uint4 P = e4_ldg128(reinterpret_cast<const uint8_t*>(
            pbase + (uint64_t)(uint32_t)max(node, 0) * ppitch) + pofs);
        // this lane's piece of Q row 0 / of the -inf row (a dead node reads -inf in every slot)
        const uint64_t qlane = qbase + pofs;
        const uint64_t nlane = negrow + pofs;
        const int32_t jdead = node >= 0 ? 0 : -1;
        // (ids outside [0, N) - an empty slot's -1, or garbage behind inputs the flags word
        // rejects - gather the -inf row)
        const uint32_t nmax = (uint32_t)p.N;
        // byte offset of the node's first row in its K chunk; the 128B swizzle XORs the 16-byte
        // piece index with row & 7 = bits 7..9 of the row offset
        const uint32_t dst0 = h_base + (uint32_t)(piece >> 3) * E4_CHUNK_BYTES;
        const uint32_t roff0 = (uint32_t)(nn * KP) * 128u;
        const uint32_t pbits = (uint32_t)(piece & 7) << 4;
        bool first = true;
#pragma unroll
        for (int s0 = 0; s0 < KP; s0 += E4_SG) {
          uint4 Q[E4_SG];
#pragma unroll
          for (int s = 0; s < E4_SG; ++s) {
            if (s0 + s < KP) {
              const int32_t j = __shfl_sync(0xffffffffu, jv, src0 + s0 + s) | jdead;
              const uint64_t qa = qlane + (uint64_t)(uint32_t)j * qpitch;
              Q[s] = e4_ldg128(reinterpret_cast<const void*>((uint32_t)j < nmax ? qa : nlane));
            }
          }
          if (first) {  // the tensor core has finished with this buffer's previous tile
            const long long t0 = timed ? clock64() : 0;
            stamp(it, 1);
            mbar_wait(&sh->h_empty[hb], hpar);
            if (timed) w1 += clock64() - t0;
            stamp(it, 2);
            first = false;
          }
#pragma unroll
          for (int s = 0; s < E4_SG; ++s) {
            if (s0 + s < KP) {
              const uint32_t roff = roff0 + (uint32_t)(s0 + s) * 128u;
              const uint32_t dst = dst0 + roff + (pbits ^ ((roff >> 3) & 0x70u));
              uint4 hv;
              hv.x = e4_act_add(P.x, Q[s].x, act);
              hv.y = e4_act_add(P.y, Q[s].y, act);
              hv.z = e4_act_add(P.z, Q[s].z, act);
              hv.w = e4_act_add(P.w, Q[s].w, act);
              if (GEN && act != ACT_RELU) {   // act(-inf) is not 0 for these: zero empty slots
                const int32_t j = __shfl_sync(0xffffffffu, jv, src0 + s0 + s) | jdead;
                if ((uint32_t)j >= nmax) hv = make_uint4(0, 0, 0, 0);
              }
              if (pok)
                asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(dst), "r"(hv.x),
                             "r"(hv.y), "r"(hv.z), "r"(hv.w)
                             : "memory");
            }
          }
        }
      }
      fence_proxy_async_smem();
      __syncwarp();
      if (lane == 0) mbar_arrive_cluster(hfull_leader0 + hb * 8);
      stamp(it, 3);
      it = it_n;
      t = t_n;
    }
  } else {
    // ====================================================================== epilogue
    // warp (q, half): TMEM lanes 32q.. = 32 of this CTA's 128 channels, columns 64*half.. = the
    // tile built by CTA `half`; KP consecutive columns per node -> plain in-register reductions;
    // the 32 lanes store 32 consecutive channels of a node
    const int q = warp & 3, half = warp >> 2;
    const int ch = (int)rank * 128 + q * 32 + lane;
    const bool ch_ok = ch < p.b;
    int crank = -1;  // position of this channel among the kNN coordinate columns
    if (p.coords != nullptr && ch_ok && ((p.coord_mask[ch >> 5] >> (ch & 31)) & 1u)) {
      crank = __popc(p.coord_mask[ch >> 5] & ((1u << (ch & 31)) - 1u));
      for (int w = 0; w < (ch >> 5); ++w) crank += __popc(p.coord_mask[w]);
    }
    const bool warp_has_coords = __any_sync(0xffffffffu, crank >= 0);
    const bool coords_packed = __all_sync(0xffffffffu, crank == (lane < p.ncoord ? lane : -1)) &&
                               NPT * p.ncoord <= 32;
    const int csel = coords_packed ? lane % p.ncoord : 0;   // (runtime divisions: once, here)
    const int nsel = coords_packed ? lane / p.ncoord : 0;
    const uint32_t lane_addr = tmem_base + ((uint32_t)(q * 32) << 16) + 64 * half;
    const uint32_t dempty_leader[2] = {map_to_cta(smem_u32(&sh->d_empty[0]), 0),
                                       map_to_cta(smem_u32(&sh->d_empty[1]), 0)};
    uint16_t* const out_ch = reinterpret_cast<uint16_t*>(p.out) + ch;
    const int aggr = GEN ? p.aggr : (int)E4_AGGR_ADD;
    const int act = GEN ? p.act : (int)ACT_RELU;
    for (int64_t it = 0; it < my_steps; ++it) {
      const int64_t tile = 2 * (step0 + it) + half;
      const int64_t grp0 = tile * NPT;
      const int nvalid = (int)max((int64_t)0, min((int64_t)NPT, num_groups - grp0));
      const uint32_t db = (uint32_t)(it & 1);
      stamp(it, 0);
      {
        const long long t0 = timed ? clock64() : 0;
        // (plain try_wait: the instruction itself suspends the warp until the phase flips; a
        // __nanosleep back-off wakes up to a microsecond late, and the late release of the
        // accumulator stalls the tensor core)
        mbar_wait(&sh->d_full[db], (uint32_t)((it >> 1) & 1));
        if (timed) w0 += clock64() - t0;
      }
      tc_fence_after();
      stamp(it, 1);
      const long long t_ld = timed ? clock64() : 0;
      const uint32_t d_addr = lane_addr + (db ? E4_TMEM_D1 : E4_TMEM_D0);
      float red[NPT];
      float2 red2[NPT], abs2[NPT];
#pragma unroll
      for (int n = 0; n < NPT; ++n) red2[n] = abs2[n] = make_float2(0.f, 0.f);
      const float init = aggr == E4_AGGR_MAX ? -INFINITY : aggr == E4_AGGR_MIN ? INFINITY : 0.f;
#pragma unroll
      for (int n = 0; n < NPT; ++n) red[n] = init;
      // slots in use per node (max / min / mean must not see the zeros of empty slots)
      int cnt[NPT];
      if (aggr != E4_AGGR_ADD) {
#pragma unroll
        for (int n = 0; n < NPT; ++n)
          cnt[n] = (n < nvalid && p.deg != nullptr) ? min(__ldg(p.deg + grp0 + n), KP) : KP;
      }
#pragma unroll
      for (int hh = 0; hh < 2; ++hh) {
        float v[32];
        tmem_ld_32x32(d_addr + 32 * hh, v);
        if (!GEN) {
          // ReLU + add without a max: sum relu(v) = (sum v + sum |v|) / 2, and |v| is an operand
          // modifier of FADD2 - two packed adds per (even, odd) column pair, all on the FMA
          // pipe (FMNMX runs on the ALU pipe the builders' integer work is already queueing
          // for).  A node's even columns go to .x and its odd columns to .y, each in column
          // order, and the halves are added at the end: the result does not depend on where
          // the node sits in its tile (a node starting on an odd column swaps the roles of x
          // and y, and x + y = y + x), so a batch gives the same bits however it is split.
#pragma unroll
          for (int c = 0; c < 32; c += 2) {
            const int col = 32 * hh + c;
            if (col + 1 < USED && col / KP == (col + 1) / KP) {
              const float2 vv = make_float2(v[c], v[c + 1]);
              red2[col / KP] = __fadd2_rn(red2[col / KP], vv);
              abs2[col / KP] = __fadd2_rn(abs2[col / KP], make_float2(fabsf(vv.x), fabsf(vv.y)));
            } else {
              if (col < USED) { red2[col / KP].x += v[c]; abs2[col / KP].x += fabsf(v[c]); }
              if (col + 1 < USED) {
                red2[(col + 1) / KP].y += v[c + 1];
                abs2[(col + 1) / KP].y += fabsf(v[c + 1]);
              }
            }
          }
        } else {
#pragma unroll
          for (int c = 0; c < 32; ++c) {
            const int col = 32 * hh + c;
            if (col < USED) {
              const float a = e4_act(v[c], act);
              if (aggr <= E4_AGGR_MEAN) {
                red[col / KP] += a;
              } else if (col % KP < cnt[col / KP]) {
                red[col / KP] = aggr == E4_AGGR_MAX ? fmaxf(red[col / KP], a)
                                                    : fminf(red[col / KP], a);
              }
            }
          }
        }
        if (MASK && !ACC && tile < num_tiles) {
          // training: the ReLU pattern of the 32 channels of this warp, one word per tile row
          // (the backward pass masks dOut with it instead of recomputing Z = H W2^T)
          uint32_t mword = 0;
#pragma unroll
          for (int c = 0; c < 32; ++c) {
            const uint32_t w = __ballot_sync(0xffffffffu, v[c] > 0.f);
            if (lane == c) mword = w;
          }
          p.relu_mask[(int64_t)((int)rank * 4 + q) * p.mask_ld + tile * E4_ROWS + 32 * hh + lane] =
              mword;
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive_cluster(dempty_leader[db]);  // values are in registers now
      const long long t_st = timed ? clock64() : 0;
      if (timed) w1 += t_st - t_ld;
      stamp(it, 2);
      if (!GEN) {
#pragma unroll
        for (int n = 0; n < NPT; ++n)
          red[n] = 0.5f * ((red2[n].x + red2[n].y) + (abs2[n].x + abs2[n].y));
      }
      if (aggr != E4_AGGR_ADD) {
#pragma unroll
        for (int n = 0; n < NPT; ++n) {
          if (aggr == E4_AGGR_MEAN) red[n] = cnt[n] > 0 ? red[n] / (float)cnt[n] : 0.f;
          else if (cnt[n] <= 0) red[n] = 0.f;   // no in-edges: 0, as PyG
        }
      }
      if (!ACC) {
        // pass 1: group g is node g -> consecutive rows of `out`
        if (ch_ok) {
          uint16_t* dst = out_ch + grp0 * p.ldo;
          if (nvalid == NPT) {
#pragma unroll
            for (int n = 0; n < NPT; ++n) { *dst = pack1_16(red[n], F16); dst += p.ldo; }
          } else {
#pragma unroll
            for (int n = 0; n < NPT; ++n) {
              if (n < nvalid) dst[(int64_t)n * p.ldo] = pack1_16(red[n], F16);
            }
          }
        }
        if (warp_has_coords && coords_packed) {
          // the coordinate columns are channels 0..ncoord-1 of this warp: the tile's
          // [nodes, ncoord] block of `coords` is contiguous - lane L writes element L
          float val = 0.f;
#pragma unroll
          for (int n = 0; n < NPT; ++n) {
            const float t = __shfl_sync(0xffffffffu, red[n], csel);
            if (nsel == n) val = t;
          }
          if (lane < nvalid * p.ncoord) p.coords[grp0 * p.ncoord + lane] = val;
        } else if (warp_has_coords) {   // general column sets
          if (crank >= 0) {
#pragma unroll
            for (int n = 0; n < NPT; ++n)
              if (n < nvalid) p.coords[(grp0 + n) * p.ncoord + crank] = red[n];
          }
        }
      } else if (ch_ok) {
#pragma unroll
        for (int n = 0; n < NPT; ++n) {
          if (n < nvalid) {
            const int64_t node = __ldg(p.group_node + grp0 + n);
            red_add1_16(out_ch + node * p.ldo, red[n], F16);
            if (crank >= 0) atomicAdd(p.coords + node * p.ncoord + crank, red[n]);
          }
        }
      }
      stamp(it, 3);
    }
  }
  if (timed && lane == 0) {
    unsigned long long* d = reinterpret_cast<unsigned long long*>(p.dbg);
    if (warp == E4_MMA_WARP) {
      if (rank == 0) {
        atomicAdd(d + 3, (unsigned long long)w0);
        atomicAdd(d + 4, (unsigned long long)w1);
        atomicAdd(d + 6, (unsigned long long)(clock64() - t_start));
        atomicAdd(d + 7, (unsigned long long)my_steps);
        atomicMax(d + 1, (unsigned long long)(clock64() - t_start));   // slowest cluster's loop
      }
    } else if (warp >= E4_EPI_WARPS) {
      atomicAdd(d + 2, (unsigned long long)w1);
    } else {
      atomicAdd(d + 5, (unsigned long long)w0);
      atomicAdd(d + 0, (unsigned long long)w1);   // TMEM loads + reduction, until the release
    }
  }
  tc_fence_before();
  cluster_sync_all();   // every MMA of the pair has been consumed by both epilogues
#ifdef GN_DEBUG_BUILD
  if (timed && threadIdx.x == 0 && rank == 0) {   // set-up and whole-kernel cycles, max over clusters
    unsigned long long* d = reinterpret_cast<unsigned long long*>(p.dbg);
    atomicMax(d + 8 + (31 * 256) + 0, (unsigned long long)(t_start - t_entry));
    atomicMax(d + 8 + (31 * 256) + 1, (unsigned long long)(clock64() - t_entry));
  }
#endif
  if (warp == E4_MMA_WARP) tmem_dealloc_pair<512>(tmem_base);
}

inline size_t edge4_smem_bytes(int Kp) {
  const int nchunk = (Kp + 63) / 64;
  return (size_t)E4_HBUFS * nchunk * E4_CHUNK_BYTES + sizeof(Edge4Shared) + 1024;
}

}  // namespace tc
}  // namespace gn
