// Fused EdgeConv message + aggregation at fp32 accuracy on the tensor cores (the reference's own
// precision, layers.py:67 / dynedge.py:193-204 in float32).
//
// Same CTA-pair structure as tc_edge_pair.cuh (read that header first); what differs:
//   * P and Q arrive in fp32; the builders form h = relu(P_i + Q_j) in fp32 and write it as TWO
//     fp16 tiles, h_hi = fp16(h) and h_lo = fp16(h - h_hi): 22 mantissa bits between them;
//   * W2 sits in tensor memory twice, W_hi = fp16(W2) and W_lo = fp16(W2 - W_hi);
//   * every K step issues three MMAs into the same fp32 accumulator,
//         D += W_hi h_hi + W_hi h_lo + W_lo h_hi          (the W_lo h_lo term is below 2^-22),
//   * outputs are fp32 (stored, or added atomically in the overflow pass).
// TMEM columns: [0, Kp/2) W_hi, [176, 176 + Kp/2) W_lo, [352, 480) the accumulator - one buffer
// only: the tensor core waits while the epilogue reads it (~15 % of a step that is three times as
// long as the 16-bit kernel's).  Shared memory: two (h_hi, h_lo) tile pairs in flight.
#pragma once

namespace gn {
namespace tc {

constexpr int X3_HB = 2;             // (h_hi, h_lo) tile pairs in flight
constexpr int X3_TMEM_WLO = 176;
constexpr int X3_TMEM_D = 352;

struct __align__(16) EdgeX3Shared {
  uint64_t h_full[X3_HB];    // used in the leader CTA only
  uint64_t h_empty[X3_HB];
  uint64_t d_full;
  uint64_t d_empty;          // used in the leader CTA only
  uint32_t tmem_base;
  uint32_t pad_;
};

__device__ __forceinline__ uint32_t x3_pack_sat(float a, float b) {   // fp16x2, lo <- a, hi <- b
  uint32_t r;
  asm("cvt.rn.satfinite.f16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(b), "f"(a));
  return r;
}

// ACC: second pass (group lists, atomic adds into out)
template <int KP, bool ACC>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(E4_THREADS, 1)
    edge_mlp_aggregate_x3_kernel(const EdgeParams p) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) &
                                             ~uintptr_t(1023));
  constexpr int NPT = E4_ROWS / KP;    // nodes (groups) per 64-row tile
  constexpr int USED = NPT * KP;
  const int ksteps = p.Kp / 16;                           // K steps (3 MMAs each) per step
  const int ppr = p.Kp / 4;                               // 16-byte fp32 pieces per row
  const int nblk = (ppr + 31) >> 5;                       // 32-piece blocks = builder tasks per node
  const int nchunk = (p.Kp + 63) / 64;                    // 128-byte K chunks of an h tile
  const int htile_bytes = nchunk * E4_CHUNK_BYTES;
  uint8_t* h_smem = smem;                                 // [X3_HB][hi, lo] tiles
  EdgeX3Shared* sh = reinterpret_cast<EdgeX3Shared*>(h_smem + 2 * X3_HB * htile_bytes);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t rank = cluster_ctarank();                // 0 = leader (issues the MMAs)
  const int64_t num_groups =
      p.group_count ? min((int64_t)__ldg(p.group_count), p.group_cap) : p.N;
  const int64_t num_tiles = (num_groups + NPT - 1) / NPT;
  const int64_t num_steps = (num_tiles + 1) / 2;          // a step = two tiles, one per CTA
  const int64_t clusters = gridDim.x >> 1, cid = blockIdx.x >> 1;
  const int64_t steps_per = (num_steps + clusters - 1) / clusters;
  const int64_t step0 = cid * steps_per;
  if (step0 >= num_steps) return;      // uniform per cluster, before any barrier
  const int64_t my_steps = min(steps_per, num_steps - step0);
  const int ntasks = NPT * nblk;       // builder tasks per tile; every task arrives once per CTA

  if (threadIdx.x == 0) {
#pragma unroll
    for (int b = 0; b < X3_HB; ++b) {
      mbar_init(&sh->h_full[b], 2 * ntasks);
      mbar_init(&sh->h_empty[b], 1);
    }
    mbar_init(&sh->d_full, 1);
    mbar_init(&sh->d_empty, 2 * E4_EPI_WARPS);
    fence_mbar_init();
  }
  if (warp == E4_MMA_WARP) tmem_alloc_pair<512>(&sh->tmem_base);
  // ---- this CTA's halves of W_hi and W_lo -> tensor memory (once), every warp taking part:
  // chunk cc < nch is 32 columns of W_hi, chunk nch + cc the same columns of W_lo
  const int wq = warp & 3, wi = warp >> 2;
  const int nwq = (E4_THREADS / 32 - wq + 3) >> 2;        // warps of this lane quadrant
  const int wch = (int)rank * 128 + wq * 32 + lane;        // output channel of this lane
  const uint4* const wrow_hi = reinterpret_cast<const uint4*>(
      reinterpret_cast<const __half*>(p.w2) + (int64_t)wch * p.Kp);
  const uint4* const wrow_lo = reinterpret_cast<const uint4*>(
      reinterpret_cast<const __half*>(p.w2lo) + (int64_t)wch * p.Kp);
  uint32_t wreg[16];
  auto load_w2_chunk = [&](int cc) {                       // 4 x 16 bytes = 32 fp16 = 16 columns
    const uint4* src = cc < p.nch ? wrow_hi + cc * 4 : wrow_lo + (cc - p.nch) * 4;
#pragma unroll
    for (int t = 0; t < 4; ++t) {
      const uint4 w = (wch < p.bnp && cc < 2 * p.nch) ? __ldg(src + t) : make_uint4(0, 0, 0, 0);
      wreg[4 * t] = w.x; wreg[4 * t + 1] = w.y; wreg[4 * t + 2] = w.z; wreg[4 * t + 3] = w.w;
    }
  };
  load_w2_chunk(wi);
  // rows that carry no slot (and the K padding of the last chunk) are never written: zero once
  for (int i = threadIdx.x * 16; i < 2 * X3_HB * htile_bytes; i += E4_THREADS * 16)
    *reinterpret_cast<uint4*>(h_smem + i) = make_uint4(0, 0, 0, 0);
  fence_proxy_async_smem();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = sh->tmem_base;
  {
    const uint32_t taddr = tmem_base + ((uint32_t)(wq * 32) << 16);
    for (int cc = wi; cc < 2 * p.nch; cc += nwq) {
      if (cc != wi) load_w2_chunk(cc);
      const uint32_t col = cc < p.nch ? (uint32_t)cc * 16u
                                      : (uint32_t)X3_TMEM_WLO + (uint32_t)(cc - p.nch) * 16u;
      tmem_st_32x32_x16(taddr + col, wreg);
    }
    tmem_st_wait();
  }
  tc_fence_before();
  cluster_sync_all();   // barriers initialised, W2 resident and h tiles zeroed in BOTH CTAs
  tc_fence_after();

  if (warp == E4_MMA_WARP) {
    // =================================================================== MMA issuer
    if (rank == 0) {
      const uint32_t idesc = make_idesc_f16(256, 2 * E4_ROWS);
      const uint32_t h_addr = smem_u32(h_smem);
      for (int64_t it = 0; it < my_steps; ++it) {
        const uint32_t hb = (uint32_t)(it & 1);
        mbar_wait(&sh->h_full[hb], (uint32_t)((it >> 1) & 1));
        mbar_wait(&sh->d_empty, (uint32_t)(it & 1) ^ 1);
        tc_fence_after();
        if (elect_one()) {
          const uint32_t d = tmem_base + X3_TMEM_D;
          uint64_t bh = make_kmajor_desc<128>(h_addr + (2 * hb) * htile_bytes);
          uint64_t bl = make_kmajor_desc<128>(h_addr + (2 * hb + 1) * htile_bytes);
          uint32_t a = tmem_base;
#pragma unroll 1
          for (int s0 = 0; s0 < ksteps; s0 += 4) {
            // K step u of this chunk: B = 32 bytes at offset 32 u of the 128-byte rows
            // (descriptor units of 16 B: +2 per step), A = TMEM columns 8 u .. 8 u + 7
#pragma unroll
            for (int u = 0; u < 4; ++u) {
              if (u < 2 || s0 + 2 < ksteps) {   // K is a multiple of 32: steps come in pairs
                mma_f16_ts_pair(d, a + 8 * u, bh + 2 * u, idesc, (s0 | u) > 0);
                mma_f16_ts_pair(d, a + 8 * u, bl + 2 * u, idesc, 1);
                mma_f16_ts_pair(d, a + X3_TMEM_WLO + 8 * u, bh + 2 * u, idesc, 1);
              }
            }
            bh += E4_CHUNK_BYTES >> 4;
            bl += E4_CHUNK_BYTES >> 4;
            a += 32;
          }
          mma_commit_pair(&sh->d_full);
          mma_commit_pair(&sh->h_empty[hb]);
        }
        __syncwarp();
      }
    }
  } else if (warp >= E4_EPI_WARPS + E4_BLD_WARPS) {
    // warp 24: W2 upload only
  } else if (warp >= E4_EPI_WARPS) {
    // ==================================================================== h builders
    // Task (node n of the tile, block pb): the node's 16-byte fp32 pieces 32 pb .. 32 pb + 31
    // (lane = piece) of all its KP slots.  An empty slot (or a node past the end) gathers a row
    // of -inf instead of Q_j: relu(P - inf) = 0 falls out of the same instructions.
    const int bw = warp - E4_EPI_WARPS;                 // 0..15
    uint64_t qbase = reinterpret_cast<uint64_t>(p.qmat);
    uint64_t pbase = reinterpret_cast<uint64_t>(p.pmat);
    uint64_t negrow = reinterpret_cast<uint64_t>(p.neg_row32);
    uint32_t qpitch = (uint32_t)p.ldq * 4, ppitch = (uint32_t)p.ldp * 4;
    asm volatile("" : "+l"(qbase), "+l"(pbase), "+l"(negrow), "+r"(qpitch), "+r"(ppitch));
    const uint32_t h_smem_addr = smem_u32(h_smem);
    const uint32_t hfull_leader0 = map_to_cta(smem_u32(&sh->h_full[0]), 0);
    // one queue of tasks over all steps, dealt round robin over the builder warps; a warp may run
    // at most X3_HB steps ahead of the tile it has just filled (parity wait on h_empty)
    const int nbw = min(E4_BLD_WARPS, X3_HB * ntasks);
    int64_t it = bw < nbw ? bw / ntasks : my_steps;
    int t = bw % ntasks;
    const int adv_it = nbw / ntasks, adv_t = nbw % ntasks;

    int32_t j_nx = -1, n_nx = -1;
    auto fetch_ids = [&](int64_t it_, int t_) {
      j_nx = -1;
      n_nx = -1;
      if (it_ >= my_steps) return;
      const int64_t tile = 2 * (step0 + it_) + rank;
      const int64_t grp = tile * NPT + t_ / nblk;
      if (grp < num_groups) {
        if (lane < KP) j_nx = e4_ldg32(p.ell + ((int32_t)grp * (int32_t)p.ell_ld + lane));
        n_nx = p.group_node ? e4_ldg32(p.group_node + grp) : (int32_t)grp;
      }
    };
    fetch_ids(it, t);
    while (it < my_steps) {
      const uint32_t hb = (uint32_t)(it & 1);
      const uint32_t hpar = (uint32_t)((it >> 1) & 1) ^ 1;
      const uint32_t hi_base = h_smem_addr + (2 * hb) * htile_bytes;
      const uint32_t lo_base = hi_base + htile_bytes;
      const int32_t jv = j_nx, node = n_nx;
      int64_t it_n = it + adv_it;
      int t_n = t + adv_t;
      if (t_n >= ntasks) { t_n -= ntasks; ++it_n; }
      fetch_ids(it_n, t_n);
      {
        const int nn = t / nblk;
        const int piece = (t - nn * nblk) * 32 + lane;       // 4 fp32 columns
        const bool pok = piece < ppr;
        const uint32_t pofs = pok ? (uint32_t)piece * 16u : 0u;
        const uint4 Pu = e4_ldg128(reinterpret_cast<const uint8_t*>(
            pbase + (uint64_t)(uint32_t)max(node, 0) * ppitch) + pofs);
        const uint64_t qlane = qbase + pofs;
        const uint64_t nlane = negrow + pofs;
        const int32_t jdead = node >= 0 ? 0 : -1;
        // (ids outside [0, N) - an empty slot's -1, or garbage behind inputs the flags word
        // rejects - gather the -inf row)
        const uint32_t nmax = (uint32_t)p.N;
        // the four columns are half of a 16-byte fp16 piece of the tile row: chunk = 64 columns,
        // the 128B swizzle XORs the piece index with row & 7
        const uint32_t cofs = (uint32_t)(piece >> 4) * E4_CHUNK_BYTES + (uint32_t)(piece & 1) * 8u;
        const uint32_t pbits = (uint32_t)((piece >> 1) & 7) << 4;
        const uint32_t roff0 = (uint32_t)(nn * KP) * 128u;
        const float p0 = __uint_as_float(Pu.x), p1 = __uint_as_float(Pu.y);
        const float p2 = __uint_as_float(Pu.z), p3 = __uint_as_float(Pu.w);
        bool first = true;
#pragma unroll
        for (int s0 = 0; s0 < KP; s0 += E4_SG) {
          uint4 Q[E4_SG];
#pragma unroll
          for (int s = 0; s < E4_SG; ++s) {
            if (s0 + s < KP) {
              const int32_t j = __shfl_sync(0xffffffffu, jv, s0 + s) | jdead;
              const uint64_t qa = qlane + (uint64_t)(uint32_t)j * qpitch;
              Q[s] = e4_ldg128(reinterpret_cast<const void*>((uint32_t)j < nmax ? qa : nlane));
            }
          }
          if (first) {  // the tensor core has finished with this buffer pair's previous tiles
            mbar_wait(&sh->h_empty[hb], hpar);
            first = false;
          }
#pragma unroll
          for (int s = 0; s < E4_SG; ++s) {
            if (s0 + s < KP) {
              const uint32_t roff = roff0 + (uint32_t)(s0 + s) * 128u;
              const uint32_t off = cofs + roff + (pbits ^ ((roff >> 3) & 0x70u));
              const float h0 = fmaxf(p0 + __uint_as_float(Q[s].x), 0.f);
              const float h1 = fmaxf(p1 + __uint_as_float(Q[s].y), 0.f);
              const float h2 = fmaxf(p2 + __uint_as_float(Q[s].z), 0.f);
              const float h3 = fmaxf(p3 + __uint_as_float(Q[s].w), 0.f);
              const uint32_t a01 = x3_pack_sat(h0, h1), a23 = x3_pack_sat(h2, h3);
              const float2 f01 = __half22float2(*reinterpret_cast<const __half2*>(&a01));
              const float2 f23 = __half22float2(*reinterpret_cast<const __half2*>(&a23));
              const uint32_t l01 = x3_pack_sat(h0 - f01.x, h1 - f01.y);
              const uint32_t l23 = x3_pack_sat(h2 - f23.x, h3 - f23.y);
              if (pok) {
                asm volatile("st.shared.v2.b32 [%0], {%1, %2};" ::"r"(hi_base + off), "r"(a01),
                             "r"(a23)
                             : "memory");
                asm volatile("st.shared.v2.b32 [%0], {%1, %2};" ::"r"(lo_base + off), "r"(l01),
                             "r"(l23)
                             : "memory");
              }
            }
          }
        }
      }
      fence_proxy_async_smem();
      __syncwarp();
      if (lane == 0) mbar_arrive_cluster(hfull_leader0 + hb * 8);
      it = it_n;
      t = t_n;
    }
  } else {
    // ====================================================================== epilogue
    // warp (q, half): TMEM lanes 32q.. = 32 of this CTA's 128 channels, columns 64*half.. = the
    // tile built by CTA `half`; KP consecutive columns per node
    const int q = warp & 3, half = warp >> 2;
    const int ch = (int)rank * 128 + q * 32 + lane;
    const bool ch_ok = ch < p.b;
    int crank = -1;  // position of this channel among the kNN coordinate columns
    if (p.coords != nullptr && ch_ok && ((p.coord_mask[ch >> 5] >> (ch & 31)) & 1u)) {
      crank = __popc(p.coord_mask[ch >> 5] & ((1u << (ch & 31)) - 1u));
      for (int w = 0; w < (ch >> 5); ++w) crank += __popc(p.coord_mask[w]);
    }
    const uint32_t d_addr = tmem_base + ((uint32_t)(q * 32) << 16) + 64 * half + X3_TMEM_D;
    const uint32_t dempty_leader = map_to_cta(smem_u32(&sh->d_empty), 0);
    float* const out_ch = reinterpret_cast<float*>(p.out) + ch;
    for (int64_t it = 0; it < my_steps; ++it) {
      const int64_t tile = 2 * (step0 + it) + half;
      const int64_t grp0 = tile * NPT;
      const int nvalid = (int)max((int64_t)0, min((int64_t)NPT, num_groups - grp0));
      mbar_wait(&sh->d_full, (uint32_t)(it & 1));
      tc_fence_after();
      float2 red2[NPT], abs2[NPT];
#pragma unroll
      for (int n = 0; n < NPT; ++n) red2[n] = abs2[n] = make_float2(0.f, 0.f);
#pragma unroll
      for (int hh = 0; hh < 2; ++hh) {
        float v[32];
        tmem_ld_32x32(d_addr + 32 * hh, v);
        // sum relu(v) = (sum v + sum |v|) / 2; a node's even columns go to .x and its odd ones
        // to .y, each in column order (see tc_edge_pair.cuh: the result does not depend on
        // where the node sits in its tile)
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
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive_cluster(dempty_leader);  // values are in registers now
      float red[NPT];
#pragma unroll
      for (int n = 0; n < NPT; ++n)
        red[n] = 0.5f * ((red2[n].x + red2[n].y) + (abs2[n].x + abs2[n].y));
      if (!ACC) {
        if (ch_ok) {
#pragma unroll
          for (int n = 0; n < NPT; ++n)
            if (n < nvalid) out_ch[(grp0 + n) * p.ldo] = red[n];
        }
        if (crank >= 0) {
#pragma unroll
          for (int n = 0; n < NPT; ++n)
            if (n < nvalid) p.coords[(grp0 + n) * p.ncoord + crank] = red[n];
        }
      } else if (ch_ok) {
#pragma unroll
        for (int n = 0; n < NPT; ++n) {
          if (n < nvalid) {
            const int64_t node = __ldg(p.group_node + grp0 + n);
            atomicAdd(out_ch + node * p.ldo, red[n]);
            if (crank >= 0) atomicAdd(p.coords + node * p.ncoord + crank, red[n]);
          }
        }
      }
    }
  }
  tc_fence_before();
  cluster_sync_all();   // every MMA of the pair has been consumed by both epilogues
  if (warp == E4_MMA_WARP) tmem_dealloc_pair<512>(tmem_base);
}

}  // namespace tc
}  // namespace gn
