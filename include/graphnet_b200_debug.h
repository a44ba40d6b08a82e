/* Diagnostic exports of the PROFILING build of libgraphnet_b200 (compiled with
 * -DGN_DEBUG_BUILD by __graft_entry__.build_debug() into libgraphnet_b200_dbg.so).  The product
 * library does not contain them, nor the counters, timelines and ablation switches
 * (GN_EDGE_DEBUG, GN_TC_DEBUG, GN_TC_HACK*) behind them.  No counterpart in the reference. */
#ifndef GRAPHNET_B200_DEBUG_H_
#define GRAPHNET_B200_DEBUG_H_

#ifdef __cplusplus
extern "C" {
#endif

/* Fused EdgeConv kernel, GN_EDGE_DEBUG=1: 8 counters of cycles spent waiting, summed over
 * warps and launches since the last call ([0] epilogue TMEM-read phase, [2] builders on a free
 * h tile, [3] / [4] MMA issuer on h tiles / on a free accumulator, [5] epilogue on a full
 * accumulator, [6] total loop cycles of the issuer, [7] steps).  Copies them out and clears
 * them. */
int gn_debug_edge_counters(unsigned long long* host8);

/* Fused EdgeConv kernel, GN_EDGE_DEBUG=1: timeline of cluster 0 of the last launch,
 * [CTA rank 0..1][warp 0..31][step 0..63][4 stamps], cycles since the kernel's set-up barrier. */
int gn_debug_edge_trace(unsigned long long* host, int words);

/* tcgen05 GEMM, GN_TC_DEBUG=1: timeline of CTA 0 of the last launch,
 * [warp 0..9][item 0..15][4 stamps] (clock64). */
int gn_debug_linear_trace(unsigned long long* host, int words);

/* Same launch, per ring stage g < 64 of CTA 0: [0][g] TMA issued, [1][g] the MMA thread saw the
 * stage full, [2][g] its commit (slot release) issued. */
int gn_debug_linear_stages(unsigned long long* host, int words);

#ifdef __cplusplus
}
#endif
#endif
