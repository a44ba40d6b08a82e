/*
 * graphnet_b200 - C ABI of the B200 (sm_100a) DynEdge hot path.
 *
 * The reference (graphnet-team/graphnet) has no FFI of its own: its DynEdge path
 * is Python that calls third-party wheels (torch_cluster, torch_scatter,
 * torch_geometric).  Each entry point below replaces one of those library calls
 * at the reference call site cited beside it (paths relative to the reference's
 * src/graphnet/).  Plain pointers and sizes only; all pointers are DEVICE
 * pointers unless marked "host"; `stream` is a cudaStream_t passed as void*.
 *
 * Every function returns GN_OK (0) or a negative GN_ERR_* code and never
 * aborts; gn_last_error() returns a message for the calling thread's last
 * failure.  Functions only enqueue work on `stream` (no synchronisation) unless
 * stated otherwise.  None of them allocates device memory: outputs and scratch
 * are provided by the caller (sizes via the *_bytes helpers).
 */
#ifndef GRAPHNET_B200_H_
#define GRAPHNET_B200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define GN_OK 0
#define GN_ERR_INVALID_ARGUMENT (-1)
#define GN_ERR_CUDA (-2)
#define GN_ERR_UNSUPPORTED (-3)

/* activation of an MLP layer (models/gnn/dynedge.py:161-168) */
#define GN_ACT_NONE 0
#define GN_ACT_RELU 1
#define GN_ACT_GELU 2
#define GN_ACT_LEAKY_RELU 3 /* torch.nn.LeakyReLU(), slope 0.01 (DynEdgeJINST) */

/* aggregation of EdgeConv messages (components/layers.py:33; dynedge.py:208 uses ADD) */
#define GN_AGGR_ADD 0
#define GN_AGGR_MEAN 1
#define GN_AGGR_MAX 2
#define GN_AGGR_MIN 3

/* global pooling schemes, GLOBAL_POOLINGS in models/gnn/dynedge.py:14-19 */
#define GN_POOL_MIN 0
#define GN_POOL_MAX 1
#define GN_POOL_SUM 2
#define GN_POOL_MEAN 3

/* element types of activations handed between kernels */
#define GN_F32 0
#define GN_BF16 1
#define GN_F16 2
/* gn_global_variables' x0 only: every value v as bf16 hi + lo (16 mantissa bits), rows laid out as
 * [hi | lo | hi], each x0_cols wide (ld >= 3 x0_cols): one bf16 GEMM against [W_hi | W_hi | W_lo]
 * then equals the fp32 product to ~2^-16 - the raw detector coordinates reach the first
 * EdgeConv's node GEMM on the tensor cores without being rounded to 8 bits */
#define GN_BF16X3 3

/* task heads, models/task/reconstruction.py */
#define GN_HEAD_IDENTITY 0
#define GN_HEAD_ENERGY 1             /* EnergyReconstruction :101-112            */
#define GN_HEAD_DIRECTION_KAPPA 2    /* DirectionReconstructionWithKappa :49-69  */
#define GN_HEAD_ZENITH 3             /* ZenithReconstruction :72-82              */
#define GN_HEAD_ZENITH_KAPPA 4       /* ZenithReconstructionWithKappa :85-98     */
#define GN_HEAD_AZIMUTH_KAPPA 5      /* AzimuthReconstructionWithKappa :11-27    */

/* flag bits written by the graph-conversion kernels */
#define GN_FLAG_UNSORTED_BATCH 1
#define GN_FLAG_UNSORTED_TARGETS 2
#define GN_FLAG_INDEX_OUT_OF_RANGE 4
#define GN_FLAG_DEGREE_OVERFLOW 8

/*
 * In-edges grouped by target node.  Two layouts:
 *   ELL (ell_stride > 0): sources of node i are col[i*ell_stride .. + deg[i])
 *   CSR (ell_stride == 0): sources of node i are col[rowptr[i] .. rowptr[i+1])
 * Within a node, sources keep the order of the reference's edge_index
 * (ascending (distance^2, index) for kNN graphs).
 */
typedef struct gn_graph {
  const int32_t* col;
  const int32_t* deg;    /* ELL: [N]   */
  const int32_t* rowptr; /* CSR: [N+1] */
  int32_t ell_stride;
  int64_t num_edges;     /* CSR: E (length of col); ignored for ELL */
} gn_graph;

const char* gn_last_error(void);
int gn_version(void);
/* Kernel launches issued by this library in this process so far (diagnostic). */
uint64_t gn_launch_count(void);

/* ---- graph construction ------------------------------------------------- */

/* Event offsets from the sorted PyG `batch` vector (what torch_cluster derives
 * internally for knn(x, x, k, batch, batch)).  ptr[e] = first node
 * of event e, ptr[B] = N.  Sets GN_FLAG_UNSORTED_BATCH / _OUT_OF_RANGE in *flags. */
int gn_batch_to_ptr(const int64_t* batch, int64_t N, int64_t B, int32_t* ptr,
                    int32_t* flags, void* stream);

/* torch_geometric.nn.knn_graph(x[:, cols], k, batch, loop) - replaces the call at
 * data_representation/graphs/edges/edges.py:78-82 and components/layers.py:70-74.
 * Per event, per query i: candidates ordered by (squared L2 distance in fp32
 * accumulated dimension by dimension without FMA, index); the first
 * min(n, k + (loop?0:1)) are kept, (i,i) is dropped unless loop.  Output is ELL:
 * nbr[i*stride + r] = r-th kept neighbour (global node id), padded with -1;
 * deg[i] in [0, k+1].  stride >= k+1.  `cols` is a HOST array of `dims` column
 * indices into the [N, ldx] matrix (dims <= 16). */
int gn_knn_graph(const float* x, int64_t ldx, const int32_t* cols, int dims,
                 const int32_t* ptr, int64_t N, int64_t B, int k, int loop,
                 int32_t* nbr, int32_t stride, int32_t* deg, void* stream);

/* Exclusive scan of the ELL degrees: rowptr[i] = sum(deg[0..i)), rowptr[N] = E.
 * `scratch` holds the per-block partial sums (gn_scan_scratch_bytes(N) bytes). */
size_t gn_scan_scratch_bytes(int64_t n);
int gn_ell_row_offsets(const int32_t* deg, int64_t N, int64_t* rowptr,
                       void* scratch, size_t scratch_bytes, void* stream);

/* Same result as gn_ell_row_offsets, for callers that may not allocate device memory while
 * running (an online filter process with a fixed GPU memory budget; a host whose device
 * memory all belongs to another framework's pool, where cudaMalloc fails): the partial sums
 * live in a 32 KiB array built into the library instead of in caller scratch, so the call
 * needs only its input and output.  No counterpart in the reference.
 * Limit: N <= 4194304 nodes per call (larger N is rejected with GN_ERR_INVALID_ARGUMENT).
 * The first call on each device looks up the array's device address (cudaGetSymbolAddress);
 * make that first call before starting any stream capture.  The array is the CURRENT device's:
 * the device that owns `stream` must be current when the call is made.  All calls on a device share the
 * array: calls whose kernels can overlap on the GPU (different streams) race - serialise them.
 * Nothing needs clearing between calls. */
int gn_ell_row_offsets_noalloc(const int32_t* deg, int64_t N, int64_t* rowptr, void* stream);

/* The int64 COO `edge_index` [2, E] the reference API returns (row 0 = source j,
 * row 1 = target i; ordered by target, then rank) from ELL + its row offsets. */
int gn_ell_to_coo(const int32_t* nbr, int32_t stride, const int32_t* deg,
                  const int64_t* rowptr, int64_t N, int64_t E, int64_t* edge_index,
                  void* stream);

/* A caller-provided PyG `edge_index` [2, E] (int64, targets non-decreasing, as
 * knn_graph / Batch.from_data_list produce) -> CSR by target.  Sets
 * GN_FLAG_UNSORTED_TARGETS / GN_FLAG_INDEX_OUT_OF_RANGE in *flags and the
 * largest in-degree in *max_degree (both device ints, may be read later). */
int gn_coo_to_csr(const int64_t* edge_index, int64_t E, int64_t N, int32_t* rowptr,
                  int32_t* col, int32_t* flags, int32_t* max_degree, void* stream);

/* ---- whole forward pass, one call ------------------------------------------ */

/* The 16-bit inference forward of DynEdge (gnn/dynedge.py:296-350: global variables, EdgeConv
 * layers with dynamic kNN graphs, skip-cat, post-processing MLP, global pooling, one readout
 * layer) enqueued by ONE call over a caller-provided workspace - the same kernels, in the same
 * order, as the per-operator entry points below.  The plan holds the repacked weights
 * (gn_edgeconv_pack_tc layouts) and the architecture; all pointers are device pointers that must
 * stay valid while a call is in flight. */
#define GN_MAX_LAYERS 8
#define GN_MAX_DENSE 4
typedef struct gn_dynedge_conv {
  const void* Wpq;      /* node GEMM: layer 0 [2Kp, 3*x0w] bf16 [W_hi | W_hi | W_lo]; else [2Kp, ldw] */
  const float* bias_pq; /* [2Kp] */
  const void* W2p;      /* [b16, Kp] fp16 */
  int32_t d_in, Kp, b, k, slots, act, off_in, off_out, ncoord, ldw;
  int32_t coord_cols[16]; /* columns of this layer's output the next kNN graph is built on */
} gn_dynedge_conv;
typedef struct gn_dynedge_dense {
  const void* W;        /* [n, ldw]: io_dtype (post-processing), fp16 x 64 (readout, tc = 1) or fp32 */
  const float* bias;
  int32_t n, K, act, ldw, tc, pad_;
} gn_dynedge_dense;
typedef struct gn_dynedge_plan {
  int32_t F;            /* input features; D0 = 2F + 5 */
  int32_t x0w;          /* D0 rounded up to 8 */
  int32_t skip_cols;    /* columns of the skip-cat buffer */
  int32_t io_dtype;     /* GN_BF16 or GN_F16 */
  int32_t n_conv, n_post, n_readout, n_pools;
  int32_t input_k;      /* k of the input graph built when edge_index is NULL */
  int32_t pad_;
  gn_dynedge_conv conv[GN_MAX_LAYERS];
  gn_dynedge_dense post[GN_MAX_DENSE];
  gn_dynedge_dense readout[GN_MAX_DENSE];
  int32_t pools[8];     /* GN_POOL_* in output order */
} gn_dynedge_plan;

/* Bytes of workspace gn_dynedge_infer16 needs for a batch of N nodes, B events and (when the data
 * set's edge_index is passed) E edges. */
size_t gn_dynedge_infer16_workspace(const gn_dynedge_plan* plan, int64_t N, int64_t B, int64_t E,
                                    int has_edge_index);

/* x [N, ldx] fp32, batch [N] int64 sorted, n_pulses [B] int32, edge_index [2, E] int64 grouped by
 * target or NULL (then KNNGraph's graph on columns 0-2 with plan->input_k neighbours is built on
 * the device); out [B, ldo] fp32 = readout output; flags: device int32 receiving the GN_FLAG_*
 * bits of the input checks.  workspace: 256-byte aligned, >= gn_dynedge_infer16_workspace bytes. */
int gn_dynedge_infer16(const gn_dynedge_plan* plan, const float* x, int64_t ldx,
                       const int64_t* batch, const int32_t* n_pulses, int64_t N, int64_t B,
                       const int64_t* edge_index, int64_t E, void* workspace,
                       size_t workspace_bytes, float* out, int64_t ldo, int32_t* flags,
                       void* stream);

/* ---- per-event summary features ----------------------------------------- */

/* DynEdge._calculate_global_variables (models/gnn/dynedge.py:267-294) with
 * calculate_xyzt_homophily (models/utils.py:17-33) and log10(n_pulses) (:303-308):
 * gv[e] = [mean of each of the F input columns, homophily of columns 0..3 over the
 * in-edges of the event's nodes (exact float equality, 0 without edges),
 * log10(n_pulses[e])]  -> [B, F+5].
 * If x0 != NULL it also performs the "distribute" step and concatenation of
 * dynedge.py:309-320: x0[i] = [x[i, 0:F], gv[batch[i]]] (row stride ldx0, element
 * type x0_dtype; columns F+F+5 .. x0_cols are zero-filled).  x0b is an optional second
 * copy in another element type (the 16-bit skip-cat buffer of the tensor-core path). */
int gn_global_variables(const float* x, int64_t ldx, int F, const gn_graph* graph,
                        const int32_t* ptr, const int32_t* n_pulses, int64_t N,
                        int64_t B, float* gv, void* x0, int64_t ldx0, int x0_cols,
                        int x0_dtype, void* x0b, int64_t ldx0b, int x0b_cols, int x0b_dtype,
                        void* stream);

/* Last post-processing Linear + ReLU (dynedge.py:222-232, 332) fused with DynEdge._global_pooling
 * (dynedge.py:252-265): relu(A[M,K] W[N,K]^T + bias) is never stored - the GEMM's epilogue reduces
 * every column over the rows of each event and leaves acc[4][B][N] 32-bit words (min, max, and
 * the sum as 64-bit fixed point, units of 2^-32: integer addition makes the result independent
 * of execution order and of the batch an event sits in), which gn_pool_finalize turns into the
 * pooled rows.  acc: 8-byte aligned, B*N even.  `batch` [M] int64 sorted (the reference's
 * `data.batch`), `ptr` [B+1] from gn_batch_to_ptr.  acc is initialised by the call. */
int gn_tc_linear_pool(const void* A, int64_t lda, const void* W, int64_t ldw, int ab_dtype,
                      const float* bias, int64_t M, int N, int K, const int64_t* batch,
                      const int32_t* ptr, int64_t B, float* acc, void* stream);

/* acc[4][B][C] of gn_tc_linear_pool -> out[e, s*C + c] for each scheme in `schemes` (HOST array;
 * the layout of gn_segment_pool, empty events give 0) and, if out16 is not NULL, the same values
 * times scale16 as IEEE fp16 (the operand of the readout GEMM, dynedge.py:347-348). */
int gn_pool_finalize(const float* acc, const int32_t* ptr, int64_t B, int C,
                     const int32_t* schemes, int num_schemes, float* out, int64_t ldo,
                     void* out16, int64_t ld16, float scale16, void* stream);

/* DynEdge._global_pooling (dynedge.py:252-265): torch_scatter.scatter_{min,max,
 * sum,mean}(x, batch, dim=0) for each scheme in `schemes` (HOST array), results
 * concatenated along features: out[e, s*C + c].  Empty events give 0. */
int gn_segment_pool(const void* x, int x_dtype, int64_t ldx, int C, const int32_t* ptr,
                    int64_t B, const int32_t* schemes, int n_schemes, float* out,
                    int64_t ldo, void* stream);

/* ---- dense layers, fp32 -------------------------------------------------- */

/* torch.nn.Linear (+ activation): C = act(A W^T + bias).  A [M,K] fp32 (row stride
 * lda), W [N,K] fp32 row-major (torch layout), fp32 FFMA accumulate; C [M,N] of c_dtype
 * (GN_F32, or GN_BF16 / GN_F16 when it feeds a tensor-core kernel; row stride ldc).  Used
 * for the post-processing MLP (dynedge.py:222-232, 332), the readout (:244-250, 348),
 * the task affine (task/task.py:293) and the first EdgeConv's node GEMM (input features
 * stay fp32). */
int gn_linear_f32(const float* A, int64_t lda, const float* W, const float* bias,
                  void* C, int64_t ldc, int c_dtype, int64_t M, int N, int K, int act,
                  void* stream);

/* Repack the two Linear layers of one DynEdgeConv MLP for gn_edgeconv_f32:
 *   W1 [a, 2d] acting on [x_i, x_j - x_i]  ->  Wpq [2a, d] = [W1a - W1b ; W1b]
 *   W2 [b, a]                               ->  W2T [a, b]
 * (W1 [x_i, x_j - x_i] = (W1a - W1b) x_i + W1b x_j).  W2/W2T may be NULL for a
 * one-layer MLP. */
int gn_edgeconv_prepare_f32(const float* W1, const float* W2, int d, int a, int b,
                            float* Wpq, float* W2T, void* stream);

/* torch_geometric EdgeConv forward as called at components/layers.py:67 with the
 * MLP built at dynedge.py:193-204 (Linear-act[-Linear-act]):
 *   out[i] = aggr_{j in N(i)} act(W2 act(W1 [x_i, x_j - x_i] + b1) + b2)
 * fp32 FFMA path.  `pq` is caller scratch [N, 2a] floats.  Nodes without
 * in-edges give 0.  W2T == NULL: one-layer MLP (then b == a). */
int gn_edgeconv_f32(const float* x, int64_t ldx, int d, const gn_graph* graph,
                    int64_t N, const float* Wpq, const float* b1, int a,
                    const float* W2T, const float* b2, int b, int act, int aggr,
                    float* pq, float* out, int64_t ldo, void* stream);

/* ---- dense layers, bf16 tensor cores (tcgen05 / TMEM / TMA) ------------------- */

/* torch.nn.Linear (+ activation) with 16-bit operands and fp32 accumulation:
 * C = act(A W^T + bias).  A [M,K] and W [N,K] of ab_dtype (GN_BF16 or GN_F16; row strides
 * lda, ldw), both 16-byte aligned with strides that are multiples of 8 elements; bias fp32 or NULL;
 * C [M,N] of c_dtype (GN_F32 / GN_BF16 / GN_F16 saturated, row stride ldc).  Same reference call sites as
 * gn_linear_f32 plus the first (factored) Linear of each EdgeConv MLP. */
int gn_tc_linear(const void* A, int64_t lda, const void* W, int64_t ldw, int ab_dtype,
                 const float* bias, void* C, int64_t ldc, int c_dtype, int64_t M, int N, int K,
                 int act, void* stream);

/* Second half of the fused EdgeConv (components/layers.py:30-36, :67):
 *   out[i] = aggr_{j in N(i)} act(W2 act(P_i + Q_j) + b2)
 * act = GN_ACT_RELU / GN_ACT_GELU (dynedge.py:161-168) / GN_ACT_LEAKY_RELU; aggr = GN_AGGR_ADD
 * (DynEdge, dynedge.py:208), MEAN, MAX or MIN (DynEdgeConv(aggr=...), PyG semantics: a node without
 * in-edges gets 0; these three need an ELL graph, i.e. any kNN graph of this library).
 * P [N, ldp] and Q [N, ldq] are fp16 (as written by gn_tc_linear / gn_linear_f32 with c_dtype
 * GN_F16), Kp columns each (Kp = hidden width a padded to a multiple of 32 with >= 2 pad
 * columns; P[:,a] = P[:,a+1] = 1, Q pads = 0).  Q is the gathered operand: keeping it in its
 * own dense array lets it stay L2 resident.  W2p [b16, Kp] fp16 row-major (b16 = b rounded
 * up to 16; columns a, a+1 hold b2 split into fp16 hi/lo parts, remaining pads 0).
 * `slots` in {8,9,16,17,32}: in-edges per node handled in the first pass (k+1 for a kNN graph);
 * further in-edges are handled by a second pass over a compacted list, any degree is correct.
 * An ELL graph whose stride is one of those values is consumed in place (slots = stride).
 * out [N, ldo] of out_dtype (GN_BF16 / GN_F16).  If ncoord > 0, the fp32 (unrounded) values of the columns listed in
 * the HOST array coord_cols (strictly increasing) are also written to coords [N, ncoord]
 * for the next kNN.  scratch: gn_tc_edge_scratch_bytes(graph, N, slots) bytes (graph lists). */
size_t gn_tc_edge_scratch_bytes(const gn_graph* graph, int64_t N, int slots);
int gn_tc_edge_mlp_aggregate(const void* P, int64_t ldp, const void* Q, int64_t ldq, int Kp,
                             const gn_graph* graph, int64_t N, int slots, const void* W2p, int b,
                             void* out,
                             int64_t ldo, int out_dtype, float* coords,
                             const int32_t* coord_cols, int ncoord, void* scratch,
                             size_t scratch_bytes, int act, int aggr, void* stream);

/* The same op at the reference's own precision (float32 EdgeConv, layers.py:67 with
 * dynedge.py:193-204): P, Q and out are fp32, ReLU + aggr='add'.  The second Linear still runs on
 * the tensor cores: the hidden activations and W2 each as a pair of fp16 matrices (value =
 * hi + lo, 22 mantissa bits), three MMAs per K step into an fp32 accumulator.  W2hi / W2lo:
 * [b16, Kp] fp16 in the layout of gn_edgeconv_pack_tc's W2p (W2lo = fp16(W2p32 - W2hi)), Kp <= 352. */
int gn_tc_edge_mlp_aggregate_f32(const float* P, int64_t ldp, const float* Q, int64_t ldq, int Kp,
                                 const gn_graph* graph, int64_t N, int slots, const void* W2hi,
                                 const void* W2lo, int b, float* out, int64_t ldo, float* coords,
                                 const int32_t* coord_cols, int ncoord, void* scratch,
                                 size_t scratch_bytes, void* stream);

/* bf16 pieces of the fp32 matrix X[M, cols] (x = x1 + x2 + x3, x1 = bf16(x), x2 = bf16(x - x1),
 * x3 = bf16(x - x1 - x2)) laid out so that ONE gn_tc_linear call evaluates x W^T beyond bf16
 * accuracy - the fp32 path's node-level Linear layers (dynedge.py:193-232) on the tensor cores:
 *   terms = 3: out[M, 3 kc] = [x1 | x2 | x1]                against W = [w1 | w1 | w2]  (16 bits)
 *   terms = 6: out[M, 6 kc] = [x1 | x1 | x1 | x2 | x2 | x3] against [w1 | w2 | w3 | w1 | w2 | w1]
 *              (every product term down to 2^-16 of the leading one: fp32-grade)
 * Blocks are kc columns wide, columns cols..kc-1 zero.  No counterpart in the reference (it
 * multiplies in fp32). */
int gn_split_bf16(const float* X, int64_t ldx, int64_t M, int cols, int kc, int terms, void* out,
                  int64_t ldo, void* stream);
/* Both layouts from one read of X: out3 as terms = 3, out6 as terms = 6 (a layer output feeds the
 * post-processing MLP with three terms and the next layer's node GEMM with six). */
int gn_split_bf16_both(const float* X, int64_t ldx, int64_t M, int cols, int kc, void* out3,
                       int64_t ldo3, void* out6, int64_t ldo6, void* stream);

/* Training forward of the same op: additionally records the ReLU pattern of the second Linear
 * (what torch autograd keeps as the saved output of the ReLU in PyG's EdgeConv.message,
 * components/layers.py:67) as bits, so that the backward pass need not recompute Z = H W2^T.
 * relu_mask [8, mask_ld] uint32: bit c of word (g, r) = [z(r, 32 g + c) > 0], where slot
 * e = i*slots + t of the ELL graph is row r = (e / used)*64 + e % used, used = (64/slots)*slots.
 * mask_ld >= gn_tc_edge_relu_mask_ld(N, slots).  Needs an ELL graph whose stride is a supported
 * slot count (every kNN graph of this library); otherwise GN_ERR_UNSUPPORTED. */
int64_t gn_tc_edge_relu_mask_ld(int64_t N, int slots);
int gn_tc_edge_mlp_aggregate_mask(const void* P, int64_t ldp, const void* Q, int64_t ldq, int Kp,
                                  const gn_graph* graph, int64_t N, int slots, const void* W2p,
                                  int b, void* out, int64_t ldo, int out_dtype, float* coords,
                                  const int32_t* coord_cols, int ncoord, void* scratch,
                                  size_t scratch_bytes, uint32_t* relu_mask, int64_t mask_ld,
                                  void* stream);

/* ---- task heads ----------------------------------------------------------- */

/* The `_forward` of a reconstruction head (models/task/reconstruction.py) applied
 * to the affine output y [M, p]; out is [M, p_out] (p_out = 4 for
 * DIRECTION_KAPPA, 2 for *_KAPPA angles, else p). */
int gn_task_head(const float* y, int64_t M, int p, int kind, float* out,
                 void* stream);

/* out = 16-bit(X * scale): operand conversion in front of gn_tc_linear (the readout Linear,
 * gnn/dynedge.py:347-348, runs with fp16 operands scaled by a power of two so that pooled sums
 * stay in range; fp16 conversions saturate at +-65504). */
int gn_convert_scale_16(const float* X, int64_t ldx, void* out, int out_dtype, int64_t ldo,
                        int64_t M, int N, float scale, void* stream);

/* ---- add_norm_layer=True (gnn/dynedge.py:202-203, :224-226) ---------------------- */

/* X[m, :] <- act(LayerNorm(X[m, :]) * gamma + beta) in place (torch.nn.LayerNorm: biased
 * variance, eps under the root; gamma / beta may be NULL), N <= 512. */
int gn_layernorm_act_f32(float* X, int64_t ldx, int64_t M, int N, const float* gamma,
                         const float* beta, float eps, int act, void* stream);

/* Pre-activations of the first EdgeConv Linear for every slot of an ELL graph:
 * out[i*stride + t, :] = P_i + Q_j with pq = [P | Q] [N, 2a] as written by gn_linear_f32 with the
 * factored weights of gn_edgeconv_prepare_f32 (zero rows for empty slots). */
int gn_edge_preact_f32(const float* pq, int a, const gn_graph* graph, int64_t N, float* out,
                       int64_t ldo, void* stream);

/* aggr="add" over materialised per-slot messages: out[i, :] = sum of Z[i*stride + t, :] over the
 * filled slots of node i. */
int gn_slot_sum_f32(const float* Z, int64_t ldz, int b, const gn_graph* graph, int64_t N,
                    float* out, int64_t ldo, void* stream);

/* The same with PyG's other aggregations (aggr = GN_AGGR_ADD / MEAN / MAX / MIN): ParticleNeT builds
 * its DynEdgeConv layers with aggr="mean" (gnn/particlenet.py:193-198).  A node without in-edges
 * gives 0. */
int gn_slot_aggregate_f32(const float* Z, int64_t ldz, int b, const gn_graph* graph, int64_t N,
                          int aggr, float* out, int64_t ldo, void* stream);

/* ---- backward pass of the 16-bit training path (BASELINE config 3) ------------- */
/* The reference holds no backward code: gradients come from torch autograd through the modules
 * cited at each entry.  All matrices row-major; 16-bit operands 16-byte aligned with row strides
 * that are multiples of 8 elements wherever they feed gn_tc_linear*. */

/* C += A W^T with fp32 reductions (red.add): weight gradients dW = dY^T X of nn.Linear
 * (gnn/dynedge.py:193-232), as A = dY^T [N_out, rows], W = X^T [K_in, rows]; also input gradients
 * added onto an existing gradient (dX += dY W).  A [M,K], W [N,K] of ab_dtype; C [M,N] fp32 must
 * hold the value to add onto (zeros for a fresh gradient).  `splits` > 1 cuts K into that many
 * ranges handled by different CTAs (0 = choose so that all SMs are busy). */
int gn_tc_linear_accumulate(const void* A, int64_t lda, const void* W, int64_t ldw, int ab_dtype,
                            float* C, int64_t ldc, int64_t M, int N, int K, int splits,
                            void* stream);

/* dst[c, r] = src[r, c] for 16-bit elements (operand layout for the weight-gradient GEMMs). */
int gn_transpose_16(const void* src, int64_t ld_src, int64_t rows, int cols, void* dst,
                    int64_t ld_dst, void* stream);

/* out[n] += sum_m X[m, n]: bias gradients of nn.Linear.  X of x_dtype (GN_F32 / GN_BF16). */
int gn_colsum_accumulate(const void* X, int x_dtype, int64_t ldx, int64_t M, int N, float* out,
                         void* stream);

/* out = dY * [Y > 0] converted to bf16: backward of the ReLU after every Linear
 * (dynedge.py:199-204, :224-232).  Y == NULL: plain conversion to bf16.  dY GN_F32 / GN_BF16. */
int gn_relu_backward(const void* dY, int dy_dtype, int64_t ld_dy, const void* Y, int y_dtype,
                     int64_t ld_y, void* out, int64_t ld_out, int64_t M, int N, void* stream);

/* Repack the weights of one EdgeConv MLP (nn.Linear W1 [a, 2d], b1, W2 [b, a], b2; reference
 * gnn/dynedge.py:193-204) for the tensor-core kernels, in one launch: Wpq [2Kp, Kd] (rows 0..a-1 =
 * W1[:, :d] - W1[:, d:], rows Kp.. = W1[:, d:]; wpq_dtype GN_F32 for the first layer, else 16-bit),
 * bias_pq [2Kp] fp32 (b1, then the two "ones" columns), W2p [b16, Kp] fp16 (W2; columns a, a+1 =
 * b2 split in fp16 hi/lo).  Optional, for the backward pass: W2b = bf16(W2p), W2bT [Kp, b16] its
 * transpose, WpqT [Kd, 2Kp] the transpose of a 16-bit Wpq.  b1, b2 may be NULL (no bias).
 * `act` is the MLP's activation (GN_ACT_RELU / GELU / LEAKY_RELU): the "ones" columns pass
 * through it in the edge kernel, so b2 is stored divided by act(1). */
int gn_edgeconv_pack_tc(const float* W1, const float* b1, const float* W2, const float* b2, int d,
                        int a, int b, int Kp, int Kd, void* Wpq, int wpq_dtype, float* bias_pq,
                        void* W2p, void* W2b, void* W2bT, void* WpqT, int act, void* stream);

/* Re-materialise the hidden activations of the EdgeConv message MLP (PyG EdgeConv.message,
 * components/layers.py:67) for every slot of an ELL graph (stride = slots per node, -1 = empty):
 * H[i*stride + t, :] = ReLU(P_i + Q_j), zero rows for empty slots; bf16.  HT (optional) receives
 * the transpose [Kp, N*stride] (row stride ldht); H may be NULL when only HT is wanted. */
int gn_edge_build_h(const void* P, int64_t ldp, const void* Q, int64_t ldq, int Kp,
                    const gn_graph* graph, int64_t N, void* H, int64_t ldh, void* HT,
                    int64_t ldht, void* stream);

/* G2[e, :] = dOut[target(e), :] * [Z[e, :] > 0] (backward of the ReLU after the second Linear and
 * of the aggr="add" sum, dynedge.py:206-211); Z [N*stride, b] bf16 recomputed pre-activations,
 * dOut [N, b] fp32; writes G2 [N*stride, b] and its transpose G2T [b, N*stride], bf16. */
int gn_edge_grad_z(const void* Z, int64_t ldz, const float* dOut, int64_t ldd, int b,
                   const gn_graph* graph, int64_t N, void* G2, int64_t ldg, void* G2T,
                   int64_t ldgt, void* stream);

/* The same from the bits recorded by gn_tc_edge_mlp_aggregate_mask (no Z): G2[e, :] =
 * dOut[target(e), :] where the forward's z(e, :) was positive, else 0. */
int gn_edge_grad_z_mask(const uint32_t* relu_mask, int64_t mask_ld, const float* dOut, int64_t ldd,
                        int b, const gn_graph* graph, int64_t N, void* G2, int64_t ldg, void* G2T,
                        int64_t ldgt, void* stream);

/* Slots of an ELL graph grouped by SOURCE node (the transposed adjacency autograd's scatter-add
 * of x_j's gradient walks implicitly, components/layers.py:67): for node j the slots e with
 * nbr[e] == j, ascending, are rslot[rstart[j] .. rstart[j] + rcount[j]).  ptr [B+1] event offsets
 * (edges never leave an event); rstart, rcount [N], rslot [N*stride] int32. */
int gn_ell_reverse(const gn_graph* graph, const int32_t* ptr, int64_t N, int64_t B,
                   int32_t* rstart, int32_t* rcount, int32_t* rslot, void* stream);

/* Backward of h = ReLU(P_i + Q_j) without global atomics: m[e] = dH[e] * [P_i + Q_j > 0] (fp16
 * sum, as the forward kernel forms it); dP[i] = sum of m over the slots of node i, dQ[j] = sum of
 * m over the slots that name j as neighbour; bf16 outputs (overwritten), optional fp32 copies
 * dP32 / dQ32.  dH bf16 [N*stride, lddh]; P, Q fp16; Kp a multiple of 8; 16-byte aligned rows.
 * rstart == NULL (default): one streaming pass over dH; dQ is accumulated in dQ32 (required,
 * zeroed by the call) with 16-byte vector reductions and then rounded to dQ (the order of those
 * additions is not fixed: results agree to fp32 rounding between runs).  rstart, rcount, rslot
 * from gn_ell_reverse: the list form - every dQ row is a plain sum in slot order
 * (bit-reproducible), at the price of reading dH twice. */
int gn_edge_grad_pq_gather(const void* dH, int64_t lddh, const void* P, int64_t ldp, const void* Q,
                           int64_t ldq, int Kp, const gn_graph* graph, const int32_t* ptr,
                           int64_t N, int64_t B, const int32_t* rstart, const int32_t* rcount,
                           const int32_t* rslot, void* dP, int64_t lddp, void* dQ, int64_t lddq,
                           float* dP32, int64_t ldp32, float* dQ32, int64_t ldq32, void* stream);

/* Backward of h = ReLU(P_i + Q_j): m = dH * [H > 0]; dP[i] = sum over the slots of node i (bf16,
 * overwritten; dP32, if not NULL, receives the same sums unrounded), dQ[j] += m (fp32, must be
 * zeroed by the caller). */
int gn_edge_grad_pq(const void* dH, int64_t lddh, const void* H, int64_t ldh, int Kp,
                    const gn_graph* graph, int64_t N, void* dP, int64_t ldp, float* dQ,
                    int64_t ldq, float* dP32, int64_t ldp32, void* stream);

/* out[c, k] += sum_n G[n, c] X[n, k], all fp32, K <= 64: weight gradient of the first EdgeConv
 * Linear (dynedge.py:193-199) with respect to the raw input features - its two halves are
 * subtracted (the Linear acts on [x_i, x_j - x_i]), so it is kept out of 16-bit arithmetic. */
int gn_outer_accumulate_f32(const float* G, int64_t ldg, int cols, const float* X, int64_t ldx,
                            int K, int64_t N, float* out, int64_t ldo, void* stream);

/* Backward of gn_segment_pool (torch_scatter.scatter_{min,max,sum,mean}, dynedge.py:252-265):
 * dX [N, C] fp32 (overwritten) from d_pooled [B, C*n_schemes]; min / max send their gradient to
 * the first row of the event that attains the pooled value. */
int gn_segment_pool_backward(const void* x, int x_dtype, int64_t ldx, int C, const int32_t* ptr,
                             int64_t B, const int32_t* schemes, int n_schemes,
                             const float* pooled, int64_t ld_pooled, const float* d_pooled,
                             int64_t ld_dpooled, float* dX, int64_t ld_dx, void* stream);

/* CSR -> slot table (ELL) for the backward kernels, which address edges as (node, slot):
 * nbr[i*stride + t] = t-th in-edge of node i in CSR order, -1 beyond its degree; deg[i] =
 * min(degree, stride).  A node with more than `stride` in-edges sets GN_FLAG_DEGREE_OVERFLOW in
 * *flags (may be NULL) and loses the excess: the caller picks stride >= the maximum degree
 * reported by gn_coo_to_csr.  Layout plumbing for the user edge_index (PyG Batch) of layer 1. */
int gn_csr_to_ell(const int32_t* rowptr, const int32_t* col, int64_t N, int stride, int32_t* nbr,
                  int32_t* deg, int32_t* flags, void* stream);

/* Zero `bytes` bytes at p on `stream` (the accumulators of gn_tc_linear_accumulate,
 * gn_colsum_accumulate, ... inside a captured launch sequence, where the caller's own fill kernels
 * would run out of order). */
int gn_zero(void* p, size_t bytes, void* stream);

/* Launch sequences composed by the caller, without idle GPU between the kernels: every gn_* call
 * between gn_graph_capture_begin and gn_graph_capture_launch is given the stream `begin` returns
 * (a private stream in capture mode: nothing runs); `launch` folds the captured sequence into the
 * executable CUDA graph cached under `key` (cudaGraphExecUpdate: same topology, new parameters -
 * batches of any size; instantiated on first use or when the topology changed) and launches it on
 * `stream`.  Only this library's calls (and no allocation, no synchronisation) may happen in
 * between; one capture per device at a time.  `launch` returns GN_ERR_UNSUPPORTED when the
 * sequence could not be captured: issue the same calls directly then.  `abort` drops a capture.
 * gn_dynedge_infer16 does this internally.  No counterpart in the reference. */
int gn_graph_capture_begin(void** capture_stream);
int gn_graph_capture_launch(const void* key, void* stream);
int gn_graph_capture_abort(void);

#ifdef __cplusplus
}
#endif
#endif /* GRAPHNET_B200_H_ */
