// Native driver of the 16-bit DynEdge inference forward (reference gnn/dynedge.py:296-350): ONE
// C-ABI call enqueues every kernel of the pass - event offsets, CSR / kNN input graph, summary
// features, four EdgeConv layers with their node GEMMs and dynamic kNN graphs, the
// post-processing MLP with the pooled epilogue, the readout.  It composes the entry points of this
// library (no new device code) over one caller-provided workspace.
//
// Why it exists: the Python orchestration of the same sequence (graphnet_b200/dynedge_bf16.py)
// costs ~0.56 ms of host time per forward (~45 launches through ctypes, ~30 torch.empty calls),
// which is no less than the kernels themselves take on a B200 (0.64 ms at C2) - the step followed
// the host.  From C++ the pass is enqueued in ~0.1 ms.
#include <stdlib.h>
#include <string.h>

#include <mutex>

#include "common.cuh"

using namespace gn;

namespace {

struct Carver {
  uint8_t* base;
  size_t off = 0;
  explicit Carver(void* b) : base(reinterpret_cast<uint8_t*>(b)) {}
  template <typename T>
  T* take(size_t count) {
    off = (off + 255) & ~size_t(255);
    T* p = base ? reinterpret_cast<T*>(base + off) : nullptr;
    off += count * sizeof(T);
    return p;
  }
};

inline int round_up(int v, int m) { return (v + m - 1) / m * m; }
inline size_t io_bytes(int dtype) { return dtype == GN_F32 ? 4 : 2; }

// every buffer of one forward; `base == nullptr` only measures
struct Buffers {
  int32_t *ptr, *rowptr, *col, *maxdeg, *nbr[2], *deg[2];
  float *gv, *coords, *acc, *pooled;
  uint8_t *x0s, *skip, *pq, *h[2], *hb16, *scratch;
  size_t scratch_bytes, total;
};

int layout(const gn_dynedge_plan* pl, int64_t N, int64_t B, int64_t E, bool has_edges, void* base,
           Buffers* b) {
  Carver c(base);
  const size_t io = io_bytes(pl->io_dtype);
  int kmax = pl->input_k, kpmax = 0, hmax = 0;
  for (int l = 0; l < pl->n_conv; ++l) {
    kmax = kmax > pl->conv[l].k ? kmax : pl->conv[l].k;
    kpmax = kpmax > pl->conv[l].Kp ? kpmax : pl->conv[l].Kp;
  }
  for (int i = 0; i + 1 < pl->n_post; ++i) hmax = hmax > round_up(pl->post[i].n, 8) ? hmax : round_up(pl->post[i].n, 8);
  b->ptr = c.take<int32_t>(B + 1);
  b->maxdeg = c.take<int32_t>(1);
  b->rowptr = has_edges ? c.take<int32_t>(N + 1) : nullptr;
  b->col = has_edges ? c.take<int32_t>(E > 0 ? E : 1) : nullptr;
  for (int i = 0; i < 2; ++i) {
    b->nbr[i] = c.take<int32_t>((size_t)N * (kmax + 1));
    b->deg[i] = c.take<int32_t>(N);
  }
  b->gv = c.take<float>((size_t)B * (pl->F + 5));
  b->x0s = c.take<uint8_t>((size_t)N * 3 * pl->x0w * 2);
  b->skip = c.take<uint8_t>((size_t)N * pl->skip_cols * io);
  b->pq = c.take<uint8_t>((size_t)N * 2 * kpmax * 2);
  b->coords = c.take<float>((size_t)N * 16);
  b->h[0] = c.take<uint8_t>((size_t)N * (hmax > 0 ? hmax : 8) * io);
  b->h[1] = c.take<uint8_t>((size_t)N * (hmax > 0 ? hmax : 8) * io);
  const int C = pl->post[pl->n_post - 1].n;
  b->acc = c.take<float>((size_t)4 * B * C);
  b->pooled = c.take<float>((size_t)B * C * pl->n_pools);
  b->hb16 = c.take<uint8_t>((size_t)B * C * pl->n_pools * 2);
  // graph lists of the edge kernel: a CSR input graph needs them, and so does a slot table whose
  // stride is not one of the kernel's slot counts (kNN graphs with k + 1 in {8, 9, 16, 17, 32} are
  // consumed in place); sized for the worst layer
  size_t sb = 256;
  for (int l = 0; l < pl->n_conv; ++l) {
    gn_graph g;
    memset(&g, 0, sizeof(g));
    g.col = reinterpret_cast<const int32_t*>(8);   // sizes only: never dereferenced
    if (l == 0 && has_edges) {
      g.rowptr = reinterpret_cast<const int32_t*>(8);
      g.num_edges = E;
    } else {
      g.deg = reinterpret_cast<const int32_t*>(8);
      g.ell_stride = (l == 0 ? pl->input_k : pl->conv[l - 1].k) + 1;
    }
    const size_t s0 = gn_tc_edge_scratch_bytes(&g, N, pl->conv[l].slots);
    sb = sb > s0 ? sb : s0;
  }
  b->scratch = c.take<uint8_t>(sb);
  b->scratch_bytes = sb;
  b->total = c.off + 256;
  return GN_OK;
}

}  // namespace

extern "C" size_t gn_dynedge_infer16_workspace(const gn_dynedge_plan* plan, int64_t N, int64_t B,
                                               int64_t E, int has_edge_index) {
  if (!plan || N < 0 || B < 0 || plan->n_post < 1) return 0;
  Buffers b;
  layout(plan, N, B, E, has_edge_index != 0, nullptr, &b);
  return b.total;
}

// every launch of the forward, in order, on `stream`
static int enqueue_forward(const gn_dynedge_plan* pl, const float* x, int64_t ldx,
                           const int64_t* batch, const int32_t* n_pulses, int64_t N, int64_t B,
                           const int64_t* edge_index, int64_t E, void* workspace,
                           size_t workspace_bytes, float* out, int64_t ldo, int32_t* flags,
                           void* stream) {
  GN_CHECK_ARG(pl && x && batch && n_pulses && out && flags && workspace,
               "gn_dynedge_infer16: null pointer");
  GN_CHECK_ARG(pl->n_conv >= 1 && pl->n_conv <= GN_MAX_LAYERS && pl->n_post >= 1 &&
                   pl->n_post <= GN_MAX_DENSE && pl->n_readout == 1 && pl->n_pools >= 1 &&
                   pl->n_pools <= 8,
               "gn_dynedge_infer16: plan out of range");
  GN_CHECK_ARG(pl->io_dtype == GN_BF16 || pl->io_dtype == GN_F16, "gn_dynedge_infer16: io_dtype");
  GN_CHECK_ARG(N >= 1 && B >= 1, "gn_dynedge_infer16: empty batch");
  const bool has_edges = edge_index != nullptr;
  Buffers b;
  layout(pl, N, B, E, has_edges, workspace, &b);
  GN_CHECK_ARG(workspace_bytes >= b.total && (reinterpret_cast<uintptr_t>(workspace) & 255) == 0,
               "gn_dynedge_infer16: workspace too small (%zu < %zu) or not 256-byte aligned",
               workspace_bytes, b.total);
  cudaStream_t st = as_stream(stream);
  const size_t io = io_bytes(pl->io_dtype);
  int rc;
#define GN_TRY(call)            \
  do {                          \
    rc = (call);                \
    if (rc != GN_OK) return rc; \
  } while (0)

  GN_CUDA(cudaMemsetAsync(flags, 0, sizeof(int32_t), st));
  GN_TRY(gn_batch_to_ptr(batch, N, B, b.ptr, flags, stream));
  // ---- input graph: the data set's edge_index as CSR, or KNNGraph's graph built here
  gn_graph graph;
  memset(&graph, 0, sizeof(graph));
  const int32_t xyz[3] = {0, 1, 2};
  if (has_edges) {
    // (zero-filled: with ids out of range part of rowptr stays unwritten - consumers must then
    // see empty rows, not garbage; the flags word reports the error)
    GN_CUDA(cudaMemsetAsync(b.rowptr, 0, (size_t)(N + 1) * sizeof(int32_t), st));
    GN_TRY(gn_coo_to_csr(edge_index, E, N, b.rowptr, b.col, flags, b.maxdeg, stream));
    graph.col = b.col;
    graph.rowptr = b.rowptr;
    graph.num_edges = E;
  } else {
    GN_TRY(gn_knn_graph(x, ldx, xyz, 3, b.ptr, N, B, pl->input_k, 0, b.nbr[1], pl->input_k + 1,
                        b.deg[1], stream));
    graph.col = b.nbr[1];
    graph.deg = b.deg[1];
    graph.ell_stride = pl->input_k + 1;
  }
  // ---- summary features, x0 = [x, gv[batch]] as the bf16 split and into the skip buffer
  bool needs_zero = false;
  for (int l = 0; l < pl->n_conv; ++l) needs_zero = needs_zero || (pl->conv[l].b % 8) != 0;
  if (needs_zero) GN_CUDA(cudaMemsetAsync(b.skip, 0, (size_t)N * pl->skip_cols * io, st));
  GN_TRY(gn_global_variables(x, ldx, pl->F, &graph, b.ptr, n_pulses, N, B, b.gv, b.x0s,
                             3 * pl->x0w, pl->x0w, GN_BF16X3, b.skip, pl->skip_cols, pl->x0w,
                             pl->io_dtype, stream));
  // ---- EdgeConv layers
  for (int l = 0; l < pl->n_conv; ++l) {
    const gn_dynedge_conv& cv = pl->conv[l];
    const int64_t ldpq = 2 * cv.Kp;
    if (l == 0) {
      GN_TRY(gn_tc_linear(b.x0s, 3 * pl->x0w, cv.Wpq, 3 * pl->x0w, GN_BF16, cv.bias_pq, b.pq, ldpq,
                          GN_F16, N, 2 * cv.Kp, 3 * pl->x0w, GN_ACT_NONE, stream));
    } else {
      GN_TRY(gn_tc_linear(b.skip + (size_t)cv.off_in * io, pl->skip_cols, cv.Wpq, cv.ldw,
                          pl->io_dtype, cv.bias_pq, b.pq, ldpq, GN_F16, N, 2 * cv.Kp, cv.d_in,
                          GN_ACT_NONE, stream));
    }
    const bool last = l + 1 == pl->n_conv;
    size_t sbytes = gn_tc_edge_scratch_bytes(&graph, N, cv.slots);
    GN_CHECK_ARG(sbytes <= b.scratch_bytes, "gn_dynedge_infer16: edge scratch");
    GN_TRY(gn_tc_edge_mlp_aggregate(b.pq, ldpq, b.pq + (size_t)cv.Kp * 2, ldpq, cv.Kp, &graph, N,
                                    cv.slots, cv.W2p, cv.b, b.skip + (size_t)cv.off_out * io,
                                    pl->skip_cols, pl->io_dtype, last ? nullptr : b.coords,
                                    cv.coord_cols, last ? 0 : cv.ncoord, b.scratch, sbytes, cv.act,
                                    GN_AGGR_ADD, stream));
    if (!last) {
      int32_t cc[16];
      for (int t = 0; t < cv.ncoord; ++t) cc[t] = t;
      GN_TRY(gn_knn_graph(b.coords, cv.ncoord, cc, cv.ncoord, b.ptr, N, B, cv.k, 0, b.nbr[l & 1],
                          cv.k + 1, b.deg[l & 1], stream));
      memset(&graph, 0, sizeof(graph));
      graph.col = b.nbr[l & 1];
      graph.deg = b.deg[l & 1];
      graph.ell_stride = cv.k + 1;
    }
  }
  // ---- post-processing MLP; its last Linear reduces per event in the epilogue
  const uint8_t* h = b.skip;
  int64_t ldh = pl->skip_cols;
  for (int i = 0; i + 1 < pl->n_post; ++i) {
    const gn_dynedge_dense& d = pl->post[i];
    const int64_t ldo_i = round_up(d.n, 8);
    uint8_t* o = b.h[i & 1];
    GN_TRY(gn_tc_linear(h, ldh, d.W, d.ldw, pl->io_dtype, d.bias, o, ldo_i, pl->io_dtype, N, d.n,
                        d.K, d.act, stream));
    h = o;
    ldh = ldo_i;
  }
  const gn_dynedge_dense& dl = pl->post[pl->n_post - 1];
  GN_CHECK_ARG(dl.act == GN_ACT_RELU, "gn_dynedge_infer16: the pooled epilogue needs ReLU");
  GN_TRY(gn_tc_linear_pool(h, ldh, dl.W, dl.ldw, pl->io_dtype, dl.bias, N, dl.n, dl.K, batch, b.ptr,
                           B, b.acc, stream));
  const gn_dynedge_dense& ro = pl->readout[0];
  const int64_t pc = (int64_t)dl.n * pl->n_pools;
  GN_CHECK_ARG(ro.K == pc, "gn_dynedge_infer16: readout width %d != pooled width %lld", ro.K,
               (long long)pc);
  GN_TRY(gn_pool_finalize(b.acc, b.ptr, B, dl.n, pl->pools, pl->n_pools, b.pooled, pc,
                          ro.tc ? b.hb16 : nullptr, pc, 1.0f / 64.0f, stream));
  // ---- readout
  if (ro.tc) {
    GN_TRY(gn_tc_linear(b.hb16, pc, ro.W, ro.ldw, GN_F16, ro.bias, out, ldo, GN_F32, B, ro.n, ro.K,
                        ro.act, stream));
  } else {
    GN_TRY(gn_linear_f32(b.pooled, pc, reinterpret_cast<const float*>(ro.W), ro.bias, out, ldo,
                         GN_F32, B, ro.n, ro.K, ro.act, stream));
  }
#undef GN_TRY
  return GN_OK;
}

// The launches above leave ~2 us of idle GPU between consecutive kernels (58 us of a 0.67 ms step at
// C2).  A CUDA graph removes those gaps, but a captured graph is only good for the shapes and
// pointers it was captured with, and a batch of events has a different N every time.  So every
// call CAPTURES the launch sequence (on a private stream: nothing runs, and the caller's stream may
// be the legacy default stream, which cannot capture) and folds it into the cached executable
// graph with cudaGraphExecUpdate - same topology, new kernel parameters, grids and memset sizes -
// then launches that on the caller's stream.  A change of topology (another model, edge_index
// present or not) re-instantiates.  GN_NO_GRAPH=1, or a caller stream that is itself capturing:
// direct launches.
namespace {
struct GraphSlot {
  const void* plan = nullptr;     // key: the caller's plan object and whether it brings edges
  bool has_edges = false;
  cudaGraphExec_t exec = nullptr;
};
struct GraphCache {
  std::mutex mu;
  cudaStream_t cap = nullptr;
  GraphSlot slots[4];             // a few models per device, replaced round robin
  int next = 0;
};
GraphCache g_graphs[64];
}  // namespace

extern "C" int gn_dynedge_infer16(const gn_dynedge_plan* pl, const float* x, int64_t ldx,
                                  const int64_t* batch, const int32_t* n_pulses, int64_t N,
                                  int64_t B, const int64_t* edge_index, int64_t E, void* workspace,
                                  size_t workspace_bytes, float* out, int64_t ldo, int32_t* flags,
                                  void* stream) {
  static const bool no_graph = getenv("GN_NO_GRAPH") != nullptr;
  cudaStream_t st = as_stream(stream);
  cudaStreamCaptureStatus cs = cudaStreamCaptureStatusNone;
  if (!no_graph) GN_CUDA(cudaStreamIsCapturing(st, &cs));
  if (no_graph || cs != cudaStreamCaptureStatusNone)
    return enqueue_forward(pl, x, ldx, batch, n_pulses, N, B, edge_index, E, workspace,
                           workspace_bytes, out, ldo, flags, stream);
  int dev = 0;
  GN_CUDA(cudaGetDevice(&dev));
  GraphCache& gc = g_graphs[dev & 63];
  std::lock_guard<std::mutex> lock(gc.mu);
  if (gc.cap == nullptr) GN_CUDA(cudaStreamCreateWithFlags(&gc.cap, cudaStreamNonBlocking));
  GN_CUDA(cudaStreamBeginCapture(gc.cap, cudaStreamCaptureModeThreadLocal));
  const int rc = enqueue_forward(pl, x, ldx, batch, n_pulses, N, B, edge_index, E, workspace,
                                 workspace_bytes, out, ldo, flags, gc.cap);
  cudaGraph_t graph = nullptr;
  const cudaError_t ce = cudaStreamEndCapture(gc.cap, &graph);
  if (rc == GN_ERR_INVALID_ARGUMENT || rc == GN_ERR_UNSUPPORTED) {   // the message is already set
    if (graph) cudaGraphDestroy(graph);
    (void)cudaGetLastError();
    return rc;
  }
  if (rc != GN_OK || ce != cudaSuccess || graph == nullptr) {
    // something in the sequence cannot be captured here (e.g. the first launch of a kernel the
    // driver has not loaded yet): run this call directly - it also serves as the warm-up
    if (graph) cudaGraphDestroy(graph);
    (void)cudaGetLastError();
    return enqueue_forward(pl, x, ldx, batch, n_pulses, N, B, edge_index, E, workspace,
                           workspace_bytes, out, ldo, flags, stream);
  }
  GraphSlot* slot = nullptr;
  for (GraphSlot& c : gc.slots)
    if (c.exec != nullptr && c.plan == pl && c.has_edges == (edge_index != nullptr)) slot = &c;
  if (slot == nullptr) {
    slot = &gc.slots[gc.next];
    gc.next = (gc.next + 1) % 4;
    if (slot->exec != nullptr) cudaGraphExecDestroy(slot->exec);
    slot->exec = nullptr;
    slot->plan = pl;
    slot->has_edges = edge_index != nullptr;
  }
  if (slot->exec != nullptr) {
    cudaGraphExecUpdateResultInfo info;
    if (cudaGraphExecUpdate(slot->exec, graph, &info) != cudaSuccess) {
      (void)cudaGetLastError();          // another topology: build a new executable graph
      cudaGraphExecDestroy(slot->exec);
      slot->exec = nullptr;
    }
  }
  if (slot->exec == nullptr) {
    const cudaError_t ie = cudaGraphInstantiate(&slot->exec, graph, 0);
    if (ie != cudaSuccess) {
      cudaGraphDestroy(graph);
      slot->exec = nullptr;
      set_error("gn_dynedge_infer16: cudaGraphInstantiate failed: %s", cudaGetErrorString(ie));
      return GN_ERR_CUDA;
    }
  }
  const cudaError_t le = cudaGraphLaunch(slot->exec, st);
  cudaGraphDestroy(graph);
  if (le != cudaSuccess) {
    set_error("gn_dynedge_infer16: cudaGraphLaunch failed: %s", cudaGetErrorString(le));
    return GN_ERR_CUDA;
  }
  return GN_OK;
}

// The same mechanism for a launch sequence the CALLER composes from this library's entry points
// (the fp32 forward of dynedge_x3.py, any pipeline of gn_* calls): between
// gn_graph_capture_begin and gn_graph_capture_launch every call must be given the returned stream;
// nothing runs until the launch, which folds the captured sequence into the executable graph
// cached under `key` (cudaGraphExecUpdate; instantiated on first use or when the topology
// changed) and launches it on `stream`.  One capture per device at a time (the pair holds a
// per-device lock).  gn_graph_capture_launch returns GN_ERR_UNSUPPORTED when the sequence could
// not be captured (e.g. first launch of a kernel the driver had not loaded): the caller then
// issues the same calls directly.
extern "C" int gn_graph_capture_begin(void** capture_stream) {
  GN_CHECK_ARG(capture_stream != nullptr, "gn_graph_capture_begin: null pointer");
  int dev = 0;
  GN_CUDA(cudaGetDevice(&dev));
  GraphCache& gc = g_graphs[dev & 63];
  gc.mu.lock();
  cudaError_t e = cudaSuccess;
  if (gc.cap == nullptr) e = cudaStreamCreateWithFlags(&gc.cap, cudaStreamNonBlocking);
  // (relaxed mode: the caller's allocator may call cudaMalloc between the launches; the capture
  // stream is private, nothing else can depend on it)
  if (e == cudaSuccess) e = cudaStreamBeginCapture(gc.cap, cudaStreamCaptureModeRelaxed);
  if (e != cudaSuccess) {
    gc.mu.unlock();
    set_error("gn_graph_capture_begin: %s", cudaGetErrorString(e));
    (void)cudaGetLastError();
    return GN_ERR_CUDA;
  }
  *capture_stream = gc.cap;
  return GN_OK;
}

static int graph_capture_finish(const void* key, cudaStream_t st, bool launch) {
  int dev = 0;
  GN_CUDA(cudaGetDevice(&dev));
  GraphCache& gc = g_graphs[dev & 63];
  cudaGraph_t graph = nullptr;
  const cudaError_t ce = cudaStreamEndCapture(gc.cap, &graph);
  int rc = GN_OK;
  if (!launch || ce != cudaSuccess || graph == nullptr) {
    if (graph) cudaGraphDestroy(graph);
    (void)cudaGetLastError();
    if (launch) {
      set_error("gn_graph_capture_launch: the sequence could not be captured (%s)",
                cudaGetErrorString(ce));
      rc = GN_ERR_UNSUPPORTED;
    }
    gc.mu.unlock();
    return rc;
  }
  GraphSlot* slot = nullptr;
  for (GraphSlot& c : gc.slots)
    if (c.exec != nullptr && c.plan == key && !c.has_edges) slot = &c;
  if (slot == nullptr) {
    slot = &gc.slots[gc.next];
    gc.next = (gc.next + 1) % 4;
    if (slot->exec != nullptr) cudaGraphExecDestroy(slot->exec);
    slot->exec = nullptr;
    slot->plan = key;
    slot->has_edges = false;
  }
  if (slot->exec != nullptr) {
    cudaGraphExecUpdateResultInfo info;
    if (cudaGraphExecUpdate(slot->exec, graph, &info) != cudaSuccess) {
      (void)cudaGetLastError();
      cudaGraphExecDestroy(slot->exec);
      slot->exec = nullptr;
    }
  }
  cudaError_t e = cudaSuccess;
  if (slot->exec == nullptr) e = cudaGraphInstantiate(&slot->exec, graph, 0);
  if (e == cudaSuccess) e = cudaGraphLaunch(slot->exec, st);
  cudaGraphDestroy(graph);
  if (e != cudaSuccess) {
    if (slot->exec) cudaGraphExecDestroy(slot->exec);
    slot->exec = nullptr;
    set_error("gn_graph_capture_launch: %s", cudaGetErrorString(e));
    (void)cudaGetLastError();
    rc = GN_ERR_CUDA;
  }
  gc.mu.unlock();
  return rc;
}

extern "C" int gn_graph_capture_launch(const void* key, void* stream) {
  return graph_capture_finish(key, as_stream(stream), true);
}

extern "C" int gn_graph_capture_abort(void) {
  return graph_capture_finish(nullptr, nullptr, false);
}
