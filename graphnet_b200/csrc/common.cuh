// Shared helpers for the graphnet_b200 CUDA library (sm_100a only).
#pragma once

#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>

#include "../../include/graphnet_b200.h"

#if defined(__CUDA_ARCH__) && (__CUDA_ARCH__ < 1000)
#error "graphnet_b200 is written for sm_100a (B200) only"
#endif

namespace gn {

// ---- error plumbing ------------------------------------------------------ //
// Every extern "C" entry point returns GN_OK or a negative GN_ERR_* code and
// leaves a message readable through gn_last_error().  Nothing aborts.
void set_error(const char* fmt, ...);

#define GN_CHECK_ARG(cond, ...)                 \
  do {                                          \
    if (!(cond)) {                              \
      gn::set_error(__VA_ARGS__);               \
      return GN_ERR_INVALID_ARGUMENT;           \
    }                                           \
  } while (0)

#define GN_CUDA(call)                                                         \
  do {                                                                        \
    cudaError_t e__ = (call);                                                 \
    if (e__ != cudaSuccess) {                                                 \
      gn::set_error("%s failed at %s:%d: %s", #call, __FILE__, __LINE__,      \
                    cudaGetErrorString(e__));                                 \
      return GN_ERR_CUDA;                                                     \
    }                                                                         \
  } while (0)

// Every kernel launch of the library is followed by this macro; it also feeds
// gn_launch_count() (bench.py reports the launches inside its timed region).
extern unsigned long long g_launch_count;
#define GN_LAUNCH_CHECK()          \
  do {                             \
    ++gn::g_launch_count;          \
    GN_CUDA(cudaGetLastError());   \
  } while (0)

inline cudaStream_t as_stream(void* s) { return reinterpret_cast<cudaStream_t>(s); }

inline int64_t ceil_div(int64_t a, int64_t b) { return (a + b - 1) / b; }

// Rows with in-edges are described to the kernels by one of two layouts:
//  ELL: neighbours of node i are col[i*stride .. i*stride + deg[i])   (our kNN output)
//  CSR: neighbours of node i are col[rowptr[i] .. rowptr[i+1])        (converted user COO)
struct GraphView {
  const int32_t* col;
  const int32_t* deg;     // ELL only
  const int32_t* rowptr;  // CSR only
  int32_t stride;         // > 0: ELL, 0: CSR
};

__device__ __forceinline__ void graph_row(const GraphView& g, int64_t i, int64_t& start, int& count) {
  if (g.stride > 0) {
    start = i * (int64_t)g.stride;
    count = g.deg[i];
  } else {
    int32_t a = g.rowptr[i], b = g.rowptr[i + 1];
    start = a;
    count = b - a;
  }
}

inline int make_graph_view(const gn_graph* g, GraphView* v) {
  if (!g || !g->col) { set_error("graph: null"); return GN_ERR_INVALID_ARGUMENT; }
  v->col = g->col;
  v->deg = g->deg;
  v->rowptr = g->rowptr;
  v->stride = g->ell_stride;
  if (g->ell_stride > 0) {
    if (!g->deg) { set_error("graph: ELL layout needs deg"); return GN_ERR_INVALID_ARGUMENT; }
  } else {
    if (!g->rowptr) { set_error("graph: CSR layout needs rowptr"); return GN_ERR_INVALID_ARGUMENT; }
  }
  return GN_OK;
}

enum Act : int {
  ACT_NONE = GN_ACT_NONE, ACT_RELU = GN_ACT_RELU, ACT_GELU = GN_ACT_GELU,
  ACT_LEAKY_RELU = GN_ACT_LEAKY_RELU
};

template <int ACT>
__device__ __forceinline__ float activate(float v) {
  if (ACT == ACT_RELU) return fmaxf(v, 0.0f);
  if (ACT == ACT_LEAKY_RELU) return v > 0.0f ? v : 0.01f * v;
  if (ACT == ACT_GELU) return 0.5f * v * (1.0f + erff(v * 0.70710678118654752440f));
  return v;
}

}  // namespace gn
