// Segmented per-event kNN graph construction and graph-layout conversions.
//
// Replaces torch_cluster.knn / torch_geometric.nn.knn_graph as reached from the
// reference at data_representation/graphs/edges/edges.py:78-82 and
// models/components/layers.py:70-74.  See include/graphnet_b200.h for the contract.
//
// Kernel shape (HBM-bound integer/float work, no tensor cores): one CTA per event.
// The event's coordinates are staged once in shared memory as SoA; every thread owns
// one query at a time and scans the candidates in index order (shared-memory
// broadcast reads), keeping the best k+1 in a register-resident sorted list.  The
// (distance^2, index) total order falls out of the scan order plus a strict '<'
// insertion.  Distances use __fsub_rn/__fmul_rn/__fadd_rn so nvcc cannot contract
// them into FMAs: the rounding is that of the CPU reference.
#include <atomic>

#include "common.cuh"

namespace gn {

// ---------------------------------------------------------------- batch -> ptr
__global__ void batch_to_ptr_kernel(const int64_t* __restrict__ batch, int64_t N, int64_t B,
                                    int32_t* __restrict__ ptr, int32_t* __restrict__ flags) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i > N) return;
  int64_t cur = (i < N) ? batch[i] : B;
  int64_t prev = (i > 0) ? batch[i - 1] : -1;
  if (i < N && (cur < 0 || cur >= B)) {
    atomicOr(flags, GN_FLAG_INDEX_OUT_OF_RANGE);
    return;
  }
  if (i > 0 && (prev < 0 || prev >= B)) return;
  if (cur < prev) {
    atomicOr(flags, GN_FLAG_UNSORTED_BATCH);
    return;
  }
  for (int64_t e = prev + 1; e <= cur; ++e) ptr[e] = (int32_t)i;
}

// ------------------------------------------------------------------------ kNN
constexpr int KNN_MAX_DIMS = 16;
constexpr int KNN_THREADS = 128;
constexpr int KNN_SMEM_FLOATS = 3072;  // 12 KB of staged coordinates per CTA
constexpr int KNN_QDEPTH = 12;         // pending candidates per thread between flushes (queued scan)
constexpr int KNN_QUEUED_MIN_N = 160;  // events at least this large take the queued scan
constexpr int KNN_MAX_SPLIT = 16;      // CTAs per event (query blocks are dealt round robin)
constexpr uint32_t KNN_EMPTY = 0xffffffffu;

struct KnnParams {
  const float* x;
  int64_t ldx;
  int cols[KNN_MAX_DIMS];
  int dims;
  const int32_t* ptr;
  int kk;    // neighbours kept before dropping self: k + (loop ? 0 : 1)
  int loop;
  int32_t* nbr;
  int32_t stride;
  int32_t* deg;
  int chunk;  // candidates staged per pass (multiple of 32)
  int group;  // events per CTA (small events), 1: one event per grid row
  int64_t B;
};

// Sorted list of the CAP best (distance, index) pairs seen so far.  Distances are kept as the
// bit pattern of the (non-negative) float: unsigned order == float order, an empty slot
// (0xffffffff) ranks after everything, a NaN distance after every real one.
template <int CAP>
struct TopList {
  uint32_t d[CAP];
  int idx[CAP];
  __device__ __forceinline__ void init() {
#pragma unroll
    for (int s = 0; s < CAP; ++s) { d[s] = KNN_EMPTY; idx[s] = -1; }
  }
  __device__ __forceinline__ uint32_t worst() const { return d[CAP - 1]; }
  // Candidates arrive in ascending index order and an equal distance never displaces an
  // earlier entry, so ties rank by index.  Every slot is updated independently of the others
  // (no carried value): slot s keeps its entry if that entry is <= the candidate, otherwise it
  // takes the candidate (if its left neighbour stays) or the left neighbour's entry.  A
  // candidate that is not better than the worst entry changes nothing.
  __device__ __forceinline__ void push(uint32_t cd, int ci) {
    bool keep[CAP];
#pragma unroll
    for (int s = 0; s < CAP; ++s) keep[s] = d[s] <= cd;
#pragma unroll
    for (int s = CAP - 1; s > 0; --s) {
      const uint32_t nd = keep[s - 1] ? cd : d[s - 1];
      const int ni = keep[s - 1] ? ci : idx[s - 1];
      d[s] = keep[s] ? d[s] : nd;
      idx[s] = keep[s] ? idx[s] : ni;
    }
    d[0] = keep[0] ? d[0] : cd;
    idx[0] = keep[0] ? idx[0] : ci;
  }
};

template <int DIMS>
__device__ __forceinline__ uint32_t knn_dist_bits(const float* qc, const float* s_coord, int chunk,
                                                  int c, int dims) {
  float d = 0.f;
#pragma unroll
  for (int t = 0; t < (DIMS > 0 ? DIMS : KNN_MAX_DIMS); ++t) {
    if (t < dims) {
      const float diff = __fsub_rn(qc[t], s_coord[t * chunk + c]);
      d = __fadd_rn(d, __fmul_rn(diff, diff));
    }
  }
  return __float_as_uint(d);
}

template <int CAP>
__device__ __forceinline__ void knn_write_row(const KnnParams& p, const TopList<CAP>& top, int lo,
                                              int q) {
  int32_t* row = p.nbr + (int64_t)(lo + q) * p.stride;
  int cnt = 0;
#pragma unroll
  for (int s = 0; s < CAP; ++s) {
    if (s < p.kk) {
      const int j = top.idx[s];
      if (j >= 0 && (p.loop || j != q)) row[cnt++] = lo + j;
    }
  }
  p.deg[lo + q] = cnt;
  for (int r = cnt; r < p.stride; ++r) row[r] = -1;
}

// grid = (events, split).  Three shapes of the same scan, chosen per event (uniform per CTA):
//  * n <= 64: two threads per query, each scans half of the candidates; the upper half's list
//    is then pushed, in order, into the lower half's (all its indices are larger, so the
//    (distance, index) order is the same as one scan would give);
//  * n < KNN_QUEUED_MIN_N: one thread per query, a candidate better than the current worst is
//    inserted at once;
//  * larger events: a passing candidate is appended to a per-thread queue in shared memory and
//    the warp inserts together when a queue fills - by then most lanes have work, where an
//    immediate insertion would run the whole warp through the insertion code for the one or
//    two lanes that need it.  The threshold is the worst entry as of the last flush, so a few
//    candidates are queued that no longer qualify; push() ignores them.
//    The query blocks of a large event are dealt to the `split` CTAs of its grid row.
template <int CAP, int DIMS>  // DIMS == 0: runtime p.dims
__device__ __forceinline__ void knn_one_event(const KnnParams& p, int e, int yb, int ny,
                                              float* s_coord) {
  const int dims = DIMS > 0 ? DIMS : p.dims;
  const int lo = p.ptr[e];
  const int n = p.ptr[e + 1] - lo;
  if (n <= 0) return;
  const int chunk = p.chunk;
  const int tid = threadIdx.x;
  constexpr int QD = DIMS > 0 ? DIMS : KNN_MAX_DIMS;

  if (n <= KNN_THREADS / 2) {
    if (yb != 0) return;
    __syncthreads();  // a previous event of this CTA may still be read from the staging buffer
    for (int t = 0; t < dims; ++t)
      for (int c = tid; c < n; c += KNN_THREADS)
        s_coord[t * chunk + c] = p.x[(int64_t)(lo + c) * p.ldx + p.cols[t]];
    __syncthreads();
    const int q = tid >> 1, h = tid & 1;
    const bool active = q < n;
    float qc[QD];
#pragma unroll
    for (int t = 0; t < QD; ++t) qc[t] = (active && t < dims) ? s_coord[t * chunk + q] : 0.f;
    TopList<CAP> top;
    top.init();
    const int mid = (n + 1) >> 1;
    const int c0 = h ? mid : 0, c1 = active ? (h ? n : mid) : 0;
    for (int c = c0; c < c1; ++c) {
      const uint32_t cd = knn_dist_bits<DIMS>(qc, s_coord, chunk, c, dims);
      if (cd < top.worst()) top.push(cd, c);
    }
#pragma unroll
    for (int s = 0; s < CAP; ++s) {
      const uint32_t od = __shfl_down_sync(0xffffffffu, top.d[s], 1);
      const int oi = __shfl_down_sync(0xffffffffu, top.idx[s], 1);
      if (h == 0) top.push(od, oi);
    }
    if (active && h == 0) knn_write_row<CAP>(p, top, lo, q);
    return;
  }

  uint2* const queue = reinterpret_cast<uint2*>(s_coord + dims * chunk);  // [KNN_QDEPTH][threads]
  const bool queued = n >= KNN_QUEUED_MIN_N;
  const int qfirst = yb * KNN_THREADS;
  for (int qbase = qfirst; qbase < n; qbase += ny * KNN_THREADS) {
    const int q = qbase + tid;
    const bool active = q < n;
    float qc[QD];
#pragma unroll
    for (int t = 0; t < QD; ++t)
      qc[t] = (active && t < dims) ? p.x[(int64_t)(lo + q) * p.ldx + p.cols[t]] : 0.f;
    TopList<CAP> top;
    top.init();
    int pending = 0;
    uint32_t thr = active ? KNN_EMPTY : 0u;  // an idle lane queues nothing

    for (int cbase = 0; cbase < n; cbase += chunk) {
      const int cn = min(chunk, n - cbase);
      if (n > chunk || qbase == qfirst) {
        __syncthreads();
        for (int t = 0; t < dims; ++t)
          for (int c = tid; c < cn; c += KNN_THREADS)
            s_coord[t * chunk + c] = p.x[(int64_t)(lo + cbase + c) * p.ldx + p.cols[t]];
        __syncthreads();
      }
      if (!queued) {
        if (active) {
          for (int c = 0; c < cn; ++c) {
            const uint32_t cd = knn_dist_bits<DIMS>(qc, s_coord, chunk, c, dims);
            if (cd < top.worst()) top.push(cd, cbase + c);
          }
        }
        continue;
      }
      for (int c4 = 0; c4 < cn; c4 += 4) {
        // four candidates per step (the staged columns are 16-byte aligned at c4)
        float dd[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
        for (int t = 0; t < QD; ++t) {
          if (t < dims) {
            const float4 cc = *reinterpret_cast<const float4*>(s_coord + t * chunk + c4);
            const float v[4] = {cc.x, cc.y, cc.z, cc.w};
#pragma unroll
            for (int u = 0; u < 4; ++u) {
              const float diff = __fsub_rn(qc[t], v[u]);
              dd[u] = __fadd_rn(dd[u], __fmul_rn(diff, diff));
            }
          }
        }
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          const uint32_t cd = __float_as_uint(dd[u]);
          if (c4 + u < cn && cd < thr) {
            queue[pending * KNN_THREADS + tid] = make_uint2(cd, (uint32_t)(cbase + c4 + u));
            ++pending;
          }
        }
        if (__any_sync(0xffffffffu, pending > KNN_QDEPTH - 4)) {
          for (int r = 0; __any_sync(0xffffffffu, r < pending); ++r) {
            if (r < pending) {
              const uint2 v = queue[r * KNN_THREADS + tid];
              top.push(v.x, (int)v.y);
            }
          }
          pending = 0;
          thr = active ? top.worst() : 0u;
        }
      }
    }
    if (queued) {
      for (int r = 0; __any_sync(0xffffffffu, r < pending); ++r) {
        if (r < pending) {
          const uint2 v = queue[r * KNN_THREADS + tid];
          top.push(v.x, (int)v.y);
        }
      }
    }
    if (active) knn_write_row<CAP>(p, top, lo, q);
  }
}

// grid = (event groups, split).  p.group events per CTA: with small events (mean below 32 pulses)
// one event cannot occupy a CTA, so several consecutive events - their rows are contiguous in x -
// are staged together and every pair of threads takes one query of the group, scanning only its
// own event's candidates.  A group that does not fit that shape (an event above 64 pulses, or
// more rows than the staging buffer) is walked event by event.
template <int CAP, int DIMS>
__global__ void __launch_bounds__(KNN_THREADS) knn_kernel(KnnParams p) {
  extern __shared__ __align__(16) float s_coord[];  // [dims][chunk], then the queues
  if (p.group <= 1) {
    knn_one_event<CAP, DIMS>(p, blockIdx.x, blockIdx.y, gridDim.y, s_coord);
    return;
  }
  const int dims = DIMS > 0 ? DIMS : p.dims;
  constexpr int QD = DIMS > 0 ? DIMS : KNN_MAX_DIMS;
  const int chunk = p.chunk;
  const int tid = threadIdx.x;
  const int e0 = blockIdx.x * p.group;
  const int e1 = min(e0 + p.group, (int)p.B);
  const int lo0 = p.ptr[e0];
  const int total = p.ptr[e1] - lo0;
  bool packed = total <= chunk;
  for (int e = e0; e < e1; ++e) packed = packed && (p.ptr[e + 1] - p.ptr[e] <= KNN_THREADS / 2);
  if (!packed) {
    for (int e = e0; e < e1; ++e) knn_one_event<CAP, DIMS>(p, e, blockIdx.y, gridDim.y, s_coord);
    return;
  }
  if (blockIdx.y != 0 || total <= 0) return;
  for (int t = 0; t < dims; ++t)
    for (int c = tid; c < total; c += KNN_THREADS)
      s_coord[t * chunk + c] = p.x[(int64_t)(lo0 + c) * p.ldx + p.cols[t]];
  __syncthreads();
  for (int qbase = 0; qbase < total; qbase += KNN_THREADS / 2) {
    const int q = qbase + (tid >> 1), h = tid & 1;   // q: row of the group
    const bool active = q < total;
    int ev = e0;                                      // the event of row q
    if (active)
      while (p.ptr[ev + 1] - lo0 <= q) ++ev;
    const int lo = active ? p.ptr[ev] - lo0 : 0;      // first row of that event in the group
    const int n = active ? p.ptr[ev + 1] - lo0 - lo : 0;
    float qc[QD];
#pragma unroll
    for (int t = 0; t < QD; ++t) qc[t] = (active && t < dims) ? s_coord[t * chunk + q] : 0.f;
    TopList<CAP> top;
    top.init();
    const int mid = (n + 1) >> 1;
    const int c0 = h ? mid : 0, c1 = h ? n : mid;     // candidate indices inside the event
    for (int c = c0; c < c1; ++c) {
      const uint32_t cd = knn_dist_bits<DIMS>(qc, s_coord, chunk, lo + c, dims);
      if (cd < top.worst()) top.push(cd, c);
    }
#pragma unroll
    for (int s = 0; s < CAP; ++s) {
      const uint32_t od = __shfl_down_sync(0xffffffffu, top.d[s], 1);
      const int oi = __shfl_down_sync(0xffffffffu, top.idx[s], 1);
      if (h == 0) top.push(od, oi);
    }
    if (active && h == 0) knn_write_row<CAP>(p, top, lo0 + lo, q - lo);
  }
}

// Runtime-capacity variant for large k (k+1 > 33): the list lives in local memory.
constexpr int KNN_MAX_KK = 128;
template <int DIMS>
__global__ void __launch_bounds__(KNN_THREADS) knn_kernel_bigk(KnnParams p) {
  extern __shared__ float s_coord[];
  const int dims = DIMS > 0 ? DIMS : p.dims;
  const int e = blockIdx.x;
  const int lo = p.ptr[e];
  const int n = p.ptr[e + 1] - lo;
  if (n <= 0) return;
  const int chunk = p.chunk;
  const int tid = threadIdx.x;
  const int cap = p.kk;
  for (int qbase = 0; qbase < n; qbase += KNN_THREADS) {
    const int q = qbase + tid;
    const bool active = q < n;
    float qc[KNN_MAX_DIMS];
    for (int t = 0; t < dims; ++t)
      qc[t] = active ? p.x[(int64_t)(lo + q) * p.ldx + p.cols[t]] : 0.f;
    float bd[KNN_MAX_KK];
    int bi[KNN_MAX_KK];
    int m = 0;
    for (int cbase = 0; cbase < n; cbase += chunk) {
      const int cn = min(chunk, n - cbase);
      if (n > chunk || qbase == 0) {
        __syncthreads();
        for (int t = 0; t < dims; ++t)
          for (int c = tid; c < cn; c += KNN_THREADS)
            s_coord[t * chunk + c] = p.x[(int64_t)(lo + cbase + c) * p.ldx + p.cols[t]];
        __syncthreads();
      }
      if (active) {
        for (int c = 0; c < cn; ++c) {
          float d = 0.f;
          for (int t = 0; t < dims; ++t) {
            float diff = __fsub_rn(qc[t], s_coord[t * chunk + c]);
            d = __fadd_rn(d, __fmul_rn(diff, diff));
          }
          if (m == cap && !(d < bd[cap - 1])) continue;
          int s = (m < cap) ? m : cap - 1;  // slot that is freed / appended
          while (s > 0 && d < bd[s - 1]) { bd[s] = bd[s - 1]; bi[s] = bi[s - 1]; --s; }
          bd[s] = d;
          bi[s] = cbase + c;
          if (m < cap) ++m;
        }
      }
    }
    if (active) {
      int32_t* row = p.nbr + (int64_t)(lo + q) * p.stride;
      int cnt = 0;
      for (int s = 0; s < m; ++s) {
        int j = bi[s];
        if (p.loop || j != q) row[cnt++] = lo + j;
      }
      p.deg[lo + q] = cnt;
      for (int r = cnt; r < p.stride; ++r) row[r] = -1;
    }
  }
}

static size_t knn_smem_bytes(const KnnParams& p) {
  return (size_t)p.dims * p.chunk * sizeof(float) + (size_t)KNN_QDEPTH * KNN_THREADS * sizeof(uint2);
}

template <int CAP>
static int launch_knn(const KnnParams& p, int64_t N, int64_t B, cudaStream_t st) {
  const size_t smem = knn_smem_bytes(p);
  // CTAs per event from the mean event size (the host knows N and B, not the largest event)
  int64_t split = ceil_div(ceil_div(N, B), (int64_t)KNN_THREADS);
  split = split < 1 ? 1 : (split > KNN_MAX_SPLIT ? KNN_MAX_SPLIT : split);
  dim3 grid((unsigned)ceil_div(B, (int64_t)p.group), (unsigned)split), block(KNN_THREADS);
  switch (p.dims) {
    case 1: knn_kernel<CAP, 1><<<grid, block, smem, st>>>(p); break;
    case 2: knn_kernel<CAP, 2><<<grid, block, smem, st>>>(p); break;
    case 3: knn_kernel<CAP, 3><<<grid, block, smem, st>>>(p); break;
    case 4: knn_kernel<CAP, 4><<<grid, block, smem, st>>>(p); break;
    default: knn_kernel<CAP, 0><<<grid, block, smem, st>>>(p); break;
  }
  return GN_OK;
}

// ------------------------------------------------------- exclusive scan (int32)
constexpr int SCAN_THREADS = 256;
constexpr int SCAN_ITEMS = 4;
constexpr int SCAN_TILE = SCAN_THREADS * SCAN_ITEMS;  // 1024 values per block

__device__ __forceinline__ int64_t block_exclusive_scan(int64_t v, int64_t* total) {
  // exclusive scan of one value per thread over a SCAN_THREADS block
  __shared__ int64_t warp_sums[SCAN_THREADS / 32];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  int64_t incl = v;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    int64_t t = __shfl_up_sync(0xffffffffu, incl, o);
    if (lane >= o) incl += t;
  }
  if (lane == 31) warp_sums[warp] = incl;
  __syncthreads();
  if (warp == 0) {
    int64_t w = (lane < SCAN_THREADS / 32) ? warp_sums[lane] : 0;
    int64_t wi = w;
#pragma unroll
    for (int o = 1; o < SCAN_THREADS / 32; o <<= 1) {
      int64_t t = __shfl_up_sync(0xffffffffu, wi, o);
      if (lane >= o) wi += t;
    }
    if (lane < SCAN_THREADS / 32) warp_sums[lane] = wi - w;  // exclusive warp offsets
    if (lane == SCAN_THREADS / 32 - 1) *total = wi;
  }
  __syncthreads();
  int64_t r = warp_sums[warp] + incl - v;
  __syncthreads();
  return r;
}

__global__ void __launch_bounds__(SCAN_THREADS) scan_block_sums_kernel(
    const int32_t* __restrict__ in, int64_t n, int64_t* __restrict__ block_sums) {
  __shared__ int64_t total;
  int64_t base = (int64_t)blockIdx.x * SCAN_TILE + threadIdx.x * SCAN_ITEMS;
  int64_t s = 0;
#pragma unroll
  for (int t = 0; t < SCAN_ITEMS; ++t)
    if (base + t < n) s += in[base + t];
  block_exclusive_scan(s, &total);
  if (threadIdx.x == 0) block_sums[blockIdx.x] = total;
}

// one block turns the per-block sums into exclusive block offsets, in place
__global__ void __launch_bounds__(SCAN_THREADS) scan_offsets_kernel(int64_t* __restrict__ block_sums,
                                                                   int64_t nblocks) {
  __shared__ int64_t total;
  int64_t carry = 0;
  for (int64_t base = 0; base < nblocks; base += SCAN_THREADS) {
    int64_t i = base + threadIdx.x;
    int64_t v = (i < nblocks) ? block_sums[i] : 0;
    int64_t ex = block_exclusive_scan(v, &total);
    if (i < nblocks) block_sums[i] = carry + ex;
    carry += total;
    __syncthreads();
  }
}

__global__ void __launch_bounds__(SCAN_THREADS) scan_apply_kernel(
    const int32_t* __restrict__ in, int64_t n, const int64_t* __restrict__ block_offsets,
    int64_t* __restrict__ out) {
  __shared__ int64_t total;
  int64_t base = (int64_t)blockIdx.x * SCAN_TILE + threadIdx.x * SCAN_ITEMS;
  int32_t v[SCAN_ITEMS];
  int64_t s = 0;
#pragma unroll
  for (int t = 0; t < SCAN_ITEMS; ++t) {
    v[t] = (base + t < n) ? in[base + t] : 0;
    s += v[t];
  }
  int64_t ex = block_exclusive_scan(s, &total) + block_offsets[blockIdx.x];
#pragma unroll
  for (int t = 0; t < SCAN_ITEMS; ++t) {
    if (base + t < n) out[base + t] = ex;
    ex += v[t];
  }
  // out[n] = grand total, written by the thread that owns the last element
  if (base <= n - 1 && n - 1 < base + SCAN_ITEMS) out[n] = ex;
}

__global__ void ell_to_coo_kernel(const int32_t* __restrict__ nbr, int32_t stride,
                                  const int32_t* __restrict__ deg,
                                  const int64_t* __restrict__ rowptr, int64_t N, int64_t E,
                                  int64_t* __restrict__ ei) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= N) return;
  int64_t o = rowptr[i];
  int d = deg[i];
  for (int r = 0; r < d; ++r) {
    ei[o + r] = nbr[i * stride + r];
    ei[E + o + r] = i;
  }
}

// ------------------------------------------------------------------ COO -> CSR
__global__ void coo_to_csr_kernel(const int64_t* __restrict__ ei, int64_t E, int64_t N,
                                  int32_t* __restrict__ rowptr, int32_t* __restrict__ col,
                                  int32_t* __restrict__ flags) {
  int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (e > E) return;
  const int64_t* src = ei;
  const int64_t* dst = ei + E;
  int64_t cur = (e < E) ? dst[e] : N;
  int64_t prev = (e > 0) ? dst[e - 1] : -1;
  if (e < E) {
    int64_t s = src[e];
    if (s < 0 || s >= N || cur < 0 || cur >= N) {
      atomicOr(flags, GN_FLAG_INDEX_OUT_OF_RANGE);
      col[e] = 0;
      return;
    }
    col[e] = (int32_t)s;
  }
  if (e > 0 && (prev < 0 || prev >= N)) return;
  if (cur < prev) {
    atomicOr(flags, GN_FLAG_UNSORTED_TARGETS);
    return;
  }
  for (int64_t v = prev + 1; v <= cur; ++v) rowptr[v] = (int32_t)e;
}

__global__ void max_degree_kernel(const int32_t* __restrict__ rowptr, int64_t N,
                                  int32_t* __restrict__ max_degree) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  int d = (i < N) ? rowptr[i + 1] - rowptr[i] : 0;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) d = max(d, __shfl_xor_sync(0xffffffffu, d, o));
  if ((threadIdx.x & 31) == 0 && d > 0) atomicMax(max_degree, d);
}

}  // namespace gn

using namespace gn;

extern "C" int gn_batch_to_ptr(const int64_t* batch, int64_t N, int64_t B, int32_t* ptr,
                               int32_t* flags, void* stream) {
  GN_CHECK_ARG(N >= 0 && B >= 0 && ptr && flags, "gn_batch_to_ptr: bad arguments");
  GN_CHECK_ARG(N == 0 || batch, "gn_batch_to_ptr: batch is null");
  GN_CHECK_ARG(N < INT32_MAX, "gn_batch_to_ptr: N must fit int32");
  int threads = 256;
  batch_to_ptr_kernel<<<(unsigned)ceil_div(N + 1, threads), threads, 0, as_stream(stream)>>>(
      batch, N, B, ptr, flags);
  GN_LAUNCH_CHECK();
  return GN_OK;
}

extern "C" int gn_knn_graph(const float* x, int64_t ldx, const int32_t* cols, int dims,
                            const int32_t* ptr, int64_t N, int64_t B, int k, int loop,
                            int32_t* nbr, int32_t stride, int32_t* deg, void* stream) {
  GN_CHECK_ARG(dims >= 1 && dims <= KNN_MAX_DIMS, "gn_knn_graph: dims=%d not in [1,%d]", dims,
               KNN_MAX_DIMS);
  GN_CHECK_ARG(k >= 1, "gn_knn_graph: k=%d must be >= 1", k);
  const int kk = k + (loop ? 0 : 1);
  GN_CHECK_ARG(kk <= KNN_MAX_KK, "gn_knn_graph: k=%d too large (max %d)", k, KNN_MAX_KK - 1);
  GN_CHECK_ARG(stride >= kk, "gn_knn_graph: stride=%d < k+1=%d", stride, kk);
  GN_CHECK_ARG(N >= 0 && B >= 0 && N < INT32_MAX, "gn_knn_graph: bad sizes");
  if (N == 0 || B == 0) return GN_OK;
  GN_CHECK_ARG(x && cols && ptr && nbr && deg, "gn_knn_graph: null pointer");
  KnnParams p;
  p.x = x; p.ldx = ldx; p.dims = dims; p.ptr = ptr; p.kk = kk; p.loop = loop;
  p.nbr = nbr; p.stride = stride; p.deg = deg;
  for (int t = 0; t < KNN_MAX_DIMS; ++t) p.cols[t] = t < dims ? cols[t] : 0;
  for (int t = 0; t < dims; ++t)
    GN_CHECK_ARG(cols[t] >= 0 && cols[t] < ldx, "gn_knn_graph: column %d out of range", cols[t]);
  p.chunk = (KNN_SMEM_FLOATS / dims) & ~31;
  // events per CTA: as many as fill the 64 query slots of a CTA at the mean event size
  const int64_t mean_n = ceil_div(N, B);
  p.group = mean_n >= 32 ? 1 : (int)(64 / (mean_n > 8 ? mean_n : 8));
  p.B = B;
  cudaStream_t st = as_stream(stream);
  if (kk <= 9) launch_knn<9>(p, N, B, st);
  else if (kk <= 17) launch_knn<17>(p, N, B, st);
  else if (kk <= 33) launch_knn<33>(p, N, B, st);
  else {
    size_t smem = (size_t)dims * p.chunk * sizeof(float);
    knn_kernel_bigk<0><<<(unsigned)B, KNN_THREADS, smem, st>>>(p);
  }
  GN_LAUNCH_CHECK();
  return GN_OK;
}

extern "C" size_t gn_scan_scratch_bytes(int64_t n) {
  return (size_t)(ceil_div(n > 0 ? n : 1, SCAN_TILE)) * sizeof(int64_t);
}

namespace gn {
int exclusive_scan_i32(const int32_t* in, int64_t n, int64_t* out, int64_t* block_sums,
                       cudaStream_t st) {
  int64_t nblocks = ceil_div(n, SCAN_TILE);
  scan_block_sums_kernel<<<(unsigned)nblocks, SCAN_THREADS, 0, st>>>(in, n, block_sums);
  GN_LAUNCH_CHECK();
  scan_offsets_kernel<<<1, SCAN_THREADS, 0, st>>>(block_sums, nblocks);
  GN_LAUNCH_CHECK();
  scan_apply_kernel<<<(unsigned)nblocks, SCAN_THREADS, 0, st>>>(in, n, block_sums, out);
  GN_LAUNCH_CHECK();
  return GN_OK;
}
}  // namespace gn

extern "C" int gn_ell_row_offsets(const int32_t* deg, int64_t N, int64_t* rowptr, void* scratch,
                                  size_t scratch_bytes, void* stream) {
  GN_CHECK_ARG(N >= 0 && rowptr, "gn_ell_row_offsets: bad arguments");
  cudaStream_t st = as_stream(stream);
  if (N == 0) {
    GN_CUDA(cudaMemsetAsync(rowptr, 0, sizeof(int64_t), st));
    return GN_OK;
  }
  GN_CHECK_ARG(deg && scratch, "gn_ell_row_offsets: null pointer");
  GN_CHECK_ARG(scratch_bytes >= gn_scan_scratch_bytes(N),
               "gn_ell_row_offsets: scratch too small (%zu < %zu)", scratch_bytes,
               gn_scan_scratch_bytes(N));
  return exclusive_scan_i32(deg, N, rowptr, reinterpret_cast<int64_t*>(scratch), st);
}

// ------------------------------------------------ row offsets without caller scratch
// Some hosts of this library are not allowed to allocate device memory once they are running:
// an online filter process at the detector is started with a fixed GPU memory budget and must
// not grow it per event, and a process that has handed the whole device to another framework's
// memory pool would see cudaMalloc fail.  For them the per-block partial sums of the scan live
// in this fixed array instead of in caller-provided scratch.  It is part of the library image:
// every GPU that loads the library holds these 32 KiB until the process exits.
constexpr int64_t STATIC_SCAN_MAX_N = 4194304;  // nodes per call (4096 blocks of 1024)
constexpr int64_t STATIC_SCAN_MAX_BLOCKS = STATIC_SCAN_MAX_N / gn::SCAN_TILE;
static __device__ __align__(256) int64_t g_static_scan_sums[STATIC_SCAN_MAX_BLOCKS];

// `g_static_scan_sums` is only a host-side handle in host code; the device address has to be
// asked for, once per device, with that device current.
static int static_scan_scratch(int64_t** out) {
  constexpr int MAX_DEVICES = 64;
  // several host threads may make their first call at once: the slots are atomics (all writers
  // of a slot store the same address)
  static std::atomic<int64_t*> addr[MAX_DEVICES];
  int dev = 0;
  GN_CUDA(cudaGetDevice(&dev));   // the CURRENT device: it must be the one `stream` belongs to
  GN_CHECK_ARG(dev >= 0 && dev < MAX_DEVICES, "device ordinal %d not supported", dev);
  int64_t* a = addr[dev].load(std::memory_order_acquire);
  if (a == nullptr) {
    void* p = nullptr;
    GN_CUDA(cudaGetSymbolAddress(&p, g_static_scan_sums));
    a = reinterpret_cast<int64_t*>(p);
    addr[dev].store(a, std::memory_order_release);
  }
  *out = a;
  return GN_OK;
}

extern "C" int gn_ell_row_offsets_noalloc(const int32_t* deg, int64_t N, int64_t* rowptr,
                                          void* stream) {
  GN_CHECK_ARG(N >= 0 && rowptr, "gn_ell_row_offsets_noalloc: bad arguments");
  GN_CHECK_ARG(N <= STATIC_SCAN_MAX_N,
               "gn_ell_row_offsets_noalloc: N=%lld exceeds the built-in scratch (max %lld nodes "
               "per call); use gn_ell_row_offsets with caller-provided scratch",
               (long long)N, (long long)STATIC_SCAN_MAX_N);
  cudaStream_t st = as_stream(stream);
  if (N == 0) {
    GN_CUDA(cudaMemsetAsync(rowptr, 0, sizeof(int64_t), st));
    return GN_OK;
  }
  GN_CHECK_ARG(deg, "gn_ell_row_offsets_noalloc: null pointer");
  int64_t* sums = nullptr;
  int rc = static_scan_scratch(&sums);
  if (rc != GN_OK) return rc;
  // every slot of `sums` that the apply pass reads is first written by the block-sum pass of
  // this same call, so nothing has to be cleared between calls
  return exclusive_scan_i32(deg, N, rowptr, sums, st);
}

extern "C" int gn_ell_to_coo(const int32_t* nbr, int32_t stride, const int32_t* deg,
                             const int64_t* rowptr, int64_t N, int64_t E, int64_t* edge_index,
                             void* stream) {
  GN_CHECK_ARG(N >= 0 && E >= 0, "gn_ell_to_coo: bad sizes");
  if (N == 0 || E == 0) return GN_OK;
  GN_CHECK_ARG(nbr && deg && rowptr && edge_index && stride > 0, "gn_ell_to_coo: null pointer");
  ell_to_coo_kernel<<<(unsigned)ceil_div(N, 256), 256, 0, as_stream(stream)>>>(
      nbr, stride, deg, rowptr, N, E, edge_index);
  GN_LAUNCH_CHECK();
  return GN_OK;
}

extern "C" int gn_coo_to_csr(const int64_t* edge_index, int64_t E, int64_t N, int32_t* rowptr,
                             int32_t* col, int32_t* flags, int32_t* max_degree, void* stream) {
  GN_CHECK_ARG(E >= 0 && N >= 0 && rowptr && flags, "gn_coo_to_csr: bad arguments");
  GN_CHECK_ARG(E < INT32_MAX && N < INT32_MAX, "gn_coo_to_csr: sizes must fit int32");
  GN_CHECK_ARG(E == 0 || (edge_index && col), "gn_coo_to_csr: null pointer");
  cudaStream_t st = as_stream(stream);
  coo_to_csr_kernel<<<(unsigned)ceil_div(E + 1, 256), 256, 0, st>>>(edge_index, E, N, rowptr, col,
                                                                    flags);
  GN_LAUNCH_CHECK();
  if (max_degree) {
    GN_CUDA(cudaMemsetAsync(max_degree, 0, sizeof(int32_t), st));
    if (N > 0) {
      max_degree_kernel<<<(unsigned)ceil_div(N, 256), 256, 0, st>>>(rowptr, N, max_degree);
      GN_LAUNCH_CHECK();
    }
  }
  return GN_OK;
}
