// libptgs A3: balanced binary (exact median) split, level synchronous.
// Reference: balanced_binary_partition, scene/ptgs/density_partition.py:9-99.
//
// Every level keeps the points of each tree node contiguous in a ping-pong
// buffer of (x,y) pairs.  A node's split value is the rank-n/2 element of its
// axis coordinate, found with an MSD radix select on the order-preserving
// integer image of the coordinate: shared-memory privatised 2048-bin
// histograms (11/11/10 bits for f32, 11/11/11/11/10/10 for f64), a
// warp-shuffle prefix scan that picks the digit bucket holding rank k, and a
// partition pass that copies each point to the closed child rectangle(s) it
// falls in (points ON the split line go to both, density_partition.py:59-64
// with the closed mask of :22-23).
#include "common.cuh"

namespace ptgs {

constexpr int kSplitTile = 4096;   // entries per CTA (256 threads x 16)
constexpr int kBins = 2048;
constexpr int kMaxNodes = 4096;    // 2^12

template <typename T>
struct alignas(2 * sizeof(T)) XY {
  T x, y;
};

template <typename T>
struct Node {
  T minx, miny, maxx, maxy;
  int64_t start, count;
  int64_t k_rem;     // rank still to resolve inside the current prefix bucket
  int64_t less;      // # keys strictly below the split value
  int64_t eq;        // # keys equal to the split value
  int64_t child_start[2];
  typename KeyOf<T>::type prefix;
  uint32_t path;
  int32_t depth, active, axis;
};

template <typename T>
struct SplitState {
  Node<T>* nodes[2];
  int32_t* n_nodes;          // [2]
  int64_t* level_total;      // [2] entries in buffer
  unsigned long long* cursors;  // [kMaxNodes*2] child fill cursors
  uint32_t* hist;            // [max_level_nodes][kBins]
  // unsorted leaves
  double* lf_bounds; uint32_t* lf_path; int32_t* lf_depth; int64_t* lf_count;
  int32_t* lf_n;
  int32_t* status;
  int64_t capacity;
};

template <typename T> struct Digits;
template <> struct Digits<float> {
  static constexpr int passes = 3;
  __host__ __device__ static int shift(int p) { return p == 0 ? 21 : (p == 1 ? 10 : 0); }
  __host__ __device__ static int nbits(int p) { return p == 2 ? 10 : 11; }
};
template <> struct Digits<double> {
  static constexpr int passes = 6;
  __host__ __device__ static int shift(int p) {
    const int s[6] = {53, 42, 31, 20, 10, 0};
    return s[p];
  }
  __host__ __device__ static int nbits(int p) { return p >= 4 ? 10 : 11; }
};

template <typename T>
__global__ void __launch_bounds__(256) pack_xy_kernel(
    const T* __restrict__ x, const T* __restrict__ y,
    const int64_t* __restrict__ n_dev, XY<T>* __restrict__ out) {
  const int64_t n = *n_dev;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    XY<T> v;
    v.x = x[i]; v.y = y[i];
    out[i] = v;
  }
}

template <typename T>
__global__ void split_init_kernel(SplitState<T> s, const int64_t* n_dev,
                                  const double* root_bounds) {
  Node<T>& nd = s.nodes[0][0];
  nd.minx = (T)root_bounds[0]; nd.miny = (T)root_bounds[1];
  nd.maxx = (T)root_bounds[2]; nd.maxy = (T)root_bounds[3];
  nd.start = 0; nd.count = *n_dev; nd.path = 0; nd.depth = 0; nd.active = 0;
  s.n_nodes[0] = 1; s.n_nodes[1] = 0;
  s.level_total[0] = *n_dev; s.level_total[1] = 0;
  *s.lf_n = 0;
  *s.status = PTGS_OK;
}

// one thread per node of the level: leaf test + select initialisation.
template <typename T>
__global__ void __launch_bounds__(1024) classify_kernel(SplitState<T> s, int cur,
                                                        double leaf_limit, int max_depth) {
  const int nn = s.n_nodes[cur];
  __shared__ int s_active;
  if (threadIdx.x == 0) s_active = 0;
  __syncthreads();
  int mine_active = 0;
  for (int j = threadIdx.x; j < nn; j += blockDim.x) {
    Node<T>& nd = s.nodes[cur][j];
    if (!((double)nd.count <= leaf_limit || nd.depth >= max_depth)) ++mine_active;
    if ((double)nd.count <= leaf_limit || nd.depth >= max_depth) {
      int o = atomicAdd(s.lf_n, 1);
      s.lf_bounds[4 * o + 0] = (double)nd.minx; s.lf_bounds[4 * o + 1] = (double)nd.miny;
      s.lf_bounds[4 * o + 2] = (double)nd.maxx; s.lf_bounds[4 * o + 3] = (double)nd.maxy;
      s.lf_path[o] = nd.path; s.lf_depth[o] = nd.depth; s.lf_count[o] = nd.count;
      nd.active = 0;
    } else {
      double xr = (double)nd.maxx - (double)nd.minx;
      double yr = (double)nd.maxy - (double)nd.miny;
      nd.axis = (xr >= yr) ? 0 : 1;            // density_partition.py:45-50
      nd.k_rem = nd.count / 2;                 // :55
      nd.less = 0; nd.eq = 0; nd.prefix = 0;
      nd.active = 1;
    }
  }
  if (mine_active) atomicAdd(&s_active, mine_active);
  __syncthreads();
  if (threadIdx.x == 0) s.n_nodes[2] = s_active;
}

template <typename T>
__device__ __forceinline__ int find_node(const Node<T>* nodes, int nn, int64_t pos) {
  // last node with start <= pos
  int lo = 0, hi = nn - 1;
  while (lo < hi) {
    int mid = (lo + hi + 1) >> 1;
    if (nodes[mid].start <= pos) lo = mid; else hi = mid - 1;
  }
  return lo;
}

template <typename T>
__global__ void __launch_bounds__(256) hist_kernel(SplitState<T> s, int cur, int pass,
                                                   const XY<T>* __restrict__ buf) {
  using K = typename KeyOf<T>::type;
  __shared__ uint32_t sh[kBins];
  const int64_t total = s.level_total[cur];
  const int64_t t0 = (int64_t)blockIdx.x * kSplitTile;
  if (t0 >= total) return;
  const int64_t t1 = min(t0 + (int64_t)kSplitTile, total);
  const Node<T>* nodes = s.nodes[cur];
  const int nn = s.n_nodes[cur];
  const int shift = Digits<T>::shift(pass), nb = Digits<T>::nbits(pass);
  const int hi_shift = shift + nb;
  int j = find_node(nodes, nn, t0);
  for (; j < nn; ++j) {
    const Node<T>& nd = nodes[j];
    if (nd.start >= t1) break;
    if (!nd.active) continue;
    const int64_t a = max(t0, nd.start), b = min(t1, nd.start + nd.count);
    if (a >= b) continue;
    for (int i = threadIdx.x; i < kBins; i += 256) sh[i] = 0;
    __syncthreads();
    const int axis = nd.axis;
    const K prefix = nd.prefix;
    // 8 independent loads in flight per thread: the pass is a pure stream and was latency bound
    // with one (ncu: 23 long-scoreboard stalls per issue, 46 % of DRAM peak)
    constexpr int kU = 8;
    for (int64_t i0 = a; i0 < b; i0 += 256 * kU) {
      XY<T> v[kU];
#pragma unroll
      for (int q = 0; q < kU; ++q) {
        const int64_t i = i0 + q * 256 + threadIdx.x;
        if (i < b) v[q] = buf[i];
      }
#pragma unroll
      for (int q = 0; q < kU; ++q) {
        const int64_t i = i0 + q * 256 + threadIdx.x;
        if (i < b) {
          const K key = KeyOf<T>::encode(axis ? v[q].y : v[q].x);
          const bool match = (hi_shift >= KeyOf<T>::bits) ? true : ((key >> hi_shift) == (prefix >> hi_shift));
          if (match) atomicAdd(&sh[(uint32_t)((key >> shift) & ((K(1) << nb) - 1))], 1u);
        }
      }
    }
    __syncthreads();
    uint32_t* gh = s.hist + (size_t)j * kBins;
    for (int i = threadIdx.x; i < kBins; i += 256) {
      uint32_t c = sh[i];
      if (c) atomicAdd(gh + i, c);
    }
    __syncthreads();
  }
}

// one CTA per node: find the digit bucket that holds rank k_rem.
template <typename T>
__global__ void __launch_bounds__(256) pick_kernel(SplitState<T> s, int cur, int pass) {
  using K = typename KeyOf<T>::type;
  __shared__ long long sm[33];
  __shared__ int s_bin;
  __shared__ long long s_below;
  __shared__ uint32_t s_cnt;
  const int j = blockIdx.x;
  if (j >= s.n_nodes[cur]) return;
  Node<T>& nd = s.nodes[cur][j];
  if (!nd.active) return;
  uint32_t* gh = s.hist + (size_t)j * kBins;
  uint32_t c[8];
  long long sum = 0;
#pragma unroll
  for (int q = 0; q < 8; ++q) { c[q] = gh[threadIdx.x * 8 + q]; sum += c[q]; }
  long long tot;
  long long ex = block_exclusive_sum<long long, 256>(sum, sm, &tot);
  const long long k = nd.k_rem;
  if (k >= ex && k < ex + sum) {
    long long run = ex;
#pragma unroll
    for (int q = 0; q < 8; ++q) {
      if (k >= run && k < run + c[q]) { s_bin = threadIdx.x * 8 + q; s_below = run; s_cnt = c[q]; }
      run += c[q];
    }
  }
#pragma unroll
  for (int q = 0; q < 8; ++q) gh[threadIdx.x * 8 + q] = 0;   // ready for next pass
  __syncthreads();
  if (threadIdx.x == 0) {
    nd.less += s_below;
    nd.k_rem = k - s_below;
    nd.prefix |= (K)s_bin << Digits<T>::shift(pass);
    if (pass == Digits<T>::passes - 1) nd.eq = s_cnt;
  }
}

// single CTA: child segment layout + next level node table.
template <typename T>
__global__ void __launch_bounds__(1024) layout_kernel(SplitState<T> s, int cur) {
  __shared__ long long sm[33];
  __shared__ long long s_carry_sz;
  __shared__ int s_carry_rank;
  const int nn = s.n_nodes[cur];
  const int nxt = cur ^ 1;
  if (threadIdx.x == 0) { s_carry_sz = 0; s_carry_rank = 0; }
  __syncthreads();
  for (int base = 0; base < nn; base += 1024) {
    const int j = base + threadIdx.x;
    long long sz = 0, act = 0;
    Node<T>* nd = nullptr;
    if (j < nn) {
      nd = &s.nodes[cur][j];
      if (nd->active) { act = 1; sz = (nd->less + nd->eq) + (nd->count - nd->less); }
    }
    long long tot_sz, tot_act;
    long long ex_sz = block_exclusive_sum<long long, 1024>(sz, sm, &tot_sz);
    long long ex_act = block_exclusive_sum<long long, 1024>(act, sm, &tot_act);
    if (act) {
      const long long st = s_carry_sz + ex_sz;
      const int r = s_carry_rank + (int)ex_act;
      const long long nL = nd->less + nd->eq, nR = nd->count - nd->less;
      const T split = KeyOf<T>::decode(nd->prefix);
      nd->child_start[0] = st; nd->child_start[1] = st + nL;
      s.cursors[2 * j] = 0; s.cursors[2 * j + 1] = 0;
      Node<T> L, R;
      L = *nd; R = *nd;
      if (nd->axis == 0) { L.maxx = split; R.minx = split; } else { L.maxy = split; R.miny = split; }
      L.start = st; L.count = nL; L.path = (nd->path << 1); L.depth = nd->depth + 1; L.active = 0;
      R.start = st + nL; R.count = nR; R.path = (nd->path << 1) | 1u; R.depth = nd->depth + 1; R.active = 0;
      if (2 * r + 1 < kMaxNodes) { s.nodes[nxt][2 * r] = L; s.nodes[nxt][2 * r + 1] = R; }
    }
    __syncthreads();
    if (threadIdx.x == 0) { s_carry_sz += tot_sz; s_carry_rank += (int)tot_act; }
    __syncthreads();
  }
  if (threadIdx.x == 0) {
    long long total = s_carry_sz;
    int nnext = 2 * s_carry_rank;
    if (total > s.capacity || nnext > kMaxNodes) {
      *s.status = PTGS_E_CAPACITY;
      total = 0; nnext = 0;
      // nothing is copied: deactivate so the partition pass does no work
      for (int j = 0; j < nn; ++j) s.nodes[cur][j].active = 0;
    }
    s.n_nodes[nxt] = nnext;
    s.level_total[nxt] = total;
  }
}

template <typename T>
__global__ void __launch_bounds__(256) partition_kernel(SplitState<T> s, int cur,
                                                        const XY<T>* __restrict__ src,
                                                        XY<T>* __restrict__ dst) {
  __shared__ unsigned long long sm[33];
  __shared__ unsigned long long s_base[2];
  // the tile's left (then right) children are compacted here and written out as one coalesced
  // run each: a thread storing its own 16 elements one by one left the pass at 31 % of DRAM peak
  extern __shared__ __align__(16) unsigned char split_dyn[];     // kSplitTile * sizeof(XY<T>) bytes
  XY<T>* s_out = reinterpret_cast<XY<T>*>(split_dyn);
  const int64_t total = s.level_total[cur];
  const int64_t t0 = (int64_t)blockIdx.x * kSplitTile;
  if (t0 >= total) return;
  const int64_t t1 = min(t0 + (int64_t)kSplitTile, total);
  const Node<T>* nodes = s.nodes[cur];
  const int nn = s.n_nodes[cur];
  int j = find_node(nodes, nn, t0);
  for (; j < nn; ++j) {
    const Node<T>& nd = nodes[j];
    if (nd.start >= t1) break;
    if (!nd.active) continue;
    const int64_t a = max(t0, nd.start), b = min(t1, nd.start + nd.count);
    if (a >= b) continue;
    const int axis = nd.axis;
    const T split = KeyOf<T>::decode(nd.prefix);
    // each thread owns up to 16 strided elements
    XY<T> v[16];
    uint32_t mL = 0, mR = 0;
#pragma unroll
    for (int q = 0; q < 16; ++q) {
      int64_t i = a + q * 256 + threadIdx.x;
      if (i < b) {
        v[q] = src[i];
        T c = axis ? v[q].y : v[q].x;
        if (c <= split) mL |= 1u << q;
        if (c >= split) mR |= 1u << q;
      }
    }
    unsigned long long packed = (unsigned long long)__popc(mL) | ((unsigned long long)__popc(mR) << 32);
    unsigned long long tot;
    unsigned long long ex = block_exclusive_sum<unsigned long long, 256>(packed, sm, &tot);
    if (threadIdx.x == 0) {
      s_base[0] = atomicAdd(&s.cursors[2 * j], tot & 0xffffffffull);
      s_base[1] = atomicAdd(&s.cursors[2 * j + 1], tot >> 32);
    }
    const int nL = (int)(tot & 0xffffffffull), nR = (int)(tot >> 32);
    // left children
    {
      int o = (int)(ex & 0xffffffffull);
#pragma unroll
      for (int q = 0; q < 16; ++q)
        if (mL & (1u << q)) s_out[o++] = v[q];
    }
    __syncthreads();                         // s_out (left) and s_base complete
    {
      XY<T>* out = dst + nd.child_start[0] + (int64_t)s_base[0];
      for (int i = threadIdx.x; i < nL; i += 256) out[i] = s_out[i];
    }
    __syncthreads();                         // s_out is free again
    {
      int o = (int)(ex >> 32);
#pragma unroll
      for (int q = 0; q < 16; ++q)
        if (mR & (1u << q)) s_out[o++] = v[q];
    }
    __syncthreads();
    {
      XY<T>* out = dst + nd.child_start[1] + (int64_t)s_base[1];
      for (int i = threadIdx.x; i < nR; i += 256) out[i] = s_out[i];
    }
    __syncthreads();                         // before the next node of this tile reuses s_out / s_base
  }
}

// single CTA: order the leaves depth-first (left child first).
template <typename T>
__global__ void __launch_bounds__(1024) finalize_kernel(SplitState<T> s, int max_depth,
                                                        double* leaf_bounds, uint32_t* leaf_path,
                                                        int32_t* leaf_depth, int64_t* leaf_count,
                                                        int32_t* n_leaves, int32_t max_leaves) {
  const int n = *s.lf_n;
  if (threadIdx.x == 0) {
    *n_leaves = n;
    if (n > max_leaves) *s.status = PTGS_E_CAPACITY;
  }
  if (n > max_leaves) return;
  for (int i = threadIdx.x; i < n; i += blockDim.x) {
    const uint32_t ki = s.lf_path[i] << (max_depth - s.lf_depth[i]);
    int rank = 0;
    for (int j = 0; j < n; ++j) {
      const uint32_t kj = s.lf_path[j] << (max_depth - s.lf_depth[j]);
      rank += (kj < ki) ? 1 : 0;
    }
    for (int q = 0; q < 4; ++q) leaf_bounds[4 * rank + q] = s.lf_bounds[4 * i + q];
    leaf_path[rank] = s.lf_path[i];
    leaf_depth[rank] = s.lf_depth[i];
    leaf_count[rank] = s.lf_count[i];
  }
}

template <typename T>
static SplitState<T> carve_state(void* workspace, int64_t capacity, int max_depth,
                                 XY<T>** bufs, size_t* used) {
  Carver cv(workspace);
  SplitState<T> s;
  s.nodes[0] = cv.take<Node<T>>(kMaxNodes);
  s.nodes[1] = cv.take<Node<T>>(kMaxNodes);
  s.n_nodes = cv.take<int32_t>(4);          // [2] = nodes of the current level that still split
  s.level_total = cv.take<int64_t>(2);
  s.cursors = cv.take<unsigned long long>(2 * kMaxNodes);
  const int level_nodes = 1 << (max_depth > 0 ? max_depth - 1 : 0);
  s.hist = cv.take<uint32_t>((size_t)level_nodes * kBins);
  s.lf_bounds = cv.take<double>(4 * kMaxNodes);
  s.lf_path = cv.take<uint32_t>(kMaxNodes);
  s.lf_depth = cv.take<int32_t>(kMaxNodes);
  s.lf_count = cv.take<int64_t>(kMaxNodes);
  s.lf_n = cv.take<int32_t>(1);
  s.status = nullptr;
  s.capacity = capacity;
  bufs[0] = cv.take<XY<T>>(capacity);
  bufs[1] = cv.take<XY<T>>(capacity);
  *used = cv.used();
  return s;
}

static int64_t split_capacity(int64_t n_capacity, double factor) {
  if (factor < 1.0) factor = 1.0;
  return (int64_t)((double)n_capacity * factor) + 4096;
}

template <typename T>
static int split_impl(const void* x, const void* y, const int64_t* n_dev, int64_t n_capacity,
                      const double* root_bounds, double leaf_limit, int max_depth,
                      double capacity_factor, double* leaf_bounds, uint32_t* leaf_path,
                      int32_t* leaf_depth, int64_t* leaf_count, int32_t* n_leaves,
                      int32_t max_leaves, int32_t* status_dev, void* workspace,
                      cudaStream_t st) {
  const int64_t cap = split_capacity(n_capacity, capacity_factor);
  XY<T>* bufs[2];
  size_t used;
  SplitState<T> s = carve_state<T>(workspace, cap, max_depth, bufs, &used);
  s.status = status_dev;
  const int level_nodes = 1 << (max_depth > 0 ? max_depth - 1 : 0);
  PTGS_CUDA_TRY(cudaMemsetAsync(s.hist, 0, (size_t)level_nodes * kBins * 4, st));
  split_init_kernel<T><<<1, 1, 0, st>>>(s, n_dev, root_bounds);
  PTGS_LAUNCH_CHECK();
  int64_t gpack64 = (n_capacity + 2047) / 2048 + 1;
  int gpack = (int)(gpack64 < (int64_t)kNumSMs * 8 ? gpack64 : (int64_t)kNumSMs * 8);
  pack_xy_kernel<T><<<gpack, 256, 0, st>>>((const T*)x, (const T*)y, n_dev, bufs[0]);
  PTGS_LAUNCH_CHECK();
  const int tiles = (int)((cap + kSplitTile - 1) / kSplitTile);
  PTGS_CUDA_TRY(cudaFuncSetAttribute(partition_kernel<T>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                     (int)(kSplitTile * sizeof(XY<T>))));
  for (int level = 0; level <= max_depth; ++level) {
    const int cur = level & 1;
    classify_kernel<T><<<1, 1024, 0, st>>>(s, cur, leaf_limit, max_depth);
    PTGS_LAUNCH_CHECK();
    if (level == max_depth) break;
    // every node of the level is a leaf: the remaining levels would only launch empty kernels
    // (one 4-byte read-back per level; c2 stops after 4 of its 10 possible levels)
    int32_t n_active = 0;
    PTGS_CUDA_TRY(cudaMemcpyAsync(&n_active, s.n_nodes + 2, 4, cudaMemcpyDeviceToHost, st));
    PTGS_CUDA_TRY(cudaStreamSynchronize(st));
    if (n_active == 0) break;
    const int max_nodes_here = 1 << level;
    for (int pass = 0; pass < Digits<T>::passes; ++pass) {
      hist_kernel<T><<<tiles, 256, 0, st>>>(s, cur, pass, bufs[cur]);
      PTGS_LAUNCH_CHECK();
      pick_kernel<T><<<max_nodes_here, 256, 0, st>>>(s, cur, pass);
      PTGS_LAUNCH_CHECK();
    }
    layout_kernel<T><<<1, 1024, 0, st>>>(s, cur);
    PTGS_LAUNCH_CHECK();
    partition_kernel<T><<<tiles, 256, kSplitTile * sizeof(XY<T>), st>>>(s, cur, bufs[cur], bufs[cur ^ 1]);
    PTGS_LAUNCH_CHECK();
  }
  finalize_kernel<T><<<1, 1024, 0, st>>>(s, max_depth, leaf_bounds, leaf_path, leaf_depth,
                                         leaf_count, n_leaves, max_leaves);
  PTGS_LAUNCH_CHECK();
  return PTGS_OK;
}

template <typename T>
static size_t split_ws_bytes(int64_t n_capacity, double factor, int max_depth) {
  XY<T>* bufs[2];
  size_t used;
  carve_state<T>(nullptr, split_capacity(n_capacity, factor), max_depth, bufs, &used);
  return used + 256;
}

}  // namespace ptgs

using namespace ptgs;

extern "C" {

size_t ptgs_balanced_split_workspace_bytes(int dtype, int64_t n_capacity,
                                           double capacity_factor, int max_depth) {
  if (max_depth < 0 || max_depth > 12) return 0;
  return dtype == PTGS_F64 ? split_ws_bytes<double>(n_capacity, capacity_factor, max_depth)
                           : split_ws_bytes<float>(n_capacity, capacity_factor, max_depth);
}

int ptgs_balanced_split(int dtype, const void* x, const void* y,
                        const int64_t* n_dev, int64_t n_capacity,
                        const double* root_bounds, double leaf_limit,
                        int max_depth, double capacity_factor,
                        double* leaf_bounds, uint32_t* leaf_path,
                        int32_t* leaf_depth, int64_t* leaf_count,
                        int32_t* n_leaves, int32_t max_leaves,
                        int32_t* status_dev, void* workspace,
                        size_t workspace_bytes, ptgs_stream_t stream) {
  PTGS_REQUIRE(dtype == PTGS_F32 || dtype == PTGS_F64, "dtype");
  PTGS_REQUIRE(max_depth >= 0 && max_depth <= 12, "max_depth must be in [0,12]");
  PTGS_REQUIRE(max_leaves >= (1 << max_depth), "max_leaves < 2^max_depth");
  PTGS_REQUIRE(n_capacity >= 0 && n_dev && root_bounds && status_dev, "arguments");
  if (workspace_bytes < ptgs_balanced_split_workspace_bytes(dtype, n_capacity, capacity_factor, max_depth)) {
    set_error("balanced_split: workspace too small");
    return PTGS_E_WORKSPACE;
  }
  if (dtype == PTGS_F32)
    return split_impl<float>(x, y, n_dev, n_capacity, root_bounds, leaf_limit, max_depth,
                             capacity_factor, leaf_bounds, leaf_path, leaf_depth, leaf_count,
                             n_leaves, max_leaves, status_dev, workspace, (cudaStream_t)stream);
  return split_impl<double>(x, y, n_dev, n_capacity, root_bounds, leaf_limit, max_depth,
                            capacity_factor, leaf_bounds, leaf_path, leaf_depth, leaf_count,
                            n_leaves, max_leaves, status_dev, workspace, (cudaStream_t)stream);
}

}  // extern "C"
