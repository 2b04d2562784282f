// libptgs A14: the camera-grid ("Vast") partitioner's device work,
// scene/ptgs/data_partition.py.  All arithmetic is float64 in NumPy's operation
// order (the reference's point_in_image, :385-424, is float64 NumPy on the
// host); the library is compiled with -fmad=false so no product is fused.
//
//   ptgs_vast_point_range   get_point_range (:285-293) of every block
//   ptgs_vast_airspace      8 projected bounding-box corners -> convex hull ->
//                           clipped by the image -> area / (W H)   (:478-488,
//                           graham_scan.py:65-105)
//   ptgs_vast_coverage      point_in_image of a block's cloud for the accepted
//                           cameras, with the reference's index quirk (:505-506)
//   ptgs_vast_unique        np.unique(axis=0, return_index=True) (:545)
#include <cub/device/device_radix_sort.cuh>
#include <cub/device/device_select.cuh>
#include <cub/iterator/counting_input_iterator.cuh>

#include "common.cuh"

namespace ptgs {

constexpr int kVastCamDoubles = PTGS_VAST_CAMERA_DOUBLES;
constexpr int kVastTile = 1024;          // points per CTA pass of the coverage kernels

struct VCam {
  double r[9], t[3], fx, fy, cx, cy, w, h;
};
__device__ __forceinline__ VCam load_vcam(const double* __restrict__ p) {
  VCam c;
#pragma unroll
  for (int i = 0; i < 9; ++i) c.r[i] = p[i];
#pragma unroll
  for (int i = 0; i < 3; ++i) c.t[i] = p[9 + i];
  c.fx = p[12]; c.fy = p[13]; c.cx = p[14]; c.cy = p[15]; c.w = p[16]; c.h = p[17];
  return c;
}
// camera-space coordinates: np.dot(w2c[:3,:3], p) + T, products summed left to right
__device__ __forceinline__ void to_camera(const VCam& c, double x, double y, double z, double& X,
                                          double& Y, double& Z) {
  X = __dadd_rn(__dadd_rn(__dadd_rn(__dmul_rn(c.r[0], x), __dmul_rn(c.r[1], y)), __dmul_rn(c.r[2], z)), c.t[0]);
  Y = __dadd_rn(__dadd_rn(__dadd_rn(__dmul_rn(c.r[3], x), __dmul_rn(c.r[4], y)), __dmul_rn(c.r[5], z)), c.t[1]);
  Z = __dadd_rn(__dadd_rn(__dadd_rn(__dmul_rn(c.r[6], x), __dmul_rn(c.r[7], y)), __dmul_rn(c.r[8], z)), c.t[2]);
}
// np.dot(K, pc) then / row 2:  u = (fx X + 0 Y + cx Z) / Z,  v = (0 X + fy Y + cy Z) / Z
__device__ __forceinline__ void to_image(const VCam& c, double X, double Y, double Z, double& u,
                                         double& v) {
  const double uh = __dadd_rn(__dadd_rn(__dmul_rn(c.fx, X), __dmul_rn(0.0, Y)), __dmul_rn(c.cx, Z));
  const double vh = __dadd_rn(__dadd_rn(__dmul_rn(0.0, X), __dmul_rn(c.fy, Y)), __dmul_rn(c.cy, Z));
  u = __ddiv_rn(uh, Z);
  v = __ddiv_rn(vh, Z);
}
// the reference compares x with H and y with W (:417-422)
__device__ __forceinline__ bool in_image_quirk(const VCam& c, double u, double v) {
  return u >= 0.0 && u < c.h && v >= 0.0 && v < c.w;
}

// ---- point range -----------------------------------------------------------------
__global__ void vast_range_init_kernel(int B, unsigned long long* __restrict__ keys) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < B * 6) keys[i] = (i % 2 == 0) ? ~0ull : 0ull;
}
template <typename T>
__global__ void __launch_bounds__(256) vast_range_kernel(
    const T* __restrict__ x, const T* __restrict__ y, const T* __restrict__ z,
    const int64_t* __restrict__ offsets, const int32_t* __restrict__ indices, int B,
    unsigned long long* __restrict__ keys) {
  using KD = KeyOf<double>;
  const int b = blockIdx.y;
  const int64_t p0 = offsets[b], p1 = offsets[b + 1];
  uint64_t lo[3] = {~0ull, ~0ull, ~0ull}, hi[3] = {0, 0, 0};
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = p0 + (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < p1; i += stride) {
    const int32_t g = indices[i];
    const uint64_t k[3] = {KD::encode((double)x[g]), KD::encode((double)y[g]), KD::encode((double)z[g])};
#pragma unroll
    for (int a = 0; a < 3; ++a) { lo[a] = min(lo[a], k[a]); hi[a] = max(hi[a], k[a]); }
  }
#pragma unroll
  for (int d = 16; d > 0; d >>= 1) {
#pragma unroll
    for (int a = 0; a < 3; ++a) {
      lo[a] = min(lo[a], __shfl_xor_sync(kFullMask, lo[a], d));
      hi[a] = max(hi[a], __shfl_xor_sync(kFullMask, hi[a], d));
    }
  }
  if (lane_id() == 0 && p1 > p0) {
#pragma unroll
    for (int a = 0; a < 3; ++a) {
      atomicMin(keys + b * 6 + 2 * a, (unsigned long long)lo[a]);
      atomicMax(keys + b * 6 + 2 * a + 1, (unsigned long long)hi[a]);
    }
  }
}
__global__ void vast_range_decode_kernel(int B, const unsigned long long* __restrict__ keys,
                                         double* __restrict__ out) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < B * 6) out[i] = KeyOf<double>::decode(keys[i]);
}

// ---- airspace rate -----------------------------------------------------------------
__device__ __forceinline__ double cross2(double ox, double oy, double ax, double ay, double bx, double by) {
  return (ax - ox) * (by - oy) - (ay - oy) * (bx - ox);
}
__device__ int clip_axis(const double* px, const double* py, int n, int axis, double bound, bool le,
                         double* qx, double* qy) {
  int m = 0;
  for (int i = 0; i < n; ++i) {
    const int k = (i + 1 == n) ? 0 : i + 1;
    const double ci = axis ? py[i] : px[i], ck = axis ? py[k] : px[k];
    const double fi = le ? (ci - bound) : (bound - ci), fk = le ? (ck - bound) : (bound - ck);
    const bool in_i = fi <= 0.0, in_k = fk <= 0.0;
    if (in_i && m < 16) { qx[m] = px[i]; qy[m] = py[i]; ++m; }
    if (in_i != in_k && m < 16) {
      const double t = fi / (fi - fk);
      qx[m] = px[i] + t * (px[k] - px[i]);
      qy[m] = py[i] + t * (py[k] - py[i]);
      ++m;
    }
  }
  return m;
}
__global__ void __launch_bounds__(64) vast_airspace_kernel(
    const double* __restrict__ cams, const double* __restrict__ boxes,
    const int32_t* __restrict__ pair_block, const int32_t* __restrict__ pair_cam, int64_t P,
    double* __restrict__ rate) {
  const int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= P) return;
  const VCam c = load_vcam(cams + (int64_t)pair_cam[p] * kVastCamDoubles);
  const double* bb = boxes + (int64_t)pair_block[p] * 6;   // x0 x1 y0 y1 z0 z1
  double ux[8], uy[8];
  int n = 0;
  for (int k = 0; k < 8; ++k) {                 // get_8_corner_points order (:368-382)
    const double x = bb[(k >> 2) & 1], y = bb[2 + ((k >> 1) & 1)], z = bb[4 + (k & 1)];
    double X, Y, Z;
    to_camera(c, x, y, z, X, Y, Z);
    if (Z > 0.0) { to_image(c, X, Y, Z, ux[n], uy[n]); ++n; }
  }
  double r = 0.0;
  if (n >= 3) {
    // sort by (x, y), drop exact duplicates (insertion sort, n <= 8)
    for (int i = 1; i < n; ++i) {
      const double kx = ux[i], ky = uy[i];
      int j = i - 1;
      while (j >= 0 && (ux[j] > kx || (ux[j] == kx && uy[j] > ky))) { ux[j + 1] = ux[j]; uy[j + 1] = uy[j]; --j; }
      ux[j + 1] = kx; uy[j + 1] = ky;
    }
    int m = 1;
    for (int i = 1; i < n; ++i)
      if (ux[i] != ux[m - 1] || uy[i] != uy[m - 1]) { ux[m] = ux[i]; uy[m] = uy[i]; ++m; }
    if (m >= 3) {
      // Andrew monotone chain, counter-clockwise, collinear vertices dropped
      double hx[16], hy[16];
      int h = 0;
      for (int i = 0; i < m; ++i) {
        while (h >= 2 && cross2(hx[h - 2], hy[h - 2], hx[h - 1], hy[h - 1], ux[i], uy[i]) <= 0.0) --h;
        hx[h] = ux[i]; hy[h] = uy[i]; ++h;
      }
      const int lower = h + 1;
      for (int i = m - 2; i >= 0; --i) {
        while (h >= lower && cross2(hx[h - 2], hy[h - 2], hx[h - 1], hy[h - 1], ux[i], uy[i]) <= 0.0) --h;
        hx[h] = ux[i]; hy[h] = uy[i]; ++h;
      }
      --h;                                       // last point == first
      if (h >= 3) {
        double qx[16], qy[16];
        int k = clip_axis(hx, hy, h, 0, 0.0, false, qx, qy);
        k = clip_axis(qx, qy, k, 0, c.w, true, hx, hy);
        k = clip_axis(hx, hy, k, 1, 0.0, false, qx, qy);
        k = clip_axis(qx, qy, k, 1, c.h, true, hx, hy);
        if (k >= 3) {
          double acc = 0.0;
          for (int i = 0; i < k; ++i) {
            const int j = (i + 1 == k) ? 0 : i + 1;
            acc += hx[i] * hy[j] - hx[j] * hy[i];
          }
          r = 0.5 * fabs(acc) / (c.w * c.h);
        }
      }
    }
  }
  rate[p] = r;
}

// ---- coverage ------------------------------------------------------------------------
// pass 1: per (camera, tile) number of points with camera z > 0
template <typename T>
__global__ void __launch_bounds__(256) vast_front_count_kernel(
    const T* __restrict__ x, const T* __restrict__ y, const T* __restrict__ z,
    const int32_t* __restrict__ idx, int64_t n, const double* __restrict__ cams,
    const int32_t* __restrict__ cam_list, int64_t tiles, uint32_t* __restrict__ tile_front) {
  const int k = blockIdx.y;
  const VCam c = load_vcam(cams + (int64_t)cam_list[k] * kVastCamDoubles);
  __shared__ int s_cnt;
  for (int64_t t = blockIdx.x; t < tiles; t += gridDim.x) {
    if (threadIdx.x == 0) s_cnt = 0;
    __syncthreads();
    int cnt = 0;
    for (int q = 0; q < kVastTile / 256; ++q) {
      const int64_t i = t * kVastTile + q * 256 + threadIdx.x;
      bool f = false;
      if (i < n) {
        const int32_t g = idx[i];
        double X, Y, Z;
        to_camera(c, (double)x[g], (double)y[g], (double)z[g], X, Y, Z);
        f = Z > 0.0;
      }
      cnt += __popc(__ballot_sync(kFullMask, f));
    }
    if (lane_id() == 0) atomicAdd(&s_cnt, cnt);
    __syncthreads();
    if (threadIdx.x == 0) tile_front[(int64_t)k * tiles + t] = (uint32_t)s_cnt;
    __syncthreads();
  }
}
// pass 2: exclusive scan of the tile counts of each camera (one CTA per camera)
__global__ void __launch_bounds__(1024) vast_tile_scan_kernel(int64_t tiles, uint32_t* __restrict__ tile_front) {
  __shared__ uint32_t sm[33];
  uint32_t* row = tile_front + (int64_t)blockIdx.x * tiles;
  uint32_t carry = 0;
  for (int64_t t0 = 0; t0 < tiles; t0 += 1024) {
    const int64_t t = t0 + threadIdx.x;
    const uint32_t v = (t < tiles) ? row[t] : 0u;
    uint32_t inc = v;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      const uint32_t o = __shfl_up_sync(kFullMask, inc, d);
      if (lane_id() >= d) inc += o;
    }
    if (lane_id() == 31) sm[threadIdx.x >> 5] = inc;
    __syncthreads();
    if (threadIdx.x < 32) {
      uint32_t s = sm[threadIdx.x], si = s;
#pragma unroll
      for (int d = 1; d < 32; d <<= 1) {
        const uint32_t o = __shfl_up_sync(kFullMask, si, d);
        if (lane_id() >= d) si += o;
      }
      sm[threadIdx.x] = si - s;
      if (threadIdx.x == 31) sm[32] = si;
    }
    __syncthreads();
    if (t < tiles) row[t] = carry + sm[threadIdx.x >> 5] + (inc - v);
    carry += sm[32];
    __syncthreads();
  }
}
// pass 3: a point that is in front and inside the image selects the block's point number
// `rank` = its position among the points in front (data_partition.py:412, :505-506: the
// mask indexes the z>0-filtered array but is applied to the unfiltered cloud); position
// key = (order number of the camera, rank) = place in the reference's concatenation
template <typename T>
__global__ void __launch_bounds__(256) vast_mark_kernel(
    const T* __restrict__ x, const T* __restrict__ y, const T* __restrict__ z,
    const int32_t* __restrict__ idx, int64_t n, const double* __restrict__ cams,
    const int32_t* __restrict__ cam_list, int32_t seq_base, int64_t tiles,
    const uint32_t* __restrict__ tile_front, unsigned long long* __restrict__ first_pos) {
  const int k = blockIdx.y;
  const VCam c = load_vcam(cams + (int64_t)cam_list[k] * kVastCamDoubles);
  __shared__ uint32_t s_warp[8];
  for (int64_t t = blockIdx.x; t < tiles; t += gridDim.x) {
    uint32_t base = tile_front[(int64_t)k * tiles + t];
    for (int q = 0; q < kVastTile / 256; ++q) {
      const int64_t i = t * kVastTile + q * 256 + threadIdx.x;
      bool f = false, hit = false;
      if (i < n) {
        const int32_t g = idx[i];
        double X, Y, Z;
        to_camera(c, (double)x[g], (double)y[g], (double)z[g], X, Y, Z);
        f = Z > 0.0;
        if (f) {
          double u, v;
          to_image(c, X, Y, Z, u, v);
          hit = in_image_quirk(c, u, v);
        }
      }
      const unsigned bal = __ballot_sync(kFullMask, f);
      if (lane_id() == 0) s_warp[threadIdx.x >> 5] = __popc(bal);
      __syncthreads();
      uint32_t before = 0, total = 0;
#pragma unroll
      for (int w = 0; w < 8; ++w) {
        const uint32_t v = s_warp[w];
        if (w < (int)(threadIdx.x >> 5)) before += v;
        total += v;
      }
      if (hit) {
        const uint32_t rank = base + before + __popc(bal & ((1u << lane_id()) - 1u));
        const int32_t g2 = idx[rank];              // rank <= i < n
        atomicMin(first_pos + g2, ((unsigned long long)(uint32_t)(seq_base + k) << 32) | rank);
      }
      base += total;
      __syncthreads();
    }
  }
}
__global__ void __launch_bounds__(256) vast_mark_own_kernel(
    const int32_t* __restrict__ idx, int64_t n, uint32_t seq, unsigned long long* __restrict__ first_pos) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride)
    atomicMin(first_pos + idx[i], ((unsigned long long)seq << 32) | (uint32_t)i);
}

// ---- unique rows ------------------------------------------------------------------------
struct IsMarked {
  const unsigned long long* fp;
  __device__ bool operator()(int32_t g) const { return fp[g] != ~0ull; }
};
__global__ void __launch_bounds__(256) vast_gather_pos_kernel(
    const unsigned long long* __restrict__ first_pos, const int32_t* __restrict__ sel, int64_t m,
    unsigned long long* __restrict__ keys) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < m; i += stride)
    keys[i] = first_pos[sel[i]];
}
template <typename T>
__global__ void __launch_bounds__(256) vast_gather_key_kernel(
    const T* __restrict__ coord, const int32_t* __restrict__ sel, int64_t m,
    typename KeyOf<T>::type* __restrict__ keys) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < m; i += stride)
    keys[i] = KeyOf<T>::encode(coord[sel[i]]);
}
template <typename T>
__global__ void __launch_bounds__(256) vast_head_flag_kernel(
    const T* __restrict__ x, const T* __restrict__ y, const T* __restrict__ z,
    const int32_t* __restrict__ sel, int64_t m, uint8_t* __restrict__ head) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < m; i += stride) {
    bool h = true;
    if (i > 0) {
      const int32_t a = sel[i - 1], b = sel[i];
      h = !(x[a] == x[b] && y[a] == y[b] && z[a] == z[b]);
    }
    head[i] = h ? 1 : 0;
  }
}

static inline int vast_grid(int64_t n, int per_block) {
  int64_t b = (n + per_block - 1) / per_block;
  if (b < 1) b = 1;
  if (b > kNumSMs * 8) b = kNumSMs * 8;
  return (int)b;
}

template <typename K>
static size_t vast_sort_temp(int64_t n) {
  size_t bytes = 0;
  cub::DeviceRadixSort::SortPairs(nullptr, bytes, (const K*)nullptr, (K*)nullptr, (const int32_t*)nullptr,
                                  (int32_t*)nullptr, (int64_t)n);
  return bytes;
}
static size_t vast_select_temp(int64_t n) {
  size_t a = 0, b = 0;
  cub::DeviceSelect::If(nullptr, a, cub::CountingInputIterator<int32_t>(0), (int32_t*)nullptr,
                        (int64_t*)nullptr, (int64_t)n, IsMarked{nullptr});
  cub::DeviceSelect::Flagged(nullptr, b, (const int32_t*)nullptr, (const uint8_t*)nullptr, (int32_t*)nullptr,
                             (int64_t*)nullptr, (int64_t)n);
  return a > b ? a : b;
}

template <typename T>
static int vast_unique_impl(const T* x, const T* y, const T* z, int64_t n,
                            const unsigned long long* first_pos, int32_t* uniq_index, int64_t capacity,
                            int64_t* n_unique_dev, void* workspace, cudaStream_t st) {
  using K = typename KeyOf<T>::type;
  Carver cv(workspace);
  int64_t* m_dev = cv.take<int64_t>(2);
  int32_t* sel_a = cv.take<int32_t>(n);
  int32_t* sel_b = cv.take<int32_t>(n);
  unsigned long long* k64_a = cv.take<unsigned long long>(n);
  unsigned long long* k64_b = cv.take<unsigned long long>(n);
  uint8_t* head = cv.take<uint8_t>(n);
  size_t tb = vast_select_temp(n);
  size_t t2 = vast_sort_temp<unsigned long long>(n);
  if (t2 > tb) tb = t2;
  void* tmp = cv.take<char>(tb);

  PTGS_CUDA_TRY(cub::DeviceSelect::If(tmp, tb, cub::CountingInputIterator<int32_t>(0), sel_a, m_dev, (int64_t)n,
                                      IsMarked{first_pos}, st));
  count_launch(2);
  int64_t m = 0;
  PTGS_CUDA_TRY(cudaMemcpyAsync(&m, m_dev, 8, cudaMemcpyDeviceToHost, st));
  PTGS_CUDA_TRY(cudaStreamSynchronize(st));
  if (m > capacity) {
    set_error("vast_unique: %lld selected points exceed the capacity %lld", (long long)m, (long long)capacity);
    return PTGS_E_CAPACITY;
  }
  if (m == 0) {
    PTGS_CUDA_TRY(cudaMemsetAsync(n_unique_dev, 0, 8, st));
    return PTGS_OK;
  }
  const int g = vast_grid(m, 1024);
  // least significant key first: concatenation position, then z, y, x (stable passes)
  vast_gather_pos_kernel<<<g, 256, 0, st>>>(first_pos, sel_a, m, k64_a);
  PTGS_LAUNCH_CHECK();
  PTGS_CUDA_TRY(cub::DeviceRadixSort::SortPairs(tmp, tb, k64_a, k64_b, sel_a, sel_b, m, 0, 64, st));
  count_launch(10);
  int32_t* cur = sel_b;
  int32_t* oth = sel_a;
  const T* axes[3] = {z, y, x};
  for (int a = 0; a < 3; ++a) {
    K* ka = reinterpret_cast<K*>(k64_a);
    K* kb = reinterpret_cast<K*>(k64_b);
    vast_gather_key_kernel<T><<<g, 256, 0, st>>>(axes[a], cur, m, ka);
    PTGS_LAUNCH_CHECK();
    PTGS_CUDA_TRY(cub::DeviceRadixSort::SortPairs(tmp, tb, ka, kb, cur, oth, m, 0, (int)sizeof(K) * 8, st));
    count_launch(2 + (int)sizeof(K));
    int32_t* t = cur; cur = oth; oth = t;
  }
  vast_head_flag_kernel<T><<<g, 256, 0, st>>>(x, y, z, cur, m, head);
  PTGS_LAUNCH_CHECK();
  PTGS_CUDA_TRY(cub::DeviceSelect::Flagged(tmp, tb, cur, head, uniq_index, n_unique_dev, m, st));
  count_launch(2);
  return PTGS_OK;
}

}  // namespace ptgs

using namespace ptgs;

extern "C" {

int ptgs_vast_point_range(int dtype, const void* x, const void* y, const void* z,
                          const int64_t* offsets, const int32_t* indices, int32_t n_blocks,
                          double* range, void* workspace, size_t workspace_bytes,
                          ptgs_stream_t stream) {
  PTGS_REQUIRE(n_blocks >= 1 && n_blocks <= 65535, "vast_point_range: n_blocks in [1, 65535]");
  PTGS_REQUIRE(x && y && z && offsets && indices && range && workspace, "vast_point_range: null pointer");
  PTGS_REQUIRE(workspace_bytes >= (size_t)n_blocks * 48, "vast_point_range: workspace of n_blocks*48 bytes");
  cudaStream_t st = (cudaStream_t)stream;
  unsigned long long* keys = (unsigned long long*)workspace;
  const int gi = (n_blocks * 6 + 127) / 128;
  vast_range_init_kernel<<<gi, 128, 0, st>>>(n_blocks, keys);
  PTGS_LAUNCH_CHECK();
  dim3 grid(kNumSMs, n_blocks);
  if (dtype == PTGS_F64)
    vast_range_kernel<double><<<grid, 256, 0, st>>>((const double*)x, (const double*)y, (const double*)z,
                                                    offsets, indices, n_blocks, keys);
  else
    vast_range_kernel<float><<<grid, 256, 0, st>>>((const float*)x, (const float*)y, (const float*)z,
                                                   offsets, indices, n_blocks, keys);
  PTGS_LAUNCH_CHECK();
  vast_range_decode_kernel<<<gi, 128, 0, st>>>(n_blocks, keys, range);
  PTGS_LAUNCH_CHECK();
  return PTGS_OK;
}

int ptgs_vast_airspace(const double* cameras, const double* boxes, const int32_t* pair_block,
                       const int32_t* pair_cam, int64_t n_pairs, double* rate, ptgs_stream_t stream) {
  PTGS_REQUIRE(n_pairs >= 0, "vast_airspace: n_pairs");
  if (n_pairs == 0) return PTGS_OK;
  PTGS_REQUIRE(cameras && boxes && pair_block && pair_cam && rate, "vast_airspace: null pointer");
  vast_airspace_kernel<<<(unsigned)((n_pairs + 63) / 64), 64, 0, (cudaStream_t)stream>>>(
      cameras, boxes, pair_block, pair_cam, n_pairs, rate);
  PTGS_LAUNCH_CHECK();
  return PTGS_OK;
}

int ptgs_vast_begin_block(uint64_t* first_pos, int64_t n, ptgs_stream_t stream) {
  PTGS_REQUIRE(n >= 0, "vast_begin_block: n");
  if (n == 0) return PTGS_OK;
  PTGS_REQUIRE(first_pos, "vast_begin_block: null pointer");
  PTGS_CUDA_TRY(cudaMemsetAsync(first_pos, 0xff, (size_t)n * 8, (cudaStream_t)stream));
  return PTGS_OK;
}

size_t ptgs_vast_coverage_workspace_bytes(int64_t n_block_points, int32_t n_cameras) {
  const int64_t tiles = (n_block_points + kVastTile - 1) / kVastTile;
  return 256 + (size_t)(tiles > 0 ? tiles : 1) * (size_t)(n_cameras > 0 ? n_cameras : 1) * 4;
}

int ptgs_vast_coverage(int dtype, const void* x, const void* y, const void* z, const int32_t* block_index,
                       int64_t n_block_points, const double* cameras, const int32_t* cam_list,
                       int32_t n_cameras, int32_t seq_base, uint64_t* first_pos, void* workspace,
                       size_t workspace_bytes, ptgs_stream_t stream) {
  PTGS_REQUIRE(n_block_points >= 0 && n_block_points < 4294967296ll && n_cameras >= 0 && seq_base >= 0,
               "vast_coverage: sizes");
  if (n_block_points == 0 || n_cameras == 0) return PTGS_OK;
  PTGS_REQUIRE(n_cameras <= 65535, "vast_coverage: at most 65535 cameras per call");
  PTGS_REQUIRE(x && y && z && block_index && cameras && cam_list && first_pos && workspace,
               "vast_coverage: null pointer");
  if (workspace_bytes < ptgs_vast_coverage_workspace_bytes(n_block_points, n_cameras)) {
    set_error("vast_coverage: workspace too small");
    return PTGS_E_WORKSPACE;
  }
  cudaStream_t st = (cudaStream_t)stream;
  const int64_t tiles = (n_block_points + kVastTile - 1) / kVastTile;
  uint32_t* tile_front = (uint32_t*)workspace;
  int gx = (int)(tiles < kNumSMs * 4 ? tiles : kNumSMs * 4);
  dim3 grid(gx, n_cameras);
  unsigned long long* fp = (unsigned long long*)first_pos;
  if (dtype == PTGS_F64) {
    vast_front_count_kernel<double><<<grid, 256, 0, st>>>((const double*)x, (const double*)y, (const double*)z,
                                                          block_index, n_block_points, cameras, cam_list, tiles,
                                                          tile_front);
    PTGS_LAUNCH_CHECK();
    vast_tile_scan_kernel<<<n_cameras, 1024, 0, st>>>(tiles, tile_front);
    PTGS_LAUNCH_CHECK();
    vast_mark_kernel<double><<<grid, 256, 0, st>>>((const double*)x, (const double*)y, (const double*)z,
                                                   block_index, n_block_points, cameras, cam_list, seq_base,
                                                   tiles, tile_front, fp);
  } else {
    vast_front_count_kernel<float><<<grid, 256, 0, st>>>((const float*)x, (const float*)y, (const float*)z,
                                                         block_index, n_block_points, cameras, cam_list, tiles,
                                                         tile_front);
    PTGS_LAUNCH_CHECK();
    vast_tile_scan_kernel<<<n_cameras, 1024, 0, st>>>(tiles, tile_front);
    PTGS_LAUNCH_CHECK();
    vast_mark_kernel<float><<<grid, 256, 0, st>>>((const float*)x, (const float*)y, (const float*)z,
                                                  block_index, n_block_points, cameras, cam_list, seq_base,
                                                  tiles, tile_front, fp);
  }
  PTGS_LAUNCH_CHECK();
  return PTGS_OK;
}

int ptgs_vast_mark_own(const int32_t* block_index, int64_t n_block_points, int32_t seq,
                       uint64_t* first_pos, ptgs_stream_t stream) {
  PTGS_REQUIRE(n_block_points >= 0 && n_block_points < 4294967296ll && seq >= 0, "vast_mark_own: sizes");
  if (n_block_points == 0) return PTGS_OK;
  PTGS_REQUIRE(block_index && first_pos, "vast_mark_own: null pointer");
  vast_mark_own_kernel<<<vast_grid(n_block_points, 1024), 256, 0, (cudaStream_t)stream>>>(
      block_index, n_block_points, (uint32_t)seq, (unsigned long long*)first_pos);
  PTGS_LAUNCH_CHECK();
  return PTGS_OK;
}

size_t ptgs_vast_unique_workspace_bytes(int64_t n) {
  if (n < 1) n = 1;
  size_t b = 512;
  b += 2 * align_up((size_t)n * 4, 256);
  b += 2 * align_up((size_t)n * 8, 256);
  b += align_up((size_t)n, 256);
  size_t tb = vast_select_temp(n), t2 = vast_sort_temp<unsigned long long>(n);
  b += align_up(tb > t2 ? tb : t2, 256) + 4096;
  return b;
}

int ptgs_vast_unique(int dtype, const void* x, const void* y, const void* z, int64_t n,
                     const uint64_t* first_pos, int32_t* uniq_index, int64_t capacity,
                     int64_t* n_unique_dev, void* workspace, size_t workspace_bytes,
                     ptgs_stream_t stream) {
  PTGS_REQUIRE(n >= 0 && n < 2147483648ll && capacity >= 0, "vast_unique: sizes");
  PTGS_REQUIRE(n_unique_dev, "vast_unique: null pointer");
  cudaStream_t st = (cudaStream_t)stream;
  if (n == 0) {
    PTGS_CUDA_TRY(cudaMemsetAsync(n_unique_dev, 0, 8, st));
    return PTGS_OK;
  }
  PTGS_REQUIRE(x && y && z && first_pos && uniq_index && workspace, "vast_unique: null pointer");
  if (workspace_bytes < ptgs_vast_unique_workspace_bytes(n)) {
    set_error("vast_unique: workspace too small");
    return PTGS_E_WORKSPACE;
  }
  if (dtype == PTGS_F64)
    return vast_unique_impl<double>((const double*)x, (const double*)y, (const double*)z, n,
                                    (const unsigned long long*)first_pos, uniq_index, capacity, n_unique_dev,
                                    workspace, st);
  return vast_unique_impl<float>((const float*)x, (const float*)y, (const float*)z, n,
                                 (const unsigned long long*)first_pos, uniq_index, capacity, n_unique_dev,
                                 workspace, st);
}

}  // extern "C"
