// libptgs A1: radius outlier mask (open3d remove_radius_outlier semantics as
// called at scene/ptgs/shen_partition_utils.py:50).
//
// Uniform XY column grid with cell = radius*(1+2^-13) (>= radius so the 3x3
// column neighbourhood is a superset of the ball).  Points are sorted by
// column key (ix*ny+iy); the three columns (ix+dx, iy-1..iy+1) are one
// contiguous run of the sorted array, found with one binary search.
// Distances are evaluated in float64 without fma:  (dx*dx + dy*dy) + dz*dz.
#include <cub/device/device_radix_sort.cuh>

#include "common.cuh"

namespace ptgs {

struct GridSpec {
  double xmin, ymin, inv_cell;
  long long nx, ny;
  // narrow (32-bit) keys: column << zbits | z bucket, bucket = floor((z - zmin) * inv_cell) >> zshift
  double zmin;
  long long nzf;        // fine z cells (cell = the XY cell size)
  int zbits, zshift;
};

template <typename T>
__global__ void __launch_bounds__(256) minmax_xy_kernel(
    const T* __restrict__ x, const T* __restrict__ y, const T* __restrict__ z, int64_t n,
    unsigned long long* __restrict__ keys /*xmin,ymin,xmax,ymax,zmin,zmax as u64 keys of double*/) {
  using KD = KeyOf<double>;
  uint64_t xmin = ~0ull, ymin = ~0ull, zmin = ~0ull, xmax = 0, ymax = 0, zmax = 0;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    uint64_t kx = KD::encode((double)x[i]), ky = KD::encode((double)y[i]), kz = KD::encode((double)z[i]);
    xmin = min(xmin, kx); xmax = max(xmax, kx);
    ymin = min(ymin, ky); ymax = max(ymax, ky);
    zmin = min(zmin, kz); zmax = max(zmax, kz);
  }
#pragma unroll
  for (int d = 16; d > 0; d >>= 1) {
    xmin = min(xmin, __shfl_xor_sync(kFullMask, xmin, d));
    ymin = min(ymin, __shfl_xor_sync(kFullMask, ymin, d));
    zmin = min(zmin, __shfl_xor_sync(kFullMask, zmin, d));
    xmax = max(xmax, __shfl_xor_sync(kFullMask, xmax, d));
    ymax = max(ymax, __shfl_xor_sync(kFullMask, ymax, d));
    zmax = max(zmax, __shfl_xor_sync(kFullMask, zmax, d));
  }
  if (lane_id() == 0) {
    atomicMin(keys + 0, (unsigned long long)xmin);
    atomicMin(keys + 1, (unsigned long long)ymin);
    atomicMax(keys + 2, (unsigned long long)xmax);
    atomicMax(keys + 3, (unsigned long long)ymax);
    atomicMin(keys + 4, (unsigned long long)zmin);
    atomicMax(keys + 5, (unsigned long long)zmax);
  }
}

__device__ __forceinline__ long long cell_of(double v, double vmin, double inv_cell, long long nmax) {
  long long c = (long long)floor((v - vmin) * inv_cell);
  return c < 0 ? 0 : (c >= nmax ? nmax - 1 : c);
}

// order-preserving 32-bit image of float(z): inside a column the points are
// sorted by height, so a query only walks the slab |dz| <= r of each column
__device__ __forceinline__ uint32_t zkey_of(float f) {
  const uint32_t b = __float_as_uint(f);
  return (b & 0x80000000u) ? ~b : (b | 0x80000000u);
}

// z bucket of the narrow key: monotone in z (every operation is a correctly rounded,
// monotone float64 function), clamped to the bucket range
__device__ __forceinline__ uint32_t zbucket_of(double z, const GridSpec& g) {
  const double f = floor((z - g.zmin) * g.inv_cell);
  const long long c = (f >= 0.0) ? ((f < (double)g.nzf) ? (long long)f : g.nzf - 1) : 0;   // NaN -> 0
  return (uint32_t)(c >> g.zshift);
}

// Key layout: column << ZB | z part.  WIDE: 64-bit keys, ZB = 32, z part = the order-preserving
// image of float(z) (exact height order inside a column).  Narrow: 32-bit keys, ZB = g.zbits, z
// part = the z bucket -- half the sort traffic and 4 instead of 7 radix passes at 50 M points.
template <typename T, typename K>
__global__ void __launch_bounds__(256) column_key_kernel(
    const T* __restrict__ x, const T* __restrict__ y, const T* __restrict__ z, int64_t n, GridSpec g,
    K* __restrict__ keys, uint32_t* __restrict__ idx) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    long long ix = cell_of((double)x[i], g.xmin, g.inv_cell, g.nx);
    long long iy = cell_of((double)y[i], g.ymin, g.inv_cell, g.ny);
    const uint64_t col = (uint64_t)(ix * g.ny + iy);
    if (sizeof(K) == 8)
      keys[i] = (K)((col << 32) | zkey_of((float)z[i]));   // float(z): monotone in z
    else
      keys[i] = (K)(((uint32_t)col << g.zbits) | zbucket_of((double)z[i], g));
    idx[i] = (uint32_t)i;
  }
}

// Sorted points as packed (x, y, z, w) records, w = the bit pattern of the key's z part: the
// neighbour walk reads ONE record per candidate (a full 16 / 32-byte sector piece) instead of
// three plane values and the key.  `aos` (may be null): the cloud as (N,3) rows -- the gather
// from file order then touches one or two sectors per point instead of three.
template <typename T> struct V4Of;
template <> struct V4Of<float> {
  using type = float4;
  __device__ static __forceinline__ float pack(uint32_t z) { return __uint_as_float(z); }
  __device__ static __forceinline__ uint32_t unpack(float w) { return __float_as_uint(w); }
};
template <> struct V4Of<double> {
  using type = double4;
  __device__ static __forceinline__ double pack(uint32_t z) { return __longlong_as_double((long long)z); }
  __device__ static __forceinline__ uint32_t unpack(double w) { return (uint32_t)__double_as_longlong(w); }
};

template <typename T, typename K>
__global__ void __launch_bounds__(256) gather_sorted_kernel(
    const T* __restrict__ x, const T* __restrict__ y, const T* __restrict__ z, const T* __restrict__ aos,
    const uint32_t* __restrict__ idx, const K* __restrict__ keys, uint32_t zmask, int64_t n,
    typename V4Of<T>::type* __restrict__ sp) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    const int64_t j = idx[i];
    typename V4Of<T>::type v;
    if (aos) { v.x = aos[3 * j]; v.y = aos[3 * j + 1]; v.z = aos[3 * j + 2]; }
    else { v.x = x[j]; v.y = y[j]; v.z = z[j]; }
    v.w = V4Of<T>::pack((uint32_t)keys[i] & zmask);
    sp[i] = v;
  }
}

__device__ __forceinline__ int64_t lower_bound_u64(const uint64_t* __restrict__ a, int64_t n, uint64_t v) {
  int64_t lo = 0, hi = n;
  while (lo < hi) {
    int64_t mid = (lo + hi) >> 1;
    if (__ldg(a + mid) < v) lo = mid + 1; else hi = mid;
  }
  return lo;
}

template <typename T>
__global__ void __launch_bounds__(256) radius_count_kernel(
    const typename V4Of<T>::type* __restrict__ sp,
    const uint64_t* __restrict__ keys, const uint32_t* __restrict__ idx,
    int64_t n, GridSpec g, double r2, int nb, uint8_t* __restrict__ keep,
    unsigned long long* __restrict__ n_keep) {
  const int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  int kept = 0;
  if (p < n) {
    const typename V4Of<T>::type me = sp[p];
    const double xi = (double)me.x, yi = (double)me.y, zi = (double)me.z;
    const uint64_t key = keys[p] >> 32;
    const long long ix = (long long)(key / (uint64_t)g.ny);
    const long long iy = (long long)(key % (uint64_t)g.ny);
    const long long y0 = iy > 0 ? iy - 1 : 0;
    const long long y1 = iy + 1 < g.ny ? iy + 1 : g.ny - 1;
    int cnt = 0;
    for (int dx = -1; dx <= 1 && cnt <= nb; ++dx) {
      long long cx = ix + dx;
      if (cx < 0 || cx >= g.nx) continue;
      const uint64_t klo = (uint64_t)(cx * g.ny + y0), khi = (uint64_t)(cx * g.ny + y1);
      int64_t j = lower_bound_u64(keys, n, klo << 32);
      for (; j < n && cnt <= nb; ++j) {
        if ((__ldg(keys + j) >> 32) > khi) break;
        const typename V4Of<T>::type q = sp[j];
        double ddx = __dsub_rn((double)q.x, xi);
        double ddy = __dsub_rn((double)q.y, yi);
        double ddz = __dsub_rn((double)q.z, zi);
        double d2 = __dadd_rn(__dadd_rn(__dmul_rn(ddx, ddx), __dmul_rn(ddy, ddy)),
                              __dmul_rn(ddz, ddz));
        cnt += (d2 < r2) ? 1 : 0;
      }
    }
    kept = cnt > nb;
    keep[idx[p]] = (uint8_t)kept;
  }
  int s = warp_sum(kept);
  if (lane_id() == 0 && s) atomicAdd(n_keep, (unsigned long long)s);
}


// cell_start[k] = first sorted position whose key is >= k (k in [0, ncell]); one
// thread per sorted point fills the gap of (possibly empty) cells that ends at
// its own key.
template <typename K>
__global__ void __launch_bounds__(256) cell_start_kernel(const K* __restrict__ keys, int64_t n, int zb,
                                                         uint64_t ncell, uint32_t* __restrict__ cs) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const uint64_t k = (uint64_t)(keys[i] >> zb);
  if (i == 0) {
    for (uint64_t c = 0; c <= k; ++c) cs[c] = 0u;
  } else {
    const uint64_t kp = (uint64_t)(keys[i - 1] >> zb);
    for (uint64_t c = kp + 1; c <= k; ++c) cs[c] = (uint32_t)i;
  }
  if (i == n - 1)
    for (uint64_t c = k + 1; c <= ncell; ++c) cs[c] = (uint32_t)n;
}

// Neighbour count with the cell table: the point's own cell first (most likely
// to hold close neighbours, so the `> nb_points` early exit comes soonest), then
// the 8 surrounding cells.
template <typename T, typename K>
__global__ void __launch_bounds__(256) radius_count_cells_kernel(
    const typename V4Of<T>::type* __restrict__ sp,
    const K* __restrict__ keys, const uint32_t* __restrict__ idx,
    const uint32_t* __restrict__ cs, int64_t n, GridSpec g, double r2, int nb,
    uint8_t* __restrict__ keep, unsigned long long* __restrict__ n_keep) {
  constexpr bool kWide = sizeof(K) == 8;
  const int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  int kept = 0;
  if (p < n) {
    const typename V4Of<T>::type me = sp[p];
    const double xi = (double)me.x, yi = (double)me.y, zi = (double)me.z;
    const int zb = kWide ? 32 : g.zbits;
    const uint32_t zmask = kWide ? 0xffffffffu : ((1u << g.zbits) - 1u);
    const uint64_t key = (uint64_t)(keys[p] >> zb);
    const long long ix = (long long)(key / (uint64_t)g.ny);
    const long long iy = (long long)(key % (uint64_t)g.ny);
    // height slab in the sort order of the column.  Wide keys: the (rounded-to-float) height
    // itself, conservative by one float step.  Narrow keys: the buckets of zi -+ r with the
    // difference rounded away from zi, so every z within r of zi falls into [zlo, zhi] (the
    // bucket function is monotone).
    const double rr = sqrt(r2) * (1.0 + 1e-12);    // never below the true root: |dz| < sqrt(r2) stays inside
    uint32_t zlo, zhi;
    if (kWide) {
      zlo = zkey_of(__double2float_rd(zi - rr)) - 1u;
      zhi = zkey_of(__double2float_ru(zi + rr)) + 1u;
    } else {
      zlo = zbucket_of(__dadd_rd(zi, -rr), g);
      zhi = zbucket_of(__dadd_ru(zi, rr), g);
    }
    int cnt = 0;
    // float32 screening of the float64 test d2 < r2 (the FP64 pipe issues at half rate and the walk
    // is issue bound).  With af = fl32(a): |fl32(bf - af) - (b - a)| <= 2^-24 (|a| + |b| + |b - a|) per
    // axis; candidates of the walk are within 2 cells in x / y and r + 4 cells in z (D <= 6), so
    // |sum df^2 - d2| <= 2 D sum_axes(err) <= 2^-24 (24 S + 432) with S = |x| + |y| + |z| of the query,
    // plus 3 roundings of the sum (<= 2^-22 d2).  M and the relative term carry a 4x / 8x margin;
    // whatever falls inside the band is decided by the exact float64 expression.
    const float xf = (float)me.x, yf = (float)me.y, zf = (float)me.z;
    const float Mtol = 2.384185791015625e-07f * (24.0f * (fabsf(xf) + fabsf(yf) + fabsf(zf)) + 432.0f);
    const float r2f = (float)r2;
    // visiting order: (0,0), then the 4 edge neighbours, then the 4 corners
    const int ox[9] = {0, 0, 0, -1, 1, -1, -1, 1, 1};
    const int oy[9] = {0, -1, 1, 0, 0, -1, 1, -1, 1};
#pragma unroll 1
    for (int c = 0; c < 9 && cnt <= nb; ++c) {
      const long long cx = ix + ox[c], cy = iy + oy[c];
      if (cx < 0 || cx >= g.nx || cy < 0 || cy >= g.ny) continue;
      const uint64_t k = (uint64_t)(cx * g.ny + cy);
      const uint32_t j1 = __ldg(cs + k + 1);
      uint32_t lo = __ldg(cs + k), hi = j1;
      while (lo < hi) {                      // first point of the column inside the slab
        const uint32_t mid = (lo + hi) >> 1;
        if (((uint32_t)__ldg(keys + mid) & zmask) < zlo) lo = mid + 1; else hi = mid;
      }
      for (uint32_t j = lo; j < j1 && cnt <= nb; ++j) {
        const typename V4Of<T>::type q = sp[j];
        if (V4Of<T>::unpack(q.w) > zhi) break;
        const float dxf = __fsub_rn((float)q.x, xf), dyf = __fsub_rn((float)q.y, yf), dzf = __fsub_rn((float)q.z, zf);
        const float d2f = __fmaf_rn(dzf, dzf, __fmaf_rn(dyf, dyf, __fmul_rn(dxf, dxf)));
        const float tol = __fmaf_rn(1.9073486328125e-06f, d2f, Mtol);
        if (d2f < __fsub_rn(r2f, tol)) {
          ++cnt;
        } else if (d2f <= __fadd_rn(r2f, tol)) {
          double ddx = __dsub_rn((double)q.x, xi);
          double ddy = __dsub_rn((double)q.y, yi);
          double ddz = __dsub_rn((double)q.z, zi);
          double d2 = __dadd_rn(__dadd_rn(__dmul_rn(ddx, ddx), __dmul_rn(ddy, ddy)),
                                __dmul_rn(ddz, ddz));
          cnt += (d2 < r2) ? 1 : 0;
        }
      }
    }
    kept = cnt > nb;
    keep[idx[p]] = (uint8_t)kept;
  }
  int s = warp_sum(kept);
  if (lane_id() == 0 && s) atomicAdd(n_keep, (unsigned long long)s);
}

// ---------------------------------------------------------------------------
// Multi-GPU A1: rank r filters the points of the X slab [lo, hi) and needs, besides them, the
// points within one cell of the slab (halo) as neighbours.  slab_select compacts the planes of
// { p : lo - halo <= x < hi + halo } (index order kept) together with their cloud indices;
// slab_scatter writes the slab's own keep flags back to the cloud-sized mask (zero elsewhere, so
// the ranks' masks combine with one MAX all-reduce).
// ---------------------------------------------------------------------------
constexpr int kSlabTile = 2048;

template <typename T>
__global__ void __launch_bounds__(256) slab_count_kernel(const T* __restrict__ x, int64_t n, double lo, double hi,
                                                         int32_t* __restrict__ tile_count) {
  __shared__ int s[33];
  const int64_t base = (int64_t)blockIdx.x * kSlabTile;
  int c = 0;
#pragma unroll
  for (int k = 0; k < 8; ++k) {
    const int64_t i = base + k * 256 + threadIdx.x;
    if (i < n) { const double v = (double)x[i]; c += (v >= lo && v < hi) ? 1 : 0; }
  }
  int tot;
  block_exclusive_sum<int, 256>(c, s, &tot);
  if (threadIdx.x == 0) tile_count[blockIdx.x] = tot;
}

__global__ void __launch_bounds__(1024) slab_scan_kernel(const int32_t* __restrict__ tile_count, int64_t n_tiles,
                                                         int64_t* __restrict__ tile_base, int64_t* __restrict__ total) {
  __shared__ long long s[33];
  long long carry = 0;
  for (int64_t start = 0; start < n_tiles; start += 1024) {
    const int64_t i = start + threadIdx.x;
    long long v = (i < n_tiles) ? tile_count[i] : 0;
    long long tot;
    long long ex = block_exclusive_sum<long long, 1024>(v, s, &tot);
    if (i < n_tiles) tile_base[i] = carry + ex;
    carry += tot;
  }
  if (threadIdx.x == 0) *total = carry;
}

template <typename T>
__global__ void __launch_bounds__(256) slab_scatter_kernel(
    const T* __restrict__ x, const T* __restrict__ y, const T* __restrict__ z, int64_t n, double lo, double hi,
    const int64_t* __restrict__ tile_base, T* __restrict__ ox, T* __restrict__ oy, T* __restrict__ oz,
    int32_t* __restrict__ oidx) {
  __shared__ int s[33];
  const int64_t base = (int64_t)blockIdx.x * kSlabTile;
  int64_t out = tile_base[blockIdx.x];
#pragma unroll 1
  for (int k = 0; k < 8; ++k) {
    const int64_t i = base + k * 256 + threadIdx.x;
    T vx = 0;
    int f = 0;
    if (i < n) { vx = x[i]; const double v = (double)vx; f = (v >= lo && v < hi) ? 1 : 0; }
    int tot;
    const int ex = block_exclusive_sum<int, 256>(f, s, &tot);
    if (f) {
      ox[out + ex] = vx; oy[out + ex] = y[i]; oz[out + ex] = z[i];
      oidx[out + ex] = (int32_t)i;
    }
    out += tot;
  }
}

template <typename T>
__global__ void __launch_bounds__(256) slab_keep_kernel(const T* __restrict__ sx, const int32_t* __restrict__ sidx,
                                                        const uint8_t* __restrict__ skeep, int64_t m, double lo,
                                                        double hi, uint8_t* __restrict__ keep) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < m; i += stride) {
    const double v = (double)sx[i];
    if (v >= lo && v < hi) keep[sidx[i]] = skeep[i];
  }
}

static inline uint64_t cell_table_capacity(int64_t n) { return (uint64_t)4 * (uint64_t)n + (1ull << 20); }

static inline int grid_for(int64_t n, int per_block, int max_blocks) {
  int64_t b = (n + per_block - 1) / per_block;
  if (b < 1) b = 1;
  if (b > max_blocks) b = max_blocks;
  return (int)b;
}

static size_t sort_temp_bytes(int64_t n) {
  size_t b64 = 0, b32 = 0;
  cub::DeviceRadixSort::SortPairs(nullptr, b64, (const uint64_t*)nullptr, (uint64_t*)nullptr,
                                  (const uint32_t*)nullptr, (uint32_t*)nullptr, (int64_t)n, 0, 64);
  cub::DeviceRadixSort::SortPairs(nullptr, b32, (const uint32_t*)nullptr, (uint32_t*)nullptr,
                                  (const uint32_t*)nullptr, (uint32_t*)nullptr, (int64_t)n, 0, 32);
  return b64 > b32 ? b64 : b32;
}

template <typename T>
static int radius_impl(const void* vx, const void* vy, const void* vz, const void* vaos, int64_t n,
                       double radius, int nb, uint8_t* keep, int64_t* n_keep,
                       void* workspace, cudaStream_t st) {
  using V4 = typename V4Of<T>::type;
  const T* x = (const T*)vx; const T* y = (const T*)vy; const T* z = (const T*)vz;
  const T* aos = (const T*)vaos;
  Carver cv(workspace);
  unsigned long long* mm = cv.take<unsigned long long>(6);
  uint64_t* keys_a = cv.take<uint64_t>(n);
  uint64_t* keys_b = cv.take<uint64_t>(n);
  uint32_t* idx_a = cv.take<uint32_t>(n);
  uint32_t* idx_b = cv.take<uint32_t>(n);
  V4* sp = cv.take<V4>(n);
  size_t tb = sort_temp_bytes(n);
  void* sort_tmp = cv.take<char>(tb);
  uint32_t* cell_start = cv.take<uint32_t>(cell_table_capacity(n) + 2);

  const unsigned long long init[6] = {~0ull, ~0ull, 0ull, 0ull, ~0ull, 0ull};
  PTGS_CUDA_TRY(cudaMemcpyAsync(mm, init, sizeof(init), cudaMemcpyHostToDevice, st));
  PTGS_CUDA_TRY(cudaMemsetAsync(n_keep, 0, 8, st));
  int g1 = grid_for(n, 2048, kNumSMs * 8);
  minmax_xy_kernel<T><<<g1, 256, 0, st>>>(x, y, z, n, mm);
  PTGS_LAUNCH_CHECK();
  unsigned long long h[6];
  // synchronises: the grid dimensions size the sort and the cell table
  PTGS_CUDA_TRY(cudaMemcpyAsync(h, mm, sizeof(h), cudaMemcpyDeviceToHost, st));
  PTGS_CUDA_TRY(cudaStreamSynchronize(st));
  GridSpec g;
  g.xmin = KeyOf<double>::decode(h[0]);
  g.ymin = KeyOf<double>::decode(h[1]);
  const double xmax = KeyOf<double>::decode(h[2]), ymax = KeyOf<double>::decode(h[3]);
  g.zmin = KeyOf<double>::decode(h[4]);
  const double zmax = KeyOf<double>::decode(h[5]);
  const double cell = radius * (1.0 + 1.0 / 8192.0);
  g.inv_cell = 1.0 / cell;
  g.nx = (long long)floor((xmax - g.xmin) * g.inv_cell) + 1;
  g.ny = (long long)floor((ymax - g.ymin) * g.inv_cell) + 1;
  if (g.nx < 1 || g.ny < 1 || (double)g.nx * (double)g.ny > 9.0e18) {
    set_error("radius_outlier: grid %lld x %lld out of range", g.nx, g.ny);
    return PTGS_E_RANGE;
  }
  if ((double)g.nx * (double)g.ny >= 2147483648.0) {
    set_error("radius_outlier: grid %lld x %lld has more than 2^31 columns", g.nx, g.ny);
    return PTGS_E_RANGE;
  }
  int colbits = 1;
  while (colbits < 32 && ((unsigned long long)(g.nx * g.ny - 1) >> colbits)) ++colbits;
  const uint64_t ncell = (uint64_t)g.nx * (uint64_t)g.ny;
  const bool table = ncell <= cell_table_capacity(n);
  // narrow keys: column | z bucket in 32 bits.  Buckets are 2^zshift cells tall; the coarser they
  // are the more candidates a query walks, so at most 4 cells (the slab then spans <= 12 of them)
  g.nzf = 1; g.zbits = 1; g.zshift = 0;
  bool narrow = false;
  {
    const double nz = floor((zmax - g.zmin) * g.inv_cell) + 1.0;
    if (table && nz >= 1.0 && nz < 1.0e9) {
      g.nzf = (long long)nz;
      for (int sh = 0; sh <= 2 && !narrow; ++sh) {
        int zb = 1;
        while (zb < 31 && (((unsigned long long)(g.nzf - 1) >> sh) >> zb)) ++zb;
        if (colbits + zb <= 32) { narrow = true; g.zshift = sh; g.zbits = zb; }
      }
    }
  }
  const int nthreads_blocks = (int)((n + 255) / 256);
  if (narrow) {
    uint32_t* k32a = reinterpret_cast<uint32_t*>(keys_a);
    uint32_t* k32b = reinterpret_cast<uint32_t*>(keys_b);
    const int bits = colbits + g.zbits;
    column_key_kernel<T, uint32_t><<<g1, 256, 0, st>>>(x, y, z, n, g, k32a, idx_a);
    PTGS_LAUNCH_CHECK();
    PTGS_CUDA_TRY(cub::DeviceRadixSort::SortPairs(sort_tmp, tb, k32a, k32b, idx_a, idx_b, (int64_t)n, 0, bits, st));
    count_launch(2 + (bits + 7) / 8);
    gather_sorted_kernel<T, uint32_t><<<g1, 256, 0, st>>>(x, y, z, aos, idx_b, k32b, (1u << g.zbits) - 1u, n, sp);
    PTGS_LAUNCH_CHECK();
    cell_start_kernel<uint32_t><<<nthreads_blocks, 256, 0, st>>>(k32b, n, g.zbits, ncell, cell_start);
    PTGS_LAUNCH_CHECK();
    radius_count_cells_kernel<T, uint32_t><<<nthreads_blocks, 256, 0, st>>>(
        sp, k32b, idx_b, cell_start, n, g, radius * radius, nb, keep, (unsigned long long*)n_keep);
    PTGS_LAUNCH_CHECK();
    return PTGS_OK;
  }
  const int bits = colbits + 32;             // column | height key
  column_key_kernel<T, uint64_t><<<g1, 256, 0, st>>>(x, y, z, n, g, keys_a, idx_a);
  PTGS_LAUNCH_CHECK();
  PTGS_CUDA_TRY(cub::DeviceRadixSort::SortPairs(sort_tmp, tb, keys_a, keys_b, idx_a, idx_b,
                                                (int64_t)n, 0, bits, st));
  count_launch(2 + (bits + 7) / 8);
  gather_sorted_kernel<T, uint64_t><<<g1, 256, 0, st>>>(x, y, z, aos, idx_b, keys_b, 0xffffffffu, n, sp);
  PTGS_LAUNCH_CHECK();
  if (table) {
    cell_start_kernel<uint64_t><<<nthreads_blocks, 256, 0, st>>>(keys_b, n, 32, ncell, cell_start);
    PTGS_LAUNCH_CHECK();
    radius_count_cells_kernel<T, uint64_t><<<nthreads_blocks, 256, 0, st>>>(
        sp, keys_b, idx_b, cell_start, n, g, radius * radius, nb, keep,
        (unsigned long long*)n_keep);
  } else {   // very sparse grid (far outliers): binary search per column
    radius_count_kernel<T><<<nthreads_blocks, 256, 0, st>>>(
        sp, keys_b, idx_b, n, g, radius * radius, nb, keep,
        (unsigned long long*)n_keep);
  }
  PTGS_LAUNCH_CHECK();
  return PTGS_OK;
}

}  // namespace ptgs

using namespace ptgs;

extern "C" {

size_t ptgs_slab_select_workspace_bytes(int64_t n) {
  int64_t tiles = (n + kSlabTile - 1) / kSlabTile + 1;
  return align_up(tiles * 4, 256) + align_up(tiles * 8, 256) + 256;
}

int ptgs_slab_select(int dtype, const void* x, const void* y, const void* z, int64_t n, double lo,
                     double hi, void* out_x, void* out_y, void* out_z, int32_t* out_index,
                     int64_t* n_out, void* workspace, size_t workspace_bytes, ptgs_stream_t stream) {
  PTGS_REQUIRE(n >= 0 && n < (1ll << 31) && n_out, "arguments");
  PTGS_REQUIRE(dtype == PTGS_F32 || dtype == PTGS_F64, "dtype");
  if (workspace_bytes < ptgs_slab_select_workspace_bytes(n)) {
    set_error("slab_select: workspace too small");
    return PTGS_E_WORKSPACE;
  }
  cudaStream_t st = (cudaStream_t)stream;
  const int64_t tiles = (n + kSlabTile - 1) / kSlabTile;
  if (tiles == 0) {
    PTGS_CUDA_TRY(cudaMemsetAsync(n_out, 0, 8, st));
    return PTGS_OK;
  }
  Carver cv(workspace);
  int32_t* tile_count = cv.take<int32_t>(tiles);
  int64_t* tile_base = cv.take<int64_t>(tiles);
  if (dtype == PTGS_F32) slab_count_kernel<float><<<(int)tiles, 256, 0, st>>>((const float*)x, n, lo, hi, tile_count);
  else slab_count_kernel<double><<<(int)tiles, 256, 0, st>>>((const double*)x, n, lo, hi, tile_count);
  PTGS_LAUNCH_CHECK();
  slab_scan_kernel<<<1, 1024, 0, st>>>(tile_count, tiles, tile_base, n_out);
  PTGS_LAUNCH_CHECK();
  if (dtype == PTGS_F32)
    slab_scatter_kernel<float><<<(int)tiles, 256, 0, st>>>((const float*)x, (const float*)y, (const float*)z, n, lo, hi,
                                                          tile_base, (float*)out_x, (float*)out_y, (float*)out_z, out_index);
  else
    slab_scatter_kernel<double><<<(int)tiles, 256, 0, st>>>((const double*)x, (const double*)y, (const double*)z, n, lo,
                                                           hi, tile_base, (double*)out_x, (double*)out_y, (double*)out_z,
                                                           out_index);
  PTGS_LAUNCH_CHECK();
  return PTGS_OK;
}

int ptgs_slab_keep(int dtype, const void* slab_x, const int32_t* slab_index, const uint8_t* slab_keep,
                   int64_t m, double lo, double hi, uint8_t* keep, ptgs_stream_t stream) {
  PTGS_REQUIRE(m >= 0 && (dtype == PTGS_F32 || dtype == PTGS_F64), "arguments");
  if (m == 0) return PTGS_OK;
  PTGS_REQUIRE(slab_x && slab_index && slab_keep && keep, "arguments");
  const int grid = grid_for(m, 256, kNumSMs * 8);
  if (dtype == PTGS_F32)
    slab_keep_kernel<float><<<grid, 256, 0, (cudaStream_t)stream>>>((const float*)slab_x, slab_index, slab_keep, m, lo, hi, keep);
  else
    slab_keep_kernel<double><<<grid, 256, 0, (cudaStream_t)stream>>>((const double*)slab_x, slab_index, slab_keep, m, lo, hi, keep);
  PTGS_LAUNCH_CHECK();
  return PTGS_OK;
}

size_t ptgs_radius_outlier_workspace_bytes(int64_t n) {
  if (n < 1) n = 1;
  size_t b = 256;
  b += 2 * align_up((size_t)n * 8, 256);      // keys
  b += 2 * align_up((size_t)n * 4, 256);      // idx
  b += align_up((size_t)n * 32, 256);         // sorted packed points (double4 worst case)
  b += align_up(sort_temp_bytes(n), 256) + 4096;
  b += align_up((size_t)(cell_table_capacity(n) + 2) * 4, 256);   // cell table
  return b;
}

int ptgs_radius_outlier_mask(int dtype, const void* x, const void* y,
                             const void* z, const void* xyz_aos, int64_t n, double radius,
                             int nb_points, uint8_t* keep, int64_t* n_keep,
                             void* workspace, size_t workspace_bytes,
                             ptgs_stream_t stream) {
  PTGS_REQUIRE(n >= 0 && n < (1ll << 32), "n");
  PTGS_REQUIRE(radius > 0.0 && nb_points >= 0, "radius / nb_points");
  PTGS_REQUIRE(n_keep && (keep || n == 0), "outputs");
  if (n == 0) {
    PTGS_CUDA_TRY(cudaMemsetAsync(n_keep, 0, 8, (cudaStream_t)stream));
    return PTGS_OK;
  }
  if (workspace_bytes < ptgs_radius_outlier_workspace_bytes(n)) {
    set_error("radius_outlier: workspace too small");
    return PTGS_E_WORKSPACE;
  }
  if (dtype == PTGS_F32)
    return radius_impl<float>(x, y, z, xyz_aos, n, radius, nb_points, keep, n_keep, workspace, (cudaStream_t)stream);
  if (dtype == PTGS_F64)
    return radius_impl<double>(x, y, z, xyz_aos, n, radius, nb_points, keep, n_keep, workspace, (cudaStream_t)stream);
  PTGS_REQUIRE(false, "dtype");
}

}  // extern "C"
