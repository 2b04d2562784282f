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
};

template <typename T>
__global__ void __launch_bounds__(256) minmax_xy_kernel(
    const T* __restrict__ x, const T* __restrict__ y, int64_t n,
    unsigned long long* __restrict__ keys /*xmin,ymin,xmax,ymax as u64 keys of double*/) {
  using KD = KeyOf<double>;
  uint64_t xmin = ~0ull, ymin = ~0ull, xmax = 0, ymax = 0;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    uint64_t kx = KD::encode((double)x[i]), ky = KD::encode((double)y[i]);
    xmin = min(xmin, kx); xmax = max(xmax, kx);
    ymin = min(ymin, ky); ymax = max(ymax, ky);
  }
#pragma unroll
  for (int d = 16; d > 0; d >>= 1) {
    xmin = min(xmin, __shfl_xor_sync(kFullMask, xmin, d));
    ymin = min(ymin, __shfl_xor_sync(kFullMask, ymin, d));
    xmax = max(xmax, __shfl_xor_sync(kFullMask, xmax, d));
    ymax = max(ymax, __shfl_xor_sync(kFullMask, ymax, d));
  }
  if (lane_id() == 0) {
    atomicMin(keys + 0, (unsigned long long)xmin);
    atomicMin(keys + 1, (unsigned long long)ymin);
    atomicMax(keys + 2, (unsigned long long)xmax);
    atomicMax(keys + 3, (unsigned long long)ymax);
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

template <typename T>
__global__ void __launch_bounds__(256) column_key_kernel(
    const T* __restrict__ x, const T* __restrict__ y, const T* __restrict__ z, int64_t n, GridSpec g,
    uint64_t* __restrict__ keys, uint32_t* __restrict__ idx) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    long long ix = cell_of((double)x[i], g.xmin, g.inv_cell, g.nx);
    long long iy = cell_of((double)y[i], g.ymin, g.inv_cell, g.ny);
    keys[i] = ((uint64_t)(ix * g.ny + iy) << 32) | zkey_of((float)z[i]);   // float(z): monotone in z
    idx[i] = (uint32_t)i;
  }
}

template <typename T>
__global__ void __launch_bounds__(256) gather_sorted_kernel(
    const T* __restrict__ x, const T* __restrict__ y, const T* __restrict__ z,
    const uint32_t* __restrict__ idx, int64_t n, T* __restrict__ sx,
    T* __restrict__ sy, T* __restrict__ sz) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    uint32_t j = idx[i];
    sx[i] = x[j]; sy[i] = y[j]; sz[i] = z[j];
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
    const T* __restrict__ sx, const T* __restrict__ sy, const T* __restrict__ sz,
    const uint64_t* __restrict__ keys, const uint32_t* __restrict__ idx,
    int64_t n, GridSpec g, double r2, int nb, uint8_t* __restrict__ keep,
    unsigned long long* __restrict__ n_keep) {
  const int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  int kept = 0;
  if (p < n) {
    const double xi = (double)sx[p], yi = (double)sy[p], zi = (double)sz[p];
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
        double ddx = __dsub_rn((double)sx[j], xi);
        double ddy = __dsub_rn((double)sy[j], yi);
        double ddz = __dsub_rn((double)sz[j], zi);
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
__global__ void __launch_bounds__(256) cell_start_kernel(const uint64_t* __restrict__ keys, int64_t n,
                                                         uint64_t ncell, uint32_t* __restrict__ cs) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const uint64_t k = keys[i] >> 32;
  if (i == 0) {
    for (uint64_t c = 0; c <= k; ++c) cs[c] = 0u;
  } else {
    const uint64_t kp = keys[i - 1] >> 32;
    for (uint64_t c = kp + 1; c <= k; ++c) cs[c] = (uint32_t)i;
  }
  if (i == n - 1)
    for (uint64_t c = k + 1; c <= ncell; ++c) cs[c] = (uint32_t)n;
}

// Neighbour count with the cell table: the point's own cell first (most likely
// to hold close neighbours, so the `> nb_points` early exit comes soonest), then
// the 8 surrounding cells.
template <typename T>
__global__ void __launch_bounds__(256) radius_count_cells_kernel(
    const T* __restrict__ sx, const T* __restrict__ sy, const T* __restrict__ sz,
    const uint64_t* __restrict__ keys, const uint32_t* __restrict__ idx,
    const uint32_t* __restrict__ cs, int64_t n, GridSpec g, double r2, int nb,
    uint8_t* __restrict__ keep, unsigned long long* __restrict__ n_keep) {
  const int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  int kept = 0;
  if (p < n) {
    const double xi = (double)sx[p], yi = (double)sy[p], zi = (double)sz[p];
    const uint64_t key = keys[p] >> 32;
    const long long ix = (long long)(key / (uint64_t)g.ny);
    const long long iy = (long long)(key % (uint64_t)g.ny);
    // height slab in the (rounded-to-float) sort order: conservative by one float step
    const double rr = sqrt(r2);
    const uint32_t zlo = zkey_of(__double2float_rd(zi - rr)) - 1u;
    const uint32_t zhi = zkey_of(__double2float_ru(zi + rr)) + 1u;
    int cnt = 0;
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
        if ((uint32_t)__ldg(keys + mid) < zlo) lo = mid + 1; else hi = mid;
      }
      for (uint32_t j = lo; j < j1 && cnt <= nb; ++j) {
        if ((uint32_t)__ldg(keys + j) > zhi) break;
        double ddx = __dsub_rn((double)sx[j], xi);
        double ddy = __dsub_rn((double)sy[j], yi);
        double ddz = __dsub_rn((double)sz[j], zi);
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

static inline uint64_t cell_table_capacity(int64_t n) { return (uint64_t)4 * (uint64_t)n + (1ull << 20); }

static inline int grid_for(int64_t n, int per_block, int max_blocks) {
  int64_t b = (n + per_block - 1) / per_block;
  if (b < 1) b = 1;
  if (b > max_blocks) b = max_blocks;
  return (int)b;
}

static size_t sort_temp_bytes(int64_t n) {
  size_t bytes = 0;
  cub::DeviceRadixSort::SortPairs(nullptr, bytes, (const uint64_t*)nullptr, (uint64_t*)nullptr,
                                  (const uint32_t*)nullptr, (uint32_t*)nullptr, (int64_t)n, 0, 64);
  return bytes;
}

template <typename T>
static int radius_impl(const void* vx, const void* vy, const void* vz, int64_t n,
                       double radius, int nb, uint8_t* keep, int64_t* n_keep,
                       void* workspace, cudaStream_t st) {
  const T* x = (const T*)vx; const T* y = (const T*)vy; const T* z = (const T*)vz;
  Carver cv(workspace);
  unsigned long long* mm = cv.take<unsigned long long>(4);
  uint64_t* keys_a = cv.take<uint64_t>(n);
  uint64_t* keys_b = cv.take<uint64_t>(n);
  uint32_t* idx_a = cv.take<uint32_t>(n);
  uint32_t* idx_b = cv.take<uint32_t>(n);
  T* sx = cv.take<T>(n); T* sy = cv.take<T>(n); T* sz = cv.take<T>(n);
  size_t tb = sort_temp_bytes(n);
  void* sort_tmp = cv.take<char>(tb);
  uint32_t* cell_start = cv.take<uint32_t>(cell_table_capacity(n) + 2);

  const unsigned long long init[4] = {~0ull, ~0ull, 0ull, 0ull};
  PTGS_CUDA_TRY(cudaMemcpyAsync(mm, init, sizeof(init), cudaMemcpyHostToDevice, st));
  PTGS_CUDA_TRY(cudaMemsetAsync(n_keep, 0, 8, st));
  int g1 = grid_for(n, 2048, kNumSMs * 8);
  minmax_xy_kernel<T><<<g1, 256, 0, st>>>(x, y, n, mm);
  PTGS_LAUNCH_CHECK();
  unsigned long long h[4];
  PTGS_CUDA_TRY(cudaMemcpyAsync(h, mm, sizeof(h), cudaMemcpyDeviceToHost, st));
  PTGS_CUDA_TRY(cudaStreamSynchronize(st));
  GridSpec g;
  g.xmin = KeyOf<double>::decode(h[0]);
  g.ymin = KeyOf<double>::decode(h[1]);
  const double xmax = KeyOf<double>::decode(h[2]), ymax = KeyOf<double>::decode(h[3]);
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
  int bits = 1;
  while (bits < 32 && ((unsigned long long)(g.nx * g.ny - 1) >> bits)) ++bits;
  bits += 32;                                // column | height key
  column_key_kernel<T><<<g1, 256, 0, st>>>(x, y, z, n, g, keys_a, idx_a);
  PTGS_LAUNCH_CHECK();
  PTGS_CUDA_TRY(cub::DeviceRadixSort::SortPairs(sort_tmp, tb, keys_a, keys_b, idx_a, idx_b,
                                                (int64_t)n, 0, bits, st));
  count_launch(2 + (bits + 7) / 8);
  gather_sorted_kernel<T><<<g1, 256, 0, st>>>(x, y, z, idx_b, n, sx, sy, sz);
  PTGS_LAUNCH_CHECK();
  const uint64_t ncell = (uint64_t)g.nx * (uint64_t)g.ny;
  if (ncell <= cell_table_capacity(n)) {
    cell_start_kernel<<<(int)((n + 255) / 256), 256, 0, st>>>(keys_b, n, ncell, cell_start);
    PTGS_LAUNCH_CHECK();
    radius_count_cells_kernel<T><<<(int)((n + 255) / 256), 256, 0, st>>>(
        sx, sy, sz, keys_b, idx_b, cell_start, n, g, radius * radius, nb, keep,
        (unsigned long long*)n_keep);
  } else {   // very sparse grid (far outliers): binary search per column
    radius_count_kernel<T><<<(int)((n + 255) / 256), 256, 0, st>>>(
        sx, sy, sz, keys_b, idx_b, n, g, radius * radius, nb, keep,
        (unsigned long long*)n_keep);
  }
  PTGS_LAUNCH_CHECK();
  return PTGS_OK;
}

}  // namespace ptgs

using namespace ptgs;

extern "C" {

size_t ptgs_radius_outlier_workspace_bytes(int64_t n) {
  if (n < 1) n = 1;
  size_t b = 256;
  b += 2 * align_up((size_t)n * 8, 256);      // keys
  b += 2 * align_up((size_t)n * 4, 256);      // idx
  b += 3 * align_up((size_t)n * 8, 256);      // sorted coords (f64 worst case)
  b += align_up(sort_temp_bytes(n), 256) + 4096;
  b += align_up((size_t)(cell_table_capacity(n) + 2) * 4, 256);   // cell table
  return b;
}

int ptgs_radius_outlier_mask(int dtype, const void* x, const void* y,
                             const void* z, int64_t n, double radius,
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
    return radius_impl<float>(x, y, z, n, radius, nb_points, keep, n_keep, workspace, (cudaStream_t)stream);
  if (dtype == PTGS_F64)
    return radius_impl<double>(x, y, z, n, radius, nb_points, keep, n_keep, workspace, (cudaStream_t)stream);
  PTGS_REQUIRE(false, "dtype");
}

}  // extern "C"
