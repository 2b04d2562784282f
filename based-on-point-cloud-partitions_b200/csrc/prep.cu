// libptgs: error state, layout helpers (AoS->SoA), Manhattan transform (A13),
// XY bounds (A2) and keep-mask compaction.
#include <stdarg.h>
#include <string.h>

#include <mutex>

#include "common.cuh"

namespace ptgs {

static thread_local char g_err[512] = "";
static thread_local int64_t g_launches = 0;

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}
void count_launch(int n) { g_launches += n; }

// ---------------------------------------------------------------------------
// AoS (N,3) f32 -> planar.  Each thread moves 4 points = 3 float4 loads.
// ---------------------------------------------------------------------------
__global__ void __launch_bounds__(256) aos_to_soa_kernel(
    const float* __restrict__ xyz, int64_t n, float* __restrict__ x,
    float* __restrict__ y, float* __restrict__ z) {
  const int64_t n4 = n / 4;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  const float4* in = reinterpret_cast<const float4*>(xyz);
  for (int64_t g = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; g < n4;
       g += stride) {
    float4 a = __ldg(in + 3 * g), b = __ldg(in + 3 * g + 1),
           c = __ldg(in + 3 * g + 2);
    reinterpret_cast<float4*>(x)[g] = make_float4(a.x, a.w, b.z, c.y);
    reinterpret_cast<float4*>(y)[g] = make_float4(a.y, b.x, b.w, c.z);
    reinterpret_cast<float4*>(z)[g] = make_float4(a.z, b.y, c.x, c.w);
  }
  // tail
  if (blockIdx.x == 0 && threadIdx.x < (n & 3)) {
    int64_t i = n4 * 4 + threadIdx.x;
    x[i] = xyz[3 * i];
    y[i] = xyz[3 * i + 1];
    z[i] = xyz[3 * i + 2];
  }
}

// float64 rows (Manhattan-rotated clouds): 24-byte rows, one double2 + one
// double load per point would misalign every other row, so each thread moves
// 2 points = 3 aligned double2 loads.
__global__ void __launch_bounds__(256) aos_to_soa_f64_kernel(
    const double* __restrict__ xyz, int64_t n, double* __restrict__ x,
    double* __restrict__ y, double* __restrict__ z) {
  const int64_t n2 = n / 2;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  const double2* in = reinterpret_cast<const double2*>(xyz);
  for (int64_t g = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; g < n2;
       g += stride) {
    const double2 a = __ldg(in + 3 * g), b = __ldg(in + 3 * g + 1),
                  c = __ldg(in + 3 * g + 2);
    reinterpret_cast<double2*>(x)[g] = make_double2(a.x, b.y);
    reinterpret_cast<double2*>(y)[g] = make_double2(a.y, c.x);
    reinterpret_cast<double2*>(z)[g] = make_double2(b.x, c.y);
  }
  if (blockIdx.x == 0 && threadIdx.x == 0 && (n & 1)) {
    const int64_t i = n - 1;
    x[i] = xyz[3 * i];
    y[i] = xyz[3 * i + 1];
    z[i] = xyz[3 * i + 2];
  }
}

// ---------------------------------------------------------------------------
// Manhattan transform, f32 -> f64 (k-ordered fma accumulation, then + T).
// ---------------------------------------------------------------------------
struct Man {
  double r[9];
  double t[3];
};
__global__ void __launch_bounds__(256) manhattan_kernel(
    const float* __restrict__ x, const float* __restrict__ y,
    const float* __restrict__ z, int64_t n, Man m, double* __restrict__ ox,
    double* __restrict__ oy, double* __restrict__ oz) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n;
       i += stride) {
    double px = x[i], py = y[i], pz = z[i];
    double a = __dmul_rn(m.r[0], px);
    a = fma(m.r[1], py, a);
    a = fma(m.r[2], pz, a);
    double b = __dmul_rn(m.r[3], px);
    b = fma(m.r[4], py, b);
    b = fma(m.r[5], pz, b);
    double c = __dmul_rn(m.r[6], px);
    c = fma(m.r[7], py, c);
    c = fma(m.r[8], pz, c);
    ox[i] = __dadd_rn(a, m.t[0]);
    oy[i] = __dadd_rn(b, m.t[1]);
    oz[i] = __dadd_rn(c, m.t[2]);
  }
}

// AoS float32 -> planar float64 with the Manhattan transform applied on the way
// (layout + A13 in ONE pass: 12 B read + 24 B written per point instead of the
// 24 + 36 B of aos_to_soa followed by manhattan_kernel).  Same arithmetic as
// manhattan_kernel.  4 points per thread: 3 float4 loads, 6 double2 stores.
__device__ __forceinline__ void man_point(const Man& m, float fx, float fy, float fz, double& ox,
                                          double& oy, double& oz) {
  const double px = fx, py = fy, pz = fz;
  double a = __dmul_rn(m.r[0], px);
  a = fma(m.r[1], py, a);
  a = fma(m.r[2], pz, a);
  double b = __dmul_rn(m.r[3], px);
  b = fma(m.r[4], py, b);
  b = fma(m.r[5], pz, b);
  double c = __dmul_rn(m.r[6], px);
  c = fma(m.r[7], py, c);
  c = fma(m.r[8], pz, c);
  ox = __dadd_rn(a, m.t[0]);
  oy = __dadd_rn(b, m.t[1]);
  oz = __dadd_rn(c, m.t[2]);
}

__global__ void __launch_bounds__(256) aos_manhattan_kernel(
    const float* __restrict__ xyz, int64_t n, Man m, double* __restrict__ ox,
    double* __restrict__ oy, double* __restrict__ oz) {
  const int64_t n4 = n / 4;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  const float4* in = reinterpret_cast<const float4*>(xyz);
  for (int64_t g = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; g < n4; g += stride) {
    const float4 a = __ldg(in + 3 * g), b = __ldg(in + 3 * g + 1), c = __ldg(in + 3 * g + 2);
    double x0, y0, z0, x1, y1, z1, x2, y2, z2, x3, y3, z3;
    man_point(m, a.x, a.y, a.z, x0, y0, z0);
    man_point(m, a.w, b.x, b.y, x1, y1, z1);
    man_point(m, b.z, b.w, c.x, x2, y2, z2);
    man_point(m, c.y, c.z, c.w, x3, y3, z3);
    double2* px = reinterpret_cast<double2*>(ox) + 2 * g;
    double2* py = reinterpret_cast<double2*>(oy) + 2 * g;
    double2* pz = reinterpret_cast<double2*>(oz) + 2 * g;
    px[0] = make_double2(x0, x1); px[1] = make_double2(x2, x3);
    py[0] = make_double2(y0, y1); py[1] = make_double2(y2, y3);
    pz[0] = make_double2(z0, z1); pz[1] = make_double2(z2, z3);
  }
  if (blockIdx.x == 0 && threadIdx.x < (n & 3)) {
    const int64_t i = n4 * 4 + threadIdx.x;
    man_point(m, xyz[3 * i], xyz[3 * i + 1], xyz[3 * i + 2], ox[i], oy[i], oz[i]);
  }
}

// ---------------------------------------------------------------------------
// XY bounds: block reduce in registers/shuffles, one atomic per block on
// order-preserving keys.
// ---------------------------------------------------------------------------
template <typename T>
__global__ void bounds_init_kernel(typename KeyOf<T>::type* keys) {
  using K = typename KeyOf<T>::type;
  keys[0] = ~K(0);
  keys[1] = ~K(0);
  keys[2] = K(0);
  keys[3] = K(0);
}

__device__ __forceinline__ void atomic_min_key(uint32_t* a, uint32_t v) { atomicMin(a, v); }
__device__ __forceinline__ void atomic_max_key(uint32_t* a, uint32_t v) { atomicMax(a, v); }
__device__ __forceinline__ void atomic_min_key(uint64_t* a, uint64_t v) {
  atomicMin(reinterpret_cast<unsigned long long*>(a), (unsigned long long)v);
}
__device__ __forceinline__ void atomic_max_key(uint64_t* a, uint64_t v) {
  atomicMax(reinterpret_cast<unsigned long long*>(a), (unsigned long long)v);
}

template <typename T>
__global__ void __launch_bounds__(256) bounds_kernel(
    const T* __restrict__ x, const T* __restrict__ y,
    const int64_t* __restrict__ n_dev, typename KeyOf<T>::type* keys) {
  using K = typename KeyOf<T>::type;
  const int64_t n = *n_dev;
  K xmin = ~K(0), ymin = ~K(0), xmax = 0, ymax = 0;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n;
       i += stride) {
    K kx = KeyOf<T>::encode(x[i]), ky = KeyOf<T>::encode(y[i]);
    xmin = min(xmin, kx);
    xmax = max(xmax, kx);
    ymin = min(ymin, ky);
    ymax = max(ymax, ky);
  }
#pragma unroll
  for (int d = 16; d > 0; d >>= 1) {
    xmin = min(xmin, __shfl_xor_sync(kFullMask, xmin, d));
    ymin = min(ymin, __shfl_xor_sync(kFullMask, ymin, d));
    xmax = max(xmax, __shfl_xor_sync(kFullMask, xmax, d));
    ymax = max(ymax, __shfl_xor_sync(kFullMask, ymax, d));
  }
  if (lane_id() == 0 && xmin <= xmax) {
    atomic_min_key(keys + 0, xmin);
    atomic_min_key(keys + 1, ymin);
    atomic_max_key(keys + 2, xmax);
    atomic_max_key(keys + 3, ymax);
  }
}

template <typename T>
__global__ void bounds_finish_kernel(const typename KeyOf<T>::type* keys,
                                     double* bounds) {
  typename KeyOf<T>::type k[4];
  for (int i = 0; i < 4; ++i) k[i] = keys[i];          // keys alias bounds
  for (int i = 0; i < 4; ++i) bounds[i] = (double)KeyOf<T>::decode(k[i]);
}

// ---------------------------------------------------------------------------
// keep-mask compaction of (x, y): per-tile counts -> scan -> scatter.
// ---------------------------------------------------------------------------
constexpr int kCompactTile = 2048;  // 256 threads x 8

__global__ void __launch_bounds__(256) compact_count_kernel(
    const uint8_t* __restrict__ keep, int64_t n, int32_t* __restrict__ tile_count) {
  __shared__ int s[33];
  const int64_t base = (int64_t)blockIdx.x * kCompactTile;
  int c = 0;
#pragma unroll
  for (int k = 0; k < 8; ++k) {
    int64_t i = base + k * 256 + threadIdx.x;
    c += (i < n && keep[i]) ? 1 : 0;
  }
  int tot;
  block_exclusive_sum<int, 256>(c, s, &tot);
  if (threadIdx.x == 0) tile_count[blockIdx.x] = tot;
}

// single-block exclusive scan of tile counts (n_tiles up to a few 100k).
__global__ void __launch_bounds__(1024) scan_tiles_kernel(
    const int32_t* __restrict__ tile_count, int64_t n_tiles,
    int64_t* __restrict__ tile_base, int64_t* __restrict__ total) {
  __shared__ long long s[33];
  long long carry = 0;
  for (int64_t start = 0; start < n_tiles; start += 1024) {
    int64_t i = start + threadIdx.x;
    long long v = (i < n_tiles) ? tile_count[i] : 0;
    long long tot;
    long long ex = block_exclusive_sum<long long, 1024>(v, s, &tot);
    if (i < n_tiles) tile_base[i] = carry + ex;
    carry += tot;
  }
  if (threadIdx.x == 0) *total = carry;
}

template <typename T>
__global__ void __launch_bounds__(256) compact_scatter_kernel(
    const T* __restrict__ x, const T* __restrict__ y,
    const uint8_t* __restrict__ keep, int64_t n,
    const int64_t* __restrict__ tile_base, T* __restrict__ ox,
    T* __restrict__ oy) {
  __shared__ int s[33];
  const int64_t base = (int64_t)blockIdx.x * kCompactTile;
  int64_t out = tile_base[blockIdx.x];
  // index order: element (k, tid) -> base + k*256 + tid ; scan per k keeps it.
#pragma unroll 1
  for (int k = 0; k < 8; ++k) {
    int64_t i = base + k * 256 + threadIdx.x;
    int f = (i < n && keep[i]) ? 1 : 0;
    int tot;
    int ex = block_exclusive_sum<int, 256>(f, s, &tot);
    if (f) {
      ox[out + ex] = x[i];
      oy[out + ex] = y[i];
    }
    out += tot;
  }
}

// ---------------------------------------------------------------------------
// N3: COLMAP points3D.bin on the fast path (reference: the per-record struct
// loop of scene/colmap_loader.py:195-224).  Records are 51 bytes + 8 bytes per
// track element; the host walks the record offsets once (ptgs_points3d_index),
// the device decodes xyz / rgb / error straight from the uploaded file image.
// ---------------------------------------------------------------------------
__device__ __forceinline__ double load_f64_unaligned(const uint8_t* p) {
  unsigned long long v = 0;
#pragma unroll
  for (int b = 0; b < 8; ++b) v |= (unsigned long long)p[b] << (8 * b);
  return __longlong_as_double((long long)v);
}

__global__ void __launch_bounds__(256) points3d_decode_kernel(
    const uint8_t* __restrict__ buf, const int64_t* __restrict__ offsets, int64_t n,
    float* __restrict__ x, float* __restrict__ y, float* __restrict__ z,
    uint8_t* __restrict__ rgb, double* __restrict__ error) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    const uint8_t* r = buf + offsets[i];
    // id u64 | xyz 3 x f64 | rgb 3 x u8 | error f64 | track length u64
    x[i] = (float)load_f64_unaligned(r + 8);      // storePly's float32 cast (shen_data_read.py:135-148)
    y[i] = (float)load_f64_unaligned(r + 16);
    z[i] = (float)load_f64_unaligned(r + 24);
    if (rgb) { rgb[3 * i] = r[32]; rgb[3 * i + 1] = r[33]; rgb[3 * i + 2] = r[34]; }
    if (error) error[i] = load_f64_unaligned(r + 35);
  }
}

static inline int grid_for(int64_t n, int per_block, int max_blocks) {
  int64_t b = (n + per_block - 1) / per_block;
  if (b < 1) b = 1;
  if (b > max_blocks) b = max_blocks;
  return (int)b;
}

template <typename T>
static int bounds_impl(const void* x, const void* y, const int64_t* n_dev,
                       int64_t n_capacity, double* bounds, cudaStream_t st) {
  using K = typename KeyOf<T>::type;
  // the 4 ordered keys (<= 32 B) are accumulated in the caller's 32-byte
  // output buffer and decoded in place by one thread: no scratch allocation.
  K* keys = reinterpret_cast<K*>(bounds);
  bounds_init_kernel<T><<<1, 1, 0, st>>>(keys);
  PTGS_LAUNCH_CHECK();
  int grid = grid_for(n_capacity, 256 * 8, kNumSMs * 8);
  bounds_kernel<T><<<grid, 256, 0, st>>>((const T*)x, (const T*)y, n_dev, keys);
  PTGS_LAUNCH_CHECK();
  bounds_finish_kernel<T><<<1, 1, 0, st>>>(keys, bounds);
  PTGS_LAUNCH_CHECK();
  return PTGS_OK;
}

// ---- allocation-free stream compaction (ptgs_compact_xy_static) ------------------------------
// The per-tile counts and bases of compact_count / scan_tiles / compact_scatter live in a
// file-scope device array instead of a caller workspace.  Every GPU that loads the library holds
// the array (384 KB) until exit.  In host code the NAME below is a host-side stand-in, not a device
// address: the address a kernel can use comes from cudaGetSymbolAddress, looked up once per device
// with that device current.
constexpr int64_t kStaticCompactMaxPoints = PTGS_COMPACT_XY_STATIC_MAX_POINTS;
constexpr int64_t kStaticCompactMaxTiles = kStaticCompactMaxPoints / kCompactTile;
constexpr size_t kStaticCountBytes = (size_t)kStaticCompactMaxTiles * sizeof(int32_t);   // 128 KB, multiple of 256
constexpr size_t kStaticCompactBytes = kStaticCountBytes + (size_t)kStaticCompactMaxTiles * sizeof(int64_t);
static_assert(kStaticCompactMaxPoints % kCompactTile == 0 && kStaticCountBytes % 256 == 0, "scratch layout");
static __device__ __align__(256) unsigned char g_compact_static_scratch[kStaticCompactBytes];

constexpr int kMaxDevices = 64;
static std::mutex g_static_mu;
static void* g_static_addr[kMaxDevices] = {};

// device address of the scratch array on the CURRENT device (set-up: first call per device)
static int compact_static_address(unsigned char** out) {
  int dev = -1;
  PTGS_CUDA_TRY(cudaGetDevice(&dev));
  PTGS_REQUIRE(dev >= 0 && dev < kMaxDevices, "device ordinal above 63");
  std::lock_guard<std::mutex> lock(g_static_mu);
  if (!g_static_addr[dev]) {
    void* ptr = nullptr;
    PTGS_CUDA_TRY(cudaGetSymbolAddress(&ptr, g_compact_static_scratch));
    PTGS_REQUIRE(ptr != nullptr, "cudaGetSymbolAddress returned a null address");
    g_static_addr[dev] = ptr;
  }
  *out = static_cast<unsigned char*>(g_static_addr[dev]);
  return PTGS_OK;
}

}  // namespace ptgs

using namespace ptgs;

extern "C" {

int ptgs_version(void) { return 200; }
int ptgs_extreme_dirs(void) { return PTGS_EXTREME_DIRS; }
const char* ptgs_last_error(void) { return g_err; }
int64_t ptgs_launch_count(void) { return g_launches; }
void ptgs_reset_launch_count(void) { g_launches = 0; }

int ptgs_aos_to_soa_f32(const float* xyz, int64_t n, float* x, float* y,
                        float* z, ptgs_stream_t stream) {
  PTGS_REQUIRE(n >= 0, "n");
  if (n == 0) return PTGS_OK;
  PTGS_REQUIRE(((uintptr_t)xyz % 16 == 0) && ((uintptr_t)x % 16 == 0) &&
                   ((uintptr_t)y % 16 == 0) && ((uintptr_t)z % 16 == 0),
               "pointers must be 16-byte aligned");
  int grid = grid_for(n / 4 + 1, 256, kNumSMs * 8);
  aos_to_soa_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(xyz, n, x, y, z);
  PTGS_LAUNCH_CHECK();
  return PTGS_OK;
}

int ptgs_aos_to_soa_f64(const double* xyz, int64_t n, double* x, double* y,
                        double* z, ptgs_stream_t stream) {
  PTGS_REQUIRE(n >= 0, "n");
  if (n == 0) return PTGS_OK;
  PTGS_REQUIRE(((uintptr_t)xyz % 16 == 0) && ((uintptr_t)x % 16 == 0) &&
                   ((uintptr_t)y % 16 == 0) && ((uintptr_t)z % 16 == 0),
               "pointers must be 16-byte aligned");
  int grid = grid_for(n / 2 + 1, 256, kNumSMs * 8);
  aos_to_soa_f64_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(xyz, n, x, y, z);
  PTGS_LAUNCH_CHECK();
  return PTGS_OK;
}

int ptgs_manhattan_apply(const float* x, const float* y, const float* z,
                         int64_t n, const double* man_trans_host, double* ox,
                         double* oy, double* oz, ptgs_stream_t stream) {
  PTGS_REQUIRE(n >= 0 && man_trans_host, "n / man_trans");
  if (n == 0) return PTGS_OK;
  Man m;
  for (int r = 0; r < 3; ++r) {
    for (int c = 0; c < 3; ++c) m.r[r * 3 + c] = man_trans_host[r * 4 + c];
    m.t[r] = man_trans_host[r * 4 + 3];
  }
  int grid = grid_for(n, 256, kNumSMs * 8);
  manhattan_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(x, y, z, n, m, ox, oy, oz);
  PTGS_LAUNCH_CHECK();
  return PTGS_OK;
}

int ptgs_aos_to_soa_manhattan(const float* xyz, int64_t n, const double* man_trans_host,
                              double* ox, double* oy, double* oz, ptgs_stream_t stream) {
  PTGS_REQUIRE(n >= 0 && man_trans_host, "n / man_trans");
  if (n == 0) return PTGS_OK;
  PTGS_REQUIRE(((uintptr_t)xyz % 16 == 0) && ((uintptr_t)ox % 16 == 0) &&
                   ((uintptr_t)oy % 16 == 0) && ((uintptr_t)oz % 16 == 0),
               "pointers must be 16-byte aligned");
  Man m;
  for (int r = 0; r < 3; ++r) {
    for (int c = 0; c < 3; ++c) m.r[r * 3 + c] = man_trans_host[r * 4 + c];
    m.t[r] = man_trans_host[r * 4 + 3];
  }
  int grid = grid_for(n / 4 + 1, 256, kNumSMs * 8);
  aos_manhattan_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(xyz, n, m, ox, oy, oz);
  PTGS_LAUNCH_CHECK();
  return PTGS_OK;
}

int64_t ptgs_points3d_index(const uint8_t* file_host, int64_t n_bytes, int64_t* offsets_host,
                            int64_t capacity) {
  // HOST function: walks the variable-length records; returns the record count or a negative status
  if (!file_host || n_bytes < 8) { set_error("points3d_index: truncated file"); return PTGS_E_INVALID; }
  uint64_t n;
  memcpy(&n, file_host, 8);
  if ((int64_t)n < 0 || (offsets_host && (int64_t)n > capacity)) {
    set_error("points3d_index: %llu records exceed the capacity", (unsigned long long)n);
    return PTGS_E_CAPACITY;
  }
  int64_t off = 8;
  for (uint64_t i = 0; i < n; ++i) {
    if (off + 51 > n_bytes) { set_error("points3d_index: truncated record %llu", (unsigned long long)i); return PTGS_E_INVALID; }
    if (offsets_host) offsets_host[i] = off;
    uint64_t tl;
    memcpy(&tl, file_host + off + 43, 8);
    if (tl > (uint64_t)(n_bytes - off - 51) / 8) { set_error("points3d_index: truncated track %llu", (unsigned long long)i); return PTGS_E_INVALID; }
    off += 51 + 8 * (int64_t)tl;
  }
  return (int64_t)n;
}

int ptgs_points3d_decode(const uint8_t* file_dev, const int64_t* offsets_dev, int64_t n, float* x,
                         float* y, float* z, uint8_t* rgb, double* error, ptgs_stream_t stream) {
  PTGS_REQUIRE(n >= 0, "n");
  if (n == 0) return PTGS_OK;
  PTGS_REQUIRE(file_dev && offsets_dev && x && y && z, "arguments");
  int grid = grid_for(n, 256, kNumSMs * 8);
  points3d_decode_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(file_dev, offsets_dev, n, x, y, z, rgb, error);
  PTGS_LAUNCH_CHECK();
  return PTGS_OK;
}

int ptgs_bounds_xy(int dtype, const void* x, const void* y,
                   const int64_t* n_dev, int64_t n_capacity, double* bounds,
                   ptgs_stream_t stream) {
  PTGS_REQUIRE(n_capacity >= 0 && n_dev && bounds, "arguments");
  if (dtype == PTGS_F32)
    return bounds_impl<float>(x, y, n_dev, n_capacity, bounds, (cudaStream_t)stream);
  if (dtype == PTGS_F64)
    return bounds_impl<double>(x, y, n_dev, n_capacity, bounds, (cudaStream_t)stream);
  PTGS_REQUIRE(false, "dtype");
}

size_t ptgs_compact_xy_workspace_bytes(int64_t n) {
  int64_t tiles = (n + kCompactTile - 1) / kCompactTile + 1;
  return align_up(tiles * 4, 256) + align_up(tiles * 8, 256) + 256;
}

int ptgs_compact_xy(int dtype, const void* x, const void* y,
                    const uint8_t* keep, int64_t n, void* out_x, void* out_y,
                    int64_t* n_out, void* workspace, size_t workspace_bytes,
                    ptgs_stream_t stream) {
  PTGS_REQUIRE(n >= 0 && n_out, "arguments");
  PTGS_REQUIRE(dtype == PTGS_F32 || dtype == PTGS_F64, "dtype");
  if (workspace_bytes < ptgs_compact_xy_workspace_bytes(n)) {
    set_error("compact_xy: workspace too small");
    return PTGS_E_WORKSPACE;
  }
  cudaStream_t st = (cudaStream_t)stream;
  int64_t tiles = (n + kCompactTile - 1) / kCompactTile;
  if (tiles == 0) {
    PTGS_CUDA_TRY(cudaMemsetAsync(n_out, 0, 8, st));
    return PTGS_OK;
  }
  Carver cv(workspace);
  int32_t* tile_count = cv.take<int32_t>(tiles);
  int64_t* tile_base = cv.take<int64_t>(tiles);
  compact_count_kernel<<<(int)tiles, 256, 0, st>>>(keep, n, tile_count);
  PTGS_LAUNCH_CHECK();
  scan_tiles_kernel<<<1, 1024, 0, st>>>(tile_count, tiles, tile_base, n_out);
  PTGS_LAUNCH_CHECK();
  if (dtype == PTGS_F32)
    compact_scatter_kernel<float><<<(int)tiles, 256, 0, st>>>(
        (const float*)x, (const float*)y, keep, n, tile_base, (float*)out_x, (float*)out_y);
  else
    compact_scatter_kernel<double><<<(int)tiles, 256, 0, st>>>(
        (const double*)x, (const double*)y, keep, n, tile_base, (double*)out_x, (double*)out_y);
  PTGS_LAUNCH_CHECK();
  return PTGS_OK;
}

int ptgs_compact_xy_static(int dtype, const void* x, const void* y, const uint8_t* keep, int64_t n,
                           void* out_x, void* out_y, int64_t* n_out, ptgs_stream_t stream) {
  PTGS_REQUIRE(n >= 0 && n_out, "arguments");
  PTGS_REQUIRE(dtype == PTGS_F32 || dtype == PTGS_F64, "dtype");
  if (n > kStaticCompactMaxPoints) {
    set_error("compact_xy_static: n = %lld exceeds the built-in scratch (PTGS_COMPACT_XY_STATIC_MAX_POINTS = %lld); "
              "use ptgs_compact_xy with a caller workspace",
              (long long)n, (long long)kStaticCompactMaxPoints);
    return PTGS_E_WORKSPACE;
  }
  cudaStream_t st = (cudaStream_t)stream;
  const int64_t tiles = (n + kCompactTile - 1) / kCompactTile;
  if (tiles == 0) {
    PTGS_CUDA_TRY(cudaMemsetAsync(n_out, 0, 8, st));
    return PTGS_OK;
  }
  PTGS_REQUIRE(x && y && keep && out_x && out_y, "arguments");
  unsigned char* scratch = nullptr;
  const int rc = compact_static_address(&scratch);
  if (rc != PTGS_OK) return rc;
  // entries [0, tiles) of both tables are written by the first two kernels before anything reads
  // them, so nothing has to start at zero and the previous call's contents do not matter
  int32_t* tile_count = reinterpret_cast<int32_t*>(scratch);
  int64_t* tile_base = reinterpret_cast<int64_t*>(scratch + kStaticCountBytes);
  compact_count_kernel<<<(int)tiles, 256, 0, st>>>(keep, n, tile_count);
  PTGS_LAUNCH_CHECK();
  scan_tiles_kernel<<<1, 1024, 0, st>>>(tile_count, tiles, tile_base, n_out);
  PTGS_LAUNCH_CHECK();
  if (dtype == PTGS_F32)
    compact_scatter_kernel<float><<<(int)tiles, 256, 0, st>>>(
        (const float*)x, (const float*)y, keep, n, tile_base, (float*)out_x, (float*)out_y);
  else
    compact_scatter_kernel<double><<<(int)tiles, 256, 0, st>>>(
        (const double*)x, (const double*)y, keep, n, tile_base, (double*)out_x, (double*)out_y);
  PTGS_LAUNCH_CHECK();
  return PTGS_OK;
}

}  // extern "C"
