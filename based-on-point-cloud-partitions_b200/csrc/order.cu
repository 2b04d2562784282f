// libptgs: spatial (3-D Morton) ordering of a planar float32 cloud.
//
// The coverage sweep (config 5 / the Vast path's camera x all-points pass,
// scene/ptgs/data_partition.py:385-424, :505) gets its points in file order.
// The sweep culls cameras per 2048-point tile and per 256-point warp run, which
// only works when consecutive points are neighbours in space: this pass sorts
// the cloud by a 32-bit Morton-type code of a cubic-cell grid over its bounding
// box (2048 cells along the longest axis, one cell size for all axes) and returns the permutation, so
// per-point outputs can be written back in the caller's order.  The order only
// affects speed: hit counts and seen flags do not depend on it.
#include <cub/device/device_radix_sort.cuh>

#include "common.cuh"

namespace ptgs {

constexpr int kOrderKeyBits = 32;   // 4 radix passes on 32-bit keys

__global__ void __launch_bounds__(256) order_minmax_kernel(
    const float* __restrict__ x, const float* __restrict__ y, const float* __restrict__ z, int64_t n,
    uint32_t* __restrict__ mm /* min xyz, max xyz as order-preserving keys */) {
  const float inf = __int_as_float(0x7f800000);
  float lo[3] = {inf, inf, inf}, hi[3] = {-inf, -inf, -inf};
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    const float v[3] = {x[i], y[i], z[i]};
#pragma unroll
    for (int a = 0; a < 3; ++a) {
      if (fabsf(v[a]) < inf) {               // NaN / inf do not shape the grid
        lo[a] = fminf(lo[a], v[a]);
        hi[a] = fmaxf(hi[a], v[a]);
      }
    }
  }
#pragma unroll
  for (int d = 16; d > 0; d >>= 1) {
#pragma unroll
    for (int a = 0; a < 3; ++a) {
      lo[a] = fminf(lo[a], __shfl_xor_sync(kFullMask, lo[a], d));
      hi[a] = fmaxf(hi[a], __shfl_xor_sync(kFullMask, hi[a], d));
    }
  }
  if (lane_id() == 0) {
#pragma unroll
    for (int a = 0; a < 3; ++a) {
      atomicMin(mm + a, KeyOf<float>::encode(lo[a]));
      atomicMax(mm + 3 + a, KeyOf<float>::encode(hi[a]));
    }
  }
}

// bit i of v (10 bits) -> bit 3 i
__device__ __forceinline__ uint32_t spread3_10(uint32_t v) {
  uint32_t x = v & 0x3ffu;
  x = (x | (x << 16)) & 0x030000ffu;
  x = (x | (x << 8)) & 0x0300f00fu;
  x = (x | (x << 4)) & 0x030c30c3u;
  x = (x | (x << 2)) & 0x09249249u;
  return x;
}

// 32-bit key: cubic cells of (largest extent / 2048); 11 bits for the two longer axes, 10 for the
// shortest (clamped).  Top two bits = the high bits of the two longer axes, then a 3-D Morton code.
__global__ void __launch_bounds__(256) order_key_kernel(
    const float* __restrict__ x, const float* __restrict__ y, const float* __restrict__ z, int64_t n,
    const uint32_t* __restrict__ mm, uint32_t* __restrict__ keys, uint32_t* __restrict__ idx) {
  float lo[3], e[3], ext = 0.0f;
#pragma unroll
  for (int a = 0; a < 3; ++a) {
    lo[a] = KeyOf<float>::decode(mm[a]);
    e[a] = KeyOf<float>::decode(mm[3 + a]) - lo[a];
    if (!(e[a] >= 0.0f)) e[a] = 0.0f;        // empty axis (no finite value)
    ext = (e[a] > ext) ? e[a] : ext;
  }
  int small = 0;                              // shortest axis
  if (e[1] < e[small]) small = 1;
  if (e[2] < e[small]) small = 2;
  const int a0 = (small == 0) ? 1 : 0, a1 = (small == 2) ? 1 : 2;   // the two longer axes
  const float cells = 2048.0f;
  const float inv = (ext > 0.0f && ext < __int_as_float(0x7f800000)) ? cells / ext : 0.0f;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    const float v[3] = {x[i], y[i], z[i]};
    uint32_t c[3];
#pragma unroll
    for (int a = 0; a < 3; ++a) {
      const float f = (v[a] - lo[a]) * inv;
      const float g = (f >= 0.0f) ? f : 0.0f;                     // NaN, -inf -> cell 0
      c[a] = (g < cells) ? (uint32_t)g : 2047u;
    }
    const uint32_t cs = c[small] > 1023u ? 1023u : c[small];
    keys[i] = ((c[a0] >> 10) << 31) | ((c[a1] >> 10) << 30) | spread3_10(c[a0] & 1023u) |
              (spread3_10(c[a1] & 1023u) << 1) | (spread3_10(cs) << 2);
    idx[i] = (uint32_t)i;
  }
}

// `aos` (may be null): the same cloud as (N,3) rows -- the random read of a point then touches one
// or two 32-byte sectors instead of one per plane.
__global__ void __launch_bounds__(256) order_gather_kernel(
    const float* __restrict__ x, const float* __restrict__ y, const float* __restrict__ z,
    const float* __restrict__ aos, const uint32_t* __restrict__ perm, int64_t n, float* __restrict__ sx,
    float* __restrict__ sy, float* __restrict__ sz) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    const int64_t j = perm[i];
    if (aos) { sx[i] = aos[3 * j]; sy[i] = aos[3 * j + 1]; sz[i] = aos[3 * j + 2]; }
    else { sx[i] = x[j]; sy[i] = y[j]; sz[i] = z[j]; }
  }
}

__global__ void __launch_bounds__(256) unpermute_u8_kernel(
    const uint8_t* __restrict__ src, const uint32_t* __restrict__ perm, int64_t n,
    uint8_t* __restrict__ dst) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride)
    dst[perm[i]] = src[i];
}

static size_t order_sort_temp_bytes(int64_t n) {
  size_t bytes = 0;
  cub::DeviceRadixSort::SortPairs(nullptr, bytes, (const uint32_t*)nullptr, (uint32_t*)nullptr,
                                  (const uint32_t*)nullptr, (uint32_t*)nullptr, (int64_t)n, 0,
                                  kOrderKeyBits);
  return bytes;
}

static inline int order_grid(int64_t n) {
  int64_t b = (n + 2047) / 2048;
  if (b < 1) b = 1;
  if (b > kNumSMs * 8) b = kNumSMs * 8;
  return (int)b;
}

}  // namespace ptgs

using namespace ptgs;

extern "C" {

size_t ptgs_spatial_order_workspace_bytes(int64_t n) {
  if (n < 1) n = 1;
  size_t b = 512;
  b += 2 * align_up((size_t)n * 4, 256);      // keys in / out
  b += align_up((size_t)n * 4, 256);          // identity index
  b += align_up(order_sort_temp_bytes(n), 256) + 4096;
  return b;
}

int ptgs_spatial_order(const float* px, const float* py, const float* pz, const float* xyz_aos, int64_t n, float* sx,
                       float* sy, float* sz, uint32_t* perm, void* workspace,
                       size_t workspace_bytes, ptgs_stream_t stream) {
  PTGS_REQUIRE(n >= 0 && n < 4294967296ll, "spatial_order: n must be in [0, 2^32)");
  if (n == 0) return PTGS_OK;
  PTGS_REQUIRE(px && py && pz && sx && sy && sz && perm && workspace, "spatial_order: null pointer");
  if (workspace_bytes < ptgs_spatial_order_workspace_bytes(n)) {
    set_error("spatial_order: workspace too small");
    return PTGS_E_WORKSPACE;
  }
  cudaStream_t st = (cudaStream_t)stream;
  Carver cv(workspace);
  uint32_t* mm = cv.take<uint32_t>(8);
  uint32_t* keys_a = cv.take<uint32_t>(n);
  uint32_t* keys_b = cv.take<uint32_t>(n);
  uint32_t* idx_a = cv.take<uint32_t>(n);
  size_t tb = order_sort_temp_bytes(n);
  void* sort_tmp = cv.take<char>(tb);
  const uint32_t init[8] = {~0u, ~0u, ~0u, 0u, 0u, 0u, 0u, 0u};
  PTGS_CUDA_TRY(cudaMemcpyAsync(mm, init, sizeof(init), cudaMemcpyHostToDevice, st));
  const int g = order_grid(n);
  order_minmax_kernel<<<g, 256, 0, st>>>(px, py, pz, n, mm);
  PTGS_LAUNCH_CHECK();
  order_key_kernel<<<g, 256, 0, st>>>(px, py, pz, n, mm, keys_a, idx_a);
  PTGS_LAUNCH_CHECK();
  PTGS_CUDA_TRY(cub::DeviceRadixSort::SortPairs(sort_tmp, tb, keys_a, keys_b, idx_a, perm,
                                                (int64_t)n, 0, kOrderKeyBits, st));
  count_launch(2 + (kOrderKeyBits + 7) / 8);
  order_gather_kernel<<<g, 256, 0, st>>>(px, py, pz, xyz_aos, perm, n, sx, sy, sz);
  PTGS_LAUNCH_CHECK();
  return PTGS_OK;
}

int ptgs_unpermute_u8(const uint8_t* src, const uint32_t* perm, int64_t n, uint8_t* dst,
                      ptgs_stream_t stream) {
  PTGS_REQUIRE(n >= 0, "unpermute_u8: n < 0");
  if (n == 0) return PTGS_OK;
  PTGS_REQUIRE(src && perm && dst, "unpermute_u8: null pointer");
  unpermute_u8_kernel<<<order_grid(n), 256, 0, (cudaStream_t)stream>>>(src, perm, n, dst);
  PTGS_LAUNCH_CHECK();
  return PTGS_OK;
}

}  // extern "C"
