// Shared helpers for libptgs (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#include "../../include/ptgs.h"

namespace ptgs {

void set_error(const char* fmt, ...);
void count_launch(int n = 1);

#define PTGS_CUDA_TRY(expr)                                                   \
  do {                                                                        \
    cudaError_t _e = (expr);                                                  \
    if (_e != cudaSuccess) {                                                  \
      ptgs::set_error("%s:%d %s -> %s", __FILE__, __LINE__, #expr,            \
                      cudaGetErrorString(_e));                                \
      return PTGS_E_CUDA;                                                     \
    }                                                                         \
  } while (0)

#define PTGS_LAUNCH_CHECK()                                                   \
  do {                                                                        \
    ptgs::count_launch();                                                     \
    cudaError_t _e = cudaGetLastError();                                      \
    if (_e != cudaSuccess) {                                                  \
      ptgs::set_error("%s:%d launch -> %s", __FILE__, __LINE__,               \
                      cudaGetErrorString(_e));                                \
      return PTGS_E_CUDA;                                                     \
    }                                                                         \
  } while (0)

#define PTGS_REQUIRE(cond, msg)                                               \
  do {                                                                        \
    if (!(cond)) {                                                            \
      ptgs::set_error("%s:%d invalid argument: %s", __FILE__, __LINE__, msg); \
      return PTGS_E_INVALID;                                                  \
    }                                                                         \
  } while (0)

constexpr int kNumSMs = 148;  // B200
constexpr unsigned kFullMask = 0xffffffffu;

static inline size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

// Bump allocator over a caller-provided workspace.
struct Carver {
  char* base;
  size_t off = 0;
  explicit Carver(void* p) : base(static_cast<char*>(p)) {}
  template <typename T>
  T* take(size_t count) {
    off = align_up(off, 256);
    T* r = reinterpret_cast<T*>(base + off);
    off += count * sizeof(T);
    return r;
  }
  size_t used() const { return align_up(off, 256); }
};

// ---- order preserving float <-> unsigned key ------------------------------
// -0.0 is canonicalised to +0.0 first so that numerically equal values share
// one key (np.sort treats them as equal).
template <typename T>
struct KeyOf;
template <>
struct KeyOf<float> {
  using type = uint32_t;
  static constexpr int bits = 32;
  __host__ __device__ static inline uint32_t encode(float f) {
    f = f + 0.0f;
#ifdef __CUDA_ARCH__
    uint32_t u = __float_as_uint(f);
#else
    uint32_t u;
    memcpy(&u, &f, 4);
#endif
    return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
  }
  __host__ __device__ static inline float decode(uint32_t k) {
    uint32_t u = (k & 0x80000000u) ? (k & 0x7fffffffu) : ~k;
#ifdef __CUDA_ARCH__
    return __uint_as_float(u);
#else
    float f;
    memcpy(&f, &u, 4);
    return f;
#endif
  }
};
template <>
struct KeyOf<double> {
  using type = uint64_t;
  static constexpr int bits = 64;
  __host__ __device__ static inline uint64_t encode(double f) {
    f = f + 0.0;
#ifdef __CUDA_ARCH__
    uint64_t u = (uint64_t)__double_as_longlong(f);
#else
    uint64_t u;
    memcpy(&u, &f, 8);
#endif
    return (u & 0x8000000000000000ull) ? ~u : (u | 0x8000000000000000ull);
  }
  __host__ __device__ static inline double decode(uint64_t k) {
    uint64_t u = (k & 0x8000000000000000ull) ? (k & 0x7fffffffffffffffull) : ~k;
#ifdef __CUDA_ARCH__
    return __longlong_as_double((long long)u);
#else
    double f;
    memcpy(&f, &u, 8);
    return f;
#endif
  }
};

// ---- warp / block primitives ----------------------------------------------
__device__ __forceinline__ int lane_id() { return threadIdx.x & 31; }
__device__ __forceinline__ unsigned lanemask_lt() {
  unsigned m;
  asm("mov.u32 %0, %%lanemask_lt;" : "=r"(m));
  return m;
}

template <typename T>
__device__ __forceinline__ T warp_inclusive_sum(T v) {
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) {
    T n = __shfl_up_sync(kFullMask, v, d);
    if (lane_id() >= d) v += n;
  }
  return v;
}

template <typename T>
__device__ __forceinline__ T warp_sum(T v) {
#pragma unroll
  for (int d = 16; d > 0; d >>= 1) v += __shfl_xor_sync(kFullMask, v, d);
  return v;
}

// Exclusive sum over a block of NT threads (NT multiple of 32, <= 1024).
// Returns the exclusive prefix of `v`; *total receives the block sum.
// smem: at least 33 ints of T.
template <typename T, int NT>
__device__ __forceinline__ T block_exclusive_sum(T v, T* smem, T* total) {
  const int w = threadIdx.x >> 5;
  T inc = warp_inclusive_sum(v);
  if (lane_id() == 31) smem[w] = inc;
  __syncthreads();
  if (w == 0) {
    T s = (lane_id() < NT / 32) ? smem[lane_id()] : T(0);
    T si = warp_inclusive_sum(s);
    smem[lane_id()] = si - s;
    if (lane_id() == 31) smem[32] = si;
  }
  __syncthreads();
  T r = inc - v + smem[w];
  *total = smem[32];
  __syncthreads();
  return r;
}

}  // namespace ptgs
