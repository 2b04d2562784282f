// libptgs A6: voxel down-sampling of every block in one pass.
// Reference: downsample_point_cloud, scene/ptgs/shen_partation_densify.py:541-614
//   min_bound = min(points)                           (:567, per block)
//   vid       = floor((p - min_bound) / voxel).int32  (:572, point dtype arithmetic)
//   one output point per occupied voxel = mean of its points, accumulated
//   sequentially in index order (np.mean(points[idx], axis=0), :588-591),
//   cast to float32 (:605); voxels in first-occurrence order (:576-581).
//
// GPU form: per-block min/max by ordered-key atomics, a packed
// (block | vx | vy | vz) 64-bit key per CSR entry, one stable LSD radix sort of
// (key, entry position), segment heads by neighbour compare + scan, and one
// thread per occupied voxel walking its (tiny) run in entry order so the sum
// is the same left-to-right float addition the reference performs.  Voxels
// come out in ascending key order (coarse cell major, see vox_key_kernel); `ds_first` (position of the voxel's first
// point inside its block) recovers the reference's first-occurrence order.
#include <cub/device/device_radix_sort.cuh>

#include "common.cuh"

namespace ptgs {

constexpr int kVoxMaxBlocks = 1024;
constexpr int kVoxCoarseShift = 6;   // coarse cell = 64 voxels (6.4 units at voxel 0.1)

struct VoxLayout {
  int32_t bits_b, bits_x, bits_y, bits_z;   // field widths
  int32_t total_bits;
  int32_t max_vx, max_vy, max_vz;
  int32_t bits_m;    // width of the coarse-cell field (Morton interleave of cx, cy when it fits)
  int32_t morton;
};

template <typename T> struct VoxArith;
template <> struct VoxArith<float> {
  __device__ static __forceinline__ long long vid(float p, float mn, float voxel) {
    // fp32: subtract, IEEE divide, floor, cast (NumPy: f32 array / weak Python float)
    return (long long)floorf(__fdiv_rn(__fsub_rn(p, mn), voxel));
  }
};
template <> struct VoxArith<double> {
  __device__ static __forceinline__ long long vid(double p, double mn, double voxel) {
    return (long long)floor(__ddiv_rn(__dsub_rn(p, mn), voxel));
  }
};

// Where an entry's coordinates come from.  With an index list the points of a block are scattered
// over a file-ordered cloud: reading three planes at a random position costs three 32-byte
// sectors per value, and the three passes that need coordinates (min / max, keys, means) would
// each pay that.  The first pass (min / max) therefore packs the entries ONCE into (x, y, z, -) records
// (`ent`, 16 / 32 bytes: one sector per random read, fully used by sequential ones); without an
// index list (entries = the points themselves) the planes are read directly.
template <typename T> struct Vec4Of;
template <> struct Vec4Of<float> { using type = float4; };
template <> struct Vec4Of<double> { using type = double4; };

template <typename T>
struct PointSrc {
  const T* x; const T* y; const T* z;
  const int32_t* indices;                       // may be null
  const typename Vec4Of<T>::type* ent;          // may be null
  const T* aos;                                 // may be null: the cloud as (N,3) rows
  // first touch of an index-list call (the min / max pass): read the point through the index list
  // and leave its packed record behind for the later passes
  __device__ __forceinline__ void load_and_pack(int64_t e, typename Vec4Of<T>::type* out, T& px, T& py,
                                                T& pz) const {
    const int64_t i = indices[e];
    if (aos) { px = aos[3 * i]; py = aos[3 * i + 1]; pz = aos[3 * i + 2]; }
    else { px = x[i]; py = y[i]; pz = z[i]; }
    typename Vec4Of<T>::type v;
    v.x = px; v.y = py; v.z = pz; v.w = T(0);
    out[e] = v;
  }
  __device__ __forceinline__ void load(int64_t e, T& px, T& py, T& pz) const {
    if (ent) {
      const typename Vec4Of<T>::type v = ent[e];
      px = v.x; py = v.y; pz = v.z;
    } else {
      const int64_t i = indices ? (int64_t)indices[e] : e;
      if (aos) { px = aos[3 * i]; py = aos[3 * i + 1]; pz = aos[3 * i + 2]; }   // rows only: no planes needed
      else { px = x[i]; py = y[i]; pz = z[i]; }
    }
  }
};

// bit i of x -> bit 2i (32 source bits) / bit 3i (10 source bits)
__device__ __forceinline__ uint64_t spread2(uint64_t x) {
  x &= 0xffffffffull;
  x = (x | (x << 16)) & 0x0000ffff0000ffffull;
  x = (x | (x << 8)) & 0x00ff00ff00ff00ffull;
  x = (x | (x << 4)) & 0x0f0f0f0f0f0f0f0full;
  x = (x | (x << 2)) & 0x3333333333333333ull;
  x = (x | (x << 1)) & 0x5555555555555555ull;
  return x;
}
__device__ __forceinline__ uint32_t spread3(uint32_t x) {
  x &= 0x3ffu;
  x = (x | (x << 16)) & 0xff0000ffu;
  x = (x | (x << 8)) & 0x0300f00fu;
  x = (x | (x << 4)) & 0x030c30c3u;
  x = (x | (x << 2)) & 0x09249249u;
  return x;
}

__device__ __forceinline__ int find_block(const int64_t* s_off, int L, int64_t e) {
  int lo = 0, hi = L - 1;  // last b with off[b] <= e
  while (lo < hi) {
    int mid = (lo + hi + 1) >> 1;
    if (s_off[mid] <= e) lo = mid; else hi = mid - 1;
  }
  return lo;
}

__device__ __forceinline__ void amin(uint32_t* a, uint32_t v) { atomicMin(a, v); }
__device__ __forceinline__ void amax(uint32_t* a, uint32_t v) { atomicMax(a, v); }
__device__ __forceinline__ void amin(uint64_t* a, uint64_t v) { atomicMin((unsigned long long*)a, (unsigned long long)v); }
__device__ __forceinline__ void amax(uint64_t* a, uint64_t v) { atomicMax((unsigned long long*)a, (unsigned long long)v); }

template <typename K>
__global__ void vox_init_kernel(K* mm, int L) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < L * 6) mm[i] = ((i % 6) < 3) ? ~K(0) : K(0);
}

// per block min (slots 0..2) and max (3..5) as ordered keys.
template <typename T>
__global__ void __launch_bounds__(256) vox_minmax_kernel(
    PointSrc<T> src, typename Vec4Of<T>::type* __restrict__ pack_out /* may be null */,
    const int64_t* __restrict__ offsets,
    int64_t n_entries, int L, typename KeyOf<T>::type* __restrict__ mm) {
  using K = typename KeyOf<T>::type;
  __shared__ int64_t s_off[kVoxMaxBlocks + 1];
  for (int i = threadIdx.x; i <= L; i += blockDim.x) s_off[i] = offsets[i];
  __syncthreads();
  // every warp walks ONE contiguous run of entries and keeps a running per-lane
  // min / max for as long as the run stays inside one block; the six atomics
  // are issued once per (warp, block) instead of once per 32 entries
  const int64_t n_warps = (int64_t)gridDim.x * (blockDim.x / 32);
  const int64_t per_warp = ((n_entries + n_warps - 1) / n_warps + 31) / 32 * 32;
  const int64_t w0 = ((int64_t)blockIdx.x * (blockDim.x / 32) + threadIdx.x / 32) * per_warp;
  const int64_t w1 = w0 + per_warp < n_entries ? w0 + per_warp : n_entries;
  const int lane = threadIdx.x & 31;
  int b_run = -1;
  K rlo[3] = {~K(0), ~K(0), ~K(0)}, rhi[3] = {0, 0, 0};
  auto flush = [&]() {
    if (b_run < 0) return;
#pragma unroll
    for (int a = 0; a < 3; ++a) {
      K lo = rlo[a], hi = rhi[a];
#pragma unroll
      for (int d = 16; d > 0; d >>= 1) {
        K o = __shfl_xor_sync(kFullMask, lo, d); lo = o < lo ? o : lo;
        K p = __shfl_xor_sync(kFullMask, hi, d); hi = p > hi ? p : hi;
      }
      if (lane == 0) {
        if (lo < __ldcg(mm + 6 * b_run + a)) amin(mm + 6 * b_run + a, lo);
        if (hi > __ldcg(mm + 6 * b_run + 3 + a)) amax(mm + 6 * b_run + 3 + a, hi);
      }
      rlo[a] = ~K(0); rhi[a] = 0;
    }
  };
  for (int64_t e0 = w0; e0 < w1; e0 += 32) {
    const int64_t e = e0 + lane;
    const bool valid = e < w1;
    int b = -1;
    K k[3] = {~K(0), ~K(0), ~K(0)};
    if (valid) {
      b = find_block(s_off, L, e);
      T px, py, pz;
      if (pack_out) src.load_and_pack(e, pack_out, px, py, pz);
      else src.load(e, px, py, pz);
      k[0] = KeyOf<T>::encode(px); k[1] = KeyOf<T>::encode(py); k[2] = KeyOf<T>::encode(pz);
    }
    const int b0 = __shfl_sync(kFullMask, b, 0);
    const bool uniform = __all_sync(kFullMask, b == b0 || !valid) && b0 >= 0;
    if (uniform) {
      if (b0 != b_run) { flush(); b_run = b0; }
      if (valid) {
#pragma unroll
        for (int a = 0; a < 3; ++a) { rlo[a] = k[a] < rlo[a] ? k[a] : rlo[a]; rhi[a] = k[a] > rhi[a] ? k[a] : rhi[a]; }
      }
    } else {
      // the 32 entries straddle a block boundary (at most L - 1 times overall)
      flush(); b_run = -1;
      if (valid) {
#pragma unroll
        for (int a = 0; a < 3; ++a) {
          if (k[a] < __ldcg(mm + 6 * b + a)) amin(mm + 6 * b + a, k[a]);
          if (k[a] > __ldcg(mm + 6 * b + 3 + a)) amax(mm + 6 * b + 3 + a, k[a]);
        }
      }
    }
  }
  flush();
}

// The host needs the key layout (number of sorted bits) before it can queue the sort.  It is read
// through pinned host memory that the layout kernel writes directly: a cudaMemcpyAsync would queue
// behind whatever large device -> host copy another stream has in flight (partition_device's early
// copy of the index list), and the stage would wait for all of it.
struct VoxHostSlot { VoxLayout lay; int32_t status; };
static VoxHostSlot* vox_host_slot() {
  thread_local VoxHostSlot* slot = nullptr;
  if (!slot) {
    void* p = nullptr;
    if (cudaHostAlloc(&p, sizeof(VoxHostSlot), cudaHostAllocDefault) == cudaSuccess) slot = (VoxHostSlot*)p;
    else (void)cudaGetLastError();
  }
  return slot;
}

// single CTA: decode per block min, global max voxel index per axis, layout.
template <typename T>
__global__ void __launch_bounds__(1024) vox_layout_kernel(
    const typename KeyOf<T>::type* __restrict__ mm, const int64_t* __restrict__ offsets, int L,
    double voxel, T* __restrict__ bmin /* [L][3] */, VoxLayout* __restrict__ lay,
    int32_t* __restrict__ status, VoxHostSlot* host_slot /* pinned host memory, may be null */) {
  __shared__ long long smax[3];
  if (threadIdx.x < 3) smax[threadIdx.x] = 0;
  __syncthreads();
  for (int b = threadIdx.x; b < L; b += blockDim.x) {
    const bool empty = offsets[b + 1] == offsets[b];
#pragma unroll
    for (int a = 0; a < 3; ++a) {
      T mn = empty ? T(0) : KeyOf<T>::decode(mm[6 * b + a]);
      T mx = empty ? T(0) : KeyOf<T>::decode(mm[6 * b + 3 + a]);
      bmin[3 * b + a] = mn;
      long long v = VoxArith<T>::vid(mx, mn, (T)voxel);
      if (v < 0) v = 0;
      atomicMax((unsigned long long*)&smax[a], (unsigned long long)v);
    }
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    auto nbits = [](long long v) { int b = 1; while (b < 63 && (v >> b)) ++b; return b; };
    VoxLayout l;
    l.bits_b = nbits(L - 1);
    l.bits_x = nbits(smax[0]); l.bits_y = nbits(smax[1]); l.bits_z = nbits(smax[2]);
    {
      // coarse cells (vx >> 6, vy >> 6) are Morton-interleaved when the padding fits in 64 bits
      const int nx = l.bits_x > kVoxCoarseShift ? l.bits_x - kVoxCoarseShift : 0;
      const int ny = l.bits_y > kVoxCoarseShift ? l.bits_y - kVoxCoarseShift : 0;
      const int nm = 2 * (nx > ny ? nx : ny);
      const int rest = l.bits_b + l.bits_x + l.bits_y + l.bits_z - nx - ny;
      l.morton = (rest + nm <= 64) ? 1 : 0;
      l.bits_m = l.morton ? nm : nx + ny;
      l.total_bits = rest + l.bits_m;
    }
    const bool ok = smax[0] < 0x7fffffffll && smax[1] < 0x7fffffffll && smax[2] < 0x7fffffffll &&
                    l.total_bits <= 64;
    l.max_vx = (int32_t)min(smax[0], 0x7fffffffll);
    l.max_vy = (int32_t)min(smax[1], 0x7fffffffll);
    l.max_vz = (int32_t)min(smax[2], 0x7fffffffll);
    *lay = l;
    *status = ok ? PTGS_OK : PTGS_E_RANGE;
    if (host_slot) {
      host_slot->lay = l;
      host_slot->status = ok ? PTGS_OK : PTGS_E_RANGE;
      __threadfence_system();
    }
  }
}

template <typename T>
__global__ void __launch_bounds__(256) vox_key_kernel(
    PointSrc<T> src, const int64_t* __restrict__ offsets,
    int64_t n_entries, int L, const T* __restrict__ bmin, double voxel,
    const VoxLayout* __restrict__ lay, uint64_t* __restrict__ keys, uint32_t* __restrict__ vals) {
  __shared__ int64_t s_off[kVoxMaxBlocks + 1];
  for (int i = threadIdx.x; i <= L; i += blockDim.x) s_off[i] = offsets[i];
  __syncthreads();
  const VoxLayout l = *lay;
  const T vs = (T)voxel;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  // 32 consecutive entries almost always lie in one block: one search per warp (lane 0's entry),
  // the lanes behind a block boundary search for themselves
  const int64_t n_pad = (n_entries + 31) & ~(int64_t)31;
  for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < n_pad; e += stride) {
    const int64_t e_first = e - (threadIdx.x & 31);
    int b = 0;
    if ((threadIdx.x & 31) == 0) b = find_block(s_off, L, e_first < n_entries ? e_first : n_entries - 1);
    b = __shfl_sync(kFullMask, b, 0);
    if (e >= n_entries) continue;
    if (e >= s_off[b + 1]) b = find_block(s_off, L, e);
    T px, py, pz;
    src.load(e, px, py, pz);
    const uint64_t vx = (uint64_t)VoxArith<T>::vid(px, bmin[3 * b], vs);
    const uint64_t vy = (uint64_t)VoxArith<T>::vid(py, bmin[3 * b + 1], vs);
    const uint64_t vz = (uint64_t)VoxArith<T>::vid(pz, bmin[3 * b + 2], vs);
    // block | coarse layer (vz >> s) | coarse cell (vx >> s, vy >> s, Morton
    // interleaved) | fine (vx, vy, vz): an injective re-arrangement of the bits
    // of (vx, vy, vz), so the voxel groups are unchanged and only their order
    // moves.  Layer-major, then Z-order over the 6.4-unit cells, makes every
    // run of consecutive voxels (a sweep tile) compact in all three axes --
    // the visibility sweep's tile culling and whole-tile accept rely on it
    // (a box that is tall in z projects to a large image footprint).
    const int lx = l.bits_x < kVoxCoarseShift ? l.bits_x : kVoxCoarseShift;
    const int ly = l.bits_y < kVoxCoarseShift ? l.bits_y : kVoxCoarseShift;
    const int lz = l.bits_z < kVoxCoarseShift ? l.bits_z : kVoxCoarseShift;
    const uint64_t cx = vx >> lx, cy = vy >> ly;
    uint64_t cell;
    if (l.morton) {
      cell = (spread2(cx) << 1) | spread2(cy);      // bit i of cx -> 2i + 1, of cy -> 2i
    } else {
      cell = (cx << (l.bits_y - ly)) | cy;
    }
    uint64_t k = (uint64_t)b;
    k = (k << (l.bits_z - lz)) | (vz >> lz);
    k = (k << l.bits_m) | cell;
    // fine bits: 3-D Morton, so that 256 / 32 consecutive voxels (a warp's share
    // of a sweep tile / one register slot of a warp) are compact cubes
    if (lx == kVoxCoarseShift && ly == kVoxCoarseShift && lz == kVoxCoarseShift) {
      // all three axes have their 6 fine bits: x5 y5 z5 ... x0 y0 z0 by bit spreading
      const uint32_t m = (1u << kVoxCoarseShift) - 1u;
      k = (k << (3 * kVoxCoarseShift)) |
          (uint64_t)((spread3((uint32_t)vx & m) << 2) | (spread3((uint32_t)vy & m) << 1) | spread3((uint32_t)vz & m));
    } else {
      for (int i = kVoxCoarseShift - 1; i >= 0; --i) {
        if (i < lx) k = (k << 1) | ((vx >> i) & 1ull);
        if (i < ly) k = (k << 1) | ((vy >> i) & 1ull);
        if (i < lz) k = (k << 1) | ((vz >> i) & 1ull);
      }
    }
    keys[e] = k;
    vals[e] = (uint32_t)e;
  }
}

constexpr int kHeadTile = 2048;

// head flag = key differs from the predecessor; per tile head counts.
__global__ void __launch_bounds__(256) vox_head_count_kernel(
    const uint64_t* __restrict__ keys, int64_t n, int32_t* __restrict__ tile_count) {
  __shared__ int s[33];
  const int64_t base = (int64_t)blockIdx.x * kHeadTile;
  int c = 0;
#pragma unroll
  for (int k = 0; k < 8; ++k) {
    const int64_t i = base + k * 256 + threadIdx.x;
    if (i < n) c += (i == 0 || keys[i] != keys[i - 1]) ? 1 : 0;
  }
  int tot;
  block_exclusive_sum<int, 256>(c, s, &tot);
  if (threadIdx.x == 0) tile_count[blockIdx.x] = tot;
}

__global__ void __launch_bounds__(1024) vox_scan_tiles_kernel(
    const int32_t* __restrict__ tile_count, int64_t n_tiles, int64_t* __restrict__ tile_base,
    int64_t* __restrict__ total) {
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

// one thread per sorted entry; heads walk their run and emit the mean.
template <typename T>
__global__ void __launch_bounds__(256) vox_mean_kernel(
    PointSrc<T> src, const int64_t* __restrict__ offsets,
    const uint64_t* __restrict__ keys, const uint32_t* __restrict__ vals, int64_t n,
    int L, const VoxLayout* __restrict__ lay, const int64_t* __restrict__ tile_base,
    float* __restrict__ ds_x, float* __restrict__ ds_y, float* __restrict__ ds_z,
    int32_t* __restrict__ ds_first, int64_t* __restrict__ ds_offsets) {
  __shared__ int s[33];
  const VoxLayout l = *lay;
  const int shift_b = l.total_bits - l.bits_b;
  const int64_t base = (int64_t)blockIdx.x * kHeadTile;
  int64_t out = tile_base[blockIdx.x];
#pragma unroll 1
  for (int k = 0; k < 8; ++k) {
    const int64_t i = base + k * 256 + threadIdx.x;
    uint64_t key = 0;
    int f = 0;
    bool first_of_block = false;
    if (i < n) {
      key = keys[i];
      if (i == 0) { f = 1; first_of_block = true; }
      else {
        const uint64_t prev = keys[i - 1];
        f = key != prev;
        first_of_block = (shift_b >= 64 ? 0 : (key >> shift_b)) != (shift_b >= 64 ? 0 : (prev >> shift_b));
      }
    }
    int tot;
    const int ex = block_exclusive_sum<int, 256>(f, s, &tot);
    if (f) {
      const int64_t rank = out + ex;
      const int b = shift_b >= 64 ? 0 : (int)(key >> shift_b);
      const uint32_t e0 = vals[i];
      // np.add.reduce over axis 0 starts from the identity +0.0, so a voxel
      // whose members are all -0.0 averages to +0.0
      T sx, sy, sz;
      {
        T px, py, pz;
        src.load((int64_t)e0, px, py, pz);
        sx = T(0) + px; sy = T(0) + py; sz = T(0) + pz;
      }
      int64_t j = i + 1;
      long long cnt = 1;
      for (; j < n && keys[j] == key; ++j) {
        T px, py, pz;
        src.load((int64_t)vals[j], px, py, pz);
        sx = sx + px; sy = sy + py; sz = sz + pz;   // sequential, entry order
        ++cnt;
      }
      const T c = (T)cnt;
      ds_x[rank] = (float)(sx / c);
      ds_y[rank] = (float)(sy / c);
      ds_z[rank] = (float)(sz / c);
      ds_first[rank] = (int32_t)((int64_t)e0 - offsets[b]);
      if (first_of_block) ds_offsets[b] = rank;   // provisional: non-empty blocks only
    }
    out += tot;
  }
}

// single CTA: fill ds_offsets for empty blocks (monotone back-fill) and [L].
__global__ void vox_fix_offsets_kernel(const int64_t* __restrict__ offsets, int L,
                                       const int64_t* __restrict__ total,
                                       int64_t* __restrict__ ds_offsets) {
  if (threadIdx.x == 0 && blockIdx.x == 0) {
    int64_t next = *total;
    ds_offsets[L] = next;
    for (int b = L - 1; b >= 0; --b) {
      if (offsets[b + 1] == offsets[b]) ds_offsets[b] = next;
      else next = ds_offsets[b];
    }
  }
}

static size_t vox_sort_temp_bytes(int64_t n) {
  size_t bytes = 0;
  cub::DeviceRadixSort::SortPairs(nullptr, bytes, (const uint64_t*)nullptr, (uint64_t*)nullptr,
                                  (const uint32_t*)nullptr, (uint32_t*)nullptr, (int64_t)n, 0, 64);
  return bytes;
}

struct VoxWs {
  void* ent;
  void* mm; void* bmin; VoxLayout* lay; uint64_t* keys_a; uint64_t* keys_b;
  uint32_t* vals_a; uint32_t* vals_b; int32_t* tile_count; int64_t* tile_base; int64_t* total;
  void* sort_tmp; size_t sort_bytes; int64_t n_tiles; size_t bytes;
};
static VoxWs carve_vox(void* ws, int64_t n, int32_t L) {
  VoxWs v;
  if (n < 1) n = 1;
  Carver cv(ws);
  v.mm = cv.take<uint64_t>((size_t)L * 6);
  v.bmin = cv.take<double>((size_t)L * 3);
  v.lay = cv.take<VoxLayout>(1);
  v.keys_a = cv.take<uint64_t>(n);
  v.keys_b = cv.take<uint64_t>(n);
  v.vals_a = cv.take<uint32_t>(n);
  v.vals_b = cv.take<uint32_t>(n);
  v.n_tiles = (n + kHeadTile - 1) / kHeadTile;
  v.tile_count = cv.take<int32_t>(v.n_tiles);
  v.tile_base = cv.take<int64_t>(v.n_tiles);
  v.total = cv.take<int64_t>(1);
  v.sort_bytes = vox_sort_temp_bytes(n);
  v.sort_tmp = cv.take<char>(v.sort_bytes);
  v.ent = cv.take<double4>(n);                 // packed entries (float4 or double4), index-list calls only
  v.bytes = cv.used() + 256;
  return v;
}

template <typename T>
static int voxel_impl(const void* vx, const void* vy, const void* vz, const void* vaos, const int64_t* offsets,
                      const int32_t* indices, int64_t n, int32_t L, double voxel,
                      float* ds_x, float* ds_y, float* ds_z, int32_t* ds_first,
                      int64_t* ds_offsets, int32_t* status_dev, void* ws, cudaStream_t st) {
  using K = typename KeyOf<T>::type;
  const T* x = (const T*)vx; const T* y = (const T*)vy; const T* z = (const T*)vz;
  VoxWs w = carve_vox(ws, n, L);
  K* mm = (K*)w.mm;
  T* bmin = (T*)w.bmin;
  vox_init_kernel<K><<<(L * 6 + 255) / 256, 256, 0, st>>>(mm, L);
  PTGS_LAUNCH_CHECK();
  int64_t g64 = (n + 2047) / 2048;
  int grid = (int)(g64 < 1 ? 1 : (g64 > kNumSMs * 8 ? kNumSMs * 8 : g64));
  using V4 = typename Vec4Of<T>::type;
  PointSrc<T> src = {x, y, z, indices, nullptr, (const T*)vaos};
  // with an index list the min / max pass packs the entries' coordinates on the way (w.ent)
  vox_minmax_kernel<T><<<grid, 256, 0, st>>>(src, indices ? (V4*)w.ent : nullptr, offsets, n, L, mm);
  if (indices) src.ent = (const V4*)w.ent;
  PTGS_LAUNCH_CHECK();
  VoxHostSlot* hs = vox_host_slot();
  vox_layout_kernel<T><<<1, 1024, 0, st>>>(mm, offsets, L, voxel, bmin, w.lay, status_dev, hs);
  PTGS_LAUNCH_CHECK();
  VoxLayout lay;
  int32_t status;
  if (hs) {
    PTGS_CUDA_TRY(cudaStreamSynchronize(st));
    lay = hs->lay;
    status = hs->status;
  } else {
    PTGS_CUDA_TRY(cudaMemcpyAsync(&lay, w.lay, sizeof(lay), cudaMemcpyDeviceToHost, st));
    PTGS_CUDA_TRY(cudaMemcpyAsync(&status, status_dev, sizeof(status), cudaMemcpyDeviceToHost, st));
    PTGS_CUDA_TRY(cudaStreamSynchronize(st));
  }
  if (status != PTGS_OK) {
    set_error("voxel_downsample: voxel grid does not fit the 64-bit key (bits b/x/y/z = %d/%d/%d/%d)",
              lay.bits_b, lay.bits_x, lay.bits_y, lay.bits_z);
    return PTGS_E_RANGE;
  }
  vox_key_kernel<T><<<grid, 256, 0, st>>>(src, offsets, n, L, bmin, voxel, w.lay, w.keys_a, w.vals_a);
  PTGS_LAUNCH_CHECK();
  PTGS_CUDA_TRY(cub::DeviceRadixSort::SortPairs(w.sort_tmp, w.sort_bytes, w.keys_a, w.keys_b,
                                                w.vals_a, w.vals_b, (int64_t)n, 0, lay.total_bits, st));
  count_launch(2 + (lay.total_bits + 7) / 8);
  vox_head_count_kernel<<<(unsigned)w.n_tiles, 256, 0, st>>>(w.keys_b, n, w.tile_count);
  PTGS_LAUNCH_CHECK();
  vox_scan_tiles_kernel<<<1, 1024, 0, st>>>(w.tile_count, w.n_tiles, w.tile_base, w.total);
  PTGS_LAUNCH_CHECK();
  vox_mean_kernel<T><<<(unsigned)w.n_tiles, 256, 0, st>>>(src, offsets, w.keys_b, w.vals_b,
                                                         n, L, w.lay, w.tile_base, ds_x, ds_y, ds_z,
                                                         ds_first, ds_offsets);
  PTGS_LAUNCH_CHECK();
  vox_fix_offsets_kernel<<<1, 32, 0, st>>>(offsets, L, w.total, ds_offsets);
  PTGS_LAUNCH_CHECK();
  return PTGS_OK;
}

__global__ void vox_empty_kernel(int64_t* ds_offsets, int L, int32_t* status) {
  for (int i = threadIdx.x; i <= L; i += blockDim.x) ds_offsets[i] = 0;
  if (threadIdx.x == 0) *status = PTGS_OK;
}

}  // namespace ptgs

using namespace ptgs;

extern "C" {

size_t ptgs_voxel_downsample_workspace_bytes(int64_t n_entries, int32_t n_blocks) {
  if (n_blocks < 1) n_blocks = 1;
  return carve_vox(nullptr, n_entries, n_blocks).bytes;
}

int ptgs_voxel_downsample(int dtype, const void* x, const void* y,
                          const void* z, const void* xyz_aos, const int64_t* offsets,
                          const int32_t* indices, int64_t n_entries,
                          int32_t n_blocks, double voxel_size, float* ds_x,
                          float* ds_y, float* ds_z, int32_t* ds_first,
                          int64_t* ds_offsets, int32_t* status_dev,
                          void* workspace, size_t workspace_bytes,
                          ptgs_stream_t stream) {
  PTGS_REQUIRE(dtype == PTGS_F32 || dtype == PTGS_F64, "dtype");
  PTGS_REQUIRE(n_entries >= 0 && n_entries < (1ll << 32), "n_entries");
  PTGS_REQUIRE(n_blocks >= 1 && n_blocks <= kVoxMaxBlocks, "n_blocks must be in [1,1024]");
  PTGS_REQUIRE(voxel_size > 0.0, "voxel_size");
  PTGS_REQUIRE(offsets && ds_offsets && status_dev, "arguments");
  PTGS_REQUIRE(n_entries == 0 || (x && y && z) || xyz_aos, "either the coordinate planes or the (N,3) rows");
  if (workspace_bytes < ptgs_voxel_downsample_workspace_bytes(n_entries, n_blocks)) {
    set_error("voxel_downsample: workspace too small");
    return PTGS_E_WORKSPACE;
  }
  cudaStream_t st = (cudaStream_t)stream;
  if (n_entries == 0) {
    vox_empty_kernel<<<1, 256, 0, st>>>(ds_offsets, n_blocks, status_dev);
    PTGS_LAUNCH_CHECK();
    return PTGS_OK;
  }
  if (dtype == PTGS_F32)
    return voxel_impl<float>(x, y, z, xyz_aos, offsets, indices, n_entries, n_blocks, voxel_size, ds_x, ds_y,
                             ds_z, ds_first, ds_offsets, status_dev, workspace, st);
  return voxel_impl<double>(x, y, z, xyz_aos, offsets, indices, n_entries, n_blocks, voxel_size, ds_x, ds_y,
                            ds_z, ds_first, ds_offsets, status_dev, workspace, st);
}

}  // extern "C"
