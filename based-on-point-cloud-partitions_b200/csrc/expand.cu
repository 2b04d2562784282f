// libptgs A4 / A5 / A9: expanded-block point selection, camera assignment and
// kept-camera compaction.
//
// A4 (reference: expand_partitions / extract_point_cloud,
// scene/ptgs/shen_partation_densify.py:182-252, mask at :201-202).
// One fused pass tests every point against every expanded rectangle (the
// reference makes one full pass over the cloud per block).  The rectangles
// live in shared memory and are walked in lock-step by the warp, so each
// test is one broadcast LDS.128 + four compares + one vote.  Index lists are
// produced in ascending point order without a sort:
//   count pass : per (CTA tile, warp, block) membership counts -> per
//                (block, tile) totals, written block-major so that the
//                per-block scan over tiles is contiguous;
//   scan       : one CTA per block, exclusive scan over the tiles;
//   fill pass  : replays the tests; rank inside the warp by ballot/popc,
//                warp bases by a shared-memory scan over the 8 warps.
#include "common.cuh"

namespace ptgs {

constexpr int kExpWarps = 8;
constexpr int kExpPerLane = 16;                              // points per lane
constexpr int kExpTile = kExpWarps * 32 * kExpPerLane;       // 4096 points per CTA
constexpr int kExpMaxBlocks = 1024;                          // 2^max_depth leaves

template <typename T> struct Rect4 { T minx, maxx, miny, maxy; };

// Shared layout: rects[L] | wcnt[kExpWarps][L]
template <typename T, bool FILL>
__global__ void __launch_bounds__(kExpWarps * 32) expand_kernel(
    const T* __restrict__ x, const T* __restrict__ y, int64_t n,
    const Rect4<T>* __restrict__ rects, int L, int64_t n_tiles,
    uint32_t* __restrict__ tile_cnt /* [L][n_tiles] */,
    const int64_t* __restrict__ tile_base /* [L][n_tiles], FILL only */,
    int32_t* __restrict__ indices) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  Rect4<T>* srect = reinterpret_cast<Rect4<T>*>(smem_raw);
  uint32_t* wcnt = reinterpret_cast<uint32_t*>(srect + L);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int i = threadIdx.x; i < L; i += blockDim.x) srect[i] = rects[i];
  for (int i = threadIdx.x; i < L * kExpWarps; i += blockDim.x) wcnt[i] = 0;
  __syncthreads();

  const int64_t tile = blockIdx.x;
  // warp w owns the contiguous run [t0 + w*512, +512); sub-group k covers
  // 32 consecutive points, so ascending (warp, k, lane) == ascending index.
  const int64_t wbase = tile * kExpTile + (int64_t)warp * (32 * kExpPerLane);
  uint32_t* mycnt = wcnt + warp * L;

  if (!FILL) {
    for (int k = 0; k < kExpPerLane; ++k) {
      const int64_t i = wbase + k * 32 + lane;
      const bool valid = i < n;
      T px = 0, py = 0;
      if (valid) { px = x[i]; py = y[i]; }
      if (!__any_sync(kFullMask, valid)) break;
      for (int r = 0; r < L; ++r) {
        const Rect4<T> q = srect[r];
        const bool in = valid && (px >= q.minx) && (px <= q.maxx) && (py >= q.miny) && (py <= q.maxy);
        const unsigned b = __ballot_sync(kFullMask, in);
        if (b && lane == 0) mycnt[r] += __popc(b);
      }
    }
    __syncthreads();
    for (int r = threadIdx.x; r < L; r += blockDim.x) {
      uint32_t s = 0;
#pragma unroll
      for (int w = 0; w < kExpWarps; ++w) s += wcnt[w * L + r];
      tile_cnt[(size_t)r * n_tiles + tile] = s;
    }
  } else {
    // pass 1 of the replay: per warp counts (same as above) to derive the
    // warp bases inside the tile.
    for (int k = 0; k < kExpPerLane; ++k) {
      const int64_t i = wbase + k * 32 + lane;
      const bool valid = i < n;
      T px = 0, py = 0;
      if (valid) { px = x[i]; py = y[i]; }
      if (!__any_sync(kFullMask, valid)) break;
      for (int r = 0; r < L; ++r) {
        const Rect4<T> q = srect[r];
        const bool in = valid && (px >= q.minx) && (px <= q.maxx) && (py >= q.miny) && (py <= q.maxy);
        const unsigned b = __ballot_sync(kFullMask, in);
        if (b && lane == 0) mycnt[r] += __popc(b);
      }
    }
    __syncthreads();
    // exclusive scan over the warps, per block (in place)
    for (int r = threadIdx.x; r < L; r += blockDim.x) {
      uint32_t run = 0;
#pragma unroll
      for (int w = 0; w < kExpWarps; ++w) {
        uint32_t c = wcnt[w * L + r];
        wcnt[w * L + r] = run;
        run += c;
      }
    }
    __syncthreads();
    for (int k = 0; k < kExpPerLane; ++k) {
      const int64_t i = wbase + k * 32 + lane;
      const bool valid = i < n;
      T px = 0, py = 0;
      if (valid) { px = x[i]; py = y[i]; }
      if (!__any_sync(kFullMask, valid)) break;
      for (int r = 0; r < L; ++r) {
        const Rect4<T> q = srect[r];
        const bool in = valid && (px >= q.minx) && (px <= q.maxx) && (py >= q.miny) && (py <= q.maxy);
        const unsigned b = __ballot_sync(kFullMask, in);
        if (b) {
          const uint32_t cur = mycnt[r];
          if (in) {
            const int64_t pos = tile_base[(size_t)r * n_tiles + tile] + cur + __popc(b & lanemask_lt());
            indices[pos] = (int32_t)i;
          }
          __syncwarp();
          if (lane == 0) mycnt[r] = cur + __popc(b);
          __syncwarp();
        }
      }
    }
  }
}

// one CTA per block: exclusive scan of its tile counts (int64 result) and the
// block total.
__global__ void __launch_bounds__(1024) expand_scan_kernel(
    const uint32_t* __restrict__ tile_cnt, int64_t n_tiles,
    int64_t* __restrict__ tile_base, int64_t* __restrict__ counts) {
  __shared__ long long sm[33];
  const int r = blockIdx.x;
  const uint32_t* c = tile_cnt + (size_t)r * n_tiles;
  int64_t* o = tile_base + (size_t)r * n_tiles;
  long long carry = 0;
  for (int64_t s = 0; s < n_tiles; s += 1024) {
    const int64_t i = s + threadIdx.x;
    long long v = i < n_tiles ? (long long)c[i] : 0;
    long long tot;
    long long ex = block_exclusive_sum<long long, 1024>(v, sm, &tot);
    if (i < n_tiles) o[i] = carry + ex;  // block-relative; CSR base added below
    carry += tot;
  }
  if (threadIdx.x == 0) counts[r] = carry;
}

// single CTA: offsets[L+1] = exclusive scan of counts.
// (counts may alias offsets: every element is read before its slot is written)
__global__ void __launch_bounds__(1024) offsets_kernel(const int64_t* counts, int L,
                                                       int64_t* offsets) {
  __shared__ long long sm[33];
  long long carry = 0;
  for (int s = 0; s < L; s += 1024) {
    const int i = s + threadIdx.x;
    long long v = i < L ? counts[i] : 0;
    long long tot;
    long long ex = block_exclusive_sum<long long, 1024>(v, sm, &tot);
    if (i < L) offsets[i] = carry + ex;
    carry += tot;
  }
  if (threadIdx.x == 0) offsets[L] = carry;
}

__global__ void __launch_bounds__(256) add_base_kernel(int64_t* __restrict__ tile_base,
                                                       int64_t n_tiles,
                                                       const int64_t* __restrict__ offsets) {
  const int r = blockIdx.y;
  const int64_t base = offsets[r];
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n_tiles;
       i += (int64_t)gridDim.x * blockDim.x)
    tile_base[(size_t)r * n_tiles + i] += base;
}

struct ExpandWs {
  uint32_t* tile_cnt;
  int64_t* tile_base;
  int64_t n_tiles;
  size_t bytes;
};
static ExpandWs carve_expand(void* ws, int64_t n, int32_t L) {
  ExpandWs e;
  e.n_tiles = (n + kExpTile - 1) / kExpTile;
  if (e.n_tiles < 1) e.n_tiles = 1;
  Carver cv(ws);
  e.tile_cnt = cv.take<uint32_t>((size_t)L * e.n_tiles);
  e.tile_base = cv.take<int64_t>((size_t)L * e.n_tiles);
  e.bytes = cv.used() + 256;
  return e;
}

template <typename T>
static int expand_count_impl(const void* x, const void* y, int64_t n, const void* rects,
                             int32_t L, int64_t* counts, void* ws, cudaStream_t st) {
  ExpandWs e = carve_expand(ws, n, L);
  const size_t smem = (size_t)L * sizeof(Rect4<T>) + (size_t)L * kExpWarps * 4;
  PTGS_CUDA_TRY(cudaFuncSetAttribute(expand_kernel<T, false>,
                                     cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  expand_kernel<T, false><<<(unsigned)e.n_tiles, kExpWarps * 32, smem, st>>>(
      (const T*)x, (const T*)y, n, (const Rect4<T>*)rects, L, e.n_tiles, e.tile_cnt, nullptr, nullptr);
  PTGS_LAUNCH_CHECK();
  expand_scan_kernel<<<L, 1024, 0, st>>>(e.tile_cnt, e.n_tiles, e.tile_base, counts);
  PTGS_LAUNCH_CHECK();
  return PTGS_OK;
}

template <typename T>
static int expand_fill_impl(const void* x, const void* y, int64_t n, const void* rects,
                            int32_t L, int64_t* offsets, int32_t* indices, void* ws,
                            const int64_t* counts_dev, cudaStream_t st) {
  ExpandWs e = carve_expand(ws, n, L);
  offsets_kernel<<<1, 1024, 0, st>>>(counts_dev, L, offsets);
  PTGS_LAUNCH_CHECK();
  dim3 g((unsigned)((e.n_tiles + 255) / 256 > 64 ? 64 : (e.n_tiles + 255) / 256), (unsigned)L);
  add_base_kernel<<<g, 256, 0, st>>>(e.tile_base, e.n_tiles, offsets);
  PTGS_LAUNCH_CHECK();
  const size_t smem = (size_t)L * sizeof(Rect4<T>) + (size_t)L * kExpWarps * 4;
  PTGS_CUDA_TRY(cudaFuncSetAttribute(expand_kernel<T, true>,
                                     cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  expand_kernel<T, true><<<(unsigned)e.n_tiles, kExpWarps * 32, smem, st>>>(
      (const T*)x, (const T*)y, n, (const Rect4<T>*)rects, L, e.n_tiles, e.tile_cnt, e.tile_base, indices);
  PTGS_LAUNCH_CHECK();
  return PTGS_OK;
}

// ---------------------------------------------------------------------------
// A4 with many blocks: the same selection through a uniform cell grid.  The caller
// supplies, per cell of a G x G grid over the rectangles' bounding box, the list of
// rectangles that can contain a point of that cell (a superset, built on the host
// with a one-cell margin); a point is tested against its cell's candidates only
// (N * L tests -> N * ~2), with the very same closed compares.  Tile / warp bookkeeping
// is that of expand_kernel; because the lanes of a warp now look at different
// rectangles, counts use shared-memory atomics on the warp's private counters and the
// fill pass ranks the 32 points of a sub-group through a per-warp bit mask per
// rectangle (bit = lane), which keeps every block's index list ascending.
struct CellGrid {
  double x0, y0, inv_x, inv_y;
  int G;
};

template <typename T>
__device__ __forceinline__ void cell_range(const CellGrid& g, const int32_t* __restrict__ cell_off, bool valid,
                                           T px, T py, int& beg, int& end) {
  beg = end = 0;
  if (valid && px == px && py == py) {          // NaN lies in no rectangle
    double fx = floor(((double)px - g.x0) * g.inv_x), fy = floor(((double)py - g.y0) * g.inv_y);
    const double gm = (double)(g.G - 1);
    fx = fx < 0.0 ? 0.0 : (fx > gm ? gm : fx);  // points outside the grid fail the exact tests anyway
    fy = fy < 0.0 ? 0.0 : (fy > gm ? gm : fy);
    const int c = (int)fy * g.G + (int)fx;
    beg = cell_off[c];
    end = cell_off[c + 1];
  }
}

template <typename T, bool FILL>
__global__ void __launch_bounds__(kExpWarps * 32) expand_grid_kernel(
    const T* __restrict__ x, const T* __restrict__ y, int64_t n,
    const Rect4<T>* __restrict__ rects, int L, CellGrid g, const int32_t* __restrict__ cell_off,
    const int32_t* __restrict__ cell_list, int64_t n_tiles, uint32_t* __restrict__ tile_cnt,
    const int64_t* __restrict__ tile_base, int32_t* __restrict__ indices) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  Rect4<T>* srect = reinterpret_cast<Rect4<T>*>(smem_raw);         // [L]
  long long* sbase = reinterpret_cast<long long*>(srect + L);      // [L] tile bases (FILL)
  uint32_t* wcnt = reinterpret_cast<uint32_t*>(sbase + L);         // [kExpWarps][L]
  uint32_t* wmask = wcnt + (size_t)L * kExpWarps;                  // [kExpWarps][L], FILL only
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int64_t tile = blockIdx.x;
  for (int i = threadIdx.x; i < L; i += blockDim.x) {
    srect[i] = rects[i];
    if (FILL) sbase[i] = tile_base[(size_t)i * n_tiles + tile];
  }
  for (int i = threadIdx.x; i < L * kExpWarps * (FILL ? 2 : 1); i += blockDim.x) wcnt[i] = 0;
  __syncthreads();
  const int64_t wbase = tile * kExpTile + (int64_t)warp * (32 * kExpPerLane);
  uint32_t* mycnt = wcnt + warp * L;
  uint32_t* mymask = wmask + warp * L;

  auto contains = [&](int r, T px, T py) {
    const Rect4<T> q = srect[r];
    return (px >= q.minx) && (px <= q.maxx) && (py >= q.miny) && (py <= q.maxy);
  };
  // per warp membership counts of the tile
  for (int k = 0; k < kExpPerLane; ++k) {
    const int64_t i = wbase + k * 32 + lane;
    const bool valid = i < n;
    T px = 0, py = 0;
    if (valid) { px = x[i]; py = y[i]; }
    if (!__any_sync(kFullMask, valid)) break;
    int beg, end;
    cell_range<T>(g, cell_off, valid, px, py, beg, end);
    for (int p = beg; p < end; ++p) {
      const int r = cell_list[p];
      if (contains(r, px, py)) atomicAdd(&mycnt[r], 1u);
    }
  }
  __syncthreads();
  if (!FILL) {
    for (int r = threadIdx.x; r < L; r += blockDim.x) {
      uint32_t s = 0;
#pragma unroll
      for (int w = 0; w < kExpWarps; ++w) s += wcnt[w * L + r];
      tile_cnt[(size_t)r * n_tiles + tile] = s;
    }
    return;
  }
  for (int r = threadIdx.x; r < L; r += blockDim.x) {    // exclusive scan over the warps, per block
    uint32_t run = 0;
#pragma unroll
    for (int w = 0; w < kExpWarps; ++w) {
      const uint32_t c = wcnt[w * L + r];
      wcnt[w * L + r] = run;
      run += c;
    }
  }
  __syncthreads();
  for (int k = 0; k < kExpPerLane; ++k) {
    const int64_t i = wbase + k * 32 + lane;
    const bool valid = i < n;
    T px = 0, py = 0;
    if (valid) { px = x[i]; py = y[i]; }
    if (!__any_sync(kFullMask, valid)) break;
    int beg, end;
    cell_range<T>(g, cell_off, valid, px, py, beg, end);
    // which of the cell's candidate rectangles contain the point: evaluated once, as a bit mask
    // over the candidate list (cells with more than 32 candidates re-test in every step)
    const bool wide = (end - beg) > 32;
    unsigned cm = 0;
    if (!wide) {
      for (int p = beg; p < end; ++p)
        if (contains(cell_list[p], px, py)) cm |= 1u << (p - beg);
    }
    auto in_rect = [&](int p, int r) { return wide ? contains(r, px, py) : ((cm >> (p - beg)) & 1u) != 0u; };
    for (int p = beg; p < end; ++p) {                    // a: who is in which rectangle
      if (!wide && !((cm >> (p - beg)) & 1u)) continue;
      const int r = cell_list[p];
      if (in_rect(p, r)) atomicOr(&mymask[r], 1u << lane);
    }
    __syncwarp();
    for (int p = beg; p < end; ++p) {                    // b: emit at (tile base + warp base + rank)
      if (!wide && !((cm >> (p - beg)) & 1u)) continue;
      const int r = cell_list[p];
      if (in_rect(p, r)) {
        const uint32_t m = mymask[r];
        const int64_t pos = sbase[r] + mycnt[r] + __popc(m & lanemask_lt());
        indices[pos] = (int32_t)i;
      }
    }
    __syncwarp();
    for (int p = beg; p < end; ++p) {                    // c: the lowest lane of each mask advances the counter
      if (!wide && !((cm >> (p - beg)) & 1u)) continue;
      const int r = cell_list[p];
      if (in_rect(p, r)) {
        const uint32_t m = mymask[r];
        if ((m & lanemask_lt()) == 0u) {                 // (a lane that reads an already cleared mask adds 0)
          atomicAdd(&mycnt[r], (uint32_t)__popc(m));
          mymask[r] = 0u;
        }
      }
    }
    __syncwarp();
  }
}

template <typename T>
static int expand_grid_impl(bool fill, const void* x, const void* y, int64_t n, const void* rects, int32_t L,
                            const CellGrid& g, const int32_t* cell_off, const int32_t* cell_list,
                            int64_t* counts, int64_t* offsets, int32_t* indices, void* ws, cudaStream_t st) {
  ExpandWs e = carve_expand(ws, n, L);
  const size_t smem = (size_t)L * (sizeof(Rect4<T>) + 8) + (size_t)L * kExpWarps * 4 * (fill ? 2 : 1);
  if (!fill) {
    PTGS_CUDA_TRY(cudaFuncSetAttribute(expand_grid_kernel<T, false>,
                                       cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    expand_grid_kernel<T, false><<<(unsigned)e.n_tiles, kExpWarps * 32, smem, st>>>(
        (const T*)x, (const T*)y, n, (const Rect4<T>*)rects, L, g, cell_off, cell_list, e.n_tiles, e.tile_cnt,
        nullptr, nullptr);
    PTGS_LAUNCH_CHECK();
    expand_scan_kernel<<<L, 1024, 0, st>>>(e.tile_cnt, e.n_tiles, e.tile_base, counts);
    PTGS_LAUNCH_CHECK();
    return PTGS_OK;
  }
  offsets_kernel<<<1, 1024, 0, st>>>(counts, L, offsets);
  PTGS_LAUNCH_CHECK();
  dim3 gb((unsigned)((e.n_tiles + 255) / 256 > 64 ? 64 : (e.n_tiles + 255) / 256), (unsigned)L);
  add_base_kernel<<<gb, 256, 0, st>>>(e.tile_base, e.n_tiles, offsets);
  PTGS_LAUNCH_CHECK();
  PTGS_CUDA_TRY(cudaFuncSetAttribute(expand_grid_kernel<T, true>,
                                     cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  expand_grid_kernel<T, true><<<(unsigned)e.n_tiles, kExpWarps * 32, smem, st>>>(
      (const T*)x, (const T*)y, n, (const Rect4<T>*)rects, L, g, cell_off, cell_list, e.n_tiles, e.tile_cnt,
      e.tile_base, indices);
  PTGS_LAUNCH_CHECK();
  return PTGS_OK;
}

// ---------------------------------------------------------------------------
// A5: camera centre strictly inside the expanded box (XY), float64.
// One CTA per block; cameras scanned in order so the list is ascending.
// ---------------------------------------------------------------------------
__global__ void __launch_bounds__(256) assign_count_kernel(
    const double* __restrict__ centers, int32_t C, const double* __restrict__ boxes,
    int64_t* __restrict__ counts) {
  __shared__ int sm[33];
  const int b = blockIdx.x;
  const double minx = boxes[4 * b], miny = boxes[4 * b + 1], maxx = boxes[4 * b + 2], maxy = boxes[4 * b + 3];
  int c = 0;
  for (int i = threadIdx.x; i < C; i += 256) {
    const double cx = centers[3 * i], cy = centers[3 * i + 1];
    c += (cx > minx && cx < maxx && cy > miny && cy < maxy) ? 1 : 0;
  }
  int tot;
  block_exclusive_sum<int, 256>(c, sm, &tot);
  if (threadIdx.x == 0) counts[b] = tot;
}

__global__ void __launch_bounds__(256) assign_fill_kernel(
    const double* __restrict__ centers, int32_t C, const double* __restrict__ boxes,
    const int64_t* __restrict__ offsets, int32_t* __restrict__ cam_index, int64_t capacity) {
  __shared__ int sm[33];
  const int b = blockIdx.x;
  const double minx = boxes[4 * b], miny = boxes[4 * b + 1], maxx = boxes[4 * b + 2], maxy = boxes[4 * b + 3];
  int64_t out = offsets[b];
  for (int s = 0; s < C; s += 256) {
    const int i = s + threadIdx.x;
    int f = 0;
    if (i < C) {
      const double cx = centers[3 * i], cy = centers[3 * i + 1];
      f = (cx > minx && cx < maxx && cy > miny && cy < maxy) ? 1 : 0;
    }
    int tot;
    int ex = block_exclusive_sum<int, 256>(f, sm, &tot);
    if (f && out + ex < capacity) cam_index[out + ex] = i;
    out += tot;
  }
}

// ---------------------------------------------------------------------------
// A9: per block compaction of the kept cameras.
// ---------------------------------------------------------------------------
__global__ void __launch_bounds__(256) keep_count_kernel(
    const uint8_t* __restrict__ keep, const int64_t* __restrict__ cam_offsets,
    int64_t* __restrict__ counts) {
  __shared__ int sm[33];
  const int b = blockIdx.x;
  const int64_t a = cam_offsets[b], e = cam_offsets[b + 1];
  int c = 0;
  for (int64_t i = a + threadIdx.x; i < e; i += 256) c += keep[i] ? 1 : 0;
  int tot;
  block_exclusive_sum<int, 256>(c, sm, &tot);
  if (threadIdx.x == 0) counts[b] = tot;
}

__global__ void __launch_bounds__(256) keep_fill_kernel(
    const uint8_t* __restrict__ keep, const int64_t* __restrict__ cam_offsets,
    const int32_t* __restrict__ cam_index, const int64_t* __restrict__ kept_offsets,
    int32_t* __restrict__ kept_index, int32_t* __restrict__ kept_pos) {
  __shared__ int sm[33];
  const int b = blockIdx.x;
  const int64_t a = cam_offsets[b], e = cam_offsets[b + 1];
  int64_t out = kept_offsets[b];
  for (int64_t s = a; s < e; s += 256) {
    const int64_t i = s + threadIdx.x;
    const int f = (i < e && keep[i]) ? 1 : 0;
    int tot;
    int ex = block_exclusive_sum<int, 256>(f, sm, &tot);
    if (f) {
      kept_index[out + ex] = cam_index[i];
      if (kept_pos) kept_pos[out + ex] = (int32_t)i;
    }
    out += tot;
  }
}

}  // namespace ptgs

using namespace ptgs;

extern "C" {

size_t ptgs_expand_workspace_bytes(int64_t n, int32_t n_blocks) {
  if (n_blocks < 1) n_blocks = 1;
  return carve_expand(nullptr, n, n_blocks).bytes;
}

int ptgs_expand_count(int dtype, const void* x, const void* y, int64_t n,
                      const void* rects, int32_t n_blocks, int64_t* counts,
                      void* workspace, size_t workspace_bytes,
                      ptgs_stream_t stream) {
  PTGS_REQUIRE(dtype == PTGS_F32 || dtype == PTGS_F64, "dtype");
  PTGS_REQUIRE(n >= 0 && n < (1ll << 31), "n must fit int32 indices");
  PTGS_REQUIRE(n_blocks >= 1 && n_blocks <= kExpMaxBlocks, "n_blocks must be in [1,1024]");
  PTGS_REQUIRE(rects && counts, "arguments");
  if (workspace_bytes < ptgs_expand_workspace_bytes(n, n_blocks)) {
    set_error("expand_count: workspace too small");
    return PTGS_E_WORKSPACE;
  }
  if (dtype == PTGS_F32)
    return expand_count_impl<float>(x, y, n, rects, n_blocks, counts, workspace, (cudaStream_t)stream);
  return expand_count_impl<double>(x, y, n, rects, n_blocks, counts, workspace, (cudaStream_t)stream);
}

int ptgs_expand_fill(int dtype, const void* x, const void* y, int64_t n,
                     const void* rects, int32_t n_blocks, const int64_t* counts,
                     int64_t* offsets, int32_t* indices, int64_t indices_capacity,
                     void* workspace, size_t workspace_bytes,
                     ptgs_stream_t stream) {
  PTGS_REQUIRE(dtype == PTGS_F32 || dtype == PTGS_F64, "dtype");
  PTGS_REQUIRE(n >= 0 && n < (1ll << 31), "n must fit int32 indices");
  PTGS_REQUIRE(n_blocks >= 1 && n_blocks <= kExpMaxBlocks, "n_blocks must be in [1,1024]");
  PTGS_REQUIRE(rects && counts && offsets && (indices || indices_capacity == 0), "arguments");
  (void)indices_capacity;  // the caller sized `indices` from counts (same stream order)
  if (workspace_bytes < ptgs_expand_workspace_bytes(n, n_blocks)) {
    set_error("expand_fill: workspace too small");
    return PTGS_E_WORKSPACE;
  }
  if (dtype == PTGS_F32)
    return expand_fill_impl<float>(x, y, n, rects, n_blocks, offsets, indices, workspace, counts, (cudaStream_t)stream);
  return expand_fill_impl<double>(x, y, n, rects, n_blocks, offsets, indices, workspace, counts, (cudaStream_t)stream);
}

static int expand_grid_check(int dtype, int64_t n, const void* rects, int32_t n_blocks, const double* grid,
                             const int32_t* cell_off, const int32_t* cell_list, size_t workspace_bytes,
                             CellGrid* g) {
  PTGS_REQUIRE(dtype == PTGS_F32 || dtype == PTGS_F64, "dtype");
  PTGS_REQUIRE(n >= 0 && n < (1ll << 31), "n must fit int32 indices");
  PTGS_REQUIRE(n_blocks >= 1 && n_blocks <= kExpMaxBlocks, "n_blocks must be in [1,1024]");
  PTGS_REQUIRE(rects && grid && cell_off && cell_list, "arguments");
  g->x0 = grid[0]; g->y0 = grid[1]; g->inv_x = grid[2]; g->inv_y = grid[3]; g->G = (int)grid[4];
  PTGS_REQUIRE(g->G >= 1 && g->G <= 4096, "grid size must be in [1,4096]");
  if (workspace_bytes < ptgs_expand_workspace_bytes(n, n_blocks)) {
    set_error("expand_grid: workspace too small");
    return PTGS_E_WORKSPACE;
  }
  return PTGS_OK;
}

int ptgs_expand_count_grid(int dtype, const void* x, const void* y, int64_t n, const void* rects,
                           int32_t n_blocks, const double* grid, const int32_t* cell_off,
                           const int32_t* cell_list, int64_t* counts, void* workspace,
                           size_t workspace_bytes, ptgs_stream_t stream) {
  CellGrid g;
  int rc = expand_grid_check(dtype, n, rects, n_blocks, grid, cell_off, cell_list, workspace_bytes, &g);
  if (rc) return rc;
  PTGS_REQUIRE(counts, "arguments");
  if (dtype == PTGS_F32)
    return expand_grid_impl<float>(false, x, y, n, rects, n_blocks, g, cell_off, cell_list, counts, nullptr,
                                   nullptr, workspace, (cudaStream_t)stream);
  return expand_grid_impl<double>(false, x, y, n, rects, n_blocks, g, cell_off, cell_list, counts, nullptr,
                                  nullptr, workspace, (cudaStream_t)stream);
}

int ptgs_expand_fill_grid(int dtype, const void* x, const void* y, int64_t n, const void* rects,
                          int32_t n_blocks, const double* grid, const int32_t* cell_off,
                          const int32_t* cell_list, const int64_t* counts, int64_t* offsets,
                          int32_t* indices, int64_t indices_capacity, void* workspace,
                          size_t workspace_bytes, ptgs_stream_t stream) {
  CellGrid g;
  int rc = expand_grid_check(dtype, n, rects, n_blocks, grid, cell_off, cell_list, workspace_bytes, &g);
  if (rc) return rc;
  PTGS_REQUIRE(counts && offsets && (indices || indices_capacity == 0), "arguments");
  if (dtype == PTGS_F32)
    return expand_grid_impl<float>(true, x, y, n, rects, n_blocks, g, cell_off, cell_list,
                                   const_cast<int64_t*>(counts), offsets, indices, workspace, (cudaStream_t)stream);
  return expand_grid_impl<double>(true, x, y, n, rects, n_blocks, g, cell_off, cell_list,
                                  const_cast<int64_t*>(counts), offsets, indices, workspace, (cudaStream_t)stream);
}

int ptgs_assign_cameras(const double* centers, int32_t n_cameras,
                        const double* boxes, int32_t n_blocks,
                        int64_t* cam_offsets, int32_t* cam_index,
                        int64_t cam_capacity, ptgs_stream_t stream) {
  PTGS_REQUIRE(n_cameras >= 0 && n_blocks >= 1 && n_blocks <= kExpMaxBlocks, "sizes");
  PTGS_REQUIRE(boxes && cam_offsets && (centers || n_cameras == 0), "arguments");
  cudaStream_t st = (cudaStream_t)stream;
  int64_t* counts = cam_offsets;   // scanned in place
  assign_count_kernel<<<n_blocks, 256, 0, st>>>(centers, n_cameras, boxes, counts);
  PTGS_LAUNCH_CHECK();
  offsets_kernel<<<1, 1024, 0, st>>>(counts, n_blocks, cam_offsets);
  PTGS_LAUNCH_CHECK();
  if (cam_index && cam_capacity > 0) {
    assign_fill_kernel<<<n_blocks, 256, 0, st>>>(centers, n_cameras, boxes, cam_offsets, cam_index, cam_capacity);
    PTGS_LAUNCH_CHECK();
  }
  return PTGS_OK;
}

int ptgs_select_cameras(const uint8_t* keep, const int64_t* cam_offsets,
                        const int32_t* cam_index, int32_t n_blocks,
                        int64_t* kept_offsets, int32_t* kept_index,
                        int32_t* kept_pos, ptgs_stream_t stream) {
  PTGS_REQUIRE(n_blocks >= 1 && n_blocks <= kExpMaxBlocks, "n_blocks");
  PTGS_REQUIRE(cam_offsets && kept_offsets, "arguments");
  cudaStream_t st = (cudaStream_t)stream;
  int64_t* counts = kept_offsets;  // scanned in place
  keep_count_kernel<<<n_blocks, 256, 0, st>>>(keep, cam_offsets, counts);
  PTGS_LAUNCH_CHECK();
  offsets_kernel<<<1, 1024, 0, st>>>(counts, n_blocks, kept_offsets);
  PTGS_LAUNCH_CHECK();
  if (kept_index) {
    keep_fill_kernel<<<n_blocks, 256, 0, st>>>(keep, cam_offsets, cam_index, kept_offsets, kept_index,
                                               kept_pos);
    PTGS_LAUNCH_CHECK();
  }
  return PTGS_OK;
}

}  // extern "C"
