// libptgs multi-GPU exchange of the per-pair visibility records.
// Reference: the loop being split is the per-camera loop of
// visibility_based_camera_selection, scene/ptgs/shen_partation_densify.py:932-944;
// the reference is single process, so the exchange has no counterpart there.
//
// With range sharding a rank owns one or a few CONTIGUOUS ranges of the pair
// list.  One kernel packs the rank's result vectors into 24-byte records, one
// all-gather (NCCL, issued by the caller on the same stream) delivers every
// rank's slot, one kernel unpacks the gathered slots straight into the full
// per-pair vectors (range table by value).  No zero-filled P x world table,
// no reduction.
#include "common.cuh"

namespace ptgs {

struct alignas(8) PairRecord {
  double ratio;
  double area;
  int32_t n_hit;
  uint8_t keep;
  uint8_t flags;
  uint16_t pad;
};
static_assert(sizeof(PairRecord) == PTGS_PAIR_RECORD_BYTES, "record layout");

constexpr int kMaxRanges = 128;     // world * chunks
struct RangeTable {
  int64_t lo[kMaxRanges + 1];       // range c = pairs [lo[c], lo[c+1])
  int64_t base[kMaxRanges];         // record position of the range's first pair in the gathered table
};

__global__ void __launch_bounds__(256) pack_records_kernel(
    const int32_t* __restrict__ n_hit, const double* __restrict__ ratio,
    const double* __restrict__ area, const uint8_t* __restrict__ keep,
    const uint8_t* __restrict__ flags, int64_t n, PairRecord* __restrict__ out) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  PairRecord r;
  r.ratio = ratio[i]; r.area = area[i]; r.n_hit = n_hit[i];
  r.keep = keep[i]; r.flags = flags[i]; r.pad = 0;
  out[i] = r;
}

__global__ void __launch_bounds__(256) unpack_records_kernel(
    const PairRecord* __restrict__ gathered, int n_ranges, RangeTable rt,
    int64_t n_pairs, int32_t* __restrict__ n_hit, double* __restrict__ ratio,
    double* __restrict__ area, uint8_t* __restrict__ keep, uint8_t* __restrict__ flags) {
  const int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= n_pairs) return;
  int lo = 0, hi = n_ranges - 1;               // last range with lo[c] <= p (empty ranges share a bound)
  while (lo < hi) {
    const int mid = (lo + hi + 1) >> 1;
    if (rt.lo[mid] <= p) lo = mid; else hi = mid - 1;
  }
  const PairRecord rec = gathered[rt.base[lo] + (p - rt.lo[lo])];
  n_hit[p] = rec.n_hit; ratio[p] = rec.ratio; area[p] = rec.area;
  keep[p] = rec.keep; flags[p] = rec.flags;
}

}  // namespace ptgs

using namespace ptgs;

extern "C" {

int ptgs_pack_records(const int32_t* n_hit, const double* ratio, const double* area,
                      const uint8_t* keep, const uint8_t* flags, int64_t n_pairs,
                      void* records, ptgs_stream_t stream) {
  PTGS_REQUIRE(n_pairs >= 0, "n_pairs");
  if (n_pairs == 0) return PTGS_OK;
  PTGS_REQUIRE(n_hit && ratio && area && keep && flags && records, "arguments");
  PTGS_REQUIRE(((uintptr_t)records & 7) == 0, "records must be 8-byte aligned");
  pack_records_kernel<<<(unsigned)((n_pairs + 255) / 256), 256, 0, (cudaStream_t)stream>>>(
      n_hit, ratio, area, keep, flags, n_pairs, (PairRecord*)records);
  PTGS_LAUNCH_CHECK();
  return PTGS_OK;
}

int ptgs_unpack_records(const void* gathered, int32_t n_ranges, int64_t table_records,
                        const int64_t* range_lo_host, const int64_t* range_base_host, int64_t n_pairs,
                        int32_t* n_hit, double* ratio, double* area, uint8_t* keep, uint8_t* flags,
                        ptgs_stream_t stream) {
  PTGS_REQUIRE(n_ranges >= 1 && n_ranges <= kMaxRanges, "n_ranges must be in [1,128]");
  PTGS_REQUIRE(n_pairs >= 0 && table_records >= 0 && range_lo_host && range_base_host, "arguments");
  if (n_pairs == 0) return PTGS_OK;
  PTGS_REQUIRE(gathered && n_hit && ratio && area && keep && flags, "arguments");
  RangeTable rt;
  for (int c = 0; c <= n_ranges; ++c) rt.lo[c] = range_lo_host[c];
  for (int c = 0; c < n_ranges; ++c) rt.base[c] = range_base_host[c];
  PTGS_REQUIRE(rt.lo[0] == 0 && rt.lo[n_ranges] == n_pairs, "ranges must tile [0, n_pairs)");
  for (int c = 0; c < n_ranges; ++c)
    PTGS_REQUIRE(rt.lo[c + 1] >= rt.lo[c] && rt.base[c] >= 0 &&
                     rt.base[c] + (rt.lo[c + 1] - rt.lo[c]) <= table_records,
                 "range lies outside the gathered table");
  unpack_records_kernel<<<(unsigned)((n_pairs + 255) / 256), 256, 0, (cudaStream_t)stream>>>(
      (const PairRecord*)gathered, n_ranges, rt, n_pairs, n_hit, ratio, area, keep, flags);
  PTGS_LAUNCH_CHECK();
  return PTGS_OK;
}

}  // extern "C"
