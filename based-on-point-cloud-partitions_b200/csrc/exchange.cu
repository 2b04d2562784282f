// libptgs multi-GPU exchange of the per-pair visibility records.
// Reference: the loop being split is the per-camera loop of
// visibility_based_camera_selection, scene/ptgs/shen_partation_densify.py:932-944;
// the reference is single process, so the exchange has no counterpart there.
//
// With range sharding rank r owns the CONTIGUOUS pair range [lo_r, lo_{r+1}).
// One kernel packs the rank's result vectors into 24-byte records, one
// all-gather (NCCL, issued by the caller on the same stream) delivers every
// rank's slice, one kernel unpacks the gathered slots straight into the full
// per-pair vectors.  No zero-filled P x world table, no reduction.
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

constexpr int kMaxRanks = 64;
struct RankBounds { int64_t lo[kMaxRanks + 1]; };

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
    const PairRecord* __restrict__ gathered, int world, int64_t slot_pairs, RankBounds rb,
    int64_t n_pairs, int32_t* __restrict__ n_hit, double* __restrict__ ratio,
    double* __restrict__ area, uint8_t* __restrict__ keep, uint8_t* __restrict__ flags) {
  const int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= n_pairs) return;
  int r = 0;                                   // last rank with lo[r] <= p (empty ranks share a bound)
  for (int k = 1; k < world; ++k) r = (rb.lo[k] <= p) ? k : r;
  const PairRecord rec = gathered[(int64_t)r * slot_pairs + (p - rb.lo[r])];
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

int ptgs_unpack_records(const void* gathered, int32_t world, int64_t slot_pairs,
                        const int64_t* rank_lo_host, int64_t n_pairs, int32_t* n_hit,
                        double* ratio, double* area, uint8_t* keep, uint8_t* flags,
                        ptgs_stream_t stream) {
  PTGS_REQUIRE(world >= 1 && world <= kMaxRanks, "world must be in [1,64]");
  PTGS_REQUIRE(n_pairs >= 0 && slot_pairs >= 0 && rank_lo_host, "arguments");
  if (n_pairs == 0) return PTGS_OK;
  PTGS_REQUIRE(gathered && n_hit && ratio && area && keep && flags, "arguments");
  RankBounds rb;
  for (int r = 0; r <= world; ++r) rb.lo[r] = rank_lo_host[r];
  PTGS_REQUIRE(rb.lo[0] == 0 && rb.lo[world] == n_pairs, "rank bounds must tile [0, n_pairs)");
  for (int r = 0; r < world; ++r)
    PTGS_REQUIRE(rb.lo[r + 1] >= rb.lo[r] && rb.lo[r + 1] - rb.lo[r] <= slot_pairs,
                 "rank range exceeds the gathered slot");
  unpack_records_kernel<<<(unsigned)((n_pairs + 255) / 256), 256, 0, (cudaStream_t)stream>>>(
      (const PairRecord*)gathered, world, slot_pairs, rb, n_pairs, n_hit, ratio, area, keep, flags);
  PTGS_LAUNCH_CHECK();
  return PTGS_OK;
}

}  // extern "C"
