// libptgs A7-A9: camera visibility sweep.
// Reference: project_points_to_camera (scene/ptgs/shen_partation_densify.py:
// 410-477), run_graham_scan (:481-539), process_camera_visibility (:890-910).
//
// One persistent kernel (3 CTAs of 256 threads per SM) pulls (block, point tile) work items from a
// global counter.  A tile of 1024 points (kSwP = 4 per thread) is staged into shared memory with 1-D
// bulk copies (cp.async.bulk, mbarrier completion, double buffered so the next tile lands while the
// current one is being swept) and moved into registers; each warp owns 128 contiguous points.
// Three box levels come before any per-point work: the CTA culls the block's cameras against the
// tile box (tiles wholly inside a camera's interior region are counted at once), every warp
// classifies the surviving cameras against its own run's box, and the (run, camera) pairs that
// still need points go to a per-CTA queue served by whichever warp is free.  Per point: the camera
// transform in the k-ordered accumulation of the reference's SGEMM, division-free homogeneous
// "surely in" / "surely out" tests, and IEEE divisions only for the band in between, compacted
// into dense 32-lane groups.  Counts are warp uniform (ballot / popc); one global atomic per
// (pass, camera).
//
//   MODE 0  hit counts (+ optional per-point "seen" flag)           -- coverage sweep
//   MODE 1  hit counts + Graham pivot + 16 directional extremes     -- sweep 1 (and its sample pass)
//   MODE 2  hull candidates of the undecided pairs                  -- sweep 2
#include <cstddef>
#include <cstdlib>

#include "common.cuh"

namespace ptgs {

constexpr int kSwThreads = 256;
constexpr int kSwP = 4;                          // points per thread
constexpr int kSwCtasPerSM = 3;
constexpr int kSwTile = kSwThreads * kSwP;       // 1024 points per tile
constexpr int kSwStageFloats = kSwTile + 8;      // + alignment slack
constexpr int kSwCamChunk = 160;                 // cameras staged per chunk (5 lane groups); MODE 2: half (wider polygon rows)
constexpr int kDirs = PTGS_EXTREME_DIRS;         // directional extremes per pair (8 or 16)
constexpr int kDirStep = 16 / kDirs;             // stride through the table of 16 directions
static_assert(kDirs == 8 || kDirs == 16, "PTGS_EXTREME_DIRS");
constexpr int kSwMaxBlocks = 1024;
constexpr int kCamFloats = PTGS_CAMERA_FLOATS;   // 20
constexpr int kPolyFloats = PTGS_POLY_FLOATS;    // 28
constexpr int kKrectFloats = PTGS_KRECT_FLOATS;  // 32

// ---- mbarrier / bulk copy (TMA) -------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return (uint32_t)__cvta_generic_to_shared(p);
}
__device__ __forceinline__ void mbar_init(uint64_t* bar, int count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)),
               "r"(bytes)
               : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t phase) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "WAIT_LOOP:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
      "@p bra.uni WAIT_DONE;\n"
      "bra.uni WAIT_LOOP;\n"
      "WAIT_DONE:\n"
      "}\n" ::"r"(smem_u32(bar)),
      "r"(phase)
      : "memory");
}
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, uint32_t bytes,
                                         uint64_t* bar) {
  asm volatile(
      "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::
          "r"(smem_u32(dst)),
      "l"(src), "r"(bytes), "r"(smem_u32(bar))
      : "memory");
}
__device__ __forceinline__ void fence_proxy_async() {
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}

// ---- one visibility test ----------------------------------------------------
struct Cam {
  float r00, r01, r02, r10, r11, r12, r20, r21, r22, tx, ty, tz;
  float fx, fy, cx, cy, w, h, front_axis, pad;
};

__device__ __forceinline__ Cam load_cam(const float* s) {
  const float4* q = reinterpret_cast<const float4*>(s);
  const float4 a = q[0], b = q[1], c = q[2], d = q[3], e = q[4];
  Cam k;
  k.r00 = a.x; k.r01 = a.y; k.r02 = a.z; k.r10 = a.w;
  k.r11 = b.x; k.r12 = b.y; k.r20 = b.z; k.r21 = b.w;
  k.r22 = c.x; k.tx = c.y; k.ty = c.z; k.tz = c.w;
  k.fx = d.x; k.fy = d.y; k.cx = d.z; k.cy = d.w;
  k.w = e.x; k.h = e.y; k.front_axis = e.z; k.pad = e.w;
  return k;
}

// camera space, k-ordered accumulation (r0*x, fma r1*y, fma r2*z, + t): the
// order of the 4x4 @ 4xN SGEMM at densify:444.
__device__ __forceinline__ float cam_row(float r0, float r1, float r2, float t, float x, float y,
                                         float z) {
  float a = __fmul_rn(r0, x);
  a = __fmaf_rn(r1, y, a);
  a = __fmaf_rn(r2, z, a);
  return __fadd_rn(a, t);
}

struct Proj {
  float uh, vh, Z;
  bool maybe;
};

// everything up to (not including) the divisions; `maybe` is a conservative
// superset of the hits (see DESIGN.md "division-free reject").
__device__ __forceinline__ Proj project_pre(const Cam& c, float w1, float h1, float x, float y,
                                            float z) {
  const float X = cam_row(c.r00, c.r01, c.r02, c.tx, x, y, z);
  const float Y = cam_row(c.r10, c.r11, c.r12, c.ty, x, y, z);
  const float Z = cam_row(c.r20, c.r21, c.r22, c.tz, x, y, z);
  const float front = (c.front_axis == 1.0f) ? Y : Z;              // densify:447
  Proj p;
  p.Z = Z;
  p.uh = __fmaf_rn(c.cx, Z, __fmul_rn(c.fx, X));                   // densify:467
  p.vh = __fmaf_rn(c.cy, Z, __fmul_rn(c.fy, Y));
  const uint32_t sgn = __float_as_uint(Z) & 0x80000000u;
  const float a = __uint_as_float(__float_as_uint(p.uh) ^ sgn);
  const float b = __uint_as_float(__float_as_uint(p.vh) ^ sgn);
  const float az = fabsf(Z);
  p.maybe = (front > 0.0f) && (a >= -az) && (a <= __fmul_rn(w1, az)) && (b >= -az) &&
            (b <= __fmul_rn(h1, az));
  return p;
}

__device__ __forceinline__ bool project_exact(const Cam& c, const Proj& p, float& u, float& v) {
  u = __fdiv_rn(p.uh, p.Z);                                        // densify:468
  v = __fdiv_rn(p.vh, p.Z);
  return (u >= 0.0f) && (u < c.w) && (v >= 0.0f) && (v < c.h);     // densify:470-471
}

__device__ __forceinline__ uint32_t order_key(float f) {
  const uint32_t b = __float_as_uint(f);
  return (b & 0x80000000u) ? ~b : (b | 0x80000000u);
}
__device__ __forceinline__ float key_to_float(uint32_t k) {
  return __uint_as_float((k & 0x80000000u) ? (k & 0x7fffffffu) : ~k);
}

// The directions (a, b) along which a pair's extreme hits are tracked, counter-clockwise from +u:
// (1,0) (2,1) (1,1) (1,2) (0,1) (-1,2) (-1,1) (-2,1) and their negatives; kDirs = 8 uses every other one
// (the axes and diagonals).  value = a*u + b*v with ONE float32 rounding (2u / 2v are exact).
__constant__ float kDirA[16] = {1, 2, 1, 1, 0, -1, -1, -2, -1, -2, -1, -1, 0, 1, 1, 2};
__constant__ float kDirB[16] = {0, 1, 1, 2, 1, 2, 1, 1, 0, -1, -1, -2, -1, -2, -1, -1};
// value of table direction t (0..15); u2 = 2u, v2 = 2v.  t is a compile-time constant at every use.
__device__ __forceinline__ float dir_value(int t, float u, float v, float u2, float v2) {
  float r;
  switch (t & 7) {
    case 0: r = u; break;
    case 1: r = __fadd_rn(u2, v); break;
    case 2: r = __fadd_rn(u, v); break;
    case 3: r = __fadd_rn(u, v2); break;
    case 4: r = v; break;
    case 5: r = __fsub_rn(v2, u); break;
    case 6: r = __fsub_rn(v, u); break;
    default: r = __fsub_rn(v, u2); break;
  }
  return (t >= 8) ? -r : r;
}

// ---- conservative tile cull ----------------------------------------------------
// A tile's points lie in the axis aligned box (centre c, half extent e).  The
// hit set of a camera is {front > 0} intersected with the double pyramid
//   Z > 0:  uh >= 0, W*Z - uh > 0, vh >= 0, H*Z - vh > 0      (nappe +)
//   Z < 0:  uh <= 0, W*Z - uh < 0, vh <= 0, H*Z - vh < 0      (nappe -, the
//           mirrored points of SURVEY.md D5)
// every constraint being a linear form g(p) = n.p + d of the world point.  Over
// the box g ranges over [g(c) - |n|.e, g(c) + |n|.e]; a nappe is impossible when
// one of its constraints fails on the whole box.  The float32 evaluation of the
// real test differs from exact arithmetic by at most ~2^-21 * S in camera space
// (S = magnitude of the coordinates involved); the margins below are 2^-16 * S,
// a 32x safety factor (DESIGN.md "tile cull").  Comparisons are written so that
// a NaN keeps the tile.
struct Form { float mid, rad; };
__device__ __forceinline__ Form form(float n0, float n1, float n2, float d, const float* b) {
  Form f;
  f.mid = fmaf(n0, b[0], fmaf(n1, b[1], fmaf(n2, b[2], d)));
  f.rad = fmaf(fabsf(n0), b[3], fmaf(fabsf(n1), b[4], fabsf(n2) * b[5]));
  return f;
}
__device__ __forceinline__ bool tile_may_hit(const Cam& c, const float* b) {
  const float S = fabsf(b[0]) + fabsf(b[1]) + fabsf(b[2]) + b[3] + b[4] + b[5] +
                  fmaxf(fabsf(c.tx), fmaxf(fabsf(c.ty), fabsf(c.tz)));
  const float mF = S * (1.0f / 65536.0f);
  const float mP = mF * (c.fx + c.fy + c.w + c.h + fabsf(c.cx) + fabsf(c.cy));
  const Form Y = form(c.r10, c.r11, c.r12, c.ty, b);
  const Form Z = form(c.r20, c.r21, c.r22, c.tz, b);
  const Form F = (c.front_axis == 1.0f) ? Y : Z;
  if (F.mid + F.rad < -mF) return false;
  // uh = fx*X + cx*Z ; vh = fy*Y + cy*Z ; B = (W-cx)*Z - fx*X ; D = (H-cy)*Z - fy*Y
  const float a0 = fmaf(c.fx, c.r00, c.cx * c.r20), a1 = fmaf(c.fx, c.r01, c.cx * c.r21),
              a2 = fmaf(c.fx, c.r02, c.cx * c.r22), ad = fmaf(c.fx, c.tx, c.cx * c.tz);
  const float c0 = fmaf(c.fy, c.r10, c.cy * c.r20), c1 = fmaf(c.fy, c.r11, c.cy * c.r21),
              c2 = fmaf(c.fy, c.r12, c.cy * c.r22), cd = fmaf(c.fy, c.ty, c.cy * c.tz);
  const Form A = form(a0, a1, a2, ad, b);
  const Form Cc = form(c0, c1, c2, cd, b);
  const Form B = form(fmaf(c.w, c.r20, -a0), fmaf(c.w, c.r21, -a1), fmaf(c.w, c.r22, -a2),
                      fmaf(c.w, c.tz, -ad), b);
  const Form D = form(fmaf(c.h, c.r20, -c0), fmaf(c.h, c.r21, -c1), fmaf(c.h, c.r22, -c2),
                      fmaf(c.h, c.tz, -cd), b);
  const bool plus = !(Z.mid + Z.rad < -mF) && !(A.mid + A.rad < -mP) && !(B.mid + B.rad < -mP) &&
                    !(Cc.mid + Cc.rad < -mP) && !(D.mid + D.rad < -mP);
  const bool minus = !(Z.mid - Z.rad > mF) && !(A.mid - A.rad > mP) && !(B.mid - B.rad > mP) &&
                     !(Cc.mid - Cc.rad > mP) && !(D.mid - D.rad > mP);
  return plus || minus;
}

// ---- division-free classification (MODE 0 / 1) ------------------------------------
// A rectangle {|u - cu| <= hu, |v - cv| <= hv} of the image plane is tested in
// homogeneous form, |uh - cu*Z| <= hu*|Z|, which holds for both nappes (Z < 0
// flips both sides).  Two rectangles per camera (der[0..3] = "in", der[4..7] =
// "out"):
//   out  = the image, slightly inflated: a point outside it cannot be a hit.
//          hit  =>  0 <= RN(uh/Z) < W  =>  |uh/Z - hW| <= W - hW  with hW =
//          pred(W)/2; the inflation 2^-20 covers W - hW = hW(1 + <=2^-22), the
//          single rounding of the fma and of the product (2^-24 each); the
//          1e-30 slack covers underflow and uh = Z = 0 (-> exact path -> NaN).
//   in   = a rectangle all of whose points are hits that need no further work:
//          MODE 0: the image deflated by 2^-20 (strict compare, so |Z| = 0 never
//          passes); MODE 1: the per-pair interior rectangle K of ptgs_visibility_krect
//          (>= 1 pixel inside the convex hull of known hits, hence inside the
//          image and never a pivot / directional extreme).
// Points in "in" are counted; points in "out" but not "in" take the exact path
// (IEEE divisions, the reference's compares); the rest are misses.
__device__ __forceinline__ void make_rects(const Cam& c, const float* krect, int mode, float* der) {
  const float pW = (c.w > 0.0f) ? __uint_as_float(__float_as_uint(c.w) - 1u) : 0.0f;
  const float pH = (c.h > 0.0f) ? __uint_as_float(__float_as_uint(c.h) - 1u) : 0.0f;
  const float hW = 0.5f * pW, hH = 0.5f * pH;
  const float up = 1.0f + 9.5367431640625e-07f, dn = 1.0f - 9.5367431640625e-07f;   // 1 +- 2^-20
  der[4] = hW; der[5] = hH; der[6] = hW * up; der[7] = hH * up;
  if (mode == 0) {
    der[0] = hW; der[1] = hH; der[2] = hW * dn; der[3] = hH * dn;
  } else if (krect) {
    der[0] = krect[0]; der[1] = krect[1]; der[2] = krect[2]; der[3] = krect[3];
  } else {
    der[0] = 0.0f; der[1] = 0.0f; der[2] = -1.0f; der[3] = -1.0f;
  }
}

// Is every point of the tile box strictly inside the "in" rectangle of the
// positive nappe and in front?  Same interval arithmetic and margins as
// tile_may_hit; the four edge forms are built from combined coefficients.
__device__ __forceinline__ bool tile_all_in(const Cam& c, const float* der, const float* b) {
  if (!(der[2] > 0.0f) || !(der[3] > 0.0f)) return false;
  const float S = fabsf(b[0]) + fabsf(b[1]) + fabsf(b[2]) + b[3] + b[4] + b[5] +
                  fmaxf(fabsf(c.tx), fmaxf(fabsf(c.ty), fabsf(c.tz)));
  const float mF = S * (1.0f / 65536.0f);
  const float mP = mF * (c.fx + c.fy + c.w + c.h + fabsf(c.cx) + fabsf(c.cy));
  const Form Y = form(c.r10, c.r11, c.r12, c.ty, b);
  const Form Z = form(c.r20, c.r21, c.r22, c.tz, b);
  const Form F = (c.front_axis == 1.0f) ? Y : Z;
  if (!(F.mid - F.rad > mF) || !(Z.mid - Z.rad > mF)) return false;
  const float u1 = der[0] - der[2], u2 = der[0] + der[2], v1 = der[1] - der[3], v2 = der[1] + der[3];
  // uh - u1*Z >= 0, u2*Z - uh >= 0, vh - v1*Z >= 0, v2*Z - vh >= 0 over the box
  const float e1 = c.cx - u1, e2 = u2 - c.cx, g1 = c.cy - v1, g2 = v2 - c.cy;
  const Form A1 = form(fmaf(c.fx, c.r00, e1 * c.r20), fmaf(c.fx, c.r01, e1 * c.r21),
                       fmaf(c.fx, c.r02, e1 * c.r22), fmaf(c.fx, c.tx, e1 * c.tz), b);
  const Form A2 = form(fmaf(-c.fx, c.r00, e2 * c.r20), fmaf(-c.fx, c.r01, e2 * c.r21),
                       fmaf(-c.fx, c.r02, e2 * c.r22), fmaf(-c.fx, c.tx, e2 * c.tz), b);
  const Form B1 = form(fmaf(c.fy, c.r10, g1 * c.r20), fmaf(c.fy, c.r11, g1 * c.r21),
                       fmaf(c.fy, c.r12, g1 * c.r22), fmaf(c.fy, c.ty, g1 * c.tz), b);
  const Form B2 = form(fmaf(-c.fy, c.r10, g2 * c.r20), fmaf(-c.fy, c.r11, g2 * c.r21),
                       fmaf(-c.fy, c.r12, g2 * c.r22), fmaf(-c.fy, c.ty, g2 * c.tz), b);
  return (A1.mid - A1.rad > mP) && (A2.mid - A2.rad > mP) && (B1.mid - B1.rad > mP) &&
         (B2.mid - B2.rad > mP);
}

// Same question for the interior polygon Q of a pair (krect record: [4] = number
// of edges, [8 + 3e ..] = (a, b, c) with a*u + b*v + c >= 0 inside, |(a, b)| = 1,
// c already reduced by the safety margin): a*uh + b*vh + c*Z >= 0 over the box.
// Returns the mask of polygon edges that are NOT satisfied by every point of the
// box (0xffffffff: no polygon); `fz` = front > 0 and Z > 0 hold on the whole box.
// Mask 0 with fz: every point of the box projects into Q.  A point with front > 0
// and Z > 0 that satisfies the masked edges projects into Q as well.
__device__ __forceinline__ unsigned poly_undecided(const Cam& c, const float* kq, const float* b,
                                                   bool& fz) {
  fz = false;
  const int ne = (int)kq[4];
  if (ne < 3) return 0xffffffffu;
  const float S = fabsf(b[0]) + fabsf(b[1]) + fabsf(b[2]) + b[3] + b[4] + b[5] +
                  fmaxf(fabsf(c.tx), fmaxf(fabsf(c.ty), fabsf(c.tz)));
  const float mF = S * (1.0f / 65536.0f);
  const float mP = mF * (c.fx + c.fy + c.w + c.h + fabsf(c.cx) + fabsf(c.cy));
  const Form Y = form(c.r10, c.r11, c.r12, c.ty, b);
  const Form Z = form(c.r20, c.r21, c.r22, c.tz, b);
  const Form F = (c.front_axis == 1.0f) ? Y : Z;
  fz = (F.mid - F.rad > mF) && (Z.mid - Z.rad > mF);
  // uh = A.p + ad, vh = C.p + cd as linear forms of the world point
  const float a0 = fmaf(c.fx, c.r00, c.cx * c.r20), a1 = fmaf(c.fx, c.r01, c.cx * c.r21),
              a2 = fmaf(c.fx, c.r02, c.cx * c.r22), ad = fmaf(c.fx, c.tx, c.cx * c.tz);
  const float c0 = fmaf(c.fy, c.r10, c.cy * c.r20), c1 = fmaf(c.fy, c.r11, c.cy * c.r21),
              c2 = fmaf(c.fy, c.r12, c.cy * c.r22), cd = fmaf(c.fy, c.ty, c.cy * c.tz);
  unsigned und = 0;
  for (int e = 0; e < ne; ++e) {
    const float ea = kq[8 + 3 * e], eb = kq[9 + 3 * e], ec = kq[10 + 3 * e];
    const Form G = form(fmaf(ea, a0, fmaf(eb, c0, ec * c.r20)), fmaf(ea, a1, fmaf(eb, c1, ec * c.r21)),
                        fmaf(ea, a2, fmaf(eb, c2, ec * c.r22)), fmaf(ea, ad, fmaf(eb, cd, ec * c.tz)), b);
    if (!(G.mid - G.rad > mP)) und |= 1u << e;
  }
  return und;
}
__device__ __forceinline__ bool tile_all_in_poly(const Cam& c, const float* kq, const float* b) {
  bool fz;
  return poly_undecided(c, kq, b, fz) == 0u && fz;
}

// MODE 2: does every point of the box project strictly inside the pair's inner
// polygon (poly record of ptgs_visibility_bounds: [0] = edges, [3] = margin,
// [4 + 3e ..] = (a, b, c), inside when a*u + b*v + c > margin)?  Such points are
// not hull candidates, so the warp skips the camera.
__device__ __forceinline__ bool inner_all_in(const Cam& c, const float* pl, const float* b) {
  const int ne = (int)pl[0];
  if (ne < 3) return false;
  const float S = fabsf(b[0]) + fabsf(b[1]) + fabsf(b[2]) + b[3] + b[4] + b[5] +
                  fmaxf(fabsf(c.tx), fmaxf(fabsf(c.ty), fabsf(c.tz)));
  const float mF = S * (1.0f / 65536.0f);
  const Form Y = form(c.r10, c.r11, c.r12, c.ty, b);
  const Form Z = form(c.r20, c.r21, c.r22, c.tz, b);
  const Form F = (c.front_axis == 1.0f) ? Y : Z;
  if (!(F.mid - F.rad > mF) || !(Z.mid - Z.rad > mF)) return false;
  const float a0 = fmaf(c.fx, c.r00, c.cx * c.r20), a1 = fmaf(c.fx, c.r01, c.cx * c.r21),
              a2 = fmaf(c.fx, c.r02, c.cx * c.r22), ad = fmaf(c.fx, c.tx, c.cx * c.tz);
  const float c0 = fmaf(c.fy, c.r10, c.cy * c.r20), c1 = fmaf(c.fy, c.r11, c.cy * c.r21),
              c2 = fmaf(c.fy, c.r12, c.cy * c.r22), cd = fmaf(c.fy, c.ty, c.cy * c.tz);
  const float margin = pl[3];
  bool in = true;
  for (int e = 0; e < ne; ++e) {
    const float ea = pl[4 + 3 * e], eb = pl[5 + 3 * e];
    // twice the margin: the per-point test evaluates f in float32 with its own error
    const float ec = pl[6 + 3 * e] - 2.0f * margin;
    const float mPe = mF * (fabsf(ea) * (c.fx + fabsf(c.cx)) + fabsf(eb) * (c.fy + fabsf(c.cy)) + fabsf(ec));
    const Form G = form(fmaf(ea, a0, fmaf(eb, c0, ec * c.r20)), fmaf(ea, a1, fmaf(eb, c1, ec * c.r21)),
                        fmaf(ea, a2, fmaf(eb, c2, ec * c.r22)), fmaf(ea, ad, fmaf(eb, cd, ec * c.tz)), b);
    in = in && (G.mid - G.rad > mPe);
  }
  return in;
}

// inner_all_in evaluated by a whole warp for ONE camera: lane e owns polygon edge e (same forms,
// same margins; all lanes must call it).
__device__ __forceinline__ bool inner_all_in_warp(const Cam& c, const float* pl, const float* b, int lane) {
  const int ne = (int)pl[0];
  if (ne < 3) return false;
  const float S = fabsf(b[0]) + fabsf(b[1]) + fabsf(b[2]) + b[3] + b[4] + b[5] +
                  fmaxf(fabsf(c.tx), fmaxf(fabsf(c.ty), fabsf(c.tz)));
  const float mF = S * (1.0f / 65536.0f);
  const Form Y = form(c.r10, c.r11, c.r12, c.ty, b);
  const Form Z = form(c.r20, c.r21, c.r22, c.tz, b);
  const Form F = (c.front_axis == 1.0f) ? Y : Z;
  if (!(F.mid - F.rad > mF) || !(Z.mid - Z.rad > mF)) return false;     // warp uniform
  bool ok = true;
  if (lane < ne) {
    const float a0 = fmaf(c.fx, c.r00, c.cx * c.r20), a1 = fmaf(c.fx, c.r01, c.cx * c.r21),
                a2 = fmaf(c.fx, c.r02, c.cx * c.r22), ad = fmaf(c.fx, c.tx, c.cx * c.tz);
    const float c0 = fmaf(c.fy, c.r10, c.cy * c.r20), c1 = fmaf(c.fy, c.r11, c.cy * c.r21),
                c2 = fmaf(c.fy, c.r12, c.cy * c.r22), cd = fmaf(c.fy, c.ty, c.cy * c.tz);
    const float margin = pl[3];
    const float ea = pl[4 + 3 * lane], eb = pl[5 + 3 * lane];
    const float ec = pl[6 + 3 * lane] - 2.0f * margin;
    const float mPe = mF * (fabsf(ea) * (c.fx + fabsf(c.cx)) + fabsf(eb) * (c.fy + fabsf(c.cy)) + fabsf(ec));
    const Form G = form(fmaf(ea, a0, fmaf(eb, c0, ec * c.r20)), fmaf(ea, a1, fmaf(eb, c1, ec * c.r21)),
                        fmaf(ea, a2, fmaf(eb, c2, ec * c.r22)), fmaf(ea, ad, fmaf(eb, cd, ec * c.tz)), b);
    ok = G.mid - G.rad > mPe;
  }
  return __all_sync(kFullMask, ok);
}

// ---- work item schedule -----------------------------------------------------
// item_prefix[b] = number of work items of the blocks before b; a block has
// tiles_b * split_b items, split_b = min(cam_split, C_b).
__global__ void __launch_bounds__(1024) sweep_schedule_kernel(
    const int64_t* __restrict__ pt_offsets, const int64_t* __restrict__ cam_offsets, int L,
    int cam_split, int64_t* __restrict__ item_prefix, int n_ctas) {
  __shared__ long long sm[33];
  long long carry = 0;
  for (int s = 0; s < L; s += 1024) {
    const int b = s + threadIdx.x;
    long long v = 0;
    if (b < L) {
      const int64_t m = pt_offsets[b + 1] - pt_offsets[b];
      const int64_t c = cam_offsets[b + 1] - cam_offsets[b];
      const int64_t sp = c < cam_split ? c : cam_split;
      v = ((m + kSwTile - 1) / kSwTile) * sp;
    }
    long long tot;
    long long ex = block_exclusive_sum<long long, 1024>(v, sm, &tot);
    if (b < L) item_prefix[b] = carry + ex;
    carry += tot;
  }
  if (threadIdx.x == 0) {
    item_prefix[L] = carry;
    item_prefix[L + 1] = n_ctas;   // dynamic work counter: CTA i starts on item i
  }
}

struct SweepArgs {
  const float* px; const float* py; const float* pz;
  const int64_t* pt_offsets;
  int32_t n_blocks;
  const float* cameras;
  const int64_t* cam_offsets;
  const int32_t* cam_index;       // nullptr: camera j of the list is camera cam_first + j
  int32_t cam_first;
  int32_t cam_split;
  const int64_t* item_prefix;
  // MODE 0/1
  int32_t* n_hit;
  uint8_t* seen;                  // MODE 0, optional
  uint64_t* pivot;                // MODE 1
  uint64_t* extreme;              // MODE 1
  const float* krect;             // MODE 1, optional: per pair (cu, cv, hu, hv) interior rectangle
  // MODE 2
  const int32_t* pair_map;        // optional: pair slot of the (sub-)CSR -> pair id for poly / cand_*
  const float* poly;
  const int64_t* cand_offsets;
  float2* cand_uv;
  int32_t* cand_count;
};

struct Item {
  int b;
  int64_t tile;
  int64_t start;     // first point (global position)
  int cnt;           // points in the tile
  int64_t cam_lo, cam_hi;  // positions inside the block's camera list
};

__device__ __forceinline__ Item decode_item(const SweepArgs& a, const int64_t* s_prefix,
                                            int64_t item) {
  int lo = 0, hi = a.n_blocks - 1;
  while (lo < hi) {
    const int mid = (lo + hi + 1) >> 1;
    if (s_prefix[mid] <= item) lo = mid; else hi = mid - 1;
  }
  // skip empty blocks that share the same prefix value: lo is the LAST block
  // with prefix <= item, which is the one that owns it.
  Item it;
  it.b = lo;
  const int64_t local = item - s_prefix[lo];
  const int64_t p0 = a.pt_offsets[lo], p1 = a.pt_offsets[lo + 1];
  const int64_t c = a.cam_offsets[lo + 1] - a.cam_offsets[lo];
  const int64_t sp = c < a.cam_split ? c : a.cam_split;
  it.tile = local / sp;
  const int64_t s = local - it.tile * sp;
  it.start = p0 + it.tile * kSwTile;
  const int64_t rem = p1 - it.start;
  it.cnt = (int)(rem < kSwTile ? rem : kSwTile);
  it.cam_lo = c * s / sp;
  it.cam_hi = c * (s + 1) / sp;
  return it;
}

// diagnostics (ptgs_debug_sweep_stats): point-camera tests before / after the
// tile cull etc., summed over every sweep launch since the last reset.  The
// counters are device-global state on the hot path, so they exist only in
// diagnostics builds (-DPTGS_SWEEP_STATS_FINE); the release library has none.
#ifdef PTGS_SWEEP_STATS_FINE
__device__ unsigned long long g_sweep_stats[8];
#define PTGS_FINE_STAT(i, v) atomicAdd(&g_sweep_stats[i], (unsigned long long)(v))
#else
#define PTGS_FINE_STAT(i, v) ((void)0)
#endif

// ---- the sweep ---------------------------------------------------------------
// Members a mode does not use shrink to one element: the coverage sweep (MODE 0) has no polygons
// (20 KB), which is what lets 3 CTAs fit one SM next to the 8 KB prefix table of its 1024 super-tiles.
template <int MODE>
struct SweepSmemT {
  alignas(128) float tile[2][3][kSwStageFloats];
  alignas(16) float cam[kSwCamChunk][kCamFloats];
  // MODE 1: kSwCamChunk krect records (kKrectFloats each); MODE 2: kSwCamChunk / 2 inner polygons
  // (rows of 2 * kKrectFloats >= kPolyFloats floats) -- the same bytes
  alignas(16) float poly[MODE == 0 ? 4 : kSwCamChunk * kKrectFloats];
  long long cand_off[MODE == 2 ? kSwCamChunk : 1];        // MODE 2
  int cand_cap[MODE == 2 ? kSwCamChunk : 1];              // MODE 2
  int pair_of[MODE == 2 ? kSwCamChunk : 1];               // MODE 2
  int act[kSwThreads];                                    // cameras of the round that survive the tile cull
  long long pair[kSwCamChunk];                            // output slot (pair id / camera) of the staged cameras
  int nact;
  int tag_b;                                              // block / first camera whose whole camera range is staged
  long long tag_lo;                                       //   (-1: none) -- see `cached` in the sweep
  long long next_item;
  int qn, qhead;                                          // (warp run, camera) work queue of the chunk
  unsigned q[(kSwThreads / 32) * kSwCamChunk];            // run << 24 | camera slot << 12 | edge mask (0xfff: none)
  int tile_seen;                                          // MODE 0: a camera contains the whole tile
  alignas(16) float der[kSwCamChunk][8];                  // MODE 0/1: fast-in / not-miss rectangles
  float wbox[kSwThreads / 32][6];                         // per warp tile AABB
  // exact path of MODE 0 / 1 (COMPACT): the pass's band points as a dense list of slots
  // (k * 32 + lane) and the pair's current extreme thresholds, per warp
  unsigned char slist[kSwThreads / 32][32 * kSwP];
  float thr[kSwThreads / 32][kDirs];
  unsigned long long pvthr[kSwThreads / 32];
  float box[6];                                           // centre xyz, half extent xyz (padded)
  alignas(8) uint64_t bar[2];
  int64_t prefix[kSwMaxBlocks + 1];
};

// Exact path (MODE 0 / 1): the band points of a pass are compacted into dense 32-lane groups
// (re-read from the staged tile); pivot / directional extremes are updated only by hits that can
// still improve the pair's current global values (thresholds read once per pass).
template <int MODE>
__global__ void __launch_bounds__(kSwThreads, kSwCtasPerSM) sweep_kernel(SweepArgs a) {
  constexpr bool kWidePoly = (MODE == 2) && (kPolyFloats > kKrectFloats);      // 16-edge inner polygons
  constexpr int kChunk = kWidePoly ? kSwCamChunk / 2 : kSwCamChunk;            // cameras staged per chunk
  constexpr int kPolyStride = kWidePoly ? 2 * kKrectFloats : kKrectFloats;     // floats per staged polygon row
  static_assert(kPolyFloats <= 2 * kKrectFloats, "inner polygon row");
  extern __shared__ __align__(128) unsigned char smem_raw[];
  using SweepSmem = SweepSmemT<MODE>;
  SweepSmem& S = *reinterpret_cast<SweepSmem*>(smem_raw);
  const int tid = threadIdx.x, lane = tid & 31;

  for (int i = tid; i <= a.n_blocks; i += kSwThreads) S.prefix[i] = a.item_prefix[i];
  if (tid == 0) {
    S.tag_b = -1;
    S.tag_lo = -1;
    mbar_init(&S.bar[0], 1);
    mbar_init(&S.bar[1], 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  const int64_t n_items = S.prefix[a.n_blocks];

  auto issue = [&](const Item& it, int stage) {
    // 16-byte granular window around [start, start + cnt)
    const int64_t a0 = it.start & ~(int64_t)3;
    const int shift = (int)(it.start - a0);
    const uint32_t bytes = (uint32_t)(((shift + it.cnt + 3) & ~3) * 4);
    mbar_expect_tx(&S.bar[stage], 3 * bytes);
    bulk_g2s(S.tile[stage][0], a.px + a0, bytes, &S.bar[stage]);
    bulk_g2s(S.tile[stage][1], a.py + a0, bytes, &S.bar[stage]);
    bulk_g2s(S.tile[stage][2], a.pz + a0, bytes, &S.bar[stage]);
  };

  int64_t item = blockIdx.x;
  Item cur;
  if (item < n_items) {
    cur = decode_item(a, S.prefix, item);
    if (tid == 0) issue(cur, 0);
  }
  uint32_t phase[2] = {0, 0};
  int stage = 0;
  unsigned long long* work_counter =
      reinterpret_cast<unsigned long long*>(const_cast<int64_t*>(a.item_prefix)) + a.n_blocks + 1;
  for (; item < n_items; stage ^= 1) {
    // ---- tile -> registers
    mbar_wait(&S.bar[stage], phase[stage]);
    phase[stage] ^= 1;
    float x[kSwP], y[kSwP], z[kSwP];
    float wb[6];        // this warp's box: centre xyz, half extent xyz (its 256 points are contiguous)
    const int wcnt = min(max(cur.cnt - (tid & ~31) * kSwP, 0), 32 * kSwP);
    int have_run = tid >> 5;   // which 256-point run of the tile the registers hold
    {
      const int shift = (int)(cur.start & 3);
      const float qnan = __int_as_float(0x7fc00000);
#pragma unroll
      for (int k = 0; k < kSwP; ++k) {
        const int i = (tid & ~31) * kSwP + k * 32 + lane;   // warp w owns points [256 w, 256 w + 256)
        const bool ok = i < cur.cnt;
        x[k] = ok ? S.tile[stage][0][shift + i] : qnan;   // NaN never hits
        y[k] = ok ? S.tile[stage][1][shift + i] : qnan;
        z[k] = ok ? S.tile[stage][2][shift + i] : qnan;
      }
    }
    {
      // tile AABB (fminf/fmaxf ignore the NaN padding)
      const float inf = __int_as_float(0x7f800000);
      float lo[3] = {inf, inf, inf}, hi[3] = {-inf, -inf, -inf};
#pragma unroll
      for (int k = 0; k < kSwP; ++k) {
        lo[0] = fminf(lo[0], x[k]); hi[0] = fmaxf(hi[0], x[k]);
        lo[1] = fminf(lo[1], y[k]); hi[1] = fmaxf(hi[1], y[k]);
        lo[2] = fminf(lo[2], z[k]); hi[2] = fmaxf(hi[2], z[k]);
      }
#pragma unroll
      for (int d = 16; d > 0; d >>= 1) {
#pragma unroll
        for (int a = 0; a < 3; ++a) {
          lo[a] = fminf(lo[a], __shfl_xor_sync(kFullMask, lo[a], d));
          hi[a] = fmaxf(hi[a], __shfl_xor_sync(kFullMask, hi[a], d));
        }
      }
      if (lane == 0) {
#pragma unroll
        for (int a = 0; a < 3; ++a) { S.wbox[tid >> 5][a] = lo[a]; S.wbox[tid >> 5][3 + a] = hi[a]; }
      }
#pragma unroll
      for (int a = 0; a < 3; ++a) {
        const float pad = (fabsf(lo[a]) + fabsf(hi[a])) * (1.0f / 1048576.0f);
        wb[a] = 0.5f * lo[a] + 0.5f * hi[a];
        wb[3 + a] = (0.5f * hi[a] - 0.5f * lo[a]) + pad;
      }
    }
    __syncthreads();   // every thread is done with stage^1 (previous item) and this stage
    if (tid < 3) {
      float lo = S.wbox[0][tid], hi = S.wbox[0][3 + tid];
#pragma unroll
      for (int w = 1; w < kSwThreads / 32; ++w) {
        lo = fminf(lo, S.wbox[w][tid]);
        hi = fmaxf(hi, S.wbox[w][3 + tid]);
      }
      // centre / half extent, padded for the rounding of these two operations
      const float pad = (fabsf(lo) + fabsf(hi)) * (1.0f / 1048576.0f);
      S.box[tid] = 0.5f * lo + 0.5f * hi;
      S.box[3 + tid] = (0.5f * hi - 0.5f * lo) + pad;
    }
    // dynamic schedule: the next item comes from a global counter (items differ
    // widely in cost); its tile is requested right away
    if (tid == 0) {
      const int64_t nx = (int64_t)atomicAdd(work_counter, 1ull);
      S.next_item = nx;
      if (nx < n_items) {
        const Item nxt0 = decode_item(a, S.prefix, nx);
        fence_proxy_async();
        issue(nxt0, stage ^ 1);
      }
    }

    const int64_t cam_base = a.cam_offsets[cur.b];
    unsigned seen_mask = 0;
    if (MODE == 0 && tid == 0) S.tile_seen = 0;   // ordered by the chunk loop's barriers
    // The item's cameras are culled against the tile box by all threads at once
    // (rounds of kSwThreads cameras), the survivors are compacted into S.act and
    // swept in chunks of kChunk; tiles that lie wholly inside a camera's "in"
    // rectangle are counted right here.
    // Stages cameras [first, first + count) of the item's list (positions S.act[c0 + j] - or
    // cur.cam_lo + j when `direct`) with everything the sweep needs per camera.
    auto stage_cameras = [&](int c0, int nc, bool direct) {
      for (int i = tid; i < nc * (kCamFloats / 4); i += kSwThreads) {
        const int j = i / (kCamFloats / 4), q = i % (kCamFloats / 4);
        const int64_t lpos = cur.cam_lo + (direct ? j : S.act[c0 + j]);
        const int64_t cam = a.cam_index ? (int64_t)a.cam_index[cam_base + lpos] : (int64_t)a.cam_first + lpos;
        reinterpret_cast<float4*>(S.cam[j])[q] =
            __ldg(reinterpret_cast<const float4*>(a.cameras + cam * kCamFloats) + q);
      }
      if (MODE == 2) {
        for (int i = tid; i < nc * (kPolyFloats / 4); i += kSwThreads) {
          const int j = i / (kPolyFloats / 4), q = i % (kPolyFloats / 4);
          const int64_t slot = cam_base + cur.cam_lo + (direct ? j : S.act[c0 + j]);
          const int64_t pair = a.pair_map ? (int64_t)a.pair_map[slot] : slot;
          reinterpret_cast<float4*>(S.poly + j * kPolyStride)[q] =
              __ldg(reinterpret_cast<const float4*>(a.poly + pair * kPolyFloats) + q);
        }
      }
      if (MODE == 1) {
        for (int i = tid; i < nc * (kKrectFloats / 4); i += kSwThreads) {
          const int j = i / (kKrectFloats / 4), q = i % (kKrectFloats / 4);
          const int64_t slot = cam_base + cur.cam_lo + (direct ? j : S.act[c0 + j]);
          reinterpret_cast<float4*>(S.poly + j * kPolyStride)[q] =
              a.krect ? __ldg(reinterpret_cast<const float4*>(a.krect + slot * kKrectFloats) + q)
                      : make_float4(0.0f, 0.0f, -1.0f, -1.0f);
        }
      }
      if (tid < nc) {
        const int64_t lpos = cur.cam_lo + (direct ? tid : S.act[c0 + tid]);
        S.pair[tid] = (MODE == 0 && !a.cam_index) ? lpos : (cam_base + lpos);
        if (MODE != 2) {
          const int64_t cam = a.cam_index ? (int64_t)a.cam_index[cam_base + lpos] : (int64_t)a.cam_first + lpos;
          const float* rp = a.cameras + cam * kCamFloats;
          Cam cj;                                  // make_rects only reads w, h
          cj.w = __ldg(rp + 16); cj.h = __ldg(rp + 17);
          float der[8];
          const float* kr = nullptr;
          if (MODE == 1 && a.krect) kr = a.krect + (cam_base + lpos) * kKrectFloats;
          make_rects(cj, kr, MODE, der);
#pragma unroll
          for (int q = 0; q < 8; ++q) S.der[tid][q] = der[q];
        }
        if (MODE == 2) {
          const int64_t slot = cam_base + lpos;
          const int64_t pair = a.pair_map ? (int64_t)a.pair_map[slot] : slot;
          const int64_t o0 = a.cand_offsets[pair], o1 = a.cand_offsets[pair + 1];
          S.cand_off[tid] = o0;
          S.cand_cap[tid] = (int)(o1 - o0);
          S.pair_of[tid] = (int)pair;
        }
      }
    };
    // An item whose whole camera range fits the staging arrays keeps it there for as long as
    // the CTA stays on the same (block, range): consecutive items are mostly tiles of the same
    // block, so the camera / interior-region records are read from global memory once per block
    // and not three dependent L2 round trips per tile.
    const bool cached = (cur.cam_hi - cur.cam_lo) <= kChunk;
    if (cached) {
      const bool fresh = (S.tag_b == cur.b) && (S.tag_lo == cur.cam_lo);   // ordered by the barrier above
      if (!fresh) stage_cameras(0, (int)(cur.cam_hi - cur.cam_lo), true);
      __syncthreads();
      if (tid == 0) { S.tag_b = cur.b; S.tag_lo = cur.cam_lo; }
    } else if (tid == 0) {
      S.tag_b = -1;                      // the staging arrays are about to be overwritten chunk by chunk
    }
    for (int64_t r0 = cur.cam_lo; r0 < cur.cam_hi; r0 += kSwThreads) {
      const int nr = (int)((cur.cam_hi - r0) < kSwThreads ? (cur.cam_hi - r0) : kSwThreads);
      if (tid == 0) S.nact = 0;
      __syncthreads();                       // S.box / S.nact visible, previous round finished
      if (tid < nr) {
        const int64_t lpos = r0 + tid;       // position inside the block's camera list
        Cam cj;
        if (cached) {
          cj = load_cam(S.cam[tid]);
        } else {
          const int64_t cam = a.cam_index ? (int64_t)a.cam_index[cam_base + lpos] : (int64_t)a.cam_first + lpos;
          float rec[kCamFloats];
#pragma unroll
          for (int q = 0; q < kCamFloats / 4; ++q)
            reinterpret_cast<float4*>(rec)[q] =
                __ldg(reinterpret_cast<const float4*>(a.cameras + cam * kCamFloats) + q);
          cj = load_cam(rec);
        }
        bool live = tile_may_hit(cj, S.box);
        if (MODE != 2 && live) {
          bool all_in;
          if (cached) {
            all_in = (MODE == 1 && a.krect) ? tile_all_in_poly(cj, S.poly + tid * kPolyStride, S.box)
                                            : tile_all_in(cj, S.der[tid], S.box);
          } else {
            float der[8];
            const float* kr = nullptr;
            if (MODE == 1 && a.krect) kr = a.krect + (cam_base + lpos) * kKrectFloats;
            make_rects(cj, kr, MODE, der);
            all_in = (MODE == 1 && kr) ? tile_all_in_poly(cj, kr, S.box) : tile_all_in(cj, der, S.box);
          }
          if (all_in) {
            // every point of the tile is a hit that needs no further work
            const int64_t pair = (MODE == 0 && !a.cam_index) ? lpos : (cam_base + lpos);
            atomicAdd(a.n_hit + pair, cur.cnt);
            PTGS_FINE_STAT(5, cur.cnt);
            if (MODE == 0) S.tile_seen = 1;
            live = false;
          }
        }
        if (live) S.act[atomicAdd(&S.nact, 1)] = (int)(lpos - cur.cam_lo);
      }
      __syncthreads();
      const int nact_all = S.nact;
      if (tid == 0) {
        PTGS_FINE_STAT(0, (unsigned long long)nr * (unsigned)cur.cnt);
        PTGS_FINE_STAT(1, (unsigned long long)nact_all * (unsigned)cur.cnt);
      }
    for (int c0 = 0; c0 < nact_all; c0 += kChunk) {
      const int nc = (nact_all - c0) < kChunk ? (nact_all - c0) : kChunk;
      // ---- stage the chunk of surviving cameras (already resident when `cached`)
      if (!cached) stage_cameras(c0, nc, false);
      if (tid == 0) { S.qn = 0; S.qhead = 0; }
      __syncthreads();
      // staged slot of the i-th survivor of the chunk
      auto slot_of = [&](int i) { return cached ? S.act[c0 + i] : i; };
      for (int g0 = 0; g0 < nc; g0 += 32) {
      // ---- per-warp classification: lane j tests camera j of the chunk against the
      // box of the warp's own 256 contiguous points (a few voxel-cells wide):
      // outside -> skipped, wholly inside the "in" rectangle -> counted, else swept.
      unsigned pm = 0;
      unsigned und = 0xffffffffu;      // MODE 1, lane j: polygon edges of camera j that cut the warp's box
      if (wcnt > 0) {
        bool may = false, allin = false;
        int sl = 0;                      // staged slot of this lane's camera
        if (g0 + lane < nc) {
          sl = slot_of(g0 + lane);
          const Cam cj = load_cam(S.cam[sl]);
          may = tile_may_hit(cj, wb);
          if (MODE == 1 && may) {
            bool fz;
            und = poly_undecided(cj, S.poly + sl * kPolyStride, wb, fz);
            allin = und == 0u && fz;
          }
          if (MODE == 0 && may) allin = tile_all_in(cj, S.der[sl], wb);
          if (MODE == 2 && may) allin = inner_all_in(cj, S.poly + sl * kPolyStride, wb);   // nothing to emit
          if (MODE != 2 && allin) atomicAdd(a.n_hit + S.pair[sl], wcnt);
        }
        pm = __ballot_sync(kFullMask, may && !allin);
        if (MODE == 0 && __ballot_sync(kFullMask, allin) != 0u) seen_mask = 0xffu;
        // the cameras that cut this run go to the CTA's queue: any warp may sweep them
        if (pm) {
          int base = 0;
          if (lane == 0) base = atomicAdd(&S.qn, __popc(pm));
          base = __shfl_sync(kFullMask, base, 0);
          if (may && !allin) {
            const unsigned um = (MODE == 1 && und != 0xffffffffu) ? (und & 0xffu) : 0xfffu;
            S.q[base + __popc(pm & lanemask_lt())] =
                ((unsigned)(tid >> 5) << 24) | ((unsigned)sl << 12) | um;
          }
        }
        if (lane == 0) {
          PTGS_FINE_STAT(MODE == 2 ? 3 : 2, (unsigned long long)__popc(pm) * (unsigned)wcnt);
        }
        {
          const unsigned am = __ballot_sync(kFullMask, allin);
          if (lane == 0 && am) PTGS_FINE_STAT(4, (unsigned long long)__popc(am) * (unsigned)wcnt);
        }
      }

      }   // lane groups (phase 1: classification)
      __syncthreads();

      // ---- phase 2: the queued (run, camera) pairs are swept by whichever warp is
      // free (runs of a tile differ widely in how many cameras cut them); a warp
      // that takes another run's pair reloads that run's points from the tile
      const int qn = S.qn;
      for (;;) {
        int qi = 0;
        if (lane == 0) qi = atomicAdd(&S.qhead, 1);
        qi = __shfl_sync(kFullMask, qi, 0);
        if (qi >= qn) break;
        const unsigned ent = S.q[qi];
        const int run = (int)(ent >> 24), j = (int)((ent >> 12) & 0xfffu);
        const unsigned uj = (ent & 0xfffu) == 0xfffu ? 0xffffffffu : (ent & 0xfffu);
        if (run != have_run) {
          const int shift = (int)(cur.start & 3);
          const float qnan = __int_as_float(0x7fc00000);
#pragma unroll
          for (int k = 0; k < kSwP; ++k) {
            const int i = run * (32 * kSwP) + k * 32 + lane;
            const bool ok = i < cur.cnt;
            x[k] = ok ? S.tile[stage][0][shift + i] : qnan;
            y[k] = ok ? S.tile[stage][1][shift + i] : qnan;
            z[k] = ok ? S.tile[stage][2][shift + i] : qnan;
          }
          have_run = run;
        }
      if (MODE != 2) {
        {
          const Cam c = load_cam(S.cam[j]);
          const int64_t pair = S.pair[j];
          // MODE 1: the pair's extremes / pivot as of now, requested before the classification
          // loop so that their L2 latency is hidden (only used when the pass has band points)
          unsigned long long gcur = 0;
          unsigned long long* gpv = reinterpret_cast<unsigned long long*>(a.pivot) + pair;
          unsigned long long* gex = reinterpret_cast<unsigned long long*>(a.extreme) + pair * kDirs;
          if (MODE == 1) {
            if (lane < kDirs) gcur = *reinterpret_cast<volatile unsigned long long*>(gex + lane);
            else if (lane == kDirs) gcur = *reinterpret_cast<volatile unsigned long long*>(gpv);
          }
          const float4 din = *reinterpret_cast<const float4*>(&S.der[j][0]);
          const float4 dout = *reinterpret_cast<const float4*>(&S.der[j][4]);
          const bool front_y = c.front_axis == 1.0f;
          int cnt = 0;
          unsigned slow = 0;
          // MODE 1: when at most two polygon edges cut the warp's box, "inside those
          // edges" replaces the interior rectangle as the count-only region
          bool use_edges = false;
          float e1a = 0.0f, e1b = 0.0f, e1c = 0.0f, e2a = 0.0f, e2b = 0.0f, e2c = 0.0f;
          if (MODE == 1) {
            if (uj != 0xffffffffu && uj != 0u && __popc(uj) <= 2) {
              use_edges = true;
              const int ea = __ffs(uj) - 1, eb = 31 - __clz(uj);
              const float* pr = S.poly + j * kPolyStride;
              e1a = pr[8 + 3 * ea]; e1b = pr[9 + 3 * ea]; e1c = pr[10 + 3 * ea];
              e2a = pr[8 + 3 * eb]; e2b = pr[9 + 3 * eb]; e2c = pr[10 + 3 * eb];
            }
          }
#pragma unroll
          for (int k = 0; k < kSwP; ++k) {
            const float X = cam_row(c.r00, c.r01, c.r02, c.tx, x[k], y[k], z[k]);
            const float Y = cam_row(c.r10, c.r11, c.r12, c.ty, x[k], y[k], z[k]);
            const float Z = cam_row(c.r20, c.r21, c.r22, c.tz, x[k], y[k], z[k]);
            const float uh = __fmaf_rn(c.cx, Z, __fmul_rn(c.fx, X));
            const float vh = __fmaf_rn(c.cy, Z, __fmul_rn(c.fy, Y));
            const float az = fabsf(Z);
            const bool front = (front_y ? Y : Z) > 0.0f;
            bool in;
            if (MODE == 1 && use_edges) {
              in = front && (Z > 0.0f) &&
                   (__fmaf_rn(e1a, uh, __fmaf_rn(e1b, vh, __fmul_rn(e1c, Z))) > 0.0f) &&
                   (__fmaf_rn(e2a, uh, __fmaf_rn(e2b, vh, __fmul_rn(e2c, Z))) > 0.0f);
            } else {
              in = front && (fabsf(__fmaf_rn(-din.x, Z, uh)) < __fmul_rn(din.z, az)) &&
                   (fabsf(__fmaf_rn(-din.y, Z, vh)) < __fmul_rn(din.w, az));
            }
            bool nm;
            if (MODE == 0) {   // same centre: one fma per axis serves both rectangles
              nm = front && (fabsf(__fmaf_rn(-din.x, Z, uh)) <= __fmaf_rn(dout.z, az, 1e-30f)) &&
                   (fabsf(__fmaf_rn(-din.y, Z, vh)) <= __fmaf_rn(dout.w, az, 1e-30f));
            } else {
              nm = front && (fabsf(__fmaf_rn(-dout.x, Z, uh)) <= __fmaf_rn(dout.z, az, 1e-30f)) &&
                   (fabsf(__fmaf_rn(-dout.y, Z, vh)) <= __fmaf_rn(dout.w, az, 1e-30f));
            }
            cnt += __popc(__ballot_sync(kFullMask, in));      // warp-uniform count, no reduction later
            if (MODE == 0 && in && a.seen) a.seen[cur.start + run * (32 * kSwP) + k * 32 + lane] = 1;
            slow |= (nm && !in) ? (1u << k) : 0u;
          }
          // ---- exact path for the points between the two regions.  Each lane keeps
          // its best pivot / directional extremes of the pass in registers; one warp
          // reduction per pass then sends at most 9 atomics (lane d owns direction d)
          // to the pair's global slots, and only for values that improve them.
          if (__any_sync(kFullMask, slow != 0)) {
            if (lane == 0) PTGS_FINE_STAT(6, 1);
            const int wp = tid >> 5;
            if (MODE == 1) {
              // a stale value is only less extreme, so the filter below lets through a superset of
              // the hits that can improve the pair's extremes
              if (lane < kDirs)
                S.thr[wp][lane] = gcur ? key_to_float((uint32_t)(gcur >> 32)) : -__int_as_float(0x7f800000);
              else if (lane == kDirs)
                S.pvthr[wp] = gcur;
            }
            int total = 0;
#pragma unroll
            for (int k = 0; k < kSwP; ++k) {
              const bool mine = (slow >> k) & 1u;
              const unsigned mk = __ballot_sync(kFullMask, mine);
              if (mine) S.slist[wp][total + __popc(mk & lanemask_lt())] = (unsigned char)(k * 32 + lane);
              total += __popc(mk);
            }
            __syncwarp();
            const int shift0 = (int)(cur.start & 3);
            const uint32_t base_idx = (uint32_t)(cur.tile * kSwTile + run * (32 * kSwP));
            for (int base = 0; base < total; base += 32) {
              if (lane == 0) PTGS_FINE_STAT(7, 1);
              bool hit = false;
              float u = 0.0f, v = 0.0f;
              int sl = 0;
              if (base + lane < total) {
                sl = S.slist[wp][base + lane];
                const int i = run * (32 * kSwP) + sl;
                const float qx = S.tile[stage][0][shift0 + i], qy = S.tile[stage][1][shift0 + i],
                            qz = S.tile[stage][2][shift0 + i];
                const float X = cam_row(c.r00, c.r01, c.r02, c.tx, qx, qy, qz);
                const float Y = cam_row(c.r10, c.r11, c.r12, c.ty, qx, qy, qz);
                Proj p;
                p.Z = cam_row(c.r20, c.r21, c.r22, c.tz, qx, qy, qz);
                p.uh = __fmaf_rn(c.cx, p.Z, __fmul_rn(c.fx, X));
                p.vh = __fmaf_rn(c.cy, p.Z, __fmul_rn(c.fy, Y));
                if (((front_y ? Y : p.Z) > 0.0f) && project_exact(c, p, u, v)) {
                  hit = true;
                  if (MODE == 0 && a.seen) a.seen[cur.start + i] = 1;
                }
              }
              cnt += __popc(__ballot_sync(kFullMask, hit));
              if (MODE == 1) {
                // Pivot and directional extremes.  Per lane: can this hit reach any of the pair's
                // current values (S.thr / S.pvthr, refreshed as the pass goes)?  Only then -- rare
                // once a pair has seen a few tiles -- the group reduces convergently and sends ONE
                // atomic per improved direction.
                u = __fadd_rn(u, 0.0f);   // -0.0 -> +0.0 (numerically equal for the reference)
                v = __fadd_rn(v, 0.0f);
                const unsigned long long pk =
                    ((unsigned long long)__float_as_uint(v) << 32) | __float_as_uint(u);
                const float u2 = __fadd_rn(u, u), v2 = __fadd_rn(v, v);
                bool imp = pk < S.pvthr[wp];
#pragma unroll
                for (int d = 0; d < kDirs; ++d)
                  imp = imp || (dir_value(d * kDirStep, u, v, u2, v2) >= S.thr[wp][d]);
                if (__any_sync(kFullMask, hit && imp)) {
                  if (__any_sync(kFullMask, hit && pk < S.pvthr[wp])) {
                    const uint32_t ph = hit ? (uint32_t)(pk >> 32) : 0xffffffffu, pl = hit ? (uint32_t)pk : 0xffffffffu;
                    const uint32_t mh = __reduce_min_sync(kFullMask, ph);
                    const uint32_t ml = __reduce_min_sync(kFullMask, ph == mh ? pl : 0xffffffffu);
                    if (lane == 0) {
                      const unsigned long long best = ((unsigned long long)mh << 32) | ml;
                      if (best < *reinterpret_cast<volatile unsigned long long*>(gpv)) atomicMin(gpv, best);
                      if (best < S.pvthr[wp]) S.pvthr[wp] = best;
                    }
                  }
                  const uint32_t idx = base_idx + (uint32_t)sl;
#pragma unroll
                  for (int d = 0; d < kDirs; ++d) {
                    const float dval = dir_value(d * kDirStep, u, v, u2, v2);
                    // >=: among equal values the larger point index wins (max of (key, index))
                    const bool reach = hit && dval >= S.thr[wp][d];
                    if (__any_sync(kFullMask, reach)) {
                      const uint32_t key = reach ? order_key(dval) : 0u;
                      const uint32_t kmax = __reduce_max_sync(kFullMask, key);
                      const uint32_t imax = __reduce_max_sync(kFullMask, (reach && key == kmax) ? idx : 0u);
                      if (lane == 0) {
                        const unsigned long long val = ((unsigned long long)kmax << 32) | imax;
                        if (val > *reinterpret_cast<volatile unsigned long long*>(gex + d)) atomicMax(gex + d, val);
                        S.thr[wp][d] = key_to_float(kmax);
                      }
                    }
                  }
                  __syncwarp();
                }
              }
            }
          }
          if (lane == 0 && cnt) atomicAdd(a.n_hit + pair, cnt);
        }
      } else {
      {
        const Cam c = load_cam(S.cam[j]);
        const float w1 = __fadd_rn(c.w, 1.0f), h1 = __fadd_rn(c.h, 1.0f);
#pragma unroll
        for (int k = 0; k < kSwP; ++k) {
          const Proj p = project_pre(c, w1, h1, x[k], y[k], z[k]);
          if (p.maybe) {
            float u, v;
            if (project_exact(c, p, u, v)) {
              if (MODE == 2) {
                u = __fadd_rn(u, 0.0f);
                v = __fadd_rn(v, 0.0f);
                const float* pl = S.poly + j * kPolyStride;
                // the pivot's row: only the pivot itself takes part (see ptgs.h)
                const bool excluded = (v == pl[1]) && (u != pl[2]);
                if (!excluded) {
                  const int ne = (int)pl[0];
                  const float margin = pl[3];
                  bool inside = ne > 0;
                  for (int e = 0; e < ne; ++e) {
                    const float f = __fadd_rn(
                        __fadd_rn(__fmul_rn(pl[4 + 3 * e], u), __fmul_rn(pl[5 + 3 * e], v)),
                        pl[6 + 3 * e]);
                    inside = inside && (f > margin);
                  }
                  if (!inside) {
                    const int slot = atomicAdd(a.cand_count + S.pair_of[j], 1);
                    if (slot < S.cand_cap[j]) a.cand_uv[S.cand_off[j] + slot] = make_float2(u, v);
                  }
                }
              }
            }
          }
        }
      }
      }
      }   // queue

      // staging arrays are rewritten by the next chunk (rare: > kChunk surviving cameras)
      if (c0 + kChunk < nact_all) __syncthreads();
    }
    }
    if (MODE == 0 && a.seen) {
      const bool all = S.tile_seen != 0;   // written before the last barrier of the chunk loop
#pragma unroll
      for (int k = 0; k < kSwP; ++k)
        if ((tid & ~31) * kSwP + k * 32 + lane < cur.cnt && (all || (seen_mask & (1u << k))))
          a.seen[cur.start + (tid & ~31) * kSwP + k * 32 + lane] = 1;
    }
    item = S.next_item;            // published by the barriers of the camera loop
    if (item < n_items) cur = decode_item(a, S.prefix, item);
  }
}

// ---- hull candidates of the undecided pairs, one warp per 128-point run ---------------------
// sweep_kernel<2> stages every 1024-point tile of every block that owns an undecided pair and
// pays the item machinery (tile staging, CTA barriers, queue) for a handful of cameras per
// block: 1.17 ms at c4 for 13 % of the pairs, half of its issue slots idle at barriers.  The
// candidate pass needs none of it: a warp loads one run of 128 consecutive points of a block
// straight from global memory (4 per lane, coalesced), builds the run's box, lets lane j test
// undecided camera j of the block against the box (outside the image -> skipped, wholly inside
// the pair's inner polygon -> nothing to emit) and evaluates its points for the survivors with
// the arithmetic of the sweep.  The emitted SET per pair is defined per point (hit, not on the
// excluded pivot row, not strictly inside the inner polygon), so it does not depend on how the
// points are grouped; the box tests are the conservative ones of the sweep.
constexpr int kCrRun = 128;
constexpr int kCrWarps = 8;

__global__ void __launch_bounds__(1024) cand_schedule_kernel(
    const int64_t* __restrict__ pt_offsets, const int64_t* __restrict__ cam_offsets, int L,
    int64_t* __restrict__ run_prefix) {
  __shared__ long long sm[33];
  long long carry = 0;
  for (int s = 0; s < L; s += 1024) {
    const int b = s + threadIdx.x;
    long long v = 0;
    if (b < L && cam_offsets[b + 1] > cam_offsets[b])
      v = (pt_offsets[b + 1] - pt_offsets[b] + kCrRun - 1) / kCrRun;
    long long tot;
    const long long ex = block_exclusive_sum<long long, 1024>(v, sm, &tot);
    if (b < L) run_prefix[b] = carry + ex;
    carry += tot;
  }
  if (threadIdx.x == 0) run_prefix[L] = carry;
}

__global__ void __launch_bounds__(kCrWarps * 32, 4) cand_runs_kernel(
    const float* __restrict__ px, const float* __restrict__ py, const float* __restrict__ pz,
    const int64_t* __restrict__ pt_offsets, int L, const float* __restrict__ cameras,
    const int64_t* __restrict__ cam_offsets, const int32_t* __restrict__ cam_index,
    const int32_t* __restrict__ pair_map, const float* __restrict__ poly,
    const int64_t* __restrict__ cand_offsets, float2* __restrict__ cand_uv,
    int32_t* __restrict__ cand_count, const int64_t* __restrict__ run_prefix) {
  __shared__ int64_t s_prefix[kSwMaxBlocks + 1];
  __shared__ __align__(16) float s_cam[kCrWarps][kCamFloats];
  __shared__ __align__(16) float s_poly[kCrWarps][kPolyFloats];
  for (int i = threadIdx.x; i <= L; i += blockDim.x) s_prefix[i] = run_prefix[i];
  __syncthreads();
  const int64_t n_runs = s_prefix[L];
  const int lane = threadIdx.x & 31, wp = threadIdx.x >> 5;
  const float qnan = __int_as_float(0x7fc00000);
  const float inf = __int_as_float(0x7f800000);
  for (int64_t run = (int64_t)blockIdx.x * kCrWarps + wp; run < n_runs; run += (int64_t)gridDim.x * kCrWarps) {
    int lo = 0, hi = L - 1;                 // last block with prefix <= run: the one that owns it
    while (lo < hi) {
      const int mid = (lo + hi + 1) >> 1;
      if (s_prefix[mid] <= run) lo = mid; else hi = mid - 1;
    }
    const int b = lo;
    const int64_t p0 = pt_offsets[b] + (run - s_prefix[b]) * kCrRun;
    const int64_t rem = pt_offsets[b + 1] - p0;
    const int cnt = (int)(rem < kCrRun ? rem : kCrRun);
    float x[4], y[4], z[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const int i = k * 32 + lane;
      const bool ok = i < cnt;
      x[k] = ok ? __ldg(px + p0 + i) : qnan;    // NaN never hits
      y[k] = ok ? __ldg(py + p0 + i) : qnan;
      z[k] = ok ? __ldg(pz + p0 + i) : qnan;
    }
    float wb[6];                             // centre xyz, half extent xyz (padded like the sweep's boxes)
    {
      float bl[3] = {inf, inf, inf}, bh[3] = {-inf, -inf, -inf};
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        bl[0] = fminf(bl[0], x[k]); bh[0] = fmaxf(bh[0], x[k]);
        bl[1] = fminf(bl[1], y[k]); bh[1] = fmaxf(bh[1], y[k]);
        bl[2] = fminf(bl[2], z[k]); bh[2] = fmaxf(bh[2], z[k]);
      }
#pragma unroll
      for (int d = 16; d > 0; d >>= 1) {
#pragma unroll
        for (int a = 0; a < 3; ++a) {
          bl[a] = fminf(bl[a], __shfl_xor_sync(kFullMask, bl[a], d));
          bh[a] = fmaxf(bh[a], __shfl_xor_sync(kFullMask, bh[a], d));
        }
      }
#pragma unroll
      for (int a = 0; a < 3; ++a) {
        const float pad = (fabsf(bl[a]) + fabsf(bh[a])) * (1.0f / 1048576.0f);
        wb[a] = 0.5f * bl[a] + 0.5f * bh[a];
        wb[3 + a] = (0.5f * bh[a] - 0.5f * bl[a]) + pad;
      }
    }
    const int64_t cam_base = cam_offsets[b];
    const int nc = (int)(cam_offsets[b + 1] - cam_base);
    for (int g0 = 0; g0 < nc; g0 += 32) {
      bool live = false;
      if (g0 + lane < nc) {
        const int64_t slot = cam_base + g0 + lane;
        float rec[kCamFloats];
#pragma unroll
        for (int q = 0; q < kCamFloats / 4; ++q)
          reinterpret_cast<float4*>(rec)[q] =
              __ldg(reinterpret_cast<const float4*>(cameras + (int64_t)cam_index[slot] * kCamFloats) + q);
        const Cam cj = load_cam(rec);
        live = tile_may_hit(cj, wb);
      }
      unsigned pm = __ballot_sync(kFullMask, live);
      while (pm) {
        const int j = __ffs(pm) - 1;
        pm &= pm - 1;
        const int64_t slot = cam_base + g0 + j;
        const int64_t pair = pair_map ? (int64_t)pair_map[slot] : slot;
        __syncwarp();                        // the previous camera's staged records are no longer read
        if (lane < kCamFloats / 4)
          reinterpret_cast<float4*>(s_cam[wp])[lane] =
              __ldg(reinterpret_cast<const float4*>(cameras + (int64_t)cam_index[slot] * kCamFloats) + lane);
        else if (lane - 8 >= 0 && lane - 8 < kPolyFloats / 4)
          reinterpret_cast<float4*>(s_poly[wp])[lane - 8] =
              __ldg(reinterpret_cast<const float4*>(poly + pair * kPolyFloats) + (lane - 8));
        __syncwarp();
        const Cam c = load_cam(s_cam[wp]);
        const float* pl = s_poly[wp];
        // the whole run projects strictly inside the pair's inner polygon: nothing to emit
        // (lane e tests edge e; few cameras get here, so one lane per camera would idle the warp)
        if (inner_all_in_warp(c, pl, wb, lane)) continue;
        const float w1 = __fadd_rn(c.w, 1.0f), h1 = __fadd_rn(c.h, 1.0f);
        const int ne = (int)pl[0];
        const float margin = pl[3];
        const int64_t o0 = cand_offsets[pair];
        const int cap = (int)(cand_offsets[pair + 1] - o0);
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          bool emit = false;
          float u = 0.0f, v = 0.0f;
          const Proj p = project_pre(c, w1, h1, x[k], y[k], z[k]);
          if (p.maybe && project_exact(c, p, u, v)) {
            u = __fadd_rn(u, 0.0f);
            v = __fadd_rn(v, 0.0f);
            // the pivot's row: only the pivot itself takes part (see ptgs.h)
            const bool excluded = (v == pl[1]) && (u != pl[2]);
            if (!excluded) {
              bool inside = ne > 0;
              for (int e = 0; e < ne; ++e) {
                const float f = __fadd_rn(
                    __fadd_rn(__fmul_rn(pl[4 + 3 * e], u), __fmul_rn(pl[5 + 3 * e], v)), pl[6 + 3 * e]);
                inside = inside && (f > margin);
              }
              emit = !inside;
            }
          }
          const unsigned em = __ballot_sync(kFullMask, emit);
          if (em) {
            int base = 0;
            if (lane == 0) base = atomicAdd(cand_count + pair, __popc(em));
            base = __shfl_sync(kFullMask, base, 0);
            const int slot_c = base + __popc(em & lanemask_lt());
            if (emit && slot_c < cap) cand_uv[o0 + slot_c] = make_float2(u, v);
          }
        }
      }
    }
  }
}

// ---- inner / outer polygon, area bounds and early decision per pair --------------
// Sutherland-Hodgman clip of a convex polygon (<= 16 vertices) by a*u + b*v <= c.
__device__ int clip_halfplane(const double* px, const double* py, int n, double a, double b,
                              double c, double* qx, double* qy) {
  int m = 0;
  for (int i = 0; i < n; ++i) {
    const int k = (i + 1 == n) ? 0 : i + 1;
    const double fi = a * px[i] + b * py[i] - c, fk = a * px[k] + b * py[k] - c;
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
__device__ double shoelace(const double* px, const double* py, int n) {
  if (n < 3) return 0.0;
  double acc = 0.0;
  for (int i = 0; i < n; ++i) {
    const int k = (i + 1 == n) ? 0 : i + 1;
    acc += px[i] * py[k] - px[k] * py[i];
  }
  return 0.5 * fabs(acc);
}

// Per pair: the inner polygon (extreme hits that belong to the point set the
// reference scans) -> lower bound of the reference's hull area; the outer
// octagon (the 8 supporting lines of sweep 1, inflated for the float32 rounding
// of u+v / u-v) -> upper bound.  A pair whose bounds fall on one side of the
// threshold by more than near_eps is decided here (flags bit4: `ratio` holds the
// deciding bound, not the exact value); the rest get need = 1 and go through
// sweep 2 + the exact hull.  exact_all != 0 sends every pair with enough hits
// through the exact path.
__global__ void __launch_bounds__(128) bounds_kernel(
    const float* __restrict__ px, const float* __restrict__ py, const float* __restrict__ pz,
    const int64_t* __restrict__ pt_offsets, int L, const float* __restrict__ cameras,
    const int64_t* __restrict__ cam_offsets, const int32_t* __restrict__ cam_index,
    const int32_t* __restrict__ n_hit, const uint64_t* __restrict__ pivot,
    const uint64_t* __restrict__ extreme, int min_hits, double thr, double near_eps,
    int exact_all, float* __restrict__ poly, double* __restrict__ ratio_lo,
    double* __restrict__ ratio_hi, uint8_t* __restrict__ need, double* __restrict__ area_out,
    double* __restrict__ ratio_out, uint8_t* __restrict__ keep_out, uint8_t* __restrict__ flags_out) {
  const int64_t n_pairs = cam_offsets[L];
  const int64_t pair = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (pair >= n_pairs) return;
  float* out = poly + pair * kPolyFloats;
  for (int i = 0; i < kPolyFloats; ++i) out[i] = 0.0f;
  out[1] = -1.0f; out[2] = -1.0f;     // no hit has a negative coordinate
  const int nh = n_hit[pair];
  need[pair] = 0; area_out[pair] = 0.0; ratio_out[pair] = 0.0; keep_out[pair] = 0; flags_out[pair] = 0;
  ratio_lo[pair] = 0.0; ratio_hi[pair] = 0.0;
  if (nh <= 0) return;
  int lo = 0, hi = L - 1;             // block of the pair
  while (lo < hi) {
    const int mid = (lo + hi + 1) >> 1;
    if (cam_offsets[mid] <= pair) lo = mid; else hi = mid - 1;
  }
  const int64_t p0 = pt_offsets[lo];
  const Cam c = load_cam(cameras + (int64_t)cam_index[pair] * kCamFloats);
  const float w1 = __fadd_rn(c.w, 1.0f), h1 = __fadd_rn(c.h, 1.0f);
  const uint64_t pv = pivot[pair];
  const float vp = __uint_as_float((uint32_t)(pv >> 32)), up = __uint_as_float((uint32_t)pv);
  out[1] = vp; out[2] = up;
  const float Dm = fmaxf(c.w, c.h);
  out[3] = 4e-6f * Dm * Dm;           // > the float32 evaluation error of the half planes
  double vx[kDirs], vy[kDirs], sup[kDirs];
  int m = 0;
  for (int d = 0; d < kDirs; ++d) {
    const uint64_t e = extreme[pair * kDirs + d];
    sup[d] = 0.0;
    if (!e) continue;
    const double hval = (double)key_to_float((uint32_t)(e >> 32));
    sup[d] = hval + fabs(hval) * (1.0 / 4194304.0) + 1e-30;   // outward, covers the one float32 rounding of a*u + b*v
    const int64_t gi = p0 + (int64_t)(e & 0xffffffffull);
    const Proj p = project_pre(c, w1, h1, px[gi], py[gi], pz[gi]);
    float u, v;
    if (!project_exact(c, p, u, v)) continue;   // cannot happen: it was a hit in sweep 1
    u = __fadd_rn(u, 0.0f); v = __fadd_rn(v, 0.0f);
    if (v == vp && u != up) continue;           // pivot-row hit: not in the scanned set
    if (m && vx[m - 1] == (double)u && vy[m - 1] == (double)v) continue;
    vx[m] = u; vy[m] = v; ++m;
  }
  while (m > 1 && vx[m - 1] == vx[0] && vy[m - 1] == vy[0]) --m;
  double a_in = 0.0;
  if (m >= 3) {
    for (int i = 0; i < m; ++i) {
      const int k = (i + 1 == m) ? 0 : i + 1;
      const double ea = -(vy[k] - vy[i]), eb = vx[k] - vx[i];
      out[4 + 3 * i] = (float)ea;
      out[5 + 3 * i] = (float)eb;
      out[6 + 3 * i] = (float)(-(ea * vx[i] + eb * vy[i]));
    }
    out[0] = (float)m;
    a_in = shoelace(vx, vy, m);
  }
  // outer polygon: the kDirs supporting half planes a_d*u + b_d*v <= sup_d.  Start from the rectangle of
  // the four axis directions, clip by the others (each clip adds at most one vertex: <= kDirs <= 16).
  double ax[16], ay[16], bx[16], by[16];
  constexpr int kQ = kDirs / 4;                      // index step between the axis directions
  ax[0] = -sup[2 * kQ]; ay[0] = -sup[3 * kQ];
  ax[1] = sup[0];       ay[1] = -sup[3 * kQ];
  ax[2] = sup[0];       ay[2] = sup[kQ];
  ax[3] = -sup[2 * kQ]; ay[3] = sup[kQ];
  int n = 4;
  {
    double* sx_ = ax; double* sy_ = ay; double* tx_ = bx; double* ty_ = by;
    for (int d = 1; d < kDirs; ++d) {
      if ((d % kQ) == 0) continue;
      n = clip_halfplane(sx_, sy_, n, (double)kDirA[d * kDirStep], (double)kDirB[d * kDirStep], sup[d], tx_, ty_);
      double* t = sx_; sx_ = tx_; tx_ = t;
      t = sy_; sy_ = ty_; ty_ = t;
    }
    if (sx_ != ax) {
      for (int i = 0; i < n; ++i) { ax[i] = sx_[i]; ay[i] = sy_[i]; }
    }
  }
  const double a_out = shoelace(ax, ay, n) * (1.0 + 1e-12);
  const double wh = (double)c.w * (double)c.h;
  const double rlo = a_in / wh, rhi = a_out / wh;
  ratio_lo[pair] = rlo; ratio_hi[pair] = rhi;
  if (nh <= min_hits) return;                                  // densify:898 -> dropped
  if (!exact_all && rlo >= thr + near_eps) {
    keep_out[pair] = 1; ratio_out[pair] = rlo; area_out[pair] = a_in; flags_out[pair] = 16;
  } else if (!exact_all && rhi < thr - near_eps) {
    ratio_out[pair] = rhi; area_out[pair] = a_out; flags_out[pair] = 16;
  } else {
    need[pair] = 1;
  }
}

// candidate capacity plan: cap = need ? (caps_in ? caps_in : min(n_hit, cap_max)) : 0,
// cand_offsets = exclusive scan (single CTA).
__global__ void __launch_bounds__(1024) cand_plan_kernel(
    const int32_t* __restrict__ n_hit, const uint8_t* __restrict__ need,
    const int32_t* __restrict__ caps_in, const float* __restrict__ poly,
    const double* __restrict__ ratio_lo, const double* __restrict__ ratio_hi, int64_t n_pairs,
    int cap_max, int64_t pool, int64_t* __restrict__ cand_offsets) {
  __shared__ long long sm[33];
  long long carry = 0;
  for (int64_t s = 0; s < n_pairs; s += 1024) {
    const int64_t i = s + threadIdx.x;
    long long v = 0;
    if (i < n_pairs && need[i]) {
      if (caps_in) {
        v = caps_in[i];
      } else {
        // candidates = the hits outside the inner polygon.  With the area bounds of the pair at
        // hand their share is about 1 - (inner area / hull area) <= 1 - ratio_lo / ratio_hi: twice
        // that share of the hits plus cap_max, never more than the hits (without the bounds: half
        // of the hits).  Overflowing pairs are re-run by the caller with their exact counts.
        const int h = n_hit[i];
        long long want = (h >> 1) > cap_max ? (h >> 1) : cap_max;
        if (ratio_lo && ratio_hi) {
          const double hi_ = ratio_hi[i], lo_ = ratio_lo[i];
          const double share = (hi_ > 0.0 && lo_ >= 0.0 && lo_ <= hi_) ? (1.0 - lo_ / hi_) : 1.0;
          want = (long long)(2.0 * share * (double)h) + cap_max;
        }
        v = h < want ? h : want;
        // no inner polygon (degenerate extremes): every hit is a candidate
        if (poly && poly[i * kPolyFloats] == 0.0f) v = h;
      }
      if (v < 1) v = 1;
    }
    long long tot;
    const long long ex = block_exclusive_sum<long long, 1024>(v, sm, &tot);
    // offsets clamped to the pool: pairs past it get a shrunken (or empty)
    // segment, overflow and are re-run by the caller
    if (i < n_pairs) cand_offsets[i] = (carry + ex) < pool ? (carry + ex) : pool;
    carry += tot;
  }
  if (threadIdx.x == 0) cand_offsets[n_pairs] = carry < pool ? carry : pool;
}

// ---- exact hull + area per pair ----------------------------------------------
// Threads per pair (template NT): 512 when there are few undecided pairs (a pair's chain of passes is
// latency bound: 4 warps per scheduler, one CTA per SM at 128 registers), 256 when there are many
// (two CTAs per SM overlap one pair's serial phases with the other's parallel ones).
constexpr int kHullMaxParts = 4;    // NT / 128 point partitions of the directional scan
constexpr int kHullThreadsMany = 256;   // NT when there are many pairs
constexpr int kHullBuf = 2048;      // points sorted per round (hull so far + new candidates)
constexpr int kHullChain = 1024;    // max vertices per monotone chain
constexpr int kHullFilterMin = 128; // candidates above which the directional pre-filter runs
constexpr int kHullTwoLevel = 192;  // sorted points above which the monotone chain runs in two levels
constexpr int kHullDynSmem = 2 * kHullBuf * 8;   // lower[] + upper[] (dynamic: the static limit is 48 KB)

__device__ __forceinline__ bool lex_less(float2 a, float2 b) {
  return a.x < b.x || (a.x == b.x && a.y < b.y);
}
// cross product of (b-a) x (c-a) evaluated like the reference's Python floats
// (densify:520-521): float64, no fma.
__device__ __forceinline__ double cross3(float2 a, float2 b, float2 c) {
  const double abx = __dsub_rn((double)b.x, (double)a.x), aby = __dsub_rn((double)b.y, (double)a.y);
  const double acx = __dsub_rn((double)c.x, (double)a.x), acy = __dsub_rn((double)c.y, (double)a.y);
  return __dsub_rn(__dmul_rn(abx, acy), __dmul_rn(aby, acx));
}

// One step of the monotone chain: pop while the turn (t2, t1, p) is not strictly convex, then push p.
// The two topmost stack entries live in registers.  The sign of the float64 cross product is first
// looked for in float32: |d32 - exact| <= 4 * 2^-24 * (|l| + |r|), the float64 value differs from the
// exact one by < 2^-50 * (|l| + |r|), so beyond 1e-6 * (|l| + |r|) (and 1e-30: no underflow) both have
// the sign of d32; the
// float64 evaluation (the reference's arithmetic) decides the rest, ties included.
struct ChainTop {
  float2 t1 = {0.f, 0.f}, t2 = {0.f, 0.f};
  int ns = 0;
  bool overflow = false;
};
__device__ __forceinline__ bool chain_pops(float2 t2, float2 t1, float2 p, bool lo_chain) {
  const float abx = t1.x - t2.x, aby = t1.y - t2.y, acx = p.x - t2.x, acy = p.y - t2.y;
  const float l = abx * acy, r = aby * acx, d = l - r;
  if (fabsf(d) > fmaxf(1e-6f * (fabsf(l) + fabsf(r)), 1e-30f)) return lo_chain ? d < 0.f : d > 0.f;
  const double c = cross3(t2, t1, p);
  return lo_chain ? (c <= 0.0) : (c >= 0.0);
}
__device__ __forceinline__ void chain_push(float2* st, ChainTop& c, float2 p, bool lo_chain, int cap) {
  while (c.ns >= 2 && chain_pops(c.t2, c.t1, p, lo_chain)) {
    --c.ns;
    c.t1 = c.t2;
    if (c.ns >= 2) c.t2 = st[c.ns - 2];
  }
  if (c.ns == 1 && c.t1.x == p.x && c.t1.y == p.y) --c.ns;
  if (c.ns < cap) {
    st[c.ns++] = p;
    c.t2 = c.t1;
    c.t1 = p;
  } else {
    c.overflow = true;
  }
}

// ---- directional pre-filter of a pair's candidates (in place) ---------------------
// One round: the extreme candidates along 64 evenly spaced directions (warp w
// owns 16 of them) form, in direction order, a closed chain; a point strictly
// to the left of every edge of that chain has winding number >= 1 about it, so
// it lies inside the convex hull of the chain's vertices and can never be a
// hull vertex (nor touch the reference's `cross <= 0` rule).  Such points are
// dropped; everything within `margin` of an edge is kept.  The float32 edge
// evaluation and margin are those of sweep 2 (ptgs_visibility_bounds).
constexpr int kFiltDirs = 64;
struct FilterSmem {
  float2 dir[kFiltDirs];
  int idx[kFiltDirs];
  float2 vert[kFiltDirs];
  float ea[kFiltDirs], eb[kFiltDirs], ec[kFiltDirs];
  int m;
  int out;
  int scan[33];
  uint32_t pkey[kHullMaxParts][kFiltDirs];
  int pidx[kHullMaxParts][kFiltDirs];
};

template <int NT>
__device__ int filter_round(float2* cu, int n, float margin, FilterSmem& F) {
  constexpr int kHullThreads = NT, kHullParts = NT / 128;
  const int tid = threadIdx.x, lane = tid & 31, w = (tid >> 5) & 3, part = tid >> 7;
  // 1. extremes: warp (w, part) scans every kHullParts-th 128-point slice for the 16 directions of group w
  float best[16];
  int bi[16];
#pragma unroll
  for (int d = 0; d < 16; ++d) { best[d] = -__int_as_float(0x7f800000); bi[d] = 0; }
  // 4 independent loads in flight per lane (the loop is latency bound otherwise)
  for (int i0 = lane + 128 * part; i0 < n; i0 += 128 * kHullParts) {
    float2 q[4];
#pragma unroll
    for (int r = 0; r < 4; ++r) {
      const int i = i0 + 32 * r;
      q[r] = i < n ? cu[i] : cu[i0];
    }
#pragma unroll
    for (int r = 0; r < 4; ++r) {
      const int i = (i0 + 32 * r) < n ? (i0 + 32 * r) : i0;
#pragma unroll
      for (int d = 0; d < 16; ++d) {
        const float2 dv = F.dir[w * 16 + d];
        const float dot = fmaf(q[r].x, dv.x, q[r].y * dv.y);
        if (dot > best[d]) { best[d] = dot; bi[d] = i; }
      }
    }
  }
#pragma unroll
  for (int d = 0; d < 16; ++d) {
    const uint32_t key = order_key(best[d]);
    const uint32_t mk = __reduce_max_sync(kFullMask, key);
    const unsigned who = __ballot_sync(kFullMask, key == mk);
    const int id = __shfl_sync(kFullMask, bi[d], __ffs(who) - 1);
    if (lane == 0) { F.pkey[part][w * 16 + d] = mk; F.pidx[part][w * 16 + d] = id; }
  }
  __syncthreads();
  if (tid < kFiltDirs) {
    uint32_t bk = F.pkey[0][tid];
    int bid = F.pidx[0][tid];
#pragma unroll
    for (int p = 1; p < kHullParts; ++p)
      if (F.pkey[p][tid] > bk) { bk = F.pkey[p][tid]; bid = F.pidx[p][tid]; }
    F.idx[tid] = bid;
  }
  __syncthreads();
  if (tid < kFiltDirs) F.vert[tid] = cu[F.idx[tid]];
  __syncthreads();
  if (tid == 0) {
    // distinct consecutive vertices, edges as float half planes a*u + b*v + c
    float2 v[kFiltDirs];
    int m = 0;
    for (int k = 0; k < kFiltDirs; ++k) {
      const float2 q = F.vert[k];
      if (m && v[m - 1].x == q.x && v[m - 1].y == q.y) continue;
      v[m++] = q;
    }
    while (m > 1 && v[m - 1].x == v[0].x && v[m - 1].y == v[0].y) --m;
    if (m >= 3) {
      for (int i = 0; i < m; ++i) {
        const int k = (i + 1 == m) ? 0 : i + 1;
        const double a = -((double)v[k].y - (double)v[i].y), b = (double)v[k].x - (double)v[i].x;
        F.ea[i] = (float)a;
        F.eb[i] = (float)b;
        F.ec[i] = (float)(-(a * (double)v[i].x + b * (double)v[i].y));
      }
    }
    F.m = m >= 3 ? m : 0;
    F.out = 0;
  }
  __syncthreads();
  const int m = F.m;
  if (m == 0) return n;
  // 2. filter + in-place compaction, 128 points per step (reads of a step finish
  //    before its writes, which land at or before the step's first position)
  // 4 consecutive points per thread and step (order preserving)
  constexpr int kPer = 4;
  for (int base = 0; base < n; base += kHullThreads * kPer) {
    float2 q[kPer];
    int kp[kPer];
    int mine = 0;
#pragma unroll
    for (int r = 0; r < kPer; ++r) {
      const int i = base + tid * kPer + r;
      q[r] = i < n ? cu[i] : make_float2(0.f, 0.f);
    }
#pragma unroll
    for (int r = 0; r < kPer; ++r) {
      const int i = base + tid * kPer + r;
      bool inside = true;
      for (int e = 0; e < m && inside; ++e) {
        const float f =
            __fadd_rn(__fadd_rn(__fmul_rn(F.ea[e], q[r].x), __fmul_rn(F.eb[e], q[r].y)), F.ec[e]);
        inside = f > margin;
      }
      kp[r] = (i < n && !inside) ? 1 : 0;
      mine += kp[r];
    }
    int tot;
    const int ex = block_exclusive_sum<int, kHullThreads>(mine, F.scan, &tot);
    const int o = F.out;
    int w0 = o + ex;
#pragma unroll
    for (int r = 0; r < kPer; ++r)
      if (kp[r]) cu[w0++] = q[r];
    __syncthreads();
    if (tid == 0) F.out = o + tot;
    __syncthreads();
  }
  return F.out;
}

// ---- multi-CTA pre-filter of the largest candidate lists ---------------------------------------
// hull_kernel gives a pair ONE CTA; a pair with tens of thousands of candidates then sets the
// duration of the whole launch (and, sharded, of the step on every rank).  Pairs with more than
// kPfMinK candidates therefore get one directional filter round up front, spread over as many CTAs
// as they have 4096-candidate chunks: (1) per chunk the extreme candidates along 64 directions,
// merged per pair with atomics; (2) per pair the polygon through those extremes; (3) per chunk the
// candidates that are not strictly inside it (same float32 edge evaluation and margin as
// filter_round) are appended to a second buffer.  hull_kernel then starts from the survivors
// (a set, not a sequence: the hull sorts its input, so the append order does not matter).
constexpr int kPfChunk = 4096;
constexpr int kPfMinK = 6144;
constexpr int kPfPolyFloats = 4 + 3 * kFiltDirs;     // [0] = edges, then (a, b, c) per edge

// single CTA: per undecided pair i (position in the sub-list) whether it is pre-filtered
// (count2[i] = 0, else -1), its slice of the second buffer (off2[i]) and the running number of
// chunks before it (chunk_prefix[i]; [n_pairs] = total).
__global__ void __launch_bounds__(1024) pf_plan_kernel(
    const int64_t* __restrict__ cand_offsets, const int32_t* __restrict__ cand_count,
    const int32_t* __restrict__ pair_map, const int64_t* __restrict__ n_sub_dev, int64_t n_pairs,
    int64_t uv2_cap, int64_t* __restrict__ chunk_prefix, int64_t* __restrict__ off2,
    int32_t* __restrict__ count2) {
  __shared__ long long sm[33];
  int64_t n_sub = n_pairs;
  if (n_sub_dev && *n_sub_dev < n_sub) n_sub = *n_sub_dev;
  long long carry_sp = 0, carry_ch = 0;
  for (int64_t s0 = 0; s0 < n_pairs; s0 += 1024) {
    const int64_t i = s0 + threadIdx.x;
    long long K = 0;
    if (i < n_sub) {
      const int64_t pair = pair_map ? (int64_t)pair_map[i] : i;
      const long long k = cand_count[pair];
      const long long cap = cand_offsets[pair + 1] - cand_offsets[pair];
      if (k > kPfMinK && k <= cap) K = k;
    }
    long long tot_sp;
    const long long sp = carry_sp + block_exclusive_sum<long long, 1024>(K, sm, &tot_sp);
    carry_sp += tot_sp;
    const bool big = K > 0 && (sp + K) <= uv2_cap;       // the buffer is granted in list order
    const long long ch = big ? (K + kPfChunk - 1) / kPfChunk : 0;
    long long tot_ch;
    const long long cp = carry_ch + block_exclusive_sum<long long, 1024>(ch, sm, &tot_ch);
    carry_ch += tot_ch;
    if (i < n_pairs) {
      chunk_prefix[i] = cp;
      off2[i] = sp;
      count2[i] = big ? 0 : -1;
    }
  }
  if (threadIdx.x == 0) chunk_prefix[n_pairs] = carry_ch;
}

struct PfItem { int64_t sub; int64_t pair; int chunk; int n; const float2* pts; };

__device__ __forceinline__ bool pf_item(int64_t item, const int64_t* __restrict__ chunk_prefix, int64_t n_pairs,
                                        const int32_t* __restrict__ pair_map,
                                        const int64_t* __restrict__ cand_offsets,
                                        const int32_t* __restrict__ cand_count, const float2* cand_uv,
                                        PfItem& it) {
  int64_t lo = 0, hi = n_pairs - 1;            // last sub index with chunk_prefix <= item that owns chunks
  while (lo < hi) {
    const int64_t mid = (lo + hi + 1) >> 1;
    if (chunk_prefix[mid] <= item) lo = mid; else hi = mid - 1;
  }
  it.sub = lo;
  it.pair = pair_map ? (int64_t)pair_map[lo] : lo;
  it.chunk = (int)(item - chunk_prefix[lo]);
  const int K = cand_count[it.pair];
  const int start = it.chunk * kPfChunk;
  if (start >= K) return false;                // cannot happen for a planned item
  it.n = (K - start) < kPfChunk ? (K - start) : kPfChunk;
  it.pts = cand_uv + cand_offsets[it.pair] + start;
  return true;
}

// (1) chunk extremes along the 64 directions -> ext[sub][64] = max (ordered key << 32 | index in the pair)
__global__ void __launch_bounds__(256) pf_extremes_kernel(
    const float2* __restrict__ cand_uv, const int64_t* __restrict__ cand_offsets,
    const int32_t* __restrict__ cand_count, const int32_t* __restrict__ pair_map, int64_t n_pairs,
    const int64_t* __restrict__ chunk_prefix, unsigned long long* __restrict__ ext) {
  __shared__ float2 s_pts[kPfChunk];
  __shared__ float s_best[4][kFiltDirs];
  __shared__ int s_idx[4][kFiltDirs];
  const int64_t n_items = chunk_prefix[n_pairs];
  const int d = threadIdx.x & (kFiltDirs - 1), part = threadIdx.x >> 6;
  float sn, cs;
  sincospif((float)d * (2.0f / kFiltDirs), &sn, &cs);
  for (int64_t item = blockIdx.x; item < n_items; item += gridDim.x) {
    PfItem it;
    const bool ok = pf_item(item, chunk_prefix, n_pairs, pair_map, cand_offsets, cand_count, cand_uv, it);
    if (ok)
      for (int j = threadIdx.x; j < it.n; j += 256) s_pts[j] = it.pts[j];   // coalesced; the scan below
    __syncthreads();                                                         // reads broadcast values
    float best = -__int_as_float(0x7f800000);
    int bi = 0;
    if (ok) {
#pragma unroll 4
      for (int j = part; j < it.n; j += 4) {
        const float2 q = s_pts[j];
        const float dot = fmaf(q.x, cs, q.y * sn);
        if (dot > best) { best = dot; bi = j; }
      }
    }
    s_best[part][d] = best;
    s_idx[part][d] = bi;
    __syncthreads();
    if (ok && part == 0) {
#pragma unroll
      for (int p = 1; p < 4; ++p)
        if (s_best[p][d] > best) { best = s_best[p][d]; bi = s_idx[p][d]; }
      if (best > -__int_as_float(0x7f800000)) {
        const unsigned long long val =
            ((unsigned long long)order_key(best) << 32) | (unsigned)(it.chunk * kPfChunk + bi);
        atomicMax(ext + it.sub * kFiltDirs + d, val);
      }
    }
    __syncthreads();
  }
}

// (2) one warp per pre-filtered pair: polygon through its 64 extremes (distinct consecutive
// vertices, edges as float half planes a*u + b*v + c -- the construction of filter_round)
__global__ void __launch_bounds__(128) pf_polygon_kernel(
    const float2* __restrict__ cand_uv, const int64_t* __restrict__ cand_offsets,
    const int32_t* __restrict__ pair_map, int64_t n_pairs, const int32_t* __restrict__ count2,
    const unsigned long long* __restrict__ ext, float* __restrict__ poly) {
  __shared__ float2 s_v[4][kFiltDirs];
  __shared__ int s_cnt[4];
  const int wi = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int64_t i = (int64_t)blockIdx.x * 4 + wi;
  if (i >= n_pairs || count2[i] < 0) return;       // warp uniform
  const int64_t pair = pair_map ? (int64_t)pair_map[i] : i;
  const float2* cu = cand_uv + cand_offsets[pair];
  float* out = poly + i * kPfPolyFloats;
  // lanes fetch the extremes of directions lane and lane + 32; a direction without one repeats its
  // predecessor (dropped as a duplicate below); direction 0 always has one (every planned pair has candidates)
  const float nanv = __int_as_float(0x7fc00000);
#pragma unroll
  for (int h = 0; h < 2; ++h) {
    const unsigned long long e = ext[i * kFiltDirs + lane + 32 * h];
    s_v[wi][lane + 32 * h] = e ? cu[(uint32_t)(e & 0xffffffffull)] : make_float2(nanv, nanv);
  }
  __syncwarp();
  if (lane == 0) {
    float2* v = s_v[wi];
    int m = 0;
    for (int k = 0; k < kFiltDirs; ++k) {
      const float2 q = v[k];
      if (q.x != q.x) continue;                                          // no extreme along k
      if (m && v[m - 1].x == q.x && v[m - 1].y == q.y) continue;
      v[m++] = q;                                                         // m <= k: in place
    }
    while (m > 1 && v[m - 1].x == v[0].x && v[m - 1].y == v[0].y) --m;
    s_cnt[wi] = m;
  }
  __syncwarp();
  const int m = s_cnt[wi];
  if (m < 3) { if (lane == 0) out[0] = 0.0f; return; }
  for (int k = lane; k < m; k += 32) {
    const int k2 = (k + 1 == m) ? 0 : k + 1;
    const float2 p = s_v[wi][k], q = s_v[wi][k2];
    const double a = -((double)q.y - (double)p.y), b = (double)q.x - (double)p.x;
    out[4 + 3 * k] = (float)a;
    out[5 + 3 * k] = (float)b;
    out[6 + 3 * k] = (float)(-(a * (double)p.x + b * (double)p.y));
  }
  if (lane == 0) out[0] = (float)m;
}

// (3) per chunk: candidates that are not strictly inside the pair's polygon -> uv2[off2 + ...]
__global__ void __launch_bounds__(256) pf_filter_kernel(
    const float2* __restrict__ cand_uv, const int64_t* __restrict__ cand_offsets,
    const int32_t* __restrict__ cand_count, const int32_t* __restrict__ pair_map, int64_t n_pairs,
    const int32_t* __restrict__ pair_cam, const float* __restrict__ cameras,
    const int64_t* __restrict__ chunk_prefix, const float* __restrict__ poly,
    const int64_t* __restrict__ off2, int32_t* __restrict__ count2, float2* __restrict__ uv2) {
  __shared__ float s_e[3 * kFiltDirs];
  __shared__ int s_m;
  const int64_t n_items = chunk_prefix[n_pairs];
  const int lane = threadIdx.x & 31;
  for (int64_t item = blockIdx.x; item < n_items; item += gridDim.x) {
    PfItem it;
    const bool ok = pf_item(item, chunk_prefix, n_pairs, pair_map, cand_offsets, cand_count, cand_uv, it);
    __syncthreads();                           // previous item's edges are no longer read
    if (ok) {
      const float* pl = poly + it.sub * kPfPolyFloats;
      if (threadIdx.x == 0) s_m = (int)pl[0];
      for (int k = threadIdx.x; k < 3 * kFiltDirs; k += 256) s_e[k] = pl[4 + k];
    }
    __syncthreads();
    if (!ok) continue;
    const int m = s_m;
    const float* rec = cameras + (int64_t)pair_cam[it.pair] * kCamFloats;
    const float Dm = fmaxf(rec[16], rec[17]);
    const float margin = 4e-6f * Dm * Dm;
    for (int j0 = 0; j0 < it.n; j0 += 256) {
      const int j = j0 + threadIdx.x;
      bool keep = false;
      float2 q = make_float2(0.f, 0.f);
      if (j < it.n) {
        q = it.pts[j];
        bool inside = m >= 3;
        for (int e = 0; e < m && inside; ++e) {
          const float f = __fadd_rn(__fadd_rn(__fmul_rn(s_e[3 * e], q.x), __fmul_rn(s_e[3 * e + 1], q.y)),
                                    s_e[3 * e + 2]);
          inside = f > margin;
        }
        keep = !inside;
      }
      const unsigned mk = __ballot_sync(kFullMask, keep);
      if (mk) {
        int base = 0;
        if (lane == 0) base = atomicAdd(count2 + it.sub, __popc(mk));
        base = __shfl_sync(kFullMask, base, 0);
        if (keep) uv2[off2[it.sub] + base + __popc(mk & lanemask_lt())] = q;
      }
    }
  }
}

template <int NT>
__global__ void __launch_bounds__(NT, NT == 256 ? 3 : 1) hull_kernel(
    float2* __restrict__ cand_uv, const int64_t* __restrict__ cand_offsets,
    const int32_t* __restrict__ cand_count, const int32_t* __restrict__ n_hit, int64_t n_pairs,
    const int32_t* __restrict__ pair_cam, const float* __restrict__ cameras, int min_hits,
    double thr, double near_eps, double* __restrict__ area_out, double* __restrict__ ratio_out,
    uint8_t* __restrict__ keep_out, uint8_t* __restrict__ flags_out,
    const int32_t* __restrict__ pair_map, const int64_t* __restrict__ n_sub_dev,
    int32_t* __restrict__ status /* [0] capacity overflows, [1] vertex-budget overflows; may be null */,
    const int32_t* __restrict__ pf_count2, const int64_t* __restrict__ pf_off2,
    float2* __restrict__ pf_uv2 /* pre-filtered lists of the largest pairs; may be null */) {
  constexpr int kHullThreads = NT;
  __shared__ float2 buf[kHullBuf];
  extern __shared__ __align__(16) unsigned char hull_dyn[];   // kHullDynSmem bytes
  float2* lower = reinterpret_cast<float2*>(hull_dyn);       // final chains hold <= kHullChain vertices;
  float2* upper = lower + kHullBuf;                          // the segment chains of the two-level pass
                                                             // use the whole array
  __shared__ int s_seg[2][32];
  __shared__ int s_h, s_nl, s_nu, s_flags, s_col;
  if (blockIdx.x >= n_pairs) return;
  if (n_sub_dev && (int64_t)blockIdx.x >= *n_sub_dev) return;
  const int64_t pair = pair_map ? (int64_t)pair_map[blockIdx.x] : (int64_t)blockIdx.x;
  const int tid = threadIdx.x;
  const int64_t off = cand_offsets[pair];
  const int cap = (int)(cand_offsets[pair + 1] - off);
  int K = cand_count[pair];
  int flags = 0;
  if (K > cap) {
    // candidate capacity overflow: the result would be invalid, the caller
    // re-runs the pair with cand_count[pair] slots
    if (tid == 0) {
      area_out[pair] = 0.0; ratio_out[pair] = 0.0; keep_out[pair] = 0; flags_out[pair] = 4;
      if (status) atomicAdd(status, 1);
    }
    return;
  }
  if (tid == 0) { s_h = 0; s_flags = 0; s_col = 0; }
  __syncthreads();
  const float inf = __int_as_float(0x7f800000);
  float2* cu = cand_uv + off;
  if (pf_count2 && pf_count2[blockIdx.x] >= 0) {     // the multi-CTA pre-filter already cut this list
    cu = pf_uv2 + pf_off2[blockIdx.x];
    K = pf_count2[blockIdx.x];
  }
  if (K > kHullFilterMin) {
    __shared__ FilterSmem F;
    const float* rec = cameras + (int64_t)pair_cam[pair] * kCamFloats;
    const float Dm = fmaxf(rec[16], rec[17]);
    const float margin = 4e-6f * Dm * Dm;
    // every round uses the 64 directions shifted by another fraction of the
    // angular step, so successive rounds cut with different polygons
    const float frac[6] = {0.0f, 0.5f, 0.25f, 0.75f, 0.125f, 0.625f};
    for (int it = 0; it < 6 && K > kHullFilterMin; ++it) {
      if (tid < kFiltDirs) {
        float sn, cs;
        sincospif(((float)tid + frac[it]) * (2.0f / kFiltDirs), &sn, &cs);
        F.dir[tid] = make_float2(cs, sn);
      }
      __syncthreads();
      const int kn = filter_round<NT>(cu, K, margin, F);
      const bool progress = kn < K - K / 16;
      K = kn;
      if (!progress && it >= 2) break;
    }
    __threadfence_block();
    __syncthreads();
  }
  int pos = 0;
  bool single_round = true;
  int rounds = 0;
  while (pos < K || rounds == 0) {
    const int h = s_h;
    int m = K - pos;
    if (m > kHullBuf - h) m = kHullBuf - h;
    const int n = h + m;
    int n2 = 1;
    while (n2 < n) n2 <<= 1;
    for (int i = tid; i < n2 - h; i += kHullThreads)
      buf[h + i] = (i < m) ? cu[pos + i] : make_float2(inf, inf);
    __syncthreads();
    // bitonic sort of buf[0..n2) by (u, v)
    for (int k = 2; k <= n2; k <<= 1) {
      for (int j = k >> 1; j > 0; j >>= 1) {
        for (int i = tid; i < n2; i += kHullThreads) {
          const int l = i ^ j;
          if (l > i) {
            const float2 a = buf[i], b = buf[l];
            const bool up = (i & k) == 0;
            if (up ? lex_less(b, a) : lex_less(a, b)) { buf[i] = b; buf[l] = a; }
          }
        }
        __syncthreads();
      }
    }
    // Andrew's monotone chain; pop on cross <= 0 (densify:522).  The lower and the upper chain are
    // independent (warp 0 / warp 1).  A chain is sequential, and with one lane walking up to 2048
    // sorted points it was what every other warp of the CTA waited for, so above kHullTwoLevel points
    // it runs in two levels: the 32 lanes of the warp each build the chain of one of 32 consecutive
    // segments of the sorted array (a hull vertex of the whole set is a hull vertex of its segment),
    // then lane 0 runs the same chain over the concatenated segment chains, in place.  The chain of
    // a point set does not depend on which non-vertices were removed first, so the result is the
    // one-level chain's.
    if (tid < 64) {
      const bool lo_chain = tid < 32;
      const int ln = tid & 31;
      float2* st = lo_chain ? lower : upper;
      int* seg_n = s_seg[lo_chain ? 0 : 1];
      const bool two = n > kHullTwoLevel;
      const int seg = two ? (n + 31) >> 5 : n;
      if (two) {
        const int i0 = ln * seg, i1 = (i0 + seg < n) ? i0 + seg : n;
        ChainTop c;
        for (int i = i0; i < i1; ++i) chain_push(st + i0, c, buf[i], lo_chain, seg);
        seg_n[ln] = c.ns;
        __syncwarp();
      }
      if (ln == 0) {
        ChainTop c;
        if (two) {
          for (int sgm = 0; sgm < 32; ++sgm) {
            const int cnt = seg_n[sgm];
            const float2* src = st + sgm * seg;
            for (int k = 0; k < cnt; ++k) chain_push(st, c, src[k], lo_chain, kHullChain);
          }
        } else {
          for (int i = 0; i < n; ++i) chain_push(st, c, buf[i], lo_chain, kHullChain);
        }
        if (c.overflow) atomicOr(&s_flags, 8);
        if (lo_chain) s_nl = c.ns; else s_nu = c.ns;
      }
    }
    __syncthreads();
    pos += m;
    ++rounds;
    if (pos < K) {
      // more candidates than one round holds: keep the hull, merge the rest
      single_round = false;
      const int nl = s_nl, nu = s_nu;
      for (int i = tid; i < nl; i += kHullThreads) buf[i] = lower[i];
      for (int i = tid; i < nu - 2; i += kHullThreads) buf[nl + i] = upper[nu - 2 - i];
      __syncthreads();
      if (tid == 0) s_h = nl + (nu > 2 ? nu - 2 : 0);
      __syncthreads();
      if (s_h > kHullBuf / 2) { if (tid == 0) s_flags |= 8; __syncthreads(); break; }
    }
  }
  // hull, counter-clockwise in (u, v): lower left->right, then upper right->left
  const int nl = s_nl, nu = s_nu;
  const int nh = (nl >= 1 && nu >= 1) ? nl + (nu > 2 ? nu - 2 : 0) : 0;
  auto vert = [&](int i) -> float2 { return i < nl ? lower[i] : upper[nu - 2 - (i - nl)]; };
  // first-ray check (flag bit1): another candidate exactly on the edge that
  // leaves the pivot counter-clockwise.  Only when one round saw every candidate.
  if (nh >= 3 && single_round) {
    // the pivot is the lowest-v, then lowest-u vertex
    int ip = 0;
    float2 pv = vert(0);
    for (int i = 1; i < nh; ++i) {
      const float2 q = vert(i);
      if (q.y < pv.y || (q.y == pv.y && q.x < pv.x)) { pv = q; ip = i; }
    }
    const float2 nx = vert(ip + 1 == nh ? 0 : ip + 1);
    const int n = K;
    int found = 0;
    for (int i = tid; i < n; i += kHullThreads) {
      const float2 q = buf[i];
      const bool same = (q.x == pv.x && q.y == pv.y) || (q.x == nx.x && q.y == nx.y);
      if (!same && cross3(pv, nx, q) == 0.0) found = 1;
    }
    if (found) atomicOr(&s_col, 1);
  }
  __syncthreads();
  if (tid == 0) {
    double area = 0.0;
    if (nh >= 3) {                                       // densify:529 (len(hull) < 3 -> 0)
      const float2 o = vert(0);
      double acc = 0.0;
      for (int i = 1; i + 1 < nh; ++i) {
        const float2 p = vert(i), q = vert(i + 1);
        const double ax = (double)p.x - (double)o.x, ay = (double)p.y - (double)o.y;
        const double bx = (double)q.x - (double)o.x, by = (double)q.y - (double)o.y;
        acc += ax * by - ay * bx;
      }
      area = 0.5 * fabs(acc);                             // densify:535
    }
    const float* rec = cameras + (int64_t)pair_cam[pair] * kCamFloats;
    const double ratio = area / ((double)rec[16] * (double)rec[17]);   // densify:537-538
    const int nhit = n_hit[pair];
    const bool enough = nhit > min_hits;                  // densify:898
    flags |= s_flags;
    if ((s_flags & 8) && status) atomicAdd(status + 1, 1);
    if (s_col) flags |= 2;
    if (enough && fabs(ratio - thr) <= near_eps) flags |= 1;
    area_out[pair] = enough ? area : 0.0;
    ratio_out[pair] = enough ? ratio : 0.0;
    keep_out[pair] = (enough && ratio >= thr) ? 1 : 0;    // densify:904
    flags_out[pair] = (uint8_t)flags;
  }
}

__global__ void __launch_bounds__(256) vis_init_kernel(int64_t n_pairs, int32_t* n_hit,
                                                       uint64_t* pivot, uint64_t* extreme,
                                                       int32_t* cand_count) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n_pairs * kDirs; i += stride) {
    if (extreme) extreme[i] = 0ull;
    if (i < n_pairs) {
      if (n_hit) n_hit[i] = 0;
      if (pivot) pivot[i] = ~0ull;
      if (cand_count) cand_count[i] = 0;
    }
  }
}

__global__ void coverage_setup_kernel(int64_t n, int32_t cam_count, int64_t* pt_off,
                                      int64_t* cam_off) {
  pt_off[0] = 0; pt_off[1] = n;
  cam_off[0] = 0; cam_off[1] = cam_count;
}

// ---- interior rectangle K per pair (MODE 1 fast path) ---------------------------------
// Sweep 1 is first run on a 1-in-`stride` sample of every block.  The extreme
// hits of the sample span a convex polygon Q inside the convex hull of all
// hits; K is an axis-parallel rectangle inside Q (the bounding box of Q scaled
// about Q's vertex centroid until a corner touches an edge), shrunk by
// `margin_px`.  Every point of K is a hit (K lies inside the image, the margin
// is >> the float32 evaluation error of the homogeneous test) and is strictly
// inside the hull of the hits, so it can be neither the Graham pivot nor an
// extreme along any direction: the full sweep only counts such points.
__global__ void __launch_bounds__(1024) sample_offsets_kernel(
    const int64_t* __restrict__ pt_offsets, int L, int stride, int64_t* __restrict__ s_offsets) {
  __shared__ long long sm[33];
  long long carry = 0;
  for (int s0 = 0; s0 < L; s0 += 1024) {
    const int b = s0 + threadIdx.x;
    long long v = 0;
    if (b < L) v = (pt_offsets[b + 1] - pt_offsets[b] + stride - 1) / stride;
    long long tot;
    const long long ex = block_exclusive_sum<long long, 1024>(v, sm, &tot);
    if (b < L) s_offsets[b] = carry + ex;
    carry += tot;
  }
  if (threadIdx.x == 0) s_offsets[L] = carry;
}

__global__ void __launch_bounds__(256) sample_gather_kernel(
    const float* __restrict__ px, const float* __restrict__ py, const float* __restrict__ pz,
    const int64_t* __restrict__ pt_offsets, const int64_t* __restrict__ s_offsets, int L, int stride,
    int64_t capacity, float* __restrict__ sx, float* __restrict__ sy, float* __restrict__ sz) {
  __shared__ int64_t so[kSwMaxBlocks + 1];
  for (int i = threadIdx.x; i <= L; i += blockDim.x) so[i] = s_offsets[i];
  __syncthreads();
  const int64_t n = so[L] < capacity ? so[L] : capacity;
  const int64_t step = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += step) {
    int lo = 0, hi = L - 1;
    while (lo < hi) {
      const int mid = (lo + hi + 1) >> 1;
      if (so[mid] <= i) lo = mid; else hi = mid - 1;
    }
    const int64_t src = pt_offsets[lo] + (i - so[lo]) * stride;
    sx[i] = px[src]; sy[i] = py[src]; sz[i] = pz[src];
  }
}

// Sweep 1 on the 1-in-stride SAMPLE, pair-centric: one warp per (block, camera) pair walks the
// block's sample points (a few thousand) with the exact arithmetic of the sweep and keeps the hit
// count, the pivot and the directional extremes in registers; one reduction per pair at the end,
// no atomics.  The tile sweep paid its whole item machinery for 1 / 128 of the points (0.24 ms at
// c4, and the same fixed cost on every rank of a sharded run).  Any valid set of extreme HITS
// gives a valid interior polygon; ties are broken like the sweep's (largest point index).
__global__ void __launch_bounds__(256) sample_pairs_kernel(
    const float* __restrict__ sx, const float* __restrict__ sy, const float* __restrict__ sz,
    const int64_t* __restrict__ s_offsets, int L, const float* __restrict__ cameras,
    const int64_t* __restrict__ cam_offsets, const int32_t* __restrict__ cam_index,
    int32_t* __restrict__ n_hit_s, uint64_t* __restrict__ pivot_s /* may be null */,
    uint64_t* __restrict__ extreme_s) {
  const int64_t n_pairs = cam_offsets[L];
  const int lane = threadIdx.x & 31;
  const int64_t pair = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (pair >= n_pairs) return;                       // warp uniform
  int lo = 0, hi = L - 1;
  while (lo < hi) {
    const int mid = (lo + hi + 1) >> 1;
    if (cam_offsets[mid] <= pair) lo = mid; else hi = mid - 1;
  }
  const int64_t p0 = s_offsets[lo];
  const int m = (int)(s_offsets[lo + 1] - p0);
  const Cam c = load_cam(cameras + (int64_t)cam_index[pair] * kCamFloats);
  const float w1 = __fadd_rn(c.w, 1.0f), h1 = __fadd_rn(c.h, 1.0f);
  int cnt = 0;
  unsigned long long pv = ~0ull;
  unsigned long long best[kDirs];
#pragma unroll
  for (int d = 0; d < kDirs; ++d) best[d] = 0ull;
  for (int i = lane; i < m; i += 32) {
    const Proj p = project_pre(c, w1, h1, sx[p0 + i], sy[p0 + i], sz[p0 + i]);
    float u, v;
    if (p.maybe && project_exact(c, p, u, v)) {
      ++cnt;
      u = __fadd_rn(u, 0.0f);   // -0.0 -> +0.0
      v = __fadd_rn(v, 0.0f);
      const unsigned long long pk = ((unsigned long long)__float_as_uint(v) << 32) | __float_as_uint(u);
      pv = pk < pv ? pk : pv;
      const float u2 = __fadd_rn(u, u), v2 = __fadd_rn(v, v);
#pragma unroll
      for (int d = 0; d < kDirs; ++d) {
        const unsigned long long val =
            ((unsigned long long)order_key(dir_value(d * kDirStep, u, v, u2, v2)) << 32) | (unsigned)i;
        best[d] = val > best[d] ? val : best[d];
      }
    }
  }
  cnt = __reduce_add_sync(kFullMask, cnt);
  {
    const uint32_t ph = (uint32_t)(pv >> 32), pl = (uint32_t)pv;
    const uint32_t mh = __reduce_min_sync(kFullMask, ph);
    const uint32_t ml = __reduce_min_sync(kFullMask, ph == mh ? pl : 0xffffffffu);
    pv = ((unsigned long long)mh << 32) | ml;
  }
#pragma unroll
  for (int d = 0; d < kDirs; ++d) {
    const uint32_t kh = (uint32_t)(best[d] >> 32), kl = (uint32_t)best[d];
    const uint32_t mh = __reduce_max_sync(kFullMask, kh);
    const uint32_t ml = __reduce_max_sync(kFullMask, kh == mh ? kl : 0u);
    best[d] = ((unsigned long long)mh << 32) | ml;
  }
  if (lane == 0) {
    n_hit_s[pair] = cnt;
    if (pivot_s) pivot_s[pair] = cnt ? pv : ~0ull;
#pragma unroll
    for (int d = 0; d < kDirs; ++d) extreme_s[pair * kDirs + d] = cnt ? best[d] : 0ull;
  }
}

__global__ void __launch_bounds__(128) krect_kernel(
    const float* __restrict__ sx, const float* __restrict__ sy, const float* __restrict__ sz,
    const int64_t* __restrict__ s_offsets, int L, const float* __restrict__ cameras,
    const int64_t* __restrict__ cam_offsets, const int32_t* __restrict__ cam_index,
    const int32_t* __restrict__ n_hit_s, const uint64_t* __restrict__ extreme_s, double margin_px,
    float* __restrict__ krect) {
  const int64_t n_pairs = cam_offsets[L];
  const int64_t pair = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (pair >= n_pairs) return;
  float* out = krect + pair * kKrectFloats;
  for (int i = 0; i < kKrectFloats; ++i) out[i] = 0.0f;
  out[2] = -1.0f; out[3] = -1.0f;
  if (n_hit_s[pair] < 3) return;
  int lo = 0, hi = L - 1;
  while (lo < hi) {
    const int mid = (lo + hi + 1) >> 1;
    if (cam_offsets[mid] <= pair) lo = mid; else hi = mid - 1;
  }
  const int64_t p0 = s_offsets[lo];
  const Cam c = load_cam(cameras + (int64_t)cam_index[pair] * kCamFloats);
  double vx[8], vy[8];
  int m = 0;
  for (int d = 0; d < 8; ++d) {
    const uint64_t e = extreme_s[pair * kDirs + d * (kDirs / 8)];      // the eight axis / diagonal directions
    if (!e) continue;
    const int64_t gi = p0 + (int64_t)(e & 0xffffffffull);
    const Proj p = project_pre(c, 0.0f, 0.0f, sx[gi], sy[gi], sz[gi]);
    float u, v;
    if (!project_exact(c, p, u, v)) continue;
    if (m && vx[m - 1] == (double)u && vy[m - 1] == (double)v) continue;
    vx[m] = u; vy[m] = v; ++m;
  }
  while (m > 1 && vx[m - 1] == vx[0] && vy[m - 1] == vy[0]) --m;
  if (m < 3) return;
  double gu = 0.0, gv = 0.0, u0 = vx[0], u1 = vx[0], v0 = vy[0], v1 = vy[0];
  for (int i = 0; i < m; ++i) {
    gu += vx[i]; gv += vy[i];
    u0 = fmin(u0, vx[i]); u1 = fmax(u1, vx[i]); v0 = fmin(v0, vy[i]); v1 = fmax(v1, vy[i]);
  }
  gu /= m; gv /= m;
  const double ha = 0.5 * (u1 - u0), hb = 0.5 * (v1 - v0);
  double t = 1.0;
  for (int i = 0; i < m; ++i) {
    const int k = (i + 1 == m) ? 0 : i + 1;
    const double nx = -(vy[k] - vy[i]), ny = vx[k] - vx[i];      // inward normal (CCW polygon)
    const double fg = nx * (gu - vx[i]) + ny * (gv - vy[i]);
    const double reach = fabs(nx) * ha + fabs(ny) * hb;
    if (!(fg > 0.0)) return;                                      // degenerate / not CCW: no K
    if (reach > 0.0) t = fmin(t, fg / reach);
  }
  // the polygon itself, every edge moved inwards by margin_px (unit normals);
  // its vertices are hits, so it lies inside the image
  for (int i = 0; i < m; ++i) {
    const int k = (i + 1 == m) ? 0 : i + 1;
    const double nx = -(vy[k] - vy[i]), ny = vx[k] - vx[i];
    const double len = sqrt(nx * nx + ny * ny);
    out[8 + 3 * i] = (float)(nx / len);
    out[9 + 3 * i] = (float)(ny / len);
    out[10 + 3 * i] = (float)(-(nx * vx[i] + ny * vy[i]) / len - margin_px - 1e-3 * (fabs(vx[i]) + fabs(vy[i])) * 1e-3);
  }
  out[4] = (float)m;
  const double hu = t * ha * (1.0 - 1e-6) - margin_px, hv = t * hb * (1.0 - 1e-6) - margin_px;
  if (!(hu > 0.0) || !(hv > 0.0)) return;
  // stay strictly inside the image whatever the rounding to float
  if (gu - hu < 1.0 || gv - hv < 1.0 || gu + hu > (double)c.w - 1.0 || gv + hv > (double)c.h - 1.0) return;
  out[0] = (float)gu; out[1] = (float)gv; out[2] = (float)hu; out[3] = (float)hv;
}

// ---- coverage sweep with many cameras: camera lists per super-tile ---------------------
// The cloud (in Morton order) is cut into <= 1024 super-tiles of S consecutive points; one
// pass tests every camera against every super-tile box and builds, per super-tile, the
// ascending list of cameras that may see it.  The sweep then treats a super-tile as a block
// with that list (typically ~100 of 16 000 cameras), so its 1024-point tiles cull a short,
// shared-memory resident list instead of every camera.
__global__ void __launch_bounds__(256) cov_box_kernel(
    const float* __restrict__ px, const float* __restrict__ py, const float* __restrict__ pz, int64_t n,
    int64_t S, int64_t* __restrict__ pt_off, float* __restrict__ box /* [B][6] centre, half extent */) {
  __shared__ float sm[8][6];
  const int b = blockIdx.x;
  const int64_t p0 = (int64_t)b * S, p1 = (p0 + S < n) ? p0 + S : n;
  const float inf = __int_as_float(0x7f800000);
  float lo[3] = {inf, inf, inf}, hi[3] = {-inf, -inf, -inf};
  for (int64_t i = p0 + threadIdx.x; i < p1; i += 256) {
    const float v[3] = {px[i], py[i], pz[i]};
#pragma unroll
    for (int a = 0; a < 3; ++a) { lo[a] = fminf(lo[a], v[a]); hi[a] = fmaxf(hi[a], v[a]); }
  }
#pragma unroll
  for (int d = 16; d > 0; d >>= 1) {
#pragma unroll
    for (int a = 0; a < 3; ++a) {
      lo[a] = fminf(lo[a], __shfl_xor_sync(kFullMask, lo[a], d));
      hi[a] = fmaxf(hi[a], __shfl_xor_sync(kFullMask, hi[a], d));
    }
  }
  if ((threadIdx.x & 31) == 0) {
#pragma unroll
    for (int a = 0; a < 3; ++a) { sm[threadIdx.x >> 5][a] = lo[a]; sm[threadIdx.x >> 5][3 + a] = hi[a]; }
  }
  __syncthreads();
  if (threadIdx.x < 3) {
    float l = sm[0][threadIdx.x], h = sm[0][3 + threadIdx.x];
    for (int w = 1; w < 8; ++w) { l = fminf(l, sm[w][threadIdx.x]); h = fmaxf(h, sm[w][3 + threadIdx.x]); }
    const float pad = (fabsf(l) + fabsf(h)) * (1.0f / 1048576.0f);
    box[b * 6 + threadIdx.x] = 0.5f * l + 0.5f * h;
    box[b * 6 + 3 + threadIdx.x] = (0.5f * h - 0.5f * l) + pad;   // NaN (all-NaN super-tile) keeps every camera
  }
  if (threadIdx.x == 0) {
    pt_off[b] = p0;
    if (b == (int)gridDim.x - 1) pt_off[b + 1] = n;
  }
}

// FILL = false: counts[b]; FILL = true: cam_idx[cam_off[b] ...] ascending.
template <bool FILL>
__global__ void __launch_bounds__(256) cov_cull_kernel(
    const float* __restrict__ cameras, int32_t cam_first, int32_t cam_count, const float* __restrict__ box,
    int64_t* __restrict__ counts, const int64_t* __restrict__ cam_off, int32_t* __restrict__ cam_idx) {
  __shared__ int sm[33];
  __shared__ float sbox[6];
  const int b = blockIdx.x;
  if (threadIdx.x < 6) sbox[threadIdx.x] = box[b * 6 + threadIdx.x];
  __syncthreads();
  int64_t out = FILL ? cam_off[b] : 0;
  int total = 0;
  for (int s0 = 0; s0 < cam_count; s0 += 256) {
    const int j = s0 + threadIdx.x;
    int f = 0;
    if (j < cam_count) {
      float rec[kCamFloats];
#pragma unroll
      for (int q = 0; q < kCamFloats / 4; ++q)
        reinterpret_cast<float4*>(rec)[q] =
            __ldg(reinterpret_cast<const float4*>(cameras + (int64_t)(cam_first + j) * kCamFloats) + q);
      f = tile_may_hit(load_cam(rec), sbox) ? 1 : 0;
    }
    int tot;
    const int ex = block_exclusive_sum<int, 256>(f, sm, &tot);
    if (FILL && f) cam_idx[out + ex] = cam_first + j;
    out += tot;
    total += tot;
    __syncthreads();
  }
  if (!FILL && threadIdx.x == 0) counts[b] = total;
}

__global__ void __launch_bounds__(1024) cov_offsets_kernel(const int64_t* __restrict__ counts, int B,
                                                           int64_t* __restrict__ cam_off) {
  __shared__ long long sm[33];
  long long carry = 0;
  for (int s = 0; s < B; s += 1024) {
    const int i = s + threadIdx.x;
    const long long v = i < B ? counts[i] : 0;
    long long tot;
    const long long ex = block_exclusive_sum<long long, 1024>(v, sm, &tot);
    if (i < B) cam_off[i] = carry + ex;
    carry += tot;
    __syncthreads();
  }
  if (threadIdx.x == 0) cam_off[B] = carry;
}

__global__ void __launch_bounds__(256) cov_scatter_kernel(
    const int32_t* __restrict__ pair_hits, const int32_t* __restrict__ cam_idx,
    const int64_t* __restrict__ cam_off, int B, int32_t cam_first, int32_t* __restrict__ n_hit) {
  const int64_t P = cam_off[B];
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; p < P; p += stride) {
    const int32_t h = pair_hits[p];
    if (h) atomicAdd(n_hit + (cam_idx[p] - cam_first), h);
  }
}

struct CovPlan {
  int64_t S;        // points per super-tile
  int32_t B;        // super-tiles
  size_t bytes;     // workspace of the super-tile path
};
static CovPlan coverage_plan(int64_t n, int32_t cam_count) {
  CovPlan c;
  int64_t S = (n + kSwMaxBlocks - 1) / kSwMaxBlocks;
  if (S < 16384) S = 16384;
  S = (S + kSwTile - 1) / kSwTile * kSwTile;
  c.S = S;
  c.B = (int32_t)((n + S - 1) / S);
  if (c.B < 1) c.B = 1;
  const size_t pairs = (size_t)c.B * (size_t)(cam_count > 0 ? cam_count : 1);
  c.bytes = align_up((size_t)(c.B + 2) * 8, 256) * 4 + align_up((size_t)c.B * 24, 256) +
            2 * align_up(pairs * 4, 256) + 1024;
  return c;
}

// ---- host side -----------------------------------------------------------------
template <int MODE>
static int launch_sweep(const SweepArgs& a, cudaStream_t st) {
  using SweepSmem = SweepSmemT<MODE>;
  static_assert(sizeof(SweepSmem) < 112 * 1024, "two CTAs per SM");
  PTGS_CUDA_TRY(cudaFuncSetAttribute(sweep_kernel<MODE>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                     (int)sizeof(SweepSmem)));
  // `prefix` is the last member: only the entries of this launch's blocks are allocated, which is
  // what lets kSwCtasPerSM CTAs fit one SM for the usual block counts
  const size_t smem = offsetof(SweepSmem, prefix) + (size_t)(a.n_blocks + 1) * sizeof(int64_t);
  sweep_kernel<MODE><<<kNumSMs * kSwCtasPerSM, kSwThreads, smem, st>>>(a);
  PTGS_LAUNCH_CHECK();
  return PTGS_OK;
}

static int schedule(const int64_t* pt_offsets, const int64_t* cam_offsets, int32_t L,
                    int32_t cam_split, int64_t* item_prefix, cudaStream_t st) {
  sweep_schedule_kernel<<<1, 1024, 0, st>>>(pt_offsets, cam_offsets, L, cam_split, item_prefix,
                                            kNumSMs * kSwCtasPerSM);
  PTGS_LAUNCH_CHECK();
  return PTGS_OK;
}

static bool aligned16(const void* p) { return ((uintptr_t)p & 15) == 0; }

}  // namespace ptgs

using namespace ptgs;

extern "C" {

size_t ptgs_visibility_workspace_bytes(int32_t n_blocks) {
  if (n_blocks < 1) n_blocks = 1;
  return align_up((size_t)(n_blocks + 1) * 8, 256) + 256;
}

int ptgs_visibility_init(int64_t n_pairs, int32_t* n_hit, uint64_t* pivot, uint64_t* extreme,
                         int32_t* cand_count, ptgs_stream_t stream) {
  PTGS_REQUIRE(n_pairs >= 0, "n_pairs");
  if (n_pairs == 0) return PTGS_OK;
  int64_t g = (n_pairs * kDirs + 255) / 256;
  if (g > kNumSMs * 8) g = kNumSMs * 8;
  vis_init_kernel<<<(int)g, 256, 0, (cudaStream_t)stream>>>(n_pairs, n_hit, pivot, extreme, cand_count);
  PTGS_LAUNCH_CHECK();
  return PTGS_OK;
}

int ptgs_visibility_count(const float* px, const float* py, const float* pz,
                          const int64_t* pt_offsets, int32_t n_blocks, const float* cameras,
                          const int64_t* cam_offsets, const int32_t* cam_index, int32_t cam_split,
                          int32_t* n_hit, uint64_t* pivot, uint64_t* extreme, const float* krect,
                          void* workspace, size_t workspace_bytes, ptgs_stream_t stream) {
  PTGS_REQUIRE(n_blocks >= 1 && n_blocks <= kSwMaxBlocks, "n_blocks must be in [1,1024]");
  PTGS_REQUIRE(pt_offsets && cameras && cam_offsets && cam_index && n_hit && pivot && extreme,
               "arguments");
  PTGS_REQUIRE(aligned16(px) && aligned16(py) && aligned16(pz) && aligned16(cameras),
               "point / camera arrays must be 16-byte aligned");
  PTGS_REQUIRE(cam_split >= 1, "cam_split");
  if (workspace_bytes < ptgs_visibility_workspace_bytes(n_blocks)) {
    set_error("visibility_count: workspace too small");
    return PTGS_E_WORKSPACE;
  }
  cudaStream_t st = (cudaStream_t)stream;
  int64_t* item_prefix = (int64_t*)workspace;
  int rc = schedule(pt_offsets, cam_offsets, n_blocks, cam_split, item_prefix, st);
  if (rc) return rc;
  SweepArgs a = {};
  a.px = px; a.py = py; a.pz = pz; a.pt_offsets = pt_offsets; a.n_blocks = n_blocks;
  a.cameras = cameras; a.cam_offsets = cam_offsets; a.cam_index = cam_index;
  a.cam_split = cam_split; a.item_prefix = item_prefix;
  a.n_hit = n_hit; a.pivot = pivot; a.extreme = extreme; a.krect = krect;
  return launch_sweep<1>(a, st);
}

int ptgs_visibility_sample(const float* px, const float* py, const float* pz,
                           const int64_t* pt_offsets, int32_t n_blocks, int32_t stride, float* sx,
                           float* sy, float* sz, int64_t* s_offsets, int64_t capacity,
                           ptgs_stream_t stream) {
  PTGS_REQUIRE(n_blocks >= 1 && n_blocks <= kSwMaxBlocks, "n_blocks must be in [1,1024]");
  PTGS_REQUIRE(stride >= 1 && capacity >= 0 && pt_offsets && s_offsets, "arguments");
  cudaStream_t st = (cudaStream_t)stream;
  sample_offsets_kernel<<<1, 1024, 0, st>>>(pt_offsets, n_blocks, stride, s_offsets);
  PTGS_LAUNCH_CHECK();
  if (capacity == 0) return PTGS_OK;
  PTGS_REQUIRE(px && py && pz && sx && sy && sz, "arguments");
  int64_t g = (capacity + 255) / 256;
  if (g > kNumSMs * 8) g = kNumSMs * 8;
  sample_gather_kernel<<<(int)g, 256, 0, st>>>(px, py, pz, pt_offsets, s_offsets, n_blocks, stride,
                                               capacity, sx, sy, sz);
  PTGS_LAUNCH_CHECK();
  return PTGS_OK;
}

int ptgs_visibility_krect(const float* sx, const float* sy, const float* sz,
                          const int64_t* s_offsets, int32_t n_blocks, const float* cameras,
                          const int64_t* cam_offsets, const int32_t* cam_index,
                          const int32_t* n_hit_s, const uint64_t* extreme_s, int64_t n_pairs,
                          double margin_px, float* krect, ptgs_stream_t stream) {
  PTGS_REQUIRE(n_blocks >= 1 && n_blocks <= kSwMaxBlocks, "n_blocks must be in [1,1024]");
  PTGS_REQUIRE(n_pairs >= 0 && margin_px >= 0.5, "n_pairs / margin_px (>= 0.5 pixel)");
  if (n_pairs == 0) return PTGS_OK;
  PTGS_REQUIRE(s_offsets && cameras && cam_offsets && cam_index && n_hit_s && extreme_s && krect,
               "arguments");
  krect_kernel<<<(unsigned)((n_pairs + 127) / 128), 128, 0, (cudaStream_t)stream>>>(
      sx, sy, sz, s_offsets, n_blocks, cameras, cam_offsets, cam_index, n_hit_s, extreme_s,
      margin_px, krect);
  PTGS_LAUNCH_CHECK();
  return PTGS_OK;
}

int ptgs_visibility_sample_count(const float* sx, const float* sy, const float* sz,
                                 const int64_t* s_offsets, int32_t n_blocks, const float* cameras,
                                 const int64_t* cam_offsets, const int32_t* cam_index, int64_t n_pairs,
                                 int32_t* n_hit_s, uint64_t* pivot_s, uint64_t* extreme_s,
                                 ptgs_stream_t stream) {
  PTGS_REQUIRE(n_blocks >= 1 && n_blocks <= kSwMaxBlocks, "n_blocks must be in [1,1024]");
  PTGS_REQUIRE(n_pairs >= 0 && n_pairs < (1ll << 31), "n_pairs");
  if (n_pairs == 0) return PTGS_OK;
  PTGS_REQUIRE(s_offsets && cameras && cam_offsets && cam_index && n_hit_s && extreme_s, "arguments");
  PTGS_REQUIRE(aligned16(cameras), "camera records must be 16-byte aligned");
  sample_pairs_kernel<<<(unsigned)((n_pairs + 7) / 8), 256, 0, (cudaStream_t)stream>>>(
      sx, sy, sz, s_offsets, n_blocks, cameras, cam_offsets, cam_index, n_hit_s, pivot_s, extreme_s);
  PTGS_LAUNCH_CHECK();
  return PTGS_OK;
}

int ptgs_visibility_bounds(const float* px, const float* py, const float* pz,
                           const int64_t* pt_offsets, int32_t n_blocks, const float* cameras,
                           const int64_t* cam_offsets, const int32_t* cam_index,
                           const int32_t* n_hit, const uint64_t* pivot, const uint64_t* extreme,
                           int64_t n_pairs, int32_t min_hits, double ratio_threshold,
                           double near_eps, int32_t exact_all, float* poly, double* ratio_lo,
                           double* ratio_hi, uint8_t* need, double* area, double* ratio,
                           uint8_t* keep, uint8_t* flags, ptgs_stream_t stream) {
  PTGS_REQUIRE(n_blocks >= 1 && n_blocks <= kSwMaxBlocks, "n_blocks must be in [1,1024]");
  PTGS_REQUIRE(n_pairs >= 0, "n_pairs");
  if (n_pairs == 0) return PTGS_OK;
  PTGS_REQUIRE(pt_offsets && cameras && cam_offsets && cam_index && n_hit && pivot && extreme && poly &&
                   ratio_lo && ratio_hi && need && area && ratio && keep && flags,
               "arguments");
  bounds_kernel<<<(unsigned)((n_pairs + 127) / 128), 128, 0, (cudaStream_t)stream>>>(
      px, py, pz, pt_offsets, n_blocks, cameras, cam_offsets, cam_index, n_hit, pivot, extreme,
      min_hits, ratio_threshold, near_eps, exact_all, poly, ratio_lo, ratio_hi, need, area, ratio,
      keep, flags);
  PTGS_LAUNCH_CHECK();
  return PTGS_OK;
}

int ptgs_visibility_plan(const int32_t* n_hit, const uint8_t* need, const int32_t* caps_in,
                         const float* poly, const double* ratio_lo, const double* ratio_hi,
                         int64_t n_pairs, int32_t cap_max,
                         int64_t pool_capacity,
                         int64_t* cand_offsets, ptgs_stream_t stream) {
  PTGS_REQUIRE(n_pairs >= 0 && n_hit && need && cand_offsets && cap_max >= 1 && pool_capacity >= 0,
               "arguments");
  cand_plan_kernel<<<1, 1024, 0, (cudaStream_t)stream>>>(n_hit, need, caps_in, poly, ratio_lo, ratio_hi,
                                                       n_pairs, cap_max, pool_capacity, cand_offsets);
  PTGS_LAUNCH_CHECK();
  return PTGS_OK;
}

int ptgs_visibility_candidates(const float* px, const float* py, const float* pz,
                               const int64_t* pt_offsets, int32_t n_blocks, const float* cameras,
                               const int64_t* cam_offsets, const int32_t* cam_index,
                               int32_t cam_split, const int32_t* pair_map, const float* poly,
                               const int64_t* cand_offsets, float* cand_uv, int32_t* cand_count,
                               void* workspace, size_t workspace_bytes, ptgs_stream_t stream) {
  PTGS_REQUIRE(n_blocks >= 1 && n_blocks <= kSwMaxBlocks, "n_blocks must be in [1,1024]");
  PTGS_REQUIRE(pt_offsets && cameras && cam_offsets && cam_index && poly && cand_offsets && cand_count,
               "arguments");
  PTGS_REQUIRE(aligned16(px) && aligned16(py) && aligned16(pz) && aligned16(cameras) && aligned16(poly),
               "point / camera / polygon arrays must be 16-byte aligned");
  PTGS_REQUIRE(((uintptr_t)cand_uv & 7) == 0, "cand_uv must be 8-byte aligned");
  PTGS_REQUIRE(cam_split >= 1, "cam_split");
  if (workspace_bytes < ptgs_visibility_workspace_bytes(n_blocks)) {
    set_error("visibility_candidates: workspace too small");
    return PTGS_E_WORKSPACE;
  }
  cudaStream_t st = (cudaStream_t)stream;
  int64_t* item_prefix = (int64_t*)workspace;
  const char* mode = getenv("PTGS_SWEEP2");
  if (!(mode && strcmp(mode, "tiles") == 0)) {
    // one warp per 128-point run of the blocks that own an undecided pair (cand_runs_kernel)
    cand_schedule_kernel<<<1, 1024, 0, st>>>(pt_offsets, cam_offsets, n_blocks, item_prefix);
    PTGS_LAUNCH_CHECK();
    cand_runs_kernel<<<kNumSMs * 16, kCrWarps * 32, 0, st>>>(px, py, pz, pt_offsets, n_blocks, cameras, cam_offsets,
                                                            cam_index, pair_map, poly, cand_offsets,
                                                            (float2*)cand_uv, cand_count, item_prefix);
    PTGS_LAUNCH_CHECK();
    return PTGS_OK;
  }
  // PTGS_SWEEP2=tiles: the tile-staged sweep (A/B runs; same candidate sets)
  int rc = schedule(pt_offsets, cam_offsets, n_blocks, cam_split, item_prefix, st);
  if (rc) return rc;
  SweepArgs a = {};
  a.px = px; a.py = py; a.pz = pz; a.pt_offsets = pt_offsets; a.n_blocks = n_blocks;
  a.cameras = cameras; a.cam_offsets = cam_offsets; a.cam_index = cam_index;
  a.cam_split = cam_split; a.item_prefix = item_prefix;
  a.pair_map = pair_map; a.poly = poly; a.cand_offsets = cand_offsets; a.cand_uv = (float2*)cand_uv; a.cand_count = cand_count;
  return launch_sweep<2>(a, st);
}

namespace ptgs {
struct PfWs {
  int64_t* chunk_prefix; int64_t* off2; int32_t* count2; unsigned long long* ext; float* poly; float2* uv2;
  int64_t uv2_cap; size_t bytes;
};
static PfWs carve_pf(void* ws, int64_t P) {
  PfWs w;
  if (P < 1) P = 1;
  w.uv2_cap = P * 2048;
  if (w.uv2_cap < (1ll << 20)) w.uv2_cap = 1ll << 20;
  if (w.uv2_cap > (16ll << 20)) w.uv2_cap = 16ll << 20;
  Carver cv(ws);
  w.chunk_prefix = cv.take<int64_t>(P + 1);
  w.off2 = cv.take<int64_t>(P);
  w.count2 = cv.take<int32_t>(P);
  w.ext = cv.take<unsigned long long>((size_t)P * kFiltDirs);
  w.poly = cv.take<float>((size_t)P * kPfPolyFloats);
  w.uv2 = cv.take<float2>((size_t)w.uv2_cap);
  w.bytes = cv.used() + 256;
  return w;
}
}  // namespace ptgs

size_t ptgs_visibility_hull_workspace_bytes(int64_t n_pairs) { return carve_pf(nullptr, n_pairs).bytes; }

int ptgs_visibility_hull(float* cand_uv, const int64_t* cand_offsets,
                         const int32_t* cand_count, const int32_t* n_hit, int64_t n_pairs,
                         const int32_t* pair_cam, const float* cameras, int32_t min_hits,
                         double ratio_threshold, double near_eps, double* area, double* ratio,
                         uint8_t* keep, uint8_t* flags, const int32_t* pair_map,
                         const int64_t* n_sub_dev, int32_t* status, void* workspace,
                         size_t workspace_bytes, ptgs_stream_t stream) {
  PTGS_REQUIRE(n_pairs >= 0 && n_pairs < (1ll << 31), "n_pairs");
  if (n_pairs == 0) return PTGS_OK;
  PTGS_REQUIRE(cand_offsets && cand_count && n_hit && pair_cam && cameras && area && ratio && keep && flags,
               "arguments");
  cudaStream_t st = (cudaStream_t)stream;
  if (status) PTGS_CUDA_TRY(cudaMemsetAsync(status, 0, 8, st));
  const int32_t* pf_count2 = nullptr;
  const int64_t* pf_off2 = nullptr;
  float2* pf_uv2 = nullptr;
  if (workspace && workspace_bytes >= ptgs_visibility_hull_workspace_bytes(n_pairs)) {
    // multi-CTA pre-filter of the pairs with the longest candidate lists (see pf_* above)
    const PfWs w = carve_pf(workspace, n_pairs);
    PTGS_CUDA_TRY(cudaMemsetAsync(w.ext, 0, (size_t)n_pairs * kFiltDirs * 8, st));
    pf_plan_kernel<<<1, 1024, 0, st>>>(cand_offsets, cand_count, pair_map, n_sub_dev, n_pairs, w.uv2_cap,
                                       w.chunk_prefix, w.off2, w.count2);
    PTGS_LAUNCH_CHECK();
    pf_extremes_kernel<<<kNumSMs * 4, 256, 0, st>>>((const float2*)cand_uv, cand_offsets, cand_count, pair_map,
                                                    n_pairs, w.chunk_prefix, w.ext);
    PTGS_LAUNCH_CHECK();
    pf_polygon_kernel<<<(unsigned)((n_pairs + 3) / 4), 128, 0, st>>>(
        (const float2*)cand_uv, cand_offsets, pair_map, n_pairs, w.count2, w.ext, w.poly);
    PTGS_LAUNCH_CHECK();
    pf_filter_kernel<<<kNumSMs * 4, 256, 0, st>>>((const float2*)cand_uv, cand_offsets, cand_count, pair_map,
                                                  n_pairs, pair_cam, cameras, w.chunk_prefix, w.poly, w.off2,
                                                  w.count2, w.uv2);
    PTGS_LAUNCH_CHECK();
    pf_count2 = w.count2; pf_off2 = w.off2; pf_uv2 = w.uv2;
  }
  // the number of undecided pairs lives on the device; n_pairs / 8 is a typical share
  const char* e = getenv("PTGS_HULL_THREADS");
  const int nt = e ? atoi(e) : (n_pairs >= 8 * 4 * kNumSMs ? kHullThreadsMany : 512);
#define PTGS_HULL_LAUNCH(NT)                                                                                  \
  do {                                                                                                        \
    PTGS_CUDA_TRY(cudaFuncSetAttribute(hull_kernel<NT>, cudaFuncAttributeMaxDynamicSharedMemorySize,          \
                                       kHullDynSmem));                                                        \
    hull_kernel<NT><<<(unsigned)n_pairs, NT, kHullDynSmem, st>>>(                                             \
        (float2*)cand_uv, cand_offsets, cand_count, n_hit, n_pairs, pair_cam, cameras, min_hits,              \
        ratio_threshold, near_eps, area, ratio, keep, flags, pair_map, n_sub_dev, status, pf_count2, pf_off2, \
        pf_uv2);                                                                                              \
  } while (0)
  if (nt == 128) PTGS_HULL_LAUNCH(128);
  else if (nt == 256) PTGS_HULL_LAUNCH(256);
  else PTGS_HULL_LAUNCH(512);
#undef PTGS_HULL_LAUNCH
  PTGS_LAUNCH_CHECK();
  return PTGS_OK;
}

int ptgs_debug_sweep_stats(uint64_t* out4, int32_t reset) {
#ifdef PTGS_SWEEP_STATS_FINE
  unsigned long long h[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  PTGS_CUDA_TRY(cudaDeviceSynchronize());
  if (out4) {
    PTGS_CUDA_TRY(cudaMemcpyFromSymbol(h, g_sweep_stats, sizeof(h)));
    for (int i = 0; i < 8; ++i) out4[i] = h[i];
  }
  if (reset) {
    unsigned long long z[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    PTGS_CUDA_TRY(cudaMemcpyToSymbol(g_sweep_stats, z, sizeof(z)));
  }
  return PTGS_OK;
#else
  (void)reset;
  if (out4)
    for (int i = 0; i < 8; ++i) out4[i] = 0;
  set_error("sweep statistics are compiled out (build with -DPTGS_SWEEP_STATS_FINE)");
  return PTGS_E_INVALID;
#endif
}

// FP32 FMA issue-rate probe: 8 independent chains per thread, no memory traffic
// inside the loop.  Diagnostics for the roofline report (bench.py).
namespace ptgs {
__global__ void __launch_bounds__(256) fma_probe_kernel(float* out, int iters) {
  float a[8];
#pragma unroll
  for (int k = 0; k < 8; ++k) a[k] = 1.0f + 1e-3f * (float)(threadIdx.x + k);
  const float m = 1.0f - 1e-7f * (float)(blockIdx.x & 3), c = 1e-6f;
  for (int i = 0; i < iters; ++i) {
#pragma unroll
    for (int k = 0; k < 8; ++k) a[k] = __fmaf_rn(a[k], m, c);
  }
  float s = 0.0f;
#pragma unroll
  for (int k = 0; k < 8; ++k) s += a[k];
  out[(size_t)blockIdx.x * blockDim.x + threadIdx.x] = s;
}
}  // namespace ptgs

int ptgs_debug_fma_probe(float* out, int32_t iters, ptgs_stream_t stream) {
  PTGS_REQUIRE(out && iters > 0, "arguments");
  fma_probe_kernel<<<kNumSMs * 8, 256, 0, (cudaStream_t)stream>>>(out, iters);
  PTGS_LAUNCH_CHECK();
  return PTGS_OK;
}

int ptgs_coverage_sweep(const float* px, const float* py, const float* pz, int64_t n,
                        const float* cameras, int32_t cam_first, int32_t cam_count, int32_t* n_hit,
                        uint8_t* seen, void* workspace, size_t workspace_bytes,
                        ptgs_stream_t stream) {
  PTGS_REQUIRE(n >= 0 && cam_first >= 0 && cam_count >= 0 && n_hit, "arguments");
  PTGS_REQUIRE(aligned16(px) && aligned16(py) && aligned16(pz) && aligned16(cameras),
               "point / camera arrays must be 16-byte aligned");
  cudaStream_t st = (cudaStream_t)stream;
  if (cam_count > 0) PTGS_CUDA_TRY(cudaMemsetAsync(n_hit, 0, sizeof(int32_t) * cam_count, st));
  if (seen && n > 0) PTGS_CUDA_TRY(cudaMemsetAsync(seen, 0, (size_t)n, st));
  if (n == 0 || cam_count == 0) return PTGS_OK;
  if (!workspace || workspace_bytes < ptgs_visibility_workspace_bytes(1)) {
    set_error("coverage_sweep: workspace too small");
    return PTGS_E_WORKSPACE;
  }
  const CovPlan plan = coverage_plan(n, cam_count);
  if (cam_count > 2 * kSwCamChunk && workspace_bytes >= plan.bytes) {
    // many cameras: per super-tile camera lists first (see cov_*_kernel)
    Carver cv(workspace);
    const int B = plan.B;
    int64_t* pt_off = cv.take<int64_t>(B + 2);
    int64_t* counts = cv.take<int64_t>(B + 2);
    int64_t* cam_off = cv.take<int64_t>(B + 2);
    int64_t* item_prefix = cv.take<int64_t>(B + 2);
    float* box = cv.take<float>((size_t)B * 6);
    int32_t* cam_idx = cv.take<int32_t>((size_t)B * cam_count);
    int32_t* pair_hits = cv.take<int32_t>((size_t)B * cam_count);
    cov_box_kernel<<<B, 256, 0, st>>>(px, py, pz, n, plan.S, pt_off, box);
    PTGS_LAUNCH_CHECK();
    cov_cull_kernel<false><<<B, 256, 0, st>>>(cameras, cam_first, cam_count, box, counts, nullptr, nullptr);
    PTGS_LAUNCH_CHECK();
    cov_offsets_kernel<<<1, 1024, 0, st>>>(counts, B, cam_off);
    PTGS_LAUNCH_CHECK();
    cov_cull_kernel<true><<<B, 256, 0, st>>>(cameras, cam_first, cam_count, box, nullptr, cam_off, cam_idx);
    PTGS_LAUNCH_CHECK();
    PTGS_CUDA_TRY(cudaMemsetAsync(pair_hits, 0, (size_t)B * cam_count * 4, st));
    int rc = schedule(pt_off, cam_off, B, 1, item_prefix, st);
    if (rc) return rc;
    SweepArgs a = {};
    a.px = px; a.py = py; a.pz = pz; a.pt_offsets = pt_off; a.n_blocks = B;
    a.cameras = cameras; a.cam_offsets = cam_off; a.cam_index = cam_idx; a.cam_first = 0;
    a.cam_split = 1; a.item_prefix = item_prefix;
    a.n_hit = pair_hits; a.seen = seen;
    rc = launch_sweep<0>(a, st);
    if (rc) return rc;
    cov_scatter_kernel<<<kNumSMs * 4, 256, 0, st>>>(pair_hits, cam_idx, cam_off, B, cam_first, n_hit);
    PTGS_LAUNCH_CHECK();
    return PTGS_OK;
  }
  int64_t* ws = (int64_t*)workspace;
  coverage_setup_kernel<<<1, 1, 0, st>>>(n, cam_count, ws, ws + 2);
  PTGS_LAUNCH_CHECK();
  // few tiles: split the camera list so that every SM has work
  const int64_t tiles = (n + kSwTile - 1) / kSwTile;
  int64_t split = 1;
  if (tiles < 8 * kNumSMs) split = (8 * kNumSMs + tiles - 1) / tiles;
  if (split > cam_count) split = cam_count;
  int rc = schedule(ws, ws + 2, 1, (int32_t)split, ws + 4, st);
  if (rc) return rc;
  SweepArgs a = {};
  a.px = px; a.py = py; a.pz = pz; a.pt_offsets = ws; a.n_blocks = 1;
  a.cameras = cameras; a.cam_offsets = ws + 2; a.cam_index = nullptr; a.cam_first = cam_first;
  a.cam_split = (int32_t)split; a.item_prefix = ws + 4;
  a.n_hit = n_hit; a.seen = seen;
  return launch_sweep<0>(a, st);
}

size_t ptgs_coverage_workspace_bytes(int64_t n, int32_t cam_count) {
  const size_t direct = ptgs_visibility_workspace_bytes(1);
  if (n <= 0 || cam_count <= 2 * kSwCamChunk) return direct;
  const size_t b = coverage_plan(n, cam_count).bytes;
  return b > direct ? b : direct;
}

}  // extern "C"
