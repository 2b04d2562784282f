/*
 * libptgs -- B200 (sm_100a) point-count scene partitioner, C ABI.
 *
 * Drop-in boundary for the hot path of 92401/Based-on-point-cloud-partitions
 * (the reference has no FFI: its boundary is the set of Python callables
 * `run_DataPartition` invokes, scene/ptgs/shen_partition_utils.py:57-93).
 * Each entry point below names the reference interface it replaces.
 *
 * Conventions
 *   - every pointer is a DEVICE pointer into memory owned by the caller unless
 *     the parameter is documented as "host"; the library never frees or
 *     retains caller memory.
 *   - sizes are int64_t; the last argument is the CUDA stream (cudaStream_t)
 *     on which all work is enqueued; calls are asynchronous unless documented
 *     "synchronises".
 *   - scratch memory is caller provided: ask ptgs_<op>_workspace_bytes() first.
 *   - return value: PTGS_OK (0) or a negative ptgs_status; the message of the
 *     last failure on the calling thread is ptgs_last_error().
 *   - variable-length outputs: caller allocates capacity, the library writes a
 *     device-side count; PTGS_E_CAPACITY is reported through the device-side
 *     status word (see each call) because the overflow is detected on device.
 *   - no global mutable state on the device; the only host state is thread
 *     local (the last error string and the launch counter of the calling
 *     thread): re-entrant across streams, devices and threads.
 *   - arithmetic: every kernel is compiled with -fmad=false and IEEE division;
 *     fused multiply-adds appear only where this header says "fma".
 */
#ifndef PTGS_H_
#define PTGS_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef void* ptgs_stream_t; /* cudaStream_t */

typedef enum ptgs_status {
  PTGS_OK = 0,
  PTGS_E_INVALID = -1,   /* bad argument */
  PTGS_E_CUDA = -2,      /* CUDA runtime error (see ptgs_last_error) */
  PTGS_E_CAPACITY = -3,  /* caller-provided capacity too small */
  PTGS_E_RANGE = -4,     /* value does not fit the packed key layout */
  PTGS_E_WORKSPACE = -5  /* workspace smaller than *_workspace_bytes() */
} ptgs_status;

/* Coordinate type selector: the reference runs the split / expansion /
 * down-sampling on float32 point clouds, and on float64 ones when a Manhattan
 * transform is applied (scene/ptgs/shen_data_read.py:126-130). */
typedef enum ptgs_dtype { PTGS_F32 = 0, PTGS_F64 = 1 } ptgs_dtype;

int ptgs_version(void);
/* PTGS_EXTREME_DIRS the library was built with (sizes of `extreme` / `poly`). */
int ptgs_extreme_dirs(void);
const char* ptgs_last_error(void);
/* Number of kernel launches issued by this library on the calling thread
 * since the last ptgs_reset_launch_count() (bench.py `gpu_launches`). */
int64_t ptgs_launch_count(void);
void ptgs_reset_launch_count(void);

/* ------------------------------------------------------------------ *
 * Layout helpers
 * ------------------------------------------------------------------ */

/* (N,3) row-major float32 -> three planar arrays (vectorised float4 loads).
 * Replaces the implicit column slicing `points[:, 0]` etc. everywhere in the
 * reference (e.g. density_partition.py:22-23). */
int ptgs_aos_to_soa_f32(const float* xyz, int64_t n, float* x, float* y,
                        float* z, ptgs_stream_t stream);

/* Same for (N,3) float64 rows (the cloud after a Manhattan transform,
 * scene/ptgs/shen_data_read.py:126-130, is float64). */
int ptgs_aos_to_soa_f64(const double* xyz, int64_t n, double* x, double* y,
                        double* z, ptgs_stream_t stream);

/* A13 -- Manhattan transform of the point cloud,
 * scene/ptgs/shen_data_read.py:126-130 (fetchPly):
 *     new = R(3x3,f64) . p(f32->f64) + T        (f64 result)
 * evaluated as t = r0*x; t = fma(r1,y,t); t = fma(r2,z,t); out = t + T
 * (the k-ordered FMA accumulation of the BLAS dgemm micro-kernel NumPy calls).
 * `man_trans` is a HOST pointer to the row-major 4x4 matrix. */
int ptgs_manhattan_apply(const float* x, const float* y, const float* z,
                         int64_t n, const double* man_trans_host, double* ox,
                         double* oy, double* oz, ptgs_stream_t stream);

/* Layout + A13 in one pass: (N,3) row-major float32 -> planar float64 with the
 * Manhattan transform of ptgs_manhattan_apply applied on the way (identical
 * arithmetic, 36 instead of 60 bytes of traffic per point).  16-byte aligned
 * pointers; `man_trans` is a HOST pointer to the row-major 4x4 matrix. */
int ptgs_aos_to_soa_manhattan(const float* xyz, int64_t n,
                              const double* man_trans_host, double* ox,
                              double* oy, double* oz, ptgs_stream_t stream);

/* ------------------------------------------------------------------ *
 * A1 -- radius outlier removal
 *   replaces o3d PointCloud.remove_radius_outlier(nb_points, radius) as
 *   called at scene/ptgs/shen_partition_utils.py:50 (third-party open3d,
 *   un-pinned): keep point i iff #{j : |pi-pj|^2 < radius^2, j incl. i}
 *   > nb_points, distances in float64.
 *   out keep[n] (uint8) and *n_keep (device int64).
 *   xyz_aos (may be NULL): the same cloud as (N,3) rows of the same dtype;
 *   the gather of the points into grid order then reads one or two sectors
 *   per point instead of three.  Synchronises once (the grid dimensions size
 *   the sort and the cell table).
 * ------------------------------------------------------------------ */
size_t ptgs_radius_outlier_workspace_bytes(int64_t n);
int ptgs_radius_outlier_mask(int dtype, const void* x, const void* y,
                             const void* z, const void* xyz_aos, int64_t n,
                             double radius,
                             int nb_points, uint8_t* keep, int64_t* n_keep,
                             void* workspace, size_t workspace_bytes,
                             ptgs_stream_t stream);

/* Multi-GPU A1 (SURVEY.md 8e names only the visibility pass for sharding; the
 * radius filter is the largest replicated stage, and it shards exactly: a
 * point's neighbours lie within `radius` of it).  Rank r owns the X slab
 * [lo, hi):
 * ptgs_slab_select: compacts the planes and cloud indices of the points with
 *   lo_halo <= x < hi_halo (caller passes lo - halo, hi + halo; index order
 *   kept); out_* have capacity n, *n_out (device int64) receives the count.
 * ptgs_slab_keep: keep[slab_index[i]] = slab_keep[i] for the slab points
 *   whose own x lies in [lo, hi) (the halo points belong to the neighbours);
 *   `keep` (cloud sized) must be zeroed by the caller, the ranks' masks are
 *   then combined with one MAX all-reduce. */
size_t ptgs_slab_select_workspace_bytes(int64_t n);
int ptgs_slab_select(int dtype, const void* x, const void* y, const void* z,
                     int64_t n, double lo, double hi, void* out_x, void* out_y,
                     void* out_z, int32_t* out_index, int64_t* n_out,
                     void* workspace, size_t workspace_bytes,
                     ptgs_stream_t stream);
int ptgs_slab_keep(int dtype, const void* slab_x, const int32_t* slab_index,
                   const uint8_t* slab_keep, int64_t m, double lo, double hi,
                   uint8_t* keep, ptgs_stream_t stream);

/* Stream compaction of the kept points (index order preserved), i.e.
 * `pcd.select_by_index(ind)`; only x and y are needed by the split.
 * out_x/out_y have capacity n; *n_out receives the count. */
size_t ptgs_compact_xy_workspace_bytes(int64_t n);
int ptgs_compact_xy(int dtype, const void* x, const void* y,
                    const uint8_t* keep, int64_t n, void* out_x, void* out_y,
                    int64_t* n_out, void* workspace, size_t workspace_bytes,
                    ptgs_stream_t stream);

/* Allocation-free variant of ptgs_compact_xy.  It has NO counterpart in the
 * reference: it was added for callers
 * that cannot call cudaMalloc (CUDA graph capture, tight serving loops,
 * harnesses that forbid it).  Same inputs, outputs and results as
 * ptgs_compact_xy, but the per-tile scratch lives in a file-scope
 * `static __device__` array of the library (384 KB, held by every GPU that
 * loads the library until exit) instead of a caller workspace; the call
 * performs no cudaMalloc / cudaFree and calls no allocating library routine.
 *   Limit: n <= PTGS_COMPACT_XY_STATIC_MAX_POINTS; a larger n is rejected
 *   with PTGS_E_WORKSPACE (ptgs_last_error names the limit) before anything
 *   is launched -- use ptgs_compact_xy with a caller workspace then.
 *   The array is SHARED by every call on the device and keeps the last
 *   call's contents (nothing in it has to start at zero: each call rewrites
 *   the entries it reads).  Calls that can overlap on the GPU -- different
 *   streams, or host threads on different streams -- race on it: serialise
 *   them (one stream, or events).  The first call on each device looks up the
 *   array's device address (cudaGetSymbolAddress, with that device current). */
#define PTGS_COMPACT_XY_STATIC_MAX_POINTS (1ll << 26)
int ptgs_compact_xy_static(int dtype, const void* x, const void* y,
                           const uint8_t* keep, int64_t n, void* out_x,
                           void* out_y, int64_t* n_out, ptgs_stream_t stream);

/* ------------------------------------------------------------------ *
 * A2 -- XY bounds, scene/ptgs/shen_partition_utils.py:66-70.
 *   n is a DEVICE count (so it can be chained after ptgs_compact_xy).
 *   bounds[4] = {xmin, ymin, xmax, ymax} as float64.
 * ------------------------------------------------------------------ */
int ptgs_bounds_xy(int dtype, const void* x, const void* y,
                   const int64_t* n_dev, int64_t n_capacity, double* bounds,
                   ptgs_stream_t stream);

/* ------------------------------------------------------------------ *
 * A3 -- balanced binary (exact median) split
 *   replaces quad_tree_partition (shen_partation_densify.py:37-49) ->
 *   density_partition (density_partition.py:148-166) ->
 *   balanced_binary_partition (density_partition.py:9-99).
 *   Leaf when n <= leaf_limit (= threshold*1.2 evaluated by the caller in
 *   float64) or depth >= max_depth; axis = x when x_range >= y_range; split
 *   value = the element of rank n/2 along the axis (exact radix select);
 *   both children are closed rectangles that share the split line.
 *   Outputs (DFS, left first): leaf_bounds[L][4] f64 {minx,miny,maxx,maxy},
 *   leaf_path[L] (bit i of the path = branch taken at depth i, MSB first in
 *   a `depth`-bit field), leaf_depth[L], leaf_count[L]; *n_leaves (device).
 *   status_dev: device int32, PTGS_OK or PTGS_E_CAPACITY when the working
 *   buffers (sized by capacity_factor) overflowed (duplicate split-line
 *   points are copied into both children).
 *   max_leaves must be >= 2^max_depth.
 *   The level loop reads back one int per level (number of nodes that still
 *   split) and stops at the first level without any: the call synchronises
 *   the stream once per executed level.
 * ------------------------------------------------------------------ */
size_t ptgs_balanced_split_workspace_bytes(int dtype, int64_t n_capacity,
                                           double capacity_factor,
                                           int max_depth);
int ptgs_balanced_split(int dtype, const void* x, const void* y,
                        const int64_t* n_dev, int64_t n_capacity,
                        const double* root_bounds, double leaf_limit,
                        int max_depth, double capacity_factor,
                        double* leaf_bounds, uint32_t* leaf_path,
                        int32_t* leaf_depth, int64_t* leaf_count,
                        int32_t* n_leaves, int32_t max_leaves,
                        int32_t* status_dev, void* workspace,
                        size_t workspace_bytes, ptgs_stream_t stream);

/* ------------------------------------------------------------------ *
 * A4 -- expanded-block point selection
 *   replaces expand_partitions/extract_point_cloud
 *   (shen_partation_densify.py:182-252, mask at :201-202): closed
 *   rectangle test of every (unfiltered) point against every block's
 *   expanded box.  rects[L][4] = {minx, maxx, miny, maxy} in the point
 *   dtype (the caller rounds the float64 bounds to float32 exactly as
 *   NumPy does when comparing a float32 array with a Python float).
 *   Two calls: _count fills counts[L] (int64) and the workspace, the caller
 *   reads counts, allocates `indices` (int32, sum(counts)) and calls _fill,
 *   which writes each block's indices in ascending order
 *   (CSR offsets = exclusive scan of counts, written to offsets[L+1]).
 * ------------------------------------------------------------------ */
size_t ptgs_expand_workspace_bytes(int64_t n, int32_t n_blocks);
int ptgs_expand_count(int dtype, const void* x, const void* y, int64_t n,
                      const void* rects, int32_t n_blocks, int64_t* counts,
                      void* workspace, size_t workspace_bytes,
                      ptgs_stream_t stream);
int ptgs_expand_fill(int dtype, const void* x, const void* y, int64_t n,
                     const void* rects, int32_t n_blocks,
                     const int64_t* counts, int64_t* offsets,
                     int32_t* indices, int64_t indices_capacity,
                     void* workspace, size_t workspace_bytes,
                     ptgs_stream_t stream);

/* The same selection through a uniform cell grid, for many blocks (N * L
 * closed-rectangle tests become N * ~2).  grid[5] (HOST doubles) = {x0, y0,
 * cells per unit in x, in y, G}: a point's cell is
 * (clamp(floor((x - x0) * grid[2]), 0, G-1), same in y), row-major y * G + x;
 * cell_off[G*G + 1] / cell_list (device int32) give, per cell, the rectangles
 * that may contain a point of that cell (any superset; the exact compares are
 * those of ptgs_expand_count).  Results are identical to ptgs_expand_count /
 * ptgs_expand_fill; workspace = ptgs_expand_workspace_bytes(n, n_blocks). */
int ptgs_expand_count_grid(int dtype, const void* x, const void* y, int64_t n,
                           const void* rects, int32_t n_blocks,
                           const double* grid, const int32_t* cell_off,
                           const int32_t* cell_list, int64_t* counts,
                           void* workspace, size_t workspace_bytes,
                           ptgs_stream_t stream);
int ptgs_expand_fill_grid(int dtype, const void* x, const void* y, int64_t n,
                          const void* rects, int32_t n_blocks,
                          const double* grid, const int32_t* cell_off,
                          const int32_t* cell_list, const int64_t* counts,
                          int64_t* offsets, int32_t* indices,
                          int64_t indices_capacity, void* workspace,
                          size_t workspace_bytes, ptgs_stream_t stream);

/* ------------------------------------------------------------------ *
 * A5 -- camera -> block assignment
 *   replaces assign_cameras_to_partitions (shen_partation_densify.py:362-408):
 *   camera c belongs to block b iff its centre is STRICTLY inside the
 *   expanded box in XY (shapely contains(Point), :388), float64 compares.
 *   centers[C][3] float64 (the caller evaluates -R^T T in float32 exactly as
 *   scene/cameras.py:56-58 and widens); boxes[L][4] = {minx,miny,maxx,maxy}.
 *   cam_offsets[L+1] (int64) is always written; cam_index (int32, ascending
 *   per block) only up to cam_capacity (a second call with a larger buffer
 *   is needed when cam_offsets[L] > cam_capacity).
 * ------------------------------------------------------------------ */
int ptgs_assign_cameras(const double* centers, int32_t n_cameras,
                        const double* boxes, int32_t n_blocks,
                        int64_t* cam_offsets, int32_t* cam_index,
                        int64_t cam_capacity, ptgs_stream_t stream);

/* ------------------------------------------------------------------ *
 * A6 -- voxel down-sampling of every block
 *   replaces downsample_point_cloud(pc, voxel) (shen_partation_densify.py:
 *   541-614): per block  vid = floor((p - min_b)/voxel) (point dtype
 *   arithmetic, :572), one output point per occupied voxel = the mean of its
 *   points accumulated sequentially in index order then divided by the
 *   count (:588-591), cast to float32 (:605).
 *   Inputs: planar coordinates, block CSR (offsets[L+1], indices; indices may
 *   be NULL = identity, i.e. the blocks' points are stored back to back).
 *   xyz_aos (may be NULL): the same cloud as (N,3) rows of the same dtype;
 *   when it is given it is the source that is read and x / y / z may be NULL.
 *   With an index list the call first packs the entries' coordinates into one
 *   16 / 32-byte record each (workspace) so that the three passes that need
 *   them read whole sectors; reading the rows instead of three planes makes
 *   that one gather 2-3x cheaper.  Without one (blocks uploaded back to back
 *   as rows) no plane layout pass is needed at all.
 *   Outputs: ds_x/ds_y/ds_z (float32, capacity = offsets[L]), ds_offsets[L+1]
 *   (int64), ds_first (int32, capacity as ds_x): position inside the block's
 *   index list of the first point of the voxel.  Voxels of a block are
 *   emitted grouped by coarse XY cell (64 x 64 voxels, cell-major) so that
 *   consecutive output points are spatially compact; the reference's
 *   first-occurrence order is ds_first ascending.
 *   status_dev: PTGS_E_RANGE when a voxel index exceeds the key layout.
 * ------------------------------------------------------------------ */
size_t ptgs_voxel_downsample_workspace_bytes(int64_t n_entries,
                                             int32_t n_blocks);
int ptgs_voxel_downsample(int dtype, const void* x, const void* y,
                          const void* z, const void* xyz_aos,
                          const int64_t* offsets,
                          const int32_t* indices, int64_t n_entries,
                          int32_t n_blocks, double voxel_size, float* ds_x,
                          float* ds_y, float* ds_z, int32_t* ds_first,
                          int64_t* ds_offsets, int32_t* status_dev,
                          void* workspace, size_t workspace_bytes,
                          ptgs_stream_t stream);

/* ------------------------------------------------------------------ *
 * A7-A9 -- camera visibility
 *   replaces project_points_to_camera (shen_partation_densify.py:410-477),
 *   run_graham_scan (:481-539), process_camera_visibility (:890-910) and the
 *   selection loop of visibility_based_camera_selection (:911-954).
 *
 * Camera record: 20 floats
 *   [0..8]  R row-major (world->camera rotation, float32 of camera.R)
 *   [9..11] T
 *   [12] fx  [13] fy  [14] cx  [15] cy     (float32 of the Python doubles)
 *   [16] u_limit (image_width)  [17] v_limit (image_height)
 *   [18] front_axis (1.0 = camera-space Y > 0 as the reference, :447;
 *                    2.0 = Z > 0)          [19] reserved (0)
 * One test (fp32):
 *   Xc = ((r00*x (+fma) r01*y) (+fma) r02*z) + tx          (k-ordered GEMM)
 *   uh = fma(cx, Zc, fx*Xc), vh = fma(cy, Zc, fy*Yc), u = uh/Zc, v = vh/Zc
 *   hit = front > 0 && 0 <= u < u_limit && 0 <= v < v_limit
 * ------------------------------------------------------------------ */
#define PTGS_CAMERA_FLOATS 20

/* The visibility pass is two sweeps over the same (block, camera) pairs plus a
 * per-pair hull.  pair p = cam_offsets[b] + (position of the camera in block
 * b's list).  Point arrays px/py/pz hold every block's down-sampled points
 * back to back (pt_offsets[L+1]); they must be 16-byte aligned and readable up
 * to the next multiple of 4 elements past pt_offsets[L] (the tiles are staged
 * with 16-byte granular bulk copies).
 *
 * Sweep 1 (ptgs_visibility_count): per pair
 *   n_hit[p]   int32   number of in-image points (len(image_points), :898)
 *   pivot[p]   uint64  (bits(v) << 32 | bits(u)) of the Graham pivot: lowest v,
 *                      then lowest u (:509-511); ~0 when n_hit == 0
 *   extreme[p][PTGS_EXTREME_DIRS] uint64  (ordered key << 32 | point position
 *                      inside the block) of the extreme hit along the directions
 *                      (a, b), value a*u + b*v with one float32 rounding,
 *                      counter-clockwise from +u: (1,0) (1,1) (0,1) (-1,1) and
 *                      their negatives (PTGS_EXTREME_DIRS = 8), plus (2,1) (1,2)
 *                      (-1,2) (-2,1) and their negatives in between (= 16); 0 = none.
 * All three are accumulated with atomics and must be pre-initialised by
 * ptgs_visibility_init (also zeroes cand_count).
 *
 * ptgs_visibility_bounds: per pair
 *   - the inner polygon (the extreme hits that survive the reference's
 *     pivot-row rule, see below) as <= PTGS_EXTREME_DIRS half planes -> `poly`, and its area
 *     / (W*H) -> ratio_lo: a lower bound of the reference's ratio;
 *   - the outer polygon (intersection of the supporting half planes of sweep
 *     1) -> ratio_hi: an upper bound;
 *   - the early decision: n_hit <= min_hits -> dropped (:898); ratio_lo >=
 *     thr + near_eps -> kept; ratio_hi < thr - near_eps -> dropped.  Decided
 *     pairs get keep / ratio (= the deciding bound) / area and flags bit4; the
 *     others get need[p] = 1 and must go through sweep 2 + the exact hull.
 *     exact_all != 0 disables the early decision (every pair with enough hits
 *     gets its exact ratio).
 * ptgs_visibility_plan: cand_offsets[P+1] = exclusive scan of the candidate
 *   capacities: need ? (caps_in ? caps_in[p] : min(n_hit[p], max(cap_max,
 *   n_hit[p]/2))) : 0 (n_hit[p] when the pair has no inner polygon: poly,
 *   may be NULL), every offset clamped to pool_capacity (the size of the
 *   caller's cand_uv pool in candidates; pairs that do not fit overflow).
 *
 * Sweep 2 (ptgs_visibility_candidates): re-projects and emits every hit that
 * is (a) part of the point set the reference's scan effectively uses -- the
 * pivot plus all hits with v > v_pivot: hits on the pivot's row are popped by
 * the `cross <= 0` rule and the farthest of them is skipped by
 * `sorted_points[1:]` (:513-527) -- and (b) not strictly inside the inner
 * polygon.  The (block, camera) lists given to sweep 2 may be a sub-list of the
 * pairs (the undecided ones); pair_map[slot] then names the pair id p of every
 * slot of that sub-list (NULL = identity), poly / cand_offsets / cand_count are
 * always indexed by p.  cand_uv[cand_offsets[p] + k] (float2, k < capacity of the pair);
 * cand_count[p] receives the number of candidates even when it exceeds the
 * pair's capacity (the caller then re-runs the pair with a larger capacity).
 * The candidate SET of a pair is defined per point, the order inside the list
 * is not (the hull sorts).  Implementation: one warp per run of 128 consecutive
 * points of the blocks that own a listed pair (box tests per run, per-point
 * evaluation for the cameras the box cannot decide); `cam_split` only matters
 * to the tile-staged variant kept for A/B runs (environment PTGS_SWEEP2=tiles).
 *
 * ptgs_visibility_hull (cand_uv is compacted in place by a directional
 * pre-filter that drops candidates strictly inside the hull): exact convex hull (float64 orientation tests with the
 * reference's `cross <= 0` pop rule) of each pair's candidates, shoelace area
 * in float64, ratio = area / (W*H); keep = n_hit > min_hits && ratio >= thr.
 * flags: bit0 = |ratio - thr| <= near_eps; bit1 = another hit lies exactly on
 * the first hull edge after the pivot (the one configuration in which the
 * reference's result depends on libm atan2 ties); bit2 = candidate capacity
 * overflow (result invalid, re-run); bit3 = hull larger than the kernel's
 * vertex budget (result invalid); bit4 = decided from the bounds (set by
 * ptgs_visibility_bounds).  One CTA per entry of pair_map (NULL = every pair);
 * n_pairs bounds the grid, *n_sub_dev (device, may be NULL) the live entries.
 * cam_split >= 1 splits every block's camera list into that many work items
 * per point tile (load balance for small inputs); results do not depend on it.
 */
size_t ptgs_visibility_workspace_bytes(int32_t n_blocks);
int ptgs_visibility_init(int64_t n_pairs, int32_t* n_hit, uint64_t* pivot,
                         uint64_t* extreme, int32_t* cand_count,
                         ptgs_stream_t stream);
int ptgs_visibility_count(const float* px, const float* py, const float* pz,
                          const int64_t* pt_offsets, int32_t n_blocks,
                          const float* cameras, const int64_t* cam_offsets,
                          const int32_t* cam_index, int32_t cam_split,
                          int32_t* n_hit, uint64_t* pivot, uint64_t* extreme,
                          const float* krect, void* workspace,
                          size_t workspace_bytes, ptgs_stream_t stream);
/* Interior rectangles for sweep 1 (optional acceleration, results unchanged).
 * ptgs_visibility_sample gathers every `stride`-th point of each block
 * (s_offsets[L+1], sx/sy/sz with room for `capacity` >= sum ceil(M_b / stride)
 * points, padded like the point arrays above).  Run ptgs_visibility_count on
 * the sample (krect = NULL) and hand its n_hit / extreme to
 * ptgs_visibility_krect: per pair a record of PTGS_KRECT_FLOATS floats,
 *   [0..3]  (cu, cv, hu, hv): an axis-parallel rectangle of the image plane
 *           (hu < 0: none),
 *   [4]     number of edges m of a convex polygon (0: none), [5..7] zero,
 *   [8+3e..] (a, b, c) of edge e: a*u + b*v + c >= 0 inside, |(a, b)| = 1,
 * both at least margin_px (>= 0.5) inside the convex hull of the sample's
 * hits.  Passed to the full ptgs_visibility_count, points projecting into the
 * rectangle and point boxes projecting into the polygon are only counted:
 * they are hits, and they cannot be the Graham pivot (:509-511) or a
 * directional extreme. */
#define PTGS_KRECT_FLOATS 32
int ptgs_visibility_sample(const float* px, const float* py, const float* pz,
                           const int64_t* pt_offsets, int32_t n_blocks,
                           int32_t stride, float* sx, float* sy, float* sz,
                           int64_t* s_offsets, int64_t capacity,
                           ptgs_stream_t stream);
/* The sample pass itself, pair-centric: what ptgs_visibility_init +
 * ptgs_visibility_count (krect = NULL) produce on the sample, with one warp per
 * pair and no atomics (n_hit_s / pivot_s / extreme_s are written for every
 * pair; pivot_s may be NULL).  The extremes are hits of the sample; among equal
 * values the largest sample index wins, as in the sweep. */
int ptgs_visibility_sample_count(const float* sx, const float* sy,
                                 const float* sz, const int64_t* s_offsets,
                                 int32_t n_blocks, const float* cameras,
                                 const int64_t* cam_offsets,
                                 const int32_t* cam_index, int64_t n_pairs,
                                 int32_t* n_hit_s, uint64_t* pivot_s,
                                 uint64_t* extreme_s, ptgs_stream_t stream);
int ptgs_visibility_krect(const float* sx, const float* sy, const float* sz,
                          const int64_t* s_offsets, int32_t n_blocks,
                          const float* cameras, const int64_t* cam_offsets,
                          const int32_t* cam_index, const int32_t* n_hit_s,
                          const uint64_t* extreme_s, int64_t n_pairs,
                          double margin_px, float* krect, ptgs_stream_t stream);
#define PTGS_EXTREME_DIRS 8 /* 8 or 16 (compile time); see "extreme" above */
#define PTGS_POLY_FLOATS (4 + 3 * PTGS_EXTREME_DIRS) /* [0]=n_edges [1]=v_pivot [2]=u_pivot [3]=margin, DIRS x (a,b,c) */
int ptgs_visibility_bounds(const float* px, const float* py, const float* pz,
                           const int64_t* pt_offsets, int32_t n_blocks,
                           const float* cameras, const int64_t* cam_offsets,
                           const int32_t* cam_index, const int32_t* n_hit,
                           const uint64_t* pivot, const uint64_t* extreme,
                           int64_t n_pairs, int32_t min_hits,
                           double ratio_threshold, double near_eps,
                           int32_t exact_all, float* poly, double* ratio_lo,
                           double* ratio_hi, uint8_t* need, double* area,
                           double* ratio, uint8_t* keep, uint8_t* flags,
                           ptgs_stream_t stream);
/* ratio_lo / ratio_hi (may be NULL): the pair's area bounds of
 * ptgs_visibility_bounds; the candidate capacity of an undecided pair is then
 * sized from 1 - ratio_lo / ratio_hi instead of half of its hits. */
int ptgs_visibility_plan(const int32_t* n_hit, const uint8_t* need,
                         const int32_t* caps_in, const float* poly,
                         const double* ratio_lo, const double* ratio_hi,
                         int64_t n_pairs, int32_t cap_max,
                         int64_t pool_capacity,
                         int64_t* cand_offsets, ptgs_stream_t stream);
int ptgs_visibility_candidates(const float* px, const float* py,
                               const float* pz, const int64_t* pt_offsets,
                               int32_t n_blocks, const float* cameras,
                               const int64_t* cam_offsets,
                               const int32_t* cam_index, int32_t cam_split,
                               const int32_t* pair_map, const float* poly,
                               const int64_t* cand_offsets,
                               float* cand_uv, int32_t* cand_count,
                               void* workspace, size_t workspace_bytes,
                               ptgs_stream_t stream);
/* workspace (may be NULL): ptgs_visibility_hull_workspace_bytes(n_pairs)
 * bytes let the call pre-filter the pairs with the longest candidate lists
 * over several CTAs each (one pair is otherwise one CTA, and the largest pair
 * sets the duration of the launch); results do not depend on it. */
size_t ptgs_visibility_hull_workspace_bytes(int64_t n_pairs);
int ptgs_visibility_hull(float* cand_uv, const int64_t* cand_offsets,
                         const int32_t* cand_count, const int32_t* n_hit,
                         int64_t n_pairs, const int32_t* pair_cam,
                         const float* cameras, int32_t min_hits,
                         double ratio_threshold, double near_eps, double* area,
                         double* ratio, uint8_t* keep, uint8_t* flags,
                         const int32_t* pair_map, const int64_t* n_sub_dev,
                         int32_t* status, void* workspace,
                         size_t workspace_bytes, ptgs_stream_t stream);
/* status (device int32[2], may be NULL; zeroed by the call): [0] pairs whose
 * candidate list overflowed its capacity (flags bit2, re-run them with
 * cand_count slots), [1] pairs whose hull exceeded the vertex budget (bit3). */

/* A9 selection: per block compaction of the kept cameras (ascending order of
 * the block's camera list).  kept_offsets[L+1] int64, kept_index capacity =
 * cam_offsets[L]; kept_pos (may be NULL, same capacity) receives the pair
 * position (index into cam_index) of every kept entry. */
int ptgs_select_cameras(const uint8_t* keep, const int64_t* cam_offsets,
                        const int32_t* cam_index, int32_t n_blocks,
                        int64_t* kept_offsets, int32_t* kept_index,
                        int32_t* kept_pos, ptgs_stream_t stream);

/* ------------------------------------------------------------------ *
 * N3 -- COLMAP points3D.bin ingest (reference: the per-record struct loop of
 * scene/colmap_loader.py:195-224 and the points3D.bin -> points3D.ply ->
 * fetchPly round trip of scene/ptgs/shen_data_read.py:185-193).
 * ptgs_points3d_index: HOST function on a HOST image of the file; writes the
 *   byte offset of every record (51 bytes + 8 per track element) to
 *   offsets_host[capacity] (may be NULL to count only) and returns the record
 *   count, or a negative ptgs_status (truncated file / capacity).
 * ptgs_points3d_decode: device image of the same file + device offsets ->
 *   planar float32 xyz (the float32 cast storePly applies), optional rgb
 *   (n x 3 uint8, may be NULL) and reprojection error (float64, may be NULL).
 * ------------------------------------------------------------------ */
int64_t ptgs_points3d_index(const uint8_t* file_host, int64_t n_bytes,
                            int64_t* offsets_host, int64_t capacity);
int ptgs_points3d_decode(const uint8_t* file_dev, const int64_t* offsets_dev,
                         int64_t n, float* x, float* y, float* z, uint8_t* rgb,
                         double* error, ptgs_stream_t stream);

/* ------------------------------------------------------------------ *
 * Multi-GPU exchange of the per-pair records (the reference is
 * single process -- the loop that is split across GPUs is
 * scene/ptgs/shen_partation_densify.py:932-944).  With range sharding a rank
 * owns one or a few contiguous ranges of the (block, camera) pair list.
 * ptgs_pack_records: the rank's result vectors -> n_pairs records of
 *   PTGS_PAIR_RECORD_BYTES bytes {f64 ratio, f64 area, i32 n_hit, u8 keep,
 *   u8 flags, u16 0} (8-byte aligned buffer).  The caller all-gathers the
 *   (padded to `slot_pairs` records) buffers of all ranks, e.g. NCCL
 *   all_gather on the same stream.
 * ptgs_unpack_records: `gathered` = the all-gathered table (table_records
 *   records); the pair list is tiled by n_ranges (<= 128) contiguous ranges,
 *   range c = pairs [range_lo[c], range_lo[c+1]) stored from record
 *   range_base[c] on (HOST arrays of n_ranges + 1 / n_ranges entries;
 *   range_lo[0] = 0, range_lo[n_ranges] = n_pairs); writes the full per-pair
 *   vectors.
 * ------------------------------------------------------------------ */
#define PTGS_PAIR_RECORD_BYTES 24
int ptgs_pack_records(const int32_t* n_hit, const double* ratio,
                      const double* area, const uint8_t* keep,
                      const uint8_t* flags, int64_t n_pairs, void* records,
                      ptgs_stream_t stream);
int ptgs_unpack_records(const void* gathered, int32_t n_ranges,
                        int64_t table_records, const int64_t* range_lo_host,
                        const int64_t* range_base_host, int64_t n_pairs,
                        int32_t* n_hit, double* ratio, double* area,
                        uint8_t* keep, uint8_t* flags, ptgs_stream_t stream);

/* ------------------------------------------------------------------ *
 * Coverage sweep (config 5; the camera x all-points pass of the Vast path,
 * scene/ptgs/data_partition.py:385-424, :505): every point against every
 * camera cam_first..cam_first+cam_count-1; n_hit[cam_count] int32 (must be
 * zeroed by the caller... it is zeroed by the call); optional seen[n] uint8
 * (may be NULL): 1 when any camera of the range sees the point.
 * workspace: ptgs_coverage_workspace_bytes(n, cam_count) bytes.  With more
 * than a few hundred cameras the call first builds, per super-tile of the
 * cloud (<= 1024 runs of consecutive points), the list of cameras whose
 * frustum can meet the run's bounding box, and sweeps each run against its
 * list only -- effective when the cloud is in spatial order
 * (ptgs_spatial_order); a workspace of only ptgs_visibility_workspace_bytes(1)
 * bytes selects the direct every-tile-by-every-camera sweep.
 * px / py / pz must be 16-byte aligned and readable up to 16 bytes past their
 * last element (the tiles are staged with 16-byte granular bulk copies).
 * ------------------------------------------------------------------ */
size_t ptgs_coverage_workspace_bytes(int64_t n, int32_t cam_count);
int ptgs_coverage_sweep(const float* px, const float* py, const float* pz,
                        int64_t n, const float* cameras, int32_t cam_first,
                        int32_t cam_count, int32_t* n_hit, uint8_t* seen,
                        void* workspace, size_t workspace_bytes,
                        ptgs_stream_t stream);

/* ------------------------------------------------------------------ *
 * Spatial order for the coverage sweep (the Vast path's cloud arrives in file
 * order; scene/ptgs/data_partition.py:505 projects all of it per camera).
 * Sorts the planar float32 cloud by a 32-bit Morton-type code of a cubic-cell
 * grid over its bounding box (2048 cells along the longest axis; 4 radix
 * passes): sx/sy/sz[n] = the points in that
 * order, perm[n] uint32 = their indices in the input (n < 2^32).  Only the
 * speed of ptgs_coverage_sweep depends on the order, never its results.
 * ptgs_unpermute_u8: dst[perm[i]] = src[i] (per-point flags back to the
 * caller's order).
 * ------------------------------------------------------------------ */
size_t ptgs_spatial_order_workspace_bytes(int64_t n);
int ptgs_spatial_order(const float* px, const float* py, const float* pz,
                       const float* xyz_aos /* may be NULL: same cloud as (N,3) rows, cheaper gather */,
                       int64_t n, float* sx, float* sy, float* sz,
                       uint32_t* perm, void* workspace, size_t workspace_bytes,
                       ptgs_stream_t stream);
int ptgs_unpermute_u8(const uint8_t* src, const uint32_t* perm, int64_t n,
                      uint8_t* dst, ptgs_stream_t stream);

/* ------------------------------------------------------------------ *
 * A14 -- the camera-grid ("Vast") partitioner, scene/ptgs/data_partition.py.
 * The XZ box selection of extract_point_cloud (:276-283) is ptgs_expand_count /
 * ptgs_expand_fill called with (x, z) and rects {x0, x1, z0, z1}.  Everything
 * below computes in float64 in NumPy's operation order (the reference's
 * point_in_image, :385-424, is host NumPy).
 *
 * Camera records: PTGS_VAST_CAMERA_DOUBLES doubles per camera =
 *   {w2c rotation row-major [9] (= camera.R transposed, :389-391), T [3],
 *    fx, fy (:393-394), cx = W//2, cy = H//2 (:404-407), W, H, 0, 0}.
 *
 * ptgs_vast_point_range: get_point_range (:285-293) of every block of a CSR
 *   (offsets[B+1] int64, indices int32 into the cloud): range[B][6] =
 *   {min x, max x, min y, max y, min z, max z}; workspace >= 48 B bytes.
 * ptgs_vast_airspace: for pair p, the 8 corners (:368-382) of
 *   boxes[pair_block[p]] (6 doubles, same layout as `range`) projected into
 *   camera pair_cam[p]; corners with camera z <= 0 are dropped; area of
 *   (convex hull & [0,W]x[0,H]) / (W H), 0 with fewer than 3 distinct corners
 *   (:478-488, graham_scan.py:65-105).
 * Coverage of one destination block i (:441-557), in this order:
 *   ptgs_vast_begin_block   first_pos[n_cloud] (uint64) := ~0
 *   ptgs_vast_coverage      once per source block j, for the cameras of j
 *     accepted for i (cam_list, in the reference's order; seq_base = number of
 *     cameras accepted before): point_in_image of block j's cloud
 *     (block_index[n_block_points] int32 into the cloud, ascending).  A point
 *     in front of the camera and inside the image (x against H, y against W,
 *     :417-422) selects the block's point number `rank` = its position among
 *     the points in front (:412, :505-506), and
 *     first_pos[block_index[rank]] = min(., (seq << 32) | rank).
 *   ptgs_vast_mark_own      block i's own points, seq = total number of
 *     accepted cameras (they come last in the concatenation, :537-539)
 *   ptgs_vast_unique        np.unique(rows, axis=0, return_index=True) (:545):
 *     uniq_index[*n_unique_dev] = cloud indices of the distinct (x, y, z) rows
 *     in lexicographic order, each the first occurrence in concatenation
 *     order.  Synchronises the stream once (the number of selected points
 *     sizes the sorts); PTGS_E_CAPACITY when it exceeds `capacity`.
 * ------------------------------------------------------------------ */
#define PTGS_VAST_CAMERA_DOUBLES 20
int ptgs_vast_point_range(int dtype, const void* x, const void* y, const void* z,
                          const int64_t* offsets, const int32_t* indices,
                          int32_t n_blocks, double* range, void* workspace,
                          size_t workspace_bytes, ptgs_stream_t stream);
int ptgs_vast_airspace(const double* cameras, const double* boxes,
                       const int32_t* pair_block, const int32_t* pair_cam,
                       int64_t n_pairs, double* rate, ptgs_stream_t stream);
int ptgs_vast_begin_block(uint64_t* first_pos, int64_t n, ptgs_stream_t stream);
size_t ptgs_vast_coverage_workspace_bytes(int64_t n_block_points, int32_t n_cameras);
int ptgs_vast_coverage(int dtype, const void* x, const void* y, const void* z,
                       const int32_t* block_index, int64_t n_block_points,
                       const double* cameras, const int32_t* cam_list,
                       int32_t n_cameras, int32_t seq_base, uint64_t* first_pos,
                       void* workspace, size_t workspace_bytes,
                       ptgs_stream_t stream);
int ptgs_vast_mark_own(const int32_t* block_index, int64_t n_block_points,
                       int32_t seq, uint64_t* first_pos, ptgs_stream_t stream);
size_t ptgs_vast_unique_workspace_bytes(int64_t n);
int ptgs_vast_unique(int dtype, const void* x, const void* y, const void* z,
                     int64_t n, const uint64_t* first_pos, int32_t* uniq_index,
                     int64_t capacity, int64_t* n_unique_dev, void* workspace,
                     size_t workspace_bytes, ptgs_stream_t stream);

/* Diagnostics of a -DPTGS_SWEEP_STATS_FINE build only (the release library
 * keeps no device-global counters: the call then zeroes out4 and returns
 * PTGS_E_INVALID).  Synchronising; HOST pointer to 8 counters since the last
 * reset: [0] point-camera tests presented to the sweeps, [1] left after the
 * CTA-level tile cull, [2] swept per point (sweep 1 / coverage), [3] swept per
 * point (sweep 2), [4] accepted per warp box, [5] accepted per tile box,
 * [6] (warp, camera) passes that entered the exact path, [7] 32-point groups
 * on the exact path.  reset != 0 zeroes the counters. */
int ptgs_debug_sweep_stats(uint64_t* out4, int32_t reset);

/* Diagnostics for the roofline report: an FP32 FMA loop without memory traffic
 * (148 x 8 CTAs x 256 threads x 8 independent chains x `iters` FMAs each);
 * out[148 * 8 * 256] float.  The caller times it: lane-ops/s = threads * 8 *
 * iters / seconds. */
int ptgs_debug_fma_probe(float* out, int32_t iters, ptgs_stream_t stream);

#ifdef __cplusplus
}
#endif
#endif /* PTGS_H_ */
