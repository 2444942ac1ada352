/*
 * dbaf_b200.h -- C ABI of libdbaf_b200.so: the sm_100a implementation of DBA-Fusion's
 * `droid_backends` operator family (dense bundle adjustment + correlation).
 *
 * Conventions
 *   - every pointer is a DEVICE pointer unless its name ends in `_host`;
 *   - tensors are dense row-major with the shape given in the comment;
 *   - `stream` is a cudaStream_t passed as void*; every call only enqueues work on it
 *     (no device synchronisation, no allocation) unless stated otherwise;
 *   - return value: 0 = ok, DBAF_EINVAL = bad argument, DBAF_ECUDA = CUDA runtime error;
 *     dbaf_last_error() returns a thread-local message for the last non-zero return;
 *   - poses are [N,7] float32 (tx,ty,tz,qx,qy,qz,qw), world->camera; intrinsics [4] float32
 *     (fx,fy,cx,cy) at the feature-map resolution; ii/jj are int64 edge endpoints.
 *
 * Each entry point names the reference interface it replaces (paths relative to the
 * DBA-Fusion checkout).  The Python module `droid_backends` in dba-fusion_b200/ binds these
 * one to one; INTEGRATION.md shows the binding.
 */
#ifndef DBAF_B200_H
#define DBAF_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define DBAF_OK 0
#define DBAF_EINVAL 1
#define DBAF_ECUDA 2

#define DBAF_F16 0
#define DBAF_F32 1

int dbaf_version(void);
/* sizeof(dbaf_ba_layout) as the library was built: a binding compiled against an older header must refuse to load
 * (the library writes the whole struct).  No counterpart in the reference. */
int dbaf_ba_layout_bytes(void);
const char* dbaf_last_error(void);

/* Kernel-selection switches for tests and profiling (no counterpart in the reference).  Process-wide; the
 * defaults come from the environment, read once (DBAF_SCHUR_IMPL, DBAF_BAND_IMPL, DBAF_BAND_SWEEP,
 * DBAF_BAND_DOMAINS, DBAF_VOLUME_IMPL, DBAF_EIJ_RECOMPUTE).  Keys and values:
 *   "schur_impl"   "auto" | "simt"                 Schur complement on tcgen05 where it pays / fp32 SIMT only
 *   "band_impl"    "auto" | "single" | "twisted"   domain-decomposed banded solve / one sweep / two sweeps
 *   "band_sweep"   "auto" | "smem"                 window of the banded sweep in registers / in shared memory
 *   "band_domains" "<n>"                           number of domains of the decomposed solve (0 = automatic)
 *   "volume_impl"  "auto" | "simt"                 correlation volume on tcgen05 / SIMT cross-check kernels
 *   "eij"          "auto" | "store" | "recompute"  pose-depth blocks of the edges kept in HBM / recomputed */
int dbaf_set_option(const char* key, const char* value);

/* Measurement aid (no counterpart in the reference): a write-only stream over `bytes` of device memory, either
 * cudaMemsetAsync (kind 0) or a kernel of 128-bit stores (kind 1).  bench.py times both to state the HBM write
 * ceiling next to the copy peak for the kernels whose traffic is almost all stores (the correlation volume). */
int dbaf_probe_fill(void* dst, size_t bytes, int kind, void* stream);

/* ------------------------------------------------------------------ geometry ---- */

/* droid_backends.projmap, src/droid.cpp:200-215, src/droid_kernels.cu:471-560,1706-1731.
 * coords [E,ht,wd,3] (x,y; channel 2 = 0), valid [E,ht,wd,1]. */
int dbaf_projmap(const float* poses, const float* disps, const float* intrinsics,
                 const int64_t* ii, const int64_t* jj, int n_edges, int ht, int wd,
                 float* coords, float* valid, void* stream);

/* DepthVideo.reproject / pops.projective_transform without Jacobians, dbaf/depth_video.py:221-229,
 * dbaf/geom/projective_ops.py:96-125 (a "next" row of the scope table: the caller that produces the lookup
 * coordinates).  intrinsics: [4] shared, or [N,4] per frame with per_frame_intrinsics = 1 (frame ii for the
 * back-projection, frame jj for the projection, as the reference indexes them).  coords [E,ht,wd,2],
 * valid [E,ht,wd] (1.0 where the transformed depth exceeds 0.2).  Stereo edges (ii == jj) use the fixed
 * baseline (-0.1, 0, 0). */
int dbaf_reproject(const float* poses, const float* disps, const float* intrinsics, int per_frame_intrinsics,
                   const int64_t* ii, const int64_t* jj, int n_edges, int ht, int wd, float* coords,
                   float* valid, void* stream);

/* droid_backends.frame_distance, src/droid.cpp:181-197, src/droid_kernels.cu:562-702.  dist [M]. */
int dbaf_frame_distance(const float* poses, const float* disps, const float* intrinsics,
                        const int64_t* ii, const int64_t* jj, int n_pairs, int ht, int wd,
                        float beta, float* dist, void* stream);

/* droid_backends.iproj, src/droid.cpp:218-227, src/droid_kernels.cu:824-895.  points [N,ht,wd,3]. */
int dbaf_iproj(const float* poses, const float* disps, const float* intrinsics,
               int n_frames, int ht, int wd, float* points, void* stream);

/* droid_backends.depth_filter, src/droid.cpp:281-295, src/droid_kernels.cu:706-820.
 * counter [M,ht,wd] is overwritten (the reference allocates it zeroed). */
int dbaf_depth_filter(const float* poses, const float* disps, const float* intrinsics,
                      const int64_t* ix, const float* thresh, int n_idx, int n_frames,
                      int ht, int wd, float* counter, void* stream);

/* ----------------------------------------------------------- frame-graph edges ---- */

/* CovisibleGraph.add_proximity_factors, dbaf/covisible_graph.py:357-441 (a "next" row of the scope table): which
 * edges to add, given the distances of the candidate pairs.  The candidates are the (t-t0) x (t-t1) grid in
 * row-major order (cc = (t-t0)*(t-t1) entries: cand_ii[k] = t0 + k / (t-t1), cand_jj[k] = t1 + k % (t-t1)) followed
 * by n_cand - cc "skip" pairs (:371-377); d1 (and d2, may be NULL) are their frame distances, d = 0.5 (d1 + d2)
 * being DepthVideo.distance's bidirectional mean (dbaf/depth_video.py:255-261).  ex_ii/ex_jj: the edges the graph
 * already holds (active, bad and inactive, :383-384).  The kernel applies the masks of :379-381, the suppression
 * stamps of :386-394, emits the unconditional temporal (and stereo) edges of :396-406, runs the greedy sweep in
 * ascending distance with its 2-D suppression (:408-433; ties go to the lower index, NaN last) and the closest
 * skip edge (:435-439).  es_ii/es_jj [es_capacity] receive the list in the reference's order, *n_es its length
 * (device int32); es_capacity >= dbaf_proximity_base_count(t0,t,rad,stereo) + 2*cc + 2; d_work: float[n_cand]
 * device scratch.  One thread block; no host synchronisation (the reference does one per candidate). */
int dbaf_proximity_base_count(int t0, int t, int rad, int stereo);
int dbaf_proximity_select(const float* d1, const float* d2, const int64_t* cand_ii, const int64_t* cand_jj,
                          int n_cand, int cc, const int64_t* ex_ii, const int64_t* ex_jj, int n_existing,
                          int t0, int t1, int t, int rad, int nms, float thresh, int stereo, int max_factors,
                          float* d_work, int64_t* es_ii, int64_t* es_jj, int es_capacity, int32_t* n_es,
                          void* stream);

/* CovisibleGraph.__filter_repeated_edges, dbaf/covisible_graph.py:61-72: the new edges (ii[k], jj[k]),
 * k < min(*n_dev, n_max) (n_dev: device int32, may be NULL = n_max), that are not among ex_ii/ex_jj (the active and
 * inactive edges), in their original order -> out_ii/out_jj [n_max], count in *n_out (device int32). */
int dbaf_filter_repeated_edges(const int64_t* ii, const int64_t* jj, const int32_t* n_dev, int n_max,
                               const int64_t* ex_ii, const int64_t* ex_jj, int n_existing, int64_t* out_ii,
                               int64_t* out_jj, int32_t* n_out, void* stream);

/* --------------------------------------------------------------- correlation ---- */

/* droid_backends.corr_index_forward, src/droid.cpp:231-239, src/correlation_kernels.cu:19-70,126-155.
 * volume [n,h1,w1,h2,w2] (dtype DBAF_F16 or DBAF_F32), coords [n,2,h1,w1] float32 (x then y),
 * corr [n,2r+1,2r+1,h1,w1] same dtype, fully overwritten.  fp16 results reproduce the
 * reference's rounding sequence bit for bit. */
int dbaf_corr_index_forward(const void* volume, int dtype, const float* coords,
                            int n, int h1, int w1, int h2, int w2, int radius,
                            void* corr, void* stream);

/* All pyramid levels in one launch: what CorrBlock.__call__ does with `levels` separate
 * calls (dbaf/modules/corr.py:40-50).  volumes[l] is [n,h1,w1,h2>>l,w2>>l]; coords are the
 * level-0 coordinates (scaled by 2^-l inside); out [n,levels*(2r+1)^2,h1,w1] is the
 * concatenated tensor `torch.cat(out_pyramid, dim=2)`.  levels <= 8. */
int dbaf_corr_lookup_pyramid(const void* const* volumes_host, int levels, int dtype,
                             const float* coords, int n, int h1, int w1, int h2, int w2,
                             int radius, void* out, void* stream);

/* droid_backends.corr_index_backward, src/droid.cpp:241-252, src/correlation_kernels.cu:73-124,157-185.
 * volume_grad [n,h1,w1,h2,w2] is fully overwritten. */
int dbaf_corr_index_backward(const float* coords, const void* corr_grad, int dtype,
                             int n, int h1, int w1, int h2, int w2, int radius,
                             void* volume_grad, void* stream);

/* droid_backends.altcorr_forward, src/droid.cpp:254-264, src/altcorr_kernel.cu:27-149,290-319.
 * fmap1 [B,H1,W1,C], fmap2 [B,H2,W2,C] float32, coords [B,S,H1,W1,2], corr [B,S,(2r+1)^2,H1,W1]. */
int dbaf_altcorr_forward(const float* fmap1, const float* fmap2, const float* coords,
                         int B, int S, int H1, int W1, int H2, int W2, int C, int radius,
                         float* corr, void* stream);

/* droid_backends.altcorr_backward, src/droid.cpp:266-279, src/altcorr_kernel.cu:152-286,321-356.
 * fmap1_grad/fmap2_grad are fully overwritten; coords_grad [B,S,H1,W1,2] is zero-filled. */
int dbaf_altcorr_backward(const float* fmap1, const float* fmap2, const float* coords,
                          const float* corr_grad, int B, int S, int H1, int W1, int H2, int W2,
                          int C, int radius, float* fmap1_grad, float* fmap2_grad,
                          float* coords_grad, void* stream);

/* CorrBlock.corr + pyramid, dbaf/modules/corr.py:24-38,63-71 (torch.matmul + 3x avg_pool2d).
 * fmap1, fmap2 [n,C,ht,wd] fp16 (channel-major, as stored by DepthVideo); out_host[l] is
 * [n,ht,wd,ht>>l,wd>>l] fp16.  The 1/4 scaling of both operands is folded in.
 * workspace: dbaf_corr_volume_workspace_bytes(n,C,ht,wd) bytes of device scratch. */
size_t dbaf_corr_volume_workspace_bytes(int n, int C, int ht, int wd);
int dbaf_corr_volume_pyramid(const void* fmap1, const void* fmap2, int n, int C, int ht, int wd,
                             int levels, void* const* out_host, void* workspace,
                             size_t workspace_bytes, void* stream);

/* The same pyramid for a list of edges over frame stacks, i.e. CorrBlock(video.fmaps[ii,0], video.fmaps[jj,c])
 * of dbaf/covisible_graph.py:129-131 with the two gathers folded in: edge e correlates fmaps1[ii[e]] with
 * fmaps2[jj[e]].  fmaps1 [n1,C,ht,wd], fmaps2 [n2,C,ht,wd] fp16 (may be the same pointer: it is then laid out
 * once); ii, jj int32[n_edges] on the device; out_host[l] is [n_edges,ht,wd,ht>>l,wd>>l] fp16.  Results are
 * bit-identical to dbaf_corr_volume_pyramid on the gathered maps.  An index (or, for the pool variant, a slot) outside
 * its range is caught on the device before anything is computed: the call raises a status word inside the workspace,
 * its remaining kernels write nothing, and nothing synchronises; dbaf_corr_volume_status() copies the word (0 = ok,
 * 1 = bad index) to the host on `stream` for the caller to read after its next synchronisation.
 * workspace: dbaf_corr_volume_indexed_workspace_bytes(n1,n2,C,ht,wd) bytes of device scratch (required). */
size_t dbaf_corr_volume_indexed_workspace_bytes(int n1, int n2, int C, int ht, int wd);
int dbaf_corr_volume_status(const void* workspace, size_t workspace_bytes, int32_t* status_host, void* stream);
int dbaf_corr_volume_pyramid_indexed(const void* fmaps1, int n1, const void* fmaps2, int n2,
                                     const int32_t* ii, const int32_t* jj, int n_edges, int C, int ht,
                                     int wd, int levels, void* const* out_host, void* workspace,
                                     size_t workspace_bytes, void* stream);

/* CorrBlock bookkeeping without copies (a "next" row of the scope table).  The reference keeps the pyramids of the
 * live edges in tensors that it re-creates on every edge change: `cat` on add, boolean-mask indexing on removal
 * (dbaf/modules/corr.py:52-60, dbaf/covisible_graph.py:131-132,166-167), i.e. the whole multi-GB pyramid is
 * copied each time.  Here the pyramids live in caller-owned pools of `capacity` volumes per level,
 * pool_host[l] = [capacity, ht, wd, ht>>l, wd>>l] fp16, and an int32 device array maps edge -> slot:
 * adding edges fills free slots in place, removing edges only edits the slot array.
 * dbaf_corr_volume_pyramid_pool: as dbaf_corr_volume_pyramid_indexed, edge e written to slot slots[e] (checked
 *   against capacity; workspace: dbaf_corr_volume_indexed_workspace_bytes).
 * dbaf_corr_lookup_pyramid_pool: as dbaf_corr_lookup_pyramid, edge e (coords [n,2,h1,w1]) reads slot slots[e]. */
int dbaf_corr_volume_pyramid_pool(const void* fmaps1, int n1, const void* fmaps2, int n2, const int32_t* ii,
                                  const int32_t* jj, const int32_t* slots, int n_edges, int capacity, int C,
                                  int ht, int wd, int levels, void* const* pool_host, void* workspace,
                                  size_t workspace_bytes, void* stream);
int dbaf_corr_lookup_pyramid_pool(const void* const* pool_host, int levels, int dtype, const float* coords,
                                  const int32_t* slots, int n, int capacity, int h1, int w1, int h2, int w2,
                                  int radius, void* out, void* stream);

/* ------------------------------------------------------- bundle adjustment ---- */

/* Workspace layout (byte offsets into one caller-allocated device buffer).  The caller owns
 * the buffer; H/v live in it as float64 so a host can read them back (BACore.hessian) or a
 * collective can reduce them in place (edge-sharded global BA). */
typedef struct dbaf_ba_layout {
  int64_t total_bytes;
  int64_t info;      /* int32[16]: [0]=K (#depth frames), [1]=#failed factorisations, [2]=bad index flag,
                                   [3]=eta row mismatch flag, [4]=#rows of E inside the Schur complement */
  int64_t H;         /* float64[(6P)^2] followed directly by v float64[6P] (one contiguous span) */
  int64_t v;
  int64_t dx;        /* float32[P*6] */
  int64_t Lfac;      /* float64 factorisation scratch, used only when 6P > 160 */
  int64_t kx;        /* int32[Kcap]   depth-frame list (sorted unique of cat(arange(t0,t1), ii)) */
  int64_t kk_exp;    /* int32[P+E]    inverse map of the unique, src/droid_kernels.cu:1420-1424 */
  int64_t csr_ptr;   /* int32[Kcap+1] edges grouped by source frame ... */
  int64_t csr_idx;   /* int32[E]      ... in ascending edge order */
  int64_t srow_ptr;  /* int32[Kcap+1] per depth frame: rows of E whose pose is inside [t0,t1) ... */
  int64_t srow_idx;  /* int32[P+E]    ... row n<P = self row of pose t0+n, row P+e = edge e
                                      (the graph/index lists of schur_block, src/droid_kernels.cu:1316-1328) */
  int64_t frame2k;   /* int32[N]      frame -> position in kx, -1 if absent */
  int64_t cursor;    /* int32[Kcap+1] scratch */
  int64_t Q;         /* float32[Kcap,HW] */
  int64_t w;         /* float32[Kcap,HW] */
  int64_t Ei;        /* float32[P,6,HW] */
  int64_t Eij;       /* float32[E,6,HW] */
  int64_t fused;     /* int32[Kcap]   1 = the frame's pose-depth blocks are recomputed, not stored (see "eij" option) */
  int64_t exf;       /* float32[E,28] per-edge relative-pose record of such frames */
  int64_t rc_ptrs;   /* 2 device pointers: the weights and intrinsics of the last such linearisation */
  int64_t Kcap;      /* capacity of the depth-frame tables = min(N, P+E) */
} dbaf_ba_layout;

int dbaf_ba_workspace_layout(int n_frames, int n_edges, int t0, int t1, int ht, int wd,
                             dbaf_ba_layout* out_host);

/* Graph bookkeeping of ba_cuda, src/droid_kernels.cu:1416-1424 plus the segment lists that
 * accum_cuda rebuilds on the host every call (:993-1031).  Once per window. */
int dbaf_ba_prepare(void* workspace, const int64_t* ii, const int64_t* jj, int n_edges,
                    int n_frames, int t0, int t1, int ht, int wd, void* stream);

/* No counterpart in the reference.  The same graph tables as dbaf_ba_prepare, for callers that may
 * not allocate device memory around the call (a real-time tracking thread whose memory budget was
 * handed out at start-up -- cudaMalloc locks the driver and synchronises the device -- or a plugin
 * whose host application owns every buffer).  The caller passes inputs and outputs only:
 *   info int32[16], kx int32[min(N,P+E)], kk_exp int32[P+E], csr_ptr int32[min(N,P+E)+1],
 *   csr_idx int32[E]      (P = t1-t0, meanings as in dbaf_ba_layout).
 * Intermediate tables live in a 1.5 MiB array inside the library image (held by every GPU that loads
 * the library, until exit).  Limits: n_frames <= 16384, n_edges <= 262144, else DBAF_EINVAL.
 * The first call on a device resolves that array's device address (host-side set-up, no
 * allocation): make it before any stream capture.  Calls on one device share the array, which
 * keeps the previous call's contents (the kernel resets all it reads); calls that overlap on the
 * GPU race on it, so keep them in one stream or otherwise serialised. */
int dbaf_ba_tables_noalloc(const int64_t* ii, const int64_t* jj, int n_edges, int n_frames,
                           int t0, int t1, int32_t* info, int32_t* kx, int32_t* kk_exp,
                           int32_t* csr_ptr, int32_t* csr_idx, void* stream);

/* One linearisation: projective_transform_kernel + SparseBlock assembly + accum_cuda + schur_block,
 * src/droid_kernels.cu:220-468,1162-1222,993-1043,1297-1391.  Zeroes H/v first when `zero_first`,
 * then adds this call's edges: A - S into H, b_A - b_S into v (undamped, float64), and stores
 * E, Q, w for the back-substitution.  Only depth frames k in [k_lo,k_hi) (positions in kx) and
 * their out-edges are processed, which is how edge-sharded global BA splits the work
 * (k_hi < 0 means all).  alpha: sensor-depth prior weight (0.05 in `ba`, 0.001 in BACore).
 * eta [eta_rows,ht,wd]; eta_rows must equal K or be 1 (broadcast); anything else raises info[3] and the
 * retraction then leaves disps and dz_out untouched (the reference fails with a broadcast error there).
 * motion_only: 0 = full, 1 = pose blocks only (no E/Q/w, no Schur), 2 = full linearisation but
 * the Schur kernel is not launched (profiling aid: times the Jacobian kernel alone). */
int dbaf_ba_linearize(void* workspace, const float* poses, const float* disps,
                      const float* intrinsics, const float* disps_sens, const float* targets,
                      const float* weights, const float* eta, int eta_rows,
                      const int64_t* ii, const int64_t* jj, int n_edges, int n_frames,
                      int t0, int t1, int ht, int wd, float alpha, int motion_only,
                      int zero_first, int k_lo, int k_hi, void* stream);

/* The Schur stage of dbaf_ba_linearize alone (schur_block + EEt6x6_kernel + Ev6x1_kernel,
 * src/droid_kernels.cu:1297-1391,1046-1138): H -= E Q E^T, v -= E Q w for the depth frames [k_lo,k_hi) from the
 * E, Q, w that the workspace holds.  Diagnostics and tests (feeding the Schur kernels with chosen blocks). */
int dbaf_ba_schur(void* workspace, const int64_t* jj, int n_edges, int n_frames, int t0, int t1, int ht, int wd,
                  int k_lo, int k_hi, void* stream);

/* SparseBlock::solve / solveDense, src/droid_kernels.cu:1248-1288,180-198: L = H (+ prior);
 * diag += ep + lm*diag; float64 Cholesky; dx -> workspace (float32), zeros if a pivot <= 0.
 * prior (float64 [prior_rows,prior_cols], may be NULL) follows ba_cuda_extend :1624-1644. */
int dbaf_ba_solve(void* workspace, int t0, int t1, float lm, float ep,
                  const double* prior, int prior_rows, int prior_cols, void* stream);

/* Same solve for large windows whose reduced system is block-banded: every non-zero 6x6 block (I,J) of H
 * satisfies |I-J| <= band_blocks (a temporal graph with edges reaching `reach` frames gives 2*reach).
 * Windows up to 6P = 160 ignore the hint and use the dense shared-memory solver; beyond that a band of
 * 1..25 blocks selects the sliding-window factorisation, anything else the slow dense fallback.
 * Chains of 64 or more poses are factorised from both ends at once (two thread blocks meeting at a separator);
 * long chains with a band of at most 20 blocks are cut into up to 16 domains that are swept concurrently, with the
 * block-tridiagonal system over the separators solved afterwards by block cyclic reduction (csrc/ba_spike.cuh).
 * The workspace must have gone through dbaf_ba_prepare (it zeroes the failure counter info[1]).
 * Replaces Eigen::SimplicialLLT on the host (src/droid_kernels.cu:1248-1269). */
int dbaf_ba_solve_banded(void* workspace, int t0, int t1, float lm, float ep, int band_blocks, void* stream);
/* Zeroes the reduced camera system before a dbaf_ba_linearize(..., zero_first = 0): only the entries within
 * `band_blocks` pose blocks of the diagonal, and v, when the caller promises (as for dbaf_ba_solve_banded) that
 * nothing else is ever written -- 12 MB instead of the 288 MB of the dense float64 matrix at 1000 poses.
 * band_blocks <= 0 clears everything.  The reference allocates fresh zero tensors every iteration
 * (src/droid_kernels.cu:1440-1447). */
int dbaf_ba_zero_system(void* workspace, int t0, int t1, int band_blocks, void* stream);
/* The band of H (row r: the columns of pose blocks r/6 - band_blocks .. r/6 + band_blocks) and v as one dense strip
 * of dbaf_ba_band_strip_doubles() float64 values: pack = 1 copies workspace -> strip, pack = 0 strip -> workspace.
 * This is what edge-sharded global BA sums over the GPUs (one NCCL all-reduce on the strip) instead of the dense
 * matrix.  No counterpart in the reference (it has no multi-GPU path). */
size_t dbaf_ba_band_strip_doubles(int n_poses, int band_blocks);
int dbaf_ba_band_strip(void* workspace, int t0, int t1, int band_blocks, double* strip, int pack, void* stream);
/* how many kernels dbaf_ba_solve_banded enqueues for that many poses and that band (bench.py counts launches with it) */
int dbaf_ba_solve_banded_launches(int n_poses, int band_blocks);

/* EvT6x1_kernel + dz + pose_retr_kernel + disp_retr_kernel, src/droid_kernels.cu:1140-1160,
 * 1486-1506,943-991.  dx_in: float32 [P,6] (NULL = the workspace's dx).  poses/disps are updated
 * in place.  dz_out [K,HW] (may be NULL; K is only known on the device, info[0], so size it with Kcap rows
 * unless K is known), dx_out [P,6] (may be NULL).  update_disps=0 gives the
 * motion-only update.  Depth frames [k_lo,k_hi) only, as in dbaf_ba_linearize. */
int dbaf_ba_retract(void* workspace, float* poses, float* disps, const float* dx_in,
                    const int64_t* ii, const int64_t* jj, int n_edges, int n_frames,
                    int t0, int t1, int ht, int wd, int update_poses, int update_disps,
                    int k_lo, int k_hi, float* dx_out, float* dz_out, void* stream);

/* droid_backends.ba, src/droid.cpp:109-138, src/droid_kernels.cu:1394-1512: prepare, then
 * `iterations` x {linearize, solve, retract}.  dx_out [P,6], dz_out [K,HW] may be NULL. */
int dbaf_ba(void* workspace, float* poses, float* disps, const float* intrinsics,
            const float* disps_sens, const float* targets, const float* weights,
            const float* eta, int eta_rows, const int64_t* ii, const int64_t* jj,
            int n_edges, int n_frames, int t0, int t1, int ht, int wd, int iterations,
            float lm, float ep, int motion_only, float* dx_out, float* dz_out, void* stream);

/* BA2GTSAM, dbaf/depth_video.py:20-29 (a "next" row of the scope table): the reduced camera system left in the
 * workspace by dbaf_ba_linearize, transformed block by block, H_out = J^T H J, v_out = J^T v with J = blockdiag(A).
 * A: float64 [6,6] row-major on the device (the reference passes the row-swapped -Ad(Tbc^-1));
 * H_out float64 [(6P),(6P)], v_out float64 [6P] on the device (they must not alias the workspace). */
int dbaf_ba_congruence(void* workspace, int t0, int t1, const double* A, double* H_out, double* v_out,
                       void* stream);

/* The same call for callers whose window lives in HOST memory (a frontend that assembles targets/weights on the
 * CPU, a process that receives them over IPC): every input is copied to `staging` (device scratch of
 * dbaf_ba_host_staging_bytes() bytes) with cudaMemcpyAsync on `stream`, dbaf_ba runs on the copies, and the updated
 * poses[t0:t1], disps (unless motion_only), dx_out_h [P,6] (may be NULL) and the status words info_out_h int32[16]
 * (may be NULL; meanings as dbaf_ba_layout.info) are copied back.  Everything is enqueued
 * on `stream`; the host buffers must stay valid, and should be pinned, until the caller has synchronised it.  No
 * reference counterpart: droid_backends.ba only accepts CUDA tensors (src/droid.cpp:16-18 CHECK_INPUT). */
size_t dbaf_ba_host_staging_bytes(int n_frames, int n_edges, int t0, int t1, int ht, int wd, int eta_rows);
int dbaf_ba_host(void* workspace, void* staging, size_t staging_bytes, float* poses_h, float* disps_h,
                 const float* intrinsics_h, const float* disps_sens_h, const float* targets_h,
                 const float* weights_h, const float* eta_h, int eta_rows, const int64_t* ii_h,
                 const int64_t* jj_h, int n_edges, int n_frames, int t0, int t1, int ht, int wd,
                 int iterations, float lm, float ep, int motion_only, float* dx_out_h, int32_t* info_out_h,
                 void* stream);

#ifdef __cplusplus
}
#endif
#endif /* DBAF_B200_H */
