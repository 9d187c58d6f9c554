/*
 * forestslam_b200.h — C ABI of the B200-native (sm_100a) matching + pose hot path of Forest-SLAM.
 *
 * This is the drop-in boundary.  The reference (si220/Forest-SLAM) has no FFI of its
 * own: its boundary is a handful of cv2 / torch call expressions inside the frame loop of
 * ros_ws/src/mono_slam.py and ros_ws/src/stereo_slam.py.  Every entry point below names the call
 * expression it replaces.  The Python host mirror (forest_slam_b200/api.py) binds these with ctypes and
 * re-exposes them under the cv2 names the scripts use; INTEGRATION.md shows the stub.
 *
 * Conventions (all entry points):
 *   - every pointer is a DEVICE pointer unless the parameter name ends in `_host`;
 *   - the caller owns every buffer (inputs, outputs, workspace); the library never allocates device
 *     memory, never synchronises the device and enqueues only on `stream` (a cudaStream_t);
 *   - a "frame bank" is a dense array [n_frames, n_max, ...] with a per-frame count array
 *     `counts[n_frames]` (int32); a "pair list" is two int32 arrays pair_a / pair_b of frame indices.
 *     Temporal sequences use (i, i+1); loop closure uses every (a, b), a < b;
 *   - a "correspondence batch" is [n_pairs, n_max, ...] with per-pair counts `n_corr[n_pairs]`;
 *   - return value: FS_OK (0) or a negative FSStatus.  Nothing throws across the ABI.
 *     Per-pair degenerate inputs (fewer than 5 correspondences for E, fewer than 6 for the PnP gate of
 *     stereo_slam.py:292) do not fail the call: they are reported in the per-pair `status` output.
 *   - results are deterministic: same inputs, same bits, independent of batch split / GPU count.
 */
#ifndef FORESTSLAM_B200_H
#define FORESTSLAM_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef void* fs_stream_t; /* cudaStream_t */

typedef enum FSStatus {
    FS_OK = 0,
    FS_ERR_BAD_ARG = -1,         /* null pointer, negative size, n_max above the compiled limit */
    FS_ERR_WORKSPACE = -2,       /* workspace_bytes smaller than fs_*_workspace_bytes() */
    FS_ERR_CUDA = -3,            /* a CUDA runtime call failed; see fs_last_error_string() */
    FS_ERR_UNSUPPORTED = -4      /* feature not available on this device (not sm_100) */
} FSStatus;

/* per-pair status codes written to `status[n_pairs]` by the RANSAC entry points */
enum {
    FS_PAIR_OK = 0,
    FS_PAIR_TOO_FEW = 1,   /* <5 (E) or <6 (PnP gate, stereo_slam.py:292): outputs zeroed, mask all 0 */
    FS_PAIR_NO_MODEL = 2   /* RANSAC found no model with >= 5 inliers (cv2 returns None / False) */
};

int fs_version(void);
const char* fs_last_error_string(void);
const char* fs_status_string(int status);
/* number of SMs of the current device (grid sizing is done inside the library; exported for benches) */
int fs_device_sm_count(void);
/* number of kernels this library has launched in this process so far (bench.py reports the delta) */
long long fs_launch_count(void);

/* ------------------------------------------------------------------------------------------------
 * A2  binary descriptor matching — replaces
 *       bf = cv2.BFMatcher(cv2.NORM_HAMMING, crossCheck=True)      stereo_slam.py:85
 *       matches = bf.match(descs0, descs1)                        stereo_slam.py:234, :242
 *     and the k=2 + Lowe-ratio variant (ORB_SLAM3/src/Frame.cc:1144-1151).
 *     Distance definition: 256-bit Hamming (ORB_SLAM3/src/ORBmatcher.cc:2058-2074).
 *
 * desc      [n_frames, n_max, 32] u8      ORB-style descriptors, row-major, 16-byte aligned rows
 * counts    [n_frames] i32                valid rows per frame (<= n_max)
 * pair_a/b  [n_pairs] i32                 query frame / train frame of each pair
 * cross_check  !=0: keep (q,t) only if t's best query is q (lowest index wins ties, both directions)
 * ratio     <=0: off; else keep only if (double)best < (double)second * (double)ratio
 *           (a query with a single train row has no second neighbour and is dropped, like knnMatch k=2)
 * outputs, each [n_pairs, n_max] i32:
 *   matches0   train index of query q, or -1       (what DMatch.queryIdx/trainIdx encode, sorted by q)
 *   dist0      best Hamming distance of q (valid even when matches0 is -1); second0 the runner-up
 *              (INT32_MAX when there is none); either may be NULL
 *   matches1   query index matched to train t, or -1; may be NULL
 * n_matches [n_pairs] i32  number of q with matches0 >= 0; may be NULL
 * ---------------------------------------------------------------------------------------------- */
size_t fs_hamming_match_workspace_bytes(int n_pairs, int n_max);
int fs_hamming_match(const uint8_t* desc, int n_frames, int n_max, const int32_t* counts,
                     const int32_t* pair_a, const int32_t* pair_b, int n_pairs,
                     int cross_check, float ratio,
                     int32_t* matches0, int32_t* dist0, int32_t* second0, int32_t* matches1,
                     int32_t* n_matches,
                     void* workspace, size_t workspace_bytes, fs_stream_t stream);

/* A2 on the tensor cores (guided/all-pairs matching of ORB_SLAM3/src/ORBmatcher.cc:648-762 is the
 * workload it is for): same reference lines, same outputs bit for bit as fs_hamming_match.  Every descriptor bit
 * becomes a +-1 fp16 operand, dot = 256 - 2 * hamming is an exact integer in the fp32 tcgen05 accumulator, and the
 * packed-key row top-3 epilogue of the float matcher orders by (distance, lowest index) like BFMatcher.
 * fs_hamming_prepare: once per FRAME, desc [n_frames, n_max, 32] u8 -> operands [n_frames, n_max + 1, 512] fp16
 *   (fs_hamming_operand_bytes); fs_hamming_match_tc: n_max <= 4096 (FS_ERR_UNSUPPORTED above: use fs_hamming_match). */
size_t fs_hamming_operand_bytes(int n_frames, int n_max);
int fs_hamming_prepare(const uint8_t* desc, int n_frames, int n_max, const int32_t* counts,
                       void* operands_fp16, fs_stream_t stream);
size_t fs_hamming_match_tc_workspace_bytes(int n_pairs, int n_max);
int fs_hamming_match_tc(const void* operands_fp16, int n_frames, int n_max, const int32_t* counts,
                        const int32_t* pair_a, const int32_t* pair_b, int n_pairs,
                        int cross_check, float ratio,
                        int32_t* matches0, int32_t* dist0, int32_t* second0, int32_t* matches1,
                        int32_t* n_matches,
                        void* workspace, size_t workspace_bytes, fs_stream_t stream);

/* ------------------------------------------------------------------------------------------------
 * A1  float descriptor matching — replaces the tail of SuperGlue.forward that mono_slam.py:100-103 and
 *     stereo_slam.py:210-213, :221-224 consume:
 *       scores = einsum('bdn,bdm->bnm', d0, d1)                   superglue.py:259   (dense contraction)
 *       max0, max1 = scores.max(2), scores.max(1) ; mutual check ; threshold ; -1 fill   superglue.py:268-278
 *     plus the Lowe ratio of the north star on L2 distance d = sqrt(max(0, 2 - 2 s)) of unit vectors.
 *
 * Two-step: fs_float_prepare() turns a frame bank of channel-major fp32 descriptors (SuperPoint layout
 * [256, n], superpoint.py:90-91) into the K-major split-fp16 operand bank the tensor-core kernels read
 * (once per FRAME, not per pair); fs_float_match() then runs ONE fp16 tcgen05 pass per frame pair with a
 * fused epilogue (row top-3 keys; column side of the mutual check as packed 16-bit maxima per slab of 16
 * rows), certifies every arg-max / mutual / threshold / ratio decision against a rigorous error bound
 * (rescoring exactly, in fp32 on hi + lo, whatever the fp16 pass cannot prove) and applies the mutual /
 * threshold / ratio finalisation.  Index outputs are those of the exact score.
 *
 * desc_f32   [n_frames, 256, n_max] f32   channel-major
 * operands   [n_frames, n_max + 1, 512] fp16: rows < n_max = (hi | lo) split, x = hi + lo to 2^-25, rows
 *            >= count are zero; row n_max = per-frame norm statistics.  fs_float_operand_bytes() is its size
 * scale      (> 0) multiplies the raw dot product (superglue.py:260 uses 1/sqrt(256) = 0.0625); 1.0 = cosine
 * threshold  keep only if scaled best score > threshold (use -inf to disable)
 * ratio      <=0: off; else keep only if d_best < ratio * d_second (float32, strict), d from raw cosine
 * outputs [n_pairs, n_max]: matches0 i32, score0 f32 (scaled best), second0 f32 (scaled runner-up,
 *   -inf when none), matches1 i32; n_matches [n_pairs].  score0/second0/matches1/n_matches may be NULL.
 *   fs_float_match: score0 / second0 are exact wherever a decision depended on them and within the
 *   bound (~3e-4 for unit descriptors) elsewhere; fs_float_match_split3 / _fp32 give fp32-grade values.
 * diag       NULL or 4 x i32 on the device: [0] rows rescored from the slab bound (or rescanned), [1] rows with two
 *            rescored candidates, [2] rows whose column check rescored slabs, [3] ... with a widened slab search
 * n_max      fs_float_match: <= 4096 (FS_ERR_UNSUPPORTED above; use fs_float_match_split3)
 * ---------------------------------------------------------------------------------------------- */
#define FS_FLOAT_DIM 256
size_t fs_float_operand_bytes(int n_frames, int n_max);
int fs_float_prepare(const float* desc_f32, int n_frames, int n_max, const int32_t* counts,
                     void* operands_fp16, fs_stream_t stream);
size_t fs_float_match_workspace_bytes(int n_pairs, int n_max);        /* fs_float_match_fp32 */
size_t fs_float_match_tc_workspace_bytes(int n_pairs, int n_max);     /* fs_float_match */
size_t fs_float_match_split3_workspace_bytes(int n_pairs, int n_max); /* fs_float_match_split3 */
int fs_float_match(const void* operands_fp16, int n_frames, int n_max, const int32_t* counts,
                   const int32_t* pair_a, const int32_t* pair_b, int n_pairs,
                   float scale, float threshold, float ratio,
                   int32_t* matches0, float* score0, float* second0, int32_t* matches1,
                   int32_t* n_matches, int32_t* diag,
                   void* workspace, size_t workspace_bytes, fs_stream_t stream);
/* the same without the mutual check: BFMatcher(NORM_L2).knnMatch(k = 2) + Lowe ratio as ORB_SLAM3/src/Frame.cc:1144-1151
 * uses it (best / runner-up per query row, strict d_best < ratio * d_second); matches1 is not produced */
int fs_float_match_knn(const void* operands_fp16, int n_frames, int n_max, const int32_t* counts,
                       const int32_t* pair_a, const int32_t* pair_b, int n_pairs,
                       float scale, float threshold, float ratio,
                       int32_t* matches0, float* score0, float* second0, int32_t* n_matches, int32_t* diag,
                       void* workspace, size_t workspace_bytes, fs_stream_t stream);
/* measurement hook: only the tensor-core kernel of fs_float_match (top-3 keys into the workspace).
 * probe = 0: the real kernel (the single-pass form where fs_float_match would run it); bit 0: issue no MMAs; bit 1: skip
 * the top-3 / column network (pipe-overlap experiments); bit 2: the two-direction form;
 * trace: NULL, or 160 x i64 on the device: clock64() at the pipeline's hand-over points of the CTA of pair probe >> 8 */
int fs_float_top3_keys(const void* operands_fp16, int n_frames, int n_max, const int32_t* counts,
                       const int32_t* pair_a, const int32_t* pair_b, int n_pairs, int probe, long long* trace,
                       void* workspace, size_t workspace_bytes, fs_stream_t stream);
/* split-fp16 variant: three MMAs per k-step (hi.hi + lo.hi + hi.lo) give fp32-grade scores directly;
 * row top-2 + column top-1 in one pass.  Slower; any n_max; the in-library cross-check of fs_float_match */
int fs_float_match_split3(const void* operands_fp16, int n_frames, int n_max, const int32_t* counts,
                          const int32_t* pair_a, const int32_t* pair_b, int n_pairs,
                          float scale, float threshold, float ratio,
                          int32_t* matches0, float* score0, float* second0, int32_t* matches1,
                          int32_t* n_matches,
                          void* workspace, size_t workspace_bytes, fs_stream_t stream);
/* exact-fp32 CUDA-core variant of the same operator working directly on desc_f32 (used for n_max that
 * are not a multiple of 128 and as the in-library cross-check of the tensor-core path) */
int fs_float_match_fp32(const float* desc_f32, int n_frames, int n_max, const int32_t* counts,
                        const int32_t* pair_a, const int32_t* pair_b, int n_pairs,
                        float scale, float threshold, float ratio,
                        int32_t* matches0, float* score0, float* second0, int32_t* matches1,
                        int32_t* n_matches,
                        void* workspace, size_t workspace_bytes, fs_stream_t stream);

/* ------------------------------------------------------------------------------------------------
 * A1'   SuperGlue score head after the contraction — replaces
 *       scores = log_optimal_transport(scores, bin_score, iters)          superglue.py:143-172, :263-265
 *       max0 / max1, mutual0 / mutual1, exp, match_threshold, -1 fill      superglue.py:268-285
 *     with the hyper-parameters of mono_slam.py:59-63 (sinkhorn_iterations 20, match_threshold 0.2).
 *
 * scores     [n_pairs, m_max, n_max] f32 row-major, 16-byte aligned, n_max % 4 == 0 and <= 4096; pair p uses
 *            its top-left m_counts[p] x n_counts[p] block (already divided by sqrt(descriptor_dim))
 * bin_score  the dust-bin parameter alpha (superglue.py:220); iters Sinkhorn iterations
 * pairs_per_group  0: as many pairs per cooperative launch as are co-resident in shared memory (the matrix is read
 *            from HBM once and stays on chip for all iterations); > 0: an upper limit; < 0: force the streaming
 *            variant (one launch per iteration, matrices kept in L2) that is otherwise only used when a single
 *            pair's bands exceed the co-resident capacity (measurement / test hook)
 * outputs    matches0 i32 / mscores0 f32 [n_pairs, m_max]; matches1 / mscores1 [n_pairs, n_max]
 *            (-1 / 0 where unmatched, exactly the dictionary of superglue.py:280-285)
 * u_out      NULL or [n_pairs, m_max] f32, v_out NULL or [n_pairs, n_max] f32: the dual potentials of the
 *            ordinary rows / columns after the last iteration (for parity checks of the iteration itself)
 * ---------------------------------------------------------------------------------------------- */
/* scores[p, i, j] = scale * <desc_a[i], desc_b[j]> on the tensor cores (split-fp16, fp32-grade), the matrix
 * superglue.py:259-260 hands to log_optimal_transport: [n_pairs, n_max, n_max] f32, n_max % 128 == 0; rows >= the
 * count of frame a are not written, columns >= the count of frame b are 0 */
int fs_float_scores(const void* operands_fp16, int n_frames, int n_max, const int32_t* counts,
                    const int32_t* pair_a, const int32_t* pair_b, int n_pairs, float scale, float* scores,
                    fs_stream_t stream);
size_t fs_sinkhorn_workspace_bytes(int n_pairs, int m_max, int n_max);
int fs_sinkhorn_match(const float* scores, int n_pairs, int m_max, int n_max, const int32_t* m_counts,
                      const int32_t* n_counts, float bin_score, int iters, float match_threshold,
                      int pairs_per_group, int32_t* matches0, int32_t* matches1, float* mscores0,
                      float* mscores1, float* u_out, float* v_out,
                      void* workspace, size_t workspace_bytes, fs_stream_t stream);

/* ------------------------------------------------------------------------------------------------
 * A3  match gather / order-preserving compaction — replaces
 *       valid = matches > -1 ; mkpts0 = kpts0[valid] ; mkpts1 = kpts1[matches[valid]]
 *     mono_slam.py:106-108, stereo_slam.py:216-218, :227-229 (ascending query index: this order defines
 *     the correspondence indices the RANSAC sampler draws).
 *
 * kpts     [n_frames, n_max, 2] f32 (x, y) pixels
 * matches0 [n_pairs, n_max] i32
 * mkpts0/1 [n_pairs, n_max, 2] f32 (first n_corr[p] rows valid), n_corr [n_pairs] i32
 * src_idx  [n_pairs, n_max] i32 query index of every compacted row; may be NULL
 * ---------------------------------------------------------------------------------------------- */
int fs_gather_matches(const float* kpts, int n_frames, int n_max, const int32_t* counts,
                      const int32_t* pair_a, const int32_t* pair_b, int n_pairs,
                      const int32_t* matches0,
                      float* mkpts0, float* mkpts1, int32_t* n_corr, int32_t* src_idx,
                      fs_stream_t stream);

/* ------------------------------------------------------------------------------------------------
 * A6  stereo "triangulation": disparity -> depth -> pin-hole back-projection -> depth gate — replaces
 *       depth = fx * baseline / disparity_map (0 and -1 -> 0.1)     stereo_slam.py:117-121, :264-271
 *       Z = depth[int(y), int(x)] ; X = (x-cx)/fx*Z ; Y = (y-cy)/fy*Z  stereo_slam.py:274-284
 *       valid = (Z > 0.1) & (Z < 1000) ; filter points3D and mkpts1    stereo_slam.py:287-289
 *
 * pts        [n_points, 2] f32
 * disparity  either per point [n_points] f32 (map_h == 0) or dense maps [n_maps, map_h, map_w] f32
 *            looked up at (int(y), int(x)) of map `map_index[i]` (NULL: map 0); disparities are the
 *            raw SGBM/16 values: exactly 0 and exactly -1 are replaced by 0.1 as the reference does
 * points3d   [n_points, 3] f32,  valid [n_points] u8 (1 inside the gate)
 * Arithmetic is float32 in the reference's operation order (NumPy-1.x value-based casting keeps the
 * f32 keypoints f32); SURVEY.md §8a row A6.
 * ---------------------------------------------------------------------------------------------- */
typedef struct FSCamera {
    double fx, fy, cx, cy;
    double dist[5]; /* k1 k2 p1 p2 k3 (stereo_slam.py:50) */
} FSCamera;

int fs_backproject_stereo(const float* pts, const float* disparity, const int32_t* map_index,
                          int64_t n_points, int map_h, int map_w,
                          const FSCamera* cam_host, double baseline, double z_min, double z_max,
                          float* points3d, uint8_t* valid, fs_stream_t stream);
/* batched + fused filter (stereo_slam.py:288-289): keeps rows with valid depth, in order, for both the
 * 3-D points and the partner 2-D points; out arrays [n_pairs, n_max, 3|2], n_out [n_pairs]. */
int fs_backproject_filter(const float* mkpts0, const float* mkpts1, const float* disparity,
                          const int32_t* n_corr, int n_pairs, int n_max,
                          const FSCamera* cam_host, double baseline, double z_min, double z_max,
                          float* points3d_out, float* pts2d_out, int32_t* n_out, fs_stream_t stream);

/* the same over a FRAME BANK, with the lookup the reference does (stereo_slam.py:274-281): `disparity` is either the
 * per-keypoint bank [n_frames, n_max] f32 (map_h == 0; correspondence i of pair p reads frame pair_a[p] at query index
 * src_idx[p, i], the index fs_gather_matches reports) or the frames' dense maps [n_frames, map_h, map_w] f32 (SGBM/16,
 * looked up at (int(y), int(x)) of the previous-left keypoint; outside the map -> invalid). */
int fs_backproject_filter_bank(const float* mkpts0, const float* mkpts1, const int32_t* n_corr,
                               int n_pairs, int n_max, const float* disparity, int n_frames, int map_h, int map_w,
                               const int32_t* pair_a, const int32_t* src_idx,
                               const FSCamera* cam_host, double baseline, double z_min, double z_max,
                               float* points3d_out, float* pts2d_out, int32_t* n_out, fs_stream_t stream);

/* distance gate on a match list (in place): matches0[p, q] = -1 where dist0[p, q] > max_distance, n_matches recounted —
 * the `bestDist <= TH` test of ORB_SLAM3/src/ORBmatcher.cc:704; the loop-closure driver (BASELINE config 5) counts only
 * such matches, because crossCheck alone pairs ~n/2 random descriptors of two unrelated keyframes. */
int fs_filter_matches_distance(int32_t* matches0, const int32_t* dist0, int n_pairs, int n_max, int max_distance,
                               int32_t* n_matches, fs_stream_t stream);

/* per-pair result record of the one gather of the path: records [n_pairs, 20] f64 =
 * n_matches, n_inliers, status, n_good (0 when NULL), R[9], t[3], iters (0 when NULL), 3 x 0 — what
 * mono_slam.py:113-118 / stereo_slam.py:298-306 keep of a frame pair. */
int fs_pack_records(const int32_t* n_matches, const int32_t* n_inliers, const int32_t* status,
                    const int32_t* n_good, const double* R, const double* t, const int32_t* iters,
                    int n_pairs, double* records, fs_stream_t stream);

/* ------------------------------------------------------------------------------------------------
 * A4  essential-matrix RANSAC — replaces
 *       E, mask = cv2.findEssentialMat(mkpts0, mkpts1, focal=K0[0,0], pp=(K0[0,2],K0[1,2]),
 *                                      method=cv2.RANSAC, prob=0.999, threshold=1.0)   mono_slam.py:111
 *     OpenCV semantics reproduced: points normalised by one focal and pp; thr /= focal; sampler =
 *     cv::RNG(2^64-1) MWC, 5 distinct indices per iteration; Nister 5-point (<=10 candidates, solver
 *     order); Sampson error cast to float32, inlier iff <= (float)thr^2; best updated on strictly
 *     greater count (and > 4); iteration cap shrinks adaptively; no refit.
 *
 * mkpts0/1 [n_pairs, n_max, 2] f32, n_corr [n_pairs]
 * E        [n_pairs, 9] f64 row-major, mask [n_pairs, n_max] u8 in {0,1},
 * n_inliers [n_pairs] i32, iters_run [n_pairs] i32 (may be NULL), status [n_pairs] i32
 * ---------------------------------------------------------------------------------------------- */
size_t fs_find_essential_workspace_bytes(int n_pairs, int n_max, int max_iters);
int fs_find_essential_ransac(const float* mkpts0, const float* mkpts1, const int32_t* n_corr,
                             int n_pairs, int n_max,
                             double focal, double pp_x, double pp_y,
                             double prob, double threshold, int max_iters,
                             double* E, uint8_t* mask, int32_t* n_inliers, int32_t* iters_run,
                             int32_t* status,
                             void* workspace, size_t workspace_bytes, fs_stream_t stream);

/* the sampler alone (OpenCV RANSACPointSetRegistrator::getSubset replay): subsets [n_pairs, max_iters, 5] */
int fs_ransac_subsets(const int32_t* n_corr, int n_pairs, int max_iters, int32_t* subsets,
                      fs_stream_t stream);

/* hypothesis scoring sweeps (BASELINE config 4: 1k..64k hypotheses per frame pair, one warp per
 * hypothesis, correspondences staged in shared memory):
 *   models_E [n_pairs, n_hyp, 9] f64 -> counts [n_pairs, n_hyp] i32 (Sampson, normalised points)
 *   models_rt [n_pairs, n_hyp, 6] f64 (rvec | tvec) -> counts (reprojection with k1,k2,p1,p2,k3)   */
int fs_score_essential(const float* mkpts0, const float* mkpts1, const int32_t* n_corr,
                       int n_pairs, int n_max, double focal, double pp_x, double pp_y, double threshold,
                       const double* models_E, int n_hyp, int32_t* counts, fs_stream_t stream);
int fs_score_pnp(const float* points3d, const float* pts2d, const int32_t* n_corr,
                 int n_pairs, int n_max, const FSCamera* cam_host, double reproj_error,
                 const double* models_rt, int n_hyp, int32_t* counts, fs_stream_t stream);

/* ------------------------------------------------------------------------------------------------
 * A5  pose from E — replaces
 *       _, R, t, _ = cv2.recoverPose(E, mkpts0, mkpts1, focal=..., pp=...)           mono_slam.py:112
 *     decomposeEssentialMat -> 4 (R,t) combos -> DLT triangulation of ALL points (no mask is passed by
 *     the reference) -> cheirality / distance (50) gates -> arg-max (first wins).
 * R [n_pairs, 9] f64, t [n_pairs, 3] f64 (unit norm), mask [n_pairs, n_max] u8 in {0,255},
 * n_good [n_pairs] i32.  Pairs whose status_in (may be NULL) is not FS_PAIR_OK are skipped (R = I, t = 0).
 * workspace: fs_recover_pose_workspace_bytes(n_pairs) bytes (candidate poses + counters per pair).
 * ---------------------------------------------------------------------------------------------- */
size_t fs_recover_pose_workspace_bytes(int n_pairs);
int fs_recover_pose(const double* E, const float* mkpts0, const float* mkpts1, const int32_t* n_corr,
                    const int32_t* status_in, int n_pairs, int n_max,
                    double focal, double pp_x, double pp_y, double distance_thresh,
                    double* R, double* t, uint8_t* mask, int32_t* n_good,
                    void* workspace, size_t workspace_bytes, fs_stream_t stream);

/* fs_recover_pose_static — fs_recover_pose without a caller workspace and without any allocation.
 * NO COUNTERPART IN THE REFERENCE: added for callers that cannot call cudaMalloc
 * (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  The per-pair scratch (candidate
 * poses + cheirality counters, 192 B per pair) lives in a file-scope `static __device__ __align__(256)` array of
 * FS_RECOVER_POSE_STATIC_MAX_PAIRS entries (768 KB, held by every GPU that loads the library until the process exits);
 * its device address is resolved once per device with cudaGetSymbolAddress.  Same arguments, outputs and bits as
 * fs_recover_pose.  n_pairs > FS_RECOVER_POSE_STATIC_MAX_PAIRS -> FS_ERR_BAD_ARG (use fs_recover_pose).  Calls no
 * cudaMalloc / cudaFree / allocating library routine.  Every call on a device shares the array (the kernels rewrite
 * what they read, nothing needs clearing between calls): calls that OVERLAP on the GPU (different streams) race. */
#define FS_RECOVER_POSE_STATIC_MAX_PAIRS 4096
int fs_recover_pose_static_max_pairs(void);
int fs_recover_pose_static(const double* E, const float* mkpts0, const float* mkpts1, const int32_t* n_corr,
                           const int32_t* status_in, int n_pairs, int n_max,
                           double focal, double pp_x, double pp_y, double distance_thresh,
                           double* R, double* t, uint8_t* mask, int32_t* n_good, fs_stream_t stream);

/* ------------------------------------------------------------------------------------------------
 * A7  PnP RANSAC — replaces
 *       _, rvec, tvec, inliers = cv2.solvePnPRansac(points3D, mkpts1_l, K0, dist_coeffs_l,
 *             reprojectionError=1.0, confidence=0.99, iterationsCount=1000,
 *             flags=cv2.SOLVEPNP_ITERATIVE)                                      stereo_slam.py:294-295
 *     OpenCV semantics: inputs as float32; same sampler as A4; EPnP on each 5-subset (points
 *     undistorted first); projectPoints with distortion -> float32 -> squared error <= thr^2 in
 *     float32; adaptive stop; final Levenberg-Marquardt refit (SOLVEPNP_ITERATIVE) on the inlier set.
 * points3d [n_pairs, n_max, 3] f32, pts2d [n_pairs, n_max, 2] f32, n_corr [n_pairs]
 * rvec, tvec [n_pairs, 3] f64; mask [n_pairs, n_max] u8 {0,1}; inlier_idx [n_pairs, n_max] i32
 * ascending (first n_inliers valid; may be NULL); status as above (FS_PAIR_TOO_FEW when n < min_points,
 * the reference's gate is 6).  confidence in (0, 1] as in cv2; a value > 1 switches the adaptive iteration cap OFF so
 * that all max_iters hypotheses are solved and scored (BASELINE config 4 sweep; not a cv2 mode).
 * ---------------------------------------------------------------------------------------------- */
size_t fs_solve_pnp_workspace_bytes(int n_pairs, int n_max, int max_iters);
int fs_solve_pnp_ransac(const float* points3d, const float* pts2d, const int32_t* n_corr,
                        int n_pairs, int n_max, const FSCamera* cam_host,
                        double reproj_error, double confidence, int max_iters, int min_points,
                        double* rvec, double* tvec, uint8_t* mask, int32_t* inlier_idx,
                        int32_t* n_inliers, int32_t* iters_run, int32_t* status,
                        void* workspace, size_t workspace_bytes, fs_stream_t stream);

/* ------------------------------------------------------------------------------------------------
 * A8  pose algebra — replaces cv2.Rodrigues + [R|t] + T_cum = T_cum @ T (stereo_slam.py:298-306,
 *     mono_slam.py:115-118) and the map-point transform T_cum @ [P;1] (stereo_slam.py:309-311).
 * fs_rodrigues:   rvec [n,3] f64 -> R [n,9] f64
 * fs_compose_trajectory: relative (R [n,9], t [n,3]) -> cumulative T [n,16] f64, T_k = T_0 ... T_k
 *     (sequential by definition; one warp walks the chain), skipping entries with status != 0; T_init_host (16 doubles,
 *     NULL = identity) is read during the call and passed to the kernel by value (no host buffer outlives the call)
 * fs_transform_points: points [n_pairs, n_max, 3] f32 by T [n_pairs, 16] f64 -> out f32
 * ---------------------------------------------------------------------------------------------- */
int fs_rodrigues(const double* rvec, int n, double* R, fs_stream_t stream);
int fs_compose_trajectory(const double* R, const double* t, const int32_t* status, int n,
                          const double* T_init_host, double* T_cum, fs_stream_t stream);
int fs_transform_points(const float* points3d, const int32_t* n_pts, int n_pairs, int n_max,
                        const double* T, float* out, fs_stream_t stream);

/* ------------------------------------------------------------------------------------------------
 * SURVEY 8f rank 4 (second half): guided / windowed ORB matching — replaces
 *       ORBmatcher::SearchForInitialization(F1, F2, vbPrevMatched, vnMatches12, windowSize)
 *     ros_ws/src/ORB_SLAM3/src/ORBmatcher.cc:648-762 (call site src/Tracking.cc:2491-2492: ORBmatcher(0.9, true),
 *     window 100), with Frame::GetFeaturesInArea / PosInGrid / AssignFeaturesToGrid (src/Frame.cc:657-735, :385-416:
 *     64 x 48 grid, candidates in (ix, iy, insertion) order), ORBmatcher::DescriptorDistance (:2058-2074),
 *     TH_LOW = 50, the ratio test, the sequential one-to-one stealing rule (:706-718) and the rotation histogram with
 *     ComputeThreeMaxima (:719-755, :2012-2053), all exactly as the reference executes them.
 * Frame bank: kpts [n_frames, n_max, 2] f32 (undistorted), angle [n_frames, n_max] f32 degrees (may be NULL when
 * check_orientation == 0), octave [n_frames, n_max] i32, desc [n_frames, n_max, 32] u8, counts [n_frames].
 * prev_matched [n_pairs, n_max, 2] f32 in/out (vbPrevMatched: search centres in, matched positions out),
 * matches12 [n_pairs, n_max] i32 (-1 = none), n_matches [n_pairs].  Image bounds = mnMinX .. mnMaxY (Frame.cc:151-160).
 * n_max <= 4096 (FS_ERR_UNSUPPORTED above). */
int fs_orb_search_for_initialization(const float* kpts, const float* angle, const int32_t* octave,
                                     const uint8_t* desc, int n_frames, int n_max, const int32_t* counts,
                                     const int32_t* pair_a, const int32_t* pair_b, int n_pairs,
                                     float* prev_matched, int window_size, float nn_ratio, int check_orientation,
                                     float min_x, float max_x, float min_y, float max_y,
                                     int32_t* matches12, int32_t* n_matches, fs_stream_t stream);

/* SURVEY 8f rank 1 (sparse variant) + rank 4 (row band): stereo correspondence feeding A6 — replaces
 *       Frame::ComputeStereoMatches()      ros_ws/src/ORB_SLAM3/src/Frame.cc:811-981
 *     row-band candidate table (:825-836), Hamming best below TH_HIGH in the disparity range (:858-897), 11 x 11 SAD
 *     sliding window +-5 with parabola sub-pixel fit (:899-945), depth = mbf / disparity (:947-960), median test on the
 *     correlation distance (:964-980).  Pyramid level 0 only (scale factor 1; the octave gates are kept).
 * Left / right: kpts [n_frames, n_max, 2] f32, octave [n_frames, n_max] i32, desc [n_frames, n_max, 32] u8,
 * counts [n_frames]; images [n_frames, rows, cols] u8.  mb = baseline (m), mbf = fx * baseline.
 * Out: u_right, depth [n_frames, n_max] f32 (-1 = no match), sad [n_frames, n_max] i32 (correlation distance, -1 = none).
 * A patch that would leave the image (the reference's cv::Mat::rowRange throws there) gives "no match". */
int fs_orb_stereo_matches(const float* kpts_l, const int32_t* octave_l, const uint8_t* desc_l,
                          const int32_t* counts_l, const float* kpts_r, const int32_t* octave_r,
                          const uint8_t* desc_r, const int32_t* counts_r, int n_frames, int n_max,
                          const uint8_t* image_l, const uint8_t* image_r, int rows, int cols, float mb, float mbf,
                          float* u_right, float* depth, int32_t* sad, fs_stream_t stream);

/* CUDA ray caster for the procedural forest (synthetic image pairs rasterised in CUDA; the reference
 * reads its images from a rosbag, stereo_slam.py:35, :185-205).  T_wc_host: 4x4 row-major camera-to-world (host),
 * trunks [n_trunks, 4] f64 on the device (x, z, radius, height), ground plane y = ground_y (y points down).
 * image [height, width] u8; disparity [height, width] f32 = fx * baseline / Z_cam of the hit (0 for sky), may be NULL:
 * a dense map for the lookup of stereo_slam.py:274-281. */
int fs_render_forest(const double* T_wc_host, const FSCamera* cam_host, double baseline, double ground_y,
                     const double* trunks, int n_trunks, int width, int height, uint8_t* image, float* disparity,
                     fs_stream_t stream);

#ifdef __cplusplus
}
#endif
#endif /* FORESTSLAM_B200_H */
