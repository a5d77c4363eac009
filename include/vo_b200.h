/*
 * vo_b200.h — C-ABI of libvo_b200.so: the B200 (sm_100a) two-view visual-odometry hot path.
 *
 * The reference (Samirez/Visual_odometry) has no FFI: its hot path is a handful of cv2 calls
 * made from Python (src/image_pair.py:234,280,304,332; src/visual_slam.py:231,243).  Each entry
 * point below replaces the arithmetic of one of those call sites; the Python mirror of the
 * reference's ImagePair (visual_odometry_b200/image_pair.py) binds them with ctypes.
 *
 * Conventions
 *   - every pointer is a DEVICE pointer unless its name ends in _h (host);
 *   - the caller owns every buffer, including the workspace (query *_workspace_bytes first);
 *   - `stream` is a cudaStream_t passed as void*; calls are asynchronous on it and never
 *     synchronise the device;
 *   - return value: 0 = VO_OK, negative = error (vo_last_error_string() has the text).
 *   - batched layout ("pairs of a sequence"): frames are rows of fixed capacity `n_cap`
 *     keypoints; pair b joins frames pairs[2b] (query, frame1) and pairs[2b+1] (train, frame2).
 */
#ifndef VO_B200_H
#define VO_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define VO_OK 0
#define VO_ERR_INVALID_ARG (-1)
#define VO_ERR_TOO_FEW_POINTS (-2)
#define VO_ERR_CUDA (-3)
#define VO_ERR_WORKSPACE (-4)

#define VO_MATCH_MUTUAL 0 /* BFMatcher(crossCheck=True).match  — src/image_pair.py:234 */
#define VO_MATCH_KNN2_RATIO 1 /* knnMatch(k=2)+Lowe ratio — src/feature_detection.py:20-26 */

int vo_version(void);
const char* vo_last_error_string(void);

/* ---- K1: ORB/Hamming matching (replaces BFMatcher(NORM_HAMMING).match, image_pair.py:234) ----
 * desc  : uint8 [n_frames, n_cap, 32]   256-bit descriptors
 * nkp   : int32 [n_frames]              valid keypoints per frame (<= n_cap)
 * kp    : float32 [n_frames, n_cap, 2]  KeyPoint.pt (may be NULL -> pts not written)
 * pairs : int32 [n_pairs, 2]
 * out   : q,t int32 [n_pairs, n_cap]; dist float32 [n_pairs, n_cap] (integer valued);
 *         t2/dist2 second neighbour (mode KNN2_RATIO; may be NULL);
 *         pts float64 [n_pairs, n_cap, 4] = (x1,y1,x2,y2) of each match (image_pair.py:294-299);
 *         n_out int32 [n_pairs] number of matches, rows sorted by queryIdx.
 * Three kernels behind one entry, identical results (bit-exact): small batches run the popc
 * kernel (csrc/hamming.cu); large ones expand every frame 0..n_frames-1 once into 4-bit (e2m1)
 * operands and rank both directions with tcgen05 kind::mxf4 MMAs, the row operand in TMEM
 * (csrc/hamming_f4.cuh); the int8 kind::i8 form of round 2 (csrc/hamming_tc.cuh) stays as a
 * cross-check.  n_frames = 0 (frame count unknown) keeps the popc kernel.
 * vo_debug_hamming(1 | 2 | 3) forces popc / int8 / fp4.
 */
size_t vo_match_workspace_bytes(int n_frames, int n_pairs, int n_cap);
int vo_match_hamming_batch(const uint8_t* desc, const int32_t* nkp, const float* kp, int n_cap,
                           int n_frames, const int32_t* pairs, int n_pairs, int mode, double ratio,
                           int32_t* q, int32_t* t, float* dist, int32_t* t2, float* dist2,
                           double* pts, int32_t* n_out, void* ws, size_t ws_bytes, void* stream);
int vo_debug_hamming(int flags);
/* debug: raw fp32 accumulator of the fp4 kernel (csrc/hamming_f4.cuh) for the pair pairs[0..1] of
 * desc uint8 [2, n_cap, 32]: acc float32 [128, 208] = rows 0..127 x columns 0..207,
 * 36 * (256 - 2 * Hamming) + (207 - column) / 4 when the contraction is right */
int vo_debug_hamming_f4_acc(const uint8_t* desc, const int32_t* nkp, int n_cap, const int32_t* pairs,
                            float* acc, void* ws, size_t ws_bytes, void* stream);
/* single pair convenience: d1 [n1,32], d2 [n2,32]; outputs sized n1 */
int vo_match_hamming(const uint8_t* d1, int n1, const uint8_t* d2, int n2, int mode, double ratio,
                     int32_t* q, int32_t* t, float* dist, int32_t* t2, float* dist2,
                     int32_t* n_out, void* ws, size_t ws_bytes, void* stream);

/* no-allocation variant of vo_match_hamming.  It has NO counterpart in the reference and was
 * added for callers that cannot call cudaMalloc (code
 * under CUDA graph capture, tight serving loops, harnesses that forbid it): the scratch of the
 * match (132 bytes per key point) lives in a file-scope device array of the library
 * (1 081 344 bytes per GPU, held until the process exits) instead of a caller workspace.
 * Limit: n1, n2 <= vo_match_hamming_static_max_keypoints() = 8192; larger sizes are rejected with
 * VO_ERR_WORKSPACE.  Same arguments, same results as vo_match_hamming without ws / ws_bytes; it
 * calls no cudaMalloc / cudaFree / allocating routine.  Every call shares the one array (a call
 * finds the previous call's contents and re-initialises what it reads): calls on one stream are
 * independent, calls that OVERLAP on the GPU (two streams, two threads) race. */
int vo_match_hamming_static_max_keypoints(void);
int vo_match_hamming_static(const uint8_t* d1, int n1, const uint8_t* d2, int n2, int mode,
                            double ratio, int32_t* q, int32_t* t, float* dist, int32_t* t2,
                            float* dist2, int32_t* n_out, void* stream);

/* ---- K6: SIFT/L2 matching on tensor cores (BFMatcher(NORM_L2), visual_slam.py:19) ----
 * desc float32 [n_frames, n_cap, 128]; same outputs; dist = float32 sqrt(sum (a-b)^2).
 * Every frame 0..n_frames-1 is converted once (bf16 tiles + norms) and shared by its pairs.
 * Bit-exact for integer-valued descriptors (cv2 SIFT: exact in bf16, dot products < 2^24).
 * Other float data (RootSIFT ...): the bf16 GEMM proposes the two best candidates per row (both
 * directions for the cross-check) and an fp32 rerank orders them with cv2's formula; a neighbour
 * that bf16 rounding (2^-9 relative per element) ranks third or lower is missed, i.e. the matcher
 * is an APPROXIMATE nearest-neighbour search for non-integer descriptors (measured: identical to
 * cv2 on > 99.9 % of the matches of near-tied float data, tests/test_gpu_match_l2.py). */
size_t vo_match_l2_workspace_bytes(int n_frames, int n_pairs, int n_cap);
int vo_match_l2_batch(const float* desc, const int32_t* nkp, const float* kp, int n_cap, int dim,
                      int n_frames, const int32_t* pairs, int n_pairs, int mode, double ratio,
                      int32_t* q, int32_t* t, float* dist, int32_t* t2, float* dist2,
                      double* pts, int32_t* n_out, void* ws, size_t ws_bytes, void* stream);

/* ---- K2/K3: essential matrix 5-point RANSAC (cv2.findEssentialMat, image_pair.py:280-286) ----
 * pts   : float64 [n_pairs, n_cap, 4] pixel correspondences, m[b] of them valid
 * K     : float64 [9] row-major camera matrix (device)
 * Replays cv::RNG(2^64-1) minimal samples, OpenCV's strict-greater best update and adaptive
 * iteration bound (max_iters; adaptive=0 -> exactly max_iters hypotheses sets, config C3).
 * out   : E float64 [n_pairs,9] (||E||_F = 1), mask uint8 [n_pairs,n_cap] in {0,1},
 *         n_inl int32 [n_pairs], n_iters int32 [n_pairs] (iterations OpenCV would have run),
 *         status int32 [n_pairs] (0 ok, VO_ERR_TOO_FEW_POINTS when m<5 / no model).
 */
size_t vo_essential_workspace_bytes(int n_pairs, int n_cap, int max_iters);
int vo_essential_ransac_batch(const double* pts, const int32_t* m, int n_cap, int n_pairs,
                              const double* K, double conf, double thr_px, int max_iters,
                              int adaptive, double* E, uint8_t* mask, int32_t* n_inl,
                              int32_t* n_iters, int32_t* status, void* ws, size_t ws_bytes,
                              void* stream);
/* the round schedule vo_essential_ransac_batch uses for these sizes (iterations solved and scored
 * per round; 4 kernels per round): chunks_h int [max_rounds] on the HOST, returns the number of
 * rounds.  For launch accounting on the host side. */
int vo_ransac_round_schedule(int max_iters, int n_pairs, int adaptive, int* chunks_h, int max_rounds);
/* five-point minimal solver alone: samples int32 [n_samples,5] index into pts_n [m,4]
 * (normalised coords). out E [n_samples,10,9], n_models int32 [n_samples]. */
int vo_fivept_solve(const double* pts_n, int m, const int32_t* samples, int n_samples,
                    double* E, int32_t* n_models, void* stream);
/* Sampson scoring alone (EMEstimatorCallback::computeError restated): E [H,9];
 * pts_n [m,4] normalised; thr2 = float32 squared threshold; counts int32 [H];
 * masks uint8 [H,m] or NULL. */
int vo_sampson_score(const double* E, int H, const double* pts_n, int m, float thr2,
                     int32_t* counts, uint8_t* masks, void* stream);
/* cv::RNG replay of RANSACPointSetRegistrator::getSubset: samples int32 [n_iters,5] */
int vo_ransac_samples(int count, int n_iters, int32_t* samples, void* stream);

/* ---- K5: recoverPose (cv2.recoverPose, image_pair.py:304-308) ----
 * uses the inliers of `mask` (mask may be NULL = all m[b] points). dist_thresh = 50.
 * out R [n_pairs,9], t [n_pairs,3], pose_mask uint8 [n_pairs,n_cap] in {0,255} (may be NULL),
 * n_good int32 [n_pairs]. */
size_t vo_pose_workspace_bytes(int n_pairs, int n_cap);
int vo_recover_pose_batch(const double* E, const double* pts, const int32_t* m,
                          const uint8_t* mask, int n_cap, int n_pairs, const double* K,
                          double dist_thresh, double* R, double* t, uint8_t* pose_mask,
                          int32_t* n_good, void* ws, size_t ws_bytes, void* stream);

/* ---- K4: DLT triangulation (cv2.triangulatePoints, image_pair.py:332-339) ----
 * P1,P2 float64 [12] row-major 3x4; x1,x2 float64 [n,2]; X float64 [n,4]
 * (unit-norm null vector; dehomog!=0 -> divided by w as image_pair.py:339). */
int vo_triangulate(const double* P1, const double* P2, const double* x1, const double* x2,
                   int n, double* X, int dehomog, void* stream);
/* batched reference flavour: for each pair, inliers (mask) of pts are triangulated with
 * P_first = K [R^T | -R^T t] (paired with frame-1 points), P_second = K [I|0]
 * (image_pair.py:319-336). X [n_pairs,n_cap,4] compacted in match order, n_x [n_pairs]. */
int vo_triangulate_pairs(const double* R, const double* t, const double* pts, const int32_t* m,
                         const uint8_t* mask, int n_cap, int n_pairs, const double* K,
                         double* X, int32_t* n_x, void* stream);

/* ---- trajectory chaining (north_star): P_0 = I, P_k = rel_k * P_{k-1},
 * rel_k = [R_k | t_k] of pair k-1. poses float64 [n_pairs+1,16]. */
int vo_chain_poses(const double* R, const double* t, int n_pairs, const double* P0,
                   double* poses, void* stream);
/* multi-GPU form of the same chaining (the one collective is an all-gather of the
 * per-pair [R|t]): vo_pack_poses writes rt float64 [n_pairs,12] = (R row-major, t) rows straight
 * into the all-gather send buffer; vo_chain_poses_packed chains the gathered rows. */
int vo_pack_poses(const double* R, const double* t, int n_pairs, double* rt, void* stream);
int vo_chain_poses_packed(const double* rt, int n_pairs, const double* P0, double* poses,
                          void* stream);

/* ---- K7: PnP-RANSAC (cv2.solvePnPRansac defaults, visual_slam.py:231-235) ----
 * X [n,3], x [n,2]; out rvec[3], tvec[3], inliers int32 [n], n_inl, ok (device ints).
 * n >= 6: RANSAC over 5-point EPnP samples + iterative refit on the inliers; n == 5: EPnP on all
 * points, n == 4: P3P (three points + the fourth to choose) - OpenCV's solvepnp.cpp takes the
 * plain-solvePnP route when npoints == model_points: no RANSAC, no refit, every point an inlier. */
size_t vo_pnp_workspace_bytes(int n, int iters);
int vo_pnp_ransac(const double* X, const double* x, int n, const double* K, int iters,
                  double reproj_px, double conf, double* rvec, double* tvec, int32_t* inliers,
                  int32_t* n_inl, int32_t* ok, void* ws, size_t ws_bytes, void* stream);
/* batched (BASELINE config C5: "PnP-RANSAC of new frames against the map"): frame f owns rows
 * offsets[f] .. offsets[f+1]-1 of X [n_total,3] / x [n_total,2] (CSR; offsets int32 [n_frames+1]
 * on the device).  One launch group for all frames, no host synchronisation.
 * out: rvec, tvec float64 [n_frames,3]; inliers int32 [n_total]: frame f's inlier indices
 * (frame-local, ascending) start at inliers[offsets[f]]; n_inl, ok int32 [n_frames]
 * (ok = 0: fewer than 4 points or no model). */
size_t vo_pnp_batch_workspace_bytes(long n_total, int n_frames, int iters);
int vo_pnp_ransac_batch(const double* X, const double* x, const int32_t* offsets, long n_total,
                        int n_frames, const double* K, int iters, double reproj_px, double conf,
                        double* rvec, double* tvec, int32_t* inliers, int32_t* n_inl, int32_t* ok,
                        void* ws, size_t ws_bytes, void* stream);

/* ---- N2: reprojection errors of all map observations (Map.calculate_reprojection_error,
 * Map.remove_observations_with_reprojection_errors_above_threshold; reference src/map.py:46-97,
 * called at src/visual_slam.py:181,266) ----
 * points float64 [n_points,3]; poses float64 [n_cams,16] = TrackedCamera.pose() row-major 4x4;
 * obs_cam / obs_pt int32 [n_obs] = ROW indices into poses / points; obs_xy float64 [n_obs,2].
 * sqerr float64 [n_obs] = |dx*dx| + |dy*dy| (NaN for an out-of-range index);
 * keep uint8 [n_obs] = sqerr < threshold (may be NULL); total float64 [1] = sum of sqerr
 * (fixed-order tree sum); n_keep int32 [1] (may be NULL). */
size_t vo_reprojection_workspace_bytes(int n_obs);
int vo_reprojection_errors(const double* points, int n_points, const double* poses, int n_cams,
                           const int32_t* obs_cam, const int32_t* obs_pt, const double* obs_xy,
                           int n_obs, const double* K, double threshold, double* sqerr,
                           uint8_t* keep, double* total, int32_t* n_keep, void* ws,
                           size_t ws_bytes, void* stream);

/* ---- N1: feature-track bookkeeping + map association (VisualSlam.feature_mapper,
 * track_feature_back_in_time, mappointdict, MatchWithMap assembly; reference
 * src/visual_slam.py:46-50, 94-99, 154-156, 183-229) ----
 * A feature id (frame_id, idx) (src/frame_generator.py:34-36) is the flat index frame*n_cap+idx.
 * parent / root / root_point int32 [n_features]: parent = feature_mapper (-1 = none), root = the
 * id track_feature_back_in_time returns, root_point[id] = row of the map point whose feature_id
 * is id (-1 = none).  vo_tracks_update inserts the links of one pair's matches (q,t int32
 * [n_match], mask uint8 or NULL, n_match a DEVICE int32); vo_tracks_match_map emits, in match
 * order, (KeyPoint.pt of frame2 [k,2], map point [k,3], feature id, point row) of every match
 * whose root owns a map point: the inputs of solvePnPRansac (visual_slam.py:222-235). */
int vo_tracks_reset(int32_t* parent, int32_t* root, int32_t* root_point, long n_features,
                    void* stream);
int vo_tracks_update(const int32_t* q, const int32_t* t, const uint8_t* mask,
                     const int32_t* n_match, int frame1, int frame2, int n_cap, int32_t* parent,
                     int32_t* root, void* stream);
int vo_tracks_set_points(const int32_t* feature_ids, const int32_t* point_rows, int n,
                         int32_t* root_point, void* stream);
int vo_tracks_match_map(const int32_t* q, const int32_t* t, const uint8_t* mask,
                        const int32_t* n_match, int frame1, int frame2, int n_cap,
                        const int32_t* root, const int32_t* root_point, const float* kp,
                        const double* map_points, double* img_xy, double* map_xyz, int32_t* feat2,
                        int32_t* point_row, int32_t* n_out, void* stream);

/* ---- N1/N2: map growth and pruning on the flat, device-resident map ----
 * vo_map_extend = the loop of VisualSlam.add_information_to_map (src/visual_slam.py:172-179,
 * with :100-150) and, on an empty map with max_norm = +inf, of initialize_map (:49-52, :69-83).
 * For the masked matches of one pair, in match order: skip if |X| > max_norm (X = row r of the
 * triangulated points, r = rank among the masked matches, x_stride doubles per row); if the
 * track root of (frame2, t[i]) owns a map point append Observation(point, cam2_row, pt2), else
 * append a point (X, feature id frame1*n_cap+q[i]) and two observations (cam1_row pt1, cam2_row
 * pt2).  n_points / n_obs are DEVICE counters, read and advanced by the call; lookups never see
 * the points appended by the same call (the reference builds mappointdict before the loop);
 * publish != 0 writes root_point[] for the new points afterwards (the rebuild at :193-196).
 * *status (device int32) is set to 1 and nothing is committed if cap_points / cap_obs is too small.
 * vo_map_prune_observations = Map.remove_observations_with_reprojection_errors_above_threshold
 * (src/map.py:46-70) given the keep flags of vo_reprojection_errors: stable compaction of the
 * observation rows into a second set of arrays (must not alias); n_out a device int32. */
int vo_map_extend(const int32_t* q, const int32_t* t, const uint8_t* mask, const int32_t* n_match,
                  int frame1, int frame2, int n_cap, const int32_t* root, int32_t* root_point,
                  const float* kp, const double* X, int x_stride, int cam1_row, int cam2_row,
                  double max_norm, double* points, int32_t* point_fid, int32_t* n_points,
                  int cap_points, double* obs_xy, int32_t* obs_cam, int32_t* obs_pt,
                  int32_t* n_obs, int cap_obs, int publish, int32_t* status, void* stream);
size_t vo_map_prune_workspace_bytes(int n_obs);
int vo_map_prune_observations(const uint8_t* keep, int n_obs, const int32_t* obs_cam_in,
                              const int32_t* obs_pt_in, const double* obs_xy_in,
                              int32_t* obs_cam_out, int32_t* obs_pt_out, double* obs_xy_out,
                              int32_t* n_out, void* ws, size_t ws_bytes, void* stream);

/* vo_map_remove_camera = Map.remove_camera_from_map (src/map.py:188-232; what
 * limit_number_of_camera_in_map(18), :301-318, does to the oldest camera): removes camera row
 * cam_row and its observations, every point left with exactly ONE observation (points left with
 * none are kept, as in the reference) and the observations of those points; order is preserved,
 * rows are renumbered, poses / fixed are shifted in place, root_point[] is rebuilt for the new
 * point rows.  Point and observation arrays are compacted into a second set (must not alias);
 * n_points_out / n_obs_out are device int32. */
size_t vo_map_remove_camera_workspace_bytes(int n_points, int n_obs);
int vo_map_remove_camera(int cam_row, double* poses, uint8_t* fixed, int n_cams,
                         const double* points_in, const int32_t* fid_in, int n_points,
                         double* points_out, int32_t* fid_out, int32_t* n_points_out,
                         const int32_t* obs_cam_in, const int32_t* obs_pt_in, const double* obs_xy_in,
                         int n_obs, int32_t* obs_cam_out, int32_t* obs_pt_out, double* obs_xy_out,
                         int32_t* n_obs_out, int32_t* root_point, void* ws, size_t ws_bytes,
                         void* stream);

/* ---- N3: local bundle adjustment (Map.optimize_map, reference src/map.py:104-186: g2o
 * Levenberg + BlockSolverSE3 (Schur complement on the points) + EdgeProjectXYZ2UV + Huber,
 * optimize(iterations)) on the flat map.  poses double [n_cams,16] (row-major 4x4, world->camera)
 * and points double [n_points,3] are updated in place; fixed uint8 [n_cams] (TrackedCamera.fixed);
 * observations as for vo_reprojection_errors; focal / cx / cy as g2o.CameraParameters (ONE focal
 * length); huber_delta = 1 is g2o's RobustKernelHuber default (<= 0: no robust kernel).
 * n_cams <= 24.  log_chi2 / log_lambda / log_trials (HOST arrays [iterations], may be NULL),
 * n_iterations and chi2_initial (HOST, may be NULL) report the run.  Sums run in a fixed order:
 * results are reproducible.  UNLIKE the other entry points this one SYNCHRONISES its stream: the
 * LM accept / reject decision is taken on the host from one 32-byte read-back per trial. */
size_t vo_bundle_adjust_workspace_bytes(int n_cams, int n_points, int n_obs);
int vo_bundle_adjust(double* poses, const uint8_t* fixed, int n_cams, double* points, int n_points,
                     const int32_t* obs_cam, const int32_t* obs_pt, const double* obs_xy, int n_obs,
                     double focal, double cx, double cy, double huber_delta, int iterations,
                     double* log_chi2, double* log_lambda, int32_t* log_trials, int* n_iterations,
                     double* chi2_initial, void* ws, size_t ws_bytes, void* stream);

/* ---- N4: ORB feature extraction (the detector handed to FrameGenerator, reference
 * src/frame_generator.py:25 `detector.detectAndCompute(frame.image, None)`, src/visual_slam.py:16
 * `cv2.ORB_create()`, src/image_and_keypoints.py:8) ----
 * = cv2.ORB_create(nfeatures, scale_factor, nlevels, edge_threshold, 0, 2, HARRIS_SCORE, 31,
 *   fast_threshold).detectAndCompute(gray, None) for a batch of equally sized 8-bit gray frames.
 * images uint8 [n_frames, height, width] (width, height <= 4095);
 * kp     float32 [n_frames, cap_out, 6] = (pt.x, pt.y, size, angle, response, octave);
 * desc   uint8 [n_frames, cap_out, 32];  n_out int32 [n_frames];
 * status int32 [n_frames]: 1 = more than cap_out key points survived (ties at a cut; rows beyond
 *        cap_out dropped) - cap_out >= nfeatures + slack is the caller's choice.
 * Key points are ordered by (octave, y, x); cv2 returns the same SET per octave in the order
 * std::nth_element leaves.  vo_orb_level_info writes the pyramid geometry to HOST arrays
 * [nlevels] (any may be NULL). */
size_t vo_orb_workspace_bytes(int n_frames, int width, int height, int nfeatures,
                              float scale_factor, int nlevels, int cap_out);
int vo_orb_level_info(int width, int height, int nfeatures, float scale_factor, int nlevels,
                      int32_t* w_h, int32_t* h_h, int32_t* nfeat_h, float* scale_h);
int vo_orb_detect_batch(const uint8_t* images, int n_frames, int width, int height, int nfeatures,
                        float scale_factor, int nlevels, int edge_threshold, int fast_threshold,
                        int cap_out, float* kp, uint8_t* desc, int32_t* n_out, int32_t* status,
                        void* ws, size_t ws_bytes, void* stream);
/* FAST stage alone on level 0 (parity hook): scores uint8 [n_frames,height,width] = response + 1
 * of cv2.FastFeatureDetector(threshold, nonmaxSuppression) before suppression (0 = no corner);
 * corners uint32 [n_frames, (width/2+1)*(height/2+1)] = y << 20 | x << 8 | (response + 1) of the
 * corners that survive suppression and the border filter, unordered; n_corners int32 [n_frames];
 * hist int32 [n_frames,256] histogram of their scores. */
int vo_orb_fast_score_map(const uint8_t* images, int n_frames, int width, int height,
                          int fast_threshold, int edge_threshold, uint8_t* scores,
                          uint32_t* corners, int32_t* n_corners, int32_t* hist, void* stream);

/* ---- N4: SIFT extraction, replaces `cv2.SIFT_create()` (/root/reference/src/visual_slam.py:17,
 * src/image_and_keypoints.py:12) + `detectAndCompute(image, None)` (src/frame_generator.py:25,
 * src/image_and_keypoints.py:46) for a batch of equally sized gray frames, with cv2's default
 * parameters (3 octave layers, contrast 0.04, edge 10, sigma 1.6, image doubled) and cv2's
 * `nfeatures` (0 = all key points; n > 0 = KeyPointsFilter::retainBest: every key point whose
 * response is >= the n-th largest one, ties kept — cv2 leaves those in std::nth_element's order,
 * here they stay in the sorted order, so the two are equal as sets).  images uint8 [n_frames, height, width] on the device.
 * kp float32 [n_frames, cap, 6] = (x, y, size, angle, response, octave word — the int cv2 packs
 * octave / layer / sub-layer into, < 2^24 so exact as float), in cv2's output order
 * (KeyPoint12_LessThan, duplicates removed); desc uint8 [n_frames, cap, 128] (cv2 returns the same
 * integers as float32); n_out int32 [n_frames]; status int32 [n_frames] = 1 when more than `cap`
 * key points were found in a frame (that frame's output is incomplete: call again with a larger
 * cap).  pyramid_out: NULL, or float32 [n_frames, sum over octaves of 6 * h_o * w_o] receiving the Gaussian
 * pyramid (6 layers of h_o x w_o per octave, octave after octave) (tests: bit-exact against cv2.GaussianBlur).  vo_sift_octaves writes the octave
 * sizes to HOST arrays (16 entries) and returns the number of octaves (0 = bad size). */
int vo_sift_octaves(int width, int height, int32_t* w_h, int32_t* h_h);
size_t vo_sift_workspace_bytes(int n_frames, int width, int height, int cap);
int vo_sift_detect_batch(const uint8_t* images, int n_frames, int width, int height, int cap,
                         int nfeatures, float* kp, uint8_t* desc, int32_t* n_out, int32_t* status,
                         float* pyramid_out, void* ws, size_t ws_bytes, void* stream);

/* ---- measurement hooks (bench.py): CUDA-event timing of the dominant kernel on its own stream,
 * and the INT/popc issue-rate ceiling K1 is judged against (MEASURED_PEAKS.json has no integer
 * entry; SURVEY.md section 8d asks for a popc microbenchmark on the box).
 * kernel_id: 0 = the Hamming N1 x N2 pass (K1 popc kernel or K1t tcgen05 kernel), 1 = ransac_score_kernel
 * (K2), 2 = l2 tensor-core GEMM (K6), 3 = ransac_solve_kernel (K3), 4 = K1t operand expansion,
 * 5 = recover_pose_kernel (K5), 6 = triangulate_pairs_kernel (K4), 7 = chain_kernel,
 * 8 = match_compact_kernel, 9 = RANSAC control kernels (init / sample / scan / mask),
 * 10 = SIFT pyramid (upsample, blurs, halving), 11 = SIFT extrema / orientation / ordering,
 * 12 = SIFT descriptors.
 * on: 0 = off, 1 = ids 0..4 only (two event records per scope: cheap enough for a timed region),
 * 2 = every scope. */
int vo_profile_enable(int on);
int vo_profile_read(int kernel_id, double* total_ms, int* launches);
int vo_measure_popc_peak(double* popc32_per_s, void* stream);

/* Page-locked host staging memory for the frame uploads of the host-buffer entry (bench.py e2e;
 * PairPipeline with VO_B200_PINNED=wc).  write_combined != 0 asks for cudaHostAllocWriteCombined:
 * the host writes it at full speed but reads it very slowly, and device reads of it are not
 * snooped — an upload-only buffer.  Free with vo_host_free.  No reference counterpart. */
int vo_host_alloc(void** ptr, size_t bytes, int write_combined);
int vo_host_free(void* ptr);

#ifdef __cplusplus
}
#endif
#endif /* VO_B200_H */
