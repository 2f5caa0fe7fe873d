/* orbfe.h — C ABI of liborbfe.so: the B200 (sm_100a) ORB front end.
 *
 * This is the drop-in boundary.  The reference has no FFI: its boundary is
 * the C++ classes ORBextractor / ORBmatcher / Frame.  The
 * compat classes (compat/) keep those signatures and call the functions
 * below; the Python harness binds the same functions with ctypes.  Every
 * entry point names the reference interface it replaces (paths relative to
 * the reference repository).
 *
 * Conventions: plain C types only; all functions return 0 on success and a
 * negative orbfe_status on failure (the reference has no error channel: its
 * calls are void / return a count, SURVEY.md 8(b) "Error conventions").
 * "host" pointers are ordinary (pageable or pinned) CPU memory, "device"
 * pointers are CUDA device memory on the handle's device.  There is no CPU
 * fallback: every function fails with ORBFE_ERR_CUDA when no device works.
 */
#ifndef ORBFE_H_
#define ORBFE_H_

#include <stddef.h>
#include <stdint.h>

#include "orbfe_types.h"

#ifdef __cplusplus
extern "C" {
#endif

typedef enum orbfe_status {
  ORBFE_OK = 0,
  ORBFE_ERR_ARG = -1,      /* bad argument */
  ORBFE_ERR_CUDA = -2,     /* CUDA runtime error (see orbfe_last_error) */
  ORBFE_ERR_CAPACITY = -3, /* an output or internal slab was too small */
  ORBFE_ERR_STATE = -4     /* call order (e.g. stereo before extract) */
} orbfe_status;

/* Text of the last error on the calling thread ("" if none). */
const char* orbfe_last_error(void);
/* Library version / build string (contains "sm_100a"). */
const char* orbfe_version(void);
/* Number of CUDA devices visible (0 when none / no driver). */
int orbfe_device_count(void);

/* ------------------------------------------------------------------ */
/* ORBextractor  (include/ORBextractor.h:45-108, src/ORBextractor.cc)  */
/* ------------------------------------------------------------------ */
typedef struct orbfe_extractor orbfe_extractor;

/* ORBextractor::ORBextractor (src/ORBextractor.cc:410-470).  The image size
 * is fixed per handle (the camera's); max_batch images can be processed by
 * one call.  Each handle owns a CUDA stream and its device arenas, so two
 * handles may be driven concurrently from two host threads like the left /
 * right extractors of src/Frame.cc:89-92.
 * Limits (ORBFE_ERR_ARG with a message otherwise): 1..8 levels, scale factor > 1, every level at most 4096 + 32
 * pixels wide / high and not taller than wide by more than the reference's round(w/h) >= 1 allows (it divides by
 * zero there, :543-545), at most 255 x 255 cells per level, at most ~2550 features on one level (the quadtree of a
 * level lives in one CTA's shared memory: 88 bytes per node; 8 levels at scale 1.2 reach that at nfeatures ~ 11 700)
 * and at most 65 535 keypoints per image. */
int orbfe_extractor_create(const orbfe_extractor_params* params, int width, int height, int max_batch, int device,
                           orbfe_extractor** out);
void orbfe_extractor_destroy(orbfe_extractor* e);
/* The six getters + mnFeaturesPerLevel + umax (include/ORBextractor.h:63-83). */
int orbfe_extractor_get_info(const orbfe_extractor* e, orbfe_extractor_info* info);

/* ORBextractor::operator() (src/ORBextractor.cc:1065-1127) for one image:
 * host image in, host keypoints / descriptors out (caller allocated,
 * `capacity` entries; info.capacity is always enough).  Synchronous. */
int orbfe_extract(orbfe_extractor* e, const uint8_t* image, int row_stride, orbfe_keypoint* keypoints,
                  uint8_t* descriptors, int capacity, int32_t* n_out);

/* Same for `batch` images: image b starts at images + b*image_stride bytes;
 * outputs for image b start at keypoints + b*capacity, descriptors +
 * b*capacity*32; counts[b] keypoints are valid.  Synchronous. */
int orbfe_extract_batch(orbfe_extractor* e, const uint8_t* images, size_t image_stride, int row_stride, int batch,
                        orbfe_keypoint* keypoints, uint8_t* descriptors, int capacity, int32_t* counts);

/* Asynchronous pieces of the same call, for pipelining:
 *   upload  : host (ideally pinned) or device images -> level 0 of the pyramids (device images: any byte
 *             alignment, row_stride >= width, any image_stride; re-pitched by a kernel on the handle's stream)
 *   run     : all kernels, results stay on the device
 *   download: results -> host
 *   sync    : wait for the handle's stream. */
int orbfe_extract_upload(orbfe_extractor* e, const uint8_t* images, size_t image_stride, int row_stride, int batch,
                         int images_on_device);
int orbfe_extract_run(orbfe_extractor* e, int batch);
int orbfe_extract_download(orbfe_extractor* e, int batch, orbfe_keypoint* keypoints, uint8_t* descriptors,
                           int capacity, int32_t* counts);
int orbfe_extractor_sync(orbfe_extractor* e);
/* download without waiting: the copies are queued on the handle's stream (use pinned host memory so they overlap other
 * streams' kernels); orbfe_extract_finish waits for them and performs the overflow / capacity checks of the
 * synchronous call.  The host arrays must stay valid until then. */
int orbfe_extract_download_async(orbfe_extractor* e, int batch, orbfe_keypoint* keypoints, uint8_t* descriptors,
                                 int capacity, int32_t* counts);
int orbfe_extract_finish(orbfe_extractor* e);
/* Make `e` issue its work on `other`'s stream from now on (e.g. the left and right extractor of a stereo rig in
 * throughput mode: one in-order stream, no cross-stream events).  `other` must outlive `e`'s use of it. */
int orbfe_extractor_share_stream(orbfe_extractor* e, orbfe_extractor* other);
/* The handle's cudaStream_t (as void*), for event timing by the caller. */
void* orbfe_extractor_stream(orbfe_extractor* e);
/* Kernels launched by this handle since creation. */
int64_t orbfe_extractor_launch_count(const orbfe_extractor* e);

/* Per-stage device timing with CUDA events on the handle's stream (off by default).  After enabling, every
 * orbfe_extract_run / orbfe_stereo_match_run records events around its stages (up to 256 runs); stage_times
 * waits for the stream and returns the summed milliseconds of ms[0..6] = pyramid, FAST, quadtree,
 * blur, orient+describe, stereo match, stereo median cut, and how many runs they cover. */
int orbfe_extractor_set_profiling(orbfe_extractor* e, int enable);
int orbfe_extractor_stage_times(orbfe_extractor* e, float* ms, int32_t* runs, int32_t* stereo_runs);

/* ORBextractor::mvImagePyramid (include/ORBextractor.h:85), device resident:
 * level `level` of image `image_index` of the last run. */
int orbfe_extractor_level(const orbfe_extractor* e, int image_index, int level, const uint8_t** device_ptr,
                          int* pitch, int* width, int* height);
/* ... and a copy of it to the host (dst_stride >= width). */
int orbfe_extractor_download_level(orbfe_extractor* e, int image_index, int level, uint8_t* dst, int dst_stride);

/* Device-resident results of the last run (for chaining without a host trip):
 * keypoints of image b at keypoints + b*capacity, descriptors likewise. */
int orbfe_extractor_device_results(const orbfe_extractor* e, const orbfe_keypoint** keypoints,
                                   const uint8_t** descriptors, const int32_t** counts, int* capacity);

/* Stage dumps of the last run, used by the parity tests only:
 * FAST candidates of a level (src/ORBextractor.cc:801-850; x,y relative to
 * minBorder) in the reference's append order, and the keypoints kept by
 * DistributeOctTree (:539-765) in its list order (level coordinates). */
int orbfe_extractor_debug_candidates(orbfe_extractor* e, int image_index, int level, int cap, int32_t* xs,
                                     int32_t* ys, int32_t* scores, int32_t* n_out);
int orbfe_extractor_debug_level_keypoints(orbfe_extractor* e, int image_index, int level, int cap, int32_t* xs,
                                          int32_t* ys, int32_t* scores, int32_t* n_out);

/* ---- input side of the pipeline on the device ---- */
/* cv::cvtColor(im, im, CV_{RGB,BGR}2GRAY / CV_{RGBA,BGRA}2GRAY) as Tracking::GrabImage* does before the Frame is
 * built (src/Tracking.cc:189-214, :231-244, :259-272), fused into the upload: pixels = batch interleaved 8-bit
 * images of the handle's size with `channels` (3 or 4) bytes per pixel, rgb != 0 when the first byte is red (mbRGB).
 * Replaces orbfe_extract_upload; follow with orbfe_extract_run. */
int orbfe_extract_upload_color(orbfe_extractor* e, const uint8_t* pixels, size_t image_stride, int row_stride,
                               int channels, int rgb, int batch);
/* cv::remap(im, imRect, M1, M2, cv::INTER_LINEAR) of the EuRoC stereo example (Examples/Stereo/stereo_euroc.cc:97-98,
 * :136-137): install the two CV_32F maps of cv::initUndistortRectifyMap once (width x height floats each, the
 * handle's size; they index a src_width x src_height camera image) ... */
int orbfe_extractor_set_rectify_maps(orbfe_extractor* e, const float* map_x, const float* map_y, int src_width,
                                     int src_height);
/* ... then upload raw (distorted) 8-bit gray frames; rectification runs on the device straight into level 0
 * (constant border 0, OpenCV's 1/32-pixel fixed-point bilinear).  Replaces orbfe_extract_upload. */
int orbfe_extract_upload_distorted(orbfe_extractor* e, const uint8_t* pixels, size_t image_stride, int row_stride,
                                   int batch);

/* ------------------------------------------------------------------ */
/* Frame::ComputeStereoMatches  (src/Frame.cc:496-686)                 */
/* ------------------------------------------------------------------ */
/* Runs on the device-resident pyramids, keypoints and descriptors of the last
 * run of `left` and `right` (same batch, same image size, same device);
 * u_right / depth are host arrays with left-capacity entries per image
 * (mvuRight, mvDepth; -1 = no match); n_matched[b] = matches kept.
 * mb is explicit (SURVEY.md 8(c) deviation 2).  Synchronous. */
int orbfe_stereo_match(orbfe_extractor* left, orbfe_extractor* right, int batch, float mbf, float mb,
                       float* u_right, float* depth, int capacity, int32_t* n_matched);
/* Asynchronous variant: results stay on the device (queued on left's stream
 * after both handles' work). */
int orbfe_stereo_match_run(orbfe_extractor* left, orbfe_extractor* right, int batch, float mbf, float mb);
int orbfe_stereo_match_download(orbfe_extractor* left, int batch, float* u_right, float* depth, int capacity,
                                int32_t* n_matched);
/* ... without waiting (complete after orbfe_extractor_sync / orbfe_extract_finish on `left`). */
int orbfe_stereo_match_download_async(orbfe_extractor* left, int batch, float* u_right, float* depth, int capacity,
                                      int32_t* n_matched);
/* The stereo Frame constructor in one call (src/Frame.cc:61-116: the two ORBextractor threads :89-92, the join,
 * ComputeStereoMatches :108): host images in, keypoints / descriptors of both images and mvuRight / mvDepth out.
 * The two extractions run on the handles' own streams, the matcher after both, and the three result downloads are
 * queued behind them without intermediate host synchronisation; returns when everything is on the host.
 * Arrays hold `capacity` entries per image (>= orbfe_extractor_info::capacity); `batch` pairs, images
 * `image_stride` bytes apart (0 = densely packed), rows `row_stride` bytes apart. */
int orbfe_stereo_frame(orbfe_extractor* left, orbfe_extractor* right, const uint8_t* left_images,
                       const uint8_t* right_images, size_t image_stride, int row_stride, int batch, float mbf, float mb,
                       orbfe_keypoint* kp_left, uint8_t* desc_left, int32_t* n_left, orbfe_keypoint* kp_right,
                       uint8_t* desc_right, int32_t* n_right, float* u_right, float* depth, int32_t* n_matched,
                       int capacity);

/* ------------------------------------------------------------------ */
/* ORBmatcher  (include/ORBmatcher.h:45-114, src/ORBmatcher.cc)        */
/* ------------------------------------------------------------------ */
typedef struct orbfe_matcher orbfe_matcher;
/* A matcher context owns a stream + workspace; one per host thread (the
 * reference constructs ORBmatcher objects on the stack per use).  A handle serves one call at a time: a second host
 * thread that enters the same handle while a call is inside it gets ORBFE_ERR_STATE (never a wrong result). */
int orbfe_matcher_create(int device, orbfe_matcher** out);
void orbfe_matcher_destroy(orbfe_matcher* m);
int64_t orbfe_matcher_launch_count(const orbfe_matcher* m);
void* orbfe_matcher_stream(orbfe_matcher* m);

/* ORBmatcher::DescriptorDistance (src/ORBmatcher.cc:1715-1731) for n pairs
 * (host arrays a, b: n x 32 bytes; out: n ints). */
int orbfe_descriptor_distance(orbfe_matcher* m, const uint8_t* a, const uint8_t* b, int n, int32_t* out);

/* ORBmatcher::SearchByProjection(Frame&, const vector<MapPoint*>&, th)
 * (src/ORBmatcher.cc:46-132).  assigned[f] = index of the map point this call
 * wrote into F.mvpMapPoints[f], -1 = untouched.  *n_matches = return value. */
int orbfe_search_by_projection_map(orbfe_matcher* m, const orbfe_frame_view* frame,
                                   const orbfe_mappoints_view* mps, float th, float nnratio, int32_t* assigned,
                                   int32_t* n_matches);

/* ORBmatcher::SearchByProjection(Frame& Cur, const Frame& Last, th, bMono)
 * (src/ORBmatcher.cc:1392-1538) incl. the rotation histogram (:1499-1535).  assigned[f] = row of `last` written into
 * Cur.mvpMapPoints[f]; -1 = untouched; -2 = assigned and then reset to NULL by the rotation histogram (:1527). */
int orbfe_search_by_projection_last(orbfe_matcher* m, const orbfe_frame_view* cur, const orbfe_camera_pose* cam,
                                    const orbfe_lastframe_view* last, float th, int mono, int check_ori,
                                    int32_t* assigned, int32_t* n_matches);

/* ORBmatcher::SearchForTriangulation (src/ORBmatcher.cc:702-872): pairs =
 * (idx1, idx2) sorted by idx1, at most cap written, *n_pairs = total. */
int orbfe_search_for_triangulation(orbfe_matcher* m, const orbfe_keyframe_view* kf1,
                                   const orbfe_keyframe_view* kf2, const float* F12, float ex, float ey,
                                   int only_stereo, int check_ori, int32_t* pairs, int cap, int32_t* n_pairs);

/* ---- the seven other ORBmatcher members ---- */

/* ORBmatcher::SearchByBoW, both overloads.  Feature vectors in CSR form as for
 * SearchForTriangulation; every feature index must appear in at most one node
 * (DBoW2 FeatureVector invariant).  has_mp = "map point present and not bad".
 *  kf_kf = 0: SearchByBoW(KeyFrame* pKF, Frame& F, vpMapPointMatches) (src/ORBmatcher.cc:182-319); kf1 = pKF,
 *             kf2 = F (its has_mp is not read); matches[f], f < kf2->n, = index of the pKF feature whose map point
 *             is written to vpMapPointMatches[f], -1 = NULL.
 *  kf_kf = 1: SearchByBoW(KeyFrame*, KeyFrame*, vpMatches12) (:563-696); matches[idx1], idx1 < kf1->n, = idx2 with
 *             vpMatches12[idx1] = vpMapPoints2[idx2], -1 = NULL.
 * *n_matches = return value. */
int orbfe_search_by_bow(orbfe_matcher* m, const orbfe_keyframe_view* kf1, const orbfe_keyframe_view* kf2, int kf_kf,
                        float nnratio, int check_ori, int32_t* matches, int32_t* n_matches);

/* ORBmatcher::SearchForInitialization (src/ORBmatcher.cc:446-561).  prev_matched = vbPrevMatched as n1 x 2 floats,
 * updated in place (:556-558); matches12 = vnMatches12. */
int orbfe_search_for_initialization(orbfe_matcher* m, const orbfe_frame_view* f1, const orbfe_frame_view* f2,
                                    float* prev_matched, int window, float nnratio, int check_ori,
                                    int32_t* matches12, int32_t* n_matches);

/* ORBmatcher::SearchByProjection(Frame& Cur, KeyFrame* pKF, sAlreadyFound, th, ORBdist) (src/ORBmatcher.cc:1540-1667).
 * pts = pKF->GetMapPointMatches() (valid = present, not bad, not in sAlreadyFound; angle = pKF->mvKeysUn[i].angle);
 * cur->mp_obs[f] >= 0 means Cur.mvpMapPoints[f] != NULL.  assigned[f] = row of pts written to Cur.mvpMapPoints[f],
 * -1 = untouched, -2 = written and reset to NULL by the rotation check (:1659). */
int orbfe_search_by_projection_reloc(orbfe_matcher* m, const orbfe_frame_view* cur, const orbfe_projection* pr,
                                     const orbfe_points3d_view* pts, float th, int orb_dist, int check_ori,
                                     int32_t* assigned, int32_t* n_matches);

/* ORBmatcher::SearchByProjection(KeyFrame* pKF, Scw, vpPoints, vpMatched, th) (src/ORBmatcher.cc:321-434).
 * kf->mp_obs[f] >= 0 means vpMatched[f] != NULL on entry; valid = !isBad() && not in spAlreadyFound.
 * assigned[f] = row written to vpMatched[f], -1 = untouched. */
int orbfe_search_by_projection_sim3(orbfe_matcher* m, const orbfe_frame_view* kf, const orbfe_projection* pr,
                                    const orbfe_points3d_view* pts, int th, int32_t* assigned, int32_t* n_matches);

/* The SEARCH of ORBmatcher::Fuse, both overloads (src/ORBmatcher.cc:879-1029 with pr->is_sim3 = 0, has_ow = 1;
 * :1033-1156 with is_sim3 = 1): best_idx[i] = bestIdx of point i when bestDist <= TH_LOW, else -1.  The search of a
 * point reads nothing that earlier points write; the map surgery that follows it (Replace / AddObservation /
 * AddMapPoint / vpReplacePoint, :1008-1023, :1140-1150) is the caller's, walking best_idx in order.
 * valid = pMP && !isBad() && !IsInKeyFrame(pKF) (resp. not in pKF->GetMapPoints()).  inv_level_sigma2 =
 * pKF->mvInvLevelSigma2 (first overload only). */
int orbfe_fuse(orbfe_matcher* m, const orbfe_frame_view* kf, const orbfe_projection* pr,
               const orbfe_points3d_view* pts, float th, const float* inv_level_sigma2, int32_t* best_idx,
               int32_t* n_fused);

/* ORBmatcher::SearchBySim3 (src/ORBmatcher.cc:1158-1382).  p1 / p2 = GetMapPointMatches() of the keyframes
 * (m == n; valid = present, not bad, not vbAlreadyMatched).  match12[i1] = idx2 of the mutually agreeing pairs
 * (vpMatches12[i1] = vpMapPoints2[idx2]), -1 otherwise.  *n_found = return value. */
int orbfe_search_by_sim3(orbfe_matcher* m, const orbfe_frame_view* kf1, const orbfe_frame_view* kf2,
                         const orbfe_points3d_view* p1, const orbfe_points3d_view* p2, const orbfe_sim3_pair* s,
                         float th, int32_t* match12, int32_t* n_found);

/* Tracking::SearchLocalPoints (src/Tracking.cc:1258-1284) in one call: Frame::isInFrustum
 * (src/Frame.cc:285-349) for every local map point on the device, then SearchByProjection(Frame&, vpMapPoints, th)
 * (src/ORBmatcher.cc:46-132) on the device-resident mTrackProj* fields.  pts->valid = the point is visited by the
 * loop (mnLastFrameSeen != frame id, !isBad()); obs = Observations() per point; pr = mTcw / calibration / mbf /
 * mfLogScaleFactor of the frame (has_ow = 0: mOw = -mRcw^T mtcw as Frame::UpdatePoseMatrices).  Outputs: assigned /
 * n_matches as orbfe_search_by_projection_map; per point in_view = isInFrustum() (the caller's IncreaseVisible) and,
 * where not NULL, mTrackProjX / mTrackProjY / mTrackProjXR / mnTrackScaleLevel / mTrackViewCos. */
int orbfe_search_local_points(orbfe_matcher* m, const orbfe_frame_view* frame, const orbfe_projection* pr,
                              const orbfe_points3d_view* pts, const int32_t* obs, float viewing_cos_limit, float th,
                              float nnratio, int32_t* assigned, int32_t* n_matches, uint8_t* in_view, float* proj_x,
                              float* proj_y, float* proj_xr, int32_t* level, float* view_cos);

/* MapPoint::ComputeDistinctiveDescriptors (src/MapPoint.cc:252-318), batched over the map points a new keyframe
 * touches.  Point p observes desc[offsets[p] .. offsets[p+1]) (32 bytes each, in the iteration order
 * of its mObservations with bad keyframes left out); best_idx[p] = index WITHIN the point's list of the descriptor
 * with the least median Hamming distance to the rest (first one on ties), -1 for an empty list (the reference then
 * keeps the old descriptor); best_desc (optional, n_points x 32) = that descriptor, zeros for -1. */
int orbfe_compute_distinctive_descriptors(orbfe_matcher* m, const uint8_t* desc, const int32_t* offsets, int n_points,
                                          int32_t* best_idx, uint8_t* best_desc);

/* Brute-force Hamming (BASELINE.json config 5): per query the two smallest
 * keys (dist << 32 | index_base + j) over the map shard (index_base + nm <= 2^32); ties -> lowest index.
 * Host arrays in/out. */
int orbfe_hamming_bruteforce(orbfe_matcher* m, const uint8_t* query, int nq, const uint8_t* map, int nm,
                             int64_t index_base, uint64_t* best, uint64_t* second);
/* Same on device pointers, asynchronous on the matcher's stream: the per-shard
 * step of the multi-GPU path (the cross-shard exchange is an NCCL all-gather /
 * min all-reduce on best/second done by the caller, then orbfe_hamming_merge). */
int orbfe_hamming_bruteforce_device(orbfe_matcher* m, const uint8_t* d_query, int nq, const uint8_t* d_map, int nm,
                                    int64_t index_base, uint64_t* d_best, uint64_t* d_second);
/* Merge `shards` per-shard (best, second) pairs (device arrays laid out
 * [shard][nq]) into the global top-2 (device arrays [nq]). */
int orbfe_hamming_merge_device(orbfe_matcher* m, const uint64_t* d_best_all, const uint64_t* d_second_all,
                               int shards, int nq, uint64_t* d_best, uint64_t* d_second);
/* The same top-2 keys as orbfe_hamming_bruteforce_device, computed on the tensor cores: 256-bit Hamming distance as a
 * {0,1} int8 GEMM (d = |q| + |m| - 2 q.m; tcgen05.mma kind::i8, TMEM accumulators, TMA operand tiles;
 * my_orb_slam2_b200/csrc/k_match_mma.cu).  OPTIONAL second path of BASELINE.json config 5 — north_star's stated choice
 * (popc) stays the default; bench.py reports both with bit-equal keys.  Asynchronous on the matcher's stream; nm >= 1;
 * scratch grows to nm x 256 bytes (the map expanded to one byte per bit).  orbfe_matcher_sync returns ORBFE_ERR_CUDA if
 * the kernel's internal pipeline timed out. */
int orbfe_hamming_bruteforce_tc_device(orbfe_matcher* m, const uint8_t* d_query, int nq, const uint8_t* d_map, int nm,
                                       int64_t index_base, uint64_t* d_best, uint64_t* d_second);
/* No counterpart in the reference: added for callers that cannot allocate
 * (code under CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc).  Same result as
 * orbfe_hamming_bruteforce_device (device pointers in and out, asynchronous on `stream`, a cudaStream_t; NULL = the
 * legacy default stream) on the CURRENT device, but no matcher handle and NO allocation: the per-slice partial results
 * live in a 4 MiB file-scope static __device__ array (my_orb_slam2_b200/csrc/noalloc.cu) that every GPU loading the
 * library holds until exit.  Limit: nq <= 32768 queries (any map size with index_base + nm <= 2^32); larger calls
 * return ORBFE_ERR_CAPACITY with a message.  All calls on a device share that array: calls that overlap on the GPU
 * (different streams / threads) race, so serialise them.  The first call on each device resolves the array's device
 * address with cudaGetSymbolAddress; make that first call outside graph capture. */
int orbfe_hamming_bruteforce_device_noalloc(void* stream, const uint8_t* d_query, int nq, const uint8_t* d_map, int nm,
                                            int64_t index_base, uint64_t* d_best, uint64_t* d_second);
int orbfe_matcher_sync(orbfe_matcher* m);

/* ---- multi-GPU exchange of the brute-force top-2 (BASELINE.json config 5; SURVEY.md 8(b)
 * `orbfe_hamming_bruteforce(..., comm?, ...)` and 8(e) "map descriptors sharded, NCCL-over-NVLink min-pair reduce") ----
 * One process (or thread) per GPU.  Rank 0 makes an id (orbfe_comm_unique_id = ncclGetUniqueId; libnccl.so.2 is
 * dlopen'ed, path override: ORBFE_NCCL_LIB), hands its 128 bytes to the other ranks by any means (MPI, a file,
 * torch.distributed ...), and every rank calls orbfe_comm_create (collective: ncclCommInitRank, then the ranks map each
 * other's exchange buffers through CUDA IPC when they share a node and peer access exists; ORBFE_COMM_NO_P2P=1
 * keeps the NCCL path).  max_queries bounds nq of later calls (the exchange buffer holds 2 x nranks x 2 x max_queries
 * u64 on every rank). */
typedef struct orbfe_comm orbfe_comm;
#define ORBFE_COMM_ID_BYTES 128
enum { ORBFE_EXCHANGE_AUTO = 0, ORBFE_EXCHANGE_NCCL = 1, ORBFE_EXCHANGE_PEER = 2,
       ORBFE_KERNEL_TENSOR = 16 /* OR into `exchange`: the shard's top-2 by the tensor-core kernel (needs nm_local >= 1) */ };
int orbfe_comm_unique_id(uint8_t* id_out /* ORBFE_COMM_ID_BYTES */);
int orbfe_comm_create(const uint8_t* id, int nranks, int rank, int device, int max_queries, orbfe_comm** out);
void orbfe_comm_destroy(orbfe_comm* c);
/* peer_memory = 1 when the ranks exchange through mapped peer memory (NVLink stores), 0 = NCCL all-gather */
int orbfe_comm_info(orbfe_comm* c, int* nranks, int* rank, int* peer_memory, int* nccl_version);
/* ORBFE_ERR_CUDA (with a message) if a consuming kernel gave up waiting for a peer (a rank did not make the matching
 * call within ORBFE_COMM_TIMEOUT_S seconds, default 30); call after orbfe_matcher_sync. */
int orbfe_comm_status(orbfe_comm* c);
/* The sharded brute force: every rank passes the SAME queries and ITS shard of the map (global row numbers
 * index_base .. index_base + nm_local); on return (asynchronous on the matcher's stream) d_best / d_second hold
 * the GLOBAL top-2 keys (dist << 32 | global index, ties -> lowest index) on EVERY rank.  Collective: all ranks call it
 * in the same order with the same nq and `exchange`.  exchange = ORBFE_EXCHANGE_AUTO picks peer memory when available:
 * the kernel that merges the rank's slices stores its keys into every rank's buffer over NVLink and publishes a flag,
 * and the consuming kernel waits for all flags and takes the top-2 (no collective call); ORBFE_EXCHANGE_NCCL = one
 * in-place ncclAllGather between the two kernels. */
int orbfe_hamming_bruteforce_sharded(orbfe_matcher* m, orbfe_comm* c, const uint8_t* d_query, int nq,
                                     const uint8_t* d_map_shard, int nm_local, int64_t index_base, uint64_t* d_best,
                                     uint64_t* d_second, int exchange);
/* optional per-call device timing (CUDA events on the matcher's stream): brute-force kernel | exchange + merge */
int orbfe_comm_set_timing(orbfe_comm* c, int enable);
int orbfe_comm_last_times(orbfe_comm* c, float* kernel_ms, float* exchange_merge_ms);
/* Measured 32-bit POPC results per second of the matcher's device (register-only probe kernel): the roofline
 * denominator of the Hamming kernels ("measure it"). */
int orbfe_measure_popc_peak(orbfe_matcher* m, double* popc32_per_second);

/* ------------------------------------------------------------------ */
/* The rest of the mono / RGB-D Frame constructors (src/Frame.cc:118-228), */
/* on the resident results of orbfe_extract_run                            */
/* ------------------------------------------------------------------ */
/* Frame::UndistortKeyPoints (src/Frame.cc:429-460): keys_un[b][i] = keypoint i of image b with pt = cv::undistortPoints(pt,
 * mK, mDistCoef, noArray(), mK) (OpenCV 4.13: five fixed-point iterations in double, bit-exact); every other field is
 * copied; when k[0] == 0 the keypoints are copied unchanged (:431-435).  Host output [batch][capacity] (may be NULL: the
 * undistorted keypoints also stay on the device for orbfe_stereo_from_rgbd).  Synchronous. */
int orbfe_undistort_keypoints(orbfe_extractor* h, int batch, const orbfe_camera* cam, orbfe_keypoint* keys_un,
                              int capacity);
/* Frame::ComputeImageBounds (src/Frame.cc:462-494): bounds = {mnMinX, mnMaxX, mnMinY, mnMaxY} of a width x height
 * image (the four undistorted corners, or the image rectangle when k[0] == 0).  Host arithmetic (four points). */
int orbfe_image_bounds(const orbfe_camera* cam, int width, int height, float* bounds);
/* Frame::ComputeStereoFromRGBD (src/Frame.cc:688-712) after orbfe_undistort_keypoints: depth = CV_32F images on the
 * host ([batch] images, image_stride / row_stride in floats; Tracking::GrabImageRGBD has already applied
 * mDepthMapFactor); per keypoint d = depth(int(v), int(u)) at the DISTORTED position, and when d > 0: mvDepth = d,
 * mvuRight = keys_un.x - mbf / d; else both -1.  Host outputs [batch][capacity].  Synchronous. */
int orbfe_stereo_from_rgbd(orbfe_extractor* h, int batch, const float* depth, size_t image_stride, int row_stride,
                           float mbf, float* u_right, float* depth_out, int capacity);

/* ------------------------------------------------------------------ */
/* DBoW2 vocabulary transform                        */
/* Thirdparty/DBoW2/DBoW2/TemplatedVocabulary.h:1127-1262, called from */
/* Frame::ComputeBoW (src/Frame.cc:420-427) and KeyFrame::ComputeBoW   */
/* (src/KeyFrame.cc:59-68)                                             */
/* ------------------------------------------------------------------ */
typedef struct orbfe_vocabulary orbfe_vocabulary;
/* The tree as arrays: node 0 is the root; node i > 0 has parent[i] < i, a 32-byte descriptor, a weight and the
 * "is a word" flag of its text line; word ids count the flagged nodes in order (TemplatedVocabulary.h:1385-1418).
 * scoring / weighting are DBoW2's enums (ScoringType 0..5, WeightingType 0..3). */
int orbfe_vocabulary_create(int device, int k, int L, int scoring, int weighting, int n_nodes, const int32_t* parent,
                            const uint8_t* is_leaf, const uint8_t* desc, const double* weight,
                            orbfe_vocabulary** out);
/* TemplatedVocabulary::loadFromTextFile (TemplatedVocabulary.h:1338-1424), e.g. Vocabulary/ORBvoc.txt. */
int orbfe_vocabulary_load_text(int device, const char* path, orbfe_vocabulary** out);
void orbfe_vocabulary_destroy(orbfe_vocabulary* v);
/* info6 = k, L, scoring, weighting, nodes, words */
int orbfe_vocabulary_info(const orbfe_vocabulary* v, int32_t* info6);
/* transform(features, BowVector&, FeatureVector&, levelsup): desc = n x 32 host bytes.  BowVector: ascending
 * (bow_ids[j], bow_vals[j]), j < *n_bow.  FeatureVector in CSR form (the layout orbfe_keyframe_view takes):
 * fv_nodes[j] ascending, features of node j = fv_features[fv_offsets[j] .. fv_offsets[j+1]) in ascending feature
 * index; fv_offsets has fv_cap + 1 entries, fv_features n.  Capacities of n always suffice.  word_of / node_of /
 * weight_of (optional, n entries) = the single-feature transform (:1216-1262).  Thread safe per handle. */
int orbfe_vocabulary_transform(orbfe_vocabulary* v, const uint8_t* desc, int n, int levelsup, uint32_t* bow_ids,
                               double* bow_vals, int bow_cap, int32_t* n_bow, uint32_t* fv_nodes, int32_t* fv_offsets,
                               int32_t* fv_features, int fv_cap, int32_t* n_fv, uint32_t* word_of, uint32_t* node_of,
                               double* weight_of);
/* Same on device-resident descriptors (orbfe_extractor_device_results of a finished run, same device): ComputeBoW
 * without a host trip of the descriptors.  Outputs are host arrays. */
int orbfe_vocabulary_transform_device(orbfe_vocabulary* v, const uint8_t* d_desc, int n, int levelsup,
                                      uint32_t* bow_ids, double* bow_vals, int bow_cap, int32_t* n_bow,
                                      uint32_t* fv_nodes, int32_t* fv_offsets, int32_t* fv_features, int fv_cap,
                                      int32_t* n_fv);

#ifdef __cplusplus
}
#endif
#endif /* ORBFE_H_ */
