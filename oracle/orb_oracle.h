/* orb_oracle.h — C API of the CPU ORACLE.
 *
 * TEST INFRASTRUCTURE ONLY.  This library is a CPU restatement of the
 * reference's front-end hot path (ORBextractor / ORBmatcher /
 * Frame::ComputeStereoMatches of zackLiuzz/MY_ORB_SLAM2) and of the OpenCV
 * primitives that path calls.  It exists to CHECK the CUDA implementation and
 * to provide the timed CPU baseline; only tests/, __graft_entry__.smoke() and
 * bench.py's cpu_baseline / --impl reference legs may load it.  The product
 * (my_orb_slam2_b200, liborbfe.so) never links, imports or calls it.
 *
 * PARITY STATUS: pinned.  (1) tests/test_ref_compiled.py holds this restatement
 * bit-equal to the UNMODIFIED reference sources compiled in place
 * (oracle/_ref/libref_orbslam2.so: /root/reference/src/{ORBextractor,Frame,
 * ORBmatcher,...}.cc against the OpenCV stand-in of oracle/ref_shim/) on the
 * full-size configs; (2) the pixel arithmetic that lives in OpenCV
 * (un-vendored, version not pinned by the reference) is pinned to OpenCV 4.13.0:
 * each primitive and the whole extraction are checked bit-for-bit against
 * committed python-cv2 4.13.0 fixtures (tests/golden/, tests/test_oracle_golden.py).
 */
#ifndef ORB_ORACLE_H_
#define ORB_ORACLE_H_

#include "../include/orbfe_types.h"

#ifdef __cplusplus
extern "C" {
#endif

/* flags for orc_extractor_create */
#define ORC_FLAG_LIBM_SINCOS 1 /* use libm cosf/sinf as the reference literally does */
#define ORC_FLAG_TIE_BY_ADDRESS 2 /* quadtree ties by heap address (as-is) instead of creation order */

/* ---- OpenCV primitive restatements ---- */
void orc_resize_linear_u8(const uint8_t* src, int sw, int sh, int sstride, uint8_t* dst, int dw, int dh, int dstride);
void orc_gaussian_blur7_u8(const uint8_t* src, int w, int h, int sstride, uint8_t* dst, int dstride);
/* FAST-9/16 with optional NMS over a ROI; returns number of keypoints (may exceed cap; only cap written). */
int orc_fast9_16(const uint8_t* roi, int cols, int rows, int stride, int threshold, int nms, int cap, int32_t* xs,
                 int32_t* ys, int32_t* scores);
float orc_fast_atan2(float y, float x);
void orc_sincos_deg(float angle_deg, int use_libm, float* c, float* s);
int orc_descriptor_distance(const uint8_t* a, const uint8_t* b);
/* input side: cv::cvtColor *2GRAY and cv::remap INTER_LINEAR with CV_32F maps */
void orc_cvt_gray_u8(const uint8_t* src, int w, int h, int sstride, int channels, int rgb, uint8_t* dst, int dstride);
void orc_remap_linear_u8(const uint8_t* src, int sw, int sh, int sstride, const float* mapx, const float* mapy, int dw,
                         int dh, uint8_t* dst, int dstride);
float orc_ic_angle(const uint8_t* img, int stride, int x, int y);
void orc_orb_descriptor(const uint8_t* blurred, int stride, int x, int y, float angle_deg, int use_libm, uint8_t* desc32);

/* ---- ORBextractor ---- */
typedef struct orc_extractor orc_extractor;
orc_extractor* orc_extractor_create(const orbfe_extractor_params* p, int flags);
void orc_extractor_destroy(orc_extractor* e);
void orc_extractor_info(const orc_extractor* e, orbfe_extractor_info* info);
/* ORBextractor::operator(); returns n keypoints (<= capacity written), <0 on error */
int orc_extract(orc_extractor* e, const uint8_t* image, int width, int height, int stride, orbfe_keypoint* kps,
                uint8_t* desc, int capacity);
/* stage dumps of the last orc_extract call */
int orc_extractor_level_size(const orc_extractor* e, int level, int* w, int* h);
const uint8_t* orc_extractor_level_ptr(const orc_extractor* e, int level, int* stride);
const uint8_t* orc_extractor_blurred_ptr(const orc_extractor* e, int level, int* stride);
/* FAST candidates of a level (coords relative to minBorder, order as appended) */
int orc_extractor_candidates(const orc_extractor* e, int level, int cap, int32_t* xs, int32_t* ys, int32_t* scores);
/* kept keypoints of a level after the quadtree (level coords incl. border), in list order */
int orc_extractor_level_keypoints(const orc_extractor* e, int level, int cap, orbfe_keypoint* kps);
/* standalone quadtree on an arbitrary candidate list (x,y relative to minX/minY) */
int orc_distribute_octtree(const int32_t* xs, const int32_t* ys, const int32_t* scores, int n, int minX, int maxX,
                           int minY, int maxY, int N, int flags, int cap, int32_t* kept_index);

/* ---- Frame::ComputeStereoMatches (src/Frame.cc:496-686) ---- */
int orc_stereo_matches(const orc_extractor* left, const orc_extractor* right, const orbfe_keypoint* kpL,
                       const uint8_t* descL, int nL, const orbfe_keypoint* kpR, const uint8_t* descR, int nR, float mbf,
                       float mb, float* u_right, float* depth);

/* ---- ORBmatcher ---- */
int orc_search_by_projection_map(const orbfe_frame_view* F, const orbfe_mappoints_view* mps, float th, float nnratio,
                                 int32_t* assigned);
int orc_search_by_projection_last(const orbfe_frame_view* cur, const orbfe_camera_pose* cam,
                                  const orbfe_lastframe_view* last, float th, int mono, int check_ori,
                                  int32_t* assigned);
int orc_search_for_triangulation(const orbfe_keyframe_view* kf1, const orbfe_keyframe_view* kf2, const float* F12,
                                 float ex, float ey, int only_stereo, int check_ori, int32_t* pairs /* 2 per match */,
                                 int cap);
/* brute force: per query best (dist<<32|idx) and second-best packed the same way; ties -> lowest index */
void orc_hamming_bruteforce(const uint8_t* query, int nq, const uint8_t* map, int nm, int64_t index_base,
                            uint64_t* best, uint64_t* second);

/* ---- the seven other ORBmatcher members (orb_matcher2.cc) ---- */
int orc_search_by_bow(const orbfe_keyframe_view* k1, const orbfe_keyframe_view* k2, int kf_kf, float nnratio,
                      int check_ori, int32_t* matches);
int orc_search_for_initialization(const orbfe_frame_view* F1, const orbfe_frame_view* F2, float* prev_matched,
                                  int window, float nnratio, int check_ori, int32_t* matches12);
int orc_search_by_projection_reloc(const orbfe_frame_view* cur, const orbfe_projection* pr,
                                   const orbfe_points3d_view* pts, float th, int orb_dist, int check_ori,
                                   int32_t* assigned);
int orc_search_by_projection_sim3(const orbfe_frame_view* kf, const orbfe_projection* pr,
                                  const orbfe_points3d_view* pts, int th, int32_t* assigned);
int orc_fuse(const orbfe_frame_view* kf, const orbfe_projection* pr, const orbfe_points3d_view* pts, float th,
             const float* inv_level_sigma2, int32_t* best_idx);
int orc_search_by_sim3(const orbfe_frame_view* kf1, const orbfe_frame_view* kf2, const orbfe_points3d_view* p1,
                       const orbfe_points3d_view* p2, const orbfe_sim3_pair* s, float th, int32_t* match12);
int orc_search_local_points(const orbfe_frame_view* F, const orbfe_projection* pr, const orbfe_points3d_view* pts,
                            const int32_t* obs, float viewingCosLimit, float th, float nnratio, int32_t* assigned,
                            uint8_t* in_view, float* proj_x, float* proj_y, float* proj_xr, int32_t* level,
                            float* view_cos);
void orc_compute_distinctive_descriptors(const uint8_t* desc, const int32_t* offsets, int n_points, int32_t* best_idx,
                                         uint8_t* best_desc);
/* shared arithmetic probes (include/orb_math_shared.h) for the known-answer tests */
float orc_log_f32(float x);
/* the rest of the mono / RGB-D Frame constructors (src/Frame.cc:429-494, :688-712) */
void orc_undistort_keypoints(const orbfe_keypoint* keys, int n, const orbfe_camera* cam, orbfe_keypoint* keys_un);
void orc_image_bounds(const orbfe_camera* cam, int cols, int rows, float* bounds);
void orc_stereo_from_rgbd(const orbfe_keypoint* keys, const orbfe_keypoint* keys_un, int n, const float* depth, int w,
                          int h, int stride, float mbf, float* u_right, float* depth_out);
int orc_predict_scale(float max_distance, float dist, float log_scale_factor, int nlevels);

/* ---- DBoW2 vocabulary transform (orb_bow.cc) ---- */
typedef struct orc_vocabulary orc_vocabulary;
orc_vocabulary* orc_vocabulary_create(int k, int L, int scoring, int weighting, int n_nodes, const int32_t* parent,
                                      const uint8_t* is_leaf, const uint8_t* desc, const double* weight);
orc_vocabulary* orc_vocabulary_load_text(const char* path);
void orc_vocabulary_destroy(orc_vocabulary* v);
void orc_vocabulary_info(const orc_vocabulary* v, int* k, int* L, int* scoring, int* weighting, int* n_nodes,
                         int* n_words);
int orc_vocabulary_transform(const orc_vocabulary* v, const uint8_t* desc, int n, int levelsup, uint32_t* bow_ids,
                             double* bow_vals, int bow_cap, int32_t* n_bow, uint32_t* fv_nodes, int32_t* fv_offsets,
                             int32_t* fv_features, int fv_cap, int32_t* n_fv, uint32_t* word_of, uint32_t* node_of,
                             double* weight_of);

#ifdef __cplusplus
}
#endif
#endif
