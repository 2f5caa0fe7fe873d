/* orbfe_types.h — plain-C data types shared by the C ABI (orbfe.h), the CUDA
 * implementation and the CPU oracle.  No OpenCV / torch types.
 *
 * Every struct mirrors a set of members of a reference class; the reference
 * location is cited next to each one (paths relative to the reference repo).
 */
#ifndef ORBFE_TYPES_H_
#define ORBFE_TYPES_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define ORBFE_MAX_LEVELS 16
#define ORBFE_DESC_BYTES 32
/* Frame grid (include/Frame.h:37-38) */
#define ORBFE_GRID_COLS 64
#define ORBFE_GRID_ROWS 48
/* Matcher constants (src/ORBmatcher.cc:37-39) */
#define ORBFE_TH_HIGH 100
#define ORBFE_TH_LOW 50
#define ORBFE_HISTO_LENGTH 30

/* Layout-compatible with cv::KeyPoint (pt.x, pt.y, size, angle, response,
 * octave, class_id) = 28 bytes, so a std::vector<cv::KeyPoint>::data() can be
 * handed to the C ABI directly (include/ORBextractor.h:59-61). */
typedef struct orbfe_keypoint {
  float x, y;
  float size;
  float angle;
  float response;
  int32_t octave;
  int32_t class_id;
} orbfe_keypoint;

/* Arguments of ORBextractor::ORBextractor (src/ORBextractor.cc:410-413). */
typedef struct orbfe_extractor_params {
  int32_t nfeatures;
  float scale_factor;
  int32_t nlevels;
  int32_t ini_th_fast;
  int32_t min_th_fast;
} orbfe_extractor_params;

/* What the six getters of ORBextractor return (include/ORBextractor.h:63-83)
 * plus the per-level quota mnFeaturesPerLevel (src/ORBextractor.cc:435-446). */
typedef struct orbfe_extractor_info {
  int32_t nlevels;
  float scale_factor;
  float scale_factors[ORBFE_MAX_LEVELS];
  float inv_scale_factors[ORBFE_MAX_LEVELS];
  float level_sigma2[ORBFE_MAX_LEVELS];
  float inv_level_sigma2[ORBFE_MAX_LEVELS];
  int32_t features_per_level[ORBFE_MAX_LEVELS];
  int32_t umax[16];
  int32_t capacity; /* max keypoints one image can return */
} orbfe_extractor_info;

/* The members of Frame that ORBmatcher::SearchByProjection reads
 * (src/ORBmatcher.cc:46-132,1392-1538; src/Frame.cc:351-417).  All pointers
 * are HOST pointers in the host entry points. */
typedef struct orbfe_frame_view {
  int32_t n;                       /* Frame::N */
  const orbfe_keypoint* keys_un;   /* Frame::mvKeysUn */
  const float* u_right;            /* Frame::mvuRight */
  const uint8_t* desc;             /* Frame::mDescriptors, n x 32 */
  const int32_t* mp_obs;           /* per feature: -1 = mvpMapPoints[i]==NULL,
                                      else mvpMapPoints[i]->Observations() */
  float min_x, min_y, max_x, max_y;     /* mnMinX, mnMinY, mnMaxX, mnMaxY */
  float grid_inv_w, grid_inv_h;         /* mfGridElementWidthInv/HeightInv */
  int32_t nlevels;
  const float* scale_factors;           /* Frame::mvScaleFactors */
} orbfe_frame_view;

/* Local map points as seen by SearchByProjection(Frame&, vector<MapPoint*>&)
 * (src/ORBmatcher.cc:46-132; fields of include/MapPoint.h:92-97). */
typedef struct orbfe_mappoints_view {
  int32_t m;
  const float* proj_x;     /* mTrackProjX */
  const float* proj_y;     /* mTrackProjY */
  const float* proj_xr;    /* mTrackProjXR */
  const int32_t* level;    /* mnTrackScaleLevel */
  const float* view_cos;   /* mTrackViewCos */
  const uint8_t* in_view;  /* mbTrackInView && !isBad() */
  const int32_t* obs;      /* Observations() */
  const uint8_t* desc;     /* GetDescriptor(), m x 32 */
} orbfe_mappoints_view;

/* LastFrame as seen by SearchByProjection(Frame&, const Frame&, th, bMono)
 * (src/ORBmatcher.cc:1392-1538). */
typedef struct orbfe_lastframe_view {
  int32_t n;
  const uint8_t* valid;    /* mvpMapPoints[i]!=NULL && !mvbOutlier[i] */
  const float* world_pos;  /* pMP->GetWorldPos(), n x 3 */
  const int32_t* obs;      /* pMP->Observations() */
  const uint8_t* desc;     /* pMP->GetDescriptor(), n x 32 */
  const int32_t* octave;   /* LastFrame.mvKeys[i].octave */
  const float* angle;      /* LastFrame.mvKeysUn[i].angle */
  float tcw[12];           /* LastFrame.mTcw rows 0..2 (3x4, row-major) */
} orbfe_lastframe_view;

/* Camera + pose of the current frame for the last-frame overload. */
typedef struct orbfe_camera_pose {
  float tcw[12]; /* CurrentFrame.mTcw rows 0..2 (3x4 row-major) */
  float fx, fy, cx, cy;
  float mbf, mb;
} orbfe_camera_pose;

/* One keyframe as seen by SearchForTriangulation (src/ORBmatcher.cc:702-872).
 * The DBoW2 FeatureVector (std::map<NodeId, vector<unsigned>>) is given in
 * CSR form: node ids ascending, offsets[n_nodes+1], concatenated indices. */
typedef struct orbfe_keyframe_view {
  int32_t n;
  const orbfe_keypoint* keys_un; /* mvKeysUn */
  const float* u_right;          /* mvuRight */
  const uint8_t* desc;           /* mDescriptors */
  const uint8_t* has_mp;         /* GetMapPoint(i)!=NULL */
  int32_t n_nodes;
  const uint32_t* node_ids;
  const int32_t* node_offsets;
  const int32_t* node_features;
  int32_t nlevels;
  const float* scale_factors;    /* mvScaleFactors */
  const float* level_sigma2;     /* mvLevelSigma2 */
} orbfe_keyframe_view;

/* Map points given by their 3-D position, as the projection searches that do
 * NOT go through Frame::isInFrustum read them (src/ORBmatcher.cc:321-434,
 * 879-1156, 1158-1382, 1540-1667; accessors src/MapPoint.cc:394-444). */
typedef struct orbfe_points3d_view {
  int32_t m;
  const uint8_t* valid;      /* row takes part: pMP!=NULL && !isBad() && not in the
                                method's own skip set (sAlreadyFound, IsInKeyFrame,
                                vbAlreadyMatched) */
  const float* world_pos;    /* GetWorldPos(), m x 3 */
  const float* normal;       /* GetNormal(), m x 3; NULL where the method does not read it */
  const float* max_dist_inv; /* GetMaxDistanceInvariance() (= 1.2f * mfMaxDistance) */
  const float* min_dist_inv; /* GetMinDistanceInvariance() (= 0.8f * mfMinDistance) */
  const float* max_distance; /* mfMaxDistance, the numerator of MapPoint::PredictScale */
  const uint8_t* desc;       /* GetDescriptor(), m x 32 */
  const float* angle;        /* relocalisation overload only: pKF->mvKeysUn[i].angle */
} orbfe_points3d_view;

/* mK and mDistCoef as Tracking builds them from the settings file (src/Tracking.cc:58-100): pinhole intrinsics and the
 * OpenCV distortion coefficients k1, k2, p1, p2, k3 (k3 = 0 when the file has four). */
typedef struct orbfe_camera {
  float fx, fy, cx, cy;
  float k[5];
} orbfe_camera;

/* Pose + calibration of the frame / keyframe the points are projected into. */
typedef struct orbfe_projection {
  float tcw[12];           /* rows 0..2 of Tcw (3x4 row-major), or of the Sim3 Scw when
                              is_sim3 (decomposed as src/ORBmatcher.cc:329-334) */
  int32_t is_sim3;
  int32_t has_ow;          /* 1: ow[] is KeyFrame::GetCameraCenter() (Fuse, :890);
                              0: Ow = -Rcw^T tcw is derived from tcw (:334, :1046, :1546) */
  float ow[3];
  float fx, fy, cx, cy, bf;
  float log_scale_factor;  /* mfLogScaleFactor of the target (MapPoint::PredictScale) */
} orbfe_projection;

/* Everything ORBmatcher::SearchBySim3 reads besides the two keyframes' features
 * and map points (src/ORBmatcher.cc:1158-1177). */
typedef struct orbfe_sim3_pair {
  float t1w[12], t2w[12];  /* pKF1 / pKF2 pose rows 0..2 */
  float s12, r12[9], t12[3];
  float fx, fy, cx, cy;    /* pKF1's intrinsics, used for BOTH directions (:1161-1164) */
  float log_scale_factor1, log_scale_factor2;
} orbfe_sim3_pair;

#ifdef __cplusplus
}
#endif
#endif /* ORBFE_TYPES_H_ */
