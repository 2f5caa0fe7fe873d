// kernels.h — host-side launchers of the sm_100a kernels (one .cu per stage).
#pragma once
#include <cuda_runtime.h>

#include "geom.h"
#include "../../include/orbfe_types.h"

namespace orbfe {

// device arenas of one extractor handle (all sized for max_batch images)
struct ExtractorArena {
  uint8_t* pyr = nullptr;          // level-major: lv[l].lvl_base + b*lv[l].lvl_bytes
  uint8_t* blur = nullptr;         // blurred levels, same layout as pyr
  uint8_t* stage = nullptr;        // [batch][h][w] landing buffer of linear H2D uploads
  XTab* xtab = nullptr;
  XWord* xword = nullptr;  // per-word resize records (levels with resize_fast)
  YTab* ytab = nullptr;
  uint32_t* cand = nullptr;        // [batch][cand_total]
  uint16_t* node_of = nullptr;     // [batch][cand_total]
  int32_t* cand_count = nullptr;   // [batch][kMaxLevels]
  uint8_t* cell_hi = nullptr;      // [batch][total_cells] cell has a FAST maximum with score >= iniThFAST
  uint32_t* kept = nullptr;        // [batch][kept_total]
  int32_t* kept_count = nullptr;   // [batch][kMaxLevels]
  orbfe_keypoint* out_kp = nullptr;  // [batch][out_cap]
  uint8_t* out_desc = nullptr;       // [batch][out_cap][32]
  int32_t* out_count = nullptr;      // [batch]
  int32_t* flags = nullptr;          // [batch] overflow / error bits
  // stereo results (left handle only)
  float* st_uright = nullptr;   // [batch][out_cap]
  float* st_depth = nullptr;    // [batch][out_cap]
  int32_t* st_sad = nullptr;    // [batch][out_cap]
  int32_t* st_count = nullptr;  // [batch]
  int4* st_band = nullptr;             // [batch][out_cap] right keypoint: minr, maxr, x bits, octave
  int32_t* st_bucket_start = nullptr;  // [batch][(rows/8) * nlevels + 1] items sorted by (8-row bucket, octave)
  int4* st_bucket_items = nullptr;     // [batch][out_cap * st_items_per_kp] band + right keypoint index per bucket entry
  int st_items_per_kp = 0;
  // TMA descriptors (CUtensorMap, 128 B each) of the pyramid levels as (x, y, image) byte tensors with a
  // kFastTileW x kFastTileH box: the FAST kernel stages a strip's image tile with ONE cp.async.bulk.tensor
  void* fast_maps = nullptr;  // [kMaxLevels]
  void* desc_maps = nullptr;  // [kMaxLevels] the BLURRED levels with a kDescBoxW x kDescBoxH box (k_describe.cu)
  void* disc_maps = nullptr;  // [kMaxLevels] the un-blurred levels with a kDiscBoxW x kDiscBoxH box (k_describe.cu, IC_Angle)
  void* pyr_maps = nullptr;   // [kMaxLevels] level l as the SOURCE of level l+1: kPyrTileW x kPyrTileH box (k_pyramid.cu)
  // the rest of the mono / RGB-D Frame constructors (frame.cu), allocated on first use
  orbfe_keypoint* kp_un = nullptr;  // [batch][out_cap] undistorted keypoints (Frame::mvKeysUn)
  float* depth_img = nullptr;       // [batch][h][w] CV_32F depth images (ComputeStereoFromRGBD)
  bool kp_un_valid = false;
};

constexpr int kFastTileW = 144;                    // bytes per tile row (multiple of 16): 136 used + 8 of start alignment
constexpr int kFastTileH = kFastOwnRows + 2 + 6;   // computed rows + the FAST ring (+-3)
constexpr int kPyrTileW = 192;  // bytes per staged source row of the resize (12 x 16): the source window of 128 destination columns
constexpr int kPyrTileH = 44;   // staged source rows (32 destination rows)
constexpr int kDiscBoxW = 48;   // bytes: the 31 columns of the IC_Angle disc, from the 16-byte aligned column at or below kx-15
constexpr int kDiscBoxH = 31;   // rows ky-15 .. ky+15 of the UN-blurred level
constexpr int kDescBoxW = 64;   // bytes: the 37 columns a rotated pattern reaches, from the 16-byte aligned column below
constexpr int kDescBoxH = 37;   // rows ky-18 .. ky+18
// one-time host setup of ExtractorArena::fast_maps / desc_maps (extractor.cu; needs the driver's cuTensorMapEncodeTiled)
int configure_describe_kernel();  // k_describe.cu: opt-in dynamic shared memory
int build_fast_tensor_maps(const PyrGeom& g, ExtractorArena& a, int max_batch);

enum : int { FLAG_CAND_OVERFLOW = 1, FLAG_NODE_OVERFLOW = 2, FLAG_OUT_OVERFLOW = 4 };

// ComputePyramid (src/ORBextractor.cc:1129-1154): level `lvl` from level lvl-1 for `batch` images
void launch_pyramid_level(const PyrGeom& g, const ExtractorArena& a, int lvl, int batch, cudaStream_t s);
// cell loop of ComputeKeyPointsOctTree (:776-850): FAST + cell-local NMS; the per-cell threshold fallback is applied
// by the quadtree kernel when it reads the candidates
void launch_fast_cells(const PyrGeom& g, const ExtractorArena& a, int batch, cudaStream_t s);
// DistributeOctTree (:539-765): one CTA per (level, image)
void launch_quadtree(const PyrGeom& g, const ExtractorArena& a, int batch, cudaStream_t s);
size_t quadtree_smem_bytes(int node_cap);
int upload_pattern();          // k_describe.cu: rBRIEF pattern -> device (once per device); 0 on success
int configure_stereo_kernels(const PyrGeom& g);
int configure_fast_kernel();  // opt in to > 48 KB dynamic shared memory; 0 on success
// GaussianBlur(7x7, sigma 2) of every level (:1107-1108), one launch
void launch_blur_levels(const PyrGeom& g, const ExtractorArena& a, int batch, cudaStream_t s);
// computeOrientation (:472-479) + computeDescriptors (:1056-1063) + rescale (:1117-1123)
void launch_describe(const PyrGeom& g, const ExtractorArena& a, int batch, cudaStream_t s);
// Frame::ComputeStereoMatches (src/Frame.cc:496-686)
void launch_stereo(const PyrGeom& g, const ExtractorArena& left, const ExtractorArena& right, int batch, float mbf,
                   float mb, cudaStream_t s, cudaEvent_t* stage_events /* 3 or nullptr */);

constexpr int kLaunchesPerExtract(int nlevels) { return (nlevels - 1) + 3; }
constexpr int kLaunchesPerStereo = 3;
int stereo_items_per_kp(const PyrGeom& g);

// ---- frame.cu: Frame::UndistortKeyPoints / ComputeStereoFromRGBD on the resident keypoints ----
void launch_undistort(const PyrGeom& g, const ExtractorArena& a, int batch, const orbfe_camera& cam, cudaStream_t s);
void launch_stereo_from_rgbd(const PyrGeom& g, const ExtractorArena& a, int batch, int w, int h, float mbf, cudaStream_t s);

// ---- k_input.cu: cvtColor *2GRAY and remap INTER_LINEAR straight into level 0 ----
// rows of w bytes at any alignment (device memory) -> level 0 of the arena
void launch_repitch(const uint8_t* src, size_t image_stride, size_t row_stride, uint8_t* dst, int pitch, size_t dst_image,
                    int w, int h, int batch, cudaStream_t s);
void launch_gray(const uint8_t* src, size_t image_stride, int row_stride, int channels, int rgb, uint8_t* dst, int pitch,
                 size_t dst_image, int w, int h, int batch, cudaStream_t s);
void launch_remap(const uint8_t* src, size_t image_stride, int row_stride, int sw, int sh, const float* mapx,
                  const float* mapy, uint8_t* dst, int pitch, size_t dst_image, int w, int h, int batch, cudaStream_t s);

}  // namespace orbfe
