// geom.h — per-level geometry of one extractor handle, computed once on the host
// (src/ORBextractor.cc:410-470, :785-799, :1133-1134) and passed BY VALUE to every
// kernel (no __constant__ state: two handles run concurrently, src/Frame.cc:89-92).
#pragma once
#include <cstddef>
#include <cstdint>

// Device-side assertions of the checked build (`make EXTRA=-DORBFE_CHECKS`, tools/checked_build.sh): every shared-memory
// index and global offset the kernels derive from data is tested against the extent of the array it addresses; a
// failure prints the expression and traps (the CUDA error fails the calling test).  compute-sanitizer is closed on
// this GPU pool, so this build + guard bands around the device arenas (extractor.cu) is the memcheck of the repo.
#ifdef ORBFE_CHECKS
#include <cstdio>
#define ORBFE_DCHECK(c)                                                                                          \
  do {                                                                                                           \
    if (!(c)) {                                                                                                  \
      printf("ORBFE_DCHECK failed: %s  (%s:%d) block (%d,%d) thread %d\n", #c, __FILE__, __LINE__, (int)blockIdx.x, \
             (int)blockIdx.y, (int)threadIdx.x);                                                                 \
      __trap();                                                                                                  \
    }                                                                                                            \
  } while (0)
#else
#define ORBFE_DCHECK(c) do { } while (0)
#endif

namespace orbfe {

constexpr int kMinBorder = 16;   // EDGE_THRESHOLD-3 (src/ORBextractor.cc:785)
constexpr int kMaxLevels = 16;
constexpr int kFastOwnCols = 120;  // k_fast.cu: a strip computes 128 x (rows+2) pixels and owns the inner 120 x rows
#ifndef ORBFE_FAST_OWN_ROWS
#define ORBFE_FAST_OWN_ROWS 14
#endif
constexpr int kFastOwnRows = ORBFE_FAST_OWN_ROWS;

struct LevelGeom {
  int w, h, pitch;          // level image; pitch multiple of 128
  int max_bx, max_by;       // w-16, h-16 (:787-788)
  int n_cols, n_rows;       // cell grid (:796-797); 0 => level has no cells
  int w_cell, h_cell;       // (:798-799)
  int cell_base;            // id of this level's first cell within one image
  int blur_tile_base;       // id of this level's first 128x128 blur tile within one image
  int blur_tile_base_small; // ... first 128x32 blur tile (launches that would not fill the GPU with 128x128 tiles)
  int fast_strip_base;      // id of this level's first FAST strip (k_fast.cu) within one image
  int fast_strips_x;        // strips per strip row (0 => level has no cells)
  int quota;                // mnFeaturesPerLevel[level] (:435-446)
  int n_ini;                // quadtree roots (:543)
  float h_x;                // root width (:545)
  int node_cap;             // max quadtree nodes alive
  int cand_off, cand_cap;   // slice of the per-image candidate array
  int kept_off, kept_cap;   // slice of the per-image kept-keypoint array
  int xtab_off, ytab_off;   // resize coefficient tables (level >= 1)
  int xw_off;               // ... and the per-word records of k_pyramid.cu's word-based kernel (resize_fast levels)
  int resize_fast;          // the 4 pixels of an aligned group take their taps within 9 bytes of an aligned address
  int resize_tile;          // ... and the source window of a 128x32 destination tile fits k_pyramid.cu's shared-memory stage
  float scale;              // mvScaleFactor[level]
  float inv_scale;          // mvInvScaleFactor[level]
  float kp_size;            // (float)(int)(PATCH_SIZE*scale) (:857)
  unsigned long long lvl_base;  // byte offset of image 0 of this level in the pyramid arena
  unsigned lvl_bytes;           // pitch*h
};

struct PyrGeom {
  int nlevels;
  int total_cells;  // per image
  int blur_tiles;   // per image
  int blur_tiles_small;
  int fast_strips;  // per image
  int cand_total;   // per image candidate capacity (sum over levels)
  int kept_total;   // per image kept capacity (sum over levels)
  int out_cap;      // per image output capacity (info.capacity)
  int ini_th, min_th;
  int max_node_cap;
  LevelGeom lv[kMaxLevels];
};

// resize tables (cv::resize INTER_LINEAR)
struct XTab { int sx0, sx1; short a0, a1; };          // per destination column
struct __align__(16) YTab { int sy0, sy1; short b0, b1; int pad; };  // per destination row; one 16-byte load
// per aligned group of 4 destination columns (k_pyramid.cu): ws = 4-byte aligned source column of the group's window;
// byte k of selA / selB = the two PRMT selectors that bring pixel k's taps into the low half-word; a01[k] = a0 | a1 << 16
struct __align__(16) XWord { int ws; unsigned selA, selB, pad; unsigned a01[4]; };

// candidate / kept keypoint record: x | y<<12 | score<<24, x,y relative to minBorder
__host__ __device__ inline uint32_t pack_cand(int x, int y, int score) {
  return (uint32_t)x | ((uint32_t)y << 12) | ((uint32_t)score << 24);
}
__host__ __device__ inline int cand_x(uint32_t c) { return (int)(c & 0xFFFu); }
__host__ __device__ inline int cand_y(uint32_t c) { return (int)((c >> 12) & 0xFFFu); }
__host__ __device__ inline int cand_score(uint32_t c) { return (int)(c >> 24); }

}  // namespace orbfe
