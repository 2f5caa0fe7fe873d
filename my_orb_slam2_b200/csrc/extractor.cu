// extractor.cu — host side of the ORBextractor entry points of include/orbfe.h:
// constructor tables (reference: src/ORBextractor.cc:410-470), per-level geometry
// (:785-799, :1133-1134), device arenas, and the kernel schedule of operator()
// (:1065-1127).  No CPU computation of results happens here: everything below the
// geometry tables runs in the sm_100a kernels of k_*.cu.
#include <algorithm>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <new>
#include <vector>

#include <cuda.h>

#include "common.cuh"
#include "extractor.h"
#include "kernels.h"

namespace orbfe {

static thread_local char g_err[512] = "";

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

static inline int cv_round(float v) { return (int)nearbyintf(v); }    // cvRound: half to even
static inline int cv_round(double v) { return (int)nearbyint(v); }

// ORBextractor::ORBextractor (src/ORBextractor.cc:410-470)
static void build_info(const orbfe_extractor_params& p, orbfe_extractor_info* info) {
  memset(info, 0, sizeof(*info));
  const int L = p.nlevels;
  const double sf = (double)p.scale_factor;  // member is `double scaleFactor` (include/ORBextractor.h:95)
  info->nlevels = L;
  info->scale_factor = p.scale_factor;
  info->scale_factors[0] = 1.0f;
  info->level_sigma2[0] = 1.0f;
  for (int i = 1; i < L; ++i) {
    info->scale_factors[i] = (float)((double)info->scale_factors[i - 1] * sf);  // :421
    info->level_sigma2[i] = info->scale_factors[i] * info->scale_factors[i];    // :422
  }
  for (int i = 0; i < L; ++i) {
    info->inv_scale_factors[i] = 1.0f / info->scale_factors[i];  // :429
    info->inv_level_sigma2[i] = 1.0f / info->level_sigma2[i];    // :430
  }
  const float factor = (float)(1.0 / sf);  // :435
  float desired = (float)p.nfeatures * (1 - factor) / (1 - (float)pow((double)factor, (double)L));
  int sum = 0;
  for (int lv = 0; lv < L - 1; ++lv) {
    info->features_per_level[lv] = cv_round(desired);  // :441
    sum += info->features_per_level[lv];
    desired *= factor;
  }
  info->features_per_level[L - 1] = p.nfeatures - sum > 0 ? p.nfeatures - sum : 0;  // :446
  // umax (:454-469)
  const int half = 15;
  int* umax = info->umax;
  const int vmax = (int)floor(half * sqrt(2.f) / 2 + 1);
  const int vmin = (int)ceil(half * sqrt(2.f) / 2);
  const double hp2 = (double)half * half;
  for (int v = 0; v <= vmax; ++v) umax[v] = cv_round(sqrt(hp2 - (double)v * v));
  for (int v = half, v0 = 0; v >= vmin; --v) {
    while (umax[v0] == umax[v0 + 1]) ++v0;
    umax[v] = v0;
    ++v0;
  }
}

// cv::resize INTER_LINEAR coefficient tables
static void build_resize_tables(int sw, int sh, int dw, int dh, std::vector<XTab>& xt, std::vector<YTab>& yt) {
  const double scale_x = (double)sw / dw, scale_y = (double)sh / dh;
  for (int x = 0; x < dw; ++x) {
    float fx = (float)((x + 0.5) * scale_x - 0.5);
    int sx = (int)floorf(fx);
    fx -= sx;
    if (sx < 0) { sx = 0; fx = 0.f; }
    if (sx >= sw - 1) { sx = sw - 1; fx = 0.f; }
    XTab t;
    t.sx0 = sx;
    t.sx1 = sx + 1 < sw ? sx + 1 : sw - 1;
    t.a0 = (short)cv_round((1.f - fx) * 2048.f);
    t.a1 = (short)cv_round(fx * 2048.f);
    xt.push_back(t);
  }
  for (int y = 0; y < dh; ++y) {
    float fy = (float)((y + 0.5) * scale_y - 0.5);
    int sy = (int)floorf(fy);
    fy -= sy;
    YTab t;
    t.pad = 0;
    t.b0 = (short)cv_round((1.f - fy) * 2048.f);
    t.b1 = (short)cv_round(fy * 2048.f);
    t.sy0 = sy < 0 ? 0 : (sy > sh - 1 ? sh - 1 : sy);
    t.sy1 = sy + 1 < 0 ? 0 : (sy + 1 > sh - 1 ? sh - 1 : sy + 1);
    yt.push_back(t);
  }
}

static int build_geometry(Extractor* e, std::vector<XTab>& xt, std::vector<YTab>& yt) {
  PyrGeom& g = e->geom;
  memset(&g, 0, sizeof(g));
  const orbfe_extractor_info& info = e->info;
  g.nlevels = info.nlevels;
  g.ini_th = e->params.ini_th_fast < 0 ? 0 : (e->params.ini_th_fast > 255 ? 255 : e->params.ini_th_fast);
  g.min_th = e->params.min_th_fast < 0 ? 0 : (e->params.min_th_fast > 255 ? 255 : e->params.min_th_fast);
  unsigned long long pyr_off = 0;
  int cells = 0, cand = 0, kept = 0, node_cap_max = 8, blur_tiles = 0, blur_tiles_small = 0, fast_strips = 0;
  int xw_total = 0;
  for (int l = 0; l < g.nlevels; ++l) {
    LevelGeom& L = g.lv[l];
    L.scale = info.scale_factors[l];
    L.inv_scale = info.inv_scale_factors[l];
    L.kp_size = (float)(int)(31 * L.scale);  // :857
    L.w = cv_round((float)e->width * L.inv_scale);  // :1133-1134
    L.h = cv_round((float)e->height * L.inv_scale);
    if (L.w < 1 || L.h < 1) { set_error("level %d of a %dx%d image is empty", l, e->width, e->height); return ORBFE_ERR_ARG; }
    if (L.w > 4096 + 2 * kMinBorder || L.h > 4096 + 2 * kMinBorder || L.h > 8000) { set_error("image too large"); return ORBFE_ERR_ARG; }
    L.pitch = (int)align_up((size_t)L.w, 128);
    L.lvl_bytes = (unsigned)L.pitch * (unsigned)L.h;
    L.lvl_base = pyr_off;
    pyr_off += (unsigned long long)L.lvl_bytes * e->max_batch;
    L.max_bx = L.w - kMinBorder;
    L.max_by = L.h - kMinBorder;
    const float width = (float)(L.max_bx - kMinBorder), height = (float)(L.max_by - kMinBorder);
    L.n_cols = width > 0 ? (int)(width / 30.f) : 0;   // :796-797
    L.n_rows = height > 0 ? (int)(height / 30.f) : 0;
    if (L.n_cols <= 0 || L.n_rows <= 0) { L.n_cols = L.n_rows = 0; L.w_cell = L.h_cell = 1; }
    else {
      L.w_cell = (int)ceilf(width / L.n_cols);  // :798-799
      L.h_cell = (int)ceilf(height / L.n_rows);
    }
    if (L.n_rows > 255 || L.n_cols > 255) { set_error("image too large"); return ORBFE_ERR_ARG; }
    L.cell_base = cells;
    cells += L.n_cols * L.n_rows;
    L.blur_tile_base = blur_tiles;
    blur_tiles += ((L.w + 127) / 128) * ((L.h + 127) / 128);
    L.blur_tile_base_small = blur_tiles_small;
    blur_tiles_small += ((L.w + 127) / 128) * ((L.h + 31) / 32);
    L.fast_strip_base = fast_strips;
    L.fast_strips_x = 0;
    if (L.n_cols > 0) {  // detection domain x in [19, w-19), y in [19, h-19); owned columns start at x = 16
      L.fast_strips_x = (L.w - 2 * kMinBorder - 3 + kFastOwnCols - 1) / kFastOwnCols;
      fast_strips += L.fast_strips_x * ((L.h - 2 * (kMinBorder + 3) + kFastOwnRows - 1) / kFastOwnRows);
    }
    L.quota = info.features_per_level[l];
    L.n_ini = 1;
    L.h_x = 1.f;
    int cap = 0;
    if (L.n_cols > 0) {
      L.n_ini = (int)roundf(width / (float)(L.max_by - kMinBorder));  // :543 (float / int)
      if (L.n_ini < 1) { set_error("level %d is taller than wide: the reference divides by zero here", l); return ORBFE_ERR_ARG; }
      L.h_x = width / (float)L.n_ini;  // :545
      // strict 8-neighbour maxima cannot be adjacent: at most ceil(w/2)*ceil(h/2) per detection window
      for (int i = 0; i < L.n_rows; ++i)
        for (int j = 0; j < L.n_cols; ++j) {
          const int x0 = kMinBorder + j * L.w_cell, y0 = kMinBorder + i * L.h_cell;
          if (y0 >= L.max_by - 3 || x0 >= L.max_bx - 6) continue;
          const int cw = std::min(x0 + L.w_cell + 6, L.max_bx) - x0, ch = std::min(y0 + L.h_cell + 6, L.max_by) - y0;
          if (cw < 7 || ch < 7) continue;
          cap += ((cw - 6 + 1) / 2) * ((ch - 6 + 1) / 2);
        }
    }
    L.cand_off = cand;
    L.cand_cap = cap;
    cand += (int)align_up((size_t)cap, 4);
    L.node_cap = std::max(L.quota + 3, 4 * L.n_ini) + 1;
    node_cap_max = std::max(node_cap_max, L.node_cap);
    L.kept_off = kept;
    L.kept_cap = L.node_cap;
    kept += (int)align_up((size_t)L.kept_cap, 4);
    if (l > 0) {
      L.xtab_off = (int)xt.size();
      L.ytab_off = (int)yt.size();
      L.xw_off = xw_total;
      xw_total += (L.w + 3) / 4;
      build_resize_tables(g.lv[l - 1].w, g.lv[l - 1].h, L.w, L.h, xt, yt);
      // k_pyramid.cu's word-based kernel needs the taps of pixels x4..x4+3 at bytes <= 8 of (sx0(x4) & ~3), and b >= 0
      L.resize_fast = 1;
      for (int x4 = 0; x4 < L.w; x4 += 4) {
        const int ws = xt[L.xtab_off + x4].sx0 & ~3;
        for (int k = 0; k < 4 && x4 + k < L.w; ++k) {
          const XTab& t = xt[L.xtab_off + x4 + k];
          if (t.sx0 - ws > 7 || t.sx0 - ws < 0 || (t.a1 != 0 && t.sx1 != t.sx0 + 1) || t.a0 < 0 || t.a1 < 0) L.resize_fast = 0;
        }
      }
      for (int y = 0; y < L.h; ++y)
        if (yt[L.ytab_off + y].b0 < 0 || yt[L.ytab_off + y].b1 < 0) L.resize_fast = 0;
      // the staged variant copies, per 128x32 destination tile, 16-byte chunks [c0, c0 + 16*n) of source rows
      // sy0(first) .. sy1(last): n <= 12 chunks inside the source pitch and <= 44 rows (k_pyramid.cu PT_CHUNKS, PT_ROWS)
      L.resize_tile = L.resize_fast;
      for (int xc = 0; xc < L.w && L.resize_tile; xc += 128) {
        const int c0 = xt[L.xtab_off + xc].sx0 & ~15;
        const int ws_last = xt[L.xtab_off + std::min(xc + 124, (L.w - 1) & ~3)].sx0 & ~3;
        const int n = ((ws_last + 11 - c0) >> 4) + 1;
        if (n > 12 || c0 + 16 * n > g.lv[l - 1].pitch) L.resize_tile = 0;
      }
      for (int yc = 0; yc < L.h && L.resize_tile; yc += 32)
        if (yt[L.ytab_off + std::min(yc + 32, L.h) - 1].sy1 - yt[L.ytab_off + yc].sy0 + 1 > 44) L.resize_tile = 0;
    }
  }
  g.total_cells = cells;
  g.blur_tiles = blur_tiles;
  g.blur_tiles_small = blur_tiles_small;
  g.fast_strips = fast_strips;
  g.cand_total = std::max(cand, 4);
  g.kept_total = std::max(kept, 4);
  g.out_cap = (int)align_up((size_t)kept + 8, 8);
  g.max_node_cap = node_cap_max;
  if (g.out_cap > 65535) { set_error("nfeatures too large (capacity %d > 65535)", g.out_cap); return ORBFE_ERR_ARG; }
  e->pyr_bytes = (size_t)pyr_off;
  e->info.capacity = g.out_cap;
  return ORBFE_OK;
}

#ifdef ORBFE_CHECKS
// Checked build: every arena sits between two 4 KiB guard bands filled with 0xA5; free_arenas() verifies them and
// aborts the process if a kernel wrote outside an arena (the memcheck of this repo: compute-sanitizer is closed on
// the GPU pool).  The arenas are 4 KiB apart instead of 256 B, nothing else changes.
constexpr size_t kGuard = 4096;
struct GuardRec { uint8_t* base; size_t bytes; };
static std::vector<GuardRec>& guard_list() { static std::vector<GuardRec> v; return v; }
static std::mutex& guard_mu() { static std::mutex m; return m; }
template <typename T>
static cudaError_t dalloc(T** p, size_t n) {
  const size_t bytes = std::max<size_t>(n, 1) * sizeof(T);
  uint8_t* raw = nullptr;
  cudaError_t e = cudaMalloc(reinterpret_cast<void**>(&raw), bytes + 2 * kGuard);
  if (e != cudaSuccess) return e;
  e = cudaMemset(raw, 0xA5, bytes + 2 * kGuard);
  *p = reinterpret_cast<T*>(raw + kGuard);
  std::lock_guard<std::mutex> lock(guard_mu());
  guard_list().push_back({raw, bytes});
  return e;
}
// verify + free an arena allocated by dalloc
static void dfree(void* p) {
  if (!p) return;
  uint8_t* raw = static_cast<uint8_t*>(p) - kGuard;
  size_t bytes = 0;
  {
    std::lock_guard<std::mutex> lock(guard_mu());
    auto& v = guard_list();
    for (size_t i = 0; i < v.size(); ++i)
      if (v[i].base == raw) { bytes = v[i].bytes; v.erase(v.begin() + i); break; }
  }
  std::vector<uint8_t> lo(kGuard), hi(kGuard);
  cudaDeviceSynchronize();
  cudaMemcpy(lo.data(), raw, kGuard, cudaMemcpyDeviceToHost);
  cudaMemcpy(hi.data(), raw + kGuard + bytes, kGuard, cudaMemcpyDeviceToHost);
  for (size_t i = 0; i < kGuard; ++i)
    if (lo[i] != 0xA5 || hi[i] != 0xA5) {
      fprintf(stderr, "ORBFE_CHECKS: guard band of a %zu-byte arena overwritten (%s guard, byte %zu)\n", bytes,
              lo[i] != 0xA5 ? "lower" : "upper", i);
      abort();
    }
  cudaFree(raw);
}
#else
template <typename T>
static cudaError_t dalloc(T** p, size_t n) {
  return cudaMalloc(reinterpret_cast<void**>(p), std::max<size_t>(n, 1) * sizeof(T));
}
static void dfree(void* p) { cudaFree(p); }
#endif

static int alloc_arenas(Extractor* e, const std::vector<XTab>& xt, const std::vector<YTab>& yt) {
  const PyrGeom& g = e->geom;
  ExtractorArena& a = e->arena;
  const size_t B = (size_t)e->max_batch;
  ORBFE_CUDA(dalloc(&a.pyr, e->pyr_bytes + 256));
  ORBFE_CUDA(dalloc(&a.blur, e->pyr_bytes + 256));
  ORBFE_CUDA(dalloc(&a.stage, B * (size_t)e->width * e->height));
  ORBFE_CUDA(dalloc(&a.xtab, xt.size()));
  ORBFE_CUDA(dalloc(&a.ytab, yt.size()));
  ORBFE_CUDA(dalloc(&a.cand, B * g.cand_total));
  ORBFE_CUDA(dalloc(&a.node_of, B * g.cand_total));
  ORBFE_CUDA(dalloc(&a.cand_count, B * kMaxLevels));
  ORBFE_CUDA(dalloc(&a.cell_hi, B * (size_t)std::max(g.total_cells, 1)));
  ORBFE_CUDA(dalloc(&a.kept, B * g.kept_total));
  ORBFE_CUDA(dalloc(&a.kept_count, B * kMaxLevels));
  {
    size_t off = 0;
    auto take = [&](size_t bytes) { const size_t o = off; off = align_up(off + bytes, 256); return o; };
    take(B * g.out_cap * sizeof(orbfe_keypoint));
    e->off_desc = take(B * g.out_cap * 32);
    e->off_count = take(B * sizeof(int32_t));
    e->off_flags = take(B * sizeof(int32_t));
    e->result_bytes_extract = off;
    e->off_uright = take(B * g.out_cap * sizeof(float));
    e->off_depth = take(B * g.out_cap * sizeof(float));
    e->off_stcount = take(B * sizeof(int32_t));
    e->result_bytes = off;
    ORBFE_CUDA(cudaMalloc(reinterpret_cast<void**>(&e->result_block), e->result_bytes));
    a.out_kp = reinterpret_cast<orbfe_keypoint*>(e->result_block);
    a.out_desc = e->result_block + e->off_desc;
    a.out_count = reinterpret_cast<int32_t*>(e->result_block + e->off_count);
    a.flags = reinterpret_cast<int32_t*>(e->result_block + e->off_flags);
    a.st_uright = reinterpret_cast<float*>(e->result_block + e->off_uright);
    a.st_depth = reinterpret_cast<float*>(e->result_block + e->off_depth);
    a.st_count = reinterpret_cast<int32_t*>(e->result_block + e->off_stcount);
    if (e->result_bytes <= (1u << 20))  // small batches (single frames): one D2H copy + a host-side scatter
      ORBFE_CUDA(cudaHostAlloc(reinterpret_cast<void**>(&e->h_result), e->result_bytes, cudaHostAllocDefault));
  }
  ORBFE_CUDA(dalloc(&a.st_sad, B * g.out_cap));
  a.st_items_per_kp = stereo_items_per_kp(g);
  ORBFE_CUDA(dalloc(&a.st_band, B * g.out_cap));
  ORBFE_CUDA(dalloc(&a.st_bucket_start, B * (((size_t)(g.lv[0].h + 7) / 8) * g.nlevels + 2)));
  ORBFE_CUDA(dalloc(&a.st_bucket_items, B * g.out_cap * a.st_items_per_kp));
  if (!xt.empty()) ORBFE_CUDA(cudaMemcpy(a.xtab, xt.data(), xt.size() * sizeof(XTab), cudaMemcpyHostToDevice));
  if (!yt.empty()) ORBFE_CUDA(cudaMemcpy(a.ytab, yt.data(), yt.size() * sizeof(YTab), cudaMemcpyHostToDevice));
  {
    // k_pyramid.cu's per-word records: what a thread (4 destination columns) needs of the column table, in two
    // 16-byte loads.  Pixels past the row end repeat the last column (computed and dropped).
    std::vector<XWord> xw;
    for (int l = 1; l < g.nlevels; ++l) {
      const LevelGeom& L = g.lv[l];
      for (int x4 = 0; x4 < L.w; x4 += 4) {
        XWord r;
        memset(&r, 0, sizeof(r));
        r.ws = xt[L.xtab_off + x4].sx0 & ~3;
        for (int k = 0; k < 4; ++k) {
          const XTab& t = xt[L.xtab_off + std::min(x4 + k, L.w - 1)];
          const int o = L.resize_fast ? t.sx0 - r.ws : 0;  // 0 .. 7 (build_geometry checked it)
          r.selA |= ((unsigned)o | ((unsigned)std::min(o + 1, 7) << 4)) << (8 * k);
          r.selB |= (o == 7 ? 0x40u : 0x10u) << (8 * k);
          r.a01[k] = (unsigned)(unsigned short)t.a0 | ((unsigned)(unsigned short)t.a1 << 16);
        }
        xw.push_back(r);
      }
    }
    ORBFE_CUDA(dalloc(&a.xword, xw.size() + 1));
    if (!xw.empty()) ORBFE_CUDA(cudaMemcpy(a.xword, xw.data(), xw.size() * sizeof(XWord), cudaMemcpyHostToDevice));
  }
  ORBFE_CUDA(cudaMemset(a.pyr, 0, e->pyr_bytes + 256));
  ORBFE_CUDA(cudaMemset(a.blur, 0, e->pyr_bytes + 256));
  ORBFE_CUDA(cudaMemset(a.out_count, 0, B * sizeof(int32_t)));
  ORBFE_CUDA(cudaMemset(a.kept_count, 0, B * kMaxLevels * sizeof(int32_t)));
  ORBFE_CUDA(cudaMemset(a.flags, 0, B * sizeof(int32_t)));
  ORBFE_CUDA(cudaHostAlloc(reinterpret_cast<void**>(&e->h_flags), B * sizeof(int32_t), cudaHostAllocDefault));
  return build_fast_tensor_maps(g, a, e->max_batch);
}

// CUtensorMap of every level that has FAST strips (only those levels have keypoints): a (pitch, h, max_batch) byte
// tensor over one of the level-major arenas with a box_w x box_h x 1 box.  cuTensorMapEncodeTiled is a driver entry
// point; it is fetched through the runtime so that liborbfe.so does not link libcuda.
static int build_level_maps(const PyrGeom& g, uint8_t* arena, int max_batch, int box_w, int box_h,
                            CUtensorMapSwizzle swizzle, void** out, bool resize_sources = false) {
  typedef CUresult (*EncodeTiled)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
  void* fn = nullptr;
  cudaDriverEntryPointQueryResult q = cudaDriverEntryPointSymbolNotFound;
  if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &q) != cudaSuccess || !fn ||
      q != cudaDriverEntryPointSuccess) {
    set_error("cuTensorMapEncodeTiled is not available from this driver");
    return ORBFE_ERR_CUDA;
  }
  alignas(64) CUtensorMap maps[kMaxLevels];
  memset(maps, 0, sizeof(maps));
  for (int l = 0; l < g.nlevels; ++l) {
    const LevelGeom& L = g.lv[l];
    // resize_sources: level l is read by the staged resize kernel of level l+1
    if (resize_sources ? !(l + 1 < g.nlevels && g.lv[l + 1].resize_tile) : L.fast_strips_x <= 0) continue;
    const cuuint64_t dims[3] = {(cuuint64_t)L.pitch, (cuuint64_t)L.h, (cuuint64_t)max_batch};
    const cuuint64_t strides[2] = {(cuuint64_t)L.pitch, (cuuint64_t)L.lvl_bytes};
    const cuuint32_t box[3] = {(cuuint32_t)box_w, (cuuint32_t)box_h, 1u};
    const cuuint32_t estr[3] = {1u, 1u, 1u};
    const CUresult r = reinterpret_cast<EncodeTiled>(fn)(
        &maps[l], CU_TENSOR_MAP_DATA_TYPE_UINT8, 3, arena + L.lvl_base, dims, strides, box, estr,
        CU_TENSOR_MAP_INTERLEAVE_NONE, swizzle, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
        CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
      set_error("cuTensorMapEncodeTiled failed for level %d (%dx%d, pitch %d): CUresult %d", l, L.w, L.h, L.pitch, (int)r);
      return ORBFE_ERR_CUDA;
    }
  }
  ORBFE_CUDA(cudaMalloc(out, sizeof(maps)));
  ORBFE_CUDA(cudaMemcpy(*out, maps, sizeof(maps), cudaMemcpyHostToDevice));
  return ORBFE_OK;
}

int build_fast_tensor_maps(const PyrGeom& g, ExtractorArena& a, int max_batch) {
  int rc = build_level_maps(g, a.pyr, max_batch, kFastTileW, kFastTileH, CU_TENSOR_MAP_SWIZZLE_NONE, &a.fast_maps);  // k_fast.cu
  // k_describe.cu: 64-byte rows written with the 64-byte swizzle (bank spread of the rBRIEF gathers)
  static_assert(kDescBoxW == 64, "CU_TENSOR_MAP_SWIZZLE_64B: the box's inner extent is the swizzle span");
  if (rc == ORBFE_OK)
    rc = build_level_maps(g, a.blur, max_batch, kDescBoxW, kDescBoxH, CU_TENSOR_MAP_SWIZZLE_64B, &a.desc_maps);
  if (rc == ORBFE_OK)  // k_describe.cu: the IC_Angle disc of a keypoint
    rc = build_level_maps(g, a.pyr, max_batch, kDiscBoxW, kDiscBoxH, CU_TENSOR_MAP_SWIZZLE_NONE, &a.disc_maps);
  if (rc == ORBFE_OK)  // k_pyramid.cu: the source window of a 128 x 32 destination tile
    rc = build_level_maps(g, a.pyr, max_batch, kPyrTileW, kPyrTileH, CU_TENSOR_MAP_SWIZZLE_NONE, &a.pyr_maps, true);
  return rc;
}

static void free_arenas(Extractor* e) {
  ExtractorArena& a = e->arena;
  dfree(a.pyr); dfree(a.blur); dfree(a.stage); dfree(a.xtab); dfree(a.xword); dfree(a.ytab); dfree(a.cand); dfree(a.node_of);
  dfree(a.cand_count); dfree(a.cell_hi); dfree(a.kept); dfree(a.kept_count); cudaFree(e->result_block);
  if (e->h_result) cudaFreeHost(e->h_result);
  dfree(a.st_sad); dfree(a.st_band); dfree(a.st_bucket_start); dfree(a.st_bucket_items);
  cudaFree(a.fast_maps); cudaFree(a.desc_maps); cudaFree(a.disc_maps); cudaFree(a.pyr_maps);
  cudaFree(a.kp_un); cudaFree(a.depth_img);
  if (e->h_flags) cudaFreeHost(e->h_flags);
  if (e->prof_ev) {
    for (int i = 0; i < Extractor::kProfRuns * 9; ++i) cudaEventDestroy(e->prof_ev[i]);
    delete[] e->prof_ev;
    e->prof_ev = nullptr;
  }
  a = ExtractorArena();
}

}  // namespace orbfe

using namespace orbfe;

extern "C" {

const char* orbfe_last_error(void) { return orbfe::g_err; }

const char* orbfe_version(void) { return "orbfe-b200 0.1 (CUDA, sm_100a)"; }

int orbfe_device_count(void) {
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return 0; }
  return n;
}

int orbfe_extractor_create(const orbfe_extractor_params* params, int width, int height, int max_batch, int device,
                           orbfe_extractor** out) {
  ORBFE_ARG(params && out);
  ORBFE_ARG(params->nlevels >= 1 && params->nlevels <= kMaxLevels && params->nfeatures >= 1);
  ORBFE_ARG(params->scale_factor > 1.0f);
  ORBFE_ARG(width > 0 && height > 0 && max_batch >= 1);
  *out = nullptr;
  int ndev = 0;
  ORBFE_CUDA(cudaGetDeviceCount(&ndev));
  ORBFE_ARG(device >= 0 && device < ndev);
  Extractor* e = new (std::nothrow) Extractor();
  ORBFE_ARG(e != nullptr);
  e->params = *params;
  e->width = width;
  e->height = height;
  e->max_batch = max_batch;
  e->device = device;
  build_info(*params, &e->info);
  std::vector<XTab> xt;
  std::vector<YTab> yt;
  int rc = build_geometry(e, xt, yt);
  if (rc != ORBFE_OK) { delete e; return rc; }
  DeviceGuard guard(device);
  rc = alloc_arenas(e, xt, yt);
  if (rc == ORBFE_OK && cudaStreamCreateWithFlags(&e->stream, cudaStreamNonBlocking) != cudaSuccess) {
    set_error("cudaStreamCreate failed");
    rc = ORBFE_ERR_CUDA;
  }
  if (rc == ORBFE_OK && cudaEventCreateWithFlags(&e->done, cudaEventDisableTiming) != cudaSuccess) rc = ORBFE_ERR_CUDA;
  if (rc == ORBFE_OK && cudaEventCreateWithFlags(&e->stereo_done, cudaEventDisableTiming) != cudaSuccess) rc = ORBFE_ERR_CUDA;
  if (rc == ORBFE_OK) rc = configure_kernels(e->geom);
  { const char* ng = getenv("ORBFE_NO_GRAPHS"); e->use_graph = !(ng && ng[0] == '1'); }
  { const char* nf = getenv("ORBFE_NO_FORK"); e->fork_graph = !(nf && nf[0] == '1'); }
  if (rc == ORBFE_OK && e->fork_graph &&
      (cudaStreamCreateWithFlags(&e->side, cudaStreamNonBlocking) != cudaSuccess ||
       cudaEventCreateWithFlags(&e->fork_ev, cudaEventDisableTiming) != cudaSuccess ||
       cudaEventCreateWithFlags(&e->join_ev, cudaEventDisableTiming) != cudaSuccess))
    rc = ORBFE_ERR_CUDA;
  if (rc != ORBFE_OK) {
    free_arenas(e);
    if (e->done) cudaEventDestroy(e->done);
    if (e->stereo_done) cudaEventDestroy(e->stereo_done);
    if (e->fork_ev) cudaEventDestroy(e->fork_ev);
    if (e->join_ev) cudaEventDestroy(e->join_ev);
    if (e->side) cudaStreamDestroy(e->side);
    if (e->stream) cudaStreamDestroy(e->stream);
    delete e;
    return rc;
  }
  *out = reinterpret_cast<orbfe_extractor*>(e);
  return ORBFE_OK;
}

void orbfe_extractor_destroy(orbfe_extractor* h) {
  if (!h) return;
  Extractor* e = reinterpret_cast<Extractor*>(h);
  DeviceGuard guard(e->device);
  // a borrowed stream (orbfe_extractor_share_stream) may already be gone: wait for the whole device instead
  if (e->owns_stream) cudaStreamSynchronize(e->stream); else cudaDeviceSynchronize();
  free_arenas(e);
  if (e->d_raw) cudaFree(e->d_raw);
  if (e->d_map_x) cudaFree(e->d_map_x);
  if (e->d_map_y) cudaFree(e->d_map_y);
  if (e->run_graph) cudaGraphExecDestroy(e->run_graph);
  if (e->stereo_graph) cudaGraphExecDestroy(e->stereo_graph);
  if (e->done) cudaEventDestroy(e->done);
  if (e->stereo_done) cudaEventDestroy(e->stereo_done);
  if (e->fork_ev) cudaEventDestroy(e->fork_ev);
  if (e->join_ev) cudaEventDestroy(e->join_ev);
  if (e->side) cudaStreamDestroy(e->side);
  if (e->owns_stream) cudaStreamDestroy(e->stream);
  delete e;
}

int orbfe_extractor_get_info(const orbfe_extractor* h, orbfe_extractor_info* info) {
  ORBFE_ARG(h && info);
  *info = reinterpret_cast<const Extractor*>(h)->info;
  return ORBFE_OK;
}

void* orbfe_extractor_stream(orbfe_extractor* h) { return h ? reinterpret_cast<Extractor*>(h)->stream : nullptr; }

int64_t orbfe_extractor_launch_count(const orbfe_extractor* h) {
  return h ? reinterpret_cast<const Extractor*>(h)->launches : 0;
}

int orbfe_extractor_share_stream(orbfe_extractor* h, orbfe_extractor* other) {
  ORBFE_ARG(h && other && h != other);
  Extractor* e = reinterpret_cast<Extractor*>(h);
  Extractor* o = reinterpret_cast<Extractor*>(other);
  ORBFE_ARG(e->device == o->device);
  DeviceGuard guard(e->device);
  ORBFE_CUDA(cudaStreamSynchronize(e->stream));
  if (e->owns_stream) cudaStreamDestroy(e->stream);
  e->stream = o->stream;
  e->owns_stream = false;
  return ORBFE_OK;
}

int orbfe_extractor_sync(orbfe_extractor* h) {
  ORBFE_ARG(h);
  Extractor* e = reinterpret_cast<Extractor*>(h);
  DeviceGuard guard(e->device);
  ORBFE_CUDA(cudaStreamSynchronize(e->stream));
  return ORBFE_OK;
}

int orbfe_extract_upload(orbfe_extractor* h, const uint8_t* images, size_t image_stride, int row_stride, int batch,
                         int images_on_device) {
  ORBFE_ARG(h && images);
  Extractor* e = reinterpret_cast<Extractor*>(h);
  ORBFE_ARG(batch >= 1 && batch <= e->max_batch && row_stride >= e->width);
  DeviceGuard guard(e->device);
  const LevelGeom& L0 = e->geom.lv[0];
  const cudaMemcpyKind kind = images_on_device ? cudaMemcpyDeviceToDevice : cudaMemcpyHostToDevice;
  uint8_t* dst = e->arena.pyr + L0.lvl_base;
  const bool packed = image_stride == (size_t)row_stride * e->height || batch == 1;
  if (!images_on_device && packed && row_stride == e->width && L0.pitch != e->width) {
    // tightly packed host images: ONE linear H2D (full DMA rate; a 2-D copy with an odd width runs several times
    // slower) into a staging buffer, then a 2-D device-to-device copy into the pitched level 0
    const size_t bytes = (size_t)e->width * e->height * batch;
    ORBFE_CUDA(cudaMemcpyAsync(e->arena.stage, images, bytes, cudaMemcpyHostToDevice, e->stream));
    launch_repitch(e->arena.stage, (size_t)e->width * e->height, (size_t)e->width, dst, L0.pitch, L0.lvl_bytes, e->width,
                   e->height, batch, e->stream);
    e->launches += 1;
    ORBFE_CUDA(cudaGetLastError());
  } else if (images_on_device) {
    // device images at any alignment / stride: the library's own re-pitch kernel (a 2-D device-to-device copy of
    // odd-width rows runs at 2.6 TB/s)
    launch_repitch(images, batch == 1 ? 0 : image_stride, (size_t)row_stride, dst, L0.pitch, L0.lvl_bytes, e->width,
                   e->height, batch, e->stream);
    e->launches += 1;
    ORBFE_CUDA(cudaGetLastError());
  } else if (packed) {
    // images are back to back: one 2-D copy of batch*height rows (level-major arena: image b at b*pitch*h)
    ORBFE_CUDA(cudaMemcpy2DAsync(dst, L0.pitch, images, row_stride, e->width, (size_t)e->height * batch, kind,
                                 e->stream));
  } else {
    for (int b = 0; b < batch; ++b)
      ORBFE_CUDA(cudaMemcpy2DAsync(dst + (size_t)b * L0.lvl_bytes, L0.pitch, images + (size_t)b * image_stride,
                                   row_stride, e->width, e->height, kind, e->stream));
  }
  return ORBFE_OK;
}

static int ensure_raw(Extractor* e, size_t bytes) {
  if (bytes <= e->raw_cap) return ORBFE_OK;
  if (e->d_raw) cudaFree(e->d_raw);
  e->d_raw = nullptr; e->raw_cap = 0;
  ORBFE_CUDA(cudaMalloc(reinterpret_cast<void**>(&e->d_raw), bytes));
  e->raw_cap = bytes;
  return ORBFE_OK;
}

int orbfe_extract_upload_color(orbfe_extractor* h, const uint8_t* pixels, size_t image_stride, int row_stride,
                               int channels, int rgb, int batch) {
  ORBFE_ARG(h && pixels);
  Extractor* e = reinterpret_cast<Extractor*>(h);
  ORBFE_ARG(channels == 3 || channels == 4);
  ORBFE_ARG(batch >= 1 && batch <= e->max_batch && row_stride >= e->width * channels);
  if (batch == 1) image_stride = (size_t)row_stride * e->height;
  ORBFE_ARG(image_stride >= (size_t)row_stride * e->height);
  DeviceGuard guard(e->device);
  const size_t bytes = image_stride * (size_t)(batch - 1) + (size_t)row_stride * e->height;
  int rc = ensure_raw(e, bytes);
  if (rc != ORBFE_OK) return rc;
  ORBFE_CUDA(cudaMemcpyAsync(e->d_raw, pixels, bytes, cudaMemcpyHostToDevice, e->stream));
  const LevelGeom& L0 = e->geom.lv[0];
  launch_gray(e->d_raw, image_stride, row_stride, channels, rgb, e->arena.pyr + L0.lvl_base, L0.pitch, L0.lvl_bytes,
              e->width, e->height, batch, e->stream);
  e->launches += 1;
  ORBFE_CUDA(cudaGetLastError());
  return ORBFE_OK;
}

int orbfe_extractor_set_rectify_maps(orbfe_extractor* h, const float* map_x, const float* map_y, int src_width,
                                     int src_height) {
  ORBFE_ARG(h && map_x && map_y && src_width >= 1 && src_height >= 1 && src_width <= 32767 && src_height <= 32767);
  Extractor* e = reinterpret_cast<Extractor*>(h);
  DeviceGuard guard(e->device);
  const size_t n = (size_t)e->width * e->height;
  if (!e->d_map_x) {
    ORBFE_CUDA(cudaMalloc(reinterpret_cast<void**>(&e->d_map_x), n * sizeof(float)));
    ORBFE_CUDA(cudaMalloc(reinterpret_cast<void**>(&e->d_map_y), n * sizeof(float)));
  }
  ORBFE_CUDA(cudaMemcpyAsync(e->d_map_x, map_x, n * sizeof(float), cudaMemcpyHostToDevice, e->stream));
  ORBFE_CUDA(cudaMemcpyAsync(e->d_map_y, map_y, n * sizeof(float), cudaMemcpyHostToDevice, e->stream));
  ORBFE_CUDA(cudaStreamSynchronize(e->stream));
  e->src_w = src_width;
  e->src_h = src_height;
  return ORBFE_OK;
}

int orbfe_extract_upload_distorted(orbfe_extractor* h, const uint8_t* pixels, size_t image_stride, int row_stride,
                                   int batch) {
  ORBFE_ARG(h && pixels);
  Extractor* e = reinterpret_cast<Extractor*>(h);
  if (!e->d_map_x) {
    set_error("orbfe_extract_upload_distorted: no rectification maps (call orbfe_extractor_set_rectify_maps first)");
    return ORBFE_ERR_ARG;
  }
  ORBFE_ARG(batch >= 1 && batch <= e->max_batch && row_stride >= e->src_w);
  if (batch == 1) image_stride = (size_t)row_stride * e->src_h;
  ORBFE_ARG(image_stride >= (size_t)row_stride * e->src_h);
  DeviceGuard guard(e->device);
  const size_t bytes = image_stride * (size_t)(batch - 1) + (size_t)row_stride * e->src_h;
  int rc = ensure_raw(e, bytes);
  if (rc != ORBFE_OK) return rc;
  ORBFE_CUDA(cudaMemcpyAsync(e->d_raw, pixels, bytes, cudaMemcpyHostToDevice, e->stream));
  const LevelGeom& L0 = e->geom.lv[0];
  launch_remap(e->d_raw, image_stride, row_stride, e->src_w, e->src_h, e->d_map_x, e->d_map_y,
               e->arena.pyr + L0.lvl_base, L0.pitch, L0.lvl_bytes, e->width, e->height, batch, e->stream);
  e->launches += 1;
  ORBFE_CUDA(cudaGetLastError());
  return ORBFE_OK;
}

// the kernel schedule of operator() (:1065-1127) for `batch` images on e->stream
static void issue_extract(Extractor* e, int batch, cudaEvent_t* ev) {
  const PyrGeom& g = e->geom;
  const ExtractorArena& a = e->arena;
  cudaMemsetAsync(a.cand_count, 0, (size_t)batch * kMaxLevels * sizeof(int32_t), e->stream);
  cudaMemsetAsync(a.flags, 0, (size_t)batch * sizeof(int32_t), e->stream);
  if (ev) cudaEventRecord(ev[0], e->stream);
  for (int l = 1; l < g.nlevels; ++l) launch_pyramid_level(g, a, l, batch, e->stream);  // ComputePyramid (:1075)
  if (ev) cudaEventRecord(ev[1], e->stream);
  launch_fast_cells(g, a, batch, e->stream);                                             // :776-850
  if (ev) cudaEventRecord(ev[2], e->stream);
  launch_quadtree(g, a, batch, e->stream);                                               // :853
  if (ev) cudaEventRecord(ev[3], e->stream);
  launch_blur_levels(g, a, batch, e->stream);                                            // :1107-1108
  if (ev) cudaEventRecord(ev[4], e->stream);
  launch_describe(g, a, batch, e->stream);                                               // :873-874, :1096-1126
  if (ev) cudaEventRecord(ev[5], e->stream);
}

// the same schedule as a graph with two branches after the pyramid: GaussianBlur needs only the pyramid, so it runs
// beside FAST + quadtree (the latency-bound quadtree leaves issue slots the blur can use) and joins before describe.
// Only used under stream capture (e->side joins the capture through the fork event).
static void issue_extract_forked(Extractor* e, int batch) {
  const PyrGeom& g = e->geom;
  const ExtractorArena& a = e->arena;
  cudaMemsetAsync(a.cand_count, 0, (size_t)batch * kMaxLevels * sizeof(int32_t), e->stream);
  cudaMemsetAsync(a.flags, 0, (size_t)batch * sizeof(int32_t), e->stream);
  for (int l = 1; l < g.nlevels; ++l) launch_pyramid_level(g, a, l, batch, e->stream);
  cudaEventRecord(e->fork_ev, e->stream);
  cudaStreamWaitEvent(e->side, e->fork_ev, 0);
  launch_fast_cells(g, a, batch, e->stream);
  launch_blur_levels(g, a, batch, e->side);
  cudaEventRecord(e->join_ev, e->side);
  launch_quadtree(g, a, batch, e->stream);
  cudaStreamWaitEvent(e->stream, e->join_ev, 0);
  launch_describe(g, a, batch, e->stream);
}

int orbfe_extract_run(orbfe_extractor* h, int batch) {
  ORBFE_ARG(h);
  Extractor* e = reinterpret_cast<Extractor*>(h);
  ORBFE_ARG(batch >= 1 && batch <= e->max_batch);
  DeviceGuard guard(e->device);
  const PyrGeom& g = e->geom;
  cudaEvent_t* ev = (e->profiling && e->prof_runs < Extractor::kProfRuns) ? e->prof_ev + 6 * e->prof_runs : nullptr;
  if (!ev && e->use_graph && (!e->run_graph || e->run_graph_batch != batch)) {
    // one cudaGraphLaunch instead of 3 memsets + 11 kernel launches: what matters for single-frame latency, where
    // the kernels are a few microseconds each.  Captured once per batch size, on the calling thread only.
    if (e->run_graph) { cudaGraphExecDestroy(e->run_graph); e->run_graph = nullptr; }
    const bool ok = capture_graph(e->stream, &e->run_graph, [&] {
      if (e->fork_graph) issue_extract_forked(e, batch); else issue_extract(e, batch, nullptr);
    });
    if (ok) e->run_graph_batch = batch; else e->use_graph = false;  // plain launches from now on
  }
  if (ev || !e->use_graph) {
    issue_extract(e, batch, ev);
    if (ev) ++e->prof_runs;
  } else {
    ORBFE_CUDA(cudaGraphLaunch(e->run_graph, e->stream));
  }
  e->launches += (g.nlevels - 1) + (g.fast_strips > 0 ? 1 : 0) + 3;
  e->last_batch = batch;
  e->arena.kp_un_valid = false;  // mvKeysUn belongs to the previous extraction
  ORBFE_CUDA(cudaGetLastError());
  return ORBFE_OK;
}

int orbfe_extractor_set_profiling(orbfe_extractor* h, int enable) {
  ORBFE_ARG(h);
  Extractor* e = reinterpret_cast<Extractor*>(h);
  DeviceGuard guard(e->device);
  if (enable && !e->prof_ev) {
    const int n = Extractor::kProfRuns * 9;
    e->prof_ev = new cudaEvent_t[n];
    for (int i = 0; i < n; ++i) ORBFE_CUDA(cudaEventCreate(&e->prof_ev[i]));
  }
  e->profiling = enable != 0;
  e->prof_runs = 0;
  e->prof_stereo_runs = 0;
  return ORBFE_OK;
}

int orbfe_extractor_stage_times(orbfe_extractor* h, float* ms, int32_t* runs, int32_t* stereo_runs) {
  ORBFE_ARG(h && ms);
  Extractor* e = reinterpret_cast<Extractor*>(h);
  DeviceGuard guard(e->device);
  ORBFE_CUDA(cudaStreamSynchronize(e->stream));
  for (int s = 0; s < Extractor::kStages; ++s) ms[s] = 0.f;
  for (int r = 0; r < e->prof_runs; ++r)
    for (int s = 0; s < 5; ++s) {
      float t = 0.f;
      ORBFE_CUDA(cudaEventElapsedTime(&t, e->prof_ev[6 * r + s], e->prof_ev[6 * r + s + 1]));
      ms[s] += t;
    }
  cudaEvent_t* sev = e->prof_ev ? e->prof_ev + 6 * Extractor::kProfRuns : nullptr;
  for (int r = 0; r < e->prof_stereo_runs; ++r)
    for (int s = 0; s < 2; ++s) {
      float t = 0.f;
      ORBFE_CUDA(cudaEventElapsedTime(&t, sev[3 * r + s], sev[3 * r + s + 1]));
      ms[5 + s] += t;
    }
  if (runs) *runs = e->prof_runs;
  if (stereo_runs) *stereo_runs = e->prof_stereo_runs;
  return ORBFE_OK;
}

int orbfe_extract_download_async(orbfe_extractor* h, int batch, orbfe_keypoint* keypoints, uint8_t* descriptors,
                                 int capacity, int32_t* counts) {
  ORBFE_ARG(h && counts && capacity >= 0);
  Extractor* e = reinterpret_cast<Extractor*>(h);
  ORBFE_ARG(batch >= 1 && batch <= e->max_batch);
  DeviceGuard guard(e->device);
  const PyrGeom& g = e->geom;
  const ExtractorArena& a = e->arena;
  const int ncopy = capacity < g.out_cap ? capacity : g.out_cap;
  ORBFE_CUDA(cudaMemcpyAsync(counts, a.out_count, (size_t)batch * sizeof(int32_t), cudaMemcpyDeviceToHost, e->stream));
  ORBFE_CUDA(cudaMemcpyAsync(e->h_flags, a.flags, (size_t)batch * sizeof(int32_t), cudaMemcpyDeviceToHost, e->stream));
  if (capacity == g.out_cap) {  // same slab layout on both sides: linear copies
    if (keypoints)
      ORBFE_CUDA(cudaMemcpyAsync(keypoints, a.out_kp, (size_t)batch * g.out_cap * sizeof(orbfe_keypoint),
                                 cudaMemcpyDeviceToHost, e->stream));
    if (descriptors)
      ORBFE_CUDA(cudaMemcpyAsync(descriptors, a.out_desc, (size_t)batch * g.out_cap * 32, cudaMemcpyDeviceToHost,
                                 e->stream));
  } else {
    if (ncopy > 0 && keypoints)
      ORBFE_CUDA(cudaMemcpy2DAsync(keypoints, (size_t)capacity * sizeof(orbfe_keypoint), a.out_kp,
                                   (size_t)g.out_cap * sizeof(orbfe_keypoint), (size_t)ncopy * sizeof(orbfe_keypoint),
                                   batch, cudaMemcpyDeviceToHost, e->stream));
    if (ncopy > 0 && descriptors)
      ORBFE_CUDA(cudaMemcpy2DAsync(descriptors, (size_t)capacity * 32, a.out_desc, (size_t)g.out_cap * 32,
                                   (size_t)ncopy * 32, batch, cudaMemcpyDeviceToHost, e->stream));
  }
  e->pending_counts = counts;
  e->pending_batch = batch;
  e->pending_capacity = capacity;
  return ORBFE_OK;
}

int orbfe_extract_finish(orbfe_extractor* h) {
  ORBFE_ARG(h);
  Extractor* e = reinterpret_cast<Extractor*>(h);
  DeviceGuard guard(e->device);
  ORBFE_CUDA(cudaStreamSynchronize(e->stream));
  const int32_t* counts = e->pending_counts;
  const int batch = e->pending_batch, capacity = e->pending_capacity;
  e->pending_counts = nullptr;
  e->pending_batch = 0;
  if (!counts) return ORBFE_OK;
  for (int b = 0; b < batch; ++b) {
    if (e->h_flags[b]) { set_error("image %d: internal slab overflow (flags=%d)", b, e->h_flags[b]); return ORBFE_ERR_CAPACITY; }
    if (counts[b] > capacity) { set_error("image %d: %d keypoints > capacity %d", b, counts[b], capacity); return ORBFE_ERR_CAPACITY; }
  }
  return ORBFE_OK;
}

int orbfe_extract_download(orbfe_extractor* h, int batch, orbfe_keypoint* keypoints, uint8_t* descriptors,
                           int capacity, int32_t* counts) {
  const int rc = orbfe_extract_download_async(h, batch, keypoints, descriptors, capacity, counts);
  if (rc != ORBFE_OK) return rc;
  return orbfe_extract_finish(h);
}

int orbfe_extract_batch(orbfe_extractor* h, const uint8_t* images, size_t image_stride, int row_stride, int batch,
                        orbfe_keypoint* keypoints, uint8_t* descriptors, int capacity, int32_t* counts) {
  ORBFE_ARG(h && counts);
  if (!images) {  // empty image: the reference returns silently with no keypoints (:1068)
    for (int b = 0; b < batch; ++b) counts[b] = 0;
    return ORBFE_OK;
  }
  int rc = orbfe_extract_upload(h, images, image_stride, row_stride, batch, 0);
  if (rc != ORBFE_OK) return rc;
  rc = orbfe_extract_run(h, batch);
  if (rc != ORBFE_OK) return rc;
  return orbfe_extract_download(h, batch, keypoints, descriptors, capacity, counts);
}

int orbfe_extract(orbfe_extractor* h, const uint8_t* image, int row_stride, orbfe_keypoint* keypoints,
                  uint8_t* descriptors, int capacity, int32_t* n_out) {
  ORBFE_ARG(n_out);
  return orbfe_extract_batch(h, image, 0, row_stride, 1, keypoints, descriptors, capacity, n_out);
}

int orbfe_extractor_level(const orbfe_extractor* h, int image_index, int level, const uint8_t** device_ptr, int* pitch,
                          int* width, int* height) {
  ORBFE_ARG(h);
  const Extractor* e = reinterpret_cast<const Extractor*>(h);
  ORBFE_ARG(level >= 0 && level < e->geom.nlevels && image_index >= 0 && image_index < e->max_batch);
  const LevelGeom& L = e->geom.lv[level];
  if (device_ptr) *device_ptr = e->arena.pyr + L.lvl_base + (size_t)image_index * L.lvl_bytes;
  if (pitch) *pitch = L.pitch;
  if (width) *width = L.w;
  if (height) *height = L.h;
  return ORBFE_OK;
}

int orbfe_extractor_download_level(orbfe_extractor* h, int image_index, int level, uint8_t* dst, int dst_stride) {
  ORBFE_ARG(h && dst);
  Extractor* e = reinterpret_cast<Extractor*>(h);
  const uint8_t* src = nullptr;
  int pitch = 0, w = 0, hh = 0;
  int rc = orbfe_extractor_level(h, image_index, level, &src, &pitch, &w, &hh);
  if (rc != ORBFE_OK) return rc;
  ORBFE_ARG(dst_stride >= w);
  DeviceGuard guard(e->device);
  ORBFE_CUDA(cudaMemcpy2DAsync(dst, dst_stride, src, pitch, w, hh, cudaMemcpyDeviceToHost, e->stream));
  ORBFE_CUDA(cudaStreamSynchronize(e->stream));
  return ORBFE_OK;
}

int orbfe_extractor_device_results(const orbfe_extractor* h, const orbfe_keypoint** keypoints,
                                   const uint8_t** descriptors, const int32_t** counts, int* capacity) {
  ORBFE_ARG(h);
  const Extractor* e = reinterpret_cast<const Extractor*>(h);
  if (keypoints) *keypoints = e->arena.out_kp;
  if (descriptors) *descriptors = e->arena.out_desc;
  if (counts) *counts = e->arena.out_count;
  if (capacity) *capacity = e->geom.out_cap;
  return ORBFE_OK;
}

static int debug_list(Extractor* e, const LevelGeom* sort_geom, const uint32_t* d_list, const int32_t* d_count,
                      int off, int lcap, int cap, int32_t* xs, int32_t* ys, int32_t* scores, int32_t* n_out) {
  DeviceGuard guard(e->device);
  int32_t n = 0;
  ORBFE_CUDA(cudaStreamSynchronize(e->stream));
  ORBFE_CUDA(cudaMemcpy(&n, d_count, sizeof(n), cudaMemcpyDeviceToHost));
  if (n > lcap) n = lcap;
  std::vector<uint32_t> tmp((size_t)std::max(n, 1));
  if (n > 0) ORBFE_CUDA(cudaMemcpy(tmp.data(), d_list + off, (size_t)n * sizeof(uint32_t), cudaMemcpyDeviceToHost));
  if (sort_geom) {  // the reference's append order: cell row, cell column, row-major inside the cell (:801-846)
    const LevelGeom& L = *sort_geom;
    auto key = [&](uint32_t c) {
      const unsigned long long x = cand_x(c), y = cand_y(c);
      return ((unsigned long long)((y - 3) / L.h_cell) << 40) | ((unsigned long long)((x - 3) / L.w_cell) << 32) |
             (y << 16) | x;
    };
    std::sort(tmp.begin(), tmp.begin() + n, [&](uint32_t p, uint32_t q) { return key(p) < key(q); });
  }
  for (int i = 0; i < n && i < cap; ++i) {
    if (xs) xs[i] = cand_x(tmp[i]);
    if (ys) ys[i] = cand_y(tmp[i]);
    if (scores) scores[i] = cand_score(tmp[i]);
  }
  *n_out = n;
  return ORBFE_OK;
}

int orbfe_extractor_debug_candidates(orbfe_extractor* h, int image_index, int level, int cap, int32_t* xs, int32_t* ys,
                                     int32_t* scores, int32_t* n_out) {
  ORBFE_ARG(h && n_out);
  Extractor* e = reinterpret_cast<Extractor*>(h);
  ORBFE_ARG(level >= 0 && level < e->geom.nlevels && image_index >= 0 && image_index < e->max_batch);
  const LevelGeom& L = e->geom.lv[level];
  return debug_list(e, &L, e->arena.cand + (size_t)image_index * e->geom.cand_total,
                    e->arena.cand_count + image_index * kMaxLevels + level, L.cand_off, L.cand_cap, cap, xs, ys, scores,
                    n_out);
}

int orbfe_extractor_debug_level_keypoints(orbfe_extractor* h, int image_index, int level, int cap, int32_t* xs,
                                          int32_t* ys, int32_t* scores, int32_t* n_out) {
  ORBFE_ARG(h && n_out);
  Extractor* e = reinterpret_cast<Extractor*>(h);
  ORBFE_ARG(level >= 0 && level < e->geom.nlevels && image_index >= 0 && image_index < e->max_batch);
  const LevelGeom& L = e->geom.lv[level];
  return debug_list(e, nullptr, e->arena.kept + (size_t)image_index * e->geom.kept_total,
                    e->arena.kept_count + image_index * kMaxLevels + level, L.kept_off, L.kept_cap, cap, xs, ys, scores,
                    n_out);
}

}  // extern "C"
