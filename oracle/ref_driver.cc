// ref_driver.cc — C API around the UNMODIFIED reference sources compiled in place from /root/reference (see
// oracle/Makefile, target _ref/libref_orbslam2.so).  TEST INFRASTRUCTURE ONLY.
//
// The reference's own translation units (src/ORBextractor.cc, ...) are compiled against the OpenCV stand-in of
// oracle/ref_shim/ and linked with this file.  Nothing of the reference's text is copied here: this file only
// constructs the reference's objects, calls their public methods and copies the results out.
//
// Two things the reference's results depend on that are NOT in its source are made switchable (ref_set_mode):
//   * REF_MODE_HEAP_TIES: the quadtree's sort of (size, node pointer) pairs (src/ORBextractor.cc:686) orders equal
//     sizes by HEAP ADDRESS.  Default (bit clear): list nodes come from a monotonic arena, so address order == creation
//     order — the deterministic rule the oracle and the CUDA path implement.  Bit set: plain malloc, as the reference
//     happens to run.
//   * REF_MODE_LIBM: rBRIEF's cos/sin (:112-113) resolve to libm's cosf/sinf whose last bit is machine dependent.
//     Default: cosf/sinf/sincosf/logf of this library are the correctly rounded functions of
//     include/orb_math_shared.h.  Bit set: the system libm.
// The library must be linked with -Wl,-Bsymbolic so that the reference's calls bind to the definitions below.
#include <dlfcn.h>

#include <atomic>
#include <cmath>
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <iostream>
#include <list>
#include <new>
#include <vector>

#include "../include/orb_math_shared.h"
#include "../include/orbfe_types.h"
#include "ORBextractor.h"
#include "ref_driver.h"

#define REF_MODE_LIBM 1
#define REF_MODE_HEAP_TIES 2

static int g_mode = 0;

// ---------------------------------------------------------------- monotonic arena for the quadtree's list nodes
namespace {
constexpr size_t kNodeSize = sizeof(std::_List_node<ORB_SLAM2::ExtractorNode>);
constexpr size_t kArenaBytes = (size_t)256 << 20;
// ONE arena for the process (objects of the node size may be allocated by one thread and freed by another, e.g. a
// std::thread state): a bump pointer advanced atomically, so that within each thread later allocations get higher
// addresses.  Rewound at the start of a reference call when nothing from the arena is alive.
unsigned char* g_arena = nullptr;
std::atomic<size_t> g_arena_used{0};
std::atomic<long> g_arena_live{0};
std::atomic<int> g_arena_on{0};  // set for the duration of a reference call

inline bool in_arena(void* p) {
  return g_arena && (unsigned char*)p >= g_arena && (unsigned char*)p < g_arena + kArenaBytes;
}
}  // namespace

void* operator new(size_t n) {
  if (n == kNodeSize && g_arena_on.load(std::memory_order_relaxed)) {
    const size_t sz = (n + 15) & ~(size_t)15;
    const size_t off = g_arena_used.fetch_add(sz);
    if (off + sz > kArenaBytes) { std::fprintf(stderr, "ref_driver: node arena exhausted\n"); std::abort(); }
    g_arena_live.fetch_add(1);
    return g_arena + off;
  }
  void* p = std::malloc(n ? n : 1);
  if (!p) throw std::bad_alloc();
  return p;
}
void operator delete(void* p) noexcept {
  if (in_arena(p)) { g_arena_live.fetch_sub(1); return; }
  std::free(p);
}
void operator delete(void* p, size_t) noexcept { operator delete(p); }

// ---------------------------------------------------------------- libm entry points the reference's path calls
namespace {
template <typename F> F libm_sym(const char* name) {
  static void* h = dlopen("libm.so.6", RTLD_NOW | RTLD_LOCAL);
  void* s = h ? dlsym(h, name) : nullptr;
  if (!s) { std::fprintf(stderr, "ref_driver: libm symbol %s not found\n", name); std::abort(); }
  return reinterpret_cast<F>(s);
}
}  // namespace

extern "C" {
float cosf(float x) noexcept {
  if (g_mode & REF_MODE_LIBM) { static auto f = libm_sym<float (*)(float)>("cosf"); return f(x); }
  float c, s; orb_sincos_rad(x, &c, &s); return c;
}
float sinf(float x) noexcept {
  if (g_mode & REF_MODE_LIBM) { static auto f = libm_sym<float (*)(float)>("sinf"); return f(x); }
  float c, s; orb_sincos_rad(x, &c, &s); return s;
}
void sincosf(float x, float* s, float* c) noexcept {
  if (g_mode & REF_MODE_LIBM) { static auto f = libm_sym<void (*)(float, float*, float*)>("sincosf"); f(x, s, c); return; }
  orb_sincos_rad(x, c, s);
}
float logf(float x) noexcept {
  if ((g_mode & REF_MODE_LIBM) || !(x > 0.f) || std::isinf(x)) { static auto f = libm_sym<float (*)(float)>("logf"); return f(x); }
  return orb_log_f32(x);
}

void ref_set_mode(int mode) { g_mode = mode; }
int ref_get_mode(void) { return g_mode; }
size_t ref_list_node_size(void) { return kNodeSize; }
}  // extern "C"

// silence the reference's per-level std::cout line (src/ORBextractor.cc:854) for the duration of a call
RefCallScope::RefCallScope() {
  mute_ = new std::ios_base::iostate(std::cout.rdstate());
  std::cout.setstate(std::ios_base::failbit);
  if (!(g_mode & REF_MODE_HEAP_TIES)) {
    if (!g_arena) g_arena = (unsigned char*)std::malloc(kArenaBytes);  // virtual until touched
    if (g_arena_live.load() == 0) g_arena_used.store(0);
    g_arena_on.store(1);
  }
}
RefCallScope::~RefCallScope() {
  g_arena_on.store(0);
  std::ios_base::iostate* st = static_cast<std::ios_base::iostate*>(mute_);
  std::cout.clear(*st);
  delete st;
}

// ---------------------------------------------------------------- ORBextractor
extern "C" {

ref_extractor* ref_extractor_create(int nfeatures, float scale_factor, int nlevels, int ini_th, int min_th) {
  ref_extractor* r = new ref_extractor();
  r->ex = new ORB_SLAM2::ORBextractor(nfeatures, scale_factor, nlevels, ini_th, min_th);
  return r;
}
void ref_extractor_destroy(ref_extractor* r) {
  if (!r) return;
  delete r->ex;
  delete r;
}

// ORBextractor::operator() (src/ORBextractor.cc:1065-1127) -> number of keypoints (all of them are kept in the handle;
// at most `capacity` are copied out)
int ref_extract(ref_extractor* r, const uint8_t* image, int width, int height, int stride, orbfe_keypoint* kps,
                uint8_t* desc, int capacity) {
  RefCallScope scope;
  cv::Mat img(height, width, CV_8UC1, const_cast<uint8_t*>(image), (size_t)stride);
  r->kps.clear();
  r->desc = cv::Mat();
  (*r->ex)(img, cv::Mat(), r->kps, r->desc);
  const int n = (int)r->kps.size();
  static_assert(sizeof(orbfe_keypoint) == sizeof(cv::KeyPoint), "keypoint layout");
  for (int i = 0; i < n && i < capacity; ++i) {
    if (kps) std::memcpy(&kps[i], &r->kps[i], sizeof(cv::KeyPoint));
    if (desc) std::memcpy(desc + (size_t)i * 32, r->desc.ptr(i), 32);
  }
  return n;
}

// mvImagePyramid[level] of the last call
const uint8_t* ref_extractor_level(ref_extractor* r, int level, int* w, int* h, int* stride) {
  const cv::Mat& m = r->ex->mvImagePyramid[level];
  *w = m.cols; *h = m.rows; *stride = (int)m.step;
  return m.data;
}

void ref_extractor_tables(ref_extractor* r, float* scale, float* inv_scale, float* sigma2, float* inv_sigma2) {
  const int n = r->ex->GetLevels();
  const std::vector<float> a = r->ex->GetScaleFactors(), b = r->ex->GetInverseScaleFactors(),
                           c = r->ex->GetScaleSigmaSquares(), d = r->ex->GetInverseScaleSigmaSquares();
  for (int i = 0; i < n; ++i) { scale[i] = a[i]; inv_scale[i] = b[i]; sigma2[i] = c[i]; inv_sigma2[i] = d[i]; }
}

}  // extern "C"
