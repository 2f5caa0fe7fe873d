// matcher_ctx.cuh — the matcher context shared by matcher.cu and matcher2.cu: a stream, one device arena with a
// pinned host mirror of its upload prefix (every call packs the caller's struct-of-array views there and uploads them
// with ONE copy), a growable candidate buffer, and the helpers that mirror a frame / keyframe view onto the device.
#pragma once
#include <algorithm>
#include <cstring>
#include <new>

#include "common.cuh"
#include "match_kernels.h"

namespace orbfe {

struct Matcher {
  int device = 0;
  cudaStream_t stream = nullptr;
  uint8_t* d_buf = nullptr;   // device arena
  uint8_t* h_buf = nullptr;   // pinned mirror of its upload prefix
  size_t cap = 0, off = 0;
  uint32_t* d_cand = nullptr; // candidate lists (grown on demand)
  size_t cand_cap = 0;
  int32_t* h_small = nullptr; // pinned, 16 ints
  int64_t launches = 0;
  uint8_t* d_tc = nullptr;     // scratch of the tensor-core brute force (expanded descriptors, partial keys)
  size_t tc_cap = 0;
  unsigned* h_tc_err = nullptr;  // pinned: error word of the last tensor-core call
  bool tc_pending = false;
  int sms = 0;
  BusyState busy;      // one call at a time per handle (common.cuh)
};

int matcher_tc_launch(Matcher* m, const uint8_t* d_query, int nq, const uint8_t* d_map, int nm, int64_t index_base,
                      uint64_t* d_best, uint64_t* d_second);  // matcher.cu

inline int arena_reserve(Matcher* m, size_t bytes) {
  m->off = 0;
  if (bytes <= m->cap) return ORBFE_OK;
  if (m->d_buf) cudaFree(m->d_buf);
  if (m->h_buf) cudaFreeHost(m->h_buf);
  m->d_buf = nullptr; m->h_buf = nullptr; m->cap = 0;
  const size_t want = align_up(bytes + bytes / 4, 1 << 16);
  ORBFE_CUDA(cudaMalloc(reinterpret_cast<void**>(&m->d_buf), want));
  ORBFE_CUDA(cudaHostAlloc(reinterpret_cast<void**>(&m->h_buf), want, cudaHostAllocDefault));
  m->cap = want;
  return ORBFE_OK;
}

// copy n elements into the staging buffer; returns the matching DEVICE pointer
template <typename T>
inline const T* arena_put(Matcher* m, const T* src, size_t n) {
  m->off = align_up(m->off, 256);
  T* d = reinterpret_cast<T*>(m->d_buf + m->off);
  if (n) memcpy(m->h_buf + m->off, src, n * sizeof(T));
  m->off += n * sizeof(T);
  return d;
}

template <typename T>
inline T* arena_alloc(Matcher* m, size_t n) {
  m->off = align_up(m->off, 256);
  T* d = reinterpret_cast<T*>(m->d_buf + m->off);
  m->off += n * sizeof(T);
  return d;
}

inline size_t padded(size_t bytes) { return align_up(bytes, 256) + 256; }

inline int ensure_cand(Matcher* m, size_t n) {
  if (n <= m->cand_cap) return ORBFE_OK;
  if (m->d_cand) cudaFree(m->d_cand);
  m->d_cand = nullptr; m->cand_cap = 0;
  const size_t want = n + n / 2 + 1024;
  ORBFE_CUDA(cudaMalloc(reinterpret_cast<void**>(&m->d_cand), want * sizeof(uint32_t)));
  m->cand_cap = want;
  return ORBFE_OK;
}

inline size_t frame_bytes(const orbfe_frame_view* f) {
  const size_t n = (size_t)f->n;
  return padded(n * sizeof(orbfe_keypoint)) + padded(n * 4) + padded(n * 32) + padded(n * 4) +
         padded((size_t)f->nlevels * 4);
}

inline void put_frame(Matcher* m, const orbfe_frame_view* f, DevFrame* d) {
  d->n = f->n;
  d->keys = arena_put(m, f->keys_un, f->n);
  d->u_right = arena_put(m, f->u_right, f->n);
  d->desc = arena_put(m, f->desc, (size_t)f->n * 32);
  d->mp_obs = arena_put(m, f->mp_obs, f->n);
  d->scale = arena_put(m, f->scale_factors, f->nlevels);
  d->min_x = f->min_x; d->min_y = f->min_y; d->max_x = f->max_x; d->max_y = f->max_y;
  d->inv_w = f->grid_inv_w; d->inv_h = f->grid_inv_h;
  d->nlevels = f->nlevels;
}

inline int check_frame(const orbfe_frame_view* f) {
  ORBFE_ARG(f && f->n >= 0 && f->n <= 65535);
  ORBFE_ARG(f->n == 0 || (f->keys_un && f->u_right && f->desc && f->mp_obs));
  ORBFE_ARG(f->scale_factors && f->nlevels >= 1 && f->nlevels <= 16);
  return ORBFE_OK;
}

// scratch shared by both SearchByProjection overloads
inline size_t scratch_bytes(int rows, int nfeat) {
  return padded(((size_t)rows + 1) * 4) + padded((size_t)rows * 4) + 3 * padded((size_t)nfeat * 4) + padded(64) +
         padded(((size_t)ORBFE_GRID_COLS * ORBFE_GRID_ROWS + 1) * 4) + padded((size_t)nfeat * 4);
}

inline void alloc_scratch(Matcher* m, int rows, int nfeat, DevFrame* f, SearchScratch* sc) {
  sc->cand_count = arena_alloc<int32_t>(m, (size_t)rows + 1);
  sc->result = arena_alloc<int32_t>(m, rows);
  sc->block_a = arena_alloc<int32_t>(m, nfeat);
  sc->block_b = arena_alloc<int32_t>(m, nfeat);
  sc->assigned = arena_alloc<int32_t>(m, nfeat);
  sc->counters = arena_alloc<int32_t>(m, 16);
  sc->cand_capacity_needed = sc->counters + 8;
  f->cell_start = arena_alloc<int32_t>(m, (size_t)ORBFE_GRID_COLS * ORBFE_GRID_ROWS + 1);
  f->cell_items = arena_alloc<int32_t>(m, nfeat);
  sc->cand = nullptr;
  sc->cand_cap = 0;
  sc->proj = nullptr;
  sc->hist = nullptr;
}

// shared tail of the two SearchByProjection overloads: scan, fill, resolve, read back
inline int finish_projection(Matcher* m, DevFrame& df, int rows, const int32_t* row_obs, const float* row_angle,
                             int mode, float nnratio, int check_ori, int accept_th, SearchScratch& sc,
                             void (*launch_fill)(void*, int), void* ctx, int32_t* assigned, int32_t* n_matches) {
  launch_scan_counts(sc.cand_count, rows, sc.cand_capacity_needed, m->stream);
  // no host round trip for the candidate total: the fill pass and the resolve step run against the buffer as it is
  // (sized by earlier calls, at least 256 K entries) and leave at once if the total does not fit; only then does the
  // host grow the buffer and repeat the two steps
  int rc = ensure_cand(m, (size_t)1 << 18);
  if (rc != ORBFE_OK) return rc;
  for (int attempt = 0; attempt < 2; ++attempt) {
    sc.cand = m->d_cand;
    sc.cand_cap = (int)std::min<size_t>(m->cand_cap, (size_t)0x7FFFFFFF);
    launch_fill(ctx, 1);
    launch_resolve(df, rows, row_obs, row_angle, mode, nnratio, check_ori, accept_th, sc, m->stream);
    m->launches += attempt == 0 ? 3 : 2;
    if (df.n > 0) ORBFE_CUDA(cudaMemcpyAsync(assigned, sc.assigned, (size_t)df.n * 4, cudaMemcpyDeviceToHost, m->stream));
    ORBFE_CUDA(cudaMemcpyAsync(m->h_small, sc.counters, 16, cudaMemcpyDeviceToHost, m->stream));
    ORBFE_CUDA(cudaMemcpyAsync(m->h_small + 8, sc.cand_capacity_needed, 4, cudaMemcpyDeviceToHost, m->stream));
    ORBFE_CUDA(cudaStreamSynchronize(m->stream));
    ORBFE_CUDA(cudaGetLastError());
    const int total = m->h_small[8];
    if (total <= sc.cand_cap) break;
    if (attempt == 1) { set_error("candidate buffer still too small (%d > %d)", total, sc.cand_cap); return ORBFE_ERR_CAPACITY; }
    rc = ensure_cand(m, (size_t)total + 16);
    if (rc != ORBFE_OK) return rc;
  }
  if (n_matches) *n_matches = m->h_small[0];
  return ORBFE_OK;
}

inline int check_kf(const orbfe_keyframe_view* k) {
  ORBFE_ARG(k && k->n >= 0 && k->n_nodes >= 0);
  ORBFE_ARG(k->n == 0 || (k->keys_un && k->u_right && k->desc && k->has_mp));
  ORBFE_ARG(k->node_offsets && (k->n_nodes == 0 || (k->node_ids && k->node_features)));
  ORBFE_ARG(k->scale_factors && k->level_sigma2 && k->nlevels >= 1 && k->nlevels <= 16);
  return ORBFE_OK;
}

inline size_t kf_bytes(const orbfe_keyframe_view* k) {
  const size_t n = (size_t)k->n, e = (size_t)k->node_offsets[k->n_nodes];
  return padded(n * sizeof(orbfe_keypoint)) + padded(n * 4) + padded(n * 32) + padded(n) +
         2 * padded(((size_t)k->n_nodes + 1) * 4) + padded(e * 4) + 2 * padded((size_t)k->nlevels * 4);
}

inline void put_kf(Matcher* m, const orbfe_keyframe_view* k, DevKeyFrame* d) {
  d->n = k->n;
  d->keys = arena_put(m, k->keys_un, k->n);
  d->u_right = arena_put(m, k->u_right, k->n);
  d->desc = arena_put(m, k->desc, (size_t)k->n * 32);
  d->has_mp = arena_put(m, k->has_mp, k->n);
  d->n_nodes = k->n_nodes;
  d->node_ids = arena_put(m, k->node_ids, k->n_nodes);
  d->node_offsets = arena_put(m, k->node_offsets, (size_t)k->n_nodes + 1);
  d->node_features = arena_put(m, k->node_features, (size_t)k->node_offsets[k->n_nodes]);
  d->scale = arena_put(m, k->scale_factors, k->nlevels);
  d->sigma2 = arena_put(m, k->level_sigma2, k->nlevels);
}

}  // namespace orbfe
