// matcher.cu — host side of the ORBmatcher entry points of include/orbfe.h.
// A matcher context owns a stream, one pinned staging buffer and one device buffer;
// every call packs the caller's struct-of-array views into the staging buffer, uploads
// them with ONE copy, runs the kernels of k_match.cu / k_search.cu and reads back the
// small result arrays.  No result is computed on the host.
#include "matcher_ctx.cuh"

using namespace orbfe;

namespace orbfe {
// the tensor-core brute force on the matcher's stream (scratch grown on demand); shared with comm.cu
int matcher_tc_launch(Matcher* m, const uint8_t* d_query, int nq, const uint8_t* d_map, int nm, int64_t index_base,
                      uint64_t* d_best, uint64_t* d_second) {
  if (m->sms == 0) ORBFE_CUDA(cudaDeviceGetAttribute(&m->sms, cudaDevAttrMultiProcessorCount, m->device));
  int tps = 0;
  const int slices = hamming_mma_slices(nq, nm, m->sms, &tps);
  const size_t need = hamming_mma_scratch_bytes(nq, nm, slices);
  if (need > m->tc_cap) {
    if (m->d_tc) { ORBFE_CUDA(cudaStreamSynchronize(m->stream)); cudaFree(m->d_tc); m->d_tc = nullptr; m->tc_cap = 0; }
    ORBFE_CUDA(cudaMalloc(reinterpret_cast<void**>(&m->d_tc), need));
    m->tc_cap = need;
  }
  if (!m->h_tc_err) ORBFE_CUDA(cudaHostAlloc(reinterpret_cast<void**>(&m->h_tc_err), 64, cudaHostAllocDefault));
  unsigned* d_err = nullptr;
  const int rc = launch_hamming_mma(d_query, nq, d_map, nm, index_base, m->d_tc, slices, tps, m->sms,
                                    reinterpret_cast<unsigned long long*>(d_best),
                                    reinterpret_cast<unsigned long long*>(d_second), &d_err, m->stream);
  if (rc != 0) {
    set_error("tensor-core brute force: set-up failed (%d): %s", rc, cudaGetErrorString(cudaGetLastError()));
    return ORBFE_ERR_CUDA;
  }
  ORBFE_CUDA(cudaMemcpyAsync(m->h_tc_err, d_err, 4, cudaMemcpyDeviceToHost, m->stream));
  m->tc_pending = true;
  m->launches += 5;
  return ORBFE_OK;
}
}  // namespace orbfe

extern "C" {

int orbfe_matcher_create(int device, orbfe_matcher** out) {
  ORBFE_ARG(out);
  *out = nullptr;
  int ndev = 0;
  ORBFE_CUDA(cudaGetDeviceCount(&ndev));
  ORBFE_ARG(device >= 0 && device < ndev);
  Matcher* m = new (std::nothrow) Matcher();
  ORBFE_ARG(m != nullptr);
  m->device = device;
  DeviceGuard guard(device);
  if (cudaStreamCreateWithFlags(&m->stream, cudaStreamNonBlocking) != cudaSuccess ||
      cudaHostAlloc(reinterpret_cast<void**>(&m->h_small), 64, cudaHostAllocDefault) != cudaSuccess) {
    set_error("matcher: stream / pinned allocation failed");
    delete m;
    return ORBFE_ERR_CUDA;
  }
  *out = reinterpret_cast<orbfe_matcher*>(m);
  return ORBFE_OK;
}

void orbfe_matcher_destroy(orbfe_matcher* h) {
  if (!h) return;
  Matcher* m = reinterpret_cast<Matcher*>(h);
  DeviceGuard guard(m->device);
  cudaStreamSynchronize(m->stream);
  if (m->d_buf) cudaFree(m->d_buf);
  if (m->h_buf) cudaFreeHost(m->h_buf);
  if (m->d_cand) cudaFree(m->d_cand);
  if (m->d_tc) cudaFree(m->d_tc);
  if (m->h_tc_err) cudaFreeHost(m->h_tc_err);
  if (m->h_small) cudaFreeHost(m->h_small);
  cudaStreamDestroy(m->stream);
  delete m;
}

int64_t orbfe_matcher_launch_count(const orbfe_matcher* h) { return h ? reinterpret_cast<const Matcher*>(h)->launches : 0; }
void* orbfe_matcher_stream(orbfe_matcher* h) { return h ? reinterpret_cast<Matcher*>(h)->stream : nullptr; }

int orbfe_matcher_sync(orbfe_matcher* h) {
  ORBFE_ARG(h);
  Matcher* m = reinterpret_cast<Matcher*>(h);
  ORBFE_MATCHER_ENTER(m);
  ORBFE_CUDA(cudaStreamSynchronize(m->stream));
  if (m->tc_pending) {  // the tensor-core brute force reports a pipeline time-out through a pinned word
    m->tc_pending = false;
    if (m->h_tc_err && *reinterpret_cast<volatile unsigned*>(m->h_tc_err) != 0) {
      set_error("orbfe_hamming_bruteforce_tc_device: the tcgen05 pipeline timed out (internal error); use the POPC path");
      return ORBFE_ERR_CUDA;
    }
  }
  return ORBFE_OK;
}

int orbfe_hamming_bruteforce_tc_device(orbfe_matcher* h, const uint8_t* d_query, int nq, const uint8_t* d_map, int nm,
                                       int64_t index_base, uint64_t* d_best, uint64_t* d_second) {
  ORBFE_ARG(h && d_query && d_best && d_second && nq >= 1 && nm >= 1 && d_map);
  ORBFE_ARG(index_base >= 0 && index_base + (int64_t)nm <= ((int64_t)1 << 32));
  Matcher* m = reinterpret_cast<Matcher*>(h);
  ORBFE_MATCHER_ENTER(m);
  return matcher_tc_launch(m, d_query, nq, d_map, nm, index_base, d_best, d_second);
}

int orbfe_descriptor_distance(orbfe_matcher* h, const uint8_t* a, const uint8_t* b, int n, int32_t* out) {
  ORBFE_ARG(h && n >= 0);
  if (n == 0) return ORBFE_OK;
  ORBFE_ARG(a && b && out);
  Matcher* m = reinterpret_cast<Matcher*>(h);
  ORBFE_MATCHER_ENTER(m);
  int rc = arena_reserve(m, 2 * padded((size_t)n * 32) + padded((size_t)n * 4));
  if (rc != ORBFE_OK) return rc;
  const uint8_t* da = arena_put(m, a, (size_t)n * 32);
  const uint8_t* db = arena_put(m, b, (size_t)n * 32);
  const size_t up = m->off;
  int32_t* dout = arena_alloc<int32_t>(m, n);
  ORBFE_CUDA(cudaMemcpyAsync(m->d_buf, m->h_buf, up, cudaMemcpyHostToDevice, m->stream));
  launch_descriptor_distance(da, db, n, dout, m->stream);
  m->launches += 1;
  ORBFE_CUDA(cudaMemcpyAsync(out, dout, (size_t)n * 4, cudaMemcpyDeviceToHost, m->stream));
  ORBFE_CUDA(cudaStreamSynchronize(m->stream));
  ORBFE_CUDA(cudaGetLastError());
  return ORBFE_OK;
}

int orbfe_compute_distinctive_descriptors(orbfe_matcher* h, const uint8_t* desc, const int32_t* offsets, int n_points,
                                          int32_t* best_idx, uint8_t* best_desc) {
  ORBFE_ARG(h && n_points >= 0);
  if (n_points == 0) return ORBFE_OK;
  ORBFE_ARG(offsets && best_idx && offsets[0] == 0);
  for (int p = 0; p < n_points; ++p) ORBFE_ARG(offsets[p + 1] >= offsets[p]);
  const size_t total = (size_t)offsets[n_points];
  ORBFE_ARG(total == 0 || desc);
  Matcher* m = reinterpret_cast<Matcher*>(h);
  ORBFE_MATCHER_ENTER(m);
  const size_t P = (size_t)n_points;
  int rc = arena_reserve(m, padded(total * 32) + padded((P + 1) * 4) + padded(P * 4) + padded(P * 32));
  if (rc != ORBFE_OK) return rc;
  const uint8_t* dd = arena_put(m, desc, total * 32);
  const int32_t* doff = arena_put(m, offsets, P + 1);
  const size_t up = m->off;
  int32_t* dbest = arena_alloc<int32_t>(m, P);
  uint8_t* dbd = best_desc ? arena_alloc<uint8_t>(m, P * 32) : nullptr;
  ORBFE_CUDA(cudaMemcpyAsync(m->d_buf, m->h_buf, up, cudaMemcpyHostToDevice, m->stream));
  if (dbd) ORBFE_CUDA(cudaMemsetAsync(dbd, 0, P * 32, m->stream));
  launch_distinctive(dd, doff, n_points, dbest, dbd, m->stream);
  m->launches += 1;
  ORBFE_CUDA(cudaMemcpyAsync(best_idx, dbest, P * 4, cudaMemcpyDeviceToHost, m->stream));
  if (dbd) ORBFE_CUDA(cudaMemcpyAsync(best_desc, dbd, P * 32, cudaMemcpyDeviceToHost, m->stream));
  ORBFE_CUDA(cudaStreamSynchronize(m->stream));
  ORBFE_CUDA(cudaGetLastError());
  return ORBFE_OK;
}

int orbfe_hamming_merge_device(orbfe_matcher* h, const uint64_t* d_best_all, const uint64_t* d_second_all, int shards,
                               int nq, uint64_t* d_best, uint64_t* d_second) {
  ORBFE_ARG(h && d_best_all && d_second_all && d_best && d_second && shards >= 1 && nq >= 1);
  Matcher* m = reinterpret_cast<Matcher*>(h);
  ORBFE_MATCHER_ENTER(m);
  launch_hamming_merge(reinterpret_cast<const unsigned long long*>(d_best_all),
                       reinterpret_cast<const unsigned long long*>(d_second_all), shards, nq,
                       reinterpret_cast<unsigned long long*>(d_best), reinterpret_cast<unsigned long long*>(d_second),
                       m->stream);
  m->launches += 1;
  ORBFE_CUDA(cudaGetLastError());
  return ORBFE_OK;
}

int orbfe_measure_popc_peak(orbfe_matcher* h, double* popc32_per_second) {
  ORBFE_ARG(h && popc32_per_second);
  Matcher* m = reinterpret_cast<Matcher*>(h);
  ORBFE_MATCHER_ENTER(m);
  ORBFE_CUDA(cudaStreamSynchronize(m->stream));
  *popc32_per_second = measure_popc_peak(m->stream);
  ORBFE_CUDA(cudaGetLastError());
  return *popc32_per_second > 0.0 ? ORBFE_OK : ORBFE_ERR_CUDA;
}

int orbfe_hamming_bruteforce_device(orbfe_matcher* h, const uint8_t* d_query, int nq, const uint8_t* d_map, int nm,
                                    int64_t index_base, uint64_t* d_best, uint64_t* d_second) {
  ORBFE_ARG(h && d_query && d_best && d_second && nq >= 1 && nm >= 0);
  ORBFE_ARG(nm == 0 || d_map);
  Matcher* m = reinterpret_cast<Matcher*>(h);
  ORBFE_MATCHER_ENTER(m);
  int slice_len = 0;
  const int slices = nm > 0 ? bruteforce_slices(nq, nm, &slice_len) : 0;
  // the partial results live in the candidate buffer (2 x slices x nq u64)
  const size_t part_words = (size_t)std::max(slices, 1) * nq * 2 /*u64 -> u32*/ * 2;
  int rc = ensure_cand(m, part_words + 64);
  if (rc != ORBFE_OK) return rc;
  unsigned long long* pb = reinterpret_cast<unsigned long long*>(m->d_cand);
  unsigned long long* ps = pb + (size_t)std::max(slices, 1) * nq;
  if (slices == 0) {
    ORBFE_CUDA(cudaMemsetAsync(d_best, 0xFF, (size_t)nq * 8, m->stream));
    ORBFE_CUDA(cudaMemsetAsync(d_second, 0xFF, (size_t)nq * 8, m->stream));
    return ORBFE_OK;
  }
  launch_hamming_bruteforce(d_query, nq, d_map, nm, slice_len, slices, index_base, pb, ps, m->stream);
  launch_hamming_merge(pb, ps, slices, nq, reinterpret_cast<unsigned long long*>(d_best),
                       reinterpret_cast<unsigned long long*>(d_second), m->stream);
  m->launches += 2;
  ORBFE_CUDA(cudaGetLastError());
  return ORBFE_OK;
}

int orbfe_hamming_bruteforce(orbfe_matcher* h, const uint8_t* query, int nq, const uint8_t* map, int nm,
                             int64_t index_base, uint64_t* best, uint64_t* second) {
  ORBFE_ARG(h && nq >= 0 && nm >= 0);
  if (nq == 0) return ORBFE_OK;
  ORBFE_ARG(query && best && second && (nm == 0 || map));
  Matcher* m = reinterpret_cast<Matcher*>(h);
  ORBFE_MATCHER_ENTER(m);
  int rc = arena_reserve(m, padded((size_t)nq * 32) + padded((size_t)nm * 32) + 2 * padded((size_t)nq * 8));
  if (rc != ORBFE_OK) return rc;
  const uint8_t* dq = arena_put(m, query, (size_t)nq * 32);
  const uint8_t* dm = arena_put(m, map, (size_t)nm * 32);
  const size_t up = m->off;
  uint64_t* db = arena_alloc<uint64_t>(m, nq);
  uint64_t* ds = arena_alloc<uint64_t>(m, nq);
  ORBFE_CUDA(cudaMemcpyAsync(m->d_buf, m->h_buf, up, cudaMemcpyHostToDevice, m->stream));
  rc = orbfe_hamming_bruteforce_device(h, dq, nq, dm, nm, index_base, db, ds);
  if (rc != ORBFE_OK) return rc;
  ORBFE_CUDA(cudaMemcpyAsync(best, db, (size_t)nq * 8, cudaMemcpyDeviceToHost, m->stream));
  ORBFE_CUDA(cudaMemcpyAsync(second, ds, (size_t)nq * 8, cudaMemcpyDeviceToHost, m->stream));
  ORBFE_CUDA(cudaStreamSynchronize(m->stream));
  return ORBFE_OK;
}

int orbfe_search_by_projection_map(orbfe_matcher* h, const orbfe_frame_view* frame, const orbfe_mappoints_view* mps,
                                   float th, float nnratio, int32_t* assigned, int32_t* n_matches) {
  ORBFE_ARG(h && mps && mps->m >= 0);
  int rc = check_frame(frame);
  if (rc != ORBFE_OK) return rc;
  ORBFE_ARG(frame->n == 0 || assigned);
  if (n_matches) *n_matches = 0;
  for (int i = 0; i < frame->n; ++i) assigned[i] = -1;
  if (mps->m == 0 || frame->n == 0) return ORBFE_OK;
  ORBFE_ARG(mps->proj_x && mps->proj_y && mps->proj_xr && mps->level && mps->view_cos && mps->in_view && mps->obs &&
            mps->desc);
  Matcher* m = reinterpret_cast<Matcher*>(h);
  ORBFE_MATCHER_ENTER(m);
  const size_t M = (size_t)mps->m;
  rc = arena_reserve(m, frame_bytes(frame) + 6 * padded(M * 4) + padded(M) + padded(M * 32) +
                            scratch_bytes(mps->m, frame->n));
  if (rc != ORBFE_OK) return rc;
  DevFrame f;
  put_frame(m, frame, &f);
  DevMapPoints d;
  d.m = mps->m;
  d.proj_x = arena_put(m, mps->proj_x, M);
  d.proj_y = arena_put(m, mps->proj_y, M);
  d.proj_xr = arena_put(m, mps->proj_xr, M);
  d.view_cos = arena_put(m, mps->view_cos, M);
  d.level = arena_put(m, mps->level, M);
  d.obs = arena_put(m, mps->obs, M);
  d.in_view = arena_put(m, mps->in_view, M);
  d.desc = arena_put(m, mps->desc, M * 32);
  const size_t up = m->off;
  SearchScratch sc;
  alloc_scratch(m, mps->m, frame->n, &f, &sc);
  ORBFE_CUDA(cudaMemcpyAsync(m->d_buf, m->h_buf, up, cudaMemcpyHostToDevice, m->stream));
  launch_build_grid(f, m->stream);
  launch_sbp_map_candidates(f, d, th, sc, 0, m->stream);
  m->launches += 2;
  struct Ctx { Matcher* m; DevFrame* f; DevMapPoints* d; float th; SearchScratch* sc; } ctx = {m, &f, &d, th, &sc};
  auto fill = [](void* p, int fl) {
    Ctx* c = static_cast<Ctx*>(p);
    launch_sbp_map_candidates(*c->f, *c->d, c->th, *c->sc, fl, c->m->stream);
  };
  return finish_projection(m, f, mps->m, d.obs, nullptr, 0, nnratio, 0, ORBFE_TH_HIGH, sc, fill, &ctx, assigned,
                           n_matches);
}

int orbfe_search_by_projection_last(orbfe_matcher* h, const orbfe_frame_view* cur, const orbfe_camera_pose* cam,
                                    const orbfe_lastframe_view* last, float th, int mono, int check_ori,
                                    int32_t* assigned, int32_t* n_matches) {
  ORBFE_ARG(h && cam && last && last->n >= 0);
  int rc = check_frame(cur);
  if (rc != ORBFE_OK) return rc;
  ORBFE_ARG(cur->n == 0 || assigned);
  if (n_matches) *n_matches = 0;
  for (int i = 0; i < cur->n; ++i) assigned[i] = -1;
  if (last->n == 0 || cur->n == 0) return ORBFE_OK;
  ORBFE_ARG(last->valid && last->world_pos && last->obs && last->desc && last->octave && last->angle);
  Matcher* m = reinterpret_cast<Matcher*>(h);
  ORBFE_MATCHER_ENTER(m);
  const size_t N = (size_t)last->n;
  rc = arena_reserve(m, frame_bytes(cur) + padded(N) + padded(N * 12) + 3 * padded(N * 4) + padded(N * 32) +
                            scratch_bytes(last->n, cur->n));
  if (rc != ORBFE_OK) return rc;
  DevFrame f;
  put_frame(m, cur, &f);
  DevLastFrame d;
  d.n = last->n;
  d.valid = arena_put(m, last->valid, N);
  d.world_pos = arena_put(m, last->world_pos, N * 3);
  d.obs = arena_put(m, last->obs, N);
  d.octave = arena_put(m, last->octave, N);
  d.angle = arena_put(m, last->angle, N);
  d.desc = arena_put(m, last->desc, N * 32);
  d.tlc_z = host_tlc_z(cam->tcw, last->tcw);  // pose algebra on 24 scalars (src/ORBmatcher.cc:1403-1413)
  const size_t up = m->off;
  SearchScratch sc;
  alloc_scratch(m, last->n, cur->n, &f, &sc);
  ORBFE_CUDA(cudaMemcpyAsync(m->d_buf, m->h_buf, up, cudaMemcpyHostToDevice, m->stream));
  launch_build_grid(f, m->stream);
  launch_sbp_last_candidates(f, d, *cam, th, mono, sc, 0, m->stream);
  m->launches += 2;
  struct Ctx { Matcher* m; DevFrame* f; DevLastFrame* d; const orbfe_camera_pose* cam; float th; int mono; SearchScratch* sc; }
      ctx = {m, &f, &d, cam, th, mono, &sc};
  auto fill = [](void* p, int fl) {
    Ctx* c = static_cast<Ctx*>(p);
    launch_sbp_last_candidates(*c->f, *c->d, *c->cam, c->th, c->mono, *c->sc, fl, c->m->stream);
  };
  return finish_projection(m, f, last->n, d.obs, d.angle, 1, 0.f, check_ori, ORBFE_TH_HIGH, sc, fill, &ctx, assigned,
                           n_matches);
}

int orbfe_search_for_triangulation(orbfe_matcher* h, const orbfe_keyframe_view* kf1, const orbfe_keyframe_view* kf2,
                                   const float* F12, float ex, float ey, int only_stereo, int check_ori, int32_t* pairs,
                                   int cap, int32_t* n_pairs) {
  ORBFE_ARG(h && F12 && n_pairs && cap >= 0 && (cap == 0 || pairs));
  int rc = check_kf(kf1);
  if (rc != ORBFE_OK) return rc;
  rc = check_kf(kf2);
  if (rc != ORBFE_OK) return rc;
  *n_pairs = 0;
  if (kf1->n == 0 || kf2->n == 0 || kf1->n_nodes == 0 || kf2->n_nodes == 0) return ORBFE_OK;
  Matcher* m = reinterpret_cast<Matcher*>(h);
  ORBFE_MATCHER_ENTER(m);
  const size_t n1 = (size_t)kf1->n;
  const int out_cap = (int)std::min<size_t>((size_t)cap, n1);
  rc = arena_reserve(m, kf_bytes(kf1) + kf_bytes(kf2) + padded(64) + 2 * padded((n1 + 1) * 4) +
                            padded((size_t)out_cap * 8 + 8) + padded(64));
  if (rc != ORBFE_OK) return rc;
  DevKeyFrame a, b;
  put_kf(m, kf1, &a);
  put_kf(m, kf2, &b);
  const float* dF = arena_put(m, F12, 9);
  const size_t up = m->off;
  int32_t* matches12 = arena_alloc<int32_t>(m, n1);
  int32_t* flags = arena_alloc<int32_t>(m, n1 + 1);
  int32_t* dpairs = arena_alloc<int32_t>(m, (size_t)out_cap * 2 + 2);
  int32_t* dn = arena_alloc<int32_t>(m, 16);
  ORBFE_CUDA(cudaMemcpyAsync(m->d_buf, m->h_buf, up, cudaMemcpyHostToDevice, m->stream));
  launch_triangulation(a, b, dF, ex, ey, only_stereo, check_ori, matches12, flags, dpairs, out_cap, dn,
                       kf1->node_offsets[kf1->n_nodes], m->stream);
  m->launches += 3;
  ORBFE_CUDA(cudaMemcpyAsync(m->h_small, dn, 4, cudaMemcpyDeviceToHost, m->stream));
  if (out_cap > 0)
    ORBFE_CUDA(cudaMemcpyAsync(pairs, dpairs, (size_t)out_cap * 8, cudaMemcpyDeviceToHost, m->stream));
  ORBFE_CUDA(cudaStreamSynchronize(m->stream));
  ORBFE_CUDA(cudaGetLastError());
  *n_pairs = m->h_small[0];
  return ORBFE_OK;
}

}  // extern "C"
