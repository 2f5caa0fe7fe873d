// matcher2.cu — host side of the seven other ORBmatcher entry points of include/orbfe.h:
// SearchByBoW x2, SearchForInitialization, the relocalisation and Sim3 SearchByProjection overloads, Fuse x2 and
// SearchBySim3.  Same shape as matcher.cu: pack the caller's views into the arena, ONE upload, kernels of
// k_search.cu / k_search2.cu, small results back.  The only host arithmetic is pose algebra on a handful of scalars
// (decomposing a Sim3, composing the two transforms of SearchBySim3), as host_tlc_z does for the last-frame overload.
#include <cmath>

#include "matcher_ctx.cuh"

namespace orbfe {

// Rcw / tcw / Ow as src/ORBmatcher.cc:329-334, :1041-1046, :1544-1546 compute them from Scw or Tcw.  cv::Mat
// conventions: Mat::dot accumulates in double, `M/s` multiplies by (float)(1.0/s), a product with a transposed
// operand accumulates in double.
static void decompose_pose(const orbfe_projection* pr, DevProjection* P) {
  const float* T = pr->tcw;
  float scale = 1.0f;
  if (pr->is_sim3) {
    const double d = ((double)T[0] * (double)T[0] + (double)T[1] * (double)T[1]) + (double)T[2] * (double)T[2];
    const float scw = (float)std::sqrt(d);
    scale = (float)(1.0 / (double)scw);
  }
  for (int r = 0; r < 3; ++r) {
    for (int c = 0; c < 3; ++c) P->R[r * 3 + c] = pr->is_sim3 ? T[r * 4 + c] * scale : T[r * 4 + c];
    P->t[r] = pr->is_sim3 ? T[r * 4 + 3] * scale : T[r * 4 + 3];
  }
  for (int i = 0; i < 3; ++i) {
    if (pr->has_ow) {
      P->Ow[i] = pr->ow[i];
    } else {
      double s = 0;
      for (int k = 0; k < 3; ++k) s += (double)P->R[k * 3 + i] * (double)P->t[k];
      P->Ow[i] = (float)(-1.0 * s);
    }
  }
  for (int i = 0; i < 9; ++i) P->R2[i] = 0.f;
  for (int i = 0; i < 3; ++i) P->t2[i] = 0.f;
  P->fx = pr->fx; P->fy = pr->fy; P->cx = pr->cx; P->cy = pr->cy; P->bf = pr->bf;
  P->log_sf = pr->log_scale_factor;
}

static int check_points(const orbfe_points3d_view* p, bool need_normal, bool need_angle) {
  ORBFE_ARG(p && p->m >= 0);
  ORBFE_ARG(p->m == 0 || (p->valid && p->world_pos && p->max_dist_inv && p->min_dist_inv && p->max_distance && p->desc));
  ORBFE_ARG(p->m == 0 || !need_normal || p->normal);
  ORBFE_ARG(p->m == 0 || !need_angle || p->angle);
  return ORBFE_OK;
}

static size_t points_bytes(const orbfe_points3d_view* p) {
  const size_t m = (size_t)p->m;
  return padded(m) + 2 * padded(m * 12) + 4 * padded(m * 4) + padded(m * 32);
}

static void put_points(Matcher* m, const orbfe_points3d_view* p, DevPoints* d) {
  const size_t n = (size_t)p->m;
  d->m = p->m;
  d->valid = arena_put(m, p->valid, n);
  d->world_pos = arena_put(m, p->world_pos, n * 3);
  d->normal = p->normal ? arena_put(m, p->normal, n * 3) : nullptr;
  d->max_inv = arena_put(m, p->max_dist_inv, n);
  d->min_inv = arena_put(m, p->min_dist_inv, n);
  d->max_d = arena_put(m, p->max_distance, n);
  d->angle = p->angle ? arena_put(m, p->angle, n) : nullptr;
  d->desc = arena_put(m, p->desc, n * 32);
}

static size_t grid_bytes(int nfeat) {
  return padded(((size_t)ORBFE_GRID_COLS * ORBFE_GRID_ROWS + 1) * 4) + padded((size_t)nfeat * 4);
}

static void alloc_grid(Matcher* m, int nfeat, DevFrame* f) {
  f->cell_start = arena_alloc<int32_t>(m, (size_t)ORBFE_GRID_COLS * ORBFE_GRID_ROWS + 1);
  f->cell_items = arena_alloc<int32_t>(m, nfeat);
}

// the two projection overloads with a claim chain
static int projection_claims(orbfe_matcher* h, int mode, const orbfe_frame_view* frame, const orbfe_projection* pr,
                             const orbfe_points3d_view* pts, float th, int accept_th, int check_ori, int32_t* assigned,
                             int32_t* n_matches) {
  ORBFE_ARG(h && pr);
  int rc = check_frame(frame);
  if (rc != ORBFE_OK) return rc;
  rc = check_points(pts, mode == PM_SIM3, mode == PM_RELOC && check_ori);
  if (rc != ORBFE_OK) return rc;
  ORBFE_ARG(frame->n == 0 || assigned);
  if (n_matches) *n_matches = 0;
  for (int i = 0; i < frame->n; ++i) assigned[i] = -1;
  if (pts->m == 0 || frame->n == 0) return ORBFE_OK;
  Matcher* m = reinterpret_cast<Matcher*>(h);
  ORBFE_MATCHER_ENTER(m);
  rc = arena_reserve(m, frame_bytes(frame) + points_bytes(pts) + scratch_bytes(pts->m, frame->n));
  if (rc != ORBFE_OK) return rc;
  DevFrame f;
  put_frame(m, frame, &f);
  DevPoints d;
  put_points(m, pts, &d);
  DevProjection P;
  decompose_pose(pr, &P);
  const size_t up = m->off;
  SearchScratch sc;
  alloc_scratch(m, pts->m, frame->n, &f, &sc);
  ORBFE_CUDA(cudaMemcpyAsync(m->d_buf, m->h_buf, up, cudaMemcpyHostToDevice, m->stream));
  launch_build_grid(f, m->stream);
  launch_proj_candidates(mode, f, d, P, th, sc, 0, m->stream);
  m->launches += 2;
  struct Ctx { Matcher* m; int mode; DevFrame* f; DevPoints* d; DevProjection* P; float th; SearchScratch* sc; }
      ctx = {m, mode, &f, &d, &P, th, &sc};
  auto fill = [](void* p, int fl) {
    Ctx* c = static_cast<Ctx*>(p);
    launch_proj_candidates(c->mode, *c->f, *c->d, *c->P, c->th, *c->sc, fl, c->m->stream);
  };
  return finish_projection(m, f, pts->m, nullptr, d.angle, 1, 0.f, mode == PM_RELOC ? check_ori : 0, accept_th, sc, fill,
                           &ctx, assigned, n_matches);
}

}  // namespace orbfe

using namespace orbfe;

extern "C" {

int orbfe_search_by_projection_reloc(orbfe_matcher* h, const orbfe_frame_view* cur, const orbfe_projection* pr,
                                     const orbfe_points3d_view* pts, float th, int orb_dist, int check_ori,
                                     int32_t* assigned, int32_t* n_matches) {
  return projection_claims(h, PM_RELOC, cur, pr, pts, th, orb_dist, check_ori, assigned, n_matches);
}

int orbfe_search_by_projection_sim3(orbfe_matcher* h, const orbfe_frame_view* kf, const orbfe_projection* pr,
                                    const orbfe_points3d_view* pts, int th, int32_t* assigned, int32_t* n_matches) {
  return projection_claims(h, PM_SIM3, kf, pr, pts, (float)th, ORBFE_TH_LOW, 0, assigned, n_matches);
}

int orbfe_search_local_points(orbfe_matcher* h, const orbfe_frame_view* frame, const orbfe_projection* pr,
                              const orbfe_points3d_view* pts, const int32_t* obs, float viewing_cos_limit, float th,
                              float nnratio, int32_t* assigned, int32_t* n_matches, uint8_t* in_view, float* proj_x,
                              float* proj_y, float* proj_xr, int32_t* level, float* view_cos) {
  ORBFE_ARG(h && pr);
  int rc = check_frame(frame);
  if (rc != ORBFE_OK) return rc;
  rc = check_points(pts, true, false);
  if (rc != ORBFE_OK) return rc;
  ORBFE_ARG(frame->n == 0 || assigned);
  ORBFE_ARG(pts->m == 0 || (obs && in_view));
  if (n_matches) *n_matches = 0;
  for (int i = 0; i < frame->n; ++i) assigned[i] = -1;
  if (pts->m == 0) return ORBFE_OK;
  Matcher* m = reinterpret_cast<Matcher*>(h);
  ORBFE_MATCHER_ENTER(m);
  const size_t M = (size_t)pts->m;
  rc = arena_reserve(m, frame_bytes(frame) + points_bytes(pts) + 7 * padded(M * 4) + scratch_bytes(pts->m, frame->n));
  if (rc != ORBFE_OK) return rc;
  DevFrame f;
  put_frame(m, frame, &f);
  DevPoints d;
  put_points(m, pts, &d);
  const int32_t* d_obs = arena_put(m, obs, M);
  DevProjection P;
  decompose_pose(pr, &P);
  const size_t up = m->off;
  uint8_t* d_in = arena_alloc<uint8_t>(m, M);
  float* d_px = arena_alloc<float>(m, M);
  float* d_py = arena_alloc<float>(m, M);
  float* d_pxr = arena_alloc<float>(m, M);
  int32_t* d_lvl = arena_alloc<int32_t>(m, M);
  float* d_vc = arena_alloc<float>(m, M);
  SearchScratch sc;
  alloc_scratch(m, pts->m, std::max(frame->n, 1), &f, &sc);
  ORBFE_CUDA(cudaMemcpyAsync(m->d_buf, m->h_buf, up, cudaMemcpyHostToDevice, m->stream));
  launch_frustum(f, d, P, viewing_cos_limit, d_in, d_px, d_py, d_pxr, d_lvl, d_vc, m->stream);
  m->launches += 1;
  int result = ORBFE_OK;
  if (frame->n > 0) {
    DevMapPoints mp;  // the fields the map overload reads, now device-resident
    mp.m = pts->m;
    mp.proj_x = d_px; mp.proj_y = d_py; mp.proj_xr = d_pxr; mp.view_cos = d_vc;
    mp.level = d_lvl; mp.obs = d_obs; mp.in_view = d_in; mp.desc = d.desc;
    launch_build_grid(f, m->stream);
    launch_sbp_map_candidates(f, mp, th, sc, 0, m->stream);
    m->launches += 2;
    struct Ctx { Matcher* m; DevFrame* f; DevMapPoints* d; float th; SearchScratch* sc; } ctx = {m, &f, &mp, th, &sc};
    auto fill = [](void* p, int fl) {
      Ctx* c = static_cast<Ctx*>(p);
      launch_sbp_map_candidates(*c->f, *c->d, c->th, *c->sc, fl, c->m->stream);
    };
    result = finish_projection(m, f, pts->m, d_obs, nullptr, 0, nnratio, 0, ORBFE_TH_HIGH, sc, fill, &ctx, assigned,
                               n_matches);
    if (result != ORBFE_OK) return result;
  }
  ORBFE_CUDA(cudaMemcpyAsync(in_view, d_in, M, cudaMemcpyDeviceToHost, m->stream));
  if (proj_x) ORBFE_CUDA(cudaMemcpyAsync(proj_x, d_px, M * 4, cudaMemcpyDeviceToHost, m->stream));
  if (proj_y) ORBFE_CUDA(cudaMemcpyAsync(proj_y, d_py, M * 4, cudaMemcpyDeviceToHost, m->stream));
  if (proj_xr) ORBFE_CUDA(cudaMemcpyAsync(proj_xr, d_pxr, M * 4, cudaMemcpyDeviceToHost, m->stream));
  if (level) ORBFE_CUDA(cudaMemcpyAsync(level, d_lvl, M * 4, cudaMemcpyDeviceToHost, m->stream));
  if (view_cos) ORBFE_CUDA(cudaMemcpyAsync(view_cos, d_vc, M * 4, cudaMemcpyDeviceToHost, m->stream));
  ORBFE_CUDA(cudaStreamSynchronize(m->stream));
  ORBFE_CUDA(cudaGetLastError());
  return ORBFE_OK;
}

int orbfe_fuse(orbfe_matcher* h, const orbfe_frame_view* kf, const orbfe_projection* pr,
               const orbfe_points3d_view* pts, float th, const float* inv_level_sigma2, int32_t* best_idx,
               int32_t* n_fused) {
  ORBFE_ARG(h && pr);
  int rc = check_frame(kf);
  if (rc != ORBFE_OK) return rc;
  rc = check_points(pts, true, false);
  if (rc != ORBFE_OK) return rc;
  ORBFE_ARG(pts->m == 0 || best_idx);
  ORBFE_ARG(pr->is_sim3 || inv_level_sigma2);
  if (n_fused) *n_fused = 0;
  for (int i = 0; i < pts->m; ++i) best_idx[i] = -1;
  if (pts->m == 0 || kf->n == 0) return ORBFE_OK;
  Matcher* m = reinterpret_cast<Matcher*>(h);
  ORBFE_MATCHER_ENTER(m);
  const size_t M = (size_t)pts->m;
  rc = arena_reserve(m, frame_bytes(kf) + points_bytes(pts) + padded(64) + grid_bytes(kf->n) + padded(M * 4) + padded(64));
  if (rc != ORBFE_OK) return rc;
  DevFrame f;
  put_frame(m, kf, &f);
  DevPoints d;
  put_points(m, pts, &d);
  float sig[ORBFE_MAX_LEVELS] = {0};
  if (inv_level_sigma2)
    for (int i = 0; i < kf->nlevels; ++i) sig[i] = inv_level_sigma2[i];
  const float* dsig = arena_put(m, sig, ORBFE_MAX_LEVELS);
  DevProjection P;
  decompose_pose(pr, &P);
  const size_t up = m->off;
  alloc_grid(m, kf->n, &f);
  int32_t* dbest = arena_alloc<int32_t>(m, M);
  int32_t* dn = arena_alloc<int32_t>(m, 16);
  ORBFE_CUDA(cudaMemcpyAsync(m->d_buf, m->h_buf, up, cudaMemcpyHostToDevice, m->stream));
  ORBFE_CUDA(cudaMemsetAsync(dn, 0, 64, m->stream));
  launch_build_grid(f, m->stream);
  launch_proj_best(pr->is_sim3 ? PM_FUSE_SIM3 : PM_FUSE, f, d, P, th, dsig, ORBFE_TH_LOW, dbest, dn, m->stream);
  m->launches += 2;
  ORBFE_CUDA(cudaMemcpyAsync(best_idx, dbest, M * 4, cudaMemcpyDeviceToHost, m->stream));
  ORBFE_CUDA(cudaMemcpyAsync(m->h_small, dn, 4, cudaMemcpyDeviceToHost, m->stream));
  ORBFE_CUDA(cudaStreamSynchronize(m->stream));
  ORBFE_CUDA(cudaGetLastError());
  if (n_fused) *n_fused = m->h_small[0];
  return ORBFE_OK;
}

int orbfe_search_by_sim3(orbfe_matcher* h, const orbfe_frame_view* kf1, const orbfe_frame_view* kf2,
                         const orbfe_points3d_view* p1, const orbfe_points3d_view* p2, const orbfe_sim3_pair* s,
                         float th, int32_t* match12, int32_t* n_found) {
  ORBFE_ARG(h && s);
  int rc = check_frame(kf1);
  if (rc != ORBFE_OK) return rc;
  rc = check_frame(kf2);
  if (rc != ORBFE_OK) return rc;
  rc = check_points(p1, false, false);
  if (rc != ORBFE_OK) return rc;
  rc = check_points(p2, false, false);
  if (rc != ORBFE_OK) return rc;
  ORBFE_ARG(p1->m == kf1->n && p2->m == kf2->n);  // GetMapPointMatches(): one slot per feature
  ORBFE_ARG(p1->m == 0 || match12);
  if (n_found) *n_found = 0;
  for (int i = 0; i < p1->m; ++i) match12[i] = -1;
  if (p1->m == 0 || p2->m == 0) return ORBFE_OK;
  Matcher* m = reinterpret_cast<Matcher*>(h);
  ORBFE_MATCHER_ENTER(m);
  const size_t n1 = (size_t)p1->m, n2 = (size_t)p2->m;
  rc = arena_reserve(m, frame_bytes(kf1) + frame_bytes(kf2) + points_bytes(p1) + points_bytes(p2) + grid_bytes(kf1->n) +
                            grid_bytes(kf2->n) + 2 * padded(n1 * 4) + padded(n2 * 4) + padded(64));
  if (rc != ORBFE_OK) return rc;
  DevFrame f1, f2;
  put_frame(m, kf1, &f1);
  put_frame(m, kf2, &f2);
  DevPoints d1, d2;
  put_points(m, p1, &d1);
  put_points(m, p2, &d2);
  // sR12 = s12*R12, sR21 = (1.0/s12)*R12.t(), t21 = -sR21*t12 (:1175-1177)
  DevProjection A, B;  // A: KF1 points into KF2; B: KF2 points into KF1
  const float inv_s = (float)(1.0 / (double)s->s12);
  for (int r = 0; r < 3; ++r) {
    for (int c = 0; c < 3; ++c) {
      A.R[r * 3 + c] = s->t1w[r * 4 + c];
      B.R[r * 3 + c] = s->t2w[r * 4 + c];
      B.R2[r * 3 + c] = s->r12[r * 3 + c] * s->s12;
      A.R2[r * 3 + c] = s->r12[c * 3 + r] * inv_s;
    }
    A.t[r] = s->t1w[r * 4 + 3];
    B.t[r] = s->t2w[r * 4 + 3];
    B.t2[r] = s->t12[r];
  }
  for (int r = 0; r < 3; ++r)
    A.t2[r] = -((A.R2[r * 3] * s->t12[0] + A.R2[r * 3 + 1] * s->t12[1]) + A.R2[r * 3 + 2] * s->t12[2]);
  for (int i = 0; i < 3; ++i) A.Ow[i] = B.Ow[i] = 0.f;
  A.fx = B.fx = s->fx; A.fy = B.fy = s->fy; A.cx = B.cx = s->cx; A.cy = B.cy = s->cy;
  A.bf = B.bf = 0.f;
  A.log_sf = s->log_scale_factor2;
  B.log_sf = s->log_scale_factor1;
  const size_t up = m->off;
  alloc_grid(m, kf1->n, &f1);
  alloc_grid(m, kf2->n, &f2);
  int32_t* dm1 = arena_alloc<int32_t>(m, n1);
  int32_t* dm2 = arena_alloc<int32_t>(m, n2);
  int32_t* dout = arena_alloc<int32_t>(m, n1);
  int32_t* dn = arena_alloc<int32_t>(m, 16);
  ORBFE_CUDA(cudaMemcpyAsync(m->d_buf, m->h_buf, up, cudaMemcpyHostToDevice, m->stream));
  ORBFE_CUDA(cudaMemsetAsync(dn, 0, 64, m->stream));
  launch_build_grid(f1, m->stream);
  launch_build_grid(f2, m->stream);
  launch_proj_best(PM_BYSIM3, f2, d1, A, th, nullptr, ORBFE_TH_HIGH, dm1, nullptr, m->stream);
  launch_proj_best(PM_BYSIM3, f1, d2, B, th, nullptr, ORBFE_TH_HIGH, dm2, nullptr, m->stream);
  launch_sim3_agree(dm1, p1->m, dm2, p2->m, dout, dn, m->stream);
  m->launches += 5;
  ORBFE_CUDA(cudaMemcpyAsync(match12, dout, n1 * 4, cudaMemcpyDeviceToHost, m->stream));
  ORBFE_CUDA(cudaMemcpyAsync(m->h_small, dn, 4, cudaMemcpyDeviceToHost, m->stream));
  ORBFE_CUDA(cudaStreamSynchronize(m->stream));
  ORBFE_CUDA(cudaGetLastError());
  if (n_found) *n_found = m->h_small[0];
  return ORBFE_OK;
}

int orbfe_search_by_bow(orbfe_matcher* h, const orbfe_keyframe_view* kf1, const orbfe_keyframe_view* kf2, int kf_kf,
                        float nnratio, int check_ori, int32_t* matches, int32_t* n_matches) {
  ORBFE_ARG(h);
  int rc = check_kf(kf1);
  if (rc != ORBFE_OK) return rc;
  rc = check_kf(kf2);
  if (rc != ORBFE_OK) return rc;
  const int n_out = kf_kf ? kf1->n : kf2->n;
  ORBFE_ARG(n_out == 0 || matches);
  if (n_matches) *n_matches = 0;
  for (int i = 0; i < n_out; ++i) matches[i] = -1;
  if (kf1->n == 0 || kf2->n == 0 || kf1->n_nodes == 0 || kf2->n_nodes == 0) return ORBFE_OK;
  Matcher* m = reinterpret_cast<Matcher*>(h);
  ORBFE_MATCHER_ENTER(m);
  rc = arena_reserve(m, kf_bytes(kf1) + kf_bytes(kf2) + padded((size_t)kf2->n * 4) + padded((size_t)n_out * 4) + padded(64));
  if (rc != ORBFE_OK) return rc;
  DevKeyFrame a, b;
  put_kf(m, kf1, &a);
  put_kf(m, kf2, &b);
  const size_t up = m->off;
  int32_t* claimed = arena_alloc<int32_t>(m, kf2->n);
  int32_t* dmatch = arena_alloc<int32_t>(m, n_out);
  int32_t* dn = arena_alloc<int32_t>(m, 16);
  ORBFE_CUDA(cudaMemcpyAsync(m->d_buf, m->h_buf, up, cudaMemcpyHostToDevice, m->stream));
  ORBFE_CUDA(cudaMemsetAsync(claimed, 0, (size_t)kf2->n * 4, m->stream));
  ORBFE_CUDA(cudaMemsetAsync(dmatch, 0xFF, (size_t)n_out * 4, m->stream));
  launch_bow(a, b, kf_kf, nnratio, check_ori, claimed, dmatch, n_out, dn, m->stream);
  m->launches += 2;
  ORBFE_CUDA(cudaMemcpyAsync(matches, dmatch, (size_t)n_out * 4, cudaMemcpyDeviceToHost, m->stream));
  ORBFE_CUDA(cudaMemcpyAsync(m->h_small, dn, 4, cudaMemcpyDeviceToHost, m->stream));
  ORBFE_CUDA(cudaStreamSynchronize(m->stream));
  ORBFE_CUDA(cudaGetLastError());
  if (n_matches) *n_matches = m->h_small[0];
  return ORBFE_OK;
}

int orbfe_search_for_initialization(orbfe_matcher* h, const orbfe_frame_view* f1, const orbfe_frame_view* f2,
                                    float* prev_matched, int window, float nnratio, int check_ori, int32_t* matches12,
                                    int32_t* n_matches) {
  ORBFE_ARG(h);
  int rc = check_frame(f1);
  if (rc != ORBFE_OK) return rc;
  rc = check_frame(f2);
  if (rc != ORBFE_OK) return rc;
  ORBFE_ARG(f1->n == 0 || (matches12 && prev_matched));
  if (n_matches) *n_matches = 0;
  for (int i = 0; i < f1->n; ++i) matches12[i] = -1;
  if (f1->n == 0 || f2->n == 0) return ORBFE_OK;
  Matcher* m = reinterpret_cast<Matcher*>(h);
  ORBFE_MATCHER_ENTER(m);
  const size_t n1 = (size_t)f1->n, n2 = (size_t)f2->n;
  rc = arena_reserve(m, frame_bytes(f2) + padded(n1 * sizeof(orbfe_keypoint)) + padded(n1 * 32) + padded(n1 * 8) +
                            scratch_bytes(f1->n, f2->n) + padded(n1 * 4) + padded(n2 * 4) + padded(64));
  if (rc != ORBFE_OK) return rc;
  DevFrame d2;
  put_frame(m, f2, &d2);
  const orbfe_keypoint* keys1 = arena_put(m, f1->keys_un, n1);
  const uint8_t* desc1 = arena_put(m, f1->desc, n1 * 32);
  float* prev = const_cast<float*>(arena_put(m, prev_matched, n1 * 2));
  const size_t up = m->off;
  SearchScratch sc;
  alloc_scratch(m, f1->n, f2->n, &d2, &sc);
  int32_t* m12 = arena_alloc<int32_t>(m, n1);
  int32_t* m21 = arena_alloc<int32_t>(m, n2);
  int32_t* dn = arena_alloc<int32_t>(m, 16);
  ORBFE_CUDA(cudaMemcpyAsync(m->d_buf, m->h_buf, up, cudaMemcpyHostToDevice, m->stream));
  launch_build_grid(d2, m->stream);
  launch_init_candidates(d2, keys1, desc1, f1->n, prev, (float)window, sc, 0, m->stream);
  launch_scan_counts(sc.cand_count, f1->n, sc.cand_capacity_needed, m->stream);
  ORBFE_CUDA(cudaMemcpyAsync(m->h_small, sc.cand_capacity_needed, 4, cudaMemcpyDeviceToHost, m->stream));
  ORBFE_CUDA(cudaStreamSynchronize(m->stream));
  rc = ensure_cand(m, (size_t)m->h_small[0] + 16);
  if (rc != ORBFE_OK) return rc;
  sc.cand = m->d_cand;
  sc.cand_cap = (int)std::min<size_t>(m->cand_cap, (size_t)0x7FFFFFFF);
  launch_init_candidates(d2, keys1, desc1, f1->n, prev, (float)window, sc, 1, m->stream);
  if (launch_init_resolve(d2, keys1, f1->n, nnratio, check_ori, sc, m12, m21, prev, dn, m->stream) != 0) {
    set_error("search_for_initialization: cannot reserve %zu bytes of shared memory", n2 * 2);
    return ORBFE_ERR_CUDA;
  }
  m->launches += 5;
  ORBFE_CUDA(cudaMemcpyAsync(matches12, m12, n1 * 4, cudaMemcpyDeviceToHost, m->stream));
  ORBFE_CUDA(cudaMemcpyAsync(prev_matched, prev, n1 * 8, cudaMemcpyDeviceToHost, m->stream));
  ORBFE_CUDA(cudaMemcpyAsync(m->h_small, dn, 4, cudaMemcpyDeviceToHost, m->stream));
  ORBFE_CUDA(cudaStreamSynchronize(m->stream));
  ORBFE_CUDA(cudaGetLastError());
  if (n_matches) *n_matches = m->h_small[0];
  return ORBFE_OK;
}

}  // extern "C"
