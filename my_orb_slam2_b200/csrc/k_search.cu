// k_search.cu — ORBmatcher::SearchByProjection (two overloads) and
// SearchForTriangulation (reference: src/ORBmatcher.cc:46-132, :1392-1538, :702-872)
// plus the Frame grid they query (src/Frame.cc:243-258, :351-417).
//
// SearchByProjection has a sequential side effect: iteration i skips features already
// given to a map point with observations by iterations j < i of the same call
// (:90-92, :126).  It is parallelised exactly:
//   1. candidates kernel: every row (map point / last-frame point) lists the features
//      that pass all STATIC gates (grid window, level window, |dx|,|dy| < r, stereo
//      gate, feature not already taken before the call) in the reference's traversal
//      order (cells ix-outer / iy-inner, index order inside a cell) with their Hamming
//      distance -> CSR (count pass, scan, fill pass);
//   2. resolve kernel (one CTA): Jacobi iteration on B[f] = first row that claims
//      feature f with Observations() > 0.  Row i ignores candidate f iff B[f] < i.  Row i
//      is exact after iteration i at the latest (rows below it are exact by induction and
//      rows above it cannot lower "exists claimer < i"), and in practice after 2-4
//      iterations; the loop stops when B does not change.  First-wins ties are kept by
//      walking the candidate list in traversal order with strict '<'.
#include "search_common.cuh"

namespace orbfe {

__global__ void __launch_bounds__(1024) build_grid_kernel(const DevFrame f) {
  __shared__ int32_t cnt[GRID_CELLS + 1];
  __shared__ int wtmp[32];
  const int tid = threadIdx.x;
  for (int c = tid; c <= GRID_CELLS; c += blockDim.x) cnt[c] = 0;
  __syncthreads();
  for (int i = tid; i < f.n; i += blockDim.x) {
    const int c = pos_in_grid(f, f.keys[i]);
    if (c >= 0) atomicAdd(&cnt[c], 1);
  }
  __syncthreads();
  const int total = cta_excl_scan(cnt, GRID_CELLS, wtmp);
  if (tid == 0) cnt[GRID_CELLS] = total;
  __syncthreads();
  for (int c = tid; c <= GRID_CELLS; c += blockDim.x) f.cell_start[c] = cnt[c];
  __syncthreads();
  for (int i = tid; i < f.n; i += blockDim.x) {
    const int c = pos_in_grid(f, f.keys[i]);
    if (c >= 0) f.cell_items[atomicAdd(&cnt[c], 1)] = i;
  }
  __syncthreads();
  // cells hold their features in insertion (index) order in the reference: sort each short list
  for (int c = tid; c < GRID_CELLS; c += blockDim.x) {
    const int b = f.cell_start[c], e = f.cell_start[c + 1];
    for (int i = b + 1; i < e; ++i) {
      const int v = f.cell_items[i];
      int j = i - 1;
      while (j >= b && f.cell_items[j] > v) { f.cell_items[j + 1] = f.cell_items[j]; --j; }
      f.cell_items[j + 1] = v;
    }
  }
}

// ---- SearchByProjection(Frame&, const vector<MapPoint*>&, th): candidates (src/ORBmatcher.cc:46-118) ----
__global__ void sbp_map_candidates_kernel(const DevFrame f, const DevMapPoints mp, float th, SearchScratch sc,
                                          int fill) {
  if (fill && sc.cand_capacity_needed[0] > sc.cand_cap) return;  // the candidate buffer is too small: the host repeats the pass
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= mp.m) return;
  int count = 0;
  if (mp.in_view[i]) {  // :53-57
    const int lvl = mp.level[i];
    float r = (double)mp.view_cos[i] > 0.998 ? 2.5f : 4.0f;  // RadiusByViewingCos (:134-140)
    if ((double)th != 1.0) r = __fmul_rn(r, th);              // bFactor (:48,:68-69)
    const float rs = __fmul_rn(r, f.scale[lvl]);
    const float pxr = mp.proj_xr[i];
    const uint8_t* dmp = mp.desc + (size_t)i * 32;
    uint32_t* out = fill ? sc.cand + sc.cand_count[i] : nullptr;
    features_in_area(f, mp.proj_x[i], mp.proj_y[i], rs, lvl - 1, lvl, [&](int idx, int oct) {
      if (f.mp_obs[idx] > 0) return;  // taken before the call (:90-92)
      const float ur = f.u_right[idx];
      if (ur > 0.f) {
        const float er = fabsf(__fsub_rn(pxr, ur));
        if (er > rs) return;  // :94-99
      }
      if (fill) out[count] = pack_sc(idx, hamming32(dmp, f.desc + (size_t)idx * 32), oct);
      ++count;
    });
  }
  if (!fill) sc.cand_count[i] = count;
}

// ---- SearchByProjection(Frame& Cur, const Frame& Last, th, bMono): candidates (:1392-1492) ----
__global__ void sbp_last_candidates_kernel(const DevFrame f, const DevLastFrame lf, const orbfe_camera_pose cam,
                                           const float tlc_z, float th, int mono, SearchScratch sc, int fill) {
  if (fill && sc.cand_capacity_needed[0] > sc.cand_cap) return;  // the candidate buffer is too small: the host repeats the pass
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= lf.n) return;
  int count = 0;
  const float* T = cam.tcw;
  const bool forward = tlc_z > cam.mb && !mono;    // :1412-1413
  const bool backward = -tlc_z > cam.mb && !mono;
  if (lf.valid[i]) {
    const float X0 = lf.world_pos[3 * i], X1 = lf.world_pos[3 * i + 1], X2 = lf.world_pos[3 * i + 2];
    const float xc = __fadd_rn(__fadd_rn(__fadd_rn(__fmul_rn(T[0], X0), __fmul_rn(T[1], X1)), __fmul_rn(T[2], X2)), T[3]);
    const float yc = __fadd_rn(__fadd_rn(__fadd_rn(__fmul_rn(T[4], X0), __fmul_rn(T[5], X1)), __fmul_rn(T[6], X2)), T[7]);
    const float zc = __fadd_rn(__fadd_rn(__fadd_rn(__fmul_rn(T[8], X0), __fmul_rn(T[9], X1)), __fmul_rn(T[10], X2)), T[11]);
    const float invzc = (float)__ddiv_rn(1.0, (double)zc);  // :1430
    bool ok = !(invzc < 0.f);
    const float u = __fadd_rn(__fmul_rn(__fmul_rn(cam.fx, xc), invzc), cam.cx);  // :1435-1436
    const float v = __fadd_rn(__fmul_rn(__fmul_rn(cam.fy, yc), invzc), cam.cy);
    if (u < f.min_x || u > f.max_x) ok = false;
    if (v < f.min_y || v > f.max_y) ok = false;
    if (ok) {
      const int oct = lf.octave[i];
      const float radius = __fmul_rn(th, f.scale[oct]);  // :1445
      int min_l, max_l;
      if (forward) { min_l = oct; max_l = -1; }          // :1453-1458
      else if (backward) { min_l = 0; max_l = oct; }
      else { min_l = oct - 1; max_l = oct + 1; }
      const uint8_t* dmp = lf.desc + (size_t)i * 32;
      uint32_t* out = fill ? sc.cand + sc.cand_count[i] : nullptr;
      const float ur_pred = __fsub_rn(u, __fmul_rn(cam.mbf, invzc));  // :1477
      features_in_area(f, u, v, radius, min_l, max_l, [&](int idx, int o2) {
        if (f.mp_obs[idx] > 0) return;  // :1471-1473
        const float ur = f.u_right[idx];
        if (ur > 0.f) {
          const float er = fabsf(__fsub_rn(ur_pred, ur));
          if (er > radius) return;  // :1475-1481
        }
        if (fill) out[count] = pack_sc(idx, hamming32(dmp, f.desc + (size_t)idx * 32), o2);
        ++count;
      });
    }
  }
  if (!fill) sc.cand_count[i] = count;
}

// tlc.z = (LastFrame.Rcw * twc + LastFrame.tcw).z with twc = -Rcw^T tcw of the current frame (:1403-1410)
__host__ __device__ inline float tlc_z_of(const float* T, const float* Tl) {
  float twc[3];
  for (int i = 0; i < 3; ++i) {
    double s = 0;  // cv::gemm with the transpose flag accumulates in double
    for (int k = 0; k < 3; ++k) s += (double)T[k * 4 + i] * (double)T[k * 4 + 3];
    twc[i] = (float)(-1.0 * s);
  }
  return ((Tl[8] * twc[0] + Tl[9] * twc[1]) + Tl[10] * twc[2]) + Tl[11];
}

__global__ void __launch_bounds__(1024) scan_counts_kernel(int32_t* counts, int rows, int32_t* total) {
  __shared__ int wtmp[32];
  __shared__ int carry;
  if (threadIdx.x == 0) carry = 0;
  __syncthreads();
  const int chunk = 1024 * 8;
  for (int base = 0; base < rows; base += chunk) {
    const int n = min(chunk, rows - base);
    const int t = cta_excl_scan(counts + base, n, wtmp);
    const int c = carry;
    __syncthreads();
    for (int i = threadIdx.x; i < n; i += blockDim.x) counts[base + i] += c;
    if (threadIdx.x == 0) carry = c + t;
    __syncthreads();
  }
  if (threadIdx.x == 0) { counts[rows] = carry; *total = carry; }
}

// ---- ordered claim resolution + rotation histogram; one CTA ----
__global__ void __launch_bounds__(1024) resolve_kernel(const DevFrame f, int rows, const int32_t* __restrict__ row_obs,
                                                       const float* __restrict__ row_angle, int mode, float nnratio,
                                                       int check_ori, int accept_th, SearchScratch sc) {
  __shared__ int s_changed, s_nmatch, s_ind[3];
  __shared__ int s_hist[HISTO];
  const int tid = threadIdx.x, T = blockDim.x;
  if (sc.cand_capacity_needed[0] > sc.cand_cap) return;  // nothing was filled: the host grows the buffer and repeats
  int32_t* b_old = sc.block_a;
  int32_t* b_new = sc.block_b;
  for (int i = tid; i < f.n; i += T) b_old[i] = INT_MAX;
  int iters = 0;
  while (true) {
    for (int i = tid; i < f.n; i += T) b_new[i] = INT_MAX;
    if (tid == 0) s_changed = 0;
    __syncthreads();
    for (int i = tid; i < rows; i += T) {
      const int beg = sc.cand_count[i], end = sc.cand_count[i + 1];
      int best = 256, best2 = 256, lvl = -1, lvl2 = -1, best_idx = -1;
      for (int p = beg; p < end; ++p) {
        const uint32_t c = sc.cand[p];
        const int idx = c & 0xFFFF, dist = (c >> 16) & 0x1FF, oct = c >> 25;
        if (b_old[idx] < i) continue;  // claimed by an earlier row whose map point has observations
        if (dist < best) { best2 = best; best = dist; lvl2 = lvl; lvl = oct; best_idx = idx; }
        else if (dist < best2) { lvl2 = oct; best2 = dist; }
      }
      int res = -1;
      if (best <= accept_th) {
        if (mode == 0) {
          if (!(lvl == lvl2 && (float)best > __fmul_rn(nnratio, (float)best2))) res = best_idx;  // :121-124
        } else {
          res = best_idx;  // :1494
        }
      }
      sc.result[i] = res;
      if (res >= 0 && (row_obs == nullptr || row_obs[i] > 0)) atomicMin(&b_new[res], i);
    }
    __syncthreads();
    int diff = 0;
    for (int i = tid; i < f.n; i += T) diff |= (b_new[i] != b_old[i]);
    if (diff) s_changed = 1;
    __syncthreads();
    ++iters;
    const int changed = s_changed;
    __syncthreads();
    if (!changed) break;
    int32_t* t = b_old; b_old = b_new; b_new = t;
  }
  // F.mvpMapPoints[bestIdx] = pMP: the last writer (largest row) stays (:126, :1496)
  for (int i = tid; i < f.n; i += T) sc.assigned[i] = -1;
  for (int i = tid; i < HISTO; i += T) s_hist[i] = 0;
  if (tid == 0) s_nmatch = 0;
  __syncthreads();
  int mine = 0;
  for (int i = tid; i < rows; i += T) {
    const int res = sc.result[i];
    if (res >= 0) {
      ++mine;
      atomicMax(&sc.assigned[res], i);
      if (check_ori && mode == 1) atomicAdd(&s_hist[rot_bin(row_angle[i], f.keys[res].angle)], 1);
    }
  }
  if (mine) atomicAdd(&s_nmatch, mine);
  __syncthreads();
  if (check_ori && mode == 1) {
    if (tid == 0) three_maxima(s_hist, s_ind[0], s_ind[1], s_ind[2]);
    __syncthreads();
    int removed = 0;
    for (int i = tid; i < rows; i += T) {
      const int res = sc.result[i];
      if (res >= 0) {
        const int bin = rot_bin(row_angle[i], f.keys[res].angle);
        if (bin != s_ind[0] && bin != s_ind[1] && bin != s_ind[2]) { sc.assigned[res] = -2; ++removed; }  // :1516-1535: set to NULL
      }
    }
    if (removed) atomicSub(&s_nmatch, removed);
    __syncthreads();
  }
  if (tid == 0) { sc.counters[0] = s_nmatch; sc.counters[2] = iters; }
}

// ---- SearchForTriangulation (:702-872) + CheckDistEpipolarLine (:147-167): one warp per kf1 CSR entry ----
__global__ void __launch_bounds__(256) triangulation_kernel(const DevKeyFrame k1, const DevKeyFrame k2,
                                                            const float* __restrict__ F12, float ex, float ey,
                                                            int only_stereo, int32_t* __restrict__ matches12) {
  const int lane = threadIdx.x & 31;
  const int p1 = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  const int total = k1.node_offsets[k1.n_nodes];
  if (p1 >= total) return;
  // node of this entry (upper bound on offsets), then the same node id in kf2 (the merge-join of :736-838)
  int lo = 0, hi = k1.n_nodes;
  while (hi - lo > 1) { const int mid = (lo + hi) >> 1; if (k1.node_offsets[mid] <= p1) lo = mid; else hi = mid; }
  const uint32_t id = k1.node_ids[lo];
  int l2 = 0, h2 = k2.n_nodes;
  while (l2 < h2) { const int mid = (l2 + h2) >> 1; if (k2.node_ids[mid] < id) l2 = mid + 1; else h2 = mid; }
  if (l2 >= k2.n_nodes || k2.node_ids[l2] != id) return;
  const int idx1 = k1.node_features[p1];
  if (k1.has_mp[idx1]) return;  // :752-755
  const bool stereo1 = k1.u_right[idx1] >= 0.f;
  if (only_stereo && !stereo1) return;
  const orbfe_keypoint kp1 = k1.keys[idx1];
  const uint8_t* d1 = k1.desc + (size_t)idx1 * 32;
  // epipolar line l = x1' F12 (:149-152)
  const float la = __fadd_rn(__fadd_rn(__fmul_rn(kp1.x, F12[0]), __fmul_rn(kp1.y, F12[3])), F12[6]);
  const float lb = __fadd_rn(__fadd_rn(__fmul_rn(kp1.x, F12[1]), __fmul_rn(kp1.y, F12[4])), F12[7]);
  const float lc = __fadd_rn(__fadd_rn(__fmul_rn(kp1.x, F12[2]), __fmul_rn(kp1.y, F12[5])), F12[8]);
  const float den = __fadd_rn(__fmul_rn(la, la), __fmul_rn(lb, lb));
  const int b2 = k2.node_offsets[l2], e2 = k2.node_offsets[l2 + 1];
  // final choice of the sequential scan (:769-811) = min distance among gate-passing candidates, LATER wins ties
  unsigned best = 0xFFFFFFFFu;
  for (int p2 = b2 + lane; p2 < e2; p2 += 32) {
    const int idx2 = k2.node_features[p2];
    if (k2.has_mp[idx2]) continue;  // vbMatched2 is never set in the reference (:722,:773)
    const bool stereo2 = k2.u_right[idx2] >= 0.f;
    if (only_stereo && !stereo2) continue;
    const int dist = hamming32(d1, k2.desc + (size_t)idx2 * 32);
    if (dist > TH_LOW) continue;
    const orbfe_keypoint kp2 = k2.keys[idx2];
    if (!stereo1 && !stereo2) {
      const float dex = __fsub_rn(ex, kp2.x), dey = __fsub_rn(ey, kp2.y);
      if (__fadd_rn(__fmul_rn(dex, dex), __fmul_rn(dey, dey)) < __fmul_rn(100.f, k2.scale[kp2.octave])) continue;
    }
    const float num = __fadd_rn(__fadd_rn(__fmul_rn(la, kp2.x), __fmul_rn(lb, kp2.y)), lc);
    if (den == 0.f) continue;
    const float dsqr = __fdiv_rn(__fmul_rn(num, num), den);
    if ((double)dsqr < 3.84 * (double)k2.sigma2[kp2.octave])
      best = min(best, ((unsigned)dist << 20) | (unsigned)(0xFFFFF - (p2 - b2)));
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) best = min(best, __shfl_xor_sync(0xffffffffu, best, o));
  if (lane == 0 && best != 0xFFFFFFFFu) matches12[idx1] = k2.node_features[b2 + (0xFFFFF - (int)(best & 0xFFFFFu))];
}

// rotation histogram (:813-823, :840-859) + the idx1-sorted pair list (:861-869); one CTA
__global__ void __launch_bounds__(1024) triangulation_finish_kernel(const DevKeyFrame k1, const DevKeyFrame k2,
                                                                    int check_ori, int32_t* matches12,
                                                                    int32_t* flags /*[k1.n+1] scratch*/, int32_t* pairs,
                                                                    int cap, int32_t* n_pairs) {
  __shared__ int s_hist[HISTO];
  __shared__ int s_ind[3];
  __shared__ int wtmp[32];
  const int tid = threadIdx.x, T = blockDim.x;
  if (check_ori) {
    for (int i = tid; i < HISTO; i += T) s_hist[i] = 0;
    __syncthreads();
    for (int i = tid; i < k1.n; i += T) {
      const int m = matches12[i];
      if (m >= 0) atomicAdd(&s_hist[rot_bin(k1.keys[i].angle, k2.keys[m].angle)], 1);
    }
    __syncthreads();
    if (tid == 0) three_maxima(s_hist, s_ind[0], s_ind[1], s_ind[2]);
    __syncthreads();
    for (int i = tid; i < k1.n; i += T) {
      const int m = matches12[i];
      if (m >= 0) {
        const int bin = rot_bin(k1.keys[i].angle, k2.keys[m].angle);
        if (bin != s_ind[0] && bin != s_ind[1] && bin != s_ind[2]) matches12[i] = -1;
      }
    }
    __syncthreads();
  }
  for (int i = tid; i < k1.n; i += T) flags[i] = matches12[i] >= 0 ? 1 : 0;
  __syncthreads();
  int carry = 0;
  const int chunk = 8192;
  for (int base = 0; base < k1.n; base += chunk) {
    const int n = min(chunk, k1.n - base);
    const int t = cta_excl_scan(flags + base, n, wtmp);
    for (int i = tid; i < n; i += T) {
      const int m = matches12[base + i];
      const int slot = carry + flags[base + i];
      if (m >= 0 && slot < cap) { pairs[2 * slot] = base + i; pairs[2 * slot + 1] = m; }
    }
    carry += t;
    __syncthreads();
  }
  if (tid == 0) *n_pairs = carry;
}

__global__ void fill_i32_kernel(int32_t* p, int n, int32_t v) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) p[i] = v;
}

// ---- launchers ----
void launch_build_grid(const DevFrame& f, cudaStream_t s) { build_grid_kernel<<<1, 1024, 0, s>>>(f); }

void launch_sbp_map_candidates(const DevFrame& f, const DevMapPoints& mp, float th, const SearchScratch& sc, int fill,
                               cudaStream_t s) {
  if (mp.m > 0) sbp_map_candidates_kernel<<<(mp.m + 127) / 128, 128, 0, s>>>(f, mp, th, sc, fill);
}

void launch_sbp_last_candidates(const DevFrame& f, const DevLastFrame& lf, const orbfe_camera_pose& cam, float th,
                                int mono, const SearchScratch& sc, int fill, cudaStream_t s) {
  if (lf.n > 0) sbp_last_candidates_kernel<<<(lf.n + 127) / 128, 128, 0, s>>>(f, lf, cam, lf.tlc_z, th, mono, sc, fill);
}

float host_tlc_z(const float* tcw_cur, const float* tcw_last) { return tlc_z_of(tcw_cur, tcw_last); }

void launch_scan_counts(int32_t* counts, int rows, int32_t* total, cudaStream_t s) {
  scan_counts_kernel<<<1, 1024, 0, s>>>(counts, rows, total);
}

void launch_resolve(const DevFrame& f, int rows, const int32_t* row_obs, const float* row_angle, int mode,
                    float nnratio, int check_ori, int accept_th, const SearchScratch& sc, cudaStream_t s) {
  resolve_kernel<<<1, 1024, 0, s>>>(f, rows, row_obs, row_angle, mode, nnratio, check_ori, accept_th, sc);
}

void launch_triangulation(const DevKeyFrame& k1, const DevKeyFrame& k2, const float* F12, float ex, float ey,
                          int only_stereo, int check_ori, int32_t* matches12, int32_t* flags, int32_t* pairs, int cap,
                          int32_t* n_pairs, int total_entries, cudaStream_t s) {
  if (k1.n > 0) fill_i32_kernel<<<(k1.n + 255) / 256, 256, 0, s>>>(matches12, k1.n, -1);
  if (total_entries > 0)
    triangulation_kernel<<<(total_entries + 7) / 8, 256, 0, s>>>(k1, k2, F12, ex, ey, only_stereo, matches12);
  triangulation_finish_kernel<<<1, 1024, 0, s>>>(k1, k2, check_ori, matches12, flags, pairs, cap, n_pairs);
}

}  // namespace orbfe
