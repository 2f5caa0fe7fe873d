// k_search2.cu — the seven other ORBmatcher members (reference: src/ORBmatcher.cc):
//   SearchByBoW x2 (:182-319, :563-696), SearchForInitialization (:446-561), the relocalisation and Sim3
//   SearchByProjection overloads (:1540-1667, :321-434), Fuse x2 (:879-1156) and SearchBySim3 (:1158-1382).
//
// How their sequential parts are made parallel without changing results:
//  * projection overloads with a claim chain (a feature given to one map point is skipped by later ones): the
//    candidate CSR + Jacobi resolve of k_search.cu, with every row's claim blocking (row_obs == nullptr);
//  * Fuse / SearchBySim3: the search of a row reads no state written by other rows -> one thread per point; the map
//    surgery of Fuse (Replace / AddObservation) stays with the caller, in row order;
//  * SearchByBoW: a feature index lives in exactly one vocabulary node (DBoW2 FeatureVector invariant), so the claim
//    chain never crosses nodes -> one WARP per common node walks its rows in order, lanes scan the node's candidates;
//  * SearchForInitialization: a later row may steal a feature with a smaller distance, so rows are resolved in order
//    by one warp over a precomputed candidate CSR (distances are computed in parallel beforehand).
#include "../../include/orb_math_shared.h"
#include "search_common.cuh"

namespace orbfe {

// ------------------------------------------------------------------------------------------------ projection
__device__ __forceinline__ void rigid3(const float* R, const float* t, float x0, float x1, float x2, float* o) {
  o[0] = __fadd_rn(__fadd_rn(__fadd_rn(__fmul_rn(R[0], x0), __fmul_rn(R[1], x1)), __fmul_rn(R[2], x2)), t[0]);
  o[1] = __fadd_rn(__fadd_rn(__fadd_rn(__fmul_rn(R[3], x0), __fmul_rn(R[4], x1)), __fmul_rn(R[5], x2)), t[1]);
  o[2] = __fadd_rn(__fadd_rn(__fadd_rn(__fmul_rn(R[6], x0), __fmul_rn(R[7], x1)), __fmul_rn(R[8], x2)), t[2]);
}

__device__ __forceinline__ float norm3(float a, float b, float c) {  // cv::norm: double accumulation
  const double s = __dadd_rn(__dadd_rn(__dmul_rn((double)a, (double)a), __dmul_rn((double)b, (double)b)),
                             __dmul_rn((double)c, (double)c));
  return (float)__dsqrt_rn(s);
}

struct Projected {
  float u, v, ur, radius;
  int lvl;
};

// Everything the reference does between "MapPoint* pMP = ..." and GetFeaturesInArea, per overload.
__device__ bool project_point(int mode, const DevFrame& f, const DevPoints& p, const DevProjection& P, float th, int i,
                              Projected& q) {
  const float X0 = p.world_pos[3 * i], X1 = p.world_pos[3 * i + 1], X2 = p.world_pos[3 * i + 2];
  float pc[3];
  rigid3(P.R, P.t, X0, X1, X2, pc);
  if (mode == PM_BYSIM3) {  // p3Dc2 = sR21*p3Dc1 + t21 (:1215-1216)
    float pa[3] = {pc[0], pc[1], pc[2]};
    rigid3(P.R2, P.t2, pa[0], pa[1], pa[2], pc);
  }
  float invz;
  if (mode == PM_RELOC) {
    invz = (float)__ddiv_rn(1.0, (double)pc[2]);  // :1570, no sign test
    q.u = __fadd_rn(__fmul_rn(__fmul_rn(P.fx, pc[0]), invz), P.cx);
    q.v = __fadd_rn(__fmul_rn(__fmul_rn(P.fy, pc[1]), invz), P.cy);
    if (q.u < f.min_x || q.u > f.max_x) return false;
    if (q.v < f.min_y || q.v > f.max_y) return false;
  } else {
    if (pc[2] < 0.0f) return false;  // depth must be positive
    invz = (mode == PM_SIM3 || mode == PM_FUSE) ? __fdiv_rn(1.0f, pc[2]) : (float)__ddiv_rn(1.0, (double)pc[2]);
    const float x = __fmul_rn(pc[0], invz), y = __fmul_rn(pc[1], invz);
    q.u = __fadd_rn(__fmul_rn(P.fx, x), P.cx);
    q.v = __fadd_rn(__fmul_rn(P.fy, y), P.cy);
    if (!(q.u >= f.min_x && q.u < f.max_x && q.v >= f.min_y && q.v < f.max_y)) return false;  // KeyFrame::IsInImage
  }
  q.ur = __fsub_rn(q.u, __fmul_rn(P.bf, invz));
  const float max_d = p.max_inv[i], min_d = p.min_inv[i];  // Get{Max,Min}DistanceInvariance()
  float dist;
  if (mode == PM_BYSIM3) {
    dist = norm3(pc[0], pc[1], pc[2]);  // :1235
    if (dist < min_d || dist > max_d) return false;
  } else {
    const float po0 = __fsub_rn(X0, P.Ow[0]), po1 = __fsub_rn(X1, P.Ow[1]), po2 = __fsub_rn(X2, P.Ow[2]);
    dist = norm3(po0, po1, po2);
    if (dist < min_d || dist > max_d) return false;
    if (mode != PM_RELOC) {  // viewing angle below 60 deg (:385, :938, :1098)
      const double dot = __dadd_rn(__dadd_rn(__dmul_rn((double)po0, (double)p.normal[3 * i]),
                                             __dmul_rn((double)po1, (double)p.normal[3 * i + 1])),
                                   __dmul_rn((double)po2, (double)p.normal[3 * i + 2]));
      if (dot < __dmul_rn(0.5, (double)dist)) return false;
    }
  }
  q.lvl = orb_predict_scale(p.max_d[i], dist, P.log_sf, f.nlevels);
  q.radius = __fmul_rn(th, f.scale[q.lvl]);
  return true;
}

// candidates of the two overloads with a claim chain (PM_RELOC, PM_SIM3); CSR as in k_search.cu
__global__ void proj_candidates_kernel(int mode, const DevFrame f, const DevPoints p, const DevProjection P, float th,
                                       SearchScratch sc, int fill) {
  if (fill && sc.cand_capacity_needed[0] > sc.cand_cap) return;  // the candidate buffer is too small: the host repeats the pass
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= p.m) return;
  int count = 0;
  Projected q;
  if (p.valid[i] && project_point(mode, f, p, P, th, i, q)) {
    const uint8_t* dmp = p.desc + (size_t)i * 32;
    uint32_t* out = fill ? sc.cand + sc.cand_count[i] : nullptr;
    const int max_l = mode == PM_RELOC ? q.lvl + 1 : q.lvl;  // :1596 vs :411
    features_in_area(f, q.u, q.v, q.radius, q.lvl - 1, max_l, [&](int idx, int oct) {
      if (f.mp_obs[idx] >= 0) return;  // slot taken before the call (:1609, :406)
      if (fill) out[count] = pack_sc(idx, hamming32(dmp, f.desc + (size_t)idx * 32), oct);
      ++count;
    });
  }
  if (!fill) sc.cand_count[i] = count;
}

// Fuse x2 and one direction of SearchBySim3: best candidate of every point, no state shared between rows
__global__ void proj_best_kernel(int mode, const DevFrame f, const DevPoints p, const DevProjection P, float th,
                                 const float* __restrict__ inv_sigma2, int accept_th, int32_t* __restrict__ best_idx,
                                 int32_t* __restrict__ n_found) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= p.m) return;
  int best = 256, bi = -1;
  Projected q;
  if (p.valid[i] && project_point(mode, f, p, P, th, i, q)) {
    const uint8_t* dmp = p.desc + (size_t)i * 32;
    features_in_area(f, q.u, q.v, q.radius, q.lvl - 1, q.lvl, [&](int idx, int oct) {
      if (mode == PM_FUSE) {  // reprojection error gate (:968-992)
        const orbfe_keypoint kp = f.keys[idx];
        const float ex = __fsub_rn(q.u, kp.x), ey = __fsub_rn(q.v, kp.y);
        const float kr = f.u_right[idx];
        const float e_xy = __fadd_rn(__fmul_rn(ex, ex), __fmul_rn(ey, ey));
        if (kr >= 0.f) {
          const float er = __fsub_rn(q.ur, kr);
          const float e2 = __fadd_rn(e_xy, __fmul_rn(er, er));
          if ((double)__fmul_rn(e2, inv_sigma2[oct]) > 7.8) return;
        } else {
          if ((double)__fmul_rn(e_xy, inv_sigma2[oct]) > 5.99) return;
        }
      }
      const int d = hamming32(dmp, f.desc + (size_t)idx * 32);
      if (d < best) { best = d; bi = idx; }
    });
  }
  const int res = best <= accept_th ? bi : -1;
  best_idx[i] = res;
  if (n_found && res >= 0) atomicAdd(n_found, 1);
}

// SearchBySim3 "check agreement" (:1363-1379)
__global__ void sim3_agree_kernel(const int32_t* __restrict__ m1, int n1, const int32_t* __restrict__ m2, int n2,
                                  int32_t* __restrict__ out12, int32_t* __restrict__ n_found) {
  const int i1 = blockIdx.x * blockDim.x + threadIdx.x;
  if (i1 >= n1) return;
  const int idx2 = m1[i1];
  const bool ok = idx2 >= 0 && idx2 < n2 && m2[idx2] == i1;
  out12[i1] = ok ? idx2 : -1;
  if (ok) atomicAdd(n_found, 1);
}

// ------------------------------------------------------------------------------------------------ warp top-2
// (key = dist << 20 | position, second = second smallest distance of the multiset) merged over the warp; the
// sequential scan's outcome: first position among the minima, second-best counts a duplicated minimum.
__device__ __forceinline__ void warp_top2(unsigned& key, int& second) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const unsigned ok = __shfl_xor_sync(0xffffffffu, key, o);
    const int os = __shfl_xor_sync(0xffffffffu, second, o);
    const int da = (int)(key >> 20), db = (int)(ok >> 20);
    second = min(min(second, os), max(da, db));
    key = min(key, ok);
  }
}

constexpr unsigned KEY_NONE = 0xFFFFFFFFu;  // dist field 4095: "no candidate"

// ------------------------------------------------------------------------------------------------ SearchByBoW
__global__ void __launch_bounds__(128) bow_match_kernel(const DevKeyFrame k1, const DevKeyFrame k2, int kf_kf,
                                                        float nnratio, int32_t* claimed2, int32_t* matches) {
  const int lane = threadIdx.x & 31;
  const int a = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (a >= k1.n_nodes) return;
  const uint32_t id = k1.node_ids[a];
  int l2 = 0, h2 = k2.n_nodes;  // the merge-join of :205-295: same node id in the other feature vector
  while (l2 < h2) { const int mid = (l2 + h2) >> 1; if (k2.node_ids[mid] < id) l2 = mid + 1; else h2 = mid; }
  if (l2 >= k2.n_nodes || k2.node_ids[l2] != id) return;
  const int b2 = k2.node_offsets[l2], e2 = k2.node_offsets[l2 + 1];
  volatile int32_t* claimed = claimed2;
  for (int p1 = k1.node_offsets[a]; p1 < k1.node_offsets[a + 1]; ++p1) {
    const int idx1 = k1.node_features[p1];
    if (!k1.has_mp[idx1]) continue;  // no map point, or a bad one (:224-228, :600-603)
    const uint8_t* d1 = k1.desc + (size_t)idx1 * 32;
    unsigned key = KEY_NONE;
    int second = 256;
    for (int p2 = b2 + lane; p2 < e2; p2 += 32) {
      const int idx2 = k2.node_features[p2];
      if (claimed[idx2]) continue;                    // :240 / :617
      if (kf_kf && !k2.has_mp[idx2]) continue;        // :617-621
      const int dist = hamming32(d1, k2.desc + (size_t)idx2 * 32);
      const unsigned k = ((unsigned)dist << 20) | (unsigned)(p2 - b2);
      const int bestd = key == KEY_NONE ? 256 : (int)(key >> 20);
      if (dist < bestd) { second = bestd; key = k; }
      else if (dist < second) second = dist;
    }
    if (key == KEY_NONE) key = (256u << 20) | 0xFFFFFu;  // distance 256 never passes and orders after real ones
    warp_top2(key, second);
    const int best = (int)(key >> 20);
    const bool pass = kf_kf ? best < TH_LOW : best <= TH_LOW;  // :639 vs :259
    if (pass && (float)best < __fmul_rn(nnratio, (float)second)) {
      const int idx2 = k2.node_features[b2 + (int)(key & 0xFFFFFu)];
      if (lane == 0) {
        claimed[idx2] = 1;
        if (kf_kf) matches[idx1] = idx2; else matches[idx2] = idx1;
      }
    }
    __syncwarp();
  }
}

// rotation histogram of SearchByBoW (:298-316, :675-693) + the return value; one CTA
__global__ void __launch_bounds__(1024) bow_finish_kernel(const DevKeyFrame k1, const DevKeyFrame k2, int kf_kf,
                                                          int check_ori, int32_t* matches, int n_out,
                                                          int32_t* n_matches) {
  __shared__ int s_hist[HISTO];
  __shared__ int s_ind[3];
  __shared__ int s_n;
  const int tid = threadIdx.x, T = blockDim.x;
  for (int i = tid; i < HISTO; i += T) s_hist[i] = 0;
  if (tid == 0) s_n = 0;
  __syncthreads();
  int mine = 0;
  for (int j = tid; j < n_out; j += T) {
    const int m = matches[j];
    if (m < 0) continue;
    ++mine;
    if (check_ori) {
      const int i1 = kf_kf ? j : m, i2 = kf_kf ? m : j;
      atomicAdd(&s_hist[rot_bin(k1.keys[i1].angle, k2.keys[i2].angle)], 1);
    }
  }
  if (mine) atomicAdd(&s_n, mine);
  __syncthreads();
  if (check_ori) {
    if (tid == 0) three_maxima(s_hist, s_ind[0], s_ind[1], s_ind[2]);
    __syncthreads();
    int removed = 0;
    for (int j = tid; j < n_out; j += T) {
      const int m = matches[j];
      if (m < 0) continue;
      const int i1 = kf_kf ? j : m, i2 = kf_kf ? m : j;
      const int bin = rot_bin(k1.keys[i1].angle, k2.keys[i2].angle);
      if (bin != s_ind[0] && bin != s_ind[1] && bin != s_ind[2]) { matches[j] = -1; ++removed; }
    }
    if (removed) atomicSub(&s_n, removed);
    __syncthreads();
  }
  if (tid == 0) *n_matches = s_n;
}

// ------------------------------------------------------------------------------------------------ SearchForInitialization
__global__ void init_candidates_kernel(const DevFrame f2, const orbfe_keypoint* __restrict__ keys1,
                                       const uint8_t* __restrict__ desc1, int n1, const float* __restrict__ prev,
                                       float window, SearchScratch sc, int fill) {
  if (fill && sc.cand_capacity_needed[0] > sc.cand_cap) return;  // the candidate buffer is too small: the host repeats the pass
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n1) return;
  int count = 0;
  const int level1 = keys1[i].octave;
  if (!(level1 > 0)) {  // :462-464
    const uint8_t* d1 = desc1 + (size_t)i * 32;
    uint32_t* out = fill ? sc.cand + sc.cand_count[i] : nullptr;
    features_in_area(f2, prev[2 * i], prev[2 * i + 1], window, level1, level1, [&](int idx, int oct) {
      if (fill) out[count] = pack_sc(idx, hamming32(d1, f2.desc + (size_t)idx * 32), oct);
      ++count;
    });
  }
  if (!fill) sc.cand_count[i] = count;
}

// rows in order (:459-528) by warp 0, then the rotation check (:530-553) and the vbPrevMatched update (:556-558) by
// the whole CTA.  vMatchedDistance lives in shared memory as u16 (0xFFFF = INT_MAX).
__global__ void __launch_bounds__(1024) init_resolve_kernel(const DevFrame f2, const orbfe_keypoint* __restrict__ keys1,
                                                            int n1, float nnratio, int check_ori, SearchScratch sc,
                                                            int32_t* m12, int32_t* m21, float* prev,
                                                            int32_t* n_matches) {
  extern __shared__ uint16_t s_md[];
  __shared__ int s_hist[HISTO];
  __shared__ int s_ind[3];
  __shared__ int s_n;
  const int tid = threadIdx.x, T = blockDim.x, lane = tid & 31;
  for (int i = tid; i < f2.n; i += T) { s_md[i] = 0xFFFF; m21[i] = -1; }
  for (int i = tid; i < n1; i += T) m12[i] = -1;
  for (int i = tid; i < HISTO; i += T) s_hist[i] = 0;
  __syncthreads();
  if (tid < 32) {
    volatile int32_t* v21 = m21;
    volatile int32_t* v12 = m12;
    volatile uint16_t* md = s_md;
    int nm = 0;
    for (int i1 = 0; i1 < n1; ++i1) {
      const int beg = sc.cand_count[i1], end = sc.cand_count[i1 + 1];
      if (beg == end) continue;
      unsigned key = KEY_NONE;
      int second = INT_MAX;
      for (int p = beg + lane; p < end; p += 32) {
        const uint32_t c = sc.cand[p];
        const int idx = c & 0xFFFF, dist = (c >> 16) & 0x1FF;
        if ((int)md[idx] <= dist) continue;  // :485
        const unsigned k = ((unsigned)dist << 20) | (unsigned)(p - beg);
        const int bestd = key == KEY_NONE ? INT_MAX : (int)(key >> 20);
        if (dist < bestd) { second = bestd; key = k; }
        else if (dist < second) second = dist;
      }
      // merge: "no candidate" must act as +inf for `second` as well
      {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
          const unsigned ok = __shfl_xor_sync(0xffffffffu, key, o);
          const int os = __shfl_xor_sync(0xffffffffu, second, o);
          const int da = key == KEY_NONE ? INT_MAX : (int)(key >> 20);
          const int db = ok == KEY_NONE ? INT_MAX : (int)(ok >> 20);
          second = min(min(second, os), max(da, db));
          key = min(key, ok);
        }
      }
      if (key == KEY_NONE) continue;
      const int best = (int)(key >> 20);
      if (best <= TH_LOW && (float)best < __fmul_rn((float)second, nnratio)) {  // :500-502
        const int i2 = sc.cand[beg + (int)(key & 0xFFFFFu)] & 0xFFFF;
        const int prev1 = v21[i2];
        if (prev1 >= 0) --nm;  // :504-508
        ++nm;
        __syncwarp();  // every lane has read the old owner before lane 0 replaces it
        if (lane == 0) {
          if (prev1 >= 0) v12[prev1] = -1;
          v12[i1] = i2;
          v21[i2] = i1;
          md[i2] = (uint16_t)best;
          if (check_ori) s_hist[rot_bin(keys1[i1].angle, f2.keys[i2].angle)] += 1;  // rotHist[bin].push_back(i1)
        }
        __syncwarp();
      }
    }
    if (lane == 0) s_n = nm;
  }
  __syncthreads();
  if (check_ori) {
    if (tid == 0) three_maxima(s_hist, s_ind[0], s_ind[1], s_ind[2]);
    __syncthreads();
    int removed = 0;
    for (int i = tid; i < n1; i += T) {
      const int m = m12[i];
      if (m < 0) continue;  // stolen rows stay in the histogram SIZES but are not matches any more (:545)
      const int bin = rot_bin(keys1[i].angle, f2.keys[m].angle);
      if (bin != s_ind[0] && bin != s_ind[1] && bin != s_ind[2]) { m12[i] = -1; ++removed; }
    }
    if (removed) atomicSub(&s_n, removed);
    __syncthreads();
  }
  for (int i = tid; i < n1; i += T) {
    const int m = m12[i];
    if (m >= 0) { prev[2 * i] = f2.keys[m].x; prev[2 * i + 1] = f2.keys[m].y; }
  }
  if (tid == 0) *n_matches = s_n;
}


// ------------------------------------------------------------------------------------------------ isInFrustum
// Frame::isInFrustum (src/Frame.cc:285-349) over the local map (Tracking::SearchLocalPoints, src/Tracking.cc:1258-1273):
// fills the mbTrackInView / mTrackProj* fields that SearchByProjection(Frame&, vector<MapPoint*>&, th) reads, on the
// device, so the projection search starts without a host pass over the map.
__global__ void frustum_kernel(const DevFrame f, const DevPoints p, const DevProjection P, float cos_limit,
                               uint8_t* __restrict__ in_view, float* __restrict__ proj_x, float* __restrict__ proj_y,
                               float* __restrict__ proj_xr, int32_t* __restrict__ level, float* __restrict__ view_cos) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= p.m) return;
  uint8_t ok = 0;
  float u = 0.f, v = 0.f, ur = 0.f, vc = 0.f;
  int lvl = 0;
  if (p.valid[i]) {
    const float X0 = p.world_pos[3 * i], X1 = p.world_pos[3 * i + 1], X2 = p.world_pos[3 * i + 2];
    float pc[3];
    rigid3(P.R, P.t, X0, X1, X2, pc);
    if (!(pc[2] < 0.0f)) {
      const float invz = __fdiv_rn(1.0f, pc[2]);
      u = __fadd_rn(__fmul_rn(__fmul_rn(P.fx, pc[0]), invz), P.cx);
      v = __fadd_rn(__fmul_rn(__fmul_rn(P.fy, pc[1]), invz), P.cy);
      if (!(u < f.min_x || u > f.max_x) && !(v < f.min_y || v > f.max_y)) {
        const float po0 = __fsub_rn(X0, P.Ow[0]), po1 = __fsub_rn(X1, P.Ow[1]), po2 = __fsub_rn(X2, P.Ow[2]);
        const float dist = norm3(po0, po1, po2);
        if (!(dist < p.min_inv[i] || dist > p.max_inv[i])) {
          const double dot = __dadd_rn(__dadd_rn(__dmul_rn((double)po0, (double)p.normal[3 * i]),
                                                 __dmul_rn((double)po1, (double)p.normal[3 * i + 1])),
                                       __dmul_rn((double)po2, (double)p.normal[3 * i + 2]));
          vc = (float)__ddiv_rn(dot, (double)dist);  // const float viewCos = PO.dot(Pn)/dist
          if (!(vc < cos_limit)) {
            lvl = orb_predict_scale(p.max_d[i], dist, P.log_sf, f.nlevels);
            ur = __fsub_rn(u, __fmul_rn(P.bf, invz));
            ok = 1;
          }
        }
      }
    }
  }
  in_view[i] = ok;  // the reference leaves the fields of rejected points stale; here they read 0
  proj_x[i] = ok ? u : 0.f; proj_y[i] = ok ? v : 0.f; proj_xr[i] = ur; level[i] = lvl; view_cos[i] = ok ? vc : 0.f;
}

void launch_frustum(const DevFrame& f, const DevPoints& p, const DevProjection& P, float cos_limit, uint8_t* in_view,
                    float* proj_x, float* proj_y, float* proj_xr, int32_t* level, float* view_cos, cudaStream_t s) {
  if (p.m > 0)
    frustum_kernel<<<(p.m + 127) / 128, 128, 0, s>>>(f, p, P, cos_limit, in_view, proj_x, proj_y, proj_xr, level, view_cos);
}

// ------------------------------------------------------------------------------------------------ launchers
void launch_proj_candidates(int mode, const DevFrame& f, const DevPoints& p, const DevProjection& P, float th,
                            const SearchScratch& sc, int fill, cudaStream_t s) {
  if (p.m > 0) proj_candidates_kernel<<<(p.m + 127) / 128, 128, 0, s>>>(mode, f, p, P, th, sc, fill);
}

void launch_proj_best(int mode, const DevFrame& f, const DevPoints& p, const DevProjection& P, float th,
                      const float* inv_sigma2, int accept_th, int32_t* best_idx, int32_t* n_found, cudaStream_t s) {
  if (p.m > 0)
    proj_best_kernel<<<(p.m + 127) / 128, 128, 0, s>>>(mode, f, p, P, th, inv_sigma2, accept_th, best_idx, n_found);
}

void launch_sim3_agree(const int32_t* m1, int n1, const int32_t* m2, int n2, int32_t* out12, int32_t* n_found,
                       cudaStream_t s) {
  if (n1 > 0) sim3_agree_kernel<<<(n1 + 255) / 256, 256, 0, s>>>(m1, n1, m2, n2, out12, n_found);
}

void launch_bow(const DevKeyFrame& k1, const DevKeyFrame& k2, int kf_kf, float nnratio, int check_ori,
                int32_t* claimed2, int32_t* matches, int n_out, int32_t* n_matches, cudaStream_t s) {
  if (k1.n_nodes > 0 && k2.n_nodes > 0)
    bow_match_kernel<<<(k1.n_nodes + 3) / 4, 128, 0, s>>>(k1, k2, kf_kf, nnratio, claimed2, matches);
  bow_finish_kernel<<<1, 1024, 0, s>>>(k1, k2, kf_kf, check_ori, matches, n_out, n_matches);
}

void launch_init_candidates(const DevFrame& f2, const orbfe_keypoint* keys1, const uint8_t* desc1, int n1,
                            const float* prev, float window, const SearchScratch& sc, int fill, cudaStream_t s) {
  if (n1 > 0) init_candidates_kernel<<<(n1 + 127) / 128, 128, 0, s>>>(f2, keys1, desc1, n1, prev, window, sc, fill);
}

int launch_init_resolve(const DevFrame& f2, const orbfe_keypoint* keys1, int n1, float nnratio, int check_ori,
                        const SearchScratch& sc, int32_t* m12, int32_t* m21, float* prev, int32_t* n_matches,
                        cudaStream_t s) {
  const size_t smem = ((size_t)f2.n * 2 + 15) & ~(size_t)15;
  if (smem > 40 * 1024 &&  // per device, cheap: set whenever the default 48 KB would not do
      cudaFuncSetAttribute(init_resolve_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 160 * 1024) != cudaSuccess)
    return -1;
  init_resolve_kernel<<<1, 1024, smem, s>>>(f2, keys1, n1, nnratio, check_ori, sc, m12, m21, prev, n_matches);
  return 0;
}

}  // namespace orbfe
