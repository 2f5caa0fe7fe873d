// match_kernels.h — launchers of the ORBmatcher kernels (k_match.cu, k_search.cu).
#pragma once
#include <cuda_runtime.h>

#include <cstdint>

#include "../../include/orbfe_types.h"

namespace orbfe {

// ---- k_match.cu ----
void launch_descriptor_distance(const uint8_t* a, const uint8_t* b, int n, int32_t* out, cudaStream_t s);
int bruteforce_slices(int nq, int nm, int* slice_len, int max_slices = 0);
double measure_popc_peak(cudaStream_t s);  // 32-bit POPC results per second of the current device (register-only probe)
void launch_hamming_bruteforce(const uint8_t* query, int nq, const uint8_t* map, int nm, int slice_len, int slices,
                               long long index_base, unsigned long long* part_best, unsigned long long* part_second,
                               cudaStream_t s);
void launch_hamming_merge(const unsigned long long* best_all, const unsigned long long* second_all, int parts, int nq,
                          unsigned long long* best, unsigned long long* second, cudaStream_t s);

// ---- k_match_mma.cu: the same brute force on the tensor cores (tcgen05 kind::i8; optional second path) ----
int hamming_mma_slices(int nq, int nm, int sms, int* tiles_per_slice);
size_t hamming_mma_scratch_bytes(int nq, int nm, int slices);
int launch_hamming_mma(const uint8_t* query, int nq, const uint8_t* map, int nm, long long index_base, uint8_t* scratch,
                       int slices, int tiles_per_slice, int sms, unsigned long long* best, unsigned long long* second,
                       unsigned** err_out, cudaStream_t s);

void launch_distinctive(const uint8_t* desc, const int32_t* offsets, int n_points, int32_t* best_idx,
                        uint8_t* best_desc, cudaStream_t s);

// ---- k_search.cu ----
// Device mirror of orbfe_frame_view + the 64x48 grid of Frame::AssignFeaturesToGrid (src/Frame.cc:243-258) in CSR form.
struct DevFrame {
  int n;
  const orbfe_keypoint* keys;
  const float* u_right;
  const uint8_t* desc;
  const int32_t* mp_obs;
  float min_x, min_y, max_x, max_y, inv_w, inv_h;
  int nlevels;
  const float* scale;        // device copy of mvScaleFactors
  int32_t* cell_start;       // [64*48 + 1]
  int32_t* cell_items;       // [n]
};

struct DevMapPoints {
  int m;
  const float *proj_x, *proj_y, *proj_xr, *view_cos;
  const int32_t *level, *obs;
  const uint8_t* in_view;
  const uint8_t* desc;
};

struct DevLastFrame {
  int n;
  const uint8_t* valid;
  const float* world_pos;
  const int32_t *obs, *octave;
  const float* angle;
  const uint8_t* desc;
  float tlc_z;  // z of LastFrame.Rcw*twc + tcw (src/ORBmatcher.cc:1410), computed by the launcher kernel
};

struct DevKeyFrame {
  int n;
  const orbfe_keypoint* keys;
  const float* u_right;
  const uint8_t* desc;
  const uint8_t* has_mp;
  int n_nodes;
  const uint32_t* node_ids;
  const int32_t *node_offsets, *node_features;
  const float *scale, *sigma2;
};

// scratch of one search call (all device pointers, sized by the caller)
struct SearchScratch {
  int32_t* cand_count;   // [rows + 1] -> exclusive offsets after the scan
  uint32_t* cand;        // [total] idx | dist<<16 | octave<<25
  int32_t* cand_capacity_needed;  // [1] total candidates (device)
  int cand_cap;          // entries `cand` can hold: the fill pass and the resolve step leave at once when the total
                         // exceeds it (the host then grows the buffer and repeats them: no mid-call sync otherwise)
  int32_t* result;       // [rows] chosen feature per row (-1 none)
  int32_t* block_a;      // [n features] first blocking claimer (iteration t)
  int32_t* block_b;      // [n features] (iteration t+1)
  int32_t* assigned;     // [n features] output
  int32_t* counters;     // [8] misc: 0 nmatches, 1 changed flag, 2 iterations
  float* proj;           // [rows*4] u, v, invz, radius (last-frame overload)
  int32_t* hist;         // [32]
};

void launch_build_grid(const DevFrame& f, cudaStream_t s);
// SearchByProjection(Frame&, vector<MapPoint*>&, th): pass 1 counts, pass 2 fills (after a scan of cand_count)
void launch_sbp_map_candidates(const DevFrame& f, const DevMapPoints& mp, float th, const SearchScratch& sc, int fill,
                               cudaStream_t s);
void launch_sbp_last_candidates(const DevFrame& f, const DevLastFrame& lf, const orbfe_camera_pose& cam, float th,
                                int mono, const SearchScratch& sc, int fill, cudaStream_t s);
void launch_scan_counts(int32_t* counts, int rows, int32_t* total, cudaStream_t s);
// ordered claim resolution + rotation histogram; mode 0 = map overload (ratio test), 1 = best only
// (last frame / relocalisation / Sim3 overloads).  row_obs == nullptr: every row's claim blocks later rows.
// accept_th: bestDist <= accept_th (TH_HIGH, TH_LOW or the caller's ORBdist).
void launch_resolve(const DevFrame& f, int rows, const int32_t* row_obs, const float* row_angle, int mode,
                    float nnratio, int check_ori, int accept_th, const SearchScratch& sc, cudaStream_t s);
// SearchForTriangulation
float host_tlc_z(const float* tcw_cur, const float* tcw_last);
void launch_triangulation(const DevKeyFrame& k1, const DevKeyFrame& k2, const float* F12 /*device, 9*/, float ex,
                          float ey, int only_stereo, int check_ori, int32_t* matches12 /*[k1.n]*/,
                          int32_t* flags /*[k1.n+1]*/, int32_t* pairs, int cap, int32_t* n_pairs, int total_entries,
                          cudaStream_t s);

// ---- k_search2.cu: the seven other ORBmatcher members ----
struct DevPoints {  // device mirror of orbfe_points3d_view
  int m;
  const uint8_t* valid;
  const float *world_pos, *normal, *max_inv, *min_inv, *max_d, *angle;
  const uint8_t* desc;
};

// Rcw / tcw / Ow of the target (already decomposed when the caller passed a Sim3) and, for SearchBySim3, the second
// transform applied after the first (p3Dc2 = sR21*(R1w*X + t1w) + t21, src/ORBmatcher.cc:1215-1216).
struct DevProjection {
  float R[9], t[3], Ow[3];
  float R2[9], t2[3];
  float fx, fy, cx, cy, bf, log_sf;
};

enum ProjMode {
  PM_RELOC = 0,      // SearchByProjection(Frame&, KeyFrame*, sAlreadyFound, th, ORBdist)   :1540-1667
  PM_SIM3 = 1,       // SearchByProjection(KeyFrame*, Scw, vpPoints, vpMatched, th)          :321-434
  PM_FUSE = 2,       // Fuse(KeyFrame*, vpMapPoints, th)                                     :879-1029
  PM_FUSE_SIM3 = 3,  // Fuse(KeyFrame*, Scw, vpPoints, th, vpReplacePoint)                   :1033-1156
  PM_BYSIM3 = 4      // one direction of SearchBySim3                                        :1204-1361
};

void launch_proj_candidates(int mode, const DevFrame& f, const DevPoints& p, const DevProjection& P, float th,
                            const SearchScratch& sc, int fill, cudaStream_t s);
void launch_proj_best(int mode, const DevFrame& f, const DevPoints& p, const DevProjection& P, float th,
                      const float* inv_sigma2, int accept_th, int32_t* best_idx, int32_t* n_found, cudaStream_t s);
void launch_frustum(const DevFrame& f, const DevPoints& p, const DevProjection& P, float cos_limit, uint8_t* in_view,
                    float* proj_x, float* proj_y, float* proj_xr, int32_t* level, float* view_cos, cudaStream_t s);
void launch_sim3_agree(const int32_t* m1, int n1, const int32_t* m2, int n2, int32_t* out12, int32_t* n_found,
                       cudaStream_t s);
// SearchByBoW: claimed2 [k2.n] zeroed, matches [n_out] = -1 on entry
void launch_bow(const DevKeyFrame& k1, const DevKeyFrame& k2, int kf_kf, float nnratio, int check_ori,
                int32_t* claimed2, int32_t* matches, int n_out, int32_t* n_matches, cudaStream_t s);
void launch_init_candidates(const DevFrame& f2, const orbfe_keypoint* keys1, const uint8_t* desc1, int n1,
                            const float* prev, float window, const SearchScratch& sc, int fill, cudaStream_t s);
int launch_init_resolve(const DevFrame& f2, const orbfe_keypoint* keys1, int n1, float nnratio, int check_ori,
                        const SearchScratch& sc, int32_t* m12, int32_t* m21, float* prev, int32_t* n_matches,
                        cudaStream_t s);

// ---- k_bow.cu: DBoW2 vocabulary transform ----
// The tree of TemplatedVocabulary (m_nodes) as device arrays: children of node i are child_list[child_start[i] ..
// child_start[i+1]) in the reference's push_back order; a node without children is a leaf (Node::isLeaf()).
struct DevVocab {
  int n_nodes;
  const int32_t* child_start;  // [n_nodes + 1]
  const int32_t* child_list;   // [n_nodes - 1]
  const uint8_t* desc;         // [n_nodes][32]
  const uint32_t* word_id;     // [n_nodes]
  const double* weight;        // [n_nodes]
};
void launch_bow_descend(const DevVocab& v, const uint8_t* desc, int n, int nid_level, uint32_t* word_of,
                        uint32_t* node_of, double* weight_of, cudaStream_t s);
int bow_build_capacity();
int launch_bow_build(int n, const uint32_t* word_of, const uint32_t* node_of, const double* weight_of, int weighting,
                     int must, int l1, uint32_t* bow_ids, double* bow_vals, uint32_t* fv_nodes, int32_t* fv_offsets,
                     int32_t* fv_features, int32_t* out_counts, cudaStream_t s);

}  // namespace orbfe
