/*
 * asd_oracle.h -- CPU ORACLE for the ASD descriptor-matching hot path.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing under oracle/ is part of the product: only tests/,
 * __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may load this
 * library, and there only as the checker or the timed CPU baseline.  The product
 * (libasdmatch_cuda.so) never links, loads or calls it.
 *
 * It restates, over flat arrays, the reference matcher of mataiyuan/ASD-SLAM
 * (src/vslam/src/ORBmatcher.cc, Frame.cc, KeyFrame.cc).  Every function cites the reference
 * file:line it follows.  Parity status: PINNED against the reference's own ORBmatcher.cc
 * compiled from where it lies (oracle/_ref, see oracle/Makefile `ref` target and
 * tests/test_oracle_vs_ref.py).  The reference holds no golden vectors for this path
 * (SURVEY.md section 4).
 *
 * Float environment: the reference was built by GCC 5.4 with only --std=c++11 on x86-64, i.e.
 * scalar SSE fp32, round-to-nearest, no FMA contraction, no reassociation.  This file must be
 * compiled with -ffp-contract=off -fno-fast-math (oracle/Makefile does).
 */
#ifndef ASD_ORACLE_H
#define ASD_ORACLE_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define ORA_DESC_DIM 128
#define ORA_HISTO_LENGTH 30 /* ORBmatcher.cc:39 */

/* cv::KeyPoint fields the matcher reads: mvKeysUn[i].pt.x/.pt.y/.angle/.octave */
typedef struct {
  float x, y, angle;
  int32_t octave;
} ora_kp;

/* Frame/KeyFrame grid geometry (Frame.h:37-38; Frame.cc:34; KeyFrame.h:217-220).
 * KeyFrame keeps integer bounds; they enter every expression converted to float, so a float
 * copy is exact. */
typedef struct {
  float min_x, min_y, max_x, max_y;
  float inv_w, inv_h; /* mfGridElementWidthInv / HeightInv */
  int32_t cols, rows; /* 64 x 48 */
} ora_grid;

/* DBoW2::FeatureVector (src/dbow2/include/FeatureVector.h:21-22) flattened to CSR:
 * node ids ascending; idx[off[k]..off[k+1]) are the feature indices of node k in
 * insertion (= ascending feature index) order. */
typedef struct {
  int32_t n_nodes;
  const uint32_t* node_ids;
  const int32_t* off;
  const int32_t* idx;
} ora_featvec;

/* a1: ORBmatcher::DescriptorDistance (ORBmatcher.cc:1629-1650) */
float ora_descriptor_distance(const float* a, const float* b);

/* a12: ORBmatcher::ComputeThreeMaxima (ORBmatcher.cc:1584-1625) on bin sizes */
void ora_compute_three_maxima(const int32_t* sizes, int L, int32_t* ind1, int32_t* ind2, int32_t* ind3);

/* rotation-bin idiom (e.g. ORBmatcher.cc:245-252) */
int ora_rot_bin(float angle1, float angle2);

/* a3: Frame::PosInGrid + AssignFeaturesToGrid (Frame.cc:123-138, 276-286; KeyFrame.cc:116-143).
 * cell_start has cols*rows+1 entries (cell = ix*rows+iy), cell_items up to n entries,
 * cell_of[i] = ix*rows+iy or -1 when the keypoint falls outside the grid. */
void ora_assign_features_to_grid(const ora_kp* kps, int n, const ora_grid* g, int32_t* cell_start,
                                 int32_t* cell_items, int32_t* cell_of);

/* a3: Frame::GetFeaturesInArea (Frame.cc:219-274).  KeyFrame::GetFeaturesInArea
 * (KeyFrame.cc:839-878) is the same scan with min_level = max_level = -1.
 * Returns the number of indices written to out (scan order). */
int ora_get_features_in_area(const ora_kp* kps, int n, const ora_grid* g, const int32_t* cell_start,
                             const int32_t* cell_items, float x, float y, float r, int min_level,
                             int max_level, int32_t* out);

/* a8: SearchByBoW(KeyFrame*, Frame&, out) (ORBmatcher.cc:156-297).
 * valid_kf[i] = KF feature i holds a map point that is not bad.
 * match_f[j] (size n_f) = KF feature index whose map point was assigned to frame feature j, or -1.
 * Returns nmatches. */
int ora_search_by_bow_kf_f(const float* desc_kf, const ora_kp* kps_kf, int n_kf, const ora_featvec* fv_kf,
                           const uint8_t* valid_kf, const float* desc_f, const ora_kp* kps_f, int n_f,
                           const ora_featvec* fv_f, float nnratio, int check_ori, int32_t* match_f,
                           float* dist_f);

/* a9: SearchByBoW(KeyFrame*, KeyFrame*, out) (ORBmatcher.cc:533-666).
 * match12[i] (size n1) = feature index in KF2 or -1. Returns nmatches. */
int ora_search_by_bow_kf_kf(const float* desc1, const ora_kp* kps1, int n1, const ora_featvec* fv1,
                            const uint8_t* valid1, const float* desc2, const ora_kp* kps2, int n2,
                            const ora_featvec* fv2, const uint8_t* valid2, float nnratio, int check_ori,
                            int32_t* match12, float* dist12);

/* a11: SearchForTriangulation (ORBmatcher.cc:669-822) with CheckDistEpipolarLine (:136-153).
 * has_mp1/2[i] = feature already holds a map point (such features are skipped).
 * F12 row-major 3x3; (ex,ey) epipole in image 2 (:675-682, computed by the caller);
 * scale_factors2 / level_sigma2_2 = pKF2->mvScaleFactors / mvLevelSigma2.
 * match12[i] (size n1) = idx2 or -1. Returns nmatches. */
int ora_search_for_triangulation(const float* desc1, const ora_kp* kps1, int n1, const ora_featvec* fv1,
                                 const uint8_t* has_mp1, const float* desc2, const ora_kp* kps2, int n2,
                                 const ora_featvec* fv2, const uint8_t* has_mp2, const float* F12, float ex,
                                 float ey, const float* scale_factors2, const float* level_sigma2_2,
                                 int check_ori, int32_t* match12, float* dist12);

/* Generic projection-gated search: the common body of the four SearchByProjection overloads
 * (ORBmatcher.cc:44-122, 300-413, 1318-1452, 1455-1582).  The caller supplies, per query, what
 * the reference derives from poses / map points before it reaches the descriptor loop. */
enum {
  ORA_PROJ_RATIO_LEVEL = 0, /* :44-122  best & second with octaves, TH_HIGH, ratio only when same octave */
  ORA_PROJ_BEST_ONLY = 1    /* :300-413, :1318-1452, :1455-1582  best only */
};
typedef struct {
  int32_t mode;
  float th_accept;     /* accept if best <= th_accept (TH_HIGH, ORBdist or TH_LOW) */
  float best_init;     /* 256 (:75, :380, :1518) or 100 (:1387) */
  float nnratio;       /* mode RATIO_LEVEL only */
  int32_t check_ori;   /* rotation histogram on (:1413-1449, :1543-1579) */
  int32_t count_twice; /* :116-117 increments nmatches twice */
  double chi2_max;     /* > 0: Fuse's gate e2*inv_level_sigma2[octave] > chi2_max -> skip (:919-926) */
  int32_t n_levels;
  int32_t reserved;
  float inv_level_sigma2[32];
} ora_proj_params;

typedef struct {
  uint8_t valid;  /* query survived the reference's pre-descriptor gates */
  uint8_t blocks; /* assigned map point has Observations()>0 (or: any non-null blocks) */
  int16_t pad;
  float u, v, r;  /* window centre and half-size handed to GetFeaturesInArea */
  int32_t min_level, max_level;
  float angle;    /* query keypoint angle for the rotation histogram */
} ora_proj_query;

/* blocked0[j] = target keypoint j is unavailable from the start.
 * assign[j] (size n_t) = query index finally assigned to keypoint j; -1 untouched; -2 cleared by the rotation check.
 * Returns nmatches (doubled when count_twice). */
int ora_search_by_projection(const float* desc_t, const ora_kp* kps_t, int n_t, const ora_grid* g,
                             const ora_proj_query* q, const float* desc_q, int n_q, const uint8_t* blocked0,
                             const ora_proj_params* p, int32_t* assign, float* dist_t);

/* Same search seen from the queries (no exclusivity needed): query_best[i] = keypoint chosen by query i or -1. */
int ora_search_by_projection_q(const float* desc_t, const ora_kp* kps_t, int n_t, const ora_grid* g,
                               const ora_proj_query* q, const float* desc_q, int n_q, const uint8_t* blocked0,
                               const ora_proj_params* p, int32_t* query_best, float* query_dist);

/* a4 helper: ORBmatcher::RadiusByViewingCos (ORBmatcher.cc:126-132) and the window of :62-68 */
float ora_radius_by_viewing_cos(float view_cos);
float ora_local_map_window(float view_cos, float th, float scale_factor);

/* MapPoint::PredictScale (MapPoint.cc:421-453) */
int ora_predict_scale(float max_distance, float current_dist, float log_scale_factor, int n_levels);

/* a10: SearchForInitialization (ORBmatcher.cc:416-531).
 * prev_matched (2*n1 floats, x,y interleaved) is read and rewritten (:526-528).
 * matches12[i] (size n1) = idx2 or -1. Returns nmatches. */
int ora_search_for_initialization(const float* desc1, const ora_kp* kps1, int n1, const float* desc2,
                                  const ora_kp* kps2, int n2, const ora_grid* g2, float* prev_matched,
                                  int window_size, float nnratio, int check_ori, int32_t* matches12,
                                  float* dist12);

/* Descriptor stage of Frame::ComputeStereoMatches (Frame.cc:360-444; dead code upstream, BASELINE config 3).
 * kps_l = mvKeys, kps_r = mvKeysRight.  best_r[i] = right index when bestDist < thOrbDist, else -1.
 * Rows outside [0,n_rows) are skipped (the reference would index out of bounds). */
void ora_stereo_row_match(const float* desc_l, const ora_kp* kps_l, int n_l, const float* desc_r,
                          const ora_kp* kps_r, int n_r, const float* scale_factors, int n_rows, float min_d,
                          float max_d, int32_t* best_r, float* best_dist);

/* Batched drivers for BASELINE configs 4 and 5 (the loop bodies of LocalMapping.cc:335-370 and
 * LoopClosing.cc:291-318): pair p uses keyframes ia[p], ib[p] out of n_kf keyframes that all
 * hold n features. threads>1 runs pairs on OpenMP threads (pairs are independent). */
typedef struct {
  const float* desc;     /* n x 128 */
  const ora_kp* kps;     /* n */
  const ora_featvec* fv;
  const uint8_t* mask;   /* valid (config 5) or has_mp (config 4) */
} ora_keyframe;

void ora_batch_bow_kf_kf(const ora_keyframe* kfs, int n, const int32_t* ia, const int32_t* ib, int n_pairs,
                         float nnratio, int check_ori, int threads, int32_t* match12 /* n_pairs*n */,
                         float* dist12, int32_t* nmatches /* n_pairs */);
void ora_batch_triangulation(const ora_keyframe* kfs, int n, const int32_t* ia, const int32_t* ib, int n_pairs,
                             const float* F12 /* 9*n_pairs */, const float* epipoles /* 2*n_pairs */,
                             const float* scale_factors, const float* level_sigma2, int check_ori, int threads,
                             int32_t* match12, float* dist12, int32_t* nmatches);

/* MapPoint::ComputeDistinctiveDescriptors (MapPoint.cc:271-338), batched over map points: all-pairs
 * DescriptorDistance between the point's observed descriptors (:306-318), per observation the median
 * vDists[0.5*(N-1)] of its sorted row INCLUDING the zero self distance (:323-327), the first observation with the
 * smallest median below 100 wins (:329-333; BestIdx stays 0 otherwise). Observations (obs_kf[e], obs_row[e]),
 * e in [obs_off[p], obs_off[p+1]), are given in the reference's iteration order (std::map<KeyFrame*, size_t>).
 * best[p] = position within the point's list, -1 when it has no observation (:285-286, :300-301). */
void ora_distinctive(const float* const* kf_desc, int n_points, const int32_t* obs_off, const int32_t* obs_kf,
                     const int32_t* obs_row, int32_t* best, float* best_median);

/* DBoW2 vocabulary tree: TemplatedVocabulary<FSift::TDescriptor, FSift>::transform(features, BowVector&,
 * FeatureVector&, levelsup) (src/dbow2/include/TemplatedVocabulary.h:1125-1192) with the per-feature descent of
 * :1216-1257 and FSift::distance (src/dbow2/DBoW2/FSift.cpp:86-101; fp32 products added to a double in index order).
 * The tree is given flat, in the state TemplatedVocabulary::load leaves it (:1455-1504): node 0 = root, children in
 * push_back order. Outputs: word_id / node_id / in_fv per feature (node_id at level L - levelsup, 0 = root; when a
 * branch ends above that level the reference reads an uninitialised NodeId (:1148) and this restatement reports the
 * leaf), and the BowVector (ascending word id) after BowVector::addWeight / addIfNotExist and normalize
 * (BowVector.cpp:35-84). Returns the BowVector size. */
int ora_voc_transform(int L, int weighting, int scoring, int n_nodes, const int32_t* child_off, const int32_t* child_idx,
                      const float* node_desc, const double* weight, const int32_t* word_of_node, const float* desc, int n,
                      int levelsup, int32_t* word_id, int32_t* node_id, uint8_t* in_fv, uint32_t* bow_id, double* bow_val);

/* chamo.map descriptor block of one frame: System::saveToVisualMap (System.cc:333-343) copies mDescriptors (N x 128
 * CV_32F) into the column-major Eigen matrix descriptors(D, N); vm::save_visual_map writes int32 rows() (= D), int32
 * cols() (= N) and then descriptors(k, j) for j outer, k inner (visual_map_seri.cc:134-146); vm::loader_visual_map
 * reads the same back (:272-285). ora_map_write_block returns the bytes written (8 + 512 N); ora_map_read_block
 * returns N (or -1 when the block is truncated) and stores D in *width; rows[j*D + k] = descriptors(k, j). */
long ora_map_write_block(const float* desc, int n, int d, uint8_t* out);
int ora_map_read_block(const uint8_t* block, long bytes, float* rows, int cap_rows, int* width);

/* ---- KeyFrameDatabase: BoW candidate detection, the caller that produces the loop-closure sweep's candidate pairs ----
 * The database is given flat, in KeyFrameDatabase::add order (KeyFrameDatabase.cc:41-47): keyframe slot s owns the
 * BowVector entries [bow_off[s], bow_off[s+1]) (ascending word id, as std::map iterates); alive[s] = 0 after
 * KeyFrameDatabase::erase (:49-70).  Every inverted list holds its keyframes in add order.
 * ora_bow_score_l1: DBoW2::L1Scoring::score (src/dbow2/DBoW2/ScoringObject.cpp:24-61): merge of the two vectors,
 * score += fabs(vi - wi) - fabs(vi) - fabs(wi) over the common words in ascending word order, result -score / 2. */
double ora_bow_score_l1(const uint32_t* w1, const double* v1, int n1, const uint32_t* w2, const double* v2, int n2);

/* KeyFrameDatabase::DetectLoopCandidates (KeyFrameDatabase.cc:80-216): walk of the inverted file over the query's words
 * (:93-113; connected keyframes never enter the list), maxCommonWords (:121-126), minCommonWords = max * 0.6f (:128),
 * L1 scores of the keyframes above it and the minScore / only_global_map gates (:133-150), accumulation over
 * GetBestCovisibilityKeyFrames(10) (neigh_off / neigh_slot, CSR per slot) in float (:156-181), 0.55f * best (:184),
 * unique best keyframes in list order (:189-200).  common / score (may be NULL) receive mnLoopWords / mLoopScore of every
 * listed keyframe (score only where it was computed; others -1).  Returns the number of candidates written to out. */
int ora_detect_loop_candidates(int n_kf, const int32_t* bow_off, const uint32_t* bow_word, const double* bow_val,
                               const uint8_t* alive, const uint32_t* q_word, const double* q_val, int nq,
                               const uint8_t* connected, const uint8_t* global_flag, const int32_t* neigh_off,
                               const int32_t* neigh_slot, float min_score, int only_global_map, int32_t* out,
                               int32_t* common, float* score);
/* KeyFrameDatabase::DetectRelocalizationCandidates (:219-320): 0.8f common-word gate (:253), no minScore, bestAccScore
 * from 0 (:276), 0.75f (:303).  The neighbour accumulation (:286-296) reads mRelocScore of EVERY neighbour that shares a
 * word with the frame, scored in this query or not: reloc_score[n_kf] is that persistent member (in/out; the reference
 * leaves it uninitialised at construction, KeyFrame.cc:150 -- this restatement starts it at what the caller passes). */
int ora_detect_reloc_candidates(int n_kf, const int32_t* bow_off, const uint32_t* bow_word, const double* bow_val,
                                const uint8_t* alive, const uint32_t* q_word, const double* q_val, int nq,
                                const uint8_t* global_flag, const int32_t* neigh_off, const int32_t* neigh_slot,
                                int only_global_map, float* reloc_score, int32_t* out, int32_t* common);

int ora_max_threads(void);

#ifdef __cplusplus
}
#endif
#endif
