// ref_kfdb_driver.cc -- flat C entry points over the reference's OWN KeyFrameDatabase.cc, compiled from where it lies
// against oracle/ref_shim/kfdb (see oracle/Makefile `_ref/libref_kfdb.so`).  Same flat database description as
// ora_detect_loop_candidates / ora_detect_reloc_candidates (asd_oracle.h).  TEST INFRASTRUCTURE ONLY.
#include <cstdint>
#include <memory>
#include <vector>

#include "KeyFrameDatabase.h"

using namespace ORB_SLAM2;

namespace {
struct World {
  ORBVocabulary voc;
  std::unique_ptr<KeyFrameDatabase> db;
  std::vector<std::unique_ptr<KeyFrame> > kfs;
};

void fill_bow(DBoW2::BowVector& bv, const uint32_t* w, const double* v, int n) {
  for (int e = 0; e < n; e++) bv.insert(bv.end(), std::make_pair((DBoW2::WordId)w[e], (DBoW2::WordValue)v[e]));
}

// KeyFrameDatabase::add for every slot in order, then KeyFrameDatabase::erase for the dead ones
void build(World& W, int n_kf, const int32_t* bow_off, const uint32_t* bow_word, const double* bow_val, const uint8_t* alive,
           const uint8_t* global_flag, const int32_t* neigh_off, const int32_t* neigh_slot, unsigned n_words) {
  W.voc.n_words = n_words;
  W.db.reset(new KeyFrameDatabase(W.voc));
  for (int s = 0; s < n_kf; s++) {
    W.kfs.emplace_back(new KeyFrame());
    KeyFrame* k = W.kfs.back().get();
    k->mnId = (long unsigned int)s + 1;
    k->slot = s;
    k->global = global_flag && global_flag[s];
    fill_bow(k->mBowVec, bow_word + bow_off[s], bow_val + bow_off[s], bow_off[s + 1] - bow_off[s]);
    W.db->add(k);
  }
  for (int s = 0; s < n_kf; s++) {
    if (neigh_off)
      for (int e = neigh_off[s]; e < neigh_off[s + 1]; e++) W.kfs[s]->ordered.push_back(W.kfs[neigh_slot[e]].get());
    if (alive && !alive[s]) W.db->erase(W.kfs[s].get());
  }
}
}  // namespace

extern "C" {

double ref_bow_score_l1(const uint32_t* w1, const double* v1, int n1, const uint32_t* w2, const double* v2, int n2) {
  DBoW2::BowVector a, b;
  fill_bow(a, w1, v1, n1);
  fill_bow(b, w2, v2, n2);
  return DBoW2::L1Scoring().score(a, b);
}

int ref_detect_loop_candidates(int n_kf, const int32_t* bow_off, const uint32_t* bow_word, const double* bow_val,
                               const uint8_t* alive, const uint32_t* q_word, const double* q_val, int nq,
                               const uint8_t* connected, const uint8_t* global_flag, const int32_t* neigh_off,
                               const int32_t* neigh_slot, float min_score, int only_global_map, unsigned n_words,
                               int32_t* out, int32_t* common, float* score) {
  World W;
  build(W, n_kf, bow_off, bow_word, bow_val, alive, global_flag, neigh_off, neigh_slot, n_words);
  KeyFrame q;
  q.mnId = (long unsigned int)n_kf + 1;
  fill_bow(q.mBowVec, q_word, q_val, nq);
  for (int s = 0; s < n_kf; s++)
    if (connected && connected[s]) q.connected.insert(W.kfs[s].get());
  for (int s = 0; s < n_kf; s++) W.kfs[s]->mLoopScore = -1.f;
  const std::vector<KeyFrame*> r = W.db->DetectLoopCandidates(&q, min_score, only_global_map != 0);
  for (size_t k = 0; k < r.size(); k++) out[k] = r[k]->slot;
  for (int s = 0; s < n_kf; s++) {
    const bool listed = W.kfs[s]->mnLoopQuery == q.mnId;
    if (common) common[s] = listed ? W.kfs[s]->mnLoopWords : 0;
    if (score) score[s] = listed ? W.kfs[s]->mLoopScore : -1.f;
  }
  return (int)r.size();
}

int ref_detect_reloc_candidates(int n_kf, const int32_t* bow_off, const uint32_t* bow_word, const double* bow_val,
                                const uint8_t* alive, const uint32_t* q_word, const double* q_val, int nq,
                                const uint8_t* global_flag, const int32_t* neigh_off, const int32_t* neigh_slot,
                                int only_global_map, unsigned n_words, float* reloc_score, int32_t* out, int32_t* common) {
  World W;
  build(W, n_kf, bow_off, bow_word, bow_val, alive, global_flag, neigh_off, neigh_slot, n_words);
  Frame F;
  F.mnId = (long unsigned int)n_kf + 1;
  fill_bow(F.mBowVec, q_word, q_val, nq);
  for (int s = 0; s < n_kf; s++) W.kfs[s]->mRelocScore = reloc_score[s];
  const std::vector<KeyFrame*> r = W.db->DetectRelocalizationCandidates(&F, only_global_map != 0);
  for (size_t k = 0; k < r.size(); k++) out[k] = r[k]->slot;
  for (int s = 0; s < n_kf; s++) {
    reloc_score[s] = W.kfs[s]->mRelocScore;
    if (common) common[s] = W.kfs[s]->mnRelocQuery == F.mnId ? W.kfs[s]->mnRelocWords : 0;
  }
  return (int)r.size();
}

}  // extern "C"
