// KeyFrameDatabase_cuda.cc -- replacement for the reference's src/vslam/src/KeyFrameDatabase.cc.
//
// Same class, same header (src/vslam/include/KeyFrameDatabase.h:40-74, NOT modified), same five methods; the bodies
// hand the work to libasdmatch_cuda.so (asdm_bowdb_*, include/asdmatch.h): the BowVectors of all keyframes live on the
// device, one kernel (k_bowdb_query) counts the common words and L1-scores every keyframe against the query, and the
// reference's list logic (KeyFrameDatabase.cc:80-216, 219-320) is finished inside the call.  What stays here is
// marshalling: KeyFrame* <-> slot, the caller-side gates (connected set, global-map flags, covisibility neighbours) and
// the bookkeeping members the reference leaves on its KeyFrames (KeyFrame.h:175-180).
//
// The header has no room for device state and is kept untouched, so the state of each database sits in a side table
// keyed by `this` (`mvInvertedFile` stays empty).  The fork calls the database from one thread (the reference's own
// lock_guards are commented out, KeyFrameDatabase.cc:43,51,86); the side table takes mMutex's place with its own lock.
// There is no CPU fallback: a failing GPU call aborts (asdm::Check).
#include "KeyFrameDatabase.h"

#include <map>
#include <mutex>
#include <vector>

#include "asdm_device.h"

namespace ORB_SLAM2 {

namespace {

struct DeviceDb {
  asdm_bowdb* db = nullptr;
  std::vector<KeyFrame*> kf_of_slot;      // nullptr once erased
  std::map<KeyFrame*, int32_t> slot_of;
};

std::mutex g_mu;
std::map<const KeyFrameDatabase*, DeviceDb> g_dbs;

DeviceDb& StateOf(const KeyFrameDatabase* self) {
  DeviceDb& st = g_dbs[self];
  if (!st.db) asdm::Check(asdm_bowdb_create(asdm::Context(), &st.db), "asdm_bowdb_create");
  return st;
}

// std::map<WordId, WordValue> (ascending word ids) -> the two flat arrays of the C ABI
void Flatten(const DBoW2::BowVector& bv, std::vector<uint32_t>& word, std::vector<double>& val) {
  word.clear();
  val.clear();
  word.reserve(bv.size());
  val.reserve(bv.size());
  for (DBoW2::BowVector::const_iterator it = bv.begin(); it != bv.end(); ++it) {
    word.push_back((uint32_t)it->first);
    val.push_back((double)it->second);
  }
}

// The caller-side inputs of both detections, one entry per slot.
struct Gates {
  std::vector<uint8_t> connected, global_flag;
  std::vector<int32_t> neigh_off, neigh_slot;
  asdm_bowdb_gates g;
};

void FillGates(DeviceDb& st, const std::set<KeyFrame*>* connected, float min_score, bool only_global_map, Gates& G) {
  const int ns = (int)st.kf_of_slot.size();
  G.connected.assign(ns, 0);
  G.global_flag.assign(ns, 0);
  G.neigh_off.assign(ns + 1, 0);
  G.neigh_slot.clear();
  for (int s = 0; s < ns; s++) {
    KeyFrame* kf = st.kf_of_slot[s];
    if (kf) {
      if (connected && connected->count(kf)) G.connected[s] = 1;
      if (only_global_map) G.global_flag[s] = kf->GetGlobalMapFlag() ? 1 : 0;
      // GetBestCovisibilityKeyFrames(10) (:160 / :287); neighbours that are not in the database can never be "listed"
      // (mnLoopQuery / mnRelocQuery is only ever set on database members), so they are dropped here
      const std::vector<KeyFrame*> neigh = kf->GetBestCovisibilityKeyFrames(10);
      for (size_t k = 0; k < neigh.size(); k++) {
        std::map<KeyFrame*, int32_t>::const_iterator it = st.slot_of.find(neigh[k]);
        if (it != st.slot_of.end()) G.neigh_slot.push_back(it->second);
      }
    }
    G.neigh_off[s + 1] = (int32_t)G.neigh_slot.size();
  }
  G.g.connected = connected ? G.connected.data() : nullptr;
  G.g.global_flag = only_global_map ? G.global_flag.data() : nullptr;
  G.g.neigh_off = G.neigh_off.data();
  G.g.neigh_slot = G.neigh_slot.data();
  G.g.min_score = min_score;
  G.g.only_global_map = only_global_map ? 1 : 0;
}

}  // namespace

KeyFrameDatabase::KeyFrameDatabase(const ORBVocabulary& voc) : mpVoc(&voc) {
  // :34-38 sizes the inverted file to the vocabulary; the device arena grows with the keyframes instead.
  // The class declares no destructor (and the header stays untouched), so the device state of a database that once
  // lived at this address is dropped here.
  std::lock_guard<std::mutex> lk(g_mu);
  std::map<const KeyFrameDatabase*, DeviceDb>::iterator it = g_dbs.find(this);
  if (it != g_dbs.end()) {
    if (it->second.db) asdm_bowdb_release(it->second.db);
    g_dbs.erase(it);
  }
}

void KeyFrameDatabase::add(KeyFrame* pKF) {  // :41-47
  std::lock_guard<std::mutex> lk(g_mu);
  DeviceDb& st = StateOf(this);
  std::vector<uint32_t> word;
  std::vector<double> val;
  Flatten(pKF->mBowVec, word, val);
  int32_t slot = -1;
  asdm::Check(asdm_bowdb_add(st.db, word.data(), val.data(), (int)word.size(), &slot), "asdm_bowdb_add");
  if ((int)st.kf_of_slot.size() <= slot) st.kf_of_slot.resize(slot + 1, nullptr);
  st.kf_of_slot[slot] = pKF;
  st.slot_of[pKF] = slot;
}

void KeyFrameDatabase::erase(KeyFrame* pKF) {  // :49-70
  std::lock_guard<std::mutex> lk(g_mu);
  DeviceDb& st = StateOf(this);
  std::map<KeyFrame*, int32_t>::iterator it = st.slot_of.find(pKF);
  if (it == st.slot_of.end()) return;  // the reference's scan finds nothing to remove
  asdm::Check(asdm_bowdb_erase(st.db, it->second), "asdm_bowdb_erase");
  st.kf_of_slot[it->second] = nullptr;
  st.slot_of.erase(it);
}

void KeyFrameDatabase::clear() {  // :72-76
  std::lock_guard<std::mutex> lk(g_mu);
  DeviceDb& st = StateOf(this);
  asdm::Check(asdm_bowdb_clear(st.db), "asdm_bowdb_clear");
  for (size_t s = 0; s < st.kf_of_slot.size(); s++) st.kf_of_slot[s] = nullptr;
  st.slot_of.clear();
}

std::vector<KeyFrame*> KeyFrameDatabase::DetectLoopCandidates(KeyFrame* pKF, float minScore, bool only_global_map) {  // :80-216
  std::lock_guard<std::mutex> lk(g_mu);
  DeviceDb& st = StateOf(this);
  const int ns = (int)st.kf_of_slot.size();
  if (ns == 0) return std::vector<KeyFrame*>();
  const std::set<KeyFrame*> spConnectedKeyFrames = pKF->GetConnectedKeyFrames();  // :83
  Gates G;
  FillGates(st, &spConnectedKeyFrames, minScore, only_global_map, G);
  std::vector<uint32_t> word;
  std::vector<double> val;
  Flatten(pKF->mBowVec, word, val);
  std::vector<int32_t> out(ns), common(ns);
  std::vector<float> score(ns);
  int32_t n_out = 0;
  asdm::Check(asdm_bowdb_detect_loop_full(st.db, word.data(), val.data(), (int)word.size(), &G.g, out.data(), ns, &n_out,
                                          common.data(), nullptr, score.data()),
              "asdm_bowdb_detect_loop");
  // the members the reference leaves behind (:96-110, :143): word counts of every keyframe sharing a word, the query
  // mark and -- above the common-word gate -- the score, for the keyframes that entered lKFsSharingWords
  int maxCommonWords = 0;
  for (int s = 0; s < ns; s++)
    if (st.kf_of_slot[s] && common[s] > 0 && !G.connected[s] && common[s] > maxCommonWords) maxCommonWords = common[s];
  const int minCommonWords = maxCommonWords * 0.6f;  // :128
  for (int s = 0; s < ns; s++) {
    KeyFrame* kf = st.kf_of_slot[s];
    if (!kf || common[s] <= 0) continue;
    kf->mnLoopWords = common[s];
    if (G.connected[s]) continue;
    kf->mnLoopQuery = pKF->mnId;
    if (common[s] > minCommonWords) kf->mLoopScore = score[s];
  }
  std::vector<KeyFrame*> vpLoopCandidates;
  vpLoopCandidates.reserve(n_out);
  for (int k = 0; k < n_out; k++) vpLoopCandidates.push_back(st.kf_of_slot[out[k]]);
  return vpLoopCandidates;
}

std::vector<KeyFrame*> KeyFrameDatabase::DetectRelocalizationCandidates(Frame* F, bool only_global_map) {  // :219-320
  std::lock_guard<std::mutex> lk(g_mu);
  DeviceDb& st = StateOf(this);
  const int ns = (int)st.kf_of_slot.size();
  if (ns == 0) return std::vector<KeyFrame*>();
  Gates G;
  FillGates(st, nullptr, 0.f, only_global_map, G);
  std::vector<uint32_t> word;
  std::vector<double> val;
  Flatten(F->mBowVec, word, val);
  // KeyFrame::mRelocScore outlives the query (:270 writes it, :294 reads what an earlier query left): the member stays
  // the source of truth, the library's copy is refreshed around the call
  std::vector<float> reloc(ns, 0.f);
  for (int s = 0; s < ns; s++)
    if (st.kf_of_slot[s]) reloc[s] = st.kf_of_slot[s]->mRelocScore;
  asdm::Check(asdm_bowdb_set_reloc_scores(st.db, reloc.data()), "asdm_bowdb_set_reloc_scores");
  std::vector<int32_t> out(ns), common(ns);
  int32_t n_out = 0;
  asdm::Check(asdm_bowdb_detect_reloc_full(st.db, word.data(), val.data(), (int)word.size(), &G.g, out.data(), ns, &n_out,
                                           common.data(), nullptr, nullptr),
              "asdm_bowdb_detect_reloc");
  asdm::Check(asdm_bowdb_reloc_scores(st.db, reloc.data()), "asdm_bowdb_reloc_scores");
  for (int s = 0; s < ns; s++) {
    KeyFrame* kf = st.kf_of_slot[s];
    if (!kf) continue;
    kf->mRelocScore = reloc[s];
    if (common[s] > 0) {  // :230-239
      kf->mnRelocQuery = F->mnId;
      kf->mnRelocWords = common[s];
    }
  }
  std::vector<KeyFrame*> vpRelocCandidates;
  vpRelocCandidates.reserve(n_out);
  for (int k = 0; k < n_out; k++) vpRelocCandidates.push_back(st.kf_of_slot[out[k]]);
  return vpRelocCandidates;
}

}  // namespace ORB_SLAM2
