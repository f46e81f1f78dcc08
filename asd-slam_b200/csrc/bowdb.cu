// bowdb.cu -- keyframe database: BoW candidate detection (KeyFrameDatabase.cc:34-320).
#include <math.h>
#include <stdio.h>
#include <string.h>

#include <algorithm>
#include <new>

#include "host_util.cuh"
#include "kernels_bowdb.cuh"

// ============================================================================ keyframe database (BoW candidates)
// KeyFrameDatabase.cc:34-320. Device: the BowVectors of every keyframe (SoA arena) + one slot record each; host: the
// mirrors the list logic needs. See kernels_bowdb.cuh for the kernel and why its sums equal the reference's.
struct asdm_bowdb {
  asdm_ctx* ctx = nullptr;
  uint32_t* d_word = nullptr;
  double* d_val = nullptr;
  BowSlot* d_slot = nullptr;
  size_t cap_entries = 0, n_entries = 0;
  int cap_slots = 0;
  std::vector<BowSlot> slots;       // host mirror of d_slot
  std::vector<float> reloc_score;   // KeyFrame::mRelocScore of every slot (persists between queries, :270 / :294)
  DevBuf query, result;             // query words | values; common | first word | score
};

static int bowdb_reserve(asdm_bowdb* db, size_t entries, int n_slots) {
  asdm_ctx* ctx = db->ctx;
  if (entries > db->cap_entries) {
    // multiple of 64 entries: k_bowdb_query_hash reads whole 16-byte quads, up to 3 words past a keyframe's last entry
    const size_t want = align_up(std::max(std::max(entries, db->cap_entries * 2), (size_t)1 << 20), 64);
    void *w = nullptr, *v = nullptr;
    CK(dev_alloc(ctx, &w, want * sizeof(uint32_t)));
    CK(dev_alloc(ctx, &v, want * sizeof(double)));
    if (db->n_entries) {
      CK(cudaMemcpyAsync(w, db->d_word, db->n_entries * sizeof(uint32_t), cudaMemcpyDeviceToDevice, ctx->stream));
      CK(cudaMemcpyAsync(v, db->d_val, db->n_entries * sizeof(double), cudaMemcpyDeviceToDevice, ctx->stream));
    }
    dev_free(ctx, db->d_word);
    dev_free(ctx, db->d_val);
    db->d_word = (uint32_t*)w;
    db->d_val = (double*)v;
    db->cap_entries = want;
  }
  if (n_slots > db->cap_slots) {
    const int want = std::max(std::max(n_slots, db->cap_slots * 2), 1024);
    void* r = nullptr;
    CK(dev_alloc(ctx, &r, (size_t)want * sizeof(BowSlot)));
    if (!db->slots.empty())
      CK(cudaMemcpyAsync(r, db->d_slot, db->slots.size() * sizeof(BowSlot), cudaMemcpyDeviceToDevice, ctx->stream));
    dev_free(ctx, db->d_slot);
    db->d_slot = (BowSlot*)r;
    db->cap_slots = want;
  }
  return ASDM_OK;
}

extern "C" int asdm_bowdb_create(asdm_ctx* ctx, asdm_bowdb** out) {
  if (!ctx || !out) return ASDM_ERR_INVALID;
  std::lock_guard<std::recursive_mutex> lk(ctx->mu);
  asdm_bowdb* db = new asdm_bowdb();
  db->ctx = ctx;
  *out = db;
  return ASDM_OK;
}

extern "C" int asdm_bowdb_release(asdm_bowdb* db) {
  if (!db) return ASDM_ERR_INVALID;
  asdm_ctx* ctx = db->ctx;
  std::lock_guard<std::recursive_mutex> lk(ctx->mu);
  cudaSetDevice(ctx->device);
  dev_free(ctx, db->d_word);
  dev_free(ctx, db->d_val);
  dev_free(ctx, db->d_slot);
  release(ctx, db->query);
  release(ctx, db->result);
  delete db;
  return ASDM_OK;
}

extern "C" int asdm_bowdb_size(asdm_bowdb* db) { return db ? (int)db->slots.size() : ASDM_ERR_INVALID; }

extern "C" int asdm_bowdb_add(asdm_bowdb* db, const uint32_t* word_id, const double* value, int n, int32_t* slot) {
  if (!db || n < 0 || (n > 0 && (!word_id || !value))) return ASDM_ERR_INVALID;
  asdm_ctx* ctx = db->ctx;
  std::lock_guard<std::recursive_mutex> lk(ctx->mu);
  CK(cudaSetDevice(ctx->device));
  for (int e = 1; e < n; e++)
    if (word_id[e] <= word_id[e - 1]) return fail(ctx, ASDM_ERR_INVALID, "BowVector word ids must ascend strictly (std::map order)");
  int rc;
  if ((rc = bowdb_reserve(db, db->n_entries + (size_t)n, (int)db->slots.size() + 1))) return rc;
  const size_t wb = align_up((size_t)n * 4, 256), vb = align_up((size_t)n * 8, 256);
  void* h = nullptr;
  if ((rc = stage_alloc(ctx, wb + vb + 256, &h))) return rc;
  char* hb = (char*)h;
  BowSlot rec;
  rec.off = (long long)db->n_entries;
  rec.len = n;
  rec.alive = 1;
  if (n) {
    memcpy(hb, word_id, (size_t)n * 4);
    memcpy(hb + wb, value, (size_t)n * 8);
    CK(cudaMemcpyAsync(db->d_word + db->n_entries, hb, (size_t)n * 4, cudaMemcpyHostToDevice, ctx->stream));
    CK(cudaMemcpyAsync(db->d_val + db->n_entries, hb + wb, (size_t)n * 8, cudaMemcpyHostToDevice, ctx->stream));
  }
  memcpy(hb + wb + vb, &rec, sizeof rec);
  CK(cudaMemcpyAsync(db->d_slot + db->slots.size(), hb + wb + vb, sizeof rec, cudaMemcpyHostToDevice, ctx->stream));
  if (slot) *slot = (int32_t)db->slots.size();
  db->slots.push_back(rec);
  db->reloc_score.push_back(0.f);
  db->n_entries += (size_t)n;
  return ASDM_OK;
}

extern "C" int asdm_bowdb_erase(asdm_bowdb* db, int32_t slot) {
  if (!db) return ASDM_ERR_INVALID;
  asdm_ctx* ctx = db->ctx;
  std::lock_guard<std::recursive_mutex> lk(ctx->mu);
  CK(cudaSetDevice(ctx->device));
  if (slot < 0 || slot >= (int)db->slots.size()) return fail(ctx, ASDM_ERR_INVALID, "unknown keyframe slot");
  if (!db->slots[slot].alive) return ASDM_OK;  // the reference's erase of an absent keyframe finds nothing (:57-67)
  db->slots[slot].alive = 0;
  void* h = nullptr;
  int rc;
  if ((rc = stage_alloc(ctx, sizeof(BowSlot), &h))) return rc;
  memcpy(h, &db->slots[slot], sizeof(BowSlot));
  CK(cudaMemcpyAsync(db->d_slot + slot, h, sizeof(BowSlot), cudaMemcpyHostToDevice, ctx->stream));
  return ASDM_OK;
}

extern "C" int asdm_bowdb_clear(asdm_bowdb* db) {
  if (!db) return ASDM_ERR_INVALID;
  asdm_ctx* ctx = db->ctx;
  std::lock_guard<std::recursive_mutex> lk(ctx->mu);
  CK(cudaSetDevice(ctx->device));
  // the inverted file is emptied (:72-76); the keyframes keep their slots (and their mRelocScore), none is listed any more
  for (size_t s = 0; s < db->slots.size(); s++) db->slots[s].alive = 0;
  if (!db->slots.empty()) {
    void* h = nullptr;
    int rc;
    if ((rc = stage_alloc(ctx, db->slots.size() * sizeof(BowSlot), &h))) return rc;
    memcpy(h, db->slots.data(), db->slots.size() * sizeof(BowSlot));
    CK(cudaMemcpyAsync(db->d_slot, h, db->slots.size() * sizeof(BowSlot), cudaMemcpyHostToDevice, ctx->stream));
  }
  return ASDM_OK;
}

// queues the device pass and waits; the three result arrays (pinned staging memory, valid until the next staged call)
static int bowdb_run(asdm_bowdb* db, const uint32_t* word_id, const double* value, int n, const int32_t** common,
                     const uint32_t** first_word, const float** score) {
  asdm_ctx* ctx = db->ctx;
  const int ns = (int)db->slots.size();
  for (int e = 1; e < n; e++)
    if (word_id[e] <= word_id[e - 1]) return fail(ctx, ASDM_ERR_INVALID, "BowVector word ids must ascend strictly (std::map order)");
  ctx->stat_launches = 0;
  const size_t wb = align_up((size_t)std::max(n, 1) * 4, 256), vb = align_up((size_t)std::max(n, 1) * 8, 256);
  const size_t rb = align_up((size_t)std::max(ns, 1) * 4, 256);
  int rc;
  if ((rc = ensure(ctx, db->query, wb + vb))) return rc;
  if ((rc = ensure(ctx, db->result, 3 * rb + 256))) return rc;  // common | first word | score | slot counter
  void* h = nullptr;
  if ((rc = stage_alloc(ctx, wb + vb + 3 * rb, &h))) return rc;
  char* hb = (char*)h;
  if (n) {
    memcpy(hb, word_id, (size_t)n * 4);
    memcpy(hb + wb, value, (size_t)n * 8);
    CK(cudaMemcpyAsync(db->query.p, hb, wb + vb, cudaMemcpyHostToDevice, ctx->stream));
  }
  char* dq = (char*)db->query.p;
  char* dr = (char*)db->result.p;
  char* hr = hb + wb + vb;
  if (ns) {
    // the query as a shared-memory hash table + Bloom filter when they fit and no word collides with the empty mark
    int log2cap = 6;
    while ((1 << log2cap) * 2 < 3 * n) log2cap++;  // load factor <= 2/3 (the filter in front keeps probes rare)
    int log2bits = 15;  // Bloom filter in front of the table: >= 128 bits per query word, 4 KB .. 64 KB
    while (log2bits < 19 && (1 << log2bits) < ctx->opt_bowdb_filter_bits * n) log2bits++;
    const size_t tab_bytes = ((size_t)8 << log2cap) + ((size_t)1 << (log2bits - 3));
    if (!ctx->opt_bowdb_search && n > 0 && tab_bytes <= (size_t)200 * 1024 && word_id[n - 1] != 0xffffffffu) {
      CK(cudaFuncSetAttribute(k_bowdb_query_hash, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));  // per device
      int per_sm = 1;
      CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_bowdb_query_hash, BOWDB_HASH_WARPS * 32, tab_bytes));
      const int grid = std::min((ns + BOWDB_HASH_WARPS - 1) / BOWDB_HASH_WARPS, std::max(per_sm, 1) * ctx->sm_count);
      CK(cudaMemsetAsync(dr + 3 * rb, 0, 4, ctx->stream));
      k_bowdb_query_hash<<<grid, BOWDB_HASH_WARPS * 32, tab_bytes, ctx->stream>>>(
          db->d_slot, ns, db->d_word, db->d_val, (const uint32_t*)dq, (const double*)(dq + wb), n, log2cap, log2bits,
          (int32_t*)dr, (uint32_t*)(dr + rb), (float*)(dr + 2 * rb), (int*)(dr + 3 * rb));
    } else {
      k_bowdb_query<<<(ns + BOWDB_WARPS - 1) / BOWDB_WARPS, BOWDB_WARPS * 32, 0, ctx->stream>>>(
          db->d_slot, ns, db->d_word, db->d_val, (const uint32_t*)dq, (const double*)(dq + wb), n, (int32_t*)dr,
          (uint32_t*)(dr + rb), (float*)(dr + 2 * rb));
    }
    LAUNCHED(ctx);
    CK(cudaGetLastError());
    CK(cudaMemcpyAsync(hr, dr, 3 * rb, cudaMemcpyDeviceToHost, ctx->stream));
  }
  CK(cudaStreamSynchronize(ctx->stream));
  // NOT synced(ctx): the results are read from the staging block by the caller of this helper
  *common = (const int32_t*)hr;
  *first_word = (const uint32_t*)(hr + rb);
  *score = (const float*)(hr + 2 * rb);
  return ASDM_OK;
}

extern "C" int asdm_bowdb_query(asdm_bowdb* db, const uint32_t* word_id, const double* value, int n, int32_t* common,
                                uint32_t* first_word, float* score) {
  if (!db || n < 0 || (n > 0 && (!word_id || !value))) return ASDM_ERR_INVALID;
  asdm_ctx* ctx = db->ctx;
  std::lock_guard<std::recursive_mutex> lk(ctx->mu);
  CK(cudaSetDevice(ctx->device));
  const int32_t* c;
  const uint32_t* f;
  const float* sc;
  int rc;
  if ((rc = bowdb_run(db, word_id, value, n, &c, &f, &sc))) return rc;
  const size_t ns = db->slots.size();
  if (common) memcpy(common, c, ns * 4);
  if (first_word) memcpy(first_word, f, ns * 4);
  if (score) memcpy(score, sc, ns * 4);
  synced(ctx);
  return ASDM_OK;
}

// lKFsSharingWords (:93-113 / :226-243): a keyframe enters the list at the smallest word it shares with the query, and the
// keyframes of one inverted list stand in add order -> sort by (first common word, slot).  Only the members above the
// common-word gate are ever looked at in list order (:137 / :264), and the gate itself (the maximum over the list) does
// not depend on the order, so the gate is found in one pass over all members and only the survivors are sorted.
static void bowdb_sharing_list(const asdm_bowdb* db, const int32_t* common, const uint32_t* first_word, const uint8_t* skip,
                               const uint8_t* need, float gate_factor, std::vector<int32_t>& list, int* min_common) {
  list.clear();
  const int ns = (int)db->slots.size();
  auto member = [&](int s) { return db->slots[s].alive && common[s] > 0 && !(skip && skip[s]) && (!need || need[s]); };
  int maxCommonWords = 0;
  bool any = false;
  for (int s = 0; s < ns; s++)
    if (member(s)) {
      any = true;
      maxCommonWords = std::max(maxCommonWords, common[s]);
    }
  *min_common = (int)(maxCommonWords * gate_factor);  // :128 / :256 (int * float -> float, truncated)
  if (!any) return;
  for (int s = 0; s < ns; s++)
    if (member(s) && common[s] > *min_common) list.push_back(s);
  std::sort(list.begin(), list.end(), [&](int32_t a, int32_t b) {
    return first_word[a] != first_word[b] ? first_word[a] < first_word[b] : a < b;
  });
}

static int bowdb_detect(asdm_bowdb* db, bool reloc, const uint32_t* word_id, const double* value, int n,
                        const asdm_bowdb_gates* g, int32_t* out, int cap, int32_t* n_out, int32_t* common_out = nullptr,
                        uint32_t* first_out = nullptr, float* score_out = nullptr) {
  if (!db || !g || !n_out || n < 0 || (n > 0 && (!word_id || !value)) || cap < 0 || (cap > 0 && !out)) return ASDM_ERR_INVALID;
  asdm_ctx* ctx = db->ctx;
  std::lock_guard<std::recursive_mutex> lk(ctx->mu);
  CK(cudaSetDevice(ctx->device));
  *n_out = 0;
  const int ns = (int)db->slots.size();
  if (g->only_global_map && !g->global_flag && ns) return fail(ctx, ASDM_ERR_INVALID, "only_global_map needs global_flag");
  if (g->neigh_off && ns) {
    if (g->neigh_off[0] != 0) return fail(ctx, ASDM_ERR_INVALID, "bad neighbour offsets");
    for (int s = 0; s < ns; s++) {
      if (g->neigh_off[s + 1] < g->neigh_off[s]) return fail(ctx, ASDM_ERR_INVALID, "neighbour offsets not ascending");
      for (int e = g->neigh_off[s]; e < g->neigh_off[s + 1]; e++)
        if (g->neigh_slot[e] < 0 || g->neigh_slot[e] >= ns) return fail(ctx, ASDM_ERR_INVALID, "neighbour names an unknown slot");
    }
  }
  const int32_t* common;
  const uint32_t* first_word;
  const float* score;
  int rc;
  if ((rc = bowdb_run(db, word_id, value, n, &common, &first_word, &score))) return rc;
  if (common_out) memcpy(common_out, common, (size_t)ns * 4);
  if (first_out) memcpy(first_out, first_word, (size_t)ns * 4);
  if (score_out) memcpy(score_out, score, (size_t)ns * 4);
  const uint8_t* glob = g->only_global_map ? g->global_flag : nullptr;
  std::vector<int32_t> sharing;  // the members of lKFsSharingWords above the common-word gate, in list order
  int minCommonWords = 0;
  // loop: connected keyframes never enter the list (:104); relocalisation: the global-map gate sits on the list (:236)
  bowdb_sharing_list(db, common, first_word, reloc ? nullptr : g->connected, reloc ? glob : nullptr, reloc ? 0.8f : 0.6f,
                     sharing, &minCommonWords);
  synced(ctx);  // (the staging block stays valid: nothing else is staged before this call returns)
  if (sharing.empty()) return ASDM_OK;
  // "listed": mnLoopQuery == pKF->mnId (the sharing list) / mnRelocQuery == F->mnId (every alive keyframe sharing a word)
  auto listed = [&](int s2) {
    if (!db->slots[s2].alive || common[s2] <= 0) return false;
    return reloc ? true : !(g->connected && g->connected[s2]);
  };
  std::vector<std::pair<float, int32_t> > scoreAndMatch, acc;
  for (int32_t s : sharing) {
    if (common[s] > minCommonWords) {
      const float si = score[s];
      if (reloc) {
        db->reloc_score[s] = si;  // :270
        scoreAndMatch.push_back(std::make_pair(si, s));
      } else if (si >= g->min_score && (!glob || glob[s])) {  // :146-148
        scoreAndMatch.push_back(std::make_pair(si, s));
      }
    }
  }
  if (scoreAndMatch.empty()) return ASDM_OK;
  float bestAccScore = reloc ? 0.f : g->min_score;  // :280 / :155
  for (const std::pair<float, int32_t>& sm : scoreAndMatch) {
    const int32_t s = sm.second;
    float bestScore = sm.first, accScore = sm.first;
    int32_t best = s;
    if (g->neigh_off)
      for (int e = g->neigh_off[s]; e < g->neigh_off[s + 1]; e++) {
        const int32_t s2 = g->neigh_slot[e];
        if (!listed(s2)) continue;
        float s2score;
        if (reloc) {
          s2score = db->reloc_score[s2];  // :294 -- fresh if scored in this query, else what an earlier query left
        } else {
          if (common[s2] <= minCommonWords) continue;  // :168
          s2score = score[s2];                         // mLoopScore, computed above for every keyframe over the gate
        }
        accScore += s2score;
        if (s2score > bestScore) {
          best = s2;
          bestScore = s2score;
        }
      }
    acc.push_back(std::make_pair(accScore, best));
    if (accScore > bestAccScore) bestAccScore = accScore;
  }
  const float minScoreToRetain = (reloc ? 0.75f : 0.55f) * bestAccScore;  // :305 / :184
  std::vector<uint8_t> added((size_t)ns, 0);
  int k = 0;
  for (const std::pair<float, int32_t>& a : acc)
    if (a.first > minScoreToRetain && !added[a.second]) {
      added[a.second] = 1;
      if (k >= cap) return fail(ctx, ASDM_ERR_INVALID, "more candidates than the output holds");
      out[k++] = a.second;
    }
  *n_out = k;
  return ASDM_OK;
}

extern "C" int asdm_bowdb_detect_loop(asdm_bowdb* db, const uint32_t* word_id, const double* value, int n,
                                      const asdm_bowdb_gates* g, int32_t* out, int cap, int32_t* n_out) {
  return bowdb_detect(db, false, word_id, value, n, g, out, cap, n_out);
}

extern "C" int asdm_bowdb_detect_reloc(asdm_bowdb* db, const uint32_t* word_id, const double* value, int n,
                                       const asdm_bowdb_gates* g, int32_t* out, int cap, int32_t* n_out) {
  return bowdb_detect(db, true, word_id, value, n, g, out, cap, n_out);
}

extern "C" int asdm_bowdb_detect_loop_full(asdm_bowdb* db, const uint32_t* word_id, const double* value, int n,
                                           const asdm_bowdb_gates* g, int32_t* out, int cap, int32_t* n_out,
                                           int32_t* common, uint32_t* first_word, float* score) {
  return bowdb_detect(db, false, word_id, value, n, g, out, cap, n_out, common, first_word, score);
}

extern "C" int asdm_bowdb_detect_reloc_full(asdm_bowdb* db, const uint32_t* word_id, const double* value, int n,
                                            const asdm_bowdb_gates* g, int32_t* out, int cap, int32_t* n_out,
                                            int32_t* common, uint32_t* first_word, float* score) {
  return bowdb_detect(db, true, word_id, value, n, g, out, cap, n_out, common, first_word, score);
}

extern "C" int asdm_bowdb_set_reloc_scores(asdm_bowdb* db, const float* score) {
  if (!db || (!score && !db->reloc_score.empty())) return ASDM_ERR_INVALID;
  std::lock_guard<std::recursive_mutex> lk(db->ctx->mu);
  if (!db->reloc_score.empty()) memcpy(db->reloc_score.data(), score, db->reloc_score.size() * sizeof(float));
  return ASDM_OK;
}

extern "C" int asdm_bowdb_reloc_scores(asdm_bowdb* db, float* score) {
  if (!db || (!score && !db->reloc_score.empty())) return ASDM_ERR_INVALID;
  std::lock_guard<std::recursive_mutex> lk(db->ctx->mu);
  if (!db->reloc_score.empty()) memcpy(score, db->reloc_score.data(), db->reloc_score.size() * sizeof(float));
  return ASDM_OK;
}

