// kernels_bowdb.cuh -- KeyFrameDatabase on the device (reference src/vslam/src/KeyFrameDatabase.cc:80-320,
// src/dbow2/DBoW2/ScoringObject.cpp:24-61): the caller side of the loop-closure sweep.  The reference walks an inverted
// file (word -> list of keyframes) to count the words each keyframe shares with the query and then merges the two
// std::map BowVectors of every keyframe above the common-word gate.  Here the BowVectors of ALL keyframes live in one
// arena (ascending word ids + values, SoA) and one warp per keyframe does both in a single pass: every lane looks its
// word up in the query (binary search, query in L1/L2), the hits of a 32-word chunk are ranked by ballot and their
// terms fabs(vi - wi) - fabs(vi) - fabs(wi) are added to a double IN ASCENDING WORD ORDER -- the order of the
// reference's merge loop (:33-54), so the sum is bit-identical.  Also returned: the smallest common word, which (with
// the add order) gives the keyframe's position in the reference's lKFsSharingWords (:93-113).
// HBM-bound: 4 B per stored word + 8 B per common word; algorithmic bytes = 12 B per BowVector entry.
#pragma once
#include <cstdint>

struct BowSlot {
  long long off;  // first entry of the keyframe in the arena
  int32_t len;    // number of BowVector entries
  int32_t alive;  // 0 after KeyFrameDatabase::erase
};

#define BOWDB_WARPS 8

__global__ void __launch_bounds__(BOWDB_WARPS * 32)
k_bowdb_query(const BowSlot* __restrict__ slots, int n_slots, const uint32_t* __restrict__ db_word,
              const double* __restrict__ db_val, const uint32_t* __restrict__ q_word, const double* __restrict__ q_val,
              int nq, int32_t* __restrict__ common, uint32_t* __restrict__ first_word, float* __restrict__ score) {
  const int lane = threadIdx.x & 31;
  const int slot = blockIdx.x * BOWDB_WARPS + (threadIdx.x >> 5);
  if (slot >= n_slots) return;
  const BowSlot rec = slots[slot];
  int cnt = 0;
  uint32_t first = 0xffffffffu;
  double acc = 0.0;
  if (rec.alive && nq > 0) {
    const uint32_t q_lo = __ldg(q_word), q_hi = __ldg(q_word + nq - 1);
    for (int base = 0; base < rec.len; base += 32) {
      const int e = base + lane;
      bool hit = false;
      uint32_t w = 0;
      double term = 0.0;
      if (e < rec.len) {
        w = __ldg(db_word + rec.off + e);
        if (w >= q_lo && w <= q_hi) {
          int lo = 0, hi = nq;
          while (lo < hi) {
            const int mid = (lo + hi) >> 1;
            if (__ldg(q_word + mid) < w) lo = mid + 1; else hi = mid;
          }
          if (lo < nq && __ldg(q_word + lo) == w) {
            hit = true;
            const double vi = __ldg(q_val + lo), wi = __ldg(db_val + rec.off + e);
            term = fabs(vi - wi) - fabs(vi) - fabs(wi);  // ScoringObject.cpp:41 (v1 = query, v2 = database keyframe)
          }
        }
      }
      unsigned m = __ballot_sync(0xffffffffu, hit);
      if (m) {
        if (first == 0xffffffffu) first = __shfl_sync(0xffffffffu, w, __ffs(m) - 1);
        cnt += __popc(m);
        while (m) {  // ascending word order: one dependent DADD per common word, as the reference's loop
          const int src = __ffs(m) - 1;
          m &= m - 1;
          acc += __shfl_sync(0xffffffffu, term, src);
        }
      }
    }
  }
  if (lane == 0) {
    common[slot] = cnt;
    first_word[slot] = first;
    score[slot] = (float)(-acc / 2.0);  // :56, then `float si = mpVoc->score(...)` (KeyFrameDatabase.cc:143)
  }
}
