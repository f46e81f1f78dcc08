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

// ---------------------------------------------------------------------------------------------------------------------
// k_bowdb_query_hash -- the same pass with the query held in SHARED memory as an open-addressing hash table
// (word -> position in the query, linear probing, load factor <= 2/3): a membership test is one or two LDS instead of
// the 10-12 dependent global loads of a binary search, so the kernel streams the word arena at close to what HBM
// delivers (the binary-search kernel above is instruction/latency bound: 47% SM throughput at 6% DRAM, profiles/
// r01l_bowdb_query.txt).  Persistent: the grid is sized to fill the SMs once, every CTA builds the table once and its
// warps walk the keyframes grid-stride.  The summation order (ascending word id, one DADD per common word) and hence
// every bit of the score is unchanged.  0xffffffff marks an empty bucket: a query holding that word id (no vocabulary
// has 2^32 words) is served by k_bowdb_query, as are queries whose table would not fit.
#define BOWDB_HASH_WARPS 8
__device__ __forceinline__ uint32_t bowdb_hash(uint32_t w, int shift) { return (w * 2654435761u) >> shift; }

// Each lane loads FOUR consecutive words with one 16-byte load (a warp chunk = 128 entries = 512 B; chunks start at
// 16-byte boundaries of the arena, entries outside the keyframe are masked) and the chunk after the current one is
// already in flight, so ~1 KB per warp is outstanding -- the kernel is bound by load latency, not by issue.  Entry
// order inside a chunk is lane-major, component-minor; the hits are summed in exactly that order.  In front of the table
// sits a one-hash Bloom filter over the same words (>= 128 bits per query word): one LDS and a bit test turn away more
// than 99 % of the entries of a keyframe that shares nothing with the query, so the probing loop (whose trip count is the
// maximum over the 32 lanes) only runs for the few chunks that hold a possible hit.  Keyframes are
// handed out through a global counter (the ones that share most of their words with the query carry a serial chain
// of ~1000 dependent DADDs, 30x the cost of a keyframe without hits): no warp is left with two of them at the end.
// Measured and dropped (profiles/r01q/r01r notes in DESIGN.md): 16 warps per table with claims pipelined two keyframes
// ahead (no gain) and four keyframes per claim (slower: fewer CTAs by registers, coarser tail).
__global__ void __launch_bounds__(BOWDB_HASH_WARPS * 32)
k_bowdb_query_hash(const BowSlot* __restrict__ slots, int n_slots, const uint32_t* __restrict__ db_word,
                   const double* __restrict__ db_val, const uint32_t* __restrict__ q_word, const double* __restrict__ q_val,
                   int nq, int log2cap, int log2bits, int32_t* __restrict__ common,
                   uint32_t* __restrict__ first_word, float* __restrict__ score, int* __restrict__ next_slot) {
  extern __shared__ uint32_t bow_tab[];  // keys[cap] | position[cap] | filter[2^log2bits / 32]
  const int cap = 1 << log2cap, shift = 32 - log2cap, fshift = 32 - log2bits;
  uint32_t* keys = bow_tab;
  uint32_t* pos = bow_tab + cap;
  uint32_t* filt = bow_tab + 2 * cap;
  for (int i = threadIdx.x; i < cap; i += blockDim.x) keys[i] = 0xffffffffu;
  for (int i = threadIdx.x; i < (1 << (log2bits - 5)); i += blockDim.x) filt[i] = 0u;
  __syncthreads();
  for (int i = threadIdx.x; i < nq; i += blockDim.x) {
    const uint32_t w = __ldg(q_word + i);
    uint32_t h = bowdb_hash(w, shift);
    while (atomicCAS(&keys[h], 0xffffffffu, w) != 0xffffffffu) h = (h + 1) & (cap - 1);  // words are distinct (std::map keys)
    pos[h] = (uint32_t)i;
    const uint32_t b = bowdb_hash(w, fshift);
    atomicOr(&filt[b >> 5], 1u << (b & 31));
  }
  __syncthreads();
  const int lane = threadIdx.x & 31;
  const uint32_t q_lo = nq > 0 ? __ldg(q_word) : 1u, q_hi = nq > 0 ? __ldg(q_word + nq - 1) : 0u;
  for (;;) {
    int slot = 0;
    if (lane == 0) slot = atomicAdd(next_slot, 1);
    slot = __shfl_sync(0xffffffffu, slot, 0);
    if (slot >= n_slots) break;
    const BowSlot rec = slots[slot];
    int cnt = 0;
    uint32_t first = 0xffffffffu;
    double acc = 0.0;
    if (rec.alive && nq > 0 && rec.len > 0) {
      const long long abase = rec.off & ~3LL;  // chunks start at 16-byte boundaries of the arena
      const int lead = (int)(rec.off - abase), span = lead + rec.len;  // entries [lead, span) of the aligned run are ours
      const uint32_t* __restrict__ wp = db_word + abase;
      int r0 = 4 * lane;  // first entry of this lane's quad, relative to abase
      uint4 nxt = r0 < span ? __ldg((const uint4*)(wp + r0)) : make_uint4(0, 0, 0, 0);
      for (int cb = 0; cb < span; cb += 128, r0 += 128) {
        const uint4 cur = nxt;
        if (r0 + 128 < span) nxt = __ldg((const uint4*)(wp + r0 + 128));
        const uint32_t w4[4] = {cur.x, cur.y, cur.z, cur.w};
        double term[4] = {0.0, 0.0, 0.0, 0.0};
        unsigned hm = 0, maybe = 0;
        // one LDS + bit test per entry settles > 99 % of them (a filter of >= 128 bits per query word)
#pragma unroll
        for (int c = 0; c < 4; c++) {
          const uint32_t b = bowdb_hash(w4[c], fshift);
          const bool in = (unsigned)(r0 + c - lead) < (unsigned)rec.len && w4[c] >= q_lo && w4[c] <= q_hi;
          if (in && ((filt[b >> 5] >> (b & 31)) & 1u)) maybe |= 1u << c;
        }
        if (__any_sync(0xffffffffu, maybe != 0)) {
#pragma unroll
          for (int c = 0; c < 4; c++) {
            if (maybe & (1u << c)) {
              const uint32_t w = w4[c];
              uint32_t h = bowdb_hash(w, shift), k;
              while ((k = keys[h]) != w && k != 0xffffffffu) h = (h + 1) & (cap - 1);
              if (k == w) {
                hm |= 1u << c;
                const double vi = __ldg(q_val + pos[h]), wi = __ldg(db_val + abase + r0 + c);
                term[c] = fabs(vi - wi) - fabs(vi) - fabs(wi);  // ScoringObject.cpp:41 (v1 = query, v2 = database keyframe)
              }
            }
          }
        }
        unsigned m = __ballot_sync(0xffffffffu, hm != 0);
        if (m) {
          if (first == 0xffffffffu) {
            const int c0 = __ffs(hm) - 1;
            const uint32_t mine = c0 == 0 ? cur.x : c0 == 1 ? cur.y : c0 == 2 ? cur.z : cur.w;
            first = __shfl_sync(0xffffffffu, mine, __ffs(m) - 1);
          }
          cnt += __reduce_add_sync(0xffffffffu, (unsigned)__popc(hm));
          while (m) {  // ascending word order: one dependent DADD per common word, as the reference's loop
            const int src = __ffs(m) - 1;
            m &= m - 1;
            const unsigned shm = __shfl_sync(0xffffffffu, hm, src);
#pragma unroll
            for (int c = 0; c < 4; c++) {
              const double t = __shfl_sync(0xffffffffu, term[c], src);
              if (shm & (1u << c)) acc += t;
            }
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
}
