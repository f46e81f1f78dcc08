// kernels_proj_fused.cuh -- the projection-gated search as ONE launch (K3f): window gather from a shared-memory-staged
// grid bucket, exact distances, the greedy fixed point, ratio / threshold tests and the rotation histogram.
// Serves the four SearchByProjection overloads (ORBmatcher.cc:44-122, 300-413, 1318-1452, 1455-1582) and the match
// phases of Fuse / SearchBySim3 (:825-1314); replaces k_proj_gather -> k_proj_dist -> k_proj_resolve for them.
//
// Phase 1 (every CTA, one warp per query):
//   * the CTA stages the CSR cell offsets of the 64x48 bucket (Frame.h:37-38) in shared memory once;
//   * a warp scans its query's window column by column (cells (ix, y0..y1) are contiguous in the CSR, so this is the
//     reference's scan order, Frame.cc:244-270), 32 keypoint records per step from the cell-ordered copy `cell_kp`
//     ({x, y, octave, index}, one LDG.128 per lane), compacts the survivors with ballots into its segment of the entry
//     pool (one atomicAdd per query);
//   * distances: the query row sits in registers (lane l owns floats 4l..4l+3); each candidate row is read by ONE
//     coalesced 512-byte load, the 128 squared differences go through shared memory ([8][132] floats per warp) and
//     lane c adds candidate c's terms in index order with one fp32 accumulator -- ORBmatcher.cc:1637-1644 to the bit.
// Phase 2 (the CTA that takes the last ticket): the greedy query order as a parallel fixed point (as k_proj_resolve),
//   own / assign arrays in shared memory, histogram by __match_any_sync, results stored straight into the caller-visible
//   page-locked result block (no copy kernels, no memcpy nodes: a call is one launch and one wait).
// Algorithmic bytes: SURVEY 8d -- queries 32 B, target keypoints 16 B, 512 B per descriptor row touched, outputs 8 B
// per keypoint; C2 = 5.4 MB.
#pragma once
#include "kernels_proj.cuh"

struct ProjQ {  // one query that survived the reference's pre-descriptor gates, 32 B
  float u, v, r, angle;
  int32_t min_level, max_level;
  int32_t qid;          // index in the caller's arrays = position in the reference's loop (the greedy order)
  uint32_t row_blocks;  // descriptor row (bits 0..30) | blocks (bit 31)
};

struct ProjCtl {  // device-resident, zero at context creation; the last CTA leaves it zero again
  int32_t total, overflow, n_active;
  uint32_t ticket;
};

// What phase 2 needs of a query with candidates, 48 B: its identity and its PF_TOPK nearest candidates in the order
// the reference's scan would rank them (distance, then scan position). The greedy replay only ever needs a query's
// best (and second best) candidate among those not taken by earlier queries, so the short list decides unless all of
// it is taken -- then, and only then, the full candidate segment is scanned again.
#define PF_TOPK 4
struct ProjRec {
  int32_t qid, off;
  uint32_t cnt_blocks;  // candidates (bits 0..30) | blocks (bit 31)
  float angle;
  int2 top[PF_TOPK];    // {keypoint | octave << 24, distance bits}; x = -1: none
};

struct ProjFusedArgs {
  const ProjQ* q;  // page-locked staging arena (zero-copy loads) or device memory
  int32_t nv, nq;
  const float* desc_q;  // rows of the queries: device memory or page-locked host memory
  const float* desc_t;
  const asdm_kp* kps_t;
  const int4* cell_kp;
  const int32_t* cell_start;
  asdm_grid g;
  int32_t n_t;
  const uint8_t* blocked0;  // page-locked staging arena, or null
  uint32_t* g_blk;           // device copy of blocked0 (words), filled by phase 1 while the queries are in flight
  int2* ent;                // {keypoint index | octave << 24, distance bits}
  int32_t ent_cap;
  ProjRec* rec;             // [nv]
  ProjCtl* ctl;
  int32_t* g_state;  // global home of the phase-2 state when it does not fit shared memory
  int32_t* g_q;      // [2 * nq] per-query results before they leave coalesced (want_q)
  int32_t smem2;     // 1 = phase-2 state lives in shared memory
  int32_t want_q;
  // result block (page-locked): hdr = {total, overflow, nmatches, n_active}
  int32_t* r_hdr;
  int32_t* r_assign;
  float* r_dist;
  int32_t* r_qbest;
  float* r_qdist;
  ProjParamsDev P;
  ProjGate G;
};

#define PF_WARPS 16
#define PF_BATCH 8
#define PF_ROW 132  // floats per candidate row in shared memory: lanes 0..7 read 16-byte pieces 528 B apart -> 32 banks once

__host__ __device__ inline size_t pf_phase1_smem(int n_cells) {
  return (((size_t)(n_cells + 1) * 4 + 15) & ~(size_t)15) + (size_t)PF_WARPS * PF_BATCH * PF_ROW * 4;
}
// own | asg | dst [n_t] x 4 B, accj | accd [nv] x 4 B, rec [nv] x 48 B, blk [n_t] bytes
__host__ __device__ inline size_t pf_phase2_bytes(int n_t, int nv) {
  return (size_t)n_t * 12 + (size_t)nv * (8 + sizeof(ProjRec)) + (((size_t)n_t + 15) & ~(size_t)15) + 64;
}

__device__ __forceinline__ unsigned long long pf_now() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
  return t;
}
__device__ __forceinline__ bool pf_before(float d, int pos, float d2, int pos2) { return d < d2 || (d == d2 && pos < pos2); }

template <bool SMEM2>
__global__ void __launch_bounds__(PF_WARPS * 32) k_proj_fused(const ProjFusedArgs A) {
  extern __shared__ __align__(16) unsigned char pf_smem[];
  __shared__ int s_last, s_changed, s_count, s_pruned;
  __shared__ int s_hist[ASDM_HISTO_LENGTH];
  __shared__ int s_keep[3];
  const int t = threadIdx.x, nt = blockDim.x, lane = t & 31, warp = t >> 5;
  const int n_cells = A.g.cols * A.g.rows;

  const unsigned long long ts0 = pf_now();
  // ------------------------------------------------------------------------------------------------ phase 1
  if (A.blocked0) {  // blocked0 crosses PCIe once, in 4-byte pieces spread over the grid, while the queries are in flight
    const int nw = (A.n_t + 3) >> 2;
    const uint32_t* src = reinterpret_cast<const uint32_t*>(A.blocked0);
    for (int w = blockIdx.x * nt + t; w < nw; w += gridDim.x * nt) A.g_blk[w] = src[w];
  }
  int32_t* s_cell = reinterpret_cast<int32_t*>(pf_smem);
  float* s_sq = reinterpret_cast<float*>(pf_smem + (((size_t)(n_cells + 1) * 4 + 15) & ~(size_t)15)) + (size_t)warp * PF_BATCH * PF_ROW;
  if ((int)blockIdx.x * PF_WARPS < A.nv) {
    for (int c = t; c <= n_cells; c += nt) s_cell[c] = __ldg(A.cell_start + c);
    __syncthreads();
    for (int s = blockIdx.x * PF_WARPS + warp; s < A.nv; s += gridDim.x * PF_WARPS) {
      ProjQ Q;
      {  // 32 B record, the same address for every lane: two broadcast loads
        const uint4* p = reinterpret_cast<const uint4*>(A.q + s);
        const uint4 a = p[0], b = p[1];
        Q.u = __uint_as_float(a.x); Q.v = __uint_as_float(a.y); Q.r = __uint_as_float(a.z); Q.angle = __uint_as_float(a.w);
        Q.min_level = (int)b.x; Q.max_level = (int)b.y; Q.qid = (int)b.z; Q.row_blocks = b.w;
      }
      // the query row starts its trip now (it may come over PCIe) and is only needed after the gather
      const float4 q4 = *(reinterpret_cast<const float4*>(A.desc_q + (size_t)(Q.row_blocks & 0x7fffffffu) * ASDM_DESC_DIM) + lane);
      int x0, x1, y0, y1;
      if (!area_cells(A.g, Q.u, Q.v, Q.r, x0, x1, y0, y1)) continue;
      const bool check_levels = (Q.min_level > 0) || (Q.max_level >= 0);  // Frame.cc:242
      int cnt = 0;
      for (int ix = x0; ix <= x1; ix++) {
        const int b = s_cell[ix * A.g.rows + y0], e = s_cell[ix * A.g.rows + y1 + 1];
        for (int p = b; p < e; p += 32) {
          bool keep = false;
          if (p + lane < e) {
            const int4 r = __ldg(A.cell_kp + p + lane);
            const asdm_kp kp{__int_as_float(r.x), __int_as_float(r.y), 0.f, r.z};
            keep = area_keep(kp, Q.u, Q.v, Q.r, check_levels, Q.min_level, Q.max_level, A.G);
          }
          cnt += __popc(__ballot_sync(0xffffffffu, keep));
        }
      }
      if (cnt == 0) continue;
      int off = 0;
      if (lane == 0) off = atomicAdd(&A.ctl->total, cnt);
      off = __shfl_sync(0xffffffffu, off, 0);
      if (off + cnt > A.ent_cap) {  // the host reads total (= the exact demand), grows the pool and searches again
        if (lane == 0) A.ctl->overflow = 1;
        continue;
      }
      int w = off;
      for (int ix = x0; ix <= x1; ix++) {
        const int b = s_cell[ix * A.g.rows + y0], e = s_cell[ix * A.g.rows + y1 + 1];
        for (int p = b; p < e; p += 32) {
          bool keep = false;
          int4 r = make_int4(0, 0, 0, 0);
          if (p + lane < e) {
            r = __ldg(A.cell_kp + p + lane);
            const asdm_kp kp{__int_as_float(r.x), __int_as_float(r.y), 0.f, r.z};
            keep = area_keep(kp, Q.u, Q.v, Q.r, check_levels, Q.min_level, Q.max_level, A.G);
          }
          const unsigned m = __ballot_sync(0xffffffffu, keep);
          if (keep) A.ent[w + __popc(m & ((1u << lane) - 1u))].x = r.w | ((r.z & 0xff) << 24);
          w += __popc(m);
        }
      }
      __syncwarp();
      // the query's PF_TOPK nearest so far, replicated in every lane, ordered by (distance, scan position)
      int tj[PF_TOPK];
      float td[PF_TOPK];
#pragma unroll
      for (int k = 0; k < PF_TOPK; k++) { tj[k] = -1; td[k] = 0.f; }
      for (int b0 = 0; b0 < cnt; b0 += PF_BATCH) {
        const int nb = min(PF_BATCH, cnt - b0);
        int jv = 0;
        if (lane < nb) jv = A.ent[off + b0 + lane].x;
        for (int c = 0; c < nb; c++) {
          const int j = __shfl_sync(0xffffffffu, jv, c) & 0xffffff;
          const float4 t4 = __ldg(reinterpret_cast<const float4*>(A.desc_t + (size_t)j * ASDM_DESC_DIM) + lane);
          float4 sq;
          float d;
          d = __fsub_rn(q4.x, t4.x); sq.x = __fmul_rn(d, d);
          d = __fsub_rn(q4.y, t4.y); sq.y = __fmul_rn(d, d);
          d = __fsub_rn(q4.z, t4.z); sq.z = __fmul_rn(d, d);
          d = __fsub_rn(q4.w, t4.w); sq.w = __fmul_rn(d, d);
          *reinterpret_cast<float4*>(s_sq + c * PF_ROW + 4 * lane) = sq;
        }
        __syncwarp();
        float acc = 0.f;
        if (lane < nb) {
          const float4* row = reinterpret_cast<const float4*>(s_sq + lane * PF_ROW);
#pragma unroll 8
          for (int k = 0; k < ASDM_DESC_DIM / 4; k++) {
            const float4 x = row[k];
            acc = __fadd_rn(acc, x.x); acc = __fadd_rn(acc, x.y); acc = __fadd_rn(acc, x.z); acc = __fadd_rn(acc, x.w);
          }
          A.ent[off + b0 + lane].y = __float_as_int(acc);
        }
        __syncwarp();
        for (int c = 0; c < nb; c++) {  // candidates arrive in scan order: a later one never passes an equal distance
          const float dc = __shfl_sync(0xffffffffu, acc, c);
          const int jc = __shfl_sync(0xffffffffu, jv, c);
          float dd = dc;
          int jj = jc;
          bool ins = false;  // once placed, everything behind shifts down by one
#pragma unroll
          for (int k = 0; k < PF_TOPK; k++) {
            if (ins || tj[k] < 0 || dd < td[k]) {
              const float xd = td[k];
              const int xj = tj[k];
              td[k] = dd; tj[k] = jj;
              dd = xd; jj = xj;
              ins = true;
            }
          }
        }
      }
      if (lane == 0) {
        const int slot = atomicAdd(&A.ctl->n_active, 1);
        uint4* out = reinterpret_cast<uint4*>(A.rec + slot);
        out[0] = make_uint4((uint32_t)Q.qid, (uint32_t)off, (uint32_t)cnt | (Q.row_blocks & 0x80000000u), __float_as_uint(Q.angle));
        out[1] = make_uint4((uint32_t)tj[0], __float_as_uint(td[0]), (uint32_t)tj[1], __float_as_uint(td[1]));
        out[2] = make_uint4((uint32_t)tj[2], __float_as_uint(td[2]), (uint32_t)tj[3], __float_as_uint(td[3]));
      }
    }
  }

  // ----------------------------------------------------------------------------------- who resolves: last ticket
  const unsigned long long ts1 = pf_now();
  __syncthreads();
  if (t == 0) {
    __threadfence();  // cumulative: orders the whole CTA's writes (observed through the barrier) before the ticket
    s_last = atomicAdd(&A.ctl->ticket, 1u) == gridDim.x - 1 ? 1 : 0;
  }
  __syncthreads();
  if (!s_last) return;
  __threadfence();
  const unsigned long long ts2 = pf_now();

  // ------------------------------------------------------------------------------------------------ phase 2
  const int n_t = A.n_t;
  const int total = __ldcg(&A.ctl->total), ovf = __ldcg(&A.ctl->overflow), n_act = __ldcg(&A.ctl->n_active);
  if (t == 0) {
    A.r_hdr[0] = total;
    A.r_hdr[1] = ovf;
    A.r_hdr[3] = n_act;
    s_count = 0;
    s_pruned = 0;
  }
  if (t < ASDM_HISTO_LENGTH) s_hist[t] = 0;
  // SMEM2: the whole replay state lives in shared memory (the compiler sees the state space: LDS / STS / ATOMS);
  // otherwise the same layout in global memory
  unsigned char* base = SMEM2 ? pf_smem : reinterpret_cast<unsigned char*>(A.g_state);
  ProjRec* rec = reinterpret_cast<ProjRec*>(base);  // 16-byte aligned first
  int32_t* own = reinterpret_cast<int32_t*>(rec + A.nv);
  int32_t* asg = own + n_t;
  float* dst = reinterpret_cast<float*>(asg + n_t);
  int32_t* accj = reinterpret_cast<int32_t*>(dst + n_t);
  float* accd = reinterpret_cast<float*>(accj + A.nv);
  uint8_t* blk = reinterpret_cast<uint8_t*>(accd + A.nv);
  if (!ovf) {
    if (SMEM2) {  // one coalesced pass brings every record next to the threads that replay them
      const uint4* src = reinterpret_cast<const uint4*>(A.rec);
      uint4* dstv = reinterpret_cast<uint4*>(rec);
#pragma unroll 4
      for (int k = t; k < n_act * 3; k += nt) dstv[k] = __ldcg(src + k);
    } else {
      rec = A.rec;
    }
    const uint8_t* gb = reinterpret_cast<const uint8_t*>(A.g_blk);
    for (int j = t; j < n_t; j += nt) {
      const uint8_t b = A.blocked0 ? __ldcg(gb + j) : 0;
      blk[j] = b;
      own[j] = b ? -1 : INT_MAX;  // unavailable from the start: "owned" by a query before all others
      asg[j] = -1;
      dst[j] = -1.f;
    }
    for (int a = t; a < n_act; a += nt) accj[a] = -1;
    if (A.want_q)
      for (int i = t; i < 2 * A.nq; i += nt) A.g_q[i] = i < A.nq ? -1 : __float_as_int(-1.f);
    __syncthreads();
    const unsigned long long ts3 = pf_now();
    int sweeps = 0;
    for (;;) {
      sweeps++;
      if (t == 0) s_changed = 0;
      __syncthreads();
      for (int a = t; a < n_act; a += nt) {
        // Branch-free over the short list: the replay runs on ONE SM, so issue slots are what it costs. The list is
        // ordered as the reference's scan ranks candidates, hence: best = first entry not taken, second best = the next.
        const ProjRec R = rec[a];
        const int i = R.qid, cnt = (int)(R.cnt_blocks & 0x7fffffffu);
        const int want = A.P.mode == 0 ? 2 : 1;
        unsigned avail = 0;
#pragma unroll
        for (int k = 0; k < PF_TOPK; k++) {
          const int jx = R.top[k].x;
          const int o = own[jx < 0 ? 0 : (jx & 0xffffff)];  // :86-88, :385, :1393-1395, :1524 -- taken earlier, or blocked0
          avail |= (jx >= 0 && o >= i) ? (1u << k) : 0u;
        }
        const unsigned rest = avail & (avail - 1u);
        const int k1 = __ffs(avail) - 1, k2 = __ffs(rest) - 1;
        int2 E1 = R.top[0], E2 = R.top[0];
#pragma unroll
        for (int k = 1; k < PF_TOPK; k++) {
          E1 = k1 == k ? R.top[k] : E1;
          E2 = k2 == k ? R.top[k] : E2;
        }
        float best = A.P.best_init, best2 = A.P.best_init;
        int lvl = -1, lvl2 = -1, bj = -1, found = 0;
        if (k1 >= 0 && __int_as_float(E1.y) < best) {  // sorted: when the first one is not nearer than the initial best, none is
          best = __int_as_float(E1.y); lvl = (E1.x >> 24) & 0xff; bj = E1.x & 0xffffff;
          found = 1;
          if (k2 >= 0) {
            found = 2;
            if (__int_as_float(E2.y) < best2) { best2 = __int_as_float(E2.y); lvl2 = (E2.x >> 24) & 0xff; }
          }
        } else if (k1 >= 0) {
          found = want;  // nothing usable anywhere in the segment
        }
        if (found < want && cnt > PF_TOPK && (found > 0 || __int_as_float(R.top[PF_TOPK - 1].y) < A.P.best_init)) {
          // the short list is (all but) taken: the reference's scan over the full segment (:77-107)
          best = A.P.best_init; best2 = A.P.best_init;
          lvl = -1; lvl2 = -1; bj = -1;
          for (int e = R.off; e < R.off + cnt; e++) {
            const int2 E = __ldcg(A.ent + e);
            const int j = E.x & 0xffffff;
            if (own[j] < i) continue;
            const float d = __int_as_float(E.y);
            if (d < best) {
              best2 = best; best = d;
              lvl2 = lvl; lvl = (E.x >> 24) & 0xff;
              bj = j;
            } else if (A.P.mode == 0 && d < best2) {  // :102-106
              lvl2 = (E.x >> 24) & 0xff;
              best2 = d;
            }
          }
        }
        int dec = -1;
        if (bj >= 0 && best <= A.P.th_accept) {  // :110, :404, :1408, :1538
          dec = bj;
          if (A.P.mode == 0 && lvl == lvl2 && best > __fmul_rn(A.P.nnratio, best2)) dec = -1;  // :112
        }
        if (dec != accj[a]) {
          accj[a] = dec;
          s_changed = 1;
        }
        accd[a] = best;
      }
      __syncthreads();
      for (int j = t; j < n_t; j += nt) own[j] = blk[j] ? -1 : INT_MAX;
      __syncthreads();
      for (int a = t; a < n_act; a += nt) {
        const int j = accj[a];
        if (j >= 0 && (rec[a].cnt_blocks >> 31)) atomicMin(own + j, rec[a].qid);
      }
      __syncthreads();
      const int changed = s_changed;
      __syncthreads();  // everybody has read the flag before thread 0 clears it for the next sweep
      if (!changed) break;
    }
    const unsigned long long ts4 = pf_now();
    // F.mvpMapPoints[bestIdx] = pMP: the last accepting query owns the keypoint
    int my_count = 0;
    for (int a0 = 0; a0 < n_act; a0 += nt) {  // whole warps stay together for the match-any reduction
      const int a = a0 + t;
      int bin = -1;
      if (a < n_act) {
        const int j = accj[a];
        if (j >= 0) {
          const int qid = rec[a].qid;
          atomicMax(asg + j, qid);
          my_count += A.P.count_twice ? 2 : 1;  // :116-117
          if (A.P.check_ori) bin = asdm_rot_bin(rec[a].angle, A.kps_t[j].angle);  // :1413-1423
          if (A.want_q) {
            A.g_q[qid] = j;
            A.g_q[A.nq + qid] = __float_as_int(accd[a]);
          }
        }
      }
      if (A.P.check_ori) {  // one shared-memory add per distinct bin of the warp
        const unsigned peers = __match_any_sync(0xffffffffu, bin);
        if (bin >= 0 && lane == __ffs(peers) - 1) atomicAdd(&s_hist[bin], __popc(peers));
      }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) my_count += __shfl_xor_sync(0xffffffffu, my_count, o);
    if (lane == 0 && my_count) atomicAdd(&s_count, my_count);
    __syncthreads();
    if (t == 0 && A.P.check_ori) {
      asdm_three_maxima(s_hist, ASDM_HISTO_LENGTH, s_keep[0], s_keep[1], s_keep[2]);
      int pruned = 0;
      for (int b = 0; b < ASDM_HISTO_LENGTH; b++)
        if (b != s_keep[0] && b != s_keep[1] && b != s_keep[2]) pruned += s_hist[b];
      s_pruned = pruned;
    }
    for (int a = t; a < n_act; a += nt) {
      const int j = accj[a];
      if (j >= 0 && asg[j] == rec[a].qid) dst[j] = accd[a];
    }
    __syncthreads();
    if (A.P.check_ori) {  // :1430-1449: every entry pushed into a pruned bin clears its keypoint
      for (int a = t; a < n_act; a += nt) {
        const int j = accj[a];
        if (j < 0) continue;
        const int b = asdm_rot_bin(rec[a].angle, A.kps_t[j].angle);
        if (b != s_keep[0] && b != s_keep[1] && b != s_keep[2]) {
          asg[j] = -2;  // cleared by the rotation check, as opposed to -1 = never assigned
          dst[j] = -1.f;
        }
      }
      __syncthreads();
    }
    // results leave for the page-locked block in full 128-byte lines per warp
    for (int j = t; j < n_t; j += nt) {
      A.r_assign[j] = asg[j];
      A.r_dist[j] = dst[j];
    }
    if (A.want_q)
      for (int i = t; i < A.nq; i += nt) {
        A.r_qbest[i] = A.g_q[i];
        A.r_qdist[i] = __int_as_float(A.g_q[A.nq + i]);
      }
    if (t == 0) {
      A.r_hdr[2] = s_count - s_pruned;
      // where the time went (ns): phase 1 of this CTA | ticket | phase-2 set-up | sweeps | selection + histogram; sweeps
      const unsigned long long ts5 = pf_now();
      A.r_hdr[8] = (int)(ts1 - ts0); A.r_hdr[9] = (int)(ts2 - ts1); A.r_hdr[10] = (int)(ts3 - ts2);
      A.r_hdr[11] = (int)(ts4 - ts3); A.r_hdr[12] = (int)(ts5 - ts4); A.r_hdr[13] = sweeps;
    }
  }
  __syncthreads();
  if (t == 0) {  // leave the control block ready for the next call on this stream
    A.ctl->total = 0;
    A.ctl->overflow = 0;
    A.ctl->n_active = 0;
    __threadfence();
    A.ctl->ticket = 0u;
  }
}
