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
  int2* ent;                // {keypoint index | octave << 24, distance bits}
  int32_t ent_cap;
  int4* act;                // per query with candidates: {qid, off, cnt, blocks}
  float* act_angle;
  ProjCtl* ctl;
  int32_t* g_state;  // global fallback for phase 2 (own | asg | dst | acc | blk) when it does not fit shared memory
  int32_t smem2;     // 1 = phase 2 state lives in shared memory
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
__host__ __device__ inline size_t pf_phase2_smem(int n_t, int nv) {
  return (size_t)n_t * 12 + (size_t)nv * 4 + (((size_t)n_t + 15) & ~(size_t)15) + 64;
}

__global__ void __launch_bounds__(PF_WARPS * 32) k_proj_fused(const ProjFusedArgs A) {
  extern __shared__ __align__(16) unsigned char pf_smem[];
  __shared__ int s_last, s_changed, s_count, s_pruned;
  __shared__ int s_hist[ASDM_HISTO_LENGTH];
  __shared__ int s_keep[3];
  const int t = threadIdx.x, nt = blockDim.x, lane = t & 31, warp = t >> 5;
  const int n_cells = A.g.cols * A.g.rows;

  // ------------------------------------------------------------------------------------------------ phase 1
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
      for (int b0 = 0; b0 < cnt; b0 += PF_BATCH) {
        const int nb = min(PF_BATCH, cnt - b0);
        int jv = 0;
        if (lane < nb) jv = A.ent[off + b0 + lane].x & 0xffffff;
        for (int c = 0; c < nb; c++) {
          const int j = __shfl_sync(0xffffffffu, jv, c);
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
        if (lane < nb) {
          const float4* row = reinterpret_cast<const float4*>(s_sq + lane * PF_ROW);
          float acc = 0.f;
#pragma unroll 8
          for (int k = 0; k < ASDM_DESC_DIM / 4; k++) {
            const float4 x = row[k];
            acc = __fadd_rn(acc, x.x); acc = __fadd_rn(acc, x.y); acc = __fadd_rn(acc, x.z); acc = __fadd_rn(acc, x.w);
          }
          A.ent[off + b0 + lane].y = __float_as_int(acc);
        }
        __syncwarp();
      }
      if (lane == 0) {
        const int slot = atomicAdd(&A.ctl->n_active, 1);
        A.act[slot] = make_int4(Q.qid, off, cnt, (int)(Q.row_blocks >> 31));
        A.act_angle[slot] = Q.angle;
      }
    }
  }

  // ----------------------------------------------------------------------------------- who resolves: last ticket
  __threadfence();
  __syncthreads();
  if (t == 0) s_last = atomicAdd(&A.ctl->ticket, 1u) == gridDim.x - 1 ? 1 : 0;
  __syncthreads();
  if (!s_last) return;
  __threadfence();

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
  int32_t* own = A.smem2 ? reinterpret_cast<int32_t*>(pf_smem) : A.g_state;
  int32_t* asg = own + n_t;
  float* dst = reinterpret_cast<float*>(asg + n_t);
  int32_t* acc = reinterpret_cast<int32_t*>(dst + n_t);
  uint8_t* blk = reinterpret_cast<uint8_t*>(acc + A.nv);
  if (!ovf) {
    for (int j = t; j < n_t; j += nt) {
      const uint8_t b = A.blocked0 ? A.blocked0[j] : 0;
      blk[j] = b;
      own[j] = b ? -1 : INT_MAX;  // unavailable from the start: "owned" by a query before all others
      asg[j] = -1;
      dst[j] = -1.f;
    }
    for (int a = t; a < n_act; a += nt) acc[a] = -1;
    if (A.want_q)
      for (int i = t; i < A.nq; i += nt) {
        A.r_qbest[i] = -1;
        A.r_qdist[i] = -1.f;
      }
    __syncthreads();
    for (;;) {
      if (t == 0) s_changed = 0;
      __syncthreads();
      for (int a = t; a < n_act; a += nt) {
        const int4 R = __ldcg(A.act + a);
        const int i = R.x;
        float best = A.P.best_init, best2 = A.P.best_init;
        int lvl = -1, lvl2 = -1, be = -1;
        for (int e = R.y; e < R.y + R.z; e++) {
          const int2 E = __ldcg(A.ent + e);
          const int j = E.x & 0xffffff;
          if (own[j] < i) continue;  // :86-88, :385, :1393-1395, :1524 -- taken by an earlier query, or blocked0
          const float d = __int_as_float(E.y);
          if (d < best) {
            best2 = best; best = d;
            lvl2 = lvl; lvl = (E.x >> 24) & 0xff;
            be = e;
          } else if (A.P.mode == 0 && d < best2) {  // :102-106
            lvl2 = (E.x >> 24) & 0xff;
            best2 = d;
          }
        }
        int dec = -1;
        if (be >= 0 && best <= A.P.th_accept) {  // :110, :404, :1408, :1538
          dec = be;
          if (A.P.mode == 0 && lvl == lvl2 && best > __fmul_rn(A.P.nnratio, best2)) dec = -1;  // :112
        }
        if (dec != acc[a]) {
          acc[a] = dec;
          s_changed = 1;
        }
      }
      __syncthreads();
      for (int j = t; j < n_t; j += nt) own[j] = blk[j] ? -1 : INT_MAX;
      __syncthreads();
      for (int a = t; a < n_act; a += nt) {
        const int e = acc[a];
        if (e < 0) continue;
        const int4 R = __ldcg(A.act + a);
        if (R.w) atomicMin(own + (__ldcg(A.ent + e).x & 0xffffff), R.x);
      }
      __syncthreads();
      if (!s_changed) break;
    }
    // F.mvpMapPoints[bestIdx] = pMP: the last accepting query owns the keypoint
    int my_count = 0;
    for (int a0 = 0; a0 < n_act; a0 += nt) {  // whole warps stay together for the match-any reduction
      const int a = a0 + t;
      int bin = -1;
      if (a < n_act) {
        const int e = acc[a];
        const int4 R = __ldcg(A.act + a);
        if (e >= 0) {
          const int2 E = __ldcg(A.ent + e);
          const int j = E.x & 0xffffff;
          atomicMax(asg + j, R.x);
          my_count += A.P.count_twice ? 2 : 1;  // :116-117
          if (A.P.check_ori) bin = asdm_rot_bin(__ldcg(A.act_angle + a), A.kps_t[j].angle);  // :1413-1423
          if (A.want_q) {
            A.r_qbest[R.x] = j;
            A.r_qdist[R.x] = __int_as_float(E.y);
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
      const int e = acc[a];
      if (e < 0) continue;
      const int2 E = __ldcg(A.ent + e);
      const int j = E.x & 0xffffff;
      if (asg[j] == __ldcg(A.act + a).x) dst[j] = __int_as_float(E.y);
    }
    __syncthreads();
    if (A.P.check_ori) {  // :1430-1449: every entry pushed into a pruned bin clears its keypoint
      for (int a = t; a < n_act; a += nt) {
        const int e = acc[a];
        if (e < 0) continue;
        const int j = __ldcg(A.ent + e).x & 0xffffff;
        const int b = asdm_rot_bin(__ldcg(A.act_angle + a), A.kps_t[j].angle);
        if (b != s_keep[0] && b != s_keep[1] && b != s_keep[2]) {
          asg[j] = -2;  // cleared by the rotation check, as opposed to -1 = never assigned
          dst[j] = -1.f;
        }
      }
      __syncthreads();
    }
    for (int j = t; j < n_t; j += nt) {
      A.r_assign[j] = asg[j];
      A.r_dist[j] = dst[j];
    }
    if (t == 0) A.r_hdr[2] = s_count - s_pruned;
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
